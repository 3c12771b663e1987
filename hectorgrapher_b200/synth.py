"""Deterministic synthetic inputs for tests and bench.py (numpy only; no compute on the hot path).

World and sensor follow the reference's own simulator recipe
(cartographer/mapping/internal/3d/local_trajectory_builder_3d_test.cc:117-221): a 30 m
axis-aligned box centred at the origin with 100 spheres of radius 0.5 m placed at radius 10 m
(directions drawn from a seeded generator), seen by a spinning lidar with elevation +-15 deg.
Grids are filled the way the reference's writers fill them (projective TSD along each ray inside
the truncation band, running weighted mean, weight cap: mapping/3d/tsdf_range_data_inserter_3d.cc:
294-343, 725-737; hit / miss odds for occupancy: mapping/3d/range_data_inserter_3d.cc:27-101),
vectorised, and encoded with the reference's float -> uint16 laws (mapping/probability_values.h:32-44,
mapping/2d/tsd_value_converter.h:39-55, mapping/3d/hybrid_grid_tsdf.h:87-93).
"""
import numpy as np

BOX_HALF = 15.0
SPHERE_RADIUS = 0.5
UPDATE_MARKER = 0x8000


def spheres(seed=42, count=100):
    rng = np.random.RandomState(seed)
    d = rng.uniform(-1, 1, size=(count, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return 10.0 * d


def quat_from_yaw(yaw):
    return np.array([np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2)])


def quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx])


def quat_rotate(q, v):
    """Rotates rows of v (n x 3) by unit quaternion q = (w, x, y, z)."""
    w, u = q[0], np.asarray(q[1:])
    uv = 2 * np.cross(u, v)
    return v + w * uv + np.cross(u, uv)


def small_rotation(rng, max_angle):
    axis = rng.uniform(-1, 1, 3)
    axis /= np.linalg.norm(axis)
    a = rng.uniform(-max_angle, max_angle)
    return np.concatenate([[np.cos(a / 2)], np.sin(a / 2) * axis])


def lidar_directions(rings, azimuths):
    el = np.deg2rad(np.linspace(-15.0, 15.0, rings))
    az = np.linspace(0.0, 2 * np.pi, azimuths, endpoint=False)
    e, a = np.meshgrid(el, az, indexing="ij")
    return np.stack([np.cos(e) * np.cos(a), np.cos(e) * np.sin(a), np.sin(e)], -1).reshape(-1, 3)


def ray_cast(origin, dirs, sph):
    """Range along each unit direction to the first hit of the box interior or a sphere."""
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (BOX_HALF - origin) / dirs
        t2 = (-BOX_HALF - origin) / dirs
    t_box = np.where(dirs > 0, t1, np.where(dirs < 0, t2, np.inf)).min(axis=1)
    best = t_box
    oc = origin[None, :] - sph  # (m, 3)
    b = dirs @ oc.T             # (n, m)
    c = (oc * oc).sum(1)[None, :] - SPHERE_RADIUS ** 2
    disc = b * b - c
    t = -b - np.sqrt(np.maximum(disc, 0.0))
    t = np.where((disc > 0) & (t > 0.05), t, np.inf).min(axis=1)
    return np.minimum(best, t)


def scan(origin, yaw, rings, azimuths, sph, rng=None, noise=0.01, max_range=40.0):
    """Returns hit points in the SENSOR frame (float32, n x 3) for a sensor at (origin, yaw)."""
    q = quat_from_yaw(yaw)
    d_sensor = lidar_directions(rings, azimuths)
    d_world = quat_rotate(q, d_sensor)
    r = ray_cast(np.asarray(origin, float), d_world, sph)
    if rng is not None and noise > 0:
        r = r + rng.normal(0.0, noise, size=r.shape)
    ok = np.isfinite(r) & (r < max_range) & (r > 0.5)
    return (d_sensor[ok] * r[ok, None]).astype(np.float32)


def corkscrew_pose(k, count):
    """Sensor poses on a corkscrew through the box (local_trajectory_builder_3d_test.cc:231-255)."""
    s = k / max(1, count - 1)
    ang = 2 * np.pi * 1.5 * s
    return np.array([4.0 * np.cos(ang), 4.0 * np.sin(ang), -2.0 + 4.0 * s]), 0.7 * ang


def round_half_away(x):
    return np.where(x >= 0, np.floor(x + 0.5), np.ceil(x - 0.5)).astype(np.int64)


def cell_index(points, resolution):
    p = np.asarray(points, np.float32) / np.float32(resolution)
    return round_half_away(p.astype(np.float64)).astype(np.int32)


def tsd_to_value(tsd, max_tsd):
    max_tsd = np.float32(max_tsd)
    res = np.float32(32766.0) / (max_tsd - (-max_tsd))
    t = np.clip(np.asarray(tsd, np.float32), -max_tsd, max_tsd)
    return (round_half_away(((t - (-max_tsd)) * res).astype(np.float64)) + 1).astype(np.uint16)


def weight_to_value(w, max_weight):
    max_weight = np.float32(max_weight)
    res = np.float32(32766.0) / max_weight
    t = np.clip(np.asarray(w, np.float32), np.float32(0), max_weight)
    return (round_half_away((t * res).astype(np.float64)) + 1).astype(np.uint16)


def probability_to_value(p):
    lo, hi = np.float32(0.1), np.float32(1.0) - np.float32(0.1)
    t = np.clip(np.asarray(p, np.float32), lo, hi)
    return (round_half_away(((t - lo) * (np.float32(32766.0) / (hi - lo))).astype(np.float64)) + 1).astype(np.uint16)


def _unique_cells(cells):
    key = (cells[:, 0].astype(np.int64) + 32768) << 34 | (cells[:, 1].astype(np.int64) + 32768) << 17 | (
        cells[:, 2].astype(np.int64) + 32768)
    uniq, inv = np.unique(key, return_inverse=True)
    first = np.zeros(len(uniq), np.int64)
    first[inv[::-1]] = np.arange(len(key))[::-1]
    return cells[first], inv, len(uniq)


def build_tsdf(resolution, sensor_poses, rings, azimuths, sph, relative_truncation_distance=2.5,
               max_weight=1000.0, seed=7):
    """Returns (xyz int32 [n,3], tsd uint16 [n], weight uint16 [n]) of a TSDF submap in the world frame."""
    rng = np.random.RandomState(seed)
    T = relative_truncation_distance * resolution
    all_cells, all_tsd, all_w = [], [], []
    steps = np.arange(-T, T + 0.5 * resolution, 0.5 * resolution)
    for origin, yaw in sensor_poses:
        origin = np.asarray(origin, float)
        pts = scan(origin, yaw, rings, azimuths, sph, rng)
        world = quat_rotate(quat_from_yaw(yaw), pts.astype(float)) + origin
        ray = world - origin
        rng_len = np.linalg.norm(ray, axis=1)
        dirs = ray / rng_len[:, None]
        samples = origin[None, None, :] + (rng_len[:, None] + steps[None, :])[:, :, None] * dirs[:, None, :]
        cells = cell_index(samples.reshape(-1, 3), resolution)
        ray_id = np.repeat(np.arange(len(pts)), len(steps))
        # one update per (ray, cell)
        key = np.concatenate([ray_id[:, None], cells], axis=1)
        _, first = np.unique(key, axis=0, return_index=True)
        cells, ray_id = cells[first], ray_id[first]
        centers = cells.astype(np.float32) * np.float32(resolution)
        dist = np.linalg.norm(centers.astype(float) - origin, axis=1)
        tsd = np.clip(rng_len[ray_id] - dist, -T, T)
        norm = tsd / T
        w = np.where(norm < -0.1, np.exp(-4.0 * (-norm - 0.1) ** 2), 1.0)
        all_cells.append(cells)
        all_tsd.append(tsd * w)
        all_w.append(w)
    cells = np.concatenate(all_cells)
    uniq, inv, n = _unique_cells(cells)
    sw = np.bincount(inv, weights=np.concatenate(all_w), minlength=n)
    st = np.bincount(inv, weights=np.concatenate(all_tsd), minlength=n)
    tsd = st / sw
    weight = np.minimum(sw, max_weight)
    tv = tsd_to_value(tsd, T) | np.uint16(UPDATE_MARKER)  # SetCell leaves the marker set
    wv = weight_to_value(weight, max_weight)
    return uniq.astype(np.int32), tv.astype(np.uint16), wv.astype(np.uint16)


def build_occupancy(resolution, sensor_poses, rings, azimuths, sph, hit=0.55, miss=0.49, free=2, seed=7):
    """Returns (xyz int32 [n,3], value uint16 [n]) of an occupancy submap in the world frame."""
    rng = np.random.RandomState(seed)
    hits, misses = [], []
    for origin, yaw in sensor_poses:
        origin = np.asarray(origin, float)
        pts = scan(origin, yaw, rings, azimuths, sph, rng)
        world = quat_rotate(quat_from_yaw(yaw), pts.astype(float)) + origin
        hc = cell_index(world, resolution)
        oc = cell_index(origin[None, :], resolution)[0]
        delta = hc - oc
        n = np.abs(delta).max(axis=1)
        hits.append(np.unique(hc, axis=0))
        ms = []
        for k in range(1, free + 1):
            pos = np.maximum(n - k, 0)
            ms.append(oc + (delta * pos[:, None]) // np.maximum(n, 1)[:, None])
        misses.append(np.unique(np.concatenate(ms), axis=0))
    cells = np.concatenate(hits + misses)
    is_hit = np.concatenate([np.ones(sum(len(h) for h in hits)), np.zeros(sum(len(m) for m in misses))])
    uniq, inv, n = _unique_cells(cells)
    nh = np.bincount(inv, weights=is_hit, minlength=n)
    nm = np.bincount(inv, weights=1 - is_hit, minlength=n)
    # hits take priority within one scan (range_data_inserter_3d.cc:92-94); approximate by odds products
    odds = (hit / (1 - hit)) ** nh * (miss / (1 - miss)) ** np.where(nh > 0, 0, nm)
    p = odds / (1 + odds)
    return uniq.astype(np.int32), probability_to_value(p)


def rotational_histogram(points, size=120):
    """RotationalScanMatcher::ComputeHistogram (rotational_scan_matcher.cc:40-170), numpy restatement
    used only to produce plausible histogram inputs."""
    pts = np.asarray(points, np.float32)
    hist = np.zeros(size, np.float32)
    slices = round_half_away((pts[:, 2] / np.float32(0.2)).astype(np.float64))
    for s in np.unique(slices):
        sl = pts[slices == s]
        c = sl.mean(axis=0)
        d = sl[:, :2] - c[:2]
        keep = np.linalg.norm(d, axis=1) >= 0.2
        sl, d = sl[keep], d[keep]
        if len(sl) == 0:
            continue
        sl = sl[np.argsort(np.arctan2(d[:, 1], d[:, 0]), kind="stable")]
        c = sl.mean(axis=0)
        last = sl[0]
        for p in sl:
            delta = (p - last)[:2]
            direction = (p - c)[:2]
            dist = np.linalg.norm(delta)
            if dist < 0.2 or np.linalg.norm(direction) < 0.2:
                continue
            if dist > 0.9:
                last = p
                continue
            angle = np.arctan2(delta[1], delta[0])
            value = max(0.0, 1.0 - abs(np.dot(delta / dist, direction / np.linalg.norm(direction))))
            while angle > np.pi:
                angle -= np.pi
            while angle < 0:
                angle += np.pi
            b = int(np.clip(round_half_away(np.array([size * angle / np.pi - 0.5]))[0], 0, size - 1))
            hist[b] += value
    return hist


def rotate_histogram(histogram, angle):
    """RotationalScanMatcher::RotateHistogram (rotational_scan_matcher.cc:121-143) in float32: how a
    submap's histogram is made from a node's (inputs for the benchmarks and tests only)."""
    h = np.asarray(histogram, np.float32)
    size = len(h)
    rotate_by = np.float32((-np.float32(angle) * size) / np.pi)
    full = int(round_half_away(np.float64(rotate_by - np.float32(0.5))))
    frac = np.float32(rotate_by - np.float32(full))
    full %= size
    idx = np.arange(size)
    return (frac * h[(idx + 1 + full) % size] + (np.float32(1) - frac) * h[(idx + full) % size]).astype(np.float32)


def voxel_downsample(points, size, max_points=None, seed=0):
    """One point per voxel of edge 'size' (the role of sensor::VoxelFilter before the matcher)."""
    c = cell_index(points, size)
    _, first = np.unique(c, axis=0, return_index=True)
    out = np.asarray(points, np.float32)[np.sort(first)]
    if max_points is not None and len(out) > max_points:
        rng = np.random.RandomState(seed)
        out = out[np.sort(rng.choice(len(out), max_points, replace=False))]
    return out
