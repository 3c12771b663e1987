#!/usr/bin/env python
"""Generates tests/golden/r1_golden.npz: inputs and outputs of the CPU oracle for one small
instance of every stage of the path, frozen so that
  * the GPU library can be checked without building or trusting the oracle at test time, and
  * a change in the oracle itself (a refactor that moves a rounding) is detected.

The reference cannot be built or imported here, so the vectors come from the
oracle, which is pinned to the reference by its own known-answer tests
(tests/test_oracle_reference_kats.py). Regenerate with:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from hectorgrapher_b200 import synth  # noqa: E402  (numpy-only world / lidar simulator)
from oracle import pyoracle  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "r1_golden.npz")

KINDS = ["occupied", "interpolated_occupied", "tsdf", "interpolated_tsdf", "interpolated_tsdf_per_point",
         "multires_tsdf", "interpolated_multires_tsdf", "interpolated_multires_tsdf_per_point"]
KIND_GRIDS = {"occupied": ["prob"], "interpolated_occupied": ["prob"], "tsdf": ["hi"], "interpolated_tsdf": ["hi"],
              "interpolated_tsdf_per_point": ["hi"], "multires_tsdf": ["hi", "lo"],
              "interpolated_multires_tsdf": ["hi", "lo"], "interpolated_multires_tsdf_per_point": ["hi", "lo"]}
TSDF_OPTS = dict(relative_truncation_distance=2.5, maximum_weight=1000.0, num_free_space_voxels=0, epsilon=0.1,
                 sigma=4.0, min_range=0.4, max_range=60.0, insertion_ratio=1.0)
FCSM_OPTS = dict(branch_and_bound_depth=5, full_resolution_depth=3, use_rotational_scan_matcher=False,
                 min_low_resolution_score=0.3, linear_xy_search_window=1.5, linear_z_search_window=0.5,
                 angular_search_window=0.1)
RESULT_KEYS = ["found", "score", "tx", "ty", "tz", "qw", "qx", "qy", "qz", "rotational_score",
               "low_resolution_score", "scan_index", "ox", "oy", "oz", "num_scans"]


def sort_voxels(xyz, *values):
    order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
    return (xyz[order],) + tuple(v[order] for v in values)


def voxel_digest(xyz, *values):
    """[count, 4 x uint64 of a SHA-256] over the voxels sorted by (z, y, x): the voxel lists themselves
    would be megabytes, the digest pins every bit of them."""
    import hashlib
    parts = sort_voxels(np.ascontiguousarray(xyz, np.int32), *[np.ascontiguousarray(v) for v in values if v is not None])
    h = hashlib.sha256()
    for a in parts:
        h.update(np.ascontiguousarray(a).tobytes())
    return np.concatenate([[len(xyz)], np.frombuffer(h.digest(), np.uint64)]).astype(np.uint64)


def main():
    pyoracle.build()
    g = {}
    sph = synth.spheres()
    rng = np.random.RandomState(2024)

    # 1. cell indices, including values on and next to half-cell boundaries
    for res in (0.05, 0.1, 0.45):
        p = rng.uniform(-60, 60, (1024, 3)).astype(np.float32)
        k = rng.randint(-800, 800, (512, 3))
        edge = ((k + 0.5) * np.float32(res)).astype(np.float32)
        edge = np.concatenate([edge, np.nextafter(edge, np.float32(1e9)), np.nextafter(edge, np.float32(-1e9))])
        pts = np.concatenate([p, edge]).astype(np.float32)
        g["cell_points_%g" % res] = pts
        g["cell_index_%g" % res] = pyoracle.get_cell_index(res, pts)

    # 2. submap grids written by the oracle's inserters from two small scans
    scans = []
    for k in range(2):
        origin, yaw = synth.corkscrew_pose(k, 2)
        cloud = synth.scan(origin, yaw, 24, 300, sph, np.random.RandomState(50 + k))
        world = (synth.quat_rotate(synth.quat_from_yaw(yaw), cloud.astype(np.float64)) + origin).astype(np.float32)
        scans.append((origin.astype(np.float32), world))
        g["scan_origin_%d" % k] = scans[-1][0]
        g["scan_points_%d" % k] = world
    hi = pyoracle.Grid.tsdf(0.10, 2.5, 1000.0)
    lo = pyoracle.Grid.tsdf(0.45, 2.5, 1000.0)
    prob = pyoracle.Grid.prob(0.10)
    for origin, pts in scans:
        hi.insert_tsdf(origin, pts, **TSDF_OPTS)
        lo.insert_tsdf(origin, pts, **TSDF_OPTS)
        prob.insert_occupancy(origin, pts, 0.55, 0.49, 2)
    grids = {"hi": hi, "lo": lo, "prob": prob}
    for name in ("hi", "lo", "prob"):
        g[name + "_digest"] = voxel_digest(*grids[name].export())
        g[name + "_grid_size"] = np.int32(grids[name].grid_size)
    # a slice of the high-resolution grid in the clear, for eyeballing a mismatch
    xyz, tsd, w = sort_voxels(*hi.export())
    g["hi_sample_xyz"], g["hi_sample_tsd"], g["hi_sample_weight"] = xyz[::97], tsd[::97], w[::97]

    # 3. the eight cost functors at one pose (pair)
    origin, yaw = synth.corkscrew_pose(1, 2)
    cloud = synth.scan(origin, yaw, 12, 40, sph, np.random.RandomState(77))[:400]
    g["cloud"] = cloud
    pose0 = np.concatenate([origin + [0.04, -0.03, 0.02], synth.quat_from_yaw(yaw + 0.015)])
    pose1 = np.concatenate([origin + [-0.2, 0.1, 0.05], synth.quat_mul(synth.quat_from_yaw(yaw - 0.1),
                                                                       synth.small_rotation(rng, 0.1))])
    g["pose0"], g["pose1"] = pose0, pose1
    for kind in KINDS:
        two = kind.startswith("interpolated")
        c = cloud[100:101] if kind.endswith("per_point") else cloud
        cost = pyoracle.Cost(kind, 0.7 / np.sqrt(len(c)), c, [grids[n] for n in KIND_GRIDS[kind]], 0.37 if two else 0.0)
        r, J = cost.evaluate(np.concatenate([pose0, pose1]) if two else pose0)
        g["cost_%s_r" % kind] = r
        for b, Jb in enumerate(J):
            g["cost_%s_J%d" % (kind, b)] = Jb

    # 4. precomputation pyramid of the high-resolution TSDF grid
    stack = pyoracle.Stack(hi, 5, 3)
    for d in range(5):
        g["stack_%d_digest" % d] = voxel_digest(*stack.export(d))

    # 5. correlative match + refine
    high = synth.voxel_downsample(cloud, 0.2, 256, seed=6)
    low = synth.voxel_downsample(cloud, 0.9, 64, seed=7)
    g["node_high"], g["node_low"] = high, low
    hist = np.zeros(8, np.float32)
    matcher = pyoracle.Fcsm(hi, lo, hist, **FCSM_OPTS)
    node = np.concatenate([origin + [0.6, -0.5, 0.1], synth.quat_from_yaw(yaw + 0.04)])
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    g["node_pose"] = node
    res = matcher.match(node, submap, high, low, hist, min_score=0.15)
    print("golden match:", res)
    g["match"] = np.array([res[k] for k in RESULT_KEYS], np.float64)
    assert res["found"] == 1.0, "the golden match must exist"
    start = np.array([res[k] for k in ("tx", "ty", "tz", "qw", "qx", "qy", "qz")])
    pose, summary, _ = pyoracle.scan_match([high, low], [hi, lo], [5.0, 30.0], 10.0, 1.0, start,
                                           max_num_iterations=10, use_nonmonotonic_steps=False)
    g["refined_pose"] = pose

    # 6. voxel filter
    g["voxel_filter_kept"] = pyoracle.voxel_filter(0.5, cloud)

    np.savez_compressed(OUT, **g)
    print("wrote %s (%d arrays, %.1f KB)" % (OUT, len(g), os.path.getsize(OUT) / 1e3))


if __name__ == "__main__":
    main()
