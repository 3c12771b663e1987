"""The reference has no usable test for its TSDF lookups and functors
(interpolated_tsdf_space_cost_function_3d_test.cc:148-194 is hollowed out), so these pin the oracle to
what the reference's CODE states, independently of the oracle's own arithmetic:

  * InterpolatedTSDF (interpolated_tsdf.h:71-116): at a voxel centre the interpolant is that voxel's
    TSD; between two observed voxels it is linear; unobserved corners are skipped by the weight rule
    (:29-46); all eight unobserved gives min_tsd (:100-104);
  * multi-resolution lookup (interpolated_multi_resolution_tsdf.h:82-137): the fine level where all
    eight corners are observed, otherwise the coarse one;
  * every functor's Jacobian (Jets in the reference) equals central differences of its residuals;
  * residual definitions: s * tsd (tsdf_space_cost_function_3d.h:76), s * (1 - p)
    (occupied_space_cost_function_3d.h:77), and the interpolated functors evaluate at
    InterpolateTransform(pose0, pose1, ratio).
CPU only."""
import numpy as np
import pytest

RES = 0.1


def dense_tsdf(oracle, fn, lo=-6, hi=7, res=RES):
    """A TSDF grid with every voxel of [lo, hi)^3 observed, tsd = fn(centre), weight 1."""
    g = oracle.Grid.tsdf(res, 2.5, 1000.0)
    r = np.arange(lo, hi)
    xyz = np.array([(x, y, z) for z in r for y in r for x in r], np.int32)
    centres = xyz.astype(np.float32) * np.float32(res)
    g.set_cell(xyz, fn(centres).astype(np.float32), np.ones(len(xyz), np.float32))
    return g, xyz


def plane(c):
    return 0.05 * c[:, 0] - 0.08 * c[:, 1] + 0.11 * c[:, 2] + 0.01


def test_interpolated_tsdf_at_voxel_centres_and_linear_between(oracle):
    g, xyz = dense_tsdf(oracle, plane)
    inner = xyz[(np.abs(xyz) <= 4).all(axis=1)][::7]
    centres = inner.astype(np.float32).astype(np.float64) * np.float64(np.float32(RES))
    stored = g.lookup("tsd", inner).astype(np.float64)
    quantum = 2 * 0.25 / 32766
    assert np.abs(g.interpolate(centres) - stored).max() < 1e-6
    # trilinear interpolation reproduces an affine field up to the uint16 quantisation of the corners
    rng = np.random.RandomState(0)
    pts = rng.uniform(-0.38, 0.38, (500, 3))
    want = plane(pts)
    assert np.abs(g.interpolate(pts) - want).max() < 1.01 * quantum
    # and is exactly linear along an axis inside one cell
    a, b = np.array([0.123, 0.21, -0.107]), np.array([0.177, 0.21, -0.107])
    va, vb, vm = g.interpolate([a, b, 0.5 * (a + b)])
    assert abs(vm - 0.5 * (va + vb)) < 1e-12


def test_interpolated_tsdf_weight_rule(oracle):
    """interpolated_tsdf.h:29-46 and :100-104."""
    g = oracle.Grid.tsdf(RES, 2.5, 1000.0)
    # one observed corner only: the interpolant is that corner's value everywhere in the cell
    g.set_cell([[2, 3, 4]], [0.125], [1.0])
    v = g.lookup("tsd", [[2, 3, 4]])[0]
    p = np.array([[0.2 + 0.03, 0.3 + 0.06, 0.4 + 0.08]])
    assert abs(g.interpolate(p)[0] - v) < 1e-9
    # a cell with no observed corner: min_tsd = -relative_truncation_distance * resolution
    assert abs(g.interpolate([[-0.53, -0.51, -0.52]])[0] - (-0.25)) < 1e-6
    # two observed corners along x, the rest unobserved: linear in x, independent of y and z
    g.set_cell([[3, 3, 4]], [-0.05], [1.0])
    v0, v1 = g.lookup("tsd", [[2, 3, 4], [3, 3, 4]]).astype(np.float64)
    x0 = float(np.float32(2) * np.float32(RES))
    x1 = float(np.float32(x0) + np.float32(RES))   # x2 = float(centre + resolution), interpolated_tsdf.h:177
    for f in (0.1, 0.5, 0.9):
        for dy, dz in ((0.01, 0.02), (0.07, 0.09)):
            got = g.interpolate([[x0 + f * (x1 - x0), 0.3 + dy, 0.4 + dz]])[0]
            assert abs(got - (v0 + f * (v1 - v0))) < 1e-7


def test_multi_resolution_falls_through_to_the_coarse_level(oracle):
    """interpolated_multi_resolution_tsdf.h:82-137, through the two-level functor."""
    fine, _ = dense_tsdf(oracle, plane, lo=-3, hi=4)               # observed only near the origin
    coarse, _ = dense_tsdf(oracle, lambda c: plane(c) + 0.05, lo=-4, hi=5, res=0.45)
    pts = np.array([[0.03, -0.04, 0.05], [0.9, 0.8, -0.7]], np.float32)  # inside / outside the fine block
    pose = np.array([0, 0, 0, 1, 0, 0, 0], float)
    r = oracle.Cost("multires_tsdf", 1.0, pts, [fine, coarse]).evaluate(pose, jacobians=False)[0]
    assert abs(r[0] - plane(pts[:1].astype(np.float64))[0]) < 1e-4
    assert abs(r[1] - (plane(pts[1:].astype(np.float64))[0] + 0.05)) < 2e-4
    only_fine = oracle.Cost("tsdf", 1.0, pts, [fine]).evaluate(pose, jacobians=False)[0]
    assert abs(only_fine[0] - r[0]) < 1e-12 and abs(only_fine[1] - (-0.25)) < 1e-6


KINDS = ["occupied", "interpolated_occupied", "tsdf", "interpolated_tsdf", "interpolated_tsdf_per_point",
         "multires_tsdf", "interpolated_multires_tsdf", "interpolated_multires_tsdf_per_point"]


@pytest.mark.parametrize("kind", KINDS)
def test_functor_jacobians_equal_central_differences(oracle, kind):
    rng = np.random.RandomState(5)
    fine, _ = dense_tsdf(oracle, lambda c: 0.2 * np.sin(3 * c[:, 0]) * np.cos(2 * c[:, 1]) + 0.1 * c[:, 2])
    coarse, _ = dense_tsdf(oracle, lambda c: 0.2 * np.sin(3 * c[:, 0]) * np.cos(2 * c[:, 1]) + 0.1 * c[:, 2], -3, 4, 0.45)
    prob = oracle.Grid.prob(RES)
    r_ = np.arange(-6, 7)
    xyz = np.array([(x, y, z) for z in r_ for y in r_ for x in r_], np.int32)
    prob.set_probability(xyz, (0.5 + 0.3 * np.sin(0.9 * xyz[:, 0]) * np.cos(0.7 * xyz[:, 1] + 0.3 * xyz[:, 2])).astype(np.float32))
    grids = [prob] if "occupied" in kind else ([fine, coarse] if "multires" in kind else [fine])
    two = kind.startswith("interpolated")
    pts = rng.uniform(-0.2, 0.2, (1 if kind.endswith("per_point") else 40, 3)).astype(np.float32)
    cost = oracle.Cost(kind, 0.7, pts, grids, 0.37 if two else 0.0)

    def pose():
        q = np.array([1.0, 0, 0, 0]) + rng.uniform(-0.05, 0.05, 4)
        return np.concatenate([rng.uniform(-0.05, 0.05, 3), q])
    p = np.concatenate([pose(), pose()]) if two else pose()
    r, J = cost.evaluate(p)
    J = np.hstack(J)
    eps = 1e-6
    worst = 0.0
    for c in range(len(p)):
        d = np.zeros(len(p))
        d[c] = eps
        num = (cost.evaluate(p + d, jacobians=False)[0] - cost.evaluate(p - d, jacobians=False)[0]) / (2 * eps)
        worst = max(worst, np.abs(J[:, c] - num).max())
    # voxel values are piecewise (tri)linear / cubic: away from cell faces the difference quotient is exact
    # to O(eps^2); a point within eps of a face can differ by the kink — none at these seeds.
    assert worst < 5e-6, worst
    assert np.abs(J).max() > 1e-3


def test_residual_definitions(oracle):
    g, _ = dense_tsdf(oracle, plane)
    pts = np.array([[0.11, -0.07, 0.05], [-0.2, 0.1, 0.13]], np.float32)
    ident = np.array([0, 0, 0, 1, 0, 0, 0], float)
    s = 0.3
    r = oracle.Cost("tsdf", s, pts, [g]).evaluate(ident, jacobians=False)[0]
    assert np.abs(r - s * g.interpolate(pts.astype(np.float64))).max() < 1e-12
    # a translated pose evaluates the field at the translated points
    shift = np.array([0.05, -0.02, 0.03, 1, 0, 0, 0], float)
    r2 = oracle.Cost("tsdf", s, pts, [g]).evaluate(shift, jacobians=False)[0]
    assert np.abs(r2 - s * g.interpolate(pts.astype(np.float64) + shift[:3])).max() < 1e-12
    # the interpolated functor evaluates at InterpolateTransform(pose0, pose1, ratio)
    p0 = np.array([0.02, 0.01, -0.03, 0.999, 0.01, -0.02, 0.03])
    p1 = np.array([-0.04, 0.05, 0.02, 0.995, -0.03, 0.02, 0.08])
    mid = oracle.interpolate_transform(p0, p1, 0.37)
    r3 = oracle.Cost("interpolated_tsdf", s, pts, [g], 0.37).evaluate(np.concatenate([p0, p1]), jacobians=False)[0]
    r4 = oracle.Cost("tsdf", s, pts, [g]).evaluate(mid, jacobians=False)[0]
    assert np.abs(r3 - r4).max() < 1e-12
    prob = oracle.Grid.prob(RES)
    prob.set_probability([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 0], [1, 0, 1], [0, 1, 1], [1, 1, 1]], 0.8)
    pr = oracle.Cost("occupied", s, np.array([[0.05, 0.05, 0.05]], np.float32), [prob]).evaluate(ident, jacobians=False)[0]
    assert abs(pr[0] - s * (1.0 - prob.interpolate([[0.05, 0.05, 0.05]])[0])) < 1e-12


def test_branch_and_bound_returns_the_best_leaf_of_an_exhaustive_search(oracle):
    """fast_correlative_scan_matcher_3d.cc:408-471 claims the branch-and-bound returns the best-scoring
    leaf (full-resolution candidate) above min_score that passes the low-resolution gate. Checked by brute
    force: every (scan, x, y, z) leaf of a small window is scored, gated and the maximum compared."""
    from hectorgrapher_b200 import synth
    sph = synth.spheres()
    hi = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    lo = oracle.Grid.tsdf(0.45, 2.5, 1000.0)
    for k in range(2):
        origin, yaw = synth.corkscrew_pose(k, 2)
        cloud = synth.scan(origin, yaw, 24, 300, sph, np.random.RandomState(50 + k))
        world = (synth.quat_rotate(synth.quat_from_yaw(yaw), cloud.astype(np.float64)) + origin).astype(np.float32)
        hi.insert_tsdf(origin.astype(np.float32), world, epsilon=1.0)
        lo.insert_tsdf(origin.astype(np.float32), world, epsilon=1.0)
    origin, yaw = synth.corkscrew_pose(1, 2)
    cloud = synth.scan(origin, yaw, 12, 40, sph, np.random.RandomState(77))[:400]
    high = synth.voxel_downsample(cloud, 0.2, 200, seed=6)
    low = synth.voxel_downsample(cloud, 0.9, 50, seed=7)
    hist = np.zeros(8, np.float32)
    opts = dict(branch_and_bound_depth=4, full_resolution_depth=2, use_rotational_scan_matcher=False,
                min_low_resolution_score=0.3, linear_xy_search_window=0.6, linear_z_search_window=0.3,
                angular_search_window=0.04)
    m = oracle.Fcsm(hi, lo, hist, **opts)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    rng = np.random.RandomState(3)
    found = 0
    for trial in range(3):
        node = np.concatenate([origin + rng.uniform(-0.3, 0.3, 3) * [1, 1, 0.3], synth.quat_from_yaw(yaw + rng.uniform(-0.02, 0.02))])
        min_score = 0.2
        res = m.match(node, submap, high, low, hist, min_score=min_score)
        wxy, wz = 6, 3
        poses, rot, cells = m.discrete_scans(wxy, wz, 0.04, node, submap, high, hist)
        S = len(poses)
        assert S == int(res["num_scans"]) and S >= 3
        cand = np.array([(s, x, y, z) for s in range(S) for z in range(-wz, wz + 1) for y in range(-wxy, wxy + 1)
                         for x in range(-wxy, wxy + 1)], np.int32)
        sums, scores, _ = m.score_candidates(wxy, wz, 0.04, node, submap, high, hist, 0, cand)
        order = np.argsort(-scores, kind="stable")
        best = None
        for i in order:
            if not scores[i] > min_score:
                break
            s, x, y, z = cand[i]
            pose = poses[s].copy()
            pose[:3] = np.float32(0.1) * cand[i][1:].astype(np.float32) + pose[:3]
            if oracle.low_resolution_score(lo, low, pose) >= 0.3:
                best = i
                break
        if best is None:
            assert res["found"] == 0.0
            continue
        assert res["found"] == 1.0
        assert np.float32(res["score"]) == scores[best]
        if (scores == scores[best]).sum() == 1:  # a unique maximum identifies the candidate
            assert (int(res["scan_index"]), int(res["ox"]), int(res["oy"]), int(res["oz"])) == tuple(cand[best])
        assert res["num_scored"] < len(cand)  # the search prunes
        found += 1
    assert found >= 2  # the comparison is not vacuous
