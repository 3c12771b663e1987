"""Pins the CPU oracle against the reference's own known-answer / property tests.

Each test restates one test of /root/reference (file:line in the docstring) on the
oracle. Where the reference draws from std::mt19937 + libstdc++ distributions the
same sizes, ranges and seeds are used with numpy's generator: the reference tests
are property tests and hold for any draw.
"""
import numpy as np
import pytest


def test_get_cell_index_known_answers(oracle):
    """cartographer/mapping/3d/hybrid_grid_test.cc:91-113 (resolution 2.0)."""
    pts = np.array([[0, 0, 0], [0, 26, 10], [14, 0, 10], [14, 26, 0],
                    [8.5, 11.5, 0.5], [7.5, 12.5, 1.5], [6.5, 14.5, 2.5], [5.5, 13.5, 3.5]], np.float32)
    want = np.array([[0, 0, 0], [0, 13, 5], [7, 0, 5], [7, 13, 0],
                     [4, 6, 0], [4, 6, 1], [3, 7, 1], [3, 7, 2]], np.int32)
    assert (oracle.get_cell_index(2.0, pts) == want).all()
    # half away from zero on the negative side as well (std::lround)
    neg = oracle.get_cell_index(2.0, -pts)
    assert (neg == -want).all()


def test_get_center_of_cell(oracle):
    """hybrid_grid_test.cc:115-125."""
    c = oracle.get_center_of_cell(2.0, [[3, 2, 1]])
    assert np.allclose(c, [[6, 4, 2]], atol=1e-6)
    assert (oracle.get_cell_index(2.0, c) == [[3, 2, 1]]).all()


def test_get_probability(oracle):
    """hybrid_grid_test.cc:71-87."""
    g = oracle.Grid.prob(1.0)
    idx = g.cell_index([[0, 1, 1]])
    g.set_probability(idx, 0.9)
    assert abs(g.lookup("probability", idx)[0] - 0.9) < 1e-6
    others = g.cell_index([[0, 2, 1], [1, 1, 1], [1, 2, 1]])
    assert (g.lookup("raw_a", others) == 0).all()


def test_apply_odds(oracle):
    """hybrid_grid_test.cc:29-69."""
    g = oracle.Grid.prob(1.0)
    for c in [(0, 0, 0), (0, 1, 0), (1, 0, 0), (1, 1, 0), (0, 0, 1), (0, 1, 1), (1, 0, 1), (1, 1, 1)]:
        assert g.lookup("raw_a", [c])[0] == 0
    g.set_probability([[1, 0, 1]], 0.5)
    g.apply_odds([1, 0, 1], 0.9)
    g.finish_update()
    assert g.lookup("probability", [[1, 0, 1]])[0] > 0.5
    g.set_probability([[0, 1, 0]], 0.5)
    g.apply_odds([0, 1, 0], 0.1)
    g.finish_update()
    assert g.lookup("probability", [[0, 1, 0]])[0] < 0.5
    g.apply_odds([1, 1, 1], 0.42)
    assert abs(g.lookup("probability", [[1, 1, 1]])[0] - 0.42) < 1e-4
    assert not g.apply_odds([1, 1, 1], 0.9)  # ignored until FinishUpdate
    assert abs(g.lookup("probability", [[1, 1, 1]])[0] - 0.42) < 1e-4
    g.finish_update()
    g.apply_odds([1, 1, 1], 0.9)
    assert g.lookup("probability", [[1, 1, 1]])[0] > 0.42


def test_random_grid_iteration_equals_value(oracle):
    """hybrid_grid_test.cc:127-183: 10 000 random cells in [-3000, 2999], seed 1285120005."""
    rng = np.random.RandomState(1285120005 % (2 ** 32))
    xyz = rng.randint(-3000, 3000, size=(10000, 3)).astype(np.int32)
    p = rng.uniform(0.1, 0.9, size=10000).astype(np.float32)
    g = oracle.Grid.prob(2.0)
    # std::map semantics: the first value for a key wins (emplace); keep unique keys.
    _, first = np.unique(xyz, axis=0, return_index=True)
    xyz, p = xyz[np.sort(first)], p[np.sort(first)]
    g.set_probability(xyz, p)
    ex, ev = g.export()
    assert len(ex) == len(xyz)
    # iterator value == value(index)
    assert (g.lookup("raw_a", ex) == ev).all()
    order_a = np.lexsort(ex.T[::-1])
    order_b = np.lexsort(xyz.T[::-1])
    assert (ex[order_a] == xyz[order_b]).all()
    assert np.abs(g.lookup("probability", ex)[order_a] - p[order_b]).max() < 1e-4
    assert g.grid_size == 8192


def test_tsdf_set_cell(oracle):
    """cartographer/mapping/3d/hybrid_grid_tsdf_test.cc:31-53."""
    g = oracle.Grid.tsdf(1.0, 0.5, 10.0)  # max_tsd = 0.5 * 1.0
    g.set_cell([[1, 0, 1]], 0.1, 0.5)
    assert abs(g.lookup("tsd", [[1, 0, 1]])[0] - 0.1) < 1e-4
    assert abs(g.lookup("weight", [[1, 0, 1]])[0] - 0.5) < 1e-3
    assert g.lookup("tsd", [[5, 5, 5]])[0] == -0.5
    assert g.lookup("weight", [[5, 5, 5]])[0] == 0.0
    # SetCell leaves the update marker in the stored TSD (hybrid_grid_tsdf.h:87-93)
    assert int(g.lookup("raw_a", [[1, 0, 1]])[0]) & 0x8000


def test_value_conversion_table_law(oracle):
    """cartographer/mapping/value_conversion_tables_test.cc:45-63 (exact equality)."""
    for unknown, lo, hi in [(0.1, 0.1, 0.9), (0.9, 0.1, 0.9), (-0.125, -0.125, 0.125), (0.0, 0.0, 1000.0)]:
        t = oracle.value_table(unknown, lo, hi)
        lo32, hi32 = np.float32(lo), np.float32(hi)
        scale = np.float32((hi32 - lo32) / np.float32(32766.0))
        assert t[0] == np.float32(unknown)
        for i in [1, 2, 100, 16384, 32766, 32767]:
            want = np.float32(np.float32(i) * scale) + np.float32(lo32 - scale)
            assert t[i] == np.float32(want)
            assert t[i + 32768] == t[i]  # update marker is masked
        assert t[32768] == np.float32(unknown)
        assert abs(t[1] - lo) < 1e-6 * max(1, abs(hi)) and abs(t[32767] - hi) < 1e-4 * max(1, abs(hi))


def test_probability_value_round_trip(oracle):
    """cartographer/mapping/probability_values_test.cc:25-60."""
    L = oracle.lib()
    assert L.orc_probability_to_value(0.1) == 1
    assert L.orc_probability_to_value(0.9) == 32767
    for v in [1, 2, 77, 12345, 32766, 32767]:
        assert L.orc_probability_to_value(L.orc_value_to_probability(v)) == v
    assert abs(L.orc_value_to_probability(0) - 0.1) < 1e-7


FIXTURE_POINTS = np.array([[-3, 2, 0], [-4, 2, 0], [-5, 2, 0], [-6, 2, 0], [-6, 3, 1], [-6, 4, 2],
                           [-7, 3, 1]], np.float32)


def _interp_fixture(oracle):
    g = oracle.Grid.prob(0.1)
    g.set_probability(g.cell_index(FIXTURE_POINTS), 1.0)
    return g


def test_interpolated_tsdf_weight(oracle):
    """interpolated_tsdf.h:129-156: at a cell centre the blend is that voxel's weight,
    halfway between two centres it is their mean, far from every voxel it is 0.
    (The reference has no test of its own for GetWeight; these are the properties
    of the trilinear formula.)"""
    g = oracle.Grid.tsdf(1.0, 0.5, 100.0)
    g.set_cell([[1, 0, 1], [2, 0, 1]], [0.1, 0.2], [8.0, 24.0])
    w = g.lookup("weight", [[1, 0, 1], [2, 0, 1]]).astype(np.float64)
    got = g.interpolate_weight([[1.0, 0.0, 1.0], [2.0, 0.0, 1.0], [1.5, 0.0, 1.0], [1.0, 0.5, 1.0], [40., 40., 40.]])
    assert got[0] == w[0] and got[1] == w[1]
    assert abs(got[2] - 0.5 * (w[0] + w[1])) < 1e-12
    assert abs(got[3] - 0.5 * w[0]) < 1e-12
    assert got[4] == 0.0


def test_interpolated_grid_at_grid_points(oracle):
    """internal/3d/scan_matching/interpolated_grid_test.cc:52-61."""
    g = _interp_fixture(oracle)
    res = np.float64(np.float32(0.1))
    zs, ys, xs = [], [], []
    z = -1.0
    while z < 3.0:
        zs.append(z); z += res
    y = 1.0
    while y < 5.0:
        ys.append(y); y += res
    x = -8.0
    while x < -2.0:
        xs.append(x); x += res
    P = np.array([(x, y, z) for z in zs[::3] for y in ys[::3] for x in xs])
    got = g.interpolate(P)
    want = g.lookup("probability", g.cell_index(P.astype(np.float32)))
    assert np.abs(got - want).max() < 1e-6


def test_interpolated_grid_monotone_in_x(oracle):
    """interpolated_grid_test.cc:63-86."""
    g = _interp_fixture(oracle)
    res = np.float64(np.float32(0.1))
    step = res / 10.0
    checked = 0
    for (y, z) in [(2.0, 0.0), (3.0, 1.0), (4.0, 2.0)]:
        x = -8.0
        while x < -2.0:
            a = g.lookup("probability", g.cell_index(np.array([[x, y, z]], np.float32)))[0]
            b = g.lookup("probability", g.cell_index(np.array([[x + res, y, z]], np.float32)))[0]
            d = float(b) - float(a)
            if abs(d) >= 1e-6:
                s = step
                while s < res - 2 * step:
                    v = g.interpolate([[x + s + step, y, z], [x + s, y, z]])
                    assert d * (v[0] - v[1]) > 0.0
                    checked += 1
                    s += step
            x += res
    assert checked > 20


def test_precomputation_grid_against_naive(oracle):
    """internal/3d/scan_matching/precomputation_grid_3d_test.cc:30-77 (seed 23847)."""
    rng = np.random.RandomState(23847)
    g = oracle.Grid.prob(2.0)
    xyz = rng.randint(-50, 50, size=(1000, 3)).astype(np.int32)
    p = rng.uniform(0.1, 0.9, size=1000).astype(np.float32)
    for i in range(1000):  # later draws overwrite earlier ones, as SetProbability does
        g.set_probability(xyz[i:i + 1], p[i:i + 1])
    L = oracle.lib()
    for depth in range(4):
        width = 1 << depth
        q = rng.randint(-50, 50, size=(100, 3)).astype(np.int32)
        got = oracle.precompute_fullres_lookup(g, depth, q)
        d = np.arange(width)
        off = np.stack(np.meshgrid(d, d, d, indexing="ij"), -1).reshape(-1, 3)
        for i in range(100):
            probs = g.lookup("probability", q[i] + off)
            # the reference test starts its running max at 0, and unknown cells read 0.1
            naive = max(0.0, float(probs.max()))
            assert abs(naive - L.orc_precomp_to_probability(float(got[i]))) < 1e-2


def _rot_z(theta):
    return np.array([np.cos(theta / 2), 0, 0, np.sin(theta / 2)])


def _quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx])


def _rotate(q, v):
    w, u = q[0], np.asarray(q[1:])
    uv = 2 * np.cross(u, v)
    return v + w * uv + np.cross(u, uv)


def _rigid_matrix(t, q):
    """Eigen::Translation * Quaternion as the 4x4 matrix rigid_transform_test_helpers.h:38-40 builds."""
    w, x, y, z = np.asarray(q, float) / np.linalg.norm(q)
    m = np.eye(4)
    m[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                 [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                 [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]]
    m[:3, 3] = t
    return m


def is_nearly(pose, expected_t, expected_q, epsilon):
    """transform::IsNearly (rigid_transform_test_helpers.h:42-46) = Eigen's isApprox on the two
    4x4 matrices: |a - b|_F <= epsilon * min(|a|_F, |b|_F)."""
    a, b = _rigid_matrix(pose[:3], pose[3:]), _rigid_matrix(expected_t, expected_q)
    return np.linalg.norm(a - b) <= epsilon * min(np.linalg.norm(a), np.linalg.norm(b))


def ceres_fixture(grid_factory, initial):
    """The fixture of ceres_scan_matcher_3d_test.cc:36-117: cloud, grid, start pose, expected pose."""
    expected_t = np.array([-1., 0., 0.])
    expected_q = np.array([1., 0., 0., 0.])
    cloud = FIXTURE_POINTS.copy()
    grid_points = (cloud + expected_t).astype(np.float32)
    t0, mode = initial
    q0 = np.array([1., 0., 0., 0.])
    if mode == "full":
        # :100-115 — the cloud is turned by 0.05 rad about z, the start is turned about x
        add = _rot_z(0.05)
        cloud = np.array([_rotate(add.astype(np.float32), p) for p in cloud], np.float32)
        expected_q = _quat_mul(expected_q, add * np.array([1, -1, -1, -1]))
        q0 = np.array([np.cos(0.025), np.sin(0.025), 0., 0.])
    return cloud, grid_points, np.concatenate([t0, q0]), expected_t, expected_q


CERES_FIXTURE_STARTS = [((-1., 0., 0.), None), ((-0.8, 0., 0.), None), ((-1., 0., -0.2), None),
                        ((-0.9, -0.2, 0.2), None), ((-0.95, -0.05, 0.05), "full")]


@pytest.mark.parametrize("initial", CERES_FIXTURE_STARTS)
def test_ceres_scan_matcher_fixture(oracle, initial):
    """internal/3d/scan_matching/ceres_scan_matcher_3d_test.cc:36-117 at the reference's own
    settings: 7-point cloud, resolution-1.0 occupancy grid, occupied_space_weight 1,
    translation_weight 0.01, rotation_weight 0.1, use_nonmonotonic_steps = true,
    max_num_iterations = 10 (:50-61); EXPECT_NEAR(0, final_cost, 1e-2) and
    IsNearly(expected_pose, 3e-2) (:73-74) for PerfectEstimate, AlongX, AlongZ, AlongXYZ and
    FullPoseCorrection.

    Ceres is absent here; the solve is Ceres 1.13's trust-region loop restated in
    include/hg_trust_region.h. On this cost surface (the residuals do not vanish at the optimum
    while their derivatives do) the outcome after 10 iterations hangs on the trust-region rules:
    FullPoseCorrection only passes with the non-monotonic rule and Ceres' order of the
    convergence tests."""
    cloud, grid_points, start, expected_t, expected_q = ceres_fixture(None, initial)
    g = oracle.Grid.prob(1.0)
    g.set_probability(g.cell_index(grid_points), 1.0)
    pose, summary, _ = oracle.scan_match([cloud], [g], [1.0], 0.01, 0.1, start, max_num_iterations=10,
                                         use_nonmonotonic_steps=True)
    assert abs(summary[1]) < 1e-2
    assert is_nearly(pose, expected_t, expected_q, 3e-2)


def test_trust_region_loop_follows_ceres_rules(oracle):
    """Properties of Ceres 1.13's loop that the AlongX trace must show
    (trust_region_minimizer.cc / levenberg_marquardt_strategy.cc): every linear solve counts as
    an iteration; a rejected step divides the radius by 2, 4, 8, ... and an accepted one resets
    that factor and scales the radius by 1 / max(1/3, 1 - (2 rho - 1)^3); with non-monotonic
    steps a step whose cost rises can be accepted, and the reported pose is the lowest-cost
    accepted iterate."""
    cloud, grid_points, start, _, _ = ceres_fixture(None, ((-0.8, 0., 0.), None))
    g = oracle.Grid.prob(1.0)
    g.set_probability(g.cell_index(grid_points), 1.0)
    rows = []
    _, summary, term = oracle.scan_match([cloud], [g], [1.0], 0.01, 0.1, start, max_num_iterations=10,
                                         use_nonmonotonic_steps=True, trace=rows)
    rows = np.array(rows)  # iteration, cost, candidate cost, model change, rho, radius, successful
    assert term == 0 and len(rows) == 11 and list(rows[:, 0]) == list(range(11))
    assert rows[0, 5] == 1e4
    factor = 2.0
    rose = False
    for prev, row in zip(rows[:-1], rows[1:]):
        if row[6]:
            assert row[4] > 1e-3
            assert row[5] == min(1e16, prev[5] / max(1. / 3., 1. - (2. * row[4] - 1.) ** 3))
            factor = 2.0
            rose = rose or row[1] > prev[1]
        else:
            assert row[4] <= 1e-3 and row[1] == prev[1]
            assert row[5] == prev[5] / factor
            factor *= 2.0
    assert rose, "the non-monotonic rule accepted no uphill step on this fixture"
    assert summary[1] == rows[rows[:, 6] == 1][:, 1].min()


FCSM_CLOUD = np.array([[4, 0, 0], [4.5, 0, 0], [5, 0, 0], [5.5, 0, 0], [0, 4, 0], [0, 4.5, 0], [0, 5, 0],
                       [0, 5.5, 0], [0, 0, 4], [0, 0, 4.5], [0, 0, 5], [0, 0, 5.5]], np.float32)
FCSM_TEST_OPTIONS = dict(branch_and_bound_depth=6, full_resolution_depth=6,
                         use_rotational_scan_matcher=True, min_rotational_score=0.1,
                         min_low_resolution_score=0.15, linear_xy_search_window=0.8,
                         linear_z_search_window=0.8, angular_search_window=0.3)


def _fcsm_fixture(oracle, rng):
    t = (0.7 * rng.uniform(-1, 1, 3)).astype(np.float32)
    theta = np.float32(0.2 * rng.uniform(-1, 1))
    q = _rot_z(theta).astype(np.float32)
    world = np.array([_rotate(q, p) + t for p in FCSM_CLOUD], np.float32)
    g = oracle.Grid.prob(0.05)
    g.insert_occupancy(t, world, 0.7, 0.4, 5)
    m = oracle.Fcsm(g, g, np.zeros(10, np.float32), **FCSM_TEST_OPTIONS)
    return m, t, q


def _check_fcsm_result(r, t, q):
    assert r["found"] == 1.0
    assert r["score"] > 0.1 and r["rotational_score"] > 0.09 and r["low_resolution_score"] > 0.14
    assert np.abs(np.array([r["tx"], r["ty"], r["tz"]]) - t).max() < 0.05
    got = np.array([r["qw"], r["qx"], r["qy"], r["qz"]])
    ang = 2 * np.arccos(min(1.0, abs(float(np.dot(got / np.linalg.norm(got), q)))))
    assert ang < 0.05


def test_fcsm_match_fixture(oracle):
    """internal/3d/scan_matching/fast_correlative_scan_matcher_3d_test.cc:141-169
    (12-point axis cloud, 5 cm grid from the occupancy writer, seed 42; 8 of the 20 poses)."""
    rng = np.random.RandomState(42)
    ident = [0, 0, 0, 1, 0, 0, 0]
    for _ in range(8):
        m, t, q = _fcsm_fixture(oracle, rng)
        r = m.match(ident, ident, FCSM_CLOUD, FCSM_CLOUD, np.zeros(10), min_score=0.1)
        _check_fcsm_result(r, t, q)
        far = m.match(ident, ident, FCSM_CLOUD, [[42., 42., 42.]], np.zeros(10), min_score=0.1)
        assert far["found"] == 0.0


def test_fcsm_match_full_submap_fixture(oracle):
    """fast_correlative_scan_matcher_3d_test.cc:171-196."""
    rng = np.random.RandomState(42)
    ident = [0, 0, 0, 1, 0, 0, 0]
    m, t, q = _fcsm_fixture(oracle, rng)
    r = m.match(ident, ident, FCSM_CLOUD, FCSM_CLOUD, np.zeros(10), min_score=0.1, full=True)
    _check_fcsm_result(r, t, q)
    far = m.match(ident, ident, FCSM_CLOUD, [[42., 42., 42.]], np.zeros(10), min_score=0.1, full=True)
    assert far["found"] == 0.0


def test_rotate_histogram(oracle):
    """internal/3d/scan_matching/rotational_scan_matcher_test.cc (RotateHistogram cases):
    a quarter-turn of a 4-bucket histogram shifts by two buckets; an eighth interpolates."""
    h = np.array([1., 0., 0., 0.], np.float32)
    r = oracle.rotate_histogram(h, np.float32(np.pi / 2))
    assert np.allclose(r, [0., 0., 1., 0.], atol=1e-5)
    r = oracle.rotate_histogram(h, np.float32(np.pi / 8))
    assert np.allclose(r, [0.5, 0.5, 0., 0.], atol=1e-5)


def test_voxel_filter_first_point_in_each_voxel(oracle):
    """cartographer/sensor/internal/voxel_filter_test.cc:30-37."""
    pts = np.array([[0, 0, 0], [0.1, -0.1, 0.1], [0.3, -0.1, 0], [0, 0, 0.1]], np.float32)
    assert oracle.voxel_filter(0.3, pts).tolist() == [0, 2]


def test_voxel_filter_large_coordinates(oracle):
    """voxel_filter_test.cc:39-46."""
    pts = np.array([[100000, 0, 0], [100000.001, -0.0001, 0.0001], [100000.003, -0.0001, 0], [-200000, 0, 0]],
                   np.float32)
    assert oracle.voxel_filter(0.01, pts).tolist() == [0, 3]


def test_voxel_filter_ignores_time(oracle):
    """voxel_filter_test.cc:48-55 (TimedPointCloud: stride 4)."""
    pts = np.array([[-100, 0.3, 0.4, float(i)] for i in range(100)], np.float32)
    assert oracle.voxel_filter(0.3, pts, stride=4).tolist() == [0]


def test_adaptive_voxel_filter_reaches_min_num_points(oracle):
    """adaptive_voxel_filter.h:45-84: the result has at least min_num_points points whenever the
    input has, is a subsequence of the input, and respects max_range."""
    rng = np.random.RandomState(0)
    pts = rng.uniform(-20, 20, (5000, 3)).astype(np.float32)
    for max_length, min_points in [(4.0, 200), (0.5, 150), (8.0, 3000)]:
        kept = oracle.adaptive_voxel_filter(max_length, min_points, 25.0, pts)
        assert len(kept) >= min_points
        assert (np.diff(kept) > 0).all()
        assert (np.linalg.norm(pts[kept], axis=1) <= 25.0 + 1e-4).all()
    few = oracle.adaptive_voxel_filter(1.0, 100, 25.0, pts[:50])
    assert few.tolist() == [i for i in range(50) if np.linalg.norm(pts[i]) <= 25.0]


def _insert_reference_point_cloud(g):
    """range_data_inserter_3d_test.cc:43-51."""
    g.insert_occupancy([0.0, 0.0, -4.0], [[-3, -1, 4], [-2, 0, 4], [-1, 1, 4], [0, 2, 4]], 0.7, 0.4, 1000)


def test_range_data_inserter_insert_point_cloud(oracle):
    """cartographer/mapping/3d/range_data_inserter_3d_test.cc:72-98 (resolution 1, hit 0.7, miss 0.4,
    num_free_space_voxels 1000)."""
    g = oracle.Grid.prob(1.0)
    _insert_reference_point_cloud(g)

    def p(x, y, z):
        return g.lookup("probability", g.cell_index([[x, y, z]]))[0]

    def known(x, y, z):
        return g.lookup("raw_a", g.cell_index([[x, y, z]]))[0] != 0

    for z in (-4.0, -3.0, -2.0):
        assert abs(p(0.0, 0.0, z) - 0.4) < 1e-4
    for x in range(-4, 5):
        for y in range(-4, 5):
            if x < -3 or x > 0 or y != x + 2:
                assert not known(x, y, 4.0)
            else:
                assert abs(p(x, y, 4.0) - 0.7) < 1e-4


def test_range_data_inserter_probability_progression(oracle):
    """range_data_inserter_3d_test.cc:100-122."""
    g = oracle.Grid.prob(1.0)
    _insert_reference_point_cloud(g)

    def p(x, y, z):
        return g.lookup("probability", g.cell_index([[x, y, z]]))[0]

    assert abs(p(-2, 0, 4) - 0.7) < 1e-4
    assert abs(p(-2, 0, 3) - 0.4) < 1e-4
    assert abs(p(0, 0, -3) - 0.4) < 1e-4
    for _ in range(1000):
        _insert_reference_point_cloud(g)
    assert abs(p(-2, 0, 4) - 0.9) < 1e-3
    assert abs(p(-2, 0, 3) - 0.1) < 1e-3
    assert abs(p(0, 0, -3) - 0.1) < 1e-3


def test_tsdf_insert_single_ray_properties(oracle):
    """No reference test exists for TSDFRangeDataInserter3D; these are the properties its code
    states (tsdf_range_data_inserter_3d.cc:294-343, 725-737): only cells within the truncation
    distance of the hit along the ray are written, the TSD is range - |cell - origin| clamped,
    the weight in front of the surface is 1 and repeated insertion keeps the TSD and adds weight."""
    res, rtd = 0.1, 2.5
    g = oracle.Grid.tsdf(res, rtd, 1000.0)
    origin = np.array([0.0, 0.0, 0.0], np.float32)
    hit = np.array([[2.0, 0.0, 0.0]], np.float32)
    g.insert_tsdf(origin, hit, rtd, 1000.0, 0, 0.1, 4.0)
    xyz, tsd_raw, w_raw = g.export()
    assert len(xyz) >= 5 and (xyz[:, 1] == 0).all() and (xyz[:, 2] == 0).all()
    centers = xyz[:, 0] * np.float32(res)
    assert centers.min() >= 2.0 - rtd * res - res and centers.max() <= 2.0 + rtd * res + res
    tsd = g.lookup("tsd", xyz)
    w = g.lookup("weight", xyz)
    want = np.clip(2.0 - centers, -rtd * res, rtd * res)
    assert np.allclose(tsd, want, atol=2 * rtd * res / 32766 + 1e-6)
    assert np.allclose(w[tsd > -0.1 * rtd * res], 1.0, atol=1e-1)
    assert (w[tsd < -0.5 * rtd * res] < 0.9).all()
    g.insert_tsdf(origin, hit, rtd, 1000.0, 0, 0.1, 4.0)
    assert np.allclose(g.lookup("tsd", xyz), tsd, atol=1e-4)
    assert np.allclose(g.lookup("weight", xyz), 2 * w, atol=1e-1)
    # The gates of Insert (:700-722): range limits and the insertion ratio.
    g2 = oracle.Grid.tsdf(res, rtd, 1000.0)
    g2.insert_tsdf(origin, hit, rtd, 1000.0, 0, 0.1, 4.0, min_range=3.0)
    assert len(g2.export()[0]) == 0
    g2.insert_tsdf(origin, hit, rtd, 1000.0, 0, 0.1, 4.0, max_range=1.0)
    assert len(g2.export()[0]) == 0
    many = np.repeat(hit, 10, axis=0)
    g2.insert_tsdf(origin, many, rtd, 1000.0, 0, 0.1, 4.0, insertion_ratio=0.3)
    # the counter admits returns 0, 4, 7 (inserted <= 0.3 * seen)
    assert abs(g2.lookup("weight", [[20, 0, 0]])[0] - 3.0) < 1e-1


def _quat_axis_angle(angle, axis):
    axis = np.asarray(axis, float)
    axis = axis / np.linalg.norm(axis)
    return np.concatenate([[np.cos(angle / 2)], np.sin(angle / 2) * axis])


def _quat_mul(a, b):
    w1, x1, y1, z1 = a
    w2, x2, y2, z2 = b
    return np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])


def test_interpolate_transform(oracle):
    """cartographer/transform/timestamped_transform_test.cc:37-80: identity -> (10,10,10) with a 2 rad
    rotation about z; the interpolant at 'factor' is Translation(factor t) * Rotation(factor 2 about z)."""
    t0 = np.array([0, 0, 0, 1, 0, 0, 0], float)
    t1 = np.concatenate([[10., 10., 10.], _quat_axis_angle(2.0, [0, 0, 1])])
    assert np.abs(oracle.interpolate_transform(t0, t1, 0.0) - t0).max() < 1e-6
    assert np.abs(oracle.interpolate_transform(t0, t1, 1.0) - t1).max() < 1e-6
    for k in range(21):
        f = k / 20.0
        want = np.concatenate([f * t1[:3], _quat_axis_angle(2.0 * f, [0, 0, 1])])
        assert np.abs(oracle.interpolate_transform(t0, t1, f) - want).max() < 1e-6


def test_rotation_delta_cost_functor(oracle):
    """mapping/internal/3d/scan_matching/rotation_delta_cost_functor_3d_test.cc:50-86 (precision 1e-8)."""
    def cost(rotation, scale, target):
        r = oracle.rotation_delta(scale, target, rotation)
        return float(r @ r)

    identity = np.array([1.0, 0, 0, 0])
    assert abs(cost(identity, 1.0, identity)) < 1e-8
    rot = _quat_axis_angle(0.9, [0.2, 0.1, 0.3])
    assert abs(cost(rot, 1.0, rot)) < 1e-8
    s, angle = 1.2, 0.8
    rotation = _quat_axis_angle(angle, [0.2, 0.1, 0.8])
    target = _quat_axis_angle(0.2, [-0.5, 0.3, 0.4])
    expected = (s * np.sin(angle / 2.0)) ** 2
    assert abs(cost(rotation, s, identity) - expected) < 1e-8
    assert abs(cost(_quat_mul(target, rotation), s, target) - expected) < 1e-8
    assert abs(cost(_quat_mul(rotation, target), s, target) - expected) < 1e-8
    # the Jacobian is the constant matrix the GPU matcher and the host solver use
    r, J = oracle.rotation_delta(s, target, rotation, jacobian=True)
    eps = 1e-6
    for c in range(4):
        d = np.zeros(4)
        d[c] = eps
        num = (oracle.rotation_delta(s, target, rotation + d) - oracle.rotation_delta(s, target, rotation - d)) / (2 * eps)
        assert np.abs(J[:, c] - num).max() < 1e-8


def test_translation_delta_cost_functor(oracle):
    """translation_delta_cost_functor_3d.h:31-45: s (t - target), Jacobian s I."""
    r, J = oracle.translation_delta(2.5, [1.0, -2.0, 0.5], [1.5, -2.0, 0.0], jacobian=True)
    assert np.allclose(r, [1.25, 0.0, -1.25]) and np.allclose(J, 2.5 * np.eye(3))
