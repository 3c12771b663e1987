"""GPU parity tests for range-data insertion: hg_grid_insert_tsdf and
hg_grid_insert_occupancy against the CPU oracle's restatement of TSDFRangeDataInserter3D /
OccupancyGridRangeDataInserter3D on the same scans. Bar: the stored uint16 voxels are identical
(bit-exact), as are grid_size and everything derived from the grid afterwards."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hg():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import hectorgrapher_b200 as hg
    hg.lib()
    return hg


def world_scan(k, count, rings=16, azimuths=360, seed=0):
    """A scan of the synthetic world expressed in the map frame, and its origin."""
    from hectorgrapher_b200 import synth
    sph = synth.spheres()
    origin, yaw = synth.corkscrew_pose(k, count)
    rng = np.random.RandomState(seed + k)
    cloud = synth.scan(origin, yaw, rings, azimuths, sph, rng)
    q = synth.quat_from_yaw(yaw)
    world = (synth.quat_rotate(q, cloud.astype(np.float64)) + origin).astype(np.float32)
    return origin.astype(np.float32), world


def sorted_voxels(xyz, a, b=None, *_):
    order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
    return xyz[order], a[order], (None if b is None else b[order])


def assert_same_grid(g, o):
    gx, ga, gb = sorted_voxels(*g.export())
    ox, oa, ob = sorted_voxels(*o.export())
    assert gx.shape == ox.shape
    assert (gx == ox).all()
    assert (ga == oa).all()
    if ob is not None:
        assert (gb == ob).all()
    assert g.grid_size == o.grid_size
    return len(gx)


TSDF_CASES = [
    dict(),                                                            # oracle defaults: epsilon 0.1, sigma 4
    dict(epsilon=1.0, min_range=0.4, max_range=15.0),                  # trajectory_builder_3d.lua:78-93
    dict(num_free_space_voxels=3),                                     # rays start at the origin
    dict(insertion_ratio=0.37, min_range=1.0, max_range=60.0),         # low-resolution inserter, :102-117
    dict(relative_truncation_distance=3.3, maximum_weight=5.0, sigma=11.0),   # weight cap reached, non-float options
]


@pytest.mark.parametrize("case", range(len(TSDF_CASES)))
def test_tsdf_insertion_matches_oracle_bit_for_bit(oracle, hg, case):
    opts = TSDF_CASES[case]
    res = 0.10 if case != 3 else 0.45
    g = hg.Grid.tsdf(res, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(res, 2.5, 1000.0)
    total = 0
    for k in range(4):
        origin, pts = world_scan(k, 4)
        total += g.insert_tsdf(origin, pts, **opts)
        o.insert_tsdf(origin, pts, **opts)
        n = assert_same_grid(g, o)
    assert n > 1000 and total > n
    # decoded lookups agree as well (the read copies were regenerated)
    xyz = g.export()[0][::7]
    for what in ("tsd", "weight"):
        assert (g.lookup(what, xyz) == o.lookup(what, xyz)).all()


def test_tsdf_insertion_repeated_scan_accumulates(oracle, hg):
    """The same scan many times: long per-voxel update chains, weights run into maximum_weight."""
    g = hg.Grid.tsdf(0.10, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    origin, pts = world_scan(1, 4, rings=8, azimuths=180)
    for _ in range(12):
        g.insert_tsdf(origin, pts, maximum_weight=6.0)
        o.insert_tsdf(origin, pts, maximum_weight=6.0)
    assert_same_grid(g, o)
    w = g.lookup("weight", g.export()[0])
    assert w.max() <= 6.0 + 1000.0 / 32766 and w.max() > 5.9  # one quantisation step of the weight


def test_tsdf_insertion_into_uploaded_grid(oracle, hg):
    """A grid that came through hg_grid_upload_tsdf is written to afterwards (the active submap)."""
    from hectorgrapher_b200 import synth
    sph = synth.spheres()
    poses = [synth.corkscrew_pose(k, 4) for k in range(2)]
    xyz, tv, wv = synth.build_tsdf(0.10, poses, 16, 180, sph)
    o = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    o.set_raw(xyz, tv, wv)
    g = hg.Grid.tsdf(0.10, 2.5, 1000.0, xyz, tv, wv, grid_size=o.grid_size)
    assert_same_grid(g, o)
    for k in (2, 3):
        origin, pts = world_scan(k, 4)
        g.insert_tsdf(origin, pts, epsilon=1.0)
        o.insert_tsdf(origin, pts, epsilon=1.0)
    assert_same_grid(g, o)


def test_tsdf_insertion_edge_cases(oracle, hg):
    g = hg.Grid.tsdf(0.10, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    origin = np.array([0.3, -0.2, 0.1], np.float32)
    # nothing to do
    assert g.insert_tsdf(origin, np.zeros((0, 3), np.float32)) == 0
    # NaN returns, returns closer than the truncation distance, a return exactly on the origin
    pts = np.array([[np.nan, 1, 1], [1, np.nan, 1], [0.35, -0.2, 0.1], [0.3, -0.2, 0.1], [2.0, 1.0, 0.5],
                    [-3.0, 0.25, -0.75]], np.float32)
    assert g.insert_tsdf(origin, pts) > 0
    o.insert_tsdf(origin, pts)
    assert_same_grid(g, o)
    # stride 4 (TimedPointCloud layout) gives the same grid
    g4 = hg.Grid.tsdf(0.10, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    p4 = np.concatenate([pts, np.full((len(pts), 1), 0.5, np.float32)], axis=1)
    g4.insert_tsdf(origin, p4, stride=4)
    assert_same_grid(g4, o)
    # far away: the grid grows like DynamicGrid::Grow
    far = np.array([[100.0, 0.0, 0.0], [-150.0, 20.0, 5.0]], np.float32)
    g.insert_tsdf(origin, far)
    o.insert_tsdf(origin, far)
    assert_same_grid(g, o)
    assert g.grid_size >= 2048
    # beyond the largest grid the reference CHECK-fails; the library reports an error
    with pytest.raises(hg.HgError):
        g.insert_tsdf(origin, np.array([[900.0, 0.0, 0.0]], np.float32))
    # wrong grid type
    p = hg.Grid.prob(0.10, np.zeros((0, 3), np.int32), [])
    with pytest.raises(hg.HgError):
        p.insert_tsdf(origin, far)


def test_inserted_grid_feeds_cost_and_pyramid(oracle, hg):
    """Everything downstream of the grid sees the inserted data: interpolated cost function
    (apron + decoded copies) and the precomputation pyramid."""
    g = hg.Grid.tsdf(0.10, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    for k in range(3):
        origin, pts = world_scan(k, 3)
        g.insert_tsdf(origin, pts, epsilon=1.0)
        o.insert_tsdf(origin, pts, epsilon=1.0)
    from hectorgrapher_b200 import synth
    origin, yaw = synth.corkscrew_pose(1, 3)
    cloud = synth.scan(origin, yaw, 16, 256, synth.spheres(), np.random.RandomState(5))
    pose = np.concatenate([origin + [0.05, -0.03, 0.02], synth.quat_from_yaw(yaw + 0.01)])
    s = 1.0 / np.sqrt(len(cloud))
    gr, gJ = hg.Cost("tsdf", s, cloud, [g]).evaluate(pose)
    orr, oJ = oracle.Cost("tsdf", s, cloud, [o]).evaluate(pose)
    assert np.allclose(gr, orr, rtol=1e-5, atol=1e-9)
    for a, b in zip(gJ, oJ):
        assert np.allclose(a, b, rtol=1e-5, atol=1e-9)
    gs = hg.Stack(g, 5, 3)
    os_ = oracle.Stack(o, 5, 3)
    for d in range(5):
        gx, gv = gs.export(d)
        ox, ov = os_.export(d)
        a = sorted_voxels(gx, gv)
        b = sorted_voxels(ox, ov)
        assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


def _insert_reference_point_cloud(g):
    g.insert_occupancy([0.0, 0.0, -4.0], [[-3, -1, 4], [-2, 0, 4], [-1, 1, 4], [0, 2, 4]], 0.7, 0.4, 1000)


def test_occupancy_reference_fixture(oracle, hg):
    """range_data_inserter_3d_test.cc:72-122 run on the device, and against the oracle."""
    g = hg.Grid.prob(1.0, np.zeros((0, 3), np.int32), [])
    o = oracle.Grid.prob(1.0)
    _insert_reference_point_cloud(g)
    _insert_reference_point_cloud(o)
    assert_same_grid(g, o)

    def p(x, y, z):
        return g.lookup("probability", hg.get_cell_index(1.0, [[x, y, z]]))[0]

    for z in (-4.0, -3.0, -2.0):
        assert abs(p(0.0, 0.0, z) - 0.4) < 1e-4
    known = {tuple(c) for c in g.export()[0]}
    for x in range(-4, 5):
        for y in range(-4, 5):
            if x < -3 or x > 0 or y != x + 2:
                assert (x, y, 4) not in known
            else:
                assert abs(p(x, y, 4.0) - 0.7) < 1e-4
    for _ in range(1000):
        _insert_reference_point_cloud(g)
    assert abs(p(-2, 0, 4) - 0.9) < 1e-3
    assert abs(p(-2, 0, 3) - 0.1) < 1e-3
    assert abs(p(0, 0, -3) - 0.1) < 1e-3
    for _ in range(1000):
        _insert_reference_point_cloud(o)
    assert_same_grid(g, o)


@pytest.mark.parametrize("free", [0, 2, 1000])
def test_occupancy_insertion_matches_oracle(oracle, hg, free):
    res = 0.10 if free < 1000 else 0.45
    g = hg.Grid.prob(res, np.zeros((0, 3), np.int32), [])
    o = oracle.Grid.prob(res)
    for k in range(4):
        origin, pts = world_scan(k, 4)
        g.insert_occupancy(origin, pts, 0.55, 0.49, free)
        o.insert_occupancy(origin, pts, 0.55, 0.49, free)
        n = assert_same_grid(g, o)
    assert n > 1000
    xyz = g.export()[0][::5]
    assert (g.lookup("probability", xyz) == o.lookup("probability", xyz)).all()
    # bad options are rejected like CreateProbabilityGridRangeDataInserterOptions3D's CHECKs
    with pytest.raises(hg.HgError):
        g.insert_occupancy(origin, pts, 0.45, 0.49, free)


def test_insertion_with_hash_indexed_bricks(oracle, hg):
    """Returns 700 m apart: the brick bounding box is far too large for a dense index, so the grid
    switches to the open-addressing brick hash (hg_internal.cuh) while it is being written."""
    res = 0.10
    g = hg.Grid.tsdf(res, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(res, 2.5, 1000.0)
    origin = np.array([0.0, 0.0, 0.0], np.float32)
    rng = np.random.RandomState(8)
    near = rng.uniform(-3, 3, (200, 3)).astype(np.float32)
    for pts in (near, np.array([[700.0, 10.0, -3.0], [-690.0, -700.0, 5.0]], np.float32), near + np.float32(0.05)):
        g.insert_tsdf(origin, pts)
        o.insert_tsdf(origin, pts)
        assert_same_grid(g, o)
    assert g.grid_size == 16384
    xyz = g.export()[0]
    assert (g.lookup("tsd", xyz) == o.lookup("tsd", xyz)).all()
    # the cost function reads the same grid through the hash
    cloud = near[:64]
    pose = np.array([0.01, -0.02, 0.005, 1, 0, 0, 0], float)
    gr, gJ = hg.Cost("tsdf", 0.1, cloud, [g]).evaluate(pose)
    orr, oJ = oracle.Cost("tsdf", 0.1, cloud, [o]).evaluate(pose)
    assert np.allclose(gr, orr, rtol=1e-5, atol=1e-9)
    assert all(np.allclose(a, b, rtol=1e-5, atol=1e-9) for a, b in zip(gJ, oJ))
    # occupancy grid likewise
    gp = hg.Grid.prob(res, np.zeros((0, 3), np.int32), [])
    op = oracle.Grid.prob(res)
    for pts in (near, np.array([[700.0, 10.0, -3.0], [-690.0, -700.0, 5.0]], np.float32)):
        gp.insert_occupancy(origin, pts, 0.55, 0.49, 2)
        op.insert_occupancy(origin, pts, 0.55, 0.49, 2)
        assert_same_grid(gp, op)


def test_cost_function_created_before_an_insertion_sees_the_new_grid(oracle, hg):
    """The active submap is matched against and written to alternately
    (optimizing_local_trajectory_builder.cc); a cost function that outlives an insertion reads the
    grid's current device buffers, also after they were reallocated for new bricks."""
    from hectorgrapher_b200 import synth
    g = hg.Grid.tsdf(0.10, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    origin, pts = world_scan(0, 3, rings=8, azimuths=90)
    g.insert_tsdf(origin, pts, epsilon=1.0)
    o.insert_tsdf(origin, pts, epsilon=1.0)
    sensor, yaw = synth.corkscrew_pose(1, 3)
    cloud = synth.scan(sensor, yaw, 8, 64, synth.spheres(), np.random.RandomState(2))
    pose = np.concatenate([sensor, synth.quat_from_yaw(yaw)])
    cost = hg.Cost("tsdf", 0.1, cloud, [g])
    before = cost.evaluate(pose)[0]
    for k in (1, 2):  # many new bricks: index, brick and decoded buffers are reallocated
        origin, pts = world_scan(k, 3)
        g.insert_tsdf(origin, pts, epsilon=1.0)
        o.insert_tsdf(origin, pts, epsilon=1.0)
    r, J = cost.evaluate(pose)
    r0, J0 = oracle.Cost("tsdf", 0.1, cloud, [o]).evaluate(pose)
    assert np.allclose(r, r0, rtol=1e-5, atol=1e-9) and not np.allclose(r, before)
    assert all(np.allclose(a, b, rtol=1e-5, atol=1e-9) for a, b in zip(J, J0))
