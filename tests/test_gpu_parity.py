"""GPU parity tests: the CUDA path, called through the C-ABI, against the CPU oracle on the
same seeded inputs. Bars (BASELINE.json north_star): bit-exact voxel indices, precomputation
grids, integer candidate sums and selected candidate; residuals / Jacobians within 1e-5
relative; solved pose within 1e-4 m / 1e-4 rad."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5   # north_star: residuals and Jacobians within 1e-5 relative of the double-precision Jets
ATOL = 1e-9   # absolute floor for entries that cancel to ~0 (values are O(1e-2..1))


@pytest.fixture(scope="module")
def hg():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import hectorgrapher_b200 as hg
    hg.lib()
    return hg


@pytest.fixture(scope="module")
def world(oracle, hg):
    """Synthetic box+spheres world: TSDF high/low grids and an occupancy grid, in both libraries."""
    from hectorgrapher_b200 import synth
    sph = synth.spheres()
    poses = [synth.corkscrew_pose(k, 4) for k in range(4)]
    w = {"sph": sph, "poses": poses, "synth": synth}
    xyz, tv, wv = synth.build_tsdf(0.10, poses, 32, 360, sph)
    w["tsdf_hi"] = (xyz, tv, wv)
    g = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    g.set_raw(xyz, tv, wv)
    w["o_tsdf_hi"] = g
    w["g_tsdf_hi"] = hg.Grid.tsdf(0.10, 2.5, 1000.0, xyz, tv, wv, grid_size=g.grid_size)
    xyz, tv, wv = synth.build_tsdf(0.45, poses, 32, 360, sph)
    g = oracle.Grid.tsdf(0.45, 2.5, 1000.0)
    g.set_raw(xyz, tv, wv)
    w["o_tsdf_lo"] = g
    w["g_tsdf_lo"] = hg.Grid.tsdf(0.45, 2.5, 1000.0, xyz, tv, wv, grid_size=g.grid_size)
    xyz, pv = synth.build_occupancy(0.10, poses, 32, 360, sph)
    g = oracle.Grid.prob(0.10)
    g.set_raw(xyz, pv)
    w["o_prob"] = g
    w["g_prob"] = hg.Grid.prob(0.10, xyz, pv, grid_size=g.grid_size)
    xyz, pv = synth.build_occupancy(0.45, poses, 32, 360, sph)
    g = oracle.Grid.prob(0.45)
    g.set_raw(xyz, pv)
    w["o_prob_lo"] = g
    w["g_prob_lo"] = hg.Grid.prob(0.45, xyz, pv, grid_size=g.grid_size)
    rng = np.random.RandomState(3)
    origin, yaw = poses[1]
    cloud = synth.scan(origin, yaw, 16, 256, sph, rng)
    w["cloud"] = cloud
    w["truth"] = np.concatenate([origin, synth.quat_from_yaw(yaw)])
    return w


def perturbed(truth, rng, dt=0.1, da=0.05):
    from hectorgrapher_b200 import synth
    t = truth[:3] + rng.uniform(-dt, dt, 3)
    q = synth.quat_mul(truth[3:], synth.small_rotation(rng, da))
    return np.concatenate([t, q])


def assert_close(a, b, what=""):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape
    err = np.abs(a - b)
    tol = ATOL + RTOL * np.maximum(np.abs(a), np.abs(b))
    bad = err > tol
    assert not bad.any(), "%s: %d / %d entries differ, worst abs %.3e (values %.6e vs %.6e)" % (
        what, bad.sum(), bad.size, err.max(), a.flat[err.argmax()], b.flat[err.argmax()])


# ------------------------------------------------------------------ storage --

def test_cell_index_bit_exact(oracle, hg):
    rng = np.random.RandomState(0)
    for res in [0.05, 0.1, 0.45, 2.0]:
        pts = rng.uniform(-60, 60, size=(20000, 3)).astype(np.float32)
        # exact half-cell cases (lround: half away from zero) on both sides of zero
        k = rng.randint(-500, 500, size=(2000, 3))
        halves = ((k + 0.5) * np.float32(res)).astype(np.float32)
        pts = np.concatenate([pts, halves, -halves])
        assert (hg.get_cell_index(res, pts) == oracle.get_cell_index(res, pts)).all()
    want = np.array([[4, 6, 0], [4, 6, 1], [3, 7, 1], [3, 7, 2]])
    got = hg.get_cell_index(2.0, [[8.5, 11.5, 0.5], [7.5, 12.5, 1.5], [6.5, 14.5, 2.5], [5.5, 13.5, 3.5]])
    assert (got == want).all()  # hybrid_grid_test.cc:104-112


def test_grid_lookup_bit_exact(oracle, hg, world):
    rng = np.random.RandomState(1)
    for name, kinds in [("tsdf_hi", ["tsd", "weight"]), ("tsdf_lo", ["tsd", "weight"]), ("prob", ["probability"])]:
        og, gg = world["o_" + name], world["g_" + name]
        ex = og.export()[0]
        near = ex[rng.randint(0, len(ex), 30000)] + rng.randint(-2, 3, size=(30000, 3))
        far = rng.randint(-9000, 9000, size=(2000, 3))
        q = np.concatenate([near, far]).astype(np.int32)
        for what in kinds:
            a, b = gg.lookup(what, q), og.lookup(what, q)
            assert (a.view(np.uint32) == b.view(np.uint32)).all(), (name, what)


def test_interpolated_weight_bit_exact(oracle, hg, world):
    # InterpolatedTSDF::GetWeight (interpolated_tsdf.h:129-156; the multi-resolution
    # variant, interpolated_multi_resolution_tsdf.h:139-171, is this on the finest grid).
    rng = np.random.RandomState(11)
    for name in ["tsdf_hi", "tsdf_lo"]:
        og, gg = world["o_" + name], world["g_" + name]
        res = 0.10 if name == "tsdf_hi" else 0.45
        ex = og.export()[0]
        centres = ex[rng.randint(0, len(ex), 20000)].astype(np.float64) * np.float32(res)
        near = centres + rng.uniform(-1.5 * res, 1.5 * res, size=centres.shape)
        on_centre = centres[:2000]                     # exactly on a cell centre (n = 0 or 1)
        on_face = centres[2000:4000] + 0.5 * np.float32(res)   # cell boundary: lround ties
        far = rng.uniform(-900, 900, size=(1000, 3))   # no brick: all weights 0
        q = np.concatenate([near, on_centre, on_face, far])
        a, b = gg.interpolated_weight(q), og.interpolate_weight(q)
        assert (a.view(np.uint64) == b.view(np.uint64)).all(), name
        assert (a[:20000] > 0).mean() > 0.5 and (a[-1000:] == 0).all()
    with pytest.raises(Exception):
        world["g_prob"].interpolated_weight([[0.0, 0.0, 0.0]])
    assert len(world["g_tsdf_hi"].interpolated_weight(np.zeros((0, 3)))) == 0


def test_grid_edge_cases(oracle, hg):
    # empty grid: everything reads the default
    g = hg.Grid.prob(0.1, np.zeros((0, 3), np.int32), np.zeros(0, np.uint16))
    assert (g.lookup("probability", [[0, 0, 0], [5, -3, 9]]) == np.float32(0.1)).all()
    t = hg.Grid.tsdf(1.0, 0.5, 10.0, np.zeros((0, 3), np.int32), np.zeros(0, np.uint16), np.zeros(0, np.uint16))
    assert (t.lookup("tsd", [[1, 2, 3]]) == np.float32(-0.5)).all()   # hybrid_grid_tsdf_test.cc:45-52
    assert (t.lookup("weight", [[1, 2, 3]]) == 0).all()
    # single voxel at the extreme corner a reference grid can hold (+-8192)
    og = oracle.Grid.tsdf(0.05, 2.5, 1000.0)
    xyz = np.array([[-8192, -8192, -8192], [8191, 8191, 8191], [0, 0, 0]], np.int32)
    og.set_cell(xyz, [0.01, -0.02, 0.1], [1.0, 2.0, 3.0])
    ex, tv, wv = og.export()
    gg = hg.Grid.tsdf(0.05, 2.5, 1000.0, ex, tv, wv)
    assert gg.grid_size == og.grid_size == 16384
    for what in ["tsd", "weight"]:
        assert (gg.lookup(what, xyz).view(np.uint32) == og.lookup(what, xyz).view(np.uint32)).all()
    # update marker on probability values is masked
    g2 = hg.Grid.prob(0.1, [[1, 1, 1]], np.array([0x8000 | 20000], np.uint16))
    o2 = oracle.Grid.prob(0.1)
    o2.set_raw([[1, 1, 1]], np.array([0x8000 | 20000], np.uint16))
    assert g2.lookup("probability", [[1, 1, 1]])[0] == o2.lookup("probability", [[1, 1, 1]])[0]


# ------------------------------------------------------------ cost functions --

KIND_GRIDS = {
    "occupied": ["prob"], "interpolated_occupied": ["prob"],
    "tsdf": ["tsdf_hi"], "interpolated_tsdf": ["tsdf_hi"], "interpolated_tsdf_per_point": ["tsdf_hi"],
    "multires_tsdf": ["tsdf_hi", "tsdf_lo"], "interpolated_multires_tsdf": ["tsdf_hi", "tsdf_lo"],
    "interpolated_multires_tsdf_per_point": ["tsdf_hi", "tsdf_lo"],
}


@pytest.mark.parametrize("kind", list(KIND_GRIDS.keys()))
def test_cost_function_parity(oracle, hg, world, kind):
    rng = np.random.RandomState(11)
    cloud = world["cloud"]
    if kind.endswith("per_point"):
        cloud = cloud[777:778]
    scale = 0.7 / np.sqrt(len(cloud))
    og = [world["o_" + n] for n in KIND_GRIDS[kind]]
    gg = [world["g_" + n] for n in KIND_GRIDS[kind]]
    two = kind.startswith("interpolated")
    ratio = 0.37 if two else 0.0
    oc = oracle.Cost(kind, scale, cloud, og, ratio)
    gc = hg.Cost(kind, scale, cloud, gg, ratio)
    assert gc.n == oc.n == len(cloud) and gc.nblocks == oc.nblocks == (4 if two else 2)
    for trial in range(3):
        p = perturbed(world["truth"], rng)
        if two:
            p = np.concatenate([p, perturbed(world["truth"], rng, 0.3, 0.15)])
        r0, J0 = oc.evaluate(p)
        r1, J1 = gc.evaluate(p)
        assert_close(r1, r0, kind + " residuals")
        for b in range(len(J0)):
            assert_close(J1[b], J0[b], "%s jacobian block %d" % (kind, b))
        if len(cloud) > 1:
            assert np.abs(np.hstack(J0)).max() > 1e-4  # the comparison is not vacuous
        # jacobians == NULL, and one block held constant
        r2, _ = gc.evaluate(p, jacobians=False)
        assert_close(r2, r0, kind + " residuals (no jacobians)")
        r3, J3 = gc.evaluate(p, skip_blocks=(1,))
        assert_close(J3[0], J0[0], kind + " jacobian with block 1 skipped")
        assert (J3[1] == 0).all()


@pytest.mark.parametrize("kind", list(KIND_GRIDS.keys()))
def test_packed_stencils_equal_apron_gather(oracle, hg, world, kind, monkeypatch):
    """The packed 2x2x2 stencil copy of a grid (GridDev::stencils) holds the same floats as the
    apron bricks: every cost kind gives bit-identical outputs with and without it."""
    rng = np.random.RandomState(5)
    cloud = world["cloud"]
    gg = [world["g_" + n] for n in KIND_GRIDS[kind]]
    two = kind.startswith("interpolated")
    monkeypatch.setenv("HG_STENCIL_MIN_POINTS", "0")  # by default only batches of >= 32768 points build them
    gc = hg.Cost(kind, 0.01, cloud, gg, 0.6 if two else 0.0)
    assert all(g.stencil_bytes > 0 for g in gg)
    p = perturbed(world["truth"], rng)
    if two:
        p = np.concatenate([p, perturbed(world["truth"], rng, 0.3, 0.15)])
    r1, J1 = gc.evaluate(p)
    monkeypatch.setenv("HG_NO_STENCILS", "1")
    r0, J0 = gc.evaluate(p)
    monkeypatch.delenv("HG_NO_STENCILS")
    assert (r0 == r1).all() and np.abs(r0).max() > 0
    for a, b in zip(J0, J1):
        assert (a == b).all()


def test_packed_stencils_follow_insertions(oracle, hg, monkeypatch):
    """A grid written after a cost function was built on it: the packed stencils are dropped with
    the old voxels and rebuilt, so the next evaluation sees the new map."""
    from hectorgrapher_b200 import synth
    sph = synth.spheres()
    rng = np.random.RandomState(2)
    g = hg.Grid.tsdf(0.10, 2.5, 1000.0, np.zeros((0, 3), np.int32), [], [])
    o = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    clouds = []
    for k in range(3):
        origin, yaw = synth.corkscrew_pose(k, 4)
        pts = synth.scan(origin, yaw, 16, 180, sph, rng)
        world_pts = (synth.quat_rotate(synth.quat_from_yaw(yaw), pts.astype(np.float64)) + origin).astype(np.float32)
        clouds.append((np.asarray(origin, np.float32), world_pts))
    origin, yaw = synth.corkscrew_pose(1, 4)
    local = synth.scan(origin, yaw, 8, 128, sph, rng)
    pose = np.concatenate([origin, synth.quat_from_yaw(yaw)])
    g.insert_tsdf(*clouds[0])
    o.insert_tsdf(*clouds[0])
    monkeypatch.setenv("HG_STENCIL_MIN_POINTS", "0")
    assert g.stencil_bytes == 0
    gc = hg.Cost("tsdf", 0.05, local, [g], 0.0)
    assert g.stencil_bytes > 0
    oc = oracle.Cost("tsdf", 0.05, local, [o], 0.0)
    assert_close(gc.evaluate(pose)[0], oc.evaluate(pose)[0], "before the second insertion")
    before = gc.evaluate(pose)[0]
    for c in clouds[1:]:
        g.insert_tsdf(*c)
        o.insert_tsdf(*c)
        assert g.stencil_bytes == 0  # stale
    r1, J1 = gc.evaluate(pose)
    assert g.stencil_bytes > 0       # rebuilt by the evaluation
    r0, J0 = oc.evaluate(pose)
    assert_close(r1, r0, "after the insertions")
    assert_close(J1[1], J0[1], "after the insertions, rotation block")
    assert np.abs(r1 - before).max() > 1e-6  # the map did change under the cost function


def test_cost_function_slerp_branches(oracle, hg, world):
    """Two-pose functors: lerp branch (identical rotations) and the d < 0 sign flip."""
    cloud = world["cloud"][:500]
    oc = oracle.Cost("interpolated_tsdf", 0.1, cloud, [world["o_tsdf_hi"]], 0.25)
    gc = hg.Cost("interpolated_tsdf", 0.1, cloud, [world["g_tsdf_hi"]], 0.25)
    t = world["truth"]
    same = np.concatenate([t, t[:3] + 0.05, t[3:]])
    flipped = np.concatenate([t, t[:3] + 0.05, -t[3:] + np.array([0, 0.01, 0, 0])])
    for p in [same, flipped]:
        r0, J0 = oc.evaluate(p)
        r1, J1 = gc.evaluate(p)
        assert_close(r1, r0)
        for b in range(4):
            assert_close(J1[b], J0[b], "block %d" % b)


def test_timed_point_cloud_stride(oracle, hg, world):
    """sensor::TimedPointCloud: 16-byte points (xyz + time)."""
    cloud = world["cloud"][:1000]
    timed = np.concatenate([cloud, np.linspace(-0.1, 0, len(cloud), dtype=np.float32)[:, None]], axis=1)
    oc = oracle.Cost("tsdf", 0.2, timed, [world["o_tsdf_hi"]], stride=4)
    gc = hg.Cost("tsdf", 0.2, timed, [world["g_tsdf_hi"]], stride=4)
    p = world["truth"]
    r0, J0 = oc.evaluate(p)
    r1, J1 = gc.evaluate(p)
    assert_close(r1, r0)
    assert_close(J1[1], J0[1])


def test_cost_on_empty_grid(oracle, hg):
    """interpolated_tsdf_space_cost_function_3d_test.cc: an empty TSDF yields constant residuals
    (the minimum TSD) and zero Jacobians."""
    og = oracle.Grid.tsdf(0.1, 2.5, 100.0)
    gg = hg.Grid.tsdf(0.1, 2.5, 100.0, np.zeros((0, 3), np.int32), np.zeros(0, np.uint16), np.zeros(0, np.uint16))
    pts = np.random.RandomState(0).uniform(-3, 3, (64, 3)).astype(np.float32)
    p = np.array([0.1, 0.2, 0.3, 1, 0, 0, 0, 0.2, 0.1, 0.0, 1, 0, 0, 0], float)
    r0, J0 = oracle.Cost("interpolated_tsdf", 0.5, pts, [og], 0.5).evaluate(p)
    r1, J1 = hg.Cost("interpolated_tsdf", 0.5, pts, [gg], 0.5).evaluate(p)
    assert (r1 == r0).all() and np.allclose(r1, 0.5 * np.float32(-0.25))
    assert all((j == 0).all() for j in J1)


@pytest.mark.parametrize("kind", ["tsdf", "interpolated_multires_tsdf", "occupied"])
def test_normal_equations(oracle, hg, world, kind):
    rng = np.random.RandomState(5)
    cloud = world["cloud"]
    og = [world["o_" + n] for n in KIND_GRIDS[kind]]
    gg = [world["g_" + n] for n in KIND_GRIDS[kind]]
    two = kind.startswith("interpolated")
    oc = oracle.Cost(kind, 0.05, cloud, og, 0.6)
    gc = hg.Cost(kind, 0.05, cloud, gg, 0.6)
    p = perturbed(world["truth"], rng)
    if two:
        p = np.concatenate([p, perturbed(world["truth"], rng)])
    r, J = oc.evaluate(p)
    Jf = np.hstack(J)
    JtJ, Jtr, cost = gc.normal_equations(p)
    scale = np.abs(Jf.T @ Jf).max()
    assert np.abs(JtJ - Jf.T @ Jf).max() < 1e-9 * scale
    assert np.abs(Jtr - Jf.T @ r).max() < 1e-9 * max(1.0, np.abs(Jf.T @ r).max())
    assert abs(cost - 0.5 * r @ r) < 1e-10 * max(1.0, 0.5 * r @ r)
    # two calls in a row give identical bits (fixed-order reduction)
    JtJ2, Jtr2, cost2 = gc.normal_equations(p)
    assert (JtJ2 == JtJ).all() and (Jtr2 == Jtr).all() and cost2 == cost


def test_batch_matches_per_block(oracle, hg, world):
    """Ragged batch (one empty block, one 1-point block) against per-block oracle evaluation."""
    rng = np.random.RandomState(8)
    sizes = [700, 0, 1, 1300, 255, 257]
    clouds = [world["cloud"][rng.choice(len(world["cloud"]), n, replace=False)] for n in sizes]
    pts = np.concatenate(clouds)
    offsets = np.concatenate([[0], np.cumsum(sizes)])
    scales = rng.uniform(0.01, 0.2, len(sizes))
    ratios = rng.uniform(0, 1, len(sizes))
    poses = np.stack([perturbed(world["truth"], rng) for _ in range(4)])
    i0 = rng.randint(0, 4, len(sizes))
    i1 = rng.randint(0, 4, len(sizes))
    kind = "interpolated_multires_tsdf"
    gb = hg.Batch(kind, pts, offsets, scales, [world["g_tsdf_hi"], world["g_tsdf_lo"]], ratios)
    r, J = gb.evaluate(poses, i0, i1)
    JtJ, Jtr, cost = gb.normal_equations(poses, i0, i1)
    for b, n in enumerate(sizes):
        if n == 0:
            assert (JtJ[b] == 0).all() and cost[b] == 0
            continue
        oc = oracle.Cost(kind, scales[b], clouds[b], [world["o_tsdf_hi"], world["o_tsdf_lo"]], ratios[b])
        r0, J0 = oc.evaluate(np.concatenate([poses[i0[b]], poses[i1[b]]]))
        sl = slice(offsets[b], offsets[b + 1])
        assert_close(r[sl], r0, "block %d residuals" % b)
        for k in range(4):
            assert_close(J[k][sl], J0[k], "block %d jacobian %d" % (b, k))
        Jf = np.hstack(J0)
        assert np.abs(JtJ[b] - Jf.T @ Jf).max() < 1e-9 * max(1e-12, np.abs(Jf.T @ Jf).max())
        assert abs(cost[b] - 0.5 * r0 @ r0) < 1e-10 * max(1.0, 0.5 * r0 @ r0)


def test_scan_match_pose_parity(oracle, hg, world):
    """CeresScanMatcher3D::Match: same optimiser over oracle vs GPU cost functions, final pose
    within 1e-4 m / 1e-4 rad (north_star)."""
    synth = world["synth"]
    rng = np.random.RandomState(21)
    hi = world["cloud"]
    lo = synth.voxel_downsample(hi, 0.6)
    for grids in [("tsdf_hi", "tsdf_lo"), ("prob", "prob_lo")]:
        og = [world["o_" + g] for g in grids]
        gg = [world["g_" + g] for g in grids]
        for trial in range(2):
            init = perturbed(world["truth"], rng, 0.08, 0.03)
            p0, s0, _ = oracle.scan_match([hi, lo], og, [5.0, 30.0], 10.0, 1.0, init, max_num_iterations=12,
                                          use_nonmonotonic_steps=False)
            p1, s1 = hg.scan_match([hi, lo], gg, [5.0, 30.0], 10.0, 1.0, init, max_num_iterations=12,
                                   use_nonmonotonic_steps=False)
            assert np.abs(p1[:3] - p0[:3]).max() < 1e-4
            q0, q1 = p0[3:] / np.linalg.norm(p0[3:]), p1[3:] / np.linalg.norm(p1[3:])
            assert 2 * np.arccos(min(1.0, abs(float(q0 @ q1)))) < 1e-4
            assert abs(s1[1] - s0[1]) <= 1e-6 * max(1.0, s0[1])
            assert s1[1] < s1[0]


# --------------------------------------------------------- precomputation grids --

@pytest.mark.parametrize("name,depth,frd", [("prob", 7, 3), ("tsdf_hi", 7, 3), ("prob", 5, 5), ("tsdf_lo", 4, 1)])
def test_precomputation_stack_bit_exact(oracle, hg, world, name, depth, frd):
    os_ = oracle.Stack(world["o_" + name], depth, frd)
    gs = hg.Stack(world["g_" + name], depth, frd)
    assert gs.depth == depth
    rng = np.random.RandomState(depth)
    for d in range(depth):
        xo, vo = os_.export(d)
        xo, vo = xo[vo != 0], vo[vo != 0]
        xg, vg = gs.export(d)
        assert len(xo) == len(xg), "level %d: %d vs %d non-zero voxels" % (d, len(xo), len(xg))
        a, b = np.lexsort(xo.T[::-1]), np.lexsort(xg.T[::-1])
        assert (xo[a] == xg[b]).all() and (vo[a] == vg[b]).all(), "level %d differs" % d
        q = np.concatenate([xo[rng.randint(0, len(xo), 4000)] + rng.randint(-9, 10, (4000, 3)),
                            rng.randint(-4000, 4000, (500, 3))]).astype(np.int32)
        assert (gs.lookup(d, q) == os_.lookup(d, q)).all()


def test_precomputation_reference_known_answer(oracle, hg):
    """precomputation_grid_3d_test.cc:30-77 on the GPU stack (full-resolution levels)."""
    rng = np.random.RandomState(23847)
    og = oracle.Grid.prob(2.0)
    xyz = rng.randint(-50, 50, size=(1000, 3)).astype(np.int32)
    p = rng.uniform(0.1, 0.9, size=1000).astype(np.float32)
    for i in range(1000):
        og.set_probability(xyz[i:i + 1], p[i:i + 1])
    ex, ev = og.export()
    gs = hg.Stack(hg.Grid.prob(2.0, ex, ev), 4, 4)
    for depth in range(4):
        width = 1 << depth
        q = rng.randint(-50, 50, size=(100, 3)).astype(np.int32)
        got = gs.lookup(depth, q)
        d = np.arange(width)
        off = np.stack(np.meshgrid(d, d, d, indexing="ij"), -1).reshape(-1, 3)
        for i in range(100):
            naive = float(og.lookup("probability", q[i] + off).max())
            assert abs(naive - (0.1 + got[i] * (0.8 / 255.0))) < 1e-2


# ------------------------------------------------------- correlative matcher --

def node_inputs(world, n_high=600, n_low=150):
    synth = world["synth"]
    hi = synth.voxel_downsample(world["cloud"], 0.25, n_high, seed=1)
    lo = synth.voxel_downsample(world["cloud"], 0.9, n_low, seed=2)
    hist = synth.rotational_histogram(hi, 60)
    return hi, lo, hist


def world_matchers(oracle, hg, world, grid="tsdf", strict_visit_order=0, **opts):
    names = ("tsdf_hi", "tsdf_lo") if grid == "tsdf" else ("prob", "prob_lo")
    hi, lo, hist = node_inputs(world)
    o = dict(branch_and_bound_depth=6, full_resolution_depth=3, min_rotational_score=0.5,
             min_low_resolution_score=0.3, linear_xy_search_window=2.0, linear_z_search_window=0.6,
             angular_search_window=0.12)
    o.update(opts)
    # the submap's histogram is the node's, rotated into the submap frame (mapping/3d/submap_3d.cc)
    truth_yaw = 2 * np.arctan2(world["truth"][6], world["truth"][3])
    submap_hist = oracle.rotate_histogram(hist, np.float32(truth_yaw))
    om = oracle.Fcsm(world["o_" + names[0]], world["o_" + names[1]], submap_hist, **o)
    gm = hg.Fcsm(world["g_" + names[0]], world["g_" + names[1]], submap_hist,
                 strict_visit_order=strict_visit_order, **o)
    return om, gm, hi, lo, hist


@pytest.mark.parametrize("grid", ["tsdf", "prob"])
def test_discrete_scans_bit_exact(oracle, hg, world, grid):
    om, gm, hi, lo, hist = world_matchers(oracle, hg, world, grid)
    rng = np.random.RandomState(4)
    node = perturbed(world["truth"], rng, 0.5, 0.05)
    submap = np.array([0.3, -0.2, 0.1, *world["synth"].small_rotation(rng, 0.2)])
    for (wxy, wz, ang) in [(20, 6, 0.12), (7, 3, 0.05)]:
        p0, r0, c0 = om.discrete_scans(wxy, wz, ang, node, submap, hi, hist)
        p1, r1, c1 = gm.discrete_scans(wxy, wz, ang, node, submap, hi, hist)
        assert p0.shape == p1.shape and len(p0) > 3
        assert (p0.view(np.uint32) == p1.view(np.uint32)).all()      # poses bit-exact
        assert (r0.view(np.uint32) == r1.view(np.uint32)).all()      # rotational scores bit-exact
        assert (c0 == c1).all()                                      # rotated-scan voxel indices


@pytest.mark.parametrize("grid", ["tsdf", "prob"])
def test_candidate_scores_bit_exact(oracle, hg, world, grid):
    om, gm, hi, lo, hist = world_matchers(oracle, hg, world, grid)
    rng = np.random.RandomState(6)
    node = perturbed(world["truth"], rng, 0.3, 0.03)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    wxy, wz, ang = 20, 6, 0.12
    S = len(om.discrete_scans(wxy, wz, ang, node, submap, hi, hist)[0])
    for depth in range(6):
        n = 3000
        cand = np.stack([rng.randint(0, S, n), rng.randint(-wxy, wxy + 1, n), rng.randint(-wxy, wxy + 1, n),
                         rng.randint(-wz, wz + 1, n)], 1).astype(np.int32)
        s0, f0, _ = om.score_candidates(wxy, wz, ang, node, submap, hi, hist, depth, cand)
        s1, f1, ms = gm.score_candidates(wxy, wz, ang, node, submap, hi, hist, depth, cand)
        assert (s0 == s1).all(), "integer sums differ at depth %d" % depth
        assert (f0.view(np.uint32) == f1.view(np.uint32)).all()
        assert s0.max() > 0 and ms > 0


@pytest.mark.parametrize("grid", ["tsdf", "prob"])
def test_low_resolution_scores_bit_exact(oracle, hg, world, grid):
    name = "tsdf_lo" if grid == "tsdf" else "prob_lo"
    rng = np.random.RandomState(9)
    lo = node_inputs(world)[1]
    poses = np.stack([perturbed(world["truth"], rng, 1.0, 0.3) for _ in range(200)]).astype(np.float32)
    got = hg.low_resolution_scores(world["g_" + name], lo, poses)
    want = np.array([oracle.low_resolution_score(world["o_" + name], lo, p) for p in poses], np.float32)
    assert (got.view(np.uint32) == want.view(np.uint32)).all()


RESULT_KEYS = ["found", "score", "tx", "ty", "tz", "qw", "qx", "qy", "qz", "rotational_score",
               "low_resolution_score", "scan_index", "ox", "oy", "oz", "num_scans"]


def assert_same_match(r0, r1):
    for k in RESULT_KEYS:
        assert np.float64(r0[k]) == np.float64(r1[k]), (k, r0[k], r1[k])


@pytest.mark.parametrize("grid", ["tsdf", "prob"])
def test_match_selects_the_same_candidate(oracle, hg, world, grid):
    om, gm, hi, lo, hist = world_matchers(oracle, hg, world, grid)
    _, gm_strict, _, _, _ = world_matchers(oracle, hg, world, grid, strict_visit_order=1)
    rng = np.random.RandomState(12)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    found = 0
    for trial in range(6):
        node = perturbed(world["truth"], rng, 1.2, 0.08)
        for min_score in [0.2, 0.32]:
            r0 = om.match(node, submap, hi, lo, hist, min_score=min_score)
            r1 = gm.match(node, submap, hi, lo, hist, min_score=min_score)
            r2 = gm_strict.match(node, submap, hi, lo, hist, min_score=min_score)
            assert_same_match(r0, r1)
            assert_same_match(r0, r2)
            assert r2["num_scored"] >= r2["num_visited"] > 0
            # strict order: the replay consumes exactly the candidates the reference's DFS scores
            assert r2["num_visited"] == r0["num_scored"]
            found += int(r1["found"])
    assert found >= 3
    # a cloud that cannot pass the low-resolution gate gives a null result in both
    far = np.array([[42., 42., 42.]], np.float32)
    r0 = om.match(world["truth"], submap, hi, far, hist, min_score=0.2)
    r1 = gm.match(world["truth"], submap, hi, far, hist, min_score=0.2)
    assert r0["found"] == r1["found"] == 0.0


def test_match_full_submap_selects_the_same_candidate(oracle, hg, world):
    om, gm, hi, lo, hist = world_matchers(oracle, hg, world, "tsdf", branch_and_bound_depth=7,
                                          use_rotational_scan_matcher=False)
    hi = hi[:200]
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    node = world["truth"]
    r0 = om.match(node, submap, hi, lo, hist, min_score=0.4, full=True)
    r1 = gm.match(node, submap, hi, lo, hist, min_score=0.4, full=True)
    assert_same_match(r0, r1)
    assert r1["found"] == 1.0
    # sharded by discrete scan: the best over shards equals the unsharded answer
    shards = [gm.match(node, submap, hi, lo, hist, min_score=0.4, full=True, shard_index=s, num_shards=3)
              for s in range(3)]
    assert sum(int(s["num_scans"]) for s in shards) == int(r1["num_scans"])
    best = max((s for s in shards if s["found"]), key=lambda s: s["score"])
    assert best["score"] == r1["score"]
    assert (best["scan_index"], best["ox"], best["oy"], best["oz"]) == (
        r1["scan_index"], r1["ox"], r1["oy"], r1["oz"])


def _rot_z(theta):
    return np.array([np.cos(theta / 2), 0, 0, np.sin(theta / 2)])


def test_reference_fcsm_fixture_on_gpu(oracle, hg):
    """fast_correlative_scan_matcher_3d_test.cc:141-196 through the GPU matcher, and identical to
    the oracle's result."""
    from hectorgrapher_b200 import synth
    cloud = np.array([[4, 0, 0], [4.5, 0, 0], [5, 0, 0], [5.5, 0, 0], [0, 4, 0], [0, 4.5, 0], [0, 5, 0],
                      [0, 5.5, 0], [0, 0, 4], [0, 0, 4.5], [0, 0, 5], [0, 0, 5.5]], np.float32)
    opts = dict(branch_and_bound_depth=6, full_resolution_depth=6, use_rotational_scan_matcher=True,
                min_rotational_score=0.1, min_low_resolution_score=0.15, linear_xy_search_window=0.8,
                linear_z_search_window=0.8, angular_search_window=0.3)
    rng = np.random.RandomState(42)
    ident = [0, 0, 0, 1, 0, 0, 0]
    for trial in range(4):
        t = (0.7 * rng.uniform(-1, 1, 3)).astype(np.float32)
        q = _rot_z(np.float32(0.2 * rng.uniform(-1, 1))).astype(np.float32)
        w = (synth.quat_rotate(q, cloud) + t).astype(np.float32)
        og = oracle.Grid.prob(0.05)
        og.insert_occupancy(t, w, 0.7, 0.4, 5)
        ex, ev = og.export()
        gg = hg.Grid.prob(0.05, ex, ev, grid_size=og.grid_size)
        om = oracle.Fcsm(og, og, np.zeros(10, np.float32), **opts)
        gm = hg.Fcsm(gg, gg, np.zeros(10, np.float32), **opts)
        for full in ([False, True] if trial == 0 else [False]):
            r0 = om.match(ident, ident, cloud, cloud, np.zeros(10), min_score=0.1, full=full)
            r1 = gm.match(ident, ident, cloud, cloud, np.zeros(10), min_score=0.1, full=full)
            assert_same_match(r0, r1)
            assert r1["found"] == 1.0 and r1["score"] > 0.1
            assert np.abs(np.array([r1["tx"], r1["ty"], r1["tz"]]) - t).max() < 0.05
            far = gm.match(ident, ident, cloud, [[42., 42., 42.]], np.zeros(10), min_score=0.1, full=full)
            assert far["found"] == 0.0


def test_constraint_batch(oracle, hg, world):
    """ConstraintBuilder3D::ComputeConstraint for a few pairs: correlative match + refine equals the
    oracle's match followed by the oracle's refine (pose within 1e-4)."""
    om, gm, hi, lo, hist = world_matchers(oracle, hg, world, "tsdf")
    rng = np.random.RandomState(31)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    jobs, nodes = [], []
    for j in range(4):
        node = perturbed(world["truth"], rng, 1.0, 0.06)
        nodes.append(node)
        jobs.append(dict(submap_index=0, full=False, node_pose=node, submap_pose=submap, high=hi, low=lo, hist=hist))
    opts = dict(min_score=0.25, global_localization_min_score=0.4, occupied_space_weights=(5.0, 30.0),
                translation_weight=10.0, rotation_weight=1.0, max_num_iterations=10)
    res = hg.constraints_run_batch([gm], [world["g_tsdf_hi"]], [world["g_tsdf_lo"]], jobs, opts)
    shard0 = hg.constraints_run_batch([gm], [world["g_tsdf_hi"]], [world["g_tsdf_lo"]], jobs, opts, 0, 2)
    for j, node in enumerate(nodes):
        r0 = om.match(node, submap, hi, lo, hist, min_score=0.25)
        assert res[j]["found"] == int(r0["found"])
        if not r0["found"]:
            continue
        init = np.array([r0[k] for k in ["tx", "ty", "tz", "qw", "qx", "qy", "qz"]])
        p0, _, _ = oracle.scan_match([hi, lo], [world["o_tsdf_hi"], world["o_tsdf_lo"]], [5.0, 30.0], 10.0, 1.0,
                                     init, max_num_iterations=10, use_nonmonotonic_steps=False)
        assert res[j]["score"] == np.float32(r0["score"])
        assert np.abs(res[j]["pose"][:3] - p0[:3]).max() < 1e-4
        if j % 2 == 0:
            assert np.abs(shard0[j]["pose"] - res[j]["pose"]).max() == 0


# ------------------------------------------------------------ BASELINE.json configs --

def test_config_c1_scan_to_submap(oracle, hg):
    """C1: one scan of 32 x 640 = 20 480 points against a 5 cm TSDF (+ 0.45 m low-resolution grid):
    one Evaluate and a full CeresScanMatcher3D::Match, GPU vs oracle."""
    from hectorgrapher_b200 import synth
    sph = synth.spheres()
    poses = [synth.corkscrew_pose(k, 3) for k in range(3)]
    hi = synth.build_tsdf(0.05, poses, 48, 512, sph)
    lo = synth.build_tsdf(0.45, poses, 48, 512, sph)
    o_hi, o_lo = oracle.Grid.tsdf(0.05, 2.5, 1000.0), oracle.Grid.tsdf(0.45, 2.5, 1000.0)
    o_hi.set_raw(*hi)
    o_lo.set_raw(*lo)
    g_hi = hg.Grid.tsdf(0.05, 2.5, 1000.0, *hi, grid_size=o_hi.grid_size)
    g_lo = hg.Grid.tsdf(0.45, 2.5, 1000.0, *lo, grid_size=o_lo.grid_size)
    rng = np.random.RandomState(1)
    origin, yaw = poses[1]
    cloud = synth.scan(origin, yaw, 32, 640, sph, rng)
    assert len(cloud) == 20480
    truth = np.concatenate([origin, synth.quat_from_yaw(yaw)])
    init = perturbed(truth, rng, 0.1, 0.05)
    scale = 5.0 / np.sqrt(len(cloud))
    r0, J0 = oracle.Cost("tsdf", scale, cloud, [o_hi]).evaluate(init)
    r1, J1 = hg.Cost("tsdf", scale, cloud, [g_hi]).evaluate(init)
    assert_close(r1, r0, "C1 residuals")
    assert_close(J1[0], J0[0], "C1 translation jacobian")
    assert_close(J1[1], J0[1], "C1 rotation jacobian")
    low = synth.voxel_downsample(cloud, 0.5)
    p0, s0, _ = oracle.scan_match([cloud, low], [o_hi, o_lo], [5.0, 30.0], 10.0, 1.0, init, max_num_iterations=12,
                                  use_nonmonotonic_steps=False)
    p1, s1 = hg.scan_match([cloud, low], [g_hi, g_lo], [5.0, 30.0], 10.0, 1.0, init, max_num_iterations=12,
                           use_nonmonotonic_steps=False)
    assert np.abs(p1[:3] - p0[:3]).max() < 1e-4
    q0, q1 = p0[3:] / np.linalg.norm(p0[3:]), p1[3:] / np.linalg.norm(p1[3:])
    assert 2 * np.arccos(min(1.0, abs(float(q0 @ q1)))) < 1e-4
    assert s1[1] < s1[0]


def test_config_c3_local_search(oracle, hg, world):
    """C3: FastCorrelativeScanMatcher3D::Match with the reference's defaults (pose_graph.lua:40-49:
    7-level pyramid, full_resolution_depth 3, windows 5 m / 1 m / 15 deg, min_rotational_score 0.77,
    min_low_resolution_score 0.55), N = 2048 points. (The small test submap is observed sparsely, so
    scores top out near 0.5; min_score is 0.25 here instead of pose_graph.lua's 0.55.)"""
    synth = world["synth"]
    rng = np.random.RandomState(3)
    origin, yaw = world["poses"][1]
    cloud = synth.scan(origin, yaw, 32, 512, world["sph"], rng)
    hi = synth.voxel_downsample(cloud, 0.15, 2048, seed=3)
    lo = synth.voxel_downsample(cloud, 0.9, 200, seed=4)
    assert len(hi) == 2048
    hist = synth.rotational_histogram(hi, 120)
    submap_hist = oracle.rotate_histogram(hist, np.float32(yaw))
    opts = dict(branch_and_bound_depth=7, full_resolution_depth=3, use_rotational_scan_matcher=True,
                min_rotational_score=0.77, min_low_resolution_score=0.55, linear_xy_search_window=5.0,
                linear_z_search_window=1.0, angular_search_window=float(np.deg2rad(15.0)))
    om = oracle.Fcsm(world["o_tsdf_hi"], world["o_tsdf_lo"], submap_hist, **opts)
    gm = hg.Fcsm(world["g_tsdf_hi"], world["g_tsdf_lo"], submap_hist, **opts)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    hits = 0
    for trial in range(2):
        node = perturbed(world["truth"], rng, 0.0, 0.15)
        node[:3] += np.array([rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(-0.5, 0.5)])
        r0 = om.match(node, submap, hi, lo, hist, min_score=0.25)
        r1 = gm.match(node, submap, hi, lo, hist, min_score=0.25)
        assert_same_match(r0, r1)
        hits += int(r1["found"])
        if r1["found"]:
            assert np.abs(np.array([r1["tx"], r1["ty"]]) - world["truth"][:2]).max() < 0.35
    assert hits >= 1


# ------------------------------------------------------------------ more edge cases --

def test_three_level_pyramid_and_yaw_only(oracle, hg, world):
    """Multi-resolution lookup over three TSDF levels (fall-through twice), and
    CeresScanMatcher3D with only_optimize_yaw (YawOnlyQuaternionPlus)."""
    synth = world["synth"]
    mid = synth.build_tsdf(0.2, world["poses"], 32, 360, world["sph"])
    o_mid = oracle.Grid.tsdf(0.2, 2.5, 1000.0)
    o_mid.set_raw(*mid)
    g_mid = hg.Grid.tsdf(0.2, 2.5, 1000.0, *mid, grid_size=o_mid.grid_size)
    og = [world["o_tsdf_hi"], o_mid, world["o_tsdf_lo"]]
    gg = [world["g_tsdf_hi"], g_mid, world["g_tsdf_lo"]]
    rng = np.random.RandomState(2)
    # shift the cloud so that many points leave the fine band and fall through to coarser levels
    cloud = (world["cloud"] * np.float32(1.01) + np.float32(0.07)).astype(np.float32)
    p = perturbed(world["truth"], rng)
    r0, J0 = oracle.Cost("multires_tsdf", 0.03, cloud, og).evaluate(p)
    r1, J1 = hg.Cost("multires_tsdf", 0.03, cloud, gg).evaluate(p)
    assert_close(r1, r0, "3-level residuals")
    assert_close(J1[0], J0[0])
    assert_close(J1[1], J0[1])
    # ascending-resolution check (interpolated_multi_resolution_tsdf.h:62-67)
    with pytest.raises(hg.HgError):
        hg.Cost("multires_tsdf", 0.03, cloud, [gg[1], gg[0]])
    init = perturbed(world["truth"], rng, 0.05, 0.02)
    p0, s0, _ = oracle.scan_match([world["cloud"]], [world["o_tsdf_hi"]], [5.0], 10.0, 1.0, init,
                                  only_optimize_yaw=True, max_num_iterations=10, use_nonmonotonic_steps=False)
    p1, s1 = hg.scan_match([world["cloud"]], [world["g_tsdf_hi"]], [5.0], 10.0, 1.0, init, only_optimize_yaw=True,
                           max_num_iterations=10, use_nonmonotonic_steps=False)
    assert np.abs(p1 - p0).max() < 1e-4


def test_hash_indexed_grid_everywhere(oracle, hg, world):
    """A grid whose bricks are so scattered that the dense brick index is replaced by the hash:
    lookups, cost evaluation and the pyramid still agree with the oracle."""
    xyz, tv, wv = world["tsdf_hi"]
    far = np.array([[8000, 8000, 8000], [-8100, -8000, 7000]], np.int32)
    xyz2 = np.concatenate([xyz, far])
    tv2 = np.concatenate([tv, np.array([0x8000 | 20000, 0x8000 | 9000], np.uint16)])
    wv2 = np.concatenate([wv, np.array([500, 700], np.uint16)])
    og = oracle.Grid.tsdf(0.10, 2.5, 1000.0)
    og.set_raw(xyz2, tv2, wv2)
    gg = hg.Grid.tsdf(0.10, 2.5, 1000.0, xyz2, tv2, wv2)
    assert gg.grid_size == og.grid_size == 16384
    rng = np.random.RandomState(7)
    q = np.concatenate([xyz2[rng.randint(0, len(xyz2), 5000)] + rng.randint(-2, 3, (5000, 3)), far + 1, far]).astype(np.int32)
    for what in ["tsd", "weight"]:
        assert (gg.lookup(what, q).view(np.uint32) == og.lookup(what, q).view(np.uint32)).all()
    p = perturbed(world["truth"], rng)
    r0, J0 = oracle.Cost("tsdf", 0.05, world["cloud"], [og]).evaluate(p)
    r1, J1 = hg.Cost("tsdf", 0.05, world["cloud"], [gg]).evaluate(p)
    assert_close(r1, r0)
    assert_close(J1[1], J0[1])
    os_, gs = oracle.Stack(og, 4, 2), hg.Stack(gg, 4, 2)
    for d in range(4):
        xo, vo = os_.export(d)
        xo, vo = xo[vo != 0], vo[vo != 0]
        xg, vg = gs.export(d)
        a, b = np.lexsort(xo.T[::-1]), np.lexsort(xg.T[::-1])
        assert len(xo) == len(xg) and (xo[a] == xg[b]).all() and (vo[a] == vg[b]).all()


def test_gravity_alignment_and_scalar_float_flavor(oracle, hg, world):
    """Non-identity gravity alignment feeds the rotational prefilter's initial yaw; float_flavor 0
    (Eigen's scalar summation orders) is matched against the oracle in the same flavor."""
    synth = world["synth"]
    rng = np.random.RandomState(14)
    node = perturbed(world["truth"], rng, 0.6, 0.05)
    submap = np.array([0.2, 0.1, -0.1, *synth.small_rotation(rng, 0.1)])
    gravity = synth.small_rotation(rng, 0.3)
    for flavor in (1, 0):
        oracle.lib().orc_set_float_flavor(flavor)
        try:
            om, gm, hi, lo, hist = world_matchers(oracle, hg, world, "tsdf", float_flavor=flavor)
            p0, r0, c0 = om.discrete_scans(20, 6, 0.12, node, submap, hi, hist, gravity=gravity)
            p1, r1, c1 = gm.discrete_scans(20, 6, 0.12, node, submap, hi, hist, gravity=gravity)
            assert p0.shape == p1.shape and (p0.view(np.uint32) == p1.view(np.uint32)).all()
            assert (r0.view(np.uint32) == r1.view(np.uint32)).all() and (c0 == c1).all()
            a = om.match(node, submap, hi, lo, hist, gravity=gravity, min_score=0.2)
            b = gm.match(node, submap, hi, lo, hist, gravity=gravity, min_score=0.2)
            assert_same_match(a, b)
        finally:
            oracle.lib().orc_set_float_flavor(1)


def test_far_and_degenerate_points(oracle, hg, world):
    """Points far outside every grid, at the origin, and a scan of one point."""
    rng = np.random.RandomState(3)
    pts = np.concatenate([world["cloud"][:100], np.array([[0, 0, 0], [1e4, -1e4, 3e3], [500.5, 0.05, -0.05]], np.float32)])
    p = perturbed(world["truth"], rng)
    for kind, names in [("occupied", ["prob"]), ("multires_tsdf", ["tsdf_hi", "tsdf_lo"])]:
        r0, J0 = oracle.Cost(kind, 0.1, pts, [world["o_" + n] for n in names]).evaluate(p)
        r1, J1 = hg.Cost(kind, 0.1, pts, [world["g_" + n] for n in names]).evaluate(p)
        assert_close(r1, r0, kind)
        assert_close(J1[0], J0[0], kind)
        assert_close(J1[1], J0[1], kind)
    one = world["cloud"][5:6]
    r0, J0 = oracle.Cost("interpolated_occupied", 1.0, one, [world["o_prob"]], 0.5).evaluate(np.concatenate([p, p]))
    r1, J1 = hg.Cost("interpolated_occupied", 1.0, one, [world["g_prob"]], 0.5).evaluate(np.concatenate([p, p]))
    assert_close(r1, r0)
    for b in range(4):
        assert_close(J1[b], J0[b])


# ------------------------------------------------------------------ voxel filters (N2) --

def test_voxel_filter_bit_exact(oracle, hg, world):
    """sensor::VoxelFilter on the device: exactly the points the CPU filter keeps, in input order."""
    rng = np.random.RandomState(0)
    clouds = [world["cloud"], rng.uniform(-30, 30, (50000, 3)).astype(np.float32),
              np.array([[0, 0, 0], [0.1, -0.1, 0.1], [0.3, -0.1, 0], [0, 0, 0.1]], np.float32),          # voxel_filter_test.cc:30-37
              np.array([[100000, 0, 0], [100000.001, -0.0001, 0.0001], [100000.003, -0.0001, 0], [-200000, 0, 0]],
                       np.float32)]                                                                      # :39-46
    for pts in clouds:
        for size in [0.01, 0.05, 0.3, 1.7]:
            a, b = hg.voxel_filter(size, pts), oracle.voxel_filter(size, pts)
            assert a.tolist() == b.tolist(), (len(pts), size)
    assert hg.voxel_filter(0.3, clouds[2]).tolist() == [0, 2]
    assert hg.voxel_filter(0.01, clouds[3]).tolist() == [0, 3]
    timed = np.array([[-100, 0.3, 0.4, float(i)] for i in range(100)], np.float32)                       # :48-55
    assert hg.voxel_filter(0.3, timed, stride=4).tolist() == [0]
    assert hg.voxel_filter(0.3, np.zeros((0, 3), np.float32)).tolist() == []
    # half-cell boundaries: points exactly between two voxels round away from zero
    k = rng.randint(-200, 200, (4000, 3))
    halves = ((k + 0.5) * np.float32(0.25)).astype(np.float32)
    assert hg.voxel_filter(0.25, halves).tolist() == oracle.voxel_filter(0.25, halves).tolist()


def test_adaptive_voxel_filter_bit_exact(oracle, hg, world):
    """sensor::AdaptiveVoxelFilter (max range, halving, bisection) — same kept indices as the CPU."""
    rng = np.random.RandomState(1)
    big = rng.uniform(-20, 20, (30000, 3)).astype(np.float32)
    cases = [(world["cloud"], 2.0, 150, 50.0),      # trajectory_builder_3d.lua:23-27 high-resolution filter
             (world["cloud"], 4.0, 200, 60.0),      # low-resolution filter
             (big, 4.0, 200, 25.0), (big, 0.5, 150, 15.0), (big, 8.0, 25000, 100.0), (big[:100], 1.0, 500, 30.0),
             (big, 0.05, 29000, 100.0)]             # cannot be met: gives up after the halving loop
    for pts, max_length, min_points, max_range in cases:
        a = hg.adaptive_voxel_filter(max_length, min_points, max_range, pts)
        b = oracle.adaptive_voxel_filter(max_length, min_points, max_range, pts)
        assert a.tolist() == b.tolist(), (len(pts), max_length, min_points, max_range, len(a), len(b))


# ------------------------------------------------ BASELINE.json configs[1] at full size --

def test_config_c2_full_size(oracle, hg):
    """C2 exactly as bench.py times it: 64 two-pose multi-resolution TSDF blocks x 100 000 points in one
    batched Evaluate. At this size the oracle takes minutes, so the check is (a) the oracle on a random
    sample of points of several blocks — rows are per-point, so a sub-cloud with the block's scaling factor
    must give the same rows —, (b) properties that do not depend on the size: the fused normal equations
    equal J'J / J'r of the materialised rows, and reversing the points of the batch reverses the rows
    bit for bit."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    w = bench.build_workload()
    N = len(w["points"])
    assert N == 64 * 100000
    g_hi = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"])
    g_lo = hg.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["lo"])
    batch = hg.Batch("interpolated_multires_tsdf", w["points"], w["offsets"], w["scales"], [g_hi, g_lo], w["ratios"])
    r, J = batch.evaluate(w["poses"], w["i0"], w["i1"])
    assert np.isfinite(r).all() and all(np.isfinite(j).all() for j in J)
    assert np.abs(J[1]).max() > 1e-4 and np.abs(J[3]).max() > 1e-4

    # (a) oracle on a sample
    o_hi = oracle.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT)
    o_hi.set_raw(*w["hi"])
    o_lo = oracle.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT)
    o_lo.set_raw(*w["lo"])
    rng = np.random.RandomState(21)
    for b in (0, 17, 40, 63):
        pick = np.sort(rng.choice(100000, 600, replace=False)) + b * 100000
        oc = oracle.Cost("interpolated_multires_tsdf", w["scales"][b], w["points"][pick], [o_hi, o_lo], w["ratios"][b])
        r0, J0 = oc.evaluate(np.concatenate([w["poses"][w["i0"][b]], w["poses"][w["i1"][b]]]))
        assert_close(r[pick], r0, "C2 block %d residuals" % b)
        for k in range(4):
            assert_close(J[k][pick], J0[k], "C2 block %d jacobian %d" % (b, k))

    # (b1) fused normal equations == J'J, J'r, 0.5 r'r of the rows, per block
    JtJ, Jtr, cost = batch.normal_equations(w["poses"], w["i0"], w["i1"])
    for b in (0, 31, 63):
        s = slice(b * 100000, (b + 1) * 100000)
        rows = np.hstack([J[0][s], J[1][s], J[2][s], J[3][s]])
        want = rows.T @ rows
        scale = np.abs(want).max()
        assert np.abs(JtJ[b] - want).max() <= 1e-9 * scale
        assert np.abs(Jtr[b] - rows.T @ r[s]).max() <= 1e-9 * max(1.0, np.abs(Jtr[b]).max())
        assert abs(cost[b] - 0.5 * float(r[s] @ r[s])) <= 1e-9 * max(1.0, cost[b])

    # (b2) reversing the points inside every block reverses the rows, bit for bit
    rev = np.ascontiguousarray(w["points"].reshape(64, 100000, 3)[:, ::-1, :].reshape(-1, 3))
    batch_rev = hg.Batch("interpolated_multires_tsdf", rev, w["offsets"], w["scales"], [g_hi, g_lo], w["ratios"])
    r2, J2 = batch_rev.evaluate(w["poses"], w["i0"], w["i1"])
    assert (r2.reshape(64, 100000)[:, ::-1] == r.reshape(64, 100000)).all()
    for k in range(4):
        width = J[k].shape[1]
        assert (J2[k].reshape(64, 100000, width)[:, ::-1] == J[k].reshape(64, 100000, width)).all()


@pytest.mark.parametrize("initial", [0, 1, 2, 3, 4])
def test_reference_ceres_fixture_on_gpu(oracle, hg, initial):
    """ceres_scan_matcher_3d_test.cc:36-117 through hg_scan_match at the reference's settings
    (use_nonmonotonic_steps, max_num_iterations = 10; final cost < 1e-2, IsNearly 3e-2): the
    one-launch device solve and the host loop over GPU costs, each against the same solver over
    the oracle's Jet costs."""
    import os
    from test_oracle_reference_kats import CERES_FIXTURE_STARTS, ceres_fixture, is_nearly
    cloud, grid_points, start, expected_t, expected_q = ceres_fixture(None, CERES_FIXTURE_STARTS[initial])
    og = oracle.Grid.prob(1.0)
    og.set_probability(og.cell_index(grid_points), 1.0)
    ex, ev = og.export()
    gg = hg.Grid.prob(1.0, ex, ev, grid_size=og.grid_size)
    kw = dict(max_num_iterations=10, use_nonmonotonic_steps=True)
    pose_o, sum_o, _ = oracle.scan_match([cloud], [og], [1.0], 0.01, 0.1, start, **kw)
    for host in (False, True):
        if host:
            os.environ["HG_SCAN_MATCH_HOST"] = "1"
        try:
            pose, summary = hg.scan_match([cloud], [gg], [1.0], 0.01, 0.1, start, **kw)
        finally:
            os.environ.pop("HG_SCAN_MATCH_HOST", None)
        assert abs(summary[1]) < 1e-2
        assert is_nearly(pose, expected_t, expected_q, 3e-2)
        assert (summary[2:] == sum_o[2:]).all(), (summary, sum_o)  # same accept / reject sequence
        assert np.abs(pose - pose_o).max() < 1e-4


def test_device_scan_match_equals_host_loop(oracle, hg, world):
    """hg_scan_match solves on the device in one launch (scan_match.cu); HG_SCAN_MATCH_HOST=1 selects
    the host loop of include/hg_dense_lm.h over the fused normal equations. Same solver, same
    problem: the poses agree to rounding, the iteration counts exactly; both agree with the oracle
    (Jet cost functions under the same host solver) to the 1e-4 bar."""
    import os
    rng = np.random.RandomState(33)
    cloud = world["cloud"]
    low = cloud[::7]
    cases = [
        (["g_tsdf_hi", "g_tsdf_lo"], ["o_tsdf_hi", "o_tsdf_lo"], dict()),
        (["g_prob", "g_prob_lo"], ["o_prob", "o_prob_lo"], dict()),
        (["g_tsdf_hi"], ["o_tsdf_hi"], dict(only_optimize_yaw=True)),
        (["g_tsdf_hi", "g_tsdf_lo"], ["o_tsdf_hi", "o_tsdf_lo"], dict(use_nonmonotonic_steps=True,
                                                                         max_num_iterations=25)),
    ]
    # more than 4096 points: the solve runs on a cluster of eight CTAs (distributed shared memory)
    big = np.concatenate([cloud + rng.normal(0, 0.004, cloud.shape).astype(np.float32) for _ in range(4)])
    assert len(big) > 3 * 4096
    cases.append((["g_tsdf_hi", "g_tsdf_lo"], ["o_tsdf_hi", "o_tsdf_lo"], dict(big=True)))
    for gnames, onames, extra in cases:
        extra = dict(extra)
        clouds = [big if extra.pop("big", False) else cloud, low][:len(gnames)]
        weights = [5.0, 30.0][:len(gnames)]
        kw = dict(max_num_iterations=12, use_nonmonotonic_steps=False)
        kw.update(extra)
        for trial in range(3):
            start = perturbed(world["truth"], rng, 0.08, 0.03)
            os.environ.pop("HG_SCAN_MATCH_HOST", None)
            pose_d, sum_d = hg.scan_match(clouds, [world[n] for n in gnames], weights, 10.0, 1.0, start, **kw)
            os.environ["HG_SCAN_MATCH_HOST"] = "1"
            try:
                pose_h, sum_h = hg.scan_match(clouds, [world[n] for n in gnames], weights, 10.0, 1.0, start, **kw)
            finally:
                os.environ.pop("HG_SCAN_MATCH_HOST", None)
            assert np.abs(pose_d - pose_h).max() < 1e-9, (gnames, extra, pose_d, pose_h)
            assert (sum_d[2:] == sum_h[2:]).all(), (sum_d, sum_h)   # iterations, steps, evaluations
            assert abs(sum_d[1] - sum_h[1]) <= 1e-9 * max(1.0, sum_h[1])
            assert sum_d[1] < sum_d[0]
            pose_o = oracle.scan_match(clouds, [world[n] for n in onames], weights, 10.0, 1.0, start, **kw)[0]
            assert np.abs(pose_d[:3] - pose_o[:3]).max() < 1e-4
            assert min(np.abs(pose_d[3:] - pose_o[3:]).max(), np.abs(pose_d[3:] + pose_o[3:]).max()) < 1e-4
    # more clouds than the kernel takes: the host loop answers
    many = [cloud[k::9] for k in range(5)]
    pose, _ = hg.scan_match(many, [world["g_tsdf_hi"]] * 5, [1.0] * 5, 10.0, 1.0, world["truth"])
    assert np.isfinite(pose).all()


@pytest.mark.parametrize("depth,frd,window", [(7, 3, 5.0), (6, 3, 3.0), (5, 5, 2.0), (8, 2, 7.0), (4, 1, 1.5)])
def test_packed_lattice_kernel_equals_brick_kernel(oracle, hg, world, depth, frd, window):
    """The lowest-resolution lattice is scored from a residue-row copy of the level with packed 16-bit
    sums (ScoreLatticePackedKernel). HG_LATTICE_CHECK=1 makes every search recompute the lattice with
    the plain brick-table kernel and compare all sums; the searches must also return the oracle's
    candidate. Shapes: half-resolution levels (cell stride 2^(frd-1)), full-resolution only (stride
    2^(depth-1)), lattice rows of 5 ... 33+ candidates."""
    import os
    hi, lo, hist = node_inputs(world)
    hi = hi[:300]
    opts = dict(branch_and_bound_depth=depth, full_resolution_depth=frd, use_rotational_scan_matcher=False,
                min_low_resolution_score=0.3, linear_xy_search_window=window, linear_z_search_window=0.6,
                angular_search_window=0.06)
    om = oracle.Fcsm(world["o_tsdf_hi"], world["o_tsdf_lo"], hist, **opts)
    gm = hg.Fcsm(world["g_tsdf_hi"], world["g_tsdf_lo"], hist, **opts)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    rng = np.random.RandomState(depth * 10 + frd)
    os.environ["HG_LATTICE_CHECK"] = "1"
    try:
        for trial in range(2):
            node = perturbed(world["truth"], rng, 0.8, 0.04)
            r1 = gm.match(node, submap, hi, lo, hist, min_score=0.25)
            r0 = om.match(node, submap, hi, lo, hist, min_score=0.25)
            assert_same_match(r0, r1)
        if depth == 7:
            r1 = gm.match(world["truth"], submap, hi[:150], lo, hist, min_score=0.4, full=True)
            r0 = om.match(world["truth"], submap, hi[:150], lo, hist, min_score=0.4, full=True)
            assert_same_match(r0, r1)
    finally:
        os.environ.pop("HG_LATTICE_CHECK", None)


def test_packed_lattice_kernel_many_points(oracle, hg, world):
    """More than 8192 points per scan: a lane of the packed lattice kernel adds more than 256 voxels
    per candidate, so its 16-bit partial sums are flushed to 32 bits between chunks."""
    import os
    hi, lo, hist = node_inputs(world)
    rng = np.random.RandomState(4)
    big = np.concatenate([hi + rng.normal(0, 0.01, hi.shape).astype(np.float32) for _ in range(16)])[:9500]
    assert len(big) > 8192 + 1000
    # depth 5: lattice step 16 cells, 7 candidates per lattice row
    opts = dict(branch_and_bound_depth=5, full_resolution_depth=3, use_rotational_scan_matcher=False,
                min_low_resolution_score=0.3, linear_xy_search_window=5.0, linear_z_search_window=0.6,
                angular_search_window=0.02)
    om = oracle.Fcsm(world["o_tsdf_hi"], world["o_tsdf_lo"], hist, **opts)
    gm = hg.Fcsm(world["g_tsdf_hi"], world["g_tsdf_lo"], hist, **opts)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    node = world["truth"] + np.array([0.3, -0.2, 0.05, 0, 0, 0, 0])
    os.environ["HG_LATTICE_CHECK"] = "1"
    try:
        r1 = gm.match(node, submap, big, lo, hist, min_score=0.1)
    finally:
        os.environ.pop("HG_LATTICE_CHECK", None)
    r0 = om.match(node, submap, big, lo, hist, min_score=0.1)
    assert_same_match(r0, r1)
    assert r1["found"] == 1.0 and r1["num_scored"] > 49


def test_handles_are_usable_from_several_host_threads(oracle, hg, world):
    """include/hg_b200.h: handles may be used from several host threads for evaluate / match calls
    (the reference's matcher is const-callable from pool threads, fast_correlative_scan_matcher_3d.h:75-100).
    Eight threads share one matcher, one cost function and one grid; every result equals the serial one."""
    from concurrent.futures import ThreadPoolExecutor
    om, gm, hi, lo, hist = world_matchers(oracle, hg, world, "tsdf")
    rng = np.random.RandomState(77)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    nodes = [perturbed(world["truth"], rng, 1.0, 0.06) for _ in range(16)]
    cost = hg.Cost("tsdf", 0.05, world["cloud"], [world["g_tsdf_hi"]])
    poses = [perturbed(world["truth"], rng) for _ in range(16)]
    serial_match = [gm.match(n, submap, hi, lo, hist, min_score=0.2) for n in nodes]
    serial_eval = [cost.evaluate(p) for p in poses]
    serial_refine = [hg.scan_match([world["cloud"]], [world["g_tsdf_hi"]], [5.0], 10.0, 1.0, p)[0] for p in poses]

    def work(i):
        m = gm.match(nodes[i], submap, hi, lo, hist, min_score=0.2)
        r, J = cost.evaluate(poses[i])
        pose = hg.scan_match([world["cloud"]], [world["g_tsdf_hi"]], [5.0], 10.0, 1.0, poses[i])[0]
        return m, r, J, pose

    with ThreadPoolExecutor(8) as ex:
        for rounds in range(3):
            out = list(ex.map(work, range(16)))
            for i, (m, r, J, pose) in enumerate(out):
                assert_same_match(serial_match[i], m)
                assert (r == serial_eval[i][0]).all()
                assert all((a == b).all() for a, b in zip(J, serial_eval[i][1]))
                assert (pose == serial_refine[i]).all()


def test_queued_device_evaluations_keep_their_own_poses(oracle, hg, world):
    """hg_batch_evaluate with device outputs returns while its block-descriptor copy and kernel are
    still queued. Two such calls with different poses behind a busy stream must each use their own
    poses (the descriptors go through a ring of staging slots), and more calls than the ring has
    slots must still be correct."""
    import torch
    rng = np.random.RandomState(91)
    cloud = world["cloud"]
    clouds = [cloud, cloud[::2], cloud[1::2]]
    pts = np.concatenate(clouds)
    offsets = np.cumsum([0] + [len(c) for c in clouds]).astype(np.int64)
    scales = np.array([0.1, 0.2, 0.3])
    ratios = np.array([0.3, 0.6, 0.9])
    batch = hg.Batch("interpolated_multires_tsdf", pts, offsets, scales, [world["g_tsdf_hi"], world["g_tsdf_lo"]], ratios)
    i0 = np.array([0, 0, 1], np.int32)
    i1 = np.array([1, 2, 2], np.int32)
    N = len(pts)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream()
    pose_sets = [np.ascontiguousarray([perturbed(world["truth"], rng, 0.2, 0.08) for _ in range(3)]) for _ in range(7)]
    expected = [batch.evaluate(p, i0, i1) for p in pose_sets]
    outs = [[torch.zeros(N * k, dtype=torch.float64, device=dev) for k in (1, 3, 4, 3, 4)] for _ in pose_sets]
    with torch.cuda.stream(stream):
        torch.cuda._sleep(400_000_000)  # ~0.2 s of busy stream: everything below queues up behind it
        for p, o in zip(pose_sets, outs):
            batch.evaluate_raw(p.ctypes.data, len(p), i0.ctypes.data, i1.ctypes.data, True,
                               *[t.data_ptr() for t in o], stream.cuda_stream)
    stream.synchronize()
    for (r, J), o in zip(expected, outs):
        assert (o[0].cpu().numpy() == r).all()
        for a, b in zip(J, o[1:]):
            assert (b.cpu().numpy().reshape(a.shape) == a).all()


@pytest.mark.parametrize("pinned", [True, False])
def test_chunked_host_evaluate_after_update_points(oracle, hg, world, pinned):
    """Large batches with host outputs run chunk by chunk (points in, kernel, results out overlap);
    hg_batch_update_points only queues its copies. Results must equal a batch created on the new
    points, for pinned and pageable host memory, repeatedly."""
    import torch
    rng = np.random.RandomState(92)
    cloud = world["cloud"]
    blocks = max(96, 600000 // len(cloud) + 1)
    sizes = [len(cloud) - (b % 5) for b in range(blocks)]  # ragged, odd sizes
    offsets = np.cumsum([0] + sizes).astype(np.int64)
    N = int(offsets[-1])
    assert N * 15 > (1 << 20) and N > 2 * (1 << 18)  # the chunked path, several chunks
    scales = rng.uniform(0.05, 0.2, blocks)
    ratios = rng.uniform(0, 1, blocks)
    i0 = rng.randint(0, 4, blocks).astype(np.int32)
    i1 = rng.randint(0, 4, blocks).astype(np.int32)
    poses = np.ascontiguousarray([perturbed(world["truth"], rng, 0.2, 0.08) for _ in range(4)])
    grids = [world["g_tsdf_hi"], world["g_tsdf_lo"]]

    def points(seed):
        r = np.random.RandomState(seed)
        return np.concatenate([cloud[:n] + r.normal(0, 0.01, (n, 3)).astype(np.float32) for n in sizes])

    batch = hg.Batch("interpolated_multires_tsdf", points(0), offsets, scales, grids, ratios)
    for seed in (1, 2):
        new = points(seed)
        host = torch.from_numpy(new.copy())
        if pinned:
            host = host.pin_memory()
        batch.update_points(host.numpy())
        r, J = batch.evaluate(poses, i0, i1)
        fresh = hg.Batch("interpolated_multires_tsdf", new, offsets, scales, grids, ratios)
        r2, J2 = fresh.evaluate(poses, i0, i1)
        assert (r == r2).all()
        for a, b in zip(J, J2):
            assert (a == b).all()
        # and against the oracle on one block
        b = 7
        oc = oracle.Cost("interpolated_multires_tsdf", scales[b], new[offsets[b]:offsets[b + 1]],
                         [world["o_tsdf_hi"], world["o_tsdf_lo"]], ratios[b])
        ro, Jo = oc.evaluate(np.concatenate([poses[i0[b]], poses[i1[b]]]))
        assert_close(r[offsets[b]:offsets[b + 1]], ro, "residuals of block 7")


@pytest.mark.parametrize("grids", [("g_tsdf_hi", "g_tsdf_lo", "o_tsdf_hi", "o_tsdf_lo"), ("g_prob", "g_prob_lo", "o_prob", "o_prob_lo")])
def test_scan_matcher_evaluate(oracle, hg, world, grids):
    """CeresScanMatcher3D::Evaluate (ceres_scan_matcher_3d.cc:97-118): cost, residual vector (clouds, then the
    translation and rotation deltas) and the tangent-space gradient of the assembled problem at a pose."""
    rng = np.random.RandomState(55)
    cloud = world["cloud"]
    low = cloud[::9]
    for yaw_only in (False, True):
        pose = perturbed(world["truth"], rng, 0.1, 0.04)
        target = pose[:3] + rng.uniform(-0.05, 0.05, 3)
        c1, r1, g1 = hg.scan_match_evaluate([cloud, low], [world[grids[0]], world[grids[1]]], [5.0, 30.0], 10.0, 1.0,
                                            pose, target_translation=target, only_optimize_yaw=yaw_only)
        c0, r0, g0 = oracle.scan_match_evaluate([cloud, low], [world[grids[2]], world[grids[3]]], [5.0, 30.0], 10.0,
                                                1.0, pose, target_translation=target, only_optimize_yaw=yaw_only)
        assert len(r1) == len(cloud) + len(low) + 6 and len(g1) == (4 if yaw_only else 6)
        assert_close(r1, r0, "residuals")
        assert abs(c1 - c0) <= 1e-9 * max(1.0, abs(c0))
        assert np.abs(g1 - g0).max() <= 1e-7 * max(1.0, np.abs(g0).max())
        assert abs(c1 - 0.5 * float(r1 @ r1)) <= 1e-12 * max(1.0, c1)
        assert np.abs(r1[-3:]).max() < 1e-15  # the rotation delta vanishes at the pose it is anchored to


def test_device_search_equals_oracle_across_pyramid_shapes(oracle, hg, world):
    """The device-resident search prunes every subtree that scores below the greedy dive's bound, which is only
    right if a parent's score bounds its children's at every pyramid shape (half-resolution levels included;
    the fold ((c + start) >> e) - (start >> e) is designed, not proven, to keep that). Randomised check over
    depths, full-resolution depths, windows and node poses: the selected candidate equals the oracle's depth-first
    search, which does not prune that way; the host replay in strict order agrees too."""
    rng = np.random.RandomState(77)
    hi, lo, hist = node_inputs(world, n_high=400, n_low=100)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    truth_yaw = 2 * np.arctan2(world["truth"][6], world["truth"][3])
    submap_hist = oracle.rotate_histogram(hist, np.float32(truth_yaw))
    checked = found = 0
    for depth, frd in [(3, 1), (4, 1), (4, 2), (5, 2), (5, 3), (6, 2), (6, 4), (7, 3), (5, 5)]:
        for grids in (("tsdf_hi", "tsdf_lo"), ("prob", "prob_lo")):
            opts = dict(branch_and_bound_depth=depth, full_resolution_depth=frd, use_rotational_scan_matcher=False,
                        min_low_resolution_score=0.25, linear_xy_search_window=float(rng.uniform(0.8, 2.5)),
                        linear_z_search_window=float(rng.uniform(0.2, 0.7)), angular_search_window=0.1)
            om = oracle.Fcsm(world["o_" + grids[0]], world["o_" + grids[1]], submap_hist, **opts)
            gm = hg.Fcsm(world["g_" + grids[0]], world["g_" + grids[1]], submap_hist, **opts)
            gs = hg.Fcsm(world["g_" + grids[0]], world["g_" + grids[1]], submap_hist, strict_visit_order=1, **opts)
            for trial in range(2):
                node = perturbed(world["truth"], rng, 0.6, 0.05)
                min_score = float(rng.choice([0.15, 0.3]))
                r0 = om.match(node, submap, hi, lo, hist, min_score=min_score)
                r1 = gm.match(node, submap, hi, lo, hist, min_score=min_score)
                r2 = gs.match(node, submap, hi, lo, hist, min_score=min_score)
                assert_same_match(r0, r1)
                assert_same_match(r0, r2)
                checked += 1
                found += int(r0["found"])
    assert checked == 36 and found >= 18
