"""Frozen vectors (tests/golden/r1_golden.npz, written by tests/golden/make_golden.py from the CPU oracle).

CPU test: the oracle as built today still produces them (drift detection).
GPU test: libhg_b200 produces them through the C-ABI, without the oracle in the loop.
Bars: integers / indices / stored voxels bit-exact (SHA-256 of the sorted voxel lists);
residuals and Jacobians rtol 1e-5 (+1e-9); refined pose 1e-4 m / 1e-4 rad."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden as mg  # noqa: E402

GOLDEN = os.path.join(HERE, "golden", "r1_golden.npz")
RTOL, ATOL = 1e-5, 1e-9


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def build_grids(lib, golden, empty_args):
    """Submap grids from the stored scans through 'lib' (oracle.pyoracle or hectorgrapher_b200)."""
    hi = lib.Grid.tsdf(0.10, 2.5, 1000.0, *empty_args["tsdf"])
    lo = lib.Grid.tsdf(0.45, 2.5, 1000.0, *empty_args["tsdf"])
    prob = lib.Grid.prob(0.10, *empty_args["prob"])
    for k in range(2):
        origin, pts = golden["scan_origin_%d" % k], golden["scan_points_%d" % k]
        hi.insert_tsdf(origin, pts, **mg.TSDF_OPTS)
        lo.insert_tsdf(origin, pts, **mg.TSDF_OPTS)
        prob.insert_occupancy(origin, pts, 0.55, 0.49, 2)
    return {"hi": hi, "lo": lo, "prob": prob}


def check_all(lib, golden, grids, exact_floats):
    """Every stage against the frozen vectors. exact_floats: residuals/Jacobians must match to the bit
    (the oracle against itself) instead of to tolerance (the GPU)."""
    for res in (0.05, 0.1, 0.45):
        got = lib.get_cell_index(res, golden["cell_points_%g" % res])
        assert (got == golden["cell_index_%g" % res]).all()
    for name in ("hi", "lo", "prob"):
        assert (mg.voxel_digest(*grids[name].export()) == golden[name + "_digest"]).all(), name
        assert grids[name].grid_size == int(golden[name + "_grid_size"])
    cloud = golden["cloud"]
    for kind in mg.KINDS:
        two = kind.startswith("interpolated")
        c = cloud[100:101] if kind.endswith("per_point") else cloud
        cost = lib.Cost(kind, 0.7 / np.sqrt(len(c)), c, [grids[n] for n in mg.KIND_GRIDS[kind]],
                        0.37 if two else 0.0)
        r, J = cost.evaluate(np.concatenate([golden["pose0"], golden["pose1"]]) if two else golden["pose0"])
        pairs = [(r, golden["cost_%s_r" % kind])] + [(J[b], golden["cost_%s_J%d" % (kind, b)]) for b in range(len(J))]
        for got, want in pairs:
            if exact_floats:
                assert (got == want).all(), kind
            else:
                assert np.allclose(got, want, rtol=RTOL, atol=ATOL), kind
        assert np.abs(golden["cost_%s_J0" % kind]).max() > 1e-4
    stack = lib.Stack(grids["hi"], 5, 3)
    for d in range(5):
        assert (mg.voxel_digest(*stack.export(d)) == golden["stack_%d_digest" % d]).all(), d
    hist = np.zeros(8, np.float32)
    matcher = lib.Fcsm(grids["hi"], grids["lo"], hist, **mg.FCSM_OPTS)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    res = matcher.match(golden["node_pose"], submap, golden["node_high"], golden["node_low"], hist, min_score=0.15)
    got = np.array([res[k] for k in mg.RESULT_KEYS], np.float64)
    assert (got == golden["match"]).all(), dict(zip(mg.RESULT_KEYS, zip(got, golden["match"])))
    start = golden["match"][2:9]
    pose = lib.scan_match([golden["node_high"], golden["node_low"]], [grids["hi"], grids["lo"]], [5.0, 30.0], 10.0,
                          1.0, start, max_num_iterations=10, use_nonmonotonic_steps=False)[0]
    want = golden["refined_pose"]
    assert np.abs(pose[:3] - want[:3]).max() < 1e-4
    dq = min(np.abs(pose[3:] - want[3:]).max(), np.abs(pose[3:] + want[3:]).max())
    assert dq < 1e-4
    assert (lib.voxel_filter(0.5, cloud) == golden["voxel_filter_kept"]).all()


def test_oracle_reproduces_golden_vectors(oracle, golden):
    grids = build_grids(oracle, golden, {"tsdf": (), "prob": ()})
    check_all(oracle, golden, grids, exact_floats=True)


@pytest.mark.gpu
def test_library_reproduces_golden_vectors(golden):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import hectorgrapher_b200 as hg
    hg.lib()
    empty = np.zeros((0, 3), np.int32)
    grids = build_grids(hg, golden, {"tsdf": (empty, [], []), "prob": (empty, [])})
    check_all(hg, golden, grids, exact_floats=False)
