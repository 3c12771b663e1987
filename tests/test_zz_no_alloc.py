"""hg_cost_normal_equations_static: the Gauss-Newton terms of a residual block computed without
any allocation at call time.

Some users of this library cannot call the CUDA allocator once they are running: a scan-matching
loop with a real-time budget (cudaMalloc / cudaFree stall the whole device for an unpredictable
time), or a program whose GPU memory is entirely owned by its own arena. For them the scratch memory
lives in a fixed array inside the library image (one per GPU, 984 KB, at most 1,048,576 points per
block) instead of being allocated. This file checks that entry point, through the C-ABI, against
hg_cost_normal_equations on the same inputs — twice in a row, because every call reuses the same
array and finds the previous call's contents in it. It is named so that pytest collects it last.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import hectorgrapher_b200 as hg
    from hectorgrapher_b200 import synth
    hg.lib()
    sph = synth.spheres()
    poses = [synth.corkscrew_pose(k, 3) for k in range(3)]
    hi = synth.build_tsdf(0.10, poses, 24, 256, sph)
    lo = synth.build_tsdf(0.45, poses, 24, 256, sph)
    g_hi = hg.Grid.tsdf(0.10, 2.5, 1000.0, *hi)
    g_lo = hg.Grid.tsdf(0.45, 2.5, 1000.0, *lo)
    origin, yaw = poses[1]
    cloud = synth.scan(origin, yaw, 16, 300, sph, np.random.RandomState(5))
    truth = np.concatenate([origin, synth.quat_from_yaw(yaw)])
    return hg, synth, g_hi, g_lo, cloud, truth


def _params(synth, truth, rng, two):
    def one():
        return np.concatenate([truth[:3] + rng.uniform(-0.1, 0.1, 3),
                               synth.quat_mul(truth[3:], synth.small_rotation(rng, 0.05))])
    return np.concatenate([one(), one()]) if two else one()


def _close(a, b):
    # same tolerance as test_normal_equations in test_gpu_parity.py
    a, b = np.asarray(a), np.asarray(b)
    return np.abs(a - b).max() <= 1e-9 * max(1.0, np.abs(b).max())


@pytest.mark.parametrize("kind,two", [("tsdf", False), ("interpolated_multires_tsdf", True)])
def test_static_scratch_matches_allocating_entry_point(setup, kind, two):
    hg, synth, g_hi, g_lo, cloud, truth = setup
    grids = [g_hi, g_lo] if "multires" in kind else [g_hi]
    rng = np.random.RandomState(17)
    # two different blocks, so that the second call finds the first one's leftovers in the array
    big = hg.Cost(kind, 0.05, cloud, grids, 0.35)
    small = hg.Cost(kind, 0.11, cloud[: len(cloud) // 3], grids, 0.8)
    for call in range(2):
        for c in (big, small):
            p = _params(synth, truth, rng, two)
            want = c.normal_equations(p)
            got = c.normal_equations(p, static_scratch=True)
            assert _close(got[0], want[0]) and _close(got[1], want[1]) and _close(got[2], want[2])
            assert np.abs(want[0]).max() > 0
            again = c.normal_equations(p, static_scratch=True)  # two calls in a row, same inputs
            assert (again[0] == got[0]).all() and (again[1] == got[1]).all() and again[2] == got[2]


def test_static_scratch_rejects_blocks_beyond_its_limit(setup):
    hg, synth, g_hi, g_lo, cloud, truth = setup
    reps = (1 << 20) // len(cloud) + 1
    too_many = np.tile(cloud, (reps, 1))
    assert len(too_many) > (1 << 20)
    c = hg.Cost("tsdf", 0.01, too_many, [g_hi])
    with pytest.raises(hg.HgError) as err:
        c.normal_equations(truth, static_scratch=True)
    assert "1048576" in str(err.value) and "hg error 3" in str(err.value)
    # the allocating entry point still handles it
    JtJ, Jtr, cost = c.normal_equations(truth)
    assert np.isfinite(JtJ).all() and cost >= 0
    # exactly at the limit is accepted
    at_limit = hg.Cost("tsdf", 0.01, too_many[: 1 << 20], [g_hi])
    a = at_limit.normal_equations(truth, static_scratch=True)
    b = at_limit.normal_equations(truth)
    assert _close(a[0], b[0]) and _close(a[2], b[2])


def test_static_scratch_calls_leave_device_memory_alone(setup):
    """The reason this entry point exists: after its first call on a device (which looks up the address of
    the static array) a call must not take or return device memory. cudaMemGetInfo's free byte count stays
    where it was over 100 calls — an allocation by the library or by the CUDA runtime would move it."""
    import torch
    hg, synth, g_hi, g_lo, cloud, truth = setup
    c = hg.Cost("interpolated_multires_tsdf", 0.05, cloud, [g_hi, g_lo], 0.35)
    rng = np.random.RandomState(23)
    p = _params(synth, truth, rng, True)
    first = c.normal_equations(p, static_scratch=True)  # set-up call
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(100):
        got = c.normal_equations(p, static_scratch=True)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before
    assert (got[0] == first[0]).all() and got[2] == first[2]
