"""Multi-process (gloo, world size 2, CPU) tests of the host-side multi-GPU logic: the shard
partitions and the best-candidate reduce that bench.py / a multi-GPU caller run over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hectorgrapher_b200 import dist as hgdist


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. discrete scans are dealt round-robin; every scan is owned exactly once
        S = 11
        mine = hgdist.shard_indices(S, rank, world)
        owned = [torch.zeros(S, dtype=torch.int64) for _ in range(world)]
        flags = torch.zeros(S, dtype=torch.int64)
        flags[mine] = 1
        dist.all_gather(owned, flags)
        assert (torch.stack(owned).sum(0) == 1).all()
        # 2. best-candidate reduce (hg_pack_best_record / hg_select_best_record around an all-gather):
        # higher leaf score wins; equal leaf scores -> the leaf whose lowest-resolution ancestor scores
        # higher (the reference visits that subtree first, fast_correlative_scan_matcher_3d.cc:384,
        # 437-440, and std::max keeps the earlier, :465-466); equal ancestors -> earlier generation
        # index, flagged as unresolved.
        def rec(found, score, scan, off, anc_score=0.9, anc_index=0):
            return dict(found=found, score=score, scan_index=scan, ox=off[0], oy=off[1], oz=off[2], tx=0.1 * scan,
                        ty=0., tz=0., qw=1., qx=0., qy=0., qz=0., rotational_score=1., low_resolution_score=0.5,
                        ancestor_score=anc_score, ancestor_index=anc_index, num_scored=10)
        cases = [
            # (per-rank result) -> (expected winner rank, order unresolved)
            ([rec(1, 0.61, 7, (1, 2, 3)), rec(1, 0.75, 4, (-1, 0, 2))], 1, False),
            ([rec(1, 0.80, 2, (0, 0, 0), 0.95, 40), rec(1, 0.80, 9, (5, 5, 5), 0.97, 900)], 1, False),
            ([rec(1, 0.80, 2, (0, 0, 0), 0.97, 40), rec(1, 0.80, 9, (5, 5, 5), 0.95, 900)], 0, False),
            ([rec(1, 0.80, 9, (0, 0, 0), 0.95, 900), rec(1, 0.80, 2, (5, 5, 5), 0.95, 40)], 1, True),
            ([rec(0, 0.0, 0, (0, 0, 0)), rec(1, 0.31, 3, (1, 1, 1))], 1, False),
            ([rec(1, 0.5, 1, (0, 0, 0)), None], 0, False),
        ]
        for records, winner, unresolved in cases:
            best, flag = hgdist.reduce_best_candidate(records[rank], dist, device="cpu")
            want = records[winner]
            assert best is not None and flag == unresolved, (best, flag)
            for k in ("score", "scan_index", "ox", "oy", "oz", "tx", "ancestor_score", "ancestor_index"):
                assert float(best[k]) == float(np.float32(want[k]) if k in ("score", "ancestor_score") else want[k]), (k, best, want)
        none, _ = hgdist.reduce_best_candidate(rec(0, 0.0, 0, (0, 0, 0)), dist, device="cpu")
        assert none is None
        assert hgdist.all_reduce_max_bound(0.25 + 0.5 * rank, dist, device="cpu") == 0.75
        # a matcher that searches on the host has no device-resident bound: nothing is exchanged, nobody waits
        assert hgdist.all_reduce_max_device_bound(0, 0, dist) is None
        # 3. constraint jobs: partition by index, gather fixed-size records, identical to the serial list
        jobs = np.arange(9)
        serial = np.stack([jobs, jobs * 0.5 + 1.0], 1)
        local = np.zeros_like(serial)
        for j in hgdist.shard_indices(len(jobs), rank, world):
            local[j] = serial[j]
        merged = hgdist.gather_job_results(local, hgdist.shard_indices(len(jobs), rank, world), dist, device="cpu")
        assert np.array_equal(merged, serial)
        with open(os.path.join(out_dir, "ok%d" % rank), "w") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_gloo_world_size_2(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")


def test_shard_indices_cover():
    for n in [0, 1, 7, 64]:
        for world in [1, 2, 3, 8]:
            all_idx = np.concatenate([hgdist.shard_indices(n, r, world) for r in range(world)])
            assert sorted(all_idx.tolist()) == list(range(n))
