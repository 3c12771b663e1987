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
        # 2. best-candidate reduce: higher score wins; equal scores -> lower global scan index
        cases = [
            # (per-rank record: found, score, scan, ox, oy, oz) -> expected winner rank
            ([(1, 0.61, 7, 1, 2, 3), (1, 0.75, 4, -1, 0, 2)], 1),
            ([(1, 0.80, 9, 0, 0, 0), (1, 0.80, 2, 5, 5, 5)], 1),
            ([(0, 0.0, 0, 0, 0, 0), (1, 0.31, 3, 1, 1, 1)], 1),
            ([(1, 0.5, 1, 0, 0, 0), (0, 0.0, 0, 0, 0, 0)], 0),
        ]
        for records, winner in cases:
            best = hgdist.reduce_best_candidate(records[rank], dist, device="cpu")
            assert best is not None
            assert tuple(best[1:]) == tuple(float(v) for v in records[winner][1:]), (best, records[winner])
        none = hgdist.reduce_best_candidate((0, 0.0, 0, 0, 0, 0), dist, device="cpu")
        assert none is None
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
