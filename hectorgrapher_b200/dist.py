"""Host-side multi-GPU plumbing (one process per GPU, torch.distributed; NCCL on GPUs, gloo in the
CPU tests). The path has no data-path collective: discrete scans of a full-submap search and
(node, submap) pairs of a constraint batch are independent, so ranks only exchange small
fixed-size result records at the end."""
import numpy as np
import torch


def shard_indices(n, rank, world):
    """Indices i in [0, n) with i % world == rank — the partition hg_fcsm_match_full_submap
    (discrete scans) and hg_constraints_run_batch (jobs) apply for (shard_index, num_shards)."""
    return np.arange(rank, n, world, dtype=np.int64)


def reduce_best_candidate(record, dist, device="cuda"):
    """record = (found, score, global_scan_index, ox, oy, oz) of this rank's search. All-gathers one
    small record per rank and returns the winning record as a tuple of floats (None if no rank found a
    match): highest score; equal scores go to the lower global scan index, then the lower rank."""
    mine = torch.tensor([float(v) for v in record], dtype=torch.float64, device=device)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        allrec = mine.cpu().numpy()[None]
    else:
        bufs = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
        dist.all_gather(bufs, mine)
        allrec = torch.stack(bufs).cpu().numpy()
    found = allrec[allrec[:, 0] > 0]
    if len(found) == 0:
        return None
    order = np.lexsort((found[:, 2], -found[:, 1]))
    return tuple(float(v) for v in found[order[0]])


def gather_job_results(local, owned, dist, device="cuda"):
    """local: [num_jobs, k] array with this rank's rows filled at 'owned' indices. Returns the merged
    array on every rank (rows are disjoint across ranks, so a sum-all-reduce of zero-padded rows merges them)."""
    t = torch.zeros(local.shape, dtype=torch.float64, device=device)
    if len(owned):
        idx = torch.as_tensor(np.asarray(owned), device=device)
        t[idx] = torch.as_tensor(np.asarray(local, dtype=np.float64)[np.asarray(owned)], device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()
