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


def all_reduce_max_bound(bound, dist, device="cuda"):
    """The optional mid-search exchange of a sharded MatchFullSubmap: the largest dive bound
    any shard found (hg_fcsm_full_submap_begin), one 4-byte all-reduce-max."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(bound)
    t = torch.tensor([float(bound)], dtype=torch.float32, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_best_candidate(result, dist, device="cuda"):
    """result: this rank's match result (dict of hectorgrapher_b200.capi, or None). All-gathers one
    fixed-size record per rank (hg_pack_best_record) and selects the winner with the library's rule
    (hg_select_best_record: leaf score, then the score of the leaf's lowest-resolution ancestor, then its
    generation index — the order in which the reference's search would have met the leaves). Returns
    (winning result dict or None, order_unresolved)."""
    from . import capi
    mine = capi.pack_best_record(result) if result is not None else np.zeros(capi.BEST_RECORD_DOUBLES)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        allrec = mine[None]
    else:
        t = torch.as_tensor(mine, dtype=torch.float64, device=device)
        bufs = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
        dist.all_gather(bufs, t)
        allrec = torch.stack(bufs).cpu().numpy()
    return capi.select_best_record(allrec)


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
