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


_BUFFERS = {}


def _buffers(dist, device):
    """Persistent staging for the two small exchanges of a sharded search: the collectives run on
    preallocated device tensors fed from / drained into pinned host tensors, so that an exchange is two
    tiny copies and one collective instead of fresh allocations, list-of-tensor gathers and pageable copies."""
    world = dist.get_world_size()
    key = (str(device), world)
    if key not in _BUFFERS:
        from . import capi
        n = capi.BEST_RECORD_DOUBLES
        pin = str(device) != "cpu" and torch.cuda.is_available()
        host = lambda shape, dtype: (torch.zeros(shape, dtype=dtype).pin_memory() if pin else torch.zeros(shape, dtype=dtype))
        _BUFFERS[key] = dict(
            bound_d=torch.zeros(1, dtype=torch.float32, device=device), bound_h=host(1, torch.float32),
            rec_d=torch.zeros(n, dtype=torch.float64, device=device), rec_h=host(n, torch.float64),
            all_d=torch.zeros(world * n, dtype=torch.float64, device=device), all_h=host(world * n, torch.float64))
    return _BUFFERS[key]


def all_reduce_max_bound(bound, dist, device="cuda"):
    """The optional mid-search exchange of a sharded MatchFullSubmap: the largest dive bound
    any shard found (hg_fcsm_full_submap_begin), one 4-byte all-reduce-max."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(bound)
    b = _buffers(dist, device)
    b["bound_h"][0] = float(bound)
    b["bound_d"].copy_(b["bound_h"], non_blocking=True)
    dist.all_reduce(b["bound_d"], op=dist.ReduceOp.MAX)
    b["bound_h"].copy_(b["bound_d"], non_blocking=True)
    if str(device) != "cpu":
        torch.cuda.current_stream().synchronize()
    return float(b["bound_h"][0])


class _DeviceFloat:
    """A float in device memory that somebody else owns, as seen by torch (zero copy)."""

    def __init__(self, address):
        self.__cuda_array_interface__ = {"shape": (1,), "typestr": "<f4", "data": (int(address), False),
                                         "version": 2, "strides": None}


def all_reduce_max_device_bound(d_bound, stream, dist):
    """The same exchange with the bound left where hg_fcsm_full_submap_begin_async put it: the
    all-reduce-max is queued behind phase A on the search's own stream and phase B is queued behind it,
    so the host never waits between the two halves of a sharded search."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return
    if not d_bound:
        # the matcher searches on the host (strict_visit_order / HG_SEARCH_HOST_REPLAY): there is no
        # device-resident bound, every rank's finish() searches under its own — nothing to exchange
        return
    t = torch.as_tensor(_DeviceFloat(d_bound), device=torch.device("cuda", torch.cuda.current_device()))
    with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
        dist.all_reduce(t, op=dist.ReduceOp.MAX)


def reduce_best_candidate(result, dist, device="cuda"):
    """result: this rank's match result (dict of hectorgrapher_b200.capi, or None). All-gathers one
    fixed-size record per rank (hg_pack_best_record) and selects the winner with the library's rule
    (hg_select_best_record: leaf score, then the score of the leaf's lowest-resolution ancestor, then its
    generation index — the order in which the reference's search would have met the leaves). Returns
    (winning result dict or None, order_unresolved)."""
    from . import capi
    mine = capi.pack_best_record(result) if result is not None else np.zeros(capi.BEST_RECORD_DOUBLES)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return capi.select_best_record(mine[None])
    b = _buffers(dist, device)
    b["rec_h"].copy_(torch.from_numpy(mine))
    b["rec_d"].copy_(b["rec_h"], non_blocking=True)
    dist.all_gather_into_tensor(b["all_d"], b["rec_d"])
    b["all_h"].copy_(b["all_d"], non_blocking=True)
    if str(device) != "cpu":
        torch.cuda.current_stream().synchronize()
    return capi.select_best_record(b["all_h"].numpy().reshape(dist.get_world_size(), -1))


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
