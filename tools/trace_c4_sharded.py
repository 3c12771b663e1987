"""Where the time of a sharded MatchFullSubmap goes on each rank (development tool):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/trace_c4_sharded.py
Wall-clock per step of bench.sharded_full_submap: begin (phase A + bound read-back), all-reduce-max of the bound,
finish (phase B + result read-back), all-gather of the best records + selection."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
from hectorgrapher_b200 import dist as hgdist  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
hg.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
d = dist if world > 1 else None
w = bench.build_workload(num_scans=1, points_per_scan=1000)
g_hi = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"])
g_lo = hg.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["lo"])
ci = bench.correlative_inputs(w, rotational=False)
m = hg.Fcsm(g_hi, g_lo, ci["submap_hist"], **ci["opts"])
rows = []
for it in range(12):
    if d is not None:
        d.barrier()
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    if "--host-bound" in sys.argv:
        handle, bound = m.full_submap_begin(ci["node"][3:], ci["submap"][3:], ci["high"], ci["low"], ci["hist"],
                                            min_score=ci["min_score"], shard_index=rank, num_shards=world)
        t.append(time.perf_counter())
        bound = hgdist.all_reduce_max_bound(bound, d)
    else:
        handle, d_bound, stream = m.full_submap_begin_async(ci["node"][3:], ci["submap"][3:], ci["high"], ci["low"],
                                                            ci["hist"], min_score=ci["min_score"], shard_index=rank,
                                                            num_shards=world)
        t.append(time.perf_counter())
        hgdist.all_reduce_max_device_bound(d_bound, stream, d)
        bound = 0.0
    t.append(time.perf_counter())
    mine = m.full_submap_finish(handle, bound)
    t.append(time.perf_counter())
    best, unresolved = hgdist.reduce_best_candidate(mine, d)
    t.append(time.perf_counter())
    if it >= 2:
        rows.append(np.diff(t) * 1e6)
med = np.median(np.array(rows), axis=0)
print("rank %d of %d: begin %.0f us, all-reduce-max %.0f us, finish %.0f us, all-gather + select %.0f us, total %.0f us; "
      "best scan %d score %.6f" % (rank, world, med[0], med[1], med[2], med[3], med.sum(), int(best["scan_index"]),
                                   best["score"]), flush=True)
if world > 1:
    dist.destroy_process_group()
