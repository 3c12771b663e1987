"""C4 (MatchFullSubmap) timing, device-resident search vs host replay, with the library's phase trace
(development tool): HG_TRACE=1 python tools/trace_c4.py"""
import os
import sys
import time
sys.path.insert(0, '.')
import numpy as np  # noqa: E402
import bench, torch  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
w = bench.build_workload(num_scans=1, points_per_scan=1000)
g_hi = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"])
g_lo = hg.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["lo"])
ci = bench.correlative_inputs(w, rotational="--configured" in sys.argv)
high, low, hist, node, submap = ci["high"], ci["low"], ci["hist"], ci["node"], ci["submap"]
m = hg.Fcsm(g_hi, g_lo, ci["submap_hist"], **ci["opts"])
for mode in (("device",) if "--device-only" in sys.argv else ("device", "host")):
    if mode == "host":
        os.environ["HG_SEARCH_HOST_REPLAY"] = "1"
    else:
        os.environ.pop("HG_SEARCH_HOST_REPLAY", None)
    for shards in (1, 8):
        ts = []
        for _ in range(6):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = m.match(node, submap, high, low, hist, min_score=ci["min_score"], full=True, shard_index=0, num_shards=shards)
            ts.append(1e3 * (time.perf_counter() - t0))
        print(mode, "shards", shards, "ms: first %.2f median %.2f min %.2f" % (ts[0], np.median(ts[1:]), min(ts)),
              {k: r[k] for k in ["found", "score", "scan_index", "ox", "oy", "oz", "num_scored", "num_scans",
                                 "lattice_kernel_ms", "dive_bound", "num_launches"]}, flush=True)
