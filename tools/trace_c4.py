"""C4 (MatchFullSubmap) timing with the library's phase trace (development tool): HG_TRACE=1 python tools/trace_c4.py"""
import sys
sys.path.insert(0, '.')
import bench, torch  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
w = bench.build_workload(num_scans=1, points_per_scan=1000)
r = bench.correlative_bench(hg, torch, w, 0, 1, None)
print({k: r[k] for k in ["value", "search_ms", "candidates_scored", "discrete_scans", "best", "lowest_resolution_kernel"]})
