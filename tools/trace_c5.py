"""C5 sample at a given number of pairs (development tool): python tools/trace_c5.py [pairs]."""
import sys
sys.path.insert(0, '.')
import bench, torch  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
w = bench.build_workload(num_scans=1, points_per_scan=1000)
pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
print(bench.constraint_bench(hg, torch, w, 0, 1, None, pairs_per_rank=pairs))
