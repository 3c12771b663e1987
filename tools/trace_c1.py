"""C1 section of bench.py alone (development tool)."""
import sys
sys.path.insert(0, '.')
import bench, torch  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
w = bench.build_workload(num_scans=1, points_per_scan=1000)
print(bench.c1_bench(hg, torch, w))
