"""C5 at a given size (development tool): python tools/trace_c5.py [submaps] [nodes_per_submap]."""
import sys
sys.path.insert(0, '.')
import bench, torch  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
w = bench.build_workload(num_scans=1, points_per_scan=1000)
submaps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nodes = int(sys.argv[2]) if len(sys.argv) > 2 else 6
print(bench.constraint_bench(hg, torch, w, 0, 1, None, num_submaps=submaps, nodes_per_submap=nodes))
