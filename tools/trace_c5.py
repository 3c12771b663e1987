import sys, time, numpy as np
sys.path.insert(0, '.')
import bench, torch
import hectorgrapher_b200 as hg
w = bench.build_workload(num_scans=1, points_per_scan=1000)
print(bench.constraint_bench(hg, torch, w, 0, 1, None, pairs_per_rank=6))
