import sys, time, numpy as np
sys.path.insert(0, '.')
import bench, torch
import hectorgrapher_b200 as hg
w = bench.build_workload(num_scans=1, points_per_scan=1000)
t=time.time()
r = bench.correlative_bench(hg, torch, w, 0, 1, None)
print({k: r[k] for k in ["value","search_ms","candidates_scored","discrete_scans","best"]})
