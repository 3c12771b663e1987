"""Time of FastCorrelativeScanMatcher3D construction (pyramid build) on the C4 submap (development tool)."""
import sys, time
sys.path.insert(0, '.')
import numpy as np
import bench, torch  # noqa: E402
import hectorgrapher_b200 as hg  # noqa: E402
from oracle import pyoracle
pyoracle.build()
w = bench.build_workload(num_scans=1, points_per_scan=1000)
g_hi = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"])
g_lo = hg.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["lo"])
hist = np.zeros(8, np.float32)
opts = dict(branch_and_bound_depth=7, full_resolution_depth=3, use_rotational_scan_matcher=False)
hg.Fcsm(g_hi, g_lo, hist, **opts)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    t0 = time.perf_counter(); m = hg.Fcsm(g_hi, g_lo, hist, **opts); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
print("GPU matcher construction (7-level pyramid, %d bricks): %.2f ms" % (g_hi.num_bricks, 1e3 * min(ts)))
t0 = time.perf_counter(); g = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"]); torch.cuda.synchronize()
print("grid upload (%d voxels): %.2f ms" % (len(w["hi"][0]), 1e3 * (time.perf_counter() - t0)))
o_hi = pyoracle.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT); o_hi.set_raw(*w["hi"])
o_lo = pyoracle.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT); o_lo.set_raw(*w["lo"])
t0 = time.perf_counter(); om = pyoracle.Fcsm(o_hi, o_lo, hist, **opts); print("CPU (oracle) matcher construction: %.1f ms" % (1e3 * (time.perf_counter() - t0)))
