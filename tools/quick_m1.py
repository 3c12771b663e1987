"""Quick M1 timing (no CPU baseline, no correlative part)."""
import sys, time, numpy as np
sys.path.insert(0, '.')
import bench, torch
import hectorgrapher_b200 as hg
w = bench.build_workload()
N = int(w["offsets"][-1])
g_hi = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"])
g_lo = hg.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["lo"])
batch = hg.Batch("interpolated_multires_tsdf", w["points"], w["offsets"], w["scales"], [g_hi, g_lo], w["ratios"])
dev = torch.device("cuda", 0)
outs = [torch.empty(N * k, dtype=torch.float64, device=dev) for k in (1, 3, 4, 3, 4)]
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
poses = np.ascontiguousarray(w["poses"]); i0, i1 = w["i0"], w["i1"]
stream = torch.cuda.current_stream()
def step():
    batch.evaluate_raw(poses.ctypes.data, len(poses), i0.ctypes.data, i1.ctypes.data, True, *[o.data_ptr() for o in outs], stream.cuda_stream)
for _ in range(5): step()
torch.cuda.synchronize()
ts = []
for _ in range(20):
    flush.fill_(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); step(); e1.record(stream); e1.synchronize(); ts.append(e0.elapsed_time(e1))
print("M1 ms/step: median %.4f min %.4f  -> %.2f Gpts/s, chk %.6e" % (np.median(ts), np.min(ts), N / np.median(ts) / 1e6, float(outs[0].abs().sum())))
t0=time.perf_counter()
for _ in range(5): batch.normal_equations(poses, i0, i1)
print("NE ms/step %.4f" % ((time.perf_counter()-t0)/5*1e3))
