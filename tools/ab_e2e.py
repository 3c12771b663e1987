#!/usr/bin/env python
"""End-to-end step of bench.py (scans H2D from pinned memory + Evaluate + residuals/Jacobians D2H into
pinned memory) under different chunk counts of the host-pointer pipeline (HG_E2E_CHUNKS, read at
batch creation), plus the raw pinned copy bandwidths of the box for reference. Development aid."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import hectorgrapher_b200 as hg
    hg.set_device(0)
    w = bench.build_workload()
    N = int(w["offsets"][-1])
    g_hi = hg.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["hi"])
    g_lo = hg.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT, *w["lo"])
    poses = np.ascontiguousarray(w["poses"])
    i0, i1 = w["i0"], w["i1"]
    h_pts = torch.from_numpy(w["points"]).pin_memory()
    h_out = [torch.empty(n, dtype=torch.float64).pin_memory() for n in (N, 3 * N, 4 * N, 3 * N, 4 * N)]
    # raw copy bandwidths
    d = torch.empty(4 * N, dtype=torch.float64, device="cuda")
    for name, fn in (("D2H", lambda: h_out[2].copy_(d, non_blocking=True)), ("H2D", lambda: d.copy_(h_out[2], non_blocking=True))):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 5
        print("%s pinned 205 MB: %.2f ms = %.1f GB/s" % (name, 1e3 * dt, 4 * N * 8 / dt / 1e9), flush=True)
    for chunks in sys.argv[1:] or ["1", "4", "8", "16"]:
        os.environ["HG_E2E_CHUNKS"] = chunks
        batch = hg.Batch("interpolated_multires_tsdf", w["points"], w["offsets"], w["scales"], [g_hi, g_lo], w["ratios"])

        def step():
            batch.update_points(h_pts.numpy())
            batch.evaluate_raw(poses.ctypes.data, len(poses), i0.ctypes.data, i1.ctypes.data, False,
                               *[o.data_ptr() for o in h_out], None)
        step(); step()
        torch.cuda.synchronize()
        ts = []
        for _ in range(8):
            t0 = time.perf_counter()
            step()
            ts.append(time.perf_counter() - t0)
        print("chunks %-3s e2e median %.3f ms  min %.3f ms" % (chunks, 1e3 * np.median(ts), 1e3 * min(ts)), flush=True)
        del batch


if __name__ == "__main__":
    main()
