#!/usr/bin/env python
"""A/B timing of the C2 cost kernel under environment switches (development aid).

  python tools/ab_cost.py [--once] "HG_TILE_OUT=1024" "HG_TILE_OUT=512" ...

Each argument is a space-separated list of VAR=VALUE settings applied before the timed calls
(the library reads them per call). Same protocol as bench.py's M1: device-resident outputs, CUDA
events on the launching stream, 256 MiB L2 flush between timed steps. --once: one warm-up and one
timed call per configuration (for ncu captures)."""
import os
import sys

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
    make_batch = lambda: hg.Batch("interpolated_multires_tsdf", w["points"], w["offsets"], w["scales"], [g_hi, g_lo],
                                  w["ratios"])
    batch = make_batch()
    dev = torch.device("cuda", 0)
    outs = [torch.empty(N * k, dtype=torch.float64, device=dev) for k in (1, 3, 4, 3, 4)]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    poses = np.ascontiguousarray(w["poses"])
    i0, i1 = w["i0"], w["i1"]
    stream = torch.cuda.current_stream()

    def step():
        nonlocal batch
        batch.evaluate_raw(poses.ctypes.data, len(poses), i0.ctypes.data, i1.ctypes.data, True,
                           *[o.data_ptr() for o in outs], stream.cuda_stream)

    usage = batch.level_usage(poses, i0, i1)
    print("points answered by level (fine, coarse, none):", usage.tolist(), "fractions",
          (usage / usage.sum()).round(4).tolist(), flush=True)
    reference = None
    once = "--once" in sys.argv
    configs = [a for a in sys.argv[1:] if a != "--once"] or [""]
    for rnd in range(1 if once else 2):
        for cfg in configs:
            for kv in cfg.split():
                k, v = kv.split("=")
                os.environ[k] = v
            batch = make_batch()  # HG_TILE_OUT is read when the batch is created
            for o in outs:
                o.zero_()
            for _ in range(1 if once else 3):
                step()
            torch.cuda.synchronize()
            times = []
            for _ in range(1 if once else 20):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                step()
                e1.record(stream)
                e1.synchronize()
                times.append(e0.elapsed_time(e1))
            got = [o.clone() for o in outs]
            if reference is None:
                reference = got
            same = all(bool((a == b).all()) for a, b in zip(got, reference))
            ms = float(np.median(times))
            print("%-40s median %.4f ms  min %.4f  -> %.3f of 6540.8 GB/s at 164 B/pt   identical to first config: %s" % (
                cfg, ms, min(times), N * 164 / (ms * 1e-3) / 1e9 / 6540.8, same), flush=True)


if __name__ == "__main__":
    main()
