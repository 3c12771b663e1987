"""Freezes the CPU oracle's answers on the configurations bench.py times, so that bench.py and the GPU
tests can assert the selected candidates without the oracle in the loop:
  C4  MatchFullSubmap on bench.correlative_inputs, rotational prefilter on (configured) and off (full sweep)
  C3  Match on the first offsets of bench.local_search_offsets, N = 2048
Run from the repository root: python tests/golden/make_c4_golden.py  (the full sweep takes minutes on one thread)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import pyoracle as o  # noqa: E402

KEYS = ["found", "score", "scan_index", "ox", "oy", "oz", "low_resolution_score", "rotational_score", "num_scans"]


def main():
    w = bench.build_workload(num_scans=1, points_per_scan=1000)
    hi = o.Grid.tsdf(bench.HI_RES, bench.REL_TRUNC, bench.MAX_WEIGHT)
    hi.set_raw(*w["hi"])
    lo = o.Grid.tsdf(bench.LO_RES, bench.REL_TRUNC, bench.MAX_WEIGHT)
    lo.set_raw(*w["lo"])
    out = {}
    only = sys.argv[1:] or ["configured", "c3", "full_sweep"]
    path = os.path.join(ROOT, "tests", "golden", "c4_golden.json")
    if os.path.exists(path):
        out = json.load(open(path))
    for name in only:
        t0 = time.time()
        if name == "c3":
            ci = bench.correlative_inputs(w, rotational=True)
            m = o.Fcsm(hi, lo, ci["submap_hist"], **ci["opts"])
            rows = []
            for node in bench.local_search_offsets(ci, 100):
                r = m.match(node, ci["submap"], ci["high"], ci["low"], ci["hist"], min_score=ci["min_score"])
                rows.append([float(r[k]) for k in KEYS])
            out["c3"] = {"keys": KEYS, "rows": rows}
        else:
            ci = bench.correlative_inputs(w, rotational=(name == "configured"))
            m = o.Fcsm(hi, lo, ci["submap_hist"], **ci["opts"])
            r = m.match(ci["node"], ci["submap"], ci["high"], ci["low"], ci["hist"], min_score=ci["min_score"], full=True)
            out[name] = {k: float(r[k]) for k in KEYS}
        print(name, "%.1f s" % (time.time() - t0), out[name] if name != "c3" else "%d rows" % len(out["c3"]["rows"]), flush=True)
        with open(path, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
