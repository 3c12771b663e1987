#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv --log-file X` launch list:
python tools/launch_table.py X [first_id]  (only launches with ID >= first_id)."""
import csv
import re
import sys
from collections import OrderedDict


def main(path, first=0):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    for r in rd:
        d = dict(zip(hdr, r))
        if d.get("Metric Name") != "gpu__time_duration.sum":
            continue
        if int(d["ID"]) < first:
            continue
        v = float(d["Metric Value"].replace(",", ""))
        unit = d["Metric Unit"]
        us = v / 1000.0 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000.0)
        name = re.sub(r"\(.*", "", d["Kernel Name"])
        name = re.sub(r"^void (hg::)?(\(anonymous namespace\)::)?", "", name)
        rows.append((name, us))
    agg = OrderedDict()
    for n, us in rows:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    print("| kernel | launches | total us | avg us | share |")
    print("|---|---|---|---|---|")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.1f | %.1f | %.1f %% |" % (n[:70], c, t, t / c, 100 * t / total))
    print("total %.1f us over %d launches" % (total, len(rows)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0)
