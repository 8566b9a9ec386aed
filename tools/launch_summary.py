#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (cold-cache, serialised launches:
read the SHARES, not the absolute times)."""
import csv
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "ns")
            ms = v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0}.get(unit, 1e-6)
            rows.append((r["Kernel Name"].split("(")[0], ms))
    inst = {}
    with open(path, newline="") as f:
        for r in csv.DictReader([l for l in f if not l.startswith("==")]):
            if r.get("Metric Name") == "smsp__inst_executed.sum":
                k = r["Kernel Name"].split("(")[0]
                a = inst.setdefault(k, [0, 0.0])
                a[0] += 1
                a[1] += float(r["Metric Value"].replace(",", ""))
    agg = OrderedDict()
    for k, ms in rows:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += ms
    total = sum(v[1] for v in agg.values())
    for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{ms:10.4f} ms  {100 * ms / total:5.1f} %  x{n:<4d} avg {ms / n:8.4f} ms  {k}")
    print(f"{total:10.4f} ms total over {len(rows)} launches")
    for k, (n, v) in inst.items():
        print(f"warp instructions per launch: {v / n:14.0f}  {k}")


if __name__ == "__main__":
    main(sys.argv[1])
