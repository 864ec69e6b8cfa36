#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv`
launch list per kernel and grid size: total / average time and, when present, DRAM bytes per launch.
usage: agg_launches.py launches.csv [last_n_launches]"""
import collections
import csv
import sys

import gzip
lines = (gzip.open(sys.argv[1], "rt") if sys.argv[1].endswith(".gz") else open(sys.argv[1])).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[start:]))
ids = sorted({int(r["ID"]) for r in rows})
if len(sys.argv) > 2:
    keep = set(ids[-int(sys.argv[2]):])
    rows = [r for r in rows if int(r["ID"]) in keep]
    ids = sorted(keep)


def val(r):
    v = float(r["Metric Value"].replace(",", ""))
    u = r.get("Metric Unit", "")
    scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    return v * scale


agg = collections.defaultdict(lambda: [0, 0.0, 0.0])      # launches, ns, dram bytes
for r in rows:
    k = r["Kernel Name"].split("(")[0] + " " + r["Grid Size"]
    m = r.get("Metric Name", "gpu__time_duration.sum")
    if m == "gpu__time_duration.sum":
        agg[k][0] += 1
        agg[k][1] += val(r)
    elif m.startswith("dram__bytes"):
        agg[k][2] += val(r)
tot = sum(v[1] for v in agg.values())
totb = sum(v[2] for v in agg.values())
print(f"launches {len(ids)}  total {tot / 1e6:.3f} ms  dram {totb / 1e9:.3f} GB")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    extra = f" dram/launch={v[2] / v[0] / 1e6:9.2f} MB  {v[2] / max(v[1], 1):6.2f} B/ns" if totb else ""
    print(f"{v[1] / 1e6:9.3f} ms {100 * v[1] / tot:5.1f}% n={v[0]:5d} avg={v[1] / v[0] / 1e3:8.1f} us{extra}  {k}")
