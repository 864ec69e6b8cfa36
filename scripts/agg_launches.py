#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel and grid size.
usage: agg_launches.py launches.csv [last_n]"""
import csv, collections, sys
lines = open(sys.argv[1]).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[start:]))
if len(sys.argv) > 2:
    rows = rows[-int(sys.argv[2]):]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    k = r['Kernel Name'].split('(')[0] + ' ' + r['Grid Size']
    agg[k][0] += 1
    agg[k][1] += float(r['Metric Value'])
tot = sum(v[1] for v in agg.values())
print(f"launches {len(rows)}  total {tot/1e6:.3f} ms")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{v[1]/1e6:9.3f} ms {100*v[1]/tot:5.1f}% n={v[0]:5d} avg={v[1]/v[0]/1e3:8.1f} us  {k}")
