#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import re
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
rows = list(csv.reader(open(path)))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] == "ns" else (v * 1000 if r[ui] == "ms" else v)
    tot[name[:80]] += v
    cnt[name[:80]] += 1
T = sum(tot.values())
print("total %.0f us over %d launches" % (T, sum(cnt.values())))
print("| kernel | launches | total us | avg us | share |\n|---|---:|---:|---:|---:|")
for k, v in tot.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 22):
    print("| `%s` | %d | %.1f | %.1f | %.1f%% |" % (k, cnt[k], v, v / cnt[k], 100 * v / T))
