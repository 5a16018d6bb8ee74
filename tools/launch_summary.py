#!/usr/bin/env python3
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections
import csv
import sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]
ki, vi, ui = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    v = float(r[vi].replace(',', '')) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(r[ui], 1e-6)
    k = r[ki].split('(')[0]
    tot[k] += v
    cnt[k] += 1
T = sum(tot.values())
print(f"{'kernel':28s} {'launches':>8s} {'total ms':>10s} {'share':>7s}")
for k, v in tot.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 20):
    print(f"{k:28s} {cnt[k]:8d} {v:10.2f} {100 * v / T:6.1f}%")
print(f"{'all':28s} {sum(cnt.values()):8d} {T:10.2f}")
