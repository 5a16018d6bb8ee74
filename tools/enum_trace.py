#!/usr/bin/env python3
"""Critical-path view of a MCAAT_ENUM_TRACE file: the largest gaps between consecutive commits."""
import sys
import numpy as np
d = np.genfromtxt(sys.argv[1], delimiter=',', skip_header=1, dtype=np.int64)
t, fate, status, steps, ncyc, us_s, us_c = d.T[:7]
acc = fate == 1
print(f"tickets {len(d)}  accepted {acc.sum()}  last commit {us_c.max()} us  steps(accepted) sum {steps[acc].sum()} max {steps[acc].max()}")
keep = us_c > 0  # tickets consumed by a neighbour's hand-over carry no timestamps
d = d[keep]
t, fate, status, steps, ncyc, us_s, us_c = d.T[:7]
gap = np.diff(np.concatenate([[0], us_c]))
big = np.argsort(-gap)[:int(sys.argv[2]) if len(sys.argv) > 2 else 15]
print("ticket gap_us fate status steps n_cycles us_searched us_committed [n_ball us_build]")
for i in big:
    print(t[i], gap[i], fate[i], status[i], steps[i], ncyc[i], us_s[i], us_c[i], *d[i][7:])
print("top gaps sum", gap[big].sum(), "of", gap.sum())
