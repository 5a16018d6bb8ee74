#!/usr/bin/env python3
"""Dumps a contracted ball from a golden graph (through tests/emu) and runs the DFS micro-benchmark on it.
usage: python tools/microbench/run_dfs_microbench.py [graph] [start] (on a GPU box; builds the exe with nvcc)"""
import os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in ("tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import emu_py as E
import golden_util as G

name = sys.argv[1] if len(sys.argv) > 1 else "biallelic"
start = int(sys.argv[2]) if len(sys.argv) > 2 else 59031
mn, mx = 27, 77
g = E.EmuGraph(G.pack_raw(G.load_graph_npz(name)))
g.prune()
hot, in4 = g.contracted_ball(start, mx)
st, _ = g.search_one(start, mn, mx, 2)
print("host emulation:", st)
here = os.path.dirname(os.path.abspath(__file__))
path = os.path.join(here, "ball.bin")
with open(path, "wb") as f:
    np.array([in4.size, mn, mx, 0], np.uint32).tofile(f)
    hot.tofile(f)
    in4.tofile(f)
exe = os.path.join(here, "dfs_microbench")
subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
                "-lineinfo", os.path.join(here, "dfs_microbench.cu"), "-o", exe], check=True)
subprocess.run([exe, path, "3"], check=True)
