#!/usr/bin/env python3
"""Dumps the navigation cache of a golden graph (through tests/emu) and runs the ball-extraction micro-benchmark on it.
usage: python tools/microbench/run_build_microbench.py [graph] [start] (on a GPU box; builds the exe with nvcc)"""
import os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in ("tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import emu_py as E
import golden_util as G

name = sys.argv[1] if len(sys.argv) > 1 else "biallelic"
start = int(sys.argv[2]) if len(sys.argv) > 2 else 59031
raw = G.load_graph_npz(name)
g = E.EmuGraph(G.pack_raw(raw))
g.prune()
here = os.path.dirname(os.path.abspath(__file__))
g.adj().tofile(os.path.join(here, "adj.bin"))
np.ascontiguousarray(raw["mult"], np.uint16).tofile(os.path.join(here, "mult.bin"))
exe = os.path.join(here, "build_microbench")
subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
                "-lineinfo", os.path.join(here, "build_microbench.cu"), "-o", exe], check=True)
subprocess.run([exe, os.path.join(here, "adj.bin"), os.path.join(here, "mult.bin"), str(start), "77", "3"], check=True)
