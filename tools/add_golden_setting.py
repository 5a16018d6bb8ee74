#!/usr/bin/env python3
"""Adds reference dumps for further parameter settings to tests/golden/ WITHOUT rebuilding the graphs (TEST
INFRASTRUCTURE; needs oracle/_ref/ref_cyclefinder, i.e. /root/reference at build time).

A committed golden graph (tests/golden/<case>.graph.npz: the arrays the reference's own SDBGBuild produced) is written
back in MEGAHIT's on-disk format — tip labels are not kept in the npz and are written as zeros: CycleFinder never
reads them (only IndexBinarySearch does) — and the UNMODIFIED reference CycleFinder runs on it at threads=1.
The writer is checked first: the dump of an already committed setting must come out byte-identical.

  python tools/add_golden_setting.py strain 20 27 200 [strain 20 27 1000 ...]
"""
import gzip
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("oracle", "tests", "tools"):
    sys.path.insert(0, os.path.join(ROOT, p))
import golden_util as G  # noqa: E402
import oracle_py as O  # noqa: E402
import synth_sdbg as S  # noqa: E402


def write_files(name, prefix):
    raw = G.load_graph_npz(name)
    n = raw["w"].size
    f = np.asarray(raw["f"], np.int64)
    cls = np.searchsorted(f[1:6], np.arange(n), side="right")  # 1..5: f[c] <= i < f[c+1]
    cls = np.clip(cls, 1, 4)
    g = dict(k=int(raw["k"]), n=n, w_unpacked=raw["w"].astype(np.uint8), last_unpacked=raw["last"].astype(np.uint8),
             tip_unpacked=raw["tip"].astype(np.uint8), mult=raw["mult"].astype(np.uint16),
             tip_label_chars=np.zeros(int(raw["tip"].sum()), np.uint64), first8=((cls - 1) * 16384).astype(np.int64))
    S.write_megahit_files(g, prefix)


def main(argv):
    exe = O.ref_binary("ref_cyclefinder")
    assert exe, "oracle/_ref/ref_cyclefinder missing: make -C oracle"
    jobs = [(argv[i], int(argv[i + 1]), int(argv[i + 2]), int(argv[i + 3])) for i in range(0, len(argv), 4)]
    for name in sorted({j[0] for j in jobs}):
        tmp = tempfile.mkdtemp(prefix="golden_")
        prefix = os.path.join(tmp, "graph")
        write_files(name, prefix)
        have = [c for c in G.cases() if c[0] == name][0]
        chk = os.path.join(tmp, "check.dump")
        subprocess.run([exe, prefix, chk, "1", str(have[1]), str(have[2]), str(have[3])], check=True, capture_output=True)
        assert open(chk, "rb").read() == gzip.open(have[4], "rb").read(), f"{name}: re-written graph files give a different dump"
        for _, thr, mn, mx in [j for j in jobs if j[0] == name]:
            out = os.path.join(tmp, "new.dump")
            subprocess.run([exe, prefix, out, "1", str(thr), str(mn), str(mx)], check=True, capture_output=True)
            dst = os.path.join(G.GOLD, f"{name}.t{thr}_{mn}_{mx}.dump.gz")
            with open(out, "rb") as fi, gzip.GzipFile(dst, "wb", mtime=0) as fo:
                fo.write(fi.read())
            print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main(sys.argv[1:])
