#!/usr/bin/env python3
"""Navigation-table digests for the golden graphs that have no full table in tests/golden/ (TEST INFRASTRUCTURE;
needs oracle/_ref/ref_cyclefinder, i.e. /root/reference at build time).

A committed golden graph is written back in MEGAHIT's on-disk format (tools/add_golden_setting.py: write_files) and the
UNMODIFIED reference SDBG API answers EdgeOutdegree / EdgeIndegree / OutgoingEdges / IncomingEdges / EdgeMultiplicity
for EVERY item (`ref_cyclefinder --nav`).  The full table of `tiny` is committed (tiny.nav.npz, 0.6 MB); for the larger
graphs only SHA-256 digests of the five arrays are kept (tests/golden/<case>.nav.sha.json), which pins every entry just
as well.  The writer is checked first: the digests of `tiny` computed this way must equal those of the committed table.

  python tools/add_golden_nav.py [case ...]     (default: repeatrich biallelic strain)
"""
import hashlib
import json
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
from add_golden_setting import write_files  # noqa: E402

REC = np.dtype([("od", "i1"), ("id", "i1"), ("m", "<u2"), ("o", "<u8", 4), ("i", "<u8", 4)])


def digests(od, idg, m, o, i_):
    """sha256 of the five arrays in fixed dtypes (int8, int8, uint16, int64[n,4], int64[n,4]; C order)"""
    arrs = dict(od=np.ascontiguousarray(od, np.int8), id=np.ascontiguousarray(idg, np.int8), m=np.ascontiguousarray(m, np.uint16),
                o=np.ascontiguousarray(o, np.int64), i=np.ascontiguousarray(i_, np.int64))
    return {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in arrs.items()}


def reference_table(name):
    exe = O.ref_binary("ref_cyclefinder")
    assert exe, "oracle/_ref/ref_cyclefinder missing: make -C oracle"
    tmp = tempfile.mkdtemp(prefix="golden_nav_")
    prefix = os.path.join(tmp, "graph")
    write_files(name, prefix)
    nav = os.path.join(tmp, "out.nav")
    subprocess.run([exe, "--nav", prefix, nav], check=True, capture_output=True)
    t = np.fromfile(nav, dtype=REC, offset=16)
    return t["od"], t["id"], t["m"], t["o"].astype(np.int64), t["i"].astype(np.int64)


def main(argv):
    z = np.load(os.path.join(G.GOLD, "tiny.nav.npz"))
    assert digests(*reference_table("tiny")) == digests(z["od"], z["id"], z["m"], z["o"], z["i"]), \
        "tiny: re-written graph files give a different navigation table"
    for name in argv or ["repeatrich", "biallelic", "strain"]:
        od, idg, m, o, i_ = reference_table(name)
        out = dict(n=int(od.size), sha256=digests(od, idg, m, o, i_),
                   source="unmodified reference SDBG API over every item (oracle/_ref/ref_cyclefinder --nav)")
        dst = os.path.join(G.GOLD, f"{name}.nav.sha.json")
        with open(dst, "w") as f:
            json.dump(out, f, indent=1)
        print("wrote", dst, out["n"], "items")


if __name__ == "__main__":
    main(sys.argv[1:])
