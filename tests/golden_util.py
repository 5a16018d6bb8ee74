"""Helpers shared by the parity tests: golden graphs (raw SDBG arrays made by the reference's own
builder) and the reference dumps that pin them — see tools/make_golden.py."""
import glob
import gzip
import os
import re
import shutil

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_graph_npz(name):
    z = np.load(os.path.join(GOLD, f"{name}.graph.npz"))
    n = int(z["n"])
    w4 = z["w4"]
    w = np.empty(w4.size * 2, np.uint8)
    w[0::2] = w4 & 15
    w[1::2] = w4 >> 4
    return dict(w=w[:n].copy(), last=np.unpackbits(z["last"], bitorder="little")[:n].copy(),
                tip=np.unpackbits(z["tip"], bitorder="little")[:n].copy(),
                mult=z["mult"].astype(np.uint16), f=z["f"].astype(np.int64), k=int(z["k"]))


def cases():
    """[(name, thr, min_len, max_len, dump_path)] for every committed reference dump."""
    out = []
    for p in sorted(glob.glob(os.path.join(GOLD, "*.dump.gz"))):
        m = re.match(r"(.+)\.t(\d+)_(\d+)_(\d+)\.dump\.gz$", os.path.basename(p))
        out.append((m.group(1), int(m.group(2)), int(m.group(3)), int(m.group(4)), p))
    return out


def unpack_sdbg_files(name, dst):
    """gunzip tests/golden/<name>.sdbg/* into dst; returns the MEGAHIT file prefix."""
    src = os.path.join(GOLD, f"{name}.sdbg")
    os.makedirs(dst, exist_ok=True)
    for fn in os.listdir(src):
        with gzip.open(os.path.join(src, fn), "rb") as fi, open(os.path.join(dst, fn[:-3]), "wb") as fo:
            shutil.copyfileobj(fi, fo)
    return os.path.join(dst, "graph")


def pack_raw(raw):
    """Unpacked golden arrays -> MEGAHIT's in-memory packing (include/mcaat_gpu.h mg_graph_desc):
    w 16 nibbles per uint64 (little-endian nibble order), last/tip 64 bits per uint64."""
    n = raw["w"].size
    w = np.zeros((n + 15) // 16 * 16, np.uint8)
    w[:n] = raw["w"]
    w8 = (w[0::2] | (w[1::2] << 4)).astype(np.uint8)
    pad = (-w8.size) % 8
    w64 = np.concatenate([w8, np.zeros(pad, np.uint8)]).view("<u8").copy()

    def bits(a):
        b = np.packbits(a.astype(np.uint8), bitorder="little")
        pad = (-b.size) % 8
        return np.concatenate([b, np.zeros(pad, np.uint8)]).view("<u8").copy()

    return dict(n=n, k=raw["k"], w=w64, last=bits(raw["last"]), tip=bits(raw["tip"]),
                mult=np.ascontiguousarray(raw["mult"], np.uint16), f=np.ascontiguousarray(raw["f"], np.int64))
