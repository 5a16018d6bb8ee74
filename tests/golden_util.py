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


def region_cases():
    """[(name, key, params or None, k)] for every bitmap in tests/golden/<name>.regions.npz (made by
    tools/make_golden.py from the UNMODIFIED reference keep_crispr_regions_extended_by_k):
    params=(thr,min,max): graph state and seed ids = that reference run's pruned graph and all its cycles;
    params=None: graph as loaded, seed ids = the file's `seeds` array."""
    out = []
    for p in sorted(glob.glob(os.path.join(GOLD, "*.regions.npz"))):
        name = os.path.basename(p)[:-len(".regions.npz")]
        for key in np.load(p).files:
            m = re.match(r"cyc_t(\d+)_(\d+)_(\d+)_k(\d+)$", key)
            if m:
                out.append((name, key, tuple(int(m.group(i)) for i in (1, 2, 3)), int(m.group(4))))
            m = re.match(r"nodes_k(\d+)$", key)
            if m:
                out.append((name, key, None, int(m.group(1))))
    return out


def region_expected(name, key, n):
    z = np.load(os.path.join(GOLD, f"{name}.regions.npz"))
    return np.unpackbits(z[key], bitorder="little")[:n].astype(bool)


def region_seeds(name, params):
    """seed ids of a region case: all cycle nodes of the reference dump, flattened in `results` iteration order
    like cycles_map_to_cycles (reference src/tmp_utils.cpp:26-38) — or the stored random ids."""
    if params is None:
        return np.load(os.path.join(GOLD, f"{name}.regions.npz"))["seeds"].astype(np.uint64)
    import oracle_py as O
    thr, mn, mx = params
    ref = O.parse_dump(os.path.join(GOLD, f"{name}.t{thr}_{mn}_{mx}.dump.gz"))
    paths = [p for _, ps in ref.searched for p in ps]
    return np.concatenate(paths).astype(np.uint64) if paths else np.zeros(0, np.uint64)


def reads_case(name):
    """tests/golden/<name>.reads.npz (tools/make_golden.py, UNMODIFIED reference get_read_from_sequence):
    (sequences as bytes, {variant: (cycle node ids, reverse flag, offsets, flat ids)})"""
    z = np.load(os.path.join(GOLD, f"{name}.reads.npz"))
    seqs = bytes(z["seqs"]).split(b"\n")
    n = load_graph_npz(name)["w"].size
    set_a = region_seeds(name, tuple(int(x) for x in z["set_a_params"]))
    set_b = np.nonzero(np.unpackbits(z["set_b"], bitorder="little")[:n])[0].astype(np.uint64)
    variants = {}
    for key, ids, rev in (("a_fwd", set_a, False), ("a_rev", set_a, True), ("b_fwd", set_b, False), ("b_rev", set_b, True)):
        variants[key] = (ids, rev, z[key + "_off"].astype(np.int64), z[key + "_ids"].astype(np.uint64))
    return seqs, variants


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


def nav_digest_cases():
    """golden graphs whose full navigation table (unmodified reference SDBG API over every item) is pinned by SHA-256
    digests, tests/golden/<case>.nav.sha.json (tools/add_golden_nav.py)"""
    import glob
    return sorted(os.path.basename(p)[:-len(".nav.sha.json")] for p in glob.glob(os.path.join(GOLD, "*.nav.sha.json")))


def nav_digests(od, idg, m, o, i_):
    """the digests of tools/add_golden_nav.py for a navigation table computed here"""
    import hashlib
    arrs = dict(od=np.ascontiguousarray(od, np.int8), id=np.ascontiguousarray(idg, np.int8), m=np.ascontiguousarray(m, np.uint16),
                o=np.ascontiguousarray(o, np.int64), i=np.ascontiguousarray(i_, np.int64))
    return {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in arrs.items()}


def nav_expected(name):
    import json
    with open(os.path.join(GOLD, f"{name}.nav.sha.json")) as f:
        return json.load(f)
