"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes binding of oracle/liboracle.so (the plain-C restatement of the reference hot path) and a
parser for the dumps written by oracle/ref_harness/ref_cyclefinder_main.cpp (the UNMODIFIED
reference).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")

u64p = C.POINTER(C.c_uint64)
i64p = C.POINTER(C.c_int64)


class OrcResult(C.Structure):
    _fields_ = [
        ("n_tips_before", C.c_uint64), ("n_mult1", C.c_uint64), ("n_tips_after", C.c_uint64),
        ("n_valid", C.c_uint64), ("n_candidates", C.c_uint64), ("n_buckets", C.c_uint64),
        ("bucket_key", i64p), ("bucket_off", u64p), ("start_ids", u64p),
        ("n_searched", C.c_uint64), ("searched_start", u64p), ("cyc_off", u64p),
        ("n_cycles", C.c_uint64), ("node_off", u64p), ("nodes", u64p),
        ("n_bound_hits", C.c_uint64), ("n_cap_hits", C.c_uint64),
        ("dls_expansions", C.c_uint64), ("enum_expansions", C.c_uint64),
    ]


class OrcCycles(C.Structure):
    _fields_ = [("n_cycles", C.c_uint64), ("node_off", u64p), ("nodes", u64p),
                ("bound_hit", C.c_int), ("cap_hit", C.c_int), ("expansions", C.c_uint64),
                ("steps", C.c_uint64)]


def build():
    """Compile liboracle.so (and oracle/_ref when the reference sources are mounted)."""
    subprocess.run(["make", "-s", "-C", HERE, "all"], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_load.restype = C.c_void_p
        L.orc_load.argtypes = [C.c_char_p]
        L.orc_from_arrays.restype = C.c_void_p
        L.orc_from_arrays.argtypes = [C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_size.restype = C.c_uint64
        L.orc_size.argtypes = [C.c_void_p]
        L.orc_k.restype = C.c_uint32
        L.orc_k.argtypes = [C.c_void_p]
        L.orc_get_f.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_raw.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.orc_get_invalid.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_set_invalid.argtypes = [C.c_void_p, C.c_void_p]
        for name in ("orc_rank_w", "orc_select_w"):
            f = getattr(L, name)
            f.restype = C.c_int64
            f.argtypes = [C.c_void_p, C.c_int, C.c_int64]
        for name in ("orc_rank_last", "orc_select_last"):
            f = getattr(L, name)
            f.restype = C.c_int64
            f.argtypes = [C.c_void_p, C.c_int64]
        for name in ("orc_forward", "orc_backward"):
            f = getattr(L, name)
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p, C.c_uint64]
        for name in ("orc_outgoing", "orc_incoming"):
            f = getattr(L, name)
            f.restype = C.c_int
            f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.orc_mult.restype = C.c_uint16
        L.orc_mult.argtypes = [C.c_void_p, C.c_uint64]
        L.orc_collect_tips.restype = C.c_uint64
        L.orc_collect_tips.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_invalidate_mult1.restype = C.c_uint64
        L.orc_invalidate_mult1.argtypes = [C.c_void_p]
        L.orc_peel.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_dls.restype = C.c_int
        L.orc_dls.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
        L.orc_find_cycle_util.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_cycles_free.argtypes = [C.c_void_p]
        L.orc_run.restype = C.c_int
        L.orc_run.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p]
        L.orc_result_free.argtypes = [C.c_void_p]
        L.orc_index_binary_search.restype = C.c_uint64
        L.orc_index_binary_search.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_read.restype = C.c_uint64
        L.orc_get_read.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p]
        L.orc_reverse_complement.argtypes = [C.c_char_p, C.c_uint64]
        L.orc_kmer_index_sizes.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_kmer_index_get.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_keep_regions.restype = C.c_uint64
        L.orc_keep_regions.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]
        L.orc_uset_order.restype = C.c_int
        L.orc_uset_order.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _np(ptr, n, dtype=np.uint64):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(int(n),)).astype(dtype, copy=True)


class Run:
    """Plain-python view of one whole search (same fields as a parsed reference dump)."""

    def __init__(self):
        self.n_items = 0
        self.invalid = None      # bool array per item
        self.buckets = []        # [(key, ids ndarray)] descending key
        self.searched = []       # [(start, [ndarray path, ...])] in search order
        self.iter_order = None   # std::unordered_map iteration order of `results` (reference dumps only)
        self.stats = {}

    def start_list(self):
        return np.concatenate([ids for _, ids in self.buckets]) if self.buckets else np.zeros(0, np.uint64)


class OracleGraph:
    def __init__(self, handle):
        if not handle:
            raise RuntimeError("oracle: could not load graph")
        self.h = C.c_void_p(handle)
        self.n = int(lib().orc_size(self.h))
        self.k = int(lib().orc_k(self.h))

    @classmethod
    def load(cls, prefix):
        return cls(lib().orc_load(os.fsencode(prefix)))

    @classmethod
    def from_arrays(cls, w, last, tip, mult, f, k=23):
        w = np.ascontiguousarray(w, np.uint8)
        last = np.ascontiguousarray(last, np.uint8)
        tip = np.ascontiguousarray(tip, np.uint8)
        mult = np.ascontiguousarray(mult, np.uint16)
        f = np.ascontiguousarray(f, np.int64)
        return cls(lib().orc_from_arrays(w.size, k, w.ctypes.data, last.ctypes.data, tip.ctypes.data,
                                         mult.ctypes.data, f.ctypes.data))

    def close(self):
        if self.h:
            lib().orc_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def raw(self):
        w = np.empty(self.n, np.uint8); last = np.empty(self.n, np.uint8)
        tip = np.empty(self.n, np.uint8); mult = np.empty(self.n, np.uint16)
        lib().orc_get_raw(self.h, w.ctypes.data, last.ctypes.data, tip.ctypes.data, mult.ctypes.data)
        f = np.empty(6, np.int64)
        lib().orc_get_f(self.h, f.ctypes.data)
        return dict(w=w, last=last, tip=tip, mult=mult, f=f, k=self.k)

    def invalid(self):
        inv = np.empty(self.n, np.uint8)
        lib().orc_get_invalid(self.h, inv.ctypes.data)
        return inv.astype(bool)

    def set_invalid(self, inv):
        inv = np.ascontiguousarray(inv, np.uint8)
        lib().orc_set_invalid(self.h, inv.ctypes.data)

    def outgoing(self, e):
        out = (C.c_uint64 * 4)()
        d = lib().orc_outgoing(self.h, e, out)
        return d, list(out)[:max(d, 0)]

    def incoming(self, e):
        out = (C.c_uint64 * 4)()
        d = lib().orc_incoming(self.h, e, out)
        return d, list(out)[:max(d, 0)]

    def nav_table(self):
        """(outdeg, indeg, out[n,4], in[n,4]) for every item; unused slots = 2**64-1."""
        n = self.n
        od = np.empty(n, np.int8); idg = np.empty(n, np.int8)
        o = np.full((n, 4), np.iinfo(np.uint64).max, np.uint64)
        i_ = np.full((n, 4), np.iinfo(np.uint64).max, np.uint64)
        L = lib()
        buf = (C.c_uint64 * 4)()
        for e in range(n):
            d = L.orc_outgoing(self.h, e, buf)
            od[e] = d
            for j in range(max(d, 0)):
                o[e, j] = buf[j]
            d = L.orc_incoming(self.h, e, buf)
            idg[e] = d
            for j in range(max(d, 0)):
                i_[e, j] = buf[j]
        return od, idg, o, i_

    def dls(self, start, limit):
        exp = C.c_uint64(0)
        return bool(lib().orc_dls(self.h, start, limit, C.byref(exp))), exp.value

    def find_cycles(self, start, min_len, max_len, visited):
        """visited: uint8 ndarray (mutated like the reference mutates `visited`)."""
        c = OrcCycles()
        lib().orc_find_cycle_util(self.h, start, min_len, max_len, visited.ctypes.data, C.byref(c))
        off = _np(c.node_off, c.n_cycles + 1)
        nodes = _np(c.nodes, off[-1] if off.size else 0)
        paths = [nodes[off[i]:off[i + 1]] for i in range(int(c.n_cycles))]
        info = dict(bound_hit=c.bound_hit, cap_hit=c.cap_hit, expansions=c.expansions, steps=c.steps)
        lib().orc_cycles_free(C.byref(c))
        return paths, info

    def kmer_index(self):
        """(tip labels as file words [n_tips, wptl], bucket first, bucket last) of a graph loaded from files"""
        t, w, nb = C.c_uint64(0), C.c_uint32(0), C.c_uint64(0)
        lib().orc_kmer_index_sizes(self.h, C.byref(t), C.byref(w), C.byref(nb))
        labels = np.zeros((t.value, w.value), np.uint32)
        first = np.zeros(nb.value, np.int64)
        second = np.zeros(nb.value, np.int64)
        lib().orc_kmer_index_get(self.h, labels.ctypes.data, first.ctypes.data, second.ctypes.data)
        return labels, first, second

    def map_reads(self, seqs, cycle_nodes, reverse=False):
        """get_read_from_sequence (reference src/reads.cpp:57-89) per sequence (bytes/str); reverse=True applies
        reverse_pair_ends_sequence first (:20-31, :119-121).  Returns a list of uint64 arrays (empty = dropped)."""
        member = np.zeros(self.n, np.uint8)
        member[np.asarray(cycle_nodes, np.int64)] = 1
        out = []
        for s in seqs:
            b = s.encode() if isinstance(s, str) else bytes(s)
            buf = C.create_string_buffer(b, len(b) + 1)
            if reverse:
                lib().orc_reverse_complement(buf, len(b))
            ids = np.zeros(max(len(b), 1), np.uint64)
            c = lib().orc_get_read(self.h, member.ctypes.data, buf, len(b), ids.ctypes.data)
            out.append(ids[:c].copy())
        return out

    def keep_regions(self, ids, k):
        """reference keep_crispr_regions_extended_by_k (src/spacer_ordering.cpp:78-139); (invalidated, kept)"""
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        kept = C.c_uint64(0)
        gone = lib().orc_keep_regions(self.h, ids.ctypes.data, ids.size, k, C.byref(kept))
        if gone == 2**64 - 1:
            raise ValueError("node id out of range")
        return int(gone), int(kept.value)

    def run(self, thr=20, min_len=27, max_len=77):
        r = OrcResult()
        lib().orc_run(self.h, thr, min_len, max_len, C.byref(r))
        out = Run()
        out.n_items = self.n
        out.invalid = self.invalid()
        boff = _np(r.bucket_off, r.n_buckets + 1)
        keys = _np(r.bucket_key, r.n_buckets, np.int64)
        ids = _np(r.start_ids, boff[-1] if boff.size else 0)
        out.buckets = [(int(keys[i]), ids[boff[i]:boff[i + 1]]) for i in range(int(r.n_buckets))]
        coff = _np(r.cyc_off, r.n_searched + 1)
        noff = _np(r.node_off, r.n_cycles + 1)
        nodes = _np(r.nodes, noff[-1] if noff.size else 0)
        ss = _np(r.searched_start, r.n_searched)
        for i in range(int(r.n_searched)):
            paths = [nodes[noff[j]:noff[j + 1]] for j in range(int(coff[i]), int(coff[i + 1]))]
            out.searched.append((int(ss[i]), paths))
        out.stats = {k: int(getattr(r, k)) for k in (
            "n_tips_before", "n_mult1", "n_tips_after", "n_valid", "n_candidates", "n_bound_hits",
            "n_cap_hits", "dls_expansions", "enum_expansions", "n_cycles")}
        lib().orc_result_free(C.byref(r))
        return out


def parse_dump(path):
    """Parse a dump written by ref_cyclefinder_main.cpp (layout: oracle/ref_harness/dump_format.h)."""
    import gzip
    op = gzip.open if str(path).endswith(".gz") else open
    with op(path, "rb") as f:
        buf = f.read()
    assert buf[:8] == b"MCAATDMP", "bad dump magic"
    a = np.frombuffer(buf, dtype=np.uint64, offset=8)
    p = 0

    def take(n=1):
        nonlocal p
        v = a[p:p + n]
        p += n
        return v

    r = Run()
    r.n_items = int(take()[0])
    nw = int(take()[0])
    words = take(nw)
    bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:r.n_items]
    r.invalid = bits.astype(bool)
    nb = int(take()[0])
    for _ in range(nb):
        key = int(take()[0].astype(np.int64))
        cnt = int(take()[0])
        r.buckets.append((key, take(cnt).copy()))
    ns = int(take()[0])
    for _ in range(ns):
        s = int(take()[0])
        nc = int(take()[0])
        paths = []
        for _ in range(nc):
            ln = int(take()[0])
            paths.append(take(ln).copy())
        r.searched.append((s, paths))
    ni = int(take()[0])
    r.iter_order = take(ni).copy()
    assert p == a.size, "trailing bytes in dump"
    return r


def same_run(a, b):
    """Bit-exact comparison of two runs; returns a list of human-readable differences."""
    diffs = []
    if a.n_items != b.n_items:
        diffs.append(f"n_items {a.n_items} != {b.n_items}")
        return diffs
    if not np.array_equal(a.invalid, b.invalid):
        diffs.append(f"invalid bitmap differs at {int(np.count_nonzero(a.invalid != b.invalid))} items")
    if [k for k, _ in a.buckets] != [k for k, _ in b.buckets]:
        diffs.append(f"bucket keys {[k for k, _ in a.buckets]} != {[k for k, _ in b.buckets]}")
    else:
        for (k, x), (_, y) in zip(a.buckets, b.buckets):
            if not np.array_equal(x, y):
                diffs.append(f"bucket {k}: start lists differ ({x.size} vs {y.size})")
    if [s for s, _ in a.searched] != [s for s, _ in b.searched]:
        diffs.append(f"searched start sequence differs ({len(a.searched)} vs {len(b.searched)})")
    else:
        for (s, pa), (_, pb) in zip(a.searched, b.searched):
            if len(pa) != len(pb) or any(not np.array_equal(x, y) for x, y in zip(pa, pb)):
                diffs.append(f"start {s}: raw cycle sequence differs ({len(pa)} vs {len(pb)})")
    return diffs


def ref_binary(name):
    p = os.path.join(REF_DIR, name)
    return p if os.path.exists(p) else None


def run_reference_regions(prefix, out_path, k, *, cycles=None, nodes_file=None):
    """Run the compiled UNMODIFIED keep_crispr_regions_extended_by_k (oracle/_ref/ref_regions); returns the
    invalid bitmap (bool per item) after the call.  cycles=(thr,min,max): seeds = all cycles the reference
    CycleFinder finds; nodes_file: seeds = the u64 ids in that file, graph as loaded."""
    exe = ref_binary("ref_regions")
    if exe is None:
        raise FileNotFoundError("oracle/_ref/ref_regions not built")
    tail = ["cycles"] + [str(x) for x in cycles] if cycles is not None else ["nodes", nodes_file]
    subprocess.run([exe, prefix, out_path, str(k)] + tail, check=True, capture_output=True)
    raw = open(out_path, "rb").read()
    assert raw[:8] == b"MCAATREG"
    n, kk, n_seed, n_words = np.frombuffer(raw, "<u8", 4, 8)
    words = np.frombuffer(raw, "<u8", int(n_words), 40)
    return np.unpackbits(words.view(np.uint8), bitorder="little")[:int(n)].astype(bool)


def run_reference(prefix, dump_path, threads=1, thr=20, min_len=27, max_len=77):
    """Run the compiled UNMODIFIED reference (oracle/_ref/ref_cyclefinder) and return its REF_JSON."""
    import json
    exe = ref_binary("ref_cyclefinder")
    if exe is None:
        raise FileNotFoundError("oracle/_ref/ref_cyclefinder not built")
    out = subprocess.run([exe, prefix, dump_path, str(threads), str(thr), str(min_len), str(max_len)],
                         check=True, capture_output=True, text=True).stdout
    for line in out.splitlines():
        if line.startswith("REF_JSON "):
            return json.loads(line[9:]), out
    raise RuntimeError("no REF_JSON line in reference output")
