"""ctypes binding of libmcaat_gpu.so (include/mcaat_gpu.h) — the only compute path of this package.

There is no CPU fallback: importing works anywhere (so the C-ABI surface can be checked without a
GPU), but every compute call fails loudly when the shared library or a CUDA device is missing.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmcaat_gpu.so")
ROOT = os.path.dirname(HERE)

u64p = C.POINTER(C.c_uint64)
i64p = C.POINTER(C.c_int64)

MG_MEM_HOST, MG_MEM_DEVICE = 0, 1

# every symbol include/mcaat_gpu.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "mg_graph_create", "mg_graph_create_from_files", "mg_graph_destroy", "mg_graph_reset", "mg_graph_size", "mg_graph_device_bytes",
    "mg_navigate", "mg_prune", "mg_collect_tips", "mg_invalid_readback", "mg_invalid_readback_noalloc", "mg_candidates", "mg_dls", "mg_select_starts",
    "mg_bucket_starts", "mg_starts_free", "mg_find_cycles", "mg_cycles_free", "mg_visited_readback",
    "mg_keep_regions", "mg_invalid_merge", "mg_valid_subgraph", "mg_subgraph_free",
    "mg_kmer_index_create", "mg_graph_has_kmer_index", "mg_map_reads", "mg_reads_free",
    "mg_timers", "mg_timers_reset", "mg_step_begin", "mg_step_end", "mg_launch_count", "mg_free", "mg_last_error", "mg_version",
]


class RegionStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("n_seed_ids", "n_added", "levels", "n_invalidated", "n_kept")]


class Subgraph(C.Structure):
    _fields_ = [("n_nodes", C.c_uint64), ("n_edges", C.c_uint64), ("ids", u64p), ("off", u64p),
                ("nbr", C.POINTER(C.c_uint32))]


class KmerIndexDesc(C.Structure):
    _fields_ = [("memory", C.c_int32), ("words_per_tip_label", C.c_uint32), ("tip", C.c_void_p), ("tip_labels", C.c_void_p),
                ("n_tips", C.c_uint64), ("bucket_first", C.c_void_p), ("bucket_last", C.c_void_p), ("n_buckets", C.c_uint64)]


class Reads(C.Structure):
    _fields_ = [("n_reads", C.c_uint64), ("n_kept", C.c_uint64), ("n_ids", C.c_uint64), ("off", u64p), ("ids", u64p)]


class MgError(RuntimeError):
    pass


class GraphDesc(C.Structure):
    _fields_ = [
        ("n_items", C.c_uint64), ("k", C.c_uint32), ("memory", C.c_int32), ("device", C.c_int32),
        ("reserved", C.c_int32),
        ("w", C.c_void_p), ("last", C.c_void_p), ("tip", C.c_void_p), ("invalid", C.c_void_p),
        ("mult16", C.c_void_p), ("mult8", C.c_void_p), ("large_ids", C.c_void_p), ("large_vals", C.c_void_p),
        ("n_large", C.c_uint64), ("f", C.c_int64 * 6),
    ]


class PruneStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("n_tips_before", "n_mult1", "n_tips_after", "n_valid", "n_peeled")]


class Starts(C.Structure):
    _fields_ = [("n_candidates", C.c_uint64), ("n_starts", C.c_uint64), ("n_buckets", C.c_uint64),
                ("bucket_key", i64p), ("bucket_off", u64p), ("start_ids", u64p), ("dls_expansions", C.c_uint64)]


class Cycles(C.Structure):
    _fields_ = [("n_searched", C.c_uint64), ("searched_start", u64p), ("cyc_off", u64p), ("n_cycles", C.c_uint64),
                ("node_off", u64p), ("nodes", u64p), ("n_skipped", C.c_uint64), ("n_bound_hits", C.c_uint64),
                ("n_cap_hits", C.c_uint64), ("n_reruns", C.c_uint64), ("n_rounds", C.c_uint64),
                ("n_searches_run", C.c_uint64), ("n_abandoned", C.c_uint64), ("expansions", C.c_uint64),
                ("n_out_of_order", C.c_uint64)]


_lib = None


def lib():
    """Load libmcaat_gpu.so; raises MgError if it has not been built (see __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MgError(f"{LIB_PATH} is missing: build it with `make -C mcaat_b200/csrc` "
                          "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.mg_graph_create.argtypes = [C.POINTER(GraphDesc), C.POINTER(C.c_void_p)]
        L.mg_graph_create_from_files.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.mg_graph_destroy.argtypes = [C.c_void_p]
        L.mg_graph_destroy.restype = None
        L.mg_graph_reset.argtypes = [C.c_void_p]
        L.mg_graph_size.argtypes = [C.c_void_p]
        L.mg_graph_size.restype = C.c_uint64
        L.mg_graph_device_bytes.argtypes = [C.c_void_p]
        L.mg_graph_device_bytes.restype = C.c_uint64
        L.mg_navigate.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64] + [C.c_void_p] * 5
        L.mg_prune.argtypes = [C.c_void_p, C.POINTER(PruneStats)]
        L.mg_invalid_readback.argtypes = [C.c_void_p, C.c_void_p]
        L.mg_invalid_readback_noalloc.argtypes = [C.c_void_p, C.c_void_p]
        L.mg_visited_readback.argtypes = [C.c_void_p, C.c_void_p]
        L.mg_candidates.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(u64p), C.POINTER(C.c_uint64)]
        L.mg_collect_tips.argtypes = [C.c_void_p, C.POINTER(u64p), C.POINTER(C.c_uint64)]
        L.mg_dls.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]
        L.mg_select_starts.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.POINTER(Starts)]
        L.mg_bucket_starts.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(Starts)]
        L.mg_starts_free.argtypes = [C.POINTER(Starts)]
        L.mg_starts_free.restype = None
        L.mg_find_cycles.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.POINTER(Cycles)]
        L.mg_cycles_free.argtypes = [C.POINTER(Cycles)]
        L.mg_cycles_free.restype = None
        L.mg_graph_has_kmer_index.argtypes = [C.c_void_p]
        L.mg_kmer_index_create.argtypes = [C.c_void_p, C.POINTER(KmerIndexDesc)]
        L.mg_map_reads.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_uint64,
                                   C.POINTER(Reads)]
        L.mg_reads_free.argtypes = [C.POINTER(Reads)]
        L.mg_reads_free.restype = None
        L.mg_valid_subgraph.argtypes = [C.c_void_p, C.POINTER(Subgraph)]
        L.mg_subgraph_free.argtypes = [C.POINTER(Subgraph)]
        L.mg_subgraph_free.restype = None
        L.mg_invalid_merge.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        L.mg_keep_regions.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(RegionStats)]
        L.mg_timers.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.mg_timers_reset.argtypes = [C.c_void_p]
        L.mg_timers_reset.restype = None
        L.mg_step_begin.argtypes = [C.c_void_p]
        L.mg_step_end.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.mg_launch_count.restype = C.c_uint64
        L.mg_free.argtypes = [C.c_void_p]
        L.mg_free.restype = None
        L.mg_last_error.restype = C.c_char_p
        L.mg_version.restype = C.c_char_p
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise MgError(f"libmcaat_gpu error {rc}: {lib().mg_last_error().decode()}")


def _np(ptr, n, dtype=np.uint64):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(int(n),)).astype(dtype, copy=True)


def launch_count():
    return int(lib().mg_launch_count())


class DeviceGraph:
    """A succinct de Bruijn graph resident in HBM (mg_graph handle).

    `packed` holds MEGAHIT's in-memory packing (mg_graph_desc): n, k, w (uint64 words, 16 nibbles each),
    last / tip (uint64 bit words), mult (uint16 per item), f (int64[6]).  Arrays are numpy (host) or, with
    memory=MG_MEM_DEVICE, objects exposing `.data_ptr()` (torch CUDA tensors, e.g. after an NCCL broadcast).
    """

    def __init__(self, packed, device=0, memory=MG_MEM_HOST):
        self._keep = packed
        d = GraphDesc()
        d.n_items = int(packed["n"])
        d.k = int(packed.get("k", 23))
        d.memory = memory
        d.device = device

        def ptr(a):
            if a is None:
                return None
            if memory == MG_MEM_DEVICE:
                return C.c_void_p(a.data_ptr())
            return C.c_void_p(a.ctypes.data)

        if memory == MG_MEM_HOST:
            n = d.n_items
            assert packed["w"].dtype == np.uint64 and packed["w"].size >= (n + 15) // 16
            assert packed["last"].dtype == np.uint64 and packed["last"].size >= (n + 63) // 64
            assert packed["tip"].dtype == np.uint64 and packed["tip"].size >= (n + 63) // 64
            if "mult8" not in packed:
                assert packed["mult"].dtype == np.uint16 and packed["mult"].size >= n
        d.w, d.last, d.tip = ptr(packed["w"]), ptr(packed["last"]), ptr(packed["tip"])
        d.invalid = ptr(packed.get("invalid"))
        if "mult8" in packed:
            # the reference's in-memory form (sdbg.h:96-105): one byte per item, sentinel 255 -> (id, value) table by id
            assert memory == MG_MEM_HOST and packed["mult8"].dtype == np.uint8 and packed["mult8"].size >= n
            assert packed["large_ids"].dtype == np.uint64 and packed["large_vals"].dtype == np.uint16
            d.mult8, d.large_ids, d.large_vals = ptr(packed["mult8"]), ptr(packed["large_ids"]), ptr(packed["large_vals"])
            d.n_large = int(packed["large_ids"].size)
        else:
            d.mult16 = ptr(packed["mult"])
        f = packed["f"]
        for i in range(6):
            d.f[i] = int(f[i])
        self.n = d.n_items
        self.h = C.c_void_p()
        _check(lib().mg_graph_create(C.byref(d), C.byref(self.h)))

    @classmethod
    def from_files(cls, prefix, device=0):
        """MEGAHIT's on-disk SDBG (<prefix>.sdbg_info, <prefix>.sdbg.<i>) decoded on the device
        (mg_graph_create_from_files; replaces SDBG::LoadFromFile, reference sdbg.h:26-61)."""
        self = cls.__new__(cls)
        self._keep = None
        self.h = C.c_void_p()
        _check(lib().mg_graph_create_from_files(str(prefix).encode(), int(device), C.byref(self.h)))
        self.n = int(lib().mg_graph_size(self.h))
        return self

    def close(self):
        if getattr(self, "h", None):
            lib().mg_graph_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- graph state
    def reset(self):
        _check(lib().mg_graph_reset(self.h))

    def device_bytes(self):
        return int(lib().mg_graph_device_bytes(self.h))

    def invalid(self):
        words = np.zeros((self.n + 63) // 64, np.uint64)
        _check(lib().mg_invalid_readback(self.h, words.ctypes.data))
        return np.unpackbits(words.view(np.uint8), bitorder="little")[:self.n].astype(bool)

    def invalid_words(self, out=None):
        """current invalid bitmap as uint64 words; `out` lets the caller own the buffer (e.g. page-locked memory)"""
        words = np.zeros((self.n + 63) // 64, np.uint64) if out is None else out
        assert words.dtype == np.uint64 and words.size >= (self.n + 63) // 64
        _check(lib().mg_invalid_readback(self.h, words.ctypes.data))
        return words

    def invalid_words_noalloc(self, out=None):
        """mg_invalid_readback_noalloc: the same words through the library's static scratch (no allocation on the
        device side; graphs of at most MG_NOALLOC_MAX_ITEMS = 2**25 items).  `out` lets the caller own the buffer."""
        words = np.zeros((self.n + 63) // 64, np.uint64) if out is None else out
        _check(lib().mg_invalid_readback_noalloc(self.h, words.ctypes.data))
        return words

    def merge_invalid(self, invalid):
        """OR a host invalid array (bool per item) into the device bitmap; returns the number of new bits."""
        bits = np.packbits(np.asarray(invalid, bool).astype(np.uint8), bitorder="little")
        words = np.zeros((self.n + 63) // 64 * 8, np.uint8)
        words[:bits.size] = bits
        new = C.c_uint64(0)
        _check(lib().mg_invalid_merge(self.h, words.ctypes.data, C.byref(new)))
        return int(new.value)

    def visited(self):
        words = np.zeros((self.n + 63) // 64, np.uint64)
        _check(lib().mg_visited_readback(self.h, words.ctypes.data))
        return np.unpackbits(words.view(np.uint8), bitorder="little")[:self.n].astype(bool)

    # ---- SDBG navigation (reference sdbg.h:328-357)
    def navigate(self, first=0, count=None):
        count = self.n - first if count is None else count
        od = np.empty(count, np.int8)
        idg = np.empty(count, np.int8)
        o = np.empty((count, 4), np.uint64)
        i_ = np.empty((count, 4), np.uint64)
        m = np.empty(count, np.uint16)
        _check(lib().mg_navigate(self.h, first, count, od.ctypes.data, idg.ctypes.data, o.ctypes.data,
                                 i_.ctypes.data, m.ctypes.data))
        return od, idg, o, i_, m

    # ---- phases
    def prune(self):
        st = PruneStats()
        _check(lib().mg_prune(self.h, C.byref(st)))
        return {k: int(getattr(st, k)) for k, _ in PruneStats._fields_}

    def collect_tips(self):
        """CollectTips (reference src/cycle_finder.cpp:346-357) on the current graph state: ascending ids"""
        ids = u64p()
        n = C.c_uint64(0)
        _check(lib().mg_collect_tips(self.h, C.byref(ids), C.byref(n)))
        out = _np(ids, n.value)
        lib().mg_free(ids)
        return out

    def candidates(self, threshold_multiplicity):
        ids = u64p()
        n = C.c_uint64(0)
        _check(lib().mg_candidates(self.h, int(threshold_multiplicity), C.byref(ids), C.byref(n)))
        out = _np(ids, n.value)
        lib().mg_free(ids)
        return out

    def keep_regions(self, ids, k):
        """keep_crispr_regions_extended_by_k (reference src/spacer_ordering.cpp:78-139) on the device graph:
        ids = flattened cycle node ids, k = rounds of neighbourhood growth; returns the stats dict."""
        ids = np.ascontiguousarray(ids, np.uint64)
        st = RegionStats()
        _check(lib().mg_keep_regions(self.h, ids.ctypes.data, ids.size, int(k), C.byref(st)))
        return {k_: int(getattr(st, k_)) for k_, _ in RegionStats._fields_}

    def set_kmer_index(self, tip_words, labels, first, second):
        """attach the k-mer lookup index to a graph made from arrays (mg_kmer_index_create): tip bitmap words, tip
        labels as file words [n_tips, words_per_label], bucket first/last item"""
        tip = np.ascontiguousarray(tip_words, np.uint64)
        labels = np.ascontiguousarray(labels, np.uint32)
        first = np.ascontiguousarray(first, np.int64)
        second = np.ascontiguousarray(second, np.int64)
        d = KmerIndexDesc(0, labels.shape[1] if labels.ndim == 2 else 0, tip.ctypes.data, labels.ctypes.data,
                          labels.shape[0], first.ctypes.data, second.ctypes.data, first.size)
        _check(lib().mg_kmer_index_create(self.h, C.byref(d)))

    def map_reads(self, seqs, cycle_ids, reverse=False):
        """get_reads' per-file loop (reference src/reads.cpp:105-127): seqs = list of bytes/str; returns
        (off, ids): read r -> ids[off[r]:off[r+1]], empty = dropped, 2**64-1 = k-mer not in the graph"""
        if isinstance(seqs, tuple):  # (uint8 array of all characters, uint64 offsets): no per-read Python objects
            chars, off = np.ascontiguousarray(seqs[0], np.uint8), np.ascontiguousarray(seqs[1], np.uint64)
            blob, n_reads = chars.ctypes.data_as(C.c_char_p), off.size - 1
        else:
            bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
            off = np.zeros(len(bs) + 1, np.uint64)
            if bs:
                off[1:] = np.cumsum([len(b) for b in bs])
            blob, n_reads = b"".join(bs), len(bs)
        cyc = np.ascontiguousarray(cycle_ids, np.uint64)
        r = Reads()
        _check(lib().mg_map_reads(self.h, blob, off.ctypes.data, n_reads, int(reverse), cyc.ctypes.data, cyc.size, C.byref(r)))
        o = _np(r.off, r.n_reads + 1)
        ids = _np(r.ids, r.n_ids)
        lib().mg_reads_free(C.byref(r))
        return o, ids

    def valid_subgraph(self):
        """(ids ascending, off, nbr): the valid items and, per item, the positions of its valid out-neighbours
        in OutgoingEdges order — what the reference's SCC split walks (src/spacer_ordering.cpp:54-76,141-175)."""
        sg = Subgraph()
        _check(lib().mg_valid_subgraph(self.h, C.byref(sg)))
        ids = _np(sg.ids, sg.n_nodes)
        off = _np(sg.off, sg.n_nodes + 1)
        nbr = _np(sg.nbr, sg.n_edges, np.uint32)
        lib().mg_subgraph_free(C.byref(sg))
        return ids, off, nbr

    def dls(self, ids, limit):
        ids = np.ascontiguousarray(ids, np.uint64)
        out = np.zeros(ids.size, np.uint8)
        exp = C.c_uint64(0)
        _check(lib().mg_dls(self.h, ids.ctypes.data, ids.size, int(limit), out.ctypes.data, C.byref(exp)))
        return out.astype(bool), int(exp.value)

    @staticmethod
    def _starts_out(s):
        boff = _np(s.bucket_off, s.n_buckets + 1)
        keys = _np(s.bucket_key, s.n_buckets, np.int64)
        ids = _np(s.start_ids, s.n_starts)
        buckets = [(int(keys[i]), ids[int(boff[i]):int(boff[i + 1])]) for i in range(int(s.n_buckets))]
        info = dict(n_candidates=int(s.n_candidates), n_starts=int(s.n_starts), dls_expansions=int(s.dls_expansions))
        lib().mg_starts_free(C.byref(s))
        return buckets, info

    def select_starts(self, threshold_multiplicity, max_len):
        s = Starts()
        _check(lib().mg_select_starts(self.h, int(threshold_multiplicity), int(max_len), C.byref(s)))
        return self._starts_out(s)

    def bucket_starts(self, ids):
        ids = np.ascontiguousarray(ids, np.uint64)
        s = Starts()
        _check(lib().mg_bucket_starts(self.h, ids.ctypes.data, ids.size, C.byref(s)))
        return self._starts_out(s)

    def find_cycles(self, starts, min_len, max_len, csr=False):
        """-> ([(start, [path ndarray, ...])] in search order, stats); csr=True returns the library's own layout
        instead — (searched_start, cyc_off, node_off, nodes) — without building a Python object per cycle"""
        starts = np.ascontiguousarray(starts, np.uint64)
        c = Cycles()
        _check(lib().mg_find_cycles(self.h, starts.ctypes.data, starts.size, int(min_len), int(max_len), C.byref(c)))
        coff = _np(c.cyc_off, c.n_searched + 1)
        noff = _np(c.node_off, c.n_cycles + 1)
        nodes = _np(c.nodes, noff[-1] if noff.size else 0)
        ss = _np(c.searched_start, c.n_searched)
        searched = (ss, coff, noff, nodes) if csr else []
        for i in range(0 if csr else int(c.n_searched)):
            searched.append((int(ss[i]), [nodes[int(noff[j]):int(noff[j + 1])]
                                          for j in range(int(coff[i]), int(coff[i + 1]))]))
        stats = {k: int(getattr(c, k)) for k in ("n_searched", "n_cycles", "n_skipped", "n_bound_hits", "n_cap_hits",
                                                  "n_reruns", "n_rounds", "n_searches_run", "n_abandoned", "expansions", "n_out_of_order")}
        lib().mg_cycles_free(C.byref(c))
        return searched, stats

    # ---- device timers
    def timers(self):
        cap = 16
        names = (C.c_char_p * cap)()
        ms = (C.c_double * cap)()
        launches = (C.c_uint64 * cap)()
        units = (C.c_uint64 * cap)()
        n = lib().mg_timers(self.h, names, ms, launches, units, cap)
        return {names[i].decode(): dict(ms=ms[i], launches=int(launches[i]), units=int(units[i])) for i in range(n)}

    def timers_reset(self):
        lib().mg_timers_reset(self.h)

    def step_begin(self):
        _check(lib().mg_step_begin(self.h))

    def step_end(self):
        ms = C.c_double(0)
        _check(lib().mg_step_end(self.h, C.byref(ms)))
        return ms.value
