"""TEST INFRASTRUCTURE: ctypes view of tests/emu/libemu.so — the product's `__host__ __device__`
search logic (mcaat_b200/csrc/*.cuh) compiled for the host, so it can be checked against the
oracle without a GPU.  Never imported by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emu", "emu.cpp")
LIB = os.path.join(HERE, "emu", "libemu.so")
HDRS = [os.path.join(HERE, "..", "mcaat_b200", "csrc", h) for h in ("sdbg_dev.cuh", "search_dev.cuh", "ball_dev.cuh", "kmer_dev.cuh")]

_lib = None


def lib():
    global _lib
    if _lib is None:
        newest = max(os.path.getmtime(p) for p in [SRC] + HDRS)
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
            gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
            subprocess.run([gxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-x", "c++", SRC, "-o", LIB], check=True)
        L = C.CDLL(LIB)
        L.emu_create.restype = C.c_void_p
        L.emu_create.argtypes = [C.c_uint64] + [C.c_void_p] * 5
        L.emu_free.argtypes = [C.c_void_p]
        L.emu_navigate.argtypes = [C.c_void_p] * 5
        L.emu_rank_select_check.argtypes = [C.c_void_p, C.c_void_p]
        L.emu_prune.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.emu_invalid.argtypes = [C.c_void_p, C.c_void_p]
        L.emu_candidates.restype = C.c_uint64
        L.emu_keep_regions.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]
        L.emu_set_kmer_index.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32,
                                         C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]
        L.emu_kmer_index_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.emu_map_read.restype = C.c_uint64
        L.emu_map_read.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
        L.emu_candidates.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.emu_bucket.restype = C.c_int
        L.emu_bucket.argtypes = [C.c_void_p, C.c_uint64]
        L.emu_dls.restype = C.c_int
        L.emu_dls.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p]
        L.emu_dls_ball.restype = C.c_int
        L.emu_dls_ball.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p]
        L.emu_dls_bitmap.restype = C.c_int
        L.emu_dls_bitmap.argtypes = [C.c_void_p, C.c_uint64, C.c_int]
        L.emu_find_cycles.restype = C.c_int
        L.emu_find_cycles.argtypes = ([C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_uint32, C.c_uint32]
                                      + [C.c_void_p] * 8)
        L.emu_set_algo.argtypes = [C.c_int, C.c_uint32]
        L.emu_set_search_limits.argtypes = [C.c_int, C.c_uint32]
        L.emu_contracted_ball.restype = C.c_uint64
        L.emu_adj.argtypes = [C.c_void_p, C.c_void_p]
        L.emu_contracted_ball.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]
        L.emu_search_one.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def set_algo(algo=2, ball_log=16):
    """0 = hash-map search (enum_run) only; 1 = extracted-ball search node by node (ball_dfs), enum_run on overflow;
    2 = the same on the contracted ball (sball_dfs) — what the device runs"""
    lib().emu_set_algo(algo, ball_log)


def set_search_limits(max_steps=0, chain_cap=0):
    """step cap of FindCycle (0 = the reference's 10 M) and record capacity of the contracted ball (0 = no limit)"""
    lib().emu_set_search_limits(max_steps, chain_cap)


class EmuGraph:
    def __init__(self, packed):
        self.p = packed
        self.n = packed["n"]
        self.h = C.c_void_p(lib().emu_create(self.n, packed["w"].ctypes.data, packed["last"].ctypes.data,
                                             packed["tip"].ctypes.data, packed["mult"].ctypes.data,
                                             packed["f"].ctypes.data))

    def __del__(self):
        if getattr(self, "h", None):
            lib().emu_free(self.h)
            self.h = None

    def navigate(self):
        n = self.n
        od = np.empty(n, np.int8)
        idg = np.empty(n, np.int8)
        o = np.empty((n, 4), np.uint64)
        i_ = np.empty((n, 4), np.uint64)
        lib().emu_navigate(self.h, od.ctypes.data, idg.ctypes.data, o.ctypes.data, i_.ctypes.data)
        return od, idg, o, i_

    def rank_select_bad(self):
        bad = C.c_int64(0)
        lib().emu_rank_select_check(self.h, C.byref(bad))
        return bad.value

    def prune(self, reverse_order=False):
        st = np.zeros(6, np.uint64)
        lib().emu_prune(self.h, st.ctypes.data, int(reverse_order))
        return dict(zip(("n_tips_before", "n_mult1", "n_peeled", "n_valid", "n_tips_after", "walk_cap_hit"), map(int, st)))

    def invalid(self):
        out = np.empty(self.n, np.uint8)
        lib().emu_invalid(self.h, out.ctypes.data)
        return out.astype(bool)

    def keep_regions(self, ids, k, reverse=False):
        ids = np.ascontiguousarray(ids, np.uint64)
        st = np.zeros(4, np.uint64)
        lib().emu_keep_regions(self.h, ids.ctypes.data, ids.size, int(k), int(reverse), st.ctypes.data)
        return dict(zip(("n_added", "levels", "n_invalidated", "n_kept"), map(int, st)))

    def set_kmer_index(self, k, tip_words, labels, first, second, use_jump=True):
        labels = np.ascontiguousarray(labels, np.uint32)
        self._kx = (np.ascontiguousarray(tip_words, np.uint64), labels, np.ascontiguousarray(first, np.int64),
                    np.ascontiguousarray(second, np.int64))
        t, l, f, s = self._kx
        lib().emu_set_kmer_index(self.h, int(k), t.ctypes.data, l.ctypes.data, l.shape[0], l.shape[1] if l.ndim == 2 else 0,
                                 f.ctypes.data, s.ctypes.data, f.size, int(use_jump))

    def kmer_index_stats(self):
        st = np.zeros(2, np.uint64)
        lib().emu_kmer_index_stats(self.h, st.ctypes.data)
        return dict(items_with_jump=int(st[0]), buckets_ok=int(st[1]))

    def map_reads(self, seqs, cycle_nodes, reverse=False):
        member = np.zeros((self.n + 31) // 32 + 1, np.uint32)
        ids = np.asarray(cycle_nodes, np.int64)
        np.bitwise_or.at(member, ids >> 5, (np.uint32(1) << (ids & 31).astype(np.uint32)))
        out = []
        for s in seqs:
            b = s.encode() if isinstance(s, str) else bytes(s)
            buf = np.zeros(max(len(b), 1), np.uint64)
            c = lib().emu_map_read(self.h, b, len(b), int(reverse), member.ctypes.data, buf.ctypes.data)
            out.append(buf[:c].copy())
        return out

    def candidates(self, thr):
        ids = np.empty(self.n, np.uint64)
        c = lib().emu_candidates(self.h, thr, ids.ctypes.data)
        return ids[:c].copy()

    def bucket(self, node):
        return lib().emu_bucket(self.h, int(node))

    def dls(self, start, limit, log2cap=11):
        exp = C.c_uint64(0)
        return lib().emu_dls(self.h, int(start), limit, log2cap, C.byref(exp)), exp.value

    def dls_ball(self, start, limit, ball_log=16):
        """(code, expansions, chains) of the admission search on the extracted, contracted ball"""
        exp, nc = C.c_uint64(0), C.c_uint64(0)
        r = lib().emu_dls_ball(self.h, int(start), limit, ball_log, C.byref(exp), C.byref(nc))
        return r, exp.value, nc.value

    def dls_bitmap(self, start, limit):
        return lib().emu_dls_bitmap(self.h, int(start), limit)

    def search_one(self, start, min_len, max_len, algo):
        """one search against an all-clear visited bitmap: (stats dict, [cycle arrays])"""
        st = np.zeros(7, np.uint64)
        cyc_len = np.zeros(501, np.uint64)
        nodes = np.zeros(501 * (max_len + 1), np.uint64)
        lib().emu_search_one(self.h, int(start), min_len, max_len, algo, st.ctypes.data, cyc_len.ctypes.data,
                             nodes.ctypes.data)
        d = dict(zip(("status", "n_cycles", "n_nodes", "steps", "expansions", "n_ball", "n_chains"), map(int, st)))
        cyc, at = [], 0
        for i in range(d["n_cycles"]):
            cyc.append(nodes[at:at + int(cyc_len[i])].copy())
            at += int(cyc_len[i])
        return d, cyc

    def contracted_ball(self, start, max_len, cap=65536):
        hot = np.zeros(2 * cap, np.uint64)
        in4 = np.zeros(cap, np.uint64)
        n = lib().emu_contracted_ball(self.h, int(start), max_len, hot.ctypes.data, in4.ctypes.data, cap)
        return hot[:2 * n].copy(), in4[:n].copy()

    def adj(self):
        """navigation cache of the pruned graph (call prune() first)"""
        out = np.zeros(self.n, np.uint64)
        lib().emu_adj(self.h, out.ctypes.data)
        return out

    def find_cycles(self, starts, min_len, max_len, window=64, lock_log=13):
        starts = np.ascontiguousarray(starts, np.uint64)
        n = starts.size
        searched = np.empty(n + 1, np.uint64)
        cyc_count = np.empty(n + 1, np.uint64)
        cap_c = 500 * (n + 1)
        cyc_len = np.empty(cap_c, np.uint64)
        nodes = np.empty(cap_c * (max_len + 1), np.uint64)
        ns, nc, nn = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        stats = np.zeros(6, np.uint64)
        r = lib().emu_find_cycles(self.h, starts.ctypes.data, n, min_len, max_len, window, lock_log,
                                  searched.ctypes.data, C.byref(ns), cyc_count.ctypes.data, cyc_len.ctypes.data,
                                  C.byref(nc), nodes.ctypes.data, C.byref(nn), stats.ctypes.data)
        assert r == 0
        out = []
        ci = ni = 0
        for i in range(ns.value):
            paths = []
            for _ in range(int(cyc_count[i])):
                ln = int(cyc_len[ci])
                ci += 1
                paths.append(nodes[ni:ni + ln].copy())
                ni += ln
            out.append((int(searched[i]), paths))
        return out, dict(zip(("skipped", "bound", "cap", "reruns", "abandoned", "overflow_reruns"), map(int, stats)))
