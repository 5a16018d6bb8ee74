"""-m gpu: the bench workloads.

* c2_sample (BASELINE.json configs[1] at 1/20 scale, 11.3 M items) against the oracle, bit for bit, at the
  default parameters and at a stress setting of configs[4] (bounds 20-100, threshold 10).
* c2 itself (224.7 M items — the oracle would need minutes) through size-independent properties of the
  reference's algorithm: idempotence after reset, visited == union of all cycle nodes, every skipped
  start covered, every cycle a simple closed walk of admissible length over valid edges, start-list order.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle_py as O  # noqa: E402
from mcaat_b200 import CycleFinder, DeviceGraph, Settings  # noqa: E402

pytestmark = pytest.mark.gpu

_wl = {}


def workload(name):
    if name not in _wl:
        import synth_sdbg as S
        _wl[name] = S.make_workload(name, device="cuda")[0]
    return _wl[name]


def unpack(g):
    n = g["n"]
    w8 = g["w"].view(np.uint8)
    w = np.empty(w8.size * 2, np.uint8)
    w[0::2] = w8 & 15
    w[1::2] = w8 >> 4
    last = np.unpackbits(g["last"].view(np.uint8), bitorder="little")[:n]
    tip = np.unpackbits(g["tip"].view(np.uint8), bitorder="little")[:n]
    return w[:n], last, tip


def run(dg, thr, mn, mx):
    s = Settings(quiet=True)
    s.sdbg = dg
    s.cycle_finder_settings.threshold_multiplicity = thr
    s.cycle_finder_settings.cycle_min_length = mn
    s.cycle_finder_settings.cycle_max_length = mx
    cf = CycleFinder(s)
    r = O.Run()
    r.n_items = dg.n
    r.invalid = dg.invalid()
    r.buckets = cf.start_nodes_chunked
    r.searched = list(cf.results.items())
    r.stats = cf.stats
    return r


@pytest.mark.parametrize("thr,mn,mx", [(20, 27, 77), (10, 20, 100)])
def test_sample_workload_matches_oracle(thr, mn, mx):
    g = workload("c2_sample")
    w, last, tip = unpack(g)
    want = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"]).run(thr, mn, mx)
    got = run(DeviceGraph(g), thr, mn, mx)
    assert O.same_run(want, got) == []
    assert len(got.searched) > 50


def test_sample_workload_regions_match_oracle():
    """SURVEY 8(f) N4 at 11.3 M items: cycles of the run -> mg_keep_regions (k = 128, a 150-bp read at k = 23) against
    the oracle's restatement of keep_crispr_regions_extended_by_k on the same pruned graph."""
    g = workload("c2_sample")
    w, last, tip = unpack(g)
    og = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"])
    want = og.run(20, 27, 77)
    seeds = np.concatenate([p for _, ps in want.searched for p in ps]).astype(np.uint64)
    gone, kept = og.keep_regions(seeds, 128)
    dg = DeviceGraph(g)
    run(dg, 20, 27, 77)
    st = dg.keep_regions(seeds, 128)
    assert (st["n_invalidated"], st["n_kept"]) == (gone, kept) and kept > seeds.size // 4
    assert np.array_equal(dg.invalid(), og.invalid())


def test_full_size_properties():
    g = workload("c2")
    dg = DeviceGraph(g)
    thr, mn, mx = 20, 27, 77
    a = run(dg, thr, mn, mx)
    starts = a.start_list()
    assert starts.size > 5000 and len(a.searched) > 500
    # start-list order: buckets by descending key, ascending ids inside (reference :414-416, :468)
    keys = [k for k, _ in a.buckets]
    assert keys == sorted(keys, reverse=True)
    assert all(np.all(np.diff(ids.astype(np.int64)) > 0) for _, ids in a.buckets)
    mult = g["mult"]
    assert all(np.all(np.ceil(np.log2(mult[ids.astype(np.int64)].astype(np.float64))) == k) for k, ids in a.buckets)
    # results follow the start-list order; searched + skipped == all
    pos = {int(s): i for i, s in enumerate(starts)}
    order = [pos[int(s)] for s, _ in a.searched]
    assert order == sorted(order)
    assert a.stats["n_searched"] + a.stats["n_skipped"] == starts.size
    # visited == union of cycle nodes (:217-222); every skipped start is covered (:476)
    visited = dg.visited()
    union = np.zeros(dg.n, bool)
    for _, paths in a.searched:
        for p in paths:
            union[p.astype(np.int64)] = True
    assert np.array_equal(visited, union)
    searched = {int(s) for s, _ in a.searched}
    skipped = np.array([s for s in starts if int(s) not in searched], np.int64)
    assert visited[skipped].all()
    # every cycle: starts at its start node, admissible length, simple, valid nodes, multiplicity filter (:45)
    inv = a.invalid
    rng = np.random.default_rng(5)
    edges = []
    for s, paths in a.searched:
        for p in paths:
            q = p.astype(np.int64)
            assert q[0] == s and mn < q.size <= mx
            assert np.unique(q).size == q.size
            assert not inv[q].any()
            assert np.all(int(mult[s]) // mult[q].astype(np.int64) <= 500)
            if rng.random() < 0.02:
                j = int(rng.integers(q.size))
                edges.append((int(q[j]), int(q[(j + 1) % q.size])))
    # sampled cycle edges exist in the pruned graph (OutgoingEdges, sdbg.h:354)
    assert len(edges) > 50
    for u, v in edges[:400]:
        od, _, out, _, _ = dg.navigate(u, 1)
        assert v in [int(x) for x in out[0][:max(int(od[0]), 0)]]
    # N4 at full size: region extraction keeps every valid cycle node, only valid items, is idempotent, and a
    # second call with a larger k on the already reduced graph cannot resurrect anything
    seeds = np.concatenate([p for _, ps in a.searched for p in ps]).astype(np.uint64)
    st = dg.keep_regions(seeds, 128)
    inv2 = dg.invalid()
    assert not inv2[seeds.astype(np.int64)].any() and not (~inv2 & inv).any()
    assert st["n_kept"] == int((~inv2).sum()) and st["n_invalidated"] == int((inv2 & ~inv).sum())
    assert st["n_kept"] >= np.unique(seeds).size and st["levels"] <= 128
    again = dg.keep_regions(seeds, 128)
    assert again["n_invalidated"] == 0 and again["n_kept"] == st["n_kept"]
    # idempotence: reset + second run gives the identical result
    dg.reset()
    b = run(dg, thr, mn, mx)
    assert O.same_run(a, b) == []
