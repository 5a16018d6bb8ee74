"""-m gpu: the bench workloads.

* c1 (BASELINE.json configs[0], 20.0 M items) and c2_sample (configs[1] at 1/20 scale, 11.3 M items) against the oracle, bit for bit, at the
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
        import gc

        import torch

        import synth_sdbg as S
        gc.collect()  # device graphs of earlier tests
        _wl[name] = S.make_workload(name, device="cuda")[0]
        torch.cuda.empty_cache()  # the generator's blocks (~100 GB at c4) go back to the driver: the library allocates beside torch
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


def test_c1_workload_matches_oracle():
    """BASELINE.json configs[0] (5 x 2 Mb genomes, 10 planted arrays, error-free; 20.0 M items — the reference's own
    CPU-runnable case) against the oracle, bit for bit."""
    g = workload("c1")
    w, last, tip = unpack(g)
    want = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"]).run(20, 27, 77)
    got = run(DeviceGraph(g), 20, 27, 77)
    assert O.same_run(want, got) == []
    # SURVEY's C1 probe: 340 candidates, 20 searches (the coverage sampling differs a little between devices)
    assert 300 < want.stats["n_candidates"] < 380 and len(got.searched) >= 10
    _wl.pop("c1", None)


def test_chain_cycles_without_splitter_match_oracle(monkeypatch, capfd):
    """Closed loops of chain nodes with neither entry nor splitter (tests/test_device_logic_host.py:chain_cycle_graph):
    same result as the oracle, and the peel recognises them without the exhaustive pointer-jumping fallback."""
    from test_device_logic_host import chain_cycle_graph
    monkeypatch.setenv("MCAAT_PEEL_VERBOSE", "1")
    g = chain_cycle_graph()
    w, last, tip = unpack(g)
    want = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"]).run(2, 27, 77)
    got = run(DeviceGraph(g), 2, 27, 77)
    assert O.same_run(want, got) == []
    assert len(got.searched) > 20 and got.stats["n_peeled"] > 1000
    assert "exhaustive pointer jumping" not in capfd.readouterr().err


@pytest.mark.parametrize("thr,mn,mx", [(20, 27, 77), (10, 20, 100)])
def test_sample_workload_matches_oracle(thr, mn, mx):
    g = workload("c2_sample")
    w, last, tip = unpack(g)
    want = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"]).run(thr, mn, mx)
    got = run(DeviceGraph(g), thr, mn, mx)
    assert O.same_run(want, got) == []
    assert len(got.searched) > 50


@pytest.mark.parametrize("name", ["c2_sample", "chain_cycles"])
def test_candidate_scan_variants_agree(monkeypatch, name):
    """ChunkStartNodes' predicate (:398-408) three ways: the literal kernel on the pruned graph (MCAAT_CAND_LITERAL),
    the streaming kernel over the nodes k_peel_final marked as having a valid non-first incoming edge, and the literal
    kernel again after something else changed the invalid bits (the marks are dropped with the navigation cache)."""
    if name == "chain_cycles":
        from test_device_logic_host import chain_cycle_graph
        g = chain_cycle_graph()
    else:
        g = workload(name)
    dg = DeviceGraph(g)
    dg.prune()
    for thr in (0, 2, 20, 200):
        fast = dg.candidates(thr)
        monkeypatch.setenv("MCAAT_CAND_LITERAL", "1")
        literal = dg.candidates(thr)
        monkeypatch.delenv("MCAAT_CAND_LITERAL")
        assert np.array_equal(fast, literal) and (thr > 2 or fast.size > 20)
    # region extraction invalidates items: the in-degree of the survivors changes, the scan must follow
    seeds = dg.candidates(2)[:50]
    st = dg.keep_regions(seeds, 40)
    assert st["n_invalidated"] > 0
    after = dg.candidates(0)
    monkeypatch.setenv("MCAAT_CAND_LITERAL", "1")
    assert np.array_equal(after, dg.candidates(0))
    dg.close()


@pytest.mark.parametrize("name", ["c2_sample", "chain_cycles", "c1"])
def test_tip_collection_variants_agree(monkeypatch, name):
    """CollectTips (:346-357) literally (MCAAT_TIPS_LITERAL: EdgeOutdegree of every item) and from the dead ends' side
    (k_tips_dead_ends: IncomingEdges of every node without a valid item): same count, same pruned graph."""
    if name == "chain_cycles":
        from test_device_logic_host import chain_cycle_graph
        g = chain_cycle_graph()
    else:
        g = workload(name)
    dg = DeviceGraph(g)
    a = dg.prune()
    inv_a = dg.invalid()
    dg.reset()
    monkeypatch.setenv("MCAAT_TIPS_LITERAL", "1")
    b = dg.prune()
    assert a == b and a["n_tips_before"] > 0
    assert np.array_equal(inv_a, dg.invalid())
    dg.close()
    _wl.pop("c1", None)


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


def test_sample_workload_read_mapping_matches_oracle(tmp_path):
    """SURVEY 8(f) N3 at 11.3 M items: 3 000 synthetic reads (both strands, substitution errors) through
    mg_map_reads — index decoded from the graph files — against the oracle's restatement of
    get_read_from_sequence, forward and reverse-complemented."""
    import synth_sdbg as S
    gs, _ = S.make_workload("c2_sample", device="cuda", want_labels=True)
    prefix = str(tmp_path / "graph")
    S.write_megahit_files(gs, prefix)
    og = O.OracleGraph.load(prefix)
    dg = DeviceGraph.from_files(prefix)
    a = run(dg, 20, 27, 77)
    cyc = np.unique(np.concatenate([p for _, ps in a.searched for p in ps])).astype(np.uint64)
    chars, off = S.sample_reads("c2_sample", 3000, seed=11)
    seqs = [bytes(chars[int(off[i]):int(off[i + 1])]) for i in range(3000)]
    # half of the reads are forced to pass the filter: a random half of all ids joins the node set
    big = np.union1d(cyc, np.nonzero(np.random.default_rng(2).random(dg.n) < 0.5)[0].astype(np.uint64))
    for ids, rev in ((cyc, False), (big, False), (big, True)):
        want = og.map_reads(seqs, ids, reverse=rev)
        got_off, got_ids = dg.map_reads((chars, off), ids, reverse=rev)
        assert [w.size for w in want] == list(np.diff(got_off.astype(np.int64)))
        assert np.array_equal(np.concatenate(want), got_ids)
    assert got_ids.size > 100000


def test_cpp_drop_in_equals_the_unmodified_reference_binary_on_the_sample(tmp_path):
    """The compiled UNMODIFIED reference (oracle/_ref/ref_cyclefinder, threads=1) and the C++ drop-in
    (mcaat_b200/host/_build/shim_cyclefinder) on the 11.3 M-item sample written in MEGAHIT's on-disk format: the two
    dumps — host SDBG invalid bitmap, ordered start lists, raw results, unordered_map iteration order — are
    byte-identical, with the graph passed through the SDBG getters and decoded on the device from the files."""
    import subprocess
    import synth_sdbg as S
    ref = O.ref_binary("ref_cyclefinder")
    shim = os.path.join(ROOT, "mcaat_b200", "host", "_build", "shim_cyclefinder")
    if ref is None or not os.path.exists(shim):
        pytest.skip("reference / drop-in binaries not built (they need the reference sources at build time)")
    gs, _ = S.make_workload("c2_sample", device="cuda", want_labels=True)
    prefix = str(tmp_path / "graph")
    S.write_megahit_files(gs, prefix)
    del gs
    want_path = str(tmp_path / "ref.dump")
    subprocess.run([ref, prefix, want_path, "1", "20", "27", "77"], check=True, capture_output=True)
    want = open(want_path, "rb").read()
    assert len(want) > 1_000_000
    for ingest in ("0", "1"):
        out = str(tmp_path / f"shim_{ingest}.dump")
        subprocess.run([shim, prefix, out, "1", "20", "27", "77"], check=True, capture_output=True,
                       env=dict(os.environ, MCAAT_GPU_INGEST_FILES=ingest))
        assert open(out, "rb").read() == want, f"ingest={ingest}"


@pytest.mark.skipif(not os.environ.get("MCAAT_TEST_C2_FULL"), reason="the reference needs minutes of one host core at 224.7 M "
                    "items; run with MCAAT_TEST_C2_FULL=1 (log under profiles/)")
@pytest.mark.parametrize("thr,mn,mx", [(20, 27, 77), (10, 20, 100)])
def test_cpp_drop_in_equals_the_unmodified_reference_binary_at_full_size(tmp_path_factory, thr, mn, mx):
    """The headline workload itself — c2, 224.7 M items, what bench.py times — against the compiled UNMODIFIED
    reference at threads=1: the two dumps (host SDBG invalid bitmap after the prune, ordered start lists per bucket, raw
    per-start cycle sequences, unordered_map iteration order) are byte-identical, at the default setting and at a
    stress setting of BASELINE.json configs[4].  The Python binding's result on the same graph equals the dump too."""
    import hashlib
    import subprocess
    import time
    import synth_sdbg as S
    ref = O.ref_binary("ref_cyclefinder")
    shim = os.path.join(ROOT, "mcaat_b200", "host", "_build", "shim_cyclefinder")
    if ref is None or not os.path.exists(shim):
        pytest.skip("reference / drop-in binaries not built (they need the reference sources at build time)")
    d = _wl.get("c2_files")
    if d is None:
        d = str(tmp_path_factory.mktemp("c2_files"))
        gs, _ = S.make_workload("c2", device="cuda", want_labels=True)
        S.write_megahit_files(gs, os.path.join(d, "graph"))
        del gs
        _wl["c2_files"] = d
    prefix = os.path.join(d, "graph")
    want_path = os.path.join(d, f"ref_{thr}_{mn}_{mx}.dump")
    t0 = time.time()
    out = subprocess.run([ref, prefix, want_path, "1", str(thr), str(mn), str(mx)], check=True, capture_output=True, text=True).stdout
    t_ref = time.time() - t0
    print([ln for ln in out.splitlines() if ln.startswith("REF_JSON ")][-1])
    want = open(want_path, "rb").read()
    for ingest in ("1", "0"):
        got_path = os.path.join(d, f"shim_{thr}_{mn}_{mx}_{ingest}.dump")
        t0 = time.time()
        out = subprocess.run([shim, prefix, got_path, "1", str(thr), str(mn), str(mx)], check=True, capture_output=True, text=True,
                             env=dict(os.environ, MCAAT_GPU_INGEST_FILES=ingest)).stdout
        t_shim = time.time() - t0
        print(f"ingest={ingest}", [ln for ln in out.splitlines() if ln.startswith("REF_JSON ")][-1])
        got = open(got_path, "rb").read()
        print(f"[c2 full size, thr={thr} bounds {mn}-{mx}, ingest={ingest}] reference process {t_ref:.1f} s, drop-in process {t_shim:.1f} s, "
              f"dump {len(want)} bytes, sha256 ref {hashlib.sha256(want).hexdigest()[:16]} drop-in {hashlib.sha256(got).hexdigest()[:16]}")
        assert got == want, f"ingest={ingest}"
        os.remove(got_path)
    # the Python binding (what bench.py calls) on the array-built graph of the same workload
    ref_run = O.parse_dump(want_path)
    os.remove(want_path)
    got_run = run(DeviceGraph(workload("c2")), thr, mn, mx)
    assert O.same_run(ref_run, got_run) == []


def test_full_size_read_mapping_properties(monkeypatch):
    """N3 at 224.7 M items (index attached to the array-built graph): kept reads carry len - k + 1 ids; mapping the
    reverse complement of a batch with reverse=True gives the ids of the batch itself (reverse_pair_ends_sequence is
    an involution, reference src/reads.cpp:20-31); the same k-mer always maps to the same id; repeatable."""
    import synth_sdbg as S
    gs, _ = S.make_workload("c2", device="cuda", want_labels=True)
    dg = DeviceGraph(gs)
    dg.set_kmer_index(gs["tip"], *S.kmer_index_arrays(gs))
    n_reads, L, k = 20000, 150, gs["k"]
    chars, off = S.sample_reads("c2", n_reads, seed=5)
    ids_all = np.nonzero(np.random.default_rng(3).random(dg.n) < 0.5)[0].astype(np.uint64)
    o1, i1 = dg.map_reads((chars, off), ids_all)
    lens = np.diff(o1.astype(np.int64))
    assert set(np.unique(lens)) <= {0, L - k + 1} and (lens > 0).sum() > n_reads // 2
    null = np.uint64(2**64 - 1)
    assert (i1 != null).mean() > 0.9 and (i1[i1 != null] < dg.n).all()
    comp = np.arange(256, dtype=np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    rc = comp[chars.reshape(n_reads, L)[:, ::-1]].reshape(-1)
    o2, i2 = dg.map_reads((rc, off), ids_all, reverse=True)
    assert np.array_equal(o1, o2) and np.array_equal(i1, i2)
    # identical k-mers -> identical ids: duplicate the batch shifted by one read
    o3, i3 = dg.map_reads((np.concatenate([chars[L:], chars[:L]]), off), ids_all)
    first = i1[int(o1[1]):int(o1[2])]
    assert np.array_equal(first, i3[int(o3[0]):int(o3[1])])
    o4, i4 = dg.map_reads((chars, off), ids_all)
    assert np.array_equal(i1, i4)
    # the Backward^8 shortcut tables change nothing: same ids from the literal search
    monkeypatch.setenv("MCAAT_KMER_JUMP", "0")
    lit = DeviceGraph(gs)
    lit.set_kmer_index(gs["tip"], *S.kmer_index_arrays(gs))
    o5, i5 = lit.map_reads((chars, off), ids_all)
    assert np.array_equal(o1, o5) and np.array_equal(i1, i5)


def test_full_size_properties():
    check_full_size_properties("c2")


@pytest.mark.skipif(not os.environ.get("MCAAT_TEST_C4"), reason="1.1 G-item graph: ~100 GB of HBM for the generator and "
                    "minutes of box time; run with MCAAT_TEST_C4=1 (results under profiles/)")
def test_c4_properties():
    """BASELINE.json configs[3]: the ~1 B-item SDBG of the high-diversity community, same properties."""
    check_full_size_properties("c4")


def check_full_size_properties(name):
    g = workload(name)
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
