"""The product's device functions (mcaat_b200/csrc/sdbg_dev.cuh, search_dev.cuh — the code the CUDA
kernels execute), compiled for the host and checked against the golden reference dumps and the
oracle.  Runs without a GPU; the -m gpu tests repeat the same comparisons through the C ABI."""
import numpy as np
import pytest

import emu_py as E
import golden_util as G
import oracle_py as O

_cache = {}


def emu(name):
    if name not in _cache:
        _cache[name] = G.pack_raw(G.load_graph_npz(name))
    return E.EmuGraph(_cache[name])


def run_emu(g, thr, mn, mx, window=64, lock_log=13, reverse_peel=False):
    """Whole driver on the emulated device functions -> oracle_py.Run"""
    st = g.prune(reverse_order=reverse_peel)
    r = O.Run()
    r.n_items = g.n
    r.invalid = g.invalid()
    cand = g.candidates(thr)
    passed = []
    for c in cand:
        code = g.dls(c, mx)[0]
        if code == -1:  # tier-1 set full -> exact bitmap tier, as run_dls (starts.cu) does
            code = g.dls_bitmap(c, mx)
        assert code in (0, 1)
        if code:
            passed.append(int(c))
    buckets = {}
    for s in passed:
        buckets.setdefault(g.bucket(s), []).append(s)
    r.buckets = [(k, np.array(buckets[k], np.uint64)) for k in sorted(buckets, reverse=True)]
    starts = r.start_list()
    r.searched, stats = g.find_cycles(starts, mn, mx, window=window, lock_log=lock_log)
    r.stats = dict(st, n_candidates=len(cand), **stats)
    return r


def test_rank_select_round_trips():
    assert emu("tiny").rank_select_bad() == 0


def test_navigation_matches_reference_api():
    z = np.load(G.GOLD + "/tiny.nav.npz")
    od, idg, o, i_ = emu("tiny").navigate()
    assert np.array_equal(od, z["od"]) and np.array_equal(idg, z["id"])
    assert np.array_equal(o.astype(np.int64), z["o"]) and np.array_equal(i_.astype(np.int64), z["i"])


@pytest.mark.parametrize("name", G.nav_digest_cases())
def test_navigation_matches_reference_api_on_every_item_of_the_larger_graphs(name):
    want = G.nav_expected(name)
    od, idg, o, i_ = emu(name).navigate()
    mult = G.load_graph_npz(name)["mult"]
    assert od.size == want["n"]
    assert G.nav_digests(od, idg, mult, o, i_) == want["sha256"]


@pytest.mark.parametrize("name,thr,mn,mx,dump", G.cases(),
                         ids=lambda v: str(v) if not isinstance(v, str) or "/" not in v else "")
def test_device_logic_matches_reference_dump(name, thr, mn, mx, dump):
    ref = O.parse_dump(dump)
    got = run_emu(emu(name), thr, mn, mx)
    assert O.same_run(ref, got) == []


@pytest.mark.parametrize("name,thr,mn,mx,dump", G.cases(),
                         ids=lambda v: str(v) if not isinstance(v, str) or "/" not in v else "")
def test_hash_map_search_matches_reference_dump(name, thr, mn, mx, dump):
    """enum_run (the overflow fallback of the extracted-ball search) on its own"""
    E.set_algo(0)
    try:
        got = run_emu(emu(name), thr, mn, mx)
    finally:
        E.set_algo()
    assert O.same_run(O.parse_dump(dump), got) == []


@pytest.mark.parametrize("name,thr,mn,mx,dump", G.cases(),
                         ids=lambda v: str(v) if not isinstance(v, str) or "/" not in v else "")
def test_node_level_ball_search_matches_reference_dump(name, thr, mn, mx, dump):
    """ball_dfs (the literal search node by node, kept as the yardstick of the contracted search) on its own"""
    E.set_algo(1)
    try:
        got = run_emu(emu(name), thr, mn, mx)
    finally:
        E.set_algo()
    assert O.same_run(O.parse_dump(dump), got) == []


@pytest.mark.parametrize("window", [1, 3, 1000])
def test_speculative_window_is_order_exact(window):
    name, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "repeatrich" and c[1] == 20][0]
    ref = O.parse_dump(dump)
    got = run_emu(emu(name), thr, mn, mx, window=window, reverse_peel=True)
    assert O.same_run(ref, got) == []


def test_small_arena_overflow_is_rerun_exactly():
    name, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "repeatrich" and c[1] == 5][0]
    ref = O.parse_dump(dump)
    E.set_algo(2, ball_log=7)  # 64-node balls: most searches overflow into the hash-map search ...
    try:
        got = run_emu(emu(name), thr, mn, mx, window=32, lock_log=6)  # ... whose first arena overflows too
    finally:
        E.set_algo()
    assert O.same_run(ref, got) == []
    assert got.stats["overflow_reruns"] > 0


@pytest.mark.parametrize("name,thr", [("strain", 10), ("repeatrich", 5), ("biallelic", 10), ("tiny", 5)])
def test_dls_on_the_contracted_ball_equals_the_literal_search(name, thr):
    """dball_build / dball_contract / dball_dfs (the tier of the long admission searches) against dls_run on the graph
    itself: verdict AND expansion count, at depth limits that cut chains in the middle; small balls overflow"""
    g = emu(name)
    g.prune()
    cand = [int(c) for c in g.candidates(thr)]
    rng = np.random.default_rng(7)
    if len(cand) > 250:
        cand = [cand[i] for i in sorted(rng.choice(len(cand), 250, replace=False))]
    found = contracted = overflow = 0
    for limit in (77, 12, 33, 120, 1000):
        for c in cand:
            want, want_exp = g.dls(c, limit, log2cap=16)
            assert want in (0, 1)
            got, got_exp, chains = g.dls_ball(c, limit)
            assert (got, got_exp) == (want, want_exp), (name, c, limit)
            found += got
            contracted += chains > 0
            if g.dls_ball(c, limit, ball_log=6)[0] == -1:
                overflow += 1
    assert found > 0 and contracted > 0 and overflow > 0


@pytest.mark.parametrize("name", ["tiny", "strain", "repeatrich"])
def test_peel_variants_agree(name):
    """splitter-based chain resolution (forward / reverse item order) == exhaustive pointer jumping (2, 3);
    with and without (4..7) the re-classification of unary branch nodes; with the relink pass (8..11) instead of
    following re-classified terminals where they are asked for (the default, as on the device)"""
    inv, tips = [], []
    for mode in range(12):  # odd modes also collect the tips literally, even ones from the dead ends' side
        g = emu(name)
        tips.append(g.prune(reverse_order=mode)["n_tips_before"])
        inv.append(g.invalid())
    for other in inv[1:]:
        assert np.array_equal(inv[0], other)
    assert len(set(tips)) == 1 and tips[0] > 0


def test_dls_tiers_agree():
    g = emu("strain")
    g.prune()
    cand = g.candidates(10)
    assert len(cand) > 50
    full = 0
    for c in cand[:400]:
        small, _ = g.dls(c, 77, log2cap=6)
        exact = g.dls_bitmap(c, 77)
        big, _ = g.dls(c, 77, log2cap=14)
        assert big == exact
        if small == -1:
            full += 1
        else:
            assert small == exact
    assert full > 0  # the tier-1 overflow path is exercised


def test_synthetic_community_matches_oracle():
    """bench-style input (tools/synth_sdbg.py 'mini'): nested bubbles, tandem repeats, error tips"""
    import synth_sdbg as S
    g, _ = S.make_workload("mini")
    n = g["n"]
    w8 = g["w"].view(np.uint8)
    w = np.empty(w8.size * 2, np.uint8)
    w[0::2] = w8 & 15
    w[1::2] = w8 >> 4
    last = np.unpackbits(g["last"].view(np.uint8), bitorder="little")[:n]
    tip = np.unpackbits(g["tip"].view(np.uint8), bitorder="little")[:n]
    want = O.OracleGraph.from_arrays(w[:n], last, tip, g["mult"], g["f"], g["k"]).run(20, 27, 77)
    got = run_emu(E.EmuGraph(g), 20, 27, 77, window=256, reverse_peel=True)
    assert O.same_run(want, got) == []
    assert got.stats["n_peeled"] > 1000


def chain_cycle_graph(n_loops=120, seed=3):
    """Closed loops of chain nodes without entry or exit (circular units of 30..90 bases: what a tandem repeat
    leaves when the multiplicity filter removes its flanks), loops entered by a tail (rho shape), and linear
    contigs whose ends are tips, so that the peel has seeds.  About half of the loops hold no 1-in-64 splitter."""
    import torch

    import synth_sdbg as S
    rng = np.random.default_rng(seed)
    k = S.K
    seqs = []
    for i in range(n_loops):
        unit = rng.integers(0, 4, int(rng.integers(30, 91)))
        loop = np.concatenate([unit, unit[:k]])                       # all (k+1)-mers of the circular unit
        seqs.append(loop if i % 3 else np.concatenate([rng.integers(0, 4, int(rng.integers(5, 200))), unit, loop]))
    seqs += [rng.integers(0, 4, int(rng.integers(60, 400))) for _ in range(40)]
    lut = np.frombuffer(b"ACGT", np.uint8)
    e = torch.cat([S.windows(S.seq_codes(lut[q]), k + 1) for q in seqs])
    u, c = S.count_kmers(S.both_strands(e))
    return S.sdbg_from_kmers(u, c * 5)


def test_peel_cycles_of_chain_nodes_without_splitter():
    """chain_resolve_item recognises a splitter-free cycle of chain nodes itself (no hop-cap fallback to the
    exhaustive pointer jumping, which costs 34 passes over the graph — seen at 1.06 G items, workload c4);
    all peel variants and the oracle agree on the result"""
    g = chain_cycle_graph()
    n = g["n"]
    w8 = g["w"].view(np.uint8)
    w = np.empty(w8.size * 2, np.uint8)
    w[0::2] = w8 & 15
    w[1::2] = w8 >> 4
    last = np.unpackbits(g["last"].view(np.uint8), bitorder="little")[:n]
    tip = np.unpackbits(g["tip"].view(np.uint8), bitorder="little")[:n]
    og = O.OracleGraph.from_arrays(w[:n], last, tip, g["mult"], g["f"], g["k"])
    want = og.run(2, 27, 77)
    inv = []
    for mode in range(10):
        eg = E.EmuGraph(g)
        st = eg.prune(reverse_order=mode)
        assert st["walk_cap_hit"] == 0
        inv.append(eg.invalid())
    assert st["n_peeled"] > 1000
    for other in inv:
        assert np.array_equal(other, want.invalid)
    got = run_emu(E.EmuGraph(g), 2, 27, 77)
    assert O.same_run(want, got) == []
    assert len(got.searched) > 20  # the loops entered by a tail have a start node (in-degree 2): their cycles are found, too


@pytest.mark.parametrize("name,key,params,k", [c for c in G.region_cases() if c[0] in ("tiny", "biallelic")],
                         ids=lambda v: str(v) if isinstance(v, (str, int)) else "")
def test_region_extraction_matches_reference(name, key, params, k):
    """region_expand_word / region_keep_word (the code k_region_level / k_region_keep run)
    against the UNMODIFIED reference keep_crispr_regions_extended_by_k; both scan directions."""
    seeds = G.region_seeds(name, params)
    want = None
    for reverse in (False, True):
        g = emu(name)
        if params is not None:
            g.prune()
        before = g.invalid()
        st = g.keep_regions(seeds, k, reverse=reverse)
        want = G.region_expected(name, key, g.n)
        assert np.array_equal(g.invalid(), want)
        assert st["n_invalidated"] == int((want & ~before).sum()) and st["n_kept"] == int((~want).sum())
        assert st["levels"] <= max(k, 0)


@pytest.mark.parametrize("name", ["tiny", "biallelic"])
def test_read_mapping_matches_reference(tmp_path, name):
    """index_binary_search / read_ends / read_id_at (kmer_dev.cuh, the code k_read_ends and
    k_read_kmers run) against the UNMODIFIED reference get_read_from_sequence (tests/golden/<name>.reads.npz)."""
    og = O.OracleGraph.load(G.unpack_sdbg_files(name, str(tmp_path)))
    labels, first, second = og.kmer_index()
    seqs, variants = G.reads_case(name)
    for use_jump in (False, True):  # literal search / with the Backward^8 shortcut tables
        g = emu(name)
        g.set_kmer_index(og.k, _cache[name]["tip"], labels, first, second, use_jump=use_jump)
        if use_jump:
            st = g.kmer_index_stats()
            assert st["items_with_jump"] > 0.9 * g.n and st["buckets_ok"] > 30000
        for key, (ids, rev, off, flat) in variants.items():
            got = g.map_reads(seqs, ids, reverse=rev)
            assert [x.size for x in got] == list(np.diff(off)), (key, use_jump)
            assert np.array_equal(np.concatenate(got), flat), (key, use_jump)


@pytest.mark.parametrize("name,thr", [("repeatrich", 5), ("biallelic", 10), ("strain", 10), ("tiny", 5)])
def test_contracted_search_equals_node_level_search(name, thr):
    """sball_dfs (chains of the ball as single records) against ball_dfs (node by node) per start node: status, raw
    cycle sequence, and the step / expansion counters the reference's loop would have reached — with length bounds
    that cut chains in the middle and with lowered step caps (the cap fires inside a chain)."""
    g = emu(name)
    g.prune()
    cand = [int(c) for c in g.candidates(thr)]
    rng = np.random.default_rng(5)
    if len(cand) > 150:
        cand = [cand[i] for i in sorted(rng.choice(len(cand), 150, replace=False))]
    n_cut = n_cap = n_bound = n_contracted = 0
    for mn, mx, cap in [(27, 77, 0), (5, 33, 0), (0, 12, 0), (20, 100, 0), (27, 77, 1000), (27, 77, 137), (10, 60, 5000)]:
        E.set_search_limits(cap, 0)
        try:
            for s in cand:
                a, ca = g.search_one(s, mn, mx, 1)
                b, cb = g.search_one(s, mn, mx, 2)
                for key in ("status", "n_cycles", "n_nodes", "steps", "expansions", "n_ball"):
                    if key == "expansions" and a["status"] == 2:
                        continue  # cap inside a chain: the statistic counts the whole chain (results are unaffected)
                    assert a[key] == b[key], (name, s, mn, mx, cap, key, a, b)
                assert len(ca) == len(cb) and all(np.array_equal(x, y) for x, y in zip(ca, cb)), (name, s, mn, mx, cap)
                n_cap += a["status"] == 2
                n_bound += a["status"] == 1
                n_contracted += b["n_chains"] < b["n_ball"]
        finally:
            E.set_search_limits(0, 0)
    assert n_contracted > 0
    if name == "repeatrich":
        assert n_cap > 0
    if name == "biallelic":
        assert n_bound > 0
