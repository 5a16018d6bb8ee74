"""Parity tests proper: the CUDA path, called through the C ABI (libmcaat_gpu.so), against the
golden dumps of the UNMODIFIED reference CycleFinder at threads=1 and against the oracle.
Bit-exact: invalid bitmap, ordered start lists per bucket, raw per-start cycle sequences."""
import numpy as np
import pytest

import golden_util as G
import oracle_py as O
from mcaat_b200 import CycleFinder, DeviceGraph, Settings, gpu

pytestmark = pytest.mark.gpu

_packed = {}


def packed(name):
    if name not in _packed:
        _packed[name] = G.pack_raw(G.load_graph_npz(name))
    return _packed[name]


def run_gpu(g, thr, mn, mx):
    s = Settings(quiet=True)
    s.sdbg = g
    s.cycle_finder_settings.threshold_multiplicity = thr
    s.cycle_finder_settings.cycle_min_length = mn
    s.cycle_finder_settings.cycle_max_length = mx
    cf = CycleFinder(s)
    r = O.Run()
    r.n_items = g.n
    r.invalid = g.invalid()
    r.buckets = cf.start_nodes_chunked
    r.searched = list(cf.results.items())
    r.stats = cf.stats
    return r


def test_navigation_matches_reference_api():
    z = np.load(G.GOLD + "/tiny.nav.npz")
    g = DeviceGraph(packed("tiny"))
    od, idg, o, i_, m = g.navigate()
    assert np.array_equal(od, z["od"]) and np.array_equal(idg, z["id"])
    assert np.array_equal(o.astype(np.int64), z["o"]) and np.array_equal(i_.astype(np.int64), z["i"])
    assert np.array_equal(m, z["m"])
    before = gpu.launch_count()
    g.navigate(0, 100)
    assert gpu.launch_count() > before


@pytest.mark.parametrize("name", G.nav_digest_cases())
def test_navigation_matches_reference_api_on_every_item_of_the_larger_graphs(name):
    """degrees, neighbour ids and multiplicity of every item against digests of the unmodified reference API's answers"""
    want = G.nav_expected(name)
    od, idg, o, i_, m = DeviceGraph(packed(name)).navigate()
    assert od.size == want["n"]
    assert G.nav_digests(od, idg, m, o, i_) == want["sha256"]


@pytest.mark.parametrize("name,thr,mn,mx,dump", G.cases(),
                         ids=lambda v: str(v) if not isinstance(v, str) or "/" not in v else "")
def test_cuda_matches_reference_dump(name, thr, mn, mx, dump):
    ref = O.parse_dump(dump)
    got = run_gpu(DeviceGraph(packed(name)), thr, mn, mx)
    assert O.same_run(ref, got) == []


def test_prune_stats_match_oracle():
    raw = G.load_graph_npz("strain")
    og = O.OracleGraph.from_arrays(raw["w"], raw["last"], raw["tip"], raw["mult"], raw["f"], raw["k"])
    want = og.run(20, 27, 77).stats
    g = DeviceGraph(packed("strain"))
    st = g.prune()
    for k in ("n_tips_before", "n_mult1", "n_tips_after", "n_valid"):
        assert st[k] == want[k], k
    cand = g.candidates(20)
    assert cand.size == want["n_candidates"]
    ok, exp = g.dls(cand, 77)
    assert exp == want["dls_expansions"]


def test_reset_gives_identical_second_run():
    g = DeviceGraph(packed("repeatrich"))
    a = run_gpu(g, 20, 27, 77)
    g.reset()
    b = run_gpu(g, 20, 27, 77)
    assert O.same_run(a, b) == []


def test_small_arenas_take_the_exact_fallback_tiers(monkeypatch):
    """tier-1 set overflow -> tier-2 hash set -> bitmap tier; ball overflow -> hash-map search alone, whose
    first arena overflows as well; few workers"""
    monkeypatch.setenv("MCAAT_DLS_LOG2CAP", "6")
    monkeypatch.setenv("MCAAT_DLS_LOG2CAP2", "8")
    monkeypatch.setenv("MCAAT_ENUM_BALL_LOG", "7")
    monkeypatch.setenv("MCAAT_ENUM_LOCK_LOG", "8")
    monkeypatch.setenv("MCAAT_ENUM_WORKERS_PER_SM", "1")
    name, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "repeatrich" and c[1] == 5][0]
    ref = O.parse_dump(dump)
    got = run_gpu(DeviceGraph(packed(name)), thr, mn, mx)
    assert O.same_run(ref, got) == []


@pytest.mark.parametrize("switches", [{"MCAAT_ENUM_CHAINS_GLOBAL": "1"}, {"MCAAT_ENUM_SMEM_PER_WORKER": "2048"},
                                      {"MCAAT_ENUM_INORDER": "1"}, {"MCAAT_ENUM_WORKERS_PER_SM": "1"},
                                      {"MCAAT_ENUM_WORKERS_PER_SM": "8", "MCAAT_DLS_WORKERS_PER_SM": "3", "MCAAT_DLS_LOG2CAP": "6"},
                                      {"MCAAT_DLS_STREAM": "2", "MCAAT_DLS_LOG2CAP": "6"}, {"MCAAT_DLS_STREAM": "0", "MCAAT_DLS_LOG2CAP": "6"},
                                      {"MCAAT_DLS_STREAM": "2", "MCAAT_DLS_LOG2CAP": "6", "MCAAT_ENUM_BALL_LOG": "7"},
                                      {"MCAAT_PEEL_RELINK": "1"}],
                         ids=lambda d: "+".join(k[6:].lower() for k in d))
@pytest.mark.parametrize("name,thr", [("repeatrich", 5), ("biallelic", 20), ("strain", 10)])
def test_search_variants_agree_with_the_reference(monkeypatch, switches, name, thr):
    """the contracted-ball search with its records in the worker's arena instead of shared memory (the default) — the
    placement also taken when a contracted ball has more chains than the shared-memory share holds (forced with a
    2 KB share, which moves the path levels to the arena as well); every commit in ticket order instead of independent
    tickets committing ahead of their turn; few workers; more enumeration workers per SM than the default with fewer
    admission-search workers, every admission search past 32 nodes on the ball tier; the ball tier fed by the first
    tier while it runs (forced: the goldens have too few candidates for the default to choose it) and after it; the same with
    64-node arenas, so that balls overflow the ball tier (bitmap tier) and the enumeration's arenas (hash-map search);
    the peel with its relink pass (the default leaves re-classified terminals to terminal_of)"""
    for k, v in switches.items():
        monkeypatch.setenv(k, v)
    _, _, mn, mx, dump = [c for c in G.cases() if c[0] == name and c[1] == thr][0]
    ref = O.parse_dump(dump)
    got = run_gpu(DeviceGraph(packed(name)), thr, mn, mx)
    assert O.same_run(ref, got) == []
    if name == "biallelic":
        assert got.stats["n_bound_hits"] >= 1


def test_cluster_bound_case():
    name, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "biallelic"][0]
    got = run_gpu(DeviceGraph(packed(name)), thr, mn, mx)
    assert got.stats["n_bound_hits"] >= 1


def test_visited_equals_union_of_cycle_nodes():
    g = DeviceGraph(packed("repeatrich"))
    r = run_gpu(g, 20, 27, 77)
    want = np.zeros(g.n, bool)
    for _, paths in r.searched:
        for p in paths:
            want[p.astype(np.int64)] = True
    assert np.array_equal(g.visited(), want)


def test_empty_and_degenerate_inputs():
    g = DeviceGraph(packed("tiny"))
    g.prune()
    searched, st = g.find_cycles([], 27, 77)
    assert searched == [] and st["n_cycles"] == 0
    ok, _ = g.dls([], 77)
    assert ok.size == 0
    with pytest.raises(gpu.MgError):
        g.find_cycles([g.n + 5], 27, 77)
    # a start that is not on any cycle yields an empty entry, like results[start] = {} in the reference
    inv = g.invalid()
    v = int(np.nonzero(~inv)[0][0])
    searched, _ = g.find_cycles([v], 27, 77)
    assert len(searched) == 1 and searched[0][0] == v


def test_cpp_drop_in_cyclefinder_matches_reference_dump(tmp_path):
    """The C++ `CycleFinder` shim (mcaat_b200/host/cycle_finder_gpu.cpp, compiled against the reference's
    unchanged header) driven by the oracle harness main: its dump — invalid bitmap of the HOST SDBG after
    write-back, ordered start lists, raw results, unordered_map iteration order — must equal the dump of
    the unmodified reference byte for byte."""
    import gzip
    import os
    import subprocess
    exe = os.path.join(gpu.ROOT, "mcaat_b200", "host", "_build", "shim_cyclefinder")
    if not os.path.exists(exe):
        pytest.skip("shim_cyclefinder not built (needs the reference headers at build time)")
    prefix = G.unpack_sdbg_files("tiny", str(tmp_path))
    for ingest in ("0", "1"):  # graph through the SDBG getters / decoded on the device from the graph files
        env = dict(os.environ, MCAAT_GPU_INGEST_FILES=ingest)
        for thr, mn, mx in ((20, 27, 77), (10, 20, 100)):
            out = str(tmp_path / f"shim_{thr}_{ingest}.dump")
            subprocess.run([exe, prefix, out, "1", str(thr), str(mn), str(mx)], check=True, capture_output=True, env=env)
            want = gzip.open(os.path.join(G.GOLD, f"tiny.t{thr}_{mn}_{mx}.dump.gz"), "rb").read()
            assert open(out, "rb").read() == want


@pytest.mark.parametrize("name", ["tiny", "biallelic"])
def test_cpp_drop_in_prints_the_reference_progress_lines(tmp_path, name):
    """SURVEY 5 ("same text -> diffable logs"): the progress lines the reference's constructor writes to std::cout
    (src/cycle_finder.cpp:381,403,423,437,452,456,459-461,466,485-490) against the drop-in's, line for line, on the same
    graph at threads=1 — counts of tips, filtered items, valid edges, buckets and cycles included."""
    import os
    import subprocess
    shim = os.path.join(gpu.ROOT, "mcaat_b200", "host", "_build", "shim_cyclefinder")
    ref = O.ref_binary("ref_cyclefinder")
    if ref is None or not os.path.exists(shim):
        pytest.skip("reference / drop-in binaries not built (they need the reference sources at build time)")
    prefix = G.unpack_sdbg_files(name, str(tmp_path))
    env = {k: v for k, v in os.environ.items() if k != "MCAAT_REF_COUNT_STARTS"}
    lines = {}
    for exe in (ref, shim):
        out = subprocess.run([exe, prefix, "-", "1", "20", "27", "77"], check=True, capture_output=True, text=True, env=env).stdout
        lines[exe] = [ln for ln in out.splitlines() if not ln.startswith("REF_JSON ")]
    assert lines[ref] == lines[shim]
    assert any(ln.startswith("Cycle enumeration completed") for ln in lines[ref]) and len(lines[ref]) >= 8


@pytest.mark.parametrize("name", ["tiny", "biallelic"])
def test_crispr_arrays_file_is_identical_end_to_end(tmp_path, name):
    """SURVEY 8(f) N1 / north_star's second correctness criterion: the reference's release pipeline from the loaded graph
    to CRISPR_Arrays.txt (src/main.cpp:520-580: CycleFinder -> read mapping -> region split -> set cover -> spacer ordering
    -> CRISPRAnalyzer), every stage the reference's own unmodified translation unit, built three times:
      oracle/_ref/ref_pipeline          with the reference's src/cycle_finder.cpp
      host/_build/shim_pipeline         with the GPU CycleFinder drop-in
      host/_build/shim_pipeline_all     with the GPU drop-ins for CycleFinder, get_reads and get_crispr_regions_extended_by_k
    on the graph files of a golden case and the reads they were built from.  The output files are byte-identical, and so
    is everything the stages print (cycle order, spacer order, repeats, confidence values) except the wall times.
    kseq++ and cft, which this image lacks, are the same deterministic stand-ins (mcaat_b200/host/absent_deps) in all three."""
    import os
    import re
    import subprocess
    import sys
    sys.path.insert(0, os.path.join(gpu.ROOT, "tools"))
    import make_golden as MG
    ref = O.ref_binary("ref_pipeline")
    shims = [os.path.join(gpu.ROOT, "mcaat_b200", "host", "_build", b) for b in ("shim_pipeline", "shim_pipeline_all")]
    if ref is None or not all(os.path.exists(b) for b in shims):
        pytest.skip("reference / drop-in pipeline binaries not built (they need the reference sources at build time)")
    prefix = G.unpack_sdbg_files(name, str(tmp_path))
    subprocess.run([sys.executable, os.path.join(gpu.ROOT, "tools", "gen_reads.py"), "--out", str(tmp_path)] + MG.CASES[name][0],
                   check=True, capture_output=True)
    outs, logs = {}, {}
    for exe in [ref] + shims:
        out = str(tmp_path / (os.path.basename(exe) + ".CRISPR_Arrays.txt"))
        p = subprocess.run([exe, prefix, str(tmp_path / "r1.fq"), str(tmp_path / "r2.fq"), out, "1", "20", "27", "77"],
                           check=True, capture_output=True, text=True, env=dict(os.environ, MCAAT_GPU_GRAPH_PREFIX=prefix))
        outs[exe] = open(out, "rb").read()
        logs[exe] = [ln for ln in p.stdout.replace(out, "<out>").splitlines() if "Time elapsed" not in ln]
    assert b"Number of Systems: " in outs[ref] and outs[ref].count(b"Number of Spacers: ") >= 2
    for exe in shims:
        assert outs[exe] == outs[ref], os.path.basename(exe)
        assert logs[exe] == logs[ref], os.path.basename(exe)
    assert int(re.search(r'"systems": (\d+)', "\n".join(logs[ref])).group(1)) >= 2


def test_collect_tips_on_the_device_matches_navigation():
    """mg_collect_tips (dead ends -> incoming edges -> compaction, all on the device) == {valid, out-degree 0} from the
    per-item navigation table, on the loaded graph and again after the prune"""
    for name in ("tiny", "strain"):
        g = DeviceGraph(packed(name))
        for stage in ("loaded", "pruned"):
            od = g.navigate()[0]
            want = np.nonzero(od == 0)[0].astype(np.uint64)
            got = g.collect_tips()
            assert np.array_equal(got, want), (name, stage)
            assert stage == "pruned" or got.size > 0
            g.prune()


def test_device_ingest_of_on_disk_sdbg(tmp_path):
    """the reference's own graph files (built by its SDBGBuild, 4 part files) decoded on
    the device give the same navigation tables and the same search result as the reference."""
    prefix = G.unpack_sdbg_files("tiny", str(tmp_path))
    g = DeviceGraph.from_files(prefix)
    raw = G.load_graph_npz("tiny")
    assert g.n == raw["w"].size
    z = np.load(G.GOLD + "/tiny.nav.npz")
    od, idg, o, i_, m = g.navigate()
    assert np.array_equal(od, z["od"]) and np.array_equal(idg, z["id"])
    assert np.array_equal(o.astype(np.int64), z["o"]) and np.array_equal(i_.astype(np.int64), z["i"])
    assert np.array_equal(m, z["m"])
    name, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "tiny" and c[1] == 20][0]
    assert O.same_run(O.parse_dump(dump), run_gpu(g, thr, mn, mx)) == []
    with pytest.raises(gpu.MgError):
        DeviceGraph.from_files(prefix + "_missing")
    # truncated part file: the decoder must notice, not read past the end
    with open(prefix + ".sdbg.3", "r+b") as f:
        f.truncate(100)
    with pytest.raises(gpu.MgError):
        DeviceGraph.from_files(prefix)


# ------------------------------------------------------------------ region extraction
@pytest.mark.parametrize("name,key,params,k", G.region_cases(), ids=lambda v: str(v) if isinstance(v, (str, int)) else "")
def test_region_extraction_matches_reference(name, key, params, k):
    """mg_keep_regions against the invalid bitmap the UNMODIFIED keep_crispr_regions_extended_by_k leaves
    (reference src/spacer_ordering.cpp:78-139; tests/golden/<name>.regions.npz)."""
    g = DeviceGraph(packed(name))
    if params is not None:
        g.prune()
    before = g.invalid()
    launches = gpu.launch_count()
    st = g.keep_regions(G.region_seeds(name, params), k)
    assert gpu.launch_count() > launches
    want = G.region_expected(name, key, g.n)
    assert np.array_equal(g.invalid(), want)
    assert st["n_invalidated"] == int((want & ~before).sum()) and st["n_kept"] == int((~want).sum())
    assert st["levels"] <= k


@pytest.mark.parametrize("qlog", [2, 6, 9])
def test_region_extraction_with_overflowing_queues(monkeypatch, qlog):
    """frontier queues of 4 / 64 / 512 ids: levels that do not fit spill into the bitmap form and the remaining
    levels run as bitmap passes — same set"""
    monkeypatch.setenv("MCAAT_REGION_QUEUE_LOG", str(qlog))
    for name, key, params, k in [c for c in G.region_cases() if c[0] == "biallelic" and c[3] in (7, 40, 30)]:
        g = DeviceGraph(packed(name))
        if params is not None:
            g.prune()
        st = g.keep_regions(G.region_seeds(name, params), k)
        assert np.array_equal(g.invalid(), G.region_expected(name, key, g.n)), (name, key)
        assert st["levels"] <= k


def test_region_extraction_edge_cases():
    g = DeviceGraph(packed("tiny"))
    g.prune()
    with pytest.raises(gpu.MgError):
        g.keep_regions([g.n], 3)
    valid = np.nonzero(~g.invalid())[0]
    # k = 0: only the seeds themselves survive; duplicates and an invalid seed are harmless
    bad = int(np.nonzero(g.invalid())[0][0])
    seeds = np.array([valid[5], valid[5], valid[9], bad], np.uint64)
    st = g.keep_regions(seeds, 0)
    assert st["n_kept"] == 2 and st["levels"] == 0
    assert set(np.nonzero(~g.invalid())[0]) == {int(valid[5]), int(valid[9])}
    # the navigation cache is stale now: the search entries refuse until mg_prune has run again
    with pytest.raises(gpu.MgError):
        g.find_cycles([int(valid[5])], 27, 77)
    # idempotent; no seeds at all -> nothing stays valid (the reference's loop invalidates every item)
    assert g.keep_regions(seeds, 0)["n_invalidated"] == 0
    st = g.keep_regions([], 5)
    assert st["n_kept"] == 0 and g.invalid().all()


def test_region_growth_is_monotone_and_saturates():
    """size-independent properties: kept(k) grows with k, a huge k stops at the component closure, and the
    closure is closed under valid in-/out-neighbours"""
    raw = packed("repeatrich")
    seeds = G.region_seeds("repeatrich", (20, 27, 77))
    kept = []
    for k in (0, 3, 20, 10**9):
        g = DeviceGraph(raw)
        g.prune()
        st = g.keep_regions(seeds, k)
        kept.append(st["n_kept"])
        if k == 10**9:
            assert st["levels"] < 100000
            od, idg, o, i_, _ = g.navigate()
            inv = g.invalid()
            for v in np.nonzero(~inv)[0][::50]:
                for x in list(o[v][:max(int(od[v]), 0)]) + list(i_[v][:max(int(idg[v]), 0)]):
                    assert not inv[int(x)]
    assert kept == sorted(kept) and kept[0] < kept[-1]


def test_host_invalid_bits_merge_into_the_device_graph():
    g = DeviceGraph(packed("tiny"))
    inv = g.invalid()
    extra = np.zeros(g.n, bool)
    picks = np.nonzero(~inv)[0][[3, 77, 500, -1]]
    extra[picks] = True
    extra[np.nonzero(inv)[0][:10]] = True  # already invalid on the device: not counted
    assert g.merge_invalid(extra) == picks.size
    assert np.array_equal(g.invalid(), inv | extra)
    assert g.merge_invalid(extra) == 0


def test_cpp_drop_in_regions_match_reference(tmp_path):
    """mcaat_b200/host/regions_gpu.cpp (the reference's signature) under the region harness main: the HOST
    SDBG's invalid bitmap after the call equals what the unmodified reference function leaves — on the
    device graph kept resident by the CycleFinder drop-in (`cycles`) and on a fresh upload (`nodes`)."""
    import os
    import subprocess
    exe = os.path.join(gpu.ROOT, "mcaat_b200", "host", "_build", "shim_regions")
    if not os.path.exists(exe):
        pytest.skip("shim_regions not built (needs the reference headers at build time)")
    prefix = G.unpack_sdbg_files("tiny", str(tmp_path))
    n = G.load_graph_npz("tiny")["w"].size
    seeds_file = str(tmp_path / "seeds.u64")
    G.region_seeds("tiny", None).astype("<u8").tofile(seeds_file)
    runs = [("cyc_t20_27_77_k7", ["7", "cycles", "20", "27", "77"]),
            ("cyc_t20_27_77_k150", ["150", "cycles", "20", "27", "77"]),
            ("nodes_k30", ["30", "nodes", seeds_file])]
    for key, tail in runs:
        out = str(tmp_path / (key + ".reg"))
        subprocess.run([exe, prefix, out] + tail, check=True, capture_output=True)
        raw = open(out, "rb").read()
        assert raw[:8] == b"MCAATREG"
        words = np.frombuffer(raw, "<u8", (n + 63) // 64, 40)
        got = np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(bool)
        assert np.array_equal(got, G.region_expected("tiny", key, n)), key


def test_valid_subgraph_matches_navigation():
    """mg_valid_subgraph (input of the reference's SCC split) against the per-item navigation tables: ascending
    valid ids; rows = valid out-neighbours in OutgoingEdges order, as positions in the id list"""
    g = DeviceGraph(packed("repeatrich"))
    g.prune()
    g.keep_regions(G.region_seeds("repeatrich", (20, 27, 77)), 40)
    ids, off, nbr = g.valid_subgraph()
    inv = g.invalid()
    assert np.array_equal(ids, np.nonzero(~inv)[0].astype(np.uint64))
    od, _, o, _, _ = g.navigate()
    deg = np.maximum(od[ids.astype(np.int64)], 0).astype(np.int64)
    assert np.array_equal(np.diff(off.astype(np.int64)), deg) and off[-1] == nbr.size
    want = np.concatenate([o[int(v)][:d] for v, d in zip(ids, deg)]) if nbr.size else np.zeros(0, np.uint64)
    assert np.array_equal(ids[nbr.astype(np.int64)], want)
    # everything invalid -> empty graph
    g.keep_regions([], 1)
    ids, off, nbr = g.valid_subgraph()
    assert ids.size == 0 and off.tolist() == [0] and nbr.size == 0


@pytest.mark.parametrize("name", ["tiny", "biallelic"])
def test_cpp_drop_in_scc_split_matches_reference(tmp_path, name):
    """get_crispr_regions_extended_by_k = keep + divide_graph_into_subgraphs (reference
    src/spacer_ordering.cpp:177-184) through the drop-ins: the harness output — invalid bitmap, components in
    Tarjan completion order, node order, iteration order of each Graph's unordered containers — equals the
    unmodified reference's byte for byte."""
    import glob
    import gzip
    import os
    import subprocess
    exe = os.path.join(gpu.ROOT, "mcaat_b200", "host", "_build", "shim_regions")
    if not os.path.exists(exe):
        pytest.skip("shim_regions not built (needs the reference headers at build time)")
    prefix = G.unpack_sdbg_files(name, str(tmp_path))
    seeds_file = str(tmp_path / "seeds.u64")
    G.region_seeds(name, None).astype("<u8").tofile(seeds_file)
    golds = sorted(glob.glob(os.path.join(G.GOLD, f"{name}.scc_*.reg.gz")))
    assert len(golds) == 3
    n_sub = []
    for gold in golds:
        key = os.path.basename(gold)[len(name) + 5:-len(".reg.gz")]
        if key.startswith("nodes_k"):
            tail = [key[7:], "nodes+scc", seeds_file]
        else:
            _, t, mn, mxk = key.split("_", 3)
            mx, k = mxk.split("_k")
            tail = [k, "cycles+scc", t[1:], mn, mx]
        out = str(tmp_path / (key + ".reg"))
        subprocess.run([exe, prefix, out] + tail, check=True, capture_output=True)
        want = gzip.open(gold, "rb").read()
        got = open(out, "rb").read()
        assert got == want, key
        at = want.index(b"MCAATSCC")
        n_sub.append(int(np.frombuffer(want, "<u8", 1, at + 8)[0]))
    assert max(n_sub) >= 2


# ------------------------------------------------------------------ read -> node ids
@pytest.mark.parametrize("name", ["tiny", "biallelic"])
@pytest.mark.parametrize("via", ["files", "arrays", "files-literal"])
def test_read_mapping_matches_reference(tmp_path, monkeypatch, name, via):
    """mg_map_reads against the UNMODIFIED reference get_read_from_sequence / reverse_pair_ends_sequence /
    IndexBinarySearch (src/reads.cpp:20-89, sdbg.h:141-207; tests/golden/<name>.reads.npz): real, mutated, random,
    N-containing and too-short sequences, forward and reverse-complemented, two node sets.  The k-mer index comes
    from the graph files (mg_graph_create_from_files) or is attached to an array-built graph."""
    prefix = G.unpack_sdbg_files(name, str(tmp_path))
    if via == "files-literal":  # without the Backward^8 shortcut tables
        monkeypatch.setenv("MCAAT_KMER_JUMP", "0")
    if via.startswith("files"):
        g = DeviceGraph.from_files(prefix)
    else:
        g = DeviceGraph(packed(name))
        with pytest.raises(gpu.MgError):
            g.map_reads([b"ACGT" * 30], [1])
        labels, first, second = O.OracleGraph.load(prefix).kmer_index()
        g.set_kmer_index(packed(name)["tip"], labels, first, second)
    seqs, variants = G.reads_case(name)
    for key, (ids, rev, off, flat) in variants.items():
        launches = gpu.launch_count()
        got_off, got_ids = g.map_reads(seqs, ids, reverse=rev)
        assert gpu.launch_count() > launches
        assert np.array_equal(got_off.astype(np.int64), off), key
        assert np.array_equal(got_ids, flat), key
    # the search path is untouched by the index: same result as the reference dump
    if via == "files" and name == "tiny":
        _, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "tiny" and c[1] == 20][0]
        assert O.same_run(O.parse_dump(dump), run_gpu(g, thr, mn, mx)) == []


def test_read_mapping_edge_cases(tmp_path):
    g = DeviceGraph.from_files(G.unpack_sdbg_files("tiny", str(tmp_path)))
    off, ids = g.map_reads([], [1, 2, 3])
    assert off.tolist() == [0] and ids.size == 0
    seqs, variants = G.reads_case("tiny")
    off, ids = g.map_reads(seqs, [])  # no cycle nodes: every read is dropped
    assert off.max() == 0 and ids.size == 0
    with pytest.raises(gpu.MgError):
        g.map_reads(seqs[:3], [g.n])


@pytest.mark.parametrize("name", ["tiny", "biallelic"])
def test_cpp_drop_in_read_mapping_matches_reference(tmp_path, name):
    """mcaat_b200/host/reads_gpu.cpp (map_sequences_all, the body of the get_reads drop-in) under the read harness
    main: same per-sequence node ids as the unmodified reference get_read_from_sequence."""
    import os
    import subprocess
    exe = os.path.join(gpu.ROOT, "mcaat_b200", "host", "_build", "shim_reads")
    if not os.path.exists(exe):
        pytest.skip("shim_reads not built (needs the reference headers at build time)")
    prefix = G.unpack_sdbg_files(name, str(tmp_path))
    seqs, variants = G.reads_case(name)
    sf = str(tmp_path / "seqs.txt")
    with open(sf, "wb") as f:
        f.write(b"\n".join(seqs) + b"\n")
    env = dict(os.environ, MCAAT_GPU_GRAPH_PREFIX=prefix)
    for key, (ids, rev, off, flat) in variants.items():
        nf = str(tmp_path / (key + ".u64"))
        np.asarray(ids, "<u8").tofile(nf)
        out = str(tmp_path / (key + ".rd"))
        subprocess.run([exe, prefix, out, sf, nf, "1" if rev else "0"], check=True, capture_output=True, env=env)
        raw = open(out, "rb").read()
        assert raw[:8] == b"MCAATRDS"
        a = np.frombuffer(raw, "<u8", offset=8)
        assert int(a[0]) == len(seqs)
        pos, got_off, got = 1, [0], []
        for _ in range(len(seqs)):
            c = int(a[pos])
            got.append(a[pos + 1:pos + 1 + c])
            pos += 1 + c
            got_off.append(got_off[-1] + c)
        assert got_off == off.tolist(), key
        assert np.array_equal(np.concatenate(got), flat), key
    # without the index source the drop-in refuses instead of falling back
    env.pop("MCAAT_GPU_GRAPH_PREFIX")
    r = subprocess.run([exe, prefix, out, sf, nf, "0"], capture_output=True, env=env)
    assert r.returncode != 0
