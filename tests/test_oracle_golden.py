"""The oracle (oracle/mcaat_oracle.c) against the pins made from the UNMODIFIED reference:
raw results of reference CycleFinder at threads=1 (src/cycle_finder.cpp:433-492) and the
reference SDBG navigation API (libs/megahit/src/sdbg/sdbg.h:328-357)."""
import numpy as np
import pytest

import golden_util as G
import oracle_py as O

_graphs = {}


def graph(name):
    if name not in _graphs:
        _graphs[name] = G.load_graph_npz(name)
    r = _graphs[name]
    return O.OracleGraph.from_arrays(r["w"], r["last"], r["tip"], r["mult"], r["f"], r["k"])


@pytest.mark.parametrize("name,thr,mn,mx,dump", G.cases(), ids=lambda v: str(v) if not isinstance(v, str) or "/" not in v else "")
def test_oracle_matches_reference_dump(name, thr, mn, mx, dump):
    ref = O.parse_dump(dump)
    got = graph(name).run(thr, mn, mx)
    assert O.same_run(ref, got) == []


def test_oracle_navigation_matches_reference_api():
    z = np.load(G.GOLD + "/tiny.nav.npz")
    g = graph("tiny")
    od, idg, o, i_ = g.nav_table()
    assert np.array_equal(od, z["od"]) and np.array_equal(idg, z["id"])
    assert np.array_equal(o.astype(np.int64), z["o"]) and np.array_equal(i_.astype(np.int64), z["i"])
    assert np.array_equal(g.raw()["mult"], z["m"])


@pytest.mark.parametrize("name", G.nav_digest_cases())
def test_oracle_navigation_matches_reference_api_on_every_item_of_the_larger_graphs(name):
    """the exhaustive table for the other golden graphs, pinned by digests of the reference's answers"""
    want = G.nav_expected(name)
    g = graph(name)
    od, idg, o, i_ = g.nav_table()
    assert od.size == want["n"]
    assert G.nav_digests(od, idg, g.raw()["mult"], o, i_) == want["sha256"]


@pytest.mark.parametrize("name,key,params,k", G.region_cases(), ids=lambda v: str(v) if isinstance(v, (str, int)) else "")
def test_oracle_keep_regions_matches_reference(name, key, params, k):
    """orc_keep_regions against the UNMODIFIED keep_crispr_regions_extended_by_k
    (reference src/spacer_ordering.cpp:78-139, oracle/_ref/ref_regions)."""
    g = graph(name)
    if params is not None:
        g.run(*params)
    before = g.invalid()
    gone, kept = g.keep_regions(G.region_seeds(name, params), k)
    want = G.region_expected(name, key, g.n)
    got = g.invalid()
    assert np.array_equal(got, want)
    assert gone == int((want & ~before).sum()) and kept == int((~want).sum())


@pytest.mark.parametrize("name", ["tiny", "biallelic"])
def test_oracle_read_mapping_matches_reference(tmp_path, name):
    """orc_index_binary_search / orc_get_read / orc_reverse_complement against the UNMODIFIED
    reference (src/reads.cpp:20-89, sdbg.h:141-207; oracle/_ref/ref_reads) — real, mutated, random, N-containing
    and too-short sequences, forward and reverse-complemented, two node sets."""
    g = O.OracleGraph.load(G.unpack_sdbg_files(name, str(tmp_path)))
    seqs, variants = G.reads_case(name)
    for key, (ids, rev, off, flat) in variants.items():
        got = g.map_reads(seqs, ids, reverse=rev)
        assert [x.size for x in got] == list(np.diff(off)), key
        assert np.array_equal(np.concatenate(got) if got else np.zeros(0, np.uint64), flat), key


def test_oracle_keep_regions_rejects_bad_ids():
    g = graph("tiny")
    with pytest.raises(ValueError):
        g.keep_regions([g.n], 3)


def test_file_loader_matches_raw_arrays(tmp_path):
    prefix = G.unpack_sdbg_files("tiny", str(tmp_path))
    a = O.OracleGraph.load(prefix).raw()
    b = G.load_graph_npz("tiny")
    for key in ("w", "last", "tip", "mult", "f"):
        assert np.array_equal(a[key], b[key]), key
    assert a["k"] == b["k"] == 23


def test_cluster_bound_case_is_exercised():
    """biallelic: the 500-cycle cluster bound (cycle_finder.cpp:162-165) must fire in the pin."""
    got = graph("biallelic").run(20, 27, 77)
    assert got.stats["n_bound_hits"] >= 1 and got.stats["n_cap_hits"] >= 1


def test_uset_order_model():
    """libstdc++ unordered_set<uint64_t> iteration order (SURVEY §4.5 probe)."""
    ids = np.array([1001, 1002, 1014, 27], np.uint64)
    out = np.zeros(8, np.uint64)
    n = O.lib().orc_uset_order(ids.ctypes.data, 4, out.ctypes.data)
    assert list(out[:n]) == [27, 1002, 1014, 1001]


def test_rank_select_edge_cases():
    """kmrns.h:287-292,313-320 edge cases."""
    g = graph("tiny")
    L = O.lib()
    n = g.n
    assert L.orc_rank_last(g.h, n) == -1
    assert L.orc_rank_last(g.h, -1) == 0
    tot = L.orc_rank_last(g.h, n - 1)
    assert L.orc_select_last(g.h, tot) == n
    assert L.orc_select_last(g.h, tot + 1) == -1
    assert L.orc_rank_last(g.h, L.orc_select_last(g.h, 5)) == 6


def test_reference_unit_tests_pass_against_the_stand_ins():
    """The reference's OWN unit tests (tests/test_reads.cpp: 5 of reverse_pair_ends_sequence; tests/test_spacer_ordering.cpp:
    7 of solve_min_cover_problem), compiled unmodified (oracle/Makefile: _ref/ref_unit_tests; a minimal gtest interface
    stands in for the absent googletest submodule) and run against the deterministic stand-ins for kseq++ and cft that the
    end-to-end CRISPR_Arrays.txt comparison uses (mcaat_b200/host/absent_deps)."""
    import subprocess
    exe = O.ref_binary("ref_unit_tests")
    if exe is None:
        pytest.skip("oracle/_ref/ref_unit_tests not built (needs the reference sources at build time)")
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "UNIT_TESTS 12 run, 0 failed expectations" in out.stdout
