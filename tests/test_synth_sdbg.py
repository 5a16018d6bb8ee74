"""tools/synth_sdbg.py (the bench's input builder) against the reference's own graph producer:
given the (k+1)-mers of the reads that made a golden graph, it must return the identical arrays
(golden graphs = oracle/_ref/build_sdbg = reference src/sdbg_build.cpp + vendored MEGAHIT)."""
import argparse
import os

import numpy as np
import pytest

import gen_reads
import golden_util as G
import make_golden
import oracle_py as O
import synth_sdbg as S


def reads_of_case(name):
    """Re-create the seeded reads of a golden case (same arguments as tools/make_golden.py)."""
    argv = ["--out", "/tmp/unused"] + make_golden.CASES[name][0]
    # mirror gen_reads.main without writing files
    p = argparse.ArgumentParser()
    for a in gen_reads_argspec():
        p.add_argument(a[0], **a[1])
    args = p.parse_args(argv)
    rng = np.random.Generator(np.random.PCG64(args.seed))
    genomes, abund, _ = gen_reads.build_community(args, rng)
    r1, r2 = gen_reads.sample_reads(args, rng, genomes, abund)
    return np.concatenate([r1, r2])


def gen_reads_argspec():
    return gen_reads.ARGSPEC


@pytest.mark.parametrize("name", ["tiny", "strain"])
def test_builder_matches_reference_graph_producer(name):
    reads = reads_of_case(name)
    edges, counts = S.kmers_from_read_matrix(reads)
    got = S.sdbg_from_kmers(edges, counts, want_labels=True)
    want = G.pack_raw(G.load_graph_npz(name))
    assert got["n"] == want["n"]
    assert np.array_equal(got["f"], want["f"])
    for key in ("w", "last", "tip", "mult"):
        assert np.array_equal(got[key], want[key]), key


def test_written_files_load_like_the_reference_format(tmp_path):
    reads = reads_of_case("tiny")
    edges, counts = S.kmers_from_read_matrix(reads)
    got = S.sdbg_from_kmers(edges, counts, want_labels=True)
    prefix = str(tmp_path / "graph")
    S.write_megahit_files(got, prefix)
    a = O.OracleGraph.load(prefix).raw()
    b = G.load_graph_npz("tiny")
    for key in ("w", "last", "tip", "mult", "f"):
        assert np.array_equal(a[key], b[key]), key
    if O.ref_binary("ref_cyclefinder"):  # the UNMODIFIED reference accepts the files and agrees with its golden dump
        dump = str(tmp_path / "out.dump")
        O.run_reference(prefix, dump, 1, 20, 27, 77)
        ref = O.parse_dump(dump)
        gold = O.parse_dump(os.path.join(G.GOLD, "tiny.t20_27_77.dump.gz"))
        assert O.same_run(gold, ref) == []
