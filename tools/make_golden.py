#!/usr/bin/env python3
"""Regenerates tests/golden/ (TEST INFRASTRUCTURE; needs /root/reference to have been compiled into
oracle/_ref by `make -C oracle`).

For each named case: seeded synthetic reads (tools/gen_reads.py) -> the reference's own SDBG builder
(oracle/_ref/build_sdbg = reference src/sdbg_build.cpp + vendored MEGAHIT, unmodified) -> the
UNMODIFIED reference CycleFinder at threads=1 (oracle/_ref/ref_cyclefinder) for each parameter set.
Stored per case:
  <case>.graph.npz      raw SDBG arrays (w nibbles packed 2/byte, last/tip bit-packed, mult u16, f[6], k)
  <case>.<params>.dump.gz   the reference's dump (layout: oracle/ref_harness/dump_format.h)
  <case>.nav.npz        (small cases) per-item degrees / neighbour ids through the reference SDBG API
  <case>.regions.npz    invalid bitmaps after the UNMODIFIED reference keep_crispr_regions_extended_by_k
                        (oracle/_ref/ref_regions): keys cyc_t<thr>_<min>_<max>_k<k> (seeds = all
                        cycles of that run) and seeds / nodes_k<k> (seeds = random ids, graph as loaded)
  <case>.scc_<key>.reg.gz   (cases with on-disk files) ref_regions output incl. the reference's SCC split
One case also keeps the on-disk MEGAHIT files (gzip'ed) to pin the file-format reader.
"""
import argparse
import gzip
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# name -> (gen_reads args, [(thr, min, max), ...], keep_nav, keep_files)
CASES = {
    "tiny": (["--seed", "1", "--genomes", "2", "--genome-len", "30000", "--arrays", "2",
              "--spacers-min", "8", "--spacers-max", "12", "--pairs", "6000"],
             [(20, 27, 77), (10, 20, 100)], True, True),
    "repeatrich": (["--seed", "2", "--genomes", "1", "--genome-len", "60000", "--arrays", "6",
                    "--spacers-min", "6", "--spacers-max", "14", "--shared-family-arrays", "4",
                    "--tandem", "12", "--pairs", "9000", "--error-rate", "0.003"],
                   [(20, 27, 77), (10, 20, 100), (5, 5, 120), (50, 27, 77)], False, False),
    "biallelic": (["--seed", "3", "--genomes", "1", "--genome-len", "40000", "--arrays", "1",
                   "--spacers-min", "150", "--spacers-max", "150", "--biallelic-arrays", "1",
                   "--sites", "2", "--spacer-min-len", "38", "--pairs", "8000"],
                  [(20, 27, 77)], False, True),
    "strain": (["--seed", "4", "--genomes", "2", "--genome-len", "40000", "--arrays", "3",
                "--spacers-min", "8", "--spacers-max", "16", "--strain-pairs", "2",
                "--pairs", "16000", "--error-rate", "0.005"],
               [(20, 27, 77), (10, 27, 77)], False, False),
}


def pack_graph(raw, path):
    w = raw["w"]
    if w.size % 2:
        w = np.append(w, np.uint8(0))
    w4 = (w[0::2] | (w[1::2] << 4)).astype(np.uint8)
    np.savez_compressed(path, n=np.uint64(raw["w"].size), k=np.uint32(raw["k"]), w4=w4,
                        last=np.packbits(raw["last"], bitorder="little"),
                        tip=np.packbits(raw["tip"], bitorder="little"),
                        mult=raw["mult"], f=raw["f"])


REGION_KS = (0, 1, 7, 40, 150)  # 150 = the reference's default extension, src/main.cpp


def regions(name, prefix, tmp, params, n_items):
    out = {}
    reg = os.path.join(tmp, "out.reg")
    thr, mn, mx = params[0]
    for k in REGION_KS:
        out[f"cyc_t{thr}_{mn}_{mx}_k{k}"] = np.packbits(
            O.run_reference_regions(prefix, reg, k, cycles=(thr, mn, mx)), bitorder="little")
    rng = np.random.default_rng(sum(map(ord, name)))
    seeds = rng.integers(0, n_items, 40).astype("<u8")
    seeds = np.concatenate([seeds, seeds[:5]])  # duplicates, like a node shared by several cycles
    sf = os.path.join(tmp, "seeds.u64")
    seeds.tofile(sf)
    out["seeds"] = seeds
    for k in (3, 30):
        out[f"nodes_k{k}"] = np.packbits(O.run_reference_regions(prefix, reg, k, nodes_file=sf), bitorder="little")
    np.savez_compressed(os.path.join(GOLD, f"{name}.regions.npz"), **out)
    if os.path.isdir(os.path.join(GOLD, f"{name}.sdbg")) or CASES[name][3]:
        # + the reference's SCC split (divide_graph_into_subgraphs, src/spacer_ordering.cpp:141-175): whole
        # harness output, compared byte for byte with the drop-in's (needs the on-disk graph files)
        runs = [(f"cyc_t{thr}_{mn}_{mx}_k{k}", [str(k), "cycles+scc", str(thr), str(mn), str(mx)]) for k in (7, 150)]
        runs.append(("nodes_k30", ["30", "nodes+scc", sf]))
        for key, tail in runs:
            subprocess.run([O.ref_binary("ref_regions"), prefix, reg] + tail, check=True, capture_output=True)
            with open(reg, "rb") as fi, gzip.GzipFile(os.path.join(GOLD, f"{name}.scc_{key}.reg.gz"), "wb", mtime=0) as fo:
                shutil.copyfileobj(fi, fo)
    print("    regions:", {k: int(np.unpackbits(v, bitorder="little")[:n_items].size - np.unpackbits(v, bitorder="little")[:n_items].sum())
                           for k, v in out.items() if k != "seeds"})


def parse_reads_dump(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"MCAATRDS"
    a = np.frombuffer(raw, "<u8", offset=8)
    n, pos, off, ids = int(a[0]), 1, [0], []
    for _ in range(n):
        c = int(a[pos])
        ids.append(a[pos + 1:pos + 1 + c])
        pos += 1 + c
        off.append(off[-1] + c)
    return np.array(off, np.uint64), (np.concatenate(ids) if ids else np.zeros(0, np.uint64)).astype(np.uint64)


def reads(name, prefix, tmp, params, n_items, k):
    """SURVEY 8(f) N3 pins: the UNMODIFIED get_read_from_sequence / reverse_pair_ends_sequence (oracle/_ref/ref_reads)
    on real reads of the case, mutated copies (substitutions, N), random and too-short sequences, against the
    cycle nodes of the reference run (set A) and a random half of all ids (set B, so that most reads are kept and
    every k-mer of them is looked up)."""
    def fastq_seqs(fq, n):
        out = []
        with open(fq) as f:
            for i, line in enumerate(f):
                if i % 4 == 1:
                    out.append(line.strip())
                if len(out) >= n:
                    break
        return out
    rng = np.random.default_rng(1000 + sum(map(ord, name)))
    real = fastq_seqs(os.path.join(tmp, "r1.fq"), 120) + fastq_seqs(os.path.join(tmp, "r2.fq"), 120)
    seqs = list(real)
    for s_ in real[:70]:
        b = list(s_)
        for _ in range(int(rng.integers(1, 4))):
            b[int(rng.integers(len(b)))] = "ACGTN"[int(rng.integers(5))]
        seqs.append("".join(b))
    seqs += ["".join("ACGT"[int(x)] for x in rng.integers(0, 4, 150)) for _ in range(15)]
    base = real[0]
    seqs += ["", base[:10], base[:2 * k], base[:2 * k + 1], base[:2 * k + 2], base[:k], "N" * 60, base.lower()[:80]]
    sf = os.path.join(tmp, "seqs.txt")
    with open(sf, "w") as f:
        f.write("\n".join(seqs) + "\n")
    thr, mn, mx = params[0]
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import golden_util as G
    set_a = G.region_seeds(name, (thr, mn, mx))
    set_b = np.nonzero(rng.random(n_items) < 0.5)[0].astype("<u8")
    out = dict(seqs=np.frombuffer("\n".join(seqs).encode(), np.uint8), set_a_params=np.array([thr, mn, mx]))
    rd = os.path.join(tmp, "out.rd")
    for key, ids, rev in (("a_fwd", set_a, 0), ("a_rev", set_a, 1), ("b_fwd", set_b, 0), ("b_rev", set_b, 1)):
        nf = os.path.join(tmp, "nodes.u64")
        np.asarray(ids, "<u8").tofile(nf)
        subprocess.run([O.ref_binary("ref_reads"), prefix, rd, sf, nf, str(rev)], check=True, capture_output=True)
        off, flat = parse_reads_dump(rd)
        out[key + "_off"], out[key + "_ids"] = off, flat
        print(f"    reads {key}: kept {int((np.diff(off.astype(np.int64)) > 0).sum())}/{len(seqs)}, "
              f"null ids {int((flat == np.uint64(2**64 - 1)).sum())}/{flat.size}")
    out["set_b_seed"] = np.array([1000 + sum(map(ord, name))])
    out["set_b"] = np.packbits(np.isin(np.arange(n_items), set_b), bitorder="little")
    np.savez_compressed(os.path.join(GOLD, f"{name}.reads.npz"), **out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("cases", nargs="*", default=list(CASES))
    ap.add_argument("--regions-only", action="store_true", help="rebuild the graphs but write only <case>.regions.npz")
    ap.add_argument("--keep-tmp", default=None)
    a = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    for name in a.cases:
        gen, params, keep_nav, keep_files = CASES[name]
        tmp = a.keep_tmp and os.path.join(a.keep_tmp, name) or tempfile.mkdtemp(prefix="golden_")
        os.makedirs(tmp, exist_ok=True)
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_reads.py"), "--out", tmp] + gen, check=True)
        gdir = os.path.join(tmp, "gdir")
        subprocess.run([O.ref_binary("build_sdbg"), f"{tmp}/r1.fq {tmp}/r2.fq", gdir, "4", "4e9"],
                       check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        prefix = os.path.join(gdir, "graph")
        g = O.OracleGraph.load(prefix)
        print(f"{name}: n_items={g.n}")
        regions(name, prefix, tmp, params, g.n)
        if keep_files:
            reads(name, prefix, tmp, params, g.n, g.k)
        if a.regions_only:
            if not a.keep_tmp:
                shutil.rmtree(tmp)
            continue
        pack_graph(g.raw(), os.path.join(GOLD, f"{name}.graph.npz"))
        for thr, mn, mx in params:
            dump = os.path.join(tmp, "out.dump")
            info, _ = O.run_reference(prefix, dump, 1, thr, mn, mx)
            with open(dump, "rb") as fi, gzip.GzipFile(os.path.join(GOLD, f"{name}.t{thr}_{mn}_{mx}.dump.gz"), "wb", mtime=0) as fo:
                shutil.copyfileobj(fi, fo)
            print("   ", (thr, mn, mx), info)
        if keep_nav:
            nav = os.path.join(tmp, "out.nav")
            subprocess.run([O.ref_binary("ref_cyclefinder"), "--nav", prefix, nav], check=True)
            rec = np.dtype([("od", "i1"), ("id", "i1"), ("m", "<u2"), ("o", "<u8", 4), ("i", "<u8", 4)])
            t = np.fromfile(nav, dtype=rec, offset=16)
            np.savez_compressed(os.path.join(GOLD, f"{name}.nav.npz"), od=t["od"], id=t["id"], m=t["m"],
                                o=t["o"].astype(np.int64), i=t["i"].astype(np.int64))
        if keep_files:
            fd = os.path.join(GOLD, f"{name}.sdbg")
            os.makedirs(fd, exist_ok=True)
            for fn in sorted(os.listdir(gdir)):
                if fn.startswith("graph.sdbg"):
                    with open(os.path.join(gdir, fn), "rb") as fi, gzip.GzipFile(os.path.join(fd, fn + ".gz"), "wb", mtime=0) as fo:
                        shutil.copyfileobj(fi, fo)
        if not a.keep_tmp:
            shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
