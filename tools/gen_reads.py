#!/usr/bin/env python3
"""Seeded synthetic metagenome + paired-end read generator (test/bench INPUT synthesis only).

Shapes follow BASELINE.json `configs` / SURVEY.md §8(d): random genomes with planted CRISPR arrays
(30 bp repeat, 32-38 bp spacers), optional strain pairs (SNP copies), tandem-repeat loci, arrays that
share a repeat family, arrays with biallelic spacers (the reference's 500-cycle cluster-bound trigger,
reference src/cycle_finder.cpp:162-165) and substitution read errors.  Output: two FASTQ files whose
mate 2 is reverse-complemented, as the reference's builder expects ("pe" library, src/sdbg_build.cpp:47-57).

Everything is drawn from numpy's PCG64 seeded with --seed, so a (args, seed) pair names one input.
"""
import argparse
import json
import os

import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGT", b"TGCA"):
    COMP[a] = b


def rand_seq(rng, n):
    return ACGT[rng.integers(0, 4, size=n)]


def mutate(rng, seq, rate):
    seq = seq.copy()
    if rate <= 0:
        return seq
    m = rng.random(seq.size) < rate
    idx = np.nonzero(m)[0]
    if idx.size:
        cur = np.searchsorted(ACGT, seq[idx])
        seq[idx] = ACGT[(cur + rng.integers(1, 4, size=idx.size)) % 4]
    return seq


def make_array(rng, repeat, n_spacers, lo, hi, alleles=1, sites=1):
    """repeat (spacer repeat)* ; returns (sequence, [alternative-allele sequences])."""
    parts = [repeat]
    alt_parts = [[repeat] for _ in range(alleles - 1)]
    for _ in range(n_spacers):
        sp = rand_seq(rng, int(rng.integers(lo, hi + 1)))
        parts += [sp, repeat]
        if alleles > 1:
            # sites == 2: two variant sites > k apart (needs spacers >= 37 bp) => independent bubbles,
            # 4 paths per spacer from 2 haplotypes while every node keeps half of the coverage
            ps = [int(rng.integers(8, sp.size - 8))] if sites == 1 else [6, sp.size - 7]
            for a in range(alleles - 1):
                sp2 = sp.copy()
                for p in ps:
                    sp2[p] = ACGT[(np.searchsorted(ACGT, sp[p]) + 1 + a) % 4]
                alt_parts[a] += [sp2, repeat]
    seq = np.concatenate(parts)
    return seq, [np.concatenate(ap) for ap in alt_parts]


def build_community(args, rng):
    genomes, truth = [], []
    for g in range(args.genomes):
        genomes.append(rand_seq(rng, args.genome_len))
    # tandem-repeat loci
    for t in range(args.tandem):
        g = int(rng.integers(0, len(genomes)))
        period = int(rng.integers(24, 71))
        copies = int(rng.integers(4, 26))
        unit = rand_seq(rng, period)
        locus = np.concatenate([mutate(rng, unit, args.tandem_div) for _ in range(copies)])
        pos = int(rng.integers(1000, genomes[g].size - locus.size - 1000))
        genomes[g][pos:pos + locus.size] = locus
    # CRISPR arrays, round-robin over genomes
    families = []
    alt_haplotypes = []
    for a in range(args.arrays):
        g = a % len(genomes)
        if a < args.shared_family_arrays and a % 2 == 1 and families:
            repeat = families[-1]
        else:
            repeat = rand_seq(rng, args.repeat_len)
            families.append(repeat)
        nsp = int(rng.integers(args.spacers_min, args.spacers_max + 1))
        bial = a >= args.arrays - args.biallelic_arrays
        seq, alts = make_array(rng, repeat, nsp, args.spacer_min_len, args.spacer_max_len,
                               args.alleles if bial else 1, args.sites)
        span = genomes[g].size - seq.size - 2000
        pos = 1000 + int(rng.integers(0, max(1, span)))
        genomes[g][pos:pos + seq.size] = seq
        truth.append({"genome": g, "pos": pos, "repeat": bytes(repeat).decode(), "spacers": nsp,
                      "biallelic": bool(bial)})
        for alt in alts:
            # further haplotypes: a copy of the flanks with the alternative spacer alleles
            lo = max(0, pos - args.frag_max)
            hi = min(genomes[g].size, pos + seq.size + args.frag_max)
            hap = genomes[g][lo:hi].copy()
            hap[pos - lo:pos - lo + alt.size] = alt
            alt_haplotypes.append(hap)
    for a in range(getattr(args, "bound_arrays", 0)):
        # large arrays whose spacers carry 2 alleles at 2 independent sites: >= 500 cycles through the
        # repeat, the reference's cluster-bound case (src/cycle_finder.cpp:162-165)
        g = (args.arrays + a) % len(genomes)
        repeat = rand_seq(rng, args.repeat_len)
        seq, alts = make_array(rng, repeat, 150, 38, 38, 2, 2)
        pos = 1000 + int(rng.integers(0, max(1, genomes[g].size - seq.size - 2000)))
        genomes[g][pos:pos + seq.size] = seq
        truth.append({"genome": g, "pos": pos, "repeat": bytes(repeat).decode(), "spacers": 150, "biallelic": True})
        for alt in alts:
            lo = max(0, pos - args.frag_max)
            hi = min(genomes[g].size, pos + seq.size + args.frag_max)
            hap = genomes[g][lo:hi].copy()
            hap[pos - lo:pos - lo + alt.size] = alt
            alt_haplotypes.append(hap)
    abund = np.ones(len(genomes))
    if args.abundance_sigma > 0:
        abund = rng.lognormal(0.0, args.abundance_sigma, size=len(genomes))
    # strain pairs: SNP copies of the first `strain_pairs` genomes
    for s in range(args.strain_pairs):
        genomes.append(mutate(rng, genomes[s], args.strain_snp))
        abund = np.append(abund, abund[s] * 0.7)
    for hap in alt_haplotypes:
        genomes.append(hap)
        abund = np.append(abund, 1.0)
    return genomes, abund, truth


def sample_reads(args, rng, genomes, abund):
    lens = np.array([g.size for g in genomes], dtype=np.float64)
    usable = np.maximum(lens - args.frag_max, 1.0)
    w = abund * usable
    w /= w.sum()
    n = args.pairs
    gi = rng.choice(len(genomes), size=n, p=w)
    frag = rng.integers(args.frag_min, args.frag_max + 1, size=n)
    L = args.read_len
    r1 = np.empty((n, L), dtype=np.uint8)
    r2 = np.empty((n, L), dtype=np.uint8)
    ar = np.arange(L)
    for g in range(len(genomes)):
        sel = np.nonzero(gi == g)[0]
        if sel.size == 0:
            continue
        G = genomes[g]
        f = np.minimum(frag[sel], G.size)
        start = (rng.random(sel.size) * (G.size - f + 1)).astype(np.int64)
        strand = rng.random(sel.size) < 0.5
        a = G[start[:, None] + ar[None, :]]
        bidx = (start + f)[:, None] - 1 - ar[None, :]
        b = COMP[G[bidx]]
        # fragment from the reverse strand: swap mates
        r1[sel] = np.where(strand[:, None], a, b)
        r2[sel] = np.where(strand[:, None], b, a)
    if args.error_rate > 0:
        for r in (r1, r2):
            m = rng.random(r.shape) < args.error_rate
            idx = np.nonzero(m)
            cur = np.searchsorted(ACGT, r[idx])
            r[idx] = ACGT[(cur + rng.integers(1, 4, size=cur.size)) % 4]
    return r1, r2


def write_fastq(path, reads, tag):
    n, L = reads.shape
    ids = np.char.zfill(np.arange(n).astype(str), 10).astype("S10")
    hdr = np.frombuffer(ids.tobytes(), dtype=np.uint8).reshape(n, 10)
    rec = np.empty((n, 1 + 10 + 2 + 1 + L + 1 + 2 + L + 1), dtype=np.uint8)
    c = 0
    rec[:, c] = ord("@"); c += 1
    rec[:, c:c + 10] = hdr; c += 10
    rec[:, c] = ord("/"); rec[:, c + 1] = ord(tag); c += 2
    rec[:, c] = 10; c += 1
    rec[:, c:c + L] = reads; c += L
    rec[:, c] = 10; c += 1
    rec[:, c] = ord("+"); rec[:, c + 1] = 10; c += 2
    rec[:, c:c + L] = ord("I"); c += L
    rec[:, c] = 10
    with open(path, "wb") as f:
        step = 1 << 18
        for i in range(0, n, step):
            rec[i:i + step].tofile(f)


ARGSPEC = [
    ("--out", dict(required=True)),
    ("--seed", dict(type=int, default=11)),
    ("--genomes", dict(type=int, default=5)),
    ("--genome-len", dict(type=int, default=2_000_000)),
    ("--arrays", dict(type=int, default=10)),
    ("--repeat-len", dict(type=int, default=30)),
    ("--spacer-min-len", dict(type=int, default=32)),
    ("--spacer-max-len", dict(type=int, default=38)),
    ("--spacers-min", dict(type=int, default=15)),
    ("--spacers-max", dict(type=int, default=40)),
    ("--shared-family-arrays", dict(type=int, default=0)),
    ("--biallelic-arrays", dict(type=int, default=0)),
    ("--alleles", dict(type=int, default=2, help="alleles per spacer in the multi-allelic arrays (<= 4)")),
    ("--sites", dict(type=int, default=1, help="variant sites per spacer in those arrays (1 or 2)")),
    ("--bound-arrays", dict(type=int, default=0, help="extra 150-spacer arrays with 2 alleles at 2 sites")),
    ("--strain-pairs", dict(type=int, default=0)),
    ("--strain-snp", dict(type=float, default=0.01)),
    ("--tandem", dict(type=int, default=0)),
    ("--tandem-div", dict(type=float, default=0.03)),
    ("--abundance-sigma", dict(type=float, default=0.0)),
    ("--pairs", dict(type=int, default=1_000_000)),
    ("--read-len", dict(type=int, default=150)),
    ("--frag-min", dict(type=int, default=300)),
    ("--frag-max", dict(type=int, default=400)),
    ("--error-rate", dict(type=float, default=0.0)),
]


def main(argv=None):
    p = argparse.ArgumentParser()
    for name, kw in ARGSPEC:
        p.add_argument(name, **kw)
    args = p.parse_args(argv)
    rng = np.random.Generator(np.random.PCG64(args.seed))
    os.makedirs(args.out, exist_ok=True)
    genomes, abund, truth = build_community(args, rng)
    r1, r2 = sample_reads(args, rng, genomes, abund)
    write_fastq(os.path.join(args.out, "r1.fq"), r1, "1")
    write_fastq(os.path.join(args.out, "r2.fq"), r2, "2")
    with open(os.path.join(args.out, "truth.json"), "w") as f:
        json.dump({"args": vars(args), "arrays": truth}, f, indent=1)
    print(f"wrote {args.pairs} pairs x {args.read_len} bp from {len(genomes)} sequences to {args.out}")


if __name__ == "__main__":
    main()
