#!/usr/bin/env python3
"""Synthetic-input construction for tests and bench.py (NOT part of the timed hot path).

`sdbg_from_kmers` builds MEGAHIT's succinct de Bruijn graph arrays directly from a (k+1)-mer
multiset, restating what the reference's graph producer emits for that multiset (reference
libs/megahit/src/sorting/read_to_sdbg_s2.cpp:343-437 which items exist, :520-612 W / last / tip /
multiplicity, libs/megahit/src/sdbg/sdbg.h:36-50 the f_ boundaries).  It is pinned against the
reference's own builder: tests/test_synth_sdbg.py feeds it the (k+1)-mers of the golden cases'
reads and compares every array with the golden graph made by oracle/_ref/build_sdbg.

Why it exists: BASELINE.json's bench workload is a ~200M-item SDBG.  The reference builder needs
hours of CPU for 50M read pairs; sampling the community's (k+1)-mer spectrum (coverage-distributed
multiplicities, strain variants, tandem repeats, CRISPR arrays, read-error bubbles and tips) and
sorting it on the GPU takes seconds and gives a graph with the same structure.

All tensor code runs on any torch device (CPU here, CUDA on the B200 box).
"""
import os

import numpy as np
import torch

K = 23  # the reference builds with -k 23 (src/sdbg_build.cpp:210-220)


# ------------------------------------------------------------------ (k+1)-mers
def seq_codes(seq, device="cpu"):
    """uint8 ASCII ACGT ndarray -> int64 tensor of 2-bit codes (A0 C1 G2 T3)."""
    lut = np.full(256, 255, np.uint8)
    for i, c in enumerate(b"ACGT"):
        lut[c] = i
    return torch.from_numpy(lut[seq]).to(device=device, dtype=torch.int64)


def windows(codes, width):
    """All `width`-mers of a 1-D code tensor as int64 (first char most significant)."""
    n = codes.numel() - width + 1
    if n <= 0:
        return codes.new_zeros(0)
    out = torch.zeros(n, dtype=torch.int64, device=codes.device)
    for j in range(width):
        out = (out << 2) | codes[j:j + n]
    return out


def revcomp_kmers(e, width):
    """Reverse complement of packed `width`-mers."""
    x = ~e & ((1 << (2 * width)) - 1)  # complement every base (3 - c)
    out = torch.zeros_like(x)
    for _ in range(width):  # reverse the order of the 2-bit groups
        out = (out << 2) | (x & 3)
        x = x >> 2
    return out


def both_strands(e, width=K + 1):
    """The builder adds every edge and its reverse complement, palindromes once (:392-437)."""
    rc = revcomp_kmers(e, width)
    return torch.cat([e, rc[rc != e]])


def count_kmers(e, width=K + 1):
    """(k+1)-mer occurrences (both strands already added) -> (sorted unique codes, counts) in the
    BUILDER'S ORIENTATION: the reference producer's item order equals plain lexicographic order on
    the base-complemented (k+1)-mers (equivalently, the multiset being reverse-complement closed,
    on the reversed ones) — established empirically against oracle/_ref/build_sdbg, see
    tests/test_synth_sdbg.py."""
    e = ~e & ((1 << (2 * width)) - 1)
    u, c = torch.unique(e, sorted=True, return_counts=True)
    return u, c


def kmers_from_read_matrix(reads, k=K, device="cpu"):
    """reads: (n, L) uint8 ASCII matrix -> (sorted unique (k+1)-mer codes, counts) over both strands."""
    codes = seq_codes(reads, device)  # (n, L)
    n, L = codes.shape
    w = k + 1
    m = L - w + 1
    out = torch.zeros((n, m), dtype=torch.int64, device=device)
    for j in range(w):
        out = (out << 2) | codes[:, j:j + m]
    return count_kmers(both_strands(out.reshape(-1), w), w)


# ------------------------------------------------------------------ SDBG from a (k+1)-mer multiset
def _isin_sorted(x, sorted_set):
    if sorted_set.numel() == 0:
        return torch.zeros_like(x, dtype=torch.bool)
    i = torch.searchsorted(sorted_set, x).clamp_(max=sorted_set.numel() - 1)
    return sorted_set[i] == x


def _pack_nibbles(w):
    n = w.numel()
    pad = (-n) % 16
    if pad:
        w = torch.cat([w, w.new_zeros(pad)])
    w = w.reshape(-1, 16).to(torch.int64)
    sh = torch.arange(16, device=w.device, dtype=torch.int64) * 4
    return (w << sh).sum(dim=1)  # bit pattern of the uint64 word (top nibble may wrap the sign)


def _pack_bits(b):
    n = b.numel()
    pad = (-n) % 64
    if pad:
        b = torch.cat([b, b.new_zeros(pad)])
    b = b.reshape(-1, 64).to(torch.int64)
    sh = torch.arange(64, device=b.device, dtype=torch.int64)
    return (b << sh).sum(dim=1)


def sdbg_from_kmers(edges, counts, k=K, want_labels=False):
    """Sorted unique (k+1)-mer codes + occurrence counts -> packed SDBG arrays (mg_graph_desc packing).

    Items (read_to_sdbg_s2.cpp:343-437): one per (k+1)-mer c.X (node X = last k chars, W char c);
    a `$` item (W = 0) for every prefix k-mer that no (k+1)-mer ends in; a tip item (label S$,
    k-1 chars) for every suffix k-mer that no (k+1)-mer starts with.  Sort key = (label chars,
    full-length flag, W char) (:441-517); W / last / tip as in Lv2Postprocess (:520-612).
    """
    if isinstance(edges, list):  # [codes, counts] handed over by the caller: freed here as soon as they are consumed
        box = edges
        edges, counts = box.pop(0), box.pop(0)
    dev = edges.device
    kk = 2 * k
    mask_k = (1 << kk) - 1
    mask_s = (1 << (kk - 2)) - 1
    # Temporaries are dropped as soon as they are consumed: at 10^9 (k+1)-mers (workload c4) every int64 array is 8 GB.
    X = edges & mask_k          # e[1..k]
    Xu = torch.unique(X)        # sorted
    P = edges >> 2              # e[0..k-1]
    Pu, p_inv = torch.unique(P, return_inverse=True)
    del P
    p_cnt = torch.zeros(Pu.numel(), dtype=torch.int64, device=dev).index_add_(0, p_inv, counts)
    del p_inv
    # `$`-labelled items: prefix k-mers with no incoming (k+1)-mer
    dmask = ~_isin_sorted(Pu, Xu)
    d_nodes, d_cnt = Pu[dmask], p_cnt[dmask]
    del p_cnt, dmask
    # tips: suffix k-mers with no outgoing (k+1)-mer
    x_inv = torch.searchsorted(Xu, X)
    x_cnt = torch.zeros(Xu.numel(), dtype=torch.int64, device=dev).index_add_(0, x_inv, counts)
    del x_inv
    tmask = ~_isin_sorted(Xu, Pu)
    del Pu
    t_nodes, t_cnt = Xu[tmask], x_cnt[tmask]
    del Xu, x_cnt, tmask
    t_S = t_nodes & mask_s
    t_b = t_nodes >> (kk - 2)
    X <<= 4
    X |= 8
    X |= edges >> kk            # e[0]
    del edges
    key = torch.cat([X, (d_nodes << 4) | 8 | 4, ((t_S << 2) << 4) | t_b])
    del X, d_nodes, t_nodes, t_S, t_b
    cnt = torch.cat([counts, d_cnt, t_cnt]).clamp_(max=65535).to(torch.int32)  # multiplicity saturates (:579)
    del counts, d_cnt, t_cnt
    key, order = torch.sort(key)
    cnt = cnt[order]
    del order
    n = key.numel()
    # W: first item of an S-group carrying char c gets c+1, later ones c+5; `$` -> 0
    sb = ((key >> 6) << 3) | (key & 7)   # (S, previous char)
    sb, o2 = torch.sort(sb, stable=True)
    first = torch.ones(n, dtype=torch.bool, device=dev)
    first[1:] = sb[1:] != sb[:-1]
    del sb
    is_first = torch.empty(n, dtype=torch.bool, device=dev)
    is_first[o2] = first
    del o2, first
    prev = (key & 7).to(torch.uint8)
    w = torch.where(prev == 4, torch.zeros_like(prev), torch.where(is_first, prev + 1, prev + 5))
    del prev, is_first
    # last: final item of a node (same label, full length); tips are never `last`
    tip = ((key >> 3) & 1) == 0
    node = key >> 3
    last = torch.ones(n, dtype=torch.bool, device=dev)
    last[:-1] = node[1:] != node[:-1]
    del node
    last &= ~tip
    mult = cnt
    first_char = key >> (4 + kk - 2)
    per_char = torch.bincount(first_char, minlength=4)[:4]
    del first_char
    f = torch.zeros(6, dtype=torch.int64)
    f[0] = -1
    f[2:] = torch.cumsum(per_char.cpu(), 0)
    out = dict(n=n, k=k, w=_pack_nibbles(w).cpu().numpy().view(np.uint64),
               last=_pack_bits(last).cpu().numpy().view(np.uint64),
               tip=_pack_bits(tip).cpu().numpy().view(np.uint64),
               mult=mult.cpu().numpy().astype(np.uint16), f=f.numpy().astype(np.int64))
    if want_labels:
        out["tip_label_chars"] = (key[tip] >> 4).cpu().numpy()  # 2k bits, S then a zero pad char
        out["first8"] = (key >> (4 + kk - 16)).cpu().numpy()
        out["w_unpacked"] = w.to(torch.uint8).cpu().numpy()
        out["last_unpacked"] = last.cpu().numpy()
        out["tip_unpacked"] = tip.cpu().numpy()
    return out


# ------------------------------------------------------------------ MEGAHIT on-disk format
def write_megahit_files(g, prefix, words_per_tip_label=None):
    """Packed SDBG (with want_labels=True) -> <prefix>.sdbg_info + <prefix>.sdbg.0 in the format
    of reference libs/megahit/src/sdbg/{sdbg_meta.cpp:50-75,sdbg_writer.cpp:25-58,sdbg_item.h:14-24},
    so that the UNMODIFIED reference (oracle/_ref/ref_cyclefinder) can load a synthetic graph."""
    k, n = g["k"], g["n"]
    wptl = words_per_tip_label or (2 * k + 31) // 32
    w, last, tip, mult = g["w_unpacked"], g["last_unpacked"], g["tip_unpacked"], g["mult"]
    small = np.minimum(mult, 255).astype(np.uint8)
    rec_len = 2 + 2 * (small == 255) + 4 * wptl * tip
    off = np.zeros(n + 1, np.int64)
    np.cumsum(rec_len, out=off[1:])
    buf = np.zeros(int(off[-1]), np.uint8)
    buf[off[:-1]] = w | (last.astype(np.uint8) << 4) | (tip.astype(np.uint8) << 5)
    buf[off[:-1] + 1] = small
    big = np.nonzero(small == 255)[0]
    buf[off[big] + 2] = mult[big] & 255
    buf[off[big] + 3] = mult[big] >> 8
    tips = np.nonzero(tip)[0]
    if tips.size:
        chars = g["tip_label_chars"].astype(np.uint64) << np.uint64(64 - 2 * k)  # left-aligned 2-bit chars
        words = np.stack([(chars >> np.uint64(32)).astype(np.uint32), (chars & np.uint64(0xFFFFFFFF)).astype(np.uint32)], 1)
        base = off[tips] + 2 + 2 * (small[tips] == 255)
        for j in range(wptl):
            wj = words[:, j] if j < 2 else np.zeros(tips.size, np.uint32)
            for byte in range(4):
                buf[base + 4 * j + byte] = (wj >> np.uint32(8 * byte)) & np.uint32(255)
    with open(prefix + ".sdbg.0", "wb") as fo:
        buf.tofile(fo)
    # bucket table: 65536 buckets keyed by the first 8 label chars (kBucketPrefixLength)
    key_chars = g["first8"]
    nb = 65536
    items = np.bincount(key_chars, minlength=nb)
    tips_b = np.bincount(key_chars, weights=tip, minlength=nb).astype(np.int64)
    large_b = np.bincount(key_chars, weights=(small == 255), minlength=nb).astype(np.int64)
    first_item = np.zeros(nb + 1, np.int64)
    np.cumsum(items, out=first_item[1:])
    with open(prefix + ".sdbg_info", "w") as fo:
        fo.write(f"k {k}\nwords_per_tip_label {wptl}\nnum_buckets {nb}\nnum_files 1\n")
        for bk in range(nb):
            fo.write(f"{bk} 0 {int(off[first_item[bk]])} {int(items[bk])} {int(tips_b[bk])} {int(large_b[bk])}\n")


def kmer_index_arrays(g, words_per_tip_label=None):
    """What mg_kmer_index_create wants for a packed SDBG made with want_labels=True, without going through files:
    (tip labels as file words [n_tips, words_per_tip_label], first item per bucket, last item per bucket) — the same
    values write_megahit_files stores and SDBG::LoadFromFile derives (sdbg.h:36-46)."""
    k = g["k"]
    wptl = words_per_tip_label or (2 * k + 31) // 32
    chars = g["tip_label_chars"].astype(np.uint64) << np.uint64(64 - 2 * k)
    labels = np.zeros((chars.size, wptl), np.uint32)
    if chars.size:
        labels[:, 0] = (chars >> np.uint64(32)).astype(np.uint32)
        if wptl > 1:
            labels[:, 1] = (chars & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    nb = 65536
    items = np.bincount(g["first8"], minlength=nb).astype(np.int64)
    first = np.zeros(nb, np.int64)
    np.cumsum(items[:-1], out=first[1:])
    return labels, first, first + items - 1


# ------------------------------------------------------------------ community -> (k+1)-mer spectrum
def community_spectrum(argv, total_bases, error_events, seed, device="cpu", read_len=150, k=K):
    """Sample the (k+1)-mer spectrum that `total_bases` of error-prone reads of read_len would give
    for the community described by tools/gen_reads.py arguments `argv`, without materialising reads.

    True (k+1)-mers: every genome window gets Poisson(depth * (read_len - k) / read_len) occurrences
    (the expected number of reads containing the window), its reverse complement the same number
    (a read contributes both, read_to_sdbg_s2.cpp:392-437).  Read errors: `error_events`
    substitutions, each placed in a random read position and contributing the (k+1)-mers of that
    read that cover it, once (and their reverse complements).  Returns (codes, counts) in the
    builder's orientation, ready for sdbg_from_kmers, plus a description."""
    import argparse

    import gen_reads
    p = argparse.ArgumentParser()
    for name, kw in gen_reads.ARGSPEC:
        p.add_argument(name, **kw)
    args = p.parse_args(["--out", "-", "--seed", str(seed)] + list(argv))
    rng = np.random.Generator(np.random.PCG64(seed))
    genomes, abund, truth = gen_reads.build_community(args, rng)
    lens = np.array([g.size for g in genomes], np.float64)
    share = abund * lens / float((abund * lens).sum())
    depth = total_bases * share / lens
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    w = k + 1
    codes_all, counts_all = [], []
    ev_per_genome = rng.multinomial(error_events, share) if error_events else np.zeros(len(genomes), np.int64)
    for gi, gseq in enumerate(genomes):
        codes = seq_codes(gseq, device)
        win = windows(codes, w)
        lam = float(depth[gi]) * (read_len - k) / read_len
        cnt = torch.poisson(torch.full((win.numel(),), lam, device=device, dtype=torch.float32), generator=gen).to(torch.int64)
        keep = cnt > 0
        e, c = win[keep], cnt[keep]
        rc = revcomp_kmers(e, w)
        pal = rc == e
        codes_all += [e, rc[~pal]]
        counts_all += [c, c[~pal]]
        ne = int(ev_per_genome[gi])
        if ne and win.numel() > 2 * read_len:
            pos = torch.randint(read_len, gseq.size - read_len, (ne,), device=device, generator=gen)
            rstart = pos - torch.randint(0, read_len, (ne,), device=device, generator=gen)  # read covers pos
            newb = (codes[pos] + torch.randint(1, 4, (ne,), device=device, generator=gen)) & 3
            j = torch.arange(w, device=device)
            s = pos[:, None] - (w - 1) + j[None, :]          # window starts covering pos
            ok = (s >= rstart[:, None]) & (s + w <= rstart[:, None] + read_len)
            idx = (pos[:, None] - s)                          # position of the error inside the window
            sh = 2 * (w - 1 - idx)
            orig = win[s.clamp(0, win.numel() - 1)]
            ecode = (orig & ~(torch.full_like(sh, 3) << sh)) | (newb[:, None] << sh)
            ecode = ecode[ok]
            erc = revcomp_kmers(ecode, w)
            codes_all += [ecode, erc[erc != ecode]]
            one = torch.ones(1, dtype=torch.int64, device=device)
            counts_all += [one.expand(ecode.numel()), one.expand(int((erc != ecode).sum()))]
        del codes, win, cnt
    e = torch.cat(codes_all)
    c = torch.cat(counts_all)
    del codes_all, counts_all
    e = ~e & ((1 << (2 * w)) - 1)  # builder orientation, see count_kmers
    u, inv = torch.unique(e, sorted=True, return_inverse=True)
    cu = torch.zeros(u.numel(), dtype=torch.int64, device=device).index_add_(0, inv, c)
    info = dict(genomes=len(genomes), bases=int(lens.sum()), arrays=len(truth), total_bases=int(total_bases),
                error_events=int(error_events), mean_depth=float((depth * lens).sum() / lens.sum()), seed=seed)
    return u, cu, info


WORKLOADS = {
    # BASELINE.json configs[1]: 50-genome community (+5 strain copies), ~200M-item SDBG
    "c2": dict(argv=["--genomes", "50", "--genome-len", "2000000", "--arrays", "100", "--shared-family-arrays", "20",
                     "--biallelic-arrays", "10", "--bound-arrays", "2", "--strain-pairs", "5", "--tandem", "200",
                     "--abundance-sigma", "1.0"],
               total_bases=50_000_000 * 2 * 150, error_events=500_000, seed=12),
    # The same item count with the error density of real data ("mostly mult-1 error k-mers"): a quarter of
    # the genomes, 7 x the substitution events -> ~3/4 of the items are multiplicity-1 error k-mers.  (configs[1] names
    # BOTH "~200M-node SDBG" and "50 M pairs at 0.5 % errors"; at k = 23 the latter alone would be ~3.4 G distinct error
    # k-mers, so "c2" keeps the item count and the community and thins the errors; this variant keeps the item count and
    # the error share instead.)
    "c2_errors": dict(argv=["--genomes", "12", "--genome-len", "2000000", "--arrays", "100", "--shared-family-arrays", "20",
                            "--biallelic-arrays", "10", "--bound-arrays", "2", "--strain-pairs", "2", "--tandem", "200",
                            "--abundance-sigma", "1.0"],
                      total_bases=12_000_000 * 2 * 150, error_events=3_500_000, seed=12),
    # BASELINE.json configs[3]: high-diversity community (250 genomes, steep abundance tail), ~1 B-item SDBG
    "c4": dict(argv=["--genomes", "250", "--genome-len", "2000000", "--arrays", "500", "--shared-family-arrays", "100",
                     "--biallelic-arrays", "50", "--bound-arrays", "4", "--strain-pairs", "10", "--tandem", "1000",
                     "--abundance-sigma", "2.0"],
               total_bases=250_000_000 * 2 * 150, error_events=1_200_000, seed=13),
    # 1/20 scale of the same recipe: the bounded sample the CPU baselines run on
    "c2_sample": dict(argv=["--genomes", "3", "--genome-len", "1600000", "--arrays", "6", "--shared-family-arrays", "2",
                            "--biallelic-arrays", "1", "--bound-arrays", "1", "--strain-pairs", "1", "--tandem", "12",
                            "--abundance-sigma", "1.0"],
                      total_bases=2_500_000 * 2 * 150, error_events=25_000, seed=12),
    # BASELINE.json configs[0] shape (5 x 2 Mb, 10 arrays, error-free)
    "c1": dict(argv=["--genomes", "5", "--genome-len", "2000000", "--arrays", "10"],
               total_bases=1_000_000 * 2 * 150, error_events=0, seed=11),
    # seconds on a CPU: tests
    "mini": dict(argv=["--genomes", "3", "--genome-len", "60000", "--arrays", "4", "--shared-family-arrays", "2",
                       "--biallelic-arrays", "1", "--strain-pairs", "1", "--tandem", "6", "--abundance-sigma", "0.5"],
                 total_bases=60_000 * 4 * 40, error_events=1500, seed=5),
}


def sample_reads(name, n_reads, seed=1, read_len=150, error_rate=0.002):
    """Reads of the community of workload `name` for the read-mapping stage: n_reads windows of
    read_len bases drawn in proportion to each genome's share of the sequencing, either strand, with substitution
    errors at `error_rate`.  Returns (uint8 ASCII array of n_reads * read_len characters, uint64 offsets)."""
    import argparse

    import gen_reads
    cfg = WORKLOADS[name]
    p = argparse.ArgumentParser()
    for arg, kw in gen_reads.ARGSPEC:
        p.add_argument(arg, **kw)
    args = p.parse_args(["--out", "-", "--seed", str(cfg["seed"])] + list(cfg["argv"]))
    genomes, abund, _ = gen_reads.build_community(args, np.random.Generator(np.random.PCG64(cfg["seed"])))
    lens = np.array([g.size for g in genomes], np.float64)
    share = abund * lens / float((abund * lens).sum())
    rng = np.random.default_rng(seed)
    which = rng.choice(len(genomes), n_reads, p=share)
    out = np.empty((n_reads, read_len), np.uint8)
    comp = np.arange(256, dtype=np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    for gi in np.unique(which):
        rows = np.nonzero(which == gi)[0]
        start = rng.integers(0, genomes[gi].size - read_len, rows.size)
        win = genomes[gi][start[:, None] + np.arange(read_len)[None, :]]
        flip = rng.random(rows.size) < 0.5
        win[flip] = comp[win[flip][:, ::-1]]
        out[rows] = win
    err = rng.random(out.shape) < error_rate
    out[err] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(err.sum()))]
    return out.reshape(-1), np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)


def make_workload(name, device="cpu", want_labels=False):
    cfg = WORKLOADS[name]
    e, c, info = community_spectrum(cfg["argv"], cfg["total_bases"], cfg["error_events"], cfg["seed"], device)
    box = [e, c]
    del e, c
    g = sdbg_from_kmers(box, None, want_labels=want_labels)
    info.update(workload=name, n_items=g["n"])
    return g, info
