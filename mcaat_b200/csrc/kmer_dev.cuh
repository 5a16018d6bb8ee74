// SURVEY 8(f) N3 — k-mer -> node id and read -> node ids, device functions (also compiled for the host by
// tests/emu).  Reference: SDBG::IndexBinarySearch (libs/megahit/src/sdbg/sdbg.h:141-207), k_mer_to_node_id,
// reverse_pair_ends_sequence and get_read_from_sequence (src/reads.cpp:20-89).
#pragma once
#include "sdbg_dev.cuh"

namespace mg {

// What IndexBinarySearch needs beyond the navigation structures: tip flags with a rank (TipLabelStartPtr,
// sdbg.h:124-127), the tip labels and the bucket table prefix_look_up_ (sdbg.h:36,44-46).
struct KmerIndex {
  const uint64_t *tip;         // n bits
  const uint32_t *tip_prefix;  // per 64-bit word of `tip`: tips before the word
  const uint32_t *labels;      // words_per_label per tip, words as stored in the graph file, tips in item order
  uint32_t words_per_label;
  uint32_t n_buckets;
  const int64_t *first, *second;  // per bucket; buckets absent from the file hold (0, 0) like the reference's resize()
  uint32_t k;
  // Optional shortcut (null = off).  Every probe of the binary search first walks kJumpSteps Backward steps whose
  // characters are the bucket's own prefix — equal to the query's by construction.  jump[x] = Backward^8(x), or
  // kJumpLiteral when one of x, Backward(x), ..., Backward^7(x) is a tip (the reference would compare a tip label
  // there).  bucket_ok: bit b set when the items of bucket b really carry prefix b (checked when the index is built;
  // an absent bucket's (0, 0) range does not).  With both, a probe starts at depth 8 with ONE load instead of 8
  // rank/select rounds; everything from there on is the literal code.
  const uint32_t *jump;
  const uint32_t *bucket_ok;
};
constexpr int kJumpSteps = 8;
constexpr uint32_t kJumpLiteral = 0x80000000u;  // flag bit (ids are < 2^31)

MG_HD bool kx_is_tip(const KmerIndex &K, uint32_t x) { return (K.tip[x >> 6] >> (x & 63)) & 1; }
// CharAtTipLabel (sdbg.h:128-133) on file words: the reference reverses the 2-bit groups of every word when
// loading (sdbg_raw_content.cpp:84-87) and reads little-endian (kmcompactvector.h:14-15,52-60) => char `off` is
// bits 30 - 2 (off % 16) of word off / 16; 1-based values.
MG_HD int kx_tip_char(const KmerIndex &K, uint32_t tip_item, unsigned off) {
  uint32_t idx = K.tip_prefix[tip_item >> 6] + (uint32_t)popc64(K.tip[tip_item >> 6] & ((1ull << (tip_item & 63)) - 1));
  const uint32_t *lab = K.labels + (uint64_t)idx * K.words_per_label;
  return (int)((lab[off >> 4] >> (30 - 2 * (off & 15))) & 3u) + 1;
}
// GetLastIndex = rs_last_.succ(x) (sdbg.h:77, kmrns.h:214-222): first item >= x with the last flag
MG_HD uint32_t kx_last_index(const GraphView &g, uint32_t x) {
  uint64_t wi = x >> 6;
  const uint64_t n_words = (g.n + 63) / 64;
  uint64_t word = g.fb[wi >> 2].last[wi & 3] & (~0ull << (x & 63));
  while (!word) {
    if (++wi >= n_words) return kNull;
    word = g.fb[wi >> 2].last[wi & 3];
  }
#if defined(__CUDA_ARCH__)
  uint32_t r = (uint32_t)(wi * 64 + (__ffsll((long long)word) - 1));
#else
  uint32_t r = (uint32_t)(wi * 64 + __builtin_ctzll(word));
#endif
  return r < g.n ? r : kNull;
}

// sdbg.h:141-207 line by line; seq[0..k) holds 1..4.  W is compared RAW (:187-195, labels 5..8 are not folded to
// 1..4 as GetLabel does), and a probe that runs into a tip compares against the stored label — both decide
// where the search goes, so both are kept.  kNull = kNullID.
MG_HD uint32_t index_binary_search(const GraphView &g, const KmerIndex &K, const uint8_t *seq) {
  const int k = (int)K.k;
  uint64_t prefix = 0;
  for (uint64_t i = 0; (1u << (i * 2)) < K.n_buckets; ++i) prefix = prefix * 4 + seq[k - 1 - i] - 1;
  int64_t l = K.first[prefix], r = K.second[prefix];
  const bool shortcut = K.jump && k > kJumpSteps && ((K.bucket_ok[prefix >> 5] >> (prefix & 31)) & 1);
  while (l <= r) {
    int cmp = 0;
    const uint32_t mid = (uint32_t)((l + r) / 2);
    uint32_t y = mid;
    int i_from = k - 1;
    if (shortcut) {
      const uint32_t j = K.jump[mid];
      if (!(j & kJumpLiteral)) {  // the first 8 characters are the bucket prefix = the query's: skip them
        y = j;
        i_from = k - 1 - kJumpSteps;
      }
    }
    for (int i = i_from; i >= 0; --i) {
      if (kx_is_tip(K, y)) {
        for (int j = 0; j < i; ++j) {
          int c = kx_tip_char(K, y, (unsigned)j);
          if (c < seq[i - j]) { cmp = -1; break; }
          else if (c > seq[i - j]) { cmp = 1; break; }
        }
        if (cmp == 0) {
          if (kx_is_tip(K, mid)) cmp = -1;
          else {
            int c = kx_tip_char(K, y, (unsigned)i);
            if (c < seq[0]) cmp = -1;
            else if (c > seq[0]) cmp = 1;
          }
        }
        break;
      }
      y = backward(g, y);
      if (y >= g.n) return kNull;  // cannot happen on a well-formed graph (f_[1] == 0: no label ends in '$')
      int c = (int)get_w(g, y);
      if (c < seq[i]) { cmp = -1; break; }
      else if (c > seq[i]) { cmp = 1; break; }
    }
    if (cmp == 0) return kx_last_index(g, mid);
    else if (cmp > 0) r = (int64_t)mid - 1;
    else l = (int64_t)mid + 1;
  }
  return kNull;
}

// Index construction helpers.  jump_init: one Backward step (or the literal flag for a tip / a broken chain).
MG_HD uint32_t jump_init(const GraphView &g, const KmerIndex &K, uint32_t x) {
  if (kx_is_tip(K, x)) return kJumpLiteral;
  const uint32_t y = backward(g, x);
  return y < g.n ? y : kJumpLiteral;
}
// doubling: steps(dst) = 2 * steps(src); the flag propagates
MG_HD uint32_t jump_double(const uint32_t *src, uint32_t x) {
  const uint32_t v = src[x];
  if (v & kJumpLiteral) return kJumpLiteral;
  const uint32_t w = src[v];
  return (w & kJumpLiteral) ? kJumpLiteral : w;
}
// prefix code of item x as IndexBinarySearch computes it from a query (sdbg.h:142-146), read off the graph by the
// literal walk; -1 when a tip is met within the first `chars` steps
MG_HD int64_t item_prefix(const GraphView &g, const KmerIndex &K, uint32_t x, int chars) {
  int64_t code = 0;
  uint32_t y = x;
  for (int i = 0; i < chars; ++i) {
    if (kx_is_tip(K, y)) return -1;
    y = backward(g, y);
    if (y >= g.n) return -1;
    code = code * 4 + (int64_t)get_w(g, y) - 1;
  }
  return code;
}
// bucket b may use the shortcut when its first and last item carry prefix b (items are sorted by reversed label, so
// everything in between does too) — or when it is empty
MG_HD bool bucket_consistent(const GraphView &g, const KmerIndex &K, uint32_t b, int chars) {
  const int64_t l = K.first[b], r = K.second[b];
  if (l > r) return true;
  if (l < 0 || (uint64_t)r >= g.n) return false;
  return item_prefix(g, K, (uint32_t)l, chars) == (int64_t)b && item_prefix(g, K, (uint32_t)r, chars) == (int64_t)b;
}

// k_mer_to_node_id's alphabet (src/reads.cpp:44-52): A, C, G -> 1, 2, 3, anything else -> 4; with `reverse` the
// complement of reverse_pair_ends_sequence (:23-30: A<->T, C<->G, other characters unchanged) is applied first.
MG_HD uint8_t base_code(char c, bool reverse) {
  if (reverse) return c == 'T' ? 1 : c == 'G' ? 2 : c == 'C' ? 3 : 4;
  return c == 'A' ? 1 : c == 'C' ? 2 : c == 'G' ? 3 : 4;
}
// code of position p of the (possibly reverse-complemented) sequence seq[0..len)
MG_HD uint8_t read_code(const char *seq, uint32_t len, uint32_t p, bool reverse) {
  return base_code(reverse ? seq[len - 1 - p] : seq[p], reverse);
}

constexpr int kMaxK = 128;  // k-mer buffer per thread; MEGAHIT's k_max is 255, mcaat uses 23

// node id of the k-mer at position pos of the (possibly reverse-complemented) read
MG_HD uint32_t lookup_at(const GraphView &g, const KmerIndex &K, const char *seq, uint32_t len, uint32_t pos, bool reverse) {
  uint8_t codes[kMaxK];
  for (uint32_t i = 0; i < K.k; ++i) codes[i] = read_code(seq, len, pos + i, reverse);
  return index_binary_search(g, K, codes);
}
MG_HD bool member_get(const uint32_t *bm, uint32_t i) { return (bm[i >> 5] >> (i & 31)) & 1; }
// get_read_from_sequence, first half (src/reads.cpp:64-77): number of ids the read maps to (0 = dropped) and the
// ids of its first and last k-mer
MG_HD uint32_t read_ends(const GraphView &g, const KmerIndex &K, const char *seq, uint32_t len, bool reverse,
                         const uint32_t *member, uint32_t *first_id, uint32_t *last_id) {
  if (len <= 2 * K.k) return 0;  // :64-66
  const uint32_t s = lookup_at(g, K, seq, len, 0, reverse);
  const uint32_t e = lookup_at(g, K, seq, len, len - K.k, reverse);
  const bool in_s = s != kNull && member_get(member, s), in_e = e != kNull && member_get(member, e);
  if (!in_s && !in_e) return 0;  // :74-77
  *first_id = s;
  *last_id = e;
  return len - K.k + 1;
}
// second half (:79-87): id number p of a kept read
MG_HD uint32_t read_id_at(const GraphView &g, const KmerIndex &K, const char *seq, uint32_t len, bool reverse, uint32_t p,
                          uint32_t n_out, uint32_t first_id, uint32_t last_id) {
  if (p == 0) return first_id;
  if (p == n_out - 1) return last_id;
  return lookup_at(g, K, seq, len, p, reverse);
}

}  // namespace mg
