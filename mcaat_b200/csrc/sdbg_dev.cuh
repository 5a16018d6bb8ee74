// Device-resident succinct de Bruijn graph: layout and navigation.
//
// Answers are bit-identical to MEGAHIT's SDBG (reference libs/megahit/src/sdbg/sdbg.h) and its
// kmlib::RankAndSelect (reference libs/megahit/src/kmlib/kmrns.h); the layout is our own and is
// chosen for B200's 128-byte lines / 32-byte sectors:
//
//   WBlock (128 B)  16 B header = absolute counts of edge labels 1..4 before the block,
//                   112 B payload = 224 edge-label nibbles.  rank_w(a, i) reads ONE line.
//   FBlock (128 B)  256 items: last[32 B] | last-or-tip[32 B] | invalid[32 B] | header (absolute
//                   rank of `last` before the block).  select_last + the group scan of
//                   ComputeOutgoings (sdbg.h:296-320) read ONE line (two at a block edge).
//   hint_last / hint_w[a]   block index holding every 64th occurrence (select entry points).
//   mult            uint16 per item.
//
// Functions marked MG_HD compile for the device and, for the host-side unit tests in
// tests/emu (which check this exact code against the oracle without a GPU), for the host.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MG_HD __host__ __device__ __forceinline__
#define MG_HD_NOINLINE static __noinline__ __host__ __device__
#else
#define MG_HD inline
#define MG_HD_NOINLINE static inline
#endif

namespace mg {

constexpr uint32_t kNull = 0xFFFFFFFFu;
constexpr int kWNib = 224;    // nibbles per WBlock
constexpr int kWWords = 14;   // payload words per WBlock
constexpr int kFItems = 256;  // items per FBlock
constexpr int kSelH = 64;     // select hint stride (occurrences)

struct alignas(128) WBlock {
  uint32_t cnt[4];  // # of labels 1..4 in items [0, 224*b)
  uint64_t w[kWWords];
};
struct alignas(128) FBlock {
  uint64_t last[4];
  uint64_t lt[4];   // last | tip            (sdbg.h:71-75)
  uint64_t inv[4];  // invalid bits, mutable (sdbg.h:88-94)
  uint32_t rank_before;  // # of last bits in items [0, 256*b)
  uint32_t pad[7];
};
static_assert(sizeof(WBlock) == 128 && sizeof(FBlock) == 128, "block layout");

struct GraphView {
  uint64_t n;          // items
  uint64_t n_wblocks;  // ceil(n / 224)
  uint64_t n_fblocks;  // ceil(n / 256)
  const WBlock *wb;
  FBlock *fb;
  const uint16_t *mult;
  const uint32_t *hint_last;  // [total_last / 64 + 1]
  const uint32_t *hint_w[4];  // per label 1..4
  uint32_t total_last;
  uint32_t total_w[4];
  int64_t f[6];       // sdbg.h:36-50
  int64_t rank_f[6];  // sdbg.h:52-54
};

// ------------------------------------------------------------------ loads
#if defined(__CUDA_ARCH__)
template <class T> MG_HD T ld_ro(const T *p) { return __ldg(p); }
// mutable data written by other SMs in the same launch: bypass the (incoherent) L1
MG_HD uint64_t ld_cg64(const uint64_t *p) { return __ldcg(p); }
MG_HD uint32_t ld_cg32(const uint32_t *p) { return __ldcg(p); }
MG_HD int popc64(uint64_t x) { return __popcll(x); }
MG_HD int popc32(uint32_t x) { return __popc(x); }
#else
template <class T> MG_HD T ld_ro(const T *p) { return *p; }
MG_HD uint64_t ld_cg64(const uint64_t *p) { return *(const volatile uint64_t *)p; }
MG_HD uint32_t ld_cg32(const uint32_t *p) { return *(const volatile uint32_t *)p; }
MG_HD int popc64(uint64_t x) { return __builtin_popcountll(x); }
MG_HD int popc32(uint32_t x) { return __builtin_popcount(x); }
#endif

// bit 4i of the result is set iff nibble i of x equals a
MG_HD uint64_t nib_eq(uint64_t x, unsigned a) {
  uint64_t y = x ^ (0x1111111111111111ull * a);
  y |= y >> 1;
  y |= y >> 2;
  return ~y & 0x1111111111111111ull;
}
// position of the r-th (0-based) set bit of x; x must have > r bits set
MG_HD int select64(uint64_t x, int r) {
#if defined(__CUDA_ARCH__)
  // branch-free binary search over popcounts (__fns is a software loop over the bytes of the word)
  uint32_t w = (uint32_t)x;
  int c = __popc(w), pos = 0;
  if (r >= c) { r -= c; w = (uint32_t)(x >> 32); pos = 32; }
  c = __popc(w & 0xFFFFu);
  if (r >= c) { r -= c; w >>= 16; pos += 16; }
  c = __popc(w & 0xFFu);
  if (r >= c) { r -= c; w >>= 8; pos += 8; }
  c = __popc(w & 0xFu);
  if (r >= c) { r -= c; w >>= 4; pos += 4; }
  c = __popc(w & 0x3u);
  if (r >= c) { r -= c; w >>= 2; pos += 2; }
  return pos + (r >= (int)(w & 1u) ? 1 : 0);
#else
  for (int i = 0; i < r; ++i) x &= x - 1;
  return __builtin_ctzll(x);
#endif
}

// ------------------------------------------------------------------ item getters
MG_HD unsigned get_w(const GraphView &g, uint32_t i) {
  uint32_t b = i / kWNib, r = i - b * kWNib;
  return (unsigned)(ld_ro(&g.wb[b].w[r >> 4]) >> ((r & 15) * 4)) & 15u;
}
MG_HD bool get_lt(const GraphView &g, uint32_t i) {
  return (ld_ro(&g.fb[i >> 8].lt[(i >> 6) & 3]) >> (i & 63)) & 1;
}
MG_HD bool get_last(const GraphView &g, uint32_t i) {
  return (ld_ro(&g.fb[i >> 8].last[(i >> 6) & 3]) >> (i & 63)) & 1;
}
// COHERENT = true inside kernels that set invalid bits while other threads read them
template <bool COHERENT = false>
MG_HD bool get_inv(const GraphView &g, uint32_t i) {
  const uint64_t *p = &g.fb[i >> 8].inv[(i >> 6) & 3];
  uint64_t wv = COHERENT ? ld_cg64(p) : *p;
  return (wv >> (i & 63)) & 1;
}
MG_HD uint32_t get_mult(const GraphView &g, uint32_t i) { return ld_ro(&g.mult[i]); }

// ------------------------------------------------------------------ rank / select
// kmrns.h:287-311 for label a in 1..4 and 0 <= pos < n: # of a in [0..pos]
MG_HD uint32_t rank_w(const GraphView &g, unsigned a, uint32_t pos) {
  uint32_t b = pos / kWNib, r = pos - b * kWNib;
  const WBlock *blk = &g.wb[b];
  uint32_t c = ld_ro(&blk->cnt[a - 1]);
  int full = r >> 4;
  for (int j = 0; j < full; ++j) c += popc64(nib_eq(ld_ro(&blk->w[j]), a));
  uint64_t m = nib_eq(ld_ro(&blk->w[full]), a);
  int keep = ((r & 15) + 1) * 4;  // bits to keep
  if (keep < 64) m &= (1ull << keep) - 1;
  return c + popc64(m);
}
// kmrns.h:313-345: position of the k-th (0-based) a; k == total -> n; k > total -> kNull
MG_HD uint32_t select_w(const GraphView &g, unsigned a, int64_t k) {
  int64_t tot = g.total_w[a - 1];
  if (k < 0 || k > tot) return kNull;
  if (k == tot) return (uint32_t)g.n;
  uint32_t b = ld_ro(&g.hint_w[a - 1][k / kSelH]);
  for (;;) {
    const WBlock *blk = &g.wb[b];
    uint32_t rem = (uint32_t)k - ld_ro(&blk->cnt[a - 1]);
    for (int j = 0; j < kWWords; ++j) {
      uint64_t m = nib_eq(ld_ro(&blk->w[j]), a);
      uint32_t c = popc64(m);
      if (rem < c) return b * kWNib + j * 16 + (select64(m, (int)rem) >> 2);
      rem -= c;
    }
    if (++b >= g.n_wblocks) return kNull;  // unreachable on a consistent index
  }
}
// kmrns.h:287-311 on `last`: pos == -1 -> 0 (sdbg.h:119 calls rank(edge_id - 1)); else # in [0..pos]
MG_HD uint32_t rank_last(const GraphView &g, int64_t pos) {
  if (pos < 0) return 0;
  uint32_t p = (uint32_t)pos;
  const FBlock *blk = &g.fb[p >> 8];
  uint32_t c = ld_ro(&blk->rank_before);
  int full = (p >> 6) & 3;
  for (int j = 0; j < full; ++j) c += popc64(ld_ro(&blk->last[j]));
  uint64_t m = ld_ro(&blk->last[full]);
  int keep = (p & 63) + 1;
  if (keep < 64) m &= (1ull << keep) - 1;
  return c + popc64(m);
}
MG_HD uint32_t select_last(const GraphView &g, int64_t k) {
  int64_t tot = g.total_last;
  if (k < 0 || k > tot) return kNull;
  if (k == tot) return (uint32_t)g.n;
  uint32_t b = ld_ro(&g.hint_last[k / kSelH]);
  for (;;) {
    const FBlock *blk = &g.fb[b];
    uint32_t rem = (uint32_t)k - ld_ro(&blk->rank_before);
    for (int j = 0; j < 4; ++j) {
      uint64_t m = ld_ro(&blk->last[j]);
      uint32_t c = popc64(m);
      if (rem < c) return (b << 8) + j * 64 + select64(m, (int)rem);
      rem -= c;
    }
    if (++b >= g.n_fblocks) return kNull;  // unreachable on a consistent index
  }
}

// ------------------------------------------------------------------ navigation (sdbg.h)
// sdbg.h:107-114.  e must be a valid edge (label 1..8); returns kNull for the `$` label.
MG_HD uint32_t forward(const GraphView &g, uint32_t e) {
  unsigned a = get_w(g, e);
  if (a > 4) a -= 4;
  if (a == 0) return kNull;
  int64_t cnt = rank_w(g, a, e);
  return select_last(g, g.rank_f[a] + cnt - 1);
}
// sdbg.h:79-86
MG_HD unsigned last_char_of(const GraphView &g, uint32_t x) {
  for (unsigned i = 1; i < 6; ++i)
    if (g.f[i] > (int64_t)x) return i - 1;
  return 6;
}
// sdbg.h:116-121
MG_HD uint32_t backward(const GraphView &g, uint32_t e) {
  unsigned a = last_char_of(g, e);
  if (a < 1 || a > 4) return kNull;
  int64_t cnt = (int64_t)rank_last(g, (int64_t)e - 1) - g.rank_f[a];
  return select_w(g, a, cnt);
}

// ComputeOutgoings<kFlagWriteOut> (sdbg.h:296-320): valid edges of the target node's group in
// DESCENDING id order; -1 if e is invalid.  `next_out`, when non-null, receives the group's last
// item (Forward(e)), the key of the navigation cache.
template <bool COHERENT = false>
MG_HD int outgoing(const GraphView &g, uint32_t e, uint32_t *out, uint32_t *next_out = nullptr) {
  if (get_inv<COHERENT>(g, e)) return -1;
  uint32_t next = forward(g, e);
  if (next_out) *next_out = next;
  if (next >= g.n) return 0;  // malformed-graph guard
  int deg = 0;
  do {
    if (!get_inv<COHERENT>(g, next)) {
      if (out && deg < 4) out[deg] = next;
      ++deg;
    }
    --next;
  } while (next != kNull && !get_lt(g, next));
  return deg;
}
// ComputeIncomings<kFlagWriteOut> (sdbg.h:245-287): ASCENDING id order; -1 if e is invalid.
// CHECK_SELF = false evaluates the neighbourhood of e regardless of e's own invalid bit.
template <bool COHERENT = false, bool CHECK_SELF = true>
MG_HD int incoming(const GraphView &g, uint32_t e, uint32_t *in) {
  if (CHECK_SELF && get_inv<COHERENT>(g, e)) return -1;
  uint32_t first = backward(g, e);
  if (first >= g.n) return 0;  // malformed-graph guard
  unsigned c = get_w(g, first);
  unsigned ones = get_lt(g, first);
  int deg = !get_inv<COHERENT>(g, first);
  if (deg > 0 && in) in[0] = first;
  for (uint64_t y = (uint64_t)first + 1; ones < 5 && y < g.n; ++y) {
    uint32_t yy = (uint32_t)y;
    ones += get_lt(g, yy);
    unsigned cur = get_w(g, yy);
    if (cur == c) break;
    if (cur == c + 4 && !get_inv<COHERENT>(g, yy)) {
      if (in && deg < 4) in[deg] = yy;
      ++deg;
    }
  }
  return deg;
}

// ------------------------------------------------------------------ index construction (per block)
// Count labels 1..4 in raw-W words [14*b, 14*b+14) restricted to items < n.
MG_HD void wblock_counts(const uint64_t *raw_w, uint64_t n, uint64_t b, uint32_t cnt[4], uint64_t words[kWWords]) {
  uint64_t n_words = (n + 15) / 16;
  cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0;
  for (int j = 0; j < kWWords; ++j) {
    uint64_t wi = b * kWWords + j;
    uint64_t x = wi < n_words ? raw_w[wi] : 0;
    uint64_t first_item = wi * 16;
    if (first_item + 16 > n) {  // mask the tail beyond n
      uint64_t keep = first_item < n ? (n - first_item) * 4 : 0;
      x = keep ? (x & ((1ull << keep) - 1)) : 0;
    }
    words[j] = x;
    for (unsigned a = 1; a <= 4; ++a) cnt[a - 1] += popc64(nib_eq(x, a));
  }
}
// 64-bit mask of items [64*wi, 64*wi+64) whose label is 0, restricted to items < n
MG_HD uint64_t w_zero_mask(const uint64_t *raw_w, uint64_t n, uint64_t wi) {
  uint64_t n_words = (n + 15) / 16;
  uint64_t out = 0;
  for (int q = 0; q < 4; ++q) {
    uint64_t widx = wi * 4 + q;
    if (widx >= n_words) break;
    uint64_t m = nib_eq(raw_w[widx], 0);  // bit 4i set iff nibble i == 0
    // compress bits 0,4,8,.. into 16 contiguous bits
    m = (m | (m >> 3)) & 0x0303030303030303ull;
    m = (m | (m >> 6)) & 0x000F000F000F000Full;
    m = (m | (m >> 12)) & 0x000000FF000000FFull;
    m = (m | (m >> 24)) & 0xFFFFull;
    out |= m << (16 * q);
  }
  uint64_t first_item = wi * 64;
  if (first_item + 64 > n) {
    uint64_t keep = first_item < n ? n - first_item : 0;
    out = keep ? (out & ((1ull << keep) - 1)) : 0;
  }
  return out;
}
MG_HD uint64_t tail_mask_word(const uint64_t *bits, uint64_t n, uint64_t wi) {
  uint64_t n_words = (n + 63) / 64;
  if (wi >= n_words) return 0;
  uint64_t x = bits[wi];
  uint64_t first_item = wi * 64;
  if (first_item + 64 > n) x &= (1ull << (n - first_item)) - 1;
  return x;
}

}  // namespace mg
