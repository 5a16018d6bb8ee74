// TEST INFRASTRUCTURE — NOT PRODUCT CODE, never loaded by mcaat_b200/.
//
// Host-side driver for the device functions of mcaat_b200/csrc/{sdbg_dev,search_dev}.cuh (they are
// written `__host__ __device__`): builds the same block layout in host memory with the same
// per-block routines and runs the same navigation / peel / admission-search / enumeration code one
// "thread" at a time, so that the exact logic the CUDA kernels execute can be checked against the
// oracle on a machine without a GPU.  The speculative window + ordered commit of enumerate.cu is
// mirrored here (emu_find_cycles) to check that it reproduces the sequential loop for any window.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../mcaat_b200/csrc/ball_dev.cuh"
#include "../../mcaat_b200/csrc/kmer_dev.cuh"

using namespace mg;

struct Emu {
  GraphView v;
  std::vector<WBlock> wb;
  std::vector<FBlock> fb;
  std::vector<uint16_t> mult;
  std::vector<uint32_t> hints;
  std::vector<uint64_t> adj;
  std::vector<uint32_t> visited;
  // k-mer lookup index (N3)
  KmerIndex kx{};
  std::vector<uint64_t> kx_tip;
  std::vector<uint32_t> kx_prefix, kx_labels;
  std::vector<int64_t> kx_first, kx_second;
  std::vector<uint32_t> kx_jump, kx_ok;
};

extern "C" {

Emu *emu_create(uint64_t n, const uint64_t *w, const uint64_t *last, const uint64_t *tip, const uint16_t *mult16,
                const int64_t *f) {
  Emu *e = new Emu();
  GraphView &v = e->v;
  memset(&v, 0, sizeof v);
  v.n = n;
  v.n_wblocks = (n + kWNib - 1) / kWNib;
  v.n_fblocks = (n + kFItems - 1) / kFItems;
  for (int i = 0; i < 6; ++i) v.f[i] = f[i];
  e->wb.resize(v.n_wblocks + 1);
  e->fb.resize(v.n_fblocks + 1);
  memset(e->wb.data(), 0, e->wb.size() * sizeof(WBlock));
  memset(e->fb.data(), 0, e->fb.size() * sizeof(FBlock));
  uint32_t run[4] = {0, 0, 0, 0};
  for (uint64_t b = 0; b < v.n_wblocks; ++b) {
    uint32_t c[4];
    uint64_t words[kWWords];
    wblock_counts(w, n, b, c, words);
    for (int j = 0; j < kWWords; ++j) e->wb[b].w[j] = words[j];
    for (int a = 0; a < 4; ++a) {
      e->wb[b].cnt[a] = run[a];
      run[a] += c[a];
    }
  }
  for (int a = 0; a < 4; ++a) v.total_w[a] = run[a];
  uint32_t lr = 0;
  for (uint64_t b = 0; b < v.n_fblocks; ++b) {
    e->fb[b].rank_before = lr;
    for (int j = 0; j < 4; ++j) {
      uint64_t wi = b * 4 + j;
      uint64_t l = tail_mask_word(last, n, wi), t = tail_mask_word(tip, n, wi);
      e->fb[b].last[j] = l;
      e->fb[b].lt[j] = l | t;
      e->fb[b].inv[j] = t | w_zero_mask(w, n, wi);
      lr += popc64(l);
    }
  }
  v.total_last = lr;
  size_t hl = (size_t)lr / kSelH + 2, off[5];
  off[0] = hl;
  for (int a = 0; a < 4; ++a) off[a + 1] = off[a] + (size_t)v.total_w[a] / kSelH + 2;
  e->hints.assign(off[4], 0);
  for (uint64_t b = 0; b < v.n_wblocks; ++b)
    for (int a = 0; a < 4; ++a) {
      uint32_t before = e->wb[b].cnt[a];
      uint32_t after = b + 1 < v.n_wblocks ? e->wb[b + 1].cnt[a] : v.total_w[a];
      for (uint32_t j = (before + kSelH - 1) / kSelH; (uint64_t)j * kSelH < after; ++j) e->hints[off[a] + j] = (uint32_t)b;
    }
  for (uint64_t b = 0; b < v.n_fblocks; ++b) {
    uint32_t before = e->fb[b].rank_before;
    uint32_t after = b + 1 < v.n_fblocks ? e->fb[b + 1].rank_before : lr;
    for (uint32_t j = (before + kSelH - 1) / kSelH; (uint64_t)j * kSelH < after; ++j) e->hints[j] = (uint32_t)b;
  }
  e->mult.assign(mult16, mult16 + n);
  v.wb = e->wb.data();
  v.fb = e->fb.data();
  v.mult = e->mult.data();
  v.hint_last = e->hints.data();
  for (int a = 0; a < 4; ++a) v.hint_w[a] = e->hints.data() + off[a];
  v.rank_f[0] = 0;
  for (int i = 1; i < 6; ++i) {
    int64_t pos = v.f[i] - 1;
    v.rank_f[i] = pos >= (int64_t)n ? -1 : pos == (int64_t)n - 1 ? (int64_t)lr : (int64_t)rank_last(v, pos);
  }
  e->visited.assign((n + 31) / 32 + 1, 0);
  return e;
}
void emu_free(Emu *e) { delete e; }

void emu_navigate(Emu *e, int8_t *outdeg, int8_t *indeg, uint64_t *out_ids, uint64_t *in_ids) {
  for (uint64_t i = 0; i < e->v.n; ++i) {
    uint32_t o[4] = {kNull, kNull, kNull, kNull}, in[4] = {kNull, kNull, kNull, kNull};
    int od = outgoing(e->v, (uint32_t)i, o), id = incoming(e->v, (uint32_t)i, in);
    outdeg[i] = (int8_t)od;
    indeg[i] = (int8_t)id;
    for (int j = 0; j < 4; ++j) {
      out_ids[i * 4 + j] = j < od ? o[j] : ~0ull;
      in_ids[i * 4 + j] = j < id ? in[j] : ~0ull;
    }
  }
}

void emu_rank_select_check(Emu *e, int64_t *bad) {  // rank/select round trips over every occurrence
  GraphView &v = e->v;
  *bad = 0;
  for (unsigned a = 1; a <= 4; ++a)
    for (uint32_t k = 0; k < v.total_w[a - 1]; ++k) {
      uint32_t p = select_w(v, a, k);
      if (p >= v.n || get_w(v, p) != a || rank_w(v, a, p) != k + 1) ++*bad;
    }
  for (uint32_t k = 0; k < v.total_last; ++k) {
    uint32_t p = select_last(v, k);
    if (p >= v.n || rank_last(v, p) != k + 1) ++*bad;
  }
}

// stats: tips_before, mult1, peeled, valid, tips_after.  Mirrors run_prune (prune.cu): the same
// per-item device functions in the same phase order; `reverse_order` permutes the work lists to
// show that the result does not depend on scheduling.
void emu_prune(Emu *e, uint64_t *stats, int reverse_order) {
  GraphView &v = e->v;
  const uint64_t n = v.n;
  const size_t words = (n + 31) / 32 + 1;
  std::vector<uint32_t> tipbits(words, 0), chain(words, 0), tpeeled(words, 0);
  std::vector<uint64_t> ent(n + 1);
  uint64_t tips0 = 0;
  if (reverse_order & 1) {  // CollectTips literally ...
    for (uint64_t i = 0; i < n; ++i)
      if (outgoing(v, (uint32_t)i, nullptr) == 0) { tipbits[i >> 5] |= 1u << (i & 31); ++tips0; }
  } else {                  // ... and from the dead ends' side (k_tips_dead_ends)
    for (uint64_t i = 0; i < n; ++i) {
      uint32_t in[4];
      const int deg = dead_end_incoming(v, (uint32_t)i, in);
      for (int j = 0; j < deg; ++j) tips0 += bm_set_atomic(tipbits.data(), in[j]) ? 1 : 0;
    }
  }
  stats[0] = tips0;
  uint64_t m1 = 0;
  for (uint64_t i = 0; i < n; ++i)
    if (v.mult[i] <= 1) { set_inv_atomic(v, (uint32_t)i); ++m1; }
  stats[1] = m1;
  e->adj.assign(n + 1, 0);
  PeelView P;
  P.adj = e->adj.data(); P.ent = ent.data(); P.chain = chain.data(); P.tpeeled = tpeeled.data();
  P.tipbits = tipbits.data(); P.dep = nullptr; P.dep_log = 0;
  P.split_ids = nullptr; P.sterm = nullptr; P.n_split = 0; P.split_shift = 25;
  std::vector<uint32_t> recl(words, 0);
  P.recl = recl.data();
  uint64_t pairs = 0;
  std::vector<uint32_t> work, next;
  for (uint64_t i = 0; i < n; ++i) {
    bool c, sd;
    int deg = chain_init_item(v, P, (uint32_t)i, c, sd);
    if (c) chain[i >> 5] |= 1u << (i & 31);
    if (sd) { tpeeled[i >> 5] |= 1u << (i & 31); work.push_back((uint32_t)i); }
    if (deg >= 2) pairs += deg;
  }
  // chain nodes -> terminals: splitter walks + jumping over splitters + resolution (as prune.cu), in
  // forward or reverse item order; reverse_order & 2 forces the exhaustive pointer jumping instead
  bool capped = false;
  if (!(reverse_order & 2)) {
    std::vector<uint32_t> split_ids;
    for (uint64_t i = 0; i < n; ++i)
      if (bm_get(P.chain, (uint32_t)i) && is_splitter_id(P, (uint32_t)i)) split_ids.push_back((uint32_t)i);
    // (the walks of the device run in any order: here ascending or descending splitter number)
    std::vector<uint32_t> sterm(split_ids.size() + 1, 0);
    P.split_ids = split_ids.data();
    P.sterm = sterm.data();
    P.n_split = (uint32_t)split_ids.size();
    const uint32_t ns = (uint32_t)split_ids.size();
    if (reverse_order & 1) { for (uint32_t k = ns; k-- > 0;) capped |= !split_walk_item(P, k); }
    else for (uint32_t k = 0; k < ns; ++k) capped |= !split_walk_item(P, k);
    for (int round = 0; round < 36; ++round) {  // pointer jumping on the dense table, as prune.cu
      bool any = false;
      if (reverse_order & 1) { for (uint32_t k = ns; k-- > 0;) any |= split_jump_dense(P, k); }
      else for (uint32_t k = 0; k < ns; ++k) any |= split_jump_dense(P, k);
      if (!any) break;
    }
    for (uint32_t k = 0; k < ns; ++k) split_settle(P, k);
    if (reverse_order & 1) { for (uint64_t i = n; i-- > 0;) capped |= !chain_resolve_item(P, (uint32_t)i); }
    else for (uint64_t i = 0; i < n; ++i) capped |= !chain_resolve_item(P, (uint32_t)i);
  }
  if (capped || (reverse_order & 2))
    for (int round = 0; round < 34; ++round) {
      bool any = false;
      if (reverse_order & 1) { for (uint64_t i = n; i-- > 0;) any |= chain_jump_item(P, (uint32_t)i); }
      else for (uint64_t i = 0; i < n; ++i) any |= chain_jump_item(P, (uint32_t)i);
      if (!any) break;
    }
  if (!(reverse_order & 4)) {  // unary branch nodes -> chain nodes (as prune.cu); reverse_order & 4 skips it
    for (int round = 0; round < 4; ++round) {
      std::vector<uint32_t> fresh(words, 0), newseed(words, 0), ids;
      uint64_t nf = 0, ns = 0;
      for (uint64_t i = 0; i < n; ++i) {
        int r = reclass_item(v, P, (uint32_t)i);
        if (r == 1) { fresh[i >> 5] |= 1u << (i & 31); ids.push_back((uint32_t)i); ++nf; }
        if (r == 2) { newseed[i >> 5] |= 1u << (i & 31); ++ns; }
      }
      if (reverse_order & 1) std::reverse(ids.begin(), ids.end());
      for (int r2 = 0; r2 < 36; ++r2) {
        bool any = false;
        for (uint32_t b : ids) any |= reclass_jump_item(P, fresh.data(), b);
        if (!any) break;
      }
      for (size_t w = 0; w < words; ++w) { chain[w] |= fresh[w]; recl[w] |= fresh[w]; tpeeled[w] |= newseed[w]; }
      if (nf + ns < 64) break;
    }
    // as prune.cu: the later passes follow re-classified terminals themselves (terminal_of); reverse_order & 8 moves the
    // pointers on in a pass of its own first (MCAAT_PEEL_RELINK=1)
    if (reverse_order & 8) {
      if (reverse_order & 1) { for (uint64_t i = n; i-- > 0;) relink_final_item(P, (uint32_t)i); }
      else for (uint64_t i = 0; i < n; ++i) relink_final_item(P, (uint32_t)i);
      P.recl = nullptr;
    }
    work.clear();
    for (uint64_t i = 0; i < n; ++i)
      if (bm_get(P.tpeeled, (uint32_t)i)) work.push_back((uint32_t)i);
  }
  uint32_t dep_log = 4;
  while (((uint64_t)1 << dep_log) < 2 * pairs + 16) ++dep_log;
  std::vector<uint64_t> dep((size_t)1 << dep_log, ~0ull);
  P.dep = dep.data(); P.dep_log = dep_log;
  for (uint64_t i = 0; i < n; ++i) dep_register_item(v, P, (uint32_t)i);
  if (reverse_order & 1) std::reverse(work.begin(), work.end());
  while (!work.empty()) {
    next.clear();
    for (uint32_t t : work) cpeel_one(P, t, [&](uint32_t b) { next.push_back(b); });
    if (reverse_order & 1) std::reverse(next.begin(), next.end());
    work.swap(next);
  }
  uint64_t peeled = 0, valid = 0, tips = 0;
  std::vector<uint8_t> pl(n + 1, 0);
  std::vector<uint64_t> fin(n + 1, 0);
  for (uint64_t i = 0; i < n; ++i) {
    bool p;
    fin[i] = peel_final_item(v, P, (uint32_t)i, p);
    pl[i] = p;
  }
  for (uint64_t i = 0; i < n; ++i) {
    bool ok = !pl[i] && !get_inv(v, (uint32_t)i);
    if (pl[i]) { set_inv_atomic(v, (uint32_t)i); ++peeled; }
    e->adj[i] = fin[i];
    valid += ok;
    tips += ok && (fin[i] >> 32) == 0;
  }
  stats[2] = peeled; stats[3] = valid; stats[4] = tips; stats[5] = capped;
}
void emu_invalid(Emu *e, uint8_t *out) {
  for (uint64_t i = 0; i < e->v.n; ++i) out[i] = get_inv(e->v, (uint32_t)i);
}

// regions.cu's host loop with the same per-word device functions; `reverse` visits the words of each level
// from the top, a different scheduling that must give the same set.  st = {added, levels, invalidated, kept}
void emu_keep_regions(Emu *e, const uint64_t *ids, uint64_t n_ids, uint64_t k, int reverse, uint64_t *st) {
  const uint64_t n = e->v.n, n_words64 = (n + 63) / 64, n_words = n_words64 * 2;
  std::vector<uint32_t> member(n_words + 4, 0), cur(n_words + 4, 0), next(n_words + 4, 0);
  for (uint64_t i = 0; i < n_ids; ++i)
    if (bm_set_atomic(member.data(), (uint32_t)ids[i])) bm_set_atomic(cur.data(), (uint32_t)ids[i]);
  uint64_t added = 0, level = 0;
  for (; level < k; ++level) {
    std::fill(next.begin(), next.end(), 0u);
    uint64_t a = 0;
    for (uint64_t j = 0; j < n_words; ++j) {
      uint64_t w = reverse ? n_words - 1 - j : j;
      if (cur[w]) a += region_expand_word(e->v, member.data(), cur.data(), next.data(), (uint32_t)w);
    }
    cur.swap(next);
    added += a;
    if (!a) { ++level; break; }
  }
  unsigned kept = 0;
  uint64_t gone = 0, kept_total = 0;
  for (uint64_t w = 0; w < n_words64; ++w) {
    kept = 0;
    gone += region_keep_word(e->v, member.data(), w, &kept);
    kept_total += kept;
  }
  st[0] = added; st[1] = level; st[2] = gone; st[3] = kept_total;
}

// reads.cu's attach_kmer_index + run_map_reads with the same device functions, one "thread" at a time
void emu_set_kmer_index(Emu *e, uint32_t k, const uint64_t *tip, const uint32_t *labels, uint64_t n_tips, uint32_t wptl,
                        const int64_t *first, const int64_t *second, uint64_t n_buckets, int use_jump) {
  const uint64_t n_words = (e->v.n + 63) / 64;
  e->kx_tip.assign(tip, tip + n_words);
  e->kx_prefix.assign(n_words + 1, 0);
  uint32_t acc = 0;
  for (uint64_t i = 0; i < n_words; ++i) { e->kx_prefix[i] = acc; acc += (uint32_t)__builtin_popcountll(tip[i]); }
  e->kx_labels.assign(labels, labels + n_tips * wptl);
  e->kx_labels.push_back(0);
  e->kx_first.assign(first, first + n_buckets);
  e->kx_second.assign(second, second + n_buckets);
  e->kx = KmerIndex{e->kx_tip.data(), e->kx_prefix.data(), e->kx_labels.data(), wptl, (uint32_t)n_buckets,
                    e->kx_first.data(), e->kx_second.data(), k, nullptr, nullptr};
  if (use_jump) {  // the same construction as attach_kmer_index (reads.cu): doubling + bucket consistency bits
    const uint64_t n = e->v.n;
    std::vector<uint32_t> a(n + 1), b(n + 1);
    for (uint64_t x = 0; x < n; ++x) b[x] = jump_init(e->v, e->kx, (uint32_t)x);
    for (uint64_t x = 0; x < n; ++x) a[x] = jump_double(b.data(), (uint32_t)x);
    for (uint64_t x = 0; x < n; ++x) b[x] = jump_double(a.data(), (uint32_t)x);
    for (uint64_t x = 0; x < n; ++x) a[x] = jump_double(b.data(), (uint32_t)x);
    e->kx_jump.swap(a);
    int chars = 0;
    while ((1u << (chars * 2)) < n_buckets) ++chars;
    e->kx_ok.assign((n_buckets + 31) / 32 + 1, 0);
    for (uint64_t bk = 0; bk < n_buckets; ++bk)
      if (bucket_consistent(e->v, e->kx, (uint32_t)bk, chars)) e->kx_ok[bk >> 5] |= 1u << (bk & 31);
    e->kx.jump = e->kx_jump.data();
    e->kx.bucket_ok = e->kx_ok.data();
  }
}
// number of items whose probe may take the shortcut / buckets marked consistent (test diagnostics)
void emu_kmer_index_stats(Emu *e, uint64_t *out) {
  uint64_t j = 0, ok = 0;
  for (uint32_t v : e->kx_jump) j += !(v & kJumpLiteral);
  for (uint32_t w : e->kx_ok) ok += (uint64_t)__builtin_popcount(w);
  out[0] = j; out[1] = ok;
}
// ids of one read -> out (capacity >= len); returns the count (0 = dropped); kNull ids become ~0
uint64_t emu_map_read(Emu *e, const char *seq, uint32_t len, int reverse, const uint32_t *member, uint64_t *out) {
  uint32_t a = kNull, b = kNull;
  uint32_t n_out = read_ends(e->v, e->kx, seq, len, reverse != 0, member, &a, &b);
  for (uint32_t p = 0; p < n_out; ++p) {
    uint32_t id = read_id_at(e->v, e->kx, seq, len, reverse != 0, p, n_out, a, b);
    out[p] = id == kNull ? ~0ull : id;
  }
  return n_out;
}

uint64_t emu_candidates(Emu *e, uint64_t thr, uint64_t *ids) {
  uint64_t c = 0;
  for (uint64_t i = 0; i < e->v.n; ++i)
    if (is_candidate(e->v, (uint32_t)i, thr)) ids[c++] = i;
  return c;
}
int emu_bucket(Emu *e, uint64_t id) { return log2_bucket(e->v.mult[id]); }

// tier 1 hash set with the given log2 capacity; returns the dls_run code
int emu_dls(Emu *e, uint64_t start, int limit, int log2cap, uint64_t *expansions) {
  std::vector<uint32_t> tab((size_t)1 << log2cap, kNull), sn(4 * limit + 8);
  std::vector<uint16_t> sd(4 * limit + 8);
  U32HashSet s{tab.data(), (uint32_t)log2cap, 0, (1u << log2cap) / 2};
  uint32_t exp = 0;
  std::vector<uint64_t> sa(sn.size());
  int r = dls_run(e->adj.data(), (uint32_t)start, limit, s, sn.data(), sd.data(), sa.data(), (uint32_t)sn.size(), exp);
  if (expansions) *expansions += exp;
  return r;
}
int emu_dls_bitmap(Emu *e, uint64_t start, int limit) {
  std::vector<uint32_t> bits((e->v.n + 31) / 32 + 1, 0), sn(4 * limit + 8);
  std::vector<uint16_t> sd(4 * limit + 8);
  BitSet s{bits.data()};
  uint32_t exp = 0;
  std::vector<uint64_t> sa(sn.size());
  return dls_run(e->adj.data(), (uint32_t)start, limit, s, sn.data(), sd.data(), sa.data(), (uint32_t)sn.size(), exp);
}

// DepthLevelSearch on the extracted, contracted ball (dball_*: the middle tier of run_dls, starts.cu); returns the
// dls_run code (kDlsSetFull when the ball does not fit 2^ball_log / 2 nodes)
int emu_dls_ball(Emu *e, uint64_t start, int limit, uint32_t ball_log, uint64_t *expansions, uint64_t *n_chains);

// 0 = hash-map search (enum_run) only, 1 = extracted-ball search node by node (ball_dfs) with enum_run as overflow
// fallback, 2 = the same on the contracted ball (sball_dfs; what k_enum_pipeline runs)
static int g_algo = 2;
static uint32_t g_ball_log = 16;  // ball index slots (node capacity = half)
static int g_max_steps = kMaxSteps;  // step cap (lowered by the differential tests)
static uint32_t g_chain_cap = 0;     // > 0: capacity of the contracted ball's record array (overflow tests)

struct BallSlot {
  std::vector<LNode> nodes;
  std::vector<LTail> tails;
  std::vector<uint64_t> index;
  std::vector<Lvl> lvl;
  std::vector<uint32_t> relax, cyc_off, spos, ckeys, cyc_pos, cyc_noff, flat, relax_top;
  std::vector<uint16_t> cyc, soff, shead;
  std::vector<SHot> shot;
  std::vector<uint64_t> sin4;
  BallArena A;
  SBall B;
  RecsPtr recs() const { return RecsPtr{B.hot, B.in4, A.lvl, A.relax_top}; }
  uint32_t live = 0;  // nodes of the last build (their index slots are still occupied)
  void init(int M, uint32_t index_log) {
    uint32_t cap = std::min<uint32_t>((1u << index_log) / 2, 65535u);
    nodes.resize(cap);
    tails.resize(cap);
    memset((void *)nodes.data(), 0x77, nodes.size() * sizeof(LNode));
    index.assign((size_t)1 << index_log, kIndexEmpty);
    relax.resize(std::max<uint32_t>(cap, 32768));
    cyc.resize((size_t)kClusterBound * (M + 1));
    cyc_off.resize(kClusterBound + 1);
    lvl.resize(M + 4);
    A.lvl = lvl.data();
    A.nodes = nodes.data(); A.tails = tails.data(); A.node_cap = cap; A.index = index.data(); A.index_log = index_log;
    A.relax = relax.data(); A.relax_cap = (uint32_t)relax.size(); A.cyc = cyc.data(); A.cyc_cap = (uint32_t)cyc.size();
    A.cyc_off = cyc_off.data();
    spos.resize(cap); ckeys.resize(cap); soff.resize(cap + 1); shead.resize(cap); cyc_pos.resize(cyc.size());
    cyc_noff.resize(kClusterBound + 1); flat.resize(cyc.size());
    shot.resize(cap + 1); sin4.resize(cap + 1);
    memset((void *)shot.data(), 0x55, shot.size() * sizeof(SHot));
    A.spos = spos.data(); A.ckeys = ckeys.data(); A.soff = soff.data(); A.shead = shead.data();
    A.cyc_pos = cyc_pos.data(); A.cyc_noff = cyc_noff.data(); A.max_steps = g_max_steps;
    relax_top.resize(3); A.relax_top = relax_top.data(); A.relax_top_cap = 3;  // tiny: the spill path is exercised
    B.hot = shot.data(); B.in4 = sin4.data(); B.cap = g_chain_cap ? g_chain_cap : cap + 1; B.n = 0;
  }
};

struct EmuSlot {
  std::vector<uint32_t> path, pent, index, cyc_nodes, cyc_off;
  std::vector<int32_t> bl;
  std::vector<USet> stk;
  std::vector<HotEnt> hot;
  std::vector<ColdEnt> cold;
  std::vector<uint64_t> relax;
  EnumArena A;
  EnumResult R;
  void init(int M, uint32_t index_log, uint32_t relax_cap, uint32_t cyc_cycles) {
    path.resize(M + 4); pent.resize(M + 4); bl.resize(M + 4); stk.resize(M + 4);
    index.assign((size_t)1 << index_log, 0x5A5A5A5Au);  // deliberately not cleared between searches
    hot.resize(((size_t)1 << index_log) / 2);
    cold.resize(hot.size());
    memset(hot.data(), 0x33, hot.size() * sizeof(HotEnt));
    relax.resize(relax_cap);
    cyc_nodes.resize((size_t)cyc_cycles * (M + 1));
    cyc_off.resize(kClusterBound + 1);
    A.path = path.data(); A.pent = pent.data(); A.bl = bl.data(); A.stk = stk.data();
    A.index = index.data(); A.index_log = index_log; A.hot = hot.data(); A.cold = cold.data();
    A.ent_cap = (uint32_t)hot.size(); A.relax = relax.data(); A.relax_cap = relax_cap;
    A.cyc_nodes = cyc_nodes.data(); A.cyc_nodes_cap = (uint32_t)cyc_nodes.size(); A.cyc_off = cyc_off.data();
  }
};

// Mirrors the ordered-commit pipeline of enumerate.cu with `window` workers: ticket t is searched
// against a LAGGING view of `visited` (the state `window` commits ago) that catches up one commit at
// every poll — i.e. the bitmap changes under the running search, as it does on the device — and is
// then committed in order against the exact bitmap (skip / conflict -> re-run / accept).
// Outputs: searched starts, per-search cycle counts, flat cycle lengths and nodes (caller-sized
// generously), stats = {skipped, bound, cap, reruns, abandoned, overflow_reruns}.
int emu_find_cycles(Emu *e, const uint64_t *starts, uint64_t n, int min_len, int max_len, uint32_t window,
                    uint32_t lock_log, uint64_t *searched, uint64_t *n_searched, uint64_t *cyc_count,
                    uint64_t *cyc_len, uint64_t *n_cycles, uint64_t *nodes, uint64_t *n_nodes, uint64_t *stats) {
  GraphView &v = e->v;
  std::fill(e->visited.begin(), e->visited.end(), 0u);
  uint32_t *vis = e->visited.data();
  std::vector<uint32_t> view(e->visited.size(), 0u);
  std::vector<std::vector<uint32_t>> marks(n + 1);  // nodes marked by the commit of ticket i
  uint64_t view_idx = 0;                            // commits [0, view_idx) are applied to `view`
  *n_searched = *n_cycles = *n_nodes = 0;
  for (int i = 0; i < 6; ++i) stats[i] = 0;
  EmuSlot sl;
  sl.init(max_len, lock_log, 32768, kClusterBound);
  BallSlot bs;
  if (g_algo >= 1) bs.init(max_len, g_ball_log);
  std::vector<uint32_t> keys, cyc_nodes_flat, cyc_lens;  // result of the last search, whatever ran it
  EnumResult RR;
  // one search against bitmap `bm`: fills RR, keys (examined nodes), cyc_lens / cyc_nodes_flat
  auto collect_legacy = [&](EmuSlot *c) {
    RR = c->R;
    keys.clear();
    for (uint32_t j = 0; j < c->R.n_keys; ++j) keys.push_back(c->hot[j].key);
    cyc_lens.clear();
    for (uint32_t j = 0; j < c->R.n_cycles; ++j) cyc_lens.push_back(c->cyc_off[j + 1] - c->cyc_off[j]);
    cyc_nodes_flat.assign(c->cyc_nodes.begin(), c->cyc_nodes.begin() + c->R.n_nodes);
  };
  auto search = [&](const uint32_t *bm, uint32_t s, auto poll) {
    if (g_algo >= 1) {
      Coop co;
      ball_clear(bs.A, bs.live, co);
      uint32_t nn = 0;
      int st = ball_build(v, e->adj.data(), bm, s, max_len, bs.A, nn, co);
      bs.live = nn;
      RR.status = st;
      RR.n_cycles = RR.n_nodes = RR.n_keys = RR.steps = RR.expansions = 0;
      if (st == kEnumDone && g_algo == 2) {
        st = sball_contract(bs.A, nn, max_len, bs.B, co);
        RR.status = st;
        if (st == kEnumDone) {
          SDfsState ds;
          while (!sball_dfs(min_len, max_len, bs.A, bs.recs(), ds, 64, RR))
            if (poll()) {
              RR.status = kEnumSkipped;
              RR.n_cycles = RR.n_nodes = 0;
              break;
            }
          keys.clear();
          cyc_lens.clear();
          cyc_nodes_flat.clear();
          if (RR.status != kEnumOverflow) {
            for (uint32_t j = 0; j < nn; ++j) keys.push_back(bs.nodes[j].key);
            RR.n_keys = nn;
            if (RR.status != kEnumSkipped) {
              sball_expand(bs.A, bs.B, ds.n_staged, bs.flat.data(), co);
              for (uint32_t j = 0; j < RR.n_cycles; ++j) cyc_lens.push_back(bs.cyc_noff[j + 1] - bs.cyc_noff[j]);
              cyc_nodes_flat.assign(bs.flat.begin(), bs.flat.begin() + RR.n_nodes);
            }
          }
          return;
        }
      }
      if (st == kEnumDone) {
        DfsState ds;
        while (!ball_dfs(min_len, max_len, bs.A, ds, 256, RR))
          if (poll()) {  // the device polls between slices: start marked -> the search is abandoned
            RR.status = kEnumSkipped;
            RR.n_cycles = RR.n_nodes = 0;
            break;
          }
      }
      keys.clear();
      cyc_lens.clear();
      cyc_nodes_flat.clear();
      if (RR.status != kEnumOverflow) {
        for (uint32_t j = 0; j < nn; ++j) keys.push_back(bs.nodes[j].key);
        RR.n_keys = nn;
        for (uint32_t j = 0; j < RR.n_cycles; ++j) cyc_lens.push_back(bs.cyc_off[j + 1] - bs.cyc_off[j]);
        for (uint32_t j = 0; j < RR.n_nodes; ++j) cyc_nodes_flat.push_back(bs.nodes[bs.cyc[j]].key);
      }
      return;
    }
    enum_run(v, e->adj.data(), bm, s, min_len, max_len, sl.A, sl.R, poll);
    collect_legacy(&sl);
  };
  auto advance_view = [&](uint64_t upto) {
    while (view_idx < upto) {
      for (uint32_t x : marks[view_idx]) view[x >> 5] |= 1u << (x & 31);
      ++view_idx;
    }
  };
  for (uint64_t t = 0; t < n; ++t) {
    uint32_t s = (uint32_t)starts[t];
    advance_view(t > window ? t - window : 0);
    EmuSlot big;
    RR.status = kEnumSkipped;
    RR.n_keys = RR.n_nodes = RR.n_cycles = 0;
    if (!visited_bit(view.data(), s)) {
      search(view.data(), s, [&]() {
        advance_view(std::min<uint64_t>(view_idx + 1, t));
        return visited_bit(view.data(), s);
      });
      if (RR.status == kEnumSkipped) ++stats[4];
    }
    // ---- commit at position t: `vis` is the exact bitmap
    if (RR.status == kEnumSkipped || visited_bit(vis, s)) { ++stats[0]; continue; }
    uint32_t ll = lock_log + 4, rc = 1u << 16;
    bool solo = false;  // after an overflow the ticket is repeated with the hash-map search in a large arena
    for (;;) {
      if (RR.status == kEnumOverflow) {
        if (ll >= 26) return -1;
        big = EmuSlot();
        big.init(max_len, ll, rc, kClusterBound);
        solo = true;
        enum_run(v, e->adj.data(), vis, s, min_len, max_len, big.A, big.R, []() { return false; });
        collect_legacy(&big);
        ++stats[5];
        ll += 2; rc *= 4;
        continue;
      }
      bool conflict = false;
      for (uint32_t k : keys)
        if (k != s && visited_bit(vis, k)) { conflict = true; break; }
      if (!conflict) break;
      ++stats[3];
      if (solo) {
        enum_run(v, e->adj.data(), vis, s, min_len, max_len, big.A, big.R, []() { return false; });
        collect_legacy(&big);
      } else {
        search(vis, s, []() { return false; });
      }
    }
    for (uint32_t x : cyc_nodes_flat) {
      vis[x >> 5] |= 1u << (x & 31);
      marks[t].push_back(x);
    }
    searched[*n_searched] = s;
    cyc_count[*n_searched] = RR.n_cycles;
    ++*n_searched;
    for (uint32_t l : cyc_lens) cyc_len[(*n_cycles)++] = l;
    for (uint32_t x : cyc_nodes_flat) nodes[(*n_nodes)++] = x;
    stats[1] += RR.status == kEnumBound;
    stats[2] += RR.status == kEnumCap;
  }
  return 0;
}

int emu_dls_ball(Emu *e, uint64_t start, int limit, uint32_t ball_log, uint64_t *expansions, uint64_t *n_chains) {
  BallSlot bs;
  bs.init(limit, ball_log);
  Coop co;
  uint32_t nn = 0;
  int r = dball_build(e->adj.data(), (uint32_t)start, limit, bs.A, nn, co);
  if (n_chains) *n_chains = 0;
  if (r == kDlsSetFull) return r;
  std::vector<DChain> C(nn + 1);
  uint32_t nc = dball_contract(bs.A, nn, C.data(), co);
  if (n_chains) *n_chains = nc;
  std::vector<uint32_t> stk(4 * limit + 8);
  uint32_t exp = 0;
  r = dball_dfs(C.data(), limit, stk.data(), (uint32_t)stk.size(), exp);
  if (expansions) *expansions += exp;
  return r;
}

void emu_set_algo(int algo, uint32_t ball_log) {
  g_algo = algo;
  g_ball_log = ball_log;
}
// The contracted ball of `start` (visited all clear): hot[2 * i], hot[2 * i + 1] = the 16 bytes of chain record i,
// in4[i] its in-set.  Returns the number of chains (0 on overflow).  For the device micro-benchmark of the search.
uint64_t emu_contracted_ball(Emu *e, uint64_t start, int max_len, uint64_t *hot, uint64_t *in4, uint64_t cap) {
  std::fill(e->visited.begin(), e->visited.end(), 0u);
  BallSlot bs;
  bs.init(max_len, g_ball_log);
  Coop co;
  uint32_t nn = 0;
  if (ball_build(e->v, e->adj.data(), e->visited.data(), (uint32_t)start, max_len, bs.A, nn, co) != kEnumDone) return 0;
  if (sball_contract(bs.A, nn, max_len, bs.B, co) != kEnumDone || bs.B.n + 1 > cap) return 0;
  memcpy(hot, bs.B.hot, (size_t)(bs.B.n + 1) * 16);  // chains + the dummy record
  memcpy(in4, bs.B.in4, (size_t)(bs.B.n + 1) * 8);
  return bs.B.n + 1;
}

// The navigation cache of the pruned graph (n entries), for the device micro-benchmark of the ball extraction.
void emu_adj(Emu *e, uint64_t *out) { memcpy(out, e->adj.data(), (size_t)e->v.n * 8); }

void emu_set_search_limits(int max_steps, uint32_t chain_cap) {
  g_max_steps = max_steps > 0 ? max_steps : kMaxSteps;
  g_chain_cap = chain_cap;
}

// One search from `start` against the current invalid/visited state (visited = all clear) with the node-by-node
// ball search (algo 1) or the contracted one (algo 2): status, counters and the flat cycles, for differential tests.
// out_stats = {status, n_cycles, n_nodes, steps, expansions, n_ball, n_chains}
int emu_search_one(Emu *e, uint64_t start, int min_len, int max_len, int algo, uint64_t *out_stats, uint64_t *cyc_len,
                   uint64_t *nodes) {
  std::fill(e->visited.begin(), e->visited.end(), 0u);
  BallSlot bs;
  bs.init(max_len, g_ball_log);
  Coop co;
  uint32_t nn = 0;
  EnumResult RR;
  RR.n_cycles = RR.n_nodes = RR.n_keys = RR.steps = RR.expansions = 0;
  RR.status = ball_build(e->v, e->adj.data(), e->visited.data(), (uint32_t)start, max_len, bs.A, nn, co);
  out_stats[5] = nn;
  out_stats[6] = 0;
  if (RR.status == kEnumDone) {
    if (algo == 2) {
      RR.status = sball_contract(bs.A, nn, max_len, bs.B, co);
      out_stats[6] = bs.B.n;
      if (RR.status == kEnumDone) {
        SDfsState ds;
        while (!sball_dfs(min_len, max_len, bs.A, bs.recs(), ds, 7, RR)) {}
        sball_expand(bs.A, bs.B, ds.n_staged, bs.flat.data(), co);
        for (uint32_t j = 0; j < RR.n_cycles; ++j) cyc_len[j] = bs.cyc_noff[j + 1] - bs.cyc_noff[j];
        for (uint32_t j = 0; j < RR.n_nodes; ++j) nodes[j] = bs.flat[j];
      }
    } else {
      DfsState ds;
      while (!ball_dfs(min_len, max_len, bs.A, ds, 7, RR)) {}
      for (uint32_t j = 0; j < RR.n_cycles; ++j) cyc_len[j] = bs.cyc_off[j + 1] - bs.cyc_off[j];
      for (uint32_t j = 0; j < RR.n_nodes; ++j) nodes[j] = bs.nodes[bs.cyc[j]].key;
    }
  }
  out_stats[0] = (uint64_t)RR.status; out_stats[1] = RR.n_cycles; out_stats[2] = RR.n_nodes;
  out_stats[3] = RR.steps; out_stats[4] = RR.expansions;
  return 0;
}

}  // extern "C"
