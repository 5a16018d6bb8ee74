// Search logic of the hot path as host+device inline functions: tip peeling, the start-node
// predicate, the depth-limited admission search and the bounded multicycle enumeration.
// Each function cites the reference lines (relative to /root/reference) whose results it
// reproduces bit-exactly.  Kernels in prune.cu / starts.cu / enumerate.cu drive these per
// thread / per warp; tests/emu drives the same code on the host against the oracle.
#pragma once
#include "sdbg_dev.cuh"

namespace mg {

// ------------------------------------------------------------------ navigation cache
// After pruning the graph is static (nothing on the path invalidates edges any more,
// src/cycle_finder.cpp:458-487), so Forward(e) and the validity of the target group are
// materialised once per valid item: low 32 bits = Forward(e), bits 32..39 = validity mask of
// items Forward(e)-i.  One 8-byte load then replaces rank + select + group scan.
MG_HD uint64_t adj_encode(const GraphView &g, uint32_t e) {
  if (get_inv(g, e)) return 0;
  uint32_t next = forward(g, e);
  if (next >= g.n) return 0;
  uint64_t mask = 0;
  uint32_t x = next;
  int i = 0;
  do {
    if (!get_inv(g, x) && i < 8) mask |= 1ull << i;
    --x;
    ++i;
  } while (x != kNull && !get_lt(g, x));
  return (uint64_t)next | (mask << 32);
}
// OutgoingEdges (sdbg.h:354) from the cache: valid out-neighbours, descending id
MG_HD int adj_outgoing(uint64_t a, uint32_t *out) {
  uint32_t next = (uint32_t)a;
  unsigned mask = (unsigned)(a >> 32) & 0xFFu;
  int deg = 0;
  for (int i = 0; mask; ++i, mask >>= 1)
    if (mask & 1) {
      if (deg < 4) out[deg] = next - i;
      ++deg;
    }
  return deg;
}

MG_HD void prefetch_l2(const void *p) {  // into L1 (and with it L2)
#if defined(__CUDA_ARCH__)
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
  (void)p;
#endif
}
MG_HD uint32_t hash_u32_fwd(uint32_t k, uint32_t log2cap) { return (k * 2654435761u) >> (32 - log2cap); }

// ------------------------------------------------------------------ atomics / fences
MG_HD bool set_inv_atomic(const GraphView &g, uint32_t i) {
  uint64_t *p = &g.fb[i >> 8].inv[(i >> 6) & 3];
  uint64_t bit = 1ull << (i & 63);
#if defined(__CUDA_ARCH__)
  unsigned long long old = atomicOr((unsigned long long *)p, (unsigned long long)bit);
  return !(old & bit);
#else
  uint64_t old = *p;
  *p = old | bit;
  return !(old & bit);
#endif
}
MG_HD void fence_sc() {
#if defined(__CUDA_ARCH__)
  asm volatile("fence.sc.gpu;" ::: "memory");
#endif
}

// ------------------------------------------------------------------ peel (RecursiveReduction)
// src/cycle_finder.cpp:359-371 driven as at :440-442 computes a closure: starting from the tips
// collected BEFORE the multiplicity filter, a valid node is removed when all its valid
// out-neighbours have been removed (the set does not depend on the visiting order).  The reference
// walks it one node per recursion level; chains of 10^5..10^6 nodes make that a serial pointer
// chase.  Here the closure is computed by chain contraction on the post-filter graph:
//   1. every valid node with exactly one valid out-neighbour is a CHAIN node pointing at it; every
//      other valid node is a terminal (sink: out-degree 0, branch: >= 2) pointing at itself;
//   2. pointer jumping (ptr[v] = ptr[ptr[v]]) resolves each chain node to its terminal in
//      O(log chain length) data-parallel rounds (unresolvable = a cycle of chain nodes, never peeled);
//   3. a chain node is peeled iff its terminal is; a sink is peeled iff it is one of the collected
//      tips; a branch node is peeled iff the terminals of all its valid out-neighbours are —
//      evaluated by an asynchronous work list over a (terminal -> dependent branch nodes) multimap.
struct PeelView {
  uint64_t *adj;            // per item: Forward | validity mask << 32 on the post-filter graph
  uint64_t *ent;            // per item: chain pointer (low word) | owner: the splitter whose walk passed the item, step 2, as
                            // its position in split_ids (bits 32..62, kNoOwner = none) | item was classified a chain node
                            // by step 1 (bit 63).  One 8-byte entry = one sector per hop of a walk.
  const uint32_t *split_ids;  // the splitters (ascending item ids)
  uint32_t *sterm;          // per splitter: its pointer after step 2b — 4 bytes per 64 chain nodes, so the table the
                            // resolution pass reads at random stays in L2 (the splitters' own entries, spread over the
                            // whole entry array, do not: 64 B of DRAM traffic per resolved item)
  uint32_t *chain;          // bitmap: chain nodes
  uint32_t *tpeeled;        // bitmap: peeled terminals
  const uint32_t *tipbits;  // bitmap: CollectTips on the loaded graph (:436)
  const uint32_t *recl;     // bitmap: branch nodes re-classified as chain nodes by step 2b (all rounds so far)
  uint64_t *dep;            // multimap entries (terminal << 32 | branch node), empty = ~0
  uint32_t dep_log;
  uint32_t n_split;         // number of splitters
  uint32_t split_shift;     // one chain node in 2^(32 - split_shift) is a splitter (25: one in 128, the default)
};
constexpr uint32_t kNoOwner = 0x7FFFFFFFu, kEntChain = 0x80000000u;
MG_HD uint32_t &peel_ptr(const PeelView &P, uint32_t v) { return ((uint32_t *)P.ent)[2 * (size_t)v]; }
MG_HD bool bm_get(const uint32_t *bm, uint32_t i) { return (bm[i >> 5] >> (i & 31)) & 1; }
MG_HD bool bm_get_coherent(const uint32_t *bm, uint32_t i) { return (ld_cg32(&bm[i >> 5]) >> (i & 31)) & 1; }
// returns true when this call set the bit
MG_HD bool bm_set_atomic(uint32_t *bm, uint32_t i) {
  uint32_t bit = 1u << (i & 31);
#if defined(__CUDA_ARCH__)
  return !(atomicOr(&bm[i >> 5], bit) & bit);
#else
  uint32_t old = bm[i >> 5];
  bm[i >> 5] = old | bit;
  return !(old & bit);
#endif
}
MG_HD int ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __ffs((int)x) - 1;
#else
  return __builtin_ctz(x);
#endif
}
// Step 1 for one item.  Outputs the per-item flags (the caller packs them into bitmap words) and
// returns the out-degree on the post-filter graph (-1 for an invalid item).
MG_HD int chain_init_item(const GraphView &g, const PeelView &P, uint32_t v, bool &is_chain, bool &is_seed) {
  uint64_t a = adj_encode(g, v);
  P.adj[v] = a;
  is_chain = is_seed = false;
  uint32_t ptr = v;
  int deg = -1;
  if (!get_inv(g, v)) {
    uint32_t mask = (uint32_t)(a >> 32) & 0xFFu;
    deg = popc32(mask);
    if (deg == 1) {
      is_chain = true;
      ptr = (uint32_t)a - (uint32_t)ctz32(mask);
    } else if (deg == 0) {
      is_seed = bm_get(P.tipbits, v);  // a collected tip that survived the multiplicity filter
    }
  }
  P.ent[v] = (uint64_t)ptr | ((uint64_t)(kNoOwner | (is_chain ? kEntChain : 0u)) << 32);
  return deg;
}
// Step 2, one round for one item; returns true while the item is still unresolved.
MG_HD bool chain_jump_item(const PeelView &P, uint32_t v) {
  if (!bm_get(P.chain, v)) return false;
  uint32_t p = peel_ptr(P, v);
  if (!bm_get(P.chain, p)) return false;
  peel_ptr(P, v) = peel_ptr(P, p);
  return true;
}
// Step 2, work-efficient form.  Pointer jumping over ALL chain nodes costs one pass over the graph
// per doubling (18 passes for the 10^5-node chains of sequencing-error tips).  Instead one chain
// node in 128 (by hash; PeelView::split_shift) is a SPLITTER:
//   a. every splitter walks forward until it meets a terminal, another splitter or an item another
//      walk has claimed, claiming the items it passes (owner[]) — every chain node is claimed at most
//      once, so this is O(n) in total, in walks of ~128 hops;
//   b. pointer jumping runs over the splitters only (1/128 of the nodes);
//   c. every other chain node takes the terminal of its owner, or — near the head of a chain, where no
//      walk passed — walks to the first item that knows it.
// An item that leads into a cycle of chain nodes ends with a chain node as its pointer, which
// terminal_of() reports as unresolvable, exactly like the exhaustive pointer jumping.  Cycles without
// a splitter are recognised by the walks themselves (step a: the walk meets its own claim; step c: Brent's
// checkpoint).  Walks are capped; hitting the cap (a splitter-free stretch of 4096+ chain nodes) makes the
// host fall back to exhaustive pointer jumping — every pointer stays a valid forward pointer throughout.
constexpr uint32_t kWalkCapSplitter = 1u << 16, kWalkCapHead = 1u << 12;
MG_HD bool is_splitter_id(const PeelView &P, uint32_t v) { return ((v * 0x9E3779B1u) >> P.split_shift) == 0; }
MG_HD uint32_t ld_cg_u32(const uint32_t *p) { return ld_cg32(p); }
MG_HD uint64_t ent_cas(uint64_t *p, uint64_t expect, uint64_t val) {
#if defined(__CUDA_ARCH__)
  return (uint64_t)atomicCAS((unsigned long long *)p, (unsigned long long)expect, (unsigned long long)val);
#else
  uint64_t old = *p;
  if (old == expect) *p = val;
  return old;
#endif
}
// number of the chain splitter with item id p (binary search in the ascending id list)
MG_HD uint32_t split_number(const PeelView &P, uint32_t p) {
  uint32_t lo = 0, hi = P.n_split;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (ld_ro(&P.split_ids[mid]) < p) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
// During step b the dense table holds, per splitter, where its walk ended: kSplitTerm | item id of a terminal, or the
// NUMBER of the chain splitter ahead (its own number: known to be unresolvable) — 4 bytes per 128 chain nodes, so the
// pointer-jumping rounds run in L2 instead of on the splitters' entries scattered over the whole entry array.
constexpr uint32_t kSplitTerm = 0x80000000u;
// a. walk of splitter number k; returns false when the hop cap was hit (ptr[s] is then left untouched)
MG_HD bool split_walk_item(const PeelView &P, uint32_t k) {
  const uint32_t s = P.split_ids[k];
  uint32_t p = peel_ptr(P, s), next = kNull;
  for (uint32_t hops = 0; hops < kWalkCapSplitter; ++hops) {
    if (is_splitter_id(P, p)) { next = p; break; }      // another splitter (or a terminal whose id hashes like one)
    const uint64_t e = ld_cg64(&P.ent[p]);
    const uint32_t hi = (uint32_t)(e >> 32);
    if (!(hi & kEntChain)) { next = p; break; }      // terminal
    if ((hi & kNoOwner) != kNoOwner) { next = P.split_ids[hi & kNoOwner]; break; }  // claimed; by k itself: a splitter-free cycle
    // nobody writes the pointer of an unclaimed non-splitter chain node during the walks, so a failed exchange
    // means another walk has claimed the item in between
    const uint64_t old = ent_cas(&P.ent[p], e, (e & 0x80000000FFFFFFFFull) | ((uint64_t)k << 32));
    if (old != e) { next = P.split_ids[(uint32_t)(old >> 32) & kNoOwner]; break; }
    p = (uint32_t)e;
  }
  if (next == kNull) { P.sterm[k] = k; return false; }  // (the caller falls back to exhaustive pointer jumping)
  peel_ptr(P, s) = next;
  P.sterm[k] = bm_get(P.chain, next) ? split_number(P, next) : (kSplitTerm | next);  // a walk only ever ends at a splitter or a terminal
  return true;
}
// b. one pointer-jumping round for splitter number k on the dense table; returns true while it is still unresolved
MG_HD bool split_jump_dense(const PeelView &P, uint32_t k) {
  const uint32_t v = ld_cg_u32(&P.sterm[k]);
  if ((v & kSplitTerm) || v == k) return false;  // terminal reached / known to be unresolvable
  uint32_t q = ld_cg_u32(&P.sterm[v]);
  if (q == v) q = k;  // the splitter ahead is unresolvable, so is this one
  P.sterm[k] = q;
  return true;
}
// after the rounds: the splitter's entry and the table take the resolved pointer as an item id (an unresolvable
// splitter points at a chain node: itself)
MG_HD void split_settle(const PeelView &P, uint32_t k) {
  const uint32_t v = P.sterm[k], s = P.split_ids[k];
  const uint32_t t = (v & kSplitTerm) ? (v & ~kSplitTerm) : P.split_ids[v];
  peel_ptr(P, s) = t;
  P.sterm[k] = t;
}
// b. on the entries themselves (the host emulation's form): one round for splitter s; true while unresolved
MG_HD bool split_jump_item(const PeelView &P, uint32_t s) {
  uint32_t p = ld_cg_u32(&peel_ptr(P, s));
  if (!bm_get(P.chain, p) || p == s) return false;  // terminal reached / known to be unresolvable
  uint32_t q = ld_cg_u32(&peel_ptr(P, p));
  if (q == p) q = s;  // the splitter ahead is unresolvable, so is this one
  peel_ptr(P, s) = q;
  return true;
}
// c. every chain node that is not a splitter; returns false when the hop cap was hit
MG_HD bool chain_resolve_item(const PeelView &P, uint32_t v) {
  if (is_splitter_id(P, v)) return true;
  const uint64_t ev = P.ent[v];
  const uint32_t hv = (uint32_t)(ev >> 32);
  if (!(hv & kEntChain)) return true;
  uint32_t o = hv & kNoOwner;
  if (o != kNoOwner) { peel_ptr(P, v) = P.sterm[o]; return true; }
  uint32_t p = (uint32_t)ev;
  // Brent's cycle detection: `mark` is the item the walk stood on 2^j hops after v.  A walk that comes back to
  // it runs round a cycle of chain nodes with neither splitter nor owner (the closed loop of a tandem repeat
  // whose entry and exit were filtered out has ~50 nodes, so about half of them hold no splitter):
  // v is unresolvable and ptr[v] stays its successor, a chain node — what terminal_of() tests.  Nobody
  // writes a pointer on such a component, so the walk is deterministic: <= 3 x (tail + cycle) hops.
  uint32_t mark = v, since = 0, power = 1;
  for (uint32_t hops = 0; hops < kWalkCapHead; ++hops) {
    if (is_splitter_id(P, p)) {
      // a chain splitter knows its terminal; an item that merely hashes like one is a terminal itself
      const uint64_t ep = ld_cg64(&P.ent[p]);
      peel_ptr(P, v) = ((uint32_t)(ep >> 32) & kEntChain) ? (uint32_t)ep : p;
      return true;
    }
    const uint64_t ep = ld_cg64(&P.ent[p]);  // one entry: chain flag, owner and successor of p
    const uint32_t hp = (uint32_t)(ep >> 32);
    if (!(hp & kEntChain)) { peel_ptr(P, v) = p; return true; }
    o = hp & kNoOwner;
    if (o != kNoOwner) { peel_ptr(P, v) = P.sterm[o]; return true; }
    if (p == mark) return true;
    if (++since == power) { mark = p; power <<= 1; since = 0; }
    p = (uint32_t)ep;  // the successor, or already p's terminal: both lie ahead on the same chain
  }
  return false;
}
// terminal of an out-neighbour (kNull when it sits on an unresolvable chain cycle).  A chain node's pointer names
// the terminal its chain ended in when it was resolved; if step 2b has re-classified that terminal as a chain node
// since, the walk goes on through it (at most one hop per round of step 2b) — this replaces a pass over all items
// per round that moved every such pointer on.
MG_HD uint32_t terminal_of(const PeelView &P, uint32_t c) {
  // (chain / recl bitmaps and the pointers of chain nodes are not written by the kernels that call this)
  if (!bm_get(P.chain, c)) return c;
  uint32_t t = ld_ro(&peel_ptr(P, c));
  if (!bm_get(P.chain, t)) return t;
  for (int hop = 0; hop < 8; ++hop) {
    if (!P.recl || !bm_get(P.recl, t)) return kNull;            // an original chain node: a cycle of chain nodes
    const uint32_t t2 = ld_ro(&peel_ptr(P, t));
    if (t2 == t) return kNull;                                  // known to be cyclic (reclass_jump_item)
    t = t2;
    if (!bm_get(P.chain, t)) return t;
  }
  return kNull;
}
// Step 2b: UNARY branch nodes.  A branch node whose out-neighbours all resolve to terminals that are
// already peeled (collected tips) — except for ONE distinct terminal T — is peeled iff T is: for the
// closure it behaves like a chain node pointing at T.  Typical: a SNP bubble between two strains (both
// arms end in the same node), or a path node with a sequencing-error tip hanging off it.  Re-classifying
// them as chain nodes (and resolving the new chains, which only involves the re-classified nodes)
// removes the long serial cascades the work list of step 3 would otherwise walk one branch node
// at a time.  Returns 0 = still a terminal, 1 = now a chain node (ptr[b] set), 2 = all out-neighbours
// are peeled already: b becomes a seed.
MG_HD int reclass_item(const GraphView &g, const PeelView &P, uint32_t b) {
  if (get_inv(g, b) || bm_get(P.chain, b) || bm_get(P.tpeeled, b)) return 0;
  uint64_t a = P.adj[b];
  uint32_t mask = (uint32_t)(a >> 32) & 0xFFu, next = (uint32_t)a;
  if (popc32(mask) < 2) return 0;
  uint32_t uniq = kNull;
  for (int i = 0; mask; ++i, mask >>= 1) {
    if (!(mask & 1)) continue;
    uint32_t tc = terminal_of(P, next - i);
    if (tc == kNull || tc == b) return 0;          // chain cycle / depends on itself: never peeled
    if (bm_get(P.tpeeled, tc)) continue;           // peeled already
    if (((uint32_t)(P.adj[tc] >> 32) & 0xFFu) == 0) return 0;  // a sink that is not a collected tip: never peeled
    if (uniq == kNull) uniq = tc;
    else if (uniq != tc) return 0;                 // two distinct open dependencies: a true AND node
  }
  if (uniq == kNull) return 2;
  peel_ptr(P, b) = uniq;
  return 1;
}
// resolution among the re-classified nodes of one round (`fresh` = their bitmap); true while unresolved
MG_HD bool reclass_jump_item(const PeelView &P, const uint32_t *fresh, uint32_t b) {
  uint32_t p = ld_cg_u32(&peel_ptr(P, b));
  if (!bm_get(fresh, p) || p == b) return false;
  uint32_t q = ld_cg_u32(&peel_ptr(P, p));
  if (q == p) q = b;  // the node ahead is known to be cyclic, so is this one
  peel_ptr(P, b) = q;
  return true;
}
// After the last round of step 2b: every chain node takes its final terminal (one pass instead of one per round;
// terminal_of needs no second hop afterwards, the caller clears P.recl).  A node on an unresolvable cycle keeps its
// pointer to a chain node.
MG_HD void relink_final_item(const PeelView &P, uint32_t v) {
  if (!bm_get(P.chain, v)) return;
  const uint32_t t = ld_cg_u32(&peel_ptr(P, v));
  if (!bm_get(P.chain, t)) return;  // resolved, terminal untouched by step 2b: the common case
  uint32_t u = t;
  for (int hop = 0; hop < 8; ++hop) {
    if (!bm_get(P.recl, u)) return;
    const uint32_t u2 = ld_cg_u32(&peel_ptr(P, u));
    if (u2 == u) return;
    u = u2;
    if (!bm_get(P.chain, u)) { peel_ptr(P, v) = u; return; }
  }
}
MG_HD void dep_insert(const PeelView &P, uint32_t term, uint32_t b) {
  uint32_t mask = (1u << P.dep_log) - 1, h = hash_u32_fwd(term, P.dep_log);
  uint64_t val = ((uint64_t)term << 32) | b;
  for (;;) {
#if defined(__CUDA_ARCH__)
    unsigned long long old = atomicCAS((unsigned long long *)&P.dep[h], ~0ull, (unsigned long long)val);
    if (old == ~0ull) return;
#else
    if (P.dep[h] == ~0ull) { P.dep[h] = val; return; }
#endif
    h = (h + 1) & mask;
  }
}
// Step 3a for one branch node: register it with the terminals of its out-neighbours, unless one of
// them can never be peeled (chain cycle, or a sink that is not a collected tip).
MG_HD void dep_register_item(const GraphView &g, const PeelView &P, uint32_t b) {
  uint64_t a = P.adj[b];
  uint32_t mask = (uint32_t)(a >> 32) & 0xFFu, next = (uint32_t)a;
  if (get_inv(g, b) || popc32(mask) < 2 || bm_get(P.chain, b) || bm_get(P.tpeeled, b)) return;
  uint32_t t[8];
  int n = 0;
  for (int i = 0; mask; ++i, mask >>= 1) {
    if (!(mask & 1)) continue;
    uint32_t tc = terminal_of(P, next - i);
    if (tc == kNull) return;
    uint32_t ma = (uint32_t)(P.adj[tc] >> 32) & 0xFFu;
    if (ma == 0 && !bm_get(P.tpeeled, tc)) return;  // sink that is not a seed
    t[n++] = tc;
  }
  for (int i = 0; i < n; ++i) dep_insert(P, t[i], b);
}
// Step 3b: terminal t has just been peeled; emit every dependent branch node that is now peeled.
// The sequentially consistent fence between setting a terminal's bit and testing its siblings
// guarantees that when the last two out-neighbours of a branch node are peeled concurrently at
// least one worker sees both bits; duplicates are arbitrated by the atomic bit set.
template <class Emit>
MG_HD void cpeel_one(const PeelView &P, uint32_t t, Emit emit) {
  uint32_t mask = (1u << P.dep_log) - 1, h = hash_u32_fwd(t, P.dep_log);
  for (;; h = (h + 1) & mask) {
    uint64_t e = ld_ro(&P.dep[h]);
    if (e == ~0ull) return;
    if ((uint32_t)(e >> 32) != t) continue;
    uint32_t b = (uint32_t)e;
    if (bm_get_coherent(P.tpeeled, b)) continue;
    uint64_t a = ld_ro(&P.adj[b]);
    uint32_t m = (uint32_t)(a >> 32) & 0xFFu, next = (uint32_t)a;
    bool all = true;
    for (int i = 0; m; ++i, m >>= 1)
      if (m & 1) {
        uint32_t tc = terminal_of(P, next - i);
        if (tc == kNull || !bm_get_coherent(P.tpeeled, tc)) { all = false; break; }
      }
    if (!all) continue;
    if (!bm_set_atomic(P.tpeeled, b)) continue;
    fence_sc();
    emit(b);
  }
}
// Step 4: is item v removed by the reduction?  (valid items only)
MG_HD bool is_peeled(const PeelView &P, uint32_t v) {
  uint32_t t = terminal_of(P, v);
  return t != kNull && bm_get(P.tpeeled, t);
}
// Step 4 for one item: final navigation-cache entry (children removed by the reduction drop out of
// the validity mask); `peeled` tells the caller to set the item's invalid bit.
MG_HD uint64_t peel_final_item(const GraphView &g, const PeelView &P, uint32_t v, bool &peeled) {
  peeled = false;
  if (get_inv(g, v)) return 0;
  if (is_peeled(P, v)) { peeled = true; return 0; }
  uint64_t a = P.adj[v];
  uint32_t mask = (uint32_t)(a >> 32) & 0xFFu, next = (uint32_t)a, keep = 0;
  // a node with ONE out-neighbour was classified a chain node: it and that neighbour end in the same terminal (or on
  // the same unresolvable cycle), so the neighbour stays exactly when the node does — no second look-up for four items
  // in five
  if ((mask & (mask - 1u)) == 0) return a;
  for (int i = 0; i < 8; ++i)
    if (((mask >> i) & 1) && !is_peeled(P, next - i)) keep |= 1u << i;
  return (uint64_t)next | ((uint64_t)keep << 32);
}

// ------------------------------------------------------------------ CollectTips from the dead ends' side
// CollectTips (src/cycle_finder.cpp:346-357) asks every item e for EdgeOutdegree(e) == 0: Forward(e) (rank + select)
// and a scan of the target node's items for a valid one (sdbg.h:296-320).  The same set, turned round: an item b that
// closes a node (`last` bit) none of whose items is valid is a dead end, and the valid items e with Forward(e) == b
// are exactly its valid incoming edges (IncomingEdges, sdbg.h:245-287, evaluated whatever b's own invalid bit says).
// Dead ends are rare (0.07 % of the items at C2, 2.5 % at C4), everything else is decided by the item's own block
// line.  Returns the number of ids written to in[] (0 when b is not a dead end).
MG_HD bool is_dead_end(const GraphView &g, uint32_t b) {
  if (!get_last(g, b)) return false;
  uint32_t x = b;
  do {  // the node's items, as ComputeOutgoings scans them
    if (!get_inv(g, x)) return false;
    --x;
  } while (x != kNull && !get_lt(g, x));
  return true;
}
MG_HD int dead_end_incoming(const GraphView &g, uint32_t b, uint32_t *in) {
  if (!is_dead_end(g, b)) return 0;
  int deg = incoming<false, false>(g, b, in);
  return deg > 4 ? 4 : deg;
}

// ------------------------------------------------------------------ start-node predicate
// src/cycle_finder.cpp:398-408: valid, indegree >= 2, multiplicity > threshold, no incoming edge
// equal to the node itself (_IncomingNotEqualToCurrentNode, :29-36).
MG_HD bool is_candidate(const GraphView &g, uint32_t v, uint64_t thr) {
  if (get_inv(g, v)) return false;
  if ((uint64_t)get_mult(g, v) <= thr) return false;
  uint32_t in[4];
  int deg = incoming(g, v, in);
  if (deg < 2) return false;
  if (deg > 4) deg = 4;
  for (int i = 0; i < deg; ++i)
    if (in[i] == v) return false;
  return true;
}
// bucket key (int)ceil(log2((double)mult)), src/cycle_finder.cpp:414, exact for mult >= 1
MG_HD int log2_bucket(uint32_t m) {
  if (m <= 1) return 0;
  uint32_t x = m - 1;
  int b = 0;
  while (x) { ++b; x >>= 1; }
  return b;
}

// ------------------------------------------------------------------ DepthLevelSearch
struct U32HashSet {  // open addressing, linear probing, empty = kNull
  uint32_t *tab;
  uint32_t log2cap;
  uint32_t count;
  uint32_t max_count;
};
MG_HD uint32_t hash_u32(uint32_t k, uint32_t log2cap) { return (k * 2654435761u) >> (32 - log2cap); }
// 1 = inserted, 0 = already present, -1 = full
MG_HD int set_insert(U32HashSet &s, uint32_t key) {
  uint32_t mask = (1u << s.log2cap) - 1;
  uint32_t h = hash_u32(key, s.log2cap);
  for (;;) {
    uint32_t cur = s.tab[h];
    if (cur == key) return 0;
    if (cur == kNull) {
      if (s.count >= s.max_count) return -1;
      s.tab[h] = key;
      ++s.count;
      return 1;
    }
    h = (h + 1) & mask;
  }
}
// Two-phase insert: set_peek() loads the first probe slot of a key, set_insert(s, key, peeked) finishes.
// A node's <= 4 neighbours are peeked together, so their probe loads overlap instead of forming a chain;
// a slot that was free when peeked is read again before it is taken (a neighbour peeked in the same
// batch may have taken it), occupied slots never change during a search.
MG_HD uint32_t set_peek(const U32HashSet &s, uint32_t key) { return s.tab[hash_u32(key, s.log2cap)]; }
MG_HD int set_insert(U32HashSet &s, uint32_t key, uint32_t peeked) {
  if (peeked == key) return 0;
  uint32_t mask = (1u << s.log2cap) - 1;
  uint32_t h = hash_u32(key, s.log2cap);
  if (peeked != kNull) h = (h + 1) & mask;
  for (;;) {
    uint32_t cur = s.tab[h];
    if (cur == key) return 0;
    if (cur == kNull) {
      if (s.count >= s.max_count) return -1;
      s.tab[h] = key;
      ++s.count;
      return 1;
    }
    h = (h + 1) & mask;
  }
}
struct BitSet {  // one bit per graph item; exact fallback for very large searches
  uint32_t *bits;
};
MG_HD int set_insert(BitSet &s, uint32_t key) {
  uint32_t w = s.bits[key >> 5], b = 1u << (key & 31);
  if (w & b) return 0;
  s.bits[key >> 5] = w | b;
  return 1;
}

MG_HD uint32_t set_peek(const BitSet &s, uint32_t key) { return s.bits[key >> 5]; }
MG_HD int set_insert(BitSet &s, uint32_t key, uint32_t peeked) {
  if ((peeked >> (key & 31)) & 1) return 0;
  return set_insert(s, key);
}

constexpr int kDlsFound = 1, kDlsNotFound = 0, kDlsSetFull = -1, kDlsStackFull = -2;

// src/cycle_finder.cpp:248-343 with target == start (the only call site, :411).  `adj` is the
// navigation cache.  The explicit stack holds at most 3*limit+5 entries.
// A stack entry carries the node's navigation entry: it is loaded together with the probes of the
// parent's expansion (all <= 4 neighbours at once), so an expansion waits for ONE round of memory
// latency instead of two dependent ones (pop -> navigation entry -> probes).
template <class Set>
MG_HD int dls_run(const uint64_t *adj, uint32_t start, int limit, Set &seen, uint32_t *stk_node,
                  uint16_t *stk_depth, uint64_t *stk_adj, uint32_t stk_cap, uint32_t &expansions) {
  uint32_t sp = 0;
  stk_node[0] = start;
  stk_depth[0] = 0;
  stk_adj[0] = ld_ro(&adj[start]);
  sp = 1;
  while (sp) {
    --sp;
    uint32_t v = stk_node[sp];
    int depth = stk_depth[sp];
    uint32_t nb[4];
    int outdeg = adj_outgoing(stk_adj[sp], nb);     // :285-302 (adj == 0 for an invalid v)
    if (outdeg == 0) continue;                      // :294
    ++expansions;
    if (depth >= limit) continue;                   // :312
    if (outdeg > 4) outdeg = 4;
    uint32_t peeked[4] = {0, 0, 0, 0};
    uint64_t nadj[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (i < outdeg) {
        peeked[i] = set_peek(seen, nb[i]);
        nadj[i] = ld_ro(&adj[nb[i]]);
      }
#pragma unroll
    for (int i = 0; i < 4; ++i) {                   // :320-334, descending ids
      if (i < outdeg) {
        uint32_t x = nb[i];
        int r = set_insert(seen, x, peeked[i]);
        if (r < 0) return kDlsSetFull;
        if (r == 1 || (x == start && depth > 0)) {
          if (sp >= stk_cap) return kDlsStackFull;
          stk_node[sp] = x;
          stk_depth[sp] = (uint16_t)(depth + 1);
          stk_adj[sp] = nadj[i];
          ++sp;
        }
      }
    }
    if (v == start && depth > 1) return kDlsFound;  // :337-339
  }
  return kDlsNotFound;
}

// ------------------------------------------------------------------ FindCycle
// libstdc++ std::unordered_set<uint64_t> iteration order for the <= 4-element neighbour sets of
// FindCycle: 13 buckets, identity hash; a new element goes to the list head when its bucket is
// empty, else to the front of its bucket's run; erase keeps the relative order.
struct USet {
  uint32_t e[7];
  uint32_t n;
};
MG_HD void uset_insert(USet &s, uint32_t x) {
  int first_same = -1;
  for (uint32_t i = 0; i < s.n; ++i) {
    if (s.e[i] == x) return;
    if (first_same < 0 && s.e[i] % 13 == x % 13) first_same = (int)i;
  }
  if (s.n >= 7) return;
  int at = first_same < 0 ? 0 : first_same;
  for (int i = (int)s.n; i > at; --i) s.e[i] = s.e[i - 1];
  s.e[at] = x;
  s.n++;
}
MG_HD void uset_erase(USet &s, uint32_t x) {
  for (uint32_t i = 0; i < s.n; ++i)
    if (s.e[i] == x) {
      for (uint32_t j = i; j + 1 < s.n; ++j) s.e[j] = s.e[j + 1];
      s.n--;
      return;
    }
}

constexpr int kMaxSteps = 10000000;  // src/cycle_finder.cpp:149
constexpr int kClusterBound = 500;   // src/cycle_finder.cpp:132
constexpr int kPollSteps = 256;      // how often a running search polls for its own start being marked
constexpr int kSliceSteps = 1024;    // extracted-ball search: steps between two polls of the warp
enum EnumStatus : int32_t {
  kEnumDone = 0,
  kEnumBound = 1,     // cluster bound reached: result is empty (:162-165), search stopped early
  kEnumCap = 2,       // step cap reached (:148-151): cycles found so far are kept
  kEnumOverflow = 3,  // arena too small: must be re-run with a larger arena
  kEnumSkipped = 4,   // start already visited (:476) — seen before or while searching
};

// The per-search lock map (std::map<uint64_t,int> lock of :140) plus what the search derives again
// and again for the same node.  Entries are allocated DENSELY in first-touch order, so the working
// set of a search is contiguous (cache resident) whatever the capacity; a hash index of entry
// numbers finds them.  An index slot h is live iff index[h] < n_keys and hot[index[h]].home == h,
// so neither the index nor the entries are ever cleared between searches, and the entry array
// doubles as the key log (the "examined" set the ordered commit re-checks).
//   HotEnt  (32 B, one sector): key, lock value, on-path flag, memoised filtered out-neighbour set
//           (_GetOutgoings(node,set,mult), :58-72, in std::unordered_set iteration order)
//   ColdEnt (32 B): memoised filtered in-neighbour set (_GetIncomings, :76-88), used by relaxations
struct HotEnt {
  uint32_t key;
  uint32_t home;     // index slot that points here
  int32_t val;       // lock value
  uint32_t nout;     // low 8 bits: size of out[] or kMemoNone8; bit 8: node is on the current path
  uint32_t out[4];
};
struct ColdEnt {
  uint32_t nin;      // kMemoNone = not computed yet
  uint32_t in[4];
  uint32_t pad[3];
};
static_assert(sizeof(HotEnt) == 32 && sizeof(ColdEnt) == 32, "lock entry layout");
constexpr uint32_t kMemoNone = 0xFFFFFFFFu;
constexpr uint32_t kMemoNone8 = 0xFFu;
constexpr uint32_t kOnPath = 0x100u;

struct EnumArena {
  uint32_t *path;       // [M+2]
  uint32_t *pent;       // [M+2] entry number of path[i]
  int32_t *bl;          // [M+2]
  USet *stk;            // [M+2]
  uint32_t *index;      // [1 << index_log] entry numbers
  uint32_t index_log;
  HotEnt *hot;          // [ent_cap]
  ColdEnt *cold;        // [ent_cap]
  uint32_t ent_cap;
  uint64_t *relax;      // relax_stack entries (bl << 32 | node), :192-193
  uint32_t relax_cap;
  uint32_t *cyc_nodes;  // staged raw cycles
  uint32_t cyc_nodes_cap;
  uint32_t *cyc_off;    // [kClusterBound + 1]
};
struct EnumResult {
  int32_t status;
  uint32_t n_cycles;
  uint32_t n_nodes;
  uint32_t n_keys;      // entries in A.hot[] = nodes the search accepted (+ its start)
  uint32_t steps;
  uint32_t expansions;
};

MG_HD bool visited_bit(const uint32_t *visited, uint32_t i) { return (visited[i >> 5] >> (i & 31)) & 1; }
// `visited` is marked by other SMs while searches run (ordered commit): read it through L2
MG_HD bool visited_live(const uint32_t *visited, uint32_t i) { return (ld_cg32(&visited[i >> 5]) >> (i & 31)) & 1; }

// entry number of `key` in the lock map, inserting (val = defval) when absent (std::map::try_emplace
// / operator[] as used at :168,:200); kNull when the arena is full.
MG_HD uint32_t lock_get(EnumArena &A, uint32_t &n_keys, uint32_t key, int32_t defval) {
  uint32_t mask = (1u << A.index_log) - 1;
  uint32_t h = hash_u32(key, A.index_log);
  for (;;) {
    uint32_t v = A.index[h];
    if (v < n_keys && A.hot[v].home == h) {
      if (A.hot[v].key == key) return v;
      h = (h + 1) & mask;
      continue;
    }
    if (n_keys >= A.ent_cap) return kNull;
    v = n_keys++;
    A.index[h] = v;
    HotEnt &e = A.hot[v];
    e.key = key;
    e.home = h;
    e.val = defval;
    e.nout = kMemoNone8;
    A.cold[v].nin = kMemoNone;
    return v;
  }
}

// _BackgroundCheck (src/cycle_finder.cpp:40-52) for up to 4 neighbours at once: all multiplicity and
// visited words are loaded before any is tested, so the loads overlap instead of forming a chain.
// (two halves, so that a caller with several nodes in hand can issue the loads of all of them before any test)
MG_HD void filtered_loads(const GraphView &g, const uint32_t *visited, const uint32_t *nb, int deg, uint32_t *m, uint32_t *vw) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = vw[i] = 0;
    if (i < deg) {
      m[i] = ld_ro(&g.mult[nb[i]]);
      vw[i] = ld_cg32(&visited[nb[i] >> 5]);
    }
  }
}
MG_HD void filtered_finish(uint32_t node, uint32_t mu, const uint32_t *nb, int deg, const uint32_t *m, const uint32_t *vw,
                           uint32_t &n_out, uint32_t *out) {
  USet s;
  s.n = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (i < deg) {
      bool ok = !((vw[i] >> (nb[i] & 31)) & 1) && m[i] != 0 && mu / m[i] <= 500 && nb[i] != node;
      if (ok) uset_insert(s, nb[i]);
    }
  n_out = s.n;
  for (uint32_t i = 0; i < s.n && i < 4; ++i) out[i] = s.e[i];
}
MG_HD void filtered_set(const GraphView &g, const uint32_t *visited, uint32_t node, uint32_t mu, const uint32_t *nb,
                        int deg, uint32_t &n_out, uint32_t *out) {
  uint32_t m[4], vw[4];
  filtered_loads(g, visited, nb, deg, m, vw);
  filtered_finish(node, mu, nb, deg, m, vw, n_out, out);
}
// _GetOutgoings(node,set,mult), src/cycle_finder.cpp:58-72, memoised in the node's entry
MG_HD void enum_outgoings(const GraphView &g, const uint64_t *adj, const uint32_t *visited, HotEnt &e, uint32_t mu,
                          USet &s, uint32_t &expansions) {
  ++expansions;
  if ((e.nout & 0xFFu) == kMemoNone8) {
    uint32_t nb[4], n = 0;
    int deg = adj_outgoing(ld_ro(&adj[e.key]), nb);
    if (deg > 4) deg = 4;
    filtered_set(g, visited, e.key, mu, nb, deg, n, e.out);
    e.nout = (e.nout & ~0xFFu) | n;
  }
  s.n = e.nout & 0xFFu;
  for (uint32_t i = 0; i < 4; ++i) s.e[i] = e.out[i];
}
// _GetIncomings(node,set,mult), src/cycle_finder.cpp:76-88, memoised likewise
MG_HD void enum_incomings(const GraphView &g, const uint32_t *visited, uint32_t key, ColdEnt &c, uint32_t mu,
                          uint32_t &expansions) {
  ++expansions;
  if (c.nin == kMemoNone) {
    uint32_t nb[4];
    int deg = incoming(g, key, nb);
    if (deg > 4) deg = 4;
    if (deg < 0) deg = 0;
    filtered_set(g, visited, key, mu, nb, deg, c.nin, c.in);
  }
}

// FindCycleUtil + FindCycle, src/cycle_finder.cpp:231-243,140-225.  Does NOT mark visited (the
// ordered commit does, :217-222).  `visited` may gain marks while the search runs; every node the
// search ACCEPTED has an entry in A.hot[], and the commit re-checks those, so a search is only kept
// when it is identical to one run against the bitmap of its position in the sequential order.
// `poll()` is called every kPollSteps steps; returning true abandons the search (its start has been
// marked by an earlier start's commit: the reference would skip it, :476).
// Exact shortcut: once the cluster bound fires every later hit clears the list again, so the result
// is empty whatever follows; the search stops there instead of spinning to the step cap.
template <class Poll>
MG_HD_NOINLINE void enum_run(const GraphView &g, const uint64_t *adj, const uint32_t *visited, uint32_t start,
                             int min_len, int max_len, EnumArena &A, EnumResult &R, Poll poll) {
  const int M = max_len;
  const uint32_t mu = get_mult(g, start);
  uint32_t n_keys = 0, expansions = 0, n_cycles = 0, n_nodes = 0;
  int plen = 0, slen = 0, blen = 0;
  int counter = 0, steps = 0;
  int32_t status = kEnumDone;
  A.cyc_off[0] = 0;
  {
    uint32_t v0 = lock_get(A, n_keys, start, 0);  // lock[start] = 0, :237
    if (v0 == kNull) { status = kEnumOverflow; goto done; }
    A.hot[v0].val = 0;
    A.hot[v0].nout |= kOnPath;
    A.path[0] = start;
    A.pent[0] = v0;
    plen = 1;
    enum_outgoings(g, adj, visited, A.hot[v0], mu, A.stk[slen++], expansions);
    A.bl[blen++] = M;
  }
  while (slen > 0) {
    if (++steps > kMaxSteps) { status = kEnumCap; break; }  // :148-151
    if ((steps & (kPollSteps - 1)) == 0 && poll()) { status = kEnumSkipped; n_cycles = n_nodes = 0; goto done; }
    USet nbs = A.stk[slen - 1];                             // :153 (copy)
    bool flag = true;
    for (uint32_t i = 0; i < nbs.n; ++i) {
      uint32_t nb = nbs.e[i];
      if (nb == start) {                                    // :157-166
        A.bl[blen - 1] = 1;
        if (plen > min_len) {
          if (++counter >= kClusterBound) {                 // cluster bound: final answer is empty
            status = kEnumBound;
            n_cycles = 0;
            n_nodes = 0;
            goto leftovers;
          }
          if (n_nodes + (uint32_t)plen > A.cyc_nodes_cap) { status = kEnumOverflow; goto done; }
          for (int j = 0; j < plen; ++j) A.cyc_nodes[n_nodes + j] = A.path[j];
          n_nodes += (uint32_t)plen;
          A.cyc_off[++n_cycles] = n_nodes;
        }
      } else {
        uint32_t v = lock_get(A, n_keys, nb, M);            // try_emplace(nb, M), :168
        if (v == kNull) { status = kEnumOverflow; goto done; }
        HotEnt &e = A.hot[v];
        if (plen < e.val) {
          A.path[plen] = nb;                                // :170
          A.pent[plen] = v;
          ++plen;
          A.bl[blen++] = M;                                 // :171
          e.val = plen;                                     // :172
          e.nout |= kOnPath;
          uset_erase(A.stk[slen - 1], nb);                  // :173
          enum_outgoings(g, adj, visited, e, mu, A.stk[slen], expansions);  // :174-176
          ++slen;
          flag = false;
          break;
        }
      }
    }
    if (flag) {                                             // :181-211
      --slen;
      --plen;
      uint32_t v = A.path[plen];
      A.hot[A.pent[plen]].nout &= ~kOnPath;
      int b = A.bl[--blen];
      if (blen > 0 && b < A.bl[blen - 1]) A.bl[blen - 1] = b;
      if (b < M) {
        uint32_t rn = 0;
        A.relax[rn++] = ((uint64_t)(uint32_t)b << 32) | v;  // :192-193
        while (rn) {
          uint64_t r = A.relax[--rn];
          int rb = (int)(r >> 32);
          uint32_t u = (uint32_t)r;
          uint32_t vu = lock_get(A, n_keys, u, M);
          if (vu == kNull) { status = kEnumOverflow; goto done; }
          if (A.hot[vu].val < M - rb + 1) {
            A.hot[vu].val = M - rb + 1;
            enum_incomings(g, visited, u, A.cold[vu], mu, expansions);
            uint32_t nin = A.cold[vu].nin;
            uint32_t in[4] = {A.cold[vu].in[0], A.cold[vu].in[1], A.cold[vu].in[2], A.cold[vu].in[3]};
            for (uint32_t i = 0; i < nin; ++i) {
              uint32_t vw = lock_get(A, n_keys, in[i], M);  // every accepted neighbour gets an entry
              if (vw == kNull) { status = kEnumOverflow; goto done; }
              if (!(A.hot[vw].nout & kOnPath)) {            // :197-199
                if (rn >= A.relax_cap) { status = kEnumOverflow; goto done; }
                A.relax[rn++] = ((uint64_t)(uint32_t)(rb + 1) << 32) | in[i];
              }
            }
          }
        }
      }
    }
  }
leftovers:
  // On an early stop the neighbour sets still on the stack hold accepted nodes that never got an
  // entry; the ordered commit must see them as examined.
  if (status == kEnumBound || status == kEnumCap) {
    for (int l = 0; l < slen; ++l)
      for (uint32_t i = 0; i < A.stk[l].n; ++i) {
        uint32_t x = A.stk[l].e[i];
        if (x == start) continue;
        if (lock_get(A, n_keys, x, M) == kNull) { status = kEnumOverflow; goto done; }
      }
  }
done:
  R.status = status;
  R.n_cycles = n_cycles;
  R.n_nodes = n_nodes;
  R.n_keys = n_keys;
  R.steps = (uint32_t)steps;
  R.expansions = expansions;
}

// ------------------------------------------------------------------ region extraction
// keep_crispr_regions_extended_by_k, src/spacer_ordering.cpp:78-139: the set of cycle node ids grows for k
// rounds by the in- and out-neighbours (IncomingEdges/OutgoingEdges, i.e. valid items) of its VALID members;
// every valid item outside the final set is invalidated.  Members found in round r contribute neighbours from
// round r+1 on and re-scanning older members adds nothing, so round r only has to expand the members that
// were new in round r-1: a level-synchronous breadth-first search over bitmaps.
// Two forms of a level: a queue of the new members (regions.cu k_region_bfs, work proportional to the
// frontier) and, when a queue overflows, bitmaps (region_expand_word: member = the set; cur = members new in
// the previous round, word `wi` is expanded; next = members new in this round; returns the number added).
// Expands one member: its valid in-/out-neighbours that were not in the set yet are added to `member` and
// returned in fresh[0..count) (count <= 8).
MG_HD int region_expand_node(const GraphView &g, uint32_t *member, uint32_t v, uint32_t fresh[8]) {
  if (get_inv(g, v)) return 0;  // :94-96 (only a seed id can be invalid)
  uint32_t nb[4];
  int c = 0;
  int d = incoming(g, v, nb);   // :98-109
  for (int i = 0; i < d && i < 4; ++i)
    if (bm_set_atomic(member, nb[i])) fresh[c++] = nb[i];
  d = outgoing(g, v, nb);       // :111-122
  for (int i = 0; i < d && i < 4; ++i)
    if (bm_set_atomic(member, nb[i])) fresh[c++] = nb[i];
  return c;
}
MG_HD unsigned region_expand_word(const GraphView &g, uint32_t *member, const uint32_t *cur, uint32_t *next,
                                  uint32_t wi) {
  uint32_t word = cur[wi];
  unsigned added = 0;
  while (word) {
#if defined(__CUDA_ARCH__)
    int b = __ffs(word) - 1;
#else
    int b = __builtin_ctz(word);
#endif
    word &= word - 1;
    uint32_t fresh[8];
    int c = region_expand_node(g, member, wi * 32u + (uint32_t)b, fresh);
    for (int i = 0; i < c; ++i) bm_set_atomic(next, fresh[i]);
    added += (unsigned)c;
  }
  return added;
}
// :131-138 for the 64 items of invalid-bitmap word `wi` (word j of FBlock b holds items 256 b + 64 j ...).
// Returns the number of items invalidated; *kept += items still valid.
MG_HD unsigned region_keep_word(const GraphView &g, const uint32_t *member, uint64_t wi, unsigned *kept) {
  uint64_t *p = &g.fb[wi >> 2].inv[wi & 3];
  uint64_t inv = *p;
  uint64_t first = wi * 64, in_range = ~0ull;
  if (first + 64 > g.n) in_range = g.n > first ? (~0ull >> (64 - (g.n - first))) : 0;
  uint64_t mem = 0;
  if (first < g.n) {
    mem = member[wi * 2];
    if (first + 32 < g.n) mem |= (uint64_t)member[wi * 2 + 1] << 32;
  }
  uint64_t drop = ~inv & ~mem & in_range;
  if (drop) *p = inv | drop;
  *kept += (unsigned)popc64(~inv & mem & in_range);
  return (unsigned)popc64(drop);
}

}  // namespace mg
