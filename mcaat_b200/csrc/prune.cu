// K1-K3 + navigation cache: graph pruning of reference src/cycle_finder.cpp:436-452
// (CollectTips :346-357, InvalidateMultiplicityOneNodes :372-382, RecursiveReduction :359-371),
// and the diagnostic navigation table (reference sdbg.h:328-357).
#include "common.cuh"
#include "search_dev.cuh"

namespace mg {

static inline unsigned grid_for(uint64_t n, int block) { return (unsigned)((n + block - 1) / block); }
// The streaming kernels below count what they do.  One global atomic per warp on the same address
// serialises in the L2 (7 M warps at c2 = milliseconds); they run grid-stride over a fixed grid
// instead, keep per-warp counts in shared memory and issue one global atomic per block and counter.
static inline unsigned stream_grid(const mg_graph *g, uint64_t n, int block) {
  uint64_t want = (n + block - 1) / block, cap = (uint64_t)g->sm_count * 16;
  return (unsigned)(want < cap ? (want ? want : 1) : cap);
}
template <int K>
struct BlockCounters {
  unsigned *sh;
  __device__ BlockCounters(unsigned *shared) : sh(shared) {
    if (threadIdx.x < K) sh[threadIdx.x] = 0;
    __syncthreads();
  }
  __device__ void add(int k, unsigned v) { if (v) atomicAdd(&sh[k], v); }
  __device__ void flush(unsigned long long *const *dst) {
    __syncthreads();
    if (threadIdx.x < K && sh[threadIdx.x]) atomicAdd(dst[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
  }
};

// CollectTips: bit i of `bits` = valid(i) && outdegree(i) == 0.  One thread per item; a warp
// covers 32 consecutive items = one bitmap word.
__global__ void k_tips(GraphView g, uint32_t *bits, unsigned long long *count) {
  __shared__ unsigned sh[1];
  BlockCounters<1> bc(sh);
  const uint64_t n32 = (g.n + 31) & ~31ull;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += (uint64_t)gridDim.x * blockDim.x) {
    bool tip = false;
    if (i < g.n) tip = outgoing(g, (uint32_t)i, nullptr) == 0;
    unsigned m = __ballot_sync(0xffffffffu, tip);
    if ((threadIdx.x & 31) == 0) {
      bits[i >> 5] = m;
      bc.add(0, (unsigned)__popc(m));
    }
  }
  unsigned long long *dst[1] = {count};
  bc.flush(dst);
}
// The same bitmap from the dead ends' side (search_dev.cuh): a streaming pass marks the dead ends (one block line per
// item), their ids are compacted, and one thread per dead end evaluates IncomingEdges — dense warps, whatever the
// share of dead ends (0.07 % of the items at C2, 2.5 % at C4).  `bits` of k_tips_of_dead_ends is zeroed by the caller.
// 64 items per thread, on the words themselves.  A node is the run of items that ends at an item with the last-or-tip
// bit (ComputeOutgoings scans downwards until it meets one, sdbg.h:296-320); whether a run holds a valid item is a
// carry: adding (valid & ~end) + ~end lets a valid bit ripple upwards through the non-end positions of its run and
// deposits it at the run's end position, where both addends are zero, so it goes no further.  A run that began in the
// previous word brings its carry along (runs are a few items long; a word without any end bit takes the literal path).
__global__ void k_dead_ends(GraphView g, uint32_t *dead) {
  const uint64_t n_words = (g.n + 63) / 64, stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += stride) {
    const FBlock *fb = &g.fb[w >> 2];
    const int j = (int)(w & 3);
    const uint64_t E = ld_ro(&fb->lt[j]), La = ld_ro(&fb->last[j]), V = ~fb->inv[j];
    uint64_t cin = 0, d = 0;
    bool literal = false;
    if (w > 0) {
      const FBlock *pb = &g.fb[(w - 1) >> 2];
      const int pj = (int)((w - 1) & 3);
      const uint64_t Ep = ld_ro(&pb->lt[pj]), Vp = ~pb->inv[pj];
      if (Ep == 0) literal = true;
      else cin = ((Vp >> (63 - __clzll((long long)Ep))) >> 1) != 0;  // valid items above the last end bit of that word
    }
    if (!literal) {
      const uint64_t nE = ~E, Z = (V & nE) + nE + cin;
      d = La & ~((Z | V) & E);
    } else {
      for (int b = 0; b < 64; ++b) {
        const uint64_t i = w * 64 + b;
        if (i < g.n && is_dead_end(g, (uint32_t)i)) d |= 1ull << b;
      }
    }
    if (w == n_words - 1 && (g.n & 63)) d &= (1ull << (g.n & 63)) - 1;
    dead[2 * w] = (uint32_t)d;
    dead[2 * w + 1] = (uint32_t)(d >> 32);  // (the bitmaps carry four spare words)
  }
}
__global__ void k_tips_of_dead_ends(GraphView g, const uint32_t *ids, uint32_t n_ids, uint32_t *bits, unsigned long long *count) {
  __shared__ unsigned sh[1];
  BlockCounters<1> bc(sh);
  const uint32_t n32 = (n_ids + 31) & ~31u;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += (uint64_t)gridDim.x * blockDim.x) {
    unsigned fresh = 0;
    if (i < n_ids) {
      uint32_t in[4];
      int deg = incoming<false, false>(g, ids[i], in);
      if (deg > 4) deg = 4;
      for (int j = 0; j < deg; ++j) fresh += bm_set_atomic(bits, in[j]) ? 1u : 0u;
    }
    for (int d = 16; d; d >>= 1) fresh += __shfl_down_sync(0xffffffffu, fresh, d);
    if ((threadIdx.x & 31) == 0) bc.add(0, fresh);
  }
  unsigned long long *dst[1] = {count};
  bc.flush(dst);
}
// InvalidateMultiplicityOneNodes: every item with multiplicity <= 1 becomes invalid; the count
// includes items that were invalid already, like the reference's counter.
// (eight items = one 16-byte load per thread and round: with two bytes per thread in flight the pass ran at a tenth of the
//  HBM rate; four lanes put their bits together into the 32-bit word of the block line that they own)
__global__ void k_mult1(GraphView g, unsigned long long *count) {
  __shared__ unsigned sh[1];
  BlockCounters<1> bc(sh);
  const uint64_t n_groups = ((g.n + 31) / 32) * 4, stride = (uint64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  for (uint64_t q0 = (uint64_t)blockIdx.x * blockDim.x; q0 < n_groups; q0 += stride) {  // (whole warps: shuffles below)
    const uint64_t q = q0 + threadIdx.x, base = q * 8;
    unsigned bits = 0;
    if (base + 8 <= g.n) {
      const uint4 v = __ldg(reinterpret_cast<const uint4 *>(g.mult + base));
      const unsigned m[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) bits |= ((m[k] & 0xFFFFu) <= 1u ? 1u : 0u) << (2 * k) | ((m[k] >> 16) <= 1u ? 1u : 0u) << (2 * k + 1);
    } else {
      for (int k = 0; k < 8; ++k)
        if (base + k < g.n && g.mult[base + k] <= 1) bits |= 1u << k;
    }
    unsigned w = bits << (8 * (lane & 3));
    w |= __shfl_xor_sync(0xffffffffu, w, 1);
    w |= __shfl_xor_sync(0xffffffffu, w, 2);
    if ((lane & 3) == 0 && w) {
      const uint64_t i = base;  // first item of the 32-bit word
      uint32_t *w32 = (uint32_t *)&g.fb[i >> 8].inv[(i >> 6) & 3] + ((i >> 5) & 1);
      *w32 |= w;  // these four lanes own the word and nobody else writes invalid bits in this phase
      bc.add(0, (unsigned)__popc(w));
    }
  }
  unsigned long long *dst[1] = {count};
  bc.flush(dst);
}
// ---- RecursiveReduction by chain contraction (search_dev.cuh, "peel") -------------------------
// step 1: navigation cache on the post-filter graph, chain / seed bitmaps, chain pointers
__global__ void k_chain_init(GraphView g, PeelView P, uint32_t *branch, uint32_t *split, unsigned long long *n_pairs) {
  __shared__ unsigned sh[1];
  BlockCounters<1> bc(sh);
  const uint64_t n32 = (g.n + 31) & ~31ull;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += (uint64_t)gridDim.x * blockDim.x) {
    bool is_chain = false, is_seed = false;
    int deg = -1;
    if (i < g.n) deg = chain_init_item(g, P, (uint32_t)i, is_chain, is_seed);
    unsigned mc = __ballot_sync(0xffffffffu, is_chain), ms = __ballot_sync(0xffffffffu, is_seed);
    unsigned mb = __ballot_sync(0xffffffffu, deg >= 2);  // branch nodes: all that steps 2b and 3a ever look at
    unsigned msp = __ballot_sync(0xffffffffu, is_chain && is_splitter_id(P, (uint32_t)i));  // the splitters of step 2
    unsigned pairs = deg >= 2 ? (unsigned)deg : 0u;
    for (int d = 16; d; d >>= 1) pairs += __shfl_down_sync(0xffffffffu, pairs, d);
    if ((threadIdx.x & 31) == 0) {
      P.chain[i >> 5] = mc;
      P.tpeeled[i >> 5] = ms;
      branch[i >> 5] = mb;
      split[i >> 5] = msp;
      bc.add(0, pairs);
    }
  }
  unsigned long long *dst[1] = {n_pairs};
  bc.flush(dst);
}
// step 2 (search_dev.cuh): splitter walks, pointer jumping over the splitters, resolution of the rest
// (the per-item passes are grid-stride loops; the light ones — split_mark, relink — run on a fixed grid, because with one
//  thread per item the 4 M blocks of a 10^9-item graph cost more in block scheduling than the few bytes each of them
//  moves; the ones that walk or insert — chain_resolve, dep_register — measured faster with one thread per item)
__global__ void k_split_walk(PeelView P, uint32_t n_ids, int *capped) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_ids && !split_walk_item(P, i)) *capped = 1;
}
// the resolved pointers into the splitters' entries; the dense table keeps them as item ids for the resolution pass
__global__ void k_split_settle(PeelView P, uint32_t n_ids) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_ids) split_settle(P, i);
}
__global__ void k_split_jump(PeelView P, uint32_t n_ids, int *unresolved) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool u = i < n_ids && split_jump_dense(P, i);
  if (__any_sync(0xffffffffu, u) && (threadIdx.x & 31) == 0) *unresolved = 1;
}
// (four items per thread and round: these passes are bound by the latency of two dependent loads per item — the entry,
//  then what it names — so the loads of several items have to be in flight per thread)
__global__ void k_chain_resolve(uint64_t n, PeelView P, int *capped) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 4 * stride) {
    uint64_t ev[4];
    uint32_t tv[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t v = i0 + q * stride;
      ev[q] = v < n && !is_splitter_id(P, (uint32_t)v) ? P.ent[v] : 0ull;  // 0: not a chain node
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {  // the common case: a walk passed the item, its splitter knows the terminal
      const uint32_t hv = (uint32_t)(ev[q] >> 32);
      tv[q] = (hv & kEntChain) && (hv & kNoOwner) != kNoOwner ? P.sterm[hv & kNoOwner] : 0u;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t v = i0 + q * stride;
      const uint32_t hv = (uint32_t)(ev[q] >> 32);
      if (!(hv & kEntChain)) continue;
      if ((hv & kNoOwner) != kNoOwner) peel_ptr(P, (uint32_t)v) = tv[q];
      else if (!chain_resolve_item(P, (uint32_t)v)) *capped = 1;  // head of a chain: walks to the first item that knows
    }
  }
}
// step 2b (search_dev.cuh): unary branch nodes become chain nodes
// (over the list of branch nodes; `fresh` / `newseed` are zeroed by the caller)
__global__ void k_reclass(GraphView g, PeelView P, const uint32_t *ids, uint32_t n_ids, uint32_t *fresh, uint32_t *newseed,
                          unsigned long long *n_fresh, unsigned long long *n_seed) {
  __shared__ unsigned sh[2];
  BlockCounters<2> bc(sh);
  const uint32_t n32 = (n_ids + 31) & ~31u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += gridDim.x * blockDim.x) {
    const uint32_t b = i < n_ids ? ids[i] : 0u;
    const int r = i < n_ids ? reclass_item(g, P, b) : 0;
    if (r == 1) atomicOr(&fresh[b >> 5], 1u << (b & 31));
    if (r == 2) atomicOr(&newseed[b >> 5], 1u << (b & 31));
    unsigned mf = __ballot_sync(0xffffffffu, r == 1), ms = __ballot_sync(0xffffffffu, r == 2);
    if ((threadIdx.x & 31) == 0) {
      bc.add(0, (unsigned)__popc(mf));
      bc.add(1, (unsigned)__popc(ms));
    }
  }
  unsigned long long *dst[2] = {n_fresh, n_seed};
  bc.flush(dst);
}
__global__ void k_reclass_jump(PeelView P, const uint32_t *fresh, const uint32_t *ids, uint32_t n_ids, int *unresolved) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool u = i < n_ids && reclass_jump_item(P, fresh, ids[i]);
  if (__any_sync(0xffffffffu, u) && (threadIdx.x & 31) == 0) *unresolved = 1;
}
__global__ void k_or_words(uint32_t *dst, const uint32_t *src, size_t n_words) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_words) dst[i] |= src[i];
}
__global__ void k_relink_final(uint64_t n, PeelView P) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 4 * stride) {
    uint32_t t[4];
    bool moved[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t v = i0 + q * stride;
      t[q] = v < n && bm_get(P.chain, (uint32_t)v) ? ld_cg_u32(&peel_ptr(P, (uint32_t)v)) : kNull;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) moved[q] = t[q] != kNull && bm_get(P.chain, t[q]);  // terminal re-classified by step 2b: rare
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (moved[q]) relink_final_item(P, (uint32_t)(i0 + q * stride));
  }
}
// exhaustive fallback: one pointer-jumping round over every item
__global__ void k_chain_jump(uint64_t n, PeelView P, int *unresolved) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool u = i < n && chain_jump_item(P, (uint32_t)i);
  if (__any_sync(0xffffffffu, u) && (threadIdx.x & 31) == 0) *unresolved = 1;
}
// step 3a: (terminal -> dependent branch node) multimap
__global__ void k_dep_register(GraphView g, PeelView P, const uint32_t *ids, uint32_t n_ids) {  // over the branch nodes
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_ids; i += gridDim.x * blockDim.x)
    dep_register_item(g, P, ids[i]);
}
// step 3b: asynchronous work list over peeled terminals; each thread follows its cascade with a
// small private stack, the rest goes to the next work list
constexpr int kPeelLocal = 24;
__global__ void k_cpeel(PeelView P, const uint32_t *work, uint32_t n_work, uint32_t *next, uint32_t *next_count,
                        uint32_t next_cap, int *overflow) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_work) return;
  uint32_t lst[kPeelLocal];
  int sp = 0;
  lst[sp++] = work[i];
  while (sp) {
    uint32_t t = lst[--sp];
    cpeel_one(P, t, [&](uint32_t b) {
      if (sp < kPeelLocal) {
        lst[sp++] = b;
      } else {
        uint32_t idx = atomicAdd(next_count, 1u);
        if (idx < next_cap) next[idx] = b;
        else *overflow = 1;
      }
    });
  }
}
// step 4: invalid bits of the removed items, final navigation cache, "After pruning" statistics
// (:443-452): valid count and tips
// `merge` (zeroed by the caller): for every item e that stays valid and carries a non-first label (5..8,
// sdbg_def.h:11-29), the bits of the valid items of the node Forward(e) — the items with a second valid incoming edge,
// which is all ChunkStartNodes' in-degree test (:398) can accept; k_candidates_merge evaluates Backward + the group
// scan for those items only.
__global__ void k_peel_final(GraphView g, PeelView P, uint32_t *merge, unsigned long long *n_peeled, unsigned long long *n_valid,
                             unsigned long long *n_tips) {
  __shared__ unsigned sh[3];
  BlockCounters<3> bc(sh);
  const uint64_t n32 = (g.n + 31) & ~31ull;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += (uint64_t)gridDim.x * blockDim.x) {
    bool peeled = false, valid = false, tip = false;
    uint64_t a = 0;
    if (i < g.n) {
      a = peel_final_item(g, P, (uint32_t)i, peeled);
      valid = !peeled && !get_inv(g, (uint32_t)i);
      tip = valid && (a >> 32) == 0;
      if (valid && a && get_w(g, (uint32_t)i) > 4) {
        // the cache entry names the node's last item nx and its valid items nx - j (mask bit j): one word, or two
        // when the node straddles a 32-item boundary
        const uint32_t nx = (uint32_t)a, pos = nx & 31u;
        const uint64_t run = (uint64_t)__brev((uint32_t)(a >> 32) & 0xFFu) << (pos + 1u);  // bit j of the mask -> 32 + pos - j
        if ((uint32_t)(run >> 32)) atomicOr(&merge[nx >> 5], (uint32_t)(run >> 32));
        if ((uint32_t)run && nx >= 32u) atomicOr(&merge[(nx >> 5) - 1u], (uint32_t)run);
      }
    }
    unsigned mp = __ballot_sync(0xffffffffu, peeled), mv = __ballot_sync(0xffffffffu, valid),
             mt = __ballot_sync(0xffffffffu, tip);
    __syncwarp();
    if (i < g.n) P.adj[i] = a;  // every lane has finished reading adj[] of its own item; children's entries
                                // are only read through chain/ptr/tpeeled, never through adj[]
    if ((threadIdx.x & 31) == 0) {
      if (mp) {
        uint32_t *w32 = (uint32_t *)&g.fb[i >> 8].inv[(i >> 6) & 3] + ((i >> 5) & 1);
        *w32 |= mp;  // this warp owns the word
      }
      bc.add(0, (unsigned)__popc(mp));
      bc.add(1, (unsigned)__popc(mv));
      bc.add(2, (unsigned)__popc(mt));
    }
  }
  unsigned long long *dst[3] = {n_peeled, n_valid, n_tips};
  bc.flush(dst);
}

// CollectTips (reference src/cycle_finder.cpp:346-357) on the CURRENT graph state: ascending ids of the valid items
// with out-degree 0, compacted on the device (dead ends -> their valid incoming edges -> bitmap -> ids).
int run_collect_tips(mg_graph *g, DevBuf &ids, uint64_t *n_tips) {
  GraphView &v = g->view;
  cudaStream_t s = g->stream;
  const uint64_t n = v.n;
  *n_tips = 0;
  if (!n) return MG_OK;
  TimedScope ts(g, T_TIPS, n);
  const size_t bm_bytes = ((n + 31) / 32 + 1) * 4;
  DevBuf tipbits, dead, dead_ids, counter;
  MG_TRY(tipbits.alloc(bm_bytes));
  MG_TRY(dead.alloc(bm_bytes));
  MG_TRY(counter.alloc(8));
  MG_CUDA(cudaMemsetAsync(counter.p, 0, 8, s));
  MG_CUDA(cudaMemsetAsync(tipbits.p, 0, bm_bytes, s));
  uint64_t n_dead = 0;
  MG_LAUNCH(k_dead_ends, stream_grid(g, n, 256), 256, 0, s, v, dead.as<uint32_t>());
  MG_TRY(bitmap_to_ids(dead.as<uint32_t>(), n, dead_ids, &n_dead, s));
  if (n_dead)
    MG_LAUNCH(k_tips_of_dead_ends, stream_grid(g, n_dead, 256), 256, 0, s, v, dead_ids.as<uint32_t>(), (uint32_t)n_dead,
              tipbits.as<uint32_t>(), counter.as<unsigned long long>());
  return bitmap_to_ids(tipbits.as<uint32_t>(), n, ids, n_tips, s);
}

int run_prune(mg_graph *g, mg_prune_stats *st) {
  GraphView &v = g->view;
  cudaStream_t s = g->stream;
  const uint64_t n = v.n;
  memset(st, 0, sizeof *st);
  // [0] tips before, [1] mult1, [2] peeled, [3] valid, [4] tips after, [5] pairs, [6] next_count(u32) | overflow, [7] unresolved
  DevBuf counters;
  MG_TRY(counters.alloc(8 * 8));
  MG_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  unsigned long long *c = counters.as<unsigned long long>();
  const unsigned grid = grid_for(n, 256);
  const size_t bm_bytes = (size_t)((n + 31) / 32) * 4 + 16;
  // --- CollectTips on the loaded graph (:436)
  DevBuf tipbits;
  MG_TRY(tipbits.alloc(bm_bytes));
  if (n) {
    TimedScope ts(g, T_TIPS, n);
    if (getenv("MCAAT_TIPS_LITERAL")) {
      MG_LAUNCH(k_tips, stream_grid(g, n, 256), 256, 0, s, v, tipbits.as<uint32_t>(), c + 0);
    } else {
      DevBuf dead, dead_ids;
      uint64_t n_dead = 0;
      MG_TRY(dead.alloc(bm_bytes));
      MG_LAUNCH(k_dead_ends, stream_grid(g, n, 256), 256, 0, s, v, dead.as<uint32_t>());
      MG_TRY(bitmap_to_ids(dead.as<uint32_t>(), n, dead_ids, &n_dead, s));
      MG_CUDA(cudaMemsetAsync(tipbits.p, 0, bm_bytes, s));
      if (n_dead)
        MG_LAUNCH(k_tips_of_dead_ends, stream_grid(g, n_dead, 256), 256, 0, s, v, dead_ids.as<uint32_t>(), (uint32_t)n_dead,
                  tipbits.as<uint32_t>(), c + 0);
    }
  }
  // --- InvalidateMultiplicityOneNodes (:439)
  if (n) {
    TimedScope ts(g, T_MULT1, n);
    MG_LAUNCH(k_mult1, stream_grid(g, n, 256), 256, 0, s, v, c + 1);
  }
  // --- RecursiveReduction for every tip (:440-442), valid-edge count (:443-449), CollectTips (:451)
  MG_TRY(g->adj.ensure((size_t)n * 8 + 16));
  if (n) {
    DevBuf ent, chain, tpeeled, dep, work, next, recl;
    MG_TRY(ent.alloc((size_t)n * 8 + 16));
    MG_TRY(chain.alloc(bm_bytes));
    MG_TRY(tpeeled.alloc(bm_bytes));
    PeelView P;
    P.adj = g->adj.as<uint64_t>();
    P.ent = ent.as<uint64_t>();
    P.chain = chain.as<uint32_t>();
    P.tpeeled = tpeeled.as<uint32_t>();
    P.tipbits = tipbits.as<uint32_t>();
    P.recl = nullptr;
    P.dep = nullptr;
    P.dep_log = 0;
    P.split_ids = nullptr;
    P.sterm = nullptr;
    P.n_split = 0;
    {
      const char *e = getenv("MCAAT_PEEL_SPLIT_LOG");  // diagnostic: splitter density 2^-k (default 7: one chain node in 128)
      const int k = e && *e ? atoi(e) : 7;
      P.split_shift = 32u - (uint32_t)(k < 2 ? 2 : k > 12 ? 12 : k);
    }
    DevBuf branch, branch_ids, split;
    MG_TRY(branch.alloc(bm_bytes));
    MG_TRY(split.alloc(bm_bytes));
    {
      TimedScope ts(g, T_ADJ, n);
      MG_LAUNCH(k_chain_init, stream_grid(g, n, 256), 256, 0, s, v, P, branch.as<uint32_t>(), split.as<uint32_t>(), c + 5);
    }
    uint64_t n_work = 0;
    MG_TRY(bitmap_to_ids(tpeeled.as<uint32_t>(), n, work, &n_work, s));  // seeds
    TimedScope ts(g, T_PEEL, n);
    uint64_t n_branch = 0;
    if (n_work) MG_TRY(bitmap_to_ids(branch.as<uint32_t>(), n, branch_ids, &n_branch, s));
    if (n_work) {
      int *unresolved = (int *)(c + 7), *capped = (int *)(c + 7) + 1;
      {  // chain nodes -> terminals through splitters (one chain node in 128)
        DevBuf split_ids, sterm;
        MG_CUDA(cudaMemsetAsync(c + 7, 0, 8, s));  // (the splitters were marked by the classification pass)
        uint64_t n_split = 0;
        MG_TRY(bitmap_to_ids(split.as<uint32_t>(), n, split_ids, &n_split, s));
        if (n_split) {
          MG_TRY(sterm.alloc(n_split * 4));
          P.split_ids = split_ids.as<uint32_t>();
          P.sterm = sterm.as<uint32_t>();
          P.n_split = (uint32_t)n_split;
          MG_LAUNCH(k_split_walk, grid_for(n_split, 128), 128, 0, s, P, (uint32_t)n_split, capped);
          for (int round = 0; round < 36; round += 4) {
            MG_CUDA(cudaMemsetAsync(unresolved, 0, 4, s));
            for (int k = 0; k < 4; ++k)
              MG_LAUNCH(k_split_jump, grid_for(n_split, 256), 256, 0, s, P, (uint32_t)n_split, unresolved);
            int h = 0;
            MG_CUDA(cudaMemcpyAsync(&h, unresolved, 4, cudaMemcpyDeviceToHost, s));
            MG_CUDA(cudaStreamSynchronize(s));
            if (!h) break;
          }
          MG_LAUNCH(k_split_settle, grid_for(n_split, 256), 256, 0, s, P, (uint32_t)n_split);
        }
        MG_LAUNCH(k_chain_resolve, stream_grid(g, n, 256), 256, 0, s, n, P, capped);
        int hc = 0;
        MG_CUDA(cudaMemcpyAsync(&hc, capped, 4, cudaMemcpyDeviceToHost, s));
        MG_CUDA(cudaStreamSynchronize(s));
        if (hc || getenv("MCAAT_PEEL_EXHAUSTIVE")) {  // a walk hit its hop cap: exhaustive pointer jumping
          if (getenv("MCAAT_PEEL_VERBOSE")) fprintf(stderr, "[peel] exhaustive pointer jumping (walk cap hit: %d)\n", hc);
          for (int round = 0; round < 34; round += 2) {
            MG_CUDA(cudaMemsetAsync(unresolved, 0, 4, s));
            MG_LAUNCH(k_chain_jump, grid, 256, 0, s, n, P, unresolved);
            MG_LAUNCH(k_chain_jump, grid, 256, 0, s, n, P, unresolved);
            int h = 0;
            MG_CUDA(cudaMemcpyAsync(&h, unresolved, 4, cudaMemcpyDeviceToHost, s));
            MG_CUDA(cudaStreamSynchronize(s));
            if (!h) break;
          }
        }
      }
      if (!getenv("MCAAT_PEEL_NO_RECLASS")) {  // unary branch nodes -> chain nodes, a few rounds
        DevBuf fresh, newseed, fresh_ids;
        MG_TRY(recl.alloc(bm_bytes));
        MG_CUDA(cudaMemsetAsync(recl.p, 0, bm_bytes, s));
        P.recl = recl.as<uint32_t>();
        MG_TRY(fresh.alloc(bm_bytes));
        MG_TRY(newseed.alloc(bm_bytes));
        const size_t n_words = (size_t)((n + 31) / 32);
        DevBuf rc;
        MG_TRY(rc.alloc(16));
        for (int round = 0; round < 4; ++round) {
          MG_CUDA(cudaMemsetAsync(rc.p, 0, 16, s));
          MG_CUDA(cudaMemsetAsync(fresh.p, 0, bm_bytes, s));
          MG_CUDA(cudaMemsetAsync(newseed.p, 0, bm_bytes, s));
          if (n_branch)
            MG_LAUNCH(k_reclass, stream_grid(g, n_branch, 256), 256, 0, s, v, P, branch_ids.as<uint32_t>(), (uint32_t)n_branch,
                      fresh.as<uint32_t>(), newseed.as<uint32_t>(), rc.as<unsigned long long>(), rc.as<unsigned long long>() + 1);
          unsigned long long hc[2];
          MG_CUDA(cudaMemcpyAsync(hc, rc.p, 16, cudaMemcpyDeviceToHost, s));
          MG_CUDA(cudaStreamSynchronize(s));
          if (hc[0]) {
            uint64_t n_fresh = 0;
            MG_TRY(bitmap_to_ids(fresh.as<uint32_t>(), n, fresh_ids, &n_fresh, s));
            for (int r2 = 0; r2 < 36; r2 += 4) {
              MG_CUDA(cudaMemsetAsync(unresolved, 0, 4, s));
              for (int k = 0; k < 4; ++k)
                MG_LAUNCH(k_reclass_jump, grid_for(n_fresh, 256), 256, 0, s, P, fresh.as<uint32_t>(),
                          fresh_ids.as<uint32_t>(), (uint32_t)n_fresh, unresolved);
              int h = 0;
              MG_CUDA(cudaMemcpyAsync(&h, unresolved, 4, cudaMemcpyDeviceToHost, s));
              MG_CUDA(cudaStreamSynchronize(s));
              if (!h) break;
            }
            // the round's nodes are chain nodes from now on; pointers that name one of them as a terminal are
            // followed through it by terminal_of (no pass over the items)
            MG_LAUNCH(k_or_words, grid_for(n_words, 256), 256, 0, s, chain.as<uint32_t>(), fresh.as<uint32_t>(), n_words);
            MG_LAUNCH(k_or_words, grid_for(n_words, 256), 256, 0, s, recl.as<uint32_t>(), fresh.as<uint32_t>(), n_words);
          }
          if (hc[1])
            MG_LAUNCH(k_or_words, grid_for(n_words, 256), 256, 0, s, tpeeled.as<uint32_t>(), newseed.as<uint32_t>(), n_words);
          if (getenv("MCAAT_PEEL_VERBOSE"))
            fprintf(stderr, "[peel] reclass round %d: %llu unary branch nodes, %llu new seeds\n", round, hc[0], hc[1]);
          if (hc[0] + hc[1] < 64) break;
        }
        // Pointers that name a re-classified node as their terminal are followed through it by terminal_of in the passes
        // below as well (one hop per round of step 2b at most, five at most).  Since the final pass looks up one terminal
        // per item instead of two, that costs less than moving every such pointer on in a pass over all items
        // (k_relink_final, 3.2 ms at C2; MCAAT_PEEL_RELINK=1 runs it: same result).
        if (getenv("MCAAT_PEEL_RELINK")) {
          MG_LAUNCH(k_relink_final, stream_grid(g, n, 256), 256, 0, s, n, P);
          P.recl = nullptr;
        }
        MG_TRY(bitmap_to_ids(tpeeled.as<uint32_t>(), n, work, &n_work, s));  // collected tips + new seeds
      }
      unsigned long long pairs = 0;
      MG_CUDA(cudaMemcpyAsync(&pairs, c + 5, 8, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaStreamSynchronize(s));
      uint32_t dep_log = 10;
      while (((uint64_t)1 << dep_log) < 2 * pairs + 16) ++dep_log;
      MG_TRY(dep.alloc((size_t)8 << dep_log));
      MG_CUDA(cudaMemsetAsync(dep.p, 0xFF, dep.bytes, s));
      P.dep = dep.as<uint64_t>();
      P.dep_log = dep_log;
      if (n_branch) MG_LAUNCH(k_dep_register, grid_for(n_branch, 256), 256, 0, s, v, P, branch_ids.as<uint32_t>(), (uint32_t)n_branch);
      uint32_t next_cap = (uint32_t)(pairs + 1024 < 0xFFFFFFF0ull ? pairs + 1024 : 0xFFFFFFF0ull);
      uint32_t *next_count = (uint32_t *)(c + 6);
      int *overflow = (int *)(c + 6) + 1;
      MG_TRY(next.alloc((size_t)next_cap * 4));
      while (n_work) {
        MG_CUDA(cudaMemsetAsync(c + 6, 0, 8, s));
        MG_LAUNCH(k_cpeel, grid_for(n_work, 128), 128, 0, s, P, work.as<uint32_t>(), (uint32_t)n_work,
                  next.as<uint32_t>(), next_count, next_cap, overflow);
        uint32_t h[2];
        MG_CUDA(cudaMemcpyAsync(h, c + 6, 8, cudaMemcpyDeviceToHost, s));
        MG_CUDA(cudaStreamSynchronize(s));
        if (h[1]) {
          set_error("peel: work list overflow (%u entries)", next_cap);
          return MG_EINTERNAL;
        }
        n_work = h[0];
        if (n_work) {
          MG_TRY(work.ensure((size_t)n_work * 4));
          MG_CUDA(cudaMemcpyAsync(work.p, next.p, (size_t)n_work * 4, cudaMemcpyDeviceToDevice, s));
        }
      }
    }
    MG_TRY(g->merge.ensure(bm_bytes));
    MG_CUDA(cudaMemsetAsync(g->merge.p, 0, bm_bytes, s));
    MG_LAUNCH(k_peel_final, stream_grid(g, n, 256), 256, 0, s, v, P, g->merge.as<uint32_t>(), c + 2, c + 3, c + 4);
    MG_CUDA(cudaStreamSynchronize(s));
  }
  unsigned long long h[5];
  MG_CUDA(cudaMemcpyAsync(h, c, sizeof h, cudaMemcpyDeviceToHost, s));
  MG_CUDA(cudaStreamSynchronize(s));
  flush_timers(g);
  st->n_tips_before = h[0];
  st->n_mult1 = h[1];
  st->n_peeled = h[2];
  st->n_valid = h[3];
  st->n_tips_after = h[4];
  g->pruned = true;
  return MG_OK;
}

// ------------------------------------------------------------------ mg_navigate
__global__ void k_navigate(GraphView g, uint64_t first, uint64_t count, int8_t *outdeg, int8_t *indeg,
                           uint64_t *out_ids, uint64_t *in_ids, uint16_t *mult) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  uint32_t e = (uint32_t)(first + t);
  uint32_t o[4] = {kNull, kNull, kNull, kNull}, in[4] = {kNull, kNull, kNull, kNull};
  int od = outgoing(g, e, o);
  int id = incoming(g, e, in);
  if (outdeg) outdeg[t] = (int8_t)od;
  if (indeg) indeg[t] = (int8_t)id;
  for (int j = 0; j < 4; ++j) {
    if (out_ids) out_ids[t * 4 + j] = (j < od) ? (uint64_t)o[j] : ~0ull;
    if (in_ids) in_ids[t * 4 + j] = (j < id) ? (uint64_t)in[j] : ~0ull;
  }
  if (mult) mult[t] = g.mult[e];
}

int run_navigate(mg_graph *g, uint64_t first, uint64_t count, int8_t *outdeg, int8_t *indeg, uint64_t *out_ids,
                 uint64_t *in_ids, uint16_t *mult) {
  cudaStream_t s = g->stream;
  const uint64_t chunk = 1u << 22;
  DevBuf d_od, d_id, d_o, d_i, d_m;
  uint64_t cap = count < chunk ? count : chunk;
  MG_TRY(d_od.alloc(cap));
  MG_TRY(d_id.alloc(cap));
  MG_TRY(d_o.alloc(cap * 32));
  MG_TRY(d_i.alloc(cap * 32));
  MG_TRY(d_m.alloc(cap * 2));
  for (uint64_t off = 0; off < count; off += chunk) {
    uint64_t c = count - off < chunk ? count - off : chunk;
    {
      TimedScope ts(g, T_NAV, c);
      MG_LAUNCH(k_navigate, grid_for(c, 128), 128, 0, s, g->view, first + off, c, d_od.as<int8_t>(), d_id.as<int8_t>(),
                d_o.as<uint64_t>(), d_i.as<uint64_t>(), d_m.as<uint16_t>());
    }
    if (outdeg) MG_CUDA(cudaMemcpyAsync(outdeg + off, d_od.p, c, cudaMemcpyDeviceToHost, s));
    if (indeg) MG_CUDA(cudaMemcpyAsync(indeg + off, d_id.p, c, cudaMemcpyDeviceToHost, s));
    if (out_ids) MG_CUDA(cudaMemcpyAsync(out_ids + off * 4, d_o.p, c * 32, cudaMemcpyDeviceToHost, s));
    if (in_ids) MG_CUDA(cudaMemcpyAsync(in_ids + off * 4, d_i.p, c * 32, cudaMemcpyDeviceToHost, s));
    if (mult) MG_CUDA(cudaMemcpyAsync(mult + off, d_m.p, c * 2, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
  }
  flush_timers(g);
  return MG_OK;
}

}  // namespace mg
