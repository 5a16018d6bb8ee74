// SURVEY 8(f) N4 — region extraction around the cycles: keep_crispr_regions_extended_by_k, reference
// src/spacer_ordering.cpp:78-139 (called from get_crispr_regions_extended_by_k :177-184 with k = read length in
// nodes, src/main_run_and_debug.cpp:40-41).  The reference grows an unordered_set by re-scanning ALL members
// each of the k rounds (O(k * |set|) hash operations and rank/select navigations on one core) and then walks
// every item of the graph.  Here: a level-synchronous multi-source breadth-first search over three n-bit
// bitmaps in HBM (member / current level / next level) on the resident device graph, followed by one streaming
// pass that ORs ~member into the invalid bitmap.  The result is a set, so it is bit-exact by construction
// whatever the scheduling.
#include <vector>

#include "common.cuh"
#include "search_dev.cuh"

namespace mg {

// seed ids -> member bitmap; ids new to the set also enter level 0
__global__ void k_region_seed(const uint32_t *ids, uint64_t n_ids, uint32_t *member, uint32_t *cur) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_ids; i += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t v = ids[i];
    if (bm_set_atomic(member, v)) bm_set_atomic(cur, v);
  }
}

// one level: every 32-item word of the current level's bitmap is scanned by one thread (grid-stride, coalesced
// 4-byte loads; almost all words are zero), set bits are expanded through the succinct navigation.
__global__ void k_region_level(GraphView g, uint32_t *member, const uint32_t *cur, uint32_t *next, uint64_t n_words,
                               unsigned long long *added) {
  __shared__ unsigned sh;
  if (threadIdx.x == 0) sh = 0;
  __syncthreads();
  unsigned mine = 0;
  for (uint64_t w = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; w < n_words; w += (uint64_t)gridDim.x * blockDim.x)
    if (cur[w]) mine += region_expand_word(g, member, cur, next, (uint32_t)w);
  if (mine) atomicAdd(&sh, mine);
  __syncthreads();
  if (threadIdx.x == 0 && sh) atomicAdd(added, (unsigned long long)sh);
}

// invalid |= ~member over the whole graph; counts[0] += invalidated, counts[1] += still valid
__global__ void k_region_keep(GraphView g, const uint32_t *member, uint64_t n_words64, unsigned long long *counts) {
  __shared__ unsigned sh[2];
  if (threadIdx.x < 2) sh[threadIdx.x] = 0;
  __syncthreads();
  unsigned gone = 0, kept = 0;
  for (uint64_t w = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; w < n_words64; w += (uint64_t)gridDim.x * blockDim.x)
    gone += region_keep_word(g, member, w, &kept);
  if (gone) atomicAdd(&sh[0], gone);
  if (kept) atomicAdd(&sh[1], kept);
  __syncthreads();
  if (threadIdx.x < 2 && sh[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}

static inline unsigned grid_for(const mg_graph *g, uint64_t n, int block) {
  uint64_t want = (n + block - 1) / block, cap = (uint64_t)g->sm_count * 16;
  return (unsigned)(want < cap ? (want ? want : 1) : cap);
}

int run_keep_regions(mg_graph *g, const uint32_t *ids_dev, uint64_t n_ids, uint64_t k, mg_region_stats *st) {
  const uint64_t n = g->view.n;
  cudaStream_t s = g->stream;
  mg_region_stats out;
  memset(&out, 0, sizeof out);
  out.n_seed_ids = n_ids;
  if (n) {
    TimedScope ts(g, T_REGION, n);
    const uint64_t n_words64 = (n + 63) / 64, n_words = n_words64 * 2;
    const size_t bm_bytes = (size_t)n_words * 4 + 16;
    DevBuf member, lvl_a, lvl_b, counters;
    MG_TRY(member.alloc(bm_bytes));
    MG_TRY(lvl_a.alloc(bm_bytes));
    MG_TRY(lvl_b.alloc(bm_bytes));
    MG_TRY(counters.alloc(4 * 8));
    MG_CUDA(cudaMemsetAsync(member.p, 0, bm_bytes, s));
    MG_CUDA(cudaMemsetAsync(lvl_a.p, 0, bm_bytes, s));
    MG_CUDA(cudaMemsetAsync(counters.p, 0, 4 * 8, s));
    unsigned long long *c = counters.as<unsigned long long>();
    if (n_ids) MG_LAUNCH(k_region_seed, grid_for(g, n_ids, 256), 256, 0, s, ids_dev, n_ids, member.as<uint32_t>(), lvl_a.as<uint32_t>());
    uint32_t *cur = lvl_a.as<uint32_t>(), *next = lvl_b.as<uint32_t>();
    // The host looks at the running total every kCheck levels only (one 8-byte read-back), so that a level costs
    // a memset and a launch, not a round trip; levels after the search has died out are no-ops on empty bitmaps.
    const uint64_t kCheck = 16;
    unsigned long long total = 0, prev_total = 0;
    uint64_t level = 0;
    while (level < k && n_ids) {
      uint64_t stop = level + kCheck < k ? level + kCheck : k;
      for (; level < stop; ++level) {
        MG_CUDA(cudaMemsetAsync(next, 0, bm_bytes, s));
        MG_LAUNCH(k_region_level, grid_for(g, n_words, 256), 256, 0, s, g->view, member.as<uint32_t>(), cur, next, n_words, c + 2);
        uint32_t *t = cur;
        cur = next;
        next = t;
      }
      MG_CUDA(cudaMemcpyAsync(&total, c + 2, 8, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaStreamSynchronize(s));
      if (total == prev_total) break;  // nothing new in the last kCheck levels: every later level is empty too
      prev_total = total;
    }
    out.levels = level;
    out.n_added = total;
    MG_LAUNCH(k_region_keep, grid_for(g, n_words64, 256), 256, 0, s, g->view, member.as<uint32_t>(), n_words64, c);
    unsigned long long h[2] = {0, 0};
    MG_CUDA(cudaMemcpyAsync(h, c, 16, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
    out.n_invalidated = h[0];
    out.n_kept = h[1];
    // the navigation cache and the chain tables of mg_prune describe the graph before this call
    if (out.n_invalidated) g->pruned = false;
  }
  MG_CUDA(cudaStreamSynchronize(s));
  flush_timers(g);
  if (st) *st = out;
  return MG_OK;
}

}  // namespace mg
