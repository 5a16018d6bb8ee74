// SURVEY 8(f) N4 — region extraction around the cycles: keep_crispr_regions_extended_by_k, reference
// src/spacer_ordering.cpp:78-139 (called from get_crispr_regions_extended_by_k :177-184 with k = read length in
// nodes, src/main_run_and_debug.cpp:40-41).  The reference grows an unordered_set by re-scanning ALL members
// each of the k rounds (O(k * |set|) hash operations and rank/select navigations on one core) and then walks
// every item of the graph.  Here: a level-synchronous multi-source breadth-first search on the resident device
// graph — only the members new in the previous round are expanded — followed by one streaming pass that ORs
// ~member into the invalid bitmap.  The result is a set, so it is bit-exact whatever the scheduling.
//
//   k_region_bfs   all k levels in ONE cooperative launch: the frontier is a queue of node ids in HBM, the
//                  grid walks it, appends the members it adds to the next queue and meets at a grid barrier.
//                  A level costs one barrier plus the dependent loads of one expansion (rank/select
//                  navigation), not a launch and a pass over n bits.
//   k_region_level the same level over bitmaps (cur -> next), taken from the level at which a queue would
//                  overflow (nodes that do not fit are recorded in a spill bitmap, so nothing is lost).
//   k_region_keep  invalid |= ~member, HBM-streaming: 3 x n/8 bytes.
#include <cooperative_groups.h>

#include <vector>

#include "common.cuh"
#include "search_dev.cuh"

namespace cg = cooperative_groups;

namespace mg {

struct RegionQueues {
  uint32_t *q[2];      // frontier queues (node ids), used alternately
  uint32_t cap;        // entries per queue
  unsigned *cnt;       // cnt[3]: entries appended to the queue of level l are counted in cnt[l % 3]
  uint32_t *spill;     // n-bit bitmap: members that did not fit into a queue (then cnt[slot] > cap)
  unsigned *levels;    // out: levels completed by k_region_bfs
  unsigned long long *added;  // out: members added
};

__device__ __forceinline__ void region_push(const RegionQueues &Q, int which, unsigned level_slot, uint32_t v) {
  unsigned pos = atomicAdd(&Q.cnt[level_slot], 1u);
  if (pos < Q.cap) {
    __stcg(&Q.q[which][pos], v);
  } else {
    bm_set_atomic(Q.spill, v);
  }
}

// seed ids -> member bitmap; ids new to the set form level 0
__global__ void k_region_seed(const uint32_t *ids, uint64_t n_ids, uint32_t *member, RegionQueues Q) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_ids; i += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t v = ids[i];
    if (bm_set_atomic(member, v)) region_push(Q, 0, 0, v);
  }
}

// queue entries -> bitmap (hand-over to the bitmap levels)
__global__ void k_region_queue_to_bits(const uint32_t *q, unsigned n_q, uint32_t *bits) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_q; i += gridDim.x * blockDim.x) bm_set_atomic(bits, q[i]);
}

__global__ void __launch_bounds__(256) k_region_bfs(GraphView g, uint32_t *member, RegionQueues Q, uint64_t k) {
  cg::grid_group grid = cg::this_grid();
  const unsigned gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
  unsigned level = 0;
  unsigned mine = 0;
  while (level < k) {
    const unsigned slot = level % 3, next_slot = (level + 1) % 3;
    // complete and the same for every thread: all appends to this level's queue happened before the last
    // barrier.  A count above the capacity means part of the level is in the spill bitmap: hand over.
    const unsigned n_cur = ld_cg32(&Q.cnt[slot]);
    if (n_cur == 0 || n_cur > Q.cap) break;
    if (gtid == 0) Q.cnt[(level + 2) % 3] = 0;     // the slot of level+2: nobody reads or appends to it now
    const uint32_t *cur = Q.q[level & 1];
    for (unsigned i = gtid; i < n_cur; i += gsize) {
      uint32_t fresh[8];
      int c = region_expand_node(g, member, __ldcg(&cur[i]), fresh);
      for (int j = 0; j < c; ++j) region_push(Q, (level + 1) & 1, next_slot, fresh[j]);
      mine += (unsigned)c;
    }
    ++level;
    grid.sync();
  }
  __shared__ unsigned sh;
  if (threadIdx.x == 0) sh = 0;
  __syncthreads();
  if (mine) atomicAdd(&sh, mine);
  __syncthreads();
  if (threadIdx.x == 0 && sh) atomicAdd(Q.added, (unsigned long long)sh);
  if (gtid == 0) *Q.levels = level;
}

// one level over bitmaps: every 32-item word of the current level's bitmap is looked at by one thread
// (grid-stride, coalesced 4-byte loads), set bits are expanded.
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
    // queue capacity: 2^22 ids (16 MB) unless the graph is smaller; MCAAT_REGION_QUEUE_LOG shrinks it (tests)
    const char *ql = getenv("MCAAT_REGION_QUEUE_LOG");
    int q_log = ql ? atoi(ql) : 22;
    if (q_log < 2) q_log = 2;
    if (q_log > 28) q_log = 28;
    uint64_t cap = 1ull << q_log;
    if (cap > n) cap = n;
    DevBuf member, spill, other, queues, ctl;
    MG_TRY(member.alloc(bm_bytes));
    MG_TRY(spill.alloc(bm_bytes));
    MG_TRY(queues.alloc(2 * cap * 4 + 16));
    MG_TRY(ctl.alloc(64));
    MG_CUDA(cudaMemsetAsync(member.p, 0, bm_bytes, s));
    MG_CUDA(cudaMemsetAsync(spill.p, 0, bm_bytes, s));
    MG_CUDA(cudaMemsetAsync(ctl.p, 0, 64, s));
    // ctl: [0,1] u64 keep counters | [2] u64 added | u32 at byte 24: cnt[3], levels
    unsigned long long *c = ctl.as<unsigned long long>();
    unsigned *cu = (unsigned *)(c + 3);
    RegionQueues Q;
    Q.q[0] = queues.as<uint32_t>();
    Q.q[1] = Q.q[0] + cap;
    Q.cap = (uint32_t)cap;
    Q.cnt = cu;
    Q.levels = cu + 3;
    Q.spill = spill.as<uint32_t>();
    Q.added = c + 2;
    uint64_t level = 0;
    unsigned long long total = 0;
    if (n_ids && k) {
      MG_LAUNCH(k_region_seed, grid_for(g, n_ids, 256), 256, 0, s, ids_dev, n_ids, member.as<uint32_t>(), Q);
      int per_sm = 0;
      MG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_region_bfs, 256, 0));
      if (per_sm < 1) {
        set_error("k_region_bfs does not fit on an SM");
        return MG_ECUDA;
      }
      if (per_sm > 4) per_sm = 4;
      GraphView view = g->view;
      uint32_t *member_p = member.as<uint32_t>();
      uint64_t k_arg = k > 0xFFFFFFF0ull ? 0xFFFFFFF0ull : k;  // the level counter is 32 bits; see the loop below
      void *args[] = {&view, &member_p, &Q, &k_arg};
      MG_CUDA(cudaLaunchCooperativeKernel((void *)k_region_bfs, dim3((unsigned)(g->sm_count * per_sm)), dim3(256), args, 0, s));
      ++g_launches;
      unsigned h[4] = {0, 0, 0, 0};
      MG_CUDA(cudaMemcpyAsync(h, cu, sizeof h, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaMemcpyAsync(&total, c + 2, 8, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaStreamSynchronize(s));
      level = h[3];
      const unsigned n_q = h[level % 3] < cap ? h[level % 3] : (unsigned)cap;  // entries of the pending level's queue
      if (level < k && n_q) {
        // a queue overflowed (or k exceeds the 32-bit level counter): the pending level = queue entries + spill
        // bitmap; continue over bitmaps.  The host reads the running total every kCheck levels only.
        MG_TRY(other.alloc(bm_bytes));
        uint32_t *cur = spill.as<uint32_t>(), *next = other.as<uint32_t>();
        if (n_q) MG_LAUNCH(k_region_queue_to_bits, grid_for(g, n_q, 256), 256, 0, s, Q.q[level & 1], n_q, cur);
        const uint64_t kCheck = 16;
        unsigned long long prev_total = total;
        while (level < k) {
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
      }
    } else if (n_ids) {
      MG_LAUNCH(k_region_seed, grid_for(g, n_ids, 256), 256, 0, s, ids_dev, n_ids, member.as<uint32_t>(), Q);
    }
    out.levels = level;
    out.n_added = total;
    MG_LAUNCH(k_region_keep, grid_for(g, n_words64, 256), 256, 0, s, g->view, member.as<uint32_t>(), n_words64, c);
    unsigned long long hk[2] = {0, 0};
    MG_CUDA(cudaMemcpyAsync(hk, c, 16, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
    out.n_invalidated = hk[0];
    out.n_kept = hk[1];
    // the navigation cache and the chain tables of mg_prune describe the graph before this call
    if (out.n_invalidated) g->pruned = false;
  }
  MG_CUDA(cudaStreamSynchronize(s));
  flush_timers(g);
  if (st) *st = out;
  return MG_OK;
}

// ------------------------------------------------------------------ valid subgraph (input of the SCC split)
// divide_graph_into_subgraphs / find_strongly_connected_components (reference src/spacer_ordering.cpp:54-76,
// 141-175) walk every item of the graph with IsValidEdge and call OutgoingEdges per valid node through hash
// maps.  After region extraction only a sliver of the graph is valid, so the device compacts it: ascending
// valid ids and, per valid node, the positions (in that list) of its valid out-neighbours in OutgoingEdges
// order.  Tarjan's visiting order, which fixes the reference's output order, is sequential and runs on the
// host over this compact form (mcaat_b200/host/regions_gpu.cpp).

// valid bits as 32-bit words + per-word popcounts
__global__ void k_valid_words(GraphView g, uint64_t n_words, uint32_t *bits, uint32_t *counts) {
  for (uint64_t w = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; w < n_words; w += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t first = w * 32;
    uint32_t v = 0;
    if (first < g.n) {
      uint64_t inv = g.fb[first >> 8].inv[(first >> 6) & 3];
      v = ~(uint32_t)(inv >> (first & 32));
      if (first + 32 > g.n) v &= 0xFFFFFFFFu >> (32 - (g.n - first));
    }
    bits[w] = v;
    counts[w] = __popc(v);
  }
}

// rows of the compact graph: one thread per bitmap word walks its valid items
__global__ void k_subgraph_rows(GraphView g, uint64_t n_words, const uint32_t *bits, const uint32_t *prefix, uint32_t *ids,
                                uint32_t *nbr4) {
  for (uint64_t w = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; w < n_words; w += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t word = bits[w];
    uint32_t row = prefix[w];
    while (word) {
      int b = __ffs(word) - 1;
      word &= word - 1;
      uint32_t v = (uint32_t)(w * 32 + b), out[4];
      int d = outgoing(g, v, out);
      ids[row] = v;
      for (int j = 0; j < 4; ++j) {
        uint32_t r = kNull;
        if (j < d) {
          uint32_t x = out[j];
          r = prefix[x >> 5] + __popc(bits[x >> 5] & ((1u << (x & 31)) - 1u));
        }
        nbr4[(uint64_t)row * 4 + j] = r;
      }
      ++row;
    }
  }
}

int run_valid_subgraph(mg_graph *g, mg_subgraph *out) {
  const uint64_t n = g->view.n;
  cudaStream_t s = g->stream;
  memset(out, 0, sizeof *out);
  std::vector<uint32_t> h_ids, h_nbr;
  uint32_t tot = 0;
  if (n) {
    TimedScope ts(g, T_REGION, n);
    const uint64_t n_words = (n + 31) / 32;
    DevBuf bits, prefix, total, ids, nbr4;
    MG_TRY(bits.alloc(n_words * 4 + 16));
    MG_TRY(prefix.alloc(n_words * 4 + 16));
    MG_TRY(total.alloc(4));
    MG_LAUNCH(k_valid_words, grid_for(g, n_words, 256), 256, 0, s, g->view, n_words, bits.as<uint32_t>(), prefix.as<uint32_t>());
    MG_TRY(exclusive_scan_u32(prefix.as<uint32_t>(), n_words, total.as<uint32_t>(), s));
    MG_CUDA(cudaMemcpyAsync(&tot, total.p, 4, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
    if (tot) {
      MG_TRY(ids.alloc((size_t)tot * 4));
      MG_TRY(nbr4.alloc((size_t)tot * 16));
      MG_LAUNCH(k_subgraph_rows, grid_for(g, n_words, 256), 256, 0, s, g->view, n_words, bits.as<uint32_t>(), prefix.as<uint32_t>(),
                ids.as<uint32_t>(), nbr4.as<uint32_t>());
      h_ids.resize(tot);
      h_nbr.resize((size_t)tot * 4);
      MG_CUDA(cudaMemcpyAsync(h_ids.data(), ids.p, (size_t)tot * 4, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaMemcpyAsync(h_nbr.data(), nbr4.p, (size_t)tot * 16, cudaMemcpyDeviceToHost, s));
    }
    MG_CUDA(cudaStreamSynchronize(s));
  }
  flush_timers(g);
  out->ids = (uint64_t *)malloc(((size_t)tot + 1) * 8);
  out->off = (uint64_t *)malloc(((size_t)tot + 1) * 8);
  size_t n_edges = 0;
  for (size_t i = 0; i < h_nbr.size(); ++i) n_edges += h_nbr[i] != kNull;
  out->nbr = (uint32_t *)malloc((n_edges + 1) * 4);
  if (!out->ids || !out->off || !out->nbr) {
    free(out->ids);
    free(out->off);
    free(out->nbr);
    memset(out, 0, sizeof *out);
    set_error("out of host memory for a subgraph of %u nodes", tot);
    return MG_ENOMEM;
  }
  size_t e = 0;
  for (uint32_t i = 0; i < tot; ++i) {
    out->ids[i] = h_ids[i];
    out->off[i] = e;
    for (int j = 0; j < 4; ++j)
      if (h_nbr[(size_t)i * 4 + j] != kNull) out->nbr[e++] = h_nbr[(size_t)i * 4 + j];
  }
  out->off[tot] = e;
  out->n_nodes = tot;
  out->n_edges = e;
  return MG_OK;
}

}  // namespace mg
