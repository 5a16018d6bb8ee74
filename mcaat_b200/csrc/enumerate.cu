// K7-K8: per-start bounded multicycle enumeration (reference src/cycle_finder.cpp:231-243,140-225)
// under the sequential `visited` coupling of the driver loop (:468-487).
//
// The reference processes start nodes one after another: a start already marked visited is
// skipped, a search sees every node marked by earlier searches.  Here a WINDOW of start nodes is
// searched speculatively in parallel against one `visited` snapshot (one warp per start, lane 0
// drives the search), then a single-block commit kernel walks the window in order:
//   visited[start] now           -> skipped (:476)
//   search examined a node that an earlier commit of this window marked -> stop; the next window
//                                    starts here and re-runs it against the updated bitmap
//   otherwise                    -> accepted: its cycles' nodes are marked visited (:217-222)
// "Examined" = every node accepted by _BackgroundCheck (:40-52) = the key set of the lock map plus
// the neighbour sets left on the stack at an early stop; nodes rejected because visited stay
// rejected under a superset, and the other two tests do not depend on `visited`.  The result is
// identical to the sequential loop for every window size.
#include <algorithm>

#include "common.cuh"
#include "search_dev.cuh"

namespace mg {

struct SlotLayout {  // per-slot arena (byte offsets inside one slot)
  uint32_t lock_log, keylog_cap, relax_cap, cyc_nodes_cap, depth_cap;
  size_t off_path, off_bl, off_stk, off_lock, off_keylog, off_relax, off_cyc_nodes, off_cyc_off, bytes;
};
static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static SlotLayout make_layout(int max_len, uint32_t lock_log, uint32_t relax_cap, uint32_t cyc_cap_cycles) {
  SlotLayout L;
  L.lock_log = lock_log;
  L.keylog_cap = (1u << lock_log) / 2;
  L.relax_cap = relax_cap;
  L.depth_cap = (uint32_t)max_len + 4;
  L.cyc_nodes_cap = cyc_cap_cycles * (uint32_t)(max_len + 1);
  size_t o = 0;
  L.off_path = o; o = align_up(o + (size_t)L.depth_cap * 4, 16);
  L.off_bl = o; o = align_up(o + (size_t)L.depth_cap * 4, 16);
  L.off_stk = o; o = align_up(o + (size_t)L.depth_cap * sizeof(USet), 16);
  L.off_lock = o; o = align_up(o + ((size_t)8 << lock_log), 16);
  L.off_keylog = o; o = align_up(o + (size_t)L.keylog_cap * 4, 16);
  L.off_relax = o; o = align_up(o + (size_t)L.relax_cap * 8, 16);
  L.off_cyc_nodes = o; o = align_up(o + (size_t)L.cyc_nodes_cap * 4, 16);
  L.off_cyc_off = o; o = align_up(o + (size_t)(kClusterBound + 1) * 4, 16);
  L.bytes = align_up(o, 128);
  return L;
}
__device__ __forceinline__ EnumArena arena_of(uint8_t *base, const SlotLayout &L) {
  EnumArena A;
  A.path = (uint32_t *)(base + L.off_path);
  A.bl = (int32_t *)(base + L.off_bl);
  A.stk = (USet *)(base + L.off_stk);
  A.lock = (uint64_t *)(base + L.off_lock);
  A.lock_log = L.lock_log;
  A.keylog = (uint32_t *)(base + L.off_keylog);
  A.keylog_cap = L.keylog_cap;
  A.relax = (uint64_t *)(base + L.off_relax);
  A.relax_cap = L.relax_cap;
  A.cyc_nodes = (uint32_t *)(base + L.off_cyc_nodes);
  A.cyc_nodes_cap = L.cyc_nodes_cap;
  A.cyc_off = (uint32_t *)(base + L.off_cyc_off);
  return A;
}

// one warp per window entry
__global__ void __launch_bounds__(128) k_enum_search(GraphView g, const uint64_t *__restrict__ adj,
                                                    const uint32_t *__restrict__ visited,
                                                    const uint32_t *__restrict__ starts, uint32_t cnt, int min_len,
                                                    int max_len, uint8_t *arena, SlotLayout L, EnumResult *results) {
  const uint32_t slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (slot >= cnt) return;
  uint8_t *base = arena + (size_t)slot * L.bytes;
  const uint32_t start = starts[slot];
  const bool pre = visited_bit(visited, start);
  if (!pre) {  // clear the lock map cooperatively
    uint4 *p = (uint4 *)(base + L.off_lock);
    const size_t n16 = ((size_t)8 << L.lock_log) / 16;
    const uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
    for (size_t i = lane; i < n16; i += 32) p[i] = ones;
  }
  __syncwarp();
  if (lane == 0) {
    EnumResult R;
    if (pre) {
      R.status = kEnumPreSkipped;
      R.n_cycles = R.n_nodes = R.n_keys = R.steps = R.expansions = 0;
    } else {
      EnumArena A = arena_of(base, L);
      enum_run(g, adj, visited, start, min_len, max_len, A, R);
    }
    results[slot] = R;
  }
}

enum SlotFate : int32_t { kFateNone = 0, kFateAccepted = 1, kFateSkipped = 2 };
struct CommitSummary {
  uint32_t n_committed;   // entries decided (accepted or skipped), in order
  int32_t stop_reason;    // 0 = window exhausted, 1 = conflict, 2 = arena overflow
  uint32_t total_cycles;  // over accepted entries
  uint32_t total_nodes;
};
struct SlotOut {
  int32_t fate;
  int32_t status;
  uint32_t n_cycles, n_nodes;
  uint32_t cyc_base, node_base;  // offsets into the window's gather buffers
  uint32_t expansions, steps;
};

// single block; walks the window in start order (the reference's loop order)
__global__ void __launch_bounds__(1024) k_enum_commit(uint32_t *visited, const uint32_t *__restrict__ starts,
                                                     uint32_t cnt, const uint8_t *arena, SlotLayout L,
                                                     const EnumResult *results, SlotOut *outs, CommitSummary *summary) {
  __shared__ uint32_t tot_c, tot_n;
  if (threadIdx.x == 0) { tot_c = 0; tot_n = 0; }
  __syncthreads();
  uint32_t i = 0;
  int reason = 0;
  for (; i < cnt; ++i) {
    const uint32_t s = starts[i];
    const EnumResult R = results[i];
    const uint8_t *base = arena + (size_t)i * L.bytes;
    // visited is written by this block only; volatile reads see the marks made after the barrier
    const volatile uint32_t *vis = visited;
    if ((vis[s >> 5] >> (s & 31)) & 1) {  // :476
      if (threadIdx.x == 0) {
        SlotOut o = {kFateSkipped, R.status, 0, 0, 0, 0, 0, 0};
        outs[i] = o;
      }
      continue;
    }
    if (R.status == kEnumOverflow) { reason = 2; break; }
    const uint32_t *keys = (const uint32_t *)(base + L.off_keylog);
    int conflict = 0;
    for (uint32_t j = threadIdx.x; j < R.n_keys; j += blockDim.x) {
      uint32_t k = keys[j];
      if (k != s && ((vis[k >> 5] >> (k & 31)) & 1)) conflict = 1;
    }
    if (__syncthreads_or(conflict)) { reason = 1; break; }
    const uint32_t *nodes = (const uint32_t *)(base + L.off_cyc_nodes);
    for (uint32_t j = threadIdx.x; j < R.n_nodes; j += blockDim.x) {  // :217-222
      uint32_t x = nodes[j];
      atomicOr(&visited[x >> 5], 1u << (x & 31));
    }
    if (threadIdx.x == 0) {
      SlotOut o = {kFateAccepted, R.status, R.n_cycles, R.n_nodes, tot_c, tot_n, R.expansions, R.steps};
      outs[i] = o;
      tot_c += R.n_cycles;
      tot_n += R.n_nodes;
    }
    __threadfence_block();
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    summary->n_committed = i;
    summary->stop_reason = reason;
    summary->total_cycles = tot_c;
    summary->total_nodes = tot_n;
  }
}

// copies the accepted entries' staged cycles into the window's contiguous output; block per entry
__global__ void k_enum_gather(uint32_t cnt, const uint8_t *arena, SlotLayout L, const SlotOut *outs,
                              uint32_t *win_nodes, uint32_t *win_cyc_len) {
  uint32_t i = blockIdx.x;
  if (i >= cnt) return;
  const SlotOut o = outs[i];
  if (o.fate != kFateAccepted) return;
  const uint8_t *base = arena + (size_t)i * L.bytes;
  const uint32_t *nodes = (const uint32_t *)(base + L.off_cyc_nodes);
  const uint32_t *off = (const uint32_t *)(base + L.off_cyc_off);
  for (uint32_t j = threadIdx.x; j < o.n_nodes; j += blockDim.x) win_nodes[o.node_base + j] = nodes[j];
  for (uint32_t j = threadIdx.x; j < o.n_cycles; j += blockDim.x) win_cyc_len[o.cyc_base + j] = off[j + 1] - off[j];
}

struct HostAccum {
  std::vector<uint64_t> searched, cyc_off, node_off, nodes;
  uint64_t skipped = 0, bound = 0, cap = 0, reruns = 0, rounds = 0, expansions = 0;
};

static int to_out(HostAccum &H, mg_cycles *out) {
  memset(out, 0, sizeof *out);
  auto dup = [](const std::vector<uint64_t> &v) -> uint64_t * {
    uint64_t *p = (uint64_t *)malloc((v.size() + 1) * 8);
    if (p && !v.empty()) memcpy(p, v.data(), v.size() * 8);
    return p;
  };
  out->n_searched = H.searched.size();
  out->searched_start = dup(H.searched);
  out->cyc_off = dup(H.cyc_off);
  out->n_cycles = H.node_off.size() - 1;
  out->node_off = dup(H.node_off);
  out->nodes = dup(H.nodes);
  if (!out->searched_start || !out->cyc_off || !out->node_off || !out->nodes) return MG_ENOMEM;
  out->n_skipped = H.skipped;
  out->n_bound_hits = H.bound;
  out->n_cap_hits = H.cap;
  out->n_reruns = H.reruns;
  out->n_rounds = H.rounds;
  out->expansions = H.expansions;
  return MG_OK;
}

// Runs one window [pos, pos+cnt) through search -> commit -> gather and appends the committed
// prefix to H.  Returns the commit summary.
static int run_window(mg_graph *g, const uint32_t *starts_dev, uint32_t cnt, int min_len, int max_len, DevBuf &arena,
                      const SlotLayout &L, DevBuf &results, DevBuf &outs, DevBuf &summary, DevBuf &win_nodes,
                      DevBuf &win_len, HostAccum &H, CommitSummary *sum_out) {
  cudaStream_t s = g->stream;
  {
    TimedScope ts(g, T_ENUM, cnt);
    MG_LAUNCH(k_enum_search, (cnt * 32 + 127) / 128, 128, 0, s, g->view, g->adj.as<uint64_t>(),
              g->visited.as<uint32_t>(), starts_dev, cnt, min_len, max_len, arena.as<uint8_t>(), L,
              results.as<EnumResult>());
  }
  CommitSummary sum;
  {
    TimedScope ts(g, T_COMMIT, cnt);
    MG_LAUNCH(k_enum_commit, 1, 1024, 0, s, g->visited.as<uint32_t>(), starts_dev, cnt, arena.as<uint8_t>(), L,
              results.as<EnumResult>(), outs.as<SlotOut>(), summary.as<CommitSummary>());
    MG_CUDA(cudaMemcpyAsync(&sum, summary.p, sizeof sum, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
    if (sum.total_nodes || sum.total_cycles) {
      MG_TRY(win_nodes.ensure((size_t)sum.total_nodes * 4 + 16));
      MG_TRY(win_len.ensure((size_t)sum.total_cycles * 4 + 16));
      MG_LAUNCH(k_enum_gather, sum.n_committed, 256, 0, s, sum.n_committed, arena.as<uint8_t>(), L, outs.as<SlotOut>(),
                win_nodes.as<uint32_t>(), win_len.as<uint32_t>());
    }
  }
  std::vector<SlotOut> ho(sum.n_committed);
  std::vector<uint32_t> hn(sum.total_nodes), hl(sum.total_cycles);
  if (sum.n_committed) MG_CUDA(cudaMemcpyAsync(ho.data(), outs.p, ho.size() * sizeof(SlotOut), cudaMemcpyDeviceToHost, s));
  if (sum.total_nodes) MG_CUDA(cudaMemcpyAsync(hn.data(), win_nodes.p, hn.size() * 4, cudaMemcpyDeviceToHost, s));
  if (sum.total_cycles) MG_CUDA(cudaMemcpyAsync(hl.data(), win_len.p, hl.size() * 4, cudaMemcpyDeviceToHost, s));
  MG_CUDA(cudaStreamSynchronize(s));
  std::vector<uint32_t> hs(sum.n_committed);
  if (sum.n_committed) MG_CUDA(cudaMemcpy(hs.data(), starts_dev, hs.size() * 4, cudaMemcpyDeviceToHost));
  for (uint32_t i = 0; i < sum.n_committed; ++i) {
    const SlotOut &o = ho[i];
    if (o.fate == kFateSkipped) { ++H.skipped; continue; }
    H.searched.push_back(hs[i]);                      // results[start] = cycles, :479 (empty stored)
    H.bound += o.status == kEnumBound;
    H.cap += o.status == kEnumCap;
    H.expansions += o.expansions;
    uint32_t np = o.node_base;
    for (uint32_t c = 0; c < o.n_cycles; ++c) {
      uint32_t len = hl[o.cyc_base + c];
      for (uint32_t j = 0; j < len; ++j) H.nodes.push_back(hn[np + j]);
      np += len;
      H.node_off.push_back(H.nodes.size());
    }
    H.cyc_off.push_back(H.node_off.size() - 1);
  }
  ++H.rounds;
  *sum_out = sum;
  return MG_OK;
}

static uint32_t env_u32(const char *name, uint32_t def, uint32_t lo, uint32_t hi) {
  const char *e = getenv(name);
  if (!e || !*e) return def;
  long v = strtol(e, nullptr, 10);
  if (v < (long)lo) v = lo;
  if (v > (long)hi) v = hi;
  return (uint32_t)v;
}

int run_find_cycles(mg_graph *g, const uint64_t *starts, uint64_t n, int min_len, int max_len, mg_cycles *out) {
  cudaStream_t s = g->stream;
  HostAccum H;
  H.cyc_off.push_back(0);
  H.node_off.push_back(0);
  if (!g->pruned || !g->adj.p) {
    set_error("mg_find_cycles: call mg_prune first (it builds the navigation cache)");
    return MG_EINVAL;
  }
  if (max_len < 1 || max_len > 4000 || min_len < 0) {
    set_error("mg_find_cycles: cycle length bounds (%d, %d] outside the supported range", min_len, max_len);
    return MG_ELIMIT;
  }
  if (n == 0) return to_out(H, out);
  std::vector<uint32_t> s32(n);
  for (uint64_t i = 0; i < n; ++i) {
    if (starts[i] >= g->view.n) {
      set_error("mg_find_cycles: start id %llu out of range", (unsigned long long)starts[i]);
      return MG_EINVAL;
    }
    s32[i] = (uint32_t)starts[i];
  }
  DevBuf starts_dev;
  MG_TRY(starts_dev.alloc(n * 4));
  MG_CUDA(cudaMemcpyAsync(starts_dev.p, s32.data(), n * 4, cudaMemcpyHostToDevice, s));

  const uint32_t max_slots = (uint32_t)std::min<uint64_t>(n, (uint64_t)g->sm_count * env_u32("MCAAT_ENUM_SLOTS_PER_SM", 16, 1, 64));
  const SlotLayout L = make_layout(max_len, env_u32("MCAAT_ENUM_LOCK_LOG", 13, 8, 24), 4096, 128);
  DevBuf arena, results, outs, summary, win_nodes, win_len;
  MG_TRY(arena.alloc((size_t)max_slots * L.bytes));
  MG_TRY(results.alloc((size_t)max_slots * sizeof(EnumResult)));
  MG_TRY(outs.alloc((size_t)max_slots * sizeof(SlotOut)));
  MG_TRY(summary.alloc(sizeof(CommitSummary)));
  // exact fallback arena for searches that overflow a slot; grown on demand
  DevBuf big_arena;
  uint32_t big_lock_log = L.lock_log + 4, big_relax = 1u << 16;

  uint64_t pos = 0;
  uint32_t window = max_slots;
  while (pos < n) {
    uint32_t cnt = (uint32_t)std::min<uint64_t>(window, n - pos);
    CommitSummary sum;
    MG_TRY(run_window(g, starts_dev.as<uint32_t>() + pos, cnt, min_len, max_len, arena, L, results, outs, summary,
                      win_nodes, win_len, H, &sum));
    pos += sum.n_committed;
    if (sum.stop_reason == 1) {  // conflict: re-run from here against the updated bitmap
      ++H.reruns;
      window = std::max<uint32_t>(16u, std::min<uint32_t>(max_slots, sum.n_committed * 2));
    } else if (sum.stop_reason == 2) {  // the entry at `pos` needs a larger arena
      for (;;) {
        SlotLayout LB = make_layout(max_len, big_lock_log, big_relax, kClusterBound);
        MG_TRY(big_arena.ensure(LB.bytes));
        CommitSummary s1;
        MG_TRY(run_window(g, starts_dev.as<uint32_t>() + pos, 1, min_len, max_len, big_arena, LB, results, outs,
                          summary, win_nodes, win_len, H, &s1));
        if (s1.stop_reason == 0 && s1.n_committed == 1) break;
        if (s1.stop_reason != 2 || big_lock_log >= 28) {
          set_error("mg_find_cycles: start %llu does not fit the largest search arena", (unsigned long long)starts[pos]);
          return MG_ELIMIT;
        }
        big_lock_log += 2;
        big_relax *= 4;
      }
      pos += 1;
    } else {
      window = std::min<uint32_t>(max_slots, window * 2);
    }
  }
  flush_timers(g);
  return to_out(H, out);
}

}  // namespace mg
