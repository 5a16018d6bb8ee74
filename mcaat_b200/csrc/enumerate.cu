// K7-K8: per-start bounded multicycle enumeration (reference src/cycle_finder.cpp:231-243,140-225)
// under the sequential `visited` coupling of the driver loop (:468-487).
//
// The reference processes start nodes one after another: a start already marked visited is
// skipped (:476), a search sees every node marked by the searches before it (:217-222).  Here the
// loop is a persistent-kernel pipeline that commits IN TICKET ORDER WHERE ORDER MATTERS:
//   * worker warps pull tickets (= positions in the start list) from an atomic counter, so the
//     start nodes in flight are always a contiguous window of the sequential order;
//   * a worker searches its start immediately, reading the live `visited` bitmap: the warp extracts
//     the ball the search can touch (parallel BFS, ball_dev.cuh) — everything a search reads or marks
//     lies inside it — claims its nodes (owner[x] = min ticket, atomicMin) and publishes "built";
//     lane 0 then runs the sequential search on the contracted ball;
//   * once every earlier ticket has published its ball (built front), the ticket knows whether any
//     EARLIER ticket's ball overlaps its own (min owner over its ball).  Two searches with disjoint
//     balls commute: neither reads what the other marks.  A ticket that no earlier ball overlaps is
//     INDEPENDENT: nothing that happens before it in the sequential order can touch its ball, later
//     tickets that overlap it see its claim and wait for it, so its speculative result is already the
//     sequential one and it commits at once, out of order (marks its cycles' nodes, :217-222);
//   * every other ticket is DEPENDENT and commits at its turn (commit front == ticket: all earlier
//     tickets are done), when it knows the exact bitmap of its position:
//       visited[start] now                      -> skipped, like :476
//       a node the search ACCEPTED is marked now -> the search saw a stale bitmap: it is re-run
//                                                   on the spot (exclusive, hence exact)
//       otherwise                               -> identical to the sequential search: its cycles'
//                                                   nodes are marked (:217-222), results appended
//     ("accepted" = the nodes of the extracted ball; nodes rejected for being visited stay rejected
//     under a superset, the other tests of _BackgroundCheck ignore `visited`);
//   * marks are only ever made by commits, and a commit never marks a node of an unfinished earlier
//     ticket's ball that it does not overlap — so a ticket that sees its own START marked, at any time,
//     is skipped for good (the marker precedes it in the order) and gives up at once.
// The committed sequence is identical to the reference's threads=1 loop for any number of workers.
#include <algorithm>
#include <chrono>

#include "common.cuh"
#include "ball_dev.cuh"
#include "ball_layout.cuh"
#include "search_dev.cuh"

namespace mg {

struct SlotLayout {  // per-worker arena in global memory (byte offsets inside one slot)
  uint32_t index_log, ent_cap, relax_cap, cyc_nodes_cap, depth_cap;
  size_t off_index, off_hot, off_cold, off_relax, off_cyc_nodes, off_cyc_off, bytes;
  uint32_t smem_bytes;  // per worker: path, pent, bl, stk live in shared memory
};
static SlotLayout make_layout(int max_len, uint32_t index_log, uint32_t relax_cap, uint32_t cyc_cap_cycles) {
  SlotLayout L;
  L.index_log = index_log;
  L.ent_cap = (1u << index_log) / 2;
  L.relax_cap = relax_cap;
  L.depth_cap = (uint32_t)max_len + 4;
  L.cyc_nodes_cap = cyc_cap_cycles * (uint32_t)(max_len + 1);
  size_t o = 0;
  L.off_hot = o; o = align_up(o + (size_t)L.ent_cap * sizeof(HotEnt), 128);
  L.off_cold = o; o = align_up(o + (size_t)L.ent_cap * sizeof(ColdEnt), 128);
  L.off_index = o; o = align_up(o + ((size_t)4 << index_log), 128);
  L.off_relax = o; o = align_up(o + (size_t)L.relax_cap * 8, 128);
  L.off_cyc_nodes = o; o = align_up(o + (size_t)L.cyc_nodes_cap * 4, 128);
  L.off_cyc_off = o; o = align_up(o + (size_t)(kClusterBound + 1) * 4, 128);
  L.bytes = align_up(o, 128);
  L.smem_bytes = (uint32_t)align_up((size_t)L.depth_cap * (4 + 4 + 4 + sizeof(USet)), 16);
  return L;
}
// hot per-level state (path, entry numbers, backtrack lengths, neighbour sets) in shared memory;
// the lock map, relax stack and staged cycles in the worker's global arena
__device__ __forceinline__ EnumArena arena_of(uint8_t *base, uint8_t *smem, const SlotLayout &L) {
  EnumArena A;
  A.stk = (USet *)smem;
  A.path = (uint32_t *)(smem + (size_t)L.depth_cap * sizeof(USet));
  A.pent = A.path + L.depth_cap;
  A.bl = (int32_t *)(A.pent + L.depth_cap);
  A.index = (uint32_t *)(base + L.off_index);
  A.index_log = L.index_log;
  A.hot = (HotEnt *)(base + L.off_hot);
  A.cold = (ColdEnt *)(base + L.off_cold);
  A.ent_cap = L.ent_cap;
  A.relax = (uint64_t *)(base + L.off_relax);
  A.relax_cap = L.relax_cap;
  A.cyc_nodes = (uint32_t *)(base + L.off_cyc_nodes);
  A.cyc_nodes_cap = L.cyc_nodes_cap;
  A.cyc_off = (uint32_t *)(base + L.off_cyc_off);
  return A;
}

enum TicketFate : int32_t { kFateNone = 0, kFateAccepted = 1, kFateSkipped = 2 };
struct TicketRec {
  int32_t fate;
  int32_t status;
  uint32_t n_cycles, n_nodes;
  uint32_t cyc_base, node_base;  // offsets into the output pools
  uint32_t expansions, steps;
  uint32_t us_searched, us_committed;  // microseconds since the worker entered the kernel
  uint32_t n_ball, us_build;           // size of the last extracted ball and the time its extraction took
  uint32_t n_chains;                   // records of its contraction
  uint32_t iters, relax_pops;          // of the last search: loop iterations on the contracted ball, relaxation entries
  uint32_t us_started, n_restarts;     // when the last search of this ticket began; searches started for it
  uint32_t us_claims, us_turn;         // when every earlier ticket had claimed its ball; when this ticket's turn came
  uint32_t root, n_watched;            // group (smallest ticket) and earlier unfinished tickets of it (0xFFFFFFFF: commit front)
};
struct PipeCtl {
  uint32_t ticket;       // next ticket to hand out
  uint32_t commit_idx;   // tickets [first, commit_idx) are committed
  uint32_t stop;         // 0 = run; else the pipeline drains: 1 = arena overflow, 2 = output pool full
  uint32_t pool_nodes;   // output cursors
  uint32_t pool_cycles;
  uint32_t n_reruns;
  uint32_t n_abandoned;  // searches given up because their start got marked
  uint32_t mark_epoch;   // number of commits that marked nodes so far (waiting workers re-check on change)
  uint32_t built_idx;    // tickets [first, built_idx) have extracted and claimed their ball (or need none)
  uint32_t ooo_limit;    // tickets beyond this one must commit in order (an earlier ball could not be claimed in full)
  uint32_t n_ooo;        // commits made out of order (independent tickets)
  unsigned long long searches_run;
};
// Moves a front (commit_idx / built_idx) from `from` over every ticket whose flag is set — all lanes, 32 flags per
// round trip — and tells the ticket it stops at that its turn has come (go[f] = 1; waiting workers poll their own
// word, so there is no hot address).  Both sides of the race "flag set" / "front arrives" write, fence and read,
// so one of them always continues the sweep.
__device__ __forceinline__ void sweep_front(uint32_t *front, const uint32_t *flags, uint32_t *go, uint32_t from,
                                            uint32_t n_end, int lane) {
  uint32_t f = from;
  for (;;) {
    const uint32_t c = f + (uint32_t)lane;
    const bool set = c < n_end && *(const volatile uint32_t *)&flags[c] != 0;
    const unsigned m = __ballot_sync(0xffffffffu, set);
    const int run = m == 0xffffffffu ? 32 : __ffs((int)~m) - 1;
    f += (uint32_t)run;
    if (run == 32) continue;
    int again = 0;
    if (lane == 0) {
      atomicMax(front, f);
      fence_sc();
      if (f < n_end && *(const volatile uint32_t *)&flags[f] != 0) again = 1;
      else if (f < n_end && go) *(volatile uint32_t *)&go[f] = 1u;
    }
    if (!__shfl_sync(0xffffffffu, again, 0)) break;
  }
}
// Union-find over tickets (parent[t] <= t; the root of a group is its smallest ticket): two tickets whose balls share a
// node are united when the later claim finds the earlier one (k_enum_pipeline).  Tickets of different groups have
// disjoint balls and commute; inside a group commits go in ticket order.
__device__ __forceinline__ uint32_t uf_find(uint32_t *parent, uint32_t x) {
  for (;;) {
    const uint32_t p = *(volatile uint32_t *)&parent[x];
    if (p == x) return x;
    const uint32_t gp = *(volatile uint32_t *)&parent[p];
    if (gp != p) atomicMin(&parent[x], gp);  // path halving; parents only ever decrease
    x = p;
  }
}
__device__ __forceinline__ void uf_union(uint32_t *parent, uint32_t a, uint32_t b) {
  for (;;) {
    const uint32_t ra = uf_find(parent, a), rb = uf_find(parent, b);
    if (ra == rb) return;
    const uint32_t hi = max(ra, rb), lo = min(ra, rb);
    if (atomicCAS(&parent[hi], hi, lo) == hi) return;
  }
}
// advances a front over every ticket whose flag is set (lane 0 of any worker; used where no turn has to be announced)
__device__ __forceinline__ void advance_front(uint32_t *front, const uint32_t *flags, uint32_t n_end) {
  for (;;) {
    const uint32_t f = *(volatile uint32_t *)front;
    if (f >= n_end || !*(const volatile uint32_t *)&flags[f]) break;
    atomicCAS(front, f, f + 1);
  }
}

__device__ __forceinline__ uint32_t ld_vol(const uint32_t *p) { return *(const volatile uint32_t *)p; }

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// One warp = one worker (kWarpsPerBlock workers per block).  See the header comment for the protocol.
// k_enum_solo searches with the hash-map search enum_run (search_dev.cuh): the host launches it with
// ONE worker and a large arena for a ticket whose ball overflowed the arenas of k_enum_pipeline below.
constexpr int kWarpsPerBlock = 4;
__global__ void __launch_bounds__(32 * kWarpsPerBlock) k_enum_solo(GraphView g, const uint64_t *__restrict__ adj, uint32_t *visited,
                                                     const uint32_t *__restrict__ starts, uint32_t n_end, int min_len,
                                                     int max_len, uint8_t *arena, SlotLayout L, PipeCtl *ctl,
                                                     TicketRec *recs, uint32_t *pool_nodes, uint32_t pool_nodes_cap,
                                                     uint32_t *pool_len, uint32_t pool_len_cap) {
  extern __shared__ __align__(16) uint8_t smem_all[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t worker = blockIdx.x * kWarpsPerBlock + warp;
  uint8_t *base = arena + (size_t)worker * L.bytes;
  EnumArena A = arena_of(base, smem_all + (size_t)warp * L.smem_bytes, L);
  unsigned long long searches = 0;
  uint32_t abandoned = 0, reruns = 0;
  const unsigned long long t_kernel = gtime_ns();
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(&ctl->ticket, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= n_end || ld_vol(&ctl->stop)) break;
    const uint32_t s = starts[t];
    EnumResult R;
    R.status = kEnumSkipped;
    R.n_cycles = R.n_nodes = R.n_keys = R.steps = R.expansions = 0;
    int stopped = 0;
    // speculative search against the live bitmap (lane 0); gives up when the start gets marked
    auto speculate = [&]() {
      ++searches;
      enum_run(g, adj, visited, s, min_len, max_len, A, R, [&]() {
        if (ld_vol(&ctl->stop)) { stopped = 1; return true; }
        return visited_live(visited, s);
      });
      if (R.status == kEnumSkipped && !stopped) ++abandoned;
    };
    // does the finished search rest on a node that has been marked since? (warp-cooperative)
    auto stale = [&](uint32_t n_keys) {
      int conflict = 0;
      for (uint32_t j = lane; j < n_keys; j += 32) {
        uint32_t k = A.hot[j].key;
        if (k != s && visited_live(visited, k)) conflict = 1;
      }
      return __any_sync(0xffffffffu, conflict) != 0;
    };
    if (lane == 0 && !visited_live(visited, s)) speculate();
    __syncwarp();  // lane 0's arena writes (key log, staged cycles) are visible to the warp
    const uint32_t us_searched = (uint32_t)((gtime_ns() - t_kernel) / 1000);
    // ---- wait for this ticket's turn; meanwhile keep the speculative result fresh, so that the
    //      repeat of a stale search happens here, in parallel, and not inside the ordered commit.
    //      `visited` only changes inside commits that mark nodes (mark_epoch), so the result is
    //      re-examined only when the epoch has moved; the poll itself is one 32-byte sector and backs
    //      off with the distance to the commit front, so waiting workers leave the L2 to the searches.
    uint32_t epoch_checked = 0xFFFFFFFFu;
    for (;;) {
      uint32_t ci = 0, ep = 0;
      int st = stopped;
      if (lane == 0) {
        ci = ld_vol(&ctl->commit_idx);
        ep = ld_vol(&ctl->mark_epoch);
        st |= (int)ld_vol(&ctl->stop);
      }
      ci = __shfl_sync(0xffffffffu, ci, 0);
      ep = __shfl_sync(0xffffffffu, ep, 0);
      st = __shfl_sync(0xffffffffu, st, 0);
      if (st) { stopped = 1; break; }
      if (ci == t) break;
      int status = __shfl_sync(0xffffffffu, R.status, 0);
      if (ep != epoch_checked && status != kEnumSkipped && status != kEnumOverflow) {
        epoch_checked = ep;  // marks made after this read are caught by the next epoch change
        __threadfence();
        if (__shfl_sync(0xffffffffu, (int)visited_live(visited, s), 0)) {
          R.status = kEnumSkipped;  // an earlier start's cycles cover this one: it will be skipped
        } else if (stale(__shfl_sync(0xffffffffu, R.n_keys, 0))) {
          if (lane == 0) speculate();
          ++reruns;
          __syncwarp();
          continue;
        }
      }
      const uint32_t dist = t - ci;
      __nanosleep(dist <= 2 ? 100u : (dist < 100 ? dist * 200u : 20000u));
    }
    stopped = __shfl_sync(0xffffffffu, stopped, 0);
    if (stopped) break;
    __threadfence();  // everything committed before this turn is visible from here on
    // ---- ordered commit: the bitmap is now exactly the one of position t in the sequential loop
    int status = __shfl_sync(0xffffffffu, R.status, 0);
    bool skip = __shfl_sync(0xffffffffu, (int)visited_live(visited, s), 0) != 0;  // :476
    if (!skip && status == kEnumSkipped) {
      // cannot happen: a search is only abandoned after its start was marked; kept as a safe path
      if (lane == 0) speculate();
      __syncwarp();
      status = __shfl_sync(0xffffffffu, R.status, 0);
    }
    if (!skip) {
      for (;;) {
        if (status == kEnumOverflow) { stopped = 1; break; }
        if (!stale(__shfl_sync(0xffffffffu, R.n_keys, 0))) break;
        if (lane == 0) {  // exclusive repeat against the final bitmap
          ++searches;
          enum_run(g, adj, visited, s, min_len, max_len, A, R, []() { return false; });
        }
        ++reruns;
        __syncwarp();
        status = __shfl_sync(0xffffffffu, R.status, 0);
      }
    }
    TicketRec rec = {kFateSkipped, status, 0, 0, 0, 0, 0, 0, us_searched, 0, 0, 0, 0, 0, 0, 0, 0};
    if (!skip && !stopped) {
      const uint32_t n_nodes = __shfl_sync(0xffffffffu, R.n_nodes, 0);
      const uint32_t n_cycles = __shfl_sync(0xffffffffu, R.n_cycles, 0);
      const uint32_t node_base = ld_vol(&ctl->pool_nodes), cyc_base = ld_vol(&ctl->pool_cycles);
      if ((uint64_t)node_base + n_nodes > pool_nodes_cap || (uint64_t)cyc_base + n_cycles > pool_len_cap) {
        stopped = 2;
      } else {
        for (uint32_t j = lane; j < n_nodes; j += 32) {  // :217-222 + results[start] = cycles (:479)
          uint32_t x = A.cyc_nodes[j];
          pool_nodes[node_base + j] = x;
          atomicOr(&visited[x >> 5], 1u << (x & 31));
        }
        for (uint32_t j = lane; j < n_cycles; j += 32) pool_len[cyc_base + j] = A.cyc_off[j + 1] - A.cyc_off[j];
        rec.fate = kFateAccepted;
        rec.n_cycles = n_cycles;
        rec.n_nodes = n_nodes;
        rec.cyc_base = cyc_base;
        rec.node_base = node_base;
        rec.expansions = __shfl_sync(0xffffffffu, R.expansions, 0);
        rec.steps = __shfl_sync(0xffffffffu, R.steps, 0);
        if (lane == 0) {
          ctl->pool_nodes = node_base + n_nodes;
          ctl->pool_cycles = cyc_base + n_cycles;
        }
      }
    }
    if (stopped) {  // this ticket needs the host (larger arena / larger pool): drain the pipeline
      if (lane == 0) atomicExch(&ctl->stop, (uint32_t)stopped);
      break;
    }
    if (lane == 0) {
      rec.us_committed = (uint32_t)((gtime_ns() - t_kernel) / 1000);
      recs[t] = rec;
    }
    __threadfence();
    __syncwarp();
    if (lane == 0) {
      if (rec.n_nodes) atomicAdd(&ctl->mark_epoch, 1u);  // after the fence: the marks are visible first
      atomicExch(&ctl->commit_idx, t + 1);               // hand over to the next ticket
    }
  }
  if (lane == 0) {
    if (searches) atomicAdd(&ctl->searches_run, searches);
    if (abandoned) atomicAdd(&ctl->n_abandoned, abandoned);
    if (reruns) atomicAdd(&ctl->n_reruns, reruns);
  }
}

// ---- the pipeline proper: extracted-ball searches (ball_dev.cuh; worker arenas: ball_layout.cuh)
__global__ void __launch_bounds__(32 * kWarpsPerBlock, 1) k_enum_pipeline(GraphView g, const uint64_t *__restrict__ adj, uint32_t *visited,
                                                     const uint32_t *__restrict__ starts, uint32_t n_end, int min_len,
                                                     int max_len, uint8_t *arena, BallLayout L, PipeCtl *ctl,
                                                     TicketRec *recs, uint32_t *pool_nodes, uint32_t pool_nodes_cap,
                                                     uint32_t *pool_len, uint32_t pool_len_cap, uint32_t *owner,
                                                     uint32_t *built, uint32_t *done, uint32_t *go, uint32_t *parent, int flags) {
  extern __shared__ __align__(16) uint8_t smem_all[];
  constexpr unsigned kFull = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t worker = blockIdx.x * kWarpsPerBlock + warp;
  uint8_t *const my_smem = smem_all + (size_t)warp * L.smem_bytes;
  const BallArena A = ball_arena_of(arena + (size_t)worker * L.bytes, my_smem, L);
  uint32_t *cyc_keys = (uint32_t *)(arena + (size_t)worker * L.bytes + L.off_cyc_keys);  // staged cycles as item ids
  // records of the contracted ball: shared memory, or the arena when there are more than L.chain_cap chains
  SBall Bs, Bg;
  Bs.hot = (SHot *)(my_smem + L.smem_lvl + L.smem_relax);
  Bs.in4 = (uint64_t *)(my_smem + L.smem_lvl + L.smem_relax + (size_t)L.chain_cap * sizeof(SHot));
  Bs.cap = L.chain_cap;
  Bg.hot = (SHot *)(arena + (size_t)worker * L.bytes + L.off_shot);
  Bg.in4 = (uint64_t *)(arena + (size_t)worker * L.bytes + L.off_sin);
  Bg.cap = L.node_cap + 1;
  Bs.n = Bg.n = 0;
  Coop co;
  co.lane = lane;
  unsigned long long searches = 0;
  uint32_t abandoned = 0, reruns = 0, n_ooo = 0;
  uint32_t live = 0;  // ball nodes whose index slots are still occupied
  const unsigned long long t_kernel = gtime_ns();
  const uint32_t poll_ns = (uint32_t)flags >> 8 ? (uint32_t)flags >> 8 : 300u;  // pause between two looks at the watched tickets
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(&ctl->ticket, 1u);
    t = __shfl_sync(kFull, t, 0);
    if (t >= n_end || __shfl_sync(kFull, ld_vol(&ctl->stop), 0)) break;
    if (__shfl_sync(kFull, ld_vol(&done[t]), 0)) {  // committed by an earlier launch of this call
      if (__shfl_sync(kFull, ld_vol(&ctl->commit_idx), 0) == t) sweep_front(&ctl->commit_idx, done, go, t + 1, n_end, lane);
      continue;
    }
    const uint32_t s = starts[t];
    EnumResult R;
    R.status = kEnumSkipped;
    R.n_cycles = R.n_nodes = R.n_keys = R.steps = R.expansions = 0;
    int stopped = 0;
    uint32_t us_build = 0, t_iters = 0, t_relax = 0, us_started = 0, t_searches = 0, n_chains = 0;
    uint32_t us_claims = 0, us_turn = 0, t_root = t, n_watched = 0;
    bool claimed = false;  // this ticket's first ball has been claimed and "built" published
    auto publish_built = [&]() {
      if (claimed) return;
      claimed = true;
      __syncwarp();
      if (lane == 0) {
        __threadfence();
        atomicExch(&built[t], 1u);
        fence_sc();
        advance_front(&ctl->built_idx, built, n_end);
      }
    };
    // the ticket is finished (committed or skipped): record, flag, move the commit front
    auto finish = [&](const TicketRec &rec) {
      __syncwarp();
      uint32_t front = 0;
      if (lane == 0) {
        __threadfence();
        if (rec.n_nodes) atomicAdd(&ctl->mark_epoch, 1u);  // after the fence: the marks are visible first
        atomicExch(&done[t], 1u);
        fence_sc();
        front = ld_vol(&ctl->commit_idx);
      }
      if (__shfl_sync(kFull, front, 0) == t) sweep_front(&ctl->commit_idx, done, go, t + 1, n_end, lane);
      if (lane == 0) recs[t] = rec;  // read by the host after the kernel: off the hand-over path
    };
    // does the search rest on a node that has been marked since its ball was built? (warp-cooperative)
    auto stale = [&](uint32_t n_keys) {
      int conflict = 0;
      for (uint32_t j0 = 1 + lane; j0 < n_keys && !conflict; j0 += 128) {  // four nodes per lane: keys, then visited words, in flight
        uint32_t k[4], w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) k[q] = j0 + 32u * q < n_keys ? A.nodes[j0 + 32u * q].key : s;
#pragma unroll
        for (int q = 0; q < 4; ++q) w[q] = k[q] != s ? ld_cg32(&visited[k[q] >> 5]) : 0u;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (k[q] != s && ((w[q] >> (k[q] & 31)) & 1u)) conflict = 1;
      }
      return __any_sync(kFull, conflict) != 0;
    };
    // Speculative search against the live bitmap (all lanes call it): the warp extracts the ball (and, the first
    // time, claims it), lane 0 searches its contraction in slices of kSliceSteps chain steps.  Between slices the
    // warp polls: pipeline draining -> stop; start marked by an earlier commit -> the ticket is skipped (:476),
    // give up; the mark epoch moved and a ball node is marked now -> the search is stale, start over at once
    // instead of finishing a search whose result would be thrown away.  `exclusive` (at the ticket's turn: the
    // bitmap is final) runs to the end without polling.
    uint32_t epoch_checked = 0xFFFFFFFFu;  // `visited` as of this epoch is accounted for in R
    auto speculate = [&](bool exclusive) {
      for (;;) {
        ++searches;
        ++t_searches;
        epoch_checked = __shfl_sync(kFull, ld_vol(&ctl->mark_epoch), 0);
        const unsigned long long t_build = gtime_ns();
        us_started = (uint32_t)((t_build - t_kernel) / 1000);
        ball_clear(A, live, co);
        const int st = ball_build(g, adj, visited, s, max_len, A, live, co);
        R.status = st;
        R.n_cycles = R.n_nodes = R.steps = R.expansions = 0;
        R.n_keys = live;
        if (!claimed) {
          // claim the ball: owner[x] = smallest ticket whose ball holds x.  A ball that did not fit is claimed only in
          // part, so no later ticket may conclude that it is independent of this one: they all commit in order.
          if (st != kEnumDone) {
            if (lane == 0) atomicMin(&ctl->ooo_limit, t);
          } else {
            uint32_t last = t;
            for (uint32_t j = lane; j < live; j += 32) {
              const uint32_t old = atomicMin(&owner[A.nodes[j].key], t);
              if (old != 0xFFFFFFFFu && old != last) {  // another ticket's ball holds the node too: same group
                uf_union(parent, t, old);
                last = old;
              }
            }
          }
          publish_built();
        }
        us_build = (uint32_t)((gtime_ns() - t_build) / 1000);
        if (st != kEnumDone) break;
        // contract the ball (unary chains become single records) and search it
        bool in_smem = !(flags & 1);
        if (in_smem) in_smem = sball_contract(A, live, max_len, Bs, co) == kEnumDone;
        if (!in_smem && sball_contract(A, live, max_len, Bg, co) != kEnumDone) {
          R.status = kEnumOverflow;  // 65535 chains: left to the hash-map search
          break;
        }
        const SBall &B = in_smem ? Bs : Bg;
        n_chains = B.n;
        SDfsState ds;
        bool restart = false;
        for (;;) {
          int fin = 0;
          if (lane == 0) {
            const int budget = exclusive ? 0x7fffffff : kSliceSteps;
            if (in_smem && L.smem_lvl) {
              const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(my_smem);
              const RecsSmem Q{sbase + L.smem_lvl + L.smem_relax, sbase + L.smem_lvl + L.smem_relax + L.chain_cap * 16u, sbase,
                               sbase + L.smem_lvl};
              fin = sball_dfs(min_len, max_len, A, Q, ds, budget, R) ? 1 : 0;
            } else {
              const RecsPtr Q{B.hot, B.in4, A.lvl, A.relax_top};
              fin = sball_dfs(min_len, max_len, A, Q, ds, budget, R) ? 1 : 0;
            }
          }
          if (__shfl_sync(kFull, fin, 0)) break;
          uint32_t c_stop = 0, c_vis = 0, c_ep = 0;
          if (lane == 0) {
            c_stop = ld_vol(&ctl->stop);
            c_vis = visited_live(visited, s);
            c_ep = ld_vol(&ctl->mark_epoch);
          }
          c_stop = __shfl_sync(kFull, c_stop, 0);
          c_vis = __shfl_sync(kFull, c_vis, 0);
          c_ep = __shfl_sync(kFull, c_ep, 0);
          if (c_stop | c_vis) {
            if (c_stop) stopped = 1;
            else ++abandoned;
            R.status = kEnumSkipped;
            R.n_cycles = R.n_nodes = 0;
            ds.n_staged = 0;
            break;
          }
          if (c_ep != epoch_checked) {
            epoch_checked = c_ep;
            __threadfence();
            if (stale(live)) { restart = true; break; }
          }
        }
        if (restart) { ++reruns; continue; }
        __syncwarp();  // lane 0's writes (staged cycles) are visible to the warp
        // item ids of the staged cycles, so that the commit only has to stream them
        const int fst = __shfl_sync(kFull, R.status, 0);
        const uint32_t n_staged = __shfl_sync(kFull, ds.n_staged, 0);
        t_iters = __shfl_sync(kFull, ds.iters, 0);
        t_relax = __shfl_sync(kFull, ds.relax_pops, 0);
        if (fst == kEnumDone || fst == kEnumCap) sball_expand(A, B, n_staged, cyc_keys, co);
        break;
      }
      __syncwarp();
    };
    TicketRec rec = {kFateSkipped, kEnumSkipped, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    // a ticket whose start is already marked is skipped for good (:476)
    if (__shfl_sync(kFull, (int)visited_live(visited, s), 0)) {
      publish_built();
      finish(rec);
      continue;
    }
    speculate(false);
    const uint32_t us_searched = (uint32_t)((gtime_ns() - t_kernel) / 1000);
    // marks the cycles' nodes (:217-222), reserves the output range and writes results[start] = cycles (:479);
    // returns false when the output pools are full (the pipeline drains and the host enlarges them)
    auto commit = [&](int status) -> bool {
      const uint32_t n_nodes = __shfl_sync(kFull, R.n_nodes, 0), n_cycles = __shfl_sync(kFull, R.n_cycles, 0);
      uint32_t node_base = 0, cyc_base = 0;
      int full = 0;
      if (lane == 0 && (n_nodes | n_cycles)) {  // (most commits of a dependent chain carry no cycle: no round trip then)
        node_base = atomicAdd(&ctl->pool_nodes, n_nodes);
        cyc_base = atomicAdd(&ctl->pool_cycles, n_cycles);
        if ((uint64_t)node_base + n_nodes > pool_nodes_cap || (uint64_t)cyc_base + n_cycles > pool_len_cap) {
          atomicSub(&ctl->pool_nodes, n_nodes);
          atomicSub(&ctl->pool_cycles, n_cycles);
          full = 1;
        }
      }
      if (__shfl_sync(kFull, full, 0)) return false;
      node_base = __shfl_sync(kFull, node_base, 0);
      cyc_base = __shfl_sync(kFull, cyc_base, 0);
      for (uint32_t j = lane; j < n_nodes; j += 128) {  // four independent loads in flight per lane
        uint32_t x[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) x[q] = j + 32u * q < n_nodes ? cyc_keys[j + 32u * q] : 0xFFFFFFFFu;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (x[q] != 0xFFFFFFFFu) atomicOr(&visited[x[q] >> 5], 1u << (x[q] & 31));
      }
      rec.fate = kFateAccepted;
      rec.status = status;
      rec.n_cycles = n_cycles;
      rec.n_nodes = n_nodes;
      rec.cyc_base = cyc_base;
      rec.node_base = node_base;
      rec.expansions = __shfl_sync(kFull, R.expansions, 0);
      rec.steps = __shfl_sync(kFull, R.steps, 0);
      rec.us_searched = us_searched;
      rec.us_committed = (uint32_t)((gtime_ns() - t_kernel) / 1000);
      rec.n_ball = live;
      rec.us_build = us_build;
      rec.n_chains = n_chains;
      rec.iters = t_iters;
      rec.relax_pops = t_relax;
      rec.us_started = us_started;
      rec.n_restarts = t_searches;
      rec.us_claims = us_claims;
      rec.us_turn = us_turn;
      rec.root = t_root;
      rec.n_watched = n_watched;
      return true;
    };
    // results[start] = cycles (:479): off the hand-over path, after the ticket has been flagged done
    auto copy_out = [&]() {
      for (uint32_t j = lane; j < rec.n_nodes; j += 32) pool_nodes[rec.node_base + j] = cyc_keys[j];
      for (uint32_t j = lane; j < rec.n_cycles; j += 32) pool_len[rec.cyc_base + j] = A.cyc_noff[j + 1] - A.cyc_noff[j];
    };
    stopped = __shfl_sync(kFull, stopped, 0);
    if (stopped) break;
    int status = __shfl_sync(kFull, R.status, 0);
    if (status == kEnumSkipped) {  // the start got marked while the search ran
      finish(rec);
      continue;
    }
    // ---- every earlier ticket has claimed its ball?  Then the owners of this ball's nodes tell whether one of them
    //      overlaps it.
    for (;;) {
      uint32_t bi = 0, st = 0;
      if (lane == 0) {
        bi = ld_vol(&ctl->built_idx);
        st = ld_vol(&ctl->stop);
      }
      bi = __shfl_sync(kFull, bi, 0);
      st = __shfl_sync(kFull, st, 0);
      if (st) { stopped = 1; break; }
      if (bi > t) break;
      __nanosleep(200u);
    }
    if (stopped) break;
    us_turn = us_claims = (uint32_t)((gtime_ns() - t_kernel) / 1000);
    bool independent = false;
    if (status != kEnumOverflow && !(flags & 4)) {
      __threadfence();
      uint32_t mn = 0xFFFFFFFFu;
      for (uint32_t j = lane; j < live; j += 32) mn = min(mn, ld_cg32(&owner[A.nodes[j].key]));
      for (int d = 16; d; d >>= 1) mn = min(mn, __shfl_xor_sync(kFull, mn, d));
      independent = mn >= t && t < __shfl_sync(kFull, ld_vol(&ctl->ooo_limit), 0);
    }
    if (independent) {
      // no earlier ticket can read or mark a node of this ball and later ones wait for this commit: the result is
      // the one of the sequential order
      if (!commit(status)) {
        if (lane == 0) atomicExch(&ctl->stop, 2u);
        break;
      }
      ++n_ooo;
      finish(rec);
      copy_out();
      continue;
    }
    // ---- dependent.  Its turn has come when every earlier ticket of its GROUP is done (tickets of other groups cannot
    //      touch its ball); the earlier tickets that are not done yet are few (the window in flight): each lane watches one.
    //      More than 32 of them, or a ball somewhere that was claimed only in part: the turn is the commit front's.
    uint32_t watch = 0xFFFFFFFFu;
    bool more = false;  // the group has further unfinished members below the watched ones
    const bool grouped = status != kEnumOverflow && !(flags & 4) && t < __shfl_sync(kFull, ld_vol(&ctl->ooo_limit), 0);
    // the (up to 32) largest unfinished tickets of this ticket's group below `t`, one per lane; windows of 32 tickets from
    // t downwards, a window is taken whole or left for the next batch
    auto collect_watch = [&]() {
      const uint32_t root = __shfl_sync(kFull, lane == 0 ? uf_find(parent, t) : 0u, 0);
      const uint32_t front = __shfl_sync(kFull, ld_vol(&ctl->commit_idx), 0);
      uint32_t count = 0;
      watch = 0xFFFFFFFFu;
      more = false;
      t_root = root;
      for (uint32_t hi = t; hi > front;) {
        const uint32_t base = hi - front >= 32u ? hi - 32u : front;
        const uint32_t u = base + (uint32_t)lane;
        const bool match = u < hi && !ld_vol(&done[u]) && uf_find(parent, u) == root;
        const unsigned m = __ballot_sync(kFull, match);
        const uint32_t c = (uint32_t)__popc(m);
        if (count + c > 32) { more = true; break; }
        if ((uint32_t)lane >= count && (uint32_t)lane < count + c) watch = base + __fns(m, 0, (int)((uint32_t)lane - count) + 1);
        count += c;
        hi = base;
      }
      return count;
    };
    if (grouped) n_watched = collect_watch();
    else n_watched = 0xFFFFFFFFu;
    // ---- meanwhile keep the speculative result fresh, so that the
    bool skipped_now = false;
    for (uint32_t polls = 0;; ++polls) {
      if (grouped) {  // every watched ticket done?
        const bool pending = watch != 0xFFFFFFFFu && !ld_vol(&done[watch]);
        if (!pending) watch = 0xFFFFFFFFu;
        if (!__any_sync(kFull, pending)) {
          if (!more) break;
          collect_watch();  // the next batch of earlier members (usually none is left unfinished: they commit in order)
          continue;
        }
      } else if (__shfl_sync(kFull, ld_vol(&go[t]), 0)) {  // own word: set by whoever moves the commit front onto this ticket
        break;
      }
      if ((polls & 15u) == 0) {  // the shared control words, less often
        uint32_t ep = 0;
        int st = stopped;
        if (lane == 0) {
          ep = ld_vol(&ctl->mark_epoch);
          st |= (int)ld_vol(&ctl->stop);
        }
        ep = __shfl_sync(kFull, ep, 0);
        st = __shfl_sync(kFull, st, 0);
        if (st) { stopped = 1; break; }
        status = __shfl_sync(kFull, R.status, 0);
        if (ep != epoch_checked && status != kEnumOverflow) {
          epoch_checked = ep;  // marks made after this read are caught by the next epoch change
          __threadfence();
          if (__shfl_sync(kFull, (int)visited_live(visited, s), 0)) {
            skipped_now = true;  // an earlier start's cycles cover this one: skipped (:476)
            break;
          }
          if (stale(__shfl_sync(kFull, R.n_keys, 0))) {
            speculate(false);
            ++reruns;
            if (__shfl_sync(kFull, stopped, 0)) { stopped = 1; break; }
            if (__shfl_sync(kFull, R.status, 0) == kEnumSkipped) { skipped_now = true; break; }
            continue;
          }
        }
      }
      __nanosleep(poll_ns);
    }
    stopped = __shfl_sync(kFull, stopped, 0);
    if (stopped) break;
    if (skipped_now) {
      finish(rec);
      continue;
    }
    us_turn = (uint32_t)((gtime_ns() - t_kernel) / 1000);
    __threadfence();  // everything committed before this turn is visible from here on
    // ---- ordered commit: the bitmap is now exactly the one of position t in the sequential loop
    status = __shfl_sync(kFull, R.status, 0);
    const bool skip = __shfl_sync(kFull, (int)visited_live(visited, s), 0) != 0;  // :476
    if (!skip) {
      // Nothing has been marked since the result was last examined (same epoch) -> it is exact.  Else
      // re-examine it; a stale result is repeated exclusively.
      const uint32_t ep_now = __shfl_sync(kFull, ld_vol(&ctl->mark_epoch), 0);
      while (status != kEnumOverflow &&
             (status == kEnumSkipped || (ep_now != epoch_checked && stale(__shfl_sync(kFull, R.n_keys, 0))))) {
        speculate(true);  // exclusive repeat against the final bitmap (sets epoch_checked = ep_now)
        ++reruns;
        status = __shfl_sync(kFull, R.status, 0);
      }
      if (status == kEnumOverflow) stopped = 1;
      else if (!commit(status)) stopped = 2;
    }
    if (stopped) {  // this ticket needs the host (hash-map search in a large arena / larger pool): drain
      if (lane == 0) atomicExch(&ctl->stop, (uint32_t)stopped);
      break;
    }
    finish(rec);
    copy_out();
  }
  __syncwarp();
  ball_clear(A, live, co);  // the index is left empty for the next launch
  if (lane == 0) {
    if (searches) atomicAdd(&ctl->searches_run, searches);
    if (abandoned) atomicAdd(&ctl->n_abandoned, abandoned);
    if (reruns) atomicAdd(&ctl->n_reruns, reruns);
    if (n_ooo) atomicAdd(&ctl->n_ooo, n_ooo);
  }
}

static uint32_t env_u32(const char *name, uint32_t def, uint32_t lo, uint32_t hi) {
  const char *e = getenv(name);
  if (!e || !*e) return def;
  long v = strtol(e, nullptr, 10);
  if (v < (long)lo) v = lo;
  if (v > (long)hi) v = hi;
  return (uint32_t)v;
}

static int grow_copy(DevBuf &buf, size_t new_bytes, size_t keep_bytes, cudaStream_t s) {
  DevBuf nb;
  MG_TRY(nb.alloc(new_bytes));
  if (keep_bytes) MG_CUDA(cudaMemcpyAsync(nb.p, buf.p, keep_bytes, cudaMemcpyDeviceToDevice, s));
  MG_CUDA(cudaStreamSynchronize(s));
  std::swap(buf.p, nb.p);
  std::swap(buf.bytes, nb.bytes);
  return MG_OK;
}

int run_find_cycles(mg_graph *g, const uint64_t *starts, uint64_t n, int min_len, int max_len, mg_cycles *out) {
  cudaStream_t s = g->stream;
  memset(out, 0, sizeof *out);
  const bool host_timing = getenv("MCAAT_ENUM_HOST_TIMING") != nullptr;  // diagnostic: where the host side of the call spends its time
  auto now_ms = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_enter = now_ms();
  double t_launched = 0, t_kernels = 0, t_read = 0;
  if (!g->pruned || !g->adj.p) {
    set_error("mg_find_cycles: call mg_prune first (it builds the navigation cache)");
    return MG_EINVAL;
  }
  if (max_len < 1 || max_len > 4000 || min_len < 0) {
    set_error("mg_find_cycles: cycle length bounds (%d, %d] outside the supported range", min_len, max_len);
    return MG_ELIMIT;
  }
  if (n >= 0xFFFFFFF0ull) return MG_ELIMIT;
  std::vector<uint32_t> s32(n);
  for (uint64_t i = 0; i < n; ++i) {
    if (starts[i] >= g->view.n) {
      set_error("mg_find_cycles: start id %llu out of range", (unsigned long long)starts[i]);
      return MG_EINVAL;
    }
    s32[i] = (uint32_t)starts[i];
  }
  // results come back through page-locked memory kept in the handle (pageable read-back of the ~4 MB at C2 cost 0.6 ms)
  TicketRec *recs = nullptr;
  const uint32_t *h_nodes = nullptr, *h_len = nullptr;
  size_t n_h_nodes = 0, n_h_len = 0;
  PipeCtl ctl;
  memset(&ctl, 0, sizeof ctl);
  if (n) {
    DevBuf starts_dev, ctl_dev, recs_dev, pool_nodes, pool_len, big_arena, owner, flags_dev;
    MG_TRY(starts_dev.alloc(n * 4));
    MG_CUDA(cudaMemcpyAsync(starts_dev.p, s32.data(), n * 4, cudaMemcpyHostToDevice, s));
    MG_TRY(ctl_dev.alloc(sizeof(PipeCtl)));
    MG_TRY(recs_dev.alloc(n * sizeof(TicketRec)));
    // ball claims (smallest ticket whose ball holds the item) and the per-ticket built / done flags
    MG_TRY(owner.alloc((size_t)g->view.n * 4));
    MG_CUDA(cudaMemsetAsync(owner.p, 0xFF, (size_t)g->view.n * 4, s));
    MG_TRY(flags_dev.alloc((n + 1) * 12));
    MG_CUDA(cudaMemsetAsync(flags_dev.p, 0, (n + 1) * 12, s));
    uint32_t *built_dev = flags_dev.as<uint32_t>(), *done_dev = built_dev + n + 1, *go_dev = done_dev + n + 1;
    DevBuf parent_dev;  // union-find over tickets: groups of start nodes whose balls overlap
    MG_TRY(parent_dev.alloc((n + 1) * 4));
    {
      std::vector<uint32_t> iota(n + 1);
      for (uint64_t i = 0; i <= n; ++i) iota[i] = (uint32_t)i;
      MG_CUDA(cudaMemcpyAsync(parent_dev.p, iota.data(), (n + 1) * 4, cudaMemcpyHostToDevice, s));
      MG_CUDA(cudaStreamSynchronize(s));
    }
    ctl.ooo_limit = 0xFFFFFFFFu;
    uint32_t nodes_cap = 1u << 22, len_cap = 1u << 18;
    MG_TRY(pool_nodes.alloc((size_t)nodes_cap * 4));
    MG_TRY(pool_len.alloc((size_t)len_cap * 4));
    // worker arenas live in the graph handle (they need no clearing between searches or calls)
    const uint32_t want =
        (uint32_t)std::min<uint64_t>(n, (uint64_t)g->sm_count * env_u32("MCAAT_ENUM_WORKERS_PER_SM", 4, 1, 32));  // see below
    const uint32_t blocks = (want + kWarpsPerBlock - 1) / kWarpsPerBlock, workers = blocks * kWarpsPerBlock;
    // shared memory: the workers of one SM share what the device offers (227 KB opt-in, 1 KB reserved per block)
    const uint32_t workers_per_sm = env_u32("MCAAT_ENUM_WORKERS_PER_SM", 4, 1, 32);
    const uint32_t smem_per_worker = std::min<uint32_t>(ball_smem_per_worker(g, workers_per_sm, kWarpsPerBlock),
                                                        env_u32("MCAAT_ENUM_SMEM_PER_WORKER", 1u << 20, 2048, 1u << 20));
    const BallLayout L = make_ball_layout(max_len, env_u32("MCAAT_ENUM_BALL_LOG", 16, 5, 16), smem_per_worker);
    // diagnostics: chain records in the arena instead of shared memory (1); every commit in ticket order (4)
    const int force_global = (int)(env_u32("MCAAT_ENUM_CHAINS_GLOBAL", 0, 0, 1) | (env_u32("MCAAT_ENUM_INORDER", 0, 0, 1) << 2) |
                                 (env_u32("MCAAT_ENUM_POLL_NS", 0, 0, 100000) << 8));
    MG_CUDA(cudaFuncSetAttribute(k_enum_pipeline, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(L.smem_bytes * kWarpsPerBlock)));
    MG_TRY(ensure_ball_arena(g, workers, L));
    uint32_t big_lock_log = env_u32("MCAAT_ENUM_LOCK_LOG", 19, 6, 26), big_relax = 1u << 20;
    uint32_t pos = 0;
    bool solo = false;  // the ticket at `pos` overflowed a worker arena: run it alone in the big arena
    t_launched = now_ms();
    while (pos < n) {
      ctl.ticket = pos;
      ctl.commit_idx = pos;
      ctl.built_idx = pos;
      ctl.stop = 0;
      MG_CUDA(cudaMemcpyAsync(ctl_dev.p, &ctl, sizeof ctl, cudaMemcpyHostToDevice, s));
      MG_CUDA(cudaMemsetAsync(go_dev + pos, 1, 4, s));  // the commit front stands on ticket `pos`
      {
        TimedScope ts(g, T_ENUM, solo ? 1 : n - pos);
        if (!solo) {
          MG_LAUNCH(k_enum_pipeline, blocks, 32 * kWarpsPerBlock, L.smem_bytes * kWarpsPerBlock, s, g->view, g->adj.as<uint64_t>(), g->visited.as<uint32_t>(),
                    starts_dev.as<uint32_t>(), (uint32_t)n, min_len, max_len, g->enum_arena.as<uint8_t>(), L,
                    ctl_dev.as<PipeCtl>(), recs_dev.as<TicketRec>(), pool_nodes.as<uint32_t>(), nodes_cap,
                    pool_len.as<uint32_t>(), len_cap, owner.as<uint32_t>(), built_dev, done_dev, go_dev, parent_dev.as<uint32_t>(), force_global);
        } else {
          const SlotLayout LB = make_layout(max_len, big_lock_log, big_relax, kClusterBound);
          MG_CUDA(cudaFuncSetAttribute(k_enum_solo, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LB.smem_bytes));
          if (big_arena.bytes < LB.bytes) {
            MG_TRY(big_arena.alloc(LB.bytes));
          }
          MG_CUDA(cudaMemsetAsync(big_arena.p, 0, LB.bytes, s));
          MG_LAUNCH(k_enum_solo, 1, 32, LB.smem_bytes, s, g->view, g->adj.as<uint64_t>(), g->visited.as<uint32_t>(),
                    starts_dev.as<uint32_t>(), pos + 1, min_len, max_len, big_arena.as<uint8_t>(), LB,
                    ctl_dev.as<PipeCtl>(), recs_dev.as<TicketRec>(), pool_nodes.as<uint32_t>(), nodes_cap,
                    pool_len.as<uint32_t>(), len_cap);
        }
      }
      MG_CUDA(cudaMemcpyAsync(&ctl, ctl_dev.p, sizeof ctl, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaStreamSynchronize(s));
      ++out->n_rounds;
      if (solo && ctl.commit_idx == pos + 1) {  // the ticket is done: flag it for the pipeline launches that follow
        MG_CUDA(cudaMemsetAsync(built_dev + pos, 1, 4, s));
        MG_CUDA(cudaMemsetAsync(done_dev + pos, 1, 4, s));
      }
      pos = ctl.commit_idx;
      if (ctl.stop == 2) {  // output pools full: double them and resume at `pos`
        if ((uint64_t)nodes_cap * 2 > 0xFFFFFFF0ull) {
          set_error("mg_find_cycles: result exceeds 2^32 node ids");
          return MG_ELIMIT;
        }
        MG_TRY(grow_copy(pool_nodes, (size_t)nodes_cap * 8, (size_t)ctl.pool_nodes * 4, s));
        MG_TRY(grow_copy(pool_len, (size_t)len_cap * 8, (size_t)ctl.pool_cycles * 4, s));
        nodes_cap *= 2;
        len_cap *= 2;
      } else if (ctl.stop == 1) {  // ticket `pos`: ball overflow -> hash-map search alone in a large arena
        if (solo) {
          if (big_lock_log >= 26) {
            set_error("mg_find_cycles: start %llu does not fit the largest search arena", (unsigned long long)starts[pos]);
            return MG_ELIMIT;
          }
          big_lock_log += 2;
          big_relax *= 4;
        }
        solo = true;
      } else {
        solo = false;
      }
    }
    t_kernels = now_ms();
    const size_t b_recs = (n * sizeof(TicketRec) + 255) & ~(size_t)255, b_nodes = ((size_t)ctl.pool_nodes * 4 + 255) & ~(size_t)255,
                 b_len = (size_t)ctl.pool_cycles * 4;
    pinned_spare_take(g->host_stage);
    MG_TRY(g->host_stage.ensure(b_recs + b_nodes + b_len + 256));
    uint8_t *hs = (uint8_t *)g->host_stage.p;
    recs = (TicketRec *)hs;
    h_nodes = (const uint32_t *)(hs + b_recs);
    h_len = (const uint32_t *)(hs + b_recs + b_nodes);
    n_h_nodes = ctl.pool_nodes;
    n_h_len = ctl.pool_cycles;
    MG_CUDA(cudaMemcpyAsync(recs, recs_dev.p, n * sizeof(TicketRec), cudaMemcpyDeviceToHost, s));
    if (ctl.pool_nodes) MG_CUDA(cudaMemcpyAsync((void *)h_nodes, pool_nodes.p, (size_t)ctl.pool_nodes * 4, cudaMemcpyDeviceToHost, s));
    if (ctl.pool_cycles) MG_CUDA(cudaMemcpyAsync((void *)h_len, pool_len.p, (size_t)ctl.pool_cycles * 4, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
  }
  t_read = now_ms();
  flush_timers(g);
  if (const char *trace = getenv("MCAAT_ENUM_TRACE")) {  // per-ticket timeline for profiling
    if (FILE *f = fopen(trace, "w")) {
      fprintf(f, "ticket,fate,status,steps,n_cycles,us_searched,us_committed,n_ball,us_build,n_chains,iters,relax_pops,us_started,n_searches,us_claims,us_turn,root,n_watched\n");
      for (uint64_t i = 0; i < n; ++i)
        fprintf(f, "%llu,%d,%d,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%d\n", (unsigned long long)i, recs[i].fate, recs[i].status, recs[i].steps,
                recs[i].n_cycles, recs[i].us_searched, recs[i].us_committed, recs[i].n_ball, recs[i].us_build, recs[i].n_chains,
                recs[i].iters, recs[i].relax_pops, recs[i].us_started, recs[i].n_restarts, recs[i].us_claims, recs[i].us_turn, recs[i].root,
                (int)recs[i].n_watched);
      fclose(f);
    }
  }
  // ---- assemble `results` in insertion order (:479; empty results are stored)
  uint64_t n_searched = 0;
  for (uint64_t i = 0; i < n; ++i) n_searched += recs[i].fate == kFateAccepted;
  out->searched_start = (uint64_t *)malloc((n_searched + 1) * 8);
  out->cyc_off = (uint64_t *)malloc((n_searched + 2) * 8);
  out->node_off = (uint64_t *)malloc((n_h_len + 2) * 8);
  out->nodes = (uint64_t *)malloc((n_h_nodes + 1) * 8);
  if (!out->searched_start || !out->cyc_off || !out->node_off || !out->nodes) return MG_ENOMEM;
  uint64_t si = 0, ci = 0, ni = 0;
  out->cyc_off[0] = 0;
  out->node_off[0] = 0;
  for (uint64_t i = 0; i < n; ++i) {
    const TicketRec &r = recs[i];
    if (r.fate != kFateAccepted) {
      ++out->n_skipped;
      continue;
    }
    out->searched_start[si] = starts[i];
    out->n_bound_hits += r.status == kEnumBound;
    out->n_cap_hits += r.status == kEnumCap;
    out->expansions += r.expansions;
    uint32_t np = r.node_base;
    for (uint32_t c = 0; c < r.n_cycles; ++c) {
      uint32_t len = h_len[r.cyc_base + c];
      for (uint32_t j = 0; j < len; ++j) out->nodes[ni++] = h_nodes[np + j];
      np += len;
      out->node_off[++ci] = ni;
    }
    out->cyc_off[++si] = ci;
  }
  out->n_searched = n_searched;
  out->n_cycles = ci;
  out->n_reruns = ctl.n_reruns;
  out->n_searches_run = ctl.searches_run;
  out->n_abandoned = ctl.n_abandoned;
  out->n_out_of_order = ctl.n_ooo;
  if (host_timing)
    fprintf(stderr, "[mg_find_cycles host ms] setup %.3f kernels+sync %.3f read-back %.3f assemble %.3f\n", t_launched - t_enter,
            t_kernels - t_launched, t_read - t_kernels, now_ms() - t_read);
  return MG_OK;
}

}  // namespace mg
