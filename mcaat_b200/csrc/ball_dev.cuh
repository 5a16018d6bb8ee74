// FindCycle (reference src/cycle_finder.cpp:140-243) on an EXTRACTED BALL.
//
// Everything a search from `start` can touch lies in the forward ball of radius M = cycle_max_length
// around it in the filtered graph of _BackgroundCheck (:40-52): a node is pushed only while
// path.size() < lock[n] <= M, so path positions — and with them BFS distances — stay below M; the
// relaxation (:191-210) only propagates through nodes whose lock value was lowered, i.e. nodes that
// have been on the path.  So the search runs in two stages:
//   1. ball_build — warp-cooperative level-synchronous BFS from `start`: 32 frontier nodes at a
//      time, one per lane, so the dependent loads of an expansion (navigation cache -> multiplicity
//      + visited words of the <= 4 neighbours) overlap 32-fold instead of forming one serial
//      chain.  Nodes get dense local ids in discovery order (hash index only here); each node's
//      filtered out-neighbour set (_GetOutgoings(node,set,mult), :58-72) is stored as local ids in
//      std::unordered_set iteration order; the filtered in-neighbour sets (_GetIncomings, :76-88) are the
//      transpose (in-neighbours outside the expanded ball can never carry a lowered lock, so they
//      never influence a result).
//   2. ball_dfs — the literal sequential search of :140-225 by one lane over the dense node
//      records: no hashing, no rank/select, one record load per examined neighbour.
// The ball is the examined-node set of the ordered commit (enumerate.cu).  Capacity overflow
// (more than 65535 ball nodes, > 4 in-neighbours, relax stack) is reported as kEnumOverflow and the
// caller repeats the search with enum_run (search_dev.cuh), which has no such limits.
#pragma once
#include "search_dev.cuh"

namespace mg {

// A ball node is split into the half every search step reads (16 bytes: two nodes per sector, so the
// working set of the search stays small enough for the L1) and the half only relaxations and the
// build need.
struct alignas(16) LNode {
  uint32_t meta;     // bits 0-15 lock value | 16-18 |out| | 20-22 |in| | 24 node is on the current path
  uint32_t key;      // graph item id
  union { uint16_t out[4]; uint64_t out4; };  // local ids (kNoLid = unused), iteration order of the std::unordered_set built at :58-72
};
struct alignas(16) LTail {
  union { uint16_t in[4]; uint64_t in4; };    // local ids, iteration order of the set built at :76-88
  uint32_t slot;     // index slot of `key` (for clearing)
  uint32_t pad;
};
static_assert(sizeof(LNode) == 16 && sizeof(LTail) == 16, "two nodes per sector");
constexpr uint32_t kMetaLock = 0xFFFFu, kMetaOnPath = 1u << 24;
constexpr int kMetaOutShift = 16, kMetaInShift = 20;
constexpr uint64_t kIndexEmpty = ~0ull;
constexpr uint32_t kLidPending = 0xFFFFFFFFu;

// One path level = path[i], backtrack_lengths[i], the remaining neighbour set stack[i] and the node's
// meta word (nobody else changes the meta of a node while it is on the path): 16 bytes, one shared-
// memory access per push or pop.
struct alignas(16) Lvl {
  uint64_t set;      // 4 x uint16 local ids in iteration order, unused slots = kNoLid
  uint32_t pmeta;
  uint16_t node;
  int16_t bl;
};
constexpr uint32_t kNoLid = 0xFFFFu;

struct BallArena {
  LNode *nodes;
  LTail *tails;
  uint32_t node_cap;    // <= 65535
  uint64_t *index;      // [1 << index_log]: key << 32 | local id; kIndexEmpty when free.  Always left empty.
  uint32_t index_log;
  uint32_t *relax;      // relax_stack entries (backtrack length << 16 | local id), :192-193
  uint32_t relax_cap;
  uint16_t *cyc;        // staged raw cycles as local ids
  uint32_t cyc_cap;
  uint32_t *cyc_off;    // [kClusterBound + 1]
  Lvl *lvl;             // [M + 2] per path level (shared memory on the device)
};

// cooperative execution policy: the device runs ball_build with the 32 lanes of a warp, the host
// emulation (tests/emu) with one
#if defined(__CUDA_ARCH__)
struct Coop {
  int lane;
  static constexpr int kLanes = 32;
  __device__ void sync() const { __syncwarp(); }
  __device__ uint32_t sum(uint32_t x) const {
    for (int d = 16; d; d >>= 1) x += __shfl_xor_sync(0xffffffffu, x, d);
    return x;
  }
  __device__ uint32_t exclusive_prefix(uint32_t x) const {
    uint32_t incl = x;
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += y;
    }
    return incl - x;
  }
  __device__ bool any(bool p) const { return __any_sync(0xffffffffu, p) != 0; }
};
MG_HD uint64_t index_cas(uint64_t *p, uint64_t expect, uint64_t val) {
  return (uint64_t)atomicCAS((unsigned long long *)p, (unsigned long long)expect, (unsigned long long)val);
}
MG_HD uint32_t meta_add(uint32_t *p, uint32_t v) { return atomicAdd(p, v); }
MG_HD void st_cg64(uint64_t *p, uint64_t v) { __stcg((unsigned long long *)p, (unsigned long long)v); }
#else
struct Coop {
  int lane = 0;
  static constexpr int kLanes = 1;
  void sync() const {}
  uint32_t sum(uint32_t x) const { return x; }
  uint32_t exclusive_prefix(uint32_t) const { return 0; }
  bool any(bool p) const { return p; }
};
MG_HD uint64_t index_cas(uint64_t *p, uint64_t expect, uint64_t val) {
  uint64_t old = *p;
  if (old == expect) *p = val;
  return old;
}
MG_HD uint32_t meta_add(uint32_t *p, uint32_t v) {
  uint32_t old = *p;
  *p = old + v;
  return old;
}
MG_HD void st_cg64(uint64_t *p, uint64_t v) { *p = v; }
#endif

// Leaves the hash index empty again (all lanes).  `n_nodes` = nodes of the previous build.
MG_HD void ball_clear(const BallArena &A, uint32_t n_nodes, const Coop &co) {
  for (uint32_t i = (uint32_t)co.lane; i < n_nodes; i += Coop::kLanes) st_cg64(&A.index[A.tails[i].slot], kIndexEmpty);
  co.sync();
}

// Stage 1.  Returns kEnumDone or kEnumOverflow; n_nodes = number of ball nodes (also on overflow, so
// that ball_clear can tidy up).  All lanes call it with identical arguments.
MG_HD_NOINLINE int ball_build(const GraphView &g, const uint64_t *adj, const uint32_t *visited, uint32_t start, int M,
                              const BallArena &A, uint32_t &n_nodes, const Coop &co) {
  const uint32_t mu = get_mult(g, start);
  const uint32_t imask = (1u << A.index_log) - 1;
  LNode *N = A.nodes;
  LTail *T = A.tails;
  if (co.lane == 0) {
    uint32_t h = hash_u32(start, A.index_log);
    st_cg64(&A.index[h], ((uint64_t)start << 32) | 0u);
    N[0].meta = (uint32_t)M;
    N[0].key = start;
    N[0].out4 = ~0ull;
    T[0].slot = h;
    A.relax[0] = 0;
  }
  co.sync();
  uint32_t total_nodes = 1, lo = 0, hi = 1;
  int status = kEnumDone;
  for (int depth = 0; depth < M && lo < hi && status == kEnumDone; ++depth) {
    for (uint32_t base = lo; base < hi; base += Coop::kLanes) {
      const uint32_t i = base + (uint32_t)co.lane;
      uint32_t okn[4], slot[4], cnt = 0, mine = 0;
      if (i < hi) {
        const uint32_t node = N[i].key;
        uint32_t nb[4], nset = 0;
        int deg = adj_outgoing(ld_ro(&adj[node]), nb);
        if (deg > 4) deg = 4;
        filtered_set(g, visited, node, mu, nb, deg, nset, okn);  // :58-72 incl. the set's iteration order
        cnt = nset;
        for (uint32_t j = 0; j < cnt; ++j) {  // find or claim the neighbour's index slot
          const uint32_t key = okn[j];
          uint32_t h = hash_u32(key, A.index_log);
          for (;;) {
            uint64_t cur = ld_cg64(&A.index[h]);
            if (cur == kIndexEmpty) {
              cur = index_cas(&A.index[h], kIndexEmpty, ((uint64_t)key << 32) | kLidPending);
              if (cur == kIndexEmpty) { mine |= 1u << j; break; }
            }
            if ((uint32_t)(cur >> 32) == key) break;
            h = (h + 1) & imask;
          }
          slot[j] = h;
        }
      }
      // dense local ids for the newly discovered nodes of this chunk
      const uint32_t n_mine = (uint32_t)popc32(mine);
      const uint32_t before = co.exclusive_prefix(n_mine), chunk_new = co.sum(n_mine);
      if (total_nodes + chunk_new > A.node_cap) {
        // no room: release the claimed slots (every lane has stopped probing) and give up
        co.sync();
        for (uint32_t j = 0; j < cnt; ++j)
          if ((mine >> j) & 1) st_cg64(&A.index[slot[j]], kIndexEmpty);
        co.sync();
        status = kEnumOverflow;
        break;
      }
      uint32_t lid = total_nodes + before;
      for (uint32_t j = 0; j < cnt; ++j)
        if ((mine >> j) & 1) {
          N[lid].meta = (uint32_t)M;  // lock default (try_emplace(n, cycle_max_length), :168), no edges yet
          N[lid].key = okn[j];
          N[lid].out4 = ~0ull;
          T[lid].slot = slot[j];
          A.relax[lid] = 0;  // in-degree counter of the transpose below (the relax stack is idle during the build)
          st_cg64(&A.index[slot[j]], ((uint64_t)okn[j] << 32) | lid);
          ++lid;
        }
      total_nodes += chunk_new;
      co.sync();
      if (i < hi) {
        uint32_t meta = N[i].meta;
        for (uint32_t j = 0; j < cnt; ++j) N[i].out[j] = (uint16_t)(uint32_t)ld_cg64(&A.index[slot[j]]);
        N[i].meta = meta | (cnt << kMetaOutShift);
      }
    }
    lo = hi;
    hi = total_nodes;
  }
  co.sync();
  n_nodes = total_nodes;
  if (status != kEnumDone) return status;
  // ---- in-neighbour sets = transpose of the out-neighbour sets.  The counters live outside the node
  //      records and are only touched by atomics / L2 loads, so that no L1 line of a record the
  //      search reads later goes stale.
  bool too_many = false;
  for (uint32_t u = (uint32_t)co.lane; u < total_nodes; u += Coop::kLanes) {
    const uint32_t nout = (N[u].meta >> kMetaOutShift) & 7u;
    for (uint32_t j = 0; j < nout; ++j) {
      const uint32_t x = N[u].out[j];
      const uint32_t pos = meta_add(&A.relax[x], 1u);
      if (pos < 4) T[x].in[pos] = (uint16_t)u;
      else too_many = true;
    }
  }
  co.sync();
  if (co.any(too_many)) return kEnumOverflow;
  // _GetIncomings inserts in IncomingEdges order (ascending id, sdbg.h:266-285); reproduce the
  // iteration order of the resulting std::unordered_set
  for (uint32_t x = (uint32_t)co.lane; x < total_nodes; x += Coop::kLanes) {
    const uint32_t nin = ld_cg32(&A.relax[x]);
    if (nin == 0) continue;
    N[x].meta |= nin << kMetaInShift;
    if (nin < 2) continue;
    uint32_t lid[4], key[4];
    for (uint32_t j = 0; j < nin; ++j) {
      lid[j] = T[x].in[j];
      key[j] = N[lid[j]].key;
    }
    for (uint32_t a = 1; a < nin; ++a)  // insertion sort by key
      for (uint32_t b = a; b > 0 && key[b - 1] > key[b]; --b) {
        uint32_t t = key[b]; key[b] = key[b - 1]; key[b - 1] = t;
        t = lid[b]; lid[b] = lid[b - 1]; lid[b - 1] = t;
      }
    USet s;
    s.n = 0;
    for (uint32_t j = 0; j < nin; ++j) uset_insert(s, key[j]);
    for (uint32_t j = 0; j < nin; ++j)
      for (uint32_t q = 0; q < nin; ++q)
        if (key[q] == s.e[j]) T[x].in[j] = (uint16_t)lid[q];
  }
  co.sync();
  return kEnumDone;
}

// first 16 bytes of a node record (meta, key, out-neighbour set) in one load
struct NodeHead {
  uint32_t meta;
  uint64_t out4;
};
MG_HD NodeHead ld_head(const LNode *n) {
  NodeHead h;
#if defined(__CUDA_ARCH__)
  const uint4 v = *reinterpret_cast<const uint4 *>(n);
  h.meta = v.x;
  h.out4 = (uint64_t)v.z | ((uint64_t)v.w << 32);
#else
  h.meta = n->meta;
  h.out4 = n->out4;
#endif
  return h;
}

// Stage 2 (one lane).  Literal FindCycleUtil + FindCycle, src/cycle_finder.cpp:231-243,140-225, on the
// ball; like enum_run (search_dev.cuh) it does not mark visited, and the cluster bound stops it at
// once (the answer is empty whatever follows).  Cycles are staged as local ids in A.cyc.
struct DfsState {  // the scalars of a paused search (everything else lives in the arena)
  int depth = 0;     // path.size() == stack.size() == backtrack_lengths.size()
  int counter = 0, steps = 0;
  uint32_t expansions = 0, n_cycles = 0, n_cyc = 0;
  bool started = false;
};
// Runs at most `budget` further steps.  Returns true when the search has ended (R is filled in) and
// false when it is merely paused: the caller polls (start marked? examined nodes marked? — see
// enumerate.cu) and calls again with the same state.
MG_HD_NOINLINE bool ball_dfs(int min_len, int max_len, const BallArena &A, DfsState &S, int budget, EnumResult &R) {
  const int M = max_len;
  LNode *N = A.nodes;
  LTail *T = A.tails;
  Lvl *L = A.lvl;
  uint32_t expansions = S.expansions, n_cycles = S.n_cycles, n_cyc = S.n_cyc;
  int plen = S.depth, counter = S.counter, steps = S.steps;
  int32_t status = kEnumDone;
  Lvl cur;  // the top level lives in registers; L[plen - 1] is written when it stops being the top
  if (!S.started) {
    S.started = true;
    A.cyc_off[0] = 0;
    const NodeHead h0 = ld_head(&N[0]);
    cur.pmeta = (h0.meta & ~kMetaLock) | kMetaOnPath;  // lock[start] = 0, :237
    N[0].meta = cur.pmeta;
    cur.node = 0;
    cur.set = h0.out4;                                 // _GetOutgoings(start), :238-240
    cur.bl = (int16_t)M;
    plen = 1;
    ++expansions;
  } else {
    cur = L[plen - 1];
  }
  const int pause_at = steps + budget;
  while (plen > 0) {
    if (steps == pause_at) {
      L[plen - 1] = cur;
      S.depth = plen; S.counter = counter; S.steps = steps;
      S.expansions = expansions; S.n_cycles = n_cycles; S.n_cyc = n_cyc;
      return false;
    }
    if (++steps > kMaxSteps) { status = kEnumCap; break; }  // :148-151
    const uint64_t set = cur.set;                           // :153 (copy)
    // records of all members first: the loads overlap, and no lock changes before the scan ends
    NodeHead hd[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t nb = (uint32_t)(set >> (16 * i)) & 0xFFFFu;
      if (nb != kNoLid) hd[i] = ld_head(&N[nb]);
    }
    bool flag = true;
#pragma unroll
    for (int i = 0; i < 4; ++i) {                           // static indices: hd[] stays in registers
      const uint32_t nb = (uint32_t)(set >> (16 * i)) & 0xFFFFu;
      if (nb != kNoLid && flag) {
        if (nb == 0) {                                      // n == start_node, :157-166
          cur.bl = 1;
          if (plen > min_len) {
            if (++counter >= kClusterBound) {               // cluster bound: final answer is empty
              status = kEnumBound;
              n_cycles = n_cyc = 0;
              goto done;
            }
            if (n_cyc + (uint32_t)plen > A.cyc_cap) { status = kEnumOverflow; goto done; }
            for (int j = 0; j + 1 < plen; ++j) A.cyc[n_cyc + j] = L[j].node;
            A.cyc[n_cyc + plen - 1] = cur.node;
            n_cyc += (uint32_t)plen;
            A.cyc_off[++n_cycles] = n_cyc;
          }
        } else if (plen < (int)(hd[i].meta & kMetaLock)) {  // :168
          {                                                 // stack.back().erase(n), :173
            const uint64_t low = i ? set & ((1ull << (16 * i)) - 1) : 0ull;
            const uint64_t high = i < 3 ? (set >> (16 * (i + 1))) << (16 * i) : 0ull;
            cur.set = low | high | (0xFFFFull << 48);
          }
          L[plen - 1] = cur;
          ++plen;
          cur.set = hd[i].out4;                             // :174-176
          cur.pmeta = (hd[i].meta & ~kMetaLock) | (uint32_t)plen | kMetaOnPath;  // :172
          cur.node = (uint16_t)nb;                          // :170
          cur.bl = (int16_t)M;                              // :171
          N[nb].meta = cur.pmeta;
          ++expansions;
          flag = false;                                     // break
        }
      }
    }
    if (flag) {                                             // :181-211
      --plen;
      const uint32_t v = cur.node;
      const uint32_t mv = cur.pmeta & ~kMetaOnPath;
      N[v].meta = mv;
      const int b = cur.bl;
      if (plen > 0) {
        cur = L[plen - 1];
        if (b < cur.bl) cur.bl = (int16_t)b;
      }
      if (b < M && (int)(mv & kMetaLock) < M - b + 1) {     // the first relax entry is (b, v), :192-196
        // A relax entry (x, u) only acts when lock[u] < M - x + 1, and locks never decrease during
        // a relaxation, so entries that fail the test when they would be pushed are dropped at once
        // (:198-199 push them, :195 discards them when popped); the test is repeated at the pop.
        // The top of the relax stack lives in a register (`r`, valid while `have`): the entry pushed last is the
        // one popped next, so a chain of single-parent relaxations never goes through memory.  Same LIFO order.
        uint32_t rn = 0;
        uint32_t r = ((uint32_t)b << 16) | v;
        bool have = true;
        while (have) {
          const int rb = (int)(r >> 16);
          const uint32_t u = r & 0xFFFFu;
          const uint32_t mu_ = N[u].meta;
          const uint64_t ins = T[u].in4;
          have = false;
          if ((int)(mu_ & kMetaLock) < M - rb + 1) {        // :195-196
            N[u].meta = (mu_ & ~kMetaLock) | (uint32_t)(M - rb + 1);
            ++expansions;                                   // _GetIncomings(u), :197
            const uint32_t nin = (mu_ >> kMetaInShift) & 7u;
            if (rn + nin > A.relax_cap) { status = kEnumOverflow; goto done; }
            uint32_t mw[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if ((uint32_t)i < nin) mw[i] = N[(uint32_t)(ins >> (16 * i)) & 0xFFFFu].meta;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if ((uint32_t)i < nin) {
                const uint32_t w = (uint32_t)(ins >> (16 * i)) & 0xFFFFu;
                if (!(mw[i] & kMetaOnPath) && (int)(mw[i] & kMetaLock) < M - rb) {  // :198-199 + early :195
                  if (have) A.relax[rn++] = r;              // the previous top goes to memory
                  r = ((uint32_t)(rb + 1) << 16) | w;
                  have = true;
                }
              }
          }
          if (!have && rn) {
            r = A.relax[--rn];
            have = true;
          }
        }
      }
    }
  }
done:
  R.status = status;
  R.n_cycles = n_cycles;
  R.n_nodes = n_cyc;
  R.steps = (uint32_t)steps;
  R.expansions = expansions;
  return true;
}

}  // namespace mg
