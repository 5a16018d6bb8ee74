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
  // ---- contracted ball (sball_*, below)
  uint32_t *spos;       // [node_cap] per ball node: chain id | position inside the chain << 16
  uint32_t *ckeys;      // [node_cap] item ids grouped by chain: chain s occupies [soff[s], soff[s] + c)
  uint16_t *soff;       // [node_cap + 1]
  uint16_t *shead;      // [node_cap] local id of the first node of chain s
  uint32_t *cyc_pos;    // [cyc_cap] per staged path level: where the chain's ids go in the flat cycle list
  uint32_t *cyc_noff;   // [kClusterBound + 1] cycle boundaries in the flat list (in nodes)
  int max_steps;        // step cap of :148-151 (kMaxSteps; the differential tests lower it)
  uint32_t *relax_top;  // first relax_top_cap entries of the contracted search's relaxation stack (shared memory)
  uint32_t relax_top_cap;
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
  // position of this lane among the lanes with p (all lanes call it)
  __device__ uint32_t rank(bool p) const { return (uint32_t)__popc(__ballot_sync(0xffffffffu, p) & ((1u << lane) - 1u)); }
  // appends v of every lane with p to list[n..) (lane order); returns the new length.  All lanes call it.
  __device__ uint32_t append(uint32_t *list, uint32_t n, bool p, uint32_t v) const {
    const unsigned m = __ballot_sync(0xffffffffu, p);
    if (p) list[n + (uint32_t)__popc(m & ((1u << lane) - 1u))] = v;
    return n + (uint32_t)__popc(m);
  }
};
MG_HD uint64_t index_cas(uint64_t *p, uint64_t expect, uint64_t val) {
  return (uint64_t)atomicCAS((unsigned long long *)p, (unsigned long long)expect, (unsigned long long)val);
}
MG_HD uint32_t meta_add(uint32_t *p, uint32_t v) { return atomicAdd(p, v); }
MG_HD void meta_or(uint32_t *p, uint32_t v) { atomicOr(p, v); }
MG_HD void st_cg64(uint64_t *p, uint64_t v) { __stcg((unsigned long long *)p, (unsigned long long)v); }
#else
struct Coop {
  int lane = 0;
  static constexpr int kLanes = 1;
  void sync() const {}
  uint32_t sum(uint32_t x) const { return x; }
  uint32_t exclusive_prefix(uint32_t) const { return 0; }
  bool any(bool p) const { return p; }
  uint32_t rank(bool) const { return 0; }
  uint32_t append(uint32_t *list, uint32_t n, bool p, uint32_t v) const {
    if (p) list[n++] = v;
    return n;
  }
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
MG_HD void meta_or(uint32_t *p, uint32_t v) { *p |= v; }
MG_HD void st_cg64(uint64_t *p, uint64_t v) { *p = v; }
#endif

// Leaves the hash index empty again (all lanes).  `n_nodes` = nodes of the previous build.
MG_HD void ball_clear(const BallArena &A, uint32_t n_nodes, const Coop &co) {
  for (uint32_t i = (uint32_t)co.lane; i < n_nodes; i += Coop::kLanes) st_cg64(&A.index[A.tails[i].slot], kIndexEmpty);
  co.sync();
}

// Finds or claims the index slots of up to 4 keys (one lane), in two halves.  index_claim_issue: the first probes of
// all keys are ONE round of compare-and-swaps in flight together — a claim is a single L2 round trip instead of a load
// followed by a CAS per key, one key after the other.  index_claim_resolve: only collisions probe on.  lid[j] = the
// local id the index already held for key j, or kLidPending when the key was claimed just now (bit j of `mine`: by this
// lane) and has no id yet.  A caller with several nodes in hand issues for all of them before it resolves any.
MG_HD void index_claim_issue(const BallArena &A, const uint32_t *key, uint32_t cnt, uint32_t *slot, uint64_t *got) {
#pragma unroll
  for (uint32_t j = 0; j < 4; ++j)
    if (j < cnt) {
      slot[j] = hash_u32(key[j], A.index_log);
      got[j] = index_cas(&A.index[slot[j]], kIndexEmpty, ((uint64_t)key[j] << 32) | kLidPending);
    }
}
MG_HD void index_claim_resolve(const BallArena &A, uint32_t imask, const uint32_t *key, uint32_t cnt, uint32_t *slot,
                               const uint64_t *got, uint32_t *lid, uint32_t &mine) {
#pragma unroll
  for (uint32_t j = 0; j < 4; ++j)
    if (j < cnt) {
      uint64_t cur = got[j];
      uint32_t h = slot[j];
      for (;;) {
        if (cur == kIndexEmpty) { mine |= 1u << j; cur = kLidPending; break; }
        if ((uint32_t)(cur >> 32) == key[j]) break;
        h = (h + 1) & imask;
        cur = index_cas(&A.index[h], kIndexEmpty, ((uint64_t)key[j] << 32) | kLidPending);
      }
      slot[j] = h;
      lid[j] = (uint32_t)cur;
    }
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

// the same through L2 (records another lane has just changed with an atomic)
MG_HD NodeHead ld_head_cg(const LNode *n) {
  NodeHead h;
#if defined(__CUDA_ARCH__)
  const uint4 v = __ldcg(reinterpret_cast<const uint4 *>(n));
  h.meta = v.x;
  h.out4 = (uint64_t)v.z | ((uint64_t)v.w << 32);
#else
  h.meta = n->meta;
  h.out4 = n->out4;
#endif
  return h;
}
#ifndef MG_BUILD_K
#define MG_BUILD_K 2
#endif
constexpr int kBuildK = MG_BUILD_K;  // frontier nodes per lane and round of the ball extraction
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
    // kBuildK frontier nodes per lane and round: their loads and claims are independent and overlap
    for (uint32_t base = lo; base < hi; base += Coop::kLanes * kBuildK) {
      uint32_t okn[kBuildK][4], slot[kBuildK][4], nlid[kBuildK][4], cnt[kBuildK], mine[kBuildK], node[kBuildK];
      uint64_t av[kBuildK];
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        cnt[k] = mine[k] = 0;
        node[k] = i < hi ? N[i].key : 0u;
      }
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        av[k] = i < hi ? ld_ro(&adj[node[k]]) : 0ull;
      }
      uint32_t nb[kBuildK][4], fm[kBuildK][4], fv[kBuildK][4];
      int deg[kBuildK];
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {  // multiplicity and visited words of every neighbour of every node: one wait
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        deg[k] = 0;
        if (i < hi) {
          deg[k] = adj_outgoing(av[k], nb[k]);
          if (deg[k] > 4) deg[k] = 4;
        }
        filtered_loads(g, visited, nb[k], deg[k], fm[k], fv[k]);
      }
      uint64_t got[kBuildK][4];
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        uint32_t nset = 0;
        filtered_finish(node[k], mu, nb[k], deg[k], fm[k], fv[k], nset, okn[k]);  // :58-72 incl. the set's iteration order
        cnt[k] = nset;
      }
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) index_claim_issue(A, okn[k], cnt[k], slot[k], got[k]);  // neighbours' index slots: one wait
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) index_claim_resolve(A, imask, okn[k], cnt[k], slot[k], got[k], nlid[k], mine[k]);
      // dense local ids for the newly discovered nodes of this round
      uint32_t n_mine = 0;
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) n_mine += (uint32_t)popc32(mine[k]);
      const uint32_t before = co.exclusive_prefix(n_mine), chunk_new = co.sum(n_mine);
      if (total_nodes + chunk_new > A.node_cap) {
        // no room: release the claimed slots (every lane has stopped probing) and give up
        co.sync();
#pragma unroll
        for (int k = 0; k < kBuildK; ++k)
          for (uint32_t j = 0; j < cnt[k]; ++j)
            if ((mine[k] >> j) & 1) st_cg64(&A.index[slot[k][j]], kIndexEmpty);
        co.sync();
        status = kEnumOverflow;
        break;
      }
      uint32_t lid = total_nodes + before;
#pragma unroll
      for (int k = 0; k < kBuildK; ++k)
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j)
          if (j < cnt[k] && ((mine[k] >> j) & 1)) {
            prefetch_l2(&adj[okn[k][j]]);  // expanded at the next level: its navigation entry is on its way by then
            N[lid].meta = (uint32_t)M;  // lock default (try_emplace(n, cycle_max_length), :168), no edges yet
            N[lid].key = okn[k][j];
            N[lid].out4 = ~0ull;
            T[lid].slot = slot[k][j];
            A.relax[lid] = 0;  // in-degree counter of the transpose below (the relax stack is idle during the build)
            st_cg64(&A.index[slot[k][j]], ((uint64_t)okn[k][j] << 32) | lid);
            nlid[k][j] = lid;
            ++lid;
          }
      total_nodes += chunk_new;
      co.sync();
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        if (i < hi) {
          uint64_t out4 = ~0ull;
#pragma unroll
          for (uint32_t j = 0; j < 4; ++j)
            if (j < cnt[k]) {  // only a neighbour that another lane claimed in this very round has to be looked up
              if (nlid[k][j] == kLidPending) nlid[k][j] = (uint32_t)ld_cg64(&A.index[slot[k][j]]);
              out4 = (out4 & ~(0xFFFFull << (16 * j))) | ((uint64_t)(nlid[k][j] & 0xFFFFu) << (16 * j));
            }
          N[i].out4 = out4;
          N[i].meta = (uint32_t)M | (cnt[k] << kMetaOutShift);  // the record still holds the default written at discovery
        }
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
  for (uint32_t u0 = (uint32_t)co.lane; u0 < total_nodes; u0 += Coop::kLanes * 2) {  // two nodes per lane: their counters overlap
    NodeHead h[2];
    uint32_t pos[2][4];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const uint32_t u = u0 + (uint32_t)q * Coop::kLanes;
      h[q].meta = 0;
      h[q].out4 = ~0ull;
      if (u < total_nodes) h[q] = ld_head(&N[u]);
    }
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (uint32_t j = 0; j < 4; ++j)
        if (j < ((h[q].meta >> kMetaOutShift) & 7u)) pos[q][j] = meta_add(&A.relax[(uint32_t)(h[q].out4 >> (16 * j)) & 0xFFFFu], 1u);
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (uint32_t j = 0; j < 4; ++j)
        if (j < ((h[q].meta >> kMetaOutShift) & 7u)) {
          const uint32_t x = (uint32_t)(h[q].out4 >> (16 * j)) & 0xFFFFu;
          if (pos[q][j] < 4) T[x].in[pos[q][j]] = (uint16_t)(u0 + (uint32_t)q * Coop::kLanes);
          else too_many = true;
        }
  }
  co.sync();
  if (co.any(too_many)) return kEnumOverflow;
  // in-degrees into the records (four nodes per lane and round); the nodes with several in-neighbours are listed
  // (A.spos is idle until the contraction) and ordered below
  uint32_t n_merge = 0;
  for (uint32_t x0 = 0; x0 < total_nodes; x0 += Coop::kLanes * 4) {
    uint32_t nin[4], meta[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t x = x0 + (uint32_t)q * Coop::kLanes + (uint32_t)co.lane;
      nin[q] = x < total_nodes ? ld_cg32(&A.relax[x]) : 0u;
      meta[q] = x < total_nodes ? N[x].meta : 0u;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t x = x0 + (uint32_t)q * Coop::kLanes + (uint32_t)co.lane;
      if (nin[q]) N[x].meta = meta[q] | (nin[q] << kMetaInShift);
      n_merge = co.append(A.spos, n_merge, nin[q] >= 2, x);
    }
  }
  co.sync();
  // _GetIncomings inserts in IncomingEdges order (ascending id, sdbg.h:266-285); reproduce the
  // iteration order of the resulting std::unordered_set
  for (uint32_t i = (uint32_t)co.lane; i < n_merge; i += Coop::kLanes) {
    const uint32_t x = A.spos[i];
    const uint32_t nin = (N[x].meta >> kMetaInShift) & 7u;
    const uint64_t in4 = T[x].in4;
    uint32_t lid[4], key[4];
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) {
      lid[j] = (uint32_t)(in4 >> (16 * j)) & 0xFFFFu;
      key[j] = j < nin ? N[lid[j]].key : 0xFFFFFFFFu;
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
    if (++steps > A.max_steps) { status = kEnumCap; break; }  // :148-151
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

// ====================================================================== contracted ball
// In a de Bruijn graph almost every node has one in- and one out-neighbour, and so has the ball: the 5 668-node ball
// of the longest search at 225 M items is ~40 spacer-long unary chains between a handful of branching nodes, and the
// literal search spends its 407 415 steps walking them node by node, twice (push, pop), plus once per relaxation.
// The search below runs on the CONTRACTED ball: a maximal chain n_1 .. n_c with OUT(n_i) = {n_i+1} and
// IN(n_i+1) = {n_i} (start node excluded: it is a chain of its own) is one record.  It is exact, not an
// approximation, because the lock values along a chain can only ever take one shape:
//     lock[n_i] = a + i  for i <= j,     lock[n_i] = M  for i > j            ("(a, j)", j = 0: untouched / all M)
//   * a push at path size P passes n_1 iff P < lock[n_1]; if it does it passes the whole affine prefix (the test
//     P + i < a + i + 1 is the same for every i) and then runs on through the M zone while P + i < M: the walk
//     pushes t = min(c, M - P) nodes and leaves (P, t) (:168-172).  t < c: the chain could not be completed inside
//     the length bound; nothing below it is reached, it is popped again at once with backtrack length M — no
//     relaxation, the state stays (P, t) — and the parent's set is scanned again.  That costs 2t - 1 further loop
//     iterations of :147 (t - 1 pushes, t pops), which only matter for the step cap;
//   * a pop with backtrack length M leaves the locks alone; with backtrack length 1 (the only other value there is:
//     :158 sets 1, :187 takes minima) every node of the chain is relaxed from its own pop to M (:192-196) — the
//     in-neighbour of an inner node is still on the path, so nothing spreads until n_1 is popped, whose
//     off-path in-neighbours are entered with level 2 (:197-199);
//   * a relaxation arrives at the LAST node n_c with level x (inner nodes have no other in-neighbour).  It proposes
//     M - x + 1 for n_c, one less for every node before it: again a + i with a' = M - x + 1 - c.  Against a full
//     affine state it raises all nodes or none (same difference everywhere); against an M zone at the end it stops
//     at n_c.  When it raises, it leaves through IN(n_1) with level x + c.
// A chain is either completely on the path or not at all whenever a relaxation runs (an incomplete chain only exists
// inside one step).  Neighbour sets keep the std::unordered_set iteration order of the node ids they came from
// (OUT(n_c), IN(n_1)); the relaxation stack keeps its LIFO order; the step and expansion counters advance by what
// the literal loop would have counted.  tests/emu compares this search with ball_dfs above (and both with the
// reference dumps), also with lowered step caps and length bounds that cut chains.
// One chain record, 16 bytes = one vector access:
//   word 0: lh | j << 16      lh = lock[n_1] = (j ? a + 1 : M): what a push tests; j as above; bit 31: the chain is on
//             | on_path << 31 the current path.  The only word the search writes.
//   word 1: c | flags << 16   flags bits 0..3: slots of OUT(n_c) that hold a chain other than the start node; bits
//                             12..14: slot of the start node in OUT(n_c) (7 = not a neighbour: the set holds the start
//                             at most once, and :157-166 never erases it)
//   words 2, 3: OUT(n_c) as four chain numbers in std::unordered_set iteration order (chain 0 = the start node);
//               unused slots hold the number of the DUMMY record (index n, lh = 0, j != c): it fails every test, so
//               the search loads and tests all four slots without asking which ones exist
struct alignas(16) SHot {
  uint16_t lh, j, c;
  uint16_t flags;
  union { uint16_t out[4]; uint64_t out4; };
};
static_assert(sizeof(SHot) == 16, "one 16-byte access per chain record");
constexpr uint32_t kSOnPath = 1u << 31, kSStartShift = 12, kSNoStart = 7, kSLevelBits = 0x700Fu;
struct SBall {          // the contracted ball; records in shared memory when they fit, else in the worker's arena
  SHot *hot;            // [n + 1]: the last one is the dummy
  uint64_t *in4;        // [n + 1] chains whose last node is in IN(n_1), iteration order; unused slots = dummy
  uint32_t cap;
  uint32_t n;           // number of chains
};

MG_HD bool ball_is_head(const LNode *N, const LTail *T, uint32_t x) {
  if (x == 0) return true;
  const uint32_t m = N[x].meta;
  if (((m >> kMetaInShift) & 7u) != 1) return true;
  const uint32_t u = T[x].in[0];
  return u == 0 || ((N[u].meta >> kMetaOutShift) & 7u) != 1;
}

// Contracts the ball of ball_build (all lanes).  Returns kEnumDone, or kEnumOverflow when B.cap is too small.
MG_HD_NOINLINE int sball_contract(const BallArena &A, uint32_t n_nodes, int M, SBall &B, const Coop &co) {
  const LNode *N = A.nodes;
  const LTail *T = A.tails;
  // 1. chain heads, numbered in local-id order (the start node is chain 0)
  uint32_t n_super = 0;
  for (uint32_t base = 0; base < n_nodes; base += Coop::kLanes) {
    const uint32_t x = base + (uint32_t)co.lane;
    const bool head = x < n_nodes && ball_is_head(N, T, x);
    const uint32_t before = co.exclusive_prefix(head ? 1u : 0u), cnt = co.sum(head ? 1u : 0u);
    if (head) A.shead[n_super + before] = (uint16_t)x;
    n_super += cnt;
  }
  B.n = n_super;
  if (n_super + 1 > B.cap || n_super >= kNoLid) return kEnumOverflow;
  co.sync();
  // 2. one walk per chain: membership of its nodes, length, last node
  for (uint32_t s = (uint32_t)co.lane; s < n_super; s += Coop::kLanes) {
    uint32_t x = A.shead[s], i = 0;
    const uint64_t in4 = T[x].in4;
    const uint32_t nin = (N[x].meta >> kMetaInShift) & 7u;
    for (;;) {
      A.spos[x] = s | (i << 16);
      ++i;
      if (s == 0 || ((N[x].meta >> kMetaOutShift) & 7u) != 1) break;
      const uint32_t y = N[x].out[0];
      if (y == 0 || ((N[y].meta >> kMetaInShift) & 7u) != 1) break;
      x = y;
    }
    SHot h;
    h.lh = (uint16_t)M;       // untouched: every lock is cycle_max_length (try_emplace(n, M), :168)
    h.j = 0;
    h.c = (uint16_t)i;
    h.flags = 0;
    h.out4 = N[x].out4;       // local ids of OUT(n_c); unused slots are kNoLid already
    B.hot[s] = h;
    uint64_t in = in4;        // IN(n_1): slots beyond |in| are not initialised by the build
    for (uint32_t q = nin; q < 4; ++q) in |= 0xFFFFull << (16 * q);
    B.in4[s] = in;
  }
  co.sync();
  // 3. where each chain's ids go (exclusive prefix of the lengths)
  uint32_t run = 0;
  for (uint32_t base = 0; base < n_super; base += Coop::kLanes) {
    const uint32_t s = base + (uint32_t)co.lane;
    const uint32_t c = s < n_super ? B.hot[s].c : 0u;
    const uint32_t before = co.exclusive_prefix(c), tot = co.sum(c);
    if (s < n_super) A.soff[s] = (uint16_t)(run + before);
    run += tot;
  }
  co.sync();
  // 4. neighbour sets as chain numbers (unused slots: the dummy); item ids grouped by chain
  for (uint32_t s = (uint32_t)co.lane; s < n_super; s += Coop::kLanes) {
    uint64_t o = B.hot[s].out4, in = B.in4[s], o2 = 0, in2 = 0;
    uint32_t sslot = kSNoStart, vmask = 0;
    for (int q = 0; q < 4; ++q) {
      const uint32_t a = (uint32_t)(o >> (16 * q)) & 0xFFFFu, b = (uint32_t)(in >> (16 * q)) & 0xFFFFu;
      const uint32_t ca = a == kNoLid ? n_super : (A.spos[a] & 0xFFFFu);
      if (ca == 0) sslot = (uint32_t)q;
      else if (a != kNoLid) vmask |= 1u << q;
      o2 |= (uint64_t)ca << (16 * q);
      in2 |= (uint64_t)(b == kNoLid ? n_super : (A.spos[b] & 0xFFFFu)) << (16 * q);
    }
    B.hot[s].out4 = o2;
    B.hot[s].flags = (uint16_t)((sslot << kSStartShift) | vmask);
    B.in4[s] = in2;
  }
  if (co.lane == 0) {  // the dummy
    SHot d;
    d.lh = 0;
    d.j = 0x8000u;  // "on the path": no relaxation enters it
    d.c = 1;
    d.flags = (uint16_t)(kSNoStart << kSStartShift);
    d.out4 = 0x0001000100010001ull * n_super;
    B.hot[n_super] = d;
    B.in4[n_super] = 0x0001000100010001ull * n_super;
  }
  for (uint32_t x = (uint32_t)co.lane; x < n_nodes; x += Coop::kLanes) {
    const uint32_t sp = A.spos[x];
    A.ckeys[(uint32_t)A.soff[sp & 0xFFFFu] + (sp >> 16)] = N[x].key;
  }
  co.sync();
  return kEnumDone;
}

struct SDfsState {  // the scalars of a paused search
  int depth = 0;     // chains on the path
  int plen = 0;      // path.size()
  int counter = 0, steps = 0;
  uint32_t expansions = 0, n_cycles = 0, n_staged = 0, n_out = 0;
  uint32_t iters = 0, relax_pops = 0;  // loop iterations on the contracted ball / relaxation entries popped (profiling)
  bool started = false;
};

// Register image of a 16-byte record.  A path level (Lvl) is stored as
//   x, y: the neighbour set being scanned (OUT of the level's chain, four chain numbers)
//   z:    chain | path.size() below this chain << 16
//   w:    bits 0..3 slots (other than the start node's) not yet erased (:173) | bit 8: backtrack length is 1 (:158;
//         the only other value is M) | bits 12..14 slot of the start node in the set — the record's flags, masked
struct Rec4 {
  uint32_t x, y, z, w;
};
constexpr uint32_t kLvlReached = 1u << 8, kLvlStartShift = 12;
// Where the records live.  RecsPtr: plain pointers (arena, host emulation, or shared memory through generic
// addresses); RecsSmem (device): everything in shared memory, addressed with ld/st.shared.
struct RecsPtr {
  SHot *hot;
  const uint64_t *in4;
  Lvl *lvl;
  uint32_t *rtop;
  MG_HD Rec4 ld_hot(uint32_t s) const {
    Rec4 h;
#if defined(__CUDA_ARCH__)
    const uint4 v = *reinterpret_cast<const uint4 *>(&hot[s]);
    h.x = v.x; h.y = v.y; h.z = v.z; h.w = v.w;
#else
    const SHot &r = hot[s];
    h.x = (uint32_t)r.lh | ((uint32_t)r.j << 16); h.y = (uint32_t)r.c | ((uint32_t)r.flags << 16);
    h.z = (uint32_t)r.out4; h.w = (uint32_t)(r.out4 >> 32);
#endif
    return h;
  }
  MG_HD void st_hot_w0(uint32_t s, uint32_t w0) const {
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<uint32_t *>(&hot[s]) = w0;
#else
    hot[s].lh = (uint16_t)w0; hot[s].j = (uint16_t)(w0 >> 16);
#endif
  }
  MG_HD void ld_in(uint32_t s, uint32_t &lo, uint32_t &hi) const { const uint64_t v = in4[s]; lo = (uint32_t)v; hi = (uint32_t)(v >> 32); }
  MG_HD Rec4 ld_lvl(int i) const {
    Rec4 l;
#if defined(__CUDA_ARCH__)
    const uint4 v = *reinterpret_cast<const uint4 *>(&lvl[i]);
    l.x = v.x; l.y = v.y; l.z = v.z; l.w = v.w;
#else
    memcpy(&l, &lvl[i], 16);
#endif
    return l;
  }
  MG_HD void st_lvl(int i, const Rec4 &l) const {
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<uint4 *>(&lvl[i]) = make_uint4(l.x, l.y, l.z, l.w);
#else
    memcpy(&lvl[i], &l, 16);
#endif
  }
  MG_HD uint32_t ld_lvl_z(int i) const {
#if defined(__CUDA_ARCH__)
    return reinterpret_cast<const uint32_t *>(&lvl[i])[2];
#else
    uint32_t z;
    memcpy(&z, (const char *)&lvl[i] + 8, 4);
    return z;
#endif
  }
  MG_HD uint32_t ld_rtop(uint32_t i) const { return rtop[i]; }
  MG_HD void st_rtop(uint32_t i, uint32_t v) const { rtop[i] = v; }
};
#if defined(__CUDACC__)
struct RecsSmem {
  uint32_t hot, in4, lvl, rtop;  // shared-window byte addresses
  __device__ __forceinline__ Rec4 ld_hot(uint32_t s) const {
    Rec4 h;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(h.x), "=r"(h.y), "=r"(h.z), "=r"(h.w) : "r"(hot + s * 16u));
    return h;
  }
  __device__ __forceinline__ void st_hot_w0(uint32_t s, uint32_t w0) const {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(hot + s * 16u), "r"(w0) : "memory");
  }
  __device__ __forceinline__ void ld_in(uint32_t s, uint32_t &lo, uint32_t &hi) const {
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(in4 + s * 8u));
  }
  __device__ __forceinline__ Rec4 ld_lvl(int i) const {
    Rec4 l;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(l.x), "=r"(l.y), "=r"(l.z), "=r"(l.w) : "r"(lvl + (uint32_t)i * 16u));
    return l;
  }
  __device__ __forceinline__ void st_lvl(int i, const Rec4 &l) const {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(lvl + (uint32_t)i * 16u), "r"(l.x), "r"(l.y), "r"(l.z), "r"(l.w) : "memory");
  }
  __device__ __forceinline__ uint32_t ld_lvl_z(int i) const {
    uint32_t z;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(z) : "r"(lvl + (uint32_t)i * 16u + 8u));
    return z;
  }
  __device__ __forceinline__ uint32_t ld_rtop(uint32_t i) const {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(rtop + i * 4u));
    return v;
  }
  __device__ __forceinline__ void st_rtop(uint32_t i, uint32_t v) const {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(rtop + i * 4u), "r"(v) : "memory");
  }
};
#endif
MG_HD int ffs32(uint32_t x) {  // 1-based position of the lowest set bit (x != 0)
#if defined(__CUDA_ARCH__)
  return __ffs((int)x);
#else
  return __builtin_ffs((int)x);
#endif
}

// Literal FindCycleUtil + FindCycle (src/cycle_finder.cpp:231-243,140-225) on the contracted ball (one lane).  Runs at
// most `budget` further chain steps; returns true when the search has ended (R filled in), false when it is paused.
// Cycles are staged as (chain, position in the flat id list) per path level; sball_expand turns them into item ids.
// The loop is bound by instruction issue of ONE thread (a dependent chain: ~4 cycles per instruction), not by memory,
// so it is written for few instructions: 32-bit words only, all four slots of a neighbour set loaded and tested
// without branches (dummy record), the first slot that passes picked with a bit scan.
// (Inlined at each call site: the device instantiates it for records in shared memory and for records in the arena.)
template <class Recs>
MG_HD bool sball_dfs(int min_len, int max_len, const BallArena &A, const Recs &Q, SDfsState &S, int budget,
                     EnumResult &R) {
  const uint32_t M = (uint32_t)max_len;
  uint32_t expansions = S.expansions, n_cycles = S.n_cycles, n_staged = S.n_staged, n_out = S.n_out;
  uint32_t plen = (uint32_t)S.plen;
  int depth = S.depth, counter = S.counter, steps = S.steps;
  int32_t status = kEnumDone;
  uint32_t iters = S.iters, relax_pops = S.relax_pops;
  const int max_steps = A.max_steps;
  Rec4 cur;  // top level in registers
  if (!S.started) {
    S.started = true;
    A.cyc_off[0] = 0;
    A.cyc_noff[0] = 0;
    const Rec4 h0 = Q.ld_hot(0);
    Q.st_hot_w0(0, kSOnPath);                        // lock[start] = 0 (:237); the start stays on the path
    cur.x = h0.z;                                    // _GetOutgoings(start), :238-240
    cur.y = h0.w;
    cur.z = 0;
    cur.w = (h0.y >> 16) & kSLevelBits;
    plen = 1;
    depth = 1;
    ++expansions;
  } else {
    cur = Q.ld_lvl(depth - 1);
  }
  for (int left = budget; depth > 0; --left) {
    if (left == 0) {
      Q.st_lvl(depth - 1, cur);
      S.depth = depth; S.plen = (int)plen; S.counter = counter; S.steps = steps;
      S.expansions = expansions; S.n_cycles = n_cycles; S.n_staged = n_staged; S.n_out = n_out;
      S.iters = iters; S.relax_pops = relax_pops;
      return false;
    }
    ++iters;
    // :148-151.  `steps` also advances in bulk below (a chain's inner pushes and pops are loop iterations of the
    // reference); nothing is recorded between such an advance and this test, so one test per chain step is exact.
    if (++steps > max_steps) { status = kEnumCap; break; }
    // stack.back() (:153), in iteration order: up to four chains; bits 0..3 of cur.w = not erased yet
    uint32_t first = 4, nb = 0, hy = 0, hz = 0, hw = 0;
    if (cur.w & 0xFu) {
      const uint32_t nb0 = cur.x & 0xFFFFu, nb1 = cur.x >> 16, nb2 = cur.y & 0xFFFFu, nb3 = cur.y >> 16;
      const Rec4 r0 = Q.ld_hot(nb0), r1 = Q.ld_hot(nb1), r2 = Q.ld_hot(nb2), r3 = Q.ld_hot(nb3);
      // :168 for n_1 of each member: path.size() < lock
      uint32_t m = (plen < (r0.x & 0xFFFFu) ? 1u : 0u) | (plen < (r1.x & 0xFFFFu) ? 2u : 0u) |
                   (plen < (r2.x & 0xFFFFu) ? 4u : 0u) | (plen < (r3.x & 0xFFFFu) ? 8u : 0u);
      m &= cur.w;
      if (m) {                                                // the first member that passes is pushed
        const uint32_t b = m & (0u - m);                      // lowest set bit: 1, 2, 4 or 8
        first = (b >> 1) - (b >> 3);
        const bool lo2 = (b & 3u) != 0, even = (b & 5u) != 0;
        nb = lo2 ? (even ? nb0 : nb1) : (even ? nb2 : nb3);
        hy = lo2 ? (even ? r0.y : r1.y) : (even ? r2.y : r3.y);
        hz = lo2 ? (even ? r0.z : r1.z) : (even ? r2.z : r3.z);
        hw = lo2 ? (even ? r0.w : r1.w) : (even ? r2.w : r3.w);
        cur.w ^= b;                                           // stack.back().erase(n), :173
      }
    }
    if (((cur.w >> kSStartShift) & 7u) < first) {             // n == start_node comes before it, :157-166
      cur.w |= kLvlReached;                                   // backtrack_lengths.back() = 1
      if ((int)plen > min_len) {
        if (++counter >= kClusterBound) {                     // cluster bound: the final answer is empty
          status = kEnumBound;
          n_cycles = n_staged = n_out = 0;
          break;
        }
        if (n_staged + (uint32_t)depth > A.cyc_cap) { status = kEnumOverflow; break; }
        for (int l = 0; l + 1 < depth; ++l) {
          const uint32_t z = Q.ld_lvl_z(l);
          A.cyc[n_staged + l] = (uint16_t)z;
          A.cyc_pos[n_staged + l] = n_out + (z >> 16);
        }
        A.cyc[n_staged + depth - 1] = (uint16_t)cur.z;
        A.cyc_pos[n_staged + depth - 1] = n_out + (cur.z >> 16);
        n_staged += (uint32_t)depth;
        n_out += plen;
        A.cyc_off[++n_cycles] = n_staged;
        A.cyc_noff[n_cycles] = n_out;
      }
    }
    if (first < 4) {
      const uint32_t c = hy & 0xFFFFu, room = M - plen, t = c < room ? c : room;
      expansions += t;                                        // one _GetOutgoings per pushed node, :174-176
      const uint32_t w0 = (plen + 1u) | (t << 16);            // locks of n_1..n_t become plen + i
      if (t < c) {                                            // cut by the length bound: popped again at once
        Q.st_hot_w0(nb, w0);
        steps += (int)(2 * t - 1);
      } else {
        Q.st_hot_w0(nb, w0 | kSOnPath);
        Q.st_lvl(depth - 1, cur);
        steps += (int)(c - 1);
        ++depth;
        cur.x = hz;
        cur.y = hw;
        cur.z = nb | (plen << 16);
        cur.w = (hy >> 16) & kSLevelBits;
        plen += c;
      }
    } else {                                                  // :181-211
      const uint32_t v = cur.z & 0xFFFFu, below = cur.z >> 16, c = plen - below;
      const uint32_t vw = cur.w;
      plen = below;
      --depth;
      steps += (int)(c - 1);                                  // the chain's other nodes are popped one step each
      if (depth > 0) {
        cur = Q.ld_lvl(depth - 1);
        cur.w |= vw & kLvlReached;                            // :187 (minimum of the two backtrack lengths)
      }
      if (!(vw & kLvlReached)) {
        Q.st_hot_w0(v, (plen + 1u) | (c << 16));              // off the path; locks stay plen + i
      } else {                                                // every node with lock < M goes to M, :192-196
        const uint32_t lim = M - plen - 1u;                   // nodes i with plen + i < M
        expansions += c < lim ? c : lim;
        Q.st_hot_w0(v, M);
        if (plen + 1u < M) {                                  // n_1 was relaxed: its off-path in-neighbours, level 2
          // relax_stack (:192-210) over chains: entry = level << 16 | chain, meaning (level, last node of the chain).
          // An entry acts iff lock[n_c] < M - level + 1 — for a fully affine chain (j == c) that is
          // lh + c + level < M + 2, and then it holds for the whole chain.  Locks never decrease during a
          // relaxation, so entries that fail when they would be pushed are dropped at once; the test is
          // repeated at the pop.
          uint32_t rn = 0, x = 2, in_lo, in_hi;
          Q.ld_in(v, in_lo, in_hi);
          for (;;) {
            const uint32_t w0 = in_lo & 0xFFFFu, w1 = in_lo >> 16, w2 = in_hi & 0xFFFFu, w3 = in_hi >> 16;
            const Rec4 q0 = Q.ld_hot(w0), q1 = Q.ld_hot(w1), q2 = Q.ld_hot(w2), q3 = Q.ld_hot(w3);
            const uint32_t ws[4] = {w0, w1, w2, w3};
            const Rec4 qs[4] = {q0, q1, q2, q3};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t lh = qs[i].x & 0xFFFFu, j = qs[i].x >> 16, cc = qs[i].y & 0xFFFFu;
              if (j == cc && lh + cc + x < M + 2u) {          // :198-199 (on the path: bit 15 of j) + early :195
                if (rn >= A.relax_cap + A.relax_top_cap) { status = kEnumOverflow; goto done; }
                const uint32_t ent = (x << 16) | ws[i];
                if (rn < A.relax_top_cap) Q.st_rtop(rn, ent);
                else A.relax[rn - A.relax_top_cap] = ent;
                ++rn;
              }
            }
            bool acted = false;
            while (rn && !acted) {
              --rn;
              ++relax_pops;
              const uint32_t r = rn < A.relax_top_cap ? Q.ld_rtop(rn) : A.relax[rn - A.relax_top_cap];
              const uint32_t rx = r >> 16, u = r & 0xFFFFu;
              const Rec4 hu = Q.ld_hot(u);
              const uint32_t lh = hu.x & 0xFFFFu, j = hu.x >> 16, cc = hu.y & 0xFFFFu;
              if (j == cc && lh + cc + rx < M + 2u) {         // :195-196 at n_c, true for the whole chain
                Q.st_hot_w0(u, (M + 2u - rx - cc) | (cc << 16));  // lock[n_i] = M - rx + 1 - (c - i)
                expansions += cc;                             // _GetIncomings per node, :197
                Q.ld_in(u, in_lo, in_hi);
                x = rx + cc;
                acted = true;
              }
            }
            if (!acted) break;
          }
        }
      }
    }
  }
done:
  if (steps > max_steps) steps = max_steps + 1;
  R.status = status;
  R.n_cycles = n_cycles;
  R.n_nodes = n_out;
  R.steps = (uint32_t)steps;
  R.expansions = expansions;
  S.n_staged = n_staged;
  S.iters = iters;
  S.relax_pops = relax_pops;
  return true;
}

// Item ids of the staged cycles (all lanes): out[0 .. R.n_nodes), cycle q = [cyc_noff[q], cyc_noff[q + 1]).
MG_HD void sball_expand(const BallArena &A, const SBall &B, uint32_t n_staged, uint32_t *out, const Coop &co) {
  for (uint32_t e = (uint32_t)co.lane; e < n_staged; e += Coop::kLanes) {
    const uint32_t s = A.cyc[e], at = A.cyc_pos[e], from = A.soff[s], c = B.hot[s].c;
    for (uint32_t i = 0; i < c; ++i) out[at + i] = A.ckeys[from + i];
  }
  co.sync();
}

// ====================================================================== DepthLevelSearch on an extracted ball
// The admission search (reference src/cycle_finder.cpp:248-343, dls_run in search_dev.cuh) is an order-dependent DFS
// with ONE global seen-set: exact emulation is sequential, and on the graph every expansion is a dependent DRAM
// access (~0.7 us).  99 % of the searches are short; the few thousand long ones (10^4 expansions) used to bound the
// whole phase by the length of the longest.  For those the same two stages as for the enumeration:
//   1. dball_build — the warp extracts the forward ball of radius `limit` (everything the DFS can reach: a node is
//      pushed at depth <= limit, BFS distance <= DFS depth) level by level, 32 nodes at a time, with each node's
//      OutgoingEdges list (order kept: descending ids, sdbg.h:297-320), its in-degree inside the ball and its first
//      in-neighbour;
//   2. dball_contract + dball_dfs — a non-start node is pushed at most once (seen-set), so a maximal chain
//      n_1 .. n_c (one out-neighbour each, entered only through n_1) is walked exactly once, right after n_1 is
//      popped: one record, one step.  The DFS itself is the literal one: LIFO stack, children pushed in
//      OutgoingEdges order, the start node pushed again whenever it is reached at depth > 0, found when the start is
//      expanded at depth > 1 (:337-339), depth test before the pushes (:312), expansion counter as dls_run counts it.
// Capacity overflow (ball > node capacity) is reported as kDlsSetFull; the caller falls back to the bitmap tier.
struct alignas(16) DChain {
  uint32_t w0;   // c | tail has an out-neighbour << 16 | head seen << 31
  uint32_t pad;
  union { uint16_t out[4]; uint64_t out4; };  // chains OUT(n_c) leads to, OutgoingEdges order; unused = kNoLid
};
static_assert(sizeof(DChain) == 16, "chain record");
constexpr uint32_t kDSeen = 1u << 31, kDTailOut = 1u << 16;
// node meta of dball_build beyond the degree fields: the node starts a chain record although it continues a chain (every
// kDChainCut-th level: bounds the serial walk of the contraction), the node continues the chain of its in-neighbour
constexpr uint32_t kDMetaCut = 1u << 25, kDMetaNonHead = 1u << 26;
constexpr int kDChainCut = 16;

MG_HD_NOINLINE int dball_build(const uint64_t *adj, uint32_t start, int limit, const BallArena &A, uint32_t &n_nodes,
                               const Coop &co) {
  const uint32_t imask = (1u << A.index_log) - 1;
  LNode *N = A.nodes;
  LTail *T = A.tails;
  if (co.lane == 0) {
    uint32_t h = hash_u32(start, A.index_log);
    st_cg64(&A.index[h], ((uint64_t)start << 32) | 0u);
    N[0].meta = 0;
    N[0].key = start;
    N[0].out4 = ~0ull;
    T[0].slot = h;
    A.relax[0] = 0;
  }
  co.sync();
  uint32_t total_nodes = 1, lo = 0, hi = 1;
  int status = kDlsNotFound;
  for (int depth = 0; depth <= limit && lo < hi && status == kDlsNotFound; ++depth) {
    const bool expand = depth < limit;  // nodes at distance `limit` are only asked for their out-degree (:294)
    const uint32_t cut_here = depth > 0 && depth % kDChainCut == 0 ? kDMetaCut : 0u;
    const uint32_t cut_next = (depth + 1) % kDChainCut == 0 ? kDMetaCut : 0u;
    for (uint32_t base = lo; base < hi; base += Coop::kLanes * kBuildK) {  // kBuildK nodes per lane, as in ball_build
      uint32_t nb[kBuildK][4], slot[kBuildK][4], nlid[kBuildK][4], cnt[kBuildK], mine[kBuildK];
      uint64_t av[kBuildK];
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        cnt[k] = mine[k] = 0;
        av[k] = i < hi ? ld_ro(&adj[N[i].key]) : 0ull;
      }
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        if (i < hi) {
          int deg = adj_outgoing(av[k], nb[k]);
          if (deg > 4) deg = 4;
          cnt[k] = (uint32_t)deg;
          if (!expand) {
            N[i].meta = (cnt[k] << kMetaOutShift) | cut_here;  // out-degree on the graph (:294); no neighbours recorded
            cnt[k] = 0;
          }
        }
      }
      uint64_t got[kBuildK][4];
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) index_claim_issue(A, nb[k], cnt[k], slot[k], got[k]);  // neighbours' index slots: one wait
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) index_claim_resolve(A, imask, nb[k], cnt[k], slot[k], got[k], nlid[k], mine[k]);
      uint32_t n_mine = 0;
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) n_mine += (uint32_t)popc32(mine[k]);
      const uint32_t before = co.exclusive_prefix(n_mine), chunk_new = co.sum(n_mine);
      if (total_nodes + chunk_new > A.node_cap) {
        co.sync();
#pragma unroll
        for (int k = 0; k < kBuildK; ++k)
          for (uint32_t j = 0; j < cnt[k]; ++j)
            if ((mine[k] >> j) & 1) st_cg64(&A.index[slot[k][j]], kIndexEmpty);
        co.sync();
        status = kDlsSetFull;
        break;
      }
      uint32_t lid = total_nodes + before;
#pragma unroll
      for (int k = 0; k < kBuildK; ++k)
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j)
          if (j < cnt[k] && ((mine[k] >> j) & 1)) {
            prefetch_l2(&adj[nb[k][j]]);  // expanded (or asked for its out-degree) at the next level
            N[lid].meta = cut_next;
            N[lid].key = nb[k][j];
            N[lid].out4 = ~0ull;
            T[lid].slot = slot[k][j];
            A.relax[lid] = 0;  // in-degree inside the ball
            st_cg64(&A.index[slot[k][j]], ((uint64_t)nb[k][j] << 32) | lid);
            nlid[k][j] = lid;
            ++lid;
          }
      total_nodes += chunk_new;
      co.sync();
#pragma unroll
      for (int k = 0; k < kBuildK; ++k) {
        const uint32_t i = base + (uint32_t)k * Coop::kLanes + (uint32_t)co.lane;
        if (i < hi && expand) {
          uint64_t out4 = ~0ull;
          uint32_t was[4];
#pragma unroll
          for (uint32_t j = 0; j < 4; ++j)
            if (j < cnt[k] && nlid[k][j] == kLidPending) nlid[k][j] = (uint32_t)ld_cg64(&A.index[slot[k][j]]);
#pragma unroll
          for (uint32_t j = 0; j < 4; ++j)
            if (j < cnt[k]) {
              out4 = (out4 & ~(0xFFFFull << (16 * j))) | ((uint64_t)(nlid[k][j] & 0xFFFFu) << (16 * j));
              was[j] = meta_add(&A.relax[nlid[k][j]], 1u);
            }
#pragma unroll
          for (uint32_t j = 0; j < 4; ++j)
            if (j < cnt[k] && was[j] == 0) T[nlid[k][j]].in[0] = (uint16_t)i;  // the in-neighbour, if it stays the only one
          N[i].out4 = out4;
          N[i].meta = (cnt[k] << kMetaOutShift) | (cnt[k] << kMetaInShift) | cut_here;  // out-degree | out-neighbours recorded
        }
      }
    }
    lo = hi;
    hi = total_nodes;
  }
  co.sync();
  n_nodes = total_nodes;
  return status;
}

// (node meta of dball_build: bits 16..18 out-degree on the graph, bits 20..22 out-neighbours recorded in the ball — the
//  same, except for nodes at distance `limit`, which are not expanded)
// chains of the ball -> C[0 .. n_chains); returns the number of chains.  A node continues the chain of its in-neighbour
// u when it is not the start, has no other in-neighbour, u is not the start and has no other out-neighbour — and the
// node does not lie on a level where chains are cut (any partition of a chain into consecutive pieces is searched
// alike: a piece is one step of the search, its tail pushes the head of the next piece).
MG_HD_NOINLINE uint32_t dball_contract(const BallArena &A, uint32_t n_nodes, DChain *C, const Coop &co) {
  LNode *N = A.nodes;
  // parents mark the children that continue their chain (four nodes per lane and round: the loads overlap)
  for (uint32_t u0 = 0; u0 < n_nodes; u0 += Coop::kLanes * 4) {
    NodeHead h[4];
    uint32_t indeg[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t u = u0 + (uint32_t)q * Coop::kLanes + (uint32_t)co.lane;
      h[q].meta = 0;
      h[q].out4 = ~0ull;
      if (u != 0 && u < n_nodes) h[q] = ld_head(&N[u]);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t x = (uint32_t)h[q].out4 & 0xFFFFu;
      indeg[q] = ((h[q].meta >> kMetaInShift) & 7u) == 1 && x != 0 ? ld_cg32(&A.relax[x]) : 0u;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (indeg[q] == 1) meta_or(&N[(uint32_t)h[q].out4 & 0xFFFFu].meta, kDMetaNonHead);
  }
  co.sync();
  uint32_t n_chains = 0;
  for (uint32_t base = 0; base < n_nodes; base += Coop::kLanes) {
    const uint32_t x = base + (uint32_t)co.lane;
    const uint32_t m = x < n_nodes ? ld_cg32(&N[x].meta) : 0u;
    const bool head = x < n_nodes && (!(m & kDMetaNonHead) || (m & kDMetaCut));
    const uint32_t r = co.rank(head);
    if (head) A.spos[x] = n_chains + r;  // chain number of a head (other nodes are never looked up)
    n_chains = co.append(A.ckeys, n_chains, head, x);
  }
  co.sync();
  for (uint32_t s = (uint32_t)co.lane; s < n_chains; s += Coop::kLanes) {
    uint32_t c = 1;
    NodeHead cur = ld_head_cg(&N[A.ckeys[s]]);
    for (;;) {  // one load per node of the chain
      if (s == 0 || ((cur.meta >> kMetaInShift) & 7u) != 1) break;
      const NodeHead nx = ld_head_cg(&N[(uint32_t)cur.out4 & 0xFFFFu]);
      if (!(nx.meta & kDMetaNonHead) || (nx.meta & kDMetaCut)) break;
      cur = nx;
      ++c;
    }
    const uint32_t deg = (cur.meta >> kMetaOutShift) & 7u;
    DChain d;
    d.w0 = c | (deg ? kDTailOut : 0u);
    d.pad = 0;
    uint64_t o = cur.out4, o2 = 0;
    for (int q = 0; q < 4; ++q) {
      const uint32_t a = (uint32_t)(o >> (16 * q)) & 0xFFFFu;
      o2 |= (uint64_t)(a == kNoLid ? kNoLid : (A.spos[a] & 0xFFFFu)) << (16 * q);  // out-neighbours of a tail are heads
    }
    d.out4 = o2;
    C[s] = d;
  }
  co.sync();
  return n_chains;
}
// Literal DepthLevelSearch(start, start, limit) on the contracted ball (one lane).  `stk` holds 4 * limit + 8 entries.
MG_HD_NOINLINE int dball_dfs(DChain *C, int limit, uint32_t *stk, uint32_t stk_cap, uint32_t &expansions) {
  uint32_t sp = 0;
  stk[sp++] = 0u;  // (chain 0 = the start node, depth 0)
  while (sp) {
    const uint32_t e = stk[--sp];
    const uint32_t sid = e & 0xFFFFu;
    const int d = (int)(e >> 16);
    const DChain R = C[sid];
    const int c = (int)(R.w0 & 0xFFFFu);
    const int room = limit - d + 1;                     // nodes of the chain popped at depth <= limit
    const int t = c < room ? c : room;
    const bool tail_out = (R.w0 & kDTailOut) != 0;
    expansions += (uint32_t)((t < c || tail_out) ? t : t - 1);  // pops with an out-neighbour (:294 comes before the count)
    if (t < c || !tail_out) continue;                   // cut by the depth limit inside the chain / dead end
    const int dt = d + c - 1;                           // depth of the last node
    if (dt >= limit) continue;                          // :312
    uint32_t hw[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t h = (uint32_t)(R.out4 >> (16 * j)) & 0xFFFFu;
      hw[j] = h != kNoLid ? C[h].w0 : kDSeen;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {                       // :320-334, OutgoingEdges order
      const uint32_t h = (uint32_t)(R.out4 >> (16 * j)) & 0xFFFFu;
      if (h == kNoLid) continue;
      if (!(hw[j] & kDSeen) || (h == 0 && dt > 0)) {
        C[h].w0 = hw[j] | kDSeen;
        if (sp >= stk_cap) return kDlsStackFull;
        stk[sp++] = h | ((uint32_t)(dt + 1) << 16);
      }
    }
    if (sid == 0 && dt > 1) return kDlsFound;            // :337-339
  }
  return kDlsNotFound;
}

}  // namespace mg
