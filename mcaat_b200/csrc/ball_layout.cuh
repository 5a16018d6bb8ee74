// Per-worker arenas of the extracted-ball searches (ball_dev.cuh): one slot per worker warp in global memory plus a
// share of the block's shared memory.  Used by the enumeration pipeline (enumerate.cu) and by the ball tier of the
// admission search (starts.cu); the two never run at the same time and share the buffer in the graph handle.
#pragma once
#include <algorithm>

#include "ball_dev.cuh"
#include "common.cuh"

namespace mg {

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct BallLayout {  // per-worker arena in global memory (byte offsets inside one slot)
  uint32_t index_log, node_cap, relax_cap, cyc_cap, depth_cap;
  size_t off_nodes, off_tails, off_index, off_relax, off_cyc, off_cyc_pos, off_cyc_keys, off_cyc_off, off_cyc_noff,
      off_spos, off_ckeys, off_soff, off_shead, off_shot, off_sin, off_lvl, bytes;
  // per worker in shared memory: the path levels (Lvl; in the arena instead when cycle_max_length makes them large),
  // the top of the relaxation stack and the records of the contracted ball (up to chain_cap chains; larger
  // contracted balls use the records in the arena)
  uint32_t smem_bytes, smem_lvl, smem_relax, chain_cap;
};
constexpr uint32_t kRelaxSmem = 128;  // relaxation stack entries held in shared memory (the rest spills to the arena)
static BallLayout make_ball_layout(int max_len, uint32_t index_log, uint32_t smem_per_worker) {
  BallLayout L;
  L.index_log = index_log;
  L.node_cap = std::min<uint32_t>((1u << index_log) / 2, 65535u);
  L.relax_cap = std::max<uint32_t>(L.node_cap, 32768u);  // also holds the in-degree counters of the build
  L.depth_cap = (uint32_t)max_len + 4;
  L.cyc_cap = (uint32_t)kClusterBound * (uint32_t)(max_len + 1);
  size_t o = 0;
  L.off_nodes = o; o = align_up(o + (size_t)L.node_cap * sizeof(LNode), 128);
  L.off_tails = o; o = align_up(o + (size_t)L.node_cap * sizeof(LTail), 128);
  L.off_index = o; o = align_up(o + ((size_t)8 << index_log), 128);
  L.off_relax = o; o = align_up(o + (size_t)L.relax_cap * 4, 128);
  L.off_cyc = o; o = align_up(o + (size_t)L.cyc_cap * 2, 128);
  L.off_cyc_pos = o; o = align_up(o + (size_t)L.cyc_cap * 4, 128);
  L.off_cyc_keys = o; o = align_up(o + (size_t)L.cyc_cap * 4, 128);
  L.off_cyc_off = o; o = align_up(o + (size_t)(kClusterBound + 1) * 4, 128);
  L.off_cyc_noff = o; o = align_up(o + (size_t)(kClusterBound + 1) * 4, 128);
  L.off_spos = o; o = align_up(o + (size_t)L.node_cap * 4, 128);
  L.off_ckeys = o; o = align_up(o + (size_t)L.node_cap * 4, 128);
  L.off_soff = o; o = align_up(o + (size_t)(L.node_cap + 1) * 2, 128);
  L.off_shead = o; o = align_up(o + (size_t)L.node_cap * 2, 128);
  L.off_shot = o; o = align_up(o + (size_t)(L.node_cap + 1) * sizeof(SHot), 128);
  L.off_sin = o; o = align_up(o + (size_t)(L.node_cap + 1) * 8, 128);
  L.off_lvl = o; o = align_up(o + (size_t)L.depth_cap * sizeof(Lvl), 128);
  L.bytes = align_up(o, 128);
  const uint32_t lvl_bytes = (uint32_t)align_up((size_t)L.depth_cap * sizeof(Lvl), 16);
  L.smem_lvl = lvl_bytes <= smem_per_worker / 4 ? lvl_bytes : 0;  // long paths: levels in the arena
  L.smem_relax = kRelaxSmem * 4;
  const uint32_t rest = smem_per_worker - L.smem_lvl - L.smem_relax;
  L.chain_cap = std::min<uint32_t>(rest / 24 / 8 * 8, L.node_cap);
  L.smem_bytes = L.smem_lvl + L.smem_relax + L.chain_cap * 24;
  return L;
}
__device__ __forceinline__ BallArena ball_arena_of(uint8_t *base, uint8_t *smem, const BallLayout &L) {
  BallArena A;
  A.nodes = (LNode *)(base + L.off_nodes);
  A.tails = (LTail *)(base + L.off_tails);
  A.node_cap = L.node_cap;
  A.index = (uint64_t *)(base + L.off_index);
  A.index_log = L.index_log;
  A.relax = (uint32_t *)(base + L.off_relax);
  A.relax_cap = L.relax_cap;
  A.cyc = (uint16_t *)(base + L.off_cyc);
  A.cyc_cap = L.cyc_cap;
  A.cyc_off = (uint32_t *)(base + L.off_cyc_off);
  A.lvl = L.smem_lvl ? (Lvl *)smem : (Lvl *)(base + L.off_lvl);
  A.relax_top = (uint32_t *)(smem + L.smem_lvl);
  A.relax_top_cap = kRelaxSmem;
  A.spos = (uint32_t *)(base + L.off_spos);
  A.ckeys = (uint32_t *)(base + L.off_ckeys);
  A.soff = (uint16_t *)(base + L.off_soff);
  A.shead = (uint16_t *)(base + L.off_shead);
  A.cyc_pos = (uint32_t *)(base + L.off_cyc_pos);
  A.cyc_noff = (uint32_t *)(base + L.off_cyc_noff);
  A.max_steps = kMaxSteps;
  return A;
}


// the worker arenas in the graph handle (grow-only; the hash indices start out empty and every kernel leaves them so)
static inline int ensure_ball_arena(mg_graph *g, uint32_t workers, const BallLayout &L) {
  // the slot layout only: the two users may run different numbers of workers over the same slots
  const uint64_t sig = ((uint64_t)L.bytes << 20) ^ ((uint64_t)L.index_log << 8);
  if (g->enum_arena.bytes < (size_t)workers * L.bytes || g->enum_arena_sig != sig) {
    MG_TRY(g->enum_arena.ensure((size_t)workers * L.bytes));
    MG_CUDA(cudaMemsetAsync(g->enum_arena.p, 0xFF, g->enum_arena.bytes, g->stream));
    g->enum_arena_sig = sig;
  }
  return MG_OK;
}
// shared memory per worker warp when `workers_per_sm` of them share an SM (227 KB opt-in, 1 KB reserved per block)
static inline uint32_t ball_smem_per_worker(const mg_graph *g, uint32_t workers_per_sm, uint32_t warps_per_block) {
  const uint32_t blocks_per_sm = (workers_per_sm + warps_per_block - 1) / warps_per_block;
  const uint32_t per_block = ((uint32_t)g->smem_per_sm - blocks_per_sm * 1024u) / blocks_per_sm;
  return std::min<uint32_t>(per_block, (uint32_t)g->smem_per_block_optin) / warps_per_block / 16 * 16;
}

}  // namespace mg
