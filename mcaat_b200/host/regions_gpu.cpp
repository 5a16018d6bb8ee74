// Drop-in for the reference's keep_crispr_regions_extended_by_k (reference src/spacer_ordering.cpp:78-139,
// declared include/spacer_ordering.h:127-131): same arguments, same effect on the host SDBG — after the call
// exactly the items the reference would leave valid are valid.  The neighbourhood growth and the
// invalidation run on the GPU (mg_keep_regions, include/mcaat_gpu.h); the host only flattens the cycles,
// merges its invalid bits into the device graph and writes the result back with SetInvalidEdge.
//
// The device graph of a live CycleFinder drop-in on the same SDBG is reused (it is still in HBM when the
// reference reaches this stage, src/main.cpp:536-556); without one the graph is uploaded for this call.
// No CPU fallback: a C-ABI error becomes std::runtime_error.
//
// Two ways to adopt it: call mcaat_gpu::get_crispr_regions_extended_by_k from
// src/main_run_and_debug.cpp:41 (or the two mcaat_gpu:: functions at src/spacer_ordering.cpp:182-183), or delete
// the reference's definitions and compile this file with -DMCAAT_GPU_DEFINE_REFERENCE_SYMBOLS, which defines
// the reference's own symbols.
//
// divide_graph_into_subgraphs (reference src/spacer_ordering.cpp:141-175 with the Tarjan search :3-76) is the
// other half of get_crispr_regions_extended_by_k (:177-184): the device compacts the valid part of the graph
// (mg_valid_subgraph: ids + out-neighbour rows, instead of the reference's walk over every item and its hash
// maps), the host runs the same depth-first search over the compact rows — same visiting order, hence the same
// components in the same order, the same node order inside each, and the same insertion sequence into each
// Graph's containers (so their iteration orders match too).
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "spacer_ordering.h"  // the reference's header: struct Graph
#include "gpu_bridge.h"

namespace {

// device graph in step with the host SDBG: the resident one (plus the host's newer invalid bits) or an upload
mg_graph *acquire(SDBG &sdbg, bool *own) {
  using namespace mcaat_gpu_bridge;
  mg_graph *g = resident_graph(&sdbg);
  *own = g == nullptr;
  if (*own) {
    g = upload(sdbg);  // carries the host's current invalid bits
  } else {
    // SetInvalidEdge calls the host made since the CycleFinder stage (none in the reference's pipeline)
    const uint64_t n = sdbg.size();
    std::vector<uint64_t> inv((n + 63) / 64 + 1, 0);
    const int64_t n_words = (int64_t)((n + 63) / 64);
#pragma omp parallel for schedule(static)
    for (int64_t wi = 0; wi < n_words; ++wi) {
      uint64_t lo = (uint64_t)wi * 64, hi = lo + 64 < n ? lo + 64 : n, w = 0;
      for (uint64_t i = lo; i < hi; ++i) w |= (uint64_t)(!sdbg.IsValidEdge(i)) << (i & 63);
      inv[wi] = w;
    }
    int rc = mg_invalid_merge(g, inv.data(), nullptr);
    if (rc != MG_OK) fail("mg_invalid_merge", rc);
  }
  return g;
}

}  // namespace

namespace mcaat_gpu {

void keep_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k, const std::vector<std::vector<uint64_t>> &cycles) {
  using namespace mcaat_gpu_bridge;
  std::vector<uint64_t> ids;
  size_t total = 0;
  for (const auto &c : cycles) total += c.size();
  ids.reserve(total);
  for (const auto &c : cycles) ids.insert(ids.end(), c.begin(), c.end());

  bool own = false;
  mg_graph *g = acquire(sdbg, &own);
  mg_region_stats st{};
  int rc = mg_keep_regions(g, ids.data(), ids.size(), (uint64_t)k, &st);
  if (rc != MG_OK) {
    if (own) mg_graph_destroy(g);
    fail("mg_keep_regions", rc);
  }
  try {
    write_back_invalid(g, sdbg);
  } catch (...) {
    if (own) mg_graph_destroy(g);
    throw;
  }
  if (own) mg_graph_destroy(g);
}

// reference find_strongly_connected_components (:54-76) + _dfs (:3-52) on the compact rows: components with
// more than one node, in completion order, nodes in stack-pop order.  Positions, not ids.
static std::vector<std::vector<uint32_t>> tarjan(const mg_subgraph &sg) {
  const uint32_t n = (uint32_t)sg.n_nodes, kNone = 0xFFFFFFFFu;
  std::vector<uint32_t> index(n, kNone), low(n, 0);
  std::vector<uint8_t> on_stack(n, 0);
  std::vector<uint32_t> stack;
  struct Frame { uint32_t v; uint64_t e; };
  std::vector<Frame> call;
  std::vector<std::vector<uint32_t>> comps;
  uint32_t counter = 0;
  for (uint32_t root = 0; root < n; ++root) {  // valid nodes in ascending id order (:60-73)
    if (index[root] != kNone) continue;
    index[root] = low[root] = counter++;
    stack.push_back(root);
    on_stack[root] = 1;
    call.push_back({root, sg.off[root]});
    while (!call.empty()) {
      Frame &f = call.back();
      const uint32_t v = f.v;
      if (f.e < sg.off[v + 1]) {
        const uint32_t w = sg.nbr[f.e++];
        if (index[w] == kNone) {  // :30-32 — the recursive call; lowlink is folded in when it returns
          index[w] = low[w] = counter++;
          stack.push_back(w);
          on_stack[w] = 1;
          call.push_back({w, sg.off[w]});
        } else if (on_stack[w]) {  // :30-32
          low[v] = std::min(low[v], index[w]);
        }
        continue;
      }
      if (low[v] == index[v]) {  // :38-51
        std::vector<uint32_t> comp;
        uint32_t w;
        do {
          w = stack.back();
          stack.pop_back();
          on_stack[w] = 0;
          comp.push_back(w);
        } while (w != v);
        if (comp.size() > 1) comps.push_back(std::move(comp));
      }
      call.pop_back();
      if (!call.empty()) low[call.back().v] = std::min(low[call.back().v], low[v]);  // :29 after the return
    }
  }
  return comps;
}

std::vector<Graph> divide_graph_into_subgraphs(const SDBG &sdbg_const) {
  using namespace mcaat_gpu_bridge;
  SDBG &sdbg = const_cast<SDBG &>(sdbg_const);  // only const getters are called
  bool own = false;
  mg_graph *g = acquire(sdbg, &own);
  mg_subgraph sg{};
  int rc = mg_valid_subgraph(g, &sg);
  if (own) mg_graph_destroy(g);
  if (rc != MG_OK) fail("mg_valid_subgraph", rc);
  std::vector<Graph> subgraphs;
  try {
    std::vector<std::vector<uint32_t>> comps = tarjan(sg);
    std::vector<uint32_t> comp_of(sg.n_nodes, 0xFFFFFFFFu);
    for (size_t c = 0; c < comps.size(); ++c)
      for (uint32_t v : comps[c]) comp_of[v] = (uint32_t)c;
    for (size_t c = 0; c < comps.size(); ++c) {  // :146-171
      Graph sub;
      for (uint32_t v : comps[c])
        for (uint64_t e = sg.off[v]; e < sg.off[v + 1]; ++e)
          if (comp_of[sg.nbr[e]] == c) sub.add_edge(sg.ids[v], sg.ids[sg.nbr[e]]);
      if (sub.nodes.size() > 0) subgraphs.push_back(std::move(sub));
    }
  } catch (...) {
    mg_subgraph_free(&sg);
    throw;
  }
  mg_subgraph_free(&sg);
  return subgraphs;
}

std::vector<Graph> get_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k,
                                                    const std::vector<std::vector<uint64_t>> &cycles) {
  mcaat_gpu::keep_crispr_regions_extended_by_k(sdbg, k, cycles);  // :182
  return mcaat_gpu::divide_graph_into_subgraphs(sdbg);            // :183
}

}  // namespace mcaat_gpu

#ifdef MCAAT_GPU_DEFINE_REFERENCE_SYMBOLS
void keep_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k, const std::vector<std::vector<uint64_t>> &cycles) {
  mcaat_gpu::keep_crispr_regions_extended_by_k(sdbg, k, cycles);
}
std::vector<Graph> divide_graph_into_subgraphs(const SDBG &sdbg) { return mcaat_gpu::divide_graph_into_subgraphs(sdbg); }
#endif
