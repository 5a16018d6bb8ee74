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
// Two ways to adopt it: call mcaat_gpu::keep_crispr_regions_extended_by_k at
// src/spacer_ordering.cpp:182, or delete the reference's definition and compile this file with
// -DMCAAT_GPU_DEFINE_REFERENCE_SYMBOLS, which also defines the reference's own symbol.
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "gpu_bridge.h"

namespace mcaat_gpu {

void keep_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k, const std::vector<std::vector<uint64_t>> &cycles) {
  using namespace mcaat_gpu_bridge;
  std::vector<uint64_t> ids;
  size_t total = 0;
  for (const auto &c : cycles) total += c.size();
  ids.reserve(total);
  for (const auto &c : cycles) ids.insert(ids.end(), c.begin(), c.end());

  mg_graph *g = resident_graph(&sdbg);
  const bool own = g == nullptr;
  if (own) {
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

}  // namespace mcaat_gpu

#ifdef MCAAT_GPU_DEFINE_REFERENCE_SYMBOLS
void keep_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k, const std::vector<std::vector<uint64_t>> &cycles) {
  mcaat_gpu::keep_crispr_regions_extended_by_k(sdbg, k, cycles);
}
#endif
