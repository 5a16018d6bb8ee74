// TEST INFRASTRUCTURE (oracle side). Binary dump written by ref_cyclefinder_main.cpp
// (the UNMODIFIED reference CycleFinder) and by the product's drop-in harness, so the
// two can be compared byte for byte.
//
// Layout (all little-endian u64 unless noted):
//   magic "MCAATDMP"            8 bytes
//   n_items
//   n_words, invalid bitmap words (state of SDBG::IsValidEdge after the constructor)
//   n_buckets, then per bucket (descending key): i64 key, n_ids, ids[]   (ChunkStartNodes order)
//   n_searched, then per searched start node, in search order:
//       start, n_cycles, then per cycle: len, ids[]
//   n_iter, start ids in std::unordered_map iteration order of `results`
//       (what cycles_map_to_cycles observes, reference src/tmp_utils.cpp:26-38)
#pragma once
#include <cstdint>
#include <cstdio>
#include <vector>
static inline void dump_u64(FILE *f, uint64_t v) { fwrite(&v, 8, 1, f); }
static inline void dump_vec(FILE *f, const std::vector<uint64_t> &v) {
  dump_u64(f, v.size());
  if (!v.empty()) fwrite(v.data(), 8, v.size(), f);
}
