// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Harness around the UNMODIFIED reference keep_crispr_regions_extended_by_k
// (reference src/spacer_ordering.cpp:78-139), compiled from the source where it lies by
// oracle/Makefile.  The translation unit also includes two un-vendored third-party headers
// (kseq++, the cft set-cover library); oracle/ref_harness/stubs/ holds declarations-only stand-ins
// so that it compiles — none of that code is reached here.
//   ref_regions <sdbg_prefix> <out.reg> <k> cycles [thr=20] [min=27] [max=77]
//       cycles = every cycle the reference CycleFinder (threads=1) finds, as src/main.cpp:536-556
//   ref_regions <sdbg_prefix> <out.reg> <k> nodes <file of little-endian u64 ids>
//       one "cycle" holding the listed ids, on the graph as loaded (nothing pruned)
// Output: magic "MCAATREG", n_items, k, n_seed_ids, n_words, invalid bitmap words after the call.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <vector>

#include "cycle_finder.h"
#include "settings.h"
#include "spacer_ordering.h"
#include "dump_format.h"

void keep_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k, const vector<vector<uint64_t>> &cycles);

int main(int argc, char **argv) {
  if (argc < 5) {
    fprintf(stderr, "usage: %s <sdbg_prefix> <out.reg> <k> cycles [thr min max] | nodes <ids.u64>\n", argv[0]);
    return 1;
  }
  size_t k = strtoull(argv[3], 0, 10);
  SDBG g;
  g.LoadFromFile(argv[1]);
  std::vector<std::vector<uint64_t>> cycles;
  if (!strcmp(argv[4], "cycles")) {
    Settings s;
    s.threads = 1;
    s.cycle_finder_settings.threshold_multiplicity = argc > 5 ? strtoull(argv[5], 0, 10) : 20;
    s.cycle_finder_settings.cycle_min_length = argc > 6 ? atoi(argv[6]) : 27;
    s.cycle_finder_settings.cycle_max_length = argc > 7 ? atoi(argv[7]) : 77;
    s.cycle_finder_settings.low_abundance = true;
    omp_set_num_threads(1);
    s.sdbg = &g;
    s.graph_folder = argv[1];
    std::cout.setstate(std::ios_base::failbit);
    CycleFinder cf(s);
    std::cout.clear();
    for (auto &kv : cf.results)
      for (auto &c : kv.second) cycles.push_back(c);
  } else if (!strcmp(argv[4], "nodes") && argc > 5) {
    FILE *f = fopen(argv[5], "rb");
    if (!f) return 2;
    std::vector<uint64_t> ids;
    uint64_t v;
    while (fread(&v, 8, 1, f) == 1) ids.push_back(v);
    fclose(f);
    cycles.push_back(ids);
  } else {
    return 1;
  }
  uint64_t n_seed = 0;
  for (auto &c : cycles) n_seed += c.size();
  keep_crispr_regions_extended_by_k(g, k, cycles);

  FILE *f = fopen(argv[2], "wb");
  if (!f) return 2;
  fwrite("MCAATREG", 8, 1, f);
  uint64_t n = g.size();
  dump_u64(f, n);
  dump_u64(f, k);
  dump_u64(f, n_seed);
  std::vector<uint64_t> inv((n + 63) / 64, 0);
  uint64_t kept = 0;
  for (uint64_t i = 0; i < n; ++i) {
    if (!g.IsValidEdge(i)) inv[i >> 6] |= 1ull << (i & 63);
    else ++kept;
  }
  dump_vec(f, inv);
  fclose(f);
  printf("REF_JSON {\"n_items\": %llu, \"k\": %zu, \"seed_ids\": %llu, \"kept\": %llu}\n", (unsigned long long)n, k,
         (unsigned long long)n_seed, (unsigned long long)kept);
  return 0;
}
