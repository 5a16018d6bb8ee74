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
// With `cycles+scc` / `nodes+scc` the reference's divide_graph_into_subgraphs (src/spacer_ordering.cpp:141-175,
// Tarjan :3-76) runs next, as in get_crispr_regions_extended_by_k (:177-184), and the output continues with
// magic "MCAATSCC", n_subgraphs, then per subgraph in the returned order: n_nodes, `nodes` in its
// unordered_set iteration order, n_adj, and per `adjacency_list` entry in its unordered_map iteration order:
// from, n_to, to[].
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "cycle_finder.h"
#include "settings.h"
#include "spacer_ordering.h"
#include "dump_format.h"

#include <sys/resource.h>
#include <unistd.h>

void keep_crispr_regions_extended_by_k(SDBG &sdbg, const size_t &k, const vector<vector<uint64_t>> &cycles);
vector<Graph> divide_graph_into_subgraphs(const SDBG &sdbg);

static void raise_stack_limit(char **argv) {  // the reference's Tarjan and RecursiveReduction recurse per node
  struct rlimit rl;
  getrlimit(RLIMIT_STACK, &rl);
  if (rl.rlim_cur != RLIM_INFINITY && !getenv("MCAAT_REF_REEXEC")) {
    rl.rlim_cur = rl.rlim_max;
    if (setrlimit(RLIMIT_STACK, &rl) == 0) {
      setenv("MCAAT_REF_REEXEC", "1", 1);
      execv("/proc/self/exe", argv);
    }
  }
}

int main(int argc, char **argv) {
  raise_stack_limit(argv);
  if (argc < 5) {
    fprintf(stderr, "usage: %s <sdbg_prefix> <out.reg> <k> cycles[+scc] [thr min max] | nodes[+scc] <ids.u64>\n", argv[0]);
    return 1;
  }
  size_t k = strtoull(argv[3], 0, 10);
  SDBG g;
  g.LoadFromFile(argv[1]);
  std::vector<std::vector<uint64_t>> cycles;
  std::string mode = argv[4];
  const bool scc = mode.size() > 4 && mode.compare(mode.size() - 4, 4, "+scc") == 0;
  if (scc) mode.resize(mode.size() - 4);
  if (mode == "cycles") {
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
  } else if (mode == "nodes" && argc > 5) {
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
  uint64_t n_sub = 0;
  if (scc) {
    std::cout.setstate(std::ios_base::failbit);
    std::vector<Graph> subs = divide_graph_into_subgraphs(g);
    std::cout.clear();
    n_sub = subs.size();
    fwrite("MCAATSCC", 8, 1, f);
    dump_u64(f, subs.size());
    for (auto &sg : subs) {
      dump_vec(f, std::vector<uint64_t>(sg.nodes.begin(), sg.nodes.end()));
      dump_u64(f, sg.adjacency_list.size());
      for (auto &kv : sg.adjacency_list) {
        dump_u64(f, kv.first);
        dump_vec(f, kv.second);
      }
    }
  }
  fclose(f);
  printf("REF_JSON {\"n_items\": %llu, \"k\": %zu, \"seed_ids\": %llu, \"kept\": %llu, \"subgraphs\": %llu}\n",
         (unsigned long long)n, k, (unsigned long long)n_seed, (unsigned long long)kept, (unsigned long long)n_sub);
  return 0;
}
