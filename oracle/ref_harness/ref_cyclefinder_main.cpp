// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Harness around the UNMODIFIED reference CycleFinder (compiled from the sources where they
// lie under /root/reference by oracle/Makefile; nothing is copied into this repo).
//   ref_cyclefinder <sdbg_prefix> <out.dump|-> [threads=1] [thr=20] [min=27] [max=77] [low_abundance=1]
//   ref_cyclefinder --bench <sdbg_prefix> <threads> <thr> <min> <max> <max_reps> <budget_s> <timeout_s>
//       loads the graph ONCE, then constructs the reference's CycleFinder in a forked child per repetition (the
//       constructor mutates the graph; a child works on its copy-on-write image) and prints one REF_JSON line per
//       repetition until max_reps are done or budget_s seconds have passed (at least one); a repetition that runs
//       longer than timeout_s is killed and reported (with more than one thread the reference races on its result
//       map, SURVEY F5, and may never return).
// Constructs the reference's own `CycleFinder(settings)` (reference src/cycle_finder.cpp:131-138),
// times the constructor, and dumps invalid bitmap / ordered start lists / raw results.
// `ref_cyclefinder --nav <sdbg_prefix> <out.nav>` dumps, for every item of the LOADED graph,
// outdegree, indegree, outgoing ids, incoming ids and multiplicity through the reference's
// public SDBG API (libs/megahit/src/sdbg/sdbg.h:328-357,96-105).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <string>
#include <vector>
#include <signal.h>
#include <sys/resource.h>
#include <sys/wait.h>
#include <unistd.h>

#include "cycle_finder.h"
#include "settings.h"
#include "dump_format.h"

static void raise_stack_limit(char **argv) {
  // reference RecursiveReduction recurses once per peeled node (cycle_finder.cpp:359-371)
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

static int nav_dump(const char *prefix, const char *out) {
  SDBG g;
  g.LoadFromFile(prefix);
  FILE *f = fopen(out, "wb");
  if (!f) return 2;
  uint64_t n = g.size();
  fwrite("MCAATNAV", 8, 1, f);
  dump_u64(f, n);
  // per item: i8 outdeg, i8 indeg, u16 mult, 4 x u64 out, 4 x u64 in (unused = ~0)
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t o[4] = {~0ull, ~0ull, ~0ull, ~0ull}, in[4] = {~0ull, ~0ull, ~0ull, ~0ull};
    int8_t od = (int8_t)g.OutgoingEdges(i, o);
    int8_t id = (int8_t)g.IncomingEdges(i, in);
    uint16_t m = g.EdgeMultiplicity(i);
    fwrite(&od, 1, 1, f); fwrite(&id, 1, 1, f); fwrite(&m, 2, 1, f);
    fwrite(o, 8, 4, f); fwrite(in, 8, 4, f);
  }
  fclose(f);
  return 0;
}

static int bench(int argc, char **argv) {
  if (argc < 10) return 1;
  Settings s;
  s.threads = atoi(argv[3]);
  s.cycle_finder_settings.threshold_multiplicity = strtoull(argv[4], 0, 10);
  s.cycle_finder_settings.cycle_min_length = atoi(argv[5]);
  s.cycle_finder_settings.cycle_max_length = atoi(argv[6]);
  s.cycle_finder_settings.low_abundance = true;
  const int max_reps = atoi(argv[7]);
  const double budget_s = atof(argv[8]), timeout_s = atof(argv[9]);
  auto t0 = std::chrono::steady_clock::now();
  SDBG g;
  g.LoadFromFile(argv[2]);
  s.sdbg = &g;
  s.graph_folder = argv[2];
  double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  printf("REF_LOAD {\"n_items\": %llu, \"load_s\": %.6f}\n", (unsigned long long)g.size(), load_s);
  fflush(stdout);
  auto b0 = std::chrono::steady_clock::now();
  for (int rep = 0; rep < max_reps; ++rep) {
    if (rep > 0 && std::chrono::duration<double>(std::chrono::steady_clock::now() - b0).count() > budget_s) break;
    pid_t pid = fork();
    if (pid < 0) return 3;
    if (pid == 0) {
      omp_set_num_threads((int)s.threads);
      auto t1 = std::chrono::steady_clock::now();
      std::cout.setstate(std::ios_base::failbit);  // the progress lines of the constructor
      CycleFinder cf(s);
      double ctor_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
      std::map<int, std::vector<uint64_t>, std::greater<int>> chunks;
      size_t n_starts = cf.ChunkStartNodes(chunks);  // public, does not read `visited` (:387-427)
      std::cout.clear();
      size_t total_cycles = 0;
      for (auto &kv : cf.results) total_cycles += kv.second.size();
      printf("REF_JSON {\"rep\": %d, \"n_items\": %llu, \"threads\": %zu, \"ctor_s\": %.6f, \"start_nodes\": %zu, "
             "\"searched\": %zu, \"cycles\": %zu}\n", rep, (unsigned long long)g.size(), s.threads, ctor_s, n_starts,
             cf.results.size(), total_cycles);
      fflush(stdout);
      _exit(0);
    }
    auto w0 = std::chrono::steady_clock::now();
    int status = 0;
    for (;;) {
      pid_t r = waitpid(pid, &status, WNOHANG);
      if (r == pid) break;
      if (std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count() > timeout_s) {
        kill(pid, SIGKILL);
        waitpid(pid, &status, 0);
        printf("REF_FAIL {\"rep\": %d, \"why\": \"timeout\"}\n", rep);
        fflush(stdout);
        return 4;
      }
      usleep(20000);
    }
    if (!WIFEXITED(status) || WEXITSTATUS(status) != 0) {
      printf("REF_FAIL {\"rep\": %d, \"why\": \"crash\"}\n", rep);
      fflush(stdout);
      return 5;
    }
  }
  return 0;
}

int main(int argc, char **argv) {
  raise_stack_limit(argv);
  if (argc >= 4 && !strcmp(argv[1], "--nav")) return nav_dump(argv[2], argv[3]);
  if (argc >= 3 && !strcmp(argv[1], "--bench")) return bench(argc, argv);
  if (argc < 3) {
    fprintf(stderr, "usage: %s <sdbg_prefix> <out.dump|-> [threads thr min max low_abundance]\n", argv[0]);
    return 1;
  }
  Settings s;
  s.threads = argc > 3 ? atoi(argv[3]) : 1;
  s.cycle_finder_settings.threshold_multiplicity = argc > 4 ? strtoull(argv[4], 0, 10) : 20;
  s.cycle_finder_settings.cycle_min_length = argc > 5 ? atoi(argv[5]) : 27;
  s.cycle_finder_settings.cycle_max_length = argc > 6 ? atoi(argv[6]) : 77;
  s.cycle_finder_settings.low_abundance = argc > 7 ? atoi(argv[7]) != 0 : true;
  omp_set_num_threads((int)s.threads);

  auto t0 = std::chrono::steady_clock::now();
  SDBG g;
  g.LoadFromFile(argv[1]);
  s.sdbg = &g;
  s.graph_folder = argv[1];  // the file prefix, as reference src/main.cpp:522-530 leaves it
  auto t1 = std::chrono::steady_clock::now();
  CycleFinder cf(s);
  auto t2 = std::chrono::steady_clock::now();
  double load_s = std::chrono::duration<double>(t1 - t0).count();
  double ctor_s = std::chrono::duration<double>(t2 - t1).count();

  // ordered start lists: ChunkStartNodes is public and does not read `visited`
  // (cycle_finder.cpp:387-427), so re-running it on the pruned graph reproduces the lists.
  std::map<int, std::vector<uint64_t>, std::greater<int>> chunks;
  size_t n_starts = 0;
  if (strcmp(argv[2], "-") != 0 || getenv("MCAAT_REF_COUNT_STARTS")) {
    std::cout.setstate(std::ios_base::failbit);
    n_starts = cf.ChunkStartNodes(chunks);
    std::cout.clear();
    if (s.threads > 1) for (auto &kv : chunks) std::sort(kv.second.begin(), kv.second.end());
  }
  size_t total_cycles = 0;
  for (auto &kv : cf.results) total_cycles += kv.second.size();
  printf("REF_JSON {\"n_items\": %llu, \"threads\": %zu, \"load_s\": %.6f, \"ctor_s\": %.6f, "
         "\"start_nodes\": %zu, \"searched\": %zu, \"cycles\": %zu}\n",
         (unsigned long long)g.size(), s.threads, load_s, ctor_s, n_starts, cf.results.size(), total_cycles);
  if (!strcmp(argv[2], "-")) return 0;

  FILE *f = fopen(argv[2], "wb");
  if (!f) return 2;
  fwrite("MCAATDMP", 8, 1, f);
  uint64_t n = g.size();
  dump_u64(f, n);
  std::vector<uint64_t> inv((n + 63) / 64, 0);
  for (uint64_t i = 0; i < n; ++i)
    if (!g.IsValidEdge(i)) inv[i >> 6] |= 1ull << (i & 63);
  dump_vec(f, inv);
  dump_u64(f, chunks.size());
  for (auto &kv : chunks) {
    int64_t key = kv.first;
    fwrite(&key, 8, 1, f);
    dump_vec(f, kv.second);
  }
  std::vector<uint64_t> searched;
  for (auto &kv : chunks)
    for (uint64_t v : kv.second)
      if (cf.results.count(v)) searched.push_back(v);
  dump_u64(f, searched.size());
  for (uint64_t v : searched) {
    auto &cyc = cf.results[v];
    dump_u64(f, v);
    dump_u64(f, cyc.size());
    for (auto &c : cyc) dump_vec(f, c);
  }
  std::vector<uint64_t> iter_order;
  for (auto &kv : cf.results) iter_order.push_back(kv.first);
  dump_vec(f, iter_order);
  fclose(f);
  return 0;
}
