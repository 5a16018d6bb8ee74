// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// The reference's release pipeline from the loaded graph to CRISPR_Arrays.txt, as src/main.cpp:520-580 runs it
// (graph construction and argument parsing left out: the graph files and the read files are given):
//   SDBG::LoadFromFile -> CycleFinder(settings) -> cycles_map_to_cycles -> run_and_debug_finding_of_relevant_reads
//   -> run_and_debug_spacer_ordering -> run_and_debug_results -> CRISPRAnalyzer::run_analysis -> <out_file>
// Every stage is the reference's own, unmodified translation unit (reads.cpp, spacer_ordering.cpp, tmp_utils.cpp,
// main_run_and_debug.cpp, evaluation.cpp, post_processing.h).  Linked once with the reference's src/cycle_finder.cpp
// (oracle/_ref/ref_pipeline) and once with the GPU drop-in (mcaat_b200/host/_build/shim_pipeline): the two output
// files must be identical.  The two third-party libraries this image lacks (kseq++, cft) are the
// deterministic stand-ins under mcaat_b200/host/absent_deps, the same for both builds.
//   ref_pipeline <sdbg_prefix> <reads_1.fastq> <reads_2.fastq|-> <out_file> [threads=1 thr=20 min=27 max=77]
#include <sys/resource.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "cycle_finder.h"
#include "main_run_and_debug.h"
#include "post_processing.h"
#include "settings.h"
#include "tmp_utils.h"

int main(int argc, char **argv) {
  {  // reference RecursiveReduction recurses once per peeled node (cycle_finder.cpp:359-371)
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
  if (argc < 5) {
    fprintf(stderr, "usage: %s <sdbg_prefix> <reads_1> <reads_2|-> <out_file> [threads thr min max]\n", argv[0]);
    return 1;
  }
  Settings settings;
  settings.input_files = std::string(argv[2]) + (strcmp(argv[3], "-") ? std::string(" ") + argv[3] : std::string());
  settings.output_file = argv[4];
  settings.threads = argc > 5 ? atoi(argv[5]) : 1;
  settings.cycle_finder_settings.threshold_multiplicity = argc > 6 ? strtoull(argv[6], 0, 10) : 20;
  settings.cycle_finder_settings.cycle_min_length = argc > 7 ? atoi(argv[7]) : 27;
  settings.cycle_finder_settings.cycle_max_length = argc > 8 ? atoi(argv[8]) : 77;
  omp_set_num_threads((int)settings.threads);
  // main.cpp:520-530
  SDBG sdbg;
  settings.graph_folder = argv[1];
  sdbg.LoadFromFile(argv[1]);
  std::cout << "Loaded the graph" << std::endl;
  settings.sdbg = &sdbg;
  // main.cpp:533-542
  std::cout << "FBCE START:" << std::endl;
  CycleFinder cycle_finder(settings);
  auto cycles_map = cycle_finder.results;
  std::cout << "Number of nodes in results: " << cycles_map.size() << std::endl;
  auto cycles = cycles_map_to_cycles(cycles_map);
  // main.cpp:544-556
  const auto reads = run_and_debug_finding_of_relevant_reads(cycles, settings, sdbg);
  std::vector<std::tuple<std::string, std::string, std::vector<std::string>, float, float>> found_systems;
  if (!reads.empty()) found_systems = run_and_debug_spacer_ordering(reads, sdbg, cycles);  // (reads.at(0) throws on an empty set)
  // main.cpp:564-569
  run_and_debug_results(found_systems);
  // main.cpp:572-580
  std::cout << "POST PROCESSING START:" << std::endl;
  std::unordered_map<std::string, std::vector<std::string>> all_systems;
  for (const auto &[_sequence, repeat, spacers, _conf_a, _conf_b] : found_systems) all_systems[repeat] = spacers;
  CRISPRAnalyzer analyzer(all_systems, settings.output_file);
  analyzer.run_analysis();
  std::cout << "Saved in: " << settings.output_file << std::endl;
  printf("PIPELINE_JSON {\"cycles\": %zu, \"reads\": %zu, \"systems\": %zu}\n", cycles.size(), reads.size(), found_systems.size());
  return 0;
}
