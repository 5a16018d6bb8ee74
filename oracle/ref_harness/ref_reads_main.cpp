// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Harness around the UNMODIFIED reference read mapping (reference src/reads.cpp:20-89, compiled from the source
// where it lies by oracle/Makefile; its un-vendored kseq++ include is satisfied by a declarations-only stub).
//   ref_reads <sdbg_prefix> <out.rd> <sequences.txt> <cycle_nodes.u64> [reverse=0]
// sequences.txt: one sequence per line (what extract_sequences_from_fastq_file would return, src/reads.cpp:3-18);
// cycle_nodes.u64: little-endian ids forming nodes_of_cycles (src/reads.cpp:98-103); reverse=1 applies the
// reference's reverse_pair_ends_sequence first, as get_reads does for the second file (:119-121).
// Per sequence the reference's get_read_from_sequence(sdbg, nodes_of_cycles, sequence) is called (:57-89).
// Output: magic "MCAATRDS", n_sequences, then per sequence: n_ids, ids[] (empty = the read is dropped, :110-112).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <unordered_set>
#include <vector>

#include "reads.h"
#include "dump_format.h"

#ifdef MCAAT_READS_VIA_GPU_DROP_IN  // mcaat_b200/host/Makefile: the same main over mcaat_b200/host/reads_gpu.cpp
namespace mcaat_gpu {
std::vector<std::vector<uint64_t>> map_sequences_all(const SDBG &sdbg, const std::vector<std::string> &seqs, bool reverse,
                                                     const std::vector<std::vector<uint64_t>> &cycles);
}
#endif

int main(int argc, char **argv) {
  if (argc < 5) {
    fprintf(stderr, "usage: %s <sdbg_prefix> <out.rd> <sequences.txt> <cycle_nodes.u64> [reverse]\n", argv[0]);
    return 1;
  }
  const bool reverse = argc > 5 && atoi(argv[5]) != 0;
  SDBG g;
  g.LoadFromFile(argv[1]);
  std::unordered_set<uint64_t> nodes_of_cycles;
  {
    FILE *f = fopen(argv[4], "rb");
    if (!f) return 2;
    uint64_t v;
    while (fread(&v, 8, 1, f) == 1) nodes_of_cycles.insert(v);
    fclose(f);
  }
  std::ifstream in(argv[3]);
  if (!in) return 2;
  std::vector<std::string> seqs;
  for (std::string line; std::getline(in, line);) seqs.push_back(line);
  FILE *f = fopen(argv[2], "wb");
  if (!f) return 2;
  fwrite("MCAATRDS", 8, 1, f);
  dump_u64(f, seqs.size());
  const size_t n_seqs = seqs.size();
  uint64_t kept = 0, ids = 0;
#ifdef MCAAT_READS_VIA_GPU_DROP_IN
  {
    std::vector<std::vector<uint64_t>> cyc(1, std::vector<uint64_t>(nodes_of_cycles.begin(), nodes_of_cycles.end()));
    for (auto &read : mcaat_gpu::map_sequences_all(g, seqs, reverse, cyc)) {
      dump_vec(f, read);
      kept += !read.empty();
      ids += read.size();
    }
    seqs.clear();
  }
#endif
  for (auto &s : seqs) {
    std::string q = s;
    if (reverse) reverse_pair_ends_sequence(q);
    std::vector<uint64_t> read = get_read_from_sequence(g, nodes_of_cycles, q);
    dump_vec(f, read);
    kept += !read.empty();
    ids += read.size();
  }
  fclose(f);
  printf("REF_JSON {\"n_items\": %llu, \"sequences\": %zu, \"kept\": %llu, \"ids\": %llu}\n", (unsigned long long)g.size(),
         n_seqs, (unsigned long long)kept, (unsigned long long)ids);
  return 0;
}
