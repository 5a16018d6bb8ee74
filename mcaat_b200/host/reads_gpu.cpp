// Drop-in for the reference's get_reads (reference src/reads.cpp:91-130, declared include/reads.h:70-75): same
// arguments, same result — for every FASTQ sequence that the reference keeps (longer than 2k and starting or
// ending on a cycle node, :64-77) the node ids of all its k-mers, file 1 in order and then file 2 reverse-
// complemented (:119-121).  The k-mer lookups (SDBG::IndexBinarySearch, sdbg.h:141-207) run on the GPU
// (mg_map_reads, include/mcaat_gpu.h); FASTQ parsing stays the reference's own extract_sequences_from_fastq_file
// (:3-18, kseq++).  No CPU fallback: a C-ABI error becomes std::runtime_error.
//
// The lookup needs the tip labels and the bucket table, which the SDBG class keeps private.  They are on the
// device when the graph was decoded from the graph files (MCAAT_GPU_INGEST_FILES=1 for the CycleFinder drop-in,
// whose resident graph is reused here); otherwise MCAAT_GPU_GRAPH_PREFIX must name those files
// (settings.graph_folder after src/main.cpp:522-530) and the graph is read from them for this call.
//
// Adoption: call mcaat_gpu::get_reads where the reference calls get_reads
// (src/main_run_and_debug.cpp:11-16), or delete src/reads.cpp:91-130 and compile this file with
// -DMCAAT_GPU_DEFINE_REFERENCE_SYMBOLS.
#include <cstdint>
#include <cstdlib>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "reads.h"  // the reference's header
#include "gpu_bridge.h"

namespace mcaat_gpu {

// per-sequence result (empty = the reference drops the read), one batch = one FASTQ file
std::vector<std::vector<uint64_t>> map_sequences_all(const SDBG &sdbg, const std::vector<std::string> &seqs, bool reverse,
                                                     const std::vector<std::vector<uint64_t>> &cycles) {
  using namespace mcaat_gpu_bridge;
  mg_graph *g = resident_graph(&sdbg);
  bool own = false;
  if (!g || !mg_graph_has_kmer_index(g)) {
    const char *prefix = std::getenv("MCAAT_GPU_GRAPH_PREFIX");
    if (!prefix || !*prefix)
      throw std::runtime_error("libmcaat_gpu: read mapping needs the graph's tip labels and bucket table; run the CycleFinder "
                               "stage with MCAAT_GPU_INGEST_FILES=1 or set MCAAT_GPU_GRAPH_PREFIX to the graph file prefix");
    const char *dev = std::getenv("MCAAT_GPU_DEVICE");
    int rc = mg_graph_create_from_files(prefix, dev ? std::atoi(dev) : 0, &g);
    if (rc != MG_OK) fail("mg_graph_create_from_files", rc);
    own = true;
    if (mg_graph_size(g) != (uint64_t)sdbg.size()) {
      mg_graph_destroy(g);
      throw std::runtime_error(std::string("libmcaat_gpu: graph files at '") + prefix + "' do not match the loaded SDBG");
    }
  }
  std::vector<uint64_t> ids;
  for (const auto &c : cycles) ids.insert(ids.end(), c.begin(), c.end());
  std::vector<std::vector<uint64_t>> out(seqs.size());
  // batches of <= 1 GiB of characters (the C ABI takes < 2^32 per call)
  size_t at = 0;
  while (at < seqs.size()) {
    std::string blob;
    std::vector<uint64_t> off{0};
    size_t end = at;
    while (end < seqs.size() && (end == at || blob.size() + seqs[end].size() <= (1ull << 30))) {
      blob += seqs[end];
      off.push_back(blob.size());
      ++end;
    }
    mg_reads r{};
    int rc = mg_map_reads(g, blob.data(), off.data(), end - at, reverse ? 1 : 0, ids.data(), ids.size(), &r);
    if (rc != MG_OK) {
      if (own) mg_graph_destroy(g);
      fail("mg_map_reads", rc);
    }
    for (size_t i = 0; i < end - at; ++i) out[at + i].assign(r.ids + r.off[i], r.ids + r.off[i + 1]);
    mg_reads_free(&r);
    at = end;
  }
  if (own) mg_graph_destroy(g);
  return out;
}

// reference get_reads, :91-130
std::vector<std::vector<uint64_t>> get_reads(const SDBG &sdbg, const std::string &fastq_file_1,
                                             const std::optional<std::string> &fastq_file_2,
                                             const std::vector<std::vector<uint64_t>> cycles) {
  std::vector<std::vector<uint64_t>> reads;
  auto append = [&](const std::string &path, bool reverse) {
    std::vector<std::string> sequences = extract_sequences_from_fastq_file(path);  // the reference's parser
    for (auto &read : map_sequences_all(sdbg, sequences, reverse, cycles))
      if (!read.empty()) reads.push_back(std::move(read));  // :110-112, :123-125
  };
  append(fastq_file_1, false);
  if (fastq_file_2.has_value()) append(fastq_file_2.value(), true);
  return reads;
}

}  // namespace mcaat_gpu

#ifdef MCAAT_GPU_DEFINE_REFERENCE_SYMBOLS
vector<vector<uint64_t>> get_reads(const SDBG &sdbg, const string &fastq_file_1, const optional<string> &fastq_file_2,
                                   const vector<vector<uint64_t>> cycles) {
  return mcaat_gpu::get_reads(sdbg, fastq_file_1, fastq_file_2, cycles);
}
#endif
