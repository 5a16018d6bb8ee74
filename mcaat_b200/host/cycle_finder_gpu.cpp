// Drop-in replacement for the reference's src/cycle_finder.cpp: same class, same header
// (reference include/cycle_finder.h:25-85, compiled UNCHANGED from where it lies), same observable
// behaviour — `CycleFinder cf(settings);` runs the whole search (reference
// src/cycle_finder.cpp:131-138), `cf.results` holds one entry per searched start node inserted in
// the reference's threads=1 order with the raw per-start cycle sequence, the host SDBG's invalid
// bits are updated (SetInvalidEdge) and the same progress lines go to std::cout.
// Everything between is libmcaat_gpu.so (include/mcaat_gpu.h): the graph is extracted once through
// the public SDBG getters, uploaded, and searched on the GPU.  No CPU fallback: a C-ABI error
// becomes std::runtime_error.
//
// Build: see mcaat_b200/host/Makefile (needs the reference's headers on the include path; nothing
// of the reference is copied).  Device: env MCAAT_GPU_DEVICE (default 0).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include <chrono>

#include "cycle_finder.h"  // the reference's header
#include "mcaat_gpu.h"
#include "gpu_bridge.h"

namespace {

// wall-clock breakdown of the constructor, printed to stderr as one "GPU_SHIM_TIMING {...}" line when MCAAT_GPU_TIMING=1
struct Timing {
  double graph_ms = 0, prune_ms = 0, writeback_ms = 0, starts_ms = 0, cycles_ms = 0, results_ms = 0;
  const char *graph_path = "";
};
Timing g_timing;
double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
struct GpuState {
  mg_graph *g = nullptr;
  bool pruned = false;
  mg_prune_stats stats{};
  const SDBG *sdbg = nullptr;
};
std::mutex g_mu;
std::unordered_map<const CycleFinder *, GpuState> g_state;  // the header has no room for a handle

}  // namespace

namespace mcaat_gpu_bridge {

[[noreturn]] void fail(const char *what, int rc) {
  throw std::runtime_error(std::string("libmcaat_gpu: ") + what + " failed (" + std::to_string(rc) +
                           "): " + mg_last_error());
}

mg_graph *resident_graph(const SDBG *sdbg) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto &kv : g_state)
    if (kv.second.sdbg == sdbg && kv.second.g) return kv.second.g;
  return nullptr;
}

}  // namespace mcaat_gpu_bridge

namespace {
using mcaat_gpu_bridge::fail;

// f_[0..5] (reference sdbg.h:36-50) is private; LastCharOf (sdbg.h:79-86, public) is monotone in
// its argument, so f_[i] = the smallest x with LastCharOf(x) >= i.
void recover_f(const SDBG &sdbg, int64_t f[6]) {
  const uint64_t n = sdbg.size();
  f[0] = -1;
  for (int i = 1; i <= 5; ++i) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
      uint64_t mid = lo + (hi - lo) / 2;
      if (sdbg.LastCharOf(mid) >= i) hi = mid;
      else lo = mid + 1;
    }
    f[i] = (int64_t)lo;
  }
}

}  // namespace

namespace mcaat_gpu_bridge {
// Extract MEGAHIT's arrays through the public getters (reference sdbg.h:63-105) and upload them.
mg_graph *upload(SDBG &sdbg) {
  const uint64_t n = sdbg.size();
  std::vector<uint64_t> w((n + 15) / 16 + 1, 0), last((n + 63) / 64 + 1, 0), tip((n + 63) / 64 + 1, 0),
      inv((n + 63) / 64 + 1, 0);
  std::vector<uint16_t> mult(n + 1, 0);
  const int64_t n_words = (int64_t)((n + 63) / 64);
#pragma omp parallel for schedule(static)
  for (int64_t wi = 0; wi < n_words; ++wi) {  // one thread owns whole 64-item words: no write sharing
    uint64_t lo = (uint64_t)wi * 64, hi = std::min<uint64_t>(lo + 64, n);
    uint64_t lw = 0, tw = 0, iw = 0;
    for (uint64_t i = lo; i < hi; ++i) {
      w[i >> 4] |= (uint64_t)(sdbg.GetW(i) & 15) << ((i & 15) * 4);
      lw |= (uint64_t)sdbg.IsLast(i) << (i & 63);
      tw |= (uint64_t)sdbg.IsTip(i) << (i & 63);
      iw |= (uint64_t)(!sdbg.IsValidEdge(i)) << (i & 63);
      mult[i] = (uint16_t)sdbg.EdgeMultiplicity(i);
    }
    last[wi] = lw;
    tip[wi] = tw;
    inv[wi] = iw;
  }
  mg_graph_desc d{};
  d.n_items = n;
  d.k = sdbg.k();
  d.memory = MG_MEM_HOST;
  const char *dev = std::getenv("MCAAT_GPU_DEVICE");
  d.device = dev ? std::atoi(dev) : 0;
  d.w = w.data();
  d.last = last.data();
  d.tip = tip.data();
  d.invalid = inv.data();
  d.mult16 = mult.data();
  recover_f(sdbg, d.f);
  mg_graph *g = nullptr;
  int rc = mg_graph_create(&d, &g);
  if (rc != MG_OK) fail("mg_graph_create", rc);
  return g;
}
}  // namespace mcaat_gpu_bridge

namespace {
using mcaat_gpu_bridge::upload;

// Fast path, taken whenever settings.graph_folder names graph files (MCAAT_GPU_INGEST_FILES=0 turns it off): the files
// the host loaded the SDBG from are decoded on the device (mg_graph_create_from_files) instead of extracting every
// item through the getters.  `prefix` = settings.graph_folder, which the reference's main has turned into the file
// prefix before constructing CycleFinder (reference src/main.cpp:522-530).  The files give the LOADED graph; whatever
// the host invalidated since SDBG::LoadFromFile is ORed in afterwards (mg_invalid_merge), so the device graph equals
// the host's current state in every case.
bool graph_files_exist(const std::string &prefix) {
  if (prefix.empty()) return false;
  FILE *f = std::fopen((prefix + ".sdbg_info").c_str(), "rb");
  if (!f) return false;
  std::fclose(f);
  return true;
}
mg_graph *ingest(const std::string &prefix, SDBG &sdbg) {
  const char *dev = std::getenv("MCAAT_GPU_DEVICE");
  mg_graph *g = nullptr;
  int rc = mg_graph_create_from_files(prefix.c_str(), dev ? std::atoi(dev) : 0, &g);
  if (rc != MG_OK) fail("mg_graph_create_from_files", rc);
  if (mg_graph_size(g) == (uint64_t)sdbg.size()) {
    const uint64_t n = sdbg.size();
    const int64_t n_words = (int64_t)((n + 63) / 64);
    std::vector<uint64_t> inv((size_t)n_words + 1, 0);
#pragma omp parallel for schedule(static)
    for (int64_t wi = 0; wi < n_words; ++wi) {
      uint64_t lo = (uint64_t)wi * 64, hi = std::min<uint64_t>(lo + 64, n), iw = 0;
      for (uint64_t i = lo; i < hi; ++i) iw |= (uint64_t)(!sdbg.IsValidEdge(i)) << (i & 63);
      inv[wi] = iw;
    }
    rc = mg_invalid_merge(g, inv.data(), nullptr);
    if (rc != MG_OK) fail("mg_invalid_merge", rc);
  }
  if (mg_graph_size(g) != (uint64_t)sdbg.size()) {
    mg_graph_destroy(g);
    throw std::runtime_error("libmcaat_gpu: graph files at '" + prefix + "' do not match the loaded SDBG");
  }
  return g;
}

GpuState &state_of(const CycleFinder *cf, SDBG &sdbg, const std::string &prefix = std::string()) {
  std::lock_guard<std::mutex> lk(g_mu);
  GpuState &s = g_state[cf];
  if (!s.g) {
    const char *fast = std::getenv("MCAAT_GPU_INGEST_FILES");
    const bool from_files = !(fast && *fast == '0') && graph_files_exist(prefix);
    const double t0 = now_ms();
    s.g = from_files ? ingest(prefix, sdbg) : upload(sdbg);
    g_timing.graph_ms = now_ms() - t0;
    g_timing.graph_path = from_files ? "device ingest of the graph files" : "SDBG getters + upload";
    s.sdbg = &sdbg;
  }
  return s;
}

}  // namespace

namespace mcaat_gpu_bridge {
// invalid bits the GPU set -> host SDBG (reference SetInvalidEdge, sdbg.h:94)
void write_back_invalid(mg_graph *g, SDBG &sdbg) {
  const uint64_t n = sdbg.size();
  std::vector<uint64_t> inv((n + 63) / 64 + 1, 0);
  int rc = mg_invalid_readback(g, inv.data());
  if (rc != MG_OK) fail("mg_invalid_readback", rc);
  const int64_t n_words = (int64_t)((n + 63) / 64);
#pragma omp parallel for schedule(static)
  for (int64_t wi = 0; wi < n_words; ++wi) {
    uint64_t word = inv[wi];
    while (word) {
      uint64_t i = (uint64_t)wi * 64 + __builtin_ctzll(word);
      word &= word - 1;
      if (i < n && sdbg.IsValidEdge(i)) sdbg.SetInvalidEdge(i);
    }
  }
}
}  // namespace mcaat_gpu_bridge

namespace {
using mcaat_gpu_bridge::write_back_invalid;

void ensure_pruned(GpuState &s, SDBG &sdbg, mg_prune_stats *out) {
  if (!s.pruned) {
    const double t0 = now_ms();
    int rc = mg_prune(s.g, &s.stats);
    if (rc != MG_OK) fail("mg_prune", rc);
    s.pruned = true;
    const double t1 = now_ms();
    write_back_invalid(s.g, sdbg);
    g_timing.prune_ms = t1 - t0;
    g_timing.writeback_ms = now_ms() - t1;
  }
  if (out) *out = s.stats;
}

}  // namespace

// ---- reference src/cycle_finder.cpp:131-138
CycleFinder::CycleFinder(Settings &settings) : settings(settings), cluster_bounds(500) {
  if (this->settings.sdbg == nullptr) {
    throw std::runtime_error("CycleFinder requires settings.sdbg to be set to a valid SDBG instance");
  }
  this->FindApproximateCRISPRArrays();
}

CycleFinder::~CycleFinder() {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_state.find(this);
  if (it != g_state.end()) {
    mg_graph_destroy(it->second.g);
    g_state.erase(it);
  }
}

// ---- reference :346-357.  Evaluated on the CURRENT graph state, like the reference.
vector<uint64_t> CycleFinder::CollectTips() {
  SDBG &sdbg = *this->settings.sdbg;
  GpuState &s = state_of(this, sdbg, this->settings.graph_folder);
  uint64_t *ids = nullptr, n_tips = 0;
  int rc = mg_collect_tips(s.g, &ids, &n_tips);  // found and compacted on the device; only the ids cross the bus
  if (rc != MG_OK) fail("mg_collect_tips", rc);
  vector<uint64_t> tips(ids, ids + n_tips);
  mg_free(ids);
  return tips;
}

// ---- reference :359-371 / :372-382.  The reduction is a whole-graph closure on the device
// (mg_prune = CollectTips -> InvalidateMultiplicityOneNodes -> RecursiveReduction of every tip);
// the per-tip and per-phase entry points run that pipeline once and are idempotent afterwards.
void CycleFinder::RecursiveReduction(uint64_t) {
  SDBG &sdbg = *this->settings.sdbg;
  ensure_pruned(state_of(this, sdbg, this->settings.graph_folder), sdbg, nullptr);
}
void CycleFinder::InvalidateMultiplicityOneNodes() {
  SDBG &sdbg = *this->settings.sdbg;
  GpuState &s = state_of(this, sdbg, this->settings.graph_folder);
  bool first = !s.pruned;
  mg_prune_stats st{};
  ensure_pruned(s, sdbg, &st);
  if (first)
    std::cout << "Pre-filter: invalidated " << st.n_mult1 << " node(s) with multiplicity <= 1." << std::endl;
}

// ---- reference :387-427
size_t CycleFinder::ChunkStartNodes(map<int, vector<uint64_t>, greater<int>> &start_nodes_chunked) {
  SDBG &sdbg = *this->settings.sdbg;
  GpuState &s = state_of(this, sdbg, this->settings.graph_folder);
  ensure_pruned(s, sdbg, nullptr);
  mg_starts st{};
  int rc = mg_select_starts(s.g, this->settings.cycle_finder_settings.threshold_multiplicity,
                            this->settings.cycle_finder_settings.cycle_max_length, &st);
  if (rc != MG_OK) fail("mg_select_starts", rc);
  for (uint64_t b = 0; b < st.n_buckets; ++b) {
    auto &v = start_nodes_chunked[(int)st.bucket_key[b]];
    v.insert(v.end(), st.start_ids + st.bucket_off[b], st.start_ids + st.bucket_off[b + 1]);
  }
  mg_starts_free(&st);
  // the scan's progress lines (:403): one per 50 M valid items at threads=1
  for (uint64_t m = 50; m * 1000000ull <= s.stats.n_valid; m += 50) std::cout << "ChunkStartNodes: scanned " << m << "M nodes" << std::endl;
  size_t total = 0;
  for (const auto &[key, value] : start_nodes_chunked) {
    std::cout << "Chunked start nodes: multiplicity bucket (log2)=" << key << ", nodes=" << value.size() << std::endl;
    total += value.size();
  }
  return total;
}

// ---- reference :248-343 (only target == start is used, :411)
bool CycleFinder::DepthLevelSearch(uint64_t start, uint64_t target, int limit, int &reached_depth) {
  if (start != target) throw std::runtime_error("DepthLevelSearch: only target == start is supported on the GPU path");
  SDBG &sdbg = *this->settings.sdbg;
  GpuState &s = state_of(this, sdbg, this->settings.graph_folder);
  ensure_pruned(s, sdbg, nullptr);
  uint8_t pass = 0;
  int rc = mg_dls(s.g, &start, 1, limit, &pass, nullptr);
  if (rc != MG_OK) fail("mg_dls", rc);
  reached_depth = 0;
  return pass != 0;
}

static vector<vector<uint64_t>> cycles_of(const mg_cycles &c, uint64_t idx) {
  vector<vector<uint64_t>> out;
  for (uint64_t j = c.cyc_off[idx]; j < c.cyc_off[idx + 1]; ++j)
    out.emplace_back(c.nodes + c.node_off[j], c.nodes + c.node_off[j + 1]);
  return out;
}

// ---- reference :231-243 (+ :140-225): one start node against the current `visited` state
vector<vector<uint64_t>> CycleFinder::FindCycleUtil(uint64_t start_node) {
  SDBG &sdbg = *this->settings.sdbg;
  GpuState &s = state_of(this, sdbg, this->settings.graph_folder);
  ensure_pruned(s, sdbg, nullptr);
  mg_cycles c{};
  int rc = mg_find_cycles(s.g, &start_node, 1, this->settings.cycle_finder_settings.cycle_min_length,
                          this->settings.cycle_finder_settings.cycle_max_length, &c);
  if (rc != MG_OK) fail("mg_find_cycles", rc);
  vector<vector<uint64_t>> out;
  if (c.n_searched) out = cycles_of(c, 0);
  mg_cycles_free(&c);
  return out;
}
// The reference's FindCycle takes the search state by value; only the FindCycleUtil entry state is
// ever passed (:242), which is what the device search starts from.
vector<vector<uint64_t>> CycleFinder::FindCycle(uint64_t start_node, vector<uint64_t>, map<uint64_t, int>,
                                                vector<unordered_set<uint64_t>>, vector<int>) {
  return this->FindCycleUtil(start_node);
}

// ---- reference :433-492
int CycleFinder::FindApproximateCRISPRArrays() {
  SDBG &sdbg = *this->settings.sdbg;
  GpuState &s = state_of(this, sdbg, this->settings.graph_folder);
  const auto &cfs = this->settings.cycle_finder_settings;
  mg_prune_stats ps{};
  ensure_pruned(s, sdbg, &ps);
  std::cout << "Graph size: " << sdbg.size() << " nodes; gathered tips: " << ps.n_tips_before << std::endl;
  std::cout << "Pre-filter: invalidated " << ps.n_mult1 << " node(s) with multiplicity <= 1." << std::endl;
  std::cout << "After pruning, tips: " << ps.n_tips_after << ", valid edges: " << (int)ps.n_valid << std::endl;
  std::cout << "Total nodes in graph: " << sdbg.size() << std::endl;
  this->visited.resize(sdbg.size(), false);
  std::cout << "Starting cycle enumeration: max_len=" << cfs.cycle_max_length << " min_len=" << cfs.cycle_min_length
            << " threads=" << this->settings.threads << std::endl;
  map<int, vector<uint64_t>, greater<int>> start_nodes_chunked;
  const double t_starts = now_ms();
  size_t start_nodes_amount = this->ChunkStartNodes(start_nodes_chunked);
  g_timing.starts_ms = now_ms() - t_starts;
  std::cout << "Start nodes found in chunks: " << start_nodes_amount << std::endl;

  vector<uint64_t> order;
  order.reserve(start_nodes_amount);
  for (auto &kv : start_nodes_chunked) order.insert(order.end(), kv.second.begin(), kv.second.end());
  mg_cycles c{};
  const double t_cyc = now_ms();
  int rc = mg_find_cycles(s.g, order.data(), order.size(), cfs.cycle_min_length, cfs.cycle_max_length, &c);
  if (rc != MG_OK) fail("mg_find_cycles", rc);
  g_timing.cycles_ms = now_ms() - t_cyc;
  const double t_res = now_ms();
  // results[start] = cycles in the reference's insertion sequence (:479; F14: downstream iterates
  // the unordered_map, whose order depends on it)
  std::unordered_map<uint64_t, uint64_t> n_cycles_of;
  for (uint64_t i = 0; i < c.n_searched; ++i) {
    this->results[c.searched_start[i]] = cycles_of(c, i);
    n_cycles_of[c.searched_start[i]] = c.cyc_off[i + 1] - c.cyc_off[i];
    for (uint64_t j = c.node_off[c.cyc_off[i]]; j < c.node_off[c.cyc_off[i + 1]]; ++j) this->visited[c.nodes[j]] = true;
  }
  mg_cycles_free(&c);
  g_timing.results_ms = now_ms() - t_res;
  if (const char *tm = std::getenv("MCAAT_GPU_TIMING"); tm && *tm == '1')
    std::fprintf(stderr, "GPU_SHIM_TIMING {\"graph\": \"%s\", \"graph_ms\": %.1f, \"prune_ms\": %.1f, \"invalid_write_back_ms\": %.1f, "
                 "\"select_starts_ms\": %.1f, \"find_cycles_ms\": %.1f, \"results_ms\": %.1f}\n", g_timing.graph_path, g_timing.graph_ms,
                 g_timing.prune_ms, g_timing.writeback_ms, g_timing.starts_ms, g_timing.cycles_ms, g_timing.results_ms);
  int cumulative = 0;
  for (auto &kv : start_nodes_chunked) {
    size_t in_bucket = 0;
    for (uint64_t v : kv.second) {
      auto it = n_cycles_of.find(v);
      if (it != n_cycles_of.end()) in_bucket += it->second;
    }
    cumulative += (int)in_bucket;
    std::cout << "Bucket log2_mult=" << kv.first << ": processed " << kv.second.size() << " nodes, found " << in_bucket
              << " cycles (cumulative " << cumulative << ")" << std::endl;
  }
  std::cout << "Cycle enumeration completed: total cycles=" << cumulative << ", result nodes=" << this->results.size()
            << std::endl;
  return cumulative;
}
