// Host-side plumbing shared by the .cu files of libmcaat_gpu: error handling, device buffers,
// the graph handle, kernel-family timers and the launch counter.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/mcaat_gpu.h"
#include "sdbg_dev.cuh"
#include "kmer_dev.cuh"

namespace mg {

void set_error(const char *fmt, ...);
extern unsigned long long g_launches;

#define MG_CUDA(call)                                                                      \
  do {                                                                                     \
    cudaError_t e__ = (call);                                                              \
    if (e__ != cudaSuccess) {                                                              \
      mg::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return MG_ECUDA;                                                                     \
    }                                                                                      \
  } while (0)
#define MG_TRY(call)          \
  do {                        \
    int r__ = (call);         \
    if (r__ != MG_OK) return r__; \
  } while (0)

// every kernel launch of the library goes through this macro (launch counter + error check)
#define MG_LAUNCH(kernel, grid, block, smem, stream, ...)                     \
  do {                                                                        \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);               \
    ++mg::g_launches;                                                         \
    cudaError_t e__ = cudaGetLastError();                                     \
    if (e__ != cudaSuccess) {                                                 \
      mg::set_error("%s:%d: launch %s -> %s", __FILE__, __LINE__, #kernel, cudaGetErrorString(e__)); \
      return MG_ECUDA;                                                        \
    }                                                                         \
  } while (0)

// Device buffers come from the device's stream-ordered memory pool (cudaMallocAsync) on the stream of
// the graph handle the running API call belongs to.  In front of the pool sits an exact-size free list per
// stream (scratch_take / scratch_give, api.cu): a pass of the hot path allocates the same ~30 scratch buffers
// (3.5 GB at 225 M items) with the same sizes every time, so after the first pass no call reaches the driver's
// allocator any more — on the B200 boxes a pool allocation of a GB-sized block sporadically took hundreds of
// milliseconds (physical memory being mapped again) in the middle of a steady-state run.  Buffers are handed out
// again only to calls on the same stream, so stream order keeps reuse safe.  MCAAT_SCRATCH_CACHE=0 disables it.
extern thread_local cudaStream_t g_alloc_stream;
void *scratch_take(cudaStream_t st, size_t bytes);            // nullptr when nothing of that size is cached
bool scratch_give(cudaStream_t st, void *p, size_t bytes);    // false: not cached, the caller frees it
void scratch_flush(cudaStream_t st);                          // frees everything cached for the stream
struct DevBuf {
  void *p = nullptr;
  size_t bytes = 0;
  cudaStream_t st = nullptr;  // stream the buffer was allocated on (and is freed on)
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  ~DevBuf() { release(); }
  void release() {
    if (p && !scratch_give(st, p, bytes)) cudaFreeAsync(p, st);
    p = nullptr;
    bytes = 0;
  }
  int alloc(size_t b) {
    release();
    if (b == 0) b = 16;
    b = (b + 255) & ~(size_t)255;
    st = g_alloc_stream;
    p = scratch_take(st, b);
    if (p) {
      bytes = b;
      return MG_OK;
    }
    cudaError_t e = cudaMallocAsync(&p, b, st);
    if (e != cudaSuccess) {
      scratch_flush(st);  // give the pool its memory back and try once more
      e = cudaMallocAsync(&p, b, st);
    }
    if (e != cudaSuccess) {
      p = nullptr;
      set_error("cudaMallocAsync(%zu) -> %s", b, cudaGetErrorString(e));
      return MG_ENOMEM;
    }
    bytes = b;
    return MG_OK;
  }
  // grow-only
  int ensure(size_t b) { return b <= bytes ? MG_OK : alloc(b); }
  template <class T> T *as() const { return (T *)p; }
};

// page-locked host memory for results on their way back (grow-only, one per graph handle)
struct PinnedBuf {
  void *p = nullptr;
  size_t bytes = 0;
  int ensure(size_t b) {
    if (b <= bytes) return MG_OK;
    release();
    b = (b + (b >> 2) + 4095) & ~(size_t)4095;
    cudaError_t e = cudaMallocHost(&p, b);
    if (e != cudaSuccess) {
      p = nullptr;
      set_error("cudaMallocHost(%zu) -> %s", b, cudaGetErrorString(e));
      return MG_ENOMEM;
    }
    bytes = b;
    return MG_OK;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
  }
};

// one spare staging buffer per process: a fresh handle takes it over instead of page-locking memory again (4 ms for 5 MB)
void pinned_spare_take(PinnedBuf &into);
void pinned_spare_give(PinnedBuf &from);

enum TimerId {
  T_BUILD = 0,     // index construction
  T_TIPS,          // CollectTips scans
  T_MULT1,         // InvalidateMultiplicityOneNodes
  T_PEEL,          // RecursiveReduction
  T_ADJ,           // navigation cache + tips-after + valid count
  T_CAND,          // candidate predicate + compaction
  T_DLS,           // DepthLevelSearch
  T_ENUM,          // FindCycle searches
  T_COMMIT,        // ordered commit + gather
  T_NAV,           // mg_navigate (diagnostic)
  T_REGION,        // mg_keep_regions
  T_READS,         // mg_map_reads
  T_COUNT
};
extern const char *kTimerNames[T_COUNT];

}  // namespace mg

// The opaque handle of the C ABI.
struct mg_graph {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;  // kernels that run beside the main stream (ball tier of the admission search), lazily created
  mg::GraphView view{};  // pointers below
  mg::DevBuf wb, fb, mult, hints, inv0;
  mg::DevBuf adj;        // navigation cache (built by mg_prune)
  mg::DevBuf merge;      // n bits: the valid items of every node that a valid non-first edge (label 5..8) leads to — the
                         // nodes that can have in-degree >= 2; written with the cache, valid while `pruned`
  mg::DevBuf visited;    // n bits, uint32 words
  mg::DevBuf enum_arena; // per-worker search arenas of the enumeration pipeline (grow-only)
  uint64_t enum_arena_sig = 0;  // layout the arenas were initialised for
  mg::PinnedBuf host_stage;     // read-back staging of mg_find_cycles (taken from / returned to a process-wide spare)
  uint32_t n_ball_tier_calls = 0;  // passes of the admission search that reached its ball tier (starts.cu)
  bool pruned = false;
  // k-mer lookup index, optional
  uint32_t k = 0;
  bool has_kx = false;
  mg::KmerIndex kx{};
  mg::DevBuf kx_tip, kx_tip_prefix, kx_labels, kx_buckets, kx_jump, kx_bucket_ok;
  uint64_t device_bytes = 0;
  int sm_count = 148;
  size_t smem_per_sm = 233472, smem_per_block_optin = 232448;  // B200: 228 KB per SM, 227 KB per block (opt-in)
  // timers
  double t_ms[mg::T_COUNT] = {0};
  uint64_t t_launches[mg::T_COUNT] = {0};
  uint64_t t_units[mg::T_COUNT] = {0};
  struct Pending { int id; cudaEvent_t a, b; };
  std::vector<Pending> pending;
  std::vector<cudaEvent_t> event_pool;
  cudaEvent_t step_a = nullptr, step_b = nullptr;
};

namespace mg {

// Records CUDA events on the graph's stream around a group of launches of one kernel family.
struct TimedScope {
  mg_graph *g;
  int id;
  cudaEvent_t a = nullptr, b = nullptr;
  unsigned long long launches0;
  TimedScope(mg_graph *g_, int id_, uint64_t units = 0) : g(g_), id(id_), launches0(g_launches) {
    a = get();
    b = get();
    if (a) cudaEventRecord(a, g->stream);
    g->t_units[id] += units;
  }
  ~TimedScope() {
    if (b) cudaEventRecord(b, g->stream);
    g->t_launches[id] += g_launches - launches0;
    if (a && b) g->pending.push_back({id, a, b});
  }
  cudaEvent_t get() {
    if (!g->event_pool.empty()) {
      cudaEvent_t e = g->event_pool.back();
      g->event_pool.pop_back();
      return e;
    }
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
    return e;
  }
};
// call after the stream has been synchronised
void flush_timers(mg_graph *g);

// in-place exclusive prefix sum over n uint32 values; *total (device pointer, may be null)
// receives the grand total.  scan.cu
int exclusive_scan_u32(uint32_t *data, size_t n, uint32_t *total_dev, cudaStream_t stream);
// ascending list of set bit positions of a bitmap of n bits (uint32 words); allocates *out_ids
// (device) and writes the count to *count.  scan.cu
int bitmap_to_ids(const uint32_t *bits, uint64_t n, DevBuf &out_ids, uint64_t *count, cudaStream_t stream);

// graph_build.cu
int build_graph(mg_graph *g, const mg_graph_desc *d);
int reset_graph(mg_graph *g);
// prune.cu
int run_prune(mg_graph *g, mg_prune_stats *st);
int run_collect_tips(mg_graph *g, DevBuf &ids, uint64_t *n_tips);
int run_navigate(mg_graph *g, uint64_t first, uint64_t count, int8_t *outdeg, int8_t *indeg, uint64_t *out_ids,
                 uint64_t *in_ids, uint16_t *mult);
// starts.cu
int run_candidates(mg_graph *g, uint64_t thr, DevBuf &ids, uint64_t *n);
int run_dls(mg_graph *g, const uint32_t *ids_dev, uint64_t n, int limit, uint8_t *pass_host, uint64_t *expansions);
int bucket_on_host(const std::vector<uint64_t> &ids, const std::vector<uint16_t> &mult, mg_starts *out);
int gather_mult(mg_graph *g, const uint32_t *ids_dev, uint64_t n, std::vector<uint16_t> &out);
// regions.cu
int run_keep_regions(mg_graph *g, const uint32_t *ids_dev, uint64_t n_ids, uint64_t k, mg_region_stats *st);
int run_valid_subgraph(mg_graph *g, mg_subgraph *out);
// reads.cu
int attach_kmer_index(mg_graph *g, const uint64_t *tip_dev, const uint32_t *labels_dev, uint64_t n_tips, uint32_t wptl,
                      const int64_t *first_host, const int64_t *second_host, uint64_t n_buckets);
int run_map_reads(mg_graph *g, const char *seqs, const uint64_t *seq_off, uint64_t n_reads, int reverse,
                  const uint32_t *cycle_ids_dev, uint64_t n_cycle_ids, mg_reads *out);
// enumerate.cu
int run_find_cycles(mg_graph *g, const uint64_t *starts, uint64_t n, int min_len, int max_len, mg_cycles *out);

}  // namespace mg
