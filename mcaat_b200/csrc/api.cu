// C ABI of libmcaat_gpu (include/mcaat_gpu.h).  No CPU fallback: without a CUDA device every
// compute entry returns MG_ECUDA.
#include <stdarg.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "search_dev.cuh"

namespace mg {

static thread_local char g_err[512] = "";
unsigned long long g_launches = 0;
thread_local cudaStream_t g_alloc_stream = nullptr;

// ---- exact-size free list in front of the stream-ordered pool (see DevBuf, common.cuh).  Entries, the byte count
// and the byte limit are kept per device: a process that holds replicas on several GPUs neither evicts one
// device's buffers for another's nor frees a block while a different device is current.
namespace {
constexpr int kScratchMaxDevices = 64;
struct ScratchEntry { int dev; cudaStream_t st; void *p; size_t bytes; };
std::mutex g_scratch_mu;
std::vector<ScratchEntry> g_scratch;
size_t g_scratch_bytes[kScratchMaxDevices] = {0};
size_t g_scratch_lim[kScratchMaxDevices] = {0};
constexpr size_t kScratchMaxEntries = 512;  // per device
bool scratch_enabled() {
  static const bool on = [] { const char *e = getenv("MCAAT_SCRATCH_CACHE"); return !(e && *e == '0'); }();
  return on;
}
int scratch_device() {  // DevBuf allocates and releases with the handle's device current (MG_CHECK_HANDLE)
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= kScratchMaxDevices) return -1;
  return d;
}
size_t scratch_limit(int dev) {  // bytes kept per device: a quarter of it, at least 1 GB (call with g_scratch_mu held)
  if (!g_scratch_lim[dev]) {
    size_t fr = 0, tot = 0;
    if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) tot = (size_t)16 << 30;
    g_scratch_lim[dev] = std::max<size_t>(tot / 4, (size_t)1 << 30);
  }
  return g_scratch_lim[dev];
}
}  // namespace

void *scratch_take(cudaStream_t st, size_t bytes) {
  if (!scratch_enabled()) return nullptr;
  const int dev = scratch_device();
  if (dev < 0) return nullptr;
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  for (size_t i = g_scratch.size(); i-- > 0;)
    if (g_scratch[i].dev == dev && g_scratch[i].st == st && g_scratch[i].bytes == bytes) {
      void *p = g_scratch[i].p;
      g_scratch_bytes[dev] -= bytes;
      g_scratch.erase(g_scratch.begin() + (long)i);
      return p;
    }
  return nullptr;
}

bool scratch_give(cudaStream_t st, void *p, size_t bytes) {
  if (!scratch_enabled() || !st) return false;
  const int dev = scratch_device();
  if (dev < 0) return false;
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  const size_t lim = scratch_limit(dev);
  if (bytes > lim) return false;
  // make room on this device: oldest entries first (sizes that no longer occur, e.g. after another graph was searched)
  size_t n_dev = 0;
  for (const ScratchEntry &e : g_scratch) n_dev += e.dev == dev;
  for (size_t i = 0; i < g_scratch.size() && (g_scratch_bytes[dev] + bytes > lim || n_dev >= kScratchMaxEntries);) {
    if (g_scratch[i].dev != dev) { ++i; continue; }
    cudaFreeAsync(g_scratch[i].p, g_scratch[i].st);
    g_scratch_bytes[dev] -= g_scratch[i].bytes;
    --n_dev;
    g_scratch.erase(g_scratch.begin() + (long)i);
  }
  g_scratch.push_back({dev, st, p, bytes});
  g_scratch_bytes[dev] += bytes;
  return true;
}

void scratch_flush(cudaStream_t st) {  // called with the stream's device current
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  size_t w = 0;
  for (size_t i = 0; i < g_scratch.size(); ++i) {
    if (g_scratch[i].st == st) {
      cudaFreeAsync(g_scratch[i].p, st);
      g_scratch_bytes[g_scratch[i].dev] -= g_scratch[i].bytes;
    } else {
      g_scratch[w++] = g_scratch[i];
    }
  }
  g_scratch.resize(w);
}
namespace {
std::mutex g_pinned_mu;
PinnedBuf g_pinned_spare;
}  // namespace
void pinned_spare_take(PinnedBuf &into) {
  std::lock_guard<std::mutex> lk(g_pinned_mu);
  if (!into.p && g_pinned_spare.p) {
    into = g_pinned_spare;
    g_pinned_spare.p = nullptr;
    g_pinned_spare.bytes = 0;
  }
}
void pinned_spare_give(PinnedBuf &from) {
  std::lock_guard<std::mutex> lk(g_pinned_mu);
  if (from.p && from.bytes > g_pinned_spare.bytes) {
    g_pinned_spare.release();
    g_pinned_spare = from;
    from.p = nullptr;
    from.bytes = 0;
  }
  from.release();
}
const char *kTimerNames[T_COUNT] = {"build_index", "collect_tips", "invalidate_mult1", "peel",      "nav_cache",
                                    "candidates",  "dls",          "enumerate",        "commit",    "navigate",
                                    "keep_regions", "map_reads"};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

void flush_timers(mg_graph *g) {
  for (auto &p : g->pending) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) g->t_ms[p.id] += ms;
    g->event_pool.push_back(p.a);
    g->event_pool.push_back(p.b);
  }
  g->pending.clear();
}

static int upload_ids(mg_graph *g, const uint64_t *ids, uint64_t n, DevBuf &dev) {
  std::vector<uint32_t> v(n);
  for (uint64_t i = 0; i < n; ++i) {
    if (ids[i] >= g->view.n) {
      set_error("node id %llu out of range (graph has %llu items)", (unsigned long long)ids[i],
                (unsigned long long)g->view.n);
      return MG_EINVAL;
    }
    v[i] = (uint32_t)ids[i];
  }
  MG_TRY(dev.alloc(n * 4 + 16));
  if (n) MG_CUDA(cudaMemcpyAsync(dev.p, v.data(), n * 4, cudaMemcpyHostToDevice, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  return MG_OK;
}

}  // namespace mg

using namespace mg;

#define MG_CHECK_HANDLE(g)                   \
  do {                                       \
    if (!(g)) {                              \
      set_error("null graph handle");        \
      return MG_EINVAL;                      \
    }                                        \
    cudaError_t e__ = cudaSetDevice((g)->device); \
    if (e__ != cudaSuccess) {                \
      set_error("cudaSetDevice(%d) -> %s", (g)->device, cudaGetErrorString(e__)); \
      return MG_ECUDA;                       \
    }                                        \
    mg::g_alloc_stream = (g)->stream;        \
  } while (0)

extern "C" {

const char *mg_last_error(void) { return g_err; }
const char *mg_version(void) { return "mcaat_b200 libmcaat_gpu 0.1 (sm_100a)"; }
uint64_t mg_launch_count(void) { return g_launches; }
void mg_free(void *p) { free(p); }

int mg_graph_create(const mg_graph_desc *d, mg_graph **out) {
  if (!d || !out) { set_error("mg_graph_create: null argument"); return MG_EINVAL; }
  *out = nullptr;
  if (!d->w || !d->last || !d->tip) { set_error("mg_graph_create: w/last/tip are required"); return MG_EINVAL; }
  if (d->n_items >= 0xFFFFFFFFull) {
    set_error("mg_graph_create: %llu items; this build addresses < 2^32-1 items per graph",
              (unsigned long long)d->n_items);
    return MG_ELIMIT;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("no CUDA device available (%s); libmcaat_gpu has no CPU fallback",
              e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    return MG_ECUDA;
  }
  if (d->device < 0 || d->device >= ndev) { set_error("mg_graph_create: device %d of %d", d->device, ndev); return MG_EINVAL; }
  MG_CUDA(cudaSetDevice(d->device));
  mg_graph *g = new (std::nothrow) mg_graph();
  if (!g) return MG_ENOMEM;
  g->device = d->device;
  g->k = d->k;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, d->device) == cudaSuccess) {
    g->sm_count = prop.multiProcessorCount;
    g->smem_per_sm = prop.sharedMemPerMultiprocessor;
    g->smem_per_block_optin = prop.sharedMemPerBlockOptin;
  }
  if (cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking) != cudaSuccess) {
    set_error("cudaStreamCreate failed");
    delete g;
    return MG_ECUDA;
  }
  g_alloc_stream = g->stream;
  {  // keep freed scratch memory in the pool instead of returning it to the driver
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, d->device) == cudaSuccess) {
      uint64_t keep = UINT64_MAX;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  }
  int r = build_graph(g, d);
  if (r != MG_OK) { mg_graph_destroy(g); return r; }
  flush_timers(g);
  *out = g;
  return MG_OK;
}

void mg_graph_destroy(mg_graph *g) {
  if (!g) return;
  cudaSetDevice(g->device);
  g_alloc_stream = g->stream;
  if (g->stream) cudaStreamSynchronize(g->stream);
  flush_timers(g);
  // buffers go back to the pool on the handle's stream, which must still exist
  for (mg::DevBuf *b : {&g->wb, &g->fb, &g->mult, &g->hints, &g->inv0, &g->adj, &g->merge, &g->visited, &g->enum_arena,
                       &g->kx_tip, &g->kx_tip_prefix, &g->kx_labels, &g->kx_buckets, &g->kx_jump, &g->kx_bucket_ok})
    b->release();
  pinned_spare_give(g->host_stage);
  scratch_flush(g->stream);  // nothing cached may outlive the stream it is ordered on
  if (g->stream) cudaStreamSynchronize(g->stream);
  for (auto e : g->event_pool) cudaEventDestroy(e);
  if (g->step_a) cudaEventDestroy(g->step_a);
  if (g->step_b) cudaEventDestroy(g->step_b);
  if (g->stream2) cudaStreamDestroy(g->stream2);
  if (g->stream) cudaStreamDestroy(g->stream);
  delete g;
}

int mg_graph_reset(mg_graph *g) {
  MG_CHECK_HANDLE(g);
  return reset_graph(g);
}
uint64_t mg_graph_size(const mg_graph *g) { return g ? g->view.n : 0; }
uint64_t mg_graph_device_bytes(const mg_graph *g) { return g ? g->device_bytes + g->adj.bytes : 0; }

int mg_navigate(mg_graph *g, uint64_t first, uint64_t count, int8_t *outdeg, int8_t *indeg, uint64_t *out_ids,
                uint64_t *in_ids, uint16_t *mult) {
  MG_CHECK_HANDLE(g);
  if (first + count > g->view.n) { set_error("mg_navigate: range past the end"); return MG_EINVAL; }
  if (count == 0) return MG_OK;
  return run_navigate(g, first, count, outdeg, indeg, out_ids, in_ids, mult);
}

int mg_prune(mg_graph *g, mg_prune_stats *stats) {
  MG_CHECK_HANDLE(g);
  mg_prune_stats tmp;
  return run_prune(g, stats ? stats : &tmp);
}

__global__ void k_inv_words(uint64_t n_fblocks, const FBlock *fb, uint64_t *out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_fblocks * 4) out[i] = fb[i >> 2].inv[i & 3];
}

int mg_invalid_readback(mg_graph *g, uint64_t *host_words) {
  MG_CHECK_HANDLE(g);
  if (!host_words) return MG_EINVAL;
  const uint64_t nb = g->view.n_fblocks;
  if (!nb) return MG_OK;
  DevBuf tmp;
  MG_TRY(tmp.alloc(nb * 32));
  MG_LAUNCH(k_inv_words, (unsigned)((nb * 4 + 255) / 256), 256, 0, g->stream, nb, g->fb.as<FBlock>(), tmp.as<uint64_t>());
  MG_CUDA(cudaMemcpyAsync(host_words, tmp.p, (size_t)((g->view.n + 63) / 64) * 8, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  return MG_OK;
}

__global__ void k_inv_merge(uint64_t n, uint64_t n_words, const uint64_t *in, FBlock *fb, unsigned long long *changed) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_words) return;
  uint64_t first = i * 64, mask = first + 64 <= n ? ~0ull : (~0ull >> (64 - (n - first)));
  uint64_t *p = &fb[i >> 2].inv[i & 3], old = *p, add = in[i] & mask & ~old;
  if (add) {
    *p = old | add;
    atomicAdd(changed, (unsigned long long)__popcll(add));
  }
}

int mg_invalid_merge(mg_graph *g, const uint64_t *host_words, uint64_t *n_new) {
  MG_CHECK_HANDLE(g);
  if (!host_words) return MG_EINVAL;
  if (n_new) *n_new = 0;
  const uint64_t n = g->view.n, nw = (n + 63) / 64;
  if (!nw) return MG_OK;
  DevBuf tmp;
  MG_TRY(tmp.alloc(nw * 8 + 8));
  MG_CUDA(cudaMemcpyAsync(tmp.p, host_words, nw * 8, cudaMemcpyHostToDevice, g->stream));
  unsigned long long *cnt = tmp.as<unsigned long long>() + nw, h = 0;
  MG_CUDA(cudaMemsetAsync(cnt, 0, 8, g->stream));
  MG_LAUNCH(k_inv_merge, (unsigned)((nw + 255) / 256), 256, 0, g->stream, n, nw, tmp.as<uint64_t>(), g->fb.as<FBlock>(), cnt);
  MG_CUDA(cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  if (h) g->pruned = false;  // the navigation cache describes the graph before the merge
  if (n_new) *n_new = h;
  return MG_OK;
}

int mg_keep_regions(mg_graph *g, const uint64_t *ids, uint64_t n_ids, uint64_t k, mg_region_stats *stats) {
  MG_CHECK_HANDLE(g);
  if (n_ids && !ids) return MG_EINVAL;
  DevBuf dev;
  MG_TRY(upload_ids(g, ids, n_ids, dev));
  return run_keep_regions(g, dev.as<uint32_t>(), n_ids, k, stats);
}

int mg_kmer_index_create(mg_graph *g, const mg_kmer_index_desc *d) {
  MG_CHECK_HANDLE(g);
  if (!d || !d->tip || !d->bucket_first || !d->bucket_last || (d->n_tips && d->words_per_tip_label && !d->tip_labels)) return MG_EINVAL;
  const uint64_t n_words = (g->view.n + 63) / 64;
  DevBuf tip, lab;
  const uint64_t *tip_dev = d->tip;
  const uint32_t *lab_dev = d->tip_labels;
  if (d->memory != MG_MEM_DEVICE) {
    MG_TRY(tip.alloc(n_words * 8 + 16));
    MG_TRY(lab.alloc((size_t)d->n_tips * d->words_per_tip_label * 4 + 16));
    if (n_words) MG_CUDA(cudaMemcpyAsync(tip.p, d->tip, n_words * 8, cudaMemcpyHostToDevice, g->stream));
    if (d->n_tips && d->words_per_tip_label)
      MG_CUDA(cudaMemcpyAsync(lab.p, d->tip_labels, (size_t)d->n_tips * d->words_per_tip_label * 4, cudaMemcpyHostToDevice, g->stream));
    tip_dev = tip.as<uint64_t>();
    lab_dev = lab.as<uint32_t>();
  }
  return attach_kmer_index(g, tip_dev, lab_dev, d->n_tips, d->words_per_tip_label, d->bucket_first, d->bucket_last, d->n_buckets);
}

int mg_graph_has_kmer_index(const mg_graph *g) { return g && g->has_kx ? 1 : 0; }

int mg_map_reads(mg_graph *g, const char *seqs, const uint64_t *seq_off, uint64_t n_reads, int reverse_complement,
                 const uint64_t *cycle_ids, uint64_t n_cycle_ids, mg_reads *out) {
  MG_CHECK_HANDLE(g);
  if (!out || (n_reads && (!seqs || !seq_off)) || (n_cycle_ids && !cycle_ids)) return MG_EINVAL;
  DevBuf dev;
  MG_TRY(upload_ids(g, cycle_ids, n_cycle_ids, dev));
  return run_map_reads(g, seqs, seq_off, n_reads, reverse_complement, dev.as<uint32_t>(), n_cycle_ids, out);
}

void mg_reads_free(mg_reads *r) {
  if (!r) return;
  free(r->off);
  free(r->ids);
  memset(r, 0, sizeof *r);
}

int mg_valid_subgraph(mg_graph *g, mg_subgraph *out) {
  MG_CHECK_HANDLE(g);
  if (!out) return MG_EINVAL;
  return run_valid_subgraph(g, out);
}

void mg_subgraph_free(mg_subgraph *sg) {
  if (!sg) return;
  free(sg->ids);
  free(sg->off);
  free(sg->nbr);
  memset(sg, 0, sizeof *sg);
}

int mg_visited_readback(mg_graph *g, uint64_t *host_words) {
  MG_CHECK_HANDLE(g);
  if (!host_words) return MG_EINVAL;
  if (!g->view.n) return MG_OK;
  size_t bytes = (size_t)((g->view.n + 31) / 32) * 4;
  host_words[(g->view.n + 63) / 64 - 1] = 0;
  MG_CUDA(cudaMemcpyAsync(host_words, g->visited.p, bytes, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  return MG_OK;
}

int mg_candidates(mg_graph *g, uint64_t thr, uint64_t **ids, uint64_t *n) {
  MG_CHECK_HANDLE(g);
  if (!ids || !n) return MG_EINVAL;
  *ids = nullptr;
  *n = 0;
  DevBuf dev;
  uint64_t cnt = 0;
  MG_TRY(run_candidates(g, thr, dev, &cnt));
  std::vector<uint32_t> h(cnt);
  if (cnt) MG_CUDA(cudaMemcpyAsync(h.data(), dev.p, cnt * 4, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  uint64_t *o = (uint64_t *)malloc((cnt + 1) * 8);
  if (!o) return MG_ENOMEM;
  for (uint64_t i = 0; i < cnt; ++i) o[i] = h[i];
  *ids = o;
  *n = cnt;
  return MG_OK;
}

int mg_collect_tips(mg_graph *g, uint64_t **ids, uint64_t *n) {
  MG_CHECK_HANDLE(g);
  if (!ids || !n) return MG_EINVAL;
  *ids = nullptr;
  *n = 0;
  DevBuf dev;
  uint64_t cnt = 0;
  MG_TRY(run_collect_tips(g, dev, &cnt));
  std::vector<uint32_t> h(cnt);
  if (cnt) MG_CUDA(cudaMemcpyAsync(h.data(), dev.p, cnt * 4, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  flush_timers(g);
  uint64_t *o = (uint64_t *)malloc((cnt + 1) * 8);
  if (!o) return MG_ENOMEM;
  for (uint64_t i = 0; i < cnt; ++i) o[i] = h[i];
  *ids = o;
  *n = cnt;
  return MG_OK;
}

int mg_dls(mg_graph *g, const uint64_t *ids, uint64_t n, int limit, uint8_t *pass, uint64_t *expansions) {
  MG_CHECK_HANDLE(g);
  if (n && (!ids || !pass)) return MG_EINVAL;
  DevBuf dev;
  MG_TRY(upload_ids(g, ids, n, dev));
  return run_dls(g, dev.as<uint32_t>(), n, limit, pass, expansions);
}

int mg_select_starts(mg_graph *g, uint64_t thr, int max_len, mg_starts *out) {
  MG_CHECK_HANDLE(g);
  if (!out) return MG_EINVAL;
  memset(out, 0, sizeof *out);
  DevBuf cand;
  uint64_t nc = 0;
  MG_TRY(run_candidates(g, thr, cand, &nc));
  std::vector<uint8_t> pass(nc);
  uint64_t exp = 0;
  MG_TRY(run_dls(g, cand.as<uint32_t>(), nc, max_len, pass.data(), &exp));
  std::vector<uint32_t> h(nc);
  std::vector<uint16_t> hm;
  if (nc) MG_CUDA(cudaMemcpyAsync(h.data(), cand.p, nc * 4, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  MG_TRY(gather_mult(g, cand.as<uint32_t>(), nc, hm));
  std::vector<uint64_t> ids;
  std::vector<uint16_t> mult;
  for (uint64_t i = 0; i < nc; ++i)
    if (pass[i]) { ids.push_back(h[i]); mult.push_back(hm[i]); }
  int r = bucket_on_host(ids, mult, out);
  out->n_candidates = nc;
  out->dls_expansions = exp;
  if (r != MG_OK) mg_starts_free(out);
  return r;
}

int mg_bucket_starts(mg_graph *g, const uint64_t *ids, uint64_t n, mg_starts *out) {
  MG_CHECK_HANDLE(g);
  if (!out || (n && !ids)) return MG_EINVAL;
  memset(out, 0, sizeof *out);
  DevBuf dev;
  MG_TRY(upload_ids(g, ids, n, dev));
  std::vector<uint16_t> hm;
  MG_TRY(gather_mult(g, dev.as<uint32_t>(), n, hm));
  std::vector<uint64_t> v(ids, ids + n);
  int r = bucket_on_host(v, hm, out);
  if (r != MG_OK) mg_starts_free(out);
  return r;
}

void mg_starts_free(mg_starts *s) {
  if (!s) return;
  free(s->bucket_key);
  free(s->bucket_off);
  free(s->start_ids);
  memset(s, 0, sizeof *s);
}

int mg_find_cycles(mg_graph *g, const uint64_t *starts, uint64_t n, int min_len, int max_len, mg_cycles *out) {
  MG_CHECK_HANDLE(g);
  if (!out || (n && !starts)) return MG_EINVAL;
  memset(out, 0, sizeof *out);
  int r = run_find_cycles(g, starts, n, min_len, max_len, out);
  if (r != MG_OK) mg_cycles_free(out);
  return r;
}

void mg_cycles_free(mg_cycles *c) {
  if (!c) return;
  free(c->searched_start);
  free(c->cyc_off);
  free(c->node_off);
  free(c->nodes);
  memset(c, 0, sizeof *c);
}

int mg_timers(mg_graph *g, const char **names, double *ms, uint64_t *launches, uint64_t *units, int cap) {
  if (!g) return 0;
  cudaSetDevice(g->device);
  cudaStreamSynchronize(g->stream);
  flush_timers(g);
  int n = cap < (int)T_COUNT ? cap : (int)T_COUNT;
  for (int i = 0; i < n; ++i) {
    if (names) names[i] = kTimerNames[i];
    if (ms) ms[i] = g->t_ms[i];
    if (launches) launches[i] = g->t_launches[i];
    if (units) units[i] = g->t_units[i];
  }
  return n;
}
int mg_step_begin(mg_graph *g) {
  MG_CHECK_HANDLE(g);
  if (!g->step_a) MG_CUDA(cudaEventCreate(&g->step_a));
  if (!g->step_b) MG_CUDA(cudaEventCreate(&g->step_b));
  MG_CUDA(cudaEventRecord(g->step_a, g->stream));
  return MG_OK;
}
int mg_step_end(mg_graph *g, double *elapsed_ms) {
  MG_CHECK_HANDLE(g);
  if (!g->step_a || !g->step_b || !elapsed_ms) { set_error("mg_step_end without mg_step_begin"); return MG_EINVAL; }
  MG_CUDA(cudaEventRecord(g->step_b, g->stream));
  MG_CUDA(cudaEventSynchronize(g->step_b));
  float ms = 0.f;
  MG_CUDA(cudaEventElapsedTime(&ms, g->step_a, g->step_b));
  *elapsed_ms = ms;
  return MG_OK;
}
void mg_timers_reset(mg_graph *g) {
  if (!g) return;
  cudaSetDevice(g->device);
  cudaStreamSynchronize(g->stream);
  flush_timers(g);
  for (int i = 0; i < (int)T_COUNT; ++i) { g->t_ms[i] = 0; g->t_launches[i] = 0; g->t_units[i] = 0; }
}

}  // extern "C"
