// device-side ingest of MEGAHIT's on-disk SDBG, straight into the device layout —
// replaces the host path SDBG::LoadFromFile -> LoadSdbgRawContent (reference
// libs/megahit/src/sdbg/sdbg_raw_content.cpp:18-96) + per-item getters + upload.
//
// Format (reference sdbg_meta.cpp:50-75, sdbg_writer.cpp:25-58, sdbg_item.h:14-24, sdbg_def.h:11-29):
//   <prefix>.sdbg_info   text: "k K", "words_per_tip_label L", "num_buckets B", "num_files F", then B lines
//                        "bucket_id file_id starting_offset num_items num_tips num_large_mul"
//                        (unused buckets carry the null id 2^64-1); items are logically ordered by
//                        bucket_id, physically by (file_id, starting_offset);
//   <prefix>.sdbg.<f>    records: 2-byte SdbgItem {w:4, last:1, tip:1, pad:2 | mul:8}, then a uint16
//                        multiplicity iff mul == 255, then L uint32 label words iff tip.
// Records have variable length, so a bucket is decoded sequentially — by ONE thread; the 65 536
// buckets are decoded in parallel.  The host only reads the files into pinned memory and parses the
// bucket table; every array the graph needs is produced on the device.
#include <algorithm>
#include <string>

#include "common.cuh"

namespace mg {

struct BucketDev {
  uint64_t byte_off;    // offset of the bucket's first record in the concatenated file bytes
  uint64_t byte_end;    // end of the bucket's file (decoding never reads past it)
  uint64_t first_item;  // accumulate_item_count
  uint32_t n_items;
  uint32_t first_tip;   // accumulate_tip_count (sdbg_raw_content.cpp:58-60)
};

// Sub-word writes at bucket borders go through atomicOr; whole 64-bit words are stored directly.
__device__ __forceinline__ void flush_or(uint64_t *dst, uint64_t v) {
  if (v) atomicOr((unsigned long long *)dst, (unsigned long long)v);
}

__global__ void k_ingest(const uint8_t *__restrict__ raw, const BucketDev *__restrict__ buckets, uint32_t n_buckets,
                         uint32_t words_per_tip_label, uint64_t *w, uint64_t *last, uint64_t *tip, uint16_t *mult,
                         uint32_t *labels, uint32_t n_tips_total, int *bad) {
  const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_buckets) return;
  const BucketDev B = buckets[b];
  uint64_t off = B.byte_off;
  uint64_t wacc = 0, lacc = 0, tacc = 0;
  uint32_t tip_no = B.first_tip;
  for (uint32_t i = 0; i < B.n_items; ++i) {
    if (off + 2 > B.byte_end) { *bad = 1; return; }
    const uint64_t idx = B.first_item + i;
    const uint32_t b0 = raw[off], b1 = raw[off + 1];
    off += 2;
    uint32_t m = b1;
    if (b1 == 255u) {  // kSmallMulSentinel: the full multiplicity follows
      if (off + 2 > B.byte_end) { *bad = 1; return; }
      m = (uint32_t)raw[off] | ((uint32_t)raw[off + 1] << 8);
      off += 2;
    }
    const uint32_t is_tip = (b0 >> 5) & 1u;
    if (is_tip) {  // the label feeds the k-mer lookup (reads.cu); words kept as stored
      if (off + (uint64_t)words_per_tip_label * 4 > B.byte_end || tip_no >= n_tips_total) { *bad = 1; return; }
      for (uint32_t j = 0; j < words_per_tip_label; ++j) {
        const uint8_t *q = raw + off + 4 * j;
        labels[(uint64_t)tip_no * words_per_tip_label + j] =
            (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
      }
      ++tip_no;
      off += (uint64_t)words_per_tip_label * 4;
    }
    mult[idx] = (uint16_t)m;
    wacc |= (uint64_t)(b0 & 15u) << ((idx & 15) * 4);
    lacc |= (uint64_t)((b0 >> 4) & 1u) << (idx & 63);
    tacc |= (uint64_t)is_tip << (idx & 63);
    if ((idx & 15) == 15) { flush_or(&w[idx >> 4], wacc); wacc = 0; }
    if ((idx & 63) == 63) { flush_or(&last[idx >> 6], lacc); flush_or(&tip[idx >> 6], tacc); lacc = tacc = 0; }
  }
  if (B.n_items) {
    const uint64_t idx = B.first_item + B.n_items - 1;
    flush_or(&w[idx >> 4], wacc);
    flush_or(&last[idx >> 6], lacc);
    flush_or(&tip[idx >> 6], tacc);
  }
  if (off > B.byte_end) *bad = 1;
}

struct BucketHost {
  uint64_t bucket_id, file_id, starting_offset, num_items, num_tips, num_large_mul;
};

static bool scan_field(FILE *f, const char *name, unsigned long long *v) {
  char buf[64];
  if (fscanf(f, "%63s %llu", buf, v) != 2) return false;
  return strcmp(buf, name) == 0;
}

}  // namespace mg

using namespace mg;

extern "C" int mg_graph_create_from_files(const char *prefix, int device, mg_graph **out) {
  if (!prefix || !out) { set_error("mg_graph_create_from_files: null argument"); return MG_EINVAL; }
  *out = nullptr;
  const uint64_t kNullId = ~0ull;
  // ---- bucket table (reference sdbg_meta.cpp:58-75)
  const std::string info = std::string(prefix) + ".sdbg_info";
  FILE *fi = fopen(info.c_str(), "r");
  if (!fi) { set_error("cannot open %s", info.c_str()); return MG_EINVAL; }
  unsigned long long k = 0, wptl = 0, nb = 0, nf = 0;
  if (!scan_field(fi, "k", &k) || !scan_field(fi, "words_per_tip_label", &wptl) || !scan_field(fi, "num_buckets", &nb) ||
      !scan_field(fi, "num_files", &nf) || nb == 0 || nb > (1u << 24) || nf > 4096) {
    fclose(fi);
    set_error("%s: malformed header", info.c_str());
    return MG_EINVAL;
  }
  std::vector<BucketHost> rec(nb);
  for (auto &r : rec) {
    unsigned long long a[6];
    if (fscanf(fi, "%llu %llu %llu %llu %llu %llu", &a[0], &a[1], &a[2], &a[3], &a[4], &a[5]) != 6) {
      fclose(fi);
      set_error("%s: malformed bucket table", info.c_str());
      return MG_EINVAL;
    }
    r = {a[0], a[1], a[2], a[3], a[4], a[5]};
  }
  fclose(fi);
  // logical order = bucket id (sdbg_meta.cpp:22-38); f_ from the bucket ids (sdbg.h:36-50)
  std::vector<BucketHost> used;
  for (auto &r : rec)
    if (r.bucket_id != kNullId && r.file_id != kNullId) used.push_back(r);
  std::sort(used.begin(), used.end(), [](const BucketHost &a, const BucketHost &b) { return a.bucket_id < b.bucket_id; });
  mg_graph_desc d;
  memset(&d, 0, sizeof d);
  for (int i = 0; i < 6; ++i) d.f[i] = 0;
  d.f[0] = -1;
  uint64_t n = 0, n_tips = 0;
  std::vector<uint64_t> first_item(used.size()), first_tip(used.size());
  std::vector<int64_t> pl_first(nb, 0), pl_second(nb, 0);  // prefix_look_up_ (sdbg.h:36,44-46)
  for (size_t i = 0; i < used.size(); ++i) {
    if (used[i].bucket_id >= nb || used[i].file_id >= nf) { set_error("%s: bucket record out of range", info.c_str()); return MG_EINVAL; }
    first_item[i] = n;
    first_tip[i] = n_tips;
    pl_first[used[i].bucket_id] = (int64_t)n;
    pl_second[used[i].bucket_id] = (int64_t)n + (int64_t)used[i].num_items - 1;
    n_tips += used[i].num_tips;
    n += used[i].num_items;
    d.f[used[i].bucket_id / (nb / 4) + 2] += (int64_t)used[i].num_items;
  }
  for (int i = 2; i < 6; ++i) d.f[i] += d.f[i - 1];
  if (n >= 0xFFFFFFFFull) { set_error("%s: %llu items; this build addresses < 2^32-1 items", info.c_str(), (unsigned long long)n); return MG_ELIMIT; }
  // ---- device, handle, stream (as mg_graph_create)
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("no CUDA device available (%s); libmcaat_gpu has no CPU fallback", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    return MG_ECUDA;
  }
  if (device < 0 || device >= ndev) { set_error("mg_graph_create_from_files: device %d of %d", device, ndev); return MG_EINVAL; }
  MG_CUDA(cudaSetDevice(device));
  // ---- file bytes -> pinned staging -> device
  std::vector<uint64_t> file_base(nf + 1, 0);
  std::vector<FILE *> fh(nf, nullptr);
  for (uint64_t f = 0; f < nf; ++f) {
    const std::string name = std::string(prefix) + ".sdbg." + std::to_string(f);
    fh[f] = fopen(name.c_str(), "rb");
    if (!fh[f]) {
      for (FILE *x : fh) if (x) fclose(x);
      set_error("cannot open %s", name.c_str());
      return MG_EINVAL;
    }
    fseek(fh[f], 0, SEEK_END);
    file_base[f + 1] = file_base[f] + (uint64_t)ftell(fh[f]);
    fseek(fh[f], 0, SEEK_SET);
  }
  const uint64_t total = file_base[nf];
  cudaStream_t s = nullptr;
  MG_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  g_alloc_stream = s;
  int rc = MG_OK;
  {
    DevBuf raw, bdev, w, last, tip, mult, bad, labels;
    const size_t chunk = 64u << 20;
    uint8_t *pin[2] = {nullptr, nullptr};
    cudaEvent_t done[2] = {nullptr, nullptr};
    auto cleanup = [&]() {
      for (FILE *x : fh) if (x) fclose(x);
      for (int i = 0; i < 2; ++i) { if (pin[i]) cudaFreeHost(pin[i]); if (done[i]) cudaEventDestroy(done[i]); }
    };
#define ING_TRY(call) do { rc = (call); if (rc != MG_OK) { cleanup(); raw.release(); bdev.release(); w.release(); last.release(); tip.release(); mult.release(); bad.release(); labels.release(); scratch_flush(s); cudaStreamSynchronize(s); cudaStreamDestroy(s); return rc; } } while (0)
    auto cuda_rc = [&](cudaError_t ce, const char *what) { if (ce != cudaSuccess) { set_error("%s -> %s", what, cudaGetErrorString(ce)); return (int)MG_ECUDA; } return (int)MG_OK; };
    ING_TRY(raw.alloc(total + 16));
    for (int i = 0; i < 2; ++i) {
      ING_TRY(cuda_rc(cudaHostAlloc((void **)&pin[i], chunk, cudaHostAllocDefault), "cudaHostAlloc"));
      ING_TRY(cuda_rc(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming), "cudaEventCreate"));
    }
    int slot = 0;
    for (uint64_t f = 0; f < nf; ++f) {  // double-buffered: the disk read of one chunk overlaps the copy of the other
      uint64_t pos = 0, len = file_base[f + 1] - file_base[f];
      while (pos < len) {
        const size_t want = (size_t)std::min<uint64_t>(chunk, len - pos);
        ING_TRY(cuda_rc(cudaEventSynchronize(done[slot]), "cudaEventSynchronize"));
        if (fread(pin[slot], 1, want, fh[f]) != want) { set_error("short read from %s.sdbg.%llu", prefix, (unsigned long long)f); ING_TRY(MG_EINVAL); }
        ING_TRY(cuda_rc(cudaMemcpyAsync(raw.as<uint8_t>() + file_base[f] + pos, pin[slot], want, cudaMemcpyHostToDevice, s), "cudaMemcpyAsync"));
        ING_TRY(cuda_rc(cudaEventRecord(done[slot], s), "cudaEventRecord"));
        pos += want;
        slot ^= 1;
      }
    }
    std::vector<BucketDev> bh(used.size());
    for (size_t i = 0; i < used.size(); ++i) {
      bh[i].byte_off = file_base[used[i].file_id] + used[i].starting_offset;
      bh[i].byte_end = file_base[used[i].file_id + 1];
      bh[i].first_item = first_item[i];
      bh[i].n_items = (uint32_t)used[i].num_items;
      bh[i].first_tip = (uint32_t)first_tip[i];
      if (used[i].num_items > 0xFFFFFFFFull || bh[i].byte_off > bh[i].byte_end) { set_error("%s: bucket %llu inconsistent", info.c_str(), (unsigned long long)used[i].bucket_id); ING_TRY(MG_EINVAL); }
    }
    const size_t w_words = (size_t)((n + 15) / 16) + 1, b_words = (size_t)((n + 63) / 64) + 1;
    ING_TRY(bdev.alloc(bh.size() * sizeof(BucketDev) + 16));
    ING_TRY(w.alloc(w_words * 8));
    ING_TRY(last.alloc(b_words * 8));
    ING_TRY(tip.alloc(b_words * 8));
    ING_TRY(mult.alloc((size_t)n * 2 + 16));
    ING_TRY(bad.alloc(16));
    if (n_tips >= 0xFFFFFFFFull) { set_error("%s: %llu tips", info.c_str(), (unsigned long long)n_tips); ING_TRY(MG_ELIMIT); }
    ING_TRY(labels.alloc((size_t)n_tips * wptl * 4 + 16));
    ING_TRY(cuda_rc(cudaMemsetAsync(w.p, 0, w.bytes, s), "memset"));
    ING_TRY(cuda_rc(cudaMemsetAsync(last.p, 0, last.bytes, s), "memset"));
    ING_TRY(cuda_rc(cudaMemsetAsync(tip.p, 0, tip.bytes, s), "memset"));
    ING_TRY(cuda_rc(cudaMemsetAsync(bad.p, 0, 16, s), "memset"));
    if (!bh.empty()) {
      ING_TRY(cuda_rc(cudaMemcpyAsync(bdev.p, bh.data(), bh.size() * sizeof(BucketDev), cudaMemcpyHostToDevice, s), "cudaMemcpyAsync"));
      k_ingest<<<(unsigned)((bh.size() + 63) / 64), 64, 0, s>>>(raw.as<uint8_t>(), bdev.as<BucketDev>(), (uint32_t)bh.size(),
                                                               (uint32_t)wptl, w.as<uint64_t>(), last.as<uint64_t>(),
                                                               tip.as<uint64_t>(), mult.as<uint16_t>(), labels.as<uint32_t>(),
                                                               (uint32_t)n_tips, (int *)bad.p);
      ++g_launches;
      ING_TRY(cuda_rc(cudaGetLastError(), "launch k_ingest"));
    }
    int hbad = 0;
    ING_TRY(cuda_rc(cudaMemcpyAsync(&hbad, bad.p, 4, cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync"));
    ING_TRY(cuda_rc(cudaStreamSynchronize(s), "cudaStreamSynchronize"));
    if (hbad) { set_error("%s: records run past the end of a file (bucket table and files disagree)", prefix); ING_TRY(MG_EINVAL); }
    raw.release();  // the raw bytes are no longer needed
    // ---- the arrays are on the device: build the resident graph from them
    mg_graph *g = new (std::nothrow) mg_graph();
    if (!g) ING_TRY(MG_ENOMEM);
    g->device = device;
    g->k = (uint32_t)k;
    g->stream = s;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) {
    g->sm_count = prop.multiProcessorCount;
    g->smem_per_sm = prop.sharedMemPerMultiprocessor;
    g->smem_per_block_optin = prop.sharedMemPerBlockOptin;
  }
    {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t keep = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      }
    }
    d.n_items = n;
    d.k = (uint32_t)k;
    d.memory = MG_MEM_DEVICE;
    d.device = device;
    d.w = w.as<uint64_t>();
    d.last = last.as<uint64_t>();
    d.tip = tip.as<uint64_t>();
    d.invalid = nullptr;  // tip | (w == 0), reference sdbg.h:34,56-60
    d.mult16 = mult.as<uint16_t>();
    rc = build_graph(g, &d);
    // the k-mer lookup index: tip flags + labels + bucket table; skipped for k beyond its limit
    if (rc == MG_OK && (uint32_t)k <= 128)
      rc = attach_kmer_index(g, tip.as<uint64_t>(), labels.as<uint32_t>(), n_tips, (uint32_t)wptl, pl_first.data(), pl_second.data(), nb);
    cleanup();
    if (rc != MG_OK) {
      w.release(); last.release(); tip.release(); mult.release(); bdev.release(); bad.release(); labels.release();
      mg_graph_destroy(g);  // also destroys the stream
      return rc;
    }
    flush_timers(g);
    *out = g;
  }
  return MG_OK;
}
