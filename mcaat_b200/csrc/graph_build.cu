// K0: upload of the raw SDBG arrays and construction of the device-resident succinct graph
// (replaces SDBG::LoadFromFile's index construction: reference libs/megahit/src/sdbg/sdbg.h:26-61,
// kmlib/kmrns.h:122-167).  Layout: sdbg_dev.cuh.
#include "common.cuh"

namespace mg {

// one thread per WBlock: copy payload, count labels 1..4
__global__ void k_wblock_fill(const uint64_t *raw_w, uint64_t n, uint64_t n_wblocks, WBlock *wb,
                              uint32_t *cnt /* [4][n_wblocks] */) {
  uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_wblocks) return;
  uint32_t c[4];
  uint64_t words[kWWords];
  wblock_counts(raw_w, n, b, c, words);
#pragma unroll
  for (int j = 0; j < kWWords; ++j) wb[b].w[j] = words[j];
#pragma unroll
  for (int a = 0; a < 4; ++a) cnt[(size_t)a * n_wblocks + b] = c[a];
}
__global__ void k_wblock_header(uint64_t n_wblocks, WBlock *wb, const uint32_t *cnt) {
  uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_wblocks) return;
#pragma unroll
  for (int a = 0; a < 4; ++a) wb[b].cnt[a] = cnt[(size_t)a * n_wblocks + b];
}
// one thread per FBlock
__global__ void k_fblock_fill(const uint64_t *raw_w, const uint64_t *raw_last, const uint64_t *raw_tip,
                              const uint64_t *raw_inv, uint64_t n, uint64_t n_fblocks, FBlock *fb, uint32_t *cnt) {
  uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_fblocks) return;
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint64_t wi = b * 4 + j;
    uint64_t l = tail_mask_word(raw_last, n, wi);
    uint64_t t = tail_mask_word(raw_tip, n, wi);
    uint64_t iv = raw_inv ? tail_mask_word(raw_inv, n, wi) : (t | w_zero_mask(raw_w, n, wi));  // sdbg.h:34,56-60
    fb[b].last[j] = l;
    fb[b].lt[j] = l | t;
    fb[b].inv[j] = iv;
    c += __popcll(l);
  }
  fb[b].rank_before = 0;
#pragma unroll
  for (int j = 0; j < 7; ++j) fb[b].pad[j] = 0;
  cnt[b] = c;
}
__global__ void k_fblock_header(uint64_t n_fblocks, FBlock *fb, const uint32_t *cnt) {
  uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b < n_fblocks) fb[b].rank_before = cnt[b];
}
// select hints: block b holds occurrences [before, before + c); every multiple of 64 in that range
// gets hint = b
__global__ void k_hints_w(uint64_t n_wblocks, const WBlock *wb, const uint32_t *totals /* [4] */, uint32_t *h0,
                          uint32_t *h1, uint32_t *h2, uint32_t *h3) {
  uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_wblocks) return;
  uint32_t *hs[4] = {h0, h1, h2, h3};
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    uint32_t before = wb[b].cnt[a];
    uint32_t after = b + 1 < n_wblocks ? wb[b + 1].cnt[a] : totals[a];
    for (uint32_t j = (before + kSelH - 1) / kSelH; (uint64_t)j * kSelH < after; ++j) hs[a][j] = (uint32_t)b;
  }
}
__global__ void k_hints_last(uint64_t n_fblocks, const FBlock *fb, const uint32_t *total, uint32_t *h) {
  uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_fblocks) return;
  uint32_t before = fb[b].rank_before;
  uint32_t after = b + 1 < n_fblocks ? fb[b + 1].rank_before : *total;
  for (uint32_t j = (before + kSelH - 1) / kSelH; (uint64_t)j * kSelH < after; ++j) h[j] = (uint32_t)b;
}
// sdbg.h:52-54: rank_f_[i] = rs_last_.rank(f_[i] - 1), i = 1..5
__global__ void k_rank_f(GraphView g, int64_t *out) {
  int i = threadIdx.x;
  if (i >= 6) return;
  if (i == 0) { out[0] = 0; return; }
  int64_t pos = g.f[i] - 1;
  int64_t r;
  if (pos >= (int64_t)g.n) r = -1;                               // kmrns.h:288
  else if (pos == (int64_t)g.n - 1) r = (int64_t)g.total_last;   // kmrns.h:289-292
  else r = (int64_t)rank_last(g, pos);
  out[i] = r;
}
// small multiplicities + sorted (id, value) table -> uint16 (sdbg.h:96-105)
__global__ void k_mult_expand(const uint8_t *m8, uint64_t n, uint16_t *out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = m8[i];
}
__global__ void k_mult_large(const uint64_t *ids, const uint16_t *vals, uint64_t n_large, uint64_t n, uint16_t *out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_large && ids[i] < n) out[ids[i]] = vals[i];
}
__global__ void k_save_inv(uint64_t n_fblocks, const FBlock *fb, uint64_t *inv0) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_fblocks * 4) inv0[i] = fb[i >> 2].inv[i & 3];
}
__global__ void k_restore_inv(uint64_t n_fblocks, FBlock *fb, const uint64_t *inv0) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_fblocks * 4) fb[i >> 2].inv[i & 3] = inv0[i];
}

static inline unsigned grid_for(uint64_t n, int block) { return (unsigned)((n + block - 1) / block); }

// Stage one raw array on the device (copy from host, or alias the caller's device pointer).
struct Staged {
  DevBuf own;
  const void *ptr = nullptr;
  int stage(const void *src, size_t bytes, int memory, cudaStream_t s) {
    if (!src) { ptr = nullptr; return MG_OK; }
    if (memory == MG_MEM_DEVICE) { ptr = src; return MG_OK; }
    MG_TRY(own.alloc(bytes));
    MG_CUDA(cudaMemcpyAsync(own.p, src, bytes, cudaMemcpyHostToDevice, s));
    ptr = own.p;
    return MG_OK;
  }
};

int build_graph(mg_graph *g, const mg_graph_desc *d) {
  const uint64_t n = d->n_items;
  cudaStream_t s = g->stream;
  GraphView &v = g->view;
  v.n = n;
  v.n_wblocks = (n + kWNib - 1) / kWNib;
  v.n_fblocks = (n + kFItems - 1) / kFItems;
  for (int i = 0; i < 6; ++i) v.f[i] = d->f[i];
  const size_t w_words = (size_t)((n + 15) / 16), b_words = (size_t)((n + 63) / 64);

  Staged sw, sl, st, si, sm16, sm8, slid, slval;
  MG_TRY(sw.stage(d->w, w_words * 8, d->memory, s));
  MG_TRY(sl.stage(d->last, b_words * 8, d->memory, s));
  MG_TRY(st.stage(d->tip, b_words * 8, d->memory, s));
  MG_TRY(si.stage(d->invalid, b_words * 8, d->memory, s));

  // The multiplicities are three quarters of the upload (2 of 2.75 bytes per item) and no kernel of the index build reads
  // them: from host memory they travel on a stream of their own, behind the three arrays the build waits for, while
  // the build runs.
  cudaStream_t mult_stream = nullptr;
  cudaEvent_t mult_ev = nullptr;
  const bool mult_aside = (d->mult16 || d->mult8) && d->memory != MG_MEM_DEVICE && n > 0;
  if (mult_aside) {
    MG_TRY(g->mult.alloc((size_t)n * 2 + 16));
    if (!d->mult16) MG_TRY(sm8.own.alloc((size_t)n));
    MG_CUDA(cudaStreamCreateWithFlags(&mult_stream, cudaStreamNonBlocking));
    MG_CUDA(cudaEventCreateWithFlags(&mult_ev, cudaEventDisableTiming));
    MG_CUDA(cudaEventRecord(mult_ev, s));                  // the allocation and the three copies above come first
    MG_CUDA(cudaStreamWaitEvent(mult_stream, mult_ev, 0));
    if (d->mult16) {
      MG_CUDA(cudaMemcpyAsync(g->mult.p, d->mult16, (size_t)n * 2, cudaMemcpyHostToDevice, mult_stream));
    } else {  // the reference's own form (sdbg.h:96-105): one byte per item, expanded below
      MG_CUDA(cudaMemcpyAsync(sm8.own.p, d->mult8, (size_t)n, cudaMemcpyHostToDevice, mult_stream));
      sm8.ptr = sm8.own.p;
    }
    MG_CUDA(cudaEventRecord(mult_ev, mult_stream));
  }
  struct MultAside {  // joins and releases on every way out
    cudaStream_t &st; cudaEvent_t &ev; cudaStream_t main;
    ~MultAside() {
      if (ev) { cudaStreamWaitEvent(main, ev, 0); cudaStreamSynchronize(st); cudaEventDestroy(ev); }
      if (st) cudaStreamDestroy(st);
    }
  } mult_guard{mult_stream, mult_ev, s};
  TimedScope ts(g, T_BUILD, n);
  // --- W blocks
  MG_TRY(g->wb.alloc((size_t)(v.n_wblocks + 1) * sizeof(WBlock)));
  MG_CUDA(cudaMemsetAsync(g->wb.p, 0, g->wb.bytes, s));
  DevBuf cnt, totals;
  MG_TRY(cnt.alloc((size_t)4 * v.n_wblocks * 4 + 16));
  MG_TRY(totals.alloc(8 * 4));
  WBlock *wb = g->wb.as<WBlock>();
  MG_LAUNCH(k_wblock_fill, grid_for(v.n_wblocks, 128), 128, 0, s, (const uint64_t *)sw.ptr, n, v.n_wblocks, wb,
            cnt.as<uint32_t>());
  for (int a = 0; a < 4; ++a)
    MG_TRY(exclusive_scan_u32(cnt.as<uint32_t>() + (size_t)a * v.n_wblocks, v.n_wblocks, totals.as<uint32_t>() + a, s));
  MG_LAUNCH(k_wblock_header, grid_for(v.n_wblocks, 256), 256, 0, s, v.n_wblocks, wb, cnt.as<uint32_t>());
  // --- F blocks
  MG_TRY(g->fb.alloc((size_t)(v.n_fblocks + 1) * sizeof(FBlock)));
  MG_CUDA(cudaMemsetAsync(g->fb.p, 0, g->fb.bytes, s));
  DevBuf fcnt;
  MG_TRY(fcnt.alloc((size_t)v.n_fblocks * 4 + 16));
  FBlock *fb = g->fb.as<FBlock>();
  MG_LAUNCH(k_fblock_fill, grid_for(v.n_fblocks, 128), 128, 0, s, (const uint64_t *)sw.ptr, (const uint64_t *)sl.ptr,
            (const uint64_t *)st.ptr, (const uint64_t *)si.ptr, n, v.n_fblocks, fb, fcnt.as<uint32_t>());
  MG_TRY(exclusive_scan_u32(fcnt.as<uint32_t>(), v.n_fblocks, totals.as<uint32_t>() + 4, s));
  MG_LAUNCH(k_fblock_header, grid_for(v.n_fblocks, 256), 256, 0, s, v.n_fblocks, fb, fcnt.as<uint32_t>());
  uint32_t tot[5];
  MG_CUDA(cudaMemcpyAsync(tot, totals.p, sizeof tot, cudaMemcpyDeviceToHost, s));
  MG_CUDA(cudaStreamSynchronize(s));
  for (int a = 0; a < 4; ++a) v.total_w[a] = tot[a];
  v.total_last = tot[4];
  // --- select hints (one allocation: last, then labels 1..4)
  size_t hl = (size_t)tot[4] / kSelH + 2, hw[4], off[5];
  off[0] = hl;
  for (int a = 0; a < 4; ++a) {
    hw[a] = (size_t)tot[a] / kSelH + 2;
    off[a + 1] = off[a] + hw[a];
  }
  MG_TRY(g->hints.alloc(off[4] * 4));
  MG_CUDA(cudaMemsetAsync(g->hints.p, 0, g->hints.bytes, s));
  uint32_t *hbase = g->hints.as<uint32_t>();
  v.hint_last = hbase;
  for (int a = 0; a < 4; ++a) v.hint_w[a] = hbase + off[a];
  MG_LAUNCH(k_hints_w, grid_for(v.n_wblocks, 256), 256, 0, s, v.n_wblocks, wb, totals.as<uint32_t>(), hbase + off[0],
            hbase + off[1], hbase + off[2], hbase + off[3]);
  MG_LAUNCH(k_hints_last, grid_for(v.n_fblocks, 256), 256, 0, s, v.n_fblocks, fb, totals.as<uint32_t>() + 4, hbase);
  // --- multiplicities
  if (!mult_aside) MG_TRY(g->mult.alloc((size_t)n * 2 + 16));
  if (mult_aside && d->mult16) {
    // on its way (above); joined before the function returns
  } else if (d->mult16) {
    MG_CUDA(cudaMemcpyAsync(g->mult.p, d->mult16, (size_t)n * 2,
                            d->memory == MG_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
  } else if (d->mult8) {
    if (mult_aside) MG_CUDA(cudaStreamWaitEvent(s, mult_ev, 0));
    else MG_TRY(sm8.stage(d->mult8, (size_t)n, d->memory, s));
    MG_TRY(slid.stage(d->large_ids, (size_t)d->n_large * 8, d->memory, s));
    MG_TRY(slval.stage(d->large_vals, (size_t)d->n_large * 2, d->memory, s));
    if (n) MG_LAUNCH(k_mult_expand, grid_for(n, 256), 256, 0, s, (const uint8_t *)sm8.ptr, n, g->mult.as<uint16_t>());
    if (d->n_large)
      MG_LAUNCH(k_mult_large, grid_for(d->n_large, 256), 256, 0, s, (const uint64_t *)slid.ptr,
                (const uint16_t *)slval.ptr, d->n_large, n, g->mult.as<uint16_t>());
  } else {
    set_error("mg_graph_create: neither mult16 nor mult8 given");
    return MG_EINVAL;
  }
  v.wb = wb;
  v.fb = fb;
  v.mult = g->mult.as<uint16_t>();
  // --- rank_f
  DevBuf rf;
  MG_TRY(rf.alloc(6 * 8));
  MG_LAUNCH(k_rank_f, 1, 32, 0, s, v, rf.as<int64_t>());
  MG_CUDA(cudaMemcpyAsync(v.rank_f, rf.p, 6 * 8, cudaMemcpyDeviceToHost, s));
  // --- pristine invalid bitmap (mg_graph_reset) and visited bits
  MG_TRY(g->inv0.alloc((size_t)v.n_fblocks * 4 * 8 + 16));
  if (v.n_fblocks)
    MG_LAUNCH(k_save_inv, grid_for(v.n_fblocks * 4, 256), 256, 0, s, v.n_fblocks, fb, g->inv0.as<uint64_t>());
  MG_TRY(g->visited.alloc((size_t)((n + 31) / 32) * 4 + 16));
  MG_CUDA(cudaMemsetAsync(g->visited.p, 0, g->visited.bytes, s));
  if (mult_aside) MG_CUDA(cudaStreamWaitEvent(s, mult_ev, 0));
  MG_CUDA(cudaStreamSynchronize(s));
  g->device_bytes = g->wb.bytes + g->fb.bytes + g->mult.bytes + g->hints.bytes + g->inv0.bytes + g->visited.bytes;
  return MG_OK;
}

int reset_graph(mg_graph *g) {
  GraphView &v = g->view;
  cudaStream_t s = g->stream;
  if (v.n_fblocks)
    MG_LAUNCH(k_restore_inv, grid_for(v.n_fblocks * 4, 256), 256, 0, s, v.n_fblocks, g->fb.as<FBlock>(),
              g->inv0.as<uint64_t>());
  MG_CUDA(cudaMemsetAsync(g->visited.p, 0, g->visited.bytes, s));
  MG_CUDA(cudaStreamSynchronize(s));
  g->pruned = false;
  return MG_OK;
}

}  // namespace mg
