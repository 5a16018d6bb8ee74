// Prefix sum and ordered stream compaction used by index construction and by the start-node
// selection (the reference's `critical { chunk.push_back(node) }` at threads=1 yields ascending
// ids, src/cycle_finder.cpp:414-416; here an ordered bitmap -> id-list compaction).
#include "common.cuh"

namespace mg {

constexpr int kScanThreads = 256;
constexpr int kScanPer = 8;
constexpr int kScanChunk = kScanThreads * kScanPer;

__global__ void k_scan_local(uint32_t *data, size_t n, uint32_t *aux) {
  __shared__ uint32_t warp_sums[kScanThreads / 32];
  size_t base = (size_t)blockIdx.x * kScanChunk + (size_t)threadIdx.x * kScanPer;
  uint32_t v[kScanPer];
  uint32_t sum = 0;
#pragma unroll
  for (int i = 0; i < kScanPer; ++i) {
    v[i] = base + i < n ? data[base + i] : 0;
    sum += v[i];
  }
  // inclusive scan of per-thread sums within the warp
  uint32_t incl = sum;
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += t;
    }
    if (lane < kScanThreads / 32) warp_sums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  uint32_t excl = incl - sum + (warp ? warp_sums[warp - 1] : 0);
#pragma unroll
  for (int i = 0; i < kScanPer; ++i) {
    if (base + i < n) data[base + i] = excl;
    excl += v[i];
  }
  if (threadIdx.x == kScanThreads - 1 && aux) aux[blockIdx.x] = excl;
}

__global__ void k_scan_add(uint32_t *data, size_t n, const uint32_t *aux) {
  size_t base = (size_t)blockIdx.x * kScanChunk + (size_t)threadIdx.x * kScanPer;
  uint32_t off = aux[blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanPer; ++i)
    if (base + i < n) data[base + i] += off;
}

__global__ void k_copy_u32(const uint32_t *src, uint32_t *dst) { *dst = *src; }

int exclusive_scan_u32(uint32_t *data, size_t n, uint32_t *total_dev, cudaStream_t stream) {
  if (n == 0) {
    if (total_dev) MG_CUDA(cudaMemsetAsync(total_dev, 0, 4, stream));
    return MG_OK;
  }
  size_t nb = (n + kScanChunk - 1) / kScanChunk;
  DevBuf aux;
  MG_TRY(aux.alloc((nb + 1) * sizeof(uint32_t)));
  MG_LAUNCH(k_scan_local, (unsigned)nb, kScanThreads, 0, stream, data, n, aux.as<uint32_t>());
  if (nb > 1) {
    MG_TRY(exclusive_scan_u32(aux.as<uint32_t>(), nb, total_dev, stream));
    MG_LAUNCH(k_scan_add, (unsigned)nb, kScanThreads, 0, stream, data, n, aux.as<uint32_t>());
  } else if (total_dev) {
    MG_LAUNCH(k_copy_u32, 1, 1, 0, stream, aux.as<uint32_t>(), total_dev);
  }
  MG_CUDA(cudaStreamSynchronize(stream));  // aux is freed on return
  return MG_OK;
}

__global__ void k_word_popc(const uint32_t *bits, size_t n_words, uint32_t *counts) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_words) counts[i] = __popc(bits[i]);
}
__global__ void k_scatter_ids(const uint32_t *bits, size_t n_words, const uint32_t *offs, uint32_t *ids) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_words) return;
  uint32_t w = bits[i];
  uint32_t o = offs[i];
  while (w) {
    int b = __ffs(w) - 1;
    ids[o++] = (uint32_t)(i * 32 + b);
    w &= w - 1;
  }
}

int bitmap_to_ids(const uint32_t *bits, uint64_t n, DevBuf &out_ids, uint64_t *count, cudaStream_t stream) {
  size_t n_words = (size_t)((n + 31) / 32);
  *count = 0;
  if (n_words == 0) return out_ids.ensure(16);
  DevBuf counts, total;
  MG_TRY(counts.alloc(n_words * 4));
  MG_TRY(total.alloc(4));
  unsigned grid = (unsigned)((n_words + 255) / 256);
  MG_LAUNCH(k_word_popc, grid, 256, 0, stream, bits, n_words, counts.as<uint32_t>());
  MG_TRY(exclusive_scan_u32(counts.as<uint32_t>(), n_words, total.as<uint32_t>(), stream));
  uint32_t tot = 0;
  MG_CUDA(cudaMemcpyAsync(&tot, total.p, 4, cudaMemcpyDeviceToHost, stream));
  MG_CUDA(cudaStreamSynchronize(stream));
  MG_TRY(out_ids.ensure((size_t)tot * 4 + 16));
  if (tot)
    MG_LAUNCH(k_scatter_ids, grid, 256, 0, stream, bits, n_words, counts.as<uint32_t>(), out_ids.as<uint32_t>());
  MG_CUDA(cudaStreamSynchronize(stream));
  *count = tot;
  return MG_OK;
}

}  // namespace mg
