// mg_invalid_readback_noalloc: the validity-bitmap read-back (mg_invalid_readback, api.cu) for callers that must
// not reach the allocator — code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  It has no counterpart in the reference.
//
// The staging buffer that api.cu takes from the stream-ordered pool is replaced by a file-scope device array of
// fixed size (kNoAllocScratchBytes, 4 MiB: graphs of up to MG_NOALLOC_MAX_ITEMS = 2^25 items).  Every GPU that
// loads the library holds those 4 MiB until the process exits.  All calls on one device share the array, and it
// keeps the last call's contents: the kernel below overwrites every word the copy reads, so nothing has to be
// cleared, but two calls that overlap on the GPU (different graphs / streams of the same device) race on it.
#include <mutex>

#include "common.cuh"

namespace mg {
namespace {

constexpr size_t kNoAllocScratchBytes = (size_t)MG_NOALLOC_MAX_ITEMS / 8;  // one bit per item
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// In host code the name g_noalloc_scratch is a stand-in of the host side, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, once per device, with that device current.
std::mutex g_noalloc_mu;
uint64_t *g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

int noalloc_scratch(int device, uint64_t **out) {
  if (device < 0 || device >= kNoAllocMaxDevices) {
    set_error("mg_invalid_readback_noalloc: device %d outside 0..%d", device, kNoAllocMaxDevices - 1);
    return MG_EINVAL;
  }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void *p = nullptr;
    MG_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_noalloc_ptr[device] = (uint64_t *)p;
  }
  *out = g_noalloc_ptr[device];
  return MG_OK;
}

// same gather as k_inv_words (api.cu), bounded by the scratch capacity as well
__global__ void k_inv_words_noalloc(uint64_t n_fblocks, const FBlock *fb, uint64_t *out, uint64_t cap_words) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_fblocks * 4 && i < cap_words) out[i] = fb[i >> 2].inv[i & 3];
}

}  // namespace
}  // namespace mg

using namespace mg;

extern "C" int mg_invalid_readback_noalloc(mg_graph *g, uint64_t *host_words) {
  if (!g) {
    set_error("null graph handle");
    return MG_EINVAL;
  }
  if (!host_words) {
    set_error("mg_invalid_readback_noalloc: null output");
    return MG_EINVAL;
  }
  if (g->view.n > (uint64_t)MG_NOALLOC_MAX_ITEMS) {
    set_error("mg_invalid_readback_noalloc: graph has %llu items, the static scratch holds %llu (MG_NOALLOC_MAX_ITEMS); "
              "use mg_invalid_readback",
              (unsigned long long)g->view.n, (unsigned long long)MG_NOALLOC_MAX_ITEMS);
    return MG_ELIMIT;
  }
  MG_CUDA(cudaSetDevice(g->device));
  const uint64_t nb = g->view.n_fblocks;
  if (!nb) return MG_OK;
  if (nb * 32 > kNoAllocScratchBytes) {
    set_error("mg_invalid_readback_noalloc: %llu validity blocks exceed the static scratch", (unsigned long long)nb);
    return MG_ELIMIT;
  }
  uint64_t *scratch = nullptr;
  MG_TRY(noalloc_scratch(g->device, &scratch));
  MG_LAUNCH(k_inv_words_noalloc, (unsigned)((nb * 4 + 255) / 256), 256, 0, g->stream, nb, g->fb.as<FBlock>(), scratch,
            (uint64_t)(kNoAllocScratchBytes / 8));
  MG_CUDA(cudaMemcpyAsync(host_words, scratch, (size_t)((g->view.n + 63) / 64) * 8, cudaMemcpyDeviceToHost, g->stream));
  MG_CUDA(cudaStreamSynchronize(g->stream));
  return MG_OK;
}
