// Micro-benchmark of the contracted-ball search (ball_dev.cuh: sball_dfs) on ONE warp:
// a contracted ball dumped by tests/emu (tools/microbench/run_dfs_microbench.py) is searched `reps` times
// with each variant; prints cycles and ns per loop iteration.  Diagnostic tool, not part of the library.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../mcaat_b200/csrc/ball_dev.cuh"
using namespace mg;

struct Out { unsigned long long cycles; uint32_t iters, steps, n_cycles, status, relax_pops, expansions; };

template <int VARIANT>  // 0 records in shared memory, 2 records in global memory
__global__ void __launch_bounds__(32) k_bench(const SHot *hot_in, const uint64_t *in4_in, uint32_t n, int min_len, int max_len,
                                              SHot *hot_g, uint64_t *in4_g, uint16_t *cyc, uint32_t *cyc_pos, uint32_t *cyc_off,
                                              uint32_t *cyc_noff, uint32_t *relax, Lvl *lvl_g, Out *out) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int lane = threadIdx.x;
  const uint32_t lvl_bytes = (uint32_t)(max_len + 4) * 16, relax_bytes = 128 * 4;
  Lvl *lvl_s = (Lvl *)smem;
  uint32_t *rtop = (uint32_t *)(smem + lvl_bytes);
  SHot *hot_s = (SHot *)(smem + lvl_bytes + relax_bytes);
  uint64_t *in4_s = (uint64_t *)(smem + lvl_bytes + relax_bytes + (size_t)n * 16);
  for (uint32_t i = lane; i < n; i += 32) {
    hot_s[i] = hot_in[i]; in4_s[i] = in4_in[i];
    hot_g[i] = hot_in[i]; in4_g[i] = in4_in[i];
  }
  __syncwarp();
  BallArena A{};
  A.cyc = cyc; A.cyc_pos = cyc_pos; A.cyc_off = cyc_off; A.cyc_noff = cyc_noff; A.cyc_cap = 500u * (max_len + 1);
  A.relax = relax; A.relax_cap = 65536; A.relax_top = rtop; A.relax_top_cap = 128; A.max_steps = kMaxSteps;
  A.lvl = (VARIANT < 2) ? lvl_s : lvl_g;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  const RecsSmem Qs{sbase + lvl_bytes + relax_bytes, sbase + lvl_bytes + relax_bytes + n * 16u, sbase, sbase + lvl_bytes};
  const RecsPtr Qp{hot_g, in4_g, lvl_g, rtop};
  SDfsState ds;
  EnumResult R{};
  const long long t0 = clock64();
  bool fin = false;
  while (!fin) {
    if (VARIANT == 0) { if (lane == 0) fin = sball_dfs(min_len, max_len, A, Qs, ds, 1024, R); fin = __shfl_sync(0xffffffffu, (int)fin, 0); }
    if (VARIANT == 2) { if (lane == 0) fin = sball_dfs(min_len, max_len, A, Qp, ds, 1024, R); fin = __shfl_sync(0xffffffffu, (int)fin, 0); }
  }
  const long long t1 = clock64();
  if (lane == 0) *out = Out{(unsigned long long)(t1 - t0), ds.iters, R.steps, R.n_cycles, (uint32_t)R.status, ds.relax_pops, R.expansions};
}

int main(int argc, char **argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s ball.bin [reps]\n", argv[0]); return 2; }
  FILE *f = fopen(argv[1], "rb");
  if (!f) return 2;
  uint32_t hdr[4];
  if (fread(hdr, 4, 4, f) != 4) return 2;
  const uint32_t n = hdr[0];
  const int min_len = (int)hdr[1], max_len = (int)hdr[2];
  std::vector<SHot> hot(n);
  std::vector<uint64_t> in4(n);
  if (fread(hot.data(), 16, n, f) != n || fread(in4.data(), 8, n, f) != n) return 2;
  fclose(f);
  const int reps = argc > 2 ? atoi(argv[2]) : 5;
  SHot *d_hot, *d_hot_g; uint64_t *d_in4, *d_in4_g; uint16_t *cyc; uint32_t *cyc_pos, *cyc_off, *cyc_noff, *relax; Lvl *lvl; Out *out;
  cudaMalloc(&d_hot, n * 16); cudaMalloc(&d_hot_g, n * 16); cudaMalloc(&d_in4, n * 8); cudaMalloc(&d_in4_g, n * 8);
  cudaMalloc(&cyc, 500 * (max_len + 1) * 2); cudaMalloc(&cyc_pos, 500 * (max_len + 1) * 4); cudaMalloc(&cyc_off, 501 * 4);
  cudaMalloc(&cyc_noff, 501 * 4); cudaMalloc(&relax, 65536 * 4); cudaMalloc(&lvl, (max_len + 4) * 16); cudaMalloc(&out, sizeof(Out));
  cudaMemcpy(d_hot, hot.data(), n * 16, cudaMemcpyHostToDevice);
  cudaMemcpy(d_in4, in4.data(), n * 8, cudaMemcpyHostToDevice);
  const size_t smem = (size_t)(max_len + 4) * 16 + 512 + (size_t)n * 24;
  int clock_khz = 0;
  cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0);
  const char *names[4] = {"smem", "", "global", ""};
  for (int v = 0; v < 4; v += 2) {
    for (int r = 0; r < reps; ++r) {
#define LAUNCH(V) cudaFuncSetAttribute(k_bench<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
                  k_bench<V><<<1, 32, smem>>>(d_hot, d_in4, n, min_len, max_len, d_hot_g, d_in4_g, cyc, cyc_pos, cyc_off, cyc_noff, relax, lvl, out)
      if (v == 0) { LAUNCH(0); } else { LAUNCH(2); }
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", names[v], cudaGetErrorString(e)); return 1; }
      Out o;
      cudaMemcpy(&o, out, sizeof o, cudaMemcpyDeviceToHost);
      if (r == reps - 1)
        printf("%-14s chains %u iters %u steps %u cycles_found %u status %u relax_pops %u expansions %u: %.1f clk/iter, %.1f ns/iter (nominal %d MHz)\n",
               names[v], n, o.iters, o.steps, o.n_cycles, o.status, o.relax_pops, o.expansions, (double)o.cycles / o.iters,
               (double)o.cycles / o.iters / (clock_khz / 1e6), clock_khz / 1000);
    }
  }
  return 0;
}
