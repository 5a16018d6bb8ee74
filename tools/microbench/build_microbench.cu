// Micro-benchmark of the ball extraction (ball_dev.cuh: ball_build, dball_build) on ONE warp of an otherwise idle
// GPU: the navigation cache and multiplicities of a golden graph (dumped by run_build_microbench.py through tests/emu)
// and a start node; prints the time of one extraction, per node and per round of the level loop.
// Diagnostic tool, not part of the library.
#include <cstdio>
#include <cstdlib>
#include <unordered_set>
#include <vector>
#include "../../mcaat_b200/csrc/ball_layout.cuh"
using namespace mg;

struct Out { unsigned long long clk_build, clk_dbuild, clk_clear; uint32_t n_nodes, n_dnodes; int st, dst; };

__global__ void __launch_bounds__(32) k_bench(GraphView g, const uint64_t *adj, const uint32_t *visited, uint32_t start,
                                              int max_len, uint8_t *arena, BallLayout L, int reps, Out *out) {
  extern __shared__ __align__(16) uint8_t smem[];
  const BallArena A = ball_arena_of(arena, smem, L);
  Coop co;
  co.lane = threadIdx.x;
  uint32_t live = 0;
  Out o{};
  for (int r = 0; r < reps; ++r) {
    long long t0 = clock64();
    ball_clear(A, live, co);
    long long t1 = clock64();
    o.st = ball_build(g, adj, visited, start, max_len, A, live, co);
    long long t2 = clock64();
    o.clk_clear = (unsigned long long)(t1 - t0);
    o.clk_build = (unsigned long long)(t2 - t1);
    o.n_nodes = live;
  }
  for (int r = 0; r < reps; ++r) {
    ball_clear(A, live, co);
    long long t1 = clock64();
    o.dst = dball_build(adj, start, max_len, A, live, co);
    long long t2 = clock64();
    o.clk_dbuild = (unsigned long long)(t2 - t1);
    o.n_dnodes = live;
  }
  ball_clear(A, live, co);
  if (threadIdx.x == 0) *out = o;
}

template <class T> static std::vector<T> slurp(const char *path) {
  FILE *f = fopen(path, "rb");
  if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
  fseek(f, 0, SEEK_END);
  long bytes = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<T> v(bytes / sizeof(T));
  if (fread(v.data(), sizeof(T), v.size(), f) != v.size()) exit(2);
  fclose(f);
  return v;
}

int main(int argc, char **argv) {
  if (argc < 5) { fprintf(stderr, "usage: %s adj.bin mult.bin start max_len [reps]\n", argv[0]); return 2; }
  std::vector<uint64_t> adj = slurp<uint64_t>(argv[1]);
  std::vector<uint16_t> mult = slurp<uint16_t>(argv[2]);
  const uint32_t start = (uint32_t)atoll(argv[3]);
  const int max_len = atoi(argv[4]), reps = argc > 5 ? atoi(argv[5]) : 3;
  const size_t n = adj.size();
  // level sizes of the ball (host BFS with the same filter, visited all clear)
  std::vector<uint32_t> vis((n + 31) / 32 + 1, 0);
  GraphView hg{};
  hg.n = n;
  hg.mult = mult.data();
  std::vector<uint32_t> level{start}, sizes;
  std::unordered_set<uint32_t> seen{start};
  const uint32_t mu = mult[start];
  for (int d = 0; d < max_len && !level.empty(); ++d) {
    sizes.push_back((uint32_t)level.size());
    std::vector<uint32_t> next;
    for (uint32_t x : level) {
      uint32_t nb[4], okn[4], ns = 0;
      int deg = adj_outgoing(adj[x], nb);
      if (deg > 4) deg = 4;
      filtered_set(hg, vis.data(), x, mu, nb, deg, ns, okn);
      for (uint32_t j = 0; j < ns; ++j)
        if (seen.insert(okn[j]).second) next.push_back(okn[j]);
    }
    level.swap(next);
  }
  uint64_t rounds = 0, widest = 0;
  for (uint32_t s : sizes) { rounds += (s + 32 * kBuildK - 1) / (32 * kBuildK); widest = s > widest ? s : widest; }
  printf("ball of %u: %zu nodes on %zu levels (widest %llu), %llu rounds of %d nodes\n", start, seen.size(), sizes.size(),
         (unsigned long long)widest, (unsigned long long)rounds, 32 * kBuildK);

  uint64_t *d_adj; uint16_t *d_mult; uint32_t *d_vis; uint8_t *arena; Out *d_out;
  cudaMalloc(&d_adj, n * 8); cudaMalloc(&d_mult, n * 2); cudaMalloc(&d_vis, vis.size() * 4); cudaMalloc(&d_out, sizeof(Out));
  cudaMemcpy(d_adj, adj.data(), n * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(d_mult, mult.data(), n * 2, cudaMemcpyHostToDevice);
  cudaMemset(d_vis, 0, vis.size() * 4);
  const BallLayout L = make_ball_layout(max_len, 16, 27 * 1024);
  cudaMalloc(&arena, L.bytes);
  cudaMemset(arena, 0xFF, L.bytes);
  GraphView g{};
  g.n = n;
  g.mult = d_mult;
  cudaFuncSetAttribute(k_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
  k_bench<<<1, 32, L.smem_bytes>>>(g, d_adj, d_vis, start, max_len, arena, L, reps, d_out);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { fprintf(stderr, "%s\n", cudaGetErrorString(e)); return 1; }
  Out o;
  cudaMemcpy(&o, d_out, sizeof o, cudaMemcpyDeviceToHost);
  int mhz = 0;
  cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, 0);
  const double ns = 1e6 / (double)mhz;  // per clock
  printf("ball_build   status %d nodes %u: %.1f us  (%.0f ns/node, %.2f us/round)   ball_clear %.1f us\n", o.st, o.n_nodes,
         o.clk_build * ns / 1e3, o.clk_build * ns / o.n_nodes, o.clk_build * ns / 1e3 / (double)rounds, o.clk_clear * ns / 1e3);
  printf("dball_build  status %d nodes %u: %.1f us  (%.0f ns/node)\n", o.dst, o.n_dnodes, o.clk_dbuild * ns / 1e3,
         o.clk_dbuild * ns / (o.n_dnodes ? o.n_dnodes : 1));
  return 0;
}
