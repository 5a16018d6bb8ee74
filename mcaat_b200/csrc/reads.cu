// SURVEY 8(f) N3 — read -> node-id mapping on the device: get_reads / get_read_from_sequence, reference
// src/reads.cpp:57-130, on top of SDBG::IndexBinarySearch (sdbg.h:141-207).  The reference handles one read after
// the other on one core: two k-mer lookups decide whether the read touches a cycle node, then every k-mer of a kept
// read is looked up; a lookup is a binary search over a bucket of the graph whose every probe walks up to k
// Backward steps (rank/select).  Here one thread does one lookup and all lookups of a batch run concurrently:
//   k_member_bits   cycle node ids -> n-bit membership bitmap (nodes_of_cycles, :98-103)
//   k_read_ends     one thread per read: length rule (:64-66), first/last k-mer (:68-72), membership (:74-77)
//   (scan)          ids per kept read = len - k + 1 -> output offsets
//   k_read_kmers    one thread per output id: the remaining k-mers (:80-85)
// Latency-bound pointer chasing through the succinct index (≈ log2(bucket) probes x up to k Backward steps of
// ≈ 3 dependent loads); the machine is filled with independent lookups instead of shortening one.
#include <vector>

#include "common.cuh"
#include "kmer_dev.cuh"
#include "search_dev.cuh"

namespace mg {

static inline unsigned grid_for(const mg_graph *g, uint64_t n, int block) {
  uint64_t want = (n + block - 1) / block, cap = (uint64_t)g->sm_count * 32;
  return (unsigned)(want < cap ? (want ? want : 1) : cap);
}

__global__ void k_tip_popc(const uint64_t *tip, uint64_t n_words, uint32_t *counts) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_words; i += (uint64_t)gridDim.x * blockDim.x)
    counts[i] = (uint32_t)__popcll(tip[i]);
}

// ---- shortcut tables of the k-mer index (kmer_dev.cuh: KmerIndex::jump, bucket_ok)
__global__ void k_jump_init(GraphView g, KmerIndex K, uint32_t *dst) {
  for (uint64_t x = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; x < g.n; x += (uint64_t)gridDim.x * blockDim.x)
    dst[x] = jump_init(g, K, (uint32_t)x);
}
__global__ void k_jump_double(uint64_t n, const uint32_t *src, uint32_t *dst) {
  for (uint64_t x = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; x < n; x += (uint64_t)gridDim.x * blockDim.x)
    dst[x] = jump_double(src, (uint32_t)x);
}
__global__ void k_bucket_ok(GraphView g, KmerIndex K, int chars, uint32_t *ok) {
  const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  const bool good = b < K.n_buckets && bucket_consistent(g, K, b, chars);
  const unsigned m = __ballot_sync(0xffffffffu, good);
  if ((threadIdx.x & 31) == 0 && b < K.n_buckets) ok[b >> 5] = m;
}

__global__ void k_member_bits(const uint32_t *ids, uint64_t n_ids, uint32_t *member) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_ids; i += (uint64_t)gridDim.x * blockDim.x)
    bm_set_atomic(member, ids[i]);
}

__global__ void k_read_ends(GraphView g, KmerIndex K, const char *seqs, const uint64_t *seq_off, uint64_t n_reads,
                            int reverse, const uint32_t *member, uint32_t *first_id, uint32_t *last_id, uint32_t *out_len) {
  for (uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; r < n_reads; r += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t a = seq_off[r], b = seq_off[r + 1];
    uint32_t s = kNull, e = kNull;
    out_len[r] = read_ends(g, K, seqs + a, (uint32_t)(b - a), reverse != 0, member, &s, &e);
    first_id[r] = s;
    last_id[r] = e;
  }
}

// out_off: exclusive prefix of out_len (n_reads + 1 entries, the last one = total)
__global__ void k_read_kmers(GraphView g, KmerIndex K, const char *seqs, const uint64_t *seq_off, uint64_t n_reads,
                             int reverse, const uint32_t *first_id, const uint32_t *last_id, const uint32_t *out_off,
                             uint64_t total, uint64_t *ids) {
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t lo = 0, hi = n_reads;  // last read r with out_off[r] <= t
    while (hi - lo > 1) {
      uint64_t mid = (lo + hi) / 2;
      if (out_off[mid] <= t) lo = mid;
      else hi = mid;
    }
    const uint64_t r = lo;
    const uint32_t p = (uint32_t)(t - out_off[r]), n_out = out_off[r + 1] - out_off[r];
    const uint64_t a = seq_off[r];
    const uint32_t id = read_id_at(g, K, seqs + a, (uint32_t)(seq_off[r + 1] - a), reverse != 0, p, n_out, first_id[r], last_id[r]);
    ids[t] = id == kNull ? ~0ull : (uint64_t)id;  // kNullID stays the reference's 64-bit value
  }
}

// ---- index attachment -----------------------------------------------------------------------------------
int attach_kmer_index(mg_graph *g, const uint64_t *tip_dev, const uint32_t *labels_dev, uint64_t n_tips, uint32_t wptl,
                      const int64_t *first_host, const int64_t *second_host, uint64_t n_buckets) {
  const uint64_t n = g->view.n, n_words = (n + 63) / 64;
  cudaStream_t s = g->stream;
  if (n_buckets == 0 || n_buckets > (1u << 24)) {
    set_error("k-mer index: %llu buckets", (unsigned long long)n_buckets);
    return MG_EINVAL;
  }
  unsigned chars = 0;
  while ((1u << (chars * 2)) < n_buckets) ++chars;
  if (g->k < chars || g->k > (uint32_t)kMaxK) {
    set_error("k-mer index: k = %u outside [%u, %d]", g->k, chars, kMaxK);
    return MG_ELIMIT;
  }
  MG_TRY(g->kx_tip.alloc(n_words * 8 + 16));
  MG_TRY(g->kx_tip_prefix.alloc(n_words * 4 + 16));
  MG_TRY(g->kx_labels.alloc((size_t)n_tips * wptl * 4 + 16));
  MG_TRY(g->kx_buckets.alloc((size_t)n_buckets * 16 + 16));
  if (n_words) {
    MG_CUDA(cudaMemcpyAsync(g->kx_tip.p, tip_dev, n_words * 8, cudaMemcpyDeviceToDevice, s));
    MG_LAUNCH(k_tip_popc, grid_for(g, n_words, 256), 256, 0, s, g->kx_tip.as<uint64_t>(), n_words, g->kx_tip_prefix.as<uint32_t>());
    DevBuf total;
    MG_TRY(total.alloc(4));
    MG_TRY(exclusive_scan_u32(g->kx_tip_prefix.as<uint32_t>(), n_words, total.as<uint32_t>(), s));
    uint32_t tot = 0;
    MG_CUDA(cudaMemcpyAsync(&tot, total.p, 4, cudaMemcpyDeviceToHost, s));
    MG_CUDA(cudaStreamSynchronize(s));
    if (tot != n_tips) {
      set_error("k-mer index: %llu tip labels for %u tip flags", (unsigned long long)n_tips, tot);
      return MG_EINVAL;
    }
  }
  if (n_tips && wptl) MG_CUDA(cudaMemcpyAsync(g->kx_labels.p, labels_dev, (size_t)n_tips * wptl * 4, cudaMemcpyDeviceToDevice, s));
  MG_CUDA(cudaMemcpyAsync(g->kx_buckets.p, first_host, n_buckets * 8, cudaMemcpyHostToDevice, s));
  MG_CUDA(cudaMemcpyAsync(g->kx_buckets.as<int64_t>() + n_buckets, second_host, n_buckets * 8, cudaMemcpyHostToDevice, s));
  MG_CUDA(cudaStreamSynchronize(s));
  KmerIndex &K = g->kx;
  K.tip = g->kx_tip.as<uint64_t>();
  K.tip_prefix = g->kx_tip_prefix.as<uint32_t>();
  K.labels = g->kx_labels.as<uint32_t>();
  K.words_per_label = wptl;
  K.n_buckets = (uint32_t)n_buckets;
  K.first = g->kx_buckets.as<int64_t>();
  K.second = K.first + n_buckets;
  K.k = g->k;
  K.jump = nullptr;
  K.bucket_ok = nullptr;
  // shortcut tables: Backward^8 per item by pointer doubling (1 navigation pass + 3 gather passes) and the per-bucket
  // consistency bits; 4 B per item.  MCAAT_KMER_JUMP=0 keeps the literal search only.
  const char *je = getenv("MCAAT_KMER_JUMP");
  if (!(je && *je == '0') && chars >= (unsigned)kJumpSteps && g->k > (uint32_t)kJumpSteps && n) {
    DevBuf tmp;
    MG_TRY(g->kx_jump.alloc((size_t)n * 4 + 16));
    MG_TRY(tmp.alloc((size_t)n * 4 + 16));
    MG_TRY(g->kx_bucket_ok.alloc((size_t)((n_buckets + 31) / 32) * 4 + 16));
    uint32_t *a = g->kx_jump.as<uint32_t>(), *b = tmp.as<uint32_t>();
    const unsigned grid = grid_for(g, n, 256);
    MG_LAUNCH(k_jump_init, grid, 256, 0, s, g->view, K, b);   // 1 step  (in tmp)
    MG_LAUNCH(k_jump_double, grid, 256, 0, s, n, b, a);        // 2 steps (in jump)
    MG_LAUNCH(k_jump_double, grid, 256, 0, s, n, a, b);        // 4 steps (in tmp)
    MG_LAUNCH(k_jump_double, grid, 256, 0, s, n, b, a);        // 8 steps (in jump)
    MG_LAUNCH(k_bucket_ok, (unsigned)((n_buckets + 255) / 256), 256, 0, s, g->view, K, (int)chars, g->kx_bucket_ok.as<uint32_t>());
    MG_CUDA(cudaStreamSynchronize(s));
    K.jump = a;
    K.bucket_ok = g->kx_bucket_ok.as<uint32_t>();
    g->device_bytes += g->kx_jump.bytes + g->kx_bucket_ok.bytes;
  }
  g->has_kx = true;
  g->device_bytes += g->kx_tip.bytes + g->kx_tip_prefix.bytes + g->kx_labels.bytes + g->kx_buckets.bytes;
  return MG_OK;
}

int run_map_reads(mg_graph *g, const char *seqs, const uint64_t *seq_off, uint64_t n_reads, int reverse,
                  const uint32_t *cycle_ids_dev, uint64_t n_cycle_ids, mg_reads *out) {
  memset(out, 0, sizeof *out);
  if (!g->has_kx) {
    set_error("mg_map_reads: the graph has no k-mer index (create it with mg_graph_create_from_files or mg_kmer_index_create)");
    return MG_EINVAL;
  }
  const uint64_t n = g->view.n;
  cudaStream_t s = g->stream;
  const uint64_t n_chars = n_reads ? seq_off[n_reads] : 0;
  for (uint64_t r = 0; r < n_reads; ++r) {
    if (seq_off[r + 1] < seq_off[r] || seq_off[r + 1] - seq_off[r] > 0x7FFFFFFFull) {
      set_error("mg_map_reads: offsets of read %llu are not ascending / read longer than 2^31", (unsigned long long)r);
      return MG_EINVAL;
    }
  }
  // the id offsets are scanned in 32 bits: at most len - k + 1 <= len ids per read
  if (n_chars >= 0xFFFFFFFFull) {
    set_error("mg_map_reads: batch of %llu characters; pass batches below 2^32", (unsigned long long)n_chars);
    return MG_ELIMIT;
  }
  std::vector<uint32_t> h_off(n_reads + 1, 0);
  std::vector<uint64_t> h_ids;
  uint64_t total = 0;
  {
    TimedScope ts(g, T_READS, n_reads);
    DevBuf d_seqs, d_off, member, first_id, last_id, out_off, d_total, d_ids;
    MG_TRY(d_seqs.alloc(n_chars + 16));
    MG_TRY(d_off.alloc((n_reads + 1) * 8));
    MG_TRY(member.alloc((size_t)((n + 31) / 32) * 4 + 16));
    MG_TRY(first_id.alloc(n_reads * 4 + 16));
    MG_TRY(last_id.alloc(n_reads * 4 + 16));
    MG_TRY(out_off.alloc((n_reads + 1) * 4 + 16));
    MG_TRY(d_total.alloc(4));
    if (n_chars) MG_CUDA(cudaMemcpyAsync(d_seqs.p, seqs, n_chars, cudaMemcpyHostToDevice, s));
    const uint64_t zero = 0;
    MG_CUDA(cudaMemcpyAsync(d_off.p, n_reads ? seq_off : &zero, (n_reads + 1) * 8, cudaMemcpyHostToDevice, s));
    MG_CUDA(cudaMemsetAsync(member.p, 0, member.bytes, s));
    MG_CUDA(cudaMemsetAsync(out_off.p, 0, out_off.bytes, s));
    if (n_cycle_ids) MG_LAUNCH(k_member_bits, grid_for(g, n_cycle_ids, 256), 256, 0, s, cycle_ids_dev, n_cycle_ids, member.as<uint32_t>());
    if (n_reads) {
      MG_LAUNCH(k_read_ends, grid_for(g, n_reads, 128), 128, 0, s, g->view, g->kx, d_seqs.as<char>(), d_off.as<uint64_t>(), n_reads,
                reverse, member.as<uint32_t>(), first_id.as<uint32_t>(), last_id.as<uint32_t>(), out_off.as<uint32_t>());
      MG_TRY(exclusive_scan_u32(out_off.as<uint32_t>(), n_reads + 1, d_total.as<uint32_t>(), s));
      uint32_t tot = 0;
      MG_CUDA(cudaMemcpyAsync(&tot, d_total.p, 4, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaMemcpyAsync(h_off.data(), out_off.p, (n_reads + 1) * 4, cudaMemcpyDeviceToHost, s));
      MG_CUDA(cudaStreamSynchronize(s));
      total = tot;
      if (total) {
        MG_TRY(d_ids.alloc(total * 8));
        MG_LAUNCH(k_read_kmers, grid_for(g, total, 128), 128, 0, s, g->view, g->kx, d_seqs.as<char>(), d_off.as<uint64_t>(), n_reads,
                  reverse, first_id.as<uint32_t>(), last_id.as<uint32_t>(), out_off.as<uint32_t>(), total, d_ids.as<uint64_t>());
        h_ids.resize(total);
        MG_CUDA(cudaMemcpyAsync(h_ids.data(), d_ids.p, total * 8, cudaMemcpyDeviceToHost, s));
      }
      MG_CUDA(cudaStreamSynchronize(s));
    }
    g->t_units[T_READS] += total;  // lookups beyond the two per read
  }
  MG_CUDA(cudaStreamSynchronize(s));
  flush_timers(g);
  out->off = (uint64_t *)malloc((n_reads + 1) * 8);
  out->ids = (uint64_t *)malloc((total + 1) * 8);
  if (!out->off || !out->ids) {
    free(out->off);
    free(out->ids);
    memset(out, 0, sizeof *out);
    set_error("out of host memory for %llu read ids", (unsigned long long)total);
    return MG_ENOMEM;
  }
  uint64_t kept = 0;
  for (uint64_t r = 0; r <= n_reads; ++r) out->off[r] = h_off[r];
  for (uint64_t r = 0; r < n_reads; ++r) kept += h_off[r + 1] > h_off[r];
  if (total) memcpy(out->ids, h_ids.data(), total * 8);
  out->n_reads = n_reads;
  out->n_kept = kept;
  out->n_ids = total;
  return MG_OK;
}

}  // namespace mg
