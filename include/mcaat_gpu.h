/* libmcaat_gpu — C ABI of the B200-native (CUDA sm_100a) implementation of MCAAT's bounded-length
 * multicycle search.  This is the drop-in boundary: the C++ `CycleFinder` shim
 * (mcaat_b200/host/cycle_finder_gpu.cpp) and the Python binding (mcaat_b200/gpu.py) call only
 * these entry points.  Plain pointers and sizes; no exceptions, no STL, no torch types.
 *
 * The reference has no FFI of its own: the interface replaced is the C++ class
 * `CycleFinder` (reference include/cycle_finder.h:25-85) and the MEGAHIT `SDBG` navigation API it
 * calls (reference libs/megahit/src/sdbg/sdbg.h:63-121,245-370).  Each entry point below cites the
 * reference code whose behaviour it reproduces bit-exactly.
 *
 * Conventions: every function returns 0 on success or a negative MG_E* code (mg_last_error() gives
 * the text); all calls are blocking; a handle is used from one thread at a time; output arrays
 * that the library allocates are released with mg_free().  Node ids are uint64_t at the ABI
 * (as in the reference) and must be < 2^32 - 1 items per graph (the reference itself is only
 * well defined below 2^31, reference src/cycle_finder.cpp:192-199).
 * There is NO CPU fallback: without a CUDA device every compute entry fails with MG_ECUDA.
 */
#ifndef MCAAT_GPU_H
#define MCAAT_GPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MG_OK 0
#define MG_EINVAL (-1)   /* bad argument */
#define MG_ECUDA (-2)    /* CUDA runtime error / no device */
#define MG_ENOMEM (-3)   /* host or device allocation failed */
#define MG_ELIMIT (-4)   /* graph or parameter outside the supported range */
#define MG_EINTERNAL (-5)

#define MG_MEM_HOST 0
#define MG_MEM_DEVICE 1

typedef struct mg_graph mg_graph;

/* Raw SDBG arrays in MEGAHIT's in-memory packing (reference
 * libs/megahit/src/sdbg/sdbg_raw_content.h:16-26, kmlib/kmcompactvector.h:53-58):
 *   w        4-bit edge labels, 16 per uint64 word, item i in bits [4*(i%16), 4*(i%16)+4) of word i/16
 *   last/tip 1 bit per item, item i = bit i%64 of word i/64
 *   invalid  optional initial invalid bitmap (same packing); NULL => tip | (w==0)
 *            (reference sdbg.h:34,56-60)
 *   mult16   uint16 multiplicity per item, or NULL in which case mult8 (sentinel 255) plus the
 *            id-sorted (large_ids, large_vals) table is used (reference sdbg.h:96-105)
 *   f        character-class boundaries f_[0..5] (reference sdbg.h:36-50)
 * With memory == MG_MEM_DEVICE all array pointers are device pointers on `device` (this is the
 * path used after an NCCL broadcast of the raw arrays); the library copies what it needs. */
typedef struct {
  uint64_t n_items;
  uint32_t k;
  int32_t memory;
  int32_t device;
  int32_t reserved;
  const uint64_t *w;
  const uint64_t *last;
  const uint64_t *tip;
  const uint64_t *invalid;
  const uint16_t *mult16;
  const uint8_t *mult8;
  const uint64_t *large_ids;
  const uint16_t *large_vals;
  uint64_t n_large;
  int64_t f[6];
} mg_graph_desc;

/* Upload + build the device-resident succinct graph (replaces SDBG::LoadFromFile's index
 * construction, reference sdbg.h:26-61 and kmlib/kmrns.h:122-167). */
int mg_graph_create(const mg_graph_desc *desc, mg_graph **out);
/* Same graph straight from MEGAHIT's on-disk SDBG (<prefix>.sdbg_info + <prefix>.sdbg.<i>): the files are
 * read into pinned memory and decoded ON THE DEVICE, bucket-parallel, into the resident layout — replaces
 * SDBG::LoadFromFile / LoadSdbgRawContent (reference sdbg.h:26-61, sdbg_raw_content.cpp:18-96,
 * sdbg_meta.cpp:50-75; record layout sdbg_item.h:14-24, sdbg_writer.cpp:25-58).  SURVEY 8(f) N2. */
int mg_graph_create_from_files(const char *prefix, int device, mg_graph **out);
void mg_graph_destroy(mg_graph *g);
/* Restore the invalid bitmap to its state at creation and clear `visited` (lets one resident
 * graph be searched repeatedly with different parameters). */
int mg_graph_reset(mg_graph *g);
uint64_t mg_graph_size(const mg_graph *g);
/* bytes of device memory held by the resident graph (arrays + rank/select samples) */
uint64_t mg_graph_device_bytes(const mg_graph *g);

/* SDBG navigation for items [first, first+count): EdgeOutdegree / EdgeIndegree / OutgoingEdges /
 * IncomingEdges / EdgeMultiplicity of reference sdbg.h:328-357,96-105 evaluated on the CURRENT
 * invalid bitmap.  Host output arrays; out_ids/in_ids are [count][4], unused slots = UINT64_MAX;
 * any output pointer may be NULL. */
int mg_navigate(mg_graph *g, uint64_t first, uint64_t count, int8_t *outdeg, int8_t *indeg,
                uint64_t *out_ids, uint64_t *in_ids, uint16_t *mult);

/* Graph pruning: CollectTips -> InvalidateMultiplicityOneNodes -> RecursiveReduction per tip ->
 * valid-edge count -> CollectTips (reference src/cycle_finder.cpp:436-452). */
typedef struct {
  uint64_t n_tips_before; /* "gathered tips" (:437) */
  uint64_t n_mult1;       /* "Pre-filter: invalidated" (:381) */
  uint64_t n_tips_after;  /* "After pruning, tips" (:452) */
  uint64_t n_valid;       /* "valid edges" (:452) */
  uint64_t n_peeled;      /* nodes invalidated by the reduction itself */
} mg_prune_stats;
int mg_prune(mg_graph *g, mg_prune_stats *stats);

/* CollectTips (reference src/cycle_finder.cpp:346-357) on the CURRENT invalid bitmap: the valid items whose
 * out-degree is 0, ascending ids, found and compacted on the device.  *ids is library allocated (mg_free). */
int mg_collect_tips(mg_graph *g, uint64_t **ids, uint64_t *n);

/* Current invalid bitmap -> host words ((n+63)/64 uint64), for write-back into the host SDBG
 * (reference SetInvalidEdge, sdbg.h:94). */
int mg_invalid_readback(mg_graph *g, uint64_t *host_words);
/* mg_invalid_readback without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The staging buffer is a static 4 MiB device array inside the library (held by every GPU that
 * loads it until the process exits), so graphs of more than MG_NOALLOC_MAX_ITEMS items are refused with
 * MG_ELIMIT.  Calls no cudaMalloc / cudaFree / pool routine.  All calls on one device share the array: calls
 * that overlap on the GPU (two graphs or streams of the same device) race on it. */
#define MG_NOALLOC_MAX_ITEMS (1ull << 25)
int mg_invalid_readback_noalloc(mg_graph *g, uint64_t *host_words);
/* The other direction: ORs the host SDBG's invalid bits (!IsValidEdge, sdbg.h:88-92; (n+63)/64 uint64 words)
 * into the device bitmap, so that a device graph kept resident across stages follows SetInvalidEdge calls the
 * host made in between.  *n_new (may be NULL) = bits that were not set on the device.  When anything changed,
 * mg_prune has to run again before mg_dls / mg_find_cycles. */
int mg_invalid_merge(mg_graph *g, const uint64_t *host_words, uint64_t *n_new);

/* Candidate start nodes: valid AND indegree >= 2 AND multiplicity > threshold AND no self-loop
 * incoming edge (reference src/cycle_finder.cpp:398-408), ascending ids.  *ids is library
 * allocated (mg_free). */
int mg_candidates(mg_graph *g, uint64_t threshold_multiplicity, uint64_t **ids, uint64_t *n);

/* DepthLevelSearch(node, node, limit) for each given node (reference src/cycle_finder.cpp:248-343);
 * pass[i] = 1/0.  expansions (optional) receives the number of expanded nodes. */
int mg_dls(mg_graph *g, const uint64_t *ids, uint64_t n, int limit, uint8_t *pass, uint64_t *expansions);

/* ChunkStartNodes at threads=1 (reference src/cycle_finder.cpp:387-427): candidates, admission
 * search, bucketing by ceil(log2(multiplicity)); buckets in descending key order, ascending ids
 * inside a bucket. */
typedef struct {
  uint64_t n_candidates;
  uint64_t n_starts;
  uint64_t n_buckets;
  int64_t *bucket_key;  /* [n_buckets] descending */
  uint64_t *bucket_off; /* [n_buckets+1] offsets into start_ids */
  uint64_t *start_ids;  /* [n_starts] */
  uint64_t dls_expansions;
} mg_starts;
int mg_select_starts(mg_graph *g, uint64_t threshold_multiplicity, int max_len, mg_starts *out);
/* Same bucketing for an externally supplied set of admitted nodes (multi-GPU: every rank ran
 * mg_dls on its shard of the candidates). */
int mg_bucket_starts(mg_graph *g, const uint64_t *ids, uint64_t n, mg_starts *out);
void mg_starts_free(mg_starts *s);

/* Per-start bounded multicycle enumeration with the sequential `visited` coupling of the
 * reference driver loop (reference src/cycle_finder.cpp:468-487, FindCycleUtil :231-243,
 * FindCycle :140-225): start nodes are taken in the given order; a start already marked visited
 * is skipped; results hold the raw per-start cycle sequence (emission order and duplicates). */
typedef struct {
  uint64_t n_searched;      /* entries of `results`, in insertion order */
  uint64_t *searched_start; /* [n_searched] */
  uint64_t *cyc_off;        /* [n_searched+1] offsets into the cycle list */
  uint64_t n_cycles;
  uint64_t *node_off;       /* [n_cycles+1] offsets into nodes */
  uint64_t *nodes;
  uint64_t n_skipped;       /* starts skipped because already visited (:476) */
  uint64_t n_bound_hits;    /* searches cleared by the 500-cycle cluster bound (:162-165) */
  uint64_t n_cap_hits;      /* searches stopped by the 10 000 000-step cap (:148-151) */
  uint64_t n_reruns;        /* speculative searches repeated at commit time (saw a stale bitmap) */
  uint64_t n_rounds;        /* pipeline launches (1 unless an arena or the result pool had to grow) */
  uint64_t n_searches_run;  /* searches started, speculative ones included */
  uint64_t n_abandoned;     /* running searches given up because an earlier commit marked their start */
  uint64_t expansions;      /* _GetOutgoings/_GetIncomings evaluations of committed searches */
  uint64_t n_out_of_order;  /* searches committed ahead of their turn: no earlier start's search could touch what they read */
} mg_cycles;
int mg_find_cycles(mg_graph *g, const uint64_t *starts, uint64_t n, int min_len, int max_len, mg_cycles *out);
void mg_cycles_free(mg_cycles *c);

/* Multi-GPU support: mark nodes visited / read the visited bitmap, so that ranks that ran
 * disjoint parts of the start list can exchange their commits. */
int mg_visited_readback(mg_graph *g, uint64_t *host_words);

/* ---- region extraction around the cycles ------------------------------------
 * Replaces keep_crispr_regions_extended_by_k(sdbg, k, cycles), reference src/spacer_ordering.cpp:78-139
 * (declared include/spacer_ordering.h:127-131; called through get_crispr_regions_extended_by_k,
 * src/spacer_ordering.cpp:177-184, with k = read length in nodes, src/main_run_and_debug.cpp:40-41).
 * `ids` = the node ids of all cycles, flattened (duplicates and already-invalid ids allowed, as in the
 * reference's set); the set grows for k rounds by the valid in-/out-neighbours of its valid members, then every
 * valid item outside it is invalidated on the device graph (read the bitmap back with mg_invalid_readback).
 * MG_EINVAL when an id is >= mg_graph_size (the reference would read out of bounds).  After a call that
 * invalidated something, mg_prune has to run again before mg_dls / mg_find_cycles (their navigation cache
 * describes the graph before the call). */
typedef struct {
  uint64_t n_seed_ids;    /* ids passed in */
  uint64_t n_added;       /* members found by the expansion (not counting the seeds) */
  uint64_t levels;        /* rounds executed (<= k; stops once a batch of rounds adds nothing) */
  uint64_t n_invalidated; /* valid items outside the set, now invalid */
  uint64_t n_kept;        /* valid items afterwards */
} mg_region_stats;
int mg_keep_regions(mg_graph *g, const uint64_t *ids, uint64_t n_ids, uint64_t k, mg_region_stats *stats);

/* The valid part of the graph in compact form — the input of the reference's SCC split
 * (find_strongly_connected_components src/spacer_ordering.cpp:54-76 walks `for node < size: IsValidEdge` and
 * calls OutgoingEdges per node; divide_graph_into_subgraphs :141-175 does so again per component).
 * ids[0..n_nodes): valid item ids, ascending.  Row i = nbr[off[i]..off[i+1]): positions in `ids` of the valid
 * out-neighbours of ids[i], in OutgoingEdges order (descending id, sdbg.h:296-320).  Arrays are library
 * allocated; release with mg_subgraph_free.  Meant for the graph after mg_keep_regions (host arrays hold
 * 8 + 8 + 4*degree bytes per valid node). */
typedef struct {
  uint64_t n_nodes, n_edges;
  uint64_t *ids;
  uint64_t *off; /* n_nodes + 1 */
  uint32_t *nbr; /* n_edges */
} mg_subgraph;
int mg_valid_subgraph(mg_graph *g, mg_subgraph *out);
void mg_subgraph_free(mg_subgraph *sg);

/* ---- read -> node-id mapping --------------------------------------------------
 * Replaces get_reads / get_read_from_sequence, reference src/reads.cpp:57-130 (declared include/reads.h:38-75),
 * and underneath SDBG::IndexBinarySearch, libs/megahit/src/sdbg/sdbg.h:141-207.
 *
 * The lookup needs three things the search path does not: tip flags with a rank, the tip labels and the bucket
 * table (prefix_look_up_, sdbg.h:36,44-46).  mg_graph_create_from_files attaches them itself; for a graph made
 * by mg_graph_create pass them here.  tip_labels: words_per_tip_label uint32 words per tip AS STORED IN THE GRAPH
 * FILE (sdbg_writer.cpp:45-52), tips in item order.  bucket_first/last[b]: first and last item of bucket b;
 * (0, 0) for a bucket absent from the file and (first, first - 1) for an empty one, as sdbg.h:36-46 leaves them. */
typedef struct {
  int32_t memory; /* MG_MEM_HOST or MG_MEM_DEVICE, for tip and tip_labels; the bucket table is always host memory */
  uint32_t words_per_tip_label;
  const uint64_t *tip; /* (n+63)/64 words, as mg_graph_desc.tip */
  const uint32_t *tip_labels;
  uint64_t n_tips;
  const int64_t *bucket_first, *bucket_last;
  uint64_t n_buckets;
} mg_kmer_index_desc;
int mg_kmer_index_create(mg_graph *g, const mg_kmer_index_desc *d);
int mg_graph_has_kmer_index(const mg_graph *g); /* 1 / 0 */

/* One batch of reads: seqs = the sequences back to back (ASCII, no separators), read r = seqs[seq_off[r] ..
 * seq_off[r+1]).  reverse_complement != 0 applies reverse_pair_ends_sequence (src/reads.cpp:20-31) to every
 * read first, as get_reads does for the second FASTQ file (:119-121).  cycle_ids = nodes_of_cycles (:98-103).
 * Result: read r maps to ids[off[r] .. off[r+1]) — len - k + 1 node ids, UINT64_MAX (kNullID) where the k-mer is
 * not in the graph; an empty range = the reference drops the read (len <= 2k, or neither its first nor its last
 * k-mer is a cycle node, :64-77).  get_reads' output is the non-empty ranges in read order, file 1 then file 2. */
typedef struct {
  uint64_t n_reads, n_kept, n_ids;
  uint64_t *off; /* n_reads + 1 */
  uint64_t *ids; /* n_ids */
} mg_reads;
int mg_map_reads(mg_graph *g, const char *seqs, const uint64_t *seq_off, uint64_t n_reads, int reverse_complement,
                 const uint64_t *cycle_ids, uint64_t n_cycle_ids, mg_reads *out);
void mg_reads_free(mg_reads *r);

/* Device timers: CUDA-event time (ms) and launch count accumulated per kernel family since the
 * last mg_timers_reset().  names[i] are static strings.  Returns the number of entries written
 * (<= cap). */
int mg_timers(mg_graph *g, const char **names, double *ms, uint64_t *launches, uint64_t *units, int cap);
void mg_timers_reset(mg_graph *g);
/* Whole-step device timing: mg_step_begin records a CUDA event on the graph's stream, mg_step_end
 * records a second one, waits for it and returns the elapsed milliseconds between the two (kernel
 * time plus the gaps in which the stream waited for the host). */
int mg_step_begin(mg_graph *g);
int mg_step_end(mg_graph *g, double *elapsed_ms);
/* kernels launched by this library in this process so far */
uint64_t mg_launch_count(void);

void mg_free(void *p);
const char *mg_last_error(void);
const char *mg_version(void);

#ifdef __cplusplus
}
#endif
#endif
