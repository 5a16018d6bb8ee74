/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement (plain C) of the reference's bounded-length multicycle search
 * (reference src/cycle_finder.cpp) and of the MEGAHIT succinct de Bruijn graph navigation it
 * calls (reference libs/megahit/src/sdbg/sdbg.h, libs/megahit/src/kmlib/kmrns.h).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library,
 * and only as the CHECKER.  The product (mcaat_b200/, libmcaat_gpu.so) never links or calls it.
 *
 * Parity status: PINNED against the reference itself — the reference ships no test or golden
 * vector for this path (its tests/ cover only reads and set cover), so the pins are dumps of the
 * UNMODIFIED reference CycleFinder compiled by oracle/Makefile into oracle/_ref/ and run at
 * threads=1 on seeded synthetic graphs; they are committed under tests/golden/ together with the
 * script that made them (tools/make_golden.py).  The "next" stages restated further down (read mapping,
 * region extraction) are pinned the same way: oracle/_ref/ref_reads and oracle/_ref/ref_regions are the
 * reference's src/reads.cpp and src/spacer_ordering.cpp compiled unmodified.
 */
#ifndef MCAAT_ORACLE_H
#define MCAAT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_graph orc_graph;

/* CSR result of one whole run (reference FindApproximateCRISPRArrays, cycle_finder.cpp:433-492). */
typedef struct {
  uint64_t n_tips_before;  /* "gathered tips" (:437) */
  uint64_t n_mult1;        /* "Pre-filter: invalidated" (:381) */
  uint64_t n_tips_after;   /* "After pruning, tips" (:452) */
  uint64_t n_valid;        /* "valid edges" (:452) */
  uint64_t n_candidates;   /* nodes passing the :404-408 predicate (= DLS calls) */
  uint64_t n_buckets;
  int64_t *bucket_key;     /* descending */
  uint64_t *bucket_off;    /* n_buckets+1 offsets into start_ids */
  uint64_t *start_ids;     /* ChunkStartNodes order at threads=1 (ascending id per bucket) */
  uint64_t n_searched;     /* entries of `results`, in insertion order */
  uint64_t *searched_start;
  uint64_t *cyc_off;       /* n_searched+1 offsets into the cycle list */
  uint64_t n_cycles;
  uint64_t *node_off;      /* n_cycles+1 offsets into nodes */
  uint64_t *nodes;
  uint64_t n_bound_hits;   /* searches whose counter reached the 500-cycle cluster bound */
  uint64_t n_cap_hits;     /* searches that ran into the 10 000 000-step cap */
  uint64_t dls_expansions; /* popped-and-expanded nodes in DepthLevelSearch */
  uint64_t enum_expansions;/* _GetOutgoings/_GetIncomings evaluations in FindCycle */
} orc_result;

/* Load the on-disk MEGAHIT SDBG (<prefix>.sdbg_info, <prefix>.sdbg.<i>);
 * restates sdbg_raw_content.cpp:18-96, sdbg_meta.cpp:10-75, sdbg.h:26-61. NULL on error. */
orc_graph *orc_load(const char *prefix);

/* Build from unpacked arrays (one byte per item for w/last/tip, u16 multiplicities) and the
 * character-class boundaries f[6] of sdbg.h:36-50.  Initial invalid = tip | (w==0), sdbg.h:34,56-60. */
orc_graph *orc_from_arrays(uint64_t n, uint32_t k, const uint8_t *w, const uint8_t *last,
                           const uint8_t *tip, const uint16_t *mult, const int64_t f[6]);
void orc_free(orc_graph *g);

uint64_t orc_size(const orc_graph *g);
uint32_t orc_k(const orc_graph *g);
void orc_get_f(const orc_graph *g, int64_t f[6]);
/* copy out raw arrays (any pointer may be NULL) */
void orc_get_raw(const orc_graph *g, uint8_t *w, uint8_t *last, uint8_t *tip, uint16_t *mult);
void orc_get_invalid(const orc_graph *g, uint8_t *invalid);
void orc_set_invalid(orc_graph *g, const uint8_t *invalid);

/* kmrns.h:287-345 semantics */
int64_t orc_rank_w(const orc_graph *g, int c, int64_t pos);
int64_t orc_select_w(const orc_graph *g, int c, int64_t k);
int64_t orc_rank_last(const orc_graph *g, int64_t pos);
int64_t orc_select_last(const orc_graph *g, int64_t k);
/* sdbg.h:107-121 */
uint64_t orc_forward(const orc_graph *g, uint64_t e);
uint64_t orc_backward(const orc_graph *g, uint64_t e);
/* sdbg.h:245-320: degree or -1; ids written to out[0..deg) */
int orc_outgoing(const orc_graph *g, uint64_t e, uint64_t out[4]);
int orc_incoming(const orc_graph *g, uint64_t e, uint64_t in[4]);
uint16_t orc_mult(const orc_graph *g, uint64_t e);

/* phases */
uint64_t orc_collect_tips(const orc_graph *g, uint64_t **tips);       /* cycle_finder.cpp:346-357 */
uint64_t orc_invalidate_mult1(orc_graph *g);                          /* :372-382 */
void orc_peel(orc_graph *g, const uint64_t *tips, uint64_t n_tips);   /* :359-371, :440-442 */
/* :248-343; *expansions += nodes popped and expanded */
int orc_dls(const orc_graph *g, uint64_t start, int limit, uint64_t *expansions);
/* :231-243,140-225 against visited[] (byte per node, updated on return like :217-222).
 * Returns number of cycles, paths appended through the callback-free buffers below. */
typedef struct {
  uint64_t n_cycles;
  uint64_t *node_off; /* n_cycles+1 */
  uint64_t *nodes;
  int bound_hit, cap_hit;
  uint64_t expansions;
  uint64_t steps;
} orc_cycles;
void orc_find_cycle_util(const orc_graph *g, uint64_t start, int min_len, int max_len,
                         uint8_t *visited, orc_cycles *out);
void orc_cycles_free(orc_cycles *c);

/* whole driver; mutates g's invalid bits like the reference mutates the host SDBG */
int orc_run(orc_graph *g, uint64_t threshold_multiplicity, int min_len, int max_len, orc_result *res);
void orc_result_free(orc_result *res);

/* SURVEY 8(f) N3 — read -> node-id mapping.  Needs a graph made by orc_load (tip labels and the bucket table
 * are not part of orc_from_arrays).  Pinned against the compiled reference functions: oracle/_ref/ref_reads,
 * tests/golden/<case>.reads.npz.
 * orc_index_binary_search: SDBG::IndexBinarySearch (sdbg.h:141-207), seq[0..k) in 1..4; ~0 = kNullID.
 * orc_get_read: get_read_from_sequence (src/reads.cpp:57-89); orc_reverse_complement: :20-31. */
uint64_t orc_index_binary_search(const orc_graph *g, const uint8_t *seq);
uint64_t orc_get_read(const orc_graph *g, const uint8_t *member, const char *seq, uint64_t len, uint64_t *out);
void orc_reverse_complement(char *s, uint64_t len);
void orc_kmer_index_sizes(const orc_graph *g, uint64_t *n_tips, uint32_t *wptl, uint64_t *n_buckets);
void orc_kmer_index_get(const orc_graph *g, uint32_t *labels, int64_t *first, int64_t *second);

/* SURVEY 8(f) N4 — keep_crispr_regions_extended_by_k (reference src/spacer_ordering.cpp:78-139): grow the set of
 * cycle node ids by k rounds of "add every in- and out-neighbour of each VALID member", then invalidate every
 * valid item outside the set.  Returns the number of items invalidated, or UINT64_MAX when an id is >= size
 * (the reference would read out of bounds).  *n_kept (may be NULL) = valid items afterwards.
 * Pinned against the compiled reference function: oracle/_ref/ref_regions, tests/golden/<case>.regions.npz. */
uint64_t orc_keep_regions(orc_graph *g, const uint64_t *ids, uint64_t n_ids, uint64_t k, uint64_t *n_kept);

/* libstdc++ std::unordered_set<uint64_t> iteration-order model used by FindCycle:
 * inserts ids[0..n) (n <= 8) into an empty set and writes the iteration order; returns the size. */
int orc_uset_order(const uint64_t *ids, int n, uint64_t *order);

#ifdef __cplusplus
}
#endif
#endif
