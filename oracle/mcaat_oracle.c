/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See mcaat_oracle.h for scope and parity status.
 *
 * Plain-C restatement of the reference hot path.  Data structures are deliberately the simplest
 * ones that give the same ANSWERS as the reference (sorted occurrence lists instead of the
 * sampled rank/select of kmrns.h; one byte per flag), so that the oracle is independent of the
 * device layout it checks.  Every function cites the reference lines it follows, relative to
 * /root/reference.
 */
#define _GNU_SOURCE
#include "mcaat_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct orc_graph {
  uint64_t n;
  uint32_t k;
  uint8_t *w, *last, *tip, *lt, *invalid; /* one byte per item */
  uint16_t *mult;
  int64_t f[6], rank_f[6];
  uint64_t *occ_w[5]; /* sorted positions of nibble c (1..4) in W */
  uint64_t cnt_w[5];
  uint64_t *occ_last;
  uint64_t cnt_last;
  /* k-mer lookup (sdbg.h:141-207), only when loaded from files */
  uint32_t wptl;          /* words per tip label */
  uint32_t *tip_labels;   /* file words, wptl per tip, in item order */
  uint64_t *tip_index;    /* per item: number of tips before it (rs_is_tip_.rank(x) - 1 for a tip x) */
  uint64_t n_buckets;
  int64_t *pl_first, *pl_second; /* prefix_look_up_ (sdbg.h:36,44-46); absent buckets stay (0,0) */
};

/* ---------------------------------------------------------------- small containers */
typedef struct { uint64_t *a; uint64_t n, cap; } vec64;
static void v_push(vec64 *v, uint64_t x) {
  if (v->n == v->cap) {
    v->cap = v->cap ? v->cap * 2 : 64;
    v->a = (uint64_t *)realloc(v->a, v->cap * 8);
  }
  v->a[v->n++] = x;
}

/* open-addressing u64 -> int map (also used as a set); key ~0 is the empty marker */
typedef struct { uint64_t *key; int *val; uint64_t cap, n; } hmap;
static uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
static void hm_init(hmap *m, uint64_t cap) {
  m->cap = cap; m->n = 0;
  m->key = (uint64_t *)malloc(cap * 8);
  m->val = (int *)malloc(cap * sizeof(int));
  memset(m->key, 0xff, cap * 8);
}
static void hm_free(hmap *m) { free(m->key); free(m->val); }
static int *hm_get(hmap *m, uint64_t k, int insert, int defval);
static void hm_grow(hmap *m) {
  hmap o = *m;
  hm_init(m, o.cap * 2);
  for (uint64_t i = 0; i < o.cap; ++i)
    if (o.key[i] != ~0ull) *hm_get(m, o.key[i], 1, 0) = o.val[i];
  hm_free(&o);
}
static int *hm_get(hmap *m, uint64_t k, int insert, int defval) {
  if (insert && (m->n + 1) * 2 > m->cap) hm_grow(m);
  uint64_t i = mix64(k) & (m->cap - 1);
  while (m->key[i] != ~0ull) {
    if (m->key[i] == k) return &m->val[i];
    i = (i + 1) & (m->cap - 1);
  }
  if (!insert) return NULL;
  m->key[i] = k; m->val[i] = defval; m->n++;
  return &m->val[i];
}

/* ---------------------------------------------------------------- graph construction */
static void build_index(orc_graph *g) {
  uint64_t n = g->n;
  g->lt = (uint8_t *)malloc(n ? n : 1);
  for (int c = 0; c < 5; ++c) { g->cnt_w[c] = 0; g->occ_w[c] = NULL; }
  g->cnt_last = 0;
  for (uint64_t i = 0; i < n; ++i) {
    if (g->w[i] >= 1 && g->w[i] <= 4) g->cnt_w[g->w[i]]++;
    g->cnt_last += g->last[i];
    g->lt[i] = g->last[i] | g->tip[i]; /* sdbg.h:71-75 */
  }
  for (int c = 1; c < 5; ++c) g->occ_w[c] = (uint64_t *)malloc((g->cnt_w[c] + 1) * 8);
  g->occ_last = (uint64_t *)malloc((g->cnt_last + 1) * 8);
  uint64_t pw[5] = {0, 0, 0, 0, 0}, pl = 0;
  for (uint64_t i = 0; i < n; ++i) {
    if (g->w[i] >= 1 && g->w[i] <= 4) g->occ_w[g->w[i]][pw[g->w[i]]++] = i;
    if (g->last[i]) g->occ_last[pl++] = i;
  }
  /* sdbg.h:52-54  rank_f_[i] = rs_last_.rank(f_[i]-1), i = 1..5; rank_f_[0] stays 0 */
  g->rank_f[0] = 0;
  for (int i = 1; i < 6; ++i) g->rank_f[i] = orc_rank_last(g, g->f[i] - 1);
  /* sdbg.h:34 invalid_ := tip bits ; :56-60 W==0 => invalid */
  for (uint64_t i = 0; i < n; ++i) g->invalid[i] = g->tip[i] | (g->w[i] == 0);
}

static orc_graph *alloc_graph(uint64_t n) {
  orc_graph *g = (orc_graph *)calloc(1, sizeof(orc_graph));
  uint64_t m = n ? n : 1;
  g->n = n;
  g->w = (uint8_t *)calloc(m, 1);
  g->last = (uint8_t *)calloc(m, 1);
  g->tip = (uint8_t *)calloc(m, 1);
  g->invalid = (uint8_t *)calloc(m, 1);
  g->mult = (uint16_t *)calloc(m, 2);
  return g;
}

orc_graph *orc_from_arrays(uint64_t n, uint32_t k, const uint8_t *w, const uint8_t *last,
                           const uint8_t *tip, const uint16_t *mult, const int64_t f[6]) {
  orc_graph *g = alloc_graph(n);
  g->k = k;
  memcpy(g->w, w, n);
  memcpy(g->last, last, n);
  memcpy(g->tip, tip, n);
  memcpy(g->mult, mult, n * 2);
  memcpy(g->f, f, sizeof(g->f));
  build_index(g);
  return g;
}

typedef struct { uint64_t bucket_id, file_id, off, num_items, num_tips, num_large, acc_items, acc_tips; } bucket_rec;
static int cmp_bucket_id(const void *a, const void *b) {
  const bucket_rec *x = (const bucket_rec *)a, *y = (const bucket_rec *)b;
  return x->bucket_id < y->bucket_id ? -1 : x->bucket_id > y->bucket_id;
}
static int cmp_file_off(const void *a, const void *b) {
  const bucket_rec *x = (const bucket_rec *)a, *y = (const bucket_rec *)b;
  if (x->file_id != y->file_id) return x->file_id < y->file_id ? -1 : 1;
  return x->off < y->off ? -1 : x->off > y->off;
}

/* sdbg_meta.cpp:61-75 (Deserialize), :10-47 (FromBucketRecord), sdbg_raw_content.cpp:18-96,
 * sdbg_item.h:14-24 (16-bit item: w:4 last:1 tip:1 pad:2 mul:8), sdbg.h:36-50 (f_). */
orc_graph *orc_load(const char *prefix) {
  char path[4096];
  snprintf(path, sizeof path, "%s.sdbg_info", prefix);
  FILE *fi = fopen(path, "r");
  if (!fi) return NULL;
  char key[64];
  unsigned long long k = 0, wptl = 0, nb = 0, nf = 0;
  if (fscanf(fi, "%63s %llu", key, &k) != 2 || fscanf(fi, "%63s %llu", key, &wptl) != 2 ||
      fscanf(fi, "%63s %llu", key, &nb) != 2 || fscanf(fi, "%63s %llu", key, &nf) != 2) {
    fclose(fi);
    return NULL;
  }
  bucket_rec *rec = (bucket_rec *)calloc(nb ? nb : 1, sizeof(bucket_rec));
  for (uint64_t i = 0; i < nb; ++i) {
    unsigned long long a, b, c, d, e, f;
    if (fscanf(fi, "%llu %llu %llu %llu %llu %llu", &a, &b, &c, &d, &e, &f) != 6) { fclose(fi); free(rec); return NULL; }
    rec[i].bucket_id = a; rec[i].file_id = b; rec[i].off = c;
    rec[i].num_items = d; rec[i].num_tips = e; rec[i].num_large = f;
  }
  fclose(fi);
  /* accumulate in bucket_id order, skipping records whose ids are the null id (SIZE_MAX) */
  qsort(rec, nb, sizeof(bucket_rec), cmp_bucket_id);
  uint64_t n = 0, n_tips_total = 0;
  for (uint64_t i = 0; i < nb; ++i) {
    if (rec[i].file_id == ~0ull || rec[i].bucket_id == ~0ull) continue;
    rec[i].acc_items = n;
    rec[i].acc_tips = n_tips_total;
    n += rec[i].num_items;
    n_tips_total += rec[i].num_tips;
  }
  qsort(rec, nb, sizeof(bucket_rec), cmp_file_off);
  orc_graph *g = alloc_graph(n);
  g->k = (uint32_t)k;
  memset(g->f, 0, sizeof g->f);
  g->f[0] = -1;
  g->wptl = (uint32_t)wptl;
  g->n_buckets = nb;
  g->pl_first = (int64_t *)calloc(nb ? nb : 1, 8);
  g->pl_second = (int64_t *)calloc(nb ? nb : 1, 8);
  g->tip_index = (uint64_t *)calloc(n ? n : 1, 8);
  g->tip_labels = (uint32_t *)calloc(n_tips_total * wptl + 1, 4);
  FILE *fd = NULL;
  uint64_t cur_file = ~0ull;
  int ok = 1;
  for (uint64_t i = 0; i < nb && ok; ++i) {
    if (rec[i].bucket_id == ~0ull) break; /* sdbg.h:40, sdbg_raw_content.cpp:42-44 */
    g->f[rec[i].bucket_id / (nb / 4) + 2] += (int64_t)rec[i].num_items;
    g->pl_first[rec[i].bucket_id] = (int64_t)rec[i].acc_items;                                  /* sdbg.h:44 */
    g->pl_second[rec[i].bucket_id] = (int64_t)rec[i].acc_items + (int64_t)rec[i].num_items - 1; /* :45-46 */
    if (rec[i].file_id != cur_file) {
      if (fd) fclose(fd);
      snprintf(path, sizeof path, "%s.sdbg.%llu", prefix, (unsigned long long)rec[i].file_id);
      fd = fopen(path, "rb");
      if (!fd) { ok = 0; break; }
      cur_file = rec[i].file_id;
    }
    uint64_t lab_n = rec[i].acc_tips * wptl; /* sdbg_raw_content.cpp:58-60 */
    for (uint64_t j = 0; j < rec[i].num_items; ++j) {
      uint64_t idx = rec[i].acc_items + j;
      uint8_t it[2];
      if (fread(it, 1, 2, fd) != 2) { ok = 0; break; }
      g->w[idx] = it[0] & 15;
      g->last[idx] = (it[0] >> 4) & 1;
      g->tip[idx] = (it[0] >> 5) & 1;
      uint16_t mul = it[1];
      if (mul == 255) { if (fread(&mul, 2, 1, fd) != 1) { ok = 0; break; } }
      g->mult[idx] = mul;
      if (g->tip[idx] && wptl) { /* sdbg_raw_content.cpp:82-88; kept as stored (see tip_char) */
        if (lab_n + wptl > n_tips_total * wptl) { ok = 0; break; }
        if (fread(g->tip_labels + lab_n, 4, wptl, fd) != wptl) { ok = 0; break; }
        lab_n += wptl;
      }
    }
  }
  if (fd) fclose(fd);
  free(rec);
  if (!ok) { orc_free(g); return NULL; }
  for (int i = 2; i < 6; ++i) g->f[i] += g->f[i - 1];
  build_index(g);
  /* labels sit at accumulate_tip_count + position inside the bucket = number of tips before the item */
  {
    uint64_t n_tips = 0;
    for (uint64_t i = 0; i < n; ++i) { g->tip_index[i] = n_tips; n_tips += g->tip[i]; }
  }
  return g;
}

void orc_free(orc_graph *g) {
  if (!g) return;
  free(g->w); free(g->last); free(g->tip); free(g->lt); free(g->invalid); free(g->mult);
  for (int c = 1; c < 5; ++c) free(g->occ_w[c]);
  free(g->occ_last);
  free(g->tip_labels); free(g->tip_index); free(g->pl_first); free(g->pl_second);
  free(g);
}

uint64_t orc_size(const orc_graph *g) { return g->n; }
uint32_t orc_k(const orc_graph *g) { return g->k; }
void orc_get_f(const orc_graph *g, int64_t f[6]) { memcpy(f, g->f, sizeof g->f); }
void orc_get_raw(const orc_graph *g, uint8_t *w, uint8_t *last, uint8_t *tip, uint16_t *mult) {
  if (w) memcpy(w, g->w, g->n);
  if (last) memcpy(last, g->last, g->n);
  if (tip) memcpy(tip, g->tip, g->n);
  if (mult) memcpy(mult, g->mult, g->n * 2);
}
void orc_get_invalid(const orc_graph *g, uint8_t *invalid) { memcpy(invalid, g->invalid, g->n); }
void orc_set_invalid(orc_graph *g, const uint8_t *invalid) { memcpy(g->invalid, invalid, g->n); }
uint16_t orc_mult(const orc_graph *g, uint64_t e) { return g->mult[e]; }

/* ---------------------------------------------------------------- rank / select (kmrns.h) */
static uint64_t upper_bound(const uint64_t *a, uint64_t n, uint64_t x) { /* #elements <= x */
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    uint64_t mid = (lo + hi) / 2;
    if (a[mid] <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
/* kmrns.h:287-311: pos >= size -> -1; pos == size-1 -> total; pos == -1 -> 0 (falls through to
 * OccValue(c,0)); else #c in [0..pos]. */
static int64_t rank_generic(const uint64_t *occ, uint64_t cnt, uint64_t size, int64_t pos) {
  if (pos >= (int64_t)size) return -1;
  if (pos == (int64_t)size - 1) return (int64_t)cnt;
  if (pos < 0) return 0;
  return (int64_t)upper_bound(occ, cnt, (uint64_t)pos);
}
/* kmrns.h:313-345: k > count -> -1; k == count -> size; else position of the k-th (0-based) c. */
static int64_t select_generic(const uint64_t *occ, uint64_t cnt, uint64_t size, int64_t k) {
  if (k > (int64_t)cnt) return -1;
  if (k == (int64_t)cnt) return (int64_t)size;
  if (k < 0) return -1; /* reference behaviour undefined; never reached on a well-formed graph */
  return (int64_t)occ[k];
}
int64_t orc_rank_w(const orc_graph *g, int c, int64_t pos) {
  if (c < 1 || c > 4) return -1; /* only chars 1..4 are ever ranked (sdbg.h:108-112,118-120) */
  return rank_generic(g->occ_w[c], g->cnt_w[c], g->n, pos);
}
int64_t orc_select_w(const orc_graph *g, int c, int64_t k) {
  if (c < 1 || c > 4) return -1;
  return select_generic(g->occ_w[c], g->cnt_w[c], g->n, k);
}
int64_t orc_rank_last(const orc_graph *g, int64_t pos) { return rank_generic(g->occ_last, g->cnt_last, g->n, pos); }
int64_t orc_select_last(const orc_graph *g, int64_t k) { return select_generic(g->occ_last, g->cnt_last, g->n, k); }

/* ---------------------------------------------------------------- navigation (sdbg.h) */
/* sdbg.h:107-114 */
uint64_t orc_forward(const orc_graph *g, uint64_t e) {
  int a = g->w[e];
  if (a > 4) a -= 4;
  int64_t cnt = orc_rank_w(g, a, (int64_t)e);
  return (uint64_t)orc_select_last(g, g->rank_f[a] + cnt - 1);
}
/* sdbg.h:79-86 */
static int last_char_of(const orc_graph *g, uint64_t x) {
  for (int i = 1; i < 6; ++i)
    if (g->f[i] > (int64_t)x) return i - 1;
  return 6;
}
/* sdbg.h:116-121 */
uint64_t orc_backward(const orc_graph *g, uint64_t e) {
  int a = last_char_of(g, e);
  int64_t cnt = orc_rank_last(g, (int64_t)e - 1) - g->rank_f[a];
  return (uint64_t)orc_select_w(g, a, cnt);
}
/* sdbg.h:296-320 with kFlagWriteOut: valid edges of the target node's group, descending id */
int orc_outgoing(const orc_graph *g, uint64_t e, uint64_t out[4]) {
  if (g->invalid[e]) return -1;
  int deg = 0;
  uint64_t next = orc_forward(g, e);
  if (next >= g->n) return 0; /* malformed graph guard; the reference would read out of bounds */
  do {
    if (!g->invalid[next]) { if (out && deg < 4) out[deg] = next; ++deg; }
    --next;
  } while (next != ~0ull && !g->lt[next]);
  return deg;
}
/* sdbg.h:245-287 with kFlagWriteOut: ascending id */
int orc_incoming(const orc_graph *g, uint64_t e, uint64_t in[4]) {
  if (g->invalid[e]) return -1;
  uint64_t first = orc_backward(g, e);
  if (first >= g->n) return 0; /* malformed graph guard */
  int c = g->w[first];
  unsigned ones = g->lt[first];
  int deg = !g->invalid[first];
  if (deg > 0 && in) in[0] = first;
  for (uint64_t y = first + 1; ones < 5 && y < g->n; ++y) {
    ones += g->lt[y];
    int cur = g->w[y];
    if (cur == c) break;
    if (cur == c + 4 && !g->invalid[y]) { if (in && deg < 4) in[deg] = y; ++deg; }
  }
  return deg;
}

/* ---------------------------------------------------------------- pruning */
/* cycle_finder.cpp:346-357 — consumed as a set; returned ascending */
uint64_t orc_collect_tips(const orc_graph *g, uint64_t **tips) {
  vec64 v = {0, 0, 0};
  for (uint64_t i = 0; i < g->n; ++i)
    if (!g->invalid[i] && orc_outgoing(g, i, NULL) == 0) v_push(&v, i);
  if (tips) *tips = v.a; else free(v.a);
  return v.n;
}
/* cycle_finder.cpp:372-382 */
uint64_t orc_invalidate_mult1(orc_graph *g) {
  uint64_t c = 0;
  for (uint64_t i = 0; i < g->n; ++i)
    if (g->mult[i] <= 1) { g->invalid[i] = 1; ++c; }
  return c;
}
/* cycle_finder.cpp:359-371 (RecursiveReduction) driven as at :440-442.  The recursion is
 * restated with an explicit stack visiting parents in the same order (ascending ids are what
 * _GetIncomings inserts; the resulting closure does not depend on the order). */
void orc_peel(orc_graph *g, const uint64_t *tips, uint64_t n_tips) {
  vec64 st = {0, 0, 0};
  for (uint64_t t = 0; t < n_tips; ++t) {
    v_push(&st, tips[t]);
    while (st.n) {
      uint64_t v = st.a[--st.n];
      if (orc_outgoing(g, v, NULL) > 0) continue; /* -1 (invalid) passes, as in the reference */
      uint64_t in[4];
      int deg = orc_incoming(g, v, in); /* _GetIncomings(node,set): empty when invalid or indeg 0 (:113-116) */
      g->invalid[v] = 1;
      for (int i = deg - 1; i >= 0; --i) v_push(&st, in[i]); /* validity re-checked at pop via outgoing()==-1 → incoming()==-1 */
    }
  }
  free(st.a);
}

/* ---------------------------------------------------------------- read -> node ids (N3) */
/* CharAtTipLabel (sdbg.h:128-133) on the words as stored in the file: the loader reverses the 2-bit groups of
 * each word (sdbg_raw_content.cpp:84-87) and CompactVector<2, uint32_t> is little-endian (kmcompactvector.h:14-15,
 * 52-60), so char `off` is bits 30-2*(off%16) of file word off/16; values are 1-based. */
static int tip_char(const orc_graph *g, uint64_t tip_item, unsigned off) {
  const uint32_t *lab = g->tip_labels + g->tip_index[tip_item] * g->wptl; /* sdbg.h:124-127 */
  return (int)((lab[off / 16] >> (30 - 2 * (off % 16))) & 3u) + 1;
}
/* sdbg.h:77 GetLastIndex = rs_last_.succ(x): first item >= x with the last flag */
static uint64_t last_index(const orc_graph *g, uint64_t x) {
  while (x < g->n && !g->last[x]) ++x;
  return x < g->n ? x : ~0ull;
}
/* sdbg.h:141-207, restated line by line — including that W is compared RAW (:187-195: labels 5..8 are not
 * folded to 1..4 here, unlike GetLabel :225-228), which decides where the binary search goes. */
uint64_t orc_index_binary_search(const orc_graph *g, const uint8_t *seq) {
  if (!g->pl_first) return ~0ull;
  const int k = (int)g->k;
  uint64_t prefix = 0;
  for (uint64_t i = 0; (1u << (i * 2)) < g->n_buckets; ++i) prefix = prefix * 4 + seq[k - 1 - i] - 1; /* :142-146 */
  int64_t l = g->pl_first[prefix], r = g->pl_second[prefix];
  while (l <= r) {
    int cmp = 0;
    uint64_t mid = (uint64_t)((l + r) / 2), y = mid;
    for (int i = k - 1; i >= 0; --i) {
      if (g->tip[y]) { /* :156-184 */
        for (int j = 0; j < i; ++j) {
          int c = tip_char(g, y, (unsigned)j);
          if (c < seq[i - j]) { cmp = -1; break; }
          else if (c > seq[i - j]) { cmp = 1; break; }
        }
        if (cmp == 0) {
          if (g->tip[mid]) cmp = -1;
          else {
            int c = tip_char(g, y, (unsigned)i);
            if (c < seq[0]) cmp = -1;
            else if (c > seq[0]) cmp = 1;
          }
        }
        break;
      }
      y = orc_backward(g, y); /* :186 */
      int c = g->w[y];        /* :187 raw W */
      if (c < seq[i]) { cmp = -1; break; }
      else if (c > seq[i]) { cmp = 1; break; }
    }
    if (cmp == 0) return last_index(g, mid); /* :198-199 */
    else if (cmp > 0) r = (int64_t)mid - 1;
    else l = (int64_t)mid + 1;
  }
  return ~0ull; /* kNullID */
}
/* raw pieces of the k-mer index, for the host build of the device code (tests/emu) */
void orc_kmer_index_sizes(const orc_graph *g, uint64_t *n_tips, uint32_t *wptl, uint64_t *n_buckets) {
  uint64_t t = 0;
  for (uint64_t i = 0; i < g->n; ++i) t += g->tip[i];
  *n_tips = g->pl_first ? t : 0; *wptl = g->wptl; *n_buckets = g->pl_first ? g->n_buckets : 0;
}
void orc_kmer_index_get(const orc_graph *g, uint32_t *labels, int64_t *first, int64_t *second) {
  uint64_t t, nb; uint32_t w;
  orc_kmer_index_sizes(g, &t, &w, &nb);
  memcpy(labels, g->tip_labels, t * w * 4);
  memcpy(first, g->pl_first, nb * 8);
  memcpy(second, g->pl_second, nb * 8);
}
/* reads.cpp:33-55 k_mer_to_node_id: A,C,G -> 1,2,3, anything else -> 4 */
static uint64_t kmer_node(const orc_graph *g, const char *kmer) {
  uint8_t seq[256];
  for (uint32_t i = 0; i < g->k; ++i) seq[i] = kmer[i] == 'A' ? 1 : kmer[i] == 'C' ? 2 : kmer[i] == 'G' ? 3 : 4;
  return orc_index_binary_search(g, seq);
}
/* reads.cpp:20-31 reverse_pair_ends_sequence, in place */
void orc_reverse_complement(char *s, uint64_t len) {
  for (uint64_t i = 0; i < len / 2; ++i) { char t = s[i]; s[i] = s[len - 1 - i]; s[len - 1 - i] = t; }
  for (uint64_t i = 0; i < len; ++i)
    s[i] = s[i] == 'A' ? 'T' : s[i] == 'T' ? 'A' : s[i] == 'C' ? 'G' : s[i] == 'G' ? 'C' : s[i];
}
/* reads.cpp:57-89 get_read_from_sequence; member = nodes_of_cycles as a byte per item.  Returns the number of
 * ids written to out (capacity >= len), 0 = the read is dropped. */
uint64_t orc_get_read(const orc_graph *g, const uint8_t *member, const char *seq, uint64_t len, uint64_t *out) {
  const uint64_t K = g->k;
  if (len <= 2 * K) return 0;                                   /* :62-64 */
  uint64_t a = kmer_node(g, seq), b = kmer_node(g, seq + len - K); /* :66-70 */
  int in_a = a < g->n && member[a], in_b = b < g->n && member[b];
  if (!in_a && !in_b) return 0;                                 /* :72-75 */
  uint64_t c = 0;
  out[c++] = a;
  for (uint64_t i = 1; i < len - K; ++i) out[c++] = kmer_node(g, seq + i); /* :78-83 */
  out[c++] = b;
  return c;
}

/* ---------------------------------------------------------------- region extraction (N4) */
/* spacer_ordering.cpp:78-139.  `in_set` is the unordered_set cycle_edges as a byte per item; `members` lists
 * it.  Every round the reference scans ALL members (:93-125), collects neighbours into new_edges and
 * inserts them afterwards (:127-129) — restated literally, members found in a round are scanned from the
 * next round on.  Neighbours come from IncomingEdges/OutgoingEdges (:100-122), i.e. valid items only. */
uint64_t orc_keep_regions(orc_graph *g, const uint64_t *ids, uint64_t n_ids, uint64_t k, uint64_t *n_kept) {
  for (uint64_t i = 0; i < n_ids; ++i)
    if (ids[i] >= g->n) return ~0ull;
  uint8_t *in_set = calloc(g->n ? g->n : 1, 1);
  vec64 members = {0, 0, 0};
  for (uint64_t i = 0; i < n_ids; ++i)
    if (!in_set[ids[i]]) { in_set[ids[i]] = 1; v_push(&members, ids[i]); }
  for (uint64_t round = 0; round < k; ++round) {
    vec64 fresh = {0, 0, 0};
    for (uint64_t j = 0; j < members.n; ++j) {
      uint64_t e = members.a[j], nb[4];
      if (g->invalid[e]) continue;                                  /* :94-96 */
      int d = orc_incoming(g, e, nb);                                /* :98-109 */
      for (int i = 0; i < d && i < 4; ++i) v_push(&fresh, nb[i]);
      d = orc_outgoing(g, e, nb);                                    /* :111-122 */
      for (int i = 0; i < d && i < 4; ++i) v_push(&fresh, nb[i]);
    }
    uint64_t added = 0;
    for (uint64_t j = 0; j < fresh.n; ++j)                           /* :127-129 */
      if (!in_set[fresh.a[j]]) { in_set[fresh.a[j]] = 1; v_push(&members, fresh.a[j]); ++added; }
    free(fresh.a);
    if (!added) break; /* a round that adds nothing makes every later round a no-op */
  }
  uint64_t gone = 0, kept = 0;
  for (uint64_t e = 0; e < g->n; ++e) {                              /* :131-138 */
    if (g->invalid[e]) continue;
    if (!in_set[e]) { g->invalid[e] = 1; ++gone; } else ++kept;
  }
  if (n_kept) *n_kept = kept;
  free(in_set);
  free(members.a);
  return gone;
}

/* ---------------------------------------------------------------- DepthLevelSearch */
/* cycle_finder.cpp:248-343 with target == start (the only call site, :411). */
int orc_dls(const orc_graph *g, uint64_t start, int limit, uint64_t *expansions) {
  typedef struct { uint64_t node; int depth; } ent;
  uint64_t cap = 64, sp = 0, exp = 0;
  ent *st = (ent *)malloc(cap * sizeof(ent));
  hmap seen;
  hm_init(&seen, 256);
  int found = 0;
  st[sp].node = start; st[sp].depth = 0; ++sp;
  while (sp) {
    ent cur = st[--sp];
    uint64_t v = cur.node;
    int depth = cur.depth;
    if (g->invalid[v]) continue;                       /* :285 */
    uint64_t nb[4];
    int outdeg = orc_outgoing(g, v, nb);               /* :294-302 (three evaluations, one answer) */
    if (outdeg == 0) continue;                         /* :294 */
    ++exp;
    if (depth >= limit) continue;                      /* :312 */
    for (int i = 0; i < outdeg; ++i) {                 /* :320-334, array order = descending id */
      uint64_t nbr = nb[i];
      if (g->invalid[nbr]) continue;
      int not_visited = hm_get(&seen, nbr, 0, 0) == NULL;
      int is_start_revisit = (nbr == start && depth > 0);
      if (not_visited || is_start_revisit) {
        hm_get(&seen, nbr, 1, 1);
        if (sp == cap) { cap *= 2; st = (ent *)realloc(st, cap * sizeof(ent)); }
        st[sp].node = nbr; st[sp].depth = depth + 1; ++sp;
      }
    }
    if (v == start && depth > 1) { found = 1; break; } /* :337-339 */
  }
  free(st);
  hm_free(&seen);
  if (expansions) *expansions += exp;
  return found;
}

/* ---------------------------------------------------------------- libstdc++ unordered_set order */
/* Model of std::unordered_set<uint64_t> as libstdc++ lays it out for <= 4 (here <= 8) elements:
 * 13 buckets after the first insertion, identity hash, a new element goes to the global list head
 * when its bucket is empty, otherwise to the front of its bucket's run; erase keeps the relative
 * order (verified against the compiled reference by tests/test_oracle_golden.py). */
typedef struct { uint64_t e[8]; int n; } uset;
static void uset_insert(uset *s, uint64_t x) {
  int first_same = -1;
  for (int i = 0; i < s->n; ++i) {
    if (s->e[i] == x) return;
    if (first_same < 0 && s->e[i] % 13 == x % 13) first_same = i;
  }
  if (s->n >= 8) return;
  int at = first_same < 0 ? 0 : first_same;
  for (int i = s->n; i > at; --i) s->e[i] = s->e[i - 1];
  s->e[at] = x;
  s->n++;
}
static void uset_erase(uset *s, uint64_t x) {
  for (int i = 0; i < s->n; ++i)
    if (s->e[i] == x) {
      for (int j = i; j + 1 < s->n; ++j) s->e[j] = s->e[j + 1];
      s->n--;
      return;
    }
}
int orc_uset_order(const uint64_t *ids, int n, uint64_t *order) {
  uset s; s.n = 0;
  for (int i = 0; i < n; ++i) uset_insert(&s, ids[i]);
  for (int i = 0; i < s.n; ++i) order[i] = s.e[i];
  return s.n;
}

/* ---------------------------------------------------------------- FindCycle */
/* cycle_finder.cpp:40-52 */
static int background_check(const orc_graph *g, const uint8_t *visited, uint64_t from, uint64_t mu, uint64_t nb) {
  uint64_t m = g->mult[nb];
  if (visited[nb]) return 0;
  if (m == 0) return 0; /* the reference would divide by zero; unreachable: mult<=1 nodes are invalid */
  if (mu / m > 500) return 0;
  if (from == nb) return 0;
  return 1;
}
/* cycle_finder.cpp:58-72 */
static void get_outgoings(const orc_graph *g, const uint8_t *visited, uint64_t node, uint64_t mu, uset *s, uint64_t *expansions) {
  s->n = 0;
  uint64_t nb[4];
  ++*expansions;
  int deg = orc_outgoing(g, node, nb);
  if (deg <= 0) return;
  for (int i = 0; i < deg; ++i)
    if (background_check(g, visited, node, mu, nb[i]) && !g->invalid[nb[i]]) uset_insert(s, nb[i]);
}
/* cycle_finder.cpp:76-88 */
static void get_incomings(const orc_graph *g, const uint8_t *visited, uint64_t node, uint64_t mu, uset *s, uint64_t *expansions) {
  s->n = 0;
  uint64_t nb[4];
  ++*expansions;
  int deg = orc_incoming(g, node, nb);
  if (deg <= 0) return;
  for (int i = 0; i < deg; ++i)
    if (background_check(g, visited, node, mu, nb[i]) && !g->invalid[nb[i]]) uset_insert(s, nb[i]);
}

void orc_cycles_free(orc_cycles *c) { free(c->node_off); free(c->nodes); memset(c, 0, sizeof *c); }

/* cycle_finder.cpp:231-243 (FindCycleUtil) + :140-225 (FindCycle). */
void orc_find_cycle_util(const orc_graph *g, uint64_t start, int min_len, int max_len,
                         uint8_t *visited, orc_cycles *out) {
  const int M = max_len;
  const uint64_t mu = g->mult[start];
  memset(out, 0, sizeof *out);
  int depth_cap = M + 4;
  if (depth_cap < 8) depth_cap = 8;
  uint64_t *path = (uint64_t *)malloc((size_t)depth_cap * 8);
  uset *stack = (uset *)malloc((size_t)depth_cap * sizeof(uset));
  int *bl = (int *)malloc((size_t)depth_cap * sizeof(int));
  int plen = 0, slen = 0, blen = 0;
  hmap lock;
  hm_init(&lock, 1024);
  vec64 nodes = {0, 0, 0}, offs = {0, 0, 0};
  v_push(&offs, 0);
  typedef struct { int bl; int32_t u; } relax_ent;
  relax_ent *rs = (relax_ent *)malloc(64 * sizeof(relax_ent));
  uint64_t rs_cap = 64;

  path[plen++] = start;
  *hm_get(&lock, start, 1, 0) = 0;                       /* :237 */
  get_outgoings(g, visited, start, mu, &stack[slen++], &out->expansions);
  bl[blen++] = M;
  int counter = 0, steps = 0;
  uint64_t n_cycles = 0;
  while (slen > 0) {
    if (++steps > 10000000) { out->cap_hit = 1; break; } /* :148-151 */
    uset nbs = stack[slen - 1];                          /* :153 copy */
    int flag = 1;
    for (int i = 0; i < nbs.n; ++i) {
      uint64_t nb = nbs.e[i];
      if (nb == start) {                                 /* :157-166 */
        bl[blen - 1] = 1;
        if (plen > min_len) {
          for (int j = 0; j < plen; ++j) v_push(&nodes, path[j]);
          v_push(&offs, nodes.n);
          ++n_cycles;
          if (++counter >= 500) {                        /* cluster_bounds, :132,:162-165 */
            nodes.n = 0; offs.n = 1; n_cycles = 0;
            flag = 0;
            out->bound_hit = 1;
          }
        }
      } else {
        int *lk = hm_get(&lock, nb, 1, M);               /* try_emplace(nb, M), :168 */
        if (plen < *lk) {
          path[plen++] = nb;                             /* :170 */
          bl[blen++] = M;                                /* :171 */
          *lk = plen;                                    /* :172 */
          uset_erase(&stack[slen - 1], nb);              /* :173 */
          get_outgoings(g, visited, nb, mu, &stack[slen], &out->expansions); /* :174-176 */
          ++slen;
          flag = 0;
          break;
        }
      }
    }
    if (flag) {                                          /* :181-211 */
      --slen;
      uint64_t v = path[--plen];
      int b = bl[--blen];
      if (blen > 0 && b < bl[blen - 1]) bl[blen - 1] = b;
      if (b < M) {
        uint64_t rn = 0;
        rs[rn].bl = b; rs[rn].u = (int32_t)v; ++rn;      /* pair<int,int>, :192-193 (F12) */
        while (rn) {
          relax_ent r = rs[--rn];
          uint64_t u = (uint64_t)(int64_t)r.u;
          int *lk = hm_get(&lock, u, 1, M);
          if (*lk < M - r.bl + 1) {
            *lk = M - r.bl + 1;
            uset inc;
            get_incomings(g, visited, u, mu, &inc, &out->expansions);
            for (int i = 0; i < inc.n; ++i) {
              int on_path = 0;
              for (int j = 0; j < plen; ++j) if (path[j] == inc.e[i]) { on_path = 1; break; }
              if (!on_path) {
                if (rn == rs_cap) { rs_cap *= 2; rs = (relax_ent *)realloc(rs, rs_cap * sizeof(relax_ent)); }
                rs[rn].bl = r.bl + 1; rs[rn].u = (int32_t)inc.e[i]; ++rn;
              }
            }
          }
        }
      }
    }
  }
  out->steps = (uint64_t)steps;
  /* :215-222 */
  for (uint64_t i = 0; i < nodes.n; ++i) visited[nodes.a[i]] = 1;
  out->n_cycles = n_cycles;
  out->node_off = offs.a;
  out->nodes = nodes.a;
  free(path); free(stack); free(bl); free(rs);
  hm_free(&lock);
}

/* ---------------------------------------------------------------- driver */
/* cycle_finder.cpp:29-36 */
static int has_self_incoming(const orc_graph *g, uint64_t node) {
  uint64_t in[4];
  int deg = orc_incoming(g, node, in);
  for (int i = 0; i < deg && i < 4; ++i) if (in[i] == node) return 1;
  return 0;
}

void orc_result_free(orc_result *r) {
  free(r->bucket_key); free(r->bucket_off); free(r->start_ids); free(r->searched_start);
  free(r->cyc_off); free(r->node_off); free(r->nodes);
  memset(r, 0, sizeof *r);
}

/* cycle_finder.cpp:433-492 at threads=1 (the only deterministic setting). */
int orc_run(orc_graph *g, uint64_t thr, int min_len, int max_len, orc_result *res) {
  memset(res, 0, sizeof *res);
  uint64_t n = g->n;
  uint64_t *tips = NULL;
  uint64_t nt = orc_collect_tips(g, &tips);              /* :436 */
  res->n_tips_before = nt;
  res->n_mult1 = orc_invalidate_mult1(g);                /* :439 */
  orc_peel(g, tips, nt);                                 /* :440-442 */
  free(tips);
  for (uint64_t i = 0; i < n; ++i) res->n_valid += !g->invalid[i];
  res->n_tips_after = orc_collect_tips(g, NULL);         /* :451 */

  /* ChunkStartNodes, :387-427 */
  enum { NB = 64 };
  vec64 bucket[NB];
  memset(bucket, 0, sizeof bucket);
  for (uint64_t v = 0; v < n; ++v) {
    if (g->invalid[v]) continue;
    int indeg = orc_incoming(g, v, NULL);
    if (!(indeg >= 2 && (uint64_t)g->mult[v] > thr)) continue;
    if (has_self_incoming(g, v)) continue;
    res->n_candidates++;
    if (!orc_dls(g, v, max_len, &res->dls_expansions)) continue;
    int key = (int)ceil(log2((double)g->mult[v]));       /* :414 */
    v_push(&bucket[key], v);
  }
  uint64_t total = 0;
  for (int b = 0; b < NB; ++b) { if (bucket[b].n) res->n_buckets++; total += bucket[b].n; }
  res->bucket_key = (int64_t *)malloc((res->n_buckets + 1) * 8);
  res->bucket_off = (uint64_t *)malloc((res->n_buckets + 1) * 8);
  res->start_ids = (uint64_t *)malloc((total + 1) * 8);
  uint64_t bi = 0, off = 0;
  for (int b = NB - 1; b >= 0; --b) {                    /* map<int,...,greater<int>>: descending */
    if (!bucket[b].n) continue;
    res->bucket_key[bi] = b;
    res->bucket_off[bi] = off;
    memcpy(res->start_ids + off, bucket[b].a, bucket[b].n * 8);
    off += bucket[b].n;
    ++bi;
  }
  res->bucket_off[bi] = off;
  for (int b = 0; b < NB; ++b) free(bucket[b].a);

  /* enumeration, :468-487 */
  uint8_t *visited = (uint8_t *)calloc(n ? n : 1, 1);    /* :458 */
  vec64 searched = {0, 0, 0}, cyc_off = {0, 0, 0}, node_off = {0, 0, 0}, nodes = {0, 0, 0};
  v_push(&cyc_off, 0);
  v_push(&node_off, 0);
  for (uint64_t i = 0; i < total; ++i) {
    uint64_t s = res->start_ids[i];
    if (visited[s]) continue;                            /* :476 */
    orc_cycles c;
    orc_find_cycle_util(g, s, min_len, max_len, visited, &c);
    res->n_bound_hits += c.bound_hit;
    res->n_cap_hits += c.cap_hit;
    res->enum_expansions += c.expansions;
    v_push(&searched, s);                                /* results[start] = cycles, :479 (empty stored) */
    for (uint64_t j = 0; j < c.n_cycles; ++j) {
      for (uint64_t p = c.node_off[j]; p < c.node_off[j + 1]; ++p) v_push(&nodes, c.nodes[p]);
      v_push(&node_off, nodes.n);
    }
    v_push(&cyc_off, node_off.n - 1);
    orc_cycles_free(&c);
  }
  free(visited);
  res->n_searched = searched.n;
  res->searched_start = searched.a ? searched.a : (uint64_t *)malloc(8);
  res->cyc_off = cyc_off.a;
  res->n_cycles = node_off.n - 1;
  res->node_off = node_off.a;
  res->nodes = nodes.a ? nodes.a : (uint64_t *)malloc(8);
  return 0;
}
