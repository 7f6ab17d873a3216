/* TEST INFRASTRUCTURE ONLY — a plain-C CPU restatement of the reference's kcount CPU path, used as the parity
 * checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg. Never linked into or called from the
 * product path (mhm2_proxy_b200/), which is CUDA only.
 *
 * Parity status: PINNED. This restatement is checked (tests/test_oracle.py) against the reference's own code compiled
 * from /root/reference (oracle/_ref/libkcount_ref.so, see ref_driver.cpp) on seeded inputs, and against the golden
 * vectors under tests/golden/ that were generated from that build (tests/golden/make_golden.py).
 *
 * Each function cites the reference file:line it follows (paths relative to /root/reference/src).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>

#define KO_MAX_LONGS 4
#define KCOUNT_QUAL_CUTOFF 20 /* CMakeDefinitions.txt:46 */
#define DYN_MIN_DEPTH 0.9     /* CMakeDefinitions.txt:60 */

/* ---------------------------------------------------------------- hashes ---------------------------------------- */

static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

/* hash_funcs.c:65-74 */
static inline uint64_t fmix64(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdULL;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ULL;
  k ^= k >> 33;
  return k;
}

/* hash_funcs.c:78-176 MurmurHash3_x64_128 with seed 313, low 64 bits (hash_funcs.c:185-190), for keys that are a
 * whole number of 8-byte words (Kmer::hash passes N_LONGS*8 bytes, kmer.cpp:465-468). */
uint64_t ko_murmur64_words(const uint64_t *w, int nwords) {
  const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  uint64_t h1 = 313, h2 = 313;
  uint32_t len = (uint32_t)nwords * 8;
  int nblocks = nwords / 2;
  for (int i = 0; i < nblocks; i++) {
    uint64_t k1 = w[2 * i], k2 = w[2 * i + 1];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  }
  if (nwords & 1) { /* tail of exactly 8 bytes: `case 8` .. `case 1` of the switch, little-endian byte assembly */
    uint64_t k1 = w[nwords - 1];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
  }
  h1 ^= len; h2 ^= len;
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2;
  return h1;
}

/* hash_funcs.c:332-342 */
uint64_t ko_quick_hash(uint64_t v) {
  v = v * 3935559000370003845ULL + 2691343689449507681ULL;
  v ^= v >> 21;
  v ^= v << 37;
  v ^= v >> 4;
  v *= 4768777513237032717ULL;
  v ^= v << 20;
  v ^= v >> 41;
  v ^= v << 5;
  return v;
}

/* ---------------------------------------------------------------- k-mers ---------------------------------------- */

typedef struct { uint64_t w[KO_MAX_LONGS]; } ko_kmer;

/* kmer.cpp:187-188 / 307-308: A0 C1 G2 T3, and N (or anything with bits like it) -> 2 */
static inline uint64_t base_code(char c) {
  uint64_t x = (uint64_t)((c & 4) >> 1);
  return x + ((x ^ (uint64_t)(c & 2)) >> 1);
}

/* kmer.cpp:297-320 set_kmer: base i at bits 2*(31 - i%32) of word i/32 */
static void kmer_from_ascii(const char *s, int k, int n_longs, ko_kmer *out) {
  memset(out, 0, sizeof(*out));
  for (int i = 0; i < k; i++) out->w[i / 32] |= base_code(s[i]) << (2 * (31 - (i % 32)));
  (void)n_longs;
}

static inline int kmer_base(const ko_kmer *km, int i) { return (int)((km->w[i / 32] >> (2 * (31 - (i % 32)))) & 3); }

/* kmer.cpp:485-505 revcomp: base i of the result is the complement (3 - code) of base k-1-i */
static void kmer_revcomp(const ko_kmer *km, int k, ko_kmer *out) {
  memset(out, 0, sizeof(*out));
  for (int i = 0; i < k; i++) {
    uint64_t c = 3 - (uint64_t)kmer_base(km, k - 1 - i);
    out->w[i / 32] |= c << (2 * (31 - (i % 32)));
  }
}

/* kmer.cpp:265-272 operator< : word-wise unsigned, word 0 first */
static inline int kmer_less(const ko_kmer *a, const ko_kmer *b, int n_longs) {
  for (int i = 0; i < n_longs; i++) {
    if (a->w[i] < b->w[i]) return 1;
    if (a->w[i] > b->w[i]) return 0;
  }
  return 0;
}
static inline int kmer_eq(const ko_kmer *a, const ko_kmer *b, int n_longs) {
  for (int i = 0; i < n_longs; i++)
    if (a->w[i] != b->w[i]) return 0;
  return 1;
}

/* kmer.cpp:345-393 get_minimizer_fast(m, revcomp): the GREATEST over p in [0,k-m] of min(fwd m-mer at p, its reverse
 * complement), m-mers held left-aligned in a u64 (ZERO_MASK, kmer.cpp:81-87). */
uint64_t ko_minimizer(const ko_kmer *km, int k, int m) {
  uint64_t best = 0;
  for (int p = 0; p + m <= k; p++) {
    uint64_t f = 0, r = 0;
    for (int j = 0; j < m; j++) {
      f |= (uint64_t)kmer_base(km, p + j) << (2 * (31 - j));
      r |= (uint64_t)(3 - kmer_base(km, p + m - 1 - j)) << (2 * (31 - j));
    }
    uint64_t least = f < r ? f : r;
    if (least > best) best = least;
  }
  return best;
}

/* kmer_dht.cpp:113-116 */
int ko_minimizer_len(int k) {
  int m = k * 2 / 3 + 1;
  if (m < 15) m = 15;
  if (m > 27) m = 27;
  return m;
}

/* kmer_dht.cpp:192-196 + kmer.cpp:455-463 */
int ko_owner(const ko_kmer *km, int k, int num_shards) {
  return (int)(ko_quick_hash(ko_minimizer(km, k, ko_minimizer_len(k))) % (uint64_t)num_shards);
}

/* ------------------------------------------------------ extension counters and the table ------------------------ */

typedef struct {          /* kcount_cpu.cpp:191-200 KmerExtsCounts (ExtCounts = 4 x u16, kcount_cpu.cpp:115-119) */
  uint16_t left[4], right[4]; /* A C G T */
  uint16_t count;
  uint8_t from_ctg;
  uint8_t used;
} ko_slot;

static int g_dmin_thres = 2; /* kmer_dht.hpp:57 _dmin_thres */

static inline uint16_t inc_with_limit(int a, int b) { /* kcount_cpu.cpp:152-155 */
  a += b;
  return (uint16_t)(a < 65535 ? a : 65535);
}
static inline void ext_inc(uint16_t *e, char ext, int count) { /* kcount_cpu.cpp:157-164: only A,C,G,T count */
  switch (ext) {
    case 'A': e[0] = inc_with_limit(e[0], count); break;
    case 'C': e[1] = inc_with_limit(e[1], count); break;
    case 'G': e[2] = inc_with_limit(e[2], count); break;
    case 'T': e[3] = inc_with_limit(e[3], count); break;
  }
}

/* kcount_cpu.cpp:135-145 get_sorted + 173-182 get_ext: counts descending, ties -> larger letter first */
char ko_get_ext(const uint16_t *e, uint16_t count, int dmin_thres) {
  static const char letters[4] = {'A', 'C', 'G', 'T'};
  int idx[4] = {0, 1, 2, 3};
  for (int i = 0; i < 4; i++) /* selection sort with the reference comparator */
    for (int j = i + 1; j < 4; j++) {
      int a = idx[i], b = idx[j];
      int b_first = (e[b] == e[a]) ? (letters[b] > letters[a]) : (e[b] > e[a]);
      if (b_first) { idx[i] = b; idx[j] = a; }
    }
  int top_count = e[idx[0]], runner_up_count = e[idx[1]];
  int dmin_dyn = (int)((1.0 - DYN_MIN_DEPTH) * count);
  if (dmin_dyn < dmin_thres) dmin_dyn = dmin_thres;
  if (top_count < dmin_dyn) return 'X';
  if (runner_up_count >= dmin_dyn) return 'F';
  return letters[idx[0]];
}

/* Growable open-addressing map. The reference's table is fixed-capacity with linear probing and drops on overflow
 * (kcount_cpu.cpp:231-267); parity is defined for runs with zero drops and zero singleton overrides,
 * where the table behaves as an exact map — which is what this is. Slot hash = the reference's (Kmer::hash). */
typedef struct {
  ko_kmer *keys;
  ko_slot *vals;
  uint64_t cap, n;
  int n_longs;
} ko_map;

static void map_init(ko_map *m, uint64_t cap, int n_longs) {
  m->cap = cap; m->n = 0; m->n_longs = n_longs;
  m->keys = (ko_kmer *)calloc(cap, sizeof(ko_kmer));
  m->vals = (ko_slot *)calloc(cap, sizeof(ko_slot));
}
static void map_free(ko_map *m) { free(m->keys); free(m->vals); }

static ko_slot *map_insert(ko_map *m, const ko_kmer *km, int *is_new);
static void map_grow(ko_map *m) {
  ko_map n;
  map_init(&n, m->cap * 2, m->n_longs);
  for (uint64_t i = 0; i < m->cap; i++)
    if (m->vals[i].used) {
      int is_new;
      ko_slot *s = map_insert(&n, &m->keys[i], &is_new);
      *s = m->vals[i];
    }
  map_free(m);
  *m = n;
}
static ko_slot *map_insert(ko_map *m, const ko_kmer *km, int *is_new) {
  if ((m->n + 1) * 10 > m->cap * 6) map_grow(m);
  uint64_t slot = ko_murmur64_words(km->w, m->n_longs) & (m->cap - 1);
  for (;;) {
    if (!m->vals[slot].used) {
      m->keys[slot] = *km;
      memset(&m->vals[slot], 0, sizeof(ko_slot));
      m->vals[slot].used = 1;
      m->n++;
      *is_new = 1;
      return &m->vals[slot];
    }
    if (kmer_eq(&m->keys[slot], km, m->n_longs)) { *is_new = 0; return &m->vals[slot]; }
    slot = (slot + 1) & (m->cap - 1);
  }
}

/* utils.cpp:121-143 comp_nucleotide on the characters that can reach it here */
static inline char comp_nt(char c) {
  switch (c) {
    case 'A': return 'T';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'T': return 'A';
    default: return c; /* 'N' -> 'N', '0' -> '0' */
  }
}

/* ------------------------------------------------------ the two passes ------------------------------------------ */

typedef struct {
  int k, n_longs, num_shards, dmin_thres;
  ko_map map;
  uint64_t instances, ctg_instances, distinct_after_reads;
  /* outputs */
  uint64_t n_surv, n_pf;
  uint64_t *keys; uint16_t *counts; char *left, *right; int32_t *shard;
  uint64_t *pf_keys; uint16_t *pf_vals;
} ko_result;

/* One sequence through SeqBlockInserter::process_seq (kcount_cpu.cpp:73-103) and, at the owner, get_kmers_and_exts
 * (kcount_cpu.cpp:307-335) + insert_supermer_from_read (337-354) / insert_supermer_from_ctg (356-406).
 * `seq` uses the reference's convention: lower case = low quality. Supermer boundaries do not affect the result
 * (each k-mer at i in [1, len-k-1] is delivered exactly once with seq[i-1], seq[i+k] as its neighbours), so the
 * restatement visits the k-mers directly. */
static void process_seq(ko_result *R, const char *seq, size_t len, uint16_t depth, int is_ctg) {
  const int k = R->k, nl = R->n_longs;
  if (len < (size_t)k + 2) return;
  if (!depth) depth = 1; /* kcount_cpu.cpp:75 */
  char up[KO_MAX_LONGS * 32 + 1];
  for (size_t i = 1; i + k < len; i++) { /* i = 1 .. len-k-1 : first and last k-mer are never counted */
    for (int j = 0; j < k; j++) up[j] = (char)toupper((unsigned char)seq[i + j]); /* kcount_cpu.cpp:313 */
    ko_kmer km, rc;
    kmer_from_ascii(up, k, nl, &km); /* N -> G (kmer.cpp:169) */
    char left = (char)toupper((unsigned char)seq[i - 1]);
    if (!isupper((unsigned char)seq[i - 1])) left = '0'; /* kcount_cpu.cpp:320-321 */
    char right = (char)toupper((unsigned char)seq[i + k]);
    if (!isupper((unsigned char)seq[i + k])) right = '0';
    kmer_revcomp(&km, k, &rc);
    if (kmer_less(&rc, &km, nl)) { /* kcount_cpu.cpp:326-331: strict, palindromes keep the forward orientation */
      km = rc;
      char t = left; left = right; right = t;
      left = comp_nt(left);
      right = comp_nt(right);
    }
    int is_new;
    ko_slot *s = map_insert(&R->map, &km, &is_new);
    if (!is_ctg) { /* kcount_cpu.cpp:343-352 */
      s->count = inc_with_limit(s->count, depth);
      ext_inc(s->left, left, depth);
      ext_inc(s->right, right, depth);
      R->instances++;
    } else { /* kcount_cpu.cpp:362-404 */
      R->ctg_instances++;
      uint16_t cnt = depth;
      int insert_it = 0;
      if (is_new) {
        insert_it = 1;
      } else if (!s->from_ctg) {
        if (s->count == 1) {
          insert_it = 1;
        } else {
          char le = ko_get_ext(s->left, s->count, g_dmin_thres), re = ko_get_ext(s->right, s->count, g_dmin_thres);
          if (le == 'X' || le == 'F' || re == 'X' || re == 'F') insert_it = 1;
        }
      } else if (s->count) {
        insert_it = 1;
        char le = ko_get_ext(s->left, s->count, g_dmin_thres), re = ko_get_ext(s->right, s->count, g_dmin_thres);
        if (le != left || re != right) cnt = 0;
        else cnt = cnt < s->count ? cnt : s->count;
      }
      if (insert_it) {
        memset(s->left, 0, sizeof(s->left));
        memset(s->right, 0, sizeof(s->right));
        s->count = cnt;
        s->from_ctg = 1;
        ext_inc(s->left, left, cnt);
        ext_inc(s->right, right, cnt);
      }
    }
  }
}

static int row_cmp_nl;
static int row_cmp(const void *a, const void *b) {
  const uint64_t *x = (const uint64_t *)a, *y = (const uint64_t *)b;
  for (int i = 0; i < row_cmp_nl; i++) {
    if (x[i] < y[i]) return -1;
    if (x[i] > y[i]) return 1;
  }
  return 0;
}

/* reads: PackedRead bytes (packed_reads.cpp:73-109: base 0..4 in the low 3 bits, min(qual,31) in the high 5) with
 * offsets[n+1]; contigs: ASCII with offsets and depths (contigs.hpp:57-66 truncation applied here). */
void *ko_run(int k, int qual_offset, int dmin_thres, int num_shards, const uint8_t *bytes, const uint64_t *offs, uint64_t n_reads,
             const char *cseqs, const uint64_t *coffs, const double *cdepths, uint64_t n_ctgs, int want_prefilter) {
  (void)qual_offset; /* (q + off) < off + 20  <=>  q < 20 : the offset cancels (kcount.cpp:81) */
  if (k < 1 || k >= KO_MAX_LONGS * 32) return NULL;
  ko_result *R = (ko_result *)calloc(1, sizeof(ko_result));
  R->k = k; R->n_longs = k / 32 + 1; /* main.cpp:170 MAX_K = (k/32+1)*32 ; N_LONGS = MAX_K/32 */
  R->num_shards = num_shards; R->dmin_thres = dmin_thres;
  g_dmin_thres = dmin_thres; /* kcount.cpp:145 */
  map_init(&R->map, 1 << 16, R->n_longs);
  static const char nucleotide_map[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'}; /* packed_reads.hpp:61 */
  size_t cap = 1024;
  char *seq = (char *)malloc(cap);
  for (uint64_t r = 0; r < n_reads; r++) { /* kcount.cpp:72-88 */
    size_t len = offs[r + 1] - offs[r];
    if (len < (size_t)k) continue;
    if (len > cap) { cap = len * 2; seq = (char *)realloc(seq, cap); }
    const uint8_t *b = bytes + offs[r];
    for (size_t i = 0; i < len; i++) {
      char c = nucleotide_map[b[i] & 7];
      if ((b[i] >> 3) < KCOUNT_QUAL_CUTOFF) c = (char)tolower(c);
      seq[i] = c;
    }
    process_seq(R, seq, len, 0, 0);
  }
  free(seq);
  R->distinct_after_reads = R->map.n;
  const int nl = R->n_longs;
  if (want_prefilter) {
    R->n_pf = R->map.n;
    R->pf_keys = (uint64_t *)malloc(sizeof(uint64_t) * nl * (R->n_pf + 1));
    R->pf_vals = (uint16_t *)malloc(sizeof(uint16_t) * 9 * (R->n_pf + 1));
    uint64_t j = 0;
    for (uint64_t i = 0; i < R->map.cap; i++)
      if (R->map.vals[i].used) {
        memcpy(R->pf_keys + j * nl, R->map.keys[i].w, 8 * nl);
        uint16_t *v = R->pf_vals + j * 9;
        v[0] = R->map.vals[i].count;
        memcpy(v + 1, R->map.vals[i].left, 8);
        memcpy(v + 5, R->map.vals[i].right, 8);
        j++;
      }
  }
  for (uint64_t c = 0; c < n_ctgs; c++) { /* kcount.cpp:123-130 */
    size_t len = coffs[c + 1] - coffs[c];
    if (len < (size_t)k + 2) continue;
    double d = cdepths[c];
    uint16_t depth = (uint16_t)(d > 65535 ? 65535 : d); /* contigs.hpp:57-66 */
    process_seq(R, cseqs + coffs[c], len, depth, 1);
  }
  /* kcount_cpu.cpp:490-528 insert_into_local_hashtable */
  uint64_t n = 0;
  size_t roww = (size_t)nl + 1; /* key words + packed (count,left,right) for sorting */
  uint64_t *rows = (uint64_t *)malloc(sizeof(uint64_t) * roww * (R->map.n + 1));
  for (uint64_t i = 0; i < R->map.cap; i++) {
    ko_slot *s = &R->map.vals[i];
    if (!s->used) continue;
    if (s->count < 2) continue;
    char le = ko_get_ext(s->left, s->count, dmin_thres), re = ko_get_ext(s->right, s->count, dmin_thres);
    if (le == 'X' && re == 'X') continue;
    memcpy(rows + n * roww, R->map.keys[i].w, 8 * nl);
    rows[n * roww + nl] = ((uint64_t)s->count << 16) | ((uint64_t)(uint8_t)le << 8) | (uint8_t)re;
    n++;
  }
  row_cmp_nl = nl;
  qsort(rows, n, sizeof(uint64_t) * roww, row_cmp);
  R->n_surv = n;
  R->keys = (uint64_t *)malloc(8 * nl * (n + 1));
  R->counts = (uint16_t *)malloc(2 * (n + 1));
  R->left = (char *)malloc(n + 1); R->right = (char *)malloc(n + 1);
  R->shard = (int32_t *)malloc(4 * (n + 1));
  for (uint64_t i = 0; i < n; i++) {
    memcpy(R->keys + i * nl, rows + i * roww, 8 * nl);
    uint64_t v = rows[i * roww + nl];
    R->counts[i] = (uint16_t)(v >> 16); R->left[i] = (char)((v >> 8) & 0xff); R->right[i] = (char)(v & 0xff);
    ko_kmer km; memset(&km, 0, sizeof(km)); memcpy(km.w, rows + i * roww, 8 * nl);
    R->shard[i] = ko_owner(&km, k, num_shards);
  }
  free(rows);
  return R;
}

int ko_n_longs(void *h) { return ((ko_result *)h)->n_longs; }
uint64_t ko_num_survivors(void *h) { return ((ko_result *)h)->n_surv; }
uint64_t ko_num_prefilter(void *h) { return ((ko_result *)h)->n_pf; }
void ko_copy_survivors(void *h, uint64_t *keys, uint16_t *counts, char *left, char *right, int32_t *shard) {
  ko_result *R = (ko_result *)h;
  memcpy(keys, R->keys, 8 * R->n_longs * R->n_surv);
  memcpy(counts, R->counts, 2 * R->n_surv);
  memcpy(left, R->left, R->n_surv); memcpy(right, R->right, R->n_surv);
  memcpy(shard, R->shard, 4 * R->n_surv);
}
void ko_copy_prefilter(void *h, uint64_t *keys, uint16_t *vals) {
  ko_result *R = (ko_result *)h;
  memcpy(keys, R->pf_keys, 8 * R->n_longs * R->n_pf);
  memcpy(vals, R->pf_vals, 18 * R->n_pf);
}
/* stats: instances, distinct after reads, dropped(0), overrides(0), supermers(0), supermer bytes(0), ctg instances, purged */
void ko_stats(void *h, uint64_t *stats8, double *times3) {
  ko_result *R = (ko_result *)h;
  memset(stats8, 0, 64);
  stats8[0] = R->instances; stats8[1] = R->distinct_after_reads; stats8[6] = R->ctg_instances;
  stats8[7] = R->map.n - R->n_surv;
  times3[0] = times3[1] = times3[2] = 0;
}
void ko_free(void *h) {
  ko_result *R = (ko_result *)h;
  map_free(&R->map);
  free(R->keys); free(R->counts); free(R->left); free(R->right); free(R->shard); free(R->pf_keys); free(R->pf_vals);
  free(R);
}

/* primitives on an ASCII k-mer, mirroring oref_kmer_prims in ref_driver.cpp */
int ko_kmer_prims(int k, const char *s, int num_shards, uint64_t *words, uint64_t *rc_words, uint64_t *minimizer, uint64_t *hash,
                  int32_t *owner, int32_t *m_out) {
  int nl = k / 32 + 1;
  ko_kmer km, rc;
  kmer_from_ascii(s, k, nl, &km);
  kmer_revcomp(&km, k, &rc);
  memcpy(words, km.w, 8 * nl); memcpy(rc_words, rc.w, 8 * nl);
  *m_out = ko_minimizer_len(k);
  *minimizer = ko_minimizer(&km, k, *m_out);
  *hash = ko_murmur64_words(km.w, nl);
  *owner = ko_owner(&km, k, num_shards);
  return nl;
}
uint64_t ko_murmur64(const void *key, uint32_t len) { return ko_murmur64_words((const uint64_t *)key, (int)(len / 8)); }
char ko_get_ext4(uint16_t a, uint16_t c, uint16_t g, uint16_t t, uint16_t count, int dmin_thres) {
  uint16_t e[4] = {a, c, g, t};
  return ko_get_ext(e, count, dmin_thres);
}
