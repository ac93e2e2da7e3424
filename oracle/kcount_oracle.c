/*
 * TEST INFRASTRUCTURE -- CPU oracle for the k-mer analysis hot path (see kcount_oracle.h).
 * Plain C restatement of the reference's CPU build; every function names the reference lines it follows
 * (paths relative to the reference tree). Never on the product path.
 */
#include "kcount_oracle.h"

#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MAXW 4
#define KEY_EMPTY 0xffffffffffffffffULL
#define HT_MAX_PROBE 100 /* KCOUNT_HT_MAX_PROBE, CMakeDefinitions.txt:67 */

/* ------------------------------------------------------------------ hashes (src/hash_funcs.c) ---- */

/* src/hash_funcs.c:332-342 */
uint64_t kco_quick_hash(uint64_t v) {
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

static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
static inline uint64_t fmix64(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdULL;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ULL;
  k ^= k >> 33;
  return k;
}

/* MurmurHash3_x64_128 with seed 313, first 64 bits: src/hash_funcs.c:60-190 (Kmer::hash, src/kmer.cpp:471-473).
 * Only whole 8-byte words are ever hashed on this path (len = 8 * N_LONGS), so the tail switch reduces
 * to the "case 8" arm. */
uint64_t kco_murmur64(const void *key, uint32_t len) {
  const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  const uint64_t *blocks = (const uint64_t *)key;
  uint32_t nblocks = len / 16;
  uint64_t h1 = 313, h2 = 313;
  for (uint32_t i = 0; i < nblocks; i++) {
    uint64_t k1 = blocks[2 * i], k2 = blocks[2 * i + 1];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  }
  const uint8_t *tail = (const uint8_t *)key + nblocks * 16;
  uint64_t k1 = 0, k2 = 0;
  int rem = len & 15;
  for (int i = rem - 1; i >= 8; i--) k2 ^= (uint64_t)tail[i] << (8 * (i - 8));
  if (rem > 8) { k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2; }
  for (int i = (rem > 8 ? 8 : rem) - 1; i >= 0; i--) k1 ^= (uint64_t)tail[i] << (8 * i);
  if (rem > 0) { k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1; }
  h1 ^= len; h2 ^= len;
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2;
  return h1;
}

/* ------------------------------------------------------------------ Kmer (src/kmer.cpp) ---------- */

/* base code of an upper-case letter: Kmer::set_kmer / get_kmers, src/kmer.cpp:191-192,313-314.
 * A0 C1 G2 T3 and, as a consequence of the bit trick, N -> 2 (G). */
static inline uint64_t base_code(char c) {
  uint64_t x = ((uint64_t)(c & 4)) >> 1;
  return x + ((x ^ (uint64_t)(c & 2)) >> 1);
}

static inline int n_words_of(int k) { return (k + 31) / 32; } /* N_LONGS of MAX_K = (k/32+1)*32, main.cpp:169 */

/* src/kmer.cpp:303-325: base i -> bits 2*(31 - i%32) of word i/32, zero padded */
void kco_pack_kmer(const char *s, int k, uint64_t *w) {
  int n = n_words_of(k);
  for (int l = 0; l < n; l++) w[l] = 0;
  for (int i = 0; i < k; i++) {
    char c = s[i];
    if (c >= 'a' && c <= 'z') c += 'A' - 'a';
    w[i / 32] |= base_code(c) << (2 * (31 - (i % 32)));
  }
}

static inline uint64_t get_base(const uint64_t *w, int i) { return (w[i / 32] >> (2 * (31 - (i % 32)))) & 3; }

/* Kmer::revcomp, src/kmer.cpp:491-510 (there: byte twin table + cross-word shift; here base by base) */
void kco_revcomp(const uint64_t *in, int k, uint64_t *out) {
  int n = n_words_of(k);
  uint64_t tmp[MAXW] = {0, 0, 0, 0};
  for (int i = 0; i < k; i++) tmp[i / 32] |= (3 - get_base(in, k - 1 - i)) << (2 * (31 - (i % 32)));
  for (int l = 0; l < n; l++) out[l] = tmp[l];
}

/* Kmer::operator<, src/kmer.cpp:271-277 */
int kco_kmer_less(const uint64_t *a, const uint64_t *b, int n) {
  for (int i = 0; i < n; i++) {
    if (a[i] < b[i]) return 1;
    if (a[i] > b[i]) return 0;
  }
  return 0;
}

/* Kmer::get_minimizer_fast(m, &revcomp), src/kmer.cpp:345-398: the GREATEST over positions i of the LEAST
 * of (forward m-mer at i, m-mer at k-m-i of the reverse complement). The latter is the reverse complement
 * of the former, so both are rolled along the k-mer here (right aligned; same order as the reference's
 * left-aligned "& ZERO_MASK[m]" form), and the winner is returned left aligned as the reference does. */
uint64_t kco_minimizer(const uint64_t *kmer, int k, int m) {
  const uint64_t mask = m < 32 ? ((1ULL << (2 * m)) - 1) : ~0ULL;
  uint64_t f = 0, r = 0, best = 0;
  for (int i = 0; i < k; i++) {
    uint64_t b = get_base(kmer, i);
    f = ((f << 2) | b) & mask;
    r = (r >> 2) | ((3 - b) << (2 * (m - 1)));
    if (i >= m - 1) {
      uint64_t least = f < r ? f : r;
      if (least > best) best = least;
    }
  }
  return best << (64 - 2 * m);
}

/* KmerDHT::get_kmer_target_rank, src/kcount/kmer_dht.cpp:193-196 with Kmer::minimizer_hash_fast, src/kmer.cpp:460-468 */
int kco_owner(const uint64_t *kmer, int k, int m, int n_ranks) {
  return (int)(kco_quick_hash(kco_minimizer(kmer, k, m)) % (uint64_t)n_ranks);
}

/* ------------------------------------------------------------------ ExtCounts (kcount_cpu.cpp:115-189) */

/* ExtCounts::get_sorted + get_ext, src/kcount/kcount_cpu.cpp:135-145,173-182 */
char kco_get_ext(const uint16_t counts[4], uint16_t count, int dmin_thres, double dyn_min_depth) {
  static const char letters[4] = {'A', 'C', 'G', 'T'};
  int idx[4] = {0, 1, 2, 3};
  /* sort descending by count, ties: higher letter first */
  for (int a = 0; a < 4; a++)
    for (int b = a + 1; b < 4; b++) {
      int ia = idx[a], ib = idx[b];
      int b_first = counts[ib] > counts[ia] || (counts[ib] == counts[ia] && letters[ib] > letters[ia]);
      if (b_first) { idx[a] = ib; idx[b] = ia; }
    }
  int top_count = counts[idx[0]], runner_up_count = counts[idx[1]];
  int dmin_dyn = (int)((1.0 - dyn_min_depth) * count);
  if (dmin_dyn < dmin_thres) dmin_dyn = dmin_thres;
  if (top_count < dmin_dyn) return 'X';
  if (runner_up_count >= dmin_dyn) return 'F';
  return letters[idx[0]];
}

/* comp_nucleotide, src/utils.cpp:132-159 */
static inline char comp_nt(char c) {
  switch (c) {
    case 'A': return 'T';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'T': return 'A';
    case 'N': return 'N';
    default: return '0';
  }
}

/* ------------------------------------------------------------------ table (kcount_cpu.cpp:206-295) */

typedef struct {
  uint16_t left[4], right[4]; /* ExtCounts A C G T */
  uint16_t count;
} vals_t;

typedef struct {
  uint64_t capacity, num_elems, num_dropped, num_supermers;
  uint64_t *keys; /* capacity * N words; empty iff last word == KEY_EMPTY */
  vals_t *vals;
} table_t;

static int is_prime(uint64_t n) {
  if (n < 4) return n > 1;
  if (n % 2 == 0) return 0;
  for (uint64_t d = 3; d * d <= n; d += 2)
    if (n % d == 0) return 0;
  return 1;
}
/* KmerMapExts::reserve, src/kcount/kcount_cpu.cpp:219-230. The reference picks the next entry of a
 * precomputed prime list (src/kcount/prime.hpp); the exact prime only matters when the table overflows,
 * and parity is only defined when nothing is dropped (kco_num_dropped() == 0), so any prime >= request do. */
static int table_reserve(table_t *t, uint64_t max_elems, int N) {
  uint64_t cap = max_elems < 11 ? 11 : max_elems;
  while (!is_prime(cap)) cap++;
  t->capacity = cap;
  t->num_elems = t->num_dropped = t->num_supermers = 0;
  t->keys = (uint64_t *)malloc(cap * N * 8);
  t->vals = (vals_t *)calloc(cap, sizeof(vals_t));
  if (!t->keys || !t->vals) return -1;
  memset(t->keys, 0xff, cap * N * 8);
  return 0;
}

/* KmerMapExts::insert(kmer, false), src/kcount/kcount_cpu.cpp:232-271: Murmur3 % capacity, linear probe <= 100 */
static vals_t *table_insert(table_t *t, const uint64_t *kmer, int N) {
  uint64_t slot = kco_murmur64(kmer, N * 8) % t->capacity;
  uint64_t max_probe = t->capacity < HT_MAX_PROBE ? t->capacity : HT_MAX_PROBE;
  for (uint64_t i = 1; i <= max_probe; i++) {
    uint64_t *k = t->keys + slot * N;
    if (k[N - 1] == KEY_EMPTY) {
      memcpy(k, kmer, N * 8);
      t->num_elems++;
      return &t->vals[slot];
    }
    if (memcmp(k, kmer, N * 8) == 0) return &t->vals[slot];
    slot = (slot + 1) % t->capacity;
  }
  t->num_dropped++;
  return NULL;
}

/* ------------------------------------------------------------------ context ---------------------- */

typedef struct {
  uint64_t key[MAXW];
  uint16_t count, ext8[8];
  char left, right;
  int rank;
} result_t;

struct kco_ctx {
  int k, m, N, qual_offset, qual_cutoff, dmin_thres, n_ranks;
  double dyn_min_depth;
  table_t *tables;
  uint64_t counted_kmers;
  result_t *results; /* all entries with count >= 2, sorted by key */
  uint64_t n_results;
  double avg_kmer_count;
  int finalized;
};

kco_ctx *kco_create(int k, int m, int qual_offset, int qual_cutoff, int dmin_thres, double dyn_min_depth, int n_ranks,
                    uint64_t slots_per_rank) {
  if (k < 2 || k > 127 || k % 32 == 0 || n_ranks < 1) return NULL;
  kco_ctx *c = (kco_ctx *)calloc(1, sizeof(kco_ctx));
  c->k = k;
  /* KmerDHT::KmerDHT, src/kcount/kmer_dht.cpp:117-119 */
  if (m <= 0) {
    m = k * 2 / 3 + 1;
    if (m < 15) m = 15;
    if (m > 27) m = 27;
  }
  if (m > k) m = k;
  c->m = m;
  c->N = n_words_of(k);
  c->qual_offset = qual_offset;
  c->qual_cutoff = qual_cutoff;
  c->dmin_thres = dmin_thres;
  c->dyn_min_depth = dyn_min_depth;
  c->n_ranks = n_ranks;
  c->tables = (table_t *)calloc(n_ranks, sizeof(table_t));
  for (int r = 0; r < n_ranks; r++)
    if (table_reserve(&c->tables[r], slots_per_rank, c->N) != 0) {
      kco_destroy(c);
      return NULL;
    }
  return c;
}

void kco_destroy(kco_ctx *c) {
  if (!c) return;
  if (c->tables)
    for (int r = 0; r < c->n_ranks; r++) {
      free(c->tables[r].keys);
      free(c->tables[r].vals);
    }
  free(c->tables);
  free(c->results);
  free(c);
}

int kco_minimizer_len(const kco_ctx *c) { return c->m; }
int kco_key_words(const kco_ctx *c) { return c->N; }

/* ------------------------------------------------------------------ owner side ------------------- */

/* HashTableInserter::insert_supermer -> insert_supermer_from_read -> get_kmers_and_exts,
 * src/kcount/kcount_cpu.cpp:478-493, 339-355, 309-336 */
int kco_insert_supermer(kco_ctx *c, int rank, const char *seq_in, int len, int sm_count) {
  const int k = c->k, N = c->N;
  table_t *t = &c->tables[rank];
  t->num_supermers++;
  char seq[65536 + 8];
  unsigned char quals[65536 + 8];
  if (len > 65536) return -1;
  for (int i = 0; i < len; i++) {
    char ch = seq_in[i];
    quals[i] = isupper((unsigned char)ch) ? 1 : 0; /* :313 */
    if (ch >= 'a' && ch <= 'z') ch += 'A' - 'a';
    if (ch != 'A' && ch != 'C' && ch != 'G' && ch != 'T' && ch != 'N') return -2; /* DIE at :485-487 */
    seq[i] = ch;
  }
  for (int i = 1; i < len - k; i++) { /* :320 */
    uint64_t kmer[MAXW], rc[MAXW];
    kco_pack_kmer(seq + i, k, kmer);
    char left = quals[i - 1] ? seq[i - 1] : '0';
    char right = quals[i + k] ? seq[i + k] : '0';
    kco_revcomp(kmer, k, rc);
    const uint64_t *key = kmer;
    if (kco_kmer_less(rc, kmer, N)) { /* :328-333 */
      key = rc;
      char tmp = left;
      left = comp_nt(right);
      right = comp_nt(tmp);
    }
    vals_t *v = table_insert(t, key, N);
    if (!v) continue; /* dropped, :346 */
    int cnt = v->count + sm_count;
    v->count = cnt > 65535 ? 65535 : (uint16_t)cnt; /* :349-351 */
    /* ExtCounts::inc with inc_with_limit, :147-164: only A C G T count, saturating */
    static const char L[4] = {'A', 'C', 'G', 'T'};
    for (int b = 0; b < 4; b++) {
      if (left == L[b]) { int x = v->left[b] + sm_count; v->left[b] = x > 65535 ? 65535 : (uint16_t)x; }
      if (right == L[b]) { int x = v->right[b] + sm_count; v->right[b] = x > 65535 ? 65535 : (uint16_t)x; }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ reader side ------------------ */

typedef void (*supermer_cb)(void *user, int target, const char *seq, int start, int len);

/* count_kmers() per-read body (src/kcount/kcount.cpp:71-90) + SeqBlockInserter::process_seq
 * (src/kcount/kcount_cpu.cpp:74-103). `seq` is already case-masked. Calls cb once per supermer with
 * [start, start+len) = its span inside the read. */
static void process_seq(const kco_ctx *c, const char *seq, int len, supermer_cb cb, void *user) {
  const int k = c->k, N = c->N;
  if (len < k + 2) return; /* fewer than one interior k-mer: the loops at :86 and :98 emit nothing */
  int nk = len - k + 1;
  int *targets = (int *)malloc(sizeof(int) * nk);
  for (int i = 1; i < len - k; i++) { /* targets of k-mers 0 and nk-1 are never used */
    uint64_t kmer[MAXW], rc[MAXW];
    kco_pack_kmer(seq + i, k, kmer); /* get_kmers upper-cases and maps N -> G, kmer.cpp:156-192 */
    kco_revcomp(kmer, k, rc);
    targets[i] = kco_owner(kco_kmer_less(rc, kmer, N) ? rc : kmer, k, c->m, c->n_ranks); /* :78-82, :87 */
  }
  int sm_start = 0, sm_len = k + 1; /* :84 seq.substr(0, k+1) */
  int prev = targets[1];           /* :85 */
  for (int i = 1; i < len - k; i++) {
    int t = targets[i];
    if (t == prev) {
      sm_len++; /* :89 append seq[i+k] */
    } else {
      cb(user, prev, seq, sm_start, sm_len); /* :92 */
      sm_start = i - 1;                      /* :93 seq.substr(i-1, k+2) */
      sm_len = k + 2;
      prev = t;
    }
  }
  if (sm_len >= k + 2) cb(user, prev, seq, sm_start, sm_len); /* :98-101 */
  free(targets);
}

/* PackedRead::unpack (src/packed_reads.cpp:196-208) + the quality mask of count_kmers (kcount.cpp:80-85) */
static int unpack_and_mask(const kco_ctx *c, const uint8_t *b, int len, char *seq) {
  static const char map[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'};
  for (int i = 0; i < len; i++) {
    char ch = map[b[i] & 7];
    int qual = c->qual_offset + (b[i] >> 3);
    if (qual < c->qual_offset + c->qual_cutoff) ch = (char)tolower(ch);
    seq[i] = ch;
  }
  return len;
}

static void cb_insert(void *user, int target, const char *seq, int start, int len) {
  kco_insert_supermer((kco_ctx *)user, target, seq + start, len, 1);
}

int kco_add_reads(kco_ctx *c, const uint8_t *bytes, const uint64_t *offsets, const uint16_t *lens, uint64_t n) {
  char *seq = (char *)malloc(65536 + 8);
  for (uint64_t r = 0; r < n; r++) {
    int len = lens[r];
    if (len < c->k) continue; /* kcount.cpp:78 */
    unpack_and_mask(c, bytes + offsets[r], len, seq);
    if (len >= c->k + 2) c->counted_kmers += len - c->k - 1;
    process_seq(c, seq, len, cb_insert, c);
  }
  free(seq);
  return 0;
}

typedef struct {
  int *targets, *first, *n, cap, cnt, k;
} sm_list;
static void cb_list(void *user, int target, const char *seq, int start, int len) {
  (void)seq;
  sm_list *l = (sm_list *)user;
  if (l->cnt < l->cap) {
    l->targets[l->cnt] = target;
    l->first[l->cnt] = start + 1;     /* first counted k-mer of a supermer starts one base in */
    l->n[l->cnt] = len - l->k - 1;    /* a supermer of len bases carries len-k-1 counted k-mers */
  }
  l->cnt++;
}
int kco_read_supermers(const kco_ctx *c, const uint8_t *bytes, int len, int *targets, int *first_kmer, int *n_kmers, int cap) {
  if (len < c->k) return 0;
  char *seq = (char *)malloc(len + 8);
  unpack_and_mask(c, bytes, len, seq);
  sm_list l = {targets, first_kmer, n_kmers, cap, 0, c->k};
  process_seq(c, seq, len, cb_list, &l);
  free(seq);
  return l.cnt;
}

/* ------------------------------------------------------------------ finalize --------------------- */

static int g_cmp_words;
static int cmp_result(const void *a, const void *b) {
  const result_t *x = (const result_t *)a, *y = (const result_t *)b;
  for (int i = 0; i < g_cmp_words; i++) {
    if (x->key[i] < y->key[i]) return -1;
    if (x->key[i] > y->key[i]) return 1;
  }
  return 0;
}

/* HashTableInserter::insert_into_local_hashtable, src/kcount/kcount_cpu.cpp:524-601 */
int kco_finalize(kco_ctx *c) {
  uint64_t n_ge2 = 0;
  for (int r = 0; r < c->n_ranks; r++)
    for (uint64_t s = 0; s < c->tables[r].capacity; s++)
      if (c->tables[r].keys[s * c->N + c->N - 1] != KEY_EMPTY && c->tables[r].vals[s].count >= 2) n_ge2++;
  free(c->results);
  c->results = (result_t *)calloc(n_ge2 ? n_ge2 : 1, sizeof(result_t));
  uint64_t at = 0, n_final = 0, sum = 0;
  for (int r = 0; r < c->n_ranks; r++) {
    table_t *t = &c->tables[r];
    for (uint64_t s = 0; s < t->capacity; s++) {
      const uint64_t *k = t->keys + s * c->N;
      if (k[c->N - 1] == KEY_EMPTY) continue;
      vals_t *v = &t->vals[s];
      if (v->count < 2) continue; /* :557 */
      result_t *e = &c->results[at++];
      memcpy(e->key, k, c->N * 8);
      e->count = v->count;
      memcpy(e->ext8, v->left, 8);
      memcpy(e->ext8 + 4, v->right, 8);
      e->left = kco_get_ext(v->left, v->count, c->dmin_thres, c->dyn_min_depth);   /* :565-568 */
      e->right = kco_get_ext(v->right, v->count, c->dmin_thres, c->dyn_min_depth);
      e->rank = r;
      if (e->left != 'X' && e->left != 'F' && e->right != 'X' && e->right != 'F') { /* :569-573 */
        n_final++;
        sum += v->count;
      }
    }
  }
  c->n_results = at;
  g_cmp_words = c->N;
  qsort(c->results, at, sizeof(result_t), cmp_result);
  c->avg_kmer_count = n_final ? (double)sum / (double)n_final : 0.0; /* :594-596 */
  c->finalized = 1;
  return 0;
}

static inline int is_final(const result_t *e) { return e->left != 'X' && e->left != 'F' && e->right != 'X' && e->right != 'F'; }

uint64_t kco_num_supermers(const kco_ctx *c, int rank) { return c->tables[rank].num_supermers; }
uint64_t kco_num_entries(const kco_ctx *c, int rank) {
  uint64_t n = 0;
  for (int r = 0; r < c->n_ranks; r++)
    if (rank < 0 || rank == r) n += c->tables[r].num_elems;
  return n;
}
uint64_t kco_num_dropped(const kco_ctx *c) {
  uint64_t n = 0;
  for (int r = 0; r < c->n_ranks; r++) n += c->tables[r].num_dropped;
  return n;
}
uint64_t kco_counted_kmers(const kco_ctx *c) { return c->counted_kmers; }
double kco_avg_kmer_count(const kco_ctx *c) { return c->avg_kmer_count; }

uint64_t kco_result_size(const kco_ctx *c, int extended, int rank) {
  uint64_t n = 0;
  for (uint64_t i = 0; i < c->n_results; i++) {
    const result_t *e = &c->results[i];
    if (rank >= 0 && e->rank != rank) continue;
    if (extended || is_final(e)) n++;
  }
  return n;
}

int kco_result_copy(const kco_ctx *c, int extended, int rank, uint64_t *keys, uint16_t *counts, char *left, char *right,
                    uint16_t *ext8) {
  uint64_t n = 0;
  for (uint64_t i = 0; i < c->n_results; i++) {
    const result_t *e = &c->results[i];
    if (rank >= 0 && e->rank != rank) continue;
    if (!(extended || is_final(e))) continue;
    if (keys) memcpy(keys + n * c->N, e->key, c->N * 8);
    if (counts) counts[n] = e->count;
    if (left) left[n] = e->left;
    if (right) right[n] = e->right;
    if (ext8) memcpy(ext8 + n * 8, e->ext8, 16);
    n++;
  }
  return 0;
}

/* order-independent checksum shared with the product (kc_result_checksum): sum of per-entry mixes */
uint64_t kco_entry_mix(const uint64_t *key, int n_words, uint16_t count, char left, char right, const uint16_t *ext8) {
  uint64_t h = 0x9e3779b97f4a7c15ULL;
  for (int i = 0; i < n_words; i++) {
    h = (h ^ key[i]) * 0xff51afd7ed558ccdULL;
    h ^= h >> 32;
  }
  uint64_t v = (uint64_t)count | ((uint64_t)(uint8_t)left << 16) | ((uint64_t)(uint8_t)right << 24);
  h = (h ^ v) * 0xc4ceb9fe1a85ec53ULL;
  h ^= h >> 29;
  if (ext8) {
    uint64_t e0 = (uint64_t)ext8[0] | ((uint64_t)ext8[1] << 16) | ((uint64_t)ext8[2] << 32) | ((uint64_t)ext8[3] << 48);
    uint64_t e1 = (uint64_t)ext8[4] | ((uint64_t)ext8[5] << 16) | ((uint64_t)ext8[6] << 32) | ((uint64_t)ext8[7] << 48);
    h = (h ^ e0) * 0xff51afd7ed558ccdULL;
    h ^= h >> 31;
    h = (h ^ e1) * 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 30;
  }
  return h;
}

uint64_t kco_result_checksum(const kco_ctx *c, int extended, int rank) {
  uint64_t sum = 0;
  for (uint64_t i = 0; i < c->n_results; i++) {
    const result_t *e = &c->results[i];
    if (rank >= 0 && e->rank != rank) continue;
    if (!(extended || is_final(e))) continue;
    sum += kco_entry_mix(e->key, c->N, e->count, e->left, e->right, extended ? e->ext8 : NULL);
  }
  return sum;
}
