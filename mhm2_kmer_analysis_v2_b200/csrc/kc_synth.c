/*
 * kc_synth.c -- seeded synthetic read streams in the reference's PackedRead byte encoding
 * (base(A0 C1 G2 T3 N4) | min(qual - offset, 31) << 3, src/packed_reads.cpp:99-126).
 *
 * Host-only helper of the C ABI (kc_synth_*, kc_reads_save in include/kc_synth.h; built as libkc_synth.so). It produces the
 * inputs BASELINE.md section 4 describes: random genomes (one, or a lognormal-abundance community),
 * uniform read starts, random strand, substitution errors, Phred 40 ('I') qualities with a low-quality
 * Phred 10 ('+') on 2 % of correct and 50 % of substituted bases; optional pairs and N bases.
 * PRNG: xoshiro256** seeded through splitmix64; read i depends only on (seed, i), so generation is
 * thread-count independent.
 */
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "kc_synth.h"

typedef struct {
  uint64_t s[4];
} rng_t;

static inline uint64_t splitmix64(uint64_t *x) {
  uint64_t z = (*x += 0x9e3779b97f4a7c15ULL);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static inline uint64_t rng_next(rng_t *r) {
  uint64_t *s = r->s;
  uint64_t result = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
  s[2] ^= s[0];
  s[3] ^= s[1];
  s[1] ^= s[2];
  s[0] ^= s[3];
  s[2] ^= t;
  s[3] = rotl(s[3], 45);
  return result;
}
static inline void rng_seed(rng_t *r, uint64_t seed, uint64_t stream) {
  uint64_t x = seed ^ (stream * 0xd1342543de82ef95ULL + 0x2545f4914f6cdd1dULL);
  for (int i = 0; i < 4; i++) r->s[i] = splitmix64(&x);
}
static inline double rng_unit(rng_t *r) { return (double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint64_t rng_below(rng_t *r, uint64_t n) { /* n > 0; multiply-shift, bias < 2^-64 * n */
  return (uint64_t)(((__uint128_t)rng_next(r) * n) >> 64);
}
static double rng_normal(rng_t *r) {
  double u1 = rng_unit(r), u2 = rng_unit(r);
  if (u1 < 1e-300) u1 = 1e-300;
  return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

struct kc_synth {
  kc_synth_config cfg;
  uint32_t n_genomes;
  uint64_t *glen, *gstart; /* per genome: length and start inside `bases` */
  double *cum;             /* cumulative read-sampling weight (coverage x length), normalised */
  uint8_t *bases;          /* all genomes, one base (0..3) per byte */
  uint64_t total_bases;
};

void kc_synth_config_default(kc_synth_config *c) {
  memset(c, 0, sizeof(*c));
  c->seed = 1001;
  c->n_genomes = 1;
  c->genome_len_min = c->genome_len_max = 5000000;
  c->abundance_sigma = 0.0;
  c->read_len = 150;
  c->read_len_min = 0;
  c->sub_rate = 0.005;
  c->lowq_rate = 0.02;
  c->lowq_rate_sub = 0.5;
  c->n_rate = 0.0;
  c->paired = 0;
  c->insert_mean = 300.0;
  c->insert_sd = 30.0;
  c->qual_offset = 33;
}

typedef struct {
  kc_synth *s;
  uint32_t g0, g1;
} genome_job;

static void *genome_worker(void *arg) {
  genome_job *j = (genome_job *)arg;
  for (uint32_t g = j->g0; g < j->g1; g++) {
    rng_t r;
    rng_seed(&r, j->s->cfg.seed, 0x47000000ULL + g);
    uint8_t *b = j->s->bases + j->s->gstart[g];
    uint64_t n = j->s->glen[g], i = 0;
    while (i < n) {
      uint64_t x = rng_next(&r);
      for (int t = 0; t < 32 && i < n; t++, i++, x >>= 2) b[i] = (uint8_t)(x & 3);
    }
  }
  return NULL;
}

int kc_synth_create(const kc_synth_config *cfg, kc_synth **out) {
  if (!cfg || !out || cfg->n_genomes == 0 || cfg->read_len == 0 || cfg->read_len > 65535 ||
      cfg->genome_len_min > cfg->genome_len_max || cfg->genome_len_min < (uint64_t)cfg->read_len * 4)
    return KC_SYNTH_ERR_ARG;
  kc_synth *s = (kc_synth *)calloc(1, sizeof(kc_synth));
  if (!s) return KC_SYNTH_ERR_NOMEM;
  s->cfg = *cfg;
  uint32_t G = s->n_genomes = cfg->n_genomes;
  s->glen = (uint64_t *)malloc(G * sizeof(uint64_t));
  s->gstart = (uint64_t *)malloc(G * sizeof(uint64_t));
  s->cum = (double *)malloc(G * sizeof(double));
  rng_t r;
  rng_seed(&r, cfg->seed, 0x434f4d4dULL);
  double tot = 0;
  uint64_t at = 0;
  for (uint32_t g = 0; g < G; g++) {
    s->glen[g] = cfg->genome_len_min + rng_below(&r, cfg->genome_len_max - cfg->genome_len_min + 1);
    double cov = cfg->abundance_sigma > 0 ? exp(cfg->abundance_sigma * rng_normal(&r)) : 1.0;
    s->gstart[g] = at;
    at += s->glen[g];
    tot += cov * (double)s->glen[g];
    s->cum[g] = tot;
  }
  for (uint32_t g = 0; g < G; g++) s->cum[g] /= tot;
  s->cum[G - 1] = 1.0;
  s->total_bases = at;
  s->bases = (uint8_t *)malloc(at);
  if (!s->bases) {
    kc_synth_destroy(s);
    return KC_SYNTH_ERR_NOMEM;
  }
  int nt = G < 8 ? (int)G : 8;
  pthread_t th[8];
  genome_job jobs[8];
  for (int t = 0; t < nt; t++) {
    jobs[t].s = s;
    jobs[t].g0 = (uint32_t)((uint64_t)G * t / nt);
    jobs[t].g1 = (uint32_t)((uint64_t)G * (t + 1) / nt);
    pthread_create(&th[t], NULL, genome_worker, &jobs[t]);
  }
  for (int t = 0; t < nt; t++) pthread_join(th[t], NULL);
  *out = s;
  return KC_SYNTH_OK;
}

void kc_synth_destroy(kc_synth *s) {
  if (!s) return;
  free(s->glen);
  free(s->gstart);
  free(s->cum);
  free(s->bases);
  free(s);
}

static inline uint16_t read_len_of(const kc_synth *s, uint64_t i) {
  const kc_synth_config *c = &s->cfg;
  if (c->read_len_min == 0 || c->read_len_min >= c->read_len) return (uint16_t)c->read_len;
  rng_t r;
  rng_seed(&r, c->seed, 0x4c000000000ULL + i);
  return (uint16_t)(c->read_len_min + rng_below(&r, c->read_len - c->read_len_min + 1));
}

static void make_read(const kc_synth *s, uint64_t i, uint8_t *out, uint16_t L) {
  const kc_synth_config *c = &s->cfg;
  /* fragment-level choices come from the fragment's stream so that both mates agree */
  uint64_t frag = c->paired ? i / 2 : i;
  rng_t fr;
  rng_seed(&fr, c->seed, 0x46000000000000ULL + frag);
  double u = rng_unit(&fr);
  uint32_t lo = 0, hi = s->n_genomes - 1;
  while (lo < hi) {
    uint32_t mid = (lo + hi) / 2;
    if (s->cum[mid] > u) hi = mid; else lo = mid + 1;
  }
  const uint8_t *g = s->bases + s->gstart[lo];
  uint64_t glen = s->glen[lo];
  int rev = (int)(rng_next(&fr) >> 63);
  uint64_t start;
  if (!c->paired) {
    start = rng_below(&fr, glen - L + 1);
  } else {
    double ins = c->insert_mean + c->insert_sd * rng_normal(&fr);
    uint64_t flen = ins < (double)c->read_len ? c->read_len : (uint64_t)(ins + 0.5);
    if (flen > glen) flen = glen;
    if (flen < L) flen = L;
    uint64_t fstart = rng_below(&fr, glen - flen + 1);
    int mate = (int)(i & 1);
    /* mate 0 reads the fragment's forward strand from its left end, mate 1 the reverse strand from
       its right end; a reversed fragment swaps the two */
    int from_right = mate ^ rev;
    start = from_right ? fstart + flen - L : fstart;
    rev = from_right;
  }
  rng_t r;
  rng_seed(&r, c->seed, 0x52000000000000ULL + i);
  const double two32 = 4294967296.0;
  uint32_t sub_thr = (uint32_t)fmin(c->sub_rate * two32, 4294967295.0);
  uint32_t lq_thr = (uint32_t)fmin(c->lowq_rate * two32, 4294967295.0);
  uint32_t lqs_thr = (uint32_t)fmin(c->lowq_rate_sub * two32, 4294967295.0);
  uint32_t n_thr = (uint32_t)fmin(c->n_rate * two32, 4294967295.0);
  const uint8_t q_hi = 31 /* min('I' - 33, 31) */, q_lo = 10 /* '+' - 33 */;
  for (uint16_t p = 0; p < L; p++) {
    uint8_t b = rev ? (uint8_t)(3 - g[start + L - 1 - p]) : g[start + p];
    uint64_t x = rng_next(&r);
    uint32_t a = (uint32_t)x, q = (uint32_t)(x >> 32);
    int sub = a < sub_thr;
    if (sub) b = (uint8_t)((b + 1 + (rng_next(&r) % 3)) & 3);
    uint8_t qual = (q < (sub ? lqs_thr : lq_thr)) ? q_lo : q_hi;
    if (n_thr && (uint32_t)(rng_next(&r) >> 32) < n_thr) b = 4;
    out[p] = (uint8_t)(b | (qual << 3));
  }
}

typedef struct {
  const kc_synth *s;
  uint64_t first, lo, hi;
  uint8_t *bytes;
  const uint64_t *offsets;
  const uint16_t *lens;
} read_job;

static void *read_worker(void *arg) {
  read_job *j = (read_job *)arg;
  for (uint64_t t = j->lo; t < j->hi; t++) make_read(j->s, j->first + t, j->bytes + j->offsets[t], j->lens[t]);
  return NULL;
}

int kc_synth_generate(const kc_synth *s, uint64_t first, uint64_t n, uint8_t *bytes, uint64_t *offsets, uint16_t *lens,
                      uint64_t *n_bytes, int n_threads) {
  if (!s || !bytes || !offsets || !lens) return KC_SYNTH_ERR_ARG;
  uint64_t at = 0;
  for (uint64_t t = 0; t < n; t++) {
    lens[t] = read_len_of(s, first + t);
    offsets[t] = at;
    at += lens[t];
  }
  if (n_bytes) *n_bytes = at;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  if ((uint64_t)n_threads > n) n_threads = n ? (int)n : 1;
  pthread_t th[64];
  read_job jobs[64];
  for (int t = 0; t < n_threads; t++) {
    jobs[t] = (read_job){s, first, n * t / n_threads, n * (t + 1) / n_threads, bytes, offsets, lens};
    pthread_create(&th[t], NULL, read_worker, &jobs[t]);
  }
  for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  return KC_SYNTH_OK;
}

int kc_reads_save(const char *path, const uint8_t *bytes, uint64_t n_bytes, const uint64_t *offsets, const uint16_t *lens,
                  uint64_t n_reads) {
  FILE *f = fopen(path, "wb");
  if (!f) return KC_SYNTH_ERR_IO;
  uint64_t hdr[4] = {0x3153444145524b43ULL /* "KCREADS1" */, n_reads, n_bytes, 0};
  uint64_t pad = 0, lens_bytes = n_reads * 2, lens_pad = ((lens_bytes + 7) & ~7ull) - lens_bytes;
  int ok = fwrite(hdr, 8, 4, f) == 4 && fwrite(offsets, 8, n_reads, f) == n_reads && fwrite(lens, 2, n_reads, f) == n_reads &&
           fwrite(&pad, 1, lens_pad, f) == lens_pad && fwrite(bytes, 1, n_bytes, f) == n_bytes;
  fclose(f);
  return ok ? KC_SYNTH_OK : KC_SYNTH_ERR_IO;
}

/* Lays a read set out the way the reference's PackedReads holds it in memory (src/packed_reads.hpp:68-79, 136): 16-byte
 * PackedRead headers {is_allocated : 1, read_id : 47, read_len : 16 | unsigned char *bytes}, the bytes bump-allocated in
 * blocks of block_bytes (4 MB there; a read never straddles two blocks, src/linear_allocator_pool.cpp). The blocks are
 * placed in `arena` in DESCENDING address order, so consecutive blocks are never adjacent in memory. */
int kc_reads_to_packed_layout(const uint8_t *bytes, const uint64_t *offsets, const uint16_t *lens, uint64_t n_reads, uint8_t *arena,
                              uint64_t arena_bytes, uint64_t block_bytes, uint64_t *headers /* [2 * n_reads] */, uint64_t *blocks_used) {
  if (!bytes || !offsets || !lens || !arena || !headers || block_bytes == 0) return -1;
  const uint64_t n_blocks_total = arena_bytes / block_bytes;
  uint64_t blk = 0, used = block_bytes; /* force a first block */
  uint8_t *base = NULL;
  for (uint64_t i = 0; i < n_reads; i++) {
    const uint64_t len = lens[i];
    if (len > block_bytes) return -1;
    if (used + len > block_bytes) {
      if (blk == n_blocks_total) return -3;
      base = arena + (n_blocks_total - 1 - blk) * block_bytes;
      blk++;
      used = 0;
    }
    memcpy(base + used, bytes + offsets[i], len);
    const int64_t id = (i & 1) ? (int64_t)(i / 2 + 1) : -(int64_t)(i / 2 + 1); /* pair 1 negative, pair 2 positive */
    headers[2 * i] = 1ull | (((uint64_t)id & ((1ull << 47) - 1)) << 1) | (len << 48);
    headers[2 * i + 1] = (uint64_t)(uintptr_t)(base + used);
    used += len;
  }
  if (blocks_used) *blocks_used = blk;
  return 0;
}
