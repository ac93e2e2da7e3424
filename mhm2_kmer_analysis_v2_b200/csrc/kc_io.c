/* kc_io.c -- host-only input side of the stage (part of libkc_synth.so, no CUDA): FASTQ -> the reference's PackedRead byte
 * encoding. Replaces, for callers that start from files, FastqReader::get_next_fq_record (src/fastq.cpp:1028-1120: four
 * lines per record -- '@' name, sequence, '+', qualities of the same length) followed by the PackedRead constructor
 * (src/packed_reads.cpp:99-126: byte = base(A0 C1 G2 T3, everything else 4) | min(qual - offset, 31) << 3). Plain or gzip
 * input (zlib reads both). kc_fastq_read keeps mates as two independent reads; kc_fastq_read_pairs delivers them as the
 * pair loop of merge_reads() sees them (src/merge_reads.cpp:340-466: FASTQ characters, names normalised, a missing mate
 * replaced by "N") for the merge kernel (kc_merge.cu). NOT here: adapter trimming (src/adapters.cpp). */
#include <fcntl.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include "kc_synth.h"

static int grow(void **p, uint64_t *cap, uint64_t need, size_t elem) {
  if (need <= *cap) return 0;
  uint64_t c = *cap ? *cap : 1024;
  while (c < need) c *= 2;
  void *q = realloc(*p, c * elem);
  if (!q) return -1;
  *p = q;
  *cap = c;
  return 0;
}

/* one line without its end-of-line characters; returns length, -1 at end of file, -2 when longer than cap */
static long read_line(gzFile f, char *buf, int cap) {
  if (!gzgets(f, buf, cap)) return -1;
  size_t n = strlen(buf);
  if (n == (size_t)cap - 1 && buf[n - 1] != '\n') return -2;
  while (n && (buf[n - 1] == '\n' || buf[n - 1] == '\r')) n--;
  buf[n] = 0;
  return (long)n;
}

int kc_fastq_read(const char *path, int qual_offset, uint64_t max_reads, kc_read_set *out) {
  if (!path || !out || (qual_offset != 33 && qual_offset != 64)) return KC_SYNTH_ERR_ARG; /* src/packed_reads.cpp:271 */
  memset(out, 0, sizeof *out);
  gzFile f = gzopen(path, "rb");
  if (!f) return KC_SYNTH_ERR_IO;
  gzbuffer(f, 1 << 20);
  const int cap = 1 << 17; /* reads are at most 65535 bases (PackedRead::read_len is 16 bits) */
  char *name = malloc(cap), *seq = malloc(cap), *plus = malloc(cap), *qual = malloc(cap);
  uint64_t cap_b = 0, cap_r = 0, cap_l = 0;
  int rc = KC_SYNTH_OK;
  if (!name || !seq || !plus || !qual) rc = KC_SYNTH_ERR_NOMEM;
  while (rc == KC_SYNTH_OK && (!max_reads || out->n_reads < max_reads)) {
    long ln = read_line(f, name, cap);
    if (ln == -1) break; /* clean end of file */
    if (ln == 0) continue; /* blank line between records */
    long ls = read_line(f, seq, cap), lp = read_line(f, plus, cap), lq = read_line(f, qual, cap);
    if (ln < 0 || ls < 0 || lp < 0 || lq < 0 || name[0] != '@' || plus[0] != '+' || ls != lq || ls > 65535) {
      rc = KC_SYNTH_ERR_FORMAT; /* src/fastq.cpp:1114-1118 */
      break;
    }
    if (grow((void **)&out->bytes, &cap_b, out->n_bytes + (uint64_t)ls + 16, 1) || grow((void **)&out->offsets, &cap_r, out->n_reads + 1, 8) ||
        grow((void **)&out->lens, &cap_l, out->n_reads + 1, 2)) {
      rc = KC_SYNTH_ERR_NOMEM;
      break;
    }
    uint8_t *dst = out->bytes + out->n_bytes;
    for (long i = 0; i < ls; i++) {
      uint8_t b;
      switch (seq[i] & ~0x20) {
        case 'A': b = 0; break;
        case 'C': b = 1; break;
        case 'G': b = 2; break;
        case 'T': b = 3; break;
        default: b = 4; break; /* N and the IUPAC ambiguity codes, src/packed_reads.cpp:105-117 */
      }
      int q = (int)(unsigned char)qual[i] - qual_offset;
      if (q < 0) q = 0;
      if (q > 31) q = 31;
      dst[i] = (uint8_t)(b | (q << 3));
    }
    out->offsets[out->n_reads] = out->n_bytes;
    out->lens[out->n_reads] = (uint16_t)ls;
    out->n_bytes += (uint64_t)ls;
    out->n_reads++;
  }
  gzclose(f);
  free(name), free(seq), free(plus), free(qual);
  if (rc != KC_SYNTH_OK) kc_read_set_free(out);
  return rc;
}

void kc_read_set_free(kc_read_set *rs) {
  if (!rs) return;
  free(rs->bytes), free(rs->offsets), free(rs->lens);
  memset(rs, 0, sizeof *rs);
}

/* ---- pairs for the merge kernel ------------------------------------------------------------------------------------------ */
typedef struct {
  char *id, *seq, *qual;
  long id_len, len;
  int cap;
} fq_rec;

static int rec_init(fq_rec *r, int cap) {
  r->cap = cap;
  r->id = malloc(cap), r->seq = malloc(cap), r->qual = malloc(cap);
  r->id_len = r->len = 0;
  return r->id && r->seq && r->qual ? 0 : -1;
}
static void rec_free(fq_rec *r) { free(r->id), free(r->seq), free(r->qual); }
static void rec_swap(fq_rec *a, fq_rec *b) {
  fq_rec t = *a;
  *a = *b;
  *b = t;
}
static void rtrim_len(char *s, long *n) {
  while (*n > 0 && (s[*n - 1] == ' ' || s[*n - 1] == '\t' || s[*n - 1] == '\r' || s[*n - 1] == '\n' || s[*n - 1] == '\v' || s[*n - 1] == '\f')) (*n)--;
  s[*n] = 0;
}

/* FastqReader::get_fq_name (src/fastq.cpp:77-135) without the comment-trimming mode: strips '@', turns "-R1" / " 1:Y:..:.."
 * into "/1", cuts a comment after "/1" or an unknown comment. */
static void fq_name(char *h, long *n) {
  memmove(h, h + 1, (size_t)*n); /* (includes the terminating 0) */
  (*n)--;
  rtrim_len(h, n);
  long len = *n;
  if (len >= 3 && h[len - 2] != '/') {
    if (h[len - 2] == 'R') {
      char r = h[len - 1];
      h[len - 3] = '/';
      h[len - 2] = r;
      *n = len - 1;
      h[*n] = 0;
      return;
    }
    char *e = strchr(h, '\t');
    if (!e) e = strchr(h, ' ');
    if (!e) return;
    long end = e - h;
    if (end > 3 && h[end - 2] == '/' && (h[end - 1] == '1' || h[end - 1] == '2')) {
      *n = end;
    } else if (len >= end + 7 && h[end + 2] == ':' && h[end + 4] == ':' && h[end + 6] == ':' && (h[end + 1] == '1' || h[end + 1] == '2')) {
      h[end] = '/';
      *n = end + 2;
    } else {
      *n = end;
    }
    h[*n] = 0;
  }
}

/* one FASTQ record (FastqReader::get_next_fq_record): 1 = got one, 0 = end of file, < 0 = error */
static int next_record(gzFile f, fq_rec *r, char *plus) {
  long ln;
  do {
    ln = read_line(f, r->id, r->cap);
    if (ln == -1) return 0;
  } while (ln == 0);
  long ls = read_line(f, r->seq, r->cap), lp = read_line(f, plus, r->cap), lq = read_line(f, r->qual, r->cap);
  if (ln < 0 || ls < 0 || lp < 0 || lq < 0) return KC_SYNTH_ERR_FORMAT;
  rtrim_len(r->id, &ln), rtrim_len(r->seq, &ls), rtrim_len(r->qual, &lq);
  if (r->id[0] != '@' || plus[0] != '+' || ls != lq || ls > 65535) return KC_SYNTH_ERR_FORMAT;
  fq_name(r->id, &ln);
  for (long i = 0; i < ln; i++)
    if (r->id[i] == ' ') r->id[i] = '_';
  r->id_len = ln;
  r->len = ls;
  return 1;
}

typedef struct {
  uint64_t seq, qual, off, len;
} pair_caps;
static int emit(kc_pair_set *out, pair_caps *c, const char *seq, const char *qual, long len) {
  if (grow((void **)&out->seq, &c->seq, out->n_bytes + (uint64_t)len + 16, 1) || grow((void **)&out->qual, &c->qual, out->n_bytes + (uint64_t)len + 16, 1) ||
      grow((void **)&out->offsets, &c->off, out->n_reads + 1, 8) || grow((void **)&out->lens, &c->len, out->n_reads + 1, 2))
    return -1;
  memcpy(out->seq + out->n_bytes, seq, (size_t)len);
  memcpy(out->qual + out->n_bytes, qual, (size_t)len);
  out->offsets[out->n_reads] = out->n_bytes;
  out->lens[out->n_reads] = (uint16_t)len;
  out->n_bytes += (uint64_t)len;
  out->n_reads++;
  return 0;
}

/* the streaming reader: one thread, plain or gzip (the parallel reader below must give exactly this) */
static int read_pairs_stream(const char *path1, const char *path2, int paired, int qual_offset, uint64_t max_pairs, kc_pair_set *out) {
  if (!path1 || !out || (qual_offset != 33 && qual_offset != 64) || (path2 && !paired)) return KC_SYNTH_ERR_ARG;
  memset(out, 0, sizeof *out);
  gzFile f1 = gzopen(path1, "rb"), f2 = path2 ? gzopen(path2, "rb") : NULL;
  if (!f1 || (path2 && !f2)) {
    if (f1) gzclose(f1);
    if (f2) gzclose(f2);
    return KC_SYNTH_ERR_IO;
  }
  gzbuffer(f1, 1 << 20);
  if (f2) gzbuffer(f2, 1 << 20);
  const int cap = 1 << 17;
  fq_rec r1, r2, tmp;
  char *plus = malloc(cap);
  int rc = (rec_init(&r1, cap) | rec_init(&r2, cap) | rec_init(&tmp, cap)) || !plus ? KC_SYNTH_ERR_NOMEM : KC_SYNTH_OK;
  const char fake_seq[2] = "N";
  char fake_qual[2] = {(char)qual_offset, 0};
  pair_caps caps = {0, 0, 0, 0};
  int turn = 0; /* two files: records come alternately from file 1 and file 2 (FastqReader with fqr2, src/fastq.cpp:1058-1070) */
#define NEXT(rec) next_record((f2 && (turn++ & 1)) ? f2 : f1, rec, plus)
  int skip_read1 = 0, skip_read2 = 0;
  tmp.id_len = 0;
  while (rc == KC_SYNTH_OK && (!max_pairs || out->n_reads < 2 * max_pairs)) {
    if (!paired) { /* src/merge_reads.cpp:352-371: an unpaired read gets the dummy mate a merged read gets */
      int g = NEXT(&r1);
      if (g < 0) rc = g;
      if (g <= 0) break;
      if (emit(out, &caps, r1.seq, r1.qual, r1.len) || emit(out, &caps, fake_seq, fake_qual, 1)) rc = KC_SYNTH_ERR_NOMEM;
      continue;
    }
    /* src/merge_reads.cpp:372-466, statement by statement */
    if (skip_read1 && skip_read2) break;
    if (!skip_read1) {
      int g = NEXT(&r1);
      if (g < 0) rc = g;
      if (g <= 0) break;
    } else {
      rec_swap(&r1, &tmp);
      tmp.id_len = 0;
      skip_read1 = 0;
    }
    r2.id_len = 0;
    if (r1.id_len > 2 && r1.id[r1.id_len - 1] == '1') {
      skip_read2 = 0;
    } else {
      /* a read 2 whose read 1 is missing: it becomes mate 2 of a faked "N" */
      out->missing_read1++;
      memcpy(r2.id, r1.id, (size_t)r1.id_len + 1);
      r2.id_len = r1.id_len;
      { char *t = r2.seq; r2.seq = r1.seq; r1.seq = t; t = r2.qual; r2.qual = r1.qual; r1.qual = t; long l = r2.len; r2.len = r1.len; r1.len = l; }
      if (r1.id_len > 0) r1.id[r1.id_len - 1] = '1';
      memcpy(r1.seq, fake_seq, 2), memcpy(r1.qual, fake_qual, 2);
      r1.len = 1;
      skip_read2 = 1;
    }
    if (!skip_read2) {
      int g = NEXT(&r2);
      if (g < 0) {
        rc = g;
        break;
      }
      if (g == 0) {
        r2.id_len = 0;
        skip_read1 = skip_read2 = 1;
      }
    } else {
      skip_read2 = 0;
    }
    if (r2.id_len > 2 && r2.id[r2.id_len - 1] == '2' && r1.id_len == r2.id_len && memcmp(r1.id, r2.id, (size_t)r1.id_len - 1) == 0) {
      skip_read1 = 0;
    } else {
      if (skip_read1 && (skip_read2 || r2.id_len == 0)) break; /* end of file */
      out->missing_read2++;
      rec_swap(&tmp, &r2); /* kept as the next read 1 (r2 now holds the emptied tmp) */
      if (r2.id_len == 0) skip_read2 = 1;
      memcpy(r2.id, r1.id, (size_t)r1.id_len + 1);
      r2.id_len = r1.id_len;
      if (r2.id_len > 0) r2.id[r2.id_len - 1] = '2';
      memcpy(r2.seq, fake_seq, 2), memcpy(r2.qual, fake_qual, 2);
      r2.len = 1;
      skip_read1 = 1;
    }
    if (emit(out, &caps, r1.seq, r1.qual, r1.len) || emit(out, &caps, r2.seq, r2.qual, r2.len)) rc = KC_SYNTH_ERR_NOMEM;
  }
#undef NEXT
  gzclose(f1);
  if (f2) gzclose(f2);
  rec_free(&r1), rec_free(&r2), rec_free(&tmp), free(plus);
  if (rc != KC_SYNTH_OK) kc_pair_set_free(out);
  return rc;
}

void kc_pair_set_free(kc_pair_set *ps) {
  if (!ps) return;
  free(ps->seq), free(ps->qual), free(ps->offsets), free(ps->lens);
  memset(ps, 0, sizeof *ps);
}

/* ---- parallel reader for plain (uncompressed) files --------------------------------------------------------------------
 * The streaming reader parses 0.8 M pairs/s on one core; the merge kernel takes 400 M pairs/s. For plain files the work is
 * split the way the reference splits it between ranks (FastqReader gives every rank a byte range and looks for the next record
 * start, src/fastq.cpp:154-330), but inside one process and with the streaming reader's exact result:
 *   1. T threads index the records whose first line starts in their byte range (line offsets, normalised name);
 *   2. ONE thread runs the pairing state machine of src/merge_reads.cpp:372-466 over the record index (no copying);
 *   3. T threads copy the sequence / quality characters of the emitted mates into the output arrays.
 * Anything unusual -- a malformed record, a thread that cannot find a record start, two neighbours that disagree on where a
 * record starts -- abandons the attempt and the streaming reader decides (so errors, too, are exactly its errors). */
#define KC_SYNTH_ERR_STATE_PAR (-6)
typedef struct {
  const char *id; /* normalised name (in the indexing thread's arena) */
  uint32_t id_len;
  char last; /* its last character (the state machine rewrites it on faked mates) */
  const char *seq, *qual;
  uint32_t len;
} prec;

typedef struct {
  const char *buf;
  uint64_t size, begin, end; /* byte range of this thread: records whose first line starts in [begin, end) */
  prec *recs;
  uint64_t n_recs, cap_recs;
  char *arena;
  uint64_t arena_used, arena_cap;
  uint64_t first_start, stop_at; /* where the first record of the range starts / where the next range's first record must start */
  int failed;
} index_job;

static inline int is_ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }

/* line [p, *eol) without its end-of-line characters; *next = start of the following line (size at the end of the buffer) */
static inline uint64_t line_at(const char *buf, uint64_t size, uint64_t p, uint64_t *next) {
  const char *nl = memchr(buf + p, '\n', size - p);
  uint64_t e = nl ? (uint64_t)(nl - buf) : size;
  *next = nl ? e + 1 : size;
  while (e > p && (buf[e - 1] == '\n' || buf[e - 1] == '\r')) e--;
  return e;
}

static void *index_thread(void *arg) {
  index_job *j = (index_job *)arg;
  const char *buf = j->buf;
  const uint64_t size = j->size;
  const uint64_t line_cap = (1u << 17) - 2; /* the streaming reader's line buffer */
  uint64_t p = j->begin;
  if (p > 0 && buf[p - 1] != '\n') { /* not at a line start: go to the next one */
    const char *nl = memchr(buf + p, '\n', size - p);
    p = nl ? (uint64_t)(nl - buf) + 1 : size;
  }
  /* first record start at or after p: a non-blank line that starts with '@' and whose second-next line starts with '+' (a
   * quality line may start with '@' too, but then the line two below is a sequence) */
  if (j->begin > 0) {
    for (;;) {
      if (p >= size) break;
      uint64_t n1, n2;
      uint64_t e0 = line_at(buf, size, p, &n1);
      if (e0 > p && buf[p] == '@' && n1 < size) {
        (void)line_at(buf, size, n1, &n2);
        if (n2 < size && buf[n2] == '+') break;
      }
      p = n1;
    }
  }
  j->first_start = p;
  while (p < size) {
    uint64_t n0, n1, n2, n3;
    uint64_t e0 = line_at(buf, size, p, &n0);
    if (e0 == p) { /* blank line between records */
      p = n0;
      continue;
    }
    if (p >= j->end) break; /* this record belongs to the next range */
    /* the other three lines must exist (a line exists when at least one byte is left: the streaming reader's gzgets) */
    if (n0 >= size) goto fail;
    uint64_t e1 = line_at(buf, size, n0, &n1);
    if (n1 >= size) goto fail;
    uint64_t e2 = line_at(buf, size, n1, &n2);
    (void)e2;
    if (n2 >= size) goto fail;
    uint64_t e3 = line_at(buf, size, n2, &n3);
    if (n0 - p > line_cap || n1 - n0 > line_cap || n2 - n1 > line_cap || n3 - n2 > line_cap) goto fail;
    /* trailing white space (rtrim_len) */
    uint64_t se = e1, qe = e3;
    while (se > n0 && is_ws(buf[se - 1])) se--;
    while (qe > n2 && is_ws(buf[qe - 1])) qe--;
    if (buf[p] != '@' || buf[n1] != '+' || se - n0 != qe - n2 || se - n0 > 65535) goto fail;
    /* the name, normalised as the streaming reader does */
    const uint64_t hl = e0 - p;
    if (j->arena_used + hl + 2 > j->arena_cap) {
      /* arenas are never reallocated (records point into them): a new one is chained by leaking the old pointer into recs */
      uint64_t cap = j->arena_cap ? j->arena_cap * 2 : (1u << 20);
      if (cap < hl + 2) cap = hl + 2;
      char *a = malloc(cap + sizeof(char *));
      if (!a) goto fail;
      *(char **)a = j->arena; /* chain for release */
      j->arena = a;
      j->arena_used = sizeof(char *);
      j->arena_cap = cap + sizeof(char *);
    }
    char *id = j->arena + j->arena_used;
    memcpy(id, buf + p, hl);
    id[hl] = 0;
    long n = (long)hl;
    rtrim_len(id, &n);
    if (id[0] != '@') goto fail;
    fq_name(id, &n);
    for (long i = 0; i < n; i++)
      if (id[i] == ' ') id[i] = '_';
    j->arena_used += hl + 2;
    if (j->n_recs == j->cap_recs) {
      uint64_t c = j->cap_recs ? j->cap_recs * 2 : 4096;
      prec *r = realloc(j->recs, c * sizeof(prec));
      if (!r) goto fail;
      j->recs = r;
      j->cap_recs = c;
    }
    prec *r = &j->recs[j->n_recs++];
    r->id = id;
    r->id_len = (uint32_t)n;
    r->last = n > 0 ? id[n - 1] : 0;
    r->seq = buf + n0;
    r->qual = buf + n2;
    r->len = (uint32_t)(se - n0);
    p = n3;
  }
  j->stop_at = p;
  return NULL;
fail:
  j->failed = 1;
  return NULL;
}

typedef struct {
  const char *buf;
  uint64_t size;
  int fd;
  index_job *jobs;
  int n_jobs;
  int cur_job; /* cursor of the state machine */
  uint64_t cur_rec;
} par_file;

static void par_file_close(par_file *f) {
  for (int t = 0; t < f->n_jobs; t++) {
    free(f->jobs[t].recs);
    char *a = f->jobs[t].arena;
    while (a) {
      char *prev = *(char **)a;
      free(a);
      a = prev;
    }
  }
  free(f->jobs);
  if (f->buf && f->size) munmap((void *)f->buf, f->size);
  if (f->fd >= 0) close(f->fd);
  memset(f, 0, sizeof *f);
  f->fd = -1;
}

/* maps and indexes one plain file with n_threads threads; 0 = indexed, 1 = not a candidate / something unusual (use the
 * streaming reader), < 0 = error */
static int par_file_open(par_file *f, const char *path, int n_threads, uint64_t min_bytes) {
  memset(f, 0, sizeof *f);
  f->fd = open(path, O_RDONLY);
  if (f->fd < 0) return KC_SYNTH_ERR_IO;
  struct stat st;
  if (fstat(f->fd, &st) != 0 || !S_ISREG(st.st_mode) || (uint64_t)st.st_size < min_bytes) return 1;
  f->size = (uint64_t)st.st_size;
  void *m = mmap(NULL, f->size, PROT_READ, MAP_PRIVATE, f->fd, 0);
  if (m == MAP_FAILED) {
    f->size = 0;
    return 1;
  }
  f->buf = (const char *)m;
  if (f->size >= 2 && (unsigned char)f->buf[0] == 0x1f && (unsigned char)f->buf[1] == 0x8b) return 1; /* gzip */
  uint64_t min_chunk = 1u << 20; /* at least 1 MB per thread (KC_FASTQ_PAR_MIN_CHUNK: tests cut small files into many ranges) */
  if (getenv("KC_FASTQ_PAR_MIN_CHUNK")) min_chunk = strtoull(getenv("KC_FASTQ_PAR_MIN_CHUNK"), NULL, 10);
  if (min_chunk < 1) min_chunk = 1;
  if ((uint64_t)n_threads > f->size / min_chunk) n_threads = f->size / min_chunk ? (int)(f->size / min_chunk) : 1;
  f->jobs = calloc((size_t)n_threads, sizeof(index_job));
  if (!f->jobs) return KC_SYNTH_ERR_NOMEM;
  f->n_jobs = n_threads;
  pthread_t *th = malloc(sizeof(pthread_t) * (size_t)n_threads);
  if (!th) return KC_SYNTH_ERR_NOMEM;
  for (int t = 0; t < n_threads; t++) {
    index_job *j = &f->jobs[t];
    j->buf = f->buf;
    j->size = f->size;
    j->begin = f->size * (uint64_t)t / (uint64_t)n_threads;
    j->end = f->size * (uint64_t)(t + 1) / (uint64_t)n_threads;
  }
  int started = 0;
  for (int t = 1; t < n_threads; t++) {
    if (pthread_create(&th[t], NULL, index_thread, &f->jobs[t]) != 0) break;
    started = t;
  }
  index_thread(&f->jobs[0]);
  for (int t = 1; t <= started; t++) pthread_join(th[t], NULL);
  free(th);
  if (started != n_threads - 1) return 1;
  /* the ranges must tile the file: every thread stopped where its neighbour found its first record */
  for (int t = 0; t < n_threads; t++) {
    if (f->jobs[t].failed) return 1;
    const uint64_t want = t + 1 < n_threads ? f->jobs[t + 1].first_start : f->size;
    if (f->jobs[t].stop_at != want) return 1;
  }
  return 0;
}

/* next record of the file in order; 0 at the end */
static int par_next(par_file *f, prec *r) {
  while (f->cur_job < f->n_jobs && f->cur_rec >= f->jobs[f->cur_job].n_recs) {
    f->cur_job++;
    f->cur_rec = 0;
  }
  if (f->cur_job >= f->n_jobs) return 0;
  *r = f->jobs[f->cur_job].recs[f->cur_rec++];
  return 1;
}

typedef struct {
  const char *seq, *qual;
  uint32_t len;
} mate_ref;
typedef struct {
  const mate_ref *m;
  const uint64_t *off;
  uint64_t begin, end;
  char *seq, *qual;
} copy_job;
static void *copy_thread(void *arg) {
  copy_job *c = (copy_job *)arg;
  for (uint64_t i = c->begin; i < c->end; i++) {
    memcpy(c->seq + c->off[i], c->m[i].seq, c->m[i].len);
    memcpy(c->qual + c->off[i], c->m[i].qual, c->m[i].len);
  }
  return NULL;
}

/* phase 2 + 3 over indexed files: the pairing logic of read_pairs_stream (src/merge_reads.cpp:340-466) on record descriptors */
static int read_pairs_indexed(par_file *f1, par_file *f2, int paired, int qual_offset, uint64_t max_pairs, int n_threads, kc_pair_set *out) {
  static const char fake_seq[2] = "N";
  char *fake_qual = malloc(2); /* lives until the copy is done */
  if (!fake_qual) return KC_SYNTH_ERR_NOMEM;
  fake_qual[0] = (char)qual_offset, fake_qual[1] = 0;
  mate_ref *m = NULL;
  uint64_t n_m = 0, cap_m = 0;
  int rc = KC_SYNTH_OK, turn = 0;
  prec r1, r2, tmp;
  memset(&r1, 0, sizeof r1), memset(&r2, 0, sizeof r2), memset(&tmp, 0, sizeof tmp);
#define PNEXT(rec) par_next((f2 && (turn++ & 1)) ? f2 : f1, rec)
#define PEMIT(R)                                                   \
  do {                                                             \
    if (n_m == cap_m) {                                            \
      uint64_t c_ = cap_m ? cap_m * 2 : 8192;                      \
      mate_ref *q_ = realloc(m, c_ * sizeof(mate_ref));            \
      if (!q_) {                                                   \
        rc = KC_SYNTH_ERR_NOMEM;                                   \
        break;                                                     \
      }                                                            \
      m = q_, cap_m = c_;                                          \
    }                                                              \
    m[n_m].seq = (R).seq, m[n_m].qual = (R).qual, m[n_m].len = (R).len; \
    n_m++;                                                         \
  } while (0)
  int skip_read1 = 0, skip_read2 = 0;
  while (rc == KC_SYNTH_OK && (!max_pairs || n_m < 2 * max_pairs)) {
    if (!paired) {
      if (!PNEXT(&r1)) break;
      PEMIT(r1);
      r2.seq = fake_seq, r2.qual = fake_qual, r2.len = 1;
      PEMIT(r2);
      continue;
    }
    if (skip_read1 && skip_read2) break;
    if (!skip_read1) {
      if (!PNEXT(&r1)) break;
    } else {
      r1 = tmp;
      tmp.id_len = 0;
      skip_read1 = 0;
    }
    r2.id_len = 0;
    if (r1.id_len > 2 && r1.last == '1') {
      skip_read2 = 0;
    } else { /* a read 2 whose read 1 is missing */
      out->missing_read1++;
      r2 = r1;
      if (r1.id_len > 0) r1.last = '1';
      r1.seq = fake_seq, r1.qual = fake_qual, r1.len = 1;
      skip_read2 = 1;
    }
    if (!skip_read2) {
      if (!PNEXT(&r2)) {
        r2.id_len = 0;
        skip_read1 = skip_read2 = 1;
      }
    } else {
      skip_read2 = 0;
    }
    if (r2.id_len > 2 && r2.last == '2' && r1.id_len == r2.id_len && memcmp(r1.id, r2.id, (size_t)r1.id_len - 1) == 0) {
      skip_read1 = 0;
    } else {
      if (skip_read1 && (skip_read2 || r2.id_len == 0)) break;
      out->missing_read2++;
      tmp = r2; /* kept as the next read 1 */
      skip_read2 = 1; /* the reference tests the name it has just swapped OUT (always empty): src/merge_reads.cpp:452-457 */
      r2.id = r1.id, r2.id_len = r1.id_len;
      if (r2.id_len > 0) r2.last = '2';
      r2.seq = fake_seq, r2.qual = fake_qual, r2.len = 1;
      skip_read1 = 1;
    }
    PEMIT(r1);
    if (rc == KC_SYNTH_OK) PEMIT(r2);
  }
#undef PNEXT
#undef PEMIT
  uint64_t *off = NULL;
  if (rc == KC_SYNTH_OK) {
    off = malloc((n_m + 1) * sizeof(uint64_t));
    out->lens = malloc((n_m + 1) * sizeof(uint16_t));
    if (!off || !out->lens) rc = KC_SYNTH_ERR_NOMEM;
  }
  if (rc == KC_SYNTH_OK) {
    uint64_t at = 0;
    for (uint64_t i = 0; i < n_m; i++) {
      off[i] = at;
      out->lens[i] = (uint16_t)m[i].len;
      at += m[i].len;
    }
    out->n_bytes = at;
    out->n_reads = n_m;
    out->seq = malloc(at + 16), out->qual = malloc(at + 16);
    if (!out->seq || !out->qual) rc = KC_SYNTH_ERR_NOMEM;
  }
  if (rc == KC_SYNTH_OK) {
    int T = n_threads < 1 ? 1 : n_threads;
    if ((uint64_t)T > n_m / 4096 + 1) T = (int)(n_m / 4096 + 1);
    copy_job *cj = calloc((size_t)T, sizeof(copy_job));
    pthread_t *th = malloc(sizeof(pthread_t) * (size_t)T);
    if (!cj || !th) {
      rc = KC_SYNTH_ERR_NOMEM;
    } else {
      for (int t = 0; t < T; t++) {
        cj[t].m = m, cj[t].off = off, cj[t].seq = out->seq, cj[t].qual = out->qual;
        cj[t].begin = n_m * (uint64_t)t / (uint64_t)T, cj[t].end = n_m * (uint64_t)(t + 1) / (uint64_t)T;
      }
      int started = 0;
      for (int t = 1; t < T; t++) {
        if (pthread_create(&th[t], NULL, copy_thread, &cj[t]) != 0) break;
        started = t;
      }
      copy_thread(&cj[0]);
      for (int t = 1; t <= started; t++) pthread_join(th[t], NULL);
      for (int t = started + 1; t < T; t++) copy_thread(&cj[t]); /* (threads that could not be started) */
    }
    free(cj), free(th);
  }
  if (rc == KC_SYNTH_OK) out->offsets = off;
  else free(off);
  free(m);
  free(fake_qual);
  if (rc != KC_SYNTH_OK) kc_pair_set_free(out);
  return rc;
}

static int host_threads(void) {
  int n = (int)sysconf(_SC_NPROCESSORS_ONLN);
  if (getenv("KC_HOST_THREADS")) n = atoi(getenv("KC_HOST_THREADS"));
  if (n < 1) n = 1;
  if (n > 32) n = 32;
  return n;
}

int kc_fastq_read_pairs(const char *path1, const char *path2, int paired, int qual_offset, uint64_t max_pairs, kc_pair_set *out) {
  if (!path1 || !out || (qual_offset != 33 && qual_offset != 64) || (path2 && !paired)) return KC_SYNTH_ERR_ARG;
  const int T = host_threads();
  uint64_t min_bytes = 4u << 20; /* smaller files are not worth the threads */
  if (getenv("KC_FASTQ_PAR_MIN_BYTES")) min_bytes = strtoull(getenv("KC_FASTQ_PAR_MIN_BYTES"), NULL, 10);
  if (T > 1 || getenv("KC_FASTQ_PAR_MIN_BYTES")) {
    par_file f1, f2;
    memset(&f2, 0, sizeof f2);
    f2.fd = -1;
    int ok = par_file_open(&f1, path1, T, min_bytes);
    if (ok == 0 && path2) ok = par_file_open(&f2, path2, T, min_bytes);
    int rc = 1;
    if (ok == 0) {
      memset(out, 0, sizeof *out);
      rc = read_pairs_indexed(&f1, path2 ? &f2 : NULL, paired, qual_offset, max_pairs, T, out);
    }
    par_file_close(&f1);
    par_file_close(&f2);
    if (ok == 0) return rc;
    if (getenv("KC_FASTQ_PAR_STRICT")) return KC_SYNTH_ERR_STATE_PAR; /* (tests: tell whether the parallel path was taken) */
    /* not a candidate (gzip, small, unusual): the streaming reader decides, errors included */
  }
  return read_pairs_stream(path1, path2, paired, qual_offset, max_pairs, out);
}
