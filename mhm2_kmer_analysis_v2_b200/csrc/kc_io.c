/* kc_io.c -- host-only input side of the stage (part of libkc_synth.so, no CUDA): FASTQ -> the reference's PackedRead byte
 * encoding. Replaces, for callers that start from files, FastqReader::get_next_fq_record (src/fastq.cpp:1028-1120: four
 * lines per record -- '@' name, sequence, '+', qualities of the same length) followed by the PackedRead constructor
 * (src/packed_reads.cpp:99-126: byte = base(A0 C1 G2 T3, everything else 4) | min(qual - offset, 31) << 3). Plain or gzip
 * input (zlib reads both). kc_fastq_read keeps mates as two independent reads; kc_fastq_read_pairs delivers them as the
 * pair loop of merge_reads() sees them (src/merge_reads.cpp:340-466: FASTQ characters, names normalised, a missing mate
 * replaced by "N") for the merge kernel (kc_merge.cu). NOT here: adapter trimming (src/adapters.cpp). */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
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

int kc_fastq_read_pairs(const char *path1, const char *path2, int paired, int qual_offset, uint64_t max_pairs, kc_pair_set *out) {
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
