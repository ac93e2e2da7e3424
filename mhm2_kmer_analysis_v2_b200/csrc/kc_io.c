/* kc_io.c -- host-only input side of the stage (part of libkc_synth.so, no CUDA): FASTQ -> the reference's PackedRead byte
 * encoding. Replaces, for callers that start from files, FastqReader::get_next_fq_record (src/fastq.cpp:1028-1120: four
 * lines per record -- '@' name, sequence, '+', qualities of the same length) followed by the PackedRead constructor
 * (src/packed_reads.cpp:99-126: byte = base(A0 C1 G2 T3, everything else 4) | min(qual - offset, 31) << 3). Plain or gzip
 * input (zlib reads both). NOT here: pair merging and adapter trimming (src/merge_reads.cpp:250-771) -- mates stay two
 * independent reads, which is what the k-mer analysis of this proxy sees for unmerged pairs anyway. */
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
