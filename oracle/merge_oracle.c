/* TEST INFRASTRUCTURE (oracle/) -- plain-C restatement of the pair loop of the reference's merge_reads()
 * (src/merge_reads.cpp:466-650) and of the PackedRead encoding of its output (src/packed_reads.cpp:99-126), one pair after the
 * other, position by position, in the reference's own order of statements. It is the checker that travels to the GPU box for
 * kc_merge.cu (tests/test_merge.py); nothing under mhm2_kmer_analysis_v2_b200/ includes, links or calls it.
 * Pinned: tests/test_merge.py holds it to the fixtures under tests/golden/merge, the outputs of the reference's own merge_reads.cpp
 * (oracle/_ref/merge_ref, made by tests/golden/make_merge_golden.py).
 * Deliberately written differently from the product (kc_merge_core.cuh works on 4-bit codes, ballots and events): this file
 * keeps strings and the reference's loop.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* src/merge_reads.cpp:74-84 */
static const double Q2Perror[] = {
    1.0,       0.7943,    0.6309,    0.5012,    0.3981,    0.3162,    0.2512,    0.1995,    0.1585,    0.1259,     0.1,
    0.07943,   0.06310,   0.05012,   0.03981,   0.03162,   0.02512,   0.01995,   0.01585,   0.01259,   0.01,       0.007943,
    0.006310,  0.005012,  0.003981,  0.003162,  0.002512,  0.001995,  0.001585,  0.001259,  0.001,     0.0007943,  0.0006310,
    0.0005012, 0.0003981, 0.0003162, 0.0002512, 0.0001995, 0.0001585, 0.0001259, 0.0001,    7.943e-05, 6.310e-05,  5.012e-05,
    3.981e-05, 3.162e-05, 2.512e-05, 1.995e-05, 1.585e-05, 1.259e-05, 1e-05,     7.943e-06, 6.310e-06, 5.012e-06,  3.981e-06,
    3.162e-06, 2.512e-06, 1.995e-06, 1.585e-06, 1.259e-06, 1e-06,     7.943e-07, 6.310e-07, 5.012e-07, 3.981e-07,  3.1622e-07,
    2.512e-07, 1.995e-07, 1.585e-07, 1.259e-07, 1e-07,     7.943e-08, 6.310e-08, 5.012e-08, 3.981e-08, 3.1622e-08, 2.512e-08,
    1.995e-08, 1.585e-08, 1.259e-08, 1e-08};
#define Q2P_SIZE ((int)(sizeof(Q2Perror) / sizeof(*Q2Perror)))

/* revcomp(), src/utils.cpp:98-130; 0 = the reference DIEs */
static char comp_char(char c) {
  switch (c) {
    case 'A': case 'a': return 'T';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    case 'T': case 't': return 'A';
    case 'N': case 'n': case 'U': case 'R': case 'Y': case 'K': case 'M': case 'S': case 'W': case 'B': case 'D': case 'H': case 'V': return 'N';
  }
  return 0;
}

/* PackedRead::PackedRead, src/packed_reads.cpp:99-126; -1 = the reference DIEs */
static int pack_read(const char *seq, const char *quals, int len, int qual_offset, uint8_t *out) {
  for (int i = 0; i < len; i++) {
    uint8_t b;
    switch (seq[i]) {
      case 'A': case 'a': b = 0; break;
      case 'C': case 'c': b = 1; break;
      case 'G': case 'g': b = 2; break;
      case 'T': case 't': b = 3; break;
      case 'N': case 'n': case 'U': case 'R': case 'Y': case 'K': case 'M': case 'S': case 'W': case 'B': case 'D': case 'H': case 'V': b = 4; break;
      default: return -1;
    }
    int q = (int)(unsigned char)quals[i] - qual_offset;
    if (q > 31) q = 31;
    out[i] = (uint8_t)(b | ((unsigned char)q << 3));
  }
  return 0;
}

/* fast_count_mismatches, src/merge_reads.cpp:186-247: the count when it is <= max, else something greater than max */
static int count_mismatches(const char *a, const char *b, int len, int max) {
  int mm = 0;
  for (int j = 0; j < len; j++) {
    if (a[j] != b[j]) mm++;
    if (mm > max) break;
  }
  return mm;
}

/* One pair, src/merge_reads.cpp:466-650. seq1 / quals1 have room for len1 + len2 characters. Returns 1 = merged (seq1 / quals1
 * hold the merged read, *len1 its length), 0 = not merged (quals1 may carry zeroed N qualities), -1 = the reference DIEs. */
static int merge_one(char *seq1, char *quals1, int *len1_io, const char *seq2, const char *quals2, int len2, int qual_offset, char *rc_seq2,
                     char *rev_quals2, uint64_t *num_ambiguous, int *overlap_out) {
  const int MIN_OVERLAP = 12, EXTRA_TEST_OVERLAP = 2, MAX_MISMATCHES = 3, EXTRA_MISMATCHES_PER_1000 = 150;
  const double MAX_PERROR = 0.025;
  const int MAX_MATCH_QUAL = 41 + qual_offset;
  const int len1 = *len1_io;
  int abort_merge = 0;
  for (int t = 0; t < len2; t++) { /* :475-478 */
    rc_seq2[t] = comp_char(seq2[len2 - 1 - t]);
    if (!rc_seq2[t]) return -1;
    rev_quals2[t] = quals2[len2 - 1 - t];
  }
  const int len = len2 < len1 ? len2 : len1; /* :480-486 */
  const int start_i = (len == len1) ? 0 : len1 - len;
  int found_i = -1, best_i = -1;
  for (int i = 0; i < len - MIN_OVERLAP + EXTRA_TEST_OVERLAP; i++) { /* :489 */
    if (abort_merge) break;
    const int overlap = len - i;
    const int this_max_mismatch = MAX_MISMATCHES + (EXTRA_MISMATCHES_PER_1000 * overlap / 1000);
    const int error_max_mismatch = this_max_mismatch * 4 / 3 + 1;
    if (count_mismatches(seq1 + start_i + i, rc_seq2, overlap, error_max_mismatch) > error_max_mismatch) continue;
    int matches = 0, mismatches = 0, bothNs = 0, Ncount = 0, overlapChecked = 0;
    double perror = 0.0;
    for (int j = 0; j < overlap; j++) { /* :499-545 */
      overlapChecked++;
      const char ps = seq1[start_i + i + j], rs = rc_seq2[j];
      if (ps == rs) {
        matches++;
        if (ps == 'N') {
          Ncount += 2;
          if (bothNs++) {
            abort_merge++;
            (*num_ambiguous)++;
            break;
          }
        }
      } else {
        mismatches++;
        if (ps == 'N') {
          mismatches++;
          Ncount++;
          quals1[start_i + i + j] = (char)qual_offset;
          const int x = (unsigned char)rev_quals2[j] - qual_offset;
          if (x < 0 || x >= Q2P_SIZE) return -1;
          perror += Q2Perror[x];
        } else if (rs == 'N') {
          Ncount++;
          mismatches++;
          rev_quals2[j] = (char)qual_offset;
          const int x = (unsigned char)quals1[start_i + i + j] - qual_offset;
          if (x < 0 || x >= Q2P_SIZE) return -1;
          perror += Q2Perror[x];
        }
        const int q1 = (unsigned char)quals1[start_i + i + j] - qual_offset, q2 = (unsigned char)rev_quals2[j] - qual_offset;
        if (q1 < 0 || q2 < 0 || q1 >= Q2P_SIZE || q2 >= Q2P_SIZE) return -1;
        const int diffq = q1 > q2 ? q1 - q2 : q2 - q1;
        if (diffq <= 2)
          perror += 0.5;
        else
          perror += Q2Perror[diffq];
      }
      if (Ncount > 3) {
        abort_merge++;
        (*num_ambiguous)++;
        break;
      }
      if (mismatches > error_max_mismatch) break;
    }
    int match_thres = overlap - this_max_mismatch; /* :546-578 */
    if (match_thres < MIN_OVERLAP) match_thres = MIN_OVERLAP;
    if (matches >= match_thres && overlapChecked == overlap && mismatches <= this_max_mismatch && perror / overlap <= MAX_PERROR) {
      if (best_i < 0 && found_i < 0) {
        best_i = i;
      } else {
        (*num_ambiguous)++;
        best_i = -1;
        break;
      }
    } else if (overlapChecked == overlap && mismatches <= error_max_mismatch && perror / overlap <= MAX_PERROR * 4 / 3) {
      found_i = i;
      if (best_i >= 0) {
        (*num_ambiguous)++;
        best_i = -1;
        break;
      }
    }
  }
  if (!(best_i >= 0 && !abort_merge)) return 0;
  const int i = best_i, overlap = len - i; /* :580-613 */
  for (int j = 0; j < overlap; j++) {
    const int p = start_i + i + j;
    if (seq1[p] == rc_seq2[j]) {
      const uint16_t newQual = (uint16_t)((unsigned char)quals1[p] + (unsigned char)rev_quals2[j] - qual_offset);
      quals1[p] = (char)((newQual > MAX_MATCH_QUAL) ? MAX_MATCH_QUAL : newQual);
    } else {
      uint8_t newQual;
      if ((unsigned char)quals1[p] < (unsigned char)rev_quals2[j]) {
        newQual = (uint8_t)((unsigned char)rev_quals2[j] - (unsigned char)quals1[p] + qual_offset);
        seq1[p] = rc_seq2[j];
      } else {
        newQual = (uint8_t)((unsigned char)quals1[p] - (unsigned char)rev_quals2[j] + qual_offset);
      }
      quals1[p] = (char)((newQual > (2 + qual_offset)) ? newQual : (2 + qual_offset));
    }
  }
  /* seq1.substr(0, start_i + i + overlap) + rc_seq2.substr(overlap): start_i + i + overlap == len1 */
  memcpy(seq1 + len1, rc_seq2 + overlap, (size_t)(len2 - overlap));
  memcpy(quals1 + len1, rev_quals2 + overlap, (size_t)(len2 - overlap));
  *len1_io = len1 + len2 - overlap;
  *overlap_out = overlap;
  return 1;
}

/* All pairs: reads 2 i, 2 i + 1 of (seq, qual, offsets, lens) are the mates of pair i (FASTQ characters). Output like
 * kc_merge_pairs: PackedRead bytes, out_offsets / out_lens per read written (2 per kept pair), *n_out reads; counters[0..3] =
 * merged, ambiguous, merged_len, overlap_len. Returns 0, or -1 where the reference would DIE. */
int kmo_merge_pairs(const char *seq, const char *qual, const uint64_t *offsets, const uint16_t *lens, uint64_t n_pairs, int qual_offset,
                    int min_kmer_len, int paired, uint8_t *out_bytes, uint64_t *out_offsets, uint16_t *out_lens, uint64_t *n_out,
                    uint64_t *counters) {
  char *s1 = malloc(2 * 65536 + 8), *q1 = malloc(2 * 65536 + 8), *rc2 = malloc(65536 + 8), *rq2 = malloc(65536 + 8);
  uint64_t at = 0, w = 0, merged = 0, ambiguous = 0, merged_len = 0, overlap_len = 0;
  int rc = 0;
  for (uint64_t pr = 0; pr < n_pairs && rc == 0; pr++) {
    int len1 = lens[2 * pr], len2 = lens[2 * pr + 1];
    const char *g2 = seq + offsets[2 * pr + 1], *gq2 = qual + offsets[2 * pr + 1];
    if (paired && len1 < min_kmer_len && len2 < min_kmer_len) continue; /* :472 */
    memcpy(s1, seq + offsets[2 * pr], (size_t)len1);
    memcpy(q1, qual + offsets[2 * pr], (size_t)len1);
    int overlap = 0;
    const int m = merge_one(s1, q1, &len1, g2, gq2, len2, qual_offset, rc2, rq2, &ambiguous, &overlap);
    if (m < 0) {
      rc = -1;
      break;
    }
    if (m == 1) { /* :617-631: the merged read and the dummy mate "N" */
      if (pack_read(s1, q1, len1, qual_offset, out_bytes + at)) rc = -1;
      out_bytes[at + len1] = 4;
      out_offsets[w] = at, out_lens[w] = (uint16_t)len1;
      out_offsets[w + 1] = at + (uint64_t)len1, out_lens[w + 1] = 1;
      at += (uint64_t)len1 + 1;
      merged++;
      merged_len += (uint64_t)len1;
      overlap_len += (uint64_t)overlap;
    } else { /* :632-641 */
      if (pack_read(s1, q1, len1, qual_offset, out_bytes + at) || pack_read(g2, gq2, len2, qual_offset, out_bytes + at + len1)) rc = -1;
      out_offsets[w] = at, out_lens[w] = (uint16_t)len1;
      out_offsets[w + 1] = at + (uint64_t)len1, out_lens[w + 1] = (uint16_t)len2;
      at += (uint64_t)len1 + (uint64_t)len2;
    }
    w += 2;
  }
  free(s1), free(q1), free(rc2), free(rq2);
  *n_out = w;
  if (counters) counters[0] = merged, counters[1] = ambiguous, counters[2] = merged_len, counters[3] = overlap_len;
  return rc;
}
