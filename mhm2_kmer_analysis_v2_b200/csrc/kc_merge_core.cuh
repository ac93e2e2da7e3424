// kc_merge_core.cuh -- the per-pair arithmetic of read-pair merging, as plain functions the merge kernel (kc_merge.cu) calls.
//
// Replaces the body of the pair loop of merge_reads() (src/merge_reads.cpp:466-650): mate 2 is reverse-complemented, slid along
// the tail of mate 1, every offset that survives a cheap mismatch count is scored base by base (matches, mismatches, Ns and an
// accumulated error probability from the quality differences), exactly one good offset and no ambiguous one means "merge",
// and the overlap is then rewritten base by base from the better of the two calls. Everything is integer arithmetic except
// the error probability, a sum of table constants in double that is compared after ONE division: the sum is taken in the
// reference's order (j ascending, one thread), so the comparison is bit-exact.
//
// The functions are __host__ __device__ only so that tests/cpp/merge_core_host.cpp can run the SAME source on the CPU against
// the reference build's output where there is no GPU; the product has no CPU path (kc_merge_pairs fails without a device).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define KC_HD __host__ __device__ __forceinline__
#else
#define KC_HD inline
#endif

namespace kc {

// error probability of a Phred score (the reference's rounded table, src/merge_reads.cpp:74-84: parity needs these digits)
#define KC_Q2PERROR_VALUES                                                                                                            \
  {1.0,       0.7943,    0.6309,    0.5012,    0.3981,    0.3162,    0.2512,    0.1995,    0.1585,    0.1259,     0.1,       0.07943,   \
   0.06310,   0.05012,   0.03981,   0.03162,   0.02512,   0.01995,   0.01585,   0.01259,   0.01,      0.007943,   0.006310,  0.005012,  \
   0.003981,  0.003162,  0.002512,  0.001995,  0.001585,  0.001259,  0.001,     0.0007943, 0.0006310, 0.0005012,  0.0003981, 0.0003162, \
   0.0002512, 0.0001995, 0.0001585, 0.0001259, 0.0001,    7.943e-05, 6.310e-05, 5.012e-05, 3.981e-05, 3.162e-05,  2.512e-05, 1.995e-05, \
   1.585e-05, 1.259e-05, 1e-05,     7.943e-06, 6.310e-06, 5.012e-06, 3.981e-06, 3.162e-06, 2.512e-06, 1.995e-06,  1.585e-06, 1.259e-06, \
   1e-06,     7.943e-07, 6.310e-07, 5.012e-07, 3.981e-07, 3.1622e-07, 2.512e-07, 1.995e-07, 1.585e-07, 1.259e-07, 1e-07,     7.943e-08, \
   6.310e-08, 5.012e-08, 3.981e-08, 3.1622e-08, 2.512e-08, 1.995e-08, 1.585e-08, 1.259e-08, 1e-08}
static const int kQ2PerrorSize = 81;

static const int kMergeMinOverlap = 12;        // MIN_OVERLAP
static const int kMergeExtraTestOverlap = 2;   // EXTRA_TEST_OVERLAP
static const int kMergeMaxMismatches = 3;      // MAX_MISMATCHES
static const int kMergeExtraPer1000 = 150;     // EXTRA_MISMATCHES_PER_1000
#define KC_MERGE_MAX_PERROR 0.025              /* MAX_PERROR */

// complement of a sequence character as revcomp() gives it (src/utils.cpp:98-130): upper case, IUPAC codes -> N; 0 = illegal
KC_HD uint8_t merge_comp_char(uint8_t c) {
  switch (c) {
    case 'A': case 'a': return 'T';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    case 'T': case 't': return 'A';
    case 'N': case 'n':
    case 'U': case 'R': case 'Y': case 'K': case 'M': case 'S': case 'W': case 'B': case 'D': case 'H': case 'V': return 'N';
  }
  return 0;
}

// PackedRead byte of a (base character, quality character): src/packed_reads.cpp:99-126; 0xff = illegal base character
KC_HD uint8_t merge_packed_byte(uint8_t base, uint8_t qual, int qual_offset) {
  uint8_t b;
  switch (base) {
    case 'A': case 'a': b = 0; break;
    case 'C': case 'c': b = 1; break;
    case 'G': case 'g': b = 2; break;
    case 'T': case 't': b = 3; break;
    case 'N': case 'n':
    case 'U': case 'R': case 'Y': case 'K': case 'M': case 'S': case 'W': case 'B': case 'D': case 'H': case 'V': b = 4; break;
    default: return 0xff;
  }
  int q = (int)qual - qual_offset;
  if (q > 31) q = 31;
  return (uint8_t)(b | ((uint8_t)q << 3));
}

// mismatch budgets of an overlap (src/merge_reads.cpp:483-485, int16 arithmetic on promoted ints)
KC_HD void merge_limits(int overlap, int &this_max_mismatch, int &error_max_mismatch) {
  this_max_mismatch = kMergeMaxMismatches + (kMergeExtraPer1000 * overlap / 1000);
  error_max_mismatch = this_max_mismatch * 4 / 3 + 1;
}

// fast_count_mismatches(...) <= error_max_mismatch (src/merge_reads.cpp:186-247, 486): plain character inequality
KC_HD bool merge_prefilter(const uint8_t *a, const uint8_t *b, int overlap, int error_max_mismatch) {
  int mm = 0;
  for (int j = 0; j < overlap; j++) {
    mm += a[j] != b[j];
    if (mm > error_max_mismatch) return false;
  }
  return true;
}

struct MergeScan {
  int best_i, found_i;
  int abort_merge;
  int stop;            // the reference's `break` out of the offset loop
  unsigned ambiguous;  // num_ambiguous increments of this pair
  int bad_qual;        // a quality character outside [offset, offset + 80]: the reference DIEs
};
KC_HD void merge_scan_init(MergeScan &st) {
  st.best_i = -1;
  st.found_i = -1;
  st.abort_merge = 0;
  st.stop = 0;
  st.ambiguous = 0;
  st.bad_qual = 0;
}

// One offset i that passed the prefilter: src/merge_reads.cpp:488-578. s1 / q1 = mate 1 and its qualities, rc2 / rq2 = reverse
// complement of mate 2 and its reversed qualities (working copies: an N that meets a base zeroes its quality for good).
KC_HD void merge_detailed(MergeScan &st, const uint8_t *s1, uint8_t *q1, const uint8_t *rc2, uint8_t *rq2, int start_i, int i, int len,
                          int qual_offset, const double *q2p) {
  const int overlap = len - i;
  int this_max_mismatch, error_max_mismatch;
  merge_limits(overlap, this_max_mismatch, error_max_mismatch);
  int matches = 0, mismatches = 0, bothNs = 0, Ncount = 0, overlapChecked = 0;
  double perror = 0.0;
  for (int j = 0; j < overlap; j++) {
    overlapChecked++;
    const int p = start_i + i + j;
    const uint8_t ps = s1[p], rs = rc2[j];
    if (ps == rs) {
      matches++;
      if (ps == 'N') {
        Ncount += 2;
        if (bothNs++) {
          st.abort_merge++;
          st.ambiguous++;
          break;  // do not match multiple Ns in the same position -- 1 is okay
        }
      }
    } else {
      mismatches++;
      if (ps == 'N') {
        mismatches++;  // N still counts as a mismatch
        Ncount++;
        q1[p] = (uint8_t)qual_offset;
        const int x = (int)rq2[j] - qual_offset;
        if (x < 0 || x >= kQ2PerrorSize) {
          st.bad_qual = 1;
          st.stop = 1;
          return;
        }
        perror += q2p[x];
      } else if (rs == 'N') {
        Ncount++;
        mismatches++;
        rq2[j] = (uint8_t)qual_offset;
        const int x = (int)q1[p] - qual_offset;
        if (x < 0 || x >= kQ2PerrorSize) {
          st.bad_qual = 1;
          st.stop = 1;
          return;
        }
        perror += q2p[x];
      }
      const int a = (int)q1[p] - qual_offset, b = (int)rq2[j] - qual_offset;
      if (a < 0 || b < 0 || a >= kQ2PerrorSize || b >= kQ2PerrorSize) {
        st.bad_qual = 1;
        st.stop = 1;
        return;
      }
      const int diffq = a > b ? a - b : b - a;
      if (diffq <= 2)
        perror += 0.5;  // cap at flipping a coin when both quality scores are close
      else
        perror += q2p[diffq];
    }
    if (Ncount > 3) {
      st.abort_merge++;
      st.ambiguous++;
      break;  // do not match reads with many Ns
    }
    if (mismatches > error_max_mismatch) break;
  }
  int match_thres = overlap - this_max_mismatch;
  if (match_thres < kMergeMinOverlap) match_thres = kMergeMinOverlap;
  if (matches >= match_thres && overlapChecked == overlap && mismatches <= this_max_mismatch && perror / overlap <= KC_MERGE_MAX_PERROR) {
    if (st.best_i < 0 && st.found_i < 0) {
      st.best_i = i;
    } else {
      st.ambiguous++;  // another good or ambiguous overlap detected
      st.best_i = -1;
      st.stop = 1;
    }
  } else if (overlapChecked == overlap && mismatches <= error_max_mismatch && perror / overlap <= KC_MERGE_MAX_PERROR * 4 / 3) {
    st.found_i = i;  // lower threshold for detection of an ambiguous overlap
    if (st.best_i >= 0) {
      st.ambiguous++;  // ambiguous mapping found after a good one was
      st.best_i = -1;
      st.stop = 1;
    }
  }
  if (st.abort_merge) st.stop = 1;  // (the reference tests abort_merge at the top of the next offset)
}

// position j of the chosen overlap (src/merge_reads.cpp:584-607): agreeing calls add their qualities (capped at 41), a
// disagreement keeps the better call at the difference of the two (at least 2)
KC_HD void merge_base(uint8_t *s1, uint8_t *q1, const uint8_t *rc2, const uint8_t *rq2, int p, int j, int qual_offset) {
  if (s1[p] == rc2[j]) {
    const unsigned nq = ((unsigned)q1[p] + (unsigned)rq2[j] - (unsigned)qual_offset) & 0xffffu;
    const unsigned cap = (unsigned)(41 + qual_offset) & 0xffu;
    q1[p] = (uint8_t)(nq > cap ? cap : nq);
  } else {
    uint8_t nq;
    if (q1[p] < rq2[j]) {
      nq = (uint8_t)(rq2[j] - q1[p] + qual_offset);
      s1[p] = rc2[j];
    } else {
      nq = (uint8_t)(q1[p] - rq2[j] + qual_offset);
    }
    q1[p] = (nq > (2 + qual_offset)) ? nq : (uint8_t)(2 + qual_offset);
  }
}

}  // namespace kc
