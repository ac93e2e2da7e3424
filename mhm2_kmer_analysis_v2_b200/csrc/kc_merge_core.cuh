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

// ---- small portable helpers (device intrinsics, plain C on the host) -----------------------------------------------------
KC_HD uint32_t mc_load32(const uint8_t *p) {  // 4-byte aligned
  return *reinterpret_cast<const uint32_t *>(p);
}
KC_HD uint32_t mc_funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {  // low word of (hi:lo) >> sh, 0 <= sh < 32
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, sh);
#else
  return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
KC_HD uint32_t mc_ne_bytes(uint32_t a, uint32_t b) {  // 0xff in every byte where a and b differ
#if defined(__CUDA_ARCH__)
  return __vcmpne4(a, b);
#else
  uint32_t x = a ^ b, r = 0;
  for (int k = 0; k < 4; k++)
    if ((x >> (8 * k)) & 0xffu) r |= 0xffu << (8 * k);
  return r;
#endif
}
KC_HD int mc_popc(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
KC_HD int mc_ffs(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __ffs(x);
#else
  return __builtin_ffs((int)x);
#endif
}

// fast_count_mismatches(...) <= error_max_mismatch (src/merge_reads.cpp:186-247, 486) is plain character inequality. Here the
// characters are first reduced to 4-bit codes, eight per 32-bit word (character 8 w + k in bits 4 k .. 4 k + 3 of word w):
// A C G T N -> 0..4, anything else -> 5. The reverse complement of mate 2 only holds A C G T N (merge_comp_char), so two codes
// are equal exactly when the two characters are (a lower-case or IUPAC character of mate 1 equals nothing, as in the reference).
KC_HD uint32_t merge_char_code(uint8_t c) { return c == 'A' ? 0u : (c == 'C' ? 1u : (c == 'G' ? 2u : (c == 'T' ? 3u : (c == 'N' ? 4u : 5u)))); }
// codes of characters [8 w, 8 w + 8) of a 4-byte aligned character buffer (readable to the next multiple of 8)
KC_HD uint32_t merge_pack_codes(const uint8_t *chars, int w) {
  const uint32_t lo = mc_load32(chars + 8 * w), hi = mc_load32(chars + 8 * w + 4);
  uint32_t r = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int k = 0; k < 4; k++) {
    r |= merge_char_code((uint8_t)(lo >> (8 * k))) << (4 * k);
    r |= merge_char_code((uint8_t)(hi >> (8 * k))) << (4 * k + 16);
  }
  return r;
}
// the same through a 256-entry table of merge_char_code (the kernel keeps one in shared memory: no branches, no divergence)
KC_HD uint32_t merge_pack_codes_lut(const uint8_t *chars, int w, const uint8_t *code_lut) {
  const uint32_t lo = mc_load32(chars + 8 * w), hi = mc_load32(chars + 8 * w + 4);
  uint32_t r = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int k = 0; k < 4; k++) {
    r |= (uint32_t)code_lut[(lo >> (8 * k)) & 0xffu] << (4 * k);
    r |= (uint32_t)code_lut[(hi >> (8 * k)) & 0xffu] << (4 * k + 16);
  }
  return r;
}
// mismatches of mate 1 from character p0 on against rc2 from character 0 on, `overlap` characters, eight per step; false as
// soon as the budget is exceeded. s1c needs one readable word past the last one that holds a character of the overlap.
KC_HD bool merge_prefilter(const uint32_t *s1c, int p0, const uint32_t *rc2c, int overlap, int error_max_mismatch) {
  const uint32_t *a_base = s1c + (p0 >> 3);
  const uint32_t sh = (uint32_t)(p0 & 7) * 4u;
  int mm = 0;
  uint32_t lo = a_base[0];
  for (int w = 0; 8 * w < overlap; w++) {
    const uint32_t hi = a_base[w + 1];
    const uint32_t x = mc_funnel_r(lo, hi, sh) ^ rc2c[w];
    lo = hi;
    uint32_t y = (x | (x >> 1) | (x >> 2)) & 0x11111111u;  // one bit per differing code (codes are < 8)
    const int rem = overlap - 8 * w;
    if (rem < 8) y &= (1u << (4 * rem)) - 1u;
    mm += mc_popc(y);
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

// Scoring of one offset that passed the prefilter (src/merge_reads.cpp:488-545), restructured around EVENTS: a position where
// the two calls agree on a base changes nothing but `matches` and `overlapChecked`, and neither of the loop's two exits can
// fire there (their conditions were tested at the last event). So the positions are classified 32 at a time (in the kernel: one
// per lane, two ballots), and only the events -- mismatches and N == N -- are walked in order with the reference's statements:
// the counters, the error-probability sum (doubles, added in the reference's order) and the breaks come out identical.
struct MergeEval {
  int matches, mismatches, bothNs, Ncount, overlapChecked, broke;
  double perror;
};
KC_HD void merge_eval_init(MergeEval &ev) {
  ev.matches = ev.mismatches = ev.bothNs = ev.Ncount = ev.overlapChecked = ev.broke = 0;
  ev.perror = 0.0;
}

// one event at overlap position j (p = its position in mate 1); returns true when the reference's loop breaks here.
// `writer`: this caller stores the zeroed qualities (in the kernel every lane walks the events, lane 0 stores).
KC_HD bool merge_event(MergeScan &st, MergeEval &ev, const uint8_t *s1, uint8_t *q1, const uint8_t *rc2, uint8_t *rq2, int p, int j,
                       int error_max_mismatch, int qual_offset, const double *q2p, bool writer) {
  const uint8_t ps = s1[p], rs = rc2[j];
  if (ps == rs) {  // N == N (any other agreement is not an event)
    ev.Ncount += 2;
    if (ev.bothNs++) {
      st.abort_merge++;
      st.ambiguous++;
      return true;  // do not match multiple Ns in the same position -- 1 is okay
    }
  } else {
    int qa = (int)q1[p], qb = (int)rq2[j];
    ev.mismatches++;
    if (ps == 'N') {
      ev.mismatches++;  // N still counts as a mismatch
      ev.Ncount++;
      qa = qual_offset;
      if (writer) q1[p] = (uint8_t)qual_offset;
      const int x = qb - qual_offset;
      if (x < 0 || x >= kQ2PerrorSize) {
        st.bad_qual = 1;
        st.stop = 1;
        return true;
      }
      ev.perror += q2p[x];
    } else if (rs == 'N') {
      ev.Ncount++;
      ev.mismatches++;
      qb = qual_offset;
      if (writer) rq2[j] = (uint8_t)qual_offset;
      const int x = qa - qual_offset;
      if (x < 0 || x >= kQ2PerrorSize) {
        st.bad_qual = 1;
        st.stop = 1;
        return true;
      }
      ev.perror += q2p[x];
    }
    const int a = qa - qual_offset, b = qb - qual_offset;
    if (a < 0 || b < 0 || a >= kQ2PerrorSize || b >= kQ2PerrorSize) {
      st.bad_qual = 1;
      st.stop = 1;
      return true;
    }
    const int diffq = a > b ? a - b : b - a;
    if (diffq <= 2)
      ev.perror += 0.5;  // cap at flipping a coin when both quality scores are close
    else
      ev.perror += q2p[diffq];
  }
  if (ev.Ncount > 3) {
    st.abort_merge++;
    st.ambiguous++;
    return true;  // do not match reads with many Ns
  }
  return ev.mismatches > error_max_mismatch;
}

// 32 positions [c, c + n_valid) of the overlap: eq_mask = positions where the characters are equal, ev_mask = events
KC_HD void merge_chunk(MergeScan &st, MergeEval &ev, uint32_t eq_mask, uint32_t ev_mask, int c, int n_valid, const uint8_t *s1, uint8_t *q1,
                       const uint8_t *rc2, uint8_t *rq2, int p0, int error_max_mismatch, int qual_offset, const double *q2p, bool writer) {
  while (ev_mask) {
    const int b = mc_ffs(ev_mask) - 1;
    ev_mask &= ev_mask - 1u;
    if (merge_event(st, ev, s1, q1, rc2, rq2, p0 + c + b, c + b, error_max_mismatch, qual_offset, q2p, writer)) {
      ev.matches += mc_popc(eq_mask & (0xffffffffu >> (31 - b)));
      ev.overlapChecked = c + b + 1;
      ev.broke = 1;
      return;
    }
  }
  ev.matches += mc_popc(eq_mask);
  ev.overlapChecked = c + n_valid;
}

// verdict on offset i once its positions are scored: src/merge_reads.cpp:546-578
KC_HD void merge_conclude(MergeScan &st, const MergeEval &ev, int i, int overlap) {
  if (st.bad_qual) return;
  int this_max_mismatch, error_max_mismatch;
  merge_limits(overlap, this_max_mismatch, error_max_mismatch);
  int match_thres = overlap - this_max_mismatch;
  if (match_thres < kMergeMinOverlap) match_thres = kMergeMinOverlap;
  if (ev.matches >= match_thres && ev.overlapChecked == overlap && ev.mismatches <= this_max_mismatch &&
      ev.perror / overlap <= KC_MERGE_MAX_PERROR) {
    if (st.best_i < 0 && st.found_i < 0) {
      st.best_i = i;
    } else {
      st.ambiguous++;  // another good or ambiguous overlap detected
      st.best_i = -1;
      st.stop = 1;
    }
  } else if (ev.overlapChecked == overlap && ev.mismatches <= error_max_mismatch && ev.perror / overlap <= KC_MERGE_MAX_PERROR * 4 / 3) {
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
