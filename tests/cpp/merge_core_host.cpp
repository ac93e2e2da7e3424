// TEST INFRASTRUCTURE -- runs the per-pair arithmetic of the merge kernel (mhm2_kmer_analysis_v2_b200/csrc/kc_merge_core.cuh,
// the very source kc_merge.cu compiles for the device) on the CPU, in the kernel's order of operations with the 32 lanes
// played one after the other, behind the product's FASTQ pair reader (kc_fastq_read_pairs). Exists so that the merge
// arithmetic can be held to the reference build (oracle/_ref/merge_ref) where there is no GPU; it is not shipped and nothing
// in the package calls it -- kc_merge_pairs itself has no CPU path.
//
// usage: merge_core_host <qual_offset> <min_kmer_len> <paired 0|1> <out.txt> <fastq> [fastq2]
// writes "seq\nquals\n" per read, qualities as PackedRead keeps them (offset + min(q - offset, 31)).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "kc_synth.h"
#include "../../mhm2_kmer_analysis_v2_b200/csrc/kc_merge_core.cuh"

using namespace kc;
static const double q2p[kQ2PerrorSize] = KC_Q2PERROR_VALUES;

static void put(FILE *f, const uint8_t *packed, int n, int qo) {
  for (int i = 0; i < n; i++) fputc("ACGTNNNN"[packed[i] & 7], f);
  fputc('\n', f);
  for (int i = 0; i < n; i++) fputc(qo + (packed[i] >> 3), f);
  fputc('\n', f);
}

int main(int argc, char **argv) {
  if (argc < 6) return 2;
  const int qo = atoi(argv[1]), min_k = atoi(argv[2]), paired = atoi(argv[3]);
  kc_pair_set ps;
  int rc = kc_fastq_read_pairs(argv[5], argc > 6 ? argv[6] : NULL, paired, qo, 0, &ps);
  if (rc) {
    fprintf(stderr, "kc_fastq_read_pairs: %d\n", rc);
    return 1;
  }
  FILE *f = fopen(argv[4], "w");
  unsigned long long merged = 0, ambiguous = 0, dropped = 0;
  std::vector<uint8_t> s1, q1, rc2, rq2, out;
  for (uint64_t pr = 0; pr < ps.n_reads / 2; pr++) {
    const int len1 = ps.lens[2 * pr], len2 = ps.lens[2 * pr + 1];
    const uint8_t *g1 = (const uint8_t *)ps.seq + ps.offsets[2 * pr], *gq1 = (const uint8_t *)ps.qual + ps.offsets[2 * pr];
    const uint8_t *g2 = (const uint8_t *)ps.seq + ps.offsets[2 * pr + 1], *gq2 = (const uint8_t *)ps.qual + ps.offsets[2 * pr + 1];
    if (paired && len1 < min_k && len2 < min_k) {
      dropped++;
      continue;
    }
    // (buffers padded to the next word + 4: the prefilter reads whole aligned words, as the kernel's shared-memory buffers allow)
    s1.assign(len1 + 12, 0), q1.assign(len1 + 12, 0);
    memcpy(s1.data(), g1, len1), memcpy(q1.data(), gq1, len1);
    rc2.assign(len2 + 12, 0), rq2.assign(len2 + 12, 0);
    for (int t = 0; t < len2; t++) {
      rc2[t] = merge_comp_char(g2[len2 - 1 - t]);
      rq2[t] = gq2[len2 - 1 - t];
      if (!rc2[t]) return 3;
    }
    std::vector<uint32_t> s1c(len1 / 8 + 3, 0xffffffffu), rc2c(len2 / 8 + 3, 0xffffffffu);  // (garbage past the end, as in the kernel)
    for (int w = 0; 8 * w < len1; w++) s1c[w] = merge_pack_codes(s1.data(), w);
    for (int w = 0; 8 * w < len2; w++) rc2c[w] = merge_pack_codes(rc2.data(), w);
    const int len = len2 < len1 ? len2 : len1;
    const int start_i = len == len1 ? 0 : len1 - len;
    const int n_off = len - kMergeMinOverlap + kMergeExtraTestOverlap;
    MergeScan st;
    merge_scan_init(st);
    for (int g = 0; g < n_off && !st.stop; g += 32) {
      unsigned m = 0;
      for (int lane = 0; lane < 32; lane++) {  // the prefilter of all 32 offsets comes before any scoring, as in the kernel
        const int i = g + lane;
        if (i >= n_off) continue;
        int tm, em;
        merge_limits(len - i, tm, em);
        if (merge_prefilter(s1c.data(), start_i + i, rc2c.data(), len - i, em)) m |= 1u << lane;
      }
      while (m && !st.stop) {
        const int oi = g + __builtin_ffs(m) - 1;
        m &= m - 1u;
        const int overlap = len - oi, p0 = start_i + oi;
        int tm, em;
        merge_limits(overlap, tm, em);
        MergeEval ev;
        merge_eval_init(ev);
        for (int c = 0; c < overlap && !ev.broke; c += 32) {
          unsigned eq_mask = 0, ev_mask = 0;  // the kernel's two ballots
          for (int lane = 0; lane < 32 && c + lane < overlap; lane++) {
            const uint8_t ps = s1[p0 + c + lane], rs = rc2[c + lane];
            if (ps == rs) eq_mask |= 1u << lane;
            if (ps != rs || ps == 'N') ev_mask |= 1u << lane;
          }
          merge_chunk(st, ev, eq_mask, ev_mask, c, overlap - c < 32 ? overlap - c : 32, s1.data(), q1.data(), rc2.data(), rq2.data(), p0, em, qo, q2p, true);
        }
        merge_conclude(st, ev, oi, overlap);
      }
    }
    if (st.bad_qual) return 4;
    ambiguous += st.ambiguous;
    out.assign(len1 + len2 + 1, 0);
    if (st.best_i >= 0 && !st.abort_merge) {
      const int overlap = len - st.best_i;
      for (int j = 0; j < overlap; j++) merge_base(s1.data(), q1.data(), rc2.data(), rq2.data(), start_i + st.best_i + j, j, qo);
      const int mlen = len1 + len2 - overlap;
      for (int t = 0; t < mlen; t++)
        out[t] = merge_packed_byte(t < len1 ? s1[t] : rc2[overlap + t - len1], t < len1 ? q1[t] : rq2[overlap + t - len1], qo);
      out[mlen] = 4;
      put(f, out.data(), mlen, qo);
      put(f, out.data() + mlen, 1, qo);
      merged++;
    } else {
      for (int t = 0; t < len1; t++) out[t] = merge_packed_byte(g1[t], q1[t], qo);
      for (int t = 0; t < len2; t++) out[len1 + t] = merge_packed_byte(g2[t], gq2[t], qo);
      put(f, out.data(), len1, qo);
      put(f, out.data() + len1, len2, qo);
    }
  }
  fclose(f);
  fprintf(stderr, "pairs %llu merged %llu ambiguous %llu dropped %llu missing1 %llu missing2 %llu\n", (unsigned long long)ps.n_reads / 2, merged,
          ambiguous, dropped, (unsigned long long)ps.missing_read1, (unsigned long long)ps.missing_read2);
  kc_pair_set_free(&ps);
  return 0;
}
