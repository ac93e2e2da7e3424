// kc_merge.cu -- read-pair merging on the GPU: the step of the proxy that runs right before the k-mer analysis.
//
// Replaces the pair loop of merge_reads() (src/merge_reads.cpp:466-650; called from run_pipeline, src/main.cpp:210-215) and the
// PackedRead encoding of what it produces (src/packed_reads.cpp:99-126): mates come in as FASTQ text (sequence + quality
// characters, as kc_fastq_read_pairs delivers them), the merged / unmerged reads come out in the one-byte-per-base PackedRead
// encoding that kc_add_reads takes -- or go straight into a stage context without leaving HBM (kc_add_read_pairs).
// Why on the GPU: the reference does this on the host cores, ~14 us per pair and core (measured: oracle/_ref/merge_ref, 200 K pairs
// in 2.9 s); for the 10 M pairs of one 20 M-read GPU share that is ~9 s on 16 cores in front of a k-mer analysis that takes 0.075 s.
//
// One warp owns a pair (no CTA barrier):
//   1. stage: mate 1 and its qualities, the reverse complement of mate 2 and its reversed qualities -> shared memory
//   2. prefilter: lane l counts the character mismatches of offset g + l (32 offsets of the slide at a time), four characters
//      per step (aligned 32-bit shared-memory loads, funnel shift, per-byte compare), giving up at the offset's mismatch budget
//   3. the few offsets that pass (normally the true overlap alone) are scored in increasing order: the lanes classify 32
//      positions at a time (two ballots), and only the EVENTS -- mismatches and N against N, a handful per overlap -- are walked
//      in order with the reference's statements (kc_merge_core.cuh), so its error-probability sum (doubles, whose order must
//      not change), its early exits and the qualities an N comparison zeroes for all later offsets come out identical
//   4. merge: the lanes rewrite the overlap position by position and pack the result; an unmerged pair is packed as it came
// Integer / byte work throughout; bound by the shared-memory pipe and the serial scoring, not by HBM (2 bytes read and 1
// written per base).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/kcount_b200.h"
#include "kc_internal.h"
#include "kc_merge_core.cuh"

namespace kc {

static const int MG_WARPS = 8;
static_assert(MG_WARPS * 32 == 256, "merge_kernel fills its 256-entry character tables with one entry per thread");
static const int MG_MAXL = KC_MERGE_MAX_READ_LEN;  // longest mate the kernel merges (shared memory per warp = 4 x this)
static const int MG_CODEW = MG_MAXL / 8 + 2;       // code words per mate (+ the word the funnel shift reads past the end)

__constant__ double c_q2p[kQ2PerrorSize] = KC_Q2PERROR_VALUES;

struct MergeCounters {
  unsigned long long merged, ambiguous, merged_len, overlap_len, max_read_len, dropped_short, too_long, sum_read_lens;
  unsigned int err_bad_char, err_bad_qual;
};

struct MergeParams {
  const uint8_t *seq, *qual;     // FASTQ characters of all mates
  const uint64_t *offsets;       // [2 n_pairs] start of mate 2 p (first) and 2 p + 1 (second) inside seq / qual
  const uint16_t *lens;          // [2 n_pairs]
  const uint64_t *out_base;      // [n_pairs] first output byte of pair p (prefix sum of len1 + len2)
  uint64_t n_pairs;
  int qual_offset, min_kmer_len, paired;
  uint8_t *out_bytes;            // PackedRead encoding
  uint64_t *out_offsets;         // [2 n_pairs]
  uint16_t *out_lens;            // [2 n_pairs]; a dropped pair has both lengths 0
  MergeCounters *ctr;
};

__global__ void __launch_bounds__(MG_WARPS * 32, 4) merge_kernel(MergeParams p) {
  __shared__ __align__(16) uint8_t s_buf[MG_WARPS][4 * MG_MAXL];
  __shared__ uint32_t s_code[MG_WARPS][2 * MG_CODEW];  // 4-bit character codes of mate 1 and of the reverse complement of mate 2
  const uint32_t lane = lane_id();
  const int warp = threadIdx.x >> 5;
  uint8_t *s1 = s_buf[warp], *q1 = s1 + MG_MAXL, *rc2 = q1 + MG_MAXL, *rq2 = rc2 + MG_MAXL;
  uint32_t *s1c = s_code[warp], *rc2c = s1c + MG_CODEW;
  // character tables, built once per CTA from the functions of kc_merge_core.cuh (the switches there diverge: a third of the
  // kernel's instructions before this): complement (0 = illegal), PackedRead base code (0xff = illegal), 4-bit compare code
  __shared__ uint8_t s_comp[256], s_base[256], s_ccode[256];
  {
    const uint8_t c = (uint8_t)threadIdx.x;
    s_comp[c] = merge_comp_char(c);
    const uint8_t v = merge_packed_byte(c, (uint8_t)p.qual_offset, p.qual_offset);
    s_base[c] = v == 0xff ? 0xff : (uint8_t)(v & 7u);
    s_ccode[c] = (uint8_t)merge_char_code(c);
  }
  __syncthreads();  // (the only CTA barrier; the warps are independent from here on)
  auto packed = [&](uint8_t base, uint8_t qual) -> uint8_t {  // merge_packed_byte through the table
    const uint8_t b = s_base[base];
    int q = (int)qual - p.qual_offset;
    q = q > 31 ? 31 : q;
    return b == 0xff ? (uint8_t)0xff : (uint8_t)(b | ((uint8_t)q << 3));
  };
  const int qo = p.qual_offset;
  unsigned long long n_merged = 0, n_amb = 0, sum_mlen = 0, sum_ovl = 0, max_len = 0, n_drop = 0, n_long = 0, sum_lens = 0;
  unsigned bad_char = 0, bad_qual = 0;

  const uint64_t n_warps = (uint64_t)gridDim.x * MG_WARPS;
  for (uint64_t pr = (uint64_t)blockIdx.x * MG_WARPS + warp; pr < p.n_pairs; pr += n_warps) {
    const int len1 = p.lens[2 * pr], len2 = p.lens[2 * pr + 1];
    const uint8_t *g1 = p.seq + p.offsets[2 * pr], *gq1 = p.qual + p.offsets[2 * pr];
    const uint8_t *g2 = p.seq + p.offsets[2 * pr + 1], *gq2 = p.qual + p.offsets[2 * pr + 1];
    const uint64_t ob = p.out_base[pr];
    uint8_t *out = p.out_bytes + ob;
    // pairs of two short mates are not kept at all (src/merge_reads.cpp:472)
    if (p.paired && len1 < p.min_kmer_len && len2 < p.min_kmer_len) {
      if (lane == 0) {
        p.out_offsets[2 * pr] = p.out_offsets[2 * pr + 1] = ob;
        p.out_lens[2 * pr] = p.out_lens[2 * pr + 1] = 0;
        n_drop++;
      }
      continue;
    }
    if (lane == 0) sum_lens += (unsigned long long)(len1 + len2);
    bool merged = false;
    int best_i = -1, len = 0;
    const bool fits = len1 <= MG_MAXL && len2 <= MG_MAXL;
    if (fits) {
      // ---- 1. stage
      // (five rounds of byte loads in flight per lane before the first store -- a 150-base mate in one trip: the mates come from DRAM)
      for (int t0 = 0; t0 < len1 || t0 < len2; t0 += 160) {
        uint8_t a[5], b[5], c[5], d[5];
#pragma unroll
        for (int u = 0; u < 5; u++) {
          const int t = t0 + 32 * u + (int)lane;
          a[u] = t < len1 ? g1[t] : 0;
          b[u] = t < len1 ? gq1[t] : 0;
          c[u] = t < len2 ? g2[len2 - 1 - t] : 'N';
          d[u] = t < len2 ? gq2[len2 - 1 - t] : 0;
        }
#pragma unroll
        for (int u = 0; u < 5; u++) {
          const int t = t0 + 32 * u + (int)lane;
          if (t < len1) {
            s1[t] = a[u];
            q1[t] = b[u];
          }
          if (t < len2) {
            const uint8_t cc = s_comp[c[u]];
            if (!cc) bad_char = 1;
            rc2[t] = cc;
            rq2[t] = d[u];
          }
        }
      }
      __syncwarp();
      for (int w = (int)lane; 8 * w < len1; w += 32) s1c[w] = merge_pack_codes_lut(s1, w, s_ccode);
      for (int w = (int)lane; 8 * w < len2; w += 32) rc2c[w] = merge_pack_codes_lut(rc2, w, s_ccode);
      __syncwarp();
      // ---- 2 + 3. slide mate 2 along the tail of mate 1
      len = len2 < len1 ? len2 : len1;
      const int start_i = len == len1 ? 0 : len1 - len;
      const int n_off = len - kMergeMinOverlap + kMergeExtraTestOverlap;  // offsets 0 .. n_off - 1 are tested
      MergeScan st;
      merge_scan_init(st);
      int stop = 0;
      for (int g = 0; g < n_off && !stop; g += 32) {
        const int i = g + (int)lane;
        bool pass = false;
        if (i < n_off) {
          int tm, em;
          merge_limits(len - i, tm, em);
          pass = merge_prefilter(s1c, start_i + i, rc2c, len - i, em);
        }
        uint32_t m = __ballot_sync(0xffffffffu, pass);
        while (m && !stop) {
          // score offset g + b: the lanes classify 32 positions at a time, then every lane walks the (few) events of the
          // chunk in order -- the state is the same in all lanes, lane 0 stores the qualities an N comparison zeroes
          const int oi = g + __ffs(m) - 1;
          m &= m - 1u;
          const int overlap = len - oi, p0 = start_i + oi;
          int tm, em;
          merge_limits(overlap, tm, em);
          MergeEval ev;
          merge_eval_init(ev);
          for (int c = 0; c < overlap && !ev.broke; c += 32) {
            const int j = c + (int)lane;
            const bool valid = j < overlap;
            const uint8_t ps = valid ? s1[p0 + j] : 0, rs = valid ? rc2[j] : 1;
            const bool eq = valid && ps == rs;
            const uint32_t eq_mask = __ballot_sync(0xffffffffu, eq);
            const uint32_t ev_mask = __ballot_sync(0xffffffffu, valid && (!eq || ps == 'N'));
            merge_chunk(st, ev, eq_mask, ev_mask, c, min(32, overlap - c), s1, q1, rc2, rq2, p0, em, qo, c_q2p, lane == 0);
          }
          merge_conclude(st, ev, oi, overlap);
          __syncwarp();  // lane 0's quality stores are visible to the next offset's scoring
          stop = st.stop;
        }
      }
      best_i = st.best_i;  // (warp-uniform: every lane walked the same events)
      const int aborted = st.abort_merge;
      if (lane == 0) {
        n_amb += st.ambiguous;
        bad_qual |= (unsigned)st.bad_qual;
      }
      merged = best_i >= 0 && !aborted;
      if (merged) {
        // ---- 4. rewrite the overlap, then mate 1 + the rest of mate 2 is the merged read
        const int overlap = len - best_i;
        for (int j = (int)lane; j < overlap; j += 32) merge_base(s1, q1, rc2, rq2, start_i + best_i + j, j, qo);
        __syncwarp();
        const int mlen = len1 + len2 - overlap;
        for (int t = (int)lane; t < mlen; t += 32) {
          const uint8_t b = t < len1 ? s1[t] : rc2[overlap + t - len1];
          const uint8_t q = t < len1 ? q1[t] : rq2[overlap + t - len1];
          const uint8_t v = packed(b, q);
          if (v == 0xff) bad_char = 1;
          out[t] = v;
        }
        if (lane == 0) {
          out[mlen] = 4;  // the dummy mate "N" with quality 0 that keeps the reads in pairs (src/merge_reads.cpp:626)
          p.out_offsets[2 * pr] = ob;
          p.out_lens[2 * pr] = (uint16_t)mlen;
          p.out_offsets[2 * pr + 1] = ob + (uint64_t)mlen;
          p.out_lens[2 * pr + 1] = 1;
          n_merged++;
          sum_mlen += (unsigned long long)mlen;
          sum_ovl += (unsigned long long)overlap;
          if ((unsigned long long)mlen > max_len) max_len = (unsigned long long)mlen;
        }
      }
    } else if (lane == 0) {
      n_long++;
    }
    if (!merged) {
      // both mates as they came -- except that mate 1 keeps the qualities an N comparison zeroed (src/merge_reads.cpp:634)
      for (int t = (int)lane; t < len1; t += 32) {
        const uint8_t v = packed(g1[t], fits ? q1[t] : gq1[t]);
        if (v == 0xff) bad_char = 1;
        out[t] = v;
      }
      for (int t = (int)lane; t < len2; t += 32) {
        const uint8_t v = packed(g2[t], gq2[t]);
        if (v == 0xff) bad_char = 1;
        out[len1 + t] = v;
      }
      if (lane == 0) {
        p.out_offsets[2 * pr] = ob;
        p.out_lens[2 * pr] = (uint16_t)len1;
        p.out_offsets[2 * pr + 1] = ob + (uint64_t)len1;
        p.out_lens[2 * pr + 1] = (uint16_t)len2;
      }
    }
    __syncwarp();  // the stage buffers are reused by the next pair
  }

  bad_char = __any_sync(0xffffffffu, bad_char != 0);
  if (lane == 0) {
    if (n_merged) atomicAdd(&p.ctr->merged, n_merged);
    if (n_amb) atomicAdd(&p.ctr->ambiguous, n_amb);
    if (sum_mlen) atomicAdd(&p.ctr->merged_len, sum_mlen);
    if (sum_ovl) atomicAdd(&p.ctr->overlap_len, sum_ovl);
    if (max_len) atomicMax(&p.ctr->max_read_len, max_len);
    if (n_drop) atomicAdd(&p.ctr->dropped_short, n_drop);
    if (n_long) atomicAdd(&p.ctr->too_long, n_long);
    if (sum_lens) atomicAdd(&p.ctr->sum_read_lens, sum_lens);
    if (bad_char) p.ctr->err_bad_char = 1u;
    if (bad_qual) p.ctr->err_bad_qual = 1u;
  }
}

// everything of one call that lives on the device
struct MergeDevice {
  uint8_t *seq = nullptr, *qual = nullptr, *out_bytes = nullptr;
  uint64_t *offsets = nullptr, *out_base = nullptr, *out_offsets = nullptr;
  uint16_t *lens = nullptr, *out_lens = nullptr;
  MergeCounters *ctr = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  void release() {
    cudaFree(seq), cudaFree(qual), cudaFree(out_bytes), cudaFree(offsets), cudaFree(out_base), cudaFree(out_offsets), cudaFree(lens),
        cudaFree(out_lens), cudaFree(ctr);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    *this = MergeDevice();
  }
};

#define MCK(call)                                                                                      \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess) {                                                                           \
      fprintf(stderr, "[kc] %s:%d %s: %s\n", __FILE__, __LINE__, #call, cudaGetErrorString(e_));        \
      d.release();                                                                                     \
      return e_ == cudaErrorMemoryAllocation ? KC_ERR_NOMEM : KC_ERR_CUDA;                             \
    }                                                                                                  \
  } while (0)

// host FASTQ characters -> merged PackedRead bytes in HBM (d.out_*); the caller copies them out or hands them on
static int merge_on_device(int device, const char *seq, const char *qual, const uint64_t *offsets, const uint16_t *lens, uint64_t n_pairs,
                           uint64_t n_bytes, int qual_offset, int min_kmer_len, int paired, MergeDevice &d, kc_merge_stats *stats,
                           MergeCounters &hc) {
  if (!seq || !qual || !offsets || !lens || (qual_offset != 33 && qual_offset != 64)) return KC_ERR_ARG;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1 || device < 0 || device >= n_dev) return KC_ERR_CUDA;  // no CPU path
  MCK(cudaSetDevice(device));
  const uint64_t n_reads = 2 * n_pairs;
  std::vector<uint64_t> base(std::max<uint64_t>(n_pairs, 1));
  uint64_t total = 0;
  for (uint64_t i = 0; i < n_pairs; i++) {
    base[i] = total;
    total += (uint64_t)lens[2 * i] + lens[2 * i + 1];
    if (offsets[2 * i] + lens[2 * i] > n_bytes || offsets[2 * i + 1] + lens[2 * i + 1] > n_bytes) return KC_ERR_ARG;
  }
  MCK(cudaMalloc(&d.seq, std::max<uint64_t>(n_bytes, 1)));
  MCK(cudaMalloc(&d.qual, std::max<uint64_t>(n_bytes, 1)));
  MCK(cudaMalloc(&d.out_bytes, total + 16));
  MCK(cudaMalloc(&d.offsets, std::max<uint64_t>(n_reads, 1) * 8));
  MCK(cudaMalloc(&d.out_base, std::max<uint64_t>(n_pairs, 1) * 8));
  MCK(cudaMalloc(&d.out_offsets, std::max<uint64_t>(n_reads, 1) * 8));
  MCK(cudaMalloc(&d.lens, std::max<uint64_t>(n_reads, 1) * 2));
  MCK(cudaMalloc(&d.out_lens, std::max<uint64_t>(n_reads, 1) * 2));
  MCK(cudaMalloc(&d.ctr, sizeof(MergeCounters)));
  MCK(cudaEventCreate(&d.e0));
  MCK(cudaEventCreate(&d.e1));
  MCK(cudaMemset(d.ctr, 0, sizeof(MergeCounters)));
  MCK(cudaMemcpy(d.seq, seq, n_bytes, cudaMemcpyDefault));
  MCK(cudaMemcpy(d.qual, qual, n_bytes, cudaMemcpyDefault));
  MCK(cudaMemcpy(d.offsets, offsets, n_reads * 8, cudaMemcpyDefault));
  MCK(cudaMemcpy(d.lens, lens, n_reads * 2, cudaMemcpyDefault));
  MCK(cudaMemcpy(d.out_base, base.data(), n_pairs * 8, cudaMemcpyHostToDevice));
  MergeParams p;
  p.seq = d.seq, p.qual = d.qual, p.offsets = d.offsets, p.lens = d.lens, p.out_base = d.out_base;
  p.n_pairs = n_pairs, p.qual_offset = qual_offset, p.min_kmer_len = min_kmer_len, p.paired = paired;
  p.out_bytes = d.out_bytes, p.out_offsets = d.out_offsets, p.out_lens = d.out_lens, p.ctr = d.ctr;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const uint64_t want = (n_pairs + MG_WARPS - 1) / MG_WARPS;
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)sms * 8));  // persistent: 8 CTAs of 8 warps per SM
  MCK(cudaEventRecord(d.e0, 0));
  if (n_pairs) merge_kernel<<<grid, MG_WARPS * 32>>>(p);
  MCK(cudaGetLastError());
  MCK(cudaEventRecord(d.e1, 0));
  MCK(cudaMemcpy(&hc, d.ctr, sizeof hc, cudaMemcpyDeviceToHost));
  float ms = 0;
  cudaEventElapsedTime(&ms, d.e0, d.e1);
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->pairs = n_pairs;
    stats->merged = hc.merged;
    stats->ambiguous = hc.ambiguous;
    stats->merged_len = hc.merged_len;
    stats->overlap_len = hc.overlap_len;
    stats->max_read_len = hc.max_read_len;
    stats->dropped_short = hc.dropped_short;
    stats->too_long = hc.too_long;
    stats->sum_read_lens = hc.sum_read_lens;
    stats->out_bytes = total;
    stats->ms_kernel = ms;
    stats->launches = n_pairs ? 1 : 0;
  }
  if (hc.err_bad_char || hc.err_bad_qual) {
    fprintf(stderr, "[kc] kc_merge_pairs: %s\n", hc.err_bad_char ? "illegal base character" : "quality character outside [offset, offset + 80]");
    d.release();
    return KC_ERR_ARG;  // the reference DIEs on both (src/utils.cpp:124, src/merge_reads.cpp:525-530)
  }
  return KC_OK;
}

}  // namespace kc

extern "C" int kc_merge_pairs(int device, const char *seq, const char *qual, const uint64_t *offsets, const uint16_t *lens, uint64_t n_pairs,
                              uint64_t n_bytes, int qual_offset, int min_kmer_len, int paired, uint8_t *out_bytes, uint64_t *out_offsets,
                              uint16_t *out_lens, uint64_t *n_out_reads, kc_merge_stats *stats) {
  if (!out_bytes || !out_offsets || !out_lens || !n_out_reads) return KC_ERR_ARG;
  kc::MergeDevice d;
  kc::MergeCounters hc;
  int rc = kc::merge_on_device(device, seq, qual, offsets, lens, n_pairs, n_bytes, qual_offset, min_kmer_len, paired, d, stats, hc);
  if (rc) return rc;
  const uint64_t n_reads = 2 * n_pairs;
  uint64_t total = 0;
  for (uint64_t i = 0; i < n_reads; i++) total += lens[i];
  cudaError_t e = cudaMemcpy(out_bytes, d.out_bytes, total, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(out_offsets, d.out_offsets, n_reads * 8, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(out_lens, d.out_lens, n_reads * 2, cudaMemcpyDeviceToHost);
  d.release();
  if (e != cudaSuccess) return KC_ERR_CUDA;
  // pairs that were dropped leave no reads behind (the byte buffer keeps its gaps: offsets say where the reads are)
  uint64_t w = 0;
  for (uint64_t pr = 0; pr < n_pairs; pr++) {
    if (paired && out_lens[2 * pr] == 0 && out_lens[2 * pr + 1] == 0 && lens[2 * pr] < min_kmer_len && lens[2 * pr + 1] < min_kmer_len) continue;
    out_offsets[w] = out_offsets[2 * pr], out_lens[w] = out_lens[2 * pr];
    out_offsets[w + 1] = out_offsets[2 * pr + 1], out_lens[w + 1] = out_lens[2 * pr + 1];
    w += 2;
  }
  *n_out_reads = w;
  return KC_OK;
}

extern "C" int kc_add_read_pairs(kc_ctx *ctx, const char *seq, const char *qual, const uint64_t *offsets, const uint16_t *lens,
                                 uint64_t n_pairs, uint64_t n_bytes, int qual_offset, int min_kmer_len, int paired, kc_merge_stats *stats) {
  if (!ctx) return KC_ERR_ARG;
  kc::MergeDevice d;
  kc::MergeCounters hc;
  kc_merge_stats local;
  int rc = kc::merge_on_device(kc::ctx_device(ctx), seq, qual, offsets, lens, n_pairs, n_bytes, qual_offset, min_kmer_len, paired, d, &local, hc);
  if (rc) return rc;
  if (stats) *stats = local;
  if (n_pairs == 0) {
    d.release();
    return KC_OK;
  }
  // the merged reads never leave HBM: kc_add_reads copies device to device (a dropped pair stays behind as two empty reads)
  rc = kc_add_reads(ctx, d.out_bytes, local.out_bytes, d.out_offsets, d.out_lens, 2 * n_pairs);
  if (rc == KC_OK) rc = kc_sync(ctx);
  d.release();
  return rc;
}
