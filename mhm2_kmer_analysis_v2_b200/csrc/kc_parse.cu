// kc_parse.cu -- kernels (1) "read parse + 2-bit pack" and (2) "minimizer / supermer extraction".
//
// Replaces, for the path of BASELINE.json: count_kmers' per-read unpack + quality mask (src/kcount/kcount.cpp:71-90),
// SeqBlockInserter::process_seq (src/kcount/kcount_cpu.cpp:74-103) and the reference's GPU kernels parse_and_pack /
// build_supermers / pack_seqs (src/kcount/kcount-gpu/parse_and_pack.cpp:127-237).
//
// One warp owns one read (in segments of <= 256 bases). The bytes (base | qual << 3, src/packed_reads.cpp:99-126)
// are turned into bit planes with three ballots per 32 bases, interleaved into the 2-bit Kmer word layout and kept in
// shared memory together with a "high quality" bit plane. N becomes (G, low quality): inside a k-mer the reference maps
// N to G (src/kmer.cpp:173,191-192) and as an extension an N is never counted (kcount_cpu.cpp:157-164).
// Canonical m-mers are taken per position with a funnel shift + bit-reversal, the reference's "greatest least-complement
// m-mer" minimizer (src/kmer.cpp:345-398) is a sliding-window maximum computed by max-doubling, the owner is
// quick_hash(minimizer) % n_ranks (src/kcount/kmer_dht.cpp:193-196) and the local bin (hash / n_ranks) & (bins-1).
// Runs of equal (owner, bin) are found with ballots; each run (cut to the record capacity) becomes one fixed-size
// supermer record written with 128-bit stores into the per-(owner, bin) area of the send buffer.
#include "kc_internal.h"

namespace kc {

static const int SEG = 256;       // bases of a read handled per warp iteration
static const int WARPS = 8;       // warps per CTA
static const int RESV = 64;       // records reserved per (warp, bin) at a time on the few-bins path

struct WarpSmem {
  uint64_t W[SEG / 32 + 2];   // 2-bit packed bases of the segment (+ zero padding for window reads)
  uint32_t Q[SEG / 32 + 2];   // high-quality bit plane, base j at bit 31 - j%32
  uint64_t C[SEG];            // canonical m-mers, then running window maxima
  uint32_t ID[SEG];           // (owner, bin) id per counted k-mer
  unsigned long long resv_base[32];
  uint32_t resv_used[32];
};

template <int R>
__device__ __forceinline__ void write_empty(uint8_t *bins, uint64_t byte_base, uint64_t slot) {
  *reinterpret_cast<uint32_t *>(bins + byte_base + slot * R + (R - 4)) = 0u;
}

// build one record for k-mers [start, start + n) of the segment (segment-relative k-mer index = index of the
// record's first base) and store it
template <int R>
__device__ __forceinline__ void write_record(const WarpSmem &sm, uint8_t *dst, int start, int n, int k) {
  constexpr int W = R / 4;
  constexpr int NBW = (R == 16 ? 32 : (R == 32 ? 80 : 160)) / 16;
  constexpr int NQW = W - NBW - (R == 32 ? 0 : 1);  // R=32: third quality word shares the header word
  const int nbases = n + k + 1;
  uint32_t w[W];
#pragma unroll
  for (int u = 0; u < NBW; u++) {
    int valid = nbases - 16 * u;
    uint32_t x = 0;
    if (valid > 0) {
      x = (uint32_t)(window64(sm.W, start + 16 * u) >> 32);
      if (valid < 16) x &= ~(0xffffffffu >> (2 * valid));
    }
    w[u] = x;
  }
#pragma unroll
  for (int v = 0; v < NQW; v++) {
    int valid = nbases - 32 * v;
    uint32_t x = 0;
    if (valid > 0) {
      int pos = start + 32 * v, q = pos >> 5, s = pos & 31;
      x = __funnelshift_l(sm.Q[q + 1], sm.Q[q], s);
      if (valid < 32) x &= ~(0xffffffffu >> valid);
    }
    w[NBW + v] = x;
  }
  if (R == 32)
    w[W - 1] = (w[W - 1] & 0xffff0000u) | (uint32_t)n;
  else
    w[W - 1] = (uint32_t)n;
#pragma unroll
  for (int i = 0; i < W / 4; i++) st_stream16(dst + 16 * i, make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]));
}

template <int R, bool FEWBINS>
__global__ void __launch_bounds__(WARPS * 32) parse_kernel(ParseParams p) {
  __shared__ WarpSmem smem[WARPS];
  const int warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  WarpSmem &sm = smem[warp];
  const int k = p.k, m = p.m;
  const int w = k - m + 1;                       // minimizer window: m-mers per k-mer
  const RecordGeom geom = record_geom(R, k);
  const int max_n = geom.max_kmers;              // counted k-mers per record
  const uint32_t inv_max_n = 65536u / (uint32_t)max_n + 1u;  // x / max_n for x < 256 (exact: checked on host)
  const int seg_kmers = SEG - k - 1;             // counted k-mers per segment
  const bool need_ids = p.n_ids > 1;
  const uint64_t mmask = (m < 32) ? ((1ULL << (2 * m)) - 1ULL) : ~0ULL;

  if (FEWBINS) {
    sm.resv_base[lane] = ~0ULL;
    sm.resv_used[lane] = RESV;
  }
  __syncwarp();

  unsigned long long n_counted = 0, n_records = 0, n_raw = 0;
  const uint64_t warp_global = (uint64_t)blockIdx.x * WARPS + warp;
  const uint64_t n_warps = (uint64_t)gridDim.x * WARPS;

  for (uint64_t ridx = warp_global * p.read_stride; ridx < p.n_reads; ridx += n_warps * p.read_stride) {
    const int L = p.lens[ridx];
    if (L < k) continue;  // kcount.cpp:78
    if (lane == 0) n_raw += L - k + 1;
    const int nk = L - k - 1;  // counted (interior) k-mers: kcount_cpu.cpp:86
    if (nk <= 0) continue;
    if (lane == 0) n_counted += nk;
    const uint8_t *rd = p.bytes + p.offsets[ridx];

    for (int j0 = 0; j0 < nk; j0 += seg_kmers) {
      const int nks = min(seg_kmers, nk - j0);  // counted k-mers in this segment
      const int nb = nks + k + 1;               // bases [j0, j0 + nb)
      const int nchunks = (nb + 31) >> 5;
      // ---- (1) parse + pack: bytes -> 2-bit words + quality plane
      for (int c = 0; c < nchunks; c++) {
        int pos = 32 * c + (int)lane;
        uint32_t code = 0, hq = 0;
        if (pos < nb) {
          uint32_t b = rd[j0 + pos];
          code = b & 7u;
          hq = ((int)(b >> 3) >= p.qual_cutoff) ? 1u : 0u;  // kcount.cpp:81
          if (code >= 4u) {                                // N: G inside k-mers, never a counted extension
            code = 2u;
            hq = 0u;
          }
        }
        uint32_t lo = __brev(__ballot_sync(0xffffffffu, code & 1u));
        uint32_t hi = __brev(__ballot_sync(0xffffffffu, code & 2u));
        uint32_t qq = __brev(__ballot_sync(0xffffffffu, hq));
        if (lane == 0) {
          sm.W[c] = (spread32(hi) << 1) | spread32(lo);
          sm.Q[c] = qq;
        }
      }
      if (lane < 2) {
        sm.W[nchunks + lane] = 0;
        sm.Q[nchunks + lane] = 0;
      }
      __syncwarp();

      if (need_ids) {
        // ---- (2a) canonical m-mer at every position (right aligned; order == the reference's left-aligned order)
        const int npos = nb - m + 1;
        for (int pos = lane; pos < npos; pos += 32) {
          uint64_t f = window64(sm.W, pos) >> (64 - 2 * m);
          uint64_t r = (revcomp_word(f) >> (64 - 2 * m)) & mmask;
          sm.C[pos] = f < r ? f : r;
        }
        __syncwarp();
        // ---- (2b) sliding maximum over w positions by doubling: after the loop C[p] = max C0[p .. p + span - 1]
        int span = 1;
        while (span * 2 <= w) {
          for (int base = 0; base < npos; base += 32) {
            int pos = base + (int)lane;
            uint64_t a = 0, b = 0;
            bool ok = pos + 2 * span - 1 < npos;
            if (ok) {
              a = sm.C[pos];
              b = sm.C[pos + span];
            }
            __syncwarp();
            if (ok) sm.C[pos] = a > b ? a : b;
            __syncwarp();
          }
          span *= 2;
        }
        // ---- (2c) minimizer -> owner rank and local bin
        for (int jr = lane; jr < nks; jr += 32) {
          uint64_t a = sm.C[jr + 1], b = sm.C[jr + 1 + w - span];
          uint64_t mz = (a > b ? a : b) << (64 - 2 * m);  // left aligned as Kmer::get_minimizer_fast returns it
          uint64_t h = quick_hash(mz), q;
          uint32_t owner;
          fast_divmod(h, p.div_ranks, q, owner);
          sm.ID[jr] = owner * (p.bins_mask + 1u) + ((uint32_t)q & p.bins_mask);
        }
        __syncwarp();
      }

      // ---- (2d) cut runs of equal id into records
      int carry_rs = 0;  // first k-mer of the run in progress (segment relative)
      for (int base = 0; base < nks; base += 32) {
        const int jr = base + (int)lane;
        const bool valid = jr < nks;
        uint32_t myid = 0, previd = 0xffffffffu, nextid = 0xfffffffeu;
        if (valid) {
          if (need_ids) {
            myid = sm.ID[jr];
            if (jr > 0) previd = sm.ID[jr - 1];
            if (jr + 1 < nks) nextid = sm.ID[jr + 1];
          } else {
            if (jr > 0) previd = 0;
            if (jr + 1 < nks) nextid = 0;
          }
        }
        const bool is_start = valid && myid != previd;
        const uint32_t smask = __ballot_sync(0xffffffffu, is_start);
        const uint32_t below = smask & (0xffffffffu >> (31 - lane));
        const int rs = below ? base + 31 - __clz(below) : carry_rs;
        if (smask) carry_rs = base + 31 - __clz(smask);
        const int in_run = jr - rs;
        const int piece = (int)(((uint32_t)in_run * inv_max_n) >> 16);  // in_run / max_n
        const bool is_end = valid && (myid != nextid || in_run + 1 == (piece + 1) * max_n);
        const uint32_t emask = __ballot_sync(0xffffffffu, is_end);
        if (is_end) {
          const int start = rs + piece * max_n;
          const int n = jr - start + 1;
          // group the records of this step by destination bin: one cursor update per (warp step, bin)
          const uint32_t peers = need_ids ? __match_any_sync(emask, myid) : emask;
          const int leader = __ffs(peers) - 1;
          const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
          const uint32_t cnt = __popc(peers);
          unsigned long long slot0 = 0;
          if ((int)lane == leader) {
            if (p.count_only) {
              atomicAdd(&p.bin_cursor[myid], (unsigned long long)cnt);
            } else if (FEWBINS) {
              uint32_t used = sm.resv_used[myid];
              unsigned long long rb = sm.resv_base[myid];
              if (used + cnt > RESV) {
                // retire the old reservation: its unused tail becomes empty records
                if (rb != ~0ULL)
                  for (uint32_t u = used; u < RESV; u++)
                    if (rb + u < p.bin_cap[myid]) write_empty<R>(p.bins, p.bin_base[myid], rb + u);
                rb = atomicAdd(&p.bin_cursor[myid], (unsigned long long)RESV);
                sm.resv_base[myid] = rb;
                used = 0;
              }
              sm.resv_used[myid] = used + cnt;
              slot0 = rb + used;
            } else {
              slot0 = atomicAdd(&p.bin_cursor[myid], (unsigned long long)cnt);
            }
          }
          slot0 = __shfl_sync(peers, slot0, leader);
          if (!p.count_only) {
            const unsigned long long slot = slot0 + rank;
            if (slot < p.bin_cap[myid])
              write_record<R>(sm, p.bins + p.bin_base[myid] + slot * R, start, n, k);
            else
              p.ctr->err_bins_full = 1u;
          }
          n_records++;
        }
      }
      __syncwarp();
    }
  }

  if (FEWBINS && !p.count_only) {
    // close this warp's open reservations
    __syncwarp();
    if (lane < p.n_ids) {
      unsigned long long rb = sm.resv_base[lane];
      if (rb != ~0ULL)
        for (uint32_t u = sm.resv_used[lane]; u < RESV; u++)
          if (rb + u < p.bin_cap[lane]) write_empty<R>(p.bins, p.bin_base[lane], rb + u);
    }
  }
  // statistics: one atomic per warp
  for (int o = 16; o > 0; o >>= 1) n_records += __shfl_down_sync(0xffffffffu, n_records, o);
  if (lane == 0 && !p.count_only) {
    atomicAdd(&p.ctr->parse_counted, n_counted);
    atomicAdd(&p.ctr->parse_records, n_records);
    atomicAdd(&p.ctr->parse_raw, n_raw);
  }
}

template <int R>
static cudaError_t launch_parse_r(const ParseParams &p, int grid, cudaStream_t s) {
  if (p.n_ids <= 32)
    parse_kernel<R, true><<<grid, WARPS * 32, 0, s>>>(p);
  else
    parse_kernel<R, false><<<grid, WARPS * 32, 0, s>>>(p);
  return cudaGetLastError();
}

int parse_grid_size(uint64_t n_reads, uint32_t read_stride, int sm_count) {
  uint64_t reads = (n_reads + read_stride - 1) / read_stride;
  uint64_t blocks = (reads + WARPS * 4 - 1) / (WARPS * 4);  // >= 4 reads per warp
  uint64_t cap = (uint64_t)sm_count * 8;                    // 8 resident CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}
int parse_resv_records() { return RESV; }
int parse_warps_per_block() { return WARPS; }

cudaError_t launch_parse(const ParseParams &p, int sm_count, cudaStream_t s, uint64_t *launches) {
  int grid = parse_grid_size(p.n_reads, p.read_stride, sm_count);
  (*launches)++;
  switch (p.rec_bytes) {
    case 16: return launch_parse_r<16>(p, grid, s);
    case 32: return launch_parse_r<32>(p, grid, s);
    case 64: return launch_parse_r<64>(p, grid, s);
  }
  return cudaErrorInvalidValue;
}

// per-destination header blocks: entry b of owner o = {first record of bin (o,b) inside the owner's record area,
// number of records actually present}
__global__ void write_headers_kernel(uint8_t *const *dst_header, const uint64_t *bin_first, const unsigned long long *bin_cursor,
                                     const uint64_t *bin_cap, int n_ranks, int local_bins) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_ranks * local_bins) return;
  int o = i / local_bins, b = i % local_bins;
  unsigned long long c = bin_cursor[i];
  if (c > bin_cap[i]) c = bin_cap[i];
  uint2 *hdr = reinterpret_cast<uint2 *>(dst_header[o]);
  hdr[b] = make_uint2((uint32_t)bin_first[i], (uint32_t)c);
}

cudaError_t launch_write_headers(uint8_t *const *dst_header, const uint64_t *bin_first, const unsigned long long *bin_cursor,
                                 const uint64_t *bin_cap, int n_ranks, int local_bins, cudaStream_t s, uint64_t *launches) {
  int n = n_ranks * local_bins;
  (*launches)++;
  write_headers_kernel<<<(n + 255) / 256, 256, 0, s>>>(dst_header, bin_first, bin_cursor, bin_cap, n_ranks, local_bins);
  return cudaGetLastError();
}

}  // namespace kc
