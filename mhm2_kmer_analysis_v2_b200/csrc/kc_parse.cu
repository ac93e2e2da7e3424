// kc_parse.cu -- kernels (1) "read parse + 2-bit pack" and (2) "minimizer / supermer extraction".
//
// Replaces, for the path of BASELINE.json: count_kmers' per-read unpack + quality mask (src/kcount/kcount.cpp:71-90),
// SeqBlockInserter::process_seq (src/kcount/kcount_cpu.cpp:74-103) and the reference's GPU kernels parse_and_pack /
// build_supermers / pack_seqs (src/kcount/kcount-gpu/parse_and_pack.cpp:127-237).
//
// A warp takes 32 consecutive reads, one per lane.
//   (1) pack: every lane reads its bytes (base | qual << 3, src/packed_reads.cpp:99-126) with aligned 128-bit loads and
//       turns 16 bytes into one 32-bit word of 2-bit codes (Kmer word layout, src/kmer.cpp:303-325) plus 16 "high
//       quality" bits, kept in shared memory. N becomes (G, low quality): inside a k-mer the reference maps N to G
//       (src/kmer.cpp:173,191-192) and as an extension an N is never counted (kcount_cpu.cpp:157-164).
//   (2) scan: the lane walks its read once with a rolling forward and reverse-complement m-mer; the reference's
//       "greatest least-complement m-mer" minimizer (src/kmer.cpp:345-398) of every k-mer is a sliding-window maximum,
//       computed with the block prefix/suffix-maximum scheme (all lanes stay in lock step, no data-dependent loops).
//       A run of k-mers with the same minimizer (cut to the record capacity) is one supermer.
//   (3) emit: finished runs go to a small per-warp queue; whenever 32 are queued every lane takes one: owner =
//       quick_hash(minimizer) % n_ranks (src/kcount/kmer_dht.cpp:193-196), local bin = mulhi(mix(hash / n_ranks), bins), one
//       cursor update per distinct bin in the batch, and the fixed-size record is cut out of the owning lane's packed
//       words with funnel shifts and written with 128-bit stores into the per-(owner, bin) area of the send buffer.
#include <algorithm>
#include "kc_internal.h"

namespace kc {

static const int SEG = 160;                            // bases of a read handled per lane and pass
static const int NW = (SEG + 16) / 16 + 2;             // packed base words per lane (16 bases each): lead (< 16) + segment + 1 look-ahead
static const int NQ = (SEG + 16) / 32 + 2;             // quality-bit words per lane
static const int QCAP = 64;                            // queue entries per warp

// 4 PackedRead bytes (first base in the lowest byte) -> 8 bits of 2-bit codes (first base highest) + 4 quality bits
__device__ __forceinline__ void pack4(uint32_t x, uint32_t cut_add, uint32_t &codes8, uint32_t &hq4) {
  const uint32_t is_n = (x >> 2) & 0x01010101u;          // base code 4 = N
  const uint32_t lo2 = (x & 0x03030303u) | (is_n << 1);  // N -> G (2)
  const uint32_t q = (x >> 3) & 0x1f1f1f1fu;
  const uint32_t hq = ((q + cut_add) >> 5) & 0x01010101u & ~is_n;  // q >= cutoff, and never for N
  codes8 = (lo2 * 0x40100401u) >> 24;
  hq4 = ((hq * 0x08040201u) >> 24) & 0xfu;
}

struct WarpQueue {
  uint64_t val[QCAP];   // minimizer (right aligned)
  uint32_t meta[QCAP];  // owner lane | lead << 5 | first k-mer << 9 | k-mers << 17
};

// build the record for k-mers [start, start + n) of the segment held by lane `owner` and store it at dst
template <int R>
__device__ __forceinline__ void write_record(const uint32_t *s_w, const uint32_t *s_q, int owner, int lead, int start, int n, int k,
                                             uint32_t sub, uint8_t *dst) {
  constexpr int W = R / 4;
  constexpr int NBW = (R == 16 ? 32 : (R == 32 ? 80 : 160)) / 16;
  constexpr int NQW = W - NBW - (R == 32 ? 0 : 1);  // R=32: third quality word shares the header word
  const int nbases = n + k + 1;
  const int p0 = lead + start;  // position of the record's first base in the lane's packed stream
  uint32_t w[W];
  {
    const int q = p0 >> 4, sh = (p0 & 15) * 2;
    uint32_t prev = s_w[q * 32 + owner];
#pragma unroll
    for (int u = 0; u < NBW; u++) {
      const int valid = nbases - 16 * u;
      uint32_t x = 0;
      if (valid > 0) {
        const uint32_t next = s_w[(q + u + 1) * 32 + owner];
        x = __funnelshift_l(next, prev, sh);
        prev = next;
        if (valid < 16) x &= ~(0xffffffffu >> (2 * valid));
      }
      w[u] = x;
    }
  }
  {
    const int q = p0 >> 5, sh = p0 & 31;
    uint32_t prev = s_q[q * 32 + owner];
#pragma unroll
    for (int v = 0; v < NQW; v++) {
      const int valid = nbases - 32 * v;
      uint32_t x = 0;
      if (valid > 0) {
        const uint32_t next = s_q[(q + v + 1) * 32 + owner];
        x = __funnelshift_l(next, prev, sh);
        prev = next;
        if (valid < 32) x &= ~(0xffffffffu >> valid);
      }
      w[NBW + v] = x;
    }
  }
  // header: k-mers in bits 0..7, sub-bin (refined by kc_refine.cu) in bits 8..15
  if (R == 32)
    w[W - 1] = (w[W - 1] & 0xffff0000u) | (uint32_t)n | (sub << 8);
  else
    w[W - 1] = (uint32_t)n | (sub << 8);
#pragma unroll
  for (int i = 0; i < W / 4; i++) st_stream16(dst + 16 * i, make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]));
}

// every lane < nd takes one queued run: bin id, slot in the bin, record
template <int R>
__device__ __forceinline__ void drain(const ParseParams &p, const uint32_t *s_w, const uint32_t *s_q, WarpQueue &Q, uint32_t &q_head,
                                      uint32_t &q_count, unsigned long long &n_records) {
  const uint32_t lane = lane_id();
  const uint32_t nd = min(32u, q_count);
  __syncwarp();
  const bool mine = lane < nd;
  uint32_t id = 0, meta = 0, sub = 0;
  if (mine) {
    const uint32_t e = (q_head + lane) & (QCAP - 1);
    meta = Q.meta[e];
    if (p.n_ids > 1) {
      const uint64_t mz = Q.val[e] << (64 - 2 * p.m);  // left aligned as Kmer::get_minimizer_fast returns it
      uint64_t h = quick_hash(mz), q;
      uint32_t owner;
      fast_divmod(h, p.div_ranks, q, owner);
      uint32_t x = ((uint32_t)q ^ (uint32_t)(q >> 32)) * 0x9e3779b1u;  // the high bits pick the bin
      x ^= x >> 15;
      // fine bin of the owner = mulhi(mix, coarse bins << fine_shift); the parse routes by the coarse bin, the low bits travel
      // in the record header and are applied by the refine pass
      const uint32_t fine = __umulhi(x * 0x85ebca6bu, p.local_bins << p.fine_shift);
      id = owner * p.local_bins + (fine >> p.fine_shift);
      sub = fine & ((1u << p.fine_shift) - 1u);
    }
  }
  // One cursor update per record. (Merging the equal bins of a batch with match.any first costs 65 SM cycles per warp on
  // B200 -- more than the few atomics it saves with hundreds of thousands of bins.)
  // With few bins the batch is merged first (one atomic per distinct bin; a single bin would serialise 32 atomics).
  // (the bin's base and capacity are requested before the cursor atomic, so that the two round trips overlap)
  unsigned long long b_cap = 0, b_base = 0;
  if (mine && !p.count_only) {
    b_cap = __ldg(p.bin_cap + id);
    b_base = __ldg(p.bin_base + id);
  }
  unsigned long long slot0 = 0;
  uint32_t peers = 1u << lane;
  if (p.n_ids >= 4096u) {
    if (mine) slot0 = atomicAdd(&p.bin_cursor[id], 1ULL);
  } else {
    peers = p.n_ids > 1 ? __match_any_sync(0xffffffffu, mine ? id : (0x80000000u | lane)) : __ballot_sync(0xffffffffu, mine);
    const int leader = mine ? __ffs(peers) - 1 : (int)lane;
    if (mine && (int)lane == leader) slot0 = atomicAdd(&p.bin_cursor[id], (unsigned long long)__popc(peers));
    slot0 = __shfl_sync(0xffffffffu, slot0, leader);
  }
  if (mine) {
    if (!p.count_only) {
      const unsigned long long slot = slot0 + __popc(peers & ((1u << lane) - 1u));
      if (slot < b_cap)
        write_record<R>(s_w, s_q, meta & 31u, (meta >> 5) & 15u, (meta >> 9) & 255u, (meta >> 17) & 255u, p.k, sub,
                        p.bins + b_base + slot * R);
      else
        p.ctr->err_bins_full = 1u;
    }
    n_records++;
  }
  q_head = (q_head + nd) & (QCAP - 1);
  q_count -= nd;
  __syncwarp();
}

template <int R>
__global__ void __launch_bounds__(256, 4) parse_kernel(ParseParams p, int vh_len) {
  extern __shared__ __align__(16) uint8_t s_raw[];
  const uint32_t lane = lane_id();
  const int warp = threadIdx.x >> 5, n_warps_cta = blockDim.x >> 5;
  // per warp: packed bases [NW][32] u32, quality bits [NQ][32] u32, window maxima [vh_len][32] u64, queue
  const size_t per_warp = (size_t)(NW + NQ) * 32 * 4 + (size_t)vh_len * 32 * 8 + sizeof(WarpQueue);
  uint8_t *base_ptr = s_raw + warp * per_warp;
  uint64_t *s_vh = reinterpret_cast<uint64_t *>(base_ptr);
  WarpQueue &Q = *reinterpret_cast<WarpQueue *>(base_ptr + (size_t)vh_len * 32 * 8);
  uint32_t *s_w = reinterpret_cast<uint32_t *>(base_ptr + (size_t)vh_len * 32 * 8 + sizeof(WarpQueue));
  uint32_t *s_q = s_w + NW * 32;

  const int k = p.k, m = p.m;
  const int w = k - m + 1;                       // minimizer window: m-mers per k-mer
  const int max_n = record_geom(R, k).max_kmers;  // counted k-mers per record
  const int seg_kmers = SEG - k - 1;             // counted k-mers per segment
  const bool need_ids = p.n_ids > 1;
  const uint64_t mmask = (1ULL << (2 * m)) - 1ULL;  // m <= 28
  const uint32_t cut_add = (uint32_t)(32 - p.qual_cutoff) * 0x01010101u;

  unsigned long long n_counted = 0, n_records = 0, n_raw = 0;
  uint32_t q_head = 0, q_count = 0;
  const uint64_t warp_global = (uint64_t)blockIdx.x * n_warps_cta + warp;
  const uint64_t n_warps = (uint64_t)gridDim.x * n_warps_cta;
  const uint64_t n_sel = (p.n_reads + p.read_stride - 1) / p.read_stride;  // reads this launch looks at

  for (uint64_t batch = warp_global * 32; batch < n_sel; batch += n_warps * 32) {
    const uint64_t sel = batch + lane;
    int nk = 0;
    const uint8_t *rd = nullptr;
    if (sel < n_sel) {
      const uint64_t ridx = sel * p.read_stride;
      const int L = p.lens[ridx];
      if (L >= k) {  // kcount.cpp:78
        n_raw += L - k + 1;
        nk = max(L - k - 1, 0);  // counted (interior) k-mers: kcount_cpu.cpp:86
        n_counted += nk;
        rd = p.bytes + p.offsets[ridx];
      }
    }
    const int n_seg = (nk + seg_kmers - 1) / seg_kmers;
    const int max_seg = __reduce_max_sync(0xffffffffu, n_seg);
    for (int sg = 0; sg < max_seg; sg++) {
      const int j0 = sg * seg_kmers;
      const int nks = sg < n_seg ? min(seg_kmers, nk - j0) : 0;  // counted k-mers of this lane's segment
      const int nb = nks ? nks + k + 1 : 0;                      // bases [j0, j0 + nb)
      const uint8_t *src = rd + j0;
      const int lead = nks ? (int)((uintptr_t)src & 15) : 0;
      // ---- (1) pack: aligned 16-byte chunks -> 32 bits of codes + 16 quality bits
      {
        const int nch = nks ? (lead + nb + 15) >> 4 : 0;
        const int max_ch = __reduce_max_sync(0xffffffffu, nch);
        const uint4 *a0 = reinterpret_cast<const uint4 *>(src - lead);
        // four chunks per trip, all four loads issued before the first is used (the reads stream from DRAM: a lane
        // that waits for every chunk in turn pays the full latency eleven times per segment)
        for (int c = 0; c < max_ch; c += 4) {
          uint4 v[4];
#pragma unroll
          for (int u = 0; u < 4; u++) v[u] = c + u < nch ? __ldg(a0 + c + u) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int u = 0; u < 4; u += 2) {
            if (c + u < nch) {
              uint32_t h16[2];
#pragma unroll
              for (int e = 0; e < 2; e++) {
                uint32_t c0, c1, c2, c3, h0, h1, h2, h3;
                pack4(v[u + e].x, cut_add, c0, h0);
                pack4(v[u + e].y, cut_add, c1, h1);
                pack4(v[u + e].z, cut_add, c2, h2);
                pack4(v[u + e].w, cut_add, c3, h3);
                if (c + u + e < nch) s_w[(c + u + e) * 32 + lane] = (c0 << 24) | (c1 << 16) | (c2 << 8) | c3;
                h16[e] = (h0 << 12) | (h1 << 8) | (h2 << 4) | h3;
              }
              s_q[((c + u) >> 1) * 32 + lane] = (h16[0] << 16) | (c + u + 1 < nch ? h16[1] : 0u);
            }
          }
        }
      }
      __syncwarp();
      // ---- (2) scan
      if (!need_ids) {
        // one owner, one bin: the segment is cut into records of max_n k-mers
        const int n_pieces = (nks + max_n - 1) / max_n;
        const int max_pieces = __reduce_max_sync(0xffffffffu, n_pieces);
        for (int r = 0; r < max_pieces; r++) {
          const bool want = r < n_pieces;
          const uint32_t mask = __ballot_sync(0xffffffffu, want);
          if (want) {
            const uint32_t e = (q_head + q_count + __popc(mask & ((1u << lane) - 1u))) & (QCAP - 1);
            const int start = r * max_n, n = min(max_n, nks - start);
            Q.val[e] = 0;
            Q.meta[e] = lane | ((uint32_t)lead << 5) | ((uint32_t)start << 9) | ((uint32_t)n << 17);
          }
          q_count += __popc(mask);
          if (q_count >= 32) drain<R>(p, s_w, s_q, Q, q_head, q_count, n_records);
        }
      } else {
        const int steps = nks ? nb - 1 : 0;  // the last base is only an extension
        const int max_steps = __reduce_max_sync(0xffffffffu, steps);
        uint64_t fw = 0, rc = 0, pref = 0, run_val = 0;
        uint32_t cw = 0;
        int t = 0, run_start = 0, run_n = 0;
        // (The run state is updated by selects AFTER the finished run has been queued: with the bookkeeping inside the
        // branches the compiler carried seven live values through every path -- a third of the loop was register moves.)
        for (int i = 0; i <= max_steps; i++) {
          bool kmer_here = false;  // a counted k-mer ends at this base
          uint64_t wmax = 0;
          if (i < steps) {
            const int pos = lead + i;
            if ((pos & 15) == 0 || i == 0) cw = s_w[(pos >> 4) * 32 + lane];
            const uint64_t b = (cw >> (30 - 2 * (pos & 15))) & 3u;
            fw = ((fw << 2) | b) & mmask;
            rc = (rc >> 2) | ((3ULL - b) << (2 * m - 2));
            if (i >= m - 1) {
              const uint64_t canon = fw < rc ? fw : rc;
              pref = (t == 0 || canon > pref) ? canon : pref;
              s_vh[t * 32 + lane] = canon;
              if (i >= k - 1) {
                // k-mer starting at base s = i - k + 1: maximum of the last w canonical m-mers
                wmax = pref;
                if (t != w - 1) {
                  const uint64_t suf = s_vh[(t + 1) * 32 + lane];
                  wmax = suf > pref ? suf : pref;
                }
                kmer_here = i >= k;  // s == 0 is the first k-mer of the read (or belongs to the previous segment)
              }
              if (t == w - 1) {
                // block of w m-mers complete: turn it into suffix maxima for the windows that start inside it
                uint64_t run = canon;
                for (int u = w - 2; u >= 0; u--) {
                  const uint64_t x = s_vh[u * 32 + lane];
                  run = x > run ? x : run;
                  s_vh[u * 32 + lane] = run;
                }
                t = 0;
              } else {
                t++;
              }
            }
          }
          const bool open = run_n != 0;
          const bool cut = kmer_here && open && (wmax != run_val || run_n == max_n);  // the k-mer starts another run
          const bool closing = i == steps && open;                                     // end of the segment
          const bool want = cut || closing;
          const uint32_t mask = __ballot_sync(0xffffffffu, want);
          if (mask) {
            if (want) {
              const uint32_t e = (q_head + q_count + __popc(mask & ((1u << lane) - 1u))) & (QCAP - 1);
              Q.val[e] = run_val;
              Q.meta[e] = lane | ((uint32_t)lead << 5) | ((uint32_t)run_start << 9) | ((uint32_t)run_n << 17);
            }
            q_count += __popc(mask);
            if (q_count >= 32) drain<R>(p, s_w, s_q, Q, q_head, q_count, n_records);
          }
          const bool restart = kmer_here && (cut || !open);
          run_n = restart ? 1 : (closing ? 0 : run_n + (int)kmer_here);
          run_start = restart ? i - k : run_start;  // (record-relative number of the k-mer: s - 1)
          run_val = restart ? wmax : run_val;
        }
      }
      // the packed words are overwritten by the next segment: everything queued must be written first
      while (q_count) drain<R>(p, s_w, s_q, Q, q_head, q_count, n_records);
      __syncwarp();
    }
  }

  // statistics: one atomic per warp
  for (int o = 16; o > 0; o >>= 1) {
    n_records += __shfl_down_sync(0xffffffffu, n_records, o);
    n_counted += __shfl_down_sync(0xffffffffu, n_counted, o);
    n_raw += __shfl_down_sync(0xffffffffu, n_raw, o);
  }
  if (lane == 0) {
    atomicAdd(&p.ctr->parse_records, n_records);  // (a count-only pass reports the records it saw: the layout's average bin)
    if (!p.count_only) {
      atomicAdd(&p.ctr->parse_counted, n_counted);
      atomicAdd(&p.ctr->parse_raw, n_raw);
    }
  }
}

struct ParseLaunch {
  int grid, threads, vh_len;
  size_t smem;
};

static ParseLaunch parse_launch_config(const ParseParams &p, int sm_count) {
  ParseLaunch L;
  L.vh_len = p.n_ids > 1 ? (p.k - p.m + 1) : 1;
  const size_t per_warp = (size_t)(NW + NQ) * 32 * 4 + (size_t)L.vh_len * 32 * 8 + sizeof(WarpQueue);
  int warps = per_warp * 8 <= 100 * 1024 ? 8 : 4;
  L.threads = warps * 32;
  L.smem = per_warp * warps;
  int per_sm = (int)((220 * 1024) / (L.smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  const uint64_t n_sel = (p.n_reads + p.read_stride - 1) / p.read_stride;
  uint64_t blocks = (n_sel + (uint64_t)warps * 32 - 1) / ((uint64_t)warps * 32);
  const uint64_t cap = (uint64_t)sm_count * per_sm;
  L.grid = (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
  return L;
}

template <int R>
static cudaError_t launch_parse_r(const ParseParams &p, int sm_count, cudaStream_t s) {
  ParseLaunch L = parse_launch_config(p, sm_count);
  cudaFuncSetAttribute(parse_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem);
  int resident = 0;  // the grid is persistent: exactly the CTAs that fit (registers may allow fewer than shared memory)
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, parse_kernel<R>, L.threads, L.smem) == cudaSuccess && resident > 0)
    L.grid = std::min(L.grid, resident * sm_count);
  parse_kernel<R><<<L.grid, L.threads, L.smem, s>>>(p, L.vh_len);
  return cudaGetLastError();
}

cudaError_t launch_parse(const ParseParams &p, int sm_count, cudaStream_t s, uint64_t *launches) {
  (*launches)++;
  switch (p.rec_bytes) {
    case 16: return launch_parse_r<16>(p, sm_count, s);
    case 32: return launch_parse_r<32>(p, sm_count, s);
    case 64: return launch_parse_r<64>(p, sm_count, s);
  }
  return cudaErrorInvalidValue;
}

// per-destination header blocks: entry b of owner o = {first record of bin (o,b) inside the owner's record area,
// number of records actually present}
__global__ void write_headers_kernel(uint8_t *const *dst_header, const uint64_t *bin_first, const unsigned long long *bin_cursor,
                                     const uint64_t *bin_cap, int n_ranks, int local_bins, unsigned long long *owner_used) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_ranks * local_bins) return;
  int o = i / local_bins, b = i % local_bins;
  unsigned long long c = bin_cursor[i];
  if (c > bin_cap[i]) c = bin_cap[i];
  uint2 *hdr = reinterpret_cast<uint2 *>(dst_header[o]);
  hdr[b] = make_uint2((uint32_t)bin_first[i], (uint32_t)c);
  // records up to the end of the last non-empty bin: the slack behind it need not travel in the exchange
  if (c) atomicMax(owner_used + o, (unsigned long long)bin_first[i] + c);
}

cudaError_t launch_write_headers(uint8_t *const *dst_header, const uint64_t *bin_first, const unsigned long long *bin_cursor,
                                 const uint64_t *bin_cap, int n_ranks, int local_bins, unsigned long long *owner_used, cudaStream_t s,
                                 uint64_t *launches) {
  int n = n_ranks * local_bins;
  (*launches)++;
  write_headers_kernel<<<(n + 255) / 256, 256, 0, s>>>(dst_header, bin_first, bin_cursor, bin_cap, n_ranks, local_bins, owner_used);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ bin layout on the device
// Capacities from the (sampled or exact) cursors and their exclusive scan -> first record of every bin inside its
// owner's record area. No per-bin data crosses to the host (there may be millions of bins).
//   1. caps + scan inside blocks of 1024 bins (grid = blocks per owner x owners), block totals
//   2. one thread: scan of the block totals per owner, owner totals, region offsets
//   3. add the block offsets; byte offset of every bin
__global__ void __launch_bounds__(1024) layout_caps_kernel(LayoutParams p) {
  __shared__ unsigned long long s_warp[32];
  const uint32_t o = blockIdx.y, P = p.local_bins;
  const uint32_t b = blockIdx.x * 1024u + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  unsigned long long cap = 0;
  if (b < P) {
    const double cnt = (double)p.bin_cursor[o * P + b];
    double est = cnt * p.scale;
    if (p.scale > 1.0) {
      est += 7.0 * p.scale * sqrt(cnt + 1.0) + 0.03 * est;  // sampling noise of the sample
      // bins are filled by a hash, so a bin whose sample came out far BELOW the average is most likely just unlucky:
      // never go under 1.25 x the average bin (a 7-sigma margin on 3 sampled records is no protection at all)
      const double avg = (double)*p.sampled_total * p.scale / ((double)P * p.n_ranks);
      if (est < 1.25 * avg) est = 1.25 * avg;
    }
    cap = (unsigned long long)est + p.pad;
    p.bin_cap[o * P + b] = cap;
  }
  unsigned long long incl = cap;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, d);
    if ((int)lane >= d) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = s_warp[lane], wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xffffffffu, wi, d);
      if ((int)lane >= d) wi += v;
    }
    s_warp[lane] = wi - w;
    if (lane == 31) p.block_sums[o * gridDim.x + blockIdx.x] = wi;
  }
  __syncthreads();
  if (b < P) p.bin_first[o * P + b] = s_warp[warp] + incl - cap;
}

__global__ void layout_regions_kernel(LayoutParams p, uint32_t blocks_per_owner) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  unsigned long long at = 0;
  for (uint32_t o = 0; o < p.n_ranks; o++) {
    unsigned long long rec = 0;
    for (uint32_t j = 0; j < blocks_per_owner; j++) {
      const unsigned long long v = p.block_sums[o * blocks_per_owner + j];
      p.block_sums[o * blocks_per_owner + j] = rec;
      rec += v;
    }
    p.owner_records[o] = rec;
    if (rec > 0xffffffffULL) *p.err_too_large = 1u;  // headers carry 32-bit record indices
    p.region_off[o] = at;
    at += p.header_bytes + rec * p.rec_bytes;
    at = (at + 255ULL) & ~255ULL;
    p.region_bytes[o] = at - p.region_off[o];
  }
  p.region_off[p.n_ranks] = at;  // total
}

__global__ void layout_bases_kernel(LayoutParams p, uint32_t blocks_per_owner) {
  const uint32_t o = blockIdx.y;
  const uint32_t b = blockIdx.x * 1024u + threadIdx.x;
  if (b >= p.local_bins) return;
  const uint32_t i = o * p.local_bins + b;
  const unsigned long long first = p.bin_first[i] + p.block_sums[o * blocks_per_owner + blockIdx.x];
  p.bin_first[i] = first;
  p.bin_base[i] = p.region_off[o] + p.header_bytes + first * p.rec_bytes;
}

// where every slice of coarse bins starts inside each owner's region (edge n_slices = end of the record area by
// capacity; the caller clips it to the used bytes)
__global__ void slice_offsets_kernel(const uint64_t *bin_first, const uint64_t *bin_cap, uint32_t n_ranks, uint32_t P, int n_slices,
                                     int header_bytes, int rec_bytes, unsigned long long *out) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t per = (uint32_t)n_slices + 1u;
  if (t >= n_ranks * per) return;
  const uint32_t o = t / per, i = t % per;
  const uint32_t e = slice_edge(P, (int)i, n_slices);
  unsigned long long first;
  if (e < P) first = bin_first[(uint64_t)o * P + e];
  else first = bin_first[(uint64_t)o * P + P - 1] + (bin_cap ? bin_cap[(uint64_t)o * P + P - 1] : 0ULL);
  out[t] = (unsigned long long)header_bytes + first * (unsigned long long)rec_bytes;
}

cudaError_t launch_slice_offsets(const uint64_t *bin_first, const uint64_t *bin_cap, uint32_t n_ranks, uint32_t local_bins, int n_slices,
                                 int header_bytes, int rec_bytes, unsigned long long *out, cudaStream_t s, uint64_t *launches) {
  const uint32_t n = n_ranks * ((uint32_t)n_slices + 1u);
  slice_offsets_kernel<<<(n + 127) / 128, 128, 0, s>>>(bin_first, bin_cap, n_ranks, local_bins, n_slices, header_bytes, rec_bytes, out);
  (*launches)++;
  return cudaGetLastError();
}

cudaError_t launch_layout(const LayoutParams &p, cudaStream_t s, uint64_t *launches) {
  (*launches) += 3;
  const uint32_t nb = layout_blocks_per_owner(p.local_bins);
  layout_caps_kernel<<<dim3(nb, p.n_ranks), 1024, 0, s>>>(p);
  layout_regions_kernel<<<1, 32, 0, s>>>(p, nb);
  layout_bases_kernel<<<dim3(nb, p.n_ranks), 1024, 0, s>>>(p, nb);
  return cudaGetLastError();
}

}  // namespace kc
