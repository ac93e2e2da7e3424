// kc_refine.cu -- second level of the supermer partition (between the exchange and the count kernel).
//
// The count kernel wants bins so small that the distinct k-mers of a bin fit a warp's shared-memory table: millions of
// bins per GPU. Scattering 32-byte records straight into millions of open bins is what HBM does worst (every bin sees a
// record every few hundred microseconds, so L2 evicts single dirty sectors instead of full lines, and per-bin capacities
// estimated from a sample are noisy). So the partition is done the radix-sort way, in two levels sized to the memory
// hierarchy:
//   level 1 (kc_parse.cu): route by owner and COARSE bin (fine bin >> fine_shift); the low bits of the fine bin travel in
//           the record header (bits 8..15). Coarse bins are few enough to be written efficiently and sampled accurately,
//           and they are the unit of the all-to-all.
//   level 2 (here): one CTA takes a coarse bin (the concatenation of every source rank's segment), stages its records in
//           shared memory (one coalesced read), histograms the sub-bin ids, and writes the records grouped by fine bin
//           into ONE region with a fine header [{first record, n records}] -- the only source the count kernel then sees.
// This replaces the receive side of the reference's supermer aggregation (src/kcount/kmer_dht.cpp:248-258: the RPC
// target appends to a per-rank buffer) -- here the "append" also sorts by fine bin.
#include <cstdlib>

#include "kc_internal.h"

namespace kc {

static const int RF_MAX_SUB = 256;
// Two launch shapes: coarse bins up to 96 KB are staged by 512-thread CTAs, two per SM; when the coarse bins are larger
// (more owners -> more fine bins per coarse bin, see configure_bins) one 1024-thread CTA per SM stages up to 208 KB.
// A coarse bin that does not fit is read twice (histogram pass, scatter pass).
static const int RF_STAGE_SMALL = 96 * 1024, RF_THREADS_SMALL = 512;
static const int RF_STAGE_LARGE = 208 * 1024, RF_THREADS_LARGE = 1024;
// ... and when they are small (one GPU: ~28 KB), four 256-thread CTAs with a 48 KB stage each share an SM: the phases of
// a bin (load, histogram, scan, scatter) are separated by barriers, and more independent CTAs keep the memory pipe busy.
static const int RF_STAGE_TINY = 48 * 1024, RF_THREADS_TINY = 256;

// ---- step A: records per coarse bin over all sources, and their exclusive prefix
__global__ void refine_totals_kernel(RefineParams p) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= p.n_coarse) return;
  unsigned long long t = 0;
  for (int s = 0; s < p.n_src; s++) {
    const uint8_t *src = p.src_ptr[s];
    if (src) t += reinterpret_cast<const uint2 *>(src)[c].y;
  }
  p.coarse_base[c] = t;
}

// single CTA: in-place exclusive scan of coarse_base[0 .. n), total to coarse_base[n] and *total
__global__ void __launch_bounds__(1024) refine_scan_kernel(RefineParams p) {
  __shared__ unsigned long long s_warp[32];
  __shared__ unsigned long long s_carry;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < p.n_coarse; base += 1024u * 4u) {
    // four consecutive entries per thread
    const uint32_t i0 = base + threadIdx.x * 4u;
    unsigned long long v[4], sum = 0;
#pragma unroll
    for (int u = 0; u < 4; u++) {
      v[u] = i0 + u < p.n_coarse ? p.coarse_base[i0 + u] : 0ULL;
      sum += v[u];
    }
    unsigned long long incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, d);
      if ((int)lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = s_warp[lane], wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, wi, d);
        if ((int)lane >= d) wi += t;
      }
      s_warp[lane] = wi - w;
    }
    __syncthreads();
    unsigned long long at = s_carry + s_warp[warp] + incl - sum;
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (i0 + u < p.n_coarse) p.coarse_base[i0 + u] = at;
      at += v[u];
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = at;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    p.coarse_base[p.n_coarse] = s_carry;
    *p.total = s_carry;
    if (s_carry > 0xffffffffULL) *p.err_too_large = 1u;
  }
}

// ---- step B
template <int R, int RF_THREADS, int RF_STAGE_BYTES>
__global__ void __launch_bounds__(RF_THREADS) refine_kernel(RefineParams p) {
  constexpr int W = R / 4;
  constexpr int U = R / 16;
  extern __shared__ __align__(16) uint8_t s_stage[];
  __shared__ uint32_t s_cnt[RF_MAX_SUB], s_cur[RF_MAX_SUB];
  __shared__ const uint8_t *s_seg_ptr[kCountMaxSrc];
  __shared__ uint32_t s_seg_start[kCountMaxSrc + 1];
  __shared__ uint32_t s_warp[RF_THREADS / 32];
  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31u, warp = tid >> 5;
  const uint32_t F = 1u << p.fine_shift;
  const uint32_t stage_cap = RF_STAGE_BYTES / R;
  uint2 *fine_hdr = reinterpret_cast<uint2 *>(p.dst);
  uint8_t *dst_rec = p.dst + p.dst_header_bytes;

  for (uint32_t c = p.coarse_begin + blockIdx.x; c < p.coarse_end; c += gridDim.x) {
    __syncthreads();  // the previous coarse bin is done with the shared state
    if (tid < p.n_src) {
      const uint8_t *src = p.src_ptr[tid];
      uint2 e = make_uint2(0u, 0u);
      if (src) e = reinterpret_cast<const uint2 *>(src)[c];
      s_seg_ptr[tid] = src ? src + p.src_header_bytes + (uint64_t)e.x * R : nullptr;
      s_seg_start[tid + 1] = src ? e.y : 0u;  // (count for now)
    }
    if (tid < RF_MAX_SUB) s_cnt[tid] = 0;
    __syncthreads();
    if (tid == 0) {
      uint32_t at = 0;
      for (int s = 0; s < p.n_src; s++) {
        const uint32_t n = s_seg_start[s + 1];
        s_seg_start[s] = at;
        at += n;
      }
      s_seg_start[p.n_src] = at;
    }
    __syncthreads();
    const uint32_t n = s_seg_start[p.n_src];
    const unsigned long long base = p.coarse_base[c];
    const bool staged = n <= stage_cap;
    // ---- pass 1: (stage and) histogram the sub-bin ids
    if (staged) {
      uint4 *st4 = reinterpret_cast<uint4 *>(s_stage);
      const uint32_t n16 = n * U;
      // four loads in flight per thread before the first one is consumed
      for (uint32_t u0 = tid; u0 < n16; u0 += 4 * RF_THREADS) {
        uint4 v[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const uint32_t u = u0 + q * RF_THREADS;
          if (u < n16) {
            const uint32_t r = u / U;
            int s = 0;
            if (p.n_src > 1)
              while (s_seg_start[s + 1] <= r) s++;
            const uint4 *g = reinterpret_cast<const uint4 *>(s_seg_ptr[s] + (uint64_t)(r - s_seg_start[s]) * R);
            v[q] = ld_stream16(g + (u % U));
          }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const uint32_t u = u0 + q * RF_THREADS;
          if (u < n16) {
            st4[u] = v[q];
            if ((int)(u % U) == U - 1) atomicAdd(&s_cnt[(v[q].w >> 8) & (F - 1u)], 1u);
          }
        }
      }
    } else {
      for (uint32_t r = tid; r < n; r += RF_THREADS) {
        int s = 0;
        while (s_seg_start[s + 1] <= r) s++;
        const uint32_t w = reinterpret_cast<const uint32_t *>(s_seg_ptr[s] + (uint64_t)(r - s_seg_start[s]) * R)[W - 1];
        atomicAdd(&s_cnt[(w >> 8) & (F - 1u)], 1u);
      }
    }
    __syncthreads();
    // ---- exclusive scan over the F sub-bins (F <= 256 = one value per thread)
    {
      const uint32_t v = (uint32_t)tid < F ? s_cnt[tid] : 0u;
      uint32_t incl = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if ((int)lane >= d) incl += t;
      }
      if (lane == 31) s_warp[warp] = incl;
      __syncthreads();
      uint32_t off = incl - v;
      for (uint32_t w = 0; w < warp; w++) off += s_warp[w];
      if ((uint32_t)tid < F) {
        s_cur[tid] = off;
        fine_hdr[((uint64_t)c << p.fine_shift) + tid] = make_uint2((uint32_t)(base + off), v);
      }
    }
    __syncthreads();
    // ---- pass 2: scatter the records to their fine bins
    if (staged) {
      const uint32_t *st = reinterpret_cast<const uint32_t *>(s_stage);
      // a group of U threads moves one record (warp-uniform trip count: the shuffle below needs every lane)
      const uint32_t n16 = n * U;
      for (uint32_t u0 = 0; u0 < n16; u0 += RF_THREADS) {
        const uint32_t u = u0 + tid;
        const bool on = u < n16;
        const uint32_t r = on ? u / U : 0u, part = u % U;
        const uint32_t hdr = st[r * W + W - 1];
        uint32_t pos = 0;
        if (on && part == 0) pos = atomicAdd(&s_cur[(hdr >> 8) & (F - 1u)], 1u);
        if (U > 1) pos = __shfl_sync(0xffffffffu, pos, (lane / U) * U);
        if (on) st_stream16(dst_rec + (base + pos) * R + 16 * part, reinterpret_cast<const uint4 *>(s_stage)[u]);
      }
    } else {
      for (uint32_t r = tid; r < n; r += RF_THREADS) {
        int s = 0;
        while (s_seg_start[s + 1] <= r) s++;
        const uint4 *g = reinterpret_cast<const uint4 *>(s_seg_ptr[s] + (uint64_t)(r - s_seg_start[s]) * R);
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = g[u];
        const uint32_t pos = atomicAdd(&s_cur[(v[U - 1].w >> 8) & (F - 1u)], 1u);
#pragma unroll
        for (int u = 0; u < U; u++) st_stream16(dst_rec + (base + pos) * R + 16 * u, v[u]);
      }
    }
  }
}

cudaError_t launch_refine_totals(const RefineParams &p, cudaStream_t s, uint64_t *launches) {
  refine_totals_kernel<<<(p.n_coarse + 255) / 256, 256, 0, s>>>(p);
  refine_scan_kernel<<<1, 1024, 0, s>>>(p);
  *launches += 2;
  return cudaGetLastError();
}

template <int R, int THREADS, int STAGE>
static cudaError_t launch_refine_shape(const RefineParams &p, int sm_count, int ctas_per_sm, cudaStream_t s) {
  uint64_t grid = (uint64_t)sm_count * ctas_per_sm;
  const uint32_t n_bins = p.coarse_end - p.coarse_begin;
  if (grid > n_bins) grid = n_bins ? n_bins : 1;
  cudaError_t e = cudaFuncSetAttribute(refine_kernel<R, THREADS, STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, STAGE);
  if (e != cudaSuccess) return e;
  refine_kernel<R, THREADS, STAGE><<<(int)grid, THREADS, STAGE, s>>>(p);
  return cudaGetLastError();
}

template <int R>
static cudaError_t launch_refine_r(const RefineParams &p, int sm_count, int shape, cudaStream_t s) {
  if (shape == 2) return launch_refine_shape<R, RF_THREADS_LARGE, RF_STAGE_LARGE>(p, sm_count, 1, s);
  if (shape == 1) return launch_refine_shape<R, RF_THREADS_SMALL, RF_STAGE_SMALL>(p, sm_count, 2, s);
  return launch_refine_shape<R, RF_THREADS_TINY, RF_STAGE_TINY>(p, sm_count, 4, s);
}

// avg_records: records per coarse bin (all sources), picks the launch shape
cudaError_t launch_refine(const RefineParams &p_in, uint64_t avg_records, int sm_count, cudaStream_t s, uint64_t *launches) {
  RefineParams p = p_in;
  if (p.coarse_begin == 0 && p.coarse_end == 0) p.coarse_end = p.n_coarse;
  if (p.coarse_end <= p.coarse_begin) return cudaSuccess;
  if (p.n_src > kCountMaxSrc || (1u << p.fine_shift) > (uint32_t)RF_MAX_SUB) return cudaErrorInvalidValue;
  const uint64_t want = avg_records * (uint64_t)p.rec_bytes * 3 / 2;  // average bin + 50 %
  int shape = want > (uint64_t)RF_STAGE_SMALL ? 2 : (want > (uint64_t)RF_STAGE_TINY ? 1 : 0);
  if (const char *e = getenv("KC_REFINE_SHAPE")) shape = atoi(e);  // (experiments)
  (*launches)++;
  switch (p.rec_bytes) {
    case 16: return launch_refine_r<16>(p, sm_count, shape, s);
    case 32: return launch_refine_r<32>(p, sm_count, shape, s);
    case 64: return launch_refine_r<64>(p, sm_count, shape, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace kc
