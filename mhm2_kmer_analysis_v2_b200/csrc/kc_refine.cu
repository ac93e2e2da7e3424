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


// SM-partitioned pipeline: every thread's writes of coarse bin c are made visible device-wide, then one thread counts the
// bin into its slice (the count kernel, running beside this one, acquires the counter before it touches the slice)
__device__ __forceinline__ void announce_bin(const RefineParams &p, uint32_t c) {
  if (p.n_slices == 0) return;
  __syncthreads();  // every thread's stores of this bin happen before thread 0's fence + atomic (cumulative release)
  if (threadIdx.x == 0) {
    int sl = 0;
    while (sl + 1 < p.n_slices && c >= p.slice_coarse_edge[sl + 1]) sl++;
    __threadfence();
    atomicAdd(&p.slice_done[sl], 1u);
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
  if (p.n_slices && tid == 0) atomicAdd(p.slice_done + kSlicesMax, 1u);  // "this CTA is resident" (gate_kernel)

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
    announce_bin(p, c);
  }
}

// ---- bulk asynchronous copies (sm_90+ / sm_100a: cp.async.bulk + mbarrier; SASS UBLKCP / SYNCS) ----------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// global (local HBM or a peer's, mapped) -> this CTA's shared memory; completion is counted in bytes on `bar`
__device__ __forceinline__ void bulk_load(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---- step B, bulk-copy variant: the coarse bins of a CTA are DOUBLE BUFFERED. While the CTA histograms and scatters bin i
// out of one stage buffer, the segments of bin i + 1 (one contiguous run per source rank, local or in a peer's memory) are
// already arriving in the other one, moved by the copy unit (cp.async.bulk -> UBLKCP) and signalled through an mbarrier:
// no thread issues a load or a shared-memory store for the staging any more (the LDG -> STS variant below spent 18 % of its
// stall samples on the STS waiting for its load and 25 % at the barrier behind it, profiles/r01_count_kernel_warp.md).
// Warp 0 is the producer: lane s owns source s, keeps that source's header entries of the next two bins in registers, and
// issues the copy of its segment. A coarse bin that does not fit a stage buffer takes the two-pass path from global memory.
template <int R, int RF_THREADS, int RF_STAGE_BYTES>
__global__ void __launch_bounds__(RF_THREADS) refine_bulk_kernel(RefineParams p) {
  constexpr int W = R / 4;
  constexpr int U = R / 16;
  extern __shared__ __align__(128) uint8_t s_dyn[];
  __shared__ uint32_t s_cnt[RF_MAX_SUB], s_cur[RF_MAX_SUB];
  __shared__ const uint8_t *s_seg_ptr[2][kCountMaxSrc];
  __shared__ uint32_t s_seg_start[2][kCountMaxSrc + 1];
  __shared__ unsigned long long s_base[2];
  __shared__ uint32_t s_warp[RF_THREADS / 32];
  __shared__ __align__(8) uint64_t s_bar[2];
  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31u, warp = tid >> 5;
  const uint32_t F = 1u << p.fine_shift;
  const uint32_t stage_cap = RF_STAGE_BYTES / R;
  uint2 *fine_hdr = reinterpret_cast<uint2 *>(p.dst);
  uint8_t *dst_rec = p.dst + p.dst_header_bytes;
  uint8_t *stage[2] = {s_dyn, s_dyn + RF_STAGE_BYTES};

  if (p.n_slices && tid == 0) atomicAdd(p.slice_done + kSlicesMax, 1u);  // "this CTA is resident" (gate_kernel)
  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  // producer state (warp 0): header entry of source `lane` (and lane + 32) for the bin about to be issued
  const uint32_t c_first = p.coarse_begin + blockIdx.x, c_step = gridDim.x;
  auto load_entry = [&](uint32_t c, int s) -> uint2 {
    if (c >= p.coarse_end || s >= p.n_src) return make_uint2(0u, 0u);
    const uint8_t *src = p.src_ptr[s];
    return src ? reinterpret_cast<const uint2 *>(src)[c] : make_uint2(0u, 0u);
  };
  // issue the copies of coarse bin c into buffer b (warp 0, all lanes); e0 / e1 = this lane's entries for sources lane, lane + 32
  auto issue = [&](uint32_t c, int b, uint2 e0, uint2 e1) {
    const int s0 = (int)lane, s1 = (int)lane + 32;
    const uint8_t *b0 = s0 < p.n_src ? p.src_ptr[s0] : nullptr, *b1 = s1 < p.n_src ? p.src_ptr[s1] : nullptr;
    const uint32_t n0 = b0 ? e0.y : 0u, n1 = b1 ? e1.y : 0u;
    // exclusive scan of the segment lengths over the sources (lanes, then the second set)
    uint32_t i0 = n0, i1 = n1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t0 = __shfl_up_sync(0xffffffffu, i0, o), t1 = __shfl_up_sync(0xffffffffu, i1, o);
      if ((int)lane >= o) i0 += t0, i1 += t1;
    }
    const uint32_t tot0 = __shfl_sync(0xffffffffu, i0, 31), tot1 = __shfl_sync(0xffffffffu, i1, 31);
    const uint32_t st0 = i0 - n0, st1 = tot0 + i1 - n1, total = tot0 + tot1;
    const uint8_t *p0 = b0 ? b0 + p.src_header_bytes + (uint64_t)e0.x * R : nullptr;
    const uint8_t *p1 = b1 ? b1 + p.src_header_bytes + (uint64_t)e1.x * R : nullptr;
    if (s0 < p.n_src) s_seg_ptr[b][s0] = p0, s_seg_start[b][s0] = st0;
    if (s1 < p.n_src) s_seg_ptr[b][s1] = p1, s_seg_start[b][s1] = st1;
    const bool staged = total <= stage_cap;
    if (lane == 0) {
      s_seg_start[b][p.n_src] = total;
      s_base[b] = p.coarse_base[c];
      mbar_arrive_expect_tx(&s_bar[b], staged ? total * (uint32_t)R : 0u);
    }
    __syncwarp();
    if (staged) {
      if (n0) bulk_load(stage[b] + (size_t)st0 * R, p0, n0 * (uint32_t)R, &s_bar[b]);
      if (n1) bulk_load(stage[b] + (size_t)st1 * R, p1, n1 * (uint32_t)R, &s_bar[b]);
    }
  };

  uint2 e_a0 = make_uint2(0u, 0u), e_a1 = e_a0, e_b0 = e_a0, e_b1 = e_a0;  // entries of the next bin to issue / the one after
  if (warp == 0) {
    e_a0 = load_entry(c_first, (int)lane), e_a1 = load_entry(c_first, (int)lane + 32);
    e_b0 = load_entry(c_first + c_step, (int)lane), e_b1 = load_entry(c_first + c_step, (int)lane + 32);
    if (c_first < p.coarse_end) issue(c_first, 0, e_a0, e_a1);
  }
  uint32_t it = 0;
  for (uint32_t c = c_first; c < p.coarse_end; c += c_step, it++) {
    const int b = (int)(it & 1u);
    if (warp == 0) {
      // the next bin of this CTA goes into the other buffer (everybody left it at the barrier that ended the last round)
      const uint32_t cn = c + c_step;
      const uint2 n0 = e_b0, n1 = e_b1;
      e_b0 = load_entry(cn + c_step, (int)lane), e_b1 = load_entry(cn + c_step, (int)lane + 32);
      if (cn < p.coarse_end) issue(cn, b ^ 1, n0, n1);
    }
    if (tid < RF_MAX_SUB) s_cnt[tid] = 0;
    __syncthreads();  // s_cnt cleared; (first round) the producer's shared state of buffer b is visible
    const uint32_t n = s_seg_start[b][p.n_src];
    const unsigned long long base = s_base[b];
    const bool staged = n <= stage_cap;
    mbar_wait(&s_bar[b], (it >> 1) & 1u);
    const uint32_t *st = reinterpret_cast<const uint32_t *>(stage[b]);
    // ---- pass 1: histogram of the sub-bin ids
    if (staged) {
      for (uint32_t r = tid; r < n; r += RF_THREADS) atomicAdd(&s_cnt[(st[r * W + W - 1] >> 8) & (F - 1u)], 1u);
    } else {
      for (uint32_t r = tid; r < n; r += RF_THREADS) {
        int s = 0;
        while (s_seg_start[b][s + 1] <= r) s++;
        const uint32_t w = reinterpret_cast<const uint32_t *>(s_seg_ptr[b][s] + (uint64_t)(r - s_seg_start[b][s]) * R)[W - 1];
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
      const uint32_t n16 = n * U;
      for (uint32_t u0 = 0; u0 < n16; u0 += RF_THREADS) {
        const uint32_t u = u0 + tid;
        const bool on = u < n16;
        const uint32_t r = on ? u / U : 0u, part = u % U;
        const uint32_t hdr = st[r * W + W - 1];
        uint32_t pos = 0;
        if (on && part == 0) pos = atomicAdd(&s_cur[(hdr >> 8) & (F - 1u)], 1u);
        if (U > 1) pos = __shfl_sync(0xffffffffu, pos, (lane / U) * U);
        if (on) st_stream16(dst_rec + (base + pos) * R + 16 * part, reinterpret_cast<const uint4 *>(stage[b])[u]);
      }
    } else {
      for (uint32_t r = tid; r < n; r += RF_THREADS) {
        int s = 0;
        while (s_seg_start[b][s + 1] <= r) s++;
        const uint4 *g = reinterpret_cast<const uint4 *>(s_seg_ptr[b][s] + (uint64_t)(r - s_seg_start[b][s]) * R);
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = g[u];
        const uint32_t pos = atomicAdd(&s_cur[(v[U - 1].w >> 8) & (F - 1u)], 1u);
#pragma unroll
        for (int u = 0; u < U; u++) st_stream16(dst_rec + (base + pos) * R + 16 * u, v[u]);
      }
    }
    announce_bin(p, c);
    __syncthreads();  // buffer b and the histogram are free again
  }
}

// SM-partitioned pipeline: holds the stream it runs on until `need` CTAs of the refine kernel (launched on another stream)
// are resident -- the count kernel queued behind it would otherwise fill every SM first and wait for a kernel that cannot
// start. One warp, no shared memory: it takes nothing a refine CTA needs. Gives up after ~1 s (the count kernel's own
// time-out and the recount in kc_finalize then keep the result right).
__global__ void gate_kernel(const unsigned int *resident, unsigned int need) {
  if (threadIdx.x != 0) return;
  for (uint32_t spins = 0; spins < (1u << 20); spins++) {
    if (ld_acquire32(resident) >= need) return;
    __nanosleep(1000);
  }
}
cudaError_t launch_gate(const unsigned int *resident, unsigned int need, cudaStream_t s, uint64_t *launches) {
  (*launches)++;
  gate_kernel<<<1, 32, 0, s>>>(resident, need);
  return cudaGetLastError();
}

cudaError_t launch_refine_totals(const RefineParams &p, cudaStream_t s, uint64_t *launches) {
  refine_totals_kernel<<<(p.n_coarse + 255) / 256, 256, 0, s>>>(p);
  refine_scan_kernel<<<1, 1024, 0, s>>>(p);
  *launches += 2;
  return cudaGetLastError();
}

// bulk-copy variant: two stage buffers per CTA
template <int R, int THREADS, int STAGE>
static cudaError_t launch_refine_bulk_shape(const RefineParams &p, int sm_count, int ctas_per_sm, cudaStream_t s) {
  uint64_t grid = (uint64_t)sm_count * ctas_per_sm;
  const uint32_t n_bins = p.coarse_end - p.coarse_begin;
  if (grid > n_bins) grid = n_bins ? n_bins : 1;
  if (p.grid_limit > 0 && grid > (uint64_t)p.grid_limit) grid = p.grid_limit;
  cudaError_t e = cudaFuncSetAttribute(refine_bulk_kernel<R, THREADS, STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * STAGE);
  if (e != cudaSuccess) return e;
  refine_bulk_kernel<R, THREADS, STAGE><<<(int)grid, THREADS, 2 * STAGE, s>>>(p);
  return cudaGetLastError();
}

template <int R, int THREADS, int STAGE>
static cudaError_t launch_refine_shape(const RefineParams &p, int sm_count, int ctas_per_sm, cudaStream_t s) {
  uint64_t grid = (uint64_t)sm_count * ctas_per_sm;
  const uint32_t n_bins = p.coarse_end - p.coarse_begin;
  if (grid > n_bins) grid = n_bins ? n_bins : 1;
  if (p.grid_limit > 0 && grid > (uint64_t)p.grid_limit) grid = p.grid_limit;
  cudaError_t e = cudaFuncSetAttribute(refine_kernel<R, THREADS, STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, STAGE);
  if (e != cudaSuccess) return e;
  refine_kernel<R, THREADS, STAGE><<<(int)grid, THREADS, STAGE, s>>>(p);
  return cudaGetLastError();
}

// shapes of the bulk variant: stage buffer (x 2), threads, CTAs per SM
static const int RB_STAGE_TINY = 44 * 1024, RB_STAGE_SMALL = 100 * 1024;
template <int R>
static cudaError_t launch_refine_bulk_r(const RefineParams &p, int sm_count, int shape, cudaStream_t s) {
  if (shape >= 1) return launch_refine_bulk_shape<R, 512, RB_STAGE_SMALL>(p, sm_count, 1, s);
  return launch_refine_bulk_shape<R, 256, RB_STAGE_TINY>(p, sm_count, 2, s);
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
  bool bulk = true;
  if (p.grid_limit > 0) {  // SM-partitioned pipeline: one CTA owns a whole SM (208 KB stage), the count kernel has the others
    shape = 2;
    bulk = false;
  }
  if (const char *e = getenv("KC_REFINE_BULK")) bulk = atoi(e) != 0;  // (experiments: 0 = the LDG -> STS staging)
  if (bulk && want > (uint64_t)RB_STAGE_SMALL && !getenv("KC_REFINE_BULK_SHAPE")) bulk = false;  // bins that need the 208 KB single stage
  if (bulk) {
    const int bshape = want > (uint64_t)RB_STAGE_TINY ? 1 : 0;
    int bs = bshape;
    if (const char *e = getenv("KC_REFINE_BULK_SHAPE")) bs = atoi(e);
    switch (p.rec_bytes) {
      case 16: return launch_refine_bulk_r<16>(p, sm_count, bs, s);
      case 32: return launch_refine_bulk_r<32>(p, sm_count, bs, s);
      case 64: return launch_refine_bulk_r<64>(p, sm_count, bs, s);
    }
    return cudaErrorInvalidValue;
  }
  switch (p.rec_bytes) {
    case 16: return launch_refine_r<16>(p, sm_count, shape, s);
    case 32: return launch_refine_r<32>(p, sm_count, shape, s);
    case 64: return launch_refine_r<64>(p, sm_count, shape, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace kc
