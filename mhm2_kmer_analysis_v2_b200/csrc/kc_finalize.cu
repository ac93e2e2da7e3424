// kc_finalize.cu -- kernel (5) "purge / min-depth / extension resolution / compaction", the result checksum and the
// batched lookup.
//
// Replaces HashTableInserter::insert_into_local_hashtable + ExtCounts::get_ext / get_sorted
// (src/kcount/kcount_cpu.cpp:524-601, 173-182, 135-145) and the reference's GPU kernels gpu_purge_invalid / gpu_compact_ht
// with the host-side X/F filter (src/kcount/kcount-gpu/gpu_hash_table.cpp:206-268, src/kcount/kcount_gpu.cpp:439-471).
// One streaming pass: the 32-bit occurrence counts are read densely (N >= 2: 16 bytes = 4 slots per thread from the
// side array C; N == 1: the first half of the slot), and only slots with count >= 2 have their 32-byte block read.
// Entries with count < 2 are purged, both extensions are resolved with the reference's rule, entries with an X or F on
// either side are dropped, survivors are appended to dense SoA arrays through one warp-aggregated cursor update per
// warp. Keys with more than 65535 occurrences are only listed here ("heavy"); they are finished by
// finalize_heavy_kernel after the exact recount of their extensions.
#include "kc_internal.h"

namespace kc {

struct FinalizeArgs {
  FinalizeParams p;
  double one_minus_dyn;
};

struct Entry {
  uint32_t cnt;
  char left, right;
  bool is_final, is_ext;
};

// key words of slot i (stored complemented)
template <int N>
__device__ __forceinline__ void load_key(const FinalizeParams &p, uint64_t i, uint64_t k0x, uint64_t w1, uint64_t (&key)[N]) {
  key[0] = ~k0x;
  if (N >= 2) key[N >= 2 ? 1 : 0] = ~w1;
  if (N >= 3) {
    uint4 b = ld_stream16(p.B + i);
    key[N >= 3 ? 2 : 0] = ~(((uint64_t)b.y << 32) | b.x);
    if (N >= 4) key[N >= 4 ? 3 : 0] = ~(((uint64_t)b.w << 32) | b.z);
  }
}

template <int N>
struct Pending {  // one resolved table entry held in registers until the block knows where it goes
  uint64_t key[N];
  uint64_t lw, rw;   // packed extension counters
  uint32_t cnt;      // 0 = nothing to write
  char left, right;
  bool is_final;
};

template <int N>
__device__ __forceinline__ void write_entry(const Pending<N> &e, uint64_t o, uint64_t cap, uint64_t *out_keys, uint16_t *out_counts,
                                            char *out_left, char *out_right, uint16_t *out_ext8) {
  if (o >= cap) return;
#pragma unroll
  for (int j = 0; j < N; j++) out_keys[o * N + j] = e.key[j];
  out_counts[o] = (uint16_t)e.cnt;
  out_left[o] = e.left;
  out_right[o] = e.right;
  if (out_ext8) {
    uint4 v;  // lA lC lG lT rA rC rG rT as 8 x u16 = the two packed words
    v.x = (uint32_t)e.lw;
    v.y = (uint32_t)(e.lw >> 32);
    v.z = (uint32_t)e.rw;
    v.w = (uint32_t)(e.rw >> 32);
    *reinterpret_cast<uint4 *>(out_ext8 + o * 8) = v;
  }
}

template <int N>
__global__ void __launch_bounds__(256) finalize_kernel(FinalizeArgs args) {
  const FinalizeParams &p = args.p;
  const uint32_t lane = lane_id();
  const int warp = threadIdx.x >> 5;
  constexpr int PER = (N == 1) ? 2 : 4;  // slots per thread and iteration
  __shared__ uint32_t s_warp[8];
  __shared__ unsigned long long s_base[2];
  unsigned long long sum_counts = 0, max_count = 0, purged = 0, ge2 = 0;
  const uint64_t n_groups = (p.n_slots + PER - 1) / PER;
  const uint64_t n_rounded = (n_groups + 255) & ~255ULL;  // whole CTAs stay together for the barriers
  for (uint64_t g0 = (uint64_t)blockIdx.x * 256; g0 < n_rounded; g0 += (uint64_t)gridDim.x * 256) {
    const uint64_t g = g0 + threadIdx.x;
    uint32_t c[PER];
    uint64_t k0x[PER];
#pragma unroll
    for (int u = 0; u < PER; u++) {
      c[u] = 0;
      k0x[u] = 0;
    }
    if (g < n_groups) {
      if (N == 1) {
#pragma unroll
        for (int u = 0; u < PER; u++)
          if (g * PER + u < p.n_slots) {
            uint4 a = ld_stream16(reinterpret_cast<const uint4 *>(p.A + g * PER + u));
            k0x[u] = ((uint64_t)a.y << 32) | a.x;
            c[u] = k0x[u] ? a.z : 0u;
          }
      } else if (g * PER + PER <= p.n_slots) {
        uint4 v = *reinterpret_cast<const uint4 *>(p.C + g * PER);
        c[0] = v.x;
        c[1] = v.y;
        c[PER > 2 ? 2 : 0] = v.z;
        c[PER > 3 ? 3 : 0] = v.w;
      } else {
        for (int u = 0; u < PER; u++)
          if (g * PER + u < p.n_slots) c[u] = p.C[g * PER + u];
      }
    }
    // ---- resolve this thread's entries
    Pending<N> e[PER];
    uint32_t n_fin = 0, n_ext = 0;
#pragma unroll
    for (int u = 0; u < PER; u++) {
      const uint64_t i = g * PER + u;
      const uint32_t cnt32 = c[u];
      e[u].cnt = 0;
      e[u].is_final = false;
      if (cnt32 > 65535u) {
        // heavy key: list it, mark the slot with its index; finished after the exact recount
        unsigned long long h = atomicAdd(p.heavy.n, 1ULL);
        if (h < p.heavy.cap) {
          p.heavy.slots[h] = i;
          *(N == 1 ? reinterpret_cast<uint32_t *>(&p.A[i].w1) : p.C + i) = kHeavyMark | (uint32_t)h;
        } else {
          p.ctr->err_heavy_full = 1u;
        }
      } else if (cnt32 == 1u) {
        purged++;  // kcount_cpu.cpp:557
      } else if (cnt32 >= 2u) {
        const uint4 *sp = reinterpret_cast<const uint4 *>(p.A + i);
        uint64_t w1 = 0, k0 = k0x[u];
        if (N >= 2) {
          uint4 a = ld_stream16(sp);
          k0 = ((uint64_t)a.y << 32) | a.x;
          w1 = ((uint64_t)a.w << 32) | a.z;
        }
        uint4 b = ld_stream16(sp + 1);
        load_key<N>(p, i, k0, w1, e[u].key);
        e[u].lw = ((uint64_t)b.y << 32) | b.x;
        e[u].rw = ((uint64_t)b.w << 32) | b.z;
        uint32_t ext[8];
#pragma unroll
        for (int f = 0; f < 4; f++) {
          ext[f] = (uint32_t)(e[u].lw >> (16 * f)) & 0xffffu;
          ext[4 + f] = (uint32_t)(e[u].rw >> (16 * f)) & 0xffffu;
        }
        e[u].cnt = cnt32;
        ge2++;
        if (cnt32 > max_count) max_count = cnt32;
        e[u].left = resolve_ext(ext, cnt32, p.dmin_thres, args.one_minus_dyn);
        e[u].right = resolve_ext(ext + 4, cnt32, p.dmin_thres, args.one_minus_dyn);
        e[u].is_final = !(e[u].left == 'X' || e[u].left == 'F' || e[u].right == 'X' || e[u].right == 'F');  // kcount_cpu.cpp:569-573
        if (e[u].is_final) {
          sum_counts += cnt32;
          n_fin++;
        } else {
          purged++;
        }
        n_ext++;
      }
    }
    if (!p.keep_ext) n_ext = 0;
    // ---- block-wide exclusive scan of (survivors, extended entries): one cursor update per CTA and iteration
    uint32_t packed = n_fin | (n_ext << 16), incl = packed;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t tot = 0;
      for (int w = 0; w < 8; w++) {
        uint32_t t = s_warp[w];
        s_warp[w] = tot;
        tot += t;
      }
      s_base[0] = (tot & 0xffffu) ? atomicAdd(&p.ctr->n_final, (unsigned long long)(tot & 0xffffu)) : 0ULL;
      s_base[1] = (tot >> 16) ? atomicAdd(&p.ctr->n_ext, (unsigned long long)(tot >> 16)) : 0ULL;
    }
    __syncthreads();
    const uint32_t excl = s_warp[warp] + incl - packed;
    uint64_t o_fin = s_base[0] + (excl & 0xffffu), o_ext = s_base[1] + (excl >> 16);
    __syncthreads();  // s_warp / s_base are rewritten by the next iteration
#pragma unroll
    for (int u = 0; u < PER; u++) {
      if (e[u].cnt == 0) continue;
      if (e[u].is_final) write_entry<N>(e[u], o_fin++, p.cap_final, p.out_keys, p.out_counts, p.out_left, p.out_right, nullptr);
      if (p.keep_ext) write_entry<N>(e[u], o_ext++, p.cap_ext, p.ext_keys, p.ext_counts, p.ext_left, p.ext_right, p.ext_ext8);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum_counts += __shfl_down_sync(0xffffffffu, sum_counts, o);
    purged += __shfl_down_sync(0xffffffffu, purged, o);
    ge2 += __shfl_down_sync(0xffffffffu, ge2, o);
    unsigned long long m = __shfl_down_sync(0xffffffffu, max_count, o);
    if (m > max_count) max_count = m;
  }
  if (lane == 0) {
    if (sum_counts) atomicAdd(&p.ctr->sum_counts, sum_counts);
    if (purged) atomicAdd(&p.ctr->purged, purged);
    if (ge2) atomicAdd(&p.ctr->ge2, ge2);
    if (max_count) atomicMax(&p.ctr->max_count, max_count);
  }
}

// heavy keys (more than 65535 occurrences): count saturates at 65535 (kcount_cpu.cpp:349-351), the extension counters
// come from the exact 32-bit recount and saturate per counter (kcount_cpu.cpp:152-155)
template <int N>
__global__ void __launch_bounds__(128) finalize_heavy_kernel(FinalizeArgs args, uint64_t n_heavy) {
  const FinalizeParams &p = args.p;
  const uint64_t h = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= n_heavy) return;
  const uint64_t i = p.heavy.slots[h];
  const Slot s = p.A[i];
  uint64_t key[N];
  load_key<N>(p, i, s.k0x, s.w1, key);
  uint32_t ext[8];
#pragma unroll
  for (int f = 0; f < 8; f++) {
    uint32_t v = p.heavy.side[h * 8 + f];
    ext[f] = v > 65535u ? 65535u : v;
  }
  const uint32_t cnt = 65535u;
  const char left = resolve_ext(ext, cnt, p.dmin_thres, args.one_minus_dyn);
  const char right = resolve_ext(ext + 4, cnt, p.dmin_thres, args.one_minus_dyn);
  const bool is_final = !(left == 'X' || left == 'F' || right == 'X' || right == 'F');
  atomicAdd(&p.ctr->ge2, 1ULL);
  atomicMax(&p.ctr->max_count, (unsigned long long)cnt);
  if (is_final) {
    atomicAdd(&p.ctr->sum_counts, (unsigned long long)cnt);
    uint64_t o = atomicAdd(&p.ctr->n_final, 1ULL);
    if (o < p.cap_final) {
      for (int j = 0; j < N; j++) p.out_keys[o * N + j] = key[j];
      p.out_counts[o] = (uint16_t)cnt;
      p.out_left[o] = left;
      p.out_right[o] = right;
    }
  } else {
    atomicAdd(&p.ctr->purged, 1ULL);
  }
  if (p.keep_ext) {
    uint64_t o = atomicAdd(&p.ctr->n_ext, 1ULL);
    if (o < p.cap_ext) {
      for (int j = 0; j < N; j++) p.ext_keys[o * N + j] = key[j];
      p.ext_counts[o] = (uint16_t)cnt;
      p.ext_left[o] = left;
      p.ext_right[o] = right;
      for (int f = 0; f < 8; f++) p.ext_ext8[o * 8 + f] = (uint16_t)ext[f];
    }
  }
}

static FinalizeArgs make_args(const FinalizeParams &p) {
  FinalizeArgs a;
  a.p = p;
  a.one_minus_dyn = 1.0 - p.dyn_min_depth;  // evaluated on the host exactly as the reference's constant expression
  return a;
}

cudaError_t launch_finalize(const FinalizeParams &p, int sm_count, cudaStream_t s, uint64_t *launches) {
  FinalizeArgs a = make_args(p);
  const int per = p.n_words == 1 ? 2 : 4;
  uint64_t blocks = ((p.n_slots + per - 1) / per + 255) / 256;
  uint64_t cap = (uint64_t)sm_count * 8 * 2;
  int grid = (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
  (*launches)++;
  switch (p.n_words) {
    case 1: finalize_kernel<1><<<grid, 256, 0, s>>>(a); break;
    case 2: finalize_kernel<2><<<grid, 256, 0, s>>>(a); break;
    case 3: finalize_kernel<3><<<grid, 256, 0, s>>>(a); break;
    case 4: finalize_kernel<4><<<grid, 256, 0, s>>>(a); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_finalize_heavy(const FinalizeParams &p, uint64_t n_heavy, cudaStream_t s, uint64_t *launches) {
  if (n_heavy == 0) return cudaSuccess;
  FinalizeArgs a = make_args(p);
  int grid = (int)((n_heavy + 127) / 128);
  (*launches)++;
  switch (p.n_words) {
    case 1: finalize_heavy_kernel<1><<<grid, 128, 0, s>>>(a, n_heavy); break;
    case 2: finalize_heavy_kernel<2><<<grid, 128, 0, s>>>(a, n_heavy); break;
    case 3: finalize_heavy_kernel<3><<<grid, 128, 0, s>>>(a, n_heavy); break;
    case 4: finalize_heavy_kernel<4><<<grid, 128, 0, s>>>(a, n_heavy); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// order-independent checksum of a result: sum over entries of entry_mix (same function as the oracle's)
__global__ void checksum_kernel(const uint64_t *keys, const uint16_t *counts, const char *left, const char *right,
                                const uint16_t *ext8, uint64_t n, int n_words, unsigned long long *out) {
  unsigned long long sum = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    uint16_t e[8];
    if (ext8)
      for (int f = 0; f < 8; f++) e[f] = ext8[i * 8 + f];
    sum += entry_mix(keys + i * n_words, n_words, counts[i], left[i], right[i], ext8 ? e : nullptr);
  }
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
  if (lane_id() == 0) atomicAdd(out, sum);
}

cudaError_t launch_checksum(const uint64_t *keys, const uint16_t *counts, const char *left, const char *right, const uint16_t *ext8,
                            uint64_t n, int n_words, unsigned long long *out, cudaStream_t s, uint64_t *launches) {
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  (*launches)++;
  checksum_kernel<<<(int)blocks, 256, 0, s>>>(keys, counts, left, right, ext8, n, n_words, out);
  return cudaGetLastError();
}

}  // namespace kc
