// kc_finalize.cu -- kernel (5) "purge / min-depth / extension resolution / compaction", the result checksum and the
// batched lookup.
//
// Replaces HashTableInserter::insert_into_local_hashtable + ExtCounts::get_ext / get_sorted
// (src/kcount/kcount_cpu.cpp:524-601, 173-182, 135-145) and the reference's GPU kernels gpu_purge_invalid / gpu_compact_ht
// with the host-side X/F filter (src/kcount/kcount-gpu/gpu_hash_table.cpp:206-268, src/kcount/kcount_gpu.cpp:439-471).
// One streaming pass over the 32-byte slot blocks: entries with count < 2 are purged, both extensions are resolved with
// the reference's rule, entries with an X or F on either side are dropped, survivors are appended to dense SoA arrays
// through one warp-aggregated cursor update per warp.
#include "kc_internal.h"

namespace kc {

// ExtCounts::get_ext (kcount_cpu.cpp:173-182) on counters A,C,G,T; ties in the sort go to the higher letter (:137-143)
__device__ __forceinline__ char resolve_ext(const uint32_t c[4], uint32_t count, int dmin_thres, double one_minus_dyn) {
  int bi = 0;
  uint32_t top = c[0];
#pragma unroll
  for (int b = 1; b < 4; b++)
    if (c[b] >= top) {
      top = c[b];
      bi = b;
    }
  uint32_t runner = 0;
#pragma unroll
  for (int b = 0; b < 4; b++)
    if (b != bi && c[b] > runner) runner = c[b];
  int dmin_dyn = (int)__dmul_rn(one_minus_dyn, (double)count);  // (int)((1.0 - DYN_MIN_DEPTH) * count), in double
  if (dmin_dyn < dmin_thres) dmin_dyn = dmin_thres;
  if ((int)top < dmin_dyn) return 'X';
  if ((int)runner >= dmin_dyn) return 'F';
  return "ACGT"[bi];
}

struct FinalizeArgs {
  FinalizeParams p;
  double one_minus_dyn;
};

template <int N>
__global__ void __launch_bounds__(256) finalize_kernel(FinalizeArgs args) {
  const FinalizeParams &p = args.p;
  const uint32_t lane = lane_id();
  unsigned long long sum_counts = 0, max_count = 0, purged = 0;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_rounded = (p.n_slots + 31) & ~31ULL;  // whole warps stay together for the ballots
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rounded; i += stride) {
    bool occupied = false, is_final = false, is_ext = false;
    uint64_t key[N];
    uint32_t cnt = 0, ext[8];
    char left = 0, right = 0;
    if (i < p.n_slots) {
      const uint4 *sp = reinterpret_cast<const uint4 *>(p.A + i);
      uint4 a = ld_stream16(sp);
      uint64_t k0 = ((uint64_t)a.y << 32) | a.x;
      occupied = (N == 1) ? (k0 != 0) : (a.w >= 2u);
      if (occupied) {
        uint4 b = ld_stream16(sp + 1);
        key[0] = (N == 1) ? ~k0 : k0;
#pragma unroll
        for (int j = 1; j < N; j++) key[j] = p.B[i * (N - 1) + (j - 1)];
        const uint32_t count32 = a.z;
        const uint64_t lw = ((uint64_t)b.y << 32) | b.x, rw = ((uint64_t)b.w << 32) | b.z;
#pragma unroll
        for (int f = 0; f < 4; f++) {
          ext[f] = (uint32_t)(lw >> (16 * f)) & 0xffffu;
          ext[4 + f] = (uint32_t)(rw >> (16 * f)) & 0xffffu;
        }
        cnt = count32 > 65535u ? 65535u : count32;
        if (count32 >= 65535u) {
          // extension increments that arrived after saturation of the count live in the overflow table
          uint64_t idx = (i * 0x9e3779b97f4a7c15ULL >> 20) & p.ovf.mask;
          for (uint64_t probe = 0; probe <= p.ovf.mask; probe++) {
            uint64_t kk = p.ovf.keys[idx];
            if (kk == 0) break;
            if (kk == i + 1) {
#pragma unroll
              for (int f = 0; f < 8; f++) {
                uint64_t v = (uint64_t)ext[f] + p.ovf.ext[idx * 8 + f];
                ext[f] = v > 65535u ? 65535u : (uint32_t)v;
              }
              break;
            }
            idx = (idx + 1) & p.ovf.mask;
          }
        }
        if (cnt > max_count) max_count = cnt;
        if (cnt < 2u) {
          purged++;  // kcount_cpu.cpp:557
        } else {
          left = resolve_ext(ext, cnt, p.dmin_thres, args.one_minus_dyn);
          right = resolve_ext(ext + 4, cnt, p.dmin_thres, args.one_minus_dyn);
          is_ext = p.keep_ext != 0;
          is_final = !(left == 'X' || left == 'F' || right == 'X' || right == 'F');  // kcount_cpu.cpp:569-573
          if (is_final)
            sum_counts += cnt;
          else
            purged++;
        }
      }
    }
    // ---- survivors
    uint32_t fm = __ballot_sync(0xffffffffu, is_final);
    if (fm) {
      unsigned long long base = 0;
      if (lane == (uint32_t)(__ffs(fm) - 1)) base = atomicAdd(&p.ctr->n_final, (unsigned long long)__popc(fm));
      base = __shfl_sync(0xffffffffu, base, __ffs(fm) - 1);
      if (is_final) {
        uint64_t o = base + __popc(fm & ((1u << lane) - 1u));
        if (o < p.cap_final) {
#pragma unroll
          for (int j = 0; j < N; j++) p.out_keys[o * N + j] = key[j];
          p.out_counts[o] = (uint16_t)cnt;
          p.out_left[o] = left;
          p.out_right[o] = right;
        }
      }
    }
    // ---- every entry with count >= 2, with its counters (dump D2)
    uint32_t em = __ballot_sync(0xffffffffu, is_ext);
    if (em) {
      unsigned long long base = 0;
      if (lane == (uint32_t)(__ffs(em) - 1)) base = atomicAdd(&p.ctr->n_ext, (unsigned long long)__popc(em));
      base = __shfl_sync(0xffffffffu, base, __ffs(em) - 1);
      if (is_ext) {
        uint64_t o = base + __popc(em & ((1u << lane) - 1u));
        if (o < p.cap_ext) {
#pragma unroll
          for (int j = 0; j < N; j++) p.ext_keys[o * N + j] = key[j];
          p.ext_counts[o] = (uint16_t)cnt;
          p.ext_left[o] = left;
          p.ext_right[o] = right;
#pragma unroll
          for (int f = 0; f < 8; f++) p.ext_ext8[o * 8 + f] = (uint16_t)ext[f];
        }
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum_counts += __shfl_down_sync(0xffffffffu, sum_counts, o);
    purged += __shfl_down_sync(0xffffffffu, purged, o);
    unsigned long long m = __shfl_down_sync(0xffffffffu, max_count, o);
    if (m > max_count) max_count = m;
  }
  if (lane == 0) {
    if (sum_counts) atomicAdd(&p.ctr->sum_counts, sum_counts);
    if (purged) atomicAdd(&p.ctr->purged, purged);
    if (max_count) atomicMax(&p.ctr->max_count, max_count);
  }
}

cudaError_t launch_finalize(const FinalizeParams &p, int sm_count, cudaStream_t s, uint64_t *launches) {
  FinalizeArgs a;
  a.p = p;
  a.one_minus_dyn = 1.0 - p.dyn_min_depth;  // evaluated on the host exactly as the reference's constant expression
  uint64_t blocks = (p.n_slots + 255) / 256;
  uint64_t cap = (uint64_t)sm_count * 8 * 4;
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
