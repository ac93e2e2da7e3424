// kc_lookup.cu -- device-resident index over the finalized result and the batched query kernel behind kc_lookup.
//
// Replaces the host map the reference fills after its GPU work (src/kcount/kcount_gpu.cpp:439-471: every surviving entry
// is inserted into KmerMap on the host, serially) and the point lookups the next stage makes against it
// (KmerDHT::get_local_kmer_counts / kmer_exists / get_kmer_count, src/kcount/kmer_dht.cpp:199-245): the result stays in
// HBM as dense arrays (keys / counts / left / right), an open-addressing index of {fingerprint, entry number} words is
// built over it once, and lookups arrive in batches.
//   index slot  = u64: fingerprint (high 32 bits of the key hash) << 32 | entry number; all-ones = empty
//   capacity    = power of two >= 2 x entries (load <= 0.5), linear probing
// A query reads on average ~1.5 slots (8 B each, one 32-byte sector) and, on a fingerprint match, the N key words and
// the 4 payload bytes of the entry: three or four random sectors per query, so the kernel is bound by random-access
// HBM sector throughput, not by instructions.
#include "kc_internal.h"

namespace kc {

static const unsigned long long IX_EMPTY = ~0ULL;

template <int N>
__global__ void __launch_bounds__(256) index_build_kernel(IndexParams p) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += stride) {
    uint64_t key[N];
#pragma unroll
    for (int w = 0; w < N; w++) key[w] = p.keys[i * N + w];
    const uint64_t h = key_hash<N>(key);
    const unsigned long long entry = (h & 0xffffffff00000000ULL) | i;
    uint64_t pos = h & p.mask;
    uint32_t probes = 0;
    while (atomicCAS(p.slots + pos, IX_EMPTY, entry) != IX_EMPTY) {  // keys of a result are distinct: no compare needed
      pos = (pos + 1) & p.mask;
      if (++probes > 4096u) {  // cannot happen at load <= 0.5 with a mixing hash; never spin
        *p.err = 1u;
        break;
      }
    }
  }
}

template <int N>
__global__ void __launch_bounds__(256) index_lookup_kernel(IndexParams p, const uint64_t *__restrict__ queries, uint64_t n_queries,
                                                           uint16_t *__restrict__ out_counts, char *__restrict__ out_left,
                                                           char *__restrict__ out_right) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_queries; q += stride) {
    uint64_t key[N];
#pragma unroll
    for (int w = 0; w < N; w++) key[w] = queries[q * N + w];
    const uint64_t h = key_hash<N>(key);
    const uint64_t fp = h >> 32;
    uint64_t pos = h & p.mask;
    uint16_t c = 0;
    char l = 0, r = 0;
    for (uint32_t probes = 0; probes <= 4096u; probes++) {
      const unsigned long long e = __ldg(p.slots + pos);
      if (e == IX_EMPTY) break;  // absent (purged, dropped as X/F, or never seen): kmer_dht.cpp:199-203 returns nullptr
      if ((e >> 32) == fp) {
        const uint64_t i = e & 0xffffffffULL;
        bool same = true;
#pragma unroll
        for (int w = 0; w < N; w++) same = same && __ldg(p.keys + i * N + w) == key[w];
        if (same) {
          c = p.counts[i];
          l = p.left[i];
          r = p.right[i];
          break;
        }
      }
      pos = (pos + 1) & p.mask;
    }
    if (out_counts) out_counts[q] = c;
    if (out_left) out_left[q] = l;
    if (out_right) out_right[q] = r;
  }
}

static int grid_for(uint64_t n, int sm_count) {
  const uint64_t want = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count * 8;  // 8 CTAs of 256 threads per SM: full occupancy, grid-stride beyond
  return (int)(want < cap ? (want ? want : 1) : cap);
}

cudaError_t launch_index_build(const IndexParams &p, int sm_count, cudaStream_t s, uint64_t *launches) {
  cudaError_t e = cudaMemsetAsync(p.slots, 0xff, (p.mask + 1) * sizeof(unsigned long long), s);
  if (e != cudaSuccess) return e;
  if (p.n == 0) return cudaSuccess;
  (*launches)++;
  const int grid = grid_for(p.n, sm_count);
  switch (p.n_words) {
    case 1: index_build_kernel<1><<<grid, 256, 0, s>>>(p); break;
    case 2: index_build_kernel<2><<<grid, 256, 0, s>>>(p); break;
    case 3: index_build_kernel<3><<<grid, 256, 0, s>>>(p); break;
    case 4: index_build_kernel<4><<<grid, 256, 0, s>>>(p); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_index_lookup(const IndexParams &p, const uint64_t *queries, uint64_t n_queries, uint16_t *out_counts, char *out_left,
                                char *out_right, int sm_count, cudaStream_t s, uint64_t *launches) {
  if (n_queries == 0) return cudaSuccess;
  (*launches)++;
  const int grid = grid_for(n_queries, sm_count);
  switch (p.n_words) {
    case 1: index_lookup_kernel<1><<<grid, 256, 0, s>>>(p, queries, n_queries, out_counts, out_left, out_right); break;
    case 2: index_lookup_kernel<2><<<grid, 256, 0, s>>>(p, queries, n_queries, out_counts, out_left, out_right); break;
    case 3: index_lookup_kernel<3><<<grid, 256, 0, s>>>(p, queries, n_queries, out_counts, out_left, out_right); break;
    case 4: index_lookup_kernel<4><<<grid, 256, 0, s>>>(p, queries, n_queries, out_counts, out_left, out_right); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace kc
