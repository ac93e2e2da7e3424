// kc_sort.cu -- orders the finalized result by key on the device (kc_result_sort).
//
// Not on the hot path: the stage produces its result in bin order, and only the text dump (KmerDHT::dump_kmers,
// src/kcount/kmer_dht.cpp:274-297, written sorted here so that two runs can be compared) and the parity tests want an
// order. Keys compare word by word, unsigned (Kmer::operator<, src/kmer.cpp:271-277): an LSD radix sort over the N key
// words, last word first, each pass a stable 64-bit cub::DeviceRadixSort::SortPairs of (word, entry number); the entry
// arrays are then gathered once through the final permutation. cub is the CUDA toolkit's own library (this is a plain
// library sort, not one of the stage's kernels).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "kc_internal.h"

namespace kc {

__global__ void sort_iota_kernel(uint32_t *perm, uint64_t n) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) perm[i] = (uint32_t)i;
}

__global__ void sort_gather_word_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ perm, int w, int n_words, uint64_t n,
                                        uint64_t *__restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = keys[(uint64_t)perm[i] * n_words + w];
}

template <typename T, int PER>
__global__ void sort_gather_kernel(const T *__restrict__ in, const uint32_t *__restrict__ perm, uint64_t n, T *__restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t o = perm[i];
#pragma unroll
    for (int j = 0; j < PER; j++) out[i * PER + j] = in[o * PER + j];
  }
}

template <typename T, int PER>
static cudaError_t gather_in_place(T *arr, const uint32_t *perm, uint64_t n, void *tmp, int grid, cudaStream_t s, uint64_t *launches) {
  (*launches)++;
  sort_gather_kernel<T, PER><<<grid, 256, 0, s>>>(arr, perm, n, reinterpret_cast<T *>(tmp));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  return cudaMemcpyAsync(arr, tmp, n * PER * sizeof(T), cudaMemcpyDeviceToDevice, s);
}

cudaError_t sort_result_by_key(uint64_t *keys, uint16_t *counts, char *left, char *right, uint16_t *ext8, uint64_t n, int n_words, int sm_count,
                               cudaStream_t s, uint64_t *launches) {
  if (n < 2) return cudaSuccess;
  if (n >= 0xffffffffULL) return cudaErrorInvalidValue;
  const int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count * 8);
  uint32_t *perm[2] = {nullptr, nullptr};
  uint64_t *word[2] = {nullptr, nullptr};
  void *temp = nullptr, *tmp = nullptr;
  size_t temp_bytes = 0;
  cudaError_t e = cudaSuccess;
  auto done = [&](cudaError_t r) {
    cudaStreamSynchronize(s);
    cudaFree(perm[0]), cudaFree(perm[1]), cudaFree(word[0]), cudaFree(word[1]), cudaFree(temp), cudaFree(tmp);
    return r;
  };
  if ((e = cudaMalloc(&perm[0], n * 4)) != cudaSuccess || (e = cudaMalloc(&perm[1], n * 4)) != cudaSuccess ||
      (e = cudaMalloc(&word[0], n * 8)) != cudaSuccess || (e = cudaMalloc(&word[1], n * 8)) != cudaSuccess)
    return done(e);
  if ((e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, word[0], word[1], perm[0], perm[1], n, 0, 64, s)) != cudaSuccess) return done(e);
  if ((e = cudaMalloc(&temp, temp_bytes ? temp_bytes : 1)) != cudaSuccess) return done(e);
  (*launches)++;
  sort_iota_kernel<<<grid, 256, 0, s>>>(perm[0], n);
  for (int w = n_words - 1; w >= 0; w--) {
    (*launches)++;
    sort_gather_word_kernel<<<grid, 256, 0, s>>>(keys, perm[0], w, n_words, n, word[0]);
    if ((e = cudaGetLastError()) != cudaSuccess) return done(e);
    if ((e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, word[0], word[1], perm[0], perm[1], n, 0, 64, s)) != cudaSuccess) return done(e);
    std::swap(perm[0], perm[1]);
  }
  // one scratch array, large enough for the widest member (keys, or the 8 extension counters)
  const size_t widest = std::max<size_t>((size_t)n_words * 8, ext8 ? 16 : 0);
  if ((e = cudaMalloc(&tmp, n * widest)) != cudaSuccess) return done(e);
  switch (n_words) {
    case 1: e = gather_in_place<uint64_t, 1>(keys, perm[0], n, tmp, grid, s, launches); break;
    case 2: e = gather_in_place<uint64_t, 2>(keys, perm[0], n, tmp, grid, s, launches); break;
    case 3: e = gather_in_place<uint64_t, 3>(keys, perm[0], n, tmp, grid, s, launches); break;
    case 4: e = gather_in_place<uint64_t, 4>(keys, perm[0], n, tmp, grid, s, launches); break;
    default: e = cudaErrorInvalidValue;
  }
  if (e == cudaSuccess) e = gather_in_place<uint16_t, 1>(counts, perm[0], n, tmp, grid, s, launches);
  if (e == cudaSuccess) e = gather_in_place<char, 1>(left, perm[0], n, tmp, grid, s, launches);
  if (e == cudaSuccess) e = gather_in_place<char, 1>(right, perm[0], n, tmp, grid, s, launches);
  if (e == cudaSuccess && ext8) e = gather_in_place<uint16_t, 8>(ext8, perm[0], n, tmp, grid, s, launches);
  return done(e);
}

// offsets of the reads inside the packed byte stream from their lengths (kc_add_packed_reads): lengths widened to 64 bits
// in place of the offsets, then the toolkit's exclusive scan over them (input-side helper, not a stage kernel).
__global__ void widen_lens_kernel(const uint16_t *__restrict__ lens, uint64_t *__restrict__ out, uint64_t n) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = lens[i];
}

cudaError_t scan_lens_to_offsets(const uint16_t *lens, uint64_t *offsets, uint64_t n, cudaStream_t s, uint64_t *launches) {
  if (n == 0) return cudaSuccess;
  if (n > 0x7fffffffULL) return cudaErrorInvalidValue;
  widen_lens_kernel<<<(int)std::min<uint64_t>((n + 255) / 256, 148 * 8), 256, 0, s>>>(lens, offsets, n);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  size_t temp_bytes = 0;
  if ((e = cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, offsets, offsets, (int)n, s)) != cudaSuccess) return e;
  void *temp = nullptr;
  if ((e = cudaMallocAsync(&temp, temp_bytes ? temp_bytes : 1, s)) != cudaSuccess) return e;
  e = cub::DeviceScan::ExclusiveSum(temp, temp_bytes, offsets, offsets, (int)n, s);
  (*launches) += 3;
  cudaFreeAsync(temp, s);
  return e;
}

}  // namespace kc
