// kc_noalloc.cu -- kc_checksum_noalloc: the result checksum for callers that must not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Same value as kc_result_checksum (sum over entries of entry_mix, kc_device.cuh), computed from arrays the CALLER
// passes (device pointers, e.g. those of kc_result_device_ptrs), with the scratch in a file-scope device array:
//   * g_noalloc_scratch: KC_NOALLOC_SCRATCH_WORDS x 8 bytes = 32 KB, resident on every GPU that loads the library until exit;
//   * stage 1: block b writes its partial sum to word b (no atomics: the value does not depend on scheduling);
//     stage 2: one warp adds the words of the blocks that ran and writes the total to the last word;
//   * every word a call reads was written by that call (stale contents of earlier calls are never summed), and ALL calls
//     share the array: two calls that overlap on the GPU (different streams / threads / processes on one context) race.
//     Calls here are issued on the legacy default stream and complete before the entry point returns.
// The entry point calls no cudaMalloc / cudaFree / event or stream creation: cudaGetSymbolAddress (once per device),
// two kernel launches, one 8-byte cudaMemcpy.
#include <cstdio>
#include <mutex>

#include "../../include/kcount_b200.h"
#include "kc_internal.h"

namespace kc {

static const int kNoallocWords = KC_NOALLOC_SCRATCH_WORDS;
static const int kNoallocMaxBlocks = kNoallocWords - 1;  // the last word holds the total
static const int kNoallocMaxDevices = 64;

static __device__ __align__(256) unsigned long long g_noalloc_scratch[KC_NOALLOC_SCRATCH_WORDS];

__global__ void __launch_bounds__(256) noalloc_partials_kernel(const uint64_t *keys, const uint16_t *counts, const char *left, const char *right,
                                                               const uint16_t *ext8, uint64_t n, int n_words, unsigned long long *partials) {
  __shared__ unsigned long long s_warp[8];
  unsigned long long sum = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    uint16_t e[8];
    if (ext8)
      for (int f = 0; f < 8; f++) e[f] = ext8[i * 8 + f];
    sum += entry_mix(keys + i * n_words, n_words, counts[i], left[i], right[i], ext8 ? e : nullptr);
  }
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
  if (lane_id() == 0) s_warp[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int w = 0; w < 8; w++) t += s_warp[w];
    partials[blockIdx.x] = t;  // (overwrites whatever an earlier call left in this word)
  }
}

__global__ void noalloc_total_kernel(unsigned long long *scratch, int n_blocks, int total_word) {
  unsigned long long sum = 0;
  for (int i = (int)threadIdx.x; i < n_blocks; i += 32) sum += scratch[i];
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
  if (threadIdx.x == 0) scratch[total_word] = sum;
}

// device address of the scratch array, looked up once per device with that device current (in host code the array's name
// is only a host-side stand-in for the symbol, not an address a kernel can use)
static unsigned long long *scratch_of(int device, cudaError_t *err) {
  static std::mutex mu;
  static unsigned long long *ptr[kNoallocMaxDevices] = {};
  std::lock_guard<std::mutex> lock(mu);
  *err = cudaSuccess;
  if (!ptr[device]) {
    void *p = nullptr;
    *err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (*err != cudaSuccess) return nullptr;
    ptr[device] = static_cast<unsigned long long *>(p);
  }
  return ptr[device];
}

}  // namespace kc

extern "C" int kc_checksum_noalloc(int device, const uint64_t *keys, const uint16_t *counts, const char *left, const char *right,
                                   const uint16_t *ext8, uint64_t n_entries, uint32_t n_words, uint64_t *checksum) {
  using namespace kc;
  if (!checksum || n_words < 1 || n_words > KC_MAX_KEY_WORDS) return KC_ERR_ARG;
  if (n_entries > KC_NOALLOC_MAX_ENTRIES) {
    fprintf(stderr, "[kc] kc_checksum_noalloc: %llu entries exceed the maximum of %llu (KC_NOALLOC_MAX_ENTRIES)\n", (unsigned long long)n_entries,
            (unsigned long long)KC_NOALLOC_MAX_ENTRIES);
    return KC_ERR_ARG;
  }
  if (n_entries && (!keys || !counts || !left || !right)) return KC_ERR_ARG;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1) return KC_ERR_CUDA;  // no CPU path
  if (device < 0 || device >= n_dev || device >= kNoallocMaxDevices) return KC_ERR_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return KC_ERR_CUDA;
  cudaError_t e;
  unsigned long long *scratch = scratch_of(device, &e);
  if (!scratch) {
    fprintf(stderr, "[kc] kc_checksum_noalloc: cudaGetSymbolAddress: %s\n", cudaGetErrorString(e));
    return KC_ERR_CUDA;
  }
  uint64_t blocks = (n_entries + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks > (uint64_t)kNoallocMaxBlocks) blocks = kNoallocMaxBlocks;
  if (blocks < 1) blocks = 1;
  noalloc_partials_kernel<<<(int)blocks, 256>>>(keys, counts, left, right, ext8, n_entries, (int)n_words, scratch);
  noalloc_total_kernel<<<1, 32>>>(scratch, (int)blocks, kNoallocWords - 1);
  if ((e = cudaGetLastError()) != cudaSuccess) {
    fprintf(stderr, "[kc] kc_checksum_noalloc: launch: %s\n", cudaGetErrorString(e));
    return KC_ERR_CUDA;
  }
  // host or device destination (unified addressing); synchronous on the default stream, so the call is complete on return
  if ((e = cudaMemcpy(checksum, scratch + kNoallocWords - 1, sizeof(uint64_t), cudaMemcpyDefault)) != cudaSuccess) {
    fprintf(stderr, "[kc] kc_checksum_noalloc: %s\n", cudaGetErrorString(e));
    return KC_ERR_CUDA;
  }
  return KC_OK;
}
