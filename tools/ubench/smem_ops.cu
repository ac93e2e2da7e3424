// Microbenchmark (B200): per-SM throughput of the shared-memory operations a bin-resident k-mer table can be built from.
// Reports SM cycles per warp-wide operation with 24 resident warps per SM (2 CTAs x 384 threads), random word addresses.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o smem_ops smem_ops.cu ; run: ./smem_ops
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

static const int THREADS = 384, ITERS = 4096, WORDS = 24 * 1024;  // 96 KB of shared memory per CTA

__device__ __forceinline__ uint4 lds128(const uint32_t *p) {
  uint4 r;
  asm volatile("ld.volatile.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
  return r;
}

__device__ __forceinline__ uint32_t rnd(uint32_t &s) {
  s = s * 1664525u + 1013904223u;
  return s >> 8;
}

template <int MODE>
__global__ void __launch_bounds__(THREADS, 2) bench(uint32_t *out, long long *cycles) {
  extern __shared__ __align__(16) uint32_t sm[];
  for (int i = threadIdx.x; i < WORDS; i += THREADS) sm[i] = 0;
  __syncthreads();
  uint32_t s = blockIdx.x * 7919u + threadIdx.x * 104729u + 1u, acc = 0;
  const long long t0 = clock64();
#pragma unroll 4
  for (int it = 0; it < ITERS; it++) {
    const uint32_t r = rnd(s);
    if (MODE == 0) {  // one ATOMS.ADD, random word
      atomicAdd(sm + r % WORDS, 1u);
    } else if (MODE == 1) {  // three ATOMS.ADD into one 32-byte slot (count, left, right)
      uint32_t *p = sm + (r % (WORDS / 8)) * 8;
      atomicAdd(p, 1u);
      atomicAdd(p + 1 + (r >> 20 & 1), 1u);
      atomicAdd(p + 3 + (r >> 21 & 1), 1u);
    } else if (MODE == 2) {  // LDS.32 + STS.32 read-modify-write
      volatile uint32_t *p = sm + r % WORDS;
      *p = *p + 1u;
    } else if (MODE == 3) {  // LDS.64 + STS.64
      volatile unsigned long long *p = reinterpret_cast<unsigned long long *>(sm) + r % (WORDS / 2);
      *p = *p + 1ull;
    } else if (MODE == 4) {  // LDS.128 only (bucket of four tags)
      const uint4 v = lds128(sm + (r % (WORDS / 4)) * 4);
      acc += v.x ^ v.y ^ v.z ^ v.w;
    } else if (MODE == 5) {  // match.any on a slot number + LDS.64/STS.64 by the group leader
      const uint32_t slot = r % (WORDS / 2);
      const uint32_t m = __match_any_sync(0xffffffffu, slot);
      if ((m & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) {
        volatile unsigned long long *p = reinterpret_cast<unsigned long long *>(sm) + slot;
        *p = *p + (unsigned long long)__popc(m);
      }
    } else if (MODE == 6) {  // ATOMS.CAS 32
      acc += atomicCAS(sm + r % WORDS, 0u, r | 1u);
    } else if (MODE == 7) {  // LDS.128 + LDS.128(key) + LDS.64/STS.64: the full warp-private step
      const uint32_t b = r % (WORDS / 16);
      const uint4 t = lds128(sm + b * 4);
      const uint4 kk = lds128(sm + WORDS / 4 + b * 4);
      acc += t.x ^ kk.y;
      volatile unsigned long long *p = reinterpret_cast<unsigned long long *>(sm + WORDS / 2) + b;
      *p = *p + 1ull;
    } else if (MODE == 8) {  // ATOMS.ADD with the result used (returning form)
      acc += atomicAdd(sm + r % WORDS, 1u);
    } else if (MODE == 9) {  // ALU only: the loop overhead to subtract
      acc += r;
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  out[blockIdx.x * THREADS + threadIdx.x] = acc + sm[threadIdx.x];
}

template <int MODE>
static void run(const char *name, int n_sm, uint32_t *out, long long *cyc) {
  const size_t smem = WORDS * 4;
  cudaFuncSetAttribute(bench<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int grid = n_sm * 2;
  bench<MODE><<<grid, THREADS, smem>>>(out, cyc);
  cudaEvent_t a, b;
  cudaEventCreate(&a), cudaEventCreate(&b);
  cudaEventRecord(a);
  bench<MODE><<<grid, THREADS, smem>>>(out, cyc);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  long long h[2048];
  cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < grid; i++) avg += (double)h[i];
  avg /= grid;
  // per SM: 2 CTAs x 12 warps x ITERS warp-steps in `avg` cycles
  printf("%-44s %8.3f ms  %8.2f SM-cycles per warp step  (%s)\n", name, ms, avg / (24.0 * ITERS), cudaGetErrorString(cudaGetLastError()));
}

int main() {
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  uint32_t *out;
  long long *cyc;
  cudaMalloc(&out, 2048 * THREADS * 4);
  cudaMalloc(&cyc, 2048 * 8);
  printf("%s, %d SMs\n", pr.name, pr.multiProcessorCount);
  run<9>("ALU only (loop overhead)", pr.multiProcessorCount, out, cyc);
  run<0>("ATOMS.ADD x1 random", pr.multiProcessorCount, out, cyc);
  run<8>("ATOMS.ADD x1 random, result used", pr.multiProcessorCount, out, cyc);
  run<1>("ATOMS.ADD x3 in one slot", pr.multiProcessorCount, out, cyc);
  run<6>("ATOMS.CAS x1 random", pr.multiProcessorCount, out, cyc);
  run<2>("LDS.32 + STS.32", pr.multiProcessorCount, out, cyc);
  run<3>("LDS.64 + STS.64", pr.multiProcessorCount, out, cyc);
  run<4>("LDS.128", pr.multiProcessorCount, out, cyc);
  run<5>("MATCH.ANY + leader LDS.64/STS.64", pr.multiProcessorCount, out, cyc);
  run<7>("LDS.128 x2 + LDS.64/STS.64", pr.multiProcessorCount, out, cyc);
  return 0;
}
