// kc_device.cuh -- device-side building blocks shared by the sm_100a kernels of the k-mer analysis stage.
//
// Data layouts:
//   * packed read / k-mer words: base i at bits 2*(31 - i%32) of u64 word i/32, A0 C1 G2 T3 -- the layout of
//     the reference's Kmer<MAX_K> (src/kmer.cpp:303-325), so key words compare and dump identically.
//   * supermer record: fixed R = 16/32/64 bytes = W u32 words:
//        [0, nbw)            bases, 16 per word, base j at bits 2*(15 - j%16) of word j/16
//        [nbw, W-1] (+ the top half of word W-1 when R == 32)   "high quality" bit per base, base j at bit 31 - j%32
//        word W-1, low 16 bits  number of counted k-mers n (0 = empty record)
//     base 0 is the left extension of the first k-mer; the record carries n + k + 1 bases.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace kc {

// ---------------------------------------------------------------------------------------------- hashes
// quick_hash of the reference (src/hash_funcs.c:332-342): owner rank = quick_hash(minimizer) % n_ranks
__host__ __device__ __forceinline__ uint64_t quick_hash(uint64_t v) {
  v = v * 3935559000370003845ULL + 2691343689449507681ULL;
  v ^= v >> 21;
  v ^= v << 37;
  v ^= v >> 4;
  v *= 4768777513237032717ULL;
  v ^= v << 20;
  v ^= v >> 41;
  v ^= v << 5;
  return v;
}

// slot hash of a key: the table layout is not part of parity, so this is a
// cheaper mixer than the reference's Murmur3 x64_128: one multiply-xorshift round per word + a finaliser.
template <int N>
__host__ __device__ __forceinline__ uint64_t key_hash(const uint64_t (&w)[N]) {
  uint64_t h = 0x9e3779b97f4a7c15ULL;
#pragma unroll
  for (int i = 0; i < N; i++) {
    h = (h ^ w[i]) * 0xff51afd7ed558ccdULL;
    h ^= h >> 32;
  }
  h *= 0xc4ceb9fe1a85ec53ULL;
  h ^= h >> 29;
  return h;
}

// per-entry mix of the order-independent result checksum (same function as oracle/kcount_oracle.c kco_entry_mix)
__host__ __device__ __forceinline__ uint64_t entry_mix(const uint64_t *key, int n_words, uint16_t count, char left, char right,
                                                       const uint16_t *ext8) {
  uint64_t h = 0x9e3779b97f4a7c15ULL;
  for (int i = 0; i < n_words; i++) {
    h = (h ^ key[i]) * 0xff51afd7ed558ccdULL;
    h ^= h >> 32;
  }
  uint64_t v = (uint64_t)count | ((uint64_t)(uint8_t)left << 16) | ((uint64_t)(uint8_t)right << 24);
  h = (h ^ v) * 0xc4ceb9fe1a85ec53ULL;
  h ^= h >> 29;
  if (ext8) {
    uint64_t e0 = (uint64_t)ext8[0] | ((uint64_t)ext8[1] << 16) | ((uint64_t)ext8[2] << 32) | ((uint64_t)ext8[3] << 48);
    uint64_t e1 = (uint64_t)ext8[4] | ((uint64_t)ext8[5] << 16) | ((uint64_t)ext8[6] << 32) | ((uint64_t)ext8[7] << 48);
    h = (h ^ e0) * 0xff51afd7ed558ccdULL;
    h ^= h >> 31;
    h = (h ^ e1) * 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 30;
  }
  return h;
}

// exact x % d and x / d for a run-time 32-bit divisor without the 64-bit divide sequence:
// q = mulhi(x, floor(2^64 / d)) under-estimates the quotient by at most 2
struct FastDiv {
  uint64_t magic;  // floor(2^64 / d), or 0 for d == 1
  uint32_t d;
};
__host__ inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  f.magic = d <= 1 ? 0 : (uint64_t)(((__uint128_t)1 << 64) / d);
  return f;
}
__device__ __forceinline__ void fast_divmod(uint64_t x, const FastDiv &f, uint64_t &q, uint32_t &r) {
  if (f.d <= 1) {
    q = x;
    r = 0;
    return;
  }
  q = __umul64hi(x, f.magic);
  uint64_t rem = x - q * f.d;
  while (rem >= f.d) {
    rem -= f.d;
    q++;
  }
  r = (uint32_t)rem;
}

// ---------------------------------------------------------------------------------------------- 2-bit words
// interleave two 32-bit planes: result bit 2i+1 = hi bit i, bit 2i = lo bit i
__device__ __forceinline__ uint64_t spread32(uint32_t x) {
  uint64_t v = x;
  v = (v | (v << 16)) & 0x0000ffff0000ffffULL;
  v = (v | (v << 8)) & 0x00ff00ff00ff00ffULL;
  v = (v | (v << 4)) & 0x0f0f0f0f0f0f0f0fULL;
  v = (v | (v << 2)) & 0x3333333333333333ULL;
  v = (v | (v << 1)) & 0x5555555555555555ULL;
  return v;
}

// reverse-complement all 32 bases of a word: base i -> position 31 - i, complemented
// (the reference does this with a 256-entry byte table, src/kmer.cpp:491-500)
__host__ __device__ __forceinline__ uint64_t revcomp_word(uint64_t w) {
#ifdef __CUDA_ARCH__
  uint64_t r = __brevll(w);
#else
  uint64_t r = w;
  r = ((r >> 1) & 0x5555555555555555ULL) | ((r & 0x5555555555555555ULL) << 1);
  r = ((r >> 2) & 0x3333333333333333ULL) | ((r & 0x3333333333333333ULL) << 2);
  r = ((r >> 4) & 0x0f0f0f0f0f0f0f0fULL) | ((r & 0x0f0f0f0f0f0f0f0fULL) << 4);
  r = __builtin_bswap64(r);
#endif
  // brev reversed the two bits inside every base as well: swap them back, then complement
  r = ((r >> 1) & 0x5555555555555555ULL) | ((r & 0x5555555555555555ULL) << 1);
  return ~r;
}

// reverse complement of an N-word k-mer (k % 32 != 0); Kmer::revcomp, src/kmer.cpp:491-510
template <int N>
__host__ __device__ __forceinline__ void revcomp_key(const uint64_t (&in)[N], int k, uint64_t (&out)[N]) {
  uint64_t t[N];
#pragma unroll
  for (int l = 0; l < N; l++) t[l] = revcomp_word(in[N - 1 - l]);
  const int sh = 2 * (32 * N - k);  // 0 < sh < 64
#pragma unroll
  for (int l = 0; l < N - 1; l++) out[l] = (t[l] << sh) | (t[l + 1] >> (64 - sh));
  out[N - 1] = t[N - 1] << sh;
}

// Kmer::operator<, src/kmer.cpp:271-277
template <int N>
__host__ __device__ __forceinline__ bool key_less(const uint64_t (&a)[N], const uint64_t (&b)[N]) {
#pragma unroll
  for (int i = 0; i < N; i++) {
    if (a[i] < b[i]) return true;
    if (a[i] > b[i]) return false;
  }
  return false;
}

// 64-bit window of 32 bases starting at base `pos` of a packed u64 array (needs w[pos/32 + 1] readable)
__device__ __forceinline__ uint64_t window64(const uint64_t *w, int pos) {
  int q = pos >> 5, s = (pos & 31) * 2;
  uint64_t hi = w[q];
  if (s == 0) return hi;
  return (hi << s) | (w[q + 1] >> (64 - s));
}

// ---------------------------------------------------------------------------------------------- record geometry
struct RecordGeom {
  int bytes;      // R
  int words;      // W = R / 4
  int nbw;        // base words
  int cap_bases;  // 32, 80 or 160
  int max_kmers;  // cap_bases - k - 1
};
__host__ __device__ inline RecordGeom record_geom(int R, int k) {
  RecordGeom g;
  g.bytes = R;
  g.words = R / 4;
  g.cap_bases = R == 16 ? 32 : (R == 32 ? 80 : 160);
  g.nbw = g.cap_bases / 16;
  g.max_kmers = g.cap_bases - k - 1;
  return g;
}
// base j and its quality bit, from a record held as u32 words
__device__ __forceinline__ uint32_t rec_base(const uint32_t *rec, int j) { return (rec[j >> 4] >> (2 * (15 - (j & 15)))) & 3u; }
__device__ __forceinline__ uint32_t rec_hq(const uint32_t *rec, int nbw, int j) { return (rec[nbw + (j >> 5)] >> (31 - (j & 31))) & 1u; }

// 32 bases (a u64 key word) starting at base `pos` of a record's base words
__device__ __forceinline__ uint64_t rec_window64(const uint32_t *rec, int pos) {
  int q = pos >> 4, s = (pos & 15) * 2;
  uint32_t w0 = rec[q], w1 = rec[q + 1], w2 = rec[q + 2];
  uint32_t hi = __funnelshift_l(w1, w0, s), lo = __funnelshift_l(w2, w1, s);
  return ((uint64_t)hi << 32) | lo;
}

// ---------------------------------------------------------------------------------------------- memory ops
__device__ __forceinline__ uint4 ld_stream16(const void *p) {  // read-once data: bypass L1
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void *p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_cg16(const void *p) {  // coherent at L2 (table probes race with atomics)
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint64_t ld_cg64(const void *p) {
  uint64_t r;
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint2 ld_cg_u2(const void *p) {
  uint2 r;
  asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint32_t ld_acquire32(const void *p) {
  uint32_t r;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_release32(void *p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add64(void *p, uint64_t v) {
  asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_add32(void *p, uint32_t v) {
  asm volatile("red.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// 128-bit compare-and-swap (sm_90+): returns the previous value of the 16-byte aligned unit at p
__device__ __forceinline__ ulonglong2 cas128(void *p, ulonglong2 cmp, ulonglong2 val) {
  ulonglong2 old;
  asm volatile(
      "{\n\t.reg .b128 c, v, o;\n\t"
      "mov.b128 c, {%2, %3};\n\t"
      "mov.b128 v, {%4, %5};\n\t"
      "atom.global.cas.b128 o, [%6], c, v;\n\t"
      "mov.b128 {%0, %1}, o;\n\t}"
      : "=l"(old.x), "=l"(old.y)
      : "l"(cmp.x), "l"(cmp.y), "l"(val.x), "l"(val.y), "l"(p)
      : "memory");
  return old;
}
__device__ __forceinline__ ulonglong2 ld_cg128(const void *p) {
  uint4 r = ld_cg16(p);
  ulonglong2 v;
  v.x = ((uint64_t)r.y << 32) | r.x;
  v.y = ((uint64_t)r.w << 32) | r.z;
  return v;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

// ---------------------------------------------------------------------------------------------- extension rule
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


}  // namespace kc
