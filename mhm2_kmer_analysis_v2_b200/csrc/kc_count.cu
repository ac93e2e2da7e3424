// kc_count.cu -- bin-resident k-mer counting: kernels (3) "hash table insert with extension counts" and (5) "purge /
// min-depth / extension resolution / compaction" fused into ONE pass over the supermer records, with the hash table
// living in shared memory.
//
// Replaces, like kc_insert.cu + kc_finalize.cu, HashTableInserter::insert_supermer -> insert_supermer_from_read ->
// get_kmers_and_exts + KmerMapExts::insert (src/kcount/kcount_cpu.cpp:478-493, 339-355, 309-336, 232-271) and
// HashTableInserter::insert_into_local_hashtable + ExtCounts::get_ext (src/kcount/kcount_cpu.cpp:524-601, 173-182), i.e. the
// reference GPU kernels gpu_insert_supermer_block / gpu_purge_invalid / gpu_compact_ht
// (src/kcount/kcount-gpu/gpu_hash_table.cpp:206-268, 427-475).
//
// Why: a table in HBM costs every k-mer occurrence at least two random 32-byte sectors (read + write back) and an L2
// atomic; with ~4-30 occurrences per distinct k-mer almost all of that traffic is avoidable. The parse kernel therefore
// routes supermers by minimizer hash into so many local bins that the DISTINCT k-mers of one bin fit a small
// shared-memory table (all k-mers sharing a minimizer land in the same bin, so the bins partition the key space).
//
// Execution model: ONE WARP OWNS A BIN. Warps never synchronise with each other (no CTA barrier in the kernel); every
// warp has a private slice of shared memory (record stage + table) and walks its bins round-robin:
//   1. stage a chunk of the bin's records (from every source rank's region) in shared memory, lane = record; a warp
//      scan of the k-mer counts plus a bitmap of record starts maps occurrence q -> (record, offset)
//   2. walk: lane l of step s takes occurrence 32 s + l and builds it FROM SCRATCH -- forward window cut out of the
//      record with funnel shifts, reverse complement by bit reversal, canonical choice, extensions, hash. Every lane
//      does identical work on every step: no rolling state, no per-record branches, no divergence.
//   3. probe the warp-private table: 4-slot buckets of 32-bit tags (one 16-byte load), key compare, CAS claim. Because
//      only the lanes of this warp touch the table and they re-converge after every probe round, a claimed slot's key is
//      always published before anybody can look at it again: no "ready" flag, no fences. Three fire-and-forget
//      ATOMS.ADD (count, left word, right word) settle the lanes that hit the same slot in the same step.
//   4. sweep: slots with count >= 2 are compacted into a list, resolved with the reference rule (purge, get_ext, X / F
//      drop) and appended to the dense result with one global cursor update per 32 entries; the table is cleared.
// HBM traffic is the records read once and the survivors written once. A bin whose distinct k-mers do not fit (or that
// holds more than kCountMaxHeavy keys beyond 65535 occurrences) is split by hash bits and its records are walked once
// per part -- exact, only slower for that bin.
#include "kc_internal.h"

namespace kc {

static const int CW_WARPS = 4;        // warps per CTA (independent of each other)
static const int CW_THREADS = CW_WARPS * 32;
static const int CW_CTAS_PER_SM = 5;  // 20 warps per SM (measured: 12 -> 54 ms, 16 -> 48, 20 -> 45.6, 24 -> 48.7 on C2; more warps hide latency, fewer mean larger tables and fewer bins)
static const int CW_ILP = 1;          // k-mer occurrences per lane and step (independent instruction streams hide latency)
static const int CW_STACK = 24;       // pending (split, residue) parts of a bin: up to 8 to start with + one per doubling up to 32768
static const int CW_TMAX = 3072;      // k-mer occurrences per staged chunk (bits of the record-start bitmap)
static const uint32_t C_PENDING = 0xfffffffcu;
static const uint32_t C_NOT_FOUND = 0xffffffffu;
static const uint32_t C_SKIP = 0xfffffffdu;

// shared memory of one warp
struct CountGeom {
  int cap_rec;      // records per staged chunk
  int stage_bytes;  // record stage (also holds the list of slots to resolve during the sweep)
  int off_pre, off_bitmap, off_nz, off_stack, off_heavy, off_table;
  uint32_t slots;
  int warp_bytes;
};

static inline int count_warp_budget() {
  // per SM: 228 KB, 1 KB of it reserved per resident CTA
  return ((228 * 1024) / CW_CTAS_PER_SM - 1024 - 256) / CW_WARPS / 16 * 16;
}
static inline size_t count_slot_bytes(int n_words) { return 24 + 8 * (size_t)n_words; }  // tag 4 + count 4 + extensions 16 + key

__host__ __device__ inline CountGeom count_geom(int n_words, int rec_bytes, int k, int warp_bytes) {
  CountGeom g;
  const int max_kmers = record_geom(rec_bytes, k).max_kmers;
  int cap = rec_bytes == 64 ? 32 : 64;
  if (cap * max_kmers > CW_TMAX) cap = CW_TMAX / (max_kmers > 0 ? max_kmers : 1);
  if (cap < 1) cap = 1;
  g.cap_rec = cap;
  g.stage_bytes = (rec_bytes == 64 ? 32 : 64) * rec_bytes;
  if (g.stage_bytes < 1024) g.stage_bytes = 1024;  // (the sweep keeps its list of u16 slot numbers here)
  g.off_pre = g.stage_bytes;                 // u16 [cap + 1]
  g.off_bitmap = g.off_pre + 144;            // u32 [CW_TMAX / 32]
  g.off_nz = g.off_bitmap + CW_TMAX / 8;     // u8 [cap]
  g.off_stack = g.off_nz + 64;               // u32 [CW_STACK]
  g.off_heavy = g.off_stack + CW_STACK * 4;  // u32 slot[kCountMaxHeavy], u32 side[kCountMaxHeavy * 8]
  g.off_table = g.off_heavy + kCountMaxHeavy * 36;
  g.off_table = (g.off_table + 15) & ~15;
  const int avail = warp_bytes - g.off_table;
  uint32_t s = avail > 0 ? (uint32_t)(avail / (24 + 8 * n_words)) : 0u;
  if (s > (uint32_t)g.stage_bytes / 2u) s = (uint32_t)g.stage_bytes / 2u;  // the sweep's list lives in the record stage
  g.slots = s & ~3u;
  g.warp_bytes = warp_bytes;
  return g;
}

uint32_t count_table_slots(int n_words, int rec_bytes, int k) { return count_geom(n_words, rec_bytes, k, count_warp_budget()).slots; }

template <int N>
struct Table {  // views into the warp's shared memory
  uint32_t *tag;  // 0 = empty, else the key's 32-bit fingerprint (odd)
  uint32_t *cnt;  // occurrences (kHeavyMark | index once a heavy key has been listed)
  uint32_t *ext;  // [w * S + slot], w = {0: lA|lC<<16, 1: lG|lT<<16, 2: rA|rC<<16, 3: rG|rT<<16}
  uint64_t *key;  // [slot * N + i]
  uint32_t n_slots;
};

// slot hash: one multiply-xorshift round per key word (the table layout is not part of parity). The low half picks the
// bucket (mulhi), the high half is the fingerprint, bits 48.. split an oversized bin.
template <int N>
__device__ __forceinline__ uint64_t count_hash(const uint64_t (&w)[N]) {
  uint64_t h = w[0] * 0xff51afd7ed558ccdULL;
  h ^= h >> 32;
#pragma unroll
  for (int i = 1; i < N; i++) {
    h = (h ^ w[i]) * 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 32;
  }
  return h;
}

__device__ __forceinline__ uint4 lds_volatile16(const void *p) {
  uint4 r;
  asm volatile("ld.volatile.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"((uint32_t)__cvta_generic_to_shared(p))
               : "memory");
  return r;
}
__device__ __forceinline__ uint2 lds_volatile8(const void *p) {
  uint2 r;
  asm volatile("ld.volatile.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
  return r;
}

template <int N>
__device__ __forceinline__ bool key_equals(const uint64_t *slot_key, const uint64_t (&key)[N]) {
  bool same = true;
  if (N % 2 == 0) {
#pragma unroll
    for (int i = 0; i < N; i += 2) {
      const uint4 v = lds_volatile16(slot_key + i);
      same = same && v.x == (uint32_t)key[i] && v.y == (uint32_t)(key[i] >> 32) && v.z == (uint32_t)key[i + 1] && v.w == (uint32_t)(key[i + 1] >> 32);
    }
  } else {
#pragma unroll
    for (int i = 0; i < N; i++) {
      const uint2 v = lds_volatile8(slot_key + i);
      same = same && v.x == (uint32_t)key[i] && v.y == (uint32_t)(key[i] >> 32);
    }
  }
  return same;
}

// Warp-converged probe over 4-slot buckets of the warp-private table: every lane of the warp calls this together; lanes
// with `mine` look their key up, claiming an empty slot of the bucket when CLAIM. The loop is driven by a warp vote, so
// the lanes re-converge (and their shared-memory writes are ordered) after every round:
//   * all pending lanes load their bucket's tags in the same instruction, BEFORE any claim of this round; a tag that is
//     seen non-zero was therefore claimed in an earlier round and its key is already in place;
//   * a lane that loses the CAS for an empty slot looks at the same bucket again in the next round and then finds the
//     winner's key (maybe its own key: equal keys racing for a bucket see the same tags and pick the same slot).
// Returns the slot, C_NOT_FOUND (find-only miss, or no room: `full` set) or C_SKIP (!mine).
// ILP keys per lane go through the rounds together (their loads and compares interleave in the instruction stream: the
// kernel is latency-bound, not issue-bound). Per round: all tag loads, then all key compares, then all claims -- so the
// argument above holds across the ILP keys as well (a claim of this round is never seen by a tag load of this round).
template <int N, bool CLAIM, int ILP>
__device__ __forceinline__ void probe_warp(const Table<N> &tb, const uint64_t (&key)[ILP][N], const uint64_t (&h)[ILP], const bool (&mine)[ILP],
                                           uint32_t (&slot)[ILP], bool (&is_new)[ILP], bool &full) {
  const uint32_t nb = tb.n_slots >> 2;
  uint32_t fp[ILP], b[ILP], probes[ILP];
#pragma unroll
  for (int u = 0; u < ILP; u++) {
    fp[u] = (uint32_t)(h[u] >> 32) | 1u;
    b[u] = __umulhi((uint32_t)h[u], nb);
    slot[u] = mine[u] ? C_PENDING : C_SKIP;
    probes[u] = 0;
    is_new[u] = false;
  }
  while (true) {
    bool pending = false;
#pragma unroll
    for (int u = 0; u < ILP; u++) pending = pending || slot[u] == C_PENDING;
    if (!__any_sync(0xffffffffu, pending)) break;
    uint4 v[ILP];
#pragma unroll
    for (int u = 0; u < ILP; u++) v[u] = lds_volatile16(tb.tag + 4 * b[u]);  // (lanes that are done re-read their last bucket: harmless)
    __syncwarp();  // every tag load of this round precedes every claim of this round -- by the programming model, not by luck
    // ---- lookups
#pragma unroll
    for (int u = 0; u < ILP; u++) {
      const bool m0 = v[u].x == fp[u], m1 = v[u].y == fp[u], m2 = v[u].z == fp[u], m3 = v[u].w == fp[u];
      if (slot[u] == C_PENDING && (m0 || m1 || m2 || m3)) {
        const int jm = m0 ? 0 : (m1 ? 1 : (m2 ? 2 : 3));
        if (key_equals<N>(tb.key + (size_t)(4 * b[u] + jm) * N, key[u])) {
          slot[u] = 4 * b[u] + jm;
        } else {
          // another key with the same 32-bit fingerprint in this bucket (about once in 10^9 lookups): check the rest
          const uint32_t tv[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
          for (int j = 1; j < 4; j++)
            if (j > jm && slot[u] == C_PENDING && tv[j] == fp[u] && key_equals<N>(tb.key + (size_t)(4 * b[u] + j) * N, key[u])) slot[u] = 4 * b[u] + j;
        }
      }
    }
    // ---- claims
#pragma unroll
    for (int u = 0; u < ILP; u++) {
      if (slot[u] == C_PENDING) {
        // first empty slot, starting at a fingerprint-dependent position so that different new keys of a bucket rarely
        // fight for the same slot (equal keys still pick the same one)
        const uint32_t r = (fp[u] >> 1) & 3u;
        const uint32_t em = (uint32_t)(v[u].x == 0u) | ((uint32_t)(v[u].y == 0u) << 1) | ((uint32_t)(v[u].z == 0u) << 2) | ((uint32_t)(v[u].w == 0u) << 3);
        if (em) {
          if (!CLAIM) {
            slot[u] = C_NOT_FOUND;
          } else {
            const uint32_t rot = ((em | (em << 4)) >> r) & 0xfu;  // bit t = slot (r + t) & 3 is empty
            const uint32_t i = 4 * b[u] + ((r + (uint32_t)__ffs(rot) - 1u) & 3u);
            if (atomicCAS(tb.tag + i, 0u, fp[u]) == 0u) {
#pragma unroll
              for (int w = 0; w < N; w++) *reinterpret_cast<volatile uint64_t *>(tb.key + (size_t)i * N + w) = key[u][w];
              is_new[u] = true;
              slot[u] = i;
            }
            // else: somebody else took that slot in this round (maybe with this very key): look at the bucket again
          }
        } else {
          b[u]++;
          if (b[u] == nb) b[u] = 0;
          if (++probes[u] >= min(nb, 8u)) {  // a chain of full buckets: the table is (as good as) full
            full = true;
            slot[u] = C_NOT_FOUND;
          }
        }
      }
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------ per-warp state
struct WarpCtx {
  uint32_t *s_rec;
  uint16_t *s_pre;
  uint32_t *s_bitmap;
  uint8_t *s_nz;
  uint32_t *s_stack;
  uint32_t *s_heavy_slot, *s_heavy_side;
  // the bin's record segments, one per source rank: lane l holds source l (and l + 32 in the second set)
  const uint8_t *seg_ptr[2];
  uint32_t seg_start[2], seg_n[2];
  uint32_t n_rec;
  uint32_t split_mask, split_res;
};

// Stage records [c0, c0 + n) of the bin (an index space over the concatenated source segments) and build the
// occurrence map. Returns the number of k-mer occurrences in the chunk; n_nonempty = records that carry any.
template <int R>
__device__ __forceinline__ uint32_t stage_chunk(const CountParams &p, WarpCtx &wc, uint32_t c0, int n, uint32_t &n_nonempty) {
  constexpr int W = R / 4;
  constexpr int U = R / 16;
  const uint32_t lane = lane_id();
  // clear the bitmap (CW_TMAX / 32 words)
  {
    uint4 *bm = reinterpret_cast<uint4 *>(wc.s_bitmap);
    if (lane < CW_TMAX / 128) bm[lane] = make_uint4(0u, 0u, 0u, 0u);
  }
  __syncwarp();
  uint32_t run = 0, rank0 = 0;
  for (int g0 = 0; g0 < n; g0 += 32) {
    const int idx = g0 + (int)lane;
    const bool valid = idx < n;
    const uint32_t r = c0 + (uint32_t)idx;
    const uint8_t *src = nullptr;
    if (p.n_src == 1) {
      src = wc.seg_ptr[0] + (uint64_t)r * R;  // (every lane holds source 0 when there is only one)
    } else {
#pragma unroll
      for (int hs = 0; hs < 2; hs++) {
        const int ns = min(32, p.n_src - 32 * hs);
        for (int s = 0; s < ns; s++) {
          const uint32_t st = __shfl_sync(0xffffffffu, wc.seg_start[hs], s);
          const uint32_t cn = __shfl_sync(0xffffffffu, wc.seg_n[hs], s);
          const unsigned long long bp = __shfl_sync(0xffffffffu, (unsigned long long)wc.seg_ptr[hs], s);
          if (r >= st && r - st < cn) src = reinterpret_cast<const uint8_t *>(bp) + (uint64_t)(r - st) * R;
        }
      }
    }
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; u++) v[u] = make_uint4(0u, 0u, 0u, 0u);
    if (valid && src) {
      if (p.wait_slices) {  // the refine kernel is still writing elsewhere in this buffer: no non-coherent path
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ld_cg16(src + 16 * u);
      } else {
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ld_stream16(src + 16 * u);
      }
    }
    const uint32_t cnt = v[U - 1].w & 0xffu;
    const uint32_t m_nz = __ballot_sync(0xffffffffu, cnt != 0u);
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    const uint32_t excl = run + incl - cnt;
    if (cnt) {
      const uint32_t rank = rank0 + (uint32_t)__popc(m_nz & ((1u << lane) - 1u));
      wc.s_pre[rank] = (uint16_t)excl;
      wc.s_nz[rank] = (uint8_t)idx;
      atomicOr(wc.s_bitmap + (excl >> 5), 1u << (excl & 31u));
    }
    if (valid) {
      uint4 *dst = reinterpret_cast<uint4 *>(wc.s_rec + idx * W);
#pragma unroll
      for (int u = 0; u < U; u++) dst[u] = v[u];
    }
    run += __shfl_sync(0xffffffffu, incl, 31);
    rank0 += (uint32_t)__popc(m_nz);
  }
  n_nonempty = rank0;
  if (lane == 0) wc.s_pre[rank0] = (uint16_t)run;  // (end marker: walk_chunk_split searches s_pre)
  __syncwarp();
  return run;
}

// part id of an occurrence inside a split bin (see "split passes" below): strand-invariant, from the centre of the k-mer
template <int N, int R>
__device__ __forceinline__ uint32_t part_id(const uint32_t *rec, int j, int k) {
  // k-mer bases are record bases j + 1 .. j + k; pairs (j + 1 + i, j + k - i), the np = min(16, k / 2) innermost ones
  const int h = k >> 1, np = h < 16 ? h : 16;
  const int pl = j + 1 + h - np, pr = j + k - h + 1;  // first base of the left / right chunk
  const uint32_t l0 = rec[pl >> 4], l1 = rec[min((pl >> 4) + 1, R / 4 - 1)];
  const uint32_t r0 = rec[pr >> 4], r1 = rec[min((pr >> 4) + 1, R / 4 - 1)];
  const uint32_t L = __funnelshift_l(l1, l0, (pl & 15) * 2), Rr = __funnelshift_l(r1, r0, (pr & 15) * 2);
  // reverse the base order of the right chunk: bit reversal, then swap the two bits of every base back
  uint32_t rv = __brev(Rr & (0xffffffffu << (32 - 2 * np)));  // np bases, now in the low bits, last base first
  rv = ((rv >> 1) & 0x55555555u) | ((rv & 0x55555555u) << 1);
  const uint32_t x = (L >> (32 - 2 * np)) ^ rv;
  return (x * 0x9E3779B1u) >> 16;
}

// One staged chunk: all lanes walk its occurrences, 32 * ILP per step (lane l takes occurrences q0 + 32 u + l).
template <int N, int R, bool HEAVY, int ILP>
__device__ __forceinline__ void walk_chunk(const CountParams &p, const Table<N> &tb, WarpCtx &wc, uint32_t total, uint32_t n_rank,
                                           uint32_t &claimed, bool &overflow) {
  constexpr int W = R / 4;
  constexpr int NBW = (R == 16 ? 32 : (R == 32 ? 80 : 160)) / 16;
  const uint32_t lane = lane_id();
  const uint32_t S = tb.n_slots;
  const int k = p.k;
  const int tail = 2 * (32 * N - k);
  const uint64_t last_mask = ~0ULL << tail;
  // the right extension is base k + 1 of the window that starts at the left extension
  const int rw = (k + 1) >> 4;
  const int rsh = 2 * (15 - ((k + 1) & 15));
  const uint32_t le_mask = 0xffffffffu >> (31u - lane);
  uint32_t rank_run = 0;
  for (uint32_t q0 = 0; q0 < total; q0 += 32 * ILP) {
    uint64_t key[ILP][N], h[ILP];
    uint32_t el[ILP], er[ILP];
    bool mine[ILP];
#pragma unroll
    for (int u = 0; u < ILP; u++) {
      const uint32_t qb = q0 + 32u * u;
      const uint32_t bits = qb < total ? wc.s_bitmap[qb >> 5] : 0u;
      const uint32_t q = qb + lane;
      const bool active = q < total;
      uint32_t rank = rank_run + (uint32_t)__popc(bits & le_mask) - 1u;
      rank_run += (uint32_t)__popc(bits);
      if (rank >= n_rank) rank = n_rank - 1u;  // lanes past the end
      const uint32_t ri = wc.s_nz[rank];
      const int j = active ? (int)(q - wc.s_pre[rank]) : 0;
      const uint32_t *rec = wc.s_rec + ri * W;
      // ---- window of k + 2 bases starting at base j (left extension, k-mer, right extension)
      const int a = j >> 4, s = (j & 15) * 2;
      uint32_t wd[2 * N + 1];
#pragma unroll
      for (int i = 0; i < 2 * N; i++) wd[i] = rec[a + i];
      wd[2 * N] = rec[min(a + 2 * N, W - 1)];
      uint32_t x[2 * N + 1];
#pragma unroll
      for (int i = 0; i < 2 * N; i++) x[i] = __funnelshift_l(wd[i + 1], wd[i], s);
      x[2 * N] = wd[2 * N] << s;
      const uint32_t lb = x[0] >> 30;
      const uint32_t rbw = rw == 2 * N - 2 ? x[2 * N - 2] : (rw == 2 * N - 1 ? x[2 * N - 1] : x[2 * N]);
      const uint32_t rb = (rbw >> rsh) & 3u;
      const uint32_t lq = (rec[NBW + (j >> 5)] >> (31 - (j & 31))) & 1u;
      const int jr = j + k + 1;
      const uint32_t rq = (rec[NBW + (jr >> 5)] >> (31 - (jr & 31))) & 1u;
      uint64_t fw[N], rc[N];
#pragma unroll
      for (int i = 0; i < N; i++) {
        const uint32_t hi = __funnelshift_l(x[2 * i + 1], x[2 * i], 2), lo = __funnelshift_l(x[2 * i + 2], x[2 * i + 1], 2);
        fw[i] = ((uint64_t)hi << 32) | lo;
      }
      fw[N - 1] &= last_mask;
      revcomp_key<N>(fw, k, rc);
      const bool sw = key_less<N>(rc, fw);  // kcount_cpu.cpp:328-333: swap and complement the extensions
#pragma unroll
      for (int i = 0; i < N; i++) key[u][i] = sw ? rc[i] : fw[i];
      // extension letters with their "counts" flag in bit 2; reverse strand: swapped and complemented
      const uint32_t la = lb | (lq << 2), ra = rb | (rq << 2);
      el[u] = sw ? (ra ^ 3u) : la;
      er[u] = sw ? (la ^ 3u) : ra;
      h[u] = count_hash<N>(key[u]);
      mine[u] = active;
      if (wc.split_mask) mine[u] = active && ((part_id<N, R>(rec, j, k) & wc.split_mask) == wc.split_res);  // else: another part of a split bin
    }
    uint32_t slot[ILP];
    bool is_new[ILP], full = false;
    if (!HEAVY) {
      probe_warp<N, true, ILP>(tb, key, h, mine, slot, is_new, full);
#pragma unroll
      for (int u = 0; u < ILP; u++) {
        if (slot[u] < C_PENDING) {
          uint32_t *e = tb.ext + slot[u];
          atomicAdd(tb.cnt + slot[u], 1u);
          if (el[u] & 4u) atomicAdd(e + ((el[u] & 2u) ? S : 0u), 1u << ((el[u] & 1u) << 4));
          if (er[u] & 4u) atomicAdd(e + ((er[u] & 2u) ? 3u * S : 2u * S), 1u << ((er[u] & 1u) << 4));
        }
        claimed += is_new[u];  // (per lane; summed over the warp at the end of the chunk)
      }
      if (__any_sync(0xffffffffu, full)) {  // no room: the pass is abandoned and the bin is split
        overflow = true;
        break;
      }
    } else {  // exact 32-bit recount of the extensions of the keys listed as heavy
      probe_warp<N, false, ILP>(tb, key, h, mine, slot, is_new, full);
#pragma unroll
      for (int u = 0; u < ILP; u++) {
        if (slot[u] < C_PENDING) {
          const uint32_t c = tb.cnt[slot[u]];
          if (c & kHeavyMark) {
            uint32_t *e = wc.s_heavy_side + (c & ~kHeavyMark) * 8;
            if (el[u] & 4u) atomicAdd(e + (el[u] & 3u), 1u);
            if (er[u] & 4u) atomicAdd(e + 4 + (er[u] & 3u), 1u);
          }
        }
      }
    }
  }
  __syncwarp();
}

// ---- split passes -----------------------------------------------------------------------------------------------
// A bin whose distinct k-mers do not fit the table is counted in parts: part (split, res) takes the keys whose PART ID has
// `res` in its low bits. Walking the bin once per part the ordinary way costs a full build + probe for every occurrence in
// every pass -- p parts, p times the work (C3 at 100 x coverage, k = 55: 1.3 M split passes per GPU, the count pass 3.5 x
// its unsplit time). Here a split pass is two stages:
//   A. filter, ~1/8 of the work: the part id of occurrence q comes from the 32 bases around the CENTRE of the k-mer alone:
//      XOR of the symmetric base pairs (b[c-1-i] ^ b[c+i]; complementing and reversing the k-mer maps every pair onto itself,
//      so the id is the same on both strands and needs no canonical k-mer), mixed by one multiply. Occurrences of this part
//      are appended to a 64-entry ring of occurrence numbers;
//   B. whenever 32 are waiting, they are built and probed like an ordinary step -- with every lane busy.
// so the build + probe work of a bin stays ~constant however many parts it needs. (The ring lives in the heavy-key side
// table's shared memory, which is idle during a counting walk; the rare exact recount of keys beyond 65535 occurrences
// walks the ordinary way, with the same part id.)
template <int N, int R>
__device__ __forceinline__ void walk_chunk_split(const CountParams &p, const Table<N> &tb, WarpCtx &wc, uint32_t total, uint32_t n_rank,
                                                 uint32_t &claimed, bool &overflow) {
  constexpr int W = R / 4;
  constexpr int NBW = (R == 16 ? 32 : (R == 32 ? 80 : 160)) / 16;
  const uint32_t lane = lane_id();
  const uint32_t S = tb.n_slots;
  const int k = p.k;
  const int tail = 2 * (32 * N - k);
  const uint64_t last_mask = ~0ULL << tail;
  const int rw = (k + 1) >> 4;
  const int rsh = 2 * (15 - ((k + 1) & 15));
  const uint32_t le_mask = 0xffffffffu >> (31u - lane);
  uint16_t *ring = reinterpret_cast<uint16_t *>(wc.s_heavy_side);  // 64 entries (the side table is 256 bytes)
  uint32_t head = 0, waiting = 0, rank_run = 0;
  const uint32_t n_steps = (total + 31u) >> 5;
  for (uint32_t st = 0; st <= n_steps && !overflow; st++) {
    if (st < n_steps) {
      // ---- A: which occurrences of this step belong to the part
      const uint32_t qb = st << 5, q = qb + lane;
      const uint32_t bits = wc.s_bitmap[qb >> 5];
      const bool active = q < total;
      uint32_t rank = rank_run + (uint32_t)__popc(bits & le_mask) - 1u;
      rank_run += (uint32_t)__popc(bits);
      if (rank >= n_rank) rank = n_rank - 1u;
      const uint32_t ri = wc.s_nz[rank];
      const int j = active ? (int)(q - wc.s_pre[rank]) : 0;
      const uint32_t pid = part_id<N, R>(wc.s_rec + ri * W, j, k);
      const bool mine = active && ((pid & wc.split_mask) == wc.split_res);
      const uint32_t m = __ballot_sync(0xffffffffu, mine);
      if (mine) ring[(head + waiting + (uint32_t)__popc(m & ((1u << lane) - 1u))) & 63u] = (uint16_t)q;
      waiting += (uint32_t)__popc(m);
      __syncwarp();
      if (waiting < 32u && st + 1 < n_steps) continue;
    }
    // ---- B: build and probe up to 32 waiting occurrences (everything that is left after the last step)
    while (waiting && !overflow && (waiting >= 32u || st + 1 >= n_steps)) {
      const uint32_t nb = min(waiting, 32u);
      const bool active = lane < nb;
      const uint32_t q = active ? ring[(head + lane) & 63u] : 0u;
      head = (head + nb) & 63u;
      waiting -= nb;
      // record of occurrence q: last rank with s_pre[rank] <= q (s_pre[n_rank] = total)
      uint32_t lo = 0, hi = n_rank;
#pragma unroll
      for (int it = 0; it < 7; it++) {
        const uint32_t mid = (lo + hi) >> 1;
        if (mid > lo && (uint32_t)wc.s_pre[mid] <= q) lo = mid;
        else if (mid > lo) hi = mid;
      }
      const uint32_t ri = wc.s_nz[lo];
      const int j = (int)(q - wc.s_pre[lo]);
      const uint32_t *rec = wc.s_rec + ri * W;
      // (the build below is walk_chunk's, for one occurrence per lane)
      const int a = j >> 4, sft = (j & 15) * 2;
      uint32_t wd[2 * N + 1];
#pragma unroll
      for (int i = 0; i < 2 * N; i++) wd[i] = rec[a + i];
      wd[2 * N] = rec[min(a + 2 * N, W - 1)];
      uint32_t x[2 * N + 1];
#pragma unroll
      for (int i = 0; i < 2 * N; i++) x[i] = __funnelshift_l(wd[i + 1], wd[i], sft);
      x[2 * N] = wd[2 * N] << sft;
      const uint32_t lb = x[0] >> 30;
      const uint32_t rbw = rw == 2 * N - 2 ? x[2 * N - 2] : (rw == 2 * N - 1 ? x[2 * N - 1] : x[2 * N]);
      const uint32_t rb = (rbw >> rsh) & 3u;
      const uint32_t lq = (rec[NBW + (j >> 5)] >> (31 - (j & 31))) & 1u;
      const int jr = j + k + 1;
      const uint32_t rq = (rec[NBW + (jr >> 5)] >> (31 - (jr & 31))) & 1u;
      uint64_t fw[N], rc[N];
#pragma unroll
      for (int i = 0; i < N; i++) {
        const uint32_t hi32 = __funnelshift_l(x[2 * i + 1], x[2 * i], 2), lo32 = __funnelshift_l(x[2 * i + 2], x[2 * i + 1], 2);
        fw[i] = ((uint64_t)hi32 << 32) | lo32;
      }
      fw[N - 1] &= last_mask;
      revcomp_key<N>(fw, k, rc);
      const bool sw = key_less<N>(rc, fw);
      uint64_t key[1][N], h[1];
#pragma unroll
      for (int i = 0; i < N; i++) key[0][i] = sw ? rc[i] : fw[i];
      const uint32_t la = lb | (lq << 2), ra = rb | (rq << 2);
      const uint32_t el = sw ? (ra ^ 3u) : la, er = sw ? (la ^ 3u) : ra;
      h[0] = count_hash<N>(key[0]);
      bool mine[1] = {active};
      uint32_t slot[1];
      bool is_new[1], full = false;
      probe_warp<N, true, 1>(tb, key, h, mine, slot, is_new, full);
      if (slot[0] < C_PENDING) {
        uint32_t *e = tb.ext + slot[0];
        atomicAdd(tb.cnt + slot[0], 1u);
        if (el & 4u) atomicAdd(e + ((el & 2u) ? S : 0u), 1u << ((el & 1u) << 4));
        if (er & 4u) atomicAdd(e + ((er & 2u) ? 3u * S : 2u * S), 1u << ((er & 1u) << 4));
      }
      claimed += is_new[0];
      if (__any_sync(0xffffffffu, full)) overflow = true;
      __syncwarp();
    }
  }
  __syncwarp();
}

// all records of the current bin, a chunk at a time
template <int N, int R, bool HEAVY>
__device__ __forceinline__ bool walk_bin(const CountParams &p, const Table<N> &tb, WarpCtx &wc, int cap_rec, uint32_t &claimed,
                                         uint32_t &n_records, unsigned long long &n_occ) {
  const uint32_t total = wc.n_rec;
  bool overflow = false;
  for (uint32_t c0 = 0; c0 < total && !overflow; c0 += (uint32_t)cap_rec) {
    uint32_t n_nz = 0;
    const uint32_t occ = stage_chunk<R>(p, wc, c0, (int)min((uint32_t)cap_rec, total - c0), n_nz);
    n_records += n_nz;
    n_occ += occ;
    if (occ) {
      if (!HEAVY && wc.split_mask) walk_chunk_split<N, R>(p, tb, wc, occ, n_nz, claimed, overflow);
      else walk_chunk<N, R, HEAVY, (HEAVY ? 1 : CW_ILP)>(p, tb, wc, occ, n_nz, claimed, overflow);
    }
    // too full to go on efficiently (every lane counts its own claims)
    if (!HEAVY && !overflow && __reduce_add_sync(0xffffffffu, claimed) >= p.claim_limit) overflow = true;
  }
  return overflow;
}

// ------------------------------------------------------------------------------------------------ kernel
template <int N>
__device__ __forceinline__ void clear_table(const Table<N> &tb) {
  // tag, cnt and the four extension words are contiguous: 6 * S words, S % 4 == 0
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  uint4 *t4 = reinterpret_cast<uint4 *>(tb.tag);
  const uint32_t n16 = tb.n_slots * 6u / 4u;
  for (uint32_t i = lane_id(); i < n16; i += 32) t4[i] = z;
  __syncwarp();
}

struct PassStats {
  uint32_t unique, ge2, purged, max_count;
  unsigned long long sum;
};

template <int N, int R>
__global__ void __launch_bounds__(CW_THREADS, CW_CTAS_PER_SM) count_kernel(CountParams p) {
  extern __shared__ __align__(16) uint8_t s_dyn[];
  const uint32_t lane = lane_id();
  const int warp = threadIdx.x >> 5;
  const CountGeom g = count_geom(N, R, p.k, p.warp_bytes);
  uint8_t *base = s_dyn + (size_t)warp * p.warp_bytes;
  const uint32_t S = p.n_slots;
  WarpCtx wc;
  wc.s_rec = reinterpret_cast<uint32_t *>(base);
  wc.s_pre = reinterpret_cast<uint16_t *>(base + g.off_pre);
  wc.s_bitmap = reinterpret_cast<uint32_t *>(base + g.off_bitmap);
  wc.s_nz = base + g.off_nz;
  wc.s_stack = reinterpret_cast<uint32_t *>(base + g.off_stack);
  wc.s_heavy_slot = reinterpret_cast<uint32_t *>(base + g.off_heavy);
  wc.s_heavy_side = wc.s_heavy_slot + kCountMaxHeavy;
  uint16_t *s_list = reinterpret_cast<uint16_t *>(wc.s_rec);  // slots to resolve; the record stage is idle during the sweep
  Table<N> tb;
  tb.n_slots = S;
  tb.tag = reinterpret_cast<uint32_t *>(base + g.off_table);
  tb.cnt = tb.tag + S;
  tb.ext = tb.cnt + S;
  tb.key = reinterpret_cast<uint64_t *>(tb.ext + (size_t)4 * S);

  // statistics of this lane over all its bins
  unsigned long long st_unique = 0, st_ge2 = 0, st_purged = 0, st_sum = 0, st_max = 0, st_occ = 0, st_rec = 0, st_split = 0;

  clear_table<N>(tb);
  const uint32_t n_warps = gridDim.x * CW_WARPS;
  // Single source (the normal case: kc_refine.cu merged the sources): the header entries of the next two bins are kept
  // in flight, and the records of the next bin are pulled into L2 while this one is counted, so that a warp never waits
  // for DRAM at the top of a bin.
  const uint8_t *src0 = p.src0 ? p.src0 : (p.n_src == 1 ? p.src_ptr[0] : nullptr);
  const uint2 *hdr0 = reinterpret_cast<const uint2 *>(src0);
  const bool single = p.n_src == 1 && hdr0 != nullptr;
  // SM-partitioned pipeline: fine bins below ready_end are known to be refined; a warp that gets past it waits for the
  // next slice's counter (all lanes acquire it). An entry that may not be read yet is marked kNotYet and fetched at its turn.
  const uint2 kNotYet = make_uint2(0u, 0xffffffffu);
  uint32_t ready_end = p.wait_slices ? 0u : 0xffffffffu;
  int next_slice = 0;
  auto wait_for = [&](uint64_t b) {
    while (b >= ready_end && next_slice < p.wait_slices) {
      const uint32_t need = p.slice_need[next_slice];
      if (lane == 0) {
        // one poller per warp with a growing pause: thousands of warps wait on this one word, and the refine kernel's
        // atomics have to get through to it (polling every 256 ns from every lane made the refine pass 100x slower)
        uint32_t pause = 1000, spins = 0;
        while (ld_acquire32(p.slice_done + next_slice) < need) {
          __nanosleep(pause);
          if (pause < 16000u) pause *= 2;
          if (++spins > (1u << 18)) {  // seconds: the refine kernel is not running beside us
            *p.err_wait = 1u;
            break;
          }
        }
      }
      __syncwarp();
      (void)ld_acquire32(p.slice_done + next_slice);  // every lane acquires what lane 0 saw
      __syncwarp();
      ready_end = p.slice_fine_edge[next_slice + 1];
      next_slice++;
    }
  };
  auto load_entry = [&](uint64_t b) -> uint2 { return p.wait_slices ? ld_cg_u2(hdr0 + b) : hdr0[b]; };
  uint2 e_pre1 = make_uint2(0u, 0u), e_pre2 = make_uint2(0u, 0u);
  {
    const uint64_t b0 = (uint64_t)p.bin_begin + blockIdx.x * CW_WARPS + warp;
    if (single && b0 < p.bin_end) {
      wait_for(b0);
      e_pre1 = load_entry(b0);
    }
    if (single && b0 + n_warps < p.bin_end) e_pre2 = b0 + n_warps < ready_end ? load_entry(b0 + n_warps) : kNotYet;
  }
  float presplit_thr = (float)p.presplit_rec;  // records beyond which a bin starts out split (warp-uniform, adapts: see below)
  for (uint64_t bin64 = (uint64_t)p.bin_begin + blockIdx.x * CW_WARPS + warp; bin64 < p.bin_end; bin64 += n_warps) {
    const uint32_t bin = (uint32_t)bin64;
    // ---- the bin's segments: header entry {first record, n records} of every source region
    uint32_t total = 0;
    if (single) {
      uint2 e = e_pre1;
      if (e.y == kNotYet.y) {  // (the slice was not finished when this entry was due)
        wait_for(bin64);
        e = load_entry(bin64);
      }
      e_pre1 = e_pre2;
      if (e_pre1.y != kNotYet.y) {
        const uint8_t *nx = src0 + p.header_bytes + (uint64_t)e_pre1.x * R;
        const uint32_t n_lines = (e_pre1.y * R + 127u) >> 7;
        for (uint32_t l = lane; l < n_lines; l += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + (size_t)l * 128));
      }
      const uint64_t b2 = (uint64_t)bin + 2ull * n_warps;
      e_pre2 = b2 < p.bin_end ? (b2 < ready_end ? load_entry(b2) : kNotYet) : make_uint2(0u, 0u);
      wc.seg_ptr[0] = src0 + p.header_bytes + (uint64_t)e.x * R;  // (every lane mirrors source 0)
      wc.seg_ptr[1] = nullptr;
      wc.seg_n[0] = e.y;
      wc.seg_n[1] = 0;
      wc.seg_start[0] = wc.seg_start[1] = 0;
      total = e.y;
    } else {
#pragma unroll
      for (int hs = 0; hs < 2; hs++) {
        const int s = hs * 32 + (int)lane;
        const uint8_t *src = (p.n_src == 1) ? p.src_ptr[0] : (s < p.n_src ? p.src_ptr[s] : nullptr);
        uint2 e = make_uint2(0u, 0u);
        if (src && (hs == 0 || p.n_src > 32)) e = reinterpret_cast<const uint2 *>(src)[bin];
        if (!src || (hs == 1 && p.n_src <= 32)) e.y = 0u;
        wc.seg_ptr[hs] = src ? src + p.header_bytes + (uint64_t)e.x * R : nullptr;
        uint32_t cn = (p.n_src == 1 && (hs == 1)) ? 0u : e.y;
        if (p.n_src == 1) {  // every lane mirrors source 0
          wc.seg_n[hs] = cn;
          wc.seg_start[hs] = 0;
          if (hs == 0) total = cn;
        } else {
          uint32_t incl = cn;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((int)lane >= o) incl += t;
          }
          wc.seg_n[hs] = cn;
          wc.seg_start[hs] = total + incl - cn;
          total += __shfl_sync(0xffffffffu, incl, 31);
        }
      }
    }
    wc.n_rec = total;
    if (total == 0) continue;
    // A bin with so many records that its distinct k-mers will not fit (p.presplit_rec: the host's estimate from the claim
    // limit, the k-mers per record and the distinct / occurrence ratio) starts out split in 2, 4 or 8 parts instead of
    // finding that out by a failed pass. Any split is exact; a wrong guess only costs time.
    // The threshold is a per-warp estimate that corrects itself (the host's figure is only where it starts): the bins whose
    // record count is NEAR it vote -- one that would have fit unsplit raises it by 2 %, one that would not lowers it by 4 % --
    // so it settles where a bin of that size overflows one time in three: a failed pass of a long-k bin (full build + probe
    // of most of its occurrences) costs about twice what a needless split does (staging, filter and sweep of the extra part;
    // measured: with 4 % / 4 % the count pass of k = 77 took 79.0 ms against 75.9 ms with the host's threshold held fixed).
    // Bins far above it that would have fit raise it too (a threshold that is far too low must not stay there).
    uint32_t s0 = 1;
    if (p.presplit_rec)
      while (s0 < 8u && (float)total > (float)s0 * presplit_thr) s0 <<= 1;
    uint32_t bin_unique = 0;
    bool bin_overflowed = false;
    int stack_n = (int)s0;
    if (lane < s0) wc.s_stack[lane] = (s0 << 16) | (s0 - 1u - lane);  // (split << 16) | residue; residue 0 on top
    __syncwarp();

    // ---- parts of the bin: keys with hash bits & (split - 1) == residue; normally just (1, 0)
    while (stack_n > 0) {
      stack_n--;
      const uint32_t part = wc.s_stack[stack_n];
      __syncwarp();
      wc.split_mask = (part >> 16) - 1u;
      wc.split_res = part & 0xffffu;
      // 2. count
      uint32_t claimed = 0, nr = 0;
      unsigned long long no = 0;
      bool overflow = walk_bin<N, R, false>(p, tb, wc, g.cap_rec, claimed, nr, no);
      // 3a. scan the table: statistics, list of the slots with count >= 2, heavy keys
      uint32_t n_list = 0, n_heavy = 0;
      PassStats ps = {0u, 0u, 0u, 0u, 0ULL};
      if (!overflow) {
        for (uint32_t i0 = 0; i0 < S; i0 += 32) {
          const uint32_t i = i0 + lane;
          const uint32_t c = i < S ? tb.cnt[i] : 0u;
          ps.unique += c != 0u;
          ps.purged += c == 1u;  // kcount_cpu.cpp:557
          const bool heavy = c > 65535u;
          const uint32_t m_h = __ballot_sync(0xffffffffu, heavy);
          if (m_h) {
            // heavy key: its packed 16-bit extension fields may have carried; list it for the exact recount
            const uint32_t hi = n_heavy + (uint32_t)__popc(m_h & ((1u << lane) - 1u));
            if (heavy && hi < (uint32_t)kCountMaxHeavy) {
              wc.s_heavy_slot[hi] = i;
              tb.cnt[i] = kHeavyMark | hi;
            }
            n_heavy += (uint32_t)__popc(m_h);
          }
          const bool ge2 = c >= 2u;
          const uint32_t m = __ballot_sync(0xffffffffu, ge2);
          if (ge2) s_list[n_list + (uint32_t)__popc(m & ((1u << lane) - 1u))] = (uint16_t)i;
          n_list += (uint32_t)__popc(m);
        }
        __syncwarp();
        if (n_heavy > (uint32_t)kCountMaxHeavy) overflow = true;  // too many for the side table: split, each half holds fewer
      }
      if (overflow) {
        // the distinct k-mers of this part do not fit: replace it by its two halves
        clear_table<N>(tb);
        const uint32_t split = wc.split_mask + 1u;
        if (stack_n + 2 > CW_STACK || split >= 32768u) {
          if (lane == 0) p.ctr->err_count_stack = 1u;
        } else {
          if (lane == 0) {
            wc.s_stack[stack_n] = ((split * 2) << 16) | (wc.split_res + split);
            wc.s_stack[stack_n + 1] = ((split * 2) << 16) | wc.split_res;
          }
          stack_n += 2;
        }
        __syncwarp();
        st_split += (lane == 0);
        bin_overflowed = true;
        continue;
      }
      if (n_heavy) {
        for (uint32_t i = lane; i < n_heavy * 8; i += 32) wc.s_heavy_side[i] = 0;
        __syncwarp();
        uint32_t c2 = 0, nr2 = 0;
        unsigned long long no2 = 0;
        walk_bin<N, R, true>(p, tb, wc, g.cap_rec, c2, nr2, no2);  // (destroys the list: it shares the record stage)
        n_list = 0;
        for (uint32_t i0 = 0; i0 < S; i0 += 32) {
          const uint32_t i = i0 + lane;
          const bool ge2 = i < S && tb.cnt[i] >= 2u;
          const uint32_t m = __ballot_sync(0xffffffffu, ge2);
          if (ge2) s_list[n_list + (uint32_t)__popc(m & ((1u << lane) - 1u))] = (uint16_t)i;
          n_list += (uint32_t)__popc(m);
        }
        __syncwarp();
      }
      // 3b. resolve the listed entries by the reference rule and append them to the dense result
      for (uint32_t e0 = 0; e0 < n_list; e0 += 32) {
        const uint32_t e = e0 + lane;
        const bool on = e < n_list;
        const uint32_t i = on ? s_list[e] : 0u;
        uint32_t c = on ? tb.cnt[i] : 0u;
        uint32_t ext[8];
        {
          const uint32_t v0 = tb.ext[i], v1 = tb.ext[S + i], v2 = tb.ext[2 * S + i], v3 = tb.ext[3 * S + i];
          ext[0] = v0 & 0xffffu, ext[1] = v0 >> 16, ext[2] = v1 & 0xffffu, ext[3] = v1 >> 16;
          ext[4] = v2 & 0xffffu, ext[5] = v2 >> 16, ext[6] = v3 & 0xffffu, ext[7] = v3 >> 16;
        }
        if (c & kHeavyMark) {
          const uint32_t *sd = wc.s_heavy_side + (c & ~kHeavyMark) * 8;
#pragma unroll
          for (int f = 0; f < 8; f++) ext[f] = sd[f] > 65535u ? 65535u : sd[f];  // kcount_cpu.cpp:152-155
          c = 65535u;                                                             // kcount_cpu.cpp:349-351
        }
        char left = 'X', right = 'X';
        bool fin = false;
        if (on) {
          ps.ge2++;
          if (c > ps.max_count) ps.max_count = c;
          left = resolve_ext(ext, c, p.dmin_thres, p.one_minus_dyn);
          right = resolve_ext(ext + 4, c, p.dmin_thres, p.one_minus_dyn);
          fin = !(left == 'X' || left == 'F' || right == 'X' || right == 'F');  // kcount_cpu.cpp:569-573
          if (fin) ps.sum += c;
          else ps.purged++;
        }
        const bool is_ext = on && p.keep_ext;
        const uint32_t m_fin = __ballot_sync(0xffffffffu, fin), m_ext = __ballot_sync(0xffffffffu, is_ext);
        unsigned long long b_fin = 0, b_ext = 0;
        if (lane == 0) {
          if (m_fin) b_fin = atomicAdd(&p.ctr->n_final, (unsigned long long)__popc(m_fin));
          if (m_ext) b_ext = atomicAdd(&p.ctr->n_ext, (unsigned long long)__popc(m_ext));
        }
        b_fin = __shfl_sync(0xffffffffu, b_fin, 0);
        b_ext = __shfl_sync(0xffffffffu, b_ext, 0);
        const uint32_t below = (1u << lane) - 1u;
        if (fin) {
          const uint64_t o = b_fin + (uint32_t)__popc(m_fin & below);
          if (o < p.cap_final) {
#pragma unroll
            for (int w = 0; w < N; w++) p.out_keys[o * N + w] = tb.key[(size_t)i * N + w];
            p.out_counts[o] = (uint16_t)c;
            p.out_left[o] = left;
            p.out_right[o] = right;
          } else {
            p.ctr->err_result_full = 1u;
          }
        }
        if (is_ext) {
          const uint64_t o = b_ext + (uint32_t)__popc(m_ext & below);
          if (o < p.cap_ext) {
#pragma unroll
            for (int w = 0; w < N; w++) p.ext_keys[o * N + w] = tb.key[(size_t)i * N + w];
            p.ext_counts[o] = (uint16_t)c;
            p.ext_left[o] = left;
            p.ext_right[o] = right;
            *reinterpret_cast<uint4 *>(p.ext_ext8 + o * 8) =
                make_uint4(ext[0] | (ext[1] << 16), ext[2] | (ext[3] << 16), ext[4] | (ext[5] << 16), ext[6] | (ext[7] << 16));  // lA lC lG lT rA rC rG rT
          } else {
            p.ctr->err_result_full = 1u;
          }
        }
      }
      __syncwarp();
      clear_table<N>(tb);
      // the pass is complete: commit its statistics
      if (p.presplit_adapt) bin_unique += __reduce_add_sync(0xffffffffu, ps.unique);
      st_unique += ps.unique;
      st_ge2 += ps.ge2;
      st_purged += ps.purged;
      st_sum += ps.sum;
      if (ps.max_count > st_max) st_max = ps.max_count;
      if (wc.split_res == 0 && lane == 0) {  // exactly one part of a bin has residue 0: count its records / occurrences once
        st_rec += nr;
        st_occ += no;
      }
    }
    if (p.presplit_adapt && (float)total > 0.75f * presplit_thr) {
      // would the unsplit bin have overflowed? (a pre-split bin tells by its distinct k-mers, any other by what happened)
      const bool would = s0 > 1u ? (bin_overflowed || bin_unique > p.claim_limit) : bin_overflowed;
      if (!would) presplit_thr = fminf(presplit_thr * 1.02f, 1e9f);
      else if ((float)total < 1.33f * presplit_thr) presplit_thr = fmaxf(presplit_thr * 0.96f, 2.0f);
    }
  }

  // ---- statistics: warp reduce, one atomic per warp and counter
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    st_unique += __shfl_down_sync(0xffffffffu, st_unique, o);
    st_ge2 += __shfl_down_sync(0xffffffffu, st_ge2, o);
    st_purged += __shfl_down_sync(0xffffffffu, st_purged, o);
    st_sum += __shfl_down_sync(0xffffffffu, st_sum, o);
    st_occ += __shfl_down_sync(0xffffffffu, st_occ, o);
    st_rec += __shfl_down_sync(0xffffffffu, st_rec, o);
    st_split += __shfl_down_sync(0xffffffffu, st_split, o);
    unsigned long long m = __shfl_down_sync(0xffffffffu, st_max, o);
    if (m > st_max) st_max = m;
  }
  if (lane == 0) {
    if (st_unique) atomicAdd(&p.ctr->unique, st_unique);
    if (st_ge2) atomicAdd(&p.ctr->ge2, st_ge2);
    if (st_purged) atomicAdd(&p.ctr->purged, st_purged);
    if (st_sum) atomicAdd(&p.ctr->sum_counts, st_sum);
    if (st_occ) atomicAdd(&p.ctr->occurrences, st_occ);
    if (st_rec) atomicAdd(&p.ctr->records_in, st_rec);
    if (st_split) atomicAdd(&p.ctr->split_passes, st_split);
    if (st_max) atomicMax(&p.ctr->max_count, st_max);
  }
}

template <int N, int R>
static cudaError_t launch_count_nr(const CountParams &p, int grid, size_t smem, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute(count_kernel<N, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  count_kernel<N, R><<<grid, CW_THREADS, smem, s>>>(p);
  return cudaGetLastError();
}

template <int N>
static cudaError_t launch_count_n(const CountParams &p, int grid, size_t smem, cudaStream_t s) {
  switch (p.rec_bytes) {
    case 16: return launch_count_nr<N, 16>(p, grid, smem, s);
    case 32: return launch_count_nr<N, 32>(p, grid, smem, s);
    case 64: return launch_count_nr<N, 64>(p, grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_count(const CountParams &p_in, int n_words, bool /*merge_runs*/, int sm_count, cudaStream_t s, uint64_t *launches) {
  if (p_in.n_src > kCountMaxSrc) return cudaErrorInvalidValue;
  CountParams p = p_in;
  p.warp_bytes = count_warp_budget();
  const CountGeom g = count_geom(n_words, p.rec_bytes, p.k, p.warp_bytes);
  if (g.slots < 8 || p.n_slots != g.slots) return cudaErrorInvalidValue;
  const size_t smem = (size_t)p.warp_bytes * CW_WARPS;
  uint64_t grid = (uint64_t)sm_count * CW_CTAS_PER_SM;
  if (p.bin_end == 0 && p.bin_begin == 0) p.bin_end = p.n_bins;
  const uint64_t need = ((uint64_t)(p.bin_end - p.bin_begin) + CW_WARPS - 1) / CW_WARPS;
  if (grid > need) grid = need ? need : 1;
  (*launches)++;
  switch (n_words) {
    case 1: return launch_count_n<1>(p, (int)grid, smem, s);
    case 2: return launch_count_n<2>(p, (int)grid, smem, s);
    case 3: return launch_count_n<3>(p, (int)grid, smem, s);
    case 4: return launch_count_n<4>(p, (int)grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace kc
