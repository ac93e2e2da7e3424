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
// routes supermers by minimizer hash into so many local bins that the DISTINCT k-mers of one bin fit a shared-memory
// table (all k-mers sharing a minimizer land in the same bin, so the bins partition the key space). One CTA owns a bin:
//   1. (the table is clean: the sweeps of the previous bin clear what they visit)
//   2. stage the bin's records (from every source rank's region) in shared memory, a chunk at a time, prefix-sum
//      their k-mer counts over the whole CTA and give every THREAD an equal contiguous run of k-mer occurrences; a
//      thread walks its run with a rolling forward / reverse-complement window and counts with shared-memory
//      atomics: a 32-bit tag CAS claims a slot, the claimant publishes the key words, three ATOMS.ADD update
//      count / left / right
//   3. sweep the table: purge count < 2, resolve both extensions with the reference rule, drop X / F, and append the
//      survivors (and, when asked, all count >= 2 entries with their 8 counters) to the dense result arrays with one
//      global cursor update per CTA and bin.
// HBM traffic is the records read once and the survivors written once. A bin whose distinct k-mers do not fit (or that
// holds more than kCountMaxHeavy keys beyond 65535 occurrences) is split by hash bits and its records are walked once
// per part -- exact, only slower for that bin.
#include "kc_internal.h"

namespace kc {

static const int CT_THREADS = 384;
static const int CT_WARPS = CT_THREADS / 32;
static const int CT_STACK = 48;  // pending (split, residue) parts of a bin
static const uint32_t C_NOT_FOUND = 0xffffffffu;
static const uint32_t C_NOT_READY = 0xfffffffeu;
static const uint32_t C_SKIP = 0xfffffffdu;

// per-slot shared memory: tag 4 + count 4 + packed extensions 16 + key 8N + entry of the occupied-slot list 2
static inline size_t count_slot_bytes(int n_words) { return 26 + 8 * (size_t)n_words; }
// staging area of a CTA: a chunk of up to CAP records of the bin plus the prefix sums of their k-mer counts
static const int CT_STAGE_BYTES = 16 * 1024;
static inline int count_chunk_records(int rec_bytes) { return CT_STAGE_BYTES / rec_bytes; }
static inline size_t count_stage_bytes(int rec_bytes) { return CT_STAGE_BYTES + ((size_t)count_chunk_records(rec_bytes) + 4) * sizeof(uint32_t); }
static const size_t CT_SMEM_BUDGET = 110 * 1024;  // two CTAs per SM

uint32_t count_table_slots(int n_words, int rec_bytes) {
  size_t avail = CT_SMEM_BUDGET - count_stage_bytes(rec_bytes) - 4608 /* static control block */;
  uint32_t s = (uint32_t)(avail / count_slot_bytes(n_words));
  if (s > 65532u) s = 65532u;
  return s & ~3u;
}

struct CountCtl {  // static shared control block of a CTA
  const uint8_t *seg_ptr[kCountMaxSrc];
  uint32_t seg_n[kCountMaxSrc];
  uint32_t seg_start[kCountMaxSrc + 1];  // exclusive prefix of seg_n: the bin's records as one index space
  uint32_t stack_split[CT_STACK], stack_res[CT_STACK];
  int stack_n;
  uint32_t bin, n_rec;
  uint32_t warp_sum[CT_WARPS];
  uint32_t split_mask, split_res;
  uint32_t claimed, overflow, n_heavy;
  uint32_t n_fin, n_ext, cur_fin, cur_ext;
  unsigned long long base_fin, base_ext;
  uint32_t heavy_slot[kCountMaxHeavy];
  uint32_t heavy_count[kCountMaxHeavy];
  uint32_t heavy_side[kCountMaxHeavy * 8];
};

template <int N>
struct Table {  // views into the dynamic shared memory
  uint32_t *tag;   // 0 = empty, fp << 1 = claimed (key being written), fp << 1 | 1 = ready; after the sweep: result code
  uint32_t *cnt;   // occurrences (kHeavyMark | index once a heavy key has been listed)
  uint32_t *ext;   // [slot * 4 + {0: lA|lC<<16, 1: lG|lT<<16, 2: rA|rC<<16, 3: rG|rT<<16}]
  uint64_t *key;   // [slot * N + i]
  uint16_t *list;  // slots claimed in the current pass, in claim order (ctl.claimed entries)
  uint32_t n_slots;
};

static const uint32_t C_PENDING = 0xfffffffcu;

// slot hash: one multiply-xorshift round per key word (the table layout is not part of parity). The low half picks the
// bucket (mulhi), bits 34.. are the fingerprint, bits 48.. split an oversized bin.
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

// Warp-converged probe over 4-slot buckets: every lane of the warp calls this together; lanes with `mine` look their
// key up, claiming the first empty slot of the bucket when CLAIM. One 16-byte load fetches the four tags of a bucket, so
// at the loads this table runs at almost every lookup ends in the first bucket -- which matters because a warp step
// costs the LONGEST probe sequence of its 32 lanes. The loop is driven by a warp vote: the lanes re-converge after
// every probe step and the table updates that follow are issued for all lanes at once. A slot is never freed during
// a pass and a claim always takes the first empty slot of the bucket, so equal keys racing for a bucket contend for
// the same slot and exactly one of them wins.
// Returns the slot, C_NOT_FOUND (find-only miss, or table full -> overflow flagged), C_NOT_READY (the claimant has not
// published the key yet: retry the occurrence on the next step) or C_SKIP (!mine).
template <int N, bool CLAIM>
__device__ __forceinline__ uint32_t probe_converged(const Table<N> &tb, CountCtl &ctl, const uint64_t (&key)[N], uint64_t h, bool mine,
                                                    bool &is_new) {
  const uint32_t fp = ((uint32_t)(h >> 34) | 0x20000000u) << 1;  // 30-bit fingerprint, never 0, bit 0 free for "ready"
  const uint32_t nb = tb.n_slots >> 2;
  uint32_t b = __umulhi((uint32_t)h, nb);
  uint32_t slot = mine ? C_PENDING : C_SKIP;
  uint32_t probes = 0;
  is_new = false;
  while (__any_sync(0xffffffffu, slot == C_PENDING)) {
    if (slot == C_PENDING) {
      const uint4 v = lds_volatile16(tb.tag + 4 * b);
      const uint32_t tv[4] = {v.x, v.y, v.z, v.w};
      // first empty slot and first slot carrying this fingerprint, branch-free
      const int empty = tv[0] == 0u ? 0 : (tv[1] == 0u ? 1 : (tv[2] == 0u ? 2 : (tv[3] == 0u ? 3 : -1)));
      const bool m0 = (tv[0] & ~1u) == fp, m1 = (tv[1] & ~1u) == fp, m2 = (tv[2] & ~1u) == fp, m3 = (tv[3] & ~1u) == fp;
      int jm = m0 ? 0 : (m1 ? 1 : (m2 ? 2 : (m3 ? 3 : -1)));
      if (jm >= 0) {
        const uint32_t t = m0 ? tv[0] : (m1 ? tv[1] : (m2 ? tv[2] : tv[3]));
        if (!(t & 1u)) {
          slot = C_NOT_READY;  // the claimant (maybe a lane of this warp) is about to publish the key
        } else {
          bool same = true;
#pragma unroll
          for (int i = 0; i < N; i++) same = same && *reinterpret_cast<volatile uint64_t *>(tb.key + (size_t)(4 * b + jm) * N + i) == key[i];
          if (same) {
            slot = 4 * b + jm;
          } else {
            // another key with the same 30-bit fingerprint (about once in 10^9 comparisons): check the remaining slots
#pragma unroll
            for (int j = 1; j < 4; j++) {
              if (j <= jm || slot != C_PENDING || (tv[j] & ~1u) != fp) continue;
              if (!(tv[j] & 1u)) {
                slot = C_NOT_READY;
              } else {
                bool eq = true;
#pragma unroll
                for (int i = 0; i < N; i++) eq = eq && *reinterpret_cast<volatile uint64_t *>(tb.key + (size_t)(4 * b + j) * N + i) == key[i];
                if (eq) slot = 4 * b + j;
              }
            }
          }
        }
      }
      if (slot == C_PENDING) {
        if (empty >= 0) {
          if (!CLAIM) {
            slot = C_NOT_FOUND;
          } else {
            const uint32_t i = 4 * b + empty;
            if (atomicCAS(tb.tag + i, 0u, fp) == 0u) {
#pragma unroll
              for (int w = 0; w < N; w++) *reinterpret_cast<volatile uint64_t *>(tb.key + (size_t)i * N + w) = key[w];
              __threadfence_block();
              *reinterpret_cast<volatile uint32_t *>(tb.tag + i) = fp | 1u;
              is_new = true;
              slot = i;
            }
            // else: somebody else took that slot (maybe with this very key): look at the bucket again
          }
        } else {
          b++;
          if (b == nb) b = 0;
          if (++probes >= nb) {  // table completely full
            *reinterpret_cast<volatile uint32_t *>(&ctl.overflow) = 1u;
            slot = C_NOT_FOUND;
          }
        }
      }
    }
  }
  return slot;
}

// ------------------------------------------------------------------------------------------------ chunk walk
// All threads of the CTA: stage records [c0, c0 + n) of the bin (an index space over the concatenated source
// segments), prefix-sum their k-mer counts, split the occurrences evenly over the threads and walk.
template <int N, bool HEAVY, bool MERGE>
__device__ __forceinline__ void walk_chunk(const CountParams &p, const Table<N> &tb, CountCtl &ctl, uint32_t *s_rec, uint32_t *s_pre,
                                           uint32_t c0, int n, int CAP, uint32_t &n_records, unsigned long long &n_occ) {
  const int tid = threadIdx.x;
  const uint32_t lane = lane_id();
  const int warp = tid >> 5;
  const int k = p.k;
  const int W = p.rec_bytes / 4;
  const int U = p.rec_bytes / 16;
  const int nbw = record_geom(p.rec_bytes, k).nbw;
  const int tail = 2 * (32 * N - k);
  const uint64_t last_mask = ~0ULL << tail;
  // ---- stage (coalesced 16-byte streaming loads; a chunk may span sources)
  {
    uint4 *dst = reinterpret_cast<uint4 *>(s_rec);
    const int n16 = n * U;
    for (int u = tid; u < n16; u += CT_THREADS) {
      const uint32_t r = c0 + (uint32_t)(u / U);
      int s = 0;
      while (ctl.seg_start[s + 1] <= r) s++;
      const uint4 *g = reinterpret_cast<const uint4 *>(ctl.seg_ptr[s] + (uint64_t)(r - ctl.seg_start[s]) * p.rec_bytes);
      dst[u] = ld_stream16(g + (u % U));
    }
  }
  __syncthreads();
  // ---- exclusive prefix sum of k-mers per record over the CTA: thread t owns records [t * rpt, t * rpt + rpt)
  const int rpt = (CAP + CT_THREADS - 1) / CT_THREADS;
  uint32_t mine = 0;
  for (int u = 0; u < rpt; u++) {
    const int rr = tid * rpt + u;
    const uint32_t c = rr < n ? (s_rec[rr * W + W - 1] & 0xffffu) : 0u;
    if (c) n_records++;
    mine += c;
  }
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += v;
  }
  if (lane == 31) ctl.warp_sum[warp] = incl;
  __syncthreads();
  uint32_t base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < CT_WARPS; w++) {
    const uint32_t v = ctl.warp_sum[w];
    if (w < warp) base += v;
    total += v;
  }
  {
    uint32_t at = base + incl - mine;
    for (int u = 0; u < rpt; u++) {
      const int rr = tid * rpt + u;
      if (rr <= CAP) s_pre[rr] = at;
      at += rr < n ? (s_rec[rr * W + W - 1] & 0xffffu) : 0u;
    }
    if (tid == CT_THREADS - 1 && CT_THREADS * rpt <= CAP) s_pre[CAP] = total;
  }
  __syncthreads();
  if (tid == 0) n_occ += total;

  // ---- this thread's contiguous run of occurrences [q, qend)
  const uint32_t per = (total + CT_THREADS - 1) / CT_THREADS;
  uint32_t q = min(total, (uint32_t)tid * per);
  const uint32_t qend = min(total, q + per);
  int j = 0, nk = 0;
  const uint32_t *rec = s_rec;
  uint64_t fw[N], rc[N];
  bool fresh = true;
  if (q < qend) {
    int lo = 0, hi = n;  // last record with s_pre[rec] <= q (skips empty records)
    while (hi - lo > 1) {
      int mid = (lo + hi) >> 1;
      if (s_pre[mid] <= q) lo = mid; else hi = mid;
    }
    j = q - s_pre[lo];
    rec = s_rec + lo * W;
    nk = rec[W - 1] & 0xffffu;
  }
  bool done = !(q < qend);
  while (true) {
    // warp-uniform exit; every lane takes part in every step
    if (__all_sync(0xffffffffu, done)) break;
    if (!HEAVY && *reinterpret_cast<volatile uint32_t *>(&ctl.overflow) != 0u) done = true;  // pass abandoned: the bin is split
    uint64_t key[N];
    int el = -1, er = -1;
#pragma unroll
    for (int i = 0; i < N; i++) key[i] = 0;
    if (!done) {
      if (fresh) {  // k-mer j of the record: bases [j+1, j+k], extensions at j and j+k+1
#pragma unroll
        for (int i = 0; i < N; i++) fw[i] = rec_window64(rec, j + 1 + 32 * i);
        fw[N - 1] &= last_mask;
        revcomp_key<N>(fw, k, rc);
        fresh = false;
      }
      const uint32_t lb = rec_base(rec, j), lq = rec_hq(rec, nbw, j);
      const uint32_t rb = rec_base(rec, j + k + 1), rq = rec_hq(rec, nbw, j + k + 1);
      const bool sw = key_less<N>(rc, fw);  // kcount_cpu.cpp:328-333: swap and complement the extensions
#pragma unroll
      for (int i = 0; i < N; i++) key[i] = sw ? rc[i] : fw[i];
      const int lf = lq ? (int)lb : -1, rf = rq ? (int)rb : -1;
      const int lr = rq ? (int)(3u - rb) : -1, rr = lq ? (int)(3u - lb) : -1;
      el = sw ? lr : lf;
      er = sw ? rr : rf;
    }
    const uint64_t h = count_hash<N>(key);
    const bool mine = !done && (((uint32_t)(h >> 48) & ctl.split_mask) == ctl.split_res);  // else: another part of a split bin
    uint32_t slot;
    bool is_new;
    if (!HEAVY) {
      slot = probe_converged<N, true>(tb, ctl, key, h, mine, is_new);
      const uint32_t m_new = __ballot_sync(0xffffffffu, is_new);
      if (m_new) {
        // one counter update per warp step; the new slots go on the occupied-slot list that the sweeps run over; the pass
        // gives up when the table gets too full
        const uint32_t nn = (uint32_t)__popc(m_new);
        uint32_t base = 0;
        if (lane == 0) {
          base = atomicAdd(&ctl.claimed, nn);
          if (base + nn >= p.claim_limit) *reinterpret_cast<volatile uint32_t *>(&ctl.overflow) = 1u;
        }
        base = __shfl_sync(0xffffffffu, base, 0);
        if (is_new) tb.list[base + __popc(m_new & ((1u << lane) - 1u))] = (uint16_t)slot;
      }
      if (slot < C_PENDING) {
        atomicAdd(tb.cnt + slot, 1u);
        if (el >= 0) atomicAdd(tb.ext + slot * 4 + (el >> 1), 1u << (16 * (el & 1)));
        if (er >= 0) atomicAdd(tb.ext + slot * 4 + 2 + (er >> 1), 1u << (16 * (er & 1)));
      }
    } else {  // exact 32-bit recount of the extensions of the keys listed as heavy
      slot = probe_converged<N, false>(tb, ctl, key, h, mine, is_new);
      if (slot < C_PENDING) {
        const uint32_t c = tb.cnt[slot];
        if (c & kHeavyMark) {
          uint32_t *e = ctl.heavy_side + (c & ~kHeavyMark) * 8;
          if (el >= 0) atomicAdd(e + el, 1u);
          if (er >= 0) atomicAdd(e + 4 + er, 1u);
        }
      }
    }
    if (!done && slot != C_NOT_READY) {  // (not ready = the slot's owner has not published the key yet: same state again)
      q++;
      j++;
      if (q >= qend) {
        done = true;
      } else if (j == nk) {  // next non-empty record
        do {
          rec += W;
          nk = rec[W - 1] & 0xffffu;
        } while (nk == 0);
        j = 0;
        fresh = true;
      } else {  // roll: base j+k enters, base j leaves
        const uint64_t b = rec_base(rec, j + k);
#pragma unroll
        for (int i = 0; i < N - 1; i++) fw[i] = (fw[i] << 2) | (fw[i + 1] >> 62);
        fw[N - 1] = (fw[N - 1] << 2) | (b << tail);
#pragma unroll
        for (int i = N - 1; i > 0; i--) rc[i] = (rc[i] >> 2) | (rc[i - 1] << 62);
        rc[0] = (rc[0] >> 2) | ((3ULL - b) << 62);
        rc[N - 1] &= last_mask;
      }
    }
  }
  __syncthreads();  // the staging area is reused
}

// all records of the current bin, a chunk at a time
template <int N, bool HEAVY, bool MERGE>
__device__ __forceinline__ void walk_bin(const CountParams &p, const Table<N> &tb, CountCtl &ctl, uint32_t *s_rec, uint32_t *s_pre, int CAP,
                                         uint32_t &n_records, unsigned long long &n_occ) {
  const uint32_t total = ctl.n_rec;
  const uint32_t n_chunks = (total + CAP - 1) / CAP;
  const uint32_t sz = n_chunks ? (total + n_chunks - 1) / n_chunks : 0;  // chunks of equal size (<= CAP)
  for (uint32_t c0 = 0; c0 < total; c0 += sz) {
    if (!HEAVY && *reinterpret_cast<volatile uint32_t *>(&ctl.overflow)) break;  // uniform: read after a barrier
    walk_chunk<N, HEAVY, MERGE>(p, tb, ctl, s_rec, s_pre, c0, (int)min(sz, total - c0), CAP, n_records, n_occ);
  }
}

// ------------------------------------------------------------------------------------------------ kernel
// result code kept in tag[] between the resolve sweep and the write sweep: left << 8 | right << 16 | flags (never 0)
static const uint32_t RC_FINAL = 1u, RC_EXT = 2u, RC_HEAVY = 4u;

template <int N>
__device__ __forceinline__ void clear_table(const Table<N> &tb) {
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  uint4 *e4 = reinterpret_cast<uint4 *>(tb.ext);
  for (uint32_t i = threadIdx.x; i < tb.n_slots; i += CT_THREADS) e4[i] = z;
  uint4 *t4 = reinterpret_cast<uint4 *>(tb.tag);  // tag and cnt are contiguous: 2 * S words, S % 4 == 0
  for (uint32_t i = threadIdx.x; i < tb.n_slots / 2; i += CT_THREADS) t4[i] = z;
}

// replace the current part of the bin by its two halves (thread 0)
__device__ __forceinline__ void push_halves(CountCtl &ctl, DevCounters *ctr) {
  const uint32_t split = ctl.split_mask + 1u;
  if (ctl.stack_n + 2 > CT_STACK || split >= 32768u) {
    ctr->err_count_stack = 1u;
    return;
  }
  ctl.stack_split[ctl.stack_n] = split * 2;
  ctl.stack_res[ctl.stack_n] = ctl.split_res + split;
  ctl.stack_split[ctl.stack_n + 1] = split * 2;
  ctl.stack_res[ctl.stack_n + 1] = ctl.split_res;
  ctl.stack_n += 2;
}

struct PassStats {
  uint32_t unique, ge2, purged, max_count;
  unsigned long long sum;
};

// resolve one entry with count c >= 2 and its 8 counters: result code, statistics
__device__ __forceinline__ uint32_t resolve_entry(const CountParams &p, const uint32_t (&ext)[8], uint32_t c, uint32_t flags, PassStats &ps,
                                                  uint32_t &my_fin, uint32_t &my_ext) {
  ps.ge2++;
  if (c > ps.max_count) ps.max_count = c;
  const char left = resolve_ext(ext, c, p.dmin_thres, p.one_minus_dyn);
  const char right = resolve_ext(ext + 4, c, p.dmin_thres, p.one_minus_dyn);
  const bool fin = !(left == 'X' || left == 'F' || right == 'X' || right == 'F');  // kcount_cpu.cpp:569-573
  uint32_t code = flags;
  if (fin) {
    ps.sum += c;
    my_fin++;
    code |= RC_FINAL;
  } else {
    ps.purged++;
  }
  if (p.keep_ext) {
    my_ext++;
    code |= RC_EXT;
  }
  return ((uint32_t)(uint8_t)left << 8) | ((uint32_t)(uint8_t)right << 16) | code;
}

template <int N, bool MERGE>
__global__ void __launch_bounds__(CT_THREADS, 2) count_kernel(CountParams p) {
  extern __shared__ __align__(16) uint8_t s_dyn[];
  __shared__ CountCtl ctl;
  const int tid = threadIdx.x;
  const uint32_t lane = lane_id();
  const uint32_t S = p.n_slots;
  const int CAP = CT_STAGE_BYTES / p.rec_bytes;
  uint32_t *s_rec = reinterpret_cast<uint32_t *>(s_dyn);
  uint32_t *s_pre = reinterpret_cast<uint32_t *>(s_dyn + CT_STAGE_BYTES);
  Table<N> tb;
  tb.n_slots = S;
  tb.ext = reinterpret_cast<uint32_t *>(s_dyn + CT_STAGE_BYTES + ((size_t)CAP + 4) * sizeof(uint32_t));  // 16-byte aligned
  tb.key = reinterpret_cast<uint64_t *>(tb.ext + (size_t)S * 4);
  tb.tag = reinterpret_cast<uint32_t *>(tb.key + (size_t)S * N);
  tb.cnt = tb.tag + S;
  tb.list = reinterpret_cast<uint16_t *>(tb.cnt + S);  // S <= 65535
  uint16_t *list = tb.list;

  // statistics of this thread over all its bins
  unsigned long long st_unique = 0, st_ge2 = 0, st_purged = 0, st_sum = 0, st_max = 0, st_occ = 0, st_rec = 0, st_split = 0;

  clear_table<N>(tb);
  // bins are dealt round-robin (they are sized alike by construction); the header entries of the next bin are loaded
  // while the current one is counted, so their latency is off the per-bin critical path
  uint2 e_next = make_uint2(0u, 0u);
  if (tid < p.n_src && blockIdx.x < p.n_bins && p.src_ptr[tid]) e_next = reinterpret_cast<const uint2 *>(p.src_ptr[tid])[blockIdx.x];
  for (uint32_t bin = blockIdx.x; bin < p.n_bins; bin += gridDim.x) {
    __syncthreads();  // the control block of the previous bin is no longer read
    if (tid < p.n_src) {
      const uint8_t *src = p.src_ptr[tid];
      ctl.seg_ptr[tid] = src ? src + p.header_bytes + (uint64_t)e_next.x * p.rec_bytes : nullptr;
      ctl.seg_n[tid] = src ? e_next.y : 0u;
      const uint32_t nb = bin + gridDim.x;
      if (src && nb < p.n_bins) e_next = reinterpret_cast<const uint2 *>(src)[nb];
    }
    __syncthreads();
    if (tid == 0) {
      uint32_t at = 0;
      for (int s = 0; s < p.n_src; s++) {
        ctl.seg_start[s] = at;
        at += ctl.seg_n[s];
      }
      ctl.seg_start[p.n_src] = at;
      ctl.n_rec = at;
      ctl.stack_n = 0;
      if (at) {
        ctl.stack_split[0] = 1;
        ctl.stack_res[0] = 0;
        ctl.stack_n = 1;
      }
    }

    // ---- parts of the bin: (split, residue) = keys with hash bits & (split - 1) == residue; normally just (1, 0)
    while (true) {
      __syncthreads();
      if (tid == 0) {
        if (ctl.stack_n > 0) {
          ctl.stack_n--;
          ctl.split_mask = ctl.stack_split[ctl.stack_n] - 1u;
          ctl.split_res = ctl.stack_res[ctl.stack_n];
        } else {
          ctl.split_mask = 0xffffffffu;  // marker: no part left
        }
        ctl.claimed = ctl.overflow = ctl.n_heavy = 0;
        ctl.n_fin = ctl.n_ext = ctl.cur_fin = ctl.cur_ext = 0;
      }
      __syncthreads();
      if (ctl.split_mask == 0xffffffffu) break;
      // 2. count
      uint32_t nr = 0;
      unsigned long long no = 0;
      walk_bin<N, false, MERGE>(p, tb, ctl, s_rec, s_pre, CAP, nr, no);  // ends with a barrier
      if (ctl.overflow) {
        // the distinct k-mers of this part do not fit: replace it by its two halves
        clear_table<N>(tb);
        if (tid == 0) push_halves(ctl, p.ctr);
        st_split += (tid == 0);
        continue;
      }
      // 3a. resolve, densely over the list of claimed slots: singletons are purged, entries with count >= 2 get their
      //     extensions by the reference rule; the result code replaces the tag (a purged slot keeps a flag-less code, so
      //     that probe chains of a heavy recount still pass through it)
      PassStats ps = {0u, 0u, 0u, 0u, 0ULL};
      uint32_t my_fin = 0, my_ext = 0;
      const uint32_t n_occ_slots = ctl.claimed;
      for (uint32_t e = tid; e < n_occ_slots; e += CT_THREADS) {
        const uint32_t i = list[e];
        const uint32_t c = tb.cnt[i];
        ps.unique++;
        if (c < 2u) {
          ps.purged++;  // kcount_cpu.cpp:557
          tb.tag[i] = 0x100u;
        } else if (c > 65535u) {
          // heavy key: its packed 16-bit extension fields may have carried; list it for the exact recount
          const uint32_t h = atomicAdd(&ctl.n_heavy, 1u);
          if (h < (uint32_t)kCountMaxHeavy) {
            ctl.heavy_slot[h] = i;
            tb.cnt[i] = kHeavyMark | h;
          }
        } else {
          const uint4 v = *reinterpret_cast<const uint4 *>(tb.ext + (size_t)i * 4);
          uint32_t ext[8];
          ext[0] = v.x & 0xffffu, ext[1] = v.x >> 16, ext[2] = v.y & 0xffffu, ext[3] = v.y >> 16;
          ext[4] = v.z & 0xffffu, ext[5] = v.z >> 16, ext[6] = v.w & 0xffffu, ext[7] = v.w >> 16;
          tb.tag[i] = resolve_entry(p, ext, c, 0u, ps, my_fin, my_ext);
        }
      }
      __syncthreads();
      if (ctl.n_heavy > (uint32_t)kCountMaxHeavy) {  // too many for the side table: split, each half holds fewer
        clear_table<N>(tb);
        if (tid == 0) push_halves(ctl, p.ctr);
        st_split += (tid == 0);
        continue;
      }
      if (ctl.n_heavy) {
        for (uint32_t i = tid; i < ctl.n_heavy * 8; i += CT_THREADS) ctl.heavy_side[i] = 0;
        __syncthreads();
        uint32_t nr2 = 0;
        unsigned long long no2 = 0;
        walk_bin<N, true, MERGE>(p, tb, ctl, s_rec, s_pre, CAP, nr2, no2);  // ends with a barrier; result codes match no tag
        if ((uint32_t)tid < ctl.n_heavy) {
          const uint32_t i = ctl.heavy_slot[tid];
          const uint32_t *sd = ctl.heavy_side + tid * 8;
          uint32_t ext[8];
#pragma unroll
          for (int f = 0; f < 8; f++) ext[f] = sd[f] > 65535u ? 65535u : sd[f];  // kcount_cpu.cpp:152-155
          tb.tag[i] = resolve_entry(p, ext, 65535u /* kcount_cpu.cpp:349-351 */, RC_HEAVY, ps, my_fin, my_ext);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        my_fin += __shfl_down_sync(0xffffffffu, my_fin, o);
        my_ext += __shfl_down_sync(0xffffffffu, my_ext, o);
      }
      if (lane == 0) {
        if (my_fin) atomicAdd(&ctl.n_fin, my_fin);
        if (my_ext) atomicAdd(&ctl.n_ext, my_ext);
      }
      // the pass is complete: commit its statistics
      st_unique += ps.unique;
      st_ge2 += ps.ge2;
      st_purged += ps.purged;
      st_sum += ps.sum;
      if (ps.max_count > st_max) st_max = ps.max_count;
      if (ctl.split_res == 0) {  // exactly one part of a bin has residue 0: count its records / occurrences once
        st_rec += nr;
        st_occ += no;
      }
      __syncthreads();
      if (tid == 0) {
        ctl.base_fin = ctl.n_fin ? atomicAdd(&p.ctr->n_final, (unsigned long long)ctl.n_fin) : 0ULL;
        ctl.base_ext = ctl.n_ext ? atomicAdd(&p.ctr->n_ext, (unsigned long long)ctl.n_ext) : 0ULL;
        if (ctl.base_fin + ctl.n_fin > p.cap_final || (p.keep_ext && ctl.base_ext + ctl.n_ext > p.cap_ext)) p.ctr->err_result_full = 1u;
      }
      __syncthreads();
      // 3b. write and clear, densely over the same list: warp-aggregated positions, the survivors of a warp land next to each other
      for (uint32_t e0 = 0; e0 < n_occ_slots; e0 += CT_THREADS) {
        const uint32_t e = e0 + tid;
        const uint32_t i = e < n_occ_slots ? list[e] : 0u;
        const uint32_t code = e < n_occ_slots ? tb.tag[i] : 0u;
        const bool is_fin = (code & RC_FINAL) != 0u, is_ext = (code & RC_EXT) != 0u;
        const uint32_t m_fin = __ballot_sync(0xffffffffu, is_fin), m_ext = __ballot_sync(0xffffffffu, is_ext);
        uint32_t b_fin = 0, b_ext = 0;
        if (lane == 0) {
          if (m_fin) b_fin = atomicAdd(&ctl.cur_fin, (uint32_t)__popc(m_fin));
          if (m_ext) b_ext = atomicAdd(&ctl.cur_ext, (uint32_t)__popc(m_ext));
        }
        b_fin = __shfl_sync(0xffffffffu, b_fin, 0);
        b_ext = __shfl_sync(0xffffffffu, b_ext, 0);
        if (!code) continue;
        const char left = (char)((code >> 8) & 0xffu), right = (char)((code >> 16) & 0xffu);
        const uint32_t c = (code & RC_HEAVY) ? 65535u : tb.cnt[i];
        const uint32_t below = (1u << lane) - 1u;
        if (is_fin) {
          const uint64_t o = ctl.base_fin + b_fin + __popc(m_fin & below);
          if (o < p.cap_final) {
#pragma unroll
            for (int j = 0; j < N; j++) p.out_keys[o * N + j] = tb.key[(size_t)i * N + j];
            p.out_counts[o] = (uint16_t)c;
            p.out_left[o] = left;
            p.out_right[o] = right;
          }
        }
        if (is_ext) {
          uint4 v = *reinterpret_cast<const uint4 *>(tb.ext + (size_t)i * 4);
          if (code & RC_HEAVY) {
            const uint32_t *sd = ctl.heavy_side + (tb.cnt[i] & ~kHeavyMark) * 8;
            uint32_t x[8];
#pragma unroll
            for (int f = 0; f < 8; f++) x[f] = sd[f] > 65535u ? 65535u : sd[f];
            v = make_uint4(x[0] | (x[1] << 16), x[2] | (x[3] << 16), x[4] | (x[5] << 16), x[6] | (x[7] << 16));
          }
          const uint64_t o = ctl.base_ext + b_ext + __popc(m_ext & below);
          if (o < p.cap_ext) {
#pragma unroll
            for (int j = 0; j < N; j++) p.ext_keys[o * N + j] = tb.key[(size_t)i * N + j];
            p.ext_counts[o] = (uint16_t)c;
            p.ext_left[o] = left;
            p.ext_right[o] = right;
            *reinterpret_cast<uint4 *>(p.ext_ext8 + o * 8) = v;  // lA lC lG lT rA rC rG rT as 8 x u16
          }
        }
        tb.tag[i] = 0;
        tb.cnt[i] = 0;
        *reinterpret_cast<uint4 *>(tb.ext + (size_t)i * 4) = make_uint4(0u, 0u, 0u, 0u);
      }
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

template <int N>
static cudaError_t launch_count_n(const CountParams &p, bool merge, int grid, size_t smem, cudaStream_t s) {
  if (merge) {
    cudaFuncSetAttribute(count_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    count_kernel<N, true><<<grid, CT_THREADS, smem, s>>>(p);
  } else {
    cudaFuncSetAttribute(count_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    count_kernel<N, false><<<grid, CT_THREADS, smem, s>>>(p);
  }
  return cudaGetLastError();
}

cudaError_t launch_count(const CountParams &p, int n_words, bool merge, int sm_count, cudaStream_t s, uint64_t *launches) {
  if (p.n_src > kCountMaxSrc) return cudaErrorInvalidValue;
  const size_t smem = count_stage_bytes(p.rec_bytes) + (size_t)p.n_slots * count_slot_bytes(n_words);
  uint64_t grid = (uint64_t)sm_count * 2;
  if (grid > p.n_bins) grid = p.n_bins ? p.n_bins : 1;
  (*launches)++;
  switch (n_words) {
    case 1: return launch_count_n<1>(p, merge, (int)grid, smem, s);
    case 2: return launch_count_n<2>(p, merge, (int)grid, smem, s);
    case 3: return launch_count_n<3>(p, merge, (int)grid, smem, s);
    case 4: return launch_count_n<4>(p, merge, (int)grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace kc
