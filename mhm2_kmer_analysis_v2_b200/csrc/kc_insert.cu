// kc_insert.cu -- kernels (3) "k-mer hash table insert with extension counts" and (4) "two-choice singleton filter",
// plus the small planning kernel that turns received supermer regions into tiles and table regions.
//
// Replaces HashTableInserter::insert_supermer -> insert_supermer_from_read -> get_kmers_and_exts + KmerMapExts::insert
// (src/kcount/kcount_cpu.cpp:478-493, 339-355, 309-336, 232-271) and the reference's GPU kernel gpu_insert_supermer_block
// with its TCF calls (src/kcount/kcount-gpu/gpu_hash_table.cpp:310-475, tcf_wrapper.hpp:96-178).
//
// Every warp works on its own: it pulls a tile of supermer records (bin-major order, so the table region being hit stays
// in L2), stages it in its slice of shared memory with 128-bit streaming loads, prefix-sums the k-mer counts with
// shuffles and gives every lane an equal, CONTIGUOUS run of k-mer occurrences. A lane walks its run with a rolling
// k-mer and a rolling reverse complement (one base in, one base out per step), canonicalises, swaps/complements the two
// extension bases (kcount_cpu.cpp:326-333), merges equal consecutive occurrences (homopolymers) and then touches the
// table: linear probing with one 16-byte load per probe, claim with a single CAS (64-bit for N == 1, 128-bit otherwise),
// three fire-and-forget REDs (count, left word, right word). No block barrier, no fence, no returning atomic.
#include "kc_internal.h"

namespace kc {

static const int THREADS = 256;
static const int WARPS_PER_CTA = THREADS / 32;
static const uint32_t MAX_PROBES = 1u << 16;
static const uint64_t NOT_FOUND = ~0ULL;
static const uint64_t NOT_READY = ~0ULL - 1;  // N >= 3: the slot's owner has not completed the key yet -- retry later

int insert_tile_records(int rec_bytes) { return 2048 / rec_bytes; }  // 2 KB of records per warp tile: 128 / 64 / 32

// ------------------------------------------------------------------------------------------------ planning
__global__ void plan_kernel(PlanParams p) {
  // single CTA: (1) records per (bin, source) from the region headers, (2) tiles prefix, (3) table regions
  // proportional to the records each bin received
  extern __shared__ unsigned long long s_bin_records[];  // [local_bins]
  const int header_bytes = ((p.local_bins * 8 + kHeaderAlign - 1) / kHeaderAlign) * kHeaderAlign;
  const int n_seg = p.local_bins * p.n_src;
  for (int b = threadIdx.x; b < p.local_bins; b += blockDim.x) {
    unsigned long long tot = 0;
    for (int s = 0; s < p.n_src; s++) {
      const uint8_t *src = p.src_ptr[s];
      uint2 e = src ? reinterpret_cast<const uint2 *>(src)[b] : make_uint2(0u, 0u);
      Segment sg;
      sg.ptr = src ? src + header_bytes + (uint64_t)e.x * p.rec_bytes : nullptr;
      sg.n_records = e.y;
      sg.bin = b;
      p.segments[b * p.n_src + s] = sg;
      tot += e.y;
    }
    s_bin_records[b] = tot;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // serial prefix sums: local_bins * n_src <= a few 10^4 entries, once per stage
    uint32_t tiles = 0;
    for (int i = 0; i < n_seg; i++) {
      p.seg_tile_start[i] = tiles;
      tiles += (p.segments[i].n_records + p.tile_records - 1) / p.tile_records;
    }
    p.seg_tile_start[n_seg] = tiles;
    unsigned long long total = 0;
    for (int b = 0; b < p.local_bins; b++) total += s_bin_records[b];
    *p.total_records = total;
    const uint64_t min_slots = 64, min_buckets = 4;
    uint64_t avail_slots = p.table_slots - min_slots * p.local_bins;
    uint64_t avail_buckets = p.filter_buckets > min_buckets * p.local_bins ? p.filter_buckets - min_buckets * p.local_bins : 0;
    uint64_t slot_at = 0, bucket_at = 0;
    for (int b = 0; b < p.local_bins; b++) {
      double frac = total ? (double)s_bin_records[b] / (double)total : 1.0 / p.local_bins;
      Region r;
      r.slot_base = slot_at;
      r.n_slots = min_slots + (uint64_t)((double)avail_slots * frac);
      r.bucket_base = bucket_at;
      r.n_buckets = p.filter_buckets ? min_buckets + (uint64_t)((double)avail_buckets * frac) : 0;
      slot_at += r.n_slots;
      bucket_at += r.n_buckets;
      p.regions[b] = r;
    }
  }
}

cudaError_t launch_plan(const PlanParams &p, cudaStream_t s, uint64_t *launches) {
  (*launches)++;
  plan_kernel<<<1, 256, p.local_bins * sizeof(unsigned long long), s>>>(p);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ filter
// Two-choice filter: buckets of 16 x 16-bit tags (one 32-byte sector). test-and-set semantics: returns true when the
// tag was already present in one of the key's two buckets ("seen before"), otherwise registers it and returns false.
// Tags are never removed, so for a given key the candidate slots (bucket 1 in order, then bucket 2 in order) fill
// monotonically and two racing first sightings of the same key always contend for the same slot: exactly one of them
// is told "new". False positives (another key with the same tag) only admit a singleton to the table, where it is
// counted exactly and purged; both buckets full also answers "seen".
__device__ __forceinline__ bool filter_test_and_set(uint32_t *filter, const Region &rg, uint64_t h) {
  uint32_t ftag = (uint32_t)(h & 0xffffu);
  if (ftag == 0) ftag = 1;
  const uint64_t nb = rg.n_buckets;
  uint64_t b = __umul64hi(h, nb);
#pragma unroll 1
  for (int which = 0; which < 2; which++) {
    uint32_t *wp = filter + (rg.bucket_base + b) * 8;
    uint4 lo = ld_cg16(wp), hi = ld_cg16(wp + 4);
    uint32_t wv[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
    for (int j = 0; j < 8; j++) {
      uint32_t v = wv[j];
      while (true) {
        uint32_t t0 = v & 0xffffu, t1 = v >> 16;
        if (t0 == ftag || t1 == ftag) return true;
        if (t0 != 0 && t1 != 0) break;  // word full, next word
        uint32_t nv = (t0 == 0) ? (v | ftag) : (v | (ftag << 16));
        uint32_t old = atomicCAS(wp + j, v, nv);
        if (old == v) return false;  // registered: first sighting
        v = old;                     // somebody changed the word: look again
      }
    }
    // bucket full: second choice (xor-style displacement by a multiple of the tag, as the TCF does)
    uint64_t step = 1 + (uint64_t)(ftag * 0x5bd1e995u) % (nb - 1);
    b = b + step;
    if (b >= nb) b -= nb;
  }
  return true;
}

// ------------------------------------------------------------------------------------------------ table
// Finds the slot of `key` inside the bin's region, claiming an empty one when CLAIM. Keys are stored complemented so
// that an all-zero unit means "empty". N >= 3: the 16-byte unit B[slot] completes the key; it is written right after
// the claim; a reader that matched the first two words and finds it still zero gets NOT_READY and retries the whole
// occurrence on its next loop iteration (no spinning inside a warp: owner and reader may be lanes of the same warp).
template <int N, bool CLAIM>
__device__ __forceinline__ uint64_t find_or_claim(Slot *A, uint4 *B, const Region &rg, const uint64_t (&key)[N], uint64_t h,
                                                  bool &is_new, DevCounters *ctr) {
  uint64_t idx = __umul64hi(h, rg.n_slots);
  is_new = false;
  ulonglong2 mine;
  mine.x = ~key[0];
  mine.y = N >= 2 ? ~key[1] : 0ULL;
  ulonglong2 mineB;
  mineB.x = N >= 3 ? ~key[N >= 3 ? 2 : 0] : 0ULL;
  mineB.y = N >= 4 ? ~key[N >= 4 ? 3 : 0] : 0ULL;
#pragma unroll 1
  for (uint32_t probe = 0; probe < MAX_PROBES && probe < rg.n_slots; probe++) {
    Slot *s = A + rg.slot_base + idx;
    ulonglong2 a = ld_cg128(s);
    bool match = false;
    if (N == 1) {
      if (a.x == mine.x) return rg.slot_base + idx;
      if (a.x == 0) {
        if (!CLAIM) return NOT_FOUND;
        unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long *>(&s->k0x), 0ULL, (unsigned long long)mine.x);
        if (old == 0ULL) {
          is_new = true;
          return rg.slot_base + idx;
        }
        if (old == mine.x) return rg.slot_base + idx;
      }
    } else {
      if (a.x == 0 && a.y == 0) {
        if (!CLAIM) return NOT_FOUND;
        ulonglong2 zero;
        zero.x = zero.y = 0;
        ulonglong2 old = cas128(s, zero, mine);
        if (old.x == 0 && old.y == 0) {
          if (N >= 3) (void)cas128(B + rg.slot_base + idx, zero, mineB);  // completes the key; only the claimant writes it
          is_new = true;
          return rg.slot_base + idx;
        }
        a = old;
      }
      match = (a.x == mine.x && a.y == mine.y);
      if (match) {
        if (N == 2) return rg.slot_base + idx;
        ulonglong2 b = ld_cg128(B + rg.slot_base + idx);
        if (b.x == 0 && b.y == 0) return NOT_READY;  // claimant is about to write it
        if (b.x == mineB.x && b.y == mineB.y) return rg.slot_base + idx;
      }
    }
    idx++;
    if (idx == rg.n_slots) idx = 0;
  }
  ctr->err_table_full = 1u;
  return NOT_FOUND;
}

template <int N>
__device__ __forceinline__ uint32_t *count_ptr(Slot *A, uint32_t *C, uint64_t slot) {
  return N == 1 ? reinterpret_cast<uint32_t *>(&A[slot].w1) : C + slot;
}

// count += cnt, left[l] += cnt, right[r] += cnt. The packed 16-bit fields are exact while the key has at most 65535
// occurrences; beyond that finalize discards them and uses the recount (MODE_HEAVY) instead.
template <int N>
__device__ __forceinline__ void add_counts(Slot *A, uint32_t *C, uint64_t slot, uint32_t cnt, int l, int r) {
  red_add32(count_ptr<N>(A, C, slot), cnt);
  if (l >= 0) red_add64(&A[slot].left, (uint64_t)cnt << (16 * l));
  if (r >= 0) red_add64(&A[slot].right, (uint64_t)cnt << (16 * r));
}

// ------------------------------------------------------------------------------------------------ insert kernel
template <int N>
struct Occurrence {
  uint64_t key[N];
  int l, r;  // extension base codes 0..3, -1 = not counted (low quality / N)
};

// returns false when the occurrence has to be retried (NOT_READY); nothing has been counted then
template <int N, int MODE>
__device__ __forceinline__ bool apply_occurrence(const InsertParams &p, const Region &rg, const Occurrence<N> &o, uint32_t cnt,
                                                 uint32_t &n_unique, uint32_t &n_filter_new) {
  const uint64_t h = key_hash<N>(o.key);
  bool is_new = false;
  if (MODE == MODE_INSERT) {
    uint64_t slot = find_or_claim<N, true>(p.A, p.B, rg, o.key, h, is_new, p.ctr);
    if (N >= 3 && slot == NOT_READY) return false;
    if (slot != NOT_FOUND) add_counts<N>(p.A, p.C, slot, cnt, o.l, o.r);
  } else if (MODE == MODE_ADMIT) {
    // (a retried occurrence finds its own tag and is treated as seen: it only admits a key early, which is harmless)
    bool seen = filter_test_and_set(p.filter, rg, h);
    if (!seen) n_filter_new++;
    if (seen || cnt >= 2u) {
      uint64_t slot = find_or_claim<N, true>(p.A, p.B, rg, o.key, h, is_new, p.ctr);
      if (N >= 3 && slot == NOT_READY) return false;
    }
  } else if (MODE == MODE_UPDATE) {
    uint64_t slot = find_or_claim<N, false>(p.A, p.B, rg, o.key, h, is_new, p.ctr);
    if (slot < NOT_READY) add_counts<N>(p.A, p.C, slot, cnt, o.l, o.r);
  } else {  // MODE_HEAVY: exact 32-bit extension counters of the keys finalize marked as heavy
    uint64_t slot = find_or_claim<N, false>(p.A, p.B, rg, o.key, h, is_new, p.ctr);
    if (slot < NOT_READY) {
      uint32_t c = *count_ptr<N>(p.A, p.C, slot);
      if (c & kHeavyMark) {
        uint32_t *e = p.heavy.side + (uint64_t)(c & ~kHeavyMark) * 8;
        if (o.l >= 0) atomicAdd(e + o.l, cnt);
        if (o.r >= 0) atomicAdd(e + 4 + o.r, cnt);
      }
    }
  }
  if (is_new) n_unique++;
  return true;
}

template <int N, int MODE, bool MERGE>
__global__ void __launch_bounds__(THREADS, 4) insert_kernel(InsertParams p) {
  extern __shared__ __align__(16) uint32_t s_all[];
  const uint32_t lane = lane_id();
  const int warp = threadIdx.x >> 5;
  const int k = p.k;
  const int W = p.rec_bytes / 4;
  const int T = p.tile_records;
  const int nbw = record_geom(p.rec_bytes, k).nbw;
  uint32_t *s_rec = s_all + warp * (T * W + T + 4);  // this warp's records, then T + 1 prefix sums
  uint32_t *s_pre = s_rec + T * W;
  const uint32_t n_tiles = p.seg_tile_start[p.n_segments];
  const int tail = 2 * (32 * N - k);                 // unused low bits of the last key word
  const uint64_t last_mask = ~0ULL << tail;
  const int rpl = T / 32;                            // records per lane in the prefix step

  uint32_t n_unique = 0, n_filter_new = 0, n_records = 0;
  unsigned long long n_occ = 0;

  while (true) {
    // ---- next tile of this warp
    unsigned long long tile = 0;
    int seg = 0;
    if (lane == 0) {
      tile = atomicAdd(&p.ctr->tile_cursor, 1ULL);
      if (tile < n_tiles) {
        int lo = 0, hi = p.n_segments;  // last segment with seg_tile_start <= tile
        while (hi - lo > 1) {
          int mid = (lo + hi) >> 1;
          if (p.seg_tile_start[mid] <= (uint32_t)tile) lo = mid; else hi = mid;
        }
        seg = lo;
      }
    }
    tile = __shfl_sync(0xffffffffu, tile, 0);
    seg = __shfl_sync(0xffffffffu, seg, 0);
    if (tile >= n_tiles) break;
    const Segment sg = p.segments[seg];
    const Region rg = p.regions[sg.bin];
    const uint32_t first = ((uint32_t)tile - p.seg_tile_start[seg]) * T;
    const int nrec = min((uint32_t)T, sg.n_records - first);
    // ---- stage the records (coalesced 16-byte streaming loads)
    {
      const uint4 *src = reinterpret_cast<const uint4 *>(sg.ptr + (uint64_t)first * p.rec_bytes);
      uint4 *dst = reinterpret_cast<uint4 *>(s_rec);
      const int n16 = nrec * (p.rec_bytes / 16);
      for (int i = lane; i < n16; i += 32) dst[i] = ld_stream16(src + i);
    }
    __syncwarp();
    // ---- exclusive prefix sum of k-mers per record: lane handles records [lane * rpl, lane * rpl + rpl)
    uint32_t mine = 0;
    for (int u = 0; u < rpl; u++) {
      int rr = lane * rpl + u;
      uint32_t n = rr < nrec ? (s_rec[rr * W + W - 1] & 0xffu) : 0u;
      if (n) n_records++;
      mine += n;
    }
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += v;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    {
      uint32_t at = incl - mine;
      for (int u = 0; u < rpl; u++) {
        int rr = lane * rpl + u;
        s_pre[rr] = at;
        at += rr < nrec ? (s_rec[rr * W + W - 1] & 0xffu) : 0u;
      }
      if (lane == 31) s_pre[T] = total;
    }
    __syncwarp();
    if (lane == 0) n_occ += total;

    // ---- this lane's contiguous run of occurrences [q, qend)
    const uint32_t per = (total + 31u) >> 5;
    uint32_t q = min(total, lane * per);
    const uint32_t qend = min(total, q + per);
    int rec_i = 0, j = 0, n = 0;
    const uint32_t *rec = s_rec;
    uint64_t fw[N], rc[N];
    bool fresh = true;
    if (q < qend) {
      int lo = 0, hi = T;  // last record with s_pre[rec] <= q (skips empty records)
      while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (s_pre[mid] <= q) lo = mid; else hi = mid;
      }
      rec_i = lo;
      j = q - s_pre[lo];
      rec = s_rec + rec_i * W;
      n = rec[W - 1] & 0xffu;
    }
    Occurrence<N> pend;
    uint32_t pend_cnt = 0;
    bool done = !(q < qend);
    while (true) {
      // warp-uniform exit; the vote also brings the lanes back together after the divergent probe loops of the
      // previous step, so that the table loads / REDs below are issued for (almost) all 32 lanes at once
      if (__all_sync(0xffffffffu, done)) break;
      if (!done) {
        const bool have = q < qend;
        Occurrence<N> cur;
        if (have) {
          if (fresh) {  // k-mer j of the record: bases [j+1, j+k], extensions at j and j+k+1
#pragma unroll
            for (int i = 0; i < N; i++) fw[i] = rec_window64(rec, j + 1 + 32 * i);
            fw[N - 1] &= last_mask;
            revcomp_key<N>(fw, k, rc);
            fresh = false;
          }
          const uint32_t lb = rec_base(rec, j), lq = rec_hq(rec, nbw, j);
          const uint32_t rb = rec_base(rec, j + k + 1), rq = rec_hq(rec, nbw, j + k + 1);
          if (key_less<N>(rc, fw)) {  // kcount_cpu.cpp:328-333: swap and complement the extensions
#pragma unroll
            for (int i = 0; i < N; i++) cur.key[i] = rc[i];
            cur.l = rq ? (int)(3u - rb) : -1;
            cur.r = lq ? (int)(3u - lb) : -1;
          } else {
#pragma unroll
            for (int i = 0; i < N; i++) cur.key[i] = fw[i];
            cur.l = lq ? (int)lb : -1;
            cur.r = rq ? (int)rb : -1;
          }
        }
        bool same = false;
        if (MERGE && have && pend_cnt) {
          same = cur.l == pend.l && cur.r == pend.r;
#pragma unroll
          for (int i = 0; i < N; i++) same = same && cur.key[i] == pend.key[i];
        }
        bool advance = false;
        if (same) {
          pend_cnt++;
          advance = true;
        } else {
          // a false return = the slot's owner has not completed the key yet: same state again next iteration
          const bool ok = pend_cnt ? apply_occurrence<N, MODE>(p, rg, pend, pend_cnt, n_unique, n_filter_new) : true;
          if (ok) {
            if (!have) {
              done = true;
            } else {
              pend = cur;
              pend_cnt = 1;
              advance = true;
            }
          }
        }
        if (advance) {
          // ---- next occurrence of the run
          q++;
          j++;
          if (q < qend) {
            if (j == n) {  // next non-empty record
              do {
                rec_i++;
                rec += W;
                n = rec[W - 1] & 0xffu;
              } while (n == 0);
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
      }
    }
    __syncwarp();  // the tile buffer is reused
  }

  // ---- statistics: warp reduce, one atomic per warp and counter
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_unique += __shfl_down_sync(0xffffffffu, n_unique, o);
    n_filter_new += __shfl_down_sync(0xffffffffu, n_filter_new, o);
    n_records += __shfl_down_sync(0xffffffffu, n_records, o);
  }
  if (lane == 0) {
    if (n_unique) atomicAdd(&p.ctr->unique, (unsigned long long)n_unique);
    if (n_filter_new) atomicAdd(&p.ctr->filter_new, (unsigned long long)n_filter_new);
    if (n_records && (MODE == MODE_INSERT || MODE == MODE_ADMIT)) atomicAdd(&p.ctr->records_in, (unsigned long long)n_records);
    if (n_occ && (MODE == MODE_INSERT || MODE == MODE_ADMIT)) atomicAdd(&p.ctr->occurrences, n_occ);
  }
}

template <int N, int MODE>
static cudaError_t launch_insert_nm(const InsertParams &p, bool merge, int grid, size_t smem, cudaStream_t s) {
  if (merge) {
    cudaFuncSetAttribute(insert_kernel<N, MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    insert_kernel<N, MODE, true><<<grid, THREADS, smem, s>>>(p);
  } else {
    cudaFuncSetAttribute(insert_kernel<N, MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    insert_kernel<N, MODE, false><<<grid, THREADS, smem, s>>>(p);
  }
  return cudaGetLastError();
}

template <int N>
static cudaError_t launch_insert_n(const InsertParams &p, int mode, bool merge, int grid, size_t smem, cudaStream_t s) {
  switch (mode) {
    case MODE_INSERT: return launch_insert_nm<N, MODE_INSERT>(p, merge, grid, smem, s);
    case MODE_ADMIT: return launch_insert_nm<N, MODE_ADMIT>(p, merge, grid, smem, s);
    case MODE_UPDATE: return launch_insert_nm<N, MODE_UPDATE>(p, merge, grid, smem, s);
    case MODE_HEAVY: return launch_insert_nm<N, MODE_HEAVY>(p, merge, grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_insert(const InsertParams &p, int n_words, int mode, bool merge, int sm_count, cudaStream_t s, uint64_t *launches) {
  const size_t per_warp = ((size_t)p.tile_records * p.rec_bytes / 4 + p.tile_records + 4) * sizeof(uint32_t);
  const size_t smem = per_warp * WARPS_PER_CTA;
  // persistent warps: as many CTAs as stay resident (register bound), tiles are pulled dynamically
  const int grid = sm_count * 4;
  (*launches)++;
  switch (n_words) {
    case 1: return launch_insert_n<1>(p, mode, merge, grid, smem, s);
    case 2: return launch_insert_n<2>(p, mode, merge, grid, smem, s);
    case 3: return launch_insert_n<3>(p, mode, merge, grid, smem, s);
    case 4: return launch_insert_n<4>(p, mode, merge, grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace kc
