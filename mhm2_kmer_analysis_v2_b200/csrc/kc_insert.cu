// kc_insert.cu -- kernels (3) "k-mer hash table insert with extension counts" and (4) "two-choice singleton filter",
// plus the small planning kernel that turns received supermer regions into tiles and table regions.
//
// Replaces HashTableInserter::insert_supermer -> insert_supermer_from_read -> get_kmers_and_exts + KmerMapExts::insert
// (src/kcount/kcount_cpu.cpp:478-493, 339-355, 309-336, 232-271) and the reference's GPU kernel gpu_insert_supermer_block
// with its TCF calls (src/kcount/kcount-gpu/gpu_hash_table.cpp:310-475, tcf_wrapper.hpp:96-178).
//
// A CTA pulls tiles of 256 supermer records (bin-major order, so the table region being hit stays L2 resident), stages
// them in shared memory with 128-bit streaming loads, prefix-sums the k-mer counts and then works on a dense list of
// k-mer occurrences: thread -> (record, k-mer) by binary search in shared memory. Per occurrence: extract the N key
// words with funnel shifts, reverse-complement arithmetically, canonicalise and swap/complement the two extension bases
// (kcount_cpu.cpp:326-333), merge equal (k-mer, extensions) inside the warp with __match_any_sync, probe the region
// (linear probing, one 16-byte load per probe), claim with a single CAS (64-bit key for N == 1, 32-bit tag otherwise),
// and add: one returning atomicAdd on the count, one 64-bit RED per extension word.
#include "kc_internal.h"

namespace kc {

static const int THREADS = 256;
static const uint32_t MAX_PROBES = 1u << 16;
static const uint64_t NOT_FOUND = ~0ULL;

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
template <int N, bool CLAIM>
__device__ __forceinline__ uint64_t find_or_claim(SlotA *A, uint64_t *B, const Region &rg, const uint64_t (&key)[N], uint64_t h,
                                                  bool &is_new, DevCounters *ctr) {
  uint64_t idx = __umul64hi(h, rg.n_slots);
  const uint32_t tag = (uint32_t)h | 2u;
  is_new = false;
#pragma unroll 1
  for (uint32_t probe = 0; probe < MAX_PROBES && probe < rg.n_slots; probe++) {
    SlotA *s = A + rg.slot_base + idx;
    uint4 a = ld_cg16(s);
    uint64_t k0 = ((uint64_t)a.y << 32) | a.x;
    if (N == 1) {
      const uint64_t mine = ~key[0];
      if (k0 == mine) return rg.slot_base + idx;
      if (k0 == 0) {
        if (!CLAIM) return NOT_FOUND;
        unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long *>(&s->key0x), 0ULL, (unsigned long long)mine);
        if (old == 0ULL) {
          is_new = true;
          return rg.slot_base + idx;
        }
        if (old == mine) return rg.slot_base + idx;
      }
    } else {
      uint32_t t = a.w;
      uint64_t *kb = B + (rg.slot_base + idx) * (N - 1);
      if (t == 0) {
        if (!CLAIM) return NOT_FOUND;
        uint32_t old = atomicCAS(&s->tag, 0u, 1u);
        if (old == 0u) {
          // slot is ours: write the key, then publish the tag (release orders the key stores before it)
          *reinterpret_cast<volatile uint64_t *>(&s->key0x) = key[0];
#pragma unroll
          for (int j = 1; j < N; j++) *reinterpret_cast<volatile uint64_t *>(&kb[j - 1]) = key[j];
          __threadfence();
          st_release32(&s->tag, tag);
          is_new = true;
          return rg.slot_base + idx;
        }
        t = old;
        k0 = ~key[0];  // force a re-read below
      }
      bool waited = false;
      while (t == 1u) {  // another thread is writing this slot's key
        t = ld_acquire32(&s->tag);
        waited = true;
      }
      if (t == tag) {
        if (waited || k0 != key[0]) k0 = ld_cg64(&s->key0x);
        if (k0 == key[0]) {
          bool same = true;
#pragma unroll
          for (int j = 1; j < N; j++) same = same && (ld_cg64(&kb[j - 1]) == key[j]);
          if (same) return rg.slot_base + idx;
        }
      }
    }
    idx++;
    if (idx == rg.n_slots) idx = 0;
  }
  ctr->err_table_full = 1u;
  return NOT_FOUND;
}

__device__ __forceinline__ void overflow_add(const Overflow &ovf, uint64_t slot, int l, int r, uint32_t v, DevCounters *ctr) {
  uint64_t idx = (slot * 0x9e3779b97f4a7c15ULL >> 20) & ovf.mask;
  for (uint64_t probe = 0; probe <= ovf.mask; probe++) {
    unsigned long long cur = atomicCAS(reinterpret_cast<unsigned long long *>(&ovf.keys[idx]), 0ULL, (unsigned long long)(slot + 1));
    if (cur == 0ULL || cur == slot + 1) {
      uint32_t *e = ovf.ext + idx * 8;
      if (l >= 0) {
        uint32_t o = atomicAdd(&e[l], v);
        if (o >= 0x80000000u) atomicSub(&e[l], v);
      }
      if (r >= 0) {
        uint32_t o = atomicAdd(&e[4 + r], v);
        if (o >= 0x80000000u) atomicSub(&e[4 + r], v);
      }
      return;
    }
    idx = (idx + 1) & ovf.mask;
  }
  ctr->err_ovf_full = 1u;
}

// count += cnt (u16 saturating), left[l] += cnt, right[r] += cnt with exact per-field saturation at 65535
// (kcount_cpu.cpp:147-164, 349-353). Only the first 65535 occurrences of a key (in atomic order) touch the packed
// 16-bit fields, so a field never carries into its neighbour; later ones go to the overflow table.
__device__ __forceinline__ void add_counts(SlotA *s, uint64_t slot, uint32_t cnt, int l, int r, const Overflow &ovf,
                                           DevCounters *ctr, uint32_t &ge2) {
  uint32_t old = atomicAdd(&s->count, cnt);
  if (old < 2u && old + cnt >= 2u) ge2++;
  uint32_t fast = old >= 65535u ? 0u : min(cnt, 65535u - old);
  uint32_t slow = cnt - fast;
  if (fast) {
    if (l >= 0) red_add64(&s->left, (uint64_t)fast << (16 * l));
    if (r >= 0) red_add64(&s->right, (uint64_t)fast << (16 * r));
  }
  if (slow) {
    atomicSub(&s->count, slow);  // keeps count within [65535, 65535 + occurrences in flight]: no 32-bit wrap
    if (l >= 0 || r >= 0) overflow_add(ovf, slot, l, r, slow, ctr);
  }
}

// ------------------------------------------------------------------------------------------------ insert kernel
template <int N, int MODE, bool DEDUP>
__global__ void __launch_bounds__(THREADS) insert_kernel(InsertParams p) {
  extern __shared__ __align__(16) uint32_t s_rec[];   // kTileRecords * W words, then kTileRecords + 1 prefix
  __shared__ uint32_t s_warp_tot[THREADS / 32];
  __shared__ unsigned long long s_tile;
  __shared__ int s_seg;
  const int tid = threadIdx.x;
  const uint32_t lane = lane_id();
  const int k = p.k;
  const int W = p.rec_bytes / 4;
  const int nbw = record_geom(p.rec_bytes, k).nbw;
  uint32_t *s_pre = s_rec + kTileRecords * W;
  const uint32_t n_tiles = p.seg_tile_start[p.n_segments];
  const uint64_t last_mask = ~0ULL << (2 * (32 * N - k));  // valid bits of the last key word

  uint32_t n_unique = 0, n_ge2 = 0, n_filter_new = 0, n_records = 0;
  unsigned long long n_occ = 0;

  while (true) {
    if (tid == 0) {
      unsigned long long t = atomicAdd(&p.ctr->tile_cursor, 1ULL);
      s_tile = t;
      if (t < n_tiles) {
        // segment of this tile: last i with seg_tile_start[i] <= t
        int lo = 0, hi = p.n_segments;
        while (hi - lo > 1) {
          int mid = (lo + hi) >> 1;
          if (p.seg_tile_start[mid] <= (uint32_t)t) lo = mid; else hi = mid;
        }
        s_seg = lo;
      }
    }
    __syncthreads();
    const unsigned long long tile = s_tile;
    if (tile >= n_tiles) break;
    const Segment sg = p.segments[s_seg];
    const Region rg = p.regions[sg.bin];
    const uint32_t first = ((uint32_t)tile - p.seg_tile_start[s_seg]) * kTileRecords;
    const int nrec = min((uint32_t)kTileRecords, sg.n_records - first);
    // ---- stage the records (coalesced 16-byte streaming loads)
    {
      const uint4 *src = reinterpret_cast<const uint4 *>(sg.ptr + (uint64_t)first * p.rec_bytes);
      uint4 *dst = reinterpret_cast<uint4 *>(s_rec);
      const int n16 = nrec * (p.rec_bytes / 16);
      for (int i = tid; i < n16; i += THREADS) dst[i] = ld_stream16(src + i);
    }
    __syncthreads();
    // ---- exclusive prefix sum of k-mers per record
    uint32_t n_mine = (tid < nrec) ? (s_rec[tid * W + W - 1] & 0xffffu) : 0u;
    if (n_mine) n_records++;
    uint32_t incl = n_mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += v;
    }
    if (lane == 31) s_warp_tot[tid >> 5] = incl;
    __syncthreads();
    uint32_t warp_off = 0, total = 0;
#pragma unroll
    for (int wv = 0; wv < THREADS / 32; wv++) {
      uint32_t t = s_warp_tot[wv];
      if (wv < (tid >> 5)) warp_off += t;
      total += t;
    }
    s_pre[tid] = warp_off + incl - n_mine;
    if (tid == 0) s_pre[kTileRecords] = total;
    __syncthreads();
    n_occ += (tid == 0) ? total : 0;

    // ---- one k-mer occurrence per thread and step
    for (uint32_t q0 = 0; q0 < total; q0 += THREADS) {
      const uint32_t q = q0 + tid;
      const bool valid = q < total;
      uint64_t key[N];
      uint64_t h = 0;
      int l = -1, r = -1;
      if (valid) {
        int lo = 0, hi = kTileRecords;  // last record with s_pre[rec] <= q
        while (hi - lo > 1) {
          int mid = (lo + hi) >> 1;
          if (s_pre[mid] <= q) lo = mid; else hi = mid;
        }
        const uint32_t *rec = s_rec + lo * W;
        const int j = q - s_pre[lo];  // k-mer j of the record: bases [j+1, j+k], extensions at j and j+k+1
        uint64_t fw[N], rc[N];
#pragma unroll
        for (int i = 0; i < N; i++) fw[i] = rec_window64(rec, j + 1 + 32 * i);
        fw[N - 1] &= last_mask;
        revcomp_key<N>(fw, k, rc);
        uint32_t lb = rec_base(rec, j), lq = rec_hq(rec, nbw, j);
        uint32_t rb = rec_base(rec, j + k + 1), rq = rec_hq(rec, nbw, j + k + 1);
        if (key_less<N>(rc, fw)) {  // kcount_cpu.cpp:328-333: swap and complement the extensions
#pragma unroll
          for (int i = 0; i < N; i++) key[i] = rc[i];
          l = rq ? (int)(3u - rb) : -1;
          r = lq ? (int)(3u - lb) : -1;
        } else {
#pragma unroll
          for (int i = 0; i < N; i++) key[i] = fw[i];
          l = lq ? (int)lb : -1;
          r = rq ? (int)rb : -1;
        }
        h = key_hash<N>(key);
      }
      uint32_t cnt = valid ? 1u : 0u;
      if (DEDUP) {
        // merge identical (k-mer, extensions) inside the warp: the first lane of each group works for all
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        if (valid) {
          const uint64_t ident = h ^ ((uint64_t)(l + 1) * 0x9e3779b97f4a7c15ULL) ^ ((uint64_t)(r + 1) * 0xc2b2ae3d27d4eb4fULL);
          const uint32_t peers = __match_any_sync(vmask, ident);
          const int leader = __ffs(peers) - 1;
          bool same = true;  // a 64-bit match is only a hint: compare the real key with the leader's
#pragma unroll
          for (int i = 0; i < N; i++) same = same && (__shfl_sync(peers, key[i], leader) == key[i]);
          same = same && (__shfl_sync(peers, l * 8 + r, leader) == l * 8 + r);
          const uint32_t diff = __ballot_sync(peers, !same);
          if (diff == 0u)
            cnt = ((int)lane == leader) ? __popc(peers) : 0u;
          // else: hash collision inside the group -- every lane inserts its own occurrence
        }
      }
      if (cnt) {
        bool is_new = false;
        uint64_t slot = NOT_FOUND;
        if (MODE == MODE_INSERT) {
          slot = find_or_claim<N, true>(p.A, p.B, rg, key, h, is_new, p.ctr);
        } else if (MODE == MODE_ADMIT) {
          bool seen = filter_test_and_set(p.filter, rg, h);
          if (!seen) n_filter_new++;
          if (seen || cnt >= 2u) (void)find_or_claim<N, true>(p.A, p.B, rg, key, h, is_new, p.ctr);
        } else {
          slot = find_or_claim<N, false>(p.A, p.B, rg, key, h, is_new, p.ctr);
        }
        if (is_new) n_unique++;
        if (MODE != MODE_ADMIT && slot != NOT_FOUND) add_counts(p.A + slot, slot, cnt, l, r, p.ovf, p.ctr, n_ge2);
      }
    }
    __syncthreads();  // tile buffer is reused
  }

  // ---- statistics: warp reduce, one atomic per warp and counter
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_unique += __shfl_down_sync(0xffffffffu, n_unique, o);
    n_ge2 += __shfl_down_sync(0xffffffffu, n_ge2, o);
    n_filter_new += __shfl_down_sync(0xffffffffu, n_filter_new, o);
    n_records += __shfl_down_sync(0xffffffffu, n_records, o);
  }
  if (lane == 0) {
    if (n_unique) atomicAdd(&p.ctr->unique, (unsigned long long)n_unique);
    if (n_ge2) atomicAdd(&p.ctr->ge2, (unsigned long long)n_ge2);
    if (n_filter_new) atomicAdd(&p.ctr->filter_new, (unsigned long long)n_filter_new);
    if (n_records && MODE != MODE_UPDATE) atomicAdd(&p.ctr->records_in, (unsigned long long)n_records);
  }
  if (tid == 0 && n_occ && MODE != MODE_UPDATE) atomicAdd(&p.ctr->occurrences, n_occ);
}

template <int N, int MODE>
static cudaError_t launch_insert_nm(const InsertParams &p, bool dedup, int grid, size_t smem, cudaStream_t s) {
  if (dedup) {
    cudaFuncSetAttribute(insert_kernel<N, MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    insert_kernel<N, MODE, true><<<grid, THREADS, smem, s>>>(p);
  } else {
    cudaFuncSetAttribute(insert_kernel<N, MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    insert_kernel<N, MODE, false><<<grid, THREADS, smem, s>>>(p);
  }
  return cudaGetLastError();
}

template <int N>
static cudaError_t launch_insert_n(const InsertParams &p, int mode, bool dedup, int grid, size_t smem, cudaStream_t s) {
  switch (mode) {
    case MODE_INSERT: return launch_insert_nm<N, MODE_INSERT>(p, dedup, grid, smem, s);
    case MODE_ADMIT: return launch_insert_nm<N, MODE_ADMIT>(p, dedup, grid, smem, s);
    case MODE_UPDATE: return launch_insert_nm<N, MODE_UPDATE>(p, dedup, grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_insert(const InsertParams &p, int n_words, int mode, bool dedup, int sm_count, cudaStream_t s, uint64_t *launches) {
  const size_t smem = (size_t)kTileRecords * p.rec_bytes + (kTileRecords + 1) * sizeof(uint32_t);
  // persistent CTAs: as many as stay resident (shared memory bound for 64-byte records), tiles are pulled dynamically
  int per_sm = (int)((200 * 1024) / (smem + 1024));
  if (per_sm > 8) per_sm = 8;
  if (per_sm < 1) per_sm = 1;
  const int grid = sm_count * per_sm;
  (*launches)++;
  switch (n_words) {
    case 1: return launch_insert_n<1>(p, mode, dedup, grid, smem, s);
    case 2: return launch_insert_n<2>(p, mode, dedup, grid, smem, s);
    case 3: return launch_insert_n<3>(p, mode, dedup, grid, smem, s);
    case 4: return launch_insert_n<4>(p, mode, dedup, grid, smem, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace kc
