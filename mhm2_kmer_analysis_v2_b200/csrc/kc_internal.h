// kc_internal.h -- structures shared between the kernels (kc_parse.cu, kc_insert.cu, kc_finalize.cu) and the
// C ABI / context code (kc_api.cu). Not installed; the public surface is include/kcount_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kc_device.cuh"

struct kc_ctx;
namespace kc {
int ctx_device(const kc_ctx *c);  // the CUDA device of a context (kc_api.cu)

// ---- hash table ---------------------------------------------------------------
// One 32-byte, sector-aligned block per slot holds what an insert touches:
//   N == 1 : k0x = ~key word 0 (0 = empty), w1 = 32-bit occurrence count (low half)
//   N >= 2 : k0x = ~key word 0, w1 = ~key word 1 (both 0 = empty; a canonical key never complements to 0 because the
//            unused trailing bits of its last word are 0), claimed with ONE 128-bit CAS; the 32-bit occurrence count
//            lives in the dense side array C[slot]; key words 2..3 live in B[slot] (16 bytes, complemented, 0 = not yet
//            written -- readers that matched k0x/w1 wait for it)
// left/right pack four 16-bit counters (A,C,G,T at bits 0,16,32,48). Every update is a fire-and-forget RED: no
// returning atomic, no fence. A key with more than 65535 occurrences ("heavy", e.g. a homopolymer) may carry between
// its packed fields; finalize detects it from the 32-bit count and recounts its extensions exactly (Heavy below).
struct alignas(32) Slot {
  uint64_t k0x;
  uint64_t w1;
  uint64_t left;
  uint64_t right;
};
static_assert(sizeof(Slot) == 32, "slot block must be one 32-byte sector");

static const uint32_t kHeavyMark = 0x80000000u;  // count field of a heavy slot after detection: kHeavyMark | heavy index

struct Heavy {
  uint64_t *slots;            // [cap] slot index of each heavy key
  uint32_t *side;             // [cap * 8] exact extension counters lA lC lG lT rA rC rG rT, filled by the recount pass
  unsigned long long *n;      // number of heavy keys found by finalize (may exceed cap: then it is an error)
  uint64_t cap;
};

struct Region {  // table / filter range owned by one local bin
  uint64_t slot_base, n_slots;
  uint64_t bucket_base, n_buckets;
};

struct Segment {  // run of supermer records of one (source rank, local bin)
  const uint8_t *ptr;
  uint32_t n_records;
  uint32_t bin;
};

enum InsertMode { MODE_INSERT = 0, MODE_ADMIT = 1, MODE_UPDATE = 2, MODE_HEAVY = 3 };

struct DevCounters {  // device-resident, zeroed by kc_reset
  unsigned long long tile_cursor;
  unsigned long long unique;        // keys that claimed a table slot
  unsigned long long ge2;           // keys with count >= 2 (counted by finalize)
  unsigned long long occurrences;   // k-mer occurrences inserted
  unsigned long long records_in;    // non-empty supermer records consumed
  unsigned long long filter_new;    // first sightings registered by the filter
  unsigned long long n_final, n_ext;  // finalize output cursors
  unsigned long long sum_counts, max_count;
  unsigned long long purged;
  unsigned int err_table_full, err_heavy_full;
  unsigned int err_result_full;     // bin-resident counting: the dense result arrays were too small (n_final / n_ext hold the need)
  unsigned int err_count_stack;     // bin-resident counting: a bin could not be split far enough
  unsigned int err_index;           // lookup index: an insert did not terminate
  unsigned int err_wait;            // SM-partitioned pipeline: a count warp gave up waiting for the refine kernel
  unsigned long long split_passes;  // extra passes over bins whose distinct k-mers did not fit shared memory
  unsigned long long parse_counted, parse_records, parse_raw;
  unsigned int err_bins_full;
};

struct InsertParams {
  int k, rec_bytes, n_segments;
  const Segment *segments;
  const uint32_t *seg_tile_start;  // [n_segments + 1] exclusive prefix of tiles
  const Region *regions;
  int tile_records;  // records per warp tile
  Slot *A;
  uint4 *B;          // N >= 3: key words 2..3 (complemented)
  uint32_t *C;       // N >= 2: occurrence counts
  uint32_t *filter;  // 8 u32 words (16 tags) per bucket
  Heavy heavy;       // MODE_HEAVY only
  DevCounters *ctr;
};

struct FinalizeParams {
  int k, n_words, dmin_thres, keep_ext;
  double dyn_min_depth;
  uint64_t n_slots;
  Slot *A;
  const uint4 *B;
  uint32_t *C;
  Heavy heavy;
  DevCounters *ctr;
  // survivors (D1)
  uint64_t *out_keys;
  uint16_t *out_counts;
  char *out_left, *out_right;
  uint64_t cap_final;
  // all entries with count >= 2 (D2), only when keep_ext
  uint64_t *ext_keys;
  uint16_t *ext_counts, *ext_ext8;
  char *ext_left, *ext_right;
  uint64_t cap_ext;
};

struct ParseParams {
  const uint8_t *bytes;
  const uint64_t *offsets;
  const uint16_t *lens;
  uint64_t n_reads;
  uint32_t read_stride;  // 1 = all reads, S = every S-th read (capacity sampling pass)
  int count_only;
  int k, m, qual_cutoff, rec_bytes;
  uint32_t n_ids;        // n_ranks * local_bins
  uint32_t local_bins;   // (coarse) bins per owner: fine bin = mulhi(mix(hash / n_ranks), local_bins << fine_shift)
  uint32_t fine_shift;   // log2 of the refine factor: fine bin = coarse bin << fine_shift | sub-bin (record header bits 8..15)
  FastDiv div_ranks;
  uint8_t *bins;                     // record storage
  const uint64_t *bin_base;          // [n_ids] first record of each bin
  const uint64_t *bin_cap;           // [n_ids] capacity in records
  unsigned long long *bin_cursor;    // [n_ids] records appended (may run past cap: then nothing is written)
  DevCounters *ctr;
};

// ---- launchers (each returns the cudaError of the launch) ---------------------------------------------
cudaError_t launch_parse(const ParseParams &p, int sm_count, cudaStream_t s, uint64_t *launches);
// writes empty-record headers over [cursor, cap) of every bin so the whole bin range can be shipped/read
cudaError_t launch_seal_bins(uint8_t *bins, const uint64_t *bin_base, const uint64_t *bin_cap, const unsigned long long *bin_cursor,
                             uint32_t n_ids, int rec_bytes, cudaStream_t s, uint64_t *launches);
// per-destination header blocks + segment/region tables for the insert kernel
struct PlanParams {
  int n_src, local_bins, rec_bytes, tile_records;
  // per source: pointer to its region for this rank = [header: local_bins x {u32 first_record, u32 n_records}][records]
  const uint8_t *const *src_ptr;
  Segment *segments;         // [local_bins * n_src], bin-major
  uint32_t *seg_tile_start;  // [local_bins * n_src + 1]
  Region *regions;           // [local_bins]
  uint64_t table_slots, filter_buckets;
  unsigned long long *total_records;  // out
};
cudaError_t launch_plan(const PlanParams &p, cudaStream_t s, uint64_t *launches);
cudaError_t launch_write_headers(uint8_t *const *dst_header, const uint64_t *bin_base, const unsigned long long *bin_cursor,
                                 const uint64_t *bin_cap, int n_ranks, int local_bins, unsigned long long *owner_used, cudaStream_t s,
                                 uint64_t *launches);
// bin layout computed on the device from the cursors of a (sampled) counting pass
struct LayoutParams {
  uint32_t n_ranks, local_bins;
  double scale;                           // all reads / reads the counting pass looked at (1 = exact)
  int rec_bytes, header_bytes;
  unsigned long long pad;                 // records added to every bin
  const unsigned long long *bin_cursor;   // [n_ids] records seen by the counting pass
  const unsigned long long *sampled_total;  // records seen by the counting pass, all bins
  uint64_t *bin_cap, *bin_first, *bin_base;  // [n_ids] out
  unsigned long long *owner_records;      // [n_ranks] out: record capacity of each owner's area
  unsigned long long *region_off;         // [n_ranks + 1] out: byte offset of each owner's region, total at the end
  unsigned long long *region_bytes;       // [n_ranks] out
  unsigned long long *block_sums;         // [n_ranks * layout_blocks_per_owner(local_bins)] scratch
  unsigned int *err_too_large;
};
inline uint32_t layout_blocks_per_owner(uint32_t local_bins) { return (local_bins + 1023u) / 1024u; }
cudaError_t launch_layout(const LayoutParams &p, cudaStream_t s, uint64_t *launches);
cudaError_t launch_insert(const InsertParams &p, int n_words, int mode, bool merge_runs, int sm_count, cudaStream_t s, uint64_t *launches);
cudaError_t launch_finalize(const FinalizeParams &p, int sm_count, cudaStream_t s, uint64_t *launches);
// second finalize step for the heavy keys listed by launch_finalize, after the MODE_HEAVY recount
cudaError_t launch_finalize_heavy(const FinalizeParams &p, uint64_t n_heavy, cudaStream_t s, uint64_t *launches);
int insert_tile_records(int rec_bytes);
cudaError_t launch_checksum(const uint64_t *keys, const uint16_t *counts, const char *left, const char *right, const uint16_t *ext8,
                            uint64_t n, int n_words, unsigned long long *out, cudaStream_t s, uint64_t *launches);

// ---- bin-resident counting (kc_count.cu) ----------------------------------------
// The default path: every local bin is small enough that the distinct k-mers of the bin fit a hash table in shared
// memory; one WARP counts a bin there and emits the purged / extension-resolved entries straight to the dense result.
static const int kCountMaxSrc = 64;    // source ranks per owner the count kernel can read from
static const int kSlicesMax = 12;      // slices of the bin range (kSlices below)
static const int kCountMaxHeavy = 8;   // keys with more than 65535 occurrences per (bin, split pass)

struct CountParams {
  int k, rec_bytes, n_src, tile_records;
  uint32_t n_bins;
  uint32_t bin_begin, bin_end;       // bins this launch counts (a slice of [0, n_bins): results stream out slice by slice)
  int header_bytes;                  // bytes of the per-source header block [n_bins x {u32 first_record, u32 n_records}]
  const uint8_t *const *src_ptr;     // [n_src] region of each source rank for this owner (nullptr = nothing received)
  const uint8_t *src0;               // the single source (the refined region) when n_src == 1: no pointer table to read
  uint32_t n_slots;                  // table slots in shared memory (per warp)
  int warp_bytes;                    // shared memory of one warp (set by launch_count)
  uint32_t claim_limit;              // claimed slots at which a pass gives up and the bin is split further
  int presplit_adapt;                // 1: every warp corrects the threshold from the bins it sees (kc_count.cu)
  uint32_t presplit_rec;             // > 0: a bin with more than s x this many records starts out split in 2 s parts (s = 1, 2, 4)
  int dmin_thres, keep_ext;
  double one_minus_dyn;
  DevCounters *ctr;
  uint64_t *out_keys;
  uint16_t *out_counts;
  char *out_left, *out_right;
  uint64_t cap_final;
  uint64_t *ext_keys;
  uint16_t *ext_counts, *ext_ext8;
  char *ext_left, *ext_right;
  uint64_t cap_ext;
  // SM-partitioned pipeline (kc_api.cu): the refine kernel runs BESIDE this launch on its own SMs and announces every
  // finished coarse bin in slice_done[slice]; a warp that reaches a fine bin of slice s first waits until
  // slice_done[s] == slice_need[s]. wait_slices = 0: everything is in place (the usual launch).
  int wait_slices;
  const unsigned int *slice_done;
  uint32_t slice_fine_edge[kSlicesMax + 1];  // first fine bin of every slice
  uint32_t slice_need[kSlicesMax];           // coarse bins of every slice
  unsigned int *err_wait;                    // set when a wait did not end (the refine kernel never became resident)
};
// slots that fit the shared-memory budget of one CTA for N key words
uint32_t count_table_slots(int n_words, int rec_bytes, int k);
cudaError_t launch_count(const CountParams &p, int n_words, bool merge_runs, int sm_count, cudaStream_t s, uint64_t *launches);

// ---- second partition level (kc_refine.cu): coarse bins of every source -> one region of fine bins
struct RefineParams {
  int n_src, rec_bytes;
  uint32_t n_coarse, fine_shift;
  uint32_t coarse_begin, coarse_end;  // the slice of coarse bins this launch refines (launch_refine; 0, 0 = all)
  int src_header_bytes;            // per-source header block [n_coarse x {u32 first_record, u32 n_records}]
  const uint8_t *const *src_ptr;   // [n_src] region of each source rank for this owner (nullptr = nothing received)
  uint8_t *dst;                    // [fine header: (n_coarse << fine_shift) x {u32 first, u32 n}][records]
  int dst_header_bytes;
  unsigned long long *coarse_base; // [n_coarse + 1] records per coarse bin -> exclusive prefix
  unsigned long long *total;       // out: records of this owner
  unsigned int *err_too_large;
  // SM-partitioned pipeline: after a coarse bin is written, slice_done[slice of the bin] += 1 (release); n_slices = 0: off
  int n_slices;
  unsigned int *slice_done;
  uint32_t slice_coarse_edge[kSlicesMax + 1];
  int grid_limit;                  // > 0: at most this many CTAs (the SMs set aside for this kernel)
};
cudaError_t launch_refine_totals(const RefineParams &p, cudaStream_t s, uint64_t *launches);
// slice_done[kSlicesMax] counts the resident CTAs of a partitioned refine launch; the gate holds a stream until `need` are
cudaError_t launch_gate(const unsigned int *resident, unsigned int need, cudaStream_t s, uint64_t *launches);
cudaError_t launch_refine(const RefineParams &p, uint64_t avg_records, int sm_count, cudaStream_t s, uint64_t *launches);

// ---- lookup index over the finalized result (kc_lookup.cu)
struct IndexParams {
  const uint64_t *keys;  // [n * n_words] result entries
  const uint16_t *counts;
  const char *left, *right;
  uint64_t n;
  int n_words;
  unsigned long long *slots;  // [mask + 1] fingerprint << 32 | entry, all-ones = empty
  uint64_t mask;
  unsigned int *err;
};
cudaError_t launch_index_build(const IndexParams &p, int sm_count, cudaStream_t s, uint64_t *launches);
cudaError_t launch_index_lookup(const IndexParams &p, const uint64_t *queries, uint64_t n_queries, uint16_t *out_counts, char *out_left,
                                char *out_right, int sm_count, cudaStream_t s, uint64_t *launches);

// ---- result order (kc_sort.cu): entries sorted by key, word by word unsigned
cudaError_t sort_result_by_key(uint64_t *keys, uint16_t *counts, char *left, char *right, uint16_t *ext8, uint64_t n, int n_words, int sm_count,
                               cudaStream_t s, uint64_t *launches);

// offsets[i] = sum of lens[0 .. i) (device arrays)
cudaError_t scan_lens_to_offsets(const uint16_t *lens, uint64_t *offsets, uint64_t n, cudaStream_t s, uint64_t *launches);

static const int kHeaderAlign = 64;

// ---- slices of the bin range: the receive copies, the refine pass, the count pass and the result hand-off all work slice
// by slice so that they overlap (kc_api.cu). Edge i of kSlices slices over n bins; the first two slices are small (1/48 and
// 1/24 of the bins) so that the pipeline fills quickly, the rest are equal. The same function on every rank.
static const int kSlices = kSlicesMax;
__host__ __device__ inline uint32_t slice_edge(uint32_t n, int i, int n_slices) {
  if (i <= 0) return 0u;
  if (i >= n_slices) return n;
  if (n_slices < 3) return (uint32_t)((uint64_t)n * (uint64_t)i / (uint64_t)n_slices);
  const uint64_t head = (uint64_t)n / 48;
  if (i == 1) return (uint32_t)head;
  const uint64_t second = 3 * head;
  return (uint32_t)(second + ((uint64_t)n - second) * (uint64_t)(i - 2) / (uint64_t)(n_slices - 2));
}
// byte offset (from the start of an owner's region) of every slice edge, for every owner: out[o * (n_slices + 1) + i]
cudaError_t launch_slice_offsets(const uint64_t *bin_first, const uint64_t *bin_cap, uint32_t n_ranks, uint32_t local_bins, int n_slices, int header_bytes,
                                 int rec_bytes, unsigned long long *out, cudaStream_t s, uint64_t *launches);

}  // namespace kc
