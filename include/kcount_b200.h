/*
 * kcount_b200.h -- C ABI of the B200-native k-mer analysis ("kcount") stage.
 *
 * This is the drop-in boundary. Plain pointers and sizes only: no CUDA, torch
 * or C++ types. Every call returns KC_OK (0) or a negative kc_status; the text of the last error of a
 * context is available from kc_last_error(). Nothing in this library falls back to the CPU: without a
 * CUDA device every compute entry point fails with KC_ERR_CUDA.
 *
 * Reference interfaces each group replaces (paths relative to the reference tree):
 *   kc_create/kc_destroy     <- KmerDHT<MAX_K>::KmerDHT / ~KmerDHT                 src/kcount/kmer_dht.cpp:104-166
 *                               + HashTableInserter<MAX_K>::init                    src/kcount/kcount_gpu.cpp:225-267
 *                               + HashTableGPUDriver<MAX_K>::init                   src/kcount/kcount-gpu/gpu_hash_table.cpp:528-650
 *   kc_add_reads             <- the read loop of count_kmers()                      src/kcount/kcount.cpp:71-90
 *                               (input = PackedRead bytes, src/packed_reads.cpp:99-126)
 *   kc_parse                 <- SeqBlockInserter<MAX_K>::process_seq/process_block  src/kcount/kcount_gpu.cpp:111-180
 *                               + ParseAndPackGPUDriver::process_seq_block          src/kcount/kcount-gpu/parse_and_pack.cpp:281-338
 *   kc_exchange_*            <- KmerDHT::add_supermer / flush_updates (UPC++ RPCs)  src/kcount/kmer_dht.cpp:248-258
 *   kc_insert                <- HashTableInserter<MAX_K>::insert_supermer           src/kcount/kcount_gpu.cpp:288-292
 *                               + HashTableGPUDriver::insert_supermer_block         src/kcount/kcount-gpu/gpu_hash_table.cpp:656-735
 *   kc_finalize              <- HashTableInserter::insert_into_local_hashtable      src/kcount/kcount_gpu.cpp:372-506
 *                               + HashTableGPUDriver::done_all_inserts              src/kcount/kcount-gpu/gpu_hash_table.cpp:737-785
 *   kc_result_*              <- begin_iterate / get_next_entry                      src/kcount/kcount-gpu/gpu_hash_table.cpp:812-827
 *                               + KmerDHT::local_kmers_begin/end, dump_kmers        src/kcount/kmer_dht.cpp:274-307
 *   kc_owner_of              <- KmerDHT::get_kmer_target_rank                       src/kcount/kmer_dht.cpp:193-196
 *   kc_lookup                <- KmerDHT::get_local_kmer_counts / kmer_exists /
 *                               get_kmer_count                                      src/kcount/kmer_dht.cpp:199-245
 *   kc_get_stats             <- HashTableGPUDriver::get_stats + the stage's reduce_one() counters
 *                                                                                   src/kcount/kcount_gpu.cpp:294-364
 *   kc_run                   <- analyze_kmers<MAX_K> (single rank)                  src/kcount/kcount.cpp:143-161
 *
 * Semantics are those of the reference's CPU build (src/kcount/kcount_cpu.cpp): N counts as G inside a
 * k-mer and never as an extension, quality only masks extension bases, the first and last k-mer of a
 * read are not counted, counts and extension counts saturate at 65535.
 */
#ifndef KCOUNT_B200_H
#define KCOUNT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KC_ABI_VERSION 2
#define KC_MAX_K 127       /* 4 x u64 key words; k must not be a multiple of 32 */
#define KC_MAX_KEY_WORDS 4

typedef enum {
  KC_OK = 0,
  KC_ERR_ARG = -1,        /* bad argument / call out of order */
  KC_ERR_CUDA = -2,       /* CUDA runtime error or no device */
  KC_ERR_NOMEM = -3,      /* device or host allocation failed */
  KC_ERR_TABLE_FULL = -4, /* a k-mer could not be placed: results undefined, re-run with a larger table */
  KC_ERR_BINS_FULL = -5,  /* internal: supermer bin overflow that the automatic retry could not fix */
  KC_ERR_STATE = -6,
  KC_ERR_IO = -7
} kc_status;

typedef struct kc_ctx kc_ctx;

/* Stage configuration. Zero-initialise, then kc_config_default(), then override. */
typedef struct {
  uint32_t abi_version;      /* KC_ABI_VERSION */
  uint32_t k;                /* options.hpp:81 kmer_lens */
  uint32_t m;                /* minimizer length; 0 = reference rule clamp(2k/3+1, 15, 27), kmer_dht.cpp:117-119 */
  int32_t qual_offset;       /* 33 or 64, options.hpp:85 */
  int32_t qual_cutoff;       /* KCOUNT_QUAL_CUTOFF = 20, CMakeDefinitions.txt:58 */
  int32_t dmin_thres;        /* options.hpp:89, default 2 */
  double dyn_min_depth;      /* DYN_MIN_DEPTH = 0.9, CMakeDefinitions.txt:70 */
  int32_t use_filter;        /* --use-qf: screen singletons with the two-choice filter before table insertion */
  int32_t rank;              /* this process' rank (owner id) ... */
  int32_t n_ranks;           /* ... out of n_ranks owners; 1 = single GPU */
  int32_t device;            /* CUDA device ordinal, -1 = current */
  uint64_t est_kmers;        /* my_num_kmers: estimated raw k-mers this rank will receive, contigging.cpp:86 */
  int32_t sequencing_depth;  /* options.hpp:106, default 4 */
  double base_error_rate;    /* BASE_ERROR_RATE = 0.005 */
  uint64_t table_slots;      /* 0 = reference sizing rule (kcount_cpu.cpp:428-470); else exact slot count */
  uint64_t mem_budget_bytes; /* 0 = 80 % of free device memory */
  uint32_t local_bins;       /* sub-bins per owner; 0 = automatic. Multi-rank: must be equal on all ranks (automatic is, when
                                est_kmers is equal on all ranks or 0). HBM-table mode: power of two */
  uint32_t record_bytes;     /* supermer record size 16/32/64; 0 = automatic */
  uint32_t flags;            /* KC_FLAG_* */
  uint32_t reserved[7];
} kc_config;

#define KC_FLAG_KEEP_EXT_COUNTS 1u /* finalize also keeps the 8 extension counters + X/F entries (dump D2) */
#define KC_FLAG_NO_WARP_DEDUP 2u   /* disable the merging of equal consecutive k-mer occurrences before a table update */
#define KC_FLAG_GLOBAL_TABLE 4u    /* count in one open-addressing table in HBM (the layout use_filter needs) instead of the
                                      default bin-resident shared-memory tables */

/* Counters the reference logs (SURVEY.md section 5 "Metrics") plus what the roofline needs. */
typedef struct {
  uint64_t n_reads;            /* reads handed in */
  uint64_t n_bases;
  uint64_t raw_kmers;          /* sum(len-k+1), kcount.cpp:86 */
  uint64_t counted_kmers;      /* sum(max(0,len-k-1)) produced locally = occurrences leaving kc_parse */
  uint64_t supermers_out;      /* supermer records produced locally */
  uint64_t supermer_bytes_out; /* bytes of those records */
  uint64_t supermers_in;       /* records inserted on this rank (KmerDHT::get_num_supermer_inserts) */
  uint64_t inserted_kmers;     /* occurrences inserted on this rank */
  uint64_t unique_kmers;       /* distinct keys that reached the table ("Number of elements") */
  uint64_t filtered_singletons;/* distinct keys stopped by the filter (use_filter only; includes none that recur) */
  uint64_t ge2_kmers;          /* keys with count >= 2 */
  uint64_t final_kmers;        /* survivors: count >= 2 and both extensions in ACGT ("Total kmers") */
  uint64_t purged_kmers;       /* unique_kmers - final_kmers */
  uint64_t sum_counts;         /* "Total kmer count sum", kcount_cpu.cpp:598 */
  uint64_t max_count;
  uint64_t dropped;            /* always 0 on success: a drop is KC_ERR_TABLE_FULL */
  uint64_t table_slots;        /* capacity actually allocated */
  uint64_t table_bytes;
  uint64_t filter_bytes;
  double avg_kmer_count;       /* sum_counts / final_kmers, rank-local; kcount_cpu.cpp:594-596 */
  double ms_parse, ms_insert, ms_finalize; /* device time of the last run of each phase (CUDA events) */
  uint64_t kernel_launches;    /* kernels launched by this context so far */
  /* device time (CUDA events on the context's stream) of the kernels alone, last run of each phase */
  double ms_parse_kernel;      /* supermer kernel(s) over all read chunks, without the capacity-sampling pass */
  double ms_insert_kernel;     /* insert kernel (both passes when the filter is on), without table clearing / planning */
  double ms_table_clear;       /* zeroing of table, filter and side table before the insert */
  uint64_t insert_kernel_launches; /* launches covered by ms_insert_kernel */
  uint64_t local_bins;         /* sub-bins per owner actually used */
  uint64_t split_passes;       /* bin-resident mode: extra passes over bins whose distinct k-mers did not fit shared memory */
  uint64_t result_bytes;       /* device bytes of the dense result arrays */
  /* the resident insert pipeline (refine -> count, slice by slice) */
  double ms_refine_kernel;     /* refine kernels (second partition level), sum over the slices, on their own stream */
  double ms_exchange_copy;     /* copy-mode exchange: first copy start to last copy end on the transfer streams */
  uint64_t exchange_copy_bytes;/* bytes the copy engines moved from the peers in that window */
  uint64_t pipeline_slices;    /* slices the last kc_insert was cut into */
} kc_stats;

/* ---- lifecycle ---------------------------------------------------------------------------------- */
void kc_config_default(kc_config *cfg);
int kc_create(const kc_config *cfg, kc_ctx **out);
void kc_destroy(kc_ctx *ctx);
const char *kc_last_error(const kc_ctx *ctx); /* ctx may be NULL: last error of a failed kc_create */
const char *kc_version(void);
int kc_device_count(void);

/* ---- input -------------------------------------------------------------------------------------- */
/* Append reads. bytes: PackedRead encoding, one byte per base = base(A0 C1 G2 T3 N4) | min(q,31) << 3.
 * offsets[i] = start of read i inside `bytes`; lens[i] = its length. n_bytes = extent of `bytes`.
 * Pointers may be host (pageable or pinned) or device memory; the data is copied to the device (async
 * on the context's input stream when pinned/device) and may be released or overwritten after kc_sync(). */
int kc_add_reads(kc_ctx *ctx, const uint8_t *bytes, uint64_t n_bytes, const uint64_t *offsets, const uint16_t *lens,
                 uint64_t n_reads);
/* Append the reads of a reference PackedReads container AS THEY SIT IN MEMORY: `headers` points at its vector<PackedRead>
 * (src/packed_reads.hpp:68-79, 136): n_reads records of header_stride bytes (0 = 16) = one 64-bit word of bit fields
 * {is_allocated : 1, read_id : 47, read_len : 16} followed by the pointer to the read's bytes inside a 4 MB allocator block
 * (src/linear_allocator_pool.cpp). Replaces the per-read unpack loop of count_kmers (src/kcount/kcount.cpp:71-77,
 * src/packed_reads.cpp:196-208, 269-276): host threads gather the bytes piece by piece into pinned staging while the
 * previous piece crosses PCIe; the offsets are built by a scan on the device. Returns when the last piece has been copied
 * (the container may be released). */
int kc_add_packed_reads(kc_ctx *ctx, const void *headers, uint64_t n_reads, uint32_t header_stride);
int kc_clear_reads(kc_ctx *ctx);

/* ---- read-pair merging (the step in front of the k-mer analysis) -------------------------------------------------
 * Replaces the pair loop of merge_reads() (src/merge_reads.cpp:466-650, called at src/main.cpp:210-215) and the PackedRead
 * encoding of its output (src/packed_reads.cpp:99-126), on the GPU (kc_merge.cu). Input: the mates as kc_fastq_read_pairs
 * (kc_synth.h) delivers them -- reads 2 i and 2 i + 1 are pair i, sequence and quality CHARACTERS, offsets[] / lens[] into
 * seq / qual (n_bytes = their extent). Output, per pair and in input order: the merged read followed by the one-base dummy
 * mate "N", or the two mates as they came (mate 1 with the qualities an N comparison zeroed, as in the reference); a pair of
 * two mates shorter than min_kmer_len is dropped (paired != 0 only). No adapter trimming (src/adapters.cpp). Mates longer than
 * KC_MERGE_MAX_READ_LEN are passed through unmerged and counted in too_long. Illegal base or quality characters: KC_ERR_ARG
 * (the reference aborts). Host or device input pointers. */
#define KC_MERGE_MAX_READ_LEN 512
typedef struct {
  uint64_t pairs, merged, ambiguous; /* "merged", "ambiguous" of the reference's summary (src/merge_reads.cpp:688-690) */
  uint64_t merged_len, overlap_len;  /* sums over the merged pairs: average merged length / overlap length */
  uint64_t max_read_len;             /* longest merged read */
  uint64_t dropped_short, too_long;
  uint64_t sum_read_lens;            /* the value merge_reads returns is sum_read_lens / (2 x pairs) */
  uint64_t out_bytes;                /* extent of the output byte buffer = sum of all input lengths */
  uint64_t launches;                 /* kernels launched */
  double ms_kernel;
} kc_merge_stats;
/* host arrays out: out_bytes [n_bytes], out_offsets / out_lens [2 n_pairs]; *n_out_reads = reads written (2 per kept pair) */
int kc_merge_pairs(int device, const char *seq, const char *qual, const uint64_t *offsets, const uint16_t *lens, uint64_t n_pairs,
                   uint64_t n_bytes, int qual_offset, int min_kmer_len, int paired, uint8_t *out_bytes, uint64_t *out_offsets,
                   uint16_t *out_lens, uint64_t *n_out_reads, kc_merge_stats *stats);
/* the same, with the result appended to a stage context without leaving HBM (a dropped pair stays behind as two empty reads) */
int kc_add_read_pairs(kc_ctx *ctx, const char *seq, const char *qual, const uint64_t *offsets, const uint16_t *lens, uint64_t n_pairs,
                      uint64_t n_bytes, int qual_offset, int min_kmer_len, int paired, kc_merge_stats *stats);

/* ---- the stage, phase by phase (all asynchronous on the context's stream unless noted) ----------- */
/* (1)+(2) parse reads, extract minimizers, cut supermers, bin them per owner rank and local bin. */
int kc_parse(kc_ctx *ctx);
/* Exchange step. After kc_parse: send_bytes[r] / send_ptr(r) describe the contiguous device buffer of
 * supermer records destined to owner r (r == rank stays local). The caller moves the data with its own
 * transport (NCCL all-to-all in this repo; UPC++ RPCs in the reference) into the buffer returned by
 * kc_exchange_recv_buffer and then calls kc_exchange_done. With n_ranks == 1 none of this is needed. */
int kc_exchange_send_sizes(kc_ctx *ctx, uint64_t *send_bytes /* [n_ranks] */); /* synchronises the stream */
int kc_exchange_send_ptr(kc_ctx *ctx, int dst_rank, void **dev_ptr);
int kc_exchange_recv_buffer(kc_ctx *ctx, const uint64_t *recv_bytes /* [n_ranks] */, void **dev_ptr);
/* byte offset of each source rank's block inside the receive buffer (256-byte aligned) */
int kc_exchange_recv_offsets(kc_ctx *ctx, uint64_t *offsets /* [n_ranks] */);
int kc_exchange_done(kc_ctx *ctx);
/* Peer-memory exchange (ranks = GPUs of one NVLink / NVSwitch node). Instead of moving every owner's buffer with an
 * all-to-all and then partitioning it, a rank EXPORTS its send buffer (CUDA IPC handle + the offset / used bytes of every
 * owner's region in it), every owner IMPORTS the regions meant for it, and kc_insert's refine kernel pulls the records
 * straight out of the peers' HBM over NVLink while it sorts them into fine bins: one kernel does the transfer and the
 * partition, only used records cross the links, nothing is staged. Replaces the same reference calls as the block
 * above (KmerDHT::add_supermer / flush_updates, src/kcount/kmer_dht.cpp:248-258).
 * Protocol: kc_parse on every rank, a barrier; kc_exchange_export; all-gather of (handle, offsets, bytes);
 * kc_exchange_import for every other rank; kc_exchange_done; kc_insert; kc_finalize; a barrier before the next kc_parse /
 * kc_reset / kc_add_reads of any rank (peers may still be reading this rank's buffer until their kc_finalize returns). */
#define KC_IPC_HANDLE_BYTES 64
int kc_exchange_export(kc_ctx *ctx, uint8_t *handle /* [KC_IPC_HANDLE_BYTES] */, uint64_t *region_offsets /* [n_ranks] */,
                       uint64_t *region_bytes /* [n_ranks] */);
int kc_exchange_import(kc_ctx *ctx, int src_rank, const uint8_t *handle, uint64_t region_offset, uint64_t region_bytes);
/* How kc_insert gets at the imported regions. KC_XCHG_PULL (default): the refine / count kernels read the peers' memory
 * directly (transfer fused into the partition kernel). KC_XCHG_COPY: copy engines move the regions slice by slice into a
 * local buffer on separate streams while the refine and count kernels work on the previous slices -- no SM does transfer
 * work, and the NVLink time hides behind the count pass. For slice-granular copies the exporter's slice table travels
 * with the handle: kc_exchange_export_slices fills offsets[n_ranks][KC_MAX_SLICES + 1] (byte offset of every slice edge
 * inside each owner's region; entries beyond n_slices repeat the end), the importer hands row [its own rank] of source s
 * to kc_exchange_import_slices. Without a table the whole region moves with the first slice. */
#define KC_XCHG_PULL 0
#define KC_XCHG_COPY 1
#define KC_MAX_SLICES 12
int kc_exchange_set_mode(kc_ctx *ctx, int mode);
int kc_exchange_export_slices(kc_ctx *ctx, uint32_t *n_slices, uint64_t *offsets /* [n_ranks * (KC_MAX_SLICES + 1)] */);
int kc_exchange_import_slices(kc_ctx *ctx, int src_rank, uint32_t n_slices, const uint64_t *offsets /* [n_slices + 1] */);
/* After that barrier: frees send buffers this rank exported earlier and has since replaced by a larger one (they stay
 * allocated until here because a peer may still have had them mapped; an importer closes a stale mapping when the
 * source's next export carries a new handle). Cheap no-op otherwise. */
int kc_exchange_release(kc_ctx *ctx);
/* (3)+(4) insert every received supermer's k-mers into the table (through the filter when enabled). */
int kc_insert(kc_ctx *ctx);
/* (5) purge count<2, resolve extensions, compact survivors. Synchronises and fills stats. */
int kc_finalize(kc_ctx *ctx, kc_stats *stats);
/* kc_parse + kc_insert + kc_finalize for n_ranks == 1 == analyze_kmers(). */
int kc_run(kc_ctx *ctx, kc_stats *stats);
/* Forget table contents and results but keep reads and allocations (bench loop / iterative k). */
int kc_reset(kc_ctx *ctx);
int kc_sync(kc_ctx *ctx);
int kc_get_stats(kc_ctx *ctx, kc_stats *stats);
/* Device-side stopwatch for callers (bench.py): kc_timer_mark records CUDA event `slot` (0..7) on the context's stream;
 * kc_timer_elapsed_ms synchronises on event `to` and returns the device time between two marks. */
int kc_timer_mark(kc_ctx *ctx, int slot);
int kc_timer_elapsed_ms(kc_ctx *ctx, int from, int to, double *ms);

/* ---- results ------------------------------------------------------------------------------------ */
uint32_t kc_key_words(const kc_ctx *ctx); /* N_LONGS = ceil(k/32) */
/* number of survivors (extended = 0) or of entries with count >= 2 incl. X/F (extended = 1, needs
 * KC_FLAG_KEEP_EXT_COUNTS). Order of entries is unspecified unless kc_result_sort was called. */
int kc_result_size(kc_ctx *ctx, int extended, uint64_t *n);
/* Sort the device-resident result by key words (lexicographic, unsigned) -- the order of dumps D1/D2. */
int kc_result_sort(kc_ctx *ctx, int extended);
/* Copy to HOST buffers (any may be NULL): keys[n*N] u64, counts[n] u16, left[n]/right[n] chars in
 * ACGT (or X/F when extended), ext8[n*8] u16 = lA lC lG lT rA rC rG rT (extended only). */
int kc_result_copy(kc_ctx *ctx, int extended, uint64_t *keys, uint16_t *counts, char *left, char *right, uint16_t *ext8);
/* Host sink for the survivors: when set before kc_insert, the entries are copied to these HOST buffers (pinned memory for
 * full speed) WHILE the count kernel is still producing the rest, instead of by a kc_result_copy afterwards -- the
 * device -> host hand-off HashTableInserter::done_all_inserts / insert_into_local_hashtable does at the end of the
 * reference's GPU path (src/kcount/kcount-gpu/gpu_hash_table.cpp:737-785, src/kcount/kcount_gpu.cpp:372-506). After
 * kc_finalize the first stats.final_kmers entries are valid (unsorted, the order of kc_result_copy without
 * kc_result_sort). `capacity` is in entries; a result that does not fit makes kc_finalize fail with KC_ERR_ARG.
 * keys == NULL removes the sink. */
int kc_set_result_sink(kc_ctx *ctx, uint64_t *keys, uint16_t *counts, char *left, char *right, uint64_t capacity);
/* Device pointers of the same arrays (valid until kc_reset/kc_destroy). */
int kc_result_device_ptrs(kc_ctx *ctx, int extended, const uint64_t **keys, const uint16_t **counts, const char **left,
                          const char **right, const uint16_t **ext8);
/* 64-bit order-independent checksum of the result (sum over entries of a mix of key words, count, L, R
 * and, when extended, the 8 ext counters); computed on the device. Same function in oracle/. */
int kc_result_checksum(kc_ctx *ctx, int extended, uint64_t *checksum);
/* The same checksum WITHOUT ANY ALLOCATION, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it). No counterpart in the reference.
 * The caller passes the result arrays as DEVICE pointers (e.g. from kc_result_device_ptrs; ext8 may be NULL) and the
 * destination (host or device). Scratch is a file-scope device array of KC_NOALLOC_SCRATCH_WORDS x 8 bytes (32 KB) that every
 * GPU which loads the library holds until exit; its address is looked up once per device (cudaGetSymbolAddress). The call makes
 * no cudaMalloc / cudaFree / stream / event call, runs on the legacy default stream and is complete on return. More than
 * KC_NOALLOC_MAX_ENTRIES entries: KC_ERR_ARG. ALL calls share the scratch array (every word a call reads it wrote itself, so
 * earlier contents never leak into a result): calls that overlap on the GPU race -- serialise them. */
#define KC_NOALLOC_SCRATCH_WORDS 4096
#define KC_NOALLOC_MAX_ENTRIES (1ull << 31)
int kc_checksum_noalloc(int device, const uint64_t *keys, const uint16_t *counts, const char *left, const char *right, const uint16_t *ext8,
                        uint64_t n_entries, uint32_t n_words, uint64_t *checksum);
/* Text dump, one "KMER count L R" line per survivor, sorted by key (kmer_dht.cpp:274-297 writes the same
 * lines unsorted and gzipped). extended adds the 8 counters. */
int kc_dump_sorted(kc_ctx *ctx, const char *path, int extended);
/* KmerDHT::dump_kmers as the reference writes it (src/kcount/kmer_dht.cpp:274-297): one "KMER count L R" line per survivor
 * of this rank, in table order (unsorted), gzip-compressed when gz != 0 (the reference's kmers-<k>.txt.gz per rank). */
int kc_dump_kmers(kc_ctx *ctx, const char *path, int gz);
/* Batched lookup in the finalized result (replaces the point lookups KmerDHT::get_local_kmer_counts / kmer_exists /
 * get_kmer_count, src/kcount/kmer_dht.cpp:199-245, and the host map they search, kcount_gpu.cpp:439-471):
 * keys[n*N] canonical k-mers -> counts (0 = absent: purged, dropped as X / F, or never seen), L, R (0 when absent).
 * The first call after kc_finalize builds a device-resident index over the result; every call runs one query kernel.
 * keys and the outputs are either all HOST pointers (staged through HBM in pieces) or all DEVICE pointers (queried in
 * place); outputs may be NULL. Works on the unsorted result; kc_result_sort invalidates the index. */
int kc_lookup(kc_ctx *ctx, const uint64_t *keys, uint64_t n, uint16_t *counts, char *left, char *right);
/* Owner rank of a canonical k-mer (key_words[N]) under this context's m and n_ranks. Host-side, no GPU. */
int kc_owner_of(const kc_ctx *ctx, const uint64_t *key_words);
uint32_t kc_minimizer_len(const kc_ctx *ctx);

/* ---- host helpers: pinned memory ------------------------------------------------------------------- */
int kc_host_alloc(void **ptr, size_t bytes); /* cudaHostAlloc */
int kc_host_free(void *ptr);

/* seeded synthetic PackedRead streams and read files: include/kc_synth.h (libkc_synth.so, host-only, no CUDA) */

#ifdef __cplusplus
}
#endif
#endif /* KCOUNT_B200_H */
