// kc_api.cu -- context, memory and phase orchestration behind the C ABI of include/kcount_b200.h.
// Host C++ only calls CUDA from here; no CUDA type crosses the ABI. There is no CPU fallback anywhere: every phase is a
// kernel launch on the context's stream.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

#include <zlib.h>

#include "kc_internal.h"
#include "kcount_b200.h"

using namespace kc;

static thread_local std::string g_create_error;

struct ReadChunk {
  uint8_t *d_bytes = nullptr;   // base of the byte buffer the offsets refer to
  uint64_t *d_offsets = nullptr;
  uint16_t *d_lens = nullptr;
  uint64_t n_reads = 0, n_bytes = 0;
  bool owner = true;            // frees the buffers (the sub-chunks of one kc_add_reads call share them)
  uint64_t cap_bytes = 0, cap_reads = 0;
  cudaEvent_t ready = nullptr;  // the input copy of this chunk has arrived (recorded on in_stream)
};

enum Phase { PH_EMPTY = 0, PH_PARSED, PH_RECV_READY, PH_INSERTED, PH_FINALIZED };

struct ResultSet {
  uint64_t *keys = nullptr;
  uint16_t *counts = nullptr, *ext8 = nullptr;
  char *left = nullptr, *right = nullptr;
  uint64_t n = 0, cap = 0;
  bool sorted = false;
  // lookup index over the entries (kc_lookup.cu), built by the first kc_lookup after a finalize
  unsigned long long *index = nullptr;
  uint64_t index_slots = 0;
  bool indexed = false;
};

static const int kSinkSlices = kSlices;
static const double kSampleMaxFrac = 0.2;   // default of KC_SAMPLE_MAX_FRAC (estimate_caps)
static const double kPresplitFactor = 1.0;  // default of KC_PRESPLIT (count_params): the per-warp threshold starts at the host's estimate; 0 = bins are only split on overflow
static_assert(KC_MAX_SLICES == kSlices, "the ABI's slice table and the pipeline disagree");
static const int kXferStreams = 4;  // copy-engine streams of the copy-mode exchange (peers are dealt round-robin)
static const int kPinnedWords = 4 * kCountMaxSrc + 64;  // pinned staging for small host <-> device parameter copies
static const int kScratchWords = 64;                    // device scratch (checksums, small reductions)

struct kc_ctx {
  kc_config cfg;
  int k = 0, m = 0, N = 0, rec_bytes = 0, P = 1, G = 1, rank = 0;
  uint32_t n_ids = 1;
  bool resident = true;       // bin-resident shared-memory counting (kc_count.cu) vs one table in HBM (kc_insert.cu)
  uint32_t count_slots = 0;   // resident: table slots per CTA
  uint8_t **d_count_src = nullptr;
  // two-level partition: P coarse bins per owner (parse, layout, exchange), each refined into 1 << fine_shift fine bins
  // (kc_refine.cu) before counting
  uint32_t fine_shift = 0;
  uint8_t *d_fine = nullptr;
  uint64_t fine_alloc = 0;
  unsigned long long *d_coarse_base = nullptr;
  uint64_t coarse_base_alloc = 0;
  // host sink for the survivors (kc_set_result_sink): the count kernel runs in slices of bins and every finished slice
  // is copied out on a second stream while the next ones are counted
  uint64_t *sink_keys = nullptr;
  uint16_t *sink_counts = nullptr;
  char *sink_left = nullptr, *sink_right = nullptr;
  uint64_t sink_cap = 0, sink_done = 0;
  bool sink_streamed = false;
  cudaStream_t copy_stream = nullptr;
  // input side: the reads of one kc_add_reads call are copied in sub-chunks on in_stream, each with its own event, so
  // that kc_parse can start on the first reads while the rest is still crossing PCIe
  cudaStream_t in_stream = nullptr;
  std::vector<cudaEvent_t> ev_in_pool;
  size_t ev_in_used = 0;
  ReadChunk spare;  // buffers of the last kc_clear_reads, reused by the next kc_add_reads
  // kc_add_packed_reads: the reads of a PackedReads container are gathered into two pinned staging buffers by host threads
  // and copied from there; ev_stage[i] = the copy out of buffer i has completed
  uint8_t *h_stage[2] = {nullptr, nullptr};
  cudaEvent_t ev_stage[2] = {nullptr, nullptr};
  bool stage_busy[2] = {false, false};
  double packed_gather_ms = 0;
  cudaEvent_t ev_slice[kSinkSlices];
  unsigned long long *h_slice_n = nullptr;  // pinned: n_final after every slice
  unsigned long long *h_pinned = nullptr;   // pinned staging [kPinnedWords]
  unsigned long long *d_scratch = nullptr;  // [kScratchWords]
  // refine / receive pipeline: slice i of the coarse bins is refined on refine_stream while slice i - 1 is counted
  cudaStream_t refine_stream = nullptr;
  cudaEvent_t ev_rs0[kSinkSlices], ev_rs1[kSinkSlices];  // refine slice i: start / end (refine_stream)
  cudaEvent_t ev_cs0[kSinkSlices], ev_cs1[kSinkSlices];  // count slice i: start / end (stream)
  int n_count_slices = 0, n_refine_slices = 0;
  unsigned int *d_slice_done = nullptr;  // [kSlices] coarse bins refined per slice (SM-partitioned pipeline)
  bool partitioned = false;              // the last insert ran refine and count side by side on disjoint SMs
  // copy-mode exchange (kc_exchange_set_mode(KC_XCHG_COPY)): copy engines move the peers' regions into d_recv
  int xmode = KC_XCHG_PULL;
  cudaStream_t xfer_stream[kXferStreams] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_x[kXferStreams], ev_xbeg[kXferStreams], ev_xend[kXferStreams];
  bool xfer_started[kXferStreams] = {false, false, false, false};
  uint64_t xfer_bytes = 0;
  std::vector<uint64_t> h_slice_off;                  // [G * (kSlices + 1)] slice edges inside each owner's region of MY send buffer
  std::vector<std::vector<uint64_t>> peer_slice_off;  // [G][n_slices + 1] the same table of every source, for my region in it
  uint64_t last_final = 0, last_ext = 0;              // result sizes of this context's previous run (result reservation)
  // pooled staging of kc_lookup's host batches
  uint64_t *d_lq = nullptr;
  uint16_t *d_lc = nullptr;
  char *d_ll = nullptr, *d_lr = nullptr;
  uint64_t lookup_cap = 0;
  int device = 0, sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t ev_pk0 = nullptr, ev_pk1 = nullptr, ev_ik0 = nullptr, ev_ik1 = nullptr, ev_clr = nullptr;
  // the parse launches of consecutive read chunks alternate between `stream` and this one, so that the CTAs of the next
  // chunk move in while the last CTAs of the previous one run out (a persistent grid idles ~15 % of a launch otherwise)
  cudaStream_t parse_stream2 = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  cudaEvent_t ev_user[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  std::string err;
  Phase phase = PH_EMPTY;

  std::vector<ReadChunk> chunks;
  uint64_t n_reads = 0, n_bases = 0;

  // supermer bins (send side)
  uint8_t *d_bins = nullptr;
  uint64_t bins_bytes = 0, bins_alloc = 0;
  int header_bytes = 0;
  std::vector<uint64_t> h_region_off, h_region_bytes;
  std::vector<unsigned long long> h_owner_used;  // per owner: records up to the end of its last non-empty bin
  unsigned long long *d_block_sums = nullptr;
  unsigned long long *d_owner_vals = nullptr;    // [owner_records | region_off (+total) | region_bytes | owner_used]
  uint64_t *d_bin_base = nullptr, *d_bin_first = nullptr, *d_bin_cap = nullptr;
  unsigned long long *d_cursor = nullptr;
  uint8_t **d_hdr_ptrs = nullptr;
  bool caps_valid = false;

  // receive side
  std::vector<uint8_t *> pull_ptr;                         // peer-memory exchange: region of source s in ITS memory (IPC-mapped)
  // opened IPC mappings, one per source rank (handle bytes -> base pointer). A source that re-allocated its send buffer
  // exports a new handle: the stale mapping is closed before the new one is opened (kc_exchange_import); the exporter
  // keeps the old allocation alive until kc_exchange_release, i.e. until every importer has let go of it.
  std::vector<std::pair<std::string, void *>> ipc_mapped;
  bool bins_exported = false;            // d_bins has been handed to peers since it was allocated
  std::vector<uint8_t *> retired_bins;   // exported send buffers replaced by a larger one, freed by kc_exchange_release
  uint8_t *d_recv = nullptr;
  uint64_t recv_alloc = 0;
  std::vector<uint64_t> recv_off, recv_bytes;
  uint8_t **d_src_ptr = nullptr;
  Segment *d_segments = nullptr;
  uint32_t *d_seg_tile_start = nullptr;
  Region *d_regions = nullptr;
  unsigned long long *d_total_records = nullptr;

  // table
  Slot *d_A = nullptr;
  uint4 *d_B = nullptr;      // N >= 3
  uint32_t *d_C = nullptr;   // N >= 2
  uint64_t table_slots = 0, table_alloc_slots = 0;
  uint32_t *d_filter = nullptr;
  uint64_t filter_buckets = 0, filter_alloc_buckets = 0;
  Heavy heavy{nullptr, nullptr, nullptr, 0};

  DevCounters *d_ctr = nullptr;
  DevCounters h_ctr;

  ResultSet res_final, res_ext;
  kc_stats stats;
  uint64_t launches = 0;
  kc_ctx() {
    for (auto &e : ev_slice) e = nullptr;
    for (auto *arr : {ev_rs0, ev_rs1, ev_cs0, ev_cs1})
      for (int i = 0; i < kSinkSlices; i++) arr[i] = nullptr;
    for (auto *arr : {ev_x, ev_xbeg, ev_xend})
      for (int i = 0; i < kXferStreams; i++) arr[i] = nullptr;
  }
};

// ------------------------------------------------------------------------------------------------ helpers
static int fail(kc_ctx *c, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (c)
    c->err = buf;
  else
    g_create_error = buf;
  return code;
}

#define CK(call)                                                                                           \
  do {                                                                                                     \
    cudaError_t e_ = (call);                                                                               \
    if (e_ != cudaSuccess)                                                                                 \
      return fail(c, e_ == cudaErrorMemoryAllocation ? KC_ERR_NOMEM : KC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                  cudaGetErrorString(e_));                                                                 \
  } while (0)

template <typename T>
static cudaError_t dev_alloc(T **p, uint64_t count) {
  return cudaMalloc((void **)p, std::max<uint64_t>(count, 1) * sizeof(T));
}
template <typename T>
static void dev_free(T *&p) {
  if (p) cudaFree(p);
  p = nullptr;
}

static int reference_minimizer_len(int k) {  // KmerDHT::KmerDHT, src/kcount/kmer_dht.cpp:117-119
  int m = k * 2 / 3 + 1;
  if (m < 15) m = 15;
  if (m > 27) m = 27;
  return m;
}

static int auto_record_bytes(int k, int m, int n_ids) {
  if (n_ids == 1) return 64;  // no owner/bin cuts: a record carries up to 160 - k - 1 k-mers of a read
  int w = k - m + 1;          // expected supermer ~ (w + 1) / 2 k-mers; leave room for about twice that
  int want = k + 1 + (w + 1);
  if (want <= 32) return 16;
  if (want <= 80) return 32;
  return 64;
}


// Number of local bins, record size and header geometry. Called at creation and again by kc_parse once the amount of
// input is known (automatic sizing). Everything that depends on n_ids is released when it changes.
static void configure_bins(kc_ctx *c) {
  const kc_config &cfg = c->cfg;
  const double depth = cfg.sequencing_depth > 0 ? cfg.sequencing_depth : 4;
  const double kmer_err = 1.0 - std::pow(1.0 - cfg.base_error_rate, (double)c->k);
  uint32_t P = cfg.local_bins;
  const int R_multi = cfg.record_bytes ? (int)cfg.record_bytes : auto_record_bytes(c->k, c->m, 2);
  if (!P) {
    if (c->resident) {
      // distinct k-mers this rank will own, by the reference's own sizing rule (kmer_dht.cpp:126-131, kcount_cpu.cpp:443-446),
      // spread over bins whose shared-memory table is then about half full
      uint64_t est = cfg.est_kmers;
      if (!est && c->G == 1) est = c->n_bases > c->n_reads * (uint64_t)(c->k - 1) ? c->n_bases - c->n_reads * (uint64_t)(c->k - 1) : 0;
      if (!est && c->G > 1) {
        P = 8192;  // no rank-independent estimate: a fixed count keeps the ranks' headers compatible; big bins are split
      } else {
        const double distinct = (double)est / depth + (double)est * kmer_err;
        double load = 0.6;
        if (const char *e = getenv("KC_BIN_LOAD")) load = std::min(std::max(atof(e), 0.1), 2.0);  // (experiments)
        const double slots = count_table_slots(c->N, R_multi, c->k);
        double per_bin = load * slots;
        // k-mers arrive in a bin in GROUPS (all k-mers of a minimizer instance, ~(k - m + 2) / 2 of them): with long windows a
        // bin holds few, big groups and its population scatters widely around the mean -- measured on C4 (k = 77 / 99): a third
        // of the bins overflowed their table and were split, i.e. walked three times. Keep mu + 3 sqrt(g mu) under the claim
        // limit (0.85 slots); the factor 1.25 is the reference estimate's own margin over what the data shows.
        if (!getenv("KC_BIN_LOAD")) {
          const double g = std::max(1.0, (c->k - c->m + 2) / 2.0), lim = 0.85 * slots, z = 3.0;
          const double x = (-z * std::sqrt(g) + std::sqrt(z * z * g + 4.0 * lim)) / 2.0;  // sqrt(mu): x^2 + z sqrt(g) x - lim = 0
          per_bin = std::min(per_bin, 1.25 * x * x);
        }
        P = (uint32_t)std::min<double>(std::ceil(distinct / per_bin), (double)(1u << 26));
        if (P < 1) P = 1;
      }
    } else {
      // HBM table: sub-bins keep the table region a bin touches L2-resident once the table is large (measured 2x on C2)
      uint64_t est = cfg.est_kmers ? cfg.est_kmers : (c->n_bases > c->n_reads * (uint64_t)(c->k - 1) ? c->n_bases - c->n_reads * (uint64_t)(c->k - 1) : 0);
      const double table_bytes = ((double)est / depth + (cfg.use_filter ? 0.0 : (double)est * kmer_err)) / 0.66 * 36.0;
      P = table_bytes > 512e6 ? 1024 : 1;
    }
  }
  // P so far is the number of FINE bins the count kernel wants. Beyond 64 K of them the parse routes by coarse bin
  // (16 fine bins each) and kc_refine.cu does the second level.
  // With G owners the parse routes into G x P coarse bins, so the refine factor grows with G (16, 32, 64) to keep
  // that product -- what the parse kernel's scattered writes and the capacity sample pay for -- where it is at G = 1.
  uint32_t fine_shift = 0;
  if (c->resident && P >= 65536u) {
    fine_shift = c->G >= 4 ? 6 : (c->G >= 2 ? 5 : 4);
    while (fine_shift < 6 && (uint64_t)(P >> fine_shift) * c->G > 700000u) fine_shift++;  // (long k: many small fine bins)
    if (const char *e = getenv("KC_FINE_SHIFT")) fine_shift = (uint32_t)std::min(std::max(atoi(e), 1), 6);  // (experiments)
    P = (P + (1u << fine_shift) - 1u) >> fine_shift;
  }
  while ((uint64_t)P * c->G > (1u << 25)) P /= 2;
  c->fine_shift = fine_shift;
  const uint32_t n_ids = P * (uint32_t)c->G;
  int R = cfg.record_bytes ? (int)cfg.record_bytes : auto_record_bytes(c->k, c->m, (int)n_ids);
  while (record_geom(R, c->k).max_kmers < 1 && R < 64) R *= 2;
  if (n_ids != c->n_ids || R != c->rec_bytes || (int)P != c->P) {
    dev_free(c->d_bin_base);
    dev_free(c->d_bin_first);
    dev_free(c->d_bin_cap);
    dev_free(c->d_cursor);
    dev_free(c->d_hdr_ptrs);
    dev_free(c->d_src_ptr);
    dev_free(c->d_segments);
    dev_free(c->d_seg_tile_start);
    dev_free(c->d_regions);
    dev_free(c->d_total_records);
    c->caps_valid = false;
  }
  c->P = (int)P;
  c->n_ids = n_ids;
  c->rec_bytes = R;
  c->header_bytes = (int)((((uint64_t)P * 8 + kHeaderAlign - 1) / kHeaderAlign) * kHeaderAlign);
  c->count_slots = c->resident ? count_table_slots(c->N, R, c->k) : 0;
}

// ------------------------------------------------------------------------------------------------ lifecycle
extern "C" void kc_config_default(kc_config *cfg) {
  memset(cfg, 0, sizeof *cfg);
  cfg->abi_version = KC_ABI_VERSION;
  cfg->k = 21;
  cfg->m = 0;
  cfg->qual_offset = 33;
  cfg->qual_cutoff = 20;
  cfg->dmin_thres = 2;
  cfg->dyn_min_depth = 0.9;
  cfg->use_filter = 0;
  cfg->rank = 0;
  cfg->n_ranks = 1;
  cfg->device = -1;
  cfg->est_kmers = 0;
  cfg->sequencing_depth = 4;
  cfg->base_error_rate = 0.005;
}

extern "C" const char *kc_version(void) { return "kcount_b200 0.1 (sm_100a)"; }

extern "C" int kc_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

extern "C" const char *kc_last_error(const kc_ctx *c) { return c ? c->err.c_str() : g_create_error.c_str(); }

extern "C" int kc_create(const kc_config *cfg, kc_ctx **out) {
  kc_ctx *c = nullptr;
  if (!cfg || !out) return fail(c, KC_ERR_ARG, "kc_create: null argument");
  if (cfg->abi_version != KC_ABI_VERSION) return fail(c, KC_ERR_ARG, "kc_create: ABI version %u != %u", cfg->abi_version, KC_ABI_VERSION);
  if (cfg->k < 2 || cfg->k > KC_MAX_K || cfg->k % 32 == 0)
    return fail(c, KC_ERR_ARG, "kc_create: k=%u unsupported (2..%d, not a multiple of 32)", cfg->k, KC_MAX_K);
  if (cfg->n_ranks < 1 || cfg->rank < 0 || cfg->rank >= cfg->n_ranks) return fail(c, KC_ERR_ARG, "kc_create: bad rank %d/%d", cfg->rank, cfg->n_ranks);
  int m = cfg->m ? (int)cfg->m : reference_minimizer_len(cfg->k);
  if (m > (int)cfg->k) m = cfg->k;
  if (m < 1 || m > 28) return fail(c, KC_ERR_ARG, "kc_create: minimizer length %d unsupported (1..28)", m);
  const bool resident = !cfg->use_filter && !(cfg->flags & KC_FLAG_GLOBAL_TABLE);
  if (!resident && (cfg->local_bins & (cfg->local_bins - 1))) return fail(c, KC_ERR_ARG, "kc_create: local_bins must be a power of two in HBM-table mode");
  if (cfg->local_bins > (1u << 26) || (!resident && (uint64_t)std::max(cfg->local_bins, 1u) * cfg->n_ranks > (1u << 25)))
    return fail(c, KC_ERR_ARG, "kc_create: too many bins");
  if (cfg->record_bytes && cfg->record_bytes != 16 && cfg->record_bytes != 32 && cfg->record_bytes != 64)
    return fail(c, KC_ERR_ARG, "kc_create: record_bytes must be 16, 32 or 64");
  if (record_geom(64, cfg->k).max_kmers < 1) return fail(c, KC_ERR_ARG, "kc_create: k too large for a 64-byte record");
  if (cfg->n_ranks > kCountMaxSrc && resident) return fail(c, KC_ERR_ARG, "kc_create: more than %d ranks", kCountMaxSrc);

  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(c, KC_ERR_CUDA, "kc_create: no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
  c = new kc_ctx();
  c->cfg = *cfg;
  c->k = cfg->k;
  c->m = m;
  c->N = (cfg->k + 31) / 32;
  c->G = cfg->n_ranks;
  c->rank = cfg->rank;
  c->resident = resident;
  memset(&c->stats, 0, sizeof c->stats);
  memset(&c->h_ctr, 0, sizeof c->h_ctr);
  int dev = cfg->device;
  if (dev < 0) cudaGetDevice(&dev);
  c->device = dev;
  auto bail = [&](cudaError_t ee, const char *what) {
    int code = fail(nullptr, KC_ERR_CUDA, "kc_create: %s: %s", what, cudaGetErrorString(ee));
    delete c;
    return code;
  };
  if ((e = cudaSetDevice(dev)) != cudaSuccess) return bail(e, "cudaSetDevice");
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, dev)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
  c->sm_count = prop.multiProcessorCount;
  // every handle is checked: a failure here must be a kc_status now, not a null-handle crash in a later phase
#define CREATE_CK(call)                                 \
  do {                                                  \
    if ((e = (call)) != cudaSuccess) {                  \
      int code_ = fail(nullptr, e == cudaErrorMemoryAllocation ? KC_ERR_NOMEM : KC_ERR_CUDA, "kc_create: %s: %s", #call, cudaGetErrorString(e)); \
      kc_destroy(c);                                    \
      return code_;                                     \
    }                                                   \
  } while (0)
  CREATE_CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  for (cudaEvent_t *ev : {&c->ev0, &c->ev1, &c->ev_pk0, &c->ev_pk1, &c->ev_ik0, &c->ev_ik1, &c->ev_clr}) CREATE_CK(cudaEventCreate(ev));
  for (auto &e2 : c->ev_user) CREATE_CK(cudaEventCreate(&e2));
  CREATE_CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  CREATE_CK(cudaStreamCreateWithFlags(&c->in_stream, cudaStreamNonBlocking));
  CREATE_CK(cudaStreamCreateWithFlags(&c->parse_stream2, cudaStreamNonBlocking));
  {
    // the refine kernel of the SM-partitioned pipeline must become resident before the count kernel that waits for it
    int lo = 0, hi = 0;
    CREATE_CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CREATE_CK(cudaStreamCreateWithPriority(&c->refine_stream, cudaStreamNonBlocking, hi));
  }
  CREATE_CK(cudaEventCreate(&c->ev_fork));
  CREATE_CK(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  for (auto &e2 : c->ev_slice) CREATE_CK(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
  for (auto *arr : {c->ev_rs0, c->ev_rs1, c->ev_cs0, c->ev_cs1})
    for (int i = 0; i < kSinkSlices; i++) CREATE_CK(cudaEventCreate(&arr[i]));
  for (int q = 0; q < kXferStreams; q++) {
    CREATE_CK(cudaStreamCreateWithFlags(&c->xfer_stream[q], cudaStreamNonBlocking));
    CREATE_CK(cudaEventCreateWithFlags(&c->ev_x[q], cudaEventDisableTiming));
    CREATE_CK(cudaEventCreate(&c->ev_xbeg[q]));
    CREATE_CK(cudaEventCreate(&c->ev_xend[q]));
  }
  CREATE_CK(cudaHostAlloc((void **)&c->h_slice_n, kSinkSlices * sizeof(unsigned long long), cudaHostAllocDefault));
  CREATE_CK(cudaHostAlloc((void **)&c->h_pinned, kPinnedWords * sizeof(unsigned long long), cudaHostAllocDefault));
  CREATE_CK(dev_alloc(&c->d_ctr, 1));
  CREATE_CK(dev_alloc(&c->d_scratch, kScratchWords));
  CREATE_CK(dev_alloc(&c->d_slice_done, kSlices + 1));
  CREATE_CK(cudaMemsetAsync(c->d_ctr, 0, sizeof(DevCounters), c->stream));
#undef CREATE_CK
  configure_bins(c);
  *out = c;
  return KC_OK;
}

static void free_results(ResultSet &r) {
  dev_free(r.keys);
  dev_free(r.counts);
  dev_free(r.ext8);
  dev_free(r.left);
  dev_free(r.right);
  dev_free(r.index);
  r.index_slots = 0;
  r.n = r.cap = 0;
  r.sorted = r.indexed = false;
}

static void free_chunk(ReadChunk &ch) {
  if (!ch.owner) return;
  dev_free(ch.d_bytes);
  dev_free(ch.d_offsets);
  dev_free(ch.d_lens);
}

extern "C" int kc_clear_reads(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  cudaSetDevice(c->device);
  if (c->in_stream) cudaStreamSynchronize(c->in_stream);
  if (c->stream) cudaStreamSynchronize(c->stream);
  for (auto &ch : c->chunks) {
    if (!ch.owner) continue;
    // keep the largest allocation for the next kc_add_reads (cudaMalloc / cudaFree of gigabytes are not free)
    if (ch.cap_bytes > c->spare.cap_bytes && ch.cap_reads >= c->spare.cap_reads) {
      free_chunk(c->spare);
      c->spare = ch;
      c->spare.owner = true;
    } else {
      free_chunk(ch);
    }
  }
  c->chunks.clear();
  c->ev_in_used = 0;
  c->n_reads = c->n_bases = 0;
  c->caps_valid = false;
  c->phase = PH_EMPTY;
  return KC_OK;
}

extern "C" void kc_destroy(kc_ctx *c) {
  if (!c) return;
  kc_clear_reads(c);
  for (auto &m : c->ipc_mapped)
    if (m.second) cudaIpcCloseMemHandle(m.second);
  for (auto p : c->retired_bins) cudaFree(p);
  free_chunk(c->spare);
  for (auto &e : c->ev_in_pool) cudaEventDestroy(e);
  if (c->in_stream) cudaStreamDestroy(c->in_stream);
  if (c->parse_stream2) cudaStreamDestroy(c->parse_stream2);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  dev_free(c->d_bins);
  dev_free(c->d_bin_base);
  dev_free(c->d_bin_first);
  dev_free(c->d_bin_cap);
  dev_free(c->d_cursor);
  dev_free(c->d_hdr_ptrs);
  dev_free(c->d_owner_vals);
  dev_free(c->d_block_sums);
  dev_free(c->d_recv);
  dev_free(c->d_src_ptr);
  dev_free(c->d_count_src);
  dev_free(c->d_fine);
  dev_free(c->d_coarse_base);
  dev_free(c->d_segments);
  dev_free(c->d_seg_tile_start);
  dev_free(c->d_regions);
  dev_free(c->d_total_records);
  dev_free(c->d_A);
  dev_free(c->d_B);
  dev_free(c->d_C);
  dev_free(c->d_filter);
  dev_free(c->heavy.slots);
  dev_free(c->heavy.side);
  dev_free(c->heavy.n);
  dev_free(c->d_ctr);
  free_results(c->res_final);
  free_results(c->res_ext);
  for (cudaEvent_t e : {c->ev0, c->ev1, c->ev_pk0, c->ev_pk1, c->ev_ik0, c->ev_ik1, c->ev_clr})
    if (e) cudaEventDestroy(e);
  for (auto &e : c->ev_user)
    if (e) cudaEventDestroy(e);
  for (auto &e : c->ev_slice)
    if (e) cudaEventDestroy(e);
  for (auto *arr : {c->ev_rs0, c->ev_rs1, c->ev_cs0, c->ev_cs1})
    for (int i = 0; i < kSinkSlices; i++)
      if (arr[i]) cudaEventDestroy(arr[i]);
  for (auto *arr : {c->ev_x, c->ev_xbeg, c->ev_xend})
    for (int i = 0; i < kXferStreams; i++)
      if (arr[i]) cudaEventDestroy(arr[i]);
  for (auto &xs : c->xfer_stream)
    if (xs) cudaStreamDestroy(xs);
  for (int i = 0; i < 2; i++) {
    if (c->h_stage[i]) cudaFreeHost(c->h_stage[i]);
    if (c->ev_stage[i]) cudaEventDestroy(c->ev_stage[i]);
  }
  if (c->h_slice_n) cudaFreeHost(c->h_slice_n);
  if (c->h_pinned) cudaFreeHost(c->h_pinned);
  dev_free(c->d_scratch);
  dev_free(c->d_slice_done);
  dev_free(c->d_lq);
  dev_free(c->d_lc);
  dev_free(c->d_ll);
  dev_free(c->d_lr);
  if (c->refine_stream) cudaStreamDestroy(c->refine_stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

extern "C" int kc_sync(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->in_stream));  // the input copies of kc_add_reads: the caller may reuse its buffers after this
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaStreamSynchronize(c->refine_stream));
  CK(cudaStreamSynchronize(c->copy_stream));
  return KC_OK;
}

extern "C" int kc_set_result_sink(kc_ctx *c, uint64_t *keys, uint16_t *counts, char *left, char *right, uint64_t capacity) {
  if (!c) return KC_ERR_ARG;
  if (keys && (!counts || !left || !right)) return fail(c, KC_ERR_ARG, "kc_set_result_sink: all four arrays are needed");
  if (keys && !c->resident) return fail(c, KC_ERR_ARG, "kc_set_result_sink: needs the bin-resident count path (no filter, no global table)");
  c->sink_keys = keys;
  c->sink_counts = keys ? counts : nullptr;
  c->sink_left = keys ? left : nullptr;
  c->sink_right = keys ? right : nullptr;
  c->sink_cap = keys ? capacity : 0;
  return KC_OK;
}

int kc::ctx_device(const kc_ctx *c) { return c ? c->device : -1; }
extern "C" uint32_t kc_key_words(const kc_ctx *c) { return c ? c->N : 0; }
extern "C" uint32_t kc_minimizer_len(const kc_ctx *c) { return c ? c->m : 0; }

extern "C" int kc_host_alloc(void **ptr, size_t bytes) {
  return cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? KC_OK : KC_ERR_NOMEM;
}
extern "C" int kc_host_free(void *ptr) { return cudaFreeHost(ptr) == cudaSuccess ? KC_OK : KC_ERR_CUDA; }

// ------------------------------------------------------------------------------------------------ input
extern "C" int kc_add_reads(kc_ctx *c, const uint8_t *bytes, uint64_t n_bytes, const uint64_t *offsets, const uint16_t *lens,
                            uint64_t n_reads) {
  if (!c || (n_reads && (!bytes || !offsets || !lens))) return fail(c, KC_ERR_ARG, "kc_add_reads: null argument");
  if (n_reads == 0) return KC_OK;
  CK(cudaSetDevice(c->device));
  ReadChunk ch;
  if (c->spare.d_bytes && c->spare.cap_bytes >= n_bytes + 16 && c->spare.cap_reads >= n_reads) {
    ch = c->spare;
    c->spare = ReadChunk();
  } else {
    ch.cap_bytes = n_bytes + 16;
    ch.cap_reads = n_reads;
    CK(dev_alloc(&ch.d_bytes, ch.cap_bytes));
    CK(dev_alloc(&ch.d_offsets, n_reads));
    CK(dev_alloc(&ch.d_lens, n_reads));
  }
  ch.owner = true;
  // cudaMemcpyDefault: host (pageable or pinned) and device sources alike
  CK(cudaMemcpyAsync(ch.d_offsets, offsets, n_reads * sizeof(uint64_t), cudaMemcpyDefault, c->in_stream));
  CK(cudaMemcpyAsync(ch.d_lens, lens, n_reads * sizeof(uint16_t), cudaMemcpyDefault, c->in_stream));
  auto next_event = [&](cudaEvent_t *ev) -> cudaError_t {
    if (c->ev_in_used == c->ev_in_pool.size()) {
      cudaEvent_t e;
      cudaError_t rc = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
      if (rc != cudaSuccess) return rc;
      c->ev_in_pool.push_back(e);
    }
    *ev = c->ev_in_pool[c->ev_in_used++];
    return cudaSuccess;
  };
  // A host source is copied in sub-chunks of whole reads, each with its own event. That needs the reads to lie in the
  // byte buffer in order (offsets ascending, no overlap: the PackedReads layout); the check runs just ahead of the copies
  // and everything after the first violation becomes one chunk.
  cudaPointerAttributes at;
  bool host_src = true;
  if (cudaPointerGetAttributes(&at, bytes) == cudaSuccess && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged)) host_src = false;
  cudaGetLastError();
  cudaPointerAttributes at2;
  if (cudaPointerGetAttributes(&at2, offsets) == cudaSuccess && (at2.type == cudaMemoryTypeDevice || at2.type == cudaMemoryTypeManaged)) host_src = false;
  cudaGetLastError();
  uint64_t sub_bytes = 128ull << 20;
  if (const char *e = getenv("KC_INPUT_CHUNK_BYTES")) sub_bytes = std::max<uint64_t>(strtoull(e, nullptr, 10), 4096);  // (tests)
  uint64_t n_sub = host_src ? std::min<uint64_t>(std::max<uint64_t>(n_bytes / sub_bytes, 1), 32) : 1;
  if (n_reads < 64 * n_sub) n_sub = 1;
  uint64_t r0 = 0, b0 = 0;
  for (uint64_t j = 0; j < n_sub && r0 < n_reads; j++) {
    uint64_t r1 = n_reads * (j + 1) / n_sub;
    bool ordered = true;
    if (n_sub > 1) {
      if (offsets[r0] < b0) ordered = false;
      for (uint64_t r = r0; ordered && r + 1 < n_reads && r < r1; r++) ordered = offsets[r + 1] >= offsets[r] + lens[r];
    }
    if (!ordered || j + 1 == n_sub) r1 = n_reads;
    const uint64_t b1 = r1 == n_reads ? n_bytes : offsets[r1];
    const uint64_t lo = ordered ? b0 : 0;  // out of order: copy everything (again) before these reads are touched
    CK(cudaMemcpyAsync(ch.d_bytes + lo, bytes + lo, (ordered ? b1 : n_bytes) - lo, cudaMemcpyDefault, c->in_stream));
    ReadChunk sub = ch;
    sub.owner = j == 0;
    sub.d_offsets = ch.d_offsets + r0;
    sub.d_lens = ch.d_lens + r0;
    sub.n_reads = r1 - r0;
    sub.n_bytes = b1 > b0 ? b1 - b0 : 0;
    CK(next_event(&sub.ready));
    CK(cudaEventRecord(sub.ready, c->in_stream));
    c->chunks.push_back(sub);
    r0 = r1;
    b0 = b1;
  }
  c->n_reads += n_reads;
  c->n_bases += n_bytes;
  c->caps_valid = false;
  c->phase = PH_EMPTY;
  return KC_OK;
}

// ---- the reference's real in-memory reads ---------------------------------------------------
// PackedReads keeps a vector<PackedRead> of 16-byte headers (src/packed_reads.hpp:68-79: one 64-bit word of bit fields --
// is_allocated : 1, read_id : 47, read_len : 16, from the low bit up -- and a pointer to the read's bytes), the bytes
// themselves bump-allocated in 4 MB blocks (src/packed_reads.hpp:136, src/linear_allocator_pool.cpp). count_kmers walks them
// with get_next_read and unpacks every read into two strings (src/packed_reads.cpp:196-208, 269-276). Here host threads
// gather the bytes of whole pieces of the container into pinned staging (two buffers: one is filled while the other is
// copied), the lengths go up as they are and the offsets are made by a scan on the device.
static const uint64_t kStageBytes = 64ull << 20;
static const uint64_t kGatherBlock = 4096;  // reads per unit of host work

struct PackedHeader {
  uint64_t bits;
  const uint8_t *bytes;
};

template <typename F>
static void parallel_blocks(uint64_t n_blocks, int n_threads, F fn) {
  if (n_threads <= 1 || n_blocks <= 1) {
    for (uint64_t b = 0; b < n_blocks; b++) fn(b);
    return;
  }
  std::atomic<uint64_t> next{0};
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; t++)
    th.emplace_back([&]() {
      for (uint64_t b = next.fetch_add(1); b < n_blocks; b = next.fetch_add(1)) fn(b);
    });
  for (auto &x : th) x.join();
}

extern "C" int kc_add_packed_reads(kc_ctx *c, const void *headers, uint64_t n_reads, uint32_t header_stride) {
  if (!c || (n_reads && !headers)) return fail(c, KC_ERR_ARG, "kc_add_packed_reads: null argument");
  if (header_stride == 0) header_stride = sizeof(PackedHeader);
  if (header_stride < sizeof(PackedHeader)) return fail(c, KC_ERR_ARG, "kc_add_packed_reads: header stride %u < 16", header_stride);
  if (n_reads == 0) return KC_OK;
  CK(cudaSetDevice(c->device));
  const auto t_begin = std::chrono::high_resolution_clock::now();
  auto hdr = [&](uint64_t i) { return reinterpret_cast<const PackedHeader *>(static_cast<const uint8_t *>(headers) + i * header_stride); };
  int n_threads = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
  if (const char *e = getenv("KC_HOST_THREADS")) n_threads = std::max(1, atoi(e));
  // pass 0: lengths and bytes per block of reads
  const uint64_t n_blocks = (n_reads + kGatherBlock - 1) / kGatherBlock;
  std::vector<uint16_t> lens(n_reads);
  std::vector<uint64_t> block_base(n_blocks + 1, 0);
  parallel_blocks(n_blocks, n_threads, [&](uint64_t b) {
    const uint64_t r1 = std::min(n_reads, (b + 1) * kGatherBlock);
    uint64_t sum = 0;
    for (uint64_t r = b * kGatherBlock; r < r1; r++) {
      const uint16_t len = (uint16_t)(hdr(r)->bits >> 48);
      lens[r] = len;
      sum += len;
    }
    block_base[b + 1] = sum;
  });
  for (uint64_t b = 0; b < n_blocks; b++) block_base[b + 1] += block_base[b];
  const uint64_t n_bytes = block_base[n_blocks];
  ReadChunk ch;
  if (c->spare.d_bytes && c->spare.cap_bytes >= n_bytes + 16 && c->spare.cap_reads >= n_reads) {
    ch = c->spare;
    c->spare = ReadChunk();
  } else {
    ch.cap_bytes = n_bytes + 16;
    ch.cap_reads = n_reads;
    CK(dev_alloc(&ch.d_bytes, ch.cap_bytes));
    CK(dev_alloc(&ch.d_offsets, n_reads));
    CK(dev_alloc(&ch.d_lens, n_reads));
  }
  ch.owner = true;
  for (int i = 0; i < 2; i++) {
    if (!c->h_stage[i]) CK(cudaHostAlloc((void **)&c->h_stage[i], kStageBytes, cudaHostAllocDefault));
    if (!c->ev_stage[i]) CK(cudaEventCreateWithFlags(&c->ev_stage[i], cudaEventDisableTiming));
  }
  CK(cudaMemcpyAsync(ch.d_lens, lens.data(), n_reads * sizeof(uint16_t), cudaMemcpyHostToDevice, c->in_stream));
  CK(scan_lens_to_offsets(ch.d_lens, ch.d_offsets, n_reads, c->in_stream, &c->launches));
  auto next_event = [&](cudaEvent_t *ev) -> cudaError_t {
    if (c->ev_in_used == c->ev_in_pool.size()) {
      cudaEvent_t e;
      cudaError_t rc = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
      if (rc != cudaSuccess) return rc;
      c->ev_in_pool.push_back(e);
    }
    *ev = c->ev_in_pool[c->ev_in_used++];
    return cudaSuccess;
  };
  // pieces of whole blocks, at most one staging buffer each
  int which = 0;
  bool first = true;
  for (uint64_t b0 = 0; b0 < n_blocks;) {
    uint64_t b1 = b0 + 1;
    while (b1 < n_blocks && block_base[b1 + 1] - block_base[b0] <= kStageBytes) b1++;
    const uint64_t piece_bytes = block_base[b1] - block_base[b0];
    if (piece_bytes > kStageBytes) return fail(c, KC_ERR_ARG, "kc_add_packed_reads: %llu reads of more than %llu bytes", (unsigned long long)kGatherBlock, (unsigned long long)kStageBytes);
    if (c->stage_busy[which]) CK(cudaEventSynchronize(c->ev_stage[which]));
    uint8_t *stage = c->h_stage[which];
    parallel_blocks(b1 - b0, n_threads, [&](uint64_t bi) {
      const uint64_t b = b0 + bi, r1 = std::min(n_reads, (b + 1) * kGatherBlock);
      uint8_t *dst = stage + (block_base[b] - block_base[b0]);
      for (uint64_t r = b * kGatherBlock; r < r1; r++) {
        const PackedHeader *h = hdr(r);
        const uint16_t len = lens[r];
        if (len) memcpy(dst, h->bytes, len);
        dst += len;
      }
    });
    if (piece_bytes) CK(cudaMemcpyAsync(ch.d_bytes + block_base[b0], stage, piece_bytes, cudaMemcpyHostToDevice, c->in_stream));
    CK(cudaEventRecord(c->ev_stage[which], c->in_stream));
    c->stage_busy[which] = true;
    ReadChunk sub = ch;
    sub.owner = first;
    const uint64_t r0 = b0 * kGatherBlock, r1 = std::min(n_reads, b1 * kGatherBlock);
    sub.d_offsets = ch.d_offsets + r0;
    sub.d_lens = ch.d_lens + r0;
    sub.n_reads = r1 - r0;
    sub.n_bytes = piece_bytes;
    CK(next_event(&sub.ready));
    CK(cudaEventRecord(sub.ready, c->in_stream));
    c->chunks.push_back(sub);
    first = false;
    which ^= 1;
    b0 = b1;
  }
  // `lens` is a local: its copy has to be out before it goes away (the staging buffers are the context's)
  CK(cudaStreamSynchronize(c->in_stream));
  c->stage_busy[0] = c->stage_busy[1] = false;
  c->n_reads += n_reads;
  c->n_bases += n_bytes;
  c->caps_valid = false;
  c->phase = PH_EMPTY;
  c->packed_gather_ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t_begin).count();
  return KC_OK;
}

// ------------------------------------------------------------------------------------------------ parse
static ParseParams parse_params(kc_ctx *c, const ReadChunk &ch, uint32_t stride, int count_only) {
  ParseParams p;
  memset(&p, 0, sizeof p);
  p.bytes = ch.d_bytes;
  p.offsets = ch.d_offsets;
  p.lens = ch.d_lens;
  p.n_reads = ch.n_reads;
  p.read_stride = stride;
  p.count_only = count_only;
  p.k = c->k;
  p.m = c->m;
  p.qual_cutoff = c->cfg.qual_cutoff;
  p.rec_bytes = c->rec_bytes;
  p.n_ids = c->n_ids;
  p.local_bins = (uint32_t)c->P;
  p.fine_shift = c->fine_shift;
  p.div_ranks = make_fastdiv(c->G);
  p.bins = c->d_bins;
  p.bin_base = c->d_bin_base;
  p.bin_cap = c->d_bin_cap;
  p.bin_cursor = c->d_cursor;
  p.ctr = c->d_ctr;
  return p;
}

// Lay the bins out on the device: per owner [header block][bin 0 records][bin 1 records]..., capacities from the
// cursors of a counting pass (stride > 1: a strided sample, scaled with a noise margin; stride == 1: exact). Only the
// per-owner region offsets / sizes (n_ranks values) come back to the host.
static int layout_bins(kc_ctx *c, double scale) {
  const uint32_t n = c->n_ids;
  LayoutParams lp;
  memset(&lp, 0, sizeof lp);
  lp.n_ranks = (uint32_t)c->G;
  lp.local_bins = (uint32_t)c->P;
  lp.scale = scale;
  lp.rec_bytes = c->rec_bytes;
  lp.header_bytes = c->header_bytes;
  lp.pad = 64;
  lp.bin_cursor = c->d_cursor;
  lp.sampled_total = &c->d_ctr->parse_records;
  lp.bin_cap = c->d_bin_cap;
  lp.bin_first = c->d_bin_first;
  lp.bin_base = c->d_bin_base;
  lp.owner_records = c->d_owner_vals;
  lp.region_off = c->d_owner_vals + kCountMaxSrc;
  lp.region_bytes = c->d_owner_vals + 2 * kCountMaxSrc + 1;
  lp.block_sums = c->d_block_sums;
  lp.err_too_large = &c->d_ctr->err_bins_full;  // (checked right below, reset before the real parse)
  (void)n;
  CK(cudaMemsetAsync(&c->d_ctr->err_bins_full, 0, sizeof(unsigned int), c->stream));
  CK(launch_layout(lp, c->stream, &c->launches));
  std::vector<unsigned long long> h(3 * kCountMaxSrc + 2);
  unsigned int too_large = 0;
  CK(cudaMemcpyAsync(h.data(), c->d_owner_vals, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(&too_large, &c->d_ctr->err_bins_full, sizeof too_large, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (too_large) return fail(c, KC_ERR_ARG, "supermer region too large for 32-bit record indices; use more ranks");
  c->h_region_off.assign(c->G, 0);
  c->h_region_bytes.assign(c->G, 0);
  for (int o = 0; o < c->G; o++) {
    c->h_region_off[o] = h[kCountMaxSrc + o];
    c->h_region_bytes[o] = h[2 * kCountMaxSrc + 1 + o];
  }
  c->bins_bytes = h[kCountMaxSrc + c->G];
  if (c->bins_bytes > c->bins_alloc) {
    if (c->bins_exported && c->d_bins) {
      c->retired_bins.push_back(c->d_bins);  // peers may still hold it mapped: freed by kc_exchange_release
      c->d_bins = nullptr;
    }
    dev_free(c->d_bins);
    c->bins_exported = false;
    uint64_t want = c->bins_bytes + c->bins_bytes / 16;
    CK(dev_alloc(&c->d_bins, want));
    c->bins_alloc = want;
  }
  std::vector<uint8_t *> hdr(c->G);
  for (int o = 0; o < c->G; o++) hdr[o] = c->d_bins + c->h_region_off[o];
  CK(cudaMemcpyAsync(c->d_hdr_ptrs, hdr.data(), c->G * sizeof(uint8_t *), cudaMemcpyHostToDevice, c->stream));
  c->h_slice_off.clear();
  if (c->G > 1) {
    // where every pipeline slice of coarse bins starts inside each owner's region: the importers of the copy-mode
    // exchange cut their transfers there (the layout is by capacity, so this is known before a record is written)
    const int S = (uint32_t)c->P >= 64u * kSlices ? kSlices : 1;
    const size_t n = (size_t)c->G * (S + 1);
    unsigned long long *d_so = nullptr;
    CK(dev_alloc(&d_so, n));
    cudaError_t e = launch_slice_offsets(c->d_bin_first, c->d_bin_cap, (uint32_t)c->G, (uint32_t)c->P, S, c->header_bytes, c->rec_bytes, d_so,
                                         c->stream, &c->launches);
    c->h_slice_off.assign(n, 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(c->h_slice_off.data(), d_so, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_so);
    CK(e);
  }
  CK(cudaStreamSynchronize(c->stream));  // hdr is a local
  return KC_OK;
}

static int ensure_bin_arrays(kc_ctx *c) {
  const uint32_t n = c->n_ids;
  if (!c->d_bin_base) {
    CK(dev_alloc(&c->d_bin_base, n));
    CK(dev_alloc(&c->d_bin_first, n));
    CK(dev_alloc(&c->d_bin_cap, n));
    CK(dev_alloc(&c->d_cursor, n));
    CK(dev_alloc(&c->d_hdr_ptrs, c->G));
    dev_free(c->d_block_sums);
    CK(dev_alloc(&c->d_block_sums, (uint64_t)c->G * layout_blocks_per_owner((uint32_t)c->P)));
  }
  if (!c->d_owner_vals) CK(dev_alloc(&c->d_owner_vals, 4 * kCountMaxSrc + 4));
  return KC_OK;
}

// capacity estimate from a strided sample of the reads (count-only run of the parse kernel)
static int estimate_caps(kc_ctx *c) {
  const uint32_t n = c->n_ids;
  int rc = ensure_bin_arrays(c);
  if (rc) return rc;
  // sample: at least 262144 reads, and about 64 records per bin (a read yields ~2 (L-k-1) / (k-m+2) + 1 of them)
  uint64_t want_reads = 262144;
  if (n > 1) {
    const double avg_len = c->n_reads ? (double)c->n_bases / (double)c->n_reads : 0.0;
    const double per_read = std::max(1.0, 2.0 * std::max(avg_len - c->k - 1, 0.0) / (double)(c->k - c->m + 2) + 1.0);
    const double per_bin = n > 400000u ? std::max(24.0, 64.0 * 400000.0 / (double)n) : 64.0;
    want_reads = std::max<uint64_t>(want_reads, (uint64_t)(per_bin * n / per_read));
    // Long k: few records per read and many bins -- the rule above asks for half of the reads (k = 99: 10.9 M of 20 M, a sample
    // pass of 11 ms in front of a 21 ms parse). Never look at more than this fraction of them; the margin of layout_caps_kernel
    // widens with the thinner sample (more slack in the bin buffer, which is small for long k), and an overflow still falls
    // back to the exact layout.
    double max_frac = kSampleMaxFrac;
    if (const char *e = getenv("KC_SAMPLE_MAX_FRAC")) max_frac = std::min(std::max(atof(e), 0.01), 1.0);  // (experiments)
    want_reads = std::min<uint64_t>(want_reads, std::max<uint64_t>(262144, (uint64_t)(max_frac * (double)c->n_reads)));
  }
  // Which reads to look at: every stride-th read of all chunks -- or, when the input arrives in several chunks (a host
  // source is copied in sub-chunks), of the FIRST chunk only, so that the parse can start while the rest is still in
  // flight. Reads come off a sequencer in no genomic order, so a prefix is as good a sample as a stride; if it is not, a
  // bin overflows and kc_parse falls back to the exact layout.
  size_t n_sample_chunks = c->chunks.size();
  uint64_t pool = c->n_reads;
  if (c->chunks.size() > 1) {
    uint64_t have = 0;
    for (size_t i = 0; i < c->chunks.size(); i++) {
      have += c->chunks[i].n_reads;
      if (have >= 2 * want_reads) {
        n_sample_chunks = i + 1;
        pool = have;
        break;
      }
    }
  }
  uint32_t stride = (uint32_t)std::max<uint64_t>(1, pool / want_reads);
  CK(cudaMemsetAsync(c->d_cursor, 0, n * sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(&c->d_ctr->parse_counted, 0, 3 * sizeof(unsigned long long) + sizeof(unsigned int), c->stream));
  uint64_t looked_at = 0;
  for (size_t i = 0; i < n_sample_chunks; i++) {
    const ReadChunk &ch = c->chunks[i];
    if (ch.ready) CK(cudaStreamWaitEvent(c->stream, ch.ready, 0));
    ParseParams p = parse_params(c, ch, stride, 1);
    CK(launch_parse(p, c->sm_count, c->stream, &c->launches));
    looked_at += (ch.n_reads + stride - 1) / stride;
  }
  const double scale = looked_at && looked_at < c->n_reads ? (double)c->n_reads / (double)looked_at : 1.0;
  rc = layout_bins(c, scale);
  if (rc) return rc;
  c->caps_valid = true;
  return KC_OK;
}

static int run_parse(kc_ctx *c) {
  const uint32_t n = c->n_ids;
  unsigned long long *owner_used = c->d_owner_vals + 3 * kCountMaxSrc + 2;
  CK(cudaMemsetAsync(c->d_cursor, 0, n * sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(owner_used, 0, kCountMaxSrc * sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(&c->d_ctr->parse_counted, 0, 3 * sizeof(unsigned long long) + sizeof(unsigned int), c->stream));
  CK(cudaEventRecord(c->ev_pk0, c->stream));
  const bool two = c->chunks.size() > 1 && c->parse_stream2;
  if (two) {
    CK(cudaEventRecord(c->ev_fork, c->stream));
    CK(cudaStreamWaitEvent(c->parse_stream2, c->ev_fork, 0));
  }
  size_t ci = 0;
  for (auto &ch : c->chunks) {
    cudaStream_t ps = (two && (ci++ & 1)) ? c->parse_stream2 : c->stream;
    if (ch.ready) CK(cudaStreamWaitEvent(ps, ch.ready, 0));
    ParseParams p = parse_params(c, ch, 1, 0);
    CK(launch_parse(p, c->sm_count, ps, &c->launches));
  }
  if (two) {
    CK(cudaEventRecord(c->ev_join, c->parse_stream2));
    CK(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
  }
  CK(cudaEventRecord(c->ev_pk1, c->stream));
  CK(launch_write_headers(c->d_hdr_ptrs, c->d_bin_first, c->d_cursor, c->d_bin_cap, c->G, c->P, owner_used, c->stream, &c->launches));
  c->h_owner_used.assign(kCountMaxSrc, 0);
  CK(cudaMemcpyAsync(c->h_owner_used.data(), owner_used, kCountMaxSrc * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(&c->h_ctr, c->d_ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  return KC_OK;
}

extern "C" int kc_parse(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  if (c->chunks.empty()) return fail(c, KC_ERR_STATE, "kc_parse: no reads");
  CK(cudaSetDevice(c->device));
  if (!c->caps_valid) configure_bins(c);
  CK(cudaEventRecord(c->ev0, c->stream));
  if (!c->caps_valid) {
    int rc = estimate_caps(c);
    if (rc) return rc;
  }
  int rc = run_parse(c);
  if (rc) return rc;
  CK(cudaStreamSynchronize(c->stream));
  if (c->h_ctr.err_bins_full) {
    // the sample under-estimated a bin: the cursors now hold the exact demand, lay out again and redo
    if (getenv("KC_DEBUG")) {
      std::vector<unsigned long long> cur(c->n_ids);
      std::vector<uint64_t> cap(c->n_ids);
      cudaMemcpy(cur.data(), c->d_cursor, c->n_ids * 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(cap.data(), c->d_bin_cap, c->n_ids * 8, cudaMemcpyDeviceToHost);
      uint64_t n_over = 0, sum = 0, sumcap = 0;
      for (uint32_t i = 0; i < c->n_ids; i++) {
        sum += cur[i];
        sumcap += cap[i];
        if (cur[i] > cap[i] && n_over++ < 8) fprintf(stderr, "[kc] bin %u: %llu records, capacity %llu\n", i, cur[i], (unsigned long long)cap[i]);
      }
      fprintf(stderr, "[kc] %llu of %u bins overflowed; %llu records, capacity %llu\n", (unsigned long long)n_over, c->n_ids,
              (unsigned long long)sum, (unsigned long long)sumcap);
    }
    rc = layout_bins(c, 1.0);
    if (rc) return rc;
    rc = run_parse(c);
    if (rc) return rc;
    CK(cudaStreamSynchronize(c->stream));
    if (c->h_ctr.err_bins_full) return fail(c, KC_ERR_BINS_FULL, "kc_parse: supermer bins overflowed after exact re-layout");
  }
  CK(cudaEventRecord(c->ev1, c->stream));
  CK(cudaEventSynchronize(c->ev1));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->stats.ms_parse = ms;
  cudaEventElapsedTime(&ms, c->ev_pk0, c->ev_pk1);
  c->stats.ms_parse_kernel = ms;
  c->stats.n_reads = c->n_reads;
  c->stats.n_bases = c->n_bases;
  c->stats.raw_kmers = c->h_ctr.parse_raw;
  c->stats.counted_kmers = c->h_ctr.parse_counted;
  c->stats.supermers_out = c->h_ctr.parse_records;
  c->stats.supermer_bytes_out = c->h_ctr.parse_records * c->rec_bytes;
  c->phase = PH_PARSED;
  c->recv_off.clear();
  c->pull_ptr.clear();
  c->peer_slice_off.clear();
  return KC_OK;
}

// ------------------------------------------------------------------------------------------------ exchange
static uint64_t used_region_bytes(kc_ctx *c, int o) {
  // header + records up to the end of the last non-empty bin (the slack behind it need not travel)
  return c->header_bytes + (uint64_t)c->h_owner_used[o] * c->rec_bytes;
}

extern "C" int kc_exchange_send_sizes(kc_ctx *c, uint64_t *send_bytes) {
  if (!c || !send_bytes) return KC_ERR_ARG;
  if (c->phase < PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_send_sizes: call kc_parse first");
  for (int o = 0; o < c->G; o++) send_bytes[o] = (o == c->rank) ? 0 : used_region_bytes(c, o);
  return KC_OK;
}

extern "C" int kc_exchange_send_ptr(kc_ctx *c, int dst, void **dev_ptr) {
  if (!c || !dev_ptr || dst < 0 || dst >= c->G) return KC_ERR_ARG;
  // the send buffers stay valid until kc_reset / kc_add_reads: peers may still be pulling after this rank moved on
  if (c->phase < PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_send_ptr: call kc_parse first");
  *dev_ptr = c->d_bins + c->h_region_off[dst];
  return KC_OK;
}

extern "C" int kc_exchange_recv_buffer(kc_ctx *c, const uint64_t *recv_bytes, void **dev_ptr) {
  if (!c || !recv_bytes || !dev_ptr) return KC_ERR_ARG;
  if (c->phase != PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_recv_buffer: call kc_parse first");
  CK(cudaSetDevice(c->device));
  c->pull_ptr.clear();
  c->recv_off.assign(c->G, 0);
  c->recv_bytes.assign(recv_bytes, recv_bytes + c->G);
  uint64_t at = 0;
  for (int s = 0; s < c->G; s++) {
    c->recv_off[s] = at;
    if (s == c->rank) c->recv_bytes[s] = 0;
    at += (c->recv_bytes[s] + 255) & ~255ULL;
  }
  if (at > c->recv_alloc) {
    dev_free(c->d_recv);
    CK(dev_alloc(&c->d_recv, at + at / 16 + 256));
    c->recv_alloc = at + at / 16 + 256;
  }
  *dev_ptr = c->d_recv;
  return KC_OK;
}

extern "C" int kc_exchange_done(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  if (c->phase != PH_PARSED || ((int)c->recv_off.size() != c->G && (int)c->pull_ptr.size() != c->G))
    return fail(c, KC_ERR_STATE, "kc_exchange_done: no receive buffer and no imported regions");
  if ((int)c->pull_ptr.size() == c->G && c->xmode == KC_XCHG_COPY) {
    // copy mode: the imported regions land in d_recv (same layout as the all-to-all's receive buffer)
    CK(cudaSetDevice(c->device));
    c->recv_off.assign(c->G, 0);
    uint64_t at = 0;
    for (int s = 0; s < c->G; s++) {
      c->recv_off[s] = at;
      if (s != c->rank) at += (c->recv_bytes[s] + 255) & ~255ULL;
    }
    if (at > c->recv_alloc) {
      CK(cudaStreamSynchronize(c->stream));
      dev_free(c->d_recv);
      c->recv_alloc = 0;
      CK(dev_alloc(&c->d_recv, at + at / 16 + 256));
      c->recv_alloc = at + at / 16 + 256;
    }
  }
  c->phase = PH_RECV_READY;
  return KC_OK;
}

// ---- peer-memory exchange (CUDA IPC; the refine / count kernels read the peers' regions over NVLink)
extern "C" int kc_exchange_export(kc_ctx *c, uint8_t *handle, uint64_t *region_offsets, uint64_t *region_bytes) {
  if (!c || !handle || !region_offsets || !region_bytes) return KC_ERR_ARG;
  if (c->phase < PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_export: call kc_parse first");
  CK(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  CK(cudaIpcGetMemHandle(&h, c->d_bins));
  static_assert(sizeof h == KC_IPC_HANDLE_BYTES, "IPC handle size");
  memcpy(handle, &h, sizeof h);
  c->bins_exported = true;
  for (int o = 0; o < c->G; o++) {
    region_offsets[o] = c->h_region_off[o];
    region_bytes[o] = used_region_bytes(c, o);
  }
  return KC_OK;
}

extern "C" int kc_exchange_release(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  CK(cudaSetDevice(c->device));
  for (auto p : c->retired_bins) cudaFree(p);
  c->retired_bins.clear();
  return KC_OK;
}

extern "C" int kc_exchange_import(kc_ctx *c, int src, const uint8_t *handle, uint64_t region_offset, uint64_t region_bytes) {
  if (!c || !handle || src < 0 || src >= c->G || src == c->rank) return KC_ERR_ARG;
  if (c->phase != PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_import: call kc_parse first");
  CK(cudaSetDevice(c->device));
  if ((int)c->pull_ptr.size() != c->G) {
    c->pull_ptr.assign(c->G, nullptr);
    c->recv_bytes.assign(c->G, 0);
    c->recv_off.clear();
    c->peer_slice_off.clear();
  }
  const std::string key(reinterpret_cast<const char *>(handle), KC_IPC_HANDLE_BYTES);
  if ((int)c->ipc_mapped.size() != c->G) c->ipc_mapped.assign(c->G, std::make_pair(std::string(), (void *)nullptr));
  auto &slot = c->ipc_mapped[src];
  if (slot.second && slot.first != key) {  // the source re-allocated its send buffer: let go of the old one
    cudaIpcCloseMemHandle(slot.second);
    slot.second = nullptr;
  }
  void *base = slot.second;
  if (!base) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(c, KC_ERR_CUDA, "kc_exchange_import: cudaIpcOpenMemHandle from rank %d: %s", src, cudaGetErrorString(e));
    }
    slot.first = key;
    slot.second = base;
  }
  c->pull_ptr[src] = region_bytes ? static_cast<uint8_t *>(base) + region_offset : nullptr;
  c->recv_bytes[src] = region_bytes;
  return KC_OK;
}

extern "C" int kc_exchange_set_mode(kc_ctx *c, int mode) {
  if (!c || (mode != KC_XCHG_PULL && mode != KC_XCHG_COPY)) return KC_ERR_ARG;
  c->xmode = mode;
  return KC_OK;
}

extern "C" int kc_exchange_export_slices(kc_ctx *c, uint32_t *n_slices, uint64_t *offsets) {
  if (!c || !n_slices || !offsets) return KC_ERR_ARG;
  if (c->phase < PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_export_slices: call kc_parse first");
  const int S = (uint32_t)c->P >= 64u * kSlices ? kSlices : 1;
  *n_slices = (uint32_t)S;
  for (int o = 0; o < c->G; o++)
    for (int i = 0; i <= KC_MAX_SLICES; i++) {
      uint64_t v = used_region_bytes(c, o);
      if (i <= S && c->h_slice_off.size() == (size_t)c->G * (S + 1)) v = std::min<uint64_t>(c->h_slice_off[(size_t)o * (S + 1) + i], v);
      if (i == 0) v = std::min<uint64_t>(v, (uint64_t)c->header_bytes);
      offsets[(size_t)o * (KC_MAX_SLICES + 1) + i] = v;
    }
  return KC_OK;
}

extern "C" int kc_exchange_import_slices(kc_ctx *c, int src, uint32_t n_slices, const uint64_t *offsets) {
  if (!c || !offsets || src < 0 || src >= c->G || n_slices < 1 || n_slices > KC_MAX_SLICES) return KC_ERR_ARG;
  if (c->phase != PH_PARSED) return fail(c, KC_ERR_STATE, "kc_exchange_import_slices: call kc_parse first");
  if ((int)c->peer_slice_off.size() != c->G) c->peer_slice_off.assign(c->G, std::vector<uint64_t>());
  c->peer_slice_off[src].assign(offsets, offsets + n_slices + 1);
  return KC_OK;
}

// byte offset of source s inside the receive buffer (the caller's all-to-all needs the displacements)
extern "C" int kc_exchange_recv_offsets(kc_ctx *c, uint64_t *offsets) {
  if (!c || !offsets || (int)c->recv_off.size() != c->G) return KC_ERR_ARG;
  for (int s = 0; s < c->G; s++) offsets[s] = c->recv_off[s];
  return KC_OK;
}

// ------------------------------------------------------------------------------------------------ insert
static uint64_t slot_bytes(int N) { return 32 + (N >= 2 ? 4 : 0) + (N >= 3 ? 16 : 0); }

static uint64_t reference_table_slots(kc_ctx *c, uint64_t est_kmers) {
  // KmerDHT::KmerDHT + HashTableInserter::init: (kmers / depth + kmers * (1 - (1 - err)^k)) / 0.66
  // (src/kcount/kmer_dht.cpp:126-131, src/kcount/kcount_cpu.cpp:443-446); with the filter only the first term needs slots
  double depth = c->cfg.sequencing_depth > 0 ? c->cfg.sequencing_depth : 4;
  double kmer_err = 1.0 - std::pow(1.0 - c->cfg.base_error_rate, (double)c->k);
  double elems = (double)est_kmers / depth;
  if (!c->cfg.use_filter) elems += (double)est_kmers * kmer_err;
  return (uint64_t)(elems / 0.66) + 1024;
}

static int ensure_table(kc_ctx *c) {
  uint64_t est = c->cfg.est_kmers ? c->cfg.est_kmers : std::max<uint64_t>(c->h_ctr.parse_raw, 1024);
  uint64_t slots = c->cfg.table_slots ? c->cfg.table_slots : reference_table_slots(c, est);
  if (c->table_slots > slots) slots = c->table_slots;  // grown by a retry
  slots = std::max<uint64_t>(slots, 128ull * c->P + 1024);
  uint64_t buckets = 0;
  if (c->cfg.use_filter) {
    double kmer_err = 1.0 - std::pow(1.0 - c->cfg.base_error_rate, (double)c->k);
    double depth = c->cfg.sequencing_depth > 0 ? c->cfg.sequencing_depth : 4;
    double distinct = (double)est / depth + (double)est * kmer_err;
    buckets = (uint64_t)(distinct / 0.75 / 16.0) + 8ull * c->P + 64;
  }
  // memory budget
  size_t free_b = 0, total_b = 0;
  CK(cudaMemGetInfo(&free_b, &total_b));
  const uint64_t per_slot = slot_bytes(c->N);
  uint64_t have = (uint64_t)c->table_alloc_slots * per_slot + c->filter_alloc_buckets * 32;
  uint64_t budget = c->cfg.mem_budget_bytes ? c->cfg.mem_budget_bytes : (uint64_t)((free_b + have) * 0.8);
  if (buckets * 32 + slots * per_slot > budget) {
    if (c->cfg.table_slots) return fail(c, KC_ERR_NOMEM, "table of %llu slots does not fit the memory budget", (unsigned long long)slots);
    uint64_t fit = (budget - std::min(budget, buckets * 32)) / per_slot;
    if (fit < 128ull * c->P + 1024) return fail(c, KC_ERR_NOMEM, "no room for the k-mer table");
    slots = fit;
  }
  if (slots > c->table_alloc_slots) {
    dev_free(c->d_A);
    dev_free(c->d_B);
    dev_free(c->d_C);
    CK(dev_alloc(&c->d_A, slots));
    if (c->N >= 2) CK(dev_alloc(&c->d_C, slots + 4));
    if (c->N >= 3) CK(dev_alloc(&c->d_B, slots));
    c->table_alloc_slots = slots;
  }
  c->table_slots = slots;
  if (buckets > c->filter_alloc_buckets) {
    dev_free(c->d_filter);
    CK(dev_alloc(&c->d_filter, buckets * 8));
    c->filter_alloc_buckets = buckets;
  }
  c->filter_buckets = buckets;
  // heavy-key list: a key needs more than 65535 occurrences to get an entry
  uint64_t occ = std::max<uint64_t>(c->h_ctr.parse_counted * (uint64_t)c->G, 1);
  uint64_t want = occ / 65536 + 64;
  if (want > c->heavy.cap) {
    dev_free(c->heavy.slots);
    dev_free(c->heavy.side);
    CK(dev_alloc(&c->heavy.slots, want));
    CK(dev_alloc(&c->heavy.side, want * 8));
    c->heavy.cap = want;
  }
  if (!c->heavy.n) CK(dev_alloc(&c->heavy.n, 1));
  return KC_OK;
}

static InsertParams insert_params(kc_ctx *c) {
  InsertParams ip;
  memset(&ip, 0, sizeof ip);
  ip.k = c->k;
  ip.rec_bytes = c->rec_bytes;
  ip.n_segments = c->P * c->G;
  ip.segments = c->d_segments;
  ip.seg_tile_start = c->d_seg_tile_start;
  ip.regions = c->d_regions;
  ip.tile_records = insert_tile_records(c->rec_bytes);
  ip.A = c->d_A;
  ip.B = c->d_B;
  ip.C = c->d_C;
  ip.filter = c->d_filter;
  ip.heavy = c->heavy;
  ip.ctr = c->d_ctr;
  return ip;
}

static int alloc_results(kc_ctx *c, ResultSet &r, uint64_t cap, bool with_ext);

static int fine_header_bytes(const kc_ctx *c) {
  return (int)(((((uint64_t)c->P << c->fine_shift) * 8) + kHeaderAlign - 1) / kHeaderAlign * kHeaderAlign);
}

// ---- the resident insert pipeline ---------------------------------------------------------
// Everything between "the supermers of every source are reachable" and "the dense result is complete" is queued without
// a single host round trip, slice by slice over the coarse bins, on three kinds of streams:
//   xfer_stream[q]   (copy mode, G > 1) copy engines pull slice i of every peer's region over NVLink into d_recv
//   refine_stream    refine_kernel sorts slice i of the coarse bins (all sources) into fine bins (d_fine)
//   stream           count_kernel counts the fine bins of slice i; with a result sink the survivors of slice i go out on
//                    copy_stream while slice i + 1 is counted
// so the NVLink transfer (DMA, no SMs), the HBM-bound refine pass and the issue-bound count pass of neighbouring slices
// overlap. Sizes the host needs (d_fine, grid shapes) come from bounds it already has: the parse-side record counts and
// the region sizes the exchange reported.
static int n_pipeline_slices(const kc_ctx *c) {
  int S = (uint32_t)c->P >= 64u * kSlices ? kSlices : 1;
  // one rank and no result sink: nothing to overlap with, and every slice boundary costs the count kernel a tail
  if (c->G == 1 && !c->sink_keys) S = 1;
  if (const char *e = getenv("KC_SLICES")) {  // (experiments)
    const int v = atoi(e);
    if (v >= 1 && v <= kSlices && (uint32_t)c->P >= 64u * (uint32_t)v) S = v;
  }
  return S;
}
// Refine slice i + 1 beside count slice i (separate streams)? Measured on one GPU (C2): the two kernels cannot share an SM
// (the count CTAs hold all of its shared memory), so refine CTAs displace count CTAs and the pair takes 57.0 ms against
// 51.7 ms back to back. With peers the refine kernel spends its time waiting for NVLink, and hiding that behind the count
// pass wins.
// SM-partitioned pipeline (experimental, KC_REFINE_SMS=n): ONE refine launch on n SMs of its own (one 1024-thread CTA with
// the 208 KB stage each) pulls the peers' bins over NVLink for the whole insert phase while ONE count launch runs on all the
// other SMs; the count warps wait on per-slice counters the refine kernel bumps (device-side, no host, no launch seams), a
// gate kernel holds the count launch back until the refine CTAs are resident.
// Why it exists: a refine CTA needs a whole SM's shared memory and so do the count CTAs, so slice-wise launches on two
// streams never co-reside -- at 8 GPUs the two passes simply add up (refine + pull ~20 ms + count ~45 ms = 65 ms insert).
// Why it is off: measured at 8 GPUs (C2), 24 SMs: 132 ms per step, 36 SMs: 99.9 ms, off: 90.4 ms. One CTA moves one 113 KB
// coarse bin at a time with barriers between load, histogram and scatter: ~20 us per bin over NVLink = 5.7 GB/s per SM,
// so the refine pass on few SMs (66-98 ms) is slower than the count pass it was meant to hide behind. It needs a pull
// that keeps many bulk copies in flight per SM.
static int refine_partition_sms(const kc_ctx *c, bool to_sink) {
  int sms = 0;
  if (const char *e = getenv("KC_REFINE_SMS")) sms = atoi(e);
  if (sms < 0 || sms > c->sm_count / 2 || !c->fine_shift || to_sink) sms = 0;
  return sms;
}
static bool refine_overlaps_count(const kc_ctx *c) {
  if (const char *e = getenv("KC_REFINE_OVERLAP")) return atoi(e) != 0;  // (experiments)
  return c->G > 1 && c->xmode == KC_XCHG_PULL;
}

// upper bound of the records this owner will consume (exact for one rank)
static uint64_t record_bound(const kc_ctx *c) {
  if (c->G == 1) return c->h_ctr.parse_records;
  uint64_t n = c->h_owner_used.size() > (size_t)c->rank ? c->h_owner_used[c->rank] : 0;
  for (int s = 0; s < c->G; s++) {
    if (s == c->rank || (int)c->recv_bytes.size() != c->G) continue;
    if (c->recv_bytes[s] > (uint64_t)c->header_bytes) n += (c->recv_bytes[s] - c->header_bytes) / c->rec_bytes;
  }
  return n;
}

// copy mode: queue the peer -> d_recv copies of slice `i` (i < 0: the header blocks) on the transfer streams and make
// `waiter` wait for them
static int queue_peer_copies(kc_ctx *c, int i, int n_slices, cudaStream_t waiter) {
  int q = 0;
  bool used[kXferStreams] = {false, false, false, false};
  for (int s = 0; s < c->G; s++) {
    if (s == c->rank || !c->pull_ptr[s] || c->recv_bytes[s] == 0) continue;
    uint64_t lo, hi;
    const uint64_t used_bytes = c->recv_bytes[s];
    if (i < 0) {
      lo = 0;
      hi = std::min<uint64_t>(c->header_bytes, used_bytes);
    } else if ((int)c->peer_slice_off.size() == c->G && (int)c->peer_slice_off[s].size() == n_slices + 1) {
      lo = std::min<uint64_t>(std::max<uint64_t>(c->peer_slice_off[s][i], c->header_bytes), used_bytes);
      hi = i + 1 == n_slices ? used_bytes : std::min<uint64_t>(std::max<uint64_t>(c->peer_slice_off[s][i + 1], c->header_bytes), used_bytes);
    } else {  // no slice table from this peer: everything with the first slice
      lo = std::min<uint64_t>(c->header_bytes, used_bytes);
      hi = i == 0 ? used_bytes : lo;
    }
    if (hi <= lo) continue;
    cudaStream_t xs = c->xfer_stream[q];
    if (!c->xfer_started[q]) {
      CK(cudaEventRecord(c->ev_xbeg[q], xs));
      c->xfer_started[q] = true;
    }
    CK(cudaMemcpyAsync(c->d_recv + c->recv_off[s] + lo, c->pull_ptr[s] + lo, hi - lo, cudaMemcpyDeviceToDevice, xs));
    c->xfer_bytes += hi - lo;
    used[q] = true;
    q = (q + 1) % kXferStreams;
  }
  for (int x = 0; x < kXferStreams; x++) {
    if (!used[x]) continue;
    CK(cudaEventRecord(c->ev_x[x], c->xfer_stream[x]));
    CK(cudaStreamWaitEvent(waiter, c->ev_x[x], 0));
  }
  return KC_OK;
}

static CountParams count_params(kc_ctx *c) {
  const bool keep_ext = (c->cfg.flags & KC_FLAG_KEEP_EXT_COUNTS) != 0;
  CountParams cp;
  memset(&cp, 0, sizeof cp);
  cp.k = c->k;
  cp.rec_bytes = c->rec_bytes;
  cp.n_src = c->fine_shift ? 1 : c->G;
  cp.tile_records = insert_tile_records(c->rec_bytes);
  cp.n_bins = (uint32_t)c->P << c->fine_shift;
  cp.header_bytes = c->fine_shift ? fine_header_bytes(c) : c->header_bytes;
  cp.src_ptr = c->d_count_src;
  cp.src0 = c->fine_shift ? c->d_fine : nullptr;
  cp.n_slots = c->count_slots;
  cp.claim_limit = (uint32_t)(0.85 * c->count_slots);
  // pre-split threshold: records at which a bin's distinct k-mers are expected to pass the claim limit, from the k-mers per
  // record this parse produced and the distinct / occurrence ratio of the sizing rule (configure_bins). KC_PRESPLIT scales
  // it (0 = off).
  cp.presplit_rec = 0;
  {
    double f = kPresplitFactor;
    if (const char *e = getenv("KC_PRESPLIT")) f = atof(e);
    const double depth = c->cfg.sequencing_depth > 0 ? c->cfg.sequencing_depth : 4;
    const double kmer_err = 1.0 - std::pow(1.0 - c->cfg.base_error_rate, (double)c->k);
    // Only where the sizing rule itself says that most distinct k-mers are error k-mers (long k): there the records of a bin
    // predict its distinct k-mers. With short k a big bin is big because of coverage and the record count says little
    // (measured on C2, k = 33: 1.2 % of the bins overflow, nothing to gain, 0.2-1 ms lost to needless splits).
    const bool gate = getenv("KC_PRESPLIT") != nullptr || kmer_err * depth >= 1.0;
    if (gate && f > 0.0 && c->stats.supermers_out > 0 && c->stats.counted_kmers > 0) {
      const double r = 1.0 / depth + kmer_err;
      const double occ_per_rec = (double)c->stats.counted_kmers / (double)c->stats.supermers_out;
      const double lim = f * (double)cp.claim_limit / (r * occ_per_rec);
      cp.presplit_rec = (uint32_t)std::min(std::max(lim, 2.0), 1e9);
      cp.presplit_adapt = 1;
      if (const char *e = getenv("KC_PRESPLIT_ADAPT")) cp.presplit_adapt = atoi(e) != 0;  // (experiments)
    }
  }
  cp.dmin_thres = c->cfg.dmin_thres;
  cp.keep_ext = keep_ext;
  cp.one_minus_dyn = 1.0 - c->cfg.dyn_min_depth;  // evaluated on the host exactly as the reference's constant expression
  cp.ctr = c->d_ctr;
  cp.out_keys = c->res_final.keys;
  cp.out_counts = c->res_final.counts;
  cp.out_left = c->res_final.left;
  cp.out_right = c->res_final.right;
  cp.cap_final = c->res_final.cap;
  cp.ext_keys = c->res_ext.keys;
  cp.ext_counts = c->res_ext.counts;
  cp.ext_ext8 = c->res_ext.ext8;
  cp.ext_left = c->res_ext.left;
  cp.ext_right = c->res_ext.right;
  cp.cap_ext = keep_ext ? c->res_ext.cap : 0;
  return cp;
}

// ---- bin-resident counting: refine + count (insert + purge + extension resolution + compaction), pipelined by slices.
// `redo`: the count pass alone is repeated over the already refined bins (result arrays were too small).
static int launch_count_pipeline(kc_ctx *c, bool redo) {
  const int S = n_pipeline_slices(c);
  const bool refine = c->fine_shift != 0;
  const bool copy = !redo && c->G > 1 && c->xmode == KC_XCHG_COPY && (int)c->pull_ptr.size() == c->G;
  const bool dedup = !(c->cfg.flags & KC_FLAG_NO_WARP_DEDUP);
  CountParams cp = count_params(c);
  CK(cudaMemsetAsync(c->d_ctr, 0, offsetof(DevCounters, parse_counted), c->stream));
  CK(cudaEventRecord(c->ev_clr, c->stream));
  CK(cudaEventRecord(c->ev_ik0, c->stream));
  c->sink_streamed = false;
  c->sink_done = 0;
  c->n_count_slices = c->n_refine_slices = 0;
  c->xfer_bytes = 0;
  for (auto &x : c->xfer_started) x = false;
  const bool to_sink = !redo && c->sink_keys && S > 1;
  cudaStream_t rs = refine_overlaps_count(c) ? c->refine_stream : c->stream;

  RefineParams rp;
  memset(&rp, 0, sizeof rp);
  uint64_t avg_records = 0;
  if (!redo) {
    // the side streams start where the main stream is now
    CK(cudaEventRecord(c->ev_fork, c->stream));
    CK(cudaStreamWaitEvent(c->refine_stream, c->ev_fork, 0));
    if (copy)
      for (int q = 0; q < kXferStreams; q++) CK(cudaStreamWaitEvent(c->xfer_stream[q], c->ev_fork, 0));
  }
  const int part_sms = (refine && !redo) ? refine_partition_sms(c, to_sink) : 0;
  c->partitioned = part_sms > 0;
  if (part_sms > 0) {
    CK(cudaMemsetAsync(c->d_slice_done, 0, (kSlices + 1) * sizeof(unsigned int), c->stream));
    CK(cudaEventRecord(c->ev_fork, c->stream));
    CK(cudaStreamWaitEvent(c->refine_stream, c->ev_fork, 0));
    rp.n_src = c->G;
    rp.rec_bytes = c->rec_bytes;
    rp.n_coarse = (uint32_t)c->P;
    rp.fine_shift = c->fine_shift;
    rp.src_header_bytes = c->header_bytes;
    rp.src_ptr = c->d_count_src;
    rp.dst_header_bytes = fine_header_bytes(c);
    rp.coarse_base = c->d_coarse_base;
    rp.total = c->d_coarse_base + c->P + 1;
    rp.err_too_large = &c->d_ctr->err_bins_full;
    rp.dst = c->d_fine;
    rp.n_slices = S;
    rp.slice_done = c->d_slice_done;
    rp.grid_limit = part_sms;
    cp.wait_slices = S;
    cp.slice_done = c->d_slice_done;
    cp.err_wait = &c->d_ctr->err_wait;
    for (int i = 0; i <= S; i++) {
      rp.slice_coarse_edge[i] = slice_edge((uint32_t)c->P, i, S);
      cp.slice_fine_edge[i] = rp.slice_coarse_edge[i] << c->fine_shift;
      if (i) cp.slice_need[i - 1] = rp.slice_coarse_edge[i] - rp.slice_coarse_edge[i - 1];
    }
    CK(launch_refine_totals(rp, c->refine_stream, &c->launches));
    CK(cudaEventRecord(c->ev_rs0[0], c->refine_stream));
    CK(launch_refine(rp, record_bound(c) / std::max<uint64_t>(rp.n_coarse, 1), c->sm_count, c->refine_stream, &c->launches));
    CK(cudaEventRecord(c->ev_rs1[0], c->refine_stream));
    c->n_refine_slices = 1;
    cp.bin_begin = 0;
    cp.bin_end = cp.n_bins;
    // the count kernel may only be launched once the refine CTAs sit on their SMs (it would take every SM otherwise)
    CK(launch_gate(c->d_slice_done + kSlices, (unsigned int)std::min<uint64_t>((uint64_t)part_sms, rp.n_coarse), c->stream, &c->launches));
    CK(cudaEventRecord(c->ev_cs0[0], c->stream));
    CK(launch_count(cp, c->N, dedup, c->sm_count - part_sms, c->stream, &c->launches));
    CK(cudaEventRecord(c->ev_cs1[0], c->stream));
    c->n_count_slices = 1;
    CK(cudaStreamWaitEvent(c->stream, c->ev_rs1[0], 0));
    CK(cudaEventRecord(c->ev_ik1, c->stream));
    return KC_OK;
  }
  if (refine && !redo) {
    rp.n_src = c->G;
    rp.rec_bytes = c->rec_bytes;
    rp.n_coarse = (uint32_t)c->P;
    rp.fine_shift = c->fine_shift;
    rp.src_header_bytes = c->header_bytes;
    rp.src_ptr = c->d_count_src;
    rp.dst_header_bytes = fine_header_bytes(c);
    rp.coarse_base = c->d_coarse_base;
    rp.total = c->d_coarse_base + c->P + 1;
    rp.err_too_large = &c->d_ctr->err_bins_full;
    rp.dst = c->d_fine;
    avg_records = record_bound(c) / std::max<uint64_t>(rp.n_coarse, 1);
    if (copy) {
      int rc = queue_peer_copies(c, -1, S, rs);  // the peers' header blocks
      if (rc) return rc;
    }
    CK(launch_refine_totals(rp, rs, &c->launches));
  } else if (copy) {
    // no refine pass: the count kernel reads every source's segments itself -- everything has to be here first
    int rc = queue_peer_copies(c, -1, S, c->stream);
    for (int i = 0; i < S && !rc; i++) rc = queue_peer_copies(c, i, S, c->stream);
    if (rc) return rc;
  }
  for (int i = 0; i < S; i++) {
    const uint32_t c0 = slice_edge((uint32_t)c->P, i, S), c1 = slice_edge((uint32_t)c->P, i + 1, S);
    if (refine && !redo) {
      if (copy) {
        int rc = queue_peer_copies(c, i, S, rs);
        if (rc) return rc;
      }
      rp.coarse_begin = c0;
      rp.coarse_end = c1;
      CK(cudaEventRecord(c->ev_rs0[i], rs));
      CK(launch_refine(rp, avg_records, c->sm_count, rs, &c->launches));
      CK(cudaEventRecord(c->ev_rs1[i], rs));
      c->n_refine_slices = i + 1;
      if (rs != c->stream) CK(cudaStreamWaitEvent(c->stream, c->ev_rs1[i], 0));
    }
    cp.bin_begin = c0 << c->fine_shift;
    cp.bin_end = c1 << c->fine_shift;
    CK(cudaEventRecord(c->ev_cs0[i], c->stream));
    if (cp.bin_end > cp.bin_begin) CK(launch_count(cp, c->N, dedup, c->sm_count, c->stream, &c->launches));
    CK(cudaEventRecord(c->ev_cs1[i], c->stream));
    c->n_count_slices = i + 1;
    if (to_sink) {
      CK(cudaMemcpyAsync(&c->h_slice_n[i], &c->d_ctr->n_final, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaEventRecord(c->ev_slice[i], c->stream));
    }
  }
  CK(cudaEventRecord(c->ev_ik1, c->stream));
  if (copy)
    for (int q = 0; q < kXferStreams; q++)
      if (c->xfer_started[q]) CK(cudaEventRecord(c->ev_xend[q], c->xfer_stream[q]));
  if (!to_sink) return KC_OK;
  // Result hand-off: after every slice the result cursor comes back through pinned memory; as soon as the host sees it,
  // the new entries [done, n) go out on the copy stream while the count stream is already busy with the next slices.
  c->sink_streamed = true;
  const int N = c->N;
  for (int i = 0; i < S; i++) {
    CK(cudaEventSynchronize(c->ev_slice[i]));
    uint64_t n = c->h_slice_n[i];
    if (n > c->res_final.cap) n = c->res_final.cap;  // (the kernel flagged err_result_full: handled by kc_finalize)
    if (n > c->sink_cap) n = c->sink_cap;
    if (n <= c->sink_done) continue;
    const uint64_t a = c->sink_done, m = n - a;
    CK(cudaMemcpyAsync(c->sink_keys + a * N, c->res_final.keys + a * N, m * N * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaMemcpyAsync(c->sink_counts + a, c->res_final.counts + a, m * sizeof(uint16_t), cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaMemcpyAsync(c->sink_left + a, c->res_final.left + a, m, cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaMemcpyAsync(c->sink_right + a, c->res_final.right + a, m, cudaMemcpyDeviceToHost, c->copy_stream));
    c->sink_done = n;
  }
  return KC_OK;
}

static int insert_resident(kc_ctx *c, const std::vector<uint8_t *> &src_in) {
  const bool keep_ext = (c->cfg.flags & KC_FLAG_KEEP_EXT_COUNTS) != 0;
  std::vector<uint8_t *> src = src_in;
  const bool copy = c->G > 1 && c->xmode == KC_XCHG_COPY && (int)c->pull_ptr.size() == c->G;
  if (copy)
    for (int s = 0; s < c->G; s++)
      if (s != c->rank) src[s] = c->pull_ptr[s] && c->recv_bytes[s] ? c->d_recv + c->recv_off[s] : nullptr;
  if (!c->d_count_src) CK(dev_alloc(&c->d_count_src, 2 * kCountMaxSrc));
  // (pinned staging: the previous use of these slots completed before the last kc_finalize returned)
  for (int s = 0; s < c->G; s++) c->h_pinned[s] = (unsigned long long)(uintptr_t)src[s];
  CK(cudaMemcpyAsync(c->d_count_src, c->h_pinned, c->G * sizeof(uint8_t *), cudaMemcpyHostToDevice, c->stream));
  // Result capacity. An entry that is written has count >= 2, so occurrences / 2 bounds it -- but reserving that much is
  // 5x what a typical run writes (24.6 GB reserved for a 4.8 GB result). Reserve what the last run of this
  // context needed plus a margin, else occurrences / 6; a short-fall is reported exactly by the kernel (err_result_full
  // with the need in the counters) and kc_finalize repeats the count pass once with the exact size.
  uint64_t occ = std::max<uint64_t>(c->h_ctr.parse_counted, 1024);
  if (c->G > 1) occ += occ / 4;
  uint64_t cap = c->last_final ? c->last_final + c->last_final / 8 + 1024 : occ / 6 + 1024;
  if (keep_ext) cap = std::max<uint64_t>(cap, c->last_ext ? c->last_ext + c->last_ext / 8 + 1024 : occ / 3 + 1024);
  cap = std::min<uint64_t>(cap, occ / 2 + 1024);
  const bool have_room = c->res_final.keys && c->res_final.cap >= cap && (!keep_ext || (c->res_ext.ext8 && c->res_ext.cap >= cap));
  if (!have_room) {  // (a driver query per pass is not free: only when something has to be allocated)
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    const uint64_t per_entry = (uint64_t)c->N * 8 + 4 + (keep_ext ? (uint64_t)c->N * 8 + 20 : 0);
    const uint64_t have = (c->res_final.cap * ((uint64_t)c->N * 8 + 4)) + (c->res_ext.cap * ((uint64_t)c->N * 8 + 20));
    const uint64_t budget = c->cfg.mem_budget_bytes ? c->cfg.mem_budget_bytes : (uint64_t)((free_b + have) * 0.7);
    if (cap * per_entry > budget) cap = std::max<uint64_t>(budget / per_entry, 1024);
  }
  if (c->res_final.cap < cap || !c->res_final.keys) {
    int rc = alloc_results(c, c->res_final, cap, false);
    if (rc) return rc;
  }
  if (keep_ext && (c->res_ext.cap < cap || !c->res_ext.ext8)) {
    int rc = alloc_results(c, c->res_ext, cap, true);
    if (rc) return rc;
  }
  if (c->fine_shift) {
    const uint64_t bound = record_bound(c);
    if (bound > 0xffffffffULL) return fail(c, KC_ERR_ARG, "more than 2^32 supermer records for one rank; use more ranks");
    const uint64_t need = (uint64_t)fine_header_bytes(c) + bound * c->rec_bytes + 256;
    if (need > c->fine_alloc) {
      CK(cudaStreamSynchronize(c->stream));
      dev_free(c->d_fine);
      c->fine_alloc = 0;
      CK(dev_alloc(&c->d_fine, need + need / 32));
      c->fine_alloc = need + need / 32;
    }
    if (c->coarse_base_alloc < (uint64_t)c->P + 2) {
      dev_free(c->d_coarse_base);
      CK(dev_alloc(&c->d_coarse_base, (uint64_t)c->P + 2));
      c->coarse_base_alloc = (uint64_t)c->P + 2;
    }
  }
  CK(cudaEventRecord(c->ev0, c->stream));
  c->stats.insert_kernel_launches = 1;
  int rc = launch_count_pipeline(c, false);
  if (rc) return rc;
  CK(cudaEventRecord(c->ev1, c->stream));
  c->phase = PH_INSERTED;
  c->res_final.n = c->res_ext.n = 0;
  c->res_final.sorted = c->res_ext.sorted = false;
  c->res_final.indexed = c->res_ext.indexed = false;
  return KC_OK;
}

static void fill_stats(kc_ctx *c);

static void pipeline_times(kc_ctx *c) {
  float ms = 0, sum = 0;
  for (int i = 0; i < c->n_count_slices; i++)
    if (cudaEventElapsedTime(&ms, c->ev_cs0[i], c->ev_cs1[i]) == cudaSuccess) sum += ms;
  c->stats.ms_insert_kernel = sum;
  sum = 0;
  for (int i = 0; i < c->n_refine_slices; i++)
    if (cudaEventElapsedTime(&ms, c->ev_rs0[i], c->ev_rs1[i]) == cudaSuccess) sum += ms;
  c->stats.ms_refine_kernel = sum;
  // transfer window of the copy engines: first copy start to last copy end, over the transfer streams
  float t0 = 1e30f, t1 = 0;
  for (int q = 0; q < kXferStreams; q++) {
    if (!c->xfer_started[q]) continue;
    if (cudaEventSynchronize(c->ev_xend[q]) != cudaSuccess) continue;
    float a = 0, b = 0;
    if (cudaEventElapsedTime(&a, c->ev_fork, c->ev_xbeg[q]) == cudaSuccess && cudaEventElapsedTime(&b, c->ev_fork, c->ev_xend[q]) == cudaSuccess) {
      t0 = std::min(t0, a);
      t1 = std::max(t1, b);
    }
  }
  c->stats.ms_exchange_copy = t1 > t0 ? t1 - t0 : 0.0;
  c->stats.exchange_copy_bytes = c->xfer_bytes;
  c->stats.pipeline_slices = c->n_count_slices;
  cudaGetLastError();
}

static int finalize_resident(kc_ctx *c, kc_stats *stats) {
  const bool keep_ext = (c->cfg.flags & KC_FLAG_KEEP_EXT_COUNTS) != 0;
  CK(cudaMemcpyAsync(&c->h_ctr, c->d_ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaStreamSynchronize(c->refine_stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->stats.ms_insert = ms;
  pipeline_times(c);
  c->stats.ms_table_clear = 0;
  c->stats.ms_finalize = 0;
  if (c->h_ctr.err_bins_full) return fail(c, KC_ERR_ARG, "more than 2^32 supermer records for one rank; use more ranks");
  if (c->h_ctr.err_wait) {
    // SM-partitioned pipeline: count warps gave up waiting (the refine kernel was not resident beside them). The fine bins
    // are complete by now (the refine stream was synchronised above): count them again, in the plain way.
    if (getenv("KC_DEBUG")) fprintf(stderr, "[kc] partitioned pipeline: count warps timed out waiting for the refine kernel, recounting\n");
    CK(cudaStreamSynchronize(c->copy_stream));
    int rc = launch_count_pipeline(c, true);
    if (rc) return rc;
    c->stats.insert_kernel_launches++;
    CK(cudaMemcpyAsync(&c->h_ctr, c->d_ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (c->h_ctr.err_wait) return fail(c, KC_ERR_STATE, "finalize: count pass could not run");
  }
  if (c->h_ctr.err_result_full) {
    // the dense result was under-sized: the counters hold the exact need, allocate that and count again
    CK(cudaEventRecord(c->ev0, c->stream));
    CK(cudaStreamSynchronize(c->copy_stream));
    int rc = alloc_results(c, c->res_final, c->h_ctr.n_final, false);
    if (rc) return rc;
    if (keep_ext && (rc = alloc_results(c, c->res_ext, c->h_ctr.n_ext, true))) return rc;
    const float first_pass = (float)c->stats.ms_insert_kernel;
    if ((rc = launch_count_pipeline(c, true))) return rc;
    c->stats.insert_kernel_launches++;
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaMemcpyAsync(&c->h_ctr, c->d_ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->stats.ms_finalize = ms;  // the repeated pass is charged to finalize
    const double refine_ms = c->stats.ms_refine_kernel;
    pipeline_times(c);
    c->stats.ms_refine_kernel = refine_ms;
    (void)first_pass;
    if (c->h_ctr.err_result_full) return fail(c, KC_ERR_STATE, "finalize: result overflow after exact re-allocation");
  }
  if (c->h_ctr.err_count_stack)
    return fail(c, KC_ERR_TABLE_FULL, "bin-resident counting: a bin could not be split far enough (use more local_bins or KC_FLAG_GLOBAL_TABLE)");
  c->res_final.n = c->h_ctr.n_final;
  c->res_ext.n = keep_ext ? c->h_ctr.n_ext : 0;
  c->last_final = c->h_ctr.n_final;
  c->last_ext = keep_ext ? c->h_ctr.n_ext : 0;
  if (c->sink_keys) {
    if (c->res_final.n > c->sink_cap)
      return fail(c, KC_ERR_ARG, "kc_finalize: %llu survivors do not fit the result sink (%llu entries)", (unsigned long long)c->res_final.n,
                  (unsigned long long)c->sink_cap);
    // whatever the slices have not sent yet (everything when the kernel was not sliced, or was repeated)
    const uint64_t a = c->sink_streamed ? c->sink_done : 0, m = c->res_final.n - a;
    const int N = c->N;
    if (m) {
      CK(cudaMemcpyAsync(c->sink_keys + a * N, c->res_final.keys + a * N, m * N * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->copy_stream));
      CK(cudaMemcpyAsync(c->sink_counts + a, c->res_final.counts + a, m * sizeof(uint16_t), cudaMemcpyDeviceToHost, c->copy_stream));
      CK(cudaMemcpyAsync(c->sink_left + a, c->res_final.left + a, m, cudaMemcpyDeviceToHost, c->copy_stream));
      CK(cudaMemcpyAsync(c->sink_right + a, c->res_final.right + a, m, cudaMemcpyDeviceToHost, c->copy_stream));
    }
    CK(cudaStreamSynchronize(c->copy_stream));
  }
  fill_stats(c);
  c->phase = PH_FINALIZED;
  if (stats) *stats = c->stats;
  return KC_OK;
}

extern "C" int kc_insert(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  CK(cudaSetDevice(c->device));
  if (c->G == 1) {
    if (c->phase < PH_PARSED) return fail(c, KC_ERR_STATE, "kc_insert: call kc_parse first");
  } else if (c->phase != PH_RECV_READY && c->phase != PH_INSERTED && c->phase != PH_FINALIZED) {
    return fail(c, KC_ERR_STATE, "kc_insert: exchange not completed (kc_exchange_done)");
  }
  std::vector<uint8_t *> src(c->G, nullptr);
  for (int s = 0; s < c->G; s++) {
    if (s == c->rank)
      src[s] = c->d_bins + c->h_region_off[c->rank];
    else if ((int)c->pull_ptr.size() == c->G)
      src[s] = c->pull_ptr[s];
    else
      src[s] = c->recv_bytes[s] ? c->d_recv + c->recv_off[s] : nullptr;
  }
  if (c->resident) return insert_resident(c, src);
  int rc = ensure_table(c);
  if (rc) return rc;
  const int n_seg = c->P * c->G;
  if (!c->d_src_ptr) {
    CK(dev_alloc(&c->d_src_ptr, c->G));
    CK(dev_alloc(&c->d_segments, n_seg));
    CK(dev_alloc(&c->d_seg_tile_start, n_seg + 1));
    CK(dev_alloc(&c->d_regions, c->P));
    CK(dev_alloc(&c->d_total_records, 1));
  }
  CK(cudaMemcpyAsync(c->d_src_ptr, src.data(), c->G * sizeof(uint8_t *), cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));

  CK(cudaEventRecord(c->ev0, c->stream));
  // clear table, filter and the insert-side counters
  CK(cudaMemsetAsync(c->d_A, 0, c->table_slots * sizeof(Slot), c->stream));
  if (c->N >= 2) CK(cudaMemsetAsync(c->d_C, 0, (c->table_slots + 4) * sizeof(uint32_t), c->stream));
  if (c->N >= 3) CK(cudaMemsetAsync(c->d_B, 0, c->table_slots * sizeof(uint4), c->stream));
  if (c->filter_buckets) CK(cudaMemsetAsync(c->d_filter, 0, c->filter_buckets * 32, c->stream));
  CK(cudaMemsetAsync(c->heavy.n, 0, sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(c->d_ctr, 0, offsetof(DevCounters, parse_counted), c->stream));

  PlanParams pp;
  pp.n_src = c->G;
  pp.local_bins = c->P;
  pp.rec_bytes = c->rec_bytes;
  pp.tile_records = insert_tile_records(c->rec_bytes);
  pp.src_ptr = c->d_src_ptr;
  pp.segments = c->d_segments;
  pp.seg_tile_start = c->d_seg_tile_start;
  pp.regions = c->d_regions;
  pp.table_slots = c->table_slots;
  pp.filter_buckets = c->filter_buckets;
  pp.total_records = c->d_total_records;
  CK(cudaEventRecord(c->ev_clr, c->stream));
  CK(launch_plan(pp, c->stream, &c->launches));

  InsertParams ip = insert_params(c);
  const bool merge = !(c->cfg.flags & KC_FLAG_NO_WARP_DEDUP);
  CK(cudaEventRecord(c->ev_ik0, c->stream));
  c->stats.insert_kernel_launches = c->cfg.use_filter ? 2 : 1;
  if (!c->cfg.use_filter) {
    CK(launch_insert(ip, c->N, MODE_INSERT, merge, c->sm_count, c->stream, &c->launches));
  } else {
    // exact singleton screen: pass A admits keys seen at least twice, pass B counts only admitted keys
    CK(launch_insert(ip, c->N, MODE_ADMIT, merge, c->sm_count, c->stream, &c->launches));
    CK(cudaMemsetAsync(&c->d_ctr->tile_cursor, 0, sizeof(unsigned long long), c->stream));
    CK(launch_insert(ip, c->N, MODE_UPDATE, merge, c->sm_count, c->stream, &c->launches));
  }
  CK(cudaEventRecord(c->ev_ik1, c->stream));
  CK(cudaEventRecord(c->ev1, c->stream));
  c->phase = PH_INSERTED;
  c->res_final.n = c->res_ext.n = 0;
  c->res_final.sorted = c->res_ext.sorted = false;
  c->res_final.indexed = c->res_ext.indexed = false;
  return KC_OK;
}

// ------------------------------------------------------------------------------------------------ finalize
// result arrays are pooled: they only grow (the bench loop and iterative-k callers reuse them)
static int alloc_results(kc_ctx *c, ResultSet &r, uint64_t cap, bool with_ext) {
  r.n = 0;
  r.sorted = r.indexed = false;
  if (cap <= r.cap && r.keys && (!with_ext || r.ext8)) return KC_OK;
  free_results(r);
  r.cap = cap + cap / 16 + 16;
  CK(dev_alloc(&r.keys, r.cap * c->N));
  CK(dev_alloc(&r.counts, r.cap));
  CK(dev_alloc(&r.left, r.cap));
  CK(dev_alloc(&r.right, r.cap));
  if (with_ext) CK(dev_alloc(&r.ext8, r.cap * 8));
  return KC_OK;
}

static void fill_stats(kc_ctx *c) {
  kc_stats &s = c->stats;
  const DevCounters &h = c->h_ctr;
  s.supermers_in = h.records_in;
  s.inserted_kmers = h.occurrences;
  s.unique_kmers = h.unique;
  s.filtered_singletons = c->cfg.use_filter ? (h.filter_new > h.unique ? h.filter_new - h.unique : 0) : 0;
  s.ge2_kmers = h.ge2;
  s.final_kmers = h.n_final;
  s.purged_kmers = h.unique - h.n_final;
  s.sum_counts = h.sum_counts;
  s.max_count = h.max_count;
  s.dropped = 0;
  s.table_slots = c->resident ? c->count_slots : c->table_slots;
  s.table_bytes = c->resident ? 0 : c->table_slots * slot_bytes(c->N);
  s.filter_bytes = c->resident ? 0 : c->filter_buckets * 32;
  s.local_bins = (uint64_t)c->P << c->fine_shift;
  s.split_passes = h.split_passes;
  s.result_bytes = c->res_final.cap * ((uint64_t)c->N * 8 + 4) + (c->res_ext.ext8 ? c->res_ext.cap * ((uint64_t)c->N * 8 + 20) : 0);
  s.avg_kmer_count = h.n_final ? (double)h.sum_counts / (double)h.n_final : 0.0;
  s.kernel_launches = c->launches;
}

extern "C" int kc_finalize(kc_ctx *c, kc_stats *stats) {
  if (!c) return KC_ERR_ARG;
  if (c->phase != PH_INSERTED) return fail(c, KC_ERR_STATE, "kc_finalize: call kc_insert first");
  CK(cudaSetDevice(c->device));
  if (c->resident) return finalize_resident(c, stats);
  CK(cudaMemcpyAsync(&c->h_ctr, c->d_ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->stats.ms_insert = ms;
  cudaEventElapsedTime(&ms, c->ev_ik0, c->ev_ik1);
  c->stats.ms_insert_kernel = ms;
  cudaEventElapsedTime(&ms, c->ev0, c->ev_clr);
  c->stats.ms_table_clear = ms;
  if (c->h_ctr.err_table_full) {
    fill_stats(c);
    return fail(c, KC_ERR_TABLE_FULL, "k-mer table full (%llu slots, %llu keys placed): results undefined, enlarge table_slots",
                (unsigned long long)c->table_slots, (unsigned long long)c->h_ctr.unique);
  }
  const bool keep_ext = (c->cfg.flags & KC_FLAG_KEEP_EXT_COUNTS) != 0;
  // capacity: every key that owns a slot could survive (count >= 2 is only known after the pass)
  int rc = alloc_results(c, c->res_final, c->h_ctr.unique, false);
  if (rc) return rc;
  if (keep_ext && (rc = alloc_results(c, c->res_ext, c->h_ctr.unique, true))) return rc;

  FinalizeParams fp;
  memset(&fp, 0, sizeof fp);
  fp.k = c->k;
  fp.n_words = c->N;
  fp.dmin_thres = c->cfg.dmin_thres;
  fp.keep_ext = keep_ext;
  fp.dyn_min_depth = c->cfg.dyn_min_depth;
  fp.n_slots = c->table_slots;
  fp.A = c->d_A;
  fp.B = c->d_B;
  fp.C = c->d_C;
  fp.heavy = c->heavy;
  fp.ctr = c->d_ctr;
  fp.out_keys = c->res_final.keys;
  fp.out_counts = c->res_final.counts;
  fp.out_left = c->res_final.left;
  fp.out_right = c->res_final.right;
  fp.cap_final = c->res_final.cap;
  fp.ext_keys = c->res_ext.keys;
  fp.ext_counts = c->res_ext.counts;
  fp.ext_ext8 = c->res_ext.ext8;
  fp.ext_left = c->res_ext.left;
  fp.ext_right = c->res_ext.right;
  fp.cap_ext = c->res_ext.cap;
  CK(cudaEventRecord(c->ev0, c->stream));
  CK(launch_finalize(fp, c->sm_count, c->stream, &c->launches));
  unsigned long long n_heavy = 0;
  CK(cudaMemcpyAsync(&n_heavy, c->heavy.n, sizeof n_heavy, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (n_heavy > c->heavy.cap) return fail(c, KC_ERR_STATE, "finalize: %llu keys with more than 65535 occurrences exceed the list of %llu",
                                          n_heavy, (unsigned long long)c->heavy.cap);
  if (n_heavy) {
    // keys with more than 65535 occurrences: recount their extensions exactly from the resident supermers
    CK(cudaMemsetAsync(c->heavy.side, 0, n_heavy * 8 * sizeof(uint32_t), c->stream));
    CK(cudaMemsetAsync(&c->d_ctr->tile_cursor, 0, sizeof(unsigned long long), c->stream));
    InsertParams ip = insert_params(c);
    CK(launch_insert(ip, c->N, MODE_HEAVY, !(c->cfg.flags & KC_FLAG_NO_WARP_DEDUP), c->sm_count, c->stream, &c->launches));
    CK(launch_finalize_heavy(fp, n_heavy, c->stream, &c->launches));
  }
  CK(cudaEventRecord(c->ev1, c->stream));
  CK(cudaMemcpyAsync(&c->h_ctr, c->d_ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  c->stats.ms_finalize = ms;
  c->res_final.n = c->h_ctr.n_final;
  c->res_ext.n = keep_ext ? c->h_ctr.n_ext : 0;
  if (c->res_final.n > c->res_final.cap || c->res_ext.n > c->res_ext.cap) return fail(c, KC_ERR_STATE, "finalize: result overflow");
  fill_stats(c);
  c->phase = PH_FINALIZED;
  if (stats) *stats = c->stats;
  return KC_OK;
}

extern "C" int kc_get_stats(kc_ctx *c, kc_stats *stats) {
  if (!c || !stats) return KC_ERR_ARG;
  c->stats.kernel_launches = c->launches;
  *stats = c->stats;
  return KC_OK;
}

extern "C" int kc_timer_mark(kc_ctx *c, int slot) {
  if (!c || slot < 0 || slot >= 8) return KC_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaEventRecord(c->ev_user[slot], c->stream));
  return KC_OK;
}

extern "C" int kc_timer_elapsed_ms(kc_ctx *c, int from, int to, double *ms) {
  if (!c || !ms || from < 0 || from >= 8 || to < 0 || to >= 8) return KC_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaEventSynchronize(c->ev_user[to]));
  float f = 0;
  CK(cudaEventElapsedTime(&f, c->ev_user[from], c->ev_user[to]));
  *ms = f;
  return KC_OK;
}

extern "C" int kc_reset(kc_ctx *c) {
  if (!c) return KC_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  c->res_final.n = c->res_ext.n = 0;
  c->phase = PH_EMPTY;
  c->caps_valid = false;  // the next kc_parse samples the reads again, as a fresh stage would
  c->recv_off.clear();
  c->pull_ptr.clear();
  return KC_OK;
}

extern "C" int kc_run(kc_ctx *c, kc_stats *stats) {
  if (!c) return KC_ERR_ARG;
  if (c->G != 1) return fail(c, KC_ERR_ARG, "kc_run: single-rank only; multi-rank callers drive kc_parse / kc_exchange_* / kc_insert / kc_finalize");
  int rc = kc_parse(c);
  if (rc) return rc;
  for (int attempt = 0; attempt < 4; attempt++) {
    rc = kc_insert(c);
    if (rc) return rc;
    rc = kc_finalize(c, stats);
    if (rc != KC_ERR_TABLE_FULL || c->cfg.table_slots) return rc;
    c->table_slots *= 2;  // automatic sizing under-estimated: grow and insert again from the resident supermers
    c->phase = PH_PARSED;
  }
  return rc;
}

// ------------------------------------------------------------------------------------------------ results
static ResultSet *pick(kc_ctx *c, int extended) {
  if (!c || c->phase != PH_FINALIZED) return nullptr;
  if (extended && !(c->cfg.flags & KC_FLAG_KEEP_EXT_COUNTS)) return nullptr;
  return extended ? &c->res_ext : &c->res_final;
}

extern "C" int kc_result_size(kc_ctx *c, int extended, uint64_t *n) {
  ResultSet *r = pick(c, extended);
  if (!r || !n) return fail(c, KC_ERR_STATE, "no %s result (finalize first%s)", extended ? "extended" : "final",
                            extended ? ", with KC_FLAG_KEEP_EXT_COUNTS" : "");
  *n = r->n;
  return KC_OK;
}

struct HostResult {
  std::vector<uint64_t> keys;
  std::vector<uint16_t> counts, ext8;
  std::vector<char> left, right;
};

static int fetch(kc_ctx *c, ResultSet *r, HostResult &h, bool want_ext) {
  h.keys.resize(r->n * c->N);
  h.counts.resize(r->n);
  h.left.resize(r->n);
  h.right.resize(r->n);
  CK(cudaMemcpyAsync(h.keys.data(), r->keys, r->n * c->N * 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(h.counts.data(), r->counts, r->n * 2, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(h.left.data(), r->left, r->n, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(h.right.data(), r->right, r->n, cudaMemcpyDeviceToHost, c->stream));
  if (want_ext && r->ext8) {
    h.ext8.resize(r->n * 8);
    CK(cudaMemcpyAsync(h.ext8.data(), r->ext8, r->n * 16, cudaMemcpyDeviceToHost, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  return KC_OK;
}

// Sorting is not on the hot path (only dumps and tests need an order): radix sort by key on the device (kc_sort.cu).
extern "C" int kc_result_sort(kc_ctx *c, int extended) {
  ResultSet *r = pick(c, extended);
  if (!r) return fail(c, KC_ERR_STATE, "kc_result_sort: no result");
  if (r->sorted || r->n == 0) {
    r->sorted = true;
    return KC_OK;
  }
  CK(cudaSetDevice(c->device));
  CK(sort_result_by_key(r->keys, r->counts, r->left, r->right, extended ? r->ext8 : nullptr, r->n, c->N, c->sm_count, c->stream, &c->launches));
  r->sorted = true;
  r->indexed = false;  // the entries moved
  return KC_OK;
}

extern "C" int kc_result_copy(kc_ctx *c, int extended, uint64_t *keys, uint16_t *counts, char *left, char *right, uint16_t *ext8) {
  ResultSet *r = pick(c, extended);
  if (!r) return fail(c, KC_ERR_STATE, "kc_result_copy: no result");
  CK(cudaSetDevice(c->device));
  if (keys) CK(cudaMemcpyAsync(keys, r->keys, r->n * c->N * 8, cudaMemcpyDeviceToHost, c->stream));
  if (counts) CK(cudaMemcpyAsync(counts, r->counts, r->n * 2, cudaMemcpyDeviceToHost, c->stream));
  if (left) CK(cudaMemcpyAsync(left, r->left, r->n, cudaMemcpyDeviceToHost, c->stream));
  if (right) CK(cudaMemcpyAsync(right, r->right, r->n, cudaMemcpyDeviceToHost, c->stream));
  if (ext8) {
    if (!r->ext8) return fail(c, KC_ERR_ARG, "kc_result_copy: ext8 only exists for the extended result");
    CK(cudaMemcpyAsync(ext8, r->ext8, r->n * 16, cudaMemcpyDeviceToHost, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  return KC_OK;
}

extern "C" int kc_result_device_ptrs(kc_ctx *c, int extended, const uint64_t **keys, const uint16_t **counts, const char **left,
                                     const char **right, const uint16_t **ext8) {
  ResultSet *r = pick(c, extended);
  if (!r) return fail(c, KC_ERR_STATE, "kc_result_device_ptrs: no result");
  if (keys) *keys = r->keys;
  if (counts) *counts = r->counts;
  if (left) *left = r->left;
  if (right) *right = r->right;
  if (ext8) *ext8 = r->ext8;
  return KC_OK;
}

extern "C" int kc_result_checksum(kc_ctx *c, int extended, uint64_t *checksum) {
  ResultSet *r = pick(c, extended);
  if (!r || !checksum) return fail(c, KC_ERR_STATE, "kc_result_checksum: no result");
  CK(cudaSetDevice(c->device));
  unsigned long long *d = c->d_scratch;  // pooled: no allocation per call, nothing to leak on the error paths
  CK(cudaMemsetAsync(d, 0, 8, c->stream));
  if (r->n) CK(launch_checksum(r->keys, r->counts, r->left, r->right, extended ? r->ext8 : nullptr, r->n, c->N, d, c->stream, &c->launches));
  unsigned long long h = 0;
  CK(cudaMemcpyAsync(&h, d, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  *checksum = h;
  return KC_OK;
}

static void key_to_string(const uint64_t *w, int k, char *s) {
  for (int i = 0; i < k; i++) s[i] = "ACGT"[(w[i / 32] >> (2 * (31 - (i % 32)))) & 3];
  s[k] = 0;
}

extern "C" int kc_dump_sorted(kc_ctx *c, const char *path, int extended) {
  ResultSet *r = pick(c, extended);
  if (!r || !path) return fail(c, KC_ERR_STATE, "kc_dump_sorted: no result");
  int rc = kc_result_sort(c, extended);
  if (rc) return rc;
  HostResult h;
  rc = fetch(c, r, h, true);
  if (rc) return rc;
  FILE *f = fopen(path, "w");
  if (!f) return fail(c, KC_ERR_IO, "cannot open %s", path);
  char s[KC_MAX_K + 1];
  for (uint64_t i = 0; i < r->n; i++) {
    key_to_string(&h.keys[i * c->N], c->k, s);
    if (extended && !h.ext8.empty()) {
      const uint16_t *e = &h.ext8[i * 8];
      fprintf(f, "%s %u %c %c %u %u %u %u %u %u %u %u\n", s, h.counts[i], h.left[i], h.right[i], e[0], e[1], e[2], e[3], e[4], e[5], e[6], e[7]);
    } else {
      fprintf(f, "%s %u %c %c\n", s, h.counts[i], h.left[i], h.right[i]);  // kmer_dht.cpp:286
    }
  }
  fclose(f);
  return KC_OK;
}

// KmerDHT::dump_kmers (src/kcount/kmer_dht.cpp:274-297): the rank's survivors, one line each, in table order, gzipped
extern "C" int kc_dump_kmers(kc_ctx *c, const char *path, int gz) {
  ResultSet *r = pick(c, 0);
  if (!r || !path) return fail(c, KC_ERR_STATE, "kc_dump_kmers: no result");
  CK(cudaSetDevice(c->device));
  HostResult h;
  int rc = fetch(c, r, h, false);
  if (rc) return rc;
  gzFile zf = nullptr;
  FILE *f = nullptr;
  if (gz)
    zf = gzopen(path, "wb");
  else
    f = fopen(path, "w");
  if (!zf && !f) return fail(c, KC_ERR_IO, "cannot open %s", path);
  std::string buf;
  buf.reserve(1 << 20);
  char s[KC_MAX_K + 1], tail[32];
  bool ok = true;
  for (uint64_t i = 0; i < r->n && ok; i++) {
    key_to_string(&h.keys[i * c->N], c->k, s);
    buf.append(s, c->k);
    snprintf(tail, sizeof tail, " %u %c %c\n", h.counts[i], h.left[i], h.right[i]);  // kmer_dht.cpp:286
    buf.append(tail);
    if (buf.size() > (1u << 20) - 256 || i + 1 == r->n) {
      ok = gz ? gzwrite(zf, buf.data(), (unsigned)buf.size()) == (int)buf.size() : fwrite(buf.data(), 1, buf.size(), f) == buf.size();
      buf.clear();
    }
  }
  if (gz) ok = gzclose(zf) == Z_OK && ok;
  else ok = fclose(f) == 0 && ok;
  return ok ? KC_OK : fail(c, KC_ERR_IO, "write to %s failed", path);
}

// KmerDHT::get_local_kmer_counts / kmer_exists / get_kmer_count (src/kcount/kmer_dht.cpp:199-245) as one batched call:
// the first lookup after a finalize builds the device index over the result (kc_lookup.cu), every call then runs one
// query kernel. keys / outputs may be host or device pointers (all of one kind); host batches are staged through HBM.
static int ensure_index(kc_ctx *c, ResultSet *r) {
  if (r->indexed) return KC_OK;
  if (r->n >= 0xffffffffULL) return fail(c, KC_ERR_STATE, "kc_lookup: more than 2^32 - 1 entries per GPU are not indexed");
  uint64_t slots = 1024;
  while (slots < 2 * r->n) slots <<= 1;
  if (slots > r->index_slots) {
    dev_free(r->index);
    r->index_slots = 0;
    CK(dev_alloc(&r->index, slots));
    r->index_slots = slots;
  }
  IndexParams ip;
  ip.keys = r->keys;
  ip.counts = r->counts;
  ip.left = r->left;
  ip.right = r->right;
  ip.n = r->n;
  ip.n_words = c->N;
  ip.slots = r->index;
  ip.mask = slots - 1;
  ip.err = &c->d_ctr->err_index;
  CK(launch_index_build(ip, c->sm_count, c->stream, &c->launches));
  unsigned int err = 0;
  CK(cudaMemcpyAsync(&err, &c->d_ctr->err_index, sizeof(err), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (err) return fail(c, KC_ERR_STATE, "kc_lookup: index build did not terminate");
  r->indexed = true;
  return KC_OK;
}

static bool on_device(const void *p) {
  cudaPointerAttributes at;
  if (!p) return false;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

extern "C" int kc_lookup(kc_ctx *c, const uint64_t *keys, uint64_t n, uint16_t *counts, char *left, char *right) {
  ResultSet *r = pick(c, 0);
  if (!r || (n && !keys)) return fail(c, KC_ERR_STATE, "kc_lookup: no result");
  if (n == 0) return KC_OK;
  CK(cudaSetDevice(c->device));
  int rc = ensure_index(c, r);
  if (rc) return rc;
  IndexParams ip;
  ip.keys = r->keys;
  ip.counts = r->counts;
  ip.left = r->left;
  ip.right = r->right;
  ip.n = r->n;
  ip.n_words = c->N;
  ip.slots = r->index;
  {  // the allocation may be pooled from a larger result: the mask is the one the index was built with
    uint64_t slots = 1024;
    while (slots < 2 * r->n) slots <<= 1;
    ip.mask = slots - 1;
  }
  ip.err = &c->d_ctr->err_index;
  const bool dev = on_device(keys);
  if (dev) {
    if ((counts && !on_device(counts)) || (left && !on_device(left)) || (right && !on_device(right)))
      return fail(c, KC_ERR_ARG, "kc_lookup: device keys need device outputs");
    CK(launch_index_lookup(ip, keys, n, counts, left, right, c->sm_count, c->stream, &c->launches));
    CK(cudaStreamSynchronize(c->stream));
    return KC_OK;
  }
  // host batch: staged in pieces through a pooled device buffer (grown on demand, kept by the context: the reference's
  // callers ask key by key, kmer_dht.cpp:199-245, so a lookup must not cost an allocation)
  const uint64_t piece = std::min<uint64_t>(n, (uint64_t)1 << 24);
  if (piece > c->lookup_cap) {
    dev_free(c->d_lq);
    dev_free(c->d_lc);
    dev_free(c->d_ll);
    dev_free(c->d_lr);
    c->lookup_cap = 0;
    const uint64_t cap = std::max<uint64_t>(piece, 4096);
    CK(dev_alloc(&c->d_lq, cap * c->N));
    CK(dev_alloc(&c->d_lc, cap));
    CK(dev_alloc(&c->d_ll, cap));
    CK(dev_alloc(&c->d_lr, cap));
    c->lookup_cap = cap;
  }
  for (uint64_t q0 = 0; q0 < n; q0 += piece) {
    const uint64_t nq = std::min(piece, n - q0);
    CK(cudaMemcpyAsync(c->d_lq, keys + q0 * c->N, nq * c->N * 8, cudaMemcpyHostToDevice, c->stream));
    CK(launch_index_lookup(ip, c->d_lq, nq, c->d_lc, c->d_ll, c->d_lr, c->sm_count, c->stream, &c->launches));
    if (counts) CK(cudaMemcpyAsync(counts + q0, c->d_lc, nq * 2, cudaMemcpyDeviceToHost, c->stream));
    if (left) CK(cudaMemcpyAsync(left + q0, c->d_ll, nq, cudaMemcpyDeviceToHost, c->stream));
    if (right) CK(cudaMemcpyAsync(right + q0, c->d_lr, nq, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
  }
  return KC_OK;
}

// KmerDHT::get_kmer_target_rank (src/kcount/kmer_dht.cpp:193-196): greatest least-complement m-mer, quick_hash, % ranks
extern "C" int kc_owner_of(const kc_ctx *c, const uint64_t *w) {
  if (!c || !w) return KC_ERR_ARG;
  const int k = c->k, m = c->m;
  const uint64_t mask = m < 32 ? ((1ULL << (2 * m)) - 1) : ~0ULL;
  uint64_t f = 0, r = 0, best = 0;
  for (int i = 0; i < k; i++) {
    uint64_t b = (w[i / 32] >> (2 * (31 - (i % 32)))) & 3;
    f = ((f << 2) | b) & mask;
    r = (r >> 2) | ((3 - b) << (2 * (m - 1)));
    if (i >= m - 1) best = std::max(best, std::min(f, r));
  }
  return (int)(quick_hash(best << (64 - 2 * m)) % (uint64_t)c->G);
}
