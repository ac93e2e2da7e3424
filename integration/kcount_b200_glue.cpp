// kcount_b200_glue.cpp -- the reference-side binding: a FOURTH alternative for the plug point that
// src/kcount/CMakeLists.txt:24-69 fills at link time with exactly one of kcount_cpu.cpp, kcount_gpu.cpp (CUDA / HIP) or
// kcount_gpu.cpp (Kokkos). It defines the two pimpl classes the reference declares
//     SeqBlockInserter<MAX_K>   src/kcount/kcount.hpp:56-68
//     HashTableInserter<MAX_K>  src/kcount/kmer_dht.hpp:95-116
// by forwarding to the C ABI of include/kcount_b200.h (libkcount_b200.so). Everything above it -- analyze_kmers /
// count_kmers (src/kcount/kcount.cpp), KmerDHT (src/kcount/kmer_dht.cpp), Kmer (src/kmer.cpp) -- is the reference's own
// code, compiled unmodified. A maintainer drops this file into src/kcount/ and selects it in CMake;
// in this repo `make -C oracle ref_b200` compiles it in place of kcount_cpu.cpp into oracle/_ref/kcount_ref_b200 and
// tests/test_glue_reference.py holds its KmerDHT::local_kmers to the CPU build's (oracle/_ref/kcount_ref).
//
// The reference drives the plug point read by read (process_seq) and supermer by supermer (insert_supermer); the stage
// behind the ABI wants a rank's reads in bulk and cuts / exchanges / counts them on the device. So process_seq only
// re-encodes the (already quality-masked) read into PackedRead bytes; done_processing hands the batch to kc_add_reads;
// flush_inserts runs kc_parse, the owner exchange over peer memory, kc_insert; insert_into_local_hashtable runs
// kc_finalize and fills the reference's KmerMap. The UPC++ supermer store stays empty; a caller that does use
// KmerDHT::add_supermer (src/kcount/kmer_dht.cpp:248) still works: a supermer with count 1 IS a read as far as counting
// goes (interior k-mers, flanking bases as extensions, lower case = low quality; src/kcount/kcount_cpu.cpp:339-355).
#include <cctype>
#include <chrono>
#include <cstring>
#include <string>
#include <vector>

#include "kcount_b200.h"

#include "upcxx_utils.hpp"
#include "kcount.hpp"
#include "kmer_dht.hpp"

using namespace std;
using namespace upcxx;
using namespace upcxx_utils;

// All-gather of a small POD record over the ranks of the job. With real UPC++ this is a dist_object<Rec> plus
// fetch(r).wait() for every r; the single-node stand-in of oracle/ref_shim offers it directly.
#ifdef KCOUNT_B200_SHIM_WORLD
static void b200_all_gather(const void *mine, size_t bytes, void *all) { shim::world().allgather_bytes(mine, bytes, all); }
#else
static void b200_all_gather(const void *mine, size_t bytes, void *all) {
  static_assert(KC_IPC_HANDLE_BYTES == 64, "record layout");
  std::vector<uint8_t> rec((const uint8_t *)mine, (const uint8_t *)mine + bytes);
  upcxx::dist_object<std::vector<uint8_t>> d(rec);
  upcxx::barrier();
  for (int r = 0; r < upcxx::rank_n(); r++) {
    auto v = d.fetch(r).wait();
    memcpy((uint8_t *)all + (size_t)r * bytes, v.data(), bytes);
  }
  upcxx::barrier();
}
#endif

namespace {

// one stage per process, like Kmer<MAX_K>::k and _dmin_thres in the reference (src/kmer.hpp:63, src/kcount/kmer_dht.hpp:57)
struct B200Stage {
  kc_ctx *ctx = nullptr;
  kc_config cfg;
  bool configured = false;
  std::vector<uint8_t> bytes;
  std::vector<uint64_t> offsets;
  std::vector<uint16_t> lens;
  double t_insert = 0, t_kernel = 0;
  bool exchanged = false;

  void append(const std::string &seq) {
    // PackedRead byte = base(A0 C1 G2 T3 N4) | quality << 3 (src/packed_reads.cpp:99-126). count_kmers has already turned
    // "quality below the cutoff" into lower case (src/kcount/kcount.cpp:80-85): upper -> 31, lower -> 0.
    offsets.push_back(bytes.size());
    lens.push_back((uint16_t)seq.size());
    for (char c : seq) {
      uint8_t b;
      switch (c & ~0x20) {
        case 'A': b = 0; break;
        case 'C': b = 1; break;
        case 'G': b = 2; break;
        case 'T': b = 3; break;
        default: b = 4; break;
      }
      bytes.push_back((uint8_t)(b | (isupper((unsigned char)c) ? 31 << 3 : 0)));
    }
  }

  kc_ctx *get(int k) {
    if (ctx) return ctx;
    if (!configured) DIE("kcount_b200 glue: HashTableInserter::init was not called");
    cfg.k = (uint32_t)k;
    cfg.dmin_thres = _dmin_thres;  // set by analyze_kmers before count_kmers runs (src/kcount/kcount.cpp:147)
    cfg.qual_offset = 0;           // the bytes built by append() carry 0 / 31 directly
    cfg.rank = rank_me();
    cfg.n_ranks = rank_n();
    const int ndev = kc_device_count();
    if (ndev < 1) DIE("kcount_b200 glue: no CUDA device (this implementation has no CPU path)");
    cfg.device = rank_me() % ndev;  // as the reference spreads ranks over devices (src/kcount/kcount-gpu/gpu-utils/gpu_utils.cpp:110)
    if (kc_create(&cfg, &ctx) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(nullptr));
    return ctx;
  }

  void flush_reads(int k) {
    if (lens.empty()) return;
    kc_ctx *c = get(k);
    if (kc_add_reads(c, bytes.data(), bytes.size(), offsets.data(), lens.data(), lens.size()) != KC_OK || kc_sync(c) != KC_OK)
      DIE("kcount_b200 glue: ", kc_last_error(c));
    bytes.clear();
    offsets.clear();
    lens.clear();
  }

  // the owner exchange over peer memory: export, all-gather of ~200 bytes, import
  void exchange(kc_ctx *c) {
    const int n = rank_n(), me = rank_me();
    struct Rec {
      uint8_t handle[KC_IPC_HANDLE_BYTES];
      uint64_t off[64], bytes[64];
    };
    if (n > 64) DIE("kcount_b200 glue: more than 64 ranks per node");
    Rec mine;
    memset(&mine, 0, sizeof mine);
    if (kc_exchange_export(c, mine.handle, mine.off, mine.bytes) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(c));
    std::vector<Rec> all(n);
    b200_all_gather(&mine, sizeof mine, all.data());
    for (int p = 0; p < n; p++)
      if (p != me && kc_exchange_import(c, p, all[p].handle, all[p].off[me], all[p].bytes[me]) != KC_OK)
        DIE("kcount_b200 glue: ", kc_last_error(c));
    if (kc_exchange_done(c) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(c));
    exchanged = true;
  }

  ~B200Stage() {
    if (ctx) kc_destroy(ctx);
  }
};

B200Stage g_stage;

}  // namespace

// ------------------------------------------------------------------------------------------------ SeqBlockInserter
template <int MAX_K>
struct SeqBlockInserter<MAX_K>::SeqBlockInserterState {
  int qual_offset, minimizer_len;
};

template <int MAX_K>
SeqBlockInserter<MAX_K>::SeqBlockInserter(int qual_offset, int minimizer_len) {
  state = new SeqBlockInserterState{qual_offset, minimizer_len};
}

template <int MAX_K>
SeqBlockInserter<MAX_K>::~SeqBlockInserter() {
  if (state) delete state;
}

template <int MAX_K>
void SeqBlockInserter<MAX_K>::process_seq(string &seq, kmer_count_t depth, dist_object<KmerDHT<MAX_K>> &kmer_dht) {
  if (depth > 1) DIE("kcount_b200 glue: contig k-mers (depth > 1) are not on this path");
  g_stage.append(seq);
  if (g_stage.bytes.size() >= (256u << 20)) g_stage.flush_reads(Kmer<MAX_K>::get_k());  // hand over in 256 MB batches
}

template <int MAX_K>
void SeqBlockInserter<MAX_K>::done_processing(dist_object<KmerDHT<MAX_K>> &kmer_dht) {
  g_stage.flush_reads(Kmer<MAX_K>::get_k());
}

// ------------------------------------------------------------------------------------------------ HashTableInserter
template <int MAX_K>
struct HashTableInserter<MAX_K>::HashTableInserterState {
  kc_stats stats;
};

template <int MAX_K>
HashTableInserter<MAX_K>::HashTableInserter() {}

template <int MAX_K>
HashTableInserter<MAX_K>::~HashTableInserter() {
  if (state) delete state;
}

template <int MAX_K>
void HashTableInserter<MAX_K>::init(size_t num_elems, size_t num_ctg_elems, size_t num_errors, bool use_qf) {
  state = new HashTableInserterState();
  memset(&state->stats, 0, sizeof state->stats);
  this->use_qf = use_qf;
  kc_config &cfg = g_stage.cfg;
  kc_config_default(&cfg);
  // KmerDHT::KmerDHT has already applied the sequencing depth and the k-mer error rate (src/kcount/kmer_dht.cpp:126-131):
  // num_elems + num_errors IS the number of distinct k-mers to size for
  cfg.est_kmers = num_elems + num_errors;
  cfg.sequencing_depth = 1;
  cfg.base_error_rate = 0.0;
  cfg.use_filter = use_qf ? 1 : 0;
  g_stage.configured = true;
}

template <int MAX_K>
void HashTableInserter<MAX_K>::init_ctg_kmers(size_t max_elems) {
  DIE("kcount_b200 glue: the contig k-mer pass is not on this path (dead code in this proxy, src/kcount/kcount.cpp:150-153)");
}

template <int MAX_K>
void HashTableInserter<MAX_K>::insert_supermer(const std::string &supermer_seq, kmer_count_t supermer_count) {
  // only reached when a caller fills the UPC++ store itself (KmerDHT::add_supermer): count it as the read it is
  if (supermer_count != 1) DIE("kcount_b200 glue: supermer count ", supermer_count, " (contig k-mers) is not on this path");
  g_stage.append(supermer_seq);
}

template <int MAX_K>
void HashTableInserter<MAX_K>::flush_inserts() {
  auto t0 = std::chrono::high_resolution_clock::now();
  g_stage.flush_reads(Kmer<MAX_K>::get_k());
  kc_ctx *c = g_stage.get(Kmer<MAX_K>::get_k());
  if (kc_parse(c) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(c));
  if (rank_n() > 1) g_stage.exchange(c);
  if (kc_insert(c) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(c));
  g_stage.t_insert += std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
}

template <int MAX_K>
double HashTableInserter<MAX_K>::insert_into_local_hashtable(dist_object<KmerMap<MAX_K>> &local_kmers) {
  auto t0 = std::chrono::high_resolution_clock::now();
  kc_ctx *c = g_stage.get(Kmer<MAX_K>::get_k());
  kc_stats &st = state->stats;
  if (kc_finalize(c, &st) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(c));
  if (g_stage.exchanged) {
    barrier();  // every owner is done reading its peers' bins
    kc_exchange_release(c);
    g_stage.exchanged = false;
  }
  uint64_t n = 0;
  kc_result_size(c, 0, &n);
  const int N = Kmer<MAX_K>::get_N_LONGS();
  std::vector<uint64_t> keys(n * N);
  std::vector<uint16_t> counts(n);
  std::vector<char> left(n), right(n);
  if (kc_result_copy(c, 0, keys.data(), counts.data(), left.data(), right.data(), nullptr) != KC_OK) DIE("kcount_b200 glue: ", kc_last_error(c));
  local_kmers->reserve(n);
  for (uint64_t i = 0; i < n; i++) {  // what src/kcount/kcount_gpu.cpp:439-471 does after its device -> host copy
    KmerCounts kc = {.uutig_frag = nullptr, .count = counts[i], .left = left[i], .right = right[i]};
    local_kmers->insert({Kmer<MAX_K>(&keys[i * N]), kc});
  }
  g_stage.t_kernel += (st.ms_parse + st.ms_insert + st.ms_finalize) / 1e3;
  g_stage.t_insert += std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
  // the counters the reference logs here (src/kcount/kcount_cpu.cpp:497-520, 586-600)
  auto all_unique = reduce_one((uint64_t)st.unique_kmers, op_fast_add, 0).wait();
  auto all_purged = reduce_one((uint64_t)st.purged_kmers, op_fast_add, 0).wait();
  SLOG_VERBOSE("B200 k-mer table: ", all_unique, " unique k-mers, purged ", all_purged, ", dropped 0\n");
  auto tot_kmers_sum = reduce_all((uint64_t)st.sum_counts, op_fast_add).wait();
  auto tot_kmers = reduce_all((uint64_t)n, op_fast_add).wait();
  return tot_kmers ? (double)tot_kmers_sum / tot_kmers : 0.0;
}

template <int MAX_K>
void HashTableInserter<MAX_K>::get_elapsed_time(double &insert_time, double &kernel_time) {
  insert_time = g_stage.t_insert;
  kernel_time = g_stage.t_kernel;
}

#define SEQ_BLOCK_INSERTER_K(KMER_LEN) template struct SeqBlockInserter<KMER_LEN>;
#define HASH_TABLE_INSERTER_K(KMER_LEN) template class HashTableInserter<KMER_LEN>;
SEQ_BLOCK_INSERTER_K(32);
HASH_TABLE_INSERTER_K(32);
#if MAX_BUILD_KMER >= 64
SEQ_BLOCK_INSERTER_K(64);
HASH_TABLE_INSERTER_K(64);
#endif
#if MAX_BUILD_KMER >= 96
SEQ_BLOCK_INSERTER_K(96);
HASH_TABLE_INSERTER_K(96);
#endif
#if MAX_BUILD_KMER >= 128
SEQ_BLOCK_INSERTER_K(128);
HASH_TABLE_INSERTER_K(128);
#endif
#undef SEQ_BLOCK_INSERTER_K
#undef HASH_TABLE_INSERTER_K
