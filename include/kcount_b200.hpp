// kcount_b200.hpp -- header-only C++ façade over the C ABI (kcount_b200.h) with the names, argument order and call
// sequence of the reference's kcount interface, so that a caller written against the reference compiles against this
// with a namespace change:
//
//   reference (src/contigging.cpp:83-104)                          here
//   -------------------------------------------------------------  ------------------------------------------------
//   Kmer<MAX_K>::set_k(k)                                          kcount_b200::Kmer<MAX_K>::set_k(k)
//   dist_object<KmerDHT<MAX_K>> kmer_dht(world(), my_num_kmers,    kcount_b200::dist_object<KmerDHT<MAX_K>> kmer_dht(
//       0, max_kmer_store, max_rpcs_in_flight, use_qf, seq_depth)      my_num_kmers, 0, max_kmer_store, max_rpcs, use_qf, seq_depth)
//   analyze_kmers(k, qual_offset, packed_reads_list, dmin_thres,   kcount_b200::analyze_kmers(k, qual_offset, packed_reads_list,
//       kmer_dht, dump_kmers)                                          dmin_thres, kmer_dht, dump_kmers)
//   kmer_dht->get_avg_kmer_count()                                 kmer_dht->get_avg_kmer_count()
//
// Declarations mirrored: src/kcount/kcount.hpp:56-72, src/kcount/kmer_dht.hpp:57-182, src/packed_reads.hpp:72-201,
// src/kmer.hpp:58-110. UPC++ is absent, so dist_object<T> is a local holder (operator-> / operator*), the supermer
// RPC store is replaced by the device-side bins + kc_exchange_* (multi-rank callers plug their transport into
// KmerDHT::exchange_hook), and errors are C++ exceptions (kc_error) instead of DIE().
#pragma once
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "kcount_b200.h"

namespace kcount_b200 {

struct kc_error : std::runtime_error {
  int code;
  kc_error(int c, const std::string &msg) : std::runtime_error("kcount_b200 error " + std::to_string(c) + ": " + msg), code(c) {}
};

using kmer_count_t = uint16_t;  // src/kcount/kmer_dht.hpp:55
inline int _dmin_thres = 2;     // process-wide in the reference as well (src/kcount/kmer_dht.hpp:57)

template <typename T>
class dist_object {  // stand-in for upcxx::dist_object: one instance per process
  std::unique_ptr<T> obj;

 public:
  template <typename... A>
  explicit dist_object(A &&...a) : obj(new T(std::forward<A>(a)...)) {}
  T *operator->() { return obj.get(); }
  T &operator*() { return *obj; }
  const T *operator->() const { return obj.get(); }
};

// ---- Kmer<MAX_K> (src/kmer.hpp:58-110): N_LONGS words, base i at bits 2*(31 - i%32) of word i/32, A0 C1 G2 T3
template <int MAX_K>
struct Kmer {
  static constexpr int N_LONGS = (MAX_K + 31) / 32;
  inline static unsigned int k = 0;
  uint64_t longs[N_LONGS] = {};

  static void set_k(unsigned int kk) { k = kk; }
  static unsigned int get_k() { return k; }
  static int get_N_LONGS() { return N_LONGS; }
  Kmer() = default;
  explicit Kmer(const char *s) {  // Kmer::set_kmer (src/kmer.cpp:303-325): case-insensitive, N -> G
    for (unsigned i = 0; i < k; i++) {
      uint64_t x = ((uint64_t)s[i] & 4) >> 1;
      x += (x ^ ((uint64_t)s[i] & 2)) >> 1;
      longs[i / 32] |= x << (2 * (31 - i % 32));
    }
  }
  explicit Kmer(const uint64_t *w) { std::memcpy(longs, w, sizeof longs); }
  bool operator<(const Kmer &o) const {  // src/kmer.cpp:271-277
    for (int i = 0; i < N_LONGS; i++) {
      if (longs[i] < o.longs[i]) return true;
      if (longs[i] > o.longs[i]) return false;
    }
    return false;
  }
  bool operator==(const Kmer &o) const { return !std::memcmp(longs, o.longs, sizeof longs); }
  Kmer revcomp() const {  // src/kmer.cpp:491-510
    Kmer r;
    for (unsigned i = 0; i < k; i++) {
      uint64_t b = (longs[i / 32] >> (2 * (31 - i % 32))) & 3;
      unsigned j = k - 1 - i;
      r.longs[j / 32] |= (3 - b) << (2 * (31 - j % 32));
    }
    return r;
  }
  std::string to_string() const {
    std::string s(k, 'A');
    for (unsigned i = 0; i < k; i++) s[i] = "ACGT"[(longs[i / 32] >> (2 * (31 - i % 32))) & 3];
    return s;
  }
};

struct KmerCounts {  // src/kcount/kmer_dht.hpp:62-68
  void *uutig_frag = nullptr;
  kmer_count_t count = 0;
  char left = 0, right = 0;
};

struct Supermer {  // src/kcount/kmer_dht.hpp:78-90 (wire unit of the reference; here only for add_supermer callers)
  std::string seq;
  kmer_count_t count;
};

// ---- PackedReads (src/packed_reads.hpp:130-201): the part a kcount caller touches. Bytes are the reference's own
// encoding (src/packed_reads.cpp:99-126): base(A0 C1 G2 T3 N4) | min(qual - offset, 31) << 3, one byte per base.
class PackedReads {
  int qual_offset;
  std::string fname;
  std::vector<uint8_t> bytes;
  std::vector<uint64_t> offsets;
  std::vector<uint16_t> lens;
  uint64_t index = 0;

 public:
  PackedReads(int qual_offset, const std::string &fname = "") : qual_offset(qual_offset), fname(fname) {}
  void add_read(const std::string & /*read_id*/, const std::string &seq, const std::string &quals) {  // packed_reads.cpp:344-352
    offsets.push_back(bytes.size());
    lens.push_back((uint16_t)seq.size());
    for (size_t i = 0; i < seq.size(); i++) {
      uint8_t b = 4;
      switch (seq[i]) {
        case 'A': case 'a': b = 0; break;
        case 'C': case 'c': b = 1; break;
        case 'G': case 'g': b = 2; break;
        case 'T': case 't': b = 3; break;
      }
      int q = (int)(unsigned char)quals[i] - qual_offset;
      bytes.push_back((uint8_t)(b | ((q > 31 ? 31 : (q < 0 ? 0 : q)) << 3)));
    }
  }
  // adopt an existing PackedRead byte stream (e.g. the reference's allocator blocks flattened by the caller)
  void add_packed(const uint8_t *b, uint64_t n_bytes, const uint64_t *off, const uint16_t *len, uint64_t n_reads) {
    const uint64_t base = bytes.size();
    bytes.insert(bytes.end(), b, b + n_bytes);
    for (uint64_t i = 0; i < n_reads; i++) {
      offsets.push_back(base + off[i]);
      lens.push_back(len[i]);
    }
  }
  uint64_t get_local_num_reads() const { return lens.size(); }
  int get_qual_offset() const { return qual_offset; }
  void reset() { index = 0; }
  const std::vector<uint8_t> &raw_bytes() const { return bytes; }
  const std::vector<uint64_t> &raw_offsets() const { return offsets; }
  const std::vector<uint16_t> &raw_lens() const { return lens; }
  static uint64_t get_total_local_num_reads(const std::vector<PackedReads *> &l) {
    uint64_t n = 0;
    for (auto *p : l) n += p->get_local_num_reads();
    return n;
  }
  // src/packed_reads.cpp:484-518: k-mers of the first 50 000 reads of each file, scaled to all reads
  static uint64_t estimate_num_kmers(unsigned kmer_len, const std::vector<PackedReads *> &l) {
    uint64_t num_kmers = 0, num_reads = 0, tot = get_total_local_num_reads(l);
    for (auto *p : l) {
      for (uint64_t i = 0; i < p->lens.size() && i < 50000; i++) {
        if (p->lens[i] < kmer_len) continue;
        num_kmers += p->lens[i] - kmer_len + 1;
        num_reads++;
      }
    }
    return num_reads ? num_kmers * tot / num_reads : 0;
  }
};
using PackedReadsList = std::vector<PackedReads *>;  // src/packed_reads.hpp:201 (a deque there)

// ---- KmerDHT<MAX_K> (src/kcount/kmer_dht.hpp:118-182). One instance per rank.
template <int MAX_K>
class KmerDHT {
  kc_ctx *ctx = nullptr;
  kc_stats stats{};
  double avg_kmer_count = 0;
  bool finished = false;
  int n_ranks = 1;
  std::chrono::time_point<std::chrono::high_resolution_clock> start_t = std::chrono::high_resolution_clock::now();
  // supermers handed in one at a time (add_supermer / HashTableInserter::insert_supermer): a supermer with count 1 IS a read
  // as far as counting goes (interior k-mers, flanking bases as extensions, lower case = low quality:
  // src/kcount/kcount_cpu.cpp:339-355), so they are buffered as PackedRead bytes and handed over in bulk
  std::vector<uint8_t> sm_bytes;
  std::vector<uint64_t> sm_offsets;
  std::vector<uint16_t> sm_lens;
  std::vector<std::pair<Kmer<MAX_K>, KmerCounts>> local_cache;  // local_kmers_begin / end
  bool local_cached = false;

  void flush_supermers() {
    if (sm_lens.empty()) return;
    ck(kc_add_reads(ctx, sm_bytes.data(), sm_bytes.size(), sm_offsets.data(), sm_lens.data(), sm_lens.size()));
    ck(kc_sync(ctx));
    sm_bytes.clear(), sm_offsets.clear(), sm_lens.clear();
  }

  void ck(int rc) const {
    if (rc != KC_OK) throw kc_error(rc, kc_last_error(ctx));
  }

 public:
  // multi-rank callers: moves every rank's per-owner regions to their owners between kc_parse and kc_insert
  // (kc_exchange_send_sizes / send_ptr / recv_buffer / recv_offsets, then the transport, then kc_exchange_done)
  std::function<void(kc_ctx *)> exchange_hook;

  // same parameter list as the reference constructor (kmer_dht.cpp:104-111) without the upcxx team; rank / n_ranks /
  // device are what upcxx::rank_me(), rank_n() and init_devices() supply there
  KmerDHT(uint64_t my_num_kmers, size_t /*my_num_ctg_kmers*/, size_t /*max_kmer_store_bytes*/, int /*max_rpcs_in_flight*/, bool use_qf,
          int sequencing_depth, int rank = 0, int n_ranks = 1, int device = -1, uint32_t flags = 0)
      : n_ranks(n_ranks) {
    if (!my_num_kmers) throw kc_error(KC_ERR_ARG, "no kmers to reserve space for");
    kc_config cfg;
    kc_config_default(&cfg);
    cfg.k = Kmer<MAX_K>::get_k();
    cfg.dmin_thres = _dmin_thres;
    cfg.use_filter = use_qf;
    cfg.rank = rank;
    cfg.n_ranks = n_ranks;
    cfg.device = device;
    cfg.est_kmers = my_num_kmers;
    cfg.sequencing_depth = sequencing_depth;
    cfg.flags = flags;
    int rc = kc_create(&cfg, &ctx);
    if (rc != KC_OK) throw kc_error(rc, kc_last_error(nullptr));
  }
  ~KmerDHT() { kc_destroy(ctx); }
  KmerDHT(const KmerDHT &) = delete;
  KmerDHT &operator=(const KmerDHT &) = delete;

  kc_ctx *context() { return ctx; }
  static size_t N() { return (size_t)Kmer<MAX_K>::N_LONGS; }
  void clear_stores() {}
  int get_minimizer_len() const { return (int)kc_minimizer_len(ctx); }
  uint64_t get_local_num_kmers() {
    uint64_t n = 0;
    if (finished) ck(kc_result_size(ctx, 0, &n));
    return n;
  }
  // sum over the ranks (kmer_dht.cpp:180-187): multi-rank callers plug their all-reduce in (upcxx::reduce_all there)
  std::function<uint64_t(uint64_t)> reduce_sum_hook;
  uint64_t get_num_kmers(bool /*all*/ = false) {
    const uint64_t n = get_local_num_kmers();
    return n_ranks > 1 && reduce_sum_hook ? reduce_sum_hook(n) : n;
  }
  void init_ctg_kmers() { throw kc_error(KC_ERR_ARG, "init_ctg_kmers: the contig k-mer pass is not on this path (src/kcount/kcount.cpp:150-153)"); }
  int32_t get_time_offset_us() const {  // kmer_dht.cpp:310-313
    return (int32_t)std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::high_resolution_clock::now() - start_t).count();
  }
  int get_kmer_target_rank(const Kmer<MAX_K> &kmer, const Kmer<MAX_K> *kmer_rc = nullptr) const {  // kmer_dht.cpp:193-196
    Kmer<MAX_K> rc = kmer_rc ? *kmer_rc : kmer.revcomp();
    return kc_owner_of(ctx, (rc < kmer ? rc : kmer).longs);
  }
  // a batch of get_local_kmer_counts in one kernel launch (count 0 = absent): what a traversal should call
  void get_local_kmer_counts_batch(const Kmer<MAX_K> *kmers, uint64_t n, KmerCounts *out) {
    std::vector<uint64_t> q(n * N());
    std::vector<uint16_t> c(n);
    std::vector<char> l(n), r(n);
    for (uint64_t i = 0; i < n; i++) std::memcpy(&q[i * N()], kmers[i].longs, N() * sizeof(uint64_t));
    if (n) ck(kc_lookup(ctx, q.data(), n, c.data(), l.data(), r.data()));
    for (uint64_t i = 0; i < n; i++) out[i] = KmerCounts{nullptr, c[i], l[i], r[i]};
  }
  KmerCounts *get_local_kmer_counts(const Kmer<MAX_K> &kmer) {  // kmer_dht.cpp:199-207; pointer valid until the next call
    static thread_local KmerCounts kc;
    uint16_t c = 0;
    char l = 0, r = 0;
    ck(kc_lookup(ctx, kmer.longs, 1, &c, &l, &r));
    if (!c) return nullptr;
    kc.count = c;
    kc.left = l;
    kc.right = r;
    return &kc;
  }
  bool kmer_exists(const Kmer<MAX_K> &kmer_fw) {  // kmer_dht.cpp:225-233
    Kmer<MAX_K> rc = kmer_fw.revcomp();
    return get_local_kmer_counts(rc < kmer_fw ? rc : kmer_fw) != nullptr;
  }
  kmer_count_t get_kmer_count(const Kmer<MAX_K> &kmer_fw) {  // kmer_dht.cpp:235-245
    Kmer<MAX_K> rc = kmer_fw.revcomp();
    KmerCounts *kc = get_local_kmer_counts(rc < kmer_fw ? rc : kmer_fw);
    return kc ? kc->count : 0;
  }
  uint64_t get_num_supermer_inserts() const { return stats.supermers_in; }
  double get_avg_kmer_count() const { return avg_kmer_count; }
  void set_avg_kmer_count(double v) { avg_kmer_count = v; }
  const kc_stats &get_stats() const { return stats; }

  // count_kmers' hand-over of the local reads (kcount.cpp:71-90); the supermer cut happens on the device
  void add_reads(const PackedReads &pr) {
    if (pr.get_local_num_reads())
      ck(kc_add_reads(ctx, pr.raw_bytes().data(), pr.raw_bytes().size(), pr.raw_offsets().data(), pr.raw_lens().data(),
                      pr.get_local_num_reads()));
  }
  // kmer_dht.cpp:248-251. The reference ships the supermer to target_rank by RPC; here the device cuts and routes by owner
  // itself, so the target is only checked
  void add_supermer(Supermer &supermer, int target_rank) {
    if (target_rank < 0 || target_rank >= n_ranks) throw kc_error(KC_ERR_ARG, "add_supermer: bad target rank");
    add_supermer_seq(supermer.seq, supermer.count);
  }
  void add_supermer_seq(const std::string &seq, kmer_count_t count) {
    if (count != 1) throw kc_error(KC_ERR_ARG, "add_supermer: count != 1 (contig k-mers) is not on this path");
    sm_offsets.push_back(sm_bytes.size());
    sm_lens.push_back((uint16_t)seq.size());
    for (char ch : seq) {
      uint8_t b = 4;
      switch (ch & ~0x20) {
        case 'A': b = 0; break;
        case 'C': b = 1; break;
        case 'G': b = 2; break;
        case 'T': b = 3; break;
      }
      sm_bytes.push_back((uint8_t)(b | ((ch >= 'A' && ch <= 'Z') ? 31 << 3 : 0)));  // upper case = high quality
    }
    if (sm_bytes.size() >= (64u << 20)) flush_supermers();
  }
  void flush_updates() {  // kmer_dht.cpp:253-258
    flush_supermers();
    ck(kc_parse(ctx));
    if (n_ranks > 1) {
      if (!exchange_hook) throw kc_error(KC_ERR_STATE, "multi-rank KmerDHT needs an exchange_hook");
      exchange_hook(ctx);
    }
    ck(kc_insert(ctx));
  }
  void finish_updates() {  // kmer_dht.cpp:261-267 -> insert_into_local_hashtable
    ck(kc_finalize(ctx, &stats));
    avg_kmer_count = stats.avg_kmer_count;
    finished = true;
    local_cached = false;
  }
  // what count_kmers / insert_into_local_hashtable log at this point (src/kcount/kcount.cpp:94-102, kcount_cpu.cpp:497-520)
  void log_stage_counters(FILE *f = stderr) const {
    fprintf(f, "Processed a total of %llu reads %llu raw kmers\n", (unsigned long long)stats.n_reads, (unsigned long long)stats.raw_kmers);
    fprintf(f, "Supermer inserts %llu (rank-local; avg / max / load come from the caller's reduction)\n", (unsigned long long)stats.supermers_in);
    fprintf(f, "Number of elements in hash table: %llu, purged %llu, dropped %llu\n", (unsigned long long)stats.unique_kmers,
            (unsigned long long)stats.purged_kmers, (unsigned long long)stats.dropped);
    fprintf(f, "Total kmers %llu, kmer count sum %llu, avg %.3f, max %llu\n", (unsigned long long)stats.final_kmers,
            (unsigned long long)stats.sum_counts, stats.avg_kmer_count, (unsigned long long)stats.max_count);
  }
  void dump_kmers(const std::string &path = "") {  // kmer_dht.cpp:274-297 (here: sorted, uncompressed)
    std::string p = path.empty() ? "kmers-" + std::to_string(Kmer<MAX_K>::get_k()) + ".txt" : path;
    ck(kc_dump_sorted(ctx, p.c_str(), 0));
  }
  // local_kmers_begin/end (kmer_dht.cpp:300-307): the dense result instead of a host hash map
  struct LocalKmers {
    std::vector<uint64_t> keys;  // [n * N_LONGS]
    std::vector<uint16_t> counts;
    std::vector<char> left, right;
    size_t size() const { return counts.size(); }
    Kmer<MAX_K> kmer(size_t i) const { return Kmer<MAX_K>(&keys[i * Kmer<MAX_K>::N_LONGS]); }
    KmerCounts counts_of(size_t i) const { return KmerCounts{nullptr, counts[i], left[i], right[i]}; }
  };
  // Caller-owned host arrays that receive the survivors WHILE the count kernel runs (kc_set_result_sink): the hand-off
  // done_all_inserts() does with a copy at the end (src/kcount/kcount-gpu/gpu_hash_table.cpp:737-785). Call before the
  // reads are analysed; get_local_num_kmers() entries are valid afterwards (unsorted).
  void set_result_sink(uint64_t *keys, uint16_t *counts, char *left, char *right, uint64_t capacity) {
    ck(kc_set_result_sink(ctx, keys, counts, left, right, capacity));
  }
  // local_kmers_begin / local_kmers_end (kmer_dht.cpp:300-307): `it->first` is the Kmer, `it->second` the KmerCounts, as with
  // the reference's map iterator; the entries are materialised on the host by the first call after finish_updates
  using local_iterator = typename std::vector<std::pair<Kmer<MAX_K>, KmerCounts>>::iterator;
  local_iterator local_kmers_begin() {
    if (!local_cached) {
      LocalKmers lk = local_kmers(false);
      local_cache.clear();
      local_cache.reserve(lk.size());
      for (size_t i = 0; i < lk.size(); i++) local_cache.emplace_back(lk.kmer(i), lk.counts_of(i));
      local_cached = true;
    }
    return local_cache.begin();
  }
  local_iterator local_kmers_end() {
    local_kmers_begin();
    return local_cache.end();
  }
  LocalKmers local_kmers(bool sorted = false) {
    LocalKmers r;
    uint64_t n = get_local_num_kmers();
    if (sorted) ck(kc_result_sort(ctx, 0));
    r.keys.resize(n * kc_key_words(ctx));
    r.counts.resize(n);
    r.left.resize(n);
    r.right.resize(n);
    if (n) ck(kc_result_copy(ctx, 0, r.keys.data(), r.counts.data(), r.left.data(), r.right.data(), nullptr));
    return r;
  }
};

// ---- the reference's own plug point (chosen at link time there, src/kcount/CMakeLists.txt:24-69), for callers that drive the
// stage read by read / supermer by supermer. integration/kcount_b200_glue.cpp is the same thing against the reference's
// real class declarations.
template <int MAX_K>
struct SeqBlockInserter {  // src/kcount/kcount.hpp:56-68
  int qual_offset, minimizer_len;
  SeqBlockInserter(int qual_offset, int minimizer_len) : qual_offset(qual_offset), minimizer_len(minimizer_len) {}
  // seq: the read with low-quality bases already in lower case (count_kmers, src/kcount/kcount.cpp:80-85)
  void process_seq(std::string &seq, kmer_count_t depth, dist_object<KmerDHT<MAX_K>> &kmer_dht) {
    if (depth > 1) throw kc_error(KC_ERR_ARG, "process_seq: depth > 1 (contig k-mers) is not on this path");
    kmer_dht->add_supermer_seq(seq, 1);
  }
  void done_processing(dist_object<KmerDHT<MAX_K>> &) {}
};

template <int MAX_K>
class HashTableInserter {  // src/kcount/kmer_dht.hpp:95-116, over an existing KmerDHT's stage
  KmerDHT<MAX_K> *dht = nullptr;
  std::chrono::duration<double> t_insert{0};

 public:
  void attach(KmerDHT<MAX_K> &d) { dht = &d; }
  void init(size_t, size_t, size_t, bool) {}  // sizing happened in the KmerDHT constructor (est_kmers)
  void init_ctg_kmers(size_t) { throw kc_error(KC_ERR_ARG, "init_ctg_kmers: not on this path"); }
  void insert_supermer(const std::string &supermer_seq, kmer_count_t supermer_count) { dht->add_supermer_seq(supermer_seq, supermer_count); }
  void flush_inserts() {
    auto t0 = std::chrono::high_resolution_clock::now();
    dht->flush_updates();
    t_insert += std::chrono::high_resolution_clock::now() - t0;
  }
  double insert_into_local_hashtable() {
    auto t0 = std::chrono::high_resolution_clock::now();
    dht->finish_updates();
    t_insert += std::chrono::high_resolution_clock::now() - t0;
    return dht->get_avg_kmer_count();
  }
  void get_elapsed_time(double &insert_time, double &kernel_time) {
    insert_time = t_insert.count();
    const kc_stats &s = dht->get_stats();
    kernel_time = (s.ms_parse + s.ms_insert + s.ms_finalize) / 1e3;
  }
};

// ---- analyze_kmers<MAX_K> (src/kcount/kcount.hpp:70-72, src/kcount/kcount.cpp:143-161)
template <int MAX_K>
void analyze_kmers(unsigned kmer_len, int qual_offset, PackedReadsList &packed_reads_list, int dmin_thres,
                   dist_object<KmerDHT<MAX_K>> &kmer_dht, bool dump_kmers) {
  if (kmer_len != Kmer<MAX_K>::get_k()) throw kc_error(KC_ERR_ARG, "analyze_kmers: kmer_len differs from Kmer<MAX_K>::k");
  if (dmin_thres != _dmin_thres) throw kc_error(KC_ERR_ARG, "analyze_kmers: set kcount_b200::_dmin_thres before constructing the KmerDHT");
  for (auto *pr : packed_reads_list) {
    if (pr->get_qual_offset() != qual_offset) throw kc_error(KC_ERR_ARG, "analyze_kmers: qual_offset differs from the reads'");
    kmer_dht->add_reads(*pr);  // count_kmers
  }
  kmer_dht->flush_updates();
  kmer_dht->finish_updates();
  if (dump_kmers) kmer_dht->dump_kmers();
  kmer_dht->clear_stores();
}

}  // namespace kcount_b200
