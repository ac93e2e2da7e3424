// TEST INFRASTRUCTURE (oracle/) -- builds oracle/_ref/kcount_ref: the reference's OWN CPU k-mer
// analysis (src/kmer.cpp, src/hash_funcs.c, src/kcount/kcount.cpp, src/kcount/kcount_cpu.cpp,
// src/kcount/kmer_dht.cpp) compiled unmodified from /root/reference behind the UPC++ stand-in in
// oracle/ref_shim/. This file is the driver: it mirrors contigging<MAX_K> (src/contigging.cpp:83-104),
// runs P forked "ranks" (one per host core when asked) and writes the sorted dumps D1/D2 that
// SURVEY.md section 8c defines. The only reference function restated here is comp_nucleotide
// (src/utils.cpp:132-159): utils.cpp itself drags in process pinning / UPC++ promises and is not built.
// Never linked into, imported by, or executed from the product path.
#include <sys/wait.h>
#include <unistd.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <numeric>
#include <string>
#include <vector>

#include "upcxx_utils.hpp"
#include "utils.hpp"

// restated: src/utils.cpp:132-159 (only the cases the kcount path can produce)
char comp_nucleotide(char ch) {
  switch (ch) {
    case 'A':
    case 'a': return 'T';
    case 'C':
    case 'c': return 'G';
    case 'G':
    case 'g': return 'C';
    case 'T':
    case 't': return 'A';
    case 'N':
    case 'n': return 'N';
    case '0': return '0';
    default: DIE("Illegal char in comp nucleotide: ", (int)ch);
  }
  return 0;
}

#ifdef KC_B200_GLUE
// the same driver over the B200 binding: integration/kcount_b200_glue.cpp takes the place of kcount_cpu.cpp at the
// reference's link-time plug point (src/kcount/CMakeLists.txt:24-69) -> oracle/_ref/kcount_ref_b200
#define KCOUNT_B200_SHIM_WORLD
#include "../integration/kcount_b200_glue.cpp"
#else
#include "kcount/kcount_cpu.cpp"  // pulls kcount.hpp / kmer_dht.hpp / kmer.hpp; instantiates both inserters
#undef SLOG_CPU_HT
#endif

// ---------------------------------------------------------------- reads file (format: see reads_io.h)
#include "reads_io.h"

// restated: PackedReads::estimate_num_kmers (src/packed_reads.cpp:484-518): first 50 000 reads, scaled.
static uint64_t estimate_num_kmers(unsigned k, const kc_reads_file &rf, uint64_t lo, uint64_t hi) {
  int64_t num_kmers = 0, num_reads = 0, tot = hi - lo;
  for (uint64_t i = lo; i < hi && i < lo + 50000; i++) {
    if (rf.lens[i] < k) continue;
    num_kmers += rf.lens[i] - k + 1;
    num_reads++;
  }
  return num_reads > 0 ? num_kmers * tot / num_reads : 0;
}

struct Args {
  std::string reads, out;
  unsigned k = 21;
  int qual_offset = 33, dmin = 2, ranks = 1, seq_depth = 4;
  uint64_t max_reads = 0;
  bool extended = false, quiet = false;
};

struct RankResult {  // lives in shared memory
  double t_dht, t_analyze;
  uint64_t n_final, n_ext, sum_count, occurrences, sum_all_counts, n_entries, max_count;
  double avg_kmer_count;
  int64_t num_supermer_inserts;
};

template <int MAX_K>
static void write_part(const std::string &path, const std::vector<const Kmer<MAX_K> *> &keys, const std::vector<uint16_t> &vals,
                       int vals_per) {
  // vals_per u16 values per entry: D1 = {count, L, R}; D2 = {count, 8 ext counts, L, R}
  const int N = Kmer<MAX_K>::get_N_LONGS();
  std::vector<size_t> order(keys.size());
  std::iota(order.begin(), order.end(), 0);
  std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return *keys[a] < *keys[b]; });
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) DIE("cannot write ", path);
  uint64_t hdr[4] = {0x31504d44434bULL /* "KCDMP1" */, (uint64_t)N, (uint64_t)keys.size(), (uint64_t)vals_per};
  fwrite(hdr, 8, 4, f);
  for (size_t i : order) fwrite(keys[i]->get_longs(), 8, N, f);
  for (size_t i : order) fwrite(&vals[i * vals_per], 2, vals_per, f);
  fclose(f);
}

template <int MAX_K>
static void run_rank(const Args &a, const kc_reads_file &rf, RankResult *res) {
  using namespace std::chrono;
  auto &w = shim::world();
  // contiguous 1/P slice of the reads per rank (data-parallel over reads)
  uint64_t n = a.max_reads && a.max_reads < rf.n_reads ? a.max_reads : rf.n_reads;
  uint64_t lo = n * w.rank / w.nranks, hi = n * (w.rank + 1) / w.nranks;
  PackedReads pr(a.qual_offset, rf.bytes, rf.offsets + lo, rf.lens + lo, hi - lo);
  PackedReadsList list{&pr};

  // --- mirrors contigging<MAX_K>, src/contigging.cpp:83-104
  Kmer<MAX_K>::set_k(a.k);
  auto t0 = high_resolution_clock::now();
  auto my_num_kmers = reduce_all(estimate_num_kmers(a.k, rf, lo, hi), op_fast_add).wait() / rank_n();
  int my_num_ctg_kmers = 0;
  size_t max_kmer_store = 50 * 1024 * 1024;
  dist_object<KmerDHT<MAX_K>> kmer_dht(world(), my_num_kmers, my_num_ctg_kmers, max_kmer_store, 100, false, a.seq_depth);
  barrier();
  auto t1 = high_resolution_clock::now();
  analyze_kmers(a.k, a.qual_offset, list, a.dmin, kmer_dht, false);
  barrier();
  auto t2 = high_resolution_clock::now();
  res->t_dht = duration<double>(t1 - t0).count();
  res->t_analyze = duration<double>(t2 - t1).count();
  res->avg_kmer_count = kmer_dht->get_avg_kmer_count();
  res->num_supermer_inserts = kmer_dht->get_num_supermer_inserts();

  uint64_t occ = 0;
  for (uint64_t i = lo; i < hi; i++)
    if (rf.lens[i] >= a.k + 2) occ += rf.lens[i] - a.k - 1;
  res->occurrences = occ;

  // --- D1: the final per-rank map (what dump_kmers would write, kmer_dht.cpp:274-297)
  {
    std::vector<const Kmer<MAX_K> *> keys;
    std::vector<uint16_t> vals;
    uint64_t sum = 0;
    for (auto it = kmer_dht->local_kmers_begin(); it != kmer_dht->local_kmers_end(); ++it) {
      keys.push_back(&it->first);
      vals.push_back(it->second.count);
      vals.push_back((uint16_t)it->second.left);
      vals.push_back((uint16_t)it->second.right);
      sum += it->second.count;
    }
    res->n_final = keys.size();
    res->sum_count = sum;
    if (!a.out.empty()) write_part<MAX_K>(a.out + ".d1.rank" + std::to_string(w.rank), keys, vals, 3);
  }
#ifdef KC_B200_GLUE
  res->n_ext = 0;
  res->sum_all_counts = res->n_entries = res->max_count = 0;
  (void)a.extended;
#else
  // --- D2: every table entry with count >= 2, with its 8 saturating ext counters and the chosen
  //     ext chars (incl. X / F). Reads the file-local KmerMapExts through -fno-access-control.
  {
    auto &tbl = *(kmer_dht->ht_inserter->state->kmers);
    std::vector<const Kmer<MAX_K> *> keys;
    std::vector<uint16_t> vals;
    uint64_t sum_all = 0, n_entries = 0, max_count = 0;
    tbl.begin_iterate();
    while (true) {
      auto [kmer, ec] = tbl.get_next();
      if (!kmer) break;
      n_entries++;
      sum_all += ec->count;
      if (ec->count > max_count) max_count = ec->count;
      if (ec->count < 2) continue;
      if (!a.extended) continue;
      keys.push_back(kmer);
      vals.push_back(ec->count);
      vals.push_back(ec->left_exts.count_A);
      vals.push_back(ec->left_exts.count_C);
      vals.push_back(ec->left_exts.count_G);
      vals.push_back(ec->left_exts.count_T);
      vals.push_back(ec->right_exts.count_A);
      vals.push_back(ec->right_exts.count_C);
      vals.push_back(ec->right_exts.count_G);
      vals.push_back(ec->right_exts.count_T);
      vals.push_back((uint16_t)ec->get_left_ext());
      vals.push_back((uint16_t)ec->get_right_ext());
    }
    res->n_ext = keys.size();
    res->sum_all_counts = sum_all;
    res->n_entries = n_entries;
    res->max_count = max_count;
    if (a.extended && !a.out.empty()) write_part<MAX_K>(a.out + ".d2.rank" + std::to_string(w.rank), keys, vals, 11);
  }
#endif
}

static void merge_parts(const std::string &out, const char *tag, int P) {
  // ranks own disjoint key sets; merge = concatenate + sort by key words
  uint64_t N = 0, per = 0, total = 0;
  std::vector<std::vector<uint64_t>> K(P);
  std::vector<std::vector<uint16_t>> V(P);
  for (int r = 0; r < P; r++) {
    std::string p = out + "." + tag + ".rank" + std::to_string(r);
    FILE *f = fopen(p.c_str(), "rb");
    if (!f) {
      fprintf(stderr, "missing %s\n", p.c_str());
      exit(1);
    }
    uint64_t hdr[4];
    if (fread(hdr, 8, 4, f) != 4) exit(1);
    N = hdr[1];
    per = hdr[3];
    K[r].resize(hdr[2] * N);
    V[r].resize(hdr[2] * per);
    if (fread(K[r].data(), 8, K[r].size(), f) != K[r].size()) exit(1);
    if (fread(V[r].data(), 2, V[r].size(), f) != V[r].size()) exit(1);
    fclose(f);
    unlink(p.c_str());
    total += hdr[2];
  }
  std::vector<std::pair<int, uint64_t>> idx;
  idx.reserve(total);
  for (int r = 0; r < P; r++)
    for (uint64_t i = 0; i < K[r].size() / N; i++) idx.push_back({r, i});
  std::sort(idx.begin(), idx.end(), [&](auto &a, auto &b) {
    const uint64_t *x = &K[a.first][a.second * N], *y = &K[b.first][b.second * N];
    return std::lexicographical_compare(x, x + N, y, y + N);
  });
  FILE *f = fopen((out + "." + tag).c_str(), "wb");
  uint64_t hdr[4] = {0x31504d44434bULL, N, total, per};
  fwrite(hdr, 8, 4, f);
  for (auto &e : idx) fwrite(&K[e.first][e.second * N], 8, N, f);
  for (auto &e : idx) fwrite(&V[e.first][e.second * per], 2, per, f);
  fclose(f);
}

int main(int argc, char **argv) {
  Args a;
  for (int i = 1; i < argc; i++) {
    std::string s = argv[i];
    auto next = [&]() { return std::string(argv[++i]); };
    if (s == "--reads") a.reads = next();
    else if (s == "--out") a.out = next();
    else if (s == "--k") a.k = std::stoi(next());
    else if (s == "--qual-offset") a.qual_offset = std::stoi(next());
    else if (s == "--dmin") a.dmin = std::stoi(next());
    else if (s == "--ranks") a.ranks = std::stoi(next());
    else if (s == "--seq-depth") a.seq_depth = std::stoi(next());
    else if (s == "--max-reads") a.max_reads = std::stoull(next());
    else if (s == "--extended") a.extended = true;
    else {
      fprintf(stderr,
              "usage: kcount_ref --reads F --k K [--out PREFIX] [--extended] [--ranks P] [--dmin D] [--qual-offset Q] "
              "[--seq-depth S] [--max-reads N]\n");
      return 2;
    }
  }
  if (a.k % 32 == 0 || a.k < 1 || a.k >= 128) {
    fprintf(stderr, "k must be in [1,127] and not a multiple of 32 (reference defect, SURVEY.md 8a)\n");
    return 2;
  }
  kc_reads_file rf;
  if (kc_reads_file_map(a.reads.c_str(), &rf) != 0) {
    fprintf(stderr, "cannot map %s\n", a.reads.c_str());
    return 1;
  }
  int P = a.ranks;
  auto &w = shim::world();
  uint64_t n = a.max_reads && a.max_reads < rf.n_reads ? a.max_reads : rf.n_reads;
  // mailbox capacity: every k-mer of a shard could in the worst case travel as its own (k+2)-char supermer
  uint64_t shard_bytes = (rf.n_bytes * (double)n / std::max<uint64_t>(rf.n_reads, 1)) / P + 1;
  uint64_t box_cap = std::max<uint64_t>(4ull << 20, shard_bytes * 12 / P * 2);
  w.init_shared(P, box_cap);
  RankResult *res = (RankResult *)shim::World::shared_alloc(sizeof(RankResult) * P);
  std::vector<pid_t> pids;
  for (int r = 0; r < P; r++) {
    pid_t pid = P == 1 ? 0 : fork();
    if (pid == 0) {
      w.rank = r;
      unsigned max_k = (a.k / 32 + 1) * 32;  // src/main.cpp:169
      switch (max_k) {
        case 32: run_rank<32>(a, rf, &res[r]); break;
        case 64: run_rank<64>(a, rf, &res[r]); break;
        case 96: run_rank<96>(a, rf, &res[r]); break;
        case 128: run_rank<128>(a, rf, &res[r]); break;
      }
      if (P > 1) _exit(0);
    } else {
      pids.push_back(pid);
    }
  }
  int bad = 0;
  for (pid_t p : pids) {
    int st;
    waitpid(p, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad++;
  }
  if (bad) {
    fprintf(stderr, "%d rank(s) failed\n", bad);
    return 1;
  }
  RankResult t{};
  for (int r = 0; r < P; r++) {
    t.t_dht = std::max(t.t_dht, res[r].t_dht);
    t.t_analyze = std::max(t.t_analyze, res[r].t_analyze);
    t.n_final += res[r].n_final;
    t.n_ext += res[r].n_ext;
    t.sum_count += res[r].sum_count;
    t.occurrences += res[r].occurrences;
    t.sum_all_counts += res[r].sum_all_counts;
    t.n_entries += res[r].n_entries;
    t.max_count = std::max(t.max_count, res[r].max_count);
    t.num_supermer_inserts += res[r].num_supermer_inserts;
  }
  if (!a.out.empty()) {
    merge_parts(a.out, "d1", P);
#ifndef KC_B200_GLUE
    if (a.extended) merge_parts(a.out, "d2", P);
#endif
  }
  // without u16 saturation, sum of all table counts == counted occurrences iff nothing was dropped
  long long dropped = t.max_count < 65535 ? (long long)t.occurrences - (long long)t.sum_all_counts : -1;
#ifdef KC_B200_GLUE
  dropped = -1;  // (the binding fails loudly instead of dropping: KC_ERR_TABLE_FULL)
#endif
  printf(
      "{\"k\": %u, \"ranks\": %d, \"reads\": %llu, \"occurrences\": %llu, \"t_dht_s\": %.6f, \"t_analyze_s\": %.6f, "
      "\"entries\": %llu, \"final_kmers\": %llu, \"ge2_kmers\": %llu, \"sum_count\": %llu, \"max_count\": %llu, "
      "\"avg_kmer_count\": %.6f, \"supermer_inserts\": %lld, \"dropped\": %lld}\n",
      a.k, P, (unsigned long long)n, (unsigned long long)t.occurrences, t.t_dht, t.t_analyze, (unsigned long long)t.n_entries,
      (unsigned long long)t.n_final, (unsigned long long)t.n_ext, (unsigned long long)t.sum_count,
      (unsigned long long)t.max_count, res[0].avg_kmer_count, (long long)t.num_supermer_inserts, dropped);
  return 0;
}
