// A caller written like the reference's contigging<MAX_K>() (src/contigging.cpp:83-104) against include/kcount_b200.hpp:
// seeded synthetic reads -> PackedReads -> KmerDHT -> analyze_kmers -> sorted `KMER count L R` dump.
//   facade_main <k> <n_reads> <seed> <dump path>
// exit 0 = ok, 42 = no CUDA device (KC_ERR_CUDA, the library has no CPU path), anything else = failure.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "kc_synth.h"
#include "kcount_b200.hpp"

using namespace kcount_b200;

template <int MAX_K>
static int run(unsigned k, uint64_t n_reads, uint64_t seed, const char *dump) {
  kc_synth_config sc;
  kc_synth_config_default(&sc);
  sc.seed = seed;
  sc.n_genomes = 3;
  sc.genome_len_min = 20000;
  sc.genome_len_max = 40000;
  sc.abundance_sigma = 1.0;
  sc.n_rate = 0.001;
  kc_synth *syn = nullptr;
  if (kc_synth_create(&sc, &syn) != KC_SYNTH_OK) return 2;
  std::vector<uint8_t> bytes(n_reads * sc.read_len);
  std::vector<uint64_t> off(n_reads);
  std::vector<uint16_t> len(n_reads);
  uint64_t nb = 0;
  if (kc_synth_generate(syn, 0, n_reads, bytes.data(), off.data(), len.data(), &nb, 4) != KC_SYNTH_OK) return 3;
  kc_synth_destroy(syn);

  const int qual_offset = 33, dmin_thres = 2;
  Kmer<MAX_K>::set_k(k);
  PackedReads reads(qual_offset, "synthetic");
  reads.add_packed(bytes.data(), nb, off.data(), len.data(), n_reads);
  PackedReadsList packed_reads_list{&reads};
  uint64_t my_num_kmers = PackedReads::estimate_num_kmers(k, packed_reads_list);
  _dmin_thres = dmin_thres;
  dist_object<KmerDHT<MAX_K>> kmer_dht(my_num_kmers, (size_t)0, (size_t)0, 100, false, 4);
  analyze_kmers(k, qual_offset, packed_reads_list, dmin_thres, kmer_dht, false);
  kmer_dht->dump_kmers(dump);
  // spot checks of the accessors the next MHM2 stage uses
  auto lk = kmer_dht->local_kmers(true);
  if (lk.size() != kmer_dht->get_num_kmers()) return 4;
  for (size_t i = 0; i < lk.size() && i < 100; i++) {
    Kmer<MAX_K> km = lk.kmer(i);
    if (!kmer_dht->kmer_exists(km.revcomp())) return 5;
    if (kmer_dht->get_kmer_count(km) != lk.counts[i]) return 6;
    if (kmer_dht->get_kmer_target_rank(km) != 0) return 7;
  }
  // the plug-point route: the same reads, one at a time through SeqBlockInserter::process_seq (lower case = low quality,
  // as count_kmers prepares them) into a second KmerDHT; its table must be the first one's
  {
    dist_object<KmerDHT<MAX_K>> dht2(my_num_kmers, (size_t)0, (size_t)0, 100, false, 4);
    SeqBlockInserter<MAX_K> sbi(qual_offset, dht2->get_minimizer_len());
    std::string seq;
    for (uint64_t i = 0; i < n_reads; i++) {
      if (len[i] < k) continue;
      seq.resize(len[i]);
      for (unsigned j = 0; j < len[i]; j++) {
        const uint8_t b = bytes[off[i] + j];
        const char c = "ACGTN"[(b & 7) > 4 ? 4 : (b & 7)];
        seq[j] = (b >> 3) < 20 ? (char)(c | 0x20) : c;
      }
      sbi.process_seq(seq, 0, dht2);
    }
    sbi.done_processing(dht2);
    HashTableInserter<MAX_K> hti;
    hti.attach(*dht2);
    hti.flush_inserts();
    const double avg = hti.insert_into_local_hashtable();
    if (dht2->get_num_kmers() != kmer_dht->get_num_kmers() || avg != kmer_dht->get_avg_kmer_count()) return 8;
    // local_kmers_begin / end, and a batch of lookups against the first table
    std::vector<Kmer<MAX_K>> ks;
    std::vector<KmerCounts> want;
    for (auto it = dht2->local_kmers_begin(); it != dht2->local_kmers_end() && ks.size() < 5000; ++it) {
      ks.push_back(it->first);
      want.push_back(it->second);
    }
    std::vector<KmerCounts> got(ks.size());
    kmer_dht->get_local_kmer_counts_batch(ks.data(), ks.size(), got.data());
    for (size_t i = 0; i < ks.size(); i++)
      if (got[i].count != want[i].count || got[i].left != want[i].left || got[i].right != want[i].right) return 9;
    if (dht2->get_time_offset_us() <= 0) return 11;
  }
  printf("{\"k\": %u, \"kmers\": %llu, \"avg_kmer_count\": %.9f, \"supermer_inserts\": %llu, \"m\": %d}\n", k,
         (unsigned long long)kmer_dht->get_num_kmers(), kmer_dht->get_avg_kmer_count(),
         (unsigned long long)kmer_dht->get_num_supermer_inserts(), kmer_dht->get_minimizer_len());
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 5) return 1;
  unsigned k = (unsigned)atoi(argv[1]);
  uint64_t n = strtoull(argv[2], nullptr, 10), seed = strtoull(argv[3], nullptr, 10);
  try {
    switch ((k / 32 + 1) * 32) {  // MAX_K dispatch of src/main.cpp:169-190
      case 32: return run<32>(k, n, seed, argv[4]);
      case 64: return run<64>(k, n, seed, argv[4]);
      case 96: return run<96>(k, n, seed, argv[4]);
      case 128: return run<128>(k, n, seed, argv[4]);
    }
    return 1;
  } catch (const kc_error &e) {
    fprintf(stderr, "%s\n", e.what());
    return e.code == KC_ERR_CUDA ? 42 : 10;
  }
}
