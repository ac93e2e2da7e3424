// TEST INFRASTRUCTURE (oracle/) -- driver of the reference's own read loading + pair merging: src/merge_reads.cpp, src/fastq.cpp,
// src/packed_reads.cpp, src/adapters.cpp, src/utils.cpp, src/linear_allocator_pool.cpp, src/ssw/*.cpp compiled UNMODIFIED from
// /root/reference behind oracle/merge_shim (one rank). Does what run_pipeline does before kcount (src/main.cpp:200-215):
//   PackedReadsList of PackedReads(qual_offset, merged name) -> merge_reads(...) -> the PackedReads hold the merged reads,
// then writes every read as FASTQ-like text "id\nseq\nquals\n" to the output file (ids are the reference's r<n>/1, r<n>/2).
//
// usage: merge_ref <qual_offset> <min_kmer_len> <out.txt> <reads file spec>...
//   a file spec is what options.cpp hands to FastqReader: "file" = interleaved pairs, "file:" = unpaired, "f1:f2" = two files.
// The golden fixtures under tests/golden/merge_*.json are made from its output by tests/golden/make_merge_golden.py.
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>
#include <upcxx/upcxx.hpp>
#include "packed_reads.hpp"
#include "utils.hpp"

using namespace std;

int merge_reads(vector<string> reads_fname_list, int qual_offset, double &elapsed_write_io_t, PackedReadsList &packed_reads_list,
                bool checkpoint, const string &adapter_fname, int min_kmer_len, int subsample_pct, bool use_blastn_scores);

int main(int argc, char **argv) {
  if (argc < 5) {
    fprintf(stderr, "usage: %s <qual_offset> <min_kmer_len> <out.txt> <reads file spec>...\n", argv[0]);
    return 2;
  }
  const int qual_offset = atoi(argv[1]), min_kmer_len = atoi(argv[2]);
  const string out = argv[3];
  vector<string> fnames;
  for (int i = 4; i < argc; i++) fnames.push_back(argv[i]);
  PackedReadsList packed_reads_list;
  for (auto const &f : fnames) packed_reads_list.push_back(new PackedReads(qual_offset, get_merged_reads_fname(f)));
  double elapsed_write_io_t = 0;
  const int avg_len = merge_reads(fnames, qual_offset, elapsed_write_io_t, packed_reads_list, false, "", min_kmer_len, 100, false);
  ofstream of(out);
  of << "#avg_read_len " << avg_len << "\n";
  for (auto pr : packed_reads_list) {
    pr->reset();
    string id, seq, quals;
    while (pr->get_next_read(id, seq, quals)) of << id << "\n" << seq << "\n" << quals << "\n";
  }
  of.close();
  return 0;
}
