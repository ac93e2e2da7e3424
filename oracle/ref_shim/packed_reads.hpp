// TEST INFRASTRUCTURE (oracle/) -- minimal PackedReads for the oracle build. The reference's loader
// (packed_reads.cpp, fastq.cpp, UPC++ futures) is out of scope; this keeps only what count_kmers
// (kcount.cpp:71-77) calls, fed from the 1-byte-per-base stream described at packed_reads.cpp:99-126:
// byte = base(3 bits: A0 C1 G2 T3 N4) | min(qual - offset, 31) << 3, and decoded as at :196-208.
#pragma once
#include <cstdint>
#include <deque>
#include <string>
#include <vector>

class PackedReads {
  const unsigned char *bytes;
  const uint64_t *offsets;
  const uint16_t *lens;
  uint64_t n_reads;
  uint64_t index = 0;
  int qual_offset;

 public:
  PackedReads(int qual_offset, const unsigned char *bytes, const uint64_t *offsets, const uint16_t *lens, uint64_t n_reads)
      : bytes(bytes), offsets(offsets), lens(lens), n_reads(n_reads), qual_offset(qual_offset) {}

  void reset() { index = 0; }

  bool get_next_read(std::string &id, std::string &seq, std::string &quals) {
    static const char nucleotide_map[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'};
    if (index == n_reads) return false;
    const unsigned char *b = bytes + offsets[index];
    int len = lens[index];
    id = "@r" + std::to_string(index) + "/1";
    seq.resize(len);
    quals.resize(len);
    for (int i = 0; i < len; i++) {
      seq[i] = nucleotide_map[b[i] & 7];
      quals[i] = qual_offset + (b[i] >> 3);
    }
    index++;
    return true;
  }

  int64_t get_local_num_reads() const { return n_reads; }

  static int64_t get_total_local_num_reads(const std::deque<PackedReads *> &l) {
    int64_t n = 0;
    for (auto p : l) n += p->get_local_num_reads();
    return n;
  }
};

using PackedReadsList = std::deque<PackedReads *>;
