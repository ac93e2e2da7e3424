/*
 * kc_synth.h -- seeded synthetic PackedRead streams and the read-file format the oracle
 * harness reads. Host-only C, no CUDA: built as its own libkc_synth.so so that the CPU arms of bench.py (the reference
 * run, the oracle) can generate their input without mapping the product library libkcount_b200.so.
 * No reference counterpart: the reference reads FASTQ (src/fastq.cpp); the byte format produced here is the
 * reference's in-memory PackedRead encoding (src/packed_reads.cpp:99-126).
 */
#ifndef KC_SYNTH_H
#define KC_SYNTH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes (the values of the like-named kc_status codes in kcount_b200.h) */
enum { KC_SYNTH_OK = 0, KC_SYNTH_ERR_ARG = -1, KC_SYNTH_ERR_NOMEM = -3, KC_SYNTH_ERR_IO = -7, KC_SYNTH_ERR_FORMAT = -8 };

typedef struct {
  uint64_t seed;
  uint32_t n_genomes;
  uint64_t genome_len_min, genome_len_max; /* uniform */
  double abundance_sigma;                  /* lognormal(0, sigma) coverage per genome; 0 = equal */
  uint32_t read_len, read_len_min;         /* read_len_min == 0 -> fixed length */
  double sub_rate;                         /* substitution probability per base */
  double lowq_rate, lowq_rate_sub;         /* P(quality '+') on correct / substituted bases (0.02 / 0.5) */
  double n_rate;                           /* P(base reported as N); 0 in every BASELINE config */
  int32_t paired;                          /* 1: reads 2i,2i+1 are mates (fwd/rev) of one fragment */
  double insert_mean, insert_sd;
  int32_t qual_offset;
  uint32_t reserved[5];
} kc_synth_config;
typedef struct kc_synth kc_synth;
void kc_synth_config_default(kc_synth_config *cfg); /* config C1 of BASELINE.md with seed 1001 */
int kc_synth_create(const kc_synth_config *cfg, kc_synth **out);
void kc_synth_destroy(kc_synth *s);
/* reads [first, first+n) of the infinite seeded stream; read i depends only on (seed, i). bytes must hold
 * n*read_len; offsets/lens hold n entries; offsets are relative to bytes. Returns bytes used in *n_bytes. */
int kc_synth_generate(const kc_synth *s, uint64_t first, uint64_t n, uint8_t *bytes, uint64_t *offsets, uint16_t *lens,
                      uint64_t *n_bytes, int n_threads);
int kc_reads_save(const char *path, const uint8_t *bytes, uint64_t n_bytes, const uint64_t *offsets, const uint16_t *lens,
                  uint64_t n_reads);
/* A read set in the reference's PackedReads memory layout (src/packed_reads.hpp:68-79, 136): headers[2 i] = bit fields
 * {is_allocated : 1, read_id : 47, read_len : 16}, headers[2 i + 1] = pointer to the read's bytes inside `arena`, which is
 * cut into blocks of block_bytes (4 MB in the reference) filled by a bump allocator, placed in descending address order.
 * Input of kc_add_packed_reads in tests and bench. Returns 0, -1 (argument) or -3 (arena too small). */
int kc_reads_to_packed_layout(const uint8_t *bytes, const uint64_t *offsets, const uint16_t *lens, uint64_t n_reads, uint8_t *arena,
                              uint64_t arena_bytes, uint64_t block_bytes, uint64_t *headers, uint64_t *blocks_used);

/* FASTQ (plain or gzip) -> PackedRead bytes + offsets + lens, the flat input of kc_add_reads. Follows
 * FastqReader::get_next_fq_record (src/fastq.cpp:1028-1120: '@' name / sequence / '+' / qualities of equal length) and the
 * PackedRead encoding (src/packed_reads.cpp:99-126). qual_offset 33 or 64; max_reads 0 = all. The arrays are malloc'ed:
 * release them with kc_read_set_free. Pair merging / adapter trimming (src/merge_reads.cpp) is not done here. */
typedef struct {
  uint8_t *bytes;
  uint64_t n_bytes;
  uint64_t *offsets;
  uint16_t *lens;
  uint64_t n_reads;
} kc_read_set;
int kc_fastq_read(const char *path, int qual_offset, uint64_t max_reads, kc_read_set *out);
void kc_read_set_free(kc_read_set *rs);

/* FASTQ -> mates as the pair loop of merge_reads() sees them (src/merge_reads.cpp:340-466), the input of the merge kernel
 * (kc_merge_pairs / kc_add_read_pairs in kcount_b200.h): reads 2 i and 2 i + 1 are the mates of pair i, sequence and quality
 * CHARACTERS as in the file (merging needs the full quality values, which the 5-bit PackedRead encoding drops). path2 = NULL:
 * one interleaved file; path2 != NULL: mates alternate between the two files. paired = 0: an unpaired file -- every read gets
 * the dummy mate "N" (quality = offset) that the reference gives it. Names are normalised as FastqReader::get_fq_name does
 * (src/fastq.cpp:77-135: "/1" "/2", "-R1" "-R2", " 1:N:0:.." " 2:N:0:.."); a read 2 without its read 1 gets a faked "N" mate,
 * and -- as in the reference -- a read 1 without its read 2 gets one too and ENDS the file (src/merge_reads.cpp:446-461: the
 * end-of-file test looks at the already swapped-out name). Names that carry no pair marker are not supported.
 * Plain (uncompressed) files of 4 MB or more are read by several threads (KC_HOST_THREADS, default = cores, at most 32): byte
 * ranges are indexed in parallel, ONE thread runs the pairing logic over the record index, the characters are copied in
 * parallel -- with exactly the streaming reader's result (anything unusual falls back to it; tests/test_merge.py holds the two
 * to each other). 630 MB of FASTQ: 0.82 s streaming, 0.19 s with 8 threads. */
typedef struct {
  char *seq, *qual;
  uint64_t n_bytes;
  uint64_t *offsets;
  uint16_t *lens;
  uint64_t n_reads; /* 2 x pairs */
  uint64_t missing_read1, missing_read2;
} kc_pair_set;
int kc_fastq_read_pairs(const char *path1, const char *path2, int paired, int qual_offset, uint64_t max_pairs, kc_pair_set *out);
void kc_pair_set_free(kc_pair_set *ps);

#ifdef __cplusplus
}
#endif
#endif /* KC_SYNTH_H */
