/*
 * TEST INFRASTRUCTURE -- CPU oracle for the k-mer analysis hot path.
 *
 * A plain-C restatement of the reference's CPU kcount (src/kcount/kcount_cpu.cpp, src/kmer.cpp,
 * src/hash_funcs.c, src/kcount/kmer_dht.cpp, src/kcount/kcount.cpp), each function citing the lines it
 * follows. Pinned against the reference's own sources compiled here (oracle/_ref/kcount_ref, see
 * oracle/Makefile and tests/golden/make_golden.py): the reference ships no tests or golden vectors
 * for this path, so the pin is "outputs of the reference itself run here".
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library, and only as the checker or the timed CPU baseline -- never the product path.
 */
#ifndef KCOUNT_ORACLE_H
#define KCOUNT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct kco_ctx kco_ctx;

/* n_ranks emulates the owner partitioning (one table per rank, supermers routed by
 * KmerDHT::get_kmer_target_rank); slots_per_rank = table capacity request (rounded up to a prime). */
kco_ctx *kco_create(int k, int m, int qual_offset, int qual_cutoff, int dmin_thres, double dyn_min_depth, int n_ranks,
                    uint64_t slots_per_rank);
void kco_destroy(kco_ctx *c);
int kco_minimizer_len(const kco_ctx *c);
int kco_key_words(const kco_ctx *c);

/* count_kmers() read loop -> process_seq -> add_supermer -> insert_supermer, for n reads */
int kco_add_reads(kco_ctx *c, const uint8_t *bytes, const uint64_t *offsets, const uint16_t *lens, uint64_t n);
/* insert_supermer() entry: a mixed-case supermer string (lower case = low quality), as the wire carries it */
int kco_insert_supermer(kco_ctx *c, int rank, const char *seq, int len, int count);
/* insert_into_local_hashtable(): purge, extension resolution */
int kco_finalize(kco_ctx *c);

uint64_t kco_num_supermers(const kco_ctx *c, int rank); /* KmerDHT::get_num_supermer_inserts */
uint64_t kco_num_entries(const kco_ctx *c, int rank);   /* table entries incl. singletons; rank -1 = all */
uint64_t kco_num_dropped(const kco_ctx *c);
uint64_t kco_counted_kmers(const kco_ctx *c);
/* results, sorted by key words; rank = -1 -> union over ranks. extended = entries with count >= 2 incl.
 * X/F plus the 8 ext counters (dump D2); otherwise survivors only (dump D1). */
uint64_t kco_result_size(const kco_ctx *c, int extended, int rank);
int kco_result_copy(const kco_ctx *c, int extended, int rank, uint64_t *keys, uint16_t *counts, char *left, char *right,
                    uint16_t *ext8);
uint64_t kco_result_checksum(const kco_ctx *c, int extended, int rank);
double kco_avg_kmer_count(const kco_ctx *c);

/* supermers of one read as process_seq cuts them: for each, target rank, first k-mer index (0-based
 * position in the read of the k-mer's first base) and number of k-mers. Returns the count (<= cap). */
int kco_read_supermers(const kco_ctx *c, const uint8_t *bytes, int len, int *targets, int *first_kmer, int *n_kmers, int cap);

/* building blocks, exported for unit tests */
uint64_t kco_quick_hash(uint64_t v);
uint64_t kco_murmur64(const void *key, uint32_t len);
void kco_pack_kmer(const char *s, int k, uint64_t *words /* ceil(k/32), zero padded */);
void kco_revcomp(const uint64_t *in, int k, uint64_t *out);
int kco_kmer_less(const uint64_t *a, const uint64_t *b, int n_words);
uint64_t kco_minimizer(const uint64_t *kmer, int k, int m); /* greatest least-complement m-mer, left aligned */
int kco_owner(const uint64_t *kmer, int k, int m, int n_ranks);
char kco_get_ext(const uint16_t counts[4], uint16_t count, int dmin_thres, double dyn_min_depth);
uint64_t kco_entry_mix(const uint64_t *key, int n_words, uint16_t count, char left, char right, const uint16_t *ext8);

#ifdef __cplusplus
}
#endif
#endif
