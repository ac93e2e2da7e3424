/* TEST INFRASTRUCTURE (oracle/) -- reader for the packed-reads exchange file shared by the oracle, the
 * reference build and the product's generator (same layout as kc_reads_save in include/kcount_b200.h):
 *   u64 magic "KCREADS1", u64 n_reads, u64 n_bytes, u64 reserved,
 *   u64 offsets[n_reads], u16 lens[n_reads] (padded to 8 bytes), u8 bytes[n_bytes]
 * bytes follow the reference's PackedRead encoding (src/packed_reads.cpp:99-126). */
#ifndef KC_READS_IO_H
#define KC_READS_IO_H
#include <fcntl.h>
#include <stdint.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

typedef struct {
  uint64_t n_reads, n_bytes;
  const uint64_t *offsets;
  const uint16_t *lens;
  const unsigned char *bytes;
  void *map;
  uint64_t map_len;
} kc_reads_file;

static inline int kc_reads_file_map(const char *path, kc_reads_file *rf) {
  int fd = open(path, O_RDONLY);
  if (fd < 0) return -1;
  struct stat st;
  if (fstat(fd, &st) != 0) {
    close(fd);
    return -1;
  }
  void *m = mmap(0, st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (m == MAP_FAILED) return -1;
  const uint64_t *h = (const uint64_t *)m;
  if (h[0] != 0x3153444145524b43ULL) return -2; /* "KCREADS1" little-endian */
  rf->n_reads = h[1];
  rf->n_bytes = h[2];
  rf->offsets = h + 4;
  rf->lens = (const uint16_t *)(rf->offsets + rf->n_reads);
  uint64_t lens_bytes = (rf->n_reads * 2 + 7) & ~7ull;
  rf->bytes = (const unsigned char *)rf->lens + lens_bytes;
  rf->map = m;
  rf->map_len = st.st_size;
  return 0;
}
#endif
