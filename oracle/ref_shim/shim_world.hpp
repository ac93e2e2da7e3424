// TEST INFRASTRUCTURE (oracle/) -- the "world" behind the UPC++ stand-in: P forked processes, one
// shared anonymous mapping holding a barrier, reduction slots and per-(src,dst) supermer mailboxes.
#pragma once
#include <sched.h>
#include <sys/mman.h>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace shim {

struct Shared {
  std::atomic<int> bar_count;
  std::atomic<int> bar_sense;
  alignas(64) unsigned char slots[256][16];  // one reduction slot per rank
};

struct Mailbox {  // append-only byte log written by one src for one dst
  std::atomic<uint64_t> len;
  uint64_t cap;
  unsigned char *data;
};

struct World {
  int rank = 0;
  int nranks = 1;
  Shared *sh = nullptr;
  Mailbox *boxes = nullptr;  // [src * nranks + dst], in shared memory
  unsigned char *gather = nullptr;  // [nranks * kGatherBytes]: all-gather of small records (the B200 glue's IPC handles)
  static constexpr size_t kGatherBytes = 2048;
  int local_sense = 0;

  static void *shared_alloc(size_t bytes) {
    void *p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (p == MAP_FAILED) {
      perror("mmap");
      abort();
    }
    return p;
  }

  // call in the parent BEFORE fork
  void init_shared(int P, uint64_t box_cap_bytes) {
    nranks = P;
    sh = (Shared *)shared_alloc(sizeof(Shared));
    new (sh) Shared();
    sh->bar_count = 0;
    sh->bar_sense = 0;
    gather = (unsigned char *)shared_alloc(kGatherBytes * P);
    boxes = (Mailbox *)shared_alloc(sizeof(Mailbox) * P * P);
    for (int i = 0; i < P * P; i++) {
      new (&boxes[i]) Mailbox();
      boxes[i].len = 0;
      boxes[i].cap = box_cap_bytes;
      boxes[i].data = P > 1 ? (unsigned char *)shared_alloc(box_cap_bytes) : nullptr;
    }
  }

  void barrier() {
    if (nranks == 1) return;
    local_sense ^= 1;
    if (sh->bar_count.fetch_add(1) == nranks - 1) {
      sh->bar_count = 0;
      sh->bar_sense.store(local_sense);
    } else {
      while (sh->bar_sense.load() != local_sense) sched_yield();
    }
  }

  template <typename T, typename Op>
  T allreduce(T v, Op op) {
    if (nranks == 1) return v;
    memcpy(sh->slots[rank], &v, sizeof(T));
    barrier();
    T acc;
    memcpy(&acc, sh->slots[0], sizeof(T));
    for (int r = 1; r < nranks; r++) {
      T x;
      memcpy(&x, sh->slots[r], sizeof(T));
      acc = op(acc, x);
    }
    barrier();
    return acc;
  }

  // every rank contributes `bytes` bytes; all[r * bytes ...] = rank r's record
  void allgather_bytes(const void *mine, size_t bytes, void *all) {
    if (bytes > kGatherBytes) {
      fprintf(stderr, "shim allgather: record of %zu bytes\n", bytes);
      abort();
    }
    if (nranks == 1) {
      memcpy(all, mine, bytes);
      return;
    }
    memcpy(gather + (size_t)rank * kGatherBytes, mine, bytes);
    barrier();
    for (int r = 0; r < nranks; r++) memcpy((unsigned char *)all + (size_t)r * bytes, gather + (size_t)r * kGatherBytes, bytes);
    barrier();
  }

  void post(int dst, const void *bytes, uint64_t n) {
    Mailbox &b = boxes[rank * nranks + dst];
    uint64_t at = b.len.load(std::memory_order_relaxed);
    if (at + n > b.cap) {
      fprintf(stderr, "shim mailbox %d->%d overflow (cap %lu)\n", rank, dst, (unsigned long)b.cap);
      abort();
    }
    memcpy(b.data + at, bytes, n);
    b.len.store(at + n, std::memory_order_release);
  }
};

inline World &world() {
  static World w;
  return w;
}

}  // namespace shim
