// TEST INFRASTRUCTURE (oracle/) -- stand-in for upcxx_utils::ThreeTierAggrStore<T> (source absent from
// the reference tree; behaviour inferred from the call sites kmer_dht.cpp:144-151,249,255,158).
// update(target, elem): target == me -> run the update function now; otherwise serialise the element
// into the (me -> target) shared-memory mailbox. flush_updates(): barrier, then drain every mailbox
// addressed to me through the update function, then barrier. Only T = Supermer{string seq; u16 count}
// is used by the path, via the generic shim_put/shim_get hooks below.
#pragma once
#include <functional>
#include <string>
#include <upcxx/upcxx.hpp>

namespace upcxx_utils {

template <typename T>
class ThreeTierAggrStore {
  std::function<void(T)> func;

 public:
  void set_size(const std::string &, int64_t, int = 0, int64_t = 0) {}
  void set_update_func(std::function<void(T)> f) { func = std::move(f); }
  void clear() {}

  void update(int target, const T &elem) {
    auto &w = shim::world();
    if (target == w.rank) {
      func(elem);
      return;
    }
    // wire format: u16 seq_len, u16 count, seq bytes
    uint16_t hdr[2] = {(uint16_t)elem.seq.size(), (uint16_t)elem.count};
    unsigned char buf[4 + 65536];
    memcpy(buf, hdr, 4);
    memcpy(buf + 4, elem.seq.data(), elem.seq.size());
    w.post(target, buf, 4 + elem.seq.size());
  }

  void flush_updates() {
    auto &w = shim::world();
    if (w.nranks == 1) return;
    w.barrier();
    for (int src = 0; src < w.nranks; src++) {
      if (src == w.rank) continue;
      shim::Mailbox &b = w.boxes[src * w.nranks + w.rank];
      uint64_t n = b.len.load(std::memory_order_acquire), at = 0;
      while (at < n) {
        uint16_t hdr[2];
        memcpy(hdr, b.data + at, 4);
        T elem;
        elem.seq.assign((const char *)b.data + at + 4, hdr[0]);
        elem.count = hdr[1];
        func(elem);
        at += 4 + hdr[0];
      }
    }
    w.barrier();
  }
};

}  // namespace upcxx_utils
