// TEST INFRASTRUCTURE (oracle/) -- single-node stand-in for the UPC++ API surface that the
// reference's kcount sources touch. Written from scratch for this repo: it
// lets /root/reference/src/{kmer.cpp,kcount/kcount.cpp,kcount/kcount_cpu.cpp,kcount/kmer_dht.cpp}
// compile UNMODIFIED into oracle/_ref. Ranks are forked processes sharing one anonymous mmap
// region (see shim_world.hpp); collectives are shared-memory slots + a sense-reversing barrier.
// Nothing under oracle/ is ever on the product path.
#pragma once
#include <cstdint>
#include <cstring>
#include <functional>
// the real upcxx.hpp pulls in most of the standard library; the reference sources lean on that
#include <algorithm>
#include <array>
#include <cassert>
#include <chrono>
#include <iomanip>
#include <iostream>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <type_traits>
#include <utility>
#include "../shim_world.hpp"

#define UPCXX_SERIALIZED_FIELDS(...)

namespace upcxx {

using intrank_t = int;

inline intrank_t rank_me() { return shim::world().rank; }
inline intrank_t rank_n() { return shim::world().nranks; }
inline void barrier() { shim::world().barrier(); }
inline void progress() {}

struct team {
  intrank_t rank_n() const { return shim::world().nranks; }
  intrank_t rank_me() const { return shim::world().rank; }
};
inline team &world() {
  static team t;
  return t;
}
inline team &local_team() { return world(); }

template <typename T>
struct global_ptr {
  T *p = nullptr;
  global_ptr() = default;
  global_ptr(std::nullptr_t) {}
};

// a future that is always ready
template <typename T>
struct future {
  T v;
  T wait() const { return v; }
  template <typename F>
  auto then(F &&f) const {
    using R = decltype(f(v));
    if constexpr (std::is_void_v<R>) {
      f(v);
      return future<int>{0};
    } else {
      return future<R>{f(v)};
    }
  }
};

struct op_add_t {
  template <typename T>
  T operator()(T a, T b) const { return a + b; }
};
struct op_max_t {
  template <typename T>
  T operator()(T a, T b) const { return a < b ? b : a; }
};
struct op_min_t {
  template <typename T>
  T operator()(T a, T b) const { return b < a ? b : a; }
};
inline constexpr op_add_t op_fast_add{};
inline constexpr op_max_t op_fast_max{};
inline constexpr op_min_t op_fast_min{};

template <typename T, typename Op>
future<T> reduce_all(T v, Op op) {
  static_assert(std::is_trivially_copyable_v<T> && sizeof(T) <= 16, "shim reduce: small POD only");
  return future<T>{shim::world().allreduce(v, op)};
}
// every rank gets the reduced value (a superset of what rank `root` is promised)
template <typename T, typename Op>
future<T> reduce_one(T v, Op op, intrank_t /*root*/) {
  return reduce_all(v, op);
}

// rpc: only reached from KmerDHT::kmer_exists / get_kmer_count (kmer_dht.cpp:216-245), which nothing in
// the proxy calls; runs the callable locally (dist_object arguments are passed by reference).
template <typename F, typename... Args>
auto rpc(intrank_t, F &&f, Args &&...args) {
  using R = decltype(f(std::forward<Args>(args)...));
  return future<R>{f(std::forward<Args>(args)...)};
}

// one instance per process; operator-> / operator* reach the local value
template <typename T>
class dist_object {
  T val;

 public:
  dist_object(T v) : val(std::move(v)) {}
  dist_object(team &, T v) : val(std::move(v)) {}
  template <typename... Args>
  dist_object(team &, Args &&...args) : val(std::forward<Args>(args)...) {}
  dist_object(const dist_object &) = delete;
  T *operator->() { return &val; }
  T &operator*() { return val; }
  const T *operator->() const { return &val; }
  const T &operator*() const { return val; }
};

}  // namespace upcxx

// upcxx_utils headers leak these names into the global namespace and kmer_dht.hpp relies on it
using upcxx::dist_object;
using upcxx::global_ptr;
