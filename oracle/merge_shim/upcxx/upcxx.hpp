// TEST INFRASTRUCTURE (oracle/) -- single-process stand-in for the UPC++ API surface that the reference's
// src/merge_reads.cpp (and the headers it pulls in: fastq.hpp, packed_reads.hpp, adapters.hpp, utils.hpp) touches, so that
// the file compiles UNMODIFIED into oracle/_ref/merge_ref. Written from scratch for this repo. One rank; every future is
// ready when it is made, so .then() runs its callback at once. Nothing under oracle/ is ever on the product path.
#pragma once
#include <algorithm>
#include <array>
#include <cassert>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <tuple>
#include <type_traits>
#include <unordered_map>
#include <utility>
#include <vector>

#define UPCXX_SERIALIZED_FIELDS(...)
#define UPCXX_SERIALIZED_VALUES(...)

namespace upcxx {

using intrank_t = int;
inline intrank_t rank_me() { return 0; }
inline intrank_t rank_n() { return 1; }
inline void barrier() {}
inline void progress() {}
inline void discharge() {}
inline bool in_progress() { return false; }

struct team {
  intrank_t rank_n() const { return 1; }
  intrank_t rank_me() const { return 0; }
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

template <typename... T>
struct future;
template <typename X>
struct is_future : std::false_type {};
template <typename... T>
struct is_future<future<T...>> : std::true_type {};

template <typename... T>
struct future {
  std::tuple<T...> v;
  future() = default;
  template <typename... U, typename = std::enable_if_t<sizeof...(U) == sizeof...(T) && (sizeof...(U) > 0)>>
  future(U &&...a) : v(std::forward<U>(a)...) {}
  explicit future(std::tuple<T...> t, int) : v(std::move(t)) {}
  bool ready() const { return true; }
  bool is_ready() const { return true; }
  auto wait() const {
    if constexpr (sizeof...(T) == 0)
      return;
    else if constexpr (sizeof...(T) == 1)
      return std::get<0>(v);
    else
      return v;
  }
  auto result() const { return wait(); }
  template <typename F>
  auto then(F &&f) const {
    using R = decltype(std::apply(f, v));
    if constexpr (std::is_void_v<R>) {
      std::apply(f, v);
      return future<>();
    } else if constexpr (is_future<R>::value) {
      return std::apply(f, v);
    } else {
      return future<R>(std::apply(f, v));
    }
  }
};

inline future<> make_future() { return {}; }
template <typename T>
future<std::decay_t<T>> make_future(T &&v) {
  return future<std::decay_t<T>>(std::forward<T>(v));
}

template <typename... A, typename... B>
future<A..., B...> shim_cat2(const future<A...> &a, const future<B...> &b) {
  return future<A..., B...>(std::tuple_cat(a.v, b.v), 0);
}
inline future<> when_all() { return {}; }
template <typename F0>
F0 when_all(F0 f) {
  return f;
}
template <typename F0, typename F1, typename... Rest>
auto when_all(F0 a, F1 b, Rest... rest) {
  return when_all(shim_cat2(a, b), rest...);
}

template <typename... T>
struct promise {
  std::tuple<T...> v{};
  future<T...> get_future() const { return future<T...>(v, 0); }
  template <typename... U>
  void fulfill_result(U &&...a) {
    v = std::tuple<T...>(std::forward<U>(a)...);
  }
  void fulfill_anonymous(int) {}
  void require_anonymous(int) {}
  future<T...> finalize() { return get_future(); }
};
template <>
struct promise<> {
  future<> get_future() const { return {}; }
  void fulfill_anonymous(int) {}
  void require_anonymous(int) {}
  void fulfill_result() {}
  future<> finalize() { return {}; }
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
future<T> reduce_all(T v, Op) {
  return future<T>(v);
}
template <typename T, typename Op>
future<T> reduce_one(T v, Op, intrank_t) {
  return future<T>(v);
}
template <typename T>
future<T> broadcast(T v, intrank_t) {
  return future<T>(v);
}

template <typename T>
class dist_object {
  T val;

 public:
  dist_object(T v) : val(std::move(v)) {}
  dist_object(team &, T v) : val(std::move(v)) {}
  template <typename... Args>
  dist_object(team &, Args &&...args) : val(std::forward<Args>(args)...) {}
  dist_object(const dist_object &) = delete;
  dist_object(dist_object &&) = default;
  T *operator->() { return &val; }
  T &operator*() { return val; }
  const T *operator->() const { return &val; }
  const T &operator*() const { return val; }
  future<T> fetch(intrank_t) const { return future<T>(val); }
};

// one rank: an rpc can only target this process
template <typename F, typename... Args>
auto rpc(intrank_t, F &&f, Args &&...args) {
  using R = decltype(f(std::forward<Args>(args)...));
  if constexpr (std::is_void_v<R>) {
    f(std::forward<Args>(args)...);
    return future<>();
  } else {
    return future<R>(f(std::forward<Args>(args)...));
  }
}
template <typename F, typename... Args>
void rpc_ff(intrank_t, F &&f, Args &&...args) {
  f(std::forward<Args>(args)...);
}

}  // namespace upcxx

using upcxx::dist_object;
using upcxx::global_ptr;
