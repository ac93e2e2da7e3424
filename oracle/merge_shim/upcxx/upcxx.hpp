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
inline void barrier(const team &) {}

template <typename T>
struct global_ptr {
  T *p = nullptr;
  global_ptr() = default;
  global_ptr(std::nullptr_t) {}
};

// ---- futures with shared state (single thread): a future handed out before its promise is fulfilled sees the value later,
// and .then() callbacks run at fulfilment -- the reference's FastqReader chains its open / seek steps that way.
template <typename... T>
struct future;
template <typename X>
struct is_future : std::false_type {};
template <typename... T>
struct is_future<future<T...>> : std::true_type {};

template <typename... T>
struct fstate {
  std::tuple<T...> v{};
  bool ready = false;
  std::vector<std::function<void()>> conts;
  void set(std::tuple<T...> t) {
    v = std::move(t);
    ready = true;
    auto c = std::move(conts);
    conts.clear();
    for (auto &f : c) f();
  }
};

template <typename... T>
struct future {
  std::shared_ptr<fstate<T...>> s;
  future() : s(std::make_shared<fstate<T...>>()) { s->ready = true; }  // (a default future is ready: value-initialised results)
  struct not_ready_t {};
  explicit future(not_ready_t) : s(std::make_shared<fstate<T...>>()) {}
  template <typename... U, typename = std::enable_if_t<sizeof...(U) == sizeof...(T) && (sizeof...(U) > 0) &&
                                                       !(is_future<std::decay_t<U>>::value || ...)>>
  future(U &&...a) : s(std::make_shared<fstate<T...>>()) {
    s->v = std::tuple<T...>(std::forward<U>(a)...);
    s->ready = true;
  }
  explicit future(std::tuple<T...> t, int) : s(std::make_shared<fstate<T...>>()) {
    s->v = std::move(t);
    s->ready = true;
  }
  bool ready() const { return s->ready; }
  bool is_ready() const { return s->ready; }
  void need_ready() const {
    if (!s->ready) {
      std::cerr << "merge_shim: wait() on a future that nothing can fulfil any more (single thread)" << std::endl;
      abort();
    }
  }
  auto wait() const {
    need_ready();
    if constexpr (sizeof...(T) == 0)
      return;
    else if constexpr (sizeof...(T) == 1)
      return std::get<0>(s->v);
    else
      return s->v;
  }
  auto result() const { return wait(); }
  template <typename F>
  auto then(F &&f) const {
    using R = decltype(std::apply(f, s->v));
    auto src = s;
    if constexpr (std::is_void_v<R>) {
      future<> rf{typename future<>::not_ready_t{}};
      auto run = [src, f, rs = rf.s]() mutable {
        std::apply(f, src->v);
        rs->set({});
      };
      if (src->ready) run(); else src->conts.push_back(run);
      return rf;
    } else if constexpr (is_future<R>::value) {
      R rf{typename R::not_ready_t{}};
      auto run = [src, f, rs = rf.s]() mutable {
        R inner = std::apply(f, src->v);
        auto is = inner.s;
        auto fin = [is, rs]() { rs->set(is->v); };
        if (is->ready) fin(); else is->conts.push_back(fin);
      };
      if (src->ready) run(); else src->conts.push_back(run);
      return rf;
    } else {
      future<R> rf{typename future<R>::not_ready_t{}};
      auto run = [src, f, rs = rf.s]() mutable { rs->set(std::tuple<R>(std::apply(f, src->v))); };
      if (src->ready) run(); else src->conts.push_back(run);
      return rf;
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
  future<A..., B...> rf{typename future<A..., B...>::not_ready_t{}};
  auto as = a.s;
  auto bs = b.s;
  auto fin = [as, bs, rs = rf.s]() {
    if (as->ready && bs->ready && !rs->ready) rs->set(std::tuple_cat(as->v, bs->v));
  };
  if (as->ready && bs->ready) {
    fin();
  } else {
    if (!as->ready) as->conts.push_back(fin);
    if (!bs->ready) bs->conts.push_back(fin);
  }
  return rf;
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

// promise: dependency count starts at 1 (UPC++); fulfill_result supplies the values and takes one dependency away
template <typename... T>
struct promise {
  future<T...> f{typename future<T...>::not_ready_t{}};
  std::shared_ptr<long> deps = std::make_shared<long>(1);
  std::shared_ptr<std::tuple<T...>> val = std::make_shared<std::tuple<T...>>();
  promise() = default;
  explicit promise(long n) { *deps = n; }
  future<T...> get_future() const { return f; }
  void dec(long n) {
    *deps -= n;
    if (*deps <= 0 && !f.s->ready) f.s->set(*val);
  }
  template <typename... U>
  void fulfill_result(U &&...a) {
    *val = std::tuple<T...>(std::forward<U>(a)...);
    dec(1);
  }
  void fulfill_anonymous(long n) { dec(n); }
  void require_anonymous(long n) { *deps += n; }
  future<T...> finalize() {
    dec(1);
    return f;
  }
};

promise(int)->promise<>;
promise(long)->promise<>;

struct persona {
  bool active_with_caller() const { return true; }
};
inline persona &master_persona() {
  static persona p;
  return p;
}
inline persona &current_persona() { return master_persona(); }

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
future<> broadcast(T *, size_t, intrank_t) {  // (in place; one rank: the buffer already holds the value)
  return {};
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
template <typename F, typename... Args>
auto rpc(const team &, intrank_t r, F &&f, Args &&...args) {
  return rpc(r, std::forward<F>(f), std::forward<Args>(args)...);
}
template <typename F, typename... Args>
void rpc_ff(const team &, intrank_t r, F &&f, Args &&...args) {
  rpc_ff(r, std::forward<F>(f), std::forward<Args>(args)...);
}

}  // namespace upcxx

using upcxx::dist_object;
using upcxx::global_ptr;
