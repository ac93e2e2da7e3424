// TEST INFRASTRUCTURE (oracle/) -- timers and the promise-reduce helper with the reference's names (one rank: every
// reduction returns its own argument).
#pragma once
#include <chrono>
#include <sstream>
#include <string>
#include <upcxx/upcxx.hpp>
#include "log.hpp"

#define CLOCK_NOW std::chrono::high_resolution_clock::now

namespace upcxx_utils {

inline std::string get_basename(const std::string &fname) {
  size_t i = fname.rfind('/');
  return i == std::string::npos ? fname : fname.substr(i + 1);
}

class BaseTimer {
  mutable std::chrono::time_point<std::chrono::high_resolution_clock> t0;
  mutable double elapsed = 0;

 public:
  BaseTimer() = default;
  BaseTimer(const std::string &) {}
  void start() const { t0 = CLOCK_NOW(); }
  void stop() const { elapsed += std::chrono::duration<double>(CLOCK_NOW() - t0).count(); }
  double get_elapsed() const { return elapsed; }
  void inc_elapsed(double s) { elapsed += s; }
  void done() const {}
  void done_all() const {}
  void clear() { elapsed = 0; }
  double get_elapsed_since_start() const { return std::chrono::duration<double>(CLOCK_NOW() - t0).count(); }
  void done_all_async() const {}
};

class IntermittentTimer : public BaseTimer {
 public:
  IntermittentTimer(const std::string & = "", const std::string & = "") {}
};
class AsyncTimer : public BaseTimer {
 public:
  AsyncTimer(const std::string & = "") {}
  upcxx::future<> initiate_construct_reduction() const { return {}; }
  upcxx::future<> initiate_start_reduction() const { return {}; }
  upcxx::future<> initiate_stop_reduction() const { return {}; }
  void report(const std::string & = "") const {}
};
class Timer : public BaseTimer {
 public:
  Timer(const std::string & = "") {}
  upcxx::future<> initiate_entrance_reduction() { return {}; }
  upcxx::future<> initiate_exit_reduction() { return {}; }
};
struct BarrierTimer {
  BarrierTimer(const std::string &, bool = true) {}
  upcxx::future<> initiate_entrance_barrier() { return {}; }
  upcxx::future<> initiate_exit_barrier() { return {}; }
};

template <typename T>
struct MinSumMax {
  T my = 0, min = 0, sum = 0, max = 0;
  double avg = 0;
  std::string to_string() const {
    std::ostringstream os;
    os << min << "/" << my << "/" << avg << "/" << max;
    return os.str();
  }
};
template <typename T>
upcxx::future<MinSumMax<T>> min_sum_max_reduce_all(T v) {
  MinSumMax<T> r;
  r.my = r.min = r.max = r.sum = v;
  r.avg = (double)v;
  return upcxx::future<MinSumMax<T>>(r);
}
template <typename T>
upcxx::future<MinSumMax<T>> min_sum_max_reduce_one(T v, int = 0) {
  return min_sum_max_reduce_all(v);
}

struct PromiseReduce {
  template <typename T, typename Op>
  upcxx::future<T> reduce_all(T v, Op) {
    return upcxx::future<T>(v);
  }
  template <typename T, typename Op>
  upcxx::future<T> reduce_one(T v, Op, int = 0) {
    return upcxx::future<T>(v);
  }
  template <typename T>
  upcxx::future<MinSumMax<T>> msm_reduce_one(T v, int = 0) {
    return min_sum_max_reduce_all(v);
  }
  template <typename T>
  upcxx::future<MinSumMax<T>> msm_reduce_all(T v) {
    return min_sum_max_reduce_all(v);
  }
  upcxx::future<> fulfill() { return {}; }
};
struct Timings {
  static void set_pending(upcxx::future<>) {}
  static upcxx::future<> get_pending() { return {}; }
  static void wait_pending() {}
  static PromiseReduce &get_promise_reduce() {
    static PromiseReduce pr;
    return pr;
  }
};

}  // namespace upcxx_utils
