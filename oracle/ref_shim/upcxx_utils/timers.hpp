// TEST INFRASTRUCTURE (oracle/) -- timers and min/sum/max reductions with the reference's names.
#pragma once
#include <chrono>
#include <sstream>
#include <string>
#include <upcxx/upcxx.hpp>
#include "log.hpp"

#define CLOCK_NOW std::chrono::high_resolution_clock::now

namespace upcxx_utils {

class BaseTimer {
  std::chrono::time_point<std::chrono::high_resolution_clock> t0;
  double elapsed = 0;

 public:
  BaseTimer() = default;
  BaseTimer(const std::string &) {}
  void start() { t0 = CLOCK_NOW(); }
  void stop() { elapsed += std::chrono::duration<double>(CLOCK_NOW() - t0).count(); }
  double get_elapsed() const { return elapsed; }
  void inc_elapsed(double s) { elapsed += s; }
};

class IntermittentTimer : public BaseTimer {
 public:
  IntermittentTimer(const std::string & = "", const std::string & = "") {}
};

struct BarrierTimer {
  BarrierTimer(const std::string &) { upcxx::barrier(); }
  ~BarrierTimer() { upcxx::barrier(); }
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
  r.my = v;
  r.min = upcxx::reduce_all(v, upcxx::op_fast_min).wait();
  r.max = upcxx::reduce_all(v, upcxx::op_fast_max).wait();
  r.sum = upcxx::reduce_all(v, upcxx::op_fast_add).wait();
  r.avg = (double)r.sum / upcxx::rank_n();
  return {r};
}
template <typename T>
upcxx::future<MinSumMax<T>> min_sum_max_reduce_one(T v, int = 0) {
  return min_sum_max_reduce_all(v);
}

}  // namespace upcxx_utils
