// TEST INFRASTRUCTURE (oracle/) -- inert ProgressBar.
#pragma once
#include <cstdint>
#include <string>
#include <upcxx/upcxx.hpp>
namespace upcxx_utils {
struct ProgressBar {
  ProgressBar(int64_t, const std::string &) {}
  bool update(int64_t = -1) { return false; }
  void done() {}
  upcxx::future<> set_done() { return {}; }
};
}  // namespace upcxx_utils
