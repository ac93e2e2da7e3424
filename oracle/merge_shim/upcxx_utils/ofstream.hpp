// TEST INFRASTRUCTURE (oracle/) -- dist_ofstream as a plain local file (one rank).
#pragma once
#include <fstream>
#include <string>
#include <upcxx/upcxx.hpp>
namespace upcxx_utils {
class dist_ofstream : public std::ofstream {
 public:
  explicit dist_ofstream(const std::string &fname, bool append = false) : std::ofstream(fname, append ? std::ios::app : std::ios::out) {}
  std::streamoff get_last_known_tellp() { return (std::streamoff)tellp(); }
  upcxx::future<> close_async() {
    close();
    return {};
  }
  upcxx::future<> report_timings() { return {}; }
};
}  // namespace upcxx_utils
