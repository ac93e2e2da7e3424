// TEST INFRASTRUCTURE (oracle/) -- the logging macros of oracle/ref_shim over this directory's upcxx stand-in, plus the
// names the real upcxx_utils headers leak into the global namespace (the reference sources rely on them) and the small
// file helpers that live in the real log / ofstream headers.
#pragma once
#include <sys/stat.h>
#include <fstream>
#include <sstream>
#include "../../ref_shim/upcxx_utils/log.hpp"

using std::ifstream;
using std::istream;
using std::ofstream;
using std::ostream;
using std::ostringstream;
using std::stringstream;
using std::string;
using std::to_string;
using std::vector;
using std::pair;
using std::min;
using std::max;
using upcxx::barrier;
using upcxx::broadcast;
using upcxx::future;
using upcxx::make_future;
using upcxx::op_fast_add;
using upcxx::op_fast_max;
using upcxx::op_fast_min;
using upcxx::progress;
using upcxx::promise;
using upcxx::rank_me;
using upcxx::rank_n;
using upcxx::reduce_all;
using upcxx::reduce_one;
using upcxx::when_all;
using upcxx::world;

#define SDIE(...) DIE(__VA_ARGS__)
#define DBGLOG(...) \
  do {              \
  } while (0)
#define SOUT(...) upcxx_utils::shim_log(true, __VA_ARGS__)
#define OUT(...) upcxx_utils::shim_log(false, __VA_ARGS__)
#define INFO(...) upcxx_utils::shim_log(false, __VA_ARGS__)

namespace upcxx_utils {
inline bool file_exists(const std::string &fname) {
  struct stat st;
  return stat(fname.c_str(), &st) == 0;
}
inline std::string remove_file_ext(const std::string &fname) {
  size_t i = fname.find_last_of('.');
  return i == std::string::npos ? fname : fname.substr(0, i);
}
inline void replace_spaces(std::string &s, char with = '_') {
  for (auto &c : s)
    if (c == ' ') c = with;
}
inline size_t get_file_size(const std::string &fname) {
  struct stat st;
  if (stat(fname.c_str(), &st) != 0) return 0;
  return (size_t)st.st_size;
}
}  // namespace upcxx_utils
