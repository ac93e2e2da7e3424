// TEST INFRASTRUCTURE (oracle/) -- logging macros with the names the reference sources use.
// Output goes to stderr only when ORACLE_REF_VERBOSE is set; DIE always aborts.
#pragma once
#include <cstdlib>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <upcxx/upcxx.hpp>

namespace upcxx_utils {

inline bool shim_verbose() {
  static int v = -1;
  if (v < 0) v = getenv("ORACLE_REF_VERBOSE") ? 1 : 0;
  return v == 1;
}

template <typename... Args>
inline std::string shim_cat(Args &&...args) {
  std::ostringstream os;
  (os << ... << args);
  return os.str();
}

template <typename... Args>
inline void shim_log(bool rank0_only, Args &&...args) {
  if (!shim_verbose()) return;
  if (rank0_only && upcxx::rank_me() != 0) return;
  std::cerr << shim_cat(std::forward<Args>(args)...);
}

template <typename... Args>
[[noreturn]] inline void shim_die(const char *file, int line, Args &&...args) {
  std::cerr << "DIE [" << file << ":" << line << "] rank " << upcxx::rank_me() << ": " << shim_cat(std::forward<Args>(args)...)
            << std::endl;
  abort();
}

}  // namespace upcxx_utils

// string helpers (the real ones live in upcxx_utils' log/ofstream headers)
namespace upcxx_utils {
inline std::string get_size_str(double bytes) {
  std::ostringstream os;
  os << std::fixed << std::setprecision(2) << bytes / (1024.0 * 1024.0) << "MB";
  return os.str();
}
inline std::string perc_str(int64_t num, int64_t tot) {
  std::ostringstream os;
  os << num << " (" << std::fixed << std::setprecision(2) << (tot ? 100.0 * num / tot : 0.0) << "%)";
  return os.str();
}
inline void get_rank_path(std::string &fname, int rank) { fname = "rank" + std::to_string(rank) + "-" + fname; }
}  // namespace upcxx_utils

// the real header leaks this name; src/stage_timers.hpp relies on it
using std::string;

#define KNORM ""
#define KLMAGENTA ""
#define KBLUE ""
#define KLBLUE ""
#define KLGREEN ""
#define KLRED ""
#define KRED ""
#define KLCYAN ""

#ifndef __FILENAME__
#define __FILENAME__ (__builtin_strrchr(__FILE__, '/') ? __builtin_strrchr(__FILE__, '/') + 1 : __FILE__)
#endif
#define __FILEFUNC__ (std::string(__FILENAME__) + ":" + __func__)

#define SLOG(...) upcxx_utils::shim_log(true, __VA_ARGS__)
#define SLOG_VERBOSE(...) upcxx_utils::shim_log(true, __VA_ARGS__)
#define LOG(...) upcxx_utils::shim_log(false, __VA_ARGS__)
#define NET_LOG(...) \
  do {               \
  } while (0)
#define DBG(...) \
  do {           \
  } while (0)
#define DBG_VERBOSE(...) \
  do {                   \
  } while (0)
#define WARN(...) (std::cerr << "WARN rank " << upcxx::rank_me() << ": " << upcxx_utils::shim_cat(__VA_ARGS__) << std::endl)
#define SWARN(...)                                                                                   \
  do {                                                                                               \
    if (!upcxx::rank_me()) std::cerr << "SWARN: " << upcxx_utils::shim_cat(__VA_ARGS__) << std::endl; \
  } while (0)
#define DIE(...) upcxx_utils::shim_die(__FILENAME__, __LINE__, __VA_ARGS__)
