// TEST INFRASTRUCTURE (oracle/) -- memory probes used by the reference's table sizing.
#pragma once
#include <unistd.h>
#include <cstdlib>
#include <fstream>
#include <string>
#include "log.hpp"

#define LOG_MEM(...) \
  do {               \
  } while (0)

namespace upcxx_utils {
// The reference sizes its table as min(3, 0.8*free/need) x the 0.66-load estimate
// (kcount_cpu.cpp:455-465); ORACLE_REF_FREE_MEM lets the harness pin that input.
inline double get_free_mem(bool = false) {
  if (const char *e = getenv("ORACLE_REF_FREE_MEM")) return atof(e);
  std::ifstream f("/proc/meminfo");
  std::string key;
  double kb;
  std::string unit;
  while (f >> key >> kb >> unit)
    if (key == "MemAvailable:") return kb * 1024.0;
  return 8e9;
}
}  // namespace upcxx_utils
