// TEST INFRASTRUCTURE (oracle/) -- inert memory log.
#pragma once
#include "log.hpp"
#define LOG_MEM(...) \
  do {               \
  } while (0)
