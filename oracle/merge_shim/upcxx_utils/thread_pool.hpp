// TEST INFRASTRUCTURE (oracle/) -- only the sleep helper of the thread pool is touched (fastq.cpp's open retry loop).
#pragma once
#include <chrono>
#include <cstdint>
#include <thread>
namespace upcxx_utils {
struct ThreadPool {
  static void sleep_ns(uint64_t ns) { std::this_thread::sleep_for(std::chrono::nanoseconds(ns)); }
};
}  // namespace upcxx_utils
