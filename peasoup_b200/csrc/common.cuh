// common.cuh -- error plumbing shared by the translation units of libpeasoup_b200.so
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <string>
#include "../../include/peasoup_b200.h"

namespace ps {

void set_error(const std::string& msg);
extern std::atomic<unsigned long long> g_launches;

// Mirrors ErrorChecker::check_cuda_error's message format (utils/exceptions.hpp:64-74)
// without its device-wide synchronisation.
inline int cuda_fail(cudaError_t e, const char* where) {
  set_error(std::string("CUDA failed with error: ") + cudaGetErrorString(e) + "\nAdditional: " + where);
  return PS_ECUDA;
}

#define PS_CUDA(call, where)                          \
  do {                                                \
    cudaError_t e__ = (call);                         \
    if (e__ != cudaSuccess) return ps::cuda_fail(e__, where); \
  } while (0)

// after a kernel launch: count it and pick up launch-configuration errors
#define PS_LAUNCHED(where)                            \
  do {                                                \
    ps::g_launches.fetch_add(1, std::memory_order_relaxed); \
    cudaError_t e__ = cudaGetLastError();             \
    if (e__ != cudaSuccess) return ps::cuda_fail(e__, where); \
  } while (0)

inline int invalid(const std::string& msg) {
  set_error(msg);
  return PS_EINVAL;
}

inline cudaStream_t S(void* s) { return (cudaStream_t)s; }

// grid size for grid-stride element-wise kernels: a whole number of waves on 148 SMs
inline unsigned ew_grid(size_t n, unsigned threads, unsigned per_thread = 1) {
  size_t need = (n + (size_t)threads * per_thread - 1) / ((size_t)threads * per_thread);
  size_t cap = 148u * 8u;
  if (need < 1) need = 1;
  return (unsigned)(need < cap ? need : cap);
}

}  // namespace ps
