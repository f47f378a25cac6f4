// Shared host-side helpers: error reporting for the C ABI.
#pragma once
#include <cuda_runtime.h>

namespace kfac {

// Records a printf-style message for kfac_last_error() and returns 1.
int fail(const char* fmt, ...);

// Kernel launches issued by this library in this process (diagnostics for bench.py's gpu_launches).
extern unsigned long long g_launches;
#define KFAC_COUNT_LAUNCH(n) (::kfac::g_launches += (n))

#define KFAC_CUDA(expr)                                                                          \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) return ::kfac::fail("%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

}  // namespace kfac
