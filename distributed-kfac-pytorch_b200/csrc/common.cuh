// Shared host-side helpers: error reporting for the C ABI.
#pragma once
#include <cuda_runtime.h>

namespace kfac {

// Records a printf-style message for kfac_last_error() and returns 1.
int fail(const char* fmt, ...);

// Kernel launches issued by this library in this process (diagnostics for bench.py's gpu_launches).
extern unsigned long long g_launches;
#define KFAC_COUNT_LAUNCH(n) (::kfac::g_launches += (n))

// Optional stage timers of the batched eigensolver (kfac_profile_enable): CUDA events recorded on the launching
// stream around the tridiagonalisation rounds, the divide & conquer and the back-transformation.
struct StageTimers {
  int on = 0;
  int device = -1;
  int recorded = 0;
  int sytrd_launches = 0;   // cooperative launches of the last profiled tridiagonalisation stage
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};
extern StageTimers g_timers;
int stage_mark(int idx, cudaStream_t stream);   // no-op unless enabled; returns 0 on success

#define KFAC_CUDA(expr)                                                                          \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) return ::kfac::fail("%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

}  // namespace kfac
