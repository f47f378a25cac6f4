// Shared host-side helpers: error reporting for the C ABI.
#pragma once
#include <cuda_runtime.h>

#include <atomic>

namespace kfac {

// Records a printf-style message for kfac_last_error() and returns 1.
int fail(const char* fmt, ...);

// Kernel launches issued by this library in this process (diagnostics for bench.py's gpu_launches).
extern std::atomic<unsigned long long> g_launches;   // hooks run on the autograd thread, step() on the caller's
#define KFAC_COUNT_LAUNCH(n) (::kfac::g_launches.fetch_add((n), std::memory_order_relaxed))

// Optional stage timers of the batched eigensolver (kfac_profile_enable): CUDA events recorded on the launching
// stream around the tridiagonalisation rounds, the divide & conquer and the back-transformation.
struct StageTimers {
  int on = 0;
  int device = -1;
  int recorded = 0;
  int sytrd_launches = 0;   // cooperative launches of the last profiled tridiagonalisation stage
  int chase_launches = 0;   // ts_chase_kernel launches of that call
  int q2_launches = 0;      // ts_q2_kernel launches of that call
  // 0..3: start / tridiagonalised / divide & conquer done / end; 4..6 (two-stage matrices only): band reduction done,
  // bulge chasing done, Q2 applied
  cudaEvent_t ev[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};
extern StageTimers g_timers;
int stage_mark(int idx, cudaStream_t stream);   // no-op unless enabled; returns 0 on success

// SMs of the current device (cached per device; 148 on a full B200, fewer under MIG / an MPS SM limit).  Cooperative
// launches size their grids from this, never from a constant.
int sm_count();
constexpr int KFAC_MAX_SMS = 192;   // sizing bound for per-SM tables
constexpr int KFAC_MAX_DEVICES = 64;

// Which tridiagonalisation path the matrices of the last kfac_eigh_batched call took (kfac_eigh_last_paths).
struct EighPaths {
  int full = 0;       // one-stage, full-matrix pass (n < 1024)
  int sym_ldg = 0;    // one-stage, lower-triangle pass, two-CTA LDG kernel
  int sym_bulk = 0;   // one-stage, lower-triangle pass, three-CTA bulk-copy kernel
  int two_stage = 0;  // band reduction on the tensor cores + bulge chasing
};
extern EighPaths g_eigh_paths;

#define KFAC_CUDA(expr)                                                                          \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) return ::kfac::fail("%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

}  // namespace kfac
