// libkfac_b200.so -- C-ABI entry points (include/kfac_b200.h) over the sm_100a kernels.
// Nothing in this file allocates device memory: scratch is caller-provided, except for kfac_factor_linear_noalloc,
// whose scratch is a fixed array compiled into the library.
#include "../../include/kfac_b200.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "gemm3x.cuh"
#include "gemm3x_launch.cuh"
#include "elementwise.cuh"
#include "eigh.cuh"

namespace kfac {

thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return 1;
}

StageTimers g_timers;
EighPaths g_eigh_paths;

int sm_count() {
  static int cached[KFAC_MAX_DEVICES] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= KFAC_MAX_DEVICES) return 148;
  if (cached[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    cached[dev] = v > KFAC_MAX_SMS ? KFAC_MAX_SMS : v;
  }
  return cached[dev];
}

int stage_mark(int idx, cudaStream_t stream) {
  if (!g_timers.on) return 0;
  int dev = 0;
  KFAC_CUDA(cudaGetDevice(&dev));
  if (g_timers.ev[0] == nullptr || g_timers.device != dev) {
    for (int i = 0; i < 7; ++i) {
      if (g_timers.ev[i] != nullptr) cudaEventDestroy(g_timers.ev[i]);
      KFAC_CUDA(cudaEventCreate(&g_timers.ev[i]));
    }
    g_timers.device = dev;
    g_timers.recorded = 0;
  }
  KFAC_CUDA(cudaEventRecord(g_timers.ev[idx], stream));
  if (idx == 0) g_timers.recorded = 0;
  g_timers.recorded |= 1 << idx;
  return 0;
}

}  // namespace kfac

using namespace kfac;

extern "C" {

const char* kfac_last_error(void) { return g_err; }
int kfac_abi_version(void) { return 1; }
unsigned long long kfac_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

void kfac_profile_enable(int on) { g_timers.on = on ? 1 : 0; }

int kfac_profile_last_ms(float* out4) {
  float* out3 = out4;
  if (out3 == nullptr) return fail("kfac_profile_last_ms: null output");
  if ((g_timers.recorded & 0xF) != 0xF) return fail("kfac_profile_last_ms: no profiled kfac_eigh_batched call has completed");
  KFAC_CUDA(cudaEventSynchronize(g_timers.ev[3]));
  for (int i = 0; i < 3; ++i) KFAC_CUDA(cudaEventElapsedTime(&out3[i], g_timers.ev[i], g_timers.ev[i + 1]));
  out4[3] = (float)g_timers.sytrd_launches;
  return 0;
}

int kfac_profile_detail_ms(float* out6) {
  if (out6 == nullptr) return fail("kfac_profile_detail_ms: null output");
  if ((g_timers.recorded & 0x7F) != 0x7F)
    return fail("kfac_profile_detail_ms: the last profiled kfac_eigh_batched call had no two-stage matrix");
  KFAC_CUDA(cudaEventSynchronize(g_timers.ev[3]));
  KFAC_CUDA(cudaEventElapsedTime(&out6[0], g_timers.ev[0], g_timers.ev[4]));   // band reduction
  KFAC_CUDA(cudaEventElapsedTime(&out6[1], g_timers.ev[4], g_timers.ev[5]));   // bulge chasing
  KFAC_CUDA(cudaEventElapsedTime(&out6[2], g_timers.ev[2], g_timers.ev[6]));   // Q2
  KFAC_CUDA(cudaEventElapsedTime(&out6[3], g_timers.ev[6], g_timers.ev[3]));   // Q1 (compact WY, all matrices)
  out6[4] = (float)g_timers.chase_launches;
  out6[5] = (float)g_timers.q2_launches;
  return 0;
}

int kfac_eigh_last_paths(int* out4) {
  if (out4 == nullptr) return fail("kfac_eigh_last_paths: null output");
  out4[0] = g_eigh_paths.full;
  out4[1] = g_eigh_paths.sym_ldg;
  out4[2] = g_eigh_paths.sym_bulk;
  out4[3] = g_eigh_paths.two_stage;
  return 0;
}

int kfac_eigh_set_mode(int mode) {
  if (mode < 0 || (mode > 2 && mode < 96))
    return fail("kfac_eigh_set_mode: mode must be 0 (by size), 1 (one-stage), 2 (two-stage) or a size threshold >= 96");
  g_eigh_mode = mode;
  return 0;
}

size_t kfac_factor_workspace_bytes(int n, long long K) {
  GemmParams p;
  init_params(p);
  p.sym = 1;
  p.M = p.N = n;
  p.K = (int)(K > 0x7fffffffLL ? 0x7fffffffLL : K);
  if (n <= 0 || K <= 0) return 0;
  plan_gemm(p, true);
  return gemm_ws_bytes(p);
}

static int check_dtype(const char* who, int dtype) {
  if (dtype < KFAC_DTYPE_F32 || dtype > KFAC_DTYPE_BF16) return fail("%s: unknown dtype %d", who, dtype);
  return 0;
}

int kfac_factor_linear(const float* x, long long M, int n_in, int has_bias, float in_scale, float alpha,
                       float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  return kfac_factor_linear_dt(x, KFAC_DTYPE_F32, M, n_in, has_bias, in_scale, alpha, beta, F, ws, ws_bytes, stream);
}

int kfac_factor_linear_dt(const void* x, int dtype, long long M, int n_in, int has_bias, float in_scale, float alpha,
                          float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  if (x == nullptr || F == nullptr) return fail("kfac_factor_linear: null pointer");
  if (check_dtype("kfac_factor_linear", dtype)) return 1;
  if (M <= 0 || n_in <= 0) return fail("kfac_factor_linear: empty input (M=%lld, n_in=%d)", M, n_in);
  Operand op = dense_operand(static_cast<const float*>(x), /*kcontig=*/false, n_in + (has_bias ? 1 : 0), n_in);
  op.elt = dtype;
  op.ones_row = has_bias ? n_in : -1;
  // the ones column is not scaled by in_scale in the reference only when in_scale == 1 (A factors); G factors
  // have no bias column, so a single scalar is exact for both.
  const float a = alpha * in_scale * in_scale / (float)M;
  return launch_factor(op, M, a, beta, F, ws, ws_bytes, (cudaStream_t)stream);
}

int kfac_factor_conv2d_a(const float* x, int N, int C, int H, int W, int kh, int kw, int sh, int sw, int ph,
                         int pw, int has_bias, float alpha, float beta, float* F, void* ws, size_t ws_bytes,
                         kfac_stream_t stream) {
  return kfac_factor_conv2d_a_dt(x, KFAC_DTYPE_F32, N, C, H, W, kh, kw, sh, sw, ph, pw, has_bias, alpha, beta, F, ws,
                                 ws_bytes, stream);
}

int kfac_factor_conv2d_a_dt(const void* x, int dtype, int N, int C, int H, int W, int kh, int kw, int sh, int sw,
                            int ph, int pw, int has_bias, float alpha, float beta, float* F, void* ws,
                            size_t ws_bytes, kfac_stream_t stream) {
  if (x == nullptr || F == nullptr) return fail("kfac_factor_conv2d_a: null pointer");
  if (check_dtype("kfac_factor_conv2d_a", dtype)) return 1;
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0 || kh <= 0 || kw <= 0 || sh <= 0 || sw <= 0 || ph < 0 || pw < 0)
    return fail("kfac_factor_conv2d_a: bad geometry");
  const int OH = (H + 2 * ph - kh) / sh + 1;
  const int OW = (W + 2 * pw - kw) / sw + 1;
  if (OH <= 0 || OW <= 0) return fail("kfac_factor_conv2d_a: kernel larger than padded input");
  Operand op;
  memset(&op, 0, sizeof(op));
  op.p = static_cast<const float*>(x);
  op.elt = dtype;
  op.mode = OP_IM2COL;
  op.R = C * kh * kw + (has_bias ? 1 : 0);
  op.ones_row = has_bias ? C * kh * kw : -1;
  op.ext = nullptr;
  op.ext_k = -1;
  op.C = C; op.H = H; op.W = W; op.KH = kh; op.KW = kw; op.SH = sh; op.SW = sw; op.PH = ph; op.PW = pw;
  op.OH = OH; op.OW = OW;
  const long long S = (long long)OH * OW;
  const long long K = (long long)N * S;
  const double a = (double)alpha / ((double)K * (double)S * (double)S);
  return launch_factor(op, K, (float)a, beta, F, ws, ws_bytes, (cudaStream_t)stream);
}

int kfac_factor_conv2d_g(const float* g, int N, int Cout, int OH, int OW, float in_scale, float alpha,
                         float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  return kfac_factor_conv2d_g_dt(g, KFAC_DTYPE_F32, N, Cout, OH, OW, in_scale, alpha, beta, F, ws, ws_bytes, stream);
}

int kfac_factor_conv2d_g_dt(const void* g, int dtype, int N, int Cout, int OH, int OW, float in_scale, float alpha,
                            float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  if (g == nullptr || F == nullptr) return fail("kfac_factor_conv2d_g: null pointer");
  if (check_dtype("kfac_factor_conv2d_g", dtype)) return 1;
  if (N <= 0 || Cout <= 0 || OH <= 0 || OW <= 0) return fail("kfac_factor_conv2d_g: bad geometry");
  Operand op;
  memset(&op, 0, sizeof(op));
  op.p = static_cast<const float*>(g);
  op.elt = dtype;
  op.mode = OP_NCS;
  op.R = Cout;
  op.ones_row = -1;
  op.ext = nullptr;
  op.ext_k = -1;
  op.C = Cout;
  op.S = OH * OW;
  op.KH = op.KW = 1;
  const long long S = (long long)OH * OW;
  const long long K = (long long)N * S;
  const double a = (double)alpha * (double)in_scale * (double)in_scale / ((double)K * (double)S * (double)S);
  return launch_factor(op, K, (float)a, beta, F, ws, ws_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ factor call that allocates nothing
// Scratch for kfac_factor_linear_noalloc lives in the library image itself: one array per GPU that loads the
// library, resident until the process exits.  The name below is only a host-side handle; kernels need the device
// address, which cudaGetSymbolAddress returns for the CURRENT device -- so it is looked up once per device.
#define KFAC_NOALLOC_SCRATCH_BYTES (16u << 20)
static __device__ __align__(256) unsigned char g_noalloc_scratch[KFAC_NOALLOC_SCRATCH_BYTES];
static void* g_noalloc_scratch_addr[64] = {nullptr};

static int noalloc_scratch_for_current_device(void** out) {
  int dev = 0;
  KFAC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail("device index %d out of range", dev);
  if (g_noalloc_scratch_addr[dev] == nullptr) {
    void* addr = nullptr;
    KFAC_CUDA(cudaGetSymbolAddress(&addr, g_noalloc_scratch));
    g_noalloc_scratch_addr[dev] = addr;
  }
  *out = g_noalloc_scratch_addr[dev];
  return 0;
}

size_t kfac_noalloc_scratch_bytes(void) { return KFAC_NOALLOC_SCRATCH_BYTES; }

int kfac_factor_linear_noalloc(const float* x, long long M, int n_in, int has_bias, float in_scale, float alpha,
                               float beta, float* F, kfac_stream_t stream) {
  if (x == nullptr || F == nullptr) return fail("kfac_factor_linear_noalloc: null pointer");
  if (M <= 0 || n_in <= 0) return fail("kfac_factor_linear_noalloc: empty input (M=%lld, n_in=%d)", M, n_in);
  const size_t need = kfac_factor_workspace_bytes(n_in + (has_bias ? 1 : 0), M);
  if (need > KFAC_NOALLOC_SCRATCH_BYTES)
    return fail("kfac_factor_linear_noalloc: a %d x %d factor from %lld samples needs %zu bytes of scratch, more "
                "than the %u bytes built into the library; use kfac_factor_linear with caller-provided scratch",
                n_in + (has_bias ? 1 : 0), n_in + (has_bias ? 1 : 0), M, need, KFAC_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (noalloc_scratch_for_current_device(&scratch)) return 1;
  // Nothing in the scratch has to start at zero: every split-K partial tile is written in full by the SYRK
  // kernel before the finalize kernel reads it, so the previous call's contents are simply overwritten.
  return kfac_factor_linear(x, M, n_in, has_bias, in_scale, alpha, beta, F, scratch, KFAC_NOALLOC_SCRATCH_BYTES,
                            stream);
}

// ------------------------------------------------------------------ elementwise
int kfac_axpby(float* y, const float* x, float a, float b, long long n, kfac_stream_t stream) {
  if (n <= 0) return 0;
  if (x == nullptr || y == nullptr) return fail("kfac_axpby: null pointer");
  return launch_axpby(y, x, a, b, n, (cudaStream_t)stream);
}
int kfac_set_identity(float* F, int n, kfac_stream_t stream) {
  if (F == nullptr || n <= 0) return fail("kfac_set_identity: bad arguments");
  return launch_set_identity(F, n, (cudaStream_t)stream);
}
int kfac_add_diag(float* F, int n, float damping, kfac_stream_t stream) {
  if (F == nullptr || n <= 0) return fail("kfac_add_diag: bad arguments");
  return launch_add_diag(F, n, damping, (cudaStream_t)stream);
}
int kfac_triu_pack(const float* F, int n, float* packed, kfac_stream_t stream) {
  if (F == nullptr || packed == nullptr || n <= 0) return fail("kfac_triu_pack: bad arguments");
  return launch_triu(const_cast<float*>(F), n, packed, /*unpack=*/false, (cudaStream_t)stream);
}
int kfac_triu_unpack(const float* packed, int n, float* F, kfac_stream_t stream) {
  if (F == nullptr || packed == nullptr || n <= 0) return fail("kfac_triu_unpack: bad arguments");
  return launch_triu(F, n, const_cast<float*>(packed), /*unpack=*/true, (cudaStream_t)stream);
}
int kfac_eigen_outer(const float* dg, const float* da, int g, int a, float damping, float* dgda,
                     kfac_stream_t stream) {
  if (dg == nullptr || da == nullptr || dgda == nullptr || g <= 0 || a <= 0)
    return fail("kfac_eigen_outer: bad arguments");
  return launch_eigen_outer(dg, da, g, a, damping, dgda, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ general GEMM
static void setup_gemm(GemmParams& p, int transa, int transb, int m, int n, int k, float alpha, const float* A,
                       long long lda, const float* B, long long ldb, float beta, float* C, long long ldc) {
  init_params(p);
  p.a = dense_operand(A, /*kcontig=*/!transa, m, lda);
  p.b = dense_operand(B, /*kcontig=*/transb != 0, n, ldb);
  p.M = m; p.N = n; p.K = k;
  p.alpha = alpha; p.beta = beta;
  p.out = C; p.ldo = ldc;
}

size_t kfac_gemm_workspace_bytes(int m, int n, int k) {
  if (m <= 0 || n <= 0 || k <= 0) return 0;
  GemmParams p;
  init_params(p);
  p.M = m; p.N = n; p.K = k;
  plan_gemm(p, true);
  return gemm_ws_bytes(p);
}

int kfac_gemm(int transa, int transb, int m, int n, int k, float alpha, const float* A, long long lda,
              const float* B, long long ldb, float beta, float* C, long long ldc, void* ws, size_t ws_bytes,
              kfac_stream_t stream) {
  if (A == nullptr || B == nullptr || C == nullptr) return fail("kfac_gemm: null pointer");
  GemmParams p;
  setup_gemm(p, transa, transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
  return launch_gemm(p, true, ws, ws_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ precondition
size_t kfac_precondition_workspace_bytes(int g, int a) {
  if (g <= 0 || a <= 0) return 0;
  return 2 * (((size_t)g * a * sizeof(float) + 255) / 256 * 256);
}

int kfac_precondition_eigen(const float* w_grad, const float* b_grad, int g, int a_w, const float* qa,
                            const float* qg, const float* da, const float* dg, const float* dgda,
                            float damping, float* out, double* vg_accum, void* ws, size_t ws_bytes,
                            kfac_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (w_grad == nullptr || qa == nullptr || qg == nullptr || out == nullptr)
    return fail("kfac_precondition_eigen: null pointer");
  if (dgda == nullptr && (da == nullptr || dg == nullptr))
    return fail("kfac_precondition_eigen: need dgda or both da and dg");
  if (g <= 0 || a_w <= 0) return fail("kfac_precondition_eigen: bad shape");
  const int a = a_w + (b_grad != nullptr ? 1 : 0);
  const size_t need = kfac_precondition_workspace_bytes(g, a);
  if (ws == nullptr || ws_bytes < need)
    return fail("kfac_precondition_eigen: workspace too small (%zu given, %zu needed)", ws_bytes, need);
  float* t1 = static_cast<float*>(ws);
  float* t2 = reinterpret_cast<float*>(static_cast<char*>(ws) + need / 2);
  GemmParams p;
  // t1 = [Wg | bg] Qa                                         (eigen.py:373, right factor)
  init_params(p);
  p.a = dense_operand(w_grad, true, g, a_w);
  p.a.ext = b_grad;
  p.a.ext_k = b_grad != nullptr ? a_w : -1;
  p.b = dense_operand(qa, false, a, a);
  p.M = g; p.N = a; p.K = a; p.out = t1; p.ldo = a;
  if (launch_gemm(p, false, nullptr, 0, stream)) return 1;
  // t2 = (Qg^T t1) (*) dgda   or   / (dg da^T + damping)       (eigen.py:373-383)
  init_params(p);
  p.a = dense_operand(qg, false, g, g);
  p.b = dense_operand(t1, false, a, a);
  p.M = g; p.N = a; p.K = g; p.out = t2; p.ldo = a;
  if (dgda != nullptr) {
    p.epi = EPI_MUL_MAT; p.e0 = dgda; p.lde = a;
  } else {
    p.epi = EPI_DIV_OUTER; p.e0 = dg; p.e1 = da; p.damping = damping;
  }
  if (launch_gemm(p, false, nullptr, 0, stream)) return 1;
  // t1 = t2 Qa^T                                               (eigen.py:384, right factor)
  init_params(p);
  p.a = dense_operand(t2, true, g, a);
  p.b = dense_operand(qa, true, a, a);
  p.M = g; p.N = a; p.K = a; p.out = t1; p.ldo = a;
  if (launch_gemm(p, false, nullptr, 0, stream)) return 1;
  // out = Qg t1, with the KL-clip dot product in the epilogue  (eigen.py:384; base_preconditioner.py:428-431)
  init_params(p);
  p.a = dense_operand(qg, true, g, g);
  p.b = dense_operand(t1, false, a, a);
  p.M = g; p.N = a; p.K = g; p.out = out; p.ldo = a;
  if (vg_accum != nullptr) {
    p.dotm = w_grad; p.ldd = a_w; p.dotb = b_grad; p.dot_nw = a_w; p.dot_out = vg_accum;
  }
  return launch_gemm(p, false, nullptr, 0, stream);
}

// All layers of one step in four grouped launches (one per GEMM stage) instead of four launches per layer.
size_t kfac_precondition_grouped_workspace_bytes(const int* g, const int* a, int count) {
  if (g == nullptr || a == nullptr || count <= 0) return 0;
  size_t total = 0;
  for (int i = 0; i < count; ++i) total += kfac_precondition_workspace_bytes(g[i], a[i]);
  total += (size_t)(count + 8) * (sizeof(GemmParams) + 8) + 4096;
  return total + 256;
}

int kfac_precondition_eigen_grouped(int count, const float* const* w_grad, const float* const* b_grad, const int* g,
                                    const int* a_w, const float* const* qa, const float* const* qg,
                                    const float* const* da, const float* const* dg, const float* const* dgda,
                                    float damping, float* const* out, double* vg_accum, void* ws, size_t ws_bytes,
                                    kfac_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (count <= 0) return 0;
  if (w_grad == nullptr || b_grad == nullptr || g == nullptr || a_w == nullptr || qa == nullptr || qg == nullptr ||
      out == nullptr)
    return fail("kfac_precondition_eigen_grouped: null pointer");
  std::vector<int> a(count);
  for (int i = 0; i < count; ++i) {
    if (w_grad[i] == nullptr || qa[i] == nullptr || qg[i] == nullptr || out[i] == nullptr || g[i] <= 0 || a_w[i] <= 0)
      return fail("kfac_precondition_eigen_grouped: bad entry %d", i);
    const bool has_dgda = dgda != nullptr && dgda[i] != nullptr;
    if (!has_dgda && (da == nullptr || dg == nullptr || da[i] == nullptr || dg[i] == nullptr))
      return fail("kfac_precondition_eigen_grouped: entry %d needs dgda or both da and dg", i);
    a[i] = a_w[i] + (b_grad[i] != nullptr ? 1 : 0);
  }
  const size_t need = kfac_precondition_grouped_workspace_bytes(g, a.data(), count);
  if (ws == nullptr || ws_bytes < need)
    return fail("kfac_precondition_eigen_grouped: workspace too small (%zu given, %zu needed)", ws_bytes, need);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  std::vector<float*> t1(count), t2(count);
  size_t off = 0;
  for (int i = 0; i < count; ++i) {
    const size_t each = kfac_precondition_workspace_bytes(g[i], a[i]);
    t1[i] = reinterpret_cast<float*>(base + off);
    t2[i] = reinterpret_cast<float*>(base + off + each / 2);
    off += each;
  }
  char* table = base + off;
  const size_t table_bytes = (size_t)(count + 8) * (sizeof(GemmParams) + 8) + 4096;
  const int target = std::max(1, 296 / count);
  for (int stage = 0; stage < 4; ++stage) {
    GemmGroup grp;
    for (int i = 0; i < count; ++i) {
      GemmParams p;
      init_params(p);
      const int gi = g[i], ai = a[i];
      if (stage == 0) {          // t1 = [Wg | bg] Qa
        p.a = dense_operand(w_grad[i], true, gi, a_w[i]);
        p.a.ext = b_grad[i];
        p.a.ext_k = b_grad[i] != nullptr ? a_w[i] : -1;
        p.b = dense_operand(qa[i], false, ai, ai);
        p.M = gi; p.N = ai; p.K = ai; p.out = t1[i]; p.ldo = ai;
      } else if (stage == 1) {   // t2 = (Qg^T t1) (*) dgda  |  / (dg da^T + damping)
        p.a = dense_operand(qg[i], false, gi, gi);
        p.b = dense_operand(t1[i], false, ai, ai);
        p.M = gi; p.N = ai; p.K = gi; p.out = t2[i]; p.ldo = ai;
        if (dgda != nullptr && dgda[i] != nullptr) {
          p.epi = EPI_MUL_MAT; p.e0 = dgda[i]; p.lde = ai;
        } else {
          p.epi = EPI_DIV_OUTER; p.e0 = dg[i]; p.e1 = da[i]; p.damping = damping;
        }
      } else if (stage == 2) {   // t1 = t2 Qa^T
        p.a = dense_operand(t2[i], true, gi, ai);
        p.b = dense_operand(qa[i], true, ai, ai);
        p.M = gi; p.N = ai; p.K = ai; p.out = t1[i]; p.ldo = ai;
      } else {                   // out = Qg t1  (+ KL-clip dot)
        p.a = dense_operand(qg[i], true, gi, gi);
        p.b = dense_operand(t1[i], false, ai, ai);
        p.M = gi; p.N = ai; p.K = gi; p.out = out[i]; p.ldo = ai;
        if (vg_accum != nullptr) {
          p.dotm = w_grad[i]; p.ldd = a_w[i]; p.dotb = b_grad[i]; p.dot_nw = a_w[i]; p.dot_out = vg_accum;
        }
      }
      if (group_add(grp, p, false, target, nullptr, 0)) return 1;
    }
    if (group_launch(grp, table, table_bytes, stream)) return 1;
  }
  return 0;
}

int kfac_precondition_inverse(const float* w_grad, const float* b_grad, int g, int a_w, const float* a_inv,
                              const float* g_inv, float* out, double* vg_accum, void* ws, size_t ws_bytes,
                              kfac_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (w_grad == nullptr || a_inv == nullptr || g_inv == nullptr || out == nullptr)
    return fail("kfac_precondition_inverse: null pointer");
  if (g <= 0 || a_w <= 0) return fail("kfac_precondition_inverse: bad shape");
  const int a = a_w + (b_grad != nullptr ? 1 : 0);
  const size_t need = kfac_precondition_workspace_bytes(g, a);
  if (ws == nullptr || ws_bytes < need)
    return fail("kfac_precondition_inverse: workspace too small (%zu given, %zu needed)", ws_bytes, need);
  float* t1 = static_cast<float*>(ws);
  GemmParams p;
  // t1 = [Wg | bg] Ainv ; out = Ginv t1                         (inverse.py:233)
  init_params(p);
  p.a = dense_operand(w_grad, true, g, a_w);
  p.a.ext = b_grad;
  p.a.ext_k = b_grad != nullptr ? a_w : -1;
  p.b = dense_operand(a_inv, false, a, a);
  p.M = g; p.N = a; p.K = a; p.out = t1; p.ldo = a;
  if (launch_gemm(p, false, nullptr, 0, stream)) return 1;
  init_params(p);
  p.a = dense_operand(g_inv, true, g, g);
  p.b = dense_operand(t1, false, a, a);
  p.M = g; p.N = a; p.K = g; p.out = out; p.ldo = a;
  if (vg_accum != nullptr) {
    p.dotm = w_grad; p.ldd = a_w; p.dotb = b_grad; p.dot_nw = a_w; p.dot_out = vg_accum;
  }
  return launch_gemm(p, false, nullptr, 0, stream);
}

// The inverse-method loop over layers (kfac/base_preconditioner.py:363-365 -> inverse.py:214-233) in two grouped
// launches: t1 = [Wg | bg] Ainv for all layers, then out = Ginv t1 (+ KL-clip dot).
int kfac_precondition_inverse_grouped(int count, const float* const* w_grad, const float* const* b_grad, const int* g,
                                      const int* a_w, const float* const* a_inv, const float* const* g_inv,
                                      float* const* out, double* vg_accum, void* ws, size_t ws_bytes,
                                      kfac_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (count <= 0) return 0;
  if (w_grad == nullptr || b_grad == nullptr || g == nullptr || a_w == nullptr || a_inv == nullptr || g_inv == nullptr ||
      out == nullptr)
    return fail("kfac_precondition_inverse_grouped: null pointer");
  std::vector<int> a(count);
  for (int i = 0; i < count; ++i) {
    if (w_grad[i] == nullptr || a_inv[i] == nullptr || g_inv[i] == nullptr || out[i] == nullptr || g[i] <= 0 || a_w[i] <= 0)
      return fail("kfac_precondition_inverse_grouped: bad entry %d", i);
    a[i] = a_w[i] + (b_grad[i] != nullptr ? 1 : 0);
  }
  const size_t need = kfac_precondition_grouped_workspace_bytes(g, a.data(), count);
  if (ws == nullptr || ws_bytes < need)
    return fail("kfac_precondition_inverse_grouped: workspace too small (%zu given, %zu needed)", ws_bytes, need);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  std::vector<float*> t1(count);
  size_t off = 0;
  for (int i = 0; i < count; ++i) {
    t1[i] = reinterpret_cast<float*>(base + off);
    off += kfac_precondition_workspace_bytes(g[i], a[i]);
  }
  char* table = base + off;
  const size_t table_bytes = (size_t)(count + 8) * (sizeof(GemmParams) + 8) + 4096;
  const int target = std::max(1, 296 / count);
  for (int stage = 0; stage < 2; ++stage) {
    GemmGroup grp;
    for (int i = 0; i < count; ++i) {
      GemmParams p;
      init_params(p);
      const int gi = g[i], ai = a[i];
      if (stage == 0) {   // t1 = [Wg | bg] Ainv
        p.a = dense_operand(w_grad[i], true, gi, a_w[i]);
        p.a.ext = b_grad[i];
        p.a.ext_k = b_grad[i] != nullptr ? a_w[i] : -1;
        p.b = dense_operand(a_inv[i], false, ai, ai);
        p.M = gi; p.N = ai; p.K = ai; p.out = t1[i]; p.ldo = ai;
      } else {            // out = Ginv t1 (+ KL-clip dot)
        p.a = dense_operand(g_inv[i], true, gi, gi);
        p.b = dense_operand(t1[i], false, ai, ai);
        p.M = gi; p.N = ai; p.K = gi; p.out = out[i]; p.ldo = ai;
        if (vg_accum != nullptr) {
          p.dotm = w_grad[i]; p.ldd = a_w[i]; p.dotb = b_grad[i]; p.dot_nw = a_w[i]; p.dot_out = vg_accum;
        }
      }
      if (group_add(grp, p, false, target, nullptr, 0)) return 1;
    }
    if (group_launch(grp, table, table_bytes, stream)) return 1;
  }
  return 0;
}

int kfac_grad_dot(const float* v, const float* w_grad, const float* b_grad, int g, int a_w, double* vg_accum,
                  kfac_stream_t stream) {
  if (v == nullptr || w_grad == nullptr || vg_accum == nullptr || g <= 0 || a_w <= 0)
    return fail("kfac_grad_dot: bad arguments");
  return launch_grad_dot(v, w_grad, b_grad, g, a_w, vg_accum, (cudaStream_t)stream);
}

int kfac_klclip_scale(const double* vg_accum, float lr, float kl_clip, float* scale_out, kfac_stream_t stream) {
  if (vg_accum == nullptr || scale_out == nullptr) return fail("kfac_klclip_scale: null pointer");
  return launch_klclip_scale(vg_accum, lr, kl_clip, scale_out, (cudaStream_t)stream);
}

int kfac_scale_write_grads(const float* pg, int g, int a_w, const float* scale_dev, float* w_grad,
                           float* b_grad, kfac_stream_t stream) {
  if (pg == nullptr || w_grad == nullptr || g <= 0 || a_w <= 0) return fail("kfac_scale_write_grads: bad arguments");
  return launch_scale_write(pg, g, a_w, scale_dev, w_grad, b_grad, (cudaStream_t)stream);
}

size_t kfac_scale_write_grouped_workspace_bytes(int count) {
  return count <= 0 ? 0 : (size_t)count * sizeof(ScaleWriteItem) + 256;
}

int kfac_scale_write_grads_grouped(int count, const float* const* pg, const int* g, const int* a_w,
                                   const float* scale_dev, float* const* w_grad, float* const* b_grad, void* ws,
                                   size_t ws_bytes, kfac_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (count <= 0) return 0;
  if (pg == nullptr || g == nullptr || a_w == nullptr || w_grad == nullptr || b_grad == nullptr)
    return fail("kfac_scale_write_grads_grouped: null pointer");
  if (ws == nullptr || ws_bytes < kfac_scale_write_grouped_workspace_bytes(count))
    return fail("kfac_scale_write_grads_grouped: workspace too small");
  std::vector<ScaleWriteItem> items(count);
  long long max_total = 0;
  for (int i = 0; i < count; ++i) {
    if (pg[i] == nullptr || w_grad[i] == nullptr || g[i] <= 0 || a_w[i] <= 0)
      return fail("kfac_scale_write_grads_grouped: bad entry %d", i);
    items[i] = ScaleWriteItem{pg[i], w_grad[i], b_grad[i], g[i], a_w[i]};
    max_total = std::max(max_total, (long long)g[i] * (a_w[i] + 1));
  }
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  KFAC_CUDA(cudaMemcpyAsync(base, items.data(), (size_t)count * sizeof(ScaleWriteItem), cudaMemcpyHostToDevice, stream));
  const int bx = (int)std::max<long long>(1, std::min<long long>((max_total / 4 + EW_THREADS - 1) / EW_THREADS,
                                                               std::max(1, 8 * sm_count() / count)));
  scale_write_group_kernel<<<dim3(bx, count), EW_THREADS, 0, stream>>>(reinterpret_cast<const ScaleWriteItem*>(base),
                                                                      scale_dev);
  KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------ second-order data
size_t kfac_eigh_workspace_bytes(const int* ns, int count) { return eigh_workspace_bytes(ns, count); }

int kfac_eigh_batched(const float* const* F, const int* ns, int count, float* const* Q, float* const* d,
                      int clamp_nonneg, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  return eigh_batched(F, ns, count, Q, d, clamp_nonneg, ws, ws_bytes, (cudaStream_t)stream);
}

size_t kfac_debug_two_stage_workspace_bytes(int n) { return debug_two_stage_workspace_bytes(n); }
int kfac_debug_two_stage_tiles(int n) { return n < 3 ? 0 : ts_total_tiles(n); }
int kfac_debug_two_stage(const float* F, int n, float* W_out, float* tau1_out, float* band_out, float* d_out,
                         float* e_out, float* V2_out, float* tau2_out, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  return debug_two_stage(F, n, W_out, tau1_out, band_out, d_out, e_out, V2_out, tau2_out, ws, ws_bytes,
                         (cudaStream_t)stream);
}

size_t kfac_inverse_workspace_bytes(int n) { return inverse_workspace_bytes(n); }

size_t kfac_inverse_batched_workspace_bytes(const int* ns, int count) { return inverse_batched_workspace_bytes(ns, count); }

int kfac_inverse_newton_schulz_batched(const float* const* F, const int* ns, int count, float damping,
                                       float* const* out, int max_iters, float tol, int* info, void* ws,
                                       size_t ws_bytes, kfac_stream_t stream) {
  return inverse_newton_schulz_batched(F, ns, count, damping, out, max_iters, tol, info, ws, ws_bytes,
                                       (cudaStream_t)stream);
}

int kfac_inverse_newton_schulz(const float* F, int n, float damping, float* out, int max_iters, float tol,
                               int* iters_out, void* ws, size_t ws_bytes, kfac_stream_t stream) {
  return inverse_newton_schulz(F, n, damping, out, max_iters, tol, iters_out, ws, ws_bytes, (cudaStream_t)stream);
}

}  // extern "C"
