/* kfac_b200.h -- C ABI of the B200-native K-FAC / KAISA preconditioning kernels (libkfac_b200.so).
 *
 * The reference (kfac_pytorch 0.4.1, /root/reference) has no FFI: every entry point below replaces a
 * torch-op call site of the reference's hot path, cited as file:line.  All pointers are DEVICE pointers
 * unless stated, matrices are row-major fp32, every call is asynchronous on `stream` (a cudaStream_t
 * passed as void*), nothing here allocates: scratch comes from the caller (`ws`, sized by the matching
 * *_workspace_bytes query); the one exception is kfac_factor_linear_noalloc, whose scratch is built into the
 * library.  Return value: 0 on success, non-zero on error with a message available from
 * kfac_last_error() (thread-local).  The Python host layer raises RuntimeError from it, matching the
 * reference's exception style (kfac/layers/base.py:239-244).
 */
#ifndef KFAC_B200_H_
#define KFAC_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* kfac_stream_t;

const char* kfac_last_error(void);
int kfac_abi_version(void);
/* Diagnostics: kernel launches issued by this library in this process so far (bench.py's gpu_launches).
 * No counterpart in the reference. */
unsigned long long kfac_launch_count(void);
/* Diagnostics: when enabled, kfac_eigh_batched records CUDA events on its launching stream around its three
 * stages; kfac_profile_last_ms waits for the last profiled call and writes {tridiagonalisation, divide & conquer,
 * back-transformation} milliseconds to out4[0..2] and the number of tridiagonalisation launches to out4[3]
 * (bench.py's roofline.achieved uses the first and the last).  The reference's
 * counterpart is the wall-clock decorator kfac/tracing.py:73-107 (@trace); no data-path effect. */
void kfac_profile_enable(int on);
int kfac_profile_last_ms(float* out4);
/* Diagnostics: finer split of the same profiled call when it had two-stage matrices (non-zero return otherwise):
 * out6 = {band reduction ms, bulge chasing ms, Q2 back-transformation ms, Q1 (compact-WY) back-transformation ms,
 * ts_chase_kernel launches, ts_q2_kernel launches}.  bench.py's roofline uses the chasing and Q2 figures. */
int kfac_profile_detail_ms(float* out6);
/* Diagnostics: how many matrices of the last kfac_eigh_batched call (this process) went through each
 * tridiagonalisation path: out4 = {one-stage full-matrix pass, one-stage lower-triangle pass (LDG kernel),
 * one-stage lower-triangle pass (bulk-copy kernel), two-stage band reduction + bulge chasing}.  Lets the tests pin
 * which kernel a given size exercises, so a dispatch change cannot silently drop coverage.  No counterpart in the
 * reference (torch.linalg.eigh hides its LAPACK driver, kfac/layers/eigen.py:309). */
int kfac_eigh_last_paths(int* out4);

/* ---------------------------------------------------------------- factor stage (hooks) ------------- */

/* Scratch needed by one factor call producing an n x n factor from K samples (upper bound). */
size_t kfac_factor_workspace_bytes(int n, long long K);

/* LinearModuleHelper.get_a_factor / get_g_factor  (kfac/layers/modules.py:123-141) fused with
 * append_bias_ones + get_cov (kfac/layers/utils.py:7-14,50-56) and the running average
 * (kfac/layers/base.py:374-404):
 *     F <- beta * F + alpha * in_scale^2 * [x, 1]^T [x, 1] / M        (ones column iff has_bias)
 * x: (M, n_in) row-major.  F: (n, n), n = n_in + has_bias.  in_scale = 1/grad_scaler() for G factors
 * under AMP (kfac/layers/base.py:364-365), else 1. */
int kfac_factor_linear(const float* x, long long M, int n_in, int has_bias, float in_scale, float alpha,
                       float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream);

/* Conv2dModuleHelper.get_a_factor (kfac/layers/modules.py:170-178) with the im2col of
 * _extract_patches (:210-237) fused into the operand loader (never materialised):
 *     F <- beta * F + alpha * P~^T P~ / (M * S^2),  P~ = [patches, 1],  S = OH*OW,  M = N*S
 * x: (N, C, H, W).  F: (n, n), n = C*kh*kw + has_bias, column order (c, kh, kw). */
int kfac_factor_conv2d_a(const float* x, int N, int C, int H, int W, int kh, int kw, int sh, int sw, int ph,
                         int pw, int has_bias, float alpha, float beta, float* F, void* ws, size_t ws_bytes,
                         kfac_stream_t stream);

/* Conv2dModuleHelper.get_g_factor (kfac/layers/modules.py:180-192):
 *     F <- beta * F + alpha * in_scale^2 * g^T g / (M * S^2),  g viewed as (M = N*S, Cout)
 * g: (N, Cout, OH, OW). */
int kfac_factor_conv2d_g(const float* g, int N, int Cout, int OH, int OW, float in_scale, float alpha,
                         float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream);

/* The same three factor kernels for activations / grad-outputs that arrive in half precision (torch.autocast: the
 * hooks of kfac/base_preconditioner.py:435-477 see fp16 / bf16 tensors; the reference then computes the covariance
 * in that dtype unless factor_dtype is given, kfac/layers/base.py:349,382-383).  Here the 16-bit values are read in
 * place (no up-cast copy), are exact in tf32, and the factor is their exact fp32 covariance.  dtype: KFAC_DTYPE_*. */
#define KFAC_DTYPE_F32 0
#define KFAC_DTYPE_F16 1
#define KFAC_DTYPE_BF16 2
int kfac_factor_linear_dt(const void* x, int dtype, long long M, int n_in, int has_bias, float in_scale, float alpha,
                          float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream);
int kfac_factor_conv2d_a_dt(const void* x, int dtype, int N, int C, int H, int W, int kh, int kw, int sh, int sw,
                            int ph, int pw, int has_bias, float alpha, float beta, float* F, void* ws,
                            size_t ws_bytes, kfac_stream_t stream);
int kfac_factor_conv2d_g_dt(const void* g, int dtype, int N, int Cout, int OH, int OW, float in_scale, float alpha,
                            float beta, float* F, void* ws, size_t ws_bytes, kfac_stream_t stream);

/* Same result as kfac_factor_linear, for callers that cannot hand over scratch memory.
 * No counterpart in the reference (which allocates freely through torch); it exists for hosts that must not
 * allocate while a step is running: a training loop whose framework reserved the whole device up front, so a late
 * cudaMalloc would fail; a latency-critical loop that cannot afford the device-wide synchronisation hidden in
 * cudaMalloc/cudaFree; a hook running where the torch allocator is out of reach.  The split-K scratch is a
 * 16 MiB array compiled into the library (kfac_noalloc_scratch_bytes()); every GPU that loads the library holds
 * that much device memory until the process exits.  Problems whose scratch need
 * (kfac_factor_workspace_bytes(n, M)) exceeds it are rejected with an error -- nothing is allocated instead.
 *   - The first call on each device looks up the array's device address (a runtime query).  The CUDA runtime loads
 *     a library's device code lazily, so this first use is also when the array itself comes into existence on
 *     that device: make the first call early -- with the device current, before the host reserves the rest of the
 *     device, and BEFORE starting any stream capture.  Later calls touch no allocator at all.
 *   - The address is cached per device for the life of the process; after cudaDeviceReset it is stale and the
 *     entry point must not be used again in that process.
 *   - All calls on a device share the one array.  Two calls that may run at the same time on the GPU (different
 *     streams, or different host threads on different streams) race on it and corrupt each other's partial sums;
 *     issue them on one stream or order them with events.  The array keeps the last call's partials; that is
 *     harmless, every call overwrites what it later reads.
 * Calls no cudaMalloc / cudaFree / allocating routine. */
size_t kfac_noalloc_scratch_bytes(void);
int kfac_factor_linear_noalloc(const float* x, long long M, int n_in, int has_bias, float in_scale, float alpha,
                               float beta, float* F, kfac_stream_t stream);

/* ---------------------------------------------------------------- memory-bound passes -------------- */

/* y <- a * x + b * y  over n floats (running average when the factor was accumulated over micro-batches,
 * kfac/layers/base.py:382-388; gradient scaling, :419-420). x may equal y. */
int kfac_axpby(float* y, const float* x, float a, float b, long long n, kfac_stream_t stream);
/* F <- I (first update_*_factor call, kfac/layers/base.py:386-387). */
int kfac_set_identity(float* F, int n, kfac_stream_t stream);
/* F <- F + damping * I (kfac/layers/inverse.py:197-200). */
int kfac_add_diag(float* F, int n, float damping, kfac_stream_t stream);
/* get_triu / fill_triu (kfac/distributed.py:416-459): row-major upper triangle, n(n+1)/2 floats. */
int kfac_triu_pack(const float* F, int n, float* packed, kfac_stream_t stream);
int kfac_triu_unpack(const float* packed, int n, float* F, kfac_stream_t stream);
/* dgda <- 1 / (outer(dg, da) + damping)  (kfac/layers/eigen.py:344-345), also clamps nothing: dg/da are
 * expected already clamped at 0 by kfac_eigh_batched. */
int kfac_eigen_outer(const float* dg, const float* da, int g, int a, float damping, float* dgda,
                     kfac_stream_t stream);

/* ---------------------------------------------------------------- general fp32-grade GEMM ---------- */

/* C <- alpha * op(A) op(B) + beta * C on the tensor cores (3xTF32 split, fp32 accumulate).
 * op(A) is m x k, op(B) is k x n; transa/transb follow BLAS ('N' = 0, 'T' = 1) for ROW-major storage.
 * Used by the inverse-method precondition (kfac/layers/inverse.py:233) and by the eigensolver. */
size_t kfac_gemm_workspace_bytes(int m, int n, int k);
int kfac_gemm(int transa, int transb, int m, int n, int k, float alpha, const float* A, long long lda,
              const float* B, long long ldb, float beta, float* C, long long ldc, void* ws, size_t ws_bytes,
              kfac_stream_t stream);

/* ---------------------------------------------------------------- precondition + KL-clip ----------- */

/* Scratch for one precondition call on a (g x a) gradient. */
size_t kfac_precondition_workspace_bytes(int g, int a);

/* KFACEigenLayer.preconditioned_grad (kfac/layers/eigen.py:349-384), reading weight.grad / bias.grad in
 * place instead of ModuleHelper.get_grad's concatenation (kfac/layers/modules.py:56-69):
 *     v1 = Qg^T [Wg | bg] Qa ;  v2 = v1 * dgda   (dgda != NULL)   or   v1 / (outer(dg, da) + damping)
 *     out = Qg v2 Qa^T                                            (g x a, a = a_w + (bg != NULL))
 * If vg_accum != NULL, sum(out * [Wg | bg]) is added to *vg_accum (the KL-clip dot product,
 * kfac/base_preconditioner.py:415-431, without its per-layer .item() syncs). */
int kfac_precondition_eigen(const float* w_grad, const float* b_grad, int g, int a_w, const float* qa,
                            const float* qg, const float* da, const float* dg, const float* dgda,
                            float damping, float* out, double* vg_accum, void* ws, size_t ws_bytes,
                            kfac_stream_t stream);

/* The reference's loop `for layer in reversed(layers): layer.preconditioned_grad()`
 * (kfac/base_preconditioner.py:363-365) as four grouped launches over all layers of this rank: host arrays of
 * per-layer device pointers / shapes, entries as for kfac_precondition_eigen (b_grad[i], dgda[i], da[i], dg[i]
 * may be NULL per layer; the arrays da / dg / dgda themselves may be NULL). */
size_t kfac_precondition_grouped_workspace_bytes(const int* g, const int* a, int count);
int kfac_precondition_eigen_grouped(int count, const float* const* w_grad, const float* const* b_grad, const int* g,
                                    const int* a_w, const float* const* qa, const float* const* qg,
                                    const float* const* da, const float* const* dg, const float* const* dgda,
                                    float damping, float* const* out, double* vg_accum, void* ws, size_t ws_bytes,
                                    kfac_stream_t stream);

/* KFACInverseLayer.preconditioned_grad (kfac/layers/inverse.py:214-233): out = Ginv [Wg | bg] Ainv. */
int kfac_precondition_inverse(const float* w_grad, const float* b_grad, int g, int a_w, const float* a_inv,
                              const float* g_inv, float* out, double* vg_accum, void* ws, size_t ws_bytes,
                              kfac_stream_t stream);

/* The same for all layers of this rank in two grouped launches (entries as for kfac_precondition_inverse; scratch
 * sized by kfac_precondition_grouped_workspace_bytes). */
int kfac_precondition_inverse_grouped(int count, const float* const* w_grad, const float* const* b_grad, const int* g,
                                      const int* a_w, const float* const* a_inv, const float* const* g_inv,
                                      float* const* out, double* vg_accum, void* ws, size_t ws_bytes,
                                      kfac_stream_t stream);

/* sum(v * [Wg | bg]) added to *vg_accum, for gradients received by broadcast (no precondition call ran on
 * this rank; kfac/base_preconditioner.py:415-431). */
int kfac_grad_dot(const float* v, const float* w_grad, const float* b_grad, int g, int a_w, double* vg_accum,
                  kfac_stream_t stream);

/* scale_out[0] <- vg == 0 ? 1 : min(1, sqrt(kl_clip / |vg * lr^2|))   (kfac/base_preconditioner.py:432-433).
 * Stays on the device: no host synchronisation. */
int kfac_klclip_scale(const double* vg_accum, float lr, float kl_clip, float* scale_out, kfac_stream_t stream);

/* KFACBaseLayer.update_grad + ModuleHelper.set_grad (kfac/layers/base.py:406-422, modules.py:87-97):
 * w_grad <- scale * pg[:, :a_w], b_grad <- scale * pg[:, a_w]; scale read from device memory (NULL = 1). */
int kfac_scale_write_grads(const float* pg, int g, int a_w, const float* scale_dev, float* w_grad,
                           float* b_grad, kfac_stream_t stream);

/* The loop `for layer in reversed(layers): layer.update_grad(scale)` (kfac/base_preconditioner.py:376-377) in ONE
 * launch: host arrays of per-layer device pointers / shapes (b_grad[i] may be NULL), one device scalar. */
size_t kfac_scale_write_grouped_workspace_bytes(int count);
int kfac_scale_write_grads_grouped(int count, const float* const* pg, const int* g, const int* a_w,
                                   const float* scale_dev, float* const* w_grad, float* const* b_grad, void* ws,
                                   size_t ws_bytes, kfac_stream_t stream);

/* ---------------------------------------------------------------- second-order data ---------------- */

/* torch.linalg.eigh(F.float()) + clamp(d, min=0) for a batch of symmetric factors
 * (kfac/layers/eigen.py:294-343).  Host arrays: F[i], Q[i], d[i] are device pointers to (ns[i] x ns[i]),
 * (ns[i] x ns[i]) and (ns[i]) buffers; eigenvalues ascending, eigenvectors in the COLUMNS of Q (row-major),
 * i.e. F = Q diag(d) Q^T.  F is not modified. */
size_t kfac_eigh_workspace_bytes(const int* ns, int count);
int kfac_eigh_batched(const float* const* F, const int* ns, int count, float* const* Q, float* const* d,
                      int clamp_nonneg, void* ws, size_t ws_bytes, kfac_stream_t stream);

/* Tridiagonalisation path of kfac_eigh_batched, process-wide: 0 = chosen per call by size and batch (default: the
 * two-stage reduction -- band reduction on the tensor cores, bulge chasing, two back-transformations -- for
 * matrices of n >= 1000 when the call holds at least four of them, for n >= 4096 otherwise; the one-stage kernels
 * below that), 1 = one-stage only, 2 = two-stage for every n >= 96, any m >= 96 = two-stage for n >= m.  For tests
 * and A/B timing; the reference has no counterpart (LAPACK's driver choice is hidden behind torch.linalg.eigh,
 * kfac/layers/eigen.py:309). */
int kfac_eigh_set_mode(int mode);

/* Diagnostics for the tests: the two-stage tridiagonalisation of ONE n x n matrix (n >= 34) with its intermediate
 * results copied out, so that each stage can be compared with the numpy model oracle/twostage_model.py:
 *   W_out (n x n)   rows k < n-33 hold the stage-1 Householder vector v_k (1 at column k+32, zeros before)
 *   tau1_out (n)    stage-1 tau
 *   band_out        n x 64 floats, band_out[j*64 + d] = band(j+d, j) after stage 1 (d <= 32)
 *   d_out, e_out    the tridiagonal matrix (n, n)
 *   V2_out, tau2_out  stage-2 reflectors, kfac_debug_two_stage_tiles(n) x 32 x 32 floats and x 32 floats, laid out
 *                   [tile offset of sweep group + step][sweep within group][32]
 * No counterpart in the reference. */
size_t kfac_debug_two_stage_workspace_bytes(int n);
int kfac_debug_two_stage_tiles(int n);
int kfac_debug_two_stage(const float* F, int n, float* W_out, float* tau1_out, float* band_out, float* d_out,
                         float* e_out, float* V2_out, float* tau2_out, void* ws, size_t ws_bytes, kfac_stream_t stream);

/* (F[i] + damping I)^-1 for a batch of symmetric positive semi-definite factors by Newton-Schulz iteration on the
 * tensor cores, replacing the loop of torch.linalg.inv calls (kfac/layers/inverse.py:185-212 called per layer from
 * kfac/base_preconditioner.py:342-359).  All matrices advance together, two grouped GEMM launches per iteration; the
 * residual max |I - M X| and the decision that a matrix is finished stay on the device (finished matrices turn the
 * remaining iterations into no-ops), so the call is asynchronous like every other entry point: NO host
 * synchronisation.  max_iters <= 0 means 32; tol bounds the residual of the result (1e-6 is the default of the host
 * layer).  info (device, nullable) receives 4 ints per matrix, valid in stream order after the call:
 *   info[4i+0] iterations applied, info[4i+1] status (0 = converged or at the fp32 floor of the matrix, 1 = max_iters
 *   reached without convergence, 2 = iterate became non-finite: singular or non-finite input -- torch.linalg.inv
 *   raises in that case), info[4i+2] last residual (float bits), info[4i+3] reserved.
 * Host arrays F, ns, out hold `count` device pointers / sizes; F is not modified; out[i] may not alias F[i]. */
size_t kfac_inverse_batched_workspace_bytes(const int* ns, int count);
int kfac_inverse_newton_schulz_batched(const float* const* F, const int* ns, int count, float damping,
                                       float* const* out, int max_iters, float tol, int* info, void* ws,
                                       size_t ws_bytes, kfac_stream_t stream);

/* One matrix, with the iteration count handed back on the host (iters_out, nullable).  Unlike every other entry point
 * this one SYNCHRONISES the stream (it has to read the count back) and returns non-zero -- with the residual in
 * kfac_last_error() -- when the iteration did not converge or the input was singular (torch.linalg.inv raises there,
 * kfac/layers/inverse.py:201).  Use the batched call above inside a training step. */
size_t kfac_inverse_workspace_bytes(int n);
int kfac_inverse_newton_schulz(const float* F, int n, float damping, float* out, int max_iters, float tol,
                               int* iters_out, void* ws, size_t ws_bytes, kfac_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* KFAC_B200_H_ */
