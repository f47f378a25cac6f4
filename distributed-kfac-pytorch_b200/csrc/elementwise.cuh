// Memory-bound passes of the K-FAC step: 128-bit coalesced accesses, grid-stride loops sized to the SM
// count, warp-shuffle reductions.  Each is bounded by HBM bandwidth (bytes moved are listed per kernel).
#pragma once
#include <cstdint>
#include "common.cuh"

namespace kfac {

constexpr int EW_THREADS = 256;
constexpr int EW_MAX_BLOCKS = 148 * 8;

static inline unsigned ew_blocks(long long work_items) {
  long long b = (work_items + EW_THREADS - 1) / EW_THREADS;
  if (b > EW_MAX_BLOCKS) b = EW_MAX_BLOCKS;
  if (b < 1) b = 1;
  return (unsigned)b;
}

// y = a*x + b*y.  12 B/elt (8 B/elt when b == 0 or x == y).
__global__ void __launch_bounds__(EW_THREADS) axpby_kernel(float* __restrict__ y, const float* __restrict__ x, float a,
                                                           float b, long long n, int vec_ok) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec_ok) {
    const long long n4 = n >> 2;
    float4* y4 = reinterpret_cast<float4*>(y);
    const float4* x4 = reinterpret_cast<const float4*>(x);
    for (long long i = tid; i < n4; i += stride) {
      const float4 xv = x4[i];
      float4 yv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b != 0.f) yv = y4[i];
      yv.x = a * xv.x + b * yv.x;
      yv.y = a * xv.y + b * yv.y;
      yv.z = a * xv.z + b * yv.z;
      yv.w = a * xv.w + b * yv.w;
      y4[i] = yv;
    }
    for (long long i = (n4 << 2) + tid; i < n; i += stride) y[i] = a * x[i] + (b != 0.f ? b * y[i] : 0.f);
  } else {
    for (long long i = tid; i < n; i += stride) y[i] = a * x[i] + (b != 0.f ? b * y[i] : 0.f);
  }
}

// In-place variant (x aliases y): y = (a + b) * y, kept separate so the aliasing-free kernel can use __restrict__.
__global__ void __launch_bounds__(EW_THREADS) scale_inplace_kernel(float* y, float s, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) y[i] *= s;
}

static int launch_axpby(float* y, const float* x, float a, float b, long long n, cudaStream_t stream) {
  if (x == y) {
    scale_inplace_kernel<<<ew_blocks(n), EW_THREADS, 0, stream>>>(y, a + b, n); KFAC_COUNT_LAUNCH(1);
  } else {
    const int vec_ok = ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(x)) & 15) == 0;
    axpby_kernel<<<ew_blocks(vec_ok ? n / 4 : n), EW_THREADS, 0, stream>>>(y, x, a, b, n, vec_ok); KFAC_COUNT_LAUNCH(1);
  }
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

__global__ void __launch_bounds__(EW_THREADS) set_identity_kernel(float* F, int n) {
  const long long total = (long long)n * n;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int r = (int)(i / n);
    const int c = (int)(i - (long long)r * n);
    F[i] = (r == c) ? 1.f : 0.f;
  }
}
static int launch_set_identity(float* F, int n, cudaStream_t stream) {
  set_identity_kernel<<<ew_blocks((long long)n * n), EW_THREADS, 0, stream>>>(F, n); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

__global__ void add_diag_kernel(float* F, int n, float damping) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) F[(long long)i * n + i] += damping;
}
static int launch_add_diag(float* F, int n, float damping, cudaStream_t stream) {
  add_diag_kernel<<<(n + 255) / 256, 256, 0, stream>>>(F, n, damping); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// Row-major upper triangle <-> packed vector; offset of (i, j >= i) is i*n - i(i-1)/2 + (j - i).
__device__ __forceinline__ long long triu_index(int n, int i, int j) {
  return (long long)i * n - ((long long)i * (i - 1)) / 2 + (j - i);
}
__global__ void __launch_bounds__(EW_THREADS) triu_kernel(float* F, int n, float* packed, int unpack) {
  const long long total = (long long)n * n;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int i = (int)(e / n);
    const int j = (int)(e - (long long)i * n);
    if (unpack) {
      F[e] = (i <= j) ? packed[triu_index(n, i, j)] : packed[triu_index(n, j, i)];
    } else if (i <= j) {
      packed[triu_index(n, i, j)] = F[e];
    }
  }
}
static int launch_triu(float* F, int n, float* packed, bool unpack, cudaStream_t stream) {
  triu_kernel<<<ew_blocks((long long)n * n), EW_THREADS, 0, stream>>>(F, n, packed, unpack ? 1 : 0); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

__global__ void __launch_bounds__(EW_THREADS) eigen_outer_kernel(const float* __restrict__ dg,
                                                                  const float* __restrict__ da, int g, int a,
                                                                  float damping, float* __restrict__ out) {
  const long long total = (long long)g * a;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int i = (int)(e / a);
    const int j = (int)(e - (long long)i * a);
    out[e] = 1.f / (dg[i] * da[j] + damping);
  }
}
static int launch_eigen_outer(const float* dg, const float* da, int g, int a, float damping, float* out,
                              cudaStream_t stream) {
  eigen_outer_kernel<<<ew_blocks((long long)g * a), EW_THREADS, 0, stream>>>(dg, da, g, a, damping, out); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// sum(v[r, c] * [Wg | bg][r, c]): 8 B/elt read, double accumulation, one atomic per block.
__global__ void __launch_bounds__(EW_THREADS) grad_dot_kernel(const float* __restrict__ v,
                                                               const float* __restrict__ wg,
                                                               const float* __restrict__ bg, int g, int a_w,
                                                               double* out) {
  const int a = a_w + (bg != nullptr ? 1 : 0);
  const long long total = (long long)g * a;
  const long long stride = (long long)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int r = (int)(e / a);
    const int c = (int)(e - (long long)r * a);
    const float gval = (c < a_w) ? wg[(long long)r * a_w + c] : bg[r];
    acc += (double)(v[e] * gval);
  }
  __shared__ double red[EW_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < EW_THREADS / 32; ++w) s += red[w];
    atomicAdd(out, s);
  }
}
static int launch_grad_dot(const float* v, const float* wg, const float* bg, int g, int a_w, double* out,
                           cudaStream_t stream) {
  const long long total = (long long)g * (a_w + (bg != nullptr ? 1 : 0));
  unsigned blocks = ew_blocks(total);
  if (blocks > 148u * 2u) blocks = 148u * 2u;
  grad_dot_kernel<<<blocks, EW_THREADS, 0, stream>>>(v, wg, bg, g, a_w, out); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

__global__ void klclip_scale_kernel(const double* vg, float lr, float kl_clip, float* scale_out) {
  const double vg_sum = vg[0] * (double)lr * (double)lr;
  float s = 1.f;
  if (vg_sum != 0.0) {
    const double r = sqrt((double)kl_clip / fabs(vg_sum));
    s = r < 1.0 ? (float)r : 1.f;
  }
  scale_out[0] = s;
}
static int launch_klclip_scale(const double* vg, float lr, float kl_clip, float* scale_out, cudaStream_t stream) {
  klclip_scale_kernel<<<1, 1, 0, stream>>>(vg, lr, kl_clip, scale_out); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// w_grad = s * pg[:, :a_w], b_grad = s * pg[:, a_w].  8 B/elt.
__global__ void __launch_bounds__(EW_THREADS) scale_write_kernel(const float* __restrict__ pg, int g, int a_w,
                                                                  const float* __restrict__ scale_dev,
                                                                  float* __restrict__ wg, float* __restrict__ bg) {
  const float s = scale_dev != nullptr ? scale_dev[0] : 1.f;
  const int a = a_w + (bg != nullptr ? 1 : 0);
  const long long total = (long long)g * a;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int r = (int)(e / a);
    const int c = (int)(e - (long long)r * a);
    const float val = s * pg[e];
    if (c < a_w)
      wg[(long long)r * a_w + c] = val;
    else
      bg[r] = val;
  }
}
static int launch_scale_write(const float* pg, int g, int a_w, const float* scale_dev, float* wg, float* bg,
                              cudaStream_t stream) {
  const long long total = (long long)g * (a_w + (bg != nullptr ? 1 : 0));
  scale_write_kernel<<<ew_blocks(total), EW_THREADS, 0, stream>>>(pg, g, a_w, scale_dev, wg, bg); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kfac
