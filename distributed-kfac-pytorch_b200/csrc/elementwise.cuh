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
// pack: row i of the upper triangle is one contiguous run in F and in `packed`; a block copies whole rows.
__global__ void __launch_bounds__(EW_THREADS) triu_pack_kernel(const float* __restrict__ F, int n,
                                                                float* __restrict__ packed) {
  for (int i = blockIdx.x; i < n; i += gridDim.x) {
    const float* src = F + (long long)i * n + i;
    float* dst = packed + triu_index(n, i, i);
    for (int t = threadIdx.x; t < n - i; t += EW_THREADS) dst[t] = __ldg(src + t);
  }
}
// unpack: 64 x 64 tiles of the upper triangle are read once (coalesced along the packed rows) and written twice,
// as they are and transposed through shared memory, so the lower triangle is written coalesced as well.  Sixteen
// loads per thread are in flight before the first barrier of a tile.
constexpr int TRIU_T = 64;
__global__ void __launch_bounds__(EW_THREADS) triu_unpack_kernel(float* __restrict__ F, int n,
                                                                  const float* __restrict__ packed) {
  __shared__ float t[TRIU_T][TRIU_T + 1];
  const int nt = (n + TRIU_T - 1) / TRIU_T;
  const long long ntiles = (long long)nt * (nt + 1) / 2;
  const int x = threadIdx.x & (TRIU_T - 1), y0 = threadIdx.x / TRIU_T;   // 4 rows of the tile per pass
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    // tile -> (bi <= bj): row bi of the tile triangle starts at bi*nt - bi(bi-1)/2
    int bi = (int)((2.0 * nt + 1.0 - sqrt((2.0 * nt + 1.0) * (2.0 * nt + 1.0) - 8.0 * (double)tile)) * 0.5);
    while ((long long)bi * nt - (long long)bi * (bi - 1) / 2 > tile) --bi;
    while ((long long)(bi + 1) * nt - (long long)(bi + 1) * bi / 2 <= tile) ++bi;
    const int bj = bi + (int)(tile - ((long long)bi * nt - (long long)bi * (bi - 1) / 2));
    const int i0 = bi * TRIU_T, j0 = bj * TRIU_T;
    __syncthreads();
#pragma unroll
    for (int y = y0; y < TRIU_T; y += EW_THREADS / TRIU_T) {
      const int i = i0 + y, j = j0 + x;
      float v = 0.f;
      if (i < n && j < n && i <= j) v = __ldg(packed + triu_index(n, i, j));
      t[y][x] = v;
    }
    __syncthreads();
#pragma unroll
    for (int y = y0; y < TRIU_T; y += EW_THREADS / TRIU_T) {
      const int i = i0 + y, j = j0 + x;
      if (i < n && j < n && i <= j) F[(long long)i * n + j] = t[y][x];
      const int r = j0 + y, c = i0 + x;                    // transposed tile: F[r][c] = U[c][r], c <= r
      if (r < n && c < n && c < r) F[(long long)r * n + c] = t[x][y];
    }
  }
}
static int launch_triu(float* F, int n, float* packed, bool unpack, cudaStream_t stream) {
  if (unpack) {
    const long long nt = (n + TRIU_T - 1) / TRIU_T;
    triu_unpack_kernel<<<ew_blocks(nt * (nt + 1) / 2 * EW_THREADS), EW_THREADS, 0, stream>>>(F, n, packed);
  } else {
    triu_pack_kernel<<<ew_blocks((long long)n * EW_THREADS), EW_THREADS, 0, stream>>>(F, n, packed);
  }
  KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// out[i, j] = 1 / (dg[i] da[j] + damping): write-only.  Rows that allow 16-byte stores are cut into chunks of 1024
// floats (one float4 per thread) and the chunks are dealt round-robin to the CTAs -- equal work per CTA whatever the
// number of rows; (row, chunk) is carried along the walk instead of divided out.  Otherwise one row per block pass.
__global__ void __launch_bounds__(EW_THREADS) eigen_outer_kernel(const float* __restrict__ dg,
                                                                  const float* __restrict__ da, int g, int a,
                                                                  float damping, float* __restrict__ out, int vec_ok) {
  if (vec_ok) {
    const int cpr = (a + 4 * EW_THREADS - 1) / (4 * EW_THREADS);   // chunks per row
    const int di = (int)(gridDim.x / cpr), dc = (int)(gridDim.x % cpr);
    int i = (int)(blockIdx.x / cpr), c = (int)(blockIdx.x % cpr);
    while (i < g) {
      const int j = (c * EW_THREADS + threadIdx.x) * 4;
      if (j < a) {
        const float d = __ldg(dg + i);
        const float4 x = __ldg(reinterpret_cast<const float4*>(da + j));
        *reinterpret_cast<float4*>(out + (long long)i * a + j) =
            make_float4(1.f / (d * x.x + damping), 1.f / (d * x.y + damping), 1.f / (d * x.z + damping),
                        1.f / (d * x.w + damping));
      }
      i += di;
      c += dc;
      if (c >= cpr) { c -= cpr; ++i; }
    }
    return;
  }
  for (int i = blockIdx.x; i < g; i += gridDim.x) {
    const float d = __ldg(dg + i);
    float* row = out + (long long)i * a;
    for (int j = threadIdx.x; j < a; j += EW_THREADS) row[j] = 1.f / (d * __ldg(da + j) + damping);
  }
}
static int launch_eigen_outer(const float* dg, const float* da, int g, int a, float damping, float* out,
                              cudaStream_t stream) {
  const int vec_ok = (a % 4 == 0) && ((reinterpret_cast<uintptr_t>(da) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  eigen_outer_kernel<<<ew_blocks(vec_ok ? (long long)g * ((a + 4 * EW_THREADS - 1) / (4 * EW_THREADS)) * EW_THREADS : (long long)g * EW_THREADS), EW_THREADS, 0, stream>>>(dg, da, g, a, damping, out, vec_ok); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// sum(v[r, c] * [Wg | bg][r, c]): 8 B/elt read, double accumulation, one atomic per block.
// Without a bias column both operands are flat arrays: 16-byte loads, four independent pairs in flight per thread.
// With one, element e of v maps to (r, c) = (e / a, e % a); the grid-stride walk carries (r, c) along instead of
// dividing per element.
__device__ __forceinline__ void grad_dot_finish(double acc, double* out) {
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
__global__ void __launch_bounds__(EW_THREADS) grad_dot_flat_kernel(const float* __restrict__ v,
                                                                    const float* __restrict__ wg, long long total,
                                                                    int vec_ok, double* out) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double acc = 0.0;
  if (vec_ok) {
    const float4* v4 = reinterpret_cast<const float4*>(v);
    const float4* w4 = reinterpret_cast<const float4*>(wg);
    const long long n4 = total >> 2;
    for (long long e = t0; e < n4; e += 4 * stride) {
      float4 a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = e + u * stride;
        a[u] = i < n4 ? __ldg(v4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        b[u] = i < n4 ? __ldg(w4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        acc += (double)(a[u].x * b[u].x) + (double)(a[u].y * b[u].y) + (double)(a[u].z * b[u].z) +
               (double)(a[u].w * b[u].w);
    }
  } else {
    for (long long e = t0; e < total; e += stride) acc += (double)(__ldg(v + e) * __ldg(wg + e));
  }
  grad_dot_finish(acc, out);
}
__global__ void __launch_bounds__(EW_THREADS) grad_dot_kernel(const float* __restrict__ v,
                                                               const float* __restrict__ wg,
                                                               const float* __restrict__ bg, int g, int a_w,
                                                               double* out) {
  const int a = a_w + 1;
  const long long total = (long long)g * a;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const int dr = (int)(stride / a), dc = (int)(stride - (long long)dr * a);
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int r = (int)(e / a), c = (int)(e - (long long)r * a);
  double acc = 0.0;
  for (; e < total; e += stride) {
    const float gval = (c < a_w) ? __ldg(wg + (long long)r * a_w + c) : __ldg(bg + r);
    acc += (double)(__ldg(v + e) * gval);
    r += dr;
    c += dc;
    if (c >= a) {
      c -= a;
      ++r;
    }
  }
  grad_dot_finish(acc, out);
}
static int launch_grad_dot(const float* v, const float* wg, const float* bg, int g, int a_w, double* out,
                           cudaStream_t stream) {
  const long long total = (long long)g * (a_w + (bg != nullptr ? 1 : 0));
  if (bg == nullptr) {
    const int vec_ok = (total % 4 == 0) && ((reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(wg)) & 15) == 0;
    grad_dot_flat_kernel<<<ew_blocks(vec_ok ? total / 16 : total), EW_THREADS, 0, stream>>>(v, wg, total, vec_ok, out);
  } else {
    grad_dot_kernel<<<ew_blocks(total), EW_THREADS, 0, stream>>>(v, wg, bg, g, a_w, out);
  }
  KFAC_COUNT_LAUNCH(1);
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

// w_grad = s * pg[:, :a_w], b_grad = s * pg[:, a_w].  8 B/elt.  Same two shapes as grad_dot.
__global__ void __launch_bounds__(EW_THREADS) scale_write_flat_kernel(const float* __restrict__ pg, long long total,
                                                                       const float* __restrict__ scale_dev,
                                                                       float* __restrict__ wg, int vec_ok) {
  const float s = scale_dev != nullptr ? scale_dev[0] : 1.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec_ok) {
    const float4* p4 = reinterpret_cast<const float4*>(pg);
    float4* w4 = reinterpret_cast<float4*>(wg);
    const long long n4 = total >> 2;
    for (long long e = t0; e < n4; e += 4 * stride) {
      float4 a[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = e + u * stride;
        if (i < n4) a[u] = __ldg(p4 + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = e + u * stride;
        if (i < n4) w4[i] = make_float4(s * a[u].x, s * a[u].y, s * a[u].z, s * a[u].w);
      }
    }
  } else {
    for (long long e = t0; e < total; e += stride) wg[e] = s * __ldg(pg + e);
  }
}
__global__ void __launch_bounds__(EW_THREADS) scale_write_kernel(const float* __restrict__ pg, int g, int a_w,
                                                                  const float* __restrict__ scale_dev,
                                                                  float* __restrict__ wg, float* __restrict__ bg) {
  const float s = scale_dev != nullptr ? scale_dev[0] : 1.f;
  const int a = a_w + 1;
  const long long total = (long long)g * a;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const int dr = (int)(stride / a), dc = (int)(stride - (long long)dr * a);
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int r = (int)(e / a), c = (int)(e - (long long)r * a);
  for (; e < total; e += stride) {
    const float val = s * __ldg(pg + e);
    if (c < a_w)
      wg[(long long)r * a_w + c] = val;
    else
      bg[r] = val;
    r += dr;
    c += dc;
    if (c >= a) {
      c -= a;
      ++r;
    }
  }
}
static int launch_scale_write(const float* pg, int g, int a_w, const float* scale_dev, float* wg, float* bg,
                              cudaStream_t stream) {
  const long long total = (long long)g * (a_w + (bg != nullptr ? 1 : 0));
  if (bg == nullptr) {
    const int vec_ok = (total % 4 == 0) && ((reinterpret_cast<uintptr_t>(pg) | reinterpret_cast<uintptr_t>(wg)) & 15) == 0;
    scale_write_flat_kernel<<<ew_blocks(vec_ok ? total / 16 : total), EW_THREADS, 0, stream>>>(pg, total, scale_dev, wg, vec_ok);
  } else {
    scale_write_kernel<<<ew_blocks(total), EW_THREADS, 0, stream>>>(pg, g, a_w, scale_dev, wg, bg);
  }
  KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// All layers of a step in one launch: blockIdx.y walks the layers, each block row grid-strides over its layer.
struct ScaleWriteItem {
  const float* pg;
  float* wg;
  float* bg;   // may be null
  int g, a_w;
};
__global__ void __launch_bounds__(EW_THREADS) scale_write_group_kernel(const ScaleWriteItem* __restrict__ items,
                                                                        const float* __restrict__ scale_dev) {
  const ScaleWriteItem it = items[blockIdx.y];
  const float s = scale_dev != nullptr ? scale_dev[0] : 1.f;
  const int a = it.a_w + (it.bg != nullptr ? 1 : 0);
  const long long total = (long long)it.g * a;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (it.bg == nullptr) {
    const bool vec = (total % 4 == 0) && ((reinterpret_cast<uintptr_t>(it.pg) | reinterpret_cast<uintptr_t>(it.wg)) & 15) == 0;
    if (vec) {
      const float4* p4 = reinterpret_cast<const float4*>(it.pg);
      float4* w4 = reinterpret_cast<float4*>(it.wg);
      for (const long long n4 = total >> 2; e < n4; e += stride) {
        const float4 v = __ldg(p4 + e);
        w4[e] = make_float4(s * v.x, s * v.y, s * v.z, s * v.w);
      }
    } else {
      for (; e < total; e += stride) it.wg[e] = s * __ldg(it.pg + e);
    }
    return;
  }
  const int dr = (int)(stride / a), dc = (int)(stride - (long long)dr * a);
  int r = (int)(e / a), c = (int)(e - (long long)r * a);
  for (; e < total; e += stride) {
    const float val = s * __ldg(it.pg + e);
    if (c < it.a_w)
      it.wg[(long long)r * it.a_w + c] = val;
    else
      it.bg[r] = val;
    r += dr;
    c += dc;
    if (c >= a) {
      c -= a;
      ++r;
    }
  }
}

}  // namespace kfac
