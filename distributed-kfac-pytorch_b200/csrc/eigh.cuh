// Batched symmetric eigensolver and Newton-Schulz inverse for K-FAC factors (sm_100a).
// Replaces torch.linalg.eigh / torch.linalg.inv at kfac/layers/eigen.py:309-343 and kfac/layers/inverse.py:185-212.
//
//   stage 1  sytrd_kernel      Householder tridiagonalisation, one persistent cooperative launch per round; each
//                              matrix is owned by a group of CTAs (rows dealt cyclically) that meet at ONE global
//                              barrier per column.  The rank-2 update of column k-1 is fused with the symv of
//                              column k, so the trailing matrix is read and written once per column (8 B/elt).
//   stage 2  dc_* kernels      Cuppen divide & conquer with 1x1 leaves, all merges of one tree level of all
//                              matrices per launch; secular equation + Loewner z-hat in fp64 (dc_core.h), deflation
//                              rotations folded into the merge matrix U, Q <- Q U on the tensor cores (gemm3x).
//   stage 3  back-transform    Q <- H Q with 128-wide compact-WY blocks: four gemm3x calls per block.
//
// oracle/eigh_model.py is the numpy model of exactly this data flow.
#pragma once
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "dc_core.h"
#include "elementwise.cuh"
#include "gemm3x_launch.cuh"

namespace kfac {

// ================================================================= stage 1: tridiagonalisation
struct SytrdJob {
  float* W;       // n x ldw working copy (ldw = n rounded up to 4: rows are 16-byte aligned); on exit row k holds
                  // v_k in W[k, k+1:]
  int n;
  int ldw;
  float* d;       // n
  float* e;       // n (n-1 used)
  float* tau;     // n (n-1 used)
  float* pbuf;    // 3 * n (zeroed before launch; triple-buffered accumulation target of the symmetric pass)
  float* colbuf;  // 2 * n: column k+1 of the trailing matrix, handed from pass k to step k+1 (symmetric path)
  int sym;        // 1: lower-triangle-only path (needs 4 | n)
  int slots;      // bulk-fed pass: row slots per warp (2 normally, 1 when that is what lets three CTAs share an SM)
  unsigned* bar;  // group barrier counter (zeroed before launch)
  int cta0, ncta;
};

// Two 8-warp CTAs per SM: while one matrix's group sits in its per-column vector phase or barrier, the CTA of another
// matrix on the same SM streams its pass (the fixed per-column cost was ~45% of the tridiagonalisation before).
constexpr int SY_THREADS = 256;
constexpr int SY_WARPS = SY_THREADS / 32;
constexpr int SY_CTAS_PER_SM = 2;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void group_barrier(unsigned* bar, unsigned target) {
  // bar.sync orders the CTA's writes before thread 0's release; the release/acquire pair is cumulative, so no
  // separate (and much slower) membar.gl + L1 invalidate is needed.  All cross-CTA data is read with ld.cg.
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
    while (ld_acquire_u32(bar) < target) __nanosleep(64);   // keep the spinning lane out of the co-resident CTA's issue slots
  }
  __syncthreads();
}

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];  // same order in every CTA: bitwise identical
  return s;
}


// Strided vector read with all loads of a chunk issued before any is used: the per-column vector loops of the
// tridiagonalisation would otherwise pay one dependent L2 round trip per 512 elements.
constexpr int SY_VB = 8;
template <int NT = 256, typename F>
__device__ __forceinline__ void strided_ldcg(const float* __restrict__ src, int begin, int end, int tid, F&& consume) {
  for (int base = begin + tid; base < end; base += SY_VB * NT) {
    float v[SY_VB];
#pragma unroll
    for (int u = 0; u < SY_VB; ++u) {
      const int j = base + u * NT;
      v[u] = (j < end) ? __ldcg(src + j) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < SY_VB; ++u) {
      const int j = base + u * NT;
      if (j < end) consume(j, v[u]);
    }
  }
}

// Two vectors over the same index range with both batches of loads in flight together (one exposed L2 round trip).
template <int NT = 256, typename F>
__device__ __forceinline__ void strided_ldcg2(const float* __restrict__ s1, const float* __restrict__ s2, int begin,
                                              int end, int tid, F&& consume) {
  for (int base = begin + tid; base < end; base += SY_VB * NT) {
    float v1[SY_VB], v2[SY_VB];
#pragma unroll
    for (int u = 0; u < SY_VB; ++u) {
      const int j = base + u * NT;
      v1[u] = (j < end) ? __ldcg(s1 + j) : 0.f;
      v2[u] = (j < end) ? __ldcg(s2 + j) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < SY_VB; ++u) {
      const int j = base + u * NT;
      if (j < end) consume(j, v1[u], v2[u]);
    }
  }
}

__device__ __forceinline__ void sytrd_full_body(const SytrdJob& job, float* sy_smem, double* red) {
  const int n = job.n;
  const int ldw = job.ldw;
  const int rank = blockIdx.x - job.cta0;
  const int G = job.ncta;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* W = job.W;
  float* va = sy_smem;          // pending update vectors (absolute index)
  float* wa = sy_smem + n;
  float* vc = sy_smem + 2 * n;  // current Householder vector
  float* wc = sy_smem + 3 * n;
  bool have_prev = false;
  unsigned nbar = 0;

  for (int k = 0; k + 1 < n; ++k) {
    // ---- part 1: row k of the (virtually) updated trailing matrix -> Householder vector (redundant per CTA)
    // (for k > 0 the raw row was brought into vc[] by part 3 of the previous step, together with p)
    const float vk = have_prev ? va[k] : 0.f, wk = have_prev ? wa[k] : 0.f;
    double nrm2 = 0.0;
    if (k == 0) {
      strided_ldcg(W, 0, n, tid, [&](int j, float r) {
        vc[j] = r;
        if (j >= 2) nrm2 += (double)r * (double)r;
      });
    } else {
      for (int j = k + tid; j < n; j += SY_THREADS) {
        const float r = vc[j] - (vk * wa[j] + wk * va[j]);
        vc[j] = r;
        if (j >= k + 2) nrm2 += (double)r * (double)r;
      }
    }
    nrm2 = block_sum(nrm2, red);  // (also orders the vc[] writes)
    const float dk = vc[k];
    const float alpha = vc[k + 1];
    float beta, tau;
    if (nrm2 == 0.0) {
      beta = alpha;
      tau = 0.f;
    } else {
      const double nb = sqrt((double)alpha * (double)alpha + nrm2);
      beta = (float)(alpha >= 0.f ? -nb : nb);
      tau = (beta - alpha) / beta;
    }
    __syncthreads();
    const float scal = (tau != 0.f) ? 1.f / (alpha - beta) : 0.f;
    for (int j = k + 1 + tid; j < n; j += SY_THREADS) vc[j] = (j == k + 1) ? 1.f : vc[j] * scal;
    __syncthreads();

    // ---- part 2: own rows: apply the pending rank-2 update, write back, p_i = (A' v)_i
    // Rows i == rank (mod G) belong to this CTA; warps take them round-robin.  The pass is bandwidth bound, so
    // each lane keeps 8 independent 16-byte loads in flight (4 KB per warp, 64 KB per SM); rows are 16-byte
    // aligned when 4 | n, otherwise a scalar path with the same depth is used.
    float* pb = job.pbuf + (size_t)(k & 1) * n;
    const int jbeg = k + 1;
    const bool vec = (n & 3) == 0;
    const int jal = vec ? ((jbeg + 3) & ~3) : n;  // first 16-byte aligned column (vector path)
    for (int i = k + 1 + ((rank - (k + 1)) % G + G) % G + warp * G; i < n; i += G * SY_WARPS) {
      float* row = W + (size_t)i * ldw;
      const float vi = have_prev ? va[i] : 0.f, wi = have_prev ? wa[i] : 0.f;
      float acc = 0.f;
      if (vec) {
        // scalar head (< 4 elements)
        if (jbeg + lane < jal) {
          const int j = jbeg + lane;
          float a0 = __ldcg(row + j);
          if (have_prev) {
            a0 -= vi * wa[j] + wi * va[j];
            row[j] = a0;
          }
          acc += a0 * vc[j];
        }
        int j = jal + 4 * lane;
        for (; j + 7 * 128 < n; j += 8 * 128) {
          float4 a[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) a[u] = __ldcg(reinterpret_cast<const float4*>(row + j + 128 * u));
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int jj = j + 128 * u;
            const float4 c4 = *reinterpret_cast<const float4*>(vc + jj);
            if (have_prev) {
              const float4 w4 = *reinterpret_cast<const float4*>(wa + jj);
              const float4 v4 = *reinterpret_cast<const float4*>(va + jj);
              a[u].x -= vi * w4.x + wi * v4.x;
              a[u].y -= vi * w4.y + wi * v4.y;
              a[u].z -= vi * w4.z + wi * v4.z;
              a[u].w -= vi * w4.w + wi * v4.w;
              *reinterpret_cast<float4*>(row + jj) = a[u];
            }
            acc += a[u].x * c4.x + a[u].y * c4.y + a[u].z * c4.z + a[u].w * c4.w;
          }
        }
        for (; j < n; j += 128) {  // 4 | n and 4 | j: a full float4 always fits
          float4 a0 = __ldcg(reinterpret_cast<const float4*>(row + j));
          const float4 c4 = *reinterpret_cast<const float4*>(vc + j);
          if (have_prev) {
            const float4 w4 = *reinterpret_cast<const float4*>(wa + j);
            const float4 v4 = *reinterpret_cast<const float4*>(va + j);
            a0.x -= vi * w4.x + wi * v4.x;
            a0.y -= vi * w4.y + wi * v4.y;
            a0.z -= vi * w4.z + wi * v4.z;
            a0.w -= vi * w4.w + wi * v4.w;
            *reinterpret_cast<float4*>(row + j) = a0;
          }
          acc += a0.x * c4.x + a0.y * c4.y + a0.z * c4.z + a0.w * c4.w;
        }
      } else {
        int j = jbeg + lane;
        for (; j + 7 * 32 < n; j += 8 * 32) {
          float a[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) a[u] = __ldcg(row + j + 32 * u);
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int jj = j + 32 * u;
            if (have_prev) {
              a[u] -= vi * wa[jj] + wi * va[jj];
              row[jj] = a[u];
            }
            acc += a[u] * vc[jj];
          }
        }
        for (; j < n; j += 32) {
          float a0 = __ldcg(row + j);
          if (have_prev) {
            a0 -= vi * wa[j] + wi * va[j];
            row[j] = a0;
          }
          acc += a0 * vc[j];
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) pb[i] = acc;
    }
    nbar += G;
    group_barrier(job.bar, nbar);

    // ---- part 3: w = tau p - (tau/2)(tau p . v) v   (redundant per CTA)
    double dot = 0.0;
    {
      // p of this step and the raw next row (written by its owner before the barrier) in one batch of loads; the row
      // goes to the storage that becomes vc after the swap below (the old pair is dead once the pass is over)
      const float* rown = W + (size_t)(k + 1) * ldw;
      strided_ldcg2(pb, rown, k + 1, n, tid, [&](int i, float pv, float rv) {
        const float pt = tau * pv;
        wc[i] = pt;
        dot += (double)pt * (double)vc[i];
        va[i] = rv;
      });
    }
    dot = block_sum(dot, red);
    const float coef = 0.5f * tau * (float)dot;
    for (int i = k + 1 + tid; i < n; i += SY_THREADS) wc[i] -= coef * vc[i];
    if (rank == 0) {
      // row k is dead from now on: store v_k there (stage 3 reads it), and the tridiagonal entries
      float* vrow = W + (size_t)k * ldw;
      for (int j = k + 1 + tid; j < n; j += SY_THREADS) vrow[j] = vc[j];
      if (tid == 0) {
        job.d[k] = dk;
        job.e[k] = beta;
        job.tau[k] = tau;
      }
    }
    __syncthreads();
    float* t = va; va = vc; vc = t;
    t = wa; wa = wc; wc = t;
    have_prev = true;
  }
  if (rank == 0 && tid == 0) {
    // last diagonal entry: fully updated already (the last reflector has tau = 0), written by its owner before the
    // final barrier; n == 1 has no steps at all
    job.d[n - 1] = __ldcg(W + (size_t)(n - 1) * ldw + (n - 1));
  }
}


// ---- symmetric variant: only the lower triangle of the trailing matrix is kept up to date and read, which halves
// the memory traffic of the pass (4 B per trailing element per column instead of 8).  The trailing matrix is walked
// in column blocks of 512; the rows of a block are split evenly over the group's CTAs.  A lane owns four 16-byte
// column chunks, so the column part of the symmetric product (p_j += A_ij v_i, j < i) accumulates in 16 registers
// per lane over ALL rows the warp streams, is combined across the CTA's warps in shared memory and flushed with one
// global add per column per CTA (a first version flushed per 8-row tile and drowned in atomics, a second one
// synchronised per 64-row tile and was latency bound); the row part (p_i += sum_{j<=i} A_ij v_j) is a warp
// reduction.  p is accumulated with global atomics (order varies, every CTA reads the same totals).
// Column k+1 of the updated matrix is written to `colbuf` during pass k so that step k+1 starts from contiguous data.
constexpr int SY_TW = 512;  // column block width

__device__ __forceinline__ void sytrd_sym_body(const SytrdJob& job, float* sy_smem, double* red) {
  const int n = job.n;
  const int ldw = job.ldw;   // == n rounded up to a multiple of 4; also the stride of the shared vectors
  const int rank = blockIdx.x - job.cta0;
  const int G = job.ncta;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* W = job.W;
  float* va = sy_smem;
  float* wa = sy_smem + ldw;
  float* vc = sy_smem + 2 * ldw;
  float* wc = sy_smem + 3 * ldw;
  if (threadIdx.x < 4 * (ldw - n)) {   // padding lanes of the four vectors: only ever touched by masked components
    const int v = threadIdx.x / (ldw - n), t = threadIdx.x % (ldw - n);
    sy_smem[v * ldw + n + t] = 0.f;
  }
  bool have_prev = false;
  unsigned nbar = 0;
  const int ncb = (n + SY_TW - 1) / SY_TW;

  for (int k = 0; k + 1 < n; ++k) {
    // ---- part 1: column k of the updated trailing matrix (contiguous in colbuf) -> Householder vector
    // (for k > 0 the raw column was brought into vc[] by part 3 of the previous step, together with p: one
    // exposed L2 round trip per column instead of two)
    const float vk = have_prev ? va[k] : 0.f, wk = have_prev ? wa[k] : 0.f;
    double nrm2 = 0.0;
    if (k == 0) {
      strided_ldcg(W, 0, n, tid, [&](int j, float r) {   // row 0 == column 0 of the input
        vc[j] = r;
        if (j >= 2) nrm2 += (double)r * (double)r;
      });
    } else {
      for (int j = k + tid; j < n; j += SY_THREADS) {
        const float r = vc[j] - (vk * wa[j] + wk * va[j]);
        vc[j] = r;
        if (j >= k + 2) nrm2 += (double)r * (double)r;
      }
    }
    nrm2 = block_sum(nrm2, red);
    const float dk = vc[k];
    const float alpha = vc[k + 1];
    float beta, tau;
    if (nrm2 == 0.0) {
      beta = alpha;
      tau = 0.f;
    } else {
      const double nb = sqrt((double)alpha * (double)alpha + nrm2);
      beta = (float)(alpha >= 0.f ? -nb : nb);
      tau = (beta - alpha) / beta;
    }
    __syncthreads();
    const float scal = (tau != 0.f) ? 1.f / (alpha - beta) : 0.f;
    for (int j = k + 1 + tid; j < n; j += SY_THREADS) vc[j] = (j == k + 1) ? 1.f : vc[j] * scal;
    __syncthreads();

    // ---- part 2: lower-triangle pass.  The work is cut into items of 4 consecutive rows x one 512-wide column block
    // (16 independent 16-byte loads per lane), enumerated column block by column block; every warp of the group takes
    // a contiguous, equally long run of items and streams it without any CTA-level synchronisation (an earlier
    // version split every column block over the CTAs and synchronised twice per block: with ~1.5 batches per warp
    // between barriers each block cost one fully exposed L2 round trip).  Column sums stay in 16 registers per lane
    // while the run stays inside one column block and are flushed with 16-byte vector reductions when it leaves it.
    float* pb = job.pbuf + (size_t)(k % 3) * ldw;
    float* cnext = job.colbuf + (size_t)((k + 1) & 1) * n;
    const int j0 = k + 1;
    const int cb0 = j0 / SY_TW;
    int t_all = 0;
    for (int cb = cb0; cb < ncb; ++cb) t_all += (n - max(j0, cb * SY_TW) + 3) >> 2;
    if (job.sym == 2) t_all = 0;
    const int gw = rank * SY_WARPS + warp;
    int it = (int)(((long long)gw * t_all) / (G * SY_WARPS));
    const int it_hi = (int)(((long long)(gw + 1) * t_all) / (G * SY_WARPS));
    int cb = cb0, base = 0;
    while (it < it_hi) {
      int rfirst = max(j0, cb * SY_TW);
      int nb = (n - rfirst + 3) >> 2;
      while (it >= base + nb) {
        base += nb;
        ++cb;
        rfirst = max(j0, cb * SY_TW);
        nb = (n - rfirst + 3) >> 2;
      }
      const int cb_end = min(it_hi, base + nb);
      const int cbase = cb * SY_TW + 4 * lane;
      const bool clean = have_prev && (cb * SY_TW >= j0 + 1) && ((cb + 1) * SY_TW <= n);
      float cacc[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) cacc[u] = 0.f;
      for (; it < cb_end; ++it) {
        const int i0 = rfirst + 4 * (it - base);
        float4 a[4][4];
        if (clean && i0 >= (cb + 1) * SY_TW && i0 + 3 < n) {
          // interior item: strictly below the diagonal block, right of the first live column -> no masks
          float* row0 = W + (size_t)i0 * ldw + cbase;
#pragma unroll
          for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int u = 0; u < 4; ++u)
              a[q][u] = __ldcg(reinterpret_cast<const float4*>(row0 + (size_t)q * ldw + 128 * u));
          float nvi[4], nwi[4], vci[4], racc[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            nvi[q] = -va[i0 + q];
            nwi[q] = -wa[i0 + q];
            vci[q] = vc[i0 + q];
            racc[q] = 0.f;
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int j = cbase + 128 * u;
            const float4 w4 = *reinterpret_cast<const float4*>(wa + j);
            const float4 v4 = *reinterpret_cast<const float4*>(va + j);
            const float4 c4 = *reinterpret_cast<const float4*>(vc + j);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float4 x = a[q][u];
              x.x = fmaf(nwi[q], v4.x, fmaf(nvi[q], w4.x, x.x));
              x.y = fmaf(nwi[q], v4.y, fmaf(nvi[q], w4.y, x.y));
              x.z = fmaf(nwi[q], v4.z, fmaf(nvi[q], w4.z, x.z));
              x.w = fmaf(nwi[q], v4.w, fmaf(nvi[q], w4.w, x.w));
              *reinterpret_cast<float4*>(row0 + (size_t)q * ldw + 128 * u) = x;
              racc[q] = fmaf(x.x, c4.x, fmaf(x.y, c4.y, fmaf(x.z, c4.z, fmaf(x.w, c4.w, racc[q]))));
              cacc[4 * u] = fmaf(x.x, vci[q], cacc[4 * u]);
              cacc[4 * u + 1] = fmaf(x.y, vci[q], cacc[4 * u + 1]);
              cacc[4 * u + 2] = fmaf(x.z, vci[q], cacc[4 * u + 2]);
              cacc[4 * u + 3] = fmaf(x.w, vci[q], cacc[4 * u + 3]);
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int q = 0; q < 4; ++q) racc[q] += __shfl_xor_sync(0xffffffffu, racc[q], o);
          }
          if (lane < 4) {
            const float r = lane == 0 ? racc[0] : lane == 1 ? racc[1] : lane == 2 ? racc[2] : racc[3];
            atomicAdd(pb + i0 + lane, r);
          }
          continue;
        }
        // general item: diagonal block, first live column block, ragged right / bottom edge
        bool on[4][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int i = i0 + q;
          const float* row = W + (size_t)i * ldw;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int j = cbase + 128 * u;
            on[q][u] = (i < n) && (j <= i) && (j + 3 >= j0) && (j < n);
            a[q][u] = on[q][u] ? __ldcg(reinterpret_cast<const float4*>(row + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int i = i0 + q;
          if (i >= n) continue;
          float* row = W + (size_t)i * ldw;
          const float vi = have_prev ? va[i] : 0.f, wi = have_prev ? wa[i] : 0.f;
          const float vci = vc[i];
          float racc = 0.f;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (!on[q][u]) continue;
            const int j = cbase + 128 * u;
            float x[4] = {a[q][u].x, a[q][u].y, a[q][u].z, a[q][u].w};
            if (have_prev) {
              const float4 w4 = *reinterpret_cast<const float4*>(wa + j);
              const float4 v4 = *reinterpret_cast<const float4*>(va + j);
              x[0] -= vi * w4.x + wi * v4.x;
              x[1] -= vi * w4.y + wi * v4.y;
              x[2] -= vi * w4.z + wi * v4.z;
              x[3] -= vi * w4.w + wi * v4.w;
              *reinterpret_cast<float4*>(row + j) = make_float4(x[0], x[1], x[2], x[3]);
            }
            const float4 c4 = *reinterpret_cast<const float4*>(vc + j);
            if (j >= j0 + 1 && j + 3 < i) {
              racc += x[0] * c4.x + x[1] * c4.y + x[2] * c4.z + x[3] * c4.w;
              cacc[4 * u] += x[0] * vci;
              cacc[4 * u + 1] += x[1] * vci;
              cacc[4 * u + 2] += x[2] * vci;
              cacc[4 * u + 3] += x[3] * vci;
            } else {
              const float cv[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const int jc = j + c;
                if (jc >= j0 && jc <= i) {
                  racc += x[c] * cv[c];
                  if (jc < i) cacc[4 * u + c] += x[c] * vci;
                  if (jc == j0) cnext[i] = x[c];
                }
              }
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) racc += __shfl_xor_sync(0xffffffffu, racc, o);
          if (lane == 0) atomicAdd(pb + i, racc);
        }
      }
      // leaving the column block: flush its column sums (pb rows are 16-byte aligned: stride ldw)
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = cbase + 128 * u;
        if (j < n && (cacc[4 * u] != 0.f || cacc[4 * u + 1] != 0.f || cacc[4 * u + 2] != 0.f || cacc[4 * u + 3] != 0.f))
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(pb + j), "f"(cacc[4 * u]),
                       "f"(cacc[4 * u + 1]), "f"(cacc[4 * u + 2]), "f"(cacc[4 * u + 3])
                       : "memory");
      }
    }
    nbar += G;
    group_barrier(job.bar, nbar);

    // ---- part 3: w = tau p - (tau/2)(tau p . v) v ; clear the accumulation buffer of step k+2
    {
      float* pz = job.pbuf + (size_t)((k + 2) % 3) * ldw;
      for (int i = rank * SY_THREADS + tid; i < ldw; i += G * SY_THREADS) pz[i] = 0.f;
    }
    double dot = 0.0;
    {
      // p of this step and the raw next column (complete since the barrier) in one batch of loads; the column goes
      // to the storage that becomes vc after the swap below (the old pair is dead once the pass is over)
      const float* coln = job.colbuf + (size_t)((k + 1) & 1) * n;
      strided_ldcg2(pb, coln, k + 1, n, tid, [&](int i, float pv, float cv) {
        const float pt = tau * pv;
        wc[i] = pt;
        dot += (double)pt * (double)vc[i];
        va[i] = cv;
      });
    }
    dot = block_sum(dot, red);
    const float coef = 0.5f * tau * (float)dot;
    for (int i = k + 1 + tid; i < n; i += SY_THREADS) wc[i] -= coef * vc[i];
    if (rank == 0) {
      float* vrow = W + (size_t)k * ldw;   // upper triangle is free in this variant: row k stores v_k for stage 3
      for (int j = k + 1 + tid; j < n; j += SY_THREADS) vrow[j] = vc[j];
      if (tid == 0) {
        job.d[k] = dk;
        job.e[k] = beta;
        job.tau[k] = tau;
      }
    }
    __syncthreads();
    float* t = va; va = vc; vc = t;
    t = wa; wa = wc; wc = t;
    have_prev = (job.sym != 3);
  }
  if (rank == 0 && tid == 0) {
    // the last reflector has tau = 0, so the (n-1, n-1) entry written to colbuf by the final pass is final
    job.d[n - 1] = (n == 1) ? __ldcg(W) : __ldcg(job.colbuf + (size_t)((n - 1) & 1) * n + (n - 1));
  }
}

// ---- symmetric variant fed by the bulk-copy engine.  LSU-issued loads (LDG or LDGSTS) top out at ~10-13 B/clk/SM on
// this part (outstanding requests x L2 latency; measured on the GEMM operand pipeline and matched by the 8.9 B/clk/SM
// the LDG version of this pass reached), so the rows of the trailing triangle are streamed into per-warp shared-memory
// slots with `cp.async.bulk` instead: one lane arms an mbarrier with the byte count and issues a <= 2 KB copy per
// (row, 512-column block); the warp then reads the slot with conflict-free 16-byte shared loads.  The work unit is
// one row of one column block; every warp of the group takes a contiguous, equally long run of units.  Inside a
// The variant needs few registers (no 4-row register tile), so THREE 8-warp CTAs fit on an SM for n <= ~2400: the
// pass is a dependent-latency chain per unit, and more resident warps is what shortens it.
constexpr int SYB_SLOTS = 2;                       // slots per warp (maximum): one row in flight behind the one in use
constexpr int SYB_SLOT_BYTES = SY_TW * 4;          // 2 KB

__host__ __device__ inline size_t sytrd_bulk_smem_bytes(int ldw, int slots = SYB_SLOTS, int warps = SY_WARPS) {
  size_t ring = (size_t)warps * slots * SYB_SLOT_BYTES;
  if ((size_t)ldw * 4 > ring) ring = (size_t)ldw * 4;   // the ring doubles as scratch for w (part 3)
  return (size_t)3 * ldw * 4 + ring;
}

template <int NT>
__device__ __forceinline__ void sytrd_bulk_body(const SytrdJob& job, float* sy_smem, double* red) {
  constexpr int NW = NT / 32;
  const int n = job.n;
  const int ldw = job.ldw;   // multiple of 4; also the stride of the shared vectors
  const int rank = blockIdx.x - job.cta0;
  const int G = job.ncta;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* W = job.W;
  float* va = sy_smem;              // previous Householder pair (pending rank-2 update)
  float* wa = sy_smem + ldw;
  float* vc = sy_smem + 2 * ldw;    // current Householder vector
  float* ring = sy_smem + 3 * ldw;  // row slots during the pass, w (wc) in part 3
  float* wc = ring;
  __shared__ __align__(8) uint64_t slot_bar[NW * SYB_SLOTS];
  if (tid < NW * SYB_SLOTS) mbar_init(smem_u32(&slot_bar[tid]), 1);
  if (tid == 0) fence_mbar_init();
  for (int j = n + tid; j < ldw; j += NT) {   // padding entries: only ever multiplied into padding columns
    va[j] = 0.f;
    wa[j] = 0.f;
    vc[j] = 0.f;
  }
  __syncthreads();
  bool have_prev = false;
  unsigned nbar = 0;
  uint32_t slot_uses = 0;           // completed uses of this warp's slots (same for all lanes): phase bookkeeping
  const int ncb = (n + SY_TW - 1) / SY_TW;
  const int nslots = job.slots;
  const uint32_t my_slots = smem_u32(ring) + (uint32_t)(warp * nslots * SYB_SLOT_BYTES);
  const uint32_t my_bars = smem_u32(&slot_bar[warp * SYB_SLOTS]);

  for (int k = 0; k + 1 < n; ++k) {
    // ---- part 1: column k of the updated trailing matrix (contiguous in colbuf) -> Householder vector
    const float* colk = (k == 0) ? W : job.colbuf + (size_t)(k & 1) * n;   // row 0 == column 0 of the input
    const float vk = have_prev ? va[k] : 0.f, wk = have_prev ? wa[k] : 0.f;
    double nrm2 = 0.0;
    strided_ldcg<NT>(colk, k, n, tid, [&](int j, float r) {
      if (have_prev) r -= vk * wa[j] + wk * va[j];
      vc[j] = r;
      if (j >= k + 2) nrm2 += (double)r * (double)r;
    });
    nrm2 = block_sum(nrm2, red);
    const float dk = vc[k];
    const float alpha = vc[k + 1];
    float beta, tau;
    if (nrm2 == 0.0) {
      beta = alpha;
      tau = 0.f;
    } else {
      const double nb = sqrt((double)alpha * (double)alpha + nrm2);
      beta = (float)(alpha >= 0.f ? -nb : nb);
      tau = (beta - alpha) / beta;
    }
    __syncthreads();
    const float scal = (tau != 0.f) ? 1.f / (alpha - beta) : 0.f;
    for (int j = k + 1 + tid; j < n; j += NT) vc[j] = (j == k + 1) ? 1.f : vc[j] * scal;
    __syncthreads();

    // ---- part 2: lower-triangle pass
    float* pb = job.pbuf + (size_t)(k % 3) * ldw;
    float* cnext = job.colbuf + (size_t)((k + 1) & 1) * n;
    const int j0 = k + 1;
    const int cb0 = j0 / SY_TW;
    int t_all = 0;
    for (int cb = cb0; cb < ncb; ++cb) t_all += n - max(j0, cb * SY_TW);
    if (job.sym == 2) t_all = 0;
    const int gw = rank * NW + warp;
    const int it_lo = (int)(((long long)gw * t_all) / (G * NW));
    const int it_hi = (int)(((long long)(gw + 1) * t_all) / (G * NW));
    // unit index -> (column block, row): one cursor walks the run for the copies, one for the arithmetic
    struct Cursor {
      int it, cb, base, rfirst, cnt;
    };
    auto seek = [&](Cursor& c) {
      while (c.it >= c.base + c.cnt) {
        c.base += c.cnt;
        ++c.cb;
        c.rfirst = max(j0, c.cb * SY_TW);
        c.cnt = n - c.rfirst;
      }
    };
    Cursor pf{it_lo, cb0, 0, j0, n - j0};
    Cursor cp = pf;
    auto issue = [&](int t) {   // lane 0: unit pf.it -> slot t % SYB_SLOTS
      seek(pf);
      const int i = pf.rfirst + (pf.it - pf.base);
      const uint32_t bytes = (uint32_t)min(SY_TW, ldw - pf.cb * SY_TW) * 4u;
      const uint32_t slot = (uint32_t)(t % nslots);
      const uint32_t bar = my_bars + slot * 8u;
      mbar_arrive_expect_tx(bar, bytes);
      bulk_load_1d(my_slots + slot * SYB_SLOT_BYTES, W + (size_t)i * ldw + pf.cb * SY_TW, bytes, bar);
      ++pf.it;
    };
    if (lane == 0) {
      fence_proxy_async_all();   // rows written by other CTAs (made visible by the group barrier) before our copies
      for (int t = 0; t < nslots && pf.it < it_hi; ++t) issue(t);
    }
    int cur_cb = -1, cbase = 0;
    bool clean = false;
    float cacc[16];
    auto flush_cols = [&]() {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = cbase + 128 * u;
        if (j < n && (cacc[4 * u] != 0.f || cacc[4 * u + 1] != 0.f || cacc[4 * u + 2] != 0.f || cacc[4 * u + 3] != 0.f))
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(pb + j), "f"(cacc[4 * u]),
                       "f"(cacc[4 * u + 1]), "f"(cacc[4 * u + 2]), "f"(cacc[4 * u + 3])
                       : "memory");
      }
    };
    for (int t = 0; cp.it < it_hi; ++t, ++cp.it) {
      seek(cp);
      const int i = cp.rfirst + (cp.it - cp.base);
      if (cp.cb != cur_cb) {
        if (cur_cb >= 0) flush_cols();
        cur_cb = cp.cb;
        cbase = cur_cb * SY_TW + 4 * lane;
        clean = have_prev && (cur_cb * SY_TW >= j0 + 1) && ((cur_cb + 1) * SY_TW <= n);
#pragma unroll
        for (int u = 0; u < 16; ++u) cacc[u] = 0.f;
      }
      const uint32_t slot = (uint32_t)(t % nslots);
      mbar_wait(my_bars + slot * 8u, (slot_uses >> slot) & 1u);
      slot_uses ^= 1u << slot;
      const uint32_t sbase = my_slots + slot * SYB_SLOT_BYTES + (uint32_t)(16 * lane);
      float4 a[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float4 x;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w)
                     : "r"(sbase + (uint32_t)(512 * u)));
        a[u] = x;
      }
      float* row = W + (size_t)i * ldw;
      const float vci = vc[i];
      float racc = 0.f;
      if (clean && i >= (cur_cb + 1) * SY_TW) {
        // interior row: strictly below the diagonal block, right of the first live column -> no masks
        const float nvi = -va[i], nwi = -wa[i];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int jv = cbase + 128 * u;
          const float4 w4u = *reinterpret_cast<const float4*>(wa + jv);
          const float4 v4u = *reinterpret_cast<const float4*>(va + jv);
          const float4 c4u = *reinterpret_cast<const float4*>(vc + jv);
          float4 x = a[u];
          x.x = fmaf(nwi, v4u.x, fmaf(nvi, w4u.x, x.x));
          x.y = fmaf(nwi, v4u.y, fmaf(nvi, w4u.y, x.y));
          x.z = fmaf(nwi, v4u.z, fmaf(nvi, w4u.z, x.z));
          x.w = fmaf(nwi, v4u.w, fmaf(nvi, w4u.w, x.w));
          *reinterpret_cast<float4*>(row + jv) = x;
          racc = fmaf(x.x, c4u.x, fmaf(x.y, c4u.y, fmaf(x.z, c4u.z, fmaf(x.w, c4u.w, racc))));
          cacc[4 * u] = fmaf(x.x, vci, cacc[4 * u]);
          cacc[4 * u + 1] = fmaf(x.y, vci, cacc[4 * u + 1]);
          cacc[4 * u + 2] = fmaf(x.z, vci, cacc[4 * u + 2]);
          cacc[4 * u + 3] = fmaf(x.w, vci, cacc[4 * u + 3]);
        }
      } else {
        // diagonal block, first live column block, ragged right edge, or the very first column step
        const float vi = have_prev ? va[i] : 0.f, wi = have_prev ? wa[i] : 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = cbase + 128 * u;
          if (!((j <= i) && (j + 3 >= j0) && (j < n))) continue;
          float x[4] = {a[u].x, a[u].y, a[u].z, a[u].w};
          const float4 c4u = *reinterpret_cast<const float4*>(vc + j);
          if (have_prev) {
            const float4 w4u = *reinterpret_cast<const float4*>(wa + j);
            const float4 v4u = *reinterpret_cast<const float4*>(va + j);
            x[0] -= vi * w4u.x + wi * v4u.x;
            x[1] -= vi * w4u.y + wi * v4u.y;
            x[2] -= vi * w4u.z + wi * v4u.z;
            x[3] -= vi * w4u.w + wi * v4u.w;
            *reinterpret_cast<float4*>(row + j) = make_float4(x[0], x[1], x[2], x[3]);
          }
          if (j >= j0 + 1 && j + 3 < i) {
            racc += x[0] * c4u.x + x[1] * c4u.y + x[2] * c4u.z + x[3] * c4u.w;
            cacc[4 * u] += x[0] * vci;
            cacc[4 * u + 1] += x[1] * vci;
            cacc[4 * u + 2] += x[2] * vci;
            cacc[4 * u + 3] += x[3] * vci;
          } else {
            const float cv[4] = {c4u.x, c4u.y, c4u.z, c4u.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const int jc = j + c;
              if (jc >= j0 && jc <= i) {
                racc += x[c] * cv[c];
                if (jc < i) cacc[4 * u + c] += x[c] * vci;
                if (jc == j0) cnext[i] = x[c];
              }
            }
          }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) racc += __shfl_xor_sync(0xffffffffu, racc, o);   // also: every lane has
      if (lane == 0) {                                  // consumed its slot data, so the slot may be refilled
        atomicAdd(pb + i, racc);
        if (pf.it < it_hi) issue(t + nslots);
      }
    }
    if (cur_cb >= 0) flush_cols();
    nbar += G;
    group_barrier(job.bar, nbar);

    // ---- part 3: w = tau p - (tau/2)(tau p . v) v (into the idle ring); clear the accumulation buffer of step k+2
    {
      float* pz = job.pbuf + (size_t)((k + 2) % 3) * ldw;
      for (int i = rank * NT + tid; i < ldw; i += G * NT) pz[i] = 0.f;
    }
    double dot = 0.0;
    strided_ldcg<NT>(pb, k + 1, n, tid, [&](int i, float pv) {
      const float pt = tau * pv;
      wc[i] = pt;
      dot += (double)pt * (double)vc[i];
    });
    dot = block_sum(dot, red);
    const float coef = 0.5f * tau * (float)dot;
    if (rank == 0) {
      float* vrow = W + (size_t)k * ldw;   // upper triangle is free in this variant: row k stores v_k for stage 3
      for (int j = k + 1 + tid; j < n; j += NT) vrow[j] = vc[j];
      if (tid == 0) {
        job.d[k] = dk;
        job.e[k] = beta;
        job.tau[k] = tau;
      }
    }
    // the pair becomes the pending update of the next step (copied, not pointer-swapped: w lives in the ring)
    for (int i = k + 1 + tid; i < n; i += NT) {
      const float v = vc[i];
      va[i] = v;
      wa[i] = wc[i] - coef * v;
    }
    __syncthreads();
    have_prev = true;
  }
  if (rank == 0 && tid == 0) {
    // the last reflector has tau = 0, so the (n-1, n-1) entry written to colbuf by the final pass is final
    job.d[n - 1] = (n == 1) ? __ldcg(W) : __ldcg(job.colbuf + (size_t)((n - 1) & 1) * n + (n - 1));
  }
}

// Three CTAs per SM (<= 80 registers): rounds whose matrices all take the bulk-fed symmetric pass or are small
constexpr int SYB_CTAS_PER_SM = 3;
__global__ void __launch_bounds__(SY_THREADS, SYB_CTAS_PER_SM) sytrd_kernel3(const SytrdJob* __restrict__ jobs,
                                                                             const int* __restrict__ cta_job) {
  extern __shared__ float sy_smem[];
  __shared__ double red[SY_WARPS];
  const SytrdJob job = jobs[cta_job[blockIdx.x]];
  if (job.sym)
    sytrd_bulk_body<SY_THREADS>(job, sy_smem, red);
  else
    sytrd_full_body(job, sy_smem, red);
}

__global__ void __launch_bounds__(SY_THREADS, SY_CTAS_PER_SM) sytrd_kernel(const SytrdJob* __restrict__ jobs,
                                                               const int* __restrict__ cta_job) {
  extern __shared__ float sy_smem[];
  __shared__ double red[SY_WARPS];
  const SytrdJob job = jobs[cta_job[blockIdx.x]];
  if (job.sym)
    sytrd_sym_body(job, sy_smem, red);
  else
    sytrd_full_body(job, sy_smem, red);
}

// ================================================================= stage 2: divide & conquer
struct DcMat {
  int n;
  float* Q[2];    // ping-pong eigenvector buffers (Q[0] is the caller's output)
  float* U;       // merge matrices (block diagonal)
  double* D[2];   // eigenvalues, per node sorted ascending
  const float* e; // off-diagonal of T
  double *z, *ds, *zs, *dkc, *zkc, *newvals, *zhat, *lam_mu, *invnorm, *rot_c, *rot_s, *rho;
  int *order, *keep, *defl, *rot_i, *rot_j, *lam_origin, *colrank, *counts;
};
struct DcNode {
  int mat, lo, mid, hi, depth;
  int blk0;   // first block of this node in the 1-D grids of the warp-per-root kernels (prefix of ceil(m / 8))
};

constexpr int DC_THREADS = 256;

// The warp-per-root kernels run on a 1-D grid with ceil(m / 8) blocks per node (8 warps per block): the nodes of one
// launch differ in size by two orders of magnitude (all matrices of a call share a tree level), and a 2-D grid sized
// for the largest node spent its time retiring empty blocks (62 000 per level for the ResNet-50 batch: ~0.4 ms per
// level and kernel whatever the level's real work).  A block finds its node by bisection over the prefix `blk0`.
struct DcWhere { int node, bx; };
__device__ __forceinline__ DcWhere dc_locate(const DcNode* __restrict__ nodes, int nn) {
  const int t = (int)blockIdx.x;
  int a = 0, b = nn - 1;   // last node with blk0 <= t
  while (a < b) {
    const int c = (a + b + 1) >> 1;
    if (nodes[c].blk0 <= t) a = c; else b = c - 1;
  }
  return DcWhere{a, t - nodes[a].blk0};
}

__global__ void dc_init_kernel(const DcMat* mats, int nmats) {
  // D = d - |e_left| - |e_right| (every split point subtracts |beta| from its two neighbours); Q buffers = I
  const DcMat M = mats[blockIdx.y];
  const int n = M.n;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    // d was written as float into the first half of D[1]'s storage by the caller? no: passed via z (double) buffer
    double v = M.z[i];
    if (i > 0) v -= fabs((double)M.e[i - 1]);
    if (i + 1 < n) v -= fabs((double)M.e[i]);
    M.D[0][i] = v;
    M.D[1][i] = v;
  }
}

__global__ void dc_load_d_kernel(const DcMat* mats, const float* const* dptr) {
  const DcMat M = mats[blockIdx.y];
  const float* d = dptr[blockIdx.y];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < M.n; i += gridDim.x * blockDim.x) M.z[i] = (double)d[i];
}

// merge step A: z, sort (children are sorted: rank by binary search), deflation scan, compact secular arrays
__global__ void __launch_bounds__(DC_THREADS) dc_prepare_kernel(const DcMat* mats, const DcNode* nodes) {
  __shared__ double red[DC_THREADS / 32];
  __shared__ DeflateCounts cnt_s;
  const DcNode nd = nodes[blockIdx.x];
  const DcMat M = mats[nd.mat];
  const int n = M.n, lo = nd.lo, mid = nd.mid, hi = nd.hi, m = hi - lo;
  const int cur = (nd.depth + 1) & 1;
  const float* Q = M.Q[cur];
  const double* D = M.D[cur];
  const float beta = M.e[mid - 1];
  const double sgn = beta >= 0.f ? 1.0 : -1.0;
  double rho = fabs((double)beta);
  const int tid = threadIdx.x;
  double nrm2 = 0.0;
  for (int i = tid; i < m; i += DC_THREADS) {
    const int g = lo + i;
    const double zi = (g < mid) ? (double)Q[(size_t)(mid - 1) * n + g] : sgn * (double)Q[(size_t)mid * n + g];
    M.z[g] = zi;
    nrm2 += zi * zi;
  }
  nrm2 = block_sum(nrm2, red);
  const double inv = nrm2 > 0.0 ? 1.0 / sqrt(nrm2) : 0.0;
  rho *= nrm2;
  // merge of two sorted runs: rank = own position + number of elements of the other run that precede
  for (int i = tid; i < m; i += DC_THREADS) {
    const int g = lo + i;
    const double di = D[g];
    int a, b, r;
    if (g < mid) {
      a = mid; b = hi;  // count elements of the right run strictly smaller
      while (a < b) { const int c = (a + b) >> 1; if (D[c] < di) a = c + 1; else b = c; }
      r = i + (a - mid);
    } else {
      a = lo; b = mid;  // count elements of the left run smaller or equal
      while (a < b) { const int c = (a + b) >> 1; if (D[c] <= di) a = c + 1; else b = c; }
      r = (g - mid) + (a - lo);
    }
    M.order[lo + r] = i;
    M.ds[lo + r] = di;
    M.zs[lo + r] = M.z[g] * inv;
  }
  __syncthreads();
  if (tid == 0) {
    cnt_s = deflate_scan(m, rho, M.ds + lo, M.zs + lo, M.keep + lo, M.defl + lo, M.rot_i + lo, M.rot_j + lo,
                         M.rot_c + lo, M.rot_s + lo);
    M.counts[3 * lo] = cnt_s.k;
    M.counts[3 * lo + 1] = cnt_s.ndefl;
    M.counts[3 * lo + 2] = cnt_s.nrot;
    M.rho[lo] = rho;
  }
  __syncthreads();
  const int k = cnt_s.k, ndefl = cnt_s.ndefl;
  for (int i = tid; i < k; i += DC_THREADS) {
    const int pos = M.keep[lo + i];
    M.dkc[lo + i] = M.ds[lo + pos];
    M.zkc[lo + i] = M.zs[lo + pos];
  }
  for (int c = tid; c < ndefl; c += DC_THREADS) M.newvals[lo + c] = M.ds[lo + M.defl[lo + c]];
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_prod_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v *= __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// merge step B: one warp per secular root
__global__ void __launch_bounds__(DC_THREADS) dc_secular_kernel(const DcMat* mats, const DcNode* nodes, int nn) {
  const DcWhere wh = dc_locate(nodes, nn);
  const DcNode nd = nodes[wh.node];
  const DcMat M = mats[nd.mat];
  const int lo = nd.lo;
  const int k = M.counts[3 * lo], ndefl = M.counts[3 * lo + 1];
  const int j = wh.bx * (DC_THREADS / 32) + (threadIdx.x >> 5);
  if (j >= k) return;
  const int lane = threadIdx.x & 31;
  const double* dk = M.dkc + lo;
  const double* zk = M.zkc + lo;
  const double rho = M.rho[lo];
  // z^2 on the fly: secular_partial wants an array; use zhat storage as scratch for z2 (filled redundantly per warp
  // would race) -> compute inline instead
  SecularState st;
  auto eval = [&](double dorigin, double sgn, double t, double& sum, double& dsum) {
    double s = 0.0, ds_ = 0.0;
    for (int i = lane; i < k; i += 32) {
      const double den = (dk[i] - dorigin) - sgn * t;
      const double q = zk[i] * zk[i] / den;
      s += q;
      ds_ += q / den;
    }
    sum = warp_sum_d(s);
    dsum = warp_sum_d(ds_);
  };
  double sum, dsum;
  if (j < k - 1) {
    const double half = 0.5 * (dk[j + 1] - dk[j]);
    eval(dk[j], 1.0, half, sum, dsum);
    const double gap = dk[j + 1] - dk[j];
    const double fmid = 1.0 + rho * sum;
    if (fmid > 0.0) {
      st.origin = j; st.sgn = 1.0; st.tlo = rho * zk[j] * zk[j] / (1.0 + 2.0 * rho / gap);
    } else {
      st.origin = j + 1; st.sgn = -1.0; st.tlo = zk[j + 1] * zk[j + 1] * half;
    }
    st.thi = half;
  } else {
    double z2 = 0.0;
    for (int i = lane; i < k; i += 32) z2 += zk[i] * zk[i];
    z2 = warp_sum_d(z2);
    st.origin = k - 1; st.sgn = 1.0; st.thi = rho * z2; st.tlo = rho * zk[k - 1] * zk[k - 1];
  }
  if (st.tlo > st.thi) st.tlo = st.thi;
  st.t = st.tlo; st.phase = 0; st.iters = 0; st.done = 0;
  while (!st.done) {
    eval(dk[st.origin], st.sgn, st.t, sum, dsum);
    secular_update(st, 1.0 + rho * sum, rho * dsum);
  }
  if (lane == 0) {
    M.lam_origin[lo + j] = st.origin;
    M.lam_mu[lo + j] = st.sgn * st.t;
    M.newvals[lo + ndefl + j] = dk[st.origin] + st.sgn * st.t;
  }
}

// merge step C: Loewner z-hat, one warp per kept entry
__global__ void __launch_bounds__(DC_THREADS) dc_zhat_kernel(const DcMat* mats, const DcNode* nodes, int nn) {
  const DcWhere wh = dc_locate(nodes, nn);
  const DcNode nd = nodes[wh.node];
  const DcMat M = mats[nd.mat];
  const int lo = nd.lo;
  const int k = M.counts[3 * lo];
  const int i = wh.bx * (DC_THREADS / 32) + (threadIdx.x >> 5);
  if (i >= k) return;
  const int lane = threadIdx.x & 31;
  const double* dk = M.dkc + lo;
  const double rho = M.rho[lo];
  const double di = dk[i];
  double prod = 1.0;
  for (int j = lane; j < k; j += 32) {
    const double diff = (dk[M.lam_origin[lo + j]] - di) + M.lam_mu[lo + j];  // lambda_j - d_i
    prod *= (j == i) ? diff / rho : diff / (dk[j] - di);
  }
  prod = warp_prod_d(prod);
  if (lane == 0) M.zhat[lo + i] = copysign(sqrt(fabs(prod)), M.zkc[lo + i]);
}

// merge step D1: column norms of the secular eigenvectors (warp per root) and rank of every new eigenvalue
__global__ void __launch_bounds__(DC_THREADS) dc_norm_rank_kernel(const DcMat* mats, const DcNode* nodes, int nn) {
  const DcWhere wh = dc_locate(nodes, nn);
  const DcNode nd = nodes[wh.node];
  const DcMat M = mats[nd.mat];
  const int lo = nd.lo, m = nd.hi - nd.lo;
  const int k = M.counts[3 * lo];
  const int lane = threadIdx.x & 31;
  const int j = wh.bx * (DC_THREADS / 32) + (threadIdx.x >> 5);
  if (j < k) {
    const double* dk = M.dkc + lo;
    const double dorg = dk[M.lam_origin[lo + j]], mu = M.lam_mu[lo + j];
    double s = 0.0;
    for (int i = lane; i < k; i += 32) {
      const double v = M.zhat[lo + i] / ((dk[i] - dorg) - mu);
      s += v * v;
    }
    s = warp_sum_d(s);
    if (lane == 0) M.invnorm[lo + j] = 1.0 / sqrt(s);
  }
  // ranks: thread per element (grid.x * 8 warps * 32 lanes >= m is guaranteed by the launcher); the values pass
  // through shared memory in chunks of one per thread, so the m^2 comparisons read on-chip broadcasts
  if (wh.bx * DC_THREADS >= m) return;   // (uniform per block)
  __shared__ double nv_s[DC_THREADS];
  const int c = wh.bx * DC_THREADS + threadIdx.x;
  const double v = c < m ? M.newvals[lo + c] : 0.0;
  int r = 0;
  for (int t0 = 0; t0 < m; t0 += DC_THREADS) {
    __syncthreads();
    if (t0 + (int)threadIdx.x < m) nv_s[threadIdx.x] = M.newvals[lo + t0 + threadIdx.x];
    __syncthreads();
    const int lim = min(DC_THREADS, m - t0);
#pragma unroll 8
    for (int tt = 0; tt < lim; ++tt) {
      const double u = nv_s[tt];
      r += (u < v || (u == v && t0 + tt < c)) ? 1 : 0;
    }
  }
  if (c < m) {
    M.colrank[lo + c] = r;
    M.D[nd.depth & 1][lo + r] = v;
  }
}

// merge step D2: U block = 0, unit columns of deflated entries, scaled secular eigenvectors
__global__ void __launch_bounds__(DC_THREADS) dc_fill_u_kernel(const DcMat* mats, const DcNode* nodes) {
  const DcNode nd = nodes[blockIdx.z];
  const DcMat M = mats[nd.mat];
  const int n = M.n, lo = nd.lo, m = nd.hi - nd.lo;
  const int k = M.counts[3 * lo], ndefl = M.counts[3 * lo + 1];
  // element (sorted position sp, new column c) ; threads along c
  const int c = blockIdx.x * DC_THREADS + threadIdx.x;
  if (c >= m) return;
  const int col = M.colrank[lo + c];
  const int rows_per = (m + gridDim.y - 1) / gridDim.y;
  const int sp0 = blockIdx.y * rows_per;
  const int sp1 = min(m, sp0 + rows_per);
  if (c < ndefl) {
    const int mine = M.defl[lo + c];
    for (int sp = sp0; sp < sp1; ++sp)
      M.U[(size_t)(lo + M.order[lo + sp]) * n + lo + col] = (sp == mine) ? 1.f : 0.f;
  } else {
    const int j = c - ndefl;
    const double* dk = M.dkc + lo;
    const double dorg = dk[M.lam_origin[lo + j]], mu = M.lam_mu[lo + j], inv = M.invnorm[lo + j];
    // rows: kept entries carry values, deflated rows are zero. Walk sorted positions; kept list is ascending.
    int ki = 0;
    {  // first kept index with position >= sp0 (binary search)
      int a = 0, b = k;
      while (a < b) { const int t = (a + b) >> 1; if (M.keep[lo + t] < sp0) a = t + 1; else b = t; }
      ki = a;
    }
    for (int sp = sp0; sp < sp1; ++sp) {
      float val = 0.f;
      if (ki < k && M.keep[lo + ki] == sp) {
        val = (float)(M.zhat[lo + ki] / ((dk[ki] - dorg) - mu) * inv);
        ++ki;
      }
      M.U[(size_t)(lo + M.order[lo + sp]) * n + lo + col] = val;
    }
  }
}

// merge step D3: undo the deflation rotations on the rows of U (reverse order), chain value carried in fp64
__global__ void __launch_bounds__(DC_THREADS) dc_rotate_kernel(const DcMat* mats, const DcNode* nodes) {
  const DcNode nd = nodes[blockIdx.y];
  const DcMat M = mats[nd.mat];
  const int n = M.n, lo = nd.lo, m = nd.hi - nd.lo;
  const int nrot = M.counts[3 * lo + 2];
  const int c = blockIdx.x * DC_THREADS + threadIdx.x;
  if (c >= m || nrot == 0) return;
  float* Ucol = M.U + (size_t)lo * n + lo + c;
  int carry_row = -1;
  double carry = 0.0;
  for (int t = nrot - 1; t >= 0; --t) {
    const int si = M.rot_i[lo + t], sj = M.rot_j[lo + t];
    const double cc = M.rot_c[lo + t], ss = M.rot_s[lo + t];
    const int ri = M.order[lo + si], rj = M.order[lo + sj];
    if (carry_row >= 0 && carry_row != sj) {
      Ucol[(size_t)M.order[lo + carry_row] * n] = (float)carry;
      carry_row = -1;
    }
    const double b = (carry_row == sj) ? carry : (double)Ucol[(size_t)rj * n];
    const double a = (double)Ucol[(size_t)ri * n];
    Ucol[(size_t)rj * n] = (float)(ss * a + cc * b);
    carry = cc * a - ss * b;
    carry_row = si;
  }
  if (carry_row >= 0) Ucol[(size_t)M.order[lo + carry_row] * n] = (float)carry;
}

// merge step E (small nodes): Qnew_blk = Qcur_blk * U_blk with U staged in shared memory
__global__ void __launch_bounds__(DC_THREADS) dc_gemm_small_kernel(const DcMat* mats, const DcNode* nodes,
                                                                   int max_small) {
  extern __shared__ float us[];
  const DcNode nd = nodes[blockIdx.x];
  const DcMat M = mats[nd.mat];
  const int n = M.n, lo = nd.lo, m = nd.hi - nd.lo;
  if (m > max_small) return;
  const float* Qc = M.Q[(nd.depth + 1) & 1] + (size_t)lo * n + lo;
  float* Qn = M.Q[nd.depth & 1] + (size_t)lo * n + lo;
  const float* U = M.U + (size_t)lo * n + lo;
  for (int e = threadIdx.x; e < m * m; e += DC_THREADS) us[e] = U[(size_t)(e / m) * n + (e % m)];
  __syncthreads();
  for (int e = threadIdx.x; e < m * m; e += DC_THREADS) {
    const int r = e / m, c = e - r * m;
    const float* qr = Qc + (size_t)r * n;
    float acc = 0.f;
    for (int t = 0; t < m; ++t) acc += qr[t] * us[t * m + c];
    Qn[(size_t)r * n + c] = acc;
  }
}

// eigenvalues out (float, optionally clamped at 0: eigen.py:320,343)
__global__ void dc_store_d_kernel(const DcMat* mats, float* const* dout, int clamp_nonneg) {
  const DcMat M = mats[blockIdx.y];
  float* d = dout[blockIdx.y];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < M.n; i += gridDim.x * blockDim.x) {
    float v = (float)M.D[0][i];
    if (clamp_nonneg && v < 0.f) v = 0.f;
    d[i] = v;
  }
}

// ================================================================= stage 3: back-transformation helpers
// zero the diagonal and lower triangle of W so that rows of W are the Householder vectors with explicit zeros
__global__ void zero_lower_kernel(float* W, int n, int ldw) {
  const long long total = (long long)n * ldw;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int r = (int)(e / ldw), c = (int)(e - (long long)r * ldw);
    if (c <= r || c >= n) W[e] = 0.f;
  }
}

// One compact-WY block of one matrix: rows k0 .. k0+nb-1 of W hold v_k in columns k0+1 .. n-1.
struct WyBlock {
  const float* Vb;   // W + k0*n + (k0+1)
  int n, nb, kk;     // row stride, reflectors in the block, non-zero columns (n - k0 - 1)
  double* S;         // 128 x 128 Gram matrix (fp64)
  const float* tau;  // tau + k0
  float* T;          // 128 x 128 triangular factor (fp32, zero padded)
};

// Gram matrices of ALL blocks of all matrices in one launch: S[i,j] = v_i . v_j in fp64, j >= i.
// 128 x 128 x kk flops per block -- little work, but its accuracy decides how orthogonal I - V T V^T is.
// One CTA per 64 x 64 tile of S (blockIdx.x: 0 = rows 0..63 x columns 0..63, 1 = 0..63 x 64..127, 2 = 64..127 x
// 64..127): the two row panels go through shared memory in chunks of 32 columns, thread (ty, tx) owns rows ty + 16 u
// and columns tx + 16 w (conflict-free with the odd pitch) and keeps its 4 x 4 sums in fp64 registers.  (The first
// version read both rows of every pair from L2: 8256 x 2 x kk floats per block, 3.7 ms per ResNet-50 step.)
__global__ void __launch_bounds__(256) wy_gram_kernel(const WyBlock* __restrict__ blocks) {
  __shared__ float Vi[64][33];
  __shared__ float Vj[64][33];
  const WyBlock b = blocks[blockIdx.y];
  const int I = blockIdx.x == 2 ? 64 : 0, J = blockIdx.x == 0 ? 0 : 64;
  if (J >= b.nb) return;   // (I <= J)
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int lr = tid >> 5, lc = tid & 31;
  double acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int w = 0; w < 4; ++w) acc[u][w] = 0.0;
  for (int c0 = 0; c0 < b.kk; c0 += 32) {
    const int c = c0 + lc;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int r = lr + 8 * p;
      Vi[r][lc] = (I + r < b.nb && c < b.kk) ? __ldg(b.Vb + (size_t)(I + r) * b.n + c) : 0.f;
      Vj[r][lc] = (J + r < b.nb && c < b.kk) ? __ldg(b.Vb + (size_t)(J + r) * b.n + c) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      double x[4], y[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) x[u] = (double)Vi[ty + 16 * u][k];
#pragma unroll
      for (int w = 0; w < 4; ++w) y[w] = (double)Vj[tx + 16 * w][k];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int w = 0; w < 4; ++w) acc[u][w] = fma(x[u], y[w], acc[u][w]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int i = I + ty + 16 * u, j = J + tx + 16 * w;
      if (i < b.nb && j < b.nb && j >= i) b.S[i * 128 + j] = acc[u][w];
    }
}

// T factors of ALL blocks in one launch (forward, columnwise, fp64):
//   T[i,i] = tau_i, T[0:i,i] = -tau_i T[0:i,0:i] S[0:i,i];  written out as fp32 for the GEMMs.
// One CTA per block; column i is a triangular mat-vec done by 32 warps (rows round-robin, lanes over the sum).
constexpr int LARFT_THREADS = 1024;
__global__ void __launch_bounds__(LARFT_THREADS) larft_kernel(const WyBlock* __restrict__ blocks) {
  extern __shared__ double Ts[];  // 128 x 129 doubles (132 KB, opt-in)
  __shared__ double col[128];
  const WyBlock b = blocks[blockIdx.x];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < 128 * 129; e += LARFT_THREADS) Ts[e] = 0.0;
  __syncthreads();
  for (int i = 0; i < b.nb; ++i) {
    const double ti = (double)b.tau[i];
    if (tid < i) col[tid] = b.S[tid * 128 + i];
    __syncthreads();
    for (int r = warp; r < i; r += LARFT_THREADS / 32) {
      double acc = 0.0;
      for (int t = r + lane; t < i; t += 32) acc += Ts[r * 129 + t] * col[t];  // T upper triangular
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) Ts[r * 129 + i] = -ti * acc;
    }
    if (tid == 0) Ts[i * 129 + i] = ti;
    __syncthreads();
  }
  for (int e = tid; e < 128 * 128; e += LARFT_THREADS) {
    const int r = e >> 7, c = e & 127;
    b.T[e] = (r < b.nb && c < b.nb) ? (float)Ts[r * 129 + c] : 0.f;
  }
}

// All per-matrix initialisation of a batched call in one launch (it used to be two memsets, a 2-D copy and two
// identity launches per matrix): working copy W <- F with zero padding, both eigenvector buffers <- I, accumulation
// buffers <- 0.  blockIdx.y walks the matrices.
struct EighInit {
  const float* F;
  float* W;
  float* Q0;
  float* Q1;
  float* pbuf;
  int n, ldw;
};
__global__ void __launch_bounds__(256) eigh_init_kernel(const EighInit* __restrict__ items) {
  const EighInit it = items[blockIdx.y];
  const int n = it.n, ldw = it.ldw;
  // rows dealt round-robin to the CTAs of the matrix (no index arithmetic per element); 16-byte accesses when the
  // rows of F and of the eigenvector buffers are aligned (n a multiple of 4 and aligned bases)
  const bool vec = (n & 3) == 0 &&
      ((reinterpret_cast<uintptr_t>(it.F) | reinterpret_cast<uintptr_t>(it.W) | reinterpret_cast<uintptr_t>(it.Q0) |
        reinterpret_cast<uintptr_t>(it.Q1)) & 15) == 0;
  for (int r = blockIdx.x; r < n; r += gridDim.x) {
    const float* f = it.F + (size_t)r * n;
    float* w = it.W + (size_t)r * ldw;
    float* q0 = it.Q0 + (size_t)r * n;
    float* q1 = it.Q1 + (size_t)r * n;
    if (vec) {   // ldw == n
      for (int c = 4 * threadIdx.x; c < n; c += 4 * 256) {
        *reinterpret_cast<float4*>(w + c) = __ldg(reinterpret_cast<const float4*>(f + c));
        const float4 v = make_float4(c == r ? 1.f : 0.f, c + 1 == r ? 1.f : 0.f, c + 2 == r ? 1.f : 0.f,
                                     c + 3 == r ? 1.f : 0.f);
        *reinterpret_cast<float4*>(q0 + c) = v;
        *reinterpret_cast<float4*>(q1 + c) = v;
      }
    } else {
      for (int c = threadIdx.x; c < ldw; c += 256) w[c] = c < n ? __ldg(f + c) : 0.f;
      for (int c = threadIdx.x; c < n; c += 256) {
        const float v = c == r ? 1.f : 0.f;
        q0[c] = v;
        q1[c] = v;
      }
    }
  }
  if (blockIdx.x == 0)
    for (int e = threadIdx.x; e < 3 * ldw; e += 256) it.pbuf[e] = 0.f;
}
// zero the diagonal and lower triangle (and the padding columns) of every working copy: the rows then hold the
// Householder vectors with explicit zeros
__global__ void __launch_bounds__(256) zero_lower_batched_kernel(const EighInit* __restrict__ items) {
  const EighInit it = items[blockIdx.y];
  const int n = it.n, ldw = it.ldw;
  for (int r = blockIdx.x; r < n; r += gridDim.x) {
    float* w = it.W + (size_t)r * ldw;
    for (int c = threadIdx.x; c <= r; c += 256) w[c] = 0.f;
    for (int c = n + threadIdx.x; c < ldw; c += 256) w[c] = 0.f;
  }
}

// ================================================================= host orchestration
struct EighLayout {
  size_t total = 0;
  size_t take(size_t bytes) {
    const size_t off = total;
    total += (bytes + 255) / 256 * 256;
    return off;
  }
};

}  // namespace kfac
#include "eigh2.cuh"
namespace kfac {

// Process-wide choice of the tridiagonalisation path (kfac_eigh_set_mode): 0 = by size and batch, 1 = one-stage
// kernels only, 2 = two-stage wherever it applies (n >= 96), m >= 96 = two-stage for n >= m.
// By size and batch (measured on B200, profiles/README.md): in a call with several large matrices the two-stage path
// wins from n = 1000 (its kernels of different matrices overlap; ResNet-50's 108 factors: 287 -> 232 ms per step); a
// matrix that is (almost) alone only gains from about n = 4096 (n = 1024: 10.4 ms one-stage, 13.3 ms two-stage;
// n = 4608: 83 / 83; n = 8193: 341 / 264).
static int g_eigh_mode = 0;
constexpr int TS_BATCH_MIN = 4;        // this many matrices of TS_MIN_N or more make a "batch"
constexpr int TS_MIN_N_ALONE = 4096;
static int two_stage_threshold(const int* ns, int count) {
  if (g_eigh_mode == 1) return 1 << 30;
  if (g_eigh_mode == 2) return 96;
  if (g_eigh_mode >= 96) return g_eigh_mode;
  int big = 0;
  for (int i = 0; i < count; ++i) big += ns[i] >= TS_MIN_N ? 1 : 0;
  return big >= TS_BATCH_MIN ? TS_MIN_N : TS_MIN_N_ALONE;
}

static int sytrd_group_size(int n) {   // nominal CTAs (8 warps each) per matrix
  int g = (n + 47) / 48;
  if (g < 1) g = 1;
  return g;
}

struct EighOffsets {
  std::vector<size_t> W, Qb, U, dvec, D0, D1, evec, tauvec, pbuf, dbl, ints;  // per matrix
  std::vector<size_t> TV, C2m;  // per matrix: T*V rows (n x n) and the 128 x n product
  std::vector<TsSlots> ts;      // per matrix: two-stage buffers (unused entries stay zero)
  std::vector<int> ts_index;    // batch positions of the two-stage matrices
  TsShared tsh;
  int ts_min;                   // matrices of this size or more take the two-stage path in this call
  size_t bars, mats, jobs, cta_job, nodes, dptr_in, dptr_out, Sall, Tall, wyblocks, gtable, partials, init;
  size_t partials_bytes, gtable_bytes;
  int total_blocks;
  size_t total;
  int max_n, total_nodes;
};

constexpr int DC_NDBL = 11;  // z ds zs dkc zkc newvals zhat lam_mu invnorm rot_c rot_s  (+ rho shares slot sizes)
constexpr int DC_NINT = 8;   // order keep defl rot_i rot_j lam_origin colrank counts(3n)

static EighOffsets eigh_layout(const int* ns, int count) {
  EighOffsets o;
  EighLayout L;
  o.max_n = 0;
  long long nodes = 0;
  for (int i = 0; i < count; ++i) {
    const size_t n = (size_t)ns[i];
    o.max_n = std::max(o.max_n, ns[i]);
    nodes += ns[i];
    o.W.push_back(L.take(n * ((n + 3) / 4 * 4) * 4));
    o.Qb.push_back(L.take(n * n * 4));
    o.U.push_back(L.take(n * n * 4));
    o.dvec.push_back(L.take(n * 4));
    o.evec.push_back(L.take(n * 4));
    o.tauvec.push_back(L.take(n * 4));
    o.pbuf.push_back(L.take((5 * n + 16) * 4));  // 3 (padded) accumulation buffers + 2n column hand-over
    o.D0.push_back(L.take(n * 8));
    o.D1.push_back(L.take(n * 8));
    o.dbl.push_back(L.take((size_t)(DC_NDBL + 1) * n * 8));
    o.ints.push_back(L.take((size_t)(DC_NINT + 2) * n * 4));
  }
  o.total_nodes = (int)nodes;
  o.bars = L.take((size_t)count * 128);
  o.mats = L.take((size_t)count * sizeof(DcMat));
  o.jobs = L.take((size_t)count * sizeof(SytrdJob));
  o.cta_job = L.take((size_t)KFAC_MAX_SMS * 3 * 4 * (size_t)std::max(count, 1));
  o.nodes = L.take((size_t)std::max<long long>(nodes, 1) * sizeof(DcNode));
  o.dptr_in = L.take((size_t)count * 8);
  o.dptr_out = L.take((size_t)count * 8);
  o.init = L.take((size_t)count * sizeof(EighInit));
  o.total_blocks = 0;
  for (int i = 0; i < count; ++i) {
    const size_t n = (size_t)ns[i];
    o.total_blocks += ns[i] >= 2 ? (ns[i] - 1 + 127) / 128 : 0;
    o.TV.push_back(L.take(n * n * 4));
    o.C2m.push_back(L.take(128 * n * 4));
  }
  o.ts.resize(count);
  o.ts_min = two_stage_threshold(ns, count);
  {
    const int ts_min = o.ts_min;
    std::vector<int> ts_ns;
    for (int i = 0; i < count; ++i) {
      memset(&o.ts[i], 0, sizeof(TsSlots));
      if (ns[i] >= ts_min) {
        ts_layout_matrix(L, ns[i], o.ts[i]);
        o.ts_index.push_back(i);
        ts_ns.push_back(ns[i]);
      }
    }
    memset(&o.tsh, 0, sizeof(o.tsh));
    if (!ts_ns.empty()) ts_layout_shared(L, ts_ns, o.tsh);
  }
  o.Sall = L.take((size_t)std::max(o.total_blocks, 1) * 128 * 128 * 8);
  o.Tall = L.take((size_t)std::max(o.total_blocks, 1) * 128 * 128 * 4);
  o.wyblocks = L.take((size_t)std::max(o.total_blocks, 1) * sizeof(WyBlock));
  o.gtable_bytes = (size_t)(std::max<long long>(nodes, o.total_blocks) + 64) * (sizeof(GemmParams) + 8) + 4096;
  o.gtable = L.take(o.gtable_bytes);
  o.partials_bytes = (size_t)96 << 20;
  o.partials = L.take(o.partials_bytes);
  o.total = L.total + 256;
  return o;
}

static size_t eigh_workspace_bytes(const int* ns, int count) {
  if (ns == nullptr || count <= 0) return 0;
  return eigh_layout(ns, count).total;
}

static void build_nodes(int mat, int lo, int hi, int depth, std::vector<std::vector<DcNode>>& levels) {
  if (hi - lo <= 1) return;
  const int mid = lo + (hi - lo) / 2;
  if ((int)levels.size() <= depth) levels.resize(depth + 1);
  levels[depth].push_back(DcNode{mat, lo, mid, hi, depth, 0});
  build_nodes(mat, lo, mid, depth + 1, levels);
  build_nodes(mat, mid, hi, depth + 1, levels);
}

constexpr int DC_SMALL = 96;  // nodes up to this size use the shared-memory SIMT product
constexpr int DC_SLICE = 65535;  // nodes per launch slice (grid.y / grid.z limit of the kernels that still use them)

static int eigh_batched(const float* const* F, const int* ns, int count, float* const* Qout, float* const* dout,
                        int clamp_nonneg, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (count <= 0) return 0;
  if (F == nullptr || ns == nullptr || Qout == nullptr || dout == nullptr) return fail("kfac_eigh_batched: null pointer");
  for (int i = 0; i < count; ++i)
    if (ns[i] <= 0 || F[i] == nullptr || Qout[i] == nullptr || dout[i] == nullptr)
      return fail("kfac_eigh_batched: bad entry %d", i);
  const EighOffsets o = eigh_layout(ns, count);
  if (ws == nullptr || ws_bytes < o.total)
    return fail("kfac_eigh_batched: workspace too small (%zu given, %zu needed)", ws_bytes, o.total);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  if (ensure_gemm_attr()) return 1;

  // ---- descriptors
  std::vector<DcMat> mats(count);
  std::vector<SytrdJob> jobs(count);
  std::vector<const float*> dptr_in(count);
  std::vector<EighInit> inits(count);
  for (int i = 0; i < count; ++i) {
    const size_t n = (size_t)ns[i];
    DcMat& M = mats[i];
    M.n = ns[i];
    M.Q[0] = Qout[i];
    M.Q[1] = reinterpret_cast<float*>(base + o.Qb[i]);
    M.U = reinterpret_cast<float*>(base + o.U[i]);
    M.D[0] = reinterpret_cast<double*>(base + o.D0[i]);
    M.D[1] = reinterpret_cast<double*>(base + o.D1[i]);
    M.e = reinterpret_cast<float*>(base + o.evec[i]);
    double* db = reinterpret_cast<double*>(base + o.dbl[i]);
    M.z = db; M.ds = db + n; M.zs = db + 2 * n; M.dkc = db + 3 * n; M.zkc = db + 4 * n; M.newvals = db + 5 * n;
    M.zhat = db + 6 * n; M.lam_mu = db + 7 * n; M.invnorm = db + 8 * n; M.rot_c = db + 9 * n; M.rot_s = db + 10 * n;
    M.rho = db + 11 * n;
    int* ib = reinterpret_cast<int*>(base + o.ints[i]);
    M.order = ib; M.keep = ib + n; M.defl = ib + 2 * n; M.rot_i = ib + 3 * n; M.rot_j = ib + 4 * n;
    M.lam_origin = ib + 5 * n; M.colrank = ib + 6 * n; M.counts = ib + 7 * n;  // 3n ints
    SytrdJob& J = jobs[i];
    J.W = reinterpret_cast<float*>(base + o.W[i]);
    J.n = ns[i];
    J.ldw = (ns[i] + 3) / 4 * 4;
    J.d = reinterpret_cast<float*>(base + o.dvec[i]);
    J.e = reinterpret_cast<float*>(base + o.evec[i]);
    J.tau = reinterpret_cast<float*>(base + o.tauvec[i]);
    J.pbuf = reinterpret_cast<float*>(base + o.pbuf[i]);
    J.colbuf = J.pbuf + 3 * (size_t)J.ldw;
    J.sym = ns[i] >= 1024 ? 1 : 0;  // below that the fixed per-column cost dominates
#ifdef KFAC_DEBUG_HOOKS   // timing experiments only (profiles/README.md): 2 skips the pass, i.e. gives WRONG results
    if (J.sym && getenv("KFAC_DEBUG_SKIP_PASS")) J.sym = atoi(getenv("KFAC_DEBUG_SKIP_PASS"));
#endif
    J.bar = reinterpret_cast<unsigned*>(base + o.bars + (size_t)i * 128);
    dptr_in[i] = J.d;
    inits[i] = EighInit{F[i], J.W, M.Q[0], M.Q[1], J.pbuf, ns[i], J.ldw};
  }
  KFAC_CUDA(cudaMemsetAsync(base + o.bars, 0, (size_t)count * 128, stream));
  const EighInit* dinits = reinterpret_cast<const EighInit*>(base + o.init);
  KFAC_CUDA(cudaMemcpyAsync(base + o.init, inits.data(), (size_t)count * sizeof(EighInit), cudaMemcpyHostToDevice, stream));
  {
    // up to 512 CTAs per matrix, rows round-robin: the big matrices get the parallelism, a CTA of a small one that
    // has no row returns at once
    const int bx = std::max(1, std::min(o.max_n, 512));
    eigh_init_kernel<<<dim3(bx, count), 256, 0, stream>>>(dinits); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
  }

  // ---- stage 1: rounds of cooperative launches; largest matrices first, each with its CTA group
  if (stage_mark(0, stream)) return 1;
  // two-stage matrices: band reduction on the tensor cores + bulge chasing (eigh2.cuh)
  std::vector<TsEntry> ts_ents;
  for (int i : o.ts_index) {
    TsEntry E;
    E.index = i;
    E.n = ns[i];
    E.ldw = jobs[i].ldw;
    E.slots = o.ts[i];
    memset(&E.dev, 0, sizeof(E.dev));
    E.dev.W = jobs[i].W;
    E.dev.tau1 = jobs[i].tau;
    E.dev.d = jobs[i].d;
    E.dev.e = jobs[i].e;
    E.dev.V2 = reinterpret_cast<float*>(base + o.TV[i]);   // free until the compact-WY stage reuses it as T V
    E.dev.Z = Qout[i];
    ts_ents.push_back(E);
  }
  if (!ts_ents.empty()) {
    if (ts_setup(ts_ents, base, o.tsh, stream)) return 1;
    if (ts_tridiagonalise(ts_ents, base, o.tsh, stream)) return 1;
  }
  std::vector<int> idx;
  for (int i = 0; i < count; ++i)
    if (ns[i] < o.ts_min) idx.push_back(i);
  std::sort(idx.begin(), idx.end(), [&](int a, int b) { return ns[a] > ns[b]; });
  std::vector<int> cta_job_all;
  struct Round { int off, ncta; size_t smem; int kind; };   // kind: 0 two-CTA LDG kernel, 1 three-CTA bulk-fed kernel
  std::vector<Round> rounds;
  EighPaths paths;
  paths.two_stage = (int)ts_ents.size();
  if (!idx.empty() && (size_t)4 * ((ns[idx[0]] + 3) / 4 * 4) * sizeof(float) > 190 * 1024)
    return fail("kfac_eigh_batched: n = %d exceeds the supported maximum (12160)", ns[idx[0]]);
  {
    static int sy_smem_set[64] = {0};
    int dev = 0;
    KFAC_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !sy_smem_set[dev]) {
      KFAC_CUDA(cudaFuncSetAttribute(sytrd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 190 * 1024));
      KFAC_CUDA(cudaFuncSetAttribute(sytrd_kernel3, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      if (dev < 64) sy_smem_set[dev] = 1;
    }
#ifdef KFAC_DEBUG_HOOKS
    static const char* no_k3 = getenv("KFAC_DEBUG_NO_K3");   // A/B timing of the three-CTA kernel
#else
    const char* no_k3 = nullptr;
#endif
    size_t pos = 0;
    while (pos < idx.size()) {
      // The round's shared-memory footprint follows its largest matrix (the first one); that decides which kernel
      // runs the round and how many co-resident CTAs the cooperative launch may have.  The pass is a dependent-
      // latency chain per work unit, so the configuration with the most resident warps per SM wins:
      //   kind 1: three 8-warp CTAs per SM, bulk-copy fed, when the vectors are small enough (n <= ~2400): rounds
      //           2..5 of the ResNet-50 batch went from 102 to 86 ms
      //   kind 0: two 8-warp CTAs per SM, LDG fed with a 4-row register tile (everything else).  For n = 4608 a
      //           one-slot bulk variant at three CTAs (135 ms) and a 24-warp bulk CTA per SM (130 ms) both lost
      //           to this kernel (104 ms for the 3 x 4608 round)
      const int first = idx[pos];
      const int ldw0 = (ns[first] + 3) / 4 * 4;
      size_t smem = (size_t)4 * ldw0 * sizeof(float);
      int kind = 0, slots = 0, unit = 1;   // unit: 8-warp CTAs one CTA of this kind is worth
      {
        const size_t smem3 = jobs[first].sym ? sytrd_bulk_smem_bytes(ldw0) : smem;
        int occ3 = 0;
        if (!no_k3 && smem3 <= 100 * 1024)
          KFAC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ3, sytrd_kernel3, SY_THREADS, smem3));
        if (occ3 >= SYB_CTAS_PER_SM) {
          kind = 1;
          smem = smem3;
          slots = sm_count() * SYB_CTAS_PER_SM;
        }
      }
      if (kind == 0) {
        int occ = 0;
        KFAC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sytrd_kernel, SY_THREADS, smem));
        if (occ < 1) return fail("kfac_eigh_batched: tridiagonalisation kernel does not fit on an SM (n = %d)", ns[first]);
        slots = sm_count() * std::min(occ, SY_CTAS_PER_SM);
      }
      auto nominal = [&](int i) { return std::min((sytrd_group_size(ns[i]) + unit - 1) / unit, slots); };
      // greedy fill with the nominal group sizes ...
      std::vector<int> members;
      int used = 0;
      while (pos < idx.size()) {
        const int i = idx[pos];
        const int g = nominal(i);
        if (used + g > slots && !members.empty()) break;
        members.push_back(i);
        used += g;
        ++pos;
      }
      // ... then grow the groups proportionally so that a round with few matrices still uses every SM
      const int off = (int)cta_job_all.size();
      int total = 0;
      for (int i : members) {
        const int g = nominal(i);
        int grown = (int)((long long)g * slots / used);
        const int cap = std::max(1, ns[i] / (16 * unit));   // keep >= 2 rows per warp at the first column
        if (grown > cap) grown = cap;
        if (grown < g) grown = g;
        if (total + grown > slots) grown = slots - total;
        jobs[i].cta0 = total;
        jobs[i].ncta = grown;
        jobs[i].slots = SYB_SLOTS;
        for (int c = 0; c < grown; ++c) cta_job_all.push_back(i);
        total += grown;
        if (!jobs[i].sym) ++paths.full;
        else if (kind == 1) ++paths.sym_bulk;
        else ++paths.sym_ldg;
      }
      rounds.push_back({off, total, smem, kind});
    }
  }
  if (cta_job_all.size() * 4 > (size_t)KFAC_MAX_SMS * 3 * 4 * (size_t)count) return fail("kfac_eigh_batched: internal cta_job overflow");
  KFAC_CUDA(cudaMemcpyAsync(base + o.jobs, jobs.data(), count * sizeof(SytrdJob), cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(base + o.cta_job, cta_job_all.data(), cta_job_all.size() * 4, cudaMemcpyHostToDevice,
                            stream));
  KFAC_CUDA(cudaMemcpyAsync(base + o.mats, mats.data(), count * sizeof(DcMat), cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(base + o.dptr_in, dptr_in.data(), count * 8, cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(base + o.dptr_out, dout, count * 8, cudaMemcpyHostToDevice, stream));
  // host vectors above are pageable: the async copies are staged before returning, so they may go out of scope
  {
    const SytrdJob* djobs = reinterpret_cast<const SytrdJob*>(base + o.jobs);
    for (auto& r : rounds) {
      const int* dcta = reinterpret_cast<const int*>(base + o.cta_job) + r.off;
      void* args[] = {(void*)&djobs, (void*)&dcta};
      void* fn = r.kind == 1 ? (void*)sytrd_kernel3 : (void*)sytrd_kernel;
      KFAC_CUDA(cudaLaunchCooperativeKernel(fn, dim3(r.ncta), dim3(SY_THREADS), args, r.smem, stream));
      KFAC_COUNT_LAUNCH(1);
    }
  }

  g_eigh_paths = paths;
  g_timers.sytrd_launches = (int)rounds.size();
  if (stage_mark(1, stream)) return 1;
  // ---- stage 2
  const DcMat* dmats = reinterpret_cast<const DcMat*>(base + o.mats);
  {
    const dim3 g((o.max_n + 255) / 256, count);
    dc_load_d_kernel<<<g, 256, 0, stream>>>(dmats, reinterpret_cast<const float* const*>(base + o.dptr_in)); KFAC_COUNT_LAUNCH(1);
    dc_init_kernel<<<g, 256, 0, stream>>>(dmats, count); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());   // (the eigenvector buffers were set to the identity by eigh_init_kernel)
  }
  std::vector<std::vector<DcNode>> levels;
  for (int i = 0; i < count; ++i) build_nodes(i, 0, ns[i], 0, levels);
  std::vector<DcNode> flat;
  std::vector<int> level_off;
  for (auto& lv : levels) {
    // block prefix of the warp-per-root kernels, restarting with every launch slice of DC_SLICE nodes
    int run = 0;
    for (size_t i = 0; i < lv.size(); ++i) {
      if (i % DC_SLICE == 0) run = 0;
      lv[i].blk0 = run;
      run += (lv[i].hi - lv[i].lo + DC_THREADS / 32 - 1) / (DC_THREADS / 32);
    }
    level_off.push_back((int)flat.size());
    flat.insert(flat.end(), lv.begin(), lv.end());
  }
  level_off.push_back((int)flat.size());
  if (!flat.empty())
    KFAC_CUDA(cudaMemcpyAsync(base + o.nodes, flat.data(), flat.size() * sizeof(DcNode), cudaMemcpyHostToDevice, stream));
  {
    // function attributes belong to the device context: set them once per DEVICE, not once per process
    static int small_attr[KFAC_MAX_DEVICES] = {0};
    int dev = 0;
    KFAC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= KFAC_MAX_DEVICES || !small_attr[dev]) {
      KFAC_CUDA(cudaFuncSetAttribute(larft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 129 * 8));
      KFAC_CUDA(cudaFuncSetAttribute(dc_gemm_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     DC_SMALL * DC_SMALL * 4));
      if (dev >= 0 && dev < KFAC_MAX_DEVICES) small_attr[dev] = 1;
    }
  }
  for (int depth = (int)levels.size() - 1; depth >= 0; --depth) {
    const int nn = (int)levels[depth].size();
    if (nn == 0) continue;
    const DcNode* dn = reinterpret_cast<const DcNode*>(base + o.nodes) + level_off[depth];
    int max_m = 0;
    bool any_small = false;
    for (auto& nd : levels[depth]) {
      max_m = std::max(max_m, nd.hi - nd.lo);
      any_small |= (nd.hi - nd.lo) <= DC_SMALL;
    }
    // grid.y / grid.z are limited to 65535: a level with more nodes than that (hundreds of matrices in one call) goes in slices
    for (int n0 = 0; n0 < nn; n0 += DC_SLICE) {
      const int nc = std::min(DC_SLICE, nn - n0);
      const DcNode* dc = dn + n0;
      dc_prepare_kernel<<<nc, DC_THREADS, 0, stream>>>(dmats, dc); KFAC_COUNT_LAUNCH(1);
      // warp per root / kept entry: ceil(m / 8) blocks per node (that is >= m / 256 blocks for the rank pass as well)
      const DcNode& last = levels[depth][n0 + nc - 1];
      const unsigned gw = (unsigned)(last.blk0 + (last.hi - last.lo + DC_THREADS / 32 - 1) / (DC_THREADS / 32));
      dc_secular_kernel<<<gw, DC_THREADS, 0, stream>>>(dmats, dc, nc); KFAC_COUNT_LAUNCH(1);
      dc_zhat_kernel<<<gw, DC_THREADS, 0, stream>>>(dmats, dc, nc); KFAC_COUNT_LAUNCH(1);
      dc_norm_rank_kernel<<<gw, DC_THREADS, 0, stream>>>(dmats, dc, nc); KFAC_COUNT_LAUNCH(1);
      const int ysplit = std::max(1, std::min(64, max_m / 64));
      dc_fill_u_kernel<<<dim3((max_m + DC_THREADS - 1) / DC_THREADS, ysplit, nc), DC_THREADS, 0, stream>>>(dmats, dc); KFAC_COUNT_LAUNCH(1);
      dc_rotate_kernel<<<dim3((max_m + DC_THREADS - 1) / DC_THREADS, nc), DC_THREADS, 0, stream>>>(dmats, dc); KFAC_COUNT_LAUNCH(1);
    }
    KFAC_CUDA(cudaGetLastError());
    if (any_small) {
      dc_gemm_small_kernel<<<nn, DC_THREADS, DC_SMALL * DC_SMALL * 4, stream>>>(dmats, dn, DC_SMALL); KFAC_COUNT_LAUNCH(1);
      KFAC_CUDA(cudaGetLastError());
    }
    {
      GemmGroup grp;
      int nbig = 0;
      for (auto& nd : levels[depth]) nbig += (nd.hi - nd.lo) > DC_SMALL ? 1 : 0;
      for (auto& nd : levels[depth]) {
        const int m = nd.hi - nd.lo;
        if (m <= DC_SMALL) continue;
        const DcMat& M = mats[nd.mat];
        const size_t off = (size_t)nd.lo * M.n + nd.lo;
        GemmParams p;
        init_params(p);
        p.a = dense_operand(M.Q[(nd.depth + 1) & 1] + off, true, m, M.n);
        p.b = dense_operand(M.U + off, false, m, M.n);
        p.M = p.N = p.K = m;
        p.out = M.Q[nd.depth & 1] + off;
        p.ldo = M.n;
        p.promote_kb = 1;
        if (group_add(grp, p, true, std::max(1, 296 / std::max(nbig, 1)), base + o.partials, o.partials_bytes)) return 1;
      }
      if (group_launch(grp, base + o.gtable, o.gtable_bytes, stream)) return 1;
    }
  }
  dc_store_d_kernel<<<dim3((o.max_n + 255) / 256, count), 256, 0, stream>>>(
      dmats, reinterpret_cast<float* const*>(base + o.dptr_out), clamp_nonneg); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());

  if (stage_mark(2, stream)) return 1;
  // ---- two-stage matrices: eigenvectors of T -> eigenvectors of the band matrix
  if (!ts_ents.empty() && ts_apply_q2(ts_ents, base, o.tsh, stream)) return 1;
  // ---- stage 3: Q <- H_0 ... H_{n-2} Q, compact-WY blocks of 128 reflectors, last block first.
  // Everything that depends only on V is batched over ALL blocks of all matrices first: Gram matrices and T
  // factors (fp64), then TV = T Vb (grouped GEMM).  A block step is then two grouped GEMM launches over all
  // matrices that still have a block at that step:  C2 = TV Z[k0+1:, :]  and  Z[k0+1:, :] -= Vb^T C2.
  {
    std::vector<WyBlock> blocks;
    std::vector<std::vector<int>> blocks_of(count);  // block indices per matrix, ascending k0
    for (int i = 0; i < count; ++i) {
      const int n = ns[i];
      if (n < 2) continue;
      const int nref = n - 1;
      for (int k0 = 0; k0 < nref; k0 += 128) {
        WyBlock b;
        b.Vb = jobs[i].W + (size_t)k0 * jobs[i].ldw + (k0 + 1);
        b.n = jobs[i].ldw;   // row stride of V
        b.nb = std::min(128, nref - k0);
        b.kk = n - k0 - 1;
        b.S = reinterpret_cast<double*>(base + o.Sall) + blocks.size() * 128 * 128;
        b.tau = jobs[i].tau + k0;
        b.T = reinterpret_cast<float*>(base + o.Tall) + blocks.size() * 128 * 128;
        blocks_of[i].push_back((int)blocks.size());
        blocks.push_back(b);
      }
    }
    {
      const int bx = std::max(1, std::min(o.max_n, 512));
      zero_lower_batched_kernel<<<dim3(bx, count), 256, 0, stream>>>(dinits); KFAC_COUNT_LAUNCH(1);
    }
    KFAC_CUDA(cudaGetLastError());
    if (!blocks.empty()) {
      const WyBlock* dblocks = reinterpret_cast<const WyBlock*>(base + o.wyblocks);
      KFAC_CUDA(cudaMemcpyAsync(base + o.wyblocks, blocks.data(), blocks.size() * sizeof(WyBlock),
                                cudaMemcpyHostToDevice, stream));
      wy_gram_kernel<<<dim3(3, (unsigned)blocks.size()), 256, 0, stream>>>(dblocks); KFAC_COUNT_LAUNCH(1);
      larft_kernel<<<(unsigned)blocks.size(), LARFT_THREADS, 128 * 129 * 8, stream>>>(dblocks); KFAC_COUNT_LAUNCH(1);
      KFAC_CUDA(cudaGetLastError());
      // TV = T Vb for every block (nb x kk, stored at row k0 of the matrix's TV buffer, same column offset as Vb)
      GemmGroup grp;
      for (int i = 0; i < count; ++i) {
        const int n = ns[i];
        for (int bi : blocks_of[i]) {
          const WyBlock& b = blocks[bi];
          const int k0 = n - 1 - b.kk;
          GemmParams p;
          init_params(p);
          p.a = dense_operand(b.T, true, b.nb, 128);
          p.b = dense_operand(b.Vb, false, b.kk, b.n);
          p.M = b.nb; p.N = b.kk; p.K = b.nb;
          p.out = reinterpret_cast<float*>(base + o.TV[i]) + (size_t)k0 * n + (k0 + 1);
          p.ldo = n;
          p.promote_kb = 1;
          if (group_add(grp, p, false, 1, nullptr, 0)) return 1;
        }
      }
      if (group_launch(grp, base + o.gtable, o.gtable_bytes, stream)) return 1;
      size_t max_blocks = 0;
      for (auto& v : blocks_of) max_blocks = std::max(max_blocks, v.size());
      for (size_t step = 0; step < max_blocks; ++step) {
        int active = 0;
        for (auto& v : blocks_of) active += step < v.size() ? 1 : 0;
        GemmGroup g1, g2;
        for (int i = 0; i < count; ++i) {
          if (step >= blocks_of[i].size()) continue;
          const int n = ns[i];
          const WyBlock& b = blocks[blocks_of[i][blocks_of[i].size() - 1 - step]];
          const int k0 = n - 1 - b.kk;
          const float* TVb = reinterpret_cast<const float*>(base + o.TV[i]) + (size_t)k0 * n + (k0 + 1);
          float* Zb = Qout[i] + (size_t)(k0 + 1) * n;
          float* C2 = reinterpret_cast<float*>(base + o.C2m[i]);
          GemmParams p;
          init_params(p);  // C2 = TV Z[k0+1:, :]   (nb x n, K = kk: skinny -> split-K)
          p.a = dense_operand(TVb, true, b.nb, n);
          p.b = dense_operand(Zb, false, n, n);
          p.M = b.nb; p.N = n; p.K = b.kk; p.out = C2; p.ldo = n; p.promote_kb = 1;
          if (group_add(g1, p, true, std::max(1, 296 / active), base + o.partials, o.partials_bytes)) return 1;
          init_params(p);  // Z[k0+1:, :] -= Vb^T C2
          p.a = dense_operand(b.Vb, false, b.kk, b.n);
          p.b = dense_operand(C2, false, n, n);
          p.M = b.kk; p.N = n; p.K = b.nb; p.alpha = -1.f; p.beta = 1.f; p.out = Zb; p.ldo = n; p.promote_kb = 1;
          if (group_add(g2, p, false, 1, nullptr, 0)) return 1;
        }
        if (group_launch(g1, base + o.gtable, o.gtable_bytes, stream)) return 1;
        if (group_launch(g2, base + o.gtable, o.gtable_bytes, stream)) return 1;
      }
    }
  }
  if (stage_mark(3, stream)) return 1;
  return 0;
}

// Diagnostics: the two-stage tridiagonalisation of ONE matrix with every intermediate result copied out, so that the
// tests can compare each stage with oracle/twostage_model.py.
static size_t debug_two_stage_workspace_bytes(int n) {
  if (n < 34) return 0;
  EighLayout L;
  const size_t nn = (size_t)n * ((n + 3) / 4 * 4) * 4;
  L.take(nn); L.take(nn); L.take((size_t)n * 4 * 3);
  TsSlots sl;
  ts_layout_matrix(L, n, sl);
  TsShared sh;
  ts_layout_shared(L, std::vector<int>{n}, sh);
  return L.total + 1024;
}

static int debug_two_stage(const float* F, int n, float* W_out, float* tau1_out, float* band_out, float* d_out,
                           float* e_out, float* V2_out, float* tau2_out, void* ws, size_t ws_bytes,
                           cudaStream_t stream) {
  if (n < 34) return fail("kfac_debug_two_stage: n must be at least 34");
  if (F == nullptr || W_out == nullptr || tau1_out == nullptr || band_out == nullptr || d_out == nullptr ||
      e_out == nullptr || V2_out == nullptr || tau2_out == nullptr)
    return fail("kfac_debug_two_stage: null pointer");
  const size_t need = debug_two_stage_workspace_bytes(n);
  if (ws == nullptr || ws_bytes < need) return fail("kfac_debug_two_stage: workspace too small (%zu given, %zu needed)", ws_bytes, need);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  const int ldw = (n + 3) / 4 * 4;
  EighLayout L;
  const size_t nn = (size_t)n * ldw * 4;
  const size_t oW = L.take(nn), oV2 = L.take(nn), ovec = L.take((size_t)n * 4 * 3);
  TsEntry E;
  ts_layout_matrix(L, n, E.slots);
  TsShared sh;
  ts_layout_shared(L, std::vector<int>{n}, sh);
  E.index = 0; E.n = n; E.ldw = ldw;
  memset(&E.dev, 0, sizeof(E.dev));
  E.dev.W = reinterpret_cast<float*>(base + oW);
  E.dev.V2 = reinterpret_cast<float*>(base + oV2);
  E.dev.d = reinterpret_cast<float*>(base + ovec);
  E.dev.e = E.dev.d + n;
  E.dev.tau1 = E.dev.d + 2 * n;
  E.dev.Z = nullptr;
  KFAC_CUDA(cudaMemsetAsync(E.dev.W, 0, nn, stream));
  KFAC_CUDA(cudaMemsetAsync(E.dev.V2, 0, nn, stream));
  KFAC_CUDA(cudaMemcpy2DAsync(E.dev.W, (size_t)ldw * 4, F, (size_t)n * 4, (size_t)n * 4, n, cudaMemcpyDeviceToDevice, stream));
  std::vector<TsEntry> ents{E};
  if (ts_setup(ents, base, sh, stream)) return 1;
  if (ts_tridiagonalise(ents, base, sh, stream, band_out)) return 1;
  const TsMat& M = ents[0].dev;
  KFAC_CUDA(cudaMemcpy2DAsync(W_out, (size_t)n * 4, M.W, (size_t)ldw * 4, (size_t)n * 4, n, cudaMemcpyDeviceToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(tau1_out, M.tau1, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(d_out, M.d, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(e_out, M.e, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
  const size_t tiles = (size_t)ts_total_tiles(n);
  KFAC_CUDA(cudaMemcpyAsync(V2_out, M.V2, tiles * TS_G * TS_B * 4, cudaMemcpyDeviceToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(tau2_out, M.tau2, tiles * TS_G * 4, cudaMemcpyDeviceToDevice, stream));
  return 0;
}

// ================================================================= Newton-Schulz inverse (batched, asynchronous)
// X_{k+1} = X_k (2I - M X_k), M = F + damping I (SPD), X_0 = I / ||M||_inf: eigenvalues of M X_0 lie in (0, 1] so the
// iteration converges quadratically; about log2(16 cond(M)) iterations of two n^3 GEMMs each.  All matrices of a call
// advance together (two grouped tensor-core launches per iteration); the residual max |I - M X| is reduced on the
// device and a per-matrix state decides there whether the matrix is finished -- finished matrices turn the remaining
// iterations into no-ops through the `skip` flag of the grouped GEMM.  No host synchronisation anywhere.
struct NsMat {
  const float* F;
  float* M;        // F + damping I
  float* P;        // M X, then 2I - M X
  float* B[2];     // iterates: B[0] is the caller's output, B[1] scratch
  float* rowsum;   // n
  int n;
};
struct NsState {   // per matrix, device memory; the first four ints are copied out as the caller's `info`
  int iters;       // iterations applied
  int status;      // 0 converged, 1 stopped at max_iters without meeting tol, 2 non-finite iterate (singular input)
  float resid;     // last residual max |I - M X| that was looked at
  int done;        // != 0: skip this matrix from now on
  float prev;      // residual of the previous iteration
  unsigned norm_bits;   // ||M||_inf as float bits (atomicMax on positive floats)
  unsigned resid_bits;  // residual of the current iteration (atomicMax), 0x7fc00000 marks a NaN / Inf
  int pad;
};

// M = F + damping I and the absolute row sums (symmetric input: row sums = column sums), one warp per row
__global__ void __launch_bounds__(256) ns_prepare_kernel(const NsMat* __restrict__ mats, NsState* __restrict__ st,
                                                         float damping) {
  const NsMat m = mats[blockIdx.y];
  const int n = m.n;
  const int lane = threadIdx.x & 31;
  for (int r = blockIdx.x * 8 + (threadIdx.x >> 5); r < n; r += gridDim.x * 8) {
    const float* src = m.F + (size_t)r * n;
    float* dst = m.M + (size_t)r * n;
    float s = 0.f;
    for (int c = lane; c < n; c += 32) {
      float v = src[c];
      if (c == r) v += damping;
      dst[c] = v;
      s += fabsf(v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) atomicMax(&st[blockIdx.y].norm_bits, __float_as_uint(s));   // s >= 0 (NaN compares high: caught later)
  }
}
// X_0 = I / ||M||_inf
__global__ void __launch_bounds__(256) ns_init_kernel(const NsMat* __restrict__ mats, const NsState* __restrict__ st) {
  const NsMat m = mats[blockIdx.y];
  const float inv = 1.f / __uint_as_float(st[blockIdx.y].norm_bits);
  const long long total = (long long)m.n * m.n;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
    const int r = (int)(e / m.n), c = (int)(e - (long long)r * m.n);
    m.B[0][e] = (r == c) ? inv : 0.f;
  }
}
// P <- 2I - P (P = M X) and max |I - P|; a non-finite entry poisons the residual (fmaxf alone would drop a NaN)
__global__ void __launch_bounds__(256) ns_residual_kernel(const NsMat* __restrict__ mats, NsState* __restrict__ st) {
  __shared__ float red[256];
  __shared__ int bad_s;
  NsState& S = st[blockIdx.y];
  if (S.done) return;
  const NsMat m = mats[blockIdx.y];
  if (threadIdx.x == 0) bad_s = 0;
  __syncthreads();
  const long long total = (long long)m.n * m.n;
  float mx = 0.f;
  int bad = 0;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
    const int r = (int)(e / m.n), c = (int)(e - (long long)r * m.n);
    const float id = (r == c) ? 1.f : 0.f;
    const float pv = m.P[e];
    const float d = fabsf(id - pv);
    if (!(d <= 3.0e38f)) bad = 1;
    mx = fmaxf(mx, d);
    m.P[e] = 2.f * id - pv;
  }
  if (bad) bad_s = 1;
  red[threadIdx.x] = mx;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] = fmaxf(red[threadIdx.x], red[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) atomicMax(&S.resid_bits, bad_s ? 0x7fc00000u : __float_as_uint(red[0]));
}
// after both GEMMs of iteration `it`: count it, look at the residual of the iterate that ENTERED it (the update just
// applied squares it once more) and decide whether this matrix is finished
__global__ void ns_decide_kernel(NsState* __restrict__ st, int count, int it, int max_iters, float tol) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  NsState& S = st[i];
  if (S.done) return;
  const unsigned bits = S.resid_bits;
  S.resid_bits = 0u;
  S.iters = it + 1;
  const float r = __uint_as_float(bits);
  S.resid = r;
  if (bits >= 0x7f800000u) {          // Inf / NaN anywhere in M X: the input was singular or not finite
    S.status = 2;
    S.done = 1;
  } else if (r * r <= tol) {           // converged: the step just taken brought the residual to r^2
    S.status = 0;
    S.done = 1;
  } else if (it > 0 && r < 0.1f && r >= 0.5f * S.prev) {   // in the quadratic regime the residual must at least halve: it
    S.status = 0;                                  // has reached the fp32 floor of this matrix (~ cond * eps)
    S.done = 1;
  } else if (it + 1 >= max_iters) {
    S.status = 1;
    S.done = 1;
  }
  S.prev = r;
}
// the final iterate of a matrix that took an odd number of iterations sits in the scratch buffer
__global__ void __launch_bounds__(256) ns_finish_kernel(const NsMat* __restrict__ mats, const NsState* __restrict__ st,
                                                        int* __restrict__ info) {
  const NsMat m = mats[blockIdx.y];
  const NsState S = st[blockIdx.y];
  if (info != nullptr && blockIdx.x == 0 && threadIdx.x < 4) {
    const int v[4] = {S.iters, S.status, __float_as_int(S.resid), 0};
    info[4 * blockIdx.y + threadIdx.x] = v[threadIdx.x];
  }
  if ((S.iters & 1) == 0) return;
  const long long total = (long long)m.n * m.n;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256)
    m.B[0][e] = m.B[1][e];
}

struct NsLayout {
  std::vector<size_t> M, P, B1, rowsum;
  size_t mats, state, tables;
  size_t tables_bytes, total;
};
static NsLayout ns_layout(const int* ns, int count) {
  NsLayout o;
  EighLayout L;
  for (int i = 0; i < count; ++i) {
    const size_t nn = (size_t)ns[i] * ns[i] * 4;
    o.M.push_back(L.take(nn));
    o.P.push_back(L.take(nn));
    o.B1.push_back(L.take(nn));
    o.rowsum.push_back(L.take((size_t)ns[i] * 4));
  }
  o.mats = L.take((size_t)count * sizeof(NsMat));
  o.state = L.take((size_t)count * sizeof(NsState));
  o.tables_bytes = (size_t)4 * ((size_t)count * (sizeof(GemmParams) + 16) + 2048);
  o.tables = L.take(o.tables_bytes);
  o.total = L.total + 512;
  return o;
}
static size_t inverse_batched_workspace_bytes(const int* ns, int count) {
  if (ns == nullptr || count <= 0) return 0;
  return ns_layout(ns, count).total;
}

// info (device, nullable): 4 ints per matrix {iterations, status, residual as float bits, 0}
static int inverse_newton_schulz_batched(const float* const* F, const int* ns, int count, float damping,
                                         float* const* out, int max_iters, float tol, int* info, void* ws,
                                         size_t ws_bytes, cudaStream_t stream) {
  if (count <= 0) return 0;
  if (F == nullptr || ns == nullptr || out == nullptr) return fail("kfac_inverse_newton_schulz_batched: null pointer");
  for (int i = 0; i < count; ++i)
    if (ns[i] <= 0 || F[i] == nullptr || out[i] == nullptr) return fail("kfac_inverse_newton_schulz_batched: bad entry %d", i);
  const NsLayout o = ns_layout(ns, count);
  if (ws == nullptr || ws_bytes < o.total)
    return fail("kfac_inverse_newton_schulz_batched: workspace too small (%zu given, %zu needed)", ws_bytes, o.total);
  if (ensure_gemm_attr()) return 1;
  if (max_iters <= 0) max_iters = 32;
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  std::vector<NsMat> mats(count);
  int max_n = 0;
  for (int i = 0; i < count; ++i) {
    NsMat& m = mats[i];
    m.F = F[i];
    m.M = reinterpret_cast<float*>(base + o.M[i]);
    m.P = reinterpret_cast<float*>(base + o.P[i]);
    m.B[0] = out[i];
    m.B[1] = reinterpret_cast<float*>(base + o.B1[i]);
    m.rowsum = reinterpret_cast<float*>(base + o.rowsum[i]);
    m.n = ns[i];
    max_n = std::max(max_n, ns[i]);
  }
  const NsMat* dmats = reinterpret_cast<const NsMat*>(base + o.mats);
  NsState* dstate = reinterpret_cast<NsState*>(base + o.state);
  KFAC_CUDA(cudaMemcpyAsync(base + o.mats, mats.data(), (size_t)count * sizeof(NsMat), cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemsetAsync(dstate, 0, (size_t)count * sizeof(NsState), stream));
  // the four grouped launches of an iteration pair (P = M X and X' = X P, for even and odd iterations), planned once
  std::vector<char> tables;
  PreparedGroup g1[2], g2[2];
  const int target = std::max(1, sm_count() / count);
  for (int par = 0; par < 2; ++par) {
    GemmGroup a, b;
    for (int i = 0; i < count; ++i) {
      const int n = ns[i];
      const int* skip = &dstate[i].done;
      GemmParams p;
      init_params(p);   // P = M X
      p.a = dense_operand(mats[i].M, true, n, n);
      p.b = dense_operand(mats[i].B[par], false, n, n);
      p.M = p.N = p.K = n; p.out = mats[i].P; p.ldo = n; p.skip = skip;
      if (group_add(a, p, false, target, nullptr, 0)) return 1;
      init_params(p);   // X' = X (2I - M X)
      p.a = dense_operand(mats[i].B[par], true, n, n);
      p.b = dense_operand(mats[i].P, false, n, n);
      p.M = p.N = p.K = n; p.out = mats[i].B[par ^ 1]; p.ldo = n; p.skip = skip;
      if (group_add(b, p, false, target, nullptr, 0)) return 1;
    }
    g1[par] = group_prepare(a, tables);
    g2[par] = group_prepare(b, tables);
  }
  if (tables.size() > o.tables_bytes) return fail("kfac_inverse_newton_schulz_batched: internal table overflow");
  KFAC_CUDA(cudaMemcpyAsync(base + o.tables, tables.data(), tables.size(), cudaMemcpyHostToDevice, stream));
  const int gx = std::max(1, std::min(sm_count() * 4 / count, (max_n * max_n + 255) / 256));
  ns_prepare_kernel<<<dim3(std::max(1, std::min(max_n / 8 + 1, 2 * sm_count() / count + 1)), count), 256, 0, stream>>>(dmats, dstate, damping); KFAC_COUNT_LAUNCH(1);
  ns_init_kernel<<<dim3(gx, count), 256, 0, stream>>>(dmats, dstate); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  for (int it = 0; it < max_iters; ++it) {
    const int par = it & 1;
    if (group_run(g1[par], base + o.tables, stream)) return 1;
    ns_residual_kernel<<<dim3(gx, count), 256, 0, stream>>>(dmats, dstate); KFAC_COUNT_LAUNCH(1);
    if (group_run(g2[par], base + o.tables, stream)) return 1;
    ns_decide_kernel<<<(count + 127) / 128, 128, 0, stream>>>(dstate, count, it, max_iters, tol); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
  }
  ns_finish_kernel<<<dim3(gx, count), 256, 0, stream>>>(dmats, dstate, info); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

static size_t inverse_workspace_bytes(int n) {
  if (n <= 0) return 0;
  return inverse_batched_workspace_bytes(&n, 1) + 1024;
}

// One matrix, with the iteration count handed back on the host: this entry point synchronises the stream.
static int inverse_newton_schulz(const float* F, int n, float damping, float* out, int max_iters, float tol,
                                 int* iters_out, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (F == nullptr || out == nullptr || n <= 0) return fail("kfac_inverse_newton_schulz: bad arguments");
  const size_t need = inverse_workspace_bytes(n);
  if (ws == nullptr || ws_bytes < need)
    return fail("kfac_inverse_newton_schulz: workspace too small (%zu given, %zu needed)", ws_bytes, need);
  int* info = reinterpret_cast<int*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  if (inverse_newton_schulz_batched(&F, &n, 1, damping, &out, max_iters, tol, info,
                                    reinterpret_cast<char*>(info) + 256, ws_bytes - 512, stream))
    return 1;
  int host[4] = {0, 0, 0, 0};
  KFAC_CUDA(cudaMemcpyAsync(host, info, sizeof(host), cudaMemcpyDeviceToHost, stream));
  KFAC_CUDA(cudaStreamSynchronize(stream));
  if (iters_out != nullptr) *iters_out = host[0];
  float resid;
  memcpy(&resid, &host[2], 4);
  if (host[1] == 2) return fail("kfac_inverse_newton_schulz: the iterate became non-finite (singular or non-finite input)");
  if (host[1] == 1)
    return fail("kfac_inverse_newton_schulz: no convergence after %d iterations (residual max|I - MX| = %g)", host[0], resid);
  return 0;
}

}  // namespace kfac
