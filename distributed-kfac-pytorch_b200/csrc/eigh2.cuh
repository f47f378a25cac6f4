// Two-stage tridiagonalisation for the large factors of kfac_eigh_batched (sm_100a).
// Same contract as stage 1 of eigh.cuh (replaces the reduction behind torch.linalg.eigh, kfac/layers/eigen.py:309-311,
// 330-332): on exit d / e hold the tridiagonal matrix and the rows of W the Householder vectors of Q1; in addition the
// bulge-chasing reflectors of Q2 are kept for the back-transformation.  oracle/twostage_model.py is the numpy model of
// exactly this data flow (same layouts, same step order).
//
//   stage 1  sy2sb   dense -> band of half-width 32.  Per panel of 32 rows:
//       ts_panel_kernel  Householder QR of the part right of the band, one 8-CTA cluster per matrix, the panel
//                        spread over the cluster's shared memory; ONE fused pass of 32 dot products and ONE cluster
//                        barrier per column (norm, products with the later columns and the compact-WY T column all
//                        come out of that pass); distributed-shared-memory stores carry the partial sums.
//       gemm3x           Y0 = A22 V on the tensor cores (split-K, all active matrices in one grouped launch)
//       ts_g0 / ts_w     G0 = V^T Y0,  W = Y0 T - 1/2 V (T^T G0 T)
//       gemm3x           A22 -= V W^T + W V^T  as ONE symmetric rank-64 GEMM  [V W] [W V]^T
//   stage 2  sb2st   band -> tridiagonal by bulge chasing in fp64: one warp per sweep, 32 x 32 blocks in registers,
//                    sweeps pipelined behind each other through per-sweep progress counters (release / acquire).
//   back     Q2      eigenvectors <- Q2 eigenvectors: a thread owns a column, keeps a 63-row window in registers and
//                    applies 32 sweeps' reflectors per row tile; sweep groups are pipelined across CTAs per column slab.
#pragma once
#include <cooperative_groups.h>

namespace kfac {
namespace cg = cooperative_groups;

constexpr int TS_B = 32;        // band half-width = panel width = bulge-chasing block
constexpr int TS_G = 32;        // sweeps per back-transformation group
constexpr int TS_CL = 8;        // CTAs per panel cluster
constexpr int TS_PT = 256;      // threads of the panel kernel
constexpr int TS_LDAB = 64;     // doubles per band column: diagonal + 32 sub-diagonals + 31 of bulge
constexpr int TS_CH_WARPS = 8;  // warps per bulge-chasing CTA
constexpr int TS_Q2_T = 128;    // threads (= eigenvector columns) per back-transformation CTA
constexpr int TS_DONE = 1 << 30;

struct TsMat {
  float* W;          // n x ldw working copy (both triangles kept up to date)
  int n, ldw;
  float* tau1;       // n (zero beyond the last reflector)
  float* X1;         // [V^T ; W^T]  (64 x ldx) operand of the rank-64 update
  float* X2;         // [W^T ; V^T]
  int ldx;
  float* Y0;         // n x 32
  float* T;          // 32 x 32 compact-WY factor of the current panel
  float* G0;         // 32 x 32 accumulator of V^T Y0
  double* AB;        // n x 64 lower band storage: AB[j * 64 + d] = band(j + d, j)
  float* V2;         // stage-2 reflectors [(goff[g] + t) * 32 + sl][32]
  float* tau2;       // [(goff[g] + t) * 32 + sl]
  const int* goff;   // ngroups + 1 tile offsets
  int ngroups;
  float* d;          // tridiagonal out
  float* e;
  int* prog;         // n: completed steps per sweep
  int* q2prog;       // [slab][ngroups]: completed tiles per group
  float* Z;          // eigenvector matrix (n x n)
  int cta0, ncta;    // bulge chasing: CTA range of this matrix
};

__device__ __forceinline__ int ld_acquire_s32(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_s32(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ================================================================= stage 1: panel QR on a cluster
// Panel p (k0 = 32 p) is the 32 x m block W[k0 : k0+32, k0+32 : n]; its rows are the columns of the m x 32 matrix
// that is QR-factorised (l = panel column = row of W, c = position along m).  CTA q of the cluster holds positions
// [q * mc, (q+1) * mc) of all 32 panel columns in shared memory.
__global__ void __cluster_dims__(TS_CL, 1, 1) __launch_bounds__(TS_PT)
ts_panel_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active, int k0) {
  extern __shared__ __align__(16) float ts_tile[];
  __shared__ float redbuf[2][TS_CL][TS_B];
  __shared__ float pivbuf[2][TS_B];
  __shared__ float sc[TS_B], betas[TS_B], taus[TS_B];
  __shared__ float Ts[TS_B][TS_B + 1];
  cg::cluster_group cluster = cg::this_cluster();
  const int q = (int)cluster.block_rank();
  const TsMat M = mats[active[blockIdx.x / TS_CL]];
  const int n = M.n, ldw = M.ldw;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m = n - k0 - TS_B;           // live positions (> 1)
  const int m4 = ldw - k0 - TS_B;        // padded to the row pitch of W (multiple of 4; the padding holds zeros)
  const int nb = min(TS_B, m - 1);       // reflectors of this panel
  const int mc = ((m4 + TS_CL - 1) / TS_CL + 3) & ~3;
  const int c_lo = q * mc;
  const int cnt = max(0, min(m4, c_lo + mc) - c_lo);   // own positions (multiple of 4)
  const int pitch = mc + 4;
  float* tile = ts_tile;

  {
    const float* src = M.W + (size_t)k0 * ldw + (k0 + TS_B) + c_lo;
    const int c4n = cnt >> 2;
    for (int idx = tid; idx < TS_B * c4n; idx += TS_PT) {
      const int l = idx / c4n, c4 = idx - l * c4n;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + (size_t)l * ldw) + c4);
      *reinterpret_cast<float4*>(tile + l * pitch + 4 * c4) = v;
    }
    for (int e = tid; e < TS_B * (TS_B + 1); e += TS_PT) (&Ts[0][0])[e] = 0.f;
    if (tid < TS_B) { sc[tid] = 0.f; betas[tid] = 0.f; taus[tid] = 0.f; }
  }
  __syncthreads();
  cluster.sync();   // every CTA of the cluster is running: its shared memory may be written remotely from here on

  const int l = tid >> 3, sub = tid & 7;   // dot products / update: 32 panel columns x 8 position slices
  for (int j = 0; j < nb; ++j) {
    const int par = j & 1;
    const int cs = max(0, j - c_lo);        // first own position that takes part (global position >= j)
    const float* pj = tile + j * pitch;
    float* pl = tile + l * pitch;
    // ---- fused pass: s_l = sum_{c >= j} x_j[c] * x_l[c] for all 32 columns l (rows l < j hold earlier reflectors'
    //      unscaled tails: their sums feed the T column)
    float acc = 0.f;
    for (int c = (cs & ~3) + 4 * sub; c < cnt; c += 32) {
      float4 a = *reinterpret_cast<const float4*>(pj + c);
      const float4 b = *reinterpret_cast<const float4*>(pl + c);
      if (c < cs) {
        if (c + 0 < cs) a.x = 0.f;
        if (c + 1 < cs) a.y = 0.f;
        if (c + 2 < cs) a.z = 0.f;
        if (c + 3 < cs) a.w = 0.f;
      }
      acc = fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, fmaf(a.w, b.w, acc))));
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    // the 8 threads of a column all hold its sum: thread `sub` delivers it to cluster rank `sub`
    *cluster.map_shared_rank(&redbuf[par][q][l], sub) = acc;
    if (j >= c_lo && j < c_lo + cnt) *cluster.map_shared_rank(&pivbuf[par][l], sub) = pl[j - c_lo];   // pivot row
    cluster.sync();

    // ---- every warp recomputes the column's scalars (lane = panel column)
    float S = 0.f;
#pragma unroll
    for (int r = 0; r < TS_CL; ++r) S += redbuf[par][r][lane];
    float pv = pivbuf[par][lane];
    if (lane < j) {   // earlier columns are stored unscaled: v_l = x_l * sc[l]
      const float s = sc[lane];
      S *= s;
      pv *= s;
    }
    const float alpha = __shfl_sync(0xffffffffu, pv, j);
    const float sjj = __shfl_sync(0xffffffffu, S, j);
    float beta, tau, scal;
    if (sjj > 0.f) {
      const float nrm = sqrtf(sjj);
      beta = alpha >= 0.f ? -nrm : nrm;
      tau = (beta - alpha) / beta;
      scal = 1.f / (alpha - beta);
    } else {
      beta = alpha; tau = 0.f; scal = 0.f;
    }
    const float wl = (S - beta * pv) * scal;   // v^T x_l for l > j
    if (warp == 0 && q == 0) {
      // compact-WY column: T[0:j, j] = -tau T[0:j, 0:j] (V[:, 0:j]^T v_j),  V[:, l]^T v_j = pv_l + (S_l - pv_l alpha) scal
      const float z = lane < j ? pv + (S - pv * alpha) * scal : 0.f;
      float tacc = 0.f;
      for (int k = 0; k < j; ++k) tacc = fmaf(Ts[lane][k], __shfl_sync(0xffffffffu, z, k), tacc);
      __syncwarp();
      if (lane < j) Ts[lane][j] = -tau * tacc;
      if (lane == j) Ts[j][j] = tau;
    }
    if (tid == 0) { sc[j] = scal; betas[j] = beta; taus[j] = tau; }
    // ---- update the later columns: x_l -= tau (v^T x_l) v,  v = [1 at position j, x_j * scal beyond]
    const float wmine = __shfl_sync(0xffffffffu, wl, l);
    if (l > j && tau != 0.f) {
      const float f = tau * wmine, g = f * scal;
      for (int c = (cs & ~3) + 4 * sub; c < cnt; c += 32) {
        const float4 a = *reinterpret_cast<const float4*>(pj + c);
        float4 b = *reinterpret_cast<float4*>(pl + c);
        const int gc = c_lo + c;
        if (gc > j) {
          b.x = fmaf(-g, a.x, b.x); b.y = fmaf(-g, a.y, b.y); b.z = fmaf(-g, a.z, b.z); b.w = fmaf(-g, a.w, b.w);
        } else {
          if (gc + 0 > j) b.x = fmaf(-g, a.x, b.x); else if (gc + 0 == j) b.x -= f;
          if (gc + 1 > j) b.y = fmaf(-g, a.y, b.y); else if (gc + 1 == j) b.y -= f;
          if (gc + 2 > j) b.z = fmaf(-g, a.z, b.z); else if (gc + 2 == j) b.z -= f;
          if (gc + 3 > j) b.w = fmaf(-g, a.w, b.w); else if (gc + 3 == j) b.w -= f;
        }
        *reinterpret_cast<float4*>(pl + c) = b;
      }
    }
    __syncthreads();
  }

  // ---- write-out: W in place (R with the betas on its diagonal, reflector tails beyond), explicit V^T for the GEMMs
  {
    float* dstW = M.W + (size_t)k0 * ldw + (k0 + TS_B) + c_lo;
    float* x1 = M.X1 + c_lo;
    float* x2 = M.X2 + (size_t)TS_B * M.ldx + c_lo;
    const int c4n = cnt >> 2;
    for (int idx = tid; idx < TS_B * c4n; idx += TS_PT) {
      const int ll = idx / c4n, c4 = idx - ll * c4n;
      const float4 t4 = *reinterpret_cast<const float4*>(tile + ll * pitch + 4 * c4);
      float t[4] = {t4.x, t4.y, t4.z, t4.w};
      float wv[4], vv[4];
      const bool refl = ll < nb;
      const float s = sc[ll], bt = betas[ll];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int gc = c_lo + 4 * c4 + i;
        if (refl && gc > ll) { wv[i] = t[i] * s; vv[i] = wv[i]; }
        else if (refl && gc == ll) { wv[i] = bt; vv[i] = 1.f; }
        else { wv[i] = t[i]; vv[i] = 0.f; }
      }
      *reinterpret_cast<float4*>(dstW + (size_t)ll * ldw + 4 * c4) = make_float4(wv[0], wv[1], wv[2], wv[3]);
      const float4 v4 = make_float4(vv[0], vv[1], vv[2], vv[3]);
      *reinterpret_cast<float4*>(x1 + (size_t)ll * M.ldx + 4 * c4) = v4;
      *reinterpret_cast<float4*>(x2 + (size_t)ll * M.ldx + 4 * c4) = v4;
    }
    if (q == 0) {
      for (int e = tid; e < TS_B * TS_B; e += TS_PT) {
        M.T[e] = Ts[e >> 5][e & 31];
        M.G0[e] = 0.f;
      }
      if (tid < TS_B) M.tau1[k0 + tid] = taus[tid];
    }
  }
  cluster.sync();   // no CTA may exit while a sibling can still write into its shared memory
}

// G0 += V^T Y0 over a chunk of 64 positions (all chunks of all active matrices in one launch).
__global__ void __launch_bounds__(256) ts_g0_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active, int k0) {
  __shared__ float vs[TS_B][65];
  __shared__ float ys[64][TS_B + 1];
  const TsMat M = mats[active[blockIdx.y]];
  const int m = M.n - k0 - TS_B;
  const int r0 = blockIdx.x * 64;
  if (r0 >= m) return;
  const int rows = min(64, m - r0);
  const int tid = threadIdx.x;
  for (int e = tid; e < TS_B * 64; e += 256) {
    const int i = e >> 6, r = e & 63;
    vs[i][r] = r < rows ? M.X1[(size_t)i * M.ldx + r0 + r] : 0.f;
  }
  for (int e = tid; e < 64 * TS_B; e += 256) {
    const int r = e >> 5, k = e & 31;
    ys[r][k] = r < rows ? M.Y0[(size_t)(r0 + r) * TS_B + k] : 0.f;
  }
  __syncthreads();
  const int i = tid >> 3, k4 = (tid & 7) * 4;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 8
  for (int r = 0; r < 64; ++r) {
    const float v = vs[i][r];
    a0 = fmaf(v, ys[r][k4], a0); a1 = fmaf(v, ys[r][k4 + 1], a1);
    a2 = fmaf(v, ys[r][k4 + 2], a2); a3 = fmaf(v, ys[r][k4 + 3], a3);
  }
  float* g = M.G0 + i * TS_B + k4;
  atomicAdd(g, a0); atomicAdd(g + 1, a1); atomicAdd(g + 2, a2); atomicAdd(g + 3, a3);
}

// W = Y0 T - 1/2 V (T^T G0 T) for a chunk of 64 positions, written transposed into both rank-64 operands.
__global__ void __launch_bounds__(256) ts_w_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active, int k0) {
  __shared__ float Tm[TS_B][TS_B + 1];
  __shared__ float Sm[TS_B][TS_B + 1];   // first G0, then G0 T, finally -1/2 T^T G0 T
  __shared__ float Gt[TS_B][TS_B + 1];
  __shared__ float vs[TS_B][65];
  __shared__ float ys[64][TS_B + 1];
  __shared__ float ws[TS_B][65];
  const TsMat M = mats[active[blockIdx.y]];
  const int m = M.n - k0 - TS_B;
  const int r0 = blockIdx.x * 64;
  if (r0 >= m) return;
  const int rows = min(64, m - r0);
  const int tid = threadIdx.x;
  for (int e = tid; e < TS_B * TS_B; e += 256) {
    Tm[e >> 5][e & 31] = M.T[e];
    Sm[e >> 5][e & 31] = __ldcg(M.G0 + e);
  }
  for (int e = tid; e < TS_B * 64; e += 256) {
    const int i = e >> 6, r = e & 63;
    vs[i][r] = r < rows ? M.X1[(size_t)i * M.ldx + r0 + r] : 0.f;
  }
  for (int e = tid; e < 64 * TS_B; e += 256) {
    const int r = e >> 5, k = e & 31;
    ys[r][k] = r < rows ? M.Y0[(size_t)(r0 + r) * TS_B + k] : 0.f;
  }
  __syncthreads();
  // Gt = G0 T
  for (int e = tid; e < TS_B * TS_B; e += 256) {
    const int i = e >> 5, k = e & 31;
    float a = 0.f;
#pragma unroll 8
    for (int t = 0; t < TS_B; ++t) a = fmaf(Sm[i][t], Tm[t][k], a);
    Gt[i][k] = a;
  }
  __syncthreads();
  // Sm = -1/2 T^T Gt
  for (int e = tid; e < TS_B * TS_B; e += 256) {
    const int i = e >> 5, k = e & 31;
    float a = 0.f;
#pragma unroll 8
    for (int t = 0; t < TS_B; ++t) a = fmaf(Tm[t][i], Gt[t][k], a);
    Sm[i][k] = -0.5f * a;
  }
  __syncthreads();
  {
    const int r = tid >> 2, k8 = (tid & 3) * 8;   // 64 positions x 4 groups of 8 columns
    float a[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) a[u] = 0.f;
#pragma unroll 4
    for (int i = 0; i < TS_B; ++i) {
      const float y = ys[r][i], v = vs[i][r];
#pragma unroll
      for (int u = 0; u < 8; ++u) a[u] = fmaf(y, Tm[i][k8 + u], fmaf(v, Sm[i][k8 + u], a[u]));
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) ws[k8 + u][r] = r < rows ? a[u] : 0.f;
  }
  __syncthreads();
  for (int e = tid; e < TS_B * 64; e += 256) {
    const int k = e >> 6, r = e & 63;
    if (r < rows) {
      const float v = ws[k][r];
      M.X1[(size_t)(TS_B + k) * M.ldx + r0 + r] = v;
      M.X2[(size_t)k * M.ldx + r0 + r] = v;
    }
  }
}

// After the last panel: band -> fp64 lower band storage (room for the bulge zeroed), and the rows of W turned into
// explicit Householder vectors (zeros up to the unit entry at column k + 32; rows without a reflector zeroed).
__global__ void __launch_bounds__(256) ts_band_kernel(const TsMat* __restrict__ mats) {
  const TsMat M = mats[blockIdx.y];
  const int n = M.n, nref = max(0, n - TS_B - 1);
  const int j = blockIdx.x * 4 + (threadIdx.x >> 6);
  const int d = threadIdx.x & 63;
  if (j >= n) return;
  float* row = M.W + (size_t)j * M.ldw;
  float v = 0.f;
  if (d <= TS_B && j + d < n) {
    v = row[j + d];
    if (d >= 1) row[j + d] = (j < nref && d == TS_B) ? 1.f : 0.f;
  }
  M.AB[(size_t)j * TS_LDAB + d] = (double)v;
}

// ================================================================= stage 2: bulge chasing
// total over the 32 lanes of val[c] for every c; lane L returns the total of index L (31 shuffles)
__device__ __forceinline__ double ts_treduce(double (&val)[32], int lane) {
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1) {
    const bool up = (lane & k) != 0;
#pragma unroll
    for (int i = 0; i < k; ++i) {
      const double send = up ? val[i] : val[i + k];
      const double keep = up ? val[i + k] : val[i];
      val[i] = keep + __shfl_xor_sync(0xffffffffu, send, k);
    }
  }
  return val[0];
}
__device__ __forceinline__ double ts_wsum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(TS_CH_WARPS * 32) ts_chase_kernel(const TsMat* __restrict__ mats,
                                                                    const int* __restrict__ cta_mat) {
  __shared__ __align__(16) double bcast[TS_CH_WARPS][2][32];
  const TsMat M = mats[cta_mat[blockIdx.x]];
  const int n = M.n;
  double* __restrict__ AB = M.AB;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int NW = M.ncta * TS_CH_WARPS;
  const int gw = ((int)blockIdx.x - M.cta0) * TS_CH_WARPS + warp;
  double* vb = bcast[warp][0];   // current reflector, one entry per lane
  double* wb = bcast[warp][1];   // z (left application) / w (two-sided update)

  for (int s = gw; s < n - 2; s += NW) {
    const int g = s / TS_G, sl = s - g * TS_G;
    const int tile0 = M.goff[g];
    double tau_prev = 0.0;
    for (int t = 0;; ++t) {
      const int r0 = s + 1 + t * TS_B;
      if (r0 > n - 1) break;
      const int ln = min(TS_B, n - r0);
      if (s > 0) {   // sweep s-1 must have finished its step t+1 (it touches one corner element of this step)
        if (lane == 0) {
          while (ld_acquire_s32(M.prog + (s - 1)) < t + 2) __nanosleep(32);
        }
        __syncwarp();
      }
      const bool rok = lane < ln;
      double e[32];
      double x;   // the column to annihilate, one entry per lane
      const int c0 = r0 - TS_B;
      if (t == 0) {
        x = rok ? __ldcg(AB + (size_t)s * TS_LDAB + 1 + lane) : 0.0;
      } else {
        // sub-diagonal block rows r0.., columns c0..: lane = row, register = column
#pragma unroll
        for (int c = 0; c < 32; ++c)
          e[c] = rok ? __ldcg(AB + (size_t)(c0 + c) * TS_LDAB + (TS_B + lane - c)) : 0.0;
        // right application of the previous reflector (still in vb)
        double y = 0.0;
#pragma unroll
        for (int c = 0; c < 32; ++c) y = fma(e[c], vb[c], y);
        y *= tau_prev;
#pragma unroll
        for (int c = 0; c < 32; ++c) e[c] = fma(-y, vb[c], e[c]);
        x = e[0];
      }
      // Householder of x
      const double alpha = __shfl_sync(0xffffffffu, x, 0);
      const double tail2 = ts_wsum(lane >= 1 ? x * x : 0.0);
      double beta, tau, scal;
      if (tail2 > 0.0) {
        const double nrm = sqrt(alpha * alpha + tail2);
        beta = alpha >= 0.0 ? -nrm : nrm;
        tau = (beta - alpha) / beta;
        scal = 1.0 / (alpha - beta);
      } else {
        beta = alpha; tau = 0.0; scal = 0.0;
      }
      const double v = lane == 0 ? 1.0 : x * scal;   // zero beyond ln (x is zero there)
      __syncwarp();            // everyone is done reading the previous reflector
      vb[lane] = v;
      __syncwarp();
      if (t == 0) {
        if (rok) __stcg(AB + (size_t)s * TS_LDAB + 1 + lane, lane == 0 ? beta : 0.0);
      } else {
        // left application to columns 1..31: z_c = sum_r v_r e_r[c]
        double p[32];
        p[0] = 0.0;
#pragma unroll
        for (int c = 1; c < 32; ++c) p[c] = v * e[c];
        const double z = ts_treduce(p, lane);
        wb[lane] = z;
        __syncwarp();
        const double tv = tau * v;
        e[0] = lane == 0 ? beta : 0.0;
#pragma unroll
        for (int c = 1; c < 32; ++c) e[c] = fma(-tv, wb[c], e[c]);
        if (rok) {
#pragma unroll
          for (int c = 0; c < 32; ++c) __stcg(AB + (size_t)(c0 + c) * TS_LDAB + (TS_B + lane - c), e[c]);
        }
        __syncwarp();
      }
      // two-sided update of the diagonal block rows / columns r0..: lane = row, register c <= lane
#pragma unroll
      for (int c = 0; c < 32; ++c)
        e[c] = (rok && c <= lane) ? __ldcg(AB + (size_t)(r0 + c) * TS_LDAB + (lane - c)) : 0.0;
      if (tau != 0.0) {
        double u = 0.0;
        double p[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          u = fma(e[c], vb[c], u);                 // lower part incl. diagonal: sum_{c <= r} D[r, c] v_c
          p[c] = c < lane ? e[c] * v : 0.0;        // strictly lower entries feed the rows above: D[r, c] v_r -> u_c
        }
        u += ts_treduce(p, lane);
        const double vdv = ts_wsum(v * u);
        const double w = tau * u - 0.5 * tau * tau * vdv * v;
        wb[lane] = w;
        __syncwarp();
#pragma unroll
        for (int c = 0; c < 32; ++c) e[c] -= v * wb[c] + w * vb[c];
        if (rok) {
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c <= lane) __stcg(AB + (size_t)(r0 + c) * TS_LDAB + (lane - c), e[c]);
        }
      }
      // keep the reflector for the back-transformation
      M.V2[((size_t)(tile0 + t) * TS_G + sl) * TS_B + lane] = (float)v;
      if (lane == 0) M.tau2[(size_t)(tile0 + t) * TS_G + sl] = (float)tau;
      tau_prev = tau;
      __syncwarp();
      if (lane == 0) st_release_s32(M.prog + s, t + 1);
    }
    __syncwarp();
    if (lane == 0) st_release_s32(M.prog + s, TS_DONE);
  }
}

__global__ void ts_store_de_kernel(const TsMat* __restrict__ mats) {
  const TsMat M = mats[blockIdx.y];
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < M.n; j += gridDim.x * blockDim.x) {
    M.d[j] = (float)M.AB[(size_t)j * TS_LDAB];
    M.e[j] = j + 1 < M.n ? (float)M.AB[(size_t)j * TS_LDAB + 1] : 0.f;
  }
}

// ================================================================= back-transformation with the stage-2 reflectors
// ctas[b] = {matrix, column slab, pipeline lane k, pipeline depth NP}: the CTA applies sweep groups
// ngroups-1-k, ngroups-1-k-NP, ... to its 128 columns of Z; group g may work on row tile t once group g+1 (applied
// before it) has written its tiles 0..t.
__global__ void __launch_bounds__(TS_Q2_T) ts_q2_kernel(const TsMat* __restrict__ mats, const int4* __restrict__ ctas) {
  __shared__ __align__(16) float vs[TS_G][TS_B];
  __shared__ float tsm[TS_G];
  const int4 me = ctas[blockIdx.x];
  const TsMat M = mats[me.x];
  const int n = M.n, slab = me.y, NP = me.w;
  const int tid = threadIdx.x;
  const int col = slab * TS_Q2_T + tid;
  const bool cok = col < n;
  float* __restrict__ Z = M.Z + (cok ? col : 0);
  int* prog = M.q2prog + (size_t)slab * M.ngroups;

  for (int g = M.ngroups - 1 - me.z; g >= 0; g -= NP) {
    const int tile0 = M.goff[g];
    const int nt = M.goff[g + 1] - tile0;
    const int nt_up = g + 1 < M.ngroups ? M.goff[g + 2] - M.goff[g + 1] : 0;
    const int base = g * TS_G + 1;
    float z[TS_B + TS_G - 1];
    for (int t = 0; t < nt; ++t) {
      if (g + 1 < M.ngroups && NP > 1) {
        if (tid == 0) {
          const int need = min(t + 1, nt_up);
          while (ld_acquire_s32(prog + g + 1) < need) __nanosleep(64);
        }
      }
      __syncthreads();   // (also: the previous tile's reflectors are no longer being read)
      {
        const float4* src = reinterpret_cast<const float4*>(M.V2 + (size_t)(tile0 + t) * TS_G * TS_B);
        float4* dst = reinterpret_cast<float4*>(&vs[0][0]);
        dst[tid] = src[tid];
        dst[tid + TS_Q2_T] = src[tid + TS_Q2_T];
        if (tid < TS_G) tsm[tid] = M.tau2[(size_t)(tile0 + t) * TS_G + tid];
      }
      const int row0 = base + t * TS_B;
      if (t == 0) {
#pragma unroll
        for (int i = 0; i < TS_G - 1; ++i) {
          const int r = row0 + i;
          z[i] = (cok && r < n) ? __ldcg(Z + (size_t)r * n) : 0.f;
        }
      }
#pragma unroll
      for (int i = TS_G - 1; i < TS_B + TS_G - 1; ++i) {
        const int r = row0 + i;
        z[i] = (cok && r < n) ? __ldcg(Z + (size_t)r * n) : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int sl = TS_G - 1; sl >= 0; --sl) {
        const float tau = tsm[sl];
        if (tau != 0.f) {
          float dot = 0.f;
#pragma unroll
          for (int i = 0; i < TS_B; ++i) dot = fmaf(vs[sl][i], z[sl + i], dot);
          dot *= tau;
#pragma unroll
          for (int i = 0; i < TS_B; ++i) z[sl + i] = fmaf(-dot, vs[sl][i], z[sl + i]);
        }
      }
      // the top 32 rows of the window are final for this group
#pragma unroll
      for (int i = 0; i < TS_B; ++i) {
        const int r = row0 + i;
        if (cok && r < n) __stcg(Z + (size_t)r * n, z[i]);
      }
#pragma unroll
      for (int i = 0; i < TS_G - 1; ++i) z[i] = z[i + TS_B];
      if (NP > 1) {
        __syncthreads();
        if (tid == 0 && t + 1 < nt) st_release_s32(prog + g, t + 1);
      }
    }
    if (nt > 0) {
      const int row0 = base + nt * TS_B;
#pragma unroll
      for (int i = 0; i < TS_G - 1; ++i) {
        const int r = row0 + i;
        if (cok && r < n) __stcg(Z + (size_t)r * n, z[i]);
      }
    }
    if (NP > 1) {
      __syncthreads();
      if (tid == 0) st_release_s32(prog + g, TS_DONE);
    }
  }
}


// ================================================================= host orchestration of the two-stage path
constexpr int TS_MIN_N = 6000;   // smaller matrices stay on the one-stage kernels (measured crossover of this version)

static int ts_tiles_of_group(int n, int g) {
  const int s0 = g * TS_G;
  if (s0 > n - 3) return 0;
  return (n - 2 - s0) / TS_B + 1;
}
static int ts_ngroups(int n) { return (std::max(n - 2, 0) + TS_G - 1) / TS_G; }
static int ts_total_tiles(int n) {
  int t = 0;
  for (int g = 0; g < ts_ngroups(n); ++g) t += ts_tiles_of_group(n, g);
  return t;
}
static int ts_panels(int n) {   // panels with something to annihilate: n - 32 p - 32 > 1
  int p = 0;
  while (n - TS_B * p - TS_B > 1) ++p;
  return p;
}

struct TsSlots {   // byte offsets into the workspace for one two-stage matrix
  size_t X1, X2, Y0, T, G0, AB, tau2, goff, prog, q2prog;
};

struct TsShared {  // offsets of the tables shared by all two-stage matrices of a call
  size_t mats, active, cta_mat, q2ctas, gtables, partials;
  size_t gtables_bytes, partials_bytes, active_ints, q2ctas_max;
};

static void ts_layout_matrix(EighLayout& L, int n, TsSlots& s) {
  const size_t ldx = (size_t)(n + 3) / 4 * 4;
  s.X1 = L.take(2 * TS_B * ldx * 4);
  s.X2 = L.take(2 * TS_B * ldx * 4);
  s.Y0 = L.take((size_t)n * TS_B * 4);
  s.T = L.take(TS_B * TS_B * 4);
  s.G0 = L.take(TS_B * TS_B * 4);
  s.AB = L.take((size_t)n * TS_LDAB * 8);
  s.tau2 = L.take((size_t)std::max(ts_total_tiles(n), 1) * TS_G * 4);
  s.goff = L.take((size_t)(ts_ngroups(n) + 2) * 4);
  s.prog = L.take((size_t)n * 4);
  s.q2prog = L.take((size_t)((n + TS_Q2_T - 1) / TS_Q2_T) * std::max(ts_ngroups(n), 1) * 4);
}

static void ts_layout_shared(EighLayout& L, const std::vector<int>& ts_ns, TsShared& sh) {
  const size_t cnt = std::max<size_t>(ts_ns.size(), 1);
  size_t steps = 0, probs = 0, slabs = 0;
  for (int n : ts_ns) {
    steps = std::max<size_t>(steps, ts_panels(n));
    probs += 2 * (size_t)ts_panels(n);
    slabs += (n + TS_Q2_T - 1) / TS_Q2_T;
  }
  sh.mats = L.take(cnt * sizeof(TsMat));
  sh.active_ints = steps * cnt + cnt;
  sh.active = L.take(sh.active_ints * 4);
  sh.cta_mat = L.take((size_t)KFAC_MAX_SMS * 4 * 4);
  sh.q2ctas_max = std::max<size_t>(slabs * 16, (size_t)KFAC_MAX_SMS * 8) + 16;
  sh.q2ctas = L.take(sh.q2ctas_max * sizeof(int4));
  sh.gtables_bytes = probs * (sizeof(GemmParams) + 16) + steps * 2 * 1024 + 4096;
  sh.gtables = L.take(sh.gtables_bytes);
  sh.partials_bytes = (size_t)64 << 20;
  sh.partials = L.take(sh.partials_bytes);
}

struct TsEntry {   // one two-stage matrix of the call (host side)
  int index;       // position in the caller's batch
  int n, ldw;
  TsSlots slots;
  TsMat dev;
};

static int ts_ensure_attrs() {
  static int done[KFAC_MAX_DEVICES] = {0};
  int dev = 0;
  KFAC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= KFAC_MAX_DEVICES || !done[dev]) {
    KFAC_CUDA(cudaFuncSetAttribute(ts_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    if (dev >= 0 && dev < KFAC_MAX_DEVICES) done[dev] = 1;
  }
  return 0;
}

// Fills the device descriptors and zeroes what must start at zero.  W / d / e / tau1 / V2 / Z come from the caller.
static int ts_setup(std::vector<TsEntry>& ents, char* base, const TsShared& sh, cudaStream_t stream) {
  for (size_t k = 0; k < ents.size(); ++k) {
    TsEntry& E = ents[k];
    TsMat& M = E.dev;
    M.n = E.n;
    M.ldw = E.ldw;
    M.ldx = E.ldw;
    M.X1 = reinterpret_cast<float*>(base + E.slots.X1);
    M.X2 = reinterpret_cast<float*>(base + E.slots.X2);
    M.Y0 = reinterpret_cast<float*>(base + E.slots.Y0);
    M.T = reinterpret_cast<float*>(base + E.slots.T);
    M.G0 = reinterpret_cast<float*>(base + E.slots.G0);
    M.AB = reinterpret_cast<double*>(base + E.slots.AB);
    M.tau2 = reinterpret_cast<float*>(base + E.slots.tau2);
    M.goff = reinterpret_cast<const int*>(base + E.slots.goff);
    M.ngroups = ts_ngroups(E.n);
    M.prog = reinterpret_cast<int*>(base + E.slots.prog);
    M.q2prog = reinterpret_cast<int*>(base + E.slots.q2prog);
    M.cta0 = 0;
    M.ncta = 1;
    std::vector<int> goff(M.ngroups + 2, 0);
    for (int g = 0; g < M.ngroups; ++g) goff[g + 1] = goff[g] + ts_tiles_of_group(E.n, g);
    goff[M.ngroups + 1] = goff[M.ngroups];
    KFAC_CUDA(cudaMemcpyAsync(base + E.slots.goff, goff.data(), goff.size() * 4, cudaMemcpyHostToDevice, stream));
    KFAC_CUDA(cudaMemsetAsync(M.tau2, 0, (size_t)std::max(ts_total_tiles(E.n), 1) * TS_G * 4, stream));
    KFAC_CUDA(cudaMemsetAsync(M.prog, 0, (size_t)E.n * 4, stream));
    KFAC_CUDA(cudaMemsetAsync(M.q2prog, 0, (size_t)((E.n + TS_Q2_T - 1) / TS_Q2_T) * std::max(M.ngroups, 1) * 4, stream));
    KFAC_CUDA(cudaMemsetAsync(M.tau1, 0, (size_t)E.n * 4, stream));
    // X1 / X2: the GEMMs read every column up to the row pitch; columns past the live length must hold finite values
    KFAC_CUDA(cudaMemsetAsync(M.X1, 0, (size_t)2 * TS_B * M.ldx * 4, stream));
    KFAC_CUDA(cudaMemsetAsync(M.X2, 0, (size_t)2 * TS_B * M.ldx * 4, stream));
  }
  return 0;
}

// Stage 1 + stage 2 for all two-stage matrices of the call: on return (stream order) d / e hold the tridiagonal
// matrices, the rows of W the stage-1 reflectors, V2 / tau2 the stage-2 reflectors.
static int ts_tridiagonalise(std::vector<TsEntry>& ents, char* base, const TsShared& sh, cudaStream_t stream,
                             double* band_copy = nullptr) {
  if (ents.empty()) return 0;
  if (ensure_gemm_attr()) return 1;
  if (ts_ensure_attrs()) return 1;
  const int cnt = (int)ents.size();
  int steps = 0;
  for (auto& E : ents) steps = std::max(steps, ts_panels(E.n));

  // ---- bulge-chasing CTA ranges (needed in the descriptors before the upload)
  int occ = 0;
  KFAC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ts_chase_kernel, TS_CH_WARPS * 32, 0));
  if (occ < 1) return fail("kfac_eigh_batched: bulge-chasing kernel does not fit on an SM");
  const int cap = std::min(sm_count() * occ, KFAC_MAX_SMS * 4);
  std::vector<int> want(cnt);
  long long total_want = 0;
  for (int k = 0; k < cnt; ++k) {
    want[k] = std::max(1, (ents[k].n / (2 * TS_B) + TS_CH_WARPS - 1) / TS_CH_WARPS);
    total_want += want[k];
  }
  std::vector<int> cta_mat;
  for (int k = 0; k < cnt; ++k) {
    int c = want[k];
    if (total_want > cap) c = std::max(1, (int)((long long)want[k] * cap / total_want));
    ents[k].dev.cta0 = (int)cta_mat.size();
    ents[k].dev.ncta = c;
    for (int i = 0; i < c; ++i) cta_mat.push_back(k);
  }
  if ((int)cta_mat.size() > cap) return fail("kfac_eigh_batched: too many matrices for one bulge-chasing launch (%d CTAs, room for %d)", (int)cta_mat.size(), cap);

  std::vector<TsMat> dev(cnt);
  for (int k = 0; k < cnt; ++k) dev[k] = ents[k].dev;
  KFAC_CUDA(cudaMemcpyAsync(base + sh.mats, dev.data(), cnt * sizeof(TsMat), cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(base + sh.cta_mat, cta_mat.data(), cta_mat.size() * 4, cudaMemcpyHostToDevice, stream));
  const TsMat* dmats = reinterpret_cast<const TsMat*>(base + sh.mats);

  // ---- per panel step: who is active, and the two grouped GEMMs, all planned up front and uploaded once
  std::vector<int> active_all;
  std::vector<int> act_off(steps + 1, 0);
  std::vector<PreparedGroup> g1(steps), g2(steps);
  std::vector<int> max_m4(steps, 0), max_m(steps, 0);
  std::vector<char> tables;
  for (int p = 0; p < steps; ++p) {
    const int k0 = p * TS_B;
    act_off[p] = (int)active_all.size();
    int nact = 0;
    for (int k = 0; k < cnt; ++k)
      if (p < ts_panels(ents[k].n)) { active_all.push_back(k); ++nact; }
    GemmGroup grp1, grp2;
    for (int a = 0; a < nact; ++a) {
      const TsEntry& E = ents[active_all[act_off[p] + a]];
      const TsMat& M = E.dev;
      const int m = E.n - k0 - TS_B, m4 = E.ldw - k0 - TS_B;
      max_m4[p] = std::max(max_m4[p], m4);
      max_m[p] = std::max(max_m[p], m);
      float* A22 = M.W + (size_t)(k0 + TS_B) * E.ldw + (k0 + TS_B);
      GemmParams q;
      init_params(q);   // Y0 = A22 V   (m x 32, reduction over the padded length: the padding columns are zero)
      q.a = dense_operand(A22, true, m, E.ldw);
      q.b = dense_operand(M.X1, true, TS_B, M.ldx);
      q.M = m; q.N = TS_B; q.K = m4; q.out = M.Y0; q.ldo = TS_B; q.promote_kb = 1;
      if (group_add(grp1, q, true, std::max(1, 2 * sm_count() / nact), base + sh.partials, sh.partials_bytes)) return 1;
      init_params(q);   // A22 -= [V W] [W V]^T   (symmetric, rank 64)
      q.a = dense_operand(M.X1, false, m, M.ldx);
      q.b = dense_operand(M.X2, false, m, M.ldx);
      q.M = m; q.N = m; q.K = 2 * TS_B; q.sym = 1; q.alpha = -1.f; q.beta = 1.f; q.out = A22; q.ldo = E.ldw;
      q.promote_kb = 1;
      if (group_add(grp2, q, false, 1, nullptr, 0)) return 1;
    }
    g1[p] = group_prepare(grp1, tables);
    g2[p] = group_prepare(grp2, tables);
  }
  act_off[steps] = (int)active_all.size();
  if (tables.size() > sh.gtables_bytes) return fail("kfac_eigh_batched: internal group-table overflow (%zu > %zu)", tables.size(), sh.gtables_bytes);
  if (active_all.size() > sh.active_ints) return fail("kfac_eigh_batched: internal active-list overflow");
  if (!tables.empty())
    KFAC_CUDA(cudaMemcpyAsync(base + sh.gtables, tables.data(), tables.size(), cudaMemcpyHostToDevice, stream));
  if (!active_all.empty())
    KFAC_CUDA(cudaMemcpyAsync(base + sh.active, active_all.data(), active_all.size() * 4, cudaMemcpyHostToDevice, stream));

  for (int p = 0; p < steps; ++p) {
    const int k0 = p * TS_B;
    const int nact = act_off[p + 1] - act_off[p];
    if (nact == 0) continue;
    const int* dact = reinterpret_cast<const int*>(base + sh.active) + act_off[p];
    const int mc = (((max_m4[p] + TS_CL - 1) / TS_CL) + 3) & ~3;
    const size_t smem = (size_t)TS_B * (mc + 4) * sizeof(float);
    if (smem > 216 * 1024) return fail("kfac_eigh_batched: panel of %d positions does not fit the cluster's shared memory", max_m4[p]);
    ts_panel_kernel<<<nact * TS_CL, TS_PT, smem, stream>>>(dmats, dact, k0); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
    if (group_run(g1[p], base + sh.gtables, stream)) return 1;
    const dim3 grid((max_m[p] + 63) / 64, nact);
    ts_g0_kernel<<<grid, 256, 0, stream>>>(dmats, dact, k0); KFAC_COUNT_LAUNCH(1);
    ts_w_kernel<<<grid, 256, 0, stream>>>(dmats, dact, k0); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
    if (group_run(g2[p], base + sh.gtables, stream)) return 1;
  }
  int max_n = 0;
  for (auto& E : ents) max_n = std::max(max_n, E.n);
  ts_band_kernel<<<dim3((max_n + 3) / 4, cnt), 256, 0, stream>>>(dmats); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  if (band_copy != nullptr)   // diagnostics (kfac_debug_two_stage): the band before the bulge chasing overwrites it
    KFAC_CUDA(cudaMemcpyAsync(band_copy, ents[0].dev.AB, (size_t)ents[0].n * TS_LDAB * 8, cudaMemcpyDeviceToDevice, stream));
  {
    const int* dcta = reinterpret_cast<const int*>(base + sh.cta_mat);
    void* args[] = {(void*)&dmats, (void*)&dcta};
    KFAC_CUDA(cudaLaunchCooperativeKernel((void*)ts_chase_kernel, dim3((unsigned)cta_mat.size()), dim3(TS_CH_WARPS * 32), args, 0, stream));
    KFAC_COUNT_LAUNCH(1);
  }
  ts_store_de_kernel<<<dim3((max_n + 255) / 256, cnt), 256, 0, stream>>>(dmats); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  return 0;
}

// Z <- Q2 Z for all two-stage matrices (between the divide & conquer and the compact-WY back-transformation).
static int ts_apply_q2(std::vector<TsEntry>& ents, char* base, const TsShared& sh, cudaStream_t stream) {
  if (ents.empty()) return 0;
  const int cnt = (int)ents.size();
  const TsMat* dmats = reinterpret_cast<const TsMat*>(base + sh.mats);
  int occ = 0;
  KFAC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ts_q2_kernel, TS_Q2_T, 0));
  if (occ < 1) return fail("kfac_eigh_batched: back-transformation kernel does not fit on an SM");
  const long long cap = (long long)sm_count() * occ;
  long long slabs = 0;
  int min_groups = 1 << 30;
  for (auto& E : ents) {
    slabs += (E.n + TS_Q2_T - 1) / TS_Q2_T;
    min_groups = std::min(min_groups, std::max(1, ts_ngroups(E.n)));
  }
  int NP = (int)std::max<long long>(1, std::min<long long>(16, cap / std::max<long long>(slabs, 1)));
  NP = std::min(NP, min_groups);
  if ((size_t)(slabs * NP) > sh.q2ctas_max) NP = 1;
  std::vector<int4> ctas;
  for (int k = 0; k < NP; ++k)   // pipeline lane 0 (the groups applied first) gets the lowest block indices
    for (int i = 0; i < cnt; ++i)
      for (int sl = 0; sl < (ents[i].n + TS_Q2_T - 1) / TS_Q2_T; ++sl) ctas.push_back(make_int4(i, sl, k, NP));
  KFAC_CUDA(cudaMemcpyAsync(base + sh.q2ctas, ctas.data(), ctas.size() * sizeof(int4), cudaMemcpyHostToDevice, stream));
  const int4* dctas = reinterpret_cast<const int4*>(base + sh.q2ctas);
  if (NP > 1) {
    void* args[] = {(void*)&dmats, (void*)&dctas};
    KFAC_CUDA(cudaLaunchCooperativeKernel((void*)ts_q2_kernel, dim3((unsigned)ctas.size()), dim3(TS_Q2_T), args, 0, stream));
  } else {
    ts_q2_kernel<<<(unsigned)ctas.size(), TS_Q2_T, 0, stream>>>(dmats, dctas);
    KFAC_CUDA(cudaGetLastError());
  }
  KFAC_COUNT_LAUNCH(1);
  return 0;
}

}  // namespace kfac
