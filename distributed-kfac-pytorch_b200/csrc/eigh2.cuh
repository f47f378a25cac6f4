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
//       ts_av_kernel     Y0 = A22 V (32 columns: one pass over A22, register-tiled fp32 FMAs, k-split partial sums)
//       ts_g0 / ts_w     G0 = V^T Y0,  W = Y0 T - 1/2 V (T^T G0 T)
//       ts_syr2k_kernel  A22 -= V W^T + W V^T  as ONE symmetric rank-64 update  [V W] [W V]^T (upper tiles + mirror)
//   stage 2  sb2st   band -> tridiagonal by bulge chasing: two warps per sweep (sub-diagonal chain / diagonal blocks),
//                    32 x 32 blocks in registers, groups of 8 consecutive sweeps per CTA exchanging the band through a
//                    shared-memory window, groups pipelined behind each other through progress counters.
//   back     Q2      eigenvectors <- Q2 eigenvectors: a thread owns a column, keeps a 63-row window in registers and
//                    applies 32 sweeps' reflectors per row tile; sweep groups are pipelined across CTAs per column slab.
#pragma once
#include <cooperative_groups.h>

namespace kfac {
namespace cg = cooperative_groups;

constexpr int TS_B = 32;        // band half-width = panel width = bulge-chasing block
#ifndef KFAC_TS_G
#define KFAC_TS_G 32
#endif
constexpr int TS_G = KFAC_TS_G; // sweeps per back-transformation group
constexpr int TS_CL = 8;        // CTAs per panel cluster
constexpr int TS_PT = 256;      // threads of the panel kernel
constexpr int TS_LDAB = 64;     // floats per band column: diagonal + 32 sub-diagonals + 31 of bulge
constexpr int TS_Q2_T = 128;    // threads (= eigenvector columns) per back-transformation CTA
constexpr int TS_DONE = 1 << 30;

struct TsMat {
  float* W;          // n x ldw working copy (both triangles kept up to date)
  int n, ldw;
  float* tau1;       // n (zero beyond the last reflector)
  float* X1;         // [V^T ; W^T]  (64 x ldx) operand of the rank-64 update
  float* X2;         // [W^T ; V^T]
  int ldx;
  float* Yp;         // TS_AV_KS_MAX x n x 32: k-split partial sums of Y0 = A22 V
  float* T;          // 32 x 32 compact-WY factor of the current panel
  float* G0;         // 32 x 32 accumulator of V^T Y0
  float* AB;         // n x 64 lower band storage: AB[j * 64 + d] = band(j + d, j)
  float* V2;         // stage-2 reflectors [(goff[g] + t) * 32 + sl][32]
  float* tau2;       // [(goff[g] + t) * 32 + sl]
  const int* goff;   // ngroups + 1 tile offsets
  int ngroups;
  float* d;          // tridiagonal out
  float* e;
  int* prog;         // 2 n: completed E / D steps per sweep
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
    if (warp == 0 && q == 0 && lane < j) {
      // V[:, l]^T v_j = pv_l + (S_l - pv_l alpha) scal, kept in the strict upper part of Ts until the loop is over (the
      // compact-WY recursion is run once per panel, off the per-column critical path)
      Ts[lane][j] = pv + (S - pv * alpha) * scal;
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
      if (warp == 0) {
        // compact-WY factor, column by column: T[0:j, j] = -tau_j T[0:j, 0:j] (V[:, 0:j]^T v_j),  T[j, j] = tau_j
        for (int j = 0; j < nb; ++j) {
          const float tj = taus[j];
          float tacc = 0.f;
          for (int k = 0; k < j; ++k) tacc = fmaf(Ts[lane][k], Ts[k][j], tacc);   // Ts[lane][k] = 0 for k < lane
          __syncwarp();
          if (lane < j) Ts[lane][j] = -tj * tacc;
          if (lane == j) Ts[j][j] = tj;
          __syncwarp();
        }
      }
      __syncthreads();
      for (int e = tid; e < TS_B * TS_B; e += TS_PT) {
        M.T[e] = Ts[e >> 5][e & 31];
        M.G0[e] = 0.f;
      }
      if (tid < TS_B) M.tau1[k0 + tid] = taus[tid];
    }
  }
  cluster.sync();   // no CTA may exit while a sibling can still write into its shared memory
}

// ---------------------------------------------------------------- the two big products of a panel step, in fp32 FMAs
// Both are skinny: Y0 = A22 V has 32 columns, the trailing update has rank 64.  On the general tensor-core GEMM their
// cost was its per-tile fixed work (measured: 850 GB/s on a pass that only has to read A22 once), so they have
// kernels of their own: register-tiled fp32 FMAs, operands staged in shared memory in their natural layouts.
constexpr int TS_AV_KS_MAX = 8;   // k-splits of Y0 = A22 V (partial sums, added up in fixed order by the readers)
constexpr int TS_AV_KC = 64;      // k chunk
#ifndef KFAC_TS_AV_SLICE
#define KFAC_TS_AV_SLICE 1024
#endif
constexpr int TS_AV_SLICE = KFAC_TS_AV_SLICE;   // a CTA of ts_av_kernel walks at most about this many columns
constexpr int TS_AV_PITCH = TS_AV_KC + 4;

// Y0 partial: rows [128 bx, +128) of A22 times V over k slice by: Yp[by][r][0..31] = sum_k A22[r][k] V[k][:].
// 256 threads: thread (ri, cg) owns rows ri + 32 u and columns 4 cg + w (u, w = 0..3); both tiles are k-contiguous in
// shared memory (A22 rows and the rows of V^T are k-contiguous in global memory too: no transposition anywhere).
// The 64-wide k chunks go through a two-deep shared-memory ring filled by 16-byte cp.async copies (zero-filled past
// the edges), so the copy of chunk c+1 is in flight while chunk c is multiplied and no register holds staged data.
constexpr int TS_AV_STAGE = (128 + TS_B) * TS_AV_PITCH;        // floats per ring slot: A tile, then V tile
constexpr int TS_AV_SMEM = 2 * TS_AV_STAGE * (int)sizeof(float);
__device__ __forceinline__ void ts_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void ts_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ts_cp16z(float* smem_dst, const float* gsrc, bool ok) {
  const int sz = ok ? 16 : 0;   // src-size 0: nothing is read, the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}
__global__ void __launch_bounds__(256) ts_av_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active,
                                                    int k0, int ksplit) {
  extern __shared__ __align__(16) float ts_av_smem[];
  const TsMat M = mats[active[blockIdx.z]];
  const int m = M.n - k0 - TS_B;
  const int m4 = M.ldw - k0 - TS_B;
  const int row0 = blockIdx.x * 128;
  if (row0 >= m) return;
  const int tid = threadIdx.x, ri = tid >> 3, cg = tid & 7;
  // k slices have the SAME length for every matrix of the launch (chosen for the largest one), so that the CTAs of a
  // launch do about equal work: a small matrix simply uses fewer slices
  const int chunk = ksplit;   // (the argument carries the slice length, a multiple of TS_AV_KC)
  const int kb = blockIdx.y * chunk, ke = min(m4, kb + chunk);
  if (kb >= m4) return;
  const float* A22 = M.W + (size_t)(k0 + TS_B) * M.ldw + (k0 + TS_B);
  float acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int w = 0; w < 4; ++w) acc[u][w] = 0.f;
  // loader mapping: 16 threads cover one row's 64 floats; 16 rows of A per pass of the CTA, 8 passes; V: 2 passes
  const int lr = tid >> 4, lc = (tid & 15) * 4;
  auto issue = [&](int kc, int slot) {
    float* As = ts_av_smem + slot * TS_AV_STAGE;
    float* Vs = As + 128 * TS_AV_PITCH;
    const int k = kc + lc;
    const bool kin = k < ke;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int r = row0 + lr + 16 * p;
      const bool ok = kin && r < m;
      ts_cp16z(As + (lr + 16 * p) * TS_AV_PITCH + lc, ok ? A22 + (size_t)r * M.ldw + k : A22, ok);
    }
#pragma unroll
    for (int p = 0; p < 2; ++p) {   // column l of V goes to row (l & 3) * 8 + (l >> 2): the 8 column groups of a warp
      const int l = lr + 16 * p;    // then read 8 rows that fall into 8 different bank groups
      ts_cp16z(Vs + (((l & 3) << 3) + (l >> 2)) * TS_AV_PITCH + lc, kin ? M.X1 + (size_t)l * M.ldx + k : M.X1, kin);
    }
    ts_cp_commit();
  };
  issue(kb, 0);
  int slot = 0;
  for (int kc = kb; kc < ke; kc += TS_AV_KC, slot ^= 1) {
    if (kc + TS_AV_KC < ke) {
      issue(kc + TS_AV_KC, slot ^ 1);   // (that slot was released by the barrier at the end of the previous round)
      ts_cp_wait<1>();
    } else {
      ts_cp_wait<0>();
    }
    __syncthreads();
    const float* As = ts_av_smem + slot * TS_AV_STAGE;
    const float* Vs = As + 128 * TS_AV_PITCH;
#pragma unroll 4
    for (int k4 = 0; k4 < TS_AV_KC; k4 += 4) {
      float4 a[4], v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] = *reinterpret_cast<const float4*>(As + (ri + 32 * u) * TS_AV_PITCH + k4);
#pragma unroll
      for (int w = 0; w < 4; ++w) v[w] = *reinterpret_cast<const float4*>(Vs + (cg + 8 * w) * TS_AV_PITCH + k4);   // column 4 cg + w
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int w = 0; w < 4; ++w)
          acc[u][w] = fmaf(a[u].x, v[w].x, fmaf(a[u].y, v[w].y, fmaf(a[u].z, v[w].z, fmaf(a[u].w, v[w].w, acc[u][w]))));
    }
    __syncthreads();   // the slot may be refilled
  }
  float* Yp = M.Yp + (size_t)blockIdx.y * M.n * TS_B;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int r = row0 + ri + 32 * u;
    if (r < m) *reinterpret_cast<float4*>(Yp + (size_t)r * TS_B + 4 * cg) = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
  }
}

// A22 -= V W^T + W V^T: upper 64 x 64 tiles (blockIdx.x walks them row by row), the strictly upper ones are mirrored
// into the lower triangle.  Operands: X1 = [V^T ; W^T], X2 = [W^T ; V^T] (64 x ldx, position-contiguous rows):
// out[i][j] -= sum_k X1[k][i] X2[k][j].  256 threads, thread (ri, cj) owns rows 4 ri + u, columns 4 cj + w.
__global__ void __launch_bounds__(256) ts_syr2k_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active,
                                                       const int* __restrict__ tile_off, int nact, int k0) {
  __shared__ __align__(16) float Ps[2 * TS_B][68];
  __shared__ __align__(16) float Qs[2 * TS_B][68];
  // 1-D grid over the tiles of all active matrices (tile_off = prefix of their tile counts): the trailing matrices of
  // one launch differ in size by up to 20 x, and a 2-D grid sized for the largest one was mostly empty blocks
  int a = 0;
  {
    int b = nact - 1;   // last matrix with tile_off <= blockIdx.x
    while (a < b) {
      const int c = (a + b + 1) >> 1;
      if (tile_off[c] <= (int)blockIdx.x) a = c; else b = c - 1;
    }
  }
  const TsMat M = mats[active[a]];
  const int m = M.n - k0 - TS_B;
  const int T = (m + 63) / 64;
  // tile index -> (I, J), J >= I: row I holds T - I tiles
  int t = (int)blockIdx.x - tile_off[a], I = 0;
  if (t >= T * (T + 1) / 2) return;
  {
    // I = largest i with i * T - i (i - 1) / 2 <= t
    const float tf = (float)T + 0.5f;
    I = (int)(tf - sqrtf(fmaxf(tf * tf - 2.f * (float)t, 0.f)));
    while (I > 0 && I * T - I * (I - 1) / 2 > t) --I;
    while ((I + 1) * T - (I + 1) * I / 2 <= t) ++I;
  }
  const int J = I + (t - (I * T - I * (I - 1) / 2));
  const int i0 = I * 64, j0 = J * 64;
  const int tid = threadIdx.x, ri = tid >> 4, cj = tid & 15;
  const int mx = M.ldw - k0 - TS_B;   // positions that exist in the operand rows
  for (int e = tid; e < 2 * TS_B * 16; e += 256) {
    const int k = e >> 4, c4 = (e & 15) * 4;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(&Ps[k][c4]) = (i0 + c4 < mx) ? __ldcg(reinterpret_cast<const float4*>(M.X1 + (size_t)k * M.ldx + i0 + c4)) : z;
    *reinterpret_cast<float4*>(&Qs[k][c4]) = (j0 + c4 < mx) ? __ldcg(reinterpret_cast<const float4*>(M.X2 + (size_t)k * M.ldx + j0 + c4)) : z;
  }
  __syncthreads();
  float acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int w = 0; w < 4; ++w) acc[u][w] = 0.f;
#pragma unroll 8
  for (int k = 0; k < 2 * TS_B; ++k) {
    const float4 a = *reinterpret_cast<const float4*>(&Ps[k][4 * ri]);
    const float4 b = *reinterpret_cast<const float4*>(&Qs[k][4 * cj]);
    const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int w = 0; w < 4; ++w) acc[u][w] = fmaf(av[u], bv[w], acc[u][w]);
  }
  float* A22 = M.W + (size_t)(k0 + TS_B) * M.ldw + (k0 + TS_B);
  __syncthreads();   // the operand tiles are dead: Ps is reused to transpose the result for the mirror image
  float (*Ts)[68] = Ps;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int r = i0 + 4 * ri + u, c = j0 + 4 * cj;
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < m && c < m) {   // (rows are 16-byte aligned and m4 is a multiple of 4: a float4 never leaves the row)
      float* p = A22 + (size_t)r * M.ldw + c;
      o = *reinterpret_cast<const float4*>(p);
      o.x -= acc[u][0]; o.y -= acc[u][1]; o.z -= acc[u][2]; o.w -= acc[u][3];
      // columns at or beyond m are padding that must stay zero
      if (c + 1 >= m) o.y = 0.f;
      if (c + 2 >= m) o.z = 0.f;
      if (c + 3 >= m) o.w = 0.f;
      *reinterpret_cast<float4*>(p) = o;
    }
    *reinterpret_cast<float4*>(&Ts[4 * ri + u][4 * cj]) = o;
  }
  if (I == J) return;   // a diagonal tile was computed in full
  __syncthreads();
  // mirror image: row (j0 + jr) of the lower triangle gets column jr of the tile
  for (int e = tid; e < 64 * 16; e += 256) {
    const int jr = e >> 4, i4 = (e & 15) * 4;
    const int r = j0 + jr, c = i0 + i4;
    if (r < m && c < m) {
      float4 o = make_float4(Ts[i4][jr], Ts[i4 + 1][jr], Ts[i4 + 2][jr], Ts[i4 + 3][jr]);
      if (c + 1 >= m) o.y = 0.f;
      if (c + 2 >= m) o.z = 0.f;
      if (c + 3 >= m) o.w = 0.f;
      *reinterpret_cast<float4*>(A22 + (size_t)r * M.ldw + c) = o;
    }
  }
}

// G0 += V^T Y0 over a chunk of 64 positions (all chunks of all active matrices in one launch).
__global__ void __launch_bounds__(256) ts_g0_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active, int k0,
                                                    int ksplit) {
  __shared__ float vs[TS_B][65];
  __shared__ float ys[64][TS_B + 1];
  const TsMat M = mats[active[blockIdx.y]];
  const int m = M.n - k0 - TS_B;
  const int r0 = blockIdx.x * 64;
  if (r0 >= m) return;
  const int rows = min(64, m - r0);
  const int tid = threadIdx.x;
  const int nsp = (M.ldw - k0 - TS_B + ksplit - 1) / ksplit;   // k slices of this matrix (ksplit = slice length)
  for (int e = tid; e < TS_B * 64; e += 256) {
    const int i = e >> 6, r = e & 63;
    vs[i][r] = r < rows ? M.X1[(size_t)i * M.ldx + r0 + r] : 0.f;
  }
  for (int e = tid; e < 64 * TS_B; e += 256) {
    const int r = e >> 5, k = e & 31;
    float y = 0.f;   // the k-split partial sums of Y0 = A22 V, added in fixed order
    if (r < rows)
      for (int sp = 0; sp < nsp; ++sp) y += __ldcg(M.Yp + ((size_t)sp * M.n + r0 + r) * TS_B + k);
    ys[r][k] = y;
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
__global__ void __launch_bounds__(256) ts_w_kernel(const TsMat* __restrict__ mats, const int* __restrict__ active, int k0,
                                                   int ksplit) {
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
  const int nsp = (M.ldw - k0 - TS_B + ksplit - 1) / ksplit;   // k slices of this matrix (ksplit = slice length)
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
    float y = 0.f;   // the k-split partial sums of Y0 = A22 V, added in fixed order
    if (r < rows)
      for (int sp = 0; sp < nsp; ++sp) y += __ldcg(M.Yp + ((size_t)sp * M.n + r0 + r) * TS_B + k);
    ys[r][k] = y;
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

// After the last panel: band -> lower band storage (room for the bulge zeroed), and the rows of W turned into
// explicit Householder vectors (zeros up to the unit entry at column k + 32; rows without a reflector zeroed).
__global__ void __launch_bounds__(256) ts_band_kernel(const TsMat* __restrict__ mats) {
  const TsMat M = mats[blockIdx.y];
  const int n = M.n, nref = max(0, n - TS_B - 1);
  const int j = blockIdx.x * 4 + (threadIdx.x >> 6);
  const int d = threadIdx.x & 63;
  if (j >= n + 64) return;
  if (j >= n) {   // zero columns beyond the matrix
    M.AB[(size_t)j * TS_LDAB + d] = 0.f;
    return;
  }
  float* row = M.W + (size_t)j * M.ldw;
  float v = 0.f;
  if (d <= TS_B && j + d < n) {
    v = row[j + d];
    if (d >= 1) row[j + d] = (j < nref && d == TS_B) ? 1.f : 0.f;
  }
  M.AB[(size_t)j * TS_LDAB + d] = v;
}

// ================================================================= stage 2: bulge chasing
// total over the 32 lanes of val[c] for every c; lane L returns the total of index L (31 shuffles)
__device__ __forceinline__ float ts_treduce(float (&val)[32], int lane) {
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1) {
    const bool up = (lane & k) != 0;
#pragma unroll
    for (int i = 0; i < k; ++i) {
      const float send = up ? val[i] : val[i + k];
      const float keep = up ? val[i + k] : val[i];
      val[i] = keep + __shfl_xor_sync(0xffffffffu, send, k);
    }
  }
  return val[0];
}
__device__ __forceinline__ float ts_wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sweep s annihilates column s below the sub-diagonal and chases the bulge down the band in steps of 32 rows.  Step t
// works on rows r0 = s+1+32t .. r0+31 and has two parts:
//   E(s,t)  the sub-diagonal block (rows r0.., columns r0-32..) gets the previous reflector from the right, its first
//           column is annihilated by a new reflector v_t, which is applied to the block from the left;
//   D(s,t)  the diagonal block (rows / columns r0..) gets v_t from both sides.
// Within a sweep only the E parts form a chain (E(s,t+1) needs v_t); across sweeps E(s+1,t) needs E(s,t+1) and D(s,t),
// and D(s+1,t) needs D(s,t+1).  A CTA therefore runs a GROUP of 8 consecutive sweeps with TWO warps each -- one for
// the E chain, one trailing it with the D parts -- staggered behind each other; the part of the band the group works
// on lives in a circular shared-memory window of 768 columns, so consecutive sweeps hand their data over on chip.
// Only the group's first sweep reads the band from global memory (what the previous group's last sweep wrote, guarded
// by that sweep's two progress counters) and only its last sweep writes it back.  Groups of one matrix are pipelined
// over the matrix's CTAs.  Each sweep writes its finished column straight to d / e.
// The band storage and the window are zero beyond the matrix, so no row / column needs a bounds check.
// 768 columns: 8 sweeps, each two steps (64 columns) behind the one in front, plus the blocks of the first and the
// last, need about 544; with 512 the throttle that keeps the group inside the window was what set the pace
#ifndef KFAC_TS_SWEEPS
#define KFAC_TS_SWEEPS 8
#endif
constexpr int TS_CH_SWEEPS = KFAC_TS_SWEEPS;
// window columns: the sweeps of a group run two steps (64 columns) apart, plus the blocks of the first and the last
constexpr int TS_WINC = TS_CH_SWEEPS == 8 ? 768 : (TS_CH_SWEEPS == 4 ? 384 : 256);
// the first sweep may be this many steps ahead of the last one's finished step before its prefetch would overwrite
// window columns still in use (E part: columns up to r0 + 31; D part, one step further ahead: up to r0 + 63)
constexpr int TS_THROTTLE_E = (TS_WINC - 57 - 32 * (TS_CH_SWEEPS - 1) / 4) / 32;
constexpr int TS_THROTTLE_D = (TS_WINC - 89 - 32 * (TS_CH_SWEEPS - 1) / 4) / 32;
constexpr int TS_WSIZE = TS_WINC * TS_LDAB;
// window position of base + offset: the base is reduced modulo the window once per step, an offset inside a block is
// less than a window, so one conditional subtraction per access does the wrap
__device__ __forceinline__ int ts_wrap(int a) { return a >= TS_WSIZE ? a - TS_WSIZE : a; }
constexpr int TS_CH_THREADS = TS_CH_SWEEPS * 64;

// Progress inside a group: one counter of finished steps per warp in shared memory.  A waiting warp polls with one
// lane and backs off with nanosleep, so that it stays out of the issue slots of the warps that work.  Every wait is
// bounded: a protocol error ends the kernel with a trap (and a message) instead of hanging the GPU.
constexpr unsigned TS_WATCHDOG = 1u << 24;
#ifdef KFAC_TS_TIMING
__device__ unsigned long long g_ts_wait_cycles[16];
__device__ long long g_ts_trace[16][160][3];
#endif
__device__ __forceinline__ void ts_wait_step(volatile int* prog, int sweep, int step, int tag) {
  if (step < 0) return;
#ifdef KFAC_TS_TIMING
  const long long w0 = clock64();
#endif
  if ((threadIdx.x & 31) == 0) {
    unsigned spins = 0;
    while (prog[sweep] <= step) {
#ifndef KFAC_TS_WAIT_MODE
#define KFAC_TS_WAIT_MODE 0
#endif
#if KFAC_TS_WAIT_MODE == 0
      if (spins > 32) __nanosleep(20);   // a hand-over normally arrives within a few polls: only long waits back off
#elif KFAC_TS_WAIT_MODE == 2
      if (spins > 256) __nanosleep(20);
#elif KFAC_TS_WAIT_MODE == 3
      __nanosleep(1);
#endif
      if (++spins > TS_WATCHDOG) {
        printf("kfac ts_chase: wait %d timed out: block %d warp %d waits for sweep %d step %d (at %d)\n", tag,
               (int)blockIdx.x, (int)(threadIdx.x >> 5), sweep, step, prog[sweep]);
        __trap();
      }
    }
    __threadfence_block();
  }
  __syncwarp();
#ifdef KFAC_TS_TIMING
  if ((threadIdx.x & 31) == 0 && blockIdx.x == 0 && (threadIdx.x >> 5) >= 2 && (threadIdx.x >> 5) < 14)
    atomicAdd(&g_ts_wait_cycles[tag], (unsigned long long)(clock64() - w0));
#endif
}
// returns the counter value seen (>= need): the counters only grow, so the caller caches it and asks global memory
// again only when the cached value no longer suffices
__device__ __forceinline__ int ts_wait_global(const int* p, int need, int tag) {
  unsigned spins = 0;
  int seen;
  while ((seen = ld_acquire_s32(p)) < need) {
    if (++spins > (1u << 21)) {
      printf("kfac ts_chase: global wait %d timed out: block %d warp %d needs %d, sees %d\n", tag, (int)blockIdx.x,
             (int)(threadIdx.x >> 5), need, seen);
      __trap();
    }
  }
  return seen;
}
__device__ __forceinline__ void ts_cp4(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}

__global__ void __launch_bounds__(TS_CH_THREADS, TS_CH_SWEEPS == 8 ? 1 : 2) ts_chase_kernel(const TsMat* __restrict__ mats,
                                                                    const int* __restrict__ cta_mat) {
  extern __shared__ __align__(16) float ts_win[];   // TS_WINC x 64
  // reflector of step t in slot t & 3 (E warp -> D warp): four slots, so that the E chain may run three steps ahead
  // of its own D parts (with two slots the hand-over latencies of that loop set the pace of the whole pipeline)
  __shared__ __align__(16) float vbuf[TS_CH_SWEEPS][4][32];
  __shared__ float taubuf[TS_CH_SWEEPS][4];
  __shared__ __align__(16) float zbuf[2 * TS_CH_SWEEPS][32];  // per-warp broadcast scratch
  __shared__ int ebar_s[TS_CH_SWEEPS], dbar_s[TS_CH_SWEEPS];   // finished E / D steps per sweep
  const TsMat M = mats[cta_mat[blockIdx.x]];
  const int n = M.n;
  float* __restrict__ AB = M.AB;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = warp >> 1;             // sweep within the group
  const bool drole = (warp & 1) != 0;  // odd warps trail with the diagonal blocks
  const int crank = (int)blockIdx.x - M.cta0, NC = M.ncta;
  const int ngr = (n - 2 + TS_CH_SWEEPS - 1) / TS_CH_SWEEPS;
  float* zb = zbuf[warp];
  volatile int* ebar = ebar_s;
  volatile int* dbar = dbar_s;

#ifdef KFAC_TS_TIMING
  long long dbg_wait = 0, dbg_work = 0, dbg_steps = 0;
  const long long dbg_t0 = clock64();
#endif
  for (int gi = crank; gi < ngr; gi += NC) {
    const int s0 = gi * TS_CH_SWEEPS;
    const int nact = min(TS_CH_SWEEPS, n - 2 - s0);
    __syncthreads();   // the previous group of this CTA is completely done with the window and the counters
    if (threadIdx.x < TS_CH_SWEEPS) { ebar_s[threadIdx.x] = 0; dbar_s[threadIdx.x] = 0; }
    __syncthreads();
    if (i >= nact) continue;
    const int s = s0 + i;
    const bool lead = i == 0, tailw = i == nact - 1;
    int* gE = M.prog + 2 * s;        // this sweep's counters in global memory (published by the group's last sweep)
    int* gD = gE + 1;
    const int* pE = gE - 2;          // the sweep in front
    const int* pD = gE - 1;

    if (!drole) {
      // ============================================================ E chain
      const int g = s / TS_G, sl = s - g * TS_G;
      const int tile0 = M.goff[g];
      float tau_prev = 0.f;
      // The group's first sweep brings the band in from global memory: block E(t+1) is copied into the window with
      // cp.async while step t is computed, so that the copy latency (and the look at the previous group's counters)
      // stays off the chain; from then on it works on the window exactly like the sweeps behind it.
      int seenE = 0, seenD = 0;   // lane 0: last values read from the counters of the sweep in front
      int peekE = 0, peekD = 0;   // lane 0: reads of those counters that are still in flight
      auto fetch_e = [&](int k) {   // needs E(s-1, k+1) and D(s-1, k) done
        if (s > 0 && lane == 0) {
          // the counters were asked for one step ago (peek*): their round trip to global memory is behind us
          seenE = max(seenE, peekE);
          seenD = max(seenD, peekD);
          if (seenE < k + 2) seenE = ts_wait_global(pE, k + 2, 1);
          if (seenD < k + 1) seenD = ts_wait_global(pD, k + 1, 2);
          peekE = ld_acquire_s32(pE);   // for the next step; not looked at before then
          peekD = ld_acquire_s32(pD);
        }
        __syncwarp();
        if (k == 0) {   // column s: the diagonal entry and the 32 entries below it
          const int sw = (s * TS_LDAB) % TS_WSIZE;
          ts_cp4(ts_win + ts_wrap(sw + 1 + lane), AB + (size_t)s * TS_LDAB + 1 + lane);
          if (lane == 0) ts_cp4(ts_win + sw, AB + (size_t)s * TS_LDAB);
        } else {
          const int eb = (s + 1 + (k - 1) * TS_B) * TS_LDAB + TS_B + lane;
          const int ebw = eb % TS_WSIZE;
#pragma unroll
          for (int c = 0; c < 32; ++c) ts_cp4(ts_win + ts_wrap(ebw + 63 * c), AB + (size_t)(eb + 63 * c));
          // the rest of each window column (rows beyond this block) is structurally zero
#pragma unroll
          for (int c = 1; c < 32; ++c)
            if (lane < c) ts_win[ts_wrap(ebw + TS_B + 63 * c)] = 0.f;
        }
        ts_cp_commit();
      };
      if (lead) fetch_e(0);
      for (int t = 0;; ++t) {
        const int r0 = s + 1 + t * TS_B;
        if (r0 > n - 1) break;
#ifdef KFAC_TS_TIMING
        const long long tw0 = clock64();
#endif
        if (lead) {
          if (r0 + TS_B <= n - 1) {
            if (nact > 1) ts_wait_step(dbar, nact - 1, t - TS_THROTTLE_E, 3);   // keep the group inside the window
            fetch_e(t + 1);
            ts_cp_wait<1>();
          } else {
            ts_cp_wait<0>();
          }
          __syncwarp();
        } else {
          ts_wait_step(ebar, i - 1, t + 1, 4);
          ts_wait_step(dbar, i - 1, t, 5);
        }
        ts_wait_step(dbar, i, t - 4, 6);   // the D warp is done with the reflector slot about to be rewritten
#ifdef KFAC_TS_TIMING
        const long long tw1 = clock64();
        dbg_wait += tw1 - tw0;
#endif
        const int ln = min(TS_B, n - r0);
        const int c0 = r0 - TS_B;
        const int EB = c0 * TS_LDAB + TS_B + lane;   // linear index of (row r0+lane, column c0); column c adds 63 c
        const int EBw = t > 0 ? EB % TS_WSIZE : 0;    // ... and its position in the window
        float e[32];
        float x;
        if (t == 0) {
          const int sw = (s * TS_LDAB) % TS_WSIZE;
          x = ts_win[ts_wrap(sw + 1 + lane)];
          if (lane == 0) M.d[s] = ts_win[sw];
          if (lane >= ln) x = 0.f;
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) e[c] = ts_win[ts_wrap(EBw + 63 * c)];
          if (ln < TS_B && lane >= ln) {   // window slots of rows beyond the matrix hold leftovers: they must read as zero
#pragma unroll
            for (int c = 0; c < 32; ++c) e[c] = 0.f;
          }
          const float* vp = vbuf[i][(t - 1) & 3];
          float y0 = 0.f, y1 = 0.f, y2 = 0.f, y3 = 0.f;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 v4 = *reinterpret_cast<const float4*>(vp + c);
            y0 = fmaf(e[c], v4.x, y0); y1 = fmaf(e[c + 1], v4.y, y1);
            y2 = fmaf(e[c + 2], v4.z, y2); y3 = fmaf(e[c + 3], v4.w, y3);
          }
          const float y = ((y0 + y1) + (y2 + y3)) * tau_prev;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 v4 = *reinterpret_cast<const float4*>(vp + c);
            e[c] = fmaf(-y, v4.x, e[c]); e[c + 1] = fmaf(-y, v4.y, e[c + 1]);
            e[c + 2] = fmaf(-y, v4.z, e[c + 2]); e[c + 3] = fmaf(-y, v4.w, e[c + 3]);
          }
          x = e[0];
        }
        // ---- Householder of x (rows beyond the matrix hold zeros)
        const float alpha = __shfl_sync(0xffffffffu, x, 0);
        const float tail2 = ts_wsum(lane >= 1 ? x * x : 0.f);
        float beta, tau, scal;
        if (tail2 > 0.f) {
          const float nrm = sqrtf(fmaf(alpha, alpha, tail2));
          beta = alpha >= 0.f ? -nrm : nrm;
          tau = (beta - alpha) / beta;
          scal = 1.f / (alpha - beta);
        } else {
          beta = alpha; tau = 0.f; scal = 0.f;
        }
        const float v = lane == 0 ? 1.f : x * scal;
        vbuf[i][t & 3][lane] = v;
        if (lane == 0) taubuf[i][t & 3] = tau;
        M.V2[((size_t)(tile0 + t) * TS_G + sl) * TS_B + lane] = v;
        if (lane == 0) M.tau2[(size_t)(tile0 + t) * TS_G + sl] = tau;
        if (t == 0) {
          if (lane == 0) M.e[s] = beta;   // column s is final: (s+1, s) = beta, zeros below
        } else {
          // left application to columns 1..31: z_c = sum_r v_r e_r[c]
          float p[32];
          p[0] = 0.f;
#pragma unroll
          for (int c = 1; c < 32; ++c) p[c] = v * e[c];
          const float z = ts_treduce(p, lane);
          zb[lane] = z;
          __syncwarp();
          const float tv = tau * v;
          e[0] = lane == 0 ? beta : 0.f;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 z4 = *reinterpret_cast<const float4*>(zb + c);
            if (c > 0) e[c] = fmaf(-tv, z4.x, e[c]);
            e[c + 1] = fmaf(-tv, z4.y, e[c + 1]);
            e[c + 2] = fmaf(-tv, z4.z, e[c + 2]);
            e[c + 3] = fmaf(-tv, z4.w, e[c + 3]);
          }
          if (tailw) {
            // the counter in global memory runs one step late: the release (which waits for this warp's earlier
            // stores) then finds the previous step's stores long since written instead of stalling on fresh ones
            if (lane == 0) st_release_s32(gE, t);
#pragma unroll
            for (int c = 0; c < 32; ++c) __stcg(AB + (size_t)(EB + 63 * c), e[c]);
          } else {
#pragma unroll
            for (int c = 0; c < 32; ++c) ts_win[ts_wrap(EBw + 63 * c)] = e[c];
            // the annihilated entries are final and lie left of everything the group's last sweep will touch: the
            // copy in global memory (which the next group reads) must learn about them here
            if (lane >= 1) __stcg(AB + (size_t)EB, 0.f);
          }
        }
        tau_prev = tau;
        __syncwarp();
        if (lane == 0) {
          __threadfence_block();
          ebar[i] = t + 1;
        }
#ifdef KFAC_TS_TIMING
        dbg_work += clock64() - tw1;
        ++dbg_steps;
        if (lane == 0 && blockIdx.x == M.cta0 && gi == crank && n > 4000 && t < 160) {
          g_ts_trace[warp][t][0] = tw0 - dbg_t0; g_ts_trace[warp][t][1] = tw1 - dbg_t0; g_ts_trace[warp][t][2] = clock64() - dbg_t0;
        }
#endif
      }
      __syncwarp();
      if (lane == 0) {
        if (tailw) st_release_s32(gE, TS_DONE);
        __threadfence_block();
        ebar[i] = TS_DONE;   // the sweep behind may ask for one step more than exists
      }
    } else {
      // ============================================================ D parts, trailing the E chain of the same sweep
      int seenE = 0, seenD = 0, peekE = 0, peekD = 0;
      auto fetch_d = [&](int k) {   // lower triangle of D(k); needs E(s-1, k+1) and D(s-1, k+1) done
        if (s > 0 && lane == 0) {
          seenE = max(seenE, peekE);
          seenD = max(seenD, peekD);
          if (seenE < k + 2) seenE = ts_wait_global(pE, k + 2, 8);
          if (seenD < k + 2) seenD = ts_wait_global(pD, k + 2, 8);
          peekE = ld_acquire_s32(pE);
          peekD = ld_acquire_s32(pD);
        }
        __syncwarp();
        const int db = (s + 1 + k * TS_B) * TS_LDAB + lane;
        const int dbw = db % TS_WSIZE;
#pragma unroll
        for (int c = 0; c < 32; ++c)
          if (c <= lane) ts_cp4(ts_win + ts_wrap(dbw + 63 * c), AB + (size_t)(db + 63 * c));
        ts_cp_commit();
      };
      if (lead) fetch_d(0);
      for (int t = 0;; ++t) {
        const int r0 = s + 1 + t * TS_B;
        if (r0 > n - 1) break;
#ifdef KFAC_TS_TIMING
        const long long tw0 = clock64();
#endif
        if (lead) {
          if (r0 + TS_B <= n - 1) {
            if (nact > 1) ts_wait_step(dbar, nact - 1, t - TS_THROTTLE_D, 3);   // keep the group inside the window
            fetch_d(t + 1);
            ts_cp_wait<1>();
          } else {
            ts_cp_wait<0>();
          }
          __syncwarp();
        } else {
          ts_wait_step(dbar, i - 1, t + 1, 9);
        }
        ts_wait_step(ebar, i, t, 7);   // v_t is there (and, through it, E(s-1, t+1) is done)
#ifdef KFAC_TS_TIMING
        const long long tw1 = clock64();
        dbg_wait += tw1 - tw0;
#endif
        const int ln = min(TS_B, n - r0);
        const float* vb = vbuf[i][t & 3];
        const float v = vb[lane];
        const float tau = taubuf[i][t & 3];
        // the block is symmetric: entry (r, c) with c > r is read from its mirror image (c, r), so every lane holds
        // its full row and the product with v needs no exchange between lanes
        const int DB = r0 * TS_LDAB + lane;        // (row r0+lane, column r0+c), c <= lane: DB + 63 c
        const int DB2 = r0 * TS_LDAB + 63 * lane;  // (row r0+c, column r0+lane), c > lane: DB2 + c
        const int DBw = DB % TS_WSIZE, DB2w = DB2 % TS_WSIZE;
        float e[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) e[c] = ts_win[ts_wrap(c <= lane ? DBw + 63 * c : DB2w + c)];
        if (ln < TS_B) {   // rows / columns beyond the matrix: the window holds leftovers there
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (lane >= ln || c >= ln) e[c] = 0.f;
        }
        if (tau != 0.f) {
          float u0 = 0.f, u1 = 0.f, u2 = 0.f, u3 = 0.f;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 v4 = *reinterpret_cast<const float4*>(vb + c);
            u0 = fmaf(e[c], v4.x, u0); u1 = fmaf(e[c + 1], v4.y, u1);
            u2 = fmaf(e[c + 2], v4.z, u2); u3 = fmaf(e[c + 3], v4.w, u3);
          }
          const float u = (u0 + u1) + (u2 + u3);
          const float vdv = ts_wsum(v * u);
          const float wv = tau * u - 0.5f * tau * tau * vdv * v;
          zb[lane] = wv;
          __syncwarp();
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 v4 = *reinterpret_cast<const float4*>(vb + c);
            const float4 w4 = *reinterpret_cast<const float4*>(zb + c);
            e[c] -= v * w4.x + wv * v4.x;
            e[c + 1] -= v * w4.y + wv * v4.y;
            e[c + 2] -= v * w4.z + wv * v4.z;
            e[c + 3] -= v * w4.w + wv * v4.w;
          }
        }
        if (tailw) {
          if (lane == 0) st_release_s32(gD, t);   // one step late, see the E chain
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c <= lane) __stcg(AB + (size_t)(DB + 63 * c), e[c]);
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c <= lane) ts_win[ts_wrap(DBw + 63 * c)] = e[c];
        }
        if (s == n - 3 && t == 0) {   // the last sweep leaves the final 2 x 2 block behind
          if (lane == 0) M.d[n - 2] = e[0];
          if (lane == 1) { M.e[n - 2] = e[0]; M.d[n - 1] = e[1]; M.e[n - 1] = 0.f; }
        }
        __syncwarp();
        if (lane == 0) {
          __threadfence_block();
          dbar[i] = t + 1;
        }
#ifdef KFAC_TS_TIMING
        dbg_work += clock64() - tw1;
        ++dbg_steps;
        if (lane == 0 && blockIdx.x == M.cta0 && gi == crank && n > 4000 && t < 160) {
          g_ts_trace[warp][t][0] = tw0 - dbg_t0; g_ts_trace[warp][t][1] = tw1 - dbg_t0; g_ts_trace[warp][t][2] = clock64() - dbg_t0;
        }
#endif
      }
      __syncwarp();
      if (lane == 0) {
        if (tailw) st_release_s32(gD, TS_DONE);
        __threadfence_block();
        dbar[i] = TS_DONE;
      }
    }
  }
#ifdef KFAC_TS_TIMING
  if (lane == 0 && blockIdx.x == M.cta0 && n > 4000)
    for (int tt = 0; tt < 160; ++tt)
      if (g_ts_trace[warp][tt][2] != 0)
        printf("TL %d %d %d %lld %lld %lld\n", warp, i, tt, g_ts_trace[warp][tt][0], g_ts_trace[warp][tt][1], g_ts_trace[warp][tt][2]);
  if (threadIdx.x == 0 && blockIdx.x == M.cta0 && n > 4000)
    printf("chase waits (block 0, sweeps 1..6, total cycles): E<-E(i-1,t+1) %llu  E<-D(i-1,t) %llu  E<-D(i,t-2) %llu  D<-E(i,t) %llu  D<-D(i-1,t+1) %llu\n",
           g_ts_wait_cycles[4], g_ts_wait_cycles[5], g_ts_wait_cycles[6], g_ts_wait_cycles[7], g_ts_wait_cycles[9]);
  if (lane == 0 && blockIdx.x == M.cta0 && n > 4000)
    printf("chase timing: warp %2d (%s sweep %d) steps %lld  work %.0f cyc/step  wait %.0f cyc/step  total %.2f ms at 1.9 GHz\n",
           warp, drole ? "D" : "E", i, dbg_steps, (double)dbg_work / (double)max(dbg_steps, 1LL),
           (double)dbg_wait / (double)max(dbg_steps, 1LL), (double)(clock64() - dbg_t0) / 1.9e6);
#endif
}

// ================================================================= back-transformation with the stage-2 reflectors
// ctas[b] = {matrix, column slab, pipeline lane k, pipeline depth NP}: the CTA applies sweep groups
// ngroups-1-k, ngroups-1-k-NP, ... to its 128 columns of Z; group g may work on row tile t once group g+1 (applied
// before it) has written its tiles 0..t.
#ifndef KFAC_TS_Q2_MINB
#define KFAC_TS_Q2_MINB 4   // CTAs per SM the register budget is sized for (measured: 5 costs spills and time)
#endif
__global__ void __launch_bounds__(TS_Q2_T, KFAC_TS_Q2_MINB) ts_q2_kernel(const TsMat* __restrict__ mats, const int4* __restrict__ ctas) {
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
          unsigned spins = 0;
          while (ld_acquire_s32(prog + g + 1) < need) {
            __nanosleep(64);
            if (++spins > (1u << 22)) {   // bounded: a protocol error traps instead of hanging the GPU
              printf("kfac ts_q2: block %d waits for group %d tile %d\n", (int)blockIdx.x, g + 1, need);
              __trap();
            }
          }
        }
      }
      __syncthreads();   // (also: the previous tile's reflectors are no longer being read)
      {
        const float4* src = reinterpret_cast<const float4*>(M.V2 + (size_t)(tile0 + t) * TS_G * TS_B);
        float4* dst = reinterpret_cast<float4*>(&vs[0][0]);
#pragma unroll
        for (int q = 0; q < TS_G * TS_B / 4 / TS_Q2_T; ++q) dst[tid + q * TS_Q2_T] = src[tid + q * TS_Q2_T];
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
          // (splitting the sum into independent partial sums was measured and is slower: the kernel is bound by the
          // traffic of the row window, not by the FMA chain)
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
constexpr int TS_MIN_N = 1000;   // smaller matrices stay on the one-stage kernels (their fixed costs are lower)

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
  size_t gtables_bytes, partials_bytes, active_ints, q2ctas_max, cta_mat_bytes;
};

static void ts_layout_matrix(EighLayout& L, int n, TsSlots& s) {
  const size_t ldx = (size_t)(n + 3) / 4 * 4;
  s.X1 = L.take(2 * TS_B * ldx * 4);
  s.X2 = L.take(2 * TS_B * ldx * 4);
  s.Y0 = L.take((size_t)8 * n * TS_B * 4);   // TS_AV_KS_MAX partial sums
  s.T = L.take(TS_B * TS_B * 4);
  s.G0 = L.take(TS_B * TS_B * 4);
  s.AB = L.take((size_t)(n + 64) * TS_LDAB * 4);   // 64 zero columns beyond the matrix: no bounds checks in the chase
  s.tau2 = L.take((size_t)std::max(ts_total_tiles(n), 1) * TS_G * 4);
  s.goff = L.take((size_t)(ts_ngroups(n) + 2) * 4);
  s.prog = L.take((size_t)n * 8);
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
  sh.active = L.take(sh.active_ints * 4 * 2);   // the lists of active matrices, then their tile prefixes (ts_syr2k)
  sh.cta_mat_bytes = ((size_t)KFAC_MAX_SMS * 4 + cnt * 32) * 4;
  sh.cta_mat = L.take(sh.cta_mat_bytes);
  sh.q2ctas_max = std::max<size_t>(slabs * 16, (size_t)KFAC_MAX_SMS * 8) + 16;
  sh.q2ctas = L.take(sh.q2ctas_max * sizeof(int4));
  sh.gtables_bytes = 4096;   // (the panel steps no longer go through the grouped GEMM: no tables, no split-K scratch)
  sh.gtables = L.take(sh.gtables_bytes);
  sh.partials_bytes = 256;
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
    KFAC_CUDA(cudaFuncSetAttribute(ts_chase_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_WINC * TS_LDAB * 4));
    KFAC_CUDA(cudaFuncSetAttribute(ts_av_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_AV_SMEM));
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
    M.Yp = reinterpret_cast<float*>(base + E.slots.Y0);
    M.T = reinterpret_cast<float*>(base + E.slots.T);
    M.G0 = reinterpret_cast<float*>(base + E.slots.G0);
    M.AB = reinterpret_cast<float*>(base + E.slots.AB);
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
    KFAC_CUDA(cudaMemsetAsync(M.prog, 0, (size_t)E.n * 8, stream));
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
                             float* band_copy = nullptr) {
  if (ents.empty()) return 0;
  if (ensure_gemm_attr()) return 1;
  if (ts_ensure_attrs()) return 1;
  const int cnt = (int)ents.size();
  int steps = 0;
  for (auto& E : ents) steps = std::max(steps, ts_panels(E.n));

  // ---- bulge-chasing CTA ranges (needed in the descriptors before the upload).  One cooperative launch holds as many
  // matrices as fit with their full pipeline depth (n / 512 CTAs each); a call with more matrices than that (BERT-
  // large: 288 factors) chases them in several launches.
  int occ = 0;
  KFAC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ts_chase_kernel, TS_CH_THREADS, TS_WINC * TS_LDAB * 4));
  if (occ < 1) return fail("kfac_eigh_batched: bulge-chasing kernel does not fit on an SM");
  const int cap = std::min(sm_count() * occ, KFAC_MAX_SMS * 4);
  std::vector<int> cta_mat;
  struct ChaseLaunch { int off, ncta; };
  std::vector<ChaseLaunch> chase_launches;
  {
    int k = 0;
    while (k < cnt) {
      const int off = (int)cta_mat.size();
      int used = 0;
      while (k < cnt) {
        // groups in flight at full pipeline depth: (n / 32 steps) / (2 steps per sweep of the group + 2)
        const int want = std::min(cap, std::max(1, (ents[k].n + 64 * TS_CH_SWEEPS - 1) / (64 * TS_CH_SWEEPS)));
        if (used > 0 && used + want > cap) break;
        ents[k].dev.cta0 = used;          // block index inside its own launch
        ents[k].dev.ncta = want;
        for (int i = 0; i < want; ++i) cta_mat.push_back(k);
        used += want;
        ++k;
      }
      chase_launches.push_back({off, used});
    }
  }
  if (cta_mat.size() * 4 > sh.cta_mat_bytes) return fail("kfac_eigh_batched: internal bulge-chasing table overflow");

  std::vector<TsMat> dev(cnt);
  for (int k = 0; k < cnt; ++k) dev[k] = ents[k].dev;
  KFAC_CUDA(cudaMemcpyAsync(base + sh.mats, dev.data(), cnt * sizeof(TsMat), cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(base + sh.cta_mat, cta_mat.data(), cta_mat.size() * 4, cudaMemcpyHostToDevice, stream));
  const TsMat* dmats = reinterpret_cast<const TsMat*>(base + sh.mats);

  // ---- per panel step: who is active (one list, uploaded once)
  std::vector<int> active_all;
  std::vector<int> act_off(steps + 1, 0);
  std::vector<int> max_m4(steps, 0), max_m(steps, 0), row_blocks(steps, 0);
  std::vector<int> tile_off_all, tiles_total(steps, 0);   // per panel step: prefix of the matrices' 64 x 64 upper tiles
  for (int p = 0; p < steps; ++p) {
    const int k0 = p * TS_B;
    act_off[p] = (int)active_all.size();
    for (int k = 0; k < cnt; ++k) {
      if (p >= ts_panels(ents[k].n)) continue;
      active_all.push_back(k);
      const int m = ents[k].n - k0 - TS_B, m4 = ents[k].ldw - k0 - TS_B;
      tile_off_all.push_back(tiles_total[p]);
      tiles_total[p] += ((m + 63) / 64) * ((m + 63) / 64 + 1) / 2;
      max_m4[p] = std::max(max_m4[p], m4);
      max_m[p] = std::max(max_m[p], m);
      row_blocks[p] += (m + 127) / 128;
    }
  }
  act_off[steps] = (int)active_all.size();
  if (active_all.size() > sh.active_ints) return fail("kfac_eigh_batched: internal active-list overflow");
  if (!active_all.empty()) {
    KFAC_CUDA(cudaMemcpyAsync(base + sh.active, active_all.data(), active_all.size() * 4, cudaMemcpyHostToDevice, stream));
    KFAC_CUDA(cudaMemcpyAsync(base + sh.active + sh.active_ints * 4, tile_off_all.data(), tile_off_all.size() * 4,
                              cudaMemcpyHostToDevice, stream));
  }

  for (int p = 0; p < steps; ++p) {
    const int k0 = p * TS_B;
    const int nact = act_off[p + 1] - act_off[p];
    if (nact == 0) continue;
    const int* dact = reinterpret_cast<const int*>(base + sh.active) + act_off[p];
    const int mc = (((max_m4[p] + TS_CL - 1) / TS_CL) + 3) & ~3;
    const size_t smem = (size_t)TS_B * (mc + 4) * sizeof(float);
    if (smem > 216 * 1024) return fail("kfac_eigh_batched: panel of %d positions does not fit the cluster's shared memory", max_m4[p]);
    ts_panel_kernel<<<nact * TS_CL, TS_PT, smem, stream>>>(dmats, dact, k0); KFAC_COUNT_LAUNCH(1);
    // Y0 = A22 V: split the reduction so that about two CTAs per SM are in flight
    // ... and so that no CTA walks more than about TS_AV_SLICE columns: the slices of the largest matrix
    int nslices = std::max(1, std::min(TS_AV_KS_MAX, (2 * sm_count() + row_blocks[p] - 1) / row_blocks[p]));
    nslices = std::max(nslices, std::min(TS_AV_KS_MAX, (max_m4[p] + TS_AV_SLICE - 1) / TS_AV_SLICE));
    const int slice = ((max_m4[p] + nslices - 1) / nslices + TS_AV_KC - 1) / TS_AV_KC * TS_AV_KC;   // columns per slice
    ts_av_kernel<<<dim3((max_m[p] + 127) / 128, nslices, nact), 256, TS_AV_SMEM, stream>>>(dmats, dact, k0, slice); KFAC_COUNT_LAUNCH(1);
    const dim3 grid((max_m[p] + 63) / 64, nact);
    ts_g0_kernel<<<grid, 256, 0, stream>>>(dmats, dact, k0, slice); KFAC_COUNT_LAUNCH(1);
    ts_w_kernel<<<grid, 256, 0, stream>>>(dmats, dact, k0, slice); KFAC_COUNT_LAUNCH(1);
    const int* dtoff = reinterpret_cast<const int*>(base + sh.active + sh.active_ints * 4) + act_off[p];
    if (tiles_total[p] > 0) {
      ts_syr2k_kernel<<<tiles_total[p], 256, 0, stream>>>(dmats, dact, dtoff, nact, k0); KFAC_COUNT_LAUNCH(1);
    }
    KFAC_CUDA(cudaGetLastError());
  }
  int max_n = 0;
  for (auto& E : ents) max_n = std::max(max_n, E.n);
  ts_band_kernel<<<dim3((max_n + 64 + 3) / 4, cnt), 256, 0, stream>>>(dmats); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  if (band_copy != nullptr)   // diagnostics (kfac_debug_two_stage): the band before the bulge chasing overwrites it
    KFAC_CUDA(cudaMemcpyAsync(band_copy, ents[0].dev.AB, (size_t)ents[0].n * TS_LDAB * 4, cudaMemcpyDeviceToDevice, stream));
  if (stage_mark(4, stream)) return 1;
  g_timers.chase_launches = (int)chase_launches.size();
  for (const ChaseLaunch& cl : chase_launches) {
    const int* dcta = reinterpret_cast<const int*>(base + sh.cta_mat) + cl.off;
    void* args[] = {(void*)&dmats, (void*)&dcta};
    KFAC_CUDA(cudaLaunchCooperativeKernel((void*)ts_chase_kernel, dim3((unsigned)cl.ncta), dim3(TS_CH_THREADS), args, TS_WINC * TS_LDAB * 4, stream));
    KFAC_COUNT_LAUNCH(1);
  }
  if (stage_mark(5, stream)) return 1;
  return 0;
}

// Z <- Q2 Z for all two-stage matrices (between the divide & conquer and the compact-WY back-transformation).
// The matrices are taken a few at a time, largest first: a launch whose eigenvector matrices fit the L2 together
// streams its row windows from there instead of from HBM (every sweep group passes over the whole matrix once), and
// the fewer column slabs a launch has, the deeper each slab's pipeline of sweep groups can be.
static int ts_apply_q2(std::vector<TsEntry>& ents, char* base, const TsShared& sh, cudaStream_t stream) {
  if (ents.empty()) return 0;
  const int cnt = (int)ents.size();
  const TsMat* dmats = reinterpret_cast<const TsMat*>(base + sh.mats);
  int occ = 0;
  KFAC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ts_q2_kernel, TS_Q2_T, 0));
  if (occ < 1) return fail("kfac_eigh_batched: back-transformation kernel does not fit on an SM");
  const long long cap = (long long)sm_count() * occ;
  std::vector<int> order(cnt);
  for (int i = 0; i < cnt; ++i) order[i] = i;
  std::sort(order.begin(), order.end(), [&](int a, int b) { return ents[a].n > ents[b].n; });
  const size_t l2_budget = (size_t)96 << 20;
  size_t pos = 0, table_off = 0;
  g_timers.q2_launches = 0;
  while (pos < order.size()) {
    // members of this launch: at least one matrix, then more while the slabs leave a pipeline depth of 8 and the
    // eigenvector matrices stay within the L2 budget
    std::vector<int> members;
    long long slabs = 0;
    size_t bytes = 0;
    int min_groups = 1 << 30;
    while (pos < order.size()) {
      const TsEntry& E = ents[order[pos]];
      const long long sl = (E.n + TS_Q2_T - 1) / TS_Q2_T;
      const size_t zb = (size_t)E.n * E.n * 4;
      if (!members.empty() && ((slabs + sl) * 8 > cap || bytes + zb > l2_budget)) break;
      members.push_back(order[pos]);
      slabs += sl;
      bytes += zb;
      min_groups = std::min(min_groups, std::max(1, ts_ngroups(E.n)));
      ++pos;
    }
    int NP = (int)std::max<long long>(1, std::min<long long>(16, cap / std::max<long long>(slabs, 1)));
    NP = std::min(NP, min_groups);
    std::vector<int4> ctas;
    for (int k = 0; k < NP; ++k)   // pipeline lane 0 (the groups applied first) gets the lowest block indices
      for (int i : members)
        for (int sl = 0; sl < (ents[i].n + TS_Q2_T - 1) / TS_Q2_T; ++sl) ctas.push_back(make_int4(i, sl, k, NP));
    if (table_off + ctas.size() > sh.q2ctas_max) table_off = 0;   // the copies and launches are stream ordered
    if (ctas.size() > sh.q2ctas_max) return fail("kfac_eigh_batched: internal back-transformation table overflow");
    int4* dctas_w = reinterpret_cast<int4*>(base + sh.q2ctas) + table_off;
    KFAC_CUDA(cudaMemcpyAsync(dctas_w, ctas.data(), ctas.size() * sizeof(int4), cudaMemcpyHostToDevice, stream));
    const int4* dctas = dctas_w;
    table_off += ctas.size();
    if (NP > 1) {
      void* args[] = {(void*)&dmats, (void*)&dctas};
      KFAC_CUDA(cudaLaunchCooperativeKernel((void*)ts_q2_kernel, dim3((unsigned)ctas.size()), dim3(TS_Q2_T), args, 0, stream));
    } else {
      ts_q2_kernel<<<(unsigned)ctas.size(), TS_Q2_T, 0, stream>>>(dmats, dctas);
      KFAC_CUDA(cudaGetLastError());
    }
    KFAC_COUNT_LAUNCH(1);
    ++g_timers.q2_launches;
  }
  return stage_mark(6, stream);
}

}  // namespace kfac
