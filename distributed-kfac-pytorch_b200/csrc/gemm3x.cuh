// gemm3x: D = A * B^T on tcgen05 tensor cores with fp32-grade accuracy (3xTF32 split), sm_100a only.
//
// Both operands are *logical* matrices Op[r, k] (r = output row/col index, k = reduction index) that the
// loader warps gather straight from the caller's tensors -- a dense matrix in either orientation, a
// Conv2d input seen through im2col (never materialised), or an (N, C, S) grad-output tensor.  The loader
// splits every fp32 value x into hi = tf32(x) and lo = x - hi and writes both into K-major / 128B-swizzled
// shared-memory tiles; one elected thread then issues  hi*hi + hi*lo + lo*hi  as three kind::tf32 UMMAs
// per K=8 slab, accumulating in TMEM (fp32).  Dropped term lo*lo and the truncation of lo are ~2^-22
// relative, so results match an fp32 GEMM to ~1e-6 (the reference computes these products with fp32
// torch ops: kfac/layers/utils.py:54, kfac/layers/eigen.py:373-384, kfac/layers/inverse.py:233).
//
// Tiling: CTA tile 128 x BN (BN multiple of 32, <= 256), K step 32 (one 128-byte swizzle row), 2-4 smem
// stages, split-K across CTAs with per-tile partials in a caller-provided workspace when the output has
// too few tiles to fill 148 SMs (tall-skinny Gram matrices).  Symmetric outputs (A == B) only visit tiles
// that touch the upper triangle and share one smem tile between the A and B roles on the diagonal.
#pragma once
#include <cuda_fp16.h>
#include "ptx.cuh"

namespace kfac {

enum OpMode : int { OP_KCONTIG = 0, OP_RCONTIG = 1, OP_IM2COL = 2, OP_NCS = 3 };
// How the loader warps fetch an operand tile (chosen on the host from alignment / geometry):
//   *_SCALAR: one fp32 per lane; *_VEC: one float4 per lane (16-byte aligned, 4 | K resp. 4 | ld)
enum LoadKind : int {
  LK_KC_SCALAR = 0, LK_KC_VEC = 1, LK_RC_SCALAR = 2, LK_RC_VEC = 3, LK_IM2COL = 4, LK_NCS_SCALAR = 5, LK_NCS_VEC = 6
};
enum EpiMode : int { EPI_PARTIAL = 0, EPI_AXPBY = 1, EPI_DIV_OUTER = 2, EPI_MUL_MAT = 3 };

struct Operand {
  const float* p;
  int mode;
  int lk;         // LoadKind, filled by the launcher
  int elt;        // element type behind p: 0 = fp32, 1 = fp16, 2 = bf16 (activations / grad-outputs under autocast;
                  // both are exact in tf32, so their lo parts are zero and the result is the fp32 product of the values)
  int R;          // logical rows, including the ones-row if present
  int ones_row;   // index of a row that is identically 1 (bias column of the A factor), or -1
  long long ld;   // OP_KCONTIG: Op[r,k] = p[r*ld + k];  OP_RCONTIG: Op[r,k] = p[k*ld + r]
  const float* ext;  // OP_KCONTIG only: Op[r, ext_k] = ext[r] (bias.grad as the last column of [Wg | bg]), or NULL
  int ext_k;
  // OP_IM2COL: p is (N, C, H, W); r = (c*KH + i)*KW + j (F.unfold order, modules.py:228-236);
  //            k = (n*OH + oh)*OW + ow; Op = p[n, c, oh*SH - PH + i, ow*SW - PW + j] or 0 outside.
  int C, H, W, KH, KW, SH, SW, PH, PW, OH, OW;
  // OP_NCS: p is (N, C, S); r = c; k = n*S + s; Op = p[n, c, s]   (grad-output of a Conv2d, modules.py:188-190)
  int S;
};

struct GemmParams {
  Operand a, b;
  int same_ab;  // a and b describe the same matrix (Gram): enables smem tile sharing
  int sym;      // output symmetric: only upper tiles are computed, the lower triangle is mirrored
  int M, N, K;  // output M x N (= a.R x b.R), reduction length K
  int BN;       // tile width
  int tiles_m, tiles_n, num_tiles;
  int splits, kb_per_split, stages;
  int promote_kb;  // k-blocks accumulated in TMEM between drains into fp32 registers (1..8)
  // epilogue
  int epi;
  float alpha, beta;
  float* out;
  long long ldo;
  float* ws;  // split-K partials: [(tile*splits + split)][128][BN]
  const float* e0;  // EPI_DIV_OUTER: row vector (len M);  EPI_MUL_MAT: matrix M x N (lde)
  const float* e1;  // EPI_DIV_OUTER: col vector (len N)
  long long lde;
  float damping;
  const float* dotm;  // optional: accumulate sum(out * dotm[r*ldd + c]) into *dot_out
  long long ldd;
  // dotm may be split as [dotm (M x Nw) | dotb (M)] to read weight.grad / bias.grad in place (modules.py:63-69)
  const float* dotb;
  int dot_nw;
  double* dot_out;
  // grouped launches only: if non-null and *skip != 0 when the CTA starts, the problem is left untouched (an iteration
  // whose matrix has already converged: batched Newton-Schulz decides that on the device, without a host round trip)
  const int* skip;
};

constexpr int BM = 128;
constexpr int BK = 32;
constexpr int NTHREADS = 512;
constexpr int NWARPS = 16;
constexpr int COL_PARTS = NWARPS / 4;  // epilogue: warp w owns TMEM lane quarter (w & 3) and column part (w >> 2)
constexpr int TILE_A_BYTES = BM * 128;
constexpr int MAX_STAGES = 4;
constexpr int NUM_BARS = 2 * MAX_STAGES + 2;  // empty[4] (count 1), full[4] (count NTHREADS), done[2] (count 1)
constexpr int MAX_BN = 256;

__host__ __device__ inline int stage_bytes(int BN) { return 2 * TILE_A_BYTES + 2 * BN * 128; }

// number of 128-row blocks that touch the upper triangle of column block J
__host__ __device__ inline int sym_rows_for_col(int N, int BN, int tiles_m, int J) {
  int last = BN * (J + 1) - 1;
  if (last > N - 1) last = N - 1;
  int cnt = last / BM + 1;
  return cnt < tiles_m ? cnt : tiles_m;
}
__host__ __device__ inline void tile_to_ij(int sym, int N, int BN, int tiles_m, int tiles_n, int t, int& I, int& J) {
  if (!sym) {
    I = t / tiles_n;
    J = t - I * tiles_n;
    return;
  }
  for (J = 0; J < tiles_n; ++J) {
    int cnt = sym_rows_for_col(N, BN, tiles_m, J);
    if (t < cnt) {
      I = t;
      return;
    }
    t -= cnt;
  }
  I = 0;
  J = 0;
}

__device__ __forceinline__ uint32_t sw128_offset(int row, int kk) {
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((kk >> 2) ^ (row & 7))) << 4) + ((kk & 3) << 2));
}
__device__ __forceinline__ void store_split(uint8_t* hi, uint8_t* lo, uint32_t off, float v) {
  float h = to_tf32(v);
  *reinterpret_cast<float*>(hi + off) = h;
  *reinterpret_cast<float*>(lo + off) = v - h;
}

// ---------------------------------------------------------------------------------------------------------------
// Operand tile loaders.  A tile is nrows x 32 (k) fp32; every thread owns NS values of it (NS = 16 for the 128-row A
// tile, 32 for a 256-row B tile) held in registers between `tile_issue` (global loads, all independent, issued
// back to back so ~48 KB per SM are in flight) and `tile_store` (hi/lo split + swizzled smem writes).
// Thread -> (row, k) maps are chosen so that the global read is sector-efficient AND the smem write is free of
// bank conflicts in the K-major 128B-swizzled layout (bank = 4 * (chunk ^ (row & 7)) + (k & 3)):
//   KC scalar : lane = k, warp = row mod 8                      (128 B global row segments, one smem row per warp)
//   KC vec    : 8 lanes x float4 cover one row's 128 B, a warp covers 4 rows (STS.128)
//   RC scalar : lane = (row & 7) + 8 * (k & 3): 8 rows x 4 k    (32 B global segments, 32 distinct banks)
//   RC vec    : lane = (row quad & 1) + 2 * (k & 15): float4 along rows, 16 k     (32 distinct banks per store)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t sw128_chunk_offset(int row, int chunk) {
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}

// one element of an operand tensor, by element offset (p is a float pointer only nominally when elt != 0)
__device__ __forceinline__ float ld_elt(const float* p, long long off, int elt) {
  if (elt == 0) return __ldg(p + off);
  const unsigned short u = __ldg(reinterpret_cast<const unsigned short*>(p) + off);
  return elt == 2 ? __uint_as_float((uint32_t)u << 16) : __half2float(__ushort_as_half(u));
}

template <int LK, int NS, bool HALF>
__device__ __forceinline__ void tile_issue(const Operand& op, int row0, int nrows, int k0, int kend, float (&v)[NS],
                                           int tid, const int2* __restrict__ tab) {
  const float* __restrict__ p = op.p;
  const int lane = tid & 31, warp = tid >> 5;
  if (LK == LK_KC_SCALAR || LK == LK_IM2COL || LK == LK_NCS_SCALAR) {
    const int k = k0 + lane;
    const bool kin = k < kend;
    long long base = k;
    int ihb = 0, iwb = 0;
    if (LK == LK_IM2COL) {
      const int S = op.OH * op.OW;
      const int n = k / S;
      const int rem = k - n * S;
      const int oh = rem / op.OW;
      const int ow = rem - oh * op.OW;
      ihb = oh * op.SH - op.PH;
      iwb = ow * op.SW - op.PW;
      base = (long long)n * op.C * op.H * op.W + (long long)ihb * op.W + iwb;
    } else if (LK == LK_NCS_SCALAR) {
      const int n = k / op.S;
      base = (long long)n * op.C * op.S + (k - n * op.S);
    }
    // Per-lane invariants of this k-block.  Under the kernel's register pressure the compiler otherwise
    // re-derives `base` (64-bit multiplies) for every element; the empty asm makes the values opaque.
    const int elt = op.elt;
    const float* pk = p + base;   // fp32 path: element pointer; 16-bit path: element offsets from p
    if (HALF)
      asm volatile("" : "+l"(base), "+r"(ihb), "+r"(iwb));
    else
      asm volatile("" : "+l"(pk), "+r"(ihb), "+r"(iwb));
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const int rr = warp + NWARPS * j;
      const int r = row0 + rr;
      float x = 0.f;
      if (LK == LK_IM2COL) {
        // (c, i, j) of the row were decoded once per CTA into `tab` (the tile's rows never change across k-blocks):
        // x = tap offset (c*H + i)*W + j, or -1 (row beyond R) / -2 (ones row); y = (i << 16) | j
        if (rr < nrows && kin) {
          const int2 t = tab[rr];
          if (t.x == -2) {
            x = 1.f;
          } else if (t.x >= 0) {
            const int ih = ihb + (t.y >> 16), iw = iwb + (t.y & 0xffff);
            if ((unsigned)ih < (unsigned)op.H && (unsigned)iw < (unsigned)op.W) x = HALF ? ld_elt(p, base + t.x, elt) : __ldg(pk + t.x);
          }
        }
      } else if (rr < nrows && kin && r < op.R) {
        if (r == op.ones_row) {
          x = 1.f;
        } else if (LK == LK_KC_SCALAR) {
          x = (op.ext != nullptr && k == op.ext_k) ? __ldg(op.ext + r)
                                                   : (HALF ? ld_elt(p, base + (long long)r * op.ld, elt)
                                                           : __ldg(pk + (long long)r * op.ld));
        } else {
          x = HALF ? ld_elt(p, base + (long long)r * op.S, elt) : __ldg(pk + (long long)r * op.S);
        }
      }
      v[j] = x;
    }
  } else if (LK == LK_KC_VEC || LK == LK_NCS_VEC) {
#pragma unroll
    for (int q = 0; q < NS / 4; ++q) {
      const int c = tid + NTHREADS * q;
      const int rr = c >> 3, kc = c & 7;
      const int r = row0 + rr;
      const int k = k0 + 4 * kc;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (rr < nrows && r < op.R && k < kend) {
        if (r == op.ones_row) {
          x = make_float4(1.f, 1.f, 1.f, 1.f);
        } else if (LK == LK_KC_VEC) {
          x = __ldg(reinterpret_cast<const float4*>(p + (long long)r * op.ld + k));
        } else {
          const int n = k / op.S;
          x = __ldg(reinterpret_cast<const float4*>(p + ((long long)n * op.C + r) * op.S + (k - n * op.S)));
        }
      }
      v[4 * q] = x.x; v[4 * q + 1] = x.y; v[4 * q + 2] = x.z; v[4 * q + 3] = x.w;
    }
  } else if (LK == LK_RC_SCALAR) {
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const int u = warp + NWARPS * j;
      const int rr = ((u >> 3) << 3) + (lane & 7);
      const int k = k0 + ((u & 7) << 2) + (lane >> 3);
      const int r = row0 + rr;
      float x = 0.f;
      if (rr < nrows && k < kend && r < op.R) x = (r == op.ones_row) ? 1.f
                                                                      : (HALF ? ld_elt(p, (long long)k * op.ld + r, op.elt)
                                                                              : __ldg(p + (long long)k * op.ld + r));
      v[j] = x;
    }
  } else {  // LK_RC_VEC
#pragma unroll
    for (int q = 0; q < NS / 4; ++q) {
      const int u = warp + NWARPS * q;
      const int rr = ((u >> 1) << 3) + ((lane & 1) << 2);
      const int k = k0 + ((u & 1) << 4) + (lane >> 1);
      const int r = row0 + rr;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (rr < nrows && k < kend && r < op.R) {
        const float* src = p + (long long)k * op.ld + r;
        if (r + 3 < op.R && (op.ones_row < r || op.ones_row > r + 3)) {
          x = __ldg(reinterpret_cast<const float4*>(src));
        } else {
          float t[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) t[i] = (r + i < op.R) ? ((r + i == op.ones_row) ? 1.f : __ldg(src + i)) : 0.f;
          x = make_float4(t[0], t[1], t[2], t[3]);
        }
      }
      v[4 * q] = x.x; v[4 * q + 1] = x.y; v[4 * q + 2] = x.z; v[4 * q + 3] = x.w;
    }
  }
}

template <int LK, int NS>
__device__ __forceinline__ void tile_store(int nrows, const float (&v)[NS], uint8_t* hi, uint8_t* lo, int tid) {
  const int lane = tid & 31, warp = tid >> 5;
  if (LK == LK_KC_SCALAR || LK == LK_IM2COL || LK == LK_NCS_SCALAR) {
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const int rr = warp + NWARPS * j;
      if (rr < nrows) store_split(hi, lo, sw128_offset(rr, lane), v[j]);
    }
  } else if (LK == LK_KC_VEC || LK == LK_NCS_VEC) {
#pragma unroll
    for (int q = 0; q < NS / 4; ++q) {
      const int c = tid + NTHREADS * q;
      const int rr = c >> 3, kc = c & 7;
      if (rr < nrows) {
        const uint32_t off = sw128_chunk_offset(rr, kc);
        const float h0 = to_tf32(v[4 * q]), h1 = to_tf32(v[4 * q + 1]), h2 = to_tf32(v[4 * q + 2]),
                    h3 = to_tf32(v[4 * q + 3]);
        *reinterpret_cast<float4*>(hi + off) = make_float4(h0, h1, h2, h3);
        *reinterpret_cast<float4*>(lo + off) =
            make_float4(v[4 * q] - h0, v[4 * q + 1] - h1, v[4 * q + 2] - h2, v[4 * q + 3] - h3);
      }
    }
  } else if (LK == LK_RC_SCALAR) {
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const int u = warp + NWARPS * j;
      const int rr = ((u >> 3) << 3) + (lane & 7);
      const int kk = ((u & 7) << 2) + (lane >> 3);
      if (rr < nrows) store_split(hi, lo, sw128_offset(rr, kk), v[j]);
    }
  } else {  // LK_RC_VEC
#pragma unroll
    for (int q = 0; q < NS / 4; ++q) {
      const int u = warp + NWARPS * q;
      const int rr = ((u >> 1) << 3) + ((lane & 1) << 2);
      const int kk = ((u & 1) << 4) + (lane >> 1);
      if (rr < nrows) {
#pragma unroll
        for (int i = 0; i < 4; ++i) store_split(hi, lo, sw128_offset(rr + i, kk), v[4 * q + i]);
      }
    }
  }
}

template <int NS, bool HALF>
__device__ __forceinline__ void issue_any(const Operand& op, int row0, int nrows, int k0, int kend, float (&v)[NS],
                                          int tid, const int2* __restrict__ tab) {
  if (HALF && op.elt != 0) {   // fp16 / bf16 tensors take the scalar loaders (the launcher never picks a vector kind for them)
    switch (op.lk) {
      case LK_RC_SCALAR: tile_issue<LK_RC_SCALAR, NS, true>(op, row0, nrows, k0, kend, v, tid, tab); break;
      case LK_IM2COL: tile_issue<LK_IM2COL, NS, true>(op, row0, nrows, k0, kend, v, tid, tab); break;
      case LK_NCS_SCALAR: tile_issue<LK_NCS_SCALAR, NS, true>(op, row0, nrows, k0, kend, v, tid, tab); break;
      default: tile_issue<LK_KC_SCALAR, NS, true>(op, row0, nrows, k0, kend, v, tid, tab); break;
    }
    return;
  }
  switch (op.lk) {
    case LK_KC_VEC: tile_issue<LK_KC_VEC, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
    case LK_RC_SCALAR: tile_issue<LK_RC_SCALAR, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
    case LK_RC_VEC: tile_issue<LK_RC_VEC, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
    case LK_IM2COL: tile_issue<LK_IM2COL, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
    case LK_NCS_SCALAR: tile_issue<LK_NCS_SCALAR, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
    case LK_NCS_VEC: tile_issue<LK_NCS_VEC, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
    default: tile_issue<LK_KC_SCALAR, NS, false>(op, row0, nrows, k0, kend, v, tid, tab); break;
  }
}
template <int NS>
__device__ __forceinline__ void store_any(int lk, int nrows, const float (&v)[NS], uint8_t* hi, uint8_t* lo, int tid) {
  switch (lk) {
    case LK_KC_VEC:
    case LK_NCS_VEC: tile_store<LK_KC_VEC, NS>(nrows, v, hi, lo, tid); break;
    case LK_RC_SCALAR: tile_store<LK_RC_SCALAR, NS>(nrows, v, hi, lo, tid); break;
    case LK_RC_VEC: tile_store<LK_RC_VEC, NS>(nrows, v, hi, lo, tid); break;
    default: tile_store<LK_KC_SCALAR, NS>(nrows, v, hi, lo, tid); break;
  }
}

__device__ __forceinline__ float dot_operand(const GemmParams& p, int r, int c) {
  if (p.dotb != nullptr && c >= p.dot_nw) return __ldg(p.dotb + r);
  return __ldg(p.dotm + (long long)r * p.ldd + c);
}

// value -> output element, shared by the direct epilogue and the split-K finalize kernel
__device__ __forceinline__ float epi_apply(const GemmParams& p, float acc, int r, int c) {
  float v = p.alpha * acc;
  if (p.epi == EPI_DIV_OUTER) {
    v = v / (__ldg(p.e0 + r) * __ldg(p.e1 + c) + p.damping);
  } else if (p.epi == EPI_MUL_MAT) {
    v = v * __ldg(p.e0 + (long long)r * p.lde + c);
  }
  if (p.beta != 0.f) v += p.beta * p.out[(long long)r * p.ldo + c];
  return v;
}

// Per-CTA decode of im2col rows: the (c, i, j) of a tile row is the same for every k-block, so the two integer
// divisions per row are done once here instead of once per element.
__device__ __forceinline__ void fill_im2col_table(const Operand& op, int row0, int nrows, int2* tab, int tid) {
  const int khw = op.KH * op.KW;
  for (int rr = tid; rr < nrows; rr += NTHREADS) {
    const int r = row0 + rr;
    int2 t = make_int2(-1, 0);
    if (r < op.R) {
      if (r == op.ones_row) {
        t.x = -2;
      } else {
        const int c = r / khw;
        const int rem = r - c * khw;
        const int i = rem / op.KW;
        const int j = rem - i * op.KW;
        t = make_int2((c * op.H + i) * op.W + j, (i << 16) | j);
      }
    }
    tab[rr] = t;
  }
}

template <bool HALF>
__device__ __forceinline__ void gemm3x_body(const GemmParams& p, const int local_block, uint8_t* smem_raw,
                                            uint64_t* bars, uint32_t* tmem_base_slot, double* dot_red, int2* rowtab) {
  // SWIZZLE_128B tiles need 1024-byte alignment (swizzle is a function of absolute address bits 7..9)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int BN = p.BN;
  const int sbytes = stage_bytes(BN);
  const bool single_pass = HALF && p.a.elt != 0 && p.b.elt != 0;   // 16-bit values on both sides: exact in tf32

  const int tile = local_block / p.splits;
  const int split = local_block - tile * p.splits;
  int I, J;
  tile_to_ij(p.sym, p.N, BN, p.tiles_m, p.tiles_n, tile, I, J);
  const int arow0 = I * BM, bcol0 = J * BN;

  // operand tile sharing for Gram matrices
  const bool a_in_b = p.same_ab && arow0 >= bcol0 && arow0 + BM <= bcol0 + BN;
  const bool b_in_a = p.same_ab && !a_in_b && bcol0 >= arow0 && bcol0 + BN <= arow0 + BM;

  const int nkb_total = (p.K + BK - 1) / BK;
  const int kb_begin = split * p.kb_per_split;
  int kb_end = kb_begin + p.kb_per_split;
  if (kb_end > nkb_total) kb_end = nkb_total;
  const int nkb = kb_end - kb_begin;
  const int kend = (kb_end * BK < p.K) ? kb_end * BK : p.K;

  // two TMEM accumulator buffers (<= 2 x 256 columns): chunk c of the K loop accumulates into buffer c & 1 while
  // buffer (c - 1) & 1 is drained into registers
  uint32_t buf_cols = 32;
  while (buf_cols < (uint32_t)BN) buf_cols <<= 1;
  const uint32_t tmem_cols = 2 * buf_cols;

  if (tid == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) {
      mbar_init(smem_u32(&bars[s]), 1);
      mbar_init(smem_u32(&bars[MAX_STAGES + s]), NTHREADS);
    }
    mbar_init(smem_u32(&bars[2 * MAX_STAGES]), 1);
    mbar_init(smem_u32(&bars[2 * MAX_STAGES + 1]), 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(tmem_base_slot), tmem_cols);
  if (p.a.lk == LK_IM2COL && !a_in_b) fill_im2col_table(p.a, arow0, BM, rowtab, tid);
  if (p.b.lk == LK_IM2COL && !b_in_a) fill_im2col_table(p.b, bcol0, BN, rowtab + BM, tid);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_d = *tmem_base_slot;
  const uint32_t idesc = make_idesc_tf32(BM, BN);

  // fp32 register accumulators: thread owns tile row (32q + lane), columns [h*BN/2, (h+1)*BN/2).
  // The tensor core adds into TMEM with truncation, which biases long chains by ~3e-8 per UMMA (measured);
  // so TMEM only ever holds promote_kb k-blocks (12 UMMAs each) before it is drained into these registers
  // with round-to-nearest adds.
  const int q = warp & 3, h = warp >> 2;
  const int half = BN / COL_PARTS;   // columns per thread (multiple of 8)
  constexpr int ACC = MAX_BN / COL_PARTS;
  float acc[ACC];
#pragma unroll
  for (int i = 0; i < ACC; ++i) acc[i] = 0.f;
  // drain one TMEM accumulator buffer into the register accumulators (round-to-nearest adds)
  auto drain = [&](uint32_t buf) {
    tc_fence_after_sync();
#pragma unroll
    for (int ci = 0; ci < ACC / 8; ++ci) {
      if (ci * 8 < half) {
        uint32_t v[8];
        tmem_ld8(tmem_d + buf * buf_cols + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * half + ci * 8), v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[ci * 8 + i] += __uint_as_float(v[i]);
      }
    }
    tc_fence_before_sync();
  };

  // operand registers of the k-block in flight
  float va[BM * BK / NTHREADS];
  float vb[MAX_BN * BK / NTHREADS];
  if (nkb > 0) {
    if (!a_in_b) issue_any<BM * BK / NTHREADS, HALF>(p.a, arow0, BM, kb_begin * BK, kend, va, tid, rowtab);
    if (!b_in_a) issue_any<MAX_BN * BK / NTHREADS, HALF>(p.b, bcol0, BN, kb_begin * BK, kend, vb, tid, rowtab + BM);
  }

  int s = 0;                 // smem stage of this k-block
  uint32_t sphase = 0;       // bit s: parity of the current use of stage s (same for its full and empty barrier)
  uint32_t used = 0;         // bit s: stage s has been used before
  int in_chunk = 0;          // k-blocks already accumulated into the current TMEM buffer
  uint32_t buf = 0;          // TMEM buffer of the current chunk
  uint32_t done_uses = 0;    // bit b: parity of the pending commit on done[b]
  int chunks_done = 0;       // chunks whose UMMAs have been issued completely
  const int promote_kb = p.promote_kb, nstages = p.stages;

  for (int kb = 0; kb < nkb; ++kb) {
    const uint32_t empty_bar = smem_u32(&bars[s]), full_bar = smem_u32(&bars[MAX_STAGES + s]);
    const uint32_t par = (sphase >> s) & 1u;
    if ((used >> s) & 1u) mbar_wait(empty_bar, par ^ 1u);  // UMMAs that read this stage's previous contents retired
    uint8_t* st = smem + (size_t)s * sbytes;
    uint8_t* a_hi = st;
    uint8_t* a_lo = st + TILE_A_BYTES;
    uint8_t* b_hi = st + 2 * TILE_A_BYTES;
    uint8_t* b_lo = b_hi + BN * 128;
    if (!a_in_b) store_any<BM * BK / NTHREADS>(p.a.lk, BM, va, a_hi, a_lo, tid);
    if (!b_in_a) store_any<MAX_BN * BK / NTHREADS>(p.b.lk, BN, vb, b_hi, b_lo, tid);
    fence_proxy_async_smem();
    mbar_arrive(full_bar);   // hand the stage over: only the issuing thread waits, everyone else moves on
    const bool last_of_chunk = (in_chunk == promote_kb - 1) || kb == nkb - 1;
    if (tid == 0) {
      mbar_wait(full_bar, par);
      tc_fence_after_sync();
      uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
      if (a_in_b) {
        ah = bh + (uint32_t)(arow0 - bcol0) * 128;
        al = bl + (uint32_t)(arow0 - bcol0) * 128;
      } else if (b_in_a) {
        bh = ah + (uint32_t)(bcol0 - arow0) * 128;
        bl = al + (uint32_t)(bcol0 - arow0) * 128;
      }
      const uint32_t td = tmem_d + buf * buf_cols;
#pragma unroll
      for (int ks = 0; ks < BK / 8; ++ks) {
        const uint64_t dah = make_kmajor_sw128_desc(ah + ks * 32), dal = make_kmajor_sw128_desc(al + ks * 32);
        const uint64_t dbh = make_kmajor_sw128_desc(bh + ks * 32), dbl = make_kmajor_sw128_desc(bl + ks * 32);
        if (HALF && single_pass) {
          // both operands are fp16 / bf16 values: exact in tf32, their lo parts are zero, so one UMMA per slab IS the
          // exact product (the low-precision factor mode: a third of the tensor work of the fp32-grade path)
          mma_tf32_ss(td, dah, dbh, idesc, !(in_chunk == 0 && ks == 0));
        } else {
          mma_tf32_ss(td, dal, dbh, idesc, !(in_chunk == 0 && ks == 0));  // first UMMA of a chunk overwrites
          mma_tf32_ss(td, dah, dbl, idesc, 1);
          mma_tf32_ss(td, dah, dbh, idesc, 1);
        }
      }
      mma_commit(empty_bar);
      if (last_of_chunk) mma_commit(smem_u32(&bars[2 * MAX_STAGES + buf]));
    }
    // next k-block's global loads go out now: they are in flight while the tensor core works and while the
    // previous chunk's accumulators are drained
    if (kb + 1 < nkb) {
      const int k0n = (kb_begin + kb + 1) * BK;
      if (!a_in_b) issue_any<BM * BK / NTHREADS, HALF>(p.a, arow0, BM, k0n, kend, va, tid, rowtab);
      if (!b_in_a) issue_any<MAX_BN * BK / NTHREADS, HALF>(p.b, bcol0, BN, k0n, kend, vb, tid, rowtab + BM);
    }
    used |= 1u << s;
    sphase ^= 1u << s;
    if (++s == nstages) s = 0;
    if (last_of_chunk) {
      // drain the chunk BEFORE this one: its commit was issued a whole chunk ago, so there is (almost) no wait,
      // and the tensor core keeps running this chunk's UMMAs meanwhile.  The buffer is re-used by the next chunk,
      // whose first UMMA is issued only after every thread has arrived on that k-block's full barrier, i.e. after
      // every thread has finished this drain.
      if (chunks_done > 0) {
        const uint32_t pb = buf ^ 1u;
        mbar_wait(smem_u32(&bars[2 * MAX_STAGES + pb]), (done_uses >> pb) & 1u);
        done_uses ^= 1u << pb;
        drain(pb);
      }
      ++chunks_done;
      buf ^= 1u;
      in_chunk = 0;
    } else {
      ++in_chunk;
    }
  }
  if (chunks_done > 0) {  // the last chunk
    const uint32_t pb = buf ^ 1u;
    mbar_wait(smem_u32(&bars[2 * MAX_STAGES + pb]), (done_uses >> pb) & 1u);
    drain(pb);
  }

  // ---- epilogue: registers -> shared memory (the operand stages are idle now: every UMMA has retired) ->
  // coalesced 16-byte global accesses, one tile row per warp iteration.  Per-thread scattered 4-byte stores cost
  // ~130 us per tile (32768 partial-sector writes); staged rows cost a few us.
  __syncthreads();  // all warps are past their last smem / TMEM use of the main loop
  float* stg = reinterpret_cast<float*>(smem);
  const int pitch = BN + 4;  // 16-byte aligned rows; 4-word skew makes the float4 writes below conflict-free
  {
    float* dst = stg + (size_t)(q * 32 + lane) * pitch + h * half;
#pragma unroll
    for (int ci = 0; ci < ACC / 4; ++ci) {
      if (ci * 4 < half)
        *reinterpret_cast<float4*>(dst + ci * 4) =
            make_float4(acc[ci * 4], acc[ci * 4 + 1], acc[ci * 4 + 2], acc[ci * 4 + 3]);
    }
  }
  __syncthreads();
  float dot_local = 0.f;
  // copies of the fields used per element (p may live in shared memory: avoid reloading them after every store)
  const int epi = p.epi, sym = p.sym, Mrows = p.M, Ncols = p.N;
  const float alpha = p.alpha, beta = p.beta, damping = p.damping;
  float* __restrict__ out = p.out;
  const long long ldo = p.ldo, lde = p.lde, ldd = p.ldd;
  const float* __restrict__ e0 = p.e0;
  const float* __restrict__ e1 = p.e1;
  const float* __restrict__ dotm = p.dotm;
  const float* __restrict__ dotb = p.dotb;
  const int dot_nw = p.dot_nw;
  const bool want_dot = p.dot_out != nullptr;
  if (epi == EPI_PARTIAL) {
    float* wsb = p.ws + (size_t)local_block * BM * BN;
    for (int rr = warp; rr < BM; rr += NWARPS)
      for (int c = lane * 4; c < BN; c += 128)
        *reinterpret_cast<float4*>(wsb + (size_t)rr * BN + c) = *reinterpret_cast<const float4*>(stg + (size_t)rr * pitch + c);
  } else {
    const bool vec_ok = ((ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) && ((bcol0 & 3) == 0);
    for (int rr = warp; rr < BM; rr += NWARPS) {
      const int grow = arow0 + rr;
      if (grow >= Mrows) break;
      float* srow = stg + (size_t)rr * pitch;
      float* orow = out + (long long)grow * ldo;
      const float e0r = (epi == EPI_DIV_OUTER) ? __ldg(e0 + grow) : 0.f;
      for (int c = lane * 4; c < BN; c += 128) {
        const int gcol = bcol0 + c;
        if (gcol >= Ncols) continue;
        if (sym && grow > gcol + 3) continue;
        float4 a4 = *reinterpret_cast<const float4*>(srow + c);
        float a[4] = {a4.x, a4.y, a4.z, a4.w};
        const bool full = vec_ok && gcol + 3 < Ncols && !(sym && grow > gcol);
        float old[4] = {0.f, 0.f, 0.f, 0.f};
        if (beta != 0.f) {
          if (full) {
            const float4 o4 = *reinterpret_cast<const float4*>(orow + gcol);
            old[0] = o4.x; old[1] = o4.y; old[2] = o4.z; old[3] = o4.w;
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (gcol + i < Ncols && !(sym && grow > gcol + i)) old[i] = orow[gcol + i];
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int gc = gcol + i;
          float v = alpha * a[i];
          if (gc < Ncols) {
            if (epi == EPI_DIV_OUTER) v = v / (e0r * __ldg(e1 + gc) + damping);
            else if (epi == EPI_MUL_MAT) v = v * __ldg(e0 + (long long)grow * lde + gc);
          }
          a[i] = v + beta * old[i];
        }
        if (full) {
          *reinterpret_cast<float4*>(orow + gcol) = make_float4(a[0], a[1], a[2], a[3]);
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (gcol + i < Ncols && !(sym && grow > gcol + i)) orow[gcol + i] = a[i];
        }
        if (want_dot) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int gc = gcol + i;
            if (gc < Ncols) {
              const float dv = (dotb != nullptr && gc >= dot_nw) ? __ldg(dotb + grow) : __ldg(dotm + (long long)grow * ldd + gc);
              dot_local += a[i] * dv;
            }
          }
        }
        if (sym) *reinterpret_cast<float4*>(srow + c) = make_float4(a[0], a[1], a[2], a[3]);  // final values for the mirror
      }
    }
    if (sym) {
      // mirror: out[gcol, grow] = out[grow, gcol] for grow < gcol, written row-wise (coalesced along grow)
      __syncthreads();
      for (int c = warp; c < BN; c += NWARPS) {
        const int gcol = bcol0 + c;
        if (gcol >= Ncols) break;
        float* mrow = out + (long long)gcol * ldo + arow0;
        for (int rr = lane; rr < BM; rr += 32) {
          const int grow = arow0 + rr;
          if (grow < Mrows && grow < gcol) mrow[rr] = stg[(size_t)rr * pitch + c];
        }
      }
    }
  }
  if (want_dot && epi != EPI_PARTIAL) {
    double d = (double)dot_local;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) dot_red[warp] = d;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tmem_dealloc(tmem_d, tmem_cols);
    if (want_dot && epi != EPI_PARTIAL && lane == 0) {
      double d = 0.0;
      for (int w = 0; w < NWARPS; ++w) d += dot_red[w];
      atomicAdd(p.dot_out, d);
    }
  }
}

// HALF: the instantiation that also carries the fp16 / bf16 operand loaders (factor kernels under autocast); the
// fp32 instantiation stays as small as it was
template <bool HALF>
__global__ void __launch_bounds__(NTHREADS, 1) gemm3x_kernel(const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[NUM_BARS];
  __shared__ uint32_t tmem_base_slot;
  __shared__ double dot_red[NWARPS];
  __shared__ int2 rowtab[BM + MAX_BN];
  gemm3x_body<HALF>(p, (int)blockIdx.x, smem_raw, bars, &tmem_base_slot, dot_red, rowtab);
}

// Grouped launch: one kernel, many independent problems (all merges of a divide & conquer level, one compact-WY step
// of every matrix, one GEMM stage of every layer's precondition).  cta_end[i] = exclusive prefix sum of CTA counts.
__device__ __forceinline__ int find_problem(const int* __restrict__ ends, int nprob, int block) {
  int a = 0, b = nprob - 1;
  while (a < b) {
    const int c = (a + b) >> 1;
    if (block < ends[c]) b = c; else a = c + 1;
  }
  return a;
}

__global__ void __launch_bounds__(NTHREADS, 1) gemm3x_group_kernel(const GemmParams* __restrict__ probs,
                                                                   const int* __restrict__ cta_end, int nprob) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[NUM_BARS];
  __shared__ uint32_t tmem_base_slot;
  __shared__ double dot_red[NWARPS];
  __shared__ __align__(16) GemmParams sp;
  __shared__ int2 rowtab[BM + MAX_BN];
  const int idx = find_problem(cta_end, nprob, (int)blockIdx.x);
  const int first = idx > 0 ? cta_end[idx - 1] : 0;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(probs + idx);
  for (int i = threadIdx.x; i < (int)(sizeof(GemmParams) / 4); i += NTHREADS) reinterpret_cast<uint32_t*>(&sp)[i] = src[i];
  __syncthreads();
  if (sp.skip != nullptr && *reinterpret_cast<const volatile int*>(sp.skip) != 0) return;   // uniform over the CTA
  gemm3x_body<false>(sp, (int)blockIdx.x - first, smem_raw, bars, &tmem_base_slot, dot_red, rowtab);
}

// Sum split-K partials and apply the AXPBY epilogue (+ symmetric mirror). One thread per 4 columns.
__device__ __forceinline__ void finalize_body(const GemmParams& p, long long gid) {
  const int BN = p.BN;
  const int per_tile = BM * BN / 4;
  const int tile = (int)(gid / per_tile);
  if (tile >= p.num_tiles) return;
  const int e = (int)(gid - (long long)tile * per_tile);
  const int row = e / (BN / 4);
  const int c4 = (e - row * (BN / 4)) * 4;
  int I, J;
  tile_to_ij(p.sym, p.N, BN, p.tiles_m, p.tiles_n, tile, I, J);
  const int grow = I * BM + row;
  const int gcol0 = J * BN + c4;
  if (grow >= p.M || gcol0 >= p.N) return;
  if (p.sym && grow > gcol0 + 3) return;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const float* src = p.ws + ((size_t)tile * p.splits * BM + row) * BN + c4;
  for (int s = 0; s < p.splits; ++s) {
    const float4 t = *reinterpret_cast<const float4*>(src + (size_t)s * BM * BN);
    acc.x += t.x;
    acc.y += t.y;
    acc.z += t.z;
    acc.w += t.w;
  }
  const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gcol = gcol0 + i;
    if (gcol < p.N && !(p.sym && grow > gcol)) {
      const float val = epi_apply(p, a4[i], grow, gcol);
      p.out[(long long)grow * p.ldo + gcol] = val;
      if (p.sym && grow != gcol) p.out[(long long)gcol * p.ldo + grow] = val;
    }
  }
}

__global__ void __launch_bounds__(256) gemm3x_finalize_kernel(const GemmParams p) {
  finalize_body(p, (long long)blockIdx.x * blockDim.x + threadIdx.x);
}

// fin_end[i] = exclusive prefix sum of finalize CTA counts (0 CTAs for problems without split-K)
__global__ void __launch_bounds__(256) gemm3x_finalize_group_kernel(const GemmParams* __restrict__ probs,
                                                                    const int* __restrict__ fin_end, int nprob) {
  __shared__ __align__(16) GemmParams sp;
  const int idx = find_problem(fin_end, nprob, (int)blockIdx.x);
  const int first = idx > 0 ? fin_end[idx - 1] : 0;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(probs + idx);
  for (int i = threadIdx.x; i < (int)(sizeof(GemmParams) / 4); i += 256) reinterpret_cast<uint32_t*>(&sp)[i] = src[i];
  __syncthreads();
  if (sp.skip != nullptr && *reinterpret_cast<const volatile int*>(sp.skip) != 0) return;
  finalize_body(sp, (long long)((int)blockIdx.x - first) * 256 + threadIdx.x);
}

}  // namespace kfac
