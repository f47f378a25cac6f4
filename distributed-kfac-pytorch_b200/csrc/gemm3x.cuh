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
#include "ptx.cuh"

namespace kfac {

enum OpMode : int { OP_KCONTIG = 0, OP_RCONTIG = 1, OP_IM2COL = 2, OP_NCS = 3 };
enum EpiMode : int { EPI_PARTIAL = 0, EPI_AXPBY = 1, EPI_DIV_OUTER = 2, EPI_MUL_MAT = 3 };

struct Operand {
  const float* p;
  int mode;
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
};

constexpr int BM = 128;
constexpr int BK = 32;
constexpr int NTHREADS = 256;
constexpr int NWARPS = 8;
constexpr int TILE_A_BYTES = BM * 128;
constexpr int MAX_STAGES = 4;
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

// Gather rows [row0, row0+nrows) x k in [k0, k0+32) of a logical operand into a (hi, lo) smem tile pair.
// Loads are issued in batches of LOAD_BATCH independent requests per thread before any is consumed, so each
// SM keeps ~16 KB in flight (the first version consumed each load immediately and was latency-bound).
constexpr int LOAD_BATCH = 16;

template <int MODE>
__device__ __forceinline__ void load_rows_kc(const Operand& op, int row0, int nrows, int k0, int kend, uint8_t* hi,
                                             uint8_t* lo, int warp, int lane) {
  // K-contiguous family: lane = k offset -> 128-byte coalesced global reads, conflict-free smem rows
  const float* __restrict__ p = op.p;
  const int k = k0 + lane;
  const bool kin = k < kend;
  long long base = 0;
  int ihb = 0, iwb = 0;
  if (MODE == OP_IM2COL) {
    const int S = op.OH * op.OW;
    const int n = k / S;
    const int rem = k - n * S;
    const int oh = rem / op.OW;
    const int ow = rem - oh * op.OW;
    ihb = oh * op.SH - op.PH;
    iwb = ow * op.SW - op.PW;
    base = (long long)n * op.C * op.H * op.W;
  } else if (MODE == OP_NCS) {
    const int n = k / op.S;
    const int s = k - n * op.S;
    base = (long long)n * op.C * op.S + s;
  } else {
    base = k;
  }
  const int khw = op.KH * op.KW;
  for (int rb = warp; rb < nrows; rb += NWARPS * LOAD_BATCH) {
    float v[LOAD_BATCH];
#pragma unroll
    for (int u = 0; u < LOAD_BATCH; ++u) {
      const int rr = rb + u * NWARPS;
      const int r = row0 + rr;
      float x = 0.f;
      if (kin && rr < nrows && r < op.R) {
        if (r == op.ones_row) {
          x = 1.f;
        } else if (MODE == OP_KCONTIG) {
          x = (op.ext != nullptr && k == op.ext_k) ? __ldg(op.ext + r) : __ldg(p + (long long)r * op.ld + base);
        } else if (MODE == OP_NCS) {
          x = __ldg(p + base + (long long)r * op.S);
        } else {
          const int c = r / khw;
          const int t = r - c * khw;
          const int i = t / op.KW;
          const int j = t - i * op.KW;
          const int ih = ihb + i, iw = iwb + j;
          if ((unsigned)ih < (unsigned)op.H && (unsigned)iw < (unsigned)op.W)
            x = __ldg(p + base + ((long long)c * op.H + ih) * op.W + iw);
        }
      }
      v[u] = x;
    }
#pragma unroll
    for (int u = 0; u < LOAD_BATCH; ++u) {
      const int rr = rb + u * NWARPS;
      if (rr < nrows) store_split(hi, lo, sw128_offset(rr, lane), v[u]);
    }
  }
}

__device__ __forceinline__ void load_rows_rc(const Operand& op, int row0, int nrows, int k0, int kend, uint8_t* hi,
                                             uint8_t* lo, int warp, int lane) {
  // lanes walk consecutive rows (contiguous in memory); the smem write is the transpose
  const float* __restrict__ p = op.p;
  const int total = ((nrows + 31) >> 5) * 32;  // (row group, k) pairs
  for (int tb = warp; tb < total; tb += NWARPS * LOAD_BATCH) {
    float v[LOAD_BATCH];
#pragma unroll
    for (int u = 0; u < LOAD_BATCH; ++u) {
      const int t = tb + u * NWARPS;
      const int rr = ((t >> 5) << 5) + lane;
      const int r = row0 + rr;
      const int k = k0 + (t & 31);
      float x = 0.f;
      if (t < total && rr < nrows && k < kend && r < op.R)
        x = (r == op.ones_row) ? 1.f : __ldg(p + (long long)k * op.ld + r);
      v[u] = x;
    }
#pragma unroll
    for (int u = 0; u < LOAD_BATCH; ++u) {
      const int t = tb + u * NWARPS;
      const int rr = ((t >> 5) << 5) + lane;
      if (t < total && rr < nrows) store_split(hi, lo, sw128_offset(rr, t & 31), v[u]);
    }
  }
}

__device__ __forceinline__ void load_rows(const Operand& op, int row0, int nrows, int k0, int kend, uint8_t* hi,
                                          uint8_t* lo, int warp, int lane) {
  switch (op.mode) {
    case OP_RCONTIG: load_rows_rc(op, row0, nrows, k0, kend, hi, lo, warp, lane); break;
    case OP_IM2COL: load_rows_kc<OP_IM2COL>(op, row0, nrows, k0, kend, hi, lo, warp, lane); break;
    case OP_NCS: load_rows_kc<OP_NCS>(op, row0, nrows, k0, kend, hi, lo, warp, lane); break;
    default: load_rows_kc<OP_KCONTIG>(op, row0, nrows, k0, kend, hi, lo, warp, lane); break;
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

__global__ void __launch_bounds__(NTHREADS, 1) gemm3x_kernel(const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment (swizzle is a function of absolute address bits 7..9)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bars[MAX_STAGES + 1];
  __shared__ uint32_t tmem_base_slot;
  __shared__ double dot_red[NWARPS];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int BN = p.BN;
  const int sbytes = stage_bytes(BN);

  const int tile = blockIdx.x / p.splits;
  const int split = blockIdx.x - tile * p.splits;
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

  uint32_t tmem_cols = 32;
  while (tmem_cols < (uint32_t)BN) tmem_cols <<= 1;

  if (tid == 0) {
    for (int s = 0; s <= MAX_STAGES; ++s) mbar_init(smem_u32(&bars[s]), 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_slot), tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_d = tmem_base_slot;
  const uint32_t idesc = make_idesc_tf32(BM, BN);
  const uint32_t done_bar = smem_u32(&bars[MAX_STAGES]);

  // fp32 register accumulators: thread owns tile row (32q + lane), columns [h*BN/2, (h+1)*BN/2).
  // The tensor core adds into TMEM with truncation, which biases long chains by ~3e-8 per UMMA (measured);
  // so TMEM only ever holds promote_kb k-blocks (12 UMMAs each) before it is drained into these registers
  // with round-to-nearest adds.
  const int q = warp & 3, h = warp >> 2;
  const int half = BN >> 1;
  float acc[MAX_BN / 2];
#pragma unroll
  for (int i = 0; i < MAX_BN / 2; ++i) acc[i] = 0.f;
  uint32_t done_phase = 0;

  for (int kb = 0; kb < nkb; ++kb) {
    const int s = kb % p.stages;
    const int use = kb / p.stages;
    if (use > 0) mbar_wait(smem_u32(&bars[s]), (use - 1) & 1);  // UMMAs that read this stage have retired
    uint8_t* st = smem + (size_t)s * sbytes;
    uint8_t* a_hi = st;
    uint8_t* a_lo = st + TILE_A_BYTES;
    uint8_t* b_hi = st + 2 * TILE_A_BYTES;
    uint8_t* b_lo = b_hi + BN * 128;
    const int k0 = (kb_begin + kb) * BK;
    if (!a_in_b) load_rows(p.a, arow0, BM, k0, kend, a_hi, a_lo, warp, lane);
    if (!b_in_a) load_rows(p.b, bcol0, BN, k0, kend, b_hi, b_lo, warp, lane);
    fence_proxy_async_smem();
    __syncthreads();
    const bool promote = ((kb % p.promote_kb) == p.promote_kb - 1) || kb == nkb - 1;
    if (tid == 0) {
      tc_fence_after_sync();
      uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
      if (a_in_b) {
        ah = bh + (uint32_t)(arow0 - bcol0) * 128;
        al = bl + (uint32_t)(arow0 - bcol0) * 128;
      } else if (b_in_a) {
        bh = ah + (uint32_t)(bcol0 - arow0) * 128;
        bl = al + (uint32_t)(bcol0 - arow0) * 128;
      }
      const bool fresh = (kb % p.promote_kb) == 0;  // first k-block after a drain overwrites TMEM
#pragma unroll
      for (int ks = 0; ks < BK / 8; ++ks) {
        const uint64_t dah = make_kmajor_sw128_desc(ah + ks * 32), dal = make_kmajor_sw128_desc(al + ks * 32);
        const uint64_t dbh = make_kmajor_sw128_desc(bh + ks * 32), dbl = make_kmajor_sw128_desc(bl + ks * 32);
        mma_tf32_ss(tmem_d, dal, dbh, idesc, !(fresh && ks == 0));
        mma_tf32_ss(tmem_d, dah, dbl, idesc, 1);
        mma_tf32_ss(tmem_d, dah, dbh, idesc, 1);
      }
      mma_commit(smem_u32(&bars[s]));
      if (promote) mma_commit(done_bar);
    }
    if (promote) {
      mbar_wait(done_bar, done_phase);
      done_phase ^= 1;
      tc_fence_after_sync();
#pragma unroll
      for (int ci = 0; ci < MAX_BN / 32; ++ci) {
        if (ci * 16 < half) {
          uint32_t v[16];
          tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * half + ci * 16), v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) acc[ci * 16 + i] += __uint_as_float(v[i]);
        }
      }
      // the next UMMA (overwriting TMEM) is only issued after the __syncthreads of the next iteration
      tc_fence_before_sync();
    }
  }

  // ---- epilogue from registers
  const int grow = arow0 + q * 32 + lane;
  float dot_local = 0.f;
#pragma unroll
  for (int ci = 0; ci < MAX_BN / 32; ++ci) {
    if (ci * 16 < half) {
      const int c0 = h * half + ci * 16;
      if (p.epi == EPI_PARTIAL) {
        float4* dst = reinterpret_cast<float4*>(p.ws + ((size_t)blockIdx.x * BM + (q * 32 + lane)) * BN + c0);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          dst[i] = make_float4(acc[ci * 16 + 4 * i], acc[ci * 16 + 4 * i + 1], acc[ci * 16 + 4 * i + 2],
                               acc[ci * 16 + 4 * i + 3]);
      } else if (grow < p.M) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int gcol = bcol0 + c0 + i;
          if (gcol < p.N && !(p.sym && grow > gcol)) {
            const float val = epi_apply(p, acc[ci * 16 + i], grow, gcol);
            p.out[(long long)grow * p.ldo + gcol] = val;
            if (p.sym && grow != gcol) p.out[(long long)gcol * p.ldo + grow] = val;
            if (p.dot_out != nullptr) dot_local += val * dot_operand(p, grow, gcol);
          }
        }
      }
    }
  }
  if (p.dot_out != nullptr && p.epi != EPI_PARTIAL) {
    double d = (double)dot_local;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) dot_red[warp] = d;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tmem_dealloc(tmem_d, tmem_cols);
    if (p.dot_out != nullptr && p.epi != EPI_PARTIAL && lane == 0) {
      double d = 0.0;
      for (int w = 0; w < NWARPS; ++w) d += dot_red[w];
      atomicAdd(p.dot_out, d);
    }
  }
}

// Sum split-K partials and apply the AXPBY epilogue (+ symmetric mirror). One thread per 4 columns.
__global__ void __launch_bounds__(256) gemm3x_finalize_kernel(const GemmParams p) {
  const int BN = p.BN;
  const int per_tile = BM * BN / 4;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
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

}  // namespace kfac
