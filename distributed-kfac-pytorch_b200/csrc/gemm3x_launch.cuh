// Host-side planning + launch of gemm3x (tile shape, split-K, stages) and operand helpers.
#pragma once
#include <cstring>
#include "common.cuh"
#include "gemm3x.cuh"

namespace kfac {

// ------------------------------------------------------------------ gemm3x launcher
static int g_smem_optin[64] = {0};

static int ensure_gemm_attr() {
  int dev = 0;
  KFAC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail("device index %d out of range", dev);
  if (!g_smem_optin[dev]) {
    int maxopt = 0;
    KFAC_CUDA(cudaDeviceGetAttribute(&maxopt, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    KFAC_CUDA(cudaFuncSetAttribute(gemm3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, maxopt - 1024));
    g_smem_optin[dev] = maxopt;
  }
  return 0;
}

static int pick_bn(int N) {
  if (N >= 256) return 256;
  return ((N + 31) / 32) * 32;
}

static int count_tiles(int sym, int M, int N, int BN, int& tiles_m, int& tiles_n) {
  tiles_m = (M + BM - 1) / BM;
  tiles_n = (N + BN - 1) / BN;
  if (!sym) return tiles_m * tiles_n;
  int t = 0;
  for (int J = 0; J < tiles_n; ++J) t += sym_rows_for_col(N, BN, tiles_m, J);
  return t;
}

// Fills the tiling fields of p. allow_split: epilogue is AXPBY so split-K partials may be used.
static void plan_gemm(GemmParams& p, bool allow_split) {
  p.BN = pick_bn(p.N);
  p.num_tiles = count_tiles(p.sym, p.M, p.N, p.BN, p.tiles_m, p.tiles_n);
  const int nkb = (p.K + BK - 1) / BK;
  int splits = 1;
  if (allow_split && p.num_tiles < 148) {
    splits = 148 / p.num_tiles;
    const int max_by_k = nkb / 4 > 1 ? nkb / 4 : 1;  // keep >= 4 K-blocks per split
    if (splits > max_by_k) splits = max_by_k;
    if (splits < 1) splits = 1;
  }
  p.kb_per_split = (nkb + splits - 1) / splits;
  p.splits = (nkb + p.kb_per_split - 1) / p.kb_per_split;
  const int budget = 227 * 1024 - 1024 - 2048;
  int stages = budget / stage_bytes(p.BN);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages > p.kb_per_split) stages = p.kb_per_split;
  if (stages < 1) stages = 1;
  p.stages = stages;
  if (p.promote_kb < 1) p.promote_kb = 2;  // default: 24 UMMAs per chain, ~7e-7 truncation bias
  if (p.promote_kb > 8) p.promote_kb = 8;
}

static size_t gemm_ws_bytes(const GemmParams& p) {
  return p.splits > 1 ? (size_t)p.num_tiles * p.splits * BM * p.BN * sizeof(float) : 0;
}

static int launch_gemm(GemmParams& p, bool allow_split, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (p.M <= 0 || p.N <= 0 || p.K <= 0) return fail("gemm3x: empty problem %d x %d x %d", p.M, p.N, p.K);
  if (ensure_gemm_attr()) return 1;
  plan_gemm(p, allow_split);
  const int final_epi = p.epi;
  if (p.splits > 1) {
    const size_t need = gemm_ws_bytes(p);
    if (ws == nullptr || ws_bytes < need)
      return fail("gemm3x: workspace too small (%zu bytes given, %zu needed)", ws_bytes, need);
    if (final_epi != EPI_AXPBY) return fail("gemm3x: split-K requires the AXPBY epilogue");
    p.ws = static_cast<float*>(ws);
    p.epi = EPI_PARTIAL;
  }
  const size_t smem = (size_t)p.stages * stage_bytes(p.BN) + 1024;
  gemm3x_kernel<<<p.num_tiles * p.splits, NTHREADS, smem, stream>>>(p); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  if (p.splits > 1) {
    const long long threads = (long long)p.num_tiles * BM * p.BN / 4;
    gemm3x_finalize_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(p); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
  }
  return 0;
}

static Operand dense_operand(const float* ptr, bool kcontig, int R, long long ld) {
  Operand o;
  memset(&o, 0, sizeof(o));
  o.p = ptr;
  o.mode = kcontig ? OP_KCONTIG : OP_RCONTIG;
  o.R = R;
  o.ones_row = -1;
  o.ld = ld;
  o.ext = nullptr;
  o.ext_k = -1;
  o.KH = o.KW = 1;
  return o;
}

static void init_params(GemmParams& p) {
  memset(&p, 0, sizeof(p));
  p.epi = EPI_AXPBY;
  p.alpha = 1.f;
  p.beta = 0.f;
}

// Gram matrix F <- beta F + alpha_total * Op Op^T for a single operand description.
static int launch_factor(const Operand& op, long long K, float alpha_total, float beta, float* F, void* ws,
                         size_t ws_bytes, cudaStream_t stream) {
  if (K > 0x7fffffffLL) return fail("factor: reduction length %lld exceeds int32", K);
  GemmParams p;
  init_params(p);
  p.a = op;
  p.b = op;
  p.same_ab = 1;
  p.sym = 1;
  p.M = p.N = op.R;
  p.K = (int)K;
  p.alpha = alpha_total;
  p.beta = beta;
  p.out = F;
  p.ldo = op.R;
  return launch_gemm(p, true, ws, ws_bytes, stream);
}

}  // namespace kfac
