// Host-side planning + launch of gemm3x (tile shape, split-K, stages, loader kinds), single and grouped.
#pragma once
#include <cstring>
#include <vector>

#include "common.cuh"
#include "gemm3x.cuh"

namespace kfac {

static int g_smem_optin[64] = {0};

static int ensure_gemm_attr() {
  int dev = 0;
  KFAC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail("device index %d out of range", dev);
  if (!g_smem_optin[dev]) {
    int maxopt = 0;
    KFAC_CUDA(cudaDeviceGetAttribute(&maxopt, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    KFAC_CUDA(cudaFuncSetAttribute(gemm3x_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxopt - 6144));
    KFAC_CUDA(cudaFuncSetAttribute(gemm3x_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxopt - 6144));
    KFAC_CUDA(cudaFuncSetAttribute(gemm3x_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, maxopt - 6144));
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

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Loader kind per operand: vector (float4) paths need 16-byte alignment and K (resp. ld, S) divisible by 4.
static void choose_load_kind(Operand& op, int K) {
  if (op.mode == OP_IM2COL && op.KH == 1 && op.KW == 1 && op.SH == 1 && op.SW == 1 && op.PH == 0 && op.PW == 0 &&
      op.ones_row < 0) {
    // a 1x1 stride-1 convolution reads its input exactly like an (N, C, S) tensor
    op.mode = OP_NCS;
    op.S = op.H * op.W;
  }
  switch (op.mode) {
    case OP_KCONTIG:
      op.lk = (op.elt == 0 && op.ext == nullptr && op.ld % 4 == 0 && K % 4 == 0 && aligned16(op.p)) ? LK_KC_VEC : LK_KC_SCALAR;
      break;
    case OP_RCONTIG:
      op.lk = (op.elt == 0 && op.ld % 4 == 0 && aligned16(op.p)) ? LK_RC_VEC : LK_RC_SCALAR;   // 16-bit inputs: scalar loaders
      break;
    case OP_NCS:
      op.lk = (op.elt == 0 && op.S % 4 == 0 && aligned16(op.p)) ? LK_NCS_VEC : LK_NCS_SCALAR;
      break;
    default:
      op.lk = LK_IM2COL;
      break;
  }
}

// Fills the tiling fields of p. allow_split: epilogue is AXPBY so split-K partials may be used; target_ctas: how many
// CTAs this problem should try to occupy (148 alone, fewer inside a large group).
static void plan_gemm(GemmParams& p, bool allow_split, int target_ctas = 148) {
  p.BN = pick_bn(p.N);
  p.num_tiles = count_tiles(p.sym, p.M, p.N, p.BN, p.tiles_m, p.tiles_n);
  const int nkb = (p.K + BK - 1) / BK;
  int splits = 1;
  if (allow_split) {
    // pick the split count that minimises (waves) x (k-blocks per CTA + fixed per-CTA cost), so that e.g. 90 tiles
    // on 148 SMs run as 270 CTAs in two short waves instead of one long, 61 %-occupied wave
    const int max_by_k = nkb / 4 > 1 ? nkb / 4 : 1;  // keep >= 4 K-blocks per split
    const int fixed_kb = 6;                          // prologue + epilogue + finalize, in k-block units
    double best = 1e300;
    for (int s = 1; s <= max_by_k && s <= 160; ++s) {
      if (s > 1 && (long long)p.num_tiles * s > 8LL * target_ctas) break;
      const int per = (nkb + s - 1) / s;
      const long long ctas = (long long)p.num_tiles * ((nkb + per - 1) / per);
      const long long waves = (ctas + target_ctas - 1) / target_ctas;
      const double cost = (double)waves * (per + fixed_kb + (s > 1 ? 2 : 0));
      if (cost < best * 0.97) {   // prefer fewer splits unless the gain is real
        best = cost;
        splits = s;
      }
    }
  }
  p.kb_per_split = (nkb + splits - 1) / splits;
  p.splits = (nkb + p.kb_per_split - 1) / p.kb_per_split;
  const int budget = 227 * 1024 - 2048 - 1024 - 4096;  // static smem: barriers, im2col row table, group params
  int stages = budget / stage_bytes(p.BN);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages > p.kb_per_split) stages = p.kb_per_split;
  if (stages < 1) stages = 1;
  p.stages = stages;
  if (p.promote_kb < 1) p.promote_kb = 2;  // default: 24 UMMAs per chain, ~7e-7 truncation bias
  if (p.promote_kb > 8) p.promote_kb = 8;
  choose_load_kind(p.a, p.K);
  choose_load_kind(p.b, p.K);
}

static size_t gemm_ws_bytes(const GemmParams& p) {
  return p.splits > 1 ? (size_t)p.num_tiles * p.splits * BM * p.BN * sizeof(float) : 0;
}

static size_t gemm_smem_bytes(const GemmParams& p) {
  const size_t pipeline = (size_t)p.stages * stage_bytes(p.BN);
  const size_t staging = (size_t)BM * (p.BN + 4) * sizeof(float);  // epilogue tile, reuses the pipeline stages
  return std::max(pipeline, staging) + 1024;
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
  if (p.a.elt != 0 || p.b.elt != 0)
    gemm3x_kernel<true><<<p.num_tiles * p.splits, NTHREADS, gemm_smem_bytes(p), stream>>>(p);
  else
    gemm3x_kernel<false><<<p.num_tiles * p.splits, NTHREADS, gemm_smem_bytes(p), stream>>>(p);
  KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  if (p.splits > 1) {
    const long long threads = (long long)p.num_tiles * BM * p.BN / 4;
    gemm3x_finalize_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(p); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
  }
  return 0;
}

// ------------------------------------------------------------------ grouped launches
// A group collects independent problems and runs them as ONE gemm3x_group_kernel launch (+ one finalize launch if
// any problem uses split-K).  The problem table and split-K partials live in caller-provided device scratch.
struct GemmGroup {
  std::vector<GemmParams> probs;
  std::vector<int> cta_end, fin_end;
  size_t partial_bytes = 0;  // split-K scratch handed out so far
  size_t max_smem = 0;
  int ctas = 0, fins = 0;

  void clear() {
    probs.clear(); cta_end.clear(); fin_end.clear();
    partial_bytes = 0; max_smem = 0; ctas = 0; fins = 0;
  }
  // table bytes for the problems added so far
  size_t table_bytes() const { return (probs.size() * sizeof(GemmParams) + 255) / 256 * 256 + 2 * ((probs.size() * 4 + 255) / 256 * 256); }
};

// partials: device scratch for split-K partial tiles of the whole group (may be null if allow_split is false)
static int group_add(GemmGroup& g, GemmParams p, bool allow_split, int target_ctas, char* partials, size_t partials_bytes) {
  if (p.M <= 0 || p.N <= 0 || p.K <= 0) return 0;  // nothing to do
  plan_gemm(p, allow_split && partials != nullptr, target_ctas);
  if (p.splits > 1) {
    const size_t need = (gemm_ws_bytes(p) + 255) / 256 * 256;
    if (g.partial_bytes + need > partials_bytes) {
      // not enough scratch for split-K: fall back to a single split
      plan_gemm(p, false, target_ctas);
    } else {
      if (p.epi != EPI_AXPBY) return fail("gemm3x group: split-K requires the AXPBY epilogue");
      p.ws = reinterpret_cast<float*>(partials + g.partial_bytes);
      p.epi = EPI_PARTIAL;
      g.partial_bytes += need;
    }
  }
  g.ctas += p.num_tiles * p.splits;
  g.cta_end.push_back(g.ctas);
  if (p.splits > 1) g.fins += (int)(((long long)p.num_tiles * BM * p.BN / 4 + 255) / 256);
  g.fin_end.push_back(g.fins);
  g.max_smem = std::max(g.max_smem, gemm_smem_bytes(p));
  g.probs.push_back(p);
  return 0;
}

static int group_launch(GemmGroup& g, char* table, size_t table_bytes, cudaStream_t stream) {
  if (g.probs.empty()) return 0;
  if (ensure_gemm_attr()) return 1;
  if (table == nullptr || table_bytes < g.table_bytes()) return fail("gemm3x group: table scratch too small");
  const size_t np = g.probs.size();
  const size_t off_cta = (np * sizeof(GemmParams) + 255) / 256 * 256;
  const size_t off_fin = off_cta + (np * 4 + 255) / 256 * 256;
  KFAC_CUDA(cudaMemcpyAsync(table, g.probs.data(), np * sizeof(GemmParams), cudaMemcpyHostToDevice, stream));
  KFAC_CUDA(cudaMemcpyAsync(table + off_cta, g.cta_end.data(), np * 4, cudaMemcpyHostToDevice, stream));
  const GemmParams* dprobs = reinterpret_cast<const GemmParams*>(table);
  gemm3x_group_kernel<<<g.ctas, NTHREADS, g.max_smem, stream>>>(dprobs, reinterpret_cast<const int*>(table + off_cta), (int)np); KFAC_COUNT_LAUNCH(1);
  KFAC_CUDA(cudaGetLastError());
  if (g.fins > 0) {
    KFAC_CUDA(cudaMemcpyAsync(table + off_fin, g.fin_end.data(), np * 4, cudaMemcpyHostToDevice, stream));
    gemm3x_finalize_group_kernel<<<g.fins, 256, 0, stream>>>(dprobs, reinterpret_cast<const int*>(table + off_fin), (int)np); KFAC_COUNT_LAUNCH(1);
    KFAC_CUDA(cudaGetLastError());
  }
  return 0;
}

// A group serialised into a host buffer that is uploaded once for a whole sequence of launches (the band reduction
// issues hundreds of small grouped GEMMs back to back: a table upload in front of each would sit on the critical path).
struct PreparedGroup {
  size_t off = 0;        // byte offset of the problem table in the uploaded buffer
  size_t off_cta = 0, off_fin = 0;
  int np = 0, ctas = 0, fins = 0;
  size_t smem = 0;
};

static PreparedGroup group_prepare(const GemmGroup& g, std::vector<char>& host) {
  PreparedGroup pg;
  pg.np = (int)g.probs.size();
  if (pg.np == 0) return pg;
  const size_t np = g.probs.size();
  pg.off = (host.size() + 255) / 256 * 256;
  pg.off_cta = (np * sizeof(GemmParams) + 255) / 256 * 256;
  pg.off_fin = pg.off_cta + (np * 4 + 255) / 256 * 256;
  const size_t total = pg.off_fin + (np * 4 + 255) / 256 * 256;
  host.resize(pg.off + total);
  memcpy(host.data() + pg.off, g.probs.data(), np * sizeof(GemmParams));
  memcpy(host.data() + pg.off + pg.off_cta, g.cta_end.data(), np * 4);
  memcpy(host.data() + pg.off + pg.off_fin, g.fin_end.data(), np * 4);
  pg.ctas = g.ctas;
  pg.fins = g.fins;
  pg.smem = g.max_smem;
  return pg;
}

static int group_run(const PreparedGroup& pg, const char* devbase, cudaStream_t stream) {
  if (pg.np == 0) return 0;
  const char* table = devbase + pg.off;
  const GemmParams* dprobs = reinterpret_cast<const GemmParams*>(table);
  gemm3x_group_kernel<<<pg.ctas, NTHREADS, pg.smem, stream>>>(dprobs, reinterpret_cast<const int*>(table + pg.off_cta), pg.np); KFAC_COUNT_LAUNCH(1);
  if (pg.fins > 0) {
    gemm3x_finalize_group_kernel<<<pg.fins, 256, 0, stream>>>(dprobs, reinterpret_cast<const int*>(table + pg.off_fin), pg.np); KFAC_COUNT_LAUNCH(1);
  }
  KFAC_CUDA(cudaGetLastError());
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
