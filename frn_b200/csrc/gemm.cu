// Strided batched fp32 GEMM on the FFMA pipe (exact fp32 products, fp32
// accumulation), used for every 1x1 projection of the fusion block and its
// gradients: conv0_rgb / conv0_evt (retinanet/fusion.py:23-24), g / theta /
// phi / W (retinanet/Attention.py:24-27,34).  fp32 parity with the reference
// (<= 1e-5 relative vs an fp64 run) rules out single-pass TF32 tensor-core
// math here; see DESIGN.md "precision".
//
// Tiling: BMxBN output tile per CTA, BK-deep smem stages (double buffered via
// registers), TMxTN register tile per thread read with 128-bit LDS.
#include <cstdlib>
#include "common.cuh"

namespace frn {

#ifdef FRN_DEBUG_ENGINES      // the FFMA engine: scaffolding of round 1, kept for A/B comparisons in debug builds only
namespace {

template <int BM, int BN, int BK, int TM, int TN, int EPI>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
sgemm_kernel(GemmDesc p, float* __restrict__ aux, int ktiles, int tiles_per_split, int nsplit) {
  constexpr int NT = (BM / TM) * (BN / TN);
  constexpr int PAD = 4;
  constexpr int LA = BM * BK / NT;
  constexpr int LB = BK * BN / NT;
  constexpr int GM = TM / 4, GN = TN / 4;          // float4 groups per thread
  constexpr int SM_STRIDE = BM / GM, SN_STRIDE = BN / GN;
  static_assert(BM * BK % NT == 0 && BK * BN % NT == 0, "tile/threads");
  __shared__ __align__(16) float As[2][BK][BM + PAD];
  __shared__ __align__(16) float Bs[2][BK][BN + PAD];

  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  int z = blockIdx.z;
  int t_begin, t_end, b_fixed;
  if (p.reduce_batch) {
    t_begin = z * tiles_per_split;
    t_end = min(t_begin + tiles_per_split, p.batch * ktiles);
    b_fixed = -1;
  } else {
    t_begin = 0; t_end = ktiles; b_fixed = z;
  }
  const bool a_mfast = (p.sAm == 1);
  const bool b_nfast = (p.sBn == 1);

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; i++)
#pragma unroll
    for (int j = 0; j < TN; j++) acc[i][j] = 0.f;

  float ra[LA], rb[LB];

  auto load_tile = [&](int t) {
    int b = b_fixed >= 0 ? b_fixed : t / ktiles;
    int k0 = (b_fixed >= 0 ? t : t % ktiles) * BK;
    const float* Ab = p.A + (long)b * p.sAb;
    const float* Bb = p.B + (long)b * p.sBb;
#pragma unroll
    for (int r = 0; r < LA; r++) {
      int idx = tid + r * NT, m, k;
      if (a_mfast) { m = idx % BM; k = idx / BM; } else { k = idx % BK; m = idx / BK; }
      int gm = m0 + m, gk = k0 + k;
      ra[r] = (gm < p.M && gk < p.K) ? __ldg(Ab + (long)gm * p.sAm + (long)gk * p.sAk) : 0.f;
    }
#pragma unroll
    for (int r = 0; r < LB; r++) {
      int idx = tid + r * NT, n, k;
      if (b_nfast) { n = idx % BN; k = idx / BN; } else { k = idx % BK; n = idx / BK; }
      int gn = n0 + n, gk = k0 + k;
      rb[r] = (gn < p.N && gk < p.K) ? __ldg(Bb + (long)gk * p.sBk + (long)gn * p.sBn) : 0.f;
    }
  };
  auto store_tile = [&](int buf) {
#pragma unroll
    for (int r = 0; r < LA; r++) {
      int idx = tid + r * NT, m, k;
      if (a_mfast) { m = idx % BM; k = idx / BM; } else { k = idx % BK; m = idx / BK; }
      As[buf][k][m] = ra[r];
    }
#pragma unroll
    for (int r = 0; r < LB; r++) {
      int idx = tid + r * NT, n, k;
      if (b_nfast) { n = idx % BN; k = idx / BN; } else { k = idx % BK; n = idx / BK; }
      Bs[buf][k][n] = rb[r];
    }
  };

  if (t_begin < t_end) {
    load_tile(t_begin);
    store_tile(0);
  }
  __syncthreads();
  int buf = 0;
  for (int t = t_begin; t < t_end; t++) {
    if (t + 1 < t_end) load_tile(t + 1);
#pragma unroll
    for (int k = 0; k < BK; k++) {
      float a[TM], b[TN];
#pragma unroll
      for (int g = 0; g < GM; g++) {
        float4 v = *reinterpret_cast<const float4*>(&As[buf][k][g * SM_STRIDE + ty * 4]);
        a[g * 4 + 0] = v.x; a[g * 4 + 1] = v.y; a[g * 4 + 2] = v.z; a[g * 4 + 3] = v.w;
      }
#pragma unroll
      for (int g = 0; g < GN; g++) {
        float4 v = *reinterpret_cast<const float4*>(&Bs[buf][k][g * SN_STRIDE + tx * 4]);
        b[g * 4 + 0] = v.x; b[g * 4 + 1] = v.y; b[g * 4 + 2] = v.z; b[g * 4 + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < TM; i++)
#pragma unroll
        for (int j = 0; j < TN; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (t + 1 < t_end) {
      store_tile(buf ^ 1);
      __syncthreads();
      buf ^= 1;
    }
  }

  if (EPI == GEMM_EPI_ROWSUMSQ) {
    // aux[(b*M + m) * gridDim.x + blockIdx.x] = sum over this tile's columns of acc^2
    constexpr int TXN = BN / TN;
    static_assert(TXN <= 32 && (32 % TXN) == 0, "row reduce within a warp");
#pragma unroll
    for (int i = 0; i < TM; i++) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < TN; j++) {
        int gn = n0 + (j / 4) * SN_STRIDE + tx * 4 + (j % 4);
        float v = p.alpha * acc[i][j];
        s += (gn < p.N) ? v * v : 0.f;
      }
#pragma unroll
      for (int o = TXN / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      int gm = m0 + (i / 4) * SM_STRIDE + ty * 4 + (i % 4);
      if (tx == 0 && gm < p.M) aux[((long)z * p.M + gm) * gridDim.x + blockIdx.x] = s;
    }
    return;
  }

  // ---- store ----
  float* Cb;
  long sCm, sCn;
  bool partial = p.reduce_batch && nsplit > 1;
  if (partial) { Cb = aux + (long)z * p.M * p.N; sCm = p.N; sCn = 1; }
  else { Cb = p.C + (p.reduce_batch ? 0 : (long)z * p.sCb); sCm = p.sCm; sCn = p.sCn; }
  const bool acc_out = !partial && p.accumulate;
  const bool vec_n = (sCn == 1) && ((sCm & 3) == 0) && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0);
  const bool vec_m = (sCm == 1) && ((sCn & 3) == 0) && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0);
#pragma unroll
  for (int i = 0; i < TM; i++) {
    int gm = m0 + (i / 4) * SM_STRIDE + ty * 4 + (i % 4);
    float bias = (!partial && p.bias_m && gm < p.M) ? __ldg(p.bias_m + (p.reduce_batch ? 0 : (long)z * p.bias_bs) + gm) : 0.f;
#pragma unroll
    for (int j = 0; j < TN; j++) {
      int gn = n0 + (j / 4) * SN_STRIDE + tx * 4 + (j % 4);
      float bn = (!partial && p.bias_n && gn < p.N) ? __ldg(p.bias_n + (p.reduce_batch ? 0 : (long)z * p.bias_bs) + gn) : 0.f;
      acc[i][j] = fmaf(p.alpha, acc[i][j], bias + bn);
    }
  }
  if (vec_n) {
#pragma unroll
    for (int i = 0; i < TM; i++) {
      int gm = m0 + (i / 4) * SM_STRIDE + ty * 4 + (i % 4);
      if (gm >= p.M) continue;
#pragma unroll
      for (int g = 0; g < GN; g++) {
        int gn = n0 + g * SN_STRIDE + tx * 4;
        float* dst = Cb + (long)gm * sCm + gn;
        if (gn + 3 < p.N) {
          float4 v = make_float4(acc[i][g * 4], acc[i][g * 4 + 1], acc[i][g * 4 + 2], acc[i][g * 4 + 3]);
          if (acc_out) { float4 o = *reinterpret_cast<float4*>(dst); v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
          *reinterpret_cast<float4*>(dst) = v;
        } else {
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (gn + q < p.N) dst[q] = acc[i][g * 4 + q] + (acc_out ? dst[q] : 0.f);
        }
      }
    }
  } else if (vec_m) {
#pragma unroll
    for (int j = 0; j < TN; j++) {
      int gn = n0 + (j / 4) * SN_STRIDE + tx * 4 + (j % 4);
      if (gn >= p.N) continue;
#pragma unroll
      for (int g = 0; g < GM; g++) {
        int gm = m0 + g * SM_STRIDE + ty * 4;
        float* dst = Cb + (long)gn * sCn + gm;
        if (gm + 3 < p.M) {
          float4 v = make_float4(acc[g * 4][j], acc[g * 4 + 1][j], acc[g * 4 + 2][j], acc[g * 4 + 3][j]);
          if (acc_out) { float4 o = *reinterpret_cast<float4*>(dst); v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
          *reinterpret_cast<float4*>(dst) = v;
        } else {
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (gm + q < p.M) dst[q] = acc[g * 4 + q][j] + (acc_out ? dst[q] : 0.f);
        }
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < TM; i++) {
      int gm = m0 + (i / 4) * SM_STRIDE + ty * 4 + (i % 4);
      if (gm >= p.M) continue;
#pragma unroll
      for (int j = 0; j < TN; j++) {
        int gn = n0 + (j / 4) * SN_STRIDE + tx * 4 + (j % 4);
        if (gn >= p.N) continue;
        float* dst = Cb + (long)gm * sCm + (long)gn * sCn;
        *dst = acc[i][j] + (acc_out ? *dst : 0.f);
      }
    }
  }
}

}  // namespace
#endif  // FRN_DEBUG_ENGINES
// C[m][n] (+)= sum_z part[z][m][n]
__global__ void splitk_reduce_kernel(const float* __restrict__ part, int nsplit, int M, int N,
                                     float* __restrict__ C, long sCm, long sCn, int accumulate) {
  long total = (long)M * N;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total;
       i += (long)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int z = 0; z < nsplit; z++) s += part[(long)z * total + i];
    int m = (int)(i / N), n = (int)(i % N);
    float* dst = C + (long)m * sCm + (long)n * sCn;
    *dst = s + (accumulate ? *dst : 0.f);
  }
}

#ifdef FRN_DEBUG_ENGINES
namespace {
struct Plan { int bm, bn, ktiles, nsplit, tiles_per_split; };
constexpr int kBK = 16;

Plan make_plan(const GemmDesc& d) {
  Plan pl;
  pl.bn = 128;
  pl.bm = d.M > 64 ? 128 : (d.M > 32 ? 64 : 32);
  pl.ktiles = cdiv(d.K, kBK);
  pl.nsplit = 1;
  pl.tiles_per_split = pl.ktiles;
  if (d.reduce_batch) {
    long T = (long)d.batch * pl.ktiles;
    long ctas = (long)cdiv(d.M, pl.bm) * cdiv(d.N, pl.bn);
    long want = cdiv(2 * 148, ctas);            // fill the 148 SMs twice over
    long min_split = cdiv(T, 256);              // <= 4096 sequential fp32 adds per partial
    long max_split = T / 8 > 0 ? T / 8 : 1;     // >= 128 k per split
    long ns = want < min_split ? min_split : want;
    if (ns > max_split) ns = max_split;
    if (ns < min_split) ns = min_split;
    if (ns < 1) ns = 1;
    pl.tiles_per_split = cdiv(T, ns);
    pl.nsplit = cdiv(T, pl.tiles_per_split);
  }
  return pl;
}

template <int EPI>
int launch(cudaStream_t st, const GemmDesc& d, const Plan& pl, float* aux) {
  dim3 grid(cdiv(d.N, pl.bn), cdiv(d.M, pl.bm), d.reduce_batch ? pl.nsplit : d.batch);
  if (pl.bm == 128) {
    sgemm_kernel<128, 128, kBK, 8, 8, EPI><<<grid, 256, 0, st>>>(d, aux, pl.ktiles, pl.tiles_per_split, pl.nsplit);
  } else if (pl.bm == 64) {
    sgemm_kernel<64, 128, kBK, 4, 8, EPI><<<grid, 256, 0, st>>>(d, aux, pl.ktiles, pl.tiles_per_split, pl.nsplit);
  } else {
    sgemm_kernel<32, 128, kBK, 4, 8, EPI><<<grid, 128, 0, st>>>(d, aux, pl.ktiles, pl.tiles_per_split, pl.nsplit);
  }
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // namespace

size_t ffma_gemm_workspace_bytes(const GemmDesc& d) {
  if (!d.reduce_batch) return 0;
  Plan pl = make_plan(d);
  return pl.nsplit > 1 ? align_up((size_t)pl.nsplit * d.M * d.N * sizeof(float)) : 0;
}

int ffma_gemm(cudaStream_t st, const GemmDesc& d, void* ws, size_t ws_bytes) {
  if (d.M <= 0 || d.N <= 0 || d.K <= 0 || d.batch <= 0) return FRN_ERR_BAD_SHAPE;
  if ((d.sAm != 1 && d.sAk != 1) || (d.sBn != 1 && d.sBk != 1)) return FRN_ERR_UNSUPPORTED;
  Plan pl = make_plan(d);
  ScopedTimer tm(st, FRN_TIME_GEMM);
  ShapeTimer sh(st, "ffma", d.M, d.N, d.K, d.batch,
                (d.reduce_batch ? 1 : 0) | (d.accumulate ? 2 : 0) | (d.sAk == 1 ? 4 : 0) | (d.sBn == 1 ? 8 : 0));
  if (d.reduce_batch && pl.nsplit > 1) {
    size_t need = (size_t)pl.nsplit * d.M * d.N * sizeof(float);
    if (!ws || ws_bytes < need) return FRN_ERR_WORKSPACE;
    FRN_TRY(launch<GEMM_EPI_STORE>(st, d, pl, (float*)ws));
    long total = (long)d.M * d.N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    splitk_reduce_kernel<<<blocks, 256, 0, st>>>((const float*)ws, pl.nsplit, d.M, d.N, d.C, d.sCm,
                                                 d.sCn, d.accumulate);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  return launch<GEMM_EPI_STORE>(st, d, pl, nullptr);
}

#else
size_t ffma_gemm_workspace_bytes(const GemmDesc&) { return 0; }
int ffma_gemm(cudaStream_t, const GemmDesc&, void*, size_t) { return FRN_ERR_UNSUPPORTED; }
#endif  // FRN_DEBUG_ENGINES

int gemm_rowsumsq_parts(int N) { return cdiv(N, 128); }

#ifdef FRN_DEBUG_ENGINES
int ffma_gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial) {
  if (d.M <= 0 || d.N <= 0 || d.K <= 0 || d.batch <= 0 || d.reduce_batch) return FRN_ERR_BAD_SHAPE;
  Plan pl = make_plan(d);
  ScopedTimer tm(st, FRN_TIME_GEMM);
  return launch<GEMM_EPI_ROWSUMSQ>(st, d, pl, out_partial);
}

#else
int ffma_gemm_rowsumsq(cudaStream_t, const GemmDesc&, float*) { return FRN_ERR_UNSUPPORTED; }
#endif

// ---- engine dispatch ---------------------------------------------------------------
namespace {
std::atomic<int> g_engine{-1};
}
int gemm_engine() {
#ifndef FRN_DEBUG_ENGINES
  return 1;        // product build: tcgen05 only
#endif
  int e = g_engine.load(std::memory_order_relaxed);
  if (e < 0) {
    const char* s = getenv("FRN_GEMM_ENGINE");
    e = (s && (s[0] == 'f' || s[0] == 'F' || s[0] == '0')) ? 0 : 1;   // default: tcgen05
    g_engine.store(e);
  }
  return e;
}
void set_gemm_engine(int e) { g_engine.store(e ? 1 : 0); }

namespace {
std::atomic<int> g_attn_engine{-1};
}
int attn_engine() {
#ifndef FRN_DEBUG_ENGINES
  return 1;
#endif
  int e = g_attn_engine.load(std::memory_order_relaxed);
  if (e < 0) {
    const char* s = getenv("FRN_ATTN_ENGINE");
    e = (s && (s[0] == 'f' || s[0] == 'F' || s[0] == '0')) ? 0 : 1;   // default: tcgen05
    g_attn_engine.store(e);
  }
  return e;
}
void set_attn_engine(int e) { g_attn_engine.store(e ? 1 : 0); }

size_t gemm_workspace_bytes(const GemmDesc& d) {
  // the larger of the two so that a workspace sized once serves either engine
  size_t a = ffma_gemm_workspace_bytes(d), b = tc_gemm_workspace_bytes(d);
  return a > b ? a : b;
}
int gemm(cudaStream_t st, const GemmDesc& d, void* ws, size_t ws_bytes) {
  return gemm_engine() ? tc_gemm(st, d, ws, ws_bytes) : ffma_gemm(st, d, ws, ws_bytes);
}
int gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial) {
  return gemm_engine() ? tc_gemm_rowsumsq(st, d, out_partial) : ffma_gemm_rowsumsq(st, d, out_partial);
}

}  // namespace frn
