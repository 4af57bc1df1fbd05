// tcgen05 (5th-gen tensor core) version of the strided batched GEMM in gemm.cu, with
// split-precision operands: every fp32 operand x is staged in shared memory as a tf32 pair
// (hi, lo) and the product is accumulated as lo*hi' + hi*lo' + hi*hi' in an fp32 TMEM
// accumulator (3xTF32), which keeps the fp32 parity gate (<= 1e-5 vs fp64) that single-pass
// TF32 (~1e-3) cannot meet.
//
// CTA = 128 x BN output tile.  Warps 0-3 stage operand tiles (global -> registers -> hi/lo split
// -> SWIZZLE_128B K-major shared memory; arbitrary element strides, so every transposed use is
// free), warp 4 lane 0 issues tcgen05.mma (M=128, N=BN, K=8 per instruction, 12 per 32-deep
// stage) and commits stage release / accumulator-ready to mbarriers, then warps 0-3 read the
// accumulator back with tcgen05.ld (thread = output row) for the epilogue.
#include "common.cuh"
#include "tc05.cuh"

namespace frn {
namespace {

using namespace tc;

constexpr int TBM = 128;      // UMMA M
constexpr int TBK = 32;       // 32 tf32 = one 128-byte swizzle row
constexpr int kLoaders = 128;
constexpr int kTcThreads = 160;

template <int BN> struct TcCfg {
  static constexpr int kStages = BN <= 64 ? 4 : 3;
  static constexpr int kABytes = TBM * TBK * 4;            // one of hi / lo
  static constexpr int kBBytes = BN * TBK * 4;
  static constexpr int kStageBytes = 2 * kABytes + 2 * kBBytes;
  static constexpr int kSmem = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr int kTmemCols = 2 * BN;               // [0, BN): hi*hi ; [BN, 2BN): lo*hi + hi*lo
};

// Stage a [ROWS x 32] K-major tile: element (r, k) = g[r*rs + k*ks], zero outside (rows, kvalid).
template <int ROWS>
__device__ __forceinline__ void stage_tile(float* __restrict__ hi, float* __restrict__ lo,
                                           const float* __restrict__ g, long rs, long ks, int rows, int kvalid,
                                           int tid) {
  constexpr int CHUNKS = ROWS * 8;
  const bool kfast = (ks == 1);
  const bool vec = kfast && ((rs & 3) == 0) && ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
#pragma unroll 4
  for (int idx = tid; idx < CHUNKS; idx += kLoaders) {
    int r, c;
    if (kfast) { c = idx & 7; r = idx >> 3; } else { r = idx % ROWS; c = idx / ROWS; }
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (r < rows) {
      const float* p = g + (long)r * rs + (long)(c * 4) * ks;
      if (vec && c * 4 + 3 < kvalid) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
      } else {
#pragma unroll
        for (int q = 0; q < 4; q++)
          if (c * 4 + q < kvalid) v[q] = __ldg(p + (long)q * ks);
      }
    }
    float h[4], l[4];
#pragma unroll
    for (int q = 0; q < 4; q++) split_tf32(v[q], h[q], l[q]);
    int o = sw128_chunk_offset(r, c);
    *reinterpret_cast<float4*>(hi + o) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(lo + o) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

template <int BN, int EPI>
__global__ void __launch_bounds__(kTcThreads, 1)
tc_gemm_kernel(GemmDesc p, float* __restrict__ aux, int ktiles, int tiles_per_split, int nsplit) {
  using Cfg = TcCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty = full + Cfg::kStages;
  uint64_t* acc_ready = empty + Cfg::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.y * TBM, n0 = blockIdx.x * BN;
  const int z = blockIdx.z;
  int t_begin, t_end, b_fixed;
  if (p.reduce_batch) {
    t_begin = z * tiles_per_split;
    t_end = min(t_begin + tiles_per_split, p.batch * ktiles);
    b_fixed = -1;
  } else {
    t_begin = 0; t_end = ktiles; b_fixed = z;
  }

  if (tid == 0) {
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&full[s], kLoaders); mbar_init(&empty[s], 1); }
    mbar_init(acc_ready, 1);
    mbar_init_fence();
  }
  if (warp == 4) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp < 4) {
    // ===== operand staging =====
    int stage = 0; uint32_t phase = 0;
    for (int t = t_begin; t < t_end; t++) {
      int b = b_fixed >= 0 ? b_fixed : t / ktiles;
      int k0 = (b_fixed >= 0 ? t : t % ktiles) * TBK;
      mbar_wait(&empty[stage], phase ^ 1);
      float* sA = reinterpret_cast<float*>(smem + stage * Cfg::kStageBytes);
      float* sAl = sA + TBM * TBK;
      float* sB = sAl + TBM * TBK;
      float* sBl = sB + BN * TBK;
      stage_tile<TBM>(sA, sAl, p.A + (long)b * p.sAb + (long)m0 * p.sAm + (long)k0 * p.sAk, p.sAm, p.sAk,
                      p.M - m0, p.K - k0, tid);
      stage_tile<BN>(sB, sBl, p.B + (long)b * p.sBb + (long)n0 * p.sBn + (long)k0 * p.sBk, p.sBn, p.sBk,
                     p.N - n0, p.K - k0, tid);
      fence_proxy_async_smem();
      mbar_arrive(&full[stage]);
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (lane == 0) {
    // ===== MMA issue (one thread) =====
    constexpr uint32_t idesc = umma_idesc_tf32(TBM, BN);
    int stage = 0; uint32_t phase = 0;
    bool acc = false;
    for (int t = t_begin; t < t_end; t++) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      uint32_t base = smem_u32(smem + stage * Cfg::kStageBytes);
      uint64_t a_hi = umma_desc_sw128(base), a_lo = umma_desc_sw128(base + Cfg::kABytes);
      uint64_t b_hi = umma_desc_sw128(base + 2 * Cfg::kABytes);
      uint64_t b_lo = umma_desc_sw128(base + 2 * Cfg::kABytes + Cfg::kBBytes);
#pragma unroll
      for (int k = 0; k < TBK / 8; k++) {   // +32 bytes per K = 8 slice: start-address field += 2
        // The tensor core truncates when it aligns addends to the accumulator, so the two
        // correction terms (2^-11 of the main term) get their own accumulator (columns [BN, 2BN)).
        umma_tf32_ss(tmem_acc + BN, a_lo + 2 * k, b_hi + 2 * k, idesc, acc);
        umma_tf32_ss(tmem_acc + BN, a_hi + 2 * k, b_lo + 2 * k, idesc, true);
        umma_tf32_ss(tmem_acc, a_hi + 2 * k, b_hi + 2 * k, idesc, acc);
        acc = true;
      }
      umma_commit(&empty[stage]);            // frees the stage once these MMAs have read it
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
    umma_commit(acc_ready);
  }

  // ===== epilogue: warps 0-3, thread = output row =====
  if (warp < 4) {
    const bool any = t_begin < t_end;
    if (any) {
      mbar_wait(acc_ready, 0);
      tc_fence_after();
    }
    const int row = warp * 32 + lane, gm = m0 + row;
    const uint32_t taddr = tmem_acc + ((uint32_t)(warp * 32) << 16);
    const bool partial = p.reduce_batch && nsplit > 1;
    if (EPI == GEMM_EPI_ROWSUMSQ) {
      float s = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 16) {
        float v[16], w[16];
        if (any) { tmem_ld16(taddr + c0, v); tmem_ld16(taddr + BN + c0, w); tmem_wait_ld(); }
#pragma unroll
        for (int j = 0; j < 16; j++) {
          float x = any ? p.alpha * (v[j] + w[j]) : 0.f;
          s += (n0 + c0 + j < p.N) ? x * x : 0.f;
        }
      }
      if (gm < p.M) aux[((long)z * p.M + gm) * gridDim.x + blockIdx.x] = s;
    } else {
      float* Cb; long sCm, sCn;
      if (partial) { Cb = aux + (long)z * p.M * p.N; sCm = p.N; sCn = 1; }
      else { Cb = p.C + (p.reduce_batch ? 0 : (long)z * p.sCb); sCm = p.sCm; sCn = p.sCn; }
      const bool acc_out = !partial && p.accumulate;
      const long boff = p.reduce_batch ? 0 : (long)z * p.bias_bs;
      const float bm = (!partial && p.bias_m && gm < p.M) ? __ldg(p.bias_m + boff + gm) : 0.f;
      const bool vec = (sCn == 1) && ((sCm & 3) == 0) && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0);
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 16) {
        float v[16], w[16];
        if (any) {
          tmem_ld16(taddr + c0, v); tmem_ld16(taddr + BN + c0, w); tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; j++) v[j] += w[j];
        } else {
#pragma unroll
          for (int j = 0; j < 16; j++) v[j] = 0.f;
        }
        if (gm >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 16; j++) {
          int gn = n0 + c0 + j;
          float bn = (!partial && p.bias_n && gn < p.N) ? __ldg(p.bias_n + boff + gn) : 0.f;
          v[j] = fmaf(p.alpha, v[j], bm + bn);
        }
        float* dst = Cb + (long)gm * sCm + (long)(n0 + c0) * sCn;
        if (vec && n0 + c0 + 15 < p.N) {
#pragma unroll
          for (int q = 0; q < 4; q++) {
            float4 o = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            float4* d4 = reinterpret_cast<float4*>(dst) + q;
            if (acc_out) { float4 c = *d4; o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w; }
            *d4 = o;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; j++)
            if (n0 + c0 + j < p.N) {
              float* d1 = dst + (long)j * sCn;
              *d1 = v[j] + (acc_out ? *d1 : 0.f);
            }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<Cfg::kTmemCols>(tmem_acc);
}


struct TcPlan { int bn, ktiles, nsplit, tiles_per_split; bool swapped; GemmDesc d; };

// The UMMA M dimension is fixed at 128: problems with few rows and many columns (the C/8-row
// projections) are run as C^T = B^T A^T, which is only a relabelling of strides.
TcPlan tc_plan(const GemmDesc& in, bool allow_swap) {
  TcPlan pl;
  pl.d = in;
  pl.swapped = false;
  if (allow_swap && in.M <= 64 && in.N > in.M) {
    GemmDesc& d = pl.d;
    d.A = in.B; d.sAm = in.sBn; d.sAk = in.sBk; d.sAb = in.sBb;
    d.B = in.A; d.sBn = in.sAm; d.sBk = in.sAk; d.sBb = in.sAb;
    d.M = in.N; d.N = in.M;
    d.sCm = in.sCn; d.sCn = in.sCm;
    d.bias_m = in.bias_n; d.bias_n = in.bias_m;
    pl.swapped = true;
  }
  const GemmDesc& d = pl.d;
  pl.bn = d.N <= 32 ? 32 : (d.N <= 64 ? 64 : 128);
  pl.ktiles = cdiv(d.K, TBK);
  pl.nsplit = 1;
  pl.tiles_per_split = pl.ktiles;
  if (d.reduce_batch) {
    long T = (long)d.batch * pl.ktiles;
    long ctas = (long)cdiv(d.M, TBM) * cdiv(d.N, pl.bn);
    long want = cdiv(2 * 148, ctas);
    long min_split = cdiv(T, 128);              // <= 4096 k per partial
    long max_split = T / 4 > 0 ? T / 4 : 1;     // >= 128 k per split
    long ns = want < min_split ? min_split : want;
    if (ns > max_split) ns = max_split;
    if (ns < min_split) ns = min_split;
    if (ns < 1) ns = 1;
    pl.tiles_per_split = cdiv(T, ns);
    pl.nsplit = cdiv(T, pl.tiles_per_split);
  }
  return pl;
}

template <int BN, int EPI>
int tc_launch_bn(cudaStream_t st, const TcPlan& pl, float* aux) {
  using Cfg = TcCfg<BN>;
  auto kern = tc_gemm_kernel<BN, EPI>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem);
  if (e != cudaSuccess) return (int)e;
  const GemmDesc& d = pl.d;
  dim3 grid(cdiv(d.N, BN), cdiv(d.M, TBM), d.reduce_batch ? pl.nsplit : d.batch);
  kern<<<grid, kTcThreads, Cfg::kSmem, st>>>(d, aux, pl.ktiles, pl.tiles_per_split, pl.nsplit);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

template <int EPI>
int tc_launch(cudaStream_t st, const TcPlan& pl, float* aux) {
  if (pl.bn == 32) return tc_launch_bn<32, EPI>(st, pl, aux);
  if (pl.bn == 64) return tc_launch_bn<64, EPI>(st, pl, aux);
  return tc_launch_bn<128, EPI>(st, pl, aux);
}

}  // namespace

size_t tc_gemm_workspace_bytes(const GemmDesc& d) {
  if (!d.reduce_batch) return 0;
  TcPlan pl = tc_plan(d, true);
  return pl.nsplit > 1 ? align_up((size_t)pl.nsplit * d.M * d.N * sizeof(float)) : 0;
}

int tc_gemm(cudaStream_t st, const GemmDesc& in, void* ws, size_t ws_bytes) {
  if (in.M <= 0 || in.N <= 0 || in.K <= 0 || in.batch <= 0) return FRN_ERR_BAD_SHAPE;
  if ((in.sAm != 1 && in.sAk != 1) || (in.sBn != 1 && in.sBk != 1)) return FRN_ERR_UNSUPPORTED;
  TcPlan pl = tc_plan(in, true);
  ScopedTimer tm(st, FRN_TIME_GEMM);
  if (in.reduce_batch && pl.nsplit > 1) {
    size_t need = (size_t)pl.nsplit * in.M * in.N * sizeof(float);
    if (!ws || ws_bytes < need) return FRN_ERR_WORKSPACE;
    FRN_TRY(tc_launch<GEMM_EPI_STORE>(st, pl, (float*)ws));
    const GemmDesc& d = pl.d;     // partials are (nsplit, d.M, d.N) in the (possibly swapped) frame
    long total = (long)d.M * d.N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    splitk_reduce_kernel<<<blocks, 256, 0, st>>>((const float*)ws, pl.nsplit, d.M, d.N, d.C, d.sCm, d.sCn,
                                                 d.accumulate);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  return tc_launch<GEMM_EPI_STORE>(st, pl, nullptr);
}

int tc_gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial) {
  if (d.M <= 0 || d.N <= 0 || d.K <= 0 || d.batch <= 0 || d.reduce_batch) return FRN_ERR_BAD_SHAPE;
  TcPlan pl = tc_plan(d, false);
  pl.bn = 128;                     // gemm_rowsumsq_parts(N) == cdiv(N, 128) column tiles
  ScopedTimer tm(st, FRN_TIME_GEMM);
  return tc_launch_bn<128, GEMM_EPI_ROWSUMSQ>(st, pl, out_partial);
}

}  // namespace frn
