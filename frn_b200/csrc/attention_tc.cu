// tcgen05 / TMEM streaming-softmax attention for CrossAttentionBlock
// (retinanet/Attention.py:29-32), split-precision (3xTF32) so that fp32 parity holds.
//
// Forward, one CTA per 128 query rows, 64-key tiles, 9 warps:
//   warps 0-3  softmax: thread = query row.  Read S from TMEM, online softmax in registers,
//              write P (tf32 hi / lo) back to TMEM, promote the previous tile's P.V product
//              from TMEM into fp32 register accumulators (rescaled there, so Y in TMEM never
//              needs a correction pass and no accumulation chain exceeds 8 tensor-core steps:
//              tcgen05 truncates when it aligns addends, ~0.4 ulp bias per step, which over
//              N/8 = 2400 steps would cost 6e-5).
//   warp  4    one thread issues tcgen05.mma: S = Q.K (A, B from shared memory) and
//              Y_tile = P.V (A = P from TMEM, B = V from shared memory).
//   warps 5-8  operand staging: global (channel-major d x N, as the projection GEMM wrote it)
//              -> tf32 hi/lo split -> SWIZZLE_128B K-major shared memory tiles.
// The 2^-11-sized correction products (lo.hi + hi.lo) accumulate in their own TMEM columns.
#include "common.cuh"
#include "tc05.cuh"

namespace frn {
namespace {

using namespace tc;

constexpr float kLog2e = 1.4426950408889634f;
constexpr int ABM = 128;   // query rows per CTA (UMMA M)
constexpr int ABN = 64;    // keys per tile
constexpr int kSoftmaxThreads = 128, kLoaderThreads = 128, kAttnThreads = 288;

// Stage a [ROWS x 32] K-major SWIZZLE_128B tile pair (hi, lo); element (r, k) = g[r*rs + k*ks].
template <int ROWS>
__device__ __forceinline__ void stage_split(float* __restrict__ hi, float* __restrict__ lo,
                                            const float* __restrict__ g, long rs, long ks, int rows, int kvalid,
                                            int tid) {
  constexpr int CHUNKS = ROWS * 8;
  const bool kfast = (ks == 1);
  const bool vec = kfast && ((rs & 3) == 0) && ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
#pragma unroll 2
  for (int idx = tid; idx < CHUNKS; idx += kLoaderThreads) {
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

template <int DPAD> struct FwdCfg {
  static constexpr int kSub = DPAD / 32;                    // 32-channel sub-tiles
  static constexpr int kQBytes = 2 * kSub * ABM * 128;      // hi + lo
  static constexpr int kKBytes = 2 * kSub * ABN * 128;
  static constexpr int kVBytes = 2 * 2 * DPAD * 128;        // two 32-key sub-tiles, hi + lo
  static constexpr int kStageBytes = kKBytes + kVBytes;
  static constexpr int kStages = DPAD == 32 ? 3 : 2;
  static constexpr int kSmem = kQBytes + kStages * kStageBytes + 1024 + 256;
  // TMEM columns
  static constexpr int kSMain = 0, kSSmall = 128, kYMain = 256, kYSmall = 256 + 2 * DPAD;
};

template <int DPAD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fwd_tc_kernel(int N, int d, const float* __restrict__ Qc, const float* __restrict__ Kc,
                   const float* __restrict__ Vc, float* __restrict__ Y, float* __restrict__ lse) {
  using Cfg = FwdCfg<DPAD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + Cfg::kQBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* q_full = bars;                 // 1
  uint64_t* kv_full = bars + 1;            // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;
  uint64_t* s_full = kv_empty + Cfg::kStages;   // 2
  uint64_t* p_full = s_full + 2;                // 2
  uint64_t* y_full = p_full + 2;                // 2
  uint64_t* y_free = y_full + 2;                // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(y_free + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, i0 = blockIdx.x * ABM;
  const int T = (N + ABN - 1) / ABN;
  const long bo = (long)b * d * N;

  if (tid == 0) {
    mbar_init(q_full, kLoaderThreads);
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&kv_full[s], kLoaderThreads); mbar_init(&kv_empty[s], 1); }
    for (int s = 0; s < 2; s++) {
      mbar_init(&s_full[s], 1); mbar_init(&p_full[s], kSoftmaxThreads);
      mbar_init(&y_full[s], 1); mbar_init(&y_free[s], kSoftmaxThreads);
    }
    mbar_init_fence();
  }
  if (warp == 4) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp >= 5) {
    // ======================= operand staging =======================
    const int lt = tid - 5 * 32;
#pragma unroll
    for (int s = 0; s < Cfg::kSub; s++) {     // Q[i][c] = Qc[c*N + i]
      float* hi = reinterpret_cast<float*>(sQ + s * ABM * 128);
      float* lo = reinterpret_cast<float*>(sQ + (Cfg::kSub + s) * ABM * 128);
      stage_split<ABM>(hi, lo, Qc + bo + (long)(s * 32) * N + i0, 1, N, N - i0, d - s * 32, lt);
    }
    fence_proxy_async_smem();
    mbar_arrive(q_full);
    int stage = 0; uint32_t phase = 0;
    for (int j = 0; j < T; j++) {
      const int j0 = j * ABN;
      mbar_wait(&kv_empty[stage], phase ^ 1);
      uint8_t* st = sKV + stage * Cfg::kStageBytes;
#pragma unroll
      for (int s = 0; s < Cfg::kSub; s++) {   // K[j][c] = Kc[c*N + j]
        float* hi = reinterpret_cast<float*>(st + s * ABN * 128);
        float* lo = reinterpret_cast<float*>(st + (Cfg::kSub + s) * ABN * 128);
        stage_split<ABN>(hi, lo, Kc + bo + (long)(s * 32) * N + j0, 1, N, N - j0, d - s * 32, lt);
      }
#pragma unroll
      for (int s = 0; s < 2; s++) {           // V^T[c][j] = Vc[c*N + j], 32 keys per sub-tile
        float* hi = reinterpret_cast<float*>(st + Cfg::kKBytes + s * DPAD * 128);
        float* lo = reinterpret_cast<float*>(st + Cfg::kKBytes + (2 + s) * DPAD * 128);
        stage_split<DPAD>(hi, lo, Vc + bo + j0 + s * 32, N, 1, d, N - j0 - s * 32, lt);
      }
      fence_proxy_async_smem();
      mbar_arrive(&kv_full[stage]);
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 4) {
    // ======================= MMA issue =======================
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_tf32(ABM, ABN);
      constexpr uint32_t idesc_y = umma_idesc_tf32(ABM, DPAD);
      const uint32_t q_base = smem_u32(sQ);
      auto issue_pv = [&](int j) {
        const int buf = j & 1, stage = j % Cfg::kStages;
        mbar_wait(&p_full[buf], (j >> 1) & 1);
        if (j >= 2) mbar_wait(&y_free[buf], ((j - 2) >> 1) & 1);
        tc_fence_after();
        const uint32_t v_base = smem_u32(sKV + stage * Cfg::kStageBytes + Cfg::kKBytes);
        const uint32_t p_hi = tmem + Cfg::kSMain + buf * ABN, p_lo = tmem + Cfg::kSSmall + buf * ABN;
        const uint32_t y_main = tmem + Cfg::kYMain + buf * DPAD, y_small = tmem + Cfg::kYSmall + buf * DPAD;
#pragma unroll
        for (int s = 0; s < 2; s++) {
          uint64_t v_hi = umma_desc_sw128(v_base + s * DPAD * 128);
          uint64_t v_lo = umma_desc_sw128(v_base + (2 + s) * DPAD * 128);
#pragma unroll
          for (int k = 0; k < 4; k++) {
            const bool acc = (s | k) != 0;
            const uint32_t col = s * 32 + k * 8;
            umma_tf32_ts(y_small, p_lo + col, v_hi + 2 * k, idesc_y, acc);
            umma_tf32_ts(y_small, p_hi + col, v_lo + 2 * k, idesc_y, true);
            umma_tf32_ts(y_main, p_hi + col, v_hi + 2 * k, idesc_y, acc);
          }
        }
        umma_commit(&y_full[buf]);
        umma_commit(&kv_empty[stage]);
      };
      mbar_wait(q_full, 0);
      for (int j = 0; j < T; j++) {
        const int buf = j & 1, stage = j % Cfg::kStages;
        mbar_wait(&kv_full[stage], (j / Cfg::kStages) & 1);
        tc_fence_after();
        const uint32_t k_base = smem_u32(sKV + stage * Cfg::kStageBytes);
        const uint32_t s_main = tmem + Cfg::kSMain + buf * ABN, s_small = tmem + Cfg::kSSmall + buf * ABN;
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          uint64_t q_hi = umma_desc_sw128(q_base + s * ABM * 128);
          uint64_t q_lo = umma_desc_sw128(q_base + (Cfg::kSub + s) * ABM * 128);
          uint64_t k_hi = umma_desc_sw128(k_base + s * ABN * 128);
          uint64_t k_lo = umma_desc_sw128(k_base + (Cfg::kSub + s) * ABN * 128);
#pragma unroll
          for (int k = 0; k < 4; k++) {
            const bool acc = (s | k) != 0;
            umma_tf32_ss(s_small, q_lo + 2 * k, k_hi + 2 * k, idesc_s, acc);
            umma_tf32_ss(s_small, q_hi + 2 * k, k_lo + 2 * k, idesc_s, true);
            umma_tf32_ss(s_main, q_hi + 2 * k, k_hi + 2 * k, idesc_s, acc);
          }
        }
        umma_commit(&s_full[buf]);
        if (j >= 1) issue_pv(j - 1);
      }
      issue_pv(T - 1);
    }
  } else {
    // ======================= softmax + promotion (thread = query row) =======================
    const int row = tid;                               // 0..127
    const uint32_t lane_base = ((uint32_t)(warp * 32)) << 16;
    float m = -INFINITY, l = 0.f, corr_prev = 0.f;
    float y[DPAD];
#pragma unroll
    for (int c = 0; c < DPAD; c++) y[c] = 0.f;

    auto promote = [&](int j, float corr) {            // y = y * corr + Y_tile(j)
      const int buf = j & 1;
      mbar_wait(&y_full[buf], (j >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < DPAD; c0 += 16) {
        float a[16], s[16];
        tmem_ld16(tmem + lane_base + Cfg::kYMain + buf * DPAD + c0, a);
        tmem_ld16(tmem + lane_base + Cfg::kYSmall + buf * DPAD + c0, s);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 16; q++) y[c0 + q] = fmaf(y[c0 + q], corr, a[q] + s[q]);
      }
      tc_fence_before();
      mbar_arrive(&y_free[buf]);
    };

    for (int j = 0; j < T; j++) {
      const int buf = j & 1, j0 = j * ABN;
      mbar_wait(&s_full[buf], (j >> 1) & 1);
      tc_fence_after();
      float s[ABN];
#pragma unroll
      for (int c0 = 0; c0 < ABN; c0 += 16) {
        float a[16], w[16];
        tmem_ld16(tmem + lane_base + Cfg::kSMain + buf * ABN + c0, a);
        tmem_ld16(tmem + lane_base + Cfg::kSSmall + buf * ABN + c0, w);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 16; q++) s[c0 + q] = a[q] + w[q];
      }
      float mx = -INFINITY;
      if (j0 + ABN > N) {
#pragma unroll
        for (int q = 0; q < ABN; q++) if (j0 + q >= N) s[q] = -INFINITY;
      }
#pragma unroll
      for (int q = 0; q < ABN; q++) mx = fmaxf(mx, s[q]);
      const float m_new = fmaxf(m, mx);
      const float corr = exp2f((m - m_new) * kLog2e);
      float sum = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < ABN; c0 += 16) {
        float hi[16], lo[16];
#pragma unroll
        for (int q = 0; q < 16; q++) {
          float p = exp2f((s[c0 + q] - m_new) * kLog2e);
          sum += p;
          split_tf32(p, hi[q], lo[q]);
        }
        tmem_st16(tmem + lane_base + Cfg::kSMain + buf * ABN + c0, hi);
        tmem_st16(tmem + lane_base + Cfg::kSSmall + buf * ABN + c0, lo);
      }
      l = fmaf(l, corr, sum);
      m = m_new;
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&p_full[buf]);
      if (j >= 1) promote(j - 1, corr_prev);
      corr_prev = corr;
    }
    promote(T - 1, corr_prev);

    const int gi = i0 + row;
    if (gi < N) {
      const float inv = 1.f / l;
      lse[(long)b * N + gi] = m + logf(l);
      float* dst = Y + ((long)b * N + gi) * d;
      if ((d & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
        for (int c = 0; c < DPAD; c += 4)
          if (c < d) *reinterpret_cast<float4*>(dst + c) = make_float4(y[c] * inv, y[c + 1] * inv, y[c + 2] * inv, y[c + 3] * inv);
      } else {
#pragma unroll
        for (int c = 0; c < DPAD; c++)
          if (c < d) dst[c] = y[c] * inv;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<512>(tmem);
}

template <int DPAD>
int launch_fwd_tc(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Kc, const float* Vc,
                  float* Y, float* lse) {
  using Cfg = FwdCfg<DPAD>;
  auto kern = attn_fwd_tc_kernel<DPAD>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem);
  if (e != cudaSuccess) return (int)e;
  dim3 grid(cdiv(N, ABM), B);
  ScopedTimer tm(st, FRN_TIME_ATTN_FWD);
  kern<<<grid, kAttnThreads, Cfg::kSmem, st>>>(N, d, Qc, Kc, Vc, Y, lse);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // namespace

int attn_forward_tc(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Kc, const float* Vc,
                    float* Y, float* lse) {
  if (d <= 32) return launch_fwd_tc<32>(st, B, N, d, Qc, Kc, Vc, Y, lse);
  if (d <= 64) return launch_fwd_tc<64>(st, B, N, d, Qc, Kc, Vc, Y, lse);
  return FRN_ERR_UNSUPPORTED;
}

}  // namespace frn
