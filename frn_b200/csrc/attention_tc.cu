// tcgen05 / TMEM streaming-softmax attention for CrossAttentionBlock
// (retinanet/Attention.py:29-32) and its backward, split-precision (3xTF32) so that fp32 parity
// holds: every fp32 operand x is carried as a tf32 pair (hi, lo), and each product is accumulated
// as hi*hi' (main accumulator) and lo*hi' + hi*lo' (a second, 2^-11-sized accumulator) in fp32
// TMEM columns.
//
// Operands reach shared memory by bulk async copies (cp.async.bulk, one elected thread, mbarrier
// transaction counts) from PACKED SPLIT IMAGES: a small pre-pass (pack_split_kernel) writes each
// (B, d, N) operand as tf32 hi / lo planes that already have the SWIZZLE_128B K-major tile layout
// the UMMA descriptors expect, zero padded to 128 positions, so a tile is one contiguous chunk:
//   pos image   rows = positions, 128-byte rows of 32 channels   (Q.K^T style products)
//   chan image  rows = channels, 128-byte rows of 32 positions   (P.V style products)
//
// tcgen05 truncates (does not round) when it aligns addends to the accumulator, a relative bias
// of ~3e-9 per K = 8 step (tools/tc_error_vs_k.py).  No accumulation chain here is longer than 8
// steps: every tile product is promoted from TMEM into fp32 register accumulators.
//
// Kernels (d <= 64; larger d goes through the GEMM engine, see fusion.cu).  All of them: CTA = one tile of 128
// rows (= TMEM lanes), the resident operand of that tile held in TMEM as tf32 hi / lo and used as the A operand
// (TS form: a tcgen05.mma then costs N / 2 cycles instead of 32 + N / 4, tools/mma_cost_probe.cu), the other
// operands streamed through shared memory, score accumulators double buffered in TMEM, separate issuing
// warps for the score products and for the output products, one bulk-copy producer lane.
//   attn_fwd_tc_kernel    128 queries, 64-key tiles, softmax groups on alternate key tiles (merged at the end)
//   attn_dq2_tc_kernel    128 queries, 32-key tiles: S, dP -> dS -> dQ           (eight compute warps)
//   attn_dkv2_tc_kernel   128 keys, 32-query tiles: S^T, dP^T -> P^T, dS^T -> dV, dK   (d <= 32)
//   attn_dkv3_tc_kernel   the same for 32 < d <= 64 (K in TMEM, V in shared memory: 512 columns are full)
//   attn_dq_tc_kernel / attn_dkv_tc_kernel   the first (ping-pong, single issuing warp) designs, kept behind
//                         FRN_DQ_DESIGN=1 / FRN_DKV_DESIGN=1 for comparison
// No atomics anywhere: the backward is deterministic.
#include <mutex>
#include <cstdlib>

#include "common.cuh"
#include "tc05.cuh"

namespace frn {
namespace {

using namespace tc;

constexpr float kLog2e = 1.4426950408889634f;

// ------------------------------------------------------------------------------------------------
// packed split images
// ------------------------------------------------------------------------------------------------
// X element (c, n) of sample b at X[b*bs + c*cs + n*ns]; sub (optional): per-(b, c) value subtracted
// first (the key operand is centred over positions, see attn_tc_pack).  grid (NP/32, B), 256 threads.
template <int DPAD>
__global__ void __launch_bounds__(256)
pack_split_kernel(const float* __restrict__ X, const float* __restrict__ sub, long bs, long cs, long ns, int d,
                  int N, int NP, float* __restrict__ pos_hi, float* __restrict__ pos_lo, float* __restrict__ chan_hi,
                  float* __restrict__ chan_lo) {
  constexpr int kSub = DPAD / 32;
  __shared__ float tile[DPAD][33];
  const int b = blockIdx.y, n0 = blockIdx.x * 32, tid = threadIdx.x;
  const float* Xb = X + (long)b * bs;
  for (int idx = tid; idx < DPAD * 32; idx += 256) {
    int c, nl;
    if (ns == 1) { c = idx >> 5; nl = idx & 31; } else { c = idx % DPAD; nl = idx / DPAD; }
    float v = 0.f;
    if (c < d && n0 + nl < N) {
      v = __ldg(Xb + (long)c * cs + (long)(n0 + nl) * ns);
      if (sub) v -= __ldg(sub + (long)b * d + c);
    }
    tile[c][nl] = v;
  }
  __syncthreads();
  if (pos_hi) {
    for (int idx = tid; idx < kSub * 32 * 8; idx += 256) {
      const int chunk = idx & 7, nl = (idx >> 3) & 31, s = idx >> 8;
      const int n = n0 + nl;
      float h[4], l[4];
#pragma unroll
      for (int q = 0; q < 4; q++) split_tf32(tile[s * 32 + chunk * 4 + q][nl], h[q], l[q]);
      const long o = (((long)b * kSub + s) * NP + n) * 32 + ((chunk ^ (n & 7)) << 2);
      *reinterpret_cast<float4*>(pos_hi + o) = make_float4(h[0], h[1], h[2], h[3]);
      if (pos_lo) *reinterpret_cast<float4*>(pos_lo + o) = make_float4(l[0], l[1], l[2], l[3]);
    }
  }
  if (chan_hi) {
    for (int idx = tid; idx < DPAD * 8; idx += 256) {
      const int chunk = idx & 7, c = idx >> 3;
      float h[4], l[4];
#pragma unroll
      for (int q = 0; q < 4; q++) split_tf32(tile[c][chunk * 4 + q], h[q], l[q]);
      const long o = (((long)b * (NP / 32) + blockIdx.x) * DPAD + c) * 32 + ((chunk ^ (c & 7)) << 2);
      *reinterpret_cast<float4*>(chan_hi + o) = make_float4(h[0], h[1], h[2], h[3]);
      if (chan_lo) *reinterpret_cast<float4*>(chan_lo + o) = make_float4(l[0], l[1], l[2], l[3]);
    }
  }
}

// delta[b][n] = sum_k dY[b][n][k] * Y[b][n][k] and a copy of lse, both padded to NP positions
// (lse = +inf, delta = 0 beyond N, so padded queries get P = exp(S - inf) = 0).
__global__ void attn_bwd_prep_kernel(int N, int NP, int d, const float* __restrict__ dYp,
                                     const float* __restrict__ Y, const float* __restrict__ lse,
                                     float* __restrict__ lse_pad, float* __restrict__ delta_pad) {
  const int b = blockIdx.y;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= NP) return;
  float l = INFINITY, dl = 0.f;
  if (n < N) {
    const float* a = dYp + ((long)b * N + n) * d;
    const float* y = Y + ((long)b * N + n) * d;
    l = lse[(long)b * N + n];
    if ((d & 3) == 0) {
      for (int k = 0; k < d; k += 4) {
        float4 u = __ldg(reinterpret_cast<const float4*>(a + k));
        float4 v = __ldg(reinterpret_cast<const float4*>(y + k));
        dl = fmaf(u.x, v.x, dl); dl = fmaf(u.y, v.y, dl); dl = fmaf(u.z, v.z, dl); dl = fmaf(u.w, v.w, dl);
      }
    } else {
      for (int k = 0; k < d; k++) dl = fmaf(__ldg(a + k), __ldg(y + k), dl);
    }
  }
  lse_pad[(long)b * NP + n] = l;
  delta_pad[(long)b * NP + n] = dl;
}

// 2^x on the special-function unit alone (ex2.approx.ftz: 2 ulp; exp2f() wraps the same instruction in a
// denormal-range rescale of four more instructions, and the softmax / dS warps -- one or two per scheduler --
// are what bounds these kernels, not the tensor pipe).  Arguments here are <= 0; -inf gives 0.
__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 3xTF32 products of one 32-deep K sub-tile (4 K = 8 slices of 32 bytes each) into ONE fp32
// accumulator.  The 2^-11-sized correction terms a_lo.b_hi + a_hi.b_lo of ALL sub-tiles are issued
// first and the main terms a_hi.b_hi after them: the tensor core truncates the smaller addend when
// it aligns it to the accumulator, so corrections added to an already large accumulator would lose
// their low bits at every step, whereas in this order they are truncated once (when the first main
// term arrives) -- the same accuracy as a separate accumulator at half the TMEM read traffic
// (tcgen05.ld bandwidth is what bounds these kernels).
__device__ __forceinline__ void mma_corr_ss(uint32_t dst, uint32_t a_hi, uint32_t a_lo, uint32_t b_hi,
                                            uint32_t b_lo, uint32_t idesc, bool first) {
  const uint64_t ah = umma_desc_sw128(a_hi), al = umma_desc_sw128(a_lo);
  const uint64_t bh = umma_desc_sw128(b_hi), bl = umma_desc_sw128(b_lo);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    umma_tf32_ss(dst, al + 2 * k, bh + 2 * k, idesc, !(first && k == 0));
    umma_tf32_ss(dst, ah + 2 * k, bl + 2 * k, idesc, true);
  }
}
__device__ __forceinline__ void mma_main_ss(uint32_t dst, uint32_t a_hi, uint32_t b_hi, uint32_t idesc) {
  const uint64_t ah = umma_desc_sw128(a_hi), bh = umma_desc_sw128(b_hi);
#pragma unroll
  for (int k = 0; k < 4; k++) umma_tf32_ss(dst, ah + 2 * k, bh + 2 * k, idesc, true);
}
// Same with A = 32 TMEM columns (hi at a_hi, lo at a_lo; lane = row), B one 32-deep sub-tile.
__device__ __forceinline__ void mma_corr_ts(uint32_t dst, uint32_t a_hi, uint32_t a_lo, uint32_t b_hi,
                                            uint32_t b_lo, uint32_t idesc, bool first) {
  const uint64_t bh = umma_desc_sw128(b_hi), bl = umma_desc_sw128(b_lo);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    umma_tf32_ts(dst, a_lo + 8 * k, bh + 2 * k, idesc, !(first && k == 0));
    umma_tf32_ts(dst, a_hi + 8 * k, bl + 2 * k, idesc, true);
  }
}
__device__ __forceinline__ void mma_main_ts(uint32_t dst, uint32_t a_hi, uint32_t b_hi, uint32_t idesc,
                                            bool first = false) {
  const uint64_t bh = umma_desc_sw128(b_hi);
#pragma unroll
  for (int k = 0; k < 4; k++) umma_tf32_ts(dst, a_hi + 8 * k, bh + 2 * k, idesc, !(first && k == 0));
}

// Two independent products issued interleaved (different accumulators), so that consecutive
// tcgen05.mma never form a read-after-write chain on one accumulator.
__device__ __forceinline__ void mma2_corr_ss(uint32_t d0, uint32_t a0_hi, uint32_t a0_lo, uint32_t b0_hi, uint32_t b0_lo,
                                             uint32_t d1, uint32_t a1_hi, uint32_t a1_lo, uint32_t b1_hi, uint32_t b1_lo,
                                             uint32_t idesc, bool first) {
  const uint64_t ah0 = umma_desc_sw128(a0_hi), al0 = umma_desc_sw128(a0_lo), bh0 = umma_desc_sw128(b0_hi),
                 bl0 = umma_desc_sw128(b0_lo);
  const uint64_t ah1 = umma_desc_sw128(a1_hi), al1 = umma_desc_sw128(a1_lo), bh1 = umma_desc_sw128(b1_hi),
                 bl1 = umma_desc_sw128(b1_lo);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    umma_tf32_ss(d0, al0 + 2 * k, bh0 + 2 * k, idesc, !(first && k == 0));
    umma_tf32_ss(d1, al1 + 2 * k, bh1 + 2 * k, idesc, !(first && k == 0));
    umma_tf32_ss(d0, ah0 + 2 * k, bl0 + 2 * k, idesc, true);
    umma_tf32_ss(d1, ah1 + 2 * k, bl1 + 2 * k, idesc, true);
  }
}
__device__ __forceinline__ void mma2_main_ss(uint32_t d0, uint32_t a0_hi, uint32_t b0_hi, uint32_t d1, uint32_t a1_hi,
                                             uint32_t b1_hi, uint32_t idesc) {
  const uint64_t ah0 = umma_desc_sw128(a0_hi), bh0 = umma_desc_sw128(b0_hi);
  const uint64_t ah1 = umma_desc_sw128(a1_hi), bh1 = umma_desc_sw128(b1_hi);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    umma_tf32_ss(d0, ah0 + 2 * k, bh0 + 2 * k, idesc, true);
    umma_tf32_ss(d1, ah1 + 2 * k, bh1 + 2 * k, idesc, true);
  }
}
__device__ __forceinline__ void mma2_ts(uint32_t d0, uint32_t a0_hi, uint32_t a0_lo, uint32_t b0_hi, uint32_t b0_lo,
                                        uint32_t d1, uint32_t a1_hi, uint32_t a1_lo, uint32_t b1_hi, uint32_t b1_lo,
                                        uint32_t idesc) {
  const uint64_t bh0 = umma_desc_sw128(b0_hi), bl0 = umma_desc_sw128(b0_lo);
  const uint64_t bh1 = umma_desc_sw128(b1_hi), bl1 = umma_desc_sw128(b1_lo);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    umma_tf32_ts(d0, a0_lo + 8 * k, bh0 + 2 * k, idesc, k != 0);
    umma_tf32_ts(d1, a1_lo + 8 * k, bh1 + 2 * k, idesc, k != 0);
    umma_tf32_ts(d0, a0_hi + 8 * k, bl0 + 2 * k, idesc, true);
    umma_tf32_ts(d1, a1_hi + 8 * k, bl1 + 2 * k, idesc, true);
  }
#pragma unroll
  for (int k = 0; k < 4; k++) {
    umma_tf32_ts(d0, a0_hi + 8 * k, bh0 + 2 * k, idesc, true);
    umma_tf32_ts(d1, a1_hi + 8 * k, bh1 + 2 * k, idesc, true);
  }
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
constexpr int ABM = 128;   // query rows per CTA (UMMA M)
constexpr int ABN = 64;    // keys per tile
// warps [4g, 4g + 4): softmax group g (key tiles j = g mod kGroups), then one MMA-issue warp and one producer warp

template <int DPAD> struct FwdCfg {
  static constexpr int kSub = DPAD / 32;                    // 32-channel sub-tiles
  static constexpr int kQBytes = 2 * kSub * ABM * 128;      // hi + lo
  static constexpr int kKBytes = 2 * kSub * ABN * 128;
  static constexpr int kVBytes = 2 * 2 * DPAD * 128;        // two 32-key blocks, hi + lo
  static constexpr int kStageBytes = kKBytes + kVBytes;
#ifndef FRN_FWD_TS32
#define FRN_FWD_TS32 1
#endif
  // Q resident in TMEM as the A operand of S (TS form: N / 2 cycles per instruction instead of 32 + N / 4);
  // it costs 2 DPAD columns, which leaves room for two softmax groups
  static constexpr bool kTS = DPAD == 64 || FRN_FWD_TS32;
  static constexpr int kGroups = (DPAD == 32 && !kTS) ? 3 : 2;   // softmax groups = S / P / Y buffers in TMEM
  static constexpr int kStages = DPAD == 32 ? 5 : 2;
  static constexpr int kSmem = kQBytes + kStages * kStageBytes + 1024 + 256;
  static constexpr int kMmaWarp = 4 * kGroups;
  static constexpr int kThreads = (4 * kGroups + 3) * 32;   // softmax groups, S issue, producer, P.V issue
  // TMEM columns (per buffer g: + g * ABN, + g * ABN, + g * DPAD)
  static constexpr int kSMain = 0 /* S, then P hi */, kSSmall = kGroups * ABN /* P lo */, kYMain = 2 * kGroups * ABN;
  static constexpr int kQh = kYMain + kGroups * DPAD, kQl = kQh + DPAD;     // kTS only
};

// FAST (inference at bf16 tolerance, frn_refusion_infer with FRN_BF16): single-pass TF32 -- only the hi planes
// are fetched and multiplied (1 tcgen05.mma per product instead of 3), P is rounded to tf32 once and the
// lo stores disappear.  Relative error ~2^-11 per product; the fp32-parity path (FAST = false) is unchanged.
template <int DPAD, bool FAST>
__global__ void __launch_bounds__(FwdCfg<DPAD>::kThreads, 1)
attn_fwd_tc_kernel(int N, int NP, int d, SplitImage Qi, SplitImage Ki, SplitImage Vi, float* __restrict__ Y,
                   float* __restrict__ lse) {
  using Cfg = FwdCfg<DPAD>;
  static_assert(!FAST || Cfg::kTS, "the single-pass variant keeps Q in TMEM");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + Cfg::kQBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* q_full = bars;                 // 1
  uint64_t* kv_full = bars + 1;            // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;
  constexpr int NG = Cfg::kGroups;
  uint64_t* s_full = kv_empty + Cfg::kStages;   // NG
  uint64_t* p_full = s_full + NG;               // NG
  uint64_t* y_full = p_full + NG;               // NG
  uint64_t* y_free = y_full + NG;               // NG
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(y_free + NG);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, i0 = blockIdx.x * ABM;
  const int T = (N + ABN - 1) / ABN;

  if (tid == 0) {
    mbar_init(q_full, Cfg::kTS ? 128 * NG : 1);
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    for (int s = 0; s < NG; s++) {
      mbar_init(&s_full[s], 1); mbar_init(&p_full[s], 128);
      mbar_init(&y_full[s], 1); mbar_init(&y_free[s], 128);
    }
    mbar_init_fence();
  }
  if (warp == Cfg::kMmaWarp) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == Cfg::kMmaWarp + 1) {
    // ======================= bulk-copy producer =======================
    if (lane == 0) {
      if (!Cfg::kTS) {
        mbar_expect_tx(q_full, Cfg::kQBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + i0) * 32;
          bulk_g2s(sQ + s * ABM * 128, Qi.pos_hi + o, ABM * 128, q_full);
          bulk_g2s(sQ + (Cfg::kSub + s) * ABM * 128, Qi.pos_lo + o, ABM * 128, q_full);
        }
      }
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < T; j++) {
        const int j0 = j * ABN;
        mbar_wait(&kv_empty[stage], phase ^ 1);
        uint8_t* st = sKV + stage * Cfg::kStageBytes;
        mbar_expect_tx(&kv_full[stage], FAST ? Cfg::kStageBytes / 2 : Cfg::kStageBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + j0) * 32;
          bulk_g2s(st + s * ABN * 128, Ki.pos_hi + o, ABN * 128, &kv_full[stage]);
          if (!FAST) bulk_g2s(st + (Cfg::kSub + s) * ABN * 128, Ki.pos_lo + o, ABN * 128, &kv_full[stage]);
        }
        const long ov = (((long)b * (NP / 32) + (j0 >> 5)) * DPAD) * 32;
        bulk_g2s(st + Cfg::kKBytes, Vi.chan_hi + ov, 2 * DPAD * 128, &kv_full[stage]);
        if (!FAST) bulk_g2s(st + Cfg::kKBytes + 2 * DPAD * 128, Vi.chan_lo + ov, 2 * DPAD * 128, &kv_full[stage]);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == Cfg::kMmaWarp) {
    // ======================= S = Q.K^T issue (whole warp waits, one elected lane issues) =======================
    // Two issuing warps (this one: S; warp kMmaWarp + 2: P.V): a single issuing thread spent as long in the
    // latencies of its barrier waits and issue prologues as the tensor pipe needed for the tile.
    constexpr uint32_t idesc_s = umma_idesc_tf32(ABM, ABN);
    const uint32_t q_base = smem_u32(sQ);
    mbar_wait(q_full, 0);
    for (int j = 0; j < T; j++) {
      const int buf = j % NG, stage = j % Cfg::kStages;
      mbar_wait(&kv_full[stage], (j / Cfg::kStages) & 1);
      if (j >= NG) mbar_wait(&y_full[buf], ((j - NG) / NG) & 1);   // P.V(j - NG) has read the P this S overwrites
      tc_fence_after();
      if (elect_one()) {
        const uint32_t k_base = smem_u32(sKV + stage * Cfg::kStageBytes);
        const uint32_t s_acc = tmem + Cfg::kSMain + buf * ABN;
        if (FAST) {
#pragma unroll
          for (int s = 0; s < Cfg::kSub; s++)
            mma_main_ts(s_acc, tmem + Cfg::kQh + 32 * s, k_base + s * ABN * 128, idesc_s, s == 0);
        } else if (Cfg::kTS) {
#pragma unroll
          for (int s = 0; s < Cfg::kSub; s++)
            mma_corr_ts(s_acc, tmem + Cfg::kQh + 32 * s, tmem + Cfg::kQl + 32 * s, k_base + s * ABN * 128,
                        k_base + (Cfg::kSub + s) * ABN * 128, idesc_s, s == 0);
#pragma unroll
          for (int s = 0; s < Cfg::kSub; s++)
            mma_main_ts(s_acc, tmem + Cfg::kQh + 32 * s, k_base + s * ABN * 128, idesc_s);
        } else {
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++)
          mma_corr_ss(s_acc, q_base + s * ABM * 128, q_base + (Cfg::kSub + s) * ABM * 128,
                      k_base + s * ABN * 128, k_base + (Cfg::kSub + s) * ABN * 128, idesc_s, s == 0);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++)
          mma_main_ss(s_acc, q_base + s * ABM * 128, k_base + s * ABN * 128, idesc_s);
        }
        umma_commit(&s_full[buf]);
      }
      __syncwarp();
    }
  } else if (warp == Cfg::kMmaWarp + 2) {
    // ======================= Y += P.V issue =======================
    constexpr uint32_t idesc_y = umma_idesc_tf32(ABM, DPAD);
    for (int j = 0; j < T; j++) {
      const int buf = j % NG, stage = j % Cfg::kStages;
      mbar_wait(&p_full[buf], (j / NG) & 1);
      if (j >= NG) mbar_wait(&y_free[buf], ((j - NG) / NG) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t v_base = smem_u32(sKV + stage * Cfg::kStageBytes + Cfg::kKBytes);
        const uint32_t p_hi = tmem + Cfg::kSMain + buf * ABN, p_lo = tmem + Cfg::kSSmall + buf * ABN;
        const uint32_t y_acc = tmem + Cfg::kYMain + buf * DPAD;
        if (!FAST) {
#pragma unroll
          for (int s = 0; s < 2; s++)
            mma_corr_ts(y_acc, p_hi + s * 32, p_lo + s * 32, v_base + s * DPAD * 128,
                        v_base + (2 + s) * DPAD * 128, idesc_y, s == 0);
        }
#pragma unroll
        for (int s = 0; s < 2; s++)
          mma_main_ts(y_acc, p_hi + s * 32, v_base + s * DPAD * 128, idesc_y, FAST && s == 0);
        umma_commit(&y_full[buf]);
        umma_commit(&kv_empty[stage]);               // S(j) (the other warp) completed before p_full(j) could fire
      }
      __syncwarp();
    }
  } else {
    // ======================= softmax + promotion (thread = query row) =======================
    // NG groups of four warps: group g owns the key tiles j = g (mod NG), i.e. TMEM buffer g, with its own
    // running max / sum / accumulator, and the partial softmax states are merged at the end.  With one
    // group the four softmax warps (one per scheduler, nothing to hide their latencies behind) took twice as
    // long per tile as the tensor pipe; the groups need no exchange inside the loop.
    const int grp = warp >> 2;
    const int row = tid & 127;                         // 0..127
    const uint32_t lane_base = ((uint32_t)((warp & 3) * 32)) << 16;
    float m = -INFINITY, l = 0.f, corr_prev = 0.f;
    float y[DPAD];
#pragma unroll
    for (int c = 0; c < DPAD; c++) y[c] = 0.f;
    const int buf = grp;
    if (Cfg::kTS) {
      // this thread's query row of Q into TMEM: group 0 the hi plane, group 1 the lo plane (packed image rows are
      // SWIZZLE_128B per 32-channel sub-tile: chunk c of row n sits at chunk position c ^ (n & 7))
      const int gq = i0 + row;
      const float* plane = grp == 0 ? Qi.pos_hi : Qi.pos_lo;
      const uint32_t col = grp == 0 ? Cfg::kQh : Cfg::kQl;
      if (grp < (FAST ? 1 : 2)) {
#pragma unroll
        for (int c16 = 0; c16 < DPAD / 16; c16++) {
          const float* rp = plane + (((long)b * Cfg::kSub + (c16 >> 1)) * NP + gq) * 32;
          float v[16];
#pragma unroll
          for (int c = 0; c < 4; c++) {
            const float4 x = __ldg(reinterpret_cast<const float4*>(rp + (((4 * (c16 & 1) + c) ^ (gq & 7)) << 2)));
            v[4 * c] = x.x; v[4 * c + 1] = x.y; v[4 * c + 2] = x.z; v[4 * c + 3] = x.w;
          }
          tmem_st16(tmem + lane_base + col + 16 * c16, v);
        }
        tmem_wait_st();
      }
      tc_fence_before();
      mbar_arrive(q_full);
    }

    auto promote = [&](int j, float corr) {            // y = y * corr + Y_tile(j)
      mbar_wait(&y_full[buf], (j / NG) & 1);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < DPAD; c0 += 16) {
        float a[16];
        tmem_ld16(tmem + lane_base + Cfg::kYMain + buf * DPAD + c0, a);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 16; q += 2) {
          const float2 r = __ffma2_rn(make_float2(y[c0 + q], y[c0 + q + 1]), make_float2(corr, corr),
                                      make_float2(a[q], a[q + 1]));
          y[c0 + q] = r.x; y[c0 + q + 1] = r.y;
        }
      }
      tc_fence_before();
      mbar_arrive(&y_free[buf]);
    };

    int jlast = -1;
    for (int j = grp; j < T; j += NG) {
      const int j0 = j * ABN;
      mbar_wait(&s_full[buf], (j / NG) & 1);
      tc_fence_after();
      float s[ABN];
#pragma unroll
      for (int c0 = 0; c0 < ABN; c0 += 16) {
        float a[16];
        tmem_ld16(tmem + lane_base + Cfg::kSMain + buf * ABN + c0, a);
#pragma unroll
        for (int q = 0; q < 16; q++) s[c0 + q] = a[q];
      }
      tmem_wait_ld();
      float mx = -INFINITY;
      if (j0 + ABN > N) {
#pragma unroll
        for (int q = 0; q < ABN; q++) if (j0 + q >= N) s[q] = -INFINITY;
      }
#pragma unroll
      for (int q = 0; q < ABN; q++) mx = fmaxf(mx, s[q]);
      const float m_new = fmaxf(m, mx);
      const float corr = ex2_fast((m - m_new) * kLog2e);
      float sum = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < ABN; c0 += 16) {
        float hi[16], lo[16];
#pragma unroll
        for (int q = 0; q < 16; q += 2) {                    // packed f32x2 arithmetic (same IEEE operations)
          const float2 x = __fmul2_rn(sub2(make_float2(s[c0 + q], s[c0 + q + 1]), make_float2(m_new, m_new)),
                                      make_float2(kLog2e, kLog2e));
          const float2 p = make_float2(ex2_fast(x.x), ex2_fast(x.y));
          sum += p.x;
          sum += p.y;
          split_tf32x2(p, hi[q], hi[q + 1], lo[q], lo[q + 1]);
        }
        tmem_st16(tmem + lane_base + Cfg::kSMain + buf * ABN + c0, hi);
        if (!FAST) tmem_st16(tmem + lane_base + Cfg::kSSmall + buf * ABN + c0, lo);
      }
      l = fmaf(l, corr, sum);
      m = m_new;
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&p_full[buf]);
      if (jlast >= 0) promote(jlast, corr_prev);
      corr_prev = corr;
      jlast = j;
    }
    if (jlast >= 0) promote(jlast, corr_prev);

    // merge the groups' states (all MMAs are complete and all stages consumed once every group has promoted
    // its last tile, so the Q tile and the K / V stages behind it are dead)
    constexpr int XS = (DPAD + 2) * 128;               // floats per exchanged state: [DPAD + 2][128]
    float* xch = reinterpret_cast<float*>(sQ);
    asm volatile("bar.sync 1, %0;" ::"n"(128 * NG) : "memory");
    if (grp > 0) {
      float* x = xch + (grp - 1) * XS;
      x[row] = m;
      x[128 + row] = l;
#pragma unroll
      for (int c = 0; c < DPAD; c++) x[(2 + c) * 128 + row] = y[c];
    }
    asm volatile("bar.sync 1, %0;" ::"n"(128 * NG) : "memory");
    const int gi = i0 + row;
    if (grp == 0 && gi < N) {
      float mm = m;
#pragma unroll
      for (int g = 1; g < NG; g++) mm = fmaxf(mm, xch[(g - 1) * XS + row]);
      const float a0 = ex2_fast((m - mm) * kLog2e);
      l *= a0;
#pragma unroll
      for (int c = 0; c < DPAD; c++) y[c] *= a0;
#pragma unroll
      for (int g = 1; g < NG; g++) {
        const float* x = xch + (g - 1) * XS;
        const float ag = ex2_fast((x[row] - mm) * kLog2e);
        l = fmaf(x[128 + row], ag, l);
#pragma unroll
        for (int c = 0; c < DPAD; c++) y[c] = fmaf(x[(2 + c) * 128 + row], ag, y[c]);
      }
      const float inv = 1.f / l;
      lse[(long)b * N + gi] = mm + logf(l);
      float* dst = Y + ((long)b * N + gi) * d;
#pragma unroll
      for (int c = 0; c < DPAD; c++) y[c] *= inv;
      if ((d & 3) == 0) {
#pragma unroll
        for (int c = 0; c < DPAD; c += 4)
          if (c < d) *reinterpret_cast<float4*>(dst + c) = make_float4(y[c], y[c + 1], y[c + 2], y[c + 3]);
      } else {
#pragma unroll
        for (int c = 0; c < DPAD; c++)
          if (c < d) dst[c] = y[c];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == Cfg::kMmaWarp) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------------------
// backward, part 1: dQ.   CTA = NQ tiles of 128 queries (one compute warpgroup each, ping-pong on
// the tensor pipe), streaming 32-key tiles:  S = Q.K^T, dP = dY.V^T  (shared-memory operands)
//   -> dS = exp(S - lse) * (dP - delta)  (thread = query row; written back to TMEM as tf32 hi/lo)
//   -> dQ_tile = dS.K  (A from TMEM), promoted into fp32 registers every tile.
// ------------------------------------------------------------------------------------------------
constexpr int BT = 32;     // keys (dQ kernel) / queries (dK,dV kernel) per streamed tile

template <int DPAD, int NQ> struct DqCfg {
  static constexpr int kSub = DPAD / 32;
  static constexpr int kOperBytes = kSub * 128 * 128;          // one 128-row pos tile, hi or lo
  static constexpr int kResBytes = 4 * kOperBytes;             // Q hi, Q lo, dY hi, dY lo
  // keys per streamed tile: 64 for d <= 32 (every tcgen05.mma costs >= ~45 cycles however small its N is, so
  // the S / dP products are issued at N = 64), 32 for d = 64 (shared memory)
  static constexpr int T = DPAD == 32 ? 64 : 32;
  static constexpr int kPosBytes = kSub * T * 128;             // streamed pos tile, hi or lo
  static constexpr int kChanBytes = (T / 32) * DPAD * 128;     // streamed chan blocks, hi or lo
  static constexpr int kStageBytes = 4 * kPosBytes + 2 * kChanBytes;   // K pos hi/lo, V pos hi/lo, K chan hi/lo
  static constexpr int kStages = 2;
  static constexpr int kSmem = NQ * kResBytes + kStages * kStageBytes + 1024 + 256;
  static constexpr int kTmemCols = 256 * NQ;
  static constexpr int kThreads = (4 * NQ + 2) * 32;
  // TMEM columns of one query tile
  static constexpr int kS = 0 /* S, then dS hi */, kSs = T /* dS lo */, kP = 2 * T /* dP */, kAcc = 3 * T;
};

template <int DPAD, int NQ>
__global__ void __launch_bounds__(DqCfg<DPAD, NQ>::kThreads, 1)
attn_dq_tc_kernel(int N, int NP, int d, SplitImage Qi, SplitImage Ki, SplitImage Vi, SplitImage dYi,
                  const float* __restrict__ lse_pad, const float* __restrict__ delta_pad,
                  float* __restrict__ dQp) {
  using Cfg = DqCfg<DPAD, NQ>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sRes = smem;
  uint8_t* sStage = smem + NQ * Cfg::kResBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStage + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* res_full = bars;                       // NQ
  uint64_t* kv_full = res_full + NQ;               // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;     // kStages
  uint64_t* sdp_full = kv_empty + Cfg::kStages;    // NQ   S, dP of the tile are in TMEM (and the previous dQ tile)
  uint64_t* ds_full = sdp_full + NQ;               // NQ   dS written, previous dQ tile promoted
  uint64_t* acc_done = ds_full + NQ;               // NQ   last dQ tile complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_done + NQ);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y;
  const int nqt = NP / 128;                          // query tiles per sample
  constexpr int T = Cfg::T;
  const int TK = (N + T - 1) / T;                    // key tiles
  const int qt0 = blockIdx.x * NQ;

  if (tid == 0) {
    for (int h = 0; h < NQ; h++) {
      mbar_init(&res_full[h], 1); mbar_init(&sdp_full[h], 1); mbar_init(&ds_full[h], 128); mbar_init(&acc_done[h], 1);
    }
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init_fence();
  }
  if (warp == 4 * NQ) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 4 * NQ + 1) {
    // ======================= bulk-copy producer =======================
    if (lane == 0) {
      for (int h = 0; h < NQ; h++) {
        if (qt0 + h >= nqt) break;
        uint8_t* r = sRes + h * Cfg::kResBytes;
        mbar_expect_tx(&res_full[h], Cfg::kResBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)(qt0 + h) * 128) * 32;
          bulk_g2s(r + s * 16384, Qi.pos_hi + o, 16384, &res_full[h]);
          bulk_g2s(r + Cfg::kOperBytes + s * 16384, Qi.pos_lo + o, 16384, &res_full[h]);
          bulk_g2s(r + 2 * Cfg::kOperBytes + s * 16384, dYi.pos_hi + o, 16384, &res_full[h]);
          bulk_g2s(r + 3 * Cfg::kOperBytes + s * 16384, dYi.pos_lo + o, 16384, &res_full[h]);
        }
      }
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < TK; j++) {
        mbar_wait(&kv_empty[stage], phase ^ 1);
        uint8_t* st = sStage + stage * Cfg::kStageBytes;
        mbar_expect_tx(&kv_full[stage], Cfg::kStageBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)j * T) * 32;
          bulk_g2s(st + s * T * 128, Ki.pos_hi + o, T * 128, &kv_full[stage]);
          bulk_g2s(st + Cfg::kPosBytes + s * T * 128, Ki.pos_lo + o, T * 128, &kv_full[stage]);
          bulk_g2s(st + 2 * Cfg::kPosBytes + s * T * 128, Vi.pos_hi + o, T * 128, &kv_full[stage]);
          bulk_g2s(st + 3 * Cfg::kPosBytes + s * T * 128, Vi.pos_lo + o, T * 128, &kv_full[stage]);
        }
        const long oc = (((long)b * (NP / 32) + (long)j * (T / 32)) * DPAD) * 32;
        bulk_g2s(st + 4 * Cfg::kPosBytes, Ki.chan_hi + oc, Cfg::kChanBytes, &kv_full[stage]);
        bulk_g2s(st + 4 * Cfg::kPosBytes + Cfg::kChanBytes, Ki.chan_lo + oc, Cfg::kChanBytes, &kv_full[stage]);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 4 * NQ) {
    // ======================= MMA issue (whole warp waits, one elected lane issues) =======================
    {
      constexpr uint32_t idesc_s = umma_idesc_tf32(128, T);
      constexpr uint32_t idesc_q = umma_idesc_tf32(128, DPAD);
      int nact = 0;
      for (int h = 0; h < NQ; h++)
        if (qt0 + h < nqt) { mbar_wait(&res_full[h], 0); nact = h + 1; }
      tc_fence_after();
      for (int j = 0; j <= TK; j++) {
        if (j < TK) {
          mbar_wait(&kv_full[j % Cfg::kStages], (j / Cfg::kStages) & 1);
          tc_fence_after();
        }
        for (int h = 0; h < nact; h++) {
          const uint32_t tb = tmem + h * 256;
          const uint32_t r = smem_u32(sRes + h * Cfg::kResBytes);
          if (j > 0) {                                   // dQ_h(j-1) = dS.K
            mbar_wait(&ds_full[h], (j - 1) & 1);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t kc =
                  smem_u32(sStage + ((j - 1) % Cfg::kStages) * Cfg::kStageBytes + 4 * Cfg::kPosBytes);
#pragma unroll
              for (int sb = 0; sb < T / 32; sb++)      // 32 keys per chan block
                mma_corr_ts(tb + Cfg::kAcc, tb + Cfg::kS + 32 * sb, tb + Cfg::kSs + 32 * sb, kc + sb * DPAD * 128,
                            kc + Cfg::kChanBytes + sb * DPAD * 128, idesc_q, sb == 0);
#pragma unroll
              for (int sb = 0; sb < T / 32; sb++)
                mma_main_ts(tb + Cfg::kAcc, tb + Cfg::kS + 32 * sb, kc + sb * DPAD * 128, idesc_q);
              if (h == nact - 1) umma_commit(&kv_empty[(j - 1) % Cfg::kStages]);
            }
            __syncwarp();
          }
          if (elect_one()) {
          if (j < TK) {                                  // S_h(j) = Q.K^T ; dP_h(j) = dY.V^T
            const uint32_t st = smem_u32(sStage + (j % Cfg::kStages) * Cfg::kStageBytes);
#pragma unroll
            for (int s = 0; s < Cfg::kSub; s++)
              mma_corr_ss(tb + Cfg::kS, r + s * 16384, r + Cfg::kOperBytes + s * 16384, st + s * T * 128,
                          st + Cfg::kPosBytes + s * T * 128, idesc_s, s == 0);
#pragma unroll
            for (int s = 0; s < Cfg::kSub; s++)
              mma_main_ss(tb + Cfg::kS, r + s * 16384, st + s * T * 128, idesc_s);
#pragma unroll
            for (int s = 0; s < Cfg::kSub; s++)
              mma_corr_ss(tb + Cfg::kP, r + 2 * Cfg::kOperBytes + s * 16384, r + 3 * Cfg::kOperBytes + s * 16384,
                          st + 2 * Cfg::kPosBytes + s * T * 128, st + 3 * Cfg::kPosBytes + s * T * 128, idesc_s,
                          s == 0);
#pragma unroll
            for (int s = 0; s < Cfg::kSub; s++)
              mma_main_ss(tb + Cfg::kP, r + 2 * Cfg::kOperBytes + s * 16384, st + 2 * Cfg::kPosBytes + s * T * 128,
                          idesc_s);
            umma_commit(&sdp_full[h]);
          } else {
            umma_commit(&acc_done[h]);
          }
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ======================= dS + promotion (thread = query row) =======================
    const int h = warp >> 2;
    if (qt0 + h < nqt) {
      const int row = (warp & 3) * 32 + lane;
      const int gi = (qt0 + h) * 128 + row;
      const uint32_t tb = tmem + h * 256 + (((uint32_t)((warp & 3) * 32)) << 16);
      const float lse_i = __ldg(lse_pad + (long)b * NP + gi);
      const float delta_i = __ldg(delta_pad + (long)b * NP + gi);
      float dq[DPAD];
#pragma unroll
      for (int c = 0; c < DPAD; c++) dq[c] = 0.f;
      auto promote = [&]() {
#pragma unroll
        for (int c0 = 0; c0 < DPAD; c0 += 16) {
          float a[16];
          tmem_ld16(tb + Cfg::kAcc + c0, a);
          tmem_wait_ld();
#pragma unroll
          for (int q = 0; q < 16; q++) dq[c0 + q] += a[q];
        }
      };
      for (int j = 0; j < TK; j++) {
        mbar_wait(&sdp_full[h], j & 1);
        tc_fence_after();
        if (j > 0) promote();
        const int j0 = j * T;
#pragma unroll
        for (int c0 = 0; c0 < T; c0 += 16) {
          float p[16], hi[16], lo[16];
          {
            float a[16], g[16];
            tmem_ld16(tb + Cfg::kS + c0, a);
            tmem_ld16(tb + Cfg::kP + c0, g);
            tmem_wait_ld();
#pragma unroll
            for (int q = 0; q < 16; q++) {
              p[q] = ex2_fast((a[q] - lse_i) * kLog2e);
              if (j0 + c0 + q >= N) p[q] = 0.f;
              split_tf32(p[q] * (g[q] - delta_i), hi[q], lo[q]);
            }
          }
          tmem_st16(tb + Cfg::kS + c0, hi);
          tmem_st16(tb + Cfg::kSs + c0, lo);
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&ds_full[h]);
      }
      mbar_wait(&acc_done[h], 0);
      tc_fence_after();
      promote();
      if (gi < N) {
        float* dst = dQp + ((long)b * N + gi) * d;
        if ((d & 3) == 0) {
#pragma unroll
          for (int c = 0; c < DPAD; c += 4)
            if (c < d) *reinterpret_cast<float4*>(dst + c) = make_float4(dq[c], dq[c + 1], dq[c + 2], dq[c + 3]);
        } else {
#pragma unroll
          for (int c = 0; c < DPAD; c++)
            if (c < d) dst[c] = dq[c];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4 * NQ) tmem_dealloc<Cfg::kTmemCols>(tmem);
}

// ------------------------------------------------------------------------------------------------
// backward, part 2: dK, dV.   CTA = NKT tiles of 128 keys (lanes = keys), streaming 32-query tiles:
//   S^T = K.Q^T, dP^T = V.dY^T  ->  P^T = exp(S^T - lse[q]), dS^T = P^T * (dP^T - delta[q])
//   -> dV_tile = P^T.dY, dK_tile = dS^T.Q  (A from TMEM), promoted into registers every tile.
// Each CTA accumulates one 32-channel group of dV / dK (blockIdx.z), so d = 64 runs two CTAs per
// key tile (S^T and dP^T are recomputed; the accumulators stay at 64 registers per thread).
// ------------------------------------------------------------------------------------------------
template <int DPAD, int NKT> struct DkvCfg {
  static constexpr int kSub = DPAD / 32;
  static constexpr int kOperBytes = kSub * 128 * 128;          // one 128-row pos tile, hi or lo
  static constexpr int kResBytes = 4 * kOperBytes;             // K hi, K lo, V hi, V lo
  static constexpr int kPosBytes = kSub * BT * 128;            // streamed pos tile, hi or lo
  static constexpr int kChanBytes = 32 * 128;                  // one 32-channel group of a chan block
  static constexpr int kStageBytes = 4 * kPosBytes + 4 * kChanBytes;  // Q pos, dY pos, dY chan, Q chan (hi/lo each)
  static constexpr int kStages = 2;
  static constexpr int kRowBytes = kStages * 2 * BT * 4;       // lse, delta of the tile's queries
  static constexpr int kSmem = NKT * kResBytes + kStages * kStageBytes + kRowBytes + 1024 + 256;
  static constexpr int kTmemCols = 256 * NKT;
  static constexpr int kThreads = (4 * NKT + 2) * 32;
  static constexpr int kS = 0 /* S^T, then P^T hi */, kSs = 32 /* P^T lo */, kP = 64 /* dP^T, then dS^T hi */,
                       kPs = 96 /* dS^T lo */, kV = 128, kK = 160;
};

template <int DPAD, int NKT>
__global__ void __launch_bounds__(DkvCfg<DPAD, NKT>::kThreads, 1)
attn_dkv_tc_kernel(int N, int NP, int d, SplitImage Qi, SplitImage Ki, SplitImage Vi, SplitImage dYi,
                   const float* __restrict__ lse_pad, const float* __restrict__ delta_pad,
                   float* __restrict__ dKp, float* __restrict__ dVp) {
  using Cfg = DkvCfg<DPAD, NKT>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sRes = smem;
  uint8_t* sStage = smem + NKT * Cfg::kResBytes;
  float* sRows = reinterpret_cast<float*>(sStage + Cfg::kStages * Cfg::kStageBytes);   // [stage][lse BT | delta BT]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sRows) + Cfg::kRowBytes);
  uint64_t* res_full = bars;                       // NKT
  uint64_t* q_full = res_full + NKT;               // kStages
  uint64_t* q_empty = q_full + Cfg::kStages;       // kStages
  uint64_t* sdp_full = q_empty + Cfg::kStages;     // NKT
  uint64_t* ds_full = sdp_full + NKT;              // NKT
  uint64_t* acc_done = ds_full + NKT;              // NKT
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_done + NKT);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, cg = blockIdx.z;          // cg: 32-channel group of dV / dK
  const int nkt = NP / 128;                          // key tiles per sample
  const int TQ = (N + BT - 1) / BT;                  // query tiles
  const int kt0 = blockIdx.x * NKT;

  if (tid == 0) {
    for (int h = 0; h < NKT; h++) {
      mbar_init(&res_full[h], 1); mbar_init(&sdp_full[h], 1); mbar_init(&ds_full[h], 128); mbar_init(&acc_done[h], 1);
    }
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
    mbar_init_fence();
  }
  if (warp == 4 * NKT) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 4 * NKT + 1) {
    // ======================= bulk-copy producer =======================
    if (lane == 0) {
      for (int h = 0; h < NKT; h++) {
        if (kt0 + h >= nkt) break;
        uint8_t* r = sRes + h * Cfg::kResBytes;
        mbar_expect_tx(&res_full[h], Cfg::kResBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)(kt0 + h) * 128) * 32;
          bulk_g2s(r + s * 16384, Ki.pos_hi + o, 16384, &res_full[h]);
          bulk_g2s(r + Cfg::kOperBytes + s * 16384, Ki.pos_lo + o, 16384, &res_full[h]);
          bulk_g2s(r + 2 * Cfg::kOperBytes + s * 16384, Vi.pos_hi + o, 16384, &res_full[h]);
          bulk_g2s(r + 3 * Cfg::kOperBytes + s * 16384, Vi.pos_lo + o, 16384, &res_full[h]);
        }
      }
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < TQ; j++) {
        mbar_wait(&q_empty[stage], phase ^ 1);
        uint8_t* st = sStage + stage * Cfg::kStageBytes;
        mbar_expect_tx(&q_full[stage], Cfg::kStageBytes + 2 * BT * 4);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)j * BT) * 32;
          bulk_g2s(st + s * BT * 128, Qi.pos_hi + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + Cfg::kPosBytes + s * BT * 128, Qi.pos_lo + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + 2 * Cfg::kPosBytes + s * BT * 128, dYi.pos_hi + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + 3 * Cfg::kPosBytes + s * BT * 128, dYi.pos_lo + o, BT * 128, &q_full[stage]);
        }
        const long oc = (((long)b * (NP / 32) + j) * DPAD + cg * 32) * 32;
        uint8_t* sc = st + 4 * Cfg::kPosBytes;
        bulk_g2s(sc, dYi.chan_hi + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + Cfg::kChanBytes, dYi.chan_lo + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + 2 * Cfg::kChanBytes, Qi.chan_hi + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + 3 * Cfg::kChanBytes, Qi.chan_lo + oc, Cfg::kChanBytes, &q_full[stage]);
        float* rw = sRows + stage * 2 * BT;
        bulk_g2s(rw, lse_pad + (long)b * NP + (long)j * BT, BT * 4, &q_full[stage]);
        bulk_g2s(rw + BT, delta_pad + (long)b * NP + (long)j * BT, BT * 4, &q_full[stage]);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 4 * NKT) {
    // ======================= MMA issue (whole warp waits, one elected lane issues) =======================
    {
      constexpr uint32_t idesc = umma_idesc_tf32(128, BT);     // every product here is 128 x 32
      int nact = 0;
      for (int h = 0; h < NKT; h++)
        if (kt0 + h < nkt) { mbar_wait(&res_full[h], 0); nact = h + 1; }
      tc_fence_after();
      for (int j = 0; j <= TQ; j++) {
        if (j < TQ) {
          mbar_wait(&q_full[j % Cfg::kStages], (j / Cfg::kStages) & 1);
          tc_fence_after();
        }
        for (int h = 0; h < nact; h++) {
          const uint32_t tb = tmem + h * 256;
          const uint32_t r = smem_u32(sRes + h * Cfg::kResBytes);
          if (j > 0) {                                   // dV_h(j-1) = P^T.dY ; dK_h(j-1) = dS^T.Q
            mbar_wait(&ds_full[h], (j - 1) & 1);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t sc =
                  smem_u32(sStage + ((j - 1) % Cfg::kStages) * Cfg::kStageBytes + 4 * Cfg::kPosBytes);
              mma2_ts(tb + Cfg::kV, tb + Cfg::kS, tb + Cfg::kSs, sc, sc + Cfg::kChanBytes,
                      tb + Cfg::kK, tb + Cfg::kP, tb + Cfg::kPs, sc + 2 * Cfg::kChanBytes, sc + 3 * Cfg::kChanBytes,
                      idesc);
              if (h == nact - 1) umma_commit(&q_empty[(j - 1) % Cfg::kStages]);
            }
            __syncwarp();
          }
          if (elect_one()) {
          if (j < TQ) {                                  // S^T_h(j) = K.Q^T ; dP^T_h(j) = V.dY^T
            const uint32_t st = smem_u32(sStage + (j % Cfg::kStages) * Cfg::kStageBytes);
#pragma unroll
            for (int s = 0; s < Cfg::kSub; s++)
              mma2_corr_ss(tb + Cfg::kS, r + s * 16384, r + Cfg::kOperBytes + s * 16384, st + s * BT * 128,
                           st + Cfg::kPosBytes + s * BT * 128,
                           tb + Cfg::kP, r + 2 * Cfg::kOperBytes + s * 16384, r + 3 * Cfg::kOperBytes + s * 16384,
                           st + 2 * Cfg::kPosBytes + s * BT * 128, st + 3 * Cfg::kPosBytes + s * BT * 128, idesc,
                           s == 0);
#pragma unroll
            for (int s = 0; s < Cfg::kSub; s++)
              mma2_main_ss(tb + Cfg::kS, r + s * 16384, st + s * BT * 128,
                           tb + Cfg::kP, r + 2 * Cfg::kOperBytes + s * 16384, st + 2 * Cfg::kPosBytes + s * BT * 128,
                           idesc);
            umma_commit(&sdp_full[h]);
          } else {
            umma_commit(&acc_done[h]);
          }
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ======================= P^T, dS^T + promotion (thread = key row) =======================
    const int h = warp >> 2;
    if (kt0 + h < nkt) {
      const int row = (warp & 3) * 32 + lane;
      const int gj = (kt0 + h) * 128 + row;
      const uint32_t tb = tmem + h * 256 + (((uint32_t)((warp & 3) * 32)) << 16);
      float dv[32], dk[32];
#pragma unroll
      for (int c = 0; c < 32; c++) { dv[c] = 0.f; dk[c] = 0.f; }
      auto promote = [&]() {
#pragma unroll
        for (int c0 = 0; c0 < 32; c0 += 16) {
          float a[16], s[16];
          tmem_ld16(tb + Cfg::kV + c0, a);
          tmem_ld16(tb + Cfg::kK + c0, s);
          tmem_wait_ld();
#pragma unroll
          for (int q = 0; q < 16; q++) { dv[c0 + q] += a[q]; dk[c0 + q] += s[q]; }
        }
      };
      for (int j = 0; j < TQ; j++) {
        mbar_wait(&sdp_full[h], j & 1);
        tc_fence_after();
        if (j > 0) promote();
        mbar_wait(&q_full[j % Cfg::kStages], (j / Cfg::kStages) & 1);   // lse / delta of the tile have landed
        const float* rw = sRows + (j % Cfg::kStages) * 2 * BT;
#pragma unroll
        for (int c0 = 0; c0 < BT; c0 += 16) {
          float p[16], hi[16], lo[16], g[16];
          {
            float a[16];
            tmem_ld16(tb + Cfg::kS + c0, a);
            tmem_ld16(tb + Cfg::kP + c0, g);
            tmem_wait_ld();
#pragma unroll
            for (int q = 0; q < 16; q++) {
              p[q] = ex2_fast((a[q] - rw[c0 + q]) * kLog2e);
              split_tf32(p[q], hi[q], lo[q]);
            }
          }
          tmem_st16(tb + Cfg::kS + c0, hi);
          tmem_st16(tb + Cfg::kSs + c0, lo);
#pragma unroll
          for (int q = 0; q < 16; q++) split_tf32(p[q] * (g[q] - rw[BT + c0 + q]), hi[q], lo[q]);
          tmem_st16(tb + Cfg::kP + c0, hi);
          tmem_st16(tb + Cfg::kPs + c0, lo);
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&ds_full[h]);
      }
      mbar_wait(&acc_done[h], 0);
      tc_fence_after();
      promote();
      if (gj < N) {
        const int cw = min(32, d - cg * 32);             // channels of this group that exist
        float* dv_dst = dVp + ((long)b * N + gj) * d + cg * 32;
        float* dk_dst = dKp + ((long)b * N + gj) * d + cg * 32;
        if ((d & 3) == 0) {
#pragma unroll
          for (int c = 0; c < 32; c += 4)
            if (c < cw) {
              *reinterpret_cast<float4*>(dv_dst + c) = make_float4(dv[c], dv[c + 1], dv[c + 2], dv[c + 3]);
              *reinterpret_cast<float4*>(dk_dst + c) = make_float4(dk[c], dk[c + 1], dk[c + 2], dk[c + 3]);
            }
        } else {
#pragma unroll
          for (int c = 0; c < 32; c++)
            if (c < cw) { dv_dst[c] = dv[c]; dk_dst[c] = dk[c]; }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4 * NKT) tmem_dealloc<Cfg::kTmemCols>(tmem);
}

// ------------------------------------------------------------------------------------------------
// backward, part 2, second design (the default): one tile of 128 keys per CTA, the S^T / dP^T
// accumulators double buffered and EIGHT compute warps.
// In the ping-pong design above every tile is a serial chain  S^T,dP^T MMAs -> exp / dS warps -> dV,dK
// MMAs -> next S^T (it overwrites P^T in place), and two tiles in flight could not hide a chain that is
// three times longer than its tensor-pipe share (measured ~4600 cycles per 32-query step against 2700 of
// tcgen05.mma time, tools/mma_cost_probe.cu).  Here S^T(j+1), dP^T(j+1) go to a second pair of TMEM
// buffers and are issued BEFORE dV(j), dK(j), so the tensor pipe works on tile j+1 while the compute warps
// handle tile j; the compute work of a tile is split over two warps per lane quadrant (16 of the 32 query
// columns each, and 16 of the 32 accumulator channels each at promotion time).
// TMEM: S^T[2] 0,32 | dP^T[2] 64,96 | P^T hi 128, lo 160 | dS^T hi 192, lo 224 | dV 256 | dK 288.
// ------------------------------------------------------------------------------------------------
#ifdef FRN_ATTN_TRACE
__device__ unsigned long long g_attn_trace[16];
#define ATR_T0() const long long atr0__ = clock64()
#define ATR_ADD(slot) do { if (blockIdx.x == 0 && blockIdx.y == 0 && (threadIdx.x & 31) == 0) atomicAdd(&g_attn_trace[slot], (unsigned long long)(clock64() - atr0__)); } while (0)
#else
#define ATR_T0() do {} while (0)
#define ATR_ADD(slot) do {} while (0)
#endif

template <int DPAD> struct Dkv2Cfg {
  static constexpr int kSub = DPAD / 32;
  static constexpr int kOperBytes = kSub * 128 * 128;          // one 128-row pos tile, hi or lo
  static constexpr int kResBytes = 4 * kOperBytes;             // K hi, K lo, V hi, V lo
  static constexpr int kPosBytes = kSub * BT * 128;            // streamed pos tile, hi or lo
  static constexpr int kChanBytes = 32 * 128;                  // one 32-channel group of a chan block
  static constexpr int kStageBytes = 4 * kPosBytes + 4 * kChanBytes;  // Q pos, dY pos, dY chan, Q chan (hi/lo each)
  // d <= 32: K and V (tf32 hi / lo) live in TMEM as the A operand of S^T / dP^T (TS form, which runs at the
  // math rate: 16 cycles per N = 32 instruction instead of the 40 of a shared-memory A operand)
  static constexpr bool kTS = DPAD == 32;
  static constexpr int kStages = DPAD == 32 ? 6 : 2;
  static constexpr int kRowBytes = kStages * 2 * BT * 4;       // lse, delta of the tile's queries
  static constexpr int kSmem = (kTS ? 0 : kResBytes) + kStages * kStageBytes + kRowBytes + 1024 + 256;
  static constexpr int kKh = 320, kKl = 352, kVh = 384, kVl = 416;   // TMEM columns of the resident operands (kTS)
  static constexpr int kThreads = 11 * 32;                     // warps 0-7 compute, 8 S^T/dP^T issue, 9 producer, 10 dV/dK issue
  static constexpr int kS = 0, kP = 64, kPhi = 128, kPlo = 160, kDhi = 192, kDlo = 224, kV = 256, kK = 288;
};

template <int DPAD>
__global__ void __launch_bounds__(Dkv2Cfg<DPAD>::kThreads, 1)
attn_dkv2_tc_kernel(int N, int NP, int d, SplitImage Qi, SplitImage Ki, SplitImage Vi, SplitImage dYi,
                    const float* __restrict__ lse_pad, const float* __restrict__ delta_pad,
                    float* __restrict__ dKp, float* __restrict__ dVp) {
  using Cfg = Dkv2Cfg<DPAD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sRes = smem;
  uint8_t* sStage = smem + (Cfg::kTS ? 0 : Cfg::kResBytes);
  float* sRows = reinterpret_cast<float*>(sStage + Cfg::kStages * Cfg::kStageBytes);   // [stage][lse BT | delta BT]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sRows) + Cfg::kRowBytes);
  uint64_t* res_full = bars;                       // 1
  uint64_t* q_full = res_full + 1;                 // kStages
  uint64_t* q_empty = q_full + Cfg::kStages;       // kStages
  uint64_t* sdp_full = q_empty + Cfg::kStages;     // 2   S^T, dP^T of a tile are in TMEM buffer (j & 1)
  uint64_t* s_free = sdp_full + 2;                 // 2   ... and have been read into registers (256 arrivals)
  uint64_t* ds_full = s_free + 2;                  // 1   P^T, dS^T written, previous dV / dK promoted (256)
  uint64_t* pv_done = ds_full + 1;                 // 1   dV, dK MMAs of a tile complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, cg = blockIdx.z;          // cg: 32-channel group of dV / dK
  const int kt = blockIdx.x;                          // key tile
  const int TQ = (N + BT - 1) / BT;                   // query tiles

  if (tid == 0) {
    mbar_init(res_full, Cfg::kTS ? 256 : 1); mbar_init(ds_full, 256); mbar_init(pv_done, 1);
    for (int s = 0; s < 2; s++) { mbar_init(&sdp_full[s], 1); mbar_init(&s_free[s], 256); }
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
    mbar_init_fence();
  }
  if (warp == 8) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 9) {
    // ======================= bulk-copy producer =======================
    if (lane == 0) {
      if (!Cfg::kTS) {
        mbar_expect_tx(res_full, Cfg::kResBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)kt * 128) * 32;
          bulk_g2s(sRes + s * 16384, Ki.pos_hi + o, 16384, res_full);
          bulk_g2s(sRes + Cfg::kOperBytes + s * 16384, Ki.pos_lo + o, 16384, res_full);
          bulk_g2s(sRes + 2 * Cfg::kOperBytes + s * 16384, Vi.pos_hi + o, 16384, res_full);
          bulk_g2s(sRes + 3 * Cfg::kOperBytes + s * 16384, Vi.pos_lo + o, 16384, res_full);
        }
      }
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < TQ; j++) {
        mbar_wait(&q_empty[stage], phase ^ 1);
        uint8_t* st = sStage + stage * Cfg::kStageBytes;
        mbar_expect_tx(&q_full[stage], Cfg::kStageBytes + 2 * BT * 4);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)j * BT) * 32;
          bulk_g2s(st + s * BT * 128, Qi.pos_hi + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + Cfg::kPosBytes + s * BT * 128, Qi.pos_lo + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + 2 * Cfg::kPosBytes + s * BT * 128, dYi.pos_hi + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + 3 * Cfg::kPosBytes + s * BT * 128, dYi.pos_lo + o, BT * 128, &q_full[stage]);
        }
        const long oc = (((long)b * (NP / 32) + j) * DPAD + cg * 32) * 32;
        uint8_t* sc = st + 4 * Cfg::kPosBytes;
        bulk_g2s(sc, dYi.chan_hi + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + Cfg::kChanBytes, dYi.chan_lo + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + 2 * Cfg::kChanBytes, Qi.chan_hi + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + 3 * Cfg::kChanBytes, Qi.chan_lo + oc, Cfg::kChanBytes, &q_full[stage]);
        float* rw = sRows + stage * 2 * BT;
        bulk_g2s(rw, lse_pad + (long)b * NP + (long)j * BT, BT * 4, &q_full[stage]);
        bulk_g2s(rw + BT, delta_pad + (long)b * NP + (long)j * BT, BT * 4, &q_full[stage]);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 8) {
    // ======================= MMA issue (whole warp waits, one elected lane issues) =======================
    constexpr uint32_t idesc = umma_idesc_tf32(128, BT);     // every product here is 128 x 32
    const uint32_t r = smem_u32(sRes);
    auto issue_sdp = [&](int t) {                            // S^T(t) = K.Q^T ; dP^T(t) = V.dY^T
      { ATR_T0();
      mbar_wait(&q_full[t % Cfg::kStages], (t / Cfg::kStages) & 1);
      ATR_ADD(0); }
      { ATR_T0();
      if (t >= 2) mbar_wait(&s_free[t & 1], ((t - 2) >> 1) & 1);
      ATR_ADD(1); }
      tc_fence_after();
      ATR_T0();
      if (elect_one()) {
        const uint32_t st = smem_u32(sStage + (t % Cfg::kStages) * Cfg::kStageBytes);
        const uint32_t dS = tmem + Cfg::kS + (t & 1) * BT, dP = tmem + Cfg::kP + (t & 1) * BT;
        if (Cfg::kTS) {
          mma2_ts(dS, tmem + Cfg::kKh, tmem + Cfg::kKl, st, st + Cfg::kPosBytes,
                  dP, tmem + Cfg::kVh, tmem + Cfg::kVl, st + 2 * Cfg::kPosBytes, st + 3 * Cfg::kPosBytes, idesc);
        } else {
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++)
          mma2_corr_ss(dS, r + s * 16384, r + Cfg::kOperBytes + s * 16384, st + s * BT * 128,
                       st + Cfg::kPosBytes + s * BT * 128,
                       dP, r + 2 * Cfg::kOperBytes + s * 16384, r + 3 * Cfg::kOperBytes + s * 16384,
                       st + 2 * Cfg::kPosBytes + s * BT * 128, st + 3 * Cfg::kPosBytes + s * BT * 128, idesc, s == 0);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++)
          mma2_main_ss(dS, r + s * 16384, st + s * BT * 128,
                       dP, r + 2 * Cfg::kOperBytes + s * 16384, st + 2 * Cfg::kPosBytes + s * BT * 128, idesc);
        }
        umma_commit(&sdp_full[t & 1]);
      }
      __syncwarp();
      ATR_ADD(2);
    };
    // Two issuing warps: this one feeds S^T / dP^T, warp 10 feeds dV / dK.  A single issuing thread spent a
    // quarter of every step in the latency of its four (already satisfied) barrier waits, with the tensor
    // pipe idle meanwhile; the two streams only depend on each other through the barriers below.
    mbar_wait(res_full, 0);
    tc_fence_after();
    for (int t = 0; t < TQ; t++) issue_sdp(t);
  } else if (warp == 10) {
    constexpr uint32_t idesc = umma_idesc_tf32(128, BT);
    for (int j = 0; j < TQ; j++) {
      { ATR_T0();
      mbar_wait(ds_full, j & 1);                             // dV(j) = P^T.dY ; dK(j) = dS^T.Q
      ATR_ADD(3); }
      tc_fence_after();
      ATR_T0();
      if (elect_one()) {
        const uint32_t sc = smem_u32(sStage + (j % Cfg::kStages) * Cfg::kStageBytes + 4 * Cfg::kPosBytes);
        mma2_ts(tmem + Cfg::kV, tmem + Cfg::kPhi, tmem + Cfg::kPlo, sc, sc + Cfg::kChanBytes,
                tmem + Cfg::kK, tmem + Cfg::kDhi, tmem + Cfg::kDlo, sc + 2 * Cfg::kChanBytes, sc + 3 * Cfg::kChanBytes,
                idesc);
        umma_commit(pv_done);
        // S^T(j), dP^T(j) (warp 8) completed before compute(j) could release ds_full(j): the stage is free
        umma_commit(&q_empty[j % Cfg::kStages]);
      }
      __syncwarp();
      ATR_ADD(4);
    }
  } else {
    // ======================= P^T, dS^T + promotion (thread = key row, half of the columns) =======================
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    const int gj = kt * 128 + row;
    const uint32_t tb = tmem + (((uint32_t)(quad * 32)) << 16) + half * 16;
    float dv[16], dk[16];
#pragma unroll
    for (int c = 0; c < 16; c++) { dv[c] = 0.f; dk[c] = 0.f; }
    if (Cfg::kTS) {
      // this thread's key row, channels [16 half, 16 half + 16) of K and V (hi, lo): packed image -> TMEM.
      // The image rows are SWIZZLE_128B: 16-byte chunk c of row n sits at chunk position c ^ (n & 7).
      const long ro = ((long)b * NP + gj) * 32;
      const float* src[4] = {Ki.pos_hi + ro, Ki.pos_lo + ro, Vi.pos_hi + ro, Vi.pos_lo + ro};
      const int col[4] = {Cfg::kKh, Cfg::kKl, Cfg::kVh, Cfg::kVl};
#pragma unroll
      for (int o = 0; o < 4; o++) {
        float v[16];
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const float4 x = __ldg(reinterpret_cast<const float4*>(src[o] + (((4 * half + c) ^ (gj & 7)) << 2)));
          v[4 * c] = x.x; v[4 * c + 1] = x.y; v[4 * c + 2] = x.z; v[4 * c + 3] = x.w;
        }
        tmem_st16(tb + col[o], v);
      }
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(res_full);
    }
    auto promote = [&]() {                                  // channels [16 half, 16 half + 16) of this row
      float a[16], s[16];
      tmem_ld16(tb + Cfg::kV, a);
      tmem_ld16(tb + Cfg::kK, s);
      tmem_wait_ld();
#pragma unroll
      for (int q = 0; q < 16; q += 2) {
        add2_into(dv[q], dv[q + 1], a[q], a[q + 1]);
        add2_into(dk[q], dk[q + 1], s[q], s[q + 1]);
      }
    };
    for (int j = 0; j < TQ; j++) {
      { ATR_T0();
      mbar_wait(&sdp_full[j & 1], (j >> 1) & 1);
      ATR_ADD(5); }
      tc_fence_after();
      ATR_T0();
      float a[16], g[16];
      tmem_ld16(tb + Cfg::kS + (j & 1) * BT, a);
      tmem_ld16(tb + Cfg::kP + (j & 1) * BT, g);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&s_free[j & 1]);
      mbar_wait(&q_full[j % Cfg::kStages], (j / Cfg::kStages) & 1);   // lse / delta of the tile have landed
      const float* rw = sRows + (j % Cfg::kStages) * 2 * BT + half * 16;
      float phi[16], plo[16], dhi[16], dlo[16];
#pragma unroll
      for (int q = 0; q < 16; q += 2) {                      // packed f32x2 arithmetic: same IEEE operations, half the
        const float2 x = __fmul2_rn(sub2(make_float2(a[q], a[q + 1]), make_float2(rw[q], rw[q + 1])),   // issue slots
                                    make_float2(kLog2e, kLog2e));
        const float2 p = make_float2(ex2_fast(x.x), ex2_fast(x.y));
        split_tf32x2(p, phi[q], phi[q + 1], plo[q], plo[q + 1]);
        const float2 ds = __fmul2_rn(p, sub2(make_float2(g[q], g[q + 1]), make_float2(rw[BT + q], rw[BT + q + 1])));
        split_tf32x2(ds, dhi[q], dhi[q + 1], dlo[q], dlo[q + 1]);
      }
      ATR_ADD(6);
      if (j > 0) {                                           // dV(j-1), dK(j-1) complete: promote, then reuse P^T / dS^T
        { ATR_T0();
        mbar_wait(pv_done, (j - 1) & 1);
        ATR_ADD(7); }
        tc_fence_after();
        promote();
      }
      { ATR_T0();
      tmem_st16(tb + Cfg::kPhi, phi);
      tmem_st16(tb + Cfg::kPlo, plo);
      tmem_st16(tb + Cfg::kDhi, dhi);
      tmem_st16(tb + Cfg::kDlo, dlo);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(ds_full);
      ATR_ADD(8); }
    }
    mbar_wait(pv_done, (TQ - 1) & 1);
    tc_fence_after();
    promote();
    if (gj < N) {
      const int c0 = cg * 32 + half * 16;                  // first channel of this thread
      float* dv_dst = dVp + ((long)b * N + gj) * d + c0;
      float* dk_dst = dKp + ((long)b * N + gj) * d + c0;
      if ((d & 3) == 0) {
#pragma unroll
        for (int c = 0; c < 16; c += 4)
          if (c0 + c < d) {
            *reinterpret_cast<float4*>(dv_dst + c) = make_float4(dv[c], dv[c + 1], dv[c + 2], dv[c + 3]);
            *reinterpret_cast<float4*>(dk_dst + c) = make_float4(dk[c], dk[c + 1], dk[c + 2], dk[c + 3]);
          }
      } else {
#pragma unroll
        for (int c = 0; c < 16; c++)
          if (c0 + c < d) { dv_dst[c] = dv[c]; dk_dst[c] = dk[c]; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------------------
// backward, part 2 for 32 < d <= 64: the second design with ALL 64 channels of dV / dK in one CTA (the
// ping-pong kernel ran one CTA per 32-channel group and recomputed S^T / dP^T in each).  TMEM holds K
// (tf32 hi / lo, 128 columns) as the A operand of S^T; V stays in shared memory (dP^T in the SS form) -
// there is no room for both:
//   S^T[2] 0,32 | dP^T[2] 64,96 | P^T hi 128, lo 160 | dS^T hi 192, lo 224 | dV 256 (64) | dK 320 (64) |
//   K hi 384 (64) | K lo 448 (64)                                                       = 512 columns.
// Per 32-query step: 24 TS (S^T) + 24 SS (dP^T) at N = 32 and 2 x 12 TS at N = 64 = 2112 cycles of
// tcgen05.mma time against 2 x 2304 in the ping-pong kernel.
// ------------------------------------------------------------------------------------------------
struct Dkv3Cfg {
  static constexpr int DPAD = 64, kSub = 2;
  static constexpr int kOperBytes = kSub * 128 * 128;          // V pos tile (128 keys), hi or lo
  static constexpr int kResBytes = 2 * kOperBytes;             // V hi, V lo
  static constexpr int kPosBytes = kSub * BT * 128;            // streamed pos tile, hi or lo
  static constexpr int kChanBytes = DPAD * 128;                // streamed chan block (64 channels x 32 queries), hi or lo
  static constexpr int kStageBytes = 4 * kPosBytes + 4 * kChanBytes;  // Q pos, dY pos, dY chan, Q chan (hi/lo each)
  static constexpr int kStages = 2;
  static constexpr int kRowBytes = kStages * 2 * BT * 4;
  static constexpr int kSmem = kResBytes + kStages * kStageBytes + kRowBytes + 1024 + 256;
  static constexpr int kThreads = 11 * 32;                     // warps 0-7 compute, 8 S^T/dP^T issue, 9 producer, 10 dV/dK issue
  static constexpr int kS = 0, kP = 64, kPhi = 128, kPlo = 160, kDhi = 192, kDlo = 224, kV = 256, kK = 320,
                       kKh = 384, kKl = 448;
};

__global__ void __launch_bounds__(Dkv3Cfg::kThreads, 1)
attn_dkv3_tc_kernel(int N, int NP, int d, SplitImage Qi, SplitImage Ki, SplitImage Vi, SplitImage dYi,
                    const float* __restrict__ lse_pad, const float* __restrict__ delta_pad,
                    float* __restrict__ dKp, float* __restrict__ dVp) {
  using Cfg = Dkv3Cfg;
  constexpr int DPAD = Cfg::DPAD;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sRes = smem;
  uint8_t* sStage = smem + Cfg::kResBytes;
  float* sRows = reinterpret_cast<float*>(sStage + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sRows) + Cfg::kRowBytes);
  uint64_t* res_full = bars;                       // 1   K is in TMEM (256 arrivals)
  uint64_t* v_full = res_full + 1;                 // 1   V is in shared memory
  uint64_t* q_full = v_full + 1;                   // kStages
  uint64_t* q_empty = q_full + Cfg::kStages;       // kStages
  uint64_t* sdp_full = q_empty + Cfg::kStages;     // 2
  uint64_t* s_free = sdp_full + 2;                 // 2
  uint64_t* ds_full = s_free + 2;                  // 1
  uint64_t* pv_done = ds_full + 1;                 // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, kt = blockIdx.x;
  const int TQ = (N + BT - 1) / BT;

  if (tid == 0) {
    mbar_init(res_full, 256); mbar_init(v_full, 1); mbar_init(ds_full, 256); mbar_init(pv_done, 1);
    for (int s = 0; s < 2; s++) { mbar_init(&sdp_full[s], 1); mbar_init(&s_free[s], 256); }
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
    mbar_init_fence();
  }
  if (warp == 8) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 9) {
    // ======================= bulk-copy producer =======================
    if (lane == 0) {
      mbar_expect_tx(v_full, Cfg::kResBytes);
#pragma unroll
      for (int s = 0; s < Cfg::kSub; s++) {
        const long o = (((long)b * Cfg::kSub + s) * NP + (long)kt * 128) * 32;
        bulk_g2s(sRes + s * 16384, Vi.pos_hi + o, 16384, v_full);
        bulk_g2s(sRes + Cfg::kOperBytes + s * 16384, Vi.pos_lo + o, 16384, v_full);
      }
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < TQ; j++) {
        mbar_wait(&q_empty[stage], phase ^ 1);
        uint8_t* st = sStage + stage * Cfg::kStageBytes;
        mbar_expect_tx(&q_full[stage], Cfg::kStageBytes + 2 * BT * 4);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)j * BT) * 32;
          bulk_g2s(st + s * BT * 128, Qi.pos_hi + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + Cfg::kPosBytes + s * BT * 128, Qi.pos_lo + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + 2 * Cfg::kPosBytes + s * BT * 128, dYi.pos_hi + o, BT * 128, &q_full[stage]);
          bulk_g2s(st + 3 * Cfg::kPosBytes + s * BT * 128, dYi.pos_lo + o, BT * 128, &q_full[stage]);
        }
        const long oc = (((long)b * (NP / 32) + j) * DPAD) * 32;
        uint8_t* sc = st + 4 * Cfg::kPosBytes;
        bulk_g2s(sc, dYi.chan_hi + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + Cfg::kChanBytes, dYi.chan_lo + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + 2 * Cfg::kChanBytes, Qi.chan_hi + oc, Cfg::kChanBytes, &q_full[stage]);
        bulk_g2s(sc + 3 * Cfg::kChanBytes, Qi.chan_lo + oc, Cfg::kChanBytes, &q_full[stage]);
        float* rw = sRows + stage * 2 * BT;
        bulk_g2s(rw, lse_pad + (long)b * NP + (long)j * BT, BT * 4, &q_full[stage]);
        bulk_g2s(rw + BT, delta_pad + (long)b * NP + (long)j * BT, BT * 4, &q_full[stage]);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 8) {
    // ======================= S^T = K.Q^T (TS), dP^T = V.dY^T (SS) issue =======================
    constexpr uint32_t idesc = umma_idesc_tf32(128, BT);
    const uint32_t r = smem_u32(sRes);
    mbar_wait(res_full, 0);
    mbar_wait(v_full, 0);
    tc_fence_after();
    for (int t = 0; t < TQ; t++) {
      mbar_wait(&q_full[t % Cfg::kStages], (t / Cfg::kStages) & 1);
      if (t >= 2) mbar_wait(&s_free[t & 1], ((t - 2) >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t st = smem_u32(sStage + (t % Cfg::kStages) * Cfg::kStageBytes);
        const uint32_t dS = tmem + Cfg::kS + (t & 1) * BT, dP = tmem + Cfg::kP + (t & 1) * BT;
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          mma_corr_ts(dS, tmem + Cfg::kKh + 32 * s, tmem + Cfg::kKl + 32 * s, st + s * BT * 128,
                      st + Cfg::kPosBytes + s * BT * 128, idesc, s == 0);
          mma_corr_ss(dP, r + s * 16384, r + Cfg::kOperBytes + s * 16384, st + 2 * Cfg::kPosBytes + s * BT * 128,
                      st + 3 * Cfg::kPosBytes + s * BT * 128, idesc, s == 0);
        }
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          mma_main_ts(dS, tmem + Cfg::kKh + 32 * s, st + s * BT * 128, idesc);
          mma_main_ss(dP, r + s * 16384, st + 2 * Cfg::kPosBytes + s * BT * 128, idesc);
        }
        umma_commit(&sdp_full[t & 1]);
      }
      __syncwarp();
    }
  } else if (warp == 10) {
    // ======================= dV = P^T.dY, dK = dS^T.Q issue (N = 64 channels) =======================
    constexpr uint32_t idesc = umma_idesc_tf32(128, DPAD);
    for (int j = 0; j < TQ; j++) {
      mbar_wait(ds_full, j & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sc = smem_u32(sStage + (j % Cfg::kStages) * Cfg::kStageBytes + 4 * Cfg::kPosBytes);
        mma2_ts(tmem + Cfg::kV, tmem + Cfg::kPhi, tmem + Cfg::kPlo, sc, sc + Cfg::kChanBytes,
                tmem + Cfg::kK, tmem + Cfg::kDhi, tmem + Cfg::kDlo, sc + 2 * Cfg::kChanBytes, sc + 3 * Cfg::kChanBytes,
                idesc);
        umma_commit(pv_done);
        umma_commit(&q_empty[j % Cfg::kStages]);
      }
      __syncwarp();
    }
  } else {
    // ======================= P^T, dS^T + promotion (thread = key row) =======================
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    const int gj = kt * 128 + row;
    const uint32_t tl = tmem + (((uint32_t)(quad * 32)) << 16);   // this thread's TMEM lane
    const uint32_t tb = tl + half * 16;                           // ... and its 16 query columns
    {
      // K (hi, lo), channels [32 half, 32 half + 32) of this key row: packed image (sub-tile = half) -> TMEM
      const float* src[2] = {Ki.pos_hi, Ki.pos_lo};
      const int col[2] = {Cfg::kKh, Cfg::kKl};
#pragma unroll
      for (int o = 0; o < 2; o++) {
        const float* rp = src[o] + (((long)b * Cfg::kSub + half) * NP + gj) * 32;
#pragma unroll
        for (int c16 = 0; c16 < 2; c16++) {
          float v[16];
#pragma unroll
          for (int c = 0; c < 4; c++) {
            const float4 x = __ldg(reinterpret_cast<const float4*>(rp + (((4 * c16 + c) ^ (gj & 7)) << 2)));
            v[4 * c] = x.x; v[4 * c + 1] = x.y; v[4 * c + 2] = x.z; v[4 * c + 3] = x.w;
          }
          tmem_st16(tl + col[o] + 32 * half + 16 * c16, v);
        }
      }
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(res_full);
    }
    float dv[32], dk[32];                                    // channels [32 half, 32 half + 32) of this row
#pragma unroll
    for (int c = 0; c < 32; c++) { dv[c] = 0.f; dk[c] = 0.f; }
    auto promote = [&]() {
#pragma unroll
      for (int c0 = 0; c0 < 32; c0 += 16) {
        float a[16], s2[16];
        tmem_ld16(tl + Cfg::kV + 32 * half + c0, a);
        tmem_ld16(tl + Cfg::kK + 32 * half + c0, s2);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 16; q += 2) {
          add2_into(dv[c0 + q], dv[c0 + q + 1], a[q], a[q + 1]);
          add2_into(dk[c0 + q], dk[c0 + q + 1], s2[q], s2[q + 1]);
        }
      }
    };
    for (int j = 0; j < TQ; j++) {
      mbar_wait(&sdp_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      float a[16], g[16];
      tmem_ld16(tb + Cfg::kS + (j & 1) * BT, a);
      tmem_ld16(tb + Cfg::kP + (j & 1) * BT, g);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&s_free[j & 1]);
      mbar_wait(&q_full[j % Cfg::kStages], (j / Cfg::kStages) & 1);   // lse / delta of the tile have landed
      const float* rw = sRows + (j % Cfg::kStages) * 2 * BT + half * 16;
      float phi[16], plo[16], dhi[16], dlo[16];
#pragma unroll
      for (int q = 0; q < 16; q += 2) {                      // packed f32x2, as in attn_dkv2_tc_kernel
        const float2 x = __fmul2_rn(sub2(make_float2(a[q], a[q + 1]), make_float2(rw[q], rw[q + 1])),
                                    make_float2(kLog2e, kLog2e));
        const float2 p = make_float2(ex2_fast(x.x), ex2_fast(x.y));
        split_tf32x2(p, phi[q], phi[q + 1], plo[q], plo[q + 1]);
        const float2 ds = __fmul2_rn(p, sub2(make_float2(g[q], g[q + 1]), make_float2(rw[BT + q], rw[BT + q + 1])));
        split_tf32x2(ds, dhi[q], dhi[q + 1], dlo[q], dlo[q + 1]);
      }
      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1);
        tc_fence_after();
        promote();
      }
      tmem_st16(tb + Cfg::kPhi, phi);
      tmem_st16(tb + Cfg::kPlo, plo);
      tmem_st16(tb + Cfg::kDhi, dhi);
      tmem_st16(tb + Cfg::kDlo, dlo);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(ds_full);
    }
    mbar_wait(pv_done, (TQ - 1) & 1);
    tc_fence_after();
    promote();
    if (gj < N) {
      const int c0 = half * 32;
      float* dv_dst = dVp + ((long)b * N + gj) * d + c0;
      float* dk_dst = dKp + ((long)b * N + gj) * d + c0;
      if ((d & 3) == 0) {
#pragma unroll
        for (int c = 0; c < 32; c += 4)
          if (c0 + c < d) {
            *reinterpret_cast<float4*>(dv_dst + c) = make_float4(dv[c], dv[c + 1], dv[c + 2], dv[c + 3]);
            *reinterpret_cast<float4*>(dk_dst + c) = make_float4(dk[c], dk[c + 1], dk[c + 2], dk[c + 3]);
          }
      } else {
#pragma unroll
        for (int c = 0; c < 32; c++)
          if (c0 + c < d) { dv_dst[c] = dv[c]; dk_dst[c] = dk[c]; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------------------
// backward, part 1, second design (d <= 32): the mirror image of attn_dkv2_tc_kernel.  CTA = 128 queries
// (lanes = queries), Q and dY resident in TMEM as the A operands of S = Q.K^T and dP = dY.V^T (TS form),
// 32-key tiles streamed through shared memory (K pos, V pos, K chan), S / dP double buffered, eight
// compute warps (16 key columns each), one warp issuing S / dP and one issuing dQ = dS.K.
// TMEM: S[2] 0,32 | dP[2] 64,96 | dS hi 128, lo 160 | dQ 192 | Q hi 224, lo 256 | dY hi 288, lo 320.
// ------------------------------------------------------------------------------------------------
template <int DPAD> struct Dq2Cfg {
  static constexpr int kSub = DPAD / 32;
  static constexpr int kPosBytes = kSub * BT * 128;            // streamed pos tile (32 keys x DPAD channels), hi or lo
  static constexpr int kChanBytes = DPAD * 128;                // streamed chan block (DPAD channels x 32 keys), hi or lo
  static constexpr int kStageBytes = 4 * kPosBytes + 2 * kChanBytes;   // K pos hi/lo, V pos hi/lo, K chan hi/lo
  static constexpr int kStages = DPAD == 32 ? 8 : 4;
  static constexpr int kSmem = kStages * kStageBytes + 1024 + 256;
  static constexpr int kThreads = 11 * 32;                     // warps 0-7 compute, 8 S/dP issue, 9 producer, 10 dQ issue
  // TMEM columns (d = 64 uses all 512)
  static constexpr int kS = 0, kP = 64, kDhi = 128, kDlo = 160, kAcc = 192, kQh = 192 + DPAD, kQl = kQh + DPAD,
                       kYh = kQl + DPAD, kYl = kYh + DPAD;
  static constexpr int kCPT = DPAD / 2;                        // channels per compute thread (half a row)
};

template <int DPAD>
__global__ void __launch_bounds__(Dq2Cfg<DPAD>::kThreads, 1)
attn_dq2_tc_kernel(int N, int NP, int d, SplitImage Qi, SplitImage Ki, SplitImage Vi, SplitImage dYi,
                   const float* __restrict__ lse_pad, const float* __restrict__ delta_pad,
                   float* __restrict__ dQp) {
  using Cfg = Dq2Cfg<DPAD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sStage = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStage + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* res_full = bars;                       // 1   Q, dY are in TMEM (256 arrivals)
  uint64_t* kv_full = res_full + 1;                // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;     // kStages
  uint64_t* sdp_full = kv_empty + Cfg::kStages;    // 2   S, dP of a tile are in TMEM buffer (j & 1)
  uint64_t* s_free = sdp_full + 2;                 // 2   ... and have been read into registers (256 arrivals)
  uint64_t* ds_full = s_free + 2;                  // 1   dS written, previous dQ tile promoted (256)
  uint64_t* dq_done = ds_full + 1;                 // 1   dQ MMAs of a tile complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dq_done + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, qt = blockIdx.x;          // query tile
  const int TK = (N + BT - 1) / BT;                   // key tiles

  if (tid == 0) {
    mbar_init(res_full, 256); mbar_init(ds_full, 256); mbar_init(dq_done, 1);
    for (int s = 0; s < 2; s++) { mbar_init(&sdp_full[s], 1); mbar_init(&s_free[s], 256); }
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init_fence();
  }
  if (warp == 8) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 9) {
    // ======================= bulk-copy producer =======================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < TK; j++) {
        mbar_wait(&kv_empty[stage], phase ^ 1);
        uint8_t* st = sStage + stage * Cfg::kStageBytes;
        mbar_expect_tx(&kv_full[stage], Cfg::kStageBytes);
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          const long o = (((long)b * Cfg::kSub + s) * NP + (long)j * BT) * 32;
          bulk_g2s(st + s * BT * 128, Ki.pos_hi + o, BT * 128, &kv_full[stage]);
          bulk_g2s(st + Cfg::kPosBytes + s * BT * 128, Ki.pos_lo + o, BT * 128, &kv_full[stage]);
          bulk_g2s(st + 2 * Cfg::kPosBytes + s * BT * 128, Vi.pos_hi + o, BT * 128, &kv_full[stage]);
          bulk_g2s(st + 3 * Cfg::kPosBytes + s * BT * 128, Vi.pos_lo + o, BT * 128, &kv_full[stage]);
        }
        const long oc = (((long)b * (NP / 32) + j) * DPAD) * 32;
        bulk_g2s(st + 4 * Cfg::kPosBytes, Ki.chan_hi + oc, Cfg::kChanBytes, &kv_full[stage]);
        bulk_g2s(st + 4 * Cfg::kPosBytes + Cfg::kChanBytes, Ki.chan_lo + oc, Cfg::kChanBytes, &kv_full[stage]);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 8) {
    // ======================= S = Q.K^T, dP = dY.V^T issue =======================
    constexpr uint32_t idesc = umma_idesc_tf32(128, BT);
    mbar_wait(res_full, 0);
    tc_fence_after();
    for (int t = 0; t < TK; t++) {
      mbar_wait(&kv_full[t % Cfg::kStages], (t / Cfg::kStages) & 1);
      if (t >= 2) mbar_wait(&s_free[t & 1], ((t - 2) >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t st = smem_u32(sStage + (t % Cfg::kStages) * Cfg::kStageBytes);
        const uint32_t dS = tmem + Cfg::kS + (t & 1) * BT, dP = tmem + Cfg::kP + (t & 1) * BT;
        // corrections of every 32-channel sub-tile first, then the main terms (one accumulation chain per tile)
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          mma_corr_ts(dS, tmem + Cfg::kQh + 32 * s, tmem + Cfg::kQl + 32 * s, st + s * BT * 128,
                      st + Cfg::kPosBytes + s * BT * 128, idesc, s == 0);
          mma_corr_ts(dP, tmem + Cfg::kYh + 32 * s, tmem + Cfg::kYl + 32 * s, st + 2 * Cfg::kPosBytes + s * BT * 128,
                      st + 3 * Cfg::kPosBytes + s * BT * 128, idesc, s == 0);
        }
#pragma unroll
        for (int s = 0; s < Cfg::kSub; s++) {
          mma_main_ts(dS, tmem + Cfg::kQh + 32 * s, st + s * BT * 128, idesc);
          mma_main_ts(dP, tmem + Cfg::kYh + 32 * s, st + 2 * Cfg::kPosBytes + s * BT * 128, idesc);
        }
        umma_commit(&sdp_full[t & 1]);
      }
      __syncwarp();
    }
  } else if (warp == 10) {
    // ======================= dQ = dS.K issue =======================
    constexpr uint32_t idesc = umma_idesc_tf32(128, DPAD);
    for (int j = 0; j < TK; j++) {
      mbar_wait(ds_full, j & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t kc = smem_u32(sStage + (j % Cfg::kStages) * Cfg::kStageBytes + 4 * Cfg::kPosBytes);
        mma_corr_ts(tmem + Cfg::kAcc, tmem + Cfg::kDhi, tmem + Cfg::kDlo, kc, kc + Cfg::kChanBytes, idesc, true);
        mma_main_ts(tmem + Cfg::kAcc, tmem + Cfg::kDhi, kc, idesc);
        umma_commit(dq_done);
        umma_commit(&kv_empty[j % Cfg::kStages]);    // S(j), dP(j) completed before ds_full(j) could fire
      }
      __syncwarp();
    }
  } else {
    // ======================= dS + promotion (thread = query row, half of the key columns) =======================
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    const int gi = qt * 128 + row;
    const uint32_t tl = tmem + (((uint32_t)(quad * 32)) << 16);          // this thread's TMEM lane
    const uint32_t tb = tl + half * 16;                                  // ... and its 16 key columns
    constexpr int CPT = Cfg::kCPT;
    {
      // this thread's query row, channels [CPT half, CPT half + CPT) of Q and dY (hi, lo): packed image -> TMEM.
      // Image rows are SWIZZLE_128B per 32-channel sub-tile: chunk c of row n sits at chunk position c ^ (n & 7).
      const float* src[4] = {Qi.pos_hi, Qi.pos_lo, dYi.pos_hi, dYi.pos_lo};
      const int col[4] = {Cfg::kQh, Cfg::kQl, Cfg::kYh, Cfg::kYl};
#pragma unroll
      for (int o = 0; o < 4; o++) {
#pragma unroll
        for (int c16 = 0; c16 < CPT / 16; c16++) {
          const int ch0 = half * CPT + c16 * 16;                         // first channel of this 16-column piece
          const int sub = ch0 >> 5, chunk0 = (ch0 & 31) >> 2;
          const float* rp = src[o] + (((long)b * Cfg::kSub + sub) * NP + gi) * 32;
          float v[16];
#pragma unroll
          for (int c = 0; c < 4; c++) {
            const float4 x = __ldg(reinterpret_cast<const float4*>(rp + (((chunk0 + c) ^ (gi & 7)) << 2)));
            v[4 * c] = x.x; v[4 * c + 1] = x.y; v[4 * c + 2] = x.z; v[4 * c + 3] = x.w;
          }
          tmem_st16(tl + col[o] + ch0, v);
        }
      }
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(res_full);
    }
    const float lse_i = __ldg(lse_pad + (long)b * NP + gi);
    const float delta_i = __ldg(delta_pad + (long)b * NP + gi);
    float dq[CPT];
#pragma unroll
    for (int c = 0; c < CPT; c++) dq[c] = 0.f;
    auto promote = [&]() {                                  // channels [CPT half, CPT half + CPT) of this row
#pragma unroll
      for (int c0 = 0; c0 < CPT; c0 += 16) {
        float a[16];
        tmem_ld16(tl + Cfg::kAcc + half * CPT + c0, a);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 16; q += 2) add2_into(dq[c0 + q], dq[c0 + q + 1], a[q], a[q + 1]);
      }
    };
    for (int j = 0; j < TK; j++) {
      mbar_wait(&sdp_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      float a[16], g[16];
      tmem_ld16(tb + Cfg::kS + (j & 1) * BT, a);
      tmem_ld16(tb + Cfg::kP + (j & 1) * BT, g);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&s_free[j & 1]);
      const int j0 = j * BT + half * 16;                      // first key of this thread's columns
      float hi[16], lo[16];
      const float2 lse2 = make_float2(lse_i, lse_i), delta2 = make_float2(delta_i, delta_i);
#pragma unroll
      for (int q = 0; q < 16; q += 2) {                      // packed f32x2 arithmetic (same IEEE operations)
        const float2 x = __fmul2_rn(sub2(make_float2(a[q], a[q + 1]), lse2), make_float2(kLog2e, kLog2e));
        float2 p = make_float2(ex2_fast(x.x), ex2_fast(x.y));
        if (j0 + q >= N) p.x = 0.f;
        if (j0 + q + 1 >= N) p.y = 0.f;
        const float2 ds = __fmul2_rn(p, sub2(make_float2(g[q], g[q + 1]), delta2));
        split_tf32x2(ds, hi[q], hi[q + 1], lo[q], lo[q + 1]);
      }
      if (j > 0) {                                           // dQ(j-1) complete: promote, then reuse the dS columns
        mbar_wait(dq_done, (j - 1) & 1);
        tc_fence_after();
        promote();
      }
      tmem_st16(tb + Cfg::kDhi, hi);
      tmem_st16(tb + Cfg::kDlo, lo);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(ds_full);
    }
    mbar_wait(dq_done, (TK - 1) & 1);
    tc_fence_after();
    promote();
    if (gi < N) {
      const int c0 = half * CPT;
      float* dst = dQp + ((long)b * N + gi) * d + c0;
      if ((d & 3) == 0) {
#pragma unroll
        for (int c = 0; c < CPT; c += 4)
          if (c0 + c < d) *reinterpret_cast<float4*>(dst + c) = make_float4(dq[c], dq[c + 1], dq[c + 2], dq[c + 3]);
      } else {
#pragma unroll
        for (int c = 0; c < CPT; c++)
          if (c0 + c < d) dst[c] = dq[c];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
template <typename K>
int set_smem(K kern, int bytes) { return set_smem_once((const void*)kern, bytes); }

template <int DPAD>
int launch_pack(cudaStream_t st, int B, int d, int N, const float* X, const float* sub, long bs, long cs, long ns,
                const SplitImage& img) {
  const int NP = attn_tc_np(N);
  dim3 grid(NP / 32, B);
  pack_split_kernel<DPAD><<<grid, 256, 0, st>>>(X, sub, bs, cs, ns, d, N, NP, img.pos_hi, img.pos_lo, img.chan_hi,
                                                img.chan_lo);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

template <int DPAD, bool FAST>
int launch_fwd_tc(cudaStream_t st, int B, int N, int d, const SplitImage& Q, const SplitImage& K,
                  const SplitImage& V, float* Y, float* lse) {
  using Cfg = FwdCfg<DPAD>;
  auto kern = attn_fwd_tc_kernel<DPAD, FAST>;
  FRN_TRY(set_smem(kern, Cfg::kSmem));
  dim3 grid(cdiv(N, ABM), B);
  ScopedTimer tm(st, FRN_TIME_ATTN_FWD);
  kern<<<grid, Cfg::kThreads, Cfg::kSmem, st>>>(N, attn_tc_np(N), d, Q, K, V, Y, lse);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

// The dQ kernel and the dK/dV kernel are independent (same inputs, different outputs) and each leaves a
// ragged last wave (600 CTAs on 148 SMs at DSEC level 1 = 4.05 waves), so the dQ kernel is forked onto a side
// stream of the caller's stream (fork_stream, common.cuh) and joined back with events: the two tails fill each
// other.  With the measurement hooks on, the kernels run back to back on the caller's stream so that their times
// stay separable.
template <int DPAD, int NT>
int launch_bwd_tc(cudaStream_t st, int B, int N, int d, const SplitImage& Q, const SplitImage& K,
                  const SplitImage& V, const SplitImage& dY, const float* lse_pad, const float* delta_pad,
                  float* dQp, float* dKp, float* dVp) {
  const int NP = attn_tc_np(N);
  using CfgK2 = Dkv2Cfg<DPAD>;
  auto kern_kv2 = attn_dkv2_tc_kernel<DPAD>;
#ifdef FRN_DEBUG_ENGINES
  // first designs (two tiles per CTA in ping-pong, operands in shared memory, one issuing warp): debug builds only,
  // FRN_DQ_DESIGN=1 / FRN_DKV_DESIGN=1
  using CfgK = DkvCfg<DPAD, NT>;
  using CfgQ = DqCfg<DPAD, NT>;
  auto kern_kv = attn_dkv_tc_kernel<DPAD, NT>;
  auto kern_q = attn_dq_tc_kernel<DPAD, NT>;
  static const int dkv_design = [] { const char* e = getenv("FRN_DKV_DESIGN"); return (e && e[0] == '1') ? 1 : 2; }();
  static const int dq_design = [] { const char* e = getenv("FRN_DQ_DESIGN"); return (e && e[0] == '1') ? 1 : 2; }();
  FRN_TRY(set_smem(kern_kv, CfgK::kSmem));
  FRN_TRY(set_smem(kern_q, CfgQ::kSmem));
  const dim3 grid_kv(cdiv(NP / 128, NT), B, cdiv(d, 32)), grid_q(cdiv(NP / 128, NT), B);
#else
  constexpr int dkv_design = 2, dq_design = 2;
#endif
  FRN_TRY(set_smem(kern_kv2, CfgK2::kSmem));
  const dim3 grid_kv2(NP / 128, B, cdiv(d, 32));
  const bool dq2 = dq_design == 2;
  if (dq2) FRN_TRY(set_smem(attn_dq2_tc_kernel<DPAD>, Dq2Cfg<DPAD>::kSmem));
  const dim3 grid_q2(NP / 128, B);
  auto launch_q = [&](cudaStream_t s_) {
    if (dq2)
      attn_dq2_tc_kernel<DPAD><<<grid_q2, Dq2Cfg<DPAD>::kThreads, Dq2Cfg<DPAD>::kSmem, s_>>>(N, NP, d, Q, K, V, dY, lse_pad,
                                                                                              delta_pad, dQp);
#ifdef FRN_DEBUG_ENGINES
    else
      kern_q<<<grid_q, CfgQ::kThreads, CfgQ::kSmem, s_>>>(N, NP, d, Q, K, V, dY, lse_pad, delta_pad, dQp);
#endif
  };
  ScopedTimer tm(st, FRN_TIME_ATTN_BWD);
  const Fork fk = fork_stream(st, 0);
  if (fk.side) {
    launch_q(fk.side);
    FRN_LAUNCH_CHECK();
  }
  {
    ScopedTimer tk(st, FRN_TIME_ATTN_DKV);
    if (dkv_design == 2 && DPAD == 64) {
      FRN_TRY(set_smem(attn_dkv3_tc_kernel, Dkv3Cfg::kSmem));
      attn_dkv3_tc_kernel<<<dim3(NP / 128, B), Dkv3Cfg::kThreads, Dkv3Cfg::kSmem, st>>>(N, NP, d, Q, K, V, dY, lse_pad,
                                                                                          delta_pad, dKp, dVp);
    } else if (dkv_design == 2)
      kern_kv2<<<grid_kv2, CfgK2::kThreads, CfgK2::kSmem, st>>>(N, NP, d, Q, K, V, dY, lse_pad, delta_pad, dKp, dVp);
#ifdef FRN_DEBUG_ENGINES
    else
      kern_kv<<<grid_kv, CfgK::kThreads, CfgK::kSmem, st>>>(N, NP, d, Q, K, V, dY, lse_pad, delta_pad, dKp, dVp);
#endif
    FRN_LAUNCH_CHECK();
  }
  if (fk.side) {
    join_stream(fk);
  } else {
    launch_q(st);
    FRN_LAUNCH_CHECK();
  }
  return FRN_OK;
}

}  // namespace

int attn_tc_trace_read(unsigned long long* out16) {
#ifdef FRN_ATTN_TRACE
  unsigned long long z[16] = {0};
  cudaMemcpyFromSymbol(out16, g_attn_trace, sizeof(z));
  cudaMemcpyToSymbol(g_attn_trace, z, sizeof(z));
  return 1;
#else
  (void)out16;
  return 0;
#endif
}

int attn_tc_np(int N) { return (N + 127) / 128 * 128; }

size_t attn_tc_plane_floats(int B, int N, int d) {
  const int dpad = d <= 32 ? 32 : 64;
  return (size_t)B * dpad * attn_tc_np(N);
}

int attn_tc_pack(cudaStream_t st, int B, int d, int N, const float* X, const float* sub, long bs, long cs,
                 long ns, const SplitImage& img) {
  if (d > 64) return FRN_ERR_UNSUPPORTED;
  if ((img.pos_lo && !img.pos_hi) || (img.chan_lo && !img.chan_hi)) return FRN_ERR_NULL_POINTER;
  ScopedTimer tm(st, FRN_TIME_ROWS);
  const int planes = (img.pos_hi ? 1 : 0) + (img.pos_lo ? 1 : 0) + (img.chan_hi ? 1 : 0) + (img.chan_lo ? 1 : 0);
  KernelTimer kt(st, "pack_split_kernel", 4.0 * B * ((double)d * N + (double)planes * (d <= 32 ? 32 : 64) * attn_tc_np(N)));
  return d <= 32 ? launch_pack<32>(st, B, d, N, X, sub, bs, cs, ns, img)
                 : launch_pack<64>(st, B, d, N, X, sub, bs, cs, ns, img);
}

int attn_forward_tc(cudaStream_t st, int B, int N, int d, const SplitImage& Q, const SplitImage& K,
                    const SplitImage& V, float* Y, float* lse, int fast) {
  if (fast) {
    if (d <= 32) return launch_fwd_tc<32, true>(st, B, N, d, Q, K, V, Y, lse);
    if (d <= 64) return launch_fwd_tc<64, true>(st, B, N, d, Q, K, V, Y, lse);
    return FRN_ERR_UNSUPPORTED;
  }
  if (!Q.pos_lo || !K.pos_lo || !V.chan_lo) return FRN_ERR_NULL_POINTER;
  if (d <= 32) return launch_fwd_tc<32, false>(st, B, N, d, Q, K, V, Y, lse);
  if (d <= 64) return launch_fwd_tc<64, false>(st, B, N, d, Q, K, V, Y, lse);
  return FRN_ERR_UNSUPPORTED;
}

int attn_bwd_prep_tc(cudaStream_t st, int B, int N, int d, const float* dYp, const float* Y, const float* lse,
                     float* lse_pad, float* delta_pad) {
  const int NP = attn_tc_np(N);
  dim3 grid(cdiv(NP, 128), B);
  ScopedTimer tm(st, FRN_TIME_ROWS);
  attn_bwd_prep_kernel<<<grid, 128, 0, st>>>(N, NP, d, dYp, Y, lse, lse_pad, delta_pad);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int attn_backward_tc(cudaStream_t st, int B, int N, int d, const SplitImage& Q, const SplitImage& K,
                     const SplitImage& V, const SplitImage& dY, const float* lse_pad, const float* delta_pad,
                     float* dQp, float* dKp, float* dVp) {
  if (d <= 32) return launch_bwd_tc<32, 2>(st, B, N, d, Q, K, V, dY, lse_pad, delta_pad, dQp, dKp, dVp);
  if (d <= 64) return launch_bwd_tc<64, 1>(st, B, N, d, Q, K, V, dY, lse_pad, delta_pad, dQp, dKp, dVp);
  return FRN_ERR_UNSUPPORTED;
}

}  // namespace frn
