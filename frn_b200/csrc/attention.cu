// Streaming-softmax cross attention for CrossAttentionBlock
// (retinanet/Attention.py:29-32):  f = phi_x @ theta_x ; softmax(f, dim=-1) ;
// y = softmax @ g_x, with NO 1/sqrt(d) scale.  The N x N matrix (1.47 GB per
// sample at DSEC level 1) is never written to HBM: each CTA owns BM query
// rows, streams BN-key tiles through shared memory and keeps the running
// row max / row sum (online softmax) in registers.  Exact fp32 FFMA math.
//
// Two regimes:
//   d <= 64  (levels 1-2, N up to 19 200): the fused kernels in this file.
//   d  > 64  (levels 3-4, N <= 1 200):     N x N tiles are small (<= 5.8 MB);
//            cab.cu composes them from the GEMM kernel plus the row-softmax
//            kernels at the bottom of this file, one sample at a time.
#include "common.cuh"

#ifdef FRN_DEBUG_ENGINES      // FFMA streaming-softmax engine: debug builds only (A/B comparisons)
namespace frn {
namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr int kThreads = 256;

// smem tile [DPAD][T] from a channel-major (d, N) matrix, zero padded.
template <int DPAD, int T>
__device__ __forceinline__ void load_cmajor(float* dst, const float* __restrict__ src, int d, int N,
                                            int col0) {
  for (int idx = threadIdx.x; idx < DPAD * T; idx += kThreads) {
    int k = idx / T, i = idx % T, g = col0 + i;
    dst[idx] = (k < d && g < N) ? __ldg(src + (long)k * N + g) : 0.f;
  }
}
// smem tile [T][DPAD] from a position-major (N, d) matrix, zero padded.
template <int DPAD, int T>
__device__ __forceinline__ void load_pmajor(float* dst, const float* __restrict__ src, int d, int N,
                                            int row0) {
  for (int idx = threadIdx.x; idx < DPAD * T; idx += kThreads) {
    int r = idx / DPAD, c = idx % DPAD, g = row0 + r;
    dst[idx] = (c < d && g < N) ? __ldg(src + (long)g * d + c) : 0.f;
  }
}

template <int T>
__device__ __forceinline__ void load_frag(float (&v)[T], const float* base, int group_stride) {
  // T/4 float4 groups, group g starts at base + g*group_stride
#pragma unroll
  for (int g = 0; g < T / 4; g++) {
    float4 x = *reinterpret_cast<const float4*>(base + g * group_stride);
    v[g * 4] = x.x; v[g * 4 + 1] = x.y; v[g * 4 + 2] = x.z; v[g * 4 + 3] = x.w;
  }
}

// XOR-swizzled [J][I] tile (I contiguous, pitch I_PITCH) written 4 rows of I at a time.
__device__ __forceinline__ int swz(int j, int i, int pitch) {
  return j * pitch + ((((i >> 2) ^ ((j >> 2) & 7)) << 2) | (i & 3));
}

// ---------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------
template <int DPAD, int BM, int BN>
__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_kernel(int N, int d, const float* __restrict__ Qc, const float* __restrict__ Kc,
                const float* __restrict__ Vp, float* __restrict__ Y, float* __restrict__ lse) {
  constexpr int TSM = BM / 16, TSN = BN / 16;     // S register tile
  constexpr int GSM = BM / (TSM / 4), GSN = BN / (TSN / 4);
  constexpr int NCG = DPAD / 4, NRG = kThreads / NCG, RY = BM / NRG;  // Y register tile RY x 4
  static_assert(TSM % 4 == 0 && TSN % 4 == 0 && RY % 4 == 0, "tile shapes");
  extern __shared__ __align__(16) float smem[];
  float* Qs = smem;               // [DPAD][BM]
  float* Ks = Qs + DPAD * BM;     // [DPAD][BN]
  float* Vs = Ks + DPAD * BN;     // [BN][DPAD]
  float* Ps = Vs + BN * DPAD;     // [BN][BM] swizzled
  float* row_scale = Ps + BN * BM;  // [BM]

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int tcx = tid % NCG, trow = tid / NCG;
  const int b = blockIdx.y, i0 = blockIdx.x * BM;
  const float* Qb = Qc + (long)b * d * N;
  const float* Kb = Kc + (long)b * d * N;
  const float* Vb = Vp + (long)b * N * d;

  load_cmajor<DPAD, BM>(Qs, Qb, d, N, i0);

  float m[TSM], l[TSM], y[RY][4];
#pragma unroll
  for (int i = 0; i < TSM; i++) { m[i] = -INFINITY; l[i] = 0.f; }
#pragma unroll
  for (int r = 0; r < RY; r++) { y[r][0] = y[r][1] = y[r][2] = y[r][3] = 0.f; }

  for (int j0 = 0; j0 < N; j0 += BN) {
    load_cmajor<DPAD, BN>(Ks, Kb, d, N, j0);
    load_pmajor<DPAD, BN>(Vs, Vb, d, N, j0);
    __syncthreads();

    float s[TSM][TSN];
#pragma unroll
    for (int i = 0; i < TSM; i++)
#pragma unroll
      for (int j = 0; j < TSN; j++) s[i][j] = 0.f;
#pragma unroll 8
    for (int k = 0; k < DPAD; k++) {
      float a[TSM], bb[TSN];
      load_frag<TSM>(a, Qs + k * BM + ty * 4, GSM);
      load_frag<TSN>(bb, Ks + k * BN + tx * 4, GSN);
#pragma unroll
      for (int i = 0; i < TSM; i++)
#pragma unroll
        for (int j = 0; j < TSN; j++) s[i][j] = fmaf(a[i], bb[j], s[i][j]);
    }
    // online softmax
#pragma unroll
    for (int i = 0; i < TSM; i++) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < TSN; j++) {
        int gn = j0 + (j / 4) * GSN + tx * 4 + (j % 4);
        if (gn >= N) s[i][j] = -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      float m_new = fmaxf(m[i], mx);
      float sc = exp2f((m[i] - m_new) * kLog2e);
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < TSN; j++) {
        s[i][j] = exp2f((s[i][j] - m_new) * kLog2e);
        sum += s[i][j];
      }
      l[i] = l[i] * sc + sum;
      m[i] = m_new;
      if (tx == 0) row_scale[(i / 4) * GSM + ty * 4 + (i % 4)] = sc;
    }
    // P -> smem, [j][i] swizzled, 4 query rows per 128-bit store
#pragma unroll
    for (int j = 0; j < TSN; j++) {
      int jj = (j / 4) * GSN + tx * 4 + (j % 4);
#pragma unroll
      for (int g = 0; g < TSM / 4; g++) {
        int ii = g * GSM + ty * 4;
        *reinterpret_cast<float4*>(Ps + swz(jj, ii, BM)) =
            make_float4(s[g * 4][j], s[g * 4 + 1][j], s[g * 4 + 2][j], s[g * 4 + 3][j]);
      }
    }
    __syncthreads();
    // Y = Y * scale + P V
#pragma unroll
    for (int r = 0; r < RY; r++) {
      float sc = row_scale[trow * RY + r];
      y[r][0] *= sc; y[r][1] *= sc; y[r][2] *= sc; y[r][3] *= sc;
    }
#pragma unroll 4
    for (int j = 0; j < BN; j++) {
      float4 v = *reinterpret_cast<const float4*>(Vs + j * DPAD + tcx * 4);
#pragma unroll
      for (int q = 0; q < RY / 4; q++) {
        float4 p = *reinterpret_cast<const float4*>(Ps + swz(j, trow * RY + q * 4, BM));
        float pr[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
        for (int r = 0; r < 4; r++) {
          y[q * 4 + r][0] = fmaf(pr[r], v.x, y[q * 4 + r][0]);
          y[q * 4 + r][1] = fmaf(pr[r], v.y, y[q * 4 + r][1]);
          y[q * 4 + r][2] = fmaf(pr[r], v.z, y[q * 4 + r][2]);
          y[q * 4 + r][3] = fmaf(pr[r], v.w, y[q * 4 + r][3]);
        }
      }
    }
    __syncthreads();
  }

  // finalise: row sums across the 16 column threads, then normalise
#pragma unroll
  for (int i = 0; i < TSM; i++) {
    float t = l[i];
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    int row = (i / 4) * GSM + ty * 4 + (i % 4);
    if (tx == 0) {
      row_scale[row] = 1.f / t;
      if (i0 + row < N) lse[(long)b * N + i0 + row] = m[i] + logf(t);
    }
  }
  __syncthreads();
  float* Yb = Y + (long)b * N * d;
  const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(Yb) & 15) == 0);
#pragma unroll
  for (int r = 0; r < RY; r++) {
    int row = trow * RY + r, gi = i0 + row;
    if (gi >= N) continue;
    float inv = row_scale[row];
    float o[4] = {y[r][0] * inv, y[r][1] * inv, y[r][2] * inv, y[r][3] * inv};
    int c = tcx * 4;
    if (vec && c + 3 < d) {
      *reinterpret_cast<float4*>(Yb + (long)gi * d + c) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
#pragma unroll
      for (int q = 0; q < 4; q++)
        if (c + q < d) Yb[(long)gi * d + c + q] = o[q];
    }
  }
}

template <int DPAD, int BM, int BN>
int launch_fwd(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Kc,
               const float* Vp, float* Y, float* lse) {
  size_t smem = (size_t)(DPAD * BM + 2 * DPAD * BN + BN * BM + BM) * sizeof(float);
  auto kern = attn_fwd_kernel<DPAD, BM, BN>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  dim3 grid(cdiv(N, BM), B);
  ScopedTimer tm(st, FRN_TIME_ATTN_FWD);
  kern<<<grid, kThreads, smem, st>>>(N, d, Qc, Kc, Vp, Y, lse);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}


// ---------------------------------------------------------------------------
// backward: one CTA per BN-key tile, looping over BM-query tiles.
//   S = Q K ; P = exp(S - lse) ; dP = dY V^T ; dS = P * (dP - delta)
//   dV += P^T dY ; dK += dS^T Q   (registers, written once)
//   dQ += dS K^T                   (fp32 atomics into a zeroed buffer)
// ---------------------------------------------------------------------------
template <int DPAD, int BM, int BN>
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_kernel(int N, int d, const float* __restrict__ Qc, const float* __restrict__ Qp,
                const float* __restrict__ Kc, const float* __restrict__ Kp,
                const float* __restrict__ Vc, const float* __restrict__ dYc,
                const float* __restrict__ dYp, const float* __restrict__ lse,
                const float* __restrict__ delta, float* __restrict__ dQp,
                float* __restrict__ dKp, float* __restrict__ dVp) {
  constexpr int TSM = BM / 16, TSN = BN / 16;
  constexpr int GSN = BN / (TSN / 4);
  constexpr int NCG = DPAD / 4, NRG = kThreads / NCG;
  constexpr int RJ = BN / NRG;   // dK/dV rows per thread
  constexpr int RI = BM / NRG;   // dQ rows per thread
  static_assert(TSM == 4 && TSN % 4 == 0 && RJ == 4 && (RI == 2 || RI == 4), "tile shapes");
  extern __shared__ __align__(16) float smem[];
  float* Kcs = smem;                 // [DPAD][BN]
  float* Kps = Kcs + DPAD * BN;      // [BN][DPAD]
  float* Vcs = Kps + BN * DPAD;      // [DPAD][BN]
  float* Qcs = Vcs + DPAD * BN;      // [DPAD][BM]
  float* Qps = Qcs + DPAD * BM;      // [BM][DPAD]
  float* dYcs = Qps + BM * DPAD;     // [DPAD][BM]
  float* dYps = dYcs + DPAD * BM;    // [BM][DPAD]
  float* Ps = dYps + BM * DPAD;      // [BM][BN]
  float* dSs = Ps + BM * BN;         // [BM][BN]
  float* dSt = dSs + BM * BN;        // [BN][BM] swizzled
  float* lse_s = dSt + BN * BM;      // [BM]
  float* del_s = lse_s + BM;         // [BM]

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int tcx = tid % NCG, trow = tid / NCG;
  const int b = blockIdx.y, j0 = blockIdx.x * BN;
  const long bo = (long)b * d * N;

  load_cmajor<DPAD, BN>(Kcs, Kc + bo, d, N, j0);
  load_pmajor<DPAD, BN>(Kps, Kp + bo, d, N, j0);
  load_cmajor<DPAD, BN>(Vcs, Vc + bo, d, N, j0);

  float dv[RJ][4], dk[RJ][4];
#pragma unroll
  for (int r = 0; r < RJ; r++)
#pragma unroll
    for (int c = 0; c < 4; c++) { dv[r][c] = 0.f; dk[r][c] = 0.f; }

  for (int i0 = 0; i0 < N; i0 += BM) {
    load_cmajor<DPAD, BM>(Qcs, Qc + bo, d, N, i0);
    load_pmajor<DPAD, BM>(Qps, Qp + bo, d, N, i0);
    load_cmajor<DPAD, BM>(dYcs, dYc + bo, d, N, i0);
    load_pmajor<DPAD, BM>(dYps, dYp + bo, d, N, i0);
    if (tid < BM) {
      int gi = i0 + tid;
      lse_s[tid] = gi < N ? lse[(long)b * N + gi] : 0.f;
      del_s[tid] = gi < N ? delta[(long)b * N + gi] : 0.f;
    }
    __syncthreads();

    float s[TSM][TSN], dp[TSM][TSN];
#pragma unroll
    for (int i = 0; i < TSM; i++)
#pragma unroll
      for (int j = 0; j < TSN; j++) { s[i][j] = 0.f; dp[i][j] = 0.f; }
#pragma unroll 4
    for (int k = 0; k < DPAD; k++) {
      float a[TSM], a2[TSM], bb[TSN], b2[TSN];
      load_frag<TSM>(a, Qcs + k * BM + ty * 4, 0);
      load_frag<TSM>(a2, dYcs + k * BM + ty * 4, 0);
      load_frag<TSN>(bb, Kcs + k * BN + tx * 4, GSN);
      load_frag<TSN>(b2, Vcs + k * BN + tx * 4, GSN);
#pragma unroll
      for (int i = 0; i < TSM; i++)
#pragma unroll
        for (int j = 0; j < TSN; j++) {
          s[i][j] = fmaf(a[i], bb[j], s[i][j]);
          dp[i][j] = fmaf(a2[i], b2[j], dp[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < TSM; i++) {
      int row = ty * 4 + i;
      float ls = lse_s[row], de = del_s[row];
      bool rv = (i0 + row) < N;
#pragma unroll
      for (int j = 0; j < TSN; j++) {
        int gn = j0 + (j / 4) * GSN + tx * 4 + (j % 4);
        float p = (rv && gn < N) ? exp2f((s[i][j] - ls) * kLog2e) : 0.f;
        s[i][j] = p;
        dp[i][j] = p * (dp[i][j] - de);
      }
#pragma unroll
      for (int g = 0; g < TSN / 4; g++) {
        int col = g * GSN + tx * 4;
        *reinterpret_cast<float4*>(Ps + row * BN + col) =
            make_float4(s[i][g * 4], s[i][g * 4 + 1], s[i][g * 4 + 2], s[i][g * 4 + 3]);
        *reinterpret_cast<float4*>(dSs + row * BN + col) =
            make_float4(dp[i][g * 4], dp[i][g * 4 + 1], dp[i][g * 4 + 2], dp[i][g * 4 + 3]);
      }
    }
#pragma unroll
    for (int j = 0; j < TSN; j++) {
      int jj = (j / 4) * GSN + tx * 4 + (j % 4);
      *reinterpret_cast<float4*>(dSt + swz(jj, ty * 4, BM)) =
          make_float4(dp[0][j], dp[1][j], dp[2][j], dp[3][j]);
    }
    __syncthreads();

    // dV += P^T dY ; dK += dS^T Q     (contraction over the BM query rows)
#pragma unroll 4
    for (int i = 0; i < BM; i++) {
      float4 p4 = *reinterpret_cast<const float4*>(Ps + i * BN + trow * 4);
      float4 s4 = *reinterpret_cast<const float4*>(dSs + i * BN + trow * 4);
      float4 y4 = *reinterpret_cast<const float4*>(dYps + i * DPAD + tcx * 4);
      float4 q4 = *reinterpret_cast<const float4*>(Qps + i * DPAD + tcx * 4);
      float pr[4] = {p4.x, p4.y, p4.z, p4.w}, sr[4] = {s4.x, s4.y, s4.z, s4.w};
      float yc[4] = {y4.x, y4.y, y4.z, y4.w}, qc[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
      for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
          dv[r][c] = fmaf(pr[r], yc[c], dv[r][c]);
          dk[r][c] = fmaf(sr[r], qc[c], dk[r][c]);
        }
    }
    // dQ tile = dS K^T            (contraction over the BN keys of this CTA)
    float dq[RI][4];
#pragma unroll
    for (int r = 0; r < RI; r++) dq[r][0] = dq[r][1] = dq[r][2] = dq[r][3] = 0.f;
#pragma unroll 4
    for (int j = 0; j < BN; j++) {
      float4 k4 = *reinterpret_cast<const float4*>(Kps + j * DPAD + tcx * 4);
      float dsr[RI];
      if (RI == 4) {
        float4 t = *reinterpret_cast<const float4*>(dSt + swz(j, trow * 4, BM));
        dsr[0] = t.x; dsr[1] = t.y; dsr[RI - 2] = t.z; dsr[RI - 1] = t.w;
      } else {
        float2 t = *reinterpret_cast<const float2*>(dSt + swz(j, trow * 2, BM));
        dsr[0] = t.x; dsr[1] = t.y;
      }
#pragma unroll
      for (int r = 0; r < RI; r++) {
        dq[r][0] = fmaf(dsr[r], k4.x, dq[r][0]);
        dq[r][1] = fmaf(dsr[r], k4.y, dq[r][1]);
        dq[r][2] = fmaf(dsr[r], k4.z, dq[r][2]);
        dq[r][3] = fmaf(dsr[r], k4.w, dq[r][3]);
      }
    }
#pragma unroll
    for (int r = 0; r < RI; r++) {
      int gi = i0 + trow * RI + r;
      if (gi >= N) continue;
#pragma unroll
      for (int c = 0; c < 4; c++) {
        int cc = tcx * 4 + c;
        if (cc < d) atomicAdd(dQp + bo + (long)gi * d + cc, dq[r][c]);
      }
    }
    __syncthreads();
  }

  const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(dKp + bo) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(dVp + bo) & 15) == 0);
#pragma unroll
  for (int r = 0; r < RJ; r++) {
    int gj = j0 + trow * RJ + r;
    if (gj >= N) continue;
    int c = tcx * 4;
    if (vec && c + 3 < d) {
      *reinterpret_cast<float4*>(dKp + bo + (long)gj * d + c) = make_float4(dk[r][0], dk[r][1], dk[r][2], dk[r][3]);
      *reinterpret_cast<float4*>(dVp + bo + (long)gj * d + c) = make_float4(dv[r][0], dv[r][1], dv[r][2], dv[r][3]);
    } else {
#pragma unroll
      for (int q = 0; q < 4; q++)
        if (c + q < d) {
          dKp[bo + (long)gj * d + c + q] = dk[r][q];
          dVp[bo + (long)gj * d + c + q] = dv[r][q];
        }
    }
  }
}

template <int DPAD, int BM, int BN>
int launch_bwd(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Qp,
               const float* Kc, const float* Kp, const float* Vc, const float* dYc,
               const float* dYp, const float* lse, const float* delta, float* dQp, float* dKp,
               float* dVp) {
  size_t smem = (size_t)(3 * DPAD * BN + 4 * DPAD * BM + 3 * BM * BN + 2 * BM) * sizeof(float);
  auto kern = attn_bwd_kernel<DPAD, BM, BN>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  dim3 grid(cdiv(N, BN), B);
  ScopedTimer tm(st, FRN_TIME_ATTN_BWD);
  kern<<<grid, kThreads, smem, st>>>(N, d, Qc, Qp, Kc, Kp, Vc, dYc, dYp, lse, delta, dQp, dKp, dVp);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // namespace

int attn_forward(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Kc,
                 const float* Vp, float* Y, float* lse) {
  if (d <= 32) return launch_fwd<32, 128, 64>(st, B, N, d, Qc, Kc, Vp, Y, lse);
  if (d <= 64) return launch_fwd<64, 128, 64>(st, B, N, d, Qc, Kc, Vp, Y, lse);
  return FRN_ERR_UNSUPPORTED;
}

}  // namespace frn

namespace frn {
int attn_backward(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Qp,
                  const float* Kc, const float* Kp, const float* Vc, const float* dYc,
                  const float* dYp, const float* lse, const float* delta, float* dQp,
                  float* dKp, float* dVp) {
  if (d <= 32)
    return launch_bwd<32, 64, 128>(st, B, N, d, Qc, Qp, Kc, Kp, Vc, dYc, dYp, lse, delta, dQp, dKp, dVp);
  if (d <= 64)
    return launch_bwd<64, 64, 64>(st, B, N, d, Qc, Qp, Kc, Kp, Vc, dYc, dYp, lse, delta, dQp, dKp, dVp);
  return FRN_ERR_UNSUPPORTED;
}
}  // namespace frn

#else
namespace frn {
int attn_forward(cudaStream_t, int, int, int, const float*, const float*, const float*, float*, float*) {
  return FRN_ERR_UNSUPPORTED;
}
int attn_backward(cudaStream_t, int, int, int, const float*, const float*, const float*, const float*, const float*,
                  const float*, const float*, const float*, const float*, float*, float*, float*) {
  return FRN_ERR_UNSUPPORTED;
}
}  // namespace frn
#endif  // FRN_DEBUG_ENGINES
