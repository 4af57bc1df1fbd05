// HBM-bound kernels of the fusion block: instance statistics and the AdaIN
// affine (retinanet/AdaIN.py:7-26), the multiplicative gate of REFusion
// (retinanet/fusion.py:26-28) and their gradients.
//
// Every kernel walks rows of N = H*W contiguous floats (one (sample, channel)
// pair per row, NCHW).  A row is owned by one warp (N < 1024) or one 256-thread
// CTA (N >= 1024); loads are 128-bit when N % 4 == 0 (rows then start 16-byte
// aligned), scalar otherwise (DDD17: N = 5655, 1452, 374, 99).  Variances are
// single-pass shifted sums merged Welford-style (see Moments below): the plain
// E[x^2]-E[x]^2 form breaks in fp32 on the near-constant maps produced by RGB
// blanking (retinanet/model.py:279-280), and a second pass over the row costs
// a second trip to HBM at the full-size levels.
#include <type_traits>
#include "common.cuh"
#include "tc05.cuh"

namespace frn {
namespace {

template <int W> struct Vec;
template <> struct Vec<1> {
  float v[1];
  __device__ __forceinline__ void load(const float* p) { v[0] = __ldg(p); }
  __device__ __forceinline__ void load_rw(const float* p) { v[0] = *p; }
  __device__ __forceinline__ void store(float* p) const { *p = v[0]; }
};
template <> struct Vec<4> {
  float v[4];
  __device__ __forceinline__ void load(const float* p) {
    float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ __forceinline__ void load_rw(const float* p) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ __forceinline__ void store(float* p) const {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};

template <int GS>
__device__ __forceinline__ float group_sum(float v, float* red) {
  if (GS == 32) return warp_sum(v);
  return block_sum(v, red);
}

// Row owned by this thread group, or -1.  GS == 32: 8 rows per 256-thread CTA.
template <int GS>
__device__ __forceinline__ long my_row(long rows, int& lane) {
  if (GS == 32) {
    lane = threadIdx.x & 31;
    long r = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    return r < rows ? r : -1;
  }
  lane = threadIdx.x;
  return blockIdx.x;
}

#define ROW_KERNEL_PROLOGUE()                 \
  __shared__ float red[33];                   \
  (void)red;                                  \
  int lane;                                   \
  long row = my_row<GS>(rows, lane);          \
  if (row < 0) return;

// ---- single-pass statistics: shifted sums per thread, Chan / Welford merge across threads ---------------------
// One pass over HBM instead of two (the second pass of a 2048-row x 154 KB working set does not stay in L2:
// ncu showed 592 MB read for 315 MB of input in gate_stats at DSEC level 1).  Each thread accumulates
// s1 = sum (x - K), s2 = sum (x - K)^2 with K = the first element it sees, so near-constant maps (|mean| = 1e3 sigma,
// the blanked-RGB case of retinanet/model.py:279-280) lose nothing to cancellation: x - K is exact there and
// M2 = s2 - s1^2 / n cancels at most the spread of the thread's own 75-odd samples.  The per-thread
// (n, mean, M2) triples are then merged with the exact pairwise formula
//   M2 = M2a + M2b + (mean_b - mean_a)^2 na nb / (na + nb)
// by warp shuffles and, for CTA-wide rows, once more through shared memory (BASELINE.json north_star:
// "warp-shuffle plus shared-memory Welford").
struct Moments {
  float n, mean, m2;
  __device__ __forceinline__ void merge(float nb, float mb, float m2b) {
    const float nt = n + nb;
    if (nt > 0.f) {
      const float dlt = mb - mean, w = nb / nt;
      mean = fmaf(dlt, w, mean);
      m2 = m2 + m2b + dlt * dlt * n * w;
    }
    n = nt;
  }
};
struct ShiftedSum {
  float k, s1, s2, n;
  __device__ __forceinline__ void init() { k = 0.f; s1 = 0.f; s2 = 0.f; n = 0.f; }
  __device__ __forceinline__ void add(float x) {
    if (n == 0.f) k = x;
    const float dlt = x - k;
    s1 += dlt; s2 = fmaf(dlt, dlt, s2); n += 1.f;
  }
  __device__ __forceinline__ Moments finish() const {
    Moments m;
    m.n = n;
    m.mean = n > 0.f ? k + s1 / n : 0.f;
    m.m2 = n > 0.f ? fmaxf(s2 - s1 * s1 / n, 0.f) : 0.f;
    return m;
  }
};
__device__ __forceinline__ Moments warp_merge(Moments m) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float nb = __shfl_xor_sync(0xffffffffu, m.n, o), mb = __shfl_xor_sync(0xffffffffu, m.mean, o),
                qb = __shfl_xor_sync(0xffffffffu, m.m2, o);
    // both lanes of a pair must arrive at the same triple: merge in a fixed (lower lane first) order
    Moments a = m;
    if ((threadIdx.x & o) == 0) { a.merge(nb, mb, qb); }
    else { a.n = nb; a.mean = mb; a.m2 = qb; a.merge(m.n, m.mean, m.m2); }
    m = a;
  }
  return m;
}
// Merge over the thread group that owns the row (GS = 32: one warp, GS = 256 / 1024: the CTA).  `sm` holds 3 * 32 floats.
template <int GS>
__device__ __forceinline__ Moments group_merge(Moments m, float* sm) {
  m = warp_merge(m);
  if (GS == 32) return m;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) { sm[w] = m.n; sm[32 + w] = m.mean; sm[64 + w] = m.m2; }
  __syncthreads();
  Moments t;
  t.n = sm[0]; t.mean = sm[32]; t.m2 = sm[64];
#pragma unroll
  for (int i = 1; i < GS / 32; i++) t.merge(sm[i], sm[32 + i], sm[64 + i]);
  return t;
}

// ---- calc_mean_std -----------------------------------------------------------
template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) row_mean_std_kernel(long rows, int N, const float* __restrict__ x,
                                                           float eps, float* __restrict__ mean,
                                                           float* __restrict__ sd) {
  __shared__ float sm[96];
  int lane;
  long row = my_row<GS>(rows, lane);
  if (row < 0) return;
  const float* xr = x + row * N;
  ShiftedSum a; a.init();
  constexpr int U = 4;
  for (int n0 = lane * W; n0 < N; n0 += U * GS * W) {
    Vec<W> v[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (n0 + u * GS * W < N) v[u].load(xr + n0 + u * GS * W);
#pragma unroll
    for (int u = 0; u < U; u++)
      if (n0 + u * GS * W < N) {
#pragma unroll
        for (int q = 0; q < W; q++) a.add(v[u].v[q]);
      }
  }
  const Moments m = group_merge<GS>(a.finish(), sm);
  if (lane == 0) {
    mean[row] = m.mean;
    sd[row] = sqrtf(m.m2 / (float)(N - 1) + eps);
  }
}

// ---- gate + statistics of both gated maps ------------------------------------
template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) gate_stats_kernel(long rows, int N, const float* __restrict__ r0,
                                                         const float* __restrict__ e0,
                                                         float* __restrict__ r1, float* __restrict__ e1,
                                                         float eps, float* __restrict__ mu_r,
                                                         float* __restrict__ sd_r,
                                                         float* __restrict__ mu_e,
                                                         float* __restrict__ sd_e) {
  __shared__ float sm[96];
  int lane;
  long row = my_row<GS>(rows, lane);
  if (row < 0) return;
  const long o = row * N;
  ShiftedSum ar, ae; ar.init(); ae.init();
  // U chunks per thread and trip, all loads issued before the first use: 2 U independent 16-byte loads in flight per
  // thread (with one chunk per trip the kernel ran at 0.65 of the HBM peak, latency- not bandwidth-bound)
  constexpr int U = 4;
  for (int n0 = lane * W; n0 < N; n0 += U * GS * W) {
    Vec<W> a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int n = n0 + u * GS * W;
      if (n < N) { a[u].load(r0 + o + n); b[u].load(e0 + o + n); }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int n = n0 + u * GS * W;
      if (n < N) {
        Vec<W> x, y;
#pragma unroll
        for (int q = 0; q < W; q++) {
          float m = a[u].v[q] * b[u].v[q];
          x.v[q] = a[u].v[q] + m; y.v[q] = b[u].v[q] + m;
          ar.add(x.v[q]); ae.add(y.v[q]);
        }
        x.store(r1 + o + n); y.store(e1 + o + n);
      }
    }
  }
  const Moments mr = group_merge<GS>(ar.finish(), sm);
  const Moments me = group_merge<GS>(ae.finish(), sm);
  if (lane == 0) {
    mu_r[row] = mr.mean; sd_r[row] = sqrtf(mr.m2 / (float)(N - 1) + eps);
    mu_e[row] = me.mean; sd_e[row] = sqrtf(me.m2 / (float)(N - 1) + eps);
  }
}

// Long rows (N >= 4096, 16-byte aligned): the two input rows are streamed through shared memory by bulk async copies
// (cp.async.bulk, completion on an mbarrier), three 8 KB chunks per map in flight per CTA.  The register-staged
// kernel above keeps 2 x 16 bytes per thread in flight: 6 CTAs x 256 threads = 48 KB per SM against the ~66 KB that
// 44 GB/s per SM x ~1.5 us of loaded latency need, which is exactly the 0.66 - 0.75 of the HBM peak it reached
// however many bytes it moved (143 us at DSEC level 1 for both the two-pass and the one-pass version; more loads per
// thread cost the occupancy they bought; one CTA per (row, 2 K-element segment) with a last-arriver merge was
// slower still: 18 K CTAs of two block-wide merges each).  With bulk copies the bytes in flight do not occupy
// registers: 4 CTAs x 3 stages x 16 KB = 192 KB per SM.
constexpr int kGsChunk = 2048;      // floats per map and stage
constexpr int kGsStages = 3;
__global__ void __launch_bounds__(256) gate_stats_bulk_kernel(int N, const float* __restrict__ r0,
                                                              const float* __restrict__ e0, float* __restrict__ r1,
                                                              float* __restrict__ e1, float eps, float* __restrict__ mu_r,
                                                              float* __restrict__ sd_r, float* __restrict__ mu_e,
                                                              float* __restrict__ sd_e) {
  using namespace tc;
  extern __shared__ __align__(128) uint8_t gs_smem[];
  float (*buf)[2][kGsChunk] = reinterpret_cast<float (*)[2][kGsChunk]>(gs_smem);
  uint64_t* full = reinterpret_cast<uint64_t*>(gs_smem + sizeof(float) * kGsStages * 2 * kGsChunk);
  __shared__ float sm[96];
  const long row = blockIdx.x, o = row * N;
  const int tid = threadIdx.x;
  const int nchunks = (N + kGsChunk - 1) / kGsChunk;
  if (tid == 0) {
    for (int s = 0; s < kGsStages; s++) mbar_init(&full[s], 1);
    mbar_init_fence();
  }
  __syncthreads();
  auto issue = [&](int c) {
    const int s = c % kGsStages, n0 = c * kGsChunk;
    const uint32_t bytes = (uint32_t)(min(kGsChunk, N - n0) * 4);
    mbar_expect_tx(&full[s], 2 * bytes);
    bulk_g2s(&buf[s][0][0], r0 + o + n0, bytes, &full[s]);
    bulk_g2s(&buf[s][1][0], e0 + o + n0, bytes, &full[s]);
  };
  if (tid == 0)
    for (int c = 0; c < kGsStages && c < nchunks; c++) issue(c);
  ShiftedSum ar, ae; ar.init(); ae.init();
  for (int c = 0; c < nchunks; c++) {
    const int s = c % kGsStages, n0 = c * kGsChunk, len = min(kGsChunk, N - n0);
    mbar_wait(&full[s], (uint32_t)((c / kGsStages) & 1));
#pragma unroll
    for (int u = 0; u < kGsChunk / 1024; u++) {
      const int i = (tid + u * 256) * 4;
      if (i < len) {
        const float4 a = *reinterpret_cast<const float4*>(&buf[s][0][i]);
        const float4 b = *reinterpret_cast<const float4*>(&buf[s][1][i]);
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
        float x[4], y[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const float m = av[q] * bv[q];
          x[q] = av[q] + m; y[q] = bv[q] + m;
          ar.add(x[q]); ae.add(y[q]);
        }
        *reinterpret_cast<float4*>(r1 + o + n0 + i) = make_float4(x[0], x[1], x[2], x[3]);
        *reinterpret_cast<float4*>(e1 + o + n0 + i) = make_float4(y[0], y[1], y[2], y[3]);
      }
    }
    fence_proxy_async_smem();        // generic reads of this stage are ordered before the bulk copy that refills it
    __syncthreads();
    if (tid == 0 && c + kGsStages < nchunks) issue(c + kGsStages);
  }
  const Moments mr = group_merge<256>(ar.finish(), sm);
  const Moments me = group_merge<256>(ae.finish(), sm);
  if (tid == 0) {
    mu_r[row] = mr.mean; sd_r[row] = sqrtf(mr.m2 / (float)(N - 1) + eps);
    mu_e[row] = me.mean; sd_e[row] = sqrtf(me.m2 / (float)(N - 1) + eps);
  }
}

// ---- AdaIN apply --------------------------------------------------------------
template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) adain_apply_kernel(long rows, int C, int N, const float* __restrict__ x,
                                                          const float* __restrict__ mu_x,
                                                          const float* __restrict__ sd_x,
                                                          const float* __restrict__ mu_t,
                                                          const float* __restrict__ sd_t,
                                                          float* __restrict__ out, long ob) {
  ROW_KERNEL_PROLOGUE();
  const float* xr = x + row * N;
  float* orow = out + (row / C) * ob + (row % C) * (long)N;
  float mx = mu_x[row], a = sd_t[row] / sd_x[row], mt = mu_t[row];
  constexpr int U = 4;
  for (int n0 = lane * W; n0 < N; n0 += U * GS * W) {
    Vec<W> v[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (n0 + u * GS * W < N) v[u].load(xr + n0 + u * GS * W);
#pragma unroll
    for (int u = 0; u < U; u++)
      if (n0 + u * GS * W < N) {
#pragma unroll
        for (int q = 0; q < W; q++) v[u].v[q] = fmaf(v[u].v[q] - mx, a, mt);
        v[u].store(orow + n0 + u * GS * W);
      }
  }
}

// bf16 output (round to nearest even); W = 4: one 8-byte store
__device__ __forceinline__ uint32_t bf16_rn(float f) {
  const uint32_t u = __float_as_uint(f);
  if ((u & 0x7f800000u) == 0x7f800000u) return (u >> 16) | ((u & 0xffffu) ? 0x40u : 0u);   // inf / nan
  return (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
}
template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) adain_apply_bf16_kernel(long rows, int C, int N, const float* __restrict__ x,
                                                               const float* __restrict__ mu_x,
                                                               const float* __restrict__ sd_x,
                                                               const float* __restrict__ mu_t,
                                                               const float* __restrict__ sd_t,
                                                               uint16_t* __restrict__ out, long ob) {
  ROW_KERNEL_PROLOGUE();
  const float* xr = x + row * N;
  uint16_t* orow = out + (row / C) * ob + (row % C) * (long)N;
  float mx = mu_x[row], a = sd_t[row] / sd_x[row], mt = mu_t[row];
  for (int n = lane * W; n < N; n += GS * W) {
    Vec<W> v; v.load(xr + n);
    uint32_t h[W];
#pragma unroll
    for (int q = 0; q < W; q++) h[q] = bf16_rn(fmaf(v.v[q] - mx, a, mt));
    if (W == 4) *reinterpret_cast<uint2*>(orow + n) = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3 % W] << 16));
    else orow[n] = (uint16_t)h[0];
  }
}

// ---- AdaIN backward -------------------------------------------------------------
template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) adain_bwd_reduce_kernel(long rows, int C, int N,
                                                               const float* __restrict__ g, long gb,
                                                               const float* __restrict__ x,
                                                               const float* __restrict__ mu_x,
                                                               const float* __restrict__ sd_x,
                                                               float* __restrict__ g1,
                                                               float* __restrict__ g2) {
  ROW_KERNEL_PROLOGUE();
  const float* xr = x + row * N;
  const float* gr = g + (row / C) * gb + (row % C) * (long)N;
  float mx = mu_x[row];
  float s1 = 0.f, s2 = 0.f;
  constexpr int U = 4;
  for (int n0 = lane * W; n0 < N; n0 += U * GS * W) {
    Vec<W> v[U], gg[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (n0 + u * GS * W < N) { v[u].load(xr + n0 + u * GS * W); gg[u].load(gr + n0 + u * GS * W); }
#pragma unroll
    for (int u = 0; u < U; u++)
      if (n0 + u * GS * W < N) {
#pragma unroll
        for (int q = 0; q < W; q++) { s1 += gg[u].v[q]; s2 = fmaf(gg[u].v[q], v[u].v[q] - mx, s2); }
      }
  }
  s1 = group_sum<GS>(s1, red);
  s2 = group_sum<GS>(s2, red);
  if (lane == 0) { g1[row] = s1; g2[row] = s2 / sd_x[row]; }
}

template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) adain_bwd_content_kernel(
    long rows, int C, int N, const float* __restrict__ g, long gb, const float* __restrict__ x,
    const float* __restrict__ mu_x, const float* __restrict__ sd_x, const float* __restrict__ sd_t,
    const float* __restrict__ g1, const float* __restrict__ g2, float* __restrict__ dx, int accumulate) {
  ROW_KERNEL_PROLOGUE();
  const float* xr = x + row * N;
  const float* gr = g + (row / C) * gb + (row % C) * (long)N;
  float* dr = dx + row * N;
  float mx = mu_x[row], isx = 1.f / sd_x[row];
  float k = sd_t[row] * isx;
  float c1 = g1[row] / (float)N, c2 = g2[row] * isx / (float)(N - 1);
  constexpr int U = 2;
  for (int n0 = lane * W; n0 < N; n0 += U * GS * W) {
    Vec<W> v[U], gg[U], o[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int n = n0 + u * GS * W;
      if (n < N) {
        v[u].load(xr + n); gg[u].load(gr + n);
        if (accumulate) o[u].load_rw(dr + n);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int n = n0 + u * GS * W;
      if (n < N) {
#pragma unroll
        for (int q = 0; q < W; q++) {
          float r = k * (gg[u].v[q] - c1 - (v[u].v[q] - mx) * c2);
          o[u].v[q] = accumulate ? o[u].v[q] + r : r;
        }
        o[u].store(dr + n);
      }
    }
  }
}

template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) adain_bwd_style_kernel(long rows, int N, const float* __restrict__ t,
                                                              const float* __restrict__ mu_t,
                                                              const float* __restrict__ sd_t,
                                                              const float* __restrict__ g1,
                                                              const float* __restrict__ g2,
                                                              float* __restrict__ dstyle) {
  ROW_KERNEL_PROLOGUE();
  const float* tr = t + row * N;
  float* dr = dstyle + row * N;
  float mt = mu_t[row];
  float al = g1[row] / (float)N, be = g2[row] / ((float)(N - 1) * sd_t[row]);
  for (int n = lane * W; n < N; n += GS * W) {
    Vec<W> v; v.load(tr + n);
#pragma unroll
    for (int q = 0; q < W; q++) v.v[q] = fmaf(be, v.v[q] - mt, al);
    v.store(dr + n);
  }
}

template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) center_rows_kernel(long rows, int N, const float* __restrict__ y,
                                                          const float* __restrict__ mean,
                                                          float* __restrict__ yc) {
  ROW_KERNEL_PROLOGUE();
  float mu = mean[row];
  for (int n = lane * W; n < N; n += GS * W) {
    Vec<W> v; v.load(y + row * N + n);
#pragma unroll
    for (int q = 0; q < W; q++) v.v[q] -= mu;
    v.store(yc + row * N + n);
  }
}

template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) add_row_bias_kernel(long rows, int N, const float* __restrict__ u,
                                                           float* __restrict__ x) {
  ROW_KERNEL_PROLOGUE();
  float b = u[row];
  for (int n = lane * W; n < N; n += GS * W) {
    Vec<W> v; v.load_rw(x + row * N + n);
#pragma unroll
    for (int q = 0; q < W; q++) v.v[q] += b;
    v.store(x + row * N + n);
  }
}

template <int GS, int W>
__global__ void __launch_bounds__(GS < 256 ? 256 : GS) row_dot_kernel(long rows, int N, const float* __restrict__ x,
                                                      const float* __restrict__ y, float* __restrict__ out) {
  ROW_KERNEL_PROLOGUE();
  float s = 0.f;
  for (int n = lane * W; n < N; n += GS * W) {
    Vec<W> a, b; a.load(x + row * N + n); b.load(y + row * N + n);
#pragma unroll
    for (int q = 0; q < W; q++) s = fmaf(a.v[q], b.v[q], s);
  }
  s = group_sum<GS>(s, red);
  if (lane == 0) out[row] = s;
}

// ---- bf16-inference fast path (frn_refusion_infer, FRN_BF16): padded rows, tf32-exact GEMM operands ------------------
// Maps live in buffers whose rows are padded to ld = round_up(N, 4) floats, so that every row is 16-byte aligned and
// TMA can describe them whatever N is (DDD17: N = 5655, 1452, 374, 99), and every value a GEMM will read is rounded
// to tf32 by the kernel that WRITES it: the v2 GEMM then feeds the raw tiles to the tensor core with no staging
// pass at all (kind::tf32 would truncate unrounded operands).
__device__ __forceinline__ float rn_tf32(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}
// bf16 (rows, N) -> fp32 (rows, ld), exact (bf16 values are tf32 values); pad columns zero.  One CTA per row
// segment of 2048 outputs: four consecutive outputs per thread and trip, one 16-byte store each (rows of ld floats
// are 16-byte aligned); the bf16 source rows are only 2-byte aligned, so they are read element-wise (coalesced).
__global__ void __launch_bounds__(256) cast_pad_bf16_kernel(long rows, int N, int ld, const uint16_t* __restrict__ src,
                                                            float* __restrict__ dst) {
  const long row = blockIdx.x;
  const uint16_t* s = src + row * N;
  float* d = dst + row * ld;
  for (int n = (blockIdx.y * 256 + threadIdx.x) * 4; n < ld; n += gridDim.y * 1024) {
    float4 v;
    v.x = n < N ? __uint_as_float((uint32_t)__ldg(s + n) << 16) : 0.f;
    v.y = n + 1 < N ? __uint_as_float((uint32_t)__ldg(s + n + 1) << 16) : 0.f;
    v.z = n + 2 < N ? __uint_as_float((uint32_t)__ldg(s + n + 2) << 16) : 0.f;
    v.w = n + 3 < N ? __uint_as_float((uint32_t)__ldg(s + n + 3) << 16) : 0.f;
    *reinterpret_cast<float4*>(d + n) = v;
  }
}
__global__ void round_tf32_kernel(long n, const float* __restrict__ src, float* __restrict__ dst) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = rn_tf32(src[i]);
}
// gate (fusion.py:26-28) + statistics of both gated maps on padded rows; r1 / e1 are stored rounded to tf32 (they are
// GEMM operands and the AdaIN content), the statistics come from the unrounded values
template <int GS>
__global__ void __launch_bounds__(256) gate_stats_pad_kernel(long rows, int N, int ld, const float* __restrict__ r0,
                                                             const float* __restrict__ e0, float* __restrict__ r1,
                                                             float* __restrict__ e1, float eps, float* __restrict__ mu_r,
                                                             float* __restrict__ sd_r, float* __restrict__ mu_e,
                                                             float* __restrict__ sd_e) {
  __shared__ float sm[96];
  int lane;
  long row = my_row<GS>(rows, lane);
  if (row < 0) return;
  const long o = row * ld;
  ShiftedSum ar, ae; ar.init(); ae.init();
  constexpr int U = 3;
  for (int n0 = lane * 4; n0 < N; n0 += U * GS * 4) {
    Vec<4> a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int n = n0 + u * GS * 4;
      if (n < N) { a[u].load(r0 + o + n); b[u].load(e0 + o + n); }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int n = n0 + u * GS * 4;
      if (n < N) {
        Vec<4> x, y;
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const bool ok = n + q < N;
          const float m = a[u].v[q] * b[u].v[q];
          const float xv = a[u].v[q] + m, yv = b[u].v[q] + m;
          if (ok) { ar.add(xv); ae.add(yv); }
          x.v[q] = ok ? rn_tf32(xv) : 0.f; y.v[q] = ok ? rn_tf32(yv) : 0.f;
        }
        x.store(r1 + o + n); y.store(e1 + o + n);
      }
    }
  }
  const Moments mr = group_merge<GS>(ar.finish(), sm);
  const Moments me = group_merge<GS>(ae.finish(), sm);
  if (lane == 0) {
    mu_r[row] = mr.mean; sd_r[row] = sqrtf(mr.m2 / (float)(N - 1) + eps);
    mu_e[row] = me.mean; sd_e[row] = sqrtf(me.m2 / (float)(N - 1) + eps);
  }
}
// yc (rows, ld) = round_tf32(y (rows, N) - mean), pad columns zero
template <int GS>
__global__ void __launch_bounds__(256) center_rows_pad_kernel(long rows, int N, int ld, const float* __restrict__ y,
                                                              const float* __restrict__ mean, float* __restrict__ yc) {
  int lane;
  long row = my_row<GS>(rows, lane);
  if (row < 0) return;
  const float mu = mean[row];
  for (int n = lane; n < ld; n += GS) yc[row * ld + n] = n < N ? rn_tf32(__ldg(y + row * N + n) - mu) : 0.f;
}
// AdaIN affine of padded content rows -> bf16 output rows of N elements (batch stride ob)
template <int GS>
__global__ void __launch_bounds__(256) adain_apply_bf16_pad_kernel(long rows, int C, int N, int ld,
                                                                   const float* __restrict__ x,
                                                                   const float* __restrict__ mu_x,
                                                                   const float* __restrict__ sd_x,
                                                                   const float* __restrict__ mu_t,
                                                                   const float* __restrict__ sd_t,
                                                                   uint16_t* __restrict__ out, long ob) {
  int lane;
  long row = my_row<GS>(rows, lane);
  if (row < 0) return;
  const float* xr = x + row * ld;
  uint16_t* orow = out + (row / C) * ob + (row % C) * (long)N;
  const float mx = mu_x[row], a = sd_t[row] / sd_x[row], mt = mu_t[row];
  for (int n = lane; n < N; n += GS) orow[n] = (uint16_t)bf16_rn(fmaf(__ldg(xr + n) - mx, a, mt));
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Dispatch on row length (warp- or CTA-per-row) and vector width.
// Dispatch on row length (warp- or CTA-per-row) and vector width.  (One 1024-thread CTA per long row was tried to
// get rid of the ragged last wave -- 2048 rows are 2.3 waves of 6 CTAs per SM -- and lost: with one resident CTA per
// SM nothing overlaps a CTA's load -> reduce -> exit phases; gate_stats went from 143 to 231 us.)
#define LAUNCH_ROWS(KERNEL, passes, rows, N, vec, ...)                                    \
  do {                                                                                    \
    ScopedTimer tm__(st, FRN_TIME_ROWS);                                                  \
    KernelTimer kt__(st, #KERNEL, 4.0 * (passes) * (double)(rows) * (N));                 \
    if ((N) >= 1024) {                                                                    \
      if (vec) KERNEL<256, 4><<<(unsigned)(rows), 256, 0, st>>>(__VA_ARGS__);             \
      else KERNEL<256, 1><<<(unsigned)(rows), 256, 0, st>>>(__VA_ARGS__);                 \
    } else {                                                                              \
      unsigned nb = (unsigned)(((rows) + 7) / 8);                                         \
      if (vec) KERNEL<32, 4><<<nb, 256, 0, st>>>(__VA_ARGS__);                            \
      else KERNEL<32, 1><<<nb, 256, 0, st>>>(__VA_ARGS__);                                \
    }                                                                                     \
    FRN_LAUNCH_CHECK();                                                                   \
  } while (0)

}  // namespace

int row_mean_std(cudaStream_t st, long rows, int N, const float* x, float eps, float* mean, float* sd) {
  bool vec = (N % 4 == 0) && aligned16(x);
  LAUNCH_ROWS(row_mean_std_kernel, 1, rows, N, vec, rows, N, x, eps, mean, sd);
  return FRN_OK;
}

size_t gate_stats_scratch_bytes(long rows, int N) { (void)rows; (void)N; return 0; }
int gate_stats(cudaStream_t st, long rows, int N, const float* r0, const float* e0, float* r1, float* e1,
               float eps, float* mu_r, float* sd_r, float* mu_e, float* sd_e, void* scratch) {
  (void)scratch;
  bool vec = (N % 4 == 0) && aligned16(r0) && aligned16(e0) && aligned16(r1) && aligned16(e1);
  if (vec && N >= 4096 && rows <= 2147483647L) {
    ScopedTimer tm(st, FRN_TIME_ROWS);
    KernelTimer kt(st, "gate_stats_bulk_kernel", 16.0 * (double)rows * N);
    constexpr int smem = (int)sizeof(float) * kGsStages * 2 * kGsChunk + 64;
    FRN_TRY(set_smem_once((const void*)gate_stats_bulk_kernel, smem));
    gate_stats_bulk_kernel<<<(unsigned)rows, 256, smem, st>>>(N, r0, e0, r1, e1, eps, mu_r, sd_r, mu_e, sd_e);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  LAUNCH_ROWS(gate_stats_kernel, 4, rows, N, vec, rows, N, r0, e0, r1, e1, eps, mu_r, sd_r, mu_e, sd_e);
  return FRN_OK;
}

int adain_apply(cudaStream_t st, int B, int C, int N, const float* x, const float* mu_x, const float* sd_x,
                const float* mu_t, const float* sd_t, float* out, long ob) {
  long rows = (long)B * C;
  bool vec = (N % 4 == 0) && (ob % 4 == 0) && aligned16(x) && aligned16(out);
  LAUNCH_ROWS(adain_apply_kernel, 2, rows, N, vec, rows, C, N, x, mu_x, sd_x, mu_t, sd_t, out, ob);
  return FRN_OK;
}

int adain_apply_bf16(cudaStream_t st, int B, int C, int N, const float* x, const float* mu_x, const float* sd_x,
                     const float* mu_t, const float* sd_t, uint16_t* out, long ob) {
  long rows = (long)B * C;
  bool vec = (N % 4 == 0) && (ob % 4 == 0) && aligned16(x) && ((reinterpret_cast<uintptr_t>(out) & 7) == 0);
  LAUNCH_ROWS(adain_apply_bf16_kernel, 1.5, rows, N, vec, rows, C, N, x, mu_x, sd_x, mu_t, sd_t, out, ob);
  return FRN_OK;
}

int adain_bwd_reduce(cudaStream_t st, int B, int C, int N, const float* g, long gb, const float* x,
                     const float* mu_x, const float* sd_x, float* g1, float* g2) {
  long rows = (long)B * C;
  bool vec = (N % 4 == 0) && (gb % 4 == 0) && aligned16(x) && aligned16(g);
  LAUNCH_ROWS(adain_bwd_reduce_kernel, 2, rows, N, vec, rows, C, N, g, gb, x, mu_x, sd_x, g1, g2);
  return FRN_OK;
}

int adain_bwd_content(cudaStream_t st, int B, int C, int N, const float* g, long gb, const float* x,
                      const float* mu_x, const float* sd_x, const float* sd_t, const float* g1,
                      const float* g2, float* dx, int accumulate) {
  long rows = (long)B * C;
  bool vec = (N % 4 == 0) && (gb % 4 == 0) && aligned16(x) && aligned16(g) && aligned16(dx);
  LAUNCH_ROWS(adain_bwd_content_kernel, (accumulate ? 4 : 3), rows, N, vec, rows, C, N, g, gb, x, mu_x, sd_x, sd_t, g1, g2, dx,
              accumulate);
  return FRN_OK;
}

int adain_bwd_style(cudaStream_t st, int B, int C, int N, const float* t, const float* mu_t,
                    const float* sd_t, const float* g1, const float* g2, float* dstyle) {
  long rows = (long)B * C;
  bool vec = (N % 4 == 0) && aligned16(t) && aligned16(dstyle);
  LAUNCH_ROWS(adain_bwd_style_kernel, 2, rows, N, vec, rows, N, t, mu_t, sd_t, g1, g2, dstyle);
  return FRN_OK;
}

#define LAUNCH_PAD(KERNEL, name, bytes, rows, N, ...)                                     \
  do {                                                                                    \
    ScopedTimer tm__(st, FRN_TIME_ROWS);                                                  \
    KernelTimer kt__(st, name, bytes);                                                    \
    if ((N) >= 1024) KERNEL<256><<<(unsigned)(rows), 256, 0, st>>>(__VA_ARGS__);          \
    else KERNEL<32><<<(unsigned)(((rows) + 7) / 8), 256, 0, st>>>(__VA_ARGS__);           \
    FRN_LAUNCH_CHECK();                                                                   \
  } while (0)

int cast_pad_bf16(cudaStream_t st, long rows, int N, int ld, const uint16_t* src, float* dst) {
  ScopedTimer tm(st, FRN_TIME_ROWS);
  int gy = (ld + 1023) / 1024;
  if (gy > 4) gy = 4;
  cast_pad_bf16_kernel<<<dim3((unsigned)rows, (unsigned)gy), 256, 0, st>>>(rows, N, ld, src, dst);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
int round_tf32_copy(cudaStream_t st, long n, const float* src, float* dst) {
  long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  round_tf32_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, src, dst);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
int gate_stats_pad(cudaStream_t st, long rows, int N, int ld, const float* r0, const float* e0, float* r1, float* e1,
                   float eps, float* mu_r, float* sd_r, float* mu_e, float* sd_e) {
  LAUNCH_PAD(gate_stats_pad_kernel, "gate_stats_pad_kernel", 16.0 * rows * N, rows, N, rows, N, ld, r0, e0, r1, e1, eps,
             mu_r, sd_r, mu_e, sd_e);
  return FRN_OK;
}
int center_rows_pad(cudaStream_t st, long rows, int N, int ld, const float* y, const float* mean, float* yc) {
  LAUNCH_PAD(center_rows_pad_kernel, "center_rows_pad_kernel", 8.0 * rows * N, rows, N, rows, N, ld, y, mean, yc);
  return FRN_OK;
}
int adain_apply_bf16_pad(cudaStream_t st, int B, int C, int N, int ld, const float* x, const float* mu_x,
                         const float* sd_x, const float* mu_t, const float* sd_t, uint16_t* out, long ob) {
  const long rows = (long)B * C;
  LAUNCH_PAD(adain_apply_bf16_pad_kernel, "adain_apply_bf16_pad_kernel", 6.0 * rows * N, rows, N, rows, C, N, ld, x, mu_x,
             sd_x, mu_t, sd_t, out, ob);
  return FRN_OK;
}

int center_rows(cudaStream_t st, long rows, int N, const float* y, const float* mean, float* yc) {
  bool vec = (N % 4 == 0) && aligned16(y) && aligned16(yc);
  LAUNCH_ROWS(center_rows_kernel, 2, rows, N, vec, rows, N, y, mean, yc);
  return FRN_OK;
}

int add_row_bias(cudaStream_t st, long rows, int N, const float* u, float* x) {
  bool vec = (N % 4 == 0) && aligned16(x);
  LAUNCH_ROWS(add_row_bias_kernel, 2, rows, N, vec, rows, N, u, x);
  return FRN_OK;
}

int row_dot(cudaStream_t st, long rows, int N, const float* x, const float* y, float* out) {
  bool vec = (N % 4 == 0) && aligned16(x) && aligned16(y);
  LAUNCH_ROWS(row_dot_kernel, 2, rows, N, vec, rows, N, x, y, out);
  return FRN_OK;
}

}  // namespace frn
