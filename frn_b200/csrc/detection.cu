// the two host-bound hops between the heads and the optimiser step.
//
//   anchors    retinanet/anchors.py:25-45 (Anchors.forward), :47-82 (generate_anchors), :120-147 (shift):
//              NumPy regenerates ~230 k boxes in float64 and copies them host -> device on EVERY step.
//              Here: the (ratios x scales) base boxes are formed on the host in double (a few dozen
//              numbers), one kernel per pyramid level adds the cell shifts in double and rounds once to
//              fp32 -- bit-identical to the NumPy result, written straight into device memory.
//   focal loss retinanet/losses.py:28-169 (FocalLoss.forward) with calc_iou (:5-24): a Python loop over
//              the batch, ~60 ATen launches and three host synchronisations per sample (boolean-mask
//              indexing, `if positive_indices.sum() > 0`).  Here: ONE forward kernel (IoU assignment,
//              classification and box terms, per-CTA partial sums in double) + one finalising CTA, and
//              ONE backward kernel; no host synchronisation, no atomics (bit-reproducible).
//
// Both are HBM-bound streaming passes: forward reads (K + 4) floats per (sample, anchor) and writes one
// int32 of assignment state; backward reads the same and writes (K + 4) floats.
#include <math.h>
#include "common.cuh"

namespace frn {
namespace {

constexpr int kThreads = 256;
constexpr int kAnch = 1024;            // anchors per CTA of the focal-loss kernels (4 per thread)
constexpr int kMaxAnn = 1024;          // annotation rows per sample staged in shared memory (6 floats each)
constexpr int kMaxBase = 32;           // ratios x scales per cell

struct BaseBoxes { double v[kMaxBase][4]; };

// out[((y * fw + x) * nA + a) * 4 + j] = float(base[a][j] + shift_j),  shift = ((x|y) + 0.5) * stride
// (anchors.py:120-147: cells row-major, the nA boxes of a cell adjacent).
__global__ void __launch_bounds__(kThreads) anchors_level_kernel(int fh, int fw, int nA, int stride, BaseBoxes base,
                                                                 float4* __restrict__ out) {
  long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
  long total = (long)fh * fw * nA;
  if (i >= total) return;
  int a = (int)(i % nA);
  long cell = i / nA;
  int x = (int)(cell % fw), y = (int)(cell / fw);
  double sx = __dmul_rn((double)x + 0.5, (double)stride);
  double sy = __dmul_rn((double)y + 0.5, (double)stride);
  float4 o;
  o.x = (float)__dadd_rn(base.v[a][0], sx);
  o.y = (float)__dadd_rn(base.v[a][1], sy);
  o.z = (float)__dadd_rn(base.v[a][2], sx);
  o.w = (float)__dadd_rn(base.v[a][3], sy);
  out[i] = o;
}

// ---- focal loss ------------------------------------------------------------------------------------
// Assignment state of one (sample, anchor): >= 0 positive, matched to that annotation row (IoU >= 0.5);
// kNegative: IoU_max < 0.4 (all targets 0); kIgnore: 0.4 <= IoU_max < 0.5 (no classification term).
constexpr int kNegative = -1, kIgnore = -2;

// calc_iou (losses.py:5-24) in the reference's own operation order, every step rounded to fp32 separately
// (no FMA contraction) so that the 0.4 / 0.5 thresholds fall exactly where the reference's fp32 ATen ops
// put them (assign_anchors below).
// Assignment of kAnch / kThreads anchors at once against the sample's annotations in shared memory
// (ann: M x 5, area: M precomputed box areas).  Pairs that do not overlap have IoU = 0 / ua = 0 exactly and
// skip the division.
template <int U>
__device__ __forceinline__ void assign_anchors(const float4 (&a)[U], const float* __restrict__ ann,
                                               const float* __restrict__ area, int M, int (&state)[U]) {
  float area_a[U], best[U];
  int arg[U];
#pragma unroll
  for (int u = 0; u < U; u++) {
    area_a[u] = __fmul_rn(__fsub_rn(a[u].z, a[u].x), __fsub_rn(a[u].w, a[u].y));
    best[u] = -1.f; arg[u] = -1;
  }
  for (int j = 0; j < M; j++) {
    if (ann[j * 5 + 4] == -1.f) continue;                          // padding rows (losses.py:45); CTA-uniform
    float4 g = make_float4(ann[j * 5], ann[j * 5 + 1], ann[j * 5 + 2], ann[j * 5 + 3]);
    float area_b = area[j];
#pragma unroll
    for (int u = 0; u < U; u++) {
      float iw = fmaxf(__fsub_rn(fminf(a[u].z, g.z), fmaxf(a[u].x, g.x)), 0.f);
      float ih = fmaxf(__fsub_rn(fminf(a[u].w, g.w), fmaxf(a[u].y, g.y)), 0.f);
      float inter = __fmul_rn(iw, ih), v = 0.f;
      if (inter > 0.f) v = __fdiv_rn(inter, fmaxf(__fsub_rn(__fadd_rn(area_a[u], area_b), inter), 1e-8f));
      if (v > best[u]) { best[u] = v; arg[u] = j; }                // first maximum wins (torch.max on CPU)
    }
  }
#pragma unroll
  for (int u = 0; u < U; u++)                                      // no annotation at all: losses.py:48-79
    state[u] = (arg[u] < 0 || best[u] < 0.4f) ? kNegative : (best[u] >= 0.5f ? arg[u] : kIgnore);
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block sums of three doubles; valid in thread 0.
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c, double (*red)[3]) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  a = warp_sum_d(a); b = warp_sum_d(b); c = warp_sum_d(c);
  if (lane == 0) { red[w][0] = a; red[w][1] = b; red[w][2] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = b = c = 0.0;
    for (int i = 0; i < nw; i++) { a += red[i][0]; b += red[i][1]; c += red[i][2]; }
  }
}

// Box regression targets of a positive anchor (losses.py:123-146) minus the prediction.
__device__ __forceinline__ float4 box_residual(float4 a, float4 g, float4 r) {
  float aw = a.z - a.x, ah = a.w - a.y;
  float acx = a.x + 0.5f * aw, acy = a.y + 0.5f * ah;
  float gw = g.z - g.x, gh = g.w - g.y;
  float gcx = g.x + 0.5f * gw, gcy = g.y + 0.5f * gh;
  gw = fmaxf(gw, 1.f); gh = fmaxf(gh, 1.f);
  float4 t;
  t.x = (gcx - acx) / aw / 0.1f;
  t.y = (gcy - acy) / ah / 0.1f;
  t.z = logf(gw / aw) / 0.2f;
  t.w = logf(gh / ah) / 0.2f;
  return make_float4(t.x - r.x, t.y - r.y, t.z - r.z, t.w - r.w);
}

__device__ __forceinline__ float smooth_l1(float d) {          // losses.py:149-156, beta = 1/9
  float v = fabsf(d);
  return v <= 1.0f / 9.0f ? 4.5f * v * v : v - 0.5f / 9.0f;
}
__device__ __forceinline__ float smooth_l1_grad(float d) {     // d/dd
  float v = fabsf(d);
  float g = v <= 1.0f / 9.0f ? 9.0f * v : 1.0f;
  return d > 0.f ? g : (d < 0.f ? -g : 0.f);
}

// A score clamped to the upper bound is 1 - 1e-4 in exact arithmetic; fp32 cannot hold that bound (0.9999f =
// 1 - 1.00017e-4), so the two quantities that depend on 1 - c take their exact values there, 1e-4 and
// -log(1e-4) (the reference's own fp32 result is off by 1.7e-4 relative on those terms).
// grid (ceil(A / kAnch), B).  CTA = kAnch consecutive anchors of one sample, 4 per thread.
__global__ void __launch_bounds__(kThreads) focal_forward_kernel(
    int A, int K, int M, const float* __restrict__ cls, const float* __restrict__ reg,
    const float4* __restrict__ anchors, const float* __restrict__ ann_all, float alpha,
    int* __restrict__ assign_out, double* __restrict__ partials) {
  extern __shared__ float s_ann[];                     // M x 5, then M box areas
  __shared__ short s_class[kAnch];                     // kIgnore / kNegative / class of the matched annotation
  __shared__ double s_red[kThreads / 32][3];
  int b = blockIdx.y;
  long a0 = (long)blockIdx.x * kAnch;
  int na = (int)min((long)kAnch, A - a0);
  const float* ann = ann_all + (long)b * M * 5;
  for (int i = threadIdx.x; i < M * 5; i += kThreads) s_ann[i] = ann[i];
  float* s_area = s_ann + M * 5;
  for (int j = threadIdx.x; j < M; j += kThreads)
    s_area[j] = __fmul_rn(__fsub_rn(ann[j * 5 + 2], ann[j * 5]), __fsub_rn(ann[j * 5 + 3], ann[j * 5 + 1]));
  __syncthreads();

  float cls_sum = 0.f, reg_sum = 0.f, npos = 0.f;
  constexpr int U = kAnch / kThreads;
  float4 a[U];
  int state[U];
#pragma unroll
  for (int u = 0; u < U; u++) {
    int t = threadIdx.x + u * kThreads;
    a[u] = t < na ? anchors[a0 + t] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  assign_anchors<U>(a, s_ann, s_area, M, state);
#pragma unroll
  for (int u = 0; u < U; u++) {
    int t = threadIdx.x + u * kThreads;
    if (t < na) {
      int st = state[u];
      assign_out[(long)b * A + a0 + t] = st;
      int code = st;
      if (st >= 0) {
        code = (int)(long)s_ann[st * 5 + 4];
        float4 g = make_float4(s_ann[st * 5], s_ann[st * 5 + 1], s_ann[st * 5 + 2], s_ann[st * 5 + 3]);
        float4 r = reinterpret_cast<const float4*>(reg)[(long)b * A + a0 + t];
        float4 d = box_residual(a[u], g, r);
        reg_sum += smooth_l1(d.x) + smooth_l1(d.y) + smooth_l1(d.z) + smooth_l1(d.w);
        npos += 1.f;
        if (code < 0 || code > 32767) code = 32767;              // never equals a class index below
      }
      s_class[t] = (short)code;
    }
  }
  __syncthreads();

  // classification terms: the CTA's na*K scores are contiguous in (B, A, K)
  const float* c0 = cls + ((long)b * A + a0) * K;
  int count = na * K;
#pragma unroll 4
  for (int i = threadIdx.x; i < count; i += kThreads) {
    int al = i / K, k = i - al * K;
    int code = s_class[al];
    float raw = __ldg(c0 + i);
    if (code == kIgnore) continue;
    float c = fminf(fmaxf(raw, 1e-4f), 1.0f - 1e-4f);                       // losses.py:46
    // alpha (1-c)^2 (-log c)  |  (1-alpha) c^2 (-log(1-c))    (losses.py:98-106, gamma = 2); one logarithm
    // per element: the target-1 branch is taken by one class of the few positive anchors only
    const bool top = raw > 1.0f - 1e-4f;
    if (code == k) {
      const float omc = top ? 1e-4f : 1.f - c;
      cls_sum += alpha * omc * omc * -logf(c);
    } else {
      cls_sum += (1.f - alpha) * c * c * (top ? 9.210340371976182f : -log1pf(-c));
    }
  }
  double s0 = cls_sum, s1 = reg_sum, s2 = npos;
  block_sum3(s0, s1, s2, s_red);
  if (threadIdx.x == 0) {
    double* p = partials + ((long)b * gridDim.x + blockIdx.x) * 3;
    p[0] = s0; p[1] = s1; p[2] = s2;
  }
}

// One CTA: per-sample sums in a fixed order, the normalisations of losses.py:113,157 and the batch mean (:169).
// per_sample[b] = { classification loss of sample b, box loss of sample b, number of positive anchors }.
__global__ void __launch_bounds__(1024) focal_finalize_kernel(int B, int nblk, const double* __restrict__ partials,
                                                              double* __restrict__ per_sample,
                                                              float* __restrict__ losses) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int b = w; b < B; b += nw) {
    double s0 = 0, s1 = 0, s2 = 0;
    for (int i = lane; i < nblk; i += 32) {
      const double* p = partials + ((long)b * nblk + i) * 3;
      s0 += p[0]; s1 += p[1]; s2 += p[2];
    }
    s0 = warp_sum_d(s0); s1 = warp_sum_d(s1); s2 = warp_sum_d(s2);
    if (lane == 0) {
      per_sample[b * 3 + 0] = s0 / fmax(s2, 1.0);
      per_sample[b * 3 + 1] = s2 > 0 ? s1 / (4.0 * s2) : 0.0;
      per_sample[b * 3 + 2] = s2;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double c = 0, r = 0;
    for (int b = 0; b < B; b++) { c += per_sample[b * 3]; r += per_sample[b * 3 + 1]; }
    losses[0] = (float)(c / B);
    losses[1] = (float)(r / B);
  }
}

__global__ void __launch_bounds__(kThreads) focal_backward_kernel(
    int B, int A, int K, int M, const float* __restrict__ cls, const float* __restrict__ reg,
    const float4* __restrict__ anchors, const float* __restrict__ ann_all, float alpha,
    const int* __restrict__ assign, const double* __restrict__ per_sample, const float* __restrict__ dloss,
    float* __restrict__ dcls, float* __restrict__ dreg) {
  __shared__ short s_class[kAnch];
  int b = blockIdx.y;
  long a0 = (long)blockIdx.x * kAnch;
  int na = (int)min((long)kAnch, A - a0);
  const float* ann = ann_all + (long)b * M * 5;
  double npos = per_sample[b * 3 + 2];
  float scale_c = (float)((double)dloss[0] / B / fmax(npos, 1.0));
  float scale_r = npos > 0 ? (float)((double)dloss[1] / B / (4.0 * npos)) : 0.f;
#pragma unroll
  for (int u = 0; u < kAnch / kThreads; u++) {
    int t = threadIdx.x + u * kThreads;
    if (t < na) {
      long ia = (long)b * A + a0 + t;
      int st = assign[ia];
      int code = st;
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      if (st >= 0) {
        const float* g5 = ann + st * 5;
        code = (int)(long)g5[4];
        if (code < 0 || code > 32767) code = 32767;
        float4 d = box_residual(anchors[a0 + t], make_float4(g5[0], g5[1], g5[2], g5[3]),
                                reinterpret_cast<const float4*>(reg)[ia]);
        // residual = target - prediction: d/d prediction = -smooth_l1'(residual)
        o = make_float4(-smooth_l1_grad(d.x) * scale_r, -smooth_l1_grad(d.y) * scale_r,
                        -smooth_l1_grad(d.z) * scale_r, -smooth_l1_grad(d.w) * scale_r);
      }
      s_class[t] = (short)code;
      reinterpret_cast<float4*>(dreg)[ia] = o;
    }
  }
  __syncthreads();
  const float* c0 = cls + ((long)b * A + a0) * K;
  float* d0 = dcls + ((long)b * A + a0) * K;
  int count = na * K;
#pragma unroll 4
  for (int i = threadIdx.x; i < count; i += kThreads) {
    int al = i / K, k = i - al * K;
    int code = s_class[al];
    float c = __ldg(c0 + i), g = 0.f;
    if (code != kIgnore && c >= 1e-4f && c <= 1.0f - 1e-4f) {              // clamp passes gradient inside only
      const float omc = 1.f - c;
      if (code == k) g = alpha * (2.f * omc * logf(c) - omc * omc / c);
      else g = (1.f - alpha) * (2.f * c * -log1pf(-c) + c * c / omc);
      g *= scale_c;
    }
    d0[i] = g;
  }
}

size_t focal_assign_bytes(int B, int A) { return align_up((size_t)B * A * sizeof(int)); }

}  // namespace
}  // namespace frn

using namespace frn;

#pragma GCC visibility push(default)
extern "C" {

long frn_anchors_count(int H, int W, int n_levels, const int* pyramid_levels, int boxes_per_cell) {
  if (H <= 0 || W <= 0 || n_levels <= 0 || !pyramid_levels || boxes_per_cell <= 0) return 0;
  long total = 0;
  for (int l = 0; l < n_levels; l++) {
    long s = 1L << pyramid_levels[l];
    total += ((H + s - 1) / s) * ((W + s - 1) / s) * boxes_per_cell;      // anchors.py:30
  }
  return total;
}

int frn_anchors_generate(void* stream, int H, int W, int n_levels, const int* pyramid_levels, const int* strides,
                         const int* sizes, int n_ratios, const double* ratios, int n_scales, const double* scales,
                         float* out) {
  if (!pyramid_levels || !strides || !sizes || !ratios || !scales || !out) return FRN_ERR_NULL_POINTER;
  int nA = n_ratios * n_scales;
  if (H <= 0 || W <= 0 || n_levels <= 0 || n_ratios <= 0 || n_scales <= 0) return FRN_ERR_BAD_SHAPE;
  if (nA > kMaxBase) return FRN_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  long offset = 0;
  for (int l = 0; l < n_levels; l++) {
    if (pyramid_levels[l] < 0 || pyramid_levels[l] > 30) return FRN_ERR_BAD_SHAPE;
    BaseBoxes base;
    // generate_anchors (anchors.py:47-82) in double: box i pairs ratio i / n_scales with scale i % n_scales
    for (int i = 0; i < nA; i++) {
      double ratio = ratios[i / n_scales], side = (double)sizes[l] * scales[i % n_scales];
      double area = side * side;
      double w = sqrt(area / ratio);
      double h = w * ratio;
      base.v[i][0] = 0.0 - w * 0.5; base.v[i][1] = 0.0 - h * 0.5;
      base.v[i][2] = w - w * 0.5;   base.v[i][3] = h - h * 0.5;
    }
    long s = 1L << pyramid_levels[l];
    int fh = (int)((H + s - 1) / s), fw = (int)((W + s - 1) / s);
    long total = (long)fh * fw * nA;
    anchors_level_kernel<<<cdiv(total, kThreads), kThreads, 0, st>>>(fh, fw, nA, strides[l], base,
                                                                     reinterpret_cast<float4*>(out) + offset);
    FRN_LAUNCH_CHECK();
    offset += total;
  }
  return FRN_OK;
}

size_t frn_focal_loss_saved_bytes(int B, int A) {
  if (B <= 0 || A <= 0) return 0;
  return focal_assign_bytes(B, A) + align_up((size_t)B * 3 * sizeof(double));
}
size_t frn_focal_loss_workspace_bytes(int B, int A) {
  if (B <= 0 || A <= 0) return 0;
  return align_up((size_t)B * cdiv(A, kAnch) * 3 * sizeof(double));
}

int frn_focal_loss_forward(void* stream, int B, int A, int K, int M, const float* classifications,
                           const float* regressions, const float* anchors, const float* annotations, float alpha,
                           float* losses, void* saved, size_t saved_bytes, void* ws, size_t ws_bytes) {
  if (B <= 0 || A <= 0 || K <= 0 || M < 0) return FRN_ERR_BAD_SHAPE;
  if (M > kMaxAnn) return FRN_ERR_UNSUPPORTED;
  if (!classifications || !regressions || !anchors || !losses || !saved || !ws || (M > 0 && !annotations))
    return FRN_ERR_NULL_POINTER;
  if (saved_bytes < frn_focal_loss_saved_bytes(B, A) || ws_bytes < frn_focal_loss_workspace_bytes(B, A))
    return FRN_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  int* assign = (int*)saved;
  double* per_sample = (double*)((char*)saved + focal_assign_bytes(B, A));
  int nblk = cdiv(A, kAnch);
  ScopedTimer tm(st, FRN_TIME_ROWS);
  focal_forward_kernel<<<dim3(nblk, B), kThreads, (size_t)M * 6 * sizeof(float), st>>>(
      A, K, M, classifications, regressions, reinterpret_cast<const float4*>(anchors), annotations, alpha, assign,
      (double*)ws);
  FRN_LAUNCH_CHECK();
  focal_finalize_kernel<<<1, 1024, 0, st>>>(B, nblk, (const double*)ws, per_sample, losses);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int frn_focal_loss_backward(void* stream, int B, int A, int K, int M, const float* classifications,
                            const float* regressions, const float* anchors, const float* annotations, float alpha,
                            const void* saved, const float* dlosses, float* dclassifications, float* dregressions) {
  if (B <= 0 || A <= 0 || K <= 0 || M < 0) return FRN_ERR_BAD_SHAPE;
  if (M > kMaxAnn) return FRN_ERR_UNSUPPORTED;
  if (!classifications || !regressions || !anchors || !saved || !dlosses || !dclassifications || !dregressions ||
      (M > 0 && !annotations))
    return FRN_ERR_NULL_POINTER;
  cudaStream_t st = (cudaStream_t)stream;
  const int* assign = (const int*)saved;
  const double* per_sample = (const double*)((const char*)saved + focal_assign_bytes(B, A));
  ScopedTimer tm(st, FRN_TIME_ROWS);
  focal_backward_kernel<<<dim3(cdiv(A, kAnch), B), kThreads, 0, st>>>(
      B, A, K, M, classifications, regressions, reinterpret_cast<const float4*>(anchors), annotations, alpha, assign,
      per_sample, dlosses, dclassifications, dregressions);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // extern "C"
#pragma GCC visibility pop
