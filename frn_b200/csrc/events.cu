// event voxelisation.
//
//   retinanet/data/dsec_data.py:347-387 (preprocess_events): a per-event Python `for` loop scatter-adds the
//   polarity (+1 / -1) into a (C, H, W) grid -- 1e5 ... 1e6 interpreted iterations per sample, the reason the
//   reference's train.log shows seconds per iteration -- and :78-93 (normalize_event_data) squashes the grid.
//
// Here: one pass finds the first / last in-range event (their timestamps define the time window), one pass
// scatters with fp32 atomics (sums of +-1 are exact in fp32 below 2^24, so the result does not depend on the
// order and is bit-identical to the sequential loop), one optional pass normalises in place.  Integer / index
// work throughout; the time bin is computed in double exactly as NumPy does.
#include <math.h>
#include "common.cuh"

namespace frn {
namespace {

constexpr int kThreads = 256;

template <typename XY>
__device__ __forceinline__ bool in_mask(XY x, XY y, int H, int W) {      // (events['y'] < H) & (events['x'] < W)
  return (long)y < (long)H && (long)x < (long)W;
}

// bounds[0] = index of the first masked-in event, bounds[1] = index of the last one (-1: none).
template <typename XY>
__global__ void __launch_bounds__(kThreads) event_bounds_kernel(long n, const XY* __restrict__ x,
                                                                const XY* __restrict__ y, int H, int W,
                                                                int* __restrict__ bounds) {
  long stride = (long)gridDim.x * blockDim.x;
  int lo = 0x7fffffff, hi = -1;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += stride)
    if (in_mask(x[i], y[i], H, W)) { lo = min(lo, (int)i); hi = max(hi, (int)i); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0 && hi >= 0) { atomicMin(&bounds[0], lo); atomicMax(&bounds[1], hi); }
}

template <typename XY>
__global__ void __launch_bounds__(kThreads) event_scatter_kernel(long n, const XY* __restrict__ x,
                                                                 const XY* __restrict__ y,
                                                                 const long long* __restrict__ t,
                                                                 const unsigned char* __restrict__ p, int C, int H,
                                                                 int W, const int* __restrict__ bounds,
                                                                 float* __restrict__ voxel) {
  int last = bounds[1];
  if (last < 0) return;                                            // no event in range: the grid stays zero
  long long t_min = t[bounds[0]], t_max = t[last];
  double span = __dadd_rn((double)(t_max - t_min), 1e-6);          // dsec_data.py:361
  double cm1 = (double)(C - 1);
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += stride) {
    XY xi = x[i], yi = y[i];
    if (!in_mask(xi, yi, H, W) || (long)xi < 0 || (long)yi < 0) continue;   // :353 and the check at :380
    double tn = __ddiv_rn((double)(t[i] - t_min), span);
    long long bin = (long long)__dmul_rn(tn, cm1);                 // astype(int64): truncation
    bin = bin < 0 ? 0 : (bin > C - 1 ? C - 1 : bin);               // np.clip (:368)
    atomicAdd(voxel + ((long)bin * H + (long)yi) * W + (long)xi, p[i] > 0 ? 1.f : -1.f);
  }
}

// normalize_event_data (dsec_data.py:78-93): 1 = tanh(v / 5), 2 = clamp to [-2, 2].
__global__ void __launch_bounds__(kThreads) voxel_normalize_kernel(long total, float* __restrict__ voxel, int mode) {
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    float v = voxel[i];
    // voxel values are small integers: tanh in double, rounded once (correctly rounded but for double's own error)
    voxel[i] = mode == 1 ? (float)tanh((double)(v / 5.0f)) : fminf(fmaxf(v, -2.f), 2.f);
  }
}

template <typename XY>
int voxelize(cudaStream_t st, long n, const void* x, const void* y, const long long* t, const unsigned char* p, int C,
             int H, int W, int normalize, float* voxel, int* bounds) {
  long total = (long)C * H * W;
  cudaError_t e = cudaMemsetAsync(voxel, 0, total * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  if (n > 0) {
    e = cudaMemsetAsync(bounds, 0x7f, sizeof(int), st);               // first = 0x7f7f7f7f (> any index)
    if (e == cudaSuccess) e = cudaMemsetAsync(bounds + 1, 0xff, sizeof(int), st);      // last = -1
    if (e != cudaSuccess) return (int)e;
    int grid = (int)min((long)148 * 8, (n + kThreads - 1) / kThreads);
    event_bounds_kernel<XY><<<grid, kThreads, 0, st>>>(n, (const XY*)x, (const XY*)y, H, W, bounds);
    FRN_LAUNCH_CHECK();
    event_scatter_kernel<XY><<<grid, kThreads, 0, st>>>(n, (const XY*)x, (const XY*)y, t, p, C, H, W, bounds, voxel);
    FRN_LAUNCH_CHECK();
  }
  if (normalize) {
    voxel_normalize_kernel<<<(int)min((long)148 * 8, (total + kThreads - 1) / kThreads), kThreads, 0, st>>>(
        total, voxel, normalize);
    FRN_LAUNCH_CHECK();
  }
  return FRN_OK;
}

}  // namespace
}  // namespace frn

using namespace frn;

#pragma GCC visibility push(default)
extern "C" {

size_t frn_voxelize_workspace_bytes(void) { return 256; }

int frn_voxelize_events(void* stream, long n_events, int xy_bytes, const void* x, const void* y, const int64_t* t,
                        const uint8_t* p, int C, int H, int W, int normalize, float* voxel, void* ws,
                        size_t ws_bytes) {
  if (n_events < 0 || n_events > 0x7f000000L || C <= 0 || H <= 0 || W <= 0) return FRN_ERR_BAD_SHAPE;
  if (xy_bytes != 2 && xy_bytes != 4) return FRN_ERR_UNSUPPORTED;
  if (normalize < 0 || normalize > 2) return FRN_ERR_UNSUPPORTED;
  if (!voxel || !ws || (n_events > 0 && (!x || !y || !t || !p))) return FRN_ERR_NULL_POINTER;
  if (ws_bytes < frn_voxelize_workspace_bytes()) return FRN_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  ScopedTimer tm(st, FRN_TIME_ROWS);
  return xy_bytes == 2
             ? voxelize<unsigned short>(st, n_events, x, y, (const long long*)t, p, C, H, W, normalize, voxel, (int*)ws)
             : voxelize<int>(st, n_events, x, y, (const long long*)t, p, C, H, W, normalize, voxel, (int*)ws);
}

int frn_voxel_normalize(void* stream, long count, int mode, float* voxel) {
  if (count < 0) return FRN_ERR_BAD_SHAPE;
  if (mode != 1 && mode != 2) return FRN_ERR_UNSUPPORTED;
  if (!voxel && count > 0) return FRN_ERR_NULL_POINTER;
  if (count == 0) return FRN_OK;
  voxel_normalize_kernel<<<(int)min((long)148 * 8, (count + kThreads - 1) / kThreads), kThreads, 0,
                           (cudaStream_t)stream>>>(count, voxel, mode);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // extern "C"
#pragma GCC visibility pop
