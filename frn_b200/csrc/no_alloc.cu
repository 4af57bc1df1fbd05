// frn_adain_backward_static: frn_adain_backward for callers that cannot hand over a workspace because they cannot
// call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  No counterpart
// in the reference.
//
// The two per-row reductions of the AdaIN backward (g1 = sum_n dout, g2 = sum_n dout (x - mu) / sd, B*C floats
// each -- the `ws` of frn_adain_backward) live in a file-scope __device__ array instead of caller memory:
//   - 2 * FRN_ADAIN_STATIC_MAX_ROWS floats = 512 KiB, held by every GPU that loads the library until exit;
//   - B*C > FRN_ADAIN_STATIC_MAX_ROWS is rejected with FRN_ERR_STATIC_SCRATCH before anything is enqueued;
//   - nothing has to start at zero: adain_bwd_reduce_kernel overwrites rows [0, B*C) of both halves before
//     the content / style kernels (same stream, later) read them, so the previous call's contents are never seen;
//   - every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads, two
//     graph branches) race.  Serialise them on one stream, or use frn_adain_backward with a workspace per caller.
// The entry point calls no cudaMalloc / cudaFree / allocating library routine; it enqueues the same three kernels as
// frn_adain_backward, so the results are bit-identical.
#include <mutex>

#include "common.cuh"

namespace frn {
namespace {

constexpr long kMaxRows = FRN_ADAIN_STATIC_MAX_ROWS;
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) float g_adain_static_scratch[2 * kMaxRows];

std::mutex g_mu;
float* g_scratch_ptr[kMaxDevices] = {};   // device address of the array, per device

// Device address of the array on the CURRENT device (looked up once per device, with that device current).
int static_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return FRN_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_adain_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!p) return FRN_ERR_NO_DEVICE;
    g_scratch_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_ptr[dev];
  return FRN_OK;
}

}  // namespace
}  // namespace frn

using namespace frn;

#pragma GCC visibility push(default)
extern "C" {

long frn_adain_static_max_rows(void) { return kMaxRows; }

int frn_adain_backward_static(void* stream, int B, int C, int N, int N_style, const float* content,
                              const float* style, const float* stats, const float* dout, float* dcontent,
                              float* dstyle) {
  if (B <= 0 || C <= 0 || N < 2 || N_style < 2) return FRN_ERR_BAD_SHAPE;
  if (!content || !style || !stats || !dout || !dcontent || !dstyle) return FRN_ERR_NULL_POINTER;
  const long rows = (long)B * C;
  if (rows > kMaxRows) return FRN_ERR_STATIC_SCRATCH;
  float* scratch = nullptr;
  FRN_TRY(static_scratch(&scratch));
  cudaStream_t st = (cudaStream_t)stream;
  const float *mu_c = stats, *sd_c = stats + rows, *mu_s = stats + 2 * rows, *sd_s = stats + 3 * rows;
  float* g1 = scratch;
  float* g2 = scratch + kMaxRows;
  const long cn = (long)C * N;
  FRN_TRY(adain_bwd_reduce(st, B, C, N, dout, cn, content, mu_c, sd_c, g1, g2));
  FRN_TRY(adain_bwd_content(st, B, C, N, dout, cn, content, mu_c, sd_c, sd_s, g1, g2, dcontent, 0));
  return adain_bwd_style(st, B, C, N_style, style, mu_s, sd_s, g1, g2, dstyle);
}

}  // extern "C"
#pragma GCC visibility pop
