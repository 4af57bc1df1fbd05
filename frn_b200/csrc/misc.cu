// Small glue kernels: per-(sample, channel) coefficient math, bias-gradient
// row sums, transposes, and the row softmax used by the materialised
// attention path for d > 64 (retinanet/Attention.py:30).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <set>
#include <tuple>
#include <string>
#include <utility>
#include <vector>
#include <nvtx3/nvToolsExt.h>
#include "common.cuh"

namespace frn {
namespace {

constexpr float kLog2e = 1.4426950408889634f;

__global__ void style_coeffs_kernel(long rows, int N, const float* __restrict__ g1,
                                    const float* __restrict__ g2, const float* __restrict__ sd_t,
                                    float* __restrict__ alpha, float* __restrict__ beta) {
  long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (i >= rows) return;
  alpha[i] = g1[i] / (float)N;
  beta[i] = g2[i] / ((float)(N - 1) * sd_t[i]);
}

__global__ void finish_style_std_kernel(long rows, int nparts, int N, const float* __restrict__ parts,
                                        float eps, float* __restrict__ sd_t) {
  long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (i >= rows) return;
  float s = 0.f;
  for (int p = 0; p < nparts; p++) s += parts[i * nparts + p];
  sd_t[i] = sqrtf(s / (float)(N - 1) + eps);
}

// out[m] = sum_b sum_n x[b*sb + m*sm + n*sn]; one CTA per m.
__global__ void __launch_bounds__(256) batch_row_sum_kernel(int B, int M, int N, const float* __restrict__ x,
                                                            long sb, long sm, long sn,
                                                            float* __restrict__ out) {
  __shared__ float red[33];
  int m = blockIdx.x;
  float s = 0.f;
  for (int b = 0; b < B; b++) {
    const float* p = x + b * sb + m * sm;
    float t = 0.f;
    for (int n = threadIdx.x; n < N; n += blockDim.x) t += __ldg(p + n * sn);
    s += t;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[m] = s;
}

// Channel-major maps (B, M, N), N contiguous: stage 1, one CTA per (m, b) row -> part[b * M + m]; stage 2 adds the
// B partial sums of a channel in a fixed order.  (One CTA per channel walking all B rows left 256 CTAs with a
// handful of loads in flight each to read 157 MB at DSEC level 1.)
__global__ void __launch_bounds__(256) row_sum_kernel(int M, int N, const float* __restrict__ x, long sb, long sm,
                                                      float* __restrict__ part) {
  __shared__ float red[33];
  const int m = blockIdx.x, b = blockIdx.y;
  const float* p = x + b * sb + m * sm;
  float s = 0.f;
  if ((N & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    const float4* p4 = reinterpret_cast<const float4*>(p);
    const int n4 = N / 4, step = blockDim.x;
    float s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int n = threadIdx.x;
    for (; n + 3 * step < n4; n += 4 * step) {        // four independent loads in flight per thread
      const float4 v0 = __ldg(p4 + n), v1 = __ldg(p4 + n + step), v2 = __ldg(p4 + n + 2 * step), v3 = __ldg(p4 + n + 3 * step);
      s += (v0.x + v0.y) + (v0.z + v0.w); s1 += (v1.x + v1.y) + (v1.z + v1.w);
      s2 += (v2.x + v2.y) + (v2.z + v2.w); s3 += (v3.x + v3.y) + (v3.z + v3.w);
    }
    for (; n < n4; n += step) {
      float4 v = __ldg(p4 + n);
      s += (v.x + v.y) + (v.z + v.w);
    }
    s = (s + s1) + (s2 + s3);
  } else {
    for (int n = threadIdx.x; n < N; n += blockDim.x) s += __ldg(p + n);
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) part[(long)b * M + m] = s;
}
__global__ void sum_parts_kernel(int M, int parts, const float* __restrict__ part, float* __restrict__ out) {
  int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  float s = 0.f;
  for (int q = 0; q < parts; q++) s += part[(long)q * M + m];
  out[m] = s;
}

__global__ void gate_backward_kernel(long count, const float* __restrict__ r0, const float* __restrict__ e0,
                                     const float* dr1, const float* de1,
                                     float* dr0, float* de0) {  // dr0/de0 may alias dr1/de1
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < count; i += stride) {
    float r = r0[i], e = e0[i], gr = dr1[i], ge = de1[i];
    dr0[i] = fmaf(gr, 1.f + e, ge * e);
    de0[i] = fmaf(ge, 1.f + r, gr * r);
  }
}
__global__ void gate_backward_kernel4(long count4, const float4* __restrict__ r0,
                                      const float4* __restrict__ e0, const float4* dr1,
                                      const float4* de1, float4* dr0, float4* de0) {
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < count4; i += stride) {
    float4 r = r0[i], e = e0[i], gr = dr1[i], ge = de1[i], a, b;
    a.x = fmaf(gr.x, 1.f + e.x, ge.x * e.x); b.x = fmaf(ge.x, 1.f + r.x, gr.x * r.x);
    a.y = fmaf(gr.y, 1.f + e.y, ge.y * e.y); b.y = fmaf(ge.y, 1.f + r.y, gr.y * r.y);
    a.z = fmaf(gr.z, 1.f + e.z, ge.z * e.z); b.z = fmaf(ge.z, 1.f + r.z, gr.z * r.z);
    a.w = fmaf(gr.w, 1.f + e.w, ge.w * e.w); b.w = fmaf(ge.w, 1.f + r.w, gr.w * r.w);
    dr0[i] = a; de0[i] = b;
  }
}

// dst (B, cols, rows) <- src (B, rows, cols), 32x32 tiles through smem.
__global__ void transpose_kernel(int rows, int cols, const float* __restrict__ src, float* __restrict__ dst) {
  __shared__ float tile[32][33];
  long bo = (long)blockIdx.z * rows * cols;
  int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? src[bo + (long)r * cols + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    int c = c0 + i, r = r0 + threadIdx.x;
    if (r < rows && c < cols) dst[bo + (long)c * rows + r] = tile[threadIdx.x][i];
  }
}

__global__ void scale_rows_kernel(long total, int C, int K, const float* __restrict__ w,
                                  const float* __restrict__ beta, float* __restrict__ scaled) {
  long stride = (long)gridDim.x * blockDim.x;
  long ck = (long)C * K;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    long b = i / ck, r = i % ck;
    scaled[i] = beta[b * C + r / K] * w[r];
  }
}

// dW[c][k] (+)= scale * sum_b alpha[b][c] * ybar[b][k]
__global__ void outer_accumulate_kernel(int B, int C, int K, const float* __restrict__ alpha,
                                        const float* __restrict__ ybar, float scale,
                                        float* __restrict__ dW) {
  long total = (long)C * K;
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    int c = (int)(i / K), k = (int)(i % K);
    float s = 0.f;
    for (int b = 0; b < B; b++) s = fmaf(alpha[(long)b * C + c], ybar[(long)b * K + k], s);
    dW[i] += scale * s;
  }
}

// Column sums of x (B, N, d) over a chunk of positions: part[(b*G + g)*d + k].  grid (G, B), 256 threads.
__global__ void __launch_bounds__(256)
col_partial_sum_kernel(int N, int d, int rows_per_block, const float* __restrict__ x, float* __restrict__ part) {
  __shared__ float red[256];
  const int b = blockIdx.y, g = blockIdx.x, G = gridDim.x;
  const int n0 = g * rows_per_block, n1 = min(N, n0 + rows_per_block);
  const int lanes = 256 / d * d;                 // threads used: a whole number of rows per pass
  const int tid = threadIdx.x, k = tid % d, r = tid / d, rstep = 256 / d;
  float s = 0.f;
  if (tid < lanes) {
    // four independent partial sums: the loads of a thread are 128 * rstep bytes apart and a single dependent
    // accumulator kept only one or two of them in flight
    const float* xp = x + (long)b * N * d + k;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int n = n0 + r;
    for (; n + 3 * rstep < n1; n += 4 * rstep) {
      s0 += __ldg(xp + (long)n * d);
      s1 += __ldg(xp + (long)(n + rstep) * d);
      s2 += __ldg(xp + (long)(n + 2 * rstep) * d);
      s3 += __ldg(xp + (long)(n + 3 * rstep) * d);
    }
    for (; n < n1; n += rstep) s0 += __ldg(xp + (long)n * d);
    s = (s0 + s1) + (s2 + s3);
  }
  red[tid] = tid < lanes ? s : 0.f;
  __syncthreads();
  if (tid < d) {
    float t = 0.f;
    for (int q = 0; q < rstep; q++) t += red[q * d + tid];
    part[((long)b * G + g) * d + tid] = t;
  }
}
// x[b][n][k] -= (sum_g part[b][g][k]) / N
__global__ void __launch_bounds__(256)
col_center_kernel(int N, int d, int G, int rows_per_block, float* __restrict__ x, const float* __restrict__ part) {
  __shared__ float mean[256];
  const int b = blockIdx.y, n0 = blockIdx.x * rows_per_block, n1 = min(N, n0 + rows_per_block);
  const int tid = threadIdx.x;
  if (tid < d) {
    float t = 0.f;
    for (int g = 0; g < G; g++) t += part[((long)b * G + g) * d + tid];
    mean[tid] = t / (float)N;
  }
  __syncthreads();
  float* xb = x + ((long)b * N + n0) * d;
  const int total = (n1 - n0) * d;
  for (int i = tid; i < total; i += 256) xb[i] -= mean[i % d];
}

__global__ void fill_zero_kernel(float* __restrict__ p, size_t count) {
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += stride) p[i] = 0.f;
}

// bf16 -> fp32 is a 16-bit shift; eight elements (one 16-byte load, two 16-byte stores) per thread and step
__global__ void cast_bf16_to_f32_kernel(size_t count, const uint16_t* __restrict__ src, float* __restrict__ dst,
                                        int vec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t done = 0;
  if (vec) {
    const size_t n8 = count / 8;
    for (size_t i = t0; i < n8; i += stride) {
      const uint4 r = __ldg(reinterpret_cast<const uint4*>(src) + i);
      float4 lo4, hi4;
      lo4.x = __uint_as_float(r.x << 16); lo4.y = __uint_as_float(r.x & 0xffff0000u);
      lo4.z = __uint_as_float(r.y << 16); lo4.w = __uint_as_float(r.y & 0xffff0000u);
      hi4.x = __uint_as_float(r.z << 16); hi4.y = __uint_as_float(r.z & 0xffff0000u);
      hi4.z = __uint_as_float(r.w << 16); hi4.w = __uint_as_float(r.w & 0xffff0000u);
      reinterpret_cast<float4*>(dst)[2 * i] = lo4;
      reinterpret_cast<float4*>(dst)[2 * i + 1] = hi4;
    }
    done = n8 * 8;
  }
  for (size_t i = done + t0; i < count; i += stride) dst[i] = __uint_as_float((uint32_t)src[i] << 16);
}

__global__ void scale_sum_batch_kernel(int B, int C, const float* __restrict__ alpha, float scale,
                                       float* __restrict__ out) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (int b = 0; b < B; b++) s += alpha[(long)b * C + c];
  out[c] = scale * s;
}

// ---- materialised attention helpers (one CTA per row of length N) -------------
__global__ void __launch_bounds__(256) softmax_rows_kernel(int N, float* __restrict__ S, float* __restrict__ lse) {
  __shared__ float red[33];
  float* r = S + (long)blockIdx.x * N;
  float mx = -INFINITY;
  for (int n = threadIdx.x; n < N; n += blockDim.x) mx = fmaxf(mx, r[n]);
  mx = warp_max(mx);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : -INFINITY;
    t = warp_max(t);
    if (threadIdx.x == 0) red[32] = t;
  }
  __syncthreads();
  mx = red[32];
  float s = 0.f;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    float p = exp2f((r[n] - mx) * kLog2e);
    r[n] = p;
    s += p;
  }
  s = block_sum(s, red);
  float inv = 1.f / s;
  for (int n = threadIdx.x; n < N; n += blockDim.x) r[n] *= inv;
  if (threadIdx.x == 0) lse[blockIdx.x] = mx + logf(s);
}

__global__ void __launch_bounds__(256) exp_sub_lse_kernel(int N, float* __restrict__ S, const float* __restrict__ lse) {
  float* r = S + (long)blockIdx.x * N;
  float l = lse[blockIdx.x];
  for (int n = threadIdx.x; n < N; n += blockDim.x) r[n] = exp2f((r[n] - l) * kLog2e);
}

__global__ void __launch_bounds__(256) ds_from_p_kernel(int N, const float* __restrict__ P, float* __restrict__ dP,
                                                        const float* __restrict__ delta) {
  const float* p = P + (long)blockIdx.x * N;
  float* g = dP + (long)blockIdx.x * N;
  float de = delta[blockIdx.x];
  for (int n = threadIdx.x; n < N; n += blockDim.x) g[n] = p[n] * (g[n] - de);
}

inline int grid_for(long total, int block = 256) {
  long g = (total + block - 1) / block;
  if (g > 148L * 16) g = 148L * 16;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

std::atomic<unsigned long long> g_launches{0};

namespace {
std::atomic<int> g_timing_on{0};
std::mutex g_timing_mu;
std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_events[FRN_TIME_CLASSES];
}  // namespace

bool timing_enabled() { return g_timing_on.load(std::memory_order_relaxed) != 0; }

ScopedTimer::ScopedTimer(cudaStream_t s, int c) : st(s), cls(c), e0(nullptr), on(false) {
  if (!g_timing_on.load(std::memory_order_relaxed)) return;
  if (cudaEventCreate(&e0) != cudaSuccess) return;
  cudaEventRecord(e0, st);
  on = true;
}
ScopedTimer::~ScopedTimer() {
  if (!on) return;
  cudaEvent_t e1;
  if (cudaEventCreate(&e1) != cudaSuccess) { cudaEventDestroy(e0); return; }
  cudaEventRecord(e1, st);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  g_events[cls].emplace_back(e0, e1);
}

namespace {
struct KernelRec { const char* name; double bytes; cudaEvent_t e0, e1; };
std::vector<KernelRec> g_kernel_log;
}  // namespace

KernelTimer::KernelTimer(cudaStream_t s, const char* name_, double bytes_)
    : st(s), e0(nullptr), on(false), name(name_), bytes(bytes_) {
  if (!g_timing_on.load(std::memory_order_relaxed)) return;
  if (cudaEventCreate(&e0) != cudaSuccess) return;
  cudaEventRecord(e0, st);
  on = true;
}
KernelTimer::~KernelTimer() {
  if (!on) return;
  cudaEvent_t e1;
  if (cudaEventCreate(&e1) != cudaSuccess) { cudaEventDestroy(e0); return; }
  cudaEventRecord(e1, st);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  g_kernel_log.push_back({name, bytes, e0, e1});
}

// "name | launches | ms | bytes | launches_large | ms_large | bytes_large" per kernel; large = launches whose
// algorithmic bytes are at least twice the 126 MB L2 (those are the ones an HBM roofline fraction is meaningful for,
// and long enough that the two event records around a launch do not distort the time).
long kernel_log_dump(char* buf, long cap) {
  std::vector<KernelRec> ev;
  {
    std::lock_guard<std::mutex> lk(g_timing_mu);
    ev.swap(g_kernel_log);
  }
  struct Agg { long n = 0; double ms = 0, bytes = 0; long nl = 0; double msl = 0, bytesl = 0; };
  std::map<std::string, Agg> agg;
  for (auto& r : ev) {
    float t = 0.f;
    cudaEventSynchronize(r.e1);
    cudaEventElapsedTime(&t, r.e0, r.e1);
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
    Agg& a = agg[r.name];
    a.n++; a.ms += t; a.bytes += r.bytes;
    if (r.bytes > 250e6) { a.nl++; a.msl += t; a.bytesl += r.bytes; }
  }
  std::string out;
  char line[256];
  for (auto& kv : agg) {
    snprintf(line, sizeof line, "%s | %ld | %.5f | %.0f | %ld | %.5f | %.0f\n", kv.first.c_str(), kv.second.n, kv.second.ms,
             kv.second.bytes, kv.second.nl, kv.second.msl, kv.second.bytesl);
    out += line;
  }
  if (buf && cap > 0) {
    long n = (long)out.size() < cap - 1 ? (long)out.size() : cap - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return (long)out.size();
}

namespace {
struct ShapeRec { std::string key; cudaEvent_t e0, e1; };
std::vector<ShapeRec> g_shape_log;
int g_shape_on = -1;
}  // namespace

ShapeTimer::ShapeTimer(cudaStream_t s, const char* engine, int M, int N, int K, int batch, int flags)
    : st(s), e0(nullptr), on(false) {
  if (!g_timing_on.load(std::memory_order_relaxed)) return;
  if (g_shape_on < 0) g_shape_on = getenv("FRN_GEMM_LOG") ? 1 : 0;
  if (!g_shape_on) return;
  snprintf(key, sizeof key, "%s M=%d N=%d K=%d b=%d f=%d", engine, M, N, K, batch, flags);
  if (cudaEventCreate(&e0) != cudaSuccess) return;
  cudaEventRecord(e0, st);
  on = true;
}
ShapeTimer::~ShapeTimer() {
  if (!on) return;
  cudaEvent_t e1;
  if (cudaEventCreate(&e1) != cudaSuccess) { cudaEventDestroy(e0); return; }
  cudaEventRecord(e1, st);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  g_shape_log.push_back({key, e0, e1});
}

NvtxRange::NvtxRange(const char* name) { nvtxRangePushA(name); }
NvtxRange::~NvtxRange() { nvtxRangePop(); }

int set_smem_once(const void* kern, int bytes) {
  static std::mutex mu;
  static std::set<std::pair<const void*, int>> done;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(mu);
  if (done.count({kern, dev})) return FRN_OK;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return (int)e;
  done.insert({kern, dev});
  return FRN_OK;
}

namespace {
struct SideSet { cudaStream_t s; cudaEvent_t fork, join; };
std::mutex g_side_mu;
std::map<std::tuple<int, cudaStream_t, int>, SideSet> g_sides;
int g_no_fork = -1;
}  // namespace

Fork fork_stream(cudaStream_t st, int slot) {
  Fork f{st, nullptr, nullptr, nullptr};
  if (g_no_fork < 0) g_no_fork = getenv("FRN_NO_FORK") ? 1 : 0;       // FRN_NO_FORK=1: everything on the caller's stream
  if (g_no_fork || timing_enabled()) return f;                        // hooks on: kernels back to back, times separable
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return f;
  SideSet set{};
  {
    std::lock_guard<std::mutex> lk(g_side_mu);
    auto key = std::make_tuple(dev, st, slot);
    auto it = g_sides.find(key);
    if (it == g_sides.end()) {
      SideSet n{};
      if (cudaStreamCreateWithFlags(&n.s, cudaStreamNonBlocking) != cudaSuccess) return f;
      if (cudaEventCreateWithFlags(&n.fork, cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&n.join, cudaEventDisableTiming) != cudaSuccess)
        return f;
      it = g_sides.emplace(key, n).first;
    }
    set = it->second;
  }
  if (cudaEventRecord(set.fork, st) != cudaSuccess || cudaStreamWaitEvent(set.s, set.fork, 0) != cudaSuccess) {
    cudaGetLastError();
    return f;
  }
  f.side = set.s; f.ev_fork = set.fork; f.ev_join = set.join;
  return f;
}

void join_stream(const Fork& f) {
  if (!f.side) return;
  cudaEventRecord(f.ev_join, f.side);
  cudaStreamWaitEvent(f.main, f.ev_join, 0);
}

long shape_log_dump(char* buf, long cap) {
  std::vector<ShapeRec> ev;
  {
    std::lock_guard<std::mutex> lk(g_timing_mu);
    ev.swap(g_shape_log);
  }
  std::map<std::string, std::pair<long, double>> agg;
  for (auto& r : ev) {
    float t = 0.f;
    cudaEventSynchronize(r.e1);
    cudaEventElapsedTime(&t, r.e0, r.e1);
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
    auto& a = agg[r.key];
    a.first += 1; a.second += t;
  }
  std::string out;
  char line[160];
  for (auto& kv : agg) {
    snprintf(line, sizeof line, "%s | %ld | %.4f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    out += line;
  }
  if (buf && cap > 0) {
    long n = (long)out.size() < cap - 1 ? (long)out.size() : cap - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return (long)out.size();
}

int style_coeffs(cudaStream_t st, long rows, int N, const float* g1, const float* g2, const float* sd_t,
                 float* alpha, float* beta) {
  style_coeffs_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(rows, N, g1, g2, sd_t, alpha, beta);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int finish_style_std(cudaStream_t st, long rows, int nparts, int N, const float* parts, float eps, float* sd_t) {
  finish_style_std_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(rows, nparts, N, parts, eps, sd_t);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int center_cols_parts(int N) { return cdiv(N, 256) < 64 ? cdiv(N, 256) : 64; }

size_t batch_row_sum_scratch(int B, int M, int N) { return (size_t)B * center_cols_parts(N) * M; }

int batch_row_sum(cudaStream_t st, int B, int M, int N, const float* x, long sb, long sm, long sn, float* out,
                  float* scratch) {
  ScopedTimer tm(st, FRN_TIME_ROWS);
  if (scratch && sm == 1 && sn == M && sb == (long)N * M && M <= 256) {
    // position-major (B, N, M): coalesced column sums over chunks of positions, then one small reduction
    const int G = center_cols_parts(N), rpb = cdiv(N, G);
    {
      KernelTimer kt(st, "col_partial_sum_kernel", 4.0 * B * N * M);
      col_partial_sum_kernel<<<dim3(G, B), 256, 0, st>>>(N, M, rpb, x, scratch);
    }
    FRN_LAUNCH_CHECK();
    batch_row_sum_kernel<<<M, 256, 0, st>>>(1, M, B * G, scratch, 0, 1, M, out);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  if (scratch && sn == 1 && N >= 64 && B <= 65535 && center_cols_parts(N) >= 1) {     // scratch holds >= B * M floats
    {
      KernelTimer kt(st, "row_sum_kernel", 4.0 * B * N * M);
      row_sum_kernel<<<dim3(M, B), 256, 0, st>>>(M, N, x, sb, sm, scratch);
    }
    FRN_LAUNCH_CHECK();
    sum_parts_kernel<<<cdiv(M, 256), 256, 0, st>>>(M, B, scratch, out);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  batch_row_sum_kernel<<<M, 256, 0, st>>>(B, M, N, x, sb, sm, sn, out);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

__global__ void stack3_kernel(long n, const float* __restrict__ a, const float* __restrict__ b,
                              const float* __restrict__ c, float* __restrict__ out) {
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < 3 * n; i += stride)
    out[i] = i < n ? a[i] : (i < 2 * n ? b[i - n] : c[i - 2 * n]);
}
// out = [a; b; c], n floats each (stacked weights / biases of the fused projection GEMM)
int stack3(cudaStream_t st, long n, const float* a, const float* b, const float* c, float* out) {
  stack3_kernel<<<grid_for(3 * n), 256, 0, st>>>(n, a, b, c, out);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

__global__ void unstack3_kernel(long n, const float* __restrict__ in, float* __restrict__ a, float* __restrict__ b,
                                float* __restrict__ c) {
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < 3 * n; i += stride) {
    float v = in[i];
    if (i < n) a[i] = v; else if (i < 2 * n) b[i - n] = v; else c[i - 2 * n] = v;
  }
}
int unstack3(cudaStream_t st, long n, const float* in, float* a, float* b, float* c) {
  unstack3_kernel<<<grid_for(3 * n), 256, 0, st>>>(n, in, a, b, c);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
// out (rows, 3 d) = [a | b | c], each (rows, d) row-major: the position-major gradients of the three projections
// that read the same map, side by side, so that their dgrads / wgrads run as one GEMM with K (resp. M) = 3 d.
__global__ void stack_cols3_kernel(long rows, int d, const float* __restrict__ a, const float* __restrict__ b,
                                   const float* __restrict__ c, float* __restrict__ out) {
  const long total = rows * 3 * d, stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    const long r = i / (3 * d);
    const int col = (int)(i - r * 3 * d), g = col / d, k = col - g * d;
    const float* src = g == 0 ? a : (g == 1 ? b : c);
    out[i] = __ldg(src + r * d + k);
  }
}
// d % 4 == 0 and 16-byte aligned buffers: 128-bit copies, one row segment (d / 4 chunks) per source
__global__ void stack_cols3_v4_kernel(long rows, int d4, const float4* __restrict__ a, const float4* __restrict__ b,
                                      const float4* __restrict__ c, float4* __restrict__ out) {
  const long total = rows * 3 * d4, stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    const long r = i / (3 * d4);
    const int col = (int)(i - r * 3 * d4), g = col / d4, k = col - g * d4;
    const float4* src = g == 0 ? a : (g == 1 ? b : c);
    out[i] = __ldg(src + r * d4 + k);
  }
}
int stack_cols3(cudaStream_t st, long rows, int d, const float* a, const float* b, const float* c, float* out) {
  ScopedTimer tm(st, FRN_TIME_ROWS);
  KernelTimer kt(st, "stack_cols3_kernel", 2.0 * 4 * rows * 3 * d);
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (d % 4 == 0 && al(a) && al(b) && al(c) && al(out)) {
    stack_cols3_v4_kernel<<<grid_for(rows * 3 * (d / 4)), 256, 0, st>>>(
        rows, d / 4, (const float4*)a, (const float4*)b, (const float4*)c, (float4*)out);
  } else {
    stack_cols3_kernel<<<grid_for(rows * 3 * d), 256, 0, st>>>(rows, d, a, b, c, out);
  }
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int gate_backward(cudaStream_t st, long count, const float* r0, const float* e0, const float* dr1,
                  const float* de1, float* dr0, float* de0) {
  ScopedTimer tm(st, FRN_TIME_ROWS);
  KernelTimer kt(st, "gate_backward_kernel", 6.0 * 4 * count);      // r0, e0, dr1, de1 read; dr0, de0 written
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (count % 4 == 0 && al(r0) && al(e0) && al(dr1) && al(de1) && al(dr0) && al(de0)) {
    gate_backward_kernel4<<<grid_for(count / 4), 256, 0, st>>>(
        count / 4, (const float4*)r0, (const float4*)e0, (const float4*)dr1, (const float4*)de1,
        (float4*)dr0, (float4*)de0);
  } else {
    gate_backward_kernel<<<grid_for(count), 256, 0, st>>>(count, r0, e0, dr1, de1, dr0, de0);
  }
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int transpose_batched(cudaStream_t st, int B, int rows, int cols, const float* src, float* dst) {
  dim3 grid(cdiv(cols, 32), cdiv(rows, 32), B), block(32, 8);
  transpose_kernel<<<grid, block, 0, st>>>(rows, cols, src, dst);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int scale_rows(cudaStream_t st, int B, int C, int K, const float* w, const float* beta, float* scaled) {
  long total = (long)B * C * K;
  scale_rows_kernel<<<grid_for(total), 256, 0, st>>>(total, C, K, w, beta, scaled);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int outer_accumulate(cudaStream_t st, int B, int C, int K, const float* alpha, const float* ybar, float scale,
                     float* dW) {
  outer_accumulate_kernel<<<grid_for((long)C * K), 256, 0, st>>>(B, C, K, alpha, ybar, scale, dW);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int center_cols(cudaStream_t st, int B, int N, int d, float* x, float* part) {
  if (d > 256) return FRN_ERR_UNSUPPORTED;
  const int G = center_cols_parts(N), rpb = cdiv(N, G);
  ScopedTimer tm(st, FRN_TIME_ROWS);
  col_partial_sum_kernel<<<dim3(G, B), 256, 0, st>>>(N, d, rpb, x, part);
  FRN_LAUNCH_CHECK();
  col_center_kernel<<<dim3(G, B), 256, 0, st>>>(N, d, G, rpb, x, part);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int fill_zero(cudaStream_t st, float* p, size_t count) {
  if (count == 0) return FRN_OK;
  fill_zero_kernel<<<grid_for((long)count), 256, 0, st>>>(p, count);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int cast_bf16_to_f32(cudaStream_t st, size_t count, const uint16_t* src, float* dst) {
  if (count == 0) return FRN_OK;
  ScopedTimer tm(st, FRN_TIME_ROWS);
  const int vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
  cast_bf16_to_f32_kernel<<<grid_for((long)(count / 8 + 1)), 256, 0, st>>>(count, src, dst, vec);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int scale_sum_batch(cudaStream_t st, int B, int C, const float* alpha, float scale, float* out) {
  scale_sum_batch_kernel<<<cdiv(C, 256), 256, 0, st>>>(B, C, alpha, scale, out);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

// ---- REFusion + FPN lateral (row f1): the AdaIN affine folded into the lateral weights -------------------------------
// z[b][c][n] = sc[b][c] * x[b][c][n] + tc[b][c] with sc = sd_t / sd_x, tc = mu_t - mu_x * sc (AdaIN.py:24-26); the two
// blocks' coefficients side by side: columns [0, C) from block A, [C, 2C) from block B.
__global__ void affine_coeffs_kernel(int B, int C, const float* __restrict__ mxa, const float* __restrict__ sxa,
                                     const float* __restrict__ mta, const float* __restrict__ sta,
                                     const float* __restrict__ mxb, const float* __restrict__ sxb,
                                     const float* __restrict__ mtb, const float* __restrict__ stb,
                                     float* __restrict__ sc, float* __restrict__ tc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C) return;
  const int b = i / C, ch = i - b * C;
  const float sa = sta[i] / sxa[i], sb = stb[i] / sxb[i];
  sc[(long)b * 2 * C + ch] = sa;      tc[(long)b * 2 * C + ch] = mta[i] - mxa[i] * sa;
  sc[(long)b * 2 * C + C + ch] = sb;  tc[(long)b * 2 * C + C + ch] = mtb[i] - mxb[i] * sb;
}
int affine_coeffs(cudaStream_t st, int B, int C, const float* mxa, const float* sxa, const float* mta, const float* sta,
                  const float* mxb, const float* sxb, const float* mtb, const float* stb, float* sc, float* tc) {
  affine_coeffs_kernel<<<cdiv((long)B * C, 256), 256, 0, st>>>(B, C, mxa, sxa, mta, sta, mxb, sxb, mtb, stb, sc, tc);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
// out[b][o][k] = w[o][k] * sc[b][k]
__global__ void scale_cols_kernel(long total, int M, int K, const float* __restrict__ w, const float* __restrict__ sc,
                                  float* __restrict__ out) {
  const long stride = (long)gridDim.x * blockDim.x, mk = (long)M * K;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    const long b = i / mk, r = i - b * mk;
    out[i] = w[r] * sc[b * K + r % K];
  }
}
int scale_cols(cudaStream_t st, int B, int M, int K, const float* w, const float* sc, float* out) {
  const long total = (long)B * M * K;
  scale_cols_kernel<<<grid_for(total), 256, 0, st>>>(total, M, K, w, sc, out);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
// out[b * M + m] = sum_n x[b][m][n]   (x contiguous (B, M, N))
int row_sums(cudaStream_t st, int B, int M, int N, const float* x, float* out) {
  ScopedTimer tm(st, FRN_TIME_ROWS);
  KernelTimer kt(st, "row_sum_kernel", 4.0 * B * N * M);
  row_sum_kernel<<<dim3(M, B), 256, 0, st>>>(M, N, x, (long)M * N, N, out);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
// dW[o][k] = sum_b sc[b][k] P[b][o][k] + tc[b][k] S[b][o];  db[o] = sum_b S[b][o]
__global__ void lateral_wgrad_combine_kernel(int B, int M, int K, const float* __restrict__ P,
                                             const float* __restrict__ S, const float* __restrict__ sc,
                                             const float* __restrict__ tc, float* __restrict__ dW,
                                             float* __restrict__ db) {
  const long total = (long)M * K, stride = (long)gridDim.x * blockDim.x;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += stride) {
    const int o = (int)(i / K), k = (int)(i - (long)o * K);
    float acc = 0.f;
    for (int b = 0; b < B; b++)
      acc += sc[(long)b * K + k] * P[(long)b * total + i] + tc[(long)b * K + k] * S[(long)b * M + o];
    dW[i] = acc;
    if (k == 0 && db) {
      float t = 0.f;
      for (int b = 0; b < B; b++) t += S[(long)b * M + o];
      db[o] = t;
    }
  }
}
int lateral_wgrad_combine(cudaStream_t st, int B, int M, int K, const float* P, const float* S, const float* sc,
                          const float* tc, float* dW, float* db) {
  lateral_wgrad_combine_kernel<<<grid_for((long)M * K), 256, 0, st>>>(B, M, K, P, S, sc, tc, dW, db);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int softmax_rows(cudaStream_t st, long rows, int N, float* S, float* lse) {
  softmax_rows_kernel<<<(unsigned)rows, 256, 0, st>>>(N, S, lse);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
int exp_sub_lse(cudaStream_t st, long rows, int N, float* S, const float* lse) {
  exp_sub_lse_kernel<<<(unsigned)rows, 256, 0, st>>>(N, S, lse);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}
int ds_from_p(cudaStream_t st, long rows, int N, const float* P, float* dP, const float* delta) {
  ds_from_p_kernel<<<(unsigned)rows, 256, 0, st>>>(N, P, dP, delta);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // namespace frn

#pragma GCC visibility push(default)
extern "C" {
void frn_timing_enable(int on) { frn::g_timing_on.store(on ? 1 : 0); }
long frn_debug_gemm_log(char* buf, long cap) { return frn::shape_log_dump(buf, cap); }
long frn_timing_kernels(char* buf, long cap) { return frn::kernel_log_dump(buf, cap); }
int frn_timing_read(int cls, double* ms, long* launches) {
  if (cls < 0 || cls >= FRN_TIME_CLASSES || !ms || !launches) return FRN_ERR_BAD_SHAPE;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;
  {
    std::lock_guard<std::mutex> lk(frn::g_timing_mu);
    ev.swap(frn::g_events[cls]);
  }
  double total = 0.0;
  for (auto& p : ev) {
    float t = 0.f;
    cudaEventSynchronize(p.second);
    if (cudaEventElapsedTime(&t, p.first, p.second) == cudaSuccess) total += t;
    cudaEventDestroy(p.first);
    cudaEventDestroy(p.second);
  }
  *ms = total;
  *launches = (long)ev.size();
  return FRN_OK;
}
}
#pragma GCC visibility pop
