// Shared helpers for the FRN fusion-block kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include "../../include/frn_b200.h"

namespace frn {

extern std::atomic<unsigned long long> g_launches;

#define FRN_LAUNCH_CHECK()                                   \
  do {                                                       \
    frn::g_launches.fetch_add(1, std::memory_order_relaxed); \
    cudaError_t e__ = cudaGetLastError();                    \
    if (e__ != cudaSuccess) return (int)e__;                 \
  } while (0)

#define FRN_TRY(expr)            \
  do {                           \
    int rc__ = (expr);           \
    if (rc__ != FRN_OK) return rc__; \
  } while (0)

// Optional per-class kernel timing (include/frn_b200.h "measurement hooks").
struct ScopedTimer {
  cudaStream_t st; int cls; cudaEvent_t e0; bool on;
  ScopedTimer(cudaStream_t s, int c);
  ~ScopedTimer();
};

bool timing_enabled();    // measurement hooks on (include/frn_b200.h)
long shape_log_dump(char* buf, long cap);
long kernel_log_dump(char* buf, long cap);

// Per-kernel timing of the HBM-bound kernels (frn_timing_kernels): `bytes` = ALGORITHMIC bytes of the launch, i.e.
// every input element read once and every output element written once.
struct KernelTimer {
  cudaStream_t st; cudaEvent_t e0; bool on; const char* name; double bytes;
  KernelTimer(cudaStream_t s, const char* name_, double bytes_);
  ~KernelTimer();
};
long shape_log_dump(char* buf, long cap);

// Debug: per-shape GEMM times (FRN_GEMM_LOG=1 and the measurement hooks on); read with frn_debug_gemm_log.
struct ShapeTimer {
  cudaStream_t st; cudaEvent_t e0; bool on; char key[96];
  ShapeTimer(cudaStream_t s, const char* engine, int M, int N, int K, int batch, int flags);
  ~ShapeTimer();
};

// NVTX range over a host-side stage: shows up as "frn.<stage>" in Nsight Systems / Compute;
// costs a few ns when no tool is attached (nvtx3 is header-only and resolves its injection library lazily).
struct NvtxRange {
  explicit NvtxRange(const char* name);
  ~NvtxRange();
};

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device) instead of per launch.
int set_smem_once(const void* kern, int bytes);

// Side streams for independent work inside one call (block A / block B of REFusion, dQ / dK,dV of the attention
// backward).  They belong to the CALLER's stream: one set per (device, caller stream), created on first use and kept
// for the life of the process, so concurrent callers on different streams never share one, and a call that is being
// captured into a CUDA graph pulls its side streams into the same capture (fork / join are event edges only).
struct Fork {
  cudaStream_t main, side;     // side == nullptr: run everything on `main` (measurement hooks on, or no stream)
  cudaEvent_t ev_fork, ev_join;
};
// slot: 0 = attention backward (dQ), 1 = REFusion block B.  Returns a Fork whose side stream waits for the work
// enqueued on `st` so far.
Fork fork_stream(cudaStream_t st, int slot);
// Make `f.main` wait for everything enqueued on the side stream.
void join_stream(const Fork& f);

static inline int cdiv(long a, long b) { return (int)((a + b - 1) / b); }
static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024; result valid in every thread.
// `red` must hold 33 floats.
__device__ __forceinline__ float block_sum(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    float t = lane < nw ? red[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

// Bump allocator over a caller-provided buffer.
struct Arena {
  char* base;
  size_t off, cap, peak;
  bool ok;
  bool dry;   // only count bytes
  Arena(void* p, size_t bytes, bool dry_ = false)
      : base((char*)p), off(0), cap(bytes), peak(0), ok(true), dry(dry_) {}
  template <typename T>
  T* take(size_t count) {
    size_t bytes = align_up(count * sizeof(T));
    size_t at = off;
    off += bytes;
    if (off > peak) peak = off;
    if (dry) return nullptr;
    if (off > cap) { ok = false; return nullptr; }
    return (T*)(base + at);
  }
  void release_to(size_t mark) { off = mark; }
};

// ---- GEMM (gemm.cu) -------------------------------------------------------
// C[b][m][n] (+)= alpha * sum_k A[b][m][k] * B[b][k][n] + bias_m[m]
// with arbitrary element strides.  reduce_batch: sums the product over the
// batch into one C (weight gradients); then sCb is ignored.
struct GemmDesc {
  const float* A; const float* B; float* C;
  int M, N, K, batch;
  long sAm, sAk, sAb;
  long sBk, sBn, sBb;
  long sCm, sCn, sCb;
  const float* bias_m;   // length M (+ b*bias_bs) or nullptr
  const float* bias_n;   // length N (+ b*bias_bs) or nullptr
  long bias_bs;          // batch stride of bias_m / bias_n (0: shared)
  float alpha;
  int accumulate;        // C += result
  int reduce_batch;
  int fast;              // tcgen05 engine only: single-pass TF32 (hi * hi), ~2^-11 relative; bf16-tolerance inference.
                         // 1: operands are rounded to tf32 by the staging warps of the first kernel; 2: the operands
                         // already hold tf32-exact values (their producers rounded them): v2 kernel, no staging at all
  // tcgen05 v2 kernel only: drop the lo*lo product (3 tensor-core products instead of 4).  lo*lo is 2^-22 of the
  // product on average (one-sided); it only matters when an operand is near-constant AND the result is normalised
  // by its own variance afterwards (conv0 of a blanked RGB map, model.py:279-280) or exponentiated (attention
  // scores).  fusion.cu's desc() sets it for every other product of the block; the C ABI's stand-alone GEMM and
  // conv entry points keep all four.
  int no_lolo;
  // tcgen05 engine only, plain stores (no reduce_batch / accumulate): rows [g * group_m, (g + 1) * group_m) of the
  // product go to the g-th of up to three output matrices (C, C1, C2; same strides) -- one GEMM for the three 1x1
  // projections that read the same feature map.  group_m == 0: a single output.
  float* C1; float* C2;
  int group_m;
  // tcgen05 engine only, with reduce_batch: K is split but the batch is NOT summed -- C[b] (batch stride sCb) =
  // A[b] B[b] with the K range of every batch entry cut into per_batch_split pieces that run as separate CTAs
  // (long-K, tiny-output products such as the d x d scatter matrix of Y: one CTA per sample would leave the GPU
  // empty).  0: classic reduce_batch.
  int per_batch_split;
};
enum { GEMM_EPI_STORE = 0, GEMM_EPI_ROWSUMSQ = 1, GEMM_EPI_GROUPS = 2 /* store with row-grouped outputs */ };

// Engine: 0 = FFMA (gemm.cu), 1 = tcgen05 3xTF32 (tc_gemm.cu).  Default from FRN_GEMM_ENGINE
// ("ffma" / "tc"), see gemm_engine().
__global__ void splitk_reduce_kernel(const float* __restrict__ part, int nsplit, int M, int N,
                                     float* __restrict__ C, long sCm, long sCn, int accumulate);
int gemm_engine();
void set_gemm_engine(int e);
size_t tc_gemm_workspace_bytes(const GemmDesc& d);
int tc_gemm(cudaStream_t st, const GemmDesc& d, void* ws, size_t ws_bytes);
int tc_gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial);
int attn_tc_trace_read(unsigned long long* out16);
int tc_gemm_ts_read(long long* out);                  // stage hand-over timestamps, only with -DFRN_GEMM_TS
int tc_gemm_trace_read(unsigned long long* out16);   // role cycle counters, only with -DFRN_GEMM_TRACE
size_t ffma_gemm_workspace_bytes(const GemmDesc& d);
int ffma_gemm(cudaStream_t st, const GemmDesc& d, void* ws, size_t ws_bytes);
int ffma_gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial);
size_t gemm_workspace_bytes(const GemmDesc& d);
// ws: split-K partials (only used when reduce_batch); may be null if bytes == 0.
int gemm(cudaStream_t st, const GemmDesc& d, void* ws, size_t ws_bytes);
// out_partial[(b*M + m) * nparts + part] = sum over the tile's columns of (A*B)^2; returns nparts.
int gemm_rowsumsq_parts(int N);
int gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial);

// ---- attention (attention.cu) ----------------------------------------------
// Qc, Kc: (B, d, N) channel-major.  Vp: (B, N, d) position-major.
// Y: (B, N, d) row-major; lse: (B, N).
int attn_forward(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Kc,
                 const float* Vp, float* Y, float* lse);
// tcgen05 engine (attention_tc.cu).  Operands are "packed split images": tf32 hi / lo planes in the
// SWIZZLE_128B tile layout, zero padded to attn_tc_np(N) positions and 32 / 64 channels; each plane
// holds attn_tc_plane_floats(B, N, d) floats.  A null pair of planes is not written by the packer.
struct SplitImage { float *pos_hi, *pos_lo, *chan_hi, *chan_lo; };
int attn_tc_np(int N);
size_t attn_tc_plane_floats(int B, int N, int d);
// X element (c, n) of sample b at X[b*bs + c*cs + n*ns]; sub (optional, (B, d)) is subtracted first.
// The tcgen05 engine packs the KEY operand centred over positions (sub = its per-channel mean):
// softmax_j(q_i.k_j) and dQ_i = sum_j dS_ij k_j are invariant under k_j -> k_j - kbar (rows of dS sum
// to zero), and without the common component the cancellation in dQ no longer amplifies rounding.
int attn_tc_pack(cudaStream_t st, int B, int d, int N, const float* X, const float* sub, long bs, long cs,
                 long ns, const SplitImage& img);
// needs Q.pos, K.pos, V.chan.  fast != 0: single-pass TF32 (bf16-tolerance inference); only the hi planes
// are read and the lo planes may be null.
int attn_forward_tc(cudaStream_t st, int B, int N, int d, const SplitImage& Q, const SplitImage& K,
                    const SplitImage& V, float* Y, float* lse, int fast = 0);
// lse_pad, delta_pad: (B, attn_tc_np(N)); delta = rowsum(dY * Y)
int attn_bwd_prep_tc(cudaStream_t st, int B, int N, int d, const float* dYp, const float* Y, const float* lse,
                     float* lse_pad, float* delta_pad);
// needs Q.pos/chan, K.pos/chan, V.pos, dY.pos/chan; writes dQp, dKp, dVp (B, N, d), no atomics
int attn_backward_tc(cudaStream_t st, int B, int N, int d, const SplitImage& Q, const SplitImage& K,
                     const SplitImage& V, const SplitImage& dY, const float* lse_pad, const float* delta_pad,
                     float* dQp, float* dKp, float* dVp);
// 0 = FFMA kernels (attention.cu), 1 = tcgen05 kernels; default from FRN_ATTN_ENGINE ("ffma" / "tc").
int attn_engine();
void set_attn_engine(int e);
// Backward needs both orientations: *c = (B, d, N), *p = (B, N, d).
// dQp must be zero-initialised by the caller (accumulated with atomics).
int attn_backward(cudaStream_t st, int B, int N, int d, const float* Qc, const float* Qp,
                  const float* Kc, const float* Kp, const float* Vc, const float* dYc,
                  const float* dYp, const float* lse, const float* delta, float* dQp,
                  float* dKp, float* dVp);

// ---- elementwise / reductions (adain.cu) ------------------------------------
// mean/std over rows of length N: x is (rows, N) contiguous.
int row_mean_std(cudaStream_t st, long rows, int N, const float* x, float eps, float* mean,
                 float* sd);
// r1 = r0 + r0*e0, e1 = e0 + r0*e0 fused with the row statistics of r1 and e1.
// scratch (gate_stats_scratch_bytes, may be null): enables the segmented kernel for long rows
int gate_stats(cudaStream_t st, long rows, int N, const float* r0, const float* e0, float* r1,
               float* e1, float eps, float* mu_r, float* sd_r, float* mu_e, float* sd_e, void* scratch = nullptr);
size_t gate_stats_scratch_bytes(long rows, int N);
// out[b][c][n] = (x - mu_x)/sd_x * sd_t + mu_t   (out has batch stride ob)
int adain_apply(cudaStream_t st, int B, int C, int N, const float* x, const float* mu_x,
                const float* sd_x, const float* mu_t, const float* sd_t, float* out, long ob);
// the same with a bf16 output (inference with bf16 activations); ob in elements
int adain_apply_bf16(cudaStream_t st, int B, int C, int N, const float* x, const float* mu_x,
                     const float* sd_x, const float* mu_t, const float* sd_t, uint16_t* out, long ob);
// bf16-inference fast path (adain.cu): padded rows (ld = round_up(N, 4)), tf32-exact values
int cast_pad_bf16(cudaStream_t st, long rows, int N, int ld, const uint16_t* src, float* dst);
int round_tf32_copy(cudaStream_t st, long n, const float* src, float* dst);
int gate_stats_pad(cudaStream_t st, long rows, int N, int ld, const float* r0, const float* e0, float* r1, float* e1,
                   float eps, float* mu_r, float* sd_r, float* mu_e, float* sd_e);
int center_rows_pad(cudaStream_t st, long rows, int N, int ld, const float* y, const float* mean, float* yc);
int adain_apply_bf16_pad(cudaStream_t st, int B, int C, int N, int ld, const float* x, const float* mu_x,
                         const float* sd_x, const float* mu_t, const float* sd_t, uint16_t* out, long ob);
// dst[i] = float(src[i]) for bf16 src
int cast_bf16_to_f32(cudaStream_t st, size_t count, const uint16_t* src, float* dst);
// g1 = sum_n g, g2 = sum_n g * xhat per (b, c) row; g has batch stride gb.
int adain_bwd_reduce(cudaStream_t st, int B, int C, int N, const float* g, long gb,
                     const float* x, const float* mu_x, const float* sd_x, float* g1, float* g2);
// dx (+)= (sd_t/sd_x) * (g - g1/N - xhat*g2/(N-1))
int adain_bwd_content(cudaStream_t st, int B, int C, int N, const float* g, long gb,
                      const float* x, const float* mu_x, const float* sd_x, const float* sd_t,
                      const float* g1, const float* g2, float* dx, int accumulate);
// dstyle = g1/N + g2*(t - mu_t)/((N-1)*sd_t)
int adain_bwd_style(cudaStream_t st, int B, int C, int N, const float* t, const float* mu_t,
                    const float* sd_t, const float* g1, const float* g2, float* dstyle);
// alpha = g1/N, beta = g2/((N-1)*sd_t)
int style_coeffs(cudaStream_t st, long rows, int N, const float* g1, const float* g2,
                 const float* sd_t, float* alpha, float* beta);
// sd_t = sqrt(sum(parts)/(N-1) + eps)
int finish_style_std(cudaStream_t st, long rows, int nparts, int N, const float* parts, float eps,
                     float* sd_t);
// yc[r][n] = y[r][n] - mean[r]
int center_rows(cudaStream_t st, long rows, int N, const float* y, const float* mean, float* yc);
// out[r] = sum_n x[r][n]*y[r][n]   (rows of length N)
int row_dot(cudaStream_t st, long rows, int N, const float* x, const float* y, float* out);
// out[m] = sum_b sum_n x[b*sb + m*sm + n*sn]
// scratch (optional, batch_row_sum_scratch floats) enables the coalesced path for position-major x
int batch_row_sum(cudaStream_t st, int B, int M, int N, const float* x, long sb, long sm, long sn,
                  float* out, float* scratch = nullptr);
size_t batch_row_sum_scratch(int B, int M, int N);
// gate backward: dr0 = dr1*(1+e0) + de1*e0 ; de0 = de1*(1+r0) + dr1*r0 (in place allowed)
int stack3(cudaStream_t st, long n, const float* a, const float* b, const float* c, float* out);
int unstack3(cudaStream_t st, long n, const float* in, float* a, float* b, float* c);
int stack_cols3(cudaStream_t st, long rows, int d, const float* a, const float* b, const float* c, float* out);
int gate_backward(cudaStream_t st, long count, const float* r0, const float* e0, const float* dr1,
                  const float* de1, float* dr0, float* de0);
// dst (B, cols, rows) = transpose of src (B, rows, cols)
int transpose_batched(cudaStream_t st, int B, int rows, int cols, const float* src, float* dst);
// scaled[b][c][k] = beta[b][c] * w[c][k]
int scale_rows(cudaStream_t st, int B, int C, int K, const float* w, const float* beta,
               float* scaled);
// dst[b][k][n] = u[b][k] + src[b][k][n]   (in place)
int add_row_bias(cudaStream_t st, long rows, int N, const float* u, float* x);
// dW[c][k] (+)= sum_b alpha[b][c] * N * ybar[b][k]
int outer_accumulate(cudaStream_t st, int B, int C, int K, const float* alpha, const float* ybar,
                     float scale, float* dW);
// materialised attention helpers (rows of length N, one CTA per row)
int softmax_rows(cudaStream_t st, long rows, int N, float* S, float* lse);        // in place
int exp_sub_lse(cudaStream_t st, long rows, int N, float* S, const float* lse);   // in place
int ds_from_p(cudaStream_t st, long rows, int N, const float* P, float* dP, const float* delta);
int fill_zero(cudaStream_t st, float* p, size_t count);
// x (B, N, d) -= its mean over the N positions, per (b, channel); part: B * center_cols_parts(N) * d floats.
int center_cols_parts(int N);
int center_cols(cudaStream_t st, int B, int N, int d, float* x, float* part);
int scale_sum_batch(cudaStream_t st, int B, int C, const float* alpha, float scale, float* out);
// REFusion + lateral conv (row f1), see misc.cu
int affine_coeffs(cudaStream_t st, int B, int C, const float* mxa, const float* sxa, const float* mta, const float* sta,
                  const float* mxb, const float* sxb, const float* mtb, const float* stb, float* sc, float* tc);
int scale_cols(cudaStream_t st, int B, int M, int K, const float* w, const float* sc, float* out);
int row_sums(cudaStream_t st, int B, int M, int N, const float* x, float* out);
int lateral_wgrad_combine(cudaStream_t st, int B, int M, int K, const float* P, const float* S, const float* sc,
                          const float* tc, float* dW, float* db);

}  // namespace frn
