/*
 * frn_b200.h -- C ABI of the B200-native FRN fusion block.
 *
 * The reference (JinkaiLiu/FRN) is pure Python and has no FFI of its own; the
 * entry points below are what a binding for its fusion path would call.  Each
 * one names the reference interface it replaces (paths relative to the
 * reference repository):
 *
 *   frn_mean_std            AdaIN_block.calc_mean_std      retinanet/AdaIN.py:7-15
 *   frn_adain_forward/...   AdaIN_block.forward            retinanet/AdaIN.py:17-26
 *   frn_cab_forward/...     CrossAttentionBlock.forward    retinanet/Attention.py:21-36
 *   frn_refusion_forward/.. REFusion.forward               retinanet/fusion.py:20-36
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer on the current CUDA device unless the
 *     parameter is documented otherwise.  Feature maps are NCHW-contiguous
 *     (the layout retinanet/model.py:305-308 hands over), N = H*W.
 *   - `stream` is a cudaStream_t passed as void*.  Calls only enqueue work; they never synchronise and never
 *     allocate device memory, so they may be captured into a CUDA graph.  Re-entrant: nn.DataParallel calls one
 *     thread per device, and concurrent callers on different streams never share anything.  What the library keeps
 *     between calls, all of it created lazily, guarded by a mutex and never freed: (1) per (device, caller stream)
 *     the side streams and fork / join events on which independent halves of a call run (block A / block B of
 *     REFusion, dQ / dK,dV of the attention backward) -- they always rejoin the caller's stream before the call
 *     returns; (2) the set of kernels whose shared-memory limit has been raised; (3) switches read ONCE from the
 *     environment at first use (FRN_GEMM_TMA, FRN_GEMM_V2, FRN_GEMM_LOLO, FRN_NO_PROJ3, FRN_NO_F4, FRN_NO_FORK, FRN_INFER_V1: debug
 *     and comparison only); (4) the launch counter and, when enabled, the measurement hooks at the end of this file.
 *     The engine switches (frn_set_*_engine) only have an effect in a -DFRN_DEBUG_ENGINES build.
 *   - The caller owns every buffer, including `saved` (forward state kept for
 *     backward) and `ws` (scratch, free to reuse after the call's work is
 *     complete on `stream`).  Sizes come from the *_bytes() queries.
 *   - Return value: FRN_OK (0) or a negative FRN_ERR_* code; CUDA launch
 *     errors are returned as the positive cudaError_t value.  Nothing throws.
 *   - dtype: FRN_F32 only for gradients and parameters.  Activations may be
 *     FRN_F32 or FRN_BF16 where a `dtype` argument exists.
 *   - Conv weights are the reference's (out, in, 1, 1) tensors viewed as
 *     row-major (out, in) matrices; biases are (out).
 */
#ifndef FRN_B200_H_
#define FRN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FRN_B200_ABI_VERSION 2

enum {
  FRN_OK = 0,
  FRN_ERR_BAD_SHAPE = -1,      /* non-positive size, C % 8 != 0, N < 2 ...        */
  FRN_ERR_NULL_POINTER = -2,
  FRN_ERR_WORKSPACE = -3,      /* ws_bytes / saved_bytes smaller than required    */
  FRN_ERR_UNSUPPORTED = -4,    /* C/8 > 256, unknown dtype ...                    */
  FRN_ERR_NO_DEVICE = -5,
  FRN_ERR_STATIC_SCRATCH = -6  /* frn_adain_backward_static: B*C > FRN_ADAIN_STATIC_MAX_ROWS */
};

enum { FRN_F32 = 0, FRN_BF16 = 1 };

/* Parameters of one CrossAttentionBlock (retinanet/Attention.py:11-15). */
typedef struct frn_cab_params {
  const float* g_w;     const float* g_b;      /* (C/8, C), (C/8) */
  const float* theta_w; const float* theta_b;  /* (C/8, C), (C/8) */
  const float* phi_w;   const float* phi_b;    /* (C/8, C), (C/8) */
  const float* W_w;     const float* W_b;      /* (C, C/8), (C)   */
} frn_cab_params;

typedef struct frn_cab_grads {
  float* g_w;     float* g_b;
  float* theta_w; float* theta_b;
  float* phi_w;   float* phi_b;
  float* W_w;     float* W_b;
} frn_cab_grads;

/* Parameters of one REFusion (retinanet/fusion.py:12-17). */
typedef struct frn_refusion_params {
  const float* conv0_rgb_w; const float* conv0_rgb_b;  /* (C, C), (C) */
  const float* conv0_evt_w; const float* conv0_evt_b;  /* (C, C), (C) */
  frn_cab_params rgb_ca;     /* rgb_cross_attention   */
  frn_cab_params event_ca;   /* event_cross_attention */
} frn_refusion_params;

typedef struct frn_refusion_grads {
  float* conv0_rgb_w; float* conv0_rgb_b;
  float* conv0_evt_w; float* conv0_evt_b;
  frn_cab_grads rgb_ca;
  frn_cab_grads event_ca;
} frn_refusion_grads;

int frn_abi_version(void);
const char* frn_error_string(int code);
/* Comma-separated list of the kernels compiled into the library (for audits). */
const char* frn_kernel_inventory(void);

/* ---- AdaIN (retinanet/AdaIN.py) ------------------------------------------ */

/* calc_mean_std: feat (B, C, N) -> mean, std each (B*C).  std = sqrt(unbiased
 * variance + eps).  AdaIN.py:7-15. */
int frn_mean_std(void* stream, int B, int C, int N, const float* feat, float eps,
                 float* mean, float* std_out);

/* AdaIN_block.forward(content, style) -> out (B, C, N).  content is (B, C, N), style is
 * (B, C, N_style): the reference only asserts that the two agree in (B, C) (AdaIN.py:20) and takes
 * the style statistics over the style map's own H*W, so N_style may differ from N.  stats
 * (4*B*C floats: mu_c, sd_c, mu_s, sd_s) is kept for backward.  AdaIN.py:17-26. */
int frn_adain_forward(void* stream, int B, int C, int N, int N_style, const float* content,
                      const float* style, float* out, float* stats);
/* Gradients of AdaIN_block.forward wrt content (B, C, N) and style (B, C, N_style).
 * ws: frn_adain_workspace_bytes. */
int frn_adain_backward(void* stream, int B, int C, int N, int N_style, const float* content,
                       const float* style, const float* stats, const float* dout,
                       float* dcontent, float* dstyle, void* ws, size_t ws_bytes);
size_t frn_adain_workspace_bytes(int B, int C, int N);

/* frn_adain_backward without a workspace argument, for callers that cannot call cudaMalloc (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * The B*C-float reductions g1, g2 that frn_adain_backward keeps in `ws` live in a
 * file-scope __device__ array of 2 * FRN_ADAIN_STATIC_MAX_ROWS floats (512 KiB, held by every GPU that loads the
 * library until exit); the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and
 * enqueues the same three kernels, so dcontent / dstyle are bit-identical to frn_adain_backward's.
 *   - B*C > FRN_ADAIN_STATIC_MAX_ROWS returns FRN_ERR_STATIC_SCRATCH and enqueues nothing.
 *   - The device that owns `stream` and the buffers must be the current device.
 *   - All calls on a device share the one array (it keeps the last call's contents; every call overwrites the rows
 *     it reads first, nothing needs clearing).  Calls that overlap on the GPU -- different streams, threads or
 *     graph branches -- RACE: serialise them on one stream, or use frn_adain_backward with a workspace each. */
#define FRN_ADAIN_STATIC_MAX_ROWS 65536
long frn_adain_static_max_rows(void);
int frn_adain_backward_static(void* stream, int B, int C, int N, int N_style, const float* content,
                              const float* style, const float* stats, const float* dout,
                              float* dcontent, float* dstyle);

/* ---- CrossAttentionBlock (retinanet/Attention.py) ------------------------- */

size_t frn_cab_saved_bytes(int B, int C, int N);
size_t frn_cab_workspace_bytes(int B, int C, int N);

/* CrossAttentionBlock.forward([x0, x1]) -> out.  out has batch stride
 * out_batch_stride elements (C*N for a stand-alone block, 2*C*N when writing
 * one half of REFusion's concatenated output).  saved may be NULL for
 * inference (then nothing is kept).  Attention.py:21-36. */
int frn_cab_forward(void* stream, int B, int C, int N, const float* x0, const float* x1,
                    const frn_cab_params* p, float* out, long out_batch_stride,
                    void* saved, size_t saved_bytes, void* ws, size_t ws_bytes);

/* dout has batch stride dout_batch_stride.  dx0 / dx1 are OVERWRITTEN when
 * accumulate == 0 and added to when accumulate != 0.  Parameter gradients are
 * overwritten (summed over the batch). */
int frn_cab_backward(void* stream, int B, int C, int N, const float* x0, const float* x1,
                     const frn_cab_params* p, const void* saved, const float* dout,
                     long dout_batch_stride, float* dx0, float* dx1, int accumulate,
                     const frn_cab_grads* grads, void* ws, size_t ws_bytes);

/* ---- REFusion (retinanet/fusion.py) -------------------------------------- */

size_t frn_refusion_saved_bytes(int B, int C, int N);
size_t frn_refusion_workspace_bytes(int B, int C, int N);

/* REFusion.forward(rgb=a, evt=e) -> out (B, 2C, N); channels [0, C) come from
 * rgb_cross_attention([rgb1, evt1]), [C, 2C) from event_cross_attention
 * ([evt1, rgb1]).  fusion.py:20-36.  saved may be NULL for inference. */
int frn_refusion_forward(void* stream, int B, int C, int N, const float* a, const float* e,
                         const frn_refusion_params* p, float* out,
                         void* saved, size_t saved_bytes, void* ws, size_t ws_bytes);

int frn_refusion_backward(void* stream, int B, int C, int N, const float* a, const float* e,
                          const frn_refusion_params* p, const void* saved, const float* dout,
                          float* da, float* de, const frn_refusion_grads* grads,
                          void* ws, size_t ws_bytes);

/* Number of kernel launches issued by this library since load, on all threads
 * (monotonic counter; bench.py reports the delta over the timed region). */
unsigned long long frn_launch_count(void);

/* ---- strided batched fp32 GEMM (stage-level entry point for unit parity) -------
 * C[b][m][n] (+)= sum_k A[b][m][k] * B[b][k][n] + bias_m[m] + bias_n[n] with element strides
 * (one of sAm/sAk and one of sBk/sBn must be 1).  reduce_batch != 0 sums the product over the
 * batch into a single C (weight gradients).  engine: 0 = FFMA kernel, 1 = tcgen05 3xTF32
 * kernel, -1 = the library default (frn_set_gemm_engine / env FRN_GEMM_ENGINE=ffma|tc).
 * Every 1x1 conv of the block and its gradients go through this routine
 * (retinanet/fusion.py:23-24, retinanet/Attention.py:24-27,34). */
typedef struct frn_gemm_desc {
  const float* A; const float* B; float* C;
  int M, N, K, batch;
  long sAm, sAk, sAb, sBk, sBn, sBb, sCm, sCn, sCb;
  const float* bias_m; const float* bias_n;
  int accumulate, reduce_batch;
} frn_gemm_desc;
size_t frn_gemm_workspace_bytes(const frn_gemm_desc* d);
int frn_gemm_f32(void* stream, int engine, const frn_gemm_desc* d, void* ws, size_t ws_bytes);
void frn_set_gemm_engine(int engine);
int frn_get_gemm_engine(void);
/* Attention core engine: 0 = FFMA streaming kernels, 1 = tcgen05/TMEM kernels (env FRN_ATTN_ENGINE). */
void frn_set_attn_engine(int engine);
int frn_get_attn_engine(void);

/* ---- attention core (stage-level entry point for unit parity) -------------------------------
 * Y = softmax(Q^T K, dim = keys) V^T of retinanet/Attention.py:29-32 (no 1/sqrt(d) scale), never
 * materialising the N x N matrix for d <= 64.  Qc = phi(x1), Kc = theta(x1), Vc = g(x0), each
 * (B, d, N) channel-major as the 1x1 convs write them; Y, dY, dQ, dK, dV are (B, N, d) row-major
 * (so the reference's y.view(B, d, H, W) is a reinterpretation of Y); lse is (B, N), the row
 * log-sum-exp of the logits as the engine formed them (the tcgen05 engine subtracts the per-channel
 * mean over positions from Kc first, which leaves the softmax and all gradients unchanged); pass the
 * forward's lse to the backward of the SAME engine.
 * engine: 0 = FFMA kernels, 1 = tcgen05 kernels, -1 = library default. */
size_t frn_attn_workspace_bytes(int B, int N, int d);
int frn_attn_forward(void* stream, int engine, int B, int N, int d, const float* Qc, const float* Kc,
                     const float* Vc, float* Y, float* lse, void* ws, size_t ws_bytes);
int frn_attn_backward(void* stream, int engine, int B, int N, int d, const float* Qc, const float* Kc,
                      const float* Vc, const float* Y, const float* lse, const float* dY, float* dQ,
                      float* dK, float* dV, void* ws, size_t ws_bytes);

/* ---- REFusion, inference (retinanet/fusion.py:20-36 under torch.no_grad(); BASELINE.json config 4) -----
 * Forward pass that keeps no state for a backward.  dtype = FRN_F32: a, e, out are float (identical results
 * to frn_refusion_forward with saved == NULL).  dtype = FRN_BF16: a, e (B, C, N) and out (B, 2C, N) are
 * bfloat16, parameters stay float; statistics, softmax and accumulation are fp32 and every tensor-core
 * product is single-pass TF32 (11 significant bits), i.e. the result is at least as accurate as the
 * reference under torch.autocast(bfloat16) (train_dsec_det_fast.py:241).  Needs the tcgen05 engines. */
size_t frn_refusion_infer_workspace_bytes(int dtype, int B, int C, int N);
int frn_refusion_infer(void* stream, int dtype, int B, int C, int N, const void* a, const void* e,
                       const frn_refusion_params* p, void* out, void* ws, size_t ws_bytes);

/* ---- 1x1 convolution (the FPN laterals that consume the fused map) -------------
 * y (B, Cout, N) = w (Cout, Cin) x (B, Cin, N) + bias  -- nn.Conv2d(Cin, Cout, kernel_size=1) of
 * retinanet/model.py:28,32,36,40 (P{5,4,3,2}_1) on NCHW fp32 tensors viewed as (B, C, N = H*W), run on the
 * same GEMM engine as the block's own projections (split TF32 on tcgen05 by default).
 * backward: dx (B, Cin, N) = w^T dy (skipped when dx is NULL), dw (Cout, Cin) = sum_b dy_b x_b^T,
 * db (Cout) = sum_b sum_n dy (skipped when db is NULL); all three are overwritten. */
size_t frn_conv1x1_workspace_bytes(int B, int Cin, int Cout, int N);
int frn_conv1x1_forward(void* stream, int B, int Cin, int Cout, int N, const float* x, const float* w,
                        const float* bias, float* y, void* ws, size_t ws_bytes);
int frn_conv1x1_backward(void* stream, int B, int Cin, int Cout, int N, const float* x, const float* w,
                         const float* dy, float* dx, float* dw, float* db, void* ws, size_t ws_bytes);

/* ---- REFusion fused with the lateral conv that consumes it --------------------------------
 * lat (B, Cout, N) = P_1(REFusion(a, e)) with P_1 = nn.Conv2d(2C, Cout, 1) of retinanet/model.py:28,32,36,40
 * (lat_w (Cout, 2C), lat_b (Cout) or NULL) applied to the concatenated map of fusion.py:34.  The AdaIN affine that
 * produces each half of that map is folded into per-sample scaled lateral weights, so the GEMM reads the gated maps
 * directly and the (B, 2C, N) map is neither written nor read; pass out != NULL to have it written as well (level 4:
 * P6 also reads C5, model.py:44).  saved (frn_refusion_saved_bytes) is required.
 * backward: dlat (B, Cout, N); dout_extra (B, 2C, N) or NULL = gradient that reached the 2C map from another
 * consumer; da, de, grads as frn_refusion_backward; dlat_w (Cout, 2C) and dlat_b (Cout, may be NULL) overwritten. */
size_t frn_refusion_lateral_workspace_bytes(int B, int C, int N, int Cout);
int frn_refusion_lateral_forward(void* stream, int B, int C, int N, int Cout, const float* a, const float* e,
                                 const frn_refusion_params* p, const float* lat_w, const float* lat_b, float* lat,
                                 float* out, void* saved, size_t saved_bytes, void* ws, size_t ws_bytes);
int frn_refusion_lateral_backward(void* stream, int B, int C, int N, int Cout, const float* a, const float* e,
                                  const frn_refusion_params* p, const float* lat_w, const void* saved, const float* dlat,
                                  const float* dout_extra, float* da, float* de, const frn_refusion_grads* grads,
                                  float* dlat_w, float* dlat_b, void* ws, size_t ws_bytes);

/* ---- anchors + focal loss ----------------------------------------------------
 * frn_anchors_generate replaces Anchors.forward (retinanet/anchors.py:25-45, generate_anchors :47-82, shift
 * :120-147): all boxes of an H x W image over the pyramid levels, (1, A, 4) fp32 (x1, y1, x2, y2), cells
 * row-major and the n_ratios * n_scales boxes of a cell adjacent, computed in double and rounded once --
 * bit-identical to the NumPy result.  pyramid_levels, strides, sizes (n_levels ints), ratios and scales are
 * HOST arrays; out is a device pointer to frn_anchors_count(...) * 4 floats.
 *
 * frn_focal_loss_forward replaces FocalLoss.forward (retinanet/losses.py:28-169, calc_iou :5-24):
 * classifications (B, A, K) in (0, 1), regressions (B, A, 4), anchors (A, 4), annotations (B, M, 5) =
 * (x1, y1, x2, y2, class) padded with class == -1 rows (0 <= M <= 1024); gamma = 2.  losses[0] = mean over the batch
 * of the classification loss, losses[1] = of the box loss (the two (1,) tensors the reference returns).
 * `saved` keeps the per-(sample, anchor) assignment as int32 (>= 0: matched annotation row, -1: background,
 * -2: ignored) followed (at the next 256-byte boundary) by B x 3 doubles {classification loss, box loss,
 * positives} per sample.  The IoU thresholds are evaluated with the reference's fp32 operation order, so the
 * assignment is bit-identical to the reference's.  backward: dlosses = 2 device floats (upstream gradients of
 * the two outputs); dclassifications and dregressions are overwritten.  No host synchronisation. */
long frn_anchors_count(int H, int W, int n_levels, const int* pyramid_levels, int boxes_per_cell);
int frn_anchors_generate(void* stream, int H, int W, int n_levels, const int* pyramid_levels, const int* strides,
                         const int* sizes, int n_ratios, const double* ratios, int n_scales, const double* scales,
                         float* out);
size_t frn_focal_loss_saved_bytes(int B, int A);
size_t frn_focal_loss_workspace_bytes(int B, int A);
int frn_focal_loss_forward(void* stream, int B, int A, int K, int M, const float* classifications,
                           const float* regressions, const float* anchors, const float* annotations, float alpha,
                           float* losses, void* saved, size_t saved_bytes, void* ws, size_t ws_bytes);
int frn_focal_loss_backward(void* stream, int B, int A, int K, int M, const float* classifications,
                            const float* regressions, const float* anchors, const float* annotations, float alpha,
                            const void* saved, const float* dlosses, float* dclassifications, float* dregressions);

/* ---- event voxelisation ---------------------------------------------------------
 * Replaces DSECDet.preprocess_events (retinanet/data/dsec_data.py:347-387) and normalize_event_data (:78-93).
 * x, y: n_events coordinates (xy_bytes = 2: uint16 as stored in the DSEC h5 files, 4: int32); t: int64
 * timestamps; p: uint8 polarity (> 0 -> +1, else -1).  Events with y >= H or x >= W are dropped; the first and
 * last remaining event define the time window, bin = trunc((t - t_first) / (t_last - t_first + 1e-6) * (C - 1))
 * clipped to [0, C-1], evaluated in double like NumPy.  voxel (C, H, W) fp32 is overwritten (zero when no event
 * is in range).  normalize: 0 none, 1 tanh(v / 5), 2 clamp to [-2, 2].  The counts are bit-identical to the
 * reference's sequential loop (sums of +-1 are exact in fp32).  ws: frn_voxelize_workspace_bytes() bytes. */
size_t frn_voxelize_workspace_bytes(void);
int frn_voxelize_events(void* stream, long n_events, int xy_bytes, const void* x, const void* y, const int64_t* t,
                        const uint8_t* p, int C, int H, int W, int normalize, float* voxel, void* ws,
                        size_t ws_bytes);
/* normalize_event_data (dsec_data.py:78-93) in place on `count` floats: mode 1 tanh(v / 5), 2 clamp to [-2, 2]. */
int frn_voxel_normalize(void* stream, long count, int mode, float* voxel);

/* ---- inference post-processing ---------------------------------------------------
 * Replaces retinanet/model.py:326-364: BBoxTransform (retinanet/utils.py:103-126, std 0.1/0.1/0.2/0.2) +
 * ClipBoxes (utils.py:135-145), then per class `scores > score_thresh`, torchvision.ops.nms(iou_thresh) and the
 * concatenation of the classes.  Batch 1 like the reference: classification (A, K), regression (A, 4),
 * anchors (A, 4).  If boxes_in is not NULL it is used instead of decoding (then regression / anchors may be NULL:
 * a drop-in for torchvision.ops.nms with K = 1).  Outputs hold up to `capacity` detections, class-major and by
 * decreasing score inside a class (score ties: lower anchor index first): out_scores, out_classes (int32),
 * out_boxes (x1, y1, x2, y2), out_index (anchor index; may be NULL); *out_count (device int) = number of
 * detections (may exceed capacity: the excess is not written).  The overlap test is torchvision's CPU kernel
 * arithmetic in the same order, so the kept set is identical for identical boxes.  No host synchronisation. */
size_t frn_detect_workspace_bytes(int A, int K);
int frn_decode_boxes(void* stream, int A, const float* anchors, const float* regression, int img_h, int img_w,
                     float* boxes);
int frn_detect_postprocess(void* stream, int A, int K, const float* classification, const float* regression,
                           const float* anchors, const float* boxes_in, int img_h, int img_w, float score_thresh,
                           float iou_thresh, int capacity, float* out_scores, int* out_classes, float* out_boxes,
                           int* out_index, int* out_count, void* ws, size_t ws_bytes);

/* ---- measurement hooks (bench.py / profiling only) --------------------------
 * When enabled, the library brackets its hot kernels with cudaEvents on the
 * launching stream.  Classes: 0 = streaming attention forward, 1 = streaming
 * attention backward, 2 = fp32 GEMM (all 1x1 projections and their gradients),
 * 3 = row kernels (statistics / AdaIN / gate).  frn_timing_read synchronises
 * the recorded events, returns their summed duration in milliseconds and the
 * number of launches, and clears the class.  Off by default; the only global
 * mutable state in the library besides the launch counter. */
enum { FRN_TIME_ATTN_FWD = 0, FRN_TIME_ATTN_BWD = 1, FRN_TIME_GEMM = 2, FRN_TIME_ROWS = 3,
       FRN_TIME_ATTN_DKV = 4,   /* the dK/dV kernel alone (also counted inside class 1) */
       FRN_TIME_CLASSES = 5 };
void frn_timing_enable(int on);
int frn_timing_read(int cls, double* ms, long* launches);
/* Per-kernel record of the HBM-bound kernels since the last call (hooks on): one text line per kernel,
 * "name | launches | ms | algorithmic bytes | launches_large | ms_large | bytes_large", where "large" restricts the
 * sums to launches whose algorithmic bytes exceed 250 MB (twice the L2).  Algorithmic = each input element read once, each
 * output element written once.  Returns the length of the full text; at most cap - 1 characters are copied. */
long frn_timing_kernels(char* buf, long cap);

#ifdef __cplusplus
}
#endif
#endif  /* FRN_B200_H_ */
