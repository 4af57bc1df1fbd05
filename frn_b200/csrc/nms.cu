// inference post-processing (retinanet/model.py:326-364).
//
//   reference: BBoxTransform (utils.py:103-126) + ClipBoxes (utils.py:135-145) as ~25 ATen launches, then a
//   Python loop over the classes: boolean-mask indexing, `if scores_over_thresh.sum() == 0` (a host sync per
//   class), torchvision.ops.nms, and torch.cat of the results.
//
// Here: one decode + clip kernel, one key-preparation kernel, a radix sort per class (CUB: library plumbing),
// ONE greedy-NMS kernel for all classes (a CTA per class walks the score-sorted candidates in chunks of 1024:
// suppression against the boxes already kept, a 1024 x 1024 overlap bit matrix in shared memory, a warp-level
// scan over it), and one gather kernel that concatenates the classes.  A single 4-byte D2H read of the
// detection count at the end sizes the result; nothing else synchronises.
//
// The overlap test is torchvision's CPU `nms_kernel_impl` arithmetic in the same operation order, so the kept
// set is bit-identical given identical boxes; ties in score are broken by anchor index (stable sort).
#include <cub/cub.cuh>
#include "common.cuh"

namespace frn {
namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 1024;

// BBoxTransform.forward + ClipBoxes.forward, each step rounded separately like the ATen ops.
__global__ void __launch_bounds__(kThreads) decode_clip_kernel(int A, const float4* __restrict__ anchors,
                                                               const float4* __restrict__ deltas, float width,
                                                               float height, float4* __restrict__ boxes) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  float4 b = anchors[i], d = deltas[i];
  float w = __fsub_rn(b.z, b.x), h = __fsub_rn(b.w, b.y);
  float cx = __fadd_rn(b.x, __fmul_rn(0.5f, w)), cy = __fadd_rn(b.y, __fmul_rn(0.5f, h));
  float dx = __fadd_rn(__fmul_rn(d.x, 0.1f), 0.f), dy = __fadd_rn(__fmul_rn(d.y, 0.1f), 0.f);
  float dw = __fadd_rn(__fmul_rn(d.z, 0.2f), 0.f), dh = __fadd_rn(__fmul_rn(d.w, 0.2f), 0.f);
  float pcx = __fadd_rn(cx, __fmul_rn(dx, w)), pcy = __fadd_rn(cy, __fmul_rn(dy, h));
  float pw = __fmul_rn(expf(dw), w), ph = __fmul_rn(expf(dh), h);
  float4 o;
  o.x = fmaxf(__fsub_rn(pcx, __fmul_rn(0.5f, pw)), 0.f);
  o.y = fmaxf(__fsub_rn(pcy, __fmul_rn(0.5f, ph)), 0.f);
  o.z = fminf(__fadd_rn(pcx, __fmul_rn(0.5f, pw)), width);
  o.w = fminf(__fadd_rn(pcy, __fmul_rn(0.5f, ph)), height);
  boxes[i] = o;
}

// keys[k][a] = score if score > thresh else -1 (sorted to the end); iota[a] = a.
__global__ void __launch_bounds__(kThreads) nms_keys_kernel(int A, int K, const float* __restrict__ cls, float thresh,
                                                            float* __restrict__ keys, int* __restrict__ iota) {
  long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (i >= (long)A * K) return;
  int a = (int)(i / K), k = (int)(i - (long)a * K);
  float s = cls[i];
  keys[(long)k * A + a] = s > thresh ? s : -1.f;
  if (k == 0) iota[a] = a;
}

__device__ __forceinline__ float box_area(float4 b) { return __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y)); }

// torchvision nms_kernel_impl: ovr = inter / (iarea + area_j - inter) > thr.  Disjoint boxes (inter == 0) give
// 0 (or NaN for two empty boxes): never above a positive threshold, so the division is skipped.
__device__ __forceinline__ bool overlaps(float4 a, float area_a, float4 b, float area_b, float thr) {
  float w = fmaxf(0.f, __fsub_rn(fminf(a.z, b.z), fmaxf(a.x, b.x)));
  float h = fmaxf(0.f, __fsub_rn(fminf(a.w, b.w), fmaxf(a.y, b.y)));
  float inter = __fmul_rn(w, h);
  if (!(inter > 0.f)) return false;
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_a, area_b), inter)) > thr;
}

// One CTA (1024 threads) per class.  keys / order: the class's scores sorted descending with the anchor index.
__global__ void __launch_bounds__(kChunk) nms_class_kernel(int A, const float* __restrict__ keys_all,
                                                           const int* __restrict__ order_all,
                                                           const float4* __restrict__ boxes, float score_thr,
                                                           float iou_thr, int* kept_idx_all, float4* kept_box_all,
                                                           int* __restrict__ kept_count) {
  extern __shared__ unsigned s_mask[];                         // [kChunk][32] words, column index swizzled by row
  float4* s_box = reinterpret_cast<float4*>(s_mask + kChunk * 32);
  float* s_area = reinterpret_cast<float*>(s_box + kChunk);
  __shared__ unsigned s_removed[32];
  int k = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const float* keys = keys_all + (long)k * A;
  const int* order = order_all + (long)k * A;
  int* kept_idx = kept_idx_all + (long)k * A;
  float4* kept_box = kept_box_all + (long)k * A;
  int kept_total = 0;
  for (int base = 0; base < A; base += kChunk) {
    if (!(keys[base] > score_thr)) break;                      // sorted descending: no candidate left (uniform)
    int i = base + t;
    bool cand = i < A && keys[i] > score_thr;
    int src = cand ? order[i] : 0;
    float4 bx = cand ? boxes[src] : make_float4(0.f, 0.f, 0.f, 0.f);
    float area = box_area(bx);
    s_box[t] = bx;
    s_area[t] = area;
    // phase 1: against everything kept so far (uniform addresses: broadcast loads)
    bool alive = cand;
    for (int q = 0; q < kept_total && alive; q++) {
      float4 kb = kept_box[q];
      if (overlaps(kb, box_area(kb), bx, area, iou_thr)) alive = false;
    }
    unsigned gone = ~__ballot_sync(0xffffffffu, alive);
    if (lane == 0) s_removed[warp] = gone;
    __syncthreads();
    // phase 2: row t of the overlap matrix against the later boxes of the chunk.  Words below the diagonal are
    // zero, the diagonal word needs the j > t guard, the words above it do not (fully unrolled, one LDS.128 and
    // ~12 ALU instructions per pair; the division only for overlapping pairs).
#pragma unroll 1
    for (int w = 0; w < 32; w++) {
      unsigned bits = 0;
      if (alive && w == warp) {
#pragma unroll
        for (int b = 0; b < 32; b++) {
          int j = w * 32 + b;
          if (b > lane && overlaps(bx, area, s_box[j], s_area[j], iou_thr)) bits |= 1u << b;
        }
      } else if (alive && w > warp) {
#pragma unroll
        for (int b = 0; b < 32; b++) {
          int j = w * 32 + b;
          if (overlaps(bx, area, s_box[j], s_area[j], iou_thr)) bits |= 1u << b;
        }
      }
      s_mask[t * 32 + (w ^ lane)] = bits;
    }
    __syncthreads();
    // phase 3: greedy scan, one warp; lane w owns the removed-bits of boxes [32 w, 32 w + 32)
    if (warp == 0) {
      unsigned r = s_removed[lane];
      for (int w = 0; w < 32; w++) {
        unsigned rw = __shfl_sync(0xffffffffu, r, w), done = 0;
        while (true) {
          unsigned pending = ~rw & ~done;
          if (!pending) break;
          int b = __ffs(pending) - 1;
          done |= (2u << b) - 1u;                              // bits 0..b are settled
          int row = w * 32 + b;                                // kept: it removes what it overlaps
          r |= s_mask[row * 32 + (lane ^ (row & 31))];
          rw = __shfl_sync(0xffffffffu, r, w);
        }
      }
      s_removed[lane] = r;
    }
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < 32; w++) {
      int c = __popc(~s_removed[w]);
      before += w < warp ? c : 0;
      total += c;
    }
    unsigned mine = ~s_removed[warp];
    if ((mine >> lane) & 1u) {
      int pos = kept_total + before + __popc(mine & ((1u << lane) - 1u));
      kept_idx[pos] = src;
      kept_box[pos] = bx;
    }
    kept_total += total;
    __syncthreads();                                           // kept_box visible to the next chunk's phase 1
  }
  if (t == 0) kept_count[k] = kept_total;
}

// Concatenate the classes (model.py:352-362): out[offset_k + pos] for pos < kept_count[k].
__global__ void __launch_bounds__(kThreads) nms_gather_kernel(int A, int K, const float* __restrict__ cls,
                                                              const float4* __restrict__ boxes,
                                                              const int* __restrict__ kept_idx,
                                                              const int* __restrict__ kept_count, int capacity,
                                                              float* __restrict__ out_scores,
                                                              int* __restrict__ out_classes,
                                                              float4* __restrict__ out_boxes,
                                                              int* __restrict__ out_index,
                                                              int* __restrict__ out_count) {
  int k = blockIdx.y, pos = blockIdx.x * blockDim.x + threadIdx.x;
  int offset = 0, total = 0;
  for (int q = 0; q < K; q++) {
    int c = kept_count[q];
    offset += q < k ? c : 0;
    total += c;
  }
  if (k == 0 && pos == 0) *out_count = total;
  if (pos >= kept_count[k] || offset + pos >= capacity) return;
  int a = kept_idx[(long)k * A + pos];
  out_scores[offset + pos] = cls[(long)a * K + k];
  out_classes[offset + pos] = k;
  out_boxes[offset + pos] = boxes[a];
  if (out_index) out_index[offset + pos] = a;
}

size_t sort_temp_bytes(int A) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, (const float*)nullptr, (float*)nullptr, (const int*)nullptr,
                                            (int*)nullptr, A);
  return bytes;
}

struct NmsBuffers {
  float4* boxes; float* keys_in; float* keys_out; int* iota; int* order; int* kept_idx; float4* kept_box;
  int* kept_count; void* sort_temp; size_t sort_bytes;
};

bool carve(Arena& ar, int A, int K, NmsBuffers& b) {
  b.boxes = ar.take<float4>(A);
  b.keys_in = ar.take<float>((size_t)A * K);
  b.keys_out = ar.take<float>((size_t)A * K);
  b.iota = ar.take<int>(A);
  b.order = ar.take<int>((size_t)A * K);
  b.kept_idx = ar.take<int>((size_t)A * K);
  b.kept_box = ar.take<float4>((size_t)A * K);
  b.kept_count = ar.take<int>(K);
  b.sort_bytes = sort_temp_bytes(A);
  b.sort_temp = ar.take<char>(b.sort_bytes);
  return ar.ok;
}

constexpr size_t kNmsSmem = (size_t)kChunk * 32 * 4 + kChunk * 16 + kChunk * 4;

int run_nms(cudaStream_t st, int A, int K, const float* cls, const NmsBuffers& b, float score_thr, float iou_thr,
            int capacity, float* out_scores, int* out_classes, float* out_boxes, int* out_index, int* out_count) {
  nms_keys_kernel<<<cdiv((long)A * K, kThreads), kThreads, 0, st>>>(A, K, cls, score_thr, b.keys_in, b.iota);
  FRN_LAUNCH_CHECK();
  for (int k = 0; k < K; k++) {
    size_t bytes = b.sort_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairsDescending(b.sort_temp, bytes, b.keys_in + (size_t)k * A,
                                                              b.keys_out + (size_t)k * A, b.iota,
                                                              b.order + (size_t)k * A, A, 0, 32, st);
    if (e != cudaSuccess) return (int)e;
  }
  cudaError_t e = cudaFuncSetAttribute(nms_class_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNmsSmem);
  if (e != cudaSuccess) return (int)e;
  nms_class_kernel<<<K, kChunk, kNmsSmem, st>>>(A, b.keys_out, b.order, b.boxes, score_thr, iou_thr, b.kept_idx,
                                               b.kept_box, b.kept_count);
  FRN_LAUNCH_CHECK();
  nms_gather_kernel<<<dim3(cdiv(A, kThreads), K), kThreads, 0, st>>>(A, K, cls, b.boxes, b.kept_idx, b.kept_count,
                                                                     capacity, out_scores, out_classes,
                                                                     reinterpret_cast<float4*>(out_boxes), out_index,
                                                                     out_count);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

}  // namespace
}  // namespace frn

using namespace frn;

#pragma GCC visibility push(default)
extern "C" {

size_t frn_detect_workspace_bytes(int A, int K) {
  if (A <= 0 || K <= 0) return 0;
  Arena ar(nullptr, 0, true);
  NmsBuffers b;
  carve(ar, A, K, b);
  return ar.peak;
}

int frn_decode_boxes(void* stream, int A, const float* anchors, const float* regression, int img_h, int img_w,
                     float* boxes) {
  if (A <= 0 || img_h <= 0 || img_w <= 0) return FRN_ERR_BAD_SHAPE;
  if (!anchors || !regression || !boxes) return FRN_ERR_NULL_POINTER;
  decode_clip_kernel<<<cdiv(A, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      A, reinterpret_cast<const float4*>(anchors), reinterpret_cast<const float4*>(regression), (float)img_w,
      (float)img_h, reinterpret_cast<float4*>(boxes));
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

int frn_detect_postprocess(void* stream, int A, int K, const float* classification, const float* regression,
                           const float* anchors, const float* boxes_in, int img_h, int img_w, float score_thresh,
                           float iou_thresh, int capacity, float* out_scores, int* out_classes, float* out_boxes,
                           int* out_index, int* out_count, void* ws, size_t ws_bytes) {
  if (A <= 0 || K <= 0 || capacity < 0) return FRN_ERR_BAD_SHAPE;
  if (!classification || !out_scores || !out_classes || !out_boxes || !out_count || !ws) return FRN_ERR_NULL_POINTER;
  if (!boxes_in && (!regression || !anchors)) return FRN_ERR_NULL_POINTER;
  if (!boxes_in && (img_h <= 0 || img_w <= 0)) return FRN_ERR_BAD_SHAPE;
  if (!(iou_thresh > 0.f)) return FRN_ERR_UNSUPPORTED;
  Arena ar(ws, ws_bytes);
  NmsBuffers b;
  if (!carve(ar, A, K, b)) return FRN_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  ScopedTimer tm(st, FRN_TIME_ROWS);
  if (boxes_in) {
    cudaError_t e = cudaMemcpyAsync(b.boxes, boxes_in, (size_t)A * 16, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return (int)e;
  } else {
    FRN_TRY(frn_decode_boxes(stream, A, anchors, regression, img_h, img_w, (float*)b.boxes));
  }
  return run_nms(st, A, K, classification, b, score_thresh, iou_thresh, capacity, out_scores, out_classes, out_boxes,
                 out_index, out_count);
}

}  // extern "C"
#pragma GCC visibility pop
