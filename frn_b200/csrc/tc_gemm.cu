// tcgen05 (5th-gen tensor core) version of the strided batched GEMM in gemm.cu, with
// split-precision operands: every fp32 operand x is staged in shared memory as a tf32 pair
// (hi, lo) and the product is accumulated as hi*hi' (main accumulator) and lo*hi' + hi*lo' + lo*lo'
// (second accumulator) in fp32 TMEM columns (split TF32), which keeps the fp32 parity gate
// (<= 1e-5 vs fp64) that single-pass TF32 (~1e-3) cannot meet.
//
// CTA = 128 x BN output tile, 17 warps:
//   warps 0-3 / 4-7  two operand-staging groups that alternate 32-deep K stages (global ->
//                    registers, all loads of a stage in flight before the first use -> hi/lo split
//                    -> SWIZZLE_128B K-major shared memory; arbitrary element strides, so every
//                    transposed use is free),
//   warp 8           tcgen05.mma issue (whole warp waits, one elected lane issues: M=128, N=BN,
//                    K=8 per instruction, 16 per stage) and commits to mbarriers,
//   warps 9-16       accumulator drain + epilogue (thread = output row, half of the columns).
// The tensor core TRUNCATES when it aligns addends to the accumulator: a relative bias of
// ~3e-9 per K = 8 step that grows linearly with K (tools/tc_error_vs_k.py: 6e-6 at K = 2048).
// So no accumulation chain is longer than kChunk = 4 stages (K = 128): the accumulators are
// double buffered in TMEM and each finished chunk is added into fp32 registers by the drain
// warps while the next chunk is being multiplied.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <set>
#include <utility>

#include "common.cuh"
#include "tc05.cuh"

namespace frn {
namespace {

using namespace tc;

constexpr int TBM = 128;      // UMMA M
constexpr int TBK = 32;       // 32 tf32 = one 128-byte swizzle row
constexpr int kLoaders = 256;  // staging threads (warps 0-7)
constexpr int kMmaWarp = 8;
constexpr int kEpiWarp0 = 9;
constexpr int kEpiThreads = 256;
constexpr int kTcThreads = (kEpiWarp0 + 8) * 32;   // 544
constexpr int kSmCount = 148;  // B200
constexpr int kChunk = 4;     // stages per accumulation chain (K = 128)

// Optional cycle accounting of the role loops (CTA 0, one thread per role): compile with -DFRN_GEMM_TRACE.
#ifdef FRN_GEMM_TRACE
__device__ unsigned long long g_trace[16];
#define TRACE_T0() const long long tr0__ = clock64()
#define TRACE_ADD(slot) do { if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) atomicAdd(&g_trace[slot], (unsigned long long)(clock64() - tr0__)); } while (0)
// whole-kernel cycles (slot 13) and nanoseconds (slot 14) of CTA 0: their ratio is the SM clock under this kernel's load
#define TRACE_KERNEL_BEGIN() const long long trk0__ = clock64(); unsigned long long trn0__; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(trn0__))
#define TRACE_KERNEL_END() do { if (blockIdx.x == 0 && threadIdx.x == 0) { unsigned long long n1__; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(n1__)); \
    atomicAdd(&g_trace[13], (unsigned long long)(clock64() - trk0__)); atomicAdd(&g_trace[14], n1__ - trn0__); } } while (0)
#else
#define TRACE_T0() do {} while (0)
#define TRACE_ADD(slot) do {} while (0)
#define TRACE_KERNEL_BEGIN() do {} while (0)
#define TRACE_KERNEL_END() do {} while (0)
#endif

// -DFRN_GEMM_TS: clock64 stamps of the hand-over points of the first 256 stages of CTA 0 (plain stores, no atomics),
// read back with frn_debug_gemm_ts: the timeline of one stage slot through producer -> split -> MMA -> producer.
#ifdef FRN_GEMM_TS
__device__ long long g_ts[8][256];
#define TS(role, it) do { if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (it) < 256) g_ts[role][(int)(it)] = clock64(); } while (0)
#define TS_ANY(role, it) do { if (blockIdx.x == 0 && (it) < 256) g_ts[role][(int)(it)] = clock64(); } while (0)
#else
#define TS(role, it) do {} while (0)
#define TS_ANY(role, it) do {} while (0)
#endif

template <int BN> struct TcCfg {
  static constexpr int kStages = BN == 128 ? 2 : 3;        // split (hi / lo) operand stages feeding the MMAs
  static constexpr int kRaw = BN == 32 ? 3 : 2;            // raw fp32 stages (cp.async ring)
  static constexpr int kEpiBytes = 8 * 32 * 32 * 4;        // one 32 x 32 transposition tile per epilogue warp
  static constexpr int kABytes = TBM * TBK * 4;            // one of hi / lo
  static constexpr int kBBytes = BN * TBK * 4;
  static constexpr int kStageBytes = 2 * kABytes + 2 * kBBytes;
  static constexpr int kRawBytes = kABytes + kBBytes;
  static constexpr int kSmem = kStages * kStageBytes + kRaw * kRawBytes + kEpiBytes + 1024 /*align*/ +
                               1024 /*barriers, scratch*/;
  static constexpr int kTmemCols = 4 * BN;               // 2 buffers x (hi*hi | lo*hi + hi*lo + lo*lo)
};

// ---- cp.async (LDGSTS) with zero fill: copies src_bytes (<= 16 or 4) and zero-fills the rest ----
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Per-thread plan of one operand's [ROWS x 32] tile, computed ONCE per kernel: the stage loop then runs
// without index arithmetic (the generic-stride address math cost ~65 instructions per 16-byte chunk when it
// lived inside the loop and made the staging warps, not the tensor pipe, the bottleneck).
// Raw layout: k contiguous ([r][32]) when ks == 1 (KFAST), else r contiguous ([k][ROWS]).  Chunk = 4
// consecutive elements along the unit-stride axis; thread t owns chunks t, t + 256, ...: everything about
// chunk i is LINEAR in i (outer index, global offset, swizzled destination), so the plan is a handful of
// registers (the first version kept four arrays per operand and spilled them into the stage loop).
template <int ROWS> struct OperandPlan {
  static constexpr int CH = ROWS * 8 / kLoaders;
  static constexpr int IPO = ROWS / 4;      // chunks per outer index when the rows are contiguous
  long goff0, gstep;    // global element offset of chunk 0 from the tile origin; step per chunk
  int outer0, ostep;    // index along the strided axis (row if KFAST else k) of chunk 0; step per chunk
  int inner;            // first index along the unit-stride axis (k if KFAST else row), same for all chunks
  uint32_t dst0;        // KFAST: byte offset of chunk 0 in the swizzled hi / lo tiles (+ 32 rows per chunk)
  bool kfast;
  __device__ __forceinline__ void init(long rs, long ks, int t) {
    kfast = (ks == 1);
    const long os = kfast ? rs : ks;
    if (kfast) { inner = (t & 7) * 4; outer0 = t >> 3; ostep = kLoaders / 8; }
    else { inner = (t % IPO) * 4; outer0 = t / IPO; ostep = kLoaders / IPO; }
    dst0 = (uint32_t)sw128_chunk_offset(t >> 3, t & 7) * 4u;
    goff0 = (long)outer0 * os + inner;
    gstep = (long)ostep * os;
  }
};

template <int ROWS>
__device__ __forceinline__ void issue_raw(uint32_t raw, const float* __restrict__ g, const OperandPlan<ROWS>& pl,
                                          long os, int rows, int kvalid, int t) {
  constexpr int CH = OperandPlan<ROWS>::CH;
  const bool vec = ((os & 3) == 0) && ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
  const bool interior = rows >= ROWS && kvalid >= TBK;
  const float* src = g + pl.goff0;
  uint32_t dst = raw + (uint32_t)t * 16u;
  if (interior && vec) {
#pragma unroll
    for (int i = 0; i < CH; i++, src += pl.gstep, dst += kLoaders * 16) cp_async16(dst, src, 16);
    return;
  }
  const int n_in = pl.kfast ? kvalid : rows, n_out = pl.kfast ? rows : kvalid;
  int valid0 = n_in - pl.inner;
  valid0 = valid0 < 0 ? 0 : (valid0 > 4 ? 4 : valid0);
#pragma unroll
  for (int i = 0; i < CH; i++, src += pl.gstep, dst += kLoaders * 16) {
    const int valid = (pl.outer0 + i * pl.ostep) < n_out ? valid0 : 0;
    if (vec) {
      cp_async16(dst, valid ? src : g, valid * 4);
    } else {
#pragma unroll
      for (int q = 0; q < 4; q++) cp_async4(dst + q * 4, q < valid ? src + q : g, q < valid ? 4 : 0);
    }
  }
}

// raw tile -> tf32 hi / lo split -> SWIZZLE_128B K-major tiles (all by 32-bit shared addresses)
template <int ROWS, bool KFAST>
__device__ __forceinline__ void convert_tile(uint32_t hi, uint32_t lo, uint32_t raw, const OperandPlan<ROWS>& pl,
                                             int t, bool with_lo) {
  constexpr int CH = OperandPlan<ROWS>::CH;
  float v[CH][4];
  uint32_t o[CH];
  if (KFAST) {
#pragma unroll
    for (int i = 0; i < CH; i++) {             // all loads of the tile first, then the arithmetic
      const float4 x = lds128(raw + (uint32_t)(t + i * kLoaders) * 16u);
      v[i][0] = x.x; v[i][1] = x.y; v[i][2] = x.z; v[i][3] = x.w;
      o[i] = pl.dst0 + (uint32_t)i * (kLoaders / 8) * 128u;      // + 32 rows of 128 bytes
    }
  } else {
#pragma unroll
    for (int i = 0; i < CH; i++) {
      const int idx = t + i * kLoaders;
      const int r = idx % ROWS, c = idx / ROWS;            // ROWS is a power of two: shifts and masks
#pragma unroll
      for (int q = 0; q < 4; q++) v[i][q] = lds32(raw + (uint32_t)((c * 4 + q) * ROWS + r) * 4u);
      o[i] = (uint32_t)sw128_chunk_offset(r, c) * 4u;
    }
  }
#pragma unroll
  for (int i = 0; i < CH; i++) {
    float h[4], l[4];
#pragma unroll
    for (int q = 0; q < 4; q++) split_tf32(v[i][q], h[q], l[q]);
    sts128(hi + o[i], h[0], h[1], h[2], h[3]);
    if (with_lo) sts128(lo + o[i], l[0], l[1], l[2], l[3]);
  }
}

// Work item of the persistent loop: output tile (m, n) of batch entry / K split z.
struct TcTile { int m0, n0, z, t_begin, ntiles, b_fixed; };

__device__ __forceinline__ TcTile tc_tile(const GemmDesc& p, int id, int gx, int gy, int ktiles,
                                          int tiles_per_split, int BN) {
  TcTile t;
  const int my = id % gy;            // m fastest: CTAs that run together share the same B (activation) tile
  const int nx = (id / gy) % gx;
  t.z = id / (gx * gy);
  t.m0 = my * TBM; t.n0 = nx * BN;
  if (p.reduce_batch && p.per_batch_split > 0) {       // splits never cross a batch entry
    const int b = t.z / p.per_batch_split, j = t.z - b * p.per_batch_split;
    t.t_begin = b * ktiles + j * tiles_per_split;
    t.ntiles = max(min(tiles_per_split, ktiles - j * tiles_per_split), 0);
    t.b_fixed = -1;
  } else if (p.reduce_batch) {
    t.t_begin = t.z * tiles_per_split;
    const int t_end = min(t.t_begin + tiles_per_split, p.batch * ktiles);
    t.ntiles = max(t_end - t.t_begin, 0);
    t.b_fixed = -1;
  } else {
    t.t_begin = 0; t.ntiles = ktiles; t.b_fixed = t.z;
  }
  return t;
}

// Epilogue of one output tile, shared by both kernels: the drain thread (row = quad * 32 + lane, columns
// [half * BN / 2, (half + 1) * BN / 2) of the tile) holds the finished sums in registers.  `tile` is the warp's own
// 4 KB transposition buffer in shared memory, `red` 128 floats shared by the 8 drain warps.
template <int BN, int EPI, bool ACC>
__device__ __forceinline__ void tc_epilogue(const GemmDesc& p, float* __restrict__ aux, const TcTile& tl,
                                            float (&acc)[BN / 2], uint32_t tile, float* red, int nsplit, int gx,
                                            int quad, int half, int lane) {
  constexpr int HC = BN / 2;
  const int row = quad * 32 + lane;
  const int gm = tl.m0 + row, z = tl.z;
  const bool partial = p.reduce_batch && (nsplit > 1 || p.per_batch_split > 0);
  const int nb = tl.n0 + half * HC;          // first global column of this thread
  if (EPI == GEMM_EPI_ROWSUMSQ) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < HC; j++) {
      const float x = p.alpha * acc[j];
      s += (nb + j < p.N) ? x * x : 0.f;
    }
    if (half == 1) red[row] = s;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (half == 0 && gm < p.M) aux[((long)z * p.M + gm) * gx + (tl.n0 / BN)] = s + red[row];
    asm volatile("bar.sync 1, 256;" ::: "memory");     // red[] is reused by the next tile
  } else if (gm < p.M) {
    float* Cb; long sCm, sCn;
    if (partial) { Cb = aux + (long)z * p.M * p.N; sCm = p.N; sCn = 1; }
    else { Cb = p.C + (p.reduce_batch ? 0 : (long)z * p.sCb); sCm = p.sCm; sCn = p.sCn; }
    const bool acc_out = ACC && !partial;
    const long boff = p.reduce_batch ? 0 : (long)z * p.bias_bs;
    const float bm = (!partial && p.bias_m) ? __ldg(p.bias_m + boff + gm) : 0.f;
    const bool vec = (sCn == 1) && ((sCm & 3) == 0) && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0) &&
                     (EPI != GEMM_EPI_GROUPS || (((reinterpret_cast<uintptr_t>(p.C1) | reinterpret_cast<uintptr_t>(p.C2) |
                                           (uintptr_t)(p.sCb * 4)) & 15) == 0));
    if (!vec) {                             // n-contiguous aligned outputs take the coalesced path below
#pragma unroll
    for (int c0 = 0; c0 < HC; c0 += 4) {
      float o[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int gn = nb + c0 + j;
        const float bn = (!partial && p.bias_n && gn < p.N) ? __ldg(p.bias_n + boff + gn) : 0.f;
        o[j] = fmaf(p.alpha, acc[c0 + j], bm + bn);
      }
      float* dst = Cb + (long)gm * sCm + (long)(nb + c0) * sCn;
      if (EPI == GEMM_EPI_GROUPS) {          // row-grouped outputs (never partial / accumulated)
        const int g = gm / p.group_m;
        dst = (g == 0 ? p.C : (g == 1 ? p.C1 : p.C2)) + (long)z * p.sCb + (long)(gm - g * p.group_m) * sCm +
              (long)(nb + c0) * sCn;
      }
      if (vec && nb + c0 + 3 < p.N) {
        float4 q = make_float4(o[0], o[1], o[2], o[3]);
        float4* d4 = reinterpret_cast<float4*>(dst);
        if (acc_out) { float4 c = *d4; q.x += c.x; q.y += c.y; q.z += c.z; q.w += c.w; }
        *d4 = q;
      } else {
#pragma unroll
        for (int j = 0; j < 4; j++)
          if (nb + c0 + j < p.N) {
            float* d1 = dst + (long)j * sCn;
            *d1 = o[j] + (acc_out ? *d1 : 0.f);
          }
      }
    }
    }
  }
  if (EPI != GEMM_EPI_ROWSUMSQ) {
    // Coalesced path for n-contiguous outputs: thread = row is the TMEM read layout, but a warp-wide store
    // of 16 bytes per ROW touches 32 different lines (measured: the epilogue, not the MMAs, bounded every
    // K = 32 GEMM of the backward pass).  Each warp transposes 32 x 32 blocks through a swizzled shared
    // tile and writes whole 128-byte row segments, four rows per instruction.
    float* Cb; long sCm, sCn;
    if (partial) { Cb = aux + (long)z * p.M * p.N; sCm = p.N; sCn = 1; }
    else { Cb = p.C + (p.reduce_batch ? 0 : (long)z * p.sCb); sCm = p.sCm; sCn = p.sCn; }
    const bool vec = (sCn == 1) && ((sCm & 3) == 0) && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0) &&
                     (EPI != GEMM_EPI_GROUPS || (((reinterpret_cast<uintptr_t>(p.C1) | reinterpret_cast<uintptr_t>(p.C2) |
                                           (uintptr_t)(p.sCb * 4)) & 15) == 0));
    if (vec) {
      constexpr int W = HC < 32 ? HC : 32;      // block width (columns)
      constexpr int CPR = W / 4;                // 16-byte chunks per row
      constexpr int RPI = 32 / CPR;             // rows per store instruction
      const bool acc_out = ACC && !partial;
      const long boff = p.reduce_batch ? 0 : (long)z * p.bias_bs;
      const float bm = (!partial && p.bias_m && gm < p.M) ? __ldg(p.bias_m + boff + gm) : 0.f;
      const int rsub = lane / CPR, csub = lane % CPR;
#pragma unroll
      for (int cb = 0; cb < HC; cb += W) {
#pragma unroll
        for (int c = 0; c < CPR; c++) {         // thread = row `lane` of the warp's 32
          sts128(tile + (uint32_t)(lane * 32 + ((c ^ (lane & 7)) << 2)) * 4u,
                 fmaf(p.alpha, acc[cb + 4 * c], bm), fmaf(p.alpha, acc[cb + 4 * c + 1], bm),
                 fmaf(p.alpha, acc[cb + 4 * c + 2], bm), fmaf(p.alpha, acc[cb + 4 * c + 3], bm));
        }
        __syncwarp();
        const int gn = nb + cb + 4 * csub;
        float4 bn4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!partial && p.bias_n && gn + 3 < p.N) bn4 = __ldg(reinterpret_cast<const float4*>(p.bias_n + boff + gn));
        // C += goes out as a vector reduction (red.global.add.v4.f32): the add happens in L2, nothing is
        // read back into the SM and no registers hold old values (a prefetch array spilled to local memory
        // and every store then waited on a local load -- ncu: the top stall of the K = 32 GEMMs).  Each
        // element receives exactly one add per launch, so the result does not depend on any ordering.
#pragma unroll
        for (int it = 0; it < 32 / RPI; it++) {
          const int rr = it * RPI + rsub;
          const int grow = tl.m0 + quad * 32 + rr;
          float4 q = lds128(tile + (uint32_t)(rr * 32 + ((csub ^ (rr & 7)) << 2)) * 4u);
          if (grow < p.M && gn < p.N) {
            float* dst = Cb + (long)grow * sCm + gn;
            if (EPI == GEMM_EPI_GROUPS) {
              const int g = grow / p.group_m;
              dst = (g == 0 ? p.C : (g == 1 ? p.C1 : p.C2)) + (long)z * p.sCb + (long)(grow - g * p.group_m) * sCm + gn;
            }
            if (gn + 3 < p.N) {
              q.x += bn4.x; q.y += bn4.y; q.z += bn4.z; q.w += bn4.w;
              if (acc_out) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(q.x), "f"(q.y),
                             "f"(q.z), "f"(q.w) : "memory");
              } else {
                *reinterpret_cast<float4*>(dst) = q;
              }
            } else {
              const float o[4] = {q.x, q.y, q.z, q.w};
              for (int j = 0; j < 4 && gn + j < p.N; j++) {
                const float bn = (!partial && p.bias_n) ? __ldg(p.bias_n + boff + gn + j) : 0.f;
                dst[j] = o[j] + bn + (acc_out ? dst[j] : 0.f);
              }
            }
          }
        }
        __syncwarp();
      }
    }
  }
}

// Persistent: gridDim.x CTAs (one per SM) walk the gx*gy*gz output tiles; the shared-memory stage ring
// and the two TMEM accumulator buffers keep running across tiles, so the staging of tile i+1 and its
// first MMAs overlap the drain / global stores of tile i.
template <int BN, int EPI, bool ACC>      // ACC: C += (a separate instantiation keeps the read-modify-write
__global__ void __launch_bounds__(kTcThreads, 1)    // registers out of the plain-store kernels)
tc_gemm_kernel(GemmDesc p, float* __restrict__ aux, int ktiles, int tiles_per_split, int nsplit, int gx, int gy,
               int gz, int use_tma, const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB) {
  using Cfg = TcCfg<BN>;
  constexpr int HC = BN / 2;          // columns per drain thread
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty = full + Cfg::kStages;
  uint64_t* acc_full = empty + Cfg::kStages;     // 2
  uint64_t* acc_free = acc_full + 2;             // 2
  uint64_t* raw_full = acc_free + 2;             // kRaw (TMA path)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(raw_full + Cfg::kRaw);
  float* red = reinterpret_cast<float*>(tmem_slot + 4);   // 128 floats (row-sum-of-squares epilogue)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total = gx * gy * gz;

  if (tid == 0) {
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&full[s], kLoaders); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; s++) { mbar_init(&acc_full[s], 1); mbar_init(&acc_free[s], kEpiThreads); }
    for (int s = 0; s < Cfg::kRaw; s++) mbar_init(&raw_full[s], 1);
    if (use_tma) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
    mbar_init_fence();
  }
  if (warp == kMmaWarp) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp < kMmaWarp) {
    // ===== operand staging: cp.async raw tiles kRaw - 1 stages ahead, then split into the MMA stage ring =====
    const uint32_t raw_base = smem_u32(smem + Cfg::kStages * Cfg::kStageBytes + 1024);
    const uint32_t stage_base = smem_u32(smem);
    OperandPlan<TBM> planA;
    OperandPlan<BN> planB;
    planA.init(p.sAm, p.sAk, tid);
    planB.init(p.sBn, p.sBk, tid);
    const long osA = planA.kfast ? p.sAm : p.sAk, osB = planB.kfast ? p.sBn : p.sBk;
    // Tile cursor.  The tile decomposition (integer divisions) runs once per TILE, and the (batch entry,
    // k tile) pair of the split-K walk advances incrementally: recomputing both per stage cost the staging
    // warps ~850 cycles per stage, as much as the conversion itself.
    struct Cursor { int id, i, b, kt; TcTile tl; };
    auto enter = [&](Cursor& c) {                  // c.id: a new tile, or past the end; skips empty tiles
      while (c.id < total) {
        c.tl = tc_tile(p, c.id, gx, gy, ktiles, tiles_per_split, BN);
        if (c.tl.ntiles > 0) break;
        c.id += gridDim.x;
      }
      c.i = 0; c.b = 0; c.kt = 0;
      if (c.id < total) {
        if (c.tl.b_fixed >= 0) c.b = c.tl.b_fixed;
        else { c.b = c.tl.t_begin / ktiles; c.kt = c.tl.t_begin - c.b * ktiles; }
      }
    };
    auto advance = [&](Cursor& c) {
      c.i++; c.kt++;
      if (c.tl.b_fixed < 0 && c.kt == ktiles) { c.kt = 0; c.b++; }
      if (c.i >= c.tl.ntiles) { c.id += gridDim.x; enter(c); }
    };
    auto issue = [&](const Cursor& c, long it) {
      const TcTile& tl = c.tl;
      const int b = c.b, k0 = c.kt * TBK;
      const int slot = (int)(it % Cfg::kRaw);
      const uint32_t rA = raw_base + (uint32_t)slot * Cfg::kRawBytes;
      const uint32_t rB = rA + Cfg::kABytes;
      if (use_tma) {
        // TMA: one thread asks for the two boxes (rows x 32 k, any strides, zero fill past the edges); the
        // 256 staging threads issue no loads at all (cp.async cost them as much time as the conversion)
        if (tid == 0) {
          mbar_expect_tx(&raw_full[slot], Cfg::kRawBytes);
          const int ba = p.sAb ? b : 0, bb = p.sBb ? b : 0;
          if (planA.kfast) tma_load_3d(rA, &tmA, k0, tl.m0, ba, &raw_full[slot]);
          else tma_load_3d(rA, &tmA, tl.m0, k0, ba, &raw_full[slot]);
          if (planB.kfast) tma_load_3d(rB, &tmB, k0, tl.n0, bb, &raw_full[slot]);
          else tma_load_3d(rB, &tmB, tl.n0, k0, bb, &raw_full[slot]);
        }
        return;
      }
      issue_raw<TBM>(rA, p.A + (long)b * p.sAb + (long)tl.m0 * p.sAm + (long)k0 * p.sAk, planA, osA,
                     p.M - tl.m0, p.K - k0, tid);
      issue_raw<BN>(rB, p.B + (long)b * p.sBb + (long)tl.n0 * p.sBn + (long)k0 * p.sBk, planB, osB,
                    p.N - tl.n0, p.K - k0, tid);
    };
    Cursor ci{(int)blockIdx.x, 0, 0, 0, {}}, cc{(int)blockIdx.x, 0, 0, 0, {}};
    enter(ci); enter(cc);
    long it_issue = 0;
    for (int r = 0; r < Cfg::kRaw - 1; r++) {      // prologue
      if (ci.id < total) { issue(ci, it_issue); advance(ci); }
      it_issue++;
      cp_async_commit();
    }
    for (long it = 0; cc.id < total; it++) {
      { TRACE_T0();
      cp_async_wait<Cfg::kRaw - 2>();              // this thread's copies of stage `it` have landed
      if (use_tma) mbar_wait(&raw_full[it % Cfg::kRaw], (uint32_t)((it / Cfg::kRaw) & 1));
      TRACE_ADD(0); }
      { TRACE_T0();
      asm volatile("bar.sync 2, 256;" ::: "memory");  // ... everyone's have, and everyone is done with stage it-1
      TRACE_ADD(1); }
      { TRACE_T0();
      if (ci.id < total) { issue(ci, it_issue); advance(ci); }
      it_issue++;
      cp_async_commit();
      TRACE_ADD(12); }
      const int stage = (int)(it % Cfg::kStages);
      const uint32_t phase = (uint32_t)((it / Cfg::kStages) & 1);
      { TRACE_T0();
      mbar_wait(&empty[stage], phase ^ 1);
      TRACE_ADD(2); }
      TRACE_T0();
      const uint32_t rA = raw_base + (uint32_t)(it % Cfg::kRaw) * Cfg::kRawBytes;
      const uint32_t rB = rA + Cfg::kABytes;
      const uint32_t sA = stage_base + (uint32_t)stage * Cfg::kStageBytes;
      const uint32_t sAl = sA + Cfg::kABytes;
      const uint32_t sB = sAl + Cfg::kABytes;
      const uint32_t sBl = sB + Cfg::kBBytes;
      { TRACE_T0();
      if (planA.kfast) convert_tile<TBM, true>(sA, sAl, rA, planA, tid, !p.fast);
      else convert_tile<TBM, false>(sA, sAl, rA, planA, tid, !p.fast);
      TRACE_ADD(9); }
      { TRACE_T0();
      if (planB.kfast) convert_tile<BN, true>(sB, sBl, rB, planB, tid, !p.fast);
      else convert_tile<BN, false>(sB, sBl, rB, planB, tid, !p.fast);
      TRACE_ADD(10); }
      { TRACE_T0();
      fence_proxy_async_smem();
      mbar_arrive(&full[stage]);
      TRACE_ADD(11); }
      TRACE_ADD(3);
      advance(cc);
    }
    cp_async_wait<0>();
  } else if (warp == kMmaWarp) {
    // ===== MMA issue =====
    constexpr uint32_t idesc = umma_idesc_tf32(TBM, BN);
    long it = 0;
    int gc = 0;                                    // running chunk (accumulation chain) counter
    for (int id = blockIdx.x; id < total; id += gridDim.x) {
      const TcTile tl = tc_tile(p, id, gx, gy, ktiles, tiles_per_split, BN);
      for (int i = 0; i < tl.ntiles; i++, it++) {
        const int stage = (int)(it % Cfg::kStages);
        const uint32_t phase = (uint32_t)((it / Cfg::kStages) & 1);
        const bool first = (i % kChunk) == 0;
        const bool last = (i % kChunk) == kChunk - 1 || i == tl.ntiles - 1;
        const int buf = gc & 1;
        { TRACE_T0();
        if (first && gc >= 2) mbar_wait(&acc_free[buf], ((gc >> 1) - 1) & 1);
        TRACE_ADD(4); }
        { TRACE_T0();
        mbar_wait(&full[stage], phase);
        TRACE_ADD(5); }
        tc_fence_after();
        TRACE_T0();
        if (elect_one()) {
          const uint32_t base = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t a_hi = umma_desc_sw128(base), a_lo = umma_desc_sw128(base + Cfg::kABytes);
          const uint64_t b_hi = umma_desc_sw128(base + 2 * Cfg::kABytes);
          const uint64_t b_lo = umma_desc_sw128(base + 2 * Cfg::kABytes + Cfg::kBBytes);
          const uint32_t d_main = tmem_acc + buf * 2 * BN, d_small = d_main + BN;
#pragma unroll
          for (int k = 0; k < TBK / 8; k++) {   // +32 bytes per K = 8 slice: start-address field += 2
            const bool acc = !(first && k == 0);
            // lo*lo (2^-22 of the product) is kept as well: when one operand is near-constant (|mean| >> sigma,
            // the blanked-RGB case of model.py:279-280) all of its variation lives in `lo`, and a product with a
            // zero-sum partner would otherwise be 2^-11 off.
            if (!p.fast) {     // p.fast: single-pass TF32 (bf16-tolerance inference): hi * hi only, no lo tiles staged
              umma_tf32_ss(d_small, a_lo + 2 * k, b_lo + 2 * k, idesc, acc);
              umma_tf32_ss(d_small, a_lo + 2 * k, b_hi + 2 * k, idesc, true);
              umma_tf32_ss(d_small, a_hi + 2 * k, b_lo + 2 * k, idesc, true);
            }
            umma_tf32_ss(d_main, a_hi + 2 * k, b_hi + 2 * k, idesc, acc);
          }
          umma_commit(&empty[stage]);            // frees the stage once these MMAs have read it
          if (last) umma_commit(&acc_full[buf]);
        }
        __syncwarp();
        TRACE_ADD(6);
        if (last) gc++;
      }
    }
  } else {
    // ===== accumulator drain + epilogue: thread = output row, HC of the BN columns =====
    const int e = warp - kEpiWarp0;
    const int quad = warp & 3, half = e >> 2;
    const int row = quad * 32 + lane;
    const uint32_t tlane = ((uint32_t)(quad * 32)) << 16;
    int gc = 0;
    for (int id = blockIdx.x; id < total; id += gridDim.x) {
      const TcTile tl = tc_tile(p, id, gx, gy, ktiles, tiles_per_split, BN);
      const int gm = tl.m0 + row, z = tl.z;
      float acc[HC];
#pragma unroll
      for (int j = 0; j < HC; j++) acc[j] = 0.f;
      const int nchunks = (tl.ntiles + kChunk - 1) / kChunk;
      for (int chunk = 0; chunk < nchunks; chunk++, gc++) {
        const int buf = gc & 1;
        { TRACE_T0();
        mbar_wait(&acc_full[buf], (gc >> 1) & 1);
        TRACE_ADD(7); }
        tc_fence_after();
        const uint32_t taddr = tmem_acc + tlane + buf * 2 * BN + half * HC;
#pragma unroll
        for (int c0 = 0; c0 < HC; c0 += 16) {
          float v[16], w[16];
          tmem_ld16(taddr + c0, v);
          if (!p.fast) {
            tmem_ld16(taddr + BN + c0, w);
          } else {
#pragma unroll
            for (int j = 0; j < 16; j++) w[j] = 0.f;
          }
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; j++) acc[c0 + j] += v[j] + w[j];
        }
        tc_fence_before();
        mbar_arrive(&acc_free[buf]);
      }
      TRACE_T0();
      tc_epilogue<BN, EPI, ACC>(p, aux, tl, acc, smem_u32(smem + Cfg::kStages * Cfg::kStageBytes + 1024 +
                                                          Cfg::kRaw * Cfg::kRawBytes) + (uint32_t)e * 4096u,
                                red, nsplit, gx, quad, half, lane);
      TRACE_ADD(8);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc<Cfg::kTmemCols>(tmem_acc);
}


// =====================================================================================================================
// v2 kernel: operands arrive by TMA *in the layout the tensor core reads* -- no staging copy of the high parts.
//
// kind::tf32 ignores the 13 low mantissa bits of a 32-bit operand (tools/umma_probe.cu, run on B200: all operand
// values are truncated, never rounded).  The raw fp32 tile is therefore its own "hi" part: hi = x with the low bits
// masked, lo = x - hi exactly.  TMA writes the raw tile with the hardware swizzle straight into the canonical
// K-major (K contiguous in memory, SWIZZLE_128B) or MN-major (the other dimension contiguous, i.e. the NCHW
// activations of a 1x1 convolution; SWIZZLE_128B with 32-byte atoms, the only MN-major layout 32-bit operands
// have) UMMA layout, and 8 split warps only compute lo = round_tf32(x - trunc_tf32(x)) element by element into a
// second tile with the SAME byte offsets: one LDS.128 + one STS.128 per four elements, no address arithmetic, no
// transposition, no bank conflicts.  (The first kernel staged raw tiles and rewrote both hi and lo, transposing
// MN-major operands with scalar loads: 1650 cycles of shared-memory traffic per stage against 1024 of MMA time.)
// |lo| < 2^-10 |x| and it is rounded to nearest at 11 bits, so x = hi + lo to 2^-22 relative, unbiased -- one bit
// better than the round-then-truncate split of the first kernel.  Products: lo.lo + lo.hi + hi.lo into the small
// accumulator, hi.hi into the main one, as before.
// accumulator, hi.hi into the main one ... except that hi.[hi | lo] and lo.[hi | lo] are issued as two N = 2 BN
// instructions (see the MMA warp), so the column halves hold hh + lh and hl + ll.
//
// 19 warps: 0 = TMA producer (one lane), 1 and 18 = MMA issue, 2-9 = lo split, 10-17 = accumulator drain + epilogue.
//
// TWO issuing warps.  Timestamps of the hand-over points (-DFRN_GEMM_TS, tools/prof_gemm.py TIMELINE) showed one
// issuing thread spending ~900 cycles per stage on the 12 MMAs and another ~900 on the commits, the loop turn-around
// and the next barrier wait, with the tensor pipe (1024 cycles of work per stage) running dry in between.  Issuer w
// takes the stages it = w (mod 2) of the CTA's global stage sequence and accumulates chains of kChunk2 of ITS stages in accumulator buffer w,
// so each issuer's bookkeeping runs under the other's MMAs, the two never touch the same accumulator, and the drain
// warps empty buffer 0 while buffer 1 fills (and vice versa).
constexpr int kV2Threads = 19 * 32;
constexpr int kV2SplitWarp0 = 2, kV2EpiWarp0 = 10, kV2Issuer1 = 18;
// Accumulation chain of the v2 kernel, in stages of ONE issuer: 4 stages = 16 truncating additions per chain, as in
// the first kernel, and together the two buffers hold 8 stages (a whole K = 256 tile), so the issuers run on while
// the drain warps store the previous tile.  (2 stages per chain: the tile-end store stalled both issuers, 149 -> 168 us
// on the 256 x 19200 x 256 convolution.)
constexpr int kChunk2 = 4;

// FAST (single product on tf32-exact operands, the bf16 inference path): no lo tiles, so a stage is half the size and
// the ring twice as deep -- this mode has nothing but the TMA latency to hide (4 MMAs per stage).
template <int BN, bool FAST> struct Tc2Cfg {
  static constexpr int kPlanes = FAST ? 1 : 2;
  static constexpr int kStages = FAST ? (BN == 128 ? 6 : 8) : (BN == 128 ? 3 : 4);
  static constexpr int kABytes = TBM * TBK * 4;            // hi or lo
  static constexpr int kBBytes = BN * TBK * 4;
  static constexpr int kStageBytes = kPlanes * (kABytes + kBBytes);     // A hi | A lo | B hi | B lo  (FAST: A | B)
  static constexpr int kEpiBytes = 8 * 4096;
  static constexpr int kSmem = kStages * kStageBytes + kEpiBytes + 1024 /*align*/ + 1024 /*barriers, scratch*/;
  static constexpr int kTmemCols = 4 * BN;
};

// shared-memory descriptor with explicit layout type / LBO / SBO (bytes)
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}
// lo part of the truncating split, rounded to nearest at tf32 precision
__device__ __forceinline__ float lo_tf32(float x) {
  const float hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  const float l = x - hi;                                   // exact
  return __uint_as_float((__float_as_uint(l) + 0x1000u) & 0xFFFFE000u);
}

// Every lane polls (one polling lane + __syncwarp was measured 14 % slower on the 256 x 19200 x 256 convolution).
__device__ __forceinline__ void mbar_wait_warp(uint64_t* bar, uint32_t parity, int lane) {
  (void)lane;
#ifdef FRN_GEMM_SPIN
  mbar_spin(bar, parity);
#else
  mbar_wait(bar, parity);
#endif
}

struct V2Args {
  int ktiles, tiles_per_split, nsplit, gx, gy, gz;
  int a_mn, b_mn;            // operand is MN-major (its non-K dimension is the contiguous one)
  uint32_t mn_layout, mn_sbo;   // descriptor layout type and 4-row (or 8-row) group stride of an MN-major tile
  int dbg;                   // -DFRN_GEMM_TRACE builds only (env FRN_GEMM_DBG): timing experiments that break the result
};
#if defined(FRN_GEMM_TRACE) || defined(FRN_GEMM_KNOBS)
#define DBG(bit) ((g.dbg >> (bit)) & 1)
#else
#define DBG(bit) 0
#endif

template <int BN, int EPI, bool ACC, bool FAST>
__global__ void __launch_bounds__(kV2Threads, 1)
tc_gemm2_kernel(GemmDesc p, float* __restrict__ aux, V2Args g, const __grid_constant__ CUtensorMap tmA,
                const __grid_constant__ CUtensorMap tmB) {
  using Cfg = Tc2Cfg<BN, FAST>;
  constexpr int HC = BN / 2;
  TRACE_KERNEL_BEGIN();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* ctrl = smem + Cfg::kStages * Cfg::kStageBytes + Cfg::kEpiBytes;
  uint64_t* raw_full = reinterpret_cast<uint64_t*>(ctrl);
  uint64_t* lo_full = raw_full + Cfg::kStages;
  uint64_t* empty = lo_full + Cfg::kStages;
  uint64_t* acc_full = empty + Cfg::kStages;     // 2
  uint64_t* acc_free = acc_full + 2;             // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_free + 2);
  float* red = reinterpret_cast<float*>(tmem_slot + 4);   // 128 floats

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total = g.gx * g.gy * g.gz;
  const int ktiles = g.ktiles;

  if (tid == 0) {
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(&raw_full[s], 1); mbar_init(&lo_full[s], 8); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; s++) { mbar_init(&acc_full[s], 1); mbar_init(&acc_free[s], 8); }
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB);
    mbar_init_fence();
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;
  const uint32_t stage_base = smem_u32(smem);

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      long it = 0;
      for (int id = blockIdx.x; id < total; id += gridDim.x) {
        const TcTile tl = tc_tile(p, id, g.gx, g.gy, ktiles, g.tiles_per_split, BN);
        int b = 0, kt = 0;
        if (tl.b_fixed >= 0) b = tl.b_fixed;
        else { b = tl.t_begin / ktiles; kt = tl.t_begin - b * ktiles; }
        for (int i = 0; i < tl.ntiles; i++, it++) {
          const int stage = (int)(it % Cfg::kStages);
          const uint32_t phase = (uint32_t)((it / Cfg::kStages) & 1);
          { TRACE_T0();
#ifdef FRN_GEMM_SPIN
          mbar_spin(&empty[stage], phase ^ 1);
#else
          mbar_wait(&empty[stage], phase ^ 1);
#endif
          TRACE_ADD(0); }
          TS(0, it);
          mbar_expect_tx(&raw_full[stage], (DBG(5) ? 0 : Cfg::kABytes) + (DBG(6) ? 0 : Cfg::kBBytes));
          const uint32_t sA = stage_base + (uint32_t)stage * Cfg::kStageBytes, sB = sA + Cfg::kPlanes * Cfg::kABytes;
          const int k0 = kt * TBK, ba = p.sAb ? b : 0, bb = p.sBb ? b : 0;
          if (DBG(5)) {
          } else if (!g.a_mn) tma_load_3d(sA, &tmA, k0, tl.m0, ba, &raw_full[stage]);
          else
#pragma unroll
            for (int j = 0; j < TBM / 32; j++) tma_load_3d(sA + j * 4096, &tmA, tl.m0 + 32 * j, k0, ba, &raw_full[stage]);
          if (DBG(6)) {
          } else if (!g.b_mn) tma_load_3d(sB, &tmB, k0, tl.n0, bb, &raw_full[stage]);
          else
#pragma unroll
            for (int j = 0; j < BN / 32; j++) tma_load_3d(sB + j * 4096, &tmB, tl.n0 + 32 * j, k0, bb, &raw_full[stage]);
          TS(1, it);
          kt++;
          if (tl.b_fixed < 0 && kt == ktiles) { kt = 0; b++; }
        }
      }
    }
  } else if (warp == 1 || warp == kV2Issuer1) {
    // ===== MMA issue (issuer w: global stages it = w mod 2, accumulator buffer w) =====
    const int w = warp == 1 ? 0 : 1;
    // [b_hi | b_lo] are adjacent in shared memory (both majors), so ONE instruction with N = 2 BN forms
    // a_hi.[b_hi | b_lo] = [hh | hl] in the column halves [main | small] of the accumulator buffer; a_lo.b_hi and
    // a_lo.b_lo follow as two N = BN instructions into the small half.  Same tensor time as four N = BN instructions
    // (N / 2 cycles each), but the a_hi slice crosses the shared-memory port once instead of twice, and the main
    // accumulator still receives ONE truncating addition per K slice (with a_lo.[b_hi | b_lo] as a second N = 2 BN
    // instruction it received two, and the phi weight gradient of the materialised attention at DSEC level 3 -- the
    // canary for truncation bias -- went from under the 1e-5 gate to 1.03e-5).
    const uint32_t majors = ((uint32_t)g.a_mn << 15) | ((uint32_t)g.b_mn << 16);
    const uint32_t idesc2 = umma_idesc_tf32(TBM, 2 * BN) | majors, idesc1 = umma_idesc_tf32(TBM, BN) | majors;
    // Descriptors of (stage 0, K slice 0).  Every other descriptor differs only in the 14-bit start-address field
    // (bytes >> 4; shared memory ends below 256 KB, so the additions never carry out of the field): the issuing
    // thread spends a handful of integer additions per K slice.  (Rebuilding the four descriptors from scratch for
    // every slice, with the major-dependent branches, was ~1000 cycles of serial scalar code per stage -- as much
    // as the MMAs themselves, measured with the tensor work switched off: tools/prof_gemm.py SWEEP.)
    const uint32_t sA0 = stage_base, sB0 = stage_base + Cfg::kPlanes * Cfg::kABytes;
    const uint64_t dA0 = g.a_mn ? umma_desc(sA0, 4096, g.mn_sbo, g.mn_layout) : umma_desc_sw128(sA0);
    const uint64_t dB0 = g.b_mn ? umma_desc(sB0, 4096, g.mn_sbo, g.mn_layout) : umma_desc_sw128(sB0);
    // K-major: + 32 bytes per K = 8 slice inside the 128-byte rows; MN-major: + 8 rows of 128 bytes
    const uint32_t ka = (g.a_mn ? 1024u : 32u) >> 4, kb = (g.b_mn ? 1024u : 32u) >> 4;
    constexpr bool split = !FAST;
    const bool lolo = !p.no_lolo;
    long it = 0;
    int gc = 0;                       // chains this issuer has finished (its buffer's phase counter)
    int cnt = 0;                      // stages in this issuer's open chain
    for (int id = blockIdx.x; id < total; id += gridDim.x) {
      const TcTile tl = tc_tile(p, id, g.gx, g.gy, ktiles, g.tiles_per_split, BN);
      for (int i = 0; i < tl.ntiles; i++, it++) {
        const int stage = (int)(it % Cfg::kStages);
        const uint32_t phase = (uint32_t)((it / Cfg::kStages) & 1);
        if ((int)(it & 1) != w) {
          // Not this issuer's stage, but it still has to SEE the stage complete: a parity wait only works for a
          // waiter that is at most one phase behind.  Stages are owned by the parity of the GLOBAL stage counter, so
          // every issuer owns every other stage whatever the tiles' K is, the in-order producer can never get more
          // than a ring ahead of either issuer's last commit, and neither can miss a phase of any slot.  (Without
          // these waits the second issuer passed a slot's stale parity in the 3-deep ring and consumed a stage that
          // had not been loaded: a launch failure in the single-product mode.)
          mbar_wait_warp(split ? &lo_full[stage] : &raw_full[stage], phase, lane);
          continue;
        }
        const bool first = cnt == 0;
        const bool last = cnt == kChunk2 - 1 || i + 2 >= tl.ntiles;    // chain full, or this issuer's last stage of the tile
        cnt = last ? 0 : cnt + 1;
        { TRACE_T0();
        if (first && gc >= 1) mbar_wait_warp(&acc_free[w], (gc - 1) & 1, lane);
        TRACE_ADD(4); }
        // the split warps arrive on lo_full only after they have seen raw_full complete, so lo_full covers both
        { TRACE_T0();
        mbar_wait_warp(split ? &lo_full[stage] : &raw_full[stage], phase, lane);
        TRACE_ADD(3); }
        tc_fence_after();
        TS(4, it);
        TRACE_T0();
        if (elect_one()) {
          const uint32_t d_acc = tmem_acc + w * 2 * BN;
          const uint64_t so = (uint64_t)(((uint32_t)stage * Cfg::kStageBytes) >> 4);
          uint64_t a_hi = dA0 + so, b_both = dB0 + so;
#pragma unroll
          for (int k = 0; k < TBK / 8; k++, a_hi += ka, b_both += kb) {
            const bool acc = !(first && k == 0);
            if (split) {
              const uint64_t a_lo = a_hi + (Cfg::kABytes >> 4);
              if (!DBG(3)) umma_tf32_ss(d_acc, a_hi, b_both, idesc2, acc);                       // [hh | hl]
              if (!DBG(2) && !DBG(3)) umma_tf32_ss(d_acc + BN, a_lo, b_both, idesc1, true);      // lh (b_both starts at b_hi)
              if (lolo && !DBG(1) && !DBG(2) && !DBG(3))
                umma_tf32_ss(d_acc + BN, a_lo, b_both + (Cfg::kBBytes >> 4), idesc1, true);      // ll
            } else {
              umma_tf32_ss(d_acc, a_hi, b_both, idesc1, acc);
            }
          }
          TS_ANY(6, it);
          umma_commit(&empty[stage]);
          if (last) umma_commit(&acc_full[w]);
          TS_ANY(7, it);
        }
        __syncwarp();
        TS(5, it);
        TRACE_ADD(6);
        if (last) gc++;
      }
    }
  } else if (warp < kV2EpiWarp0) {
    // ===== lo split: same byte offsets as the raw tile, whatever its layout =====
    if (!FAST) {
      const int t = tid - kV2SplitWarp0 * 32;          // 0..255
      long it = 0;
      for (int id = blockIdx.x; id < total; id += gridDim.x) {
        const TcTile tl = tc_tile(p, id, g.gx, g.gy, ktiles, g.tiles_per_split, BN);
        for (int i = 0; i < tl.ntiles; i++, it++) {
          const int stage = (int)(it % Cfg::kStages);
          const uint32_t phase = (uint32_t)((it / Cfg::kStages) & 1);
          { TRACE_T0();
          mbar_wait_warp(&raw_full[stage], phase, lane);
          TRACE_ADD(1); }
          if (warp == kV2SplitWarp0) TS(2, it);
          TRACE_T0();
          const uint32_t sA = stage_base + (uint32_t)stage * Cfg::kStageBytes + (uint32_t)t * 16u;
          const uint32_t sB = sA + Cfg::kPlanes * Cfg::kABytes;
          float4 va[Cfg::kABytes / 4096], vb[Cfg::kBBytes / 4096];
          if (!DBG(0)) {
#pragma unroll
          for (int q = 0; q < Cfg::kABytes / 4096; q++) va[q] = lds128(sA + q * 4096);
#pragma unroll
          for (int q = 0; q < Cfg::kBBytes / 4096; q++) vb[q] = lds128(sB + q * 4096);
#pragma unroll
          for (int q = 0; q < Cfg::kABytes / 4096; q++)
            sts128(sA + Cfg::kABytes + q * 4096, lo_tf32(va[q].x), lo_tf32(va[q].y), lo_tf32(va[q].z), lo_tf32(va[q].w));
#pragma unroll
          for (int q = 0; q < Cfg::kBBytes / 4096; q++)
            sts128(sB + Cfg::kBBytes + q * 4096, lo_tf32(vb[q].x), lo_tf32(vb[q].y), lo_tf32(vb[q].z), lo_tf32(vb[q].w));
          }
          if (!DBG(8)) fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&lo_full[stage]);
          if (warp == kV2SplitWarp0) TS(3, it);
          TRACE_ADD(2);
        }
      }
    }
  } else {
    // ===== accumulator drain + epilogue: thread = output row, HC of the BN columns =====
    const int e = warp - kV2EpiWarp0;
    const int quad = warp & 3, half = e >> 2;
    const uint32_t tlane = ((uint32_t)(quad * 32)) << 16;
    const uint32_t tile = smem_u32(smem + Cfg::kStages * Cfg::kStageBytes) + (uint32_t)e * 4096u;
    int gcb[2] = {0, 0};              // chains drained from each buffer
    int cnt[2] = {0, 0};              // stages in each issuer's open chain
    long it = 0;                      // global stage counter at the start of the tile
    for (int id = blockIdx.x; id < total; id += gridDim.x) {
      const TcTile tl = tc_tile(p, id, g.gx, g.gy, ktiles, g.tiles_per_split, BN);
      float acc[HC];
#pragma unroll
      for (int j = 0; j < HC; j++) acc[j] = 0.f;
      // the chains of this tile in the order they finish, by the issuers' own rule
      for (int i = 0; i < tl.ntiles; i++) {
        const int buf = (int)((it + i) & 1);
        const bool last = cnt[buf] == kChunk2 - 1 || i + 2 >= tl.ntiles;
        cnt[buf] = last ? 0 : cnt[buf] + 1;
        if (!last) continue;
        { TRACE_T0();
        mbar_wait_warp(&acc_full[buf], (uint32_t)(gcb[buf] & 1), lane);
        TRACE_ADD(7); }
        gcb[buf]++;
        tc_fence_after();
        const uint32_t taddr = tmem_acc + tlane + buf * 2 * BN + half * HC;
        if (!DBG(7))
#pragma unroll
        for (int c0 = 0; c0 < HC; c0 += 16) {
          float v[16], w[16];
          tmem_ld16(taddr + c0, v);
          if (!FAST) {
            tmem_ld16(taddr + BN + c0, w);
          } else {
#pragma unroll
            for (int j = 0; j < 16; j++) w[j] = 0.f;
          }
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; j++) acc[c0 + j] += v[j] + w[j];
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_free[buf]);
      }
      it += tl.ntiles;
      TRACE_T0();
      if (!DBG(4)) tc_epilogue<BN, EPI, ACC>(p, aux, tl, acc, tile, red, g.nsplit, g.gx, quad, half, lane);
      TRACE_ADD(8);
    }
  }
  tc_fence_before();
  __syncthreads();
  TRACE_KERNEL_END();
  if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_acc);
}

// C[b][m][n] = sum_j part[b * s + j][m][n]   (per_batch_split products: K split, batch kept)
__global__ void splitk_reduce_grouped_kernel(const float* __restrict__ part, int s, int batch, int M, int N,
                                             float* __restrict__ C, long sCm, long sCn, long sCb) {
  const long mn = (long)M * N, total = mn * batch;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long b = i / mn, r = i - b * mn;
    float acc = 0.f;
    for (int j = 0; j < s; j++) acc += part[(b * s + j) * mn + r];
    const int m = (int)(r / N), n = (int)(r % N);
    C[b * sCb + (long)m * sCm + (long)n * sCn] = acc;
  }
}

struct TcPlan { int bn, ktiles, nsplit, tiles_per_split; bool swapped; GemmDesc d; };

// The UMMA M dimension is fixed at 128: problems with few rows and many columns (the C/8-row
// projections) are run as C^T = B^T A^T, which is only a relabelling of strides.
struct GemmEnv { int tma, v2; uint32_t mn_layout, mn_sbo; CUtensorMapSwizzle mn_swizzle; int lolo; };
const GemmEnv& gemm_env() {
  static const GemmEnv env = [] {
    GemmEnv e{1, 2, 1u, 512u, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, -1};
    if (const char* s = getenv("FRN_GEMM_LOLO")) e.lolo = s[0] != '0';       // 0 / 1: drop / keep lo.lo everywhere (study)
    if (const char* s = getenv("FRN_GEMM_TMA")) e.tma = s[0] != '0';
    if (const char* s = getenv("FRN_GEMM_V2")) e.v2 = s[0] == '0' ? 0 : (s[0] == 'k' ? 1 : 2);
    if (const char* s = getenv("FRN_GEMM_MN")) {
      unsigned a = 1, b = 512, c = 4;
      if (sscanf(s, "%u,%u,%u", &a, &b, &c) >= 2) { e.mn_layout = a; e.mn_sbo = b; e.mn_swizzle = (CUtensorMapSwizzle)c; }
    }
    return e;
  }();
  return env;
}

TcPlan tc_plan(const GemmDesc& in, bool allow_swap) {
  TcPlan pl;
  pl.d = in;
  pl.swapped = false;
  if (gemm_env().lolo >= 0) pl.d.no_lolo = !gemm_env().lolo;
  if (allow_swap && in.M <= 64 && in.N > in.M && in.group_m == 0) {
    GemmDesc& d = pl.d;
    d.A = in.B; d.sAm = in.sBn; d.sAk = in.sBk; d.sAb = in.sBb;
    d.B = in.A; d.sBn = in.sAm; d.sBk = in.sAk; d.sBb = in.sAb;
    d.M = in.N; d.N = in.M;
    d.sCm = in.sCn; d.sCn = in.sCm;
    d.bias_m = in.bias_n; d.bias_n = in.bias_m;
    pl.swapped = true;
  }
  const GemmDesc& d = pl.d;
  pl.bn = d.N <= 32 ? 32 : (d.N <= 64 ? 64 : 128);
  pl.ktiles = cdiv(d.K, TBK);
  pl.nsplit = 1;
  pl.tiles_per_split = pl.ktiles;
  if (d.reduce_batch && d.per_batch_split > 0) {
    pl.tiles_per_split = cdiv(pl.ktiles, d.per_batch_split);
    pl.nsplit = d.batch * d.per_batch_split;
  } else if (d.reduce_batch) {
    long T = (long)d.batch * pl.ktiles;
    long ctas = (long)cdiv(d.M, TBM) * cdiv(d.N, pl.bn);
    long want = cdiv(2 * 148, ctas);
    long min_split = cdiv(T, 1024);             // <= 32768 k per partial (chunk sums are added in fp32 registers)
    long max_split = T / 4 > 0 ? T / 4 : 1;     // >= 128 k per split
    long ns = want < min_split ? min_split : want;
    if (ns > max_split) ns = max_split;
    if (ns < min_split) ns = min_split;
    if (ns < 1) ns = 1;
    pl.tiles_per_split = cdiv(T, ns);
    pl.nsplit = cdiv(T, pl.tiles_per_split);
  }
  return pl;
}

// ---- TMA tensor maps for the operands -------------------------------------------------
// An operand is a (rows, K, batch) view with element strides (rs, ks, bs), one of rs / ks being 1.  TMA needs
// a 16-byte aligned base and byte strides that are multiples of 16; anything else (odd map sizes such as
// 17 positions) takes the cp.async path of the same kernel.
PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(f);
  }();
  return fn;
}

bool make_operand_map(CUtensorMap* tm, const float* base, int rows, int K, int batch, long rs, long ks, long bs,
                      int box_rows) {
  auto enc = tensor_map_encoder();
  if (!enc) return false;
  if (reinterpret_cast<uintptr_t>(base) & 15) return false;
  const bool kfast = (ks == 1);
  const long os = kfast ? rs : ks;                       // stride of the non-contiguous axis
  const int nb = (bs == 0 || batch == 1) ? 1 : batch;
  if (os <= 0 || (os & 3) || (nb > 1 && (bs <= 0 || (bs & 3)))) return false;
  cuuint64_t dims[3] = {(cuuint64_t)(kfast ? K : rows), (cuuint64_t)(kfast ? rows : K), (cuuint64_t)nb};
  // an unused batch axis still needs a legal stride
  cuuint64_t strides[2] = {(cuuint64_t)os * 4, (cuuint64_t)(nb > 1 ? bs : os * (long)dims[1]) * 4};
  cuuint32_t box[3] = {(cuuint32_t)(kfast ? TBK : box_rows), (cuuint32_t)(kfast ? box_rows : TBK), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  if (strides[0] >= (1ull << 40) || strides[1] >= (1ull << 40)) return false;
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// v2 operand map: the box lands directly in the UMMA tile layout.  K-major (ks == 1): 32 k (128 bytes) x box_rows
// rows, SWIZZLE_128B.  MN-major (rs == 1): 32 rows (128 bytes) x 32 k, SWIZZLE_128B with 32-byte atoms.
bool make_operand_map_v2(CUtensorMap* tm, const float* base, int rows, int K, int batch, long rs, long ks, long bs,
                         int box_rows, CUtensorMapSwizzle mn_swizzle) {
  auto enc = tensor_map_encoder();
  if (!enc) return false;
  if (reinterpret_cast<uintptr_t>(base) & 15) return false;
  const bool kfast = (ks == 1);
  if (!kfast && rs != 1) return false;
  const long os = kfast ? rs : ks;
  const int nb = (bs == 0 || batch == 1) ? 1 : batch;
  if (os <= 0 || (os & 3) || (nb > 1 && (bs <= 0 || (bs & 3)))) return false;
  cuuint64_t dims[3] = {(cuuint64_t)(kfast ? K : rows), (cuuint64_t)(kfast ? rows : K), (cuuint64_t)nb};
  cuuint64_t strides[2] = {(cuuint64_t)os * 4, (cuuint64_t)(nb > 1 ? bs : os * (long)dims[1]) * 4};
  cuuint32_t box[3] = {32u, (cuuint32_t)(kfast ? box_rows : TBK), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  if (strides[0] >= (1ull << 40) || strides[1] >= (1ull << 40)) return false;
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, kfast ? CU_TENSOR_MAP_SWIZZLE_128B : mn_swizzle,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Library-wide GEMM switches, read from the environment ONCE (first launch):
//   FRN_GEMM_TMA=0   cp.async staging in the first kernel (debug / comparison)
//   FRN_GEMM_V2=0    never use the v2 kernel;  FRN_GEMM_V2=k  v2 for K-major operands only
//   FRN_GEMM_MN=layout,sbo,swizzle   MN-major descriptor parameters of the v2 kernel (probing)

template <int BN, int EPI, bool ACC>
int tc_launch_acc(cudaStream_t st, const TcPlan& pl, float* aux) {
  const GemmDesc& d = pl.d;
  const GemmEnv& env = gemm_env();
  const int gx = cdiv(d.N, BN), gy = cdiv(d.M, TBM), gz = d.reduce_batch ? pl.nsplit : d.batch;
  const long total = (long)gx * gy * gz;
  const int ctas = (int)(total < kSmCount ? total : kSmCount);      // persistent: one CTA per SM
  alignas(64) CUtensorMap tmA, tmB;
  memset(&tmA, 0, sizeof tmA);
  memset(&tmB, 0, sizeof tmB);
  const bool a_mn = d.sAk != 1, b_mn = d.sBk != 1;
  // single-pass mode with arbitrary fp32 operands (fast == 1) stays on the first kernel: its staging ROUNDS the
  // operands to tf32, raw operands would be truncated (2 x the error, and biased).  fast == 2: the caller guarantees
  // tf32-exact operands, so the raw tiles go straight to the tensor core.
  if (env.tma && env.v2 && d.fast != 1 && (env.v2 == 2 || (!a_mn && !b_mn)) &&
      make_operand_map_v2(&tmA, d.A, d.M, d.K, d.batch, d.sAm, d.sAk, d.sAb, TBM, env.mn_swizzle) &&
      make_operand_map_v2(&tmB, d.B, d.N, d.K, d.batch, d.sBn, d.sBk, d.sBb, BN, env.mn_swizzle)) {
    const bool fast = d.fast != 0;
    auto kern2 = fast ? tc_gemm2_kernel<BN, EPI, ACC, true> : tc_gemm2_kernel<BN, EPI, ACC, false>;
    const int smem2 = fast ? Tc2Cfg<BN, true>::kSmem : Tc2Cfg<BN, false>::kSmem;
    FRN_TRY(set_smem_once((const void*)kern2, smem2));
    V2Args g{pl.ktiles, pl.tiles_per_split, pl.nsplit, gx, gy, gz, a_mn ? 1 : 0, b_mn ? 1 : 0, env.mn_layout, env.mn_sbo, 0};
#if defined(FRN_GEMM_TRACE) || defined(FRN_GEMM_KNOBS)
    if (const char* s = getenv("FRN_GEMM_DBG")) g.dbg = atoi(s);      // read per launch: tools/prof_gemm.py sweeps it
#endif
    kern2<<<ctas, kV2Threads, smem2, st>>>(d, aux, g, tmA, tmB);
    FRN_LAUNCH_CHECK();
#if defined(FRN_GEMM_KNOBS)
    if (getenv("FRN_GEMM_SYNC")) {          // debugging: name the launch that faults
      cudaError_t e = cudaStreamSynchronize(st);
      fprintf(stderr, "v2 BN=%d EPI=%d ACC=%d M=%d N=%d K=%d batch=%d a_mn=%d b_mn=%d fast=%d reduce=%d pbs=%d group_m=%d ktiles=%d tps=%d nsplit=%d grid=%d: %s\n",
              BN, EPI, (int)ACC, d.M, d.N, d.K, d.batch, (int)a_mn, (int)b_mn, d.fast, d.reduce_batch, d.per_batch_split,
              d.group_m, pl.ktiles, pl.tiles_per_split, pl.nsplit, ctas, cudaGetErrorString(e));
      if (e != cudaSuccess) return (int)e;
    }
#endif
    return FRN_OK;
  }
  using Cfg = TcCfg<BN>;
  auto kern = tc_gemm_kernel<BN, EPI, ACC>;
  FRN_TRY(set_smem_once((const void*)kern, Cfg::kSmem));
  const int use_tma = env.tma &&
                      make_operand_map(&tmA, d.A, d.M, d.K, d.batch, d.sAm, d.sAk, d.sAb, TBM) &&
                      make_operand_map(&tmB, d.B, d.N, d.K, d.batch, d.sBn, d.sBk, d.sBb, BN);
  kern<<<ctas, kTcThreads, Cfg::kSmem, st>>>(d, aux, pl.ktiles, pl.tiles_per_split, pl.nsplit, gx, gy, gz, use_tma,
                                             tmA, tmB);
  FRN_LAUNCH_CHECK();
  return FRN_OK;
}

template <int BN, int EPI>
int tc_launch_bn(cudaStream_t st, const TcPlan& pl, float* aux) {
  // split-K partials are plain stores; the accumulation then happens in splitk_reduce_kernel
  const bool acc = EPI == GEMM_EPI_STORE && pl.d.accumulate && !(pl.d.reduce_batch && pl.nsplit > 1);
  if (EPI == GEMM_EPI_STORE && acc) return tc_launch_acc<BN, GEMM_EPI_STORE, true>(st, pl, aux);
  return tc_launch_acc<BN, EPI, false>(st, pl, aux);
}

template <int EPI>
int tc_launch(cudaStream_t st, const TcPlan& pl, float* aux) {
  if (pl.bn == 32) return tc_launch_bn<32, EPI>(st, pl, aux);
  if (pl.bn == 64) return tc_launch_bn<64, EPI>(st, pl, aux);
  return tc_launch_bn<128, EPI>(st, pl, aux);
}

}  // namespace

int tc_gemm_trace_read(unsigned long long* out16) {
#ifdef FRN_GEMM_TRACE
  unsigned long long z[16] = {0};
  cudaMemcpyFromSymbol(out16, g_trace, sizeof(z));
  cudaMemcpyToSymbol(g_trace, z, sizeof(z));
  return 1;
#else
  (void)out16;
  return 0;
#endif
}

int tc_gemm_ts_read(long long* out) {      // 8 x 256 stamps
#ifdef FRN_GEMM_TS
  cudaMemcpyFromSymbol(out, g_ts, sizeof(long long) * 8 * 256);
  return 1;
#else
  (void)out;
  return 0;
#endif
}

size_t tc_gemm_workspace_bytes(const GemmDesc& d) {
  if (!d.reduce_batch) return 0;
  TcPlan pl = tc_plan(d, true);
  return (pl.nsplit > 1 || d.per_batch_split > 0) ? align_up((size_t)pl.nsplit * d.M * d.N * sizeof(float)) : 0;
}

int tc_gemm(cudaStream_t st, const GemmDesc& in, void* ws, size_t ws_bytes) {
  if (in.M <= 0 || in.N <= 0 || in.K <= 0 || in.batch <= 0) return FRN_ERR_BAD_SHAPE;
  if ((in.sAm != 1 && in.sAk != 1) || (in.sBn != 1 && in.sBk != 1)) return FRN_ERR_UNSUPPORTED;
  if (in.group_m > 0 && (in.reduce_batch || in.accumulate || !in.C1 || !in.C2 || in.M > 3 * in.group_m || in.bias_n))
    return FRN_ERR_UNSUPPORTED;
  TcPlan pl = tc_plan(in, true);
  if (in.group_m > 0) {
    ScopedTimer tm(st, FRN_TIME_GEMM);
    ShapeTimer sh(st, "tc-groups", in.M, in.N, in.K, in.batch, 0);
    pl.bn = 128;
    return tc_launch_acc<128, GEMM_EPI_GROUPS, false>(st, pl, nullptr);
  }
  ScopedTimer tm(st, FRN_TIME_GEMM);
  ShapeTimer sh(st, "tc", in.M, in.N, in.K, in.batch,
                (in.reduce_batch ? 1 : 0) | (in.accumulate ? 2 : 0) | (in.sAk == 1 ? 4 : 0) | (in.sBn == 1 ? 8 : 0));
  if (in.reduce_batch && in.per_batch_split > 0) {
    if (in.accumulate || in.bias_m || in.bias_n) return FRN_ERR_UNSUPPORTED;
    size_t need = (size_t)pl.nsplit * in.M * in.N * sizeof(float);
    if (!ws || ws_bytes < need) return FRN_ERR_WORKSPACE;
    FRN_TRY(tc_launch<GEMM_EPI_STORE>(st, pl, (float*)ws));
    const GemmDesc& d = pl.d;     // partials are (batch * per_batch_split, d.M, d.N) in the (possibly swapped) frame
    const long total = (long)d.batch * d.M * d.N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    splitk_reduce_grouped_kernel<<<blocks, 256, 0, st>>>((const float*)ws, d.per_batch_split, d.batch, d.M, d.N, d.C,
                                                         d.sCm, d.sCn, d.sCb);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  if (in.reduce_batch && pl.nsplit > 1) {
    size_t need = (size_t)pl.nsplit * in.M * in.N * sizeof(float);
    if (!ws || ws_bytes < need) return FRN_ERR_WORKSPACE;
    FRN_TRY(tc_launch<GEMM_EPI_STORE>(st, pl, (float*)ws));
    const GemmDesc& d = pl.d;     // partials are (nsplit, d.M, d.N) in the (possibly swapped) frame
    long total = (long)d.M * d.N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    splitk_reduce_kernel<<<blocks, 256, 0, st>>>((const float*)ws, pl.nsplit, d.M, d.N, d.C, d.sCm, d.sCn,
                                                 d.accumulate);
    FRN_LAUNCH_CHECK();
    return FRN_OK;
  }
  return tc_launch<GEMM_EPI_STORE>(st, pl, nullptr);
}

int tc_gemm_rowsumsq(cudaStream_t st, const GemmDesc& d, float* out_partial) {
  if (d.M <= 0 || d.N <= 0 || d.K <= 0 || d.batch <= 0 || d.reduce_batch) return FRN_ERR_BAD_SHAPE;
  TcPlan pl = tc_plan(d, false);
  pl.bn = 128;                     // gemm_rowsumsq_parts(N) == cdiv(N, 128) column tiles
  ScopedTimer tm(st, FRN_TIME_GEMM);
  ShapeTimer sh(st, "tc-rowsumsq", d.M, d.N, d.K, d.batch, 0);
  return tc_launch_bn<128, GEMM_EPI_ROWSUMSQ>(st, pl, out_partial);
}

}  // namespace frn
