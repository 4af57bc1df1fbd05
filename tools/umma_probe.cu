// Correctness probes for the tcgen05 operand paths the round-2 GEMM relies on (run on the B200 box):
//   1. what kind::tf32 does with the 13 low mantissa bits of a 32-bit operand (truncate / round), for A and B;
//   2. MN-major operands (the non-K dimension contiguous in memory) fetched by TMA with SWIZZLE_128B boxes of
//      128 bytes x KT rows, for kind::tf32 and kind::f16 (bf16), in all four major combinations, against an
//      exact integer-valued host product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I frn_b200/csrc tools/umma_probe.cu -o tools/build/umma_probe -lcuda
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "tc05.cuh"

using namespace frn::tc;

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void umma_f16_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc),
      "r"((uint32_t)acc)
      : "memory");
}
// shared-memory descriptor, SWIZZLE_128B, explicit LBO / SBO (bytes)
__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout = 2) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         (1ull << 46) | ((uint64_t)layout << 61);
}

// ---- probe 1: low mantissa bits ------------------------------------------------------------------
// A (128 x 8, K-major): A[m][0] = xa[m], else 0.  B (128 x 8, K-major): B[n][0] = xb[n], else 0.  D[m][n] = xa[m] * xb[n].
__global__ void __launch_bounds__(128, 1) probe_bits(const float* xa, const float* xb, float* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* A = reinterpret_cast<float*>(smem);
  float* B = reinterpret_cast<float*>(smem + 16384);
  for (int i = tid; i < 8192; i += 128) reinterpret_cast<float*>(smem)[i] = 0.f;
  __syncthreads();
  A[sw128_chunk_offset(tid, 0)] = xa[tid];
  B[sw128_chunk_offset(tid, 0)] = xb[tid];
  if (tid == 0) { mbar_init(&bar, 1); mbar_init_fence(); }
  if (warp == 0) tmem_alloc<128>(&slot);
  tc_fence_before();
  fence_proxy_async_smem();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 0) {
    if (elect_one()) {
      umma_tf32_ss(tmem, umma_desc_sw128(smem_u32(A)), umma_desc_sw128(smem_u32(B)), umma_idesc_tf32(128, 128), false);
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < 128; c0 += 16) {
    float v[16];
    tmem_ld16(tmem + (((uint32_t)(warp * 32)) << 16) + c0, v);
    tmem_wait_ld();
    for (int j = 0; j < 16; j++) out[(warp * 32 + lane) * 128 + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<128>(tmem);
}

// ---- probe 2: TMA-fed operands in both majors -------------------------------------------------------
// es: element size (4: tf32, 2: bf16).  KT = 128 / es * ... : one 128-byte swizzle row of K for K-major = 32 (tf32) / 64 (bf16).
// MN-major: blocks of W = 128 / es non-K elements x KT k rows.
__global__ void __launch_bounds__(128, 1) probe_major(const __grid_constant__ CUtensorMap tmA,
                                                      const __grid_constant__ CUtensorMap tmB, int es, int a_mn, int b_mn,
                                                      float* out, uint32_t mn_layout = 2, uint32_t mn_lbo = 0,
                                                      uint32_t mn_sbo = 1024) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar, full;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = 128 / es, KT = W, KI = 32 / es;          // 32 bytes of K per instruction
  const uint32_t sA = smem_u32(smem), sB = sA + 128 * 128;
  if (tid == 0) { mbar_init(&bar, 1); mbar_init(&full, 1); mbar_init_fence(); }
  if (warp == 0) tmem_alloc<128>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (tid == 0) {
    mbar_expect_tx(&full, 2 * 128 * 128);
    if (a_mn) for (int j = 0; j < 128 / W; j++) tma_load_2d(sA + j * KT * 128, &tmA, j * W, 0, &full);
    else tma_load_2d(sA, &tmA, 0, 0, &full);
    if (b_mn) for (int j = 0; j < 128 / W; j++) tma_load_2d(sB + j * KT * 128, &tmB, j * W, 0, &full);
    else tma_load_2d(sB, &tmB, 0, 0, &full);
  }
  mbar_wait(&full, 0);
  tc_fence_after();
  if (warp == 0) {
    if (elect_one()) {
      // instruction descriptor: fp32 accumulate, formats, majors (bit 15: A, bit 16: B), N, M
      const uint32_t fmt = es == 4 ? 2u : 1u;      // TF32 : BF16
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
                             ((128u >> 3) << 17) | ((128u >> 4) << 24);
      for (int s = 0; s < KT / KI; s++) {
        // K-major: 8-row atoms 1024 B apart (SBO), + 32 bytes per K slice.  MN-major: W-wide blocks KT * 128 B apart
        // (LBO), 8-k groups 1024 B apart (SBO), one K slice = KI rows of 128 bytes.
        const uint32_t lbo = mn_lbo ? mn_lbo : KT * 128;
        const uint64_t da = a_mn ? desc_sw128(sA + s * KI * 128, lbo, mn_sbo, mn_layout) : desc_sw128(sA + s * 32, 16, 1024);
        const uint64_t db = b_mn ? desc_sw128(sB + s * KI * 128, lbo, mn_sbo, mn_layout) : desc_sw128(sB + s * 32, 16, 1024);
        if (es == 4) umma_tf32_ss(tmem, da, db, idesc, s != 0);
        else umma_f16_ss(tmem, da, db, idesc, s != 0);
      }
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < 128; c0 += 16) {
    float v[16];
    tmem_ld16(tmem + (((uint32_t)(warp * 32)) << 16) + c0, v);
    tmem_wait_ld();
    for (int j = 0; j < 16; j++) out[(warp * 32 + lane) * 128 + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<128>(tmem);
}

static PFN_cuTensorMapEncodeTiled_v12000 encoder() {
  void* f = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
  return reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(f);
}

static uint16_t to_bf16(float f) { uint32_t u; memcpy(&u, &f, 4); return (uint16_t)(u >> 16); }

int main() {
  float* dout;
  cudaMalloc(&dout, 128 * 128 * 4);
  std::vector<float> out(128 * 128);
  // ---- probe 1 ----
  {
    std::vector<float> xa(128), xb(128);
    for (int i = 0; i < 128; i++) {
      xa[i] = 1.f + (float)(i + 1) * ldexpf(1.f, -16);      // bits below the tf32 lsb (2^-10) set in many patterns
      xb[i] = 1.f;
    }
    xb[1] = 1.f + ldexpf(1.f, -11) + ldexpf(1.f, -12);       // B side: 1 + 0.75 ulp(tf32)
    xb[2] = 1.f + ldexpf(1.f, -11);                          // exactly half an ulp
    xb[3] = 1.f + ldexpf(1.f, -11) - ldexpf(1.f, -20);       // just below half
    float *da, *db;
    cudaMalloc(&da, 512); cudaMalloc(&db, 512);
    cudaMemcpy(da, xa.data(), 512, cudaMemcpyHostToDevice);
    cudaMemcpy(db, xb.data(), 512, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe_bits, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
    probe_bits<<<1, 128, 40 * 1024>>>(da, db, dout);
    cudaError_t e = cudaMemcpy(out.data(), dout, out.size() * 4, cudaMemcpyDeviceToHost);
    printf("probe_bits: %s\n", cudaGetErrorString(e));
    int n_trunc = 0, n_rna = 0, n_other = 0;
    for (int m = 0; m < 128; m++) {
      uint32_t u; memcpy(&u, &xa[m], 4);
      uint32_t t = u & 0xFFFFE000u, r = (u + 0x1000u) & 0xFFFFE000u;
      float ft, fr; memcpy(&ft, &t, 4); memcpy(&fr, &r, 4);
      float got = out[m * 128 + 0];
      if (got == ft && got != fr) n_trunc++; else if (got == fr && got != ft) n_rna++; else if (got != ft) n_other++;
      if (m < 8 || got != ft) if (m < 40) printf("  A[%3d] = %.9g  trunc %.9g  rna %.9g  got %.9g\n", m, xa[m], ft, fr, got);
    }
    printf("A operand low bits: truncated %d, rounded-to-nearest %d, other %d (of the rows where the two differ)\n", n_trunc, n_rna, n_other);
    for (int n = 1; n <= 3; n++) printf("  B[%d] = %.9g  -> D[0][%d] / xa[0]_trunc = %.9g (1 = truncated, 1.0009765625 = rounded up)\n", n, xb[n], n, out[n] / 1.0f);
  }
  // ---- probe 2 ----
  auto enc = encoder();
  if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
  cudaFuncSetAttribute(probe_major, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
  for (int es = 4; es >= 2; es -= 2) {
    const int W = 128 / es, KT = W;
    // integer-valued operands: A (128 x KT), B (128 x KT) as [row][k]; exact product in fp32
    std::vector<float> A(128 * KT), B(128 * KT), ref(128 * 128);
    srand(1234 + es);
    for (auto& v : A) v = (float)(rand() % 9 - 4);
    for (auto& v : B) v = (float)(rand() % 9 - 4);
    for (int m = 0; m < 128; m++)
      for (int n = 0; n < 128; n++) {
        float s = 0.f;
        for (int k = 0; k < KT; k++) s += A[m * KT + k] * B[n * KT + k];
        ref[m * 128 + n] = s;
      }
    // tf32 MN-major variants: {TMA swizzle mode, descriptor layout type, LBO (0 = block stride), SBO}
    struct Var { CUtensorMapSwizzle sw; uint32_t layout, lbo, sbo; const char* name; };
    const Var vars[] = {
        {CU_TENSOR_MAP_SWIZZLE_128B, 2, 0, 1024, "SW128 / layout 2 / SBO 1024"},
        {CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, 1, 0, 512, "SW128_ATOM_32B / layout 1 / SBO 512"},
        {CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, 1, 0, 1024, "SW128_ATOM_32B / layout 1 / SBO 1024"},
        {CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, 1, 512, 4096, "SW128_ATOM_32B / layout 1 / LBO 512 SBO 4096"},
        {CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, 1, 0, 256, "SW128_ATOM_32B / layout 1 / SBO 256"},
        {CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, 2, 0, 512, "SW128_ATOM_32B / layout 2 / SBO 512"},
    };
    const int nvar = es == 4 ? 6 : 1;
    for (int vi = 0; vi < nvar; vi++)
    for (int a_mn = 0; a_mn < 2; a_mn++)
      for (int b_mn = 0; b_mn < 2; b_mn++) {
        const Var& V = vars[vi];
        if (vi > 0 && !a_mn && !b_mn) continue;
        // global images: K-major = [row][k] (k contiguous); MN-major = [k][row] (row contiguous)
        auto image = [&](const std::vector<float>& X, int mn) {
          std::vector<uint8_t> img((size_t)128 * KT * es);
          for (int r = 0; r < 128; r++)
            for (int k = 0; k < KT; k++) {
              const size_t idx = mn ? (size_t)k * 128 + r : (size_t)r * KT + k;
              if (es == 4) memcpy(&img[idx * 4], &X[r * KT + k], 4);
              else { uint16_t h = to_bf16(X[r * KT + k]); memcpy(&img[idx * 2], &h, 2); }
            }
          return img;
        };
        std::vector<uint8_t> ia = image(A, a_mn), ib = image(B, b_mn);
        void *ga, *gb;
        cudaMalloc(&ga, ia.size()); cudaMalloc(&gb, ib.size());
        cudaMemcpy(ga, ia.data(), ia.size(), cudaMemcpyHostToDevice);
        cudaMemcpy(gb, ib.data(), ib.size(), cudaMemcpyHostToDevice);
        auto make = [&](void* g, int mn, CUtensorMap* tm) {
          const CUtensorMapSwizzle swz = mn ? V.sw : CU_TENSOR_MAP_SWIZZLE_128B;
          cuuint64_t dims[2] = {(cuuint64_t)(mn ? 128 : KT), (cuuint64_t)(mn ? KT : 128)};
          cuuint64_t strides[1] = {(cuuint64_t)(mn ? 128 : KT) * es};
          cuuint32_t box[2] = {(cuuint32_t)W, (cuuint32_t)(mn ? KT : 128)};
          cuuint32_t estr[2] = {1, 1};
          return enc(tm, es == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, g, dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        };
        alignas(64) CUtensorMap tmA, tmB;
        CUresult ra = make(ga, a_mn, &tmA), rb = make(gb, b_mn, &tmB);
        cudaMemset(dout, 0xff, 128 * 128 * 4);
        probe_major<<<1, 128, 40 * 1024>>>(tmA, tmB, es, a_mn, b_mn, dout, V.layout, V.lbo, V.sbo);
        cudaError_t e = cudaMemcpy(out.data(), dout, out.size() * 4, cudaMemcpyDeviceToHost);
        double worst = 0; int bad = 0;
        for (int i = 0; i < 128 * 128; i++) {
          double d = fabs((double)out[i] - ref[i]);
          if (!(d <= 0)) bad++;
          if (d > worst || d != d) worst = d;
        }
        printf("probe_major %s [%s] A %s-major, B %s-major: encode %d/%d, %s, mismatches %d / 16384, worst |diff| %.3g\n",
               es == 4 ? "tf32" : "bf16", V.name, a_mn ? "MN" : "K", b_mn ? "MN" : "K", (int)ra, (int)rb, cudaGetErrorString(e), bad, worst);
        cudaFree(ga); cudaFree(gb);
      }
  }
  return 0;
}
