// Micro-benchmark: issue cost of one tcgen05.mma (M = 128, one 32-byte K slice per row) as a function of
// N, operand source (A from shared memory or from TMEM) and kind (tf32: K = 8, f16: K = 16).
// One CTA, one elected thread issues R back-to-back MMAs, commits, and the cycle counter is read after the
// commit barrier fires.  Operands are zeros (values do not change the timing).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I frn_b200/csrc tools/mma_cost_probe.cu -o tools/build/mma_cost_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc05.cuh"

using namespace frn::tc;

__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_f16_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc),
      "r"((uint32_t)acc)
      : "memory");
}
__device__ __forceinline__ void umma_f16_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc),
      "r"((uint32_t)acc)
      : "memory");
}

// mode: 0 tf32 SS, 1 tf32 TS, 2 f16 SS, 3 f16 TS.  nacc: accumulators used round-robin.
template <int N, int MODE>
__global__ void __launch_bounds__(128, 1) probe(int R, int nacc, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (128 * 128 + 256 * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) { mbar_init(&bar, 1); mbar_init_fence(); }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  fence_proxy_async_smem();
  __syncthreads();
  if (warp == 0) {
    const uint64_t a = umma_desc_sw128(smem_u32(smem)), b = umma_desc_sw128(smem_u32(smem + 128 * 128));
    constexpr uint32_t id = (MODE < 2) ? umma_idesc_tf32(128, N) : idesc_f16(128, N);
    long long t0 = 0;
    for (int rep = 0; rep < 2; rep++) {        // rep 0 warms up
      t0 = clock64();
      if (elect_one()) {
        const uint32_t d0 = tmem, d1 = tmem + (nacc > 1 ? N : 0), at = tmem + 448;
#pragma unroll 1
        for (int r = 0; r < R; r += 8) {       // 8 unrolled issues per trip: the loop itself costs nothing
#pragma unroll
          for (int u = 0; u < 8; u++) {
            const uint32_t d = (u & 1) ? d1 : d0;
            if (MODE == 0) umma_tf32_ss(d, a + 2 * (u & 3), b + 2 * (u & 3), id, true);
            if (MODE == 1) umma_tf32_ts(d, at + 8 * (u & 3), b + 2 * (u & 3), id, true);
            if (MODE == 2) umma_f16_ss(d, a + 2 * (u & 3), b + 2 * (u & 3), id, true);
            if (MODE == 3) umma_f16_ts(d, at + 8 * (u & 3), b + 2 * (u & 3), id, true);
          }
        }
        umma_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, rep & 1);
      tc_fence_after();
    }
    const long long t1 = clock64();
    if (tid == 0) *out = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

template <int N, int MODE>
void run(const char* name, long long* dout) {
  auto k = probe<N, MODE>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int nacc = 1; nacc <= 2; nacc++) {
    const int R = 4096;
    k<<<1, 128, 64 * 1024>>>(R, nacc, dout);
    long long c = 0;
    cudaError_t e = cudaMemcpy(&c, dout, sizeof c, cudaMemcpyDeviceToHost);
    printf("%-10s N=%3d accumulators=%d : %7.1f cycles / MMA  (%s)\n", name, N, nacc, (double)c / R,
           cudaGetErrorString(e));
  }
}

int main() {
  long long* dout;
  cudaMalloc(&dout, sizeof(long long));
  run<32, 0>("tf32 SS", dout);
  run<64, 0>("tf32 SS", dout);
  run<128, 0>("tf32 SS", dout);
  run<256, 0>("tf32 SS", dout);
  run<32, 1>("tf32 TS", dout);
  run<64, 1>("tf32 TS", dout);
  run<128, 1>("tf32 TS", dout);
  run<32, 2>("f16 SS", dout);
  run<64, 2>("f16 SS", dout);
  run<128, 2>("f16 SS", dout);
  run<32, 3>("f16 TS", dout);
  run<64, 3>("f16 TS", dout);
  return 0;
}
