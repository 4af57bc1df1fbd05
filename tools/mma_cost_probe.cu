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


// The instruction mix of one dK/dV step: 24 SS products (two accumulators, two different A tiles, hi / lo
// variants) followed by 24 TS products into two more accumulators, all N = 32.
__global__ void __launch_bounds__(128, 1) probe_mix(int R, int variant, long long* out, int ncommit = 0,
                                                    const float* gsrc = nullptr) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ volatile int stop;
  __shared__ uint64_t dummy[4];
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) for (int i = 0; i < 4; i++) mbar_init(&dummy[i], 1);
  const int bg = variant >> 2;       // background traffic from warps 1-3: 1 = tcgen05.ld, 2 = tcgen05.st, 3 = both
  variant &= 3;
  if (tid == 0) stop = 0;
  for (int i = tid; i < (96 * 1024) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) { mbar_init(&bar, 1); mbar_init_fence(); }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  fence_proxy_async_smem();
  __syncthreads();
  if (warp == 0) {
    const uint32_t base = smem_u32(smem);
    // A tiles: K hi, K lo, V hi, V lo (16 KB each); B tiles: Q hi, Q lo, dY hi, dY lo (4 KB each), chan blocks 4 x 4 KB
    const uint64_t kh = umma_desc_sw128(base), kl = umma_desc_sw128(base + 16384), vh = umma_desc_sw128(base + 32768),
                   vl = umma_desc_sw128(base + 49152);
    const uint32_t st = base + 65536;
    const uint64_t qh = umma_desc_sw128(st), ql = umma_desc_sw128(st + 4096), yh = umma_desc_sw128(st + 8192),
                   yl = umma_desc_sw128(st + 12288);
    const uint64_t c0 = umma_desc_sw128(st + 16384), c1 = umma_desc_sw128(st + 20480), c2 = umma_desc_sw128(st + 24576),
                   c3 = umma_desc_sw128(st + 28672);
    constexpr uint32_t id = umma_idesc_tf32(128, 32);
    long long t0 = 0;
    for (int rep = 0; rep < 2; rep++) {
      t0 = clock64();
      if (elect_one()) {
#pragma unroll 1
        for (int r = 0; r < R; r++) {
          const uint32_t dS = tmem + (r & 1) * 32, dP = tmem + 64 + (r & 1) * 32;
          if (variant != 2) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
              umma_tf32_ss(dS, kl + 2 * k, qh + 2 * k, id, k != 0);
              umma_tf32_ss(dP, vl + 2 * k, yh + 2 * k, id, k != 0);
              umma_tf32_ss(dS, kh + 2 * k, ql + 2 * k, id, true);
              umma_tf32_ss(dP, vh + 2 * k, yl + 2 * k, id, true);
            }
#pragma unroll
            for (int k = 0; k < 4; k++) {
              umma_tf32_ss(dS, kh + 2 * k, qh + 2 * k, id, true);
              umma_tf32_ss(dP, vh + 2 * k, yh + 2 * k, id, true);
            }
            if (ncommit >= 1) umma_commit(&dummy[0]);      // sdp_full
          }
          if (variant != 1) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
              umma_tf32_ts(tmem + 256, tmem + 160 + 8 * k, c0 + 2 * k, id, k != 0);
              umma_tf32_ts(tmem + 288, tmem + 224 + 8 * k, c2 + 2 * k, id, k != 0);
              umma_tf32_ts(tmem + 256, tmem + 128 + 8 * k, c1 + 2 * k, id, true);
              umma_tf32_ts(tmem + 288, tmem + 192 + 8 * k, c3 + 2 * k, id, true);
            }
#pragma unroll
            for (int k = 0; k < 4; k++) {
              umma_tf32_ts(tmem + 256, tmem + 128 + 8 * k, c0 + 2 * k, id, true);
              umma_tf32_ts(tmem + 288, tmem + 192 + 8 * k, c2 + 2 * k, id, true);
            }
            if (ncommit >= 2) umma_commit(&dummy[1]);      // pv_done
            if (ncommit >= 3) umma_commit(&dummy[2]);      // q_empty
          }
        }
        umma_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, rep & 1);
      tc_fence_after();
    }
    const long long t1 = clock64();
    if (tid == 0) { *out = t1 - t0; stop = 1; }
  } else if (bg == 5 || bg == 6) {
    // background: warps 1-3 poll an mbarrier that never completes (5: bare try_wait loop, 6: with nanosleep back-off)
    long long n = 0;
    while (!stop) {
      if (mbar_try_wait(&dummy[3], 0)) break;
      if (bg == 6) __nanosleep(100);
      n++;
    }
    if ((tid & 31) == 0) out[warp] = n;
  } else if (bg == 4) {
    // background bulk copies global -> shared (32 KB per trip into an unused region), warp 1 only
    if (warp == 1 && (tid & 31) == 0) {
      long long n = 0;
      uint32_t ph = 0;
      while (!stop) {
        mbar_expect_tx(&dummy[3], 32768);
        for (int i = 0; i < 8; i++) bulk_g2s(smem + 100 * 1024 + i * 4096, gsrc + ((n * 8 + i) % 4096) * 1024, 4096, &dummy[3]);
        mbar_wait(&dummy[3], ph);
        ph ^= 1;
        n++;
      }
      out[1] = n; out[2] = 0; out[3] = 0;
    }
  } else if (bg) {
    const uint32_t tb = tmem + (((uint32_t)(warp * 32)) << 16) + 384;     // columns 384.. are unused by the mix
    float v[16];
    for (int i = 0; i < 16; i++) v[i] = 0.f;
    long long n = 0;
    while (!stop) {
      if (bg & 1) { tmem_ld16(tb, v); tmem_ld16(tb + 16, v); tmem_wait_ld(); }
      if (bg & 2) { tmem_st16(tb + 32, v); tmem_st16(tb + 48, v); tmem_wait_st(); }
      n++;
    }
    if ((tid & 31) == 0) out[warp] = n;
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
  cudaMalloc(&dout, 4 * sizeof(long long));
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
  cudaFuncSetAttribute(probe_mix, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* names[3] = {"24 SS + 24 TS", "24 SS only", "24 TS only"};
  for (int nc = 0; nc <= 3; nc++) {
    probe_mix<<<1, 128, 100 * 1024>>>(512, 0, dout, nc);
    long long c = 0;
    cudaError_t e = cudaMemcpy(&c, dout, sizeof c, cudaMemcpyDeviceToHost);
    printf("dK/dV step mix with %d tcgen05.commit per step: %8.1f cycles / step  (%s)\n", nc, (double)c / 512,
           cudaGetErrorString(e));
  }
  {
    float* gsrc;
    cudaMalloc(&gsrc, 4096 * 4096 + 65536);
    cudaMemset(gsrc, 0, 4096 * 4096 + 65536);
    cudaFuncSetAttribute(probe_mix, cudaFuncAttributeMaxDynamicSharedMemorySize, 140 * 1024);
    for (int v = 0; v < 3; v++) {
      probe_mix<<<1, 128, 140 * 1024>>>(512, v | (4 << 2), dout, 0, gsrc);
      long long c[4] = {0, 0, 0, 0};
      cudaError_t e = cudaMemcpy(c, dout, sizeof c, cudaMemcpyDeviceToHost);
      printf("dK/dV step mix variant %d with background bulk copies: %8.1f cycles / step, %.1f KB copied / step  (%s)\n", v,
             (double)c[0] / 512, 32.0 * c[1] / 512, cudaGetErrorString(e));
    }
  }
  for (int bgm = 5; bgm <= 6; bgm++)
    for (int v = 0; v < 3; v++) {
      probe_mix<<<1, 128, 140 * 1024>>>(512, v | (bgm << 2), dout, 0, nullptr);
      long long c[4] = {0, 0, 0, 0};
      cudaError_t e = cudaMemcpy(c, dout, sizeof c, cudaMemcpyDeviceToHost);
      printf("dK/dV step mix variant %d, warps 1-3 polling an mbarrier%s: %8.1f cycles / step, %.1f polls / step  (%s)\n", v,
             bgm == 6 ? " with nanosleep(100)" : "", (double)c[0] / 512, (double)(c[1] + c[2] + c[3]) / 512,
             cudaGetErrorString(e));
    }
  const char* bgs[4] = {"idle", "tcgen05.ld", "tcgen05.st", "tcgen05.ld + st"};
  for (int bg = 0; bg < 4; bg++)
    for (int v = 0; v < 3; v++) {
      probe_mix<<<1, 128, 100 * 1024>>>(512, v | (bg << 2), dout);
      long long c[4] = {0, 0, 0, 0};
      cudaError_t e = cudaMemcpy(c, dout, sizeof c, cudaMemcpyDeviceToHost);
      printf("dK/dV step mix (%s), warps 1-3 %s: %8.1f cycles / step, background trips / step %.1f  (%s)\n", names[v],
             bgs[bg], (double)c[0] / 512, bg ? (double)(c[1] + c[2] + c[3]) / 512 : 0.0, cudaGetErrorString(e));
    }
  return 0;
}
