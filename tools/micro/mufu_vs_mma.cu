// Does tcgen05.mma activity slow MUFU / FMA issue on the same SM?  (the softmax warps of the prefill kernel sit
// beside MMAs that read Q and K from shared memory at more than 128 B/clk)
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/mufu_vs_mma tools/micro/mufu_vs_mma.cu
#include <cstdio>
#include "../../light_vllm_b200/csrc/umma.cuh"
using namespace b200;
using namespace b200::umma;

__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// warps 0..nmath-1: math loop (mode 0: MUFU+FADD, 1: FFMA+FMUL); warp 8: MMA stream (mma_mode 0: none, 1: SS N=64,
// 2: SS N=128, 3: TS N=128 (A from TMEM)); block = 288 threads
__global__ void __launch_bounds__(288, 1) k(float* out, long long* cyc, int iters, int mode, int mma_mode, int nmath) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + 65536, slot = bar + 16, flag = bar + 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0x3c003c00u;
  __shared__ int fin;
  if (threadIdx.x == 0) { fin = 0; mbar_init(bar, 1); mbar_fence_init(); asm volatile("st.shared.u32 [%0], %1;" :: "r"(flag), "r"(0)); }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  uint32_t tmem; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  if (warp == 8) {
    if (mma_mode != 0) {
      const int n = mma_mode == 1 ? 64 : 128;
      const uint32_t idesc = idesc_f16(0, 128, n);
      const uint64_t da = sw128_desc(base), db = sw128_desc(base + 32768);
      uint32_t done = 0;
      int batches = 0;
      while (!done && batches < 200000) {
        if (elect_one()) {
          for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
              const uint32_t adv = (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
              if (mma_mode == 3) mma_ts(tmem + 256, tmem + kk * 8, desc_advance(db, adv), idesc, 1);
              else mma_ss(tmem, desc_advance(da, adv), desc_advance(db, adv), idesc, 1);
            }
          commit(bar);
        }
        __syncwarp();
        mbar_wait_wd(bar, (uint32_t)(batches & 1));
        ++batches;
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(done) : "r"(flag));
      }
      if (lane == 0) cyc[blockIdx.x * 16 + 8] = batches;
    }
  } else if (warp < nmath) {
    float v[16];
    for (int i = 0; i < 16; ++i) v[i] = -0.001f * (threadIdx.x + i);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        if (mode == 0) v[i] = ex2f(v[i]) - 1.0f;
        else v[i] = fmaf(v[i], 0.999f, -0.001f) * 1.0001f;
      }
    }
    const long long t1 = clock64();
    float s = 0; for (int i = 0; i < 16; ++i) s += v[i];
    out[blockIdx.x * 256 + threadIdx.x] = s;
    if (lane == 0) cyc[blockIdx.x * 16 + warp] = t1 - t0;
    // the last math warp to finish stops the MMA stream
    if (lane == 0) { if (atomicAdd(&fin, 1) == nmath - 1) asm volatile("st.volatile.shared.u32 [%0], %1;" :: "r"(flag), "r"(1)); }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) { fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&cyc, 148 * 16 * 8);
  const int iters = 1000, smem = 65536 + 1024 + 64;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* mm[] = {"no MMA", "SS 128x64x16 stream", "SS 128x128x16 stream", "TS 128x128x16 stream"};
  for (int nmath : {4, 8})
    for (int mode = 0; mode < 2; ++mode)
      for (int mma_mode = 0; mma_mode < 4; ++mma_mode) {
        cudaMemset(cyc, 0, 148 * 16 * 8);
        k<<<148, 288, smem>>>(out, cyc, iters, mode, mma_mode, nmath);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        long long h[148 * 16]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0, b = 0; for (int i = 0; i < 148; ++i) { for (int w = 0; w < nmath; ++w) mx = h[i * 16 + w] > mx ? h[i * 16 + w] : mx; b += h[i * 16 + 8]; }
        const double rate = (double)nmath * 32 * 16 * iters / (double)mx;
        const double mma_cyc = b ? (double)mx / ((double)b / 148 * 32) : 0;
        printf("%d math warps, %s, %-22s: %.2f ops/clk/SM; MMA stream: %.1f cycles per MMA\n", nmath,
               mode == 0 ? "MUFU.EX2+FADD" : "FFMA+FMUL    ", mm[mma_mode], rate, mma_cyc);
        fflush(stdout);
      }
  return 0;
}
