// MUFU.EX2 issue rate per SM sub-partition on this GPU (the exponentials bound the tcgen05 prefill softmax).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/mufu_rate tools/micro/mufu_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
  float v[16];
  for (int i = 0; i < 16; ++i) v[i] = -0.001f * (threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) v[i] = ex2(v[i]) - 1.0f;                 // MUFU + FADD
      else v[i] = fmaf(v[i], 0.999f, -0.001f) * 1.0001f;      // FFMA + FMUL (no MUFU)
    }
  }
  const long long t1 = clock64();
  float s = 0; for (int i = 0; i < 16; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x % 32 == 0) cyc[blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32] = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 32 * 8);
  const int iters = 2000;
  for (int warps : {4, 8, 16, 32}) {
    for (int mode = 0; mode < 2; ++mode) {
      if (mode == 0) k<0><<<148, warps * 32>>>(out, cyc, iters); else k<1><<<148, warps * 32>>>(out, cyc, iters);
      cudaDeviceSynchronize();
      long long h[148 * 32]; cudaMemcpy(h, cyc, 148 * warps * 8, cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < 148 * warps; ++i) mx = h[i] > mx ? h[i] : mx;
      const double per_sm_per_clk = (double)warps * 32 * 16 * iters / (double)mx;
      printf("%s warps/SM %2d: %lld cycles, %.2f ops/clk/SM (%.2f cycles per warp instruction per scheduler)\n",
             mode == 0 ? "MUFU.EX2+FADD" : "FFMA+FMUL    ", warps, mx, per_sm_per_clk, 32.0 * 4 / per_sm_per_clk);
    }
  }
  return 0;
}
