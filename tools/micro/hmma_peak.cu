// Ceiling of the legacy tensor path (mma.sync.m16n8k16 bf16 -> f32) on this GPU: independent
// accumulator chains, operands in registers, no memory traffic.  Build: nvcc -O3 -gencode
// arch=compute_100a,code=sm_100a -o hmma_peak hmma_peak.cu ; run: ./hmma_peak
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void __launch_bounds__(256) hmma_loop(float* out, int iters) {
  float acc[CHAINS][4];
#pragma unroll
  for (int c = 0; c < CHAINS; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.f;
  uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int c = 0; c < CHAINS; ++c)
      asm volatile(
          "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
          "{%0,%1,%2,%3};"
          : "+f"(acc[c][0]), "+f"(acc[c][1]), "+f"(acc[c][2]), "+f"(acc[c][3])
          : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < CHAINS; ++c) s += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
  if (s == 123.456f) out[0] = s;
}

template <int CHAINS>
void run(int warps_per_sm, int sms) {
  const int iters = 20000;
  float* out;
  cudaMalloc(&out, 4);
  const int threads = 256;
  const int blocks = sms * warps_per_sm * 32 / threads;
  hmma_loop<CHAINS><<<blocks, threads>>>(out, 100);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  hmma_loop<CHAINS><<<blocks, threads>>>(out, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double flops = (double)blocks * (threads / 32) * iters * CHAINS * 4096.0;
  printf("chains %d, warps/SM %2d: %.1f TFLOP/s\n", CHAINS, warps_per_sm, flops / ms / 1e9);
  cudaFree(out);
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  for (int w : {8, 16, 32, 64}) run<1>(w, sms);
  for (int w : {8, 16, 32}) run<2>(w, sms);
  for (int w : {8, 16, 32}) run<4>(w, sms);
  for (int w : {8, 16}) run<8>(w, sms);
  return 0;
}
