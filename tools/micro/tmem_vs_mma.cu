// Latency of tcgen05.st / tcgen05.ld (issue -> wait::st / wait::ld) while another warp keeps the tensor pipe busy
// with tcgen05.mma streams of different batch lengths.  Question behind it: does a softmax thread's P store queue
// behind the MMAs already issued?
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/tmem_vs_mma tools/micro/tmem_vs_mma.cu
#include <cstdio>
#include "../../light_vllm_b200/csrc/umma.cuh"
using namespace b200;
using namespace b200::umma;

__global__ void __launch_bounds__(288, 1) k(long long* res, int iters, int mma_batch, int ts_form) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + 65536, slot = bar + 16, flag = bar + 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0x3c003c00u;
  __shared__ int fin;
  if (threadIdx.x == 0) { fin = 0; mbar_init(bar, 1); mbar_fence_init(); asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(0)); }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  uint32_t tmem; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  if (warp == 8) {
    if (mma_batch > 0) {
      const uint32_t idesc = idesc_f16(0, 128, 128);
      const uint64_t da = sw128_desc(base), db = sw128_desc(base + 32768);
      uint32_t done = 0; int batches = 0;
      while (!done && batches < 1000000) {
        if (elect_one()) {
          for (int i = 0; i < mma_batch; ++i) {
            const uint32_t adv = (uint32_t)((i >> 2) & 1) * 16384u + (uint32_t)(i & 3) * 32u;
            if (ts_form) mma_ts(tmem + 256, tmem + 384 + (i & 7) * 8, desc_advance(db, adv), idesc, 1);
            else mma_ss(tmem + 256, desc_advance(da, adv), desc_advance(db, adv), idesc, 1);
          }
          commit(bar);
        }
        __syncwarp();
        mbar_wait_wd(bar, (uint32_t)(batches & 1));
        ++batches;
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(done) : "r"(flag));
      }
    }
  } else if (warp < 4) {
    // softmax-like warps: store 32 columns of "P" to columns [0, 128) (not touched by the MMAs), then load them back
    const uint32_t t = tmem + ((uint32_t)(warp * 32) << 16) + 0;
    uint32_t r[32];
    for (int i = 0; i < 32; ++i) r[i] = i + lane;
    long long st_sum = 0, ld_sum = 0, st_max = 0;
    for (int it = 0; it < iters; ++it) {
      long long t0 = clock64();
      tmem_st32(t, r);
      tmem_wait_st();
      long long t1 = clock64();
      tmem_ld32(t, r);
      tmem_wait_ld();
      long long t2 = clock64();
      st_sum += t1 - t0; ld_sum += t2 - t1; st_max = (t1 - t0) > st_max ? (t1 - t0) : st_max;
      for (int spin = 0; spin < 20; ++spin) r[spin] = r[spin] * 3 + 1;   // a little work between probes
    }
    if (lane == 0) { res[blockIdx.x * 16 + warp * 3] = st_sum / iters; res[blockIdx.x * 16 + warp * 3 + 1] = ld_sum / iters; res[blockIdx.x * 16 + warp * 3 + 2] = st_max; }
    if (lane == 0 && atomicAdd(&fin, 1) == 3) asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(flag), "r"(1));
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) { fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* res; cudaMalloc(&res, 148 * 16 * 8);
  const int smem = 65536 + 1024 + 64;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int ts = 0; ts < 2; ++ts)
    for (int batch : {0, 1, 4, 8, 24, 64}) {
      cudaMemset(res, 0, 148 * 16 * 8);
      k<<<148, 288, smem>>>(res, 2000, batch, ts);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long h[16]; cudaMemcpy(h, res, sizeof(h), cudaMemcpyDeviceToHost);
      printf("%s MMA batch %2d (128x128x16 each): tcgen05.st+wait %lld cycles (max %lld), tcgen05.ld+wait %lld cycles\n",
             ts ? "TS" : "SS", batch, h[0], h[2], h[1]);
      fflush(stdout);
    }
  return 0;
}
