// Bring-up probe for the tcgen05 building blocks of csrc/prefix_prefill_umma.cuh, run on the GPU box:
//   S = Q K^T   (SS form: both operands K-major SWIZZLE_128B tiles gathered with 16-byte cp.async)
//   O = P V     (TS form: P = fp16/bf16(S / 16) written to TMEM by the threads, V^T tile in shared memory)
// checked against a CPU computation.  Not part of the product.  Build:
//   nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -o tools/micro/umma_probe tools/micro/umma_probe.cu
// Result on a B200 (round 2): fp16 S 4.8e-06 / O 1.9e-06, bf16 S 2.3e-06 / O 1.4e-06 max-abs error -> PASS.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#include "../../light_vllm_b200/csrc/umma.cuh"

using namespace b200;
using namespace b200::umma;

template <typename T> T from_f32_host(float v);
template <> __half from_f32_host<__half>(float v) { return __float2half_rn(v); }
template <> __nv_bfloat16 from_f32_host<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
static float to_f32_host(__half v) { return __half2float(v); }
static float to_f32_host(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
__global__ void __launch_bounds__(128) probe_kernel(const T* q, const T* k, const T* vt, float* s_out,
                                                    float* o_out, int fmt) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_s = base, k_s = base + 32768, v_s = base + 65536, bar = base + 98304, slot = bar + 16;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // gather: 128 rows x 16 chunks per operand
  for (int id = tid; id < 128 * 16; id += 128) {
    const int row = id >> 4, c = id & 15;
    const uint32_t off = (uint32_t)(c >> 3) * 16384u + sw128_offset(row, c & 7);
    cp_async16(q_s + off, q + row * 128 + c * 8, 16);
    cp_async16(k_s + off, k + row * 128 + c * 8, 16);
    cp_async16(v_s + off, vt + row * 128 + c * 8, 16);
  }
  cp_async_commit();
  cp_async_wait_all();
  fence_proxy_async_smem();
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  const uint32_t idesc = idesc_f16(fmt, 128, 128);
  if (tid == 0) {
    const uint64_t da = sw128_desc(q_s), db = sw128_desc(k_s);
    for (int kk = 0; kk < 8; ++kk) {
      const uint32_t adv = (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
      mma_ss(tmem, desc_advance(da, adv), desc_advance(db, adv), idesc, kk > 0);
    }
    commit(bar);
  }
  mbar_wait_wd(bar, 0);
  fence_after_sync();
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  const int row = warp * 32 + lane;
  uint32_t p[64];
  for (int c4 = 0; c4 < 4; ++c4) {
    uint32_t r[32];
    tmem_ld32(lane_base + c4 * 32, r);
    tmem_wait_ld();
    for (int i = 0; i < 32; ++i) s_out[row * 128 + c4 * 32 + i] = __uint_as_float(r[i]);
    for (int i = 0; i < 16; ++i)
      p[c4 * 16 + i] = MmaOp<T>::pack(__uint_as_float(r[2 * i]) * 0.0625f, __uint_as_float(r[2 * i + 1]) * 0.0625f);
  }
  {
    uint32_t lo[32], hi[32];
    for (int i = 0; i < 32; ++i) {
      lo[i] = p[i];
      hi[i] = p[32 + i];
    }
    tmem_st32(lane_base, lo);
    tmem_st32(lane_base + 32, hi);
    tmem_wait_st();
  }
  fence_before_sync();
  __syncthreads();
  if (tid == 0) {
    fence_after_sync();
    const uint64_t db = sw128_desc(v_s);
    for (int kk = 0; kk < 8; ++kk) {
      const uint32_t adv = (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
      mma_ts(tmem + 256, tmem + kk * 8, desc_advance(db, adv), idesc, kk > 0);
    }
    commit(bar);
  }
  mbar_wait_wd(bar, 1);
  fence_after_sync();
  for (int c4 = 0; c4 < 4; ++c4) {
    uint32_t r[32];
    tmem_ld32(lane_base + 256 + c4 * 32, r);
    tmem_wait_ld();
    for (int i = 0; i < 32; ++i) o_out[row * 128 + c4 * 32 + i] = __uint_as_float(r[i]);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <typename T>
int run(int fmt, const char* name) {
  const int N = 128 * 128;
  std::vector<T> q(N), k(N), vt(N);
  std::vector<float> qf(N), kf(N), vf(N);
  srand(1234 + fmt);
  auto rnd = []() { return (float)(rand() % 2001 - 1000) / 1000.f; };
  for (int i = 0; i < N; ++i) {
    q[i] = from_f32_host<T>(rnd()); k[i] = from_f32_host<T>(rnd()); vt[i] = from_f32_host<T>(rnd());
    qf[i] = to_f32_host(q[i]); kf[i] = to_f32_host(k[i]); vf[i] = to_f32_host(vt[i]);
  }
  T *dq, *dk, *dv;
  float *ds, *dout;
  cudaMalloc(&dq, N * sizeof(T)); cudaMalloc(&dk, N * sizeof(T)); cudaMalloc(&dv, N * sizeof(T));
  cudaMalloc(&ds, N * 4); cudaMalloc(&dout, N * 4);
  cudaMemcpy(dq, q.data(), N * sizeof(T), cudaMemcpyHostToDevice);
  cudaMemcpy(dk, k.data(), N * sizeof(T), cudaMemcpyHostToDevice);
  cudaMemcpy(dv, vt.data(), N * sizeof(T), cudaMemcpyHostToDevice);
  cudaMemset(ds, 0xff, N * 4); cudaMemset(dout, 0xff, N * 4);
  const int smem = 98304 + 64 + 1024;
  cudaFuncSetAttribute(probe_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe_kernel<T><<<1, 128, smem>>>(dq, dk, dv, ds, dout, fmt);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("%s: CUDA error %s\n", name, cudaGetErrorString(e));
    return 1;
  }
  std::vector<float> s(N), o(N);
  cudaMemcpy(s.data(), ds, N * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(o.data(), dout, N * 4, cudaMemcpyDeviceToHost);
  double es = 0, eo = 0;
  std::vector<float> pf(N);
  for (int r = 0; r < 128; ++r)
    for (int c = 0; c < 128; ++c) {
      double acc = 0;
      for (int d = 0; d < 128; ++d) acc += (double)qf[r * 128 + d] * kf[c * 128 + d];
      es = fmax(es, fabs(acc - s[r * 128 + c]));
      pf[r * 128 + c] = to_f32_host(from_f32_host<T>(s[r * 128 + c] * 0.0625f));
    }
  for (int r = 0; r < 128; ++r)
    for (int d = 0; d < 128; ++d) {
      double acc = 0;
      for (int c = 0; c < 128; ++c) acc += (double)pf[r * 128 + c] * vf[d * 128 + c];
      eo = fmax(eo, fabs(acc - o[r * 128 + d]));
    }
  const bool ok = es < 1e-3 && eo < 1e-2 && es == es && eo == eo;
  printf("%s: S max-abs err %.3e, O max-abs err %.3e -> %s\n", name, es, eo, ok ? "PASS" : "FAIL");
  return ok ? 0 : 1;
}

int main() {
  int rc = 0;
  rc |= run<__half>(0, "fp16");
  rc |= run<__nv_bfloat16>(1, "bf16");
  return rc;
}
