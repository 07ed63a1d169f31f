// Probe: can the paged K / V cache feed tcgen05.mma straight from TMA tensor copies (no per-thread gather)?
//   K page [D/8][BS][8]  -> per 8-token half page ONE box (64 elems x D/8 rows) = [chunk][8 tokens][16 B] in shared
//                           memory: the canonical NO-swizzle K-major layout (core matrix = 8 rows x 16 B contiguous),
//                           LBO = 128 B (next chunk), SBO = 2048 B (next 8 tokens)
//   V page [D][BS]       -> ONE box (16 tokens x D rows) with SWIZZLE_32B = the canonical 32-byte-swizzle K-major
//                           layout of V^T for one k-step of 16 keys, SBO = 256 B
// S = Q K^T over 64 keys (4 pages), O = P V with P = bf16(S / 16), checked on the host.  Not part of the product.
//   nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -o tools/micro/umma_tma_probe \
//        tools/micro/umma_tma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#include <cuda.h>
#include <cudaTypedefs.h>

#include "../../light_vllm_b200/csrc/umma.cuh"

using namespace b200;
using namespace b200::umma;

constexpr int D = 128, BS = 16, KH = 2, NBLK = 9, KT = 64;

__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  const uint32_t lo = ((saddr & 0x3FFFFu) >> 4) | ((lbo >> 4) << 16);
  const uint32_t hi = (sbo >> 4) | (1u << 14) | (layout << 29);
  return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, int c3,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
      : "memory");
}

__global__ void __launch_bounds__(128) probe_kernel(const __nv_bfloat16* q, const __grid_constant__ CUtensorMap kmap,
                                                    const __grid_constant__ CUtensorMap vmap, const int* blocks,
                                                    int kvh, float* s_out, float* o_out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_s = base, k_s = base + 32768, v_s = base + 49152, bar = base + 65536, tbar = bar + 8, slot = bar + 16;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int id = tid; id < 128 * 16; id += 128) {
    const int row = id >> 4, c = id & 15;
    cp_async16(q_s + (uint32_t)(c >> 3) * 16384u + sw128_offset(row, c & 7), q + row * 128 + c * 8, 16);
  }
  cp_async_commit();
  cp_async_wait_all();
  fence_proxy_async_smem();
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_init(tbar, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(slot, 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  if (tid == 0) {
    mbar_arrive_expect_tx(tbar, 2u * KT * D * 2u);
    for (int pg = 0; pg < KT / BS; ++pg) {
      const int blk = blocks[pg];
      for (int h = 0; h < 2; ++h)  // 8 tokens x D: [D/8 chunks][8 tokens][16 B] = 2 KiB
        tma_load_4d(k_s + (uint32_t)(2 * pg + h) * 2048u, &kmap, h * 64, 0, kvh, blk, tbar);
      tma_load_4d(v_s + (uint32_t)pg * 4096u, &vmap, 0, 0, kvh, blk, tbar);  // [D rows][32 B], 32-byte swizzle
    }
  }
  mbar_wait_wd(tbar, 0);
  if (tid == 0) {
    const uint64_t da = sw128_desc(q_s);
    const uint32_t idesc = idesc_f16(1, 128, KT);
    for (int kk = 0; kk < D / 16; ++kk) {
      const uint32_t adv = (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
      const uint64_t db = desc_kmajor(k_s + (uint32_t)kk * 256u, 128u, 2048u, 0u);
      mma_ss(tmem, desc_advance(da, adv), db, idesc, kk > 0);
    }
    commit(bar);
  }
  mbar_wait_wd(bar, 0);
  fence_after_sync();
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  const int row = warp * 32 + lane;
  uint32_t p[32];
  for (int c2 = 0; c2 < 2; ++c2) {
    uint32_t r[32];
    tmem_ld32(lane_base + c2 * 32, r);
    tmem_wait_ld();
    for (int i = 0; i < 32; ++i) s_out[row * KT + c2 * 32 + i] = __uint_as_float(r[i]);
    for (int i = 0; i < 16; ++i)
      p[c2 * 16 + i] = MmaOp<__nv_bfloat16>::pack(__uint_as_float(r[2 * i]) * 0.0625f,
                                                   __uint_as_float(r[2 * i + 1]) * 0.0625f);
  }
  tmem_st32(lane_base, p);
  tmem_wait_st();
  fence_before_sync();
  __syncthreads();
  if (tid == 0) {
    fence_after_sync();
    const uint32_t idesc = idesc_f16(1, 128, D);
    for (int kk = 0; kk < KT / 16; ++kk)
      mma_ts(tmem + 256, tmem + kk * 8, desc_kmajor(v_s + (uint32_t)kk * 4096u, 16u, 256u, 6u), idesc, kk > 0);
    commit(bar);
  }
  mbar_wait_wd(bar, 1);
  fence_after_sync();
  for (int c4 = 0; c4 < 4; ++c4) {
    uint32_t r[32];
    tmem_ld32(lane_base + 256 + c4 * 32, r);
    tmem_wait_ld();
    for (int i = 0; i < 32; ++i) o_out[row * D + c4 * 32 + i] = __uint_as_float(r[i]);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

static PFN_cuTensorMapEncodeTiled encode_fn() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult st;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &st) != cudaSuccess ||
      st != cudaDriverEntryPointSuccess)
    return nullptr;
  return (PFN_cuTensorMapEncodeTiled)fn;
}

int main() {
  const int kpage = D * BS, NK = NBLK * KH * kpage;
  std::vector<__nv_bfloat16> q(128 * D), kc(NK), vc(NK);
  std::vector<float> qf(128 * D), kf(NK), vf(NK);
  srand(7);
  auto rnd = []() { return (float)(rand() % 2001 - 1000) / 1000.f; };
  for (int i = 0; i < 128 * D; ++i) { q[i] = __float2bfloat16_rn(rnd()); qf[i] = __bfloat162float(q[i]); }
  for (int i = 0; i < NK; ++i) {
    kc[i] = __float2bfloat16_rn(rnd()); kf[i] = __bfloat162float(kc[i]);
    vc[i] = __float2bfloat16_rn(rnd()); vf[i] = __bfloat162float(vc[i]);
  }
  const int blocks[4] = {5, 2, 8, 0}, kvh = 1;
  __nv_bfloat16 *dq, *dk, *dv;
  float *ds, *dout;
  int* dblk;
  cudaMalloc(&dq, q.size() * 2); cudaMalloc(&dk, NK * 2); cudaMalloc(&dv, NK * 2);
  cudaMalloc(&ds, 128 * KT * 4); cudaMalloc(&dout, 128 * D * 4); cudaMalloc(&dblk, 16);
  cudaMemcpy(dq, q.data(), q.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dk, kc.data(), NK * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dv, vc.data(), NK * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dblk, blocks, 16, cudaMemcpyHostToDevice);
  cudaMemset(ds, 0xff, 128 * KT * 4); cudaMemset(dout, 0xff, 128 * D * 4);

  PFN_cuTensorMapEncodeTiled enc = encode_fn();
  if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
  CUtensorMap kmap, vmap;
  const cuuint32_t ones[4] = {1, 1, 1, 1};
  {  // K cache [NB][KH][D/8][BS][8]: dim0 = (token, elem) flattened (BS*8 contiguous), dim1 = chunk, dim2 = head, dim3 = block
    const cuuint64_t dims[4] = {(cuuint64_t)BS * 8, D / 8, KH, NBLK};
    const cuuint64_t strides[3] = {(cuuint64_t)BS * 16, (cuuint64_t)kpage * 2, (cuuint64_t)KH * kpage * 2};
    const cuuint32_t box[4] = {64, D / 8, 1, 1};
    CUresult r = enc(&kmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, dk, dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("K map encode failed: %d\n", (int)r); return 1; }
  }
  {  // V cache [NB][KH][D][BS]: dim0 = token, dim1 = d, dim2 = head, dim3 = block
    const cuuint64_t dims[4] = {BS, D, KH, NBLK};
    const cuuint64_t strides[3] = {(cuuint64_t)BS * 2, (cuuint64_t)kpage * 2, (cuuint64_t)KH * kpage * 2};
    const cuuint32_t box[4] = {BS, D, 1, 1};
    CUresult r = enc(&vmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, dv, dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("V map encode failed: %d\n", (int)r); return 1; }
  }
  const int smem = 65536 + 64 + 1024;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe_kernel<<<1, 128, smem>>>(dq, kmap, vmap, dblk, kvh, ds, dout);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  std::vector<float> s(128 * KT), o(128 * D), pf(128 * KT);
  cudaMemcpy(s.data(), ds, s.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost);
  auto K = [&](int key, int d) {  // key in [0, 64): page key / 16
    const int blk = blocks[key / BS], t = key % BS;
    return kf[((size_t)(blk * KH + kvh) * (D / 8) + d / 8) * BS * 8 + t * 8 + d % 8];
  };
  auto V = [&](int key, int d) {
    const int blk = blocks[key / BS], t = key % BS;
    return vf[((size_t)(blk * KH + kvh) * D + d) * BS + t];
  };
  double es = 0, eo = 0;
  for (int r = 0; r < 128; ++r)
    for (int c = 0; c < KT; ++c) {
      double acc = 0;
      for (int d = 0; d < D; ++d) acc += (double)qf[r * D + d] * K(c, d);
      es = fmax(es, fabs(acc - s[r * KT + c]));
      pf[r * KT + c] = __bfloat162float(__float2bfloat16_rn(s[r * KT + c] * 0.0625f));
    }
  for (int r = 0; r < 128; ++r)
    for (int d = 0; d < D; ++d) {
      double acc = 0;
      for (int c = 0; c < KT; ++c) acc += (double)pf[r * KT + c] * V(c, d);
      eo = fmax(eo, fabs(acc - o[r * D + d]));
    }
  const bool ok = es < 1e-3 && eo < 1e-2 && es == es && eo == eo;
  printf("TMA-fed paged tiles: S max-abs err %.3e, O max-abs err %.3e -> %s\n", es, eo, ok ? "PASS" : "FAIL");
  return ok ? 0 : 1;
}
