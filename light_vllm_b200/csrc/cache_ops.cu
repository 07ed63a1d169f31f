// Cache-write and cache-management kernels of libb200_paged.so (sm_100a).
//
//   b200_reshape_and_cache        <- csrc/cache_kernels.cu:152-204, 263-289
//   b200_reshape_and_cache_flash  <- csrc/cache_kernels.cu:206-247, 304-330
//   b200_copy_blocks              <- csrc/cache_kernels.cu:68-94, 101-148
//   b200_swap_blocks              <- csrc/cache_kernels.cu:24-63
//   b200_convert_fp8              <- csrc/cache_kernels.cu:334-405
//   b200_get_device_attribute ... <- csrc/cuda_utils_kernels.cu:8-29
//
// All of these are byte / index work bounded by HBM (or the host link for swap): the design is
// 16-byte vector accesses, the fp8 quantisation fused into the single store pass, and one
// launch per call with no host<->device synchronisation.
#include "common.cuh"

#include <algorithm>
#include <atomic>
#include <vector>

namespace b200 {

// -------------------------------------------------------------------------------------
// host-side error / bookkeeping
// -------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// -------------------------------------------------------------------------------------
// element codecs: SRC -> cache element
// -------------------------------------------------------------------------------------
template <typename T, typename CT, int KV>
__device__ __forceinline__ CT encode_elem(T v, float scale) {
  if constexpr (KV == B200_KV_AUTO) {
    return v;  // T == CT: plain copy, bit-exact
  } else {
    return fp8_encode<KV>(to_f32<T>(v), scale);
  }
}

// N elements moved as one unit; alignment capped at 16 bytes so that wider packs become
// several 16-byte accesses instead of demanding 32/64-byte aligned rows.
template <typename T, int N>
struct alignas((sizeof(T) * N) < 16 ? (sizeof(T) * N) : 16) Pack {
  T v[N];
};

// -------------------------------------------------------------------------------------
// reshape_and_cache (paged layout K [nb, kh, d/x, bs, x], V [nb, kh, d, bs])
//   one CTA per token.  Fast path (x == 16 / sizeof(CT), aligned rows): each thread moves one
//   x-element chunk: K is one 16-byte store, V is x strided element stores (layout-imposed).
// -------------------------------------------------------------------------------------
template <typename T, typename CT, int KV, bool VEC>
__global__ void __launch_bounds__(256)
reshape_and_cache_kernel(const T* __restrict__ key, const T* __restrict__ value,
                         CT* __restrict__ key_cache, CT* __restrict__ value_cache,
                         const int64_t* __restrict__ slot_mapping, int64_t key_stride,
                         int64_t value_stride, int num_heads, int head_size, int block_size, int x,
                         float k_scale, float v_scale) {
  // The attention launch that follows in a decode step is a programmatic dependent (paged_attention_inst.inc):
  // let it become resident now and park in griddepcontrol.wait -- it reads nothing before this grid has completed.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padding token (cache_kernels.cu:165-169)
  const int64_t blk = slot / block_size;
  const int64_t off = slot % block_size;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  const int64_t page = (int64_t)head_size * block_size;  // elements per (block, head)
  CT* kdst_blk = key_cache + blk * num_heads * page;
  CT* vdst_blk = value_cache + blk * num_heads * page;

  if constexpr (VEC) {
    constexpr int X = 16 / sizeof(CT);  // elements per 16-byte cache chunk
    const int chunks_per_head = head_size / X;
    const int n = num_heads * chunks_per_head;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int h = i / chunks_per_head;
      const int c = i - h * chunks_per_head;
      const Pack<T, X> kin = *reinterpret_cast<const Pack<T, X>*>(ksrc + (int64_t)i * X);
      const Pack<T, X> vin = *reinterpret_cast<const Pack<T, X>*>(vsrc + (int64_t)i * X);
      Pack<CT, X> kout;
#pragma unroll
      for (int e = 0; e < X; ++e) kout.v[e] = encode_elem<T, CT, KV>(kin.v[e], k_scale);
      *reinterpret_cast<Pack<CT, X>*>(kdst_blk + h * page + ((int64_t)c * block_size + off) * X) =
          kout;
      CT* vdst = vdst_blk + h * page + (int64_t)c * X * block_size + off;
#pragma unroll
      for (int e = 0; e < X; ++e)
        vdst[(int64_t)e * block_size] = encode_elem<T, CT, KV>(vin.v[e], v_scale);
    }
  } else {
    const int n = num_heads * head_size;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int h = i / head_size;
      const int d = i - h * head_size;
      const int xi = d / x, xo = d - xi * x;
      kdst_blk[h * page + ((int64_t)xi * block_size + off) * x + xo] =
          encode_elem<T, CT, KV>(ksrc[i], k_scale);
      vdst_blk[h * page + (int64_t)d * block_size + off] = encode_elem<T, CT, KV>(vsrc[i], v_scale);
    }
  }
}

// -------------------------------------------------------------------------------------
// reshape_and_cache_flash (layout [nb, bs, kh, d], explicit block stride): one CTA per token,
// a token row is contiguous in both source and cache -> 16-byte loads, 16/8-byte stores.
// -------------------------------------------------------------------------------------
template <typename T, typename CT, int KV, bool VEC>
__global__ void __launch_bounds__(256)
reshape_and_cache_flash_kernel(const T* __restrict__ key, const T* __restrict__ value,
                               CT* __restrict__ key_cache, CT* __restrict__ value_cache,
                               const int64_t* __restrict__ slot_mapping, int64_t block_stride,
                               int64_t key_stride, int64_t value_stride, int num_heads,
                               int head_size, int block_size, float k_scale, float v_scale) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // see reshape_and_cache_kernel
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t blk = slot / block_size;
  const int64_t off = slot % block_size;
  const int n = num_heads * head_size;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  CT* kdst = key_cache + blk * block_stride + off * n;
  CT* vdst = value_cache + blk * block_stride + off * n;
  if constexpr (VEC) {
    constexpr int N = 16 / sizeof(T);  // elements per 16-byte source vector
    for (int i = threadIdx.x * N; i < n; i += blockDim.x * N) {
      const Pack<T, N> kin = *reinterpret_cast<const Pack<T, N>*>(ksrc + i);
      const Pack<T, N> vin = *reinterpret_cast<const Pack<T, N>*>(vsrc + i);
      Pack<CT, N> ko, vo;
#pragma unroll
      for (int e = 0; e < N; ++e) {
        ko.v[e] = encode_elem<T, CT, KV>(kin.v[e], k_scale);
        vo.v[e] = encode_elem<T, CT, KV>(vin.v[e], v_scale);
      }
      *reinterpret_cast<Pack<CT, N>*>(kdst + i) = ko;
      *reinterpret_cast<Pack<CT, N>*>(vdst + i) = vo;
    }
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      kdst[i] = encode_elem<T, CT, KV>(ksrc[i], k_scale);
      vdst[i] = encode_elem<T, CT, KV>(vsrc[i], v_scale);
    }
  }
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T, typename CT, int KV>
static int launch_reshape(const void* key, const void* value, void* key_cache, void* value_cache,
                          const int64_t* slot_mapping, int num_tokens, int num_heads, int head_size,
                          int block_size, int x, int64_t key_stride, int64_t value_stride,
                          float k_scale, float v_scale, cudaStream_t stream) {
  constexpr int X = 16 / sizeof(CT);
  const bool vec = (x == X) && (head_size % X == 0) && aligned16(key) && aligned16(value) &&
                   aligned16(key_cache) && (key_stride * sizeof(T)) % 16 == 0 &&
                   (value_stride * sizeof(T)) % 16 == 0 && (X * sizeof(T)) % 16 == 0;
  const int work = vec ? num_heads * (head_size / X) : num_heads * head_size;
  const int threads = std::min(256, std::max(32, (work + 31) / 32 * 32));
  auto kp = static_cast<const T*>(key);
  auto vp = static_cast<const T*>(value);
  auto kc = static_cast<CT*>(key_cache);
  auto vc = static_cast<CT*>(value_cache);
  if (vec)
    reshape_and_cache_kernel<T, CT, KV, true><<<num_tokens, threads, 0, stream>>>(
        kp, vp, kc, vc, slot_mapping, key_stride, value_stride, num_heads, head_size, block_size, x,
        k_scale, v_scale);
  else
    reshape_and_cache_kernel<T, CT, KV, false><<<num_tokens, threads, 0, stream>>>(
        kp, vp, kc, vc, slot_mapping, key_stride, value_stride, num_heads, head_size, block_size, x,
        k_scale, v_scale);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

template <typename T, typename CT, int KV>
static int launch_reshape_flash(const void* key, const void* value, void* key_cache,
                                void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                int num_heads, int head_size, int block_size, int64_t block_stride,
                                int64_t key_stride, int64_t value_stride, float k_scale,
                                float v_scale, cudaStream_t stream) {
  constexpr int N = 16 / sizeof(T);
  const int n = num_heads * head_size;
  const bool vec = (n % N == 0) && aligned16(key) && aligned16(value) && aligned16(key_cache) &&
                   aligned16(value_cache) && (key_stride * sizeof(T)) % 16 == 0 &&
                   (value_stride * sizeof(T)) % 16 == 0 && (block_stride * sizeof(CT)) % 16 == 0 &&
                   ((int64_t)n * sizeof(CT)) % 16 == 0;
  const int work = vec ? n / N : n;
  const int threads = std::min(256, std::max(32, (work + 31) / 32 * 32));
  auto kp = static_cast<const T*>(key);
  auto vp = static_cast<const T*>(value);
  auto kc = static_cast<CT*>(key_cache);
  auto vc = static_cast<CT*>(value_cache);
  if (vec)
    reshape_and_cache_flash_kernel<T, CT, KV, true><<<num_tokens, threads, 0, stream>>>(
        kp, vp, kc, vc, slot_mapping, block_stride, key_stride, value_stride, num_heads, head_size,
        block_size, k_scale, v_scale);
  else
    reshape_and_cache_flash_kernel<T, CT, KV, false><<<num_tokens, threads, 0, stream>>>(
        kp, vp, kc, vc, slot_mapping, block_stride, key_stride, value_stride, num_heads, head_size,
        block_size, k_scale, v_scale);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

// dispatch (src dtype, kv dtype) -> (T, CT, KV); mirrors DISPATCH_BY_KV_CACHE_DTYPE
// (quant_utils.cuh:532-569).
#define B200_DISPATCH_SRC_KV(src_dtype, kv_dtype, FN, ...)                                   \
  do {                                                                                       \
    if ((kv_dtype) == B200_KV_AUTO) {                                                        \
      if ((src_dtype) == B200_DT_F32) return FN<float, float, B200_KV_AUTO>(__VA_ARGS__);     \
      if ((src_dtype) == B200_DT_F16) return FN<__half, __half, B200_KV_AUTO>(__VA_ARGS__);   \
      if ((src_dtype) == B200_DT_BF16)                                                       \
        return FN<__nv_bfloat16, __nv_bfloat16, B200_KV_AUTO>(__VA_ARGS__);                  \
    } else if ((kv_dtype) == B200_KV_FP8_E4M3) {                                             \
      if ((src_dtype) == B200_DT_F32) return FN<float, uint8_t, B200_KV_FP8_E4M3>(__VA_ARGS__); \
      if ((src_dtype) == B200_DT_F16) return FN<__half, uint8_t, B200_KV_FP8_E4M3>(__VA_ARGS__); \
      if ((src_dtype) == B200_DT_BF16)                                                       \
        return FN<__nv_bfloat16, uint8_t, B200_KV_FP8_E4M3>(__VA_ARGS__);                    \
    } else if ((kv_dtype) == B200_KV_FP8_E5M2) {                                             \
      if ((src_dtype) == B200_DT_F32) return FN<float, uint8_t, B200_KV_FP8_E5M2>(__VA_ARGS__); \
      if ((src_dtype) == B200_DT_F16) return FN<__half, uint8_t, B200_KV_FP8_E5M2>(__VA_ARGS__); \
      if ((src_dtype) == B200_DT_BF16)                                                       \
        return FN<__nv_bfloat16, uint8_t, B200_KV_FP8_E5M2>(__VA_ARGS__);                    \
    } else {                                                                                 \
      ::b200::set_error("Unsupported data type of kv cache: %d", (int)(kv_dtype));           \
      return B200_ERR_UNSUPPORTED;                                                           \
    }                                                                                        \
    ::b200::set_error("Unsupported input type of kv cache: %d", (int)(src_dtype));           \
    return B200_ERR_UNSUPPORTED;                                                             \
  } while (0)

// -------------------------------------------------------------------------------------
// copy_blocks: (layer, K|V, pair, 16 KiB chunk) -> one CTA; pointer table lives in the
// kernel parameter block, so no host->device copy and no synchronisation.
// -------------------------------------------------------------------------------------
constexpr int kCopyMaxLayers = 120;
struct CopyPtrs {
  void* ptr[2 * kCopyMaxLayers];  // [layer*2 + 0] = K base, [layer*2 + 1] = V base
};
constexpr int kCopyThreads = 256;
constexpr int kCopyUnroll = 4;
constexpr int kCopyChunk = kCopyThreads * kCopyUnroll * 16;  // 16 KiB per CTA

__device__ __forceinline__ void copy_chunk(char* __restrict__ dst, const char* __restrict__ src,
                                           int64_t block_bytes, int chunk, bool vec) {
  const int64_t base = (int64_t)chunk * kCopyChunk;
  if (vec) {
    int4 tmp[kCopyUnroll];
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) {
      const int64_t o = base + ((int64_t)u * kCopyThreads + threadIdx.x) * 16;
      if (o < block_bytes) tmp[u] = *reinterpret_cast<const int4*>(src + o);
    }
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) {
      const int64_t o = base + ((int64_t)u * kCopyThreads + threadIdx.x) * 16;
      if (o < block_bytes) *reinterpret_cast<int4*>(dst + o) = tmp[u];
    }
  } else {
    const int64_t end = min(base + (int64_t)kCopyChunk, block_bytes);
    for (int64_t o = base + threadIdx.x; o < end; o += kCopyThreads) dst[o] = src[o];
  }
}

__global__ void __launch_bounds__(kCopyThreads)
copy_blocks_kernel(const __grid_constant__ CopyPtrs ptrs, const int64_t* __restrict__ block_mapping,
                   int64_t block_bytes, bool vec) {
  char* base = static_cast<char*>(ptrs.ptr[blockIdx.z]);
  const int64_t src_blk = block_mapping[2 * (int64_t)blockIdx.y];
  const int64_t dst_blk = block_mapping[2 * (int64_t)blockIdx.y + 1];
  copy_chunk(base + dst_blk * block_bytes, base + src_blk * block_bytes, block_bytes, blockIdx.x,
             vec);
}

// -------------------------------------------------------------------------------------
// swap_blocks: gather/scatter kernel over a batch of (src, dst) block pairs carried in the
// kernel parameter block; src/dst may be device memory or pinned host memory (UVA).
// -------------------------------------------------------------------------------------
constexpr int kSwapPairsPerLaunch = 224;
struct SwapBatch {
  int64_t src_off[kSwapPairsPerLaunch];  // byte offsets
  int64_t dst_off[kSwapPairsPerLaunch];
};
__global__ void __launch_bounds__(kCopyThreads)
swap_blocks_kernel(const __grid_constant__ SwapBatch batch, const char* __restrict__ src,
                   char* __restrict__ dst, int64_t block_bytes, bool vec) {
  copy_chunk(dst + batch.dst_off[blockIdx.y], src + batch.src_off[blockIdx.y], block_bytes,
             blockIdx.x, vec);
}

// -------------------------------------------------------------------------------------
// convert_fp8: elementwise, 8 elements per thread.
// -------------------------------------------------------------------------------------
template <typename T, int KV, bool TO_FP8>
__global__ void __launch_bounds__(256)
convert_fp8_kernel(void* __restrict__ dst_, const void* __restrict__ src_, int64_t numel,
                   float scale, bool vec) {
  constexpr int N = 8;
  const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * N;
  if (i0 >= numel) return;
  if constexpr (TO_FP8) {
    const T* src = static_cast<const T*>(src_);
    uint8_t* dst = static_cast<uint8_t*>(dst_);
    if (vec && i0 + N <= numel) {
      Pack<T, N> in = *reinterpret_cast<const Pack<T, N>*>(src + i0);
      Pack<uint8_t, N> out;
#pragma unroll
      for (int e = 0; e < N; ++e) out.v[e] = fp8_encode<KV>(to_f32<T>(in.v[e]), scale);
      *reinterpret_cast<Pack<uint8_t, N>*>(dst + i0) = out;
    } else {
      for (int64_t i = i0; i < min(i0 + N, numel); ++i)
        dst[i] = fp8_encode<KV>(to_f32<T>(src[i]), scale);
    }
  } else {
    const uint8_t* src = static_cast<const uint8_t*>(src_);
    T* dst = static_cast<T*>(dst_);
    if (vec && i0 + N <= numel) {
      Pack<uint8_t, N> in = *reinterpret_cast<const Pack<uint8_t, N>*>(src + i0);
      Pack<T, N> out;
#pragma unroll
      for (int e = 0; e < N; ++e) out.v[e] = from_f32<T>(fp8_decode<KV>(in.v[e], scale));
      *reinterpret_cast<Pack<T, N>*>(dst + i0) = out;
    } else {
      for (int64_t i = i0; i < min(i0 + N, numel); ++i)
        dst[i] = from_f32<T>(fp8_decode<KV>(src[i], scale));
    }
  }
}

template <typename T, int KV>
static int launch_convert(void* dst, const void* src, int64_t numel, float scale, int to_fp8,
                          cudaStream_t stream) {
  if (numel == 0) return B200_OK;
  const bool vec = aligned16(dst) && aligned16(src);
  const int64_t threads_total = (numel + 7) / 8;
  const int64_t grid = (threads_total + 255) / 256;
  if (grid > 0x7fffffffLL) {
    set_error("convert_fp8: tensor too large");
    return B200_ERR_INVALID;
  }
  if (to_fp8)
    convert_fp8_kernel<T, KV, true><<<(unsigned)grid, 256, 0, stream>>>(dst, src, numel, scale, vec);
  else
    convert_fp8_kernel<T, KV, false><<<(unsigned)grid, 256, 0, stream>>>(dst, src, numel, scale, vec);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

// =====================================================================================
// C ABI
// =====================================================================================
using namespace b200;

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_abi_version(void) { return B200_PAGED_ABI_VERSION; }
extern "C" int64_t b200_launch_count(void) { return (int64_t)g_launches.load(); }

extern "C" int b200_parse_kv_cache_dtype(const char* s) {
  if (!s) return -1;
  auto eq = [&](const char* t) {
    const char* a = s;
    while (*a && *t && *a == *t) ++a, ++t;
    return *a == 0 && *t == 0;
  };
  if (eq("auto")) return B200_KV_AUTO;
  if (eq("fp8") || eq("fp8_e4m3")) return B200_KV_FP8_E4M3;
  if (eq("fp8_e5m2")) return B200_KV_FP8_E5M2;
  return -1;
}

extern "C" int b200_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                      void* value_cache, const int64_t* slot_mapping,
                                      int num_tokens, int num_heads, int head_size, int block_size,
                                      int x, int64_t key_stride, int64_t value_stride,
                                      float k_scale, float v_scale, int src_dtype, int kv_dtype,
                                      int device, b200_stream_t stream) {
  B200_CHECK_ARG(num_tokens >= 0 && num_heads > 0 && head_size > 0 && block_size > 0 && x > 0,
                 "reshape_and_cache: bad shape (tokens=%d heads=%d head_size=%d block_size=%d x=%d)",
                 num_tokens, num_heads, head_size, block_size, x);
  B200_CHECK_ARG(head_size % x == 0, "reshape_and_cache: head_size %d not divisible by x %d",
                 head_size, x);
  if (num_tokens == 0) return B200_OK;
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_DISPATCH_SRC_KV(src_dtype, kv_dtype, launch_reshape, key, value, key_cache, value_cache,
                       slot_mapping, num_tokens, num_heads, head_size, block_size, x, key_stride,
                       value_stride, k_scale, v_scale, (cudaStream_t)stream);
}

extern "C" int b200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                            void* value_cache, const int64_t* slot_mapping,
                                            int num_tokens, int num_heads, int head_size,
                                            int block_size, int64_t block_stride,
                                            int64_t key_stride, int64_t value_stride, float k_scale,
                                            float v_scale, int src_dtype, int kv_dtype, int device,
                                            b200_stream_t stream) {
  B200_CHECK_ARG(num_tokens >= 0 && num_heads > 0 && head_size > 0 && block_size > 0,
                 "reshape_and_cache_flash: bad shape");
  if (num_tokens == 0) return B200_OK;
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_DISPATCH_SRC_KV(src_dtype, kv_dtype, launch_reshape_flash, key, value, key_cache,
                       value_cache, slot_mapping, num_tokens, num_heads, head_size, block_size,
                       block_stride, key_stride, value_stride, k_scale, v_scale,
                       (cudaStream_t)stream);
}

extern "C" int b200_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs,
                                int num_layers, const int64_t* block_mapping, int num_pairs,
                                int64_t block_bytes, int device, b200_stream_t stream) {
  B200_CHECK_ARG(num_layers >= 0 && num_pairs >= 0 && block_bytes > 0, "copy_blocks: bad sizes");
  if (num_layers == 0 || num_pairs == 0) return B200_OK;
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  bool vec = (block_bytes % 16 == 0);
  for (int l = 0; l < num_layers; ++l)
    vec = vec && aligned16(key_cache_ptrs[l]) && aligned16(value_cache_ptrs[l]);
  const unsigned chunks = (unsigned)((block_bytes + kCopyChunk - 1) / kCopyChunk);
  for (int l0 = 0; l0 < num_layers; l0 += kCopyMaxLayers) {
    const int nl = std::min(kCopyMaxLayers, num_layers - l0);
    CopyPtrs p;
    for (int l = 0; l < nl; ++l) {
      p.ptr[2 * l] = key_cache_ptrs[l0 + l];
      p.ptr[2 * l + 1] = value_cache_ptrs[l0 + l];
    }
    for (int p0 = 0; p0 < num_pairs; p0 += 65535) {
      const int np = std::min(65535, num_pairs - p0);
      dim3 grid(chunks, np, 2 * nl);
      copy_blocks_kernel<<<grid, kCopyThreads, 0, (cudaStream_t)stream>>>(
          p, block_mapping + 2 * (int64_t)p0, block_bytes, vec);
      count_launch();
    }
  }
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}


// Resolve a pointer to something a kernel running on `device` can dereference, or nullptr.
static const void* device_addressable(const void* p, int p_device) {
  if (p_device >= 0) return p;
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged)
    return attr.devicePointer;
  return nullptr;  // pageable host memory
}

// mode: 0 = auto (gather kernel when both sides are device-addressable, else memcpy), 1 = force
// run-merged cudaMemcpyAsync, 2 = force the kernel.  A per-call argument; no state is kept.
extern "C" int b200_swap_blocks_mode(const void* src, void* dst, const int64_t* block_mapping_host,
                                     int num_pairs, int64_t block_bytes, int src_device,
                                     int dst_device, int mode, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK_ARG(num_pairs >= 0 && block_bytes > 0, "swap_blocks: bad sizes");
  cudaMemcpyKind kind;
  if (src_device >= 0 && dst_device >= 0) {
    B200_CHECK_ARG(src_device == dst_device, "src and dst must be on the same GPU");
    kind = cudaMemcpyDeviceToDevice;
  } else if (src_device >= 0) {
    kind = cudaMemcpyDeviceToHost;
  } else if (dst_device >= 0) {
    kind = cudaMemcpyHostToDevice;
  } else {
    set_error("Invalid device combination");
    return B200_ERR_INVALID;
  }
  if (num_pairs == 0) return B200_OK;
  DeviceGuard guard(src_device >= 0 ? src_device : dst_device);
  B200_CUDA_OK(guard.err);

  B200_CHECK_ARG(mode >= 0 && mode <= 2, "swap_blocks: mode must be 0, 1 or 2");
  const void* src_dev = device_addressable(src, src_device);
  void* dst_dev = const_cast<void*>(device_addressable(dst, dst_device));
  const bool kernel_ok = src_dev && dst_dev;
  if (mode == 2 && !kernel_ok) {
    set_error("swap_blocks: kernel mode forced but host memory is not pinned");
    return B200_ERR_INVALID;
  }
  if (kernel_ok && mode != 1) {
    const bool vec = (block_bytes % 16 == 0) && aligned16(src_dev) && aligned16(dst_dev);
    const unsigned chunks = (unsigned)((block_bytes + kCopyChunk - 1) / kCopyChunk);
    for (int p0 = 0; p0 < num_pairs; p0 += kSwapPairsPerLaunch) {
      const int np = std::min(kSwapPairsPerLaunch, num_pairs - p0);
      SwapBatch b;
      for (int i = 0; i < np; ++i) {
        b.src_off[i] = block_mapping_host[2 * (int64_t)(p0 + i)] * block_bytes;
        b.dst_off[i] = block_mapping_host[2 * (int64_t)(p0 + i) + 1] * block_bytes;
      }
      dim3 grid(chunks, np);
      swap_blocks_kernel<<<grid, kCopyThreads, 0, stream>>>(b, static_cast<const char*>(src_dev),
                                                            static_cast<char*>(dst_dev),
                                                            block_bytes, vec);
      count_launch();
    }
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
  }
  // Copy-engine path: merge runs of consecutive (src+1, dst+1) pairs into one memcpy each.
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  int i = 0;
  while (i < num_pairs) {
    const int64_t s0 = block_mapping_host[2 * (int64_t)i];
    const int64_t d0 = block_mapping_host[2 * (int64_t)i + 1];
    int run = 1;
    while (i + run < num_pairs && block_mapping_host[2 * (int64_t)(i + run)] == s0 + run &&
           block_mapping_host[2 * (int64_t)(i + run) + 1] == d0 + run)
      ++run;
    B200_CUDA_OK(cudaMemcpyAsync(d + d0 * block_bytes, s + s0 * block_bytes,
                                 (size_t)run * block_bytes, kind, stream));
    i += run;
  }
  return B200_OK;
}

extern "C" int b200_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host,
                                int num_pairs, int64_t block_bytes, int src_device, int dst_device,
                                b200_stream_t stream) {
  return b200_swap_blocks_mode(src, dst, block_mapping_host, num_pairs, block_bytes, src_device,
                               dst_device, 0, stream);
}

extern "C" int b200_convert_fp8(void* dst, const void* src, int64_t numel, float scale,
                                int wide_dtype, int kv_dtype, int to_fp8, int device,
                                b200_stream_t stream) {
  B200_CHECK_ARG(numel >= 0, "convert_fp8: negative numel");
  if (kv_dtype != B200_KV_FP8_E4M3 && kv_dtype != B200_KV_FP8_E5M2) {
    set_error("Unsupported data type: convert_fp8 needs an fp8 kv_cache_dtype (got %d)", kv_dtype);
    return B200_ERR_UNSUPPORTED;
  }
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  cudaStream_t st = (cudaStream_t)stream;
#define B200_CONV(T)                                                                         \
  return kv_dtype == B200_KV_FP8_E4M3                                                        \
             ? launch_convert<T, B200_KV_FP8_E4M3>(dst, src, numel, scale, to_fp8, st)       \
             : launch_convert<T, B200_KV_FP8_E5M2>(dst, src, numel, scale, to_fp8, st)
  if (wide_dtype == B200_DT_F32) { B200_CONV(float); }
  if (wide_dtype == B200_DT_F16) { B200_CONV(__half); }
  if (wide_dtype == B200_DT_BF16) { B200_CONV(__nv_bfloat16); }
#undef B200_CONV
  set_error("convert_fp8: unsupported dtype %d", wide_dtype);
  return B200_ERR_UNSUPPORTED;
}

extern "C" int64_t b200_get_device_attribute(int64_t attribute, int64_t device_id) {
  int device = (int)device_id;
  if (device_id < 0) {
    if (cudaGetDevice(&device) != cudaSuccess) return -1;
  }
  int value = 0;
  if (cudaDeviceGetAttribute(&value, static_cast<cudaDeviceAttr>(attribute), device) !=
      cudaSuccess) {
    set_error("cudaDeviceGetAttribute(%lld, %d) failed", (long long)attribute, device);
    return -1;
  }
  return value;
}

extern "C" int64_t b200_get_max_shared_memory_per_block_device_attribute(int64_t device_id) {
  // cudaDevAttrMaxSharedMemoryPerBlockOptin (cuda_utils_kernels.cu:17-29)
  return b200_get_device_attribute((int64_t)cudaDevAttrMaxSharedMemoryPerBlockOptin, device_id);
}
