// Per-step ops either side of the paged-attention call, sm_100a:
//
//   b200_advance_step                  <- csrc/prepare_inputs/advance_step.cu:15-130
//   b200_rotary_embedding              <- csrc/pos_encoding_kernels.cu:9-200 (plain and batched)
//   b200_rotary_reshape_and_cache      =  rotary_embedding followed by reshape_and_cache
//                                         (csrc/cache_kernels.cu:152-204) in ONE pass over K:
//                                         q and k are rotated in place and the token's CTA then
//                                         writes the rotated k (an L1/L2 hit) and v straight into
//                                         the paged cache, fp8 quantisation included.
//
// These are launch-latency-bound index / elementwise kernels (a decode step touches ~1 MiB): the
// point is fewer launches and no host round trip, so a whole decode step can sit in a CUDA graph.
// Rounding follows the reference exactly: its arithmetic on c10::Half / c10::BFloat16 is
// "multiply in float, round to T; subtract in float, round to T" (three roundings per output).
#include "common.cuh"

#include <algorithm>

namespace b200 {

// -------------------------------------------------------------------------------------
// advance_step: one thread per query (the reference launches #SM CTAs x 256 threads and lets the
// surplus exit; same result).
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
advance_step_kernel(int num_queries, int block_size, int64_t* __restrict__ input_tokens,
                    const int64_t* __restrict__ sampled_token_ids,
                    int64_t* __restrict__ input_positions, int32_t* __restrict__ seq_lens,
                    int64_t* __restrict__ slot_mapping, const int32_t* __restrict__ block_tables,
                    int64_t block_tables_stride) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= num_queries) return;
  input_tokens[q] = sampled_token_ids[q];
  const int next_seq_len = seq_lens[q] + 1;
  const int next_pos = next_seq_len - 1;
  seq_lens[q] = next_seq_len;
  input_positions[q] = next_pos;
  const int32_t* bt = block_tables + block_tables_stride * q;
  // int arithmetic as in the reference (advance_step.cu:52-55)
  const int slot = bt[next_pos / block_size] * block_size + next_pos % block_size;
  slot_mapping[q] = slot;
}

// -------------------------------------------------------------------------------------
// rotary embedding
// -------------------------------------------------------------------------------------
// x*cos - y*sin and y*cos + x*sin with the reference's rounding points.
template <typename T>
__device__ __forceinline__ void rotate_pair(T x, T y, T c, T s, T& ox, T& oy) {
  if constexpr (sizeof(T) == 4) {
    // what nvcc makes of the reference's `x * cos - y * sin` / `y * cos + x * sin` in float
    // (SASS of pos_encoding_kernels.cu for sm_100a: new_x = fma(x, cos, -(y*sin)), new_y =
    // fma(x, sin, y*cos) in all four float instantiations), written out so it cannot drift
    ox = __fmaf_rn(x, c, -__fmul_rn(y, s));
    oy = __fmaf_rn(x, s, __fmul_rn(y, c));
  } else {
    const float xf = to_f32<T>(x), yf = to_f32<T>(y), cf = to_f32<T>(c), sf = to_f32<T>(s);
    const float xc = to_f32<T>(from_f32<T>(xf * cf)), ys = to_f32<T>(from_f32<T>(yf * sf));
    const float yc = to_f32<T>(from_f32<T>(yf * cf)), xs = to_f32<T>(from_f32<T>(xf * sf));
    ox = from_f32<T>(xc - ys);
    oy = from_f32<T>(yc + xs);
  }
}

// Rotate the heads of one token's row `arr` ([heads, head_size]) in place; elements beyond rot_dim
// stay untouched (partial rotary).
template <typename T, bool NEOX>
__device__ __forceinline__ void rotate_row(T* __restrict__ arr, const T* __restrict__ cache_ptr,
                                           int heads, int head_size, int rot_dim) {
  const int embed = rot_dim / 2;
  const T* cos_ptr = cache_ptr;
  const T* sin_ptr = cache_ptr + embed;
  const int n = heads * embed;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int h = i / embed, r = i - h * embed;
    const int xi = NEOX ? r : 2 * r, yi = NEOX ? embed + r : 2 * r + 1;
    T* a = arr + (int64_t)h * head_size;
    T ox, oy;
    rotate_pair<T>(a[xi], a[yi], cos_ptr[r], sin_ptr[r], ox, oy);
    a[xi] = ox;
    a[yi] = oy;
  }
}

template <typename T, bool NEOX>
__global__ void __launch_bounds__(512)
rotary_embedding_kernel(const int64_t* __restrict__ positions, T* __restrict__ query,
                        T* __restrict__ key, const T* __restrict__ cos_sin_cache,
                        const int64_t* __restrict__ offsets, int rot_dim, int64_t query_stride,
                        int64_t key_stride, int num_heads, int num_kv_heads, int head_size) {
  const int64_t token = blockIdx.x;
  const int64_t pos = positions[token] + (offsets ? offsets[token] : 0);
  const T* cache_ptr = cos_sin_cache + pos * rot_dim;
  rotate_row<T, NEOX>(query + token * query_stride, cache_ptr, num_heads, head_size, rot_dim);
  rotate_row<T, NEOX>(key + token * key_stride, cache_ptr, num_kv_heads, head_size, rot_dim);
}

// -------------------------------------------------------------------------------------
// fused rotary + reshape_and_cache: one CTA per token.
//   phase 1: rotate q in place; rotate k in place;            (global, same thread owns x and y)
//   phase 2: after a CTA barrier the threads scatter the just-rotated key row (written by this
//            CTA: L1/L2 hit) and the value row into the paged layout; consecutive threads fill
//            one 16-byte K chunk; fp8 quantisation in the same pass.
// -------------------------------------------------------------------------------------
template <typename T, typename CT, int KV>
__device__ __forceinline__ CT enc(T v, float scale) {
  if constexpr (KV == B200_KV_AUTO)
    return v;
  else
    return fp8_encode<KV>(to_f32<T>(v), scale);
}

template <typename T, typename CT, int KV, bool NEOX>
__global__ void __launch_bounds__(512)
rotary_reshape_and_cache_kernel(const int64_t* __restrict__ positions, T* __restrict__ query,
                                T* __restrict__ key, const T* __restrict__ value,
                                const T* __restrict__ cos_sin_cache,
                                const int64_t* __restrict__ offsets, CT* __restrict__ key_cache,
                                CT* __restrict__ value_cache,
                                const int64_t* __restrict__ slot_mapping, int rot_dim,
                                int64_t query_stride, int64_t key_stride, int64_t value_stride,
                                int num_heads, int num_kv_heads, int head_size, int block_size,
                                int x, float k_scale, float v_scale) {
  // like reshape_and_cache: the attention launch behind this kernel is a programmatic dependent that reads nothing
  // (query, cache) before this grid has completed -- let it become resident now
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t token = blockIdx.x;
  const int64_t pos = positions[token] + (offsets ? offsets[token] : 0);
  const T* cache_ptr = cos_sin_cache + pos * rot_dim;
  T* krow = key + token * key_stride;
  rotate_row<T, NEOX>(query + token * query_stride, cache_ptr, num_heads, head_size, rot_dim);
  rotate_row<T, NEOX>(krow, cache_ptr, num_kv_heads, head_size, rot_dim);

  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padding token: rotated, not cached (cache_kernels.cu:165-169)
  __syncthreads();       // the rotated key row (global writes of this CTA) is visible CTA-wide
  const int64_t blk = slot / block_size, off = slot % block_size;
  const T* vrow = value + token * value_stride;
  const int64_t page = (int64_t)head_size * block_size;
  CT* kdst_blk = key_cache + blk * num_kv_heads * page;
  CT* vdst_blk = value_cache + blk * num_kv_heads * page;
  const int n = num_kv_heads * head_size;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int h = i / head_size, d = i - h * head_size;
    const int xi = d / x, xo = d - xi * x;
    kdst_blk[h * page + ((int64_t)xi * block_size + off) * x + xo] = enc<T, CT, KV>(krow[i], k_scale);
    vdst_blk[h * page + (int64_t)d * block_size + off] = enc<T, CT, KV>(vrow[i], v_scale);
  }
}


// -------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------
template <typename T>
static int launch_rotary(const int64_t* positions, void* query, void* key, const void* cache,
                         const int64_t* offsets, int num_tokens, int num_heads, int num_kv_heads,
                         int head_size, int rot_dim, int64_t query_stride, int64_t key_stride,
                         int is_neox, cudaStream_t stream) {
  const int threads = std::min(512, std::max(32, (num_heads * rot_dim / 2 + 31) / 32 * 32));
  auto q = static_cast<T*>(query);
  auto k = static_cast<T*>(key);
  auto c = static_cast<const T*>(cache);
  if (is_neox)
    rotary_embedding_kernel<T, true><<<num_tokens, threads, 0, stream>>>(
        positions, q, k, c, offsets, rot_dim, query_stride, key_stride, num_heads, num_kv_heads,
        head_size);
  else
    rotary_embedding_kernel<T, false><<<num_tokens, threads, 0, stream>>>(
        positions, q, k, c, offsets, rot_dim, query_stride, key_stride, num_heads, num_kv_heads,
        head_size);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

template <typename T, typename CT, int KV>
static int launch_fused(const int64_t* positions, void* query, void* key, const void* value,
                        const void* cache, const int64_t* offsets, void* key_cache,
                        void* value_cache, const int64_t* slot_mapping, int num_tokens,
                        int num_heads, int num_kv_heads, int head_size, int rot_dim, int block_size,
                        int x, int64_t query_stride, int64_t key_stride, int64_t value_stride,
                        int is_neox, float k_scale, float v_scale, cudaStream_t stream) {
  const int work = std::max(num_heads * rot_dim / 2, num_kv_heads * head_size);
  const int threads = std::min(512, std::max(32, (work + 31) / 32 * 32));
  auto q = static_cast<T*>(query);
  auto k = static_cast<T*>(key);
  auto v = static_cast<const T*>(value);
  auto c = static_cast<const T*>(cache);
  auto kc = static_cast<CT*>(key_cache);
  auto vc = static_cast<CT*>(value_cache);
  if (is_neox)
    rotary_reshape_and_cache_kernel<T, CT, KV, true><<<num_tokens, threads, 0, stream>>>(
        positions, q, k, v, c, offsets, kc, vc, slot_mapping, rot_dim, query_stride, key_stride,
        value_stride, num_heads, num_kv_heads, head_size, block_size, x, k_scale, v_scale);
  else
    rotary_reshape_and_cache_kernel<T, CT, KV, false><<<num_tokens, threads, 0, stream>>>(
        positions, q, k, v, c, offsets, kc, vc, slot_mapping, rot_dim, query_stride, key_stride,
        value_stride, num_heads, num_kv_heads, head_size, block_size, x, k_scale, v_scale);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

template <typename T>
static int dispatch_fused_kv(int kv_dtype, const int64_t* positions, void* query, void* key,
                             const void* value, const void* cache, const int64_t* offsets,
                             void* key_cache, void* value_cache, const int64_t* slot_mapping,
                             int num_tokens, int num_heads, int num_kv_heads, int head_size,
                             int rot_dim, int block_size, int x, int64_t query_stride,
                             int64_t key_stride, int64_t value_stride, int is_neox, float k_scale,
                             float v_scale, cudaStream_t stream) {
#define B200_FUSED(CT, KV)                                                                        \
  return launch_fused<T, CT, KV>(positions, query, key, value, cache, offsets, key_cache,         \
                                 value_cache, slot_mapping, num_tokens, num_heads, num_kv_heads,  \
                                 head_size, rot_dim, block_size, x, query_stride, key_stride,     \
                                 value_stride, is_neox, k_scale, v_scale, stream)
  switch (kv_dtype) {
    case B200_KV_AUTO: B200_FUSED(T, B200_KV_AUTO);
    case B200_KV_FP8_E4M3: B200_FUSED(uint8_t, B200_KV_FP8_E4M3);
    case B200_KV_FP8_E5M2: B200_FUSED(uint8_t, B200_KV_FP8_E5M2);
    default: set_error("Unsupported data type of kv cache: %d", kv_dtype); return B200_ERR_UNSUPPORTED;
  }
#undef B200_FUSED
}

}  // namespace b200

using namespace b200;

extern "C" int b200_advance_step(int num_seqs, int num_queries, int block_size,
                                 int64_t* input_tokens, const int64_t* sampled_token_ids,
                                 int64_t* input_positions, int32_t* seq_lens,
                                 int64_t* slot_mapping, const int32_t* block_tables,
                                 int64_t block_tables_stride, int device, b200_stream_t stream) {
  B200_CHECK_ARG(num_queries >= 0 && num_queries <= num_seqs && block_size > 0,
                 "advance_step: need 0 <= num_queries (%d) <= num_seqs (%d), block_size > 0",
                 num_queries, num_seqs);
  if (num_queries == 0) return B200_OK;
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  advance_step_kernel<<<(num_queries + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      num_queries, block_size, input_tokens, sampled_token_ids, input_positions, seq_lens,
      slot_mapping, block_tables, block_tables_stride);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_rotary_embedding(const int64_t* positions, void* query, void* key,
                                     const void* cos_sin_cache, const int64_t* cos_sin_cache_offsets,
                                     int num_tokens, int num_heads, int num_kv_heads, int head_size,
                                     int rot_dim, int64_t query_stride, int64_t key_stride,
                                     int is_neox, int dtype, int device, b200_stream_t stream) {
  B200_CHECK_ARG(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size,
                 "rotary_embedding: rot_dim (%d) must be even and <= head_size (%d)", rot_dim,
                 head_size);
  if (num_tokens == 0) return B200_OK;
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case B200_DT_F32:
      return launch_rotary<float>(positions, query, key, cos_sin_cache, cos_sin_cache_offsets,
                                  num_tokens, num_heads, num_kv_heads, head_size, rot_dim,
                                  query_stride, key_stride, is_neox, st);
    case B200_DT_F16:
      return launch_rotary<__half>(positions, query, key, cos_sin_cache, cos_sin_cache_offsets,
                                   num_tokens, num_heads, num_kv_heads, head_size, rot_dim,
                                   query_stride, key_stride, is_neox, st);
    case B200_DT_BF16:
      return launch_rotary<__nv_bfloat16>(positions, query, key, cos_sin_cache,
                                          cos_sin_cache_offsets, num_tokens, num_heads, num_kv_heads,
                                          head_size, rot_dim, query_stride, key_stride, is_neox, st);
    default: set_error("Unsupported data type: %d", dtype); return B200_ERR_UNSUPPORTED;
  }
}

extern "C" int b200_rotary_reshape_and_cache(
    const int64_t* positions, void* query, void* key, const void* value, const void* cos_sin_cache,
    const int64_t* cos_sin_cache_offsets, void* key_cache, void* value_cache,
    const int64_t* slot_mapping, int num_tokens, int num_heads, int num_kv_heads, int head_size,
    int rot_dim, int block_size, int x, int64_t query_stride, int64_t key_stride,
    int64_t value_stride, int is_neox, float k_scale, float v_scale, int dtype, int kv_dtype,
    int device, b200_stream_t stream) {
  B200_CHECK_ARG(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size,
                 "rotary_reshape_and_cache: rot_dim (%d) must be even and <= head_size (%d)",
                 rot_dim, head_size);
  B200_CHECK_ARG(block_size > 0 && x > 0 && head_size % x == 0,
                 "rotary_reshape_and_cache: head_size (%d) must be a multiple of x (%d)", head_size, x);
  if (num_tokens == 0) return B200_OK;
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  cudaStream_t st = (cudaStream_t)stream;
#define B200_ARGS                                                                                \
  kv_dtype, positions, query, key, value, cos_sin_cache, cos_sin_cache_offsets, key_cache,       \
      value_cache, slot_mapping, num_tokens, num_heads, num_kv_heads, head_size, rot_dim,        \
      block_size, x, query_stride, key_stride, value_stride, is_neox, k_scale, v_scale, st
  switch (dtype) {
    case B200_DT_F32: return dispatch_fused_kv<float>(B200_ARGS);
    case B200_DT_F16: return dispatch_fused_kv<__half>(B200_ARGS);
    case B200_DT_BF16: return dispatch_fused_kv<__nv_bfloat16>(B200_ARGS);
    default: set_error("Unsupported data type: %d", dtype); return B200_ERR_UNSUPPORTED;
  }
#undef B200_ARGS
}
