// Thin torch binding over the C ABI of libb200_paged.so (include/b200_paged.h).
//
// Registers the reference's operator surface with the reference's exact schemas so that
// light_vllm/backends/_custom_ops.py works unmodified against this library:
//   torch.ops._C.paged_attention_v1 / _v2                       csrc/torch_bindings.cpp:24-48
//   torch.ops._C_cache_ops.{swap_blocks,copy_blocks,reshape_and_cache,
//                           reshape_and_cache_flash,convert_fp8}  csrc/torch_bindings.cpp:227-265
//   torch.ops._C_cuda_utils.{get_device_attribute,
//                            get_max_shared_memory_per_block_device_attribute}  :267-280
// All compute ops are registered for the CUDA dispatch key only (as in the reference): there is
// no CPU implementation behind them.  This file unpacks tensors, validates what the reference
// validates, reads the current CUDA stream, calls the C ABI and turns non-zero status codes into
// TORCH_CHECK failures (-> RuntimeError in Python, the reference's error convention).
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>
#include <torch/types.h>

#include <optional>
#include <string>
#include <vector>

#include "../../include/b200_paged.h"

namespace {

int dt_code(at::ScalarType t, const char* what) {
  switch (t) {
    case at::ScalarType::Float: return B200_DT_F32;
    case at::ScalarType::Half: return B200_DT_F16;
    case at::ScalarType::BFloat16: return B200_DT_BF16;
    default: TORCH_CHECK(false, "Unsupported data type: ", what, " is ", c10::toString(t));
  }
  return -1;
}

int kv_code(const std::string& s) {
  const int kv = b200_parse_kv_cache_dtype(s.c_str());
  // reference message: quant_utils.cuh:566-567
  TORCH_CHECK(kv >= 0, "Unsupported data type of kv cache: ", s);
  return kv;
}

inline void check(int rc) { TORCH_CHECK(rc == B200_OK, b200_last_error()); }

inline b200_stream_t cur_stream() {
  return reinterpret_cast<b200_stream_t>(at::cuda::getCurrentCUDAStream().stream());
}

void check_cache_dtype(const torch::Tensor& q, const torch::Tensor& cache, int kv) {
  if (kv == B200_KV_AUTO) {
    TORCH_CHECK(cache.scalar_type() == q.scalar_type(),
                "kv_cache_dtype 'auto' needs a cache of the query dtype, got ",
                c10::toString(cache.scalar_type()), " vs ", c10::toString(q.scalar_type()));
  } else {
    TORCH_CHECK(cache.element_size() == 1, "fp8 kv cache must be a 1-byte dtype (uint8), got ",
                c10::toString(cache.scalar_type()));
  }
}

struct AttnArgs {
  int num_seqs, num_heads, head_size, max_blocks, q_dtype, kv_dtype, device;
  int64_t q_stride, kv_block_stride, kv_head_stride;
  const float* alibi;
  b200_blocksparse_t bs;
};

AttnArgs unpack_attn(const torch::Tensor& out, const torch::Tensor& query,
                     const torch::Tensor& key_cache, const torch::Tensor& value_cache,
                     const torch::Tensor& block_tables, const torch::Tensor& seq_lens,
                     const std::optional<torch::Tensor>& alibi_slopes,
                     const std::string& kv_cache_dtype, int64_t tp_rank, int64_t local_blocks,
                     int64_t vert_stride, int64_t bs_block_size, int64_t head_sliding_step) {
  AttnArgs a{};
  TORCH_CHECK(query.dim() == 3, "query must be [num_seqs, num_heads, head_size]");
  TORCH_CHECK(query.is_cuda() && out.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() &&
                  block_tables.is_cuda() && seq_lens.is_cuda(),
              "paged_attention: all tensors must be CUDA tensors");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(out.scalar_type() == query.scalar_type(), "out and query dtypes differ");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query must be dense in its last two dims");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && block_tables.dim() == 2 &&
                  block_tables.is_contiguous(),
              "block_tables must be a contiguous int32 [num_seqs, max_blocks] tensor");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int && seq_lens.is_contiguous(),
              "seq_lens must be a contiguous int32 tensor");
  a.num_seqs = (int)query.size(0);
  a.num_heads = (int)query.size(1);
  a.head_size = (int)query.size(2);
  TORCH_CHECK(block_tables.size(0) >= a.num_seqs && seq_lens.numel() >= a.num_seqs,
              "block_tables / seq_lens shorter than the batch");
  a.max_blocks = (int)block_tables.size(1);
  a.q_stride = query.stride(0);
  a.kv_block_stride = key_cache.stride(0);
  a.kv_head_stride = key_cache.stride(1);
  a.q_dtype = dt_code(query.scalar_type(), "query");
  a.kv_dtype = kv_code(kv_cache_dtype);
  check_cache_dtype(query, key_cache, a.kv_dtype);
  check_cache_dtype(query, value_cache, a.kv_dtype);
  a.alibi = nullptr;
  if (alibi_slopes.has_value() && alibi_slopes->defined()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::ScalarType::Float && alibi_slopes->is_cuda() &&
                    alibi_slopes->is_contiguous() && alibi_slopes->numel() >= a.num_heads,
                "alibi_slopes must be a contiguous float32 CUDA tensor of num_heads entries");
    a.alibi = alibi_slopes->data_ptr<float>();
  }
  a.bs.tp_rank = (int32_t)tp_rank;
  a.bs.local_blocks = (int32_t)local_blocks;
  a.bs.vert_stride = (int32_t)vert_stride;
  a.bs.block_size = (int32_t)bs_block_size;
  a.bs.head_sliding_step = (int32_t)head_sliding_step;
  a.device = query.get_device();
  return a;
}

// torch.ops._C.paged_attention_v1 (csrc/ops.h:8-16)
void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
                        torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double k_scale, double v_scale,
                        const int64_t tp_rank, const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride,
                        const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  const AttnArgs a = unpack_attn(out, query, key_cache, value_cache, block_tables, seq_lens,
                                 alibi_slopes, kv_cache_dtype, tp_rank, blocksparse_local_blocks,
                                 blocksparse_vert_stride, blocksparse_block_size,
                                 blocksparse_head_sliding_step);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check(b200_paged_attention_v1(
      out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      block_tables.data_ptr<int32_t>(), seq_lens.data_ptr<int32_t>(), a.alibi, a.num_seqs,
      a.num_heads, (int)num_kv_heads, a.head_size, (int)block_size, (int)max_seq_len, a.max_blocks,
      a.q_stride, a.kv_block_stride, a.kv_head_stride, (float)scale, (float)k_scale,
      (float)v_scale, a.q_dtype, a.kv_dtype, &a.bs, a.device, cur_stream()));
}

// torch.ops._C.paged_attention_v2 (csrc/ops.h:18-27)
void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
                        torch::Tensor& tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
                        torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double k_scale, double v_scale,
                        const int64_t tp_rank, const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride,
                        const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  const AttnArgs a = unpack_attn(out, query, key_cache, value_cache, block_tables, seq_lens,
                                 alibi_slopes, kv_cache_dtype, tp_rank, blocksparse_local_blocks,
                                 blocksparse_vert_stride, blocksparse_block_size,
                                 blocksparse_head_sliding_step);
  TORCH_CHECK(exp_sums.scalar_type() == at::ScalarType::Float &&
                  max_logits.scalar_type() == at::ScalarType::Float &&
                  exp_sums.is_contiguous() && max_logits.is_contiguous() &&
                  tmp_out.is_contiguous() && tmp_out.scalar_type() == query.scalar_type(),
              "paged_attention_v2: exp_sums/max_logits must be contiguous float32 and tmp_out a "
              "contiguous tensor of the query dtype");
  TORCH_CHECK(exp_sums.dim() == 3 && max_logits.dim() == 3 && tmp_out.dim() == 4 &&
                  exp_sums.size(2) == max_logits.size(2) && exp_sums.size(2) == tmp_out.size(2),
              "paged_attention_v2: scratch shapes must be [B,H,P], [B,H,P], [B,H,P,D]");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check(b200_paged_attention_v2(
      out.data_ptr(), exp_sums.data_ptr<float>(), max_logits.data_ptr<float>(), tmp_out.data_ptr(),
      query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      block_tables.data_ptr<int32_t>(), seq_lens.data_ptr<int32_t>(), a.alibi, a.num_seqs,
      a.num_heads, (int)num_kv_heads, a.head_size, (int)block_size, (int)max_seq_len, a.max_blocks,
      (int)exp_sums.size(2), a.q_stride, a.kv_block_stride, a.kv_head_stride, (float)scale,
      (float)k_scale, (float)v_scale, a.q_dtype, a.kv_dtype, &a.bs, a.device, cur_stream()));
}

// torch.ops._C_cache_ops.swap_blocks (csrc/cache.h:8-9, cache_kernels.cu:24-63)
void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const torch::Tensor& block_mapping) {
  const torch::Device sd = src.device(), dd = dst.device();
  int s_dev, d_dev;
  if (sd.is_cuda() && dd.is_cuda()) {
    TORCH_CHECK(sd.index() == dd.index(), "src and dst must be on the same GPU");
    s_dev = sd.index();
    d_dev = dd.index();
  } else if (sd.is_cuda() && dd.is_cpu()) {
    s_dev = sd.index();
    d_dev = -1;
  } else if (sd.is_cpu() && dd.is_cuda()) {
    s_dev = -1;
    d_dev = dd.index();
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  TORCH_CHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long,
              "block_mapping must be an int64 tensor");
  const int64_t n = block_mapping.dim() >= 1 ? block_mapping.size(0) : 0;
  if (n == 0) return;
  TORCH_CHECK(block_mapping.dim() == 2 && block_mapping.size(1) == 2,
              "block_mapping must be [num_pairs, 2]");
  TORCH_CHECK(src.is_contiguous() && dst.is_contiguous(), "swap_blocks: caches must be contiguous");
  const torch::Tensor map = block_mapping.contiguous();
  const int64_t block_bytes = src.element_size() * src[0].numel();
  const at::cuda::OptionalCUDAGuard guard(sd.is_cuda() ? sd : dd);
  check(b200_swap_blocks(src.data_ptr(), dst.data_ptr(), map.data_ptr<int64_t>(), (int)n,
                         block_bytes, s_dev, d_dev, cur_stream()));
}

// torch.ops._C_cache_ops.copy_blocks (csrc/cache.h:14-16, cache_kernels.cu:101-148)
void copy_blocks(std::vector<torch::Tensor> const& key_caches,
                 std::vector<torch::Tensor> const& value_caches,
                 const torch::Tensor& block_mapping) {
  const int num_layers = (int)key_caches.size();
  TORCH_CHECK(num_layers == (int)value_caches.size());
  if (num_layers == 0) return;
  const torch::Device dev = key_caches[0].device();
  TORCH_CHECK(dev.is_cuda());
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long,
              "block_mapping must be an int64 tensor");
  const int64_t n = block_mapping.dim() >= 1 ? block_mapping.size(0) : 0;
  if (n == 0) return;
  TORCH_CHECK(block_mapping.dim() == 2 && block_mapping.size(1) == 2 &&
                  block_mapping.is_contiguous(),
              "block_mapping must be a contiguous [num_pairs, 2] tensor");
  // The kernel dereferences block_mapping on the device (as the reference does, :146): it must be
  // a CUDA tensor or pinned host memory (what light-vllm passes, model_input_builder.py:93-96).
  TORCH_CHECK(block_mapping.is_cuda() || block_mapping.is_pinned(),
              "copy_blocks: block_mapping must be a CUDA tensor or a pinned CPU tensor");
  std::vector<void*> kp(num_layers), vp(num_layers);
  const int64_t block_bytes = key_caches[0].element_size() * key_caches[0][0].numel();
  for (int l = 0; l < num_layers; ++l) {
    TORCH_CHECK(key_caches[l].device() == dev && value_caches[l].device() == dev,
                "copy_blocks: all caches must live on one device");
    kp[l] = key_caches[l].data_ptr();
    vp[l] = value_caches[l].data_ptr();
  }
  const at::cuda::OptionalCUDAGuard guard(dev);
  check(b200_copy_blocks(kp.data(), vp.data(), num_layers, block_mapping.data_ptr<int64_t>(),
                         (int)n, block_bytes, dev.index(), cur_stream()));
}

// torch.ops._C_cache_ops.reshape_and_cache (csrc/cache.h:18-22, cache_kernels.cu:263-289)
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                       torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, const double k_scale,
                       const double v_scale) {
  TORCH_CHECK(key.dim() == 3 && value.dim() == 3 && key_cache.dim() == 5 && value_cache.dim() == 4,
              "reshape_and_cache: key/value [T,H,D], key_cache [NB,H,D/x,BS,x], value_cache "
              "[NB,H,D,BS]");
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() &&
                  slot_mapping.is_cuda(),
              "reshape_and_cache: all tensors must be CUDA tensors");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(),
              "slot_mapping must be a contiguous int64 tensor");
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous(),
              "reshape_and_cache: caches must be contiguous");
  TORCH_CHECK(key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key/value must be dense in their last two dims");
  const int kv = kv_code(kv_cache_dtype);
  check_cache_dtype(key, key_cache, kv);
  check_cache_dtype(key, value_cache, kv);
  const int num_tokens = (int)std::min<int64_t>(key.size(0), slot_mapping.numel());
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  check(b200_reshape_and_cache(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(),
                               value_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(), num_tokens,
                               (int)key.size(1), (int)key.size(2), (int)key_cache.size(3),
                               (int)key_cache.size(4), key.stride(0), value.stride(0),
                               (float)k_scale, (float)v_scale, dt_code(key.scalar_type(), "key"),
                               kv, key.get_device(), cur_stream()));
}

// torch.ops._C_cache_ops.reshape_and_cache_flash (csrc/cache.h:24-29, cache_kernels.cu:304-330)
void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                             torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                             const std::string& kv_cache_dtype, const double k_scale,
                             const double v_scale) {
  TORCH_CHECK(key.dim() == 3 && value.dim() == 3 && key_cache.dim() == 4 && value_cache.dim() == 4,
              "reshape_and_cache_flash: key/value [T,H,D], caches [NB,BS,H,D]");
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() &&
                  slot_mapping.is_cuda(),
              "reshape_and_cache_flash: all tensors must be CUDA tensors");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(),
              "slot_mapping must be a contiguous int64 tensor");
  TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0));
  TORCH_CHECK(key_cache[0].is_contiguous() && value_cache[0].is_contiguous(),
              "reshape_and_cache_flash: blocks must be dense");
  TORCH_CHECK(key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key/value must be dense in their last two dims");
  const int kv = kv_code(kv_cache_dtype);
  check_cache_dtype(key, key_cache, kv);
  check_cache_dtype(key, value_cache, kv);
  const int num_tokens = (int)std::min<int64_t>(key.size(0), slot_mapping.numel());
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  check(b200_reshape_and_cache_flash(
      key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      slot_mapping.data_ptr<int64_t>(), num_tokens, (int)key.size(1), (int)key.size(2),
      (int)key_cache.size(1), key_cache.stride(0), key.stride(0), value.stride(0), (float)k_scale,
      (float)v_scale, dt_code(key.scalar_type(), "key"), kv, key.get_device(), cur_stream()));
}

// torch.ops._C_cache_ops.convert_fp8 (csrc/cache.h:31-33, cache_kernels.cu:355-405)
void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache, const double scale,
                 const std::string& kv_cache_dtype) {
  const torch::Device sd = src_cache.device(), dd = dst_cache.device();
  TORCH_CHECK(sd.is_cuda(), "src must be on a GPU");
  TORCH_CHECK(dd.is_cuda(), "dst must be on a GPU");
  TORCH_CHECK(sd.index() == dd.index(), "src and dst must be on the same GPU");
  // the reference only converts for "fp8" / "fp8_e4m3" ("auto" reaches an assert(false) stub,
  // quant_utils.cuh:515-526; anything else raises, cache_kernels.cu:403)
  TORCH_CHECK(kv_cache_dtype == "fp8" || kv_cache_dtype == "fp8_e4m3",
              "Unsupported data type: ", kv_cache_dtype);
  TORCH_CHECK(src_cache.is_contiguous() && dst_cache.is_contiguous() &&
                  src_cache.numel() == dst_cache.numel(),
              "convert_fp8: src and dst must be contiguous and of equal numel");
  const bool src_wide = src_cache.element_size() > 1;
  const bool dst_wide = dst_cache.element_size() > 1;
  TORCH_CHECK(src_wide != dst_wide, "convert_fp8: exactly one of src/dst must be an fp8 (uint8) cache");
  const int wide = dt_code(src_wide ? src_cache.scalar_type() : dst_cache.scalar_type(),
                           "convert_fp8 wide side");
  const at::cuda::OptionalCUDAGuard guard(sd);
  check(b200_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(),
                         (float)scale, wide, B200_KV_FP8_E4M3, src_wide ? 1 : 0, sd.index(),
                         cur_stream()));
}

// torch.ops._C.advance_step (csrc/ops.h:57-60, csrc/prepare_inputs/advance_step.cu:75-130)
void verify_tensor(const char* name, const torch::Tensor& t, int64_t size_0, int64_t size_1,
                   c10::ScalarType type) {
  const bool ok = (size_0 == -1 || t.size(0) == size_0) &&
                  (size_1 == -1 || (t.dim() > 1 && t.size(1) == size_1)) && t.is_contiguous() &&
                  t.scalar_type() == type && t.is_cuda();
  TORCH_CHECK(ok, "tensor: name = ", name, ", shape = ", t.sizes(), " is_cont = ",
              t.is_contiguous(), ", type = ", t.dtype(), " is not as expected: shape = [", size_0,
              ", ", size_1, "], type = ", type);
}

void advance_step(int64_t num_seqs, int64_t num_queries, int64_t block_size,
                  torch::Tensor& input_tokens, torch::Tensor& sampled_token_ids,
                  torch::Tensor& input_positions, torch::Tensor& seq_lens,
                  torch::Tensor& slot_mapping, torch::Tensor& block_tables) {
  verify_tensor("input_tokens", input_tokens, num_seqs, -1, at::kLong);
  verify_tensor("sampled_token_ids", sampled_token_ids, num_queries, 1, at::kLong);
  verify_tensor("input_positions", input_positions, num_seqs, -1, at::kLong);
  verify_tensor("seq_lens", seq_lens, num_seqs, -1, at::kInt);
  verify_tensor("slot_mapping", slot_mapping, num_seqs, -1, at::kLong);
  verify_tensor("block_tables", block_tables, num_seqs, -1, at::kInt);
  const at::cuda::OptionalCUDAGuard guard(device_of(sampled_token_ids));
  check(b200_advance_step((int)num_seqs, (int)num_queries, (int)block_size,
                          input_tokens.data_ptr<int64_t>(), sampled_token_ids.data_ptr<int64_t>(),
                          input_positions.data_ptr<int64_t>(), seq_lens.data_ptr<int32_t>(),
                          slot_mapping.data_ptr<int64_t>(), block_tables.data_ptr<int32_t>(),
                          block_tables.stride(0), sampled_token_ids.get_device(), cur_stream()));
}

struct RopeArgs {
  int num_tokens, num_heads, num_kv_heads, rot_dim, dtype;
  int64_t query_stride, key_stride;
};

RopeArgs unpack_rope(const torch::Tensor& positions, const torch::Tensor& query,
                     const torch::Tensor& key, int64_t head_size,
                     const torch::Tensor& cos_sin_cache, int64_t num_tokens) {
  TORCH_CHECK(positions.is_cuda() && query.is_cuda() && key.is_cuda() && cos_sin_cache.is_cuda(),
              "rotary_embedding: all tensors must be CUDA tensors");
  TORCH_CHECK(positions.scalar_type() == at::kLong && positions.is_contiguous(),
              "positions must be a contiguous int64 tensor");
  TORCH_CHECK(query.scalar_type() == key.scalar_type() &&
                  query.scalar_type() == cos_sin_cache.scalar_type(),
              "query, key and cos_sin_cache must share one dtype");
  TORCH_CHECK(cos_sin_cache.dim() == 2 && cos_sin_cache.is_contiguous(),
              "cos_sin_cache must be a contiguous [max_position, rot_dim] tensor");
  TORCH_CHECK(query.stride(-1) == 1 && key.stride(-1) == 1 && head_size > 0 &&
                  query.size(-1) % head_size == 0 && key.size(-1) % head_size == 0,
              "query/key last dim must be dense and a multiple of head_size");
  RopeArgs a{};
  a.num_tokens = (int)num_tokens;
  a.num_heads = (int)(query.size(-1) / head_size);
  a.num_kv_heads = (int)(key.size(-1) / head_size);
  a.rot_dim = (int)cos_sin_cache.size(1);
  a.query_stride = query.stride(-2);
  a.key_stride = key.stride(-2);
  a.dtype = dt_code(query.scalar_type(), "query");
  return a;
}

// torch.ops._C.rotary_embedding (csrc/ops.h:35-37, pos_encoding_kernels.cu:119-160)
void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key,
                      int64_t head_size, torch::Tensor& cos_sin_cache, bool is_neox) {
  const RopeArgs a = unpack_rope(positions, query, key, head_size, cos_sin_cache,
                                 query.numel() / query.size(-1));
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check(b200_rotary_embedding(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(),
                              cos_sin_cache.data_ptr(), nullptr, a.num_tokens, a.num_heads,
                              a.num_kv_heads, (int)head_size, a.rot_dim, a.query_stride,
                              a.key_stride, is_neox ? 1 : 0, a.dtype, query.get_device(),
                              cur_stream()));
}

// torch.ops._C.batched_rotary_embedding (csrc/ops.h:39-43, pos_encoding_kernels.cu:166-200)
void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key,
                              int64_t head_size, torch::Tensor& cos_sin_cache, bool is_neox,
                              int64_t rot_dim, torch::Tensor& cos_sin_cache_offsets) {
  TORCH_CHECK(cos_sin_cache_offsets.is_cuda() && cos_sin_cache_offsets.scalar_type() == at::kLong &&
                  cos_sin_cache_offsets.is_contiguous(),
              "cos_sin_cache_offsets must be a contiguous int64 CUDA tensor");
  RopeArgs a = unpack_rope(positions, query, key, head_size, cos_sin_cache,
                           cos_sin_cache_offsets.size(0));
  a.rot_dim = (int)rot_dim;  // the batched op takes rot_dim explicitly (:176,:183)
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check(b200_rotary_embedding(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(),
                              cos_sin_cache.data_ptr(), cos_sin_cache_offsets.data_ptr<int64_t>(),
                              a.num_tokens, a.num_heads, a.num_kv_heads, (int)head_size, a.rot_dim,
                              a.query_stride, a.key_stride, is_neox ? 1 : 0, a.dtype,
                              query.get_device(), cur_stream()));
}

// torch.ops._C_cache_ops.rotary_embedding_reshape_and_cache: B200 extension = rotary_embedding +
// reshape_and_cache in one launch (arguments = the union of the two reference ops).
void rotary_embedding_reshape_and_cache(torch::Tensor& positions, torch::Tensor& query,
                                        torch::Tensor& key, torch::Tensor& value, int64_t head_size,
                                        torch::Tensor& cos_sin_cache, bool is_neox,
                                        torch::Tensor& key_cache, torch::Tensor& value_cache,
                                        torch::Tensor& slot_mapping,
                                        const std::string& kv_cache_dtype, const double k_scale,
                                        const double v_scale) {
  const RopeArgs a = unpack_rope(positions, query, key, head_size, cos_sin_cache,
                                 query.numel() / query.size(-1));
  TORCH_CHECK(value.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() &&
                  slot_mapping.is_cuda(), "rotary_embedding_reshape_and_cache: CUDA tensors only");
  TORCH_CHECK(value.scalar_type() == key.scalar_type() && value.stride(-1) == 1 &&
                  value.size(-1) == key.size(-1), "value must match key in dtype and width");
  TORCH_CHECK(slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous() &&
                  slot_mapping.numel() >= a.num_tokens,
              "slot_mapping must be a contiguous int64 tensor of num_tokens entries");
  TORCH_CHECK(key_cache.dim() == 5 && value_cache.dim() == 4 && key_cache.is_contiguous() &&
                  value_cache.is_contiguous(),
              "caches must be dense [nb, kh, d/x, bs, x] and [nb, kh, d, bs]");
  const int kv = kv_code(kv_cache_dtype);
  check_cache_dtype(key, key_cache, kv);
  check_cache_dtype(key, value_cache, kv);
  TORCH_CHECK(key_cache.size(1) == a.num_kv_heads && value_cache.size(2) == head_size,
              "cache shape does not match key heads / head_size");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check(b200_rotary_reshape_and_cache(
      positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(), value.data_ptr(),
      cos_sin_cache.data_ptr(), nullptr, key_cache.data_ptr(), value_cache.data_ptr(),
      slot_mapping.data_ptr<int64_t>(), a.num_tokens, a.num_heads, a.num_kv_heads, (int)head_size,
      a.rot_dim, (int)key_cache.size(3), (int)key_cache.size(4), a.query_stride, a.key_stride,
      value.stride(-2), is_neox ? 1 : 0, (float)k_scale, (float)v_scale, a.dtype, kv,
      query.get_device(), cur_stream()));
}

int64_t get_device_attribute(int64_t attribute, int64_t device_id) {
  const int64_t v = b200_get_device_attribute(attribute, device_id);
  TORCH_CHECK(v >= 0, b200_last_error());
  return v;
}
int64_t get_max_shared_memory_per_block_device_attribute(int64_t device_id) {
  const int64_t v = b200_get_max_shared_memory_per_block_device_attribute(device_id);
  TORCH_CHECK(v >= 0, b200_last_error());
  return v;
}

}  // namespace

TORCH_LIBRARY_FRAGMENT(_C, ops) {
  ops.def(
      "paged_attention_v1("
      "    Tensor! out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, float k_scale, float v_scale,"
      "    int tp_rank, int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def(
      "paged_attention_v2("
      "    Tensor! out, Tensor exp_sums, Tensor max_logits,"
      "    Tensor tmp_out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, float k_scale, float v_scale,"
      "    int tp_rank, int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);
  // adjacent per-step ops (csrc/torch_bindings.cpp:75-77, 93-108)
  ops.def("advance_step", &advance_step);
  ops.impl("advance_step", torch::kCUDA, &advance_step);
  ops.def(
      "rotary_embedding(Tensor positions, Tensor! query,"
      "                 Tensor! key, int head_size,"
      "                 Tensor cos_sin_cache, bool is_neox) -> ()");
  ops.impl("rotary_embedding", torch::kCUDA, &rotary_embedding);
  ops.def(
      "batched_rotary_embedding(Tensor positions, Tensor! query,"
      "                         Tensor! key, int head_size,"
      "                         Tensor cos_sin_cache, bool is_neox,"
      "                         int rot_dim,"
      "                         Tensor cos_sin_cache_offsets) -> ()");
  ops.impl("batched_rotary_embedding", torch::kCUDA, &batched_rotary_embedding);
}

TORCH_LIBRARY_FRAGMENT(_C_cache_ops, cache_ops) {
  cache_ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  cache_ops.impl("swap_blocks", torch::kCUDA, &swap_blocks);
  cache_ops.def(
      "copy_blocks(Tensor[]! key_caches, Tensor[]! value_caches, Tensor block_mapping) -> ()");
  cache_ops.impl("copy_blocks", torch::kCUDA, &copy_blocks);
  cache_ops.def(
      "reshape_and_cache(Tensor key, Tensor value,"
      "                  Tensor! key_cache, Tensor! value_cache,"
      "                  Tensor slot_mapping,"
      "                  str kv_cache_dtype,"
      "                  float k_scale, float v_scale) -> ()");
  cache_ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  cache_ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value,"
      "                        Tensor! key_cache,"
      "                        Tensor! value_cache,"
      "                        Tensor slot_mapping,"
      "                        str kv_cache_dtype,"
      "                        float k_scale, float v_scale) -> ()");
  cache_ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  cache_ops.def(
      "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()");
  cache_ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
  // B200 extension: rotary_embedding + reshape_and_cache in one launch
  cache_ops.def(
      "rotary_embedding_reshape_and_cache(Tensor positions, Tensor! query, Tensor! key,"
      "                                   Tensor value, int head_size, Tensor cos_sin_cache,"
      "                                   bool is_neox, Tensor! key_cache, Tensor! value_cache,"
      "                                   Tensor slot_mapping, str kv_cache_dtype,"
      "                                   float k_scale, float v_scale) -> ()");
  cache_ops.impl("rotary_embedding_reshape_and_cache", torch::kCUDA,
                 &rotary_embedding_reshape_and_cache);
}

TORCH_LIBRARY_FRAGMENT(_C_cuda_utils, cuda_utils) {
  cuda_utils.def("get_device_attribute(int attribute, int device_id) -> int");
  cuda_utils.impl("get_device_attribute", &get_device_attribute);
  cuda_utils.def("get_max_shared_memory_per_block_device_attribute(int device_id) -> int");
  cuda_utils.impl("get_max_shared_memory_per_block_device_attribute",
                  &get_max_shared_memory_per_block_device_attribute);
}
