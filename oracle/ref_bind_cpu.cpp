// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product path.
//
// Harness-owned torch binding that exposes the *unmodified* reference CPU
// implementation (compiled in place from /root/reference/csrc/cpu/attention.cpp
// and cache.cpp by oracle/Makefile) under the namespace `_ref_cpu`, so tests
// and the cpu_baseline leg of bench.py can call the reference's own code.
//
// Schemas mirror the reference's CPU registration
// (csrc/cpu/torch_bindings.cpp:16-40 and :101-115).  No reference source is
// copied; only its public declarations (csrc/ops.h, csrc/cache.h) are included.
#include "cache.h"
#include "ops.h"

#include <torch/library.h>

TORCH_LIBRARY(_ref_cpu, ops) {
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
  ops.impl("paged_attention_v1", torch::kCPU, &paged_attention_v1);
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
  ops.impl("paged_attention_v2", torch::kCPU, &paged_attention_v2);
  ops.def(
      "copy_blocks(Tensor[]! key_caches, Tensor[]! value_caches, Tensor "
      "block_mapping) -> ()");
  ops.impl("copy_blocks", torch::kCPU, &copy_blocks);
  ops.def(
      "reshape_and_cache(Tensor key, Tensor value,"
      "                  Tensor! key_cache, Tensor! value_cache,"
      "                  Tensor slot_mapping,"
      "                  str kv_cache_dtype,"
      "                  float k_scale, float v_scale) -> ()");
  ops.impl("reshape_and_cache", torch::kCPU, &reshape_and_cache);
}
