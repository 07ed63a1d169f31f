// Parameter block shared by the prefix / chunked-prefill kernels (prefix_prefill.cu: mma.sync tile loop;
// prefix_prefill_umma.cuh: tcgen05 / TMEM tile loop).
#pragma once

#include "common.cuh"

namespace b200 {

struct PrefixParams {
  const void *q, *k, *v;
  void* out;
  const void *k_cache, *v_cache;
  const int32_t *block_tables, *start_loc, *seq_lens, *ctx_lens;
  const float* alibi_slopes;
  const unsigned char* kv_scratch;  // new k/v pre-staged as pages (prefix_stage_kv_kernel) or NULL
  int num_heads, num_kv_heads, group, max_blocks_per_seq, sliding_window, tok_tile, num_stages,
      stage_shift;
  int64_t scratch_slices;  // capacity of kv_scratch in (slice, all kv heads) units: staged indices are guarded
  int64_t q_s0, q_s1, k_s0, k_s1, v_s0, v_s1, o_s0, o_s1, kc_bs, kc_hs, vc_bs, vc_hs;
  float scale_log2;
};

// Four CUtensorMap objects (opaque, 128 bytes, 64-byte aligned) for the TMA-fed tcgen05 tile loop: K and V of the
// cache, K and V halves of the staged new-token slices.  Encoded on the host per call (prefix_prefill.cu).
struct alignas(64) PrefixTmaMaps {
  unsigned char k[128], v[128], sk[128], sv[128];
};

}  // namespace b200
