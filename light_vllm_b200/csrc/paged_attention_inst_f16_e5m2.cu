// Instantiations of the paged-attention kernel for query dtype f16, KV cache format e5m2.
#define INST_T __half
#define INST_KV B200_KV_FP8_E5M2
#define INST_KV_IS_AUTO 0
#define INST_NAME attn_lookup_f16_e5m2
#include "paged_attention_inst.inc"
