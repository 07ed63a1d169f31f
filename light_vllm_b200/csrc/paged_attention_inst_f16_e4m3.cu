// Instantiations of the paged-attention kernel for query dtype f16, KV cache format e4m3.
#define INST_T __half
#define INST_KV B200_KV_FP8_E4M3
#define INST_KV_IS_AUTO 0
#define INST_NAME attn_lookup_f16_e4m3
#include "paged_attention_inst.inc"
