// Instantiations of the paged-attention kernel for query dtype bf16, KV cache format auto.
#define INST_T __nv_bfloat16
#define INST_KV B200_KV_AUTO
#define INST_KV_IS_AUTO 1
#define INST_NAME attn_lookup_bf16_auto
#include "paged_attention_inst.inc"
