// Instantiations of the paged-attention kernel for query dtype f16, KV cache format auto.
#define INST_T __half
#define INST_KV B200_KV_AUTO
#define INST_KV_IS_AUTO 1
#define INST_NAME attn_lookup_f16_auto
#include "paged_attention_inst.inc"
