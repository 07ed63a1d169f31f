/*
 * paged_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the algorithm of light-vllm's decode hot path, used as the parity
 * checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 * Nothing in light_vllm_b200/ (the product) imports, links or executes this file.
 *
 * Each function cites the reference file:line it follows (paths relative to the reference
 * checkout).  Pinning: this restatement is checked (tests/test_oracle.py, tests/golden/) against
 * outputs of the reference's OWN compiled code -- csrc/cpu/{attention,cache}.cpp built in place by
 * oracle/Makefile (oracle/_ref/libref_cpu.so, fp32/bf16, block 16) and, on the GPU box,
 * csrc/attention/attention_kernels.cu + csrc/cache_kernels.cu built for sm_100a
 * (oracle/_ref/libref_cuda.so, fp16 / fp8 / block 8,32 / blocksparse).  The reference's own test
 * suite holds no golden vectors for this path.
 *
 * Numerics: integer / byte / fp8 work is restated bit-exactly.  Attention keeps the CUDA
 * reference's rounding points that matter for the error budget (fp32 logits, exp(x - max),
 * 1/(sum + 1e-6), probabilities rounded to T before P.V, per-partition outputs rounded to T in
 * v2) but accumulates P.V in fp32 where the reference multiplies in T (dtype_float16.cuh:445-457).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { DT_F32 = 0, DT_F16 = 1, DT_BF16 = 2 };
enum { KV_AUTO = 0, KV_FP8_E4M3 = 1, KV_FP8_E5M2 = 2 };
#define PARTITION_SIZE 512 /* attention_kernels.cu:850 */

/* ------------------------------------------------------------------ scalar codecs ---- */
static inline float f16_to_f32(uint16_t h) {
  _Float16 v;
  memcpy(&v, &h, 2);
  return (float)v;
}
static inline uint16_t f32_to_f16(float f) { /* round-to-nearest-even (cvt.rn.f16.f32) */
  _Float16 v = (_Float16)f;
  uint16_t h;
  memcpy(&h, &v, 2);
  return h;
}
static inline float bf16_to_f32(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_bf16(float f) { /* __float2bfloat16_rn */
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fff; /* NaN */
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static inline float load_elem(const void* p, int64_t i, int dt) {
  if (dt == DT_F32) return ((const float*)p)[i];
  if (dt == DT_F16) return f16_to_f32(((const uint16_t*)p)[i]);
  return bf16_to_f32(((const uint16_t*)p)[i]);
}
static inline void store_elem(void* p, int64_t i, int dt, float v) {
  if (dt == DT_F32) ((float*)p)[i] = v;
  else if (dt == DT_F16) ((uint16_t*)p)[i] = f32_to_f16(v);
  else ((uint16_t*)p)[i] = f32_to_bf16(v);
}
static inline float round_to(float v, int dt) { /* value after an RN store to T */
  if (dt == DT_F32) return v;
  if (dt == DT_F16) return f16_to_f32(f32_to_f16(v));
  return bf16_to_f32(f32_to_bf16(v));
}
static inline int elem_size(int dt) { return dt == DT_F32 ? 4 : 2; }

/* fp8 (OCP E4M3 "fn" / E5M2).  decode = __nv_cvt_fp8_to_halfraw (exact) then half->float
 * (quant_utils.cuh:294-315).  encode = __nv_cvt_float_to_fp8(x, __NV_SATFINITE, kind): round to
 * nearest even, saturate to +-max finite (quant_utils.cuh:458-489). */
static double fp8_value(int code7, int kv) { /* magnitude of the 7-bit code; <0 for NaN/Inf */
  if (kv == KV_FP8_E4M3) {
    int e = code7 >> 3, m = code7 & 7;
    if (code7 == 0x7f) return -1.0;
    return e == 0 ? ldexp((double)m, -9) : ldexp(1.0 + m / 8.0, e - 7);
  } else {
    int e = code7 >> 2, m = code7 & 3;
    if (e == 31) return -1.0;
    return e == 0 ? ldexp((double)m, -16) : ldexp(1.0 + m / 4.0, e - 15);
  }
}
static float fp8_decode(uint8_t v, int kv) {
  double mag = fp8_value(v & 0x7f, kv);
  if (mag < 0) return (kv == KV_FP8_E5M2 && (v & 3) == 0) ? ((v & 0x80) ? -INFINITY : INFINITY) : NAN;
  return (float)((v & 0x80) ? -mag : mag);
}
static uint8_t fp8_encode(float x, int kv) {
  const int max_code = kv == KV_FP8_E4M3 ? 0x7e : 0x7b;
  uint8_t sign = signbit(x) ? 0x80 : 0;
  if (isnan(x)) return 0x7f; /* cvt.rn.satfinite.e4m3x2/e5m2x2: NaN -> 0x7f */
  double a = fabs((double)x);
  if (a >= fp8_value(max_code, kv)) return sign | (uint8_t)max_code;
  int lo = 0, hi = max_code; /* largest code with value <= a */
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (fp8_value(mid, kv) <= a) lo = mid; else hi = mid - 1;
  }
  double below = fp8_value(lo, kv), above = fp8_value(lo + 1, kv);
  double dl = a - below, dh = above - a;
  int code = (dl < dh) ? lo : (dl > dh) ? lo + 1 : ((lo & 1) ? lo + 1 : lo);
  return sign | (uint8_t)code;
}
/* fp8::scaled_convert<fp8, T>(a, scale): fp32 DIVISION then saturating RN (quant_utils.cuh:458-489) */
static inline uint8_t fp8_quant(float x, float scale, int kv) { return fp8_encode(x / scale, kv); }
/* fp8::scaled_convert<T, fp8>(a, scale): half(fp8) -> float * scale -> RN to T (:294-359,:406-415) */
static inline float fp8_dequant_f32(uint8_t v, float scale, int kv) { return fp8_decode(v, kv) * scale; }

/* Exported scalar helpers so the tests can pin the codecs exhaustively. */
uint8_t oracle_fp8_encode(float x, int kv) { return fp8_encode(x, kv); }
float oracle_fp8_decode(uint8_t v, int kv) { return fp8_decode(v, kv); }
uint16_t oracle_f32_to_f16(float f) { return f32_to_f16(f); }
uint16_t oracle_f32_to_bf16(float f) { return f32_to_bf16(f); }

/* ------------------------------------------------------------------ cache ops -------- */
/* One source element -> its cache bytes (copy for "auto", scaled fp8 quantisation otherwise). */
static inline void put_cache(void* cache, int64_t idx, const void* src, int64_t sidx, int dt, int kv,
                             float scale) {
  if (kv == KV_AUTO) {
    const int es = elem_size(dt);
    memcpy((char*)cache + idx * es, (const char*)src + sidx * es, es);
  } else {
    ((uint8_t*)cache)[idx] = fp8_quant(load_elem(src, sidx, dt), scale, kv);
  }
}

/* reshape_and_cache_kernel, csrc/cache_kernels.cu:152-204:
 * key_cache [nb, H, D/x, BS, x], value_cache [nb, H, D, BS]; slot < 0 => skip (:165-169). */
int oracle_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                             const int64_t* slot_mapping, int num_tokens, int num_heads,
                             int head_size, int block_size, int x, int64_t key_stride,
                             int64_t value_stride, float k_scale, float v_scale, int dt, int kv) {
  const int n = num_heads * head_size;
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue;
    const int64_t blk = slot / block_size, off = slot % block_size;
    for (int i = 0; i < n; ++i) {
      const int h = i / head_size, d = i % head_size;
      const int xi = d / x, xo = d % x;
      const int64_t kidx = blk * num_heads * (head_size / x) * block_size * x +
                           (int64_t)h * (head_size / x) * block_size * x +
                           (int64_t)xi * block_size * x + off * x + xo; /* :184-189 */
      const int64_t vidx = blk * num_heads * head_size * block_size +
                           (int64_t)h * head_size * block_size + (int64_t)d * block_size + off; /* :190-193 */
      put_cache(key_cache, kidx, key, t * key_stride + i, dt, kv, k_scale);
      put_cache(value_cache, vidx, value, t * value_stride + i, dt, kv, v_scale);
    }
  }
  return 0;
}

/* reshape_and_cache_flash_kernel, csrc/cache_kernels.cu:206-247: caches [nb, BS, H, D]. */
int oracle_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                   void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                   int num_heads, int head_size, int block_size,
                                   int64_t block_stride, int64_t key_stride, int64_t value_stride,
                                   float k_scale, float v_scale, int dt, int kv) {
  const int n = num_heads * head_size;
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue; /* :222-225 */
    const int64_t blk = slot / block_size, off = slot % block_size;
    for (int i = 0; i < n; ++i) {
      const int64_t idx = blk * block_stride + off * n + i; /* :233-236 */
      put_cache(key_cache, idx, key, t * key_stride + i, dt, kv, k_scale);
      put_cache(value_cache, idx, value, t * value_stride + i, dt, kv, v_scale);
    }
  }
  return 0;
}

/* copy_blocks_kernel, csrc/cache_kernels.cu:68-94: per layer, per (src,dst) pair, K then V.
 * Pairs are applied in order (the CUDA grid runs them concurrently; callers never chain them). */
int oracle_copy_blocks(void* const* key_caches, void* const* value_caches, int num_layers,
                       const int64_t* block_mapping, int num_pairs, int64_t block_bytes) {
  for (int l = 0; l < num_layers; ++l)
    for (int p = 0; p < num_pairs; ++p) {
      const int64_t s = block_mapping[2 * p], d = block_mapping[2 * p + 1];
      memmove((char*)key_caches[l] + d * block_bytes, (char*)key_caches[l] + s * block_bytes, block_bytes);
      memmove((char*)value_caches[l] + d * block_bytes, (char*)value_caches[l] + s * block_bytes, block_bytes);
    }
  return 0;
}

/* swap_blocks, csrc/cache_kernels.cu:24-63: dst[d] = src[s] for every mapping row, in order. */
int oracle_swap_blocks(const void* src, void* dst, const int64_t* block_mapping, int num_pairs,
                       int64_t block_bytes) {
  for (int p = 0; p < num_pairs; ++p)
    memcpy((char*)dst + block_mapping[2 * p + 1] * block_bytes,
           (const char*)src + block_mapping[2 * p] * block_bytes, block_bytes);
  return 0;
}

/* convert_fp8_kernel, csrc/cache_kernels.cu:334-345 with fp8::scaled_convert
 * (quant_utils.cuh:287-526).  to_fp8: dst u8 = quant(src T); else dst T = RN_T(dequant(src u8)). */
int oracle_convert_fp8(void* dst, const void* src, int64_t numel, float scale, int wide_dt, int kv,
                       int to_fp8) {
  if (kv != KV_FP8_E4M3 && kv != KV_FP8_E5M2) return 2;
  for (int64_t i = 0; i < numel; ++i) {
    if (to_fp8) ((uint8_t*)dst)[i] = fp8_quant(load_elem(src, i, wide_dt), scale, kv);
    else store_elem(dst, i, wide_dt, fp8_dequant_f32(((const uint8_t*)src)[i], scale, kv));
  }
  return 0;
}

/* ------------------------------------------------------------------ attention -------- */
typedef struct {
  const void *q, *k_cache, *v_cache;
  const int32_t *block_tables, *seq_lens;
  const float* alibi;
  int num_heads, num_kv_heads, head_size, block_size, max_blocks_per_seq;
  int64_t q_stride, kv_block_stride, kv_head_stride;
  float scale, k_scale, v_scale;
  int dt, kv;
  int bs_on, bs_tp_rank, bs_local_blocks, bs_vert_stride, bs_block_size, bs_head_sliding_step;
} attn_args;

/* K / V element as the kernel sees it after the load (+ fp8 dequant to T, attention_kernels.cu
 * :277-290 and :404-419: scaled_convert rounds the dequantised value to scalar_t). */
static inline float cache_elem(const void* cache, int64_t idx, const attn_args* a, float sc) {
  if (a->kv == KV_AUTO) return load_elem(cache, idx, a->dt);
  return round_to(fp8_dequant_f32(((const uint8_t*)cache)[idx], sc, a->kv), a->dt);
}

/* paged_attention_kernel (attention_kernels.cu:86-496) for one (seq, head) over tokens
 * [tok_begin, tok_end) of the sequence: returns the partition's max logit and exp sum and writes
 * the partition-normalised output (rounded to T by the caller). logits: scratch of >= n floats. */
static void attend_range(const attn_args* a, int seq, int head, int tok_begin, int tok_end,
                         float* logits, float* acc, float* max_out, float* sum_out) {
  const int D = a->head_size, BS = a->block_size;
  const int x = 16 / (a->kv == KV_AUTO ? elem_size(a->dt) : 1); /* :142 */
  const int seq_len = a->seq_lens[seq];
  const int kvh = head / (a->num_heads / a->num_kv_heads);       /* :150-153 */
  const int32_t* bt = a->block_tables + (int64_t)seq * a->max_blocks_per_seq;
  const float slope = a->alibi ? a->alibi[head] : 0.f;
  const int n = tok_end - tok_begin;
  int q_bs_block = 0, bs_off = 0;
  if (a->bs_on) { /* :209-225 */
    q_bs_block = (seq_len - 1) / a->bs_block_size;
    bs_off = a->bs_head_sliding_step >= 0
                 ? (a->bs_tp_rank * a->num_heads + head) * a->bs_head_sliding_step + 1
                 : (a->bs_tp_rank * a->num_kv_heads + kvh) * (-a->bs_head_sliding_step) + 1;
  }
  float qk_max = -3.402823466e+38f; /* -FLT_MAX, :143 */
  for (int t = tok_begin; t < tok_end; ++t) {
    const int blk_idx = t / BS, off = t % BS;
    if (a->bs_on) { /* :235-256: skipped blocks get -FLT_MAX logits and do not move the max */
      const int kb = blk_idx * BS / a->bs_block_size;
      const int remote = ((kb + bs_off) % a->bs_vert_stride) == 0;
      const int local = kb > q_bs_block - a->bs_local_blocks;
      if (!remote && !local) { logits[t - tok_begin] = -3.402823466e+38f; continue; }
    }
    const int64_t base = (int64_t)bt[blk_idx] * a->kv_block_stride + (int64_t)kvh * a->kv_head_stride;
    float dot = 0.f;
    for (int d = 0; d < D; ++d) { /* element (d, t) at (d/x)*BS*x + t*x + d%x, :273-286 */
      const float kx = cache_elem(a->k_cache, base + (int64_t)(d / x) * BS * x + off * x + d % x, a, a->k_scale);
      dot += load_elem(a->q, (int64_t)seq * a->q_stride + (int64_t)head * D + d, a->dt) * kx;
    }
    float qk = a->scale * dot;                                        /* :294 */
    qk += (slope != 0.f) ? slope * (float)(t - seq_len + 1) : 0.f;    /* :296 */
    logits[t - tok_begin] = qk;
    qk_max = fmaxf(qk_max, qk);                                       /* :299-306 */
  }
  float exp_sum = 0.f;
  for (int i = 0; i < n; ++i) { logits[i] = expf(logits[i] - qk_max); exp_sum += logits[i]; } /* :333-338 */
  const float inv = 1.f / (exp_sum + 1e-6f);                          /* :342 */
  for (int d = 0; d < D; ++d) acc[d] = 0.f;
  for (int t = tok_begin; t < tok_end; ++t) {
    const int blk_idx = t / BS, off = t % BS;
    if (a->bs_on) { /* :387-393 */
      const int vb = blk_idx * BS / a->bs_block_size;
      if (((vb + bs_off) % a->bs_vert_stride) != 0 && !(vb > q_bs_block - a->bs_local_blocks)) continue;
    }
    const float p = round_to(logits[t - tok_begin] * inv, a->dt);    /* from_float(logits_vec), :398-400 */
    const int64_t base = (int64_t)bt[blk_idx] * a->kv_block_stride + (int64_t)kvh * a->kv_head_stride;
    for (int d = 0; d < D; ++d) /* element (d, t) at d*BS + t, :402-412 */
      acc[d] += p * cache_elem(a->v_cache, base + (int64_t)d * BS + off, a, a->v_scale);
  }
  *max_out = qk_max;
  *sum_out = exp_sum;
}

static void fill_args(attn_args* a, const void* q, const void* kc, const void* vc, const int32_t* bt,
                      const int32_t* sl, const float* alibi, int num_heads, int num_kv_heads,
                      int head_size, int block_size, int max_blocks_per_seq, int64_t q_stride,
                      int64_t kv_block_stride, int64_t kv_head_stride, float scale, float k_scale,
                      float v_scale, int dt, int kv, const int32_t* bs5) {
  memset(a, 0, sizeof(*a));
  a->q = q; a->k_cache = kc; a->v_cache = vc; a->block_tables = bt; a->seq_lens = sl; a->alibi = alibi;
  a->num_heads = num_heads; a->num_kv_heads = num_kv_heads; a->head_size = head_size;
  a->block_size = block_size; a->max_blocks_per_seq = max_blocks_per_seq;
  a->q_stride = q_stride; a->kv_block_stride = kv_block_stride; a->kv_head_stride = kv_head_stride;
  a->scale = scale; a->k_scale = k_scale; a->v_scale = v_scale; a->dt = dt; a->kv = kv;
  if (bs5 && bs5[2] > 1) { /* is_block_sparse = blocksparse_vert_stride > 1, :825 */
    a->bs_on = 1; a->bs_tp_rank = bs5[0]; a->bs_local_blocks = bs5[1]; a->bs_vert_stride = bs5[2];
    a->bs_block_size = bs5[3]; a->bs_head_sliding_step = bs5[4];
  }
}

/* paged_attention_v1 (attention_kernels.cu:499-527, 808-829): one pass over the whole context.
 * bs5 = {tp_rank, local_blocks, vert_stride, block_size, head_sliding_step} or NULL. */
int oracle_paged_attention_v1(void* out, const void* q, const void* kc, const void* vc,
                              const int32_t* bt, const int32_t* sl, const float* alibi, int num_seqs,
                              int num_heads, int num_kv_heads, int head_size, int block_size,
                              int max_seq_len, int max_blocks_per_seq, int64_t q_stride,
                              int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                              float k_scale, float v_scale, int dt, int kv, const int32_t* bs5) {
  attn_args a;
  fill_args(&a, q, kc, vc, bt, sl, alibi, num_heads, num_kv_heads, head_size, block_size,
            max_blocks_per_seq, q_stride, kv_block_stride, kv_head_stride, scale, k_scale, v_scale,
            dt, kv, bs5);
  (void)max_seq_len;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int s = 0; s < num_seqs; ++s)
    for (int h = 0; h < num_heads; ++h) {
      const int seq_len = sl[s];
      float* logits = (float*)malloc(sizeof(float) * (size_t)(seq_len > 0 ? seq_len : 1));
      float* acc = (float*)malloc(sizeof(float) * (size_t)head_size);
      float m, l;
      attend_range(&a, s, h, 0, seq_len, logits, acc, &m, &l);
      for (int d = 0; d < head_size; ++d) /* from_float(*(out_ptr + row_idx), accs[i]), :484-495 */
        store_elem(out, ((int64_t)s * num_heads + h) * head_size + d, dt, acc[d]);
      free(logits);
      free(acc);
    }
  return 0;
}

/* paged_attention_v2 (attention_kernels.cu:530-562 + reduce :565-669, 848-934): 512-token
 * partitions write tmp_out / max_logits / exp_sums, then the merge
 * out = sum_j tmp_out_j * exp_sum_j e^(m_j - m) / (sum_j exp_sum_j e^(m_j - m) + 1e-6). */
int oracle_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                              const void* q, const void* kc, const void* vc, const int32_t* bt,
                              const int32_t* sl, const float* alibi, int num_seqs, int num_heads,
                              int num_kv_heads, int head_size, int block_size, int max_seq_len,
                              int max_blocks_per_seq, int max_num_partitions, int64_t q_stride,
                              int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                              float k_scale, float v_scale, int dt, int kv, const int32_t* bs5) {
  attn_args a;
  fill_args(&a, q, kc, vc, bt, sl, alibi, num_heads, num_kv_heads, head_size, block_size,
            max_blocks_per_seq, q_stride, kv_block_stride, kv_head_stride, scale, k_scale, v_scale,
            dt, kv, bs5);
  (void)max_seq_len;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int s = 0; s < num_seqs; ++s)
    for (int h = 0; h < num_heads; ++h) {
      const int seq_len = sl[s];
      const int np = (seq_len + PARTITION_SIZE - 1) / PARTITION_SIZE;
      float logits[PARTITION_SIZE];
      float* acc = (float*)malloc(sizeof(float) * (size_t)head_size);
      const int64_t hp = ((int64_t)s * num_heads + h) * max_num_partitions;
      for (int j = 0; j < np; ++j) { /* partition kernel; empty partitions exit early (:116-119) */
        const int t0 = j * PARTITION_SIZE;
        const int t1 = t0 + PARTITION_SIZE < seq_len ? t0 + PARTITION_SIZE : seq_len;
        float m, l;
        attend_range(&a, s, h, t0, t1, logits, acc, &m, &l);
        max_logits[hp + j] = m; /* :348-357 */
        exp_sums[hp + j] = l;
        for (int d = 0; d < head_size; ++d) store_elem(tmp_out, (hp + j) * head_size + d, dt, acc[d]);
      }
      float* o = acc;
      if (np == 1) { /* plain copy, :582-594 */
        for (int d = 0; d < head_size; ++d)
          store_elem(out, ((int64_t)s * num_heads + h) * head_size + d, dt,
                     load_elem(tmp_out, hp * head_size + d, dt));
      } else if (np > 1) {
        float m = -3.402823466e+38f, gsum = 0.f;
        for (int j = 0; j < np; ++j) m = fmaxf(m, max_logits[hp + j]);          /* :611-631 */
        for (int j = 0; j < np; ++j) gsum += exp_sums[hp + j] * expf(max_logits[hp + j] - m); /* :640-650 */
        const float inv = 1.f / (gsum + 1e-6f);                                 /* :651 */
        for (int d = 0; d < head_size; ++d) {                                   /* :660-667 */
          float v = 0.f;
          for (int j = 0; j < np; ++j)
            v += load_elem(tmp_out, (hp + j) * head_size + d, dt) *
                 (exp_sums[hp + j] * expf(max_logits[hp + j] - m)) * inv;
          o[d] = v;
        }
        for (int d = 0; d < head_size; ++d)
          store_elem(out, ((int64_t)s * num_heads + h) * head_size + d, dt, o[d]);
      }
      free(acc);
    }
  return 0;
}

/* ------------------------------------------------------------------ per-step ops ------ */
/* advance_step_kernel (csrc/prepare_inputs/advance_step.cu:15-58): per query, next token in,
 * seq_len + 1, position = new seq_len - 1, slot from the block table -- 32-bit int arithmetic. */
int oracle_advance_step(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                        const int64_t* sampled_token_ids, int64_t* input_positions,
                        int32_t* seq_lens, int64_t* slot_mapping, const int32_t* block_tables,
                        int64_t block_tables_stride) {
  (void)num_seqs;
  for (int q = 0; q < num_queries; ++q) {
    input_tokens[q] = sampled_token_ids[q];                       /* :35 */
    const int next_seq_len = seq_lens[q] + 1;                     /* :37-39 */
    const int next_pos = next_seq_len - 1;
    seq_lens[q] = next_seq_len;                                   /* :42 */
    input_positions[q] = next_pos;                                /* :44 */
    const int32_t* bt = block_tables + block_tables_stride * q;   /* :46-47 */
    const int slot = bt[next_pos / block_size] * block_size + next_pos % block_size; /* :49-53 */
    slot_mapping[q] = slot;                                       /* :55 */
  }
  return 0;
}

/* apply_token_rotary_embedding (csrc/pos_encoding_kernels.cu:10-36).  scalar_t arithmetic of the
 * reference: for c10::Half / c10::BFloat16 every operator computes in float and rounds back to T
 * (three roundings per output); for float nvcc contracts `x*cos - y*sin` to fma(x, cos, -(y*sin))
 * and `y*cos + x*sin` to fma(x, sin, y*cos) (SASS of all four float instantiations of the file
 * compiled for sm_100a). */
static void rotate_pair(void* arr, int64_t xi, int64_t yi, float c, float s, int dt) {
  const float x = load_elem(arr, xi, dt), y = load_elem(arr, yi, dt);
  float ox, oy;
  if (dt == DT_F32) {
    ox = fmaf(x, c, -(y * s));
    oy = fmaf(x, s, y * c);
  } else {
    ox = round_to(x * c, dt) - round_to(y * s, dt);
    oy = round_to(y * c, dt) + round_to(x * s, dt);
  }
  store_elem(arr, xi, dt, ox);
  store_elem(arr, yi, dt, oy);
}

/* rotary_embedding_kernel / batched_rotary_embedding_kernel (pos_encoding_kernels.cu:73-116):
 * in place on query [T, H, head_size] and key [T, KH, head_size] (token strides in elements);
 * cos_sin_cache [max_position, rot_dim] = cos | sin halves; offsets may be NULL. */
int oracle_rotary_embedding(const int64_t* positions, void* query, void* key,
                            const void* cos_sin_cache, const int64_t* offsets, int num_tokens,
                            int num_heads, int num_kv_heads, int head_size, int rot_dim,
                            int64_t query_stride, int64_t key_stride, int is_neox, int dt) {
  const int embed = rot_dim / 2;
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t pos = positions[t] + (offsets ? offsets[t] : 0);
    const int64_t cbase = pos * rot_dim;
    for (int pass = 0; pass < 2; ++pass) {
      void* arr = pass == 0 ? query : key;
      const int heads = pass == 0 ? num_heads : num_kv_heads;
      const int64_t stride = pass == 0 ? query_stride : key_stride;
      for (int h = 0; h < heads; ++h)
        for (int r = 0; r < embed; ++r) {
          const int64_t base = t * stride + (int64_t)h * head_size;
          const int64_t xi = is_neox ? r : 2 * r, yi = is_neox ? embed + r : 2 * r + 1; /* :16-28 */
          rotate_pair(arr, base + xi, base + yi, load_elem(cos_sin_cache, cbase + r, dt),
                      load_elem(cos_sin_cache, cbase + embed + r, dt), dt);
        }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ prefix prefill --- */
/* context_attention_fwd (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811;
 * kernels _fwd_kernel :13-243 and _fwd_kernel_alibi :437-672).  For sequence b, query token i
 * (position ctx_len + i, packed at row start_loc[b] + i) attends the ctx_len cached tokens read
 * through the block table (K layout [NB,KH,D/x,BS,x] :112-117, V layout [NB,KH,D,BS] :119-123,
 * no mask :125-131) and the new tokens j <= i read from the packed k/v tensors (causal mask
 * :205-207).  sm_scale = 1/sqrt(D) (:702).  Sliding window: keys with pos_q - pos_k >= window get
 * the logit -10000 (:132-148, :208-211).  ALiBi: bias slope * (pos_k - pos_q), entries with a
 * positive bias are masked (:551-557, :622-628); the ALiBi kernel takes no window, so with slopes
 * given the sliding window is ignored (:712-760).  The Triton kernel runs the softmax online over
 * 128-token tiles with fp32 statistics and rounds P to the value dtype before P.V (:170,:236);
 * this restatement computes the same softmax in one pass (fp32 logits, P rounded to T, fp32
 * accumulation) -- the tile order only changes fp32 summation order.
 * Pinned by tests/golden/prefix.pt = outputs of the reference kernel itself (Triton interpreter,
 * tests/golden/make_golden_prefix.py).  start_loc: int32 [B]. */
int oracle_context_attention_fwd(const void* q, const void* k, const void* v, void* out,
                                 const void* k_cache, const void* v_cache,
                                 const int32_t* block_tables, const int32_t* start_loc,
                                 const int32_t* seq_lens, const int32_t* ctx_lens, int num_seqs,
                                 int num_heads, int num_kv_heads, int head_size, int block_size,
                                 int max_blocks_per_seq, int64_t q_stride0, int64_t q_stride1,
                                 int64_t k_stride0, int64_t k_stride1, int64_t v_stride0,
                                 int64_t v_stride1, int64_t o_stride0, int64_t o_stride1,
                                 int64_t kc_block_stride, int64_t kc_head_stride,
                                 int64_t vc_block_stride, int64_t vc_head_stride,
                                 const float* alibi_slopes, int sliding_window, int dt) {
  const int D = head_size, BS = block_size;
  const int x = 16 / elem_size(dt);
  const float sm_scale = 1.0f / sqrtf((float)D);
  const int group = num_heads / num_kv_heads;
  int rc = 0;
#pragma omp parallel for collapse(2) schedule(dynamic)
  for (int b = 0; b < num_seqs; ++b) {
    for (int h = 0; h < num_heads; ++h) {
      const int ctx = ctx_lens[b], qlen = seq_lens[b] - ctx_lens[b];
      if (qlen <= 0) continue;
      const int kvh = h / group;
      const int32_t* bt = block_tables + (int64_t)b * max_blocks_per_seq;
      float* logits = (float*)malloc(sizeof(float) * (size_t)(ctx + qlen));
      float* acc = (float*)malloc(sizeof(float) * (size_t)D);
      if (!logits || !acc) { rc = 1; free(logits); free(acc); continue; }
      const float slope = alibi_slopes ? alibi_slopes[h] : 0.f;
      for (int i = 0; i < qlen; ++i) {
        const int64_t qrow = (int64_t)(start_loc[b] + i) * q_stride0 + (int64_t)h * q_stride1;
        const int pos_q = ctx + i, n = ctx + i + 1;
        float mx = -INFINITY;
        for (int t = 0; t < n; ++t) {
          float dot = 0.f;
          if (t < ctx) {
            const int64_t base = (int64_t)bt[t / BS] * kc_block_stride + (int64_t)kvh * kc_head_stride;
            for (int d = 0; d < D; ++d)
              dot += load_elem(q, qrow + d, dt) *
                     load_elem(k_cache, base + (int64_t)(d / x) * BS * x + (int64_t)(t % BS) * x + d % x, dt);
          } else {
            const int64_t krow = (int64_t)(start_loc[b] + t - ctx) * k_stride0 + (int64_t)kvh * k_stride1;
            for (int d = 0; d < D; ++d) dot += load_elem(q, qrow + d, dt) * load_elem(k, krow + d, dt);
          }
          float s = dot * sm_scale;
          if (alibi_slopes) s += slope * (float)(t - pos_q);
          if (!alibi_slopes && sliding_window > 0 && pos_q - t >= sliding_window) s = -10000.f;
          logits[t] = s;
          mx = fmaxf(mx, s);
        }
        float sum = 0.f;
        for (int t = 0; t < n; ++t) { logits[t] = expf(logits[t] - mx); sum += logits[t]; }
        for (int d = 0; d < D; ++d) acc[d] = 0.f;
        for (int t = 0; t < n; ++t) {
          const float p = round_to(logits[t] / sum, dt);
          if (p == 0.f) continue;
          if (t < ctx) {
            const int64_t base = (int64_t)bt[t / BS] * vc_block_stride + (int64_t)kvh * vc_head_stride;
            for (int d = 0; d < D; ++d) acc[d] += p * load_elem(v_cache, base + (int64_t)d * BS + t % BS, dt);
          } else {
            const int64_t vrow = (int64_t)(start_loc[b] + t - ctx) * v_stride0 + (int64_t)kvh * v_stride1;
            for (int d = 0; d < D; ++d) acc[d] += p * load_elem(v, vrow + d, dt);
          }
        }
        const int64_t orow = (int64_t)(start_loc[b] + i) * o_stride0 + (int64_t)h * o_stride1;
        for (int d = 0; d < D; ++d) store_elem(out, orow + d, dt, acc[d]);
      }
      free(logits);
      free(acc);
    }
  }
  return rc;
}
