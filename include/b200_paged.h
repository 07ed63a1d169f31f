/*
 * b200_paged.h — C ABI of libb200_paged.so: the B200-native (sm_100a) decode hot path of
 * light-vllm: paged attention over the block-table KV cache plus the cache-write /
 * cache-management ops that feed it.
 *
 * This is the drop-in boundary.  Every entry point takes plain pointers, sizes and strides
 * (no torch types) and launches on the caller's stream without synchronising.  Each function
 * names the reference interface it replaces (paths relative to the reference checkout).
 *
 * Conventions
 *   - return value: 0 on success, a B200_ERR_* code otherwise; b200_last_error() returns a
 *     thread-local human-readable message for the last failing call on this thread
 *     (the reference raises via TORCH_CHECK -> RuntimeError; the torch binding converts
 *     non-zero codes to exactly that).
 *   - `device`: CUDA ordinal that owns the device pointers; the call makes it current for the
 *     duration of the call (reference: at::cuda::OptionalCUDAGuard, attention_kernels.cu:736).
 *   - `stream`: a cudaStream_t (reference: at::cuda::getCurrentCUDAStream(), :737).
 *   - strides are in ELEMENTS of the tensor they index and are 64-bit (the reference uses
 *     int, attention_kernels.cu:705-707; pools above 2^31 elements need 64 bits).
 *   - there is no CPU implementation behind any of these symbols: without a CUDA device they
 *     fail with B200_ERR_CUDA.
 *   - launch ordering: every call is ordered behind the work already queued on `stream`, as the
 *     reference's launches are.  The attention kernels are launched with programmatic stream
 *     serialization and start with griddepcontrol.wait: when the preceding kernel on the stream is
 *     b200_reshape_and_cache[_flash] (or another kernel that executes griddepcontrol.launch_dependents)
 *     their launch latency overlaps it, behind anything else they are ordinary launches.  Nothing is
 *     required of the caller, and the calls can be captured in CUDA graphs.
 */
#ifndef B200_PAGED_H_
#define B200_PAGED_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_PAGED_ABI_VERSION 2

typedef struct CUstream_st* b200_stream_t; /* == cudaStream_t */

/* Query / activation dtype (reference: query.dtype(), quant_utils.cuh:532-569). */
enum { B200_DT_F32 = 0, B200_DT_F16 = 1, B200_DT_BF16 = 2 };

/* KV-cache storage format (reference: Fp8KVCacheDataType, csrc/attention/dtype_fp8.cuh:14-18). */
enum { B200_KV_AUTO = 0, B200_KV_FP8_E4M3 = 1, B200_KV_FP8_E5M2 = 2 };

enum {
  B200_OK = 0,
  B200_ERR_INVALID = 1,     /* bad argument / inconsistent shapes                         */
  B200_ERR_UNSUPPORTED = 2, /* head size / block size / dtype combo not instantiated      */
  B200_ERR_CUDA = 3         /* CUDA runtime error (message carries cudaGetErrorString)    */
};

/* Blocksparse tail of the paged_attention schema (csrc/torch_bindings.cpp:31-34;
 * semantics attention_kernels.cu:209-256,387-393).  Pass NULL or vert_stride <= 1 for dense. */
typedef struct {
  int32_t tp_rank;
  int32_t local_blocks;
  int32_t vert_stride;
  int32_t block_size;
  int32_t head_sliding_step;
} b200_blocksparse_t;

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* b200_last_error(void);

int b200_abi_version(void);

/* Maps the reference's kv_cache_dtype strings "auto" | "fp8" | "fp8_e4m3" | "fp8_e5m2"
 * (quant_utils.cuh:533,544,555) to B200_KV_*; returns -1 for anything else. */
int b200_parse_kv_cache_dtype(const char* s);

/*
 * paged_attention_v1 — replaces torch.ops._C.paged_attention_v1
 *   (csrc/torch_bindings.cpp:24-35, csrc/attention/attention_kernels.cu:808-829, :690-770).
 *
 *   out          [num_seqs, num_heads, head_size]            T, dense
 *   query        [num_seqs, num_heads, head_size]            T, row stride q_stride
 *   key_cache    [num_blocks, num_kv_heads, head_size/x, block_size, x]   x = 16 / sizeof(cache elem)
 *   value_cache  [num_blocks, num_kv_heads, head_size, block_size]
 *                cache elem = T for B200_KV_AUTO, uint8 (fp8) otherwise;
 *                kv_block_stride / kv_head_stride index dim 0 / dim 1 of BOTH caches
 *   block_tables [num_seqs, max_num_blocks_per_seq] int32;  seq_lens [num_seqs] int32
 *   alibi_slopes [num_heads] float32 or NULL
 */
int b200_paged_attention_v1(void* out, const void* query, const void* key_cache,
                            const void* value_cache, const int32_t* block_tables,
                            const int32_t* seq_lens, const float* alibi_slopes, int num_seqs,
                            int num_heads, int num_kv_heads, int head_size, int block_size,
                            int max_seq_len, int max_num_blocks_per_seq, int64_t q_stride,
                            int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                            float k_scale, float v_scale, int q_dtype, int kv_dtype,
                            const b200_blocksparse_t* blocksparse, int device,
                            b200_stream_t stream);

/*
 * paged_attention_v2 — replaces torch.ops._C.paged_attention_v2
 *   (csrc/torch_bindings.cpp:37-48, attention_kernels.cu:973-997, :848-934, reduce :565-669).
 *   Split-context: 512-token partitions write
 *     tmp_out    [num_seqs, num_heads, max_num_partitions, head_size]  T
 *     max_logits [num_seqs, num_heads, max_num_partitions]  float32
 *     exp_sums   [num_seqs, num_heads, max_num_partitions]  float32
 *   (caller-owned scratch, paged_attn.py:156-166) and a second kernel merges them into `out`.
 *   max_num_partitions must be >= ceil(max_seq_len / 512).
 */
int b200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                            const void* query, const void* key_cache, const void* value_cache,
                            const int32_t* block_tables, const int32_t* seq_lens,
                            const float* alibi_slopes, int num_seqs, int num_heads,
                            int num_kv_heads, int head_size, int block_size, int max_seq_len,
                            int max_num_blocks_per_seq, int max_num_partitions, int64_t q_stride,
                            int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                            float k_scale, float v_scale, int q_dtype, int kv_dtype,
                            const b200_blocksparse_t* blocksparse, int device,
                            b200_stream_t stream);

/*
 * paged_attention_v2_static — NO counterpart in the reference.
 *   b200_paged_attention_v2 for callers that may not allocate (CUDA graph capture, tight serving loops,
 *   harnesses that forbid cudaMalloc): the caller passes inputs and `out` only; exp_sums / max_logits /
 *   tmp_out live in ONE library-owned static device array of B200_STATIC_SCRATCH_BYTES (16 MiB) that every GPU
 *   loading the library holds until exit.  Calls no cudaMalloc / cudaFree / allocating routine.
 *   Limit: 2 * align256(4 * cells) + cells * head_size * sizeof(T) <= B200_STATIC_SCRATCH_BYTES with
 *   cells = num_seqs * num_heads * ceil(max_seq_len / 512); larger calls are rejected (B200_ERR_INVALID, message
 *   names the sizes) -- use b200_paged_attention_v2 with caller scratch for those.
 *   Every call on a device shares the array and it keeps the last call's contents (nothing in it needs to
 *   start at zero, so nothing is cleared): calls that overlap on the GPU (different streams) RACE on it.
 *   The array's device address is resolved once per device at the first call (cudaGetSymbolAddress); make
 *   that first call outside stream capture.
 */
#define B200_STATIC_SCRATCH_BYTES (16 * 1024 * 1024)
int b200_paged_attention_v2_static(void* out, const void* query, const void* key_cache,
                                   const void* value_cache, const int32_t* block_tables,
                                   const int32_t* seq_lens, const float* alibi_slopes, int num_seqs,
                                   int num_heads, int num_kv_heads, int head_size, int block_size,
                                   int max_seq_len, int max_num_blocks_per_seq, int64_t q_stride,
                                   int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                                   float k_scale, float v_scale, int q_dtype, int kv_dtype,
                                   const b200_blocksparse_t* blocksparse, int device,
                                   b200_stream_t stream);

/*
 * reshape_and_cache — replaces torch.ops._C_cache_ops.reshape_and_cache
 *   (csrc/torch_bindings.cpp:241-247, csrc/cache_kernels.cu:263-289, kernel :152-204).
 *   key/value [num_tokens, num_heads, head_size] T with token strides key_stride/value_stride;
 *   caches dense in the paged layout above; slot_mapping [num_tokens] int64, slot < 0 = skip.
 *   kv_dtype != AUTO: stored value = fp8(float(x) / scale), saturating (quant_utils.cuh:458-489).
 */
int b200_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                           const int64_t* slot_mapping, int num_tokens, int num_heads,
                           int head_size, int block_size, int x, int64_t key_stride,
                           int64_t value_stride, float k_scale, float v_scale, int src_dtype,
                           int kv_dtype, int device, b200_stream_t stream);

/*
 * reshape_and_cache_flash — replaces torch.ops._C_cache_ops.reshape_and_cache_flash
 *   (csrc/torch_bindings.cpp:250-258, csrc/cache_kernels.cu:304-330, kernel :206-247).
 *   caches [num_blocks, block_size, num_heads, head_size] with dim-0 stride block_stride.
 */
int b200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                 void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                 int num_heads, int head_size, int block_size,
                                 int64_t block_stride, int64_t key_stride, int64_t value_stride,
                                 float k_scale, float v_scale, int src_dtype, int kv_dtype,
                                 int device, b200_stream_t stream);

/*
 * copy_blocks — replaces torch.ops._C_cache_ops.copy_blocks
 *   (csrc/torch_bindings.cpp:235-238, csrc/cache_kernels.cu:101-148, kernel :68-94).
 *   key_cache_ptrs/value_cache_ptrs: HOST arrays of num_layers device base pointers;
 *   block_mapping: [num_pairs, 2] int64 (src, dst), readable from the device (device memory or
 *   pinned host memory — light-vllm passes a pinned CPU tensor, model_input_builder.py:93-96);
 *   block_bytes = bytes of one block of one layer's K (== V) cache.  No host<->device sync.
 */
int b200_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int num_layers,
                     const int64_t* block_mapping, int num_pairs, int64_t block_bytes, int device,
                     b200_stream_t stream);

/*
 * swap_blocks — replaces torch.ops._C_cache_ops.swap_blocks
 *   (csrc/torch_bindings.cpp:230-232, csrc/cache_kernels.cu:24-63).
 *   block_mapping_host: HOST [num_pairs, 2] int64 (src block, dst block) — the reference requires a
 *   CPU tensor (:44).  src_device / dst_device: CUDA ordinal, or -1 for host memory.
 *   Host<->device pairs over pinned host memory run as one gather/scatter kernel per 224 pairs;
 *   pageable host memory falls back to run-merged cudaMemcpyAsync.  Stream-ordered, no sync.
 */
int b200_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host, int num_pairs,
                     int64_t block_bytes, int src_device, int dst_device, b200_stream_t stream);

/*
 * convert_fp8 — replaces torch.ops._C_cache_ops.convert_fp8
 *   (csrc/torch_bindings.cpp:261-264, csrc/cache_kernels.cu:355-405, kernel :334-345).
 *   Direction from dtypes: exactly one of src/dst is fp8 (uint8); `wide_dtype` is the B200_DT_* of
 *   the other side.  to_fp8 != 0: dst = fp8(float(src)/scale); else dst = T(float(half(src))*scale).
 */
int b200_convert_fp8(void* dst, const void* src, int64_t numel, float scale, int wide_dtype,
                     int kv_dtype, int to_fp8, int device, b200_stream_t stream);

/*
 * advance_step — replaces torch.ops._C.advance_step
 *   (csrc/torch_bindings.cpp:75-77, csrc/prepare_inputs/advance_step.cu:15-130).
 *   For every query i < num_queries: input_tokens[i] = sampled_token_ids[i]; seq_lens[i] += 1;
 *   input_positions[i] = seq_lens[i] - 1; slot_mapping[i] = block_tables[i][pos / block_size] *
 *   block_size + pos % block_size.  All device pointers; block_tables row stride in elements.
 *   Lets the next decode step's inputs be produced on the GPU (no host round trip, graph-capturable).
 */
int b200_advance_step(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                      const int64_t* sampled_token_ids, int64_t* input_positions,
                      int32_t* seq_lens, int64_t* slot_mapping, const int32_t* block_tables,
                      int64_t block_tables_stride, int device, b200_stream_t stream);

/*
 * rotary_embedding — replaces torch.ops._C.rotary_embedding and, with cos_sin_cache_offsets != NULL,
 *   torch.ops._C.batched_rotary_embedding (csrc/torch_bindings.cpp:93-108,
 *   csrc/pos_encoding_kernels.cu:9-200).  In-place GPT-NeoX (is_neox) or GPT-J rotation of the first
 *   rot_dim elements of every head of query [num_tokens, num_heads, head_size] and key
 *   [num_tokens, num_kv_heads, head_size] (token strides in elements); cos_sin_cache
 *   [max_position, rot_dim] of the same dtype; positions int64 [num_tokens].
 *   Rounding as the reference: products and the sum/difference are each rounded to the tensor dtype.
 */
int b200_rotary_embedding(const int64_t* positions, void* query, void* key,
                          const void* cos_sin_cache, const int64_t* cos_sin_cache_offsets,
                          int num_tokens, int num_heads, int num_kv_heads, int head_size,
                          int rot_dim, int64_t query_stride, int64_t key_stride, int is_neox,
                          int dtype, int device, b200_stream_t stream);

/*
 * rotary_reshape_and_cache — rotary_embedding followed by reshape_and_cache (both above) as ONE
 *   launch: the two reference ops light-vllm issues back to back for every layer
 *   (decode_only/modelzoo/qwen2.py:150-154 -> layers/rotary_embedding.py, then
 *   cache_kernels.cu:152-204).  query/key are rotated in place; the rotated key and `value` are
 *   scattered into the paged caches at slot_mapping (slot < 0: rotated but not cached), with the
 *   fp8 quantisation of reshape_and_cache.  Bit-identical to the two-op sequence.
 */
int b200_rotary_reshape_and_cache(const int64_t* positions, void* query, void* key,
                                  const void* value, const void* cos_sin_cache,
                                  const int64_t* cos_sin_cache_offsets, void* key_cache,
                                  void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                  int num_heads, int num_kv_heads, int head_size, int rot_dim,
                                  int block_size, int x, int64_t query_stride, int64_t key_stride,
                                  int64_t value_stride, int is_neox, float k_scale, float v_scale,
                                  int dtype, int kv_dtype, int device, b200_stream_t stream);

/*
 * get_device_attribute / get_max_shared_memory_per_block_device_attribute — replace
 *   torch.ops._C_cuda_utils.* (csrc/torch_bindings.cpp:271-279, csrc/cuda_utils_kernels.cu:8-29).
 *   Return the attribute value, or -1 on error.
 */
int64_t b200_get_device_attribute(int64_t attribute, int64_t device_id);
int64_t b200_get_max_shared_memory_per_block_device_attribute(int64_t device_id);

/*
 * context_attention_fwd — replaces the Triton kernels behind PagedAttention.forward_prefix
 *   (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811; caller
 *   ops/paged_attn.py:193-225): prefill of new tokens whose sequence already has ctx_len tokens in
 *   the paged cache (prefix-cache hit, chunked prefill).
 *     q / out      [num_tokens, num_heads, head_size]     T (fp16 | bf16 on the tensor-core kernel,
 *                  fp32 on a CUDA-core compatibility kernel), strides *_stride0/1
 *     k / v        [num_tokens, num_kv_heads, head_size]  T: the NEW tokens, packed per sequence
 *     key_cache    [num_blocks, num_kv_heads, head_size/x, block_size, x], value_cache
 *                  [num_blocks, num_kv_heads, head_size, block_size]  T (the reference kernel has no
 *                  fp8 path); only the first ctx_len tokens of each sequence are read from them
 *     block_tables [num_seqs, max_blocks_per_seq] int32; start_loc [num_seqs] int32 (first packed
 *                  row of each sequence); seq_lens = ctx_lens + query length, ctx_lens: int32
 *     alibi_slopes [num_heads] float32 or NULL; sliding_window <= 0 disables the window (ignored
 *                  with ALiBi, as in the reference dispatch :713-811)
 *     scale        the reference fixes 1/sqrt(head_size) (:702); passed explicitly here
 *     kv_scratch   optional device scratch of >= b200_context_attention_scratch_bytes(num_tokens =
 *                  rows of k/v, ...) bytes, 16-byte aligned: the new k/v are first rewritten into
 *                  page layout there (one extra launch) and then stream through the same bulk-copy
 *                  pipeline as the cached context; NULL (or too small) = each CTA stages them itself.
 *                  The staged mode relies on start_loc being the packed prefix sums of the query
 *                  lengths (non-decreasing in the sequence index), as query_start_loc[:-1] is.
 *   Query token i attends all ctx_len cached tokens and new tokens j <= i.
 */
int b200_context_attention_fwd(
    const void* q, const void* k, const void* v, void* out, const void* key_cache,
    const void* value_cache, const int32_t* block_tables, const int32_t* start_loc,
    const int32_t* seq_lens, const int32_t* ctx_lens, int num_seqs, int num_heads, int num_kv_heads,
    int head_size, int block_size, int max_blocks_per_seq, int max_query_len, int64_t q_stride0,
    int64_t q_stride1, int64_t k_stride0, int64_t k_stride1, int64_t v_stride0, int64_t v_stride1,
    int64_t o_stride0, int64_t o_stride1, int64_t kc_block_stride, int64_t kc_head_stride,
    int64_t vc_block_stride, int64_t vc_head_stride, const float* alibi_slopes, int sliding_window,
    float scale, void* kv_scratch, int64_t kv_scratch_bytes, int64_t num_tokens, int dtype,
    int device, b200_stream_t stream);
/* Same call with the tile loop chosen by the caller (tests, benchmarks, autotuning; per call, no
 * global state).  AUTO = b200_context_attention_fwd.  MMA_SYNC: csrc/prefix_prefill.cu (every head size,
 * ALiBi, sliding window, fp32 on CUDA cores).  TCGEN05: csrc/prefix_prefill_umma.cuh (tcgen05.mma with
 * TMEM accumulators, 256-row tiles; head size 64 or 128, fp16 / bf16, no ALiBi / window, needs kv_scratch and
 * 16-byte aligned q / out rows -- B200_ERR_UNSUPPORTED otherwise). */
#define B200_PREFIX_PATH_AUTO 0
#define B200_PREFIX_PATH_MMA_SYNC 1
#define B200_PREFIX_PATH_TCGEN05 2        /* tcgen05 tile loop; K/V tiles by TMA tensor copies when block size is 16 */
#define B200_PREFIX_PATH_TCGEN05_GATHER 3 /* tcgen05 tile loop with the per-thread 16-byte gather (any block size) */
int b200_context_attention_fwd_path(
    const void* q, const void* k, const void* v, void* out, const void* key_cache,
    const void* value_cache, const int32_t* block_tables, const int32_t* start_loc,
    const int32_t* seq_lens, const int32_t* ctx_lens, int num_seqs, int num_heads, int num_kv_heads,
    int head_size, int block_size, int max_blocks_per_seq, int max_query_len, int64_t q_stride0,
    int64_t q_stride1, int64_t k_stride0, int64_t k_stride1, int64_t v_stride0, int64_t v_stride1,
    int64_t o_stride0, int64_t o_stride1, int64_t kc_block_stride, int64_t kc_head_stride,
    int64_t vc_block_stride, int64_t vc_head_stride, const float* alibi_slopes, int sliding_window,
    float scale, void* kv_scratch, int64_t kv_scratch_bytes, int64_t num_tokens, int dtype,
    int device, b200_stream_t stream, int path);
int64_t b200_context_attention_scratch_bytes(int64_t num_tokens, int num_seqs, int num_kv_heads,
                                             int head_size, int block_size);

/*
 * Context-sharded decode (one sequence whose KV pages are spread over several
 * GPUs).  No reference twin: the reference keeps a sequence's pool on one device; the merge is
 * the reference's own v2 formula (attention_kernels.cu:565-669) applied across ranks.
 *
 *   b200_paged_attention_ctx_shard — every rank calls it in the same order with the pages IT owns:
 *   block_tables / local_seq_lens describe the local slice (0 tokens is fine), the scratch tensors
 *   are v2's ([B,H,P], [B,H,P], [B,H,P,D]; P >= ceil(max_local_seq_len / 512), >= 1).  The rank's
 *   partitions are folded into one (max, sum, out) triple per (seq, head), stored straight into
 *   every peer's exchange buffer over NVLink, and merged after an epoch-flag handshake — one kernel
 *   after the partition pass, no NCCL, no host synchronisation.  `out` on EVERY rank receives the
 *   attention over the whole context.  ALiBi and blocksparse are not defined for shards.
 *     exchange_bufs [world] host array: this process's addresses of every rank's exchange buffer
 *                   (own buffer from b200_peer_alloc, peers' from b200_peer_open)
 *     max_pairs     (seq, head) capacity the buffers were sized for (b200_ctx_exchange_bytes)
 *     epoch         1, 2, 3, ... incremented by the caller on every call, same on all ranks; or 0 on
 *                   every rank = device-managed (a counter in the exchange buffer, advanced by the
 *                   merge phase), which is what lets the call be captured in a CUDA graph
 *     phases        3 = partition pass, then publish + wait + merge in one kernel (the normal call);
 *                   1 = partition pass + publish only, 2 = wait + merge only (same epoch): a
 *                   split-phase form that lets the caller put other work between the two
 *   A peer that never arrives makes the wait time out (~2 s) and sets the status word
 *   (b200_ctx_exchange_status) instead of hanging the device.
 */
int64_t b200_ctx_exchange_bytes(int world, int max_pairs, int head_size);
int b200_paged_attention_ctx_shard(
    void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
    const void* key_cache, const void* value_cache, const int32_t* block_tables,
    const int32_t* local_seq_lens, int num_seqs, int num_heads, int num_kv_heads, int head_size,
    int block_size, int max_local_seq_len, int max_num_blocks_per_seq, int max_num_partitions,
    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride, float scale, float k_scale,
    float v_scale, int q_dtype, int kv_dtype, void* const* exchange_bufs, int rank, int world,
    int max_pairs, uint32_t epoch, int phases, int device, b200_stream_t stream);
int b200_ctx_exchange_status(const void* local_buf, int world, int32_t* status, int device);
/* A wait that times out is LOUD: the waiting rank's output rows of that call are NaN, status word 0
 * holds 1 + the rank that never published and word 3 the epoch of the failing call (sticky until
 * b200_ctx_exchange_clear_status).  b200_ctx_exchange_status_async copies the 4 status words into
 * pinned host memory on `stream`, so that the host can check call n when it issues call n+1 without
 * synchronising; b200_ctx_exchange_set_timeout sets this group's timeout (ms, 0 = 2000 ms default).
 * All state lives in the group's own exchange buffer. */
int b200_ctx_exchange_status_async(const void* local_buf, int world, int32_t* pinned_host4, int device,
                                   b200_stream_t stream);
int b200_ctx_exchange_set_timeout(void* local_buf, int world, int timeout_ms, int device);
/* Zero-copy failure mirror: 4 pinned (mapped) host words {status, -, -, failing epoch} that the exchange kernel
 * writes when a wait times out, so that the host can notice at its next call without copying the status word
 * back every step (NULL detaches). */
int b200_ctx_exchange_set_host_mirror(void* local_buf, int world, void* pinned_host4, int device);
/* Diagnostics (synchronises): CTA 0's timestamps of the last two exchange launches, out32[2][8][2] =
 * (globaltimer ns, SM clock) at {entry, dependency resolved, first statistics loaded, first row folded,
 * rows published, first pair arrived, merged, end}; slot = epoch & 1. */
int b200_ctx_exchange_trace(const void* local_buf, int world, uint64_t* out32, int device);
int b200_ctx_exchange_clear_status(void* local_buf, int world, int device);

/* Peer-visible device memory for the exchange buffers: plain cudaMalloc (zero-filled) exported /
 * opened through CUDA IPC handles (64 bytes) that the host side passes between the ranks. */
int b200_peer_alloc(void** ptr, int64_t bytes, int device);
int b200_peer_free(void* ptr, int device);
int b200_peer_export(const void* ptr, unsigned char* handle64, int device);
int b200_peer_open(const unsigned char* handle64, void** ptr, int device);
int b200_peer_close(void* ptr, int device);

/* Introspection used by bench.py / tests: launch geometry the library would pick. */
typedef struct {
  int32_t cluster_size;   /* v1: CTAs per (seq, kv-head) cooperating through DSMEM */
  int32_t num_warps;      /* warps per CTA                                        */
  int32_t num_stages;     /* K+V page pairs in flight per warp                    */
  int32_t smem_bytes;     /* dynamic shared memory per CTA                        */
  int32_t ctas_per_sm;    /* occupancy reported by the runtime                    */
  int32_t grid_x, grid_y, grid_z;
} b200_attn_plan_t;
/* Per-call launch overrides for benchmarking / autotuning (every field: 0 = library default).  The
 * library keeps NO mutable state between calls and reads no environment variables (the reference's
 * ops are stateless, and DP replicas call them from one thread per GPU,
 * prefill_only/executor/gpu_data_parallelism_executor.py:41-72): overrides travel as arguments. */
typedef struct {
  int32_t num_warps;   /* warps per CTA (1..8)                                              */
  int32_t num_stages;  /* K+V page pairs in flight per warp                                 */
  int32_t split;       /* v1: cluster size (context split through DSMEM); v2: partitions    */
                       /* per sequence at max_seq_len (each >= 512 tokens, balanced pages)  */
  int32_t head_rows;   /* q-heads of a GQA group per CTA (1..group; fewer than the group =   */
                       /* several CTAs share each KV page through L2: small-batch grids)     */
} b200_attn_tuning_t;
int b200_paged_attention_plan(int version, int num_seqs, int num_heads, int num_kv_heads,
                              int head_size, int block_size, int max_seq_len, int q_dtype,
                              int kv_dtype, int device, const b200_attn_tuning_t* tuning_or_null,
                              b200_attn_plan_t* plan);

/* b200_paged_attention_v1 (version 1; scratch arguments ignored) / _v2 (version 2) with an explicit
 * launch plan.  Same results as the plain entry points for every plan (tools/sweep_attn.py). */
int b200_paged_attention_tuned(
    int version, void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
    const void* key_cache, const void* value_cache, const int32_t* block_tables,
    const int32_t* seq_lens, const float* alibi_slopes, int num_seqs, int num_heads,
    int num_kv_heads, int head_size, int block_size, int max_seq_len, int max_num_blocks_per_seq,
    int max_num_partitions, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    float scale, float k_scale, float v_scale, int q_dtype, int kv_dtype,
    const b200_blocksparse_t* blocksparse, const b200_attn_tuning_t* tuning_or_null, int device,
    b200_stream_t stream);

/* b200_swap_blocks with the copy mechanism forced: mode 0 = auto (what b200_swap_blocks does),
 * 1 = run-merged cudaMemcpyAsync, 2 = the gather/scatter kernel (needs pinned host memory). */
int b200_swap_blocks_mode(const void* src, void* dst, const int64_t* block_mapping_host,
                          int num_pairs, int64_t block_bytes, int src_device, int dst_device,
                          int mode, b200_stream_t stream);

/* Number of kernels this library has launched from the calling process (for gpu_launches). */
int64_t b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_PAGED_H_ */
