// C entry points for paged attention: argument validation, launch planning, dispatch to the
// instantiated kernels, and the v2 partition-merge kernel
// (replaces attention_kernels.cu:565-669, 690-770, 808-829, 848-934, 973-997).
#include <mutex>
#include "paged_attention.cuh"

#include <atomic>
#include <cstdlib>
#include <cstring>

namespace b200 {

AttnEntry attn_lookup_f16_auto(int, int, int);
AttnEntry attn_lookup_f16_e4m3(int, int, int);
AttnEntry attn_lookup_f16_e5m2(int, int, int);
AttnEntry attn_lookup_bf16_auto(int, int, int);
AttnEntry attn_lookup_bf16_e4m3(int, int, int);
AttnEntry attn_lookup_bf16_e5m2(int, int, int);
AttnEntry attn_lookup_f32(int head_size, int block_size, int kv_dtype, int group_rows);
int attn_f32_group_rows(int group, int head_size);

constexpr int kPartitionSize = 512;  // fixed by the op contract (paged_attn.py:14)

// context_shard.cu: fold this rank's partitions, exchange with the peers over NVLink, merge
int launch_ctx_exchange(void* out, const float* exp_sums, const float* max_logits,
                        const void* tmp_out, const int32_t* seq_lens, int num_seqs, int num_heads,
                        int head_size, int max_num_partitions, int partition_tokens, int parts,
                        int q_dtype, void* const* exchange_bufs, int rank, int world, int max_pairs,
                        uint32_t epoch, int phases, cudaStream_t stream);
struct CtxShardArgs {
  void* const* exchange_bufs;
  int rank, world, max_pairs;
  uint32_t epoch;
  int phases;  // 1 = partition pass + publish, 2 = wait + merge, 3 = both in one kernel
};

// ---------------------------------------------------------------------------------------
// v2 merge: out = sum_j tmp_out_j * exp_sum_j * e^(m_j - m) / (sum_j exp_sum_j e^(m_j - m) + 1e-6)
// (reference: paged_attention_v2_reduce_kernel, attention_kernels.cu:565-669, one 128-thread CTA
// per (seq, head)).  Here: also one CTA per (seq, head), launched as a programmatic dependent of the
// partition pass; statistics and the first rows are requested in one round trip, the four warps fold
// interleaved quarters of the partitions (coalesced 4-byte row loads, four short chains instead of one
// long one) and the CTA adds the partial rows.  The same fold runs inside the context-sharded exchange
// kernel (context_shard.cu).
// `partition_tokens` is the context length one partition covers (a multiple of 512 chosen by
// make_plan; the scratch tensors' partition stride stays max_num_partitions).
// ---------------------------------------------------------------------------------------
constexpr int kReducePrefetch = 8;  // partitions per warp requested before their weights exist

template <typename T>
__global__ void __launch_bounds__(128, 4)  // <= 128 registers: parks beside the partition pass's resident CTAs
paged_attention_v2_reduce_kernel(T* __restrict__ out, const float* __restrict__ exp_sums,
                                 const float* __restrict__ max_logits,
                                 const T* __restrict__ tmp_out, const int32_t* __restrict__ seq_lens,
                                 int num_heads, int head_size, int max_num_partitions,
                                 int partition_tokens, int parts_cap, int64_t total_pairs) {
  // One CTA per (seq, head).  Launched with programmatic stream serialisation behind the partition pass
  // (which triggers its dependents once all of its CTAs run): the grid is resident and parked here when
  // the partials land.  The first kReducePrefetch rows of every warp, the statistics and seq_lens are then
  // requested together -- one memory round trip -- and the four warps fold interleaved quarters of the
  // partitions (independent coalesced row loads, short dependency chains); the CTA adds the four partial rows.
  extern __shared__ float wsm[];  // [4][head_size] partial rows | [parts_cap] max logits | [parts_cap] exp sums
  float* part = wsm;
  float* sm_ml = wsm + 4 * head_size;
  float* sm_es = sm_ml + parts_cap;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  for (int64_t pair = blockIdx.x; pair < total_pairs; pair += gridDim.x) {
    const int seq = (int)(pair / num_heads);
    const int seq_len = seq_lens[seq];
    T* o = out + pair * head_size;
    const T* t = tmp_out + pair * max_num_partitions * head_size;
    const float* ml = max_logits + pair * max_num_partitions;
    const float* es = exp_sums + pair * max_num_partitions;
    float pre[kReducePrefetch][2][2];  // [partition warp + 4u][64-wide slice of the first 128 elements][2]
#pragma unroll
    for (int u = 0; u < kReducePrefetch; ++u) {
      const int j = warp + 4 * u;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int d0 = 2 * lane + 64 * c;
        // rows below parts_cap exist in the scratch tensor whether or not this sequence filled them: the
        // guard does not depend on seq_lens (unused rows are dropped by a select, never multiplied by 0)
        const bool ok = j < parts_cap && d0 < head_size;
        pre[u][c][0] = ok ? to_f32<T>(t[(int64_t)j * head_size + d0]) : 0.f;
        pre[u][c][1] = ok ? to_f32<T>(t[(int64_t)j * head_size + d0 + 1]) : 0.f;
      }
    }
    for (int j = threadIdx.x; j < parts_cap; j += blockDim.x) {
      sm_ml[j] = ml[j];
      sm_es[j] = es[j];
    }
    const int np = min((seq_len + partition_tokens - 1) / partition_tokens, parts_cap);
    __syncthreads();
    if (np <= 1) {
      // np == 1: plain copy (attention_kernels.cu:582-594).  np == 0 (empty sequence): the
      // partition kernel wrote nothing; define the output as zeros.
      for (int d = threadIdx.x; d < head_size; d += blockDim.x) o[d] = np == 1 ? t[d] : from_f32<T>(0.f);
      __syncthreads();
      continue;
    }
    float mx = -__builtin_huge_valf();
    for (int j = lane; j < np; j += 32) mx = fmaxf(mx, sm_ml[j]);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, s));
    float sum = 0.f;
    for (int j = lane; j < np; j += 32) sum += sm_es[j] * expf(sm_ml[j] - mx);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, s);
    const float inv = __fdividef(1.f, sum + 1e-6f);
    float acc[4][2];
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[c][0] = acc[c][1] = 0.f;
#pragma unroll
    for (int u = 0; u < kReducePrefetch; ++u) {
      const int j = warp + 4 * u;
      const float wj = j < np ? sm_es[j] * expf(sm_ml[j] - mx) : 0.f;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        acc[c][0] += j < np ? pre[u][c][0] * wj : 0.f;
        acc[c][1] += j < np ? pre[u][c][1] * wj : 0.f;
      }
    }
    // whatever the prefetch did not cover: partitions beyond 4 * kReducePrefetch, elements beyond 128
    if (np > 4 * kReducePrefetch || head_size > 128) {
      for (int j = warp; j < np; j += 4) {
        const float wj = sm_es[j] * expf(sm_ml[j] - mx);
        const bool fetched = j < 4 * kReducePrefetch;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int d0 = 2 * lane + 64 * c;
          if (d0 < head_size && !(fetched && c < 2)) {
            acc[c][0] += to_f32<T>(t[(int64_t)j * head_size + d0]) * wj;
            acc[c][1] += to_f32<T>(t[(int64_t)j * head_size + d0 + 1]) * wj;
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int d0 = 2 * lane + 64 * c;
      if (d0 < head_size) *reinterpret_cast<float2*>(part + warp * head_size + d0) = make_float2(acc[c][0], acc[c][1]);
    }
    __syncthreads();
    for (int d = threadIdx.x; d < head_size; d += blockDim.x)
      o[d] = from_f32<T>((part[d] + part[head_size + d] + part[2 * head_size + d] + part[3 * head_size + d]) * inv);
    __syncthreads();  // the shared buffers are reused by the next pair
  }
}

// ---------------------------------------------------------------------------------------
// launch planning
// ---------------------------------------------------------------------------------------
// Per-call launch overrides (b200_attn_tuning_t; 0 = library default).  They arrive as an argument of
// b200_paged_attention_tuned / b200_paged_attention_plan only: the library keeps no mutable state
// between calls and reads no environment variables (thread-per-GPU callers).
struct Knobs {
  int warps = 0, stages = 0, cluster = 0, rows = 0;
};

struct Plan {
  AttnEntry entry;
  int grows, gchunks, nwarps, nstages, cluster, ctas_per_sm, num_sms;
  int rows;                 // q-heads per CTA (<= grows; < group = head split)
  int v2_parts;             // v2: partitions per sequence at max_seq_len (1 = single pass)
  int v2_partition_tokens;  // v2: context tokens per partition (>= 512, a multiple of the block size)
  size_t smem;
};

static int make_plan(Plan& pl, int version, int num_seqs, int num_heads, int num_kv_heads,
                     int head_size, int block_size, int max_seq_len, int q_dtype, int kv_dtype,
                     int device, const Knobs& kn) {
  B200_CHECK_ARG(num_kv_heads > 0 && num_heads > 0 && num_heads % num_kv_heads == 0,
                 "num_heads (%d) must be a positive multiple of num_kv_heads (%d)", num_heads,
                 num_kv_heads);
  const int group = num_heads / num_kv_heads;
  pl.grows = (group > 8 && head_size <= 128) ? 16 : 8;
  pl.rows = group < pl.grows ? group : pl.grows;
  pl.gchunks = (group + pl.rows - 1) / pl.rows;
  AttnEntry e{nullptr, nullptr, nullptr};
  if (q_dtype == B200_DT_F16) {
    e = kv_dtype == B200_KV_AUTO       ? attn_lookup_f16_auto(head_size, block_size, pl.grows)
        : kv_dtype == B200_KV_FP8_E4M3 ? attn_lookup_f16_e4m3(head_size, block_size, pl.grows)
        : kv_dtype == B200_KV_FP8_E5M2 ? attn_lookup_f16_e5m2(head_size, block_size, pl.grows)
                                       : e;
  } else if (q_dtype == B200_DT_BF16) {
    e = kv_dtype == B200_KV_AUTO       ? attn_lookup_bf16_auto(head_size, block_size, pl.grows)
        : kv_dtype == B200_KV_FP8_E4M3 ? attn_lookup_bf16_e4m3(head_size, block_size, pl.grows)
        : kv_dtype == B200_KV_FP8_E5M2 ? attn_lookup_bf16_e5m2(head_size, block_size, pl.grows)
                                       : e;
  } else if (q_dtype == B200_DT_F32) {
    // fp32 queries run on the CUDA-core kernel (paged_attention_f32.cu): 1, 2 or 4 q-heads per CTA
    pl.grows = attn_f32_group_rows(group, head_size);
    pl.rows = group < pl.grows ? group : pl.grows;
    pl.gchunks = (group + pl.rows - 1) / pl.rows;
    e = attn_lookup_f32(head_size, block_size, kv_dtype, pl.grows);
  } else {
    set_error("Unsupported data type: query dtype %d", q_dtype);
    return B200_ERR_UNSUPPORTED;
  }
  if (kv_dtype < 0 || kv_dtype > B200_KV_FP8_E5M2) {
    set_error("Unsupported data type of kv cache: %d", kv_dtype);
    return B200_ERR_UNSUPPORTED;
  }
  if (!e.launch) {
    if (block_size != 8 && block_size != 16 && block_size != 32)
      set_error("Unsupported block size: %d", block_size);
    else
      set_error("Unsupported head size: %d", head_size);
    return B200_ERR_UNSUPPORTED;
  }
  pl.entry = e;

  int num_sms = 148;
  cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device);
  pl.num_sms = num_sms;

  // Pipeline geometry.  Registers allow three 4-warp CTAs per SM (<= 168 regs x 128 threads), and
  // three CTAs with a short ring beat two with a deep one on every measured config
  // (profiles/r1_sweep_grid.md): size the per-warp ring so that three CTAs fit in 227 KiB of
  // shared memory, between 2 and 4 stages (one stage = one K page + one V page).
  const int es = kv_dtype == B200_KV_AUTO ? (q_dtype == B200_DT_F32 ? 4 : 2) : 1;
  const size_t stage_bytes = (size_t)2 * head_size * block_size * es;
  pl.nwarps = kn.warps;
  if (pl.nwarps < 1) pl.nwarps = 4;
  if (pl.nwarps > 8) pl.nwarps = 8;
  if (q_dtype == B200_DT_F32) {  // 128-thread kernel; fp32 pages are up to 32 KiB each
    if (pl.nwarps > 4) pl.nwarps = 4;
    while (pl.nwarps > 1 && (size_t)pl.nwarps * 2 * stage_bytes > (size_t)200 * 1024) pl.nwarps /= 2;
  }
  if (kn.rows > 0) {  // forced head split
    pl.rows = kn.rows < pl.rows ? kn.rows : pl.rows;
    pl.gchunks = (group + pl.rows - 1) / pl.rows;
  }
  long long base = (long long)num_seqs * num_kv_heads * pl.gchunks;  // CTAs before any split
  if (base < 1) base = 1;
  // One-wave regime (profiles/r2_small_plan_sweep.md): when the unsplit grid has at most one CTA per SM, the
  // batch-oriented plan below (4-warp CTAs, three per SM, split -- and for B <= 2 head-split -- until the slots
  // are full) pays 2-3 prologues and a cluster / scratch merge per SM inside a kernel that lasts one wave, and
  // the head split multiplies the L2 traffic.  Measured better on every such shape: ONE 8-warp CTA per SM (twice
  // the pages in flight per CTA), no head split, context split only while the grid stays within the SM count;
  // 2 stages when most SMs have a CTA, 3 when fewer do.  v1: cluster 8 at one CTA per SM co-schedules fewer
  // than 16 clusters (two per GPC at best), so it is used up to half the SMs only.  v2 runs unsplit from a
  // third of the SMs up, which also drops its reduce launch.
  // v1 1 x 8192: 15.7 -> 12.6 us, 1 x 32768: 45.2 -> 27.8 us, 16 x 2048: 25.6 -> 23.2 us, 16 x 4096: 45.0 -> 42.2 us
  // (0.97 of the HBM peak); v2 16 x 2048: 27.3 -> 23.1 us.
  int wide_split = 0;
  if (kn.warps <= 0 && kn.stages <= 0 && kn.cluster <= 0 && kn.rows <= 0 && q_dtype != B200_DT_F32 &&
      base <= num_sms && e.smem(8, 2) <= (size_t)227 * 1024) {  // (head 256 x block 32: an 8-warp ring does not fit)
    const int pages = (max_seq_len + block_size - 1) / block_size;
    int sp = 1;
    if (version == 1) {
      auto cap = [&](int c) { return c >= 8 ? num_sms / 2 : (c >= 4 ? num_sms * 9 / 10 : num_sms); };
      while (sp < 8 && base * sp * 2 <= cap(sp * 2) && pages / (2 * sp) >= 2 * 8) sp *= 2;
    } else if (base * 3 < num_sms) {
      sp = (int)(num_sms / base);  // partitions: as many as keep the grid within one CTA per SM
      const int most = pages / (kPartitionSize / block_size > 0 ? kPartitionSize / block_size : 1);
      if (sp > most) sp = most;
      if (sp < 1) sp = 1;
    }
    wide_split = sp;
    pl.nwarps = 8;
  }
  int ns = kn.stages;
  if (wide_split > 0) {
    ns = base * wide_split * 4 >= 3LL * num_sms ? 2 : 3;
  } else if (ns <= 0) {
    // candidate A: ring that lets three CTAs share an SM (2..4 stages); candidate B: a deeper ring
    // with two CTAs per SM.  A wins whenever the grid runs for many waves or fits in one; B wins
    // when a 1-3 wave grid would leave A with a nearly empty last wave (c2: 512 CTAs on 444 slots).
    auto fit = [&](size_t budget, int cap) {
      int n = (int)(budget / ((size_t)pl.nwarps * stage_bytes));
      return n > cap ? cap : (n < 2 ? 2 : n);
    };
    const int ns_a = fit((size_t)74 * 1024, 4), ns_b = fit((size_t)104 * 1024, 6);
    ns = ns_a;
    if (ns_b > ns_a) {
      const long long slots_a = (long long)num_sms * 3, slots_b = (long long)num_sms * 2;
      const long long waves_a = (base + slots_a - 1) / slots_a, waves_b = (base + slots_b - 1) / slots_b;
      if (waves_a >= 2 && waves_a <= 3) {
        const double eff_a = (double)base / (double)(waves_a * slots_a);
        const double eff_b = (double)base / (double)(waves_b * slots_b);
        if (eff_b > eff_a + 0.1) ns = ns_b;
      }
    }
  }
  if (ns > 32) ns = 32;
  while (ns > 1 && e.smem(pl.nwarps, ns) > (size_t)227 * 1024) --ns;
  pl.nstages = ns;
  pl.smem = e.smem(pl.nwarps, ns);
  B200_CHECK_ARG(pl.smem <= (size_t)227 * 1024, "shared memory plan does not fit (%zu bytes)",
                 pl.smem);
  pl.ctas_per_sm = e.occupancy(pl.nwarps, ns);
  if (pl.ctas_per_sm < 1) pl.ctas_per_sm = 1;

  // ---- context split and head split ----
  // v1 splits through a thread-block cluster (<= 8 CTAs, DSMEM merge), v2 through the caller's scratch.
  // Measured on B200 (profiles/r1_sweep_c2.md): an unsplit grid of >= one CTA per residency slot already
  // reaches ~97 % of the HBM peak, and clusters cost co-scheduling, so v1 splits only when the grid cannot
  // fill the machine.  The op contract fixes the v2 scratch SHAPE ([B,H,ceil(max_seq_len/512)(,D)]) but its
  // contents are scratch (never read by the caller, paged_attn.py:156-191): a partition may be any whole
  // number of pages >= 512 tokens, so the partition count is chosen for the MACHINE -- one resident wave of
  // ~1.5 CTAs per SM, pages dealt evenly -- not fixed at 512 tokens.  When even the
  // finest split leaves most slots empty (B <= 4 at 8 kv heads), the q-heads of a GQA group are spread over
  // several CTAs (head split): they read the same pages at the same time, so HBM still sees each page once
  // and L2 serves the copies.
  const long long slots = (long long)num_sms * pl.ctas_per_sm;
  const int max_pages = (max_seq_len + block_size - 1) / block_size;
  const int min_part_pages = kPartitionSize / block_size > 0 ? kPartitionSize / block_size : 1;
  const int parts_max = max_pages > 0 ? (max_pages + min_part_pages - 1) / min_part_pages : 1;
  pl.cluster = 1;
  pl.v2_parts = 1;
  pl.v2_partition_tokens = kPartitionSize;
  auto v1_split = [&](long long b) {
    int s = 1;
    while (s < 8 && 2 * b * s < slots && max_pages / (2 * s) >= 2 * pl.nwarps) s *= 2;
    return s;
  };
  const long long finest = version == 1 ? (q_dtype == B200_DT_F32 ? 1 : 8) : parts_max;
  // (measured, profiles/r2_sweep_small.md: the head split pays for v1 -- one launch, B = 1 x 8 Ki tokens 23.9 ->
  // 18.6 us -- and changes nothing for v2, whose two launches set the floor at these sizes)
  if (kn.rows <= 0 && q_dtype != B200_DT_F32 && version == 1 && wide_split == 0) {
    while (pl.rows > 1 && 2 * base * finest <= slots) {  // less than half the slots even at the finest split
      pl.rows = (pl.rows + 1) / 2;
      pl.gchunks = (group + pl.rows - 1) / pl.rows;
      base = (long long)num_seqs * num_kv_heads * pl.gchunks;
      if (base < 1) base = 1;
    }
  }
  if (version == 1) {
    const int forced = kn.cluster;
    if (q_dtype == B200_DT_F32) {
      pl.cluster = 1;  // the fp32 kernel has no cluster merge: v1 runs unsplit
    } else if (forced > 0) {
      pl.cluster = forced;
    } else if (wide_split > 0) {
      pl.cluster = wide_split;
    } else {
      pl.cluster = v1_split(base);
    }
  } else {
    int parts;
    if (kn.cluster > 0) {  // tuning: "split" = forced partition count for v2
      parts = kn.cluster < parts_max ? kn.cluster : parts_max;
    } else if (wide_split > 0) {
      parts = wide_split < parts_max ? wide_split : parts_max;
    } else {
      // Measured (profiles/r2_sweep_small.md, tools/sweep_ctx_plan.sh): for grids that need a split at all
      // the best partition count puts ~1.5 CTAs on every SM in ONE resident wave (4 x 16 Ki tokens: 7
      // partitions 51.7 us, 27 partitions 56.5 us; 1 x 64 Ki: 32 partitions 53.1 us, 111 partitions 67.7 us);
      // more, shorter partitions only add prologues, partial rows and reduce work, and counts that land
      // just above a multiple of the residency (14-16 partitions x 32 CTAs on 444 slots) lose a wave.
      long long want = (3LL * num_sms / 2 + base / 2) / base;
      if (want < 1) want = 1;
      if (want > parts_max) want = parts_max;
      parts = (int)want;
      (void)slots;
    }
    if (parts < 1) parts = 1;
    int pp = max_pages > 0 ? (max_pages + parts - 1) / parts : min_part_pages;
    if (pp < min_part_pages) pp = min_part_pages;
    pl.v2_partition_tokens = pp * block_size;
    pl.v2_parts = max_pages > 0 ? (max_pages + pp - 1) / pp : 1;
  }
  return B200_OK;
}

static int run_attention(int version, void* out, float* exp_sums, float* max_logits, void* tmp_out,
                         const void* query, const void* key_cache, const void* value_cache,
                         const int32_t* block_tables, const int32_t* seq_lens,
                         const float* alibi_slopes, int num_seqs, int num_heads, int num_kv_heads,
                         int head_size, int block_size, int max_seq_len,
                         int max_num_blocks_per_seq, int max_num_partitions, int64_t q_stride,
                         int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                         float k_scale, float v_scale, int q_dtype, int kv_dtype,
                         const b200_blocksparse_t* bsp, int device, cudaStream_t stream,
                         const CtxShardArgs* ctx = nullptr, const Knobs& kn = Knobs()) {
  B200_CHECK_ARG(num_seqs >= 0 && head_size > 0 && block_size > 0 && max_seq_len >= 0,
                 "paged_attention: bad sizes");
  if (num_seqs == 0) return B200_OK;
  if (ctx) version = 2;  // context shard = v2 partition pass, partials always kept
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  int dev = device;
  if (dev < 0) B200_CUDA_OK(cudaGetDevice(&dev));

  Plan pl;
  int rc = make_plan(pl, version, num_seqs, num_heads, num_kv_heads, head_size, block_size,
                     max_seq_len, q_dtype, kv_dtype, dev, kn);
  if (rc != B200_OK) return rc;

  const int es = kv_dtype == B200_KV_AUTO ? (q_dtype == B200_DT_F32 ? 4 : 2) : 1;
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(key_cache) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(value_cache) & 15) == 0 &&
                     (kv_block_stride * es) % 16 == 0 && (kv_head_stride * es) % 16 == 0,
                 "paged_attention: KV cache base/strides must be 16-byte aligned");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(query) & 3) == 0 && q_stride % 2 == 0,
                 "paged_attention: query rows must be 4-byte aligned");
  B200_CHECK_ARG(kv_head_stride >= (int64_t)head_size * block_size,
                 "paged_attention: kv_head_stride smaller than one page");
  if (version == 2) {
    const int need = (max_seq_len + kPartitionSize - 1) / kPartitionSize;
    B200_CHECK_ARG(max_num_partitions >= need && max_num_partitions >= 1,
                   "paged_attention_v2: scratch has %d partitions, need %d", max_num_partitions,
                   need);
    B200_CHECK_ARG(kPartitionSize % block_size == 0, "partition size not a multiple of block size");
  }

  AttnParams p{};
  p.out = out;
  p.exp_sums = exp_sums;
  p.max_logits = max_logits;
  p.tmp_out = tmp_out;
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_heads / num_kv_heads;
  p.gchunks = pl.gchunks;
  p.rows_per_cta = pl.rows;
  p.max_blocks_per_seq = max_num_blocks_per_seq;
  p.max_num_partitions = max_num_partitions;
  const bool split_v2 = (version == 2 && pl.v2_parts > 1) || ctx != nullptr;
  p.partition_pages = split_v2 ? pl.v2_partition_tokens / block_size : 0;
  p.num_stages = pl.nstages;
  p.q_stride = q_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  const float log2e = 1.4426950408889634f;
  p.scale_log2 = scale * log2e * (kv_dtype == B200_KV_AUTO ? 1.f : k_scale);
  p.out_scale = kv_dtype == B200_KV_AUTO ? 1.f : v_scale;
  if (bsp && bsp->vert_stride > 1) {  // is_block_sparse (attention_kernels.cu:825)
    B200_CHECK_ARG(bsp->block_size > 0, "blocksparse_block_size must be positive");
    p.bs_enabled = 1;
    p.bs_tp_rank = bsp->tp_rank;
    p.bs_local_blocks = bsp->local_blocks;
    p.bs_vert_stride = bsp->vert_stride;
    p.bs_block_size = bsp->block_size;
    p.bs_head_sliding_step = bsp->head_sliding_step;
  }

  const int splits = version == 2 ? pl.v2_parts : pl.cluster;
  B200_CHECK_ARG((long long)num_kv_heads * pl.gchunks <= 65535 && num_seqs <= 65535,
                 "paged_attention: grid too large");
  dim3 grid(splits, num_kv_heads * pl.gchunks, num_seqs);
  if (!(ctx && ctx->phases == 2)) {  // merge-only call of the split-phase form: partials exist already
    rc = pl.entry.launch(p, grid, version == 2 ? 1 : pl.cluster, pl.nwarps, pl.nstages, dev, stream);
    if (rc != B200_OK) return rc;
  }

  if (ctx)
    return launch_ctx_exchange(out, exp_sums, max_logits, tmp_out, seq_lens, num_seqs, num_heads,
                               head_size, max_num_partitions, pl.v2_partition_tokens, pl.v2_parts,
                               q_dtype, ctx->exchange_bufs, ctx->rank, ctx->world, ctx->max_pairs,
                               ctx->epoch, ctx->phases, stream);
  if (split_v2) {
    const int64_t pairs = (int64_t)num_seqs * num_heads;
    const int64_t cap = (int64_t)pl.num_sms * 4;  // grid-stride beyond four CTAs per SM
    const unsigned blocks = (unsigned)(pairs < cap ? pairs : cap);
    const int parts_cap = pl.v2_parts > 1 ? pl.v2_parts : 1;
    const size_t wsmem = ((size_t)2 * parts_cap + (size_t)4 * head_size) * sizeof(float);
    B200_CHECK_ARG(head_size <= 256, "paged_attention_v2: the reduce kernel folds rows of up to 256 elements");
    B200_CHECK_ARG(wsmem <= 48 * 1024, "paged_attention_v2: %d partitions per sequence exceed the reduce kernel's shared memory",
                   parts_cap);
    // Programmatic dependent launch: resident and parked in griddepcontrol.wait while the partition pass drains.
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = wsmem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
#define B200_REDUCE_LAUNCH(T)                                                                               \
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, paged_attention_v2_reduce_kernel<T>, static_cast<T*>(out),          \
                                  (const float*)exp_sums, (const float*)max_logits,                        \
                                  static_cast<const T*>(tmp_out), seq_lens, num_heads, head_size,          \
                                  max_num_partitions, pl.v2_partition_tokens, parts_cap, pairs))
    if (q_dtype == B200_DT_F16) B200_REDUCE_LAUNCH(__half);
    else if (q_dtype == B200_DT_F32) B200_REDUCE_LAUNCH(float);
    else B200_REDUCE_LAUNCH(__nv_bfloat16);
#undef B200_REDUCE_LAUNCH
    count_launch();
    B200_CUDA_OK(cudaGetLastError());
  }
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_paged_attention_v1(void* out, const void* query, const void* key_cache,
                                       const void* value_cache, const int32_t* block_tables,
                                       const int32_t* seq_lens, const float* alibi_slopes,
                                       int num_seqs, int num_heads, int num_kv_heads, int head_size,
                                       int block_size, int max_seq_len, int max_num_blocks_per_seq,
                                       int64_t q_stride, int64_t kv_block_stride,
                                       int64_t kv_head_stride, float scale, float k_scale,
                                       float v_scale, int q_dtype, int kv_dtype,
                                       const b200_blocksparse_t* blocksparse, int device,
                                       b200_stream_t stream) {
  return run_attention(1, out, nullptr, nullptr, nullptr, query, key_cache, value_cache,
                       block_tables, seq_lens, alibi_slopes, num_seqs, num_heads, num_kv_heads,
                       head_size, block_size, max_seq_len, max_num_blocks_per_seq, 0, q_stride,
                       kv_block_stride, kv_head_stride, scale, k_scale, v_scale, q_dtype, kv_dtype,
                       blocksparse, device, (cudaStream_t)stream);
}

extern "C" int b200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                       const void* query, const void* key_cache,
                                       const void* value_cache, const int32_t* block_tables,
                                       const int32_t* seq_lens, const float* alibi_slopes,
                                       int num_seqs, int num_heads, int num_kv_heads, int head_size,
                                       int block_size, int max_seq_len, int max_num_blocks_per_seq,
                                       int max_num_partitions, int64_t q_stride,
                                       int64_t kv_block_stride, int64_t kv_head_stride, float scale,
                                       float k_scale, float v_scale, int q_dtype, int kv_dtype,
                                       const b200_blocksparse_t* blocksparse, int device,
                                       b200_stream_t stream) {
  B200_CHECK_ARG(exp_sums && max_logits && tmp_out, "paged_attention_v2: scratch tensors required");
  return run_attention(2, out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       block_tables, seq_lens, alibi_slopes, num_seqs, num_heads, num_kv_heads,
                       head_size, block_size, max_seq_len, max_num_blocks_per_seq,
                       max_num_partitions, q_stride, kv_block_stride, kv_head_stride, scale,
                       k_scale, v_scale, q_dtype, kv_dtype, blocksparse, device,
                       (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------
// paged_attention_v2 without caller scratch and without any allocation (no counterpart in the reference;
// added for callers that may not allocate: code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).
// The partition scratch (exp_sums | max_logits | tmp_out) lives in ONE file-scope device array of
// B200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Every call on a device
// shares it and it keeps the last call's contents: nothing in it has to start at zero (the partition pass
// writes every (pair, partition) cell the reduce kernel weights; cells it only prefetches are dropped by a
// select), so nothing is cleared -- but two calls that overlap on the GPU (different streams) race on it.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[B200_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_scratch_mu;
void* g_static_scratch_ptr[kMaxDevices] = {};  // device address of g_static_scratch per device (set-up cache)

// In host code `g_static_scratch` is a host-side stand-in, not a device address: resolve it once per device,
// with that device current.
int static_scratch_address(int dev, unsigned char** ptr) {
  B200_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "paged_attention_v2_static: device %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    B200_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
    B200_CHECK_ARG(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0,
                   "paged_attention_v2_static: bad scratch address");
    g_static_scratch_ptr[dev] = p;
  }
  *ptr = static_cast<unsigned char*>(g_static_scratch_ptr[dev]);
  return B200_OK;
}
}  // namespace

extern "C" int b200_paged_attention_v2_static(
    void* out, const void* query, const void* key_cache, const void* value_cache,
    const int32_t* block_tables, const int32_t* seq_lens, const float* alibi_slopes, int num_seqs,
    int num_heads, int num_kv_heads, int head_size, int block_size, int max_seq_len,
    int max_num_blocks_per_seq, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    float scale, float k_scale, float v_scale, int q_dtype, int kv_dtype,
    const b200_blocksparse_t* blocksparse, int device, b200_stream_t stream) {
  B200_CHECK_ARG(num_seqs >= 0 && num_heads > 0 && head_size > 0 && max_seq_len >= 0,
                 "paged_attention_v2_static: bad sizes");
  if (num_seqs == 0) return B200_OK;
  const int64_t parts = max_seq_len > 0 ? (max_seq_len + kPartitionSize - 1) / kPartitionSize : 1;
  const int64_t cells = (int64_t)num_seqs * num_heads * parts;
  const int64_t elem = q_dtype == B200_DT_F32 ? 4 : 2;
  const int64_t stat_bytes = (cells * 4 + 255) & ~(int64_t)255;  // exp_sums, max_logits: 256-byte aligned sections
  const int64_t need = 2 * stat_bytes + cells * head_size * elem;
  B200_CHECK_ARG(need <= (int64_t)B200_STATIC_SCRATCH_BYTES,
                 "paged_attention_v2_static: %d seqs x %d heads x %d partitions x head size %d needs %lld scratch "
                 "bytes, the built-in array holds %lld (use b200_paged_attention_v2 with caller scratch)",
                 num_seqs, num_heads, (int)parts, head_size, (long long)need,
                 (long long)B200_STATIC_SCRATCH_BYTES);
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  int dev = device;
  if (dev < 0) B200_CUDA_OK(cudaGetDevice(&dev));
  unsigned char* base = nullptr;
  int rc = static_scratch_address(dev, &base);
  if (rc != B200_OK) return rc;
  float* exp_sums = reinterpret_cast<float*>(base);
  float* max_logits = reinterpret_cast<float*>(base + stat_bytes);
  void* tmp_out = base + 2 * stat_bytes;
  return run_attention(2, out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       block_tables, seq_lens, alibi_slopes, num_seqs, num_heads, num_kv_heads,
                       head_size, block_size, max_seq_len, max_num_blocks_per_seq, (int)parts, q_stride,
                       kv_block_stride, kv_head_stride, scale, k_scale, v_scale, q_dtype, kv_dtype,
                       blocksparse, dev, (cudaStream_t)stream);
}

extern "C" int b200_paged_attention_ctx_shard(
    void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
    const void* key_cache, const void* value_cache, const int32_t* block_tables,
    const int32_t* local_seq_lens, int num_seqs, int num_heads, int num_kv_heads, int head_size,
    int block_size, int max_local_seq_len, int max_num_blocks_per_seq, int max_num_partitions,
    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride, float scale, float k_scale,
    float v_scale, int q_dtype, int kv_dtype, void* const* exchange_bufs, int rank, int world,
    int max_pairs, uint32_t epoch, int phases, int device, b200_stream_t stream) {
  B200_CHECK_ARG(exp_sums && max_logits && tmp_out, "paged_attention_ctx_shard: scratch tensors required");
  CtxShardArgs ctx{exchange_bufs, rank, world, max_pairs, epoch, phases};
  return run_attention(2, out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       block_tables, local_seq_lens, nullptr, num_seqs, num_heads, num_kv_heads,
                       head_size, block_size, max_local_seq_len, max_num_blocks_per_seq,
                       max_num_partitions, q_stride, kv_block_stride, kv_head_stride, scale,
                       k_scale, v_scale, q_dtype, kv_dtype, nullptr, device, (cudaStream_t)stream,
                       &ctx);
}

static Knobs to_knobs(const b200_attn_tuning_t* t) {
  Knobs k;
  if (t) {
    k.warps = t->num_warps;
    k.stages = t->num_stages;
    k.cluster = t->split;
    k.rows = t->head_rows;
  }
  return k;
}

extern "C" int b200_paged_attention_tuned(
    int version, void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
    const void* key_cache, const void* value_cache, const int32_t* block_tables,
    const int32_t* seq_lens, const float* alibi_slopes, int num_seqs, int num_heads,
    int num_kv_heads, int head_size, int block_size, int max_seq_len, int max_num_blocks_per_seq,
    int max_num_partitions, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    float scale, float k_scale, float v_scale, int q_dtype, int kv_dtype,
    const b200_blocksparse_t* blocksparse, const b200_attn_tuning_t* tuning, int device,
    b200_stream_t stream) {
  B200_CHECK_ARG(version == 1 || version == 2, "paged_attention_tuned: version must be 1 or 2");
  if (version == 2)
    B200_CHECK_ARG(exp_sums && max_logits && tmp_out, "paged_attention_v2: scratch tensors required");
  return run_attention(version, out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       block_tables, seq_lens, alibi_slopes, num_seqs, num_heads, num_kv_heads,
                       head_size, block_size, max_seq_len, max_num_blocks_per_seq,
                       version == 2 ? max_num_partitions : 0, q_stride, kv_block_stride,
                       kv_head_stride, scale, k_scale, v_scale, q_dtype, kv_dtype, blocksparse,
                       device, (cudaStream_t)stream, nullptr, to_knobs(tuning));
}

extern "C" int b200_paged_attention_plan(int version, int num_seqs, int num_heads, int num_kv_heads,
                                         int head_size, int block_size, int max_seq_len,
                                         int q_dtype, int kv_dtype, int device,
                                         const b200_attn_tuning_t* tuning,
                                         b200_attn_plan_t* plan) {
  B200_CHECK_ARG(plan != nullptr, "plan must not be null");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  int dev = device;
  if (dev < 0) B200_CUDA_OK(cudaGetDevice(&dev));
  Plan pl;
  int rc = make_plan(pl, version, num_seqs, num_heads, num_kv_heads, head_size, block_size,
                     max_seq_len, q_dtype, kv_dtype, dev, to_knobs(tuning));
  if (rc != B200_OK) return rc;
  plan->cluster_size = pl.cluster;
  plan->num_warps = pl.nwarps;
  plan->num_stages = pl.nstages;
  plan->smem_bytes = (int32_t)pl.smem;
  plan->ctas_per_sm = pl.ctas_per_sm;
  plan->grid_x = version == 2 ? pl.v2_parts : pl.cluster;
  plan->grid_y = num_kv_heads * pl.gchunks;
  plan->grid_z = num_seqs;
  return B200_OK;
}

