// Main kernel: per-warp bulk-copy pipeline around process_page(), warp merge, split merge.
#pragma once

#include "paged_attention.cuh"

#ifndef B200_ATTN_MAXT
#define B200_ATTN_MAXT 256
#endif
#ifndef B200_ATTN_MINB
#define B200_ATTN_MINB 1
#endif

namespace b200 {

// Load this lane's A-operand fragments of Q for the q-head group (rows = heads, padded to GROWS).
// Storage order per k-step: [0]=a0 (row g, k lo), [1]=a2 (row g, k hi), [2]=a1 (row g+8, k lo),
// [3]=a3 (row g+8, k hi).  For fp8 caches the k slots follow the K-page byte order (see
// process_page): lane q4 owns d = 16ks + 4*q4 + {0,1 | 2,3}.
template <typename T, int D, int KV, int GROWS>
__device__ __forceinline__ void load_q(uint32_t (&qa)[(D + 15) / 16][GROWS / 4], const T* q_seq,
                                       int head0, int gcount, int lane) {
  using MT = typename MmaType<T, KV>::type;
  constexpr int KS = (D + 15) / 16;
  const int g = lane >> 2, q4 = lane & 3;
#pragma unroll
  for (int rs = 0; rs < GROWS / 8; ++rs) {
    const int row = g + 8 * rs;
    const bool on = row < gcount;
    const T* qrow = q_seq + (int64_t)(head0 + (on ? row : 0)) * D;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t lo = 0u, hi = 0u;
      if (on) {
        if constexpr (KV == B200_KV_AUTO) {
          const int d0 = 16 * ks + 2 * q4;
          if (d0 < D) lo = *reinterpret_cast<const uint32_t*>(qrow + d0);
          if (d0 + 8 < D) hi = *reinterpret_cast<const uint32_t*>(qrow + d0 + 8);
        } else {
          const int d0 = 16 * ks + 4 * q4;
          if (d0 < D) {
            const float f0 = to_f32<T>(qrow[d0]), f1 = to_f32<T>(qrow[d0 + 1]);
            const float f2 = to_f32<T>(qrow[d0 + 2]), f3 = to_f32<T>(qrow[d0 + 3]);
            lo = MmaOp<MT>::pack(f0, f1);
            hi = MmaOp<MT>::pack(f2, f3);
          }
        }
      }
      qa[ks][2 * rs] = lo;
      qa[ks][2 * rs + 1] = hi;
    }
  }
}

// Grid: (splits, KH * gchunks, num_seqs).  v1: `splits` CTAs form one cluster (DSMEM merge).
// v2: splits = max_num_partitions, each CTA owns one 512-token partition (scratch merge).
template <typename T, int D, int BS, int KV, int GROWS, bool PLAIN>
__global__ void __launch_bounds__(B200_ATTN_MAXT, B200_ATTN_MINB) paged_attention_kernel(const __grid_constant__ AttnParams p) {
  constexpr int ES = KvTraits<KV>::ES;
  constexpr int PAGE_BYTES = D * BS * ES;        // one K (or V) page of one KV head
  constexpr int STAGE_BYTES = 2 * PAGE_BYTES;    // K page + V page
  constexpr int RS = GROWS / 8;
  constexpr int NDT = D / 8;
  constexpr int ROWF = D + kMergePad;            // floats per merged row
  static_assert(PAGE_BYTES % 16 == 0, "bulk copies need 16-byte multiples");

  // A kernel queued behind this one with programmatic stream serialisation (the context-sharded exchange
  // kernel) may become resident now and park in griddepcontrol.wait; a no-op for ordinary successors.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // ... and this grid is itself launched as a programmatic dependent (paged_attention_inst.inc): wait for the
  // predecessor's completion before the first global read (a no-op after an ordinary predecessor).
  asm volatile("griddepcontrol.wait;" ::: "memory");
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  const int NS = p.num_stages;
  const bool is_v2 = p.partition_pages > 0;

  const int seq = blockIdx.z;
  const int kvh = blockIdx.y / p.gchunks;
  const int gchunk = blockIdx.y - kvh * p.gchunks;
  const int head0 = kvh * p.group + gchunk * p.rows_per_cta;           // first q-head of this CTA
  const int gcount = min(p.rows_per_cta, p.group - gchunk * p.rows_per_cta);  // real q-heads in this CTA
  const int seq_len = p.seq_lens[seq];
  const int n_pages = (seq_len + BS - 1) / BS;

  int page_begin, page_end;
  if (is_v2) {
    page_begin = blockIdx.x * p.partition_pages;
    if (page_begin >= n_pages) return;  // empty partition (attention_kernels.cu:116-119)
    page_end = min(page_begin + p.partition_pages, n_pages);
  } else {
    const int splits = gridDim.x;
    const int per = (n_pages + splits - 1) / splits;
    page_begin = min((int)blockIdx.x * per, n_pages);
    page_end = min(page_begin + per, n_pages);
  }

  // ---- shared memory carve-up ----
  const uint32_t ring_bytes = (uint32_t)nwarps * NS * STAGE_BYTES;
  const uint32_t merge_bytes = (uint32_t)(nwarps + 1) * GROWS * (ROWF + 2) * 4;
  const uint32_t data_bytes = ring_bytes > merge_bytes ? ring_bytes : merge_bytes;
  unsigned char* ring = smem + (size_t)warp * NS * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ((data_bytes + 127) & ~127u)) + warp * NS;
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t bars_u32 = smem_u32(bars);

  if (lane == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(bars_u32 + 8 * s, 1);
    mbar_fence_init();
  }
  __syncwarp();

  // ---- per-warp page schedule: local pages warp, warp + nwarps, ... ----
  const int n_local = page_end - page_begin;
  const int n_my = n_local > warp ? (n_local - warp + nwarps - 1) / nwarps : 0;
  const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
  const unsigned char* kbase =
      static_cast<const unsigned char*>(p.k_cache) + (int64_t)kvh * p.kv_head_stride * ES;
  const unsigned char* vbase =
      static_cast<const unsigned char*>(p.v_cache) + (int64_t)kvh * p.kv_head_stride * ES;
  const int64_t blk_bytes = p.kv_block_stride * ES;
  const uint64_t policy = l2_policy_evict_first();

  // block-table window: lane L holds the physical block of my page index (win0 + L)
  int win0 = 0;
  auto load_tbl = [&](int first) -> int {
    const int i = first + lane;
    return (i < n_my) ? bt[page_begin + warp + i * nwarps] : 0;
  };
  int tblA = load_tbl(0), tblB = load_tbl(32);

  auto issue = [&](int i, int s) {  // all lanes call; lane 0 issues page i into ring stage s
    const int src_lane = i & 31;
    const int a = __shfl_sync(0xffffffffu, tblA, src_lane);
    const int b = __shfl_sync(0xffffffffu, tblB, src_lane);
    const int64_t blk = (i < win0 + 32) ? a : b;
    if (elect_one()) {
      const uint32_t bar = bars_u32 + 8 * s;
      const uint32_t dst = ring_u32 + (uint32_t)s * STAGE_BYTES;
      mbar_arrive_expect_tx(bar, STAGE_BYTES);
      bulk_g2s(dst, kbase + blk * blk_bytes, PAGE_BYTES, bar, policy);
      bulk_g2s(dst + PAGE_BYTES, vbase + blk * blk_bytes, PAGE_BYTES, bar, policy);
    }
  };

  const int pre = min(NS, n_my);
  for (int i = 0; i < pre; ++i) issue(i, i);

  // ---- Q fragments, running state ----
  uint32_t qa[(D + 15) / 16][GROWS / 4];
  load_q<T, D, KV, GROWS>(qa, static_cast<const T*>(p.q) + (int64_t)seq * p.q_stride, head0,
                          gcount, lane);
  float o[NDT][GROWS / 4];
#pragma unroll
  for (int j = 0; j < NDT; ++j)
#pragma unroll
    for (int e = 0; e < GROWS / 4; ++e) o[j][e] = 0.f;
  float m[RS], l[RS], slope[RS];
  const int g = lane >> 2;
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    m[rs] = kNegInf;
    l[rs] = 0.f;
    const int row = g + 8 * rs;
    slope[rs] = (p.alibi_slopes != nullptr && row < gcount)
                    ? p.alibi_slopes[head0 + row] * 1.4426950408889634f
                    : 0.f;
  }

  // blocksparse per-row offsets (attention_kernels.cu:209-225)
  int bs_off[RS];
  int q_bs_block_id = 0;
  if (p.bs_enabled) {
    q_bs_block_id = (seq_len - 1) / p.bs_block_size;
#pragma unroll
    for (int rs = 0; rs < RS; ++rs) {
      const int head = head0 + min(g + 8 * rs, gcount - 1);
      bs_off[rs] = (p.bs_head_sliding_step >= 0)
                       ? (p.bs_tp_rank * p.num_heads + head) * p.bs_head_sliding_step + 1
                       : (p.bs_tp_rank * p.num_kv_heads + kvh) * (-p.bs_head_sliding_step) + 1;
    }
  }

  // ---- main loop ----
  int stage = 0;
  uint32_t parity = 0;
  for (int i = 0; i < n_my; ++i) {
    if (i == win0 + 32) {  // slide the block-table window
      win0 += 32;
      tblA = tblB;
      tblB = load_tbl(win0 + 32);
    }
    const int page = page_begin + warp + i * nwarps;
    const int tok0 = page * BS;
    const uint32_t ksm = ring_u32 + (uint32_t)stage * STAGE_BYTES;
    bool row_on[RS];
#pragma unroll
    for (int rs = 0; rs < RS; ++rs) row_on[rs] = true;
    if (!PLAIN && p.bs_enabled) {
      const int k_bs_block_id = tok0 / p.bs_block_size;
#pragma unroll
      for (int rs = 0; rs < RS; ++rs) {
        const bool is_remote = ((k_bs_block_id + bs_off[rs]) % p.bs_vert_stride) == 0;
        const bool is_local = k_bs_block_id > q_bs_block_id - p.bs_local_blocks;
        row_on[rs] = is_remote || is_local;
      }
    }
    mbar_wait(bars_u32 + 8 * stage, parity);
    const float tok_rel0 = (float)(tok0 - seq_len + 1);
    if (tok0 + BS > seq_len) {
      process_page<T, D, BS, KV, GROWS, true, PLAIN>(ksm, ksm + PAGE_BYTES, qa, o, m, l, lane,
                                              seq_len - tok0, p.scale_log2, slope, tok_rel0,
                                              row_on);
    } else {
      process_page<T, D, BS, KV, GROWS, false, PLAIN>(ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, BS,
                                               p.scale_log2, slope, tok_rel0, row_on);
    }
    __syncwarp();  // every lane is done reading this stage before it is refilled
    if (i + NS < n_my) issue(i + NS, stage);  // page i + NS lands in the stage just consumed
    if (++stage == NS) {
      stage = 0;
      parity ^= 1u;
    }
  }

  // ---- merge warps through shared memory (ring is idle now) ----
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    l[rs] += __shfl_xor_sync(0xffffffffu, l[rs], 1);
    l[rs] += __shfl_xor_sync(0xffffffffu, l[rs], 2);
  }
  __syncthreads();
  float* mo = reinterpret_cast<float*>(smem);                 // [nwarps][GROWS][ROWF]
  float* mm = mo + (size_t)nwarps * GROWS * ROWF;             // [nwarps][GROWS]
  float* ml = mm + (size_t)nwarps * GROWS;                    // [nwarps][GROWS]
  float* fo = ml + (size_t)nwarps * GROWS;                    // [GROWS][ROWF]  CTA result
  float* fm = fo + (size_t)GROWS * ROWF;                      // [GROWS]
  float* fl = fm + GROWS;                                     // [GROWS]
  {
    const int q4 = lane & 3;
#pragma unroll
    for (int rs = 0; rs < RS; ++rs) {
      const int row = g + 8 * rs;
      float* dst = mo + ((size_t)warp * GROWS + row) * ROWF + 2 * q4;
#pragma unroll
      for (int j = 0; j < NDT; ++j)
        *reinterpret_cast<float2*>(dst + 8 * j) = make_float2(o[j][2 * rs], o[j][2 * rs + 1]);
      if (q4 == 0) {
        mm[warp * GROWS + row] = m[rs];
        ml[warp * GROWS + row] = l[rs];
      }
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < gcount * D; idx += blockDim.x) {
    const int row = idx / D, d = idx - row * D;
    float M = kNegInf;
    for (int w = 0; w < nwarps; ++w) M = fmaxf(M, mm[w * GROWS + row]);
    float L = 0.f, O = 0.f;
    if (M != kNegInf) {
      for (int w = 0; w < nwarps; ++w) {
        const float sc = fast_exp2(mm[w * GROWS + row] - M);
        L += ml[w * GROWS + row] * sc;
        O += mo[((size_t)w * GROWS + row) * ROWF + d] * sc;
      }
    }
    fo[row * ROWF + d] = O;
    if (d == 0) {
      fm[row] = M;
      fl[row] = L;
    }
  }
  __syncthreads();

  T* out = static_cast<T*>(p.out);
  if (is_v2) {
    // partition result, normalised per partition like the reference (:339-357, :484-495)
    T* tmp = static_cast<T*>(p.tmp_out);
    for (int idx = threadIdx.x; idx < gcount * D; idx += blockDim.x) {
      const int row = idx / D, d = idx - row * D;
      const int64_t hp = ((int64_t)seq * p.num_heads + head0 + row) * p.max_num_partitions +
                         blockIdx.x;
      const float inv = __fdividef(1.f, fl[row] + 1e-6f);
      tmp[hp * D + d] = from_f32<T>(fo[row * ROWF + d] * inv * p.out_scale);
      if (d == 0) {
        p.max_logits[hp] = fm[row] * 0.6931471805599453f;  // back to natural-log units
        p.exp_sums[hp] = fl[row];
      }
    }
    return;
  }

  const uint32_t csize = cluster_nctarank();
  if (csize == 1) {
    for (int idx = threadIdx.x; idx < gcount * D; idx += blockDim.x) {
      const int row = idx / D, d = idx - row * D;
      const float inv = __fdividef(1.f, fl[row] + 1e-6f);
      out[((int64_t)seq * p.num_heads + head0 + row) * D + d] =
          from_f32<T>(fo[row * ROWF + d] * inv * p.out_scale);
    }
    return;
  }
  // cluster merge: every CTA published (fo, fm, fl) in its own shared memory; rank 0 pulls.
  cluster_sync_all();
  if (cluster_ctarank() == 0) {
    const uint32_t fo_u32 = smem_u32(fo), fm_u32 = smem_u32(fm), fl_u32 = smem_u32(fl);
    for (int idx = threadIdx.x; idx < gcount * D; idx += blockDim.x) {
      const int row = idx / D, d = idx - row * D;
      float M = kNegInf;
      for (uint32_t r = 0; r < csize; ++r)
        M = fmaxf(M, ld_dsmem_f32(dsmem_addr(fm_u32 + 4 * row, r)));
      float L = 0.f, O = 0.f;
      if (M != kNegInf) {
        for (uint32_t r = 0; r < csize; ++r) {
          const float sc = fast_exp2(ld_dsmem_f32(dsmem_addr(fm_u32 + 4 * row, r)) - M);
          L += ld_dsmem_f32(dsmem_addr(fl_u32 + 4 * row, r)) * sc;
          O += ld_dsmem_f32(dsmem_addr(fo_u32 + 4 * (row * ROWF + d), r)) * sc;
        }
      }
      const float inv = __fdividef(1.f, L + 1e-6f);
      out[((int64_t)seq * p.num_heads + head0 + row) * D + d] =
          from_f32<T>(O * inv * p.out_scale);
    }
  }
  cluster_sync_all();  // peers must stay resident until rank 0 has read their shared memory
}

// Shared memory a CTA of this instantiation needs.
template <int D, int BS, int KV, int GROWS>
inline size_t attn_smem_bytes(int nwarps, int nstages) {
  constexpr int ES = KvTraits<KV>::ES;
  const size_t ring = (size_t)nwarps * nstages * 2 * D * BS * ES;
  const size_t merge = (size_t)(nwarps + 1) * GROWS * (D + kMergePad + 2) * 4;
  const size_t data = ((ring > merge ? ring : merge) + 127) & ~(size_t)127;
  return data + (size_t)nwarps * nstages * 8 + 1024;  // + mbarriers + slack for padded ldmatrix rows
}

}  // namespace b200
