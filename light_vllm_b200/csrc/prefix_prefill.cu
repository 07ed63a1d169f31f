// Prefix / chunked-prefill attention over the paged cache: replaces context_attention_fwd
// (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811, Triton kernels :13-243 and
// :437-672) behind PagedAttention.forward_prefix (ops/paged_attn.py:193-225).
//
// Query token i of sequence b (position ctx_len + i) attends (1) the ctx_len tokens already in the
// paged cache, unmasked, and (2) the new tokens j <= i, read from the packed k/v tensors.
//
// One 4-warp CTA owns (sequence, KV head, tile of query tokens).  The M dimension of the MMAs is
// (query token, q-head of the GQA group): row r = token r / G, head r % G, 16 (or 8) rows per warp,
// so all q-heads of a group share ONE pass over K/V (the reference launches one program per
// q-head).  Phase 1 streams the context pages through a CTA-wide ring of bulk copies (TMA engine,
// full/empty mbarriers, no CTA barrier); every warp runs process_page_ex() on the page for its
// rows.  Phase 2 stages BS-token slices of the new k/v into the same page layouts in shared memory
// (registers prefetched one slice ahead) and applies the causal frontier per row.  Online softmax
// in the exp2 domain, fp32 statistics and accumulation, mma.sync m16n8k16 (legacy tensor path).
#include "paged_attention.cuh"
#include "prefix_prefill.cuh"
#include "prefix_prefill_umma.cuh"

#include <cuda.h>
#include <cudaTypedefs.h>

#include <atomic>
#include <cstring>

namespace b200 {

constexpr int kPrefixWarps = 4;

template <typename T, int D, int BS, int GROWS, bool PLAIN>
__global__ void __launch_bounds__(kPrefixWarps * 32, (D <= 128 ? 3 : 1))
context_attention_kernel(const __grid_constant__ PrefixParams p) {
  constexpr int KV = B200_KV_AUTO;
  constexpr int PAGE_BYTES = D * BS * 2;
  constexpr int STAGE_BYTES = 2 * PAGE_BYTES;
  constexpr int RS = GROWS / 8;
  constexpr int NDT = D / 8;
  constexpr int KS = (D + 15) / 16;
  constexpr int NTHREADS = kPrefixWarps * 32;
  constexpr int VEC_PER_PAGE = BS * NDT;                           // 16-byte vectors in a K (or V) slice
  constexpr int VPT = (VEC_PER_PAGE + NTHREADS - 1) / NTHREADS;    // per thread

  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int NS = p.num_stages;

  const int seq = blockIdx.z;
  const int kvh = blockIdx.y;
  const int ctx = p.ctx_lens[seq];
  const int qlen = p.seq_lens[seq] - ctx;
  const int m0 = blockIdx.x * p.tok_tile;  // first query token of this CTA
  if (m0 >= qlen) return;
  const int64_t start = p.start_loc[seq];
  const int G = p.group;
  const int tile_tokens = min(p.tok_tile, qlen - m0);
  const int W = p.sliding_window;
  // Staged mode indexes the scratch by start_loc (device data the host cannot verify): a start_loc that
  // is not the packed prefix sum would point past the scratch.  Never read out of bounds -- poison this
  // CTA's output rows with NaN instead (loud in the data) and leave.
  if (p.kv_scratch != nullptr &&
      (start < 0 || (int64_t)(start / BS) + seq + (m0 + tile_tokens - 1) / BS >= p.scratch_slices)) {
    for (int r = threadIdx.x; r < tile_tokens * G; r += blockDim.x) {
      T* orow = static_cast<T*>(p.out) + (start + m0 + r / G) * p.o_s0 + (int64_t)(kvh * G + r % G) * p.o_s1;
      for (int d = 0; d < D; ++d) orow[d] = from_f32<T>(__int_as_float(0x7fc00000));
    }
    return;
  }

  unsigned char* ring = smem;  // [NS stages][K page | V page], then the mbarriers, then slack
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)NS * STAGE_BYTES);
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t full_u32 = smem_u32(bars);
  const uint32_t empty_u32 = full_u32 + 8 * NS;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(full_u32 + 8 * s, 1);
      mbar_init(empty_u32 + 8 * s, kPrefixWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  // ---- rows of this lane: (token, head) pairs ----
  const int g = lane >> 2, q4 = lane & 3;
  int row_tok[RS];    // query index within the sequence (m0 + ...), or -1 for a padding row
  int row_head[RS];
  float slope[RS];
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    const int r = warp * GROWS + g + 8 * rs;
    const int ti = r / G;
    const bool on = ti < tile_tokens;
    row_tok[rs] = on ? m0 + ti : -1;
    row_head[rs] = kvh * G + (r - ti * G);
    slope[rs] = (!PLAIN && on && p.alibi_slopes) ? p.alibi_slopes[row_head[rs]] * 1.4426950408889634f : 0.f;
  }
  // warp-uniform range of query indices this warp holds (padding rows excluded)
  const int w_r0 = warp * GROWS, w_r1 = w_r0 + GROWS - 1;
  const int w_tok_lo = m0 + w_r0 / G;
  const int w_tok_hi = m0 + min(w_r1 / G, tile_tokens - 1);
  const bool warp_on = (w_r0 / G) < tile_tokens;

  // ---- Q fragments ----
  uint32_t qa[KS][GROWS / 4];
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    const bool on = row_tok[rs] >= 0;
    const T* qrow = static_cast<const T*>(p.q) + (start + (on ? row_tok[rs] : 0)) * p.q_s0 +
                    (int64_t)row_head[rs] * p.q_s1;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t lo = 0u, hi = 0u;
      const int d0 = 16 * ks + 2 * q4;
      if (on) {
        if (d0 < D) lo = *reinterpret_cast<const uint32_t*>(qrow + d0);
        if (d0 + 8 < D) hi = *reinterpret_cast<const uint32_t*>(qrow + d0 + 8);
      }
      qa[ks][2 * rs] = lo;
      qa[ks][2 * rs + 1] = hi;
    }
  }
  float o[NDT][GROWS / 4];
#pragma unroll
  for (int j = 0; j < NDT; ++j)
#pragma unroll
    for (int e = 0; e < GROWS / 4; ++e) o[j][e] = 0.f;
  float m[RS], l[RS];
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    m[rs] = kNegInf;
    l[rs] = 0.f;
  }
  bool row_on[RS];
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) row_on[rs] = true;

  // =========================== phase 1: cached context ===========================
  {
    const int n_pages = (ctx + BS - 1) / BS;
    int page0 = 0;
    if (W > 0) page0 = max(0, ctx + m0 - W + 1) / BS;  // nothing older is visible to any row
    const int n = max(0, n_pages - page0);
    const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq + page0;
    const unsigned char* kbase =
        static_cast<const unsigned char*>(p.k_cache) + (int64_t)kvh * p.kc_hs * 2;
    const unsigned char* vbase =
        static_cast<const unsigned char*>(p.v_cache) + (int64_t)kvh * p.vc_hs * 2;
    // every query tile of this (sequence, kv head) re-reads the same pages: keep them in L2
    const uint64_t policy = l2_policy_evict_last();
    // The producer role rotates over the warps (page j is issued by warp j % 4), so no warp carries
    // the whole issue overhead; every warp keeps its own 32-entry block-table window.
    int win0 = 0;
    int tbl = (lane < n) ? bt[lane] : 0;
    // Staged mode: the new tokens' k/v were rewritten into page layout (one K page + V page per
    // (slice, kv head), contiguous), so they follow the context pages through the SAME ring:
    // pages n .. n_all-1 are slices slice0 .. of the new tokens, one bulk copy each.
    const bool staged = p.kv_scratch != nullptr;
    const int last_tok = m0 + tile_tokens - 1;  // newest query of this CTA: keys j <= last_tok
    const int slice0 = W > 0 ? max(0, m0 - W + 1) / BS : 0;
    const int n_all = n + (staged ? last_tok / BS + 1 - slice0 : 0);
    const unsigned char* sbase =
        staged ? p.kv_scratch + (((int64_t)(start / BS) + seq + slice0) * p.num_kv_heads + kvh) * STAGE_BYTES
               : nullptr;
    auto issue = [&](int i, int s) {  // one warp, all lanes: page i -> stage s
      const uint32_t bar = full_u32 + 8 * s;
      const uint32_t dst = ring_u32 + (uint32_t)s * STAGE_BYTES;
      if (i >= n) {
        if (elect_one()) {
          mbar_arrive_expect_tx(bar, STAGE_BYTES);
          bulk_g2s(dst, sbase + (int64_t)(i - n) * p.num_kv_heads * STAGE_BYTES, STAGE_BYTES, bar, policy);
        }
        return;
      }
      while (i >= win0 + 32) {
        win0 += 32;
        tbl = (win0 + lane < n) ? bt[win0 + lane] : 0;
      }
      const int64_t blk = __shfl_sync(0xffffffffu, tbl, i - win0);
      if (elect_one()) {
        mbar_arrive_expect_tx(bar, STAGE_BYTES);
        bulk_g2s(dst, kbase + blk * p.kc_bs * 2, PAGE_BYTES, bar, policy);
        bulk_g2s(dst + PAGE_BYTES, vbase + blk * p.vc_bs * 2, PAGE_BYTES, bar, policy);
      }
    };
    if (warp == 0) {
      const int pre = min(NS, n_all);
      for (int i = 0; i < pre; ++i) issue(i, i);
    }
    // NS is a power of two: stage of page j = j & (NS - 1), phase parity = (j >> stage_shift) & 1.
    const int smask = NS - 1, sshift = p.stage_shift;
    constexpr bool kPairOk = PLAIN && BS == 16;  // two pages per step on the unmasked fast path
    int i = 0, i_prev = 0, next_issue = min(NS, n_all);
    if constexpr (kPairOk) {
      // Interior of the context (full pages, no window, nothing masked): a tight loop over page
      // pairs.  i stays even here and NS is even, so the pair's stages are adjacent (no wrap) and
      // share one phase parity.
      if (W == 0 && warp_on) {
        const int n_fast = ctx / BS;  // full context pages (page0 == 0 without a window)
        const int zi[RS] = {};
        const float zf[RS] = {};
        while (i + 2 <= n_fast) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {  // refill at most the two stages left one step ago
            const int j = next_issue;
            if (j < n_all && j - NS < i_prev) {
              if (warp == (j & (kPrefixWarps - 1))) {
                mbar_wait(empty_u32 + 8 * (j & smask), (uint32_t)(((j >> sshift) - 1) & 1));
                issue(j, j & smask);
              }
              ++next_issue;
            }
          }
          const int s = i & smask;
          const uint32_t par = (uint32_t)((i >> sshift) & 1);
          mbar_wait(full_u32 + 8 * s, par);
          mbar_wait(full_u32 + 8 * s + 8, par);
          const uint32_t ksm = ring_u32 + (uint32_t)s * STAGE_BYTES;
          process_page_ex<T, D, BS, KV, GROWS, false, true, false, true, 2>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, 2 * BS, p.scale_log2, slope, 0.f, row_on, zi, zi,
              zf, (uint32_t)STAGE_BYTES);
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(empty_u32 + 8 * s);
            mbar_arrive(empty_u32 + 8 * s + 8);
          }
          i_prev = i;
          i += 2;
        }
      }
    }
    if constexpr (PLAIN && BS == 32) {
      // Same interior loop for 32-token pages: one page is already a full 32-token step.
      if (W == 0 && warp_on) {
        const int n_fast = ctx / BS;
        const int zi[RS] = {};
        const float zf[RS] = {};
        while (i < n_fast) {
          const int j = next_issue;
          if (j < n_all && j - NS < i_prev) {
            if (warp == (j & (kPrefixWarps - 1))) {
              mbar_wait(empty_u32 + 8 * (j & smask), (uint32_t)(((j >> sshift) - 1) & 1));
              issue(j, j & smask);
            }
            ++next_issue;
          }
          const int s = i & smask;
          mbar_wait(full_u32 + 8 * s, (uint32_t)((i >> sshift) & 1));
          const uint32_t ksm = ring_u32 + (uint32_t)s * STAGE_BYTES;
          process_page_ex<T, D, BS, KV, GROWS, false, true, false, true>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, BS, p.scale_log2, slope, 0.f, row_on, zi, zi, zf);
          __syncwarp();
          if (lane == 0) mbar_arrive(empty_u32 + 8 * s);
          i_prev = i;
          ++i;
        }
      }
    }
    while (i < n_all) {
      if constexpr (kPairOk) {
        // Staged new-token slices entirely below this warp's causal frontier: same tight pair loop
        // (i may be odd here, so the two stages can wrap and carry different parities).
        if (W == 0 && warp_on && i >= n) {
          const int zi[RS] = {};
          const float zf[RS] = {};
          while (i + 1 < n_all && (slice0 + i - n) * BS + 2 * BS - 1 <= w_tok_lo) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int j = next_issue;
              if (j < n_all && j - NS < i_prev) {
                if (warp == (j & (kPrefixWarps - 1))) {
                  mbar_wait(empty_u32 + 8 * (j & smask), (uint32_t)(((j >> sshift) - 1) & 1));
                  issue(j, j & smask);
                }
                ++next_issue;
              }
            }
            const int s = i & smask, s1 = (i + 1) & smask;
            mbar_wait(full_u32 + 8 * s, (uint32_t)((i >> sshift) & 1));
            mbar_wait(full_u32 + 8 * s1, (uint32_t)(((i + 1) >> sshift) & 1));
            const uint32_t ksm = ring_u32 + (uint32_t)s * STAGE_BYTES;
            process_page_ex<T, D, BS, KV, GROWS, false, true, false, true, 2>(
                ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, 2 * BS, p.scale_log2, slope, 0.f, row_on, zi,
                zi, zf, (uint32_t)(s1 - s) * (uint32_t)STAGE_BYTES);
            __syncwarp();
            if (lane == 0) {
              mbar_arrive(empty_u32 + 8 * s);
              mbar_arrive(empty_u32 + 8 * s1);
            }
            i_prev = i;
            i += 2;
          }
          if (i >= n_all) break;
        }
      }
      // refill the stages every warp left at least one step ago (pages below i_prev)
      while (next_issue < n_all && next_issue - NS < i_prev) {
        const int j = next_issue++;
        if (warp == (j & (kPrefixWarps - 1))) {
          mbar_wait(empty_u32 + 8 * (j & smask), (uint32_t)(((j >> sshift) - 1) & 1));
          issue(j, j & smask);
        }
      }
      const int s = i & smask;
      const int s1 = (i + 1) & smask;
      const bool is_new = i >= n;
      // Position of the page's first key relative to the sequence start, in "key position" units:
      // context pages start at (page0 + i) * BS, staged slices at ctx + (slice0 + i - n) * BS.
      const int tok0 = is_new ? ctx + (slice0 + i - n) * BS : (page0 + i) * BS;
      const int hi_q = ctx + w_tok_hi, lo_q = ctx + w_tok_lo;  // query positions of this warp
      bool pair = false;
      if (kPairOk && W == 0 && warp_on && i + 1 < n_all)
        pair = is_new ? (tok0 + 2 * BS - 1 <= lo_q) : (tok0 + 2 * BS <= ctx);
      mbar_wait(full_u32 + 8 * s, (uint32_t)((i >> sshift) & 1));
      if (pair) mbar_wait(full_u32 + 8 * s1, (uint32_t)(((i + 1) >> sshift) & 1));
      if (warp_on && tok0 <= hi_q) {
        const uint32_t ksm = ring_u32 + (uint32_t)s * STAGE_BYTES;
        int row_lo[RS], row_hi[RS];
        float row_rel0[RS];
#pragma unroll
        for (int rs = 0; rs < RS; ++rs) {
          const int pos_q = ctx + max(row_tok[rs], 0);
          row_lo[rs] = W > 0 ? pos_q - W + 1 - tok0 : 0;
          row_hi[rs] = is_new ? pos_q - tok0 + 1 : BS;  // causal frontier inside the new tokens
          row_rel0[rs] = (float)(tok0 - pos_q);
        }
        const bool tail = !is_new && tok0 + BS > ctx;
        const bool masked = (W > 0 && (hi_q - W + 1 > tok0)) || (is_new && tok0 + BS - 1 > lo_q);
        if constexpr (kPairOk) {
          if (pair) {
            const uint32_t stride = (uint32_t)(s1 - s) * (uint32_t)STAGE_BYTES;  // wraps for s1 < s
            process_page_ex<T, D, BS, KV, GROWS, false, true, false, true, 2>(
                ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, 2 * BS, p.scale_log2, slope, 0.f, row_on,
                row_lo, row_hi, row_rel0, stride);
          }
        }
        if (pair) {
        } else if (tail)
          process_page_ex<T, D, BS, KV, GROWS, true, PLAIN, true, true>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, ctx - tok0, p.scale_log2, slope, 0.f, row_on,
              row_lo, row_hi, row_rel0);
        else if (masked || !PLAIN)
          process_page_ex<T, D, BS, KV, GROWS, false, PLAIN, true, true>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, BS, p.scale_log2, slope, 0.f, row_on, row_lo,
              row_hi, row_rel0);
        else
          process_page_ex<T, D, BS, KV, GROWS, false, true, false, true>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, BS, p.scale_log2, slope, 0.f, row_on, row_lo,
              row_hi, row_rel0);
      }
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(empty_u32 + 8 * s);
        if (pair) mbar_arrive(empty_u32 + 8 * s1);
      }
      i_prev = i;
      i += pair ? 2 : 1;
    }
  }
  __syncthreads();  // the ring becomes the staging area of phase 2

  // ============ phase 2 (no scratch given): new tokens staged by the CTA itself ============
  if (p.kv_scratch == nullptr) {
    const int last_tok = m0 + tile_tokens - 1;  // newest query of this CTA: keys j <= last_tok
    int slice0 = 0;
    if (W > 0) slice0 = max(0, m0 - W + 1) / BS;
    const int n = last_tok / BS + 1 - slice0;
    const T* kg = static_cast<const T*>(p.k) + (int64_t)kvh * p.k_s1;
    const T* vg = static_cast<const T*>(p.v) + (int64_t)kvh * p.v_s1;
    uint4 kr[VPT], vr[VPT];
    auto fetch = [&](int i) {  // slice i (tokens (slice0+i)*BS ..) -> registers
      const int t0 = (slice0 + i) * BS;
#pragma unroll
      for (int u = 0; u < VPT; ++u) {
        const int idx = threadIdx.x + u * NTHREADS;
        const int t = idx % BS, c = idx / BS;
        kr[u] = make_uint4(0u, 0u, 0u, 0u);
        vr[u] = make_uint4(0u, 0u, 0u, 0u);
        if (idx < VEC_PER_PAGE && t0 + t < qlen) {
          kr[u] = *reinterpret_cast<const uint4*>(kg + (start + t0 + t) * p.k_s0 + c * 8);
          vr[u] = *reinterpret_cast<const uint4*>(vg + (start + t0 + t) * p.v_s0 + c * 8);
        }
      }
    };
    auto stash = [&](int buf) {  // registers -> K page [D/8][BS][8], V page [D][BS]
      unsigned char* kd = ring + (size_t)buf * STAGE_BYTES;
      unsigned char* vd = kd + PAGE_BYTES;
#pragma unroll
      for (int u = 0; u < VPT; ++u) {
        const int idx = threadIdx.x + u * NTHREADS;
        const int t = idx % BS, c = idx / BS;
        if (idx < VEC_PER_PAGE) {
          *reinterpret_cast<uint4*>(kd + (size_t)(c * BS + t) * 16) = kr[u];
          const uint16_t* e = reinterpret_cast<const uint16_t*>(&vr[u]);
#pragma unroll
          for (int x = 0; x < 8; ++x)
            *reinterpret_cast<uint16_t*>(vd + (size_t)((c * 8 + x) * BS + t) * 2) = e[x];
        }
      }
    };
    if (n > 0) {
      fetch(0);
      stash(0);
    }
    __syncthreads();
    for (int i = 0; i < n; ++i) {
      if (i + 1 < n) fetch(i + 1);
      const int t0 = (slice0 + i) * BS;
      if (warp_on && t0 <= w_tok_hi) {
        const uint32_t ksm = ring_u32 + (uint32_t)(i & 1) * STAGE_BYTES;
        int row_lo[RS], row_hi[RS];
        float row_rel0[RS];
#pragma unroll
        for (int rs = 0; rs < RS; ++rs) {
          const int iq = max(row_tok[rs], 0);
          row_lo[rs] = W > 0 ? iq - W + 1 - t0 : 0;
          row_hi[rs] = iq - t0 + 1;  // causal: key j <= query i
          row_rel0[rs] = (float)(t0 - iq);
        }
        const bool frontier = t0 + BS - 1 > w_tok_lo;
        const bool windowed = W > 0 && (w_tok_hi - W + 1 > t0);
        if (frontier || windowed || !PLAIN)
          process_page_ex<T, D, BS, KV, GROWS, false, PLAIN, true, true>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, BS, p.scale_log2, slope, 0.f, row_on, row_lo,
              row_hi, row_rel0);
        else
          process_page_ex<T, D, BS, KV, GROWS, false, true, false, true>(
              ksm, ksm + PAGE_BYTES, qa, o, m, l, lane, BS, p.scale_log2, slope, 0.f, row_on, row_lo,
              row_hi, row_rel0);
      }
      if (i + 1 < n) stash((i + 1) & 1);
      __syncthreads();
    }
  }

  // ---- epilogue: normalise and store (the reference divides by the plain sum, :160-166) ----
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    float L = l[rs];
    L += __shfl_xor_sync(0xffffffffu, L, 1);
    L += __shfl_xor_sync(0xffffffffu, L, 2);
    if (row_tok[rs] < 0) continue;
    const float inv = L > 0.f ? 1.f / L : 0.f;
    T* orow = static_cast<T*>(p.out) + (start + row_tok[rs]) * p.o_s0 + (int64_t)row_head[rs] * p.o_s1;
#pragma unroll
    for (int j = 0; j < NDT; ++j) {
      const uint32_t packed = MmaOp<T>::pack(o[j][2 * rs] * inv, o[j][2 * rs + 1] * inv);
      *reinterpret_cast<uint32_t*>(orow + 8 * j + 2 * q4) = packed;
    }
  }
}

// Pre-pass of the staged mode: rewrites the packed new-token k/v of every (sequence, kv head) into
// page layout -- per BS-token slice one K page [D/8][BS][8] followed by one V page [D][BS], zero
// filled past the sequence's last new token -- so that the attention kernel can pull them through
// the same bulk-copy ring as the cached context.  Slice `sl` of sequence b lives at slice index
// start_loc[b] / BS + b + sl (sequences never overlap: sum of ceil(q/BS) <= start/BS + b).
// Traffic: 2x the new k/v bytes, a few MB; one launch.
template <typename T, int D, int BS>
__global__ void __launch_bounds__(128) prefix_stage_kv_kernel(const __grid_constant__ PrefixParams p,
                                                              unsigned char* scratch) {
  constexpr int PAGE_BYTES = D * BS * 2;
  constexpr int NDT = D / 8;
  constexpr int VEC_PER_PAGE = BS * NDT;
  // The tcgen05 tile loop behind this pre-pass is a programmatic dependent: it may start now (Q load, context tiles)
  // and waits for this grid only before its first read of the staged pages.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  __shared__ __align__(16) unsigned char sm[2 * PAGE_BYTES];
  const int seq = blockIdx.z, kvh = blockIdx.y, sl = blockIdx.x;
  const int qlen = p.seq_lens[seq] - p.ctx_lens[seq];
  const int t0 = sl * BS;
  if (t0 >= qlen) return;
  const int64_t start = p.start_loc[seq];
  const T* kg = static_cast<const T*>(p.k) + (int64_t)kvh * p.k_s1;
  const T* vg = static_cast<const T*>(p.v) + (int64_t)kvh * p.v_s1;
  for (int idx = threadIdx.x; idx < VEC_PER_PAGE; idx += blockDim.x) {
    const int t = idx % BS, c = idx / BS;
    uint4 kr = make_uint4(0u, 0u, 0u, 0u), vr = kr;
    if (t0 + t < qlen) {
      kr = *reinterpret_cast<const uint4*>(kg + (start + t0 + t) * p.k_s0 + c * 8);
      vr = *reinterpret_cast<const uint4*>(vg + (start + t0 + t) * p.v_s0 + c * 8);
    }
    *reinterpret_cast<uint4*>(sm + (size_t)(c * BS + t) * 16) = kr;
    const uint16_t* e = reinterpret_cast<const uint16_t*>(&vr);
#pragma unroll
    for (int x = 0; x < 8; ++x)
      *reinterpret_cast<uint16_t*>(sm + PAGE_BYTES + (size_t)((c * 8 + x) * BS + t) * 2) = e[x];
  }
  __syncthreads();
  const int64_t slice = (int64_t)(start / BS) + seq + sl;
  if (start < 0 || slice >= p.scratch_slices) return;  // non-packed start_loc: never write out of bounds
  unsigned char* dst = scratch + (slice * p.num_kv_heads + kvh) * (2 * PAGE_BYTES);
  for (int idx = threadIdx.x; idx < 2 * PAGE_BYTES / 16; idx += blockDim.x)
    reinterpret_cast<uint4*>(dst)[idx] = reinterpret_cast<const uint4*>(sm)[idx];
}

// float32 prefix prefill (the reference accepts fp32 q/k/v and caches with x = 4, halving its tile,
// prefix_prefill.py:690-693).  Compatibility path on the CUDA cores, exact fp32 products: one warp per
// (query token, head); 32 keys per round -- each lane owns one key for Q.K^T (16-byte loads; cached
// keys of a page are consecutive 16-byte rows), the warp takes the online-softmax step, then the
// lanes switch to owning d = lane + 32 i for P.V with the probabilities broadcast by shuffle.
__global__ void __launch_bounds__(128)
context_attention_f32_kernel(const __grid_constant__ PrefixParams p, int head_size, int block_size,
                             float scale) {
  extern __shared__ __align__(16) float q_s[];  // [4 warps][head_size]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int seq = blockIdx.z, head = blockIdx.y;
  const int ctx = p.ctx_lens[seq];
  const int qlen = p.seq_lens[seq] - ctx;
  const int i = blockIdx.x * 4 + warp;
  if (i >= qlen) return;
  const int D = head_size, BS = block_size;
  const int kvh = head / p.group;
  const int64_t start = p.start_loc[seq];
  const int pos_q = ctx + i;
  const float* q = static_cast<const float*>(p.q) + (start + i) * p.q_s0 + (int64_t)head * p.q_s1;
  float* qs = q_s + warp * D;
  for (int d = lane; d < D; d += 32) qs[d] = q[d];
  __syncwarp();
  const float* kc = static_cast<const float*>(p.k_cache) + (int64_t)kvh * p.kc_hs;
  const float* vc = static_cast<const float*>(p.v_cache) + (int64_t)kvh * p.vc_hs;
  const float* kn = static_cast<const float*>(p.k) + (int64_t)kvh * p.k_s1;
  const float* vn = static_cast<const float*>(p.v) + (int64_t)kvh * p.v_s1;
  const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
  const float slope = p.alibi_slopes ? p.alibi_slopes[head] : 0.f;
  const int W = p.sliding_window;
  const int n = pos_q + 1;  // visible keys: positions 0 .. pos_q
  int first = 0;
  if (W > 0) first = max(0, pos_q - W + 1) & ~31;
  float acc[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) acc[u] = 0.f;
  float m = kNegInf, l = 0.f;
  for (int g0 = first; g0 < n; g0 += 32) {
    const int key = g0 + lane;
    const bool vis = key < n && !(W > 0 && pos_q - key >= W);
    int64_t blk = 0;
    float logit = kNegInf;
    if (key < n) {
      float dot = 0.f;
      if (key < ctx) {
        blk = bt[key / BS];
        const float* base = kc + blk * p.kc_bs + (int64_t)(key % BS) * 4;
        for (int c = 0; c < D / 4; ++c) {
          const float4 kv = *reinterpret_cast<const float4*>(base + (int64_t)c * BS * 4);
          const float4 qv = *reinterpret_cast<const float4*>(qs + 4 * c);
          dot += qv.x * kv.x + qv.y * kv.y + qv.z * kv.z + qv.w * kv.w;
        }
      } else {
        const float* row = kn + (start + key - ctx) * p.k_s0;
        for (int c = 0; c < D / 4; ++c) {
          const float4 kv = *reinterpret_cast<const float4*>(row + 4 * c);
          const float4 qv = *reinterpret_cast<const float4*>(qs + 4 * c);
          dot += qv.x * kv.x + qv.y * kv.y + qv.z * kv.z + qv.w * kv.w;
        }
      }
      if (vis) logit = dot * scale + slope * (float)(key - pos_q);
    }
    float mx = logit;
#pragma unroll
    for (int sft = 16; sft >= 1; sft >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, sft));
    const float m_new = fmaxf(m, mx);
    if (m_new == kNegInf) continue;  // nothing visible yet (sliding window)
    const float alpha = expf(m - m_new);
    const float pr = vis ? expf(logit - m_new) : 0.f;
    float ps = pr;
#pragma unroll
    for (int sft = 16; sft >= 1; sft >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, sft);
    l = l * alpha + ps;
    m = m_new;
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] *= alpha;
    const int cnt = min(32, n - g0);
    for (int jj = 0; jj < cnt; ++jj) {
      const float pj = __shfl_sync(0xffffffffu, pr, jj);
      const int64_t bj = __shfl_sync(0xffffffffu, blk, jj);
      if (pj == 0.f) continue;  // warp-uniform
      const int kj = g0 + jj;
      if (kj < ctx) {
        const float* vb = vc + bj * p.vc_bs + kj % BS;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int d = lane + 32 * u;
          if (d < D) acc[u] += pj * vb[(int64_t)d * BS];
        }
      } else {
        const float* vr = vn + (start + kj - ctx) * p.v_s0;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int d = lane + 32 * u;
          if (d < D) acc[u] += pj * vr[d];
        }
      }
    }
  }
  float* o = static_cast<float*>(p.out) + (start + i) * p.o_s0 + (int64_t)head * p.o_s1;
  const float inv = l > 0.f ? 1.f / l : 0.f;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int d = lane + 32 * u;
    if (d < D) o[d] = acc[u] * inv;
  }
}

namespace {

template <int D, int BS>
constexpr size_t prefix_smem(int ns) {
  // + 1 KiB slack: ldmatrix.x4 over an odd number of 8-wide d chunks (head 120) reads past the page
  return (size_t)ns * 2 * D * BS * 2 + (size_t)ns * 16 + 1024;
}

// The staging pre-pass shared by both tile loops: packed new k/v -> page layout in the scratch.
template <typename T, int D, int BS>
int launch_prefix_stage(const PrefixParams& p, int num_seqs, int max_query_len, cudaStream_t stream) {
  dim3 sgrid((max_query_len + BS - 1) / BS, p.num_kv_heads, num_seqs);
  prefix_stage_kv_kernel<T, D, BS><<<sgrid, 128, 0, stream>>>(p, const_cast<unsigned char*>(p.kv_scratch));
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

// Can the tcgen05 tile loop (prefix_prefill_umma.cuh) serve this call?  Head 64 or 128, no ALiBi / window,
// staged new tokens, 16-byte aligned q / out rows, a GQA group that fits one 128-row M tile.
template <int D>
bool umma_eligible(const PrefixParams& p) {
  auto a16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0; };
  return (D == 128 || D == 64) && p.alibi_slopes == nullptr && p.sliding_window == 0 && p.kv_scratch != nullptr &&
         p.group <= 128 && a16(p.q) && a16(p.out) && p.q_s0 % 8 == 0 && p.q_s1 % 8 == 0 && p.o_s0 % 8 == 0 &&
         p.o_s1 % 8 == 0;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (no link against libcuda).
PFN_cuTensorMapEncodeTiled tensor_map_encoder() {
  static std::atomic<void*> cached{nullptr};
  void* fn = cached.load(std::memory_order_acquire);
  if (fn == nullptr) {
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &st) != cudaSuccess ||
        st != cudaDriverEntryPointSuccess)
      return nullptr;
    cached.store(fn, std::memory_order_release);
  }
  return reinterpret_cast<PFN_cuTensorMapEncodeTiled>(fn);
}

// Tensor maps of the TMA-fed tile loop (16-bit elements, block size 16).  A K page is [D/8][BS][8]: dim0 = the
// BS * 8 contiguous elements of one chunk, dim1 = chunk; the box is 8 tokens x all chunks, no swizzle.  A V page is
// [D][BS]: dim0 = token, dim1 = d; the box is the page, SWIZZLE_32B.  dim2 = kv head, dim3 = block (cache) or
// staged slice (scratch: slices of [K page | V page] per kv head).  Block ids come from the block table and are
// not range-checked by the copy engine: dim3 is declared as large as the format allows.
template <int D, int BS>
bool encode_prefix_maps(const PrefixParams& p, bool bf16, PrefixTmaMaps& m) {
  PFN_cuTensorMapEncodeTiled enc = tensor_map_encoder();
  if (enc == nullptr) return false;
  const CUtensorMapDataType dt = bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const cuuint32_t ones[4] = {1, 1, 1, 1};
  const cuuint64_t many = (cuuint64_t)1 << 31;
  const cuuint64_t slice_bytes = (cuuint64_t)2 * D * BS * 2;
  auto make = [&](unsigned char* out, const void* base, bool is_k, cuuint64_t head_bytes, cuuint64_t blk_bytes) {
    const cuuint64_t dims[4] = {is_k ? (cuuint64_t)BS * 8 : (cuuint64_t)BS, is_k ? (cuuint64_t)D / 8 : (cuuint64_t)D,
                                (cuuint64_t)p.num_kv_heads, many};
    const cuuint64_t strides[3] = {is_k ? (cuuint64_t)BS * 16 : (cuuint64_t)BS * 2, head_bytes, blk_bytes};
    const cuuint32_t box[4] = {is_k ? 64u : (cuuint32_t)BS, is_k ? (cuuint32_t)D / 8 : (cuuint32_t)D, 1, 1};
    return enc(reinterpret_cast<CUtensorMap*>(out), dt, 4, const_cast<void*>(base), dims, strides, box, ones,
               CU_TENSOR_MAP_INTERLEAVE_NONE, is_k ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_32B,
               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  };
  return make(m.k, p.k_cache, true, (cuuint64_t)p.kc_hs * 2, (cuuint64_t)p.kc_bs * 2) &&
         make(m.v, p.v_cache, false, (cuuint64_t)p.vc_hs * 2, (cuuint64_t)p.vc_bs * 2) &&
         make(m.sk, p.kv_scratch, true, slice_bytes, slice_bytes * p.num_kv_heads) &&
         make(m.sv, p.kv_scratch + (size_t)D * BS * 2, false, slice_bytes, slice_bytes * p.num_kv_heads);
}

// `gather_only`: keep the per-thread LDGSTS gather (B200_PREFIX_PATH_TCGEN05_GATHER; also what block sizes other
// than 16, caches whose strides are not 16-byte multiples and drivers without the encoder entry point get).
template <typename T, int D, int BS>
int launch_prefix_umma(PrefixParams& p, int num_seqs, int max_query_len, int device, cudaStream_t stream,
                       bool gather_only) {
  p.tok_tile = kUmmaRows / p.group;
  static std::atomic<int> configured[64];
  const int slot = device & 63;
  if (!configured[slot].load(std::memory_order_acquire)) {
    B200_CUDA_OK(cudaFuncSetAttribute(context_attention_umma_kernel<T, D, BS, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUmmaSmem));
    if constexpr (BS == 16 && D == 128)
      B200_CUDA_OK(cudaFuncSetAttribute(context_attention_umma_kernel<T, D, BS, true>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUmmaSmem));
    configured[slot].store(1, std::memory_order_release);
  }
  const int rc = launch_prefix_stage<T, D, BS>(p, num_seqs, max_query_len, stream);
  if (rc != B200_OK) return rc;
  const int tiles = (max_query_len + p.tok_tile - 1) / p.tok_tile;
  B200_CHECK_ARG(p.num_kv_heads <= 65535 && num_seqs <= 65535, "context_attention_fwd: grid too large");
  dim3 grid(tiles, p.num_kv_heads, num_seqs);
  PrefixTmaMaps maps;
  bool tma = false;
  // (head size 64: the boxes shrink to 1-2 KiB and the copies no longer beat the gather -- 538 vs 547 TFLOP/s)
  if constexpr (BS == 16 && D == 128) {
    auto a16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0; };
    tma = !gather_only && a16(p.k_cache) && a16(p.v_cache) && p.kc_hs % 8 == 0 && p.kc_bs % 8 == 0 &&
          p.vc_hs % 8 == 0 && p.vc_bs % 8 == 0 &&
          encode_prefix_maps<D, BS>(p, std::is_same<T, __nv_bfloat16>::value, maps);
  }
  // Programmatic dependent of the staging pre-pass launched above (same stream): the kernel executes
  // griddepcontrol.wait right before its first read of the staged pages and nowhere else -- everything it touches
  // earlier (q, block tables, the cache) was complete before the pre-pass itself could start.
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kUmmaThreads);
  cfg.dynamicSmemBytes = kUmmaSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if constexpr (BS == 16 && D == 128) {
    if (tma) B200_CUDA_OK(cudaLaunchKernelEx(&cfg, context_attention_umma_kernel<T, D, BS, true>, p, maps));
  }
  if (!tma) {
    memset(&maps, 0, sizeof(maps));
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, context_attention_umma_kernel<T, D, BS, false>, p, maps));
  }
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

template <typename T, int D, int BS>
int launch_prefix(PrefixParams& p, int num_seqs, int max_query_len, int device, cudaStream_t stream, int path) {
  if constexpr (D == 128 || D == 64) {
    const bool can = umma_eligible<D>(p);
    const bool forced = path == B200_PREFIX_PATH_TCGEN05 || path == B200_PREFIX_PATH_TCGEN05_GATHER;
    if (forced && !can) {
      set_error("context_attention_fwd: the tcgen05 path needs head size 64 or 128, no ALiBi / sliding window, "
                "a staging scratch and 16-byte aligned q / out rows");
      return B200_ERR_UNSUPPORTED;
    }
    // auto: the 256-row tile pays once a sequence brings at least one full M tile of rows
    if (can && (forced || (path == B200_PREFIX_PATH_AUTO && (int64_t)max_query_len * p.group >= 128)))
      return launch_prefix_umma<T, D, BS>(p, num_seqs, max_query_len, device, stream,
                                          path == B200_PREFIX_PATH_TCGEN05_GATHER);
  } else if (path == B200_PREFIX_PATH_TCGEN05 || path == B200_PREFIX_PATH_TCGEN05_GATHER) {
    set_error("context_attention_fwd: the tcgen05 path needs head size 64 or 128 (got %d)", D);
    return B200_ERR_UNSUPPORTED;
  }
  constexpr int GROWS = D <= 128 ? 16 : 8;
  constexpr int ROWS = kPrefixWarps * GROWS;
  B200_CHECK_ARG(p.group <= ROWS, "context_attention_fwd: %d q-heads per kv-head exceed the %d-row tile",
                 p.group, ROWS);
  p.tok_tile = ROWS / p.group;
  int ns = 8;  // ring depth (a power of two): three CTAs per SM must still fit in 227 KiB
  while (ns > 2 && prefix_smem<D, BS>(ns) > (size_t)72 * 1024) ns /= 2;
  p.num_stages = ns;
  p.stage_shift = ns == 8 ? 3 : ns == 4 ? 2 : 1;
  const size_t smem = prefix_smem<D, BS>(ns);
  const bool plain = p.alibi_slopes == nullptr;
  auto kern = plain ? context_attention_kernel<T, D, BS, GROWS, true>
                    : context_attention_kernel<T, D, BS, GROWS, false>;
  static std::atomic<int> configured[2][64];
  const int slot = device & 63;
  if (!configured[plain][slot].load(std::memory_order_acquire)) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
    configured[plain][slot].store(1, std::memory_order_release);
  }
  if (p.kv_scratch != nullptr) {
    const int rc = launch_prefix_stage<T, D, BS>(p, num_seqs, max_query_len, stream);
    if (rc != B200_OK) return rc;
  }
  const int tiles = (max_query_len + p.tok_tile - 1) / p.tok_tile;
  B200_CHECK_ARG(p.num_kv_heads <= 65535 && num_seqs <= 65535, "context_attention_fwd: grid too large");
  // longest tiles first would need the per-sequence lengths on the host; x = tile keeps the CTAs of
  // one (sequence, kv head) adjacent so that they share the context pages in L2
  dim3 grid(tiles, p.num_kv_heads, num_seqs);
  kern<<<grid, kPrefixWarps * 32, smem, stream>>>(p);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

template <typename T>
int dispatch_prefix(PrefixParams& p, int head_size, int block_size, int num_seqs, int max_query_len,
                    int device, cudaStream_t stream, int path) {
#define B200_PFX(DD, BB) \
  if (head_size == DD && block_size == BB) \
    return launch_prefix<T, DD, BB>(p, num_seqs, max_query_len, device, stream, path);
#ifdef B200_FAST_BUILD
  B200_PFX(64, 16) B200_PFX(128, 16)
#else
#define B200_PFX_D(DD) B200_PFX(DD, 8) B200_PFX(DD, 16) B200_PFX(DD, 32)
  B200_PFX_D(64) B200_PFX_D(80) B200_PFX_D(96) B200_PFX_D(112) B200_PFX_D(120) B200_PFX_D(128)
  B200_PFX_D(192) B200_PFX_D(256)
#undef B200_PFX_D
#endif
#undef B200_PFX
  if (block_size != 8 && block_size != 16 && block_size != 32)
    set_error("Unsupported block size: %d", block_size);
  else
    set_error("Unsupported head size: %d", head_size);
  return B200_ERR_UNSUPPORTED;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int64_t b200_context_attention_scratch_bytes(int64_t num_tokens, int num_seqs,
                                                        int num_kv_heads, int head_size,
                                                        int block_size) {
  if (num_tokens < 0 || num_seqs < 0 || block_size <= 0) return -1;
  return (num_tokens / block_size + num_seqs + 1) * (int64_t)num_kv_heads * 2 * head_size * block_size * 2;
}

extern "C" int b200_context_attention_fwd_path(
    const void* q, const void* k, const void* v, void* out, const void* key_cache,
    const void* value_cache, const int32_t* block_tables, const int32_t* start_loc,
    const int32_t* seq_lens, const int32_t* ctx_lens, int num_seqs, int num_heads, int num_kv_heads,
    int head_size, int block_size, int max_blocks_per_seq, int max_query_len, int64_t q_stride0,
    int64_t q_stride1, int64_t k_stride0, int64_t k_stride1, int64_t v_stride0, int64_t v_stride1,
    int64_t o_stride0, int64_t o_stride1, int64_t kc_block_stride, int64_t kc_head_stride,
    int64_t vc_block_stride, int64_t vc_head_stride, const float* alibi_slopes, int sliding_window,
    float scale, void* kv_scratch, int64_t kv_scratch_bytes, int64_t num_tokens, int dtype,
    int device, b200_stream_t stream, int path) {
  B200_CHECK_ARG(path == B200_PREFIX_PATH_AUTO || path == B200_PREFIX_PATH_MMA_SYNC ||
                     path == B200_PREFIX_PATH_TCGEN05 || path == B200_PREFIX_PATH_TCGEN05_GATHER,
                 "context_attention_fwd: unknown path %d", path);
  B200_CHECK_ARG(num_seqs >= 0 && max_query_len >= 0 && head_size > 0 && block_size > 0,
                 "context_attention_fwd: bad sizes");
  if (num_seqs == 0 || max_query_len == 0) return B200_OK;
  B200_CHECK_ARG(num_kv_heads > 0 && num_heads % num_kv_heads == 0,
                 "num_heads (%d) must be a positive multiple of num_kv_heads (%d)", num_heads, num_kv_heads);
  if (dtype != B200_DT_F16 && dtype != B200_DT_BF16 && dtype != B200_DT_F32) {
    set_error("Unsupported data type: context_attention_fwd runs fp32 / fp16 / bf16 (got %d)", dtype);
    return B200_ERR_UNSUPPORTED;
  }
  const int esz = dtype == B200_DT_F32 ? 4 : 2;
  auto a16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0; };
  const int va = 16 / esz;  // elements per 16 bytes
  B200_CHECK_ARG(a16(k) && a16(v) && a16(key_cache) && a16(value_cache) && k_stride0 % va == 0 &&
                     k_stride1 % va == 0 && v_stride0 % va == 0 && v_stride1 % va == 0 &&
                     kc_block_stride % va == 0 && kc_head_stride % va == 0 && vc_block_stride % va == 0 &&
                     vc_head_stride % va == 0,
                 "context_attention_fwd: k/v/cache rows must be 16-byte aligned");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(q) & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0 &&
                     q_stride0 % 2 == 0 && q_stride1 % 2 == 0 && o_stride0 % 2 == 0 && o_stride1 % 2 == 0,
                 "context_attention_fwd: q/out rows must be 4-byte aligned");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  int dev = device;
  if (dev < 0) B200_CUDA_OK(cudaGetDevice(&dev));
  PrefixParams p{};
  p.q = q; p.k = k; p.v = v; p.out = out; p.k_cache = key_cache; p.v_cache = value_cache;
  p.block_tables = block_tables; p.start_loc = start_loc; p.seq_lens = seq_lens; p.ctx_lens = ctx_lens;
  p.alibi_slopes = alibi_slopes;
  p.num_heads = num_heads; p.num_kv_heads = num_kv_heads; p.group = num_heads / num_kv_heads;
  p.max_blocks_per_seq = max_blocks_per_seq;
  p.sliding_window = (sliding_window > 0 && alibi_slopes == nullptr) ? sliding_window : 0;  // :713-811
  p.q_s0 = q_stride0; p.q_s1 = q_stride1; p.k_s0 = k_stride0; p.k_s1 = k_stride1;
  p.v_s0 = v_stride0; p.v_s1 = v_stride1; p.o_s0 = o_stride0; p.o_s1 = o_stride1;
  p.kc_bs = kc_block_stride; p.kc_hs = kc_head_stride; p.vc_bs = vc_block_stride; p.vc_hs = vc_head_stride;
  p.scale_log2 = scale * 1.4426950408889634f;
  // staged mode needs room for num_tokens / block_size + num_seqs + 1 slices of every kv head
  const int64_t need = b200_context_attention_scratch_bytes(num_tokens, num_seqs, num_kv_heads, head_size,
                                                            block_size);
  p.kv_scratch = (kv_scratch != nullptr && kv_scratch_bytes >= need &&
                  (reinterpret_cast<uintptr_t>(kv_scratch) & 15) == 0)
                     ? static_cast<const unsigned char*>(kv_scratch)
                     : nullptr;
  p.scratch_slices = p.kv_scratch ? kv_scratch_bytes / ((int64_t)num_kv_heads * 2 * head_size * block_size * 2) : 0;
  if (dtype == B200_DT_F32 && (path == B200_PREFIX_PATH_TCGEN05 || path == B200_PREFIX_PATH_TCGEN05_GATHER)) {
    set_error("context_attention_fwd: the tcgen05 path runs fp16 / bf16 only");
    return B200_ERR_UNSUPPORTED;
  }
  if (dtype == B200_DT_F32) {
    B200_CHECK_ARG(head_size % 4 == 0 && head_size <= 256 && (reinterpret_cast<uintptr_t>(q) & 15) == 0 &&
                       q_stride0 % 4 == 0 && q_stride1 % 4 == 0,
                   "context_attention_fwd(fp32): head size must be a multiple of 4 (<= 256), q rows 16-byte aligned");
    B200_CHECK_ARG(num_heads <= 65535 && num_seqs <= 65535, "context_attention_fwd: grid too large");
    dim3 grid((max_query_len + 3) / 4, num_heads, num_seqs);
    context_attention_f32_kernel<<<grid, 128, (size_t)4 * head_size * sizeof(float), (cudaStream_t)stream>>>(
        p, head_size, block_size, scale);
    count_launch();
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
  }
  if (dtype == B200_DT_F16)
    return dispatch_prefix<__half>(p, head_size, block_size, num_seqs, max_query_len, dev, (cudaStream_t)stream,
                                   path);
  return dispatch_prefix<__nv_bfloat16>(p, head_size, block_size, num_seqs, max_query_len, dev,
                                        (cudaStream_t)stream, path);
}

extern "C" int b200_context_attention_fwd(
    const void* q, const void* k, const void* v, void* out, const void* key_cache,
    const void* value_cache, const int32_t* block_tables, const int32_t* start_loc,
    const int32_t* seq_lens, const int32_t* ctx_lens, int num_seqs, int num_heads, int num_kv_heads,
    int head_size, int block_size, int max_blocks_per_seq, int max_query_len, int64_t q_stride0,
    int64_t q_stride1, int64_t k_stride0, int64_t k_stride1, int64_t v_stride0, int64_t v_stride1,
    int64_t o_stride0, int64_t o_stride1, int64_t kc_block_stride, int64_t kc_head_stride,
    int64_t vc_block_stride, int64_t vc_head_stride, const float* alibi_slopes, int sliding_window,
    float scale, void* kv_scratch, int64_t kv_scratch_bytes, int64_t num_tokens, int dtype,
    int device, b200_stream_t stream) {
  return b200_context_attention_fwd_path(
      q, k, v, out, key_cache, value_cache, block_tables, start_loc, seq_lens, ctx_lens, num_seqs, num_heads,
      num_kv_heads, head_size, block_size, max_blocks_per_seq, max_query_len, q_stride0, q_stride1, k_stride0,
      k_stride1, v_stride0, v_stride1, o_stride0, o_stride1, kc_block_stride, kc_head_stride, vc_block_stride,
      vc_head_stride, alibi_slopes, sliding_window, scale, kv_scratch, kv_scratch_bytes, num_tokens, dtype,
      device, stream, B200_PREFIX_PATH_AUTO);
}

#ifdef B200_UMMA_TRACE
extern "C" int b200_debug_umma_trace(long long* out) {  // debug build only: [4][128][8] clock64 stamps
  return (int)cudaMemcpyFromSymbol(out, b200::g_umma_trace, sizeof(long long) * 4 * 128 * 8);
}
#endif
