// Prefix / chunked-prefill attention on the 5th-generation tensor cores (tcgen05.mma, accumulators in
// TMEM): the compute-bound tile loop of context_attention_fwd
// (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811) for head size 128,
// fp16 / bf16, no ALiBi, no sliding window.  Everything else stays on prefix_prefill.cu.
//
// One CTA (16 warps, one per SM) owns (sequence, KV head, tile of 256 / G query tokens): its 256 rows
// (query token, q-head of the GQA group) are two 128-row M tiles that share every K/V tile.
//
//   warps 0-3 / 4-7   softmax warpgroup of M tile 0 / 1: one thread per row (= TMEM lane).  S comes back
//                     with tcgen05.ld, the row maximum needs no shuffle, P = exp2(S*c - m) is written
//                     back over S as 16-bit pairs (tcgen05.st) and becomes the A operand of P.V straight
//                     from TMEM.  The output accumulator is only rescaled when a row's maximum grew by
//                     more than 2^8 (the stale maximum stays the reference point otherwise), so the
//                     TMEM round trip of O is rare.
//   warps 8-11        gather: the cache keeps a K page as [D/8][BS][8] and a V page as [D][BS]; neither is
//                     an operand layout tcgen05.mma can address across pages.  Both ARE permutations of
//                     16-byte chunks of the K-major SWIZZLE_128B operand tiles, so 128 threads move a
//                     128-key K tile and V tile with 32 cp.async (LDGSTS, 16 B) each -- one block-table
//                     lookup per thread and tile, global reads in 128..512-byte runs, bank-conflict-free
//                     shared-memory writes, no register staging, no extra pass over HBM.  New tokens
//                     arrive as pages too (prefix_stage_kv_kernel), so there is one gather path.
//   warp 12           one elected thread issues every tcgen05.mma: S_w = Q_w K^T (SS form, 8 k-steps of
//                     128x128x16) and O_w += P_w V (TS form), and signals mbarriers with tcgen05.commit.
//
// TMEM (512 columns): S_0 | S_1 | O_0 | O_1, 128 fp32 columns each; P_w aliases the first 64 of S_w.
// Shared memory: Q 2 x 32 KiB, a ring of three K tiles and a ring of two V tiles (32 KiB each), all in
// 64-element-of-K slabs of 128-byte rows.  The MMA thread runs one step ahead on the S side:
// P.V_w(j) is followed at once by S_w(j+1) = Q_w K(j+1)^T, so softmax_w(j+1) overlaps the other
// M tile's P.V -- the two warpgroups ping-pong on the tensor pipe and on the MUFU pipe.
#pragma once

#include <type_traits>

#include "prefix_prefill.cuh"
#include "umma.cuh"

namespace b200 {

constexpr int kUmmaThreads = 512;
constexpr int kUmmaRows = 256;         // rows per CTA (two M tiles)
constexpr int kUmmaKeys = 128;         // keys per K/V tile
constexpr uint32_t kUmmaTile = 32768;  // bytes of one 128 x 128 16-bit operand tile
constexpr int kUmmaKStages = 3, kUmmaVStages = 2;
constexpr uint32_t kUmmaSmem = (2 + kUmmaKStages + kUmmaVStages) * kUmmaTile + 128 + 1024;

template <typename T, int BS>
__global__ void __launch_bounds__(kUmmaThreads, 1)
context_attention_umma_kernel(const __grid_constant__ PrefixParams p) {
  using namespace umma;
  constexpr int D = 128;
  constexpr int PAGE_BYTES = D * BS * 2;
  constexpr int SLICE_BYTES = 2 * PAGE_BYTES;
  constexpr int FMT = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
  constexpr uint32_t FULL = 0xffffffffu;
  constexpr float kNInf = -__builtin_huge_valf();

  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  const int seq = blockIdx.z, kvh = blockIdx.y;
  const int ctx = p.ctx_lens[seq];
  const int qlen = p.seq_lens[seq] - ctx;
  const int G = p.group;
  const int m0 = (int)(gridDim.x - 1 - blockIdx.x) * p.tok_tile;  // heaviest (latest) query tiles first
  if (m0 >= qlen) return;
  const int64_t start = p.start_loc[seq];
  const int tile_tokens = min(p.tok_tile, qlen - m0);
  // the staged new-token pages are indexed by start_loc (device data): never read outside the scratch
  if (start < 0 || (int64_t)(start / BS) + seq + (m0 + tile_tokens - 1) / BS >= p.scratch_slices) {
    for (int r = tid; r < tile_tokens * G; r += kUmmaThreads) {
      T* orow = static_cast<T*>(p.out) + (start + m0 + r / G) * p.o_s0 + (int64_t)(kvh * G + r % G) * p.o_s1;
      for (int d = 0; d < D; ++d) orow[d] = from_f32<T>(__int_as_float(0x7fc00000));
    }
    return;
  }
  const int rows_valid = tile_tokens * G;
  const int nq = rows_valid > 128 ? 2 : 1;  // M tiles that hold at least one real row
  const int n_ctx_tiles = (ctx + kUmmaKeys - 1) / kUmmaKeys;
  const int last_tok = m0 + tile_tokens - 1;
  const int nt = n_ctx_tiles + last_tok / kUmmaKeys + 1;

  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_s = base, k_ring = base + 2 * kUmmaTile, v_ring = k_ring + kUmmaKStages * kUmmaTile,
                 bars = v_ring + kUmmaVStages * kUmmaTile;
  const uint32_t k_full = bars, k_empty = bars + 24, v_full = bars + 48, v_empty = bars + 64, s_full = bars + 80,
                 p_full = bars + 96, slot = bars + 112;

  if (tid == 0) {
    for (int i = 0; i < kUmmaKStages; ++i) {
      mbar_init(k_full + 8 * i, 128);
      mbar_init(k_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(v_full + 8 * i, 128);
      mbar_init(v_empty + 8 * i, 1);
      mbar_init(s_full + 8 * i, 1);
      mbar_init(p_full + 8 * i, 128);
    }
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(slot, 512);

  // ---- Q: rows (token, head) -> two K-major SWIZZLE_128B tiles (padding rows repeat the last token) ----
  {
    const T* qg = static_cast<const T*>(p.q);
#pragma unroll
    for (int i = 0; i < kUmmaRows * 16 / kUmmaThreads; ++i) {
      const int id = tid + kUmmaThreads * i;
      const int R = id >> 4, c = id & 15;
      if (R < 128 * nq) {
        const int ti = min(R / G, tile_tokens - 1);
        const int head = kvh * G + R % G;
        const T* src = qg + (start + m0 + ti) * p.q_s0 + (int64_t)head * p.q_s1 + c * 8;
        const int r = R & 127;
        cp_async16(q_s + (uint32_t)(R >> 7) * kUmmaTile + (uint32_t)(c >> 3) * 16384u + sw128_offset(r, c & 7),
                   src, 16);
      }
    }
    cp_async_commit();
    cp_async_wait_all();
    fence_proxy_async_smem();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));

  if (warp < 8) {
    setmaxnreg_inc<200>();
  } else {
    setmaxnreg_dec<56>();
  }

  if (warp < 8) {
    // ======================= softmax warpgroups: one thread per row =======================
    const int w = warp >> 2;
    if (w < nq) {
      const int r = (warp & 3) * 32 + lane;
      const int R = w * 128 + r;
      const int ti = R / G;
      const bool valid = ti < tile_tokens;
      const int iq = m0 + min(ti, tile_tokens - 1);  // query index among the new tokens
      const uint32_t lane_bits = (uint32_t)((warp & 3) * 32) << 16;
      const uint32_t t_s = tmem + lane_bits + (uint32_t)w * 128u;
      const uint32_t t_o = tmem + lane_bits + 256u + (uint32_t)w * 128u;
      const float c2 = p.scale_log2;
      float m_ref = kNInf, l = 0.f;
      for (int j = 0; j < nt; ++j) {
        mbar_wait_wd(s_full + 8 * w, (uint32_t)(j & 1));
        fence_after_sync();
        uint32_t s[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(t_s + 32 * c, s[c]);
        tmem_wait_ld();
        // visible keys of this row in the tile: a prefix [0, lim)
        int lim;
        if (j < n_ctx_tiles)
          lim = min(kUmmaKeys, ctx - kUmmaKeys * j);
        else
          lim = min(kUmmaKeys, max(0, iq - kUmmaKeys * (j - n_ctx_tiles) + 1));
        if (__any_sync(FULL, lim < kUmmaKeys)) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (32 * c + i >= lim) s[c][i] = __float_as_uint(kNInf);
        }
        float mx0 = kNInf, mx1 = kNInf, mx2 = kNInf, mx3 = kNInf;
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            mx0 = fmaxf(mx0, __uint_as_float(s[c][i]));
            mx1 = fmaxf(mx1, __uint_as_float(s[c][i + 1]));
            mx2 = fmaxf(mx2, __uint_as_float(s[c][i + 2]));
            mx3 = fmaxf(mx3, __uint_as_float(s[c][i + 3]));
          }
        const float mloc = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * c2;
        // lazy rescale: keep the stale reference maximum unless the row maximum outgrew it by 2^8
        float alpha = 1.f;
        bool resc = false;
        if (m_ref == kNInf) {
          m_ref = mloc;  // first visible keys of this row (nothing accumulated yet: l == 0, O row all zero)
        } else if (mloc > m_ref + 8.f) {
          alpha = fast_exp2(m_ref - mloc);
          l *= alpha;
          m_ref = mloc;
          resc = true;
        }
        if (j > 0 && __any_sync(FULL, resc)) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            tmem_ld32(t_o + 32 * c, o);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(t_o + 32 * c, o);
          }
        }
        const float neg_m = m_ref == kNInf ? 0.f : -m_ref;
        float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t pk[32];
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            const int c = 2 * h + cc;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float e0 = fast_exp2(fmaf(__uint_as_float(s[c][i]), c2, neg_m));
              const float e1 = fast_exp2(fmaf(__uint_as_float(s[c][i + 1]), c2, neg_m));
              const float e2 = fast_exp2(fmaf(__uint_as_float(s[c][i + 2]), c2, neg_m));
              const float e3 = fast_exp2(fmaf(__uint_as_float(s[c][i + 3]), c2, neg_m));
              l0 += e0;
              l1 += e1;
              l2 += e2;
              l3 += e3;
              pk[16 * cc + i / 2] = MmaOp<T>::pack(e0, e1);
              pk[16 * cc + i / 2 + 1] = MmaOp<T>::pack(e2, e3);
            }
          }
          tmem_st32(t_s + 32 * h, pk);  // P over the first 64 columns of S (all of S is in registers)
        }
        l += (l0 + l1) + (l2 + l3);
        tmem_wait_st();
        fence_before_sync();
        mbar_arrive(p_full + 8 * w);
      }
      // ---- epilogue: O / l -> out ----
      mbar_wait_wd(s_full + 8 * w, (uint32_t)(nt & 1));
      fence_after_sync();
      const float inv = l > 0.f ? 1.f / l : 0.f;
      T* orow = static_cast<T*>(p.out) + (start + m0 + min(ti, tile_tokens - 1)) * p.o_s0 +
                (int64_t)(kvh * G + R % G) * p.o_s1;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        tmem_ld32(t_o + 32 * c, o);
        tmem_wait_ld();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 v;
            v.x = MmaOp<T>::pack(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv);
            v.y = MmaOp<T>::pack(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
            v.z = MmaOp<T>::pack(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
            v.w = MmaOp<T>::pack(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + 32 * c + i) = v;
          }
        }
      }
    }
  } else if (warp < 12) {
    // ======================= gather warps: pages -> swizzled operand tiles =======================
    const int pt = tid - 256;
    const int gidx = pt >> 3, l8 = pt & 7;  // group of 8 keys inside the tile, lane inside the group
    const unsigned char* kbase = static_cast<const unsigned char*>(p.k_cache) + (int64_t)kvh * p.kc_hs * 2;
    const unsigned char* vbase = static_cast<const unsigned char*>(p.v_cache) + (int64_t)kvh * p.vc_hs * 2;
    const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
    const unsigned char* sbase =
        p.kv_scratch + (((int64_t)(start / BS) + seq) * p.num_kv_heads + kvh) * SLICE_BYTES;
    const int64_t slice_stride = (int64_t)p.num_kv_heads * SLICE_BYTES;
    const uint32_t k_off = (uint32_t)(8 * gidx + l8) * 128u;  // K row of this thread's key
    const uint32_t v_off =
        (uint32_t)(gidx >> 3) * 16384u + (uint32_t)l8 * 128u + (uint32_t)(((gidx & 7) ^ l8) << 4);
    int ks = 0, kround = 0;  // K ring position of tile j, number of times the ring wrapped
    for (int j = 0; j < nt; ++j) {
      const bool is_ctx = j < n_ctx_tiles;
      const int pos = kUmmaKeys * (is_ctx ? j : j - n_ctx_tiles) + 8 * gidx;  // first key of the group
      const int nval = min(8, max(0, (is_ctx ? ctx : qlen) - pos));           // its valid keys
      const int toff = pos % BS;
      const unsigned char *kpage = kbase, *vpage = vbase;
      if (nval > 0) {
        if (is_ctx) {
          const int64_t blk = bt[pos / BS];
          kpage = kbase + blk * p.kc_bs * 2;
          vpage = vbase + blk * p.vc_bs * 2;
        } else {
          kpage = sbase + (int64_t)(pos / BS) * slice_stride;
          vpage = kpage + PAGE_BYTES;
        }
      }
      {
        if (kround > 0) mbar_wait_wd(k_empty + 8 * ks, (uint32_t)((kround - 1) & 1));
        const unsigned char* src = kpage + (toff + l8) * 16;
        const uint32_t nb = l8 < nval ? 16u : 0u;
        const uint32_t dst = k_ring + (uint32_t)ks * kUmmaTile + k_off;
#pragma unroll
        for (int c = 0; c < 16; ++c)
          cp_async16(dst + (uint32_t)(c >> 3) * 16384u + (uint32_t)(((c & 7) ^ l8) << 4), src + c * BS * 16, nb);
        cp_async_arrive_noinc(k_full + 8 * ks);
        if (++ks == kUmmaKStages) {
          ks = 0;
          ++kround;
        }
      }
      {
        const int vs = j & 1;
        if (j >= 2) mbar_wait_wd(v_empty + 8 * vs, (uint32_t)(((j >> 1) - 1) & 1));
        const unsigned char* src = vpage + toff * 2 + l8 * BS * 2;
        const uint32_t nb = nval > 0 ? 16u : 0u;
        const uint32_t dst = v_ring + (uint32_t)vs * kUmmaTile + v_off;
#pragma unroll
        for (int i = 0; i < 16; ++i) cp_async16(dst + (uint32_t)i * 1024u, src + i * 8 * BS * 2, nb);
        if (is_ctx && nval > 0 && nval < 8) {
          // last context page: slots past ctx_len may hold anything (NaN included); P is 0 there, but
          // 0 * NaN is not -- clear them in the V tile (staged new-token pages are already zero filled)
          cp_async_commit();
          cp_async_wait_all();
          uint32_t mk[4];
#pragma unroll
          for (int wd = 0; wd < 4; ++wd)
            mk[wd] = (2 * wd < nval ? 0x0000ffffu : 0u) | (2 * wd + 1 < nval ? 0xffff0000u : 0u);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint32_t a = dst + (uint32_t)i * 1024u;
            uint32_t x, y, z, u;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(u) : "r"(a));
            x &= mk[0];
            y &= mk[1];
            z &= mk[2];
            u &= mk[3];
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(u)
                         : "memory");
          }
          fence_proxy_async_smem();
          mbar_arrive(v_full + 8 * vs);
        } else {
          cp_async_arrive_noinc(v_full + 8 * vs);
        }
      }
    }
    cp_async_commit();
    cp_async_wait_all();  // nothing of this thread is still in flight when the CTA winds down
  } else if (warp == 12) {
    // ======================= MMA issuer =======================
    constexpr uint32_t idesc = idesc_f16(FMT, 128, 128);
    // S_w = Q_w K(tile in ring slot ks)^T, then "S_w ready"
    auto issue_qk = [&](int w, int ks) {
      const uint64_t da = sw128_desc(q_s + (uint32_t)w * kUmmaTile);
      const uint64_t db = sw128_desc(k_ring + (uint32_t)ks * kUmmaTile);
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t adv = (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
        mma_ss(tmem + (uint32_t)w * 128u, desc_advance(da, adv), desc_advance(db, adv), idesc, kk > 0);
      }
      commit(s_full + 8 * w);
    };
    // the gather wrote the tiles through the generic proxy: order those writes (acquired through the
    // mbarrier) before this thread's async-proxy reads
    mbar_wait_wd(k_full, 0u);
    fence_proxy_async_smem();
    fence_after_sync();
    if (lane == 0) {
      for (int w = 0; w < nq; ++w) issue_qk(w, 0);
      commit(k_empty);
    }
    __syncwarp();
    int ks = 1 % kUmmaKStages, kround = 0;  // ring position of tile j + 1
    for (int j = 0; j < nt; ++j) {
      const int vs = j & 1;
      mbar_wait_wd(v_full + 8 * vs, (uint32_t)((j >> 1) & 1));
      fence_proxy_async_smem();
      const bool more = j + 1 < nt;
      for (int w = 0; w < nq; ++w) {
        mbar_wait_wd(p_full + 8 * w, (uint32_t)(j & 1));
        if (more && w == 0) {
          mbar_wait_wd(k_full + 8 * ks, (uint32_t)(kround & 1));
          fence_proxy_async_smem();
        }
        fence_after_sync();
        if (lane == 0) {
          const uint64_t db = sw128_desc(v_ring + (uint32_t)vs * kUmmaTile);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t adv = (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
            mma_ts(tmem + 256u + (uint32_t)w * 128u, tmem + (uint32_t)w * 128u + (uint32_t)kk * 8u,
                   desc_advance(db, adv), idesc, (j > 0 || kk > 0) ? 1u : 0u);
          }
          if (w == nq - 1) commit(v_empty + 8 * vs);
          if (more) {
            issue_qk(w, ks);  // overwrites P_w(j): ordered behind P.V_w(j) in the tensor pipe
            if (w == nq - 1) commit(k_empty + 8 * ks);
          }
        }
        __syncwarp();
      }
      if (++ks == kUmmaKStages) {
        ks = 0;
        ++kround;
      }
    }
    if (lane == 0)
      for (int w = 0; w < nq; ++w) commit(s_full + 8 * w);  // "O_w is final" (tracks every MMA above)
    __syncwarp();
  }

  fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    fence_after_sync();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace b200
