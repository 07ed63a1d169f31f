// Prefix / chunked-prefill attention on the 5th-generation tensor cores (tcgen05.mma, accumulators in
// TMEM): the compute-bound tile loop of context_attention_fwd
// (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811) for head size 64 or 128,
// fp16 / bf16, no ALiBi, no sliding window.  Everything else stays on prefix_prefill.cu.
//
// One CTA (16 warps, one per SM) owns (sequence, KV head, tile of 256 / G query tokens): its 256 rows
// (query token, q-head of the GQA group) are two 128-row M tiles that share every K/V tile.  Keys are
// consumed in steps of 64.
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
//                     64-key K tile and V tile with 16 cp.async (LDGSTS, 16 B) each -- one block-table
//                     lookup per thread and tile, global reads in 128..512-byte runs, bank-conflict-free
//                     shared-memory writes, no register staging, no extra pass over HBM; completion is
//                     signalled by cp.async.mbarrier.arrive, so the gather never blocks on its own
//                     copies.  New tokens arrive as pages too (prefix_stage_kv_kernel): one gather path.
//   warps 12 / 13     one elected thread per M tile issues its tcgen05.mma stream: S_w = Q_w K^T (SS form,
//                     128x64x16) and O_w += P_w V (TS form, 128x128x16), and signals mbarriers with
//                     tcgen05.commit.  Issue sits under elect.sync so that ptxas keeps descriptors in
//                     uniform registers (a lane == 0 branch costs a uniform-loop per instruction).
//
// The softmax of an M tile must never wait for the tensor pipe: S_w is double buffered in TMEM and the
// MMA thread runs TWO steps ahead on the S side -- P.V_w(j) is followed at once by S_w(j+2) into the
// buffer P_w(j) just vacated -- so S_w(j+1) is already complete when softmax_w(j) ends.  The
// exponentials (MUFU, 16 / clock / SM: as many cycles per step as the two MMAs) are then the only
// serial resource of a warpgroup, and the two warpgroups share it back to back.
//
// TMEM (512 columns): S_0a S_0b S_1a S_1b (64 fp32 columns each) | O_0 | O_1 (128 each); P aliases the
// first 32 columns of its S buffer.  Shared memory: Q 2 x 32 KiB, rings of K tiles and V tiles
// (16 KiB each), all in 64-element-of-K slabs of 128-byte rows.
#pragma once

#include <type_traits>

#include "prefix_prefill.cuh"
#include "umma.cuh"

namespace b200 {

#ifdef B200_UMMA_TRACE
// Debug build only: clock64 stamps of one CTA (blockIdx 0,0,0): [role][step][stamp]; roles 0/1 = softmax warp 0 of
// M tile 0/1, 2/3 = MMA issuer of M tile 0/1.
__device__ long long g_umma_trace[4][128][8];
#define UMMA_STAMP(role, step, i)                                                                    \
  {                                                                                                  \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (step) < 128)        \
      g_umma_trace[role][step][i] = clock64();                                                       \
  }
#else
#define UMMA_STAMP(role, step, i) \
  {}
#endif

constexpr int kUmmaThreads = 512;
constexpr int kUmmaRows = 256;          // rows per CTA (two M tiles)
constexpr int kUmmaKeys = 64;           // keys per step
constexpr uint32_t kUmmaQTile = 32768;  // bytes of one 128-row x 128-d Q tile
constexpr uint32_t kUmmaKVTile = 16384; // bytes of one 64-key K (or V) tile
constexpr int kUmmaKStages = 5, kUmmaVStages = 4;
constexpr uint32_t kUmmaBarBytes = 8 * (2 * kUmmaKStages + 2 * kUmmaVStages + 4 + 4 + 2) + 32;
constexpr uint32_t kUmmaSmem = 2 * kUmmaQTile + (kUmmaKStages + kUmmaVStages) * kUmmaKVTile + kUmmaBarBytes + 1024;

template <typename T, int D, int BS, bool TMA>
__global__ void __launch_bounds__(kUmmaThreads, 1)
context_attention_umma_kernel(const __grid_constant__ PrefixParams p, const __grid_constant__ PrefixTmaMaps maps) {
  using namespace umma;
  static_assert(D == 64 || D == 128, "head size 64 or 128");
  constexpr int KT = kUmmaKeys;
  constexpr int KCH = D / 8;  // 16-byte chunks per key (K) / per 8-key chunk column (Q rows)
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
  const int n_ctx_tiles = (ctx + KT - 1) / KT;
  const int last_tok = m0 + tile_tokens - 1;
  const int nt = n_ctx_tiles + last_tok / KT + 1;

  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_s = base, k_ring = base + 2 * kUmmaQTile, v_ring = k_ring + kUmmaKStages * kUmmaKVTile,
                 bars = v_ring + kUmmaVStages * kUmmaKVTile;
  const uint32_t k_full = bars, k_empty = k_full + 8 * kUmmaKStages, v_full = k_empty + 8 * kUmmaKStages,
                 v_empty = v_full + 8 * kUmmaVStages, s_full = v_empty + 8 * kUmmaVStages,  // [w][buffer]
                 p_full = s_full + 32, o_ready = p_full + 32, slot = o_ready + 16,  // p_full: [w][buffer]
                 aux = slot + 8;  // TMA variant: landing barrier of V pages that need patching

  if (tid == 0) {
    for (int i = 0; i < kUmmaKStages; ++i) {
      mbar_init(k_full + 8 * i, TMA ? 1 : 128);
      mbar_init(k_empty + 8 * i, nq);
    }
    for (int i = 0; i < kUmmaVStages; ++i) {
      mbar_init(v_full + 8 * i, TMA ? 1 : 128);
      mbar_init(v_empty + 8 * i, nq);
    }
    mbar_init(aux, 1);
    for (int i = 0; i < 4; ++i) {
      mbar_init(s_full + 8 * i, 1);
      mbar_init(p_full + 8 * i, 128);
    }
    for (int i = 0; i < 2; ++i) mbar_init(o_ready + 8 * i, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(slot, 512);

  // ---- Q: rows (token, head) -> two K-major SWIZZLE_128B tiles (padding rows repeat the last token) ----
  {
    const T* qg = static_cast<const T*>(p.q);
#pragma unroll
    for (int i = 0; i < kUmmaRows * KCH / kUmmaThreads; ++i) {
      const int id = tid + kUmmaThreads * i;
      const int R = id / KCH, c = id % KCH;
      if (R < 128 * nq) {
        const int ti = min(R / G, tile_tokens - 1);
        const int head = kvh * G + R % G;
        const T* src = qg + (start + m0 + ti) * p.q_s0 + (int64_t)head * p.q_s1 + c * 8;
        const int r = R & 127;
        cp_async16(q_s + (uint32_t)(R >> 7) * kUmmaQTile + (uint32_t)(c >> 3) * 16384u + sw128_offset(r, c & 7),
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
    setmaxnreg_inc<192>();
  } else {
    setmaxnreg_dec<64>();
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
        const int b = j & 1;
        if ((warp & 3) == 0) UMMA_STAMP(w, j, 0)
        mbar_wait_wd(s_full + 8 * (2 * w + b), (uint32_t)((j >> 1) & 1));
        fence_after_sync();
        if ((warp & 3) == 0) UMMA_STAMP(w, j, 1)
        uint32_t s[2][32];
        tmem_ld32(t_s + 64 * b, s[0]);
        tmem_ld32(t_s + 64 * b + 32, s[1]);
        tmem_wait_ld();
        if ((warp & 3) == 0) UMMA_STAMP(w, j, 2)
        // visible keys of this row in the tile: a prefix [0, lim)
        int lim;
        if (j < n_ctx_tiles)
          lim = min(KT, ctx - KT * j);
        else
          lim = min(KT, max(0, iq - KT * (j - n_ctx_tiles) + 1));
        if (__any_sync(FULL, lim < KT)) {
#pragma unroll
          for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (32 * c + i >= lim) s[c][i] = __float_as_uint(kNInf);
        }
        float mx0 = kNInf, mx1 = kNInf, mx2 = kNInf, mx3 = kNInf;
#pragma unroll
        for (int c = 0; c < 2; ++c)
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
          // O_w must hold every product up to step j - 1 before it is scaled
          mbar_wait_wd(o_ready + 8 * w, (uint32_t)((j - 1) & 1));
          fence_after_sync();
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            tmem_ld32(t_o + 32 * c, o);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(t_o + 32 * c, o);
          }
        }
        const float neg_m = m_ref == kNInf ? 0.f : -m_ref;
        if ((warp & 3) == 0) UMMA_STAMP(w, j, 3)
        float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
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
            pk[16 * c + i / 2] = MmaOp<T>::pack(e0, e1);
            pk[16 * c + i / 2 + 1] = MmaOp<T>::pack(e2, e3);
          }
        }
        if ((warp & 3) == 0) UMMA_STAMP(w, j, 4)
        tmem_st32(t_s + 64 * b, pk);  // P over the first 32 columns of this S buffer (S is in registers)
        l += (l0 + l1) + (l2 + l3);
        tmem_wait_st();
        fence_before_sync();
        if ((warp & 3) == 0) UMMA_STAMP(w, j, 5)
        // per S buffer: the softmax may finish step j + 1 before the MMA thread has looked at step j
        mbar_arrive(p_full + 8 * (2 * w + b));
      }
      // ---- epilogue: O / l -> out ----
      // o_ready has completed nt - 2 phases for certain (S(nt-1) was issued behind P.V(nt-3)); walk the
      // last two one by one -- a parity wait can only tell "this phase" from "the next"
      if (nt >= 2) mbar_wait_wd(o_ready + 8 * w, (uint32_t)((nt - 2) & 1));
      mbar_wait_wd(o_ready + 8 * w, (uint32_t)((nt - 1) & 1));
      fence_after_sync();
      const float inv = l > 0.f ? 1.f / l : 0.f;
      T* orow = static_cast<T*>(p.out) + (start + m0 + min(ti, tile_tokens - 1)) * p.o_s0 +
                (int64_t)(kvh * G + R % G) * p.o_s1;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
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
  } else if (TMA && warp < 12) {
    // ======================= TMA producer (warp 8): pages -> operand tiles, no per-thread copies =======================
    // BS = 16.  A K page [D/8][16][8] is fetched as two boxes of 8 tokens x D: each lands as [chunk][8 tokens][16 B],
    // the canonical NO-swizzle K-major layout (core matrix = 8 keys x 16 bytes; next chunk +128 B, next 8 keys
    // +16 D bytes).  A V page [D][16] is fetched as ONE box with SWIZZLE_32B: the canonical 32-byte-swizzle K-major
    // layout of V^T for one 16-key k-step.  Twelve tensor copies replace the 4096 16-byte LDGSTS of a tile, whose
    // time in flight backed up the MIO queue the softmax warps' MUFU go through (910 -> ~1040 TFLOP/s).
    // Lanes 0-3 each own one page of the tile.  Keys past the end of the context / the new tokens: K may hold
    // anything (S is masked before the row maximum), V must be finite there because 0 * NaN is not 0 -- such tiles
    // (at most two per CTA) take the slow road: land on `aux`, clear by hand, then signal.
    if constexpr (TMA) {
      if (warp == 8) {
        static_assert(!TMA || BS == 16, "one page = one 16-key k-step");
        constexpr uint32_t KGRP = (uint32_t)D * 16u;       // bytes of 8 keys of K
        constexpr uint32_t VPAGE = (uint32_t)D * BS * 2u;  // bytes of one V page tile
        const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
        const int slice0 = (int)(start / BS) + seq;  // first staged slice of this sequence (guarded above)
        const bool owner = lane < KT / BS;
        int ks = 0, kround = 0, vs = 0, vround = 0;
        uint32_t aux_phase = 0;
        if (lane == 0) {
          tma_prefetch_desc(maps.k);
          tma_prefetch_desc(maps.v);
        }
        // (valid keys, block id / staged slice) of this lane's page of tile jj.  The block-table entry of tile j + 1 is
        // requested while tile j is being issued: one warp runs this loop, and a dependent global load per tile
        // would sit in a chain that has to stay ahead of the tensor pipe.
        auto lookup = [&](int jj, int& nv, int& id) {
          const bool in_ctx = jj < n_ctx_tiles;
          const int pos = KT * (in_ctx ? jj : jj - n_ctx_tiles) + BS * lane;  // first key of my page
          nv = (owner && jj < nt) ? min(BS, max(0, (in_ctx ? ctx : qlen) - pos)) : 0;
          id = 0;
          if (nv > 0) id = in_ctx ? bt[pos / BS] : slice0 + pos / BS;
        };
        int nval, c3, nval_next, c3_next;
        lookup(0, nval, c3);
        for (int j = 0; j < nt; ++j, nval = nval_next, c3 = c3_next) {
          lookup(j + 1, nval_next, c3_next);
          const bool is_ctx = j < n_ctx_tiles;
          // first tile of new tokens: the staging pre-pass (prefix_stage_kv_kernel, this grid's programmatic
          // predecessor) must have completed before its pages are fetched
          if (j == n_ctx_tiles) asm volatile("griddepcontrol.wait;" ::: "memory");
          const void* kmap = is_ctx ? (const void*)maps.k : (const void*)maps.sk;
          const void* vmap = is_ctx ? (const void*)maps.v : (const void*)maps.sv;
          const uint32_t have = __ballot_sync(FULL, nval > 0);
          const bool special = __any_sync(FULL, owner && nval < BS);
          const uint32_t npages = (uint32_t)__popc(have);
          {
            if (kround > 0) mbar_wait_wd(k_empty + 8 * ks, (uint32_t)((kround - 1) & 1));
            const uint32_t dst = k_ring + (uint32_t)ks * kUmmaKVTile + (uint32_t)(2 * lane) * KGRP;
            if (lane == 0) mbar_arrive_expect_tx(k_full + 8 * ks, npages * 2u * KGRP);
            __syncwarp();
            if (nval > 0) {
              tma_load_4d(dst, kmap, 0, 0, kvh, c3, k_full + 8 * ks);
              tma_load_4d(dst + KGRP, kmap, 64, 0, kvh, c3, k_full + 8 * ks);
            }
            if (++ks == kUmmaKStages) {
              ks = 0;
              ++kround;
            }
          }
          {
            if (vround > 0) mbar_wait_wd(v_empty + 8 * vs, (uint32_t)((vround - 1) & 1));
            const uint32_t tile = v_ring + (uint32_t)vs * kUmmaKVTile;
            if (!special) {
              if (lane == 0) mbar_arrive_expect_tx(v_full + 8 * vs, npages * VPAGE);
              __syncwarp();
              if (nval > 0) tma_load_4d(tile + (uint32_t)lane * VPAGE, vmap, 0, 0, kvh, c3, v_full + 8 * vs);
            } else {
              if (npages > 0) {
                if (lane == 0) mbar_arrive_expect_tx(aux, npages * VPAGE);
                __syncwarp();
                if (nval > 0) tma_load_4d(tile + (uint32_t)lane * VPAGE, vmap, 0, 0, kvh, c3, aux);
                mbar_wait_wd(aux, aux_phase);
                aux_phase ^= 1u;
              }
              for (int pg = 0; pg < KT / BS; ++pg) {
                const int nv = __shfl_sync(FULL, nval, pg);
                const uint32_t page = tile + (uint32_t)pg * VPAGE;
                if (nv == 0) {  // no such page: whatever the slot held (nothing yet, perhaps) must not reach P.V
                  for (uint32_t o = (uint32_t)lane * 16u; o < VPAGE; o += 512u)
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(page + o), "r"(0u) : "memory");
                } else if (nv < BS && is_ctx) {
                  // last context page: slots past ctx_len may hold anything; row d is 32 bytes whose 16-byte halves
                  // are swapped when (d >> 2) & 1 (SWIZZLE_32B).  Staged new-token pages are zero filled already.
                  for (int d = lane; d < D; d += 32) {
                    const uint32_t row = page + (uint32_t)d * 32u;
                    const int sw = (d >> 2) & 1;
#pragma unroll
                    for (int ph = 0; ph < 2; ++ph) {
                      const int t0 = 8 * (ph ^ sw);  // first token of this physical half
                      uint32_t x[4];
                      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                                   : "=r"(x[0]), "=r"(x[1]), "=r"(x[2]), "=r"(x[3])
                                   : "r"(row + 16u * ph));
#pragma unroll
                      for (int wd = 0; wd < 4; ++wd)
                        x[wd] &= (t0 + 2 * wd < nv ? 0x0000ffffu : 0u) | (t0 + 2 * wd + 1 < nv ? 0xffff0000u : 0u);
                      asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(row + 16u * ph), "r"(x[0]), "r"(x[1]),
                                   "r"(x[2]), "r"(x[3])
                                   : "memory");
                    }
                  }
                }
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) mbar_arrive(v_full + 8 * vs);
            }
            if (++vs == kUmmaVStages) {
              vs = 0;
              ++vround;
            }
          }
        }
      }
    }
  } else if (warp < 12) {
    // ======================= gather warps: pages -> swizzled operand tiles =======================
    // 16 groups of 8 threads: group g serves the 8 keys [8 (g & 7), +8) of the tile; for K it moves half of
    // the d chunks of its key (lane = key; half g >> 3), for V the rows d = (D/2) (g >> 3) + 8 i + lane of
    // its 8-key chunk.  Every thread therefore touches one page per tile.
    const int pt = tid - 256;
    const int g = pt >> 3, l8 = pt & 7, kg = g & 7, hi = g >> 3;
    const unsigned char* kbase = static_cast<const unsigned char*>(p.k_cache) + (int64_t)kvh * p.kc_hs * 2;
    const unsigned char* vbase = static_cast<const unsigned char*>(p.v_cache) + (int64_t)kvh * p.vc_hs * 2;
    const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
    const unsigned char* sbase =
        p.kv_scratch + (((int64_t)(start / BS) + seq) * p.num_kv_heads + kvh) * SLICE_BYTES;
    const int64_t slice_stride = (int64_t)p.num_kv_heads * SLICE_BYTES;
    const uint32_t k_off = (uint32_t)(8 * kg + l8) * 128u;
    const uint32_t v_off = (uint32_t)((D / 2) * hi + l8) * 128u + (uint32_t)((kg ^ l8) << 4);
    int ks = 0, kround = 0, vs = 0, vround = 0;  // ring positions of tile j, times each ring wrapped
    for (int j = 0; j < nt; ++j) {
      const bool is_ctx = j < n_ctx_tiles;
      if (j == n_ctx_tiles) asm volatile("griddepcontrol.wait;" ::: "memory");  // staged pages: see the TMA producer
      const int pos = KT * (is_ctx ? j : j - n_ctx_tiles) + 8 * kg;  // first key of the group
      const int nval = min(8, max(0, (is_ctx ? ctx : qlen) - pos));  // its valid keys
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
        const uint32_t dst = k_ring + (uint32_t)ks * kUmmaKVTile + k_off;
#pragma unroll
        for (int i = 0; i < KCH / 2; ++i) {
          const int c = (KCH / 2) * hi + i;  // d chunk: slab c / 8, swizzled position (c % 8) ^ (row % 8)
          cp_async16(dst + (uint32_t)(c >> 3) * 8192u + (uint32_t)(((c & 7) ^ l8) << 4), src + c * BS * 16, nb);
        }
        cp_async_arrive_noinc(k_full + 8 * ks);
        if (++ks == kUmmaKStages) {
          ks = 0;
          ++kround;
        }
      }
      {
        if (vround > 0) mbar_wait_wd(v_empty + 8 * vs, (uint32_t)((vround - 1) & 1));
        const unsigned char* src = vpage + toff * 2 + ((D / 2) * hi + l8) * BS * 2;
        const uint32_t nb = nval > 0 ? 16u : 0u;
        const uint32_t dst = v_ring + (uint32_t)vs * kUmmaKVTile + v_off;
#pragma unroll
        for (int i = 0; i < D / 16; ++i) cp_async16(dst + (uint32_t)i * 1024u, src + i * 8 * BS * 2, nb);
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
          for (int i = 0; i < D / 16; ++i) {
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
        if (++vs == kUmmaVStages) {
          vs = 0;
          ++vround;
        }
      }
    }
    cp_async_commit();
    cp_async_wait_all();  // nothing of this thread is still in flight when the CTA winds down
  } else if (warp < 12 + nq) {
    // ======================= MMA issuers: warp 12 for M tile 0, warp 13 for M tile 1 =======================
    // Two independent instruction streams (disjoint TMEM, shared read-only K/V tiles): while one thread
    // polls its barriers the other one's MMAs keep the tensor pipe fed.  A K/V ring slot is free once
    // BOTH have committed past it (empty barriers count nq arrivals).
    const int w = warp - 12;
    constexpr uint32_t idesc_s = idesc_f16(FMT, 128, KT);
    constexpr uint32_t idesc_o = idesc_f16(FMT, 128, D);
    const uint32_t t_sw = tmem + (uint32_t)w * 128u, t_ow = tmem + 256u + (uint32_t)w * 128u;
    const uint64_t da = sw128_desc(q_s + (uint32_t)w * kUmmaQTile);
    // S_w(step in K ring slot ks) = Q_w K^T into S buffer b, then "S_w[b] ready" and "done with the slot"
    auto issue_qk = [&](int b, int ks) {
      const uint32_t k_tile = k_ring + (uint32_t)ks * kUmmaKVTile;
      // gathered tile: SWIZZLE_128B slabs; TMA tile: [8 keys][chunk][8 keys x 16 B], no swizzle (k-step = 2 chunks)
      const uint64_t db = TMA ? desc_kmajor(k_tile, 128u, (uint32_t)D * 16u, 0u) : sw128_desc(k_tile);
#pragma unroll
      for (int kk = 0; kk < D / 16; ++kk) {
        const uint32_t in_slab = (uint32_t)(kk & 3) * 32u;
        mma_ss(t_sw + (uint32_t)b * 64u, desc_advance(da, (uint32_t)(kk >> 2) * 16384u + in_slab),
               desc_advance(db, TMA ? (uint32_t)kk * 256u : (uint32_t)(kk >> 2) * 8192u + in_slab), idesc_s, kk > 0);
      }
      commit(s_full + 8 * (2 * w + b));
      commit(k_empty + 8 * ks);
    };
    // Prologue: S(0) and S(1).  The gather wrote the tiles through the generic proxy: order those writes
    // (acquired through the mbarrier) before this thread's async-proxy reads.
    int ks = 0, kround = 0;  // K ring position of the next step whose S has not been issued
    const int pre = min(2, nt);
    for (int jj = 0; jj < pre; ++jj) {
      mbar_wait_wd(k_full + 8 * ks, 0u);
      fence_proxy_async_smem();
      fence_after_sync();
      if (elect_one()) issue_qk(jj, ks);
      __syncwarp();
      ++ks;
    }
    int vs = 0, vround = 0;
    for (int j = 0; j < nt; ++j) {
      const bool more = j + 2 < nt;
      UMMA_STAMP(2 + w, j, 0)
      mbar_wait_wd(p_full + 8 * (2 * w + (j & 1)), (uint32_t)((j >> 1) & 1));
      UMMA_STAMP(2 + w, j, 1)
      mbar_wait_wd(v_full + 8 * vs, (uint32_t)(vround & 1));
      UMMA_STAMP(2 + w, j, 2)
      if (more) mbar_wait_wd(k_full + 8 * ks, (uint32_t)(kround & 1));
      UMMA_STAMP(2 + w, j, 3)
      fence_proxy_async_smem();
      fence_after_sync();
      if (elect_one()) {
        const uint32_t v_tile = v_ring + (uint32_t)vs * kUmmaKVTile;
        // gathered tile: V^T rows of 64 keys, SWIZZLE_128B; TMA tile: one SWIZZLE_32B page (16 keys) per k-step
        const uint64_t db = TMA ? desc_kmajor(v_tile, 16u, 256u, 6u) : sw128_desc(v_tile);
        const uint32_t t_p = t_sw + (uint32_t)(j & 1) * 64u;
#pragma unroll
        for (int kk = 0; kk < KT / 16; ++kk)
          mma_ts(t_ow, t_p + (uint32_t)kk * 8u, desc_advance(db, TMA ? (uint32_t)kk * (uint32_t)(D * BS * 2) : (uint32_t)kk * 32u),
                 idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
        commit(o_ready + 8 * w);
        commit(v_empty + 8 * vs);
        if (more) issue_qk(j & 1, ks);  // overwrites P_w(j): ordered behind P.V_w(j) in the tensor pipe
      }
      __syncwarp();
      UMMA_STAMP(2 + w, j, 4)
      if (more && ++ks == kUmmaKStages) {
        ks = 0;
        ++kround;
      }
      if (++vs == kUmmaVStages) {
        vs = 0;
        ++vround;
      }
    }
  }

  fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    fence_after_sync();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace b200
