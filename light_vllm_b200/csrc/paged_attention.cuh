// Paged-attention decode kernel for sm_100a — replaces the reference's
// paged_attention_kernel / _v1_kernel / _v2_kernel (csrc/attention/attention_kernels.cu:86-562).
//
// Design (B200-first, not a translation of the reference's SIMT kernel):
//   * one CTA per (sequence, KV head, context split); ALL q-heads of the GQA group are computed
//     from ONE read of each KV page (the reference launches one CTA per q-head and re-reads the
//     page H/KH times, attention_kernels.cu:150-153,734).
//   * every warp streams its own interleaved subset of the split's pages through a private
//     multi-stage shared-memory ring.  A K page and a V page are each one contiguous
//     head_size*block_size run in the reference layout, so each is fetched with ONE
//     cp.async.bulk (TMA engine, SASS UBLKCP) that signals an mbarrier with complete_tx.
//     No register staging, no per-thread address math, no CTA-wide barrier in the main loop.
//   * QK^T and PV run on tensor cores (mma.sync m16n8k16, fp32 accumulate) with the q-head group
//     as the (padded) M dimension: A = Q[16 x D] from registers, B = K page via ldmatrix straight
//     from the reference layout [D/8][BS][8]; P stays in registers and becomes the A operand of
//     PV, B = V page [D][BS] via ldmatrix.  At HBM speed the CUDA-core formulation would need
//     35-100+% of the FP32 issue rate; this needs ~32 HMMA + ~16 LDSM per page.
//   * streaming (online) softmax per warp in the exp2 domain; warps merge through shared memory;
//     context splits merge either through distributed shared memory inside a thread-block
//     cluster (v1: no scratch in the op schema) or through the caller's scratch + the reduce
//     kernel (v2, 512-token partitions as in attention_kernels.cu:850,888-897).
#pragma once

#include "common.cuh"

namespace b200 {

struct AttnParams {
  void* out;               // v1: [B, H, D] T
  float* exp_sums;         // v2: [B, H, P]
  float* max_logits;       // v2: [B, H, P]
  void* tmp_out;           // v2: [B, H, P, D] T
  const void* q;           // [B, H, D] T, row stride q_stride
  const void* k_cache;     // [NB, KH, D/x, BS, x]
  const void* v_cache;     // [NB, KH, D, BS]
  const int32_t* block_tables;  // [B, max_blocks_per_seq]
  const int32_t* seq_lens;      // [B]
  const float* alibi_slopes;    // [H] or null
  int num_heads, num_kv_heads, group;  // group = H / KH
  int gchunks;                  // ceil(group / rows_per_cta)
  int rows_per_cta;             // q-heads of the group one CTA computes (<= GROWS; fewer = "head split"
                                // for grids too small to fill the machine: more CTAs share each page via L2)
  int max_blocks_per_seq;
  int max_num_partitions;       // v2 only
  int partition_pages;          // v2: pages per 512-token partition; v1: 0
  int num_stages;               // ring depth per warp
  int64_t q_stride, kv_block_stride, kv_head_stride;  // elements
  float scale_log2;             // scale * log2(e) (* k_scale for fp8 caches)
  float out_scale;              // v_scale for fp8 caches, else 1
  int bs_enabled, bs_tp_rank, bs_local_blocks, bs_vert_stride, bs_block_size,
      bs_head_sliding_step;
};

// Dispatch table entry exported by each instantiation TU (paged_attention_inst_*.cu).
using AttnLaunchFn = int (*)(const AttnParams&, dim3, int, int, int, int, cudaStream_t);
using AttnSmemFn = size_t (*)(int, int);
using AttnOccFn = int (*)(int, int);
struct AttnEntry {
  AttnLaunchFn launch;     // (params, grid, cluster, nwarps, nstages, device, stream)
  AttnSmemFn smem;         // (nwarps, nstages) -> dynamic shared memory bytes
  AttnOccFn occupancy;     // (nwarps, nstages) -> resident CTAs per SM
};

constexpr float kNegInf = -__builtin_huge_valf();
constexpr int kMergePad = 8;  // floats of padding per merged row (bank spread)

template <int KV>
struct KvTraits {
  static constexpr int ES = (KV == B200_KV_AUTO) ? 2 : 1;  // bytes per cache element
};

// MMA element type: the cache's own 16-bit type for "auto"; fp16 for fp8 caches (e4m3/e5m2 -> f16
// is exact, and q is converted to f16 once per CTA).
template <typename T, int KV>
struct MmaType {
  using type = T;
};
template <typename T>
struct MmaType<T, B200_KV_FP8_E4M3> {
  using type = __half;
};
template <typename T>
struct MmaType<T, B200_KV_FP8_E5M2> {
  using type = __half;
};

template <int KV>
__device__ __forceinline__ uint32_t fp8x2_to_f16x2(uint16_t v) {
  uint32_t r;
  if constexpr (KV == B200_KV_FP8_E5M2)
    asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(r) : "h"(v));
  else
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"(v));
  return r;
}
// Four fp8 bytes -> two f16x2 registers (bytes 0,1 -> lo; bytes 2,3 -> hi).  Written so that ptxas
// feeds the two halves of `w` to F2FP through its .H0/.H1 operand selectors: no PRMT/unpack
// instructions (the cuda_fp8.h struct round trip cost ~3 PRMT per conversion, see
// profiles/r1_c4_fp8_before.md).
template <int KV>
__device__ __forceinline__ void fp8x4_to_f16x2x2(uint32_t w, uint32_t& lo, uint32_t& hi) {
  if constexpr (KV == B200_KV_FP8_E5M2)
    asm("{\n.reg .b16 a, b;\nmov.b32 {a, b}, %2;\ncvt.rn.f16x2.e5m2x2 %0, a;\n"
        "cvt.rn.f16x2.e5m2x2 %1, b;\n}" : "=r"(lo), "=r"(hi) : "r"(w));
  else
    asm("{\n.reg .b16 a, b;\nmov.b32 {a, b}, %2;\ncvt.rn.f16x2.e4m3x2 %0, a;\n"
        "cvt.rn.f16x2.e4m3x2 %1, b;\n}" : "=r"(lo), "=r"(hi) : "r"(w));
}

// Token (within the page) of the score this thread holds in S-accumulator n-tile `nt`, column pair
// of lane-quad index `qq`, element `e`; equally the token whose K row the lane with column
// c = 2*qq + e supplies to that n-tile.
//   16-bit caches / 8-token pages: the natural order, token = 8*nt + c.
//   fp8 caches with >= 16-token pages: within each 16-token group the columns are permuted so that
//   a thread's four scores (tiles 2kk, 2kk+1) are the four CONSECUTIVE tokens 4*qq..4*qq+3.  P.V
//   then needs V[d][4*qq..4*qq+3] = one 32-bit shared load per d-tile instead of two 16-bit ones,
//   and both K-row sets {0,1,4,5,10,11,14,15} / {2,3,6,7,8,9,12,13} still cover all 8 16-byte
//   bank groups, so the K loads stay conflict-free.
template <int KV, int BS, bool PERM16 = false>
__device__ __forceinline__ int page_token(int nt, int qq, int e) {
  if constexpr ((KV != B200_KV_AUTO || PERM16) && BS >= 16)
    return (nt >> 1) * 16 + 4 * qq + 2 * ((nt & 1) ^ (qq >> 1)) + e;
  else
    return nt * 8 + 2 * qq + e;
}

// One KV page (BS tokens) against the warp's q-group: S = QK^T, online softmax, O += P V.
// PLAIN: no ALiBi and no blocksparse mask (the common case) -- the logit is just s * scale.
// ROWB (prefix prefill): every row has its own window of visible tokens [row_lo, row_hi) inside the
// page (causal frontier / sliding window) and its own ALiBi origin row_rel0 = (first token of the
// page) - (the row's query position); decode rows all share `valid` and `tok_rel0`.
// PFX (prefill kernel, 16-bit caches, >= 16-token pages): the compute-bound variant.  Tokens take the
// permuted order of page_token<.., true> so that P.V reads V[d][4*q4..4*q4+3] with ONE conflict-free
// 64-bit shared load per d-tile (ldmatrix over the 32-byte V rows is 2-way bank conflicted), and
// Q.K^T accumulates even / odd k-steps in separate chains (half the dependent-HMMA latency).
// PGS (PFX only): pages consumed per step.  PGS = 2 treats two 16-token pages, `pg_stride` bytes
// apart in shared memory, as one 32-token step: twice the independent HMMA work per softmax round.
template <typename T, int D, int BS, int KV, int GROWS, bool TAIL, bool PLAIN, bool ROWB,
          bool PFX = false, int PGS = 1>
__device__ __forceinline__ void process_page_ex(uint32_t k_smem, uint32_t v_smem,
                                                const uint32_t (&qa)[(D + 15) / 16][GROWS / 4],
                                                float (&o)[D / 8][GROWS / 4], float (&m)[GROWS / 8],
                                                float (&l)[GROWS / 8], int lane, int valid,
                                                float scale_log2, const float (&slope)[GROWS / 8],
                                                float tok_rel0,  // (first token of page) - seq_len + 1
                                                const bool (&row_on)[GROWS / 8],
                                                const int (&row_lo)[GROWS / 8],
                                                const int (&row_hi)[GROWS / 8],
                                                const float (&row_rel0)[GROWS / 8],
                                                uint32_t pg_stride = 0) {
  using MT = typename MmaType<T, KV>::type;
  static_assert(PGS == 1 || (PFX && KV == B200_KV_AUTO && BS >= 16), "multi-page steps: PFX only");
  constexpr int ES = KvTraits<KV>::ES;
  constexpr int KS = (D + 15) / 16;  // k-steps over head dim
  constexpr int TB = BS * PGS;       // tokens per step
  constexpr int GPP = BS >= 16 ? BS / 16 : 1;  // 16-token groups per page
  constexpr int NT = TB / 8;         // 8-token n-tiles in a step
  constexpr int RS = GROWS / 8;      // row sets (rows g, g+8)
  constexpr int NDT = D / 8;         // 8-wide d tiles of the output
  const int q4 = lane & 3;
  const int mi = lane >> 3, r8 = lane & 7;

  // ---------------- S = Q K^T ----------------
  float s[NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;

  if constexpr (KV == B200_KV_AUTO) {
    // K page layout [D/8][BS][8] (16-byte rows).  ldmatrix row addresses for this lane.
    if constexpr (BS >= 16) {
      const int krow = PFX ? page_token<KV, BS, true>(mi >> 1, r8 >> 1, r8 & 1) : (mi >> 1) * 8 + r8;
      const uint32_t base = k_smem + (uint32_t)((mi & 1) * BS * 16 + krow * 16);
      constexpr bool SPLIT = PFX && PGS == 1;  // two accumulation chains per n-tile
      float s2[SPLIT ? NT : 1][4];
      if constexpr (SPLIT) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) s2[nt][0] = s2[nt][1] = s2[nt][2] = s2[nt][3] = 0.f;
      }
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
        for (int th = 0; th < NT / 2; ++th) {  // 16 tokens per ldmatrix.x4
          uint32_t b00, b01, b10, b11;
          ldsm_x4(base + (uint32_t)(ks * 2 * BS * 16 + (th % GPP) * 16 * 16) + (th / GPP) * pg_stride,
                  b00, b01, b10, b11);
          if (2 * ks + 1 >= NDT) b01 = b11 = 0u;  // odd chunk count (head 120): no such chunk
          const uint32_t a1 = (RS == 2) ? qa[ks][2] : 0u, a3 = (RS == 2) ? qa[ks][3] : 0u;
          if (SPLIT && (ks & 1)) {
            MmaOp<MT>::run(s2[SPLIT ? 2 * th : 0], qa[ks][0], a1, qa[ks][1], a3, b00, b01);
            MmaOp<MT>::run(s2[SPLIT ? 2 * th + 1 : 0], qa[ks][0], a1, qa[ks][1], a3, b10, b11);
          } else {
            MmaOp<MT>::run(s[2 * th], qa[ks][0], a1, qa[ks][1], a3, b00, b01);
            MmaOp<MT>::run(s[2 * th + 1], qa[ks][0], a1, qa[ks][1], a3, b10, b11);
          }
        }
      }
      if constexpr (SPLIT) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e) s[nt][e] += s2[nt][e];
      }
    } else {  // BS == 8: one n-tile, ldmatrix.x2 (lanes 0-15 supply addresses)
      const uint32_t base = k_smem + (uint32_t)((mi & 1) * BS * 16 + r8 * 16);
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t b0, b1;
        ldsm_x2(base + (uint32_t)(ks * 2 * BS * 16), b0, b1);
        if (2 * ks + 1 >= NDT) b1 = 0u;
        const uint32_t a1 = (RS == 2) ? qa[ks][2] : 0u, a3 = (RS == 2) ? qa[ks][3] : 0u;
        MmaOp<MT>::run(s[0], qa[ks][0], a1, qa[ks][1], a3, b0, b1);
      }
    }
  } else {
    // fp8 K page layout [D/16][BS][16].  Lane (column c = lane/4, q4) reads 4 bytes:
    // d = 16ks + 4*q4 + 0..3 of token page_token(nt, c/2, c%2); (d0,d1) -> b0, (d2,d3) -> b1.
    // Q fragments are loaded with the matching d permutation (see load_q).
    const int c = lane >> 2;
    uint32_t koff[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
      koff[nt] = k_smem + (uint32_t)(page_token<KV, BS>(nt, c >> 1, c & 1) * 16 + q4 * 4);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        uint32_t w, b0, b1;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(koff[nt] + (uint32_t)(ks * BS * 16)));
        fp8x4_to_f16x2x2<KV>(w, b0, b1);
        const uint32_t a1 = (RS == 2) ? qa[ks][2] : 0u, a3 = (RS == 2) ? qa[ks][3] : 0u;
        MmaOp<MT>::run(s[nt], qa[ks][0], a1, qa[ks][1], a3, b0, b1);
      }
    }
  }

  // ---------------- online softmax (exp2 domain) ----------------
  float alpha[RS];
  bool changed = false;
  // FASTSM (prefill fast path, no mask / bias): max over the raw scores, then one FFMA + ex2 per
  // score (scale > 0, so max(s * scale) = max(s) * scale).
  constexpr bool FASTSM = PFX && PLAIN && !TAIL && !ROWB;
#pragma unroll
  for (int rs = 0; rs < RS; ++rs) {
    if constexpr (FASTSM) {
      float mx = kNegInf;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) mx = fmaxf(mx, fmaxf(s[nt][2 * rs], s[nt][2 * rs + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      // Lazy reference point: m only has to keep exp2(x - m) in range, so it moves only when the
      // step's maximum exceeds it by more than 2^8 (p <= 256 fits fp16/bf16; o and l are fp32);
      // the accumulator rescale below then runs a handful of times per row instead of per new max.
      const float m_cand = mx * scale_log2;
      const float m_new = (m_cand > m[rs] + 8.f) ? m_cand : m[rs];
      alpha[rs] = fast_exp2(m[rs] - m_new);  // m_new is finite here: nothing is masked
      changed |= (m_new != m[rs]);
      m[rs] = m_new;
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float pv = fast_exp2(fmaf(s[nt][2 * rs + e], scale_log2, -m_new));
          s[nt][2 * rs + e] = pv;
          sum += pv;
        }
      }
      l[rs] = l[rs] * alpha[rs] + sum;
      continue;
    }
    float mx = kNegInf;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int t = page_token<KV, TB, PFX>(nt, q4, e);  // token within the step
        float x = s[nt][2 * rs + e] * scale_log2;
        if constexpr (!PLAIN) x += slope[rs] * ((ROWB ? row_rel0[rs] : tok_rel0) + (float)t);
        if (TAIL && t >= valid) x = kNegInf;
        if constexpr (ROWB) {
          if (t < row_lo[rs] || t >= row_hi[rs]) x = kNegInf;
        } else if constexpr (!PLAIN) {
          if (!row_on[rs]) x = kNegInf;
        }
        s[nt][2 * rs + e] = x;
        mx = fmaxf(mx, x);
      }
    }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m[rs], mx);
    const float m_safe = (m_new == kNegInf) ? 0.f : m_new;
    alpha[rs] = fast_exp2(m[rs] - m_safe);
    changed |= (m_new != m[rs]);
    m[rs] = m_new;
    float sum = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float pv = fast_exp2(s[nt][2 * rs + e] - m_safe);
        s[nt][2 * rs + e] = pv;
        sum += pv;
      }
    }
    l[rs] = l[rs] * alpha[rs] + sum;
  }
  if (__any_sync(0xffffffffu, changed)) {  // warp-uniform: the running max rarely moves
#pragma unroll
    for (int j = 0; j < NDT; ++j) {
#pragma unroll
      for (int rs = 0; rs < RS; ++rs) {
        o[j][2 * rs] *= alpha[rs];
        o[j][2 * rs + 1] *= alpha[rs];
      }
    }
  }

  // ---------------- O += P V ----------------
  // tail masks: V columns past the sequence end may hold NaN; 0 * NaN must not reach the MMA.
  uint32_t vmask[TB >= 16 ? TB / 8 : 1];
  if constexpr (TAIL) {
#pragma unroll
    for (int h = 0; h < (TB >= 16 ? TB / 8 : 1); ++h) {
      const int t0 = page_token<KV, TB, PFX>(h, q4, 0);
      vmask[h] = (t0 < valid ? 0x0000ffffu : 0u) | (t0 + 1 < valid ? 0xffff0000u : 0u);
    }
  }
  constexpr int KSV = (TB >= 16) ? TB / 16 : 1;  // k-steps over tokens
#pragma unroll
  for (int kk = 0; kk < KSV; ++kk) {
    uint32_t pa0, pa1, pa2, pa3;
    if constexpr (BS >= 16) {
      pa0 = MmaOp<MT>::pack(s[2 * kk][0], s[2 * kk][1]);
      pa2 = MmaOp<MT>::pack(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa1 = (RS == 2) ? MmaOp<MT>::pack(s[2 * kk][2], s[2 * kk][3]) : 0u;
      pa3 = (RS == 2) ? MmaOp<MT>::pack(s[2 * kk + 1][2], s[2 * kk + 1][3]) : 0u;
    } else {
      pa0 = MmaOp<MT>::pack(s[0][0], s[0][1]);
      pa1 = (RS == 2) ? MmaOp<MT>::pack(s[0][2], s[0][3]) : 0u;
      pa2 = pa3 = 0u;
    }
    if constexpr (KV == B200_KV_AUTO && PFX && BS >= 16) {
      // For q4 >= 2 tile 2kk holds the UPPER token pair of this lane's four tokens.  The k-slot a
      // fragment sits in is only a label shared by all lanes with the same q4, so the two P
      // fragments are swapped (4 selects) instead of the V words (2 selects per d-tile).
      if (q4 & 2) {
        uint32_t tmp = pa0; pa0 = pa2; pa2 = tmp;
        tmp = pa1; pa1 = pa3; pa3 = tmp;
      }
      uint32_t vm0 = 0xffffffffu, vm1 = 0xffffffffu;
      if constexpr (TAIL) {
        const int t0 = kk * 16 + 4 * q4;
        vm0 = (t0 < valid ? 0x0000ffffu : 0u) | (t0 + 1 < valid ? 0xffff0000u : 0u);
        vm1 = (t0 + 2 < valid ? 0x0000ffffu : 0u) | (t0 + 3 < valid ? 0xffff0000u : 0u);
      }
      // V page [D][BS] 16-bit: this lane's P fragments cover tokens 4*q4..4*q4+3 of the 16-token
      // group (page_token<.., true>): one 64-bit load per d-tile, a warp reads 256 contiguous bytes.
      const uint32_t base = v_smem + (uint32_t)((lane >> 2) * BS * 2 + q4 * 8 + (kk % GPP) * 32) +
                            (kk / GPP) * pg_stride;
#pragma unroll
      for (int j = 0; j < NDT; ++j) {
        uint32_t w0, w1;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(base + (uint32_t)(j * 8 * BS * 2)));
        uint32_t b0 = w0, b1 = w1;
        if constexpr (TAIL) {
          b0 &= vm0;
          b1 &= vm1;
        }
        float acc[4] = {o[j][0], o[j][1], RS == 2 ? o[j][2 * (RS - 1)] : 0.f,
                        RS == 2 ? o[j][2 * (RS - 1) + 1] : 0.f};
        MmaOp<MT>::run(acc, pa0, pa1, pa2, pa3, b0, b1);
        o[j][0] = acc[0]; o[j][1] = acc[1];
        if constexpr (RS == 2) { o[j][2] = acc[2]; o[j][3] = acc[3]; }
      }
    } else if constexpr (KV == B200_KV_AUTO) {
      if constexpr (BS >= 16) {
        // V page [D][BS]: 8x8 matrices = 8 d-rows x 8 tokens (16 B); x4 = 2 d-tiles x 2 token halves
        const uint32_t base = v_smem + (uint32_t)((((mi >> 1) * 8 + r8) * BS * 2) + (mi & 1) * 16 +
                                                  kk * 32);
#pragma unroll
        for (int jj = 0; jj < (NDT + 1) / 2; ++jj) {
          uint32_t b00, b01, b10, b11;
          ldsm_x4(base + (uint32_t)(jj * 16 * BS * 2), b00, b01, b10, b11);
          if constexpr (TAIL) {
            b00 &= vmask[2 * kk]; b10 &= vmask[2 * kk];
            b01 &= vmask[2 * kk + 1]; b11 &= vmask[2 * kk + 1];
          }
          {
            float acc[4] = {o[2 * jj][0], o[2 * jj][1], RS == 2 ? o[2 * jj][2 * (RS - 1)] : 0.f,
                            RS == 2 ? o[2 * jj][2 * (RS - 1) + 1] : 0.f};
            MmaOp<MT>::run(acc, pa0, pa1, pa2, pa3, b00, b01);
            o[2 * jj][0] = acc[0]; o[2 * jj][1] = acc[1];
            if constexpr (RS == 2) { o[2 * jj][2] = acc[2]; o[2 * jj][3] = acc[3]; }
          }
          if (2 * jj + 1 < NDT) {
            float acc[4] = {o[2 * jj + 1][0], o[2 * jj + 1][1],
                            RS == 2 ? o[2 * jj + 1][2 * (RS - 1)] : 0.f,
                            RS == 2 ? o[2 * jj + 1][2 * (RS - 1) + 1] : 0.f};
            MmaOp<MT>::run(acc, pa0, pa1, pa2, pa3, b10, b11);
            o[2 * jj + 1][0] = acc[0]; o[2 * jj + 1][1] = acc[1];
            if constexpr (RS == 2) { o[2 * jj + 1][2] = acc[2]; o[2 * jj + 1][3] = acc[3]; }
          }
        }
      } else {  // BS == 8: rows are 16 B; x4 = 4 d-tiles, tokens 8..15 of the k-step are zero
        const uint32_t base = v_smem + (uint32_t)((mi * 8 + r8) * 16);
#pragma unroll
        for (int jj = 0; jj < (NDT + 3) / 4; ++jj) {
          uint32_t b[4];
          ldsm_x4(base + (uint32_t)(jj * 32 * 16), b[0], b[1], b[2], b[3]);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (4 * jj + u < NDT) {
              uint32_t bb = b[u];
              if constexpr (TAIL) bb &= vmask[0];
              float acc[4] = {o[4 * jj + u][0], o[4 * jj + u][1],
                              RS == 2 ? o[4 * jj + u][2 * (RS - 1)] : 0.f,
                              RS == 2 ? o[4 * jj + u][2 * (RS - 1) + 1] : 0.f};
              MmaOp<MT>::run(acc, pa0, pa1, pa2, pa3, bb, 0u);
              o[4 * jj + u][0] = acc[0]; o[4 * jj + u][1] = acc[1];
              if constexpr (RS == 2) { o[4 * jj + u][2] = acc[2]; o[4 * jj + u][3] = acc[3]; }
            }
          }
        }
      }
    } else {
      if constexpr (BS >= 16) {
        // fp8 V page [D][BS] bytes.  Lane (n = lane/4 -> d = 8j + n, q4) needs the tokens of its
        // two P fragments = 4*q4..4*q4+3 of this 16-token group (see page_token): one 32-bit load.
        // For q4 >= 2 tile 2kk holds the upper token pair, so the halves are swapped first.
        const uint32_t base = v_smem + (uint32_t)((lane >> 2) * BS + q4 * 4 + kk * 16);
        const uint32_t rot = (q4 & 2) ? 16u : 0u;
#pragma unroll
        for (int j = 0; j < NDT; ++j) {
          uint32_t w, b0, b1;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(base + (uint32_t)(j * 8 * BS)));
          w = __funnelshift_l(w, w, rot);
          fp8x4_to_f16x2x2<KV>(w, b0, b1);
          if constexpr (TAIL) {
            b0 &= vmask[2 * kk];
            b1 &= vmask[2 * kk + 1];
          }
          float acc[4] = {o[j][0], o[j][1], RS == 2 ? o[j][2 * (RS - 1)] : 0.f,
                          RS == 2 ? o[j][2 * (RS - 1) + 1] : 0.f};
          MmaOp<MT>::run(acc, pa0, pa1, pa2, pa3, b0, b1);
          o[j][0] = acc[0]; o[j][1] = acc[1];
          if constexpr (RS == 2) { o[j][2] = acc[2]; o[j][3] = acc[3]; }
        }
      } else {
        // 8-token fp8 pages: tokens (2q4, 2q4+1) of row d = 8j + n; upper half of the k-step is zero
        const uint32_t base = v_smem + (uint32_t)((lane >> 2) * BS + q4 * 2);
#pragma unroll
        for (int j = 0; j < NDT; ++j) {
          uint16_t w0;
          asm volatile("ld.shared.u16 %0, [%1];" : "=h"(w0) : "r"(base + (uint32_t)(j * 8 * BS)));
          uint32_t b0 = fp8x2_to_f16x2<KV>(w0);
          if constexpr (TAIL) b0 &= vmask[0];
          float acc[4] = {o[j][0], o[j][1], RS == 2 ? o[j][2 * (RS - 1)] : 0.f,
                          RS == 2 ? o[j][2 * (RS - 1) + 1] : 0.f};
          MmaOp<MT>::run(acc, pa0, pa1, pa2, pa3, b0, 0u);
          o[j][0] = acc[0]; o[j][1] = acc[1];
          if constexpr (RS == 2) { o[j][2] = acc[2]; o[j][3] = acc[3]; }
        }
      }
    }
  }
  (void)ES;
}

// Decode form: uniform tail, optional per-row on/off (blocksparse).
template <typename T, int D, int BS, int KV, int GROWS, bool TAIL, bool PLAIN>
__device__ __forceinline__ void process_page(uint32_t k_smem, uint32_t v_smem,
                                             const uint32_t (&qa)[(D + 15) / 16][GROWS / 4],
                                             float (&o)[D / 8][GROWS / 4], float (&m)[GROWS / 8],
                                             float (&l)[GROWS / 8], int lane, int valid,
                                             float scale_log2, const float (&slope)[GROWS / 8],
                                             float tok_rel0, const bool (&row_on)[GROWS / 8]) {
  const int zi[GROWS / 8] = {};
  const float zf[GROWS / 8] = {};
  process_page_ex<T, D, BS, KV, GROWS, TAIL, PLAIN, false>(k_smem, v_smem, qa, o, m, l, lane, valid,
                                                           scale_log2, slope, tok_rel0, row_on, zi,
                                                           zi, zf);
}

}  // namespace b200
