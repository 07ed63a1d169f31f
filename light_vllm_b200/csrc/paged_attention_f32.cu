// float32-query paged attention (the reference dispatches float alongside half/bfloat16:
// quant_utils.cuh:532-569, attention_kernels.cu:738-769 with scalar_t = float; caches are float
// for "auto" (x = 4) or uint8 for fp8).
//
// Same CTA / pipeline structure as the 16-bit kernel (paged_attention_main.cuh): one CTA per
// (sequence, KV head[, group chunk], context split), every warp streams its own interleaved pages
// through a private ring filled by ONE cp.async.bulk per K page and per V page, no CTA-wide
// barrier in the main loop, online softmax, warp merge through shared memory, v2 partitions
// through the caller's scratch.  The arithmetic is plain fp32 FMA on the CUDA cores: there is no
// fp32 tensor-core MMA (tf32 would drop 13 mantissa bits of q and k), and full-precision serving
// is a compatibility path of the reference, not a bandwidth target, so exact fp32 products matter
// more than issue rate here.
//
// Lane roles per page (BS tokens):
//   QK^T : lane = (token t = lane % BS, part = lane / BS); the 32/BS parts split the 16-byte
//          chunks of the head dim (K layout [D/x][BS][x]: lane t of one part reads consecutive
//          16-byte rows -> conflict-free), partial dots are xor-reduced across parts.
//   P.V  : lane = (row r = lane / 4, token quarter tl = lane % 4); pass j covers d = 8j + r; each
//          lane keeps the partial sums of ITS BS/4 tokens in registers for the whole sequence and
//          the four quarters are reduced once, after the last page.
#include "paged_attention.cuh"

#include <atomic>

namespace b200 {

namespace {

template <int KV>
__device__ __forceinline__ float fp8_to_f32_raw(uint8_t v) {  // exact: fp8 -> half -> float
  constexpr __nv_fp8_interpretation_t kind = (KV == B200_KV_FP8_E5M2) ? __NV_E5M2 : __NV_E4M3;
  __half_raw h = __nv_cvt_fp8_to_halfraw(v, kind);
  return __half2float(*reinterpret_cast<__half*>(&h));
}

template <int D, int BS, int KV, int GR>
__global__ void __launch_bounds__(128)
paged_attention_f32_kernel(const __grid_constant__ AttnParams p) {
  constexpr int ES = (KV == B200_KV_AUTO) ? 4 : 1;
  constexpr int X = 16 / ES;                 // cache elements per 16-byte K chunk
  constexpr int NCH = D / X;                 // K chunks over the head dim
  constexpr int PAGE_BYTES = D * BS * ES;
  constexpr int STAGE_BYTES = 2 * PAGE_BYTES;
  constexpr int NPART = 32 / BS;             // lane groups splitting the K chunks
  constexpr int TPL = BS / 4;                // tokens per lane in P.V
  constexpr int NPASS = D / 8;               // d = 8j + r
  constexpr int ROWF = D + kMergePad;
  static_assert(D % X == 0 && D % 8 == 0, "head size must be a multiple of the K chunk");

  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  const int NS = p.num_stages;
  const bool is_v2 = p.partition_pages > 0;

  const int seq = blockIdx.z;
  const int kvh = blockIdx.y / p.gchunks;
  const int gchunk = blockIdx.y - kvh * p.gchunks;
  const int head0 = kvh * p.group + gchunk * p.rows_per_cta;
  const int gcount = min(p.rows_per_cta, p.group - gchunk * p.rows_per_cta);
  const int seq_len = p.seq_lens[seq];
  const int n_pages = (seq_len + BS - 1) / BS;

  int page_begin = 0, page_end = n_pages;
  if (is_v2) {
    page_begin = blockIdx.x * p.partition_pages;
    if (page_begin >= n_pages) return;
    page_end = min(page_begin + p.partition_pages, n_pages);
  }

  const uint32_t ring_bytes = (uint32_t)nwarps * NS * STAGE_BYTES;
  const uint32_t merge_bytes = (uint32_t)(nwarps + 1) * GR * (ROWF + 2) * 4;
  const uint32_t data_bytes = ((ring_bytes > merge_bytes ? ring_bytes : merge_bytes) + 127) & ~127u;
  unsigned char* ring = smem + (size_t)warp * NS * STAGE_BYTES;
  float* qs = reinterpret_cast<float*>(smem + data_bytes);                       // [GR][D]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + data_bytes + GR * D * 4) + warp * NS;
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t bars_u32 = smem_u32(bars);

  if (lane == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(bars_u32 + 8 * s, 1);
    mbar_fence_init();
  }
  __syncwarp();

  const int n_local = page_end - page_begin;
  const int n_my = n_local > warp ? (n_local - warp + nwarps - 1) / nwarps : 0;
  const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
  const unsigned char* kbase =
      static_cast<const unsigned char*>(p.k_cache) + (int64_t)kvh * p.kv_head_stride * ES;
  const unsigned char* vbase =
      static_cast<const unsigned char*>(p.v_cache) + (int64_t)kvh * p.kv_head_stride * ES;
  const int64_t blk_bytes = p.kv_block_stride * ES;
  const uint64_t policy = l2_policy_evict_first();

  auto issue = [&](int i) {
    if (elect_one()) {
      const int64_t blk = bt[page_begin + warp + i * nwarps];
      const int s = i % NS;
      const uint32_t bar = bars_u32 + 8 * s;
      const uint32_t dst = ring_u32 + (uint32_t)s * STAGE_BYTES;
      mbar_arrive_expect_tx(bar, STAGE_BYTES);
      bulk_g2s(dst, kbase + blk * blk_bytes, PAGE_BYTES, bar, policy);
      bulk_g2s(dst + PAGE_BYTES, vbase + blk * blk_bytes, PAGE_BYTES, bar, policy);
    }
  };
  const int pre = min(NS, n_my);
  for (int i = 0; i < pre; ++i) issue(i);

  // q rows of this CTA's heads -> shared memory (zeros for padded rows)
  {
    const float* q_seq = static_cast<const float*>(p.q) + (int64_t)seq * p.q_stride;
    for (int idx = threadIdx.x; idx < GR * D; idx += blockDim.x) {
      const int h = idx / D, d = idx - h * D;
      qs[idx] = h < gcount ? q_seq[(int64_t)(head0 + h) * D + d] : 0.f;
    }
  }
  __syncthreads();

  float o[GR][NPASS];
#pragma unroll
  for (int h = 0; h < GR; ++h)
#pragma unroll
    for (int j = 0; j < NPASS; ++j) o[h][j] = 0.f;
  float m[GR], l[GR], slope[GR];
  int bs_off[GR];
#pragma unroll
  for (int h = 0; h < GR; ++h) {
    m[h] = kNegInf;
    l[h] = 0.f;
    slope[h] = (p.alibi_slopes != nullptr && h < gcount)
                   ? p.alibi_slopes[head0 + h] * 1.4426950408889634f
                   : 0.f;
    const int head = head0 + min(h, gcount - 1);
    bs_off[h] = (p.bs_head_sliding_step >= 0)
                    ? (p.bs_tp_rank * p.num_heads + head) * p.bs_head_sliding_step + 1
                    : (p.bs_tp_rank * p.num_kv_heads + kvh) * (-p.bs_head_sliding_step) + 1;
  }
  const int q_bs_block_id = p.bs_enabled ? (seq_len - 1) / p.bs_block_size : 0;

  const int t = lane % BS, part = lane / BS;
  const int tl = lane & 3, r = lane >> 2;

  int stage = 0;
  uint32_t parity = 0;
  for (int i = 0; i < n_my; ++i) {
    const int page = page_begin + warp + i * nwarps;
    const int tok0 = page * BS;
    const unsigned char* kp = ring + (size_t)stage * STAGE_BYTES;
    const unsigned char* vp = kp + PAGE_BYTES;
    mbar_wait(bars_u32 + 8 * stage, parity);

    // ---- S = q . k for token t, chunks part, part + NPART, ... ----
    float acc[GR];
#pragma unroll
    for (int h = 0; h < GR; ++h) acc[h] = 0.f;
#pragma unroll 2
    for (int c = part; c < NCH; c += NPART) {
      float kf[X];
      const uint4 raw = *reinterpret_cast<const uint4*>(kp + ((size_t)c * BS + t) * 16);
      if constexpr (KV == B200_KV_AUTO) {
        kf[0] = __uint_as_float(raw.x); kf[1] = __uint_as_float(raw.y);
        kf[2] = __uint_as_float(raw.z); kf[3] = __uint_as_float(raw.w);
      } else {
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int e = 0; e < 16; ++e) kf[e] = fp8_to_f32_raw<KV>((uint8_t)(w[e >> 2] >> (8 * (e & 3))));
      }
#pragma unroll
      for (int h = 0; h < GR; ++h) {
        const float4* qv = reinterpret_cast<const float4*>(qs + h * D + c * X);
#pragma unroll
        for (int e4 = 0; e4 < X / 4; ++e4) {
          const float4 qq = qv[e4];
          acc[h] = fmaf(qq.x, kf[4 * e4], acc[h]);
          acc[h] = fmaf(qq.y, kf[4 * e4 + 1], acc[h]);
          acc[h] = fmaf(qq.z, kf[4 * e4 + 2], acc[h]);
          acc[h] = fmaf(qq.w, kf[4 * e4 + 3], acc[h]);
        }
      }
    }
#pragma unroll
    for (int off = BS; off < 32; off <<= 1)
#pragma unroll
      for (int h = 0; h < GR; ++h) acc[h] += __shfl_xor_sync(0xffffffffu, acc[h], off);

    // ---- online softmax (exp2 domain); m, l are warp-uniform ----
    const int token = tok0 + t;
    const bool tok_valid = token < seq_len;
    const int k_bs_block_id = p.bs_enabled ? tok0 / p.bs_block_size : 0;
    float prob[GR], alpha[GR];
    bool changed = false;
#pragma unroll
    for (int h = 0; h < GR; ++h) {
      float x = acc[h] * p.scale_log2 + slope[h] * (float)(token - seq_len + 1);
      bool on = tok_valid;
      if (p.bs_enabled) {
        const bool is_remote = ((k_bs_block_id + bs_off[h]) % p.bs_vert_stride) == 0;
        const bool is_local = k_bs_block_id > q_bs_block_id - p.bs_local_blocks;
        on = on && (is_remote || is_local);
      }
      if (!on) x = kNegInf;
      float mx = x;
#pragma unroll
      for (int off = BS / 2; off >= 1; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float m_new = fmaxf(m[h], mx);
      const float m_safe = (m_new == kNegInf) ? 0.f : m_new;
      alpha[h] = fast_exp2(m[h] - m_safe);
      changed |= (m_new != m[h]);
      m[h] = m_new;
      const float pr = fast_exp2(x - m_safe);
      float sum = pr;
#pragma unroll
      for (int off = BS / 2; off >= 1; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      l[h] = l[h] * alpha[h] + sum;
      prob[h] = pr;
    }
    if (changed) {  // warp-uniform
#pragma unroll
      for (int h = 0; h < GR; ++h)
#pragma unroll
        for (int j = 0; j < NPASS; ++j) o[h][j] *= alpha[h];
    }

    // ---- O += P V over this lane's TPL tokens ----
    float pt[GR][TPL];
    bool vok[TPL];
#pragma unroll
    for (int e = 0; e < TPL; ++e) {
      vok[e] = tok0 + tl * TPL + e < seq_len;  // V past the end may hold NaN: never multiply it
#pragma unroll
      for (int h = 0; h < GR; ++h) pt[h][e] = __shfl_sync(0xffffffffu, prob[h], tl * TPL + e);
    }
#pragma unroll
    for (int j = 0; j < NPASS; ++j) {
      const int d = 8 * j + r;
      float vv[TPL];
#pragma unroll
      for (int e = 0; e < TPL; ++e) {
        const size_t idx = (size_t)d * BS + tl * TPL + e;
        float v;
        if constexpr (KV == B200_KV_AUTO)
          v = reinterpret_cast<const float*>(vp)[idx];
        else
          v = fp8_to_f32_raw<KV>(vp[idx]);
        vv[e] = vok[e] ? v : 0.f;
      }
#pragma unroll
      for (int h = 0; h < GR; ++h)
#pragma unroll
        for (int e = 0; e < TPL; ++e) o[h][j] = fmaf(pt[h][e], vv[e], o[h][j]);
    }

    __syncwarp();
    if (i + NS < n_my) issue(i + NS);
    if (++stage == NS) {
      stage = 0;
      parity ^= 1u;
    }
  }

  // ---- reduce the four token quarters, then merge warps through shared memory ----
#pragma unroll
  for (int h = 0; h < GR; ++h)
#pragma unroll
    for (int j = 0; j < NPASS; ++j) {
      o[h][j] += __shfl_xor_sync(0xffffffffu, o[h][j], 1);
      o[h][j] += __shfl_xor_sync(0xffffffffu, o[h][j], 2);
    }
  __syncthreads();
  float* mo = reinterpret_cast<float*>(smem);       // [nwarps][GR][ROWF]
  float* mm = mo + (size_t)nwarps * GR * ROWF;      // [nwarps][GR]
  float* ml = mm + (size_t)nwarps * GR;             // [nwarps][GR]
  float* fo = ml + (size_t)nwarps * GR;             // [GR][ROWF]
  float* fm = fo + (size_t)GR * ROWF;
  float* fl = fm + GR;
#pragma unroll
  for (int h = 0; h < GR; ++h) {
    if (tl == 0) {
#pragma unroll
      for (int j = 0; j < NPASS; ++j) mo[((size_t)warp * GR + h) * ROWF + 8 * j + r] = o[h][j];
    }
    if (lane == 0) {
      mm[warp * GR + h] = m[h];
      ml[warp * GR + h] = l[h];
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < gcount * D; idx += blockDim.x) {
    const int row = idx / D, d = idx - row * D;
    float M = kNegInf;
    for (int w = 0; w < nwarps; ++w) M = fmaxf(M, mm[w * GR + row]);
    float L = 0.f, O = 0.f;
    if (M != kNegInf) {
      for (int w = 0; w < nwarps; ++w) {
        const float sc = fast_exp2(mm[w * GR + row] - M);
        L += ml[w * GR + row] * sc;
        O += mo[((size_t)w * GR + row) * ROWF + d] * sc;
      }
    }
    fo[row * ROWF + d] = O;
    if (d == 0) {
      fm[row] = M;
      fl[row] = L;
    }
  }
  __syncthreads();

  for (int idx = threadIdx.x; idx < gcount * D; idx += blockDim.x) {
    const int row = idx / D, d = idx - row * D;
    const float inv = __fdividef(1.f, fl[row] + 1e-6f);
    const float val = fo[row * ROWF + d] * inv * p.out_scale;
    if (is_v2) {
      const int64_t hp =
          ((int64_t)seq * p.num_heads + head0 + row) * p.max_num_partitions + blockIdx.x;
      static_cast<float*>(p.tmp_out)[hp * D + d] = val;
      if (d == 0) {
        p.max_logits[hp] = fm[row] * 0.6931471805599453f;
        p.exp_sums[hp] = fl[row];
      }
    } else {
      static_cast<float*>(p.out)[((int64_t)seq * p.num_heads + head0 + row) * D + d] = val;
    }
  }
}

template <int D, int BS, int KV, int GR>
size_t smem_f32(int nwarps, int nstages) {
  constexpr int ES = (KV == B200_KV_AUTO) ? 4 : 1;
  const size_t ring = (size_t)nwarps * nstages * 2 * D * BS * ES;
  const size_t merge = (size_t)(nwarps + 1) * GR * (D + kMergePad + 2) * 4;
  const size_t data = ((ring > merge ? ring : merge) + 127) & ~(size_t)127;
  return data + (size_t)GR * D * 4 + (size_t)nwarps * nstages * 8 + 128;
}

template <int D, int BS, int KV, int GR>
int launch_f32(const AttnParams& p, dim3 grid, int /*cluster*/, int nwarps, int nstages, int device,
               cudaStream_t stream) {
  auto kern = paged_attention_f32_kernel<D, BS, KV, GR>;
  const size_t smem = smem_f32<D, BS, KV, GR>(nwarps, nstages);
  static std::atomic<size_t> configured[64];
  const int slot = device & 63;
  if (configured[slot].load(std::memory_order_acquire) < smem) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(227 * 1024)));
    configured[slot].store((size_t)227 * 1024, std::memory_order_release);
  }
  kern<<<grid, nwarps * 32, smem, stream>>>(p);
  B200_CUDA_OK(cudaGetLastError());
  count_launch();
  return B200_OK;
}

template <int D, int BS, int KV, int GR>
int occ_f32(int nwarps, int nstages) {
  auto kern = paged_attention_f32_kernel<D, BS, KV, GR>;
  const size_t smem = smem_f32<D, BS, KV, GR>(nwarps, nstages);
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, nwarps * 32, smem) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

template <int D, int BS, int KV>
AttnEntry pick_gr(int gr) {
  constexpr int GMAX = D <= 128 ? 4 : 2;
  if (gr == 1) return {launch_f32<D, BS, KV, 1>, smem_f32<D, BS, KV, 1>, occ_f32<D, BS, KV, 1>};
  if (gr == GMAX)
    return {launch_f32<D, BS, KV, GMAX>, smem_f32<D, BS, KV, GMAX>, occ_f32<D, BS, KV, GMAX>};
  return {nullptr, nullptr, nullptr};
}

template <int D, int KV>
AttnEntry pick_bs(int bs, int gr) {
#ifndef B200_FAST_BUILD
  if (bs == 8) return pick_gr<D, 8, KV>(gr);
  if (bs == 32) return pick_gr<D, 32, KV>(gr);
#endif
  if (bs == 16) return pick_gr<D, 16, KV>(gr);
  return {nullptr, nullptr, nullptr};
}

template <int KV>
AttnEntry pick_head(int head_size, int bs, int gr) {
  switch (head_size) {
    case 64: return pick_bs<64, KV>(bs, gr);
    case 128: return pick_bs<128, KV>(bs, gr);
#ifndef B200_FAST_BUILD
    case 80: return pick_bs<80, KV>(bs, gr);
    case 96: return pick_bs<96, KV>(bs, gr);
    case 112: return pick_bs<112, KV>(bs, gr);
    case 120:
      if constexpr (KV == B200_KV_AUTO) return pick_bs<120, KV>(bs, gr);
      return {nullptr, nullptr, nullptr};  // fp8 pages need head_size % 16 == 0
    case 192: return pick_bs<192, KV>(bs, gr);
    case 256: return pick_bs<256, KV>(bs, gr);
#endif
    default: return {nullptr, nullptr, nullptr};
  }
}

}  // namespace

// Group rows one CTA of the fp32 kernel covers for a GQA group of `group` q-heads.
int attn_f32_group_rows(int group, int head_size) {
  return group == 1 ? 1 : (head_size <= 128 ? 4 : 2);
}

AttnEntry attn_lookup_f32(int head_size, int block_size, int kv_dtype, int gr) {
  switch (kv_dtype) {
    case B200_KV_AUTO: return pick_head<B200_KV_AUTO>(head_size, block_size, gr);
    case B200_KV_FP8_E4M3: return pick_head<B200_KV_FP8_E4M3>(head_size, block_size, gr);
    case B200_KV_FP8_E5M2: return pick_head<B200_KV_FP8_E5M2>(head_size, block_size, gr);
    default: return {nullptr, nullptr, nullptr};
  }
}

}  // namespace b200
