// Context-sharded decode: ONE sequence's KV pages live on several GPUs (the
// long-context case sequence sharding cannot cover).  Every rank runs the paged-attention
// partition kernel over the pages it owns, then ONE kernel per rank
//   (1) folds the rank's partitions into a single (m, l, o) triple per (seq, head),
//   (2) all-gathers the triples by storing them straight into every peer's exchange buffer over
//       NVLink (peer pointers from CUDA IPC; no NCCL, no host round trip) as 8-byte words
//       {fp32 value, epoch}: the flag travels INSIDE every word (the "LL" idea of NCCL's low-latency
//       protocol), so there is no system fence and no separate flag round trip on the critical path,
//   (3) polls the words of its own buffer until they carry this call's epoch, and
//   (4) merges the `world` triples with the v2 formula (attention_kernels.cu:565-669) into `out`.
// Every rank ends with the full output, as the callers of paged_attention expect.
//
// Exchange buffer layout (identical on every rank; b200_ctx_exchange_bytes):
//   u32   status[32]                           [0] != 0: a wait timed out (1 + rank that never arrived);
//                                              [1] device-managed epoch, [2] its completion counter,
//                                              [3] epoch of the call that timed out, [4] timeout in ms
//                                              (0 = 2000; b200_ctx_exchange_set_timeout),
//                                              [6..7] device-visible pointer to 4 pinned host words that
//                                              mirror [0] and [3] on a timeout (0 = none)
//   u64   trace[2][8][2]                       CTA 0's (globaltimer ns, SM clock) at eight points of the
//                                              exchange kernel, for the last two calls (slot = epoch & 1;
//                                              b200_ctx_exchange_trace)
//   u64   data[2][world][max_pairs][D + 2]     row = {m, l, o[0..D)}, word = value | epoch << 32; slot = epoch & 1
// Two data slots suffice: a rank can only be one call ahead of its slowest peer, because call
// n+1 cannot finish before every peer has published call n+1, which it does after reading call n.
// A word of slot n & 1 still holding epoch n-2 never equals n, so stale rows are never merged; the
// comparison is for equality, hence safe across the 32-bit wrap.
#include "paged_attention.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {

constexpr int kCtxMaxWorld = 16;
constexpr int kCtxMaxCtas = 148;                               // one per SM: all CTAs co-resident
constexpr int kCtxRowExtra = 2;                                // m, l in front of the D outputs
constexpr int kCtxTraceStamps = 8;                             // per call: (globaltimer ns, SM clock) pairs
constexpr int kCtxStatusWords = 32 + 2 * kCtxTraceStamps * 4;  // status[32] + trace[2 calls][8][2] u64
constexpr int kCtxPrefetch = 8;                                // partitions per warp loaded before the weights exist
constexpr int kCtxChunk = 8;                                   // ranks polled together

struct CtxPeers {
  unsigned char* buf[kCtxMaxWorld];
};

__host__ __device__ inline size_t ctx_header_bytes() {
  return ((size_t)kCtxStatusWords * 4 + 127) & ~(size_t)127;
}

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ctx_word(float v, uint32_t epoch) {
  return (unsigned long long)__float_as_uint(v) | ((unsigned long long)epoch << 32);
}

template <typename T>
__global__ void __launch_bounds__(128, 4)  // <= 128 registers: must fit beside the partition pass's resident CTAs
ctx_exchange_merge_kernel(T* __restrict__ out, const float* __restrict__ exp_sums,
                          const float* __restrict__ max_logits, const T* __restrict__ tmp_out,
                          const int32_t* __restrict__ seq_lens, int num_heads, int head_size,
                          int max_num_partitions, int partition_tokens, int parts_cap,
                          int total_pairs, const __grid_constant__ CtxPeers peers, int rank,
                          int world, int max_pairs, uint32_t epoch, int phases,
                          long long timeout_cycles) {
  extern __shared__ float wsm[];  // [4 warps][head_size] partial folds | [parts_cap] max logits | [parts_cap] exp sums
  __shared__ uint32_t epoch_s;
  __shared__ int timed_out_s;
  __shared__ long long timeout_s;
  unsigned long long stamp_g[kCtxTraceStamps], stamp_c[kCtxTraceStamps];  // CTA 0 / thread 0 only
#define CTX_STAMP(i)                                                       \
  if (blockIdx.x == 0 && threadIdx.x == 0) {                               \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(stamp_g[i]));         \
    stamp_c[i] = (unsigned long long)clock64();                            \
  }
  CTX_STAMP(0)
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  // epoch == 0: device-managed.  The call's epoch is the counter in this rank's own header + 1; the
  // last CTA to finish the merge phase advances it, so that a CUDA graph holding this launch can be
  // replayed (a host-supplied epoch would be frozen into the graph).
  unsigned char* mine = peers.buf[rank];
  uint32_t* status = reinterpret_cast<uint32_t*>(mine);
  uint32_t* hdr_epoch = status + 1;
  const bool device_epoch = epoch == 0;
  if (threadIdx.x == 0) {  // no barrier here: the first barrier of either phase publishes these
    epoch_s = device_epoch ? __ldcg(hdr_epoch) + 1u : epoch;
    timed_out_s = 0;
    const uint32_t ms = __ldcg(status + 4);  // per-group timeout, 0 = default
    timeout_s = ms ? (long long)ms * 2000000LL : timeout_cycles;  // ~2 GHz SM clock
  }
  const int row_words = head_size + kCtxRowExtra;
  const size_t hdr = ctx_header_bytes();
  const size_t slot_words = (size_t)world * max_pairs * row_words;
  // Launched with programmatic stream serialisation: the partition pass (which triggers its dependents as
  // soon as all of its CTAs run) lets this grid become resident early and park in griddepcontrol.wait.  Only
  // this group's own header -- last written by the previous exchange launch, which the partition pass already
  // waited for -- is read above.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  CTX_STAMP(1)

  // ---- phase 1: fold my partitions, publish the triple to every rank (incl. myself) ----
  // One CTA per (seq, head) pair.  The four warps fold interleaved quarters of the partitions; the first
  // kCtxPrefetch rows of every warp are requested BEFORE the statistics they will be weighted with have
  // arrived, so that the fold costs one memory round trip instead of three.
  float* part = wsm;                          // [4 warps][head_size]
  float* sm_ml = wsm + 4 * head_size;         // [parts_cap]
  float* sm_es = sm_ml + parts_cap;           // [parts_cap]
  for (int pair = blockIdx.x; (phases & 1) && pair < total_pairs; pair += gridDim.x) {
    const int seq = pair / num_heads;
    const int seq_len = seq_lens[seq];
    const float* ml = max_logits + (int64_t)pair * max_num_partitions;
    const float* es = exp_sums + (int64_t)pair * max_num_partitions;
    const T* t = tmp_out + (int64_t)pair * max_num_partitions * head_size;
    float pre[kCtxPrefetch][2][2];  // [partition warp + 4u][64-wide slice of the row (the first 128 elements)][2]
#pragma unroll
    for (int u = 0; u < kCtxPrefetch; ++u) {
      const int j = warp + 4 * u;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int d0 = 2 * lane + 64 * c;
        // rows up to parts_cap exist in the scratch tensor whether or not this sequence filled them: the guard
        // does not depend on seq_lens, so these loads leave together with that one (unused rows get weight 0
        // through a select, never through a multiplication)
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
    if (pair == blockIdx.x) CTX_STAMP(2)
    float mx = kNegInf;
    for (int j = lane; j < np; j += 32) mx = fmaxf(mx, sm_ml[j]);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, s));
    float sum = 0.f;
    for (int j = lane; j < np; j += 32) sum += (mx == kNegInf) ? 0.f : sm_es[j] * expf(sm_ml[j] - mx);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, s);
    const float inv = sum > 0.f ? __fdividef(1.f, sum) : 0.f;
    float acc[4][2];
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[c][0] = acc[c][1] = 0.f;
#pragma unroll
    for (int u = 0; u < kCtxPrefetch; ++u) {
      const int j = warp + 4 * u;
      const float wj = (j < np && mx != kNegInf) ? sm_es[j] * expf(sm_ml[j] - mx) : 0.f;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        acc[c][0] += j < np ? pre[u][c][0] * wj : 0.f;
        acc[c][1] += j < np ? pre[u][c][1] * wj : 0.f;
      }
    }
    // whatever the prefetch did not cover: partitions beyond 4 * kCtxPrefetch, elements beyond 128
    if (np > 4 * kCtxPrefetch || head_size > 128) {
      for (int j = warp; j < np; j += 4) {
        const float wj = (mx != kNegInf) ? sm_es[j] * expf(sm_ml[j] - mx) : 0.f;
        const bool fetched = j < 4 * kCtxPrefetch;
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
    if (pair == blockIdx.x) CTX_STAMP(3)
    const uint32_t ep = epoch_s;
    const size_t row_off = (size_t)(ep & 1u) * slot_words + ((size_t)rank * max_pairs + pair) * row_words;
    for (int w = threadIdx.x; w < row_words; w += blockDim.x) {
      float v;
      if (w < kCtxRowExtra) {
        v = w == 0 ? mx : sum;
      } else {
        const int d = w - kCtxRowExtra;
        v = (part[d] + part[head_size + d] + part[2 * head_size + d] + part[3 * head_size + d]) * inv;
      }
      const unsigned long long word = ctx_word(v, ep);
      for (int r = 0; r < world; ++r)
        st_relaxed_sys_u64(reinterpret_cast<unsigned long long*>(peers.buf[r] + hdr) + row_off + w, word);
    }
    __syncthreads();  // part[] / sm_ml[] / sm_es[] are reused by the next pair
  }
  CTX_STAMP(4)

  // ---- phase 2: wait for every rank's row of this epoch, merge ----
  if (!(phases & 2)) return;
  __syncthreads();  // epoch_s / timeout_s (merge-only launches run no phase-1 barrier)
  const uint32_t epoch_now = epoch_s;
  timeout_cycles = timeout_s;
  const unsigned long long* data =
      reinterpret_cast<const unsigned long long*>(mine + hdr) + (size_t)(epoch_now & 1u) * slot_words;
  bool poisoned = false;
  for (int pair = blockIdx.x; pair < total_pairs; pair += gridDim.x) {
    T* o = out + (int64_t)pair * head_size;
    // thread -> output elements d = threadIdx.x (+ 128): needs {m, l, o[d]} of every rank
    float M = kNegInf, S = 0.f, acc0 = 0.f, acc1 = 0.f;
    const int d_a = threadIdx.x, d_b = threadIdx.x + blockDim.x;
    bool late = false;
    for (int r0 = 0; r0 < world && !poisoned; r0 += kCtxChunk) {
      unsigned long long wm[kCtxChunk], wl[kCtxChunk], wa[kCtxChunk], wb[kCtxChunk];
      const unsigned long long want = (unsigned long long)epoch_now << 32;
      auto load = [&](int i) {
        const unsigned long long* row = data + ((size_t)(r0 + i) * max_pairs + pair) * row_words;
        wm[i] = ld_relaxed_sys_u64(row);
        wl[i] = ld_relaxed_sys_u64(row + 1);
        wa[i] = d_a < head_size ? ld_relaxed_sys_u64(row + kCtxRowExtra + d_a) : want;
        wb[i] = d_b < head_size ? ld_relaxed_sys_u64(row + kCtxRowExtra + d_b) : want;
      };
      auto ready = [&](int i) {
        return (((wm[i] ^ want) | (wl[i] ^ want) | (wa[i] ^ want) | (wb[i] ^ want)) >> 32) == 0;
      };
#pragma unroll
      for (int i = 0; i < kCtxChunk; ++i)
        if (r0 + i < world) load(i);
#pragma unroll
      for (int i = 0; i < kCtxChunk; ++i) {
        if (r0 + i >= world || late) continue;
        if (!ready(i)) {
          const long long t0 = clock64();
          do {
            if (clock64() - t0 > timeout_cycles) {
              if (atomicCAS(status, 0u, 1u + (uint32_t)(r0 + i)) == 0u) {  // first failure: who and when
                status[3] = epoch_now;
                // mirror into the caller's pinned host words (b200_ctx_exchange_set_host_mirror): the host
                // sees the failure at its next call without any per-step device-to-host copy
                uint32_t* host = *reinterpret_cast<uint32_t* volatile*>(status + 6);
                if (host != nullptr) {
                  host[3] = epoch_now;
                  __threadfence_system();
                  host[0] = 1u + (uint32_t)(r0 + i);
                }
              }
              late = true;
              break;
            }
            __nanosleep(32);
            load(i);
          } while (!ready(i));
        }
        if (late) continue;
        const float m = __uint_as_float((uint32_t)wm[i]), l = __uint_as_float((uint32_t)wl[i]);
        if (l > 0.f) {  // a rank that owns no page of this sequence contributes nothing
          const float Mn = fmaxf(M, m);
          const float keep = expf(M - Mn), w = l * expf(m - Mn);  // M = -inf: keep = 0 (S, acc are 0 anyway)
          S = S * keep + w;
          acc0 = acc0 * keep + __uint_as_float((uint32_t)wa[i]) * w;
          acc1 = acc1 * keep + __uint_as_float((uint32_t)wb[i]) * w;
          M = Mn;
        }
      }
      if (late) break;
    }
    // A peer never published this epoch: its row is stale or half written.  Fail loudly in the data --
    // NaN rows for this and every later pair of the CTA -- instead of merging it (the status word says
    // who and which epoch).
    if (__syncthreads_or(late ? 1 : 0)) poisoned = true;
    if (pair == blockIdx.x) CTX_STAMP(5)
    const float inv = __fdividef(1.f, S + 1e-6f);
    const float nan = __int_as_float(0x7fc00000);
    if (d_a < head_size) o[d_a] = from_f32<T>(poisoned ? nan : acc0 * inv);
    if (d_b < head_size) o[d_b] = from_f32<T>(poisoned ? nan : acc1 * inv);
  }
  epoch = epoch_now;
  CTX_STAMP(6)
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    CTX_STAMP(7)
    unsigned long long* tr =
        reinterpret_cast<unsigned long long*>(status + 32) + (epoch & 1u) * 2 * kCtxTraceStamps;
#pragma unroll
    for (int i = 0; i < kCtxTraceStamps; ++i) {
      tr[2 * i] = stamp_g[i];
      tr[2 * i + 1] = stamp_c[i];
    }
  }
#undef CTX_STAMP
  if (device_epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int* done = reinterpret_cast<unsigned int*>(hdr_epoch) + 1;
      __threadfence();
      if (atomicAdd(done, 1u) == gridDim.x - 1) {  // every CTA of this launch has read hdr_epoch
        *done = 0u;
        *hdr_epoch = epoch;
        __threadfence();
      }
    }
  }
}

int launch_ctx_exchange(void* out, const float* exp_sums, const float* max_logits,
                        const void* tmp_out, const int32_t* seq_lens, int num_seqs, int num_heads,
                        int head_size, int max_num_partitions, int partition_tokens, int parts,
                        int q_dtype, void* const* exchange_bufs, int rank, int world, int max_pairs,
                        uint32_t epoch, int phases, cudaStream_t stream) {
  B200_CHECK_ARG(world >= 1 && world <= kCtxMaxWorld && rank >= 0 && rank < world,
                 "ctx_shard: bad rank/world (%d/%d)", rank, world);
  const long long pairs = (long long)num_seqs * num_heads;
  B200_CHECK_ARG(pairs <= max_pairs, "ctx_shard: %lld (seq, head) pairs exceed the exchange buffer's %d",
                 pairs, max_pairs);
  B200_CHECK_ARG(head_size % 2 == 0 && head_size <= 256, "ctx_shard: head size must be even and <= 256");
  B200_CHECK_ARG(phases >= 1 && phases <= 3, "ctx_shard: phases must be 1 (publish), 2 (merge) or 3");
  CtxPeers peers{};
  for (int r = 0; r < world; ++r) {
    B200_CHECK_ARG(exchange_bufs && exchange_bufs[r], "ctx_shard: exchange buffer of rank %d is null", r);
    peers.buf[r] = static_cast<unsigned char*>(exchange_bufs[r]);
  }
  int blocks = (int)pairs;
  if (blocks > kCtxMaxCtas) blocks = kCtxMaxCtas;  // all CTAs co-resident: a waiting CTA never keeps a publishing one out
  const int parts_cap = parts > 1 ? parts : 1;
  const size_t wsmem = ((size_t)4 * head_size + (size_t)2 * parts_cap) * sizeof(float);
  B200_CHECK_ARG(wsmem <= 48 * 1024, "ctx_shard: %d partitions per sequence exceed the exchange kernel's shared memory",
                 parts_cap);
  const long long timeout = 4000000000LL;  // ~2 s of SM clocks: a missing peer must not hang the GPU
  // Programmatic dependent launch: the grid is scheduled while the partition pass (same stream) drains and
  // parks in griddepcontrol.wait, which takes the launch latency off the step's critical path.
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
  const int pairs_i = (int)pairs;
#define B200_CTX_LAUNCH(T)                                                                                    \
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, ctx_exchange_merge_kernel<T>, static_cast<T*>(out), exp_sums,          \
                                  max_logits, static_cast<const T*>(tmp_out), seq_lens, num_heads, head_size,  \
                                  max_num_partitions, partition_tokens, parts_cap, pairs_i, peers, rank, world, \
                                  max_pairs, epoch, phases, timeout))
  if (q_dtype == B200_DT_F16) B200_CTX_LAUNCH(__half);
  else if (q_dtype == B200_DT_BF16) B200_CTX_LAUNCH(__nv_bfloat16);
  else B200_CTX_LAUNCH(float);
#undef B200_CTX_LAUNCH
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int64_t b200_ctx_exchange_bytes(int world, int max_pairs, int head_size) {
  if (world < 1 || world > kCtxMaxWorld || max_pairs < 1 || head_size < 2) return -1;
  return (int64_t)(ctx_header_bytes() +
                   (size_t)2 * world * max_pairs * (head_size + kCtxRowExtra) * sizeof(unsigned long long));
}

// Peer-visible device memory.  cudaMalloc (not a pool / VMM allocation) so that it can be exported
// with cudaIpcGetMemHandle; zero-filled so that every word carries epoch 0, "nothing published".
extern "C" int b200_peer_alloc(void** ptr, int64_t bytes, int device) {
  B200_CHECK_ARG(ptr && bytes > 0, "peer_alloc: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaMalloc(ptr, (size_t)bytes));
  B200_CUDA_OK(cudaMemset(*ptr, 0, (size_t)bytes));
  B200_CUDA_OK(cudaDeviceSynchronize());
  return B200_OK;
}

extern "C" int b200_peer_free(void* ptr, int device) {
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaFree(ptr));
  return B200_OK;
}

extern "C" int b200_peer_export(const void* ptr, unsigned char* handle64, int device) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  B200_CHECK_ARG(ptr && handle64, "peer_export: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
  memcpy(handle64, &h, 64);
  return B200_OK;
}

extern "C" int b200_peer_open(const unsigned char* handle64, void** ptr, int device) {
  B200_CHECK_ARG(ptr && handle64, "peer_open: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}

extern "C" int b200_peer_close(void* ptr, int device) {
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return B200_OK;
}

// Asynchronous copy of {status, -, -, epoch of the failing call} (4 words) into pinned host memory on
// `stream`: lets the host check the PREVIOUS call's outcome at the next call without synchronising.
extern "C" int b200_ctx_exchange_status_async(const void* local_buf, int world, int32_t* pinned_host4,
                                              int device, b200_stream_t stream) {
  B200_CHECK_ARG(local_buf && pinned_host4, "ctx_exchange_status_async: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaMemcpyAsync(pinned_host4, static_cast<const unsigned char*>(local_buf) +
                                                 (size_t)0,
                               16, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return B200_OK;
}

// Pinned (mapped) host words {status, -, -, failing epoch} that the exchange kernel writes on a timeout: the
// zero-copy alternative to b200_ctx_exchange_status_async on every step.  NULL detaches the mirror.
extern "C" int b200_ctx_exchange_set_host_mirror(void* local_buf, int world, void* pinned_host4, int device) {
  B200_CHECK_ARG(local_buf, "ctx_exchange_set_host_mirror: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  void* dptr = nullptr;
  if (pinned_host4 != nullptr) B200_CUDA_OK(cudaHostGetDevicePointer(&dptr, pinned_host4, 0));
  B200_CUDA_OK(cudaMemcpy(static_cast<unsigned char*>(local_buf) + (size_t)6 * 4, &dptr, 8,
                          cudaMemcpyHostToDevice));
  return B200_OK;
}

// Per-group wait timeout (milliseconds at ~2 GHz; 0 restores the 2 s default) and status reset.
// State lives in the group's own exchange buffer, not in the library.
extern "C" int b200_ctx_exchange_set_timeout(void* local_buf, int world, int timeout_ms, int device) {
  B200_CHECK_ARG(local_buf && timeout_ms >= 0, "ctx_exchange_set_timeout: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  const uint32_t ms = (uint32_t)timeout_ms;
  B200_CUDA_OK(cudaMemcpy(static_cast<unsigned char*>(local_buf) + (size_t)4 * 4,
                          &ms, 4, cudaMemcpyHostToDevice));
  return B200_OK;
}
extern "C" int b200_ctx_exchange_clear_status(void* local_buf, int world, int device) {
  B200_CHECK_ARG(local_buf, "ctx_exchange_clear_status: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  const uint32_t z = 0;
  unsigned char* st = static_cast<unsigned char*>(local_buf) + (size_t)0;
  B200_CUDA_OK(cudaMemcpy(st, &z, 4, cudaMemcpyHostToDevice));
  B200_CUDA_OK(cudaMemcpy(st + 12, &z, 4, cudaMemcpyHostToDevice));
  return B200_OK;
}

// Diagnostics: CTA 0's timestamps of the last two exchange launches, out[2][8][2] = (globaltimer ns, SM clock)
// at {entry, dependency resolved, first statistics loaded, first row folded, rows published, first pair arrived, merged, end};
// slot = epoch & 1.  Synchronises the device.
extern "C" int b200_ctx_exchange_trace(const void* local_buf, int world, uint64_t* out32, int device) {
  B200_CHECK_ARG(local_buf && out32, "ctx_exchange_trace: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaDeviceSynchronize());
  B200_CUDA_OK(cudaMemcpy(out32, static_cast<const unsigned char*>(local_buf) + (size_t)32 * 4,
                          (size_t)2 * kCtxTraceStamps * 2 * 8, cudaMemcpyDeviceToHost));
  return B200_OK;
}

// Synchronous read of the exchange buffer's timeout word (tests / diagnostics; not on the hot path).
extern "C" int b200_ctx_exchange_status(const void* local_buf, int world, int32_t* status, int device) {
  B200_CHECK_ARG(local_buf && status, "ctx_exchange_status: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaMemcpy(status, static_cast<const unsigned char*>(local_buf) +
                                       (size_t)0,
                          4, cudaMemcpyDeviceToHost));
  return B200_OK;
}
