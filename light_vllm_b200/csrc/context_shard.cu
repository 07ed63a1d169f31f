// Context-sharded decode: ONE sequence's KV pages live on several GPUs (the
// long-context case sequence sharding cannot cover).  Every rank runs the paged-attention
// partition kernel over the pages it owns, then ONE kernel per rank
//   (1) folds the rank's partitions into a single (m, l, o) triple per (seq, head),
//   (2) all-gathers the triples by storing them straight into every peer's exchange buffer over
//       NVLink (peer pointers from CUDA IPC; no NCCL, no host round trip),
//   (3) signals / waits on per-CTA epoch flags (st.release.sys / ld.acquire.sys), and
//   (4) merges the `world` triples with the v2 formula (attention_kernels.cu:565-669) into `out`.
// Every rank ends with the full output, as the callers of paged_attention expect.
//
// Exchange buffer layout (identical on every rank; b200_ctx_exchange_bytes):
//   u32   flags[world][kCtxMaxCtas]            epoch last published by (source rank, CTA)
//   u32   status[32]                           [0] != 0: a wait timed out (1 + rank that never arrived);
//                                              [1] device-managed epoch, [2] its completion counter,
//                                              [3] epoch of the call that timed out, [4] timeout in ms
//                                              (0 = 2000; b200_ctx_exchange_set_timeout)
//   float data[2][world][max_pairs][D + 4]     row = {m, l, 0, 0, o[0..D)}; slot = epoch & 1
// Two data slots suffice: a rank can only be one call ahead of its slowest peer, because call
// n+1 cannot finish before every peer has published call n+1, which it does after reading call n.
#include "paged_attention.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {

constexpr int kCtxMaxCtas = 160;
constexpr int kCtxMaxWorld = 16;
constexpr int kCtxRowPad = 4;

struct CtxPeers {
  unsigned char* buf[kCtxMaxWorld];
};

__host__ __device__ inline size_t ctx_header_bytes(int world) {
  return (((size_t)world * kCtxMaxCtas + 32) * 4 + 127) & ~(size_t)127;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <typename T>
__global__ void __launch_bounds__(128)
ctx_exchange_merge_kernel(T* __restrict__ out, const float* __restrict__ exp_sums,
                          const float* __restrict__ max_logits, const T* __restrict__ tmp_out,
                          const int32_t* __restrict__ seq_lens, int num_heads, int head_size,
                          int max_num_partitions, int partition_tokens, int weights_stride,
                          int total_pairs, const __grid_constant__ CtxPeers peers, int rank,
                          int world, int max_pairs, uint32_t epoch, int phases,
                          long long timeout_cycles) {
  extern __shared__ float wsm[];  // [4 warps][head_size] partial folds
  __shared__ uint32_t epoch_s;
  // Launched with programmatic stream serialisation: this grid may be scheduled while the partition pass
  // drains; nothing of its output is read before the dependency resolves.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __shared__ int timed_out_s;
  __shared__ long long timeout_s;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  // epoch == 0: device-managed.  The call's epoch is the counter in this rank's own header + 1; the
  // last CTA to finish the merge phase advances it, so that a CUDA graph holding this launch can be
  // replayed (a host-supplied epoch would be frozen into the graph).
  uint32_t* hdr_epoch = reinterpret_cast<uint32_t*>(peers.buf[rank]) + world * kCtxMaxCtas + 1;
  const bool device_epoch = epoch == 0;
  if (threadIdx.x == 0) {
    epoch_s = device_epoch ? __ldcg(hdr_epoch) + 1u : epoch;
    timed_out_s = 0;
    const uint32_t ms = __ldcg(hdr_epoch + 3);  // status[4]: per-group timeout, 0 = default
    timeout_s = ms ? (long long)ms * 2000000LL : timeout_cycles;  // ~2 GHz SM clock
  }
  __syncthreads();
  timeout_cycles = timeout_s;
  epoch = epoch_s;
  const int row_floats = head_size + kCtxRowPad;
  const size_t hdr = ctx_header_bytes(world);
  const size_t slot_floats = (size_t)world * max_pairs * row_floats;
  const size_t slot_off = (size_t)(epoch & 1u) * slot_floats;

  // ---- phase 1: fold my partitions, publish the triple to every rank (incl. myself) ----
  // One CTA per (seq, head) pair: every warp derives the pair's (max, sum) from the partition
  // statistics, the four warps fold interleaved quarters of the partitions (independent loads,
  // short dependency chains) into shared memory, and the CTA stores the normalised row to the peers.
  float* part = wsm;  // [4 warps][head_size]
  for (int pair = blockIdx.x; (phases & 1) && pair < total_pairs; pair += gridDim.x) {
    const int seq = pair / num_heads;
    const int seq_len = seq_lens[seq];
    const int np = (seq_len + partition_tokens - 1) / partition_tokens;
    const float* ml = max_logits + (int64_t)pair * max_num_partitions;
    const float* es = exp_sums + (int64_t)pair * max_num_partitions;
    const T* t = tmp_out + (int64_t)pair * max_num_partitions * head_size;
    float mx = kNegInf;
    for (int j = lane; j < np; j += 32) mx = fmaxf(mx, ml[j]);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, s));
    float sum = 0.f;
    for (int j = lane; j < np; j += 32) sum += (mx == kNegInf) ? 0.f : es[j] * expf(ml[j] - mx);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, s);
    const float inv = sum > 0.f ? __fdividef(1.f, sum) : 0.f;
    for (int d0 = 2 * lane; d0 < head_size; d0 += 64) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll 4
      for (int j = warp; j < np; j += 4) {
        const float wj = es[j] * expf(ml[j] - mx);
        a0 += to_f32<T>(t[(int64_t)j * head_size + d0]) * wj;
        a1 += to_f32<T>(t[(int64_t)j * head_size + d0 + 1]) * wj;
      }
      *reinterpret_cast<float2*>(part + warp * head_size + d0) = make_float2(a0, a1);
    }
    __syncthreads();
    const size_t row_off = slot_off + ((size_t)rank * max_pairs + pair) * row_floats;
    // 16-byte peer stores: row = {m, l, 0, 0 | o[0..D)}; thread t carries float4 number t of the row
    for (int v4 = threadIdx.x; v4 < (head_size + kCtxRowPad) / 4; v4 += blockDim.x) {
      float4 val;
      if (v4 == 0) {
        val = make_float4(mx, sum, 0.f, 0.f);
      } else {
        const int d = 4 * (v4 - 1);
        val.x = (part[d] + part[head_size + d] + part[2 * head_size + d] + part[3 * head_size + d]) * inv;
        val.y = (part[d + 1] + part[head_size + d + 1] + part[2 * head_size + d + 1] + part[3 * head_size + d + 1]) * inv;
        val.z = (part[d + 2] + part[head_size + d + 2] + part[2 * head_size + d + 2] + part[3 * head_size + d + 2]) * inv;
        val.w = (part[d + 3] + part[head_size + d + 3] + part[2 * head_size + d + 3] + part[3 * head_size + d + 3]) * inv;
      }
      for (int r = 0; r < world; ++r)
        reinterpret_cast<float4*>(reinterpret_cast<float*>(peers.buf[r] + hdr) + row_off)[v4] = val;
    }
    __syncthreads();  // part[] is reused by the next pair
  }
  __threadfence_system();
  __syncthreads();
  if ((phases & 1) && threadIdx.x < world) {
    uint32_t* f = reinterpret_cast<uint32_t*>(peers.buf[threadIdx.x]) + rank * kCtxMaxCtas + blockIdx.x;
    st_release_sys(f, epoch);
  }

  // ---- phase 2: wait for every rank's publication of this epoch, merge ----
  if (!(phases & 2)) return;
  unsigned char* mine = peers.buf[rank];
  if (threadIdx.x < world) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(mine) + threadIdx.x * kCtxMaxCtas + blockIdx.x;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if (clock64() - t0 > timeout_cycles) {
        unsigned int* st = reinterpret_cast<unsigned int*>(mine) + world * kCtxMaxCtas;
        if (atomicCAS(st, 0u, 1u + threadIdx.x) == 0u) st[3] = epoch;  // first failure: who and when
        timed_out_s = 1;
        break;
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
  const float* data = reinterpret_cast<const float*>(mine + hdr) + slot_off;
  const bool poisoned = timed_out_s != 0;
  for (int pair = blockIdx.x; pair < total_pairs; pair += gridDim.x) {
    if (poisoned) {
      // A peer never published this epoch: its slot is stale or half written.  Fail loudly in the
      // data -- NaN rows -- instead of merging it (the status word says who and which epoch).
      T* o = out + (int64_t)pair * head_size;
      for (int d = threadIdx.x; d < head_size; d += blockDim.x) o[d] = from_f32<T>(__int_as_float(0x7fc00000));
      continue;
    }
    float M = kNegInf;
    for (int r = 0; r < world; ++r) {
      const float2 st = __ldcg(reinterpret_cast<const float2*>(data + ((size_t)r * max_pairs + pair) * row_floats));
      if (st.y > 0.f) M = fmaxf(M, st.x);
    }
    float S = 0.f;
    for (int r = 0; r < world; ++r) {
      const float2 st = __ldcg(reinterpret_cast<const float2*>(data + ((size_t)r * max_pairs + pair) * row_floats));
      if (st.y > 0.f) S += st.y * expf(st.x - M);
    }
    const float inv = __fdividef(1.f, S + 1e-6f);
    T* o = out + (int64_t)pair * head_size;
    for (int d = threadIdx.x; d < head_size; d += blockDim.x) {
      float acc = 0.f;
      for (int r = 0; r < world; ++r) {
        const float* row = data + ((size_t)r * max_pairs + pair) * row_floats;
        const float2 st = __ldcg(reinterpret_cast<const float2*>(row));
        if (st.y > 0.f) acc += __ldcg(row + kCtxRowPad + d) * (st.y * expf(st.x - M) * inv);
      }
      o[d] = from_f32<T>(acc);
    }
  }
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
  B200_CHECK_ARG(head_size % 4 == 0, "ctx_shard: head size must be a multiple of 4");
  B200_CHECK_ARG(phases >= 1 && phases <= 3, "ctx_shard: phases must be 1 (publish), 2 (merge) or 3");
  CtxPeers peers{};
  for (int r = 0; r < world; ++r) {
    B200_CHECK_ARG(exchange_bufs && exchange_bufs[r], "ctx_shard: exchange buffer of rank %d is null", r);
    peers.buf[r] = static_cast<unsigned char*>(exchange_bufs[r]);
  }
  int blocks = (int)pairs;
  const int cap = 148;  // same on every rank (flags are per CTA); all CTAs co-resident
  if (blocks > cap) blocks = cap;
  const int wstride = head_size;
  const size_t wsmem = (size_t)4 * head_size * sizeof(float);
  (void)parts;
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
#define B200_CTX_LAUNCH(T)                                                                                   \
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, ctx_exchange_merge_kernel<T>, static_cast<T*>(out), exp_sums,         \
                                  max_logits, static_cast<const T*>(tmp_out), seq_lens, num_heads, head_size, \
                                  max_num_partitions, partition_tokens, wstride, pairs_i, peers, rank, world, \
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
  return (int64_t)(ctx_header_bytes(world) +
                   (size_t)2 * world * max_pairs * (head_size + kCtxRowPad) * sizeof(float));
}

// Peer-visible device memory.  cudaMalloc (not a pool / VMM allocation) so that it can be exported
// with cudaIpcGetMemHandle; zero-filled so that epoch 0 flags read "nothing published".
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
                                                 (size_t)world * kCtxMaxCtas * 4,
                               16, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return B200_OK;
}

// Per-group wait timeout (milliseconds at ~2 GHz; 0 restores the 2 s default) and status reset.
// State lives in the group's own exchange buffer, not in the library.
extern "C" int b200_ctx_exchange_set_timeout(void* local_buf, int world, int timeout_ms, int device) {
  B200_CHECK_ARG(local_buf && timeout_ms >= 0, "ctx_exchange_set_timeout: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  const uint32_t ms = (uint32_t)timeout_ms;
  B200_CUDA_OK(cudaMemcpy(static_cast<unsigned char*>(local_buf) + ((size_t)world * kCtxMaxCtas + 4) * 4,
                          &ms, 4, cudaMemcpyHostToDevice));
  return B200_OK;
}
extern "C" int b200_ctx_exchange_clear_status(void* local_buf, int world, int device) {
  B200_CHECK_ARG(local_buf, "ctx_exchange_clear_status: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  const uint32_t z = 0;
  unsigned char* st = static_cast<unsigned char*>(local_buf) + (size_t)world * kCtxMaxCtas * 4;
  B200_CUDA_OK(cudaMemcpy(st, &z, 4, cudaMemcpyHostToDevice));
  B200_CUDA_OK(cudaMemcpy(st + 12, &z, 4, cudaMemcpyHostToDevice));
  return B200_OK;
}

// Synchronous read of the exchange buffer's timeout word (tests / diagnostics; not on the hot path).
extern "C" int b200_ctx_exchange_status(const void* local_buf, int world, int32_t* status, int device) {
  B200_CHECK_ARG(local_buf && status, "ctx_exchange_status: bad arguments");
  DeviceGuard guard(device);
  B200_CUDA_OK(guard.err);
  B200_CUDA_OK(cudaMemcpy(status, static_cast<const unsigned char*>(local_buf) +
                                       (size_t)world * kCtxMaxCtas * 4,
                          4, cudaMemcpyDeviceToHost));
  return B200_OK;
}
