// One-shot all-reduce over NVLink peer memory for the tensor-parallel o_proj / down_proj outputs of a
// decode step (reference: layers/linear.py:149-153 `dist.all_reduce` after the row-parallel
// GEMM).  A decode-size message ([B, hidden] bf16: 16 KB .. 1 MB) is latency-bound: NCCL's ring / tree
// costs several hops and a proxy wake-up; here every rank reads all peers' buffers directly through
// NVSwitch (900 GB/s per direction, any-to-any) and sums them in rank order in fp32, so every rank
// produces the same bits.  Prefill-size messages stay on NCCL (bandwidth-bound, NVLS in-switch reduce).
//
// Buffers are cudaMalloc'd here and shared with cudaIpc handles (exchanged by the caller through
// torch.distributed); layout of a rank's buffer: [ epochs + arrival words | payload ].
// Barrier per block (a block starts reading as soon as the same block of every peer has arrived): one remote
// store of a device-side epoch + a spin on local memory; a second one after the reads lets the next call
// overwrite the payload.  No host-side counters, so a captured call replays unchanged.
#include <string.h>

#include "common.cuh"

namespace cvllm {

constexpr int kArMaxWorld = 8;
constexpr int kArBlocks = 64;
constexpr int kArThreads = 512;
constexpr int kArPhases = 3;  // one-shot uses phases 0-1, two-shot 0-2
constexpr size_t kArFlagBytes = (kArBlocks + kArPhases * kArBlocks * kArMaxWorld) * sizeof(uint32_t) + 256;  // 6.5 KB, payload 256-B aligned

struct ArPeers {
  void* base[kArMaxWorld];  // base[r] = rank r's buffer as mapped in THIS process
};

// Memory model.  Every barrier that GUARDS DATA a peer GPU reads is a system-scope release / acquire pair:
//   signaller: __syncthreads (block's prior accesses happen-before thread t) -> fence.acq_rel.sys -> st.relaxed.sys flag
//   waiter   : ld.acquire.sys poll -> __syncthreads -> the block's loads of the peer's data
// This covers (a) the start barrier: the payload was written by EARLIER kernels of the producer's stream, which the
// kernel boundary orders before this kernel's fence, so the release is cumulative over it; (b) the two-shot publish
// barrier: slices written inside this kernel; (c) the end barrier: the peer's loads of my payload are ordered before
// its flag store by the same fence, and my acquire orders the next call's overwrite after them (write-after-read).
// The round-1 form (volatile flags, gpu-scope fence on publish only: 1-3 us faster per call, correct on this
// hardware in every stress run but a data race under the PTX model) is kept behind -DCVLLM_AR_RELAXED_FLAGS for A/B.
__device__ __forceinline__ void st_flag(uint32_t* p, uint32_t v) {
#ifdef CVLLM_AR_RELAXED_FLAGS
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#else
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#endif
}
__device__ __forceinline__ uint32_t ld_flag(const uint32_t* p) {
  uint32_t v;
#ifdef CVLLM_AR_RELAXED_FLAGS
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
#else
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
#endif
  return v;
}

// Flag words of a rank's buffer: epoch[kArBlocks] | arrive[kArPhases][kArBlocks][8].  A block's epoch counts its
// calls; ranks run the same launches in the same order, so epochs agree without any host-side counter (a CUDA
// graph replays the same kernel arguments).  Signal = one remote store of the epoch into the peer's word for
// (phase, block, me); wait = spin on my own words until they reach it.  A peer can be at most one call ahead,
// hence the >= comparison.  PUBLISH marks the barrier behind which peers read data written by THIS kernel (only
// the relaxed-flags build treats it differently).
template <bool PUBLISH = false>
__device__ __forceinline__ void ar_barrier(const ArPeers& peers, int rank, int world, int phase, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < world) {
    const int peer = threadIdx.x;
    const int word = kArBlocks + (phase * kArBlocks + blockIdx.x) * kArMaxWorld;
#ifdef CVLLM_AR_RELAXED_FLAGS
    if (PUBLISH) __threadfence();
#else
    asm volatile("fence.acq_rel.sys;" ::: "memory");
#endif
    st_flag(reinterpret_cast<uint32_t*>(peers.base[peer]) + word + rank, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(peers.base[rank]) + word + peer;
    // bounded: a peer that never arrives (crashed rank, mismatched launch order) traps this context after ~1 min
    // of polling instead of hanging the GPU until the driver's watchdog
    uint32_t spins = 0;
    while ((int32_t)(ld_flag(mine) - epoch) < 0)
      if (++spins > (1u << 27)) __trap();
  }
  __syncthreads();
}

__device__ __forceinline__ uint4 ld_volatile_v4(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <bool BF16>
__global__ void __launch_bounds__(kArThreads) allreduce_oneshot_kernel(ArPeers peers, void* __restrict__ out,
                                                                       int64_t n_vec, int rank, int world) {
  uint32_t* my_epoch = reinterpret_cast<uint32_t*>(peers.base[rank]) + blockIdx.x;
  const uint32_t epoch = *my_epoch + 1;
  ar_barrier(peers, rank, world, 0, epoch);
  const int64_t stride = (int64_t)gridDim.x * kArThreads;
  for (int64_t i = (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < n_vec; i += stride) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    uint4 v[kArMaxWorld];
#pragma unroll
    for (int r = 0; r < kArMaxWorld; ++r)
      if (r < world) v[r] = ld_volatile_v4(reinterpret_cast<const char*>(peers.base[r]) + kArFlagBytes + i * 16);
#pragma unroll
    for (int r = 0; r < kArMaxWorld; ++r) {
      if (r < world) {  // rank order on every rank: identical bits everywhere
        const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = unpack2<BF16>(w[e]);
          acc[2 * e] += f.x;
          acc[2 * e + 1] += f.y;
        }
      }
    }
    uint4 o;
    o.x = pack2<BF16>(acc[0], acc[1]); o.y = pack2<BF16>(acc[2], acc[3]);
    o.z = pack2<BF16>(acc[4], acc[5]); o.w = pack2<BF16>(acc[6], acc[7]);
    reinterpret_cast<uint4*>(out)[i] = o;
  }
  ar_barrier(peers, rank, world, 1, epoch);  // every peer has read my payload: the next call may overwrite it
  if (threadIdx.x == 0) *my_epoch = epoch;
}

// Two-shot: rank r reduces slice r of the message (reads 1/world of every peer's payload), leaves it in the RESULT
// area of its own buffer, and after a publishing barrier every rank gathers the world slices from their owners.
// Per rank it moves 2 x (world-1)/world x message over NVLink instead of the one-shot's (world-1) x message: the
// form for messages of a few MB at 4-8 ranks.  Same rank-order fp32 sum, so the bits equal the one-shot's.
// Block b works on the same sub-range of every slice on every rank, so the per-block barriers suffice.
template <bool BF16>
__global__ void __launch_bounds__(kArThreads) allreduce_twoshot_kernel(ArPeers peers, void* __restrict__ out,
                                                                       int64_t n_vec, int64_t result_off, int rank,
                                                                       int world) {
  uint32_t* my_epoch = reinterpret_cast<uint32_t*>(peers.base[rank]) + blockIdx.x;
  const uint32_t epoch = *my_epoch + 1;
  const int64_t slice = (n_vec + world - 1) / world;
  const int64_t stride = (int64_t)gridDim.x * kArThreads;
  ar_barrier(peers, rank, world, 0, epoch);  // every peer's payload is there (written by earlier kernels)
  {
    const int64_t lo = rank * slice, hi = min(n_vec, lo + slice);
    uint4* res = reinterpret_cast<uint4*>(reinterpret_cast<char*>(peers.base[rank]) + kArFlagBytes + result_off);
    for (int64_t i = lo + (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < hi; i += stride) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      uint4 v[kArMaxWorld];
#pragma unroll
      for (int r = 0; r < kArMaxWorld; ++r)
        if (r < world) v[r] = ld_volatile_v4(reinterpret_cast<const char*>(peers.base[r]) + kArFlagBytes + i * 16);
#pragma unroll
      for (int r = 0; r < kArMaxWorld; ++r) {
        if (r < world) {
          const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = unpack2<BF16>(w[e]);
            acc[2 * e] += f.x;
            acc[2 * e + 1] += f.y;
          }
        }
      }
      uint4 o;
      o.x = pack2<BF16>(acc[0], acc[1]); o.y = pack2<BF16>(acc[2], acc[3]);
      o.z = pack2<BF16>(acc[4], acc[5]); o.w = pack2<BF16>(acc[6], acc[7]);
      res[i] = o;
    }
  }
  ar_barrier<true>(peers, rank, world, 1, epoch);  // slices published
  // gather: the loads from all owners are issued before the first store — one NVLink round trip per trip of the
  // loop, not `world` dependent ones (the serial form cost ~12 us at 8 ranks whatever the message size)
  for (int64_t j = (int64_t)blockIdx.x * kArThreads + threadIdx.x; j < slice; j += stride) {
    uint4 v[kArMaxWorld];
#pragma unroll
    for (int r = 0; r < kArMaxWorld; ++r) {
      const int64_t i = r * slice + j;
      if (r < world && i < n_vec)
        v[r] = ld_volatile_v4(reinterpret_cast<const char*>(peers.base[r]) + kArFlagBytes + result_off + i * 16);
    }
#pragma unroll
    for (int r = 0; r < kArMaxWorld; ++r) {
      const int64_t i = r * slice + j;
      if (r < world && i < n_vec) reinterpret_cast<uint4*>(out)[i] = v[r];
    }
  }
  ar_barrier(peers, rank, world, 2, epoch);  // every peer has read my slice: payload and result may be rewritten
  if (threadIdx.x == 0) *my_epoch = epoch;
}

}  // namespace cvllm

using namespace cvllm;

extern "C" size_t cvllm_allreduce_flag_bytes(void) { return kArFlagBytes; }

// cudaMalloc + zeroed flags + IPC handle (64 bytes written to handle_out_host).  Returns the device pointer.
extern "C" int cvllm_ipc_alloc(size_t payload_bytes, void** ptr_out_host, void* handle_out_host) {
  CVLLM_CHECK_ARG(ptr_out_host && handle_out_host && payload_bytes > 0);
  void* p = nullptr;
  CVLLM_CHECK_CUDA(cudaMalloc(&p, kArFlagBytes + payload_bytes));
  CVLLM_CHECK_CUDA(cudaMemset(p, 0, kArFlagBytes));
  cudaIpcMemHandle_t h;
  CVLLM_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle_out_host, &h, sizeof(h));
  CVLLM_CHECK_CUDA(cudaDeviceSynchronize());
  *ptr_out_host = p;
  return CVLLM_OK;
}

extern "C" int cvllm_ipc_open(const void* handle_host, void** ptr_out_host) {
  CVLLM_CHECK_ARG(handle_host && ptr_out_host);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle_host, sizeof(h));
  CVLLM_CHECK_CUDA(cudaIpcOpenMemHandle(ptr_out_host, h, cudaIpcMemLazyEnablePeerAccess));
  return CVLLM_OK;
}

extern "C" int cvllm_ipc_close(void* peer_ptr) {
  CVLLM_CHECK_CUDA(cudaIpcCloseMemHandle(peer_ptr));
  return CVLLM_OK;
}

extern "C" int cvllm_ipc_free(void* own_ptr) {
  CVLLM_CHECK_CUDA(cudaFree(own_ptr));
  return CVLLM_OK;
}

// out[i] = sum_r payload_r[i], i < numel (16-bit elements, numel % 8 == 0).  peer_bases_host[r] = rank r's buffer
// as mapped in this process (own buffer for r == rank).  Every rank must launch with the same numel.
extern "C" int cvllm_allreduce_oneshot(void* const* peer_bases_host, int rank, int world, void* out, int64_t numel,
                                       int dtype, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(peer_bases_host && out && world >= 1 && world <= kArMaxWorld && rank >= 0 && rank < world);
  CVLLM_CHECK_ARG(numel >= 0 && numel % 8 == 0);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_ALIGN(aligned16(out));
  if (numel == 0) return CVLLM_OK;
  ArPeers peers;
  for (int r = 0; r < kArMaxWorld; ++r) peers.base[r] = r < world ? peer_bases_host[r] : nullptr;
  const int64_t n_vec = numel / 8;
  int blocks = (int)((n_vec + kArThreads - 1) / kArThreads);
  blocks = blocks < 1 ? 1 : (blocks > kArBlocks ? kArBlocks : blocks);
  auto st = static_cast<cudaStream_t>(stream);
  if (dtype == CVLLM_BF16)
    allreduce_oneshot_kernel<true><<<blocks, kArThreads, 0, st>>>(peers, out, n_vec, rank, world);
  else
    allreduce_oneshot_kernel<false><<<blocks, kArThreads, 0, st>>>(peers, out, n_vec, rank, world);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

// Two-shot form of the same sum for messages of a few MB: result_offset_bytes = distance from the payload start to a
// result area of at least numel * 2 bytes inside every rank's buffer (same offset everywhere; the buffers must have
// been allocated with room for it).  Same bits as cvllm_allreduce_oneshot.
extern "C" int cvllm_allreduce_twoshot(void* const* peer_bases_host, int rank, int world, void* out, int64_t numel,
                                       int dtype, int64_t result_offset_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(peer_bases_host && out && world >= 1 && world <= kArMaxWorld && rank >= 0 && rank < world);
  CVLLM_CHECK_ARG(numel >= 0 && numel % 8 == 0);
  CVLLM_CHECK_ARG(result_offset_bytes >= numel * 2 && result_offset_bytes % 16 == 0);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_ALIGN(aligned16(out));
  if (numel == 0) return CVLLM_OK;
  ArPeers peers;
  for (int r = 0; r < kArMaxWorld; ++r) peers.base[r] = r < world ? peer_bases_host[r] : nullptr;
  const int64_t n_vec = numel / 8;
  const int64_t slice = (n_vec + world - 1) / world;
  int blocks = (int)((slice + kArThreads - 1) / kArThreads);
  blocks = blocks < 1 ? 1 : (blocks > kArBlocks ? kArBlocks : blocks);
  auto st = static_cast<cudaStream_t>(stream);
  if (dtype == CVLLM_BF16)
    allreduce_twoshot_kernel<true><<<blocks, kArThreads, 0, st>>>(peers, out, n_vec, result_offset_bytes, rank, world);
  else
    allreduce_twoshot_kernel<false><<<blocks, kArThreads, 0, st>>>(peers, out, n_vec, result_offset_bytes, rank, world);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
