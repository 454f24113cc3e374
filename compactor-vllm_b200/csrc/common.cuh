// Shared host/device helpers for libcompactor_sm100.so (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/cvllm.h"

namespace cvllm {

// ---------------------------------------------------------------------------
// host-side error plumbing
// ---------------------------------------------------------------------------
void set_last_error(const char* msg);

#define CVLLM_CHECK_ARG(cond)                                                  \
  do {                                                                         \
    if (!(cond)) {                                                             \
      ::cvllm::set_last_error("argument check failed: " #cond);                \
      return CVLLM_E_ARG;                                                      \
    }                                                                          \
  } while (0)

#define CVLLM_CHECK_SUPPORTED(cond)                                            \
  do {                                                                         \
    if (!(cond)) {                                                             \
      ::cvllm::set_last_error("unsupported configuration: " #cond);            \
      return CVLLM_E_UNSUPPORTED;                                              \
    }                                                                          \
  } while (0)

#define CVLLM_CHECK_ALIGN(cond)                                                \
  do {                                                                         \
    if (!(cond)) {                                                             \
      ::cvllm::set_last_error("alignment check failed: " #cond);               \
      return CVLLM_E_ALIGN;                                                    \
    }                                                                          \
  } while (0)

#define CVLLM_CHECK_CUDA(expr)                                                 \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess) {                                                   \
      ::cvllm::set_last_error(cudaGetErrorString(_e));                         \
      return CVLLM_E_CUDA;                                                     \
    }                                                                          \
  } while (0)

// every kernel launch of the library is counted (cvllm_launch_count) so callers can state how
// many of OUR kernels ran inside a timed region
void note_launches(int n);
#define CVLLM_CHECK_LAUNCHES(n)                                                \
  do {                                                                         \
    ::cvllm::note_launches(n);                                                 \
    CVLLM_CHECK_CUDA(cudaPeekAtLastError());                                   \
  } while (0)
#define CVLLM_CHECK_LAUNCH() CVLLM_CHECK_LAUNCHES(1)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE property of a kernel: raise it once per
// (kernel, device), so one process can drive several GPUs.  `mask` (one static per call site) has a
// bit per device ordinal.  Not a stream operation: safe under CUDA-graph capture.
inline cudaError_t ensure_max_dyn_smem(const void* func, int bytes, std::atomic<uint64_t>& mask) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const uint64_t bit = 1ull << (dev & 63);
  if (mask.load(std::memory_order_acquire) & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) mask.fetch_or(bit, std::memory_order_release);
  return e;
}

// Test / profiling switches, set only through cvllm_set_test_option (include/cvllm.h); the product path
// never reads the environment.
struct TestOptions {
  std::atomic<int> prefill_v1{0};             // 1: force the one-head-per-CTA prefill kernel
  std::atomic<float> prefill_threshold{-1.f}; // >= 0: lazy-rescale threshold of the prefill softmax
  std::atomic<int> decode_mode{0};            // decode_attn.cu profiling modes (0 = product)
  std::atomic<int> decode_slots{0};           // >= 2: cap the decode TMA ring depth
  std::atomic<int> eig_trace{0};
  std::atomic<int> snap_v1{0};                // 1: the mma.sync SnapKV kernels
  std::atomic<int> chunk_scores_v1{0};        // 1: the mma.sync chunk scorer (A/B against the tcgen05 one)
  std::atomic<int> pdl{0};                    // 1: launch the selection chain with programmatic dependent launch (A/B; off in the product)
  std::atomic<int> grs_pairs{0};              // > 0: cap the CTA pairs of gemm_reduce (several virtual ranks on one GPU)
  std::atomic<int> grs_probe{0};              // gemm_reduce timing probes (bit mask; results are wrong)
  std::atomic<uint64_t> grs_debug_host_ptr{0};  // pinned host buffer for gemm_reduce's timeout records
  std::atomic<int> grs_rows_per_round{0};     // > 0: rows per reduce round of gemm_reduce (multiple of 256)
};
TestOptions& test_options();

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int sm_count();
int launch_decode_store(const void* key, const void* value, int64_t sk_b, int64_t sk_h, int64_t sv_b,
                        int64_t sv_h, const int32_t* batch_mapping, int32_t* bh_lens,
                        const int32_t* page_table, void* k_cache, void* v_cache, int B, int HKV, int D,
                        int PS, int NLP, int elem_bytes, int reserved_batch, int lens_by_slot,
                        cvllm_stream_t stream);
uint64_t launches();

// ---------------------------------------------------------------------------
// Programmatic dependent launch (a chain of short kernels on one stream): `launch_pdl` lets the NEXT kernel of the
// stream be scheduled while this one is still running; a kernel launched that way calls `pdl_begin()` as its first
// statement, which (a) allows ITS successor to be scheduled and (b) waits until every earlier kernel of the stream
// has completed and its writes are visible.  Ordering and results are those of plain stream order.
// OFF by default (test option "pdl" = 1 turns it on): measured on the selection chain it takes 15-19 us off the eager
// chain run alone and nothing off its CUDA-graph replay, but INSIDE the prefill step — where every SM is held by an
// attention CTA that uses all its shared memory and registers — the early-scheduled CTAs sit on SMs while they wait,
// and the step gets 1-2 % slower (tools/ab_pdl.py, tools/ab_overlap.py).  Without the attribute `pdl_begin()` is a no-op.
// ---------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_begin() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = test_options().pdl.load() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------
__host__ __device__ __forceinline__ int cdiv(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ void stg_na_v4(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// float pair -> packed 16-bit pair (lo = a, hi = b)
template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  uint32_t r;
  if constexpr (BF16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  } else {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  }
  return r;
}

template <bool BF16>
__device__ __forceinline__ float2 unpack2(uint32_t v) {
  if constexpr (BF16) {
    return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
  } else {
    __half2 h = *reinterpret_cast<__half2*>(&v);
    return __half22float2(h);
  }
}

// order-preserving fp32 -> uint32 key; NaN ranks above +inf (torch.topk order)
__host__ __device__ __forceinline__ uint32_t score_key(uint32_t bits) {
  if ((bits & 0x7fffffffu) > 0x7f800000u) return 0xffffffffu;  // NaN
  return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}

}  // namespace cvllm
