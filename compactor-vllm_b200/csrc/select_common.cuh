// Helpers shared by the ranking / index-store path (select.cu) and the fused multi-CTA
// select + compaction path (select_store.cu).
#pragma once
#include "common.cuh"

namespace cvllm {

constexpr int kSelThreads = 1024;

__device__ __forceinline__ uint32_t load_key(const float* scores, int64_t stride_n, int n, int h) {
  return score_key(__float_as_uint(__ldg(scores + (int64_t)n * stride_n + h)));
}

// CTA-wide exclusive prefix of a 0/1 flag in thread order; total returned through `total`.
__device__ __forceinline__ int block_excl_count(bool flag, int* s_warp /*[32]*/, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t m = __ballot_sync(0xffffffffu, flag);
  if (lane == 0) s_warp[warp] = __popc(m);
  __syncthreads();
  int before = 0, tot = 0;
  const int nw = blockDim.x >> 5;
  for (int w = 0; w < nw; ++w) {
    const int c = s_warp[w];
    if (w < warp) before += c;
    tot += c;
  }
  __syncthreads();
  total = tot;
  return before + __popc(m & ((1u << lane) - 1u));
}

// pad rule of store_kv_cache.py:205-248 in closed form
__device__ __forceinline__ int pad_amount(int L_after_topk, int ctx_len, int PS) {
  const int mod = L_after_topk % PS;
  if (mod == 0) return 0;
  const int need = PS - mod;
  const int max_add = ctx_len - L_after_topk;
  return max(0, min(need, max_add));
}

}  // namespace cvllm

// host: row geometry / alignment checks shared by the store entry points
int cvllm_check_store_args(const void* keys, const void* vals, int64_t sk_n, int64_t sk_h, int64_t sv_n,
                           int64_t sv_h, void* k_cache, void* v_cache, int D, int elem_bytes, int H,
                           int* vpr_out, int64_t* per16_out);
