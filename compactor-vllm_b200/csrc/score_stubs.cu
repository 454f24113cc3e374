// Temporary: scorer entry points not implemented yet (fail loudly).
#include "common.cuh"
using namespace cvllm;
extern "C" size_t cvllm_leverage_scores_workspace_bytes(int, int, int, int) { return 256; }
extern "C" int cvllm_leverage_scores(const void*, int64_t, int64_t, const int32_t*, int, int, const void*,
                                     void*, int, int, int, int, float, int, int, void*, size_t,
                                     cvllm_stream_t) {
  set_last_error("cvllm_leverage_scores: not implemented");
  return CVLLM_E_UNSUPPORTED;
}
extern "C" size_t cvllm_chunk_attn_scores_workspace_bytes(int, int, int, int) { return 256; }
extern "C" int cvllm_chunk_attn_scores(const void*, const void*, int64_t, int64_t, const int32_t*, int, int,
                                       int, int, int, int, int, float, int, const void*, int, int64_t, float,
                                       const int32_t*, const int32_t*, float*, int, void*, size_t,
                                       cvllm_stream_t) {
  set_last_error("cvllm_chunk_attn_scores: not implemented");
  return CVLLM_E_UNSUPPORTED;
}
extern "C" size_t cvllm_snapkv_scores_workspace_bytes(int, int, int, int, int) { return 256; }
extern "C" int cvllm_snapkv_scores(const void*, const void*, int64_t, int64_t, const int32_t*, const int32_t*,
                                   int, int, int, int, int, int, int, float, int, int, float*, int, void*,
                                   size_t, cvllm_stream_t) {
  set_last_error("cvllm_snapkv_scores: not implemented");
  return CVLLM_E_UNSUPPORTED;
}
