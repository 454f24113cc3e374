// Host-side runtime glue: error string, device properties, TMA descriptor
// encoding through the driver entry point (no link-time libcuda dependency).
#include <atomic>
#include <mutex>
#include <string.h>

#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

static thread_local char g_last_error[256] = "";

void set_last_error(const char* msg) {
  strncpy(g_last_error, msg ? msg : "", sizeof(g_last_error) - 1);
  g_last_error[sizeof(g_last_error) - 1] = 0;
}

static std::atomic<uint64_t> g_launches{0};
void note_launches(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
uint64_t launches() { return g_launches.load(std::memory_order_relaxed); }

TestOptions& test_options() {
  static TestOptions opts;
  return opts;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

// 16-bit elements, rank <= 3, inner box = 64 elements (128 B) with 128B swizzle.
int make_tmap_16b(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes /* rank-1 entries */, const uint32_t* box,
                  bool bf16) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled driver entry point unavailable");
    return CVLLM_E_CUDA;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i + 1 < rank) gstr[i] = strides_bytes[i];
  }
  CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                   (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[96];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed: CUresult %d", (int)r);
    set_last_error(buf);
    return CVLLM_E_CUDA;
  }
  return CVLLM_OK;
}

}  // namespace cvllm

extern "C" int cvllm_abi_version(void) { return CVLLM_ABI_VERSION; }
extern "C" const char* cvllm_last_error(void) { return cvllm::g_last_error; }
extern "C" uint64_t cvllm_launch_count(void) { return cvllm::launches(); }

extern "C" int cvllm_set_test_option(const char* name, double value) {
  if (!name) return CVLLM_E_ARG;
  auto& o = cvllm::test_options();
  if (!strcmp(name, "prefill_v1")) o.prefill_v1 = (int)value;
  else if (!strcmp(name, "prefill_threshold")) o.prefill_threshold = (float)value;
  else if (!strcmp(name, "decode_mode")) o.decode_mode = (int)value;
  else if (!strcmp(name, "decode_slots")) o.decode_slots = (int)value;
  else if (!strcmp(name, "chunk_scores_v1")) o.chunk_scores_v1 = (int)value;
  else if (!strcmp(name, "snap_v1")) o.snap_v1 = (int)value;
  else if (!strcmp(name, "pdl")) o.pdl = (int)value;
  else if (!strcmp(name, "grs_pairs")) o.grs_pairs = (int)value;
  else if (!strcmp(name, "grs_probe")) o.grs_probe = (int)value;
  else if (!strcmp(name, "grs_debug_host_ptr")) o.grs_debug_host_ptr = (uint64_t)value;
  else if (!strcmp(name, "grs_rows_per_round")) o.grs_rows_per_round = (int)value;
  else {
    cvllm::set_last_error("unknown test option");
    return CVLLM_E_ARG;
  }
  return CVLLM_OK;
}
