// NVLink peer-WRITE bandwidth by store method (one process, two GPUs, cudaDeviceEnablePeerAccess): what the epilogue
// and the reducers of csrc/gemm_reduce.cu can expect from the wire.  GPU 0 writes 512 MiB into GPU 1's memory with
//   A  st.global.v4, a warp writes 512 contiguous bytes per instruction               (the reducers' all-gather)
//   B  TMA tensor store of [128 x 64] 16-bit boxes (128-byte rows, 16 KB row pitch)   (the epilogue today)
//   C  cp.async.bulk.global.shared::cta of 16 KB contiguous                           (a tile-major slot layout)
// and, for reference, the same three into GPU 0's own memory.  148 CTAs, up to `depth` bulk groups in flight per CTA.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I compactor-vllm_b200/csrc -o tools/ubench/peer_store_bw tools/ubench/peer_store_bw.cu
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include "tma.cuh"

using namespace cvllm;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int kRows = 32768, kCols = 8192;  // 16-bit elements: 512 MiB
constexpr size_t kBytes = (size_t)kRows * kCols * 2;
constexpr int kBoxBytes = 128 * 128;

__global__ void __launch_bounds__(256) k_st(uint4* dst, size_t n_vec) {
  const uint4 v = make_uint4(threadIdx.x, blockIdx.x, 3, 4);
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n_vec; i += (size_t)gridDim.x * 256)
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int DEPTH>
__global__ void __launch_bounds__(128) k_tma(const __grid_constant__ CUtensorMap map, int n_boxes) {
  extern __shared__ __align__(1024) unsigned char raw[];
  unsigned char* buf = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  for (int i = threadIdx.x; i < DEPTH * kBoxBytes / 16; i += 128) reinterpret_cast<uint4*>(buf)[i] = make_uint4(i, 1, 2, 3);
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = 0;
    for (int b = blockIdx.x; b < n_boxes; b += gridDim.x, ++n) {
      const int row0 = (b / (kCols / 64)) * 128, col0 = (b % (kCols / 64)) * 64;
      bulk_wait_read<DEPTH - 1>();
      tma_store_2d(&map, buf + (n % DEPTH) * kBoxBytes, col0, row0);
      bulk_commit();
    }
    bulk_wait<0>();
  }
}

template <int DEPTH>
__global__ void __launch_bounds__(128) k_bulk(unsigned char* dst, int n_boxes) {
  extern __shared__ __align__(1024) unsigned char raw[];
  unsigned char* buf = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  for (int i = threadIdx.x; i < DEPTH * kBoxBytes / 16; i += 128) reinterpret_cast<uint4*>(buf)[i] = make_uint4(i, 1, 2, 3);
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = 0;
    for (int b = blockIdx.x; b < n_boxes; b += gridDim.x, ++n) {
      bulk_wait_read<DEPTH - 1>();
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (size_t)b * kBoxBytes),
                   "r"(smem_u32(buf + (n % DEPTH) * kBoxBytes)), "r"(kBoxBytes) : "memory");
      bulk_commit();
    }
    bulk_wait<0>();
  }
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                            const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(void* base) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  CUtensorMap m;
  cuuint64_t dims[2] = {(cuuint64_t)kCols, (cuuint64_t)kRows}, str[1] = {(cuuint64_t)kCols * 2};
  cuuint32_t box[2] = {64, 128}, es[2] = {1, 1};
  CUresult r = ((PFN_enc)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

template <class F>
static double timed_gbs(F launch) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  for (int i = 0; i < 2; ++i) launch();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < 5; ++i) launch();
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  CK(cudaGetLastError());
  return kBytes * 5.0 / (ms * 1e-3) / 1e9;
}

int main() {
  int n_dev = 0;
  CK(cudaGetDeviceCount(&n_dev));
  void* local = nullptr; void* peer = nullptr;
  CK(cudaSetDevice(0));
  CK(cudaMalloc(&local, kBytes));
  if (n_dev >= 2) {
    CK(cudaSetDevice(1));
    CK(cudaMalloc(&peer, kBytes));
    CK(cudaSetDevice(0));
    CK(cudaDeviceEnablePeerAccess(1, 0));
  }
  const int n_boxes = (int)(kBytes / kBoxBytes);
  const int smem = 4 * kBoxBytes + 1024;
  CK(cudaFuncSetAttribute(k_tma<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_tma<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_bulk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_bulk<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int t = 0; t < 2; ++t) {
    void* dst = t == 0 ? local : peer;
    if (!dst) { printf("one GPU only: peer rows skipped\n"); break; }
    const CUtensorMap map = make_map(dst);
    printf("%s:\n", t == 0 ? "into OWN memory" : "into PEER memory over NVLink");
    printf("  A st.global.v4 (512 B / warp instr), 148 x 8 CTAs : %7.1f GB/s\n",
           timed_gbs([&] { k_st<<<148 * 8, 256>>>((uint4*)dst, kBytes / 16); }));
    printf("  B TMA tensor store 128 x 64 boxes, depth 2         : %7.1f GB/s\n",
           timed_gbs([&] { k_tma<2><<<148, 128, smem>>>(map, n_boxes); }));
    printf("  B TMA tensor store 128 x 64 boxes, depth 4         : %7.1f GB/s\n",
           timed_gbs([&] { k_tma<4><<<148, 128, smem>>>(map, n_boxes); }));
    printf("  C bulk copy 16 KB contiguous, depth 2              : %7.1f GB/s\n",
           timed_gbs([&] { k_bulk<2><<<148, 128, smem>>>((unsigned char*)dst, n_boxes); }));
    printf("  C bulk copy 16 KB contiguous, depth 4              : %7.1f GB/s\n",
           timed_gbs([&] { k_bulk<4><<<148, 128, smem>>>((unsigned char*)dst, n_boxes); }));
  }
  return 0;
}
