// Bring-up of tcgen05.mma.cta_group::2 (a CTA pair on one TPC): D[256 x 128] = A[256 x 128] . B, each CTA of the pair
// holding its own 128 rows of A and D and HALF of B's N dimension in shared memory.
//   mode 0 (the S = Q.K^T shape): A from smem K-major, B = K tile [128 keys x 128 d] K-major, CTA r holds keys 64r..64r+63
//   mode 1 (the O = P.V shape):   A from TMEM (16-bit P rows), B = V tile [128 keys x 128 d] MN-major, CTA r holds d 64r..64r+63
// Checks every accumulator entry against an fp32 dot product on the host.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I compactor-vllm_b200/csrc -o tools/ubench/umma_2cta tools/ubench/umma_2cta.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "tma.cuh"

using namespace cvllm;

struct __align__(1024) Smem {
  uint8_t a[128 * 256];  // A: [128 rows x 128 k] K-major SW128, two [128 x 64] boxes
  uint8_t b[128 * 128];  // B half: mode 0 [64 keys x 128 d] K-major (two [64 x 64] boxes); mode 1 [128 keys x 64 d] MN-major
  uint64_t done;
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t sw_off(int r, int c, int R) {  // K-major: row r, column c of an R-row operand
  const int box = c >> 6, cc = c & 63;
  return box * R * 128 + r * 128 + ((((cc >> 3) ^ (r & 7)) << 4) | ((cc & 7) << 1));
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// A: [2][128][128] bf16 bits (per CTA rank), B: [128 keys][128 d] bf16 bits; out: [2][128][128] fp32
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) k2(const unsigned short* A, const unsigned short* B,
                                                                     float* out, int mode) {
  extern __shared__ uint8_t raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t rank = cluster_rank();
  const unsigned short* a = A + (size_t)rank * 128 * 128;
  if (mode == 0) {
    for (int c = 0; c < 128; c += 8)
      *reinterpret_cast<uint4*>(sm.a + sw_off(tid, c, 128)) = *reinterpret_cast<const uint4*>(a + tid * 128 + c);
    if (tid < 64)  // my half of B: keys 64*rank + tid
      for (int c = 0; c < 128; c += 8)
        *reinterpret_cast<uint4*>(sm.b + sw_off(tid, c, 64)) =
            *reinterpret_cast<const uint4*>(B + (size_t)(64 * rank + tid) * 128 + c);
  } else {
    // MN-major B half: row = key (k index), 64 contiguous d (my half) = one 128-byte swizzled row
    for (int c = 0; c < 64; c += 8)
      *reinterpret_cast<uint4*>(sm.b + tid * 128 + ((((c >> 3) ^ (tid & 7)) << 4))) =
          *reinterpret_cast<const uint4*>(B + (size_t)tid * 128 + 64 * rank + c);
  }
  if (tid == 0) { mbar_init(&sm.done, 1); mbar_fence_init(); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  if (mode == 1) {  // A rows (16-bit) -> TMEM columns [128, 192): word w holds elements (2w, 2w+1)
    const uint32_t* arow = reinterpret_cast<const uint32_t*>(a) + tid * 64;
    for (int c = 0; c < 2; ++c) {
      uint32_t w[32];
      for (int i = 0; i < 32; ++i) w[i] = arow[c * 32 + i];
      tmem_st32(tmem + 128 + lane_base + c * 32, w);
    }
    tc_wait_st();
    tc_fence_before();
  }
  cluster_sync();  // both CTAs' operands are in place
  tc_fence_after();
  if (rank == 0 && tid == 0) {
    const uint32_t a_addr = smem_u32(sm.a), b_addr = smem_u32(sm.b);
    if (mode == 0) {
      const uint32_t idesc = umma_idesc(true, 256, 128, false, false);
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t offa = (kk >> 2) * 16384 + (kk & 3) * 32, offb = (kk >> 2) * 8192 + (kk & 3) * 32;
        const uint64_t ad = umma_smem_desc(a_addr + offa, 16, 1024), bd = umma_smem_desc(b_addr + offb, 16, 1024);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(ad), "l"(bd),
                     "r"(idesc), "r"(kk > 0 ? 1u : 0u) : "memory");
      }
    } else {
      const uint32_t idesc = umma_idesc(true, 256, 128, false, true);
      for (int kk = 0; kk < 8; ++kk) {
        const uint64_t bd = umma_smem_desc(b_addr + kk * 2048, 16, 1024);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem), "r"(tmem + 128 + kk * 8),
                     "l"(bd), "r"(idesc), "r"(kk > 0 ? 1u : 0u) : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(&sm.done)), "h"((unsigned short)3) : "memory");
  }
  mbar_wait(&sm.done, 0);
  tc_fence_after();
  for (int c = 0; c < 4; ++c) {
    uint32_t o[32];
    tmem_ld32(tmem + lane_base + c * 32, o);
    tc_wait_ld();
    for (int i = 0; i < 32; ++i) out[((size_t)rank * 128 + tid) * 128 + c * 32 + i] = __uint_as_float(o[i]);
  }
  tc_fence_before();
  cluster_sync();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

static float bf(unsigned short h) { unsigned u = (unsigned)h << 16; float f; memcpy(&f, &u, 4); return f; }
static unsigned short tobf(float f) { __nv_bfloat16 h = __float2bfloat16(f); unsigned short u; memcpy(&u, &h, 2); return u; }

int main() {
  std::vector<unsigned short> A(2 * 128 * 128), B(128 * 128);
  srand(1);
  for (auto& x : A) x = tobf((rand() % 2001 - 1000) / 500.f);
  for (auto& x : B) x = tobf((rand() % 2001 - 1000) / 500.f);
  unsigned short *dA, *dB; float* dO;
  cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dO, 2 * 128 * 128 * 4);
  cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
  const int smem = sizeof(Smem) + 1024;
  cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> O(2 * 128 * 128);
  for (int mode = 0; mode < 2; ++mode) {
    cudaMemset(dO, 0xff, O.size() * 4);
    k2<<<2, 128, smem>>>(dA, dB, dO, mode);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(O.data(), dO, O.size() * 4, cudaMemcpyDeviceToHost);
    double max_err = 0; long bad = 0;
    for (int r = 0; r < 256; ++r)
      for (int n = 0; n < 128; ++n) {
        double ref = 0;
        for (int k = 0; k < 128; ++k)  // mode 0: B[n][k] (n = key), mode 1: B[k][n] (k = key, n = d)
          ref += (double)bf(A[r * 128 + k]) * (mode == 0 ? bf(B[n * 128 + k]) : bf(B[k * 128 + n]));
        const double err = fabs(O[r * 128 + n] - ref);
        if (!(err <= 1e-2)) ++bad;
        if (err > max_err || err != err) max_err = err;
      }
    printf("mode %d (%s): max |err| %.3g, %ld of 32768 entries off  [%s]\n", mode, mode == 0 ? "SS, K-major B split over keys" : "TS, MN-major B split over d",
           max_err, bad, cudaGetErrorString(e));
  }
}
