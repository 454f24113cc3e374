// What slows the prefill issuer's UMMA stream from the nominal 64 cycles per 128 x 128 x 16 to the ~86 seen by the
// in-kernel trace (tools/trace_prefill.py)?  One CTA per SM, thread 0 issues the kernel's mix (8 TS into O_t, 8 SS into
// S_t, t alternating) while the other warps do ONE of the things the kernel's other warps do:
//   commits   tcgen05.commit after every group of 8 (as the kernel: s_full / o_done / empty)
//   spin      N warps polling mbarrier.try_wait on a barrier that completes only at the end
//   ldtm      N warps reading 128 TMEM columns (tcgen05.ld x4 + wait) in a loop, as the softmax does with S
//   sttm      ... and writing 64 columns back (tcgen05.st), as the softmax does with P
// The operands are the constant 1.0 unless `random_data`; random operand data does NOT slow the stream (64.0 either way
// over 20 M cycles), so the gap to cuBLAS's measured 1658 TFLOP/s is not a data-dependent throttle at this time scale.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I compactor-vllm_b200/csrc -o tools/ubench/umma_mix tools/ubench/umma_mix.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tma.cuh"
using namespace cvllm;
struct __align__(1024) Smem {
  uint8_t a[32768], b[32768], v[32768];
  uint64_t done, never, ca, cb;
  uint32_t tmem_base;
  volatile int stop;
};

__global__ void __launch_bounds__(384) k(long long* clk, int commits, int spin_warps, int ldtm_warps, int sttm, int reps, int random_data) {
  extern __shared__ uint8_t raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 3 * 32768 / 4; i += 384) {
    uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    // random_data: two fp16 values with random sign / mantissa and exponents in [2^-3, 2^0]; else the constant 1.0
    const uint32_t rnd = (h & 0x83ff83ffu) | 0x30003000u | ((h >> 4) & 0x0c000c00u);
    reinterpret_cast<uint32_t*>(sm.a)[i] = random_data ? rnd : 0x3c003c00u;
  }
  if (tid == 0) {
    mbar_init(&sm.done, 1); mbar_init(&sm.never, 1); mbar_init(&sm.ca, 1); mbar_init(&sm.cb, 1);
    sm.stop = 0;
    mbar_fence_init();
  }
  if (warp == 0) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  if (warp == 0) {
    if (tid == 0) {
      const uint32_t a = smem_u32(sm.a), b = smem_u32(sm.b), v = smem_u32(sm.v);
      const uint32_t id_k = umma_idesc(false, 128, 128, false, false), id_mn = umma_idesc(false, 128, 128, false, true);
      const long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        const int t = r & 1;
        for (int kk = 0; kk < 8; ++kk)
          umma_ts(tmem + 256 + t * 128, tmem + t * 128 + kk * 8, umma_smem_desc(v + kk * 2048, 16384, 1024), id_mn, 1);
        if (commits) umma_commit(&sm.ca);
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          umma_ss(tmem + t * 128, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(b + off, 16, 1024), id_k, 1);
        }
        if (commits) { umma_commit(&sm.cb); if (commits > 1) umma_commit(&sm.ca); }
      }
      umma_commit(&sm.done);
      mbar_wait(&sm.done, 0);
      const long long t1 = clock64();
      if (blockIdx.x == 0) clk[0] = t1 - t0;
      sm.stop = 1;
      mbar_arrive(&sm.never);
    }
  } else if (warp >= 4 && warp < 4 + spin_warps) {
    while (!mbar_try_wait(&sm.never, 0)) { }
  } else if (warp >= 4 && warp < 4 + ldtm_warps) {
    // a warp may only touch its own TMEM lane quadrant: warp % 4
    const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + ((warp >> 2) & 1) * 128;
    uint32_t acc = 0;
    while (!sm.stop) {
      for (int c = 0; c < 4; ++c) {
        uint32_t x[32];
        tmem_ld32(base + c * 32, x);
        tc_wait_ld();
        acc += x[0] ^ x[31];
      }
      if (sttm) {
        uint32_t y[16];
        for (int i = 0; i < 16; ++i) y[i] = 0x3c003c00u;
        for (int c = 0; c < 4; ++c) tmem_st16(base + c * 16, y);
        tc_wait_st();
      }
    }
    if (acc == 0x12345) clk[1] = acc;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  const int smem = sizeof(Smem) + 1024, reps = 2000;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  struct Case { const char* name; int commits, spin, ldtm, sttm, rnd; } cases[] = {
      {"mix alone", 0, 0, 0, 0},
      {"mix alone, RANDOM operand data", 0, 0, 0, 0, 1},
      {"+ commits + 8 warps ld/st, RANDOM data", 2, 0, 8, 1, 1},
      {"+ 2 commits per group pair", 1, 0, 0, 0},
      {"+ 3 commits per group pair", 2, 0, 0, 0},
      {"+ 8 warps polling mbarrier.try_wait", 0, 8, 0, 0},
      {"+ 4 warps polling mbarrier.try_wait", 0, 4, 0, 0},
      {"+ 8 warps tcgen05.ld 128 cols in a loop", 0, 0, 8, 0},
      {"+ 8 warps tcgen05.ld + tcgen05.st", 0, 0, 8, 1},
      {"+ commits + 8 warps ld/st", 2, 0, 8, 1},
  };
  for (int rep = 0; rep < 2; ++rep)
    for (auto& c : cases) {
      k<<<148, 384, smem>>>(d, c.commits, c.spin, c.ldtm, c.sttm, c.rnd ? 20000 : reps, c.rnd);
      cudaError_t e = cudaDeviceSynchronize();
      long long cyc; cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
      printf("%-44s %6.1f cycles per 128x128x16  [%s]\n", c.name, cyc / ((double)(c.rnd ? 20000 : reps) * 16), cudaGetErrorString(e));
    }
}
