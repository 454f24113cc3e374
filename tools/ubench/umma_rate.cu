// Throughput of tcgen05.mma from ONE CTA per SM (cta_group::1), operands resident in shared memory / TMEM, the
// accumulator rotating over independent TMEM buffers so that the figure is the issue rate, not the latency of a
// dependent accumulate chain (a single accumulator measures 75-230 cycles per instruction depending on the box).
// Nominal floor for 128 x 128 x 16: 64 cycles (8192 flop / clk / SM).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I compactor-vllm_b200/csrc -o tools/ubench/umma_rate tools/ubench/umma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tma.cuh"
using namespace cvllm;
struct __align__(1024) Smem { uint8_t a[32768], b[32768], v[32768]; uint64_t done; uint32_t tmem_base; };
// mode 0: SS K-major B (S = Q.K^T)   1: TS, MN-major B (O += P.V, P in TMEM cols [0,64))   2: SS, MN-major B (P from smem)
// mode 3: the prefill issuer's mix: PV0 S0 PV1 S1 on its four accumulators
__global__ void __launch_bounds__(128) k(long long* clk, int mode, int reps, int n_acc) {
  extern __shared__ uint8_t raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 3 * 32768 / 4; i += 128) reinterpret_cast<uint32_t*>(sm.a)[i] = 0x3c003c00u;
  if (tid == 0) { mbar_init(&sm.done, 1); mbar_fence_init(); }
  if (warp == 0) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  if (tid == 0) {
    const uint32_t a = smem_u32(sm.a), b = smem_u32(sm.b), v = smem_u32(sm.v);
    const uint32_t id_k = umma_idesc(false, 128, 128, false, false), id_mn = umma_idesc(false, 128, 128, false, true);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (mode == 3) {
        const int t = r & 1;
        for (int kk = 0; kk < 8; ++kk) umma_ts(tmem + 256 + t * 128, tmem + t * 128 + kk * 8, umma_smem_desc(v + kk * 2048, 16384, 1024), id_mn, 1);
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          umma_ss(tmem + t * 128, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(b + off, 16, 1024), id_k, 1);
        }
        continue;
      }
      // accumulators: SS modes rotate over all four 128-column buffers; TS keeps A in [0,64) and rotates D over 128..511
      const uint32_t d = (mode == 1) ? tmem + 128 + (r % (n_acc > 3 ? 3 : n_acc)) * 128 : tmem + (r % n_acc) * 128;
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
        if (mode == 0) umma_ss(d, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(b + off, 16, 1024), id_k, 1);
        if (mode == 1) umma_ts(d, tmem + kk * 8, umma_smem_desc(v + kk * 2048, 16384, 1024), id_mn, 1);
        if (mode == 2) umma_ss(d, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(v + kk * 2048, 16384, 1024), id_mn, 1);
      }
    }
    umma_commit(&sm.done);
    mbar_wait(&sm.done, 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0) clk[0] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  const int smem = sizeof(Smem) + 1024, reps = 2000;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"SS, K-major B   (S = Q.K^T)", "TS, MN-major B  (O += P.V, P in TMEM)", "SS, MN-major B  (P.V with P in smem)", "prefill mix     (PV_t then S_t, t alternating)"};
  for (int rep = 0; rep < 2; ++rep)
    for (int mode = 0; mode < 4; ++mode)
      for (int n_acc = 1; n_acc <= 4; n_acc += (mode == 3 ? 4 : 3)) {
        k<<<148, 128, smem>>>(d, mode, reps, n_acc); cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        const double per = c / ((double)reps * (mode == 3 ? 16 : 8));
        printf("%-48s %d accumulator(s): %6.1f cycles per 128x128x16 (%.0f flop/clk/SM)\n", names[mode], mode == 3 ? 4 : n_acc, per,
               128.0 * 128 * 16 * 2 / per);
      }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
