// Issue-rate ceiling of tcgen05.mma from ONE CTA per SM (cta_group::1), operands resident in shared memory / TMEM:
// cycles per 128x128x16 instruction for  SS (A and B from smem: the S = Q.K^T shape),  TS (A from TMEM: O += P.V),
// and the prefill kernel's mix (S, PV alternating on two accumulators).  Nominal floor: 64 cycles (8192 flop/clk/SM).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I compactor-vllm_b200/csrc -o tools/ubench/umma_rate tools/ubench/umma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tma.cuh"
using namespace cvllm;
struct __align__(1024) Smem { uint8_t a[32768], b[32768], v[32768]; uint64_t done; uint32_t tmem_base; };
__global__ void __launch_bounds__(128) k(long long* clk, int mode, int reps, int a_col, int d_col, int alt) {
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
    const uint32_t id_ss = umma_idesc(false, 128, 128, false, false), id_ts = umma_idesc(false, 128, 128, false, true);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
        if (mode == 0) umma_ss(tmem, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(b + off, 16, 1024), id_ss, 1);
        if (mode == 1) umma_ts(tmem + d_col + ((alt && (kk & 1)) ? 128 : 0), tmem + a_col + kk * 8, umma_smem_desc(v + kk * 2048, 16384, 1024), id_ts, 1);
        if (mode == 3) umma_ss(tmem + d_col, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(v + kk * 2048, 16384, 1024), umma_idesc(false, 128, 128, false, true), 1);
        if (mode == 2) {  // S0, PV0, S1, PV1 interleaved like the prefill issuer
          umma_ts(tmem + 256 + (r & 1) * 128, tmem + (r & 1) * 128 + kk * 8, umma_smem_desc(v + kk * 2048, 16384, 1024), id_ts, 1);
        }
      }
      if (mode == 2)
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          umma_ss(tmem + (r & 1) * 128, umma_smem_desc(a + off, 16, 1024), umma_smem_desc(b + off, 16, 1024), id_ss, 1);
        }
    }
    umma_commit(&sm.done);
    mbar_wait(&sm.done, 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0) clk[mode] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  const int smem = sizeof(Smem) + 1024, reps = 2000;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"SS  (A, B from smem)", "TS  (A from TMEM, B from smem)", "mix (8 TS + 8 SS per round)"};
  for (int mode = 0; mode < 3; ++mode) {
    k<<<148, 128, smem>>>(d, mode, reps, 0, 256, 0); cudaDeviceSynchronize();
    k<<<148, 128, smem>>>(d, mode, reps, 0, 256, 0); cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, d + mode, 8, cudaMemcpyDeviceToHost);
    const double n = (double)reps * (mode == 2 ? 16 : 8);
    printf("%-34s %.1f cycles per 128x128x16 instruction (%.0f flop/clk/SM)  [%s]\n", names[mode], c / n, 128.0 * 128 * 16 * 2 / (c / n),
           cudaGetErrorString(cudaGetLastError()));
  }
  // TS placement variants: where A (P, 64 columns) and D (O, 128 columns) sit in TMEM, and alternating accumulators
  const int var[][3] = {{0, 256, 0}, {128, 256, 0}, {384, 0, 0}, {64, 128, 0}, {0, 256, 1}, {448, 0, 1}};
  for (auto& v : var) {
    k<<<148, 128, smem>>>(d, 1, reps, v[0], v[1], v[2]); cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, d + 1, 8, cudaMemcpyDeviceToHost);
    printf("TS  A cols [%3d,+64)  D cols [%3d,+128)%s  %.1f cycles per instruction\n", v[0], v[1], v[2] ? " alternating with D+128" : "", c / (reps * 8.0));
  }
  k<<<148, 128, smem>>>(d, 3, reps, 0, 256, 0); cudaDeviceSynchronize();
  { long long c; cudaMemcpy(&c, d + 3, 8, cudaMemcpyDeviceToHost); printf("SS with MN-major B (P from smem, the PV shape)  %.1f cycles per instruction\n", c / (reps * 8.0)); }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
