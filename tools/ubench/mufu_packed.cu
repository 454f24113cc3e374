// Throughput of the exponential on the MUFU pipe: ex2.approx.ftz.f32 (one result per instruction) against the packed
// 16-bit forms ex2.approx.ftz.bf16x2 / ex2.approx.f16x2 (two results per instruction).  Results per clock per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench/mufu_packed tools/ubench/mufu_packed.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(unsigned* out, int iters, unsigned seed) {
  unsigned a[8];
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 8 + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(a[i]));
      if (MODE == 2) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(a[i]));
    }
  }
  long long t1 = clock64();
  unsigned s = 0;
  for (int i = 0; i < 8; ++i) s ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[gridDim.x * blockDim.x] = (unsigned)(t1 - t0);
}
int main() {
  unsigned* d; cudaMalloc(&d, (148 * 1024 + 1) * 4);
  const int iters = 4096;
  const char* names[] = {"ex2.approx.ftz.f32", "ex2.approx.ftz.bf16x2", "ex2.approx.f16x2"};
  for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      if (mode == 0) k<0><<<148, 1024>>>(d, iters, 1u);
      if (mode == 1) k<1><<<148, 1024>>>(d, iters, 1u);
      if (mode == 2) k<2><<<148, 1024>>>(d, iters, 1u);
      cudaDeviceSynchronize();
    }
    unsigned clk; cudaMemcpy(&clk, d + 148 * 1024, 4, cudaMemcpyDeviceToHost);
    const double instr_per_sm = 1024.0 * 8 * iters;  // thread-instructions on one SM
    printf("%-24s %.2f thread-instr/clk/SM = %.2f results/clk/SM\n", names[mode], instr_per_sm / clk, instr_per_sm / clk * (mode ? 2 : 1));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
