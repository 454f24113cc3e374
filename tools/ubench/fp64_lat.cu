// Microbenchmark: fp64 dependent-chain latency and issue throughput on this GPU (one warp, one CTA).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* clk, double x0, int iters, int mode, int active_lanes) {
  const int lane = threadIdx.x & 31;
  double a = x0 + lane * 1e-9, b = 1.0000001, c = 1e-9;
  double a1 = a + 1, a2 = a + 2, a3 = a + 3, a4 = a + 4, a5 = a + 5, a6 = a + 6, a7 = a + 7;
  __syncthreads();
  long long t0 = clock64();
  if (lane < active_lanes) {
    if (mode == 0) {  // dependent DFMA chain
      for (int i = 0; i < iters; ++i) { a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); }
    } else if (mode == 1) {  // 8 independent DFMA chains
      for (int i = 0; i < iters; ++i) {
        a = fma(a, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
      }
    } else if (mode == 2) {  // dependent rsqrt.approx.f64 chain (+1 DADD to stay in range)
      for (int i = 0; i < iters; ++i) {
        double y; asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a)); a = y + 1.0;
        asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a)); a = y + 1.0;
      }
    } else if (mode == 3) {  // dependent FFMA chain (fp32) for reference
      float f = (float)a, g = 1.0000001f, h = 1e-9f;
      for (int i = 0; i < iters; ++i) { f = fmaf(f, g, h); f = fmaf(f, g, h); f = fmaf(f, g, h); f = fmaf(f, g, h); }
      a = f;
    } else if (mode == 4) {  // dependent DMUL chain
      for (int i = 0; i < iters; ++i) { a = a * b; a = a * b; a = a * b; a = a * b; }
    } else if (mode == 5) {  // dependent DADD chain
      for (int i = 0; i < iters; ++i) { a = a + c; a = a + c; a = a + c; a = a + c; }
    }
  }
  long long t1 = clock64();
  if (lane == 0) clk[0] = t1 - t0;
  out[threadIdx.x] = a + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
int main() {
  double* out; long long* clk;
  cudaMalloc(&out, 1024 * 8); cudaMalloc(&clk, 8);
  const char* names[] = {"DFMA dependent", "DFMA 8 independent chains", "RSQ64H+DADD dependent", "FFMA dependent", "DMUL dependent", "DADD dependent"};
  const int per_iter[] = {4, 8, 2, 4, 4, 4};
  for (int lanes : {32, 1}) for (int warps : {1, 4}) for (int mode = 0; mode < 6; ++mode) {
    const int iters = 2000;
    k<<<1, 32 * warps>>>(out, clk, 1.5, iters, mode, lanes);
    k<<<1, 32 * warps>>>(out, clk, 1.5, iters, mode, lanes);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    printf("lanes=%2d warps=%d %-28s %.1f cycles per op\n", lanes, warps, names[mode], (double)c / (iters * per_iter[mode]));
  }
  cudaError_t e = cudaGetLastError(); if (e) printf("err %s\n", cudaGetErrorString(e));
  return 0;
}
