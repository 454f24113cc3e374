// Does a tcgen05 (UMMA) contraction reproduce cuBLAS's bf16 GEMM X = K.PHI bit for bit?  Companion of
// sketch_order.cu (which tries mma.sync orders).  Reads /tmp/sk_{k,phi,x}.bin written by sketch_order.py
// (K [N,128] bf16 of one head, PHI [128,48] bf16, X_ref [N,48] bf16 from torch.matmul).
//   mode 0: M=128 tokens x N=48, 8 x k16 accumulated in TMEM in ascending k
//   mode 1: same, descending k
//   mode 2: transposed problem (A = PHI^T zero-padded to 128 rows, N = 128 tokens), ascending k
//   mode 3: two k-halves in separate accumulators, added in fp32 registers
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I compactor-vllm_b200/csrc -o tools/ubench/sketch_umma tools/ubench/sketch_umma.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "tma.cuh"

using namespace cvllm;

struct __align__(1024) Smem {
  uint8_t a[2 * 128 * 128];  // A operand: [rows x 64] K-major SW128 boxes, second box = columns 64..127
  uint8_t b[2 * 128 * 128];
  uint64_t done;
  uint32_t tmem_base;
};

// byte offset of element (r, c) of a K-major 128B-swizzled operand with R rows and 128 columns
__device__ __forceinline__ uint32_t sw_off(int r, int c, int R) {
  const int box = c >> 6, cc = c & 63;
  return box * R * 128 + r * 128 + ((((cc >> 3) ^ (r & 7)) << 4) | ((cc & 7) << 1));
}

__global__ void __launch_bounds__(128) sk_umma(const unsigned short* K, const unsigned short* PHI, unsigned short* X,
                                                int N, int mode) {
  extern __shared__ uint8_t raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  const int t0 = blockIdx.x * 128;
  const bool transposed = mode == 2;
  const int RA = 128, RB = transposed ? 128 : 48;
  uint8_t* tok = transposed ? sm.b : sm.a;   // the 128 token rows
  uint8_t* phi = transposed ? sm.a : sm.b;   // PHI^T rows (48; zero-padded to 128 when it is the A operand)
  const int RT = 128, RP = transposed ? 128 : 48;
  for (int c = 0; c < 128; c += 8) {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (t0 + tid < N) v = *reinterpret_cast<const uint4*>(K + (size_t)(t0 + tid) * 128 + c);
    *reinterpret_cast<uint4*>(tok + sw_off(tid, c, RT)) = v;
  }
  if (tid < RP)
    for (int c = 0; c < 128; ++c)
      *reinterpret_cast<unsigned short*>(phi + sw_off(tid, c, RP)) = tid < 48 ? PHI[c * 48 + tid] : (unsigned short)0;
  if (tid == 0) { mbar_init(&sm.done, 1); mbar_fence_init(); }
  if (warp == 0) { tmem_alloc(&sm.tmem_base, 256); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  if (tid == 0) {
    const uint32_t a_addr = smem_u32(sm.a), b_addr = smem_u32(sm.b);
    const uint32_t idesc = umma_idesc(true, RA, RB, false, false);
    for (int i = 0; i < 8; ++i) {
      const int kk = mode == 1 ? 7 - i : i;
      const uint32_t offa = (kk >> 2) * RA * 128 + (kk & 3) * 32, offb = (kk >> 2) * RB * 128 + (kk & 3) * 32;
      const uint32_t d = mode == 3 ? tmem + (kk >> 2) * 128 : tmem;
      const uint32_t acc = mode == 3 ? ((kk & 3) > 0) : (i > 0);
      umma_ss(d, umma_smem_desc(a_addr + offa, 16, 1024), umma_smem_desc(b_addr + offb, 16, 1024), idesc, acc);
    }
    umma_commit(&sm.done);
  }
  mbar_wait(&sm.done, 0);
  tc_fence_after();
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  if (!transposed) {
    uint32_t o[32], o2[32], p[32], p2[32];
    tmem_ld32(tmem + lane_base, o);
    tmem_ld32(tmem + lane_base + 32, o2);
    if (mode == 3) { tmem_ld32(tmem + lane_base + 128, p); tmem_ld32(tmem + lane_base + 160, p2); }
    tc_wait_ld();
    if (t0 + tid < N)
      for (int n = 0; n < 48; ++n) {
        float f = __uint_as_float(n < 32 ? o[n] : o2[n - 32]);
        if (mode == 3) f += __uint_as_float(n < 32 ? p[n] : p2[n - 32]);
        __nv_bfloat16 h = __float2bfloat16_rn(f);
        X[(size_t)(t0 + tid) * 48 + n] = *reinterpret_cast<unsigned short*>(&h);
      }
  } else if (warp < 2) {  // D is [128 x 128]: row = PHI column n (lanes 0..47 used), column = token
    const int n = warp * 32 + (tid & 31);
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      tmem_ld32(tmem + lane_base + c * 32, o);
      tc_wait_ld();
      if (n < 48)
        for (int i = 0; i < 32; ++i)
          if (t0 + c * 32 + i < N) {
            __nv_bfloat16 h = __float2bfloat16_rn(__uint_as_float(o[i]));
            X[(size_t)(t0 + c * 32 + i) * 48 + n] = *reinterpret_cast<unsigned short*>(&h);
          }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <class T> std::vector<T> rd(const char* f) {
  FILE* fp = fopen(f, "rb");
  if (!fp) { printf("missing %s\n", f); exit(1); }
  fseek(fp, 0, SEEK_END); long n = ftell(fp); fseek(fp, 0, SEEK_SET);
  std::vector<T> v(n / sizeof(T));
  if (fread(v.data(), 1, n, fp) != (size_t)n) exit(1);
  fclose(fp);
  return v;
}

int main() {
  auto k = rd<unsigned short>("/tmp/sk_k.bin"), phi = rd<unsigned short>("/tmp/sk_phi.bin"), x = rd<unsigned short>("/tmp/sk_x.bin");
  const int N = (int)(k.size() / 128);
  unsigned short *dk, *dp, *dx;
  cudaMalloc(&dk, k.size() * 2); cudaMalloc(&dp, phi.size() * 2); cudaMalloc(&dx, x.size() * 2);
  cudaMemcpy(dk, k.data(), k.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dp, phi.data(), phi.size() * 2, cudaMemcpyHostToDevice);
  const size_t smem = sizeof(Smem) + 1024;
  cudaFuncSetAttribute(sk_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const char* names[] = {"umma 128x48 k ascending", "umma 128x48 k descending", "umma transposed 128x128", "umma two k-halves + fadd"};
  std::vector<unsigned short> out(x.size());
  for (int mode = 0; mode < 4; ++mode) {
    cudaMemset(dx, 0, x.size() * 2);
    sk_umma<<<(N + 127) / 128, 128, smem>>>(dk, dp, dx, N, mode);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(out.data(), dx, x.size() * 2, cudaMemcpyDeviceToHost);
    long diff = 0, big = 0;
    for (size_t i = 0; i < x.size(); ++i) {
      diff += out[i] != x[i];
      int d = (int)out[i] - (int)x[i];
      big += (d > 1 || d < -1);
    }
    printf("%-28s differs from torch.matmul in %ld of %zu entries (%ld by more than one bf16 ulp) [%s]\n", names[mode], diff,
           x.size(), big, cudaGetErrorString(e));
  }
}
