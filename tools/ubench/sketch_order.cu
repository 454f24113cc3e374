// Which fp32 accumulation order reproduces cuBLAS's bf16 GEMM X = K.PHI bit for bit?  Reads /tmp/sk_{k,phi,x}.bin
// written by tools/ubench/sketch_order.py (K [N,128] bf16 of one head, PHI [128,48] bf16, X_ref [N,48] bf16).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <mma.h>
__device__ __forceinline__ void mma16816(float* c, unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma1688(float* c, unsigned a0, unsigned a1, unsigned b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
// one warp per 16 tokens; mode 0: k16 sequential; 1: k8 sequential; 2: two k-halves summed; 3: four quarters;
// 4: scalar fp32 fma sequential; 5: k16 sequential in reverse order; 6: k32 pairs (two k16 into separate acc then add, sequential over pairs)
__global__ void sk(const __nv_bfloat16* K, const __nv_bfloat16* PHI, __nv_bfloat16* X, int N, int mode) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int t0 = warp * 16;
  if (t0 >= N) return;
  const int r = lane >> 2, j = lane & 3;
  const unsigned short* k16 = reinterpret_cast<const unsigned short*>(K);
  const unsigned short* p16 = reinterpret_cast<const unsigned short*>(PHI);
  auto ka = [&](int row, int col) -> unsigned { return *reinterpret_cast<const unsigned*>(k16 + (size_t)(t0 + row) * 128 + col); };
  auto pb = [&](int kk, int n) -> unsigned { return (unsigned)p16[kk * 48 + n] | ((unsigned)p16[(kk + 1) * 48 + n] << 16); };
  for (int nn = 0; nn < 6; ++nn) {
    float acc[4] = {0, 0, 0, 0};
    if (mode == 4) {
      for (int e = 0; e < 4; ++e) {
        const int row = r + (e >> 1) * 8, col = nn * 8 + 2 * j + (e & 1);
        float a = 0;
        for (int kk = 0; kk < 128; ++kk)
          a = fmaf(__bfloat162float(K[(size_t)(t0 + row) * 128 + kk]), __bfloat162float(PHI[kk * 48 + col]), a);
        acc[e] = a;
      }
    } else if (mode == 1) {
      for (int ks = 0; ks < 16; ++ks) {
        const int k0 = ks * 8;
        mma1688(acc, ka(r, k0 + 2 * j), ka(r + 8, k0 + 2 * j), pb(k0 + 2 * j, nn * 8 + r));
      }
    } else {
      const int parts = mode == 2 ? 2 : mode == 3 ? 4 : mode == 6 ? 4 : 1;
      float tot[4] = {0, 0, 0, 0};
      for (int part = 0; part < parts; ++part) {
        float a4[4] = {0, 0, 0, 0};
        const int per = 8 / parts;
        for (int q = 0; q < per; ++q) {
          int ks = part * per + q;
          if (mode == 5) ks = 7 - ks;
          const int k0 = ks * 16;
          mma16816(mode == 6 ? tot : a4, ka(r, k0 + 2 * j), ka(r + 8, k0 + 2 * j), ka(r, k0 + 8 + 2 * j), ka(r + 8, k0 + 8 + 2 * j),
                   pb(k0 + 2 * j, nn * 8 + r), pb(k0 + 8 + 2 * j, nn * 8 + r));
        }
        if (mode != 6) for (int e = 0; e < 4; ++e) tot[e] += a4[e];
      }
      for (int e = 0; e < 4; ++e) acc[e] = tot[e];
    }
    for (int e = 0; e < 4; ++e) {
      const int row = r + (e >> 1) * 8, col = nn * 8 + 2 * j + (e & 1);
      X[(size_t)(t0 + row) * 48 + col] = __float2bfloat16_rn(acc[e]);
    }
  }
}
template <class T> std::vector<T> rd(const char* f) { FILE* fp = fopen(f, "rb"); if (!fp) { printf("missing %s\n", f); exit(1); } fseek(fp, 0, SEEK_END); long n = ftell(fp); fseek(fp, 0, SEEK_SET); std::vector<T> v(n / sizeof(T)); if (fread(v.data(), 1, n, fp) != (size_t)n) exit(1); fclose(fp); return v; }
int main() {
  auto k = rd<unsigned short>("/tmp/sk_k.bin"), phi = rd<unsigned short>("/tmp/sk_phi.bin"), x = rd<unsigned short>("/tmp/sk_x.bin");
  const int N = (int)(k.size() / 128);
  __nv_bfloat16 *dk, *dp, *dx; cudaMalloc(&dk, k.size() * 2); cudaMalloc(&dp, phi.size() * 2); cudaMalloc(&dx, x.size() * 2);
  cudaMemcpy(dk, k.data(), k.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(dp, phi.data(), phi.size() * 2, cudaMemcpyHostToDevice);
  const char* names[] = {"k16 sequential", "k8 sequential", "2 halves", "4 quarters", "scalar fmaf sequential", "k16 reverse", "k16 seq (same acc)"};
  std::vector<unsigned short> out(x.size());
  for (int mode = 0; mode < 7; ++mode) {
    sk<<<(N / 16 + 3) / 4, 128>>>(dk, dp, dx, N, mode);
    cudaMemcpy(out.data(), dx, x.size() * 2, cudaMemcpyDeviceToHost);
    long diff = 0; for (size_t i = 0; i < x.size(); ++i) diff += out[i] != x[i];
    printf("%-26s differs from torch.matmul in %ld of %zu entries\n", names[mode], diff, x.size());
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
