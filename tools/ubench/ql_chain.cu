// Microbenchmark of the QL rotation recurrence (score_leverage.cu producer loop): cycles per rotation for variants.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ double rsqrt64(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x * y, y, 1.0);
  return fma(y * e, fma(0.375, e, 0.5), y);
}
struct Sm { double d[64], e[64]; double2 ring[1024]; };
template <int MODE>
__global__ void k(double* out, long long* clk, int sweeps) {
  __shared__ Sm s;
  const int lane = threadIdx.x & 31;
  for (int i = lane; i < 64; i += 32) { s.d[i] = 1.0 + 0.37 * i; s.e[i] = 0.21 + 0.01 * i; }
  __syncwarp();
  double acc = 0;
  long long t0 = clock64();
  int slot = 0;
  for (int sw = 0; sw < sweeps; ++sw) {
    const int m = 47, l = 0;
    double p = s.d[m];
    double c = 1.0, c2 = 1.0, c3 = 1.0, sn = 0.0, s2 = 0.0;
    double ei = s.e[m - 1], di = s.d[m - 1];
    for (int i = m - 1; i >= l; --i) {
      const double ein = (i > l) ? s.e[i - 1] : 0.0, din = (i > l) ? s.d[i - 1] : 0.0;
      c3 = c2; c2 = c; s2 = sn;
      const double t = ei * ei;
      const double g = c * ei, h = c * p;
      const double r2 = fma(p, p, t);
      const double qv = fma(p, di, -(t * c));
      double ri;
      if (MODE & 1) ri = rsqrt64(r2); else ri = (r2 > 0.0) ? rsqrt64(r2) : 0.0;
      const double r = r2 * ri;
      sn = ei * ri;
      c = p * ri;
      p = ri * qv;
      if (!(MODE & 2)) {
        if (lane == 0) {
          s.e[i + 1] = s2 * r;
          s.d[i + 1] = h + sn * fma(c, g, sn * di);
          s.ring[slot] = make_double2(c, sn);
        }
      } else {
        acc += s2 * r + h + sn * fma(c, g, sn * di);
      }
      slot = (slot + 1) & 1023;
      ei = ein; di = din;
    }
    acc += p + c3;
    __syncwarp();
    if (lane == 0) { s.e[l] = 0.3; s.d[l] = 1.1; }
    __syncwarp();
  }
  long long t1 = clock64();
  if (lane == 0) clk[0] = t1 - t0;
  out[threadIdx.x] = acc + s.d[3];
}
// software-pipelined: the d/e outputs of rotation i are formed during rotation i-1's latency shadow
template <int MODE>
__global__ void kp(double* out, long long* clk, int sweeps) {
  __shared__ Sm s;
  const int lane = threadIdx.x & 31;
  for (int i = lane; i < 64; i += 32) { s.d[i] = 1.0 + 0.37 * i; s.e[i] = 0.21 + 0.01 * i; }
  __syncwarp();
  double acc = 0;
  long long t0 = clock64();
  int slot = 0;
  for (int sw = 0; sw < sweeps; ++sw) {
    const int m = 47, l = 0;
    double p = s.d[m];
    double c = 1.0, sn = 0.0;
    double ei = s.e[m - 1], di = s.d[m - 1];
    // deferred state of the previous rotation
    double pr_h = 0, pr_g = 0, pr_di = 0, pr_sn = 0, pr_c = 1, pr_s2 = 0, pr_r2 = 0, pr_ri = 0;
    double c2 = 1.0, c3 = 1.0, s2 = 0.0;
    for (int i = m - 1; i >= l; --i) {
      const double ein = (i > l) ? s.e[i - 1] : 0.0, din = (i > l) ? s.d[i - 1] : 0.0;
      const double t = ei * ei;
      const double r2 = fma(p, p, (MODE & 1) ? t + 1e-300 : t);
      const double qv = fma(p, di, -(t * c));
      double y;
      asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r2));
      // ---- deferred outputs of rotation i+1 (inputs all ready) ----
      if (i < m - 1) {
        const double r = pr_r2 * pr_ri;
        const double eo = pr_s2 * r;
        const double dd = pr_h + pr_sn * fma(pr_c, pr_g, pr_sn * pr_di);
        if (lane == 0) { s.e[i + 2] = eo; s.d[i + 2] = dd; s.ring[slot] = make_double2(pr_c, pr_sn); }
        slot = (slot + 1) & 1023;
      }
      const double g = c * ei, h = c * p;
      const double e = fma(-r2 * y, y, 1.0);
      const double ri = fma(y * e, fma(0.375, e, 0.5), y);
      c3 = c2; c2 = c; s2 = sn;
      pr_h = h; pr_g = g; pr_di = di; pr_s2 = sn; pr_r2 = r2; pr_ri = ri;
      sn = ei * ri;
      c = p * ri;
      p = ri * qv;
      pr_sn = sn; pr_c = c;
      ei = ein; di = din;
    }
    {
      const double r = pr_r2 * pr_ri;
      if (lane == 0) { s.e[l + 1] = pr_s2 * r; s.d[l + 1] = pr_h + pr_sn * fma(pr_c, pr_g, pr_sn * pr_di); s.ring[slot] = make_double2(pr_c, pr_sn); }
      slot = (slot + 1) & 1023;
    }
    acc += p + c3 + s2;
    __syncwarp();
    if (lane == 0) { s.e[l] = 0.3; s.d[l] = 1.1; }
    __syncwarp();
  }
  long long t1 = clock64();
  if (lane == 0) clk[0] = t1 - t0;
  out[threadIdx.x] = acc + s.d[3];
}
// lean loop: only the recurrence; (p, 1/r) per rotation go to smem, everything else is a parallel post-pass
template <int MODE>
__global__ void kl(double* out, long long* clk, int sweeps, int mm) {
  __shared__ Sm s;
  __shared__ double2 pr[64];
  const int lane = threadIdx.x & 31;
  for (int i = lane; i < 64; i += 32) { s.d[i] = 1.0 + 0.37 * i; s.e[i] = 0.21 + 0.01 * i; }
  __syncwarp();
  double acc = 0;
  long long t0 = clock64();
  for (int sw = 0; sw < sweeps; ++sw) {
    const int m = mm, l = 0;
    double p = s.d[m];
    double c = 1.0;
    double ei = s.e[m - 1], di = s.d[m - 1];
    for (int i = m - 1; i >= l; --i) {
      const double ein = (i > l) ? s.e[i - 1] : 0.0, din = (i > l) ? s.d[i - 1] : 0.0;
      const double t = ei * ei;
      const double r2 = fma(p, p, fma(ei, ei, 1e-300));
      const double qv = fma(p, di, -(t * c));
      double y;
      asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r2));
      const double e = fma(-r2 * y, y, 1.0);
      const double ri = fma(y * e, fma(0.375, e, 0.5), y);
      if (lane == 0) pr[i] = make_double2(p, ri);
      c = p * ri;
      p = ri * qv;
      ei = ein; di = din;
    }
    __syncwarp();
    // post-pass: rotation i handled by lane (m-1-i) (+32)
    for (int k = lane; k < m - l; k += 32) {
      const int i = m - 1 - k;
      const double2 a = pr[i];
      const double e_i = s.e[i], d_i = s.d[i];
      const double c_i = a.x * a.y, s_i = e_i * a.y;
      double c_p = 1.0, s_p = 0.0;
      if (i < m - 1) { const double2 b = pr[i + 1]; c_p = b.x * b.y; s_p = s.e[i + 1] * b.y; }
      const double g = c_p * e_i, h = c_p * a.x;
      const double r = fma(a.x, a.x, e_i * e_i) * a.y;
      acc += c_i;
      s.ring[(sw * 47 + k) & 1023] = make_double2(c_i, s_i);
      // (stores to d/e omitted from the benchmark's data flow to keep inputs stable)
      if (MODE == 99) { s.e[i + 1] = s_p * r; s.d[i + 1] = h + s_i * fma(c_i, g, s_i * d_i); }
    }
    __syncwarp();
    acc += p;
  }
  long long t1 = clock64();
  if (lane == 0) clk[0] = t1 - t0;
  out[threadIdx.x] = acc + s.d[3];
}
int main() {
  double* out; long long* clk;
  cudaMalloc(&out, 1024 * 8); cudaMalloc(&clk, 8);
  const int sweeps = 200;
  auto run = [&](auto kern, const char* name) {
    kern<<<1, 32>>>(out, clk, sweeps); kern<<<1, 32>>>(out, clk, sweeps);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    printf("%-40s %.1f cycles per rotation\n", name, (double)c / (sweeps * 47));
  };
  run(k<0>, "guarded rsqrt, smem stores");
  run(k<1>, "unguarded rsqrt, smem stores");
  run(k<2>, "guarded rsqrt, no stores");
  run(k<3>, "unguarded rsqrt, no stores");
  {
    kl<0><<<1, 32>>>(out, clk, sweeps, 47); kl<0><<<1, 32>>>(out, clk, sweeps, 47);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    printf("%-40s %.1f cycles per rotation\n", "lean loop + parallel post-pass", (double)c / (sweeps * 47));
  }
  run(kp<0>, "pipelined, stores");
  run(kp<1>, "pipelined, stores, +tiny guard");
  return 0;
}
