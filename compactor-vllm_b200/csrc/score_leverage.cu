// Compactor pre-RoPE approximate leverage scores.
//
// Reference: compression/compactor.py:113-221 (torch: batched GEMM K.PHI,
// python loops over chunks, cuSOLVER gesvda on the 48x48 Grams, second GEMM,
// Triton z-score :224-269).  Here: ONE CTA per (chunk, kv head) does the whole
// chain in shared memory, then one small launch z-scores each chunk.
//
//   X   = round16(K_h . PHI)                mma.sync, K rows straight from global
//   Xc  = round16(X - round16(mean_t X))    per chunk
//   G   = fp32(round16(Xc^T Xc)) + reg*I    mma.sync on the transposed sketch
//   G   = V diag(lam) V^T                   Householder + implicit QL, fp32, one warp per matrix
//   u   = round16(Xc . round16(V |lam|^-1/2));  score = round16(sum_j round16(u_j^2))
//
// The reference keeps X, Xc, G, V S^-1/2, U and the row sums in K's 16-bit
// dtype; rounding G entrywise perturbs it by far more than the
// 5e-3 regulariser (it can even go indefinite: the SVD then yields |lam|), so
// those rounding points are part of the reference's scores and are reproduced
// here — an exact fp32 x^T G^-1 x only overlaps ~80 % with the reference's
// kept set (tests/test_oracle_golden.py).  Work is ~170 KFLOP/token: latency
// and launch count, not throughput, is what the fusion removes.
#include <stdlib.h>

#include "common.cuh"
#include "mma.cuh"

namespace cvllm {

constexpr int kLevThreads = 256;
constexpr int kSk = 48;            // sketch columns
constexpr int kTokTile = 512;      // tokens per smem tile: the engine's chunk (512) is resident, one sketch pass
constexpr int kXtStride = kTokTile + 8;  // 16-bit elements per sketch row (pad: bank spread)

struct LevParams {
  const void* k;
  int64_t k_stride_n, k_stride_h;  // elements
  const int32_t* chunk_offsets;
  const void* phi;  // [128, 48] 16-bit
  const void* sketch;  // optional [N, HKV, 48] 16-bit = round16(K.PHI) from rope.cu; null: derive it from k
  float* raw;       // [N, HKV] fp32 (values already rounded to 16-bit)
  float* gram;      // [n_chunks*HKV][48][48] fp32
  float* mean;      // [n_chunks*HKV][48] fp32
  float* sv;        // [n_chunks*HKV][48][48] fp32 (values rounded to 16-bit)
  uint16_t* xc;     // centred sketch, 16-bit: chunk ch / head h at (c_beg*HKV + h*c_len)*48, [48][c_len] sketch-major
  int HKV;
  float reg;
};

template <bool BF16>
__device__ __forceinline__ float round16(float x) {
  if constexpr (BF16) return __bfloat162float(__float2bfloat16_rn(x));
  else return __half2float(__float2half_rn(x));
}
template <bool BF16>
__device__ __forceinline__ uint16_t to16(float x) {
  if constexpr (BF16) return __bfloat16_as_ushort(__float2bfloat16_rn(x));
  else return __half_as_ushort(__float2half_rn(x));
}
template <bool BF16>
__device__ __forceinline__ float from16(uint16_t x) {
  if constexpr (BF16) return __uint_as_float((uint32_t)x << 16);
  else return __half2float(__ushort_as_half(x));
}

struct LevSmem {
  uint16_t phiT[kSk][128 + 8];        // PHI^T: sketch-major, d contiguous (68-word rows: conflict-free LDS.32 fragments)
  uint16_t xt[kSk][kXtStride];        // X^T tile: sketch-major, token contiguous
  float g[kSk][kSk + 1];              // Gram accumulator (gram kernel) / unused (score kernel)
  float sv[kSk][kSk];                 // round16(V |lam|^-1/2) (score kernel) / staging (gram kernel)
  float colsum[kSk];
  float mean[kSk];
};

// X^T tile for tokens [t0, t0+nt): xt[c][t] = round16(sum_d K[t][d] * PHI[d][c]) (optionally centred)
template <bool BF16>
__device__ __forceinline__ void sketch_tile(LevSmem& sm, const LevParams& p, int h, int tok0, int nt,
                                            bool centre) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, r = lane >> 2, j = lane & 3;
  const char* kb = reinterpret_cast<const char*>(p.k);
  const int n_mt = (nt + 15) / 16;
  if (p.sketch) {
    // precomputed sketch rows (96 B per (token, head)): transpose into xt, zero the m-tile padding
    const uint16_t* xs = reinterpret_cast<const uint16_t*>(p.sketch);
    const int nt16 = n_mt * 16;
    for (int i = threadIdx.x; i < nt16 * (kSk / 8); i += kLevThreads) {
      const int cchunk = i / nt16, t = i - cchunk * nt16;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (t < nt) v = *reinterpret_cast<const uint4*>(xs + ((int64_t)(tok0 + t) * p.HKV + h) * kSk + cchunk * 8);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        uint16_t bits = (uint16_t)(w[e >> 1] >> ((e & 1) * 16));
        if (centre && t < nt)
          bits = to16<BF16>(round16<BF16>(from16<BF16>(bits) - sm.mean[cchunk * 8 + e]));
        sm.xt[cchunk * 8 + e][t] = bits;
      }
    }
    return;
  }
  for (int mt = warp; mt < n_mt; mt += kLevThreads / 32) {
    const int t_lo = mt * 16 + r, t_hi = t_lo + 8;
    auto krow = [&](int t) -> const char* {
      return t < nt ? kb + ((int64_t)(tok0 + t) * p.k_stride_n + (int64_t)h * p.k_stride_h) * 2 : nullptr;
    };
    // natural-k chain (mma.cuh): the bits of the reference's cuBLAS GEMM at compactor.py:158
    FragNat128 alo, ahi;
    alo.load(reinterpret_cast<const uint16_t*>(krow(t_lo)), j);
    ahi.load(reinterpret_cast<const uint16_t*>(krow(t_hi)), j);
#pragma unroll
    for (int nn = 0; nn < kSk / 8; ++nn) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const uint32_t* prow = reinterpret_cast<const uint32_t*>(sm.phiT[nn * 8 + r]);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)
        mma16816<BF16>(acc, alo.w[2 * ks], ahi.w[2 * ks], alo.w[2 * ks + 1], ahi.w[2 * ks + 1], prow[8 * ks + j],
                       prow[8 * ks + 4 + j]);
      // acc: rows t_lo / t_hi, sketch cols nn*8 + 2j, +1
      const int c0 = nn * 8 + 2 * j;
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float x0 = round16<BF16>(acc[e]), x1 = round16<BF16>(acc[2 + e]);
        if (centre) {
          x0 = round16<BF16>(x0 - sm.mean[c0 + e]);
          x1 = round16<BF16>(x1 - sm.mean[c0 + e]);
        }
        sm.xt[c0 + e][t_lo] = (t_lo < nt) ? to16<BF16>(x0) : 0;
        sm.xt[c0 + e][t_hi] = (t_hi < nt) ? to16<BF16>(x1) : 0;
      }
    }
  }
}

__device__ __forceinline__ void load_phi(LevSmem& sm, const LevParams& p) {
  const uint16_t* phi = reinterpret_cast<const uint16_t*>(p.phi);
  for (int i = threadIdx.x; i < 128 * kSk; i += kLevThreads) {
    const int d = i / kSk, c = i - d * kSk;
    sm.phiT[c][d] = phi[i];
  }
}

// ---- kernel 1: grid (n_chunks, HKV): column means + Gram of the centred sketch ----
// out: gram[(ch*HKV+h)][48][48] = fp32(round16(Xc^T Xc)) + reg*I ; mean[(ch*HKV+h)][48]
template <bool BF16>
__global__ void __launch_bounds__(kLevThreads, 2) leverage_gram_kernel(const LevParams p) {
  extern __shared__ __align__(16) unsigned char lev_smem_raw[];
  LevSmem& sm = *reinterpret_cast<LevSmem*>(lev_smem_raw);
  const int ch = blockIdx.x, h = blockIdx.y;
  const int c_beg = p.chunk_offsets[ch], c_len = p.chunk_offsets[ch + 1] - c_beg;
  if (c_len <= 0) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (!p.sketch) load_phi(sm, p);
  if (tid < kSk) sm.colsum[tid] = 0.f;
  for (int i = tid; i < kSk * (kSk + 1); i += kLevThreads) (&sm.g[0][0])[i] = 0.f;
  __syncthreads();

  // ---- pass A: column means of X over the chunk ----
  for (int t0 = 0; t0 < c_len; t0 += kTokTile) {
    const int nt = min(kTokTile, c_len - t0);
    sketch_tile<BF16>(sm, p, h, c_beg + t0, nt, false);
    __syncthreads();
    // 48 columns x 256 threads: warp w sums columns w, w+8, ...
    for (int c = warp; c < kSk; c += kLevThreads / 32) {
      float s = 0.f;
      for (int t = lane; t < nt; t += 32) s += from16<BF16>(sm.xt[c][t]);
      s = warp_sum(s);
      if (lane == 0) sm.colsum[c] += s;
    }
    __syncthreads();
  }
  if (tid < kSk) sm.mean[tid] = round16<BF16>(sm.colsum[tid] / (float)c_len);
  __syncthreads();

  // ---- pass B: Gram of the centred sketch, G[i][j] = sum_t xc[t][i] xc[t][j] ----
  for (int t0 = 0; t0 < c_len; t0 += kTokTile) {
    const int nt = min(kTokTile, c_len - t0);
    // Centre (one tile: the uncentred sketch is still resident; several tiles: re-derived centred), zero the
    // padded token columns (partial k16 steps must contribute nothing) and hand the centred sketch to the score
    // kernel — one vectorised pass, a warp per sketch row, 8 tokens per lane.  (A flat `i / nt` element loop
    // here was 65 % of the kernel's instructions: ncu r02g.)
    const bool resident = c_len <= kTokTile;
    if (!resident) {
      sketch_tile<BF16>(sm, p, h, c_beg + t0, nt, true);
      __syncthreads();
    }
    const int nt16 = (nt + 15) & ~15;
    {
      uint16_t* xc = p.xc + ((int64_t)c_beg * p.HKV + (int64_t)h * c_len) * kSk;
      const bool vec_ok = (c_len & 7) == 0;
      for (int c = warp; c < kSk; c += kLevThreads / 32) {
        const float mu = resident ? sm.mean[c] : 0.f;
        for (int t8 = lane * 8; t8 < nt16; t8 += 256) {
          uint4 v = *reinterpret_cast<const uint4*>(&sm.xt[c][t8]);
          uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float lo = from16<BF16>((uint16_t)(w4[e] & 0xffffu)), hi = from16<BF16>((uint16_t)(w4[e] >> 16));
            if (resident) { lo = round16<BF16>(lo - mu); hi = round16<BF16>(hi - mu); }
            const uint32_t blo = (t8 + 2 * e < nt) ? to16<BF16>(lo) : 0u;
            const uint32_t bhi = (t8 + 2 * e + 1 < nt) ? to16<BF16>(hi) : 0u;
            w4[e] = blo | (bhi << 16);
          }
          v = make_uint4(w4[0], w4[1], w4[2], w4[3]);
          *reinterpret_cast<uint4*>(&sm.xt[c][t8]) = v;
          uint16_t* dst = xc + (int64_t)c * c_len + t0 + t8;
          if (vec_ok && t8 + 8 <= nt) {
            *reinterpret_cast<uint4*>(dst) = v;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e)
              if (t8 + e < nt) dst[e] = (uint16_t)(w4[e >> 1] >> ((e & 1) * 16));
          }
        }
      }
    }
    __syncthreads();
    const int r = lane >> 2, j = lane & 3;
    // 3 m-tiles (16 sketch rows) x 6 n-tiles (8 sketch cols) = 18 output tiles over 8 warps
    for (int tile = warp; tile < 18; tile += kLevThreads / 32) {
      const int mt = tile / 6, nn = tile - mt * 6;
      // the accumulator is CARRIED across token tiles (not summed per tile and added): one sequential
      // k16 chain over the whole chunk, the order a non-split-K library GEMM accumulates in — the Gram is
      // rounded to 16 bit next, and on low-rank keys one flipped ulp there moves a chunk's scores by O(1)
      const int gi = mt * 16 + r, gj = nn * 8 + 2 * j;
      float acc[4] = {sm.g[gi][gj], sm.g[gi][gj + 1], sm.g[gi + 8][gj], sm.g[gi + 8][gj + 1]};
      const uint16_t* a_lo = sm.xt[mt * 16 + r];
      const uint16_t* a_hi = sm.xt[mt * 16 + r + 8];
      const uint16_t* b_r = sm.xt[nn * 8 + r];
      for (int k0 = 0; k0 < nt16; k0 += 16) {
        const uint32_t a0 = *reinterpret_cast<const uint32_t*>(a_lo + k0 + 2 * j);
        const uint32_t a1 = *reinterpret_cast<const uint32_t*>(a_hi + k0 + 2 * j);
        const uint32_t a2 = *reinterpret_cast<const uint32_t*>(a_lo + k0 + 8 + 2 * j);
        const uint32_t a3 = *reinterpret_cast<const uint32_t*>(a_hi + k0 + 8 + 2 * j);
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(b_r + k0 + 2 * j);
        const uint32_t b1 = *reinterpret_cast<const uint32_t*>(b_r + k0 + 8 + 2 * j);
        mma16816<BF16>(acc, a0, a1, a2, a3, b0, b1);
      }
      sm.g[gi][gj] = acc[0];
      sm.g[gi][gj + 1] = acc[1];
      sm.g[gi + 8][gj] = acc[2];
      sm.g[gi + 8][gj + 1] = acc[3];
    }
    __syncthreads();
  }
  // round the Gram to the 16-bit dtype (compactor.py:166,171,173), add the regulariser.
  // Symmetrise from the upper triangle so rounding cannot make G asymmetric.
  float* gout = p.gram + (int64_t)(ch * p.HKV + h) * kSk * kSk;
  for (int i = tid; i < kSk * kSk; i += kLevThreads) {
    const int a = i / kSk, b = i - a * kSk;
    gout[i] = round16<BF16>(a <= b ? sm.g[a][b] : sm.g[b][a]) + ((a == b) ? p.reg : 0.f);
  }
  if (tid < kSk) p.mean[(int64_t)(ch * p.HKV + h) * kSk + tid] = sm.mean[tid];
}

// ---- kernel 2: symmetric eigendecomposition in fp64, one 128-thread CTA per 48x48 matrix ----
// Why fp64: the reference's gesvda result agrees with the EXACT (fp64) eigendecomposition of the same
// fp32 Gram to 99.3 % of the scores within 2e-2 on low-rank keys, while any norm-wise backward-stable
// fp32 solver (cuSOLVER gesvd / gesvdj, and round 1's fp32 QL here) only reaches 60-85 %
// (profiles/r02a_leverage_selfdistance_before.json): |lam| ranges over 1e-2 .. 1e5, and an absolute
// error of eps32 * |lam|_max ~ 1e-2 destroys the small eigenpairs that dominate x^T |G|^-1 x.
//
// Structure (EISPACK tred2 / tql2 restated for a CTA):
//   phase 1  all 4 warps: Householder reduction to tridiagonal form (matvec and rank-2 update spread over
//            128 threads, lanes own COLUMNS so every shared-memory access is conflict-free).
//   phase 2  warps 1-3: accumulate the reflectors into Z (in place)        } concurrently
//   phase 3  warp 0: implicit-shift QL on (d, e) ONLY — the rotations do   }
//            not depend on Z, so they are streamed through a ring in shared memory as (c, s) pairs, one
//            sweep at a time, and warps 1-2 apply them to the rows of Z behind the producer.
// The QL recurrence is the latency chain (~2000 dependent rotations).  It is rewritten so that one
// rotation costs fma -> rsqrt -> mul on the dependent path:  with r^2 = p^2 + e_i^2, c = p/r, s = e_i/r,
//   p_next = c d_i - s (c_prev e_i) = (p d_i - e_i^2 c_prev) / r,
// and the numerator does not wait for 1/r.  1/r = rsqrt.approx.f64 + one third-order (Halley) step.
// out: sv[mat][48][48] = round16(V |lam|^-1/2)   (compactor.py:197: S of a symmetric G is |lam|);
// the column (eigenpair) order is arbitrary, the score sums over all of them.
constexpr int kEigThreads = 128;
constexpr int kEigRing = 1024;  // (c, s) entries; one sweep needs <= 48
struct EigSmem {
  double a[kSk][kSk + 1];  // A, then Z in place
  double d[kSk], e[kSk], hh[kSk], hinv[kSk], u[kSk], q[kSk];
  double part[2][64];
  double2 ring[kEigRing];
  double2 pr[kSk];  // (p_i, 1/r_i) of the current QL sweep
  int produced, done;
  int consumed[2];
};

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double rsqrt64(double x) {  // x > 0, normal: seed 2^-22, Halley -> fp64
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x * y, y, 1.0);
  return fma(y * e, fma(0.375, e, 0.5), y);
}
__device__ __forceinline__ double rcp64(double x) {  // x != 0, normal
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  return fma(y, e, y);
}
__device__ __forceinline__ void bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ int ld_volatile_s(const int* p) { return *reinterpret_cast<const volatile int*>(p); }
__device__ __forceinline__ void st_volatile_s(int* p, int v) { *reinterpret_cast<volatile int*>(p) = v; }

template <bool BF16>
__global__ void __launch_bounds__(kEigThreads) leverage_eig_kernel(const float* __restrict__ gram,
                                                                   float* __restrict__ sv_out,
                                                                   const int32_t* __restrict__ chunk_offsets,
                                                                   int HKV, int mats, double* __restrict__ dbg_lam,
                                                                   double* __restrict__ dbg_vec,
                                                                   long long* __restrict__ dbg_clk) {
  __shared__ __align__(16) EigSmem s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int mat = blockIdx.x;
  if (mat >= mats) return;
  if (chunk_offsets) {
    const int ch = mat / HKV;
    if (chunk_offsets[ch + 1] - chunk_offsets[ch] <= 0) return;
  }
  constexpr int n = kSk;
  constexpr int R1 = kEigRing - 1;
  const float* gin = gram + (int64_t)mat * n * n;
  for (int i = tid; i < n * n; i += kEigThreads) s.a[i / n][i % n] = (double)gin[i];
  if (tid == 0) { s.produced = 0; s.done = 0; s.consumed[0] = 0; s.consumed[1] = 0; }
  __syncthreads();
  if (dbg_clk && tid == 0) dbg_clk[mat * 6 + 0] = clock64();

  // ---- phase 1: Householder reduction to tridiagonal form; u^(i) is parked in column i (rows < i) ----
  {
    const int col = tid & 63, half = tid >> 6;
    for (int i = n - 1; i >= 2; --i) {
      // every warp forms the reflector of row i redundantly (no broadcast round trip)
      double x0 = (lane < i) ? s.a[i][lane] : 0.0;
      double x1 = (lane + 32 < i) ? s.a[i][lane + 32] : 0.0;
      const double h = warp_sum_d(fma(x0, x0, x1 * x1));
      if (!(h > 1e-280)) {  // row already reduced (or not finite): H_i = I.  Uniform over the CTA.
        if (tid == 0) { s.e[i] = 0.0; s.hh[i] = 0.0; s.hinv[i] = 0.0; s.d[i] = s.a[i][i]; }
        continue;
      }
      const double f = s.a[i][i - 1];
      const double g = -copysign(h * rsqrt64(h), f);
      const double hi = h - f * g;
      const double rh = rcp64(hi);
      if (lane == ((i - 1) & 31)) {
        if (i - 1 < 32) x0 = f - g; else x1 = f - g;
      }
      if (warp == 0) {
        if (lane < i) { s.u[lane] = x0; s.a[lane][i] = x0; }
        if (lane + 32 < i) { s.u[lane + 32] = x1; s.a[lane + 32][i] = x1; }
        if (lane == 0) { s.e[i] = g; s.hh[i] = hi; s.hinv[i] = rh; s.d[i] = s.a[i][i]; }
      }
      __syncthreads();
      // p = A u / h: thread (col, half) sums rows k = half, half+2, ... of column col (A is symmetric)
      if (col < i) {
        double acc0 = 0.0, acc1 = 0.0;
        int k = half;
        for (; k + 2 < i; k += 4) {
          acc0 = fma(s.a[k][col], s.u[k], acc0);
          acc1 = fma(s.a[k + 2][col], s.u[k + 2], acc1);
        }
        for (; k < i; k += 2) acc0 = fma(s.a[k][col], s.u[k], acc0);
        s.part[half][col] = acc0 + acc1;
      }
      __syncthreads();
      const double p0 = (lane < i) ? (s.part[0][lane] + s.part[1][lane]) * rh : 0.0;
      const double p1 = (lane + 32 < i) ? (s.part[0][lane + 32] + s.part[1][lane + 32]) * rh : 0.0;
      const double K = warp_sum_d(fma(x0, p0, x1 * p1)) * 0.5 * rh;
      if (warp == 0) {
        if (lane < i) s.q[lane] = fma(-K, x0, p0);
        if (lane + 32 < i) s.q[lane + 32] = fma(-K, x1, p1);
      }
      __syncthreads();
      // A <- A - u q^T - q u^T on the leading i x i block (kept fully symmetric)
      if (col < i) {
        const double uc = s.u[col], qc = s.q[col];
        for (int j = half; j < i; j += 2) s.a[j][col] -= fma(s.u[j], qc, s.q[j] * uc);
      }
      __syncthreads();
    }
    if (tid == 0) {
      s.d[0] = s.a[0][0]; s.d[1] = s.a[1][1];
      s.e[1] = s.a[1][0]; s.e[0] = 0.0;
      s.a[0][0] = 1.0;
    }
    __syncthreads();
  }

  if (dbg_clk && tid == 0) dbg_clk[mat * 6 + 1] = clock64();
  if (warp == 0) {
    // ---- phase 3 (producer): implicit-shift QL on (d, e); rotations -> ring ----
    {
      const double e0 = s.e[lane + 1];
      const double e1 = (lane + 33 < n) ? s.e[lane + 33] : 0.0;
      __syncwarp();
      s.e[lane] = e0;
      if (lane + 32 < n) s.e[lane + 32] = e1;
      __syncwarp();
    }
    double fsh = 0.0, tst1 = 0.0;
    int prod = 0;
    const double eps = 2.3e-13;  // 2^-42: |lam| is needed to ~1e-4 relative at |lam| >= reg = 5e-3
    for (int l = 0; l < n; ++l) {
      tst1 = fmax(tst1, fabs(s.d[l]) + fabs(s.e[l]));
      // first m >= l with a negligible e[m] (e[n-1] == 0 ends the scan): two ballots instead of a serial walk
      int m;
      {
        const double thr = eps * tst1;
        const bool n0 = (lane >= l) && !(fabs(s.e[lane]) > thr);
        const bool n1 = (lane + 32 >= l) && (lane + 32 >= n || !(fabs(s.e[lane + 32 < n ? lane + 32 : n - 1]) > thr));
        const unsigned b0 = __ballot_sync(0xffffffffu, n0), b1 = __ballot_sync(0xffffffffu, n1);
        m = b0 ? (__ffs(b0) - 1) : (32 + __ffs(b1) - 1);
        if (m > n - 1) m = n - 1;
      }
      if (m > l) {
        int iter = 0;
        double el_abs;
        do {
          ++iter;
          const double g0 = s.d[l], el = s.e[l], el1 = s.e[l + 1];
          const double pp = (s.d[l + 1] - g0) * rcp64(2.0 * el);
          const double x2 = fma(pp, pp, 1.0);
          double rr = x2 * rsqrt64(x2);
          if (pp < 0.0) rr = -rr;
          const double ppr = pp + rr;
          const double inv = rcp64(ppr);
          const double dl = el * inv, dl1 = el * ppr;
          const double hs = g0 - dl;
          const double dj0 = s.d[lane], dj1 = (lane + 32 < n) ? s.d[lane + 32] : 0.0;
          __syncwarp();
          if (lane >= l + 2) s.d[lane] = dj0 - hs;
          if (lane + 32 >= l + 2 && lane + 32 < n) s.d[lane + 32] = dj1 - hs;
          if (lane == 0) { s.d[l] = dl; s.d[l + 1] = dl1; }
          __syncwarp();
          fsh += hs;
          double p = s.d[m];
          // ring space for this sweep: header + (m - l) rotations
          const int need = m - l + 1;
          while (prod + need - min(ld_volatile_s(&s.consumed[0]), ld_volatile_s(&s.consumed[1])) > kEigRing)
            __nanosleep(64);
          if (lane == 0) s.ring[prod & R1] = make_double2((double)l, (double)m);
          // Warps issue in order, so anything between two links of the dependent chain delays it.  The loop
          // therefore carries ONLY the recurrence (12 fp64 ops, p -> r^2 -> 1/r -> p on the chain) and parks
          // (p_i, 1/r_i) in shared memory; the new d / e entries and the (c, s) ring records of the whole sweep
          // are then formed by the lanes in parallel, one rotation per lane.  198 -> 86 cycles per rotation
          // (tools/ubench/ql_chain.cu).
          {
            double c = 1.0;
            double ei = s.e[m - 1], di = s.d[m - 1];
            for (int i = m - 1; i >= l; --i) {
              const double ein = (i > l) ? s.e[i - 1] : 0.0, din = (i > l) ? s.d[i - 1] : 0.0;  // prefetch
              const double t = ei * ei;
              const double r2 = fma(p, p, fma(ei, ei, 1e-300));  // > 0: no guard on the chain
              const double qv = fma(p, di, -(t * c));            // r (c_new d_i - s_new g): does not wait for 1/r
              double y;
              asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r2));
              const double er = fma(-r2 * y, y, 1.0);
              const double ri = fma(y * er, fma(0.375, er, 0.5), y);
              if (lane == 0) s.pr[i] = make_double2(p, ri);
              c = p * ri;
              p = ri * qv;
              ei = ein; di = din;
            }
          }
          __syncwarp();
          // rotation i = m-1-k on lane k (and k+32): reads first, then writes (e[i+1] is another lane's e_i)
          double o_e[2], o_d[2], o_c[2], o_s[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int k = lane + 32 * u, i = m - 1 - k;
            o_e[u] = o_d[u] = o_c[u] = o_s[u] = 0.0;
            if (i >= l) {
              const double2 a = s.pr[i];
              const double e_i = s.e[i], d_i = s.d[i];
              const double c_i = a.x * a.y, s_i = e_i * a.y;
              double c_p = 1.0, s_p = 0.0;
              if (i < m - 1) {
                const double2 b = s.pr[i + 1];
                c_p = b.x * b.y;
                s_p = s.e[i + 1] * b.y;
              }
              const double g = c_p * e_i, h = c_p * a.x;
              const double r = fma(a.x, a.x, e_i * e_i) * a.y;
              o_e[u] = s_p * r;
              o_d[u] = h + s_i * fma(c_i, g, s_i * d_i);
              o_c[u] = c_i;
              o_s[u] = s_i;
            }
          }
          __syncwarp();
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int k = lane + 32 * u, i = m - 1 - k;
            if (i >= l) {
              s.e[i + 1] = o_e[u];
              s.d[i + 1] = o_d[u];
              s.ring[(prod + 1 + k) & R1] = make_double2(o_c[u], o_s[u]);
            }
          }
          __syncwarp();
          // s_l, s_{l+1}, c_l, c_{l+2} of this sweep (absent rotations: s = 0, c = 1)
          const int nrot = m - l;
          const double2 q0 = s.ring[(prod + nrot) & R1];
          const double sn = q0.y, c = q0.x;
          const double s2 = (nrot >= 2) ? s.ring[(prod + nrot - 1) & R1].y : 0.0;
          const double c3 = (nrot >= 3) ? s.ring[(prod + nrot - 2) & R1].x : 1.0;
          const double pe = -sn * s2 * c3 * el1 * inv;  // = -s s2 c3 e[l+1] e[l] / d[l+1] with d[l+1] = e[l] (pp + rr)
          el_abs = fabs(sn * pe);
          __syncwarp();
          if (lane == 0) { s.e[l] = sn * pe; s.d[l] = c * pe; }
          prod += need;
          __threadfence_block();
          if (lane == 0) st_volatile_s(&s.produced, prod);
          __syncwarp();
        } while (el_abs > eps * tst1 && iter < 60);
      }
      __syncwarp();
      if (lane == 0) { s.d[l] += fsh; s.e[l] = 0.0; }
      __syncwarp();
    }
    __threadfence_block();
    if (lane == 0) st_volatile_s(&s.done, 1);
    if (dbg_clk && lane == 0) { dbg_clk[mat * 6 + 2] = clock64(); dbg_clk[mat * 6 + 5] = prod; }
  } else {
    // ---- phase 2 (warps 1-3): Z = H_{n-1} ... H_2, built in place; thread owns (column, row parity) ----
    const int t2 = tid - 32, zc = t2 % n, zp = t2 / n;
    for (int i = 2; i <= n; ++i) {
      // row / column i-1 of the growing block become the identity (column i-1 held u^(i-1), consumed)
      for (int idx = t2; idx < i - 1; idx += 96) { s.a[i - 1][idx] = 0.0; s.a[idx][i - 1] = 0.0; }
      if (t2 == 0) s.a[i - 1][i - 1] = 1.0;
      bar_sync(1, 96);
      if (i < n && s.hh[i] != 0.0) {
        if (zc < i) {
          double acc0 = 0.0, acc1 = 0.0;
          int k = zp;
          for (; k + 2 < i; k += 4) {
            acc0 = fma(s.a[k][i], s.a[k][zc], acc0);
            acc1 = fma(s.a[k + 2][i], s.a[k + 2][zc], acc1);
          }
          for (; k < i; k += 2) acc0 = fma(s.a[k][i], s.a[k][zc], acc0);
          s.part[zp][zc] = acc0 + acc1;
        }
        bar_sync(1, 96);
        if (zc < i) {
          const double g = (s.part[0][zc] + s.part[1][zc]) * s.hinv[i];
          for (int k = zp; k < i; k += 2) s.a[k][zc] = fma(-g, s.a[k][i], s.a[k][zc]);
        }
        bar_sync(1, 96);
      }
    }
    if (dbg_clk && t2 == 0) dbg_clk[mat * 6 + 3] = clock64();
    // ---- phase 3 (consumers, warps 1-2): apply the rotation stream to the rows of Z ----
    if (warp <= 2) {
      const bool active = t2 < n;
      const int row = active ? t2 : 0;
      int cons = 0;
      for (;;) {
        const int dn = ld_volatile_s(&s.done);
        const int pr = ld_volatile_s(&s.produced);
        __threadfence_block();
        if (pr == cons) {
          if (dn) break;
          __nanosleep(100);
          continue;
        }
        while (cons < pr) {  // whole sweeps are published at once
          const double2 hd = s.ring[cons & R1];
          const int l = (int)hd.x, m = (int)hd.y, cnt = m - l;
          double z = s.a[row][m];
          int slot = (cons + 1) & R1;
          for (int k = 0, i = m - 1; k < cnt; ++k, --i) {
            const double2 cs = s.ring[slot];
            slot = (slot + 1) & R1;
            const double y = s.a[row][i];
            if (active) s.a[row][i + 1] = fma(cs.y, y, cs.x * z);
            z = fma(cs.x, y, -cs.y * z);
          }
          if (active) s.a[row][l] = z;
          cons += cnt + 1;
        }
        __syncwarp();
        __threadfence_block();
        if (lane == 0) st_volatile_s(&s.consumed[warp - 1], cons);
      }
    }
  }
  __syncthreads();
  if (tid < n) s.u[tid] = rsqrt(fabs(s.d[tid]));
  __syncthreads();
  float* out = sv_out + (int64_t)mat * n * n;
  for (int i = tid; i < n * n; i += kEigThreads) {
    const int a = i / n, b = i - a * n;
    out[i] = round16<BF16>((float)(s.a[a][b] * s.u[b]));
    if (dbg_vec) dbg_vec[(int64_t)mat * n * n + i] = s.a[a][b];
  }
  if (dbg_lam && tid < n) dbg_lam[(int64_t)mat * n + tid] = s.d[tid];
  if (dbg_clk && tid == 0) dbg_clk[mat * 6 + 4] = clock64();
}

// ---- kernel 3: grid (n_chunks, HKV): score_t = round16(sum_j round16(round16(xc_t . sv_j)^2)) ----
template <bool BF16>
__global__ void __launch_bounds__(kLevThreads, 2) leverage_score_kernel(const LevParams p) {
  extern __shared__ __align__(16) unsigned char lev_smem_raw[];
  LevSmem& sm = *reinterpret_cast<LevSmem*>(lev_smem_raw);
  const int ch = blockIdx.x, h = blockIdx.y;
  const int c_beg = p.chunk_offsets[ch], c_len = p.chunk_offsets[ch + 1] - c_beg;
  if (c_len <= 0) return;
  const int tid = threadIdx.x;
  {
    const float* svin = p.sv + (int64_t)(ch * p.HKV + h) * kSk * kSk;
    for (int i = tid; i < kSk * kSk; i += kLevThreads) (&sm.sv[0][0])[i] = svin[i];
    if (tid < kSk) sm.mean[tid] = p.mean[(int64_t)(ch * p.HKV + h) * kSk + tid];
  }
  __syncthreads();
  // B fragments of U = Xc . SV (k = sketch column c, n = eigen-column): SV's entries are 16-bit values held in fp32
  const int warp = tid >> 5, lane = tid & 31, r = lane >> 2, j = lane & 3;
  uint32_t bfr[3][6][2];
#pragma unroll
  for (int ks = 0; ks < 3; ++ks)
#pragma unroll
    for (int nn = 0; nn < 6; ++nn) {
      bfr[ks][nn][0] = pack2<BF16>(sm.sv[16 * ks + 2 * j][nn * 8 + r], sm.sv[16 * ks + 2 * j + 1][nn * 8 + r]);
      bfr[ks][nn][1] = pack2<BF16>(sm.sv[16 * ks + 8 + 2 * j][nn * 8 + r], sm.sv[16 * ks + 9 + 2 * j][nn * 8 + r]);
    }
  const bool vec_ok = (c_len & 7) == 0;
  for (int t0 = 0; t0 < c_len; t0 += kTokTile) {
    const int nt = min(kTokTile, c_len - t0);
    {
      const uint16_t* xc = p.xc + ((int64_t)c_beg * p.HKV + (int64_t)h * c_len) * kSk;
      for (int c = warp; c < kSk; c += kLevThreads / 32) {
        const uint16_t* src = xc + (int64_t)c * c_len + t0;
        for (int t8 = lane * 8; t8 < nt; t8 += 256) {
          if (vec_ok) {
            *reinterpret_cast<uint4*>(&sm.xt[c][t8]) = *reinterpret_cast<const uint4*>(src + t8);
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) sm.xt[c][t8 + e] = (t8 + e < nt) ? src[t8 + e] : (uint16_t)0;
          }
        }
      }
    }
    __syncthreads();
    // U[16 tokens x 48] per m-tile on mma.sync (the scalar form was 48 x 48 FFMA per token); the A operand is read
    // straight out of the sketch-major tile with ldmatrix.trans
    const int n_mt = (nt + 15) / 16;
    for (int mt = warp; mt < n_mt; mt += kLevThreads / 32) {
      float acc[6][4];
#pragma unroll
      for (int nn = 0; nn < 6; ++nn) acc[nn][0] = acc[nn][1] = acc[nn][2] = acc[nn][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 3; ++ks) {
        const int mid = lane >> 3;  // 8x8 matrix this lane addresses: (k half, token half)
        const uint16_t* rowp = &sm.xt[16 * ks + (mid >> 1) * 8 + (lane & 7)][mt * 16 + (mid & 1) * 8];
        uint32_t a0, a1, a2, a3;
        asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                     : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3)
                     : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(rowp))));
#pragma unroll
        for (int nn = 0; nn < 6; ++nn) mma16816<BF16>(acc[nn], a0, a1, a2, a3, bfr[ks][nn][0], bfr[ks][nn][1]);
      }
      // score = round16(sum_j round16(round16(u_j)^2)) (compactor.py:209-210), rows r and r + 8 of the m-tile
      float s_lo = 0.f, s_hi = 0.f;
#pragma unroll
      for (int nn = 0; nn < 6; ++nn) {
        const float u0 = round16<BF16>(acc[nn][0]), u1 = round16<BF16>(acc[nn][1]);
        const float u2 = round16<BF16>(acc[nn][2]), u3 = round16<BF16>(acc[nn][3]);
        s_lo += round16<BF16>(u0 * u0) + round16<BF16>(u1 * u1);
        s_hi += round16<BF16>(u2 * u2) + round16<BF16>(u3 * u3);
      }
      s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1);
      s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2);
      s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1);
      s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);
      if (j == 0) {
        const int t_lo = mt * 16 + r, t_hi = t_lo + 8;
        if (t_lo < nt) p.raw[(int64_t)(c_beg + t0 + t_lo) * p.HKV + h] = fmaxf(round16<BF16>(s_lo), 0.f);
        if (t_hi < nt) p.raw[(int64_t)(c_beg + t0 + t_hi) * p.HKV + h] = fmaxf(round16<BF16>(s_hi), 0.f);
      }
    }
    __syncthreads();
  }
}

// grid (n_chunks): z-score a chunk over all heads (compactor.py:224-269) and cast to the output dtype
template <bool BF16>
__global__ void __launch_bounds__(256) leverage_finalize_kernel(const float* __restrict__ raw,
                                                                const int32_t* __restrict__ chunk_offsets,
                                                                void* __restrict__ out, int HKV, int normalize) {
  const int ch = blockIdx.x;
  const int beg = chunk_offsets[ch], len = chunk_offsets[ch + 1] - beg;
  if (len <= 0) return;
  const int total = len * HKV;
  const float* src = raw + (int64_t)beg * HKV;
  uint16_t* dst = reinterpret_cast<uint16_t*>(out) + (int64_t)beg * HKV;
  __shared__ float s_red[2][8];
  __shared__ float s_stat[2];
  float mean = 0.f, invstd = 1.f;
  if (normalize) {
    float a1 = 0.f, a2 = 0.f;
    for (int i = threadIdx.x; i < total; i += 256) {
      const float v = src[i];
      a1 += v;
      a2 += v * v;
    }
    a1 = warp_sum(a1);
    a2 = warp_sum(a2);
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = a1; s_red[1][threadIdx.x >> 5] = a2; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float t1 = 0.f, t2 = 0.f;
      for (int w = 0; w < 8; ++w) { t1 += s_red[0][w]; t2 += s_red[1][w]; }
      const float cnt = (float)total;
      const float m = t1 / cnt;
      const float var = fmaxf(t2 / cnt - m * m, 0.f);
      s_stat[0] = m;
      s_stat[1] = 1.0f / sqrtf(var);
    }
    __syncthreads();
    mean = s_stat[0];
    invstd = s_stat[1];
  }
  for (int i = threadIdx.x; i < total; i += 256) {
    float v = src[i];
    if (normalize) v = (v - mean) * invstd;
    dst[i] = to16<BF16>(v);
  }
}

}  // namespace cvllm

using namespace cvllm;

static size_t lev_align(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" size_t cvllm_leverage_scores_workspace_bytes(int N, int HKV, int sketch, int n_chunks) {
  (void)sketch;
  if (N <= 0 || HKV <= 0) return 256;
  const size_t mats = (size_t)(n_chunks > 0 ? n_chunks : 0) * HKV;
  return lev_align((size_t)N * HKV * sizeof(float)) + 2 * lev_align(mats * kSk * kSk * sizeof(float)) +
         lev_align(mats * kSk * sizeof(float)) + lev_align((size_t)N * HKV * kSk * sizeof(uint16_t));
}

static int leverage_launch(const void* k_pre, int64_t k_stride_n, int64_t k_stride_h, const void* sketch_rows,
                           const int32_t* chunk_offsets, int n_chunks, int max_chunk,
                           const void* phi, void* out_scores, int N, int HKV, int D, int sketch,
                           float regularizer, int normalize, int dtype, void* workspace,
                           size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(N >= 0 && HKV > 0 && n_chunks >= 0 && max_chunk >= 0);
  if (N == 0 || n_chunks == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG((sketch_rows || (k_pre && phi)) && chunk_offsets && out_scores && workspace);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_SUPPORTED(D == 128 && sketch == kSk);
  CVLLM_CHECK_ALIGN(aligned16(k_pre) && aligned16(sketch_rows) && k_stride_n % 8 == 0 && k_stride_h % 8 == 0);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_leverage_scores_workspace_bytes(N, HKV, sketch, n_chunks));
  auto st = static_cast<cudaStream_t>(stream);
  const size_t mats = (size_t)n_chunks * HKV;
  char* wsb = reinterpret_cast<char*>(workspace);
  LevParams p;
  p.k = k_pre; p.k_stride_n = k_stride_n; p.k_stride_h = k_stride_h; p.chunk_offsets = chunk_offsets;
  p.phi = phi; p.sketch = sketch_rows; p.HKV = HKV; p.reg = regularizer;
  p.raw = reinterpret_cast<float*>(wsb);
  wsb += lev_align((size_t)N * HKV * sizeof(float));
  p.gram = reinterpret_cast<float*>(wsb);
  wsb += lev_align(mats * kSk * kSk * sizeof(float));
  p.sv = reinterpret_cast<float*>(wsb);
  wsb += lev_align(mats * kSk * kSk * sizeof(float));
  p.mean = reinterpret_cast<float*>(wsb);
  wsb += lev_align(mats * kSk * sizeof(float));
  p.xc = reinterpret_cast<uint16_t*>(wsb);
  const size_t smem = sizeof(LevSmem);
  static std::atomic<uint64_t> attr_g[2] = {{0}, {0}}, attr_s[2] = {{0}, {0}};
  const dim3 grid(n_chunks, HKV);
  if (dtype == CVLLM_BF16) {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)leverage_gram_kernel<true>, (int)smem, attr_g[1]));
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)leverage_score_kernel<true>, (int)smem, attr_s[1]));
    leverage_gram_kernel<true><<<grid, kLevThreads, smem, st>>>(p);
    leverage_eig_kernel<true><<<(unsigned)mats, kEigThreads, 0, st>>>(p.gram, p.sv, chunk_offsets, HKV, (int)mats, nullptr, nullptr, nullptr);
    leverage_score_kernel<true><<<grid, kLevThreads, smem, st>>>(p);
    leverage_finalize_kernel<true><<<n_chunks, 256, 0, st>>>(p.raw, chunk_offsets, out_scores, HKV, normalize);
  } else {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)leverage_gram_kernel<false>, (int)smem, attr_g[0]));
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)leverage_score_kernel<false>, (int)smem, attr_s[0]));
    leverage_gram_kernel<false><<<grid, kLevThreads, smem, st>>>(p);
    leverage_eig_kernel<false><<<(unsigned)mats, kEigThreads, 0, st>>>(p.gram, p.sv, chunk_offsets, HKV, (int)mats, nullptr, nullptr, nullptr);
    leverage_score_kernel<false><<<grid, kLevThreads, smem, st>>>(p);
    leverage_finalize_kernel<false><<<n_chunks, 256, 0, st>>>(p.raw, chunk_offsets, out_scores, HKV, normalize);
  }
  CVLLM_CHECK_LAUNCHES(4);
  return CVLLM_OK;
}

extern "C" int cvllm_leverage_scores(const void* k_pre, int64_t k_stride_n, int64_t k_stride_h,
                                     const int32_t* chunk_offsets, int n_chunks, int max_chunk,
                                     const void* phi, void* out_scores, int N, int HKV, int D, int sketch,
                                     float regularizer, int normalize, int dtype, void* workspace,
                                     size_t workspace_bytes, cvllm_stream_t stream) {
  if (N > 0 && n_chunks > 0) CVLLM_CHECK_ARG(k_pre != nullptr);
  return leverage_launch(k_pre, k_stride_n, k_stride_h, nullptr, chunk_offsets, n_chunks, max_chunk, phi,
                         out_scores, N, HKV, D, sketch, regularizer, normalize, dtype, workspace,
                         workspace_bytes, stream);
}

extern "C" int cvllm_leverage_scores_from_sketch(const void* sketch_rows, const int32_t* chunk_offsets,
                                                 int n_chunks, int max_chunk, const void* phi,
                                                 void* out_scores, int N, int HKV, int sketch,
                                                 float regularizer, int normalize, int dtype,
                                                 void* workspace, size_t workspace_bytes,
                                                 cvllm_stream_t stream) {
  if (N > 0 && n_chunks > 0) CVLLM_CHECK_ARG(sketch_rows != nullptr);
  return leverage_launch(nullptr, 0, 0, sketch_rows, chunk_offsets, n_chunks, max_chunk, phi, out_scores, N,
                         HKV, 128, sketch, regularizer, normalize, dtype, workspace, workspace_bytes, stream);
}

// Test entry (tests/test_gpu_scores.py): the eigensolver alone.  gram: fp32 [mats, 48, 48] symmetric;
// lam_out: fp64 [mats, 48]; vec_out: fp64 [mats, 48, 48] (columns = eigenvectors, arbitrary order);
// sv_out: fp32 [mats, 48, 48] = round16(V |lam|^-1/2), the product output.
extern "C" int cvllm_selftest_eig(const float* gram, double* lam_out, double* vec_out, float* sv_out, int mats,
                                  int dtype, long long* clk_out, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(gram && lam_out && vec_out && sv_out && mats > 0);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  auto st = static_cast<cudaStream_t>(stream);
  if (dtype == CVLLM_BF16)
    leverage_eig_kernel<true><<<mats, kEigThreads, 0, st>>>(gram, sv_out, nullptr, 1, mats, lam_out, vec_out, clk_out);
  else
    leverage_eig_kernel<false><<<mats, kEigThreads, 0, st>>>(gram, sv_out, nullptr, 1, mats, lam_out, vec_out, clk_out);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
