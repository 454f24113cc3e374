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
constexpr int kTokTile = 256;      // tokens per smem tile
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
  uint16_t phiT[kSk][128 + 32];       // PHI^T: sketch-major, d contiguous (+64 B pad: conflict-free LDS.128)
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
    AFrag128 a;
    a.load(krow(t_lo), krow(t_hi), j);
#pragma unroll
    for (int nn = 0; nn < kSk / 8; ++nn) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const uint16_t* prow = sm.phiT[nn * 8 + r];
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const uint4 pb = *reinterpret_cast<const uint4*>(prow + (j + 4 * cc) * 8);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int ks = 2 * cc + half;
          uint32_t a0, a1, a2, a3, b0, b1;
          a.regs(ks, a0, a1, a2, a3);
          bfrag_regs(pb, ks, b0, b1);
          mma16816<BF16>(acc, a0, a1, a2, a3, b0, b1);
        }
      }
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
    // one tile and the uncentred sketch is still resident: centre it in place
    if (c_len <= kTokTile) {
      for (int i = tid; i < kSk * nt; i += kLevThreads) {
        const int c = i / nt, t = i - c * nt;
        sm.xt[c][t] = to16<BF16>(round16<BF16>(from16<BF16>(sm.xt[c][t]) - sm.mean[c]));
      }
    } else {
      sketch_tile<BF16>(sm, p, h, c_beg + t0, nt, true);
    }
    __syncthreads();
    // zero the padded token columns so partial k16 steps contribute nothing
    const int nt16 = (nt + 15) & ~15;
    for (int i = tid; i < kSk * (nt16 - nt); i += kLevThreads) {
      const int c = i / (nt16 - nt), t = nt + i % (nt16 - nt);
      sm.xt[c][t] = 0;
    }
    __syncthreads();
    const int r = lane >> 2, j = lane & 3;
    // 3 m-tiles (16 sketch rows) x 6 n-tiles (8 sketch cols) = 18 output tiles over 8 warps
    for (int tile = warp; tile < 18; tile += kLevThreads / 32) {
      const int mt = tile / 6, nn = tile - mt * 6;
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
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
      const int gi = mt * 16 + r, gj = nn * 8 + 2 * j;
      sm.g[gi][gj] += acc[0];
      sm.g[gi][gj + 1] += acc[1];
      sm.g[gi + 8][gj] += acc[2];
      sm.g[gi + 8][gj + 1] += acc[3];
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

// ---- kernel 2: symmetric eigendecomposition, one WARP per matrix — Householder tridiagonalisation, accumulation of the
// reflectors and implicit-shift QL (the EISPACK tred2 / tql2 pair, restated for a warp: lanes own
// matrix rows, the scalar recurrences are computed redundantly by every lane).  ~1900 dependent
// rotations of ~50 cycles each; a block-wide parallel cyclic Jacobi (the first version of this kernel)
// moved 27 MB through shared memory per matrix, this moves 2.7 MB.
// out: sv[(ch*HKV+h)][48][48] = round16(V |lam|^-1/2)   (compactor.py:197: S of a symmetric G is |lam|);
// the column (eigenpair) order is arbitrary, the score sums over all of them.
constexpr int kQlWarps = 16;  // matrices per CTA: few, fat CTAs block few SMs while attention runs beside them
struct QlSmem {
  float a[kSk][kSk + 1];
  float d[kSk], e[kSk], hh[kSk], u[kSk], q[kSk];
};

__device__ __forceinline__ float rsqrt_nr(float x) {  // MUFU.RSQ + one Newton step; 0 -> 0
  if (!(x > 0.f)) return 0.f;
  const float y = rsqrtf(x);
  return y * fmaf(-0.5f * x * y, y, 1.5f);
}

template <bool BF16>
__global__ void __launch_bounds__(kQlWarps * 32) leverage_eigql_kernel(const float* __restrict__ gram,
                                                                      float* __restrict__ sv_out,
                                                                      const int32_t* __restrict__ chunk_offsets,
                                                                      int HKV, int mats) {
  extern __shared__ __align__(16) unsigned char ql_smem_raw[];
  QlSmem* sm_all = reinterpret_cast<QlSmem*>(ql_smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int mat = blockIdx.x * kQlWarps + warp;
  if (mat >= mats) return;
  const int ch = mat / HKV;
  if (chunk_offsets[ch + 1] - chunk_offsets[ch] <= 0) return;
  QlSmem& s = sm_all[warp];
  constexpr int n = kSk;
  const float* gin = gram + (int64_t)mat * n * n;
  for (int i = lane; i < n * n; i += 32) s.a[i / n][i % n] = gin[i];
  __syncwarp();
  const int j0 = lane, j1 = lane + 32;  // the two rows (or columns) of this lane; j1 valid for lane < 16
  const bool has1 = j1 < n;

  // ---- Householder reduction to tridiagonal form; u^(i) is parked in column i (rows < i) ----
  for (int i = n - 1; i >= 2; --i) {
    float x0 = (j0 < i) ? s.a[i][j0] : 0.f;
    float x1 = (has1 && j1 < i) ? s.a[i][j1] : 0.f;
    const float scale = warp_sum(fabsf(x0) + fabsf(x1));
    if (scale == 0.f) {  // row already reduced: H_i = I
      if (lane == 0) { s.e[i] = 0.f; s.hh[i] = 0.f; s.d[i] = s.a[i][i]; }
      __syncwarp();
      continue;
    }
    const float inv = 1.f / scale;
    x0 *= inv; x1 *= inv;
    float h = warp_sum(x0 * x0 + x1 * x1);
    const float f = (i - 1 < 32) ? __shfl_sync(0xffffffffu, x0, i - 1) : __shfl_sync(0xffffffffu, x1, i - 33);
    const float g = -copysignf(sqrtf(h), f);
    h -= f * g;
    if (j0 == i - 1) x0 = f - g;
    if (j1 == i - 1) x1 = f - g;
    if (j0 < i) { s.u[j0] = x0; s.a[j0][i] = x0; }
    if (has1 && j1 < i) { s.u[j1] = x1; s.a[j1][i] = x1; }
    if (lane == 0) { s.e[i] = scale * g; s.hh[i] = h; s.d[i] = s.a[i][i]; }
    __syncwarp();
    // p = A u / h  (rows >= i compute garbage that is never used)
    float p0 = 0.f, p1 = 0.f;
    for (int k = 0; k < i; ++k) {
      const float uk = s.u[k];
      p0 = fmaf(s.a[j0][k], uk, p0);
      if (has1) p1 = fmaf(s.a[j1][k], uk, p1);
    }
    const float rh = 1.f / h;
    p0 *= rh; p1 *= rh;
    const float K = warp_sum(((j0 < i) ? x0 * p0 : 0.f) + ((has1 && j1 < i) ? x1 * p1 : 0.f)) * 0.5f * rh;
    const float q0 = p0 - K * x0, q1 = p1 - K * x1;
    if (j0 < i) s.q[j0] = q0;
    if (has1 && j1 < i) s.q[j1] = q1;
    __syncwarp();
    // A <- A - u q^T - q u^T on the leading i x i block (kept fully symmetric)
    for (int k = 0; k < i; ++k) {
      const float uk = s.u[k], qk = s.q[k];
      if (j0 < i) s.a[j0][k] -= fmaf(x0, qk, q0 * uk);
      if (has1 && j1 < i) s.a[j1][k] -= fmaf(x1, qk, q1 * uk);
    }
    __syncwarp();
  }
  if (lane == 0) {
    s.d[0] = s.a[0][0]; s.d[1] = s.a[1][1];
    s.e[1] = s.a[1][0]; s.e[0] = 0.f;
    s.a[0][0] = 1.f;
  }
  __syncwarp();

  // ---- Z = H_{n-1} ... H_2, built in place: lanes own columns ----
  for (int i = 2; i <= n; ++i) {
    // row / column i-1 of the growing block become the identity (column i-1 held u^(i-1), consumed)
    if (j0 < i - 1) { s.a[i - 1][j0] = 0.f; s.a[j0][i - 1] = 0.f; }
    if (has1 && j1 < i - 1) { s.a[i - 1][j1] = 0.f; s.a[j1][i - 1] = 0.f; }
    if (lane == 0) s.a[i - 1][i - 1] = 1.f;
    __syncwarp();
    if (i < n) {
      const float h = s.hh[i];
      if (h != 0.f) {
        float g0 = 0.f, g1 = 0.f;
        for (int k = 0; k < i; ++k) {
          const float uk = s.a[k][i];
          g0 = fmaf(uk, s.a[k][j0], g0);
          if (has1) g1 = fmaf(uk, s.a[k][j1], g1);
        }
        const float rh = 1.f / h;
        g0 *= rh; g1 *= rh;
        for (int k = 0; k < i; ++k) {
          const float uk = s.a[k][i];
          if (j0 < i) s.a[k][j0] = fmaf(-g0, uk, s.a[k][j0]);
          if (has1 && j1 < i) s.a[k][j1] = fmaf(-g1, uk, s.a[k][j1]);
        }
      }
      __syncwarp();
    }
  }

  // ---- implicit-shift QL on (d, e); rotations accumulated into Z (lanes own rows) ----
  {
    const float e0 = s.e[j0 + 1 < n ? j0 + 1 : 0];
    const float e1 = (has1 && j1 + 1 < n) ? s.e[j1 + 1] : 0.f;
    __syncwarp();
    s.e[j0] = e0;  // j0 + 1 < n always holds for j0 <= 31
    if (has1) s.e[j1] = e1;
    __syncwarp();
  }
  float fsh = 0.f, tst1 = 0.f;
  const float eps = 1.1920929e-7f;
  for (int l = 0; l < n; ++l) {
    tst1 = fmaxf(tst1, fabsf(s.d[l]) + fabsf(s.e[l]));
    int m = l;
    while (m < n && fabsf(s.e[m]) > eps * tst1) ++m;  // e[n-1] == 0 ends the scan
    if (m > l) {
      int iter = 0;
      float el_abs;
      do {
        ++iter;
        const float g0 = s.d[l], el = s.e[l], el1 = s.e[l + 1];
        float pp = (s.d[l + 1] - g0) / (2.f * el);
        float rr = sqrtf(fmaf(pp, pp, 1.f));
        if (pp < 0.f) rr = -rr;
        const float dl = el / (pp + rr), dl1 = el * (pp + rr);
        const float hs = g0 - dl;
        const float dj0 = s.d[j0], dj1 = has1 ? s.d[j1] : 0.f;
        float pcur = s.d[m];
        __syncwarp();
        if (j0 >= l + 2) s.d[j0] = dj0 - hs;
        if (has1 && j1 >= l + 2) s.d[j1] = dj1 - hs;
        if (lane == 0) { s.d[l] = dl; s.d[l + 1] = dl1; }
        __syncwarp();
        if (m >= l + 2 && m < n) pcur = s.d[m];  // d[m] was shifted by hs above
        else if (m == l + 1) pcur = dl1;
        fsh += hs;
        float c = 1.f, c2 = 1.f, c3 = 1.f, sn = 0.f, s2 = 0.f;
        // column m of Z is carried in registers from one rotation to the next
        float z0 = s.a[j0][m], z1 = has1 ? s.a[j1][m] : 0.f;
        float ei = s.e[m - 1], di = s.d[m - 1];
        for (int i = m - 1; i >= l; --i) {
          const float ein = (i > l) ? s.e[i - 1] : 0.f, din = (i > l) ? s.d[i - 1] : 0.f;  // prefetch
          c3 = c2; c2 = c; s2 = sn;
          const float g = c * ei, h = c * pcur;
          const float x2 = fmaf(pcur, pcur, ei * ei);
          const float ri = rsqrt_nr(x2);
          const float r = x2 * ri;
          sn = ei * ri;
          c = pcur * ri;
          pcur = c * di - sn * g;
          if (lane == 0) { s.e[i + 1] = s2 * r; s.d[i + 1] = h + sn * (c * g + sn * di); }
          // Z[:, i+1] = sn Z[:, i] + c Z[:, i+1];  Z[:, i] = c Z[:, i] - sn Z[:, i+1]
          const float y0 = s.a[j0][i];
          s.a[j0][i + 1] = fmaf(sn, y0, c * z0);
          z0 = fmaf(c, y0, -sn * z0);
          if (has1) {
            const float y1 = s.a[j1][i];
            s.a[j1][i + 1] = fmaf(sn, y1, c * z1);
            z1 = fmaf(c, y1, -sn * z1);
          }
          ei = ein; di = din;
        }
        s.a[j0][l] = z0;
        if (has1) s.a[j1][l] = z1;
        const float pe = -sn * s2 * c3 * el1 * s.e[l] / dl1;
        el_abs = fabsf(sn * pe);
        __syncwarp();
        if (lane == 0) { s.e[l] = sn * pe; s.d[l] = c * pe; }
        __syncwarp();
      } while (el_abs > eps * tst1 && iter < 60);
    }
    __syncwarp();
    if (lane == 0) { s.d[l] += fsh; s.e[l] = 0.f; }
    __syncwarp();
  }

  float* out = sv_out + (int64_t)mat * n * n;
  for (int i = lane; i < n * n; i += 32) {
    const int a = i / n, b = i - a * n;
    out[i] = round16<BF16>(s.a[a][b] * rsqrtf(fabsf(s.d[b])));
  }
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
  if (!p.sketch) load_phi(sm, p);
  {
    const float* svin = p.sv + (int64_t)(ch * p.HKV + h) * kSk * kSk;
    for (int i = tid; i < kSk * kSk; i += kLevThreads) (&sm.sv[0][0])[i] = svin[i];
    if (tid < kSk) sm.mean[tid] = p.mean[(int64_t)(ch * p.HKV + h) * kSk + tid];
  }
  __syncthreads();
  for (int t0 = 0; t0 < c_len; t0 += kTokTile) {
    const int nt = min(kTokTile, c_len - t0);
    sketch_tile<BF16>(sm, p, h, c_beg + t0, nt, true);
    __syncthreads();
    for (int t = tid; t < nt; t += kLevThreads) {
      float x[kSk];
#pragma unroll
      for (int c = 0; c < kSk; ++c) x[c] = from16<BF16>(sm.xt[c][t]);
      float score = 0.f;
#pragma unroll 4
      for (int jc = 0; jc < kSk; jc += 4) {
        float u0 = 0.f, u1 = 0.f, u2 = 0.f, u3 = 0.f;
#pragma unroll
        for (int c = 0; c < kSk; ++c) {
          const float4 s4 = *reinterpret_cast<const float4*>(&sm.sv[c][jc]);
          u0 = fmaf(x[c], s4.x, u0); u1 = fmaf(x[c], s4.y, u1);
          u2 = fmaf(x[c], s4.z, u2); u3 = fmaf(x[c], s4.w, u3);
        }
        u0 = round16<BF16>(u0); u1 = round16<BF16>(u1); u2 = round16<BF16>(u2); u3 = round16<BF16>(u3);
        score += round16<BF16>(u0 * u0) + round16<BF16>(u1 * u1) + round16<BF16>(u2 * u2) +
                 round16<BF16>(u3 * u3);
      }
      p.raw[(int64_t)(c_beg + t0 + t) * p.HKV + h] = fmaxf(round16<BF16>(score), 0.f);
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
         lev_align(mats * kSk * sizeof(float));
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
  const size_t smem = sizeof(LevSmem);
  static bool attr_set = false;
  if (!attr_set) {
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(leverage_gram_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(leverage_gram_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(leverage_score_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(leverage_score_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(leverage_eigql_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(QlSmem) * kQlWarps)));
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(leverage_eigql_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(QlSmem) * kQlWarps)));
    attr_set = true;
  }
  const dim3 grid(n_chunks, HKV);
  if (dtype == CVLLM_BF16) {
    leverage_gram_kernel<true><<<grid, kLevThreads, smem, st>>>(p);
    leverage_eigql_kernel<true><<<(unsigned)((mats + kQlWarps - 1) / kQlWarps), kQlWarps * 32, sizeof(QlSmem) * kQlWarps, st>>>(p.gram, p.sv, chunk_offsets, HKV, (int)mats);
    leverage_score_kernel<true><<<grid, kLevThreads, smem, st>>>(p);
    leverage_finalize_kernel<true><<<n_chunks, 256, 0, st>>>(p.raw, chunk_offsets, out_scores, HKV, normalize);
  } else {
    leverage_gram_kernel<false><<<grid, kLevThreads, smem, st>>>(p);
    leverage_eigql_kernel<false><<<(unsigned)((mats + kQlWarps - 1) / kQlWarps), kQlWarps * 32, sizeof(QlSmem) * kQlWarps, st>>>(p.gram, p.sv, chunk_offsets, HKV, (int)mats);
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
