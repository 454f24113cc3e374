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
//   G   = V diag(lam) V^T                   parallel cyclic Jacobi, fp32, in smem
//   u   = round16(Xc . round16(V |lam|^-1/2));  score = round16(sum_j round16(u_j^2))
//
// The reference keeps X, Xc, G, V S^-1/2, U and the row sums in K's 16-bit
// dtype; rounding G entrywise perturbs it by far more than the
// 5e-3 regulariser (it can even go indefinite: the SVD then yields |lam|), so
// those rounding points are part of the reference's scores and are reproduced
// here — an exact fp32 x^T G^-1 x only overlaps ~80 % with the reference's
// kept set (tests/test_oracle_golden.py).  Work is ~170 KFLOP/token: latency
// and launch count, not throughput, is what the fusion removes.
#include "common.cuh"
#include "mma.cuh"

namespace cvllm {

constexpr int kLevThreads = 256;
constexpr int kSk = 48;            // sketch columns
constexpr int kTokTile = 256;      // tokens per smem tile
constexpr int kXtStride = kTokTile + 8;  // 16-bit elements per sketch row (pad: bank spread)
constexpr int kJacobiSweeps = 8;     // upper bound (8 == 10 in score accuracy, tools/leverage_sweeps.py); sweeps stop once no rotation moves G any more
constexpr int kEigThreads = 128;

struct LevParams {
  const void* k;
  int64_t k_stride_n, k_stride_h;  // elements
  const int32_t* chunk_offsets;
  const void* phi;  // [128, 48] 16-bit
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

  load_phi(sm, p);
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

// ---- kernel 2: grid (n_chunks * HKV): symmetric eigendecomposition, parallel cyclic Jacobi ----
// 24 disjoint (p,q) pairs per round, 47 rounds per sweep (round-robin tournament).  A round is
//   phase 1  24 threads: rotation (c,s) of each pair from the current G
//   phase 2  all threads: G <- J^T G J as 24x24 independent 2x2 blocks, V <- V J
// in place (every element belongs to exactly one block), two barriers per round.
// out: sv[(ch*HKV+h)][48][48] = round16(V |lam|^-1/2)   (compactor.py:197: S of a symmetric G is |lam|)
struct EigSmem {
  float g[kSk][kSk + 1];
  float v[kSk][kSk + 1];
  float4 rot[kSk / 2];  // (c, s, p, q) of each pair; p, q as int bit patterns
  int any_rot;
};

__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sqrt_approx(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <bool BF16>
__global__ void __launch_bounds__(kEigThreads, 4) leverage_eig_kernel(const float* __restrict__ gram,
                                                                   float* __restrict__ sv_out,
                                                                   const int32_t* __restrict__ chunk_offsets,
                                                                   int HKV) {
  __shared__ EigSmem sm;
  const int ch = blockIdx.x / HKV;
  if (chunk_offsets[ch + 1] - chunk_offsets[ch] <= 0) return;
  const int tid = threadIdx.x;
  const float* gin = gram + (int64_t)blockIdx.x * kSk * kSk;
  for (int i = tid; i < kSk * kSk; i += kEigThreads) {
    const int a = i / kSk, b = i - a * kSk;
    sm.g[a][b] = gin[i];
    sm.v[a][b] = (a == b) ? 1.f : 0.f;
  }
  if (tid == 0) sm.any_rot = 0;
  __syncthreads();

  constexpr int NP = kSk / 2;                                      // 24 pairs
  constexpr int GB = (NP * NP + kEigThreads - 1) / kEigThreads;    // G blocks per thread (5)
  constexpr int VB = (NP * kSk + kEigThreads - 1) / kEigThreads;   // V items per thread (9)
  for (int sweep = 0; sweep < kJacobiSweeps; ++sweep) {
    for (int rnd = 0; rnd < kSk - 1; ++rnd) {
      if (tid < NP) {
        // round-robin tournament: position 0 fixed, the others rotate
        const int n = kSk;
        int a = (tid == 0) ? 0 : 1 + (tid - 1 + rnd) % (n - 1);
        int b = 1 + (n - 2 - tid + rnd) % (n - 1);
        if (tid == 0) b = 1 + (n - 2 + rnd) % (n - 1);
        const int pp = min(a, b), qq = max(a, b);
        const float app = sm.g[pp][pp], aqq = sm.g[qq][qq], apq = sm.g[pp][qq];
        float c = 1.f, s = 0.f;
        if (fabsf(apq) > 1e-30f) {
          // the rotation only has to be orthogonal (c^2 + s^2 = 1 to an ulp: c = rsqrt(1 + t^2),
          // s = t c), not the exact annihilator: approximate rcp / sqrt keep the chain short
          const float tau = (aqq - app) * 0.5f * rcp_approx(apq);
          const float t = copysignf(rcp_approx(fabsf(tau) + sqrt_approx(fmaf(tau, tau, 1.f))), tau);
          c = rsqrtf(fmaf(t, t, 1.f));
          s = t * c;
          // still moving? (below this the rotation is the identity in fp32)
          if (apq * apq > 1e-14f * fabsf(app * aqq)) sm.any_rot = 1;
        }
        sm.rot[tid] = make_float4(c, s, __int_as_float(pp), __int_as_float(qq));
      }
      __syncthreads();
      {
        // The kernel is bound by shared-memory wavefronts, so the item -> lane maps are chosen for
        // the banks: G block (i, j) with j fastest (g[p_i][p_j] over distinct p_j: <= 2-way), V item
        // (i, row) with row fastest (v[row][p_i], stride 49: conflict-free).  Every load is issued
        // before the first in-place store (the stores would otherwise serialise the iterations).
        float4 gi[GB], gj[GB];
        float gpp[GB], gpq[GB], gqp[GB], gqq[GB];
#pragma unroll
        for (int u = 0; u < GB; ++u) {
          const int it = tid + u * kEigThreads;
          if (it < NP * NP) {
            const int i = it / NP, j = it - i * NP;
            gi[u] = sm.rot[i];
            gj[u] = sm.rot[j];
            const int pi = __float_as_int(gi[u].z), qi = __float_as_int(gi[u].w);
            const int pj = __float_as_int(gj[u].z), qj = __float_as_int(gj[u].w);
            gpp[u] = sm.g[pi][pj]; gpq[u] = sm.g[pi][qj];
            gqp[u] = sm.g[qi][pj]; gqq[u] = sm.g[qi][qj];
          }
        }
        float4 vi[VB];
        float vp[VB], vq[VB];
#pragma unroll
        for (int u = 0; u < VB; ++u) {
          const int it = tid + u * kEigThreads;
          if (it < NP * kSk) {
            const int i = it / kSk, row = it - i * kSk;
            vi[u] = sm.rot[i];
            vp[u] = sm.v[row][__float_as_int(vi[u].z)];
            vq[u] = sm.v[row][__float_as_int(vi[u].w)];
          }
        }
        // G blocks: rows {p_i, q_i} x cols {p_j, q_j}:  J_i^T . G . J_j
#pragma unroll
        for (int u = 0; u < GB; ++u) {
          const int it = tid + u * kEigThreads;
          if (it < NP * NP) {
            const int pi = __float_as_int(gi[u].z), qi = __float_as_int(gi[u].w);
            const int pj = __float_as_int(gj[u].z), qj = __float_as_int(gj[u].w);
            const float ci = gi[u].x, si = gi[u].y, cj = gj[u].x, sj = gj[u].y;
            const float rpp = ci * gpp[u] - si * gqp[u], rpq = ci * gpq[u] - si * gqq[u];
            const float rqp = si * gpp[u] + ci * gqp[u], rqq = si * gpq[u] + ci * gqq[u];
            sm.g[pi][pj] = cj * rpp - sj * rpq;
            sm.g[pi][qj] = sj * rpp + cj * rpq;
            sm.g[qi][pj] = cj * rqp - sj * rqq;
            sm.g[qi][qj] = sj * rqp + cj * rqq;
          }
        }
        // V <- V J_i
#pragma unroll
        for (int u = 0; u < VB; ++u) {
          const int it = tid + u * kEigThreads;
          if (it < NP * kSk) {
            const int i = it / kSk, row = it - i * kSk;
            const float c = vi[u].x, s = vi[u].y;
            sm.v[row][__float_as_int(vi[u].z)] = c * vp[u] - s * vq[u];
            sm.v[row][__float_as_int(vi[u].w)] = s * vp[u] + c * vq[u];
          }
        }
      }
      __syncthreads();
    }
    const bool done = sm.any_rot == 0;
    __syncthreads();
    if (done) break;
    if (tid == 0) sm.any_rot = 0;
    __syncwarp();
  }
  float* out = sv_out + (int64_t)blockIdx.x * kSk * kSk;
  for (int i = tid; i < kSk * kSk; i += kEigThreads) {
    const int a = i / kSk, b = i - a * kSk;
    out[i] = round16<BF16>(sm.v[a][b] * rsqrtf(fabsf(sm.g[b][b])));
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
  load_phi(sm, p);
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

extern "C" int cvllm_leverage_scores(const void* k_pre, int64_t k_stride_n, int64_t k_stride_h,
                                     const int32_t* chunk_offsets, int n_chunks, int max_chunk,
                                     const void* phi, void* out_scores, int N, int HKV, int D, int sketch,
                                     float regularizer, int normalize, int dtype, void* workspace,
                                     size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(N >= 0 && HKV > 0 && n_chunks >= 0 && max_chunk >= 0);
  if (N == 0 || n_chunks == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(k_pre && chunk_offsets && phi && out_scores && workspace);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_SUPPORTED(D == 128 && sketch == kSk);
  CVLLM_CHECK_ALIGN(aligned16(k_pre) && k_stride_n % 8 == 0 && k_stride_h % 8 == 0);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_leverage_scores_workspace_bytes(N, HKV, sketch, n_chunks));
  auto st = static_cast<cudaStream_t>(stream);
  const size_t mats = (size_t)n_chunks * HKV;
  char* wsb = reinterpret_cast<char*>(workspace);
  LevParams p;
  p.k = k_pre; p.k_stride_n = k_stride_n; p.k_stride_h = k_stride_h; p.chunk_offsets = chunk_offsets;
  p.phi = phi; p.HKV = HKV; p.reg = regularizer;
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
    attr_set = true;
  }
  const dim3 grid(n_chunks, HKV);
  if (dtype == CVLLM_BF16) {
    leverage_gram_kernel<true><<<grid, kLevThreads, smem, st>>>(p);
    leverage_eig_kernel<true><<<(unsigned)mats, kEigThreads, 0, st>>>(p.gram, p.sv, chunk_offsets, HKV);
    leverage_score_kernel<true><<<grid, kLevThreads, smem, st>>>(p);
    leverage_finalize_kernel<true><<<n_chunks, 256, 0, st>>>(p.raw, chunk_offsets, out_scores, HKV, normalize);
  } else {
    leverage_gram_kernel<false><<<grid, kLevThreads, smem, st>>>(p);
    leverage_eig_kernel<false><<<(unsigned)mats, kEigThreads, 0, st>>>(p.gram, p.sv, chunk_offsets, HKV);
    leverage_score_kernel<false><<<grid, kLevThreads, smem, st>>>(p);
    leverage_finalize_kernel<false><<<n_chunks, 256, 0, st>>>(p.raw, chunk_offsets, out_scores, HKV, normalize);
  }
  CVLLM_CHECK_LAUNCHES(4);
  return CVLLM_OK;
}
