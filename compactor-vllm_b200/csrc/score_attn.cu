// Attention-mass key scorers: Compactor's chunk-local non-causal scorer and
// SnapKV's observation-window scorer.
//
// Reference: compression/compactor.py:338-599 (K11 + z-score epilogue + blend
// + protection) and compression/snapkv.py:39-276,332-448 (K12 writes ALL
// logits to a [N, HKV, w*G] fp32 scratch, K13 re-reads it).  Both are
// "softmax column sums" over a small set of query rows, ~1 % of the causal
// attention's FLOPs; they run on the side stream next to the tcgen05 prefill
// kernel, so they are kept narrow: mma.sync m16n8k16 tiles, K tile staged once
// per CTA in shared memory, Q rows read straight from global in fragment
// order (mma.cuh), probabilities never leave registers, and no logits scratch.
#include "common.cuh"
#include "mma.cuh"

namespace cvllm {

constexpr int kScThreads = 256;   // 8 warps
constexpr int kKeyTile = 128;     // keys per CTA tile
constexpr int kKRowBytes = 320;   // 256 B of keys + 64 B pad: conflict-free LDS.128 of B fragments
constexpr int kRefBlockM = 64;    // the reference's only K11 config (compactor.py:338-347)

// stage a [<=128 keys x 128 d] tile of 16-bit keys into padded smem; rows >= n_keys are zero
__device__ __forceinline__ void stage_k_tile(unsigned char* s_k, const char* k_base, int64_t k_stride_bytes,
                                             int n_keys) {
  for (int idx = threadIdx.x; idx < kKeyTile * 16; idx += blockDim.x) {
    const int r = idx >> 4, c = idx & 15;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < n_keys) v = ldg_nc_v4(k_base + (int64_t)r * k_stride_bytes + c * 16);
    *reinterpret_cast<uint4*>(s_k + r * kKRowBytes + c * 16) = v;
  }
}

// S[16 rows x 128 keys] for one m-tile: acc[nt][4], nt = key n-tile (8 keys)
template <bool BF16>
__device__ __forceinline__ void qk_tile(float (&acc)[16][4], const AFrag128& a, const unsigned char* s_k,
                                        int r, int j) {
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) {
    acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
    const unsigned char* krow = s_k + (nt * 8 + r) * kKRowBytes;
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const uint4 kb = *reinterpret_cast<const uint4*>(krow + (j + 4 * cc) * 16);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int ks = 2 * cc + half;
        uint32_t a0, a1, a2, a3, b0, b1;
        a.regs(ks, a0, a1, a2, a3);
        bfrag_regs(kb, ks, b0, b1);
        mma16816<BF16>(acc[nt], a0, a1, a2, a3, b0, b1);
      }
    }
  }
}

// ===========================================================================
// Compactor chunk scorer
// ===========================================================================
struct ChunkParams {
  const void* q;
  const void* k;
  int64_t q_stride_n, k_stride_n;  // elements
  const int32_t* cu;
  float* raw;        // [N, HKV] column sums (pre z-score)
  float* partials;   // [B, max_chunks, HKV, 2] (sum, sumsq) of raw per CTA
  int B, HQ, HKV, G, chunk, max_chunks, n_ci;  // n_ci = chunks of the longest sequence
  float scale_log2;
};

// async copy of a [<=128 keys x 128 d] tile into padded smem (rows >= n_keys zero-filled)
__device__ __forceinline__ void stage_k_tile_async(unsigned char* s_k, const char* k_base,
                                                   int64_t k_stride_bytes, int n_keys) {
  for (int idx = threadIdx.x; idx < kKeyTile * 16; idx += blockDim.x) {
    const int r = idx >> 4, c = idx & 15;
    const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(s_k + r * kKRowBytes + c * 16));
    const bool ok = r < n_keys;
    const char* src = k_base + (int64_t)(ok ? r : 0) * k_stride_bytes + c * 16;
    const int nbytes = ok ? 16 : 0;  // src-size 0: the 16 destination bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(nbytes) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// Persistent: grid = min(items, #SM); item = (b, kv head, chunk).  The K tile of the NEXT item is
// in flight (cp.async, second smem buffer) and the Q fragments of the next m-tile are in
// registers while the current one is multiplied: the kernel is otherwise latency-bound (one
// 8-warp CTA per SM, 12 KB between consecutive query tokens).
template <bool BF16>
__global__ void __launch_bounds__(kScThreads, 1) chunk_scores_kernel(const ChunkParams p) {
  extern __shared__ __align__(16) unsigned char s_kbuf[];  // 2 x [128][kKRowBytes]
  __shared__ float s_col[kScThreads / 32][kKeyTile];
  __shared__ float s_red[2][kScThreads / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, r = lane >> 2, j = lane & 3;
  const int G = p.G;
  const int total = p.B * p.HKV * p.n_ci;
  const char* qb = reinterpret_cast<const char*>(p.q);

  // item -> (b, h, ci); returns false when the chunk lies beyond its sequence
  auto decode = [&](int idx, int& b, int& h, int& ci, int& cs, int& M) -> bool {
    ci = idx % p.n_ci;
    const int t = idx / p.n_ci;
    h = t % p.HKV;
    b = t / p.HKV;
    const int seq_beg = p.cu[b], seq_end = p.cu[b + 1];
    cs = seq_beg + ci * p.chunk;
    M = min(p.chunk, seq_end - cs);
    return cs < seq_end;
  };
  auto next_valid = [&](int idx) -> int {
    int b, h, ci, cs, M;
    while (idx < total && !decode(idx, b, h, ci, cs, M)) idx += gridDim.x;
    return idx;
  };
  auto prefetch = [&](int idx, int buf) {
    int b, h, ci, cs, M;
    decode(idx, b, h, ci, cs, M);
    stage_k_tile_async(s_kbuf + buf * (kKeyTile * kKRowBytes),
                       reinterpret_cast<const char*>(p.k) + ((int64_t)cs * p.k_stride_n + h * 128) * 2,
                       p.k_stride_n * 2, M);
  };

  int cur = next_valid(blockIdx.x);
  int buf = 0;
  if (cur < total) prefetch(cur, 0);
  while (cur < total) {
    const int nxt = next_valid(cur + gridDim.x);
    if (nxt < total) {
      prefetch(nxt, buf ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const unsigned char* s_k = s_kbuf + buf * (kKeyTile * kKRowBytes);
    int b, h, ci, cs, M;
    decode(cur, b, h, ci, cs, M);

    // rows = (token, g): row R -> token R / G, query head h*G + R % G
    const int n_rows = M * G;
    const int n_mt = (n_rows + 15) / 16;
    float colacc[16][2];
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) colacc[nt][0] = colacc[nt][1] = 0.f;
    auto qrow = [&](int R) -> const char* {
      if (R >= n_rows) return nullptr;
      const int tok = R / G, g = R - tok * G;
      return qb + ((int64_t)(cs + tok) * p.q_stride_n + (int64_t)(h * G + g) * 128) * 2;
    };
    AFrag128 a_next;
    if (warp < n_mt) a_next.load(qrow(warp * 16 + r), qrow(warp * 16 + r + 8), j);
    for (int mt = warp; mt < n_mt; mt += kScThreads / 32) {
      const int R0 = mt * 16 + r, R1 = R0 + 8;
      const AFrag128 a = a_next;
      const int mt2 = mt + kScThreads / 32;
      if (mt2 < n_mt) a_next.load(qrow(mt2 * 16 + r), qrow(mt2 * 16 + r + 8), j);
      float acc[16][4];
      qk_tile<BF16>(acc, a, s_k, r, j);

      // softmax over the chunk's keys, per row; lanes j=0..3 of a row share it
      float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        const int k0 = nt * 8 + 2 * j;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool ok = (k0 + e) < M;
          acc[nt][e] = ok ? acc[nt][e] * p.scale_log2 : -INFINITY;
          acc[nt][2 + e] = ok ? acc[nt][2 + e] * p.scale_log2 : -INFINITY;
          m0 = fmaxf(m0, acc[nt][e]);
          m1 = fmaxf(m1, acc[nt][2 + e]);
        }
      }
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          acc[nt][e] = ex2(acc[nt][e] - m0);
          acc[nt][2 + e] = ex2(acc[nt][2 + e] - m1);
          l0 += acc[nt][e];
          l1 += acc[nt][2 + e];
        }
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
      l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      const float i0 = (R0 < n_rows) ? 1.f / l0 : 0.f;
      const float i1 = (R1 < n_rows) ? 1.f / l1 : 0.f;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        colacc[nt][0] += acc[nt][0] * i0 + acc[nt][2] * i1;
        colacc[nt][1] += acc[nt][1] * i0 + acc[nt][3] * i1;
      }
    }
    // column sums: over the 8 row-lanes of the warp, then over warps in fixed order
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float v = colacc[nt][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (r == 0) s_col[warp][nt * 8 + 2 * j + e] = v;
      }
    }
    __syncthreads();
    // rows of the reference's last 64-row query tile that fall beyond the chunk add 1/chunk each
    // (compactor.py:411-414,477-479)
    const float pad = (float)(((M + kRefBlockM - 1) / kRefBlockM * kRefBlockM - M) * G) / (float)p.chunk;
    float v = 0.f;
    if (threadIdx.x < kKeyTile && (int)threadIdx.x < M) {
#pragma unroll
      for (int w = 0; w < kScThreads / 32; ++w) v += s_col[w][threadIdx.x];
      v += pad;
      p.raw[(int64_t)(cs + threadIdx.x) * p.HKV + h] = v;
    }
    // per-item (sum, sumsq) for the per-sequence z-score
    float s1 = ((int)threadIdx.x < M) ? v : 0.f, s2 = s1 * s1;
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) { s_red[0][warp] = s1; s_red[1][warp] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float a1 = 0.f, a2 = 0.f;
      for (int w = 0; w < kScThreads / 32; ++w) { a1 += s_red[0][w]; a2 += s_red[1][w]; }
      float* dst = p.partials + (((int64_t)b * p.max_chunks + ci) * p.HKV + h) * 2;
      dst[0] = a1;
      dst[1] = a2;
    }
    // the next iteration's prefetch overwrites buffer `buf` and its epilogue reuses s_col / s_red
    __syncthreads();
    cur = nxt;
    buf ^= 1;
  }
}

struct ChunkFinalizeParams {
  float* out;  // in: raw, out: final
  const float* partials;
  const int32_t* cu;
  const void* accum;
  int accum_dtype;  // 0 f16, 1 bf16, 2 f32
  int64_t accum_stride_n;
  float blend;
  const int32_t* prot_first;
  const int32_t* prot_last;
  int HKV, chunk, max_chunks, normalize;
};

// grid (tiles, B): every CTA re-reduces its sequence's partials in a fixed order (deterministic),
// then finishes a tile of tokens: z-score, + blend*accum, protected tokens <- +inf
__global__ void __launch_bounds__(256) chunk_finalize_kernel(const ChunkFinalizeParams p) {
  const int b = blockIdx.y;
  const int seq_beg = p.cu[b], seq_end = p.cu[b + 1], len = seq_end - seq_beg;
  if (len <= 0) return;
  __shared__ double s_sum[2][256];
  __shared__ float s_stat[2];
  float mean = 0.f, invstd = 1.f;
  if (p.normalize) {
    const int n_part = ((len + p.chunk - 1) / p.chunk) * p.HKV;
    const float* part = p.partials + (int64_t)b * p.max_chunks * p.HKV * 2;
    double a1 = 0.0, a2 = 0.0;
    for (int i = threadIdx.x; i < n_part; i += 256) {
      a1 += (double)part[2 * i];
      a2 += (double)part[2 * i + 1];
    }
    s_sum[0][threadIdx.x] = a1;
    s_sum[1][threadIdx.x] = a2;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if ((int)threadIdx.x < o) {
        s_sum[0][threadIdx.x] += s_sum[0][threadIdx.x + o];
        s_sum[1][threadIdx.x] += s_sum[1][threadIdx.x + o];
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      // compactor.py:258-260: var = max(E[x^2] - mean^2, 0), no eps
      const float cnt = (float)len * (float)p.HKV;
      const float m = (float)(s_sum[0][0] / (double)cnt);
      const float var = fmaxf((float)(s_sum[1][0] / (double)cnt) - m * m, 0.f);
      s_stat[0] = m;
      s_stat[1] = 1.0f / sqrtf(var);
    }
    __syncthreads();
    mean = s_stat[0];
    invstd = s_stat[1];
  }
  const int pf = p.prot_first ? min(max(p.prot_first[b], 0), len) : 0;
  const int pl = p.prot_last ? min(max(p.prot_last[b], 0), len) : 0;
  const int total = len * p.HKV;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const int tok = i / p.HKV, h = i - tok * p.HKV;
    const int64_t n = seq_beg + tok;
    float v = p.out[n * p.HKV + h];
    if (p.normalize) v = (v - mean) * invstd;
    if (p.accum) {
      float a;
      if (p.accum_dtype == 2) {
        a = reinterpret_cast<const float*>(p.accum)[n * p.accum_stride_n + h];
      } else if (p.accum_dtype == 1) {
        a = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.accum)[n * p.accum_stride_n + h]);
      } else {
        a = __half2float(reinterpret_cast<const __half*>(p.accum)[n * p.accum_stride_n + h]);
      }
      v += a * p.blend;
    }
    if (tok < pf || tok >= len - pl) v = INFINITY;
    p.out[n * p.HKV + h] = v;
  }
}

// ===========================================================================
// SnapKV
// ===========================================================================
struct SnapParams {
  const void* q;
  const void* k;
  int64_t q_stride_n, k_stride_n;
  const int32_t* cu_q;
  const int32_t* cu_k;
  float* stats;  // [B, HKV, n_splits, rows_pad, 2] partial (m, l) per row per key split
  float* out;    // [Nk, HKV]
  int HQ, HKV, G, w, rows_pad, n_splits, keys_per_split, kpool, pool_segment;
  float scale_log2;
};

// rows = (q_off, g): row R -> query token q_end - w + R / G, head hk*G + R % G (snapkv.py:106-113)
__device__ __forceinline__ const char* snap_qrow(const SnapParams& p, int R, int q_end, int hk) {
  if (R >= p.w * p.G) return nullptr;
  const int qo = R / p.G, g = R - qo * p.G;
  return reinterpret_cast<const char*>(p.q) +
         ((int64_t)(q_end - p.w + qo) * p.q_stride_n + (int64_t)(hk * p.G + g) * 128) * 2;
}

// pass 1, grid (n_splits, HKV, B): running (max, sum) of every row over this split's keys
template <bool BF16>
__global__ void __launch_bounds__(kScThreads, 1) snap_stats_kernel(const SnapParams p) {
  extern __shared__ __align__(16) unsigned char s_k[];
  const int sp = blockIdx.x, hk = blockIdx.y, b = blockIdx.z;
  const int kb = p.cu_k[b], keff = p.cu_k[b + 1] - p.w, q_end = p.cu_q[b + 1];
  const int n_keys_seq = keff - kb;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, r = lane >> 2, j = lane & 3;
  const int n_mt = p.rows_pad / 16;
  float* st = p.stats + ((((int64_t)b * p.HKV + hk) * p.n_splits + sp) * p.rows_pad) * 2;
  const int k_lo = sp * p.keys_per_split;
  const int k_hi = min(n_keys_seq, k_lo + p.keys_per_split);
  if (n_keys_seq <= 0 || k_lo >= k_hi) {
    for (int R = threadIdx.x; R < p.rows_pad; R += kScThreads) { st[2 * R] = -INFINITY; st[2 * R + 1] = 0.f; }
    return;
  }
  // each warp keeps the stats of its m-tiles in registers across key tiles (<= 2 m-tiles per warp)
  float m_run[2][2], l_run[2][2];
#pragma unroll
  for (int t = 0; t < 2; ++t) m_run[t][0] = m_run[t][1] = -INFINITY, l_run[t][0] = l_run[t][1] = 0.f;

  for (int kt = k_lo; kt < k_hi; kt += kKeyTile) {
    const int nk = min(kKeyTile, k_hi - kt);
    __syncthreads();
    stage_k_tile(s_k, reinterpret_cast<const char*>(p.k) + ((int64_t)(kb + kt) * p.k_stride_n + hk * 128) * 2,
                 p.k_stride_n * 2, nk);
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const int mt = warp + t * (kScThreads / 32);
      if (mt >= n_mt) continue;
      AFrag128 a;
      a.load(snap_qrow(p, mt * 16 + r, q_end, hk), snap_qrow(p, mt * 16 + r + 8, q_end, hk), j);
      float acc[16][4];
      qk_tile<BF16>(acc, a, s_k, r, j);
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool ok = (nt * 8 + 2 * j + e) < nk;
          acc[nt][e] = ok ? acc[nt][e] * p.scale_log2 : -INFINITY;
          acc[nt][2 + e] = ok ? acc[nt][2 + e] * p.scale_log2 : -INFINITY;
          mx0 = fmaxf(mx0, acc[nt][e]);
          mx1 = fmaxf(mx1, acc[nt][2 + e]);
        }
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float n0 = fmaxf(m_run[t][0], mx0), n1 = fmaxf(m_run[t][1], mx1);
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        s0 += ex2(acc[nt][0] - n0) + ex2(acc[nt][1] - n0);
        s1 += ex2(acc[nt][2] - n1) + ex2(acc[nt][3] - n1);
      }
      s0 += __shfl_xor_sync(0xffffffffu, s0, 1);
      s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
      s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
      s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
      l_run[t][0] = l_run[t][0] * ex2(m_run[t][0] - n0) + s0;
      l_run[t][1] = l_run[t][1] * ex2(m_run[t][1] - n1) + s1;
      m_run[t][0] = n0;
      m_run[t][1] = n1;
    }
  }
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int mt = warp + t * (kScThreads / 32);
    if (mt >= n_mt || j != 0) continue;
    const int R0 = mt * 16 + r, R1 = R0 + 8;
    st[2 * R0] = m_run[t][0]; st[2 * R0 + 1] = l_run[t][0];
    st[2 * R1] = m_run[t][1]; st[2 * R1 + 1] = l_run[t][1];
  }
}

// pass 2, grid (key tiles, HKV, B): column sums of exp2(s - m_row) / l_row, pooled, written out
template <bool BF16>
__global__ void __launch_bounds__(kScThreads, 1) snap_scores_kernel(const SnapParams p) {
  extern __shared__ __align__(16) unsigned char s_k[];
  __shared__ float s_col[kScThreads / 32][kKeyTile];
  __shared__ float s_raw[kKeyTile];
  __shared__ float s_m[256], s_il[256];
  const int tile = blockIdx.x, hk = blockIdx.y, b = blockIdx.z;
  const int kb = p.cu_k[b], ke = p.cu_k[b + 1], keff = ke - p.w, q_end = p.cu_q[b + 1];
  const int len = ke - kb;
  const int k0 = tile * kKeyTile;  // relative to the sequence start
  if (k0 >= len) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, r = lane >> 2, j = lane & 3;
  const int n_keys_seq = keff - kb;
  if (p.w <= 0 || n_keys_seq <= 0) {
    // L <= w: the reference leaves these rows uninitialised (snapkv.py:94-95,204-205); zeros here
    if ((int)threadIdx.x < min(kKeyTile, len - k0)) p.out[(int64_t)(kb + k0 + threadIdx.x) * p.HKV + hk] = 0.f;
    return;
  }
  const int nk = max(0, min(kKeyTile, n_keys_seq - k0));  // scored keys in this tile
  if (nk > 0) {
    // combine the per-split row stats
    for (int R = threadIdx.x; R < p.rows_pad; R += kScThreads) {
      const float* st = p.stats + (((int64_t)b * p.HKV + hk) * p.n_splits * p.rows_pad + R) * 2;
      float M = -INFINITY;
      for (int s = 0; s < p.n_splits; ++s) M = fmaxf(M, st[(int64_t)s * p.rows_pad * 2]);
      float L = 0.f;
      for (int s = 0; s < p.n_splits; ++s) {
        const float ms = st[(int64_t)s * p.rows_pad * 2];
        if (ms > -INFINITY) L += st[(int64_t)s * p.rows_pad * 2 + 1] * ex2(ms - M);
      }
      s_m[R] = M;
      s_il[R] = (R < p.w * p.G && L > 0.f) ? 1.f / L : 0.f;
    }
    stage_k_tile(s_k, reinterpret_cast<const char*>(p.k) + ((int64_t)(kb + k0) * p.k_stride_n + hk * 128) * 2,
                 p.k_stride_n * 2, nk);
    __syncthreads();
    float colacc[16][2];
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) colacc[nt][0] = colacc[nt][1] = 0.f;
    const int n_mt = p.rows_pad / 16;
    for (int mt = warp; mt < n_mt; mt += kScThreads / 32) {
      const int R0 = mt * 16 + r, R1 = R0 + 8;
      AFrag128 a;
      a.load(snap_qrow(p, R0, q_end, hk), snap_qrow(p, R1, q_end, hk), j);
      float acc[16][4];
      qk_tile<BF16>(acc, a, s_k, r, j);
      const float m0 = s_m[R0], m1 = s_m[R1], i0 = s_il[R0], i1 = s_il[R1];
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        colacc[nt][0] += ex2(acc[nt][0] * p.scale_log2 - m0) * i0 + ex2(acc[nt][2] * p.scale_log2 - m1) * i1;
        colacc[nt][1] += ex2(acc[nt][1] * p.scale_log2 - m0) * i0 + ex2(acc[nt][3] * p.scale_log2 - m1) * i1;
      }
    }
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float v = colacc[nt][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (r == 0) s_col[warp][nt * 8 + 2 * j + e] = v;
      }
    }
    __syncthreads();
    if (threadIdx.x < kKeyTile) {
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < kScThreads / 32; ++w) v += s_col[w][threadIdx.x];
      s_raw[threadIdx.x] = v;
    }
    __syncthreads();
  }
  // causal kpool-tap mean inside pool_segment-key segments (snapkv.py:253-262); window <- +inf
  if (threadIdx.x < kKeyTile) {
    const int i = threadIdx.x;
    const int kabs = k0 + i;
    if (kabs < len) {
      float v;
      if (i < nk) {
        const int lo = max(i - (p.kpool - 1), (i / p.pool_segment) * p.pool_segment);
        float s = 0.f;
        for (int t = lo; t <= i; ++t) s += s_raw[t];
        v = s / (float)(i + 1 - lo);
      } else {
        v = INFINITY;
      }
      p.out[(int64_t)(kb + kabs) * p.HKV + hk] = v;
    }
  }
}

}  // namespace cvllm

using namespace cvllm;

int cvllm_launch_chunk_scores_umma(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                                   const int32_t* cu, int B, int N, int HQ, int HKV, int chunk, int max_chunks,
                                   int n_ci, float scale_log2, float* raw, float* partials, bool bf16,
                                   cudaStream_t st);

size_t cvllm_snap_umma_workspace_bytes(int B, int max_seqlen_k, int HKV);
int cvllm_launch_snap_umma(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                           const int32_t* cu_q, const int32_t* cu_k, int w, int B, int Nq, int Nk, int HQ, int HKV,
                           int max_seqlen_k, float scale_log2, int kpool, int pool_segment, float* out, float* stats,
                           bool bf16, cudaStream_t st);

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static const int kScSmem = kKeyTile * kKRowBytes;  // 40 KB

extern "C" size_t cvllm_chunk_attn_scores_workspace_bytes(int B, int N, int HKV, int chunk) {
  if (B <= 0 || N <= 0 || HKV <= 0 || chunk <= 0) return 256;
  const size_t max_chunks = (size_t)(N + chunk - 1) / chunk + 1;
  return al256((size_t)B * max_chunks * HKV * 2 * sizeof(float));
}

static int chunk_scores_impl(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                             const int32_t* cu_seqlens, int B, int N, int HQ, int HKV, int D,
                             int chunk, int max_seqlen, float sm_scale, int normalize,
                             const void* accum, int accum_dtype, int64_t accum_stride_n,
                             float blend, const int32_t* prot_first, const int32_t* prot_last,
                             float* out, int dtype, void* workspace, size_t workspace_bytes,
                             cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && N >= 0 && HQ > 0 && HKV > 0 && chunk > 0 && max_seqlen >= 0);
  if (N == 0 || max_seqlen == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(q && k && cu_seqlens && out && workspace);
  CVLLM_CHECK_ARG(HQ % HKV == 0 && (dtype == CVLLM_F16 || dtype == CVLLM_BF16));
  CVLLM_CHECK_ARG(accum_dtype >= 0 && accum_dtype <= 2);
  CVLLM_CHECK_SUPPORTED(D == 128 && chunk <= kKeyTile);
  CVLLM_CHECK_ALIGN(aligned16(q) && aligned16(k) && q_stride_n % 8 == 0 && k_stride_n % 8 == 0);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_chunk_attn_scores_workspace_bytes(B, N, HKV, chunk));
  const int max_chunks = (N + chunk - 1) / chunk + 1;
  const int grid_chunks = (max_seqlen + chunk - 1) / chunk;
  CVLLM_CHECK_ARG(grid_chunks <= max_chunks);
  auto st = static_cast<cudaStream_t>(stream);
  ChunkParams p;
  p.q = q; p.k = k; p.q_stride_n = q_stride_n; p.k_stride_n = k_stride_n; p.cu = cu_seqlens;
  p.raw = out; p.partials = reinterpret_cast<float*>(workspace);
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = HQ / HKV; p.chunk = chunk; p.max_chunks = max_chunks;
  p.n_ci = grid_chunks;
  p.scale_log2 = sm_scale * 1.4426950408889634f;
  const int64_t items = (int64_t)B * HKV * grid_chunks;
  CVLLM_CHECK_SUPPORTED(items < (1ll << 31));
  if (test_options().chunk_scores_v1.load() == 0) {
    // product path: tcgen05 tiles fed by TMA (score_chunk_umma.cu)
    const int rc = cvllm_launch_chunk_scores_umma(q, k, q_stride_n, k_stride_n, cu_seqlens, B, N, HQ, HKV, chunk,
                                                  max_chunks, grid_chunks, p.scale_log2, out, p.partials,
                                                  dtype == CVLLM_BF16, st);
    if (rc != CVLLM_OK) return rc;
  } else {
    const int grid = (int)(items < sm_count() ? items : sm_count());
    static std::atomic<uint64_t> attr_bf{0}, attr_h{0};  // 2 K buffers = 80 KB dynamic smem
    if (dtype == CVLLM_BF16) CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)chunk_scores_kernel<true>, 2 * kScSmem, attr_bf));
    else CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)chunk_scores_kernel<false>, 2 * kScSmem, attr_h));
    if (dtype == CVLLM_BF16) chunk_scores_kernel<true><<<grid, kScThreads, 2 * kScSmem, st>>>(p);
    else chunk_scores_kernel<false><<<grid, kScThreads, 2 * kScSmem, st>>>(p);
    CVLLM_CHECK_LAUNCH();
  }
  ChunkFinalizeParams f;
  f.out = out; f.partials = p.partials; f.cu = cu_seqlens; f.accum = accum; f.accum_dtype = accum_dtype;
  f.accum_stride_n = accum_stride_n; f.blend = blend; f.prot_first = prot_first; f.prot_last = prot_last;
  f.HKV = HKV; f.chunk = chunk; f.max_chunks = max_chunks; f.normalize = normalize;
  const int tiles = max(1, min((max_seqlen * HKV + 255) / 256, 2 * sm_count()));
  chunk_finalize_kernel<<<dim3(tiles, B), 256, 0, st>>>(f);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

extern "C" int cvllm_chunk_attn_scores(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                                       const int32_t* cu_seqlens, int B, int N, int HQ, int HKV, int D,
                                       int chunk, int max_seqlen, float sm_scale, int normalize,
                                       const void* accum, int accum_dtype, int64_t accum_stride_n,
                                       float blend, const int32_t* prot_first, const int32_t* prot_last,
                                       float* out, int dtype, void* workspace, size_t workspace_bytes,
                                       cvllm_stream_t stream) {
  return chunk_scores_impl(q, k, q_stride_n, k_stride_n, cu_seqlens, B, N, HQ, HKV, D, chunk, max_seqlen, sm_scale,
                           normalize, accum, accum_dtype, accum_stride_n, blend, prot_first, prot_last, out, dtype,
                           workspace, workspace_bytes, stream);
}

static void snap_geometry(int max_seqlen_k, int HQ, int HKV, int w, int* rows_pad, int* n_splits,
                          int* keys_per_split) {
  const int G = HQ / HKV;
  *rows_pad = ((w * G + 15) / 16) * 16;
  // ~2048 keys per pass-1 CTA keeps the partial-stat traffic small and still fills the chip at 40K
  int kps = 2048;
  int ns = (max_seqlen_k + kps - 1) / kps;
  if (ns < 1) ns = 1;
  *n_splits = ns;
  *keys_per_split = kps;
}

extern "C" size_t cvllm_snapkv_scores_workspace_bytes(int B, int N, int HQ, int HKV, int w) {
  if (B <= 0 || N <= 0 || HKV <= 0 || HQ <= 0) return 256;
  int rows_pad, n_splits, kps;
  snap_geometry(N, HQ, HKV, w, &rows_pad, &n_splits, &kps);
  const size_t v1 = (size_t)B * HKV * n_splits * rows_pad * 2 * sizeof(float);
  const size_t v2 = cvllm_snap_umma_workspace_bytes(B, N, HKV);  // tcgen05 route (w * G <= 128)
  return al256(v1 > v2 ? v1 : v2);
}

extern "C" int cvllm_snapkv_scores(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                                   const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k, int w, int B,
                                   int Nk, int HQ, int HKV, int D, int max_seqlen_k, float sm_scale,
                                   int kpool, int pool_segment, float* out, int dtype, void* workspace,
                                   size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && Nk >= 0 && HQ > 0 && HKV > 0 && w >= 0 && kpool >= 1 && max_seqlen_k >= 0);
  if (Nk == 0 || max_seqlen_k == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(q && k && cu_seqlens_q && cu_seqlens_k && out && workspace);
  CVLLM_CHECK_ARG(HQ % HKV == 0 && (dtype == CVLLM_F16 || dtype == CVLLM_BF16));
  CVLLM_CHECK_SUPPORTED(D == 128);
  CVLLM_CHECK_SUPPORTED(pool_segment > 0 && kKeyTile % pool_segment == 0);
  CVLLM_CHECK_SUPPORTED(w * (HQ / HKV) <= 256);
  CVLLM_CHECK_ALIGN(aligned16(q) && aligned16(k) && q_stride_n % 8 == 0 && k_stride_n % 8 == 0);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_snapkv_scores_workspace_bytes(B, max_seqlen_k, HQ, HKV, w));
  if (w > 0 && w * (HQ / HKV) <= 128 && test_options().snap_v1.load() == 0) {
    // product path for the engine's window (w = 32, G <= 4): tcgen05 tiles fed by TMA (score_snap_umma.cu).
    // Nq: the q tensor has at least cu_seqlens_q[B] = Nk rows on the prefill path (queries and keys are the same tokens)
    return cvllm_launch_snap_umma(q, k, q_stride_n, k_stride_n, cu_seqlens_q, cu_seqlens_k, w, B, Nk, Nk, HQ, HKV,
                                  max_seqlen_k, sm_scale * 1.4426950408889634f, kpool, pool_segment, out,
                                  reinterpret_cast<float*>(workspace), dtype == CVLLM_BF16,
                                  static_cast<cudaStream_t>(stream));
  }
  SnapParams p;
  p.q = q; p.k = k; p.q_stride_n = q_stride_n; p.k_stride_n = k_stride_n;
  p.cu_q = cu_seqlens_q; p.cu_k = cu_seqlens_k; p.stats = reinterpret_cast<float*>(workspace); p.out = out;
  p.HQ = HQ; p.HKV = HKV; p.G = HQ / HKV; p.w = w; p.kpool = kpool; p.pool_segment = pool_segment;
  snap_geometry(max_seqlen_k, HQ, HKV, w, &p.rows_pad, &p.n_splits, &p.keys_per_split);
  p.scale_log2 = sm_scale * 1.4426950408889634f;
  auto st = static_cast<cudaStream_t>(stream);
  const dim3 g1(p.n_splits, HKV, B), g2((max_seqlen_k + kKeyTile - 1) / kKeyTile, HKV, B);
  if (dtype == CVLLM_BF16) {
    snap_stats_kernel<true><<<g1, kScThreads, kScSmem, st>>>(p);
    snap_scores_kernel<true><<<g2, kScThreads, kScSmem, st>>>(p);
  } else {
    snap_stats_kernel<false><<<g1, kScThreads, kScSmem, st>>>(p);
    snap_scores_kernel<false><<<g2, kScThreads, kScSmem, st>>>(p);
  }
  CVLLM_CHECK_LAUNCHES(2);
  return CVLLM_OK;
}
