// Head-sparse split-KV decode attention over the paged cache.
//
// Replaces attention/sparse_decode_kernel.py:10-165 (stage 1 :247-388, stage 2
// :392-435) with ONE launch:
//   * work = 128-token units of every (b, kv-head); the flattened unit list
//     is cut into gridDim.x equal ranges (stream-K style), so ragged per-head
//     lengths after compression are balanced exactly — the reference gives
//     every head the same KEY_SPLIT regardless of its length (:303-306)
//   * a warp owns 16 keys of a unit; K and V rows go global -> registers as
//     16-byte vectors laid out directly in mma.sync m16n8k16 fragment order
//     (the contraction dim is permuted identically on both operands, so no
//     shared-memory staging / ldmatrix is needed and every 32-byte sector
//     fetched is used)
//   * S^T = K.Q^T (keys on M, the G<=8 query heads on N), softmax stats per
//     column with warp shuffles, P^T goes through movmatrix to the B-operand
//     layout, O^T += V^T.P^T
//   * partial (m, l, O) of a range that does not cover its whole head goes to
//     a scratch slot; the CTA that completes a head merges (atomic ticket)
// HBM-bound: 4*D bytes per cached (token, head) — see DESIGN.md.
#include "common.cuh"
#include "mma.cuh"

namespace cvllm {

constexpr int kUnit = 128;       // tokens per work unit
constexpr int kDecWarps = 8;     // 8 warps x 16 keys = one unit
constexpr int kDecThreads = kDecWarps * 32;

struct DecodeParams {
  const void* q;
  void* out;
  const void* k_cache;
  const void* v_cache;
  const int32_t* seq_lens_bh;
  const int32_t* page_table;
  const int32_t* batch_mapping;
  float* scratch;     // [(BH + C)] slots of G*(D+2) floats
  int32_t* counters;  // [BH]
  int B, HQ, HKV, G, PS, NLP;
  float scale_log2;
};

// One warp's 16-key slice: K rows r / r+8 (4 chunks each), V for 4 keys x 2 chunks.
struct Slice {
  uint4 k_lo[4], k_hi[4];  // rows r, r+8; chunk (j + 4c)
  uint4 v[4][2];           // keys {2j,2j+1,2j+8,2j+9}; chunks r, 8+r
};

template <int D>
__device__ __forceinline__ void load_slice(Slice& s, const char* __restrict__ kc,
                                           const char* __restrict__ vc, int64_t row0, int n_valid,
                                           int r, int j) {
  constexpr int ROWB = D * 2;
  const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    s.k_lo[c] = (r < n_valid) ? ldg_nc_v4(kc + (row0 + r) * ROWB + (j + 4 * c) * 16) : z;
    s.k_hi[c] = (r + 8 < n_valid) ? ldg_nc_v4(kc + (row0 + r + 8) * ROWB + (j + 4 * c) * 16) : z;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int key = 2 * j + (i & 1) + ((i >> 1) << 3);
    const bool ok = key < n_valid;
    s.v[i][0] = ok ? ldg_nc_v4(vc + (row0 + key) * ROWB + r * 16) : z;
    s.v[i][1] = ok ? ldg_nc_v4(vc + (row0 + key) * ROWB + (8 + r) * 16) : z;
  }
}

template <bool BF16, int D>
__global__ void __launch_bounds__(kDecThreads, 1) decode_attn_kernel(const DecodeParams p) {
  static_assert(D == 128, "decode kernel is specialised for head_dim 128");
  constexpr int KSTEPS = D / 16;  // 8
  constexpr int MT = D / 16;      // 8 output m-tiles (16 d each)
  constexpr int ROW = D + 2;      // merged row: O[D], m, l

  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int BH = p.B * p.HKV;
  int32_t* s_prefix = reinterpret_cast<int32_t*>(smem_raw);  // [BH + 1]
  float* s_merge = reinterpret_cast<float*>(smem_raw + (((BH + 1) * 4 + 15) & ~15));
  // s_merge: [kDecWarps][8][ROW]  then s_final [8][ROW]
  float* s_final = s_merge + kDecWarps * 8 * ROW;
  __shared__ int32_t s_warp_tot[kDecWarps];
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r = lane >> 2, j = lane & 3;
  const int C = gridDim.x, c = blockIdx.x;
  const int G = p.G;

  // ---- phase 0: units per head, exclusive prefix over the BH heads ---------
  {
    const int per = cdiv(BH, kDecThreads);
    const int beg = tid * per, end = min(BH, beg + per);
    int local = 0;
    for (int i = beg; i < end; ++i) local += cdiv(p.seq_lens_bh[i], kUnit);
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp_tot[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < warp; ++w) base += s_warp_tot[w];
    int run = base + incl - local;
    for (int i = beg; i < end; ++i) {
      s_prefix[i] = run;
      run += cdiv(p.seq_lens_bh[i], kUnit);
    }
    if (tid == kDecThreads - 1) s_prefix[BH] = run;
    __syncthreads();
  }
  const int U = s_prefix[BH];

  // zero rows for empty heads (reference leaves them uninitialised)
  for (int i = c; i < BH; i += C) {
    if (s_prefix[i + 1] == s_prefix[i]) {
      const int b = i / p.HKV, h = i % p.HKV;
      uint32_t* o = reinterpret_cast<uint32_t*>(p.out) + ((int64_t)b * p.HQ + h * G) * (D / 2);
      for (int e = tid; e < G * D / 2; e += kDecThreads) o[e] = 0u;
    }
  }
  if (U == 0) return;

  const int u_begin = (int)((int64_t)U * c / C);
  const int u_end = (int)((int64_t)U * (c + 1) / C);
  if (u_begin >= u_end) return;

  // head containing u_begin: largest i with s_prefix[i] <= u_begin
  int head;
  {
    int lo = 0, hi = BH;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_prefix[mid] <= u_begin) lo = mid; else hi = mid;
    }
    head = lo;
  }

  const char* kc = reinterpret_cast<const char*>(p.k_cache);
  const char* vc = reinterpret_cast<const char*>(p.v_cache);

  int u = u_begin;
  while (u < u_end) {
    while (s_prefix[head + 1] <= u) ++head;  // skip finished / empty heads
    const int h_beg = s_prefix[head], h_end = s_prefix[head + 1];
    const int seg_end = min(u_end, h_end);
    const int ua = u - h_beg, ub = seg_end - h_beg;  // units of this head
    const int b = head / p.HKV, h = head % p.HKV;
    const int L = p.seq_lens_bh[head];
    const int32_t* pt = p.page_table + ((int64_t)p.batch_mapping[b] * p.HKV + h) * p.NLP;

    // ---- Q^T as B operand (n = query r, k = permuted d) ----
    uint32_t qb[KSTEPS][2];
    {
      const uint4 z = make_uint4(0, 0, 0, 0);
      const char* qrow = reinterpret_cast<const char*>(p.q) + ((int64_t)b * p.HQ + h * G + r) * (D * 2);
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const uint4 v = (r < G) ? *reinterpret_cast<const uint4*>(qrow + (j + 4 * cc) * 16) : z;
        qb[2 * cc][0] = v.x; qb[2 * cc][1] = v.y;
        qb[2 * cc + 1][0] = v.z; qb[2 * cc + 1][1] = v.w;
      }
    }

    float acc[MT][4];
#pragma unroll
    for (int t = 0; t < MT; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    // first slice of this warp inside [ua, ub)
    int uu = ua;
    auto slice_pos = [&](int unit, int64_t& row0, int& n_valid) {
      const int t0 = unit * kUnit + warp * 16;
      n_valid = min(16, L - t0);
      if (n_valid > 0) {
        const int pg = t0 / p.PS;
        row0 = (int64_t)pt[pg] * p.PS + (t0 - pg * p.PS);
      } else {
        row0 = 0;
      }
    };
    Slice cur, nxt;
    int64_t row0;
    int n_valid;
    slice_pos(uu, row0, n_valid);
    if (n_valid > 0) load_slice<D>(cur, kc, vc, row0, n_valid, r, j);

    while (uu < ub && n_valid > 0) {
      // prefetch the warp's slice of the next unit
      int n_valid_next = 0;
      if (uu + 1 < ub) {
        int64_t row_next;
        slice_pos(uu + 1, row_next, n_valid_next);
        if (n_valid_next > 0) load_slice<D>(nxt, kc, vc, row_next, n_valid_next, r, j);
      }

      // ---- S^T[16 keys x 8 queries] ----
      float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        const int cc = ks >> 1;
        const uint32_t a0 = (ks & 1) ? cur.k_lo[cc].z : cur.k_lo[cc].x;
        const uint32_t a1 = (ks & 1) ? cur.k_hi[cc].z : cur.k_hi[cc].x;
        const uint32_t a2 = (ks & 1) ? cur.k_lo[cc].w : cur.k_lo[cc].y;
        const uint32_t a3 = (ks & 1) ? cur.k_hi[cc].w : cur.k_hi[cc].y;
        mma16816<BF16>(s, a0, a1, a2, a3, qb[ks][0], qb[ks][1]);
      }
      const bool ok_lo = r < n_valid, ok_hi = r + 8 < n_valid;
      const float x0 = ok_lo ? s[0] * p.scale_log2 : -INFINITY;
      const float x1 = ok_lo ? s[1] * p.scale_log2 : -INFINITY;
      const float x2 = ok_hi ? s[2] * p.scale_log2 : -INFINITY;
      const float x3 = ok_hi ? s[3] * p.scale_log2 : -INFINITY;
      float mx0 = fmaxf(x0, x2), mx1 = fmaxf(x1, x3);
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
      }
      float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      // a NaN/-inf column max (garbage q) must not poison the rest
      if (!(mn0 > -INFINITY)) mn0 = 0.f;
      if (!(mn1 > -INFINITY)) mn1 = 0.f;
      const float al0 = ex2(m0 - mn0), al1 = ex2(m1 - mn1);
      m0 = mn0; m1 = mn1;
      const float p0 = ex2(x0 - mn0), p1 = ex2(x1 - mn1), p2 = ex2(x2 - mn0), p3 = ex2(x3 - mn1);
      l0 = l0 * al0 + p0 + p2;
      l1 = l1 * al1 + p1 + p3;
      if (__any_sync(0xffffffffu, (al0 != 1.f) | (al1 != 1.f))) {
#pragma unroll
        for (int t = 0; t < MT; ++t) {
          acc[t][0] *= al0; acc[t][1] *= al1; acc[t][2] *= al0; acc[t][3] *= al1;
        }
      }
      const uint32_t pb0 = movmatrix_trans(pack2<BF16>(p0, p1));
      const uint32_t pb1 = movmatrix_trans(pack2<BF16>(p2, p3));

      // ---- O^T[D x 8] += V^T[D x 16] . P^T[16 x 8] ----
#pragma unroll
      for (int t = 0; t < MT; ++t) {
        const uint32_t sel = (t & 1) ? 0x7632u : 0x5410u;
        const int w = t >> 1;
        const uint32_t a0 = __byte_perm(u4_get(cur.v[0][0], w), u4_get(cur.v[1][0], w), sel);
        const uint32_t a1 = __byte_perm(u4_get(cur.v[0][1], w), u4_get(cur.v[1][1], w), sel);
        const uint32_t a2 = __byte_perm(u4_get(cur.v[2][0], w), u4_get(cur.v[3][0], w), sel);
        const uint32_t a3 = __byte_perm(u4_get(cur.v[2][1], w), u4_get(cur.v[3][1], w), sel);
        mma16816<BF16>(acc[t], a0, a1, a2, a3, pb0, pb1);
      }

      ++uu;
      n_valid = n_valid_next;
      if (n_valid > 0) cur = nxt;
    }

    // ---- per-warp state -> smem ----
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      l0 += __shfl_xor_sync(0xffffffffu, l0, o);
      l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    }
    {
      float* mw = s_merge + warp * 8 * ROW;
      const int q0 = 2 * j, q1 = 2 * j + 1;
#pragma unroll
      for (int t = 0; t < MT; ++t) {
        mw[q0 * ROW + 8 * r + t] = acc[t][0];
        mw[q1 * ROW + 8 * r + t] = acc[t][1];
        mw[q0 * ROW + 64 + 8 * r + t] = acc[t][2];
        mw[q1 * ROW + 64 + 8 * r + t] = acc[t][3];
      }
      if (r == 0) {
        mw[q0 * ROW + D] = m0; mw[q0 * ROW + D + 1] = l0;
        mw[q1 * ROW + D] = m1; mw[q1 * ROW + D + 1] = l1;
      }
    }
    __syncthreads();

    // ---- merge the 8 warps -> s_final[q][ROW] (unnormalised O, m, l) ----
    for (int e = tid; e < G * ROW; e += kDecThreads) {
      const int q = e / ROW, d = e - q * ROW;
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < kDecWarps; ++w) M = fmaxf(M, s_merge[(w * 8 + q) * ROW + D]);
      float val = 0.f;
      if (d < D) {
#pragma unroll
        for (int w = 0; w < kDecWarps; ++w) {
          const float mw = s_merge[(w * 8 + q) * ROW + D];
          const float sc = (mw > -INFINITY) ? ex2(mw - M) : 0.f;
          val += s_merge[(w * 8 + q) * ROW + d] * sc;
        }
      } else if (d == D) {
        val = M;
      } else {
#pragma unroll
        for (int w = 0; w < kDecWarps; ++w) {
          const float mw = s_merge[(w * 8 + q) * ROW + D];
          const float sc = (mw > -INFINITY) ? ex2(mw - M) : 0.f;
          val += s_merge[(w * 8 + q) * ROW + D + 1] * sc;
        }
      }
      s_final[e] = val;
    }
    __syncthreads();

    const bool whole = (ua == 0) && (ub == h_end - h_beg);
    uint32_t* orow = reinterpret_cast<uint32_t*>(p.out) + ((int64_t)b * p.HQ + h * G) * (D / 2);
    if (whole) {
      for (int e = tid; e < G * D / 2; e += kDecThreads) {
        const int q = e / (D / 2), d2 = e - q * (D / 2);
        const float lq = s_final[q * ROW + D + 1];
        const float inv = lq > 0.f ? 1.f / lq : 0.f;
        orow[e] = pack2<BF16>(s_final[q * ROW + 2 * d2] * inv, s_final[q * ROW + 2 * d2 + 1] * inv);
      }
    } else {
      // publish the partial, then take a ticket; the CTA completing the head merges
      float* slot = p.scratch + (int64_t)(head + c) * G * ROW;
      for (int e = tid; e < G * ROW; e += kDecThreads) __stcg(slot + e, s_final[e]);
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        const int done = atomicAdd(p.counters + head, ub - ua) + (ub - ua);
        s_last = (done == h_end - h_beg);
      }
      __syncthreads();
      if (s_last) {
        __threadfence();
        const int c_first = (int)((((int64_t)h_beg + 1) * C - 1) / U);
        const int c_last = (int)(((int64_t)h_end * C - 1) / U);
        for (int e = tid; e < G * D / 2; e += kDecThreads) {
          const int q = e / (D / 2), d2 = e - q * (D / 2);
          // CTAs with an empty unit range (U < C) never wrote their slot
          auto empty_cta = [&](int cc) {
            return (int)((int64_t)U * cc / C) >= (int)((int64_t)U * (cc + 1) / C);
          };
          float M = -INFINITY;
          for (int cc = c_first; cc <= c_last; ++cc) {
            if (empty_cta(cc)) continue;
            M = fmaxf(M, __ldcg(p.scratch + ((int64_t)(head + cc) * G + q) * ROW + D));
          }
          float o0 = 0.f, o1 = 0.f, lsum = 0.f;
          for (int cc = c_first; cc <= c_last; ++cc) {
            if (empty_cta(cc)) continue;
            const float* sl = p.scratch + ((int64_t)(head + cc) * G + q) * ROW;
            const float mw = __ldcg(sl + D);
            const float sc = (mw > -INFINITY) ? ex2(mw - M) : 0.f;
            const float2 ov = __ldcg(reinterpret_cast<const float2*>(sl + 2 * d2));
            o0 += ov.x * sc; o1 += ov.y * sc;
            lsum += __ldcg(sl + D + 1) * sc;
          }
          const float inv = lsum > 0.f ? 1.f / lsum : 0.f;
          orow[e] = pack2<BF16>(o0 * inv, o1 * inv);
        }
        if (tid == 0) p.counters[head] = 0;
      }
    }
    __syncthreads();  // s_merge / s_final reused by the next segment
    u = seg_end;
  }
}

}  // namespace cvllm

using namespace cvllm;

static int decode_ctas(int num_ctas) { return num_ctas > 0 ? num_ctas : sm_count(); }

extern "C" size_t cvllm_decode_attention_workspace_bytes(int B, int HQ, int HKV, int D, int num_ctas) {
  if (B <= 0 || HKV <= 0 || HQ <= 0) return 0;
  const int G = HQ / HKV;
  const size_t BH = (size_t)B * HKV;
  const size_t slots = BH + (size_t)decode_ctas(num_ctas);
  const size_t counters = (BH * 4 + 255) & ~(size_t)255;
  return counters + slots * G * (D + 2) * sizeof(float);
}

extern "C" int cvllm_decode_attention(const void* q, void* out, const void* k_cache,
                                      const void* v_cache, const int32_t* seq_lens_bh,
                                      const int32_t* page_table, const int32_t* batch_mapping,
                                      void* workspace, size_t workspace_bytes, int B, int HQ, int HKV,
                                      int D, int PS, int NLP, float sm_scale, int dtype, int num_ctas,
                                      cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B >= 0 && HQ > 0 && HKV > 0 && PS > 0 && NLP > 0);
  if (B == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(q && out && k_cache && v_cache && seq_lens_bh && page_table && batch_mapping &&
                  workspace);
  CVLLM_CHECK_ARG(HQ % HKV == 0);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  const int G = HQ / HKV;
  CVLLM_CHECK_SUPPORTED(D == 128);
  CVLLM_CHECK_SUPPORTED(G <= 8);
  CVLLM_CHECK_SUPPORTED(PS % 16 == 0);
  CVLLM_CHECK_ALIGN(aligned16(q) && aligned16(out) && aligned16(k_cache) && aligned16(v_cache) &&
                    aligned16(workspace));
  const int C = decode_ctas(num_ctas);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_decode_attention_workspace_bytes(B, HQ, HKV, D, num_ctas));
  const size_t BH = (size_t)B * HKV;
  const size_t counters_bytes = (BH * 4 + 255) & ~(size_t)255;
  auto st = static_cast<cudaStream_t>(stream);
  CVLLM_CHECK_CUDA(cudaMemsetAsync(workspace, 0, counters_bytes, st));

  DecodeParams p;
  p.q = q; p.out = out; p.k_cache = k_cache; p.v_cache = v_cache;
  p.seq_lens_bh = seq_lens_bh; p.page_table = page_table; p.batch_mapping = batch_mapping;
  p.counters = reinterpret_cast<int32_t*>(workspace);
  p.scratch = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + counters_bytes);
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = G; p.PS = PS; p.NLP = NLP;
  p.scale_log2 = sm_scale * 1.4426950408889634f;

  const size_t smem = (((BH + 1) * 4 + 15) & ~(size_t)15) + (size_t)(kDecWarps + 1) * 8 * (D + 2) * 4;
  constexpr int kMaxSmem = 200 * 1024;
  CVLLM_CHECK_SUPPORTED(smem <= (size_t)kMaxSmem);
  // raise the dynamic-smem cap once per process (not a stream op; kept out of graph capture)
  static bool attr_set = false;
  if (!attr_set) {
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(decode_attn_kernel<true, 128>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem));
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(decode_attn_kernel<false, 128>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem));
    attr_set = true;
  }
  if (dtype == CVLLM_BF16) {
    decode_attn_kernel<true, 128><<<C, kDecThreads, smem, st>>>(p);
  } else {
    decode_attn_kernel<false, 128><<<C, kDecThreads, smem, st>>>(p);
  }
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
