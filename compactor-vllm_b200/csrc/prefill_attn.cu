// Causal head-sparse varlen prefill attention over [paged prefix || new K/V].
//
// Replaces attention/sparse_varlen_kernel.py:11-197,265-519 (Triton, tl.dot
// tiles <= 256x64 fed by pointer gathers) with a tcgen05 / TMEM / TMA kernel:
//
//   CTA            one 128-row query tile of one query head (rows = tokens)
//   warp 0         TMA producer: Q once, then K and V tiles of 128 keys through
//                  a 3-stage mbarrier ring.  Prefix tiles come straight out of
//                  the paged cache (a 128-row tile never crosses a page:
//                  PS % 128 == 0), new-token tiles from the packed k / v.
//   warp 1         UMMA issuer: S = Q.K^T (SS, both K-major, 128x128x128) into
//                  one of two TMEM S buffers; O += P.V (TS: P read from TMEM,
//                  V MN-major from smem).  QK(j+1) is issued before PV(j) so
//                  the tensor pipe works on the next S while softmax(j) runs.
//   warps 2-5      softmax: one thread per query row (no shuffles), S row from
//                  TMEM, running max with lazy rescale (O is only touched when
//                  the max grows by > 8 in log2 units), P written back to TMEM
//                  as bf16 over the S columns it came from; epilogue O / l.
//
// TMEM: S0 [0,128) S1 [128,256) O [256,384)  (fp32 columns, 128 lanes).
// Tensor-pipe bound: 4*D flops per (query row, key) pair — see DESIGN.md.
#include <stdlib.h>

#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

constexpr int kBM = 128;  // query rows per CTA
constexpr int kBN = 128;  // keys per tile
constexpr int kHD = 128;  // head dim
constexpr int kStages = 3;
constexpr int kTileBytes = kBN * kHD * 2;  // 32 KB
constexpr int kBoxBytes = kTileBytes / 2;  // one [128 x 64] swizzled box
constexpr int kPrefillThreads = 192;
constexpr float kRescaleThreshold = 8.0f;  // log2 units

struct PrefillParams {
  void* out;
  const int32_t* seq_lens_bh;
  const int32_t* page_table;
  const int32_t* batch_mapping;
  const int32_t* cu_seqlens_q;
  int B, HQ, HKV, G, PS, NLP;
  int64_t o_stride_n;
  float scale_log2;
  float rescale_threshold;
};

struct PrefillSmem {
  // data (all 1024-aligned: SWIZZLE_128B atoms are 8 rows x 128 B)
  unsigned char q[kTileBytes];
  unsigned char k[kStages][kTileBytes];
  unsigned char v[kStages][kTileBytes];
  uint64_t q_full;
  uint64_t k_full[kStages], k_empty[kStages];
  uint64_t v_full[kStages], v_empty[kStages];
  uint64_t s_full[2], p_full[2];
  uint64_t o_done;   // one phase per PV(i): conditional O-correction waits (at most one phase behind)
  uint64_t o_final;  // single phase: last PV complete (the epilogue may be >1 phase behind o_done)
  uint32_t tmem_base;
};

template <bool BF16>
__global__ void __launch_bounds__(kPrefillThreads, 1)
prefill_attn_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                    const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_kc,
                    const __grid_constant__ CUtensorMap tm_vc, const PrefillParams p) {
  extern __shared__ unsigned char smem_raw[];
  PrefillSmem& sm = *reinterpret_cast<PrefillSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, hq = blockIdx.y, hkv = hq / p.G;
  const int q_beg = p.cu_seqlens_q[b];
  const int La = p.cu_seqlens_q[b + 1] - q_beg;
  const int n_qt = (La + kBM - 1) / kBM;
  const int qt = n_qt - 1 - (int)blockIdx.x;  // heaviest (latest) tiles first
  if (qt < 0) return;
  const int q0 = qt * kBM;
  const int Lc = p.seq_lens_bh[b * p.HKV + hkv];
  const int n_pre = (Lc + kBN - 1) / kBN;
  const int n_tiles = n_pre + qt + 1;

  // ---------------- setup ----------------
  if (threadIdx.x == 0) {
    mbar_init(&sm.q_full, 1);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sm.k_full[s], 1);
      mbar_init(&sm.k_empty[s], 1);
      mbar_init(&sm.v_full[s], 1);
      mbar_init(&sm.v_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sm.s_full[s], 1);
      mbar_init(&sm.p_full[s], 128);
    }
    mbar_init(&sm.o_done, 1);
    mbar_init(&sm.o_final, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(&sm.tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_kc);
    tma_prefetch_desc(&tm_vc);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_s0 = tmem, tmem_o = tmem + 256;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      mbar_expect_tx(&sm.q_full, kTileBytes);
      tma_load_3d(sm.q, &tm_q, &sm.q_full, 0, hq, q_beg + q0);
      tma_load_3d(sm.q + kBoxBytes, &tm_q, &sm.q_full, 64, hq, q_beg + q0);
      const int32_t* pt = p.page_table + ((int64_t)p.batch_mapping[b] * p.HKV + hkv) * p.NLP;
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j % kStages;
        const uint32_t ph = (j / kStages) & 1;
        int row = 0;
        const bool pre = j < n_pre;
        if (pre) {
          const int t0 = j * kBN;
          const int pg = t0 / p.PS;
          row = pt[pg] * p.PS + (t0 - pg * p.PS);
        } else {
          row = q_beg + (j - n_pre) * kBN;
        }
        mbar_wait(&sm.k_empty[st], ph ^ 1);
        mbar_expect_tx(&sm.k_full[st], kTileBytes);
        if (pre) {
          tma_load_2d(sm.k[st], &tm_kc, &sm.k_full[st], 0, row);
          tma_load_2d(sm.k[st] + kBoxBytes, &tm_kc, &sm.k_full[st], 64, row);
        } else {
          tma_load_3d(sm.k[st], &tm_k, &sm.k_full[st], 0, hkv, row);
          tma_load_3d(sm.k[st] + kBoxBytes, &tm_k, &sm.k_full[st], 64, hkv, row);
        }
        mbar_wait(&sm.v_empty[st], ph ^ 1);
        mbar_expect_tx(&sm.v_full[st], kTileBytes);
        if (pre) {
          tma_load_2d(sm.v[st], &tm_vc, &sm.v_full[st], 0, row);
          tma_load_2d(sm.v[st] + kBoxBytes, &tm_vc, &sm.v_full[st], 64, row);
        } else {
          tma_load_3d(sm.v[st], &tm_v, &sm.v_full[st], 0, hkv, row);
          tma_load_3d(sm.v[st] + kBoxBytes, &tm_v, &sm.v_full[st], 64, hkv, row);
        }
      }
    }
  } else if (warp == 1) {
    // ============================ UMMA issuer ============================
    constexpr uint32_t idesc_qk = umma_idesc(BF16, kBM, kBN, false, false);
    constexpr uint32_t idesc_pv = umma_idesc(BF16, kBM, kHD, false, true);
    const uint64_t q_desc = umma_smem_desc(smem_u32(sm.q), 16, 1024);
    // warp-uniform copies for the issuing thread (no R2UR loop per UMMA; see prefill_attn2.cu)
    const uint32_t tmem_s0u = __shfl_sync(0xffffffffu, tmem_s0, 0), tmem_ou = tmem_s0u + 256;

    auto issue_pv = [&](int i) {
      const int sb = i & 1, st = i % kStages;
      mbar_wait(&sm.p_full[sb], (i >> 1) & 1);
      mbar_wait(&sm.v_full[st], (i / kStages) & 1);
      if (i == n_pre - 1 && (Lc % kBN) != 0) {
        // ragged last prefix tile: rows >= Lc % 128 of the page are uninitialised memory;
        // P is exactly 0 there but 0 * NaN would poison O, so zero those V rows.
        const int valid = Lc % kBN;
        uint4* v4 = reinterpret_cast<uint4*>(sm.v[st]);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int idx = lane; idx < (kBN - valid) * 16; idx += 32) {
          const int r = valid + idx / 16, c = idx % 16;  // 16 x 16B per row across the 2 boxes
          v4[(c >> 3) * (kBoxBytes / 16) + r * 8 + (c & 7)] = z;
        }
        fence_proxy_async_smem();
        __syncwarp();
      }
      tc_fence_after();
      if (lane == 0) {
        const uint64_t v_desc = umma_smem_desc(smem_u32(sm.v[st]), kBoxBytes, 1024);
        const uint32_t p_tmem = tmem_s0u + sb * kBN;
#pragma unroll
        for (int kk = 0; kk < kBN / 16; ++kk)
          umma_ts(tmem_ou, p_tmem + kk * 8, umma_desc_add(v_desc, kk * 2048), idesc_pv, (i > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&sm.v_empty[st]);
        umma_commit(&sm.o_done);
        if (i == n_tiles - 1) umma_commit(&sm.o_final);
      }
      __syncwarp();
    };

    mbar_wait(&sm.q_full, 0);
    for (int j = 0; j < n_tiles; ++j) {
      const int st = j % kStages;
      mbar_wait(&sm.k_full[st], (j / kStages) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint64_t k_desc = umma_smem_desc(smem_u32(sm.k[st]), 16, 1024);
        const uint32_t s_tmem = tmem_s0u + (j & 1) * kBN;
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk) {
          const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
          umma_ss(s_tmem, umma_desc_add(q_desc, off), umma_desc_add(k_desc, off), idesc_qk, kk > 0 ? 1u : 0u);
        }
        umma_commit(&sm.k_empty[st]);
        umma_commit(&sm.s_full[j & 1]);
      }
      __syncwarp();
      if (j >= 1) issue_pv(j - 1);
    }
    issue_pv(n_tiles - 1);
  } else {
    // ====================== softmax / correction / epilogue ======================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const bool row_valid = (q0 + row) < La;
    float m_ref = -INFINITY, l = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
      const int sb = j & 1;
      mbar_wait(&sm.s_full[sb], (j >> 1) & 1);
      tc_fence_after();
      const uint32_t s_tmem = tmem_s0 + sb * kBN + lane_base;
      float s[kBN];
#pragma unroll
      for (int c = 0; c < kBN / 32; ++c) {
        uint32_t tmp[32];
        tmem_ld32(s_tmem + c * 32, tmp);
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(tmp[i]);
      }
      tc_wait_ld();

      // ---- masks (boundary tiles only) ----
      if (j < n_pre) {
        const int valid = Lc - j * kBN;
        if (valid < kBN) {
#pragma unroll
          for (int c = 0; c < kBN; ++c)
            if (c >= valid) s[c] = -INFINITY;
        }
      } else if (j == n_tiles - 1) {
#pragma unroll
        for (int c = 0; c < kBN; ++c)
          if (c > row) s[c] = -INFINITY;
      }

      float mx = s[0];
#pragma unroll
      for (int c = 1; c < kBN; ++c) mx = fmaxf(mx, s[c]);
      const float m_new = fmaxf(m_ref, mx * p.scale_log2);
      const bool need = (j == 0) || (m_new > m_ref + p.rescale_threshold);
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = ex2(m_ref - m_new);  // m_ref == -inf on the first tile -> 0
        m_ref = m_new;
        l *= alpha;
        if (j > 0) {
          // O holds PV(0..j-1); PV(j-1) may still be in flight
          mbar_wait(&sm.o_done, (j - 1) & 1);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < kHD / 32; ++c) {
            uint32_t o[32];
            tmem_ld32(tmem_o + lane_base + c * 32, o);
            tc_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(tmem_o + lane_base + c * 32, o);
          }
        }
      }

      // ---- P = exp2(S*scale - m_ref), row sum, bf16 pack, back to TMEM ----
      const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
      // packed fp32 pairs (FFMA2 / FADD2: half the FMA-pipe issue slots of the scalar form, same roundings)
      const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), nm2 = make_float2(neg_m, neg_m);
      float2 l2a = make_float2(0.f, 0.f), l2b = make_float2(0.f, 0.f);
#pragma unroll
      for (int c = 0; c < kBN / 64; ++c) {
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float2 t0 = __ffma2_rn(make_float2(s[c * 64 + 2 * i], s[c * 64 + 2 * i + 1]), sc2, nm2);
          const float2 t1 = __ffma2_rn(make_float2(s[c * 64 + 2 * i + 2], s[c * 64 + 2 * i + 3]), sc2, nm2);
          const float2 e0 = make_float2(ex2(t0.x), ex2(t0.y)), e1 = make_float2(ex2(t1.x), ex2(t1.y));
          l2a = __fadd2_rn(l2a, e0);
          l2b = __fadd2_rn(l2b, e1);
          pk[i] = pack2<BF16>(e0.x, e0.y);
          pk[i + 1] = pack2<BF16>(e1.x, e1.y);
        }
        tmem_st32(s_tmem + c * 32, pk);
      }
      l += (l2a.x + l2b.x) + (l2a.y + l2b.y);
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(&sm.p_full[sb]);
    }

    // ---- epilogue: O / l -> out ----
    // NOT o_done: only PV(n-3) is known complete here, and a parity wait two phases ahead
    // of the barrier passes spuriously.
    mbar_wait(&sm.o_final, 0);
    tc_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    char* orow = reinterpret_cast<char*>(p.out) +
                 ((int64_t)(q_beg + q0 + row) * p.o_stride_n + (int64_t)hq * kHD) * 2;
#pragma unroll
    for (int c = 0; c < kHD / 32; ++c) {
      uint32_t o[32];
      tmem_ld32(tmem_o + lane_base + c * 32, o);
      tc_wait_ld();
      if (row_valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 w;
          w.x = pack2<BF16>(__uint_as_float(o[8 * i + 0]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
          w.y = pack2<BF16>(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
          w.z = pack2<BF16>(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
          w.w = pack2<BF16>(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + c * 64 + i * 16) = w;
        }
      }
    }
  }

  // ---------------- teardown ----------------
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------
// UMMA self test: one 128x128x128 product through each operand path.
//   mode 0: D = A . B^T   A [128 x 128] K-major smem, B [128 x 128] K-major smem
//   mode 1: D = A . B     A [128 x 128] from TMEM (16-bit packed), B [128(k) x 128(n)] MN-major smem
// lbo / sbo override the B-operand descriptor so encodings can be swept from the host.
// ---------------------------------------------------------------------------
struct SelfTestSmem {
  unsigned char a[kTileBytes];
  unsigned char b[kTileBytes];
  uint64_t full, done;
  uint32_t tmem_base;
};

template <bool BF16>
__global__ void __launch_bounds__(128, 1)
umma_selftest_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                     const void* a_gmem, float* d_out, int mode, uint32_t lbo, uint32_t sbo) {
  extern __shared__ unsigned char smem_raw[];
  SelfTestSmem& sm = *reinterpret_cast<SelfTestSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&sm.full, 1);
    mbar_init(&sm.done, 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    tmem_alloc(&sm.tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const int row = warp * 32 + lane;

  if (threadIdx.x == 0) {
    mbar_expect_tx(&sm.full, mode == 0 ? 2 * kTileBytes : kTileBytes);
    if (mode == 0) {
      tma_load_2d(sm.a, &tm_a, &sm.full, 0, 0);
      tma_load_2d(sm.a + kBoxBytes, &tm_a, &sm.full, 64, 0);
    }
    tma_load_2d(sm.b, &tm_b, &sm.full, 0, 0);
    tma_load_2d(sm.b + kBoxBytes, &tm_b, &sm.full, 64, 0);
  }
  if (mode == 1) {
    // A row -> TMEM columns [256, 320): word w holds elements (2w, 2w+1)
    const uint32_t* arow = reinterpret_cast<const uint32_t*>(a_gmem) + row * 64;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t w[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) w[i] = arow[c * 32 + i];
      tmem_st32(tmem + 256 + lane_base + c * 32, w);
    }
    tc_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 0) {
    mbar_wait(&sm.full, 0);
    tc_fence_after();
    const uint32_t a_addr = smem_u32(sm.a), b_addr = smem_u32(sm.b);
    if (mode == 0) {
      constexpr uint32_t idesc = umma_idesc(BF16, 128, 128, false, false);
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
        umma_ss(tmem, umma_smem_desc(a_addr + off, 16, 1024), umma_smem_desc(b_addr + off, lbo, sbo),
                idesc, kk > 0 ? 1u : 0u);
      }
    } else {
      constexpr uint32_t idesc = umma_idesc(BF16, 128, 128, false, true);
      for (int kk = 0; kk < 8; ++kk)
        umma_ts(tmem, tmem + 256 + kk * 8, umma_smem_desc(b_addr + kk * 2048, lbo, sbo), idesc,
                kk > 0 ? 1u : 0u);
    }
    umma_commit(&sm.done);
  }
  mbar_wait(&sm.done, 0);
  tc_fence_after();
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t o[32];
    tmem_ld32(tmem + lane_base + c * 32, o);
    tc_wait_ld();
#pragma unroll
    for (int i = 0; i < 32; ++i) d_out[row * 128 + c * 32 + i] = __uint_as_float(o[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace cvllm

using namespace cvllm;

int cvllm_launch_prefill2(const CUtensorMap& tm_q, const CUtensorMap& tm_k, const CUtensorMap& tm_v,
                          const CUtensorMap& tm_kc, const CUtensorMap& tm_vc, void* out,
                          const int32_t* seq_lens_bh, const int32_t* page_table,
                          const int32_t* batch_mapping, const int32_t* cu_seqlens_q, int B, int HQ,
                          int HKV, int PS, int NLP, int64_t o_stride_n, int max_seqlen_q,
                          float scale_log2, float rescale_threshold, bool bf16, cudaStream_t st);

extern "C" int cvllm_prefill_attention(const void* q, const void* k, const void* v, void* out,
                                       const void* k_cache, const void* v_cache,
                                       const int32_t* seq_lens_bh, const int32_t* page_table,
                                       const int32_t* batch_mapping, const int32_t* cu_seqlens_q, int B,
                                       int N, int HQ, int HKV, int D, int PS, int NLP, int64_t cache_rows,
                                       int64_t q_stride_n, int64_t k_stride_n, int64_t v_stride_n,
                                       int64_t o_stride_n, int max_seqlen_q, float sm_scale, int dtype,
                                       cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && N >= 0 && HQ > 0 && HKV > 0 && PS > 0 && NLP > 0 && max_seqlen_q >= 0);
  if (N == 0 || max_seqlen_q == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(q && k && v && out && k_cache && v_cache && seq_lens_bh && page_table &&
                  batch_mapping && cu_seqlens_q);
  CVLLM_CHECK_ARG(HQ % HKV == 0 && cache_rows > 0);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_SUPPORTED(D == kHD);
  CVLLM_CHECK_SUPPORTED(PS % kBN == 0);
  CVLLM_CHECK_ALIGN(aligned16(q) && aligned16(k) && aligned16(v) && aligned16(out) &&
                    aligned16(k_cache) && aligned16(v_cache));
  CVLLM_CHECK_ALIGN(q_stride_n % 8 == 0 && k_stride_n % 8 == 0 && v_stride_n % 8 == 0 &&
                    o_stride_n % 8 == 0);
  const bool bf16 = dtype == CVLLM_BF16;

  CUtensorMap tm_q, tm_k, tm_v, tm_kc, tm_vc;
  {
    const uint32_t box3[3] = {64, 1, (uint32_t)kBM};
    const uint64_t dq[3] = {(uint64_t)D, (uint64_t)HQ, (uint64_t)N};
    const uint64_t sq[2] = {(uint64_t)D * 2, (uint64_t)q_stride_n * 2};
    int rc = make_tmap_16b(&tm_q, q, 3, dq, sq, box3, bf16);
    if (rc) return rc;
    const uint64_t dk[3] = {(uint64_t)D, (uint64_t)HKV, (uint64_t)N};
    const uint64_t sk[2] = {(uint64_t)D * 2, (uint64_t)k_stride_n * 2};
    rc = make_tmap_16b(&tm_k, k, 3, dk, sk, box3, bf16);
    if (rc) return rc;
    const uint64_t sv[2] = {(uint64_t)D * 2, (uint64_t)v_stride_n * 2};
    rc = make_tmap_16b(&tm_v, v, 3, dk, sv, box3, bf16);
    if (rc) return rc;
    const uint32_t box2[2] = {64, (uint32_t)kBN};
    const uint64_t dc[2] = {(uint64_t)D, (uint64_t)cache_rows};
    const uint64_t sc[1] = {(uint64_t)D * 2};
    rc = make_tmap_16b(&tm_kc, k_cache, 2, dc, sc, box2, bf16);
    if (rc) return rc;
    rc = make_tmap_16b(&tm_vc, v_cache, 2, dc, sc, box2, bf16);
    if (rc) return rc;
  }

  PrefillParams p;
  p.out = out; p.seq_lens_bh = seq_lens_bh; p.page_table = page_table;
  p.batch_mapping = batch_mapping; p.cu_seqlens_q = cu_seqlens_q;
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = HQ / HKV; p.PS = PS; p.NLP = NLP;
  p.o_stride_n = o_stride_n;
  p.scale_log2 = sm_scale * 1.4426950408889634f;
  p.rescale_threshold = kRescaleThreshold;
  // test hook: 0 forces an O rescale whenever a row max grows (exercises the correction path)
  {
    const float t = test_options().prefill_threshold.load(std::memory_order_relaxed);
    if (t >= 0.f) p.rescale_threshold = t;
  }

  auto st = static_cast<cudaStream_t>(stream);
  // even GQA group: two query heads of one KV head per CTA, ping-ponged S accumulators
  // (prefill_attn2.cu).  Test option "prefill_v1" forces the one-head-per-CTA kernel (tests cover both).
  if ((HQ / HKV) % 2 == 0 && !test_options().prefill_v1.load(std::memory_order_relaxed))
    return cvllm_launch_prefill2(tm_q, tm_k, tm_v, tm_kc, tm_vc, out, seq_lens_bh, page_table,
                                 batch_mapping, cu_seqlens_q, B, HQ, HKV, PS, NLP, o_stride_n,
                                 max_seqlen_q, p.scale_log2, p.rescale_threshold, bf16, st);
  const size_t smem = sizeof(PrefillSmem) + 1024;
  const dim3 grid((max_seqlen_q + kBM - 1) / kBM, HQ, B);
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn_kernel<true>, (int)smem, attr_bf));
    prefill_attn_kernel<true><<<grid, kPrefillThreads, smem, st>>>(tm_q, tm_k, tm_v, tm_kc, tm_vc, p);
  } else {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn_kernel<false>, (int)smem, attr_h));
    prefill_attn_kernel<false><<<grid, kPrefillThreads, smem, st>>>(tm_q, tm_k, tm_v, tm_kc, tm_vc, p);
  }
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

extern "C" int cvllm_selftest_umma(const void* a, const void* b, float* d, int mode, uint32_t lbo,
                                   uint32_t sbo, int dtype, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(a && b && d && (mode == 0 || mode == 1));
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  const bool bf16 = dtype == CVLLM_BF16;
  CUtensorMap tm_a, tm_b;
  const uint32_t box2[2] = {64, 128};
  const uint64_t dims[2] = {128, 128};
  const uint64_t strides[1] = {256};
  int rc = make_tmap_16b(&tm_a, a, 2, dims, strides, box2, bf16);
  if (rc) return rc;
  rc = make_tmap_16b(&tm_b, b, 2, dims, strides, box2, bf16);
  if (rc) return rc;
  const size_t smem = sizeof(SelfTestSmem) + 1024;
  auto st = static_cast<cudaStream_t>(stream);
  if (bf16) {
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(umma_selftest_kernel<true>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma_selftest_kernel<true><<<1, 128, smem, st>>>(tm_a, tm_b, a, d, mode, lbo, sbo);
  } else {
    CVLLM_CHECK_CUDA(cudaFuncSetAttribute(umma_selftest_kernel<false>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma_selftest_kernel<false><<<1, 128, smem, st>>>(tm_a, tm_b, a, d, mode, lbo, sbo);
  }
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
