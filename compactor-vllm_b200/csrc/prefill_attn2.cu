// Causal head-sparse varlen prefill attention, two query tiles per CTA (GQA pairs).
//
// Same contract as prefill_attn.cu (attention/sparse_varlen_kernel.py:11-197,265-519), used when
// the GQA group size is even.  One CTA = 128 query tokens x TWO query heads of the same KV head:
// the K / V tiles are loaded once and feed both heads, and — the point — the two S = Q.K^T
// accumulators ping-pong so the tensor pipe never waits for a softmax:
//
//   tensor pipe :  S0(0) S1(0) | PV0(0) S0(1) PV1(0) S1(1) | PV0(1) S0(2) PV1(1) S1(2) | ...
//   softmax WG0 :        [ S0(0) -> P0(0) ]          [ S0(1) -> P0(1) ]
//   softmax WG1 :              [ S1(0) -> P1(0) ]          [ S1(1) -> P1(1) ]
//
// i.e. each softmax warpgroup has two UMMA tiles (PV of the other head + the other head's next
// QK^T) of time to turn S into P before its own PV is due.
//
//   warps 0-3     softmax of head 0: one thread per query row, S row TMEM -> registers, running
//   warps 4-7     softmax of head 1  max with lazy O rescale (only when the max grows by > 8 in
//                 log2 units), P written back over S as packed 16-bit, epilogue O / l.
//   warp 8        TMA producer: Q0, Q1 once; K_j, V_j through ONE ring of 5 x 32 KB slots
//                 (K0 V0 K1 V1 ...).  Prefix tiles straight from the paged cache (a 128-row tile
//                 never crosses a page), new-token tiles from the packed k / v.
//   warp 9        UMMA issuer (single thread): QK^T as SS, PV as TS (P read from TMEM).
//   warps 10-11   idle (registers are handed out per 4-warp group: the third group gives its
//                 registers to the softmax groups with setmaxnreg, 56 vs 224 per thread).
//
// TMEM (512 columns): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512).
#include <stdlib.h>

#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

namespace p2 {
constexpr int kBM = 128, kBN = 128, kHD = 128;
constexpr int kSlots = 5;
constexpr int kTileBytes = kBN * kHD * 2;  // 32 KB
constexpr int kBoxBytes = kTileBytes / 2;  // one [128 x 64] swizzled box
constexpr int kThreads = 384;              // 3 warpgroups
constexpr int kRegsSoftmax = 224, kRegsOther = 56;  // 2*128*224 + 128*56 = 64512 <= 65536
}  // namespace p2

struct Prefill2Params {
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

struct Prefill2Smem {
  unsigned char q[2][p2::kTileBytes];
  unsigned char kv[p2::kSlots][p2::kTileBytes];
  uint64_t q_full;
  uint64_t full[p2::kSlots], empty[p2::kSlots];
  uint64_t s_full[2], p_full[2];
  uint64_t p_half[2];   // P of the tile's first 64 keys is in TMEM: the PV UMMAs of those keys may start
  uint64_t o_done[2];   // one phase per PV_t(i)
  uint64_t o_final[2];  // last PV_t complete
  uint32_t tmem_base;
};

// kPolyPairs = e: e of every 8 logit PAIRS take their exponentials on the FMA pipe (ex2_poly2) instead of the MUFU.
// A warp that issues MUFU.EX2 is held for the 8 cycles its scheduler's 4-lane unit needs (59 % of a softmax warp's
// active samples in the ncu source view); ten packed fp32x2 instructions do a PAIR.  Measured with the in-kernel trace
// (cycles per KV tile, together with the converged-warp UMMA issue below, P handed over at once): e = 0: 3161, 1: 3050,
// 2: 2881, 3: 2887, 4: 3218 — beyond a third the FMA-pipe work outweighs the MUFU time it frees.
#ifndef CVLLM_PREFILL_POLY_PAIRS
#define CVLLM_PREFILL_POLY_PAIRS 3
#endif
// P is handed to the issuer in two parts (the first kPSplit 32-key pieces, then the rest): the first PV UMMAs run while
// the softmax group is still on the second part.  With the ELECT-loop issuer this LOST 3 % (the issuer stalled in the
// middle of a group); with the converged-warp issue it takes another 230 cycles off the tile: trace, cycles per KV tile,
// polynomial pairs 3 of 8: split 0: 2887, 1: 2678, 2: 2647, 3: 2668 (and 3021 with 4 of 8 pairs).
#ifndef CVLLM_PREFILL_P_SPLIT
#define CVLLM_PREFILL_P_SPLIT 2
#endif
constexpr int kPSplit = CVLLM_PREFILL_P_SPLIT;  // 32-key pieces of P handed to the issuer ahead of the rest (0: all at once)
constexpr int kPolyPairs = CVLLM_PREFILL_POLY_PAIRS;
// 2^x for two values x <= ~8 with packed fp32x2 arithmetic: round-to-nearest split x = n + f (|f| <= 0.5), degree-3
// polynomial in f (relative error ~1e-4, far below the 2^-9 rounding of P to 16 bits), exponent added as an integer.
// Inputs below -125 (masked logits are -inf) flush to 2^-125, which the 16-bit P and the fp32 row sum take as 0.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  x.x = fmaxf(x.x, -125.f);
  x.y = fmaxf(x.y, -125.f);
  const float2 magic = make_float2(12582912.f, 12582912.f);
  const float2 t = __fadd2_rn(x, magic);
  const float2 n = __fadd2_rn(t, make_float2(-12582912.f, -12582912.f));
  const float2 f = __ffma2_rn(n, make_float2(-1.f, -1.f), x);
  float2 p = __ffma2_rn(f, make_float2(0.05550357f, 0.05550357f), make_float2(0.24022652f, 0.24022652f));
  p = __ffma2_rn(p, f, make_float2(0.69314575f, 0.69314575f));
  p = __ffma2_rn(p, f, make_float2(1.f, 1.f));
  return make_float2(__int_as_float(__float_as_int(p.x) + (__float_as_int(t.x) << 23)),
                     __int_as_float(__float_as_int(p.y) + (__float_as_int(t.y) << 23)));
}

#ifdef CVLLM_PREFILL_TRACE  // profiling build only (tools/trace_prefill.py): clock64 stamps of CTA (0,0,0), first 256 KV tiles
__device__ long long g_p2_trace[8][256];
#define P2_TRACE(rowi, jj)                                                                              \
  do {                                                                                                  \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (jj) < 256) g_p2_trace[rowi][jj] = clock64(); \
  } while (0)
#else
#define P2_TRACE(rowi, jj) do { } while (0)
#endif

template <bool BF16>
__global__ void __launch_bounds__(p2::kThreads, 1)
prefill_attn2_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                     const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_kc,
                     const __grid_constant__ CUtensorMap tm_vc, const Prefill2Params p) {
  using namespace p2;
  extern __shared__ unsigned char smem_raw[];
  Prefill2Smem& sm = *reinterpret_cast<Prefill2Smem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z;
  const int hq0 = 2 * blockIdx.y;  // heads hq0, hq0 + 1 share KV head hkv (G even)
  const int hkv = hq0 / p.G;
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
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&sm.s_full[t], 1);
      mbar_init(&sm.p_full[t], 128);
      mbar_init(&sm.p_half[t], 128);
      mbar_init(&sm.o_done[t], 1);
      mbar_init(&sm.o_final[t], 1);
    }
    mbar_fence_init();
  }
  if (warp == 9) {
    tmem_alloc(&sm.tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == 8 && lane == 0) {
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

  if (warp == 8) {
    // =========================== TMA producer ===========================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    if (lane == 0) {
      mbar_expect_tx(&sm.q_full, 2 * kTileBytes);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        tma_load_3d(sm.q[t], &tm_q, &sm.q_full, 0, hq0 + t, q_beg + q0);
        tma_load_3d(sm.q[t] + kBoxBytes, &tm_q, &sm.q_full, 64, hq0 + t, q_beg + q0);
      }
      const int32_t* pt = p.page_table + ((int64_t)p.batch_mapping[b] * p.HKV + hkv) * p.NLP;
      for (int n = 0; n < 2 * n_tiles; ++n) {
        const int j = n >> 1;
        const bool is_v = n & 1;
        const int slot = n % kSlots;
        const uint32_t ph = (n / kSlots) & 1;
        const bool pre = j < n_pre;
        int row;
        if (pre) {
          const int t0 = j * kBN;
          const int pg = t0 / p.PS;
          row = pt[pg] * p.PS + (t0 - pg * p.PS);
        } else {
          row = q_beg + (j - n_pre) * kBN;
        }
        mbar_wait(&sm.empty[slot], ph ^ 1);
#ifdef CVLLM_PROBE_NO_TMA  // timing probe only (wrong results): K / V tiles are never fetched
        mbar_arrive(&sm.full[slot]);
        continue;
#endif
        mbar_expect_tx(&sm.full[slot], kTileBytes);
        unsigned char* dst = sm.kv[slot];
        if (pre) {
          const CUtensorMap* m = is_v ? &tm_vc : &tm_kc;
          tma_load_2d(dst, m, &sm.full[slot], 0, row);
          tma_load_2d(dst + kBoxBytes, m, &sm.full[slot], 64, row);
        } else {
          const CUtensorMap* m = is_v ? &tm_v : &tm_k;
          tma_load_3d(dst, m, &sm.full[slot], 0, hkv, row);
          tma_load_3d(dst + kBoxBytes, m, &sm.full[slot], 64, hkv, row);
        }
      }
    }
  } else if (warp == 9) {
    // ============================ UMMA issuer ============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    constexpr uint32_t idesc_qk = umma_idesc(BF16, kBM, kBN, false, false);
    constexpr uint32_t idesc_pv = umma_idesc(BF16, kBM, kHD, false, true);

    auto slot_of = [](int n) { return n % kSlots; };
    auto phase_of = [](int n) { return (uint32_t)((n / kSlots) & 1); };

    // The issuing thread is on the critical path of every hand-off (tools/trace_prefill.py: 86 cycles per UMMA in the
    // loop against 64 on the pipe, 13 SASS instructions each).  Two things keep its per-UMMA work to a few uniform-
    // datapath instructions: the TMEM base is made provably warp-uniform (a shuffle from lane 0: otherwise every UMMA
    // is wrapped in an ELECT / R2UR / branch loop), and a descriptor is its tile's base descriptor plus a constant
    // (the 14-bit start-address field cannot carry: shared memory ends below 256 KB) instead of a shift and two masks.
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
    auto desc_at = [](uint64_t base, uint32_t byte_off) {
      return (base & 0xffffffff00000000ull) | (uint64_t)((uint32_t)base + (byte_off >> 4));
    };
    const uint64_t q_desc[2] = {umma_smem_desc(smem_u32(sm.q[0]), 16, 1024), umma_smem_desc(smem_u32(sm.q[1]), 16, 1024)};

    // S_t = Q_t . K_j^T  (K_j in ring entry 2j, already waited for)
    auto issue_qk = [&](int t, int j) {
      const uint64_t k_desc = umma_smem_desc(smem_u32(sm.kv[slot_of(2 * j)]), 16, 1024);
      const uint32_t s_tmem = tmem_u + t * kBN;
#pragma unroll
      for (int kk = 0; kk < kHD / 16; ++kk) {
        const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
        umma_ss_warp(s_tmem, desc_at(q_desc[t], off), desc_at(k_desc, off), idesc_qk, kk > 0 ? 1u : 0u);
      }
      umma_commit_warp(&sm.s_full[t]);
      if (t == 1) umma_commit_warp(&sm.empty[slot_of(2 * j)]);
    };
    auto issue_pv = [&](int t, int j) {
      const uint64_t v_desc = umma_smem_desc(smem_u32(sm.kv[slot_of(2 * j + 1)]), kBoxBytes, 1024);
      const uint32_t p_tmem = tmem_u + t * kBN;
      const uint32_t o_tmem = tmem_u + 256 + t * kHD;
#if CVLLM_PREFILL_P_SPLIT > 0
      mbar_wait(&sm.p_half[t], j & 1);
      P2_TRACE(2 * t, j);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < 2 * kPSplit; ++kk)
        umma_ts_warp(o_tmem, p_tmem + kk * 8, desc_at(v_desc, kk * 2048), idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
      mbar_wait(&sm.p_full[t], j & 1);
      tc_fence_after();
#pragma unroll
      for (int kk = 2 * kPSplit; kk < kBN / 16; ++kk)
        umma_ts_warp(o_tmem, p_tmem + kk * 8, desc_at(v_desc, kk * 2048), idesc_pv, 1u);
#else
      mbar_wait(&sm.p_full[t], j & 1);
      P2_TRACE(2 * t, j);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < kBN / 16; ++kk)
        umma_ts_warp(o_tmem, p_tmem + kk * 8, desc_at(v_desc, kk * 2048), idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
#endif
      if (t == 1) umma_commit_warp(&sm.empty[slot_of(2 * j + 1)]);
      umma_commit_warp(&sm.o_done[t]);
      if (j == n_tiles - 1) umma_commit_warp(&sm.o_final[t]);
    };
    auto wait_v = [&](int j) {
      const int n = 2 * j + 1;
      mbar_wait(&sm.full[slot_of(n)], phase_of(n));
      if (j == n_pre - 1 && (Lc % kBN) != 0) {
        // ragged last prefix tile: rows >= Lc % 128 of the page are uninitialised memory; P is
        // exactly 0 there but 0 * NaN would poison O, so zero those V rows.
        const int valid = Lc % kBN;
        uint4* v4 = reinterpret_cast<uint4*>(sm.kv[slot_of(n)]);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int idx = lane; idx < (kBN - valid) * 16; idx += 32) {
          const int r = valid + idx / 16, c = idx % 16;  // 16 x 16B per row across the 2 boxes
          v4[(c >> 3) * (kBoxBytes / 16) + r * 8 + (c & 7)] = z;
        }
        fence_proxy_async_smem();
        __syncwarp();
      }
      tc_fence_after();
    };
    auto wait_k = [&](int j) {
      const int n = 2 * j;
      mbar_wait(&sm.full[slot_of(n)], phase_of(n));
      tc_fence_after();
    };

    mbar_wait(&sm.q_full, 0);
    wait_k(0);
    issue_qk(0, 0);
    issue_qk(1, 0);
    for (int j = 0; j < n_tiles; ++j) {
      const bool more = j + 1 < n_tiles;
      wait_v(j);
      issue_pv(0, j);
      if (more) {
        wait_k(j + 1);
        issue_qk(0, j + 1);
      }
      P2_TRACE(1, j);
      issue_pv(1, j);
      if (more) issue_qk(1, j + 1);
      P2_TRACE(3, j);
    }
  } else if (warp < 8) {
    // ====================== softmax / correction / epilogue ======================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmax));
    const int t = warp >> 2;        // which head of the pair
    const int quad = warp & 3;      // TMEM lane quadrant this warp may touch
    const int row = quad * 32 + lane;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const bool row_valid = (q0 + row) < La;
    const uint32_t s_tmem = tmem + t * kBN + lane_base;
    const uint32_t o_tmem = tmem + 256 + t * kHD + lane_base;
    float m_ref = -INFINITY, l = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
      mbar_wait(&sm.s_full[t], j & 1);
      if (quad == 0) P2_TRACE(4 + 2 * t, j);
      tc_fence_after();
#ifdef CVLLM_PROBE_NO_SOFTMAX  // timing probe only (wrong results): the issuer / TMA side alone — 6.4 ms at 32K
      tc_fence_before();
      mbar_arrive(&sm.p_half[t]);
      mbar_arrive(&sm.p_full[t]);
      if (quad == 0) P2_TRACE(5 + 2 * t, j);
      continue;
#endif
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

      auto row_max = [&]() -> float {
        float mx0 = s[0], mx1 = s[1], mx2 = s[2], mx3 = s[3];
#pragma unroll
        for (int c = 4; c < kBN; c += 4) {
          mx0 = fmaxf(mx0, s[c]);
          mx1 = fmaxf(mx1, s[c + 1]);
          mx2 = fmaxf(mx2, s[c + 2]);
          mx3 = fmaxf(mx3, s[c + 3]);
        }
        return fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;
      };
      // O *= alpha, l *= alpha, m_ref <- m_to.  O holds PV(0..j-1); PV(j-1) may still be in flight.
      auto rescale_to = [&](float m_to) {
        const float alpha = ex2(m_ref - m_to);  // m_ref == -inf on the first tile -> 0
        m_ref = m_to;
        l *= alpha;
        if (j > 0) {
          mbar_wait(&sm.o_done[t], (j - 1) & 1);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < kHD / 32; ++c) {
            uint32_t o[32];
            tmem_ld32(o_tmem + c * 32, o);
            tc_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(o_tmem + c * 32, o);
          }
        }
      };
      // P = exp2(S*scale - m_ref) packed to 16 bit over the S columns; returns the row sum of this tile.
      // Packed fp32 pairs: one FFMA2 scales and shifts two logits, one FADD2 adds two probabilities (A/B on B200,
      // r02d: 7.62-7.68 ms vs 7.95-8.05 ms scalar at 32K).
      auto write_p = [&]() -> float {
        const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
        const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), nm2 = make_float2(neg_m, neg_m);
        float2 l2a = make_float2(0.f, 0.f), l2b = make_float2(0.f, 0.f);
#pragma unroll
        for (int c = 0; c < kBN / 32; ++c) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; i += 2) {
            const float2 t0 = __ffma2_rn(make_float2(s[c * 32 + 2 * i], s[c * 32 + 2 * i + 1]), sc2, nm2);
            const float2 t1 = __ffma2_rn(make_float2(s[c * 32 + 2 * i + 2], s[c * 32 + 2 * i + 3]), sc2, nm2);
#ifdef CVLLM_PROBE_NO_EXP  // timing probe only (wrong results): the loop without its MUFU work — 7.6 ms, unchanged
            const float2 e0 = t0, e1 = t1;
#else
            // kPolyPairs = e: e of every 8 logit pairs (two loop trips, four pairs each... slots b0 b2 b1 b3 a1 a3 a0 a2)
            const int slot = (i >> 1) & 3;
            const bool p1 = (kPolyPairs >= 1 && slot == 0) || (kPolyPairs >= 2 && slot == 2) ||
                            (kPolyPairs >= 3 && slot == 1) || (kPolyPairs >= 4 && slot == 3);
            const bool p0 = (kPolyPairs >= 5 && slot == 1) || (kPolyPairs >= 6 && slot == 3) ||
                            (kPolyPairs >= 7 && slot == 0) || (kPolyPairs >= 8 && slot == 2);
            const float2 e0 = p0 ? ex2_poly2(t0) : make_float2(ex2(t0.x), ex2(t0.y));
            const float2 e1 = p1 ? ex2_poly2(t1) : make_float2(ex2(t1.x), ex2(t1.y));
#endif
            l2a = __fadd2_rn(l2a, e0);
            l2b = __fadd2_rn(l2b, e1);
            pk[i] = pack2<BF16>(e0.x, e0.y);
            pk[i + 1] = pack2<BF16>(e1.x, e1.y);
          }
          tmem_st16(s_tmem + c * 16, pk);
#if CVLLM_PREFILL_P_SPLIT > 0
          if (c == kPSplit - 1) {  // the first part of P (and any O rescale before it) is in TMEM
            tc_wait_st();
            tc_fence_before();
            mbar_arrive(&sm.p_half[t]);
          }
#endif
        }
        return (l2a.x + l2b.x) + (l2a.y + l2b.y);
      };

      // The row maximum of the tile is known before the first exponential.  Taking it off the critical path
      // (exponentials against the previous tiles' reference, rescale deferred by one tile) was measured on B200
      // and changed nothing: 7.70-7.91 ms vs 7.60-7.80 ms at 32K (r02i) — the FMNMX3 chain already hides under the
      // TMEM load latency.
      {
        const float m_new = fmaxf(m_ref, row_max());
        const bool need = (j == 0) || (m_new > m_ref + p.rescale_threshold);
        if (__any_sync(0xffffffffu, need)) rescale_to(m_new);
        l += write_p();
      }
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(&sm.p_full[t]);
      if (quad == 0) P2_TRACE(5 + 2 * t, j);
    }

    // ---- epilogue: O / l -> out ----
    mbar_wait(&sm.o_final[t], 0);
    tc_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    char* orow = reinterpret_cast<char*>(p.out) +
                 ((int64_t)(q_beg + q0 + row) * p.o_stride_n + (int64_t)(hq0 + t) * kHD) * 2;
#pragma unroll
    for (int c = 0; c < kHD / 32; ++c) {
      uint32_t o[32];
      tmem_ld32(o_tmem + c * 32, o);
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

  else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));  // warps 10, 11
  }

  // ---------------- teardown ----------------
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, 512);
}

}  // namespace cvllm

using namespace cvllm;

// Launch the paired-head kernel.  Called by cvllm_prefill_attention (prefill_attn.cu) after the
// argument checks and tensor-map construction; returns a CVLLM_* code.
int cvllm_launch_prefill2(const CUtensorMap& tm_q, const CUtensorMap& tm_k, const CUtensorMap& tm_v,
                          const CUtensorMap& tm_kc, const CUtensorMap& tm_vc, void* out,
                          const int32_t* seq_lens_bh, const int32_t* page_table,
                          const int32_t* batch_mapping, const int32_t* cu_seqlens_q, int B, int HQ,
                          int HKV, int PS, int NLP, int64_t o_stride_n, int max_seqlen_q,
                          float scale_log2, float rescale_threshold, bool bf16, cudaStream_t st) {
  Prefill2Params p;
  p.out = out; p.seq_lens_bh = seq_lens_bh; p.page_table = page_table;
  p.batch_mapping = batch_mapping; p.cu_seqlens_q = cu_seqlens_q;
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = HQ / HKV; p.PS = PS; p.NLP = NLP;
  p.o_stride_n = o_stride_n;
  p.scale_log2 = scale_log2;
  p.rescale_threshold = rescale_threshold;
  const size_t smem = sizeof(Prefill2Smem) + 1024;
  const dim3 grid((max_seqlen_q + p2::kBM - 1) / p2::kBM, HQ / 2, B);
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn2_kernel<true>, (int)smem, attr_bf));
  else CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn2_kernel<false>, (int)smem, attr_h));
  if (bf16) prefill_attn2_kernel<true><<<grid, p2::kThreads, smem, st>>>(tm_q, tm_k, tm_v, tm_kc, tm_vc, p);
  else prefill_attn2_kernel<false><<<grid, p2::kThreads, smem, st>>>(tm_q, tm_k, tm_v, tm_kc, tm_vc, p);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

#ifdef CVLLM_PREFILL_TRACE
// profiling build only: copies the 8 x 256 clock64 stamps of the last launch's CTA (0,0,0) to host memory
extern "C" int cvllm_debug_prefill_trace_read(void* host_out) {
  CVLLM_CHECK_CUDA(cudaDeviceSynchronize());
  CVLLM_CHECK_CUDA(cudaMemcpyFromSymbol(host_out, cvllm::g_p2_trace, sizeof(long long) * 8 * 256));
  return CVLLM_OK;
}
#endif
