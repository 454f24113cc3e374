// Causal head-sparse varlen prefill attention, FOUR S streams per CTA (64-key half tiles).
//
// prefill_attn2.cu keeps two streams in flight per CTA (two GQA heads, one 128-key S tile each).  Its timing
// probes show the kernel is bound neither by the MUFU (exponentials removed: same time) nor by the K / V
// feed (TMA removed: same time) nor by the UMMA issue rate (the kernel's instruction mix runs at the nominal 64
// cycles per 128x128x16 when nothing waits, tools/ubench/umma_rate.cu): with the softmax skipped AND the loads
// skipped it still takes 6.2 of 7.4 ms.  What is left is the round trip of every hand-off — UMMA completion ->
// tcgen05.commit -> mbarrier -> softmax warps -> mbarrier -> issuer -> UMMA start, ~1.2 us — of which only the other
// head's two UMMAs (0.5 us at the nominal rate) are there to hide it.
// Here every head has TWO S buffers of 64 keys (a K tile is consumed as two N = 64 UMMAs, a V tile as two K = 64
// chains), so four (head, half) streams alternate: while the softmax group of head t exponentiates half a of a tile,
// the tensor pipe already holds S of half b and the P -> PV hand-off of half a overlaps the softmax of half b.
// Same FLOPs, same TMEM budget (4 x 64 S columns + 2 x 128 O columns), twice as many, half as long hand-offs.
//
// TMEM: S[t][h] at columns t*128 + h*64 (fp32, 64 columns; P[t][h] packed to 16 bit over its first 32 columns),
// O0 [256,384), O1 [384,512).  Issue order (u = 2*tile + half):
//   S(0,0) S(0,1) S(1,0) S(1,1) | PV(0,u) S(0,u+2) PV(1,u) S(1,u+2) | ...
// Everything else (TMA ring of 32 KB K / V tiles, paged prefix, ragged last prefix tile, lazy rescale, packed
// FFMA2 / FADD2 softmax math, epilogue) is prefill_attn2.cu's.
#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

namespace p4 {
constexpr int kBM = 128, kBN = 128, kHD = 128;
constexpr int kSlots = 5;
constexpr int kTileBytes = kBN * kHD * 2;  // 32 KB
constexpr int kBoxBytes = kTileBytes / 2;  // one [128 x 64] swizzled box
constexpr int kThreads = 384;              // 3 warpgroups
constexpr int kRegsSoftmax = 224, kRegsOther = 56;  // 2*128*224 + 128*56 = 64512 <= 65536
}  // namespace p4

struct Prefill4Params {
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

struct Prefill4Smem {
  unsigned char q[2][p4::kTileBytes];
  unsigned char kv[p4::kSlots][p4::kTileBytes];
  uint64_t q_full;
  uint64_t full[p4::kSlots], empty[p4::kSlots];
  uint64_t s_full[2][2], p_full[2][2];  // [head][half]
  uint64_t o_done[2];   // one phase per PV_t(i)
  uint64_t o_final[2];  // last PV_t complete
  uint32_t tmem_base;
};

template <bool BF16>
__global__ void __launch_bounds__(p4::kThreads, 1)
prefill_attn4_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                     const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_kc,
                     const __grid_constant__ CUtensorMap tm_vc, const Prefill4Params p) {
  using namespace p4;
  extern __shared__ unsigned char smem_raw[];
  Prefill4Smem& sm = *reinterpret_cast<Prefill4Smem*>(
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
      for (int h = 0; h < 2; ++h) {
        mbar_init(&sm.s_full[t][h], 1);
        mbar_init(&sm.p_full[t][h], 128);
      }
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
    constexpr int kHN = kBN / 2;  // 64 keys per half tile
    constexpr uint32_t idesc_qk = umma_idesc(BF16, kBM, kHN, false, false);
    constexpr uint32_t idesc_pv = umma_idesc(BF16, kBM, kHD, false, true);

    auto slot_of = [](int n) { return n % kSlots; };
    auto phase_of = [](int n) { return (uint32_t)((n / kSlots) & 1); };
    const int n_half = 2 * n_tiles;

    // S(t, u) = Q_t . K[keys of half u]^T  (K tile in ring entry 2*(u/2), already waited for)
    auto issue_qk = [&](int t, int u) {
      if (lane == 0) {
        const int j = u >> 1, h = u & 1;
        const uint32_t q_addr = smem_u32(sm.q[t]);
        const uint32_t k_addr = smem_u32(sm.kv[slot_of(2 * j)]) + h * (kHN * 128);  // rows 64h.. of each [128 x 64] box
        const uint32_t s_tmem = tmem + t * kBN + h * kHN;
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk) {
          const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
          umma_ss(s_tmem, umma_smem_desc(q_addr + off, 16, 1024), umma_smem_desc(k_addr + off, 16, 1024),
                  idesc_qk, kk > 0 ? 1u : 0u);
        }
        umma_commit(&sm.s_full[t][h]);
        if (t == 1 && h == 1) umma_commit(&sm.empty[slot_of(2 * j)]);  // K_j consumed by both heads, both halves
      }
      __syncwarp();
    };
    // O_t += P(t, u) . V[keys of half u]  (V tile in ring entry 2*(u/2) + 1)
    auto issue_pv = [&](int t, int u) {
      const int j = u >> 1, h = u & 1;
      mbar_wait(&sm.p_full[t][h], j & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t v_addr = smem_u32(sm.kv[slot_of(2 * j + 1)]) + h * (kHN * 128);  // key rows 64h..
        const uint32_t p_tmem = tmem + t * kBN + h * kHN;
        const uint32_t o_tmem = tmem + 256 + t * kHD;
#pragma unroll
        for (int kk = 0; kk < kHN / 16; ++kk) {
          const uint64_t vdesc = umma_smem_desc(v_addr + kk * 2048, kBoxBytes, 1024);
          umma_ts(o_tmem, p_tmem + kk * 8, vdesc, idesc_pv, (u > 0 || kk > 0) ? 1u : 0u);
        }
        if (t == 1 && h == 1) umma_commit(&sm.empty[slot_of(2 * j + 1)]);  // V_j consumed
        umma_commit(&sm.o_done[t]);
        if (u == n_half - 1) umma_commit(&sm.o_final[t]);
      }
      __syncwarp();
    };
    auto wait_v = [&](int j) {
      const int n = 2 * j + 1;
      mbar_wait(&sm.full[slot_of(n)], phase_of(n));
      if (j == n_pre - 1 && (Lc % kBN) != 0) {
        // ragged last prefix tile: rows >= Lc % 128 of the page are uninitialised memory; P is exactly 0 there but
        // 0 * NaN would poison O, so zero those V rows.
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
    issue_qk(0, 1);
    issue_qk(1, 0);
    issue_qk(1, 1);
    for (int u = 0; u < n_half; ++u) {
      const bool more = u + 2 < n_half;
      if ((u & 1) == 0) wait_v(u >> 1);
      issue_pv(0, u);
      if (more) {
        if ((u & 1) == 0) wait_k((u >> 1) + 1);
        issue_qk(0, u + 2);
      }
      issue_pv(1, u);
      if (more) issue_qk(1, u + 2);
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

    constexpr int kHN = kBN / 2;  // 64 keys per half tile
    const int n_half = 2 * n_tiles;
    for (int u = 0; u < n_half; ++u) {
      const int j = u >> 1, h = u & 1;
      const uint32_t sh_tmem = s_tmem + h * kHN;
      mbar_wait(&sm.s_full[t][h], j & 1);
      tc_fence_after();
#ifdef CVLLM_PROBE_NO_SOFTMAX  // timing probe only (wrong results): the issuer / TMA side alone
      tc_fence_before();
      mbar_arrive(&sm.p_full[t][h]);
      continue;
#endif
      float s[kHN];
#pragma unroll
      for (int c = 0; c < kHN / 32; ++c) {
        uint32_t tmp[32];
        tmem_ld32(sh_tmem + c * 32, tmp);
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(tmp[i]);
      }
      tc_wait_ld();

      // ---- masks (boundary tiles only); key index inside the tile = 64 * h + c ----
      if (j < n_pre) {
        const int valid = Lc - j * kBN - h * kHN;
        if (valid < kHN) {
#pragma unroll
          for (int c = 0; c < kHN; ++c)
            if (c >= valid) s[c] = -INFINITY;
        }
      } else if (j == n_tiles - 1) {
        const int lim = row - h * kHN;
#pragma unroll
        for (int c = 0; c < kHN; ++c)
          if (c > lim) s[c] = -INFINITY;
      }

      float mx0 = s[0], mx1 = s[1], mx2 = s[2], mx3 = s[3];
#pragma unroll
      for (int c = 4; c < kHN; c += 4) {
        mx0 = fmaxf(mx0, s[c]);
        mx1 = fmaxf(mx1, s[c + 1]);
        mx2 = fmaxf(mx2, s[c + 2]);
        mx3 = fmaxf(mx3, s[c + 3]);
      }
      const float m_new = fmaxf(m_ref, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2);
      // u == 0 always holds a live key for every row (key 0 of the sequence or of the prefix), so m_ref is finite
      // from the first half tile on; later halves may be fully masked (m_new == m_ref, P = 0)
      const bool need = (u == 0) || (m_new > m_ref + p.rescale_threshold);
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = ex2(m_ref - m_new);  // m_ref == -inf on the first half tile -> 0
        m_ref = m_new;
        l *= alpha;
        if (u > 0) {
          // O holds PV(0..u-1); PV(u-1) may still be in flight
          mbar_wait(&sm.o_done[t], (u - 1) & 1);
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
      }

      // ---- P = exp2(S*scale - m_ref), row sum, 16-bit pack, back to TMEM over the half's S columns ----
      const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
      const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), nm2 = make_float2(neg_m, neg_m);
      float2 l2a = make_float2(0.f, 0.f), l2b = make_float2(0.f, 0.f);
#pragma unroll
      for (int c = 0; c < kHN / 32; ++c) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const float2 t0 = __ffma2_rn(make_float2(s[c * 32 + 2 * i], s[c * 32 + 2 * i + 1]), sc2, nm2);
          const float2 t1 = __ffma2_rn(make_float2(s[c * 32 + 2 * i + 2], s[c * 32 + 2 * i + 3]), sc2, nm2);
          const float2 e0 = make_float2(ex2(t0.x), ex2(t0.y)), e1 = make_float2(ex2(t1.x), ex2(t1.y));
          l2a = __fadd2_rn(l2a, e0);
          l2b = __fadd2_rn(l2b, e1);
          pk[i] = pack2<BF16>(e0.x, e0.y);
          pk[i + 1] = pack2<BF16>(e1.x, e1.y);
        }
        tmem_st16(sh_tmem + c * 16, pk);
      }
      l += (l2a.x + l2b.x) + (l2a.y + l2b.y);
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(&sm.p_full[t][h]);
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
int cvllm_launch_prefill4(const CUtensorMap& tm_q, const CUtensorMap& tm_k, const CUtensorMap& tm_v,
                          const CUtensorMap& tm_kc, const CUtensorMap& tm_vc, void* out,
                          const int32_t* seq_lens_bh, const int32_t* page_table,
                          const int32_t* batch_mapping, const int32_t* cu_seqlens_q, int B, int HQ,
                          int HKV, int PS, int NLP, int64_t o_stride_n, int max_seqlen_q,
                          float scale_log2, float rescale_threshold, bool bf16, cudaStream_t st) {
  Prefill4Params p;
  p.out = out; p.seq_lens_bh = seq_lens_bh; p.page_table = page_table;
  p.batch_mapping = batch_mapping; p.cu_seqlens_q = cu_seqlens_q;
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = HQ / HKV; p.PS = PS; p.NLP = NLP;
  p.o_stride_n = o_stride_n;
  p.scale_log2 = scale_log2;
  p.rescale_threshold = rescale_threshold;
  const size_t smem = sizeof(Prefill4Smem) + 1024;
  const dim3 grid((max_seqlen_q + p4::kBM - 1) / p4::kBM, HQ / 2, B);
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn4_kernel<true>, (int)smem, attr_bf));
  else CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn4_kernel<false>, (int)smem, attr_h));
  if (bf16) prefill_attn4_kernel<true><<<grid, p4::kThreads, smem, st>>>(tm_q, tm_k, tm_v, tm_kc, tm_vc, p);
  else prefill_attn4_kernel<false><<<grid, p4::kThreads, smem, st>>>(tm_q, tm_k, tm_v, tm_kc, tm_vc, p);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
