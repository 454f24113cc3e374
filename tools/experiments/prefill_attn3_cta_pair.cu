// Causal head-sparse varlen prefill attention on CTA PAIRS (tcgen05.mma.cta_group::2).
//
// prefill_attn2.cu (one CTA = 128 query rows x two GQA heads, cta_group::1) spends ~85 % of its time on the issuer /
// TMA side: with its softmax skipped altogether it still takes 6.4 of 7.5 ms at 32K.  Every UMMA there
// reads its whole B operand (the K or V tile) out of the CTA's own shared memory, and every CTA fills its own copy
// of every K / V tile.  Here two CTAs of one TPC (a cluster of 2 along the query-tile axis) work on two consecutive
// 128-row query tiles of the same head pair and SHARE each K / V tile: each CTA loads HALF of it (K: 64 of the
// 128 keys, V: 64 of the 128 head dims — the halves `tcgen05.mma.cta_group::2` reads from the peer's shared
// memory), so B-operand shared-memory reads and TMA fills per FLOP are halved.  M = 256 per instruction; each CTA
// keeps its own 128 rows of S / P / O in its own TMEM.  Bring-up of the instruction forms: tools/ubench/umma_2cta.cu.
//
// Per CTA (384 threads):
//   warp 8   TMA producer: own Q tiles, own halves of K_j / V_j through a ring of 16 KB slots (local mbarriers)
//   warp 10  forwarder: waits for the local half to land (zeroes the rows of a ragged last prefix tile), then
//            arrives on the LEADER's pair barrier for that slot (count 2: one arrival per CTA)
//   warp 9   UMMA issuer — in the leader CTA (cluster rank 0) only: S_t = Q_t.K^T (SS), O_t += P_t.V (TS) for both
//            CTAs at once; commits are multicast to both CTAs' barriers
//   warps 0-7  softmax of head 0 / 1 exactly as in prefill_attn2.cu; "P ready" is one arrival per warp on the
//            leader's barrier (remote for the peer CTA, release at cluster scope)
// Causality: the pair runs the tile count of its UPPER query tile; the lower tile sees one extra KV tile beyond
// its diagonal, fully masked.  A sequence with an odd number of query tiles gets a ghost upper CTA (rows invalid).
#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

namespace p3 {
constexpr int kBM = 128, kBN = 128, kHD = 128;
constexpr int kSlots = 8;                  // ring of half tiles
constexpr int kHalfBytes = kBN * kHD;      // 16 KB: K half = 64 keys x 128 d, V half = 128 keys x 64 d
constexpr int kQBytes = kBM * kHD * 2;     // 32 KB per head
constexpr int kQBox = kQBytes / 2;         // [128 x 64] box
constexpr int kKBox = kHalfBytes / 2;      // [64 x 64] box of a K half
constexpr int kThreads = 384;
constexpr int kRegsSoftmax = 224, kRegsOther = 56;
}  // namespace p3

struct Prefill3Params {
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

struct Prefill3Smem {
  unsigned char q[2][p3::kQBytes];
  unsigned char kv[p3::kSlots][p3::kHalfBytes];
  uint64_t q_local, q_pair;                              // own Q landed / both CTAs' Q landed (leader's copy is used)
  uint64_t full[p3::kSlots], empty[p3::kSlots];          // local: own half landed / slot consumed (multicast commit)
  uint64_t pair_full[p3::kSlots];                        // leader's copy: both halves landed (2 arrivals)
  uint64_t s_full[2], p_full[2];                         // p_full: leader's copy, 8 warp arrivals (4 per CTA)
  uint64_t o_done[2], o_final[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(rank));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
// wait with acquire at cluster scope (the arrivals may come from the peer CTA)
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return;
    if (++spins > CVLLM_MBAR_SPIN_LIMIT) __trap();
  }
}
__device__ __forceinline__ void umma2_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc),
               "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma2_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc),
               "r"(idesc), "r"(acc) : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs once every UMMA issued so far has completed
__device__ __forceinline__ void umma2_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)), "h"((unsigned short)3) : "memory");
}

template <bool BF16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(p3::kThreads, 1)
prefill_attn3_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k64,
                     const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_kc64,
                     const __grid_constant__ CUtensorMap tm_vc, const Prefill3Params p) {
  using namespace p3;
  extern __shared__ unsigned char smem_raw[];
  Prefill3Smem& sm = *reinterpret_cast<Prefill3Smem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int b = blockIdx.z;
  const int hq0 = 2 * blockIdx.y;  // heads hq0, hq0 + 1 share KV head hkv (G even)
  const int hkv = hq0 / p.G;
  const int q_beg = p.cu_seqlens_q[b];
  const int La = p.cu_seqlens_q[b + 1] - q_beg;
  const int n_qt = (La + kBM - 1) / kBM;
  const int n_pairs = (n_qt + 1) / 2;
  const int pc = (int)(blockIdx.x >> 1);
  if (pc >= n_pairs) return;  // both CTAs of the cluster
  const int pr = n_pairs - 1 - pc;  // heaviest (latest) pairs first
  const int qt = 2 * pr + (int)rank;  // this CTA's query tile; == n_qt for the ghost upper CTA of an odd sequence
  const int qt_max = min(2 * pr + 1, n_qt - 1);
  const int q0 = qt * kBM;
  const int Lc = p.seq_lens_bh[b * p.HKV + hkv];
  const int n_pre = (Lc + kBN - 1) / kBN;
  const int n_tiles = n_pre + qt_max + 1;  // the pair's tile count
  const int jd = n_pre + qt;               // this CTA's diagonal tile; tiles beyond it are fully masked

  // ---------------- setup ----------------
  if (threadIdx.x == 0) {
    mbar_init(&sm.q_local, 1);
    mbar_init(&sm.q_pair, 2);
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
      mbar_init(&sm.pair_full[s], 2);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&sm.s_full[t], 1);
      mbar_init(&sm.p_full[t], 8);
      mbar_init(&sm.o_done[t], 1);
      mbar_init(&sm.o_final[t], 1);
    }
    mbar_fence_init();
  }
  if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                 "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k64);
    tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_kc64);
    tma_prefetch_desc(&tm_vc);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers exist before anyone arrives remotely
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 8) {
    // =========================== TMA producer (own Q, own halves) ===========================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    if (lane == 0) {
      mbar_expect_tx(&sm.q_local, 2 * kQBytes);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        tma_load_3d(sm.q[t], &tm_q, &sm.q_local, 0, hq0 + t, q_beg + q0);
        tma_load_3d(sm.q[t] + kQBox, &tm_q, &sm.q_local, 64, hq0 + t, q_beg + q0);
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
#ifdef CVLLM_PROBE_NO_TMA  // timing probe only (wrong results): K / V halves are never fetched
        mbar_arrive(&sm.full[slot]);
        continue;
#endif
        mbar_expect_tx(&sm.full[slot], kHalfBytes);
        unsigned char* dst = sm.kv[slot];
        if (!is_v) {  // K half: keys 64*rank .. 64*rank+63, all 128 d (two [64 x 64] boxes)
          const int r0 = row + 64 * (int)rank;
          if (pre) {
            tma_load_2d(dst, &tm_kc64, &sm.full[slot], 0, r0);
            tma_load_2d(dst + kKBox, &tm_kc64, &sm.full[slot], 64, r0);
          } else {
            tma_load_3d(dst, &tm_k64, &sm.full[slot], 0, hkv, r0);
            tma_load_3d(dst + kKBox, &tm_k64, &sm.full[slot], 64, hkv, r0);
          }
        } else {  // V half: all 128 keys, d 64*rank .. 64*rank+63 (one [128 x 64] box)
          if (pre) tma_load_2d(dst, &tm_vc, &sm.full[slot], 64 * (int)rank, row);
          else tma_load_3d(dst, &tm_v, &sm.full[slot], 64 * (int)rank, hkv, row);
        }
      }
    }
  } else if (warp == 10) {
    // =========================== forwarder: local landing -> the leader's pair barrier ===========================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    mbar_wait(&sm.q_local, 0);
    if (lane == 0) mbar_arrive_cluster(&sm.q_pair, 0);
    for (int n = 0; n < 2 * n_tiles; ++n) {
      const int j = n >> 1;
      const int slot = n % kSlots;
      mbar_wait(&sm.full[slot], (n / kSlots) & 1);
      if ((n & 1) && j == n_pre - 1 && (Lc % kBN) != 0) {
        // ragged last prefix tile: rows >= Lc % 128 of the page are uninitialised memory; P is exactly 0 there but
        // 0 * NaN would poison O, so zero those rows of my V half (128 B per key row)
        const int valid = Lc % kBN;
        uint4* v4 = reinterpret_cast<uint4*>(sm.kv[slot]);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int idx = lane; idx < (kBN - valid) * 8; idx += 32) v4[(valid + idx / 8) * 8 + (idx & 7)] = z;
        fence_proxy_async_smem();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&sm.pair_full[slot], 0);
    }
  } else if (warp == 9) {
    // ============================ UMMA issuer (leader CTA only) ============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    if (leader) {
      constexpr uint32_t idesc_qk = umma_idesc(BF16, 2 * kBM, kBN, false, false);
      constexpr uint32_t idesc_pv = umma_idesc(BF16, 2 * kBM, kHD, false, true);
      auto slot_of = [](int n) { return n % kSlots; };
      auto phase_of = [](int n) { return (uint32_t)((n / kSlots) & 1); };
      // S_t = Q_t . K_j^T  (K_j halves in ring entry 2j of both CTAs, already waited for)
      auto issue_qk = [&](int t, int j) {
        if (lane == 0) {
          const uint32_t q_addr = smem_u32(sm.q[t]);
          const uint32_t k_addr = smem_u32(sm.kv[slot_of(2 * j)]);
          const uint32_t s_tmem = tmem + t * kBN;
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk) {
            const uint32_t offq = (kk >> 2) * kQBox + (kk & 3) * 32, offk = (kk >> 2) * kKBox + (kk & 3) * 32;
            umma2_ss(s_tmem, umma_smem_desc(q_addr + offq, 16, 1024), umma_smem_desc(k_addr + offk, 16, 1024), idesc_qk,
                     kk > 0 ? 1u : 0u);
          }
          umma2_commit(&sm.s_full[t]);
          if (t == 1) umma2_commit(&sm.empty[slot_of(2 * j)]);  // K_j consumed by both heads
        }
        __syncwarp();
      };
      // O_t += P_t . V_j  (V_j halves in ring entry 2j + 1)
      auto issue_pv = [&](int t, int j) {
        mbar_wait_cluster(&sm.p_full[t], j & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t v_addr = smem_u32(sm.kv[slot_of(2 * j + 1)]);
          const uint32_t p_tmem = tmem + t * kBN;
          const uint32_t o_tmem = tmem + 256 + t * kHD;
#pragma unroll
          for (int kk = 0; kk < kBN / 16; ++kk)
            umma2_ts(o_tmem, p_tmem + kk * 8, umma_smem_desc(v_addr + kk * 2048, 16, 1024), idesc_pv,
                     (j > 0 || kk > 0) ? 1u : 0u);
          if (t == 1) umma2_commit(&sm.empty[slot_of(2 * j + 1)]);  // V_j consumed by both heads
          umma2_commit(&sm.o_done[t]);
          if (j == n_tiles - 1) umma2_commit(&sm.o_final[t]);
        }
        __syncwarp();
      };
      auto wait_entry = [&](int n) {
        mbar_wait_cluster(&sm.pair_full[slot_of(n)], phase_of(n));
        tc_fence_after();
      };

      mbar_wait_cluster(&sm.q_pair, 0);
      wait_entry(0);
      issue_qk(0, 0);
      issue_qk(1, 0);
      for (int j = 0; j < n_tiles; ++j) {
        const bool more = j + 1 < n_tiles;
        wait_entry(2 * j + 1);
        issue_pv(0, j);
        if (more) {
          wait_entry(2 * j + 2);
          issue_qk(0, j + 1);
        }
        issue_pv(1, j);
        if (more) issue_qk(1, j + 1);
      }
    }
  } else if (warp < 8) {
    // ====================== softmax / correction / epilogue (own 128 rows) ======================
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
      tc_fence_after();
#ifdef CVLLM_PROBE_NO_SOFTMAX  // timing probe only (wrong results): the issuer / TMA side alone
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&sm.p_full[t], 0);
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
      } else if (j >= jd) {
        const int lim = (j == jd) ? row : -1;  // beyond my diagonal: every key of the tile is in my future
#pragma unroll
        for (int c = 0; c < kBN; ++c)
          if (c > lim) s[c] = -INFINITY;
      }

      float mx0 = s[0], mx1 = s[1], mx2 = s[2], mx3 = s[3];
#pragma unroll
      for (int c = 4; c < kBN; c += 4) {
        mx0 = fmaxf(mx0, s[c]);
        mx1 = fmaxf(mx1, s[c + 1]);
        mx2 = fmaxf(mx2, s[c + 2]);
        mx3 = fmaxf(mx3, s[c + 3]);
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      const float m_new = fmaxf(m_ref, mx * p.scale_log2);
      const bool need = (j == 0) || (m_new > m_ref + p.rescale_threshold);
      if (__any_sync(0xffffffffu, need)) {
        const float alpha = ex2(m_ref - m_new);  // m_ref == -inf on the first tile -> 0
        m_ref = m_new;
        l *= alpha;
        if (j > 0) {
          // O holds PV(0..j-1); PV(j-1) may still be in flight
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
      }

      // ---- P = exp2(S*scale - m_ref), row sum, 16-bit pack, back to TMEM over S ----
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
          const float2 e0 = make_float2(ex2(t0.x), ex2(t0.y)), e1 = make_float2(ex2(t1.x), ex2(t1.y));
          l2a = __fadd2_rn(l2a, e0);
          l2b = __fadd2_rn(l2b, e1);
          pk[i] = pack2<BF16>(e0.x, e0.y);
          pk[i + 1] = pack2<BF16>(e1.x, e1.y);
        }
        tmem_st16(s_tmem + c * 16, pk);
      }
      l += (l2a.x + l2b.x) + (l2a.y + l2b.y);
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&sm.p_full[t], 0);  // one arrival per warp on the leader's barrier
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
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));  // warp 11
  }

  // ---------------- teardown: nobody leaves while the peer may still touch this CTA's smem / TMEM ----------------
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 9)
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

}  // namespace cvllm

using namespace cvllm;

// Launch the CTA-pair kernel.  tm_k64 / tm_kc64: K tensor maps with 64-row boxes; tm_v / tm_vc: the 128-row V maps.
int cvllm_launch_prefill3(const CUtensorMap& tm_q, const CUtensorMap& tm_k64, const CUtensorMap& tm_v,
                          const CUtensorMap& tm_kc64, const CUtensorMap& tm_vc, void* out,
                          const int32_t* seq_lens_bh, const int32_t* page_table,
                          const int32_t* batch_mapping, const int32_t* cu_seqlens_q, int B, int HQ,
                          int HKV, int PS, int NLP, int64_t o_stride_n, int max_seqlen_q,
                          float scale_log2, float rescale_threshold, bool bf16, cudaStream_t st) {
  Prefill3Params p;
  p.out = out; p.seq_lens_bh = seq_lens_bh; p.page_table = page_table;
  p.batch_mapping = batch_mapping; p.cu_seqlens_q = cu_seqlens_q;
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = HQ / HKV; p.PS = PS; p.NLP = NLP;
  p.o_stride_n = o_stride_n;
  p.scale_log2 = scale_log2;
  p.rescale_threshold = rescale_threshold;
  const size_t smem = sizeof(Prefill3Smem) + 1024;
  const int n_qt = (max_seqlen_q + p3::kBM - 1) / p3::kBM;
  const dim3 grid(2 * ((n_qt + 1) / 2), HQ / 2, B);
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn3_kernel<true>, (int)smem, attr_bf));
  else CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)prefill_attn3_kernel<false>, (int)smem, attr_h));
  if (bf16) prefill_attn3_kernel<true><<<grid, p3::kThreads, smem, st>>>(tm_q, tm_k64, tm_v, tm_kc64, tm_vc, p);
  else prefill_attn3_kernel<false><<<grid, p3::kThreads, smem, st>>>(tm_q, tm_k64, tm_v, tm_kc64, tm_vc, p);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
