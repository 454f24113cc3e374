// SnapKV observation-window scorer on tcgen05 (reference: compression/snapkv.py:39-276,332-448; K12 + K13).
//
// Rows = the last w query tokens x the G query heads of one kv head (w * G <= 128: one UMMA M tile; the mma.sync
// kernels of score_attn.cu serve larger windows), columns = 128-key tiles of the sequence's non-window keys:
//   S = Qwin [128 x 128 d] . K_tile^T   one `tcgen05.mma` chain of 8 x (128 x 128 x 16) per key tile.
// Two launches over the same (key split, kv head, sequence) grid, one CTA per split of kSnapSplit keys:
//   pass 1  running (max, sum) of every row over the split's keys          -> stats[part][row]
//   pass 2  column sums of exp2(s - m_row) / l_row with the FINAL row statistics (all parts combined), causal
//           kpool-tap mean inside pool_segment-key segments, window keys <- +inf                 -> out[key, hk]
// Roles per CTA as in score_chunk_umma.cu: warp 8 TMA producer (Q window once: a 3-D box {64 d, G heads, w tokens}
// lands in shared memory already in row order (token, g); K tiles through a 4-slot ring), warp 9 UMMA issuer (four
// 128-column TMEM buffers), warps 0-7 two warpgroups that each own one half of the key columns of every tile: a
// thread keeps its row's 64 logits in registers from one tcgen05.ld on.  The two halves never exchange anything in
// pass 1 (each writes its own partial statistics: parts = 2 x splits) and only meet in pass 2's column sums.
#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

namespace sn {
constexpr int kTile = 128, kHD = 128, kHalf = 64;
constexpr int kTileBytes = kTile * kHD * 2;  // 32 KB
constexpr int kBoxBytes = kTileBytes / 2;
constexpr int kSlots = 4;
constexpr int kThreads = 384;
constexpr int kRegsSoftmax = 224, kRegsOther = 56;
constexpr int kSplit = 1024;  // keys per CTA
}  // namespace sn

struct SnapUmmaParams {
  const int32_t* cu_q;
  const int32_t* cu_k;
  float* stats;  // [B, HKV, parts, 128, 2] partial (m, l) per row; parts = 2 * n_splits
  float* out;    // [Nk, HKV]
  int HKV, G, w, n_splits, kpool, pool_segment;
  float scale_log2;
};

struct SnapUmmaSmem {
  unsigned char q[sn::kTileBytes];
  unsigned char ring[sn::kSlots][sn::kTileBytes];
  float col[4][sn::kTile];
  float raw[sn::kTile];
  uint64_t q_full;
  uint64_t full[sn::kSlots], empty[sn::kSlots];
  uint64_t s_full[4], s_empty[4];
  uint32_t tmem_base;
};

template <int N, int O>
__device__ __forceinline__ void snap_colsum_stage(float (&acc)[sn::kHalf], int lane) {
  const bool hi = (lane & O) != 0;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float keep = hi ? acc[i + N] : acc[i];
    const float send = hi ? acc[i] : acc[i + N];
    acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, O);
  }
}

__device__ __forceinline__ void snap_bar_256() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

template <bool BF16, bool PASS2>
__global__ void __launch_bounds__(sn::kThreads, 1)
snap_umma_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const SnapUmmaParams p) {
  using namespace sn;
  extern __shared__ unsigned char smem_raw[];
  SnapUmmaSmem& sm = *reinterpret_cast<SnapUmmaSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sp = blockIdx.x, hk = blockIdx.y, b = blockIdx.z;
  const int kb = p.cu_k[b], len = p.cu_k[b + 1] - kb, q_end = p.cu_q[b + 1];
  const int rows = p.w * p.G;
  const int n_scored = len - p.w;  // keys [0, L - w) are scored, the window keys get +inf
  const int k_lo = sp * kSplit;
  const int k_hi = min(n_scored, k_lo + kSplit);  // scored keys of this split
  const int n_tiles = k_hi > k_lo ? (k_hi - k_lo + kTile - 1) / kTile : 0;
  const int parts = 2 * p.n_splits;
  float* st_base = p.stats + ((int64_t)b * p.HKV + hk) * parts * kTile * 2;

  if (n_tiles == 0) {  // uniform over the CTA; covers splits beyond the sequence (k_lo >= len) too
    // no scored key here: pass 1 publishes empty statistics; pass 2 writes +inf for window keys (0 when the whole
    // sequence is no longer than the window: the reference leaves those rows uninitialised, snapkv.py:94-95,204-205)
    if (!PASS2) {
      for (int i = threadIdx.x; i < 2 * kTile; i += kThreads) {
        float* st = st_base + ((int64_t)(2 * sp + i / kTile) * kTile + (i % kTile)) * 2;
        st[0] = -INFINITY;
        st[1] = 0.f;
      }
    } else {
      const float v = n_scored > 0 ? INFINITY : 0.f;
      for (int k = k_lo + threadIdx.x; k < min(len, k_lo + kSplit); k += kThreads)
        p.out[(int64_t)(kb + k) * p.HKV + hk] = v;
    }
    return;
  }

  if (threadIdx.x == 0) {
    mbar_init(&sm.q_full, 1);
    for (int i = 0; i < kSlots; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.s_full[i], 1); mbar_init(&sm.s_empty[i], 256); }
    mbar_fence_init();
  }
  if (warp == 9) {
    tmem_alloc(&sm.tmem_base, 512);
    tmem_relinquish();
  }
  // rows >= w * G of the Q tile are never loaded: zero them (their logits are then 0, not NaN from stale smem)
  for (int i = threadIdx.x; i < (kTile - rows) * 16; i += kThreads) {
    const int r = rows + i / 16, c = i % 16;
    *reinterpret_cast<uint4*>(sm.q + (c >> 3) * kBoxBytes + r * 128 + (c & 7) * 16) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 8) {
    // =========================== TMA producer ===========================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    if (lane == 0) {
      mbar_expect_tx(&sm.q_full, 2 * rows * 128);
      tma_load_3d(sm.q, &tm_q, &sm.q_full, 0, hk * p.G, q_end - p.w);
      tma_load_3d(sm.q + kBoxBytes, &tm_q, &sm.q_full, 64, hk * p.G, q_end - p.w);
      for (int t = 0; t < n_tiles; ++t) {
        const int slot = t % kSlots;
        mbar_wait(&sm.empty[slot], ((t / kSlots) & 1) ^ 1);
        mbar_expect_tx(&sm.full[slot], kTileBytes);
        tma_load_3d(sm.ring[slot], &tm_k, &sm.full[slot], 0, hk, kb + k_lo + t * kTile);
        tma_load_3d(sm.ring[slot] + kBoxBytes, &tm_k, &sm.full[slot], 64, hk, kb + k_lo + t * kTile);
      }
    }
  } else if (warp == 9) {
    // ============================ UMMA issuer ============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    constexpr uint32_t idesc = umma_idesc(BF16, kTile, kTile, false, false);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);  // warp-uniform for the compiler: no R2UR per UMMA
    mbar_wait(&sm.q_full, 0);
    for (int t = 0; t < n_tiles; ++t) {
      const int slot = t % kSlots, buf = t & 3;
      mbar_wait(&sm.full[slot], (t / kSlots) & 1);
      mbar_wait(&sm.s_empty[buf], ((t >> 2) & 1) ^ 1);
      tc_fence_after();
      if (lane == 0) {
        const uint64_t q_desc = umma_smem_desc(smem_u32(sm.q), 16, 1024);
        const uint64_t k_desc = umma_smem_desc(smem_u32(sm.ring[slot]), 16, 1024);
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk) {
          const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
          umma_ss(tmem_u + buf * kTile, umma_desc_add(q_desc, off), umma_desc_add(k_desc, off), idesc, kk > 0 ? 1u : 0u);
        }
        umma_commit(&sm.s_full[buf]);
        umma_commit(&sm.empty[slot]);
      }
      __syncwarp();
    }
  } else if (warp < 8) {
    // ====================== softmax statistics / column sums ======================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmax));
    const int wg = warp >> 2, quad = warp & 3;
    const int row = quad * 32 + lane;
    const int st = threadIdx.x;  // 0..255
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f;  // pass 1: this half's running statistics of the row
    float neg_m = 0.f, inv_l = 0.f;        // pass 2: the row's final statistics
    if (PASS2) {
      float M = -INFINITY;
      for (int s = 0; s < parts; ++s) M = fmaxf(M, st_base[((int64_t)s * kTile + row) * 2]);
      float L = 0.f;
      for (int s = 0; s < parts; ++s) {
        const float ms = st_base[((int64_t)s * kTile + row) * 2];
        if (ms > -INFINITY) L += st_base[((int64_t)s * kTile + row) * 2 + 1] * ex2(ms - M);
      }
      neg_m = (M > -INFINITY) ? -M : 0.f;
      inv_l = (row < rows && L > 0.f) ? 1.f / L : 0.f;
    }
    for (int t = 0; t < n_tiles; ++t) {
      const int buf = t & 3;
      const int nk = min(kTile, k_hi - (k_lo + t * kTile));  // scored keys in this tile
      mbar_wait(&sm.s_full[buf], (t >> 2) & 1);
      tc_fence_after();
      uint32_t r0[32], r1[32];
      tmem_ld32(tmem + buf * kTile + wg * kHalf + lane_base, r0);
      tmem_ld32(tmem + buf * kTile + wg * kHalf + lane_base + 32, r1);
      tc_wait_ld();
      tc_fence_before();
      mbar_arrive(&sm.s_empty[buf]);
      float sv[kHalf];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        sv[i] = __uint_as_float(r0[i]) * p.scale_log2;
        sv[32 + i] = __uint_as_float(r1[i]) * p.scale_log2;
      }
      if (nk < kTile) {
#pragma unroll
        for (int i = 0; i < kHalf; ++i)
          if (wg * kHalf + i >= nk) sv[i] = -INFINITY;
      }
      if (!PASS2) {
        float m0 = sv[0], m1 = sv[1], m2 = sv[2], m3 = sv[3];
#pragma unroll
        for (int i = 4; i < kHalf; i += 4) {
          m0 = fmaxf(m0, sv[i]); m1 = fmaxf(m1, sv[i + 1]); m2 = fmaxf(m2, sv[i + 2]); m3 = fmaxf(m3, sv[i + 3]);
        }
        const float m_new = fmaxf(m_run, fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)));
        if (m_new > -INFINITY) {  // else: every key of this half is masked so far
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int i = 0; i < kHalf; i += 2) {
            s0 += ex2(sv[i] - m_new);
            s1 += ex2(sv[i + 1] - m_new);
          }
          l_run = l_run * ex2(m_run - m_new) + (s0 + s1);  // ex2(-inf) = 0 on the first live tile
          m_run = m_new;
        }
      } else {
        float acc[kHalf];
#pragma unroll
        for (int i = 0; i < kHalf; ++i) acc[i] = ex2(sv[i] + neg_m) * inv_l;
        snap_colsum_stage<32, 16>(acc, lane);
        snap_colsum_stage<16, 8>(acc, lane);
        snap_colsum_stage<8, 4>(acc, lane);
        snap_colsum_stage<4, 2>(acc, lane);
        snap_colsum_stage<2, 1>(acc, lane);
        *reinterpret_cast<float2*>(&sm.col[quad][wg * kHalf + 2 * lane]) = make_float2(acc[0], acc[1]);
        snap_bar_256();
        if (st < kTile) sm.raw[st] = ((sm.col[0][st] + sm.col[1][st]) + sm.col[2][st]) + sm.col[3][st];
        snap_bar_256();
        // causal kpool-tap mean inside pool_segment-key segments (snapkv.py:253-262); window keys <- +inf
        if (st < kTile) {
          const int kabs = k_lo + t * kTile + st;
          if (kabs < len) {
            float v = INFINITY;
            if (st < nk) {
              const int lo = max(st - (p.kpool - 1), (st / p.pool_segment) * p.pool_segment);
              float s = 0.f;
              for (int u = lo; u <= st; ++u) s += sm.raw[u];
              v = s / (float)(st + 1 - lo);
            }
            p.out[(int64_t)(kb + kabs) * p.HKV + hk] = v;
          }
        }
        // sm.raw / sm.col are rewritten only behind the next tile's first barrier, which every reader passes first
      }
    }
    if (!PASS2) {
      float* stp = st_base + ((int64_t)(2 * sp + wg) * kTile + row) * 2;
      stp[0] = m_run;
      stp[1] = l_run;
    } else {
      // window keys of this split that lie beyond its last scored tile
      for (int k = k_lo + n_tiles * kTile + st; k < min(len, k_lo + kSplit); k += 256)
        p.out[(int64_t)(kb + k) * p.HKV + hk] = INFINITY;
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));  // warps 10, 11
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, 512);
}

}  // namespace cvllm

using namespace cvllm;

size_t cvllm_snap_umma_workspace_bytes(int B, int max_seqlen_k, int HKV) {
  const size_t n_splits = (size_t)(max_seqlen_k + sn::kSplit - 1) / sn::kSplit;
  return (size_t)B * HKV * 2 * (n_splits ? n_splits : 1) * sn::kTile * 2 * sizeof(float);
}

// Launcher used by cvllm_snapkv_scores (score_attn.cu) when w * G <= 128.
int cvllm_launch_snap_umma(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                           const int32_t* cu_q, const int32_t* cu_k, int w, int B, int Nq, int Nk, int HQ, int HKV,
                           int max_seqlen_k, float scale_log2, int kpool, int pool_segment, float* out, float* stats,
                           bool bf16, cudaStream_t st) {
  const int G = HQ / HKV;
  CUtensorMap tm_q, tm_k;
  const uint32_t boxq[3] = {64, (uint32_t)G, (uint32_t)w};
  const uint64_t dq[3] = {(uint64_t)sn::kHD, (uint64_t)HQ, (uint64_t)Nq};
  const uint64_t sq[2] = {(uint64_t)sn::kHD * 2, (uint64_t)q_stride_n * 2};
  int rc = make_tmap_16b(&tm_q, q, 3, dq, sq, boxq, bf16);
  if (rc) return rc;
  const uint32_t boxk[3] = {64, 1, (uint32_t)sn::kTile};
  const uint64_t dk[3] = {(uint64_t)sn::kHD, (uint64_t)HKV, (uint64_t)Nk};
  const uint64_t sk[2] = {(uint64_t)sn::kHD * 2, (uint64_t)k_stride_n * 2};
  rc = make_tmap_16b(&tm_k, k, 3, dk, sk, boxk, bf16);
  if (rc) return rc;
  SnapUmmaParams p;
  p.cu_q = cu_q; p.cu_k = cu_k; p.stats = stats; p.out = out; p.HKV = HKV; p.G = G; p.w = w;
  p.n_splits = (max_seqlen_k + sn::kSplit - 1) / sn::kSplit;
  p.kpool = kpool; p.pool_segment = pool_segment; p.scale_log2 = scale_log2;
  const dim3 grid(p.n_splits, HKV, B);
  const int smem = (int)sizeof(SnapUmmaSmem) + 1024;
  static std::atomic<uint64_t> attr[4] = {{0}, {0}, {0}, {0}};
  if (bf16) {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)snap_umma_kernel<true, false>, smem, attr[0]));
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)snap_umma_kernel<true, true>, smem, attr[1]));
    snap_umma_kernel<true, false><<<grid, sn::kThreads, smem, st>>>(tm_q, tm_k, p);
    snap_umma_kernel<true, true><<<grid, sn::kThreads, smem, st>>>(tm_q, tm_k, p);
  } else {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)snap_umma_kernel<false, false>, smem, attr[2]));
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)snap_umma_kernel<false, true>, smem, attr[3]));
    snap_umma_kernel<false, false><<<grid, sn::kThreads, smem, st>>>(tm_q, tm_k, p);
    snap_umma_kernel<false, true><<<grid, sn::kThreads, smem, st>>>(tm_q, tm_k, p);
  }
  CVLLM_CHECK_LAUNCHES(2);
  return CVLLM_OK;
}
