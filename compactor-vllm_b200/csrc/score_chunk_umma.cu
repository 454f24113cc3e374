// Compactor chunk scorer on tcgen05 (reference: compression/compactor.py:338-599, kernel K11 :350-482).
//
// One item = (sequence b, kv head h, 128-token chunk).  Its scores are the column sums of softmax(Q_rows K^T) over
// the chunk's own keys, rows = the chunk's tokens x the G query heads of h.  That is G UMMA tiles
// S_g = Q_g [128 tok x 128 d] . K^T [128 d x 128 keys] sharing one K tile:
//
//   warp 8      TMA producer: K through two dedicated 32 KB buffers, Q_g tiles through a 4-slot ring
//               (cp.async.bulk.tensor, 128B swizzle, 3-D maps [D, H, N] over the caller's strided q / k)
//   warp 9      UMMA issuer: 8 x (128x128x16) per tile, accumulator = one of four 128-column TMEM buffers
//   warps 0-3 / 4-7   two softmax warpgroups, alternating items; a thread owns one token row.  Per tile three
//               passes over the TMEM tile (max | exp2, row sum, e written back | acc += e / l), so the 128
//               per-column accumulators of the item are the only long-lived registers; after the G tiles the
//               accumulators are summed across the 32 rows of each warp by a 124-shuffle transpose-reduce and
//               across the four warps through shared memory.
// Persistent: grid = min(items, #SM); 192 KB of tiles in flight per SM keep it on the HBM roofline
// (10 KB per token: 335 MB at 32K tokens) instead of the mma.sync version's issue / latency limit.
#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

namespace cs {
constexpr int kTile = 128, kHD = 128;
constexpr int kTileBytes = kTile * kHD * 2;  // 32 KB
constexpr int kBoxBytes = kTileBytes / 2;    // one [128 x 64] swizzled box
constexpr int kQSlots = 4;
constexpr int kThreads = 384;                // 3 warpgroups
constexpr int kRegsSoftmax = 224, kRegsOther = 56;
constexpr int kRefBlockM = 64;               // the reference's only K11 config (compactor.py:338-347)
}  // namespace cs

struct ChunkUmmaParams {
  const int32_t* cu;
  float* raw;       // [N, HKV] column sums (pre z-score)
  float* partials;  // [B, max_chunks, HKV, 2] (sum, sumsq) of raw per item
  int B, HKV, G, chunk, max_chunks, n_ci;
  float scale_log2;
};

struct ChunkUmmaSmem {
  unsigned char kbuf[2][cs::kTileBytes];
  unsigned char qring[cs::kQSlots][cs::kTileBytes];
  float col[2][4][cs::kTile];  // [softmax group][warp][key column]
  float red[2][2][4];
  uint64_t k_full[2], k_empty[2];
  uint64_t q_full[cs::kQSlots], q_empty[cs::kQSlots];
  uint64_t s_full[4], s_empty[4];
  uint32_t tmem_base;
};

// One stage of the transpose-reduce: N columns survive; lanes with bit O set keep the upper N, the others the lower
// N, and each hands the half it drops to its partner (lane ^ O).
template <int N, int O>
__device__ __forceinline__ void colsum_stage(float (&acc)[cs::kTile], int lane) {
  const bool hi = (lane & O) != 0;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float keep = hi ? acc[i + N] : acc[i];
    const float send = hi ? acc[i] : acc[i + N];
    acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, O);
  }
}

__device__ __forceinline__ void wg_bar(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

template <bool BF16>
__global__ void __launch_bounds__(cs::kThreads, 1)
chunk_scores_umma_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                         const ChunkUmmaParams p) {
  using namespace cs;
  extern __shared__ unsigned char smem_raw[];
  ChunkUmmaSmem& sm = *reinterpret_cast<ChunkUmmaSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = p.G;
  const int total = p.B * p.HKV * p.n_ci;

  // item -> (b, h, ci, chunk start, tokens); false when the chunk lies beyond its sequence
  auto decode = [&](int idx, int& b, int& h, int& ci, int& c0, int& M) -> bool {
    ci = idx % p.n_ci;
    const int t = idx / p.n_ci;
    h = t % p.HKV;
    b = t / p.HKV;
    const int seq_beg = p.cu[b], seq_end = p.cu[b + 1];
    c0 = seq_beg + ci * p.chunk;
    M = min(p.chunk, seq_end - c0);
    return c0 < seq_end;
  };
  auto next_valid = [&](int idx) -> int {
    int b, h, ci, c0, M;
    while (idx < total && !decode(idx, b, h, ci, c0, M)) idx += gridDim.x;
    return idx;
  };

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], 1); }
    for (int i = 0; i < kQSlots; ++i) { mbar_init(&sm.q_full[i], 1); mbar_init(&sm.q_empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.s_full[i], 1); mbar_init(&sm.s_empty[i], 128); }
    mbar_fence_init();
  }
  if (warp == 9) {
    tmem_alloc(&sm.tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 8) {
    // =========================== TMA producer ===========================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    if (lane == 0) {
      int it = 0, qn = 0;
      for (int idx = next_valid(blockIdx.x); idx < total; idx = next_valid(idx + gridDim.x), ++it) {
        int b, h, ci, c0, M;
        decode(idx, b, h, ci, c0, M);
        const int kb = it & 1;
        mbar_wait(&sm.k_empty[kb], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&sm.k_full[kb], kTileBytes);
        tma_load_3d(sm.kbuf[kb], &tm_k, &sm.k_full[kb], 0, h, c0);
        tma_load_3d(sm.kbuf[kb] + kBoxBytes, &tm_k, &sm.k_full[kb], 64, h, c0);
        for (int g = 0; g < G; ++g, ++qn) {
          const int slot = qn % kQSlots;
          mbar_wait(&sm.q_empty[slot], ((qn / kQSlots) & 1) ^ 1);
          mbar_expect_tx(&sm.q_full[slot], kTileBytes);
          tma_load_3d(sm.qring[slot], &tm_q, &sm.q_full[slot], 0, h * G + g, c0);
          tma_load_3d(sm.qring[slot] + kBoxBytes, &tm_q, &sm.q_full[slot], 64, h * G + g, c0);
        }
      }
    }
  } else if (warp == 9) {
    // ============================ UMMA issuer ============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    constexpr uint32_t idesc = umma_idesc(BF16, kTile, kTile, false, false);
    int it = 0, qn = 0;
    int wg_tiles[2] = {0, 0};  // tiles handed to each softmax group so far
    for (int idx = next_valid(blockIdx.x); idx < total; idx = next_valid(idx + gridDim.x), ++it) {
      const int kb = it & 1, w = it & 1;
      mbar_wait(&sm.k_full[kb], (it >> 1) & 1);
      for (int g = 0; g < G; ++g, ++qn) {
        const int slot = qn % kQSlots;
        const int k = wg_tiles[w]++;
        const int buf = 2 * w + (k & 1);
        mbar_wait(&sm.q_full[slot], (qn / kQSlots) & 1);
        mbar_wait(&sm.s_empty[buf], ((k >> 1) & 1) ^ 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t q_addr = smem_u32(sm.qring[slot]), k_addr = smem_u32(sm.kbuf[kb]);
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk) {
            const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
            umma_ss(tmem + buf * kTile, umma_smem_desc(q_addr + off, 16, 1024),
                    umma_smem_desc(k_addr + off, 16, 1024), idesc, kk > 0 ? 1u : 0u);
          }
          umma_commit(&sm.s_full[buf]);
          umma_commit(&sm.q_empty[slot]);
          if (g == G - 1) umma_commit(&sm.k_empty[kb]);
        }
        __syncwarp();
      }
    }
  } else if (warp < 8) {
    // ====================== softmax + column sums ======================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmax));
    const int w = warp >> 2, quad = warp & 3;
    const int row = quad * 32 + lane;  // token row of the chunk == key column in the epilogue
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    int it = 0, k = 0;
    for (int idx = next_valid(blockIdx.x); idx < total; idx = next_valid(idx + gridDim.x), ++it) {
      if ((it & 1) != w) continue;
      int b, h, ci, c0, M;
      decode(idx, b, h, ci, c0, M);
      const bool ragged = M < kTile;
      float acc[kTile];
#pragma unroll
      for (int c = 0; c < kTile; ++c) acc[c] = 0.f;
      for (int g = 0; g < G; ++g, ++k) {
        const int buf = 2 * w + (k & 1);
        const uint32_t s_t = tmem + buf * kTile + lane_base;
        mbar_wait(&sm.s_full[buf], (k >> 1) & 1);
        tc_fence_after();
        // Three passes over the TMEM tile in 16-column pieces, software-pipelined: the load of piece c+1 is in
        // flight while piece c is processed (tcgen05.wait::ld waits for every outstanding load, so it sits
        // AFTER the work on the current piece).  pass 1: row max | pass 2: e = exp2(s*scale - m), row sum,
        // e written back over S | pass 3: acc += e / l.
        uint32_t pa[16], pb[16];
        float mx = -INFINITY;
        tmem_ld16(s_t, pa);
        tc_wait_ld();
#pragma unroll
        for (int c = 0; c < kTile / 16; ++c) {
          uint32_t(&cur)[16] = (c & 1) ? pb : pa;
          uint32_t(&nxt)[16] = (c & 1) ? pa : pb;
          if (c + 1 < kTile / 16) tmem_ld16(s_t + (c + 1) * 16, nxt);
          else tmem_ld16(s_t, nxt);  // first piece of pass 2
          float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
          for (int i = 0; i < 16; i += 4) {
            float a0 = __uint_as_float(cur[i]), a1 = __uint_as_float(cur[i + 1]);
            float a2 = __uint_as_float(cur[i + 2]), a3 = __uint_as_float(cur[i + 3]);
            if (ragged) {
              if (c * 16 + i >= M) a0 = -INFINITY;
              if (c * 16 + i + 1 >= M) a1 = -INFINITY;
              if (c * 16 + i + 2 >= M) a2 = -INFINITY;
              if (c * 16 + i + 3 >= M) a3 = -INFINITY;
            }
            m0 = fmaxf(m0, a0); m1 = fmaxf(m1, a1); m2 = fmaxf(m2, a2); m3 = fmaxf(m3, a3);
          }
          mx = fmaxf(mx, fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)));
          tc_wait_ld();
        }
        const float neg_m = -mx * p.scale_log2;  // M >= 1: every row has at least one live key
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int c = 0; c < kTile / 16; ++c) {  // kTile / 16 is even: pass 2 starts on pa again
          uint32_t(&cur)[16] = (c & 1) ? pb : pa;
          uint32_t(&nxt)[16] = (c & 1) ? pa : pb;
          if (c + 1 < kTile / 16) tmem_ld16(s_t + (c + 1) * 16, nxt);
#pragma unroll
          for (int i = 0; i < 16; i += 2) {
            float e0 = ex2(fmaf(__uint_as_float(cur[i]), p.scale_log2, neg_m));
            float e1 = ex2(fmaf(__uint_as_float(cur[i + 1]), p.scale_log2, neg_m));
            if (ragged) {
              if (c * 16 + i >= M) e0 = 0.f;
              if (c * 16 + i + 1 >= M) e1 = 0.f;
            }
            l0 += e0;
            l1 += e1;
            cur[i] = __float_as_uint(e0);
            cur[i + 1] = __float_as_uint(e1);
          }
          tmem_st16(s_t + c * 16, cur);
          tc_wait_ld();
        }
        tc_wait_st();
        const float inv = (row < M) ? 1.f / (l0 + l1) : 0.f;  // rows beyond the chunk contribute nothing
        tmem_ld16(s_t, pa);
        tc_wait_ld();
#pragma unroll
        for (int c = 0; c < kTile / 16; ++c) {
          uint32_t(&cur)[16] = (c & 1) ? pb : pa;
          uint32_t(&nxt)[16] = (c & 1) ? pa : pb;
          if (c + 1 < kTile / 16) tmem_ld16(s_t + (c + 1) * 16, nxt);
#pragma unroll
          for (int i = 0; i < 16; ++i) acc[c * 16 + i] = fmaf(__uint_as_float(cur[i]), inv, acc[c * 16 + i]);
          tc_wait_ld();
        }
        tc_fence_before();
        mbar_arrive(&sm.s_empty[buf]);
      }
      // column sums over the warp's 32 rows: each stage keeps one half of the columns and hands the other half
      // to the partner lane; lane L ends with columns 4L .. 4L+3
      colsum_stage<64, 16>(acc, lane);
      colsum_stage<32, 8>(acc, lane);
      colsum_stage<16, 4>(acc, lane);
      colsum_stage<8, 2>(acc, lane);
      colsum_stage<4, 1>(acc, lane);
      *reinterpret_cast<float4*>(&sm.col[w][quad][4 * lane]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      wg_bar(1 + w);
      // rows of the reference's last 64-row query tile that fall beyond the chunk add 1/chunk each
      // (compactor.py:411-414,477-479)
      const float pad = (float)(((M + kRefBlockM - 1) / kRefBlockM * kRefBlockM - M) * G) / (float)p.chunk;
      float v = 0.f;
      if (row < M) {
        v = ((sm.col[w][0][row] + sm.col[w][1][row]) + sm.col[w][2][row]) + sm.col[w][3][row] + pad;
        p.raw[(int64_t)(c0 + row) * p.HKV + h] = v;
      }
      const float s1 = warp_sum(v), s2 = warp_sum(v * v);
      if (lane == 0) { sm.red[w][0][quad] = s1; sm.red[w][1][quad] = s2; }
      wg_bar(1 + w);
      if (row == 0) {
        float* dst = p.partials + (((int64_t)b * p.max_chunks + ci) * p.HKV + h) * 2;
        dst[0] = ((sm.red[w][0][0] + sm.red[w][0][1]) + sm.red[w][0][2]) + sm.red[w][0][3];
        dst[1] = ((sm.red[w][1][0] + sm.red[w][1][1]) + sm.red[w][1][2]) + sm.red[w][1][3];
      }
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

// Launcher used by cvllm_chunk_attn_scores (score_attn.cu).  q / k: [N, H, 128] 16-bit, head stride 128 elements,
// token strides in elements.
int cvllm_launch_chunk_scores_umma(const void* q, const void* k, int64_t q_stride_n, int64_t k_stride_n,
                                   const int32_t* cu, int B, int N, int HQ, int HKV, int chunk, int max_chunks,
                                   int n_ci, float scale_log2, float* raw, float* partials, bool bf16,
                                   cudaStream_t st) {
  CUtensorMap tm_q, tm_k;
  const uint32_t box[3] = {64, 1, (uint32_t)cs::kTile};
  const uint64_t dq[3] = {(uint64_t)cs::kHD, (uint64_t)HQ, (uint64_t)N};
  const uint64_t sq[2] = {(uint64_t)cs::kHD * 2, (uint64_t)q_stride_n * 2};
  int rc = make_tmap_16b(&tm_q, q, 3, dq, sq, box, bf16);
  if (rc) return rc;
  const uint64_t dk[3] = {(uint64_t)cs::kHD, (uint64_t)HKV, (uint64_t)N};
  const uint64_t sk[2] = {(uint64_t)cs::kHD * 2, (uint64_t)k_stride_n * 2};
  rc = make_tmap_16b(&tm_k, k, 3, dk, sk, box, bf16);
  if (rc) return rc;
  ChunkUmmaParams p;
  p.cu = cu; p.raw = raw; p.partials = partials; p.B = B; p.HKV = HKV; p.G = HQ / HKV; p.chunk = chunk;
  p.max_chunks = max_chunks; p.n_ci = n_ci; p.scale_log2 = scale_log2;
  const int64_t items = (int64_t)B * HKV * n_ci;
  const int grid = (int)(items < sm_count() ? items : sm_count());
  const int smem = (int)sizeof(ChunkUmmaSmem) + 1024;
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)chunk_scores_umma_kernel<true>, smem, attr_bf));
    chunk_scores_umma_kernel<true><<<grid, cs::kThreads, smem, st>>>(tm_q, tm_k, p);
  } else {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)chunk_scores_umma_kernel<false>, smem, attr_h));
    chunk_scores_umma_kernel<false><<<grid, cs::kThreads, smem, st>>>(tm_q, tm_k, p);
  }
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
