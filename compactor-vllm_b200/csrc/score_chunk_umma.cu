// Compactor chunk scorer on tcgen05 (reference: compression/compactor.py:338-599, kernel K11 :350-482).
//
// One item = (sequence b, kv head h, 128-token chunk).  Its scores are the column sums of softmax(Q_rows K^T) over
// the chunk's own keys, rows = the chunk's tokens x the G query heads of h.  That is G UMMA tiles
// S_g = Q_g [128 tok x 128 d] . K^T [128 d x 128 keys] sharing one K tile:
//
//   warp 8      TMA producer: K through two dedicated 32 KB buffers, Q_g tiles through a 4-slot ring
//               (cp.async.bulk.tensor, 128B swizzle, 3-D maps [D, H, N] over the caller's strided q / k)
//   warp 9      UMMA issuer: 8 x (128x128x16) per tile, accumulator = one of four 128-column TMEM buffers
//   warps 0-3 / 4-7   two softmax warpgroups on every tile: a thread owns one token row and one half of the key
//               columns, keeps its 64 logits in registers from one tcgen05.ld to the accumulation (row max and
//               row sum are completed across the halves through shared memory), and carries 64 per-column
//               accumulators over the item's G tiles; these are then summed across the 32 rows of each warp by
//               a 62-shuffle transpose-reduce and across the four warps through shared memory.
// Persistent: grid = min(items, #SM); 192 KB of tiles in flight per SM keep it on the HBM roofline
// (10 KB per token: 335 MB at 32K tokens) instead of the mma.sync version's issue / latency limit.
#include "common.cuh"
#include "tma.cuh"

namespace cvllm {

namespace cs {
constexpr int kTile = 128, kHD = 128;
constexpr int kTileBytes = kTile * kHD * 2;  // 32 KB
constexpr int kBoxBytes = kTileBytes / 2;    // one [128 x 64] swizzled box
constexpr int kMaxQSlots = 4;  // 2 K buffers + 4 Q slots = 192 KB of tiles in flight per SM
constexpr int kThreads = 384;                // 3 warpgroups
constexpr int kRegsSoftmax = 224, kRegsOther = 56;
constexpr int kRefBlockM = 64;               // the reference's only K11 config (compactor.py:338-347)
}  // namespace cs

struct ChunkUmmaParams {
  const int32_t* cu;
  float* raw;       // [N, HKV] column sums (pre z-score)
  float* partials;  // [B, max_chunks, HKV, 2] (sum, sumsq) of raw per item
  int B, HKV, G, chunk, max_chunks, n_ci;
  int k_bufs, q_slots;  // tile buffers of this launch (the 32 KB tiles follow the struct below in shared memory)
  float scale_log2;
};

struct ChunkUmmaSmem {
  float col[4][cs::kTile];       // [warp quadrant][key column]: column sums over the quadrant's 32 rows
  float xmax[2][cs::kTile];      // row maximum / row sum of each column half, exchanged between the warpgroups
  float xsum[2][cs::kTile];
  float red[2][4];
  uint64_t k_full[2], k_empty[2];
  uint64_t q_full[cs::kMaxQSlots], q_empty[cs::kMaxQSlots];
  uint64_t s_full[4], s_empty[4];
  uint32_t tmem_base;
};

// One stage of the transpose-reduce: N columns survive; lanes with bit O set keep the upper N, the others the lower
// N, and each hands the half it drops to its partner (lane ^ O).
template <int N, int O>
__device__ __forceinline__ void colsum_stage(float (&acc)[cs::kTile / 2], int lane) {
  const bool hi = (lane & O) != 0;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float keep = hi ? acc[i + N] : acc[i];
    const float send = hi ? acc[i] : acc[i + N];
    acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, O);
  }
}

__device__ __forceinline__ void sm_bar_256() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

template <bool BF16>
__global__ void __launch_bounds__(cs::kThreads, 1)
chunk_scores_umma_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                         const ChunkUmmaParams p) {
  using namespace cs;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* tiles = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int kKBufs = p.k_bufs, kQSlots = p.q_slots;
  unsigned char* const kbuf0 = tiles;                                  // [k_bufs][32 KB]
  unsigned char* const qring0 = tiles + (size_t)kKBufs * kTileBytes;    // [q_slots][32 KB]
  ChunkUmmaSmem& sm = *reinterpret_cast<ChunkUmmaSmem*>(qring0 + (size_t)kQSlots * kTileBytes);
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
    for (int i = 0; i < kMaxQSlots; ++i) { mbar_init(&sm.q_full[i], 1); mbar_init(&sm.q_empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.s_full[i], 1); mbar_init(&sm.s_empty[i], 256); }
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
        const int kb = it % kKBufs;
        unsigned char* kdst = kbuf0 + (size_t)kb * kTileBytes;
        mbar_wait(&sm.k_empty[kb], ((it / kKBufs) & 1) ^ 1);
        mbar_expect_tx(&sm.k_full[kb], kTileBytes);
        tma_load_3d(kdst, &tm_k, &sm.k_full[kb], 0, h, c0);
        tma_load_3d(kdst + kBoxBytes, &tm_k, &sm.k_full[kb], 64, h, c0);
        for (int g = 0; g < G; ++g, ++qn) {
          const int slot = qn % kQSlots;
          mbar_wait(&sm.q_empty[slot], ((qn / kQSlots) & 1) ^ 1);
          mbar_expect_tx(&sm.q_full[slot], kTileBytes);
          unsigned char* qdst = qring0 + (size_t)slot * kTileBytes;
          tma_load_3d(qdst, &tm_q, &sm.q_full[slot], 0, h * G + g, c0);
          tma_load_3d(qdst + kBoxBytes, &tm_q, &sm.q_full[slot], 64, h * G + g, c0);
        }
      }
    }
  } else if (warp == 9) {
    // ============================ UMMA issuer ============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    constexpr uint32_t idesc = umma_idesc(BF16, kTile, kTile, false, false);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);  // warp-uniform for the compiler: no R2UR per UMMA
    int it = 0, qn = 0;  // qn doubles as the CTA-local tile counter: TMEM buffer qn % 4
    for (int idx = next_valid(blockIdx.x); idx < total; idx = next_valid(idx + gridDim.x), ++it) {
      const int kb = it % kKBufs;
      mbar_wait(&sm.k_full[kb], (it / kKBufs) & 1);
      for (int g = 0; g < G; ++g, ++qn) {
        const int slot = qn % kQSlots;
        const int buf = qn & 3;
        mbar_wait(&sm.q_full[slot], (qn / kQSlots) & 1);
        mbar_wait(&sm.s_empty[buf], ((qn >> 2) & 1) ^ 1);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t q_desc = umma_smem_desc(smem_u32(qring0 + (size_t)slot * kTileBytes), 16, 1024);
          const uint64_t k_desc = umma_smem_desc(smem_u32(kbuf0 + (size_t)kb * kTileBytes), 16, 1024);
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk) {
            const uint32_t off = (kk >> 2) * kBoxBytes + (kk & 3) * 32;
            umma_ss(tmem_u + buf * kTile, umma_desc_add(q_desc, off), umma_desc_add(k_desc, off), idesc,
                    kk > 0 ? 1u : 0u);
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
    // Both warpgroups work on EVERY tile: a thread owns one token row and one HALF of the key columns
    // (warpgroup 0: keys 0-63, warpgroup 1: keys 64-127).  Its 64 logits stay in registers from the single
    // TMEM load to the accumulation, so a tile costs one tcgen05.ld per element and no write-back; the row
    // maximum and the row sum are completed across the two halves through shared memory (two 256-thread
    // named barriers per tile).
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmax));
    constexpr int kHalf = kTile / 2;
    const int w = warp >> 2, quad = warp & 3;
    const int row = quad * 32 + lane;
    const int st = threadIdx.x;  // 0..255
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    int t = 0;  // CTA-local tile counter: TMEM buffer t % 4
    for (int idx = next_valid(blockIdx.x); idx < total; idx = next_valid(idx + gridDim.x)) {
      int b, h, ci, c0, M;
      decode(idx, b, h, ci, c0, M);
      const bool ragged = M < kTile;
      float acc[kHalf];
#pragma unroll
      for (int c = 0; c < kHalf; ++c) acc[c] = 0.f;
      for (int g = 0; g < G; ++g, ++t) {
        const int buf = t & 3;
        const uint32_t s_t = tmem + buf * kTile + w * kHalf + lane_base;
        mbar_wait(&sm.s_full[buf], (t >> 2) & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld32(s_t, r0);
        tmem_ld32(s_t + 32, r1);
        tc_wait_ld();
        tc_fence_before();
        mbar_arrive(&sm.s_empty[buf]);  // the logits are in registers: the issuer may overwrite the buffer
        float sv[kHalf];
#pragma unroll
        for (int i = 0; i < 32; ++i) { sv[i] = __uint_as_float(r0[i]); sv[32 + i] = __uint_as_float(r1[i]); }
        if (ragged) {
#pragma unroll
          for (int i = 0; i < kHalf; ++i)
            if (w * kHalf + i >= M) sv[i] = -INFINITY;
        }
        float m0 = sv[0], m1 = sv[1], m2 = sv[2], m3 = sv[3];
#pragma unroll
        for (int i = 4; i < kHalf; i += 4) {
          m0 = fmaxf(m0, sv[i]); m1 = fmaxf(m1, sv[i + 1]); m2 = fmaxf(m2, sv[i + 2]); m3 = fmaxf(m3, sv[i + 3]);
        }
        float mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
        sm.xmax[w][row] = mx;
        sm_bar_256();
        mx = fmaxf(mx, sm.xmax[w ^ 1][row]);
        const float neg_m = -mx * p.scale_log2;  // M >= 1: key 0 is live in every row, so mx is finite
        const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), nm2 = make_float2(neg_m, neg_m);
        float2 la = make_float2(0.f, 0.f), lb = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < kHalf; i += 4) {
          const float2 a = __ffma2_rn(make_float2(sv[i], sv[i + 1]), sc2, nm2);
          const float2 c = __ffma2_rn(make_float2(sv[i + 2], sv[i + 3]), sc2, nm2);
          sv[i] = ex2(a.x); sv[i + 1] = ex2(a.y); sv[i + 2] = ex2(c.x); sv[i + 3] = ex2(c.y);  // ex2(-inf) = 0
          la = __fadd2_rn(la, make_float2(sv[i], sv[i + 1]));
          lb = __fadd2_rn(lb, make_float2(sv[i + 2], sv[i + 3]));
        }
        const float l_half = (la.x + lb.x) + (la.y + lb.y);
        sm.xsum[w][row] = l_half;
        sm_bar_256();
        const float l_row = w == 0 ? l_half + sm.xsum[1][row] : sm.xsum[0][row] + l_half;  // same bits in both halves
        const float inv = (row < M) ? 1.f / l_row : 0.f;  // rows beyond the chunk contribute nothing
        const float2 inv2 = make_float2(inv, inv);
#pragma unroll
        for (int i = 0; i < kHalf; i += 2) {
          const float2 r = __ffma2_rn(make_float2(sv[i], sv[i + 1]), inv2, make_float2(acc[i], acc[i + 1]));
          acc[i] = r.x;
          acc[i + 1] = r.y;
        }
      }
      // column sums over the warp's 32 rows: each stage keeps one half of the columns and hands the other half
      // to the partner lane; lane L ends with columns 2L, 2L+1 of this warpgroup's half
      colsum_stage<32, 16>(acc, lane);
      colsum_stage<16, 8>(acc, lane);
      colsum_stage<8, 4>(acc, lane);
      colsum_stage<4, 2>(acc, lane);
      colsum_stage<2, 1>(acc, lane);
      *reinterpret_cast<float2*>(&sm.col[quad][w * kHalf + 2 * lane]) = make_float2(acc[0], acc[1]);
      sm_bar_256();
      // rows of the reference's last 64-row query tile that fall beyond the chunk add 1/chunk each
      // (compactor.py:411-414,477-479)
      const float pad = (float)(((M + kRefBlockM - 1) / kRefBlockM * kRefBlockM - M) * G) / (float)p.chunk;
      float v = 0.f;
      if (st < M) {  // st < 128: warpgroup 0, one thread per key column
        v = ((sm.col[0][st] + sm.col[1][st]) + sm.col[2][st]) + sm.col[3][st] + pad;
        p.raw[(int64_t)(c0 + st) * p.HKV + h] = v;
      }
      if (w == 0) {
        const float s1 = warp_sum(v), s2 = warp_sum(v * v);
        if (lane == 0) { sm.red[0][quad] = s1; sm.red[1][quad] = s2; }
      }
      sm_bar_256();
      if (st == 0) {
        float* dst = p.partials + (((int64_t)b * p.max_chunks + ci) * p.HKV + h) * 2;
        dst[0] = ((sm.red[0][0] + sm.red[0][1]) + sm.red[0][2]) + sm.red[0][3];
        dst[1] = ((sm.red[1][0] + sm.red[1][1]) + sm.red[1][2]) + sm.red[1][3];
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
  p.k_bufs = 2;
  p.q_slots = cs::kMaxQSlots;
  const int64_t items = (int64_t)B * HKV * n_ci;
  const int grid = (int)(items < sm_count() ? items : sm_count());
  const int smem = (p.k_bufs + p.q_slots) * cs::kTileBytes + (int)sizeof(ChunkUmmaSmem) + 1024;
  const int smem_max = (2 + cs::kMaxQSlots) * cs::kTileBytes + (int)sizeof(ChunkUmmaSmem) + 1024;
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)chunk_scores_umma_kernel<true>, smem_max, attr_bf));
    chunk_scores_umma_kernel<true><<<grid, cs::kThreads, smem, st>>>(tm_q, tm_k, p);
  } else {
    CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)chunk_scores_umma_kernel<false>, smem_max, attr_h));
    chunk_scores_umma_kernel<false><<<grid, cs::kThreads, smem, st>>>(tm_q, tm_k, p);
  }
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
