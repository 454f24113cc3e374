// Row-parallel linear FUSED with its tensor-parallel all-reduce (reference: layers/linear.py:133-153
// `RowParallelLinear.forward` = `F.linear(x, W)` then `dist.all_reduce(y)`, the o_proj of models/llama3.py:111).
//
//   y[M, N] = sum over ranks r of  bf16( x_r[M, K_r] . W_r[N, K_r]^T )        (fp32 sum in rank order, rounded once)
//
// ONE kernel per rank does the GEMM, the reduce-scatter and the all-gather over NVLink peer memory:
//   * the GEMM runs on CTA PAIRS (tcgen05.mma.cta_group::2, 256 x 256 tile per pair, fp32 accumulators double-buffered
//     in TMEM, operands staged by TMA through a 6-deep ring; each CTA of a pair loads its 128 rows of x and HALF of
//     the W tile, so a pair moves 64 B/clk/SM from L2 instead of the 96 B/clk of a lone 128 x 256 tile);
//   * rows are owned by ranks in blocks of Mo = ceil(M / world) rows.  The epilogue rounds a finished 128 x 256
//     accumulator to 16 bits and TMA-STORES it straight into the OWNER's receive slot for this rank (a peer GPU's
//     memory unless the owner is this rank), then adds 1 to a counter in the owner's memory (system-scope release).
//     Ranks walk the owners in rotated order (rank+1, rank+2, ...), so at any time every NVLink port has one sender
//     and one receiver, and the transfer of tile i overlaps the math of tile i+1;
//   * owner blocks are cut into ROUNDS of R rows.  Eight reducer warps per CTA wait until every rank's counter for a
//     round is complete, sum the `world` slots in rank order in fp32, and store the 16-bit result into y on EVERY rank
//     (the all-gather, also by peer stores) — while the tensor pipe is already working on later rounds;
//   * a call ends when every peer has signalled that all its stores into this rank's y are performed.
// y and the slots live in one cudaIpc-shared allocation per rank (cvllm_ipc_alloc), so `world` processes map each
// other's buffers; a single process may also pass plain device pointers (several virtual ranks on one GPU: the tests).
//
// Memory model: every flag or counter that guards data read by another GPU is written after `fence.acq_rel.sys` and
// read with `ld.acquire.sys` (csrc/allreduce.cu has the argument).  TMA stores are async-proxy writes: the storing
// thread waits for the bulk group to COMPLETE (not only to have read shared memory) before the release.
// Write-after-read across calls: (a) slots — a rank leaves call n only after every peer's DONE(n), which a peer sends
// after its reducers have read their slots; (b) y — a reducer pushes into peer q's y only after q's STARTED(n+1),
// which q sends from inside call n+1, i.e. after the kernels that consumed y of call n on q's stream.
#include <string.h>

#include "common.cuh"
#include "tma.cuh"

namespace cvllm {
namespace grs {
constexpr int kStages = 6;
constexpr int kBM = 128;  // rows per CTA (256 per pair)
constexpr int kBN = 256;  // columns per pair tile
constexpr int kBK = 64;
constexpr int kABytes = kBM * kBK * 2;         // 16 KB: x rows of this CTA
constexpr int kBBytes = (kBN / 2) * kBK * 2;   // 16 KB: this CTA's half of the W tile
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kBoxBytes = 128 * 128;           // [128 rows x 64 cols] 16-bit, 128-byte swizzled rows
constexpr int kThreads = 512;                  // warps: 0 TMA, 1 UMMA, 3 TMEM, 4-7 epilogue, 8-15 reducer
constexpr int kReducers = 256;
constexpr int kMaxWorld = 8;
constexpr int kMaxRounds = 128;
// flag words at the start of a rank's buffer (the kArFlagBytes header of cvllm_ipc_alloc, zeroed at allocation)
constexpr int kWEpoch = 0, kWFin = 1, kWStarted = 8, kWDone = 16, kWCnt = 32;  // kWCnt + round * 8 + source rank
constexpr size_t kHeaderBytes = 6656;  // == cvllm_allreduce_flag_bytes()
static_assert((kWCnt + kMaxRounds * 8) * 4 <= kHeaderBytes, "flag words fit the header");
}  // namespace grs

struct alignas(64) GrsMaps {
  CUtensorMap a, b, slot[grs::kMaxWorld];
};

struct GrsParams {
  void* base[grs::kMaxWorld];  // rank r's buffer as mapped here
  int64_t slot_off, y_off;     // bytes from a buffer's start
  int M, N, K;
  int Mo, R, C;                // rows per owner, rows per round, rounds (Mo = C * R)
  int NT, KS, mc;              // column tiles, k steps, pair tiles (256 rows) per round
  int rank, world, n_pairs;
  uint32_t* dbg;               // pinned host words for the timeout records (test option), or null
  int probe;                   // timing probes (test option grs_probe; wrong results): 1 no TMA loads, 2 no UMMA, 4 no stores, 8 no reduce, 16 no L2 hints
};

struct __align__(1024) GrsSmem {
  uint8_t ab[grs::kStages][grs::kStageBytes];
  uint8_t out[2][grs::kBoxBytes];
  uint64_t full[grs::kStages], empty[grs::kStages];  // full: the LEADER's counts both CTAs' bytes
  uint64_t acc_full[2], acc_empty[2];
  uint32_t tmem_base, epoch;
};

__device__ __forceinline__ void grs_st_flag(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t grs_ld_flag(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void grs_fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
// Every wait of this kernel is bounded by TIME (%globaltimer; 4 s for a pair's own barriers, 20 s for flags other ranks write): a protocol bug or a missing peer ends in a trap,
// not a hung GPU.  With the `grs_debug_host_ptr` test option (a pinned host buffer, 64 words per CTA) each wait that
// gave up leaves {code, thread, a, b, observed} at word 8 * code; codes 1 producer/empty, 2 issuer/acc_empty, 3 issuer/full,
// 4 epilogue/acc_full, 5 STARTED, 6 round counter, 7 DONE.
__device__ __forceinline__ uint64_t grs_now() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __noinline__ void grs_fail(uint32_t* dbg, uint32_t code, uint32_t a, uint32_t b, uint32_t seen) {
  if (dbg) {
    volatile uint32_t* d = dbg + 64 * blockIdx.x + 8 * code;
    d[1] = threadIdx.x; d[2] = a; d[3] = b; d[4] = seen; d[0] = code;
    __threadfence_system();
    // let the other stuck waits of the grid write their records before the trap kills it
    const uint64_t t0 = grs_now();
    while (grs_now() - t0 < 200000000ull) __nanosleep(1000000);
  }
  __trap();
}
constexpr uint64_t kGrsTimeoutNs = 4000000000ull;        // shared-memory barriers: this CTA pair only
constexpr uint64_t kGrsPeerTimeoutNs = 20000000000ull;   // flags written by other ranks (launch skew included)
__device__ __forceinline__ void grs_mbar_wait(uint64_t* bar, uint32_t parity, uint32_t* dbg, uint32_t code, uint32_t a,
                                              uint32_t b) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = grs_now();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity))
    if ((++spins & 15) == 0 && grs_now() - t0 > kGrsTimeoutNs) grs_fail(dbg, code, a, b, 0);
}
__device__ __forceinline__ void grs_mbar_wait_cluster(uint64_t* bar, uint32_t parity, uint32_t* dbg, uint32_t code,
                                                      uint32_t a, uint32_t b) {
  const uint64_t t0 = grs_now();
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
    if ((++spins & 15) == 0 && grs_now() - t0 > kGrsTimeoutNs) grs_fail(dbg, code, a, b, 0);
  }
}
// spin on a word of THIS rank's memory until it reaches `want` (peers write it)
__device__ __forceinline__ void grs_wait_ge(const uint32_t* p, uint32_t want, uint32_t* dbg, uint32_t code, uint32_t a) {
  uint32_t seen = grs_ld_flag(p);
  if ((int32_t)(seen - want) >= 0) return;
  const uint64_t t0 = grs_now();
  uint32_t spins = 0;
  while ((int32_t)((seen = grs_ld_flag(p)) - want) < 0) {
    __nanosleep(100);
    if ((++spins & 255) == 0 && grs_now() - t0 > kGrsPeerTimeoutNs) grs_fail(dbg, code, a, want, seen);
  }
}
// A slot line is written once per call (by a TMA store, maybe a peer's) and read once, after the acquire on its
// round's counter (ordered by that acquire and the reducers' barrier): streaming policy, so that the GEMM's operands
// stay in L2, not the slots.
__device__ __forceinline__ uint4 grs_ld_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void grs_st_v4(void* p, const uint4& v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// job j of this rank -> pair tile.  Order: round c, then owner (rank+1, rank+2, ..., rank), then the round's pair
// tiles of that owner, columns fastest.
struct GrsTile {
  int m_pair0, n0, owner, c;
};
__device__ __forceinline__ GrsTile grs_decode(const GrsParams& p, int j) {
  const int per_owner = p.mc * p.NT, per_round = p.world * per_owner;
  GrsTile t;
  t.c = j / per_round;
  int rem = j - t.c * per_round;
  const int s = rem / per_owner;
  rem -= s * per_owner;
  const int i = rem / p.NT, n = rem - i * p.NT;
  t.owner = p.rank + 1 + s;
  if (t.owner >= p.world) t.owner -= p.world;
  t.m_pair0 = t.owner * p.Mo + t.c * p.R + i * 256;
  t.n0 = n * grs::kBN;
  return t;
}

// Vectors [lo, hi) of my owner block (16-byte units): sum the W slots in rank order, store to y on every rank.
// U vectors per thread per trip with all U x W loads issued before the first use (W x U = 8 loads in flight).
template <bool BF16, int W, int U>
__device__ __forceinline__ void grs_reduce_span(const GrsParams& p, const char* slots, int64_t slot_stride,
                                                int64_t y_row0, int64_t lo, int64_t hi, int rt) {
  const int64_t stride = (int64_t)gridDim.x * grs::kReducers;
  for (int64_t i0 = lo + (int64_t)blockIdx.x * grs::kReducers + rt; i0 < hi; i0 += stride * U) {
    uint4 v[U][W];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int r = 0; r < W; ++r)
        if (r < p.world && i0 + u * stride < hi) v[u][r] = grs_ld_v4(slots + r * slot_stride + (i0 + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      if (i >= hi) break;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int r = 0; r < W; ++r) {
        if (r < p.world) {  // rank order on every rank: identical bits everywhere
          const uint32_t w[4] = {v[u][r].x, v[u][r].y, v[u][r].z, v[u][r].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = unpack2<BF16>(w[e]);
            acc[2 * e] += f.x;
            acc[2 * e + 1] += f.y;
          }
        }
      }
      uint4 o;
      o.x = pack2<BF16>(acc[0], acc[1]); o.y = pack2<BF16>(acc[2], acc[3]);
      o.z = pack2<BF16>(acc[4], acc[5]); o.w = pack2<BF16>(acc[6], acc[7]);
#pragma unroll
      for (int r = 0; r < W; ++r)
        if (r < p.world) grs_st_v4(reinterpret_cast<uint4*>(reinterpret_cast<char*>(p.base[r]) + p.y_off) + y_row0 + i, o);
    }
  }
}

template <bool BF16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(grs::kThreads, 1)
gemm_reduce_kernel(const __grid_constant__ GrsMaps maps, const GrsParams p) {
  using namespace grs;
  extern __shared__ unsigned char smem_raw[];
  GrsSmem& sm = *reinterpret_cast<GrsSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t cr = cluster_ctarank();
  const int pair = (int)(blockIdx.x >> 1);
  uint32_t* flags = reinterpret_cast<uint32_t*>(p.base[p.rank]);

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&sm.acc_full[t], 1);
      mbar_init(&sm.acc_empty[t], 8);  // four epilogue warps of each CTA of the pair
    }
    mbar_fence_init();
    sm.epoch = *reinterpret_cast<volatile uint32_t*>(flags + kWEpoch) + 1;  // written by the last CTA of the previous call
  }
  if (warp == 3) tmem_alloc_pair(&sm.tmem_base, 512);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a);
    tma_prefetch_desc(&maps.b);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers exist before anyone arrives remotely
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const uint32_t epoch = sm.epoch;
  const int jobs = p.C * p.world * p.mc * p.NT;

  if (warp == 0) {
    // ===================== TMA producer: my 128 rows of x, my half of the W tile =====================
    // The bytes of BOTH CTAs are counted on the leader's `full` barrier (cp.async.bulk.tensor.cta_group::2): the
    // issuer's single wait covers the pair's four boxes; a forwarding hop through a warp of each CTA cost a
    // release.cluster arrive per stage (~540 clk, serialised: as long as the stage's four UMMAs).
    if (lane == 0 && !((p.probe & 1) && cr != 0)) {
      const uint64_t pol_a = (p.probe & 16) ? kL2EvictNormal : kL2EvictFirst;
      const uint64_t pol_b = (p.probe & 16) ? kL2EvictNormal : kL2EvictLast;
      int n_it = 0;
      for (int j = pair; j < jobs; j += p.n_pairs) {
        const GrsTile t = grs_decode(p, j);
        if (t.m_pair0 >= p.M) continue;
        for (int ks = 0; ks < p.KS; ++ks, ++n_it) {
          const int s = n_it % kStages;
          grs_mbar_wait(&sm.empty[s], ((n_it / kStages) & 1) ^ 1, p.dbg, 1, n_it, j);
          if (p.probe & 1) {  // the leader free-runs; its peer is not part of the handshake and must not follow it
            mbar_arrive(&sm.full[s]);  // (a follower that falls a lap behind misses a phase flip and waits for ever)
            continue;
          }
          if (cr == 0) mbar_expect_tx(&sm.full[s], 2 * kStageBytes);
          const uint32_t bar = mapa_u32(&sm.full[s], 0);
          // x rows are used by one sweep over the column tiles and never again; W is re-read by every row tile
          tma_load_2d_pair_hint(sm.ab[s], &maps.a, bar, ks * kBK, t.m_pair0 + kBM * (int)cr, pol_a);
          tma_load_2d_pair_hint(sm.ab[s] + kABytes, &maps.b, bar, ks * kBK, t.n0 + (kBN / 2) * (int)cr, pol_b);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer (leader CTA only) =====================
    // The whole warp walks the schedule (every value below is warp-uniform: kernel parameters, blockIdx, loop counters,
    // the TMEM base shuffled from lane 0) and lane 0 only executes the tcgen05 instructions: the compiler then keeps
    // addresses and descriptors on the uniform datapath.  Computed inside a one-lane region they live in vector
    // registers and every UMMA is preceded by five R2UR moves inside an ELECT loop (21 SASS instructions per UMMA; the
    // same finding as the prefill issuer, tools/trace_prefill.py).  A descriptor is its stage's base descriptor plus a
    // constant (the 14-bit start-address field cannot carry: shared memory ends below 256 KB).
    if (cr == 0) {
      constexpr uint32_t idesc = umma_idesc(BF16, 2 * kBM, kBN, false, false);
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
      const uint32_t ab0 = smem_u32(sm.ab[0]);
      int n_it = 0, t_it = 0;
      for (int j = pair; j < jobs; j += p.n_pairs) {
        const GrsTile t = grs_decode(p, j);
        if (t.m_pair0 >= p.M) continue;
        const int buf = t_it & 1;
        grs_mbar_wait_cluster(&sm.acc_empty[buf], ((t_it >> 1) & 1) ^ 1, p.dbg, 2, t_it, j);
        tc_fence_after();
        const uint32_t d_tmem = tmem_u + buf * kBN;
        for (int ks = 0; ks < p.KS; ++ks, ++n_it) {
          const int s = n_it % kStages;
          grs_mbar_wait(&sm.full[s], (n_it / kStages) & 1, p.dbg, 3, n_it, j);
          tc_fence_after();
          if (lane == 0) {
            const uint64_t a_desc = umma_smem_desc(ab0 + s * kStageBytes, 16, 1024);
            const uint64_t hi = a_desc & 0xffffffff00000000ull;
            const uint32_t a_lo = (uint32_t)a_desc, b_lo = a_lo + (kABytes >> 4);
            if (!(p.probe & 2))
#pragma unroll
            for (int kk = 0; kk < kBK / 16; ++kk)
              umma2_ss(d_tmem, hi | (uint64_t)(a_lo + 2 * kk), hi | (uint64_t)(b_lo + 2 * kk), idesc,
                       (ks > 0 || kk > 0) ? 1u : 0u);
            umma2_commit(&sm.empty[s]);  // both CTAs' stage s may be refilled
          }
          __syncwarp();
        }
        if (lane == 0) umma2_commit(&sm.acc_full[buf]);
        __syncwarp();
        ++t_it;
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== epilogue: TMEM -> 16 bit -> shared -> TMA store into the owner's slot =====================
    const int q = warp & 3, row = 32 * q + lane;
    const bool is_e = tid == 128;
    const uint32_t lane_base = (uint32_t)(32 * q) << 16;
    int t_it = 0;
    uint32_t n_box = 0;
    uint32_t* pending = nullptr;  // counter of the last stored tile, not yet signalled
    for (int j = pair; j < jobs; j += p.n_pairs) {
      const GrsTile t = grs_decode(p, j);
      if (t.m_pair0 >= p.M) continue;
      const int buf = t_it & 1;
      grs_mbar_wait(&sm.acc_full[buf], (t_it >> 1) & 1, p.dbg, 4, t_it, j);
      tc_fence_after();
      const int m0 = t.m_pair0 + kBM * (int)cr;
      const bool valid = m0 < p.M && !(p.probe & 4);  // the upper CTA of the last pair tile may hold only padding rows
      const int dest_row = p.rank * p.Mo + (m0 - t.owner * p.Mo);
#pragma unroll 1
      for (int ch = 0; ch < 4; ++ch) {
        uint8_t* box = sm.out[n_box & 1];
        if (valid) {
          if (is_e) bulk_wait_read<1>();  // the store that last read this box has drained it
          asm volatile("bar.sync 1, 128;" ::: "memory");
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t v[32];
          __syncwarp();
          tmem_ld32(tmem + lane_base + buf * kBN + ch * 64 + h * 32, v);
          tc_wait_ld();
          if (valid) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              uint4 o;
              o.x = pack2<BF16>(__uint_as_float(v[8 * k + 0]), __uint_as_float(v[8 * k + 1]));
              o.y = pack2<BF16>(__uint_as_float(v[8 * k + 2]), __uint_as_float(v[8 * k + 3]));
              o.z = pack2<BF16>(__uint_as_float(v[8 * k + 4]), __uint_as_float(v[8 * k + 5]));
              o.w = pack2<BF16>(__uint_as_float(v[8 * k + 6]), __uint_as_float(v[8 * k + 7]));
              *reinterpret_cast<uint4*>(box + row * 128 + (((4 * h + k) ^ (row & 7)) << 4)) = o;
            }
          }
        }
        if (ch == 3) {  // the accumulator is in registers / shared memory: hand the TMEM buffer back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&sm.acc_empty[buf], 0);
        }
        if (valid) {
          fence_proxy_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (is_e) {
            tma_store_2d(&maps.slot[t.owner], box, t.n0 + ch * 64, dest_row);
            bulk_commit();
          }
          ++n_box;
        }
      }
      if (valid && is_e) {
        // this tile's four groups are in flight; the previous tile's are complete once at most four are pending
        if (pending) {
          bulk_wait<4>();
          asm volatile("fence.proxy.async.global;" ::: "memory");
          grs_fence_sys();
          asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(pending) : "memory");
        }
        pending = reinterpret_cast<uint32_t*>(p.base[t.owner]) + kWCnt + t.c * 8 + p.rank;
      }
      ++t_it;
    }
    if (is_e && pending) {
      bulk_wait<0>();
      asm volatile("fence.proxy.async.global;" ::: "memory");
      grs_fence_sys();
      asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(pending) : "memory");
    }
  } else if (warp >= 8) {
    // ===================== reducer: sum the slots of my rows in rank order, all-gather by peer stores =====================
    const int rt = tid - 256;
    if (blockIdx.x == 0 && rt < p.world) {  // STARTED: my stream has finished with y of the previous call
      grs_fence_sys();
      grs_st_flag(reinterpret_cast<uint32_t*>(p.base[rt]) + kWStarted + p.rank, epoch);
    }
    if (rt < p.world) grs_wait_ge(flags + kWStarted + rt, epoch, p.dbg, 5, rt);
    asm volatile("bar.sync 2, 256;" ::: "memory");
    const int my_rows = max(0, min(p.M - p.rank * p.Mo, p.Mo));
    const int64_t row_vecs = p.N / 8;
    const char* slots = reinterpret_cast<const char*>(p.base[p.rank]) + p.slot_off;
    const int64_t slot_stride = (int64_t)p.Mo * p.N * 2;
    const int64_t y_row0 = (int64_t)p.rank * p.Mo * row_vecs;
    for (int c = 0; c < p.C; ++c) {
      const int r_lo = c * p.R;
      if (r_lo >= my_rows || (p.probe & 12)) break;
      const int r_hi = min(r_lo + p.R, my_rows);
      const uint32_t expected = (uint32_t)((r_hi - r_lo + kBM - 1) / kBM) * (uint32_t)p.NT;
      if (rt < p.world) grs_wait_ge(flags + kWCnt + c * 8 + rt, expected, p.dbg, 6, c);
      asm volatile("bar.sync 2, 256;" ::: "memory");
      const int64_t lo = r_lo * row_vecs, hi = r_hi * row_vecs;
      if (p.world <= 2) grs_reduce_span<BF16, 2, 4>(p, slots, slot_stride, y_row0, lo, hi, rt);
      else if (p.world <= 4) grs_reduce_span<BF16, 4, 2>(p, slots, slot_stride, y_row0, lo, hi, rt);
      else grs_reduce_span<BF16, 8, 1>(p, slots, slot_stride, y_row0, lo, hi, rt);
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");
    if (rt == 0) {
      grs_fence_sys();
      uint32_t old;
      asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(flags + kWFin) : "memory");
      if (old == gridDim.x - 1) {  // every CTA of this rank has read its slots and issued its stores
        for (int w = 0; w < p.C * 8; ++w) *reinterpret_cast<volatile uint32_t*>(flags + kWCnt + w) = 0;
        *reinterpret_cast<volatile uint32_t*>(flags + kWFin) = 0;
        *reinterpret_cast<volatile uint32_t*>(flags + kWEpoch) = epoch;
        grs_fence_sys();
        for (int r = 0; r < p.world; ++r)
          grs_st_flag(reinterpret_cast<uint32_t*>(p.base[r]) + kWDone + p.rank, epoch);
      }
    }
    if (rt < p.world) grs_wait_ge(flags + kWDone + rt, epoch, p.dbg, 7, rt);  // every peer's stores into my y are performed
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 3) tmem_dealloc_pair(tmem, 512);
}

// Rows per owner / per round for an [M, N] output over `world` ranks (same arithmetic on every rank).
static void grs_layout(int64_t M, int64_t N, int world, int rows_per_round, int* Mo, int* R, int* C, int64_t* slot_bytes,
                       int64_t* y_off) {
  const int64_t per = (M + world - 1) / world;
  int64_t r = rows_per_round > 0 ? rows_per_round : 256;
  r = (r + 255) / 256 * 256;
  while ((per + r - 1) / r > grs::kMaxRounds) r += 256;
  int64_t c = (per + r - 1) / r;
  if (c < 1) c = 1;
  *R = (int)r;
  *C = (int)c;
  *Mo = (int)(c * r);
  *slot_bytes = (int64_t)world * (c * r) * N * 2;
  *y_off = (int64_t)grs::kHeaderBytes + ((*slot_bytes + 1023) / 1024) * 1024;
}

}  // namespace cvllm

using namespace cvllm;

// Payload bytes (after the flag header of cvllm_ipc_alloc) a rank's buffer needs for an [M, N] 16-bit output.
extern "C" int64_t cvllm_gemm_allreduce_payload_bytes(int64_t M, int64_t N, int world) {
  if (M <= 0 || N <= 0 || world < 1 || world > grs::kMaxWorld) return -1;
  int Mo, R, C;
  int64_t slot_bytes, y_off;
  grs_layout(M, N, world, test_options().grs_rows_per_round.load(), &Mo, &R, &C, &slot_bytes, &y_off);
  return y_off - (int64_t)grs::kHeaderBytes + M * N * 2;
}

// y = sum_r x_r . W_r^T over the `world` ranks, left in EVERY rank's buffer; *y_out_host = where (inside
// peer_bases_host[rank]).  x: [M, K] (row stride lda elements), w: [N, K] (row stride ldw) — this rank's shards.
// Every rank must call with the same M, N, dtype in the same order; K may differ per rank.
extern "C" int cvllm_gemm_allreduce(const void* x, int64_t lda, const void* w, int64_t ldw, int64_t M, int64_t N,
                                    int64_t K, void* const* peer_bases_host, int rank, int world, int dtype,
                                    int64_t payload_capacity_bytes, void** y_out_host, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(x && w && peer_bases_host && y_out_host);
  CVLLM_CHECK_ARG(world >= 1 && world <= grs::kMaxWorld && rank >= 0 && rank < world);
  CVLLM_CHECK_ARG(M > 0 && N > 0 && K > 0 && M < (1ll << 30) && N < (1ll << 30) && K < (1ll << 30));
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_SUPPORTED(N % 8 == 0 && K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K);
  CVLLM_CHECK_ALIGN(aligned16(x) && aligned16(w));
  GrsParams p;
  int64_t slot_bytes, y_off;
  grs_layout(M, N, world, test_options().grs_rows_per_round.load(), &p.Mo, &p.R, &p.C, &slot_bytes, &y_off);
  CVLLM_CHECK_ARG(y_off - (int64_t)grs::kHeaderBytes + M * N * 2 <= payload_capacity_bytes);
  for (int r = 0; r < grs::kMaxWorld; ++r) p.base[r] = r < world ? peer_bases_host[r] : nullptr;
  for (int r = 0; r < world; ++r) CVLLM_CHECK_ARG(p.base[r] != nullptr);
  p.slot_off = (int64_t)grs::kHeaderBytes;
  p.y_off = y_off;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.NT = (int)((N + grs::kBN - 1) / grs::kBN);
  p.KS = (int)((K + grs::kBK - 1) / grs::kBK);
  p.mc = p.R / 256;
  p.rank = rank; p.world = world;
  int pairs = sm_count() / 2;
  const int cap = test_options().grs_pairs.load();
  if (cap > 0 && cap < pairs) pairs = cap;
  p.n_pairs = pairs;
  p.probe = test_options().grs_probe.load();
  p.dbg = reinterpret_cast<uint32_t*>(static_cast<uintptr_t>(test_options().grs_debug_host_ptr.load()));

  GrsMaps maps;
  const bool bf16 = dtype == CVLLM_BF16;
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M}, str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, 128};
    int rc = make_tmap_16b(&maps.a, x, 2, dims, str, box, bf16);
    if (rc != CVLLM_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N}, str[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {64, 128};
    int rc = make_tmap_16b(&maps.b, w, 2, dims, str, box, bf16);
    if (rc != CVLLM_OK) return rc;
  }
  for (int r = 0; r < grs::kMaxWorld; ++r) {
    const int q = r < world ? r : 0;  // unused entries repeat a valid map
    const uint64_t dims[2] = {(uint64_t)N, (uint64_t)world * (uint64_t)p.Mo}, str[1] = {(uint64_t)N * 2};
    const uint32_t box[2] = {64, 128};
    int rc = make_tmap_16b(&maps.slot[r], static_cast<char*>(p.base[q]) + p.slot_off, 2, dims, str, box, bf16);
    if (rc != CVLLM_OK) return rc;
  }
  const int smem = (int)sizeof(GrsSmem) + 1024;
  static std::atomic<uint64_t> attr_bf16{0}, attr_f16{0};
  if (bf16) CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)gemm_reduce_kernel<true>, smem, attr_bf16));
  else CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)gemm_reduce_kernel<false>, smem, attr_f16));
  auto st = static_cast<cudaStream_t>(stream);
  if (bf16) gemm_reduce_kernel<true><<<2 * pairs, grs::kThreads, smem, st>>>(maps, p);
  else gemm_reduce_kernel<false><<<2 * pairs, grs::kThreads, smem, st>>>(maps, p);
  CVLLM_CHECK_LAUNCH();
  *y_out_host = static_cast<char*>(p.base[rank]) + p.y_off;
  return CVLLM_OK;
}
