// Head-sparse split-KV decode attention over the paged cache.
//
// Replaces attention/sparse_decode_kernel.py:10-165 (stage 1 :247-388, stage 2
// :392-435) with ONE persistent launch, one CTA per SM:
//   * work = 128-token units of every (b, kv-head); the flattened unit list
//     is cut into gridDim.x equal ranges (stream-K style), so ragged per-head
//     lengths after compression are balanced exactly — the reference gives
//     every head the same KEY_SPLIT regardless of its length (:303-306)
//   * warp 8 is the TMA producer: it resolves (head, page) for 32 units at a
//     time (one lane per unit/page piece), then streams K and V units through
//     a ring of kSlots x 32 KB shared-memory slots (cp.async.bulk.tensor, 128B
//     swizzle, mbarrier full/empty pairs).  Up to 160 KB per SM are in flight,
//     which is what HBM3e needs (Little: 6.5 TB/s x ~1 us / 148 SMs ~ 45 KB)
//     with a wide margin; the previous register-staged version peaked at 64 KB
//     and measured 4.5 TB/s
//   * warps 0-7 consume: a warp owns 16 keys of a unit, reads its K rows and V
//     rows from the swizzled slot as conflict-free 16-byte vectors laid out
//     directly in mma.sync m16n8k16 fragment order (the contraction dim is
//     permuted identically on both operands, so no ldmatrix is needed)
//   * S^T = K.Q^T (keys on M, the G<=8 query heads on N), softmax stats per
//     column with warp shuffles, P^T goes through movmatrix to the B-operand
//     layout, O^T += V^T.P^T
//   * partial (m, l, O) of a range that does not cover its whole head goes to
//     a scratch slot; the CTA that completes a head merges (atomic ticket),
//     every warp folding a strided subset of the slots with independent loads
// HBM-bound: 4*D bytes per cached (token, head) — see DESIGN.md.
#include <stdlib.h>

#include "common.cuh"
#include "mma.cuh"
#include "tma.cuh"

namespace cvllm {

constexpr int kUnit = 128;        // tokens per work unit
constexpr int kDecWarps = 8;      // consumer warps: 8 x 16 keys = one unit
constexpr int kDecCons = kDecWarps * 32;
constexpr int kDecThreads = kDecCons + 32;  // + producer warp
constexpr int kMaxSlots = 6;      // ring slots (one K unit or one V unit each): 6 when G <= 4, else 5
constexpr int kSlotBytes = kUnit * 256;     // 32 KB at D = 128, 16-bit
constexpr int kHalfBytes = kUnit * 128;     // one 64-column half of a slot

struct DecodeParams {
  const void* q;
  void* out;
  const int32_t* seq_lens_bh;
  const int32_t* page_table;
  const int32_t* batch_mapping;
  float* scratch;     // [(BH + C)] slots of G*(D+2) floats
  int32_t* counters;  // [BH], zero on entry, left zero on exit
  int B, HQ, HKV, G, PS, NLP;
  int box_rows;       // min(PS, 128): rows of one TMA box (a box never crosses a page)
  int slots;          // ring depth (<= kMaxSlots)
  int lens_by_slot;   // seq_lens_bh is the cache's [Bmax+1, HKV] table, indexed through batch_mapping
  int mode;           // 0 = normal; 1 = consumers only drain the ring (load ceiling, profiling only)
  float scale_log2;
#ifdef CVLLM_DECODE_TRACE
  unsigned long long* trace;  // profiling build only (tools/trace_decode.py): 8 globaltimer stamps per CTA
#endif
};

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// The stamps cost ~2 us at B = 1 (measured), so they exist only in the -DCVLLM_DECODE_TRACE variant build.
#ifdef CVLLM_DECODE_TRACE
#define DEC_TRACE(slot) do { if (p.trace) p.trace[blockIdx.x * 8 + (slot)] = gtime(); } while (0)
#else
#define DEC_TRACE(slot) do { } while (0)
#endif

// One warp's 16-key slice: K rows r / r+8 (4 chunks each), V for 4 keys x 2 chunks.
struct Slice {
  uint4 k_lo[4], k_hi[4];  // rows r, r+8; 16-byte chunk kchunk(j, c)
  uint4 v[4][2];           // keys {2j,2j+1,2j+8,2j+9}; chunks r (d 8r..8r+7), 8+r
};

// Lane j owns the 16-byte chunks {2j, 2j+1} of each 128-byte half of a row; c = 0..3 walks them.
// With the 128B swizzle (chunk ^= row & 7) the eight lanes of a quarter-warp (rows r0, r0+1 x
// j = 0..3) then hit eight distinct bank groups.  Q uses the same chunk order, so the
// contraction is unchanged.
__device__ __forceinline__ int kchunk(int j, int c) { return ((c >> 1) << 3) + 2 * j + (c & 1); }

__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}
// address of 16-byte chunk `ch` (0..15) of row t (0..127) inside a swizzled slot
__device__ __forceinline__ uint32_t slot_addr(uint32_t slot, int t, int ch) {
  return slot + (uint32_t)((ch >> 3) * kHalfBytes + t * 128 + (((ch & 7) ^ (t & 7)) << 4));
}

__device__ __forceinline__ void cons_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kDecCons) : "memory"); }

template <bool BF16, int D>
__global__ void __launch_bounds__(kDecThreads, 1)
decode_attn_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                   const DecodeParams p) {
  static_assert(D == 128, "decode kernel is specialised for head_dim 128");
  constexpr int KSTEPS = D / 16;  // 8
  constexpr int MT = D / 16;      // 8 output m-tiles (16 d each)
  constexpr int ROW = D + 2;      // merged row: O[D], m, l
  constexpr int NW = kDecThreads / 32;

  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem_raw =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
  const int BH = p.B * p.HKV;
  unsigned char* ring = smem_raw;
  const int kSlots = p.slots;
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem_raw + kSlots * kSlotBytes);
  uint64_t* bar_empty = bar_full + kMaxSlots;
  float* s_merge = reinterpret_cast<float*>(bar_empty + kMaxSlots);   // [kDecWarps][G][ROW]
  float* s_final = s_merge + kDecWarps * p.G * ROW;                   // [G][ROW]
  int32_t* s_prefix = reinterpret_cast<int32_t*>(s_final + p.G * ROW);  // [BH + 1]
  __shared__ int32_t s_warp_tot[NW];
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int C = gridDim.x, c = blockIdx.x;
  const int G = p.G;
  // rows cached for flattened head i = b * HKV + h
  auto len_of = [&](int i) -> int {
    if (!p.lens_by_slot) return p.seq_lens_bh[i];
    const int b = i / p.HKV;
    return p.seq_lens_bh[p.batch_mapping[b] * p.HKV + (i - b * p.HKV)];
  };

  if (tid == 0) DEC_TRACE(0);
  if (tid == kDecCons) {  // producer warp: descriptor fetch overlaps the prologue
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
  }
  if (tid == 0) {
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], kDecWarps);
    }
    mbar_fence_init();
  }

  // ---- phase 0: units per head, exclusive prefix over the BH heads ---------
  {
    const int per = cdiv(BH, kDecThreads);
    const int beg = min(BH, tid * per), end = min(BH, beg + per);
    int local = 0;
    for (int i = beg; i < end; ++i) local += cdiv(len_of(i), kUnit);
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
      run += cdiv(len_of(i), kUnit);
    }
    if (tid == kDecThreads - 1) s_prefix[BH] = run;
    __syncthreads();
  }
  const int U = s_prefix[BH];
  if (tid == 0) DEC_TRACE(1);

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

  // largest i with s_prefix[i] <= u
  auto find_head = [&](int u) {
    int lo = 0, hi = BH;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_prefix[mid] <= u) lo = mid; else hi = mid;
    }
    return lo;
  };

  // =========================== producer warp ================================
  if (warp == kDecWarps) {
    const int BR = p.box_rows;
    const int pieces = kUnit / BR;      // TMA boxes per (unit, half): 1 when PS % 128 == 0
    const int upb = 32 / pieces;        // units resolved per batch (one lane per piece)
    int p_slot = 0;                     // ring position of the next item (K and V units alternate)
    uint32_t p_par = 0;                 // and its phase parity
    // first cache row of this lane's (unit, piece) of the batch starting at unit ub
    auto resolve = [&](int ub, int& row, int& nvp) {
      const int my_unit = ub + lane / pieces, my_piece = lane % pieces;
      row = 0;
      nvp = 0;
      if (my_unit < u_end) {
        const int head = find_head(my_unit);
        const int t0 = (my_unit - s_prefix[head]) * kUnit;
        const int L = len_of(head);
        nvp = min(pieces, cdiv(L - t0, BR));  // pages past the head's length are not mapped
        if (my_piece < nvp) {
          const int b = head / p.HKV, h = head % p.HKV;
          const int tok = t0 + my_piece * BR;
          const int pg = tok / p.PS;
          const int32_t* pt = p.page_table + ((int64_t)p.batch_mapping[b] * p.HKV + h) * p.NLP;
          row = pt[pg] * p.PS + (tok - pg * p.PS);
        }
      }
    };
    int row, nvp;
    resolve(u_begin, row, nvp);
    for (int ub = u_begin; ub < u_end; ub += upb) {
      // the next batch's page-table loads are in flight while this batch is issued
      int row_next = 0, nvp_next = 0;
      if (ub + upb < u_end) resolve(ub + upb, row_next, nvp_next);
      const int cnt = min(upb, u_end - ub);
      for (int i = 0; i < cnt; ++i) {
        const int nv = __shfl_sync(0xffffffffu, nvp, i * pieces);
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {
          const int slot = p_slot;
          const uint32_t par = p_par;
          if (++p_slot == kSlots) { p_slot = 0; p_par ^= 1u; }
          unsigned char* dst = ring + slot * kSlotBytes;
          const CUtensorMap* tm = kv ? &tm_v : &tm_k;
          if (lane == 0) {
            mbar_wait(&bar_empty[slot], par ^ 1u);
            mbar_expect_tx(&bar_full[slot], (uint32_t)(nv * BR * 256));
          }
          for (int pc = 0; pc < nv; ++pc) {
            const int r0 = __shfl_sync(0xffffffffu, row, i * pieces + pc);
            if (lane == 0) {
              tma_load_2d(dst + pc * BR * 128, tm, &bar_full[slot], 0, r0);
              tma_load_2d(dst + kHalfBytes + pc * BR * 128, tm, &bar_full[slot], 64, r0);
            }
          }
        }
      }
      row = row_next;
      nvp = nvp_next;
    }
    return;
  }

  // =========================== consumer warps ===============================
  const int r = lane >> 2, j = lane & 3;
  const uint32_t ring_u32 = smem_u32(ring);
  int c_slot = 0;        // ring position of the next item to consume
  uint32_t c_par = 0;    // and its mbarrier phase parity
  int head = find_head(u_begin);
  int u = u_begin;
  while (u < u_end) {
    while (s_prefix[head + 1] <= u) ++head;  // skip finished / empty heads
    const int h_beg = s_prefix[head], h_end = s_prefix[head + 1];
    const int seg_end = min(u_end, h_end);
    const int ua = u - h_beg, ub = seg_end - h_beg;  // units of this head
    const int b = head / p.HKV, h = head % p.HKV;
    const int L = len_of(head);

    // ---- Q^T as B operand (n = query r, k = permuted d) ----
    uint32_t qb[KSTEPS][2];
    {
      const uint4 z = make_uint4(0, 0, 0, 0);
      const char* qrow = reinterpret_cast<const char*>(p.q) + ((int64_t)b * p.HQ + h * G + r) * (D * 2);
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const uint4 v = (r < G) ? *reinterpret_cast<const uint4*>(qrow + kchunk(j, cc) * 16) : z;
        qb[2 * cc][0] = v.x; qb[2 * cc][1] = v.y;
        qb[2 * cc + 1][0] = v.z; qb[2 * cc + 1][1] = v.w;
      }
    }

    float acc[MT][4];
#pragma unroll
    for (int t = 0; t < MT; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    for (int uu = ua; uu < ub; ++uu) {
      // ring positions of this unit's K and V items (advanced without divisions)
      const int ks_slot = c_slot;
      const uint32_t k_par = c_par;
      if (++c_slot == kSlots) { c_slot = 0; c_par ^= 1u; }
      const int vs_slot = c_slot;
      const uint32_t v_par = c_par;
      if (++c_slot == kSlots) { c_slot = 0; c_par ^= 1u; }
      const uint32_t k_base = ring_u32 + ks_slot * kSlotBytes, v_base = ring_u32 + vs_slot * kSlotBytes;
      const int t0 = warp * 16;                               // this warp's rows inside the unit
      const int n_valid = min(16, L - (uu * kUnit + t0));     // <= 0: nothing for this warp

      // every warp waits (so nobody can run a full ring ahead) and releases the slots
      mbar_wait(&bar_full[ks_slot], k_par);
      if (tid == 0 && u == u_begin && uu == ua) DEC_TRACE(2);
      if (n_valid > 0 && p.mode == 0) {
        Slice cur;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          cur.k_lo[cc] = lds_v4(slot_addr(k_base, t0 + r, kchunk(j, cc)));
          cur.k_hi[cc] = lds_v4(slot_addr(k_base, t0 + r + 8, kchunk(j, cc)));
        }
        // V rows of keys past the head's length are uninitialised memory: P is 0 there but
        // 0 * NaN would poison O, so those rows are read as zeros.  V is fetched before the
        // S chain so its shared-memory latency hides behind the mma / softmax dependency chain.
        mbar_wait(&bar_full[vs_slot], v_par);
        {
          const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int key = 2 * j + (i & 1) + ((i >> 1) << 3);
            const bool ok = key < n_valid;
            cur.v[i][0] = ok ? lds_v4(slot_addr(v_base, t0 + key, r)) : z;
            cur.v[i][1] = ok ? lds_v4(slot_addr(v_base, t0 + key, 8 + r)) : z;
          }
        }
        // ---- S^T[16 keys x 8 queries]: two accumulators halve the dependent mma chain ----
        float s[4] = {0.f, 0.f, 0.f, 0.f}, s_b[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < KSTEPS; ++ks) {
          const int cc = ks >> 1;
          const uint32_t a0 = (ks & 1) ? cur.k_lo[cc].z : cur.k_lo[cc].x;
          const uint32_t a1 = (ks & 1) ? cur.k_hi[cc].z : cur.k_hi[cc].x;
          const uint32_t a2 = (ks & 1) ? cur.k_lo[cc].w : cur.k_lo[cc].y;
          const uint32_t a3 = (ks & 1) ? cur.k_hi[cc].w : cur.k_hi[cc].y;
          if (ks & 1) mma16816<BF16>(s_b, a0, a1, a2, a3, qb[ks][0], qb[ks][1]);
          else mma16816<BF16>(s, a0, a1, a2, a3, qb[ks][0], qb[ks][1]);
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) s[e] += s_b[e];
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_empty[ks_slot]);  // K rows are in registers: free the slot now
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
      } else {
        if (lane == 0) mbar_arrive(&bar_empty[ks_slot]);
        mbar_wait(&bar_full[vs_slot], v_par);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_empty[vs_slot]);
    }

    if (tid == 0) DEC_TRACE(3);
    // ---- per-warp state -> smem ----
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      l0 += __shfl_xor_sync(0xffffffffu, l0, o);
      l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    }
    {
      float* mw = s_merge + warp * G * ROW;
      const int q0 = 2 * j, q1 = 2 * j + 1;  // query columns of this lane; only q < G exist
      if (q0 < G) {
#pragma unroll
        for (int t = 0; t < MT; ++t) {
          mw[q0 * ROW + 8 * r + t] = acc[t][0];
          mw[q0 * ROW + 64 + 8 * r + t] = acc[t][2];
        }
        if (r == 0) { mw[q0 * ROW + D] = m0; mw[q0 * ROW + D + 1] = l0; }
      }
      if (q1 < G) {
#pragma unroll
        for (int t = 0; t < MT; ++t) {
          mw[q1 * ROW + 8 * r + t] = acc[t][1];
          mw[q1 * ROW + 64 + 8 * r + t] = acc[t][3];
        }
        if (r == 0) { mw[q1 * ROW + D] = m1; mw[q1 * ROW + D + 1] = l1; }
      }
    }
    cons_bar();

    // ---- merge the 8 warps' rows -> s_final[q][ROW] (unnormalised O, m, l) ----
    auto fold_warps = [&]() {
      for (int e = tid; e < G * ROW; e += kDecCons) {
        const int q = e / ROW, d = e - q * ROW;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kDecWarps; ++w) M = fmaxf(M, s_merge[(w * G + q) * ROW + D]);
        float val = 0.f;
        if (d == D) {
          val = M;
        } else {
          const int src = d < D ? d : D + 1;
#pragma unroll
          for (int w = 0; w < kDecWarps; ++w) {
            const float mw = s_merge[(w * G + q) * ROW + D];
            const float sc = (mw > -INFINITY) ? ex2(mw - M) : 0.f;
            val += s_merge[(w * G + q) * ROW + src] * sc;
          }
        }
        s_final[e] = val;
      }
    };
    auto write_out = [&]() {
      uint32_t* orow = reinterpret_cast<uint32_t*>(p.out) + ((int64_t)b * p.HQ + h * G) * (D / 2);
      for (int e = tid; e < G * D / 2; e += kDecCons) {
        const int q = e / (D / 2), d2 = e - q * (D / 2);
        const float lq = s_final[q * ROW + D + 1];
        const float inv = lq > 0.f ? 1.f / lq : 0.f;
        orow[e] = pack2<BF16>(s_final[q * ROW + 2 * d2] * inv, s_final[q * ROW + 2 * d2 + 1] * inv);
      }
    };
    fold_warps();
    cons_bar();

    const bool whole = (ua == 0) && (ub == h_end - h_beg);
    if (whole) {
      write_out();
    } else {
      // publish the partial, then take a ticket; the CTA completing the head merges
      float* slot = p.scratch + (int64_t)(head + c) * G * ROW;
      for (int e = tid; e < G * ROW; e += kDecCons) __stcg(slot + e, s_final[e]);
      __threadfence();
      cons_bar();
      if (tid == 0) {
        const int done = atomicAdd(p.counters + head, ub - ua) + (ub - ua);
        s_last = (done == h_end - h_beg);
      }
      cons_bar();
      if (tid == 0) DEC_TRACE(6);
      if (s_last) {
        __threadfence();
        const int c_first = (int)((((int64_t)h_beg + 1) * C - 1) / U);
        const int c_last = (int)(((int64_t)h_end * C - 1) / U);
        // Each warp folds the slots c_first + warp, + 8, ... into a running (m, l, O) with all
        // loads of a slot independent; lane owns d = 2*lane, 2*lane+1, 64 + 2*lane, 65 + 2*lane.
        float* mw = s_merge + warp * G * ROW;
        float M[8], ls[8];
        float2 oa[8], ob[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          M[q] = -INFINITY; ls[q] = 0.f;
          oa[q] = make_float2(0.f, 0.f); ob[q] = make_float2(0.f, 0.f);
        }
        for (int cc = c_first + warp; cc <= c_last; cc += kDecWarps) {
          // CTAs with an empty unit range (U < C) never wrote their slot
          if ((int)((int64_t)U * cc / C) >= (int)((int64_t)U * (cc + 1) / C)) continue;
          const float* sl = p.scratch + (int64_t)(head + cc) * G * ROW;
          float2 va[8], vb[8], ml[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (q < G) {
              va[q] = __ldcg(reinterpret_cast<const float2*>(sl + q * ROW + 2 * lane));
              vb[q] = __ldcg(reinterpret_cast<const float2*>(sl + q * ROW + 64 + 2 * lane));
              ml[q] = __ldcg(reinterpret_cast<const float2*>(sl + q * ROW + D));
            }
          }
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (q < G) {
              const float Mn = fmaxf(M[q], ml[q].x);
              const float so = (M[q] > -INFINITY) ? ex2(M[q] - Mn) : 0.f;
              const float sn = (ml[q].x > -INFINITY) ? ex2(ml[q].x - Mn) : 0.f;
              oa[q].x = oa[q].x * so + va[q].x * sn; oa[q].y = oa[q].y * so + va[q].y * sn;
              ob[q].x = ob[q].x * so + vb[q].x * sn; ob[q].y = ob[q].y * so + vb[q].y * sn;
              ls[q] = ls[q] * so + ml[q].y * sn;
              M[q] = Mn;
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < G) {
            *reinterpret_cast<float2*>(mw + q * ROW + 2 * lane) = oa[q];
            *reinterpret_cast<float2*>(mw + q * ROW + 64 + 2 * lane) = ob[q];
            if (lane == 0) { mw[q * ROW + D] = M[q]; mw[q * ROW + D + 1] = ls[q]; }
          }
        }
        cons_bar();
        if (tid == 0) DEC_TRACE(7);
        fold_warps();
        cons_bar();
        write_out();
        if (tid == 0) p.counters[head] = 0;
      }
    }
    cons_bar();  // s_merge / s_final reused by the next segment
#ifdef CVLLM_DECODE_TRACE
    if (tid == 0 && p.trace) {
      DEC_TRACE(4);
      p.trace[blockIdx.x * 8 + 5] = (unsigned long long)(whole ? 0 : (s_last ? 2 : 1));
    }
#endif
    u = seg_end;
  }
}

}  // namespace cvllm

using namespace cvllm;

static unsigned long long* g_decode_trace = nullptr;
// profiling only: device buffer of 8 x #CTAs uint64 that the next decode launches stamp (nullptr: off).
// Returns 0 when the library was built with -DCVLLM_DECODE_TRACE, -1 when the stamps are compiled out.
extern "C" int cvllm_debug_decode_trace(void* dev_buf) {
  g_decode_trace = reinterpret_cast<unsigned long long*>(dev_buf);
#ifdef CVLLM_DECODE_TRACE
  return 0;
#else
  return -1;
#endif
}

static int decode_ctas(int num_ctas) { return num_ctas > 0 ? num_ctas : sm_count(); }
static size_t decode_counter_bytes(size_t BH) { return (BH * 4 + 255) & ~(size_t)255; }

extern "C" size_t cvllm_decode_attention_workspace_bytes(int B, int HQ, int HKV, int D, int num_ctas) {
  if (B <= 0 || HKV <= 0 || HQ <= 0) return 0;
  const int G = HQ / HKV;
  const size_t BH = (size_t)B * HKV;
  const size_t slots = BH + (size_t)decode_ctas(num_ctas);
  return decode_counter_bytes(BH) + slots * G * (D + 2) * sizeof(float);
}

static int launch_decode_attn(const void* q, void* out, const void* k_cache, const void* v_cache,
                              const int32_t* seq_lens_bh, const int32_t* page_table,
                              const int32_t* batch_mapping, void* workspace, size_t workspace_bytes, int B,
                              int HQ, int HKV, int D, int PS, int NLP, float sm_scale, int dtype,
                              int num_ctas, int lens_by_slot, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B >= 0 && HQ > 0 && HKV > 0 && PS > 0 && NLP > 0);
  if (B == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(q && out && k_cache && v_cache && seq_lens_bh && page_table && batch_mapping &&
                  workspace);
  CVLLM_CHECK_ARG(HQ % HKV == 0);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  const int G = HQ / HKV;
  CVLLM_CHECK_SUPPORTED(D == 128);
  CVLLM_CHECK_SUPPORTED(G <= 8);
  // a TMA box (min(PS,128) rows) must tile the 128-token unit and never cross a page
  CVLLM_CHECK_SUPPORTED(PS % kUnit == 0 || PS == 64 || PS == 32 || PS == 16);
  CVLLM_CHECK_ALIGN(aligned16(q) && aligned16(out) && aligned16(k_cache) && aligned16(v_cache) &&
                    aligned16(workspace));
  const int C = decode_ctas(num_ctas);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_decode_attention_workspace_bytes(B, HQ, HKV, D, num_ctas));
  const size_t BH = (size_t)B * HKV;
  const size_t counters_bytes = decode_counter_bytes(BH);
  auto st = static_cast<cudaStream_t>(stream);

  DecodeParams p;
  p.q = q; p.out = out;
  p.seq_lens_bh = seq_lens_bh; p.page_table = page_table; p.batch_mapping = batch_mapping;
  p.counters = reinterpret_cast<int32_t*>(workspace);
  p.scratch = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + counters_bytes);
  p.B = B; p.HQ = HQ; p.HKV = HKV; p.G = G; p.PS = PS; p.NLP = NLP;
  p.box_rows = PS < kUnit ? PS : kUnit;
  p.lens_by_slot = lens_by_slot;
  p.scale_log2 = sm_scale * 1.4426950408889634f;
#ifdef CVLLM_DECODE_TRACE
  p.trace = g_decode_trace;
#endif
  p.mode = test_options().decode_mode.load(std::memory_order_relaxed);

  // 2-D maps over the flat [rows, D] caches; the row bound is not known to this ABI (only
  // mapped pages are ever addressed), so it is the largest int32 coordinate
  const bool bf16 = dtype == CVLLM_BF16;
  CUtensorMap tm_k, tm_v;
  {
    const uint32_t box[2] = {64, (uint32_t)p.box_rows};
    const uint64_t dims[2] = {(uint64_t)D, (uint64_t)0x7fffffff};
    const uint64_t strides[1] = {(uint64_t)D * 2};
    int rc = make_tmap_16b(&tm_k, k_cache, 2, dims, strides, box, bf16);
    if (rc) return rc;
    rc = make_tmap_16b(&tm_v, v_cache, 2, dims, strides, box, bf16);
    if (rc) return rc;
  }

  // ring depth: as many 32 KB slots as fit beside the merge rows (6 for G <= 4, else 5)
  constexpr int kMaxSmem = 224 * 1024;  // 227 KB opt-in limit minus static smem
  const size_t fixed = 1024 + 2 * kMaxSlots * 8 + (size_t)(kDecWarps + 1) * G * (D + 2) * 4 + (BH + 1) * 4;
  CVLLM_CHECK_SUPPORTED(fixed + 2 * (size_t)kSlotBytes <= (size_t)kMaxSmem);
  int slots = (int)(((size_t)kMaxSmem - fixed) / kSlotBytes);
  if (slots > kMaxSlots) slots = kMaxSlots;
  {
    const int cap = test_options().decode_slots.load(std::memory_order_relaxed);  // profiling knob
    if (cap >= 2 && cap < slots) slots = cap;
  }
  p.slots = slots;
  const size_t smem = fixed + (size_t)slots * kSlotBytes;
  // raise the dynamic-smem cap once per (kernel, device)
  static std::atomic<uint64_t> attr_bf{0}, attr_h{0};
  if (bf16) CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)decode_attn_kernel<true, 128>, kMaxSmem, attr_bf));
  else CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)decode_attn_kernel<false, 128>, kMaxSmem, attr_h));
  if (bf16) {
    decode_attn_kernel<true, 128><<<C, kDecThreads, smem, st>>>(tm_k, tm_v, p);
  } else {
    decode_attn_kernel<false, 128><<<C, kDecThreads, smem, st>>>(tm_k, tm_v, p);
  }
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

extern "C" int cvllm_decode_attention(const void* q, void* out, const void* k_cache,
                                      const void* v_cache, const int32_t* seq_lens_bh,
                                      const int32_t* page_table, const int32_t* batch_mapping,
                                      void* workspace, size_t workspace_bytes, int B, int HQ, int HKV,
                                      int D, int PS, int NLP, float sm_scale, int dtype, int num_ctas,
                                      cvllm_stream_t stream) {
  return launch_decode_attn(q, out, k_cache, v_cache, seq_lens_bh, page_table, batch_mapping, workspace,
                            workspace_bytes, B, HQ, HKV, D, PS, NLP, sm_scale, dtype, num_ctas, 0, stream);
}

// One decode step of Attention.forward (layers/attention.py:133-160): append the new (b, h) rows, then attend,
// both indexing the cache's own bh_seq_lens table [Bmax+1, HKV] through batch_mapping — the reference's
// index_select / clone / index_copy_ of a gathered length copy around the two kernels disappear.
extern "C" int cvllm_decode_step(const void* key, const void* value, int64_t sk_b, int64_t sk_h,
                                 int64_t sv_b, int64_t sv_h, const void* q, void* out, void* k_cache,
                                 void* v_cache, int32_t* bh_seq_lens, const int32_t* page_table,
                                 const int32_t* batch_mapping, void* workspace, size_t workspace_bytes,
                                 int B, int HQ, int HKV, int D, int PS, int NLP, float sm_scale, int dtype,
                                 int reserved_batch, int num_ctas, cvllm_stream_t stream) {
  int rc = launch_decode_store(key, value, sk_b, sk_h, sv_b, sv_h, batch_mapping, bh_seq_lens, page_table,
                               k_cache, v_cache, B, HKV, D, PS, NLP, 2, reserved_batch, 1, stream);
  if (rc) return rc;
  return launch_decode_attn(q, out, k_cache, v_cache, bh_seq_lens, page_table, batch_mapping, workspace,
                            workspace_bytes, B, HQ, HKV, D, PS, NLP, sm_scale, dtype, num_ctas, 1, stream);
}
