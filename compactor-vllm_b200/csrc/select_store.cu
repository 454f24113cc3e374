// Fused top-k eviction + paged-cache compaction, multi-CTA.
//
// Reference: compression/common.py:126-168 (scores -> torch.topk full sort -> K4 atomic-slot store
// -> K5 serial pad walk, kv_cache/store_kv_cache.py:9-248).  The closed form (SURVEY
// 7.3-4): after the per-sequence global top-`retain`, head h holds c_h pairs; it is padded to
// n_h = min(c_h + pad, len) and its kept set is exactly the top-n_h tokens OF THAT HEAD.  So:
//
//   phase 1  4 radix passes (8 bits each) over all len*H keys of a sequence  -> threshold T, ties
//   phase 2  count per head: keys > T, ties at T taken in ascending flat index  -> c_h -> n_h
//   phase 3  4 radix passes per (sequence, head)                              -> T_h, ties_h
//   phase 4  per tile / head: #keys > T_h and #keys == T_h
//   phase 5  ordered compaction: rank inside the head = (kept before this tile) + (kept before
//            this token inside the tile); rows copied straight into the paged cache
//
// Every phase is a grid of (token tiles x sequences) CTAs: the keys of a 32K-token sequence are
// 1 MB and live in L2, each pass is a few microseconds instead of a single-SM walk.  The CTA that
// takes the last ticket of a sequence folds the global histogram into the next digit.
// Order everywhere: key descending, ties by ascending flat index tok*H + head, NaN above +inf.
// Rows land in ascending token order per head (the reference's order is atomic-nondeterministic).
#include "common.cuh"
#include "select_common.cuh"

namespace cvllm {

constexpr int kTileTok = 256;   // tokens per tile (one thread per token in phase 5)
constexpr int kPassThreads = 256;
constexpr int kMaxH = 16;

struct SegState {
  uint32_t prefix;     // decided high digits of the threshold key
  uint32_t remaining;  // rank still to resolve inside the prefix bucket (after 4 passes: ties)
  uint32_t active;     // 0: nothing is selected in this segment
  uint32_t pad_;
};

struct SelWs {
  uint32_t* hist_g;    // [B][256]
  uint32_t* hist_h;    // [B][H][256]
  uint32_t* ticket;    // [B][2]
  uint32_t* gt_h;      // [B][H]   keys > T per head
  SegState* state_g;   // [B]
  SegState* state_h;   // [B][H]
  int32_t* n_keep;     // [B][H]
  int32_t* len0;       // [B][H]   bh_lens before this call
  uint32_t* tile_tie;  // [B][tiles][H]
  uint2* tile_cnt;     // [B][tiles][H]  (gt, tie) w.r.t. T_h
  int tiles;           // tile slots per sequence
};

// ---------------------------------------------------------------------------
// init: k per sequence, copy of the starting lengths
// ---------------------------------------------------------------------------
__global__ void sel_init_kernel(const int32_t* __restrict__ cu, const int32_t* __restrict__ num_retain,
                                const int32_t* __restrict__ batch_mapping,
                                const int32_t* __restrict__ bh_lens, int B, int H, int reserved_batch,
                                SelWs w) {
  pdl_begin();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) {
    const int cnt = (cu[i + 1] - cu[i]) * H;
    const int k = max(0, min(num_retain[i], cnt));
    SegState s;
    s.prefix = 0;
    s.remaining = (uint32_t)k;
    s.active = (batch_mapping[i] != reserved_batch && k > 0) ? 1u : 0u;
    s.pad_ = 0;
    w.state_g[i] = s;
  }
  if (i < B * H) w.len0[i] = bh_lens[i];
}

// digit pick: 256-bin histogram (descending) -> digit holding the `remaining`-th entry.
// Executed by one full warp; returns through lane-uniform refs.
__device__ __forceinline__ void pick_digit(const uint32_t* hist /*global, 256*/, uint32_t remaining,
                                           uint32_t& digit, uint32_t& new_remaining) {
  const int lane = threadIdx.x & 31;
  uint32_t h8[8];
  uint32_t own = 0;
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    h8[t] = __ldcg(hist + 255 - 8 * lane - t);
    own += h8[t];
  }
  uint32_t incl = own;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const uint32_t above = incl - own;
  const bool mine = (above < remaining) && (remaining <= incl);
  uint32_t d = 0, rem = 0;
  if (mine) {
    uint32_t acc = above;
    int t = 0;
    for (; t < 8; ++t) {
      if (remaining <= acc + h8[t]) break;
      acc += h8[t];
    }
    d = (uint32_t)(255 - 8 * lane - t);
    rem = remaining - acc;
  }
  const uint32_t who = __ballot_sync(0xffffffffu, mine);
  const int src = who ? (__ffs(who) - 1) : 0;
  digit = __shfl_sync(0xffffffffu, d, src);
  new_remaining = __shfl_sync(0xffffffffu, rem, src);
}

// ---------------------------------------------------------------------------
// one radix pass.  PER_HEAD == false: segment = sequence; true: segment = (sequence, head)
// grid (tiles, B)
// ---------------------------------------------------------------------------
template <bool PER_HEAD>
__global__ void __launch_bounds__(kPassThreads) sel_pass_kernel(const float* __restrict__ scores,
                                                                int64_t stride_n,
                                                                const int32_t* __restrict__ cu, int H,
                                                                int pass, SelWs w) {
  pdl_begin();
  __shared__ uint32_t s_hist[PER_HEAD ? kMaxH * 256 : 256];
  __shared__ uint32_t s_prefix[kMaxH];
  __shared__ uint32_t s_active[kMaxH];
  __shared__ int s_last;
  const int b = blockIdx.y, tile = blockIdx.x;
  const int beg = cu[b], len = cu[b + 1] - beg;
  const int tiles_b = cdiv(len, kTileTok);
  if (tile >= tiles_b) return;
  const int nseg = PER_HEAD ? H : 1;
  SegState* st = PER_HEAD ? (w.state_h + (int64_t)b * H) : (w.state_g + b);
  uint32_t* hist = PER_HEAD ? (w.hist_h + (int64_t)b * H * 256) : (w.hist_g + (int64_t)b * 256);
  const int tid = threadIdx.x;
  if (tid < nseg) {
    s_prefix[tid] = st[tid].prefix;
    s_active[tid] = st[tid].active;
  }
  for (int i = tid; i < nseg * 256; i += kPassThreads) s_hist[i] = 0;
  __syncthreads();
  bool any_active = false;
  for (int s = 0; s < nseg; ++s) any_active |= (s_active[s] != 0);
  if (!any_active) return;  // uniform over the sequence's CTAs: no ticket needed

  const int shift = 24 - 8 * pass;
  const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (32 - 8 * pass));
  const int t0 = tile * kTileTok;
  const int nk = min(kTileTok, len - t0) * H;
  const float* src = scores + (int64_t)(beg + t0) * stride_n;
  for (int i = tid; i < nk; i += kPassThreads) {
    const int tok = i / H, h = i - tok * H;
    const int seg = PER_HEAD ? h : 0;
    if (!s_active[seg]) continue;
    const uint32_t key = score_key(__float_as_uint(__ldg(src + (int64_t)tok * stride_n + h)));
    if ((key & mask) == s_prefix[seg]) atomicAdd(&s_hist[seg * 256 + ((key >> shift) & 255u)], 1u);
  }
  __syncthreads();
  for (int i = tid; i < nseg * 256; i += kPassThreads) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(hist + i, c);
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const uint32_t t = atomicAdd(w.ticket + 2 * b + (PER_HEAD ? 1 : 0), 1u);
    s_last = (t == (uint32_t)(tiles_b - 1));
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last CTA of the sequence: fold the histogram(s) into the next digit
  const int warp = tid >> 5, lane = tid & 31;
  for (int s = warp; s < nseg; s += kPassThreads / 32) {
    if (s_active[s]) {
      uint32_t digit, rem;
      pick_digit(hist + s * 256, st[s].remaining, digit, rem);
      if (lane == 0) {
        st[s].prefix = s_prefix[s] | (digit << shift);
        st[s].remaining = rem;
      }
    }
    __syncwarp();
    for (int i = lane; i < 256; i += 32) hist[s * 256 + i] = 0;
  }
  if (tid == 0) w.ticket[2 * b + (PER_HEAD ? 1 : 0)] = 0;
}

// ---------------------------------------------------------------------------
// phase 2a: per tile, per head: keys > T (-> global per-head counters), ties at T (-> per tile)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kPassThreads) sel_count_global_kernel(const float* __restrict__ scores,
                                                                        int64_t stride_n,
                                                                        const int32_t* __restrict__ cu,
                                                                        int H, SelWs w) {
  pdl_begin();
  __shared__ uint32_t s_gt[kMaxH], s_tie[kMaxH];
  const int b = blockIdx.y, tile = blockIdx.x;
  const int beg = cu[b], len = cu[b + 1] - beg;
  if (tile >= cdiv(len, kTileTok)) return;
  const SegState sg = w.state_g[b];
  const int tid = threadIdx.x;
  if (tid < H) s_gt[tid] = s_tie[tid] = 0;
  __syncthreads();
  if (sg.active) {
    const int t0 = tile * kTileTok;
    const int nk = min(kTileTok, len - t0) * H;
    const float* src = scores + (int64_t)(beg + t0) * stride_n;
    for (int i = tid; i < nk; i += kPassThreads) {
      const int tok = i / H, h = i - tok * H;
      const uint32_t key = score_key(__float_as_uint(__ldg(src + (int64_t)tok * stride_n + h)));
      if (key > sg.prefix) atomicAdd(&s_gt[h], 1u);
      else if (key == sg.prefix) atomicAdd(&s_tie[h], 1u);
    }
  }
  __syncthreads();
  if (tid < H) {
    if (s_gt[tid]) atomicAdd(w.gt_h + b * H + tid, s_gt[tid]);
    w.tile_tie[((int64_t)b * w.tiles + tile) * H + tid] = s_tie[tid];
  }
}

// ---------------------------------------------------------------------------
// phase 2b: one CTA per sequence: allocate the ties in flat-index order, c_h -> n_h, seed phase 3
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kPassThreads) sel_resolve_kernel(const float* __restrict__ scores,
                                                                   int64_t stride_n,
                                                                   const int32_t* __restrict__ cu,
                                                                   const int32_t* __restrict__ batch_mapping,
                                                                   int H, int PS, int pad_to_page,
                                                                   int reserved_batch, SelWs w) {
  pdl_begin();
  __shared__ uint32_t s_tiesel[kMaxH];
  __shared__ int s_warp[32];
  __shared__ int s_btile;
  __shared__ uint32_t s_before;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int beg = cu[b], len = cu[b + 1] - beg;
  const int tiles_b = cdiv(len, kTileTok);
  const SegState sg = w.state_g[b];
  const uint32_t* tt = w.tile_tie + (int64_t)b * w.tiles * H;
  if (tid < H) s_tiesel[tid] = 0;
  if (tid == 0) { s_btile = -1; s_before = 0; }
  __syncthreads();
  if (sg.active) {
    // boundary tile: the first one whose cumulative tie count (flat order) exceeds `remaining`
    {
      const int lane = tid & 31, warp = tid >> 5;
      uint32_t carry = 0;
      if (tid == 0) s_btile = tiles_b;  // default: every tie is taken
      __syncthreads();
      for (int base = 0; base < tiles_b; base += kPassThreads) {
        const int t = base + tid;
        uint32_t c = 0;
        if (t < tiles_b)
          for (int h = 0; h < H; ++h) c += tt[t * H + h];
        // block-wide exclusive prefix of c
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += x;
        }
        if (lane == 31) s_warp[warp] = (int)incl;
        __syncthreads();
        uint32_t wbase = 0, total = 0;
        for (int wv = 0; wv < kPassThreads / 32; ++wv) {
          if (wv < warp) wbase += (uint32_t)s_warp[wv];
          total += (uint32_t)s_warp[wv];
        }
        const uint32_t run = carry + wbase + incl - c;
        if (t < tiles_b && run <= sg.remaining && run + c > sg.remaining) {
          s_btile = t;  // unique: the prefix is monotone
          s_before = run;
        }
        carry += total;
        __syncthreads();
      }
      if (tid == 0 && s_btile == tiles_b) s_before = carry;
      __syncthreads();
    }
    const int bt = s_btile;
    // whole tiles before the boundary: all their ties are taken
    for (int i = tid; i < min(bt, tiles_b) * H; i += kPassThreads) {
      const uint32_t c = tt[i];
      if (c) atomicAdd(&s_tiesel[i % H], c);
    }
    __syncthreads();
    if (bt < tiles_b) {
      // ordered pass inside the boundary tile (flat order = token major, head minor)
      const int need = (int)(sg.remaining - s_before);
      const int t0 = bt * kTileTok;
      const int nk = min(kTileTok, len - t0) * H;
      const float* src = scores + (int64_t)(beg + t0) * stride_n;
      int seen = 0;
      for (int base = 0; base < nk && seen < need; base += kPassThreads) {
        const int i = base + tid;
        bool tie = false;
        int h = 0;
        if (i < nk) {
          const int tok = i / H;
          h = i - tok * H;
          tie = score_key(__float_as_uint(__ldg(src + (int64_t)tok * stride_n + h))) == sg.prefix;
        }
        int tot;
        const int rank = seen + block_excl_count(tie, s_warp, tot);
        if (tie && rank < need) atomicAdd(&s_tiesel[h], 1u);
        seen += tot;
      }
      __syncthreads();
    }
  }
  if (tid < H) {
    const int h = tid;
    int n = 0;
    if (sg.active) {
      const int c = (int)(w.gt_h[b * H + h] + s_tiesel[h]);
      n = c;
      if (pad_to_page) n += pad_amount(w.len0[b * H + h] + c, len, PS);
      n = min(n, len);
    }
    w.n_keep[b * H + h] = n;
    SegState s;
    s.prefix = 0;
    s.remaining = (uint32_t)n;
    s.active = (n > 0 && batch_mapping[b] != reserved_batch) ? 1u : 0u;
    s.pad_ = 0;
    w.state_h[b * H + h] = s;
    w.gt_h[b * H + h] = 0;  // leave the workspace clean for the next call on this stream
  }
}

// ---------------------------------------------------------------------------
// phase 4: per tile / head: (#keys > T_h, #keys == T_h)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kPassThreads) sel_count_heads_kernel(const float* __restrict__ scores,
                                                                       int64_t stride_n,
                                                                       const int32_t* __restrict__ cu,
                                                                       int H, SelWs w) {
  pdl_begin();
  __shared__ uint32_t s_gt[kMaxH], s_tie[kMaxH], s_thr[kMaxH], s_act[kMaxH];
  const int b = blockIdx.y, tile = blockIdx.x;
  const int beg = cu[b], len = cu[b + 1] - beg;
  if (tile >= cdiv(len, kTileTok)) return;
  const int tid = threadIdx.x;
  if (tid < H) {
    s_gt[tid] = s_tie[tid] = 0;
    s_thr[tid] = w.state_h[b * H + tid].prefix;
    s_act[tid] = w.state_h[b * H + tid].active;
  }
  __syncthreads();
  const int t0 = tile * kTileTok;
  const int nk = min(kTileTok, len - t0) * H;
  const float* src = scores + (int64_t)(beg + t0) * stride_n;
  for (int i = tid; i < nk; i += kPassThreads) {
    const int tok = i / H, h = i - tok * H;
    if (!s_act[h]) continue;
    const uint32_t key = score_key(__float_as_uint(__ldg(src + (int64_t)tok * stride_n + h)));
    if (key > s_thr[h]) atomicAdd(&s_gt[h], 1u);
    else if (key == s_thr[h]) atomicAdd(&s_tie[h], 1u);
  }
  __syncthreads();
  if (tid < H) w.tile_cnt[((int64_t)b * w.tiles + tile) * H + tid] = make_uint2(s_gt[tid], s_tie[tid]);
}

// ---------------------------------------------------------------------------
// phase 5: ordered compaction + row copy.  One thread per token of the tile.
// ---------------------------------------------------------------------------
template <int VEC_PER_ROW>
__global__ void __launch_bounds__(kTileTok) sel_gather_kernel(
    const float* __restrict__ scores, int64_t stride_n, const uint4* __restrict__ keys,
    const uint4* __restrict__ vals, int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
    const int32_t* __restrict__ cu, const int32_t* __restrict__ batch_mapping,
    const int32_t* __restrict__ page_table, int32_t* __restrict__ bh_lens, uint4* __restrict__ k_cache,
    uint4* __restrict__ v_cache, int H, int PS, int NLP, SelWs w) {
  pdl_begin();
  __shared__ uint32_t s_gt_before[kMaxH], s_tie_before[kMaxH], s_gt_all[kMaxH], s_tie_all[kMaxH];
  __shared__ uint32_t s_thr[kMaxH], s_ties[kMaxH], s_act[kMaxH];
  __shared__ uint32_t s_wsel[kMaxH][kTileTok / 32], s_wtie[kMaxH][kTileTok / 32];
  __shared__ int64_t s_dst[kTileTok * kMaxH];  // destination row in the cache, -1 = dropped
  const int b = blockIdx.y, tile = blockIdx.x;
  const int beg = cu[b], len = cu[b + 1] - beg;
  const int tiles_b = cdiv(len, kTileTok);
  if (tile >= tiles_b) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < H) {
    s_gt_before[tid] = s_tie_before[tid] = s_gt_all[tid] = s_tie_all[tid] = 0;
    const SegState s = w.state_h[b * H + tid];
    s_thr[tid] = s.prefix;
    s_ties[tid] = s.remaining;
    s_act[tid] = s.active;
  }
  __syncthreads();
  // prefix of the per-tile counts over the tiles before this one (and totals, for bh_lens)
  {
    const uint2* tc = w.tile_cnt + (int64_t)b * w.tiles * H;
    for (int i = tid; i < tiles_b * H; i += kTileTok) {
      const int t = i / H, h = i - t * H;
      if (!s_act[h]) continue;
      const uint2 c = tc[i];
      if (t < tile) {
        if (c.x) atomicAdd(&s_gt_before[h], c.x);
        if (c.y) atomicAdd(&s_tie_before[h], c.y);
      }
      if (tile == 0) {
        if (c.x) atomicAdd(&s_gt_all[h], c.x);
        if (c.y) atomicAdd(&s_tie_all[h], c.y);
      }
    }
  }
  __syncthreads();
  const int tok = tile * kTileTok + tid;
  const bool valid = tok < len;
  const float* srow = scores + (int64_t)(beg + tok) * stride_n;
  // per-head flags of this token
  uint32_t gt_mask = 0, tie_mask = 0;
  if (valid) {
    for (int h = 0; h < H; ++h) {
      if (!s_act[h]) continue;
      const uint32_t key = score_key(__float_as_uint(__ldg(srow + h)));
      if (key > s_thr[h]) gt_mask |= 1u << h;
      else if (key == s_thr[h]) tie_mask |= 1u << h;
    }
  }
  // ties first (their rank decides whether they are kept)
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t tie_lane[kMaxH];
#pragma unroll
  for (int h = 0; h < kMaxH; ++h) {
    if (h < H) {
      const uint32_t m = __ballot_sync(0xffffffffu, (tie_mask >> h) & 1u);
      tie_lane[h] = __popc(m & lt_mask);
      if (lane == 0) s_wtie[h][warp] = __popc(m);
    }
  }
  __syncthreads();
  uint32_t sel_mask = gt_mask;
#pragma unroll
  for (int h = 0; h < kMaxH; ++h) {
    if (h < H && ((tie_mask >> h) & 1u)) {
      uint32_t before = s_tie_before[h] + tie_lane[h];
      for (int wv = 0; wv < warp; ++wv) before += s_wtie[h][wv];
      if (before < s_ties[h]) sel_mask |= 1u << h;
    }
  }
  uint32_t sel_lane[kMaxH];
#pragma unroll
  for (int h = 0; h < kMaxH; ++h) {
    if (h < H) {
      const uint32_t m = __ballot_sync(0xffffffffu, (sel_mask >> h) & 1u);
      sel_lane[h] = __popc(m & lt_mask);
      if (lane == 0) s_wsel[h][warp] = __popc(m);
    }
  }
  __syncthreads();
  const int32_t* pt = page_table + (int64_t)batch_mapping[b] * H * NLP;
#pragma unroll
  for (int h = 0; h < kMaxH; ++h) {
    if (h < H) {
      int64_t dst = -1;
      if ((sel_mask >> h) & 1u) {
        uint32_t rank = s_gt_before[h] + min(s_tie_before[h], s_ties[h]) + sel_lane[h];
        for (int wv = 0; wv < warp; ++wv) rank += s_wsel[h][wv];
        const int p = w.len0[b * H + h] + (int)rank;
        const int lp = p / PS;
        dst = (int64_t)pt[h * NLP + lp] * PS + (p - lp * PS);
      }
      s_dst[tid * H + h] = dst;
    }
  }
  if (tile == 0 && blockIdx.z == 0 && tid < H && s_act[tid])
    bh_lens[b * H + tid] = w.len0[b * H + tid] + (int)(s_gt_all[tid] + min(s_tie_all[tid], s_ties[tid]));
  __syncthreads();
  // cooperative row copy: VEC_PER_ROW lanes per (token, head) row
  constexpr int ROWS_PER_IT = kTileTok / VEC_PER_ROW;
  const int vec = tid % VEC_PER_ROW, rin = tid / VEC_PER_ROW;
  const int n_rows = min(kTileTok, len - tile * kTileTok) * H;
  const int64_t tok0 = beg + tile * kTileTok;
  // gridDim.z CTAs share a tile's rows (each recomputes the destinations above, ~1 us, and copies every z-th
  // group): the copy is 134 MB of pure bandwidth at 32K and one 256-thread CTA per 256 tokens (128 CTAs, 32 KB in
  // flight each) reached 2.4 TB/s
  for (int r0 = blockIdx.z * 4 * ROWS_PER_IT; r0 < n_rows; r0 += gridDim.z * 4 * ROWS_PER_IT) {
    uint4 kk[4], vv[4];
    int64_t d[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * ROWS_PER_IT + rin;
      d[u] = r < n_rows ? s_dst[r] : -1;
      if (d[u] >= 0) {
        const int t = r / H, h = r - t * H;
        kk[u] = ldg_nc_v4(keys + (tok0 + t) * sk_n + h * sk_h + vec);
        vv[u] = ldg_nc_v4(vals + (tok0 + t) * sv_n + h * sv_h + vec);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (d[u] >= 0) {
        stg_na_v4(k_cache + d[u] * VEC_PER_ROW + vec, kk[u]);
        stg_na_v4(v_cache + d[u] * VEC_PER_ROW + vec, vv[u]);
      }
    }
  }
}

}  // namespace cvllm

using namespace cvllm;

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct SelLayout {
  size_t hist_g, hist_h, ticket, gt_h, zero_end, state_g, state_h, n_keep, len0, tile_tie, tile_cnt, total;
  int tiles;
};

static SelLayout sel_layout(int B, int H, int N) {
  SelLayout l;
  l.tiles = (N + kTileTok - 1) / kTileTok + 1;
  size_t o = 0;
  l.hist_g = o; o += align256((size_t)B * 256 * 4);
  l.hist_h = o; o += align256((size_t)B * H * 256 * 4);
  l.ticket = o; o += align256((size_t)B * 2 * 4);
  l.gt_h = o; o += align256((size_t)B * H * 4);
  l.zero_end = o;
  l.state_g = o; o += align256((size_t)B * sizeof(SegState));
  l.state_h = o; o += align256((size_t)B * H * sizeof(SegState));
  l.n_keep = o; o += align256((size_t)B * H * 4);
  l.len0 = o; o += align256((size_t)B * H * 4);
  l.tile_tie = o; o += align256((size_t)B * l.tiles * H * 4);
  l.tile_cnt = o; o += align256((size_t)B * l.tiles * H * 8);
  l.total = o;
  return l;
}

extern "C" size_t cvllm_select_store_workspace_bytes(int B, int H, int N) {
  if (B <= 0 || H <= 0 || N < 0) return 0;
  return sel_layout(B, H, N).total;
}

template <int V>
static void launch_sel_gather(dim3 grid, cudaStream_t st, const float* scores, int64_t s_stride_n,
                              const void* keys, const void* vals, int64_t sk_n, int64_t sk_h,
                              int64_t sv_n, int64_t sv_h, const int32_t* cu, const int32_t* bm,
                              const int32_t* pt, int32_t* bh_lens, void* kc, void* vc, int H, int PS,
                              int NLP, const SelWs& w) {
  launch_pdl(sel_gather_kernel<V>, grid, dim3(kTileTok), 0, st, scores, s_stride_n, (const uint4*)keys,
             (const uint4*)vals, sk_n, sk_h, sv_n, sv_h, cu, bm, pt, bh_lens, (uint4*)kc, (uint4*)vc, H, PS, NLP, w);
}

extern "C" int cvllm_select_store_topk(const float* scores, int64_t s_stride_n, const void* keys,
                                       const void* vals, int64_t sk_n, int64_t sk_h, int64_t sv_n,
                                       int64_t sv_h, const int32_t* num_retain,
                                       const int32_t* cu_seqlens_k, const int32_t* batch_mapping,
                                       int32_t* bh_lens, const int32_t* page_table, void* k_cache,
                                       void* v_cache, int B, int N, int HKV, int D, int PS, int NLP,
                                       int elem_bytes, int pad_to_page, int reserved_batch,
                                       void* workspace, size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && N >= 0 && HKV > 0 && D > 0 && PS > 0 && NLP > 0);
  CVLLM_CHECK_ARG(scores && keys && vals && num_retain && cu_seqlens_k && batch_mapping && bh_lens &&
                  page_table && k_cache && v_cache && workspace);
  if (N == 0) return CVLLM_OK;
  int vpr;
  int64_t per16;
  const int rc = cvllm_check_store_args(keys, vals, sk_n, sk_h, sv_n, sv_h, k_cache, v_cache, D,
                                        elem_bytes, HKV, &vpr, &per16);
  if (rc != CVLLM_OK) return rc;
  CVLLM_CHECK_SUPPORTED(HKV <= kMaxH);
  const SelLayout l = sel_layout(B, HKV, N);
  CVLLM_CHECK_ARG(workspace_bytes >= l.total);
  CVLLM_CHECK_ALIGN(aligned16(workspace));
  char* base = reinterpret_cast<char*>(workspace);
  SelWs w;
  w.hist_g = reinterpret_cast<uint32_t*>(base + l.hist_g);
  w.hist_h = reinterpret_cast<uint32_t*>(base + l.hist_h);
  w.ticket = reinterpret_cast<uint32_t*>(base + l.ticket);
  w.gt_h = reinterpret_cast<uint32_t*>(base + l.gt_h);
  w.state_g = reinterpret_cast<SegState*>(base + l.state_g);
  w.state_h = reinterpret_cast<SegState*>(base + l.state_h);
  w.n_keep = reinterpret_cast<int32_t*>(base + l.n_keep);
  w.len0 = reinterpret_cast<int32_t*>(base + l.len0);
  w.tile_tie = reinterpret_cast<uint32_t*>(base + l.tile_tie);
  w.tile_cnt = reinterpret_cast<uint2*>(base + l.tile_cnt);
  w.tiles = l.tiles;
  auto st = static_cast<cudaStream_t>(stream);
  CVLLM_CHECK_CUDA(cudaMemsetAsync(base, 0, l.zero_end, st));
  const dim3 grid((N + kTileTok - 1) / kTileTok, B);
  sel_init_kernel<<<(B * HKV + 255) / 256, 256, 0, st>>>(cu_seqlens_k, num_retain, batch_mapping, bh_lens,
                                                         B, HKV, reserved_batch, w);
  // eleven short kernels in a row, each dependent on the one before; launch_pdl is a plain launch unless the test
  // option "pdl" is set (common.cuh has the measurements)
  const dim3 pb(kPassThreads);
  for (int pass = 0; pass < 4; ++pass)
    CVLLM_CHECK_CUDA(launch_pdl(sel_pass_kernel<false>, grid, pb, 0, st, scores, s_stride_n, cu_seqlens_k, HKV, pass, w));
  CVLLM_CHECK_CUDA(launch_pdl(sel_count_global_kernel, grid, pb, 0, st, scores, s_stride_n, cu_seqlens_k, HKV, w));
  CVLLM_CHECK_CUDA(launch_pdl(sel_resolve_kernel, dim3(B), pb, 0, st, scores, s_stride_n, cu_seqlens_k, batch_mapping,
                              HKV, PS, pad_to_page, reserved_batch, w));
  for (int pass = 0; pass < 4; ++pass)
    CVLLM_CHECK_CUDA(launch_pdl(sel_pass_kernel<true>, grid, pb, 0, st, scores, s_stride_n, cu_seqlens_k, HKV, pass, w));
  CVLLM_CHECK_CUDA(launch_pdl(sel_count_heads_kernel, grid, pb, 0, st, scores, s_stride_n, cu_seqlens_k, HKV, w));
  CVLLM_CHECK_LAUNCHES(12);
  // enough copy CTAs to fill the chip a few times over: 4 per tile at long prompts, 1 when the grid is already wide
  const int64_t gtiles = (int64_t)grid.x * grid.y;
  const dim3 ggrid(grid.x, grid.y, gtiles >= 4 * sm_count() ? 1 : 4);
#define G(V)                                                                                            \
  launch_sel_gather<V>(ggrid, st, scores, s_stride_n, keys, vals, sk_n / per16, sk_h / per16, sv_n / per16, \
                       sv_h / per16, cu_seqlens_k, batch_mapping, page_table, bh_lens, k_cache, v_cache,   \
                       HKV, PS, NLP, w)
  switch (vpr) {
    case 1: G(1); break;
    case 2: G(2); break;
    case 4: G(4); break;
    case 8: G(8); break;
    case 16: G(16); break;
    case 32: G(32); break;
    default: return CVLLM_E_UNSUPPORTED;
  }
#undef G
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
