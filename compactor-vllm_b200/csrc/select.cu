// Score ranking, top-k eviction and paged-cache compaction.
//
// Reference: compression/common.py:126-243 (scores -> torch.topk full sort ->
// store) and kv_cache/store_kv_cache.py:9-248 (K4 atomic-slot store + K5
// serial pad walk).  Three entry points:
//   cvllm_rank_scores            full descending ranking (stable LSD radix sort)
//   cvllm_prefill_store_topk_idx ranking -> keep flags -> pad -> compaction
//   cvllm_select_store_topk      scores -> radix-select (no sort) -> compaction
// Kept sets are identical between the two store flavours given the same
// scores; ordering rule everywhere: key descending, ties by ascending flat
// index (tok*H + head), NaN above +inf.  Rows land in ascending token order.
#include "common.cuh"
#include "select_common.cuh"

namespace cvllm {


// ===========================================================================
// A. full ranking: stable LSD radix sort, 4 x 8-bit passes, multi-block
// ===========================================================================
constexpr int kSortTile = 2048;  // elements per block (256 threads x 8 rounds)
constexpr int kSortThreads = 256;

__global__ void sort_init_kernel(const float* __restrict__ scores, int64_t stride_n,
                                 const int32_t* __restrict__ cu, int H, int total,
                                 uint32_t* __restrict__ keys, uint32_t* __restrict__ idx) {
  const int b = blockIdx.y;
  const int n_real = (cu[b + 1] - cu[b]) * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    uint32_t k;
    if (i < n_real) {
      const int tok = i / H, h = i - tok * H;
      k = ~load_key(scores, stride_n, cu[b] + tok, h);
    } else {
      k = ~score_key(0xff800000u);  // padded slot: -inf, after every real entry
    }
    keys[(int64_t)b * total + i] = k;
    idx[(int64_t)b * total + i] = (uint32_t)i;
  }
}

// hist[b][digit][tile]
__global__ void sort_hist_kernel(const uint32_t* __restrict__ keys, int total, int tiles, int shift,
                                 uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_h[256];
  const int b = blockIdx.y, tile = blockIdx.x;
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t* kb = keys + (int64_t)b * total;
  for (int rnd = 0; rnd < kSortTile / kSortThreads; ++rnd) {
    const int i = tile * kSortTile + rnd * kSortThreads + threadIdx.x;
    if (i < total) atomicAdd(&s_h[(kb[i] >> shift) & 255u], 1u);
  }
  __syncthreads();
  hist[((int64_t)b * 256 + threadIdx.x) * tiles + tile] = s_h[threadIdx.x];
}

// exclusive scan of hist[b][256*tiles] in place (one block per sequence)
__global__ void sort_scan_kernel(uint32_t* __restrict__ hist, int tiles) {
  __shared__ uint32_t s_part[kSelThreads];
  const int n = 256 * tiles;
  uint32_t* hb = hist + (int64_t)blockIdx.x * n;
  const int per = cdiv(n, kSelThreads);
  const int beg = threadIdx.x * per, end = min(n, beg + per);
  uint32_t local = 0;
  for (int i = beg; i < end; ++i) local += hb[i];
  s_part[threadIdx.x] = local;
  __syncthreads();
  // Hillis-Steele over 1024 partials
  for (int o = 1; o < kSelThreads; o <<= 1) {
    uint32_t t = (threadIdx.x >= o) ? s_part[threadIdx.x - o] : 0u;
    __syncthreads();
    s_part[threadIdx.x] += t;
    __syncthreads();
  }
  uint32_t run = s_part[threadIdx.x] - local;
  for (int i = beg; i < end; ++i) {
    const uint32_t v = hb[i];
    hb[i] = run;
    run += v;
  }
}

__global__ void sort_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                    const uint32_t* __restrict__ idx_in, uint32_t* __restrict__ keys_out,
                                    uint32_t* __restrict__ idx_out, int total, int tiles, int shift,
                                    const uint32_t* __restrict__ offs) {
  __shared__ uint32_t s_base[256];                       // running global offset per digit
  __shared__ uint32_t s_wcnt[kSortThreads / 32][256];    // per-warp digit counts of this round
  const int b = blockIdx.y, tile = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  s_base[threadIdx.x] = offs[((int64_t)b * 256 + threadIdx.x) * tiles + tile];
  const uint32_t* kb = keys_in + (int64_t)b * total;
  const uint32_t* ib = idx_in + (int64_t)b * total;
  uint32_t* ko = keys_out + (int64_t)b * total;
  uint32_t* io = idx_out + (int64_t)b * total;
  for (int rnd = 0; rnd < kSortTile / kSortThreads; ++rnd) {
    for (int d = threadIdx.x; d < (kSortThreads / 32) * 256; d += kSortThreads)
      (&s_wcnt[0][0])[d] = 0;
    __syncthreads();
    const int i = tile * kSortTile + rnd * kSortThreads + threadIdx.x;
    const bool valid = i < total;
    uint32_t key = 0, id = 0, digit = 0;
    if (valid) { key = kb[i]; id = ib[i]; digit = (key >> shift) & 255u; }
    // rank among same-digit lanes of this warp (invalid lanes use digit 256+lane: unique)
    const uint32_t mkey = valid ? digit : (256u + lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, mkey);
    const int wrank = __popc(peers & ((1u << lane) - 1u));
    if (valid && wrank == 0) s_wcnt[warp][digit] = __popc(peers);
    __syncthreads();
    uint32_t pos = 0;
    if (valid) {
      uint32_t before = 0;
      for (int w = 0; w < warp; ++w) before += s_wcnt[w][digit];
      pos = s_base[digit] + before + wrank;
    }
    __syncthreads();
    // advance the running base by this round's totals
    {
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < kSortThreads / 32; ++w) tot += s_wcnt[w][threadIdx.x];
      s_base[threadIdx.x] += tot;
    }
    if (valid) { ko[pos] = key; io[pos] = id; }
    __syncthreads();
  }
}

__global__ void sort_emit_kernel(const uint32_t* __restrict__ idx, const int32_t* __restrict__ cu,
                                 int H, int total, int k_eff, int64_t* __restrict__ out) {
  const int b = blockIdx.y;
  const int64_t base = (int64_t)cu[b] * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k_eff; i += gridDim.x * blockDim.x)
    out[(int64_t)b * k_eff + i] = base + idx[(int64_t)b * total + i];
}

// ===========================================================================
// shared: CTA-wide radix select of the k-th largest key
// ===========================================================================
struct SelectResult {
  uint32_t thresh;  // key of the k-th largest
  uint32_t ties;    // how many entries == thresh belong to the top-k
};

// f(i) -> key for i in [0, cnt); all kSelThreads threads call this.
template <class F>
__device__ SelectResult block_radix_select(F f, int cnt, uint32_t k, uint32_t* s_hist /*[256]*/,
                                           uint32_t* s_bcast /*[2]*/) {
  uint32_t prefix = 0, mask = 0, remaining = k;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    if (threadIdx.x < 256) s_hist[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      const uint32_t key = f(i);
      if ((key & mask) == prefix) atomicAdd(&s_hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 32) {
      // lane l owns digits [255-8l-7, 255-8l]; suffix-sum from the top
      const int lane = threadIdx.x;
      uint32_t own = 0;
#pragma unroll
      for (int t = 0; t < 8; ++t) own += s_hist[255 - 8 * lane - t];
      uint32_t incl = own;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const uint32_t above = incl - own;  // entries in digits above this lane's block
      const bool mine = (above < remaining) && (remaining <= incl);
      if (mine) {
        uint32_t acc = above;
        int d = 255 - 8 * lane;
        for (int t = 0; t < 8; ++t, --d) {
          const uint32_t c = s_hist[d];
          if (remaining <= acc + c) break;
          acc += c;
        }
        s_bcast[0] = (uint32_t)d;
        s_bcast[1] = remaining - acc;
      }
    }
    __syncthreads();
    prefix |= s_bcast[0] << shift;
    mask |= 0xffu << shift;
    remaining = s_bcast[1];
    __syncthreads();
  }
  SelectResult r;
  r.thresh = prefix;
  r.ties = remaining;
  return r;
}

// ===========================================================================
// C. ordered compaction from keep flags (index flavour) + row gather
// ===========================================================================
// one CTA per (head, sequence): PRED == 0: threshold the head's own scores at
// n_keep[b,h]; PRED == 1: use the keep flags.  Ordered pass writes the
// destination logical position of every kept (token, head) into pos[n*H+h]
// (-1 = dropped) and bumps bh_lens.
template <int PRED>
__global__ void __launch_bounds__(kSelThreads) compact_head_kernel(
    const float* __restrict__ scores, int64_t stride_n, const uint8_t* __restrict__ keep,
    const int32_t* __restrict__ cu, const int32_t* __restrict__ batch_mapping,
    const int32_t* __restrict__ n_keep, int32_t* __restrict__ bh_lens, int H, int reserved_batch,
    int32_t* __restrict__ pos) {
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_bcast[2];
  __shared__ int s_warp[32];
  const int h = blockIdx.x, b = blockIdx.y;
  const int beg = cu[b], len = cu[b + 1] - beg;
  const bool skip = batch_mapping[b] == reserved_batch;
  const int L0 = bh_lens[b * H + h];
  SelectResult sr;
  sr.thresh = 0xffffffffu;
  sr.ties = 0;
  int k = 0;
  if (PRED == 0) {
    k = skip ? 0 : min(n_keep[b * H + h], len);
    if (k > 0) {
      auto f = [&](int i) { return load_key(scores, stride_n, beg + i, h); };
      sr = block_radix_select(f, len, (uint32_t)k, s_hist, s_bcast);
    }
  }
  int run = 0, ties_seen = 0;
  for (int base = 0; base < len; base += kSelThreads) {
    const int i = base + threadIdx.x;
    const bool valid = i < len;
    bool sel = false;
    if (PRED == 0) {
      uint32_t key = 0;
      if (valid && k > 0) key = load_key(scores, stride_n, beg + i, h);
      const bool tie = valid && k > 0 && key == sr.thresh;
      int tt;
      const int trank = ties_seen + block_excl_count(tie, s_warp, tt);
      ties_seen += tt;
      sel = valid && k > 0 && (key > sr.thresh || (tie && trank < (int)sr.ties));
    } else {
      sel = valid && !skip && keep[(int64_t)(beg + i) * H + h] != 0;
    }
    int tot;
    const int rank = run + block_excl_count(sel, s_warp, tot);
    run += tot;
    if (valid) pos[(int64_t)(beg + i) * H + h] = sel ? (L0 + rank) : -1;
  }
  if (threadIdx.x == 0 && !skip) bh_lens[b * H + h] = L0 + run;
}

// gather kept rows into the paged cache: one row per VEC_PER_ROW lanes
template <int VEC_PER_ROW>
__global__ void gather_rows_kernel(const uint4* __restrict__ keys, const uint4* __restrict__ vals,
                                   int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
                                   const int32_t* __restrict__ pos, const int32_t* __restrict__ cu,
                                   const int32_t* __restrict__ batch_mapping,
                                   const int32_t* __restrict__ page_table, uint4* __restrict__ k_cache,
                                   uint4* __restrict__ v_cache, int B, int N, int H, int PS, int NLP) {
  constexpr int ROWS_PER_CTA = 256 / VEC_PER_ROW;
  const int vec = threadIdx.x % VEC_PER_ROW;
  const int rin = threadIdx.x / VEC_PER_ROW;
  const int64_t total_rows = (int64_t)N * H;
  for (int64_t row = (int64_t)blockIdx.x * ROWS_PER_CTA + rin; row < total_rows;
       row += (int64_t)gridDim.x * ROWS_PER_CTA) {
    const int n = (int)(row / H), h = (int)(row % H);
    if (n >= cu[B]) continue;  // rows past the packed batch carry no position
    const int p = pos[row];
    if (p < 0) continue;
    int lo = 0, hi = B;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cu[mid] <= n) lo = mid; else hi = mid;
    }
    const int lp = p / PS, off = p - lp * PS;
    const int64_t phys = page_table[((int64_t)batch_mapping[lo] * H + h) * NLP + lp];
    const int64_t dst = (phys * PS + off) * VEC_PER_ROW + vec;
    const uint4 kk = ldg_nc_v4(keys + n * sk_n + h * sk_h + vec);
    const uint4 vv = ldg_nc_v4(vals + n * sv_n + h * sv_h + vec);
    stg_na_v4(k_cache + dst, kk);
    stg_na_v4(v_cache + dst, vv);
  }
}

// ===========================================================================
// B. index flavour: mark the first num_retain entries, count per head
// ===========================================================================
template <typename IDX>
__global__ void mark_topk_kernel(const IDX* __restrict__ indices, int max_sel,
                                 const int32_t* __restrict__ num_retain, const int32_t* __restrict__ cu,
                                 const int32_t* __restrict__ batch_mapping, int H, int reserved_batch,
                                 uint8_t* __restrict__ keep, int32_t* __restrict__ cnt /*[B,H]*/) {
  const int b = blockIdx.y;
  if (batch_mapping[b] == reserved_batch) return;
  const int k = min(num_retain[b], max_sel);
  const int64_t lo = (int64_t)cu[b] * H, hi = (int64_t)cu[b + 1] * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k; i += gridDim.x * blockDim.x) {
    const int64_t sel = (int64_t)indices[(int64_t)b * max_sel + i];
    if (sel < lo || sel >= hi) continue;  // padded slot: out of contract, skipped
    keep[sel] = 1;
    atomicAdd(&cnt[b * H + (int)(sel % H)], 1);
  }
}

// one CTA per sequence: walk the ranking from num_retain[b] in 1024-wide steps
// and mark the next entries of each head until it reaches a page multiple.
template <typename IDX>
__global__ void __launch_bounds__(kSelThreads) pad_walk_kernel(
    const IDX* __restrict__ indices, int max_sel, const int32_t* __restrict__ num_retain,
    const int32_t* __restrict__ cu, const int32_t* __restrict__ batch_mapping,
    const int32_t* __restrict__ bh_lens, const int32_t* __restrict__ cnt, int H, int PS,
    int reserved_batch, uint8_t* __restrict__ keep) {
  __shared__ int s_need[64];
  __shared__ int s_have[64];
  __shared__ int s_wcnt[32][17];
  __shared__ int s_pending;
  const int b = blockIdx.x;
  if (batch_mapping[b] == reserved_batch) return;
  const int len = cu[b + 1] - cu[b];
  const int64_t lo = (int64_t)cu[b] * H, hi = (int64_t)cu[b + 1] * H;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < H) {
    s_need[threadIdx.x] = pad_amount(bh_lens[b * H + threadIdx.x] + cnt[b * H + threadIdx.x], len, PS);
    s_have[threadIdx.x] = 0;
  }
  __syncthreads();
  for (int base = max(0, min(num_retain[b], max_sel)); base < max_sel; base += kSelThreads) {
    if (threadIdx.x == 0) {
      int pend = 0;
      for (int h = 0; h < H; ++h) pend += (s_have[h] < s_need[h]);
      s_pending = pend;
    }
    __syncthreads();
    if (s_pending == 0) break;
    const int i = base + threadIdx.x;
    int64_t sel = -1;
    int head = -1;
    if (i < max_sel) {
      sel = (int64_t)indices[(int64_t)b * max_sel + i];
      if (sel >= lo && sel < hi) head = (int)(sel % H); else sel = -1;
    }
    // per-head rank inside this step (H <= 16 ballots)
    int wrank = 0;
    for (int h = 0; h < H; ++h) {
      const uint32_t m = __ballot_sync(0xffffffffu, head == h);
      if (head == h) wrank = __popc(m & ((1u << lane) - 1u));
      if (lane == 0) s_wcnt[warp][h] = __popc(m);
    }
    __syncthreads();
    if (head >= 0) {
      int before = s_have[head];
      for (int w = 0; w < warp; ++w) before += s_wcnt[w][head];
      if (before + wrank < s_need[head]) keep[sel] = 1;
    }
    __syncthreads();
    if (threadIdx.x < H) {
      int tot = 0;
      for (int w = 0; w < 32; ++w) tot += s_wcnt[w][threadIdx.x];
      s_have[threadIdx.x] += tot;
    }
    __syncthreads();
  }
}

}  // namespace cvllm

using namespace cvllm;

// ---------------------------------------------------------------------------
// host wrappers
// ---------------------------------------------------------------------------
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" size_t cvllm_rank_scores_workspace_bytes(int B, int H, int max_k_len) {
  if (B <= 0 || H <= 0 || max_k_len <= 0) return 0;
  const size_t total = (size_t)max_k_len * H;
  const size_t tiles = (total + kSortTile - 1) / kSortTile;
  return 4 * align256((size_t)B * total * 4) + align256((size_t)B * 256 * tiles * 4);
}

extern "C" int cvllm_rank_scores(const float* scores, int64_t s_stride_n, const int32_t* cu_seqlens_k,
                                 int B, int H, int max_k_len, int top_k, int64_t* out_indices,
                                 void* workspace, size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && H > 0 && max_k_len >= 0 && top_k >= 0);
  CVLLM_CHECK_ARG(scores && cu_seqlens_k);
  const int64_t total64 = (int64_t)max_k_len * H;
  CVLLM_CHECK_SUPPORTED(total64 < (1ll << 30));
  const int total = (int)total64;
  const int k_eff = top_k < total ? top_k : total;
  if (k_eff == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(out_indices && workspace);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_rank_scores_workspace_bytes(B, H, max_k_len));
  const int tiles = cdiv(total, kSortTile);
  const size_t buf = align256((size_t)B * total * 4);
  char* w = reinterpret_cast<char*>(workspace);
  uint32_t* keys0 = reinterpret_cast<uint32_t*>(w);
  uint32_t* idx0 = reinterpret_cast<uint32_t*>(w + buf);
  uint32_t* keys1 = reinterpret_cast<uint32_t*>(w + 2 * buf);
  uint32_t* idx1 = reinterpret_cast<uint32_t*>(w + 3 * buf);
  uint32_t* hist = reinterpret_cast<uint32_t*>(w + 4 * buf);
  auto st = static_cast<cudaStream_t>(stream);
  const int gx = max(1, min(cdiv(total, 256), 4 * sm_count()));
  sort_init_kernel<<<dim3(gx, B), 256, 0, st>>>(scores, s_stride_n, cu_seqlens_k, H, total, keys0, idx0);
  CVLLM_CHECK_LAUNCH();
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 8 * pass;
    sort_hist_kernel<<<dim3(tiles, B), kSortThreads, 0, st>>>(keys0, total, tiles, shift, hist);
    sort_scan_kernel<<<B, kSelThreads, 0, st>>>(hist, tiles);
    sort_scatter_kernel<<<dim3(tiles, B), kSortThreads, 0, st>>>(keys0, idx0, keys1, idx1, total, tiles,
                                                                  shift, hist);
    CVLLM_CHECK_LAUNCHES(3);
    uint32_t* t = keys0; keys0 = keys1; keys1 = t;
    t = idx0; idx0 = idx1; idx1 = t;
  }
  sort_emit_kernel<<<dim3(max(1, min(cdiv(k_eff, 256), 4 * sm_count())), B), 256, 0, st>>>(
      idx0, cu_seqlens_k, H, total, k_eff, out_indices);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// cvllm_rank_scores_static: the same ranking with NO caller workspace and no allocation anywhere.  Added
// for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
// harnesses that forbid it); it has no counterpart in the reference.  The scratch of the radix sort (two key / index
// ping-pong buffers + the per-tile histograms) lives in a file-scope device array of kRankStaticBytes = 8 MiB, which
// every GPU that loads the library holds until exit.  8 MiB covers B x max_k_len x H <= ~510 K scores (e.g. one
// 32768-token sequence x 8 KV heads = 4.1 MiB, or 16 sequences x 4096 tokens x 8 heads); larger problems are
// refused with CVLLM_E_UNSUPPORTED and must use cvllm_rank_scores with their own workspace.
// Every word of the scratch that a call reads it has written first (sort_init fills keys0 / idx0, sort_hist writes
// every histogram cell), so nothing needs clearing between calls.  ALL calls on a device share the one array: two calls
// that overlap on the GPU (different streams, or graphs replayed concurrently) race; calls on one stream are ordered.
// In HOST code `g_rank_static` is only the symbol's host-side stand-in, never a device address: the address is taken
// once per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels get.
// ---------------------------------------------------------------------------------------------------------------
constexpr size_t kRankStaticBytes = (size_t)8 << 20;
static __device__ __align__(256) unsigned char g_rank_static[kRankStaticBytes];

static int rank_static_scratch(void** out) {
  constexpr int kMaxDev = 64;
  static std::atomic<void*> cached[kMaxDev];
  int dev = 0;
  CVLLM_CHECK_CUDA(cudaGetDevice(&dev));
  CVLLM_CHECK_ARG(dev >= 0 && dev < kMaxDev);
  void* p = cached[dev].load(std::memory_order_acquire);
  if (!p) {
    CVLLM_CHECK_CUDA(cudaGetSymbolAddress(&p, g_rank_static));  // this device's copy of the array
    CVLLM_CHECK_ARG(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0);
    cached[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return CVLLM_OK;
}

extern "C" size_t cvllm_rank_scores_static_max_bytes(void) { return kRankStaticBytes; }

extern "C" int cvllm_rank_scores_static(const float* scores, int64_t s_stride_n, const int32_t* cu_seqlens_k, int B,
                                        int H, int max_k_len, int top_k, int64_t* out_indices,
                                        cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && H > 0 && max_k_len >= 0 && top_k >= 0);
  if ((int64_t)max_k_len * H == 0 || top_k == 0) return CVLLM_OK;
  const size_t need = cvllm_rank_scores_workspace_bytes(B, H, max_k_len);
  if (need > kRankStaticBytes) {
    char msg[200];
    snprintf(msg, sizeof(msg),
             "cvllm_rank_scores_static: B=%d x max_k_len=%d x H=%d needs %zu bytes of scratch, the built-in array "
             "holds %zu; use cvllm_rank_scores with a caller workspace",
             B, max_k_len, H, need, kRankStaticBytes);
    set_last_error(msg);
    return CVLLM_E_UNSUPPORTED;
  }
  void* scratch = nullptr;
  const int rc = rank_static_scratch(&scratch);
  if (rc != CVLLM_OK) return rc;
  return cvllm_rank_scores(scores, s_stride_n, cu_seqlens_k, B, H, max_k_len, top_k, out_indices, scratch,
                           kRankStaticBytes, stream);
}

template <int V>
static void launch_gather(const void* keys, const void* vals, int64_t sk_n, int64_t sk_h, int64_t sv_n,
                          int64_t sv_h, const int32_t* pos, const int32_t* cu, const int32_t* bm,
                          const int32_t* pt, void* kc, void* vc, int B, int N, int H, int PS, int NLP,
                          cudaStream_t st) {
  const int rows_per_cta = 256 / V;
  const int64_t total_rows = (int64_t)N * H;
  int grid = (int)((total_rows + rows_per_cta - 1) / rows_per_cta);
  const int cap = sm_count() * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  gather_rows_kernel<V><<<grid, 256, 0, st>>>((const uint4*)keys, (const uint4*)vals, sk_n, sk_h, sv_n,
                                              sv_h, pos, cu, bm, pt, (uint4*)kc, (uint4*)vc, B, N, H,
                                              PS, NLP);
}

static int gather_dispatch(int vpr, const void* keys, const void* vals, int64_t sk_n, int64_t sk_h,
                           int64_t sv_n, int64_t sv_h, const int32_t* pos, const int32_t* cu,
                           const int32_t* bm, const int32_t* pt, void* kc, void* vc, int B, int N, int H,
                           int PS, int NLP, cudaStream_t st) {
#define G(V) launch_gather<V>(keys, vals, sk_n, sk_h, sv_n, sv_h, pos, cu, bm, pt, kc, vc, B, N, H, PS, NLP, st)
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
  return CVLLM_OK;
}

int cvllm_check_store_args(const void* keys, const void* vals, int64_t sk_n, int64_t sk_h, int64_t sv_n,
                            int64_t sv_h, void* k_cache, void* v_cache, int D, int elem_bytes, int H,
                            int* vpr_out, int64_t* per16_out) {
  CVLLM_CHECK_ARG(elem_bytes == 1 || elem_bytes == 2 || elem_bytes == 4);
  CVLLM_CHECK_SUPPORTED(H <= 16);
  const int row_bytes = D * elem_bytes;
  CVLLM_CHECK_SUPPORTED(row_bytes % 16 == 0 && row_bytes / 16 <= 32 &&
                        ((row_bytes / 16) & (row_bytes / 16 - 1)) == 0);
  const int64_t per16 = 16 / elem_bytes;
  CVLLM_CHECK_ALIGN(aligned16(keys) && aligned16(vals) && aligned16(k_cache) && aligned16(v_cache));
  CVLLM_CHECK_ALIGN(sk_n % per16 == 0 && sk_h % per16 == 0 && sv_n % per16 == 0 && sv_h % per16 == 0);
  *vpr_out = row_bytes / 16;
  *per16_out = per16;
  return CVLLM_OK;
}

extern "C" size_t cvllm_store_topk_workspace_bytes(int B, int H, int N, int max_sel) {
  (void)max_sel;
  if (B <= 0 || H <= 0 || N < 0) return 0;
  return align256((size_t)N * H) + align256((size_t)N * H * 4) + align256((size_t)B * H * 4);
}

extern "C" int cvllm_prefill_store_topk_idx(
    const void* keys, const void* vals, int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
    const void* indices, int idx_bytes, int max_sel, const int32_t* num_retain,
    const int32_t* cu_seqlens_k, const int32_t* batch_mapping, int32_t* bh_lens,
    const int32_t* page_table, void* k_cache, void* v_cache, int B, int N, int HKV, int D, int PS,
    int NLP, int elem_bytes, int pad_to_page, int reserved_batch, void* workspace,
    size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(B > 0 && N >= 0 && HKV > 0 && D > 0 && PS > 0 && NLP > 0 && max_sel >= 0);
  CVLLM_CHECK_ARG(keys && vals && num_retain && cu_seqlens_k && batch_mapping && bh_lens && page_table &&
                  k_cache && v_cache && workspace);
  CVLLM_CHECK_ARG(idx_bytes == 4 || idx_bytes == 8);
  CVLLM_CHECK_ARG(max_sel == 0 || indices);
  if (N == 0) return CVLLM_OK;
  int vpr;
  int64_t per16;
  const int rc = cvllm_check_store_args(keys, vals, sk_n, sk_h, sv_n, sv_h, k_cache, v_cache, D, elem_bytes,
                                  HKV, &vpr, &per16);
  if (rc != CVLLM_OK) return rc;
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_store_topk_workspace_bytes(B, HKV, N, max_sel));
  char* w = reinterpret_cast<char*>(workspace);
  uint8_t* keep = reinterpret_cast<uint8_t*>(w);
  int32_t* pos = reinterpret_cast<int32_t*>(w + align256((size_t)N * HKV));
  int32_t* cnt = reinterpret_cast<int32_t*>(w + align256((size_t)N * HKV) + align256((size_t)N * HKV * 4));
  auto st = static_cast<cudaStream_t>(stream);
  CVLLM_CHECK_CUDA(cudaMemsetAsync(keep, 0, (size_t)N * HKV, st));
  CVLLM_CHECK_CUDA(cudaMemsetAsync(cnt, 0, (size_t)B * HKV * 4, st));
  if (max_sel > 0) {
    const int gx = max(1, min(cdiv(max_sel, 256), 2 * sm_count()));
    if (idx_bytes == 8) {
      mark_topk_kernel<int64_t><<<dim3(gx, B), 256, 0, st>>>((const int64_t*)indices, max_sel, num_retain,
                                                             cu_seqlens_k, batch_mapping, HKV,
                                                             reserved_batch, keep, cnt);
      if (pad_to_page)
        pad_walk_kernel<int64_t><<<B, kSelThreads, 0, st>>>((const int64_t*)indices, max_sel, num_retain,
                                                            cu_seqlens_k, batch_mapping, bh_lens, cnt,
                                                            HKV, PS, reserved_batch, keep);
    } else {
      mark_topk_kernel<int32_t><<<dim3(gx, B), 256, 0, st>>>((const int32_t*)indices, max_sel, num_retain,
                                                             cu_seqlens_k, batch_mapping, HKV,
                                                             reserved_batch, keep, cnt);
      if (pad_to_page)
        pad_walk_kernel<int32_t><<<B, kSelThreads, 0, st>>>((const int32_t*)indices, max_sel, num_retain,
                                                            cu_seqlens_k, batch_mapping, bh_lens, cnt,
                                                            HKV, PS, reserved_batch, keep);
    }
    CVLLM_CHECK_LAUNCHES(pad_to_page ? 2 : 1);
  }
  compact_head_kernel<1><<<dim3(HKV, B), kSelThreads, 0, st>>>(nullptr, 0, keep, cu_seqlens_k,
                                                               batch_mapping, nullptr, bh_lens, HKV,
                                                               reserved_batch, pos);
  CVLLM_CHECK_LAUNCH();
  const int g = gather_dispatch(vpr, keys, vals, sk_n / per16, sk_h / per16, sv_n / per16, sv_h / per16,
                                pos, cu_seqlens_k, batch_mapping, page_table, k_cache, v_cache, B, N,
                                HKV, PS, NLP, st);
  if (g != CVLLM_OK) return g;
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

