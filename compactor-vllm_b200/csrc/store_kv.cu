// Paged-cache store kernels: decode append, uncompressed prefill store, and
// the page-table append used by the allocator.
//
// Reference behaviour: kv_cache/store_kv_cache.py:251-466 and
// kv_cache/write_page_table.py:6-95.  All three are pure HBM-bound row copies
// (a row = D elements of one (token, head)); rows move as 16-byte vectors,
// one row per group of D*elem/16 lanes, so a warp moves full 128-byte lines.
#include "common.cuh"

namespace cvllm {

// ---------------------------------------------------------------------------
// decode_store_kv: one row per (b, h)
// ---------------------------------------------------------------------------
template <int VEC_PER_ROW>
__global__ void decode_store_kernel(const uint4* __restrict__ key, const uint4* __restrict__ value,
                                    int64_t sk_b, int64_t sk_h, int64_t sv_b, int64_t sv_h,
                                    const int32_t* __restrict__ batch_mapping,
                                    int32_t* __restrict__ bh_lens,
                                    const int32_t* __restrict__ page_table, uint4* __restrict__ k_cache,
                                    uint4* __restrict__ v_cache, int B, int HKV, int PS, int NLP,
                                    int reserved_batch, int lens_by_slot) {
  // strides here are already in uint4 units
  const int rows = B * HKV;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int row = t / VEC_PER_ROW;
  const int vec = t % VEC_PER_ROW;
  if (row >= rows) return;
  const int b = row / HKV, h = row % HKV;
  const int mapped = batch_mapping[b];
  if (mapped == reserved_batch) return;
  // lens_by_slot: bh_lens is the cache's [Bmax+1, HKV] table (no gathered copy to scatter back)
  const int li = lens_by_slot ? mapped * HKV + h : row;
  const int len = bh_lens[li];
  const int lp = len / PS, off = len - lp * PS;
  const int64_t phys = page_table[((int64_t)mapped * HKV + h) * NLP + lp];
  const int64_t dst = (phys * PS + off) * VEC_PER_ROW + vec;
  k_cache[dst] = key[b * sk_b + h * sk_h + vec];
  v_cache[dst] = value[b * sv_b + h * sv_h + vec];
  // every lane of the row group has read `len` before any of them writes it:
  // the group lives inside one warp (VEC_PER_ROW <= 32).
  __syncwarp();
  if (vec == 0) bh_lens[li] = len + 1;
}

// ---------------------------------------------------------------------------
// prefill_store_all_kv: every (token, head) row of the packed batch
// ---------------------------------------------------------------------------
template <int VEC_PER_ROW>
__global__ void store_all_kernel(const uint4* __restrict__ keys, const uint4* __restrict__ vals,
                                 int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
                                 const int32_t* __restrict__ cu, const int32_t* __restrict__ batch_mapping,
                                 const int32_t* __restrict__ bh_lens,
                                 const int32_t* __restrict__ page_table, uint4* __restrict__ k_cache,
                                 uint4* __restrict__ v_cache, int B, int N, int HKV, int PS, int NLP) {
  constexpr int ROWS_PER_CTA = 256 / VEC_PER_ROW;
  const int vec = threadIdx.x % VEC_PER_ROW;
  const int rin = threadIdx.x / VEC_PER_ROW;
  const int64_t total_rows = (int64_t)N * HKV;
  for (int64_t row = (int64_t)blockIdx.x * ROWS_PER_CTA + rin; row < total_rows;
       row += (int64_t)gridDim.x * ROWS_PER_CTA) {
    const int n = (int)(row / HKV), h = (int)(row % HKV);
    // sequence lookup: binary search over cu (B is small)
    int lo = 0, hi = B;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cu[mid] <= n) lo = mid; else hi = mid;
    }
    const int b = lo;
    if (n >= cu[B]) continue;
    const int tok = n - cu[b];
    const int pos = bh_lens[b * HKV + h] + tok;
    const int lp = pos / PS, off = pos - lp * PS;
    const int64_t phys = page_table[((int64_t)batch_mapping[b] * HKV + h) * NLP + lp];
    const int64_t dst = (phys * PS + off) * VEC_PER_ROW + vec;
    const uint4 kk = ldg_nc_v4(keys + n * sk_n + h * sk_h + vec);
    const uint4 vv = ldg_nc_v4(vals + n * sv_n + h * sv_h + vec);
    stg_na_v4(k_cache + dst, kk);
    stg_na_v4(v_cache + dst, vv);
  }
}

__global__ void bump_lens_kernel(const int32_t* __restrict__ cu, int32_t* __restrict__ bh_lens, int B,
                                 int HKV) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * HKV) return;
  const int b = i / HKV;
  bh_lens[i] += cu[b + 1] - cu[b];
}

// ---------------------------------------------------------------------------
// scatter_to_page_table: one CTA, exclusive scan of add_pages then append
// ---------------------------------------------------------------------------
__global__ void scatter_pages_kernel(const int32_t* __restrict__ add_pages,
                                     const int32_t* __restrict__ new_phys,
                                     const int32_t* __restrict__ curr_pages,
                                     int32_t* __restrict__ page_table, int64_t s_l, int64_t s_h,
                                     int64_t s_p, int L, int H, int max_pages) {
  extern __shared__ int32_t s_cum[];  // [L*H + 1]
  const int n = L * H;
  // serial-per-chunk scan: n = layers*heads is a few hundred
  __shared__ int32_t s_chunk[1024];
  const int per = cdiv(n, blockDim.x);
  int local = 0;
  const int beg = threadIdx.x * per, end = min(n, beg + per);
  for (int i = beg; i < end; ++i) local += max(add_pages[i], 0);
  s_chunk[threadIdx.x] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int i = 0; i < (int)blockDim.x; ++i) {
      const int v = s_chunk[i];
      s_chunk[i] = run;
      run += v;
    }
  }
  __syncthreads();
  int run = s_chunk[threadIdx.x];
  for (int i = beg; i < end; ++i) {
    s_cum[i] = run;
    run += max(add_pages[i], 0);
  }
  __syncthreads();
  for (int lh = threadIdx.x; lh < n; lh += blockDim.x) {
    const int ap = add_pages[lh];
    if (ap <= 0) continue;
    const int l = lh / H, h = lh % H;
    const int base = s_cum[lh], cp = curr_pages[lh];
    for (int i = 0; i < ap; ++i) {
      const int lp = cp + i;
      if (lp < max_pages) page_table[l * s_l + h * s_h + lp * s_p] = new_phys[base + i];
    }
  }
}

}  // namespace cvllm

using namespace cvllm;

namespace cvllm {
int launch_decode_store(const void* key, const void* value, int64_t sk_b, int64_t sk_h, int64_t sv_b,
                        int64_t sv_h, const int32_t* batch_mapping, int32_t* bh_lens,
                        const int32_t* page_table, void* k_cache, void* v_cache, int B, int HKV, int D,
                        int PS, int NLP, int elem_bytes, int reserved_batch, int lens_by_slot,
                        cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(key && value && batch_mapping && bh_lens && page_table && k_cache && v_cache);
  CVLLM_CHECK_ARG(B >= 0 && HKV > 0 && D > 0 && PS > 0 && NLP > 0);
  CVLLM_CHECK_ARG(elem_bytes == 1 || elem_bytes == 2 || elem_bytes == 4);
  if (B == 0) return CVLLM_OK;
  const int row_bytes = D * elem_bytes;
  CVLLM_CHECK_SUPPORTED(row_bytes % 16 == 0 && row_bytes / 16 <= 32 &&
                        ((row_bytes / 16) & (row_bytes / 16 - 1)) == 0);
  const int64_t per16 = 16 / elem_bytes;
  CVLLM_CHECK_ALIGN(aligned16(key) && aligned16(value) && aligned16(k_cache) && aligned16(v_cache));
  CVLLM_CHECK_ALIGN(sk_b % per16 == 0 && sk_h % per16 == 0 && sv_b % per16 == 0 && sv_h % per16 == 0);
  const int vpr = row_bytes / 16;
  const int threads = 128;
  const int total = B * HKV * vpr;
  const int grid = cdiv(total, threads);
  auto st = static_cast<cudaStream_t>(stream);
#define LAUNCH(V)                                                                                  \
  decode_store_kernel<V><<<grid, threads, 0, st>>>(                                                \
      (const uint4*)key, (const uint4*)value, sk_b / per16, sk_h / per16, sv_b / per16,            \
      sv_h / per16, batch_mapping, bh_lens, page_table, (uint4*)k_cache, (uint4*)v_cache, B, HKV,  \
      PS, NLP, reserved_batch, lens_by_slot)
  switch (vpr) {
    case 1: LAUNCH(1); break;
    case 2: LAUNCH(2); break;
    case 4: LAUNCH(4); break;
    case 8: LAUNCH(8); break;
    case 16: LAUNCH(16); break;
    case 32: LAUNCH(32); break;
    default: return CVLLM_E_UNSUPPORTED;
  }
#undef LAUNCH
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
}  // namespace cvllm

extern "C" int cvllm_decode_store_kv(const void* key, const void* value, int64_t sk_b, int64_t sk_h,
                                     int64_t sv_b, int64_t sv_h, const int32_t* batch_mapping,
                                     int32_t* bh_lens, const int32_t* page_table, void* k_cache,
                                     void* v_cache, int B, int HKV, int D, int PS, int NLP,
                                     int elem_bytes, int reserved_batch, cvllm_stream_t stream) {
  return cvllm::launch_decode_store(key, value, sk_b, sk_h, sv_b, sv_h, batch_mapping, bh_lens, page_table,
                                    k_cache, v_cache, B, HKV, D, PS, NLP, elem_bytes, reserved_batch, 0, stream);
}

extern "C" int cvllm_prefill_store_all(const void* keys, const void* vals, int64_t sk_n, int64_t sk_h,
                                       int64_t sv_n, int64_t sv_h, const int32_t* cu_seqlens_k,
                                       const int32_t* batch_mapping, int32_t* bh_lens,
                                       const int32_t* page_table, void* k_cache, void* v_cache, int B,
                                       int N, int HKV, int D, int PS, int NLP, int elem_bytes,
                                       cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(keys && vals && cu_seqlens_k && batch_mapping && bh_lens && page_table && k_cache &&
                  v_cache);
  CVLLM_CHECK_ARG(B > 0 && N >= 0 && HKV > 0 && D > 0 && PS > 0 && NLP > 0);
  CVLLM_CHECK_ARG(elem_bytes == 1 || elem_bytes == 2 || elem_bytes == 4);
  const int row_bytes = D * elem_bytes;
  CVLLM_CHECK_SUPPORTED(row_bytes % 16 == 0 && row_bytes / 16 <= 32 &&
                        ((row_bytes / 16) & (row_bytes / 16 - 1)) == 0);
  const int64_t per16 = 16 / elem_bytes;
  CVLLM_CHECK_ALIGN(aligned16(keys) && aligned16(vals) && aligned16(k_cache) && aligned16(v_cache));
  CVLLM_CHECK_ALIGN(sk_n % per16 == 0 && sk_h % per16 == 0 && sv_n % per16 == 0 && sv_h % per16 == 0);
  auto st = static_cast<cudaStream_t>(stream);
  if (N > 0) {
    const int vpr = row_bytes / 16;
    const int rows_per_cta = 256 / vpr;
    const int64_t total_rows = (int64_t)N * HKV;
    int grid = (int)((total_rows + rows_per_cta - 1) / rows_per_cta);
    const int cap = sm_count() * 8;
    if (grid > cap) grid = cap;
#define LAUNCH(V)                                                                                  \
  store_all_kernel<V><<<grid, 256, 0, st>>>((const uint4*)keys, (const uint4*)vals, sk_n / per16,   \
                                            sk_h / per16, sv_n / per16, sv_h / per16, cu_seqlens_k, \
                                            batch_mapping, bh_lens, page_table, (uint4*)k_cache,    \
                                            (uint4*)v_cache, B, N, HKV, PS, NLP)
    switch (vpr) {
      case 1: LAUNCH(1); break;
      case 2: LAUNCH(2); break;
      case 4: LAUNCH(4); break;
      case 8: LAUNCH(8); break;
      case 16: LAUNCH(16); break;
      case 32: LAUNCH(32); break;
      default: return CVLLM_E_UNSUPPORTED;
    }
#undef LAUNCH
    CVLLM_CHECK_LAUNCH();
  }
  bump_lens_kernel<<<cdiv(B * HKV, 128), 128, 0, st>>>(cu_seqlens_k, bh_lens, B, HKV);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

extern "C" int cvllm_scatter_pages(const int32_t* add_pages, const int32_t* new_phys,
                                   const int32_t* curr_pages, int32_t* page_table, int64_t pt_stride_l,
                                   int64_t pt_stride_h, int64_t pt_stride_p, int L, int H,
                                   int max_pages_per_head, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(L >= 0 && H >= 0);
  if (L == 0 || H == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(add_pages && new_phys && curr_pages && page_table);
  const size_t smem = (size_t)(L * H + 1) * sizeof(int32_t);
  CVLLM_CHECK_SUPPORTED(smem <= 40 * 1024);
  scatter_pages_kernel<<<1, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      add_pages, new_phys, curr_pages, page_table, pt_stride_l, pt_stride_h, pt_stride_p, L, H,
      max_pages_per_head);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
