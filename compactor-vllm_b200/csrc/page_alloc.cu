// Device-side page allocator: the reference keeps per-layer min-heaps of free physical
// pages on the HOST (kv_cache/page_table.py:111-115) and pays two device->host syncs (`.all()`,
// `.tolist()`), a Python heap loop and a host->device copy per reservation (:144-198), three syncs per
// reclaim (:200-267).  Here the free set of a layer is a bitmap in HBM and "pop the n smallest pages"
// is a popcount scan over it, so reserve / reclaim / free are stream-ordered launches with no host
// round trip: capturable in a CUDA graph next to the decode step.
//
// Same results as the heaps, bit for bit (tests replay the reference's trace): a min-heap hands out the
// smallest free ids in ascending order; heads take consecutive pops in head order
// (kv_cache/write_page_table.py:62-95), layers are all-or-nothing (:171-176), sequences of one call are
// served in order, each seeing the pages the earlier ones left.
//
//   plan  (1 CTA)        per sequence: pages to add per (layer, head), status, per-layer rank bases
//   take  (1 CTA/layer)  n smallest set bits of the layer's bitmap -> page_table rows, bh_num_pages
//   give  (1 CTA/layer)  reclaim / free: pages beyond ceil((len+future)/PS) back into the bitmap
#include "common.cuh"

namespace cvllm {

constexpr int kPgThreads = 256;

struct PagePlan {            // workspace layout
  int32_t* add;              // [nb][L][H] pages to add
  int32_t* base;             // [nb][L]    pages of layer l taken by sequences before i
  int32_t* total;            // [L]        pages of layer l taken by the whole call
  int32_t* taken;            // [L][cap]   the page ids, ascending
  int cap;
};

__device__ __forceinline__ int pg_cdiv(int a, int b) { return (a + b - 1) / b; }

// A slot named twice in one call is served once (its first occurrence): the second reserve of the same tokens
// is a no-op on the host allocator too, and counting its pages twice would desynchronise free_count from the
// bitmap.  nb is a decode batch (<= a few hundred), so the scan is cheaper than a visited table.
__device__ __forceinline__ bool pg_first_occurrence(const int32_t* __restrict__ slots, int i) {
  const int b = slots[i];
  for (int j = 0; j < i; ++j)
    if (slots[j] == b) return false;
  return true;
}

__global__ void __launch_bounds__(kPgThreads) pages_plan_kernel(
    const int32_t* __restrict__ free_count, const int32_t* __restrict__ bh_seq_lens,
    const int32_t* __restrict__ bh_num_pages, const int32_t* __restrict__ slots, int nb, int add_tokens,
    int32_t* __restrict__ status, PagePlan plan, int L, int Bs, int H, int P, int PS, int reserved) {
  extern __shared__ int32_t pg_sh[];  // avail[L] | need[L]
  int32_t* avail = pg_sh;
  int32_t* need = pg_sh + L;
  __shared__ int s_short, s_over, s_verdict;
  const int tid = threadIdx.x;
  for (int l = tid; l < L; l += kPgThreads) { avail[l] = free_count[l]; plan.total[l] = 0; }
  __syncthreads();
  for (int i = 0; i < nb; ++i) {
    const int b = slots[i];
    const bool live = b != reserved && b >= 0 && b < Bs && pg_first_occurrence(slots, i);
    if (tid == 0) { s_short = 0; s_over = 0; }
    for (int l = tid; l < L; l += kPgThreads) need[l] = 0;
    __syncthreads();
    for (int e = tid; e < L * H; e += kPgThreads) {
      const int l = e / H, h = e - l * H;
      int add = 0;
      if (live) {
        const int64_t o = ((int64_t)l * Bs + b) * H + h;
        const int pages = bh_num_pages[o];
        const int missing = bh_seq_lens[o] + add_tokens - pages * PS;
        if (missing > 0) {
          add = pg_cdiv(missing, PS);
          s_short = 1;
          if (pages + add > P) s_over = 1;
          atomicAdd(&need[l], add);
        }
      }
      plan.add[((int64_t)i * L + l) * H + h] = add;
    }
    __syncthreads();
    if (tid == 0) {
      int v = CVLLM_PAGES_OK;
      if (s_short) {
        if (s_over) v = CVLLM_PAGES_EXCEEDS_MAX_SEQUENCE_LENGTH;
        else
          for (int l = 0; l < L; ++l)
            if (need[l] > avail[l]) { v = CVLLM_PAGES_EXCEEDS_AVAILABLE_PAGES; break; }
      }
      s_verdict = v;
      status[i] = v;
    }
    __syncthreads();
    const bool ok = s_verdict == CVLLM_PAGES_OK;
    for (int l = tid; l < L; l += kPgThreads) {
      plan.base[(int64_t)i * L + l] = plan.total[l];
      if (ok) { plan.total[l] += need[l]; avail[l] -= need[l]; }
    }
    if (!ok)
      for (int e = tid; e < L * H; e += kPgThreads) plan.add[(int64_t)i * L * H + e] = 0;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kPgThreads) pages_take_kernel(
    uint32_t* __restrict__ free_bits, int32_t* __restrict__ free_count, int32_t* __restrict__ page_table,
    int32_t* __restrict__ bh_num_pages, const int32_t* __restrict__ slots, int nb, PagePlan plan, int L,
    int Bs, int H, int P, int W) {
  const int l = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int total = plan.total[l];
  if (total == 0) return;
  uint32_t* bits = free_bits + (int64_t)l * W;
  int32_t* taken = plan.taken + (int64_t)l * plan.cap;
  // thread t owns words [t*cw, (t+1)*cw): ascending thread order = ascending page id
  const int cw = (W + kPgThreads - 1) / kPgThreads;
  const int w0 = min(tid * cw, W), w1 = min(w0 + cw, W);
  int mine = 0;
  for (int w = w0; w < w1; ++w) mine += __popc(bits[w]);
  // exclusive block scan of `mine`
  __shared__ int s_warp[kPgThreads / 32];
  int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int before = incl - mine;
  for (int w = 0; w < warp; ++w) before += s_warp[w];
  // take this thread's bits while their rank is below total
  int rank = before;
  for (int w = w0; w < w1 && rank < total; ++w) {
    uint32_t word = bits[w], keep = word;
    while (word && rank < total) {
      const int bit = __ffs(word) - 1;
      word &= word - 1;
      keep &= ~(1u << bit);
      taken[rank++] = w * 32 + bit;
    }
    bits[w] = keep;
  }
  __syncthreads();
  // hand the ranks out: sequence i, head h, logical slot pages+j <- taken[base_i + sum_{h'<h} add + j]
  for (int e = tid; e < nb * H; e += kPgThreads) {
    const int i = e / H, h = e - i * H;
    const int32_t* add = plan.add + ((int64_t)i * L + l) * H;
    const int n = add[h];
    if (n == 0) continue;
    int r = plan.base[(int64_t)i * L + l];
    for (int hh = 0; hh < h; ++hh) r += add[hh];
    const int b = slots[i];
    const int64_t o = ((int64_t)l * Bs + b) * H + h;
    const int pages = bh_num_pages[o];
    int32_t* row = page_table + o * P;
    for (int j = 0; j < n; ++j) row[pages + j] = taken[r + j];
    bh_num_pages[o] = pages + n;
  }
  if (tid == 0) free_count[l] -= total;
}

__global__ void __launch_bounds__(kPgThreads) pages_give_kernel(
    uint32_t* __restrict__ free_bits, int32_t* __restrict__ free_count, const int32_t* __restrict__ page_table,
    int32_t* __restrict__ bh_seq_lens, int32_t* __restrict__ bh_num_pages, const int32_t* __restrict__ slots,
    int nb, int future_tokens, int free_all, unsigned long long* __restrict__ freed_out, int L, int Bs, int H,
    int P, int PS, int W, int reserved) {
  const int l = blockIdx.x, tid = threadIdx.x;
  uint32_t* bits = free_bits + (int64_t)l * W;
  int cnt = 0;
  for (int e = tid; e < nb * H; e += kPgThreads) {
    const int i = e / H, h = e - i * H;
    const int b = slots[i];
    if (b == reserved || b < 0 || b >= Bs || !pg_first_occurrence(slots, i)) continue;
    const int64_t o = ((int64_t)l * Bs + b) * H + h;
    const int alloc = bh_num_pages[o];
    const int used = free_all ? 0 : min(pg_cdiv(bh_seq_lens[o] + future_tokens, PS), alloc);
    const int32_t* row = page_table + o * P;
    for (int p = used; p < alloc; ++p) {
      const int page = row[p];
      atomicOr(&bits[page >> 5], 1u << (page & 31));
      ++cnt;
    }
    bh_num_pages[o] = used;
    if (free_all) bh_seq_lens[o] = 0;
  }
  cnt = (int)warp_sum((float)cnt);  // <= nb*H*P per thread group: exact in fp32 below 2^24
  __shared__ int s_cnt;
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  if ((tid & 31) == 0 && cnt) atomicAdd(&s_cnt, cnt);
  __syncthreads();
  if (tid == 0 && s_cnt) {
    free_count[l] += s_cnt;
    if (freed_out) atomicAdd(freed_out, (unsigned long long)s_cnt);
  }
}

__global__ void pages_init_kernel(uint32_t* __restrict__ free_bits, int32_t* __restrict__ free_count, int L,
                                  int W, int n_pages) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < (int64_t)L * W) {
    const int w = (int)(idx % W);
    const int lo = w * 32;
    uint32_t v = 0;
    if (lo + 32 <= n_pages) v = 0xffffffffu;
    else if (lo < n_pages) v = (1u << (n_pages - lo)) - 1u;
    free_bits[idx] = v;
  }
  if (idx < L) free_count[idx] = n_pages;
}

}  // namespace cvllm

using namespace cvllm;

static size_t pg_align(size_t x) { return (x + 255) & ~(size_t)255; }
static int pg_cap(int nb, int H, int P, int n_pages) {
  const long long worst = (long long)nb * H * P;
  return (int)(worst < n_pages ? worst : n_pages);
}

extern "C" size_t cvllm_pages_bitmap_words(int n_pages) { return (size_t)((n_pages + 31) / 32); }

extern "C" size_t cvllm_pages_workspace_bytes(int L, int H, int P, int n_pages, int nb) {
  if (L <= 0 || H <= 0 || nb <= 0) return 256;
  return pg_align((size_t)nb * L * H * 4) + pg_align((size_t)nb * L * 4) + pg_align((size_t)L * 4) +
         pg_align((size_t)L * pg_cap(nb, H, P, n_pages) * 4);
}

extern "C" int cvllm_pages_init(uint32_t* free_bits, int32_t* free_count, int L, int n_pages,
                                cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(free_bits && free_count && L > 0 && n_pages > 0);
  const int W = (n_pages + 31) / 32;
  const int64_t n = (int64_t)L * W;
  pages_init_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      free_bits, free_count, L, W, n_pages);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

extern "C" int cvllm_pages_reserve(uint32_t* free_bits, int32_t* free_count, int32_t* page_table,
                                   const int32_t* bh_seq_lens, int32_t* bh_num_pages, const int32_t* slots,
                                   int nb, int add_tokens, int32_t* status, int L, int num_slots, int H,
                                   int P, int PS, int n_pages, int reserved_batch, void* workspace,
                                   size_t workspace_bytes, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(nb >= 0 && L > 0 && num_slots > 0 && H > 0 && P > 0 && PS > 0 && n_pages > 0 && add_tokens >= 0);
  if (nb == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(free_bits && free_count && page_table && bh_seq_lens && bh_num_pages && slots && status && workspace);
  CVLLM_CHECK_ARG(workspace_bytes >= cvllm_pages_workspace_bytes(L, H, P, n_pages, nb));
  CVLLM_CHECK_SUPPORTED(2 * (size_t)L * 4 <= 48 * 1024);
  char* ws = reinterpret_cast<char*>(workspace);
  PagePlan plan;
  plan.add = reinterpret_cast<int32_t*>(ws);    ws += pg_align((size_t)nb * L * H * 4);
  plan.base = reinterpret_cast<int32_t*>(ws);   ws += pg_align((size_t)nb * L * 4);
  plan.total = reinterpret_cast<int32_t*>(ws);  ws += pg_align((size_t)L * 4);
  plan.taken = reinterpret_cast<int32_t*>(ws);
  plan.cap = pg_cap(nb, H, P, n_pages);
  auto st = static_cast<cudaStream_t>(stream);
  pages_plan_kernel<<<1, kPgThreads, 2 * (size_t)L * 4, st>>>(free_count, bh_seq_lens, bh_num_pages, slots, nb,
                                                              add_tokens, status, plan, L, num_slots, H, P, PS,
                                                              reserved_batch);
  pages_take_kernel<<<L, kPgThreads, 0, st>>>(free_bits, free_count, page_table, bh_num_pages, slots, nb, plan, L,
                                              num_slots, H, P, (n_pages + 31) / 32);
  CVLLM_CHECK_LAUNCHES(2);
  return CVLLM_OK;
}

extern "C" int cvllm_pages_release(uint32_t* free_bits, int32_t* free_count, const int32_t* page_table,
                                   int32_t* bh_seq_lens, int32_t* bh_num_pages, const int32_t* slots, int nb,
                                   int future_tokens, int free_all, unsigned long long* freed_pages_out, int L,
                                   int num_slots, int H, int P, int PS, int n_pages, int reserved_batch,
                                   cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(nb >= 0 && L > 0 && num_slots > 0 && H > 0 && P > 0 && PS > 0 && n_pages > 0 && future_tokens >= 0);
  if (nb == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(free_bits && free_count && page_table && bh_seq_lens && bh_num_pages && slots);
  pages_give_kernel<<<L, kPgThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      free_bits, free_count, page_table, bh_seq_lens, bh_num_pages, slots, nb, future_tokens, free_all,
      freed_pages_out, L, num_slots, H, P, PS, (n_pages + 31) / 32, reserved_batch);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
