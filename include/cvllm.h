/*
 * cvllm.h — C ABI of libcompactor_sm100.so
 *
 * Drop-in boundary for the attention + KV-compression hot path of
 * vnchari/compactor-vllm, rebuilt as hand-written sm_100a CUDA.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`
 *   - every launch function takes the CUDA stream (a cudaStream_t passed as
 *     void*) last, launches on it and returns without synchronising
 *   - return value: 0 = launched; negative = argument error detected on the
 *     host before any launch (see CVLLM_E_*); nothing is thrown
 *   - no torch types; TMA descriptors are rebuilt per call.  Process state is
 *     limited to: the launch counter, a per-(kernel, device) "dynamic shared
 *     memory limit raised" bit (so one process can drive several GPUs), a
 *     per-device SM-count cache, and the explicit test options below.  The
 *     library never reads the environment.
 *   - all index tensors are int32 unless stated; scores are fp32
 *   - strides are in ELEMENTS of the tensor's dtype
 *   - `dtype`: CVLLM_F16 or CVLLM_BF16 (the arithmetic is always fp32
 *     accumulate)
 *
 * Each entry point cites the reference interface it replaces
 * (paths relative to the reference's src/compactor_vllm/).
 */
#ifndef CVLLM_H_
#define CVLLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CVLLM_ABI_VERSION 1

enum { CVLLM_F16 = 0, CVLLM_BF16 = 1 };

enum {
  CVLLM_OK = 0,
  CVLLM_E_ARG = -1,         /* bad shape / size / null pointer            */
  CVLLM_E_UNSUPPORTED = -2, /* head_dim / group size / page size combo    */
  CVLLM_E_WORKSPACE = -3,   /* workspace too small                        */
  CVLLM_E_CUDA = -4,        /* launch or driver-entry-point failure       */
  CVLLM_E_ALIGN = -5        /* pointer / stride not 16-byte aligned       */
};

typedef void* cvllm_stream_t;

/* ABI version + last CUDA error string (host, static storage). */
int cvllm_abi_version(void);
const char* cvllm_last_error(void);
/* Number of kernels this library has launched in this process (monotonic).
 * Lets a caller report how many of its kernels ran inside a timed region. */
uint64_t cvllm_launch_count(void);
/* Test / profiling switches (never needed by a product caller; defaults = product behaviour):
 *   "prefill_v1" 1|0         force the one-query-head-per-CTA prefill kernel (the path odd GQA groups take)
 *   "prefill_threshold" t    lazy-rescale threshold of the prefill softmax in log2 units (0 exercises the
 *                            O-correction path on every growing row max); negative = built-in default
 *   "decode_mode" m          decode_attn.cu profiling modes (0 = product), "decode_slots" n  TMA ring depth cap
 *   "chunk_scores_v1" / "snap_v1" 1|0   the mma.sync scorers instead of the tcgen05 ones (A/B)
 *   "pdl" 1|0                launch the selection chain with programmatic dependent launch (A/B; off by default)
 *   "grs_pairs" n            cap the CTA pairs of cvllm_gemm_allreduce (several virtual ranks on one GPU: the tests)
 *   "grs_rows_per_round" r   rows per reduce round of cvllm_gemm_allreduce (multiple of 256; 0 = default 256)
 *   "grs_probe" mask         timing probes of cvllm_gemm_allreduce (results are wrong on purpose: 1 no TMA loads,
 *                            2 no UMMA, 4 no stores, 8 no reduce, 16 no L2 hints)
 *   "grs_debug_host_ptr" p   pinned host buffer (64 words per CTA) for cvllm_gemm_allreduce's timeout records
 * Returns CVLLM_E_ARG for an unknown name. */
int cvllm_set_test_option(const char* name, double value);
/* Profiling only: in a library built with -DCVLLM_DECODE_TRACE (tools/trace_decode.py) the next decode launches
 * stamp 8 globaltimer values per CTA (entry, lengths read, first unit landed, streaming done, merge done, role,
 * ticket taken, slots folded) into dev_buf (8 * #CTAs uint64); NULL = off.  Returns -1 in the normal build. */
int cvllm_debug_decode_trace(void* dev_buf);

/* ------------------------------------------------------------------------
 * Store path
 * --------------------------------------------------------------------- */

/* decode_store_kv — kv_cache/store_kv_cache.py:374-466 (kernel :374-416).
 * Append key[b,h,:], value[b,h,:] at logical position bh_lens[b,h] of head
 * (batch_mapping[b], h); bh_lens[b,h] += 1.  Rows with
 * batch_mapping[b] == reserved_batch are skipped. */
int cvllm_decode_store_kv(const void* key, const void* value,
                          int64_t sk_b, int64_t sk_h, int64_t sv_b, int64_t sv_h,
                          const int32_t* batch_mapping, int32_t* bh_lens,
                          const int32_t* page_table, void* k_cache, void* v_cache,
                          int B, int HKV, int D, int PS, int NLP, int elem_bytes,
                          int reserved_batch, cvllm_stream_t stream);

/* prefill_store_all_kv — kv_cache/store_kv_cache.py:251-371.
 * Scatter every (token, head) row of the packed batch to logical position
 * bh_lens[b,h] + t, then bh_lens[b,:] += len_b (the reference's
 * `bh_lens += cu_seqlens_k.diff()[:,None]`, :371). */
int cvllm_prefill_store_all(const void* keys, const void* vals,
                            int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
                            const int32_t* cu_seqlens_k, const int32_t* batch_mapping,
                            int32_t* bh_lens, const int32_t* page_table,
                            void* k_cache, void* v_cache,
                            int B, int N, int HKV, int D, int PS, int NLP,
                            int elem_bytes, cvllm_stream_t stream);

/* scatter_to_page_table — kv_cache/write_page_table.py:6-95.
 * page_table[l,h,curr[l,h]+i] = new_phys[cum[l,h]+i] for i < add[l,h];
 * the exclusive prefix sum `cum` is computed inside (replaces torch.cumsum,
 * :42-44).  pt_stride_* in int32 elements. */
int cvllm_scatter_pages(const int32_t* add_pages, const int32_t* new_phys,
                        const int32_t* curr_pages, int32_t* page_table,
                        int64_t pt_stride_l, int64_t pt_stride_h, int64_t pt_stride_p,
                        int L, int H, int max_pages_per_head, cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * Device-side page allocator — PagedKVCache.reserve_tokens /
 * reclaim_pages / free_batch, kv_cache/page_table.py:144-291, without the
 * host heaps: the free set of a layer is a bitmap (bit set = page free), "pop
 * the n smallest ids" is a popcount scan.  Same page ids as the reference's
 * min-heaps.  page_table int32 [L, num_slots, H, P]; bh_seq_lens /
 * bh_num_pages int32 [L, num_slots, H]; free_bits uint32 [L, words];
 * free_count int32 [L]; slots int32 [nb] (entries == reserved_batch are
 * skipped); sequences of one call are served in order.
 * --------------------------------------------------------------------- */
enum {
  CVLLM_PAGES_OK = 0,
  CVLLM_PAGES_EXCEEDS_MAX_SEQUENCE_LENGTH = 1,
  CVLLM_PAGES_EXCEEDS_AVAILABLE_PAGES = 2
};
size_t cvllm_pages_bitmap_words(int n_pages);
size_t cvllm_pages_workspace_bytes(int L, int H, int P, int n_pages, int nb);
int cvllm_pages_init(uint32_t* free_bits, int32_t* free_count, int L, int n_pages,
                     cvllm_stream_t stream);
/* status int32 [nb] <- CVLLM_PAGES_*; a sequence that fails takes nothing
 * (all-or-nothing across layers, page_table.py:171-176). */
int cvllm_pages_reserve(uint32_t* free_bits, int32_t* free_count, int32_t* page_table,
                        const int32_t* bh_seq_lens, int32_t* bh_num_pages, const int32_t* slots,
                        int nb, int add_tokens, int32_t* status, int L, int num_slots, int H,
                        int P, int PS, int n_pages, int reserved_batch, void* workspace,
                        size_t workspace_bytes, cvllm_stream_t stream);
/* free_all == 0: reclaim_pages (keep ceil((len + future_tokens)/PS) pages per
 * head); free_all != 0: free_batch (all pages back, lens and counts zeroed).
 * freed_pages_out (may be NULL) is incremented by the pages returned. */
int cvllm_pages_release(uint32_t* free_bits, int32_t* free_count, const int32_t* page_table,
                        int32_t* bh_seq_lens, int32_t* bh_num_pages, const int32_t* slots, int nb,
                        int future_tokens, int free_all, unsigned long long* freed_pages_out,
                        int L, int num_slots, int H, int P, int PS, int n_pages,
                        int reserved_batch, cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * The step after the path: the tensor-parallel all-reduce
 * behind o_proj / down_proj — layers/linear.py:149-153 `dist.all_reduce(y)`.
 * One-shot all-reduce over NVLink peer memory for decode-size messages;
 * every rank reads all peers' buffers and sums in rank order (fp32), so all
 * ranks hold identical bits.  Buffers come from cvllm_ipc_alloc (cudaMalloc +
 * cudaIpc handle, flags zeroed); the 64-byte handles are exchanged by the
 * caller (torch.distributed) and opened with cvllm_ipc_open.  A rank's payload
 * starts cvllm_allreduce_flag_bytes() into its buffer.  *_host = host pointers.
 * --------------------------------------------------------------------- */
size_t cvllm_allreduce_flag_bytes(void);
int cvllm_ipc_alloc(size_t payload_bytes, void** ptr_out_host, void* handle_out_host);
int cvllm_ipc_open(const void* handle_host, void** ptr_out_host);
int cvllm_ipc_close(void* peer_ptr);
int cvllm_ipc_free(void* own_ptr);
/* out[i] = sum over ranks of payload_r[i]; numel 16-bit elements, % 8 == 0,
 * the same on every rank; world <= 8.  peer_bases_host[r] = rank r's buffer as
 * mapped in this process.  The payload may be rewritten as soon as the kernel
 * has finished on this rank's stream. */
int cvllm_allreduce_oneshot(void* const* peer_bases_host, int rank, int world, void* out,
                            int64_t numel, int dtype, cvllm_stream_t stream);
/* The same sum in two steps for messages of a few MB at 4-8 ranks: rank r reduces slice r into a RESULT area of its
 * own buffer (result_offset_bytes past the payload start, >= numel * 2, the same on every rank; allocate
 * cvllm_ipc_alloc(payload + result)), publishes it, and every rank gathers the slices.  NVLink traffic per rank
 * 2 x (world-1)/world x message instead of (world-1) x message; identical bits to the one-shot form. */
int cvllm_allreduce_twoshot(void* const* peer_bases_host, int rank, int world, void* out,
                            int64_t numel, int dtype, int64_t result_offset_bytes, cvllm_stream_t stream);

/* RowParallelLinear.forward as ONE kernel — layers/linear.py:133-153 (`F.linear` then `dist.all_reduce`), the o_proj of
 * models/llama3.py:111.  y[M, N] = sum over ranks r of round16(x_r[M, K] . w_r[N, K]^T), fp32 sum in rank order,
 * rounded once, identical bits on every rank.  The GEMM runs on tcgen05 CTA pairs; its epilogue TMA-stores each
 * finished tile into the owning rank's buffer over NVLink (reduce-scatter), reducer warps of the same kernel sum a
 * round of rows as soon as every rank's tiles for it have landed and store the result into y on every rank
 * (all-gather).  Buffers: one cvllm_ipc_alloc(cvllm_gemm_allreduce_payload_bytes(M_max, N, world)) per rank, mapped
 * by every rank (cvllm_ipc_open); a buffer is private to this entry (its flag words are not the all-reduce's).
 * Every rank calls with the same M, N, dtype, in the same order.  *y_out_host = y inside peer_bases_host[rank]; it
 * stays valid until the next call on this buffer.  Needs N, K, lda, ldw multiples of 8 (else CVLLM_E_UNSUPPORTED). */
int64_t cvllm_gemm_allreduce_payload_bytes(int64_t M, int64_t N, int world);
int cvllm_gemm_allreduce(const void* x, int64_t lda, const void* w, int64_t ldw, int64_t M, int64_t N, int64_t K,
                         void* const* peer_bases_host, int rank, int world, int dtype,
                         int64_t payload_capacity_bytes, void** y_out_host, cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * Selection + compaction
 * --------------------------------------------------------------------- */

/* scores_to_retain_indices — compression/common.py:171-243.
 * Per sequence b: full descending ranking of its len_b*H scores, ties broken
 * by ascending flat index, NaN ranked above +inf (torch.topk order).  Output
 * int64 [B, k_eff], k_eff = min(top_k, max_k_len*H); entry = global flat
 * index (cu[b]+tok)*H + head.  Positions >= len_b*H hold the padded-slot
 * indices the reference would produce (cu[b]*H + padded position). */
size_t cvllm_rank_scores_workspace_bytes(int B, int H, int max_k_len);
int cvllm_rank_scores(const float* scores, int64_t s_stride_n,
                      const int32_t* cu_seqlens_k, int B, int H, int max_k_len,
                      int top_k, int64_t* out_indices, void* workspace,
                      size_t workspace_bytes, cvllm_stream_t stream);
/* The same ranking with NO caller workspace and no allocation of any kind (no cudaMalloc / cudaFree / allocating
 * library call).  No counterpart in the reference: added for callers that cannot
 * allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The sort's scratch is a
 * file-scope device array of cvllm_rank_scores_static_max_bytes() = 8 MiB that every GPU loading the library holds
 * until exit; problems whose cvllm_rank_scores_workspace_bytes(B, H, max_k_len) exceeds it (more than ~510 K scores)
 * return CVLLM_E_UNSUPPORTED with the sizes in cvllm_last_error().  All calls on a device SHARE the array: calls that
 * overlap on the GPU (different streams) race; nothing needs clearing between calls.  Output identical to
 * cvllm_rank_scores. */
size_t cvllm_rank_scores_static_max_bytes(void);
int cvllm_rank_scores_static(const float* scores, int64_t s_stride_n,
                             const int32_t* cu_seqlens_k, int B, int H, int max_k_len,
                             int top_k, int64_t* out_indices, cvllm_stream_t stream);

/* prefill_store_topk_kv — kv_cache/store_kv_cache.py:9-248 (K4 + K5), index
 * flavour: `indices` is the [B, max_sel] ranking (int32 or int64 by
 * idx_bytes).  The first num_retain[b] entries are stored, then (if
 * pad_to_page) each head is topped up to a page multiple with the next
 * entries of that head (:205-248).  Rows are written in ascending token order
 * per head (the reference's order is atomic-nondeterministic). */
size_t cvllm_store_topk_workspace_bytes(int B, int H, int N, int max_sel);
int cvllm_prefill_store_topk_idx(const void* keys, const void* vals,
                                 int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
                                 const void* indices, int idx_bytes, int max_sel,
                                 const int32_t* num_retain, const int32_t* cu_seqlens_k,
                                 const int32_t* batch_mapping, int32_t* bh_lens,
                                 const int32_t* page_table, void* k_cache, void* v_cache,
                                 int B, int N, int HKV, int D, int PS, int NLP,
                                 int elem_bytes, int pad_to_page, int reserved_batch,
                                 void* workspace, size_t workspace_bytes,
                                 cvllm_stream_t stream);

/* extract_and_store_top_kv — compression/common.py:126-168, fused: scores ->
 * radix-select (global top num_retain[b], then per-head top n_h) -> gather
 * into the paged cache.  Same kept sets as rank + store_topk_idx with the full
 * ranking (top_k >= max_len*H, the engine's setting, utils/arguments.py
 * :208-211). */
size_t cvllm_select_store_workspace_bytes(int B, int H, int N);
int cvllm_select_store_topk(const float* scores, int64_t s_stride_n,
                            const void* keys, const void* vals,
                            int64_t sk_n, int64_t sk_h, int64_t sv_n, int64_t sv_h,
                            const int32_t* num_retain, const int32_t* cu_seqlens_k,
                            const int32_t* batch_mapping, int32_t* bh_lens,
                            const int32_t* page_table, void* k_cache, void* v_cache,
                            int B, int N, int HKV, int D, int PS, int NLP,
                            int elem_bytes, int pad_to_page, int reserved_batch,
                            void* workspace, size_t workspace_bytes,
                            cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * Attention
 * --------------------------------------------------------------------- */

/* head_sparse_decode_attention — attention/sparse_decode_kernel.py:10-165
 * (K2 :247-388 + K3 :392-435).  One query token per sequence against the
 * per-(b,h) paged rows; balanced split over 128-token units with an in-kernel
 * LSE merge.  Rows with L==0 are written as zeros.  `num_ctas` <= 0 picks
 * one persistent CTA per SM.  PS must be 16, 32, 64 or a multiple of 128 (a
 * TMA box of min(PS,128) rows never crosses a page).  Workspace contract: the
 * first 256-byte-rounded B*HKV*4 bytes are per-head tickets that must be ZERO
 * the first time a workspace is used; every launch leaves them zero again, so
 * a workspace zeroed once at allocation can be reused forever (and a captured
 * graph is the single kernel node).  Graph-capturable: no host sync, fixed
 * grid. */
size_t cvllm_decode_attention_workspace_bytes(int B, int HQ, int HKV, int D, int num_ctas);
int cvllm_decode_attention(const void* q, void* out, const void* k_cache, const void* v_cache,
                           const int32_t* seq_lens_bh, const int32_t* page_table,
                           const int32_t* batch_mapping, void* workspace,
                           size_t workspace_bytes, int B, int HQ, int HKV, int D,
                           int PS, int NLP, float sm_scale, int dtype, int num_ctas,
                           cvllm_stream_t stream);

/* The decode branch of Attention.forward (layers/attention.py:133-160) as one call: decode_store_kv followed by
 * head_sparse_decode_attention, both indexing the cache's own `bh_seq_lens` table [Bmax+1, HKV] through
 * `batch_mapping` (rows mapped to `reserved_batch` are skipped by the store and read length 0).  Replaces the
 * reference's index_select / clone / index_copy_ of a gathered length copy: two kernel nodes per layer-step.
 * 16-bit caches only; workspace contract as for cvllm_decode_attention. */
int cvllm_decode_step(const void* key, const void* value, int64_t sk_b, int64_t sk_h,
                      int64_t sv_b, int64_t sv_h, const void* q, void* out, void* k_cache,
                      void* v_cache, int32_t* bh_seq_lens, const int32_t* page_table,
                      const int32_t* batch_mapping, void* workspace, size_t workspace_bytes,
                      int B, int HQ, int HKV, int D, int PS, int NLP, float sm_scale, int dtype,
                      int reserved_batch, int num_ctas, cvllm_stream_t stream);

/* causal_sparse_varlen_with_cache — attention/sparse_varlen_kernel.py:11-197
 * (K1 :265-519).  Causal attention of the packed new tokens over
 * [cached prefix of (b,h) || new K/V of b]; tcgen05/TMEM/TMA kernel. */
int cvllm_prefill_attention(const void* q, const void* k, const void* v, void* out,
                            const void* k_cache, const void* v_cache,
                            const int32_t* seq_lens_bh, const int32_t* page_table,
                            const int32_t* batch_mapping, const int32_t* cu_seqlens_q,
                            int B, int N, int HQ, int HKV, int D, int PS, int NLP,
                            int64_t cache_rows, int64_t q_stride_n, int64_t k_stride_n,
                            int64_t v_stride_n, int64_t o_stride_n, int max_seqlen_q,
                            float sm_scale, int dtype, cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * Scoring
 * --------------------------------------------------------------------- */

/* approximate_leverage_scores — compression/compactor.py:113-221 (+ z-score
 * :224-269).  chunk_offsets[n_chunks+1] are packed-token offsets of the
 * (<= chunk_size) chunks produced by split_into_chunks (:62-110).  out is
 * [N, HKV] in `dtype` (the reference returns K's dtype). */
size_t cvllm_leverage_scores_workspace_bytes(int N, int HKV, int sketch, int n_chunks);
int cvllm_leverage_scores(const void* k_pre, int64_t k_stride_n, int64_t k_stride_h,
                          const int32_t* chunk_offsets, int n_chunks, int max_chunk,
                          const void* phi, void* out_scores, int N, int HKV, int D,
                          int sketch, float regularizer, int normalize, int dtype,
                          void* workspace, size_t workspace_bytes, cvllm_stream_t stream);
/* Same chain from a precomputed sketch: sketch_rows [N, HKV, 48] in `dtype`
 * = round16(K_pre . PHI) — compression/compactor.py:153-158 — as written by
 * cvllm_rope_qk; bit-identical scores to cvllm_leverage_scores on the same
 * keys.  phi may be NULL. */
int cvllm_leverage_scores_from_sketch(const void* sketch_rows, const int32_t* chunk_offsets,
                                      int n_chunks, int max_chunk, const void* phi,
                                      void* out_scores, int N, int HKV, int sketch,
                                      float regularizer, int normalize, int dtype,
                                      void* workspace, size_t workspace_bytes,
                                      cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * The step in front of the path
 * --------------------------------------------------------------------- */

/* RotaryEmbedding.forward — layers/rotary_embedding.py:69-80 (neox
 * rotate-half, :8-16), optionally preceded by the per-head q_norm / k_norm
 * RMSNorm of models/qwen3.py:90-91 (layers/layernorm.py:16-26) and emitting
 * the Compactor sketch of the (normed) pre-RoPE keys
 * (compression/compactor.py:153-158) in the same pass.
 *   q [N,HQ,128], k [N,HKV,128] strided; q_out / k_out may alias q / k
 *   positions int64 [N]; cos_sin_cache fp32 [max_position, 128] = cos | sin
 *   q_norm_weight / k_norm_weight: [128] in `dtype` or NULL (no norm)
 *   phi [128,48] in `dtype` + sketch_out [N,HKV,48] in `dtype`, or both NULL
 *   (sketch needs HKV <= 8) */
int cvllm_rope_qk(const void* q, const void* k, void* q_out, void* k_out, int64_t q_stride_n,
                  int64_t q_stride_h, int64_t k_stride_n, int64_t k_stride_h,
                  int64_t qo_stride_n, int64_t qo_stride_h, int64_t ko_stride_n,
                  int64_t ko_stride_h, const int64_t* positions, const float* cos_sin_cache,
                  int max_position, const void* q_norm_weight, const void* k_norm_weight,
                  float eps, const void* phi, void* sketch_out, int N, int HQ, int HKV, int D,
                  int sketch, int dtype, cvllm_stream_t stream);

/* RMSNorm.forward — layers/layernorm.py:16-49.  residual == NULL:
 * out = round16(x * rsqrt(mean(x^2) + eps)) * weight.  residual != NULL
 * (add_rms_forward): x += residual in fp32, residual <- round16(x) in place,
 * then the same norm.  width % 8 == 0, width <= 8192. */
int cvllm_rmsnorm(const void* x, void* residual, const void* weight, void* out, int64_t x_stride,
                  int64_t residual_stride, int64_t out_stride, int rows, int width, float eps,
                  int dtype, cvllm_stream_t stream);

/* non_causal_attn_scores — compression/compactor.py:338-599.  Chunk-local
 * softmax column sums, per-sequence z-score, + blend*accum, protected
 * first/last tokens <- +inf.  out fp32 [N, HKV]. accum may be NULL;
 * accum_dtype: 2 = fp32, else CVLLM_F16/BF16.  prot_first/prot_last may be
 * NULL (device int32 [B]). */
size_t cvllm_chunk_attn_scores_workspace_bytes(int B, int N, int HKV, int chunk);
int cvllm_chunk_attn_scores(const void* q, const void* k, int64_t q_stride_n,
                            int64_t k_stride_n, const int32_t* cu_seqlens,
                            int B, int N, int HQ, int HKV, int D, int chunk,
                            int max_seqlen, float sm_scale, int normalize,
                            const void* accum, int accum_dtype, int64_t accum_stride_n,
                            float blend, const int32_t* prot_first,
                            const int32_t* prot_last, float* out, int dtype,
                            void* workspace, size_t workspace_bytes,
                            cvllm_stream_t stream);

/* query_aware_key_scores — compression/snapkv.py:39-276,332-448.  SnapKV:
 * last-w queries x G heads against keys [0, L-w), softmax over those keys,
 * column sums, causal `kpool`-tap mean pool inside `pool_segment`-key
 * segments (the reference's autotuned BLOCK_K, :160-168,253-262), window
 * keys <- +inf.  No logits scratch. */
size_t cvllm_snapkv_scores_workspace_bytes(int B, int N, int HQ, int HKV, int w);
int cvllm_snapkv_scores(const void* q, const void* k, int64_t q_stride_n,
                        int64_t k_stride_n, const int32_t* cu_seqlens_q,
                        const int32_t* cu_seqlens_k, int w, int B, int Nk, int HQ,
                        int HKV, int D, int max_seqlen_k, float sm_scale, int kpool,
                        int pool_segment, float* out, int dtype, void* workspace,
                        size_t workspace_bytes, cvllm_stream_t stream);

/* Test entry: the fp64 symmetric eigensolver of the leverage chain alone (csrc/score_leverage.cu).
 * gram fp32 [mats,48,48]; lam_out fp64 [mats,48]; vec_out fp64 [mats,48,48] (columns = eigenvectors, any
 * order); sv_out fp32 [mats,48,48] = round16(V |lam|^-1/2), what the score kernel consumes
 * (reference compression/compactor.py:177,197).  clk_out (optional, int64 [mats,6]): SM clock at kernel entry,
 * after the tridiagonal reduction, at the end of the QL producer, at the end of the reflector accumulation, at
 * exit, and the number of ring entries produced (profiling). */
int cvllm_selftest_eig(const float* gram, double* lam_out, double* vec_out, float* sv_out, int mats,
                       int dtype, long long* clk_out, cvllm_stream_t stream);

/* ------------------------------------------------------------------------
 * Self tests (device sanity for descriptor encodings; used by tests only)
 * --------------------------------------------------------------------- */
int cvllm_selftest_umma(const void* a, const void* b, float* d, int mode,
                        uint32_t lbo, uint32_t sbo, int dtype, cvllm_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CVLLM_H_ */
