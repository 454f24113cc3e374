// The step in front of the hot path, fused: per-head q/k RMSNorm (Qwen3) -> Compactor sketch
// X = K_pre . PHI -> rotary embedding of q and k, one pass over q and k.
//
// Reference (three torch.compile'd ops + one cuBLAS GEMM, each a full read and write of q / k):
//   layers/layernorm.py:16-26          RMSNorm.rms_forward   (models/qwen3.py:90-91: q_norm, k_norm)
//   compression/compactor.py:153-158   X = key_states @ PHI  (pre-RoPE keys, K's dtype)
//   layers/rotary_embedding.py:8-16,69-80   neox rotate-half with the fp32 cos|sin cache
//
// HBM-bound: 2 x (HQ+HKV) x 256 B per token read + written (20 KB at HQ=32, HKV=8) + 96 B per
// (token, kv head) of sketch.  The normed pre-RoPE keys never reach HBM: the leverage chain reads
// the 48-wide sketch instead of re-deriving it from K three times (score_leverage.cu).
//
// Layout: a CTA owns 16-token tiles (one mma m-tile of the sketch).  Eight lanes share a head row:
// lane l holds d = 8l..8l+7 and d+64 — both halves of every rotation pair, two 16-byte loads.
// Rounding points are the reference's eager ones (no FMA contraction), so the rotation is
// bit-exact against the oracle; the norm differs only by the order of the 128-term sum of squares.
#include "common.cuh"
#include "mma.cuh"

namespace cvllm {

constexpr int kRopeThreads = 256;
constexpr int kRopeTok = 16;
constexpr int kRopeSk = 48;
constexpr int kRopeBatch = 4;  // head rows in flight per 8-lane group
constexpr int kRopeMaxHkv = 8;  // kv heads whose pre-RoPE rows are staged for the sketch

struct RopeParams {
  const void* q;
  const void* k;
  void* q_out;
  void* k_out;
  int64_t q_sn, q_sh, k_sn, k_sh, qo_sn, qo_sh, ko_sn, ko_sh;  // elements
  const int64_t* positions;  // [N]
  const float* cos_sin;      // [max_pos, 128]: cos | sin
  const void* q_w;           // [128] 16-bit or null
  const void* k_w;
  float eps;
  float one;  // 1.0f (see add2_rn)
  const void* phi;  // [128, 48] 16-bit or null
  void* sketch;     // [N, HKV, 48] 16-bit or null
  int N, HQ, HKV, max_pos;
};

struct RopeSmem {
  uint16_t phiT[kRopeSk][128 + 8];                     // PHI^T, 68-word rows: conflict-free LDS.32 fragments
  // normed pre-RoPE key rows of the tile, double-buffered: one __syncthreads per tile (the barrier of tile i+1
  // orders the sketch reads of tile i before the writes of tile i+2 into the same buffer)
  uint16_t kpre[2][kRopeMaxHkv][kRopeTok][128 + 8];
};

template <bool BF16>
__device__ __forceinline__ float r16(float x) {
  if constexpr (BF16) return __bfloat162float(__float2bfloat16_rn(x));
  else return __half2float(__float2half_rn(x));
}

template <bool BF16>
__device__ __forceinline__ void unpack8(const uint4& v, float* x) {
  float2 a = unpack2<BF16>(v.x), b = unpack2<BF16>(v.y), c = unpack2<BF16>(v.z), d = unpack2<BF16>(v.w);
  x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y; x[4] = c.x; x[5] = c.y; x[6] = d.x; x[7] = d.y;
}
template <bool BF16>
__device__ __forceinline__ uint4 pack8(const float* x) {
  return make_uint4(pack2<BF16>(x[0], x[1]), pack2<BF16>(x[2], x[3]), pack2<BF16>(x[4], x[5]),
                    pack2<BF16>(x[6], x[7]));
}

// packed fp32 pairs (FMUL2 / FADD2 / FFMA2 of sm_100): half the issue slots of the scalar forms, each
// lane still rounded to nearest, and the _rn intrinsics are never contracted into an FMA
// explicit .rn PTX: ptxas never fuses a mul.rn with an add.rn (the sm_100_rt intrinsics were seen contracted)
__device__ __forceinline__ float2 mul2_rn(float2 a, float2 b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
// a + b as fma(a, ones, b) with `ones` = (1, 1) from a kernel parameter: ptxas contracts mul.rn.f32x2 +
// add.rn.f32x2 into one FFMA2 (seen in SASS even with -fmad=false), which would drop the reference's rounding
// of the products; it cannot fold an fma whose multiplier it does not know.  a*1+b rounds once: exact.
__device__ __forceinline__ float2 add2_rn(float2 a, float2 b, float2 ones) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<uint64_t*>(&a)),
      "l"(*reinterpret_cast<uint64_t*>(&ones)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
template <bool BF16>
__device__ __forceinline__ void unpack8v(const uint4& v, float2* x) {
  x[0] = unpack2<BF16>(v.x); x[1] = unpack2<BF16>(v.y); x[2] = unpack2<BF16>(v.z); x[3] = unpack2<BF16>(v.w);
}
template <bool BF16>
__device__ __forceinline__ uint4 pack8v(const float2* x) {
  return make_uint4(pack2<BF16>(x[0].x, x[0].y), pack2<BF16>(x[1].x, x[1].y), pack2<BF16>(x[2].x, x[2].y),
                    pack2<BF16>(x[3].x, x[3].y));
}
template <bool BF16>
__device__ __forceinline__ float2 r16v(float2 x) { return unpack2<BF16>(pack2<BF16>(x.x, x.y)); }

template <bool BF16, bool NORM, bool SKETCH>
__global__ void __launch_bounds__(kRopeThreads, 2) rope_qk_kernel(const RopeParams p) {
  extern __shared__ __align__(16) unsigned char rope_smem_raw[];
  RopeSmem& sm = *reinterpret_cast<RopeSmem*>(rope_smem_raw);
  static_assert(kRopeThreads / 8 == 2 * kRopeTok, "two 8-lane groups per token of the tile");
  const int tid = threadIdx.x, grp = tid >> 3, l = tid & 7;
  const int rows_per_tok = p.HQ + p.HKV;
  const int n_tiles = (p.N + kRopeTok - 1) / kRopeTok;

  if constexpr (SKETCH) {
    const uint16_t* phi = reinterpret_cast<const uint16_t*>(p.phi);
    for (int i = tid; i < 128 * kRopeSk; i += kRopeThreads) {
      const int d = i / kRopeSk, c = i - d * kRopeSk;
      sm.phiT[c][d] = phi[i];
    }
  }
  __shared__ __align__(16) float s_w[2][128];  // q / k norm weights as fp32
  if constexpr (NORM) {
    if (tid < 128) {
      const uint16_t* qw = reinterpret_cast<const uint16_t*>(p.q_w);
      const uint16_t* kw = reinterpret_cast<const uint16_t*>(p.k_w);
      s_w[0][tid] = qw ? unpack2<BF16>((uint32_t)qw[tid]).x : 1.f;
      s_w[1][tid] = kw ? unpack2<BF16>((uint32_t)kw[tid]).x : 1.f;
    }
    __syncthreads();
  }

  // group (8 lanes) = one token of the tile and every second head row of it: cos / sin are loaded once per
  // tile, kRopeBatch rows are in flight per group before any of them is touched
  const int t = grp & (kRopeTok - 1), half = grp >> 4;
  const float2 ones = make_float2(p.one, p.one);
  int buf = 0;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, buf ^= 1) {
    const int tok0 = tile * kRopeTok;
    const int nt = min(kRopeTok, p.N - tok0);
    const bool tvalid = t < nt;
    const int64_t tok = tok0 + t;
    int64_t pos = tvalid ? p.positions[tok] : 0;
    pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);  // memory safety only
    const float* cs = p.cos_sin + pos * 128;
    const float4 c0 = *reinterpret_cast<const float4*>(cs + 8 * l);
    const float4 c1 = *reinterpret_cast<const float4*>(cs + 8 * l + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(cs + 64 + 8 * l);
    const float4 s1 = *reinterpret_cast<const float4*>(cs + 64 + 8 * l + 4);
    const float2 cc[4] = {{c0.x, c0.y}, {c0.z, c0.w}, {c1.x, c1.y}, {c1.z, c1.w}};
    const float2 sn[4] = {{s0.x, s0.y}, {s0.z, s0.w}, {s1.x, s1.y}, {s1.z, s1.w}};
    const float2 ns[4] = {{-s0.x, -s0.y}, {-s0.z, -s0.w}, {-s1.x, -s1.y}, {-s1.z, -s1.w}};
    const uint16_t* q_tok = reinterpret_cast<const uint16_t*>(p.q) + tok * p.q_sn;
    const uint16_t* k_tok = reinterpret_cast<const uint16_t*>(p.k) + tok * p.k_sn;
    uint16_t* qo_tok = reinterpret_cast<uint16_t*>(p.q_out) + tok * p.qo_sn;
    uint16_t* ko_tok = reinterpret_cast<uint16_t*>(p.k_out) + tok * p.ko_sn;
    for (int hh0 = half; hh0 < rows_per_tok; hh0 += 2 * kRopeBatch) {
      uint4 lo[kRopeBatch], hi[kRopeBatch];
#pragma unroll
      for (int r = 0; r < kRopeBatch; ++r) {
        const int hh = hh0 + 2 * r;
        lo[r] = make_uint4(0, 0, 0, 0);
        hi[r] = lo[r];
        if (tvalid && hh < rows_per_tok) {
          const uint16_t* src = hh >= p.HQ ? k_tok + (int64_t)(hh - p.HQ) * p.k_sh : q_tok + (int64_t)hh * p.q_sh;
          lo[r] = *reinterpret_cast<const uint4*>(src + 8 * l);  // plain loads: q_out / k_out may alias q / k
          hi[r] = *reinterpret_cast<const uint4*>(src + 64 + 8 * l);
        }
      }
#pragma unroll
      for (int r = 0; r < kRopeBatch; ++r) {
        const int hh = hh0 + 2 * r;
        const bool valid = tvalid && hh < rows_per_tok;
        const bool is_k = hh >= p.HQ;
        const int h = is_k ? hh - p.HQ : hh;
        float2 x[8];  // x[0..3] = d 8l..8l+7, x[4..7] = d 64+8l..
        unpack8v<BF16>(lo[r], x);
        unpack8v<BF16>(hi[r], x + 4);
        if constexpr (NORM) {
          const bool has_w = is_k ? (p.k_w != nullptr) : (p.q_w != nullptr);
          float2 ss2 = make_float2(0.f, 0.f);
#pragma unroll
          for (int i = 0; i < 8; ++i) ss2 = __ffma2_rn(x[i], x[i], ss2);
          float ss = ss2.x + ss2.y;
          ss += __shfl_xor_sync(0xffffffffu, ss, 1);
          ss += __shfl_xor_sync(0xffffffffu, ss, 2);
          ss += __shfl_xor_sync(0xffffffffu, ss, 4);
          if (has_w) {
            // x.float() * rsqrt(mean(x^2) + eps) -> 16 bit -> * weight -> 16 bit (layernorm.py:21-25);
            // rsqrt = MUFU + one Newton step (<= 1 ulp of the reference's 1/sqrt)
            const float var = __fadd_rn(ss * (1.0f / 128.0f), p.eps);
            float rs = rsqrtf(var);
            rs = rs * fmaf(-0.5f * var * rs, rs, 1.5f);
            const float2 rs2 = make_float2(rs, rs);
            const float* wrow = s_w[is_k ? 1 : 0];
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              const float4 w0 = *reinterpret_cast<const float4*>(wrow + 64 * hf + 8 * l);
              const float4 w1 = *reinterpret_cast<const float4*>(wrow + 64 * hf + 8 * l + 4);
              const float2 ww[4] = {{w0.x, w0.y}, {w0.z, w0.w}, {w1.x, w1.y}, {w1.z, w1.w}};
#pragma unroll
              for (int i = 0; i < 4; ++i)
                x[4 * hf + i] = r16v<BF16>(mul2_rn(r16v<BF16>(mul2_rn(x[4 * hf + i], rs2)), ww[i]));
            }
          }
        }
        if constexpr (SKETCH) {
          if (valid && is_k) {
            *reinterpret_cast<uint4*>(&sm.kpre[buf][h][t][8 * l]) = pack8v<BF16>(x);
            *reinterpret_cast<uint4*>(&sm.kpre[buf][h][t][64 + 8 * l]) = pack8v<BF16>(x + 4);
          }
        }
        // y1 = x1*cos - x2*sin ; y2 = x2*cos + x1*sin, fp32, each product rounded (rotary_embedding.py:13-15)
        float2 y[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          y[i] = add2_rn(mul2_rn(x[i], cc[i]), mul2_rn(x[4 + i], ns[i]), ones);
          y[4 + i] = add2_rn(mul2_rn(x[4 + i], cc[i]), mul2_rn(x[i], sn[i]), ones);
        }
        if (valid) {
          uint16_t* dst = is_k ? ko_tok + (int64_t)h * p.ko_sh : qo_tok + (int64_t)h * p.qo_sh;
          stg_na_v4(dst + 8 * l, pack8v<BF16>(y));
          stg_na_v4(dst + 64 + 8 * l, pack8v<BF16>(y + 4));
        }
      }
    }
    if constexpr (SKETCH) {
      __syncthreads();
      // warp w = kv head w: X[16 tokens, 48] = Kpre_h . PHI, the fragment order of
      // score_leverage.cu's sketch_tile so both routes give the same bits
      const int warp = tid >> 5, lane = tid & 31, r = lane >> 2, j = lane & 3;
      for (int h = warp; h < p.HKV; h += kRopeThreads / 32) {
        const int t_lo = r, t_hi = r + 8;
        FragNat128 alo, ahi;  // natural-k chain (mma.cuh): cuBLAS's bits for K.PHI
        alo.load(t_lo < nt ? sm.kpre[buf][h][t_lo] : nullptr, j);
        ahi.load(t_hi < nt ? sm.kpre[buf][h][t_hi] : nullptr, j);
        uint16_t* xo = reinterpret_cast<uint16_t*>(p.sketch);
#pragma unroll
        for (int nn = 0; nn < kRopeSk / 8; ++nn) {
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
          const uint32_t* prow = reinterpret_cast<const uint32_t*>(sm.phiT[nn * 8 + r]);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            mma16816<BF16>(acc, alo.w[2 * ks], ahi.w[2 * ks], alo.w[2 * ks + 1], ahi.w[2 * ks + 1],
                           prow[8 * ks + j], prow[8 * ks + 4 + j]);
          const int cx = nn * 8 + 2 * j;
          if (t_lo < nt)
            *reinterpret_cast<uint32_t*>(xo + ((int64_t)(tok0 + t_lo) * p.HKV + h) * kRopeSk + cx) =
                pack2<BF16>(acc[0], acc[1]);
          if (t_hi < nt)
            *reinterpret_cast<uint32_t*>(xo + ((int64_t)(tok0 + t_hi) * p.HKV + h) * kRopeSk + cx) =
                pack2<BF16>(acc[2], acc[3]);
        }
      }
    }
  }
}


// ---- generic RMSNorm over rows of `width` (layers/layernorm.py:16-49), one CTA per row ----
// residual != null: x += residual (fp32), residual <- round16(x) (add_rms_forward :28-40)
constexpr int kNormThreads = 256;
constexpr int kNormMaxChunks = 4;  // 16-byte chunks per thread: width <= 8192

template <bool BF16>
__global__ void __launch_bounds__(kNormThreads) rmsnorm_kernel(const void* __restrict__ x_, void* res_,
                                                               const void* __restrict__ w_, void* out_,
                                                               int64_t x_stride, int64_t res_stride,
                                                               int64_t out_stride, int width, float eps) {
  const int row = blockIdx.x, tid = threadIdx.x;
  const uint16_t* x = reinterpret_cast<const uint16_t*>(x_) + (int64_t)row * x_stride;
  uint16_t* res = res_ ? reinterpret_cast<uint16_t*>(res_) + (int64_t)row * res_stride : nullptr;
  const uint16_t* w = reinterpret_cast<const uint16_t*>(w_);
  uint16_t* out = reinterpret_cast<uint16_t*>(out_) + (int64_t)row * out_stride;
  const int n_chunks = width >> 3;
  float v[kNormMaxChunks][8];
  float ss = 0.f;
#pragma unroll
  for (int c = 0; c < kNormMaxChunks; ++c) {
    const int ch = tid + c * kNormThreads;
    if (ch < n_chunks) {
      unpack8<BF16>(*reinterpret_cast<const uint4*>(x + 8 * ch), v[c]);
      if (res) {
        float rr[8];
        unpack8<BF16>(*reinterpret_cast<const uint4*>(res + 8 * ch), rr);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[c][i] = __fadd_rn(v[c][i], rr[i]);
        *reinterpret_cast<uint4*>(res + 8 * ch) = pack8<BF16>(v[c]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ss = fmaf(v[c][i], v[c][i], ss);
    }
  }
  __shared__ float s_part[kNormThreads / 32];
  __shared__ float s_rs;
  ss = warp_sum(ss);
  if ((tid & 31) == 0) s_part[tid >> 5] = ss;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int i = 0; i < kNormThreads / 32; ++i) t += s_part[i];
    s_rs = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(t, (float)width), eps)));
  }
  __syncthreads();
  const float rs = s_rs;
#pragma unroll
  for (int c = 0; c < kNormMaxChunks; ++c) {
    const int ch = tid + c * kNormThreads;
    if (ch < n_chunks) {
      float ww[8], y[8];
      unpack8<BF16>(*reinterpret_cast<const uint4*>(w + 8 * ch), ww);
#pragma unroll
      for (int i = 0; i < 8; ++i) y[i] = __fmul_rn(r16<BF16>(__fmul_rn(v[c][i], rs)), ww[i]);
      *reinterpret_cast<uint4*>(out + 8 * ch) = pack8<BF16>(y);
    }
  }
}

}  // namespace cvllm

using namespace cvllm;

template <bool BF16>
static int launch_rope(const RopeParams& p, bool norm, bool sketch, int n_tiles, cudaStream_t st) {
  const size_t smem = sketch ? sizeof(RopeSmem) : 0;
#define CVLLM_ROPE_LAUNCH(NORM, SK)                                                                \
  do {                                                                                             \
    static std::atomic<uint64_t> attr{0};                                                          \
    if (SK)                                                                                        \
      CVLLM_CHECK_CUDA(ensure_max_dyn_smem((const void*)rope_qk_kernel<BF16, NORM, SK>, (int)smem, attr)); \
    int per_sm = 1;                                                                                \
    CVLLM_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(                                \
        &per_sm, rope_qk_kernel<BF16, NORM, SK>, kRopeThreads, smem));                             \
    const int g = n_tiles < sm_count() * per_sm ? n_tiles : sm_count() * per_sm;                   \
    rope_qk_kernel<BF16, NORM, SK><<<g, kRopeThreads, smem, st>>>(p);                              \
  } while (0)
  if (norm && sketch) CVLLM_ROPE_LAUNCH(true, true);
  else if (norm) CVLLM_ROPE_LAUNCH(true, false);
  else if (sketch) CVLLM_ROPE_LAUNCH(false, true);
  else CVLLM_ROPE_LAUNCH(false, false);
#undef CVLLM_ROPE_LAUNCH
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}

extern "C" int cvllm_rope_qk(const void* q, const void* k, void* q_out, void* k_out, int64_t q_stride_n,
                             int64_t q_stride_h, int64_t k_stride_n, int64_t k_stride_h,
                             int64_t qo_stride_n, int64_t qo_stride_h, int64_t ko_stride_n,
                             int64_t ko_stride_h, const int64_t* positions, const float* cos_sin_cache,
                             int max_position, const void* q_norm_weight, const void* k_norm_weight,
                             float eps, const void* phi, void* sketch_out, int N, int HQ, int HKV, int D,
                             int sketch, int dtype, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(N >= 0 && HQ >= 0 && HKV >= 0 && max_position > 0);
  if (N == 0 || HQ + HKV == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(positions && cos_sin_cache);
  CVLLM_CHECK_ARG((HQ == 0 || (q && q_out)) && (HKV == 0 || (k && k_out)));
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_SUPPORTED(D == 128);
  CVLLM_CHECK_ARG((phi == nullptr) == (sketch_out == nullptr));
  const bool do_sketch = sketch_out != nullptr;
  if (do_sketch) CVLLM_CHECK_SUPPORTED(sketch == kRopeSk && HKV >= 1 && HKV <= kRopeMaxHkv);
  CVLLM_CHECK_ALIGN(aligned16(q) && aligned16(k) && aligned16(q_out) && aligned16(k_out) &&
                    aligned16(cos_sin_cache) && aligned16(q_norm_weight) && aligned16(k_norm_weight) &&
                    (reinterpret_cast<uintptr_t>(sketch_out) & 3) == 0);
  CVLLM_CHECK_ALIGN(q_stride_n % 8 == 0 && q_stride_h % 8 == 0 && k_stride_n % 8 == 0 &&
                    k_stride_h % 8 == 0 && qo_stride_n % 8 == 0 && qo_stride_h % 8 == 0 &&
                    ko_stride_n % 8 == 0 && ko_stride_h % 8 == 0);
  RopeParams p;
  p.q = q; p.k = k; p.q_out = q_out; p.k_out = k_out;
  p.q_sn = q_stride_n; p.q_sh = q_stride_h; p.k_sn = k_stride_n; p.k_sh = k_stride_h;
  p.qo_sn = qo_stride_n; p.qo_sh = qo_stride_h; p.ko_sn = ko_stride_n; p.ko_sh = ko_stride_h;
  p.positions = positions; p.cos_sin = cos_sin_cache; p.q_w = q_norm_weight; p.k_w = k_norm_weight;
  p.eps = eps; p.one = 1.0f; p.phi = phi; p.sketch = sketch_out; p.N = N; p.HQ = HQ; p.HKV = HKV; p.max_pos = max_position;
  const bool norm = q_norm_weight != nullptr || k_norm_weight != nullptr;
  const int n_tiles = (N + kRopeTok - 1) / kRopeTok;
  auto st = static_cast<cudaStream_t>(stream);
  return dtype == CVLLM_BF16 ? launch_rope<true>(p, norm, do_sketch, n_tiles, st)
                             : launch_rope<false>(p, norm, do_sketch, n_tiles, st);
}

extern "C" int cvllm_rmsnorm(const void* x, void* residual, const void* weight, void* out, int64_t x_stride,
                             int64_t residual_stride, int64_t out_stride, int rows, int width, float eps,
                             int dtype, cvllm_stream_t stream) {
  CVLLM_CHECK_ARG(rows >= 0 && width > 0);
  if (rows == 0) return CVLLM_OK;
  CVLLM_CHECK_ARG(x && weight && out);
  CVLLM_CHECK_ARG(dtype == CVLLM_F16 || dtype == CVLLM_BF16);
  CVLLM_CHECK_SUPPORTED(width % 8 == 0 && width <= 8 * kNormThreads * kNormMaxChunks);
  CVLLM_CHECK_ALIGN(aligned16(x) && aligned16(residual) && aligned16(weight) && aligned16(out) &&
                    x_stride % 8 == 0 && residual_stride % 8 == 0 && out_stride % 8 == 0);
  auto st = static_cast<cudaStream_t>(stream);
  if (dtype == CVLLM_BF16)
    rmsnorm_kernel<true><<<rows, kNormThreads, 0, st>>>(x, residual, weight, out, x_stride, residual_stride,
                                                        out_stride, width, eps);
  else
    rmsnorm_kernel<false><<<rows, kNormThreads, 0, st>>>(x, residual, weight, out, x_stride, residual_stride,
                                                         out_stride, width, eps);
  CVLLM_CHECK_LAUNCH();
  return CVLLM_OK;
}
