// mma.sync m16n8k16 helpers shared by the memory-bound / small-GEMM kernels
// (decode attention and the scorers).  The big contraction (prefill) is
// tcgen05 — see prefill_attn.cu.
//
// "Permuted-k" fragments: an operand whose contraction dim (k) is contiguous
// in memory is read as 16-byte chunks; lane j (= lane & 3) owns chunks
// j, j+4, j+8, ... and maps chunk cc's four 32-bit words to k-steps 2cc
// (words x,y) and 2cc+1 (words z,w).  As long as BOTH operands use the same
// mapping the contraction is unchanged, every load is a full 16-byte vector
// and no ldmatrix / shared-memory transpose is needed.
#pragma once
#include "common.cuh"

namespace cvllm {

template <bool BF16>
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  if constexpr (BF16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
}

__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

__device__ __forceinline__ uint32_t u4_get(const uint4& v, int i) {
  return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w));
}

// A-operand (16 rows x 128 k) of rows `row_lo` / `row_hi` (this lane's rows r and r+8), k
// contiguous with 256-byte rows: a[cc] = chunk (j + 4cc) of each row.  Null pointer = zero row.
struct AFrag128 {
  uint4 lo[4], hi[4];
  __device__ __forceinline__ void load(const char* row_lo, const char* row_hi, int j) {
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      lo[cc] = row_lo ? *reinterpret_cast<const uint4*>(row_lo + (j + 4 * cc) * 16) : z;
      hi[cc] = row_hi ? *reinterpret_cast<const uint4*>(row_hi + (j + 4 * cc) * 16) : z;
    }
  }
  // registers of k-step ks (0..7)
  __device__ __forceinline__ void regs(int ks, uint32_t& a0, uint32_t& a1, uint32_t& a2, uint32_t& a3) const {
    const int cc = ks >> 1;
    a0 = (ks & 1) ? lo[cc].z : lo[cc].x;
    a1 = (ks & 1) ? hi[cc].z : hi[cc].x;
    a2 = (ks & 1) ? lo[cc].w : lo[cc].y;
    a3 = (ks & 1) ? hi[cc].w : hi[cc].y;
  }
};

// B-operand registers of k-step ks from the chunk (j + 4*(ks>>1)) of the n-row this lane owns
__device__ __forceinline__ void bfrag_regs(const uint4& chunk, int ks, uint32_t& b0, uint32_t& b1) {
  b0 = (ks & 1) ? chunk.z : chunk.x;
  b1 = (ks & 1) ? chunk.w : chunk.y;
}

// "Natural-k" operands: k-step ks contracts k = 16ks .. 16ks+15 with the instruction's own slot order (a0/a1: k =
// 2j, 2j+1; a2/a3: k = 8+2j, 9+2j; likewise b0/b1), accumulated over ks = 0..7 in ONE accumulator.  That is the
// order a library GEMM accumulates in: cuBLAS's nvjet (tcgen05) kernel for X = K.PHI gives the same bits as this
// chain in all 1.57 M entries of a 32768 x 48 product, the permuted-k chain above differs in ~50 per million
// (tools/ubench/sketch_order.cu, sketch_umma.cu; profiles/r02c_sketch_bits.txt).  Used where the result is rounded
// to 16 bit and the reference's bits matter (the Compactor sketch).  Rows: 16-bit, k contiguous, 128 long.
// w[2*ks] = k-slots (2j, 2j+1), w[2*ks+1] = k-slots (8+2j, 9+2j).
struct FragNat128 {
  uint32_t w[16];
  __device__ __forceinline__ void load(const uint16_t* row, int j) {
    const uint32_t* r32 = reinterpret_cast<const uint32_t*>(row);
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = row ? r32[8 * (i >> 1) + 4 * (i & 1) + j] : 0u;
  }
};

}  // namespace cvllm
