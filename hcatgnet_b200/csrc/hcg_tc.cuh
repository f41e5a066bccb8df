// Building blocks of the tensor-core (tcgen05, TF32 x 3 split) dense product used inside the
// graph-resident kernel:  D[128 x 64] = A[128 x K] . W[64 x K]^T  with fp32-level accuracy.
//
// fp32 operands are split  v = hi + lo,  hi = v rounded to TF32 (13 low mantissa bits zero),
// lo = (v - hi) rounded to TF32 as well; D accumulates hi.hi + hi.lo + lo.hi in fp32.
// The dropped lo.lo term and the truncation of lo are O(2^-22) relative per product — the same
// order as fp32 rounding — which is what the 1e-5 parity bound of the GCN path needs; plain TF32
// (2^-11) would miss it by two orders of magnitude.
//
// A lives in tensor memory (thread r of the 128-thread group owns row r = TMEM lane r; column k
// holds A[r][k]), written with tcgen05.st straight from the registers that produced the row.
// W^T tiles live in shared memory in the K-major / 128-byte-swizzle layout the MMA reads.
#pragma once

#include "hcg_ptx.cuh"

namespace hcg {
namespace tc {

constexpr int kRows = 128;       // M of the MMA = rows (nodes) per tile = threads per group
constexpr int kN = 64;           // N of the MMA = hidden width
constexpr int kChunkFloats = kN * 32;  // one K-chunk (32 floats = 128 B per row) of a W tile

// round-to-nearest on the magnitude (bit pattern + half an ulp of TF32, then clear): |lo| <= 2^-12 |v|
__device__ __forceinline__ uint32_t tf32_hi(float v) { return (__float_as_uint(v) + 0x1000u) & 0xffffe000u; }
__device__ __forceinline__ uint32_t tf32_lo(float v, uint32_t hi) {
  return (__float_as_uint(v - __uint_as_float(hi)) + 0x1000u) & 0xffffe000u;  // rounded too: |err| <= 2^-24 |v|
}

// float offset of element (n, k) inside a W tile of [K/32 chunks][64 rows][32 floats], swizzled
__device__ __forceinline__ int w_tile_offset(int n, int k) {
  const int chunk = k >> 5, kk = k & 31;
  return chunk * kChunkFloats + (n >> 3) * 256 + (n & 7) * 32 + ((((kk >> 2) ^ (n & 7)) << 2) | (kk & 3));
}

// W: global [64][Kin] row-major (state_dict layout [out][in]).  Writes the hi / lo tiles for K
// padded to Kpad (32 or 64).  Cooperative over `nthreads` threads; tiles must be 1024-B aligned.
__device__ __forceinline__ void prep_weight_tiles(const float* __restrict__ W, int Kin, int Kpad, float* s_hi,
                                                  float* s_lo, int tid, int nthreads) {
  constexpr int U = 4;   // loads of a thread in flight at once (the loop is a chain of memory latencies otherwise)
  for (int base = tid; base < kN * Kpad; base += U * nthreads) {
    float v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads, n = idx / Kpad, k = idx - n * Kpad;
      v[u] = idx < kN * Kpad && k < Kin ? __ldg(W + n * Kin + k) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads, n = idx / Kpad, k = idx - n * Kpad;
      if (idx >= kN * Kpad) break;
      const uint32_t hi = tf32_hi(v[u]), lo = tf32_lo(v[u], hi);
      const int off = w_tile_offset(n, k);
      s_hi[off] = __uint_as_float(hi);
      s_lo[off] = __uint_as_float(lo);
    }
  }
}

// One layer's product: ksteps = K/8 MMAs of K = 8 for each of the three split terms, all into ONE fp32
// accumulator.  Issued by a single thread; the caller commits.  a_*: TMEM addresses (lane 0 of the
// group, first column).  The tensor core truncates (rounds toward zero) when it adds into its
// accumulator, a bias that grows with the number of accumulation steps at the accumulator's magnitude.
// The small cross terms (hi.lo + lo.hi, 2^-11 of the result) are therefore accumulated FIRST, while the
// accumulator is still small, and the hi.hi chain goes on top: the truncation bias is that of `ksteps`
// steps at full magnitude, not of 3 x ksteps.
template <int KSTEPS>
__device__ __forceinline__ void issue_split_mma(uint32_t d, uint32_t a_hi, uint32_t a_lo, uint32_t b_hi_smem,
                                                uint32_t b_lo_smem, uint32_t accumulate_first = 0u) {
  constexpr uint32_t idesc = ptx::make_idesc_tf32(kRows, kN);
  // descriptors of K-step s = descriptor of step 0 + a compile-time offset in its 14-bit address field
  // (16-byte units; shared-memory addresses are below 2^18, so the field never carries)
  const uint64_t bh = ptx::make_kmajor_sw128_desc(b_hi_smem), bl = ptx::make_kmajor_sw128_desc(b_lo_smem);
#pragma unroll
  for (int s = 0; s < KSTEPS; ++s) {
    const uint64_t boff = (uint64_t)(((s >> 2) * (kChunkFloats * 4) + (s & 3) * 32) >> 4);
    ptx::mma_tf32_ts(d, a_lo + s * 8, bh + boff, idesc, s > 0 ? 1u : accumulate_first);  // (continues an earlier chain?)
    ptx::mma_tf32_ts(d, a_hi + s * 8, bl + boff, idesc, 1u);
  }
#pragma unroll
  for (int s = 0; s < KSTEPS; ++s) {
    const uint64_t boff = (uint64_t)(((s >> 2) * (kChunkFloats * 4) + (s & 3) * 32) >> 4);
    ptx::mma_tf32_ts(d, a_hi + s * 8, bh + boff, idesc, 1u);
  }
}

// Row r's 16 accumulator columns [c0, c0+16), as floats.
__device__ __forceinline__ void load_acc16(uint32_t d, uint32_t trow, int c0, float (&out)[16]) {
  uint32_t rb[16];
  ptx::tmem_ld16(d + trow + c0, rb);
  ptx::tmem_wait_ld();
#pragma unroll
  for (int q = 0; q < 16; ++q) out[q] = __uint_as_float(rb[q]);
}

// Splits 16 consecutive values of this thread's row and stores them to tensor memory columns
// [col, col+16) of the hi and lo operands.  `trow` = TMEM lane base of this warp (32 * warp-in-group) << 16.
__device__ __forceinline__ void store_split16(const float (&v)[16], uint32_t a_hi, uint32_t a_lo, uint32_t trow, int col) {
  uint32_t h[16], l[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    h[q] = tf32_hi(v[q]);
    // the residual goes in unrounded: the tensor core reads only the TF32 bits of an operand (the 13 low
    // mantissa bits are ignored), i.e. truncates it: |err| <= 2^-10 |lo| <= 2^-21 |v|, typically 2^-23
    l[q] = __float_as_uint(v[q] - __uint_as_float(h[q]));
  }
  ptx::tmem_st16(a_hi + trow + col, h);
  ptx::tmem_st16(a_lo + trow + col, l);
}

// Eight consecutive values: columns [col, col+8).
__device__ __forceinline__ void store_split8(const float (&v)[8], uint32_t a_hi, uint32_t a_lo, uint32_t trow, int col) {
  uint32_t h[8], l[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    h[q] = tf32_hi(v[q]);
    l[q] = __float_as_uint(v[q] - __uint_as_float(h[q]));
  }
  ptx::tmem_st8(a_hi + trow + col, h);
  ptx::tmem_st8(a_lo + trow + col, l);
}

}  // namespace tc
}  // namespace hcg
