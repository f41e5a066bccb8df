// Building blocks of the tensor-core weight-gradient product  G[64 x Nc] = X[K x 64]^T . Y[K x Nc]  whose reduction
// index K is the NODE (tile row) index: both operands are read straight from row-per-node activation buffers in
// shared memory as MN-major tcgen05 operands, so nothing is transposed and no fragment is loaded by a thread.
//
// Buffer layout ("planes"): a [128 rows x 64 floats] activation tile is stored as two planes of [128 rows x 32 floats]
// (columns 0..31 and 32..63), each row 128 bytes, with the MN-major TF32 swizzle (the 32-byte unit index of a row XOR
// (row & 3)): one plane is a stack of canonical atoms (32 MN-elements x 4 K-rows = 512 bytes).  K-groups of 4 rows
// are 512 bytes apart (stride byte offset), the two planes kPlaneBytes apart (leading byte offset).  The same layout
// serves 128-bit row gathers (a row's 32-byte units are permuted, every 16-byte chunk stays aligned).
//
// fp32 accuracy: the tensor core reads the TF32 part of an fp32 word (truncation), so the "hi" operand is the raw
// fp32 buffer itself and the "lo" operand a buffer of residuals v - trunc_tf32(v); G = Xlo.Yhi + Xhi.Ylo + Xhi.Yhi.
#pragma once

#include "hcg_ptx.cuh"

namespace hcg {
namespace tcb {

constexpr int kRows = 128;                         // K of the product = node rows of a tile
constexpr int kPlaneFloats = kRows * 32;           // one 32-column plane
constexpr uint32_t kPlaneBytes = kPlaneFloats * 4;

// float offset of element (row r, column c) of a two-plane activation tile
__device__ __forceinline__ int plane_off(int r, int c) {
  return (c >> 5) * kPlaneFloats + r * 32 + (((((c & 31) >> 3) ^ (r & 3)) << 3) | (c & 7));
}
// float offset of the 16-byte chunk q (0..15: four columns each) of row r
__device__ __forceinline__ int plane_chunk_off(int r, int q) {
  return (q >> 3) * kPlaneFloats + r * 32 + (((((q & 7) >> 1) ^ (r & 3)) << 3) | ((q & 1) << 2));
}

__device__ __forceinline__ float tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

// G[M = 128 (rows 64..127 are don't-care) x N] (+)= X^T . Y over `ksteps` groups of 8 node rows, three split terms.
// x_hi / x_lo / y_hi / y_lo: shared-memory byte addresses of the two-plane tiles (1024-byte aligned).  The A operand
// claims four MN atoms (M = 128): atoms 2 and 3 read whatever follows the X tile (must be inside shared memory) and
// only pollute accumulator rows 64..127, which nobody reads.  Issued by one thread; the caller commits.
template <int N>
__device__ __forceinline__ void issue_xty(uint32_t d_tmem, uint32_t x_hi, uint32_t x_lo, uint32_t y_hi, uint32_t y_lo, int ksteps,
                                          uint32_t accumulate_first) {
  constexpr uint32_t idesc = ptx::make_idesc_tf32_mn(128, N, true, true);
  const uint64_t xh = ptx::make_mnmajor_tf32_desc(x_hi, kPlaneBytes, 512), xl = ptx::make_mnmajor_tf32_desc(x_lo, kPlaneBytes, 512);
  const uint64_t yh = ptx::make_mnmajor_tf32_desc(y_hi, kPlaneBytes, 512), yl = ptx::make_mnmajor_tf32_desc(y_lo, kPlaneBytes, 512);
  for (int s = 0; s < ksteps; ++s) {
    const uint64_t off = (uint64_t)(s * 1024 >> 4);     // next K group: + 1024 bytes in the 14-bit address field
    ptx::mma_tf32_ss(d_tmem, xl + off, yh + off, idesc, s > 0 ? 1u : accumulate_first);
    ptx::mma_tf32_ss(d_tmem, xh + off, yl + off, idesc, 1u);
  }
  for (int s = 0; s < ksteps; ++s) {
    const uint64_t off = (uint64_t)(s * 1024 >> 4);
    ptx::mma_tf32_ss(d_tmem, xh + off, yh + off, idesc, 1u);
  }
}

}  // namespace tcb
}  // namespace hcg
