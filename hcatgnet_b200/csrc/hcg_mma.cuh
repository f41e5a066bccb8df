// Warp-level TF32 tensor-core building blocks (mma.sync m16n8k8) with the 3-term error-compensated split,
// and the X^T . Y product over XOR-swizzled [rows][64] shared-memory planes.  Used by the graph-resident
// backward (hcg_fused_bwd.cu) and by the weight-gradient kernel of the layer-wise backward (hcg_net.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace hcg {

__device__ __forceinline__ int swz(int r) { return ((r & 3) << 3) | (r & 4); }
__device__ __forceinline__ int at(int r, int c) { return r * 64 + (c ^ swz(r)); }

__device__ __forceinline__ void split_tf32(float v, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;  // round to nearest TF32 on the magnitude
  lo = __float_as_uint(v - __uint_as_float(hi));      // exact residual; the tensor core reads its TF32 bits
}

// D[16x8] += A[16x8] . B[8x8]   (A row-major fragment, B column-major fragment, fp32 accumulators)
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_split(float (&d)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4],
                                          uint32_t bh0, uint32_t bh1, uint32_t bl0, uint32_t bl1) {
  mma_tf32(d, al, bh0, bh1);  // small cross terms first
  mma_tf32(d, ah, bl0, bl1);
  mma_tf32(d, ah, bh0, bh1);
}

// acc[j] += (Q^T . Y)[16 mt .. +16][8 (nb + j) .. +8]  over rows 0..npad-1 (pad rows of Q are zero)
template <int NT>
__device__ __forceinline__ void xty_tiles(const float* __restrict__ Q, const float* __restrict__ Y, int npad, int mt,
                                          int nb, float (&acc)[NT][4]) {
  const int lane = threadIdx.x & 31, g8 = lane >> 2, t4 = lane & 3;
  float d[NT][4];
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) d[j][q] = 0.f;
  // fragment rows are r0 + t4 (swizzle t4 << 3) and r0 + t4 + 4 (swizzle (t4 << 3) | 4) with r0 a multiple of 8:
  // every offset is a per-thread constant plus r0 * 64
  const int s0 = t4 << 3, s1 = s0 | 4, oc = 16 * mt + g8;
  const float* qa = Q + t4 * 64 + (oc ^ s0);
  const float* qb = Q + t4 * 64 + ((oc + 8) ^ s0);
  const float* qc = Q + (t4 + 4) * 64 + (oc ^ s1);
  const float* qd = Q + (t4 + 4) * 64 + ((oc + 8) ^ s1);
  const float* yr0 = Y + t4 * 64;
  const float* yr1 = Y + (t4 + 4) * 64;
  int off[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) off[j] = (8 * (nb + j) + g8) ^ s0;  // column of tile j in row r0 + t4; row + 4: ^ 4
  for (int r0 = 0; r0 < npad; r0 += 8) {
    const int o = r0 * 64;
    uint32_t ah[4], al[4];
    split_tf32(qa[o], ah[0], al[0]);
    split_tf32(qb[o], ah[1], al[1]);
    split_tf32(qc[o], ah[2], al[2]);
    split_tf32(qd[o], ah[3], al[3]);
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      uint32_t bh0, bl0, bh1, bl1;
      split_tf32(yr0[o + off[j]], bh0, bl0);
      split_tf32(yr1[o + (off[j] ^ 4)], bh1, bl1);
      mma_split(d[j], ah, al, bh0, bh1, bl0, bl1);
    }
  }
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[j][q] += d[j][q];
}

// The same product over planes that were split into TF32 hi / lo parts when they were staged (each element is
// split once, by the thread that stages it, instead of once per warp that consumes it): the loop is LDS + MMA.
template <int NT>
__device__ __forceinline__ void xty_tiles_presplit(const float* __restrict__ Qh, const float* __restrict__ Ql,
                                                   const float* __restrict__ Yh, const float* __restrict__ Yl, int npad,
                                                   int mt, int nb, float (&acc)[NT][4]) {
  const int lane = threadIdx.x & 31, g8 = lane >> 2, t4 = lane & 3;
  float d[NT][4];
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) d[j][q] = 0.f;
  const int s0 = t4 << 3, s1 = s0 | 4, oc = 16 * mt + g8;
  const int a0 = t4 * 64 + (oc ^ s0), a1 = t4 * 64 + ((oc + 8) ^ s0);
  const int a2 = (t4 + 4) * 64 + (oc ^ s1), a3 = (t4 + 4) * 64 + ((oc + 8) ^ s1);
  int off[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) off[j] = t4 * 64 + ((8 * (nb + j) + g8) ^ s0);  // row r0 + t4; row + 4: + 256, column ^ 4
  const uint32_t* qh = reinterpret_cast<const uint32_t*>(Qh);
  const uint32_t* ql = reinterpret_cast<const uint32_t*>(Ql);
  const uint32_t* yh = reinterpret_cast<const uint32_t*>(Yh);
  const uint32_t* yl = reinterpret_cast<const uint32_t*>(Yl);
  for (int r0 = 0; r0 < npad; r0 += 8) {
    const int o = r0 * 64;
    const uint32_t ah[4] = {qh[o + a0], qh[o + a1], qh[o + a2], qh[o + a3]};
    const uint32_t al[4] = {ql[o + a0], ql[o + a1], ql[o + a2], ql[o + a3]};
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int b0 = o + off[j], b1 = o + 256 + (off[j] ^ 4);
      mma_split(d[j], ah, al, yh[b0], yh[b1], yl[b0], yl[b1]);
    }
  }
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[j][q] += d[j][q];
}

}  // namespace hcg
