// The default readout head (emb [*, 128] -> 64 -> C) for a TILE of up to 32 graphs, by 256 threads:
//   hid = lrelu(emb . W1^T + b1)     out = hid . W2^T + b2          (reference model/gcn.py:68-71)
// One fmaf chain per output element with k ascending (then + bias): every caller - k_readout2 (a block per 32
// consecutive graphs), k_readout2_ens, the tail of the edge-list forward kernels (each slot runs the head of the graphs
// it has just pooled) - produces the same bits.
#pragma once

#include "hcg_common.cuh"

namespace hcg {
namespace ro {

constexpr int IN = 128, HID = 64, KC = 32, FG = 32;
constexpr int kEsPitch = FG + 4, kWsPitch = HID + 4, kHsPitch = HID + 1;
constexpr int kSmemFloats = KC * kEsPitch + KC * kWsPitch + FG * kHsPitch;   // 21.6 KB, 16-byte aligned base

__device__ __forceinline__ float lrelu_ro(float v, float slope) { return v > 0.f ? v : v * slope; }

// gid(r): global graph id of tile row r (r < n);  sync(): barrier of the 256 calling threads;  tid: 0 .. 255.
// 16 x 16 threads, each a 2 (graphs) x 4 (hidden units) micro-tile.
template <class GId, class Sync>
__device__ __forceinline__ void head_tile(const float* __restrict__ emb, const float* __restrict__ W1, const float* __restrict__ b1,
                                          const float* __restrict__ W2, const float* __restrict__ b2, float* __restrict__ hid_save,
                                          float* __restrict__ out, int n, GId gid, int C, float slope, float* sm, int tid, Sync sync) {
  float* Es = sm;                       // [k][graph]
  float* Ws = Es + KC * kEsPitch;       // [k][unit]
  float* Hs = Ws + KC * kWsPitch;       // [graph][unit]
  const int tx = tid & 15, ty = tid >> 4;
  float acc[2][4];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
  for (int k0 = 0; k0 < IN; k0 += KC) {
    for (int idx = tid; idx < FG * KC; idx += 256) {
      const int r = idx / KC, c = idx % KC;
      Es[c * kEsPitch + r] = r < n ? emb[(int64_t)gid(r) * IN + k0 + c] : 0.f;
    }
    for (int idx = tid; idx < HID * KC; idx += 256) {
      const int r = idx / KC, c = idx % KC;
      Ws[c * kWsPitch + r] = W1[r * IN + k0 + c];
    }
    sync();
#pragma unroll 8
    for (int kk = 0; kk < KC; ++kk) {
      const float2 e2 = *reinterpret_cast<const float2*>(Es + kk * kEsPitch + ty * 2);
      const float4 w4 = *reinterpret_cast<const float4*>(Ws + kk * kWsPitch + tx * 4);
      const float e[2] = {e2.x, e2.y}, w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(e[a], w[b], acc[a][b]);
    }
    sync();
  }
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int g = ty * 2 + a, o = tx * 4 + b;
      const float v = lrelu_ro(acc[a][b] + b1[o], slope);
      Hs[g * kHsPitch + o] = v;
      if (hid_save && g < n) hid_save[(int64_t)gid(g) * HID + o] = v;
    }
  sync();
  for (int idx = tid; idx < FG * C; idx += 256) {
    const int g = idx % FG, c = idx / FG;
    if (g >= n) continue;
    float s = 0.f;
#pragma unroll 8
    for (int o = 0; o < HID; ++o) s = fmaf(Hs[g * kHsPitch + o], W2[c * HID + o], s);
    out[(int64_t)gid(g) * C + c] = s + b2[c];
  }
  sync();   // (the tile's shared memory may be refilled)
}

}  // namespace ro
}  // namespace hcg
