// Graph-resident forward of the GCN hot path (H = 64): one CTA owns one reaction graph at a time
// and keeps every intermediate of that graph on chip.
//
//   x[n,F] --stage--> smem
//   layer 0 :  h1 = lrelu( (Â x) W0^T + b0 )         Â x aggregated per 16-row warp tile, then a
//   layer l :  h  = lrelu( (Â h) Wl^T + bl )         register-tiled 16x64xK product against W in smem
//   pooling :  emb[g] = [ max_i h_i || mean_i h_i ]  taken from the last layer's accumulators
//
// (reference model/gcn.py:58-66; GCNConv computes Â(XW) — the re-association (ÂX)W is algebraically
// identical, moves fp32 rounding by ~1e-7, and lets the pooling run in the product's epilogue.)
// HBM traffic per graph is the compulsory x + CSR + dis read and the 2H-float embedding write;
// nothing of size [N,H] is written unless the caller asks for the activations (training).
// Graphs that do not fit the shared-memory capacity run the same code on global scratch buffers.
// All reductions are order-fixed: results are run-to-run identical.
#include "hcg_fused.cuh"

namespace hcg {

constexpr int kStatusCrossEdge = 4;  // an edge leaves its graph (cannot come out of PyG collate)

struct GraphCtx {
  int n0, n, e0, e;          // first node, node count, first CSR slot, slot count
  const int32_t* rp;         // [n+1] raw rowptr values (global slot numbers)
  const int32_t* cl;         // [e]   raw col values (global node numbers)
  const float* ds;           // [n]   deg^-1/2
};

// ---- aggregation of one 16-row tile: tile[r][0..W) = sum_j norm_ij * src[j][0..W) -------------
// W == 64: a half-warp owns a row, each lane one float4.  Neighbours in CSR order, self-loop last.
__device__ __forceinline__ void agg_tile64(const GraphCtx& g, const float* src, int row0,
                                           float* __restrict__ tile, float fill, int lane, int* status) {
  const int half = lane >> 4, l16 = lane & 15;
#pragma unroll 1
  for (int m = 0; m < kTileRows / 2; ++m) {
    const int r = 2 * m + half;
    const int i = row0 + r;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < g.n) {
      const int beg = g.rp[i] - g.e0, end = g.rp[i + 1] - g.e0;
      const float di = g.ds[i];
      int k = beg;
      for (; k + 2 <= end; k += 2) {
        int j0 = g.cl[k] - g.n0, j1 = g.cl[k + 1] - g.n0;
        bool ok0 = (unsigned)j0 < (unsigned)g.n, ok1 = (unsigned)j1 < (unsigned)g.n;
        if (!(ok0 && ok1)) { if (status) atomicOr(status, kStatusCrossEdge); j0 = ok0 ? j0 : i; j1 = ok1 ? j1 : i; }
        float w0 = ok0 ? g.ds[j0] * di : 0.f, w1 = ok1 ? g.ds[j1] * di : 0.f;
        float4 v0 = *reinterpret_cast<const float4*>(src + j0 * kFusedH + 4 * l16);
        float4 v1 = *reinterpret_cast<const float4*>(src + j1 * kFusedH + 4 * l16);
        acc.x = fmaf(w0, v0.x, acc.x); acc.y = fmaf(w0, v0.y, acc.y); acc.z = fmaf(w0, v0.z, acc.z); acc.w = fmaf(w0, v0.w, acc.w);
        acc.x = fmaf(w1, v1.x, acc.x); acc.y = fmaf(w1, v1.y, acc.y); acc.z = fmaf(w1, v1.z, acc.z); acc.w = fmaf(w1, v1.w, acc.w);
      }
      if (k < end) {
        int j0 = g.cl[k] - g.n0;
        bool ok0 = (unsigned)j0 < (unsigned)g.n;
        if (!ok0) { if (status) atomicOr(status, kStatusCrossEdge); j0 = i; }
        float w0 = ok0 ? g.ds[j0] * di : 0.f;
        float4 v0 = *reinterpret_cast<const float4*>(src + j0 * kFusedH + 4 * l16);
        acc.x = fmaf(w0, v0.x, acc.x); acc.y = fmaf(w0, v0.y, acc.y); acc.z = fmaf(w0, v0.z, acc.z); acc.w = fmaf(w0, v0.w, acc.w);
      }
      const float ws = (di * fill) * di;
      float4 vs = *reinterpret_cast<const float4*>(src + i * kFusedH + 4 * l16);
      acc.x = fmaf(ws, vs.x, acc.x); acc.y = fmaf(ws, vs.y, acc.y); acc.z = fmaf(ws, vs.z, acc.z); acc.w = fmaf(ws, vs.w, acc.w);
    }
    *reinterpret_cast<float4*>(tile + r * kTileStride + 4 * l16) = acc;
  }
}

// Input layer: rows of F floats at stride F (F <= 64, any alignment); lanes run over features.
// Columns F..Fp-1 of the tile are zeroed so that the product can run over Fp (a multiple of 4).
__device__ __forceinline__ void agg_tile_x(const GraphCtx& g, const float* __restrict__ xs, int F, int Fp, int row0,
                                           float* __restrict__ tile, float fill, int lane, int* status) {
  const bool two = F > 32;
#pragma unroll 1
  for (int r = 0; r < kTileRows; ++r) {
    const int i = row0 + r;
    float a0 = 0.f, a1 = 0.f;
    if (i < g.n) {
      const int beg = g.rp[i] - g.e0, end = g.rp[i + 1] - g.e0;
      const float di = g.ds[i];
      const int c0 = lane < F ? lane : 0;
      const int c1 = (two && lane + 32 < F) ? lane + 32 : 0;
      for (int k = beg; k < end; ++k) {
        int j = g.cl[k] - g.n0;
        bool ok = (unsigned)j < (unsigned)g.n;
        if (!ok) { if (status) atomicOr(status, kStatusCrossEdge); j = i; }
        float w = ok ? g.ds[j] * di : 0.f;
        a0 = fmaf(w, xs[j * F + c0], a0);
        if (two) a1 = fmaf(w, xs[j * F + c1], a1);
      }
      const float ws = (di * fill) * di;
      a0 = fmaf(ws, xs[i * F + c0], a0);
      if (two) a1 = fmaf(ws, xs[i * F + c1], a1);
    }
    if (lane < Fp) tile[r * kTileStride + lane] = lane < F ? a0 : 0.f;
    if (two && lane + 32 < Fp) tile[r * kTileStride + lane + 32] = lane + 32 < F ? a1 : 0.f;
  }
}

// ---- 16 x 64 x K product of a warp tile against W^T (smem, [K][64]) ---------------------------
// lane = (rg, cg): rows rg + 2m (m = 0..7), columns 4cg..4cg+3  ->  32 accumulators per lane.
__device__ __forceinline__ void tile_gemm(const float* __restrict__ tile, const float* __restrict__ Wt, int K,
                                          int lane, float (&acc)[8][4]) {
  const int rg = lane >> 4, cg = lane & 15;
#pragma unroll
  for (int m = 0; m < 8; ++m)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[m][q] = 0.f;
  const float* trow = tile + rg * kTileStride;
  const float* wcol = Wt + 4 * cg;
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    const float4 w0 = *reinterpret_cast<const float4*>(wcol + (k + 0) * kFusedH);
    const float4 w1 = *reinterpret_cast<const float4*>(wcol + (k + 1) * kFusedH);
    const float4 w2 = *reinterpret_cast<const float4*>(wcol + (k + 2) * kFusedH);
    const float4 w3 = *reinterpret_cast<const float4*>(wcol + (k + 3) * kFusedH);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const float4 a = *reinterpret_cast<const float4*>(trow + 2 * m * kTileStride + k);
      acc[m][0] = fmaf(a.x, w0.x, acc[m][0]); acc[m][1] = fmaf(a.x, w0.y, acc[m][1]);
      acc[m][2] = fmaf(a.x, w0.z, acc[m][2]); acc[m][3] = fmaf(a.x, w0.w, acc[m][3]);
      acc[m][0] = fmaf(a.y, w1.x, acc[m][0]); acc[m][1] = fmaf(a.y, w1.y, acc[m][1]);
      acc[m][2] = fmaf(a.y, w1.z, acc[m][2]); acc[m][3] = fmaf(a.y, w1.w, acc[m][3]);
      acc[m][0] = fmaf(a.z, w2.x, acc[m][0]); acc[m][1] = fmaf(a.z, w2.y, acc[m][1]);
      acc[m][2] = fmaf(a.z, w2.z, acc[m][2]); acc[m][3] = fmaf(a.z, w2.w, acc[m][3]);
      acc[m][0] = fmaf(a.w, w3.x, acc[m][0]); acc[m][1] = fmaf(a.w, w3.y, acc[m][1]);
      acc[m][2] = fmaf(a.w, w3.z, acc[m][2]); acc[m][3] = fmaf(a.w, w3.w, acc[m][3]);
    }
  }
}

// ---- one graph, all layers --------------------------------------------------------------------
// xs: the graph's node features (row stride F); bufA/bufB: [n,64] ping-pong buffers for the hidden
// activations; all three live in shared memory on the fast path and in global scratch otherwise.
struct Smem {
  float* W0t;      // [Fp][64]
  float* Wt;       // [(L-1)][64][64]
  float* bias;     // [L][64]
  float* xs;       // [ncap*F]
  float* bufA;     // [ncap][64]
  float* bufB;     // [ncap][64] (L > 2 only)
  float* tiles;    // [warps][16][68]
  float* pool;     // [warps][128]
  int32_t* rp;     // [ncap+1]
  int32_t* cl;     // [ecap]
  float* ds;       // [ncap]
};

__device__ __forceinline__ void run_graph(const FusedFwdArgs& a, const Smem& s, const GraphCtx& g,
                                          const float* __restrict__ xs, float* __restrict__ bufA,
                                          float* __restrict__ bufB, int64_t gid) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rg = lane >> 4, cg = lane & 15;
  float* tile = s.tiles + warp * (kTileRows * kTileStride);
  const int ntiles = (g.n + kTileRows - 1) / kTileRows;
  float pmax[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
  float psum[4] = {0.f, 0.f, 0.f, 0.f};

  const float* src = xs;
  float* dst = bufA;
  for (int l = 0; l < a.L; ++l) {
    const bool last = l == a.L - 1;
    const float* Wt = l == 0 ? s.W0t : s.Wt + (size_t)(l - 1) * kFusedH * kFusedH;
    const int K = l == 0 ? a.Fp : kFusedH;
    const float4 b4 = *reinterpret_cast<const float4*>(s.bias + l * kFusedH + 4 * cg);
    float* save = a.save_act[l];
    for (int t = warp; t < ntiles; t += kFusedWarps) {
      const int row0 = t * kTileRows;
      if (l == 0) agg_tile_x(g, src, a.F, a.Fp, row0, tile, a.fill, lane, a.status);
      else agg_tile64(g, src, row0, tile, a.fill, lane, a.status);
      __syncwarp();
      float acc[8][4];
      tile_gemm(tile, Wt, K, lane, acc);
      __syncwarp();  // the tile is rewritten by this warp's next aggregation
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        const int i = row0 + rg + 2 * m;
        float4 v;
        v.x = lrelu(acc[m][0] + b4.x, a.slope); v.y = lrelu(acc[m][1] + b4.y, a.slope);
        v.z = lrelu(acc[m][2] + b4.z, a.slope); v.w = lrelu(acc[m][3] + b4.w, a.slope);
        if (i < g.n) {
          if (!last) *reinterpret_cast<float4*>(dst + i * kFusedH + 4 * cg) = v;
          if (save) *reinterpret_cast<float4*>(save + ((int64_t)g.n0 + i) * kFusedH + 4 * cg) = v;
          if (last) {
            pmax[0] = fmaxf(pmax[0], v.x); pmax[1] = fmaxf(pmax[1], v.y);
            pmax[2] = fmaxf(pmax[2], v.z); pmax[3] = fmaxf(pmax[3], v.w);
            psum[0] += v.x; psum[1] += v.y; psum[2] += v.z; psum[3] += v.w;
          }
        }
      }
    }
    if (!last) {
      __syncthreads();  // every row of this layer's output is in place (block-uniform branch)
      src = dst;
      dst = (dst == bufA) ? bufB : bufA;
    }
  }
  // pooling: lanes (0,cg) and (1,cg) hold the even / odd rows of the same 4 columns
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    pmax[q] = fmaxf(pmax[q], __shfl_xor_sync(0xffffffffu, pmax[q], 16));
    psum[q] += __shfl_xor_sync(0xffffffffu, psum[q], 16);
  }
  if (rg == 0) {
    *reinterpret_cast<float4*>(s.pool + warp * 2 * kFusedH + 4 * cg) = make_float4(pmax[0], pmax[1], pmax[2], pmax[3]);
    *reinterpret_cast<float4*>(s.pool + warp * 2 * kFusedH + kFusedH + 4 * cg) = make_float4(psum[0], psum[1], psum[2], psum[3]);
  }
  __syncthreads();
  if (tid < 2 * kFusedH) {
    const bool is_sum = tid >= kFusedH;
    float v = is_sum ? 0.f : -INFINITY;
#pragma unroll
    for (int w = 0; w < kFusedWarps; ++w) {
      float p = s.pool[w * 2 * kFusedH + tid];
      v = is_sum ? v + p : fmaxf(v, p);
    }
    if (g.n <= 0) v = 0.f;  // an empty graph cannot come out of PyG collate; defined as zeros
    else if (is_sum) v = v / (float)g.n;
    a.emb[gid * 2 * kFusedH + tid] = v;
  }
}

__global__ void __launch_bounds__(kFusedThreads, 2)
k_fused_fwd(const FusedFwdArgs a) {
  extern __shared__ __align__(16) float smem_f[];
  const int tid = threadIdx.x;
  const int64_t total = a.list ? (int64_t)*a.list_count : a.B;
  if ((int64_t)blockIdx.x >= total) return;  // nothing for this CTA (an empty overflow list costs ~nothing)
  Smem s;
  {
    float* p = smem_f;
    s.W0t = p; p += a.Fp * kFusedH;
    s.Wt = p; p += (a.L - 1) * kFusedH * kFusedH;
    s.bias = p; p += a.L * kFusedH;
    s.bufA = p; p += a.ncap * kFusedH;
    s.bufB = p; p += (a.L > 2 ? a.ncap * kFusedH : 0);
    s.tiles = p; p += kFusedWarps * kTileRows * kTileStride;
    s.pool = p; p += kFusedWarps * 2 * kFusedH;
    s.xs = p; p += (a.ncap * a.F + 3) / 4 * 4;
    s.ds = p; p += a.ncap;
    s.rp = reinterpret_cast<int32_t*>(p); p += (a.ncap + 1 + 3) / 4 * 4;
    s.cl = reinterpret_cast<int32_t*>(p);
  }
  // weights: W[out][in] (state_dict layout) -> W^T[in][out] in shared memory, once per CTA
  for (int idx = tid; idx < a.Fp * kFusedH; idx += kFusedThreads) {
    int k = idx / kFusedH, o = idx % kFusedH;
    s.W0t[idx] = k < a.F ? __ldg(a.conv_w[0] + o * a.F + k) : 0.f;
  }
  for (int l = 1; l < a.L; ++l)
    for (int idx = tid; idx < kFusedH * kFusedH; idx += kFusedThreads) {
      int k = idx / kFusedH, o = idx % kFusedH;
      s.Wt[(size_t)(l - 1) * kFusedH * kFusedH + idx] = __ldg(a.conv_w[l] + o * kFusedH + k);
    }
  for (int idx = tid; idx < a.L * kFusedH; idx += kFusedThreads)
    s.bias[idx] = __ldg(a.conv_b[idx / kFusedH] + idx % kFusedH);
  __syncthreads();

  for (int64_t it = blockIdx.x; it < total; it += gridDim.x) {
    const int64_t gid = a.list ? (int64_t)a.list[it] : it;
    GraphCtx g;
    g.n0 = a.graph_ptr[gid];
    g.n = a.graph_ptr[gid + 1] - g.n0;
    g.e0 = g.n > 0 ? a.rowptr[g.n0] : 0;
    g.e = g.n > 0 ? a.rowptr[g.n0 + g.n] - g.e0 : 0;
    if (g.n <= a.ncap && g.e <= a.ecap) {
      // stage the graph: x rows, CSR slice and deg^-1/2 are contiguous ranges of the batch arrays
      const float* xg = a.x + (int64_t)g.n0 * a.F;
      const int nx = g.n * a.F;
      for (int idx = tid; idx < nx; idx += kFusedThreads) s.xs[idx] = __ldg(xg + idx);
      for (int idx = tid; idx <= g.n; idx += kFusedThreads) s.rp[idx] = __ldg(a.rowptr + g.n0 + idx);
      for (int idx = tid; idx < g.e; idx += kFusedThreads) s.cl[idx] = __ldg(a.col + g.e0 + idx);
      for (int idx = tid; idx < g.n; idx += kFusedThreads) s.ds[idx] = __ldg(a.dis + g.n0 + idx);
      __syncthreads();
      g.rp = s.rp; g.cl = s.cl; g.ds = s.ds;
      run_graph(a, s, g, s.xs, s.bufA, s.bufB, gid);
    } else {
      g.rp = a.rowptr + g.n0; g.cl = a.col + g.e0; g.ds = a.dis + g.n0;
      run_graph(a, s, g, a.x + (int64_t)g.n0 * a.F, a.scratch_a + (int64_t)g.n0 * kFusedH,
                a.scratch_b + (int64_t)g.n0 * kFusedH, gid);
    }
    __syncthreads();  // pool / staging buffers are reused by the next graph
  }
}

size_t fused_fwd_smem_bytes(int F, int L, int ncap, int ecap) {
  const int Fp = (F + 3) / 4 * 4;
  size_t fl = (size_t)Fp * kFusedH + (size_t)(L - 1) * kFusedH * kFusedH + (size_t)L * kFusedH +
              (size_t)ncap * kFusedH * (L > 2 ? 2 : 1) + (size_t)kFusedWarps * kTileRows * kTileStride +
              (size_t)kFusedWarps * 2 * kFusedH + (size_t)(ncap * F + 3) / 4 * 4 + ncap + (ncap + 1 + 3) / 4 * 4 + ecap;
  return fl * 4;
}

int launch_fused_fwd(const FusedFwdArgs& a, size_t smem, cudaStream_t st, int64_t max_ctas) {
  if (a.B <= 0) return HCG_OK;
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_fused_fwd, smem, configured)) return rc;
  int per_sm = 0;
  HCG_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fused_fwd, kFusedThreads, smem));
  if (per_sm < 1) return set_err(HCG_ERR_UNSUPPORTED, "fused forward does not fit on an SM (%zu B smem)", smem);
  int64_t grid = (int64_t)kNumSMs * per_sm;
  if (grid > a.B) grid = a.B;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  k_fused_fwd<<<(unsigned)grid, kFusedThreads, smem, st>>>(a);
  HCG_LAUNCH_OK(st, "k_fused_fwd");
  return HCG_OK;
}

}  // namespace hcg
