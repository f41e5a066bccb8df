// General layer-wise path of the GCN hot path (any graph size, any H/F): forward and backward.
//
//   conv layer l :  T = h_in · W_l^T            k_gemm      (GCNConv.lin,  reference model/gcn.py:18-20,27-29)
//                   h = lrelu(Â·T + b_l)         k_agg       (GCNConv.propagate + bias + LeakyReLU, :58-63)
//   pooling      :  emb = [max ‖ mean]           k_pool      (global_max_pool / global_mean_pool, :65-66)
//   readout      :  z = lrelu(z·W^T + b) ...     k_gemm      (readout MLP, :70-71)
//
// Â = D^-1/2 (A + I) D^-1/2 lives in the dst-sorted CSR built by hcg_collate; every reduction
// (neighbour sum, segment max/mean, weight-gradient sums) walks its operands in a fixed order,
// so there are no floating-point atomics and results are run-to-run identical.
#include <atomic>

#include "hcg_common.cuh"
#include "hcg_fused.cuh"
#include "hcg_mma.cuh"

namespace hcg {

// ==========================================================================================
// Dense tile GEMM:  C[M,Nc] = A[M,K] · op(B)  (+ epilogue)
//   B_T = true : B is [Nc, K] row-major (a weight stored [out, in])      C = A · B^T
//   B_T = false: B is [K, Nc] row-major                                    C = A · B
// ==========================================================================================
enum { EPI_NONE = 0, EPI_BIAS = 1, EPI_BIAS_LRELU = 2, EPI_MASK = 3, EPI_ADD = 4 };

constexpr int GM = 128, GN = 64, GK = 32;

template <bool B_T, int EPI>
__global__ void __launch_bounds__(256)
k_gemm(const float* __restrict__ A, const float* __restrict__ Bm, float* __restrict__ C,
       int64_t M, int K, int Nc, const float* __restrict__ aux, float slope) {
  __shared__ __align__(16) float As[GK][GM + 4];
  __shared__ __align__(16) float Bs[GK][GN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, each 8 rows x 4 cols
  const int64_t m0 = (int64_t)blockIdx.x * GM;
  const int n0 = blockIdx.y * GN;

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += GK) {
    const int kmax = min(GK, K - k0);
    // A tile: a warp reads one row's GK consecutive k (rows of A are contiguous)
#pragma unroll 4
    for (int idx = tid; idx < GM * GK; idx += 256) {
      int r = idx / GK, c = idx % GK;
      float v = 0.f;
      if (m0 + r < M && c < kmax) v = A[(m0 + r) * K + k0 + c];
      As[c][r] = v;
    }
    if (B_T) {
#pragma unroll 4
      for (int idx = tid; idx < GN * GK; idx += 256) {
        int n = idx / GK, c = idx % GK;
        float v = 0.f;
        if (n0 + n < Nc && c < kmax) v = Bm[(int64_t)(n0 + n) * K + k0 + c];
        Bs[c][n] = v;
      }
    } else {
#pragma unroll 4
      for (int idx = tid; idx < GN * GK; idx += 256) {
        int c = idx / GN, n = idx % GN;
        float v = 0.f;
        if (n0 + n < Nc && c < kmax) v = Bm[(int64_t)(k0 + c) * Nc + n0 + n];
        Bs[c][n] = v;
      }
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < kmax; ++kk) {
      float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 8]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 8 + 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int64_t m = m0 + ty * 8 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= Nc) continue;
      float v = acc[i][j];
      if (EPI == EPI_BIAS) v += aux[n];
      if (EPI == EPI_BIAS_LRELU) v = lrelu(v + aux[n], slope);
      if (EPI == EPI_MASK) v *= lrelu_grad_from_out(aux[m * Nc + n], slope);
      if (EPI == EPI_ADD) v += aux[m * Nc + n];
      C[m * Nc + n] = v;
    }
  }
}

// hcg_tcgemm.cu: the same product on the tensor cores (split-TF32 tcgen05.mma), Nc in {64,128,192,256}
size_t tc_gemm_image_floats(int K, int Nc);
bool tc_gemm_applicable(int64_t M, int K, int Nc);
int launch_tc_gemm(const float* A, const float* W, float* C, int64_t M, int K, int Nc, bool transposed, int epi,
                   const float* aux, float slope, float* img, cudaStream_t st);

// wimg: scratch for the tensor-core path's operand image (NULL = fp32 FMA tiles only)
template <bool B_T>
static int launch_gemm(const float* A, const float* Bm, float* C, int64_t M, int K, int Nc, int epi,
                       const float* aux, float slope, cudaStream_t st, float* wimg = nullptr) {
  if (M == 0 || Nc == 0) return HCG_OK;
  {
    // HCG_CFG_GEMM_PATH: 0 = tensor cores for output widths 64..256 on batches with enough row tiles (0.40 -> 0.23 ms at
    // Nc = 64, 3.96 -> 0.99 ms at Nc = 256 for 1.31 M rows), 1 = never, 2 = whenever applicable
    const int gp = config_get(HCG_CFG_GEMM_PATH);
    const bool shape_ok = wimg != nullptr && (epi == EPI_NONE || epi == EPI_MASK) && tc_gemm_applicable(M, K, Nc) &&
                          (reinterpret_cast<uintptr_t>(C) & 15) == 0 && (epi != EPI_MASK || (reinterpret_cast<uintptr_t>(aux) & 15) == 0);
    if (shape_ok && (gp == 2 || (gp == 0 && Nc >= 64 && M >= 128 * 64)))
      return launch_tc_gemm(A, Bm, C, M, K, Nc, !B_T, epi == EPI_MASK ? 1 : 0, aux, slope, wimg, st);
  }
  dim3 grid((unsigned)cdiv(M, GM), (unsigned)cdiv(Nc, GN));
  switch (epi) {
    case EPI_NONE: k_gemm<B_T, EPI_NONE><<<grid, 256, 0, st>>>(A, Bm, C, M, K, Nc, aux, slope); break;
    case EPI_BIAS: k_gemm<B_T, EPI_BIAS><<<grid, 256, 0, st>>>(A, Bm, C, M, K, Nc, aux, slope); break;
    case EPI_BIAS_LRELU: k_gemm<B_T, EPI_BIAS_LRELU><<<grid, 256, 0, st>>>(A, Bm, C, M, K, Nc, aux, slope); break;
    case EPI_MASK: k_gemm<B_T, EPI_MASK><<<grid, 256, 0, st>>>(A, Bm, C, M, K, Nc, aux, slope); break;
    case EPI_ADD: k_gemm<B_T, EPI_ADD><<<grid, 256, 0, st>>>(A, Bm, C, M, K, Nc, aux, slope); break;
    default: return set_err(HCG_ERR_INVALID_ARG, "bad gemm epilogue %d", epi);
  }
  HCG_LAUNCH_OK(st, "k_gemm");
  return HCG_OK;
}

// ==========================================================================================
// Sparse aggregation  out[i] = act( bias + sum_{k in row i} norm_k · T[col[k]] + selfnorm_i · T[i] )
// A group of `lpr` lanes owns one destination row; each lane VEC consecutive features.
// Neighbours are visited in CSR order (= original edge order), the self-loop last, exactly the
// order in which the reference's index_add_ accumulates.
// ==========================================================================================
template <int VEC>
struct Vec;
template <>
struct Vec<4> {
  float4 v;
  __device__ __forceinline__ void zero() { v = make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ __forceinline__ void load(const float* p) { v = ldg4(p); }
  __device__ __forceinline__ void fma(float w, const Vec<4>& o) {
    v.x = fmaf(w, o.v.x, v.x); v.y = fmaf(w, o.v.y, v.y); v.z = fmaf(w, o.v.z, v.z); v.w = fmaf(w, o.v.w, v.w);
  }
  __device__ __forceinline__ float get(int i) const { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }
  __device__ __forceinline__ void set(int i, float f) { if (i == 0) v.x = f; else if (i == 1) v.y = f; else if (i == 2) v.z = f; else v.w = f; }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
};
template <>
struct Vec<1> {
  float v;
  __device__ __forceinline__ void zero() { v = 0.f; }
  __device__ __forceinline__ void load(const float* p) { v = __ldg(p); }
  __device__ __forceinline__ void fma(float w, const Vec<1>& o) { v = fmaf(w, o.v, v); }
  __device__ __forceinline__ float get(int) const { return v; }
  __device__ __forceinline__ void set(int, float f) { v = f; }
  __device__ __forceinline__ void store(float* p) const { *p = v; }
};

template <int VEC>
__global__ void __launch_bounds__(256)
k_agg(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ dis,
      const float* __restrict__ ew, const float* __restrict__ selfw, float fill,
      const float* __restrict__ T, const float* __restrict__ bias, float* __restrict__ out,
      int64_t N, int H, int lpr, int act, float slope, const int32_t* __restrict__ eid,
      const float* __restrict__ mask) {
  const int rows_per_block = 256 / lpr;
  const int64_t i = (int64_t)blockIdx.x * rows_per_block + threadIdx.x / lpr;
  const int sub = threadIdx.x % lpr;
  if (i >= N) return;
  const int beg = rowptr[i], end = rowptr[i + 1];
  const float di = dis[i];
  const float sw = selfw ? selfw[i] : fill;
  const float self_norm = (di * sw) * di;
  for (int c = sub * VEC; c < H; c += lpr * VEC) {
    Vec<VEC> acc;
    acc.zero();
    int k = beg;
    for (; k + 4 <= end; k += 4) {  // 4 independent gathers in flight
      int j0 = col[k], j1 = col[k + 1], j2 = col[k + 2], j3 = col[k + 3];
      float w0 = dis[j0], w1 = dis[j1], w2 = dis[j2], w3 = dis[j3];
      if (ew) { w0 *= ew[k]; w1 *= ew[k + 1]; w2 *= ew[k + 2]; w3 *= ew[k + 3]; }
      if (mask) { w0 *= mask[eid[k]]; w1 *= mask[eid[k + 1]]; w2 *= mask[eid[k + 2]]; w3 *= mask[eid[k + 3]]; }
      Vec<VEC> t0, t1, t2, t3;
      t0.load(T + (int64_t)j0 * H + c); t1.load(T + (int64_t)j1 * H + c);
      t2.load(T + (int64_t)j2 * H + c); t3.load(T + (int64_t)j3 * H + c);
      acc.fma(w0 * di, t0); acc.fma(w1 * di, t1); acc.fma(w2 * di, t2); acc.fma(w3 * di, t3);
    }
    for (; k < end; ++k) {
      int j = col[k];
      float w = dis[j];
      if (ew) w *= ew[k];
      if (mask) w *= mask[eid[k]];
      Vec<VEC> t;
      t.load(T + (int64_t)j * H + c);
      acc.fma(w * di, t);
    }
    Vec<VEC> ts;
    ts.load(T + i * H + c);
    acc.fma(self_norm, ts);
#pragma unroll
    for (int q = 0; q < VEC; ++q) {
      float v = acc.get(q);
      if (bias) v += bias[c + q];
      if (act) v = lrelu(v, slope);
      acc.set(q, v);
    }
    acc.store(out + i * H + c);
  }
}

// Gradient w.r.t. the soft edge mask of one conv layer:  g_mask[e] (+)= norm_e <gA[dst_e], T[src_e]>
// (out_i = sum_e mask_e norm_e T_src(e) + ...).  A warp per destination row, lanes over the features, fixed
// shuffle tree; every edge owns exactly one CSR slot, so the accumulation over layers needs no atomics.
__global__ void __launch_bounds__(256)
k_edge_dot(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const int32_t* __restrict__ eid,
           const float* __restrict__ dis, const float* __restrict__ ew, const float* __restrict__ gA,
           const float* __restrict__ T, int64_t N, int H, int accumulate, float* __restrict__ g_mask) {
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  const int beg = rowptr[i], end = rowptr[i + 1];
  const float di = dis[i];
  for (int k = beg; k < end; ++k) {
    const int j = col[k];
    float s = 0.f;
    for (int c = lane; c < H; c += 32) s = fmaf(gA[i * H + c], __ldg(T + (int64_t)j * H + c), s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      float w = dis[j] * di;
      if (ew) w *= ew[k];
      const float v = w * s;
      const int e = eid[k];
      g_mask[e] = accumulate ? g_mask[e] + v : v;
    }
  }
}

static int lanes_per_row(int H, int vec) {
  int need = (H + vec - 1) / vec;
  int l = 1;
  while (l < need && l < 32) l <<= 1;
  return l;
}

// ------------------------------------------------------------------------------------------
// Wide-row aggregation (H >= 128, unweighted, unmasked): k_agg above spends ~400 instructions per 256-wide
// row (the neighbour list is walked once per 128-column pass, runtime switches inside the loops, scalar
// bias loads) and is issue-bound at ~40 % of the HBM rate.  Here a warp owns a whole row: the neighbour
// list is walked ONCE, each lane carries NV float4 accumulators (columns 4 lane + 128 v), the weights are
// formed once per neighbour, bias / LeakyReLU / store are vector operations.  Same weights, same order of
// the sums as k_agg: bit-identical rows.
// POOL: a block owns one graph (rows round-robin over its 8 warps) and also produces emb = [max || mean]
// of the finished rows (reference model/gcn.py:65-66) - the last layer's k_pool pass over [N, H] disappears,
// and in inference (out == NULL) so does the last layer's [N, H] store.
// ------------------------------------------------------------------------------------------
template <int NV, bool EXACT>
__device__ __forceinline__ void agg_wide_fma(float4 (&acc)[NV], float w, const float* __restrict__ p, int H, int lane) {
  float4 t[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) t[v] = (EXACT || 128 * v + 4 * lane < H) ? ldg4(p + 128 * v) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    acc[v].x = fmaf(w, t[v].x, acc[v].x); acc[v].y = fmaf(w, t[v].y, acc[v].y);
    acc[v].z = fmaf(w, t[v].z, acc[v].z); acc[v].w = fmaf(w, t[v].w, acc[v].w);
  }
}

// A warp finishes `rows` (1..R) consecutive rows starting at i0.  The rows' CSR slots are contiguous, so the
// row pointers, the first 32 column indices and their dis[] values are fetched by ONE coalesced load each and
// handed out by shuffles: three dependent memory latencies per tile instead of three per row.
template <int NV, int R, bool EXACT, bool POOL>
__device__ __forceinline__ void agg_wide_tile(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                              const float* __restrict__ dis, float fill, const float* __restrict__ T,
                                              int64_t i0, int rows, int H, int lane, const float4 (&b4)[NV], int act,
                                              float slope, float* __restrict__ out, float4 (&pmax)[NV], float4 (&psum)[NV]) {
  constexpr unsigned kFull = 0xffffffffu;
  int rp = 0;
  float drow = 0.f;
  if (lane <= rows) rp = __ldg(rowptr + i0 + lane);
  if (lane < rows) drow = __ldg(dis + i0 + lane);
  const int beg0 = __shfl_sync(kFull, rp, 0), end_all = __shfl_sync(kFull, rp, rows);
  int jl = 0;
  float wl = 0.f;
  if (beg0 + lane < end_all) { jl = __ldg(col + beg0 + lane); wl = __ldg(dis + jl); }
  const float* Tl = T + 4 * lane;
  const int HH = EXACT ? NV * 128 : H;  // a compile-time row pitch turns every row address into one IMAD.WIDE
#pragma unroll 1
  for (int r = 0; r < rows; ++r) {
    const int beg = __shfl_sync(kFull, rp, r), end = __shfl_sync(kFull, rp, r + 1);
    const float di = __shfl_sync(kFull, drow, r);
    float4 acc[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int cut = min(end, beg0 + 32);  // slots below it were preloaded (all of them for molecule-like rows)
    int k = beg;
#pragma unroll 2
    for (; k < cut; ++k) {
      const int j = __shfl_sync(kFull, jl, k - beg0);
      const float w = __shfl_sync(kFull, wl, k - beg0);
      agg_wide_fma<NV, EXACT>(acc, w * di, Tl + (int64_t)j * HH, H, lane);
    }
    for (; k < end; ++k) {
      const int j = __ldg(col + k);
      const float w = __ldg(dis + j);
      agg_wide_fma<NV, EXACT>(acc, w * di, Tl + (int64_t)j * HH, H, lane);
    }
    const int64_t i = i0 + r;
    agg_wide_fma<NV, EXACT>(acc, (di * fill) * di, Tl + i * HH, H, lane);
#pragma unroll
    for (int v = 0; v < NV; ++v)
      if (EXACT || 128 * v + 4 * lane < H) {
        acc[v].x += b4[v].x; acc[v].y += b4[v].y; acc[v].z += b4[v].z; acc[v].w += b4[v].w;
        if (act == 2) {  // 0 <= slope <= 1: lrelu(v) = max(v, slope v), same bits, one instruction less
          acc[v].x = fmaxf(acc[v].x, acc[v].x * slope); acc[v].y = fmaxf(acc[v].y, acc[v].y * slope);
          acc[v].z = fmaxf(acc[v].z, acc[v].z * slope); acc[v].w = fmaxf(acc[v].w, acc[v].w * slope);
        } else if (act) {
          acc[v].x = lrelu(acc[v].x, slope); acc[v].y = lrelu(acc[v].y, slope);
          acc[v].z = lrelu(acc[v].z, slope); acc[v].w = lrelu(acc[v].w, slope);
        }
        if (!POOL || out) *reinterpret_cast<float4*>(out + i * HH + 128 * v + 4 * lane) = acc[v];
        if (POOL) {
          pmax[v].x = fmaxf(pmax[v].x, acc[v].x); pmax[v].y = fmaxf(pmax[v].y, acc[v].y);
          pmax[v].z = fmaxf(pmax[v].z, acc[v].z); pmax[v].w = fmaxf(pmax[v].w, acc[v].w);
          psum[v].x += acc[v].x; psum[v].y += acc[v].y; psum[v].z += acc[v].z; psum[v].w += acc[v].w;
        }
      }
  }
}

constexpr int kAggTileRows = 8;   // rows per warp tile, row-parallel variant

template <int NV, bool EXACT, bool POOL>
__global__ void __launch_bounds__(256)
k_agg_wide(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ dis, float fill,
           const float* __restrict__ T, const float* __restrict__ bias, float* __restrict__ out, int64_t N, int H, int act,
           float slope, const int32_t* __restrict__ graph_ptr, float* __restrict__ emb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float4 b4[NV], pmax[NV], psum[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    b4[v] = (bias && 128 * v + 4 * lane < H) ? ldg4(bias + 128 * v + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    pmax[v] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    psum[v] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (!POOL) {
    const int64_t i0 = ((int64_t)blockIdx.x * 8 + warp) * kAggTileRows;
    if (i0 >= N) return;
    const int rows = (int)min((int64_t)kAggTileRows, N - i0);
    agg_wide_tile<NV, kAggTileRows, EXACT, false>(rowptr, col, dis, fill, T, i0, rows, H, lane, b4, act, slope, out, pmax, psum);
    return;
  }
  __shared__ float4 s_max[8][NV][32], s_sum[8][NV][32];
  const int64_t gid = blockIdx.x;
  const int n0 = graph_ptr[gid], n1 = graph_ptr[gid + 1];
  // every warp takes a contiguous share of ceil(n / 8) rows (balanced to one row), walked in tiles of <= 8 rows
  const int share = (n1 - n0 + 7) >> 3;
  const int lo = n0 + warp * share, hi = min(n1, lo + share);
  for (int i0 = lo; i0 < hi; i0 += kAggTileRows)
    agg_wide_tile<NV, kAggTileRows, EXACT, true>(rowptr, col, dis, fill, T, i0, min(kAggTileRows, hi - i0), H, lane, b4, act,
                                                 slope, out, pmax, psum);
#pragma unroll
  for (int v = 0; v < NV; ++v) { s_max[warp][v][lane] = pmax[v]; s_sum[warp][v][lane] = psum[v]; }
  __syncthreads();
  // thread t combines column t of every 128-column chunk over the 8 warps in warp order (= node order of the shares)
  const int n = n1 - n0;
  for (int c = threadIdx.x; c < NV * 128; c += 256) {
    const int v = c >> 7, cc = c & 127;
    if (128 * v + cc >= H) continue;
    float mx = -INFINITY, sm = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      mx = fmaxf(mx, reinterpret_cast<const float*>(&s_max[w][v][0])[cc]);
      sm += reinterpret_cast<const float*>(&s_sum[w][v][0])[cc];
    }
    emb[gid * 2 * H + 128 * v + cc] = n > 0 ? mx : 0.f;
    emb[gid * 2 * H + H + 128 * v + cc] = n > 0 ? sm / (float)n : 0.f;
  }
}

// The same tile scheme for narrow layers (H <= 64, H even - the reference's default embedding_dim = 64): a lane
// owns columns 2 lane, 2 lane + 1, so a 256-byte row of T is ONE coalesced 8-byte-per-lane access and all 32
// lanes work (k_agg_wide would idle half of them).  Row-parallel only; k_pool follows.
template <bool EXACT>
__global__ void __launch_bounds__(256)
k_agg_narrow(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ dis, float fill,
             const float* __restrict__ T, const float* __restrict__ bias, float* __restrict__ out, int64_t N, int H, int act,
             float slope) {
  constexpr unsigned kFull = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t i0 = ((int64_t)blockIdx.x * 8 + warp) * kAggTileRows;
  if (i0 >= N) return;
  const int rows = (int)min((int64_t)kAggTileRows, N - i0);
  const bool on = EXACT || 2 * lane < H;
  float2 b2 = make_float2(0.f, 0.f);
  if (bias && on) b2 = __ldg(reinterpret_cast<const float2*>(bias + 2 * lane));
  int rp = 0;
  float drow = 0.f;
  if (lane <= rows) rp = __ldg(rowptr + i0 + lane);
  if (lane < rows) drow = __ldg(dis + i0 + lane);
  const int beg0 = __shfl_sync(kFull, rp, 0), end_all = __shfl_sync(kFull, rp, rows);
  int jl = 0;
  float wl = 0.f;
  if (beg0 + lane < end_all) { jl = __ldg(col + beg0 + lane); wl = __ldg(dis + jl); }
  const float* Tl = T + 2 * lane;
  const int HH = EXACT ? 64 : H;  // compile-time row pitch for the default width
#pragma unroll 1
  for (int r = 0; r < rows; ++r) {
    const int beg = __shfl_sync(kFull, rp, r), end = __shfl_sync(kFull, rp, r + 1);
    const float di = __shfl_sync(kFull, drow, r);
    float2 acc = make_float2(0.f, 0.f);
    const int cut = min(end, beg0 + 32);  // slots below it were preloaded
    int k = beg;
#pragma unroll 2
    for (; k < cut; ++k) {
      const int j = __shfl_sync(kFull, jl, k - beg0);
      const float w = __shfl_sync(kFull, wl, k - beg0) * di;
      if (on) {
        const float2 t = __ldg(reinterpret_cast<const float2*>(Tl + (int64_t)j * HH));
        acc.x = fmaf(w, t.x, acc.x); acc.y = fmaf(w, t.y, acc.y);
      }
    }
    for (; k < end; ++k) {
      const int j = __ldg(col + k);
      const float w = __ldg(dis + j) * di;
      if (on) {
        const float2 t = __ldg(reinterpret_cast<const float2*>(Tl + (int64_t)j * HH));
        acc.x = fmaf(w, t.x, acc.x); acc.y = fmaf(w, t.y, acc.y);
      }
    }
    const int64_t i = i0 + r;
    if (on) {
      const float ws = (di * fill) * di;
      const float2 t = __ldg(reinterpret_cast<const float2*>(Tl + i * HH));
      acc.x = fmaf(ws, t.x, acc.x) + b2.x; acc.y = fmaf(ws, t.y, acc.y) + b2.y;
      if (act == 2) { acc.x = fmaxf(acc.x, acc.x * slope); acc.y = fmaxf(acc.y, acc.y * slope); }
      else if (act) { acc.x = lrelu(acc.x, slope); acc.y = lrelu(acc.y, slope); }
      *reinterpret_cast<float2*>(out + i * HH + 2 * lane) = acc;
    }
  }
}

// HCG_CFG_AGG_PATH: 0 = tile kernels when H >= 128 (wide) or 32 <= H <= 64 (narrow), 1 = k_agg / k_pool only,
// 2 = tile kernels whenever their constraints hold
enum { AGG_ROWS = 0, AGG_WIDE = 1, AGG_NARROW = 2 };
static int agg_kind(const HcgGraph* g, int H, const float* T, const float* out) {
  const int ap = config_get(HCG_CFG_AGG_PATH);
  if (ap == 1 || g->edge_mask != nullptr) return AGG_ROWS;
  const uintptr_t al = reinterpret_cast<uintptr_t>(T) | reinterpret_cast<uintptr_t>(out);
  if (H <= 64 && (H & 1) == 0 && al % 8 == 0 && (ap == 2 || H >= 32)) return AGG_NARROW;
  if (H > 64 && (H & 3) == 0 && H <= 512 && al % 16 == 0 && (ap == 2 || H >= 128)) return AGG_WIDE;
  return AGG_ROWS;
}

static int launch_agg_narrow(const int32_t* rowptr, const int32_t* col, const float* dis, float fill, const float* T,
                             const float* bias, float* out, int64_t N, int H, int act, float slope, cudaStream_t st) {
  if (N == 0) return HCG_OK;
  if (act && slope >= 0.f && slope <= 1.f) act = 2;
  const unsigned grid = (unsigned)cdiv(N, 8 * kAggTileRows);
  if (H == 64) k_agg_narrow<true><<<grid, 256, 0, st>>>(rowptr, col, dis, fill, T, bias, out, N, H, act, slope);
  else k_agg_narrow<false><<<grid, 256, 0, st>>>(rowptr, col, dis, fill, T, bias, out, N, H, act, slope);
  HCG_LAUNCH_OK(st, "k_agg_narrow");
  return HCG_OK;
}

// out may be NULL only with emb (pooling variant: grid = graphs); emb == NULL: grid = rows / 8
static int launch_agg_wide(const int32_t* rowptr, const int32_t* col, const float* dis, float fill, const float* T,
                           const float* bias, float* out, int64_t N, int H, int act, float slope,
                           const int32_t* graph_ptr, int64_t B, float* emb, cudaStream_t st) {
  if (N == 0 || (emb && B == 0)) return HCG_OK;
  if (act && slope >= 0.f && slope <= 1.f) act = 2;
  const int nv = (H + 127) / 128;
  const unsigned grid = (unsigned)(emb ? B : cdiv(N, 8 * kAggTileRows));
  const bool exact = H % 128 == 0;
#define HCG_AGG_WIDE2(NVV, EX)                                                                                          \
  do {                                                                                                                 \
    if (emb) k_agg_wide<NVV, EX, true><<<grid, 256, 0, st>>>(rowptr, col, dis, fill, T, bias, out, N, H, act, slope, graph_ptr, emb); \
    else k_agg_wide<NVV, EX, false><<<grid, 256, 0, st>>>(rowptr, col, dis, fill, T, bias, out, N, H, act, slope, graph_ptr, emb);    \
  } while (0)
#define HCG_AGG_WIDE(NVV)                                                                                              \
  do {                                                                                                                 \
    if (exact) HCG_AGG_WIDE2(NVV, true); else HCG_AGG_WIDE2(NVV, false);                                               \
  } while (0)
  switch (nv) {
    case 1: HCG_AGG_WIDE(1); break;
    case 2: HCG_AGG_WIDE(2); break;
    case 3: HCG_AGG_WIDE(3); break;
    default: HCG_AGG_WIDE(4); break;
  }
#undef HCG_AGG_WIDE2
#undef HCG_AGG_WIDE
  if (emb) HCG_LAUNCH_OK(st, "k_agg_wide_pool");
  else HCG_LAUNCH_OK(st, "k_agg_wide");
  return HCG_OK;
}

static int launch_agg(const int32_t* rowptr, const int32_t* col, const float* dis, const float* ew,
                      const float* selfw, float fill, const float* T, const float* bias, float* out,
                      int64_t N, int H, int act, float slope, cudaStream_t st, const int32_t* eid = nullptr,
                      const float* mask = nullptr) {
  if (N == 0) return HCG_OK;
  const bool vec4 = (H % 4 == 0) && ((reinterpret_cast<uintptr_t>(T) | reinterpret_cast<uintptr_t>(out)) % 16 == 0);
  if (vec4) {
    int lpr = lanes_per_row(H, 4);
    k_agg<4><<<(unsigned)cdiv(N, 256 / lpr), 256, 0, st>>>(rowptr, col, dis, ew, selfw, fill, T, bias, out, N, H, lpr, act, slope, eid, mask);
  } else {
    int lpr = lanes_per_row(H, 1);
    k_agg<1><<<(unsigned)cdiv(N, 256 / lpr), 256, 0, st>>>(rowptr, col, dis, ew, selfw, fill, T, bias, out, N, H, lpr, act, slope, eid, mask);
  }
  HCG_LAUNCH_OK(st, "k_agg");
  return HCG_OK;
}

// ==========================================================================================
// Segmented pooling: emb[g] = [ max_i h[i] ‖ mean_i h[i] ] over the nodes of graph g
// (a lane group per graph, nodes walked in order: the mean's sum order is the node order).
// ==========================================================================================
__global__ void __launch_bounds__(256)
k_pool(const float* __restrict__ h, const int32_t* __restrict__ graph_ptr, int64_t B, int H, int lpr,
       float* __restrict__ emb) {
  const int gpb = 256 / lpr;
  const int64_t g = (int64_t)blockIdx.x * gpb + threadIdx.x / lpr;
  const int sub = threadIdx.x % lpr;
  if (g >= B) return;
  const int beg = graph_ptr[g], end = graph_ptr[g + 1];
  const int n = end - beg;
  for (int c = sub; c < H; c += lpr) {
    float mx = -INFINITY, sum = 0.f;
    int i = beg;
    for (; i + 4 <= end; i += 4) {
      float v0 = __ldg(h + (int64_t)i * H + c), v1 = __ldg(h + (int64_t)(i + 1) * H + c);
      float v2 = __ldg(h + (int64_t)(i + 2) * H + c), v3 = __ldg(h + (int64_t)(i + 3) * H + c);
      mx = fmaxf(fmaxf(mx, v0), fmaxf(v1, fmaxf(v2, v3)));
      sum += v0; sum += v1; sum += v2; sum += v3;
    }
    for (; i < end; ++i) {
      float v = __ldg(h + (int64_t)i * H + c);
      mx = fmaxf(mx, v);
      sum += v;
    }
    // an empty graph cannot come out of PyG collate; define it as zeros rather than -inf / NaN
    emb[g * 2 * H + c] = n > 0 ? mx : 0.f;
    emb[g * 2 * H + H + c] = n > 0 ? sum / (float)n : 0.f;
  }
}

// ==========================================================================================
// Fused two-layer readout head (the default network: emb[B,128] -> 64 -> C), one launch:
//   hid = lrelu(emb . W1^T + b1)   (optionally saved for the backward pass)
//   out = hid . W2^T + b2
// A block owns RO_FG = 64 graphs and keeps the whole problem in shared memory - its embedding rows [64][128], W1
// [64][132] and, afterwards, the hidden tile: one round of coalesced 16-byte loads and ONE barrier before the product
// (the first version walked k in four load -> barrier -> FMA -> barrier rounds of 32 and was latency-bound at 27 us per
// 16384 graphs).  16 x 16 threads, each a 4 (graphs 4 ty ..) x 4 (hidden units tx + 16 b) micro-tile fed by 16-byte
// shared-memory reads along k; per output the k loop is ONE fmaf chain in ascending order, like k_gemm and the head of
// k_readout2_bwd<true>, so all of them produce the same bits.
// ==========================================================================================
constexpr int RO_G = 64, RO_IN = 128, RO_HID = 64, RO_KC = 32;
constexpr int RO_FG = 64;            // graphs per block of the forward head
constexpr int RO_FW1P = RO_IN + 4;   // pitch of its W1 tile: 16-byte reads of rows tx + 16 b hit eight different bank groups
static inline size_t ro2_fwd_smem_bytes() { return sizeof(float) * (size_t)(RO_HID * RO_FW1P + RO_FG * RO_IN + RO_FG * (RO_HID + 1)); }

__device__ __forceinline__ void
readout2_block(const float* __restrict__ emb, const float* __restrict__ W1, const float* __restrict__ b1,
               const float* __restrict__ W2, const float* __restrict__ b2, float* __restrict__ hid_save,
               float* __restrict__ out, int64_t B, int C, float slope) {
  extern __shared__ __align__(16) float rof_smem[];
  float* W1s = rof_smem;                   // [64][132]  (unit j, input k)
  float* Es = W1s + RO_HID * RO_FW1P;      // [64][128]  (graph g, input k)
  float* Hs = Es + RO_FG * RO_IN;          // [64][65]   hidden tile
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t g0 = (int64_t)blockIdx.x * RO_FG;
  {   // all sixteen 16-byte loads of a thread in flight at once: one memory latency, not sixteen
    constexpr int NL = RO_HID * RO_IN / 4 / 256;   // = RO_FG * RO_IN / 4 / 256 = 8
    static_assert(RO_FG == RO_HID, "one load count for both tiles");
    float4 wv[NL], ev[NL];
#pragma unroll
    for (int u = 0; u < NL; ++u) wv[u] = ldg4(W1 + 4 * (tid + 256 * u));
#pragma unroll
    for (int u = 0; u < NL; ++u) {
      const int idx = tid + 256 * u;
      ev[u] = g0 + idx / (RO_IN / 4) < B ? ldg4(emb + g0 * RO_IN + 4 * (int64_t)idx) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < NL; ++u) {
      const int idx = tid + 256 * u, j = idx / (RO_IN / 4), k4 = idx - j * (RO_IN / 4);
      *reinterpret_cast<float4*>(W1s + j * RO_FW1P + 4 * k4) = wv[u];
    }
#pragma unroll
    for (int u = 0; u < NL; ++u) reinterpret_cast<float4*>(Es)[tid + 256 * u] = ev[u];
  }
  __syncthreads();
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 2
  for (int k = 0; k < RO_IN; k += 4) {
    float4 e[4], w[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) e[a] = *reinterpret_cast<const float4*>(Es + (4 * ty + a) * RO_IN + k);
#pragma unroll
    for (int b = 0; b < 4; ++b) w[b] = *reinterpret_cast<const float4*>(W1s + (tx + 16 * b) * RO_FW1P + k);
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        float v = acc[a][b];
        v = fmaf(e[a].x, w[b].x, v); v = fmaf(e[a].y, w[b].y, v); v = fmaf(e[a].z, w[b].z, v); v = fmaf(e[a].w, w[b].w, v);
        acc[a][b] = v;
      }
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int g = 4 * ty + a, j = tx + 16 * b;
      const float v = lrelu(acc[a][b] + __ldg(b1 + j), slope);
      Hs[g * (RO_HID + 1) + j] = v;
      if (hid_save && g0 + g < B) hid_save[(g0 + g) * RO_HID + j] = v;
    }
  __syncthreads();
  for (int idx = tid; idx < RO_FG * C; idx += 256) {
    const int g = idx % RO_FG, c = idx / RO_FG;
    if (g0 + g >= B) continue;
    float s = 0.f;
#pragma unroll 8
    for (int o = 0; o < RO_HID; ++o) s = fmaf(Hs[g * (RO_HID + 1) + o], __ldg(W2 + c * RO_HID + o), s);
    out[(g0 + g) * C + c] = s + __ldg(b2 + c);
  }
}

__global__ void __launch_bounds__(256)
k_readout2(const float* __restrict__ emb, const float* __restrict__ W1, const float* __restrict__ b1,
           const float* __restrict__ W2, const float* __restrict__ b2, float* __restrict__ hid_save,
           float* __restrict__ out, int64_t B, int C, float slope) {
  readout2_block(emb, W1, b1, W2, b2, hid_save, out, B, C, slope);
}

// the same head for the M models of an ensemble in one launch: blockIdx.y = model; stacked weights
// (W1 [M,64,128], b1 [M,64], W2 [M,C,64], b2 [M,C]), emb [M,B,128], out [M,B,C]
__global__ void __launch_bounds__(256)
k_readout2_ens(const float* __restrict__ emb, const float* __restrict__ W1, const float* __restrict__ b1,
               const float* __restrict__ W2, const float* __restrict__ b2, float* __restrict__ out, int64_t B, int C, float slope) {
  const int64_t m = blockIdx.y;
  readout2_block(emb + m * B * RO_IN, W1 + m * RO_HID * RO_IN, b1 + m * RO_HID, W2 + m * C * RO_HID, b2 + m * C, nullptr,
                 out + m * B * C, B, C, slope);
}

// ==========================================================================================
// Backward of the default two-layer readout head in one launch (+ one reduce):
//   gpre1 = (g_out . W2) * lrelu'(hid)                      [B,64]
//   g_emb = gpre1 . W1 (+ extra gradient into emb)          [B,128]  -> feeds the pooling backward
//   gW1 = gpre1^T emb, gb1 = colsum(gpre1), gW2 = g_out^T hid, gb2 = colsum(g_out)
// A block owns 64 graphs and writes its partial parameter gradients; k_readout2_bwd_reduce sums the
// blocks in order (deterministic).  Replaces 11 launches of the general chain.
// ==========================================================================================
constexpr int RB_P = 68;  // pitch of the 64-wide shared tiles

__host__ __device__ inline int ro2_partial_elems(int C) { return RO_HID * RO_IN + RO_HID + C * RO_HID + C; }
__host__ __device__ inline int ro2_partial_stride(int C) { return (ro2_partial_elems(C) + 3) & ~3; }  // 16-byte aligned blocks
constexpr int RO_W1P = RO_IN + 4;  // pitch of the W1 tile: the head's forward product reads it by columns (bank = 4 j + k)
static inline size_t ro2_bwd_smem_bytes(int C, bool train = false) {
  return sizeof(float) * (size_t)(RO_HID * RO_W1P + RO_G * RO_IN + 2 * RO_G * RB_P + RO_G * C + (train ? RO_G * (RO_HID + 1) + 256 : 0));
}

// TRAIN = true: the whole head of a training step in one kernel - its forward (hid = lrelu(emb W1^T + b1),
// out = hid W2^T + b2, the arithmetic and order of k_readout2), the loss seed (g_out = 2 (out - y), the block's share of
// sum (out - y)^2 -> sse_partial[block]; hcg_rmse_seed) and the backward below, with hid and g_out kept in shared
// memory (`hid`, `g_out` are not read).
template <bool TRAIN>
__global__ void __launch_bounds__(256)
k_readout2_bwd(const float* __restrict__ emb, const float* __restrict__ hid, const float* __restrict__ g_out,
               const float* __restrict__ g_emb_extra, const float* __restrict__ W1, const float* __restrict__ W2,
               int64_t B, int C, float slope, float* __restrict__ g_emb, float* __restrict__ partial,
               const float* __restrict__ b1, const float* __restrict__ b2, const float* __restrict__ y,
               float* __restrict__ sse_partial) {
  extern __shared__ __align__(16) float ro_smem[];
  float* W1s = ro_smem;                    // [64][132]  (unit j, input k)
  float* Es = W1s + RO_HID * RO_W1P;       // [64][128]  (graph g, input k)
  float* G = Es + RO_G * RO_IN;            // [64][68]   gpre1[g][j]
  float* GT = G + RO_G * RB_P;             // [64][68]   gpre1[j][g]
  float* Go = GT + RO_G * RB_P;            // [64][C]    g_out tile
  float* Hs = Go + RO_G * C;               // TRAIN: [64][65] hid tile, then 256 floats of reduction scratch
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t g0 = (int64_t)blockIdx.x * RO_G;
  float* part = partial + (int64_t)blockIdx.x * ro2_partial_stride(C);

  {   // all sixteen 16-byte loads of a thread in flight at once: one memory latency, not sixteen
    constexpr int NL = RO_HID * RO_IN / 4 / 256;   // = RO_G * RO_IN / 4 / 256 = 8
    static_assert(RO_G == RO_HID, "one load count for both tiles");
    float4 wv[NL], ev[NL];
#pragma unroll
    for (int u = 0; u < NL; ++u) wv[u] = ldg4(W1 + 4 * (tid + 256 * u));
#pragma unroll
    for (int u = 0; u < NL; ++u) {
      const int idx = tid + 256 * u;
      ev[u] = g0 + idx / (RO_IN / 4) < B ? ldg4(emb + g0 * RO_IN + 4 * (int64_t)idx) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < NL; ++u) {
      const int idx = tid + 256 * u, j = idx / (RO_IN / 4), k4 = idx - j * (RO_IN / 4);
      *reinterpret_cast<float4*>(W1s + j * RO_W1P + 4 * k4) = wv[u];
    }
#pragma unroll
    for (int u = 0; u < NL; ++u) reinterpret_cast<float4*>(Es)[tid + 256 * u] = ev[u];
  }
  if (!TRAIN)
    for (int idx = tid; idx < RO_G * C; idx += 256) Go[idx] = g0 + idx / C < B ? g_out[g0 * C + idx] : 0.f;
  __syncthreads();
  if (TRAIN) {
    // forward head: thread = graphs 4 ty .. + 3, units tx + 16 b; k ascending with one fmaf chain per output, as k_readout2
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 2
    for (int k = 0; k < RO_IN; k += 4) {   // 16-byte reads along k; per output still one chain in ascending k
      float4 e[4], w[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) e[a] = *reinterpret_cast<const float4*>(Es + (4 * ty + a) * RO_IN + k);
#pragma unroll
      for (int b = 0; b < 4; ++b) w[b] = *reinterpret_cast<const float4*>(W1s + (tx + 16 * b) * RO_W1P + k);
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          float v = acc[a][b];
          v = fmaf(e[a].x, w[b].x, v); v = fmaf(e[a].y, w[b].y, v); v = fmaf(e[a].z, w[b].z, v); v = fmaf(e[a].w, w[b].w, v);
          acc[a][b] = v;
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int j = tx + 16 * b;
        Hs[(4 * ty + a) * (RO_HID + 1) + j] = lrelu(acc[a][b] + __ldg(b1 + j), slope);
      }
    __syncthreads();
    float sq = 0.f;
    for (int idx = tid; idx < RO_G * C; idx += 256) {
      const int g = idx / C, c = idx - g * C;
      float o = 0.f;
#pragma unroll 8
      for (int j = 0; j < RO_HID; ++j) o = fmaf(Hs[g * (RO_HID + 1) + j], __ldg(W2 + c * RO_HID + j), o);
      o += __ldg(b2 + c);
      const float d = g0 + g < B ? o - __ldg(y + g0 + g) : 0.f;
      Go[idx] = 2.f * d;
      sq = fmaf(d, d, sq);
    }
    float* red = Hs + RO_G * (RO_HID + 1);
    red[tid] = sq;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
      if (tid < w) red[tid] += red[tid + w];
      __syncthreads();
    }
    if (tid == 0) sse_partial[blockIdx.x] = red[0];
  }

  // gpre1 (both orientations)
  for (int idx = tid; idx < RO_G * RO_HID; idx += 256) {
    const int g = idx >> 6, j = idx & 63;
    float s = 0.f;
    for (int c = 0; c < C; ++c) s = fmaf(Go[g * C + c], __ldg(W2 + c * RO_HID + j), s);
    const float h = TRAIN ? Hs[g * (RO_HID + 1) + j] : (g0 + g < B ? hid[(g0 + g) * RO_HID + j] : 0.f);
    s *= lrelu_grad_from_out(h, slope);
    G[g * RB_P + j] = s;
    GT[j * RB_P + g] = s;
  }
  __syncthreads();

  // small gradients: gb1[j], gW2[c][j], gb2[c]
  if (tid < RO_HID) {
    float s = 0.f;
    for (int g = 0; g < RO_G; ++g) s += G[g * RB_P + tid];
    part[RO_HID * RO_IN + tid] = s;
    for (int c = 0; c < C; ++c) {
      float w = 0.f;
      for (int g = 0; g < RO_G; ++g)
        if (g0 + g < B) w = fmaf(Go[g * C + c], TRAIN ? Hs[g * (RO_HID + 1) + tid] : hid[(g0 + g) * RO_HID + tid], w);
      part[RO_HID * RO_IN + RO_HID + c * RO_HID + tid] = w;
    }
  } else if (tid < RO_HID + C) {
    const int c = tid - RO_HID;
    float s = 0.f;
    for (int g = 0; g < RO_G; ++g) s += Go[g * C + c];
    part[RO_HID * RO_IN + RO_HID + C * RO_HID + c] = s;
  }

  // g_emb[g][k] = sum_j gpre1[g][j] W1[j][k]   ; thread: graphs 4 ty .. +4, inputs 8 tx .. +8
  {
    float acc[4][8];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
#pragma unroll 4
    for (int j = 0; j < RO_HID; ++j) {
      const float4 a4 = *reinterpret_cast<const float4*>(GT + j * RB_P + 4 * ty);
      const float4 w0 = *reinterpret_cast<const float4*>(W1s + j * RO_W1P + 8 * tx);
      const float4 w1 = *reinterpret_cast<const float4*>(W1s + j * RO_W1P + 8 * tx + 4);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w};
      const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = fmaf(av[a], wv[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int64_t g = g0 + 4 * ty + a;
      if (g >= B) continue;
      float* dst = g_emb + g * RO_IN + 8 * tx;
      float4 o0 = make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]);
      float4 o1 = make_float4(acc[a][4], acc[a][5], acc[a][6], acc[a][7]);
      if (g_emb_extra) {
        const float4 e0 = ldg4(g_emb_extra + g * RO_IN + 8 * tx), e1 = ldg4(g_emb_extra + g * RO_IN + 8 * tx + 4);
        o0.x += e0.x; o0.y += e0.y; o0.z += e0.z; o0.w += e0.w;
        o1.x += e1.x; o1.y += e1.y; o1.z += e1.z; o1.w += e1.w;
      }
      *reinterpret_cast<float4*>(dst) = o0;
      *reinterpret_cast<float4*>(dst + 4) = o1;
    }
  }
  // gW1[j][k] = sum_g gpre1[g][j] emb[g][k]     ; thread: units 4 ty .. +4, inputs 8 tx .. +8
  {
    float acc[4][8];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
#pragma unroll 4
    for (int g = 0; g < RO_G; ++g) {
      const float4 a4 = *reinterpret_cast<const float4*>(G + g * RB_P + 4 * ty);
      const float4 e0 = *reinterpret_cast<const float4*>(Es + g * RO_IN + 8 * tx);
      const float4 e1 = *reinterpret_cast<const float4*>(Es + g * RO_IN + 8 * tx + 4);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w};
      const float ev[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = fmaf(av[a], ev[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      float* dst = part + (4 * ty + a) * RO_IN + 8 * tx;
      *reinterpret_cast<float4*>(dst) = make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[a][4], acc[a][5], acc[a][6], acc[a][7]);
    }
  }
}

__global__ void __launch_bounds__(256)
k_readout2_bwd_reduce(const float* __restrict__ partial, int blocks, int C, float* __restrict__ gW1, float* __restrict__ gb1,
                      float* __restrict__ gW2, float* __restrict__ gb2, const float* __restrict__ sse_partial = nullptr,
                      float* __restrict__ tail = nullptr, float count = 0.f) {
  const int pe = ro2_partial_stride(C);
  const int e = blockIdx.x * 32 + (threadIdx.x & 31);
  if (sse_partial != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {   // the training head's loss seed: [sum of squared errors, count], blocks in order
    float s = 0.f;
    for (int k = 0; k < blocks; ++k) s += sse_partial[k];
    tail[0] = s;
    tail[1] = count;
  }
  __shared__ float s_red[256];
  const float v = det_colsum32(partial, blocks, pe, e, e < ro2_partial_elems(C), s_red);
  if (e >= ro2_partial_elems(C) || threadIdx.x >= 32) return;
  const int n1 = RO_HID * RO_IN;
  if (e < n1) gW1[e] = v;
  else if (e < n1 + RO_HID) gb1[e - n1] = v;
  else if (e < n1 + RO_HID + C * RO_HID) gW2[e - n1 - RO_HID] = v;
  else gb2[e - n1 - RO_HID - C * RO_HID] = v;
}

// Backward of pooling fused with the LeakyReLU derivative of the last conv layer:
//   gA[i][c] = ( g_mean[g][c] / n_g + [h[i][c] == max] · g_max[g][c] / ties ) · lrelu'(h[i][c])
// Ties share the gradient evenly, as torch's scatter_reduce(amax) backward does.
__global__ void __launch_bounds__(256)
k_pool_bwd(const float* __restrict__ h, const float* __restrict__ emb, const float* __restrict__ g_emb,
           const int32_t* __restrict__ graph_ptr, int64_t B, int H, int lpr, float slope,
           float* __restrict__ gA) {
  const int gpb = 256 / lpr;
  const int64_t g = (int64_t)blockIdx.x * gpb + threadIdx.x / lpr;
  const int sub = threadIdx.x % lpr;
  if (g >= B) return;
  const int beg = graph_ptr[g], end = graph_ptr[g + 1];
  const int n = end - beg;
  if (n <= 0) return;
  for (int c = sub; c < H; c += lpr) {
    const float mx = emb[g * 2 * H + c];
    const float gmax = g_emb[g * 2 * H + c];
    const float gmean = g_emb[g * 2 * H + H + c] / (float)n;
    int ties = 0;
    for (int i = beg; i < end; ++i) ties += (__ldg(h + (int64_t)i * H + c) == mx) ? 1 : 0;
    const float gm = ties > 0 ? gmax / (float)ties : 0.f;
    for (int i = beg; i < end; ++i) {
      float v = __ldg(h + (int64_t)i * H + c);
      float gg = gmean + (v == mx ? gm : 0.f);
      gA[(int64_t)i * H + c] = gg * lrelu_grad_from_out(v, slope);
    }
  }
}

// ==========================================================================================
// Deterministic reduction over rows:  out[P,Q] = sum_n X[n][p] · Y[n][q]   (weight gradients)
// stage 1: each block owns a contiguous chunk of rows and one 64x64 tile of [P,Q] -> partials
// stage 2: partials summed in chunk order.
// ==========================================================================================
constexpr int XR = 32;  // rows staged per step

__global__ void __launch_bounds__(256)
k_xty_partial(const float* __restrict__ X, const float* __restrict__ Y, int64_t Nrows, int P, int Q,
              int64_t rows_per_chunk, float* __restrict__ partial) {
  __shared__ __align__(16) float Xs[XR][64 + 4];
  __shared__ __align__(16) float Ys[XR][64 + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int p0 = blockIdx.y * 64, q0 = blockIdx.z * 64;
  const int64_t r_beg = (int64_t)blockIdx.x * rows_per_chunk;
  const int64_t r_end = min(Nrows, r_beg + rows_per_chunk);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int64_t r0 = r_beg; r0 < r_end; r0 += XR) {
    for (int idx = tid; idx < XR * 64; idx += 256) {
      int r = idx >> 6, c = idx & 63;
      int64_t row = r0 + r;
      Xs[r][c] = (row < r_end && p0 + c < P) ? X[row * P + p0 + c] : 0.f;
      Ys[r][c] = (row < r_end && q0 + c < Q) ? Y[row * Q + q0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int r = 0; r < XR; ++r) {
      float4 a = *reinterpret_cast<const float4*>(&Xs[r][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Ys[r][tx * 4]);
      float aa[4] = {a.x, a.y, a.z, a.w}, bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* dst = partial + (int64_t)blockIdx.x * P * Q;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int p = p0 + ty * 4 + i;
    if (p >= P) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int q = q0 + tx * 4 + j;
      if (q < Q) dst[(int64_t)p * Q + q] = acc[i][j];
    }
  }
}

// The same contract on the tensor cores (warp-level mma.sync m16n8k8, TF32 with the 3-term split: fp32-grade
// accuracy, see hcg_mma.cuh).  The reduction dimension is the ROW index, which is the row of the staged planes:
// fragments come straight out of shared memory.  32 rows are staged per step, split into TF32 hi / lo parts
// by the staging thread (once per element, not once per consuming warp) into four XOR-swizzled [32][64] planes
// (32 KB, several CTAs per SM overlap staging and MMAs): the inner loop is LDS + MMA.  Each slab's product
// starts from a zero accumulator and is added to the thread's running fp32 sums.  At P = Q = 256 (stress configuration) the fp32
// FMA kernel above is 63 % of the training step.
constexpr int XM = 32;

__device__ __forceinline__ void split4(const float4 v, float4& hi, float4& lo) {
  uint32_t h, l;
  split_tf32(v.x, h, l); hi.x = __uint_as_float(h); lo.x = __uint_as_float(l);
  split_tf32(v.y, h, l); hi.y = __uint_as_float(h); lo.y = __uint_as_float(l);
  split_tf32(v.z, h, l); hi.z = __uint_as_float(h); lo.z = __uint_as_float(l);
  split_tf32(v.w, h, l); hi.w = __uint_as_float(h); lo.w = __uint_as_float(l);
}

__global__ void __launch_bounds__(256)
k_xty_mma(const float* __restrict__ X, const float* __restrict__ Y, int64_t Nrows, int P, int Q,
          int64_t rows_per_chunk, float* __restrict__ partial) {
  __shared__ __align__(16) float Xh[XM * 64], Xl[XM * 64], Yh[XM * 64], Yl[XM * 64];  // TF32 hi / lo planes
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int p0 = blockIdx.y * 64, q0 = blockIdx.z * 64;
  const int64_t r_beg = (int64_t)blockIdx.x * rows_per_chunk;
  const int64_t r_end = min(Nrows, r_beg + rows_per_chunk);
  const bool vecX = (P & 3) == 0, vecY = (Q & 3) == 0;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // this thread stages float4 (r, c4) and (r + 16, c4) of both planes
  const int sr = tid >> 4, c4 = (tid & 15) * 4;
  const bool xin = vecX && p0 + c4 + 3 < P, yin = vecY && q0 + c4 + 3 < Q;

  for (int64_t r0 = r_beg; r0 < r_end; r0 += XM) {
    const int rows = (int)min((int64_t)XM, r_end - r0);
    const int npad = (rows + 7) & ~7;
    float4 xv[2], yv[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = sr + 16 * h;
      xv[h] = make_float4(0.f, 0.f, 0.f, 0.f);
      yv[h] = xv[h];
      if (r < rows) {
        const float* xp = X + (r0 + r) * P + p0 + c4;
        const float* yp = Y + (r0 + r) * Q + q0 + c4;
        if (xin) xv[h] = ldg4(xp);
        else {
          if (p0 + c4 < P) xv[h].x = __ldg(xp);
          if (p0 + c4 + 1 < P) xv[h].y = __ldg(xp + 1);
          if (p0 + c4 + 2 < P) xv[h].z = __ldg(xp + 2);
          if (p0 + c4 + 3 < P) xv[h].w = __ldg(xp + 3);
        }
        if (yin) yv[h] = ldg4(yp);
        else {
          if (q0 + c4 < Q) yv[h].x = __ldg(yp);
          if (q0 + c4 + 1 < Q) yv[h].y = __ldg(yp + 1);
          if (q0 + c4 + 2 < Q) yv[h].z = __ldg(yp + 2);
          if (q0 + c4 + 3 < Q) yv[h].w = __ldg(yp + 3);
        }
      }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = sr + 16 * h;
      if (r < npad) {
        const int o = r * 64 + (c4 ^ swz(r));
        float4 hi, lo;
        split4(xv[h], hi, lo);
        *reinterpret_cast<float4*>(Xh + o) = hi;
        *reinterpret_cast<float4*>(Xl + o) = lo;
        split4(yv[h], hi, lo);
        *reinterpret_cast<float4*>(Yh + o) = hi;
        *reinterpret_cast<float4*>(Yl + o) = lo;
      }
    }
    __syncthreads();
    xty_tiles_presplit<4>(Xh, Xl, Yh, Yl, npad, warp & 3, (warp >> 2) * 4, acc);
    __syncthreads();
  }
  // fragment of tile j: rows 16 mt + g8 (+8), columns 8 (nb + j) + 2 t4 (+1)
  float* dst = partial + (int64_t)blockIdx.x * P * Q;
  const int g8 = lane >> 2, t4 = lane & 3;
  const int pr = p0 + 16 * (warp & 3) + g8;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int qc = q0 + 8 * ((warp >> 2) * 4 + j) + 2 * t4;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int p = pr + 8 * h;
      if (p >= P) continue;
      if (qc < Q) dst[(int64_t)p * Q + qc] = acc[j][2 * h];
      if (qc + 1 < Q) dst[(int64_t)p * Q + qc + 1] = acc[j][2 * h + 1];
    }
  }
}

// column sums of X[n][p] over a chunk of rows (bias gradients)
__global__ void __launch_bounds__(256)
k_colsum_partial(const float* __restrict__ X, int64_t Nrows, int P, int64_t rows_per_chunk,
                 float* __restrict__ partial) {
  const int64_t r_beg = (int64_t)blockIdx.x * rows_per_chunk;
  const int64_t r_end = min(Nrows, r_beg + rows_per_chunk);
  for (int p = threadIdx.x; p < P; p += blockDim.x) {
    float s = 0.f;
    for (int64_t r = r_beg; r < r_end; ++r) s += X[r * P + p];
    partial[(int64_t)blockIdx.x * P + p] = s;
  }
}

__global__ void __launch_bounds__(256)
k_reduce_partials(const float* __restrict__ partial, int chunks, int64_t elems, float* __restrict__ out) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= elems) return;
  float s = 0.f;
  for (int c = 0; c < chunks; ++c) s += partial[(int64_t)c * elems + e];
  out[e] = s;
}

constexpr int kMaxChunks = kNumSMs * 4;

static int64_t pick_chunks(int64_t rows, int tiles) {
  int64_t want = kMaxChunks / (tiles > 0 ? tiles : 1);
  if (want < 1) want = 1;
  int64_t by_rows = cdiv(rows, 256);  // at least 256 rows per chunk
  if (by_rows < 1) by_rows = 1;
  return want < by_rows ? want : by_rows;
}

static size_t xty_partial_elems(int64_t rows, int P, int Q) {
  int tiles = (int)(cdiv(P, 64) * cdiv(Q, 64));
  return (size_t)pick_chunks(rows, tiles) * P * Q;
}

static int launch_xty(const float* X, const float* Y, int64_t rows, int P, int Q, float* partial, float* out,
                      cudaStream_t st) {
  int tiles = (int)(cdiv(P, 64) * cdiv(Q, 64));
  int64_t chunks = pick_chunks(rows, tiles);
  int64_t rpc = cdiv(rows > 0 ? rows : 1, chunks);
  dim3 grid((unsigned)chunks, (unsigned)cdiv(P, 64), (unsigned)cdiv(Q, 64));
  // HCG_CFG_GEMM_PATH = 1 keeps every dense product on the fp32 FMA pipe (A/B reference for the tests)
  if (config_get(HCG_CFG_GEMM_PATH) != 1 && P >= 32) {
    k_xty_mma<<<grid, 256, 0, st>>>(X, Y, rows, P, Q, rpc, partial);
    HCG_LAUNCH_OK(st, "k_xty_mma");
  } else {
    k_xty_partial<<<grid, 256, 0, st>>>(X, Y, rows, P, Q, rpc, partial);
    HCG_LAUNCH_OK(st, "k_xty_partial");
  }
  k_reduce_partials<<<(unsigned)cdiv((int64_t)P * Q, 256), 256, 0, st>>>(partial, (int)chunks, (int64_t)P * Q, out);
  HCG_LAUNCH_OK(st, "k_reduce_partials");
  return HCG_OK;
}

static int launch_colsum(const float* X, int64_t rows, int P, float* partial, float* out, cudaStream_t st) {
  int64_t chunks = pick_chunks(rows, 1);
  int64_t rpc = cdiv(rows > 0 ? rows : 1, chunks);
  k_colsum_partial<<<(unsigned)chunks, 256, 0, st>>>(X, rows, P, rpc, partial);
  HCG_LAUNCH_OK(st, "k_colsum_partial");
  k_reduce_partials<<<(unsigned)cdiv(P, 256), 256, 0, st>>>(partial, (int)chunks, P, out);
  HCG_LAUNCH_OK(st, "k_reduce_partials");
  return HCG_OK;
}

// elementwise helpers of the readout backward
__global__ void __launch_bounds__(256)
k_mask_mul(const float* __restrict__ g, const float* __restrict__ act_out, int64_t n, float slope,
           float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = g[i] * lrelu_grad_from_out(act_out[i], slope);
}

// ------------------------------------------------------------------------------------------
// argument checking shared by forward / backward
// ------------------------------------------------------------------------------------------
static int check_params(const HcgParams* p) {
  HCG_REQUIRE(p != nullptr, HCG_ERR_INVALID_ARG, "params is NULL");
  HCG_REQUIRE(p->F >= 1 && p->H >= 1 && p->C >= 1, HCG_ERR_INVALID_ARG, "bad dims F=%d H=%d C=%d", p->F, p->H, p->C);
  HCG_REQUIRE(p->L >= 1 && p->L <= HCG_MAX_CONV, HCG_ERR_UNSUPPORTED, "n_convolutions=%d outside [1,%d]", p->L, HCG_MAX_CONV);
  HCG_REQUIRE(p->R >= 1 && p->R <= HCG_MAX_READOUT, HCG_ERR_UNSUPPORTED, "readout_layers=%d outside [1,%d]", p->R, HCG_MAX_READOUT);
  int d = 2 * p->H;
  for (int k = 0; k < p->R - 1; ++k) {
    d /= 2;
    HCG_REQUIRE(d >= 1, HCG_ERR_UNSUPPORTED, "readout width reaches 0 at layer %d", k);
  }
  for (int l = 0; l < p->L; ++l)
    HCG_REQUIRE(p->conv_w[l] && p->conv_b[l], HCG_ERR_INVALID_ARG, "conv layer %d parameter pointer is NULL", l);
  for (int k = 0; k < p->R; ++k)
    HCG_REQUIRE(p->ro_w[k] && p->ro_b[k], HCG_ERR_INVALID_ARG, "readout layer %d parameter pointer is NULL", k);
  return HCG_OK;
}

static int check_graph(const HcgGraph* g, bool need_transpose) {
  HCG_REQUIRE(g != nullptr, HCG_ERR_INVALID_ARG, "graph is NULL");
  HCG_REQUIRE(g->N >= 0 && g->E >= 0 && g->B >= 0, HCG_ERR_INVALID_ARG, "negative graph sizes");
  HCG_REQUIRE(g->rowptr && g->dis && g->graph_ptr, HCG_ERR_INVALID_ARG, "graph rowptr/dis/graph_ptr is NULL");
  HCG_REQUIRE(g->E == 0 || g->col, HCG_ERR_INVALID_ARG, "graph col is NULL");
  if (need_transpose)
    HCG_REQUIRE(g->t_rowptr && (g->E == 0 || g->t_col), HCG_ERR_INVALID_ARG,
                "backward needs the transposed CSR (collate with HCG_COLLATE_TRANSPOSE)");
  if (g->ew0) HCG_REQUIRE(g->selfw0 && g->dis0, HCG_ERR_INVALID_ARG, "weighted layer-0 override is incomplete");
  if (g->edge_mask && g->E > 0)
    HCG_REQUIRE(g->eid && (!need_transpose || g->t_eid), HCG_ERR_INVALID_ARG, "edge_mask needs the eid / t_eid arrays of hcg_collate");
  return HCG_OK;
}

static inline int ro_in_dim(const HcgParams* p, int k) {
  int d = 2 * p->H;
  for (int i = 0; i < k; ++i) d /= 2;
  return d;
}
static inline int ro_out_dim(const HcgParams* p, int k) { return k == p->R - 1 ? p->C : ro_in_dim(p, k) / 2; }
static inline int max_ro_dim(const HcgParams* p) {
  int m = 2 * p->H;
  return m > p->C ? m : p->C;
}

struct FwdWs {
  float* T;
  float* act[2];
  float* hid[HCG_MAX_READOUT];
  char* tc;
  float* wimg;  // operand image of the tensor-core GEMM (layer-wise path, wide layers)
  size_t bytes;
};

static FwdWs carve_fwd(void* ws, int64_t N, int64_t B, const HcgParams* p) {
  Carver c(ws, (size_t)-1);
  FwdWs w{};
  w.T = c.take<float>((size_t)N * p->H + 4);
  w.act[0] = c.take<float>((size_t)N * p->H + 4);
  w.act[1] = c.take<float>((size_t)N * p->H + 4);
  for (int k = 0; k < p->R - 1; ++k) w.hid[k] = c.take<float>((size_t)B * ro_out_dim(p, k) + 4);
  w.tc = c.take<char>(tc_fwd_workspace_bytes(B));
  if (p->H % 64 == 0 && p->H <= 256) w.wimg = c.take<float>(tc_gemm_image_floats(p->F > p->H ? p->F : p->H, p->H) + 4);
  w.bytes = c.off;
  return w;
}

struct BwdWs {
  char* tcb;                   // tensor-core graph-resident backward: tile work list
  float *gA, *gT, *partial, *gz[2], *gemb;
  float *fpartial, *fscratch;  // graph-resident backward: per-CTA partials, oversize-graph scratch
  float *ropartial;            // fused readout backward: per-block partials
  float *wimg;                 // operand image of the tensor-core GEMM
  size_t bytes;
};

// the graph-resident backward applies (hcg_fused_bwd.cu); g_x and the config switch are checked by the caller
static bool fused_bwd_shape_ok(const HcgGraph* g, const HcgParams* p) {
  return p->H == kFusedH && p->L == 2 && p->F <= 32 && g->ew0 == nullptr && g->edge_mask == nullptr && g->B > 0 && g->N > 0 &&
         !(g->flags & HCG_GRAPH_CROSS_EDGES) && g->N / g->B <= 512;
}

static BwdWs carve_bwd(void* ws, int64_t N, int64_t B, const HcgParams* p) {
  Carver c(ws, (size_t)-1);
  BwdWs w{};
  // graph-resident scratch: three [N + 16 B + 16, 64] planes; gA / gT of the layer-wise path alias the
  // first two when the shape allows the graph-resident kernel (only one of the two paths runs per call)
  const bool fshape = p->H == kFusedH && p->L == 2 && p->F <= 32 && B > 0 && N > 0;
  if (fshape) {
    w.fscratch = c.take<float>(fused_bwd_scratch_elems(N, B) + 4);
    w.fpartial = c.take<float>(fused_bwd_partial_elems(B, p->F) + 4);
    w.tcb = c.take<char>(tc_bwd_workspace_bytes(B));
    w.gA = w.fscratch;
    w.gT = w.fscratch + (size_t)N * p->H + 64;
  } else {
    w.gA = c.take<float>((size_t)N * p->H + 4);
    w.gT = c.take<float>((size_t)N * p->H + 4);
  }
  size_t pe = 0;
  for (int l = 0; l < p->L; ++l) {
    size_t e = xty_partial_elems(N, p->H, l == 0 ? p->F : p->H);
    pe = e > pe ? e : pe;
  }
  for (int k = 0; k < p->R; ++k) {
    size_t e = xty_partial_elems(B, ro_out_dim(p, k), ro_in_dim(p, k));
    pe = e > pe ? e : pe;
  }
  size_t cs = (size_t)kMaxChunks * (size_t)(p->H > max_ro_dim(p) ? p->H : max_ro_dim(p));
  pe = cs > pe ? cs : pe;
  w.partial = c.take<float>(pe + 4);
  w.gz[0] = c.take<float>((size_t)B * max_ro_dim(p) + 4);
  w.gz[1] = c.take<float>((size_t)B * max_ro_dim(p) + 4);
  w.gemb = c.take<float>((size_t)B * 2 * p->H + 4);
  if (p->H % 64 == 0 && p->H <= 256) w.wimg = c.take<float>(tc_gemm_image_floats(p->H, p->H) + 4);
  if (p->R == 2 && 2 * p->H == RO_IN && p->C <= 16 && B > 0)
    w.ropartial = c.take<float>((size_t)cdiv(B, RO_G) * ro2_partial_stride(p->C) + 4);
  w.bytes = c.off;
  return w;
}

}  // namespace hcg

using namespace hcg;

extern "C" size_t hcg_forward_workspace_bytes(int64_t N, int64_t E, int64_t B, const HcgParams* p) {
  (void)E;
  if (!p || check_params(p)) return 0;
  return carve_fwd(nullptr, N, B, p).bytes;
}

extern "C" size_t hcg_backward_workspace_bytes(int64_t N, int64_t E, int64_t B, const HcgParams* p) {
  (void)E;
  if (!p || check_params(p)) return 0;
  return carve_bwd(nullptr, N, B, p).bytes;
}

// readout MLP on the pooled embeddings: the default head (128 -> 64 -> C) in one launch, anything else layer by layer
static int run_readout(const HcgParams* p, const float* emb, float* out, const HcgSaved* saved, float* const* hid_ws, int64_t B,
                       cudaStream_t st) {
  if (p->R == 2 && 2 * p->H == RO_IN && B > 0) {
    float* hs = saved ? saved->hid[0] : nullptr;
    if (saved) HCG_REQUIRE(hs != nullptr, HCG_ERR_INVALID_ARG, "saved->hid[0] is NULL");
    static SmemConfigured configured;
    if (int rc = ensure_dyn_smem(k_readout2, ro2_fwd_smem_bytes(), configured)) return rc;
    k_readout2<<<(unsigned)cdiv(B, RO_FG), 256, ro2_fwd_smem_bytes(), st>>>(emb, p->ro_w[0], p->ro_b[0], p->ro_w[1], p->ro_b[1], hs,
                                                                           out, B, p->C, p->slope);
    HCG_LAUNCH_OK(st, "k_readout2");
    return HCG_OK;
  }
  const float* z = emb;
  for (int k = 0; k < p->R; ++k) {
    const bool last = k == p->R - 1;
    float* zo = last ? out : (saved ? saved->hid[k] : hid_ws[k]);
    HCG_REQUIRE(zo != nullptr, HCG_ERR_INVALID_ARG, "saved->hid[%d] is NULL", k);
    const int rc = launch_gemm<true>(z, p->ro_w[k], zo, B, ro_in_dim(p, k), ro_out_dim(p, k),
                                     last ? EPI_BIAS : EPI_BIAS_LRELU, p->ro_b[k], p->slope, st);
    if (rc) return rc;
    z = zo;
  }
  return HCG_OK;
}

extern "C" int hcg_gcn_forward(const HcgGraph* g, const float* x, const HcgParams* p, float* out, float* emb,
                               const HcgSaved* saved, void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_params(p);
  if (rc) return rc;
  rc = check_graph(g, false);
  if (rc) return rc;
  HCG_REQUIRE(out && emb && ws, HCG_ERR_INVALID_ARG, "NULL out/emb/workspace");
  HCG_REQUIRE(g->N == 0 || x, HCG_ERR_INVALID_ARG, "x is NULL");
  FwdWs w = carve_fwd(ws, g->N, g->B, p);
  HCG_REQUIRE(ws_bytes >= w.bytes, HCG_ERR_WORKSPACE, "forward workspace too small: %zu < %zu", ws_bytes, w.bytes);
  const int64_t N = g->N, B = g->B;
  const int H = p->H;
  note_region(st);

  // ---- graph-resident path: conv layers + pooling in one kernel, then the readout --------------
  const int path = config_get(HCG_CFG_FORWARD_PATH);
  const bool fusable = H == kFusedH && p->F <= kMaxFusedF && g->ew0 == nullptr && g->edge_mask == nullptr && B > 0 && N > 0 &&
                       !(g->flags & HCG_GRAPH_CROSS_EDGES);
  // molecule-sized graphs only: one CTA walks a graph, so a batch of few huge graphs stays layer-wise
  const bool molecule_like = N / (B > 0 ? B : 1) <= 512;
  if (path >= 2 && !fusable)
    return set_err(HCG_ERR_UNSUPPORTED, "graph-resident forward needs H == %d, F <= %d, no edge weights and a disjoint batch", kFusedH, kMaxFusedF);
  bool fused_done = false;
  if (path >= 2 || (path == 0 && fusable && molecule_like)) {
    FusedFwdArgs a{};
    a.rowptr = g->rowptr; a.col = g->col; a.dis = g->dis; a.graph_ptr = g->graph_ptr; a.x = x;
    for (int l = 0; l < p->L; ++l) {
      a.conv_w[l] = p->conv_w[l]; a.conv_b[l] = p->conv_b[l];
      a.save_act[l] = saved ? saved->acts[l] : nullptr;
      if (saved) HCG_REQUIRE(saved->acts[l] != nullptr, HCG_ERR_INVALID_ARG, "saved->acts[%d] is NULL", l);
    }
    a.emb = emb; a.scratch_a = w.act[0]; a.scratch_b = w.act[1]; a.status = nullptr;
    a.B = B; a.F = p->F; a.Fp = (p->F + 3) / 4 * 4; a.L = p->L; a.slope = p->slope; a.fill = g->fill;
    // FMA kernel capacity: 128 nodes covers RhCAA (max 127) and the synthetic screening graphs (<= 120)
    // with two CTAs per SM; a hint above that selects 256 (biaryl: max 184) if it fits, at one CTA per SM
    int ncap = 128;
    if (g->max_graph_nodes > 128) ncap = 256;
    size_t smem = fused_fwd_smem_bytes(p->F, p->L, ncap, 6 * ncap);
    if (smem > 227 * 1024 && ncap > 128) { ncap = 128; smem = fused_fwd_smem_bytes(p->F, p->L, ncap, 6 * ncap); }
    const bool fma_fits = smem <= 227 * 1024;
    a.ncap = ncap; a.ecap = 6 * ncap;
    // tensor-core kernel: F <= 32, its smem must fit, the FMA kernel must exist for the overflow list,
    // and the batch should mostly fit the 128-node slots
    const bool tc_ok = p->F <= 32 && tc_fwd_smem_bytes(p->F, p->L) <= 227 * 1024 && fma_fits &&
                       p->slope >= 0.f && p->slope <= 1.f;  // its LeakyReLU is max(v, slope * v)
    if (path == 3 && !tc_ok)
      return set_err(HCG_ERR_UNSUPPORTED, "tensor-core forward needs F <= 32 and %d conv layers to fit in shared memory", p->L);
    const bool want_tc = path == 3 || (path == 0 && tc_ok && (g->max_graph_nodes == 0 || g->max_graph_nodes <= 128));
    if (want_tc) {
      const bool all_fit = g->max_graph_nodes > 0 && g->max_graph_nodes <= tc_fwd_row_cap() && g->max_graph_edges > 0 &&
                           g->max_graph_edges <= tc_fwd_edge_cap();
      rc = launch_tc_fwd(a, N, g->E, w.tc, st, all_fit);
      if (rc) return rc;
      fused_done = true;
    } else if (fma_fits) {
      rc = launch_fused_fwd(a, smem, st);
      if (rc) return rc;
      fused_done = true;
    } else if (path >= 2) {
      return set_err(HCG_ERR_UNSUPPORTED, "graph-resident forward: %d conv layers do not fit in shared memory", p->L);
    }
  }

  const float* h_in = x;
  int k_in = p->F;
  float* h_out = nullptr;
  bool pooled = false;
  for (int l = 0; l < p->L && !fused_done; ++l) {
    rc = launch_gemm<true>(h_in, p->conv_w[l], w.T, N, k_in, H, EPI_NONE, nullptr, p->slope, st, w.wimg);
    if (rc) return rc;
    h_out = saved ? saved->acts[l] : w.act[l & 1];
    HCG_REQUIRE(h_out != nullptr, HCG_ERR_INVALID_ARG, "saved->acts[%d] is NULL", l);
    const bool weighted0 = (l == 0 && g->ew0 != nullptr);
    const int kind = weighted0 ? AGG_ROWS : agg_kind(g, H, w.T, h_out);
    if (kind == AGG_WIDE) {
      // whole-row kernel; the last layer pools in its epilogue and keeps its rows only for a backward
      const bool last = l == p->L - 1 && B > 0;
      rc = launch_agg_wide(g->rowptr, g->col, g->dis, g->fill, w.T, p->conv_b[l], last && !saved ? nullptr : h_out, N, H, 1,
                           p->slope, g->graph_ptr, B, last ? emb : nullptr, st);
      pooled = last;
    } else if (kind == AGG_NARROW) {
      rc = launch_agg_narrow(g->rowptr, g->col, g->dis, g->fill, w.T, p->conv_b[l], h_out, N, H, 1, p->slope, st);
    } else {
      rc = launch_agg(g->rowptr, g->col, weighted0 ? g->dis0 : g->dis, weighted0 ? g->ew0 : nullptr,
                      weighted0 ? g->selfw0 : nullptr, g->fill, w.T, p->conv_b[l], h_out, N, H, 1, p->slope, st,
                      g->eid, g->edge_mask);
    }
    if (rc) return rc;
    h_in = h_out;
    k_in = H;
  }
  if (B > 0 && !fused_done && !pooled) {
    int lpr = lanes_per_row(H, 1);
    k_pool<<<(unsigned)cdiv(B, 256 / lpr), 256, 0, st>>>(h_out, g->graph_ptr, B, H, lpr, emb);
    HCG_LAUNCH_OK(st, "k_pool");
  }
  return run_readout(p, emb, out, saved, w.hid, B, st);
}

// ---- forward from the raw edge list (no collate) ---------------------------------------------------
struct CooWs {
  char* tc;
  float* hid[HCG_MAX_READOUT];
  size_t bytes;
};

static CooWs carve_coo(void* ws, int64_t B, const HcgParams* p) {
  Carver c(ws, (size_t)-1);
  CooWs w{};
  w.tc = c.take<char>(tc_fwd_coo_workspace_bytes(B));   // first: hcg_forward_coo_status finds the flags from B alone
  for (int k = 0; k < p->R - 1; ++k) w.hid[k] = c.take<float>((size_t)B * ro_out_dim(p, k) + 4);
  w.bytes = c.off;
  return w;
}

extern "C" int hcg_forward_coo_supported(const HcgParams* p) {
  if (!p || check_params(p)) return 0;
  return p->H == kFusedH && p->F <= 32 && p->slope >= 0.f && p->slope <= 1.f && tc_fwd_coo_smem_bytes(p->F, p->L) <= 227 * 1024;
}

extern "C" size_t hcg_forward_coo_workspace_bytes(int64_t B, const HcgParams* p) {
  if (!p || check_params(p) || B < 0) return 0;
  return carve_coo(nullptr, B, p).bytes;
}

extern "C" int hcg_gcn_forward_coo(const float* x, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr,
                                   int64_t N, int64_t E, int64_t B, float fill, const HcgParams* p, float* out, float* emb,
                                   void* ws, size_t ws_bytes, void* stream) {
  return gcn_forward_coo_saved(x, edge_index, node_ptr, edge_ptr, N, E, B, fill, p, out, emb, nullptr, ws, ws_bytes, stream);
}

// the same with the activations saved for hcg_gcn_backward (the training step's forward: hcg_train.cu)
int hcg::gcn_forward_coo_saved(const float* x, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N,
                               int64_t E, int64_t B, float fill, const HcgParams* p, float* out, float* emb, const HcgSaved* saved,
                               void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_params(p);
  if (rc) return rc;
  HCG_REQUIRE(N >= 0 && E >= 0 && B >= 0 && N < INT32_MAX && E < INT32_MAX, HCG_ERR_INVALID_ARG, "bad sizes N=%lld E=%lld B=%lld",
              (long long)N, (long long)E, (long long)B);
  HCG_REQUIRE((out || saved) && emb && ws, HCG_ERR_INVALID_ARG, "NULL out/emb/workspace");
  HCG_REQUIRE((N == 0 || x) && (E == 0 || edge_index) && node_ptr && edge_ptr, HCG_ERR_INVALID_ARG, "NULL x/edge_index/offsets");
  if (!hcg_forward_coo_supported(p))
    return set_err(HCG_ERR_UNSUPPORTED, "edge-list forward needs H == %d, F <= 32, 0 <= slope <= 1 and conv layers that fit shared memory", kFusedH);
  CooWs w = carve_coo(ws, B, p);
  HCG_REQUIRE(ws_bytes >= w.bytes, HCG_ERR_WORKSPACE, "edge-list forward workspace too small: %zu < %zu", ws_bytes, w.bytes);
  if (B == 0) return HCG_OK;
  note_region(st);
  FusedFwdArgs a{};
  a.x = x;
  for (int l = 0; l < p->L; ++l) {
    a.conv_w[l] = p->conv_w[l]; a.conv_b[l] = p->conv_b[l];
    a.save_act[l] = saved ? saved->acts[l] : nullptr;
    if (saved) HCG_REQUIRE(saved->acts[l] != nullptr, HCG_ERR_INVALID_ARG, "saved->acts[%d] is NULL", l);
  }
  a.emb = emb;
  a.B = B; a.F = p->F; a.Fp = (p->F + 3) / 4 * 4; a.L = p->L; a.slope = p->slope; a.fill = fill;
  rc = launch_tc_fwd_coo(a, edge_index, node_ptr, edge_ptr, N, E, w.tc, st);
  if (rc || out == nullptr) return rc;     // (out == NULL: embeddings only - the training step runs the head itself)
  return run_readout(p, emb, out, saved, w.hid, B, st);
}

// ---- the training step's head in one kernel (+ the reduce): forward, loss seed, backward of the default readout ------
bool hcg::readout2_train_supported(const HcgParams* p) { return p->R == 2 && 2 * p->H == RO_IN && p->C >= 1 && p->C <= 16; }
size_t hcg::readout2_train_partial_floats(int64_t B, int C) {
  return (size_t)cdiv(B, RO_G) * (size_t)ro2_partial_stride(C) + (size_t)cdiv(B, RO_G) + 8;
}
int hcg::launch_readout2_train(const float* emb, const HcgParams* p, const float* y, int64_t B, float* g_emb, float* partial,
                               const HcgGrads* grads, float* tail, cudaStream_t st) {
  const size_t smem = ro2_bwd_smem_bytes(p->C, true);
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_readout2_bwd<true>, smem, configured)) return rc;
  const int blocks = (int)cdiv(B, RO_G);
  float* sse_partial = partial + (size_t)blocks * ro2_partial_stride(p->C);
  k_readout2_bwd<true><<<blocks, 256, smem, st>>>(emb, nullptr, nullptr, nullptr, p->ro_w[0], p->ro_w[1], B, p->C, p->slope, g_emb,
                                                 partial, p->ro_b[0], p->ro_b[1], y, sse_partial);
  HCG_LAUNCH_OK(st, "k_readout2_train");
  k_readout2_bwd_reduce<<<(unsigned)cdiv(ro2_partial_elems(p->C), 32), 256, 0, st>>>(
      partial, blocks, p->C, grads->ro_w[0], grads->ro_b[0], grads->ro_w[1], grads->ro_b[1], sse_partial, tail, (float)(B * p->C));
  HCG_LAUNCH_OK(st, "k_readout2_bwd_reduce");
  return HCG_OK;
}

extern "C" int hcg_forward_coo_status(const void* ws, int64_t B, void* stream, int32_t* flags) {
  HCG_REQUIRE(ws && flags && B >= 0, HCG_ERR_INVALID_ARG, "NULL workspace / flags");
  int f = 0;
  const int rc = tc_fwd_coo_flags(ws, B, static_cast<cudaStream_t>(stream), &f);   // (the tile work list opens the workspace)
  *flags = f;
  return rc;
}

// ---- forward straight from a packed graph image (the on-disk / on-wire format) -------------
extern "C" int hcg_forward_packed_supported(const HcgPackedStore* store, const HcgParams* p) {
  if (!store || !hcg_forward_coo_supported(p)) return 0;
  if (store->F != p->F || store->Fb < 0 || store->Fu + store->Ff + store->Fb != store->F) return 0;
  if (store->Fb > 0 && (store->xbits == nullptr || store->bit_stride * 8 < store->Fb)) return 0;
  return tc_fwd_packed_fits(*store) ? 1 : 0;
}

extern "C" int hcg_gcn_forward_packed(const HcgPackedStore* store, int64_t first, int64_t B, float fill, const HcgParams* p,
                                      float* out, float* emb, void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_params(p);
  if (rc) return rc;
  HCG_REQUIRE(store != nullptr && out && emb && ws, HCG_ERR_INVALID_ARG, "NULL store/out/emb/workspace");
  HCG_REQUIRE(first >= 0 && B >= 0 && first + B <= store->G, HCG_ERR_INVALID_ARG, "graphs [%lld, %lld) outside the image (G=%lld)",
              (long long)first, (long long)(first + B), (long long)store->G);
  HCG_REQUIRE(store->N < INT32_MAX && store->E < INT32_MAX, HCG_ERR_UNSUPPORTED, "image above 2^31 nodes or edges");
  if (!hcg_forward_packed_supported(store, p))
    return set_err(HCG_ERR_UNSUPPORTED, "packed forward needs the edge-list forward's network shape, byte edge endpoints, F == the "
                                        "image's F <= 32 and at most ~50 packed bytes per node (use hcg_unpack_batch + hcg_gcn_forward_coo)");
  HCG_REQUIRE(store->node_ptr && store->edge_ptr && store->col_slot && (store->E == 0 || store->edges) &&
              (store->Fu == 0 || store->xu8) && (store->Ff == 0 || store->xf32), HCG_ERR_INVALID_ARG, "NULL section pointer");
  CooWs w = carve_coo(ws, B, p);
  HCG_REQUIRE(ws_bytes >= w.bytes, HCG_ERR_WORKSPACE, "packed forward workspace too small: %zu < %zu", ws_bytes, w.bytes);
  if (B == 0) return HCG_OK;
  note_region(st);
  FusedFwdArgs a{};
  for (int l = 0; l < p->L; ++l) { a.conv_w[l] = p->conv_w[l]; a.conv_b[l] = p->conv_b[l]; }
  a.emb = emb;
  a.B = B; a.F = p->F; a.Fp = (p->F + 3) / 4 * 4; a.L = p->L; a.slope = p->slope; a.fill = fill;
  rc = launch_tc_fwd_packed(a, *store, first, w.tc, st);
  if (rc) return rc;
  return run_readout(p, emb, out, nullptr, w.hid, B, st);
}

// ---- one-pass ensemble forward ------------------------------------------------
struct EnsWs {
  char* tc;
  float* wimg;
  size_t bytes;
};
static EnsWs carve_ens(void* ws, int64_t B, int M) {
  Carver c(ws, (size_t)-1);
  EnsWs w{};
  w.tc = c.take<char>(tc_fwd_coo_workspace_bytes(B));   // first: hcg_forward_coo_status finds the flags from B alone
  w.wimg = c.take<float>(tc_fwd_ens_image_floats(M) + 4);
  w.bytes = c.off;
  return w;
}

static int check_ensemble(const HcgEnsemble* e) {
  HCG_REQUIRE(e != nullptr, HCG_ERR_INVALID_ARG, "ensemble is NULL");
  HCG_REQUIRE(e->M >= 1 && e->M <= 4096, HCG_ERR_INVALID_ARG, "ensemble of %d models", (int)e->M);
  HCG_REQUIRE(e->H == kFusedH && e->F >= 1 && e->F <= 32 && e->C >= 1 && e->C <= 64, HCG_ERR_UNSUPPORTED,
              "one-pass ensemble forward needs H == %d, F <= 32, C <= 64 (got H=%d F=%d C=%d)", kFusedH, (int)e->H, (int)e->F, (int)e->C);
  HCG_REQUIRE(e->slope >= 0.f && e->slope <= 1.f, HCG_ERR_UNSUPPORTED, "LeakyReLU slope %g outside [0, 1]", (double)e->slope);
  HCG_REQUIRE(e->conv_w0 && e->conv_b0 && e->conv_w1 && e->conv_b1 && e->ro_w0 && e->ro_b0 && e->ro_w1 && e->ro_b1,
              HCG_ERR_INVALID_ARG, "NULL weight pointer in the ensemble");
  return HCG_OK;
}

extern "C" size_t hcg_ensemble_workspace_bytes(int64_t B, int32_t M) {
  if (B < 0 || M < 1) return 0;
  return carve_ens(nullptr, B, M).bytes;
}

extern "C" int hcg_gcn_forward_coo_ensemble(const float* x, const int64_t* edge_index, const int64_t* node_ptr,
                                            const int64_t* edge_ptr, int64_t N, int64_t E, int64_t B, float fill,
                                            const HcgEnsemble* ens, float* out, float* emb, void* ws, size_t ws_bytes,
                                            void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_ensemble(ens);
  if (rc) return rc;
  HCG_REQUIRE(N >= 0 && E >= 0 && B >= 0 && N < INT32_MAX && E < INT32_MAX, HCG_ERR_INVALID_ARG, "bad sizes N=%lld E=%lld B=%lld",
              (long long)N, (long long)E, (long long)B);
  HCG_REQUIRE(out && emb && ws, HCG_ERR_INVALID_ARG, "NULL out/emb/workspace");
  HCG_REQUIRE((N == 0 || x) && (E == 0 || edge_index) && node_ptr && edge_ptr, HCG_ERR_INVALID_ARG, "NULL x/edge_index/offsets");
  EnsWs w = carve_ens(ws, B, ens->M);
  HCG_REQUIRE(ws_bytes >= w.bytes, HCG_ERR_WORKSPACE, "ensemble workspace too small: %zu < %zu", ws_bytes, w.bytes);
  if (B == 0) return HCG_OK;
  note_region(st);
  rc = launch_tc_fwd_ens(x, edge_index, node_ptr, edge_ptr, N, E, B, ens->M, ens->F, ens->slope, fill, ens->conv_w0, ens->conv_b0,
                         ens->conv_w1, ens->conv_b1, w.wimg, emb, w.tc, st);
  if (rc) return rc;
  static SmemConfigured ro_configured;
  if (int rc2 = ensure_dyn_smem(k_readout2_ens, ro2_fwd_smem_bytes(), ro_configured)) return rc2;
  k_readout2_ens<<<dim3((unsigned)cdiv(B, RO_FG), (unsigned)ens->M), 256, ro2_fwd_smem_bytes(), st>>>(
      emb, ens->ro_w0, ens->ro_b0, ens->ro_w1, ens->ro_b1, out, B, ens->C, ens->slope);
  HCG_LAUNCH_OK(st, "k_readout2_ens");
  return HCG_OK;
}

// ---- the same forward with NO caller workspace and no allocation of any kind -------------------------
// (no counterpart in the reference: added for callers that cannot allocate -
//  code under CUDA graph capture, tight serving loops.)  The scratch of hcg_gcn_forward_coo - the tile work list
// k_tile_desc writes and k_tc_fwd reads, the per-slot refusal flags, the hidden rows of a readout deeper than the
// default - lives in one file-scope device array.  Every GPU that loads the library holds its kCooStaticBytes until
// exit.  The array is shared by every call on a device and keeps the last call's contents: nothing in it has to start
// at zero (every word a call reads, that call's own kernels have written first), but two calls that overlap on the
// GPU (different streams, no ordering between them) race on it.
constexpr int64_t kCooStaticMaxGraphs = 16384;
constexpr size_t kCooStaticBytes = 6u << 20;     // 6 MiB: a 16384-graph batch of the default network needs 5.1 MiB
static __device__ __align__(256) unsigned char g_coo_static_scratch[kCooStaticBytes];

// Device address of the array on the CURRENT device (in host code the array's name is only a handle for
// cudaGetSymbolAddress, not an address a kernel could use); looked up once per device.
static int coo_static_scratch(void** out) {
  constexpr int kMaxDevices = 64;
  static std::atomic<void*> cache[kMaxDevices];
  int dev = 0;
  HCG_CUDA_OK(cudaGetDevice(&dev));
  HCG_REQUIRE(dev >= 0 && dev < kMaxDevices, HCG_ERR_UNSUPPORTED, "device ordinal %d outside [0, %d)", dev, kMaxDevices);
  void* ptr = cache[dev].load(std::memory_order_acquire);
  if (ptr == nullptr) {
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_coo_static_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return set_err(HCG_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d failed: %s", dev, cudaGetErrorString(e));
    cache[dev].store(ptr, std::memory_order_release);   // (two threads racing here store the same address)
  }
  *out = ptr;
  return HCG_OK;
}

extern "C" int hcg_gcn_forward_coo_static(const float* x, const int64_t* edge_index, const int64_t* node_ptr,
                                          const int64_t* edge_ptr, int64_t N, int64_t E, int64_t B, float fill,
                                          const HcgParams* p, float* out, float* emb, int32_t* flags, void* stream) {
  int rc = check_params(p);
  if (rc) return rc;
  HCG_REQUIRE(B >= 0 && B <= kCooStaticMaxGraphs, HCG_ERR_WORKSPACE,
              "hcg_gcn_forward_coo_static takes at most %lld graphs per call, got %lld (use hcg_gcn_forward_coo with a workspace)",
              (long long)kCooStaticMaxGraphs, (long long)B);
  const size_t need = carve_coo(nullptr, B, p).bytes;
  HCG_REQUIRE(need <= kCooStaticBytes, HCG_ERR_WORKSPACE,
              "hcg_gcn_forward_coo_static: this network needs %zu bytes of scratch for %lld graphs, the static array holds %zu",
              need, (long long)B, kCooStaticBytes);
  void* scratch = nullptr;
  rc = coo_static_scratch(&scratch);
  if (rc) return rc;
  rc = hcg_gcn_forward_coo(x, edge_index, node_ptr, edge_ptr, N, E, B, fill, p, out, emb, scratch, kCooStaticBytes, stream);
  if (rc || flags == nullptr) return rc;
  return hcg_forward_coo_status(scratch, B, stream, flags);
}

// event the NEXT hcg_gcn_backward of this thread waits for before it reads the graph arrays (set by hcg_train_step)
static thread_local cudaEvent_t g_backward_prereq = nullptr;
static thread_local cudaStream_t g_backward_side = nullptr;
static thread_local const int64_t* g_backward_node_ptr = nullptr;   // the batch's int64 offset vectors (training step)
static thread_local const int64_t* g_backward_edge_ptr = nullptr;
static thread_local bool g_backward_keep_readout = false;
// the next hcg_gcn_backward of this thread (called with g_out == NULL) leaves the readout gradients as they are
void hcg::set_backward_keep_readout_grads(bool keep) { g_backward_keep_readout = keep; }
void hcg::set_backward_prereq(cudaEvent_t ev, cudaStream_t side, const int64_t* node_ptr, const int64_t* edge_ptr) {
  g_backward_prereq = ev; g_backward_side = ev ? side : nullptr;
  g_backward_node_ptr = ev ? node_ptr : nullptr; g_backward_edge_ptr = ev ? edge_ptr : nullptr;
}

extern "C" int hcg_gcn_backward(const HcgGraph* g, const float* x, const HcgParams* p, const HcgSaved* saved,
                                const float* emb, const float* g_out, const float* g_emb,
                                const HcgGrads* grads, float* g_x, void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_params(p);
  if (rc) return rc;
  rc = check_graph(g, true);
  if (rc) return rc;
  HCG_REQUIRE(saved && emb && grads && ws, HCG_ERR_INVALID_ARG, "NULL saved/emb/grads/workspace");
  HCG_REQUIRE(g_out || g_emb, HCG_ERR_INVALID_ARG, "both g_out and g_emb are NULL");
  for (int l = 0; l < p->L; ++l)
    HCG_REQUIRE(saved->acts[l] && grads->conv_w[l] && grads->conv_b[l], HCG_ERR_INVALID_ARG, "NULL saved act / grad pointer at conv %d", l);
  for (int k = 0; k < p->R; ++k)
    HCG_REQUIRE(grads->ro_w[k] && grads->ro_b[k] && (k == p->R - 1 || saved->hid[k]), HCG_ERR_INVALID_ARG,
                "NULL saved hid / grad pointer at readout %d", k);
  BwdWs w = carve_bwd(ws, g->N, g->B, p);
  HCG_REQUIRE(ws_bytes >= w.bytes, HCG_ERR_WORKSPACE, "backward workspace too small: %zu < %zu", ws_bytes, w.bytes);
  const int64_t N = g->N, B = g->B;
  const int H = p->H;
  note_region(st);

  // ---- readout MLP backward -------------------------------------------------------------
  const bool keep_readout_grads = g_backward_keep_readout;
  g_backward_keep_readout = false;
  const float* gemb_total = g_emb;
  if (g_out && w.ropartial != nullptr && config_get(HCG_CFG_BACKWARD_PATH) != 1) {
    // default head (128 -> 64 -> C): one kernel + one reduce
    const size_t smem = ro2_bwd_smem_bytes(p->C);
    static SmemConfigured configured;
    if (int rc2 = ensure_dyn_smem(k_readout2_bwd<false>, smem, configured)) return rc2;
    const int blocks = (int)cdiv(B, RO_G);
    k_readout2_bwd<false><<<blocks, 256, smem, st>>>(emb, saved->hid[0], g_out, g_emb, p->ro_w[0], p->ro_w[1], B, p->C, p->slope,
                                                    w.gemb, w.ropartial, nullptr, nullptr, nullptr, nullptr);
    HCG_LAUNCH_OK(st, "k_readout2_bwd");
    k_readout2_bwd_reduce<<<(unsigned)cdiv(ro2_partial_elems(p->C), 32), 256, 0, st>>>(
        w.ropartial, blocks, p->C, grads->ro_w[0], grads->ro_b[0], grads->ro_w[1], grads->ro_b[1]);
    HCG_LAUNCH_OK(st, "k_readout2_bwd_reduce");
    gemb_total = w.gemb;
  } else if (g_out) {
    const float* gz = g_out;  // gradient w.r.t. the (post-activation) output of layer k
    int pp = 0;
    for (int k = p->R - 1; k >= 0; --k) {
      const int din = ro_in_dim(p, k), dout = ro_out_dim(p, k);
      const float* zin = k == 0 ? emb : saved->hid[k - 1];
      const float* gpre = gz;
      if (k != p->R - 1) {  // through the LeakyReLU of this hidden layer
        float* buf = w.gz[pp]; pp ^= 1;
        if (B * dout > 0) {
          k_mask_mul<<<(unsigned)cdiv(B * dout, 256), 256, 0, st>>>(gz, saved->hid[k], B * dout, p->slope, buf);
          HCG_LAUNCH_OK(st, "k_mask_mul");
        }
        gpre = buf;
      }
      rc = launch_xty(gpre, zin, B, dout, din, w.partial, grads->ro_w[k], st);
      if (rc) return rc;
      rc = launch_colsum(gpre, B, dout, w.partial, grads->ro_b[k], st);
      if (rc) return rc;
      // gradient w.r.t. this layer's input
      float* gin = k == 0 ? w.gemb : w.gz[pp];
      if (k != 0) pp ^= 1;
      if (k == 0 && g_emb)
        rc = launch_gemm<false>(gpre, p->ro_w[k], gin, B, dout, din, EPI_ADD, g_emb, p->slope, st);
      else
        rc = launch_gemm<false>(gpre, p->ro_w[k], gin, B, dout, din, EPI_NONE, nullptr, p->slope, st);
      if (rc) return rc;
      gz = gin;
    }
    gemb_total = w.gemb;
  } else if (!keep_readout_grads) {
    for (int k = 0; k < p->R; ++k) {
      HCG_CUDA_OK(cudaMemsetAsync(grads->ro_w[k], 0, sizeof(float) * (size_t)ro_in_dim(p, k) * ro_out_dim(p, k), st));
      HCG_CUDA_OK(cudaMemsetAsync(grads->ro_b[k], 0, sizeof(float) * (size_t)ro_out_dim(p, k), st));
    }
  }

  // (hcg_train_step builds the transposed CSR on a second stream while the readout part above runs: it is joined here,
  //  or - tensor-core path - after the backward's work list has been built on that stream as well)
  const cudaEvent_t side_ev = g_backward_prereq;
  const cudaStream_t side_st = g_backward_side;
  const int64_t* const np64 = g_backward_node_ptr;
  const int64_t* const ep64 = g_backward_edge_ptr;
  g_backward_prereq = nullptr; g_backward_side = nullptr; g_backward_node_ptr = nullptr; g_backward_edge_ptr = nullptr;
  bool joined = side_ev == nullptr;
  auto join_side = [&]() -> int {
    if (!joined) { joined = true; HCG_CUDA_OK(cudaStreamWaitEvent(st, side_ev, 0)); }
    return HCG_OK;
  };
  // ---- graph-resident path: pooling backward + both conv layers in one kernel ----------------
  {
    const int bpath = config_get(HCG_CFG_BACKWARD_PATH);
    const bool ok = fused_bwd_shape_ok(g, p) && g_x == nullptr && w.fscratch != nullptr;
    if ((bpath == 2 || bpath == 3) && !ok) {
      join_side();
      return set_err(HCG_ERR_UNSUPPORTED, "graph-resident backward needs H == %d, L == 2, F <= 32, no edge weights, a disjoint batch and g_x == NULL", kFusedH);
    }
    if (ok && bpath != 1) {
      FusedBwdArgs a{};
      a.t_rowptr = g->t_rowptr; a.t_col = g->t_col; a.dis = g->dis; a.graph_ptr = g->graph_ptr;
      a.x = x; a.h1 = saved->acts[0]; a.h2 = saved->acts[1]; a.emb = emb; a.gemb = gemb_total;
      a.w2 = p->conv_w[1]; a.scratch = w.fscratch; a.partial = w.fpartial;
      a.B = B; a.F = p->F; a.slope = p->slope; a.fill = g->fill;
      // bpath 3 = the mma.sync kernel only; otherwise the tcgen05 kernel takes every graph that fits a tile
      if (bpath != 3 && w.tcb != nullptr) {
        const bool all_fit = g->max_graph_nodes > 0 && g->max_graph_nodes <= tc_fwd_row_cap() && g->max_graph_edges > 0 &&
                             g->max_graph_edges <= tc_fwd_edge_cap();
        if (!joined) {   // the work list goes out on the second stream, behind the collate; `st` joins after it
          joined = true;
          return launch_tc_bwd(a, N, g->E, w.tcb, all_fit, grads->conv_w[1], grads->conv_w[0], grads->conv_b[1], grads->conv_b[0], st,
                               side_st, side_ev, np64, ep64);
        }
        return launch_tc_bwd(a, N, g->E, w.tcb, all_fit, grads->conv_w[1], grads->conv_w[0], grads->conv_b[1], grads->conv_b[0], st);
      }
      if (int rcj = join_side()) return rcj;
      return launch_fused_bwd(a, N, grads->conv_w[1], grads->conv_w[0], grads->conv_b[1], grads->conv_b[0], st);
    }
  }
  if (int rcj = join_side()) return rcj;

  // ---- pooling backward (+ LeakyReLU' of the last conv layer) -----------------------------
  if (N > 0) {  // nodes of no graph
    HCG_CUDA_OK(cudaMemsetAsync(w.gA, 0, sizeof(float) * (size_t)N * H, st));
    note_other("memset(gA)", st);
  }
  if (B > 0) {
    int lpr = lanes_per_row(H, 1);
    k_pool_bwd<<<(unsigned)cdiv(B, 256 / lpr), 256, 0, st>>>(saved->acts[p->L - 1], emb, gemb_total, g->graph_ptr,
                                                              B, H, lpr, p->slope, w.gA);
    HCG_LAUNCH_OK(st, "k_pool_bwd");
  }

  // ---- conv layers backward ---------------------------------------------------------------
  if (grads->edge_mask && g->E > 0)  // dropped self-loop edges and a call without a mask leave zeros
    HCG_CUDA_OK(cudaMemsetAsync(grads->edge_mask, 0, sizeof(float) * (size_t)g->E, st));
  for (int l = p->L - 1; l >= 0; --l) {
    const float* h_in = l == 0 ? x : saved->acts[l - 1];
    const int k_in = l == 0 ? p->F : H;
    rc = launch_colsum(w.gA, N, H, w.partial, grads->conv_b[l], st);
    if (rc) return rc;
    const bool weighted0 = (l == 0 && g->ew0 != nullptr);
    if (grads->edge_mask && g->edge_mask && g->E > 0 && N > 0) {
      // T_l = h_in W_l^T again (the forward does not keep it), into the not yet used gT buffer
      rc = launch_gemm<true>(h_in, p->conv_w[l], w.gT, N, k_in, H, EPI_NONE, nullptr, p->slope, st, w.wimg);
      if (rc) return rc;
      k_edge_dot<<<(unsigned)cdiv(N, 8), 256, 0, st>>>(g->rowptr, g->col, g->eid, weighted0 ? g->dis0 : g->dis,
                                                     weighted0 ? g->ew0 : nullptr, w.gA, w.gT, N, H, 1, grads->edge_mask);
      HCG_LAUNCH_OK(st, "k_edge_dot");
    }
    // gT = Â^T · gA  (transposed CSR; no bias, no activation)
    const int tkind = weighted0 ? AGG_ROWS : agg_kind(g, H, w.gA, w.gT);
    if (tkind == AGG_WIDE)
      rc = launch_agg_wide(g->t_rowptr, g->t_col, g->dis, g->fill, w.gA, nullptr, w.gT, N, H, 0, p->slope, g->graph_ptr, B,
                           nullptr, st);
    else if (tkind == AGG_NARROW)
      rc = launch_agg_narrow(g->t_rowptr, g->t_col, g->dis, g->fill, w.gA, nullptr, w.gT, N, H, 0, p->slope, st);
    else
      rc = launch_agg(g->t_rowptr, g->t_col, weighted0 ? g->dis0 : g->dis, weighted0 ? g->t_ew0 : nullptr,
                      weighted0 ? g->selfw0 : nullptr, g->fill, w.gA, nullptr, w.gT, N, H, 0, p->slope, st,
                      g->t_eid, g->edge_mask);
    if (rc) return rc;
    rc = launch_xty(w.gT, h_in, N, H, k_in, w.partial, grads->conv_w[l], st);
    if (rc) return rc;
    if (l > 0) {
      rc = launch_gemm<false>(w.gT, p->conv_w[l], w.gA, N, H, k_in, EPI_MASK, saved->acts[l - 1], p->slope, st, w.wimg);
      if (rc) return rc;
    } else if (g_x) {
      rc = launch_gemm<false>(w.gT, p->conv_w[0], g_x, N, H, k_in, EPI_NONE, nullptr, p->slope, st);
      if (rc) return rc;
    }
  }
  return HCG_OK;
}
