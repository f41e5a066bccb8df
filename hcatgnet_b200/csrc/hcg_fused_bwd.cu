// Graph-resident backward of the GCN hot path (H = 64, two conv layers): one CTA owns one reaction
// graph at a time and keeps every [n,64] intermediate of that graph on chip.  It replaces, per batch,
// the layer-wise chain  k_pool_bwd -> (k_colsum, k_agg^T, k_xty, k_gemm) x 2  and its ~1 GB of [N,64]
// round trips through HBM by one pass that reads h1, h2 (saved by the forward), x and the transposed
// CSR once and writes only per-CTA partial sums of the parameter gradients.
//
// With z_l = (Â h_{l-1}) W_l^T + b_l, h_l = lrelu(z_l), emb = [max_i h_2 || mean_i h_2]
// (reference model/gcn.py:58-66) and g_emb = d loss / d emb, per graph:
//
//   P1  gA2 = pool'(g_emb) . lrelu'(h2)          max routed to the arg-max rows (ties split evenly,
//                                                 torch scatter_reduce(amax) convention), mean / n
//       g_b2 += colsum(gA2)
//   P2  gT2 = Â^T gA2                            transposed CSR, neighbours in slot order, self-loop last
//   P3  g_W2 += gT2^T h1                         [64 x n] . [n x 64]
//   P4  gA1 = (gT2 W2) . lrelu'(h1)              [n x 64] . [64 x 64]
//   P5  g_b1 += colsum(gA1)
//   P6  gT1 = Â^T gA1
//   P7  g_W1 += gT1^T x                          [64 x n] . [n x F]
//
// The three dense products run on the tensor cores with warp-level mma.sync (m16n8k8, TF32 operands,
// fp32 accumulate) and the 3-term error-compensated split  a.b ~ a_lo.b_hi + a_hi.b_lo + a_hi.b_hi
// (fp32-level accuracy: the dropped a_lo.b_lo term is 2^-22 relative).  Their shapes are tiny and their
// reduction dimension is the NODE index (P3, P7), which is the row index of every buffer on chip — a
// fragment load straight out of shared memory, no transposition, no TMEM choreography.  Each graph's
// product starts from a zero accumulator and is then added to the thread's running fp32 sum, so the
// tensor core's accumulator rounding never compounds over the batch.
//
// Activation buffers are [rows][64] floats with the column index XOR-swizzled by the row
// (swz(r) = ((r & 3) << 3) | (r & 4)): conflict-free for both fragment patterns (8 rows x 4 columns and
// 4 rows x 8 columns), for column owners and for 128-bit row gathers.
// Parameter-gradient partials are summed over CTAs in CTA order by k_fused_bwd_reduce: no floating
// point atomics, run-to-run identical.  Graphs above the shared-memory capacity run the same code on a
// private slice of a global scratch buffer.
#include "hcg_fused.cuh"
#include "hcg_mma.cuh"

namespace hcg {

extern __shared__ __align__(16) float smem_b[];

constexpr int kBwdThreads = 256;
constexpr int kBwdNcap = 128;   // nodes per graph held in shared memory
constexpr int kBwdEcap = 640;   // transposed-CSR slots per graph held in shared memory
constexpr int kBwdSlack = 16;   // rows of slack per graph in the global scratch (pad rows of the MMA tiles)

// dst[i] = sum_k norm_ik src[k] over the transposed CSR row of i, self-loop last; rows n..npad-1 := 0.
// A half-warp per row, a lane per 16-byte chunk.
__device__ __forceinline__ void agg_rows(const float* __restrict__ src, float* __restrict__ dst, const int32_t* trp,
                                         const int32_t* tcl, const float* ds, int n0, int n, int e0, int npad,
                                         float fill) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int half = lane >> 4, c0 = 4 * (lane & 15);
  for (int i = 2 * warp + half; i < npad; i += 2 * (kBwdThreads / 32)) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n) {
      const int beg = trp[i] - e0, end = trp[i + 1] - e0;
      const float di = ds[i];
      for (int k = beg; k < end; ++k) {
        int j = tcl[k] - n0;
        const bool ok = (unsigned)j < (unsigned)n;  // an edge leaving its graph: skipped (cannot come out of PyG collate)
        if (!ok) j = i;
        const float w = ok ? ds[j] * di : 0.f;
        const float4 v = *reinterpret_cast<const float4*>(src + (((j << 6) | swz(j)) ^ c0));
        acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y); acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
      }
      const float ws = (di * fill) * di;
      const float4 v = *reinterpret_cast<const float4*>(src + i * 64 + (c0 ^ swz(i)));
      acc.x = fmaf(ws, v.x, acc.x); acc.y = fmaf(ws, v.y, acc.y); acc.z = fmaf(ws, v.z, acc.z); acc.w = fmaf(ws, v.w, acc.w);
    }
    *reinterpret_cast<float4*>(dst + i * 64 + (c0 ^ swz(i))) = acc;
  }
}

struct BwdState {       // per-thread running sums over the graphs this CTA walks
  float w2[4][4];       // g_W2 tiles: rows 16 mt + g8 (+8), columns 8 (nb + j) + 2 t4 (+1)
  float w1[2][4];       // g_W1 tiles, same mapping over the (zero padded) 32 input columns
  float b2, b1;         // column (tid & 63), rows = tid >> 6 (mod 4)
  float wreg[8][2];     // W2 fragments of this warp's 8 output columns (B operand of P4)
};

__device__ __forceinline__ void bwd_graph(const FusedBwdArgs& a, BwdState& s, float* __restrict__ P, float* __restrict__ Q,
                                          float* __restrict__ Hx, const int32_t* trp, const int32_t* tcl,
                                          const float* ds, int* s_cnt, int n0, int n, int e0, int64_t gid) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g8 = lane >> 2, t4 = lane & 3;
  const int npad = (n + 7) & ~7;
  const int c = tid & 63, qg = tid >> 6;

  // ---- stage h2 -> P and h1 -> Hx (pad rows zero) -------------------------------------------------
  {
    const float4* h2g = reinterpret_cast<const float4*>(a.h2 + (int64_t)n0 * 64);
    const float4* h1g = reinterpret_cast<const float4*>(a.h1 + (int64_t)n0 * 64);
    for (int idx = tid; idx < npad * 16; idx += kBwdThreads) {
      const int r = idx >> 4, q = idx & 15;
      float4 v2 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v2;
      if (r < n) { v2 = __ldg(h2g + idx); v1 = __ldg(h1g + idx); }
      const int o = r * 64 + ((4 * q) ^ swz(r));
      *reinterpret_cast<float4*>(P + o) = v2;
      *reinterpret_cast<float4*>(Hx + o) = v1;
    }
  }
  const float mx = __ldg(a.emb + gid * 128 + c);
  const float gmax = __ldg(a.gemb + gid * 128 + c);
  const float gmean = __ldg(a.gemb + gid * 128 + 64 + c) / (float)n;
  __syncthreads();

  // ---- P1: pooling backward . lrelu'(h2), in place in P; bias gradient of layer 2 -----------------
  {
    int ties = 0;
    for (int r = qg; r < n; r += 4) ties += (P[at(r, c)] == mx) ? 1 : 0;
    s_cnt[qg * 64 + c] = ties;
    __syncthreads();
    ties = s_cnt[c] + s_cnt[64 + c] + s_cnt[128 + c] + s_cnt[192 + c];
    const float gm = ties > 0 ? gmax / (float)ties : 0.f;
    for (int r = qg; r < n; r += 4) {
      const int o = at(r, c);
      const float v = P[o];
      const float gg = (gmean + (v == mx ? gm : 0.f)) * lrelu_grad_from_out(v, a.slope);
      P[o] = gg;
      s.b2 += gg;
    }
  }
  __syncthreads();

  // ---- P2: gT2 = Â^T gA2 -> Q --------------------------------------------------------------------
  agg_rows(P, Q, trp, tcl, ds, n0, n, e0, npad, a.fill);
  __syncthreads();

  // ---- P3: g_W2 += gT2^T h1 ----------------------------------------------------------------------
  xty_tiles<4>(Q, Hx, npad, warp & 3, (warp >> 2) * 4, s.w2);

  // ---- P4: gA1 = (gT2 W2) . lrelu'(h1) -> P ; this warp owns output columns 8 warp .. +8 ----------
  {
    const int i0 = 8 * warp;
    const int mtiles = (n + 15) >> 4;
    // element (16 mt + g8 (+8), 8 ks + t4 (+4)) sits at row * 64 + ((8 ks) ^ v), v = t4 ^ swz(g8) (^ 4):
    // with ks = j + 4 h that is (v ^ 8 j) + 32 h — four per-thread constants and immediates
    const int v0 = t4 ^ swz(g8), v1 = v0 ^ 4;
    for (int mt = 0; mt < mtiles; ++mt) {
      const int ra = 16 * mt + g8, rb = ra + 8;
      const float* qrow = Q + ra * 64;
      float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const float* pa = qrow + (v0 ^ (8 * (ks & 3))) + 32 * (ks >> 2);
        const float* pb = qrow + (v1 ^ (8 * (ks & 3))) + 32 * (ks >> 2);
        uint32_t ah[4], al[4], bh0, bl0, bh1, bl1;
        split_tf32(pa[0], ah[0], al[0]);
        split_tf32(pa[512], ah[1], al[1]);
        split_tf32(pb[0], ah[2], al[2]);
        split_tf32(pb[512], ah[3], al[3]);
        split_tf32(s.wreg[ks][0], bh0, bl0);
        split_tf32(s.wreg[ks][1], bh1, bl1);
        mma_split(d, ah, al, bh0, bh1, bl0, bl1);
      }
      const int cc = i0 + 2 * t4;
      if (ra < n) {
        const int o = at(ra, cc);
        const float2 h = *reinterpret_cast<const float2*>(Hx + o);
        *reinterpret_cast<float2*>(P + o) =
            make_float2(d[0] * lrelu_grad_from_out(h.x, a.slope), d[1] * lrelu_grad_from_out(h.y, a.slope));
      }
      if (rb < n) {
        const int o = at(rb, cc);
        const float2 h = *reinterpret_cast<const float2*>(Hx + o);
        *reinterpret_cast<float2*>(P + o) =
            make_float2(d[2] * lrelu_grad_from_out(h.x, a.slope), d[3] * lrelu_grad_from_out(h.y, a.slope));
      }
    }
  }
  __syncthreads();

  // ---- x -> Hx (h1 is dead): columns F..31 and pad rows zero --------------------------------------
  {
    const float* xg = a.x + (int64_t)n0 * a.F;
    for (int idx = tid; idx < npad * 32; idx += kBwdThreads) {
      const int r = idx >> 5, cx = idx & 31;
      Hx[at(r, cx)] = (r < n && cx < a.F) ? __ldg(xg + r * a.F + cx) : 0.f;
    }
  }
  // ---- P5: bias gradient of layer 1 ; P6: gT1 = Â^T gA1 -> Q --------------------------------------
  for (int r = qg; r < n; r += 4) s.b1 += P[at(r, c)];
  agg_rows(P, Q, trp, tcl, ds, n0, n, e0, npad, a.fill);
  __syncthreads();

  // ---- P7: g_W1 += gT1^T x -----------------------------------------------------------------------
  xty_tiles<2>(Q, Hx, npad, warp & 3, (warp >> 2) * 2, s.w1);
  __syncthreads();  // the buffers are restaged by the next graph
}

__global__ void __launch_bounds__(kBwdThreads, 2)
k_fused_bwd(const FusedBwdArgs a) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g8 = lane >> 2, t4 = lane & 3;
  float* P = smem_b;
  float* Q = smem_b + kBwdNcap * 64;
  float* Hx = smem_b + 2 * kBwdNcap * 64;
  float* s_ds = smem_b + 3 * kBwdNcap * 64;
  int32_t* s_rp = reinterpret_cast<int32_t*>(s_ds + kBwdNcap);
  int32_t* s_cl = s_rp + (kBwdNcap + 4);
  int* s_cnt = reinterpret_cast<int*>(s_cl + kBwdEcap);

  BwdState s;
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) s.w2[j][q] = 0.f;
#pragma unroll
  for (int j = 0; j < 2; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) s.w1[j][q] = 0.f;
  s.b2 = s.b1 = 0.f;
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) {  // B[k = o][n = i] = W2[o][i]  (state_dict layout [out][in])
    s.wreg[ks][0] = __ldg(a.w2 + (8 * ks + t4) * 64 + 8 * warp + g8);
    s.wreg[ks][1] = __ldg(a.w2 + (8 * ks + t4 + 4) * 64 + 8 * warp + g8);
  }
  // shared memory starts undefined: pad rows take part in products whose results are discarded, but
  // must never hold NaN patterns where a zero operand meets them
  for (int idx = tid; idx < 3 * kBwdNcap * 64; idx += kBwdThreads) smem_b[idx] = 0.f;
  __syncthreads();

  const int64_t total = (a.only && *a.only_count == 0) ? 0 : a.B;
  for (int64_t gid = blockIdx.x; gid < total; gid += gridDim.x) {
    if (a.only && a.only[gid] == 0) continue;  // block-uniform
    const int n0 = a.graph_ptr[gid];
    const int n = a.graph_ptr[gid + 1] - n0;
    if (n <= 0) continue;  // block-uniform
    const int e0 = a.t_rowptr[n0];
    const int e = a.t_rowptr[n0 + n] - e0;
    if (n <= kBwdNcap && e <= kBwdEcap) {
      for (int idx = tid; idx <= n; idx += kBwdThreads) s_rp[idx] = __ldg(a.t_rowptr + n0 + idx);
      for (int idx = tid; idx < e; idx += kBwdThreads) s_cl[idx] = __ldg(a.t_col + e0 + idx);
      for (int idx = tid; idx < n; idx += kBwdThreads) s_ds[idx] = __ldg(a.dis + n0 + idx);
      // (the first __syncthreads of bwd_graph orders these writes before their readers)
      bwd_graph(a, s, P, Q, Hx, s_rp, s_cl, s_ds, s_cnt, n0, n, e0, gid);
    } else {
      // private slice of the global scratch: rows [n0 + 16 gid, n0 + 16 gid + n + 16)
      const int64_t row = (int64_t)n0 + (int64_t)kBwdSlack * gid;
      float* gP = a.scratch + row * 64;
      bwd_graph(a, s, gP, gP + a.scratch_stride, gP + 2 * a.scratch_stride, a.t_rowptr + n0, a.t_col + e0,
                a.dis + n0, s_cnt, n0, n, e0, gid);
    }
  }

  // ---- this CTA's partial sums -> partial[blockIdx.x][...] ----------------------------------------
  const int F = a.F;
  const int pe = 64 * 64 + 64 * F + 128;
  float* out = a.partial + (int64_t)blockIdx.x * pe;
  {
    const int mt = warp & 3, nb = (warp >> 2) * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = 16 * mt + g8, i = 8 * (nb + j) + 2 * t4;
      out[o * 64 + i] = s.w2[j][0];
      out[o * 64 + i + 1] = s.w2[j][1];
      out[(o + 8) * 64 + i] = s.w2[j][2];
      out[(o + 8) * 64 + i + 1] = s.w2[j][3];
    }
    const int nb1 = (warp >> 2) * 2;
    float* o1 = out + 64 * 64;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int o = 16 * mt + g8, i = 8 * (nb1 + j) + 2 * t4;
      if (i < F) { o1[o * F + i] = s.w1[j][0]; o1[(o + 8) * F + i] = s.w1[j][2]; }
      if (i + 1 < F) { o1[o * F + i + 1] = s.w1[j][1]; o1[(o + 8) * F + i + 1] = s.w1[j][3]; }
    }
  }
  __syncthreads();
  float* s_red = smem_b;  // [2][4][64]
  s_red[(tid >> 6) * 64 + (tid & 63)] = s.b2;
  s_red[256 + (tid >> 6) * 64 + (tid & 63)] = s.b1;
  __syncthreads();
  if (tid < 128) {
    const float* p = s_red + (tid >> 6) * 256 + (tid & 63);
    out[64 * 64 + 64 * F + tid] = ((p[0] + p[64]) + p[128]) + p[192];  // [b2 (64) | b1 (64)]
  }
}

// grads = sum over CTAs of the partials, in a fixed order (det_colsum32)
__global__ void __launch_bounds__(256)
k_fused_bwd_reduce(const float* __restrict__ partial, int ctas, int F, float* __restrict__ g_w2, float* __restrict__ g_w1,
                   float* __restrict__ g_b2, float* __restrict__ g_b1) {
  __shared__ float s_red[256];
  const int pe = 64 * 64 + 64 * F + 128;
  const int e = blockIdx.x * 32 + (threadIdx.x & 31);
  const float v = det_colsum32(partial, ctas, pe, e, e < pe, s_red);
  if (e >= pe || threadIdx.x >= 32) return;
  if (e < 64 * 64) g_w2[e] = v;
  else if (e < 64 * 64 + 64 * F) g_w1[e - 64 * 64] = v;
  else if (e < 64 * 64 + 64 * F + 64) g_b2[e - 64 * 64 - 64 * F] = v;
  else g_b1[e - 64 * 64 - 64 * F - 64] = v;
}

static size_t fused_bwd_smem_bytes() {
  return (size_t)(3 * kBwdNcap * 64 + kBwdNcap + (kBwdNcap + 4) + kBwdEcap + 256) * 4;
}

int fused_bwd_grid(int64_t B) {
  int64_t grid = 2 * (int64_t)kNumSMs;
  return (int)(grid < B ? grid : (B > 0 ? B : 1));
}

size_t fused_bwd_partial_elems(int64_t B, int F) {
  return (size_t)(fused_bwd_grid(B) + tc_bwd_grid(B)) * (size_t)(64 * 64 + 64 * F + 128);   // + the tensor-core kernel's rows
}

size_t fused_bwd_scratch_elems(int64_t N, int64_t B) { return 3 * (size_t)(N + kBwdSlack * B + kBwdSlack) * 64; }

int launch_fused_bwd(FusedBwdArgs a, int64_t N, float* g_w2, float* g_w1, float* g_b2, float* g_b1, cudaStream_t st,
                     int max_ctas, const float* reduce_base, int extra_rows) {
  if (a.B <= 0) return HCG_OK;
  const size_t smem = fused_bwd_smem_bytes();
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_fused_bwd, smem, configured)) return rc;
  a.scratch_stride = (int64_t)(N + kBwdSlack * a.B + kBwdSlack) * 64;
  int grid = fused_bwd_grid(a.B);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  k_fused_bwd<<<grid, kBwdThreads, smem, st>>>(a);
  HCG_LAUNCH_OK(st, "k_fused_bwd");
  const int pe = 64 * 64 + 64 * a.F + 128;
  k_fused_bwd_reduce<<<(unsigned)cdiv(pe, 32), 256, 0, st>>>(reduce_base ? reduce_base : a.partial, grid + extra_rows, a.F, g_w2,
                                                             g_w1, g_b2, g_b1);
  HCG_LAUNCH_OK(st, "k_fused_bwd_reduce");
  return HCG_OK;
}

}  // namespace hcg
