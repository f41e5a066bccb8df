// Graph-resident backward of the GCN hot path on the 5th-generation tensor cores (H = 64, two conv layers, F <= 32):
// the tcgen05 counterpart of hcg_fused_bwd.cu (same mathematics, same per-CTA partial layout, same reduce kernel).
//
// With z_l = (Â h_{l-1}) W_l^T + b_l, h_l = lrelu(z_l), emb = [max_i h_2 || mean_i h_2] (reference model/gcn.py:58-66)
// and g_emb = d loss / d emb, per TILE of 128 node rows holding one or two whole graphs (the forward's work list):
//
//   P1  gA2 = pool'(g_emb) . lrelu'(h2)           max routed to the arg-max rows (ties split evenly), mean / n
//       g_b2 += colsum(gA2)
//   P2  gT2 = Â^T gA2                              transposed CSR as (tile row, norm) records, self-loop last
//   P3  g_W2 += gT2^T h1        tcgen05.mma, BOTH operands MN-major from shared memory: the reduction index is the
//                               node row, which is the row of every activation tile on chip - nothing is transposed
//                               and no thread loads a fragment (hcg_tcbwd_blocks.cuh)
//   P4  gA1 = (gT2 W2) . lrelu'(h1)               tcgen05.mma, A = the gathered rows in tensor memory (as the forward)
//   P5  g_b1 += colsum(gA1)
//   P6  gT1 = Â^T gA1
//   P7  g_W1 += gT1^T x         tcgen05.mma, MN-major operands, N = 32
//
// All three products keep fp32-grade accuracy with the three-term TF32 split.  P3 and P7 run ASYNCHRONOUSLY: they are
// committed to their own mbarriers and only waited for when their operand tiles are about to be overwritten, so their
// ~2000 tensor-pipe cycles per tile hide under the gathers.  Each tile's weight-gradient product starts from a zero
// accumulator in tensor memory and is added to fp32 registers afterwards (the tensor core's accumulator truncation
// never compounds over the batch); per-CTA sums go to `partial` and are summed in CTA order by k_fused_bwd_reduce:
// no floating-point atomics, run-to-run identical.
//
// One CTA of 512 threads per SM (four threads per tile row, 16 columns each); shared memory holds four two-plane
// operand tiles (gT hi / lo, h1-or-x hi / lo), the gA tile in plain rows (pitch 68 floats), W2^T tiles, the neighbour
// records and two CSR stages: ~215 KB.
// Graphs above the tile capacity are flagged and served by the mma.sync kernel of hcg_fused_bwd.cu.
#include "hcg_fused.cuh"
#include "hcg_tc.cuh"
#include "hcg_tc_tiles.cuh"
#include "hcg_tcbwd_blocks.cuh"

namespace hcg {

#ifndef HCG_TB_THREADS
#define HCG_TB_THREADS 512
#endif
constexpr int kTbThreads = HCG_TB_THREADS;     // 256 or 512
constexpr int kTbTpr = kTbThreads / 128;        // threads per tile row
constexpr int kTbCw = 64 / kTbTpr;              // columns per thread (32 or 16): units of 8 floats inside one 32-column plane
constexpr int kTbUnits = kTbCw / 8;
constexpr int kTbPq = kTbThreads / 64;          // row groups of the column owners (P1, P5)
constexpr int kTbLd = 2048 / kTbThreads;        // 16-byte chunks of a 128 x 64 tile per thread
constexpr int kTbLx = 4096 / kTbThreads;        // elements of a 128 x 32 x tile per thread
// The gA tile (P) is no MMA operand: plain rows with a pitch of 68 floats.  A warp whose lanes are 32 consecutive rows, or
// 32 consecutive columns of one row, reads and writes it without bank conflicts, and random neighbour rows spread over
// eight 16-byte bank groups instead of the four 32-byte slots of the operand layout.
#ifndef HCG_TB_PPITCH
#define HCG_TB_PPITCH 68
#endif
constexpr int kTbPp = HCG_TB_PPITCH;
static_assert(kTbThreads == 256 || kTbThreads == 512, "k_tc_bwd: two or four threads per tile row");

#ifdef HCG_TC_TIMING
__device__ unsigned long long g_tb_timing[16];
#define TB_TICK(slot_) do { if (tid == 0 && blockIdx.x == 0) { long long t_ = clock64(); g_tb_timing[slot_] += (unsigned long long)(t_ - t_prev); t_prev = t_; } } while (0)
#else
#define TB_TICK(slot_) do { } while (0)
#endif

struct TbArgs {
  FusedBwdArgs b;
  const TcTile* tiles;      // [grid][cap] dense per-CTA work lists (k_tile_desc over the TRANSPOSED CSR)
  const int* counts;        // [grid]
  int cap;
  int* status;
};

struct TbSmem {   // byte offsets from the 1024-aligned base
  uint32_t q_hi, q_lo, hx_hi, hx_lo, p, w_hi, w_lo, nbr, cst, misc;
  uint32_t rw, cw, dw, csr_words, total;
};
__host__ __device__ inline TbSmem tb_layout() {
  TbSmem y;
  uint32_t o = 0;
  // order matters: an MN-major A operand with M = 128 reads two more planes behind its tile (don't-care rows), so
  // every gT tile must be followed by initialised shared memory
  y.q_hi = o; o += 2 * tcb::kPlaneBytes;
  y.q_lo = o; o += 2 * tcb::kPlaneBytes;
  y.hx_hi = o; o += 2 * tcb::kPlaneBytes;
  y.hx_lo = o; o += 2 * tcb::kPlaneBytes;
  y.p = o; o += (uint32_t)kTcRows * kTbPp * 4;
  y.w_hi = o; o += tc::kN * 64 * 4;
  y.w_lo = o; o += tc::kN * 64 * 4;
  y.nbr = o; o += (uint32_t)kTcEcap * 8;
  y.rw = ((uint32_t)kTcRows + 2 + 16 + 3) & ~3u;
  y.cw = ((uint32_t)kTcEcap + 16 + 3) & ~3u;
  y.dw = ((uint32_t)kTcRows + 16 + 3) & ~3u;
  y.csr_words = y.rw + y.cw + y.dw;
  y.cst = o; o += 2 * y.csr_words * 4;
  y.misc = o; o += (2 * 3 * 64 + 2 * kTbPq * 64 + 2 * kTbPq * 64) * 4;   // [gb][mx | gmax | gmean/n][64], tie counts [gb][kTbPq][64], column sums [2][kTbPq][64]
  y.total = o + 1024;
  return y;
}

// W: global [64][64] (state_dict layout [out o][in i]).  Tiles of W^T for  D[r][i] = sum_o A[r][o] W[o][i]:
// tile element (n = i, k = o) = W[o][i].
__device__ __forceinline__ void prep_weight_tiles_t(const float* __restrict__ W, float* s_hi, float* s_lo, int tid, int nthreads) {
  constexpr int U = 4;   // loads of a thread in flight at once
  for (int base = tid; base < 64 * 64; base += U * nthreads) {
    float v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = base + u * nthreads < 64 * 64 ? __ldg(W + base + u * nthreads) : 0.f;   // coalesced read of W[o][i]
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads, o = idx >> 6, i = idx & 63;
      if (idx >= 64 * 64) break;
      const uint32_t hi = tc::tf32_hi(v[u]), lo = tc::tf32_lo(v[u], hi);
      const int off = tc::w_tile_offset(i, o);
      s_hi[off] = __uint_as_float(hi);
      s_lo[off] = __uint_as_float(lo);
    }
  }
}

__global__ void __launch_bounds__(kTbThreads, 1)
k_tc_bwd(const TbArgs a) {
#ifdef HCG_TC_TIMING
  long long t_prev = clock64();
#endif
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage[2];
  __shared__ uint64_t bar_g, bar_w2, bar_w1;
  __shared__ uint32_t tmem_base_s;
  __shared__ TcMeta metas[2];
  __shared__ int4 recs[2][3];

  const int tid = threadIdx.x, lane = tid & 31, wslot = tid >> 5;
  const int cg = tid >> 7;            // column group: columns kTbCw cg .. + kTbCw - 1
  const int pl = (cg * kTbCw) >> 5;   // ... which sit in plane pl of the two-plane tiles,
  const int ub = ((cg * kTbCw) & 31) >> 3;   // ... from 8-float unit ub on
  const bool hswap = ((tid >> 2) & 1) != 0;  // row-owner accesses of operand tiles: 16-byte halves in swapped order
  const int i = tid & 127;            // tile row
  const int i_warp_min = tid & 96;
  const bool ctl_warp = wslot == 3, pf_warp = wslot == 7;
  const int F = a.b.F;
  const float slope = a.b.slope;

  const TbSmem lay = tb_layout();
  unsigned char* base = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  float* smf = reinterpret_cast<float*>(base);
  int32_t* smi = reinterpret_cast<int32_t*>(base);
  float* Qh = reinterpret_cast<float*>(base + lay.q_hi);
  float* Ql = reinterpret_cast<float*>(base + lay.q_lo);
  float* Hh = reinterpret_cast<float*>(base + lay.hx_hi);
  float* Hl = reinterpret_cast<float*>(base + lay.hx_lo);
  float* P = reinterpret_cast<float*>(base + lay.p);
  float* w_hi = reinterpret_cast<float*>(base + lay.w_hi);
  float* w_lo = reinterpret_cast<float*>(base + lay.w_lo);
  const int nb_w = (int)(lay.nbr >> 2), cs_w0 = (int)(lay.cst >> 2);
  const int rw = (int)lay.rw, cw = (int)lay.cw, csw = (int)lay.csr_words;
  float* s_pool = reinterpret_cast<float*>(base + lay.misc);          // [gb][3][64]
  int* s_cnt = reinterpret_cast<int*>(s_pool + 2 * 3 * 64);           // [gb][kTbPq][64]
  float* s_red = reinterpret_cast<float*>(s_cnt + 2 * kTbPq * 64);    // [2][kTbPq][64]

  // ---- one-time setup: every byte of shared memory initialised (don't-care operand rows must never hold NaN) ----
  for (int idx = tid; idx < (int)(lay.total - 1024) / 4; idx += kTbThreads) smf[idx] = 0.f;
  if (tid < 32) ptx::tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) {
    ptx::mbar_init(&bar_stage[0], 6);   // the six feeding lanes (rowptr, col, dis of up to two graphs)
    ptx::mbar_init(&bar_stage[1], 6);
    ptx::mbar_init(&bar_g, 1);
    ptx::mbar_init(&bar_w2, 1);
    ptx::mbar_init(&bar_w1, 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  prep_weight_tiles_t(a.b.w2, w_hi, w_lo, tid, kTbThreads);
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();

  const uint32_t tb = tmem_base_s;
  const uint32_t d_g = tb, a_hi = tb + 64, a_lo = tb + 128, d_w2 = tb + 192, d_w1 = tb + 256;
  const uint32_t trow = (uint32_t)((wslot & 3) * 32) << 16;
  const bool wrow = (wslot & 3) < 2;    // this thread's TMEM lane (= i) is one of the 64 weight-gradient rows

  const int ntiles = a.counts[blockIdx.x];
  const TcTile* my_tiles = a.tiles + (size_t)blockIdx.x * a.cap;

  auto fetch_rec = [&](int t) {
    if (t < ntiles && lane < 3) ptx::cp_async16(&recs[t & 1][lane], reinterpret_cast<const int4*>(my_tiles + t) + lane);
  };
  // lanes 1..3 / 5..7 of pf_warp: lane = 4 gb + q copies slice q (1 t_rowptr, 2 t_col, 3 dis) of graph gb into stage t & 1
  auto issue = [&](int t) {
    const int s = t & 1;
    TcMeta* m = &metas[s];
    uint64_t* bar = &bar_stage[s];
    ptx::cp_async_wait_all();
    __syncwarp();
    if (lane >= 8 || (lane & 3) == 0) return;
    if (t >= ntiles) {
      if (lane == 1) m->gA = -1;
      ptx::mbar_arrive(bar);
      return;
    }
    const int4 r0 = recs[s][0], r1 = recs[s][1], r2 = recs[s][2];
    const int gb = lane >> 2, q = lane & 3;
    const int n0 = gb ? r1.y : r0.y, n = gb ? r1.z : r0.z, e0 = gb ? r1.w : r0.w, e = gb ? r2.y : r2.x;
    const bool have = gb == 0 || r1.x >= 0;
    const bool manual = (r2.z & kTileManual) != 0;
    uintptr_t pA, pS;
    int cA, cS;
    if (q == 1) { pA = reinterpret_cast<uintptr_t>(a.b.t_rowptr + r0.y); cA = r0.z + 1;
                  pS = reinterpret_cast<uintptr_t>(a.b.t_rowptr + n0); cS = n + 1; }
    else if (q == 2) { pA = reinterpret_cast<uintptr_t>(a.b.t_col + r0.w); cA = r2.x;
                       pS = reinterpret_cast<uintptr_t>(a.b.t_col + e0); cS = e; }
    else { pA = reinterpret_cast<uintptr_t>(a.b.dis + r0.y); cA = r0.z;
           pS = reinterpret_cast<uintptr_t>(a.b.dis + n0); cS = n; }
    uintptr_t loA, loS;
    uint32_t bA, bS;
    int leadA, leadS;
    copy_window(pA, cA, &loA, &bA, &leadA);
    copy_window(pS, cS, &loS, &bS, &leadS);
    if (manual) { leadA = 0; leadS = 0; bA = (uint32_t)((cA * 4 + 15) & ~15); }
    const int arr_w = cs_w0 + s * csw + (q == 1 ? 0 : q == 2 ? rw : rw + cw);
    const int dst_w = arr_w + (gb ? (int)(bA >> 2) : 0);
    if (lane == 1) {
      m->gA = r0.x; m->n0A = r0.y; m->nA = r0.z; m->e0A = r0.w;
      m->gB = r1.x; m->n0B = r1.y; m->nB = r1.x >= 0 ? r1.z : 0; m->e0B = r1.w;
      m->eA = r2.x; m->eB = r2.y; m->flags = r2.z;
    }
    m->off[lane] = (gb ? (int)(bA >> 2) : 0) + leadS;
    if (!have || manual) {
      ptx::mbar_arrive(bar);
    } else {
      ptx::mbar_expect_tx(bar, bS);
      ptx::bulk_g2s(smf + dst_w, reinterpret_cast<const void*>(loS), bS, bar);
    }
  };

  // per-thread running sums over the tiles of this CTA
  float acc_w2[kTbCw], acc_w1[16];      // (g_W1: the column groups below 32 only)
#pragma unroll
  for (int q = 0; q < kTbCw; ++q) acc_w2[q] = 0.f;
#pragma unroll
  for (int q = 0; q < 16; ++q) acc_w1[q] = 0.f;
  float b2acc = 0.f, b1acc = 0.f;      // column (tid & 63), rows (tid >> 6) mod kTbPq
  const int pc = tid & 63, pq = tid >> 6;

  // adds the finished weight-gradient accumulators of a tile to the register sums (rows 0..63 = TMEM lanes 0..63)
  auto take_w2 = [&]() {
    if (wrow) {
#pragma unroll
      for (int part = 0; part < kTbCw / 16; ++part) {
        uint32_t r0[16];
        ptx::tmem_ld16(d_w2 + trow + cg * kTbCw + part * 16, r0);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 16; ++q) acc_w2[part * 16 + q] += __uint_as_float(r0[q]);
      }
    }
  };
  auto take_w1 = [&]() {
    if (wrow && cg < 2) {
      uint32_t r0[16];
      ptx::tmem_ld16(d_w1 + trow + cg * 16, r0);
      ptx::tmem_wait_ld();
#pragma unroll
      for (int q = 0; q < 16; ++q) acc_w1[q] += __uint_as_float(r0[q]);
    }
  };

  if (pf_warp) {
    fetch_rec(0);
    issue(0);
    fetch_rec(1);
  }
  uint32_t g_phase = 0, w2_phase = 0, w1_phase = 0;
  bool w1_pending = false;
  for (int k = 0;; ++k) {
    const int s = k & 1;
    TB_TICK(0);
    ptx::mbar_wait(&bar_stage[s], (uint32_t)(k >> 1) & 1u);
    TB_TICK(1);
    const TcMeta* mp = &metas[s];
    const int gA = mp->gA;
    if (gA < 0) break;
    if (pf_warp) {
      issue(k + 1);
      fetch_rec(k + 2);
    }
    const int nA = mp->nA, nB = mp->nB, gB = mp->gB, eA = mp->eA, n0A = mp->n0A, n0B = mp->n0B;
    const int ntot = nA + nB;
    const int ksteps = (ntot + 7) >> 3;
    const int srp = cs_w0 + s * csw, scl = srp + rw, sds = scl + cw;
    if (mp->flags & kTileManual) {
      for (int gb = 0; gb < (nB > 0 ? 2 : 1); ++gb) {
        const int n0 = gb ? n0B : n0A, n = gb ? nB : nA, e0 = gb ? mp->e0B : mp->e0A, e = gb ? mp->eB : eA;
        const int o1 = mp->off[4 * gb + 1], o2 = mp->off[4 * gb + 2], o3 = mp->off[4 * gb + 3];
        for (int idx = tid; idx <= n; idx += kTbThreads) smi[srp + o1 + idx] = __ldg(a.b.t_rowptr + n0 + idx);
        for (int idx = tid; idx < e; idx += kTbThreads) smi[scl + o2 + idx] = __ldg(a.b.t_col + e0 + idx);
        for (int idx = tid; idx < n; idx += kTbThreads) smf[sds + o3 + idx] = __ldg(a.b.dis + n0 + idx);
      }
    }
    // ---- h2 -> P (16-byte chunks, coalesced), per-graph pooling data -> shared memory -------------------------------
    {   // all (up to eight) loads of a thread in flight at once: one memory latency per tile, not eight
      float4 v[kTbLd];
#pragma unroll
      for (int u = 0; u < kTbLd; ++u) {
        const int idx = tid + u * kTbThreads, r = idx >> 4, q = idx & 15;
        const int64_t grow = r < nA ? (int64_t)n0A + r : (int64_t)n0B + (r - nA);
        if (r < ntot) v[u] = ldg4(a.b.h2 + grow * 64 + 4 * q);
      }
#pragma unroll
      for (int u = 0; u < kTbLd; ++u) {
        const int idx = tid + u * kTbThreads, r = idx >> 4, q = idx & 15;
        if (r < ntot) *reinterpret_cast<float4*>(P + r * kTbPp + 4 * q) = v[u];
      }
    }
    if (tid < 128) {
      const int gb = tid >> 6, c = tid & 63;
      if (gb == 0 || nB > 0) {
        const int64_t g = gb ? gB : gA;
        const float n = (float)(gb ? nB : nA);
        s_pool[(gb * 3 + 0) * 64 + c] = __ldg(a.b.emb + g * 128 + c);
        s_pool[(gb * 3 + 1) * 64 + c] = __ldg(a.b.gemb + g * 128 + c);
        s_pool[(gb * 3 + 2) * 64 + c] = __ldg(a.b.gemb + g * 128 + 64 + c) / n;
      }
    }
    TB_TICK(2);
    __syncthreads();   // (also orders the manual loads above before their readers)
    TB_TICK(3);

    // h1 of this tile: the loads go out now and are consumed after P1 (their latency hides under the records and P1)
    float4 h1v[kTbLd];
#pragma unroll
    for (int u = 0; u < kTbLd; ++u) {
      const int idx = tid + u * kTbThreads, r = idx >> 4, q = idx & 15;
      const int64_t grow = r < nA ? (int64_t)n0A + r : (int64_t)n0B + (r - nA);
      if (r < ntot) h1v[u] = ldg4(a.b.h1 + grow * 64 + 4 * q);
    }
    // ---- this thread's row + its transposed-CSR records ---------------------------------------------------------------
    const bool row_ok = i < ntot;
    const bool warp_ok = i_warp_min < ntot;
    const bool inB = row_ok && i >= nA;
    const int n0 = inB ? n0B : n0A, ng = inB ? nB : nA, e0 = inB ? mp->e0B : mp->e0A;
    const int rowbase = inB ? nA : 0;
    const int li = i - rowbase, jb = n0 - rowbase;
    const int nbo = nb_w + 2 * ((inB ? eA : 0) - e0);
    int beg = 0, deg = 0;
    float di = 0.f;
    {
      const int rp_o = srp + (inB ? mp->off[5] : mp->off[1]);
      const int cl_o = scl + (inB ? mp->off[6] : mp->off[2]) - e0;
      const int ds_o = sds + (inB ? mp->off[7] : mp->off[3]) - n0;
      if (row_ok) { beg = smi[rp_o + li]; deg = smi[rp_o + li + 1] - beg; di = smf[ds_o + n0 + li]; }
      for (int t = cg; t < deg; t += kTbTpr) {
        const int j = smi[cl_o + beg + t];
        const bool ok = (unsigned)(j - n0) < (unsigned)ng;
        if (!ok && a.status) atomicOr(a.status, kStatusCrossEdgeTc);
        const float w = ok ? smf[ds_o + (ok ? j : n0)] * di : 0.f;
        *reinterpret_cast<int2*>(smi + nbo + 2 * (beg + t)) = make_int2(ok ? j - jb : i, __float_as_int(w));
      }
    }
    const float wself = (di * a.b.fill) * di;
    const int iself = row_ok ? i : 0;

    // ---- P1: pooling backward . lrelu'(h2), in place in P; bias gradient of layer 2 (column owners) ---------------
    // (column owner pc: element (r, pc) sits at pcol[r kTbPp]; a warp = 32 consecutive columns of one row)
    float* const pcol = P + pc;
    for (int gb = 0; gb < (nB > 0 ? 2 : 1); ++gb) {
      const int rb = gb ? nA : 0, n = gb ? nB : nA;
      const float mx = s_pool[(gb * 3 + 0) * 64 + pc];
      int ties = 0, r = rb + pq;
      for (; r + 3 * kTbPq < rb + n; r += 4 * kTbPq) {
        const float v0 = pcol[r * kTbPp], v1 = pcol[(r + kTbPq) * kTbPp], v2 = pcol[(r + 2 * kTbPq) * kTbPp], v3 = pcol[(r + 3 * kTbPq) * kTbPp];
        ties += (v0 == mx) + (v1 == mx) + (v2 == mx) + (v3 == mx);
      }
      for (; r < rb + n; r += kTbPq) ties += (pcol[r * kTbPp] == mx) ? 1 : 0;
      s_cnt[(gb * kTbPq + pq) * 64 + pc] = ties;
    }
    __syncthreads();   // tie counts complete; the records above are visible
    for (int gb = 0; gb < (nB > 0 ? 2 : 1); ++gb) {
      const int rb = gb ? nA : 0, n = gb ? nB : nA;
      const float mx = s_pool[(gb * 3 + 0) * 64 + pc], gmax = s_pool[(gb * 3 + 1) * 64 + pc], gmean = s_pool[(gb * 3 + 2) * 64 + pc];
      int ties = 0;
#pragma unroll
      for (int u = 0; u < kTbPq; ++u) ties += s_cnt[(gb * kTbPq + u) * 64 + pc];
      const float gm = ties > 0 ? gmax / (float)ties : 0.f;
      int r = rb + pq;
      for (; r + kTbPq < rb + n; r += 2 * kTbPq) {
        const float v0 = pcol[r * kTbPp], v1 = pcol[(r + kTbPq) * kTbPp];
        const float g0 = (gmean + (v0 == mx ? gm : 0.f)) * lrelu_grad_from_out(v0, slope);
        const float g1 = (gmean + (v1 == mx ? gm : 0.f)) * lrelu_grad_from_out(v1, slope);
        pcol[r * kTbPp] = g0; pcol[(r + kTbPq) * kTbPp] = g1;
        b2acc += g0; b2acc += g1;
      }
      for (; r < rb + n; r += kTbPq) {
        const float v = pcol[r * kTbPp];
        const float gg = (gmean + (v == mx ? gm : 0.f)) * lrelu_grad_from_out(v, slope);
        pcol[r * kTbPp] = gg;
        b2acc += gg;
      }
    }
    TB_TICK(4);
    // ---- the previous tile's P7 (g_W1 += gT1^T x) has to be over before its operand tiles are refilled -----------------
    if (w1_pending) {
      ptx::mbar_wait(&bar_w1, w1_phase);
      w1_phase ^= 1u;
      ptx::tc_fence_after_sync();
      take_w1();
    }
    TB_TICK(5);
    // ---- h1 -> Hh (raw) and Hl (residual of the TF32 truncation): the B operand of P3, and lrelu'(h1) for P4 -------
    {
#pragma unroll
      for (int u = 0; u < kTbLd; ++u) {
        const int idx = tid + u * kTbThreads, r = idx >> 4, q = idx & 15;
        if (r < ntot) {
          const float4 v = h1v[u];
          const int o = tcb::plane_chunk_off(r, q);
          *reinterpret_cast<float4*>(Hh + o) = v;
          *reinterpret_cast<float4*>(Hl + o) = make_float4(v.x - tcb::tf32_trunc(v.x), v.y - tcb::tf32_trunc(v.y),
                                                           v.z - tcb::tf32_trunc(v.z), v.w - tcb::tf32_trunc(v.w));
        }
      }
    }
    TB_TICK(6);
    __syncthreads();   // gA2 complete in P
    TB_TICK(7);

    const int maxdeg = __reduce_max_sync(0xffffffffu, deg);
    const int2* nbp = reinterpret_cast<const int2*>(smi + nbo) + beg;
    auto nb_at = [&](int t) { return t < deg ? nbp[t] : make_int2(iself, 0); };
    // gT = Â^T g over the records: this thread's kTbCw columns (units ub .. of plane pl) of row i; self-loop last
    auto gather_plane = [&](float (&acc)[kTbCw]) {
#pragma unroll
      for (int c = 0; c < kTbCw; ++c) acc[c] = 0.f;
      if (!warp_ok) return;
      const float* src = P + cg * kTbCw;
      auto row_fma = [&](int r, float w) {
        const float* row = src + r * kTbPp;
#pragma unroll
        for (int u = 0; u < kTbUnits; ++u) {
          const float4 v0 = *reinterpret_cast<const float4*>(row + 8 * u);
          const float4 v1 = *reinterpret_cast<const float4*>(row + 8 * u + 4);
          acc[8 * u + 0] = fmaf(w, v0.x, acc[8 * u + 0]); acc[8 * u + 1] = fmaf(w, v0.y, acc[8 * u + 1]);
          acc[8 * u + 2] = fmaf(w, v0.z, acc[8 * u + 2]); acc[8 * u + 3] = fmaf(w, v0.w, acc[8 * u + 3]);
          acc[8 * u + 4] = fmaf(w, v1.x, acc[8 * u + 4]); acc[8 * u + 5] = fmaf(w, v1.y, acc[8 * u + 5]);
          acc[8 * u + 6] = fmaf(w, v1.z, acc[8 * u + 6]); acc[8 * u + 7] = fmaf(w, v1.w, acc[8 * u + 7]);
        }
      };
      int2 nb = nb_at(0);
#pragma unroll 1
      for (int t = 0; t < maxdeg; ++t) {
        const int2 cur = nb;
        nb = nb_at(t + 1);
        if (t < deg) row_fma(cur.x, __int_as_float(cur.y));
      }
      row_fma(iself, wself);
    };
    // row i, this thread's units of an operand tile pair (raw, residual) <- acc
    auto store_planes = [&](const float (&acc)[kTbCw]) {
      if (!warp_ok) return;
      float* qh = Qh + pl * tcb::kPlaneFloats + i * 32;
      float* ql = Ql + pl * tcb::kPlaneFloats + i * 32;
      // rows i and i + 4 of a quarter warp share a 32-byte slot of the operand layout: the upper one writes the two
      // 16-byte halves in the opposite order, so the eight lanes of a store cover eight different bank groups
#pragma unroll
      for (int u = 0; u < kTbUnits; ++u) {
        const int o = ((ub + u) ^ (i & 3)) << 3;
        const float4 lo4 = make_float4(acc[8 * u], acc[8 * u + 1], acc[8 * u + 2], acc[8 * u + 3]);
        const float4 hi4 = make_float4(acc[8 * u + 4], acc[8 * u + 5], acc[8 * u + 6], acc[8 * u + 7]);
        const float4 f0 = hswap ? hi4 : lo4, f1 = hswap ? lo4 : hi4;
        const int o0 = o + (hswap ? 4 : 0), o1 = o + (hswap ? 0 : 4);
        *reinterpret_cast<float4*>(qh + o0) = f0;
        *reinterpret_cast<float4*>(qh + o1) = f1;
        *reinterpret_cast<float4*>(ql + o0) = make_float4(f0.x - tcb::tf32_trunc(f0.x), f0.y - tcb::tf32_trunc(f0.y),
                                                          f0.z - tcb::tf32_trunc(f0.z), f0.w - tcb::tf32_trunc(f0.w));
        *reinterpret_cast<float4*>(ql + o1) = make_float4(f1.x - tcb::tf32_trunc(f1.x), f1.y - tcb::tf32_trunc(f1.y),
                                                          f1.z - tcb::tf32_trunc(f1.z), f1.w - tcb::tf32_trunc(f1.w));
      }
    };

    // ---- P2: gT2 = Â^T gA2 -> Q tiles (P3's A operand) and tensor memory (P4's A operand) ---------------------------
    {
      float acc[kTbCw];
      gather_plane(acc);
      store_planes(acc);
      if (warp_ok) {
#pragma unroll
        for (int cc = 0; cc < kTbCw; cc += 16) {
          float v[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) v[q] = acc[cc + q];
          tc::store_split16(v, a_hi, a_lo, trow, cg * kTbCw + cc);
        }
      }
    }
    TB_TICK(8);
    ptx::fence_proxy_async_smem();   // the Q / H tiles were written by ordinary stores; the MMAs read them through the async proxy
    ptx::tmem_wait_st();
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (ctl_warp && lane == 0) {
      ptx::tc_fence_after_sync();
      tc::issue_split_mma<8>(d_g, a_hi, a_lo, ptx::smem_u32(w_hi), ptx::smem_u32(w_lo));
      ptx::mma_commit(&bar_g);
    }
    if (ctl_warp) { __syncwarp(); ptx::mbar_wait(&bar_g, g_phase); }
    g_phase ^= 1u;
    __syncthreads();
    ptx::tc_fence_after_sync();
    if (ctl_warp && lane == 0) {   // P3 goes out behind P4: it runs under the epilogue and nobody waits for its issue
      tcb::issue_xty<64>(d_w2, ptx::smem_u32(Qh), ptx::smem_u32(Ql), ptx::smem_u32(Hh), ptx::smem_u32(Hl), ksteps, 0u);
      ptx::mma_commit(&bar_w2);
    }
    TB_TICK(9);

    // ---- P4 epilogue: gA1 = (gT2 W2) . lrelu'(h1) -> P (own row: every gather from P is over) -------------------------
    if (warp_ok) {
      float* prow = P + i * kTbPp + cg * kTbCw;
      const float* hrow = Hh + pl * tcb::kPlaneFloats + i * 32;
#pragma unroll
      for (int part = 0; part < kTbCw / 16; ++part) {
        uint32_t q0[16];
        ptx::tmem_ld16(d_g + trow + cg * kTbCw + part * 16, q0);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int o = (((ub + 2 * part + u) ^ (i & 3)) << 3);
          const float4 ha = *reinterpret_cast<const float4*>(hrow + o + (hswap ? 4 : 0));   // (half order: see store_planes)
          const float4 hb = *reinterpret_cast<const float4*>(hrow + o + (hswap ? 0 : 4));
          const float4 h0 = hswap ? hb : ha, h1v = hswap ? ha : hb;
          float4 g0, g1;
          g0.x = __uint_as_float(q0[8 * u + 0]) * lrelu_grad_from_out(h0.x, slope);
          g0.y = __uint_as_float(q0[8 * u + 1]) * lrelu_grad_from_out(h0.y, slope);
          g0.z = __uint_as_float(q0[8 * u + 2]) * lrelu_grad_from_out(h0.z, slope);
          g0.w = __uint_as_float(q0[8 * u + 3]) * lrelu_grad_from_out(h0.w, slope);
          g1.x = __uint_as_float(q0[8 * u + 4]) * lrelu_grad_from_out(h1v.x, slope);
          g1.y = __uint_as_float(q0[8 * u + 5]) * lrelu_grad_from_out(h1v.y, slope);
          g1.z = __uint_as_float(q0[8 * u + 6]) * lrelu_grad_from_out(h1v.z, slope);
          g1.w = __uint_as_float(q0[8 * u + 7]) * lrelu_grad_from_out(h1v.w, slope);
          if (!row_ok) { g0 = make_float4(0.f, 0.f, 0.f, 0.f); g1 = g0; }
          *reinterpret_cast<float4*>(prow + 8 * (2 * part + u)) = g0;
          *reinterpret_cast<float4*>(prow + 8 * (2 * part + u) + 4) = g1;
        }
      }
    }
    TB_TICK(10);
    ptx::tc_fence_before_sync();
    __syncthreads();   // gA1 complete in P

    // x of this tile: the loads go out now and are consumed once P3 has released the H tiles
    float xv[kTbLx];
#pragma unroll
    for (int u = 0; u < kTbLx; ++u) {
      const int idx = tid + u * kTbThreads, r = idx >> 5, c = idx & 31;
      const int64_t grow = r < nA ? (int64_t)n0A + r : (int64_t)n0B + (r - nA);
      xv[u] = (r < ntot && c < F) ? __ldg(a.b.x + grow * F + c) : 0.f;
    }
    // ---- P5: bias gradient of layer 1 (column owners) ----------------------------------------------------------------
    {
      int r = pq;
      float s0 = 0.f, s1 = 0.f;
      for (; r + 3 * kTbPq < ntot; r += 4 * kTbPq) {
        s0 += pcol[r * kTbPp]; s1 += pcol[(r + kTbPq) * kTbPp]; s0 += pcol[(r + 2 * kTbPq) * kTbPp]; s1 += pcol[(r + 3 * kTbPq) * kTbPp];
      }
      for (; r < ntot; r += kTbPq) s0 += pcol[r * kTbPp];
      b1acc += s0 + s1;
    }

    TB_TICK(11);
    // ---- P3 must be over before its operand tiles are refilled; its accumulator -> registers --------------------------
    ptx::mbar_wait(&bar_w2, w2_phase);
    w2_phase ^= 1u;
    ptx::tc_fence_after_sync();
    take_w2();

    TB_TICK(12);
    // ---- x -> Hh plane 0 (columns >= F zero) and Hl: the B operand of P7 -----------------------------------------------
#pragma unroll
    for (int u = 0; u < kTbLx; ++u) {
      const int idx = tid + u * kTbThreads, r = idx >> 5, c = idx & 31;
      if (r < ntot) {
        const int o = tcb::plane_off(r, c);
        Hh[o] = xv[u];
        Hl[o] = xv[u] - tcb::tf32_trunc(xv[u]);
      }
    }
    TB_TICK(13);
    // ---- P6: gT1 = Â^T gA1 -> Q tiles --------------------------------------------------------------------------------
    {
      float acc[kTbCw];
      gather_plane(acc);
      store_planes(acc);
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    // ---- P7: g_W1 (this tile) = gT1^T x, asynchronous ----------------------------------------------------------------
    if (ctl_warp && lane == 0) {
      ptx::tc_fence_after_sync();
      tcb::issue_xty<32>(d_w1, ptx::smem_u32(Qh), ptx::smem_u32(Ql), ptx::smem_u32(Hh), ptx::smem_u32(Hl), ksteps, 0u);
      ptx::mma_commit(&bar_w1);
    }
    w1_pending = true;
    TB_TICK(14);
  }
  if (w1_pending) {
    ptx::mbar_wait(&bar_w1, w1_phase);
    ptx::tc_fence_after_sync();
    take_w1();
  }

  // ---- this CTA's partial sums -> partial[blockIdx.x][g_W2 (64 x 64) | g_W1 (64 x F) | g_b2 (64) | g_b1 (64)] ---------
  const int pe = 64 * 64 + 64 * F + 128;
  float* out = a.b.partial + (int64_t)blockIdx.x * pe;
  if (wrow) {
    float* o2 = out + i * 64 + cg * kTbCw;
#pragma unroll
    for (int q = 0; q < kTbCw; q += 4) *reinterpret_cast<float4*>(o2 + q) = make_float4(acc_w2[q], acc_w2[q + 1], acc_w2[q + 2], acc_w2[q + 3]);
    float* o1 = out + 64 * 64 + i * F + cg * 16;
#pragma unroll
    for (int q = 0; q < 16; ++q)
      if (cg < 2 && cg * 16 + q < F) o1[q] = acc_w1[q];
  }
  __syncthreads();
  s_red[pq * 64 + pc] = b2acc;
  s_red[kTbPq * 64 + pq * 64 + pc] = b1acc;
  __syncthreads();
  if (tid < 128) {
    const float* p = s_red + (tid >> 6) * kTbPq * 64 + (tid & 63);
    float sum = p[0];
#pragma unroll
    for (int u = 1; u < kTbPq; ++u) sum += p[u * 64];
    out[64 * 64 + 64 * F + tid] = sum;  // [b2 (64) | b1 (64)]
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (tid < 32) ptx::tmem_dealloc(tmem_base_s, 512);
}

int tc_bwd_grid(int64_t B) { return tc_plan(B, 1).grid; }

size_t tc_bwd_workspace_bytes(int64_t B) {
  Carver c(nullptr, (size_t)-1);
  c.take<TcTile>(tc_tiles_capacity(B));
  c.take<int>((size_t)kNumSMs + 1);
  c.take<int>(64);
  c.take<int>((size_t)B + 1);
  c.take<int>((size_t)B + 1);
  return c.off;
}

// `a.partial` has room for tc_bwd_grid(B) + fused_bwd_grid(B) rows of 64*64 + 64*F + 128 floats: the tensor-core CTAs
// write the first rows, the overflow kernel the rest; both are summed by k_fused_bwd_reduce.
int launch_tc_bwd(FusedBwdArgs a, int64_t N, int64_t E, void* ws, bool all_fit, float* g_w2, float* g_w1, float* g_b2, float* g_b1,
                  cudaStream_t st, cudaStream_t desc_st, cudaEvent_t desc_ev, const int64_t* node_ptr64, const int64_t* edge_ptr64) {
  if (a.B <= 0) return HCG_OK;
  Carver c(ws, (size_t)-1);
  TcTile* tiles = c.take<TcTile>(tc_tiles_capacity(a.B));
  int* counts = c.take<int>((size_t)kNumSMs + 1);
  int* overflow_count = c.take<int>(64);
  int* overflow_list = c.take<int>((size_t)a.B + 1);
  int* overflow_flag = c.take<int>((size_t)a.B + 1);
  const TcPlan p = tc_plan(a.B, 1);
  // desc_st / desc_ev: the work list is built on another stream (the one that produced the transposed CSR), `st` joins it.
  // With the batch's int64 offset vectors the work list needs nothing of the collate: it goes out on `st` BEFORE `st`
  // waits for the other stream (desc_ev), i.e. beside the collate instead of behind it.
  const bool ptr64 = node_ptr64 != nullptr && edge_ptr64 != nullptr;
  const bool side = !ptr64 && desc_st != nullptr && desc_ev != nullptr;
  const int rc_desc = launch_tile_desc(a.graph_ptr, a.t_rowptr, a.t_col, a.dis, a.x, a.F, N, E, a.B, p, tiles, counts, overflow_count,
                                       overflow_list, side ? desc_st : st, overflow_flag, ptr64 ? node_ptr64 : nullptr,
                                       ptr64 ? edge_ptr64 : nullptr);
  if (side) {
    HCG_CUDA_OK(cudaEventRecord(desc_ev, desc_st));
    HCG_CUDA_OK(cudaStreamWaitEvent(st, desc_ev, 0));
  } else if (desc_ev != nullptr) {
    HCG_CUDA_OK(cudaStreamWaitEvent(st, desc_ev, 0));
  }
  if (rc_desc) return rc_desc;
  const size_t smem = tb_layout().total;
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_tc_bwd, smem, configured)) return rc;
  TbArgs t{};
  t.b = a;
  t.tiles = tiles; t.counts = counts; t.cap = p.cap; t.status = nullptr;
  k_tc_bwd<<<(unsigned)p.grid, kTbThreads, smem, st>>>(t);
  HCG_LAUNCH_OK(st, "k_tc_bwd");
  // graphs above the tile capacity: the mma.sync kernel walks the overflow list, its CTAs write the following rows
  FusedBwdArgs o = a;
  o.only = overflow_flag;
  o.only_count = overflow_count;
  o.partial = a.partial + (size_t)p.grid * (64 * 64 + 64 * a.F + 128);
  return launch_fused_bwd(o, N, g_w2, g_w1, g_b2, g_b1, st, all_fit ? 1 : 0, a.partial, p.grid);
}

}  // namespace hcg

#ifdef HCG_TC_TIMING
extern "C" int hcg_debug_tb_timing(unsigned long long* out16, int reset) {
  if (out16) cudaMemcpyFromSymbol(out16, hcg::g_tb_timing, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(hcg::g_tb_timing, z, sizeof(z)); }
  return 0;
}
#endif
