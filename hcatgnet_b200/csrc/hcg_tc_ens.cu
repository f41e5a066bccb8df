// One-pass ensemble forward ("run all models over a screening set in one pass, weights batched
// in shared memory"; the reference loops `predict_network` over the trained models one after the other,
// scripts_experiments/predict_test.py:47-70 -> utils/utils_model.py:82-111).
//
// k_tc_fwd_ens walks the batch ONCE: a slot stages a tile's x rows and edge slices, builds the normalised neighbour
// records and gathers the layer-0 operand  A x  (all of it model-independent) one time, keeps that operand in tensor
// memory, and then loops over the M models of the ensemble: per model two tcgen05 products, two epilogues, one
// 64-wide gather and the pooling.  The models' weights arrive as pre-split TF32 hi / lo operand images (k_ens_prep,
// once per ensemble) through bulk copies into ONE weight buffer per slot: W0 of model m + 1 is fetched as soon as
// model m's layer-0 epilogue is over, W1 of model m + 1 after its layer-1 epilogue, so the copies hide behind the
// other layer's work and no double buffer is needed.
// Same arithmetic, in the same order, as k_tc_fwd<.., COO> per model: identical bits per model.
#include <cuda_runtime.h>

#include "../../include/hcg.h"
#include "hcg_common.cuh"
#include "hcg_fused.cuh"
#include "hcg_ptx.cuh"
#include "hcg_tc.cuh"
#include "hcg_tc_tiles.cuh"

namespace hcg {

// operand image of one model, in floats: [W0 hi 64x32 | W0 lo | b0 64] [W1 hi 64x64 | W1 lo | b1 64]
constexpr int kEnsPart0 = 2 * tc::kN * 32 + 64;     // 4160 floats = 16640 bytes
constexpr int kEnsPart1 = 2 * tc::kN * 64 + 64;     // 8256 floats = 33024 bytes
constexpr int kEnsImage = kEnsPart0 + kEnsPart1;    // 12416 floats per model
constexpr uint32_t kEnsBuf0 = 17 * 1024, kEnsBuf1 = 33 * 1024;   // shared-memory homes of the two parts (1024-byte aligned)

struct EnsArgs {
  const float* x;
  const int64_t* ei;        // [2, E]
  int64_t N, E, B;
  const TcTile* tiles;
  const int* counts;
  int cap;
  const float* wimg;        // [M][kEnsImage]
  float* emb;               // [M][B][128]
  int M, F;
  float slope, fill;
  int* kflags;
};

struct EnsSmemLayout {
  uint32_t w0, w1, hbuf, nbr, deg, slot_bytes, total;
  uint32_t xw, ew;
};
__host__ __device__ inline EnsSmemLayout ens_layout(int F) {
  EnsSmemLayout y;
  y.xw = ((uint32_t)kTcRows * F + 16 + 3) & ~3u;   // words of the x stage
  y.ew = 2u * kTcEcap + 16;                        // words of one edge-row slice (int64 entries)
  y.w0 = 0;
  y.w1 = kEnsBuf0;
  y.hbuf = kEnsBuf0 + kEnsBuf1;                    // activation rows; before the model loop: x stage | sources | targets
  y.nbr = y.hbuf + (uint32_t)kTcRows * kTcHPitch * 4;
  y.deg = y.nbr + (uint32_t)kTcCooDeg * kTcRows * 8;
  y.slot_bytes = (y.deg + (uint32_t)kTcRows * 4 + 1023u) & ~1023u;
  y.total = 2 * y.slot_bytes + 1024;               // + slack for the 1024-byte alignment of the base
  return y;
}
static_assert((128 * 32 + 16 + 2 * (2 * kTcEcap + 16)) * 4 <= kTcRows * kTcHPitch * 4, "the stages must fit the activation buffer");

// conv weights of M models (stacked state_dict tensors) -> operand images
__global__ void __launch_bounds__(256)
k_ens_prep(const float* __restrict__ w0, const float* __restrict__ b0, const float* __restrict__ w1, const float* __restrict__ b1,
           int F, float* __restrict__ img) {
  const int m = blockIdx.x;
  float* o = img + (size_t)m * kEnsImage;
  tc::prep_weight_tiles(w0 + (size_t)m * 64 * F, F, 32, o, o + tc::kN * 32, threadIdx.x, 256);
  tc::prep_weight_tiles(w1 + (size_t)m * 64 * 64, 64, 64, o + kEnsPart0, o + kEnsPart0 + tc::kN * 64, threadIdx.x, 256);
  if (threadIdx.x < 64) {
    o[2 * tc::kN * 32 + threadIdx.x] = b0[(size_t)m * 64 + threadIdx.x];
    o[kEnsPart0 + 2 * tc::kN * 64 + threadIdx.x] = b1[(size_t)m * 64 + threadIdx.x];
  }
}

template <int F_>
__global__ void __launch_bounds__(512, 1)
k_tc_fwd_ens(const EnsArgs a) {
  constexpr int NS = 2, HP = kTcHPitch, kSlotThreads = 256, NW = 8;
  constexpr int CW = 32;   // hidden columns per thread
  constexpr int XW = 16;   // input columns per thread
  constexpr int NBS = kTcRows;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage[NS];
  __shared__ uint64_t bar_w0[NS], bar_w1[NS];
  __shared__ uint64_t bar_mma[NS];
  __shared__ uint32_t tmem_base_s;
  __shared__ TcMeta metas[NS];
  __shared__ float s_lut[32];
  __shared__ int s_bad[NS];

  const int tid = threadIdx.x;
  const int slot = tid / kSlotThreads, gt = tid % kSlotThreads;
  const int cg = gt >> 7;
  const int wslot = gt >> 5;
  const bool rev = slot == 1;        // (see k_tc_fwd: the slots fill the lane quadrants from opposite ends)
  const int i = rev ? 127 - (gt & 127) : (gt & 127);
  const int i_warp_min = rev ? 96 - (gt & 96) : (gt & 96);
  const bool ctl_warp = wslot == (rev ? 0 : 3);
  const bool pf_warp = wslot == (rev ? NW - 4 : NW - 1);
  const int lane = tid & 31;
  const int F = F_ > 0 ? F_ : a.F;
  const int M = a.M;

  const EnsSmemLayout lay = ens_layout(F);
  unsigned char* base = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u) + (uint32_t)slot * lay.slot_bytes;
  float* smf = reinterpret_cast<float*>(base);
  int32_t* smi = reinterpret_cast<int32_t*>(base);
  float* wb0 = reinterpret_cast<float*>(base + lay.w0);      // W0 hi | W0 lo | b0
  float* wb1 = reinterpret_cast<float*>(base + lay.w1);      // W1 hi | W1 lo | b1
  const float* s_b0 = wb0 + 2 * tc::kN * 32;
  const float* s_b1 = wb1 + 2 * tc::kN * 64;
  const int hb_w = (int)(lay.hbuf >> 2);
  const int nb_w = (int)(lay.nbr >> 2);
  const int dg_w = (int)(lay.deg >> 2);
  const int sx = hb_w, ssr = hb_w + (int)lay.xw, sds = ssr + (int)lay.ew;   // stages inside the activation buffer

  if (tid < 32) ptx::tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) {
      ptx::mbar_init(&bar_stage[s], 8);
      ptx::mbar_init(&bar_w0[s], 1);
      ptx::mbar_init(&bar_w1[s], 1);
      ptx::mbar_init(&bar_mma[s], 1);
    }
    ptx::fence_mbar_init();
  }
  if (tid < 32) {
    const float r = __fdiv_rn(1.0f, __fsqrt_rn((float)tid + a.fill));   // as hcg_collate computes deg^-1/2
    s_lut[tid] = isinf(r) ? 0.f : r;
  }
  if (tid < NS) s_bad[tid] = 0;
  if (gt < kTcRows) smi[dg_w + gt] = 0;
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();

  const uint32_t tbase = tmem_base_s + (uint32_t)slot * 256;
  const uint32_t d_t = tbase, ax_hi = tbase + 64, ax_lo = tbase + 96, a_hi = tbase + 128, a_lo = tbase + 192;
  const uint32_t trow = (uint32_t)((wslot & 3) * 32) << 16;
  const int bar_full = 1 + slot, bar_aop = 1 + NS + slot;
  const float slope = a.slope;

  const int slot_id = blockIdx.x * NS + slot;
  const int ntiles = a.counts[slot_id];
  const TcTile* my_tiles = a.tiles + (size_t)slot_id * a.cap;
  const int total_it = ntiles * M;      // (tile, model) iterations of this slot

  // weights of iteration `it` (model it % M): one lane issues the bulk copy of one part
  auto fetch_w0 = [&](int it) {
    if (it >= total_it) return;
    const float* src = a.wimg + (size_t)(it % M) * kEnsImage;
    ptx::fence_proxy_async_smem();
    ptx::mbar_expect_tx(&bar_w0[slot], kEnsPart0 * 4);
    ptx::bulk_g2s(wb0, src, kEnsPart0 * 4, &bar_w0[slot]);
  };
  auto fetch_w1 = [&](int it) {
    if (it >= total_it) return;
    const float* src = a.wimg + (size_t)(it % M) * kEnsImage + kEnsPart0;
    ptx::fence_proxy_async_smem();
    ptx::mbar_expect_tx(&bar_w1[slot], kEnsPart1 * 4);
    ptx::bulk_g2s(wb1, src, kEnsPart1 * 4, &bar_w1[slot]);
  };

  // feeding lanes (lane < 8 of pf_warp): lane = 4 gb + q copies slice q (0 x rows, 1 sources, 2 targets, 3 nothing)
  // of graph gb into the (free) activation buffer; all eight arrive on the slot's stage barrier once per tile
  auto issue = [&](int t) {
    if (lane >= 8) return;
    TcMeta* m = &metas[slot];
    uint64_t* bar = &bar_stage[slot];
    const int4* rec = reinterpret_cast<const int4*>(my_tiles + t);
    const int4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2);
    const int gb = lane >> 2, q = lane & 3;
    const int n0 = gb ? r1.y : r0.y, n = gb ? r1.z : r0.z, e0 = gb ? r1.w : r0.w, e = gb ? r2.y : r2.x;
    const bool have = gb == 0 || r1.x >= 0;
    const bool manual = (r2.z & kTileManual) != 0;
    uintptr_t pA, pS;
    int cA, cS;
    if (q == 0) { pA = reinterpret_cast<uintptr_t>(a.x + (int64_t)r0.y * F); cA = r0.z * F;
                  pS = reinterpret_cast<uintptr_t>(a.x + (int64_t)n0 * F); cS = n * F; }
    else {
      const int64_t* row = a.ei + (q == 2 ? a.E : 0);
      pA = reinterpret_cast<uintptr_t>(row + r0.w); cA = 2 * r2.x;
      pS = reinterpret_cast<uintptr_t>(row + e0); cS = 2 * e;
    }
    uintptr_t loA, loS;
    uint32_t bA, bS;
    int leadA, leadS;
    copy_window(pA, cA, &loA, &bA, &leadA);
    copy_window(pS, cS, &loS, &bS, &leadS);
    if (manual) { leadA = 0; leadS = 0; bA = (uint32_t)((cA * 4 + 15) & ~15); }
    const int arr_w = q == 0 ? sx : q == 1 ? ssr : sds;
    const int dst_w = arr_w + (gb ? (int)(bA >> 2) : 0);
    if (lane == 1) {
      m->gA = r0.x; m->n0A = r0.y; m->nA = r0.z; m->e0A = r0.w;
      m->gB = r1.x; m->n0B = r1.y; m->nB = r1.x >= 0 ? r1.z : 0; m->e0B = r1.w;
      m->eA = r2.x; m->eB = r2.y; m->flags = r2.z;
    }
    m->off[lane] = (gb ? (int)(bA >> 2) : 0) + leadS;
    if (!have || manual || q == 3 || cS == 0) {
      ptx::mbar_arrive(bar);
    } else {
      ptx::fence_proxy_async_smem();
      ptx::mbar_expect_tx(bar, bS);
      ptx::bulk_g2s(smf + dst_w, reinterpret_cast<const void*>(loS), bS, bar);
    }
  };

  // pooling of the rows in the activation buffer (one model of one tile): warp w owns columns 4 w + cc and
  // 32 + 4 w + cc; the eight lanes of a column take rows eight apart and are combined by shuffles (see k_tc_fwd)
  auto pool_tile = [&](float* emb_m, int gA, int gB, int nA, int nB) {
    constexpr int CPL = 64 / (4 * NW);
    const int cc = lane & 3, q = lane >> 2;
    const int c0 = 4 * wslot + cc;
    const float* hp = smf + hb_w + c0;
    for (int gb = 0; gb < (nB > 0 ? 2 : 1); ++gb) {
      const int rb = gb ? nA : 0, n = gb ? nB : nA;
      float mx[CPL], sm[CPL];
#pragma unroll
      for (int u = 0; u < CPL; ++u) { mx[u] = -INFINITY; sm[u] = 0.f; }
      int r = rb + q;
      for (; r + 8 < rb + n; r += 16) {
#pragma unroll
        for (int u = 0; u < CPL; ++u) {
          const float v0 = hp[r * HP + 32 * u], v1 = hp[(r + 8) * HP + 32 * u];
          mx[u] = fmaxf(mx[u], fmaxf(v0, v1));
          sm[u] += v0; sm[u] += v1;
        }
      }
      if (r < rb + n) {
#pragma unroll
        for (int u = 0; u < CPL; ++u) {
          const float v0 = hp[r * HP + 32 * u];
          mx[u] = fmaxf(mx[u], v0);
          sm[u] += v0;
        }
      }
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
#pragma unroll
        for (int u = 0; u < CPL; ++u) {
          mx[u] = fmaxf(mx[u], __shfl_xor_sync(0xffffffffu, mx[u], o));
          sm[u] += __shfl_xor_sync(0xffffffffu, sm[u], o);
        }
      }
      if (q == 0) {
        float* eo = emb_m + (int64_t)(gb ? gB : gA) * 128;
        const float inv = 1.f / (float)n;
#pragma unroll
        for (int u = 0; u < CPL; ++u) {
          eo[c0 + 32 * u] = mx[u];
          eo[64 + c0 + 32 * u] = sm[u] * inv;
        }
      }
    }
  };

  uint32_t mma_phase = 0;
  // hand-off to the tensor core: every thread of the slot has stored its part of the A operand; the control warp's
  // first lane waits for the weights of this step (`wbar`, parity `wph`), issues and commits
  auto mma_go = [&](uint64_t* wbar, uint32_t wph, auto&& issue_fn) {
    ptx::tc_fence_before_sync();
    if (ctl_warp) {
      ptx::named_bar_sync(bar_aop, kSlotThreads);
      if (lane == 0) {
        ptx::mbar_wait(wbar, wph);
        ptx::tc_fence_after_sync();
        issue_fn();
        ptx::mma_commit(&bar_mma[slot]);
      }
      __syncwarp();
    } else {
      ptx::named_bar_arrive(bar_aop, kSlotThreads);
    }
  };
  auto mma_done = [&]() {
    if (ctl_warp) ptx::mbar_wait(&bar_mma[slot], mma_phase);   // one warp polls; the others sleep in the barrier
    mma_phase ^= 1u;
    ptx::named_bar_sync(bar_full, kSlotThreads);
    ptx::tc_fence_after_sync();
  };

  if (pf_warp && ntiles > 0) {
    issue(0);
    if (lane == 0) { fetch_w0(0); fetch_w1(0); }
  }
  for (int k = 0; k < ntiles; ++k) {
    ptx::mbar_wait(&bar_stage[slot], (uint32_t)k & 1u);
    const TcMeta* mp = &metas[slot];
    const int gA = mp->gA, nA = mp->nA, nB = mp->nB, gB = mp->gB, eA = mp->eA;
    const int ntot = nA + nB;
    if (mp->flags & kTileManual) {  // flagged tile: fetch the slices with ordinary loads
      for (int gb = 0; gb < (nB > 0 ? 2 : 1); ++gb) {
        const int n0 = gb ? mp->n0B : mp->n0A, n = gb ? nB : nA, e0 = gb ? mp->e0B : mp->e0A, e = gb ? mp->eB : eA;
        const int o0 = mp->off[4 * gb + 0], o1 = mp->off[4 * gb + 1], o2 = mp->off[4 * gb + 2];
        for (int idx = gt; idx < n * F; idx += kSlotThreads) smf[sx + o0 + idx] = __ldg(a.x + (int64_t)n0 * F + idx);
        const int2* es = reinterpret_cast<const int2*>(a.ei + e0);
        const int2* ed = reinterpret_cast<const int2*>(a.ei + a.E + e0);
        for (int idx = gt; idx < e; idx += kSlotThreads) {
          *reinterpret_cast<int2*>(smi + ssr + o1 + 2 * idx) = __ldg(es + idx);
          *reinterpret_cast<int2*>(smi + sds + o2 + 2 * idx) = __ldg(ed + idx);
        }
      }
      ptx::named_bar_sync(bar_full, kSlotThreads);
    }
    const bool row_ok = i < ntot;
    const bool warp_ok = i_warp_min < ntot;
    const bool inB = row_ok && i >= nA;
    const int rowbase = inB ? nA : 0;
    const int xt_o = sx + (inB ? mp->off[4] : mp->off[0]) - rowbase * F;   // x[tile row r][c] at smf[xt_o + r F + c]
    const int nbo = nb_w + 2 * i;                                         // record t of this row at smi[nbo + 2 t kTcRows]

    // ---- A. every edge claims a record of its target row ---------------------------------------------
    {
      int bad = 0;
      const int etot = eA + mp->eB;
      for (int e = gt; e < etot; e += kSlotThreads) {
        const int eb = e >= eA;
        const int kk = eb ? e - eA : e;
        const int2 sv = *reinterpret_cast<const int2*>(smi + ssr + mp->off[4 * eb + 1] + 2 * kk);
        const int2 dv = *reinterpret_cast<const int2*>(smi + sds + mp->off[4 * eb + 2] + 2 * kk);
        const int gn0 = eb ? mp->n0B : mp->n0A, gn = eb ? nB : nA, rb = eb ? nA : 0;
        const int ts = sv.x - gn0, td = dv.x - gn0;
        if (sv.y != 0 || dv.y != 0 || (unsigned)ts >= (unsigned)gn || (unsigned)td >= (unsigned)gn || ts == td) {
          bad |= kCooBadEdge;
          continue;
        }
        const int pos = atomicAdd(smi + dg_w + rb + td, 1);
        if (pos < kTcCooDeg) smi[nb_w + 2 * (pos * kTcRows + rb + td)] = (e << 8) | (rb + ts);
        else bad |= kCooDegree;
      }
      if (bad) atomicOr(&s_bad[slot], bad);
    }
    ptx::named_bar_sync(bar_full, kSlotThreads);
    // ---- B. order the row's records by edge number, attach the norms ------------------------------------
    const int dtrue = row_ok ? smi[dg_w + i] : 0;
    const int deg = dtrue < kTcCooDeg ? dtrue : kTcCooDeg;
    const float di = row_ok ? s_lut[dtrue < 31 ? dtrue : 31] : 0.f;
    {
      const int wmax = __reduce_max_sync(0xffffffffu, deg);
      if (cg == 0 && wmax > 0) {
        int* kp = smi + nbo;
        auto finish = [&](int t, int key) {
          const int j = key & 0xff;
          const int dj = smi[dg_w + j];
          const float w = s_lut[dj < 31 ? dj : 31] * di;
          *reinterpret_cast<int2*>(kp + 2 * t * kTcRows) = make_int2(j, __float_as_int(w));
        };
        if (wmax <= 8) {
          int r[8];
#pragma unroll
          for (int t = 0; t < 8; ++t) r[t] = (t < deg && (t < 4 || wmax > 4)) ? kp[2 * t * kTcRows] : 0x7fffffff;
          if (wmax <= 4) {
            key_cx(r[0], r[1]); key_cx(r[2], r[3]); key_cx(r[0], r[2]); key_cx(r[1], r[3]); key_cx(r[1], r[2]);
          } else {
            key_cx(r[0], r[1]); key_cx(r[2], r[3]); key_cx(r[4], r[5]); key_cx(r[6], r[7]);
            key_cx(r[0], r[2]); key_cx(r[1], r[3]); key_cx(r[4], r[6]); key_cx(r[5], r[7]);
            key_cx(r[1], r[2]); key_cx(r[5], r[6]); key_cx(r[0], r[4]); key_cx(r[3], r[7]);
            key_cx(r[1], r[5]); key_cx(r[2], r[6]);
            key_cx(r[1], r[4]); key_cx(r[3], r[6]);
            key_cx(r[2], r[4]); key_cx(r[3], r[5]);
            key_cx(r[3], r[4]);
          }
#pragma unroll
          for (int t = 0; t < 8; ++t)
            if (t < deg) finish(t, r[t]);
        } else {
          for (int u = 1; u < deg; ++u) {
            const int key = kp[2 * u * kTcRows];
            int c = u - 1;
            while (c >= 0 && kp[2 * c * kTcRows] > key) { kp[2 * (c + 1) * kTcRows] = kp[2 * c * kTcRows]; --c; }
            kp[2 * (c + 1) * kTcRows] = key;
          }
          for (int t = 0; t < deg; ++t) finish(t, kp[2 * t * kTcRows]);
        }
      }
    }
    const float wself = (di * a.fill) * di;
    const int iself = row_ok ? i : 0;
    ptx::named_bar_sync(bar_full, kSlotThreads);
    if (cg == 0) smi[dg_w + i] = 0;   // every norm has been taken: counters ready for the next tile
    const int maxdeg = __reduce_max_sync(0xffffffffu, deg);
    const int2* nbp = reinterpret_cast<const int2*>(smi + nbo);
    auto nb_at = [&](int t) { return t < deg ? nbp[t * NBS] : make_int2(iself, 0); };

    // ---- layer-0 operand, once per tile: a_r = sum_j norm_rj x_j -> tensor memory (stays for all models) -----
    if (warp_ok) {
      float acc[XW];
#pragma unroll
      for (int c = 0; c < XW; ++c) acc[c] = 0.f;
      const int xc = xt_o + XW * cg;
      const int ncol = F - XW * cg;
      if (F_ > 0) {
        constexpr int N0 = F_ < XW ? F_ : XW;
        constexpr int N1 = F_ - XW < 0 ? 0 : (F_ - XW < XW ? F_ - XW : XW);
        if (cg == 0) gather_x<N0, XW, NBS>(acc, smf + xc, F, nbp, deg, maxdeg, iself, wself);
        else gather_x<N1, XW, NBS>(acc, smf + xc, F, nbp, deg, maxdeg, iself, wself);
      } else if (ncol > 0) {
        gather_x<-1, XW, NBS>(acc, smf + xc, F, nbp, deg, maxdeg, iself, wself, ncol < XW ? ncol : XW);
      }
      tc::store_split16(acc, ax_hi, ax_lo, trow, cg * 16);
    }
    ptx::tmem_wait_st();

    for (int m = 0; m < M; ++m) {
      const int it = k * M + m;
      const uint32_t wph = (uint32_t)it & 1u;
      float* emb_m = a.emb + (int64_t)m * a.B * 128;
      // ---- layer 0: D = (A x) W0_m^T ---------------------------------------------------------------
      mma_go(&bar_w0[slot], wph, [&] {
        tc::issue_split_mma<4>(d_t, ax_hi, ax_lo, ptx::smem_u32(wb0), ptx::smem_u32(wb0) + tc::kN * 32 * 4);
      });
      if (m > 0) pool_tile(emb_m - (int64_t)a.B * 128, gA, gB, nA, nB);   // the previous model's rows, while the MMAs run
      mma_done();
      if (warp_ok) {
#pragma unroll
        for (int part = 0; part < CW / 16; ++part) {
          uint32_t q0[16];
          ptx::tmem_ld16(d_t + trow + cg * CW + part * 16, q0);
          ptx::tmem_wait_ld();
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const int c0 = cg * CW + part * 16 + 4 * q4;
            const float4 b4 = *reinterpret_cast<const float4*>(s_b0 + c0);
            float4 v;
            v.x = __uint_as_float(q0[4 * q4 + 0]) + b4.x; v.x = fmaxf(v.x, v.x * slope);
            v.y = __uint_as_float(q0[4 * q4 + 1]) + b4.y; v.y = fmaxf(v.y, v.y * slope);
            v.z = __uint_as_float(q0[4 * q4 + 2]) + b4.z; v.z = fmaxf(v.z, v.z * slope);
            v.w = __uint_as_float(q0[4 * q4 + 3]) + b4.w; v.w = fmaxf(v.w, v.w * slope);
            *reinterpret_cast<float4*>(smf + hb_w + i * HP + c0) = v;
          }
        }
      }
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_full, kSlotThreads);   // h1 complete; W0_m and b0_m are no longer needed
      if (pf_warp && lane == 0) fetch_w0(it + 1);

      // ---- layer-1 operand: a_r = sum_j norm_rj h1_j ---------------------------------------------------
      {
        float acc[CW];
#pragma unroll
        for (int c = 0; c < CW; ++c) acc[c] = 0.f;
        if (warp_ok) {
          const float* hcol = smf + hb_w + CW * cg;
          auto row_fma = [&](int r, float w) {
            const float4* row = reinterpret_cast<const float4*>(hcol + r * HP);
#pragma unroll
            for (int q = 0; q < CW / 4; ++q) {
              const float4 v = row[q];
              acc[4 * q + 0] = fmaf(w, v.x, acc[4 * q + 0]);
              acc[4 * q + 1] = fmaf(w, v.y, acc[4 * q + 1]);
              acc[4 * q + 2] = fmaf(w, v.z, acc[4 * q + 2]);
              acc[4 * q + 3] = fmaf(w, v.w, acc[4 * q + 3]);
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
#pragma unroll
          for (int cc = 0; cc < CW; cc += 16) {
            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = acc[cc + q];
            tc::store_split16(v, a_hi, a_lo, trow, cg * CW + cc);
          }
        }
      }
      ptx::tmem_wait_st();
      // ---- layer 1: D = (A h1) W1_m^T ----------------------------------------------------------------
      mma_go(&bar_w1[slot], wph, [&] {
        tc::issue_split_mma<8>(d_t, a_hi, a_lo, ptx::smem_u32(wb1), ptx::smem_u32(wb1) + tc::kN * 64 * 4);
      });
      mma_done();                                    // (every gather of the slot is over: the rows may be overwritten)
      if (warp_ok) {
#pragma unroll
        for (int part = 0; part < CW / 16; ++part) {
          uint32_t q0[16];
          ptx::tmem_ld16(d_t + trow + cg * CW + part * 16, q0);
          ptx::tmem_wait_ld();
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const int c0 = cg * CW + part * 16 + 4 * q4;
            const float4 b4 = *reinterpret_cast<const float4*>(s_b1 + c0);
            float4 v;
            v.x = __uint_as_float(q0[4 * q4 + 0]) + b4.x; v.x = fmaxf(v.x, v.x * slope);
            v.y = __uint_as_float(q0[4 * q4 + 1]) + b4.y; v.y = fmaxf(v.y, v.y * slope);
            v.z = __uint_as_float(q0[4 * q4 + 2]) + b4.z; v.z = fmaxf(v.z, v.z * slope);
            v.w = __uint_as_float(q0[4 * q4 + 3]) + b4.w; v.w = fmaxf(v.w, v.w * slope);
            *reinterpret_cast<float4*>(smf + hb_w + i * HP + c0) = v;
          }
        }
      }
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_full, kSlotThreads);   // h2 complete; W1_m and b1_m are no longer needed
      if (pf_warp && lane == 0) fetch_w1(it + 1);
    }
    pool_tile(a.emb + (int64_t)(M - 1) * a.B * 128, gA, gB, nA, nB);
    ptx::named_bar_sync(bar_full, kSlotThreads);     // the activation buffer is free: the next tile's slices may land
    if (pf_warp && k + 1 < ntiles) issue(k + 1);
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (a.kflags && gt == 0) a.kflags[slot_id] = s_bad[slot];
  if (tid < 32) ptx::tmem_dealloc(tmem_base_s, 512);
}

size_t tc_fwd_ens_smem_bytes(int F) { return ens_layout(F).total; }
size_t tc_fwd_ens_image_floats(int M) { return (size_t)M * kEnsImage; }

template <int F_>
static int launch_ens_variant(const EnsArgs& a, int grid, size_t smem, cudaStream_t st) {
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_tc_fwd_ens<F_>, smem, configured)) return rc;
  k_tc_fwd_ens<F_><<<(unsigned)grid, 512, smem, st>>>(a);
  return HCG_OK;
}

int launch_tc_fwd_ens(const float* x, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N, int64_t E,
                      int64_t B, int M, int F, float slope, float fill, const float* w0, const float* b0, const float* w1,
                      const float* b1, float* wimg, float* emb, void* ws, cudaStream_t st) {
  if (B <= 0 || M <= 0) return HCG_OK;
  k_ens_prep<<<(unsigned)M, 256, 0, st>>>(w0, b0, w1, b1, F, wimg);
  HCG_LAUNCH_OK(st, "k_ens_prep");
  TcCooWs w = carve_tc_coo(ws, B);
  const TcPlan p = tc_plan(B, 2);
  if (int rc = launch_tile_desc_coo(x, F, edge_index, node_ptr, edge_ptr, N, E, B, p, w, emb, 128, st)) return rc;
  EnsArgs a{};
  a.x = x; a.ei = edge_index; a.N = N; a.E = E; a.B = B;
  a.tiles = w.tiles; a.counts = w.counts; a.cap = p.cap;
  a.wimg = wimg; a.emb = emb; a.M = M; a.F = F; a.slope = slope; a.fill = fill;
  a.kflags = w.kflags;
  const size_t smem = ens_layout(F).total;
  const int rc = F == 25 ? launch_ens_variant<25>(a, p.grid, smem, st)
               : F == 32 ? launch_ens_variant<32>(a, p.grid, smem, st) : launch_ens_variant<0>(a, p.grid, smem, st);
  if (rc) return rc;
  HCG_LAUNCH_OK(st, "k_tc_fwd_ens");
  return HCG_OK;
}

}  // namespace hcg
