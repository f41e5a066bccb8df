// Tensor-core graph-resident forward (see hcg_tc.cuh for the split-TF32 product) and a small
// self-test entry point that exercises exactly the same building blocks on one 128 x 64 x K tile.
#include "hcg_common.cuh"
#include "hcg_fused.cuh"
#include "hcg_tc.cuh"
#include "hcg_tcbwd_blocks.cuh"
#include "hcg_tc_tiles.cuh"
#include <stdlib.h>

#ifndef HCG_TC_DYNAMIC
#define HCG_TC_DYNAMIC 1   // 0: every slot walks its own list (the first scheduling; kept for A/B builds)
#endif

namespace hcg {

// ---- self-test: D[128,64] = A[128,K] . W[64,K]^T through TMEM-resident A and tcgen05.mma -------
__global__ void __launch_bounds__(128, 1)
k_tc_selftest(const float* __restrict__ A, const float* __restrict__ W, int K, float* __restrict__ D) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  // 1024-byte aligned weight tiles
  float* s_hi = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int Kpad = K <= 32 ? 32 : 64;
  float* s_lo = s_hi + tc::kN * Kpad;
  if (warp == 0) ptx::tmem_alloc(&tmem_base_s, 256);
  if (tid == 0) {
    ptx::mbar_init(&bar_mma, 1);
    ptx::fence_mbar_init();
  }
  tc::prep_weight_tiles(W, K, Kpad, s_hi, s_lo, tid, 128);
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t d_t = tbase, a_hi = tbase + 128, a_lo = tbase + 192;
  const uint32_t trow = (uint32_t)(warp * 32) << 16;

  for (int c0 = 0; c0 < Kpad; c0 += 16) {
    float v[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) v[q] = (c0 + q) < K ? A[tid * K + c0 + q] : 0.f;
    tc::store_split16(v, a_hi, a_lo, trow, c0);
  }
  ptx::tmem_wait_st();
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (tid == 0) {
    ptx::tc_fence_after_sync();
    if (Kpad == 32) tc::issue_split_mma<4>(d_t, a_hi, a_lo, ptx::smem_u32(s_hi), ptx::smem_u32(s_lo));
    else tc::issue_split_mma<8>(d_t, a_hi, a_lo, ptx::smem_u32(s_hi), ptx::smem_u32(s_lo));
    ptx::mma_commit(&bar_mma);
  }
  ptx::mbar_wait(&bar_mma, 0);
  ptx::tc_fence_after_sync();
  for (int c0 = 0; c0 < tc::kN; c0 += 16) {
    float r[16];
    tc::load_acc16(d_t, trow, c0, r);
#pragma unroll
    for (int q = 0; q < 16; ++q) D[tid * tc::kN + c0 + q] = r[q];
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tbase, 256);
}

// ---- self-test: G[64, 64] = X[128, 64]^T . Y[128, 64] (the node index is the reduction index) through MN-major
//      shared-memory operands; rows >= K of X and Y are zero ------------------------------------------------------
__global__ void __launch_bounds__(128, 1)
k_tc_selftest_xty(const float* __restrict__ X, const float* __restrict__ Y, int K, float* __restrict__ G) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  float* base = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // X hi | X lo | (two junk planes = the Y tiles below serve as the A operand's atoms 2, 3) ... keep it simple:
  float* x_hi = base;                          // 2 planes
  float* pad = x_hi + 2 * tcb::kPlaneFloats;   // 2 planes read as atoms 2, 3 of X hi (zeroed)
  float* x_lo = pad + 2 * tcb::kPlaneFloats;
  float* pad2 = x_lo + 2 * tcb::kPlaneFloats;  // atoms 2, 3 of X lo (zeroed)
  float* y_hi = pad2 + 2 * tcb::kPlaneFloats;
  float* y_lo = y_hi + 2 * tcb::kPlaneFloats;
  if (warp == 0) ptx::tmem_alloc(&tmem_base_s, 64);
  if (tid == 0) {
    ptx::mbar_init(&bar_mma, 1);
    ptx::fence_mbar_init();
  }
  for (int idx = tid; idx < 128 * 64; idx += 128) {
    const int r = idx >> 6, c = idx & 63;
    const float xv = r < K ? X[idx] : 0.f, yv = r < K ? Y[idx] : 0.f;
    const int o = tcb::plane_off(r, c);
    x_hi[o] = xv; x_lo[o] = xv - tcb::tf32_trunc(xv);
    y_hi[o] = yv; y_lo[o] = yv - tcb::tf32_trunc(yv);
    pad[o] = 0.f; pad2[o] = 0.f;
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t d_t = tmem_base_s;
  if (tid == 0) {
    tcb::issue_xty<64>(d_t, ptx::smem_u32(x_hi), ptx::smem_u32(x_lo), ptx::smem_u32(y_hi), ptx::smem_u32(y_lo), 16, 0u);
    ptx::mma_commit(&bar_mma);
  }
  ptx::mbar_wait(&bar_mma, 0);
  ptx::tc_fence_after_sync();
  const uint32_t trow = (uint32_t)(warp * 32) << 16;
  for (int c0 = 0; c0 < 64; c0 += 16) {
    float r[16];
    tc::load_acc16(d_t, trow, c0, r);
    if (tid < 64) {
#pragma unroll
      for (int q = 0; q < 16; ++q) G[tid * 64 + c0 + q] = r[q];
    }
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tmem_base_s, 64);
}

}  // namespace hcg

using namespace hcg;

extern "C" int hcg_selftest_tc_xty(const float* X, const float* Y, int K, float* G, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(X && Y && G, HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(K >= 1 && K <= 128, HCG_ERR_UNSUPPORTED, "K=%d outside [1,128]", K);
  const size_t smem = 12 * (size_t)tcb::kPlaneBytes + 1024;
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_tc_selftest_xty, smem, configured)) return rc;
  k_tc_selftest_xty<<<1, 128, smem, st>>>(X, Y, K, G);
  HCG_LAUNCH_OK(st, "k_tc_selftest_xty");
  return HCG_OK;
}

extern "C" int hcg_selftest_tc_gemm(const float* A, const float* W, int K, float* D, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(A && W && D, HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(K >= 1 && K <= 64, HCG_ERR_UNSUPPORTED, "K=%d outside [1,64]", K);
  const size_t smem = 2 * 64 * 64 * sizeof(float) + 1024;
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_tc_selftest, smem, configured)) return rc;
  k_tc_selftest<<<1, 128, smem, st>>>(A, W, K, D);
  HCG_LAUNCH_OK(st, "k_tc_selftest");
  return HCG_OK;
}

// =================================================================================================
// Graph-resident forward on the tensor cores.
//
// A CTA holds two (or three) independent "slots" of 256 threads; a slot walks its share of the batch one
// TILE at a time.  A tile is 128 node rows = the M of one tcgen05.mma = the 128 lanes of tensor memory,
// and holds ONE OR TWO whole reaction graphs (k_tile_desc pairs graphs of a small window of the batch
// whose node counts fit together: at 40..120 atoms a tile carries 1.37 graphs and ~85 % of its rows are
// real nodes instead of 62 %).  The adjacency of a tile is block diagonal, so nothing below cares
// how many graphs it holds except the pooling.  Two threads own a row (half of the columns each):
//
//   prefetch   the small CSR slices (rowptr, col, deg^-1/2) of the NEXT tile go out a whole tile ahead
//              into the other of two stages; its x rows, which have one buffer, as soon as this tile's
//              layer-0 gather is over: up to eight bulk async copies (cp.async.bulk -> one mbarrier),
//              issued by eight lanes of one warp; the work-list record itself arrives by cp.async
//   records    per CSR slot (tile row of the neighbour, norm_rj = dis_r dis_j) -> shared memory, once per tile
//   layer 0    a_r = sum_j norm_rj x_j gathered straight from the copied x rows (scalar loads: the rows
//              keep their F-float pitch, no re-pitch pass); all lanes of a warp take the same number of
//              neighbour steps (the warp's largest degree), a lane past its own degree idles;
//              split into TF32 hi / lo and stored as row r of the A operand in tensor memory
//   MMA        producers bar.arrive, the control warp bar.sync's, its first lane issues the split-TF32
//              tcgen05.mma chain against W_l^T (shared memory, 128-byte swizzle) and commits to an
//              mbarrier; the control warp polls it, the other warps sleep in a hardware barrier
//   epilogue   row r of the fp32 accumulator (tcgen05.ld) + bias, LeakyReLU -> activation buffer
//   layer l>0  a_r = sum_j norm_rj h_j from the activation buffer (128-bit loads), as above
//   pooling    column owners scan the last layer's rows of each graph of the tile; the eight row
//              interleaves of a column sit in one warp and are combined by shuffles:
//              emb[g] = [max || mean].  Runs one tile late, inside the next tile's layer-0 MMA window.
//
// The slots of a CTA run out of phase, so one slot's gathers overlap the other's MMAs.  Nothing of size
// [N,H] touches HBM unless the caller wants the activations (training forward).  Graphs above the tile
// capacity (128 nodes / kTcEcap CSR slots) are appended to an overflow list that the FMA graph-resident
// kernel processes afterwards.
// =================================================================================================
namespace hcg {

constexpr int kTcDefaultSlots = 2;    // tile pipelines per CTA (HCG_TC_SLOTS=3 selects the three-slot variant: measured slower,
                                      // 441 vs 398 us per 16384-graph batch - the kernel is bound by shared-memory throughput, not by latency)
// ---- work list ----------------------------------------------------------------------------------
// A warp per slot of the forward kernel; lane w takes the w-th window of W consecutive graphs that the slot
// owns (windows are dealt round-robin to the slots), pairs its graphs greedily (first unpaired graph + the
// largest later graph of the window that still fits the tile), and the lanes' tiles are written densely,
// in window order, through a warp scan.  Graphs above the tile capacity and graphs without nodes or edges
// go to the overflow list (the FMA graph-resident kernel).
__global__ void __launch_bounds__(128)
k_tile_desc(const int32_t* __restrict__ graph_ptr, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
            const float* __restrict__ dis, const float* __restrict__ x, int F, int64_t N, int64_t E, int64_t B, int W,
            int nslots, int nwin, int cap, TcTile* __restrict__ tiles, int* __restrict__ counts,
            int* __restrict__ overflow_count, int* __restrict__ overflow_list, int* __restrict__ overflow_flag,
            const int64_t* __restrict__ node_ptr64, const int64_t* __restrict__ edge_ptr64, const int64_t* __restrict__ ei,
            int* __restrict__ dflags, float* __restrict__ emb, int packed) {
  const int slot = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (slot >= nslots) return;  // whole warps leave together
  if (slot == 0 && lane == 0) counts[nslots] = 0;   // the tile queue of the kernel that walks the lists (k_tc_fwd)
  // COO mode (ei != NULL): node / edge offsets come from the batch's own int64 offset vectors, a graph may have
  // no edges, the copy windows are those of the two edge_index rows, and a graph that does not fit a tile is
  // REFUSED (flag in dflags[slot], embedding row zeroed) - there is no fallback kernel behind this path.
  // packed mode: the same, for the kernel variant that reads a packed graph image (byte-granular slices of 64-byte
  // aligned sections: no copy window can leave the image, no tile is flagged for ordinary loads)
  const bool coo = ei != nullptr || packed != 0;
  // CSR mode with the batch's int64 offset vectors (the training step's backward): the same tiles as from graph_ptr and
  // rowptr[graph_ptr[g]] - for a batch the collate accepts they hold the same numbers - without reading the collate's
  // outputs, so the work list does not have to wait for the collate
  const bool ptr64 = !coo && node_ptr64 != nullptr && edge_ptr64 != nullptr;
  int refused = 0;
  TcTile* out = tiles + (size_t)slot * cap;
  int total = 0;
  const int wps = (nwin + nslots - 1) / nslots;
  for (int w0 = 0; w0 < wps; w0 += 32) {
    const int win = slot + (w0 + lane) * nslots;
    int gp[kTcWinMax + 1], ee[kTcWinMax + 1];
    const int64_t g0 = (int64_t)win * W;
    const bool live = (w0 + lane) < wps && win < nwin;
    int wcnt = 0;
    if (live) wcnt = (int)((B - g0) < (int64_t)W ? (B - g0) : (int64_t)W);
#pragma unroll
    for (int k = 0; k <= kTcWinMax; ++k) {
      gp[k] = 0;
      if (live && k <= wcnt) {
        if (coo || ptr64) {
          const int64_t v = node_ptr64[g0 + k];
          gp[k] = (int)v;
          if (v < 0 || v > N) { gp[k] = 0; refused |= kCooBadOffsets; }
        } else {
          gp[k] = graph_ptr[g0 + k];
        }
      }
    }
#pragma unroll
    for (int k = 0; k <= kTcWinMax; ++k) {
      ee[k] = 0;
      if (live && k <= wcnt) {
        if (coo || ptr64) {
          const int64_t v = edge_ptr64[g0 + k];
          ee[k] = (int)v;
          if (v < 0 || v > E) { ee[k] = 0; refused |= kCooBadOffsets; }
        } else {
          ee[k] = rowptr[gp[k]];
        }
      }
    }
    unsigned used = 0, manual = 0;  // bit k: graph k of the window is not available / needs ordinary loads
#pragma unroll
    for (int k = 0; k < kTcWinMax; ++k) {
      const int n = gp[k + 1] - gp[k], e = ee[k + 1] - ee[k];
      if (k >= wcnt) { used |= 1u << k; continue; }
      const bool over = n <= 0 || (coo ? e < 0 : e <= 0) || n > kTcRows || e > kTcEcap;
      if (overflow_flag) overflow_flag[g0 + k] = over ? 1 : 0;   // (a per-graph flag: an order that does not depend on timing)
      if (over) {
        used |= 1u << k;
        if (coo) {
          refused |= kCooTooLarge;
          if (emb) for (int c = 0; c < 128; ++c) emb[(g0 + k) * 128 + c] = 0.f;
        } else {
          overflow_list[atomicAdd(overflow_count, 1)] = (int)(g0 + k);
        }
        continue;
      }
      if (packed) continue;
      uintptr_t lo; uint32_t bytes; int lead;
      bool m = false;
      copy_window(reinterpret_cast<uintptr_t>(x + (int64_t)gp[k] * F), n * F, &lo, &bytes, &lead);
      m |= lo + bytes > reinterpret_cast<uintptr_t>(x + N * F);
      if (coo) {   // slices of the two edge_index rows, as 4-byte words
        copy_window(reinterpret_cast<uintptr_t>(ei + ee[k]), 2 * e, &lo, &bytes, &lead);
        m |= lo + bytes > reinterpret_cast<uintptr_t>(ei + 2 * E);
        copy_window(reinterpret_cast<uintptr_t>(ei + E + ee[k]), 2 * e, &lo, &bytes, &lead);
        m |= lo + bytes > reinterpret_cast<uintptr_t>(ei + 2 * E);
        if (m) manual |= 1u << k;
        continue;
      }
      copy_window(reinterpret_cast<uintptr_t>(rowptr + gp[k]), n + 1, &lo, &bytes, &lead);
      m |= lo + bytes > reinterpret_cast<uintptr_t>(rowptr + N + 1);
      copy_window(reinterpret_cast<uintptr_t>(col + ee[k]), e, &lo, &bytes, &lead);
      m |= lo + bytes > reinterpret_cast<uintptr_t>(col + E);
      copy_window(reinterpret_cast<uintptr_t>(dis + gp[k]), n, &lo, &bytes, &lead);
      m |= lo + bytes > reinterpret_cast<uintptr_t>(dis + N);
      if (m) manual |= 1u << k;
    }
    // greedy pairing inside the window
    unsigned leader = 0;
    int partner[kTcWinMax];
#pragma unroll
    for (int a = 0; a < kTcWinMax; ++a) {
      partner[a] = -1;
      if (used & (1u << a)) continue;
      used |= 1u << a;
      leader |= 1u << a;
      const int na = gp[a + 1] - gp[a], ea = ee[a + 1] - ee[a];
      int best = -1, bn = 0;
#pragma unroll
      for (int b = a + 1; b < kTcWinMax; ++b) {
        const int nb = gp[b + 1] - gp[b], eb = ee[b + 1] - ee[b];
        if (!(used & (1u << b)) && na + nb <= kTcRows && ea + eb <= kTcEcap && nb > bn) { best = b; bn = nb; }
      }
      if (best >= 0) used |= 1u << best;
      partner[a] = best;
    }
    const int cnt = __popc(leader);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    int pos = total + incl - cnt;
    total += __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
    for (int a = 0; a < kTcWinMax; ++a) {
      if (!(leader & (1u << a))) continue;
      TcTile t;
      t.gA = (int)(g0 + a); t.n0A = gp[a]; t.nA = gp[a + 1] - gp[a]; t.e0A = ee[a]; t.eA = ee[a + 1] - ee[a];
      t.gB = -1; t.n0B = 0; t.nB = 0; t.e0B = 0; t.eB = 0;
      t.flags = (manual >> a) & 1u ? kTileManual : 0;
      const int b = partner[a];
#pragma unroll
      for (int k = 1; k < kTcWinMax; ++k) {
        if (k == b) {
          t.gB = (int)(g0 + k); t.n0B = gp[k]; t.nB = gp[k + 1] - gp[k]; t.e0B = ee[k]; t.eB = ee[k + 1] - ee[k];
          if ((manual >> k) & 1u) t.flags |= kTileManual;
        }
      }
      t.pad0 = 0; t.pad1[0] = t.pad1[1] = t.pad1[2] = t.pad1[3] = 0;
      int4* o4 = reinterpret_cast<int4*>(out + pos);
      o4[0] = make_int4(t.gA, t.n0A, t.nA, t.e0A);
      o4[1] = make_int4(t.gB, t.n0B, t.nB, t.e0B);
      o4[2] = make_int4(t.eA, t.eB, t.flags, 0);
      ++pos;
    }
  }
  if (lane == 0) counts[slot] = total;
  if (dflags) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) refused |= __shfl_xor_sync(0xffffffffu, refused, o);
    if (lane == 0) dflags[slot] = refused;
  }
}

// shared-memory carve-up, identical on host (size query) and device; weights first, then per slot the
// activation buffer, the neighbour records, ONE x stage and two stages of the small CSR slices.  Offsets in
// BYTES from the 1024-aligned base; stage array sizes in 4-byte words (multiples of 4, so every array is
// 16-byte aligned).
struct TcSmemLayout {
  uint32_t w0_hi, w0_lo, wl, bias, hbuf, nbr, deg, xst, cst;
  uint32_t xw, rw, cw, dw, csr_words, slot_bytes, total;
};
// coo: the stage arrays are the int64 source / target slices (2 words per edge) and there is no dis array; the
// neighbour records are kTcCooDeg planes of kTcRows (record t of row r at [t][r]) and a per-row in-degree
// counter follows them.
__host__ __device__ inline TcSmemLayout tc_layout(int F, int L, int nslots, bool coo = false) {
  TcSmemLayout y;
  uint32_t p = 0;
  y.w0_hi = p; p += tc::kN * 32 * 4;
  y.w0_lo = p; p += tc::kN * 32 * 4;
  y.wl = p; p += (uint32_t)(L - 1) * 2 * tc::kN * 64 * 4;  // per layer: hi tile, lo tile
  y.bias = p; p += (uint32_t)L * 64 * 4;
  // two slices per array, each a 16-byte granular window: up to 12 + 15 bytes of slack per slice
  y.xw = ((uint32_t)kTcRows * F + 16 + 3) & ~3u;
  y.rw = coo ? (2u * kTcEcap + 16) : (((uint32_t)kTcRows + 2 + 16 + 3) & ~3u);
  y.cw = coo ? (2u * kTcEcap + 16) : (((uint32_t)kTcEcap + 16 + 3) & ~3u);
  y.dw = coo ? 0u : (((uint32_t)kTcRows + 16 + 3) & ~3u);
  y.csr_words = y.rw + y.cw + y.dw;
  y.hbuf = p;
  y.nbr = y.hbuf + (uint32_t)kTcRows * kTcHPitch * 4;   // slot 0's (tile row, norm) record per CSR slot
  y.deg = y.nbr + (coo ? (uint32_t)kTcCooDeg * kTcRows * 8 : (uint32_t)kTcEcap * 8);   // coo: in-degree per tile row
  y.xst = y.deg + (coo ? (uint32_t)kTcRows * 4 : 0u);   // slot 0's x rows
  y.cst = y.xst + y.xw * 4;                             // slot 0's first CSR stage (rowptr | col | dis)
  y.slot_bytes = (y.cst - y.hbuf) + 2 * y.csr_words * 4;
  p += (uint32_t)nslots * y.slot_bytes;
  y.total = p + 1024;  // slack for the 1024-byte alignment of the base
  return y;
}

#ifdef HCG_TC_TIMING
__device__ unsigned long long g_tc_timing[16];
#define TC_TICK(slot_) do { if (tid == 0 && blockIdx.x == 0) { long long t_ = clock64(); g_tc_timing[slot_] += (unsigned long long)(t_ - t_prev); t_prev = t_; } } while (0)
#else
#define TC_TICK(slot_) do { } while (0)
#endif

// F_ = compile-time number of input features (25: the RhCAA featuriser, 32: biaryl) or 0 = run-time F <= 32.
// NS = tile pipelines ("slots") per CTA, 2 or 3.  Two threads own a tile row (32 of the 64 hidden columns each,
// in two warps of the same lane quadrant), so a slot has 256 threads and the CTA 256 NS.
//   NS = 2: 128 registers per thread; tensor memory per slot: accumulator 64 | - | A hi 64 | A lo 64 columns.
//   NS = 3:  80 registers per thread; tensor memory per slot: accumulator 64 | A hi 32 | A lo 32 columns, so a
//            64-wide layer runs as two K halves: the threads of column half 0 store their operand half and its
//            MMAs are issued; the threads of half 1 (their gather ran meanwhile) wait for those MMAs to have
//            read the operand (tcgen05.commit -> mbarrier), store theirs, and the second chain accumulates on top.
//
// COO = true: the kernel reads the raw edge list (a.ei: int64 [2, E], each graph's edges contiguous, located by the
// work list) instead of a collated CSR, and builds the tile's neighbour records itself:
//   A. every staged edge (s, d) claims the next record of row d (shared-memory atomic on the row's in-degree counter)
//      and leaves (tile row of s, edge number) there;
//   B. the row's first thread orders its records by edge number (sorting network in registers on the packed keys
//      edge number << 8 | tile row: the claim order is arbitrary, the summation order must not be) and rewrites
//      them as (tile row, norm_ij), norm from the two
//      in-degrees through a small table of (deg + fill)^-1/2 - IEEE sqrt and divide, as hcg_collate computes them.
// The records of a row are kTcRows apart (plane t holds record t of every row): the gathers' record loads are
// conflict-free.  Same arithmetic in the same order as the CSR variant on hcg_collate's output: identical bits.
// Refused input (kCoo* flags): see k_tile_desc; the caller falls back to hcg_collate + hcg_gcn_forward.
//
// MODE 2 (PK): the same from a PACKED graph image resident in device memory (hcatgnet_b200/packed.py; a.pk): the
// feeding lanes copy each graph's slices of the bit / byte / f32 feature sections and of the local byte edge pairs
// (a few bytes per node instead of 4 F) into the stage, the slot expands them to fp32 rows in its x buffer (what
// hcg_unpack_batch would have written to HBM and this kernel read back), and the records come from the byte pairs.
template <int F_, int NS, int MODE = 0>
__global__ void __launch_bounds__(NS * 256, 1)
k_tc_fwd(const TcArgs a) {
  constexpr bool COO = MODE == 1 || MODE == 2 || MODE == 4, PK = MODE == 2;
  constexpr bool SAVE = MODE == 3 || MODE == 4;   // the training forward (every layer's rows are saved): 3 from the CSR, 4 from the edge list
#ifdef HCG_TC_TIMING
  long long t_prev = clock64();
#endif
  constexpr int HP = kTcHPitch;
  constexpr int TPR = 2;
  constexpr int kSlotThreads = 128 * TPR, kThreads = NS * kSlotThreads, NW = 4 * TPR;
  constexpr int CW = 64 / TPR;    // hidden columns per thread (gather of layers >= 1, epilogue)
  constexpr int XW = 32 / TPR;    // input columns per thread (layer-0 gather)
  constexpr bool KSPLIT = NS == 3;
  constexpr int NBS = COO ? kTcRows : 1;   // records between consecutive neighbour records of a row
  __shared__ float s_lut[32];              // COO: (deg + fill)^-1/2 for deg = 0..31
  __shared__ int s_bad[NS];                // COO: refusal flags of the slot
  __shared__ int s_slot[32];               // PK: where column c of x lives in the image (HcgPackedStore.col_slot)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage[NS][2];
  __shared__ uint64_t bar_mma[NS];
  __shared__ uint64_t bar_free[NS];        // K-split: the first half's MMAs have read the A operand
  __shared__ uint32_t tmem_base_s;
  __shared__ TcMeta metas[NS][2];
  __shared__ int4 recs[NS][2][3];          // work-list records in flight (cp.async), two deep
#if HCG_TC_DYNAMIC
  __shared__ int s_pref[3 * kNumSMs + 1];  // exclusive prefix sums of the per-slot list lengths: tile g of the launch is
  __shared__ int s_rec_ok[NS][2];          //   entry g - s_pref[l] of list l; s_rec_ok: the record in flight is a tile
#endif

  const int tid = threadIdx.x;
  const int slot = tid / kSlotThreads, gt = tid % kSlotThreads;
  const int cg = gt >> 7;            // which half of the columns this thread works on
  const int wslot = gt >> 5;         // warp inside the slot
  // Tile row owned by this thread.  Its TMEM lane is fixed by the hardware (lane quadrant = warp id % 4 =
  // the warp scheduler), and tiles fill rows from 0 up, so slot 1 maps its rows in reverse: slots 0 and 2 load
  // schedulers 0,1 the most, slot 1 schedulers 3,2.
  const bool rev = slot == 1;
  const int i = rev ? 127 - (gt & 127) : (gt & 127);
  const int i_warp_min = rev ? 96 - (gt & 96) : (gt & 96);  // smallest row owned by this warp
  // the warp of each slot whose quadrant is filled last issues the MMAs; its sibling in the other column
  // half feeds the slot (bulk copies of the next tile)
  const bool ctl_warp = wslot == (rev ? 0 : 3);
  const bool pf_warp = wslot == (rev ? NW - 4 : NW - 1);
  const int lane = tid & 31;
  const int F = F_ > 0 ? F_ : a.f.F;
  const int L = a.f.L;

  // ---- carve shared memory (one base pointer, word offsets: keeps every access an LDS / STS) -------
  const TcSmemLayout lay = tc_layout(F, L, NS, COO);
  unsigned char* base = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  float* smf = reinterpret_cast<float*>(base);
  int32_t* smi = reinterpret_cast<int32_t*>(base);
  float* w0_hi = reinterpret_cast<float*>(base + lay.w0_hi);
  float* w0_lo = reinterpret_cast<float*>(base + lay.w0_lo);
  float* wl_base = reinterpret_cast<float*>(base + lay.wl);
  float* s_bias = reinterpret_cast<float*>(base + lay.bias);
  const int hb_w = (int)((lay.hbuf + (uint32_t)slot * lay.slot_bytes) >> 2);   // activation buffer, word offset
  const int nb_w = (int)((lay.nbr + (uint32_t)slot * lay.slot_bytes) >> 2);    // neighbour records
  const int dg_w = (int)((lay.deg + (uint32_t)slot * lay.slot_bytes) >> 2);    // COO: in-degree counters
  const int sx = (int)((lay.xst + (uint32_t)slot * lay.slot_bytes) >> 2);      // x rows (one stage)
  const int cs_w0 = (int)((lay.cst + (uint32_t)slot * lay.slot_bytes) >> 2);   // CSR stage 0
  const int rw = (int)lay.rw, cw = (int)lay.cw, csw = (int)lay.csr_words;

  // ---- one-time setup ----------------------------------------------------------------------------
  if (tid < 32) ptx::tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) {
      ptx::mbar_init(&bar_stage[s][0], 8);   // the eight feeding lanes: arrive (+ expect_tx) once per tile each
      ptx::mbar_init(&bar_stage[s][1], 8);
      ptx::mbar_init(&bar_mma[s], 1);
      ptx::mbar_init(&bar_free[s], 1);
    }
    ptx::fence_mbar_init();
  }
  tc::prep_weight_tiles(a.f.conv_w[0], F, 32, w0_hi, w0_lo, tid, kThreads);
  for (int l = 1; l < L; ++l)
    tc::prep_weight_tiles(a.f.conv_w[l], 64, 64, wl_base + (size_t)(l - 1) * 2 * tc::kN * 64,
                          wl_base + (size_t)(l - 1) * 2 * tc::kN * 64 + tc::kN * 64, tid, kThreads);
  for (int idx = tid; idx < L * 64; idx += kThreads) s_bias[idx] = __ldg(a.f.conv_b[idx >> 6] + (idx & 63));
  if (COO) {
    if (tid < 32) {
      const float r = __fdiv_rn(1.0f, __fsqrt_rn((float)tid + a.f.fill));
      s_lut[tid] = isinf(r) ? 0.f : r;
    }
    if (tid < NS) s_bad[tid] = 0;
    if (gt < kTcRows) smi[dg_w + gt] = 0;
    if (PK && tid < 32) s_slot[tid] = tid < F ? a.pk.col_slot[tid] : 0;
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();

  const uint32_t tbase = tmem_base_s + (uint32_t)slot * (KSPLIT ? 128 : 256);
  const uint32_t d_t = tbase, a_hi = tbase + (KSPLIT ? 64 : 128), a_lo = tbase + (KSPLIT ? 96 : 192);
  const uint32_t trow = (uint32_t)((wslot & 3) * 32) << 16;  // warps w and w + 4 share a lane quadrant
  const int bar_full = 1 + slot;        // all threads of the slot
  const int bar_aop = 1 + NS + slot;    // A operand complete: the other warps arrive, the control warp waits
  const int bar_a1 = 1 + 2 * NS + slot; // K-split: second operand half complete (the warps of half 1 arrive)
  const float slope = a.f.slope;

  const int slot_id = blockIdx.x * NS + slot;
#if HCG_TC_DYNAMIC
  // The work list is per-slot lists of tiles (k_tile_desc).  Walking only its own list leaves a slot with the largest
  // graphs (or one window more) working alone at the end: with 16384 graphs of 40..120 nodes the longest list has 52
  // tiles against 44.4 on average.  No result depends on which slot computes a tile, so the slots take tiles from ONE
  // queue over all lists instead: a global counter (zeroed by k_tile_desc, behind the list lengths) hands out tile
  // numbers; the feeding warp finds the list by a binary search in the prefix sums and fetches the record two tiles ahead.
  const int nslots_all = (int)gridDim.x * NS;
  if (tid < 32) {
    int run = 0;
    for (int b0 = 0; b0 < nslots_all; b0 += 32) {
      const int v = b0 + lane < nslots_all ? a.counts[b0 + lane] : 0;
      int incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
      }
      if (b0 + lane < nslots_all) s_pref[b0 + lane] = run + incl - v;
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) s_pref[nslots_all] = run;
  }
  __syncthreads();
  const int total_tiles = s_pref[nslots_all];
  int* const queue = const_cast<int*>(a.counts) + nslots_all;
#else
  const int ntiles = a.counts[slot_id];
  const TcTile* my_tiles = a.tiles + (size_t)slot_id * a.cap;
#endif

  // PK: byte offsets, inside a stage, of the four staged arrays (q = 1 edge pairs | 0 bit rows | 2 byte rows | 3 f32 rows);
  // every region leaves room for the 16-byte granular windows of two slices
  const int pk_bs = PK ? a.pk.bit_stride : 0, pk_fu = PK ? a.pk.Fu : 0, pk_ff = PK ? a.pk.Ff : 0;
  const int pk_ob = 2 * kTcEcap + 64;     // (two windows: up to 15 bytes of lead and 15 of rounding each)
  const int pk_ou = pk_ob + ((kTcRows * pk_bs + 79) & ~15);
  const int pk_of = pk_ou + ((kTcRows * pk_fu + 79) & ~15);
  // Feeding lanes (lane < 8 of pf_warp): lane = 4 gb + q copies slice q (0 x, 1 rowptr, 2 col, 3 dis) of graph gb.
  // The record of tile t travels global -> recs[t & 1] by cp.async (three lanes, 16 bytes each), a tile ahead.
#if HCG_TC_DYNAMIC
  auto fetch_rec = [&](int t) {   // (whole feeding warp)
    int g = 0;
    if (lane == 0) g = atomicAdd(queue, 1);
    g = __shfl_sync(0xffffffffu, g, 0);
    const bool ok = g < total_tiles;
    if (ok) {
      int lo = 0, hi = nslots_all;          // largest l with s_pref[l] <= g (empty lists repeat their prefix: never chosen)
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_pref[mid] <= g) lo = mid; else hi = mid;
      }
      const TcTile* src = a.tiles + (size_t)lo * a.cap + (g - s_pref[lo]);
      if (lane < 3) ptx::cp_async16(&recs[slot][t & 1][lane], reinterpret_cast<const int4*>(src) + lane);
    }
    if (lane == 0) s_rec_ok[slot][t & 1] = ok ? 1 : 0;
  };
#else
  auto fetch_rec = [&](int t) {
    if (t < ntiles && lane < 3) ptx::cp_async16(&recs[slot][t & 1][lane], reinterpret_cast<const int4*>(my_tiles + t) + lane);
  };
#endif
  // Issue the copies of tile t (or mark "no more tiles").  The small CSR slices are double-buffered and go out
  // a whole tile ahead (want_x = false: the six lanes with q >= 1); the x rows have ONE buffer, so the two x
  // lanes go later, once the previous tile's layer-0 gather is over (want_x = true).  All eight lanes arrive
  // on the mbarrier of stage t & 1, each exactly once per tile.
  auto issue = [&](int t, bool want_x) {
    const int s = t & 1;
    TcMeta* m = &metas[slot][s];
    uint64_t* bar = &bar_stage[slot][s];
    if (!want_x) {
      ptx::cp_async_wait_all();   // this lane's part of the record ...
      __syncwarp();               // ... and the other lanes'
    }
    if (lane >= 8 || (PK ? want_x : ((lane & 3) == 0) != want_x)) return;   // (PK: nothing waits for the x buffer: all eight go out early)
#if HCG_TC_DYNAMIC
    if (!s_rec_ok[slot][s]) {
#else
    if (t >= ntiles) {
#endif
      if (lane == 1) m->gA = -1;
      ptx::mbar_arrive(bar);
      return;
    }
    const int4 r0 = recs[slot][s][0], r1 = recs[slot][s][1], r2 = recs[slot][s][2];
    const int gb = lane >> 2, q = lane & 3;
    const int n0 = gb ? r1.y : r0.y, n = gb ? r1.z : r0.z, e0 = gb ? r1.w : r0.w, e = gb ? r2.y : r2.x;
    const bool have = gb == 0 || r1.x >= 0;
    const bool manual = (r2.z & kTileManual) != 0;
    if (PK) {
      // byte slices of the image: q = 0 bit rows, 1 local (src, dst) byte pairs, 2 byte rows, 3 f32 rows
      const unsigned char* arr = q == 0 ? a.pk.xbits : q == 1 ? static_cast<const unsigned char*>(a.pk.edges)
                               : q == 2 ? a.pk.xu8 : reinterpret_cast<const unsigned char*>(a.pk.xf32);
      const int64_t unit = q == 0 ? pk_bs : q == 1 ? 2 : q == 2 ? pk_fu : 4 * pk_ff;
      const uintptr_t pA = reinterpret_cast<uintptr_t>(arr) + (uintptr_t)((q == 1 ? (int64_t)r0.w : (int64_t)r0.y) * unit);
      const uintptr_t pS = reinterpret_cast<uintptr_t>(arr) + (uintptr_t)((q == 1 ? (int64_t)e0 : (int64_t)n0) * unit);
      const uint32_t cA = (uint32_t)((q == 1 ? r2.x : r0.z) * unit), cS = (uint32_t)((q == 1 ? e : n) * unit);
      const uintptr_t loA = pA & ~uintptr_t(15), loS = pS & ~uintptr_t(15);
      const uint32_t bA = (uint32_t)((pA - loA) + cA + 15) & ~15u, bS = (uint32_t)((pS - loS) + cS + 15) & ~15u;
      const int reg = q == 0 ? pk_ob : q == 1 ? 0 : q == 2 ? pk_ou : pk_of;
      if (lane == 1) {
        m->gA = r0.x; m->n0A = r0.y; m->nA = r0.z; m->e0A = r0.w;
        m->gB = r1.x; m->n0B = r1.y; m->nB = r1.x >= 0 ? r1.z : 0; m->e0B = r1.w;
        m->eA = r2.x; m->eB = r2.y; m->flags = 0;
      }
      m->off[lane] = reg + (gb ? (int)bA : 0) + (int)(pS - loS);     // BYTES from the stage's start to element 0 of the slice
      if (!have || cS == 0) {
        ptx::mbar_arrive(bar);
      } else {
        ptx::mbar_expect_tx(bar, bS);
        ptx::bulk_g2s(reinterpret_cast<unsigned char*>(smf + cs_w0 + s * csw) + reg + (gb ? bA : 0u), reinterpret_cast<const void*>(loS), bS, bar);
      }
      return;
    }
    // slice q of graph A too (every lane needs its window size: graph B's slice lands right behind it)
    uintptr_t pA, pS;
    int cA, cS;
    if (q == 0) { pA = reinterpret_cast<uintptr_t>(a.f.x + (int64_t)r0.y * F); cA = r0.z * F;
                  pS = reinterpret_cast<uintptr_t>(a.f.x + (int64_t)n0 * F); cS = n * F; }
    else if (COO) {
      // q = 1: sources, q = 2: targets of the graph's edges (int64: two words each); q = 3 has nothing to copy
      const int64_t* row = a.ei + (q == 2 ? a.E : 0);
      pA = reinterpret_cast<uintptr_t>(row + r0.w); cA = 2 * r2.x;
      pS = reinterpret_cast<uintptr_t>(row + e0); cS = 2 * e;
    }
    else if (q == 1) { pA = reinterpret_cast<uintptr_t>(a.f.rowptr + r0.y); cA = r0.z + 1;
                       pS = reinterpret_cast<uintptr_t>(a.f.rowptr + n0); cS = n + 1; }
    else if (q == 2) { pA = reinterpret_cast<uintptr_t>(a.f.col + r0.w); cA = r2.x;
                       pS = reinterpret_cast<uintptr_t>(a.f.col + e0); cS = e; }
    else { pA = reinterpret_cast<uintptr_t>(a.f.dis + r0.y); cA = r0.z;
           pS = reinterpret_cast<uintptr_t>(a.f.dis + n0); cS = n; }
    uintptr_t loA, loS;
    uint32_t bA, bS;
    int leadA, leadS;
    copy_window(pA, cA, &loA, &bA, &leadA);
    copy_window(pS, cS, &loS, &bS, &leadS);
    if (manual) { leadA = 0; leadS = 0; bA = (uint32_t)((cA * 4 + 15) & ~15); }   // ordinary loads place the slices themselves
    const int arr_w = q == 0 ? sx : cs_w0 + s * csw + (q == 1 ? 0 : q == 2 ? rw : rw + cw);
    const int dst_w = arr_w + (gb ? (int)(bA >> 2) : 0);
    if (lane == 1) {
      m->gA = r0.x; m->n0A = r0.y; m->nA = r0.z; m->e0A = r0.w;
      m->gB = r1.x; m->n0B = r1.y; m->nB = r1.x >= 0 ? r1.z : 0; m->e0B = r1.w;
      m->eA = r2.x; m->eB = r2.y; m->flags = r2.z;
    }
    m->off[lane] = (gb ? (int)(bA >> 2) : 0) + leadS;
    // the metadata must be in place before the arrival that can complete the phase (release / acquire)
    if (!have || manual || (COO && (q == 3 || cS == 0))) {
      ptx::mbar_arrive(bar);
    } else {
      ptx::mbar_expect_tx(bar, bS);   // this lane's arrival for the phase + its bytes in flight
      ptx::bulk_g2s(smf + dst_w, reinterpret_cast<const void*>(loS), bS, bar);
    }
  };

  // ---- pooling: warp w owns columns 4 w + cc and 32 + 4 w + cc, cc = lane & 3; the eight lanes of a column
  //      take the rows of a graph eight apart (bank = 4 row + column: conflict-free) and are combined by
  //      shuffles.  Runs one tile late, inside the next tile's layer-0 MMA window. ------------------------
  auto pool_tile = [&](int gA, int gB, int nA, int nB) {
    constexpr int CPL = 64 / (4 * NW);   // columns per lane
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
        float* eo = a.f.emb + (int64_t)(gb ? gB : gA) * 128;
        const float inv = 1.f / (float)n;   // one division per graph; the mean is sum * (1 / n)
#pragma unroll
        for (int u = 0; u < CPL; ++u) {
          eo[c0 + 32 * u] = mx[u];
          eo[64 + c0 + 32 * u] = sm[u] * inv;
        }
      }
    }
  };
  int p_gA = -1, p_gB = -1, p_nA = 0, p_nB = 0;   // the tile whose last-layer rows sit in the activation buffer, not yet pooled
  int p_n0A = 0, p_n0B = 0;
  // (training forward) the rows of a finished layer -> global memory, from the activation buffer, in 16-byte chunks with
  // sixteen lanes per row: a warp writes two whole rows per instruction.  (One 256-byte bulk copy per row costs ~21 issue
  // cycles in the copy engine - 2 700 per layer and tile, shared by the two slots.)  Runs one MMA window late, like the
  // pooling: the buffer keeps the rows until the next epilogue, which every thread reaches through a barrier behind this.
  auto save_rows = [&](int layer, int nA_, int nB_, int n0A_, int n0B_) {
    float* const dst = a.f.save_act[layer];
    if (dst == nullptr) return;
    const int nt = nA_ + nB_;
    for (int idx = gt; idx < nt * 16; idx += kSlotThreads) {
      const int r = idx >> 4, q = idx & 15;
      const int64_t grow = r < nA_ ? (int64_t)n0A_ + r : (int64_t)n0B_ + (r - nA_);
      __stcs(reinterpret_cast<float4*>(dst + grow * 64 + 4 * q), *reinterpret_cast<const float4*>(smf + hb_w + r * HP + 4 * q));
    }
  };

  // PK: the x rows of a tile, expanded from the staged bit / byte / f32 slices of stage `st` into dense fp32 rows of the
  // slot's x buffer.  Runs at the top of a tile, or - from the second conv layer on - a tile EARLY, inside the previous
  // tile's layer-1 MMA window (the x buffer is free once that tile's layer-0 gather is over).
  bool pre_expanded = false;
  auto expand_x = [&](const TcMeta* m, int st) {
      // ---- the tile's x rows, expanded from the staged bit / byte / f32 slices into dense fp32 rows ---------
      // (two threads per row, half of the columns each; a thread's columns are walked from a row-dependent start so
      //  that the lanes of a warp - consecutive rows, pitch F floats - write different banks for any F)
      const unsigned char* st8 = reinterpret_cast<const unsigned char*>(smf + cs_w0 + st * csw);
      const int r = gt & 127, half = gt >> 7;
      if (r < m->nA + m->nB) {
        const int gb = r >= m->nA, lr = gb ? r - m->nA : r;
        const unsigned char* pb = st8 + m->off[4 * gb + 0] + lr * pk_bs;
        uint32_t bits = 0;                       // F <= 32: at most 32 bit columns, four bytes
        for (int q = 0; q < pk_bs; ++q) bits |= (uint32_t)pb[q] << (8 * q);
        const unsigned char* pu = st8 + m->off[4 * gb + 2] + lr * pk_fu;
        const float* pf = reinterpret_cast<const float*>(st8 + m->off[4 * gb + 3]) + lr * pk_ff;
        const int c0 = half ? (F + 1) >> 1 : 0, w = half ? F - ((F + 1) >> 1) : (F + 1) >> 1;
        float* xo = smf + sx + r * F + c0;
        const int cs = r % (w > 0 ? w : 1);
        constexpr int WMAX = F_ > 0 ? (F_ + 1) / 2 : 16;
        int sl[WMAX];
#pragma unroll
        for (int j = 0; j < WMAX; ++j) {          // (unrolled: the table reads, then the data reads, go out together)
          int c = cs + j; c = c >= w ? c - w : c;
          sl[j] = j < w ? s_slot[c0 + c] : 65536;
        }
        float v[WMAX];
#pragma unroll
        for (int j = 0; j < WMAX; ++j)
          v[j] = sl[j] >= 65536 ? (float)((bits >> ((sl[j] - 65536) & 31)) & 1u) : sl[j] >= 0 ? (float)pu[sl[j]] : pf[~sl[j]];
#pragma unroll
        for (int j = 0; j < WMAX; ++j) {
          int c = cs + j; c = c >= w ? c - w : c;
          if (j < w) xo[c] = v[j];
        }
      }
  };

  if (pf_warp) {
    fetch_rec(0);
    issue(0, false);
    issue(0, true);
    fetch_rec(1);
  }
  uint32_t mma_phase = 0, free_phase = 0;
  for (int k = 0;; ++k) {
    const int s = k & 1;
    TC_TICK(0);
    ptx::mbar_wait(&bar_stage[slot][s], (uint32_t)(k >> 1) & 1u);
    TC_TICK(1);
    const TcMeta* mp = &metas[slot][s];
    const int gA = mp->gA;
    if (gA < 0) break;  // slot-uniform: no more tiles
    // next tile's CSR slices -> the other CSR stage: its readers (the record pass of tile k - 1) are long done,
    // and the record slot recs[(k + 2) & 1] was last read when tile k's x rows were issued
    if (pf_warp) {
      issue(k + 1, false);
      fetch_rec(k + 2);
    }
    const int nA = mp->nA, nB = mp->nB, gB = mp->gB, eA = mp->eA;   // (stage s's record is rewritten during tile k + 1)
    const int n0A_t = mp->n0A, n0B_t = mp->n0B;
    const int ntot = nA + nB;
    const int srp = cs_w0 + s * csw, scl = srp + rw, sds = scl + cw;
    if (mp->flags & kTileManual) {  // flagged tile: fetch the slices with ordinary loads
      for (int gb = 0; gb < (nB > 0 ? 2 : 1); ++gb) {
        const int n0 = gb ? mp->n0B : mp->n0A, n = gb ? nB : nA, e0 = gb ? mp->e0B : mp->e0A, e = gb ? mp->eB : eA;
        const int o0 = mp->off[4 * gb + 0], o1 = mp->off[4 * gb + 1], o2 = mp->off[4 * gb + 2], o3 = mp->off[4 * gb + 3];
        for (int idx = gt; idx < n * F; idx += kSlotThreads) smf[sx + o0 + idx] = __ldg(a.f.x + (int64_t)n0 * F + idx);
        if (COO) {
          const int2* es = reinterpret_cast<const int2*>(a.ei + e0);
          const int2* ed = reinterpret_cast<const int2*>(a.ei + a.E + e0);
          for (int idx = gt; idx < e; idx += kSlotThreads) {
            *reinterpret_cast<int2*>(smi + srp + o1 + 2 * idx) = __ldg(es + idx);
            *reinterpret_cast<int2*>(smi + scl + o2 + 2 * idx) = __ldg(ed + idx);
          }
          continue;
        }
        for (int idx = gt; idx <= n; idx += kSlotThreads) smi[srp + o1 + idx] = __ldg(a.f.rowptr + n0 + idx);
        for (int idx = gt; idx < e; idx += kSlotThreads) smi[scl + o2 + idx] = __ldg(a.f.col + e0 + idx);
        for (int idx = gt; idx < n; idx += kSlotThreads) smf[sds + o3 + idx] = __ldg(a.f.dis + n0 + idx);
      }
      ptx::named_bar_sync(bar_full, kSlotThreads);
    }
    if (PK && !pre_expanded) expand_x(mp, s);
    pre_expanded = false;
    // ---- this thread's row: which graph of the tile, where its slices start -------------------------
    const bool row_ok = i < ntot;
    const bool warp_ok = i_warp_min < ntot;  // this warp owns at least one row of the tile
    const bool inB = row_ok && i >= nA;      // (rows past the tile use graph A's addresses and weight 0)
    const int n0 = inB ? mp->n0B : mp->n0A, ng = inB ? nB : nA, e0 = inB ? mp->e0B : mp->e0A;
    const int rowbase = inB ? nA : 0;
    const int li = i - rowbase;               // row inside its graph
    const int jb = n0 - rowbase;              // tile row of batch node j = j - jb
    const int xt_o = PK ? sx : sx + (inB ? mp->off[4] : mp->off[0]) - rowbase * F;   // x[tile row r][c] at smf[xt_o + r F + c]
    const int nbo = COO ? nb_w + 2 * i                                    // COO: record t of this row at smi[nbo + 2 t kTcRows]
                        : nb_w + 2 * ((inB ? eA : 0) - e0);               // record of CSR slot pos at smi[nbo + 2 pos]
    int beg = 0, deg = 0;
    float di = 0.f;
    if (COO) {
      // ---- A. every edge claims a record of its target row -------------------------------------------
      int bad = 0;
      const int etot = eA + mp->eB;
      for (int k = gt; k < etot; k += kSlotThreads) {
        const int eb = k >= eA;                    // edge of graph B
        const int kk = eb ? k - eA : k;
        int2 sv, dv;
        int gn0 = eb ? mp->n0B : mp->n0A;
        const int gn = eb ? nB : nA, rb = eb ? nA : 0;
        if (PK) {     // a pair of bytes local to the graph
          const unsigned short pr = *reinterpret_cast<const unsigned short*>(reinterpret_cast<const unsigned char*>(smf + srp) +
                                                                             mp->off[4 * eb + 1] + 2 * kk);
          sv = make_int2(pr & 0xff, 0); dv = make_int2(pr >> 8, 0); gn0 = 0;
        } else {
          sv = *reinterpret_cast<const int2*>(smi + srp + mp->off[4 * eb + 1] + 2 * kk);
          dv = *reinterpret_cast<const int2*>(smi + scl + mp->off[4 * eb + 2] + 2 * kk);
        }
        const int ts = sv.x - gn0, td = dv.x - gn0;
        if (sv.y != 0 || dv.y != 0 || (unsigned)ts >= (unsigned)gn || (unsigned)td >= (unsigned)gn || ts == td) {
          bad |= kCooBadEdge;     // (a self-loop would be dropped and replaced by the fill-weight one: not handled here)
          continue;
        }
        const int pos = atomicAdd(smi + dg_w + rb + td, 1);
        if (pos < kTcCooDeg) smi[nb_w + 2 * (pos * kTcRows + rb + td)] = (k << 8) | (rb + ts);   // key: edge number, then the source's tile row
        else bad |= kCooDegree;
      }
      if (bad) atomicOr(&s_bad[slot], bad);
      ptx::named_bar_sync(bar_full, kSlotThreads);
      // ---- B. order the row's records by edge number, attach the norms ----------------------------------
      const int dtrue = row_ok ? smi[dg_w + i] : 0;
      deg = dtrue < kTcCooDeg ? dtrue : kTcCooDeg;
      di = row_ok ? s_lut[dtrue < 31 ? dtrue : 31] : 0.f;
      const int wmax = __reduce_max_sync(0xffffffffu, deg);
      if (cg == 0 && wmax > 0) {
        int* kp = smi + nbo;                        // key of record t at kp[2 t kTcRows]
        auto finish = [&](int t, int key) {        // (tile row of the source, norm)
          const int j = key & 0xff;
          const int dj = smi[dg_w + j];
          const float w = s_lut[dj < 31 ? dj : 31] * di;
          *reinterpret_cast<int2*>(kp + 2 * t * kTcRows) = make_int2(j, __float_as_int(w));
        };
        if (wmax <= 8) {   // sorting network in registers, sized by the warp's largest row
          int r[8];
#pragma unroll
          for (int t = 0; t < 8; ++t) r[t] = (t < deg && (t < 4 || wmax > 4)) ? kp[2 * t * kTcRows] : 0x7fffffff;
          if (wmax <= 4) {
            key_cx(r[0], r[1]); key_cx(r[2], r[3]); key_cx(r[0], r[2]); key_cx(r[1], r[3]); key_cx(r[1], r[2]);
          } else {         // 19 exchanges for eight keys
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
        } else {           // a row with 9..16 records somewhere in the warp: insertion sort in place
          for (int u = 1; u < deg; ++u) {
            const int key = kp[2 * u * kTcRows];
            int c = u - 1;
            while (c >= 0 && kp[2 * c * kTcRows] > key) { kp[2 * (c + 1) * kTcRows] = kp[2 * c * kTcRows]; --c; }
            kp[2 * (c + 1) * kTcRows] = key;
          }
          for (int t = 0; t < deg; ++t) finish(t, kp[2 * t * kTcRows]);
        }
      }
    } else {
      const int rp_o = srp + (inB ? mp->off[5] : mp->off[1]);             // rowptr[n0 + r] at smi[rp_o + r]
      const int cl_o = scl + (inB ? mp->off[6] : mp->off[2]) - e0;        // col[pos] at smi[cl_o + pos]
      const int ds_o = sds + (inB ? mp->off[7] : mp->off[3]) - n0;        // dis[j] at smf[ds_o + j]
      if (row_ok) { beg = smi[rp_o + li]; deg = smi[rp_o + li + 1] - beg; di = smf[ds_o + n0 + li]; }
      // ---- per CSR slot: (tile row of the neighbour, norm_ij); the two threads of a row take alternate slots ----
      for (int t = cg; t < deg; t += TPR) {
        const int j = smi[cl_o + beg + t];
        const bool ok = (unsigned)(j - n0) < (unsigned)ng;
        if (!ok && a.f.status) atomicOr(a.f.status, kStatusCrossEdgeTc);
        const float w = ok ? smf[ds_o + (ok ? j : n0)] * di : 0.f;
        *reinterpret_cast<int2*>(smi + nbo + 2 * (beg + t)) = make_int2(ok ? j - jb : i, __float_as_int(w));
      }
    }
    const float wself = (di * a.f.fill) * di;
    const int iself = row_ok ? i : 0;
    TC_TICK(2);
    ptx::named_bar_sync(bar_full, kSlotThreads);
    if (COO && cg == 0) smi[dg_w + i] = 0;   // every norm has been taken: counters ready for the next tile
    TC_TICK(3);
    // every lane of the warp walks the same number of neighbour steps (the warp's largest degree)
    const int maxdeg = __reduce_max_sync(0xffffffffu, deg);
    const int2* nbp = reinterpret_cast<const int2*>(smi + nbo) + beg;
    auto nb_at = [&](int t) { return t < deg ? nbp[t * NBS] : make_int2(iself, 0); };

    // ---- layer 0: a_r = sum_j norm_rj x_j ; this thread's XW of the (up to) 32 columns ---------------
    if (warp_ok) {
      float acc[XW];
#pragma unroll
      for (int c = 0; c < XW; ++c) acc[c] = 0.f;
      const int xc = xt_o + XW * cg;
      const int ncol = F - XW * cg;    // columns of this thread that exist (<= 0: none)
      if (F_ > 0) {
        // compile-time column count of every column group: min(XW, F - XW g), dispatched on the (warp-uniform) group
        constexpr int N0 = F_ < XW ? F_ : XW;
        constexpr int N1 = F_ - XW < 0 ? 0 : (F_ - XW < XW ? F_ - XW : XW);
        if (cg == 0) gather_x<N0, XW, NBS>(acc, smf + xc, F, nbp, deg, maxdeg, iself, wself);
        else gather_x<N1, XW, NBS>(acc, smf + xc, F, nbp, deg, maxdeg, iself, wself);
      } else if (ncol > 0) {
        gather_x<-1, XW, NBS>(acc, smf + xc, F, nbp, deg, maxdeg, iself, wself, ncol < XW ? ncol : XW);
      }
      tc::store_split16(acc, a_hi, a_lo, trow, cg * 16);
    }
    TC_TICK(4);

    for (int l = 0; l < L; ++l) {
      const bool last = l == L - 1;
      const uint32_t wt_hi = l == 0 ? ptx::smem_u32(w0_hi) : ptx::smem_u32(wl_base + (size_t)(l - 1) * 2 * tc::kN * 64);
      const uint32_t wt_lo = l == 0 ? ptx::smem_u32(w0_lo) : wt_hi + tc::kN * 64 * 4;
      if (!KSPLIT || l == 0) {
        ptx::tmem_wait_st();
        ptx::tc_fence_before_sync();
        if (ctl_warp) {
          ptx::named_bar_sync(bar_aop, kSlotThreads);  // A operand complete; every reader of the activation buffer is done
          if (lane == 0) {
            ptx::tc_fence_after_sync();
            if (l == 0) tc::issue_split_mma<4>(d_t, a_hi, a_lo, wt_hi, wt_lo);  // K = 32, zero padded
            else tc::issue_split_mma<8>(d_t, a_hi, a_lo, wt_hi, wt_lo);
            ptx::mma_commit(&bar_mma[slot]);
          }
          __syncwarp();
        } else {
          ptx::named_bar_arrive(bar_aop, kSlotThreads);
        }
      } else {
        // K-split layer: (the operand halves were gathered below, at the end of the previous layer)
        if (cg == 0) {
          ptx::tmem_wait_st();
          ptx::tc_fence_before_sync();
          if (ctl_warp) {
            ptx::named_bar_sync(bar_aop, 128);   // the four warps of half 0
            if (lane == 0) {
              ptx::tc_fence_after_sync();
              tc::issue_split_mma<4>(d_t, a_hi, a_lo, wt_hi, wt_lo);                         // k = 0..31
              ptx::mma_commit(&bar_free[slot]);
            }
            __syncwarp();
            ptx::named_bar_sync(bar_a1, 160);    // + the four warps of half 1
            if (lane == 0) {
              ptx::tc_fence_after_sync();
              tc::issue_split_mma<4>(d_t, a_hi, a_lo, wt_hi + tc::kChunkFloats * 4, wt_lo + tc::kChunkFloats * 4, 1u);  // k = 32..63, on top
              ptx::mma_commit(&bar_mma[slot]);
            }
            __syncwarp();
          } else {
            ptx::named_bar_arrive(bar_aop, 128);
          }
        } else {
          ptx::named_bar_arrive(bar_a1, 160);    // (half 1 stored its operand inside the gather block below)
        }
      }
      if (SAVE) {   // the rows the activation buffer holds: this tile's previous layer, or the previous tile's last one
        if (l > 0) save_rows(l - 1, nA, nB, n0A_t, n0B_t);
        else if (p_gA >= 0) save_rows(L - 1, p_nA, p_nB, p_n0A, p_n0B);
      }
      if (l == 0 && p_gA >= 0) pool_tile(p_gA, p_gB, p_nA, p_nB);   // the previous tile, while the MMAs run
      if (PK && l == 1) {   // the NEXT tile's x rows, while this layer's MMAs run
        ptx::mbar_wait(&bar_stage[slot][(k + 1) & 1], (uint32_t)((k + 1) >> 1) & 1u);
        const TcMeta* m2 = &metas[slot][(k + 1) & 1];
        if (m2->gA >= 0) expand_x(m2, (k + 1) & 1);
        pre_expanded = true;
      }
      if (ctl_warp) ptx::mbar_wait(&bar_mma[slot], mma_phase);      // one warp polls; the others sleep in the barrier below
      mma_phase ^= 1u;
      TC_TICK(5);
      ptx::named_bar_sync(bar_full, kSlotThreads);
      ptx::tc_fence_after_sync();
      if (l == 0 && pf_warp) issue(k + 1, true);   // every layer-0 gather of the slot is over: the x stage is free
      TC_TICK(6);

      // ---- epilogue: h_r = lrelu(D_r + b_l) -> activation buffer (+ optional save) ------------------
      if (warp_ok) {
        // (training forward: the finished row goes out below, as ONE bulk copy per row from the activation buffer)
#pragma unroll
        for (int part = 0; part < CW / 16; ++part) {
          uint32_t q0[16];
          ptx::tmem_ld16(d_t + trow + cg * CW + part * 16, q0);
          ptx::tmem_wait_ld();
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const int c0 = cg * CW + part * 16 + 4 * q4;
            const float4 b4 = *reinterpret_cast<const float4*>(s_bias + l * 64 + c0);
            float4 v;
            v.x = __uint_as_float(q0[4 * q4 + 0]) + b4.x; v.x = fmaxf(v.x, v.x * slope);  // LeakyReLU, 0 <= slope <= 1
            v.y = __uint_as_float(q0[4 * q4 + 1]) + b4.y; v.y = fmaxf(v.y, v.y * slope);
            v.z = __uint_as_float(q0[4 * q4 + 2]) + b4.z; v.z = fmaxf(v.z, v.z * slope);
            v.w = __uint_as_float(q0[4 * q4 + 3]) + b4.w; v.w = fmaxf(v.w, v.w * slope);
            *reinterpret_cast<float4*>(smf + hb_w + i * HP + c0) = v;
          }
        }
      }
      TC_TICK(7);
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_full, kSlotThreads);  // h complete in shared memory; D may be overwritten
      TC_TICK(8);

      if (!last) {
        // ---- next layer's A operand: a_r = sum_j norm_rj h_j ; this thread's CW of the 64 columns -----
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
        }
        if (KSPLIT && cg == 1) {
          // the first half's MMAs must have read the operand before it is overwritten
          ptx::mbar_wait(&bar_free[slot], free_phase);
          ptx::tc_fence_after_sync();
        }
        free_phase ^= 1u;
        if (warp_ok) {
#pragma unroll
          for (int cc = 0; cc < CW; cc += 16) {
            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = acc[cc + q];
            tc::store_split16(v, a_hi, a_lo, trow, KSPLIT ? cc : cg * CW + cc);
          }
        }
        if (KSPLIT && cg == 1) {
          ptx::tmem_wait_st();
          ptx::tc_fence_before_sync();
        }
      }
      TC_TICK(9);
    }
    p_gA = gA; p_gB = gB; p_nA = nA; p_nB = nB; p_n0A = n0A_t; p_n0B = n0B_t;
    TC_TICK(10);
  }
  if (SAVE && p_gA >= 0) save_rows(L - 1, p_nA, p_nB, p_n0A, p_n0B);
  if (p_gA >= 0) pool_tile(p_gA, p_gB, p_nA, p_nB);

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (COO && a.kflags && gt == 0) a.kflags[slot_id] = s_bad[slot];
  if (tid < 32) ptx::tmem_dealloc(tmem_base_s, 512);
}

size_t tc_fwd_smem_bytes(int F, int L) { return tc_layout(F, L, 2).total; }
int tc_fwd_row_cap() { return kTcRows; }
int tc_fwd_edge_cap() { return kTcEcap; }

}  // namespace hcg

namespace hcg {

TcPlan tc_plan(int64_t B, int ns) {
  TcPlan p{};
  int64_t grid = cdiv(B, ns);
  if (grid > kNumSMs) grid = kNumSMs;
  if (grid < 1) grid = 1;
  p.grid = (int)grid;
  p.nslots = p.grid * ns;
  // graphs per slot -> windows of at most kTcWinMax graphs, as equal as possible
  const int64_t gps = cdiv(B, p.nslots);
  const int64_t wins = cdiv(gps, kTcWinMax);
  int64_t W = cdiv(gps, wins);
  if (W < 1) W = 1;
  p.W = (int)W;
  p.nwin = (int)cdiv(B, W);
  p.cap = (int)(cdiv(p.nwin, p.nslots) * W);
  return p;
}

size_t tc_tiles_capacity(int64_t B) {
  size_t best = 0;
  for (int ns = 1; ns <= 3; ++ns) {
    const TcPlan p = tc_plan(B, ns);
    const size_t t = (size_t)p.nslots * p.cap;
    best = t > best ? t : best;
  }
  return best + 1;
}

int launch_tile_desc(const int32_t* graph_ptr, const int32_t* rowptr, const int32_t* col, const float* dis, const float* x, int F,
                     int64_t N, int64_t E, int64_t B, const TcPlan& p, TcTile* tiles, int* counts, int* overflow_count,
                     int* overflow_list, cudaStream_t st, int* overflow_flag, const int64_t* node_ptr64, const int64_t* edge_ptr64) {
  HCG_CUDA_OK(cudaMemsetAsync(overflow_count, 0, sizeof(int), st));
  k_tile_desc<<<(unsigned)cdiv((int64_t)p.nslots * 32, 128), 128, 0, st>>>(graph_ptr, rowptr, col, dis, x, F, N, E, B, p.W, p.nslots,
                                                                          p.nwin, p.cap, tiles, counts, overflow_count, overflow_list,
                                                                          overflow_flag, node_ptr64, edge_ptr64, nullptr, nullptr, nullptr, 0);
  HCG_LAUNCH_OK(st, "k_tile_desc");
  return HCG_OK;
}

struct TcWs {
  TcTile* tiles;
  int* counts;
  int* overflow_count;
  int* overflow_list;
  size_t bytes;
};

static TcWs carve_tc(void* ws, int64_t B) {
  Carver c(ws, (size_t)-1);
  TcWs w{};
  w.tiles = c.take<TcTile>(tc_tiles_capacity(B));
  w.counts = c.take<int>((size_t)3 * kNumSMs + 1);
  w.overflow_count = c.take<int>(64);
  w.overflow_list = c.take<int>((size_t)B + 1);
  w.bytes = c.off;
  return w;
}

size_t tc_fwd_workspace_bytes(int64_t B) { return carve_tc(nullptr, B).bytes; }

// slots per CTA: three when their shared memory fits (two conv layers), else two; HCG_TC_SLOTS=2|3 in the
// environment overrides (tests / tuning)
static int tc_slots(int F, int L) {
  static int env = -1;
  if (env < 0) {
    const char* e = getenv("HCG_TC_SLOTS");
    env = (e && e[0] == '2') ? 2 : (e && e[0] == '3') ? 3 : 0;
  }
  const bool fits3 = tc_layout(F, L, 3).total <= 227 * 1024;
  if (env == 2 || !fits3) return 2;
  return env == 3 ? 3 : kTcDefaultSlots;
}

template <int F_, int NS, int MODE = 0>
static int launch_tc_variant(const TcArgs& a, int grid, size_t smem, cudaStream_t st) {
  static SmemConfigured configured;
  if (int rc = ensure_dyn_smem(k_tc_fwd<F_, NS, MODE>, smem, configured)) return rc;
  k_tc_fwd<F_, NS, MODE><<<(unsigned)grid, NS * 256, smem, st>>>(a);
  return HCG_OK;
}

int launch_tc_fwd(const FusedFwdArgs& f, int64_t N, int64_t E, void* ws, cudaStream_t st, bool all_fit) {
  if (f.B <= 0) return HCG_OK;
  TcWs w = carve_tc(ws, f.B);
  const int ns = tc_slots(f.F, f.L);
  const TcPlan p = tc_plan(f.B, ns);
  const size_t smem = tc_layout(f.F, f.L, ns).total;
  if (int rc = launch_tile_desc(f.graph_ptr, f.rowptr, f.col, f.dis, f.x, f.F, N, E, f.B, p, w.tiles, w.counts, w.overflow_count,
                                w.overflow_list, st)) return rc;

  TcArgs a{};
  a.f = f;
  a.tiles = w.tiles;
  a.counts = w.counts;
  a.cap = p.cap;
  a.N = N; a.E = E;
  bool save = false;    // training forward: MODE 3 (the inference instantiations carry no trace of the row saving)
  for (int l = 0; l < f.L; ++l) save |= f.save_act[l] != nullptr;
  int rc;
  if (ns == 2 && save)
    rc = f.F == 25 ? launch_tc_variant<25, 2, 3>(a, p.grid, smem, st) : f.F == 32 ? launch_tc_variant<32, 2, 3>(a, p.grid, smem, st)
                                                                                  : launch_tc_variant<0, 2, 3>(a, p.grid, smem, st);
  else if (ns == 2)
    rc = f.F == 25 ? launch_tc_variant<25, 2>(a, p.grid, smem, st) : f.F == 32 ? launch_tc_variant<32, 2>(a, p.grid, smem, st)
                                                                               : launch_tc_variant<0, 2>(a, p.grid, smem, st);
  else if (save)
    rc = f.F == 25 ? launch_tc_variant<25, 3, 3>(a, p.grid, smem, st) : f.F == 32 ? launch_tc_variant<32, 3, 3>(a, p.grid, smem, st)
                                                                                  : launch_tc_variant<0, 3, 3>(a, p.grid, smem, st);
  else
    rc = f.F == 25 ? launch_tc_variant<25, 3>(a, p.grid, smem, st) : f.F == 32 ? launch_tc_variant<32, 3>(a, p.grid, smem, st)
                                                                               : launch_tc_variant<0, 3>(a, p.grid, smem, st);
  if (rc) return rc;
  HCG_LAUNCH_OK(st, "k_tc_fwd");

  // graphs above the tile capacity: the FMA graph-resident kernel walks the overflow list.  When the caller's hints
  // say that every graph fits, one CTA is enough to find the list empty (or to serve a wrong hint, slowly).
  FusedFwdArgs o = f;
  o.list = w.overflow_list;
  o.list_count = w.overflow_count;
  const size_t osmem = fused_fwd_smem_bytes(f.F, f.L, o.ncap, o.ecap);
  return launch_fused_fwd(o, osmem, st, all_fit ? 1 : 0);
}

// ---- COO variant: edge list in, embeddings out --------------------------------------------------------
TcCooWs carve_tc_coo(void* ws, int64_t B) {
  Carver c(ws, (size_t)-1);
  TcCooWs w{};
  w.tiles = c.take<TcTile>(tc_tiles_capacity(B));
  w.counts = c.take<int>((size_t)3 * kNumSMs + 1);
  w.dflags = c.take<int>((size_t)3 * kNumSMs + 1);
  w.kflags = c.take<int>((size_t)3 * kNumSMs + 1);
  w.bytes = c.off;
  return w;
}

size_t tc_fwd_coo_workspace_bytes(int64_t B) { return carve_tc_coo(nullptr, B).bytes; }
size_t tc_fwd_coo_smem_bytes(int F, int L) { return tc_layout(F, L, 2, true).total; }
int tc_fwd_coo_degree_cap() { return kTcCooDeg; }

int launch_tile_desc_coo(const float* x, int F, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N,
                         int64_t E, int64_t B, const TcPlan& p, const TcCooWs& w, float* emb, int emb_zero_cols, cudaStream_t st) {
  k_tile_desc<<<(unsigned)cdiv((int64_t)p.nslots * 32, 128), 128, 0, st>>>(nullptr, nullptr, nullptr, nullptr, x, F, N, E, B, p.W, p.nslots,
                                                                          p.nwin, p.cap, w.tiles, w.counts, nullptr, nullptr, nullptr,
                                                                          node_ptr, edge_ptr, edge_index, w.dflags, emb_zero_cols > 0 ? emb : nullptr,
                                                                          edge_index == nullptr ? 1 : 0);
  HCG_LAUNCH_OK(st, "k_tile_desc");
  return HCG_OK;
}

int launch_tc_fwd_coo(const FusedFwdArgs& f, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N,
                      int64_t E, void* ws, cudaStream_t st) {
  if (f.B <= 0) return HCG_OK;
  TcCooWs w = carve_tc_coo(ws, f.B);
  const TcPlan p = tc_plan(f.B, 2);
  const size_t smem = tc_layout(f.F, f.L, 2, true).total;
  if (int rc = launch_tile_desc_coo(f.x, f.F, edge_index, node_ptr, edge_ptr, N, E, f.B, p, w, f.emb, 128, st)) return rc;
  TcArgs a{};
  a.f = f;
  a.tiles = w.tiles;
  a.counts = w.counts;
  a.cap = p.cap;
  a.N = N; a.E = E;
  a.ei = edge_index;
  a.kflags = w.kflags;
  bool save = false;
  for (int l = 0; l < f.L; ++l) save |= f.save_act[l] != nullptr;
  const int rc = save ? (f.F == 25 ? launch_tc_variant<25, 2, 4>(a, p.grid, smem, st)
                       : f.F == 32 ? launch_tc_variant<32, 2, 4>(a, p.grid, smem, st)
                                   : launch_tc_variant<0, 2, 4>(a, p.grid, smem, st))
                      : (f.F == 25 ? launch_tc_variant<25, 2, 1>(a, p.grid, smem, st)
                       : f.F == 32 ? launch_tc_variant<32, 2, 1>(a, p.grid, smem, st)
                                   : launch_tc_variant<0, 2, 1>(a, p.grid, smem, st));
  if (rc) return rc;
  HCG_LAUNCH_OK(st, "k_tc_fwd_coo");
  return HCG_OK;
}

// ---- packed-image variant ------------------------------------------------------------------------------
// the staged slices of a tile must fit one stage: edge pairs + bit rows + byte rows + f32 rows, each with window slack
bool tc_fwd_packed_fits(const HcgPackedStore& pk) {
  const int ob = 2 * kTcEcap + 64;
  const int ou = ob + ((kTcRows * pk.bit_stride + 79) & ~15);
  const int of = ou + ((kTcRows * pk.Fu + 79) & ~15);
  const int end = of + ((kTcRows * pk.Ff * 4 + 79) & ~15);
  return pk.edge_kind == 1 && pk.F >= 1 && pk.F <= 32 && end <= (int)tc_layout(pk.F, 2, 2, true).csr_words * 4;
}

int launch_tc_fwd_packed(const FusedFwdArgs& f, const HcgPackedStore& pk, int64_t first, void* ws, cudaStream_t st) {
  if (f.B <= 0) return HCG_OK;
  TcCooWs w = carve_tc_coo(ws, f.B);
  const TcPlan p = tc_plan(f.B, 2);
  const size_t smem = tc_layout(f.F, f.L, 2, true).total;
  if (int rc = launch_tile_desc_coo(nullptr, f.F, nullptr, pk.node_ptr + first, pk.edge_ptr + first, pk.N, pk.E, f.B, p, w, f.emb, 128, st))
    return rc;
  TcArgs a{};
  a.f = f;
  a.tiles = w.tiles;
  a.counts = w.counts;
  a.cap = p.cap;
  a.N = pk.N; a.E = pk.E;
  a.kflags = w.kflags;
  a.pk = pk;
  const int rc = f.F == 25 ? launch_tc_variant<25, 2, 2>(a, p.grid, smem, st)
               : f.F == 32 ? launch_tc_variant<32, 2, 2>(a, p.grid, smem, st)
                           : launch_tc_variant<0, 2, 2>(a, p.grid, smem, st);
  if (rc) return rc;
  HCG_LAUNCH_OK(st, "k_tc_fwd_packed");
  return HCG_OK;
}

// OR of the refusal flags the last launch_tc_fwd_coo with this workspace left (synchronises the stream)
int tc_fwd_coo_flags(const void* ws, int64_t B, cudaStream_t st, int* flags) {
  *flags = 0;
  if (B <= 0) return HCG_OK;
  TcCooWs w = carve_tc_coo(const_cast<void*>(ws), B);
  const TcPlan p = tc_plan(B, 2);
  static thread_local int host[2 * (3 * kNumSMs + 1)];
  // dflags and kflags are adjacent up to the carver's alignment: two copies keep this independent of it
  HCG_CUDA_OK(cudaMemcpyAsync(host, w.dflags, sizeof(int) * p.nslots, cudaMemcpyDeviceToHost, st));
  HCG_CUDA_OK(cudaMemcpyAsync(host + p.nslots, w.kflags, sizeof(int) * p.nslots, cudaMemcpyDeviceToHost, st));
  HCG_CUDA_OK(cudaStreamSynchronize(st));
  for (int k = 0; k < 2 * p.nslots; ++k) *flags |= host[k];
  return HCG_OK;
}

}  // namespace hcg

#ifdef HCG_TC_TIMING
extern "C" int hcg_debug_tc_timing(unsigned long long* out16, int reset) {
  if (out16) cudaMemcpyFromSymbol(out16, hcg::g_tc_timing, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(hcg::g_tc_timing, z, sizeof(z)); }
  return 0;
}
#endif
