// Tensor-core graph-resident forward (see hcg_tc.cuh for the split-TF32 product) and a small
// self-test entry point that exercises exactly the same building blocks on one 128 x 64 x K tile.
#include "hcg_common.cuh"
#include "hcg_fused.cuh"
#include "hcg_tc.cuh"

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

}  // namespace hcg

using namespace hcg;

extern "C" int hcg_selftest_tc_gemm(const float* A, const float* W, int K, float* D, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(A && W && D, HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(K >= 1 && K <= 64, HCG_ERR_UNSUPPORTED, "K=%d outside [1,64]", K);
  const size_t smem = 2 * 64 * 64 * sizeof(float) + 1024;
  HCG_CUDA_OK(cudaFuncSetAttribute(k_tc_selftest, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_tc_selftest<<<1, 128, smem, st>>>(A, W, K, D);
  HCG_LAUNCH_OK(st, "k_tc_selftest");
  return HCG_OK;
}

// =================================================================================================
// Graph-resident forward on the tensor cores.
//
// A CTA holds two independent "slots" of 128 threads; a slot walks its share of the batch one graph
// at a time with thread r of the slot owning node r of the graph (= row r of the MMA tile = lane r
// of tensor memory):
//
//   prefetch   x rows, rowptr/col slice, deg^-1/2 of the NEXT graph: four bulk async copies
//              (cp.async.bulk -> mbarrier), issued one graph ahead, double-buffered
//   layer l    thread r gathers  a_r = sum_j norm_rj * h_j  (neighbours' rows from shared memory),
//              splits it into TF32 hi/lo and stores it as row r of the A operand in tensor memory;
//              one elected thread issues the split-TF32 tcgen05.mma chain against W_l^T (shared
//              memory, swizzled) and commits to an mbarrier; every thread then pulls row r of the
//              fp32 accumulator back (tcgen05.ld), adds bias, applies LeakyReLU and writes h_r to
//              the slot's activation buffer (16-byte chunks XOR-swizzled by row: conflict-free for
//              row-owner stores, column readers and neighbour gathers alike)
//   pooling    column owners scan the last layer's rows: emb[g] = [max || mean]
//
// The two slots of a CTA run out of phase, so one slot's gathers overlap the other's MMAs.
// Nothing of size [N,H] touches HBM unless the caller wants the activations (training forward).
// Graphs above the slot capacity (128 nodes / kTcEcap CSR slots) are appended to an overflow list
// that the FMA graph-resident kernel processes afterwards.
// =================================================================================================
namespace hcg {

constexpr int kTcSlots = 2;           // independent graph pipelines per CTA (tensor memory: 256 columns each)
constexpr int kTcSlotThreads = 256;  // two threads per node row: each owns half of the columns
constexpr int kTcThreads = kTcSlots * kTcSlotThreads;
constexpr int kTcNcap = 128;
constexpr int kTcEcap = 640;
constexpr int kTcHPitch = 68;        // floats per activation row: 8 consecutive rows cover all banks for 128-bit access
constexpr int kStatusCrossEdgeTc = 4;

struct TcSlotMeta {
  int n0, n, e0, e;
  int lead_x, lead_rp, lead_cl, lead_ds;  // element offsets of the data inside the 16-byte aligned copies
  int skip, pad0, pad1, pad2;
};

struct TcStage {
  float* x;       // [ncap*F + 4]
  int32_t* rp;    // [ncap + 1 + 4]
  int32_t* cl;    // [ecap + 4]
  float* ds;      // [ncap + 4]
};

struct TcArgs {
  FusedFwdArgs f;
  const int4* gdesc;     // [B] {n0, n, e0, e}; e < 0: not for this kernel (on the overflow list)
};

// The 16-byte granular window [lo, lo + bytes) the bulk copy of `count` 4-byte elements at `p` reads.
__host__ __device__ __forceinline__ void copy_window(uintptr_t p, int count, uintptr_t* lo, uint32_t* bytes, int* lead) {
  *lo = p & ~uintptr_t(15);
  *lead = (int)((p - *lo) >> 2);
  *bytes = (uint32_t)(((p - *lo) + (uintptr_t)count * 4 + 15) & ~uintptr_t(15));
}

// Per-graph descriptor + routing: graphs above the slot capacity and graphs without nodes or edges go to
// the overflow list (the FMA graph-resident kernel); everything else runs on the tensor-core kernel.
// A graph whose 16-byte granular copy windows would reach past the end of an input array (the last
// graph of a batch) is flagged (n stored negated): its slot fetches it with ordinary loads.
__global__ void __launch_bounds__(256)
k_graph_desc(const int32_t* __restrict__ graph_ptr, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
             const float* __restrict__ dis, const float* __restrict__ x, int F, int64_t N, int64_t E, int64_t B,
             int4* __restrict__ gdesc, int* __restrict__ overflow_count, int* __restrict__ overflow_list) {
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= B) return;
  const int n0 = graph_ptr[g], n = graph_ptr[g + 1] - n0;
  const int e0 = rowptr[n0], e = rowptr[n0 + n] - e0;
  const bool skip = n <= 0 || e <= 0 || n > kTcNcap || e > kTcEcap;
  bool manual = false;  // a 16-byte granular copy window would reach past the end of an input array
  if (!skip) {
    uintptr_t lo; uint32_t bytes; int lead;
    copy_window(reinterpret_cast<uintptr_t>(x + (int64_t)n0 * F), n * F, &lo, &bytes, &lead);
    manual |= lo + bytes > reinterpret_cast<uintptr_t>(x + N * F);
    copy_window(reinterpret_cast<uintptr_t>(rowptr + n0), n + 1, &lo, &bytes, &lead);
    manual |= lo + bytes > reinterpret_cast<uintptr_t>(rowptr + N + 1);
    copy_window(reinterpret_cast<uintptr_t>(col + e0), e, &lo, &bytes, &lead);
    manual |= lo + bytes > reinterpret_cast<uintptr_t>(col + E);
    copy_window(reinterpret_cast<uintptr_t>(dis + n0), n, &lo, &bytes, &lead);
    manual |= lo + bytes > reinterpret_cast<uintptr_t>(dis + N);
  }
  if (skip) overflow_list[atomicAdd(overflow_count, 1)] = (int)g;
  gdesc[g] = make_int4(n0, manual ? -n : n, e0, skip ? -1 : e);
}

// thread 0 of a slot: start the four copies of one graph into a stage (branch-free for routed graphs)
__device__ __forceinline__ void tc_prefetch(const TcArgs& a, int4 d, const TcStage& st, TcSlotMeta* meta, uint64_t* bar) {
  meta->n0 = d.x; meta->n = d.y < 0 ? -d.y : d.y; meta->e0 = d.z; meta->e = d.w;
  meta->skip = d.w < 0 ? 1 : (d.y < 0 ? 2 : 0);
  meta->lead_x = meta->lead_rp = meta->lead_cl = meta->lead_ds = 0;
  if (d.w < 0 || d.y < 0) {
    ptx::mbar_arrive(bar);
    return;
  }
  uintptr_t lo0, lo1, lo2, lo3;
  uint32_t b0, b1, b2, b3;
  int l0, l1, l2, l3;
  copy_window(reinterpret_cast<uintptr_t>(a.f.x + (int64_t)d.x * a.f.F), d.y * a.f.F, &lo0, &b0, &l0);
  copy_window(reinterpret_cast<uintptr_t>(a.f.rowptr + d.x), d.y + 1, &lo1, &b1, &l1);
  copy_window(reinterpret_cast<uintptr_t>(a.f.col + d.z), d.w, &lo2, &b2, &l2);
  copy_window(reinterpret_cast<uintptr_t>(a.f.dis + d.x), d.y, &lo3, &b3, &l3);
  // the metadata must be in place before the arrival that can complete the phase (release/acquire)
  meta->lead_x = l0; meta->lead_rp = l1; meta->lead_cl = l2; meta->lead_ds = l3;
  ptx::mbar_expect_tx(bar, b0 + b1 + b2 + b3);  // the slot's single arrival for this phase + the bytes in flight
  ptx::bulk_g2s(st.x, reinterpret_cast<const void*>(lo0), b0, bar);
  ptx::bulk_g2s(st.rp, reinterpret_cast<const void*>(lo1), b1, bar);
  ptx::bulk_g2s(st.cl, reinterpret_cast<const void*>(lo2), b2, bar);
  ptx::bulk_g2s(st.ds, reinterpret_cast<const void*>(lo3), b3, bar);
}

// shared-memory carve-up, identical on host (size query) and device
struct TcSmemLayout {
  uint32_t w0_hi, w0_lo, wl, bias, hbuf, wbuf, pool, stage;  // byte offsets from the 1024-aligned base
  uint32_t x_bytes, rp_bytes, cl_bytes, ds_bytes, stage_bytes, total;
};
__host__ __device__ inline TcSmemLayout tc_layout(int F, int L) {
  TcSmemLayout y;
  uint32_t p = 0;
  y.w0_hi = p; p += tc::kN * 32 * 4;
  y.w0_lo = p; p += tc::kN * 32 * 4;
  y.wl = p; p += (uint32_t)(L - 1) * 2 * tc::kN * 64 * 4;  // per layer: hi tile, lo tile
  y.bias = p; p += (uint32_t)L * 64 * 4;
  y.hbuf = p; p += (uint32_t)kTcSlots * kTcNcap * kTcHPitch * 4;
  y.wbuf = p; p += (uint32_t)kTcSlots * kTcEcap * 4;
  y.pool = p; p += (uint32_t)kTcSlots * 4 * 128 * 4;
  y.x_bytes = ((uint32_t)kTcNcap * F * 4 + 16 + 15) & ~15u;
  y.rp_bytes = ((uint32_t)(kTcNcap + 1) * 4 + 16 + 15) & ~15u;
  y.cl_bytes = ((uint32_t)kTcEcap * 4 + 16 + 15) & ~15u;
  y.ds_bytes = ((uint32_t)kTcNcap * 4 + 16 + 15) & ~15u;
  y.stage_bytes = y.x_bytes + y.rp_bytes + y.cl_bytes + y.ds_bytes;
  y.stage = p; p += (uint32_t)kTcSlots * 2 * y.stage_bytes;
  y.total = p + 1024;  // slack for the 1024-byte alignment of the base
  return y;
}

// One thread's share of a gather: NCH 16-byte chunks, starting at chunk q0, of the rows of `src` (pitch
// P floats):  acc = sum_k w_k * src[j_k] over the row's CSR slots, then the self-loop (last, as in the
// reference).  `jl` holds graph-local neighbour ids, `wn` the matching norm_ij (both per CSR slot).
template <int NCH, int P, int NACC>
__device__ __forceinline__ void gather_rows(float (&acc)[NACC], const float* src, const int32_t* jl,
                                            const float* wn, int beg, int end, int i, float wself, int q0) {
#pragma unroll
  for (int c = 0; c < NACC; ++c) acc[c] = 0.f;
  for (int kk = beg; kk <= end; ++kk) {
    int j = i;
    float w = wself;
    if (kk < end) { j = jl[kk]; w = wn[kk]; }
    const float4* row = reinterpret_cast<const float4*>(src + j * P) + q0;
#pragma unroll
    for (int q = 0; q < NCH; ++q) {
      const float4 v = row[q];
      acc[4 * q + 0] = fmaf(w, v.x, acc[4 * q + 0]);
      acc[4 * q + 1] = fmaf(w, v.y, acc[4 * q + 1]);
      acc[4 * q + 2] = fmaf(w, v.z, acc[4 * q + 2]);
      acc[4 * q + 3] = fmaf(w, v.w, acc[4 * q + 3]);
    }
  }
}

#ifdef HCG_TC_TIMING
__device__ unsigned long long g_tc_timing[16];
#define TC_TICK(slot_) do { if (tid == 0) { long long t_ = clock64(); g_tc_timing[slot_] += (unsigned long long)(t_ - t_prev); t_prev = t_; } } while (0)
#else
#define TC_TICK(slot_) do { } while (0)
#endif

// NQ = 16-byte chunks per re-pitched x row (7: F <= 28, 8: F <= 32); the pitch P keeps P/4 odd so that
// eight consecutive rows cover all 32 banks for 128-bit reads.
template <int NQ>
__global__ void __launch_bounds__(kTcThreads, 1)
k_tc_fwd(const TcArgs a) {
#ifdef HCG_TC_TIMING
  long long t_prev = clock64();
#endif
  constexpr int P = (NQ & 1) ? 4 * NQ : 4 * NQ + 4;
  constexpr int HP = kTcHPitch;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage[kTcSlots][2];
  __shared__ uint64_t bar_mma[kTcSlots];
  __shared__ uint32_t tmem_base_s;
  __shared__ TcSlotMeta metas[kTcSlots][2];

  const int tid = threadIdx.x;
  const int slot = tid / kTcSlotThreads, gt = tid % kTcSlotThreads;
  const int half = gt >> 7;          // which half of the columns this thread works on
  // Node (row of the graph) owned by this thread.  Its TMEM lane is fixed by the hardware (lane quadrant
  // = warp id % 4 = the warp scheduler), and graphs fill rows from 0 up, so slot 1 maps its nodes in
  // reverse: slot 0 loads schedulers 0,1 the most, slot 1 schedulers 3,2.
  const int i = slot == 0 ? (gt & 127) : 127 - (gt & 127);
  const int i_warp_min = slot == 0 ? (gt & 96) : 96 - (gt & 96);  // smallest node id owned by this warp
  // the warp of each slot whose quadrant is filled last does the waiting and the MMA issue for the slot
  const bool ctl_warp = (gt >> 5) == (slot == 0 ? 3 : 0);
  const int lane = tid & 31;
  // the control warp's first lane also feeds the slot: it issues the next graph's bulk copies while the
  // layer-0 MMAs run (it would only be polling their mbarrier), off the other warps' critical path
  const bool pf_thread = ctl_warp && lane == 0;
  const int F = a.f.F, L = a.f.L;

  // ---- carve shared memory (pointer arithmetic on the shared array only: keeps the address space) --
  const TcSmemLayout lay = tc_layout(F, L);
  unsigned char* base = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  float* w0_hi = reinterpret_cast<float*>(base + lay.w0_hi);
  float* w0_lo = reinterpret_cast<float*>(base + lay.w0_lo);
  float* wl_base = reinterpret_cast<float*>(base + lay.wl);
  float* s_bias = reinterpret_cast<float*>(base + lay.bias);
  float* Hbuf = reinterpret_cast<float*>(base + lay.hbuf) + (size_t)slot * kTcNcap * HP;
  float* wn = reinterpret_cast<float*>(base + lay.wbuf) + (size_t)slot * kTcEcap;
  float* s_pool = reinterpret_cast<float*>(base + lay.pool) + slot * 4 * 128;
  unsigned char* stage0 = base + lay.stage + (size_t)slot * 2 * lay.stage_bytes;
  auto stage_of = [&](int s) {
    TcStage t;
    unsigned char* q = stage0 + (size_t)s * lay.stage_bytes;
    t.x = reinterpret_cast<float*>(q); q += lay.x_bytes;
    t.rp = reinterpret_cast<int32_t*>(q); q += lay.rp_bytes;
    t.cl = reinterpret_cast<int32_t*>(q); q += lay.cl_bytes;
    t.ds = reinterpret_cast<float*>(q);
    return t;
  };

  // ---- one-time setup ----------------------------------------------------------------------------
  if (tid < 32) ptx::tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) {
    for (int s = 0; s < kTcSlots; ++s) {
      ptx::mbar_init(&bar_stage[s][0], 1);
      ptx::mbar_init(&bar_stage[s][1], 1);
      ptx::mbar_init(&bar_mma[s], 1);
    }
    ptx::fence_mbar_init();
  }
  tc::prep_weight_tiles(a.f.conv_w[0], F, 32, w0_hi, w0_lo, tid, kTcThreads);
  for (int l = 1; l < L; ++l)
    tc::prep_weight_tiles(a.f.conv_w[l], 64, 64, wl_base + (size_t)(l - 1) * 2 * tc::kN * 64,
                          wl_base + (size_t)(l - 1) * 2 * tc::kN * 64 + tc::kN * 64, tid, kTcThreads);
  for (int idx = tid; idx < L * 64; idx += kTcThreads) s_bias[idx] = __ldg(a.f.conv_b[idx >> 6] + (idx & 63));
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();

  const uint32_t tbase = tmem_base_s + (uint32_t)slot * 256;  // accumulator | (free) | A hi | A lo, 64 columns each
  const uint32_t d_t = tbase, a_hi = tbase + 128, a_lo = tbase + 192;
  const uint32_t trow = (uint32_t)(((gt >> 5) & 3) * 32) << 16;  // warps w and w+4 share a lane quadrant
  const int bar_id = 1 + slot;
  const float slope = a.f.slope;

  const int64_t first = (int64_t)blockIdx.x * kTcSlots + slot;
  const int64_t stride = (int64_t)gridDim.x * kTcSlots;
  int4 d_next = make_int4(0, 0, 0, -1);
  if (pf_thread && first < a.f.B) {
    tc_prefetch(a, a.gdesc[first], stage_of(0), &metas[slot][0], &bar_stage[slot][0]);
    if (first + stride < a.f.B) d_next = a.gdesc[first + stride];
  }
  uint32_t mma_phase = 0;
  int k = 0;
  for (int64_t gid = first; gid < a.f.B; gid += stride, ++k) {
    const int s = k & 1;
    // next graph -> the other stage; its readers finished with the previous iteration.  Issued by pf_thread
    // either right here (this graph is skipped) or after the layer-0 MMAs have been committed.
    auto issue_next = [&]() {
      if (gid + stride < a.f.B) {
        tc_prefetch(a, d_next, stage_of(s ^ 1), &metas[slot][s ^ 1], &bar_stage[slot][s ^ 1]);
        if (gid + 2 * stride < a.f.B) d_next = a.gdesc[gid + 2 * stride];
      }
    };
    TC_TICK(0);
    ptx::mbar_wait(&bar_stage[slot][s], (uint32_t)(k >> 1) & 1u);
    TC_TICK(1);
    const TcSlotMeta m = metas[slot][s];
    if (m.skip == 1) {  // slot-uniform: this graph is on the overflow list
      if (pf_thread) issue_next();
      continue;
    }
    const int n = m.n, n0 = m.n0, e0 = m.e0;
    const TcStage st = stage_of(s);
    if (m.skip == 2) {  // flagged graph: fetch the four slices with ordinary loads
      for (int idx = gt; idx < n * F; idx += kTcSlotThreads) st.x[idx] = __ldg(a.f.x + (int64_t)n0 * F + idx);
      for (int idx = gt; idx <= n; idx += kTcSlotThreads) st.rp[idx] = __ldg(a.f.rowptr + n0 + idx);
      for (int idx = gt; idx < m.e; idx += kTcSlotThreads) st.cl[idx] = __ldg(a.f.col + e0 + idx);
      for (int idx = gt; idx < n; idx += kTcSlotThreads) st.ds[idx] = __ldg(a.f.dis + n0 + idx);
      ptx::named_bar_sync(bar_id, kTcSlotThreads);
    }
    const float* xs = st.x + m.lead_x;
    const int32_t* rp = st.rp + m.lead_rp;
    int32_t* cl = st.cl + m.lead_cl;
    const float* ds = st.ds + m.lead_ds;
    const bool row_ok = i < n;
    const bool warp_ok = i_warp_min < n;  // this warp owns at least one node of the graph

    // ---- re-pitch the x rows to P floats (zero padded) inside the idle activation buffer: a warp per
    //      row, a lane per column ------------------------------------------------------------------------
    for (int r0 = gt >> 5; r0 < n; r0 += 4 * (kTcSlotThreads / 32)) {  // four rows in flight per warp
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int r = r0 + u * (kTcSlotThreads / 32);
        v[u] = (r < n && lane < F) ? xs[r * F + lane] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int r = r0 + u * (kTcSlotThreads / 32);
        if (r < n && lane < (P < 32 ? P : 32)) Hbuf[r * P + lane] = v[u];
      }
    }
    // ---- per CSR slot: graph-local neighbour id (in place) and norm_ij; the two threads of a row take
    //      alternate slots --------------------------------------------------------------------------------
    int beg = 0, end = -1;
    float di = 0.f;
    if (row_ok) {
      beg = rp[i] - e0; end = rp[i + 1] - e0; di = ds[i];
      for (int kk = beg + half; kk < end; kk += 2) {
        int j = cl[kk] - n0;
        const bool ok = (unsigned)j < (unsigned)n;
        if (!ok) { if (a.f.status) atomicOr(a.f.status, kStatusCrossEdgeTc); j = i; }
        cl[kk] = j;
        wn[kk] = ok ? ds[j] * di : 0.f;
      }
    }
    const float wself = (di * a.f.fill) * di;
    TC_TICK(2);
    ptx::named_bar_sync(bar_id, kTcSlotThreads);
    TC_TICK(3);

    // ---- layer 0: a_r = sum_j norm_rj x_j ; this thread's 16 of the (up to) 32 columns ---------------
    if (warp_ok) {
      float acc[16];
      if (half == 0) gather_rows<(NQ < 4 ? NQ : 4), P, 16>(acc, Hbuf, cl, wn, beg, end, i, wself, 0);
      else           gather_rows<(NQ > 4 ? NQ - 4 : 0), P, 16>(acc, Hbuf, cl, wn, beg, end, i, wself, 4);
      tc::store_split16(acc, a_hi, a_lo, trow, half * 16);
    }

    for (int l = 0; l < L; ++l) {
      const bool last = l == L - 1;
      TC_TICK(4);
      ptx::tmem_wait_st();
      TC_TICK(5);
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_id, kTcSlotThreads);  // A operand complete; previous readers of Hbuf / D are done
      if (ctl_warp) {
       if (lane == 0) {
        ptx::tc_fence_after_sync();
        if (l == 0)
          tc::issue_split_mma<4>(d_t, a_hi, a_lo, ptx::smem_u32(w0_hi), ptx::smem_u32(w0_lo));  // K = 32, zero padded
        else
          tc::issue_split_mma<8>(d_t, a_hi, a_lo, ptx::smem_u32(wl_base + (size_t)(l - 1) * 2 * tc::kN * 64),
                                 ptx::smem_u32(wl_base + (size_t)(l - 1) * 2 * tc::kN * 64 + tc::kN * 64));
        ptx::mma_commit(&bar_mma[slot]);
        if (l == 0) issue_next();
       }
       __syncwarp();
       ptx::mbar_wait(&bar_mma[slot], mma_phase);  // one warp polls; the others sleep in the barrier below
      }
      mma_phase ^= 1u;
      TC_TICK(6);
      ptx::named_bar_sync(bar_id, kTcSlotThreads);
      ptx::tc_fence_after_sync();
      TC_TICK(7);

      // ---- epilogue: h_r = lrelu(D_r + b_l) -> activation buffer (+ optional save) ------------------
      if (warp_ok) {
        float* save = a.f.save_act[l];
        uint32_t r0[16], r1[16];
        ptx::tmem_ld16(d_t + trow + half * 32, r0);
        ptx::tmem_ld16(d_t + trow + half * 32 + 16, r1);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int q4 = 0; q4 < 8; ++q4) {
          const int c0 = half * 32 + 4 * q4;
          const float4 b4 = *reinterpret_cast<const float4*>(s_bias + l * 64 + c0);
          const uint32_t* rr = q4 < 4 ? &r0[4 * q4] : &r1[4 * (q4 - 4)];
          float4 v;
          v.x = __uint_as_float(rr[0]) + b4.x; v.x = fmaxf(v.x, v.x * slope);  // LeakyReLU, 0 <= slope <= 1
          v.y = __uint_as_float(rr[1]) + b4.y; v.y = fmaxf(v.y, v.y * slope);
          v.z = __uint_as_float(rr[2]) + b4.z; v.z = fmaxf(v.z, v.z * slope);
          v.w = __uint_as_float(rr[3]) + b4.w; v.w = fmaxf(v.w, v.w * slope);
          *reinterpret_cast<float4*>(Hbuf + i * HP + c0) = v;
          if (save && row_ok) *reinterpret_cast<float4*>(save + ((int64_t)n0 + i) * 64 + c0) = v;
        }
      }
      TC_TICK(8);
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_id, kTcSlotThreads);  // h complete in shared memory; D may be overwritten
      TC_TICK(9);

      if (!last && warp_ok) {
        // ---- next layer's A operand: a_r = sum_j norm_rj h_j ; this thread's 32 of the 64 columns ----
        float acc[32];
        gather_rows<8, HP, 32>(acc, Hbuf, cl, wn, beg, end, i, wself, half * 8);
#pragma unroll
        for (int cc = 0; cc < 32; cc += 16) {
          float v[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) v[q] = acc[cc + q];
          tc::store_split16(v, a_hi, a_lo, trow, half * 32 + cc);
        }
      }
    }

    // ---- pooling over the graph's rows (column owners; four interleaved quarters of the rows) ------
    {
      const int c = gt & 63, quarter = gt >> 6;
      float mx = -INFINITY, sm = 0.f;
      const float* hp = Hbuf + c;
      int r = quarter;
      for (; r + 12 < n; r += 16) {
        const float v0 = hp[r * HP], v1 = hp[(r + 4) * HP], v2 = hp[(r + 8) * HP], v3 = hp[(r + 12) * HP];
        mx = fmaxf(fmaxf(mx, v0), fmaxf(v1, fmaxf(v2, v3)));
        sm += v0; sm += v1; sm += v2; sm += v3;
      }
      for (; r < n; r += 4) {
        const float v0 = hp[r * HP];
        mx = fmaxf(mx, v0);
        sm += v0;
      }
      TC_TICK(10);
      s_pool[quarter * 128 + c] = mx;
      s_pool[quarter * 128 + 64 + c] = sm;
      ptx::named_bar_sync(bar_id, kTcSlotThreads);
      if (gt < 64) {
        mx = fmaxf(fmaxf(s_pool[c], s_pool[128 + c]), fmaxf(s_pool[256 + c], s_pool[384 + c]));
        sm = ((s_pool[64 + c] + s_pool[192 + c]) + s_pool[320 + c]) + s_pool[448 + c];
        a.f.emb[gid * 128 + c] = mx;
        a.f.emb[gid * 128 + 64 + c] = sm / (float)n;
      }
      TC_TICK(11);
    }
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (tid < 32) ptx::tmem_dealloc(tmem_base_s, 512);
}

size_t tc_fwd_smem_bytes(int F, int L) { return tc_layout(F, L).total; }

}  // namespace hcg

namespace hcg {

struct TcWs {
  int4* gdesc;
  int* overflow_count;
  int* overflow_list;
  size_t bytes;
};

static TcWs carve_tc(void* ws, int64_t B) {
  Carver c(ws, (size_t)-1);
  TcWs w{};
  w.gdesc = c.take<int4>((size_t)B + 1);
  w.overflow_count = c.take<int>(64);
  w.overflow_list = c.take<int>((size_t)B + 1);
  w.bytes = c.off;
  return w;
}

size_t tc_fwd_workspace_bytes(int64_t B) { return carve_tc(nullptr, B).bytes; }

int launch_tc_fwd(const FusedFwdArgs& f, int64_t N, int64_t E, void* ws, cudaStream_t st) {
  if (f.B <= 0) return HCG_OK;
  TcWs w = carve_tc(ws, f.B);
  const size_t smem = tc_fwd_smem_bytes(f.F, f.L);
  static thread_local size_t configured[2] = {0, 0};
  const int v = f.F <= 28 ? 0 : 1;
  if (smem > configured[v]) {
    HCG_CUDA_OK(cudaFuncSetAttribute(v == 0 ? k_tc_fwd<7> : k_tc_fwd<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[v] = smem;
  }
  HCG_CUDA_OK(cudaMemsetAsync(w.overflow_count, 0, sizeof(int), st));
  k_graph_desc<<<(unsigned)cdiv(f.B, 256), 256, 0, st>>>(f.graph_ptr, f.rowptr, f.col, f.dis, f.x, f.F, N, E, f.B, w.gdesc,
                                                         w.overflow_count, w.overflow_list);
  HCG_LAUNCH_OK(st, "k_graph_desc");

  TcArgs a{};
  a.f = f;
  a.gdesc = w.gdesc;
  int64_t grid = cdiv(f.B, kTcSlots);
  if (grid > kNumSMs) grid = kNumSMs;
  if (v == 0) k_tc_fwd<7><<<(unsigned)grid, kTcThreads, smem, st>>>(a);
  else k_tc_fwd<8><<<(unsigned)grid, kTcThreads, smem, st>>>(a);
  HCG_LAUNCH_OK(st, "k_tc_fwd");

  // graphs above the slot capacity: the FMA graph-resident kernel walks the overflow list
  FusedFwdArgs o = f;
  o.list = w.overflow_list;
  o.list_count = w.overflow_count;
  const size_t osmem = fused_fwd_smem_bytes(f.F, f.L, o.ncap, o.ecap);
  return launch_fused_fwd(o, osmem, st);
}

}  // namespace hcg

#ifdef HCG_TC_TIMING
extern "C" int hcg_debug_tc_timing(unsigned long long* out16, int reset) {
  if (out16) cudaMemcpyFromSymbol(out16, hcg::g_tc_timing, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(hcg::g_tc_timing, z, sizeof(z)); }
  return 0;
}
#endif
