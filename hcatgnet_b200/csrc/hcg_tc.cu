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
    tc::issue_split_mma(d_t, a_hi, a_lo, ptx::smem_u32(s_hi), ptx::smem_u32(s_lo), Kpad / 8);
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
constexpr int kTcMaxF = 32;
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

__device__ __forceinline__ int h_off(int r, int c) {  // float offset of (row, col) in a swizzled [128][64] buffer
  const int q = c >> 2;
  return r * 64 + ((((q ^ r) & 7) | (q & 8)) << 2) + (c & 3);
}

// one contiguous slice global -> shared through the bulk-copy engine; returns the bytes in flight.
// `count` elements of 4 bytes starting at `src`; data lands at dst + lead (elements).  A tail the
// 16-byte granularity would read past `src_end` is fetched with ordinary loads instead.
__device__ __forceinline__ uint32_t bulk_slice(void* dst, const void* src, int count, const void* src_end, int* lead,
                                               uint64_t* bar) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(src);
  const uintptr_t a_al = a & ~uintptr_t(15);
  *lead = (int)((a - a_al) >> 2);
  if (count <= 0) return 0;
  uint32_t bytes = (uint32_t)(((a - a_al) + (uintptr_t)count * 4 + 15) & ~uintptr_t(15));
  const uintptr_t avail = (reinterpret_cast<uintptr_t>(src_end) - a_al) & ~uintptr_t(15);
  if (bytes > avail) {  // last elements of the array: copy them by hand
    const int done = (int)((avail - (a - a_al)) >> 2);
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
    uint32_t* d32 = reinterpret_cast<uint32_t*>(dst) + *lead;
    for (int i = done < 0 ? 0 : done; i < count; ++i) d32[i] = __ldg(s32 + i);
    bytes = (uint32_t)avail;
  }
  if (bytes) ptx::bulk_g2s(dst, reinterpret_cast<const void*>(a_al), bytes, bar);
  return bytes;
}

struct TcArgs {
  FusedFwdArgs f;
  const int4* gdesc;     // [B] {n0, n, e0, e}
  int* overflow_count;
  int* overflow_list;    // [B]
  const void* x_end;
  const void* rp_end;
  const void* cl_end;
  const void* ds_end;
};

__global__ void __launch_bounds__(256)
k_graph_desc(const int32_t* __restrict__ graph_ptr, const int32_t* __restrict__ rowptr, int64_t B, int4* __restrict__ gdesc) {
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= B) return;
  int n0 = graph_ptr[g], n1 = graph_ptr[g + 1];
  int e0 = rowptr[n0], e1 = rowptr[n1];
  gdesc[g] = make_int4(n0, n1 - n0, e0, e1 - e0);
}

// thread 0 of a slot: start the copies of one graph into a stage (or mark it skipped / overflowed)
__device__ __forceinline__ void tc_prefetch(const TcArgs& a, int64_t gid, int4 d, const TcStage& st, TcSlotMeta* meta,
                                            uint64_t* bar) {
  meta->n0 = d.x; meta->n = d.y; meta->e0 = d.z; meta->e = d.w;
  if (d.y > kTcNcap || d.w > kTcEcap) {
    meta->skip = 1;
    const int pos = atomicAdd(a.overflow_count, 1);
    a.overflow_list[pos] = (int)gid;
    ptx::mbar_arrive(bar);
    return;
  }
  meta->skip = 0;
  uint32_t bytes = 0;
  bytes += bulk_slice(st.x, a.f.x + (int64_t)d.x * a.f.F, d.y * a.f.F, a.x_end, &meta->lead_x, bar);
  bytes += bulk_slice(st.rp, a.f.rowptr + d.x, d.y + 1, a.rp_end, &meta->lead_rp, bar);
  bytes += bulk_slice(st.cl, a.f.col + d.z, d.w, a.cl_end, &meta->lead_cl, bar);
  bytes += bulk_slice(st.ds, a.f.dis + d.x, d.y, a.ds_end, &meta->lead_ds, bar);
  ptx::mbar_expect_tx(bar, bytes);  // the slot's single arrival for this phase + the bytes in flight
}

// shared-memory carve-up, identical on host (size query) and device
struct TcSmemLayout {
  uint32_t w0_hi, w0_lo, wl, bias, hbuf, pool, stage;  // byte offsets from the 1024-aligned base
  uint32_t x_bytes, rp_bytes, cl_bytes, ds_bytes, stage_bytes, total;
};
__host__ __device__ inline TcSmemLayout tc_layout(int F, int L) {
  TcSmemLayout y;
  uint32_t p = 0;
  y.w0_hi = p; p += tc::kN * 32 * 4;
  y.w0_lo = p; p += tc::kN * 32 * 4;
  y.wl = p; p += (uint32_t)(L - 1) * 2 * tc::kN * 64 * 4;  // per layer: hi tile, lo tile
  y.bias = p; p += (uint32_t)L * 64 * 4;
  y.hbuf = p; p += (uint32_t)kTcSlots * kTcNcap * 64 * 4;
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

// Layer-0 gather of one thread: NCH 16-byte chunks of the rows of `xp` (pitch P floats), starting at
// chunk `q0`: acc += w * x_j for every neighbour, then the self-loop.
template <int NCH, int P>
__device__ __forceinline__ void gather_x(float (&acc)[16], const float* xp, const int32_t* cl, const float* ds,
                                         int beg, int end, int i, int n, int n0, float di, float wself, int q0,
                                         int* status) {
#pragma unroll
  for (int c = 0; c < 16; ++c) acc[c] = 0.f;
  for (int kk = beg; kk <= end; ++kk) {  // kk == end: the self-loop, last as in the reference
    int j = i;
    float w = wself;
    if (kk < end) {
      j = cl[kk] - n0;
      const bool ok = (unsigned)j < (unsigned)n;
      if (!ok) { if (status) atomicOr(status, kStatusCrossEdgeTc); j = i; }
      w = ok ? ds[j] * di : 0.f;
    }
    const float4* xr = reinterpret_cast<const float4*>(xp + j * P) + q0;
#pragma unroll
    for (int q = 0; q < NCH; ++q) {
      const float4 v = xr[q];
      acc[4 * q + 0] = fmaf(w, v.x, acc[4 * q + 0]);
      acc[4 * q + 1] = fmaf(w, v.y, acc[4 * q + 1]);
      acc[4 * q + 2] = fmaf(w, v.z, acc[4 * q + 2]);
      acc[4 * q + 3] = fmaf(w, v.w, acc[4 * q + 3]);
    }
  }
}

// NQ = 16-byte chunks per re-pitched x row (7: F <= 28, 8: F <= 32); the pitch P keeps P/4 odd so that
// eight consecutive rows cover all 32 banks for 128-bit reads.
template <int NQ>
__global__ void __launch_bounds__(kTcThreads, 1)
k_tc_fwd(const TcArgs a) {
  constexpr int P = (NQ & 1) ? 4 * NQ : 4 * NQ + 4;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage[kTcSlots][2];
  __shared__ uint64_t bar_mma[kTcSlots];
  __shared__ uint32_t tmem_base_s;
  __shared__ TcSlotMeta metas[kTcSlots][2];

  const int tid = threadIdx.x;
  const int slot = tid / kTcSlotThreads, gt = tid % kTcSlotThreads;
  const int half = gt >> 7;          // which half of the columns this thread works on
  const int i = gt & 127;            // node (row) of the graph = TMEM lane
  const int F = a.f.F, L = a.f.L;

  // ---- carve shared memory (pointer arithmetic on the shared array only: keeps the address space) --
  const TcSmemLayout lay = tc_layout(F, L);
  unsigned char* base = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  float* w0_hi = reinterpret_cast<float*>(base + lay.w0_hi);
  float* w0_lo = reinterpret_cast<float*>(base + lay.w0_lo);
  float* wl_base = reinterpret_cast<float*>(base + lay.wl);
  float* s_bias = reinterpret_cast<float*>(base + lay.bias);
  float* Hbuf = reinterpret_cast<float*>(base + lay.hbuf) + (size_t)slot * kTcNcap * 64;
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

  const uint32_t tbase = tmem_base_s + (uint32_t)slot * 256;  // big acc | small acc | A hi | A lo, 64 columns each
  const uint32_t d_t = tbase, a_hi = tbase + 128, a_lo = tbase + 192;
  const uint32_t trow = (uint32_t)(((gt >> 5) & 3) * 32) << 16;  // warps w and w+4 share a lane quadrant
  const int bar_id = 1 + slot;
  const int ksteps0 = (F + 7) >> 3;

  const int64_t first = (int64_t)blockIdx.x * kTcSlots + slot;
  const int64_t stride = (int64_t)gridDim.x * kTcSlots;
  int4 d_next = make_int4(0, 0, 0, 0);
  if (gt == 0 && first < a.f.B) {
    tc_prefetch(a, first, a.gdesc[first], stage_of(0), &metas[slot][0], &bar_stage[slot][0]);
    if (first + stride < a.f.B) d_next = a.gdesc[first + stride];
  }
  uint32_t mma_phase = 0;
  int k = 0;
  for (int64_t gid = first; gid < a.f.B; gid += stride, ++k) {
    const int s = k & 1;
    if (gt == 0 && gid + stride < a.f.B) {  // next graph -> the other stage (its readers finished last iteration)
      tc_prefetch(a, gid + stride, d_next, stage_of(s ^ 1), &metas[slot][s ^ 1], &bar_stage[slot][s ^ 1]);
      if (gid + 2 * stride < a.f.B) d_next = a.gdesc[gid + 2 * stride];
    }
    ptx::mbar_wait(&bar_stage[slot][s], (uint32_t)(k >> 1) & 1u);
    const TcSlotMeta m = metas[slot][s];
    if (m.skip) continue;  // slot-uniform
    const int n = m.n, n0 = m.n0, e0 = m.e0;
    const TcStage st = stage_of(s);
    const float* xs = st.x + m.lead_x;
    const int32_t* rp = st.rp + m.lead_rp;
    const int32_t* cl = st.cl + m.lead_cl;
    const float* ds = st.ds + m.lead_ds;
    const bool row_ok = i < n;
    const bool warp_ok = (i & ~31) < n;  // this warp owns at least one node of the graph

    // ---- re-pitch the x rows to P floats (zero padded) inside the idle activation buffer ----------
    for (int idx = gt; idx < n * P; idx += kTcSlotThreads) {
      const int r = idx / P, c = idx - r * P;
      Hbuf[idx] = c < F ? xs[r * F + c] : 0.f;
    }
    int beg = 0, end = -1;
    float di = 0.f;
    if (row_ok) { beg = rp[i] - e0; end = rp[i + 1] - e0; di = ds[i]; }
    const float wself = (di * a.f.fill) * di;
    ptx::named_bar_sync(bar_id, kTcSlotThreads);

    // ---- layer 0: a_r = sum_j norm_rj x_j ; this thread's 16 of the (up to) 32 columns ---------------
    if (warp_ok && half * 16 < ksteps0 * 8) {
      float acc[16];
      if (half == 0) gather_x<(NQ < 4 ? NQ : 4), P>(acc, Hbuf, cl, ds, beg, end, i, n, n0, di, wself, 0, a.f.status);
      else           gather_x<(NQ > 4 ? NQ - 4 : 0), P>(acc, Hbuf, cl, ds, beg, end, i, n, n0, di, wself, 4, nullptr);
      tc::store_split16(acc, a_hi, a_lo, trow, half * 16);
    }

    for (int l = 0; l < L; ++l) {
      const bool last = l == L - 1;
      ptx::tmem_wait_st();
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_id, kTcSlotThreads);  // A operand complete; previous readers of Hbuf / D are done
      if (gt == 0) {
        ptx::tc_fence_after_sync();
        if (l == 0)
          tc::issue_split_mma(d_t, a_hi, a_lo, ptx::smem_u32(w0_hi), ptx::smem_u32(w0_lo), ksteps0);
        else
          tc::issue_split_mma(d_t, a_hi, a_lo, ptx::smem_u32(wl_base + (size_t)(l - 1) * 2 * tc::kN * 64),
                              ptx::smem_u32(wl_base + (size_t)(l - 1) * 2 * tc::kN * 64 + tc::kN * 64), 8);
        ptx::mma_commit(&bar_mma[slot]);
      }
      ptx::mbar_wait(&bar_mma[slot], mma_phase);
      mma_phase ^= 1u;
      ptx::tc_fence_after_sync();

      // ---- epilogue: h_r = lrelu(D_r + b_l) -> swizzled activation buffer (+ optional save) ------
      if (warp_ok) {
        float* save = a.f.save_act[l];
#pragma unroll
        for (int cc = 0; cc < 32; cc += 16) {
          const int c0 = half * 32 + cc;
          float r[16];
          tc::load_acc16(d_t, trow, c0, r);
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const float4 b4 = *reinterpret_cast<const float4*>(s_bias + l * 64 + c0 + 4 * q4);
            float4 v;
            v.x = lrelu(r[4 * q4 + 0] + b4.x, a.f.slope);
            v.y = lrelu(r[4 * q4 + 1] + b4.y, a.f.slope);
            v.z = lrelu(r[4 * q4 + 2] + b4.z, a.f.slope);
            v.w = lrelu(r[4 * q4 + 3] + b4.w, a.f.slope);
            *reinterpret_cast<float4*>(Hbuf + h_off(i, c0 + 4 * q4)) = v;
            if (save && row_ok) *reinterpret_cast<float4*>(save + ((int64_t)n0 + i) * 64 + c0 + 4 * q4) = v;
          }
        }
      }
      ptx::tc_fence_before_sync();
      ptx::named_bar_sync(bar_id, kTcSlotThreads);  // h complete in shared memory; D may be overwritten

      if (!last && warp_ok) {
        // ---- next layer's A operand: a_r = sum_j norm_rj h_j ; this thread's 32 of the 64 columns ----
        float acc[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) acc[c] = 0.f;
        for (int kk = beg; kk <= end; ++kk) {  // kk == end: the self-loop, last as in the reference
          int j = i;
          float w = wself;
          if (kk < end) {
            j = cl[kk] - n0;
            const bool ok = (unsigned)j < (unsigned)n;
            if (!ok) j = i;
            w = ok ? ds[j] * di : 0.f;
          }
          const float* hr = Hbuf + j * 64 + half * 32;
          const int sw = j & 7;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 h4 = *reinterpret_cast<const float4*>(hr + ((q ^ sw) << 2));
            acc[4 * q + 0] = fmaf(w, h4.x, acc[4 * q + 0]);
            acc[4 * q + 1] = fmaf(w, h4.y, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(w, h4.z, acc[4 * q + 2]);
            acc[4 * q + 3] = fmaf(w, h4.w, acc[4 * q + 3]);
          }
        }
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
      int r = quarter;
      for (; r + 12 < n; r += 16) {
        const float v0 = Hbuf[h_off(r, c)], v1 = Hbuf[h_off(r + 4, c)];
        const float v2 = Hbuf[h_off(r + 8, c)], v3 = Hbuf[h_off(r + 12, c)];
        mx = fmaxf(fmaxf(mx, v0), fmaxf(v1, fmaxf(v2, v3)));
        sm += v0; sm += v1; sm += v2; sm += v3;
      }
      for (; r < n; r += 4) {
        const float v0 = Hbuf[h_off(r, c)];
        mx = fmaxf(mx, v0);
        sm += v0;
      }
      s_pool[quarter * 128 + c] = mx;
      s_pool[quarter * 128 + 64 + c] = sm;
      ptx::named_bar_sync(bar_id, kTcSlotThreads);
      if (gt < 64) {
        mx = fmaxf(fmaxf(s_pool[c], s_pool[128 + c]), fmaxf(s_pool[256 + c], s_pool[384 + c]));
        sm = ((s_pool[64 + c] + s_pool[192 + c]) + s_pool[320 + c]) + s_pool[448 + c];
        a.f.emb[gid * 128 + c] = n > 0 ? mx : 0.f;
        a.f.emb[gid * 128 + 64 + c] = n > 0 ? sm / (float)n : 0.f;
      }
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
  k_graph_desc<<<(unsigned)cdiv(f.B, 256), 256, 0, st>>>(f.graph_ptr, f.rowptr, f.B, w.gdesc);
  HCG_LAUNCH_OK(st, "k_graph_desc");

  TcArgs a{};
  a.f = f;
  a.gdesc = w.gdesc;
  a.overflow_count = w.overflow_count;
  a.overflow_list = w.overflow_list;
  a.x_end = f.x + N * f.F;
  a.rp_end = f.rowptr + N + 1;
  a.cl_end = f.col + E;
  a.ds_end = f.dis + N;
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
