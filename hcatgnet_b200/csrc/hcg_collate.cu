// Batch collate: COO edge_index (int64) + batch vector -> dst-sorted CSR, graph offsets,
// symmetric-normalisation factors.  Replaces the index work PyG redoes inside every GCNConv
// call (gcn_norm / add_remaining_self_loops / scatter degree; call sites reference
// model/gcn.py:58,62) and the segment bookkeeping of global_{max,mean}_pool (model/gcn.py:65-66).
//
// Integer contract (bit-exact vs oracle/ref_collate.py): self-loop edges dropped, kept edges
// ordered by (dst, original edge id).  Histogram and slot claiming use INTEGER atomics (their
// results are order-independent once each row is re-sorted by edge id); no float atomics.
#include "hcg_common.cuh"

namespace hcg {

constexpr int kStatusBadIndex = 1;
constexpr int kStatusBadBatch = 2;
constexpr int kStatusCrossEdge = 4;

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048 elements per block

// ------------------------------------------------------------------------------------------
// 1. in-degree (and out-degree) histogram over kept edges
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_count(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, const int64_t* __restrict__ batch,
        int64_t E, int64_t N, int32_t* __restrict__ cnt_dst, int32_t* __restrict__ cnt_src, int32_t* __restrict__ loop_eid,
        int* __restrict__ status) {
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
    int64_t s = src[e], d = dst[e];
    if (s < 0 || s >= N || d < 0 || d >= N) {
      atomicOr(status, kStatusBadIndex);
      continue;
    }
    if (s == d) {
      if (loop_eid) atomicMax(&loop_eid[s], (int32_t)e);  // "last one wins" like index assignment
      continue;
    }
    if (batch && batch[s] != batch[d]) atomicOr(status, kStatusCrossEdge);
    atomicAdd(&cnt_dst[d], 1);
    if (cnt_src) atomicAdd(&cnt_src[s], 1);
  }
}

// ------------------------------------------------------------------------------------------
// 2. exclusive scan (three phases; tile-local scan, scan of tile sums, add back)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int warp_incl_scan(int v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) >= o) v += t;
  }
  return v;
}

// block-wide exclusive scan of one value per thread (kScanThreads threads); returns total in *total
__device__ __forceinline__ int block_excl_scan(int v, int* total) {
  __shared__ int warp_sums[kScanThreads / 32];
  __shared__ int block_total;
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int incl = warp_incl_scan(v);
  if (lane == 31) warp_sums[w] = incl;
  __syncthreads();
  if (w == 0) {
    int s = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
    int si = warp_incl_scan(s);
    if (lane < kScanThreads / 32) warp_sums[lane] = si - s;
    if (lane == kScanThreads / 32 - 1) block_total = si;
  }
  __syncthreads();
  int r = incl - v + warp_sums[w];
  *total = block_total;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kScanThreads)
k_scan_tile(const int32_t* __restrict__ in, int64_t n, int32_t* __restrict__ out, int32_t* __restrict__ sums) {
  int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int v[kScanItems];
  int local = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? in[base + i] : 0;
    local += v[i];
  }
  int total;
  int excl = block_excl_scan(local, &total);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = excl;
    excl += v[i];
  }
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// single block: exclusive scan of the tile sums, in place, sequential over chunks with carry
__global__ void __launch_bounds__(kScanThreads)
k_scan_sums(int32_t* __restrict__ sums, int64_t nblk, int32_t* __restrict__ grand_total) {
  int carry = 0;
  for (int64_t c = 0; c < nblk; c += kScanThreads) {
    int64_t i = c + threadIdx.x;
    int v = i < nblk ? sums[i] : 0;
    int total;
    int excl = block_excl_scan(v, &total);
    if (i < nblk) sums[i] = excl + carry;
    carry += total;
  }
  if (threadIdx.x == 0) *grand_total = carry;
}

__global__ void __launch_bounds__(256)
k_scan_add(int32_t* __restrict__ out, int64_t n, const int32_t* __restrict__ sums,
           const int32_t* __restrict__ grand_total) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] += sums[i / kScanTile];
  if (i == n) out[n] = *grand_total;
}

// ------------------------------------------------------------------------------------------
// 3. claim a slot per kept edge (order inside a row fixed up by k_rowsort)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_fill(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t E, int64_t N,
       const int32_t* __restrict__ rowptr, int32_t* __restrict__ cnt_dst, int2* __restrict__ tmp,
       const int32_t* __restrict__ t_rowptr, int32_t* __restrict__ cnt_src, int2* __restrict__ t_tmp) {
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
    int64_t s = src[e], d = dst[e];
    if (s < 0 || s >= N || d < 0 || d >= N || s == d) continue;
    int pos = atomicSub(&cnt_dst[d], 1) - 1;  // counts return to zero: no cursor array needed
    tmp[rowptr[d] + pos] = make_int2((int)e, (int)s);
    if (t_tmp) {
      int tpos = atomicSub(&cnt_src[s], 1) - 1;
      t_tmp[t_rowptr[s] + tpos] = make_int2((int)e, (int)d);
    }
  }
}

// ------------------------------------------------------------------------------------------
// 4. restore original edge order inside every row; emit col / eid / ew
// ------------------------------------------------------------------------------------------
__device__ void heap_sort_by_eid(int2* a, int n) {
  auto sift = [&](int start, int end) {
    int root = start;
    while (2 * root + 1 <= end) {
      int child = 2 * root + 1, sw = root;
      if (a[sw].x < a[child].x) sw = child;
      if (child + 1 <= end && a[sw].x < a[child + 1].x) sw = child + 1;
      if (sw == root) return;
      int2 t = a[root]; a[root] = a[sw]; a[sw] = t;
      root = sw;
    }
  };
  for (int s = (n - 2) / 2; s >= 0; --s) sift(s, n - 1);
  for (int end = n - 1; end > 0; --end) {
    int2 t = a[0]; a[0] = a[end]; a[end] = t;
    sift(0, end - 1);
  }
}

constexpr int kSmallRow = 8;

__global__ void __launch_bounds__(256)
k_rowsort(const int32_t* __restrict__ rowptr, int64_t N, int2* __restrict__ tmp,
          int32_t* __restrict__ col, int32_t* __restrict__ eid,
          const float* __restrict__ edge_weight, float* __restrict__ ew_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  int beg = rowptr[i], end = rowptr[i + 1];
  int d = end - beg;
  if (d <= 0) return;
  if (d <= kSmallRow) {
    int2 r[kSmallRow];
#pragma unroll
    for (int k = 0; k < kSmallRow; ++k) r[k] = k < d ? tmp[beg + k] : make_int2(0x7fffffff, 0);
    // fixed odd-even transposition network: no dynamic indexing, stays in registers
#pragma unroll
    for (int pass = 0; pass < kSmallRow; ++pass) {
#pragma unroll
      for (int k = pass & 1; k + 1 < kSmallRow; k += 2) {
        if (r[k].x > r[k + 1].x) { int2 t = r[k]; r[k] = r[k + 1]; r[k + 1] = t; }
      }
    }
#pragma unroll
    for (int k = 0; k < kSmallRow; ++k) {
      if (k < d) {
        eid[beg + k] = r[k].x;
        col[beg + k] = r[k].y;
        if (ew_out) ew_out[beg + k] = edge_weight[r[k].x];
      }
    }
  } else {
    heap_sort_by_eid(tmp + beg, d);
    for (int k = beg; k < end; ++k) {
      int2 v = tmp[k];
      eid[k] = v.x;
      col[k] = v.y;
      if (ew_out) ew_out[k] = edge_weight[v.x];
    }
  }
}

// ------------------------------------------------------------------------------------------
// 5. deg^-1/2 (IEEE sqrt + divide, to track torch's deg.pow(-0.5) on the CPU reference)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float inv_sqrt_ieee(float deg) {
  float r = __fdiv_rn(1.0f, __fsqrt_rn(deg));
  return isinf(r) ? 0.f : r;
}

__global__ void __launch_bounds__(256)
k_dis(const int32_t* __restrict__ rowptr, int64_t N, float fill, float* __restrict__ dis,
      int weighted, float fill_w, const float* __restrict__ ew_csr, const int32_t* __restrict__ loop_eid,
      const float* __restrict__ edge_weight, float* __restrict__ selfw0, float* __restrict__ dis0) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  int beg = rowptr[i], end = rowptr[i + 1];
  dis[i] = inv_sqrt_ieee((float)(end - beg) + fill);
  if (weighted) {
    float sw = fill_w;
    int le = loop_eid[i];
    if (le >= 0) sw = edge_weight[le];
    selfw0[i] = sw;
    float deg = 0.f;
    for (int k = beg; k < end; ++k) deg += ew_csr[k];  // edge order, self-loop last (as index_add)
    deg += sw;
    dis0[i] = inv_sqrt_ieee(deg);
  }
}

// ------------------------------------------------------------------------------------------
// 6. graph offsets from the sorted batch vector
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_graph_ptr(const int64_t* __restrict__ batch, int64_t N, int64_t B, int32_t* __restrict__ graph_ptr,
            int* __restrict__ status) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (batch == nullptr) {  // single graph
    if (i == 0) { graph_ptr[0] = 0; for (int64_t g = 1; g <= B; ++g) graph_ptr[g] = (int32_t)N; }
    return;
  }
  if (i > N) return;
  if (i == N) {  // tail: graphs after the last node's graph are empty
    int64_t last = N > 0 ? batch[N - 1] : -1;
    if (last >= B) last = B - 1;
    if (last < -1) last = -1;
    for (int64_t g = last + 1; g <= B; ++g) graph_ptr[g] = (int32_t)N;
    return;
  }
  int64_t b = batch[i];
  int64_t prev = i > 0 ? batch[i - 1] : -1;
  if (b < 0 || b >= B || b < prev) {
    atomicOr(status, kStatusBadBatch);
    return;
  }
  if (prev < -1) prev = -1;
  for (int64_t g = prev + 1; g <= b; ++g) graph_ptr[g] = (int32_t)i;
}

// eptr[g] = first edge whose dst belongs to graph g or a later one, eptr[B] = E.  A warp per graph runs a
// 32-ary search: the lanes probe 32 evenly spaced positions of the current range at once, a ballot finds the
// first probe at or past the graph, and the range shrinks 32-fold per step (3 dependent loads inside the
// +-2048 window around the proportional guess node * E / N, 5 over a whole 4 M-edge list) instead of one load
// per halving.  Edge lists grouped by graph make "dst[e] >= first node of g" monotone in e; on other inputs
// the result is some position in [0, E] and the per-edge checks of the build kernels refuse the batch.
__global__ void __launch_bounds__(256)
k_edge_ptr(const int64_t* __restrict__ dst, int64_t E, int64_t N, int64_t B, const int32_t* __restrict__ graph_ptr,
           int32_t* __restrict__ eptr) {
  const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (g > B) return;  // whole warps leave together
  if (g == B) { if (lane == 0) eptr[B] = (int32_t)E; return; }
  const int64_t node = graph_ptr[g];
  if (node <= 0) { if (lane == 0) eptr[g] = 0; return; }
  int64_t lo = 0, hi = E;
  {
    constexpr int64_t kWin = 2048;
    const int64_t guess = (int64_t)((double)node * (double)E / (double)N);
    const int64_t a = guess - kWin, b = guess + kWin;
    // lane 0 tests the lower fence, lane 1 the upper one
    bool ok = false;
    if (lane == 0) ok = a > 0 && a < E && dst[a - 1] < node;
    if (lane == 1) ok = b > 0 && b < E && dst[b] >= node;
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (m & 1u) lo = a;
    if (m & 2u) hi = b;
    if (lo > hi) { lo = 0; hi = E; }
  }
  while (hi > lo) {  // answer in [lo, hi]; dst[e] >= node is monotone on grouped input
    const int64_t step = (hi - lo + 31) >> 5;
    const int64_t pos = lo + (int64_t)lane * step;
    const bool ge = pos >= hi ? true : dst[pos] >= node;
    const unsigned m = __ballot_sync(0xffffffffu, ge);
    const int f = m ? __ffs(m) - 1 : 32;  // first probe at or past the graph
    if (f == 0) { hi = lo; break; }
    const int64_t nlo = lo + (int64_t)(f - 1) * step + 1;
    int64_t nhi = f < 32 ? lo + (int64_t)f * step : hi;
    if (nhi > hi) nhi = hi;
    lo = nlo < nhi ? nlo : nhi;
    hi = nhi;
  }
  if (lane == 0) eptr[g] = (int32_t)lo;
}

// Offsets handed in by the caller (PyG's Batch carries them: `ptr` and `_slice_dict['edge_index']`): graph_ptr / eptr
// are their int32 copies, clamped into range (a wrong offset is then caught by the per-edge checks of the build
// kernels).  Without edge offsets the warp searches for its graph's slice exactly as k_edge_ptr does.
__global__ void __launch_bounds__(256)
k_offsets_from_ptr(const int64_t* __restrict__ node_ptr, const int64_t* __restrict__ edge_ptr, int64_t E, int64_t N, int64_t B,
                   int32_t* __restrict__ graph_ptr, int32_t* __restrict__ eptr, int* __restrict__ status) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g == 0) { status[0] = 0; status[1] = 0; }   // (nothing else writes the status words in this kernel)
  if (g > B) return;
  int64_t n = node_ptr[g];
  n = n < 0 ? 0 : (n > N ? N : n);
  graph_ptr[g] = (int32_t)n;
  int64_t e = edge_ptr[g];
  e = e < 0 ? 0 : (e > E ? E : e);
  eptr[g] = (int32_t)e;
}

// ==========================================================================================
// Graph-local fast path (HCG_COLLATE_GROUPED).  PyG's collate concatenates the edge lists graph
// by graph, so the edges of graph g are one contiguous slice of edge_index: find the slices
// (k_edge_ptr: a search on dst per graph), then build each graph's CSR entirely in shared memory
// (k_collate_graph_warp / k_collate_graph, which also check every edge against its slice) and write
// it out coalesced.  Without dropped self-loops the CSR slots of graph g are exactly its edge slice,
// so no batch-wide scan is needed.  Every precondition is checked on the device.
// ==========================================================================================
constexpr int kStatusNotGrouped = 8;
constexpr int kGrpNodes = 1024;  // per-graph capacity of the shared-memory CSR build
constexpr int kGrpEdges = 2048;
constexpr int kGrpThreads = 128;

// Per-edge check of the graph-local path, made by the thread that places the edge into its dst row: both
// endpoints must be nodes of the graph whose slice [eptr[g], eptr[g+1]) holds the edge.  The slices
// partition [0, E) (a non-monotone eptr is refused by the kernels below), so every edge is checked once.
__device__ __forceinline__ void grp_check_edge(int64_t s, int64_t d, int n0, int n, int64_t N, int* status) {
  if (s < 0 || s >= N || d < 0 || d >= N) { atomicOr(status, kStatusBadIndex | kStatusNotGrouped); return; }
  if (s == d) { atomicOr(status, kStatusNotGrouped); return; }  // dropped self-loops shift the slots
  if ((uint64_t)(d - n0) >= (uint64_t)n) { atomicOr(status, kStatusNotGrouped); return; }  // edge outside its graph's slice
  if ((uint64_t)(s - n0) >= (uint64_t)n) atomicOr(status, kStatusCrossEdge | kStatusNotGrouped);
}

// block-wide exclusive scan of v[0..n) in shared memory (n <= kGrpNodes), kGrpThreads threads;
// out[0..n] receives the offsets, out[n] the total.  `warp_tot` is scratch of kGrpThreads/32 ints.
__device__ __forceinline__ void grp_scan(const int* v, int n, int* out, int* warp_tot) {
  constexpr int PER = kGrpNodes / kGrpThreads;  // 8 consecutive items per thread
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  int local[PER], sum = 0;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int idx = t * PER + q;
    local[q] = idx < n ? v[idx] : 0;
    sum += local[q];
  }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) warp_tot[w] = incl;
  __syncthreads();
  int base = 0;
  for (int q = 0; q < w; ++q) base += warp_tot[q];
  int run = base + incl - sum;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int idx = t * PER + q;
    if (idx < n) out[idx] = run;
    run += local[q];
  }
  if (t == kGrpThreads - 1) out[n] = run;  // the last thread's running sum is the total
  __syncthreads();
}

// One CSR (by_dst = rows are dst, values src) or its transpose (rows src, values dst) of one graph,
// built in shared memory; slots of a row end up in original edge order.
__device__ __forceinline__ void grp_build(const int64_t* __restrict__ rowkey, const int64_t* __restrict__ val,
                                          int e_lo, int e_cnt, int n0, int n, int* cnt, int* rp, int* slot_eid,
                                          int* slot_val, int* warp_tot, int32_t* __restrict__ rowptr_out,
                                          int32_t* __restrict__ col_out, int32_t* __restrict__ eid_out,
                                          float* __restrict__ dis_out, float fill, int64_t N, int* check_status) {
  const int t = threadIdx.x;
  for (int i = t; i < n; i += kGrpThreads) cnt[i] = 0;
  __syncthreads();
  for (int k = t; k < e_cnt; k += kGrpThreads) {
    const int64_t key = rowkey[e_lo + k];
    if ((uint64_t)(key - n0) < (uint64_t)n) atomicAdd(&cnt[(int)(key - n0)], 1);
  }
  __syncthreads();
  grp_scan(cnt, n, rp, warp_tot);
  for (int i = t; i < n; i += kGrpThreads) {
    rowptr_out[n0 + i] = e_lo + rp[i];
    if (dis_out) {
      const float r = __fdiv_rn(1.0f, __fsqrt_rn((float)cnt[i] + fill));
      dis_out[n0 + i] = isinf(r) ? 0.f : r;
    }
  }
  __syncthreads();
  for (int i = t; i < n; i += kGrpThreads) cnt[i] = rp[i];  // cursors
  __syncthreads();
  for (int k = t; k < e_cnt; k += kGrpThreads) {
    const int64_t key = rowkey[e_lo + k], v = val[e_lo + k];
    if (check_status) grp_check_edge(v, key, n0, n, N, check_status);  // by-dst pass: key = dst, v = src
    if ((uint64_t)(key - n0) >= (uint64_t)n) continue;
    const int pos = atomicAdd(&cnt[(int)(key - n0)], 1);
    slot_eid[pos] = e_lo + k;
    slot_val[pos] = (int)v;
  }
  __syncthreads();
  for (int i = t; i < n; i += kGrpThreads) {  // insertion sort by edge id (rows are a handful of slots)
    const int b = rp[i], e = rp[i + 1];
    for (int a = b + 1; a < e; ++a) {
      const int ke = slot_eid[a], kv = slot_val[a];
      int c = a - 1;
      while (c >= b && slot_eid[c] > ke) { slot_eid[c + 1] = slot_eid[c]; slot_val[c + 1] = slot_val[c]; --c; }
      slot_eid[c + 1] = ke; slot_val[c + 1] = kv;
    }
  }
  __syncthreads();
  const int kept = rp[n];
  for (int k = t; k < e_cnt; k += kGrpThreads) {  // kept == e_cnt unless the batch was flagged
    eid_out[e_lo + k] = k < kept ? slot_eid[k] : 0;
    col_out[e_lo + k] = k < kept ? slot_val[k] : 0;
  }
  __syncthreads();
}

// ---- warp-per-graph variant for molecule-sized graphs (<= kWgNodes nodes, <= kWgEdges edges): no
//      block-wide barriers, eight graphs in flight per CTA ----------------------------------------
constexpr int kWgNodes = 256, kWgEdges = 512, kWgWarps = 4;

// The graph's edges are staged ONCE in registers (lane l holds edges l, l + 32, ...; sixteen 64-bit loads in flight
// at a time instead of one per loop trip) as words (dst - n0) << 16 | (src - n0) with a bit per edge for "key inside the
// graph"; both CSR directions are then built from the registers.
constexpr int kWgPer = kWgEdges / 32;   // staged edges per lane

__device__ __forceinline__ void wg_build(const unsigned (&pk)[kWgPer], unsigned okmask, bool by_src, int e_lo, int e_cnt, int n0,
                                         int n, int* cnt, int* rp, unsigned* slot, int32_t* __restrict__ rowptr_out,
                                         int32_t* __restrict__ col_out, int32_t* __restrict__ eid_out,
                                         float* __restrict__ dis_out, float fill) {
  const int lane = threadIdx.x & 31;
  constexpr int PER = kWgNodes / 32;
  for (int i = lane; i < n; i += 32) cnt[i] = 0;
  __syncwarp();
#pragma unroll
  for (int q = 0; q < kWgPer; ++q) {
    const int key = by_src ? (int)(pk[q] & 0xffffu) : (int)(pk[q] >> 16);
    if (lane + 32 * q < e_cnt && (okmask >> q & 1u)) atomicAdd(&cnt[key], 1);
  }
  __syncwarp();
  int local[PER], sum = 0;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int idx = lane * PER + q;
    local[q] = idx < n ? cnt[idx] : 0;
    sum += local[q];
  }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  int run = incl - sum;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int idx = lane * PER + q;
    if (idx < n) rp[idx] = run;
    run += local[q];
  }
  if (lane == 31) rp[n] = run;
  __syncwarp();
  for (int i = lane; i < n; i += 32) {
    const int b = rp[i];
    rowptr_out[n0 + i] = e_lo + b;
    if (dis_out) {
      const float r = __fdiv_rn(1.0f, __fsqrt_rn((float)cnt[i] + fill));
      dis_out[n0 + i] = isinf(r) ? 0.f : r;
    }
    cnt[i] = b;  // cursor of the row from here on
  }
  __syncwarp();
  // A slot is one word: (edge number inside the slice) << 16 | (value node - n0).  Both fit 16 bits for an
  // accepted graph (<= kWgEdges edges, <= kWgNodes nodes); sorting the words sorts a row by edge order.
#pragma unroll
  for (int q = 0; q < kWgPer; ++q) {
    const int k = lane + 32 * q;
    if (k >= e_cnt || !(okmask >> q & 1u)) continue;
    const int key = by_src ? (int)(pk[q] & 0xffffu) : (int)(pk[q] >> 16);
    const unsigned v = by_src ? pk[q] >> 16 : pk[q] & 0xffffu;
    const int pos = atomicAdd(&cnt[key], 1);
    slot[pos] = ((unsigned)k << 16) | v;
  }
  __syncwarp();
  for (int i = lane; i < n; i += 32) {  // insertion sort by edge number inside each row
    const int b = rp[i], e = rp[i + 1];
    for (int a = b + 1; a < e; ++a) {
      const unsigned ks = slot[a];
      int c = a - 1;
      while (c >= b && slot[c] > ks) { slot[c + 1] = slot[c]; --c; }
      slot[c + 1] = ks;
    }
  }
  __syncwarp();
  const int kept = rp[n];
  for (int k = lane; k < e_cnt; k += 32) {  // kept == e_cnt unless the batch was flagged
    const unsigned sv = k < kept ? slot[k] : 0u;
    const int rel = (int)(sv & 0xffffu);
    eid_out[e_lo + k] = k < kept ? e_lo + (int)(sv >> 16) : 0;
    col_out[e_lo + k] = k < kept ? n0 + (rel < n ? rel : 0) : 0;  // a flagged batch stays memory-safe
  }
  __syncwarp();
}

__global__ void __launch_bounds__(kWgWarps * 32, 8)
k_collate_graph_warp(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t E, int64_t N, int64_t B,
                     const int32_t* __restrict__ graph_ptr, const int32_t* __restrict__ eptr, float fill,
                     int32_t* __restrict__ rowptr, int32_t* __restrict__ col, int32_t* __restrict__ eid,
                     float* __restrict__ dis, int32_t* __restrict__ t_rowptr, int32_t* __restrict__ t_col,
                     int32_t* __restrict__ t_eid, int* __restrict__ status) {
  int* const big_count = status + 1;
  __shared__ int cnt[kWgWarps][kWgNodes], rp[kWgWarps][kWgNodes + 1];
  __shared__ unsigned slot[kWgWarps][kWgEdges];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * kWgWarps;
  for (int64_t g = (int64_t)blockIdx.x * kWgWarps + w; g < B; g += stride) {
    const int n0 = graph_ptr[g], n = graph_ptr[g + 1] - n0;
    const int e_lo = eptr[g], e_cnt = eptr[g + 1] - e_lo;
    if (g == B - 1 && lane == 0) {
      rowptr[N] = (int32_t)E;
      if (t_rowptr) t_rowptr[N] = (int32_t)E;
    }
    if (n < 0 && lane == 0) atomicOr(status, kStatusBadBatch);   // offsets that do not ascend
    if (n <= 0) continue;
    if (n > kWgNodes || e_cnt > kWgEdges || e_cnt < 0 || e_lo < 0 || (int64_t)e_lo + e_cnt > E) {
      if (lane == 0) atomicAdd(big_count, 1);  // left to the block-per-graph kernel
      continue;
    }
    unsigned pk[kWgPer], okd = 0u, oks = 0u;
#pragma unroll
    for (int h = 0; h < kWgPer; h += 8) {   // eight edges of the lane = sixteen loads in flight
      int64_t sv[8], dv[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int k = lane + 32 * (h + q);
        sv[q] = k < e_cnt ? __ldg(src + e_lo + k) : 0;
        dv[q] = k < e_cnt ? __ldg(dst + e_lo + k) : 0;
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int k = lane + 32 * (h + q);
        pk[h + q] = 0u;
        if (k < e_cnt) {
          grp_check_edge(sv[q], dv[q], n0, n, N, status);
          const int64_t sr = sv[q] - n0, dr = dv[q] - n0;
          if ((uint64_t)dr < (uint64_t)n) okd |= 1u << (h + q);
          if ((uint64_t)sr < (uint64_t)n) oks |= 1u << (h + q);
          pk[h + q] = ((unsigned)dr & 0xffffu) << 16 | ((unsigned)sr & 0xffffu);
        }
      }
    }
    wg_build(pk, okd, false, e_lo, e_cnt, n0, n, cnt[w], rp[w], slot[w], rowptr, col, eid, dis, fill);
    if (t_rowptr) wg_build(pk, oks, true, e_lo, e_cnt, n0, n, cnt[w], rp[w], slot[w], t_rowptr, t_col, t_eid, nullptr, fill);
  }
}

__global__ void __launch_bounds__(kGrpThreads)
k_collate_graph(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t E, int64_t N, int64_t B,
                const int32_t* __restrict__ graph_ptr, const int32_t* __restrict__ eptr, float fill,
                int32_t* __restrict__ rowptr, int32_t* __restrict__ col, int32_t* __restrict__ eid,
                float* __restrict__ dis, int32_t* __restrict__ t_rowptr, int32_t* __restrict__ t_col,
                int32_t* __restrict__ t_eid, int* __restrict__ status) {
  __shared__ int cnt[kGrpNodes + 1], rp[kGrpNodes + 1], slot_eid[kGrpEdges], slot_val[kGrpEdges];
  __shared__ int warp_tot[kGrpThreads / 32];
  if (status[1] == 0) return;  // k_collate_graph_warp left nothing behind
  for (int64_t g = blockIdx.x; g < B; g += gridDim.x) {
    const int n0 = graph_ptr[g], n = graph_ptr[g + 1] - n0;
    int e_lo = eptr[g], e_cnt = eptr[g + 1] - e_lo;
    if (n <= 0) continue;
    if (n <= kWgNodes && e_cnt <= kWgEdges && e_cnt >= 0 && e_lo >= 0 && (int64_t)e_lo + e_cnt <= E) continue;  // done by a warp
    if (n > kGrpNodes || e_cnt > kGrpEdges || e_cnt < 0 || e_lo < 0 || (int64_t)e_lo + e_cnt > E) {
      // Refused graph: the status word says so and the caller is expected to collate again without the flag.  A
      // caller that never reads the status (validate = false) must still stay in bounds: leave this graph's rows
      // EMPTY (every row starts at the slice's first slot) and point the slice's column entries at the graph's own
      // first node, instead of leaving uninitialised offsets behind.
      if (threadIdx.x == 0) atomicOr(status, kStatusNotGrouped);
      const int64_t lo = e_lo < 0 ? 0 : (e_lo > E ? E : (int64_t)e_lo);
      int64_t hi = (int64_t)e_lo + (e_cnt > 0 ? e_cnt : 0);
      hi = hi < lo ? lo : (hi > E ? E : hi);
      const float d1 = __fdiv_rn(1.0f, __fsqrt_rn(fill));
      for (int i = threadIdx.x; i < n; i += kGrpThreads) {
        rowptr[n0 + i] = (int32_t)lo;
        dis[n0 + i] = isinf(d1) ? 0.f : d1;
        if (t_rowptr) t_rowptr[n0 + i] = (int32_t)lo;
      }
      for (int64_t k = lo + threadIdx.x; k < hi; k += kGrpThreads) {
        col[k] = n0; eid[k] = (int32_t)k;
        if (t_rowptr) { t_col[k] = n0; t_eid[k] = (int32_t)k; }
      }
      continue;
    }
    grp_build(dst, src, e_lo, e_cnt, n0, n, cnt, rp, slot_eid, slot_val, warp_tot, rowptr, col, eid, dis, fill, N, status);
    if (t_rowptr)
      grp_build(src, dst, e_lo, e_cnt, n0, n, cnt, rp, slot_eid, slot_val, warp_tot, t_rowptr, t_col, t_eid, nullptr, fill,
                N, nullptr);
  }
}

static int scan_i32(const int32_t* in, int64_t n, int32_t* out, int32_t* sums, cudaStream_t st) {
  // out has n+1 entries; out[n] = total
  int64_t nblk = cdiv(n > 0 ? n : 1, kScanTile);
  k_scan_tile<<<(unsigned)nblk, kScanThreads, 0, st>>>(in, n, out, sums);
  HCG_LAUNCH_OK(st, "k_scan_tile");
  k_scan_sums<<<1, kScanThreads, 0, st>>>(sums, nblk, sums + nblk);
  HCG_LAUNCH_OK(st, "k_scan_sums");
  k_scan_add<<<(unsigned)cdiv(n + 1, 256), 256, 0, st>>>(out, n, sums, sums + nblk);
  HCG_LAUNCH_OK(st, "k_scan_add");
  return HCG_OK;
}

struct CollateWs {
  int* status;
  int32_t *cnt_dst, *cnt_src, *loop_eid, *sums, *t_sums, *eptr;
  int2 *tmp, *t_tmp;
  size_t bytes;
};

// one carving routine for both the size query (base == nullptr) and the real call
static CollateWs carve_collate(void* ws, int64_t N, int64_t E, int64_t B, int flags) {
  Carver c(ws, (size_t)-1);
  CollateWs w{};
  int64_t nblk = cdiv(N > 0 ? N : 1, kScanTile) + 1;
  w.status = c.take<int>(64);
  w.eptr = c.take<int32_t>(B + 2);   // status and eptr are adjacent: one memset clears both
  w.cnt_dst = c.take<int32_t>(N + 1);   // status, cnt_dst, cnt_src are adjacent: one memset clears them
  w.cnt_src = c.take<int32_t>(N + 1);
  w.loop_eid = c.take<int32_t>(N + 1);
  w.sums = c.take<int32_t>(nblk);
  w.t_sums = c.take<int32_t>(nblk);
  w.tmp = c.take<int2>(E + 1);
  w.t_tmp = (flags & HCG_COLLATE_TRANSPOSE) ? c.take<int2>(E + 1) : nullptr;
  w.bytes = c.off;
  return w;
}

}  // namespace hcg

using namespace hcg;

extern "C" size_t hcg_collate_workspace_bytes(int64_t N, int64_t E, int64_t B, int flags) {
  return carve_collate(nullptr, N, E, B, flags).bytes;
}

extern "C" int hcg_collate(const int64_t* edge_index, int64_t E, const int64_t* batch, int64_t N, int64_t B,
                           float fill, const float* edge_weight, float fill_w, int flags,
                           int32_t* rowptr, int32_t* col, int32_t* eid, float* dis, int32_t* graph_ptr,
                           int32_t* t_rowptr, int32_t* t_col, int32_t* t_eid,
                           float* ew0, float* t_ew0, float* selfw0, float* dis0,
                           void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(N >= 0 && E >= 0 && B >= 0, HCG_ERR_INVALID_ARG, "negative size N=%lld E=%lld B=%lld",
              (long long)N, (long long)E, (long long)B);
  HCG_REQUIRE(N < INT32_MAX && E < INT32_MAX && B < INT32_MAX, HCG_ERR_UNSUPPORTED,
              "sizes must fit int32 (N=%lld E=%lld B=%lld)", (long long)N, (long long)E, (long long)B);
  HCG_REQUIRE(E == 0 || edge_index != nullptr, HCG_ERR_INVALID_ARG, "edge_index is NULL with E=%lld", (long long)E);
  HCG_REQUIRE(rowptr && dis && graph_ptr && ws, HCG_ERR_INVALID_ARG, "NULL output or workspace");
  HCG_REQUIRE(E == 0 || (col && eid), HCG_ERR_INVALID_ARG, "NULL col/eid output");
  HCG_REQUIRE(batch != nullptr || B <= 1, HCG_ERR_INVALID_ARG, "batch is NULL with B=%lld", (long long)B);
  const bool transpose = (flags & HCG_COLLATE_TRANSPOSE) != 0;
  const bool weighted = edge_weight != nullptr;
  if (transpose) HCG_REQUIRE(t_rowptr && (E == 0 || (t_col && t_eid)), HCG_ERR_INVALID_ARG, "NULL transposed outputs");
  if (weighted)
    HCG_REQUIRE((E == 0 || ew0) && selfw0 && dis0 && (!transpose || E == 0 || t_ew0), HCG_ERR_INVALID_ARG,
                "edge_weight given but ew0/selfw0/dis0 outputs missing");
  const size_t need = hcg_collate_workspace_bytes(N, E, B, flags);
  HCG_REQUIRE(ws_bytes >= need, HCG_ERR_WORKSPACE, "collate workspace too small: %zu < %zu", ws_bytes, need);

  CollateWs w = carve_collate(ws, N, E, B, flags);
  note_region(st);
  const int64_t* src = edge_index;
  const int64_t* dst = edge_index + E;

  // ---- graph-local fast path (checked on the device; see HCG_COLLATE_GROUPED in hcg.h) -----------
  if ((flags & HCG_COLLATE_GROUPED) && batch != nullptr && !weighted && B > 0 && N > 0) {
    HCG_CUDA_OK(cudaMemsetAsync(w.status, 0, (size_t)((char*)w.cnt_dst - (char*)w.status), st));
    HCG_CUDA_OK(cudaMemsetAsync(graph_ptr, 0, (size_t)(B + 1) * sizeof(int32_t), st));
    note_other("memset(collate)", st);
    k_graph_ptr<<<(unsigned)cdiv(N + 1, 256), 256, 0, st>>>(batch, N, B, graph_ptr, w.status);
    HCG_LAUNCH_OK(st, "k_graph_ptr");
    if (E > 0) {  // E == 0 leaves the cleared eptr
      k_edge_ptr<<<(unsigned)cdiv((B + 1) * 32, 256), 256, 0, st>>>(dst, E, N, B, graph_ptr, w.eptr);
      HCG_LAUNCH_OK(st, "k_edge_ptr");
    }
    int64_t wb = cdiv(B, kWgWarps);
    if (wb > (int64_t)kNumSMs * 12) wb = (int64_t)kNumSMs * 12;
    k_collate_graph_warp<<<(unsigned)wb, kWgWarps * 32, 0, st>>>(src, dst, E, N, B, graph_ptr, w.eptr, fill, rowptr, col, eid, dis,
                                                                transpose ? t_rowptr : nullptr, t_col, t_eid, w.status);
    HCG_LAUNCH_OK(st, "k_collate_graph_warp");
    int64_t gb = B < (int64_t)kNumSMs * 8 ? B : (int64_t)kNumSMs * 8;
    k_collate_graph<<<(unsigned)gb, kGrpThreads, 0, st>>>(src, dst, E, N, B, graph_ptr, w.eptr, fill, rowptr, col, eid, dis,
                                                        transpose ? t_rowptr : nullptr, t_col, t_eid, w.status);
    HCG_LAUNCH_OK(st, "k_collate_graph");
    return HCG_OK;
  }

  // status + both histograms are adjacent: one memset; loop_eid = -1; graph_ptr = 0 so that a
  // rejected batch vector still leaves memory-safe offsets behind
  size_t zero_bytes = (size_t)((char*)w.loop_eid - (char*)w.status);
  HCG_CUDA_OK(cudaMemsetAsync(w.status, 0, zero_bytes, st));
  if (weighted) HCG_CUDA_OK(cudaMemsetAsync(w.loop_eid, 0xff, (size_t)(N + 1) * sizeof(int32_t), st));
  HCG_CUDA_OK(cudaMemsetAsync(graph_ptr, 0, (size_t)(B + 1) * sizeof(int32_t), st));
  note_other("memset(collate)", st);

  int64_t eb = cdiv(E, 256);
  if (eb > (int64_t)kNumSMs * 16) eb = (int64_t)kNumSMs * 16;
  const unsigned eblocks = (unsigned)eb;
  if (E > 0) {
    k_count<<<eblocks, 256, 0, st>>>(src, dst, batch, E, N, w.cnt_dst, transpose ? w.cnt_src : nullptr,
                                     weighted ? w.loop_eid : nullptr, w.status);
    HCG_LAUNCH_OK(st, "k_count");
  }
  int rc = scan_i32(w.cnt_dst, N, rowptr, w.sums, st);
  if (rc) return rc;
  if (transpose) {
    rc = scan_i32(w.cnt_src, N, t_rowptr, w.t_sums, st);
    if (rc) return rc;
  }
  if (E > 0) {
    k_fill<<<eblocks, 256, 0, st>>>(src, dst, E, N, rowptr, w.cnt_dst, w.tmp,
                                    transpose ? t_rowptr : nullptr, transpose ? w.cnt_src : nullptr,
                                    transpose ? w.t_tmp : nullptr);
    HCG_LAUNCH_OK(st, "k_fill");
  }
  if (N > 0) {
    const unsigned nb = (unsigned)cdiv(N, 256);
    if (E > 0) {
      k_rowsort<<<nb, 256, 0, st>>>(rowptr, N, w.tmp, col, eid, edge_weight, weighted ? ew0 : nullptr);
      HCG_LAUNCH_OK(st, "k_rowsort");
      if (transpose) {
        k_rowsort<<<nb, 256, 0, st>>>(t_rowptr, N, w.t_tmp, t_col, t_eid, edge_weight, weighted ? t_ew0 : nullptr);
        HCG_LAUNCH_OK(st, "k_rowsort(t)");
      }
    }
    k_dis<<<nb, 256, 0, st>>>(rowptr, N, fill, dis, weighted ? 1 : 0, fill_w, ew0, w.loop_eid, edge_weight, selfw0, dis0);
    HCG_LAUNCH_OK(st, "k_dis");
  }
  k_graph_ptr<<<(unsigned)cdiv(N + 1, 256), 256, 0, st>>>(batch, N, B, graph_ptr, w.status);
  HCG_LAUNCH_OK(st, "k_graph_ptr");
  return HCG_OK;
}

extern "C" int hcg_collate_offsets(const int64_t* edge_index, int64_t E, const int64_t* node_ptr, const int64_t* edge_ptr,
                                   int64_t N, int64_t B, float fill, int flags, int32_t max_graph_nodes, int32_t max_graph_edges,
                                   int32_t* rowptr, int32_t* col, int32_t* eid, float* dis, int32_t* graph_ptr,
                                   int32_t* t_rowptr, int32_t* t_col, int32_t* t_eid, void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(N > 0 && E > 0 && B > 0, HCG_ERR_INVALID_ARG, "hcg_collate_offsets needs N, E, B > 0");
  HCG_REQUIRE(N < INT32_MAX && E < INT32_MAX && B < INT32_MAX, HCG_ERR_UNSUPPORTED, "sizes must fit int32");
  HCG_REQUIRE(edge_index && node_ptr && edge_ptr && rowptr && col && eid && dis && graph_ptr && ws, HCG_ERR_INVALID_ARG, "NULL pointer");
  const bool transpose = (flags & HCG_COLLATE_TRANSPOSE) != 0;
  if (transpose) HCG_REQUIRE(t_rowptr && t_col && t_eid, HCG_ERR_INVALID_ARG, "NULL transposed outputs");
  const size_t need = hcg_collate_workspace_bytes(N, E, B, flags | HCG_COLLATE_GROUPED);
  HCG_REQUIRE(ws_bytes >= need, HCG_ERR_WORKSPACE, "collate workspace too small: %zu < %zu", ws_bytes, need);
  CollateWs w = carve_collate(ws, N, E, B, flags);
  note_region(st);
  const int64_t* src = edge_index;
  const int64_t* dst = edge_index + E;
  k_offsets_from_ptr<<<(unsigned)cdiv(B + 1, 256), 256, 0, st>>>(node_ptr, edge_ptr, E, N, B, graph_ptr, w.eptr, w.status);
  HCG_LAUNCH_OK(st, "k_offsets_from_ptr");
  int64_t wb = cdiv(B, kWgWarps);
  if (wb > (int64_t)kNumSMs * 12) wb = (int64_t)kNumSMs * 12;
  k_collate_graph_warp<<<(unsigned)wb, kWgWarps * 32, 0, st>>>(src, dst, E, N, B, graph_ptr, w.eptr, fill, rowptr, col, eid, dis,
                                                              transpose ? t_rowptr : nullptr, t_col, t_eid, w.status);
  HCG_LAUNCH_OK(st, "k_collate_graph_warp");
  // graphs the warp kernel leaves behind (above kWgNodes / kWgEdges): when the caller's hints say there are none, one
  // CTA is enough to find that out (and to serve a wrong hint, slowly)
  const bool none_big = max_graph_nodes > 0 && max_graph_nodes <= kWgNodes && max_graph_edges > 0 && max_graph_edges <= kWgEdges;
  int64_t gb = none_big ? 1 : (B < (int64_t)kNumSMs * 8 ? B : (int64_t)kNumSMs * 8);
  k_collate_graph<<<(unsigned)gb, kGrpThreads, 0, st>>>(src, dst, E, N, B, graph_ptr, w.eptr, fill, rowptr, col, eid, dis,
                                                      transpose ? t_rowptr : nullptr, t_col, t_eid, w.status);
  HCG_LAUNCH_OK(st, "k_collate_graph");
  return HCG_OK;
}

extern "C" int hcg_collate_status(const void* ws, void* stream, int32_t* graph_flags) {
  int status = 0;
  if (graph_flags) *graph_flags = 0;
  HCG_REQUIRE(ws != nullptr, HCG_ERR_INVALID_ARG, "NULL workspace");
  HCG_CUDA_OK(cudaMemcpyAsync(&status, ws, sizeof(int), cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
  HCG_CUDA_OK(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  if (status & kStatusBadIndex) return set_err(HCG_ERR_BAD_INDEX, "edge_index holds a node index outside [0, N)");
  if (status & kStatusBadBatch) return set_err(HCG_ERR_BAD_BATCH, "batch vector is not sorted or holds an id outside [0, B)");
  if (status & kStatusNotGrouped)
    return set_err(HCG_ERR_NOT_GROUPED, "HCG_COLLATE_GROUPED: edges are not grouped by graph, or self-loops / cross-graph edges / "
                                        "a graph above %d nodes or %d edges are present; collate again without the flag", kGrpNodes, kGrpEdges);
  if (graph_flags && (status & kStatusCrossEdge)) *graph_flags |= HCG_GRAPH_CROSS_EDGES;
  return HCG_OK;
}

// ---- batch vector from graph offsets ------------------------------------------------------------------
// PyG's Batch carries both `batch` [N] and `ptr` [B+1]; they hold the same information.  A host batch can
// therefore cross PCIe without its `batch` vector (8 bytes per node, 5 % of the bytes of a reaction graph)
// and have it rebuilt here.
namespace hcg {
__global__ void __launch_bounds__(256)
k_batch_from_ptr(const int64_t* __restrict__ ptr, int64_t B, int64_t N, int64_t* __restrict__ batch) {
  // a warp per graph: lanes stride over the graph's node range
  const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (g >= B) return;
  int64_t lo = ptr[g], hi = ptr[g + 1];
  lo = lo < 0 ? 0 : lo;
  hi = hi > N ? N : hi;
  for (int64_t i = lo + lane; i < hi; i += 32) batch[i] = g;
}
}  // namespace hcg

extern "C" int hcg_batch_from_ptr(const int64_t* ptr, int64_t B, int64_t N, int64_t* batch, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(ptr && (batch || N == 0), HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(B >= 0 && N >= 0, HCG_ERR_INVALID_ARG, "negative sizes");
  if (B == 0 || N == 0) return HCG_OK;
  hcg::k_batch_from_ptr<<<(unsigned)hcg::cdiv(B * 32, 256), 256, 0, st>>>(ptr, B, N, batch);
  HCG_LAUNCH_OK(st, "k_batch_from_ptr");
  return HCG_OK;
}
