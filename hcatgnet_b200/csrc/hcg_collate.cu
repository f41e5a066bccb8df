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
  int32_t *cnt_dst, *cnt_src, *loop_eid, *sums, *t_sums;
  int2 *tmp, *t_tmp;
  size_t bytes;
};

// one carving routine for both the size query (base == nullptr) and the real call
static CollateWs carve_collate(void* ws, int64_t N, int64_t E, int flags) {
  Carver c(ws, (size_t)-1);
  CollateWs w{};
  int64_t nblk = cdiv(N > 0 ? N : 1, kScanTile) + 1;
  w.status = c.take<int>(64);
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
  (void)B;
  return carve_collate(nullptr, N, E, flags).bytes;
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

  CollateWs w = carve_collate(ws, N, E, flags);
  note_region(st);
  const int64_t* src = edge_index;
  const int64_t* dst = edge_index + E;

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

extern "C" int hcg_collate_status(const void* ws, void* stream, int32_t* graph_flags) {
  int status = 0;
  if (graph_flags) *graph_flags = 0;
  HCG_REQUIRE(ws != nullptr, HCG_ERR_INVALID_ARG, "NULL workspace");
  HCG_CUDA_OK(cudaMemcpyAsync(&status, ws, sizeof(int), cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
  HCG_CUDA_OK(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  if (status & kStatusBadIndex) return set_err(HCG_ERR_BAD_INDEX, "edge_index holds a node index outside [0, N)");
  if (status & kStatusBadBatch) return set_err(HCG_ERR_BAD_BATCH, "batch vector is not sorted or holds an id outside [0, B)");
  if (graph_flags && (status & kStatusCrossEdge)) *graph_flags |= HCG_GRAPH_CROSS_EDGES;
  return HCG_OK;
}
