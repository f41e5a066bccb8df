// Tile work list shared by the tensor-core graph-resident kernels (forward: hcg_tc.cu, backward: hcg_tc_bwd.cu):
// a tile is 128 node rows holding one or two whole graphs; k_tile_desc (hcg_tc.cu) builds one dense list per slot.
#pragma once

#include "hcg_common.cuh"
#include "hcg_fused.cuh"

namespace hcg {

constexpr int kTcRows = 128;          // node rows per tile
constexpr int kTcEcap = 512;          // CSR slots per tile (4 per node row: molecules)
constexpr int kTcHPitch = 68;         // floats per activation row: 8 consecutive rows cover all banks for 128-bit access
constexpr int kTcWinMax = 8;          // graphs per pairing window (at most)
constexpr int kStatusCrossEdgeTc = 4;
// COO variant of the forward kernel: refusal flags (HCG_COO_* of include/hcg.h)
constexpr int kTcCooDeg = 16;         // neighbour records per node row (in-degree cap)
constexpr int kCooTooLarge = 1;       // a graph has no nodes, more than kTcRows nodes or more than kTcEcap edges
constexpr int kCooBadEdge = 2;        // an edge leaves its graph's node range, or is a self-loop
constexpr int kCooDegree = 4;         // a node has more than kTcCooDeg incoming edges
constexpr int kCooBadOffsets = 8;     // node / edge offsets outside [0, N] / [0, E]
constexpr int kTileManual = 1;        // TcTile.flags: a 16-byte granular copy window would leave an input array

struct __align__(16) TcTile {   // 64 bytes: one tile of the work list
  int gA, n0A, nA, e0A;         // first graph: id, first node, nodes, first CSR slot
  int gB, n0B, nB, e0B;         // second graph (gB < 0: none)
  int eA, eB, flags, pad0;
  int pad1[4];
};

struct __align__(16) TcMeta {   // a tile as the slot's threads see it
  int gA, n0A, nA, e0A;
  int gB, n0B, nB, e0B;
  int eA, eB, flags, pad0;
  int off[8];                   // [4 gb + q]: word offset, inside stage array q (0 x, 1 rowptr, 2 col, 3 dis),
                                // of element 0 of graph gb's slice
};

// The 16-byte granular window [lo, lo + bytes) the bulk copy of `count` 4-byte elements at `p` reads.
__host__ __device__ __forceinline__ void copy_window(uintptr_t p, int count, uintptr_t* lo, uint32_t* bytes, int* lead) {
  *lo = p & ~uintptr_t(15);
  *lead = (int)((p - *lo) >> 2);
  *bytes = (uint32_t)(((p - *lo) + (uintptr_t)count * 4 + 15) & ~uintptr_t(15));
}

struct TcArgs {
  FusedFwdArgs f;
  const TcTile* tiles;      // [nslots][cap] dense per-slot work lists
  const int* counts;        // [nslots]
  int cap;
  int64_t N, E;
  // edge-list variants (k_tc_fwd<.., COO = true>, k_tc_fwd_ens): the raw edge list instead of the CSR of f.rowptr / f.col / f.dis
  const int64_t* ei;        // [2, E]: row 0 sources, row 1 targets, grouped by graph
  int* kflags;              // [nslots]: per-slot refusal flags (plain stores, no memset needed)
  HcgPackedStore pk;        // packed-image variant (k_tc_fwd<.., 2>): the image the graphs are read from
};

struct TcPlan {     // launch geometry of the work list for a batch of B graphs
  int grid, nslots, W, nwin, cap;
};
// ns = tile pipelines per CTA; at most max_grid CTAs (the SM count)
TcPlan tc_plan(int64_t B, int ns);
// element count of the TcTile array a batch of B graphs can need for any ns in 1..3
size_t tc_tiles_capacity(int64_t B);
// builds the per-slot tile lists + overflow list (rowptr / col: the CSR the kernel will copy slices of)
int launch_tile_desc(const int32_t* graph_ptr, const int32_t* rowptr, const int32_t* col, const float* dis, const float* x, int F,
                     int64_t N, int64_t E, int64_t B, const TcPlan& p, TcTile* tiles, int* counts, int* overflow_count,
                     int* overflow_list, cudaStream_t st, int* overflow_flag = nullptr,   // flag[g] = 1: graph g is on the list
                     const int64_t* node_ptr64 = nullptr, const int64_t* edge_ptr64 = nullptr);   // offsets from the batch's own vectors

// workspace of the edge-list variants: work list + per-slot refusal flags
struct TcCooWs {
  TcTile* tiles;
  int* counts;
  int* dflags;     // refusal flags of k_tile_desc, one per slot
  int* kflags;     // refusal flags of the tile kernel, one per slot
  size_t bytes;
};
TcCooWs carve_tc_coo(void* ws, int64_t B);
// work list from the batch's own int64 offsets (two slots per CTA); a graph that does not fit a tile is refused
// (flag) and, when emb_zero_cols > 0, its 128-float row of `emb` zeroed
int launch_tile_desc_coo(const float* x, int F, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N,
                         int64_t E, int64_t B, const TcPlan& p, const TcCooWs& w, float* emb, int emb_zero_cols, cudaStream_t st);

#ifdef __CUDACC__
// Layer-0 gather of one thread: acc[c] = sum_j w_j x[j][c] over its NC columns (NC < 0: `ncol` at run time), rows
// walked as (tile row, norm) records; every lane takes `maxdeg` steps (a lane past its own degree idles:
// fewer shared-memory wavefronts than adding 0 x its own row), then the self-loop, last, as in the reference.
// `xcol`: the thread's first column of tile row 0.
// NBS: distance, in records, between consecutive records of a row (1: CSR slots; kTcRows: COO planes).
template <int NC, int XW, int NBS>
__device__ __forceinline__ void gather_x(float (&acc)[XW], const float* xcol, int F, const int2* nbp, int deg, int maxdeg,
                                         int iself, float wself, int ncol = 0) {
  if (NC == 0) return;
  auto row_fma = [&](int r, float w) {
    const float* xr = xcol + r * F;
#pragma unroll
    for (int c = 0; c < XW; ++c)
      if (NC < 0 ? c < ncol : c < NC) acc[c] = fmaf(w, xr[c], acc[c]);
  };
  int2 nb = 0 < deg ? nbp[0] : make_int2(iself, 0);
  for (int t = 0; t < maxdeg; ++t) {
    const int2 cur = nb;
    nb = t + 1 < deg ? nbp[(t + 1) * NBS] : make_int2(iself, 0);
    if (t < deg) row_fma(cur.x, __int_as_float(cur.y));
  }
  row_fma(iself, wself);
}

// compare-exchange of two (edge number << 8 | tile row) keys
__device__ __forceinline__ void key_cx(int& p, int& q) {
  const int lo = min(p, q), hi = max(p, q);
  p = lo; q = hi;
}
#endif  // __CUDACC__

}  // namespace hcg
