// Device-side loader of the packed graph image (hcatgnet_b200/packed.py):
// replaces the reference's per-batch Python collate of pickled Data objects (reference call_methods.py:41-46,
// data/datasets.py:74-78) for a store that is resident in device memory.
//
//   hcg_unpack_batch   selected graphs g_b = graph_ids[b] (or first + b)  ->  x f32 [N_out, F],
//                      edge_index i64 [2, E_out] (local endpoints + the graph's node offset in the batch),
//                      batch i64 [N_out], y f32 [B]          — what Batch.from_data_list builds, bit for bit
//
// One block per selected graph.  HBM-bound byte shuffling: reads ~28 B/node + 2 B/edge of packed data, writes
// 4F + 8 B/node and 16 B/edge; the u8 -> f32 conversion is exact (values 0..255).
#include "hcg_common.cuh"

namespace hcg {

constexpr int kUnpackThreads = 128;
constexpr int kUnpackMaxF = 256;
constexpr int kBitSlot0 = 65536;   // col_slot values from here on name a bit of the node's xbits row

template <typename ET>
__global__ void __launch_bounds__(kUnpackThreads)
k_unpack_batch(const HcgPackedStore s, const int64_t* __restrict__ graph_ids, int64_t first,
               const int64_t* __restrict__ out_ptr, const int64_t* __restrict__ out_eptr, float* __restrict__ x,
               int64_t* __restrict__ edge_index, int64_t E_out, int64_t* __restrict__ batch, float* __restrict__ y) {
  __shared__ int32_t s_slot[kUnpackMaxF];
  const int64_t b = blockIdx.x;
  const int64_t g = graph_ids ? graph_ids[b] : first + b;
  const int tid = threadIdx.x;
  const int F = s.F;
  for (int c = tid; c < F; c += kUnpackThreads) s_slot[c] = s.col_slot[c];
  __syncthreads();
  if (g < 0 || g >= s.G) return;  // ids are range-checked by the host wrapper; never write from a bad one
  const int64_t s0 = s.node_ptr[g], n = s.node_ptr[g + 1] - s0, d0 = out_ptr[b];
  const int64_t se0 = s.edge_ptr[g], e = s.edge_ptr[g + 1] - se0, de0 = out_eptr[b];
  // node features: consecutive threads write consecutive floats of the [n, F] block
  // (a single graph is far below 2^31 elements: 32-bit index arithmetic inside the graph)
  const int total = (int)n * F;
  const uint8_t* xu = s.xu8 + s0 * s.Fu;
  const float* xf = s.xf32 + s0 * s.Ff;
  const uint8_t* xb = s.xbits + s0 * s.bit_stride;
  float* xo = x + d0 * F;
  int r = tid / F, c = tid - r * F;                      // element tid, then steps of kUnpackThreads
  const int dr = kUnpackThreads / F, dc = kUnpackThreads - dr * F;
  for (int t = tid; t < total; t += kUnpackThreads) {
    const int slot = s_slot[c];
    float v;
    if (slot >= kBitSlot0) {
      const int k = slot - kBitSlot0;
      v = (float)((xb[r * s.bit_stride + (k >> 3)] >> (k & 7)) & 1);
    } else {
      v = slot >= 0 ? (float)xu[r * s.Fu + slot] : xf[r * s.Ff + (~slot)];
    }
    xo[t] = v;
    r += dr; c += dc;
    if (c >= F) { c -= F; ++r; }
  }
  const ET* ed = static_cast<const ET*>(s.edges) + 2 * se0;
  int64_t* e_src = edge_index + de0;
  int64_t* e_dst = edge_index + E_out + de0;
  for (int k = tid; k < (int)e; k += kUnpackThreads) {
    e_src[k] = d0 + (int64_t)ed[2 * k];
    e_dst[k] = d0 + (int64_t)ed[2 * k + 1];
  }
  if (batch)
    for (int rr = tid; rr < (int)n; rr += kUnpackThreads) batch[d0 + rr] = b;
  if (y && tid == 0) y[b] = s.y[g];
}

}  // namespace hcg

using namespace hcg;

extern "C" int hcg_unpack_batch(const HcgPackedStore* store, const int64_t* graph_ids, int64_t first, int64_t B,
                                const int64_t* out_ptr, const int64_t* out_eptr, float* x, int64_t* edge_index,
                                int64_t E_out, int64_t* batch, float* y, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(store != nullptr, HCG_ERR_INVALID_ARG, "NULL store");
  HCG_REQUIRE(B >= 0 && E_out >= 0 && first >= 0, HCG_ERR_INVALID_ARG, "negative sizes");
  if (B == 0) return HCG_OK;
  HCG_REQUIRE(graph_ids != nullptr || first + B <= store->G, HCG_ERR_INVALID_ARG, "graphs [%lld, %lld) outside the store (G=%lld)",
              (long long)first, (long long)(first + B), (long long)store->G);
  HCG_REQUIRE(store->F >= 1 && store->F <= kUnpackMaxF && store->Fb >= 0 && store->Fu + store->Ff + store->Fb == store->F,
              HCG_ERR_INVALID_ARG, "bad feature split F=%d Fu=%d Ff=%d Fb=%d", store->F, store->Fu, store->Ff, store->Fb);
  HCG_REQUIRE(store->Fb == 0 || (store->xbits && store->bit_stride * 8 >= store->Fb), HCG_ERR_INVALID_ARG,
              "bit columns without a bit section (Fb=%d, bit_stride=%d)", store->Fb, store->bit_stride);
  HCG_REQUIRE(store->node_ptr && store->edge_ptr && store->col_slot && store->y && out_ptr && out_eptr && x,
              HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE((store->Fu == 0 || store->xu8) && (store->Ff == 0 || store->xf32) && (E_out == 0 || (store->edges && edge_index)),
              HCG_ERR_INVALID_ARG, "NULL data section");
  note_region(st);
  switch (store->edge_kind) {
    case 1: k_unpack_batch<uint8_t><<<(unsigned)B, kUnpackThreads, 0, st>>>(*store, graph_ids, first, out_ptr, out_eptr, x, edge_index, E_out, batch, y); break;
    case 2: k_unpack_batch<uint16_t><<<(unsigned)B, kUnpackThreads, 0, st>>>(*store, graph_ids, first, out_ptr, out_eptr, x, edge_index, E_out, batch, y); break;
    case 4: k_unpack_batch<int32_t><<<(unsigned)B, kUnpackThreads, 0, st>>>(*store, graph_ids, first, out_ptr, out_eptr, x, edge_index, E_out, batch, y); break;
    default: return set_err(HCG_ERR_INVALID_ARG, "edge_kind must be 1, 2 or 4 bytes, got %d", store->edge_kind);
  }
  HCG_LAUNCH_OK(st, "k_unpack_batch");
  return HCG_OK;
}
