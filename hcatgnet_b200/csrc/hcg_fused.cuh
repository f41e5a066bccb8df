// Declarations shared between hcg_net.cu (dispatch) and hcg_fused.cu (graph-resident kernels).
#pragma once

#include "hcg_common.cuh"

namespace hcg {

constexpr int kFusedH = 64;        // hidden width the graph-resident kernels are specialised for
constexpr int kFusedThreads = 256; // 8 warps per CTA
constexpr int kFusedWarps = kFusedThreads / 32;
constexpr int kTileRows = 16;      // node rows per warp tile
constexpr int kTileStride = 68;    // floats per tile row (64 + 4: adjacent rows land 4 banks apart)
constexpr int kMaxFusedF = 64;     // widest input feature vector kept in shared memory

// Everything the graph-resident forward needs, passed by value (fits the 4 KB parameter space).
struct FusedFwdArgs {
  const int32_t* rowptr;
  const int32_t* col;
  const float* dis;
  const int32_t* graph_ptr;
  const float* x;
  const float* conv_w[HCG_MAX_CONV];
  const float* conv_b[HCG_MAX_CONV];
  float* save_act[HCG_MAX_CONV];  // optional: post-activation output of every conv layer, [N, H]
  float* emb;                     // [B, 2H]
  float* scratch_a;               // [N, H] global scratch for graphs that exceed the smem capacity
  float* scratch_b;               // [N, H]
  int* status;                    // device status word (may be NULL)
  const int* list;                // optional: only these graph ids are processed ...
  const int* list_count;          // ... (device-resident count), e.g. the tensor-core kernel's overflow list
  int64_t B;
  int F, Fp, L;
  int ncap, ecap;                 // per-graph node / edge capacity of the shared-memory path
  float slope, fill;
};

// bytes of dynamic shared memory the kernel needs for a given capacity
size_t fused_fwd_smem_bytes(int F, int L, int ncap, int ecap);
// launches the kernel; grid sized from the SM count and the achievable residency
// (max_ctas > 0 caps the grid: one CTA is enough to find an overflow list empty)
int launch_fused_fwd(const FusedFwdArgs& a, size_t smem, cudaStream_t st, int64_t max_ctas = 0);

// tensor-core graph-resident forward (hcg_tc.cu): H == 64, F <= 32
size_t tc_fwd_smem_bytes(int F, int L);
size_t tc_fwd_workspace_bytes(int64_t B);
int tc_fwd_row_cap();    // tile capacity: node rows ...
int tc_fwd_edge_cap();   // ... and CSR slots
// `a.scratch_*`, `a.ncap/ecap` describe the FMA kernel that picks up the overflow list; ws holds
// the per-graph descriptors and that list
// all_fit: the caller's hints say that every graph fits a tile (the overflow pass then runs as a single CTA)
int launch_tc_fwd(const FusedFwdArgs& a, int64_t N, int64_t E, void* ws, cudaStream_t st, bool all_fit = false);

// COO variant: the kernel builds each tile's neighbour lists from the raw edge list (edge_index int64 [2, E], every
// graph's edges contiguous; node_ptr / edge_ptr: int64 [B + 1] offsets).  No fallback kernel: input it cannot take
// (a graph above the tile capacity, an in-degree above tc_fwd_coo_degree_cap(), a self-loop, an edge that leaves its
// graph) is reported through tc_fwd_coo_flags and the embeddings of the call are then unspecified (finite memory
// accesses only).  a.rowptr / col / dis / graph_ptr / scratch are not used.
size_t tc_fwd_coo_smem_bytes(int F, int L);
size_t tc_fwd_coo_workspace_bytes(int64_t B);
int tc_fwd_coo_degree_cap();
int launch_tc_fwd_coo(const FusedFwdArgs& a, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N,
                      int64_t E, void* ws, cudaStream_t st);
int tc_fwd_coo_flags(const void* ws, int64_t B, cudaStream_t st, int* flags);

// hcg_gcn_forward_coo with the activations saved for the backward pass (hcg_net.cu; used by hcg_train_step)
int gcn_forward_coo_saved(const float* x, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N,
                          int64_t E, int64_t B, float fill, const HcgParams* p, float* out, float* emb, const HcgSaved* saved,
                          void* ws, size_t ws_bytes, void* stream);

// the next hcg_gcn_backward of the calling thread makes its stream wait for `ev` after the readout part, before the
// first kernel that reads the graph arrays (rowptr / col / dis / transposed CSR)
// (`side`: the stream `ev` was recorded on; the tensor-core backward builds its work list there too and re-records `ev`)
// (`node_ptr` / `edge_ptr`: the batch's int64 offset vectors; with them the tensor-core backward builds its work list on its
//  own stream before it waits for `ev` - the list then needs nothing the other stream produces)
void set_backward_prereq(cudaEvent_t ev, cudaStream_t side = nullptr, const int64_t* node_ptr = nullptr, const int64_t* edge_ptr = nullptr);

// The head of a training step in one kernel + reduce (hcg_net.cu): forward of the default readout, loss seed
// (tail = [sum (out - y)^2, B C]) and its backward: g_emb [B, 2H] and the readout gradients into `grads`.
// partial: readout2_train_partial_floats(B, C) floats of scratch.
bool readout2_train_supported(const HcgParams* p);
size_t readout2_train_partial_floats(int64_t B, int C);
int launch_readout2_train(const float* emb, const HcgParams* p, const float* y, int64_t B, float* g_emb, float* partial,
                          const HcgGrads* grads, float* tail, cudaStream_t st);
void set_backward_keep_readout_grads(bool keep);

// packed-image variant: graphs [first, first + a.B) of a packed image resident in device memory (a.x is not used);
// same workspace and flags as the COO variant.  tc_fwd_packed_fits: the image's per-node bytes fit the kernel's stage.
bool tc_fwd_packed_fits(const HcgPackedStore& pk);
int launch_tc_fwd_packed(const FusedFwdArgs& a, const HcgPackedStore& pk, int64_t first, void* ws, cudaStream_t st);

// one-pass ensemble variant (hcg_tc_ens.cu): M models (stacked conv weights: w0 [M,64,F], b0 [M,64], w1 [M,64,64],
// b1 [M,64]) over one batch; emb [M,B,128]; wimg: scratch of tc_fwd_ens_image_floats(M) floats; ws as the COO variant's
// (tc_fwd_coo_workspace_bytes, flags by tc_fwd_coo_flags).  Two conv layers, H = 64, F <= 32.
size_t tc_fwd_ens_smem_bytes(int F);
size_t tc_fwd_ens_image_floats(int M);
int launch_tc_fwd_ens(const float* x, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr, int64_t N, int64_t E,
                      int64_t B, int M, int F, float slope, float fill, const float* w0, const float* b0, const float* w1,
                      const float* b1, float* wimg, float* emb, void* ws, cudaStream_t st);

// ---- graph-resident backward (hcg_fused_bwd.cu): H == 64, two conv layers, F <= 32 ---------------
struct FusedBwdArgs {
  const int32_t* t_rowptr;        // transposed (src-sorted) CSR
  const int32_t* t_col;
  const float* dis;
  const int32_t* graph_ptr;
  const float* x;                 // [N, F]
  const float* h1;                // [N, 64] saved output of conv layer 0
  const float* h2;                // [N, 64] saved output of conv layer 1
  const float* emb;               // [B, 128] forward embedding (max || mean)
  const float* gemb;              // [B, 128] gradient flowing into the embedding
  const float* w2;                // conv_w[1], [64, 64] (out, in)
  float* scratch;                 // 3 planes of [(N + 16 B + 16), 64] for graphs above the smem capacity
  int64_t scratch_stride;         // floats between the planes (set by launch_fused_bwd)
  float* partial;                 // [grid][64*64 + 64*F + 128] per-CTA partial parameter gradients
  const int* only;                // optional per-graph flags: only graphs with only[g] != 0 are processed (the tensor-core
  const int* only_count;          // kernel's overflow graphs; a fixed graph -> CTA assignment keeps the sums reproducible) ...
                                  // ... and their number (device-resident): 0 = nothing to do
  int64_t B;
  int F;
  float slope, fill;
};
int fused_bwd_grid(int64_t B);
size_t fused_bwd_partial_elems(int64_t B, int F);
size_t fused_bwd_scratch_elems(int64_t N, int64_t B);
// writes g_w2 [64,64], g_w1 [64,F], g_b2 [64], g_b1 [64]
// max_ctas > 0 caps the grid; the reduce kernel sums `extra_rows` rows starting at reduce_base (the tensor-core kernel's
// partials, reduce_base + extra_rows * row == a.partial) followed by this kernel's own rows
int launch_fused_bwd(FusedBwdArgs a, int64_t N, float* g_w2, float* g_w1, float* g_b2, float* g_b1, cudaStream_t st,
                     int max_ctas = 0, const float* reduce_base = nullptr, int extra_rows = 0);

// ---- tensor-core graph-resident backward (hcg_tc_bwd.cu): same contract, tiles of one or two graphs ----------------
int tc_bwd_grid(int64_t B);
size_t tc_bwd_workspace_bytes(int64_t B);
int launch_tc_bwd(FusedBwdArgs a, int64_t N, int64_t E, void* ws, bool all_fit, float* g_w2, float* g_w1, float* g_b2, float* g_b1,
                  cudaStream_t st, cudaStream_t desc_st = nullptr, cudaEvent_t desc_ev = nullptr, const int64_t* node_ptr64 = nullptr,
                  const int64_t* edge_ptr64 = nullptr);

}  // namespace hcg
