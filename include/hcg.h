/*
 * hcg.h — C-ABI of libhcg.so, the B200 (sm_100a) implementation of HCat-GNet's GCN hot path.
 *
 * The reference (EdAguilarB/hcatgnet) has no FFI of its own: its boundary is the Python
 * torch.nn.Module `GCN` (reference model/gcn.py:9-76) whose arithmetic is delegated to
 * torch_geometric.  Each entry point below replaces one group of PyG/ATen calls made on that
 * path; the host-side mirror of the reference interface (hcatgnet_b200/model/gcn.py) binds
 * them through ctypes.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends
 *    in `_host`;
 *  - nothing is allocated inside: the caller supplies outputs and one scratch workspace whose
 *    size comes from the matching `*_workspace_bytes` query;
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*) and returns 0
 *    (HCG_OK) or an error code; `hcg_last_error()` gives the thread-local message;
 *  - input-dependent errors (bad node index, unsorted batch vector) are recorded in a device
 *    status word and surfaced by `hcg_collate_status` (one 4-byte read-back, synchronises);
 *    the kernels themselves stay memory-safe by skipping offending entries;
 *  - all reductions are atomic-free in floating point and therefore run-to-run deterministic.
 */
#ifndef HCG_H_
#define HCG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HCG_MAX_CONV 8
#define HCG_MAX_READOUT 8

enum {
  HCG_OK = 0,
  HCG_ERR_INVALID_ARG = 1,
  HCG_ERR_UNSUPPORTED = 2,
  HCG_ERR_WORKSPACE = 3,
  HCG_ERR_CUDA = 4,
  HCG_ERR_BAD_INDEX = 5,      /* edge_index entry outside [0, N) */
  HCG_ERR_BAD_BATCH = 6,      /* batch vector not sorted / outside [0, B) */
  HCG_ERR_NOT_GROUPED = 7     /* HCG_COLLATE_GROUPED was requested but the batch does not qualify: call again without it */
};

/* HcgGraph.flags */
enum {
  HCG_GRAPH_CROSS_EDGES = 1   /* some edge joins two different graphs (reported by hcg_collate_status):
                                 the batch is not a disjoint union, use the layer-wise kernels */
};

/* flags of hcg_collate */
enum {
  HCG_COLLATE_TRANSPOSE = 1,  /* also build the src-sorted (transposed) CSR for the backward pass */
  HCG_COLLATE_GROUPED = 2     /* graph-local fast path: the caller expects edge_index to list the edges graph by
                                 graph in batch order (what PyG's Batch.from_data_list produces), without self-loop
                                 edges, graphs of at most 1024 nodes / 2048 edges, no edge weights.  Same outputs,
                                 bit for bit.  The expectation is CHECKED on the device: if it does not hold the
                                 status word says so (hcg_collate_status returns HCG_ERR_NOT_GROUPED), the outputs
                                 are unusable but every access stayed in bounds; call again without the flag. */
};

/*
 * Batch topology produced by hcg_collate and consumed by forward/backward.
 * Replaces what PyG recomputes in every GCNConv call: add_remaining_self_loops + scatter
 * degree + deg^-1/2 + per-edge norm (gcn_norm; call sites reference model/gcn.py:58,62), and
 * the implicit `batch`->segment bookkeeping of global_max_pool / global_mean_pool
 * (reference model/gcn.py:65-66).
 */
typedef struct HcgGraph {
  int64_t N;                 /* nodes in the batch */
  int64_t E;                 /* edges as stored in edge_index (self-loops included) */
  int64_t B;                 /* graphs in the batch */
  const int32_t* rowptr;     /* [N+1] dst-sorted CSR over kept (non self-loop) edges */
  const int32_t* col;        /* [E]   src of each CSR slot, stable in original edge order */
  const int32_t* eid;        /* [E]   original edge id of each CSR slot */
  const float* dis;          /* [N]   (in-degree + fill)^-1/2 of the UNWEIGHTED graph, 0 where infinite */
  const int32_t* graph_ptr;  /* [B+1] first node of each graph (PyG Batch.ptr) */
  const int32_t* t_rowptr;   /* [N+1] transposed (src-sorted) CSR, or NULL if not built */
  const int32_t* t_col;      /* [E]   dst of each transposed slot */
  const int32_t* t_eid;      /* [E] */
  float fill;                /* self-loop weight of the unweighted graph (1.0) */
  /* Layer-0 override, all NULL unless an explicit edge_weight was collated.  The reference
   * hands edge_weight to conv1 only (model/gcn.py:127 vs :131), so conv layers >= 1 always
   * use the unweighted normalisation above. */
  const float* ew0;          /* [E] edge weight per CSR slot */
  const float* t_ew0;        /* [E] edge weight per transposed slot */
  const float* selfw0;       /* [N] self-loop weight per node */
  const float* dis0;         /* [N] weighted-degree^-1/2 */
  /* Optional hint: node count of the largest graph of the batch (0 = unknown).  Only used to size
   * the shared-memory capacity of the graph-resident kernels; a wrong hint costs speed, never
   * correctness (graphs above the capacity run on global scratch). */
  int32_t max_graph_nodes;
  int32_t flags;             /* HCG_GRAPH_* bits */
  /* Optional soft edge mask, [E] in ORIGINAL edge order (indexed through eid / t_eid), NULL = none: every
   * message of every conv layer is multiplied by its edge's mask value, added self-loops are not masked and
   * the degree normalisation is unchanged — what PyG's Explainer hooks do to a MessagePassing layer
   * (edge_mask_type='object'; reference scripts_experiments/explain_gnn.py:39-50).  Masked calls run on the
   * layer-wise kernels. */
  const float* edge_mask;
  /* Optional hint like max_graph_nodes: directed edge count of the largest graph of the batch (0 = unknown). */
  int32_t max_graph_edges;
  int32_t reserved0;
} HcgGraph;

/*
 * Network parameters, laid out exactly as the reference state_dict tensors
 * (reference model/gcn.py:18-45): conv weights [H, F_l] (out,in) row-major, conv biases [H],
 * readout weights [d_out, d_in], readout biases [d_out]; d_0 = 2H, d_{k+1} = d_k / 2,
 * last layer -> C.
 */
typedef struct HcgParams {
  int32_t F;                 /* n_node_features */
  int32_t H;                 /* embedding_dim */
  int32_t L;                 /* n_convolutions (>= 1) */
  int32_t R;                 /* readout_layers (>= 1) */
  int32_t C;                 /* n_classes */
  float slope;               /* LeakyReLU negative slope (0.01) */
  const float* conv_w[HCG_MAX_CONV];
  const float* conv_b[HCG_MAX_CONV];
  const float* ro_w[HCG_MAX_READOUT];
  const float* ro_b[HCG_MAX_READOUT];
} HcgParams;

/* Same shape as HcgParams' pointer arrays; receives d(loss)/d(param).  Written, not accumulated. */
typedef struct HcgGrads {
  float* conv_w[HCG_MAX_CONV];
  float* conv_b[HCG_MAX_CONV];
  float* ro_w[HCG_MAX_READOUT];
  float* ro_b[HCG_MAX_READOUT];
  float* edge_mask;          /* [E] d(loss)/d(HcgGraph.edge_mask) in original edge order, or NULL to skip it */
} HcgGrads;

/*
 * Activations kept by the forward pass for the backward pass (what autograd would save).
 * acts: L buffers of [N, H] (post-LeakyReLU output of each conv layer);
 * emb:  [B, 2H] graph embedding (max || mean)  — also a forward output;
 * hid:  R-1 buffers [B, d_{k+1}] (post-LeakyReLU readout hidden layers).
 */
typedef struct HcgSaved {
  float* acts[HCG_MAX_CONV];
  float* hid[HCG_MAX_READOUT];
} HcgSaved;

const char* hcg_last_error(void);
const char* hcg_version(void);

/*
 * Instrumentation (used by bench.py; never on the product's critical path).
 * hcg_launch_count: kernels this library has launched in this process so far.
 * hcg_profile_enable(1) starts recording a CUDA event after every kernel launch on the launch
 * stream, hcg_profile_enable(0) stops; hcg_profile_read synchronises and aggregates device time
 * per kernel name into name_buf (cap rows of 64 bytes), total_ms and counts.  It returns
 * -(rows written) on success (<= 0) or a positive HCG_ERR_* code.
 */
uint64_t hcg_launch_count(void);
int hcg_profile_enable(int on);
int hcg_profile_read(char* name_buf, float* total_ms, int64_t* counts, int cap);

/*
 * Library-wide switches (tests and benchmarks; defaults are what production uses).
 * HCG_CFG_FORWARD_PATH: 0 = auto (graph-resident kernels when H == 64, no edge weights and the batch
 * looks like molecules: the tensor-core kernel for F <= 32, the FMA kernel for F <= 64; layer-wise
 * kernels otherwise), 1 = force layer-wise, 2 = force graph-resident FMA, 3 = force graph-resident
 * tensor-core (HCG_ERR_UNSUPPORTED if not applicable).
 * HCG_CFG_BACKWARD_PATH: 0 = auto (the graph-resident backward kernel when H == 64, L == 2, F <= 32, no
 * edge weights, a disjoint molecule-like batch and no gradient w.r.t. x is wanted; layer-wise kernels
 * otherwise), 1 = force layer-wise, 2 = force graph-resident (HCG_ERR_UNSUPPORTED if not applicable), 3 = force the
 * graph-resident mma.sync kernel for every graph (the default graph-resident path runs the tcgen05 kernel and leaves
 * only graphs above its tile capacity to the mma.sync kernel).
 * HCG_CFG_GEMM_PATH (dense X.W products of the layer-wise kernels): 0 = auto (split-TF32 tcgen05 kernel for
 * output widths 64/128/192/256 on batches of >= 8192 rows, fp32 FMA tiles otherwise), 1 = FMA tiles only,
 * 2 = the tensor-core kernel wherever its shape constraints hold (output width a multiple of 64, <= 256).
 * The weight-gradient products of the layer-wise backward run on mma.sync TF32 tiles (3-term split) unless 1.
 * HCG_CFG_AGG_PATH (neighbour aggregation of the layer-wise kernels): 0 = auto (warp-tile kernels without edge
 * weights / edge mask: whole-row kernel with the pooling fused into the last layer for H >= 128, H % 4 == 0,
 * H <= 512; narrow kernel for even 32 <= H <= 64; column-pass gather + k_pool otherwise), 1 = column-pass
 * gather + k_pool only, 2 = the warp-tile kernels whenever their shape constraints hold.
 * All are CUDA paths; there is no CPU path.  Returns the previous value, or -1 for a bad key.
 */
enum { HCG_CFG_FORWARD_PATH = 0, HCG_CFG_BACKWARD_PATH = 1, HCG_CFG_GEMM_PATH = 2, HCG_CFG_AGG_PATH = 3, HCG_CFG_COUNT_ = 4 };
int hcg_config_set(int key, int value);
int hcg_config_get(int key);

/*
 * Self-test of the tensor-core building block (split-TF32 tcgen05.mma with the A operand in tensor
 * memory): D[128,64] = A[128,K] . W[64,K]^T, 1 <= K <= 64, fp32 in / fp32 out.  Test-suite only.
 */
int hcg_selftest_tc_gemm(const float* A, const float* W, int K, float* D, void* stream);
/* The weight-gradient building block: G[64,64] = X[K,64]^T . Y[K,64], 1 <= K <= 128 (node rows), both operands read from
 * row-per-node shared-memory tiles as MN-major tcgen05 operands, split-TF32.  Test-suite only. */
int hcg_selftest_tc_xty(const float* X, const float* Y, int K, float* G, void* stream);

/* sizeof() of the structs above, for bindings that mirror them (ctypes, cgo, JNI ...) */
size_t hcg_sizeof_graph(void);
size_t hcg_sizeof_params(void);
size_t hcg_sizeof_grads(void);
size_t hcg_sizeof_saved(void);

/* ---- collate (replaces gcn_norm index work; north_star "batch-collate kernel") ------------ */

size_t hcg_collate_workspace_bytes(int64_t N, int64_t E, int64_t B, int flags);

/*
 * edge_index: int64 [2, E] row-major (row 0 = src j, row 1 = dst i), as PyG stores it;
 * batch:      int64 [N], sorted node->graph ids (PyG Batch.batch); may be NULL when B <= 1;
 * fill:       self-loop weight of the unweighted normalisation (1.0 reproduces the reference);
 * edge_weight: float [E] in original edge order or NULL.  With weights, `fill_w` is the weight
 *             of added self-loops (2.0 when improved) and an existing self-loop edge keeps its
 *             own weight (add_remaining_self_loops semantics).
 * Outputs: rowptr[N+1], col[E], eid[E], dis[N], graph_ptr[B+1]; t_rowptr/t_col/t_eid only with
 * HCG_COLLATE_TRANSPOSE; ew0[E], selfw0[N], dis0[N] (and t_ew0[E]) only when edge_weight != NULL.
 * The first 4 bytes of `ws` hold the device status word read by hcg_collate_status.
 */
int hcg_collate(const int64_t* edge_index, int64_t E, const int64_t* batch, int64_t N, int64_t B,
                float fill, const float* edge_weight, float fill_w, int flags,
                int32_t* rowptr, int32_t* col, int32_t* eid, float* dis, int32_t* graph_ptr,
                int32_t* t_rowptr, int32_t* t_col, int32_t* t_eid,
                float* ew0, float* t_ew0, float* selfw0, float* dis0,
                void* ws, size_t ws_bytes, void* stream);

/*
 * The graph-local path of hcg_collate (HCG_COLLATE_GROUPED: same preconditions, same device-side checks, same
 * outputs bit for bit) for callers that hold the graph offsets: node_ptr int64 [B+1] (PyG `Batch.ptr`) and
 * edge_ptr int64 [B+1] (first edge of each graph; PyG keeps it as `Batch._slice_dict['edge_index']`).  Nothing is
 * derived from the N-entry batch vector or searched in the edge list: two kernels instead of five and two memsets.
 * max_graph_nodes / max_graph_edges: sizes of the largest graph if known (0 = unknown); they only decide how many
 * CTAs look for graphs the warp-per-graph kernel cannot take.  flags: HCG_COLLATE_TRANSPOSE or 0.
 */
int hcg_collate_offsets(const int64_t* edge_index, int64_t E, const int64_t* node_ptr, const int64_t* edge_ptr,
                        int64_t N, int64_t B, float fill, int flags, int32_t max_graph_nodes, int32_t max_graph_edges,
                        int32_t* rowptr, int32_t* col, int32_t* eid, float* dis, int32_t* graph_ptr,
                        int32_t* t_rowptr, int32_t* t_col, int32_t* t_eid, void* ws, size_t ws_bytes, void* stream);

/*
 * batch[i] = g for ptr[g] <= i < ptr[g+1]: rebuilds PyG's `Batch.batch` [N] from `Batch.ptr` [B+1] on the
 * device, so that a host batch need not carry the redundant vector across PCIe.
 */
int hcg_batch_from_ptr(const int64_t* ptr, int64_t B, int64_t N, int64_t* batch, void* stream);

/*
 * Packed graph image resident in device memory (layout: hcatgnet_b200/packed.py).  Replaces the reference's
 * one-pickle-per-reaction store + Python collate (reference data/datasets.py:74-78, call_methods.py:41-46).
 * All pointers are device pointers into the image.
 */
typedef struct HcgPackedStore {
  int64_t G, N, E;            /* graphs, nodes, directed edges in the image */
  int32_t F, Fu, Ff;          /* feature columns: total, stored as u8, stored as f32 (Fu + Ff + Fb == F) */
  int32_t edge_kind;          /* bytes per local edge endpoint: 1, 2 or 4 */
  const int32_t* col_slot;    /* [F]  0 .. 65535: column of xu8;  >= 65536: bit (slot - 65536) of the node's xbits row;  < 0: ~column of xf32 */
  const int64_t* node_ptr;    /* [G+1] */
  const int64_t* edge_ptr;    /* [G+1] */
  const float* y;             /* [G] */
  const uint8_t* xu8;         /* [N, Fu] */
  const float* xf32;          /* [N, Ff] */
  const void* edges;          /* [E, 2] (src, dst) local to the graph */
  int32_t Fb, bit_stride;     /* columns whose every value is 0 or 1, stored as one bit each; bytes per node of xbits ((Fb + 7) / 8) */
  const uint8_t* xbits;       /* [N, bit_stride]  bit k of a node: byte k / 8, bit k % 8 */
} HcgPackedStore;

/*
 * Assemble the batch of graphs graph_ids[0..B) (device int64; NULL = first, first+1, ...) exactly as PyG's
 * collate would: x f32 [N_out, F], edge_index i64 [2, E_out] (row 0 at edge_index, row 1 at edge_index +
 * E_out) with node offsets added, batch i64 [N_out] (may be NULL), y f32 [B] (may be NULL).  out_ptr /
 * out_eptr: device int64 [B+1] exclusive prefix sums of the selected graphs' node / edge counts.
 */
int hcg_unpack_batch(const HcgPackedStore* store, const int64_t* graph_ids, int64_t first, int64_t B,
                     const int64_t* out_ptr, const int64_t* out_eptr, float* x, int64_t* edge_index, int64_t E_out,
                     int64_t* batch, float* y, void* stream);

/*
 * Inference straight from the packed image: graphs [first, first + B) of `store`, in order, through the edge-list
 * forward without hcg_unpack_batch - the conv kernel stages each graph's bit / byte / f32 feature slices and its local
 * byte edge pairs (a few bytes per node instead of 4 F + 16 per edge) and expands them in shared memory, so x and
 * edge_index never exist in HBM.  Same bits as hcg_unpack_batch + hcg_gcn_forward_coo.  Workspace:
 * hcg_forward_coo_workspace_bytes(B, p); refusal flags (HCG_COO_*): hcg_forward_coo_status on the same `ws`.
 * hcg_forward_packed_supported: 1 when the network (as hcg_forward_coo_supported) and the image (byte edge endpoints,
 * F == p->F <= 32, at most ~50 packed bytes per node) can take this path; otherwise unpack first.
 * Replaces, for a resident or streamed screening set: DataLoader collate (call_methods.py:41-46) + model(batch)
 * (utils/utils_model.py:83-96).
 */
int hcg_forward_packed_supported(const HcgPackedStore* store, const HcgParams* p);
int hcg_gcn_forward_packed(const HcgPackedStore* store, int64_t first, int64_t B, float fill, const HcgParams* p,
                           float* out, float* emb, void* ws, size_t ws_bytes, void* stream);

/* Synchronises `stream`, reads the status word of the last hcg_collate that used `ws`.  Returns
 * HCG_ERR_BAD_INDEX / HCG_ERR_BAD_BATCH for unusable input; `graph_flags` (may be NULL) receives
 * the HCG_GRAPH_* bits to put into HcgGraph.flags. */
int hcg_collate_status(const void* ws, void* stream, int32_t* graph_flags);

/* ---- forward (replaces GCNConv x L, LeakyReLU, global max/mean pool, readout MLP) ---------- */

size_t hcg_forward_workspace_bytes(int64_t N, int64_t E, int64_t B, const HcgParams* p);

/*
 * x: [N, F] fp32 row-major.  out: [B, C].  emb: [B, 2H] (max first, then mean).
 * saved: NULL for inference; otherwise its buffers are filled for hcg_gcn_backward.
 */
int hcg_gcn_forward(const HcgGraph* g, const float* x, const HcgParams* p,
                    float* out, float* emb, const HcgSaved* saved,
                    void* ws, size_t ws_bytes, void* stream);

/*
 * Inference straight from the batch as torch_geometric holds it (`Batch.x`, `Batch.edge_index`, `Batch.ptr`; the
 * call of the reference this replaces is `model(batch)` under torch.no_grad, utils/utils_model.py:83-96): no
 * hcg_collate, no HcgGraph.  The conv kernel stages each graph's slice of the int64 edge list and builds the
 * normalised neighbour lists in shared memory (add_remaining_self_loops + the symmetric degree norm of
 * GCNConv.forward), so the CSR never exists in HBM.  Results carry the same bits as hcg_collate + hcg_gcn_forward.
 *
 * edge_index: int64 [2, E], the edges of graph g contiguous at [edge_ptr[g], edge_ptr[g+1]) (what
 *             Batch.from_data_list produces);  node_ptr, edge_ptr: int64 [B + 1] device arrays.
 * Preconditions, checked ON THE DEVICE and reported by hcg_forward_coo_status as HCG_COO_* bits: every graph has
 * 1..128 nodes and at most 512 edges, every node at most 16 incoming edges, no self-loops in the input, no edge
 * between two graphs.  When a bit is set the outputs of that call are unspecified (all memory accesses stay in
 * bounds): run hcg_collate + hcg_gcn_forward instead.  hcg_forward_coo_supported: 1 when the network's shape
 * (H = 64, F <= 32, conv layers that fit shared memory) can take this path at all.
 */
enum {
  HCG_COO_TOO_LARGE = 1,    /* a graph without nodes, with more than 128 nodes or more than 512 edges */
  HCG_COO_BAD_EDGE = 2,     /* self-loop, or an endpoint outside the graph that owns the edge slice */
  HCG_COO_DEGREE = 4,       /* a node with more than 16 incoming edges */
  HCG_COO_BAD_OFFSETS = 8   /* node_ptr / edge_ptr entry outside [0, N] / [0, E] */
};
int hcg_forward_coo_supported(const HcgParams* p);
size_t hcg_forward_coo_workspace_bytes(int64_t B, const HcgParams* p);
int hcg_gcn_forward_coo(const float* x, const int64_t* edge_index, const int64_t* node_ptr, const int64_t* edge_ptr,
                        int64_t N, int64_t E, int64_t B, float fill, const HcgParams* p, float* out, float* emb,
                        void* ws, size_t ws_bytes, void* stream);
/* Synchronises `stream` and ORs the HCG_COO_* bits of the last hcg_gcn_forward_coo that used `ws` into *flags. */
int hcg_forward_coo_status(const void* ws, int64_t B, void* stream, int32_t* flags);

/*
 * One-pass ENSEMBLE inference (replaces the loop over trained models of scripts_experiments/predict_test.py:47-70, each
 * iteration a whole predict_network, utils/utils_model.py:82-111; the reference's own ensemble script
 * scripts_experiments/predict.py:11-60 is dead code): M networks of one architecture (the default one: two conv layers,
 * H = 64, readout 128 -> 64 -> C) over one batch in ONE pass.  Per tile of the batch the kernel stages the graphs,
 * builds the normalised neighbour lists and gathers the layer-0 operand A.x once, then streams the models' weights
 * through shared memory; per model only the two tensor-core products, the 64-wide gather and the pooling remain.
 * Weights are the state_dict tensors of the M models STACKED along a new first axis.  Per model the results carry the
 * same bits as hcg_gcn_forward_coo.  Input preconditions and HCG_COO_* flags as hcg_gcn_forward_coo
 * (hcg_forward_coo_status on the same `ws`).
 * out: [M, B, C];  emb: [M, B, 2H].
 */
typedef struct HcgEnsemble {
  int32_t M, F, H, C;
  float slope;
  int32_t reserved0;
  const float* conv_w0;   /* [M, H, F]   conv1.lin.weight            */
  const float* conv_b0;   /* [M, H]      conv1.bias                  */
  const float* conv_w1;   /* [M, H, H]   conv_layers.0.lin.weight    */
  const float* conv_b1;   /* [M, H]      conv_layers.0.bias          */
  const float* ro_w0;     /* [M, H, 2H]  readout.0.0.weight          */
  const float* ro_b0;     /* [M, H]      readout.0.0.bias            */
  const float* ro_w1;     /* [M, C, H]   readout.1.weight            */
  const float* ro_b1;     /* [M, C]      readout.1.bias              */
} HcgEnsemble;
size_t hcg_sizeof_ensemble(void);   /* sizeof(HcgEnsemble), for bindings */
size_t hcg_ensemble_workspace_bytes(int64_t B, int32_t M);
int hcg_gcn_forward_coo_ensemble(const float* x, const int64_t* edge_index, const int64_t* node_ptr,
                                 const int64_t* edge_ptr, int64_t N, int64_t E, int64_t B, float fill,
                                 const HcgEnsemble* ens, float* out, float* emb, void* ws, size_t ws_bytes,
                                 void* stream);

/*
 * hcg_gcn_forward_coo for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as hcg_gcn_forward_coo, but no workspace argument: the scratch (tile work list,
 * refusal flags, hidden rows of a deeper readout) is a file-scope device array of 6 MiB inside the library, held by
 * every GPU that loads it until the process exits.  The call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine (with hcg_profile_enable off, the default).
 * Limits: B <= 16384 graphs per call and a scratch need (hcg_forward_coo_workspace_bytes) of at most 6 MiB - the
 * default network at 16384 graphs needs 5.1 MiB; beyond either the call returns HCG_ERR_WORKSPACE.
 * The array is shared by all calls on a device and keeps the last call's contents (nothing in it needs clearing:
 * each call writes every word it later reads); calls that OVERLAP on the GPU - different streams without an
 * ordering between them - race on it: serialise them or use hcg_gcn_forward_coo with one workspace per stream.
 * flags: NULL (no read-back, fully asynchronous), or a host pointer that receives the HCG_COO_* bits of this call
 * (synchronises `stream`).
 */
int hcg_gcn_forward_coo_static(const float* x, const int64_t* edge_index, const int64_t* node_ptr,
                               const int64_t* edge_ptr, int64_t N, int64_t E, int64_t B, float fill,
                               const HcgParams* p, float* out, float* emb, int32_t* flags, void* stream);

/* ---- backward (replaces autograd over the ops above) -------------------------------------- */

size_t hcg_backward_workspace_bytes(int64_t N, int64_t E, int64_t B, const HcgParams* p);

/*
 * g_out: [B, C] gradient w.r.t. the network output, or NULL (= 0);
 * g_emb: [B, 2H] extra gradient flowing into the graph embedding, or NULL;
 * grads: receives parameter gradients;  g_x: [N, F] gradient w.r.t. x, or NULL to skip it.
 * Needs the transposed CSR in `g` (HCG_COLLATE_TRANSPOSE).
 */
int hcg_gcn_backward(const HcgGraph* g, const float* x, const HcgParams* p, const HcgSaved* saved,
                     const float* emb, const float* g_out, const float* g_emb,
                     const HcgGrads* grads, float* g_x,
                     void* ws, size_t ws_bytes, void* stream);

/* ---- training step around forward/backward (replaces the loss / optimizer lines of the reference's
 *      train loop, utils/utils_model.py:64-66, and torch.optim.Adam of model/networks.py:36-44) ------ */

/*
 * Seeds the backward pass of loss = sqrt(MSELoss(out, y.unsqueeze(1))) in its shard-additive form:
 * g_out[B,C] = 2 (out - y)  and  tail[0] = sum (out - y)^2, tail[1] = B*C.   y: [B] (broadcast over C).
 * The 1 / (2 count rmse) factor of d sqrt(sse/count) is applied by hcg_adam_step after the (optional)
 * all-reduce of gradient slab + tail, so data-parallel ranks take exactly the single-device step on the
 * concatenated batch.
 */
int hcg_rmse_seed(const float* out, const float* y, int64_t B, int C, float* g_out, float* tail, void* stream);

/*
 * One Adam step (torch.optim.Adam semantics: amsgrad off, weight_decay 0) over a flat parameter slab.
 * tail: NULL, or the [sse, count] pair written by hcg_rmse_seed (summed over ranks): the gradient is then
 * scaled by 1 / (2 count sqrt(sse / count)) and the RMSE is written to loss_out (may be NULL).
 * step: 1-based step number (bias correction).  The hyper-parameters arrive as doubles, as torch holds them:
 * 1 - beta and lr / bias_correction1 are formed in double and rounded to fp32 once (1.f - 0.999f is off by 1e-5).
 */
int hcg_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n,
                  const float* tail, double lr, double beta1, double beta2, double eps, int64_t step,
                  float* loss_out, void* stream);

/*
 * One whole optimisation step of the reference's train loop (utils/utils_model.py:60-66) in ONE call: collate
 * (hcg_collate_offsets with the transposed CSR) -> forward -> hcg_rmse_seed -> backward -> hcg_adam_step, over one
 * caller-owned workspace (hcg_train_step_workspace_bytes).  For the reference's own regime (batches of 40 reactions)
 * the step is bound by the host's per-call cost, not by the kernels.  `p` / `grads` point into `param_slab` /
 * `grad_slab` (n floats + the [SSE, count] tail).  Single rank; preconditions of the graph-local collate apply.
 * step = 0: stop after the backward (grad_slab = [gradient of the shard's SSE | SSE | count]; no optimizer state needed):
 * data-parallel ranks all-reduce grad_slab and call hcg_adam_step themselves.
 * max_graph_nodes / max_graph_edges / max_in_degree: host-verified facts about the batch (0 / 0 / -1 = unknown).  When
 * they show that the edge-list forward applies (<= 128 nodes and <= 512 edges per graph, in-degree <= 16), the forward
 * runs straight from the edge list and the collate - whose transposed CSR only the backward reads - runs beside it on
 * a second, library-owned stream that is joined into `stream` before the backward: same results, bit for bit.
 */
size_t hcg_train_step_workspace_bytes(int64_t N, int64_t E, int64_t B, const HcgParams* p);
int hcg_train_step(const float* x, const int64_t* edge_index, int64_t E, const int64_t* node_ptr, const int64_t* edge_ptr,
                   const float* y, int64_t N, int64_t B, int32_t max_graph_nodes, int32_t max_graph_edges,
                   int32_t max_in_degree, float fill,
                   const HcgParams* p, const HcgGrads* grads, float* param_slab, float* grad_slab, float* exp_avg,
                   float* exp_avg_sq, int64_t n, double lr, double beta1, double beta2, double eps, int64_t step,
                   float* loss_out, void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HCG_H_ */
