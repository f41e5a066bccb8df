// Device-resident pieces of the training step around forward/backward (reference
// utils/utils_model.py:60-66: loss = sqrt(MSELoss(out, y.unsqueeze(1))); loss.backward(); optimizer.step()
// with the Adam of reference model/networks.py:36-44):
//
//   hcg_rmse_seed  g_out = 2 (out - y), tail = [ sum (out - y)^2 , count ]      (one block, fixed order)
//   hcg_adam_step  scale = 1 / (2 count rmse), rmse = sqrt(sse / count)  -> the gradient of sqrt(MSE);
//                  Adam update of one flat parameter slab from one flat gradient slab
//
// sqrt(MSE) is not additive over data-parallel shards, (sse, count) are: every rank back-propagates
// 2 (out - y), the flat gradient travels with the tail through ONE all-reduce, and the rescale happens
// inside the optimizer kernel.  With one rank the same code gives the reference's single-device step.
// Also here: the fused backward of the default two-layer readout head.
#include <mutex>

#include "hcg_common.cuh"
#include "hcg_fused.cuh"

namespace hcg {

constexpr int kSeedThreads = 1024;

// y: [B] (broadcast over the C outputs, as MSELoss does for y.unsqueeze(1) against out[B,C])
__global__ void __launch_bounds__(kSeedThreads)
k_rmse_seed(const float* __restrict__ out, const float* __restrict__ y, int64_t B, int C, float* __restrict__ g_out,
            float* __restrict__ tail) {
  __shared__ float s_part[kSeedThreads];
  const int tid = threadIdx.x;
  const int64_t total = B * C;
  float s = 0.f;
  for (int64_t i = tid; i < total; i += kSeedThreads) {
    const float d = out[i] - y[i / C];
    g_out[i] = 2.f * d;
    s = fmaf(d, d, s);
  }
  s_part[tid] = s;
  __syncthreads();
  for (int w = kSeedThreads / 2; w > 0; w >>= 1) {
    if (tid < w) s_part[tid] += s_part[tid + w];
    __syncthreads();
  }
  if (tid == 0) {
    tail[0] = s_part[0];
    tail[1] = (float)total;
  }
}

// torch.optim.Adam (amsgrad = False, weight_decay = 0, maximize = False), same operation order:
//   m = m + (g - m)(1 - b1);  v = v b2 + (1 - b2) g g;  p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
__global__ void __launch_bounds__(256)
k_adam(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
       const float* __restrict__ tail, float one_minus_b1, float b2, float one_minus_b2, float step_size,
       float bc2_sqrt, float eps, float* __restrict__ loss_out) {
  float scale = 1.f;
  if (tail) {
    const float sse = tail[0], cnt = tail[1];
    const float rmse = sqrtf(sse / cnt);
    scale = 1.f / (2.f * cnt * rmse);
    if (loss_out && blockIdx.x == 0 && threadIdx.x == 0) *loss_out = rmse;
  }
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float gi = g[i] * scale;
  const float mi = m[i] + (gi - m[i]) * one_minus_b1;
  const float vi = v[i] * b2 + one_minus_b2 * (gi * gi);
  m[i] = mi;
  v[i] = vi;
  const float denom = sqrtf(vi) / bc2_sqrt + eps;   // torch: (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
  p[i] = p[i] - step_size * (mi / denom);
}

}  // namespace hcg

using namespace hcg;

extern "C" int hcg_rmse_seed(const float* out, const float* y, int64_t B, int C, float* g_out, float* tail, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(out && y && g_out && tail, HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(B >= 1 && C >= 1, HCG_ERR_INVALID_ARG, "bad sizes B=%lld C=%d", (long long)B, C);
  note_region(st);
  k_rmse_seed<<<1, kSeedThreads, 0, st>>>(out, y, B, C, g_out, tail);
  HCG_LAUNCH_OK(st, "k_rmse_seed");
  return HCG_OK;
}

extern "C" int hcg_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n,
                             const float* tail, double lr, double beta1, double beta2, double eps, int64_t step,
                             float* loss_out, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HCG_REQUIRE(param && grad && exp_avg && exp_avg_sq, HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(n >= 0 && step >= 1, HCG_ERR_INVALID_ARG, "bad n=%lld / step=%lld", (long long)n, (long long)step);
  if (n == 0) return HCG_OK;
  // bias corrections in double on the host, exactly as torch's Python code computes them
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  const float step_size = (float)(lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  note_region(st);
  k_adam<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(param, grad, exp_avg, exp_avg_sq, n, tail, (float)(1.0 - beta1), (float)beta2,
                                                 (float)(1.0 - beta2), step_size, bc2_sqrt, (float)eps, loss_out);
  HCG_LAUNCH_OK(st, "k_adam");
  return HCG_OK;
}

// ---- one whole optimisation step in ONE library call ----------------------------------------------------------
// The reference's regime is tiny: 668 reactions in batches of 40 (reference scripts_experiments/train_GNN.py:73-111:
// ~14 steps per epoch, ~100 epochs, 90 runs).  There a step is bound by the HOST: six C-ABI calls, ~25 tensor
// allocations and their Python glue cost more than the ~14 kernels they launch.  hcg_train_step chains
// collate -> forward -> loss seed -> backward -> Adam inside the library over one caller-owned workspace, so the
// host pays for one call.  Single-rank only (with several ranks the all-reduce sits between backward and Adam:
// use the separate entry points).
namespace hcg {

struct StepWs {
  char* collate;
  int32_t *rowptr, *col, *eid, *graph_ptr, *t_rowptr, *t_col, *t_eid;
  float* dis;
  char* fwd;
  float* acts[HCG_MAX_CONV];
  float* hid[HCG_MAX_READOUT];
  float *out, *emb, *g_out;
  char* bwd;
  char* coo;
  float *head_gemb, *head_partial;
  size_t collate_bytes, fwd_bytes, bwd_bytes, coo_bytes, bytes;
};

static StepWs carve_step(void* ws, int64_t N, int64_t E, int64_t B, const HcgParams* p) {
  Carver c(ws, (size_t)-1);
  StepWs w{};
  w.collate_bytes = hcg_collate_workspace_bytes(N, E, B, HCG_COLLATE_TRANSPOSE | HCG_COLLATE_GROUPED);
  w.fwd_bytes = hcg_forward_workspace_bytes(N, E, B, p);
  w.bwd_bytes = hcg_backward_workspace_bytes(N, E, B, p);
  w.collate = c.take<char>(w.collate_bytes);
  w.rowptr = c.take<int32_t>(N + 1); w.col = c.take<int32_t>(E + 1); w.eid = c.take<int32_t>(E + 1);
  w.graph_ptr = c.take<int32_t>(B + 1);
  w.t_rowptr = c.take<int32_t>(N + 1); w.t_col = c.take<int32_t>(E + 1); w.t_eid = c.take<int32_t>(E + 1);
  w.dis = c.take<float>(N + 1);
  w.fwd = c.take<char>(w.fwd_bytes);
  for (int l = 0; l < p->L; ++l) w.acts[l] = c.take<float>((size_t)N * p->H + 4);
  int d = 2 * p->H;
  for (int k = 0; k < p->R - 1; ++k) { d /= 2; w.hid[k] = c.take<float>((size_t)B * d + 4); }
  w.out = c.take<float>((size_t)B * p->C + 4);
  w.emb = c.take<float>((size_t)B * 2 * p->H + 4);
  w.g_out = c.take<float>((size_t)B * p->C + 4);
  w.bwd = c.take<char>(w.bwd_bytes);
  w.coo_bytes = hcg_forward_coo_workspace_bytes(B, p);
  w.coo = c.take<char>(w.coo_bytes);
  w.head_gemb = c.take<float>((size_t)B * 2 * p->H + 4);
  w.head_partial = c.take<float>(readout2_train_partial_floats(B, p->C > 0 ? p->C : 1));
  w.bytes = c.off;
  return w;
}

// A second stream (and two events) per device for the part of the step that the forward does not depend on.
struct SideStream {
  cudaStream_t st = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
static int side_stream(SideStream** out) {
  constexpr int kMaxDevices = 64;
  static SideStream side[kMaxDevices];
  static std::mutex mu;
  int dev = 0;
  HCG_CUDA_OK(cudaGetDevice(&dev));
  HCG_REQUIRE(dev >= 0 && dev < kMaxDevices, HCG_ERR_UNSUPPORTED, "device ordinal %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  SideStream& s = side[dev];
  if (s.st == nullptr) {
    HCG_CUDA_OK(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    HCG_CUDA_OK(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
    HCG_CUDA_OK(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
  }
  *out = &s;
  return HCG_OK;
}

}  // namespace hcg

extern "C" size_t hcg_train_step_workspace_bytes(int64_t N, int64_t E, int64_t B, const HcgParams* p) {
  if (!p || p->L < 1 || p->L > HCG_MAX_CONV || p->R < 1 || p->R > HCG_MAX_READOUT) return 0;
  return carve_step(nullptr, N, E, B, p).bytes;
}

extern "C" int hcg_train_step(const float* x, const int64_t* edge_index, int64_t E, const int64_t* node_ptr,
                              const int64_t* edge_ptr, const float* y, int64_t N, int64_t B, int32_t max_graph_nodes,
                              int32_t max_graph_edges, int32_t max_in_degree, float fill, const HcgParams* p, const HcgGrads* grads,
                              float* param_slab, float* grad_slab, float* exp_avg, float* exp_avg_sq, int64_t n,
                              double lr, double beta1, double beta2, double eps, int64_t step, float* loss_out,
                              void* ws, size_t ws_bytes, void* stream) {
  HCG_REQUIRE(x && edge_index && node_ptr && edge_ptr && y && p && grads && grad_slab && ws, HCG_ERR_INVALID_ARG, "NULL pointer");
  HCG_REQUIRE(step <= 0 || (param_slab && exp_avg && exp_avg_sq), HCG_ERR_INVALID_ARG, "NULL optimizer state");
  HCG_REQUIRE(N > 0 && E > 0 && B > 0 && n > 0, HCG_ERR_INVALID_ARG, "hcg_train_step needs N, E, B, n > 0");
  HCG_REQUIRE(p->L >= 1 && p->L <= HCG_MAX_CONV && p->R >= 1 && p->R <= HCG_MAX_READOUT, HCG_ERR_UNSUPPORTED, "bad layer counts");
  StepWs w = carve_step(ws, N, E, B, p);
  HCG_REQUIRE(ws_bytes >= w.bytes, HCG_ERR_WORKSPACE, "train-step workspace too small: %zu < %zu", ws_bytes, w.bytes);
  // When the caller's (host-verified) size facts allow the edge-list forward, the forward does not need the collate:
  // the collate (whose transposed CSR only the backward reads) leaves the critical path (see below).
  const bool coo_fwd = hcg_forward_coo_supported(p) && max_graph_nodes > 0 && max_graph_nodes <= 128 && max_graph_edges >= 0 &&
                       max_graph_edges <= 512 && max_in_degree >= 0 && max_in_degree <= 16 &&
                       hcg_config_get(HCG_CFG_FORWARD_PATH) == 0;
  cudaStream_t main_st = static_cast<cudaStream_t>(stream);
  HcgGraph g{};
  g.N = N; g.E = E; g.B = B;
  g.rowptr = w.rowptr; g.col = w.col; g.eid = w.eid; g.dis = w.dis; g.graph_ptr = w.graph_ptr;
  g.t_rowptr = w.t_rowptr; g.t_col = w.t_col; g.t_eid = w.t_eid;
  g.fill = fill; g.max_graph_nodes = max_graph_nodes; g.max_graph_edges = max_graph_edges;
  HcgSaved saved{};
  for (int l = 0; l < p->L; ++l) saved.acts[l] = w.acts[l];
  for (int k = 0; k < p->R - 1; ++k) saved.hid[k] = w.hid[k];
  int rc;
  bool fused_head = false;   // forward head + loss seed + backward head in ONE kernel (default readout)
  if (coo_fwd) {
    // forward first, on the caller's stream; then the collate on the second stream, beside the loss seed and the readout
    // part of the backward (small kernels that leave most of the GPU idle); hcg_gcn_backward joins it before its conv part
    SideStream* side = nullptr;
    if (int rc0 = side_stream(&side)) return rc0;
    fused_head = readout2_train_supported(p);
    rc = gcn_forward_coo_saved(x, edge_index, node_ptr, edge_ptr, N, E, B, fill, p, fused_head ? nullptr : w.out, w.emb, &saved, w.coo,
                               w.coo_bytes, stream);
    if (rc) return rc;
    HCG_CUDA_OK(cudaEventRecord(side->fork, main_st));     // (orders the side stream behind the previous step's backward, too)
    HCG_CUDA_OK(cudaStreamWaitEvent(side->st, side->fork, 0));
    rc = hcg_collate_offsets(edge_index, E, node_ptr, edge_ptr, N, B, fill, HCG_COLLATE_TRANSPOSE, max_graph_nodes, max_graph_edges,
                             w.rowptr, w.col, w.eid, w.dis, w.graph_ptr, w.t_rowptr, w.t_col, w.t_eid, w.collate, w.collate_bytes,
                             side->st);
    HCG_CUDA_OK(cudaEventRecord(side->join, side->st));
    if (rc) {
      cudaStreamWaitEvent(main_st, side->join, 0);          // nothing of this call may outlive it out of stream order
      return rc;
    }
    set_backward_prereq(side->join, side->st, node_ptr, edge_ptr);
  } else {
    rc = hcg_collate_offsets(edge_index, E, node_ptr, edge_ptr, N, B, fill, HCG_COLLATE_TRANSPOSE, max_graph_nodes, max_graph_edges,
                             w.rowptr, w.col, w.eid, w.dis, w.graph_ptr, w.t_rowptr, w.t_col, w.t_eid, w.collate, w.collate_bytes, stream);
    if (rc) return rc;
    rc = hcg_gcn_forward(&g, x, p, w.out, w.emb, &saved, w.fwd, w.fwd_bytes, stream);
    if (rc) return rc;
  }
  // (error paths below: drop the pending join and let both streams drain, so that nothing of this call is still
  //  running out of stream order when it returns)
  auto bail = [&](int code) {
    if (coo_fwd) { set_backward_prereq(nullptr); cudaDeviceSynchronize(); }
    return code;
  };
  float* tail = grad_slab + n;
  if (fused_head) {
    rc = launch_readout2_train(w.emb, p, y, B, w.head_gemb, w.head_partial, grads, tail, main_st);
    if (rc) return bail(rc);
    set_backward_keep_readout_grads(true);
    rc = hcg_gcn_backward(&g, x, p, &saved, w.emb, nullptr, w.head_gemb, grads, nullptr, w.bwd, w.bwd_bytes, stream);
    set_backward_keep_readout_grads(false);
  } else {
    rc = hcg_rmse_seed(w.out, y, B, p->C, w.g_out, tail, stream);
    if (rc) return bail(rc);
    rc = hcg_gcn_backward(&g, x, p, &saved, w.emb, w.g_out, nullptr, grads, nullptr, w.bwd, w.bwd_bytes, stream);
  }
  if (rc) return bail(rc);
  if (step <= 0) return HCG_OK;   // gradients only: data-parallel ranks all-reduce grad_slab, then call hcg_adam_step
  return hcg_adam_step(param_slab, grad_slab, exp_avg, exp_avg_sq, n, tail, lr, beta1, beta2, eps, step, loss_out, stream);
}
