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
#include "hcg_common.cuh"

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
