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
  const uint32_t d_t = tbase, a_hi = tbase + 64, a_lo = tbase + 128;
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
    uint32_t r[16];
    ptx::tmem_ld16(d_t + trow + c0, r);
    ptx::tmem_wait_ld();
#pragma unroll
    for (int q = 0; q < 16; ++q) D[tid * tc::kN + c0 + q] = __uint_as_float(r[q]);
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
