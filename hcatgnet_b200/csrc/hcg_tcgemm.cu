// Dense layer product of the layer-wise path on the 5th-generation tensor cores:
//
//     C[M, Nc] = A[M, K] . B^T        B(n, k) = W[n][k]  (forward:  T = h W^T, W = [out, in])
//                                     B(n, k) = W[k][n]  (backward: g_in = gT W)
//
// for hidden widths the graph-resident kernels do not cover (H = 128, 192, 256: the reference's
// `embedding_dim` option; BASELINE's stress configuration is H = 256, 4 conv layers), where X.W is a real
// dense contraction: 88 % of the FLOPs and 87 % of the step time on the fp32 FMA pipe.
//
// Same error-compensated scheme as the graph-resident forward (hcg_tc.cuh): fp32 operands are split into
// TF32 hi + lo, D accumulates lo.hi + hi.lo + hi.hi in fp32 (fp32-grade accuracy; plain TF32 misses the
// 1e-5 parity bound by two orders of magnitude).
//
//  * k_w_image turns W into the exact shared-memory image the MMA reads — per 32-wide K chunk a hi and a lo
//    plane of all Nc rows in the K-major / 128-byte-swizzle layout — once per call;
//  * k_tc_gemm: one persistent CTA of 128 threads per SM walks 128-row tiles.  Thread r owns row r: it
//    loads the row's K chunk from global memory (one chunk ahead, in registers), splits it and stores it
//    as lane r of the A operand in tensor memory (tcgen05.st, double-buffered); the W chunk image (up to
//    64 KB) arrives by ONE bulk async copy into one of three shared-memory slots (cp.async.bulk ->
//    mbarrier, issued as soon as the slot's previous MMAs have completed); one thread issues the 12
//    tcgen05.mma (M = 128, N = Nc, K = 8; three split terms x four K steps) of the chunk into the
//    [128 x Nc] fp32 accumulator in tensor memory and commits them to the mbarrier that frees the A buffer
//    and the W slot two chunks later.  Epilogue: tcgen05.ld of the accumulator row, optional LeakyReLU'
//    mask, row store.
// Measured (B200, N = 1.31 M rows, K = Nc = 256): 0.99 ms = 173 TFLOP/s fp32-equivalent, 3 x 173 = 520 TFLOP/s
// on the TF32 pipe (fp32 FMA tiles: 3.96 ms).  Sharing the W stream across 2-CTA clusters (multicast) and a
// deeper A prefetch were both measured without gain (profiles/experiments/README.md): the kernel runs at the
// TF32 rate the tensor pipe sustains for this shape; what is left is the arithmetic of the 3-term split.
#include "hcg_common.cuh"
#include "hcg_tc.cuh"

namespace hcg {

constexpr int kGemmThreads = 128;
constexpr int kGemmKC = 32;                      // K chunk
constexpr int kGemmMaxNc = 256;                  // fp32 accumulator columns in tensor memory (+ 128 for the A operand)

// ---- W -> per-chunk image: [kc][hi | lo][Nc x 32 swizzled] -----------------------------------------------
__global__ void __launch_bounds__(256)
k_w_image(const float* __restrict__ W, int Nc, int K, int transposed, int nkc, float* __restrict__ img) {
  const int Kpad = nkc * kGemmKC;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Nc * Kpad) return;
  const int n = idx / Kpad, k = idx - n * Kpad;
  float v = 0.f;
  if (k < K) v = transposed ? __ldg(W + (int64_t)k * Nc + n) : __ldg(W + (int64_t)n * K + k);
  const uint32_t hi = tc::tf32_hi(v), lo = tc::tf32_lo(v, hi);
  const int kc = k >> 5, kk = k & 31;
  const int off = kc * 2 * Nc * kGemmKC + tc::w_tile_offset(n, kk);  // 8-row groups 1024 bytes apart, all Nc rows in one plane
  img[off] = __uint_as_float(hi);
  img[off + Nc * kGemmKC] = __uint_as_float(lo);
}

// one chunk's MMAs (M = 128, N = Nc, K = 32 in four steps of 8 per split term), accumulated on top of D unless `fresh`
__device__ __forceinline__ void issue_chunk_mma(uint32_t d, uint32_t a_hi, uint32_t a_lo, uint32_t b_hi_smem, uint32_t b_lo_smem,
                                                bool fresh, int Nc) {
  const uint32_t idesc = ptx::make_idesc_tf32(tc::kRows, Nc);
  const uint64_t bh = ptx::make_kmajor_sw128_desc(b_hi_smem), bl = ptx::make_kmajor_sw128_desc(b_lo_smem);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const uint64_t boff = (uint64_t)((s * 32) >> 4);
    ptx::mma_tf32_ts(d, a_lo + s * 8, bh + boff, idesc, (s > 0 || !fresh) ? 1u : 0u);
    ptx::mma_tf32_ts(d, a_hi + s * 8, bl + boff, idesc, 1u);
  }
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const uint64_t boff = (uint64_t)((s * 32) >> 4);
    ptx::mma_tf32_ts(d, a_hi + s * 8, bh + boff, idesc, 1u);
  }
}

// this thread's 32 floats of chunk kc of its row (zero beyond K / beyond M)
__device__ __forceinline__ void load_a_chunk(const float* __restrict__ A, int64_t row, bool valid, int K, int kc, bool vec4,
                                             float (&v)[32]) {
  const int k0 = kc * kGemmKC;
  if (valid && vec4 && k0 + kGemmKC <= K) {
    const float4* p = reinterpret_cast<const float4*>(A + row * K + k0);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 t = __ldg(p + q);
      v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int q = 0; q < 32; ++q) v[q] = (valid && k0 + q < K) ? __ldg(A + row * K + k0 + q) : 0.f;
  }
}

enum { TCG_EPI_NONE = 0, TCG_EPI_MASK = 1 };

template <int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
k_tc_gemm(const float* __restrict__ A, const float* __restrict__ img, float* __restrict__ C, int64_t M, int K, int Nc,
          int nkc, const float* __restrict__ aux, float slope) {
  extern __shared__ __align__(1024) unsigned char gemm_smem[];
  __shared__ uint64_t bar_w[3];
  __shared__ uint64_t bar_mma[2];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t chunk_bytes = 2u * (uint32_t)Nc * kGemmKC * 4u;  // hi + lo plane of one K chunk
  unsigned char* base = gemm_smem + ((1024u - (ptx::smem_u32(gemm_smem) & 1023u)) & 1023u);

  if (warp == 0) ptx::tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) {
    ptx::mbar_init(&bar_w[0], 1);
    ptx::mbar_init(&bar_w[1], 1);
    ptx::mbar_init(&bar_w[2], 1);
    ptx::mbar_init(&bar_mma[0], 1);
    ptx::mbar_init(&bar_mma[1], 1);
    ptx::fence_mbar_init();
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t d_t = tbase;  // A operand of chunk g: columns 256 + 64 (g & 1): hi (32) | lo (32)
  const uint32_t trow = (uint32_t)(warp * 32) << 16;
  const bool vec4 = (K & 3) == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0;

  const int64_t ntiles = (M + 127) >> 7;
  const int64_t my_tiles = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int64_t total_chunks = my_tiles * nkc;

  // W chunk g (image chunk g % nkc) -> slot g % 3
  auto copy_chunk = [&](int64_t g) {
    const int b = (int)(g % 3);
    ptx::mbar_expect_tx(&bar_w[b], chunk_bytes);
    ptx::bulk_g2s(base + (size_t)b * chunk_bytes, img + (size_t)(g % nkc) * (chunk_bytes / 4), chunk_bytes, &bar_w[b]);
  };
  if (tid == 0) {
    for (int64_t c = 0; c < 3 && c < total_chunks; ++c) copy_chunk(c);
  }

  int64_t g = 0;
  float v[32];
  if (my_tiles > 0) load_a_chunk(A, (int64_t)blockIdx.x * 128 + tid, (int64_t)blockIdx.x * 128 + tid < M, K, 0, vec4, v);
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t row = tile * 128 + tid;
    const bool valid = row < M;
    for (int kc = 0; kc < nkc; ++kc, ++g) {
      const uint32_t a_hi = tbase + 256 + 64 * (uint32_t)(g & 1), a_lo = a_hi + 32;
      if (g >= 2) {
        // the MMAs of chunk g - 2 are done: this A operand buffer and W slot (g + 1) % 3 are free again
        ptx::mbar_wait(&bar_mma[g & 1], (uint32_t)(((g - 2) >> 1) & 1));
        if (tid == 0 && g + 1 < total_chunks) copy_chunk(g + 1);
      }
      ptx::tc_fence_after_sync();
      {
        float h[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) h[q] = v[q];
        tc::store_split16(h, a_hi, a_lo, trow, 0);
#pragma unroll
        for (int q = 0; q < 16; ++q) h[q] = v[16 + q];
        tc::store_split16(h, a_hi, a_lo, trow, 16);
      }
      // next chunk of this row (or the first chunk of this CTA's next tile) -> registers, under the MMAs
      {
        const bool last = kc + 1 == nkc;
        const int64_t nrow = last ? row + (int64_t)gridDim.x * 128 : row;
        const bool more = !last || tile + gridDim.x < ntiles;
        if (more) load_a_chunk(A, nrow, nrow < M, K, last ? 0 : kc + 1, vec4, v);
      }
      ptx::tmem_wait_st();
      ptx::tc_fence_before_sync();
      __syncthreads();
      if (tid == 0) {
        ptx::tc_fence_after_sync();
        ptx::mbar_wait(&bar_w[g % 3], (uint32_t)((g / 3) & 1));
        const uint32_t wb = ptx::smem_u32(base + (size_t)(g % 3) * chunk_bytes);
        issue_chunk_mma(d_t, a_hi, a_lo, wb, wb + chunk_bytes / 2, kc == 0, Nc);
        ptx::mma_commit(&bar_mma[g & 1]);
      }
    }
    // ---- epilogue of the tile: accumulator row -> C ------------------------------------------------
    ptx::mbar_wait(&bar_mma[(g - 1) & 1], (uint32_t)(((g - 1) >> 1) & 1));  // the last chunk's commit covers all before it
    ptx::tc_fence_after_sync();
    for (int c0 = 0; c0 < Nc; c0 += 16) {
      float r[16];
      tc::load_acc16(d_t, trow, c0, r);
      if (valid) {
        float* dst = C + row * Nc + c0;
#pragma unroll
        for (int q = 0; q < 16; q += 4) {
          float4 o = make_float4(r[q], r[q + 1], r[q + 2], r[q + 3]);
          if (EPI == TCG_EPI_MASK) {
            const float4 hm = ldg4(aux + row * Nc + c0 + q);
            o.x *= lrelu_grad_from_out(hm.x, slope); o.y *= lrelu_grad_from_out(hm.y, slope);
            o.z *= lrelu_grad_from_out(hm.z, slope); o.w *= lrelu_grad_from_out(hm.w, slope);
          }
          *reinterpret_cast<float4*>(dst + q) = o;
        }
      }
    }
    ptx::tc_fence_before_sync();
    __syncthreads();  // every row has left tensor memory before the next tile's first MMA overwrites D
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tbase, 512);
}

size_t tc_gemm_image_floats(int K, int Nc) {
  const int nkc = (K + kGemmKC - 1) / kGemmKC;
  return (size_t)nkc * 2 * (size_t)Nc * kGemmKC;
}

bool tc_gemm_applicable(int64_t M, int K, int Nc) {
  return M >= 1 && K >= 1 && Nc >= 64 && (Nc & 63) == 0 && Nc <= kGemmMaxNc && K <= 4096;
}

// C = A . B^T with the B operand taken from W as described at the top.  `img`: tc_gemm_image_floats(K, Nc)
// floats of scratch, 16-byte aligned.  aux (mask epilogue): [M, Nc].
int launch_tc_gemm(const float* A, const float* W, float* C, int64_t M, int K, int Nc, bool transposed, int epi,
                   const float* aux, float slope, float* img, cudaStream_t st) {
  if (M == 0) return HCG_OK;
  HCG_REQUIRE(tc_gemm_applicable(M, K, Nc), HCG_ERR_UNSUPPORTED, "tensor-core GEMM needs Nc in {64,128,192,256}, got Nc=%d K=%d", Nc, K);
  HCG_REQUIRE((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (reinterpret_cast<uintptr_t>(img) & 15) == 0 &&
                  (epi != TCG_EPI_MASK || (aux && (reinterpret_cast<uintptr_t>(aux) & 15) == 0)),
              HCG_ERR_INVALID_ARG, "tensor-core GEMM: misaligned C / image / aux pointer");
  const int nkc = (K + kGemmKC - 1) / kGemmKC;
  k_w_image<<<(unsigned)cdiv((int64_t)Nc * nkc * kGemmKC, 256), 256, 0, st>>>(W, Nc, K, transposed ? 1 : 0, nkc, img);
  HCG_LAUNCH_OK(st, "k_w_image");
  const size_t smem = 3 * 2 * (size_t)Nc * kGemmKC * 4 + 1024;
  static SmemConfigured configured[2];
  const int e = epi == TCG_EPI_MASK ? 1 : 0;
  if (int rc = e ? ensure_dyn_smem(k_tc_gemm<TCG_EPI_MASK>, smem, configured[1])
                 : ensure_dyn_smem(k_tc_gemm<TCG_EPI_NONE>, smem, configured[0])) return rc;
  int64_t grid = (M + 127) >> 7;
  if (grid > kNumSMs) grid = kNumSMs;
  if (e) k_tc_gemm<TCG_EPI_MASK><<<(unsigned)grid, kGemmThreads, smem, st>>>(A, img, C, M, K, Nc, nkc, aux, slope);
  else k_tc_gemm<TCG_EPI_NONE><<<(unsigned)grid, kGemmThreads, smem, st>>>(A, img, C, M, K, Nc, nkc, aux, slope);
  HCG_LAUNCH_OK(st, "k_tc_gemm");
  return HCG_OK;
}

}  // namespace hcg
