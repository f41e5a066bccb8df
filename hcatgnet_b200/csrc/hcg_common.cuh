// Shared helpers for libhcg (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/hcg.h"

namespace hcg {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// thread-local error text behind hcg_last_error()
char* err_buf();
int set_err(int code, const char* fmt, ...);

// hcg_runtime.cu: launch counter + event profiler hooks (cheap no-ops unless profiling is on)
void note_launch(const char* name, cudaStream_t st);   // a kernel of ours was launched
void note_region(cudaStream_t st);                     // an API call starts issuing work
void note_other(const char* name, cudaStream_t st);    // non-kernel stream work (memset)
int config_get(int key);                               // hcg_config_get

#define HCG_CUDA_OK(expr)                                                             \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess)                                                            \
      return ::hcg::set_err(HCG_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,             \
                            cudaGetErrorString(_e), __FILE__, __LINE__);              \
  } while (0)

// after every kernel launch: error check, launch counter, optional profiler mark
#define HCG_LAUNCH_OK(stream, name)                                                   \
  do {                                                                                \
    cudaError_t _e = cudaGetLastError();                                              \
    if (_e != cudaSuccess)                                                            \
      return ::hcg::set_err(HCG_ERR_CUDA, "launch of %s failed: %s", name,            \
                            cudaGetErrorString(_e));                                  \
    ::hcg::note_launch(name, stream);                                                 \
  } while (0)

#define HCG_REQUIRE(cond, code, ...)                                                  \
  do {                                                                                \
    if (!(cond)) return ::hcg::set_err(code, __VA_ARGS__);                            \
  } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE setting: the "already configured" note is kept
// per device (a process may drive several GPUs), indexed by cudaGetDevice().
constexpr int kMaxDevices = 64;
struct SmemConfigured { size_t bytes[kMaxDevices] = {}; };
template <typename K>
inline int ensure_dyn_smem(K kernel, size_t smem, SmemConfigured& done) {
  int dev = 0;
  HCG_CUDA_OK(cudaGetDevice(&dev));
  const bool known = dev >= 0 && dev < kMaxDevices;
  if (!known || smem > done.bytes[dev]) {
    HCG_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (known) done.bytes[dev] = smem;
  }
  return HCG_OK;
}

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }
inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Carves aligned sub-buffers out of the caller's workspace.
struct Carver {
  char* base;
  size_t off;
  size_t cap;
  Carver(void* p, size_t bytes) : base(static_cast<char*>(p)), off(0), cap(bytes) {}
  template <typename T>
  T* take(size_t count) {
    size_t bytes = align_up(count * sizeof(T));
    T* r = reinterpret_cast<T*>(base + off);
    off += bytes;
    return r;
  }
  bool ok() const { return off <= cap; }
};

__device__ __forceinline__ float lrelu(float v, float slope) { return v > 0.f ? v : v * slope; }
// derivative of LeakyReLU recovered from the activation's OUTPUT (sign is preserved; 0 -> slope,
// as torch's leaky_relu_backward does for input == 0)
__device__ __forceinline__ float lrelu_grad_from_out(float h, float slope) { return h > 0.f ? 1.f : slope; }

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

#ifdef __CUDACC__
// Deterministic column sum of `rows` partial rows of pitch `pe` for a block of 32 consecutive elements: 256 threads =
// 8 row slices x 32 elements (coalesced 128-byte loads); slice q adds rows q, q + 8, ... in order into two interleaved
// sums, the eight slices are combined in a fixed order.  ~rows / 16 dependent loads instead of rows / 4.
// Returns the sum for element blockIdx.x * 32 + (threadIdx.x & 31) in the threads of slice 0 (others: undefined).
__device__ __forceinline__ float det_colsum32(const float* __restrict__ partial, int rows, int pe, int e, bool valid, float* s_red) {
  const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float s0 = 0.f, s1 = 0.f;
  if (valid) {
    int k = q;
    for (; k + 8 < rows; k += 16) {
      s0 += partial[(int64_t)k * pe + e];
      s1 += partial[(int64_t)(k + 8) * pe + e];
    }
    if (k < rows) s0 += partial[(int64_t)k * pe + e];
  }
  s_red[q * 32 + lane] = s0 + s1;
  __syncthreads();
  float v = 0.f;
  if (q == 0) {
    v = ((s_red[lane] + s_red[32 + lane]) + (s_red[64 + lane] + s_red[96 + lane])) +
        ((s_red[128 + lane] + s_red[160 + lane]) + (s_red[192 + lane] + s_red[224 + lane]));
  }
  return v;
}
#endif

}  // namespace hcg
