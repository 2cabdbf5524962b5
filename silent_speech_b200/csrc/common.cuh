// Shared helpers for the sm_100a kernels behind libssb200.so.
//
// Error convention: every extern "C" entry point returns
// 0 on success or a non-zero code, with a thread-local message retrievable via
// ssb_last_error().  Nothing throws or aborts across the ABI.  All entry points
// only enqueue work on the caller's stream and return.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define SSB_OK 0
#define SSB_ERR_ARG 1
#define SSB_ERR_CUDA 2
#define SSB_ERR_UNSUPPORTED 3

extern "C" void ssb_set_error_(const char* msg);

#define SSB_REQUIRE(cond, msg)                                                        \
  do {                                                                                \
    if (!(cond)) {                                                                    \
      char b_[512];                                                                   \
      snprintf(b_, sizeof b_, "%s:%d: %s (%s)", __FILE__, __LINE__, msg, #cond);      \
      ssb_set_error_(b_);                                                             \
      return SSB_ERR_ARG;                                                             \
    }                                                                                 \
  } while (0)

#define SSB_CUDA(call)                                                                \
  do {                                                                                \
    cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess) {                                                          \
      char b_[512];                                                                   \
      snprintf(b_, sizeof b_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call,           \
               cudaGetErrorString(e_));                                               \
      ssb_set_error_(b_);                                                             \
      return SSB_ERR_CUDA;                                                            \
    }                                                                                 \
  } while (0)

#define SSB_LAUNCH_CHECK() SSB_CUDA(cudaGetLastError())

namespace ssb {

constexpr int kSMs = 148;          // B200: 2 dies x 74 SMs
constexpr int kHidden = 192;       // GRU hidden size (train_model_official.py:402)
constexpr int kGates = 3 * kHidden;
constexpr int kRoiH = 48, kRoiW = 96, kRoiEmb = 32;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// Counter-based RNG for dropout masks: the same (seed, index) gives the same
// keep/drop decision in forward and backward, on any grid shape.
__device__ __forceinline__ uint32_t mix32(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (uint32_t)(z >> 32);
}
__device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t idx, float p) {
  // keep with probability 1-p
  return (mix32(seed, idx) >> 8) * (1.0f / 16777216.0f) >= p;
}

inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace ssb

// A zeroed 32-bit device counter for one kernel launch on `stream` (dynamic work hand-out of persistent
// kernels).  The only device memory the library owns: one 4 KB ring of counters per device, allocated on first
// use and never freed; consecutive launches rotate through the ring, so launches on different streams do not
// share a counter.  Returns SSB_OK or an error code.
extern "C" int ssb_scratch_counter_(cudaStream_t stream, unsigned int** counter);
