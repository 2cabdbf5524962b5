// Library-level entry points of libssb200.so: version, error channel, device probe.
#include "common.cuh"
#include <string.h>

static thread_local char g_err[512] = "";

extern "C" void ssb_set_error_(const char* msg) {
  strncpy(g_err, msg, sizeof g_err - 1);
  g_err[sizeof g_err - 1] = 0;
}

// Last error message of the calling thread ("" if none).
extern "C" const char* ssb_last_error(void) { return g_err; }

// ABI version: bumped whenever a signature in include/ssb200.h changes.
extern "C" int ssb_version(void) { return 100; }

// Compute capability and SM count of the current device; returns SSB_ERR_UNSUPPORTED
// unless it is an sm_100 part (the kernels are built for sm_100a only -- no fallback).
extern "C" int ssb_device_info(int32_t* sm_count, int32_t* cc_major, int32_t* cc_minor) {
  int dev = 0;
  SSB_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp p;
  SSB_CUDA(cudaGetDeviceProperties(&p, dev));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (p.major != 10) {
    ssb_set_error_("libssb200 is built for sm_100a (B200) only");
    return SSB_ERR_UNSUPPORTED;
  }
  return SSB_OK;
}

// Per-device ring of launch counters (see common.cuh).
#include <atomic>
extern "C" int ssb_scratch_counter_(cudaStream_t stream, unsigned int** counter) {
  constexpr int kMaxDev = 64, kRing = 1024;
  static unsigned int* ring[kMaxDev] = {};
  static std::atomic<unsigned int> next[kMaxDev];
  static std::atomic_flag lock = ATOMIC_FLAG_INIT;
  int dev = 0;
  SSB_CUDA(cudaGetDevice(&dev));
  SSB_REQUIRE(dev >= 0 && dev < kMaxDev, "device index out of range");
  if (!ring[dev]) {
    while (lock.test_and_set(std::memory_order_acquire)) {}
    cudaError_t e = cudaSuccess;
    if (!ring[dev]) {
      unsigned int* p = nullptr;
      e = cudaMalloc(&p, kRing * sizeof(unsigned int));
      if (e == cudaSuccess) ring[dev] = p;
    }
    lock.clear(std::memory_order_release);
    SSB_CUDA(e);
  }
  unsigned int* c = ring[dev] + next[dev].fetch_add(1) % kRing;
  SSB_CUDA(cudaMemsetAsync(c, 0, sizeof(unsigned int), stream));
  *counter = c;
  return SSB_OK;
}
