// Micro-benchmarks that size the ROI-CNN / GRU kernels' design (not part of the product library):
//   (1) tcgen05.ld throughput per SM as a function of the number of reading warps,
//   (2) tcgen05.mma issue->complete time for the small-N shapes the fused kernels use (M=128, K=16, no swizzle).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I silent_speech_b200/csrc -o profiles/microbench/tc_micro profiles/microbench/tc_micro.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tc.cuh"
using namespace ssb;

__global__ void tmem_ld_kernel(int iters, long long* cycles, float* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tc::tmem_alloc<512>(&tmem_base);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t t = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  float acc = 0.f;
  __syncthreads();
  const long long c0 = clock64();
  for (int i = 0; i < iters; ++i) {
    float v[32];
    tc::tmem_ld32(t + ((i * 32 + (warp >> 2) * 64) & 255), v);
    tc::tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 32; q += 8) acc += v[q];
  }
  __syncthreads();
  const long long c1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = c1 - c0;
  if (acc == 123.456f) sink[0] = acc;
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem_base);
}

// two loads in flight per warp before the wait
__global__ void tmem_ld2_kernel(int iters, long long* cycles, float* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tc::tmem_alloc<512>(&tmem_base);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t t = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  float acc = 0.f;
  __syncthreads();
  const long long c0 = clock64();
  for (int i = 0; i < iters; ++i) {
    float v[32], w[32];
    tc::tmem_ld32(t + ((i * 64) & 255), v);
    tc::tmem_ld32(t + ((i * 64 + 32) & 255), w);
    tc::tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 32; q += 8) acc += v[q] + w[q];
  }
  __syncthreads();
  const long long c1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = c1 - c0;
  if (acc == 123.456f) sink[0] = acc;
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem_base);
}

template <int N>
__global__ void mma_kernel(int iters, int per_commit, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_base;
  __shared__ uint64_t bar;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) tc::tmem_alloc<512>(&tmem_base);
  if (threadIdx.x == 32) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = tc::make_idesc_bf16(128, N);
    const uint32_t a = tc::smem_u32(smem), b = a + 32768;
    uint32_t phase = 0;
    const long long c0 = clock64();
    for (int i = 0; i < iters; ++i) {
      for (int j = 0; j < per_commit; ++j) {
        const uint64_t da = tc::make_desc(a + (j & 3) * 4096, 2048, 128), db = tc::make_desc(b + (j & 3) * 4096, 2048, 128);
        tc::mma_bf16(tmem_base + (j & 1) * 256, da, db, idesc, j > 1);
      }
      tc::mma_commit(&bar);
      tc::mbar_wait(&bar, phase & 1);
      ++phase;
    }
    const long long c1 = clock64();
    cycles[blockIdx.x] = c1 - c0;
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem_base);
}

template <int N>
void run_mma(long long* d_cyc, int per_commit) {
  const int iters = 200;
  cudaFuncSetAttribute(mma_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  mma_kernel<N><<<148, 64, 65536>>>(iters, per_commit, d_cyc);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d_cyc, sizeof h, cudaMemcpyDeviceToHost);
  double per = (double)h[0] / iters;
  printf("mma M=128 N=%3d K=16: %3d MMAs per commit: %8.1f cyc per round, %6.1f cyc per MMA (floor %d)  err=%s\n", N, per_commit, per,
         per / per_commit, 128 * N / 256, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d_cyc; float* d_sink;
  cudaMalloc(&d_cyc, 148 * sizeof(long long));
  cudaMalloc(&d_sink, 4);
  const int iters = 2000;
  for (int warps : {1, 2, 4, 8, 12, 16}) {
    tmem_ld_kernel<<<148, warps * 32>>>(iters, d_cyc, d_sink);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d_cyc, sizeof h, cudaMemcpyDeviceToHost);
    const double bytes = (double)warps * iters * 32 * 32 * 4;
    printf("tmem ld x32, %2d warps, 1 in flight: %8.1f cyc/iter  -> %7.1f B/cyc/SM  err=%s\n", warps, (double)h[0] / iters, bytes / h[0],
           cudaGetErrorString(cudaGetLastError()));
    tmem_ld2_kernel<<<148, warps * 32>>>(iters, d_cyc, d_sink);
    cudaDeviceSynchronize();
    cudaMemcpy(h, d_cyc, sizeof h, cudaMemcpyDeviceToHost);
    printf("tmem ld x32, %2d warps, 2 in flight: %8.1f cyc/iter  -> %7.1f B/cyc/SM  err=%s\n", warps, (double)h[0] / iters, 2 * bytes / h[0],
           cudaGetErrorString(cudaGetLastError()));
  }
  for (int pc : {1, 4, 16, 64}) {
    run_mma<32>(d_cyc, pc); run_mma<64>(d_cyc, pc); run_mma<128>(d_cyc, pc); run_mma<144>(d_cyc, pc); run_mma<256>(d_cyc, pc);
  }
  return 0;
}
