// Micro-benchmark 3 (r2): the per-MMA cost once the issue path is warp-uniform (elect.sync, descriptors in uniform
// registers: UIADD3 + UTCHMMA per MMA) -- the r1 numbers of tc_micro2 were taken with `threadIdx.x == 0` issue, which the
// compiler wraps in a 14-instruction ELECT / R2UR waterfall per MMA.  M = 128, K = 16, no swizzle; N and the number of
// concurrently issuing warps vary; the A operand advances by 4 KB per MMA like a conv tile sweep.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I silent_speech_b200/csrc -o profiles/microbench/tc_micro3 profiles/microbench/tc_micro3.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tc.cuh"
using namespace ssb;

template <int N, int PER>
__global__ void __launch_bounds__(128) mma_kernel(int iters, int issuers, int same_a, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_base;
  __shared__ uint64_t bar[4];
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  for (int i = threadIdx.x; i < 196608 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) tc::tmem_alloc<512>(&tmem_base);
  if (threadIdx.x == 32) { for (int i = 0; i < 4; ++i) tc::mbar_init(&bar[i], 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = __shfl_sync(0xffffffffu, tmem_base, 0);
  if (warp < issuers) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_f16(128, N);
      const uint64_t da0 = tc::make_desc(tc::smem_u32(smem) + warp * 32768, 2048, 128);
      const uint64_t db0 = tc::make_desc(tc::smem_u32(smem) + 131072 + warp * 16384, 2048, 128);
      const uint32_t d = tmem + warp * 128;
      uint32_t phase = 0;
      const long long c0 = clock64();
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < PER; ++j)
          tc::mma_f16(d, tc::desc_add(da0, same_a ? 0 : (j & 7) * 4096), tc::desc_add(db0, (j & 1) * 4096), idesc, j > 0);
        tc::mma_commit(&bar[warp]);
        tc::mbar_wait(&bar[warp], phase & 1);
        ++phase;
      }
      const long long c1 = clock64();
      if (warp == 0) cycles[blockIdx.x] = c1 - c0;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem_base);
}

template <int N, int PER>
void run(long long* d_cyc, int issuers, int same_a) {
  const int iters = 100;
  cudaFuncSetAttribute(mma_kernel<N, PER>, cudaFuncAttributeMaxDynamicSharedMemorySize, 196608);
  mma_kernel<N, PER><<<148, 128, 196608>>>(iters, issuers, same_a, d_cyc);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d_cyc, sizeof h, cudaMemcpyDeviceToHost);
  const double per = (double)h[0] / iters / PER;
  printf("N=%3d  %3d MMAs/commit  issuers=%d  %s: %6.1f cyc per MMA per issuer -> %6.1f cyc per MMA per SM  (math floor %d, A fetch 32)  %s\n",
         N, PER, issuers, same_a ? "same A " : "A sweeps", per, per / issuers, N / 2, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d_cyc;
  cudaMalloc(&d_cyc, 148 * sizeof(long long));
  run<32, 1>(d_cyc, 1, 0);
  run<32, 64>(d_cyc, 1, 0); run<64, 64>(d_cyc, 1, 0); run<128, 64>(d_cyc, 1, 0);
  run<32, 64>(d_cyc, 1, 1); run<64, 64>(d_cyc, 1, 1); run<128, 64>(d_cyc, 1, 1);
  run<32, 64>(d_cyc, 2, 0); run<64, 64>(d_cyc, 2, 0); run<128, 64>(d_cyc, 2, 0);
  run<32, 64>(d_cyc, 4, 0); run<64, 64>(d_cyc, 4, 0); run<128, 64>(d_cyc, 4, 0);
  run<32, 16>(d_cyc, 4, 0); run<64, 16>(d_cyc, 4, 0); run<128, 16>(d_cyc, 4, 0);
  run<32, 6>(d_cyc, 4, 0); run<64, 8>(d_cyc, 4, 0);
  return 0;
}
