// Micro-benchmark 2: what bounds small-N tcgen05.mma (M=128 or 64, K=16, no swizzle) -- the per-instruction floor,
// the operand layout, or the issuing thread?  Variants: M, number of concurrently issuing warps, CTAs per SM, layout.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I silent_speech_b200/csrc -o profiles/microbench/tc_micro2 profiles/microbench/tc_micro2.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tc.cuh"
using namespace ssb;

template <int M, int N>
__global__ void __launch_bounds__(128) mma_kernel(int iters, int per_commit, int issuers, uint32_t a_lbo, uint32_t a_sbo, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_base;
  __shared__ uint64_t bar[4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 98304 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) tc::tmem_alloc<256>(&tmem_base);
  if (threadIdx.x == 32) { for (int i = 0; i < 4; ++i) tc::mbar_init(&bar[i], 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  if (warp < issuers && lane == 0) {
    constexpr uint32_t idesc = tc::make_idesc_bf16(M, N);
    const uint32_t a = tc::smem_u32(smem) + warp * 16384, b = tc::smem_u32(smem) + 65536 + warp * 8192;
    const uint32_t d = tmem_base + warp * 64;      // N <= 64 when issuers > 1
    uint32_t phase = 0;
    const long long c0 = clock64();
    for (int i = 0; i < iters; ++i) {
      for (int j = 0; j < per_commit; ++j) {
        const uint64_t da = tc::make_desc(a + (j & 3) * 32, a_lbo, a_sbo), db = tc::make_desc(b + (j & 1) * 4096, 2048, 128);
        tc::mma_bf16(d, da, db, idesc, j > 0);
      }
      tc::mma_commit(&bar[warp]);
      tc::mbar_wait(&bar[warp], phase & 1);
      ++phase;
    }
    const long long c1 = clock64();
    if (warp == 0) cycles[blockIdx.x] = c1 - c0;
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<256>(tmem_base);
}

template <int M, int N>
void run(long long* d_cyc, int per_commit, int issuers, int ctas_per_sm, uint32_t lbo, uint32_t sbo, const char* tag) {
  const int iters = 100;
  cudaFuncSetAttribute(mma_kernel<M, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304);
  mma_kernel<M, N><<<148 * ctas_per_sm, 128, 98304>>>(iters, per_commit, issuers, lbo, sbo, d_cyc);
  cudaDeviceSynchronize();
  long long h[148 * 2];
  cudaMemcpy(h, d_cyc, sizeof(long long) * 148 * ctas_per_sm, cudaMemcpyDeviceToHost);
  double per = (double)h[0] / iters / per_commit;
  printf("M=%3d N=%3d  %d MMAs/commit  issuers=%d ctas/sm=%d  A(lbo=%5u,sbo=%4u) %-10s: %6.1f cyc per MMA per issuer -> %6.1f cyc per MMA per SM   %s\n",
         M, N, per_commit, issuers, ctas_per_sm, lbo, sbo, tag, per, per / (issuers * ctas_per_sm), cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d_cyc;
  cudaMalloc(&d_cyc, 148 * 2 * sizeof(long long));
  // baseline layout: k-chunks 2048 B apart, 8-row groups contiguous
  run<128, 32>(d_cyc, 32, 1, 1, 2048, 128, "canonical");
  run<128, 64>(d_cyc, 32, 1, 1, 2048, 128, "canonical");
  run<64, 32>(d_cyc, 32, 1, 1, 2048, 128, "canonical");
  run<64, 64>(d_cyc, 32, 1, 1, 2048, 128, "canonical");
  run<64, 128>(d_cyc, 32, 1, 1, 2048, 128, "canonical");
  // several issuers in one CTA, several CTAs per SM
  run<128, 32>(d_cyc, 32, 2, 1, 2048, 128, "canonical");
  run<128, 32>(d_cyc, 32, 4, 1, 2048, 128, "canonical");
  run<128, 64>(d_cyc, 32, 2, 1, 2048, 128, "canonical");
  run<128, 32>(d_cyc, 32, 1, 2, 2048, 128, "canonical");
  run<128, 64>(d_cyc, 32, 2, 2, 2048, 128, "canonical");
  // layouts of the fused kernels: k-chunks adjacent (LBO=16..), row groups strided
  run<128, 32>(d_cyc, 32, 1, 1, 16, 256, "lbo16");
  run<128, 32>(d_cyc, 32, 1, 1, 128, 256, "kmajor-blk");
  run<128, 64>(d_cyc, 32, 1, 1, 5200, 128, "conv2-like");
  run<128, 32>(d_cyc, 32, 1, 1, 7488, 416, "conv3-like");
  run<128, 32>(d_cyc, 32, 1, 1, 2048, 144, "sbo144");
  // long chains (amortise the commit round trip)
  run<128, 32>(d_cyc, 128, 1, 1, 2048, 128, "canonical");
  run<128, 96>(d_cyc, 128, 1, 1, 2048, 128, "canonical");
  run<128, 128>(d_cyc, 128, 1, 1, 2048, 128, "canonical");
  run<128, 192>(d_cyc, 128, 1, 1, 2048, 128, "canonical");
  run<128, 256>(d_cyc, 128, 1, 1, 2048, 128, "canonical");
  return 0;
}
