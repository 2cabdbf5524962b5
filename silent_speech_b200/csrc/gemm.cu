// fp32 CUDA-core GEMM used by the reference-precision path (exact fp32 FMA; the
// tensor-core path lives in gemm_tc.cu).  One kernel covers the three shapes the
// path needs through element strides:
//
//   C[i,j] (+)= sum_k A(i,k) * B(k,j) (+ bias[j]),   A(i,k) = A[i*sai + k*sak],  B(k,j) = B[k*sbk + j*sbj]
//
//   input projection  gi = Z W_ih^T + b_ih      A k-contiguous, B k-contiguous  (train_model_official.py:261-267, nn.GRU)
//   data gradient     dZ = dgi W_ih             A k-contiguous, B j-contiguous
//   weight gradient   dW = dgi^T Z              A i-contiguous, B j-contiguous, split over k with atomics
//
// Optional row mask: when (lengths, T) are given, A/C rows are (clip, frame) pairs and
// 64-row tiles made only of padded frames (t >= lengths[b]) are skipped -- their
// outputs are never read downstream.
#include "common.cuh"

namespace ssb {

constexpr int GBM = 64, GBN = 64, GBK = 16, GTHREADS = 256;

template <bool A_KCONTIG, bool B_KCONTIG>
__global__ void __launch_bounds__(GTHREADS)
sgemm_kernel(const float* __restrict__ A, int64_t sai, int64_t sak,
             const float* __restrict__ Bm, int64_t sbk, int64_t sbj,
             const float* __restrict__ bias, float* __restrict__ C, int64_t ldc,
             int M, int N, int K, int k_per_split, int accumulate, int atomic_out,
             const int64_t* __restrict__ lengths, int T) {
  __shared__ float As[GBK][GBM + 4];
  __shared__ float Bs[GBK][GBN + 4];
  const int tid = threadIdx.x;
  const int i0 = blockIdx.x * GBM, j0 = blockIdx.y * GBN;
  const int kbeg = blockIdx.z * k_per_split, kend = min(K, kbeg + k_per_split);

  if (lengths) {      // skip tiles of padded frames only
    int live = 0;
    if (tid < GBM) { int m = i0 + tid; if (m < M) live = (m % T) < (int)lengths[m / T]; }
    if (!__syncthreads_or(live)) return;
  }

  const int ti = tid / 16, tj = tid % 16;       // 16x16 threads, 4x4 outputs each
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  for (int k0 = kbeg; k0 < kend; k0 += GBK) {
    // ---- stage tiles (1024 elements each, 4 per thread), coalesced along the contiguous index
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int idx = tid + e * GTHREADS;
      int ii, kk;
      if (A_KCONTIG) { kk = idx % GBK; ii = idx / GBK; } else { ii = idx % GBM; kk = idx / GBM; }
      const int gi = i0 + ii, gk = k0 + kk;
      As[kk][ii] = (gi < M && gk < kend) ? A[gi * sai + gk * sak] : 0.f;
      int jj, kb;
      if (B_KCONTIG) { kb = idx % GBK; jj = idx / GBK; } else { jj = idx % GBN; kb = idx / GBN; }
      const int gj = j0 + jj, gkb = k0 + kb;
      Bs[kb][jj] = (gj < N && gkb < kend) ? Bm[gkb * sbk + gj * sbj] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GBK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ti * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tj * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] = fmaf(av[x], bv[y], acc[x][y]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int x = 0; x < 4; ++x) {
    const int gi = i0 + ti * 4 + x;
    if (gi >= M) continue;
#pragma unroll
    for (int y = 0; y < 4; ++y) {
      const int gj = j0 + tj * 4 + y;
      if (gj >= N) continue;
      float v = acc[x][y];
      if (bias && blockIdx.z == 0) v += bias[gj];
      float* c = C + (int64_t)gi * ldc + gj;
      if (atomic_out) atomicAdd(c, v);
      else *c = accumulate ? (*c + v) : v;
    }
  }
}

}  // namespace ssb

// Generic strided fp32 GEMM; see the header comment.  split_k > 1 accumulates with
// atomics into C (which the caller has zeroed, or wants accumulated into).
extern "C" int ssb_sgemm(const float* A, int64_t sai, int64_t sak, const float* Bm, int64_t sbk, int64_t sbj,
                         const float* bias, float* C, int64_t ldc, int64_t M, int64_t N, int64_t K,
                         int32_t accumulate, int32_t split_k, const int64_t* lengths, int64_t T,
                         void* stream) {
  using namespace ssb;
  SSB_REQUIRE(A && Bm && C, "null pointer");
  SSB_REQUIRE(M >= 0 && N >= 0 && K >= 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "bad shape");
  SSB_REQUIRE(sak == 1 || sai == 1, "A must be contiguous along i or k");
  SSB_REQUIRE(sbk == 1 || sbj == 1, "B must be contiguous along k or j");
  SSB_REQUIRE(split_k >= 1, "split_k >= 1");
  SSB_REQUIRE(!lengths || T > 0, "row mask needs T");
  if (M == 0 || N == 0) return SSB_OK;
  if (split_k > 1) SSB_REQUIRE(accumulate, "split_k > 1 accumulates into C");
  int k_per_split = (int)((K + split_k - 1) / split_k);
  k_per_split = ((k_per_split + GBK - 1) / GBK) * GBK;
  if (k_per_split == 0) k_per_split = GBK;
  const int splits = K == 0 ? 1 : (int)((K + k_per_split - 1) / k_per_split);
  dim3 grid(cdiv(M, GBM), cdiv(N, GBN), splits), block(GTHREADS);
  SSB_REQUIRE(grid.y <= 65535u && grid.z <= 65535u, "N or split_k too large");
  const int atomic_out = splits > 1;
  cudaStream_t st = (cudaStream_t)stream;
  const bool ak = (sak == 1), bk = (sbk == 1);
#define SSB_GEMM_LAUNCH(AK, BK)                                                                      \
  sgemm_kernel<AK, BK><<<grid, block, 0, st>>>(A, sai, sak, Bm, sbk, sbj, bias, C, ldc, (int)M, (int)N, \
                                               (int)K, k_per_split, accumulate, atomic_out, lengths, (int)T)
  if (ak && bk) SSB_GEMM_LAUNCH(true, true);
  else if (ak && !bk) SSB_GEMM_LAUNCH(true, false);
  else if (!ak && bk) SSB_GEMM_LAUNCH(false, true);
  else SSB_GEMM_LAUNCH(false, false);
#undef SSB_GEMM_LAUNCH
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
