// K3a -- GRU input projection on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM).
//
//   C[M,N] = [A1 | A2] (M x (K1+K2), fp32) . W[N, K1+K2]^T (fp32) + bias        (nn.GRU's x W_ih^T + b_ih,
//   train_model_official.py:261-267; A1 = point features, A2 = ROI embedding: the concat of train:297
//   is two K-segments of one product and is never materialised)
//
// Operands stay fp32 in HBM (the state-dict contract); each CTA converts its tiles to fp16 on the way
// into shared memory, so there is no separate cast pass.  passes = 1: plain fp16 (rel. error ~2^-12 per
// product); passes = 3: split-fp16, x = hi + lo, D += Ahi.Bhi + Ahi.Blo + Alo.Bhi (error ~2^-22, i.e. fp32
// parity grade).  One CTA = one 128 x 192 output tile (UMMA M=128, N=192, fp32 accumulators = 192 TMEM
// columns), K in chunks of 64 through a 2-stage shared-memory ring; one elected thread issues the MMAs
// and tcgen05.commit signals an mbarrier when a stage may be overwritten.
#include "common.cuh"
#include "tc.cuh"

namespace ssb {
namespace gemmtc {

constexpr int BM = 128, BN = 192, BK = 32, THREADS = 256, STAGES = 2;   // 40 KB per stage -> 2 CTAs per SM
constexpr int LBO = 128, SBO = (BK / 8) * 128;            // BK/8 k-chunks of 16 B per 8-row group
constexpr int A_BYTES = (BM / 8) * SBO, B_BYTES = (BN / 8) * SBO;   // 16 KB, 24 KB
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;              // hi + lo of both = 80 KB
constexpr int TMEM_COLS = 256;

struct Tail {
  uint64_t bar[STAGES];
  uint32_t tmem_base;
};

// rows x BK fp32 (global, row stride ld) -> fp16 hi (and lo) in the canonical no-swizzle K-major layout
template <int ROWS, bool kSplit>
__device__ __forceinline__ void stage_operand(const float* __restrict__ src, int64_t ld, int rows_valid, int k_valid,
                                              uint8_t* hi, uint8_t* lo, int tid) {
#pragma unroll
  for (int i = 0; i < ROWS * (BK / 8) / THREADS; ++i) {
    const int p = tid + i * THREADS;
    const int r8 = p & 7, kc = (p >> 3) & (BK / 8 - 1), rg = p / (8 * (BK / 8));
    const int row = rg * 8 + r8, k = kc * 8;
    float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
    if (row < rows_valid) {
      const float* g = src + (int64_t)row * ld + k;
      if (k < k_valid) v0 = __ldg(reinterpret_cast<const float4*>(g));
      if (k + 4 < k_valid) v1 = __ldg(reinterpret_cast<const float4*>(g + 4));
    }
    const float x[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    const uint32_t off = rg * SBO + kc * LBO + r8 * 16;
    if (kSplit) {
      float h[8], l[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) tc::split_f16(x[j], h[j], l[j]);
      *reinterpret_cast<uint4*>(hi + off) = make_uint4(tc::pack_f16(h[0], h[1]), tc::pack_f16(h[2], h[3]),
                                                       tc::pack_f16(h[4], h[5]), tc::pack_f16(h[6], h[7]));
      *reinterpret_cast<uint4*>(lo + off) = make_uint4(tc::pack_f16(l[0], l[1]), tc::pack_f16(l[2], l[3]),
                                                       tc::pack_f16(l[4], l[5]), tc::pack_f16(l[6], l[7]));
    } else {
      *reinterpret_cast<uint4*>(hi + off) = make_uint4(tc::pack_f16(x[0], x[1]), tc::pack_f16(x[2], x[3]),
                                                       tc::pack_f16(x[4], x[5]), tc::pack_f16(x[6], x[7]));
    }
  }
}

template <bool kSplit>
__global__ void __launch_bounds__(THREADS, 2)
gemm_tc_kernel(const float* __restrict__ A1, int64_t lda1, int K1, const float* __restrict__ A2, int64_t lda2, int K2,
               const float* __restrict__ W, int64_t ldw, const float* __restrict__ bias,
               float* __restrict__ C, int64_t ldc, int M, int N, int accumulate,
               const int64_t* __restrict__ lengths, int T) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Tail& tail = *reinterpret_cast<Tail*>(smem + STAGES * STAGE_BYTES);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // known warp-uniform (MMA descriptors stay in uniform registers)
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;

  if (lengths) {          // tiles made only of padded frames are never read downstream
    int live = 0;
    if (tid < BM) { const int m = m0 + tid; if (m < M) live = (m % T) < (int)lengths[m / T]; }
    if (!__syncthreads_or(live)) return;
  }
  if (warp == 0) tc::tmem_alloc<TMEM_COLS>(&tail.tmem_base);
  if (tid == 32) {
    for (int s = 0; s < STAGES; ++s) tc::mbar_init(&tail.bar[s], 1);
    tc::mbar_fence_init();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = __shfl_sync(0xffffffffu, tail.tmem_base, 0);
  constexpr uint32_t idesc = tc::make_idesc_f16(BM, BN);

  const int chunks1 = (K1 + BK - 1) / BK, chunks2 = A2 ? (K2 + BK - 1) / BK : 0;
  const int chunks = chunks1 + chunks2;
  bool first = true;
  for (int c = 0; c < chunks; ++c) {
    const int s = c % STAGES;
    uint8_t* st = smem + s * STAGE_BYTES;
    uint8_t *a_hi = st, *a_lo = st + A_BYTES, *b_hi = st + 2 * A_BYTES, *b_lo = st + 2 * A_BYTES + B_BYTES;
    if (c >= STAGES) tc::mbar_wait(&tail.bar[s], ((c / STAGES) - 1) & 1);     // MMAs that read this stage are done
    const bool seg2 = c >= chunks1;
    const int k0 = (seg2 ? c - chunks1 : c) * BK;
    const int kseg = seg2 ? K2 : K1;
    const int k_valid = min(BK, kseg - k0);
    const float* a_src = seg2 ? A2 + (int64_t)m0 * lda2 + k0 : A1 + (int64_t)m0 * lda1 + k0;
    const float* w_src = W + (int64_t)n0 * ldw + (seg2 ? K1 : 0) + k0;
    stage_operand<BM, kSplit>(a_src, seg2 ? lda2 : lda1, M - m0, k_valid, a_hi, a_lo, tid);
    stage_operand<BN, kSplit>(w_src, ldw, N - n0, k_valid, b_hi, b_lo, tid);
    tc::fence_async_smem();
    __syncthreads();
    if (warp == 0 && tc::elect_one()) {
      tc::fence_after_sync();
      const int ksteps = (k_valid + 15) / 16;
      for (int j = 0; j < ksteps; ++j) {
        const uint64_t dah = tc::make_desc(tc::smem_u32(a_hi) + j * 2 * LBO, LBO, SBO);
        const uint64_t dbh = tc::make_desc(tc::smem_u32(b_hi) + j * 2 * LBO, LBO, SBO);
        tc::mma_f16(tmem, dah, dbh, idesc, !first);
        first = false;
        if (kSplit) {
          const uint64_t dal = tc::make_desc(tc::smem_u32(a_lo) + j * 2 * LBO, LBO, SBO);
          const uint64_t dbl = tc::make_desc(tc::smem_u32(b_lo) + j * 2 * LBO, LBO, SBO);
          tc::mma_f16(tmem, dah, dbl, idesc, true);
          tc::mma_f16(tmem, dal, dbh, idesc, true);
        }
      }
      tc::mma_commit(&tail.bar[s]);
    }
  }
  // ---- epilogue: wait for the last commit (commits complete in order), TMEM -> regs -> + bias -> C
  {
    const int c = chunks - 1;
    tc::mbar_wait(&tail.bar[c % STAGES], (c / STAGES) & 1);
    tc::fence_after_sync();
    // all MMAs have completed, so the operand stages are free: each warp transposes its 32x32 blocks
    // through a private padded smem tile so that global stores are full 128-byte row segments
    float* tile = reinterpret_cast<float*>(smem) + warp * (32 * 33);
    const int row0 = m0 + (warp & 3) * 32;
    const int jbeg = (warp >> 2) * 3;
#pragma unroll 1
    for (int j = jbeg; j < jbeg + 3; ++j) {
      float v[32];
      tc::tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + j * 32, v);
      tc::tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 32; ++q) tile[lane * 33 + q] = v[q];
      __syncwarp();
      const int n = n0 + j * 32 + lane;
      const float bv = (bias && n < N) ? __ldg(bias + n) : 0.f;
      if (n < N) {
#pragma unroll 4
        for (int r = 0; r < 32; ++r) {
          if (row0 + r < M) {
            float* c = C + (int64_t)(row0 + r) * ldc + n;
            float o = tile[r * 33 + lane] + bv;
            if (accumulate) o += *c;
            *c = o;
          }
        }
      }
      __syncwarp();
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<TMEM_COLS>(tmem);
}

}  // namespace gemmtc
}  // namespace ssb

// Tensor-core projection GEMM; see the header comment.  passes = 1 (fp16) or 3 (split-fp16, fp32 grade).
// Requirements: lda*, ldw, K1, K2 multiples of 4 floats and 16-byte aligned bases (true for 180/32/212/384).
extern "C" int ssb_gemm_tc(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                           const float* W, int64_t ldw, const float* bias, float* C, int64_t ldc,
                           int64_t M, int64_t N, int32_t passes, int32_t accumulate,
                           const int64_t* lengths, int64_t T, void* stream) {
  using namespace ssb;
  using namespace ssb::gemmtc;
  SSB_REQUIRE(A1 && W && C, "null pointer");
  SSB_REQUIRE(passes == 1 || passes == 3, "passes must be 1 or 3");
  SSB_REQUIRE(K1 > 0 && K1 % 4 == 0 && lda1 % 4 == 0 && ldw % 4 == 0, "K1/lda1/ldw must be multiples of 4");
  SSB_REQUIRE(!A2 || (K2 > 0 && K2 % 4 == 0 && lda2 % 4 == 0), "K2/lda2 must be multiples of 4");
  SSB_REQUIRE(((uintptr_t)A1 & 15) == 0 && ((uintptr_t)W & 15) == 0 && ((uintptr_t)A2 & 15) == 0, "operands must be 16-byte aligned");
  SSB_REQUIRE(!lengths || T > 0, "row mask needs T");
  if (M == 0 || N == 0) return SSB_OK;
  const size_t smem = STAGES * STAGE_BYTES + sizeof(Tail);
  dim3 grid(cdiv(N, BN), cdiv(M, BM)), block(THREADS);
  SSB_REQUIRE(grid.y <= 65535u, "M too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (passes == 3) {
    SSB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gemm_tc_kernel<true><<<grid, block, smem, st>>>(A1, lda1, (int)K1, A2, lda2, (int)K2, W, ldw, bias, C, ldc,
                                                    (int)M, (int)N, accumulate, lengths, (int)T);
  } else {
    SSB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gemm_tc_kernel<false><<<grid, block, smem, st>>>(A1, lda1, (int)K1, A2, lda2, (int)K2, W, ldw, bias, C, ldc,
                                                     (int)M, (int)N, accumulate, lengths, (int)T);
  }
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
