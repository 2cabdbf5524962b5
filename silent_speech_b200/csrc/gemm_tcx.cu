// Backward-pass GEMMs of the GRU on the tensor cores (tcgen05.mma, accumulators in TMEM).
//
// The weight and data gradients of nn.GRU's two linear maps (train_model_official.py:261-267, autograd of
// `loss.backward()` train:437) are plain matrix products whose operands sit in HBM in the "wrong" orientation for a
// K-major MMA operand:
//
//   weight gradient   dW[i,j] += sum_m dgi[m, i] * Z[m, j]      contraction over the B*T frames: BOTH operands are
//                                                               contiguous along their output index (i / j), not along m
//   data gradient     dZ[m,j]  = sum_i dgi[m, i] * W_ih[i, j]   A is contraction-contiguous, B is j-contiguous
//
// So the general entry point takes element strides like ssb_sgemm,
//   C[i,j] (+)= sum_k A[i*sai + k*sak] * B[k*sbk + j*sbj],
// and each operand is staged by one of two loaders: contraction-contiguous rows are read as float4 along k; an
// output-contiguous operand is read with one coalesced 128-byte row segment per warp and frame and TRANSPOSED on the
// way into shared memory (each thread collects 8 consecutive k of one output index = one 16-byte row of a core
// matrix), so the MMA always sees the canonical no-swizzle K-major layout and no transposed copy ever exists in HBM.
//
// Precision: operands are split x = hi + lo in fp16 (22 significant bits, like the forward tc3 mode) and
// D += Ahi.Bhi + Ahi.Blo + Alo.Bhi in fp32.  Back-propagated gate gradients reach 1e-8 and below -- under fp16's
// subnormal range -- so the A operand is multiplied by a power of two chosen from max|A| (ssb_absmax, one pass over
// the gradient tensor per GRU layer; the scale is read from device memory, no host synchronisation) that puts its
// largest element in [2048, 4096); elements more than 2^-23 below the maximum keep an absolute error of 2^-25 of it.
// The epilogue multiplies by the inverse power of two (exact).  A first version used a bf16 split (fp32's exponent
// range, no scaling): its 2^-17 representation error passed the 1e-4 per-tensor gradient test but moved weights whose
// gradient is a cancellation residue by 10 % of an Adam step (test_three_train_steps_vs_oracle), so it was dropped.
//
// One CTA = one 128 x 192 output tile (fp32 accumulators = 192 TMEM columns) over one K split; K in chunks of 32
// through a 2-stage ring; the global loads of chunk c+1 are issued before the wait for the stage they will overwrite;
// 2 CTAs per SM.  Split-K partial tiles are combined with fp32 atomics (128-byte row segments per warp), exactly as the
// fp32 CUDA-core GEMM it replaces did.
#include "common.cuh"
#include "tc.cuh"
#include <algorithm>

namespace ssb {
namespace gemmx {

constexpr int BM = 128, BN = 192, BK = 32, THREADS = 256, STAGES = 2;
constexpr int LBO = 128, SBO = (BK / 8) * 128;
constexpr int A_BYTES = (BM / 8) * SBO, B_BYTES = (BN / 8) * SBO;   // 8 KB, 12 KB
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;              // hi + lo of both = 40 KB
constexpr int TMEM_COLS = 256;

struct Tail {
  uint64_t bar[STAGES];
  uint32_t tmem_base;
};

// 8 consecutive-k values of one operand row (times a power-of-two scale) -> 16 bytes of hi and of lo
__device__ __forceinline__ void split_store(const float (&x)[8], float scale, uint8_t* hi, uint8_t* lo, uint32_t off) {
  float h[8], l[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) tc::split_f16(x[j] * scale, h[j], l[j]);
  *reinterpret_cast<uint4*>(hi + off) = make_uint4(tc::pack_f16(h[0], h[1]), tc::pack_f16(h[2], h[3]),
                                                   tc::pack_f16(h[4], h[5]), tc::pack_f16(h[6], h[7]));
  *reinterpret_cast<uint4*>(lo + off) = make_uint4(tc::pack_f16(l[0], l[1]), tc::pack_f16(l[2], l[3]),
                                                   tc::pack_f16(l[4], l[5]), tc::pack_f16(l[6], l[7]));
}

// One operand tile of ROWS output indices x BK contraction indices, held in registers between load and store.
//   kMN = false: element (r, k) at src[r*ld + k]   (k-contiguous; float4 loads; ld, k0 multiples of 4, 16-byte aligned)
//   kMN = true : element (r, k) at src[k*ld + r]   (r-contiguous; one 128-byte segment per warp and k; any alignment)
template <int ROWS, bool kMN>
struct Loader {
  static constexpr int ITEMS = ROWS * (BK / 8) / THREADS;
  float x[ITEMS][8];

  __device__ __forceinline__ void load(const float* __restrict__ src, int64_t ld, int rows_valid, int k_valid, int tid) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int p = tid + i * THREADS;
      if (kMN) {
        const int r = p % ROWS, kc = p / ROWS;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int k = kc * 8 + j;
          x[i][j] = (r < rows_valid && k < k_valid) ? __ldg(src + (int64_t)k * ld + r) : 0.f;
        }
      } else {
        const int r8 = p & 7, kc = (p >> 3) & (BK / 8 - 1), rg = p / (8 * (BK / 8));
        const int row = rg * 8 + r8, k = kc * 8;
        float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
        if (row < rows_valid) {
          const float* g = src + (int64_t)row * ld + k;
          if (k < k_valid) v0 = __ldg(reinterpret_cast<const float4*>(g));
          if (k + 4 < k_valid) v1 = __ldg(reinterpret_cast<const float4*>(g + 4));
        }
        x[i][0] = v0.x; x[i][1] = v0.y; x[i][2] = v0.z; x[i][3] = v0.w;
        x[i][4] = v1.x; x[i][5] = v1.y; x[i][6] = v1.z; x[i][7] = v1.w;
      }
    }
  }
  __device__ __forceinline__ void store(uint8_t* hi, uint8_t* lo, float scale, int tid) const {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int p = tid + i * THREADS;
      uint32_t off;
      if (kMN) {
        const int r = p % ROWS, kc = p / ROWS;
        off = (r >> 3) * SBO + kc * LBO + (r & 7) * 16;
      } else {
        const int r8 = p & 7, kc = (p >> 3) & (BK / 8 - 1), rg = p / (8 * (BK / 8));
        off = rg * SBO + kc * LBO + r8 * 16;
      }
      split_store(x[i], scale, hi, lo, off);
    }
  }
};

template <bool kAmn, bool kBmn>
__global__ void __launch_bounds__(THREADS, 2)
gemm_tcx_kernel(const float* __restrict__ A, int64_t lda, const float* __restrict__ Bm, int64_t ldb,
                float* __restrict__ C, int64_t ldc, int M, int N, int K, int k_per_split, int accumulate, int atomic_out,
                const int64_t* __restrict__ lengths, int T, const float* __restrict__ a_amax) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Tail& tail = *reinterpret_cast<Tail*>(smem + STAGES * STAGE_BYTES);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // known warp-uniform (MMA descriptors stay in uniform registers)
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
  const int kbeg = blockIdx.z * k_per_split, kend = min(K, kbeg + k_per_split);

  if (lengths) {          // output rows are (clip, frame) pairs: tiles made only of padded frames are never read downstream
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
  float a_scale = 1.f, c_scale = 1.f;                   // power of two that lifts max|A| into [2048, 4096), and its inverse
  if (a_amax) {
    const float mx = __ldg(a_amax);
    if (mx > 0.f && mx < 3.0e38f) {
      int e;
      frexpf(mx, &e);
      e = max(-100, min(100, e));
      a_scale = ldexpf(1.f, 12 - e);
      c_scale = ldexpf(1.f, e - 12);
    }
  }

  const float* a_base = kAmn ? A + m0 : A + (int64_t)m0 * lda;
  const float* b_base = kBmn ? Bm + n0 : Bm + (int64_t)n0 * ldb;
  const int chunks = (kend - kbeg + BK - 1) / BK;
  Loader<BM, kAmn> la;
  Loader<BN, kBmn> lb;
  bool first = true;
  for (int c = 0; c < chunks; ++c) {
    const int s = c % STAGES;
    uint8_t* st = smem + s * STAGE_BYTES;
    uint8_t *a_hi = st, *a_lo = st + A_BYTES, *b_hi = st + 2 * A_BYTES, *b_lo = st + 2 * A_BYTES + B_BYTES;
    const int k0 = kbeg + c * BK, k_valid = min(BK, kend - k0);
    la.load(kAmn ? a_base + (int64_t)k0 * lda : a_base + k0, lda, M - m0, k_valid, tid);
    lb.load(kBmn ? b_base + (int64_t)k0 * ldb : b_base + k0, ldb, N - n0, k_valid, tid);
    if (c >= STAGES) tc::mbar_wait(&tail.bar[s], ((c / STAGES) - 1) & 1);     // MMAs that read this stage are done
    la.store(a_hi, a_lo, a_scale, tid);
    lb.store(b_hi, b_lo, 1.f, tid);
    tc::fence_async_smem();
    __syncthreads();
    if (warp == 0 && tc::elect_one()) {
      tc::fence_after_sync();
      const int ksteps = (k_valid + 15) / 16;
      for (int j = 0; j < ksteps; ++j) {
        const uint64_t dah = tc::make_desc(tc::smem_u32(a_hi) + j * 2 * LBO, LBO, SBO);
        const uint64_t dbh = tc::make_desc(tc::smem_u32(b_hi) + j * 2 * LBO, LBO, SBO);
        const uint64_t dal = tc::make_desc(tc::smem_u32(a_lo) + j * 2 * LBO, LBO, SBO);
        const uint64_t dbl = tc::make_desc(tc::smem_u32(b_lo) + j * 2 * LBO, LBO, SBO);
        tc::mma_f16(tmem, dah, dbh, idesc, !first);
        tc::mma_f16(tmem, dah, dbl, idesc, true);
        tc::mma_f16(tmem, dal, dbh, idesc, true);
        first = false;
      }
      tc::mma_commit(&tail.bar[s]);
    }
  }
  // ---- epilogue: wait for the last commit (commits complete in order), TMEM -> regs -> smem transpose -> C
  if (chunks > 0) {
    const int c = chunks - 1;
    tc::mbar_wait(&tail.bar[c % STAGES], (c / STAGES) & 1);
    tc::fence_after_sync();
    float* tile = reinterpret_cast<float*>(smem) + warp * (32 * 33);      // operand stages are free now
    const int row0 = m0 + (warp & 3) * 32;
    const int jbeg = (warp >> 2) * 3;
#pragma unroll 1
    for (int j = jbeg; j < jbeg + 3; ++j) {
      if (n0 + j * 32 >= N) break;                                         // warp-uniform
      float v[32];
      tc::tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + j * 32, v);
      tc::tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 32; ++q) tile[lane * 33 + q] = v[q];
      __syncwarp();
      const int n = n0 + j * 32 + lane;
      if (n < N) {
#pragma unroll 4
        for (int r = 0; r < 32; ++r) {
          if (row0 + r < M) {
            float* cp = C + (int64_t)(row0 + r) * ldc + n;
            const float o = tile[r * 33 + lane] * c_scale;
            if (atomic_out) atomicAdd(cp, o);
            else *cp = accumulate ? (*cp + o) : o;
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


// max |x| over a tensor (non-negative floats order like their bit patterns, so one integer atomicMax combines blocks)
__global__ void __launch_bounds__(256) absmax_kernel(const float* __restrict__ x, int64_t n, float* __restrict__ out) {
  float m = 0.f;
  const int64_t n4 = n >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(x4 + i);
    m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) m = fmaxf(m, fabsf(__ldg(x + (n4 << 2) + threadIdx.x)));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<int*>(out), __float_as_int(m));
}

}  // namespace gemmx
}  // namespace ssb

// out[0] = max |x[i]| (0 for an empty tensor; NaNs are ignored).  x must be 16-byte aligned.
extern "C" int ssb_absmax(const float* x, int64_t n, float* out, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(out && (x || n == 0) && n >= 0, "bad arguments");
  SSB_REQUIRE(((uintptr_t)x & 15) == 0, "x must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SSB_CUDA(cudaMemsetAsync(out, 0, sizeof(float), st));
  if (n == 0) return SSB_OK;
  const int blocks = (int)std::min<int64_t>(4 * kSMs, (n / 4 + 255) / 256 + 1);
  gemmx::absmax_kernel<<<blocks, 256, 0, st>>>(x, n, out);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// Strided tensor-core GEMM (split-fp16 x3, fp32 accumulate); same argument meaning as ssb_sgemm (no bias).
// a_amax (nullable, device): max|A| as written by ssb_absmax -- required whenever A holds values outside fp16's range.
//   C[i,j] (+)= sum_k A[i*sai + k*sak] * B[k*sbk + j*sbj]
// Each operand must be contiguous along k (then its other stride and K must be multiples of 4 floats and its base
// 16-byte aligned) or along its output index.  split_k > 1 accumulates into C with atomics (accumulate must be 1).
extern "C" int ssb_gemm_tcx(const float* A, int64_t sai, int64_t sak, const float* Bm, int64_t sbk, int64_t sbj,
                            float* C, int64_t ldc, int64_t M, int64_t N, int64_t K,
                            int32_t accumulate, int32_t split_k, const int64_t* lengths, int64_t T,
                            const float* a_amax, void* stream) {
  using namespace ssb;
  using namespace ssb::gemmx;
  SSB_REQUIRE(A && Bm && C, "null pointer");
  SSB_REQUIRE(M >= 0 && N >= 0 && K >= 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "bad shape");
  SSB_REQUIRE(sak == 1 || sai == 1, "A must be contiguous along i or k");
  SSB_REQUIRE(sbk == 1 || sbj == 1, "B must be contiguous along k or j");
  SSB_REQUIRE(split_k >= 1, "split_k >= 1");
  SSB_REQUIRE(!lengths || T > 0, "row mask needs T");
  const bool amn = (sak != 1), bmn = (sbk != 1);       // a unit stride along k wins when both are 1
  const int64_t lda = amn ? sak : sai, ldb = bmn ? sbk : sbj;
  if (!amn) SSB_REQUIRE(lda % 4 == 0 && K % 4 == 0 && ((uintptr_t)A & 15) == 0, "k-contiguous A: stride, K multiples of 4 and 16-byte aligned");
  if (!bmn) SSB_REQUIRE(ldb % 4 == 0 && K % 4 == 0 && ((uintptr_t)Bm & 15) == 0, "k-contiguous B: stride, K multiples of 4 and 16-byte aligned");
  if (M == 0 || N == 0) return SSB_OK;
  if (split_k > 1) SSB_REQUIRE(accumulate, "split_k > 1 accumulates into C");
  int k_per_split = (int)((K + split_k - 1) / split_k);
  k_per_split = ((k_per_split + BK - 1) / BK) * BK;
  if (k_per_split == 0) k_per_split = BK;
  const int splits = K == 0 ? 1 : (int)((K + k_per_split - 1) / k_per_split);
  if (K == 0) {                                          // empty contraction: C = 0 unless accumulating
    if (!accumulate) SSB_CUDA(cudaMemset2DAsync(C, ldc * 4, 0, N * 4, M, (cudaStream_t)stream));
    return SSB_OK;
  }
  const size_t smem = STAGES * STAGE_BYTES + sizeof(Tail);
  dim3 grid(cdiv(N, BN), cdiv(M, BM), splits), block(THREADS);
  SSB_REQUIRE(grid.y <= 65535u && grid.z <= 65535u, "M or split_k too large");
  const int atomic_out = splits > 1;
  cudaStream_t st = (cudaStream_t)stream;
#define SSB_TCX_LAUNCH(AMN, BMN)                                                                                   \
  do {                                                                                                             \
    SSB_CUDA(cudaFuncSetAttribute(gemm_tcx_kernel<AMN, BMN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    gemm_tcx_kernel<AMN, BMN><<<grid, block, smem, st>>>(A, lda, Bm, ldb, C, ldc, (int)M, (int)N, (int)K, k_per_split, \
                                                         accumulate, atomic_out, lengths, (int)T, a_amax);        \
  } while (0)
  if (amn && bmn) SSB_TCX_LAUNCH(true, true);
  else if (amn && !bmn) SSB_TCX_LAUNCH(true, false);
  else if (!amn && bmn) SSB_TCX_LAUNCH(false, true);
  else SSB_TCX_LAUNCH(false, false);
#undef SSB_TCX_LAUNCH
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
