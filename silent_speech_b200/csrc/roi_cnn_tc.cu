// K2 on the tensor cores -- fused mouth-ROI CNN forward with conv2 and conv3 as implicit GEMMs on
// tcgen05 (87 % of the network's FLOPs), everything else of the frame still fused around them.
//
// Same arithmetic contract as roi_cnn_fwd.cu (train_model_official.py:286-294, 209-229).  Per frame:
//   P0  u8 load, /255, per-frame standardise                          -> fp32 plane in smem
//   P1  conv1(1->8)+ReLU+pool2 on CUDA cores (K = 9: not a dense contraction) -> a1 as bf16, PIXEL-MAJOR
//       (8 channels = 16 bytes per pixel), zero border
//   P2  conv2(8->16) as implicit GEMM: NO im2col copy.  An output tile is 16 image rows x 8 columns
//       (UMMA M = 128); for tap (dy,dx) the A operand is the a1 plane itself, addressed through the
//       shared-memory descriptor: core-matrix rows = 8 consecutive pixels (16 B apart), 8-row groups =
//       image rows (SBO = row pitch), and the two K-halves of a K=16 MMA are TWO TAPS (LBO = distance
//       between the taps' base addresses).  9 taps -> 5 MMAs per tile, fp32 accumulators in TMEM.
//   P3  epilogue: TMEM -> registers, +bias, ReLU, 2x2 max-pool by warp shuffles (a warp holds 4 image
//       rows x 8 columns), pooled pixels -> a2 as bf16 pixel-major in two 8-channel planes
//   P4  conv3(16->24): same trick, K=16 per tap = the two channel planes (LBO = plane stride), 9 MMAs/tile
//   P5  epilogue: +bias, ReLU, mean over the 288 positions, Linear(24->32)
// kSplit: a = a_hi + a_lo, w = w_hi + w_lo in bf16; w_hi and w_lo are stacked along N so that
// a_hi.[w_hi|w_lo] is one MMA and a_lo.w_hi a second -> fp32-parity grade at 2 MMAs per K-step.
#include "common.cuh"
#include "tc.cuh"

namespace ssb {
namespace cnntc {

constexpr int THREADS = 256;
constexpr int H0 = 48, W0 = 96, NPIX = H0 * W0;
constexpr int H1 = 24, W1 = 48, C1 = 8;
constexpr int H2 = 12, W2 = 24, C2 = 16;
constexpr int C3 = 24, C3P = 32, E = 32;
constexpr int IN_LD = 100, IN_ROWS = 50;
constexpr int PITCH1 = 50, ROWS1 = 26;                 // a1: [26][50] pixels x 16 B
constexpr int PITCH2 = 26, ROWS2 = 18;                 // a2: 2 planes x [18][26] pixels x 16 B
constexpr int A1_BYTES = ROWS1 * PITCH1 * 16;          // 20,800
constexpr int PLANE2 = ROWS2 * PITCH2 * 16;            // 7,488
constexpr int A2_BYTES = 2 * PLANE2;                   // 14,976
constexpr int IN_BYTES = IN_ROWS * IN_LD * 4;          // 20,000
template <bool kSplit> constexpr int kTmemCols = kSplit ? 256 : 128;   // x1: 6x16 / 3x32 = 96 columns in use

template <bool kSplit>
struct Layout {
  static constexpr int P = kSplit ? 2 : 1;
  static constexpr int W2_ROWS = kSplit ? 32 : 16;     // hi rows then lo rows (stacked along N)
  static constexpr int W3_ROWS = kSplit ? 64 : 32;
  static constexpr int W2_CHUNK = W2_ROWS * 16, W3_CHUNK = W3_ROWS * 16;
  static constexpr int W2_BYTES = 10 * W2_CHUNK, W3_BYTES = 18 * W3_CHUNK;
  // region R0 holds the fp32 input plane during P0/P1 and a2 (hi, lo) afterwards
  static constexpr int R0_BYTES = (P * A2_BYTES > IN_BYTES) ? P * A2_BYTES : IN_BYTES;
  static constexpr int OFF_A1 = R0_BYTES;
  static constexpr int OFF_W2 = OFF_A1 + P * A1_BYTES;
  static constexpr int OFF_W3 = OFF_W2 + W2_BYTES;
  static constexpr int OFF_MISC = OFF_W3 + W3_BYTES;
};

struct Misc {
  float w1[9][C1];
  float b1[C1], b2[C2], b3[C3P];
  float wfc[E][C3];
  float bfc[E];
  float sum[C3P];
  float red[16];
  uint64_t bar;
  uint32_t tmem_base;
};

struct Weights {
  const float *w1, *b1, *w2, *b2, *w3, *b3, *wfc, *bfc;
};

__device__ __forceinline__ float block_sum(float v, float* red, int tid) {
  v = warp_sum(v);
  __syncthreads();
  if ((tid & 31) == 0) red[tid >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < THREADS / 32; ++i) t += red[i];
  return t;
}

__device__ __forceinline__ uint4 pack8(const float* v) {
  return make_uint4(tc::pack_bf16(v[0], v[1]), tc::pack_bf16(v[2], v[3]), tc::pack_bf16(v[4], v[5]), tc::pack_bf16(v[6], v[7]));
}

// conv weights -> bf16 B operands.  Chunk j (8 k-values = one tap [x one channel plane]) holds ROWS rows x 16 B;
// rows [0, n_real) = hi of output channels, rows [half, half + n_real) = lo (kSplit), other rows zero.
template <bool kSplit>
__device__ __forceinline__ void build_weight_operands(uint8_t* smem, const Weights& g, int tid) {
  using L = Layout<kSplit>;
  uint8_t* w2op = smem + L::OFF_W2;
  uint8_t* w3op = smem + L::OFF_W3;
  for (int i = tid; i < (L::W2_BYTES + L::W3_BYTES) / 16; i += THREADS) reinterpret_cast<uint4*>(w2op)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  // conv2: [16,8,3,3]; chunk = tap (0..8), chunk 9 stays zero (the dummy tap that pads K to 80)
  for (int i = tid; i < C2 * 9; i += THREADS) {
    const int co = i / 9, tap = i % 9;
    float hi[8], lo[8];
#pragma unroll
    for (int ci = 0; ci < 8; ++ci) tc::split_bf16(g.w2[(co * C1 + ci) * 9 + tap], hi[ci], lo[ci]);
    uint8_t* row = w2op + tap * L::W2_CHUNK + (co >> 3) * 128 + (co & 7) * 16;
    *reinterpret_cast<uint4*>(row) = pack8(hi);
    if (kSplit) *reinterpret_cast<uint4*>(row + 16 * 16) = pack8(lo);          // row co + 16
  }
  // conv3: [24,16,3,3]; chunk = tap*2 + plane (channels 8*plane..+7)
  for (int i = tid; i < C3 * 18; i += THREADS) {
    const int co = i / 18, ch = i % 18, tap = ch >> 1, pl = ch & 1;
    float hi[8], lo[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) tc::split_bf16(g.w3[(co * C2 + pl * 8 + c) * 9 + tap], hi[c], lo[c]);
    uint8_t* row = w3op + ch * L::W3_CHUNK + (co >> 3) * 128 + (co & 7) * 16;
    *reinterpret_cast<uint4*>(row) = pack8(hi);
    if (kSplit) *reinterpret_cast<uint4*>(row + 32 * 16) = pack8(lo);          // row co + 32
  }
}


template <bool kSplit>
__global__ void __launch_bounds__(THREADS, kSplit ? 2 : 3)
roi_cnn_tc_kernel(const uint8_t* __restrict__ R, const int64_t* __restrict__ lengths, int B, int T,
                  Weights gw, int standardize, float* __restrict__ out, int64_t out_stride) {
  using L = Layout<kSplit>;
  extern __shared__ __align__(1024) uint8_t smem[];
  float* s_in = reinterpret_cast<float*>(smem);                 // region R0, P0/P1
  uint8_t* a2 = smem;                                           // region R0, P3..P5 (hi, then lo at +A2_BYTES)
  uint8_t* a1 = smem + L::OFF_A1;                               // hi, then lo at +A1_BYTES
  uint8_t* w2op = smem + L::OFF_W2;
  uint8_t* w3op = smem + L::OFF_W3;
  Misc& mz = *reinterpret_cast<Misc*>(smem + L::OFF_MISC);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- one-time setup: weights, zero planes, TMEM, barrier
  build_weight_operands<kSplit>(smem, gw, tid);
  for (int i = tid; i < C1 * 9; i += THREADS) mz.w1[i % 9][i / 9] = gw.w1[i];
  for (int i = tid; i < E * C3; i += THREADS) mz.wfc[i / C3][i % C3] = gw.wfc[i];
  if (tid < C1) mz.b1[tid] = gw.b1[tid];
  if (tid < C2) mz.b2[tid] = gw.b2[tid];
  if (tid < C3P) mz.b3[tid] = tid < C3 ? gw.b3[tid] : 0.f;
  if (tid < E) mz.bfc[tid] = gw.bfc[tid];
  for (int i = tid; i < L::P * A1_BYTES / 16; i += THREADS) reinterpret_cast<uint4*>(a1)[i] = make_uint4(0, 0, 0, 0);
  if (warp == 0) tc::tmem_alloc<kTmemCols<kSplit>>(&mz.tmem_base);
  if (tid == 32) { tc::mbar_init(&mz.bar, 1); tc::mbar_fence_init(); }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = mz.tmem_base;
  const uint32_t a1_u = tc::smem_u32(a1), a2_u = tc::smem_u32(a2), w2_u = tc::smem_u32(w2op), w3_u = tc::smem_u32(w3op);
  constexpr uint32_t idesc2_hi = tc::make_idesc_bf16(128, kSplit ? 32 : 16), idesc2_lo = tc::make_idesc_bf16(128, 16);
  constexpr uint32_t idesc3_hi = tc::make_idesc_bf16(128, kSplit ? 64 : 32), idesc3_lo = tc::make_idesc_bf16(128, 32);
  constexpr int COLS2 = kSplit ? 32 : 16, COLS3 = kSplit ? 64 : 32;
  uint32_t phase = 0;
  const int q = warp & 3, wg = warp >> 2;                       // TMEM lane quarter, warp group
  const int xo = lane & 7, rl = lane >> 3;                      // position of this lane inside a 4x8 patch of a tile

  const int64_t frames = (int64_t)B * T;
  for (int64_t f = blockIdx.x; f < frames; f += gridDim.x) {
    const int bidx = (int)(f / T), t = (int)(f % T);
    float* o = out + f * out_stride;
    if (lengths && t >= (int)lengths[bidx]) {
      if (tid < E) o[tid] = 0.f;
      continue;
    }
    // ================= P0: load + standardise (train_model_official.py:286-291)
    // mean and unbiased variance from EXACT integer sums of the u8 pixels (dp4a), one block reduction:
    //   (u/255 - mu)/max(sd, 1e-6)  ==  (u - S1/N) / max(sqrt((S2 - S1^2/N)/(N-1)), 255e-6)
    {
      uint32_t wv[5];
      uint32_t s1 = 0, s2 = 0;
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const int word = tid + j * THREADS;
        wv[j] = word < NPIX / 4 ? __ldg(reinterpret_cast<const uint32_t*>(R + f * NPIX) + word) : 0u;
        s1 = __dp4a(wv[j], 0x01010101u, s1);
        s2 = __dp4a(wv[j], wv[j], s2);
      }
      float mean_u = 0.f, inv_s = 1.0f / 255.0f;
      if (standardize) {
        s1 = __reduce_add_sync(0xffffffffu, s1);
        s2 = __reduce_add_sync(0xffffffffu, s2);
        uint32_t* redu = reinterpret_cast<uint32_t*>(mz.red);
        if (lane == 0) { redu[warp] = s1; redu[8 + warp] = s2; }
        __syncthreads();
        uint32_t t1 = 0, t2 = 0;
#pragma unroll
        for (int i = 0; i < THREADS / 32; ++i) { t1 += redu[i]; t2 += redu[8 + i]; }
        const double S1 = (double)t1, S2 = (double)t2;
        const double var_u = fmax(S2 - S1 * S1 / (double)NPIX, 0.0) / (double)(NPIX - 1);
        mean_u = (float)(S1 / (double)NPIX);
        inv_s = (float)(1.0 / fmax(sqrt(var_u), 255.0e-6));
      }
      // the border of the fp32 plane must read as zero padding: region R0 was a2 a moment ago
      for (int i = tid; i < IN_ROWS; i += THREADS) { s_in[i * IN_LD] = 0.f; s_in[i * IN_LD + W0 + 1] = 0.f; }
      for (int i = tid; i < IN_LD; i += THREADS) { s_in[i] = 0.f; s_in[(H0 + 1) * IN_LD + i] = 0.f; }
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const int word = tid + j * THREADS;
        if (word < NPIX / 4) {
          const int y = word / (W0 / 4), x = (word % (W0 / 4)) * 4;
          float* d = s_in + (y + 1) * IN_LD + x + 1;
#pragma unroll
          for (int k = 0; k < 4; ++k) d[k] = ((float)((wv[j] >> (8 * k)) & 0xffu) - mean_u) * inv_s;
        }
      }
    }
    if (tid < C3P) mz.sum[tid] = 0.f;
    __syncthreads();

    // ================= P1: conv1 + ReLU + pool on CUDA cores -> a1 (bf16, pixel-major, 16 B per pixel)
    for (int item = tid; item < H1 * W1; item += THREADS) {
      const int py = item / W1, px = item % W1;
      float P[4][4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float2 a = *reinterpret_cast<const float2*>(s_in + (2 * py + r) * IN_LD + 2 * px);
        const float2 b = *reinterpret_cast<const float2*>(s_in + (2 * py + r) * IN_LD + 2 * px + 2);
        P[r][0] = a.x; P[r][1] = a.y; P[r][2] = b.x; P[r][3] = b.y;
      }
      float acc[C1][4];
#pragma unroll
      for (int co = 0; co < C1; ++co) acc[co][0] = acc[co][1] = acc[co][2] = acc[co][3] = 0.f;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int dy = tap / 3, dx = tap % 3;
        float wv[C1];
        *reinterpret_cast<float4*>(wv) = *reinterpret_cast<const float4*>(&mz.w1[tap][0]);
        *reinterpret_cast<float4*>(wv + 4) = *reinterpret_cast<const float4*>(&mz.w1[tap][4]);
#pragma unroll
        for (int co = 0; co < C1; ++co) {
          acc[co][0] = fmaf(wv[co], P[dy][dx], acc[co][0]);
          acc[co][1] = fmaf(wv[co], P[dy][dx + 1], acc[co][1]);
          acc[co][2] = fmaf(wv[co], P[dy + 1][dx], acc[co][2]);
          acc[co][3] = fmaf(wv[co], P[dy + 1][dx + 1], acc[co][3]);
        }
      }
      float hi[C1], lo[C1];
#pragma unroll
      for (int co = 0; co < C1; ++co) {
        const float m4 = fmaxf(fmaxf(acc[co][0], acc[co][1]), fmaxf(acc[co][2], acc[co][3]));
        tc::split_bf16(fmaxf(m4 + mz.b1[co], 0.f), hi[co], lo[co]);
      }
      const uint32_t off = ((py + 1) * PITCH1 + px + 1) * 16;
      *reinterpret_cast<uint4*>(a1 + off) = pack8(hi);
      if (kSplit) *reinterpret_cast<uint4*>(a1 + A1_BYTES + off) = pack8(lo);
    }
    __syncthreads();                                            // s_in is dead from here on
    // region R0 becomes a2: clear it (its zero border is conv3's padding)
    for (int i = tid; i < L::P * A2_BYTES / 16; i += THREADS) reinterpret_cast<uint4*>(a2)[i] = make_uint4(0, 0, 0, 0);
    tc::fence_async_smem();
    __syncthreads();

    // ================= P2/P3: conv2 on the tensor cores, two rounds of 6 tiles (16 rows x 8 cols each)
#pragma unroll 1
    for (int ty = 0; ty < 2; ++ty) {
      const int y0 = ty * 8;                                    // rows 0..15, then 8..23 (rows 8..15 recomputed)
      if (tid == 0) {
        tc::fence_after_sync();
        for (int tx = 0; tx < 6; ++tx) {
          const uint32_t d = tmem + tx * COLS2;
#pragma unroll
          for (int p = 0; p < 5; ++p) {
            const int t0 = 2 * p, dy = t0 / 3, dx = t0 % 3;
            const uint32_t a_off = ((y0 + dy) * PITCH1 + tx * 8 + dx) * 16;
            // second K-half = the next tap: same row (+16 B) or the first tap of the next row
            const uint32_t lbo = (t0 % 3 == 2 && p < 4) ? (PITCH1 - 2) * 16 : 16;
            const uint64_t db = tc::make_desc(w2_u + 2 * p * L::W2_CHUNK, L::W2_CHUNK, 128);
            tc::mma_bf16(d, tc::make_desc(a1_u + a_off, lbo, PITCH1 * 16), db, idesc2_hi, p > 0);
            if (kSplit) tc::mma_bf16(d, tc::make_desc(a1_u + A1_BYTES + a_off, lbo, PITCH1 * 16), db, idesc2_lo, true);
          }
        }
        tc::mma_commit(&mz.bar);
      }
      tc::mbar_wait_bounded(&mz.bar, phase & 1);
      ++phase;
      tc::fence_after_sync();
      // epilogue: this warp's 4 rows x 8 cols patch of tiles wg, wg+2, wg+4; round 1 only adds rows 16..23
      if (ty == 0 || q >= 2) {
#pragma unroll 1
        for (int tx = wg; tx < 6; tx += 2) {
          float v[16];
          tc::tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + tx * COLS2, v);
          if (kSplit) {
            float w[16];
            tc::tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + tx * COLS2 + 16, w);
            tc::tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 16; ++c) v[c] += w[c];
          } else {
            tc::tmem_ld_wait();
          }
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            float x = fmaxf(v[c] + mz.b2[c], 0.f);
            x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
            x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 8));
            v[c] = x;
          }
          if ((xo & 1) == 0 && (rl & 1) == 0) {
            const int py = (y0 + 4 * q + rl) >> 1, px = (tx * 8 + xo) >> 1;
            const uint32_t off = ((py + 1) * PITCH2 + px + 1) * 16;
            float hi[16], lo[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) tc::split_bf16(v[c], hi[c], lo[c]);
            *reinterpret_cast<uint4*>(a2 + off) = pack8(hi);
            *reinterpret_cast<uint4*>(a2 + PLANE2 + off) = pack8(hi + 8);
            if (kSplit) {
              *reinterpret_cast<uint4*>(a2 + A2_BYTES + off) = pack8(lo);
              *reinterpret_cast<uint4*>(a2 + A2_BYTES + PLANE2 + off) = pack8(lo + 8);
            }
          }
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      __syncthreads();
    }

    // ================= P4/P5: conv3 on the tensor cores: 3 tiles (rows 0..15 of which 12 are real)
    if (tid == 0) {
      tc::fence_after_sync();
      for (int tx = 0; tx < 3; ++tx) {
        const uint32_t d = tmem + tx * COLS3;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int dy = tap / 3, dx = tap % 3;
          const uint32_t a_off = (dy * PITCH2 + tx * 8 + dx) * 16;
          const uint64_t db = tc::make_desc(w3_u + 2 * tap * L::W3_CHUNK, L::W3_CHUNK, 128);
          tc::mma_bf16(d, tc::make_desc(a2_u + a_off, PLANE2, PITCH2 * 16), db, idesc3_hi, tap > 0);
          if (kSplit) tc::mma_bf16(d, tc::make_desc(a2_u + A2_BYTES + a_off, PLANE2, PITCH2 * 16), db, idesc3_lo, true);
        }
      }
      tc::mma_commit(&mz.bar);
    }
    tc::mbar_wait_bounded(&mz.bar, phase & 1);
    ++phase;
    tc::fence_after_sync();
    if (q < 3) {                                                // rows 12..15 of the tiles are padding
#pragma unroll 1
      for (int tx = wg; tx < 3; tx += 2) {
        float v[32];
        tc::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + tx * COLS3, v);
        if (kSplit) {
          float w[32];
          tc::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + tx * COLS3 + 32, w);
          tc::tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < C3; ++c) v[c] += w[c];
        } else {
          tc::tmem_ld_wait();
        }
#pragma unroll
        for (int c = 0; c < C3P; ++c) v[c] = c < C3 ? fmaxf(v[c] + mz.b3[c], 0.f) : 0.f;
        // transpose-reduce: 31 shuffles leave the sum over the warp's 32 positions of channel `lane` in v[0]
#pragma unroll
        for (int off = 16, n = 16; off >= 1; off >>= 1, n >>= 1) {
          const bool up = (lane & off) != 0;
#pragma unroll
          for (int i = 0; i < n; ++i) {
            const float send = up ? v[i] : v[i + n], keep = up ? v[i + n] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
          }
        }
        if (lane < C3) atomicAdd(&mz.sum[lane], v[0]);
      }
    }
    tc::fence_before_sync();
    __syncthreads();
    // ================= P6: mean + Linear(24->32)
    if (tid < E) {
      float a = mz.bfc[tid];
#pragma unroll
      for (int c = 0; c < C3; ++c) a = fmaf(mz.wfc[tid][c], mz.sum[c] / (float)(H2 * W2), a);
      o[tid] = a;
    }
    __syncthreads();                                            // region R0 and mz.sum are reused by the next frame
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<kTmemCols<kSplit>>(tmem);
}

}  // namespace cnntc
}  // namespace ssb

// Tensor-core build of ssb_roi_cnn_fwd (same arguments); passes = 1 (bf16) or 3 (split-bf16, fp32-parity grade).
extern "C" int ssb_roi_cnn_fwd_tc(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                                  int32_t roi_h, int32_t roi_w, const float* const* weights, int32_t standardize,
                                  float* out, int64_t out_stride, int32_t passes, void* stream) {
  using namespace ssb;
  using namespace ssb::cnntc;
  SSB_REQUIRE(R && weights && out, "null pointer");
  SSB_REQUIRE(roi_h == H0 && roi_w == W0, "ROI must be 48x96 (train_model_official.py:38)");
  SSB_REQUIRE(out_stride >= E, "out_stride < 32");
  SSB_REQUIRE(passes == 1 || passes == 3, "passes must be 1 or 3");
  for (int i = 0; i < 8; ++i) SSB_REQUIRE(weights[i], "null weight pointer");
  SSB_REQUIRE(((uintptr_t)R & 3) == 0, "R must be 4-byte aligned");
  if (B * T == 0) return SSB_OK;
  Weights gw{weights[0], weights[1], weights[2], weights[3], weights[4], weights[5], weights[6], weights[7]};
  const int64_t frames = B * T;
  const int per_sm = passes == 3 ? 2 : 3;
  const int grid = (int)(frames < per_sm * kSMs ? frames : per_sm * kSMs);
  cudaStream_t st = (cudaStream_t)stream;
  if (passes == 3) {
    const size_t smem = Layout<true>::OFF_MISC + sizeof(Misc);
    SSB_CUDA(cudaFuncSetAttribute(roi_cnn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    roi_cnn_tc_kernel<true><<<grid, THREADS, smem, st>>>(R, lengths, (int)B, (int)T, gw, standardize, out, out_stride);
  } else {
    const size_t smem = Layout<false>::OFF_MISC + sizeof(Misc);
    SSB_CUDA(cudaFuncSetAttribute(roi_cnn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    roi_cnn_tc_kernel<false><<<grid, THREADS, smem, st>>>(R, lengths, (int)B, (int)T, gw, standardize, out, out_stride);
  }
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
