// Shared pieces of the fused mouth-ROI CNN kernels (forward and backward).
//
// The network (train_model_official.py:209-229) on one 48x96 grayscale frame:
//   conv(1->8,3x3,p1)+ReLU -> maxpool2 -> conv(8->16)+ReLU -> maxpool2
//   -> conv(16->24)+ReLU -> mean over 12x24 -> Linear(24->32)
// preceded by /255 and the per-frame standardise of train_model_official.py:286-291.
//
// One CTA of 288 threads (9 warps) keeps a whole frame's activations in shared
// memory; nothing but the 4,608 input bytes and the 32 output floats touches HBM.
// 9 warps because every stage's work splits into a multiple of 9 warp-sized
// items (1152 / 288 / 288 positions).
#pragma once
#include "common.cuh"

namespace ssb {
namespace cnn {

constexpr int kThreads = 288;
constexpr int H0 = 48, W0 = 96;              // input
constexpr int H1 = 24, W1 = 48, C1 = 8;      // after conv1+pool
constexpr int H2 = 12, W2 = 24, C2 = 16;     // after conv2+pool
constexpr int C3 = 24, E = 32;               // conv3 channels, embedding
constexpr int NPIX = H0 * W0;

// padded shared-memory planes (1-pixel zero border; even row strides for float2)
constexpr int IN_LD = 100, IN_ROWS = H0 + 2;         // [50][100]
constexpr int A1_LD = 52, A1_ROWS = H1 + 2;          // [8][26][52]
constexpr int A2_LD = 28, A2_ROWS = H2 + 2;          // [16][14][28]
// plane strides: a1 = 1352+4 (== 12 mod 32), a2 = 392+26 (== 2 mod 32) so that threads that differ only
// in the input channel hit different banks in the weight-gradient passes of the backward kernel
constexpr int A1_PLANE = A1_ROWS * A1_LD + 4, A2_PLANE = A2_ROWS * A2_LD + 26;
constexpr int A3_LD = A2_LD, A3_PLANE = A2_ROWS * A2_LD;      // conv3 output kept padded [24][14][28] when saved

struct Weights {            // device pointers in the reference's state-dict layout
  const float *w1, *b1;     // roi_cnn.net.0  [8,1,3,3],  [8]
  const float *w2, *b2;     // roi_cnn.net.3  [16,8,3,3], [16]
  const float *w3, *b3;     // roi_cnn.net.6  [24,16,3,3],[24]
  const float *wfc, *bfc;   // roi_cnn.fc     [32,24],    [32]
};

// shared-memory weight copies, re-laid so that the output channels a thread
// accumulates together are contiguous (vector broadcast loads)
struct alignas(16) SmemWeights {
  float w1[9][C1];            // [tap][co]
  float b1[C1];
  float w2[C1][9][C2];        // [ci][tap][co]
  float b2[C2];
  float w3[C2][9][C3];        // [ci][tap][co]
  float b3[C3];
  float wfc[E][C3];
  float bfc[E];
};

__device__ __forceinline__ void load_weights(SmemWeights& s, const Weights& g, int tid) {
  for (int i = tid; i < C1 * 9; i += kThreads) { int co = i / 9, tap = i % 9; s.w1[tap][co] = g.w1[i]; }
  for (int i = tid; i < C2 * C1 * 9; i += kThreads) {
    int co = i / (C1 * 9), ci = (i / 9) % C1, tap = i % 9; s.w2[ci][tap][co] = g.w2[i];
  }
  for (int i = tid; i < C3 * C2 * 9; i += kThreads) {
    int co = i / (C2 * 9), ci = (i / 9) % C2, tap = i % 9; s.w3[ci][tap][co] = g.w3[i];
  }
  for (int i = tid; i < E * C3; i += kThreads) s.wfc[i / C3][i % C3] = g.wfc[i];
  if (tid < C1) s.b1[tid] = g.b1[tid];
  if (tid < C2) s.b2[tid] = g.b2[tid];
  if (tid < C3) s.b3[tid] = g.b3[tid];
  if (tid < E) s.bfc[tid] = g.bfc[tid];
}

// block-wide sum over kThreads threads; red must hold >= 9 floats; result to all threads
__device__ __forceinline__ float block_sum(float v, float* red, int tid) {
  v = warp_sum(v);
  __syncthreads();
  if ((tid & 31) == 0) red[tid >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < kThreads / 32; ++i) t += red[i];
  return t;
}

// u8 frame -> standardised float plane in padded smem (interior only; the border is
// zeroed once by the caller).  Returns (mean, 1/std) for the backward pass's benefit.
__device__ __forceinline__ void load_frame(const uint8_t* __restrict__ src, float* s_in, float* red,
                                           int tid, bool standardize) {
  // 4608 bytes = 1152 words; thread i takes words i, i+288, i+576, i+864
  float v[16];
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(src) + tid + j * kThreads);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[4 * j + k] = (float)((w >> (8 * k)) & 0xffu) / 255.0f;        // R.float()/255.0
      sum += v[4 * j + k];
    }
  }
  float mu = 0.f;
  if (standardize) {
    mu = block_sum(sum, red, tid) / (float)NPIX;
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) { float d = v[k] - mu; sq += d * d; }
    const float var = block_sum(sq, red, tid) / (float)(NPIX - 1);     // unbiased
    const float sd = fmaxf(sqrtf(var), 1e-6f);                         // clamp_min(1e-6)
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = (v[k] - mu) / sd;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int word = tid + j * kThreads;          // 24 words per image row
    const int y = word / (W0 / 4), x = (word % (W0 / 4)) * 4;
    float* d = s_in + (y + 1) * IN_LD + x + 1;
    d[0] = v[4 * j]; d[1] = v[4 * j + 1]; d[2] = v[4 * j + 2]; d[3] = v[4 * j + 3];
  }
}


// ---- stage 1: conv(1->8)+ReLU+maxpool2.  One thread = one pooled position, all 8 channels.
// kSave also records which of the 2x2 window won (first maximum in row-major order, the
// element ATen's max-pool backward routes the gradient to).
template <bool kSave>
__device__ __forceinline__ void conv1_pool(const float* __restrict__ s_in, float* __restrict__ s_a1,
                                           const SmemWeights& w, uint8_t* __restrict__ idx1, int tid) {
  for (int item = tid; item < H1 * W1; item += kThreads) {
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
    for (int co = 0; co < C1; ++co) { acc[co][0] = acc[co][1] = acc[co][2] = acc[co][3] = 0.f; }
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int dy = tap / 3, dx = tap % 3;
      float wv[C1];
      *reinterpret_cast<float4*>(wv) = *reinterpret_cast<const float4*>(&w.w1[tap][0]);
      *reinterpret_cast<float4*>(wv + 4) = *reinterpret_cast<const float4*>(&w.w1[tap][4]);
#pragma unroll
      for (int co = 0; co < C1; ++co) {
        acc[co][0] = fmaf(wv[co], P[dy][dx], acc[co][0]);
        acc[co][1] = fmaf(wv[co], P[dy][dx + 1], acc[co][1]);
        acc[co][2] = fmaf(wv[co], P[dy + 1][dx], acc[co][2]);
        acc[co][3] = fmaf(wv[co], P[dy + 1][dx + 1], acc[co][3]);
      }
    }
#pragma unroll
    for (int co = 0; co < C1; ++co) {
      float m = acc[co][0]; int arg = 0;
#pragma unroll
      for (int k = 1; k < 4; ++k) if (acc[co][k] > m) { m = acc[co][k]; arg = k; }
      s_a1[co * A1_PLANE + (py + 1) * A1_LD + px + 1] = fmaxf(m + w.b1[co], 0.f);
      if (kSave) idx1[co * (H1 * W1) + item] = (uint8_t)arg;
    }
  }
}

// ---- stage 2: conv(8->16)+ReLU+maxpool2.  One thread = one pooled position x 8 channels
// (two rounds over the channel halves; the half is uniform across the block -> broadcasts).
template <bool kSave>
__device__ __forceinline__ void conv2_pool(const float* __restrict__ s_a1, float* __restrict__ s_a2,
                                           const SmemWeights& w, uint8_t* __restrict__ idx2, int tid) {
  const int py = tid / W2, px = tid % W2;
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    float acc[8][4];
#pragma unroll
    for (int co = 0; co < 8; ++co) { acc[co][0] = acc[co][1] = acc[co][2] = acc[co][3] = 0.f; }
#pragma unroll 1
    for (int ci = 0; ci < C1; ++ci) {
      float P[4][4];
      const float* base = s_a1 + ci * A1_PLANE + (2 * py) * A1_LD + 2 * px;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float2 a = *reinterpret_cast<const float2*>(base + r * A1_LD);
        const float2 b = *reinterpret_cast<const float2*>(base + r * A1_LD + 2);
        P[r][0] = a.x; P[r][1] = a.y; P[r][2] = b.x; P[r][3] = b.y;
      }
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int dy = tap / 3, dx = tap % 3;
        float wv[8];
        *reinterpret_cast<float4*>(wv) = *reinterpret_cast<const float4*>(&w.w2[ci][tap][half * 8]);
        *reinterpret_cast<float4*>(wv + 4) = *reinterpret_cast<const float4*>(&w.w2[ci][tap][half * 8 + 4]);
#pragma unroll
        for (int co = 0; co < 8; ++co) {
          acc[co][0] = fmaf(wv[co], P[dy][dx], acc[co][0]);
          acc[co][1] = fmaf(wv[co], P[dy][dx + 1], acc[co][1]);
          acc[co][2] = fmaf(wv[co], P[dy + 1][dx], acc[co][2]);
          acc[co][3] = fmaf(wv[co], P[dy + 1][dx + 1], acc[co][3]);
        }
      }
    }
#pragma unroll
    for (int co = 0; co < 8; ++co) {
      const int c = half * 8 + co;
      float m = acc[co][0]; int arg = 0;
#pragma unroll
      for (int k = 1; k < 4; ++k) if (acc[co][k] > m) { m = acc[co][k]; arg = k; }
      s_a2[c * A2_PLANE + (py + 1) * A2_LD + px + 1] = fmaxf(m + w.b2[c], 0.f);
      if (kSave) idx2[c * (H2 * W2) + tid] = (uint8_t)arg;
    }
  }
}

// ---- stage 3: conv(16->24)+ReLU, mean over the 288 positions, Linear(24->32).
// One thread = a vertical pair of positions x 12 channels.  s_part: >= 18*12 floats,
// s_mean: >= 24 floats.  kSave keeps the ReLU output a3 (padded [24][14][28]) for the backward pass.
template <bool kSave>
__device__ __forceinline__ void conv3_mean(const float* __restrict__ s_a2, const SmemWeights& w,
                                           float* __restrict__ s_part, float* __restrict__ s_mean,
                                           float* __restrict__ a3, int tid) {
  const int q = tid % 144, half = tid / 144;
  const int y = 2 * (q / W2), x = q % W2;
  float acc[12][2];
#pragma unroll
  for (int co = 0; co < 12; ++co) acc[co][0] = acc[co][1] = 0.f;
#pragma unroll 1
  for (int ci = 0; ci < C2; ++ci) {
    float Q[4][3];
    const float* base = s_a2 + ci * A2_PLANE + y * A2_LD + x;
#pragma unroll
    for (int r = 0; r < 4; ++r) { Q[r][0] = base[r * A2_LD]; Q[r][1] = base[r * A2_LD + 1]; Q[r][2] = base[r * A2_LD + 2]; }
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int dy = tap / 3, dx = tap % 3;
      float wv[12];
      const float4* wp = reinterpret_cast<const float4*>(&w.w3[ci][tap][half * 12]);
      *reinterpret_cast<float4*>(wv) = wp[0];
      *reinterpret_cast<float4*>(wv + 4) = wp[1];
      *reinterpret_cast<float4*>(wv + 8) = wp[2];
#pragma unroll
      for (int co = 0; co < 12; ++co) {
        acc[co][0] = fmaf(wv[co], Q[dy][dx], acc[co][0]);
        acc[co][1] = fmaf(wv[co], Q[dy + 1][dx], acc[co][1]);
      }
    }
  }
  float s[12];
#pragma unroll
  for (int co = 0; co < 12; ++co) {
    const int c = half * 12 + co;
    const float r0 = fmaxf(acc[co][0] + w.b3[c], 0.f), r1 = fmaxf(acc[co][1] + w.b3[c], 0.f);
    if (kSave) { a3[c * A3_PLANE + (y + 1) * A3_LD + x + 1] = r0; a3[c * A3_PLANE + (y + 2) * A3_LD + x + 1] = r1; }
    float v = r0 + r1;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);     // 16-lane segments
    s[co] = v;
  }
  if ((tid & 15) == 0) {
#pragma unroll
    for (int co = 0; co < 12; ++co) s_part[(tid >> 4) * 12 + co] = s[co];
  }
  __syncthreads();
  if (tid < C3) {
    const int h = tid / 12, co = tid % 12;
    float t = 0.f;
#pragma unroll
    for (int seg = 0; seg < 9; ++seg) t += s_part[(h * 9 + seg) * 12 + co];
    s_mean[tid] = t / (float)(H2 * W2);
  }
  __syncthreads();
}

__device__ __forceinline__ float fc_out(const SmemWeights& w, const float* s_mean, int e) {
  float t = w.bfc[e];
#pragma unroll
  for (int c = 0; c < C3; ++c) t = fmaf(w.wfc[e][c], s_mean[c], t);
  return t;
}

}  // namespace cnn
}  // namespace ssb
