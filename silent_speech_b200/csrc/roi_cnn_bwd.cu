// K2-bwd (fp32 build) -- weight gradients of the fused mouth-ROI CNN.
//
// What loss.backward() (train_model_official.py:437) does through TinyROICNN
// (train_model_official.py:209-229) and the standardise of 289-291, per frame and entirely
// in shared memory: the forward pass is recomputed (keeping the two max-pool argmax maps and
// the conv3 ReLU output), then
//   B0  Linear(24->32) and the mean:      dWfc, dbfc, g3 = relu'(a3) * dmean/288, db3
//   B1  conv3 weight gradient             dW3[co,ci,tap] += sum_pos g3[co,pos] a2[ci,pos+tap]
//   B2  conv3 data gradient (gather)      da2, masked by relu'(a2)
//   B3  conv2 weight gradient through the pool argmax     dW2, db2   (one 4x4 patch load serves all 16 channels)
//   B4  conv2 data gradient (gather through the argmax)   da1, masked by relu'(a1)
//   B5  conv1 weight gradient through the pool argmax     dW1, db1
// The ROI pixels are inputs, not parameters, so nothing flows past conv1.  Every thread keeps
// its slice of the weight gradients in registers across all frames its CTA visits and adds
// them to the global gradient once, at the end.
#include "roi_cnn.cuh"

namespace ssb {
namespace cnn {

constexpr int A3_COL0 = 2;     // interior of the saved conv3 plane starts at column 2 (float2-aligned pairs)
// per-frame record written by the split-mode forward kernel (roi_cnn_tc3.cu: SV_*; ssb_roi_cnn_saved_bytes)
constexpr int SV_A1 = 0, SV_A2 = 36864, SV_IDX1 = 55296, SV_IDX2 = 57600, SV_MASK3 = 58752, SV_MEAN3 = 59904, SV_BYTES = 60000;
static_assert(SV_A2 == H1 * W1 * C1 * 4 && SV_IDX1 == SV_A2 + H2 * W2 * C2 * 4 && SV_IDX2 == SV_IDX1 + H1 * W1 * 2 &&
              SV_MASK3 == SV_IDX2 + H2 * W2 * 4 && SV_MEAN3 == SV_MASK3 + H2 * W2 * 4 && SV_BYTES == SV_MEAN3 + C3 * 4 + 0 &&
              kThreads == H2 * W2, "saved-activation record layout");

struct BwdSmem {
  SmemWeights w;
  float w2t[C2][9][C1];            // [co][tap][ci]: conv2 weights for the data gradient (B4)
  float dW2s[C2 * C1 * 9 + C2];    // conv2 weight and bias gradient of this CTA, summed over its frames
  float in[IN_ROWS * IN_LD];
  float a1[C1 * A1_PLANE];
  float a2[C2 * A2_PLANE];
  float g3[C3 * A3_PLANE];
  float da2m[C2 * H2 * W2];
  float da1m[C1 * H1 * W1];
  uint8_t idx1[C1 * H1 * W1];
  uint8_t idx2[C2 * H2 * W2];
  float part[18 * 12];
  float mean[C3];
  float red[16];
  float dE[E];
  float dmean[C3];
  int cnt[C3];
};

// conv3 forward that keeps the ReLU output in the padded, column-shifted plane used by B1/B2
__device__ __forceinline__ void conv3_save(const float* __restrict__ s_a2, const SmemWeights& w,
                                           float* __restrict__ s_part, float* __restrict__ s_mean,
                                           float* __restrict__ g3, int tid) {
  // a3 is written through conv3_mean<true> at [c][y+1][x+1]; shift the base by one column
  conv3_mean<true>(s_a2, w, s_part, s_mean, g3 + (A3_COL0 - 1), tid);
}

__global__ void __launch_bounds__(kThreads, 1)
roi_cnn_bwd_kernel(const uint8_t* __restrict__ R, const int64_t* __restrict__ lengths, int B, int T,
                   Weights gw, int standardize, const float* __restrict__ d_out, int64_t d_out_stride,
                   float* __restrict__ gW1, float* __restrict__ gb1, float* __restrict__ gW2, float* __restrict__ gb2,
                   float* __restrict__ gW3, float* __restrict__ gb3, float* __restrict__ gWfc, float* __restrict__ gbfc,
                   const uint8_t* __restrict__ saved) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BwdSmem& s = *reinterpret_cast<BwdSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31;
  load_weights(s.w, gw, tid);
  for (int i = tid; i < IN_ROWS * IN_LD; i += kThreads) s.in[i] = 0.f;
  for (int i = tid; i < C1 * A1_PLANE; i += kThreads) s.a1[i] = 0.f;
  for (int i = tid; i < C2 * A2_PLANE; i += kThreads) s.a2[i] = 0.f;
  for (int i = tid; i < C3 * A3_PLANE; i += kThreads) s.g3[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < C2 * 9 * C1; i += kThreads) {
    const int co = i / (9 * C1), tap = (i / C1) % 9, ci = i % C1;
    s.w2t[co][tap][ci] = s.w.w2[ci][tap][co];
  }
  for (int i = tid; i < C2 * C1 * 9 + C2; i += kThreads) s.dW2s[i] = 0.f;
  __syncthreads();

  // ---- per-thread slices of the weight gradients
  const int b1_third = tid / 96, b1_cog = (tid % 96) / 16, b1_ci = tid % 16;          // dW3
  float acc3[4][9];
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int t = 0; t < 9; ++t) acc3[k][t] = 0.f;
  const int b3_ci = lane & 7, b3_cog = lane >> 3, b3_w = tid >> 5;   // dW2: 4 output channels x 1 input channel; warp = 32 cells
  const int b5_co = tid / 36, b5_grp = tid % 36;                                      // dW1: 32 pooled cells of one channel
  float acc1[9], accb1 = 0.f, accb3 = 0.f, accfc[3] = {0.f, 0.f, 0.f};

#pragma unroll
  for (int t = 0; t < 9; ++t) acc1[t] = 0.f;
  const int64_t frames = (int64_t)B * T;
  for (int64_t f = blockIdx.x; f < frames; f += gridDim.x) {
    const int b = (int)(f / T), t = (int)(f % T);
    if (lengths && t >= (int)lengths[b]) continue;

    // ================= forward recompute
    load_frame(R + f * NPIX, s.in, s.red, tid, standardize != 0);
    __syncthreads();
    if (saved) {
      // the forward call kept what the recompute would produce (a third of this kernel's arithmetic): pooled activations,
      // pool argmax maps, the conv3 ReLU mask and the channel means -- 60 KB per frame, read once, coalesced
      const uint8_t* rec = saved + f * SV_BYTES;
#pragma unroll
      for (int j = 0; j < H1 * W1 / kThreads; ++j) {             // a1 + idx1: 4 pooled pixels per thread
        const int item = tid + j * kThreads, py = item / W1, px = item % W1;
        const float4 v0 = __ldg(reinterpret_cast<const float4*>(rec + SV_A1) + 2 * item);
        const float4 v1 = __ldg(reinterpret_cast<const float4*>(rec + SV_A1) + 2 * item + 1);
        const uint32_t am = __ldg(reinterpret_cast<const uint16_t*>(rec + SV_IDX1) + item);
        const float a[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int co = 0; co < C1; ++co) {
          s.a1[co * A1_PLANE + (py + 1) * A1_LD + px + 1] = a[co];
          s.idx1[co * (H1 * W1) + item] = (uint8_t)((am >> (2 * co)) & 3u);
        }
      }
      {                                                          // a2 + idx2 + conv3 mask: one pooled pixel per thread
        const int py = tid / W2, px = tid % W2;
        const uint32_t am = __ldg(reinterpret_cast<const uint32_t*>(rec + SV_IDX2) + tid);
        const uint32_t on = __ldg(reinterpret_cast<const uint32_t*>(rec + SV_MASK3) + tid);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(rec + SV_A2) + 4 * tid + q);
          const float a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = 4 * q + e;
            s.a2[c * A2_PLANE + (py + 1) * A2_LD + px + 1] = a[e];
            s.idx2[c * (H2 * W2) + tid] = (uint8_t)((am >> (2 * c)) & 3u);
          }
        }
        float* g = s.g3 + (py + 1) * A3_LD + px + A3_COL0;
#pragma unroll
        for (int c = 0; c < C3; ++c) g[c * A3_PLANE] = ((on >> c) & 1u) ? 1.f : 0.f;    // B0 only tests g > 0
      }
      if (tid < C3) s.mean[tid] = __ldg(reinterpret_cast<const float*>(rec + SV_MEAN3) + tid);
      __syncthreads();
    } else {
      conv1_pool<true>(s.in, s.a1, s.w, s.idx1, tid);
      __syncthreads();
      conv2_pool<true>(s.a1, s.a2, s.w, s.idx2, tid);
      __syncthreads();
      conv3_save(s.a2, s.w, s.part, s.mean, s.g3, tid);          // ends with a barrier; s.mean valid
    }

    // ================= B0: fc, mean, relu mask of conv3
    if (tid < E) s.dE[tid] = d_out[f * d_out_stride + tid];
    __syncthreads();
    if (tid < C3) {
      float a = 0.f;
#pragma unroll
      for (int e = 0; e < E; ++e) a = fmaf(s.w.wfc[e][tid], s.dE[e], a);
      s.dmean[tid] = a / (float)(H2 * W2);
      s.cnt[tid] = 0;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int i = tid + k * kThreads;
      if (i < E * C3) accfc[k] = fmaf(s.dE[i / C3], s.mean[i % C3], accfc[k]);
      else if (i < E * C3 + E) accfc[k] += s.dE[i - E * C3];
    }
    __syncthreads();
    {
      const int y = tid / W2, x = tid % W2;
      float* g = s.g3 + (y + 1) * A3_LD + x + A3_COL0;
#pragma unroll 4
      for (int c = 0; c < C3; ++c) {
        const bool on = g[c * A3_PLANE] > 0.f;
        g[c * A3_PLANE] = on ? s.dmean[c] : 0.f;
        const unsigned bal = __ballot_sync(0xffffffffu, on);
        if (lane == 0) atomicAdd(&s.cnt[c], __popc(bal));
      }
    }
    __syncthreads();
    if (tid < C3) accb3 = fmaf(s.dmean[tid], (float)s.cnt[tid], accb3);

    // ================= B1: dW3
    {
      const float* gbase = s.g3 + (b1_cog * 4) * A3_PLANE + A3_COL0;
      const float* abase = s.a2 + b1_ci * A2_PLANE;
#pragma unroll 1
      for (int yy = 0; yy < 4; ++yy) {
        const int y = b1_third * 4 + yy;
#pragma unroll 2
        for (int xp = 0; xp < W2 / 2; ++xp) {
          const int x = 2 * xp;
          float2 g[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) g[k] = *reinterpret_cast<const float2*>(gbase + k * A3_PLANE + (y + 1) * A3_LD + x);
          float P[3][4];
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const float2 p0 = *reinterpret_cast<const float2*>(abase + (y + r) * A2_LD + x);
            const float2 p1 = *reinterpret_cast<const float2*>(abase + (y + r) * A2_LD + x + 2);
            P[r][0] = p0.x; P[r][1] = p0.y; P[r][2] = p1.x; P[r][3] = p1.y;
          }
#pragma unroll
          for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int dy = 0; dy < 3; ++dy)
#pragma unroll
              for (int dx = 0; dx < 3; ++dx)
                acc3[k][dy * 3 + dx] = fmaf(g[k].y, P[dy][dx + 1], fmaf(g[k].x, P[dy][dx], acc3[k][dy * 3 + dx]));
        }
      }
    }

    // ================= B2: da2 = conv3^T(g3), masked by relu'(a2)
    {
      const int q = tid % 144, half = tid / 144;
      const int y = 2 * (q / W2), x = q % W2;
      float acc[8][2];
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[c][0] = acc[c][1] = 0.f;
#pragma unroll 1
      for (int tap = 0; tap < 9; ++tap) {
        const int dy = tap / 3, dx = tap % 3;
        // position (y,x) receives from conv3 output (y-dy+1, x-dx+1): padded row +1, col +A3_COL0
        const float* gp = s.g3 + (y - dy + 2) * A3_LD + (x - dx + 1 + A3_COL0);
        float g0[C3], g1[C3];
#pragma unroll
        for (int co = 0; co < C3; ++co) { g0[co] = gp[co * A3_PLANE]; g1[co] = gp[co * A3_PLANE + A3_LD]; }
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4* wp = reinterpret_cast<const float4*>(&s.w.w3[half * 8 + c][tap][0]);
          float a0 = acc[c][0], a1v = acc[c][1];
#pragma unroll
          for (int v = 0; v < C3 / 4; ++v) {
            const float4 wv = wp[v];
            a0 = fmaf(wv.x, g0[4 * v], a0); a0 = fmaf(wv.y, g0[4 * v + 1], a0);
            a0 = fmaf(wv.z, g0[4 * v + 2], a0); a0 = fmaf(wv.w, g0[4 * v + 3], a0);
            a1v = fmaf(wv.x, g1[4 * v], a1v); a1v = fmaf(wv.y, g1[4 * v + 1], a1v);
            a1v = fmaf(wv.z, g1[4 * v + 2], a1v); a1v = fmaf(wv.w, g1[4 * v + 3], a1v);
          }
          acc[c][0] = a0; acc[c][1] = a1v;
        }
      }
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int ci = half * 8 + c;
        const float v0 = s.a2[ci * A2_PLANE + (y + 1) * A2_LD + x + 1], v1 = s.a2[ci * A2_PLANE + (y + 2) * A2_LD + x + 1];
        s.da2m[ci * (H2 * W2) + y * W2 + x] = v0 > 0.f ? acc[c][0] : 0.f;
        s.da2m[ci * (H2 * W2) + (y + 1) * W2 + x] = v1 > 0.f ? acc[c][1] : 0.f;
      }
    }
    __syncthreads();

    // ================= B3: dW2, db2 (gradient sits at the pooled cell's argmax position)
    // The 4x4 patch of a1 under a pooled cell holds the 3x3 window of every one of its four argmax positions, and it is
    // the same for all 16 output channels: a lane loads it once for its input channel and picks the window of each of
    // its four output channels with selects on the argmax bits.
    {
      float acc2[4][9], accb2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int t = 0; t < 9; ++t) acc2[k][t] = 0.f;
      const float* abase = s.a1 + b3_ci * A1_PLANE;
#pragma unroll 1
      for (int i = 0; i < 32; ++i) {
        const int cell = b3_w * 32 + i;
        const int py = cell / W2, px = cell % W2;
        float P[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float2 a = *reinterpret_cast<const float2*>(abase + (2 * py + r) * A1_LD + 2 * px);
          const float2 b = *reinterpret_cast<const float2*>(abase + (2 * py + r) * A1_LD + 2 * px + 2);
          P[r][0] = a.x; P[r][1] = a.y; P[r][2] = b.x; P[r][3] = b.y;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int co = b3_cog * 4 + k;
          const float d = s.da2m[co * (H2 * W2) + cell];
          const int id = s.idx2[co * (H2 * W2) + cell];
          const bool dn = (id & 2) != 0, rt = (id & 1) != 0;
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            float R[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) R[c] = dn ? P[dy + 1][c] : P[dy][c];
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) acc2[k][dy * 3 + dx] = fmaf(d, rt ? R[dx + 1] : R[dx], acc2[k][dy * 3 + dx]);
          }
          accb2[k] += d;
        }
      }
      // the 9 warps' partial sums meet in the (now dead) a2/g3 region: slot [warp][cog*296 + k*72 + ci*9 + tap]
      // (296: the four channel groups of a warp land in different banks), biases behind the slots
      constexpr int SLOT = 4 * 296, NW2 = C2 * C1 * 9;
      float* red = s.a2;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int t = 0; t < 9; ++t) red[b3_w * SLOT + b3_cog * 296 + k * 72 + b3_ci * 9 + t] = acc2[k][t];
        if (b3_ci == 0) red[(kThreads / 32) * SLOT + b3_w * C2 + b3_cog * 4 + k] = accb2[k];
      }
      __syncthreads();
      for (int i = tid; i < NW2 + C2; i += kThreads) {
        const int at = i < NW2 ? (i / 288) * 296 + i % 288 : (kThreads / 32) * SLOT + (i - NW2);
        const int stride = i < NW2 ? SLOT : C2;
        float a = 0.f;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) a += red[at + w * stride];
        s.dW2s[i] += a;
      }
    }

    // ================= B4: da1 = conv2^T(dc2), masked by relu'(a1)
    // dc2 (the gradient at conv2's 24x48 outputs: da2m at each pooled cell's argmax, zero elsewhere) is written out
    // DENSE, 8 channels at a time, as zero-bordered planes in a1's layout over the dead a2/g3 region; the data gradient
    // is then a plain 3x3 convolution with flipped taps -- one thread = one 2x2 block of positions x all 8 input
    // channels, vector loads, no per-tap index tests.
    {
      const int py = tid / W2, px = tid % W2;
      float* const dp = s.a2;                                   // [8][A1_PLANE], spills into g3
      float acc[C1][4];
#pragma unroll
      for (int c = 0; c < C1; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.f;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        __syncthreads();                                        // B2/B3 (h = 0) or the previous half's reads are done
        if (h == 0) {
          constexpr int BORDER = A1_PLANE - H1 * W1;            // 204 floats per plane
          for (int i = tid; i < 8 * BORDER; i += kThreads) {
            const int pl = i / BORDER, r = i % BORDER;
            int off;
            if (r < A1_LD) off = r;
            else if (r < 2 * A1_LD) off = (A1_ROWS - 1) * A1_LD + (r - A1_LD);
            else if (r < 2 * A1_LD + 4 * H1) { const int k = r - 2 * A1_LD, e = k & 3; off = (1 + (k >> 2)) * A1_LD + (e == 0 ? 0 : W1 + e); }
            else off = A1_ROWS * A1_LD + (r - 2 * A1_LD - 4 * H1);
            dp[pl * A1_PLANE + off] = 0.f;
          }
        }
#pragma unroll
        for (int co = 0; co < 8; ++co) {
          const int c = h * 8 + co;
          const float d = s.da2m[c * (H2 * W2) + tid];
          const int id = s.idx2[c * (H2 * W2) + tid];
          float* q = dp + co * A1_PLANE + (2 * py + 1) * A1_LD + 2 * px + 1;
          q[0] = id == 0 ? d : 0.f; q[1] = id == 1 ? d : 0.f;
          q[A1_LD] = id == 2 ? d : 0.f; q[A1_LD + 1] = id == 3 ? d : 0.f;
        }
        __syncthreads();
#pragma unroll 1
        for (int co = 0; co < 8; ++co) {
          float P[4][4];
          const float* base = dp + co * A1_PLANE + (2 * py) * A1_LD + 2 * px;
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const float2 a = *reinterpret_cast<const float2*>(base + r * A1_LD);
            const float2 b = *reinterpret_cast<const float2*>(base + r * A1_LD + 2);
            P[r][0] = a.x; P[r][1] = a.y; P[r][2] = b.x; P[r][3] = b.y;
          }
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const int dy = tap / 3, dx = tap % 3;
            float wv[C1];
            *reinterpret_cast<float4*>(wv) = *reinterpret_cast<const float4*>(&s.w2t[h * 8 + co][tap][0]);
            *reinterpret_cast<float4*>(wv + 4) = *reinterpret_cast<const float4*>(&s.w2t[h * 8 + co][tap][4]);
#pragma unroll
            for (int c = 0; c < C1; ++c) {                      // position (a,b) of the block takes dc2 at (a+2-dy, b+2-dx)
              acc[c][0] = fmaf(wv[c], P[2 - dy][2 - dx], acc[c][0]);
              acc[c][1] = fmaf(wv[c], P[2 - dy][3 - dx], acc[c][1]);
              acc[c][2] = fmaf(wv[c], P[3 - dy][2 - dx], acc[c][2]);
              acc[c][3] = fmaf(wv[c], P[3 - dy][3 - dx], acc[c][3]);
            }
          }
        }
      }
#pragma unroll
      for (int c = 0; c < C1; ++c)
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          const float* m = s.a1 + c * A1_PLANE + (2 * py + a + 1) * A1_LD + 2 * px + 1;
          float2 v;
          v.x = m[0] > 0.f ? acc[c][2 * a] : 0.f;
          v.y = m[1] > 0.f ? acc[c][2 * a + 1] : 0.f;
          *reinterpret_cast<float2*>(&s.da1m[c * (H1 * W1) + (2 * py + a) * W1 + 2 * px]) = v;
        }
    }
    __syncthreads();
    // the dense planes lay over a2 and the head of g3: give both their zero borders back (interiors are rewritten
    // by the next frame's forward pass)
    {
      float4* z = reinterpret_cast<float4*>(s.a2);
      for (int i = tid; i < 8 * A1_PLANE / 4; i += kThreads) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }

    // ================= B5: dW1, db1 -- one thread = 32 pooled cells of one channel x all 9 taps
    {
      const float* dptr = s.da1m + b5_co * (H1 * W1);
      const uint8_t* iptr = s.idx1 + b5_co * (H1 * W1);
#pragma unroll 2
      for (int k = 0; k < (H1 * W1) / 36; ++k) {
        const int cell = b5_grp + 36 * k;
        const float d = dptr[cell];
        const int id = iptr[cell];
        const int py = cell / W1, px = cell % W1;
        const float* p = s.in + (2 * py + (id >> 1)) * IN_LD + 2 * px + (id & 1);
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) acc1[dy * 3 + dx] = fmaf(d, p[dy * IN_LD + dx], acc1[dy * 3 + dx]);
        accb1 += d;
      }
    }
    __syncthreads();      // all reads of this frame's planes are done before the next frame overwrites them
  }

  // ---- one global accumulation per CTA
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int t = 0; t < 9; ++t) atomicAdd(gW3 + ((b1_cog * 4 + k) * C2 + b1_ci) * 9 + t, acc3[k][t]);
  for (int i = tid; i < C2 * C1 * 9 + C2; i += kThreads) {
    if (i < C2 * C1 * 9) atomicAdd(gW2 + i, s.dW2s[i]); else atomicAdd(gb2 + (i - C2 * C1 * 9), s.dW2s[i]);
  }
  {                                                             // dW1/db1: 36 partial sums per channel meet in shared memory
    float* red = s.da1m;                                        // [288][10]; the frame loop ended with a barrier
#pragma unroll
    for (int t = 0; t < 9; ++t) red[tid * 10 + t] = acc1[t];
    red[tid * 10 + 9] = accb1;
    __syncthreads();
    if (tid < C1 * 10) {
      const int co = tid / 10, t = tid % 10;
      float a = 0.f;
      for (int g = 0; g < 36; ++g) a += red[(co * 36 + g) * 10 + t];
      if (t < 9) atomicAdd(gW1 + co * 9 + t, a); else atomicAdd(gb1 + co, a);
    }
  }
  if (tid < C3) atomicAdd(gb3 + tid, accb3);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int i = tid + k * kThreads;
    if (i < E * C3) atomicAdd(gWfc + i, accfc[k]);
    else if (i < E * C3 + E) atomicAdd(gbfc + (i - E * C3), accfc[k]);
  }
}

}  // namespace cnn
}  // namespace ssb

// Accumulates (+=) the 8 parameter gradients {net.0.w,b, net.3.w,b, net.6.w,b, fc.w,b} of the fused
// ROI CNN given d_out[(b*T+t)*d_out_stride + 0..31] = dL/d(embedding).  saved = NULL: R is re-read and the forward
// recomputed in shared memory (nothing kept by the forward call).  saved = the records written by
// ssb_roi_cnn_fwd_tc3(passes = 3, saved) for the SAME R / lengths / weights: the recompute is skipped (R is still read
// for the conv1 weight gradient).
extern "C" int ssb_roi_cnn_bwd(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                               int32_t roi_h, int32_t roi_w, const float* const* weights, int32_t standardize,
                               const float* d_out, int64_t d_out_stride, float* const* grads, const void* saved,
                               void* stream) {
  using namespace ssb;
  using namespace ssb::cnn;
  SSB_REQUIRE(R && weights && d_out && grads, "null pointer");
  SSB_REQUIRE(roi_h == H0 && roi_w == W0, "ROI must be 48x96");
  SSB_REQUIRE(d_out_stride >= E, "d_out_stride < 32");
  SSB_REQUIRE(((uintptr_t)saved & 15) == 0, "saved must be 16-byte aligned");
  for (int i = 0; i < 8; ++i) SSB_REQUIRE(weights[i] && grads[i], "null weight/grad pointer");
  if (B * T == 0) return SSB_OK;
  Weights gw{weights[0], weights[1], weights[2], weights[3], weights[4], weights[5], weights[6], weights[7]};
  SSB_CUDA(cudaFuncSetAttribute(roi_cnn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
  const int64_t frames = B * T;
  const int grid = (int)(frames < kSMs ? frames : kSMs);
  roi_cnn_bwd_kernel<<<grid, kThreads, sizeof(BwdSmem), (cudaStream_t)stream>>>(
      R, lengths, (int)B, (int)T, gw, standardize, d_out, d_out_stride,
      grads[0], grads[1], grads[2], grads[3], grads[4], grads[5], grads[6], grads[7], reinterpret_cast<const uint8_t*>(saved));
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
