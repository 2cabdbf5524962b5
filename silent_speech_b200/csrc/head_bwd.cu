// K4-bwd -- backward of the pooled head, split at its two matrix products so that the
// weight gradients (dW2 = dlogits^T hid, dW1 = dhid^T ln) and dln = dhid W1 run as GEMMs.
//
//   head_bwd_logits   CE(label_smoothing) gradient at the logits (train_model_official.py:405,434)
//                     or a caller-provided dlogits; back through Linear(128,C), Dropout, ReLU
//   head_bwd_pool     back through LayerNorm and AttnPool (train_model_official.py:240-248, 272)
//                     to the GRU output sequence; accumulates d(ln affine), d(pool.score)
//   colsum            bias gradients
#include "common.cuh"

namespace ssb {
namespace head {

constexpr int D = 2 * kHidden;      // 384
constexpr int F = 128;
constexpr int PER_LANE = D / 32;    // 12
constexpr int CPB = 4;

// grid: B blocks of 128 threads.  dlogits_in == nullptr -> fused CE gradient:
//   dlogit_c = gscale * (softmax_c - (1-eps)[c==y] - eps/C)
__global__ void __launch_bounds__(F)
head_bwd_logits_kernel(const float* __restrict__ logits, const float* __restrict__ dlogits_in,
                       const int64_t* __restrict__ y, float label_smoothing, float gscale,
                       const float* __restrict__ w2, const float* __restrict__ hid, float drop_p,
                       int C, float* __restrict__ dlogits, float* __restrict__ dhid) {
  extern __shared__ float s_d[];       // [C]
  __shared__ float s_red[8];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (dlogits_in) {
    for (int c = tid; c < C; c += F) s_d[c] = dlogits_in[(int64_t)b * C + c];
  } else {
    const float* lg = logits + (int64_t)b * C;
    float mx = -INFINITY;
    for (int c = tid; c < C; c += F) mx = fmaxf(mx, lg[c]);
    mx = warp_max(mx);
    if ((tid & 31) == 0) s_red[tid >> 5] = mx;
    __syncthreads();
    mx = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
    float se = 0.f;
    for (int c = tid; c < C; c += F) se += expf(lg[c] - mx);
    se = warp_sum(se);
    if ((tid & 31) == 0) s_red[4 + (tid >> 5)] = se;
    __syncthreads();
    se = s_red[4] + s_red[5] + s_red[6] + s_red[7];
    const int yy = (int)y[b];
    for (int c = tid; c < C; c += F) {
      const float pc = expf(lg[c] - mx) / se;
      const float tgt = (c == yy ? 1.0f - label_smoothing : 0.f) + label_smoothing / (float)C;
      s_d[c] = gscale * (pc - tgt);
    }
  }
  __syncthreads();
  for (int c = tid; c < C; c += F) dlogits[(int64_t)b * C + c] = s_d[c];
  // dhid_j = sum_c dlogit_c W2[c,j], through dropout and ReLU (hid > 0 <=> kept and active)
  float a = 0.f;
  for (int c = 0; c < C; ++c) a = fmaf(s_d[c], __ldg(w2 + (int64_t)c * F + tid), a);
  const float hv = hid[(int64_t)b * F + tid];
  dhid[(int64_t)b * F + tid] = hv > 0.f ? a / (1.0f - drop_p) : 0.f;
}

// One warp per clip, 4 clips per CTA.  s_da: dynamic [CPB][T].
__global__ void __launch_bounds__(32 * CPB)
head_bwd_pool_kernel(const float* __restrict__ dln, const float* __restrict__ xhat, const float* __restrict__ rstd,
                     const float* __restrict__ attn, const float* __restrict__ seq,
                     const int64_t* __restrict__ lengths, int B, int T,
                     const float* __restrict__ ln_g, const float* __restrict__ ws,
                     float* __restrict__ dseq, float* __restrict__ d_ln_g, float* __restrict__ d_ln_b,
                     float* __restrict__ d_ws, float* __restrict__ d_bs) {
  extern __shared__ float s_da[];
  __shared__ float s_acc[3][D];        // d_ln_g, d_ln_b, d_ws partials of this CTA
  __shared__ float s_dbs;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * CPB + warp;
  const bool live = b < B;
  for (int i = threadIdx.x; i < 3 * D; i += blockDim.x) (&s_acc[0][0])[i] = 0.f;
  if (threadIdx.x == 0) s_dbs = 0.f;
  __syncthreads();
  if (live) {
    int len = (int)lengths[b];
    len = max(0, min(len, T));
    float dp[PER_LANE], wsv[PER_LANE], dws[PER_LANE];
    float c1 = 0.f, c2 = 0.f;
    float dxh[PER_LANE], xh[PER_LANE];
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int k = lane + 32 * i;
      const float d = dln[(int64_t)b * D + k];
      xh[i] = xhat[(int64_t)b * D + k];
      atomicAdd(&s_acc[0][k], d * xh[i]);
      atomicAdd(&s_acc[1][k], d);
      dxh[i] = d * __ldg(ln_g + k);
      c1 += dxh[i]; c2 += dxh[i] * xh[i];
      wsv[i] = __ldg(ws + k);
      dws[i] = 0.f;
    }
    c1 = warp_sum(c1) / (float)D; c2 = warp_sum(c2) / (float)D;
    const float rs = rstd[b];
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) dp[i] = rs * (dxh[i] - c1 - xh[i] * c2);     // d pooled

    float* da = s_da + warp * T;
    float S = 0.f;
    for (int t = 0; t < len; ++t) {
      const float* row = seq + ((int64_t)b * T + t) * D;
      float dot = 0.f;
#pragma unroll
      for (int i = 0; i < PER_LANE; ++i) dot = fmaf(dp[i], __ldg(row + lane + 32 * i), dot);
      dot = warp_sum(dot);
      if (lane == 0) da[t] = dot;
      S = fmaf(attn[(int64_t)b * T + t], dot, S);
    }
    __syncwarp();
    float dbs = 0.f;
    for (int t = 0; t < T; ++t) {
      float* orow = dseq + ((int64_t)b * T + t) * D;
      if (t >= len) {
#pragma unroll
        for (int i = 0; i < PER_LANE; ++i) orow[lane + 32 * i] = 0.f;
        continue;
      }
      const float a = attn[(int64_t)b * T + t];
      const float ds = a * (da[t] - S);                       // d score_t
      dbs += ds;
      const float* row = seq + ((int64_t)b * T + t) * D;
#pragma unroll
      for (int i = 0; i < PER_LANE; ++i) {
        const float hv = __ldg(row + lane + 32 * i);
        dws[i] = fmaf(ds, hv, dws[i]);
        orow[lane + 32 * i] = fmaf(a, dp[i], ds * wsv[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) atomicAdd(&s_acc[2][lane + 32 * i], dws[i]);
    if (lane == 0) atomicAdd(&s_dbs, dbs);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < D; k += blockDim.x) {
    atomicAdd(d_ln_g + k, s_acc[0][k]);
    atomicAdd(d_ln_b + k, s_acc[1][k]);
    atomicAdd(d_ws + k, s_acc[2][k]);
  }
  if (threadIdx.x == 0) atomicAdd(d_bs, s_dbs);
}

// out[j] += sum_i A[i*lda + j]; grid (ceil(N/32), row_blocks)
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ A, int64_t lda, int64_t M, int N, float* __restrict__ out) {
  __shared__ float s[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + tx;
  float a = 0.f;
  if (j < N)
    for (int64_t i = (int64_t)blockIdx.y * 8 + ty; i < M; i += (int64_t)gridDim.y * 8) a += A[i * lda + j];
  s[ty][tx] = a;
  __syncthreads();
  if (ty == 0 && j < N) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += s[r][tx];
    atomicAdd(out + j, t);
  }
}

}  // namespace head
}  // namespace ssb

extern "C" int ssb_head_bwd_logits(const float* logits, const float* dlogits_in, const int64_t* y,
                                   float label_smoothing, float grad_scale, const float* w2, const float* hid,
                                   float dropout_p, int64_t B, int32_t num_classes,
                                   float* dlogits, float* dhid, void* stream) {
  using namespace ssb;
  using namespace ssb::head;
  SSB_REQUIRE((dlogits_in || (logits && y)) && w2 && hid && dlogits && dhid, "null pointer");
  SSB_REQUIRE(num_classes >= 1 && num_classes <= 1024, "num_classes out of range");
  if (B == 0) return SSB_OK;
  head_bwd_logits_kernel<<<(unsigned)B, F, sizeof(float) * num_classes, (cudaStream_t)stream>>>(
      logits, dlogits_in, y, label_smoothing, grad_scale, w2, hid, dropout_p, num_classes, dlogits, dhid);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// Accumulates (+=) into d_ln_g[384], d_ln_b[384], d_ws[384], d_bs[1]; overwrites dseq [B,T,384].
extern "C" int ssb_head_bwd_pool(const float* dln, const float* xhat, const float* rstd, const float* attn,
                                 const float* seq, const int64_t* lengths, int64_t B, int64_t T,
                                 const float* ln_g, const float* ws, float* dseq,
                                 float* d_ln_g, float* d_ln_b, float* d_ws, float* d_bs, void* stream) {
  using namespace ssb;
  using namespace ssb::head;
  SSB_REQUIRE(dln && xhat && rstd && attn && seq && lengths && ln_g && ws && dseq && d_ln_g && d_ln_b && d_ws && d_bs,
              "null pointer");
  SSB_REQUIRE(T * CPB * sizeof(float) <= 40000, "T too large for the pooled-head backward");
  if (B == 0) return SSB_OK;
  head_bwd_pool_kernel<<<cdiv(B, CPB), 32 * CPB, sizeof(float) * CPB * T, (cudaStream_t)stream>>>(
      dln, xhat, rstd, attn, seq, lengths, (int)B, (int)T, ln_g, ws, dseq, d_ln_g, d_ln_b, d_ws, d_bs);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

extern "C" int ssb_colsum(const float* A, int64_t lda, int64_t M, int64_t N, float* out, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(A && out, "null pointer");
  if (M == 0 || N == 0) return SSB_OK;
  int rb = (int)((M + 255) / 256);
  if (rb > 256) rb = 256;
  if (rb < 1) rb = 1;
  dim3 grid(cdiv(N, 32), rb);
  head::colsum_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, lda, M, (int)N, out);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
