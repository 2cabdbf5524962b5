// K4 -- masked attention pool + LayerNorm + Linear/ReLU/(Dropout)/Linear, with the
// inference epilogue (softmax, top-k) and the training epilogue (label-smoothed CE) fused.
//
//   AttnPool   train_model_official.py:231-248   score = h.w_s + b_s ; softmax over t < len ; sum_t a_t h_t
//   head       train_model_official.py:271-277   LayerNorm(384, eps 1e-5) -> Linear(384,128) -> ReLU
//                                                 -> Dropout(0.2) -> Linear(128,C)
//   top-k      live_infer_official.py:223-226     softmax, descending argsort
//   loss       train_model_official.py:405        CrossEntropyLoss(label_smoothing=0.05), per-clip terms
//
// masked_fill(-1e9)+softmax equals a softmax restricted to t < len in fp32 (exp underflows to
// exactly 0), which is what is computed.  One CTA = 4 clips, one warp per clip for the pooling
// (single pass, online softmax), the four warps share the fc1 weight rows.
#include "common.cuh"

namespace ssb {
namespace head {

constexpr int D = 2 * kHidden;      // 384
constexpr int F = 128;              // head.1 out features
constexpr int CPB = 4;              // clips per CTA
constexpr int THREADS = 32 * CPB;
constexpr int PER_LANE = D / 32;    // 12
constexpr int MAXC = 1024;

struct Params {
  const float *ws, *bs;             // pool.score  [1,384],[1]
  const float *ln_g, *ln_b;         // head.0      [384],[384]
  const float *w1, *b1;             // head.1      [128,384],[128]
  const float *w2, *b2;             // head.4      [C,128],[C]
};

struct Saved {                      // all nullable (inference)
  float* attn;                      // [B,T]   softmax weights (0 past len)
  float* pooled;                    // [B,384]
  float* xhat;                      // [B,384] normalised pooled vector (before affine)
  float* rstd;                      // [B]
  float* ln;                        // [B,384] LayerNorm output
  float* hid;                       // [B,128] after ReLU and dropout
};

__global__ void __launch_bounds__(THREADS)
head_fwd_kernel(const float* __restrict__ seq, const int64_t* __restrict__ lengths, int B, int T, int C,
                Params p, Saved sv, float* __restrict__ logits, float* __restrict__ probs,
                int32_t* __restrict__ topk_ids, float* __restrict__ topk_probs, int k_top,
                const int64_t* __restrict__ y, float label_smoothing, float* __restrict__ loss,
                float drop_p, uint64_t seed) {
  __shared__ float s_ln[CPB][D];
  __shared__ float s_hid[CPB][F];
  extern __shared__ float s_logit[];          // [CPB][C]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * CPB + warp;
  const bool live = b < B;
  int len = live ? (int)lengths[b] : 0;
  len = max(0, min(len, T));

  // ---- attention pooling, one pass (online softmax)
  float wsv[PER_LANE];
#pragma unroll
  for (int i = 0; i < PER_LANE; ++i) wsv[i] = __ldg(p.ws + lane + 32 * i);
  const float bsv = __ldg(p.bs);
  float m = -INFINITY, l = 0.f, acc[PER_LANE];
#pragma unroll
  for (int i = 0; i < PER_LANE; ++i) acc[i] = 0.f;
  for (int t = 0; t < len; ++t) {
    const float* row = seq + ((int64_t)b * T + t) * D;
    float hv[PER_LANE], dot = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) { hv[i] = __ldg(row + lane + 32 * i); dot = fmaf(hv[i], wsv[i], dot); }
    const float sc = warp_sum(dot) + bsv;
    if (sv.attn && lane == 0) sv.attn[(int64_t)b * T + t] = sc;      // raw score, normalised below
    const float mn = fmaxf(m, sc);
    const float corr = expf(m - mn), e = expf(sc - mn);              // exp(-inf) = 0 on the first frame
    l = l * corr + e;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) acc[i] = acc[i] * corr + e * hv[i];
    m = mn;
  }
  const float inv_l = len > 0 ? 1.0f / l : 0.f;
  if (sv.attn && live) {
    __syncwarp();
    for (int t = lane; t < T; t += 32) {
      float* a = sv.attn + (int64_t)b * T + t;
      *a = t < len ? expf(*a - m) * inv_l : 0.f;
    }
  }
  // ---- LayerNorm (biased variance, eps 1e-5)
  float x[PER_LANE], sum = 0.f;
#pragma unroll
  for (int i = 0; i < PER_LANE; ++i) { x[i] = acc[i] * inv_l; sum += x[i]; }
  const float mean = warp_sum(sum) / (float)D;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < PER_LANE; ++i) { const float d = x[i] - mean; sq += d * d; }
  const float rstd = rsqrtf(warp_sum(sq) / (float)D + 1e-5f);
#pragma unroll
  for (int i = 0; i < PER_LANE; ++i) {
    const int kk = lane + 32 * i;
    const float xh = (x[i] - mean) * rstd;
    const float yv = xh * __ldg(p.ln_g + kk) + __ldg(p.ln_b + kk);
    s_ln[warp][kk] = yv;
    if (live) {
      if (sv.pooled) sv.pooled[(int64_t)b * D + kk] = x[i];
      if (sv.xhat) sv.xhat[(int64_t)b * D + kk] = xh;
      if (sv.ln) sv.ln[(int64_t)b * D + kk] = yv;
    }
  }
  if (sv.rstd && live && lane == 0) sv.rstd[b] = rstd;
  __syncthreads();

  // ---- fc1 + ReLU (+ dropout): each warp takes 32 of the 128 rows and serves all 4 clips
  for (int j = warp * (F / CPB); j < (warp + 1) * (F / CPB); ++j) {
    float part[CPB];
#pragma unroll
    for (int c = 0; c < CPB; ++c) part[c] = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int kk = lane + 32 * i;
      const float w = __ldg(p.w1 + (int64_t)j * D + kk);
#pragma unroll
      for (int c = 0; c < CPB; ++c) part[c] = fmaf(w, s_ln[c][kk], part[c]);
    }
    const float bj = __ldg(p.b1 + j);
#pragma unroll
    for (int c = 0; c < CPB; ++c) {
      float v = fmaxf(warp_sum(part[c]) + bj, 0.f);
      const int bc = blockIdx.x * CPB + c;
      if (drop_p > 0.f) v = dropout_keep(seed, (uint64_t)bc * F + j, drop_p) ? v / (1.0f - drop_p) : 0.f;
      if (lane == 0) {
        s_hid[c][j] = v;
        if (sv.hid && bc < B) sv.hid[(int64_t)bc * F + j] = v;
      }
    }
  }
  __syncthreads();

  // ---- fc2: warp = clip
  float* lg = s_logit + warp * C;
  for (int c = 0; c < C; ++c) {
    float part = 0.f;
#pragma unroll
    for (int i = 0; i < F / 32; ++i) part = fmaf(__ldg(p.w2 + (int64_t)c * F + lane + 32 * i), s_hid[warp][lane + 32 * i], part);
    const float v = warp_sum(part) + __ldg(p.b2 + c);
    if (lane == 0) { lg[c] = v; if (live) logits[(int64_t)b * C + c] = v; }
  }
  __syncwarp();
  if (!live || !(probs || topk_ids || loss)) return;

  // ---- softmax over classes
  float mx = -INFINITY;
  for (int c = lane; c < C; c += 32) mx = fmaxf(mx, lg[c]);
  mx = warp_max(mx);
  float se = 0.f;
  for (int c = lane; c < C; c += 32) se += expf(lg[c] - mx);
  se = warp_sum(se);
  const float lse = mx + logf(se);
  if (loss) {
    // (1-eps) * (-log p_y) + eps * mean_c(-log p_c)
    float nl = 0.f;
    for (int c = lane; c < C; c += 32) nl += lse - lg[c];
    nl = warp_sum(nl);
    if (lane == 0) {
      const int yy = (int)y[b];
      loss[b] = (1.0f - label_smoothing) * (lse - lg[yy]) + label_smoothing * nl / (float)C;
    }
  }
  if (probs) for (int c = lane; c < C; c += 32) probs[(int64_t)b * C + c] = expf(lg[c] - mx) / se;
  if (topk_ids) {
    // k rounds of arg-max; ties resolve to the lower class id
    for (int r = 0; r < k_top; ++r) {
      float bv = -INFINITY; int bi = 0x7fffffff;
      for (int c = lane; c < C; c += 32) { const float v = lg[c]; if (v > bv) { bv = v; bi = c; } }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
      }
      if (lane == 0) {
        topk_ids[(int64_t)b * k_top + r] = bi;
        if (topk_probs) topk_probs[(int64_t)b * k_top + r] = expf(bv - mx) / se;
        lg[bi] = -INFINITY;
      }
      __syncwarp();
    }
  }
}

}  // namespace head
}  // namespace ssb

// weights: 8 pointers {pool.score.weight, pool.score.bias, head.0.weight, head.0.bias,
//                      head.1.weight, head.1.bias, head.4.weight, head.4.bias}
// saved:   6 pointers {attn[B,T], pooled[B,384], xhat[B,384], rstd[B], ln[B,384], hid[B,128]} or NULL
extern "C" int ssb_head_fwd(const float* seq, const int64_t* lengths, int64_t B, int64_t T, int32_t dim,
                            const float* const* weights, int32_t num_classes, float* const* saved,
                            float* logits, float* probs, int32_t* topk_ids, float* topk_probs, int32_t k_top,
                            const int64_t* y, float label_smoothing, float* loss,
                            float dropout_p, uint64_t seed, void* stream) {
  using namespace ssb;
  using namespace ssb::head;
  SSB_REQUIRE(seq && lengths && weights && logits, "null pointer");
  SSB_REQUIRE(dim == D, "pooled width must be 384 (2 x hidden 192)");
  SSB_REQUIRE(num_classes >= 1 && num_classes <= MAXC, "num_classes out of range");
  SSB_REQUIRE(!loss || y, "loss needs labels");
  SSB_REQUIRE(!topk_ids || (k_top >= 1 && k_top <= num_classes), "bad k");
  SSB_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "bad dropout p");
  for (int i = 0; i < 8; ++i) SSB_REQUIRE(weights[i], "null weight pointer");
  if (B == 0) return SSB_OK;
  Params p{weights[0], weights[1], weights[2], weights[3], weights[4], weights[5], weights[6], weights[7]};
  Saved sv{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  if (saved) sv = Saved{saved[0], saved[1], saved[2], saved[3], saved[4], saved[5]};
  const size_t dyn = sizeof(float) * CPB * num_classes;
  head_fwd_kernel<<<cdiv(B, CPB), THREADS, dyn, (cudaStream_t)stream>>>(
      seq, lengths, (int)B, (int)T, num_classes, p, sv, logits, probs, topk_ids, topk_probs, k_top, y,
      label_smoothing, loss, dropout_p, seed);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
