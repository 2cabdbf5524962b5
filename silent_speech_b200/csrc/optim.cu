// K5 -- global-norm clip + Adam over one flat fp32 bucket, and the dropout mask kernel.
//
//   clip   nn.utils.clip_grad_norm_(params, 1.0)   train_model_official.py:438
//          coef = min(1, max_norm / (||g||_2 + 1e-6)) over ALL parameters jointly
//   Adam   torch.optim.Adam(lr=3e-4), defaults      train_model_official.py:403, 439
//          m += (1-b1)(g-m);  v = b2 v + (1-b2) g^2;  p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
//   dropout  nn.GRU(dropout=0.1) between the layers (train:266), counter-based mask so that
//          forward and backward regenerate the same keep/drop decisions from (seed, index)
// HBM-bound: 4 reads + 3 writes of 4 B per parameter.
#include "common.cuh"

namespace ssb {

__global__ void __launch_bounds__(256)
sqnorm_kernel(const float* __restrict__ g, int64_t n, double* __restrict__ out) {
  __shared__ double red[8];
  double a = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = g[i];
    a += (double)v * (double)v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; ++i) t += red[i];
    atomicAdd(out, t);
  }
}

__global__ void __launch_bounds__(256)
clip_adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 int64_t n, const double* __restrict__ sqnorm, float max_norm, float step_size,
                 float beta1, float beta2, float bc2_sqrt, float eps, int write_back_grad,
                 float* __restrict__ norm_out) {
  const float total = (float)sqrt(*sqnorm);
  const float coef = fminf(max_norm / (total + 1e-6f), 1.0f);
  if (norm_out && blockIdx.x == 0 && threadIdx.x == 0) *norm_out = total;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * coef;
    float mi = m[i], vi = v[i];
    mi = mi + (1.0f - beta1) * (gi - mi);                 // exp_avg.lerp_(grad, 1-beta1)
    vi = vi * beta2 + (1.0f - beta2) * gi * gi;           // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
    m[i] = mi; v[i] = vi;
    if (write_back_grad) g[i] = gi;
  }
}

__global__ void __launch_bounds__(256)
dropout_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n, float p, uint64_t seed) {
  const float scale = 1.0f / (1.0f - p);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = dropout_keep(seed, (uint64_t)i, p) ? x[i] * scale : 0.f;
}

}  // namespace ssb

// sqnorm[0] += sum g^2 (double; the caller zeroes it once per step, may call per bucket)
extern "C" int ssb_grad_sqnorm(const float* g, int64_t n, double* sqnorm, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(g && sqnorm, "null pointer");
  if (n == 0) return SSB_OK;
  int blocks = cdiv(n, 256 * 8);
  if (blocks > 4 * kSMs) blocks = 4 * kSMs;
  sqnorm_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(g, n, sqnorm);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// One Adam step with the joint-norm clip folded in.  `step` is 1-based.  norm_out (nullable)
// receives the pre-clip global norm (what clip_grad_norm_ returns).
extern "C" int ssb_clip_adam_step(float* p, float* g, float* m, float* v, int64_t n, const double* sqnorm,
                                  float max_norm, float lr, float beta1, float beta2, float eps, int64_t step,
                                  int32_t write_back_grad, float* norm_out, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(p && g && m && v && sqnorm, "null pointer");
  SSB_REQUIRE(step >= 1, "step is 1-based");
  if (n == 0) return SSB_OK;
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  int blocks = cdiv(n, 256 * 4);
  if (blocks > 8 * kSMs) blocks = 8 * kSMs;
  clip_adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, n, sqnorm, max_norm, step_size, beta1, beta2,
                                                             bc2_sqrt, eps, write_back_grad, norm_out);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// y = x * keep / (1-p); the same (seed, p) on the upstream gradient is the backward.
extern "C" int ssb_dropout(const float* x, float* y, int64_t n, float p, uint64_t seed, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(x && y, "null pointer");
  SSB_REQUIRE(p >= 0.f && p < 1.f, "bad p");
  if (n == 0) return SSB_OK;
  int blocks = cdiv(n, 256 * 4);
  if (blocks > 8 * kSMs) blocks = 8 * kSMs;
  dropout_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, y, n, p, seed);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
