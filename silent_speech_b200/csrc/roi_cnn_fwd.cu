// K2 (fp32 reference-precision build) -- fused mouth-ROI CNN forward.
//
// Replaces, per batch, train_model_official.py:286-294 (normalise, standardise,
// TinyROICNN.forward at 223-229) -- and live_infer_official.py:126-127 when
// standardize == 0.  Persistent CTAs stride over the B*T frames; frames at
// t >= lengths[b] are never read (their embedding is written as zeros, which is
// what the packed GRU downstream ignores anyway).
#include "roi_cnn.cuh"

namespace ssb {
namespace cnn {

struct FwdSmem {
  SmemWeights w;
  float in[IN_ROWS * IN_LD];
  float a1[C1 * A1_PLANE];
  float a2[C2 * A2_PLANE];
  float part[18 * 12];
  float mean[C3];
  float red[16];
};

__global__ void __maxnreg__(112)
roi_cnn_fwd_kernel(const uint8_t* __restrict__ R, const int64_t* __restrict__ lengths, int B, int T,
                   Weights gw, int standardize, float* __restrict__ out, int64_t out_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FwdSmem& s = *reinterpret_cast<FwdSmem*>(smem_raw);
  const int tid = threadIdx.x;
  load_weights(s.w, gw, tid);
  for (int i = tid; i < IN_ROWS * IN_LD; i += kThreads) s.in[i] = 0.f;
  for (int i = tid; i < C1 * A1_PLANE; i += kThreads) s.a1[i] = 0.f;
  for (int i = tid; i < C2 * A2_PLANE; i += kThreads) s.a2[i] = 0.f;
  __syncthreads();

  const int64_t frames = (int64_t)B * T;
  for (int64_t f = blockIdx.x; f < frames; f += gridDim.x) {
    const int b = (int)(f / T), t = (int)(f % T);
    float* o = out + f * out_stride;
    if (lengths && t >= (int)lengths[b]) {
      if (tid < E) o[tid] = 0.f;
      continue;
    }
    load_frame(R + f * NPIX, s.in, s.red, tid, standardize != 0);
    __syncthreads();
    conv1_pool<false>(s.in, s.a1, s.w, nullptr, tid);
    __syncthreads();
    conv2_pool<false>(s.a1, s.a2, s.w, nullptr, tid);
    __syncthreads();
    conv3_mean<false>(s.a2, s.w, s.part, s.mean, nullptr, tid);
    if (tid < E) o[tid] = fc_out(s.w, s.mean, tid);
    // next iteration's first smem write (s.in interior) is ordered after everyone's conv1
    // reads by the barriers above; s.mean/s.part are rewritten only after two more barriers.
  }
}

}  // namespace cnn
}  // namespace ssb

extern "C" int ssb_roi_cnn_fwd(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                               int32_t roi_h, int32_t roi_w, const float* const* weights,
                               int32_t standardize, float* out, int64_t out_stride, void* stream) {
  using namespace ssb;
  using namespace ssb::cnn;
  SSB_REQUIRE(R && weights && out, "null pointer");
  SSB_REQUIRE(roi_h == H0 && roi_w == W0, "ROI must be 48x96 (train_model_official.py:38)");
  SSB_REQUIRE(out_stride >= E, "out_stride < 32");
  for (int i = 0; i < 8; ++i) SSB_REQUIRE(weights[i], "null weight pointer");
  SSB_REQUIRE(((uintptr_t)R & 3) == 0, "R must be 4-byte aligned");
  if (B * T == 0) return SSB_OK;
  Weights gw{weights[0], weights[1], weights[2], weights[3], weights[4], weights[5], weights[6], weights[7]};
  static bool configured = false;
  if (!configured) {
    SSB_CUDA(cudaFuncSetAttribute(roi_cnn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(FwdSmem)));
    configured = true;
  }
  const int64_t frames = B * T;
  const int grid = (int)(frames < 2 * kSMs ? frames : 2 * kSMs);
  roi_cnn_fwd_kernel<<<grid, kThreads, sizeof(FwdSmem), (cudaStream_t)stream>>>(
      R, lengths, (int)B, (int)T, gw, standardize, out, out_stride);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
