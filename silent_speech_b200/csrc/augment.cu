// On-device training augmentation of the point features, the policy of
// NPZWordDataset.__getitem__, train_model_official.py:143-152 + 154-172:
//   with probability 0.7:                       X += N(0, NOISE_STD = 0.01)            (all T rows of the clip)
//   if T > 12, with probability 0.35:           k = randint(1, DROP_FRAMES_MAX = 2); k distinct frames drawn from
//                                               [1, T-2] are REMOVED from X (rows close up, T -= k)
//   then X is zero-padded to max_t and -- a quirk of the reference kept as is -- the ROI clip is NOT thinned: it is
//   only cut to the new length T_use = min(T - k, Tr, max_t) and zero-padded, so after a drop X and R are offset by
//   one or two frames from the drop point on.
// The reference draws from Python's / numpy's global CPU generators, so bit parity is impossible by construction;
// the kernel uses the counter-based generator of common.cuh (decisions depend only on (seed, clip)), writes the
// decisions it took to `decisions` [B,4] = {noise applied, k, drop_0, drop_1} so that tests replay them exactly, and
// is validated statistically (rates, index range, noise moments) plus exactly (compaction, padding, ROI cut).
// One CTA per clip; pure data movement + one Box-Muller pair per two elements: HBM-bound.
#include "common.cuh"

namespace ssb {
namespace aug {

__device__ __forceinline__ float u01(uint64_t seed, uint64_t idx) { return (mix32(seed, idx) >> 8) * (1.0f / 16777216.0f); }

__global__ void augment_kernel(const float* __restrict__ X, const int64_t* __restrict__ lengths, int T, int D,
                               uint8_t* __restrict__ R, int64_t roi_frame_bytes, uint64_t seed, float p_noise, float noise_std,
                               float p_drop, int drop_max, float* __restrict__ Xo, int64_t* __restrict__ lengths_out,
                               int32_t* __restrict__ decisions) {
  const int b = blockIdx.x, tid = threadIdx.x;
  const int len = max(0, min((int)lengths[b], T));
  __shared__ int s_dec[4];
  if (tid == 0) {
    const uint64_t base = (uint64_t)b << 8;
    const int noise = u01(seed, base + 0) < p_noise;
    int k = 0, d0 = -1, d1 = -1;
    if (len > 12 && u01(seed, base + 1) < p_drop) {
      k = 1 + min((int)(u01(seed, base + 2) * drop_max), drop_max - 1);
      k = min(k, 2);
      const int n = len - 2;                                   // candidates 1 .. len-2
      d0 = 1 + min((int)(u01(seed, base + 3) * n), n - 1);
      if (k == 2) {                                            // second index among the remaining n-1, without replacement
        int j = 1 + min((int)(u01(seed, base + 4) * (n - 1)), n - 2);
        if (j >= d0) ++j;
        d1 = j;
        if (d1 < d0) { const int t = d0; d0 = d1; d1 = t; }
      }
    }
    s_dec[0] = noise; s_dec[1] = k; s_dec[2] = d0; s_dec[3] = d1;
    if (decisions) { decisions[4 * b] = noise; decisions[4 * b + 1] = k; decisions[4 * b + 2] = d0; decisions[4 * b + 3] = d1; }
    if (lengths_out) lengths_out[b] = len - k;
  }
  __syncthreads();
  const int noise = s_dec[0], k = s_dec[1], d0 = s_dec[2], d1 = s_dec[3];
  const int new_len = len - k;
  const float* xin = X + (int64_t)b * T * D;
  float* xout = Xo + (int64_t)b * T * D;
  for (int i = tid; i < T * D; i += blockDim.x) {
    const int t = i / D, c = i - t * D;
    float v = 0.f;
    if (t < new_len) {
      int ts = t;                                              // source frame of output row t
      if (k >= 1 && ts >= d0) ++ts;
      if (k == 2 && ts >= d1) ++ts;
      v = xin[ts * D + c];
      if (noise) {                                             // noise is drawn for the SOURCE element (added before the drop)
        const uint64_t e = ((uint64_t)b * T + ts) * D + c;
        const float u1 = fmaxf(u01(seed ^ 0xA5A5A5A5ull, e >> 1 << 1), 5.9604645e-8f), u2 = u01(seed ^ 0x5A5A5A5Aull, e >> 1 << 1);
        const float r = sqrtf(-2.0f * logf(u1));
        float sn, cs;
        sincospif(2.0f * u2, &sn, &cs);
        v += noise_std * r * ((e & 1) ? sn : cs);
      }
    }
    xout[i] = v;
  }
  // ROI: frames past the new length become zero frames (roi_pad_trim, train:105-117); nothing is thinned
  if (R && k > 0) {
    uint32_t* r = reinterpret_cast<uint32_t*>(R + ((int64_t)b * T + new_len) * roi_frame_bytes);
    const int64_t words = (int64_t)k * roi_frame_bytes / 4;
    for (int64_t i = tid; i < words; i += blockDim.x) r[i] = 0u;
  }
}

}  // namespace aug
}  // namespace ssb

// X f32 [B,T,D] (zero rows past lengths) -> Xo (may not alias X) with the reference's augmentation policy; lengths_out
// [B] = lengths - frames dropped; R u8 [B,T,roi_frame_bytes] (nullable) is cut IN PLACE to the new length; decisions
// i32 [B,4] (nullable) = {noise applied, k, drop_0, drop_1}.  Defaults of the reference: p_noise 0.7, noise_std 0.01,
// p_drop 0.35, drop_max 2 (train_model_official.py:40-43, 143-152).
extern "C" int ssb_augment_clips(const float* X, const int64_t* lengths, int64_t B, int64_t T, int64_t D,
                                 uint8_t* R, int64_t roi_frame_bytes, uint64_t seed, float p_noise, float noise_std,
                                 float p_drop, int32_t drop_max, float* Xo, int64_t* lengths_out, int32_t* decisions,
                                 void* stream) {
  using namespace ssb;
  SSB_REQUIRE(X && lengths && Xo, "null pointer");
  SSB_REQUIRE(X != Xo, "Xo must not alias X (rows move)");
  SSB_REQUIRE(drop_max >= 1 && drop_max <= 2, "drop_max must be 1 or 2 (DROP_FRAMES_MAX, train_model_official.py:43)");
  SSB_REQUIRE(!R || (roi_frame_bytes % 4 == 0 && ((uintptr_t)R & 3) == 0), "ROI frames must be 4-byte multiples");
  SSB_REQUIRE(T * D < ((int64_t)1 << 31), "clip too large");
  if (B == 0) return SSB_OK;
  aug::augment_kernel<<<(unsigned)B, 256, 0, (cudaStream_t)stream>>>(X, lengths, (int)T, (int)D, R, roi_frame_bytes, seed, p_noise,
                                                                   noise_std, p_drop, drop_max, Xo, lengths_out, decisions);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
