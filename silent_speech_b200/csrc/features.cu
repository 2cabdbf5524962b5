// K1 -- per-frame landmark features (HBM-bound gather + small reductions).
//
// Replaces, for a whole batch of clips at once, the per-frame Python of
//   extract_feature   live_infer_official.py:146-169 (== record_landmarks_official.py:62-100)
//   mouth_width_px    live_infer_official.py:141-144
// and the capture loop's gating / velocity-reset policy
//   live_infer_official.py:275-296.
//
// Layout: landmarks f32 [B,T,n_lm,lm_stride] (x,y first), ids i32 [K] sorted.
// One warp owns a run of `frames_per_warp` consecutive frames of one clip and
// carries the previous frame's normalised points in registers, so the only
// cross-frame dependency (velocity, t-1 -> t) never leaves the warp; the frame
// just before the run is recomputed (an L2 hit: the neighbouring warp streams it).
// Per valid frame the compulsory traffic is K*8 B gathered + (2K+4)*4 B written
// = 1,424 B at K = 88.
#include "common.cuh"

namespace ssb {

constexpr int kMaxPtsPerLane = 4;      // K <= 128

struct FrameGeom {
  float cx, cy;       // subset mean, pixels
  float mouth_w;      // |p61 - p291|, pixels (float32 arithmetic, the live flavour)
  float open_px;      // |p13 - p14|
  bool usable;        // present && inside the distance band
};

// (x, y) of one landmark; two 4-byte loads so any landmark stride (3 for x,y,z) is legal.
__device__ __forceinline__ float2 ld_xy(const float* __restrict__ f, int64_t idx, int lm_stride) {
  const float* p = f + idx * lm_stride;
  return make_float2(__ldg(p), __ldg(p + 1));
}

__device__ __forceinline__ float dist2f(float ax, float ay, float bx, float by) {
  float dx = ax - bx, dy = ay - by;
  return sqrtf(dx * dx + dy * dy);
}

// Loads one frame: gathers this lane's points (pixel units) and the anchors.
__device__ __forceinline__ FrameGeom load_frame(const float* __restrict__ f, int lm_stride,
                                                const int* __restrict__ ids, int K, int lane,
                                                int4 anchors, float w, float h, bool present,
                                                float gate_lo, float gate_hi,
                                                float (&px)[kMaxPtsPerLane], float (&py)[kMaxPtsPerLane]) {
  float sx = 0.f, sy = 0.f;
#pragma unroll
  for (int j = 0; j < kMaxPtsPerLane; ++j) {
    int k = lane + 32 * j;
    px[j] = 0.f; py[j] = 0.f;
    if (k < K) {
      const float2 v = ld_xy(f, __ldg(ids + k), lm_stride);
      px[j] = v.x * w; py[j] = v.y * h;
      sx += px[j]; sy += py[j];
    }
  }
  // anchors: every lane reads the same 4 landmarks (one broadcast transaction each)
  const float2 a = ld_xy(f, anchors.x, lm_stride);
  const float2 b = ld_xy(f, anchors.y, lm_stride);
  const float2 u = ld_xy(f, anchors.z, lm_stride);
  const float2 l = ld_xy(f, anchors.w, lm_stride);
  FrameGeom g;
  g.mouth_w = dist2f(a.x * w, a.y * h, b.x * w, b.y * h);
  g.open_px = dist2f(u.x * w, u.y * h, l.x * w, l.y * h);
  g.cx = warp_sum(sx) / (float)K;
  g.cy = warp_sum(sy) / (float)K;
  g.usable = present && g.mouth_w >= gate_lo && g.mouth_w <= gate_hi;
  return g;
}

__global__ void __launch_bounds__(128)
features_kernel(const float* __restrict__ lm, int B, int T, int n_lm, int lm_stride,
                const int* __restrict__ ids, int K, int4 anchors,
                const uint8_t* __restrict__ present, float w, float h,
                float gate_lo, float gate_hi, int compact, int frames_per_warp,
                float* __restrict__ X, float* __restrict__ center, float* __restrict__ mouth_w,
                uint8_t* __restrict__ valid, int64_t* __restrict__ lengths) {
  const int lane = threadIdx.x & 31;
  const int runs_per_clip = (T + frames_per_warp - 1) / frames_per_warp;
  const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (warp_id >= (int64_t)B * runs_per_clip) return;
  const int b = (int)(warp_id / runs_per_clip);
  const int t0 = (int)(warp_id % runs_per_clip) * frames_per_warp;
  const int t1 = min(T, t0 + frames_per_warp);
  const int D = 2 * K + 4;
  const float* clip = lm + (int64_t)b * T * n_lm * lm_stride;
  const uint8_t* pres = present ? present + (int64_t)b * T : nullptr;

  // Output row of frame t0 when compacting = number of usable frames before it.
  int row = t0;
  if (compact) {
    int cnt = 0;
    for (int t = lane; t < t0; t += 32) {
      const float* f = clip + (int64_t)t * n_lm * lm_stride;
      const float2 a = ld_xy(f, anchors.x, lm_stride);
      const float2 c = ld_xy(f, anchors.y, lm_stride);
      float mw = dist2f(a.x * w, a.y * h, c.x * w, c.y * h);
      bool ok = (!pres || pres[t]) && mw >= gate_lo && mw <= gate_hi;
      cnt += ok ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    row = cnt;
  }

  float qx[kMaxPtsPerLane], qy[kMaxPtsPerLane];     // previous frame, normalised
  bool have_prev = false;
  if (t0 > 0) {
    float px[kMaxPtsPerLane], py[kMaxPtsPerLane];
    const int t = t0 - 1;
    FrameGeom g = load_frame(clip + (int64_t)t * n_lm * lm_stride, lm_stride, ids, K, lane, anchors,
                             w, h, !pres || pres[t], gate_lo, gate_hi, px, py);
    if (g.usable) {
      const float scale = (float)((double)g.mouth_w + 1e-6);
#pragma unroll
      for (int j = 0; j < kMaxPtsPerLane; ++j) { qx[j] = (px[j] - g.cx) / scale; qy[j] = (py[j] - g.cy) / scale; }
      have_prev = true;
    }
  }

  for (int t = t0; t < t1; ++t) {
    float px[kMaxPtsPerLane], py[kMaxPtsPerLane];
    FrameGeom g = load_frame(clip + (int64_t)t * n_lm * lm_stride, lm_stride, ids, K, lane, anchors,
                             w, h, !pres || pres[t], gate_lo, gate_hi, px, py);
    if (valid && lane == 0) valid[(int64_t)b * T + t] = g.usable ? 1 : 0;
    if (!g.usable) {
      have_prev = false;                              // live_infer_official.py:294-296
      if (!compact) {
        float* xr = X + ((int64_t)b * T + t) * D;
        for (int i = lane; i < D; i += 32) xr[i] = 0.f;
        if (lane == 0) {
          if (center) { center[((int64_t)b * T + t) * 2] = 0.f; center[((int64_t)b * T + t) * 2 + 1] = 0.f; }
          if (mouth_w) mouth_w[(int64_t)b * T + t] = 0.f;
        }
      }
      continue;
    }
    const float scale = (float)((double)g.mouth_w + 1e-6);        // live:151
    float vel_part = 0.f;
    const int64_t orow = (int64_t)b * T + (compact ? row : t);
    float* xr = X + orow * D;
#pragma unroll
    for (int j = 0; j < kMaxPtsPerLane; ++j) {
      const int k = lane + 32 * j;
      const float nx = (px[j] - g.cx) / scale, ny = (py[j] - g.cy) / scale;   // live:153
      if (k < K) {
        *reinterpret_cast<float2*>(xr + 2 * k) = make_float2(nx, ny);
        if (have_prev) vel_part += dist2f(nx, ny, qx[j], qy[j]);
      }
      qx[j] = nx; qy[j] = ny;
    }
    const float vel = have_prev ? warp_sum(vel_part) / (float)K : 0.f;        // live:155-158
    if (lane == 0) {
      const float aspect = (float)((double)g.open_px / ((double)g.mouth_w + 1e-6));   // live:163
      xr[2 * K] = vel; xr[2 * K + 1] = g.open_px; xr[2 * K + 2] = g.mouth_w; xr[2 * K + 3] = aspect;   // live:167
      if (center) { center[orow * 2] = g.cx; center[orow * 2 + 1] = g.cy; }
      if (mouth_w) mouth_w[orow] = g.mouth_w;
    }
    have_prev = true;
    ++row;
  }

  if (compact && t1 == T) {
    // the warp owning the clip's last run knows the total; it zero-fills the tail rows
    if (lengths && lane == 0) lengths[b] = row;
    for (int r = row; r < T; ++r) {
      float* xr = X + ((int64_t)b * T + r) * D;
      for (int i = lane; i < D; i += 32) xr[i] = 0.f;
      if (lane == 0) {
        if (center) { center[((int64_t)b * T + r) * 2] = 0.f; center[((int64_t)b * T + r) * 2 + 1] = 0.f; }
        if (mouth_w) mouth_w[(int64_t)b * T + r] = 0.f;
      }
    }
  }
}

}  // namespace ssb

extern "C" int ssb_features_fwd(const float* landmarks, int64_t B, int64_t T, int64_t n_lm, int64_t lm_stride,
                                const int32_t* ids, int32_t K, const int32_t anchor_ids[4],
                                const uint8_t* present, float frame_w, float frame_h,
                                float gate_min_px, float gate_max_px, int32_t compact,
                                float* X, float* center, float* mouth_w, uint8_t* valid,
                                int64_t* lengths, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(landmarks && ids && X && anchor_ids, "null pointer");
  SSB_REQUIRE(B >= 0 && T >= 1 && n_lm >= 1, "bad shape");
  SSB_REQUIRE(lm_stride >= 2, "landmark stride must hold at least x,y");
  SSB_REQUIRE(K >= 1 && K <= 32 * kMaxPtsPerLane, "K must be in 1..128");
  for (int i = 0; i < 4; ++i) SSB_REQUIRE(anchor_ids[i] >= 0 && anchor_ids[i] < n_lm, "anchor id out of range");
  if (B == 0) return SSB_OK;
  const int frames_per_warp = 6;
  const int64_t warps = B * ((T + frames_per_warp - 1) / frames_per_warp);
  const int warps_per_block = 4;
  dim3 grid((unsigned)cdiv(warps, warps_per_block)), block(32 * warps_per_block);
  int4 anchors = make_int4(anchor_ids[0], anchor_ids[1], anchor_ids[2], anchor_ids[3]);
  features_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(
      landmarks, (int)B, (int)T, (int)n_lm, (int)lm_stride, ids, K, anchors, present, frame_w, frame_h,
      gate_min_px, gate_max_px, compact, frames_per_warp, X, center, mouth_w, valid, lengths);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
