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
// cross-frame dependency (velocity, previous used frame -> t) never leaves the warp; the
// last detected frame before the run is recomputed (an L2 hit: the neighbouring warp streams it).
// Per valid frame the compulsory traffic is K*8 B gathered + (2K+4)*4 B written
// = 1,424 B at K = 88; at the (x,y,z) stride of 12 B the 88 points touch 4.2 KB of 32-byte
// sectors of the 5,736-byte landmark row, and that sector traffic is what the kernel is bound by
// (ncu: 6.3 TB/s of DRAM traffic, profiles/r2a_features_ncu.txt).
//
// Velocity history (live_infer_official.py:272-296, record_landmarks_official.py:178-201): a frame with NO
// detected face leaves prev_xy untouched (the reset sits inside `if res.face_landmarks:`); a detected face outside
// the 60..150 px band resets it.  So the predecessor of frame t is the last DETECTED frame before t, and it counts
// only if it was inside the band.
#include "common.cuh"

namespace ssb {

constexpr int kMaxPtsPerLane = 4;      // K <= 128

struct FrameGeom {
  float cx, cy;       // subset mean, pixels
  float mouth_w;      // |p61 - p291|, pixels (float32 arithmetic, the live flavour)
  float open_px;      // |p13 - p14|
  bool usable;        // present && inside the distance band
};

// (x, y) of one landmark: one 8-byte load when the landmark stride keeps (x, y) 8-byte aligned (kVec2: even stride and
// an 8-byte aligned tensor), two 4-byte loads otherwise (stride 3 = MediaPipe's x,y,z).  A compile-time switch: a
// run-time branch around every load kept the compiler from batching a frame's loads (0.31 -> 0.52 ms per 4096 clips).
template <bool kVec2>
__device__ __forceinline__ float2 ld_xy(const float* __restrict__ f, int64_t idx, int lm_stride) {
  const float* p = f + idx * lm_stride;
  if (kVec2) return __ldg(reinterpret_cast<const float2*>(p));
  return make_float2(__ldg(p), __ldg(p + 1));
}

__device__ __forceinline__ float dist2f(float ax, float ay, float bx, float by) {
  float dx = ax - bx, dy = ay - by;
  return sqrtf(dx * dx + dy * dy);
}

// Loads one frame: gathers this lane's points (pixel units) and the anchors.
template <bool kVec2>
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
      const float2 v = ld_xy<kVec2>(f, __ldg(ids + k), lm_stride);
      px[j] = v.x * w; py[j] = v.y * h;
      sx += px[j]; sy += py[j];
    }
  }
  // anchors: every lane reads the same 4 landmarks (one broadcast transaction each)
  const float2 a = ld_xy<kVec2>(f, anchors.x, lm_stride);
  const float2 b = ld_xy<kVec2>(f, anchors.y, lm_stride);
  const float2 u = ld_xy<kVec2>(f, anchors.z, lm_stride);
  const float2 l = ld_xy<kVec2>(f, anchors.w, lm_stride);
  FrameGeom g;
  g.mouth_w = dist2f(a.x * w, a.y * h, b.x * w, b.y * h);
  g.open_px = dist2f(u.x * w, u.y * h, l.x * w, l.y * h);
  g.cx = warp_sum(sx) / (float)K;
  g.cy = warp_sum(sy) / (float)K;
  g.usable = present && g.mouth_w >= gate_lo && g.mouth_w <= gate_hi;
  return g;
}

// Pass 0 of compact mode: usable[b,t] = detected && mouth width inside the band (two landmarks per frame), so that the
// main kernel finds a run's first output row by summing T bytes instead of re-deriving every earlier frame.
__global__ void __launch_bounds__(256)
features_valid_kernel(const float* __restrict__ lm, int64_t frames, int n_lm, int lm_stride, int4 anchors,
                      const uint8_t* __restrict__ present, float w, float h, float gate_lo, float gate_hi,
                      uint8_t* __restrict__ valid) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= frames) return;
  bool ok = !present || present[i];
  if (ok) {
    const float* f = lm + i * n_lm * lm_stride;
    const float2 a = ld_xy<false>(f, anchors.x, lm_stride), c = ld_xy<false>(f, anchors.y, lm_stride);
    const float mw = dist2f(a.x * w, a.y * h, c.x * w, c.y * h);
    ok = mw >= gate_lo && mw <= gate_hi;
  }
  valid[i] = ok ? 1 : 0;
}

template <bool kVec2>
__global__ void __launch_bounds__(128)
features_kernel(const float* __restrict__ lm, int B, int T, int n_lm, int lm_stride,
                const int* __restrict__ ids, int K, int4 anchors,
                const uint8_t* __restrict__ present, float w, float h,
                float gate_lo, float gate_hi, int compact, int frames_per_warp,
                float* __restrict__ X, float* __restrict__ center, float* __restrict__ mouth_w,
                uint8_t* __restrict__ valid, int64_t* __restrict__ lengths, float* __restrict__ xy_px) {
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

  // Output row of frame t0 when compacting = number of usable frames before it (pass 0 wrote the flags).
  int row = t0;
  if (compact) {
    int cnt = 0;
    for (int t = lane; t < t0; t += 32) cnt += valid[(int64_t)b * T + t];
    row = __reduce_add_sync(0xffffffffu, cnt);
  }

  float qx[kMaxPtsPerLane], qy[kMaxPtsPerLane];     // previous used frame, normalised
  bool have_prev = false;
  // last DETECTED frame before the run (absent frames are transparent): one coalesced byte load + a ballot per 32 frames
  int tp = pres ? -1 : t0 - 1;
  if (pres) {
    for (int base = t0 - 1; base >= 0 && tp < 0; base -= 32) {
      const int t = base - lane;
      const unsigned m = __ballot_sync(0xffffffffu, t >= 0 && pres[t] != 0);
      if (m) tp = base - (__ffs(m) - 1);
    }
  }
  if (tp >= 0) {
    float px[kMaxPtsPerLane], py[kMaxPtsPerLane];
    const int t = tp;
    FrameGeom g = load_frame<kVec2>(clip + (int64_t)t * n_lm * lm_stride, lm_stride, ids, K, lane, anchors,
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
    FrameGeom g = load_frame<kVec2>(clip + (int64_t)t * n_lm * lm_stride, lm_stride, ids, K, lane, anchors,
                             w, h, !pres || pres[t], gate_lo, gate_hi, px, py);
    const bool detected = !pres || pres[t];
    if (valid && !compact && lane == 0) valid[(int64_t)b * T + t] = g.usable ? 1 : 0;
    if (!g.usable) {
      if (detected) have_prev = false;                // live_infer_official.py:294-296: only a detected, out-of-band face resets
      if (!compact) {
        float* xr = X + ((int64_t)b * T + t) * D;
        for (int i = lane; i < D; i += 32) xr[i] = 0.f;
        if (xy_px) for (int i = lane; i < 2 * K; i += 32) xy_px[((int64_t)b * T + t) * 2 * K + i] = 0.f;
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
        if (xy_px) *reinterpret_cast<float2*>(xy_px + (orow * K + k) * 2) = make_float2(px[j], py[j]);   // the gather itself (live:147)
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
      if (xy_px) for (int i = lane; i < 2 * K; i += 32) xy_px[((int64_t)b * T + r) * 2 * K + i] = 0.f;
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
                                int64_t* lengths, float* xy_px, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(landmarks && ids && X && anchor_ids, "null pointer");
  SSB_REQUIRE(B >= 0 && T >= 1 && n_lm >= 1, "bad shape");
  SSB_REQUIRE(lm_stride >= 2, "landmark stride must hold at least x,y");
  SSB_REQUIRE(K >= 1 && K <= 32 * kMaxPtsPerLane, "K must be in 1..128");
  for (int i = 0; i < 4; ++i) SSB_REQUIRE(anchor_ids[i] >= 0 && anchor_ids[i] < n_lm, "anchor id out of range");
  SSB_REQUIRE(!compact || valid, "compact mode needs the valid buffer (pass 0 writes the flags it counts)");
  if (B == 0) return SSB_OK;
  int4 anchors = make_int4(anchor_ids[0], anchor_ids[1], anchor_ids[2], anchor_ids[3]);
  if (compact) {
    const int64_t frames = B * T;
    features_valid_kernel<<<(unsigned)cdiv(frames, 256), 256, 0, (cudaStream_t)stream>>>(
        landmarks, frames, (int)n_lm, (int)lm_stride, anchors, present, frame_w, frame_h, gate_min_px, gate_max_px, valid);
    SSB_LAUNCH_CHECK();
  }
  const int frames_per_warp = 6;
  const int64_t warps = B * ((T + frames_per_warp - 1) / frames_per_warp);
  const int warps_per_block = 4;
  dim3 grid((unsigned)cdiv(warps, warps_per_block)), block(32 * warps_per_block);
  const bool vec2 = (lm_stride & 1) == 0 && (reinterpret_cast<uintptr_t>(landmarks) & 7) == 0;
  auto kern = vec2 ? features_kernel<true> : features_kernel<false>;
  kern<<<grid, block, 0, (cudaStream_t)stream>>>(
      landmarks, (int)B, (int)T, (int)n_lm, (int)lm_stride, ids, K, anchors, present, frame_w, frame_h,
      gate_min_px, gate_max_px, compact, frames_per_warp, X, center, mouth_w, valid, lengths, xy_px);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
