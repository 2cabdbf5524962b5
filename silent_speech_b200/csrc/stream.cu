// Sliding-window state for concurrent camera streams (windowing policy of
// inactive/live_feed.py:155-213: a deque(maxlen=max_t) per stream, every frame appended, the window zero-padded
// to max_t and classified every PRED_EVERY-th frame once WARMUP_MIN frames are buffered).
//
// Device state: one ring of rows per stream, ring[S][W][D] (D = 180 point features + 32 ROI-CNN embedding: the CNN
// is frame-local, so a frame's embedding is computed once, when the frame arrives, and re-used by every window that
// contains it).  The host keeps the tiny integer state (head / count / frame counter per stream) and passes
// positions; these two kernels move the rows.  Pure data movement: bit-exact by construction, HBM-bound.
#include "common.cuh"

namespace ssb {
namespace stream {

// ring[sid[i]][pos[i]][:] = [x[i][0:dx] | e[i][0:de]]
__global__ void ring_push_kernel(float* __restrict__ ring, int W, int D, const int32_t* __restrict__ sid,
                                 const int32_t* __restrict__ pos, const float* __restrict__ x, int dx,
                                 const float* __restrict__ e, int de, int n) {
  const int i = blockIdx.x;
  if (i >= n) return;
  float* dst = ring + ((size_t)sid[i] * W + pos[i]) * D;
  for (int c = threadIdx.x; c < dx; c += blockDim.x) dst[c] = x[(size_t)i * dx + c];
  if (e) for (int c = threadIdx.x; c < de; c += blockDim.x) dst[dx + c] = e[(size_t)i * de + c];
}

// window i = rows start[i], start[i]+1, ... (mod W), count[i] of them, oldest first; zero rows after that.
// Xw [m][T][dx], Ew [m][T][de] (nullable), lengths[i] = count[i].
__global__ void ring_gather_kernel(const float* __restrict__ ring, int W, int D, const int32_t* __restrict__ sid,
                                   const int32_t* __restrict__ start, const int32_t* __restrict__ count,
                                   float* __restrict__ Xw, int dx, float* __restrict__ Ew, int de, int T,
                                   int64_t* __restrict__ lengths) {
  const int i = blockIdx.x, t = blockIdx.y;
  const int cnt = min(count[i], T);
  if (t == 0 && threadIdx.x == 0) lengths[i] = cnt;
  float* xo = Xw + ((size_t)i * T + t) * dx;
  float* eo = Ew ? Ew + ((size_t)i * T + t) * de : nullptr;
  if (t < cnt) {
    const float* src = ring + ((size_t)sid[i] * W + (start[i] + t) % W) * D;
    for (int c = threadIdx.x; c < dx; c += blockDim.x) xo[c] = src[c];
    if (eo) for (int c = threadIdx.x; c < de; c += blockDim.x) eo[c] = src[dx + c];
  } else {
    for (int c = threadIdx.x; c < dx; c += blockDim.x) xo[c] = 0.f;
    if (eo) for (int c = threadIdx.x; c < de; c += blockDim.x) eo[c] = 0.f;
  }
}

}  // namespace stream
}  // namespace ssb

extern "C" int ssb_ring_push(float* ring, int64_t W, int64_t D, const int32_t* sid, const int32_t* pos,
                             const float* x, int64_t dx, const float* e, int64_t de, int64_t n, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(ring && sid && pos && x, "null pointer");
  SSB_REQUIRE(dx + (e ? de : 0) <= D, "row wider than the ring");
  if (n == 0) return SSB_OK;
  stream::ring_push_kernel<<<(unsigned)n, 64, 0, (cudaStream_t)stream>>>(ring, (int)W, (int)D, sid, pos, x, (int)dx, e, (int)de, (int)n);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

extern "C" int ssb_ring_gather(const float* ring, int64_t W, int64_t D, const int32_t* sid, const int32_t* start,
                               const int32_t* count, int64_t m, float* Xw, int64_t dx, float* Ew, int64_t de,
                               int64_t T, int64_t* lengths, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(ring && sid && start && count && Xw && lengths, "null pointer");
  SSB_REQUIRE(dx + (Ew ? de : 0) <= D && T > 0 && T <= 65535, "bad shape");
  if (m == 0) return SSB_OK;
  dim3 grid((unsigned)m, (unsigned)T);
  stream::ring_gather_kernel<<<grid, 64, 0, (cudaStream_t)stream>>>(ring, (int)W, (int)D, sid, start, count, Xw, (int)dx, Ew, (int)de,
                                                                  (int)T, lengths);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
