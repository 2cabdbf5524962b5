// K6 -- mouth-ROI crop: crop box around the mouth centre, BGR -> gray, resize to 96 x 48.
//
// Replaces crop_roi_gray (live_infer_official.py:171-187, flavor 0: box in double arithmetic, skipped when
// x2 <= x1+2 or y2 <= y1+2, cv2.INTER_AREA) and crop_roi (record_landmarks_official.py:102-118, flavor 1: box in
// float32 arithmetic, skipped when x2 <= x1 or y2 <= y1, cv2 default INTER_LINEAR).  The image arithmetic is
// OpenCV's, reproduced bit for bit (oracle/ref_roi_crop.py is the restatement pinned against cv2):
//   gray   = (3735 B + 19235 G + 9798 R + 2^14) >> 15
//   AREA   both axes shrinking: cell-overlap weights (float32, products and sums in OpenCV's order, no FMA);
//          integer ratios: block sums ((s+2)>>2 for 2x2, rint(s * (1.f/area)) otherwise);
//          an axis growing: the fixed-point bilinear below with cell-overlap fractions
//   LINEAR 11-bit fixed point: rows = S[x0] c0 + S[x1] c1; out = ((b0 (r0>>4))>>16 + (b1 (r1>>4))>>16 + 2) >> 2;
//          x fractions forced to 0 outside the image, y keeps the fraction and clips the row index; exact 2x
//          shrinking is routed to AREA.
// Four CTAs per crop (12 output rows each, so that ~5 CTAs are resident per SM and hide each other's latency): the
// band of gray source rows a CTA needs is staged in shared memory (each source byte read about once), then one
// thread per output pixel.  HBM-bound: 3 x crop area bytes in, 4,608 bytes out.
#include "common.cuh"

namespace ssb {
namespace crop {

constexpr int DW = 96, DH = 48, THREADS = 256;
constexpr int SPLIT = 4, ROWS = DH / SPLIT;                 // a crop is shared by 4 CTAs of 12 output rows each
constexpr int STAGE_MAX = 40 * 1024;                       // gray band staged in smem up to this many pixels (else global reads)

struct Tab {                                               // one axis of the resize
  // area (shrinking): source cell of destination index d: [optional left partial] full cells [optional right partial]
  int s0[DW]; int n[DW]; float a_first[DW], a_mid[DW], a_last[DW]; unsigned char has_first[DW], has_last[DW];
  // linear: two taps and their 11-bit coefficients
  int i0[DW], i1[DW]; int c0[DW], c1[DW];
};

struct Sh {
  int box[4];            // x1, y1, x2, y2
  int ok, mode;          // mode 0: area general, 1: area integer, 2: fixed-point bilinear
  int ix, iy;
  Tab x, y;
};

__device__ __forceinline__ int gray_of(const unsigned char* p) {
  return (3735 * p[0] + 19235 * p[1] + 9798 * p[2] + (1 << 14)) >> 15;
}

// cv::computeResizeAreaTab for one destination index
__device__ void area_entry(Tab& t, int d, int ssize, double scale) {
  const double fsx1 = d * scale, fsx2 = fsx1 + scale;
  const double cell = fmin(scale, ssize - fsx1);
  int sx1 = (int)ceil(fsx1), sx2 = (int)floor(fsx2);
  sx2 = min(sx2, ssize - 1);
  sx1 = min(sx1, sx2);
  const bool first = sx1 - fsx1 > 1e-3, last = fsx2 - sx2 > 1e-3;
  t.has_first[d] = first; t.has_last[d] = last;
  t.a_first[d] = first ? (float)((sx1 - fsx1) / cell) : 0.f;
  t.a_mid[d] = (float)(1.0 / cell);
  t.a_last[d] = last ? (float)(fmin(fmin(fsx2 - sx2, 1.0), cell) / cell) : 0.f;
  t.s0[d] = first ? sx1 - 1 : sx1;
  t.n[d] = (first ? 1 : 0) + (sx2 - sx1) + (last ? 1 : 0);
}

// cv::resize bilinear coefficient for one destination index (x: clamp_frac, y: index clipping only)
__device__ void linear_entry(Tab& t, int d, int ssize, int dsize, bool clamp_frac, bool area_mode) {
  const double inv_scale = (double)dsize / ssize, scale = 1.0 / inv_scale;
  int s; float f;
  if (area_mode) {
    s = (int)floor(d * scale);
    f = (float)((d + 1) - (s + 1) * inv_scale);
    f = f <= 0.f ? 0.f : f - floorf(f);
  } else {
    f = (float)((d + 0.5) * scale - 0.5);
    s = (int)floorf(f);
    f -= (float)s;
  }
  if (clamp_frac) {
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
  }
  t.i0[d] = min(max(s, 0), ssize - 1);
  t.i1[d] = min(max(s + 1, 0), ssize - 1);
  t.c0[d] = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));
  t.c1[d] = __float2int_rn(__fmul_rn(f, 2048.f));
}

__global__ void __launch_bounds__(THREADS) roi_crop_kernel(
    const unsigned char* __restrict__ frames, const int* __restrict__ frame_index, int Hf, int Wf,
    const float* __restrict__ centers, const double* __restrict__ mouth_w, int flavor,
    unsigned char* __restrict__ out, unsigned char* __restrict__ ok_out) {
  extern __shared__ __align__(16) unsigned char smem[];
  Sh& sh = *reinterpret_cast<Sh*>(smem);
  unsigned char* gray_s = smem + sizeof(Sh);
  const int n = blockIdx.x, tid = threadIdx.x;
  const int dy0 = blockIdx.y * ROWS, dy1 = dy0 + ROWS;   // this CTA's output rows
  const unsigned char* frame = frames + (size_t)(frame_index ? frame_index[n] : n) * Hf * Wf * 3;
  unsigned char* o = out + (size_t)n * DW * DH;

  if (tid == 0) {
    int x1, x2, y1, y2, ok;
    if (flavor == 0) {                                     // crop_roi_gray: Python float (double) arithmetic
      const double cx = (double)centers[2 * n], cy = (double)centers[2 * n + 1], mw = mouth_w[n];
      const double hw = 1.2 * mw, hh = 1.0 * mw;
      x1 = (int)fmax(0.0, cx - hw); x2 = (int)fmin((double)Wf, cx + hw);
      y1 = (int)fmax(0.0, cy - hh); y2 = (int)fmin((double)Hf, cy + hh);
      ok = !(x2 <= x1 + 2 || y2 <= y1 + 2);
    } else {                                               // crop_roi: float32 (numpy weak-scalar promotion)
      const float cx = centers[2 * n], cy = centers[2 * n + 1];
      const float hw = (float)(1.2 * mouth_w[n]), hh = (float)(1.0 * mouth_w[n]);
      x1 = (int)fmaxf(0.f, __fsub_rn(cx, hw)); x2 = (int)fminf((float)Wf, __fadd_rn(cx, hw));
      y1 = (int)fmaxf(0.f, __fsub_rn(cy, hh)); y2 = (int)fminf((float)Hf, __fadd_rn(cy, hh));
      ok = !(x2 <= x1 || y2 <= y1);
    }
    sh.box[0] = x1; sh.box[1] = y1; sh.box[2] = x2; sh.box[3] = y2; sh.ok = ok;
    if (ok) {
      const int sw = x2 - x1, shh = y2 - y1;
      const double scale_x = 1.0 / ((double)DW / sw), scale_y = 1.0 / ((double)DH / shh);
      const int ix = (int)floor(scale_x + 0.5), iy = (int)floor(scale_y + 0.5);
      const bool fast = fabs(scale_x - ix) < 2.220446049250313e-16 && fabs(scale_y - iy) < 2.220446049250313e-16;
      int mode;
      if (flavor == 0) mode = (scale_x >= 1.0 && scale_y >= 1.0) ? (fast ? 1 : 0) : 2;
      else mode = (sw == 2 * DW && shh == 2 * DH) ? 1 : 2;
      sh.mode = mode; sh.ix = ix; sh.iy = iy;
    }
    if (ok_out && blockIdx.y == 0) ok_out[n] = (unsigned char)ok;
  }
  __syncthreads();
  if (!sh.ok) {                                            // the callers substitute a zero frame (live:290-293)
    for (int i = tid; i < DW * ROWS / 4; i += THREADS) reinterpret_cast<uint32_t*>(o + dy0 * DW)[i] = 0u;
    return;
  }
  const int x1 = sh.box[0], y1 = sh.box[1], sw = sh.box[2] - sh.box[0], shh = sh.box[3] - sh.box[1];
  const int mode = sh.mode;
  // ---- per-axis tables
  if (tid < DW) {
    if (mode == 0) area_entry(sh.x, tid, sw, 1.0 / ((double)DW / sw));
    else if (mode == 2) linear_entry(sh.x, tid, sw, DW, true, flavor == 0);
  } else if (tid >= 128 && tid < 128 + DH) {
    const int d = tid - 128;
    if (mode == 0) area_entry(sh.y, d, shh, 1.0 / ((double)DH / shh));
    else if (mode == 2) linear_entry(sh.y, d, shh, DH, false, flavor == 0);
  }
  __syncthreads();
  // ---- band of source rows this CTA's output rows read
  int ylo, yhi;
  if (mode == 0) { ylo = sh.y.s0[dy0]; yhi = sh.y.s0[dy1 - 1] + sh.y.n[dy1 - 1]; }
  else if (mode == 1) { ylo = dy0 * sh.iy; yhi = dy1 * sh.iy; }
  else {
    ylo = shh; yhi = 0;
    for (int d = dy0; d < dy1; ++d) { ylo = min(ylo, min(sh.y.i0[d], sh.y.i1[d])); yhi = max(yhi, max(sh.y.i0[d], sh.y.i1[d]) + 1); }
  }
  const int band = yhi - ylo;
  // ---- gray band -> shared memory (rows are contiguous 3*sw-byte runs; a warp takes two rows per pass, a lane four
  // pixels per row: 24 independent byte loads in flight per thread)
  const bool staged = sw * band <= STAGE_MAX;
  if (staged) {
    constexpr int NW = THREADS / 32;
    const int warp = tid >> 5, lane = tid & 31;
    for (int y = warp; y < band; y += 2 * NW) {
      const int yb = y + NW;
      const unsigned char* row_a = frame + ((size_t)(y1 + ylo + y) * Wf + x1) * 3;
      const unsigned char* row_b = frame + ((size_t)(y1 + ylo + min(yb, band - 1)) * Wf + x1) * 3;
      for (int xs = 0; xs < sw; xs += 128) {
        int ga[4], gb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int x = min(xs + lane + 32 * u, sw - 1);
          ga[u] = gray_of(row_a + 3 * x);
          gb[u] = gray_of(row_b + 3 * x);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int x = xs + lane + 32 * u;
          if (x < sw) {
            gray_s[y * sw + x] = (unsigned char)ga[u];
            if (yb < band) gray_s[yb * sw + x] = (unsigned char)gb[u];
          }
        }
      }
    }
  }
  __syncthreads();
  auto G = [&](int y, int x) -> int {
    return staged ? (int)gray_s[(y - ylo) * sw + x] : gray_of(frame + ((size_t)(y1 + y) * Wf + x1 + x) * 3);
  };
  for (int p = dy0 * DW + tid; p < dy1 * DW; p += THREADS) {
    const int dy = p / DW, dx = p % DW;
    int res;
    if (mode == 1) {
      const int ix = sh.ix, iy = sh.iy;
      int s = 0;
      for (int yy = 0; yy < iy; ++yy)
        for (int xx = 0; xx < ix; ++xx) s += G(dy * iy + yy, dx * ix + xx);
      if (ix == 2 && iy == 2) res = (s + 2) >> 2;
      else res = __float2int_rn(__fmul_rn((float)s, __fdiv_rn(1.f, (float)(ix * iy))));
    } else if (mode == 0) {
      float sum = 0.f;
      const int ys0 = sh.y.s0[dy], yn = sh.y.n[dy], xs0 = sh.x.s0[dx], xn = sh.x.n[dx];
      for (int j = 0; j < yn; ++j) {
        const float beta = (j == 0 && sh.y.has_first[dy]) ? sh.y.a_first[dy]
                         : (j == yn - 1 && sh.y.has_last[dy]) ? sh.y.a_last[dy] : sh.y.a_mid[dy];
        float buf = 0.f;
        for (int k = 0; k < xn; ++k) {
          const float alpha = (k == 0 && sh.x.has_first[dx]) ? sh.x.a_first[dx]
                            : (k == xn - 1 && sh.x.has_last[dx]) ? sh.x.a_last[dx] : sh.x.a_mid[dx];
          buf = __fadd_rn(buf, __fmul_rn((float)G(ys0 + j, xs0 + k), alpha));
        }
        sum = j == 0 ? __fmul_rn(beta, buf) : __fadd_rn(sum, __fmul_rn(beta, buf));
      }
      res = __float2int_rn(sum);
    } else {
      const int xa = sh.x.i0[dx], xb = sh.x.i1[dx], ca = sh.x.c0[dx], cb = sh.x.c1[dx];
      const int ya = sh.y.i0[dy], yb = sh.y.i1[dy];
      const int r0 = G(ya, xa) * ca + G(ya, xb) * cb, r1 = G(yb, xa) * ca + G(yb, xb) * cb;
      res = (((sh.y.c0[dy] * (r0 >> 4)) >> 16) + ((sh.y.c1[dy] * (r1 >> 4)) >> 16) + 2) >> 2;
    }
    o[p] = (unsigned char)min(max(res, 0), 255);
  }
}

}  // namespace crop
}  // namespace ssb

// K6: batched mouth-ROI crop (see the header comment).  frames u8 [n_frames, Hf, Wf, 3] (BGR, contiguous);
// frame_index i32 [N] (nullable: crop n reads frame n); centers f32 [N,2] (pixel x,y); mouth_w f64 [N] (the 4th
// value extract_feature returns: mouth width in px for flavor 0 = live, width + 1e-6 for flavor 1 = record);
// out u8 [N,48,96] (zeros where the reference returns None); ok u8 [N] (nullable).
extern "C" int ssb_roi_crop(const uint8_t* frames, const int32_t* frame_index, int64_t Hf, int64_t Wf,
                            const float* centers, const double* mouth_w, int64_t N, int32_t flavor,
                            uint8_t* out, uint8_t* ok, void* stream) {
  using namespace ssb;
  using namespace ssb::crop;
  SSB_REQUIRE(frames && centers && mouth_w && out, "null pointer");
  SSB_REQUIRE(flavor == 0 || flavor == 1, "flavor: 0 = live (crop_roi_gray), 1 = record (crop_roi)");
  SSB_REQUIRE(Hf > 0 && Wf > 0 && Hf * Wf < ((int64_t)1 << 28), "bad frame size");
  SSB_REQUIRE(((uintptr_t)out & 3) == 0, "out must be 4-byte aligned");
  if (N == 0) return SSB_OK;
  const size_t smem = sizeof(Sh) + STAGE_MAX;
  SSB_CUDA(cudaFuncSetAttribute(roi_crop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  roi_crop_kernel<<<dim3((unsigned)N, SPLIT), THREADS, smem, (cudaStream_t)stream>>>(frames, frame_index, (int)Hf, (int)Wf, centers, mouth_w,
                                                                      flavor, out, ok);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
