// K7 -- GRU recurrence for generic (small) hidden sizes, with the time-mean pooling fused.
//
// The reference's earlier model generation, GRUClassifier (inactive/train_reduced.py:129-145; shipped weights
// inactive/word_model_5.pt: input 83, hidden 64, one bidirectional layer, max_t 60):
//     out, _ = nn.GRU(in, H, 1, batch_first=True, bidirectional=True)(x)     # NO packing: all T rows, padding included
//     logits = Linear(2H, C)(out.mean(dim=1))
// K3b (gru_fwd.cu / gru_tc_fwd.cu) is specialised for H = 192 with W_hh spread over a cluster; here W_hh (3H x H fp32)
// fits ONE SM's shared memory for H <= 128 (H = 64: 48 KB), so a CTA keeps it resident, owns a tile of clips for one
// direction and runs all T steps with one barrier per step.  Thread (u, q) owns hidden unit u of clips 8q..8q+7 for all
// three gates (24 accumulators); weights are stored k-major so a warp reads 32 consecutive floats, h rows are warp-wide
// broadcasts.  Gate order (r, z, n), h0 = 0; with `lengths` the packed-sequence semantics of K3b apply (clip b runs
// lengths[b] steps, reverse starts at its last frame, zeros past the end), without it every clip runs T steps.
// pooled[b, d*H + u] = (1/T) * sum_t out[b, t, d*H + u]  (== out.mean(dim=1) of the padded output).
#include "common.cuh"

namespace ssb {
namespace grusmall {

constexpr int CQ = 8;                       // clips per thread
constexpr int BT = 32;                      // clips per CTA

__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void gru_small_fwd_kernel(const float* __restrict__ gi, int64_t gi_ld,
                                     const float* __restrict__ whh_f, const float* __restrict__ bhh_f,
                                     const float* __restrict__ whh_r, const float* __restrict__ bhh_r,
                                     const int64_t* __restrict__ lengths, int B, int T, int H, int ndir,
                                     float* __restrict__ out, float* __restrict__ pooled) {
  extern __shared__ __align__(16) float sm[];
  const int HP = H + 4;                                        // padded h row: 16-byte aligned, conflict-free across clips
  float* wt = sm;                                              // [H][3H]   wt[k][g*H + u] = W_hh[g*H + u][k]
  float* hb = wt + (size_t)H * 3 * H;                          // [2][BT][HP]
  int* slen = reinterpret_cast<int*>(hb + 2 * BT * HP);        // [BT]
  const int tid = threadIdx.x, u = tid % H, q = tid / H;
  const int dir = blockIdx.y, b0 = blockIdx.x * BT;
  const float* __restrict__ whh = dir ? whh_r : whh_f;
  const float* __restrict__ bhh = dir ? bhh_r : bhh_f;
  for (int i = tid; i < 3 * H * H; i += blockDim.x) {
    const int row = i / H, k = i - row * H;                    // coalesced global read, transposed shared write
    wt[k * 3 * H + row] = whh[i];
  }
  for (int i = tid; i < 2 * BT * HP; i += blockDim.x) hb[i] = 0.f;
  if (tid < BT) {
    const int b = b0 + tid;
    slen[tid] = b < B ? (lengths ? max(0, min((int)lengths[b], T)) : T) : 0;
  }
  __syncthreads();
  int len[CQ], maxlen = 0;
#pragma unroll
  for (int c = 0; c < CQ; ++c) len[c] = slen[q * CQ + c];
  for (int i = 0; i < BT; ++i) maxlen = max(maxlen, slen[i]);
  const float br = bhh[u], bz = bhh[H + u], bn = bhh[2 * H + u];
  float h[CQ], hsum[CQ];
#pragma unroll
  for (int c = 0; c < CQ; ++c) { h[c] = 0.f; hsum[c] = 0.f; }
  const int ocol = dir * H + u, ow = ndir * H;

  for (int s = 0; s < maxlen; ++s) {
    const float* hcur = hb + (s & 1) * BT * HP + q * CQ * HP;
    float* hnext = hb + ((s + 1) & 1) * BT * HP + q * CQ * HP;
    float ar[CQ], az[CQ], an[CQ];
#pragma unroll
    for (int c = 0; c < CQ; ++c) { ar[c] = 0.f; az[c] = 0.f; an[c] = 0.f; }
    for (int k = 0; k < H; k += 4) {
      float wr[4], wz[4], wn[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float* wrow = wt + (size_t)(k + j) * 3 * H + u;
        wr[j] = wrow[0]; wz[j] = wrow[H]; wn[j] = wrow[2 * H];
      }
#pragma unroll
      for (int c = 0; c < CQ; ++c) {
        const float4 hv = *reinterpret_cast<const float4*>(hcur + c * HP + k);
        ar[c] = fmaf(hv.x, wr[0], ar[c]); az[c] = fmaf(hv.x, wz[0], az[c]); an[c] = fmaf(hv.x, wn[0], an[c]);
        ar[c] = fmaf(hv.y, wr[1], ar[c]); az[c] = fmaf(hv.y, wz[1], az[c]); an[c] = fmaf(hv.y, wn[1], an[c]);
        ar[c] = fmaf(hv.z, wr[2], ar[c]); az[c] = fmaf(hv.z, wz[2], az[c]); an[c] = fmaf(hv.z, wn[2], an[c]);
        ar[c] = fmaf(hv.w, wr[3], ar[c]); az[c] = fmaf(hv.w, wz[3], az[c]); an[c] = fmaf(hv.w, wn[3], an[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < CQ; ++c) {
      if (s < len[c]) {
        const int b = b0 + q * CQ + c;
        const int t = dir ? (len[c] - 1 - s) : s;
        const float* g = gi + ((int64_t)b * T + t) * gi_ld + (int64_t)dir * 3 * H + u;
        const float r = sigm(g[0] + ar[c] + br);
        const float z = sigm(g[H] + az[c] + bz);
        const float n = tanhf(g[2 * H] + r * (an[c] + bn));
        h[c] = (1.0f - z) * n + z * h[c];
        hsum[c] += h[c];
        if (out) out[((int64_t)b * T + t) * ow + ocol] = h[c];
      }
      hnext[c * HP + u] = h[c];
    }
    __syncthreads();
  }
#pragma unroll
  for (int c = 0; c < CQ; ++c) {
    const int b = b0 + q * CQ + c;
    if (b < B) {
      if (pooled) pooled[(int64_t)b * ow + ocol] = hsum[c] / (float)T;
      if (out)
        for (int t = len[c]; t < T; ++t) out[((int64_t)b * T + t) * ow + ocol] = 0.f;      // pad_packed_sequence zeros
    }
  }
}

}  // namespace grusmall
}  // namespace ssb

// gi f32 [B,T,gi_ld]: x W_ih^T + b_ih for all steps, direction d in columns [d*3H, (d+1)*3H), gate order (r,z,n).
// whh_* [3H,H], bhh_* [3H]; whh_r / bhh_r ignored when ndir == 1.  lengths i64 [B] or NULL (every clip runs T steps,
// GRUClassifier's un-packed call).  out f32 [B,T,ndir*H] or NULL; pooled f32 [B,ndir*H] or NULL (time mean over T).
extern "C" int ssb_gru_small_fwd(const float* gi, int64_t gi_ld, const float* whh_f, const float* bhh_f,
                                 const float* whh_r, const float* bhh_r, const int64_t* lengths,
                                 int64_t B, int64_t T, int64_t H, int32_t ndir, float* out, float* pooled, void* stream) {
  using namespace ssb;
  using namespace ssb::grusmall;
  SSB_REQUIRE(gi && whh_f && bhh_f, "null pointer");
  SSB_REQUIRE(ndir == 1 || ndir == 2, "ndir must be 1 or 2");
  SSB_REQUIRE(ndir == 1 || (whh_r && bhh_r), "reverse-direction weights missing");
  SSB_REQUIRE(out || pooled, "nothing to write");
  SSB_REQUIRE(H >= 32 && H <= 128 && H % 32 == 0, "hidden size must be 32, 64, 96 or 128 (W_hh resident in one SM); H = 192 is K3b");
  SSB_REQUIRE(gi_ld >= ndir * 3 * H, "gi_ld too small");
  SSB_REQUIRE(T > 0 && B >= 0, "bad sizes");
  if (B == 0) return SSB_OK;
  const size_t smem = ((size_t)3 * H * H + 2 * BT * (H + 4)) * sizeof(float) + BT * sizeof(int);
  SSB_CUDA(cudaFuncSetAttribute(gru_small_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)cdiv(B, BT), (unsigned)ndir);
  gru_small_fwd_kernel<<<grid, (unsigned)(H * (BT / CQ)), smem, (cudaStream_t)stream>>>(
      gi, gi_ld, whh_f, bhh_f, whh_r, bhh_r, lengths, (int)B, (int)T, (int)H, ndir, out, pooled);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Mouth-open detector of important_landmarks.py:130-144 for many streams: per frame with a face
//     openness = |y[14] - y[13]| / (dist2d(p[33], p[263]) + 1e-6)          (normalised landmark coordinates)
//     ema      = (1 - alpha) * ema + alpha * openness                       (EMA_ALPHA = 0.25)
//     open     = open ? !(ema < close_thr) : (ema > open_thr)               (hysteresis, both 0.02 in the reference)
// in double precision like the reference's Python floats.  A frame without a face leaves the state untouched
// (important_landmarks.py:127).  The scan over a stream's frames is sequential, streams are independent: one thread
// per stream.
namespace ssb {
namespace mouth {
__global__ void mouth_open_kernel(const float* __restrict__ lm, const uint8_t* __restrict__ valid, int S, int F, int n_lm,
                                  int i_top, int i_bot, int i_le, int i_re, double alpha, double open_thr, double close_thr,
                                  double* __restrict__ ema_state, uint8_t* __restrict__ open_state,
                                  double* __restrict__ ema_out, uint8_t* __restrict__ open_out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  double ema = ema_state[s];
  bool open = open_state[s] != 0;
  for (int f = 0; f < F; ++f) {
    const int64_t fr = (int64_t)s * F + f;
    if (!valid || valid[fr]) {
      const float* p = lm + fr * n_lm * 3;
      const double lip_gap = fabs((double)p[i_bot * 3 + 1] - (double)p[i_top * 3 + 1]);
      const double dx = (double)p[i_le * 3] - (double)p[i_re * 3], dy = (double)p[i_le * 3 + 1] - (double)p[i_re * 3 + 1];
      const double openness = lip_gap / (sqrt(dx * dx + dy * dy) + 1e-6);
      ema = (1.0 - alpha) * ema + alpha * openness;
      if (open) { if (ema < close_thr) open = false; }
      else      { if (ema > open_thr) open = true; }
    }
    if (ema_out) ema_out[fr] = ema;
    if (open_out) open_out[fr] = open ? 1 : 0;
  }
  ema_state[s] = ema;
  open_state[s] = open ? 1 : 0;
}
}  // namespace mouth
}  // namespace ssb

// lm f32 [S,F,n_lm,3]; valid u8 [S,F] or NULL; ema_state f64 [S] and open_state u8 [S] are read and updated;
// ema_out f64 [S,F] / open_out u8 [S,F] (nullable) receive the per-frame values the reference overlays on the video.
extern "C" int ssb_mouth_open_ema(const float* lm, const uint8_t* valid, int64_t S, int64_t F, int64_t n_lm,
                                  double alpha, double open_thr, double close_thr, double* ema_state, uint8_t* open_state,
                                  double* ema_out, uint8_t* open_out, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(lm && ema_state && open_state, "null pointer");
  SSB_REQUIRE(n_lm > 263, "landmark rows must include ids 13, 14, 33 and 263 (important_landmarks.py:49-55)");
  SSB_REQUIRE(S >= 0 && F >= 0, "bad sizes");
  if (S == 0) return SSB_OK;
  mouth::mouth_open_kernel<<<(unsigned)cdiv(S, 128), 128, 0, (cudaStream_t)stream>>>(
      lm, valid, (int)S, (int)F, (int)n_lm, 13, 14, 33, 263, alpha, open_thr, close_thr, ema_state, open_state, ema_out, open_out);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
