// K3-bwd (fp32 build) -- back-propagation through time of one bidirectional GRU layer.
//
// Backward of the packed-sequence nn.GRU of train_model_official.py:261-267 / 301-305 (what
// loss.backward() at train:437 runs inside ATen).  With the gates saved by the forward kernel
// (r, z, n, hn = W_hn h + b_hn) one step of direction d at frame t is
//   dh   = dh_carry + dout[t]
//   dn   = dh (1-z)           dz = dh (h_prev - n)
//   dn~  = dn (1-n^2)         dz~ = dz z(1-z)        dr~ = dn~ hn r(1-r)
//   dgi[t]  = (dr~, dz~, dn~)                  -> feeds dW_ih, db_ih, d(input) as GEMMs
//   dgh[t]  = (dr~, dz~, dn~ r)                -> feeds dW_hh, db_hh as GEMMs (with h_prev[t])
//   dh_carry = dh z + dgh[t] W_hh              <- the only serial part, done here
// Work split: one CTA = BT clips x one direction, 8 BT threads (BT = 64, 16 or 8, picked so that the grid covers
// the SMs: the serial product is FMA-issue-bound per CTA, so small batches want MANY small tiles), thread
// (cg, ug) owns units 3ug..3ug+2 of clips 8cg..8cg+7; dh stays in registers; dgh of the step is
// shared through smem; W_hh^T streams from L2 in k chunks (cp.async double buffer).
#include "common.cuh"

namespace ssb {
namespace gru {

constexpr int H = kHidden, G3 = 3 * H;
constexpr int KC = 16, WLD = KC + 4;
constexpr int DLD = G3 + 4;            // padded dgh row
constexpr int NCHUNK_B = G3 / KC;      // 36

template <int BT>
struct SmemB {
  float dgh[BT][DLD];
  float w[2][H][WLD];
  int len[BT];
  int maxlen;
};

__device__ __forceinline__ void cp_async16b(void* smem, const void* gmem) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}

// whhT: [H][G3] (transposed recurrent weights): chunk = rows j (192) x 16 k
template <int THREADS>
__device__ __forceinline__ void stage_wT(float (*dst)[WLD], const float* __restrict__ whhT, int chunk, int tid) {
  for (int p = tid; p < H * (KC / 4); p += THREADS) {
    const int row = p >> 2, q = p & 3;
    cp_async16b(&dst[row][q * 4], whhT + (int64_t)row * G3 + chunk * KC + q * 4);
  }
}

template <int BT>
__global__ void __launch_bounds__(BT * 8, 1)
gru_rec_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ gates,
                   const float* __restrict__ whhT_f, const float* __restrict__ whhT_r,
                   const int64_t* __restrict__ lengths, int B, int T,
                   float* __restrict__ dgi, float* __restrict__ dghn, float* __restrict__ hprev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int THREADS = BT * 8;
  SmemB<BT>& s = *reinterpret_cast<SmemB<BT>*>(smem_raw);
  const int tid = threadIdx.x, dir = blockIdx.y, b0 = blockIdx.x * BT;
  const float* __restrict__ whhT = dir ? whhT_r : whhT_f;
  const int ug = tid & 63, cg = tid >> 6, j0 = ug * 3, c0 = cg * 8;

  if (tid < BT) {
    const int b = b0 + tid;
    int l = b < B ? (int)lengths[b] : 0;
    s.len[tid] = max(0, min(l, T));
  }
  if (tid == 0) s.maxlen = 0;
  __syncthreads();
  if (tid < BT) atomicMax(&s.maxlen, s.len[tid]);
  __syncthreads();
  const int maxlen = s.maxlen;

  float dh[3][8];
#pragma unroll
  for (int u = 0; u < 3; ++u)
#pragma unroll
    for (int c = 0; c < 8; ++c) dh[u][c] = 0.f;

  for (int step = maxlen - 1; step >= 0; --step) {
    // ---- element-wise part: this thread's (clip, unit) pairs
    float dhz[3][8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const int cl = c0 + c, b = b0 + cl, len = s.len[cl];
      const bool act = step < len;
      const int t = dir ? (len - 1 - step) : step;
      const int tp = dir ? t + 1 : t - 1;
      const int64_t bt = (int64_t)b * T + t;
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        float g_r = 0.f, g_z = 0.f, g_n = 0.f;
        dhz[u][c] = 0.f;
        if (act) {
          const int j = j0 + u;
          const float* gs = gates + (bt * 2 + dir) * (4 * H) + j;
          const float r = gs[0], z = gs[H], n = gs[2 * H], hn = gs[3 * H];
          const float hp = step > 0 ? out[((int64_t)b * T + tp) * (2 * H) + dir * H + j] : 0.f;
          const float d = dh[u][c] + dout[bt * (2 * H) + dir * H + j];
          const float dn = d * (1.f - z), dz = d * (hp - n);
          const float dnp = dn * (1.f - n * n), dzp = dz * z * (1.f - z);
          const float drp = dnp * hn * r * (1.f - r);
          float* gi_row = dgi + bt * (2 * G3) + dir * G3 + j;
          gi_row[0] = drp; gi_row[H] = dzp; gi_row[2 * H] = dnp;
          g_r = drp; g_z = dzp; g_n = dnp * r;
          dghn[(bt * 2 + dir) * H + j] = g_n;
          hprev[(bt * 2 + dir) * H + j] = hp;
          dhz[u][c] = d * z;
        }
        s.dgh[cl][j0 + u] = g_r; s.dgh[cl][H + j0 + u] = g_z; s.dgh[cl][2 * H + j0 + u] = g_n;
      }
    }
    __syncthreads();

    // ---- dh_carry = dh z + dgh W_hh   ([64 x 576] . [576 x 192])
    float acc[3][8];
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[u][c] = 0.f;
    stage_wT<THREADS>(s.w[0], whhT, 0, tid);
    asm volatile("cp.async.commit_group;\n" ::);
#pragma unroll 1
    for (int ch = 0; ch < NCHUNK_B; ++ch) {
      if (ch + 1 < NCHUNK_B) {
        stage_wT<THREADS>(s.w[(ch + 1) & 1], whhT, ch + 1, tid);
        asm volatile("cp.async.commit_group;\n" ::);
        asm volatile("cp.async.wait_group 1;\n" ::);
      } else {
        asm volatile("cp.async.wait_group 0;\n" ::);
      }
      __syncthreads();
      const float (*wb)[WLD] = s.w[ch & 1];
#pragma unroll
      for (int kq = 0; kq < KC / 4; ++kq) {
        float4 dv[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) dv[c] = *reinterpret_cast<const float4*>(&s.dgh[c0 + c][ch * KC + kq * 4]);
#pragma unroll
        for (int u = 0; u < 3; ++u) {
          const float4 wv = *reinterpret_cast<const float4*>(&wb[j0 + u][kq * 4]);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            float a = acc[u][c];
            a = fmaf(wv.x, dv[c].x, a); a = fmaf(wv.y, dv[c].y, a);
            a = fmaf(wv.z, dv[c].z, a); a = fmaf(wv.w, dv[c].w, a);
            acc[u][c] = a;
          }
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int c = 0; c < 8; ++c) dh[u][c] = dhz[u][c] + acc[u][c];
    // inactive clips: dgh = 0 and dhz = 0 -> dh stays 0 until their last frame is reached
  }

  // zero rows past each clip's length so the weight-gradient GEMMs can run over all B*T rows
  for (int c = 0; c < 8; ++c) {
    const int cl = c0 + c, b = b0 + cl;
    if (b >= B) continue;
    for (int t = s.len[cl]; t < T; ++t) {
      const int64_t bt = (int64_t)b * T + t;
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int j = j0 + u;
        float* gi_row = dgi + bt * (2 * G3) + dir * G3 + j;
        gi_row[0] = 0.f; gi_row[H] = 0.f; gi_row[2 * H] = 0.f;
        dghn[(bt * 2 + dir) * H + j] = 0.f;
        hprev[(bt * 2 + dir) * H + j] = 0.f;
      }
    }
  }
}

// ---- small-batch variant: 8 clips per CTA, the serial product split FOUR ways along K.
// At B = 128 per GPU (BASELINE configs[2]) the kernel above runs 32 CTAs of 64 threads: 13.8 k dependent FMAs per thread
// and step, 26 us per step.  Here thread (ks, ug) accumulates chunks ks, ks+4, .. of dgh . W_hh for all 8 clips, the four
// partial sums meet in shared memory, and the element-wise part is dealt out by clip (thread (ks, ug) owns clips 2ks, 2ks+1
// of units 3ug..3ug+2, so the carried dh stays in its registers).  Same arithmetic, different summation order.
constexpr int KS = 4, BTK = 8, THREADS_KS = 64 * KS;
struct SmemKS {
  float dgh[BTK][DLD];
  float part[KS][BTK][H];
  float w[2][KS][H][WLD];
  int len[BTK];
  int maxlen;
};

__global__ void __launch_bounds__(THREADS_KS, 1)
gru_rec_bwd_ks_kernel(const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ gates,
                      const float* __restrict__ whhT_f, const float* __restrict__ whhT_r,
                      const int64_t* __restrict__ lengths, int B, int T,
                      float* __restrict__ dgi, float* __restrict__ dghn, float* __restrict__ hprev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmemKS& s = *reinterpret_cast<SmemKS*>(smem_raw);
  const int tid = threadIdx.x, dir = blockIdx.y, b0 = blockIdx.x * BTK;
  const float* __restrict__ whhT = dir ? whhT_r : whhT_f;
  const int ug = tid & 63, ks = tid >> 6, j0 = ug * 3, cA = 2 * ks;      // element-wise clips cA, cA + 1

  if (tid < BTK) {
    const int b = b0 + tid;
    int l = b < B ? (int)lengths[b] : 0;
    s.len[tid] = max(0, min(l, T));
  }
  if (tid == 0) s.maxlen = 0;
  __syncthreads();
  if (tid < BTK) atomicMax(&s.maxlen, s.len[tid]);
  __syncthreads();
  const int maxlen = s.maxlen;

  auto stage_round = [&](int buf, int r) {                    // chunks 4r .. 4r+3, one per K slice
    for (int p = tid; p < KS * H * (KC / 4); p += THREADS_KS) {
      const int k = p / (H * (KC / 4)), rem = p % (H * (KC / 4)), row = rem >> 2, q = rem & 3;
      cp_async16b(&s.w[buf][k][row][q * 4], whhT + (int64_t)row * G3 + (r * KS + k) * KC + q * 4);
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };

  float dh[3][2];
#pragma unroll
  for (int u = 0; u < 3; ++u) dh[u][0] = dh[u][1] = 0.f;

  for (int step = maxlen - 1; step >= 0; --step) {
    stage_round(0, 0);                                         // weights of the first round travel during the gate math
    float dhz[3][2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int cl = cA + c, b = b0 + cl, len = s.len[cl];
      const bool act = step < len;
      const int t = dir ? (len - 1 - step) : step;
      const int tp = dir ? t + 1 : t - 1;
      const int64_t bt = (int64_t)b * T + t;
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        float g_r = 0.f, g_z = 0.f, g_n = 0.f;
        dhz[u][c] = 0.f;
        if (act) {
          const int j = j0 + u;
          const float* gs = gates + (bt * 2 + dir) * (4 * H) + j;
          const float r = gs[0], z = gs[H], n = gs[2 * H], hn = gs[3 * H];
          const float hp = step > 0 ? out[((int64_t)b * T + tp) * (2 * H) + dir * H + j] : 0.f;
          const float d = dh[u][c] + dout[bt * (2 * H) + dir * H + j];
          const float dn = d * (1.f - z), dz = d * (hp - n);
          const float dnp = dn * (1.f - n * n), dzp = dz * z * (1.f - z);
          const float drp = dnp * hn * r * (1.f - r);
          float* gi_row = dgi + bt * (2 * G3) + dir * G3 + j;
          gi_row[0] = drp; gi_row[H] = dzp; gi_row[2 * H] = dnp;
          g_r = drp; g_z = dzp; g_n = dnp * r;
          dghn[(bt * 2 + dir) * H + j] = g_n;
          hprev[(bt * 2 + dir) * H + j] = hp;
          dhz[u][c] = d * z;
        }
        s.dgh[cl][j0 + u] = g_r; s.dgh[cl][H + j0 + u] = g_z; s.dgh[cl][2 * H + j0 + u] = g_n;
      }
    }
    __syncthreads();

    float acc[3][BTK];
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int c = 0; c < BTK; ++c) acc[u][c] = 0.f;
    constexpr int ROUNDS = NCHUNK_B / KS;                      // 9
#pragma unroll 1
    for (int r = 0; r < ROUNDS; ++r) {
      if (r + 1 < ROUNDS) {
        stage_round((r + 1) & 1, r + 1);
        asm volatile("cp.async.wait_group 1;\n" ::);
      } else {
        asm volatile("cp.async.wait_group 0;\n" ::);
      }
      __syncthreads();
      const float (*wb)[WLD] = s.w[r & 1][ks];
      const int k0 = (r * KS + ks) * KC;
#pragma unroll
      for (int kq = 0; kq < KC / 4; ++kq) {
        float4 dv[BTK];
#pragma unroll
        for (int c = 0; c < BTK; ++c) dv[c] = *reinterpret_cast<const float4*>(&s.dgh[c][k0 + kq * 4]);
#pragma unroll
        for (int u = 0; u < 3; ++u) {
          const float4 wv = *reinterpret_cast<const float4*>(&wb[j0 + u][kq * 4]);
#pragma unroll
          for (int c = 0; c < BTK; ++c) {
            float a = acc[u][c];
            a = fmaf(wv.x, dv[c].x, a); a = fmaf(wv.y, dv[c].y, a);
            a = fmaf(wv.z, dv[c].z, a); a = fmaf(wv.w, dv[c].w, a);
            acc[u][c] = a;
          }
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int c = 0; c < BTK; ++c) s.part[ks][c][j0 + u] = acc[u][c];
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float a = dhz[u][c];
#pragma unroll
        for (int k = 0; k < KS; ++k) a += s.part[k][cA + c][j0 + u];
        dh[u][c] = a;
      }
    // s.part is rewritten a whole step (>= 10 barriers) later; s.dgh after the barrier that ended the last round
  }

  for (int c = 0; c < 2; ++c) {
    const int cl = cA + c, b = b0 + cl;
    if (b >= B) continue;
    for (int t = s.len[cl]; t < T; ++t) {
      const int64_t bt = (int64_t)b * T + t;
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int j = j0 + u;
        float* gi_row = dgi + bt * (2 * G3) + dir * G3 + j;
        gi_row[0] = 0.f; gi_row[H] = 0.f; gi_row[2 * H] = 0.f;
        dghn[(bt * 2 + dir) * H + j] = 0.f;
        hprev[(bt * 2 + dir) * H + j] = 0.f;
      }
    }
  }
}

__global__ void transpose_kernel(const float* __restrict__ in, float* __restrict__ outp, int rows, int cols) {
  __shared__ float tile[32][33];
  const int x = blockIdx.x * 32 + threadIdx.x, y0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y)
    if (y0 + i < rows && x < cols) tile[i][threadIdx.x] = in[(int64_t)(y0 + i) * cols + x];
  __syncthreads();
  const int xo = blockIdx.y * 32 + threadIdx.x, yo0 = blockIdx.x * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y)
    if (yo0 + i < cols && xo < rows) outp[(int64_t)(yo0 + i) * rows + xo] = tile[threadIdx.x][i];
}

}  // namespace gru
}  // namespace ssb

// out[cols][rows] = in[rows][cols]^T (derived weight layout for the backward recurrence)
extern "C" int ssb_transpose(const float* in, float* out, int64_t rows, int64_t cols, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(in && out, "null pointer");
  if (rows == 0 || cols == 0) return SSB_OK;
  dim3 grid(cdiv(cols, 32), cdiv(rows, 32)), block(32, 8);
  gru::transpose_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(in, out, (int)rows, (int)cols);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

extern "C" int ssb_gru_layer_bwd_cluster_(const float* dout, const float* out, const float* gates,
                                          const float* whhT_f, const float* whhT_r, const int64_t* lengths,
                                          int64_t B, int64_t T, float* dgi, float* dghn, float* hprev, void* stream);

// dout/out [B,T,384], gates [B,T,2,4,192] (from ssb_gru_layer_fwd), whhT_* [192,576] = W_hh^T.
// Writes dgi [B,T,1152], dghn [B,T,2,192] (n-gate part of dgh) and hprev [B,T,2,192].
extern "C" int ssb_gru_layer_bwd(const float* dout, const float* out, const float* gates,
                                 const float* whhT_f, const float* whhT_r, const int64_t* lengths,
                                 int64_t B, int64_t T, int32_t hidden,
                                 float* dgi, float* dghn, float* hprev, void* stream) {
  using namespace ssb;
  using namespace ssb::gru;
  SSB_REQUIRE(dout && out && gates && whhT_f && whhT_r && lengths && dgi && dghn && hprev, "null pointer");
  SSB_REQUIRE(hidden == H, "hidden must be 192");
  SSB_REQUIRE(((uintptr_t)whhT_f & 15) == 0 && ((uintptr_t)whhT_r & 15) == 0, "W_hh^T must be 16-byte aligned");
  if (B == 0) return SSB_OK;
  // small batches (at most two waves of 33 clusters): W_hh resident across a cluster of 4 CTAs, no weight streaming
  if (cdiv(B, 8) * 2 <= 66) return ssb_gru_layer_bwd_cluster_(dout, out, gates, whhT_f, whhT_r, lengths, B, T, dgi, dghn, hprev, stream);
  // tile size: the largest one whose grid still covers the SMs (smaller tiles = more CTAs sharing the serial FMA work)
#define SSB_LAUNCH_GRUBWD(BT_)                                                                                         \
  do {                                                                                                                  \
    SSB_CUDA(cudaFuncSetAttribute(gru_rec_bwd_kernel<BT_>, cudaFuncAttributeMaxDynamicSharedMemorySize,                 \
                                  (int)sizeof(SmemB<BT_>)));                                                            \
    dim3 grid(cdiv(B, BT_), 2), block(BT_ * 8);                                                                         \
    gru_rec_bwd_kernel<BT_><<<grid, block, sizeof(SmemB<BT_>), (cudaStream_t)stream>>>(                                 \
        dout, out, gates, whhT_f, whhT_r, lengths, (int)B, (int)T, dgi, dghn, hprev);                                   \
  } while (0)
  if (cdiv(B, 64) * 2 >= kSMs) SSB_LAUNCH_GRUBWD(64);
  else if (cdiv(B, 16) * 2 >= kSMs) SSB_LAUNCH_GRUBWD(16);
  else {                                                     // small batch: 8 clips per CTA, K split four ways
    SSB_CUDA(cudaFuncSetAttribute(gru_rec_bwd_ks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SmemKS)));
    dim3 grid(cdiv(B, BTK), 2), block(THREADS_KS);
    gru_rec_bwd_ks_kernel<<<grid, block, sizeof(SmemKS), (cudaStream_t)stream>>>(
        dout, out, gates, whhT_f, whhT_r, lengths, (int)B, (int)T, dgi, dghn, hprev);
  }
#undef SSB_LAUNCH_GRUBWD
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
