// K3b (fp32 reference-precision build) -- GRU time recurrence, one layer, both directions.
//
// Semantics are those of nn.GRU on a packed sequence (train_model_official.py:261-267,
// 301-305): gate order (r, z, n); h0 = 0; clip b runs exactly lengths[b] steps; the
// reverse direction starts at t = lengths[b]-1; outputs at t >= lengths[b] are zero.
//   r = sigma(gi_r + W_hr h + b_hr)     z = sigma(gi_z + W_hz h + b_hz)
//   n = tanh(gi_n + r * (W_hn h + b_hn))        h' = (1-z) n + z h
// gi = x W_ih^T + b_ih for all timesteps comes from the projection GEMM (K3a).
//
// One CTA = 64 clips x one direction, 512 threads; thread (cg, ug) owns hidden units
// 3ug..3ug+2 of clips 8cg..8cg+7 for all three gates (72 accumulators), so the gate
// non-linearities and the state update happen in registers.  h lives in shared memory;
// W_hh (442 KB fp32 -- does not fit one SM) streams from L2 in 16-wide k chunks through a
// cp.async double buffer.  Per step and CTA: 7.1 M FMA against 442 KB of L2 reads.
#include "common.cuh"

namespace ssb {
namespace gru {

constexpr int H = kHidden;            // 192
constexpr int G3 = 3 * H;             // 576
constexpr int BT = 64;                // clips per CTA
constexpr int THREADS = 512;
constexpr int KC = 16;                // k chunk
constexpr int WLD = KC + 4;           // padded row of a W chunk (conflict-free LDS.128)
constexpr int HLD = H + 4;
constexpr int NCHUNK = H / KC;        // 12

struct Smem {
  float h[BT][HLD];
  float w[2][G3][WLD];
  int len[BT];
  int maxlen;
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void stage_w(float (*dst)[WLD], const float* __restrict__ whh, int chunk, int tid) {
  // 576 rows x 16 floats = 2304 16-byte pieces
  for (int p = tid; p < G3 * (KC / 4); p += THREADS) {
    const int row = p >> 2, q = p & 3;
    cp_async16(&dst[row][q * 4], whh + (int64_t)row * H + chunk * KC + q * 4);
  }
}

// gates (optional, training): [B][T][2][4][H] = r, z, n, (W_hn h + b_hn)
template <bool kSave>
__global__ void __launch_bounds__(THREADS, 1)
gru_rec_fwd_kernel(const float* __restrict__ gi, int64_t gi_ld,
                   const float* __restrict__ whh_f, const float* __restrict__ bhh_f,
                   const float* __restrict__ whh_r, const float* __restrict__ bhh_r,
                   const int64_t* __restrict__ lengths, int B, int T,
                   float* __restrict__ out, float* __restrict__ gates) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x;
  const int dir = blockIdx.y;
  const int b0 = blockIdx.x * BT;
  const float* __restrict__ whh = dir ? whh_r : whh_f;
  const float* __restrict__ bhh = dir ? bhh_r : bhh_f;
  const int ug = tid & 63, cg = tid >> 6;
  const int j0 = ug * 3, c0 = cg * 8;

  if (tid < BT) {
    const int b = b0 + tid;
    int l = b < B ? (int)lengths[b] : 0;
    l = max(0, min(l, T));
    s.len[tid] = l;
  }
  if (tid == 0) s.maxlen = 0;
  for (int i = tid; i < BT * HLD; i += THREADS) (&s.h[0][0])[i] = 0.f;
  __syncthreads();
  if (tid < BT) atomicMax(&s.maxlen, s.len[tid]);
  __syncthreads();
  const int maxlen = s.maxlen;

  for (int step = 0; step < maxlen; ++step) {
    float acc[3][3][8];
#pragma unroll
    for (int g = 0; g < 3; ++g)
#pragma unroll
      for (int u = 0; u < 3; ++u)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[g][u][c] = 0.f;

    stage_w(s.w[0], whh, 0, tid);
    cp_async_commit();
#pragma unroll 1
    for (int ch = 0; ch < NCHUNK; ++ch) {
      if (ch + 1 < NCHUNK) { stage_w(s.w[(ch + 1) & 1], whh, ch + 1, tid); cp_async_commit(); cp_async_wait<1>(); }
      else cp_async_wait<0>();
      __syncthreads();
      const float (*wb)[WLD] = s.w[ch & 1];
#pragma unroll
      for (int kq = 0; kq < KC / 4; ++kq) {
        float4 hv[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) hv[c] = *reinterpret_cast<const float4*>(&s.h[c0 + c][ch * KC + kq * 4]);
#pragma unroll
        for (int g = 0; g < 3; ++g)
#pragma unroll
          for (int u = 0; u < 3; ++u) {
            const float4 wv = *reinterpret_cast<const float4*>(&wb[g * H + j0 + u][kq * 4]);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              float a = acc[g][u][c];
              a = fmaf(wv.x, hv[c].x, a); a = fmaf(wv.y, hv[c].y, a);
              a = fmaf(wv.z, hv[c].z, a); a = fmaf(wv.w, hv[c].w, a);
              acc[g][u][c] = a;
            }
          }
      }
      __syncthreads();          // everyone done with this buffer (and, on the last chunk, with h)
    }

    // ---- gates + state update in registers; h is rewritten after the barrier above
    float bh[3][3];
#pragma unroll
    for (int g = 0; g < 3; ++g)
#pragma unroll
      for (int u = 0; u < 3; ++u) bh[g][u] = __ldg(bhh + g * H + j0 + u);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const int cl = c0 + c, b = b0 + cl;
      const int len = s.len[cl];
      if (step >= len) continue;
      const int t = dir ? (len - 1 - step) : step;
      const float* g_row = gi + ((int64_t)b * T + t) * gi_ld + dir * G3 + j0;
      float* o_row = out + ((int64_t)b * T + t) * (2 * H) + dir * H + j0;
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const float hprev = s.h[cl][j0 + u];
        const float r = sigmoidf_(g_row[u] + acc[0][u][c] + bh[0][u]);
        const float z = sigmoidf_(g_row[H + u] + acc[1][u][c] + bh[1][u]);
        const float hn = acc[2][u][c] + bh[2][u];
        const float n = tanhf(g_row[2 * H + u] + r * hn);
        const float hnew = (1.0f - z) * n + z * hprev;
        s.h[cl][j0 + u] = hnew;          // only this thread ever touches h[cl][j0+u] outside the k loop
        o_row[u] = hnew;
        if (kSave) {
          float* gs = gates + (((int64_t)b * T + t) * 2 + dir) * (4 * H) + j0 + u;
          gs[0] = r; gs[H] = z; gs[2 * H] = n; gs[3 * H] = hn;
        }
      }
    }
    __syncthreads();            // new h visible before the next step's k loop
  }

  // zeros past each clip's length (pad_packed_sequence semantics)
  for (int c = 0; c < 8; ++c) {
    const int cl = c0 + c, b = b0 + cl;
    if (b >= B) continue;
    for (int t = s.len[cl]; t < T; ++t) {
      float* o_row = out + ((int64_t)b * T + t) * (2 * H) + dir * H + j0;
      o_row[0] = 0.f; o_row[1] = 0.f; o_row[2] = 0.f;
    }
  }
}

}  // namespace gru
}  // namespace ssb

// One bidirectional GRU layer's recurrence.  gi [B,T,gi_ld] holds the forward
// direction's 576 projected gates at column 0 and the reverse direction's at 576.
extern "C" int ssb_gru_layer_fwd(const float* gi, int64_t gi_ld, const int64_t* lengths, int64_t B, int64_t T,
                                 int32_t hidden, const float* whh_f, const float* bhh_f,
                                 const float* whh_r, const float* bhh_r,
                                 float* out, float* gates, void* stream) {
  using namespace ssb;
  using namespace ssb::gru;
  SSB_REQUIRE(gi && lengths && whh_f && bhh_f && whh_r && bhh_r && out, "null pointer");
  SSB_REQUIRE(hidden == H, "hidden must be 192 (train_model_official.py:402)");
  SSB_REQUIRE(gi_ld >= 2 * G3, "gi row must hold both directions' gates");
  SSB_REQUIRE(((uintptr_t)whh_f & 15) == 0 && ((uintptr_t)whh_r & 15) == 0, "W_hh must be 16-byte aligned");
  if (B == 0) return SSB_OK;
  static bool configured = false;
  if (!configured) {
    SSB_CUDA(cudaFuncSetAttribute(gru_rec_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
    SSB_CUDA(cudaFuncSetAttribute(gru_rec_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
    configured = true;
  }
  dim3 grid(cdiv(B, BT), 2), block(THREADS);
  if (gates)
    gru_rec_fwd_kernel<true><<<grid, block, sizeof(Smem), (cudaStream_t)stream>>>(
        gi, gi_ld, whh_f, bhh_f, whh_r, bhh_r, lengths, (int)B, (int)T, out, gates);
  else
    gru_rec_fwd_kernel<false><<<grid, block, sizeof(Smem), (cudaStream_t)stream>>>(
        gi, gi_ld, whh_f, bhh_f, whh_r, bhh_r, lengths, (int)B, (int)T, out, gates);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
