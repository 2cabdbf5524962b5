// K3-bwd for small batches -- BPTT of one bidirectional GRU layer with W_hh resident across a thread-block cluster.
//
// Same arithmetic and outputs as gru_rec_bwd_kernel (gru_bwd.cu; backward of the packed nn.GRU of
// train_model_official.py:261-267 / 301-305).  At the train step's 128 clips per GPU (BASELINE configs[2]) the
// streaming kernels run 32 CTAs that each re-read W_hh^T (442 KB) from L2 for every one of the T serial steps.  Here a
// cluster of 4 CTAs owns 8 clips x one direction; CTA `rank` keeps the 576 x 48 slice of W_hh that produces ITS 48
// hidden units of dh (110 KB of shared memory, loaded once), computes the element-wise gate gradients of those units,
// and publishes them -- 3 x 48 values per clip -- into all four CTAs' copies of dgh with distributed-shared-memory
// stores.  One cluster barrier per step, no weight traffic, 128 CTAs busy instead of 32.
//   dh   = dh_carry + dout[t];  dn = dh (1-z);  dz = dh (h_prev - n)
//   dn~  = dn (1-n^2);  dz~ = dz z(1-z);  dr~ = dn~ hn r(1-r)
//   dgi[t] = (dr~, dz~, dn~);  dgh[t] = (dr~, dz~, dn~ r);  dh_carry = dh z + dgh[t] W_hh
// Thread t = (c, u): clip c (0..7), unit u (0..47) for the element-wise part and the carried dh; the same thread is
// (ks, u) in the product: K slice ks (72 of the 576 gate rows) of unit u for all 8 clips; the 8 partial sums per
// (clip, unit) meet in shared memory.
#include "common.cuh"
#include "tc.cuh"

namespace ssb {
namespace grubc {

constexpr int H = kHidden, G3 = 3 * H;
constexpr int CL = 4, U = H / CL;          // 48 units per CTA
constexpr int BT = 8;                      // clips per cluster
constexpr int KS = 8, KPER = G3 / KS;      // 72 gate rows per K slice
constexpr int THREADS = BT * U;            // 384
constexpr int DLD = G3 + 4;

struct Smem {
  float w[G3][U];                          // w[k][u] = W_hh[k][rank*48 + u]           110,592 B
  float dgh[2][BT][DLD];                   // all 576 gate gradients of the step, double-buffered
  float part[KS][BT][U];
  int len[BT];
  int maxlen;
};

__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(addr), "f"(v) : "memory");
}

__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(THREADS, 1)
gru_rec_bwd_cluster_kernel(const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ gates,
                           const float* __restrict__ whhT_f, const float* __restrict__ whhT_r,
                           const int64_t* __restrict__ lengths, int B, int T,
                           float* __restrict__ dgi, float* __restrict__ dghn, float* __restrict__ hprev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, dir = blockIdx.y;
  const uint32_t rank = tc::cluster_ctarank();
  const int b0 = (blockIdx.x / CL) * BT;
  const float* __restrict__ whhT = dir ? whhT_r : whhT_f;     // [H][G3]: whhT[j][k] = W_hh[k][j]
  const int u = tid % U, c = tid / U;                          // element-wise role; product role: ks = c
  const int j = (int)rank * U + u;                             // global hidden unit

  // ---- one-time: this CTA's slice of W_hh, k-major (row j of whhT is contiguous in k)
  for (int i = tid; i < U * (G3 / 4); i += THREADS) {
    const int uu = i / (G3 / 4), k4 = (i % (G3 / 4)) * 4;
    const float4 v = __ldg(reinterpret_cast<const float4*>(whhT + (int64_t)(rank * U + uu) * G3 + k4));
    s.w[k4][uu] = v.x; s.w[k4 + 1][uu] = v.y; s.w[k4 + 2][uu] = v.z; s.w[k4 + 3][uu] = v.w;
  }
  if (tid < BT) {
    const int b = b0 + tid;
    int l = b < B ? (int)lengths[b] : 0;
    s.len[tid] = max(0, min(l, T));
  }
  if (tid == 0) s.maxlen = 0;
  __syncthreads();
  if (tid < BT) atomicMax(&s.maxlen, s.len[tid]);
  __syncthreads();
  const int maxlen = s.maxlen;                                 // identical in the four CTAs (same clips)
  const int b = b0 + c, len = s.len[c];
  // addresses of this thread's three dgh slots in the four CTAs (buffer 0; buffer 1 is BT*DLD floats further)
  uint32_t slot[CL];
#pragma unroll
  for (int p = 0; p < CL; ++p) slot[p] = tc::mapa(tc::smem_u32(&s.dgh[0][c][j]), (uint32_t)p);
  tc::cluster_sync_all();                                      // every CTA is resident before any remote store

  float dh = 0.f;
  for (int step = maxlen - 1; step >= 0; --step) {
    const int buf = step & 1;
    // ---- element-wise part for (clip c, unit j)
    const bool act = step < len;
    float g_r = 0.f, g_z = 0.f, g_n = 0.f, dhz = 0.f;
    if (act) {
      const int t = dir ? (len - 1 - step) : step;
      const int tp = dir ? t + 1 : t - 1;
      const int64_t bt = (int64_t)b * T + t;
      const float* gs = gates + (bt * 2 + dir) * (4 * H) + j;
      const float r = gs[0], z = gs[H], n = gs[2 * H], hn = gs[3 * H];
      const float hp = step > 0 ? out[((int64_t)b * T + tp) * (2 * H) + dir * H + j] : 0.f;
      const float d = dh + dout[bt * (2 * H) + dir * H + j];
      const float dn = d * (1.f - z), dz = d * (hp - n);
      const float dnp = dn * (1.f - n * n), dzp = dz * z * (1.f - z);
      const float drp = dnp * hn * r * (1.f - r);
      float* gi_row = dgi + bt * (2 * G3) + dir * G3 + j;
      gi_row[0] = drp; gi_row[H] = dzp; gi_row[2 * H] = dnp;
      g_r = drp; g_z = dzp; g_n = dnp * r;
      dghn[(bt * 2 + dir) * H + j] = g_n;
      hprev[(bt * 2 + dir) * H + j] = hp;
      dhz = d * z;
    }
    // publish (r, z, n) gradients of (c, j) into every CTA's dgh[buf] (inactive clips publish zeros)
    const uint32_t boff = (uint32_t)buf * (BT * DLD * 4);
#pragma unroll
    for (int p = 0; p < CL; ++p) {
      st_cluster_f32(slot[p] + boff, g_r);
      st_cluster_f32(slot[p] + boff + H * 4, g_z);
      st_cluster_f32(slot[p] + boff + 2 * H * 4, g_n);
    }
    __syncwarp();
    tc::cluster_sync_all();                                    // release/acquire: all four slices of dgh[buf] are in place

    // ---- partial product: K slice ks = c of dgh[8 x 576] . W_hh[576 x 48], unit u
    float acc[BT];
#pragma unroll
    for (int q = 0; q < BT; ++q) acc[q] = 0.f;
    const int k0 = c * KPER;
#pragma unroll 2
    for (int k = k0; k < k0 + KPER; k += 4) {
      const float w0 = s.w[k][u], w1 = s.w[k + 1][u], w2 = s.w[k + 2][u], w3 = s.w[k + 3][u];
#pragma unroll
      for (int q = 0; q < BT; ++q) {
        const float4 dv = *reinterpret_cast<const float4*>(&s.dgh[buf][q][k]);
        acc[q] = fmaf(w3, dv.w, fmaf(w2, dv.z, fmaf(w1, dv.y, fmaf(w0, dv.x, acc[q]))));
      }
    }
#pragma unroll
    for (int q = 0; q < BT; ++q) s.part[c][q][u] = acc[q];
    __syncthreads();
    float a = dhz;
#pragma unroll
    for (int k = 0; k < KS; ++k) a += s.part[k][c][u];
    dh = a;
    // s.part is rewritten only after the next step's cluster barrier; dgh[buf] after two of them
  }

  // zero rows past each clip's length so the weight-gradient GEMMs can run over all B*T rows
  if (b < B) {
    for (int t = len; t < T; ++t) {
      const int64_t bt = (int64_t)b * T + t;
      float* gi_row = dgi + bt * (2 * G3) + dir * G3 + j;
      gi_row[0] = 0.f; gi_row[H] = 0.f; gi_row[2 * H] = 0.f;
      dghn[(bt * 2 + dir) * H + j] = 0.f;
      hprev[(bt * 2 + dir) * H + j] = 0.f;
    }
  }
  tc::cluster_sync_all();                                      // no CTA leaves while a peer may still store into it
}

}  // namespace grubc
}  // namespace ssb

// Same contract as ssb_gru_layer_bwd (which dispatches here for small batches).
extern "C" int ssb_gru_layer_bwd_cluster_(const float* dout, const float* out, const float* gates,
                                          const float* whhT_f, const float* whhT_r, const int64_t* lengths,
                                          int64_t B, int64_t T, float* dgi, float* dghn, float* hprev, void* stream) {
  using namespace ssb;
  using namespace ssb::grubc;
  SSB_CUDA(cudaFuncSetAttribute(gru_rec_bwd_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
  dim3 grid(CL * cdiv(B, BT), 2), block(THREADS);
  gru_rec_bwd_cluster_kernel<<<grid, block, sizeof(Smem), (cudaStream_t)stream>>>(
      dout, out, gates, whhT_f, whhT_r, lengths, (int)B, (int)T, dgi, dghn, hprev);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
