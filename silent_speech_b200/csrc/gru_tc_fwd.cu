// K3b on the tensor cores -- GRU time recurrence as a persistent thread-block-cluster kernel.
//
// Same semantics as gru_fwd.cu (packed-sequence nn.GRU, train_model_official.py:261-267, 301-305).
// W_hh (576x192) does not fit one SM, so a cluster of 4 CTAs splits the 192 hidden units: CTA `rank`
// owns units 48*rank..+47, i.e. 144 gate rows of W_hh, resident in shared memory as fp16 for the whole
// sequence.  Per step every CTA computes D[128 clips x 144] = h[128 x 192] . W_slice^T with tcgen05.mma
// (fp32 accumulators in TMEM), applies the gate non-linearities to its 48 units in registers (fp32
// master copy of h stays in registers), writes the new h slice as fp16 into its own A-operand buffer and
// pushes that 12 KB slice to the three peers with ONE bulk DSMEM copy each (cp.async.bulk
// shared::cta -> shared::cluster) that completes on the peer's mbarrier.  No cluster-wide barrier in the
// loop: `full` (bytes landed) and `empty` (peer's MMA finished reading) mbarriers carry the dependencies.
//
// kSplit = false: fp16 operands (fast mode).  kSplit = true: split-fp16 (h = hi + lo, W = hi + lo,
// D += Ahi.Bhi + Ahi.Blo + Alo.Bhi) -- fp32-parity grade.
//
// gi must carry b_ih + (b_hr, b_hz, 0): the r/z parts of b_hh are folded into the projection bias
// (ssb_gru_fold_bias); b_hn stays separate because it sits inside r * (W_hn h + b_hn).
#include "common.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>

namespace ssb {
namespace grutc {

constexpr int H = kHidden, G3 = 3 * H;
constexpr int CL = 4;                      // cluster size
constexpr int U = H / CL;                  // 48 units per CTA
constexpr int NG = 3 * U;                  // 144 gate columns per CTA (UMMA N)
constexpr int BM = 128;                    // clips per cluster (UMMA M)
constexpr int KCH = H / 8;                 // 24 k-chunks of 8
constexpr int A_LBO = (BM / 8) * 128;      // 2048: one k-chunk holds all 128 rows
constexpr int B_LBO = (NG / 8) * 128;      // 2304
constexpr int A_BYTES = KCH * A_LBO;       // 49,152
constexpr int B_BYTES = KCH * B_LBO;       // 55,296
constexpr int SLICE = (U / 8) * A_LBO;     // 12,288: this CTA's 6 k-chunks of the A operand
constexpr int EPI_THREADS = 384, THREADS = EPI_THREADS + 32;   // 12 gate warps (3 per TMEM lane quarter) + 1 MMA warp
constexpr int UT = 16;                     // hidden units per gate thread
constexpr int TMEM_COLS = 256;

struct Ctl {
  uint64_t full, mma_done, empty, ready, gi_full[2], gi_rd[4];
  uint32_t tmem_base;
  int maxlen;
};

// single-pass mode has shared memory to spare: the next step's input projections are staged in shared memory by bulk
// copies that complete on the `gi_full` mbarrier, so their HBM latency never sits on the step's critical path.  A bulk
// copy costs its issuing warp ~90 cycles whatever its size (the per-lane UBLKCP loop), so the fewer the better:
//   kBlk (gi written by the projection GEMM in the "GRU-blocked" column order [dir][rank][gate][48 units], see
//        ssb_proj_prepack): a row's 144 values for this CTA are ONE contiguous 576-byte segment.  The three gate warps
//        of a TMEM lane quarter share 32 rows: once all three have read step s (the `gi_rd` mbarrier of the quarter,
//        96 arrivals; only the refilling lanes wait on it, then fence.proxy.async) each issues the copies of every
//        third row for step s+1 -- 128 copies per step and CTA, ~11 per warp.
//   !kBlk (standard column order [dir][gate][192]): every gate thread issues one 192-byte copy (gate `ug` of its row),
//        384 per step, 32 per warp, after a named barrier over all gate threads + fence.proxy.async.
// Rows are padded (592 B / 208 B) so that a quarter-warp's LDS.128 of eight consecutive rows hits eight different bank
// groups.  Split mode (two operand copies resident) prefetches into L2 only.
constexpr int GROW = 52;                                     // !kBlk: floats per staged (gate,row): 48 + 4 pad
constexpr int GROWB = 148;                                   // kBlk: floats per staged row: 144 + 4 pad
constexpr int GI_STAGE_BYTES = 3 * BM * GROW * 4;            // 79,872 (>= 128 * 148 * 4 = 75,776)
// kGi == 2: gi stored in fp16 (GRU-blocked): a row is 288 B, padded to 304 B in the staging area; two buffers of 38,912 B
// fit where one fp32 buffer did, so the copies for step s+2 are issued during step s and a whole step of HBM latency is
// covered (with one fp32 buffer 14 % of the gate warps' time was spent waiting for the bytes of the next step).
constexpr int GI16_ROW = NG * 2 + 16, GI16_BUF = BM * GI16_ROW;
static_assert(2 * GI16_BUF <= GI_STAGE_BYTES, "fp16 staging buffers");
template <bool kSplit>
constexpr size_t smem_bytes() { return (kSplit ? 2 : 1) * (A_BYTES + B_BYTES) + (kSplit ? 0 : GI_STAGE_BYTES) + sizeof(Ctl); }

// fast, ~1e-7-accurate gate functions: one ex2 + one rcp each (the recurrence is MUFU/latency-bound)
// single-pass mode: the hardware tanh (MUFU.TANH, |err| <= 2^-11, the same grade as the fp16 operands) serves both
// non-linearities -- sigmoid(x) = 0.5 tanh(0.5 x) + 0.5 -- so a GRU unit costs 3 MUFU ops instead of 6
__device__ __forceinline__ float tanh_hw(float x) { float y; asm("tanh.approx.f32 %0, %1;\n" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigm_hw(float x) { return fmaf(tanh_hw(0.5f * x), 0.5f, 0.5f); }
__device__ __forceinline__ float sigm(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) {
  const float e = __expf(-2.0f * fabsf(x));
  return copysignf(__fdividef(1.0f - e, 1.0f + e), x);
}

// Global accesses of the gate warps.  A TMEM lane is a clip, so the 32 lanes of a warp own 32 different rows of gi / out /
// gates, and a float4 access per lane touches 32 different 128-byte lines per instruction -- the LSU spends one cycle per
// line, and with 12 + 4 (+ 16 when the gates are saved) such instructions per thread and step the split-mode recurrence was
// bound by exactly that (13.2 us per step with saved gates, 8.4 without).  Instead the four lanes of a quad read / write the
// four consecutive float4 of ONE row per instruction (a quarter-warp covers 2 x 64 contiguous bytes: 8 lines per
// instruction instead of 32) and a 4x4 transpose over the quad (two shuffle butterflies; an involution, so the same call
// serves loads and stores) moves every item to / from the lane that owns the row.
__device__ __forceinline__ void quad_transpose(float4 (&v)[4], int lane) {
#pragma unroll
  for (int mask = 1; mask <= 2; mask <<= 1) {
    const bool up = (lane & mask) != 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i & mask) continue;
      const int i2 = i | mask;
      const float4 send = up ? v[i] : v[i2];
      float4 recv;
      recv.x = __shfl_xor_sync(0xffffffffu, send.x, mask);
      recv.y = __shfl_xor_sync(0xffffffffu, send.y, mask);
      recv.z = __shfl_xor_sync(0xffffffffu, send.z, mask);
      recv.w = __shfl_xor_sync(0xffffffffu, send.w, mask);
      if (up) v[i] = recv; else v[i2] = recv;
    }
  }
}
// the row pointer of quad member i (nullptr = that row is inactive); all 32 lanes must call
template <typename P>
__device__ __forceinline__ P* quad_ptr(P* mine, int lane, int i) {
  return reinterpret_cast<P*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(mine), (lane & ~3) | i));
}
// v[0..3] = the four float4 of this lane's row segment (64 bytes at `mine`); zeros when the row is inactive
__device__ __forceinline__ void quad_load64(const float* mine, float4 (&v)[4], int lane) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float* p = quad_ptr(mine, lane, i);
    v[i] = p ? __ldg(reinterpret_cast<const float4*>(p) + (lane & 3)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  quad_transpose(v, lane);
}
// this lane's 16 floats -> the 64-byte row segment at `mine` (skipped when nullptr)
__device__ __forceinline__ void quad_store64(float* mine, const float* x16, int lane) {
  float4 v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = make_float4(x16[4 * i], x16[4 * i + 1], x16[4 * i + 2], x16[4 * i + 3]);
  quad_transpose(v, lane);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float* p = quad_ptr(mine, lane, i);
    if (p) reinterpret_cast<float4*>(p)[lane & 3] = v[i];
  }
}

template <bool kSplit, bool kSave, int kGi>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(THREADS, 1)
gru_tc_fwd_kernel(const float* __restrict__ gi, int64_t gi_ld,
                  const float* __restrict__ whh_f, const float* __restrict__ whh_r,
                  const float* __restrict__ bhn_f, const float* __restrict__ bhn_r,
                  const int64_t* __restrict__ lengths, const int32_t* __restrict__ order, int B, int T,
                  float* __restrict__ out, float* __restrict__ gates, int rows) {
  constexpr bool kBlk = kGi != 0;                           // 0: standard fp32 columns, 1: GRU-blocked fp32, 2: GRU-blocked fp16
  static_assert(kGi == 0 || !kSplit, "blocked gi staging exists in single-pass mode only");
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a_hi = smem;
  uint8_t* b_hi = smem + A_BYTES;
  uint8_t* a_lo = smem + A_BYTES + B_BYTES;                 // only when kSplit
  uint8_t* b_lo = smem + 2 * A_BYTES + B_BYTES;
  float* gstage = reinterpret_cast<float*>(smem + (kSplit ? 2 : 1) * (A_BYTES + B_BYTES));     // !kSplit only
  Ctl& ctl = *reinterpret_cast<Ctl*>(smem + (kSplit ? 2 : 1) * (A_BYTES + B_BYTES) + (kSplit ? 0 : GI_STAGE_BYTES));

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // known warp-uniform: role branch and MMA descriptors go uniform
  const uint32_t rank = tc::cluster_ctarank();
  const int dir = blockIdx.y;
  const int b0 = (blockIdx.x / CL) * rows;       // rows <= BM clips per cluster: the tile's first `rows` rows are real
  const float* __restrict__ whh = dir ? whh_r : whh_f;
  const float* __restrict__ bhn = dir ? bhn_r : bhn_f;

  // ---- one-time: W_hh slice -> fp16 (hi, lo) in the canonical K-major layout; h = 0
  for (int p = tid; p < NG * KCH; p += THREADS) {
    const int n = p / KCH, kc = p % KCH;
    const int g = n / U, u = n % U;
    const float* src = whh + (int64_t)(g * H + rank * U + u) * H + kc * 8;
    const float4 v0 = __ldg(reinterpret_cast<const float4*>(src)), v1 = __ldg(reinterpret_cast<const float4*>(src + 4));
    const float x[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    const uint32_t off = kc * B_LBO + (n >> 3) * 128 + (n & 7) * 16;
    float hh[8], ll[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) tc::split_f16(x[j], hh[j], ll[j]);
    *reinterpret_cast<uint4*>(b_hi + off) = make_uint4(tc::pack_f16(hh[0], hh[1]), tc::pack_f16(hh[2], hh[3]),
                                                       tc::pack_f16(hh[4], hh[5]), tc::pack_f16(hh[6], hh[7]));
    if (kSplit)
      *reinterpret_cast<uint4*>(b_lo + off) = make_uint4(tc::pack_f16(ll[0], ll[1]), tc::pack_f16(ll[2], ll[3]),
                                                         tc::pack_f16(ll[4], ll[5]), tc::pack_f16(ll[6], ll[7]));
  }
  for (int i = tid; i < A_BYTES / 16; i += THREADS) {
    reinterpret_cast<uint4*>(a_hi)[i] = make_uint4(0, 0, 0, 0);
    if (kSplit) reinterpret_cast<uint4*>(a_lo)[i] = make_uint4(0, 0, 0, 0);
  }
  if (tid == 0) {
    tc::mbar_init(&ctl.full, 1);
    tc::mbar_init(&ctl.mma_done, 1);
    tc::mbar_init(&ctl.ready, EPI_THREADS);
    tc::mbar_init(&ctl.gi_full[0], EPI_THREADS);     // every gate thread arrives once per step, with or without bytes
    tc::mbar_init(&ctl.gi_full[1], EPI_THREADS);     // kGi == 2: second staging buffer
    for (int i = 0; i < 4; ++i) tc::mbar_init(&ctl.gi_rd[i], 96);   // kBlk: the three gate warps of a TMEM lane quarter
    tc::mbar_init(&ctl.empty, CL);                 // one multicast tcgen05.commit arrive per CTA of the cluster (incl. this one)
    tc::mbar_fence_init();
    ctl.maxlen = 0;
  }
  if (warp == EPI_THREADS / 32) tc::tmem_alloc<TMEM_COLS>(&ctl.tmem_base);
  __syncthreads();

  // ---- this thread's clip (gate threads: one row each, three threads per row split the 48 units)
  const int m = (warp & 3) * 32 + lane;          // row of the tile == TMEM lane
  const int ug = (warp >> 2) % 3;                // which 16 of the CTA's 48 units
  int b = -1, len = 0;
  if (tid < EPI_THREADS && m < rows && b0 + m < B) {
    b = order ? order[b0 + m] : b0 + m;
    len = max(0, min((int)lengths[b], T));
  }
  if (tid < EPI_THREADS) atomicMax(&ctl.maxlen, len);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const int maxlen = ctl.maxlen;
  const uint32_t tmem = __shfl_sync(0xffffffffu, ctl.tmem_base, 0);
  tc::cluster_sync_all();                        // every CTA's barriers exist before any remote operation

  constexpr uint32_t idesc = tc::make_idesc_f16(BM, NG);
  constexpr uint32_t kTxBytes = (CL - 1) * SLICE * (kSplit ? 2 : 1);

  if (warp == EPI_THREADS / 32) {
    // =============================== MMA issuer warp ===============================
    // The issuer warp also runs the exchange, so no gate thread sits on the step's serial chain:
    // ready (all gate threads wrote h(s-1)) -> push this CTA's slice to the three peers -> full (all four slices here)
    // -> MMA #s -> commits.  Every active gate thread awaited `empty`(s-2) before rewriting the slice, so the peers
    // have consumed h(s-2) by the time `ready` completes.
    const uint32_t src_hi = tc::smem_u32(a_hi) + rank * SLICE, src_lo = tc::smem_u32(a_lo) + rank * SLICE;
    for (int s = 1; s < maxlen; ++s) {
      tc::mbar_wait_bounded(&ctl.ready, (s - 1) & 1);
      if (lane == 0) {
        tc::mbar_arrive_expect_tx(&ctl.full, kTxBytes);
        for (uint32_t p = 1; p < CL; ++p) {
          const uint32_t peer = (rank + p) % CL;
          const uint32_t bar = tc::mapa(tc::smem_u32(&ctl.full), peer);
          tc::bulk_copy_to_peer(tc::mapa(src_hi, peer), src_hi, SLICE, bar);
          if (kSplit) tc::bulk_copy_to_peer(tc::mapa(src_lo, peer), src_lo, SLICE, bar);
        }
      }
      tc::mbar_wait_bounded(&ctl.full, (s - 1) & 1);            // all four h slices of step s-1 are in place
      tc::fence_after_sync();
      if (tc::elect_one()) {
#pragma unroll 1
        for (int j = 0; j < H / 16; ++j) {
          const uint64_t dah = tc::make_desc(tc::smem_u32(a_hi) + j * 2 * A_LBO, A_LBO, 128);
          const uint64_t dbh = tc::make_desc(tc::smem_u32(b_hi) + j * 2 * B_LBO, B_LBO, 128);
          tc::mma_f16(tmem, dah, dbh, idesc, j > 0);
          if (kSplit) {
            const uint64_t dal = tc::make_desc(tc::smem_u32(a_lo) + j * 2 * A_LBO, A_LBO, 128);
            const uint64_t dbl = tc::make_desc(tc::smem_u32(b_lo) + j * 2 * B_LBO, B_LBO, 128);
            tc::mma_f16(tmem, dah, dbl, idesc, true);
            tc::mma_f16(tmem, dal, dbh, idesc, true);
          }
        }
        tc::mma_commit(&ctl.mma_done);
        tc::mma_commit_multicast(&ctl.empty, (uint16_t)((1u << CL) - 1));   // every CTA: "rank's tensor core has consumed h(s-1)"
      }
      __syncwarp();
    }
  } else {
    // =============================== gate / state warps ===============================
    const int j0 = (int)rank * U + ug * UT;                      // first global hidden unit of this thread
    float h[UT], bn[UT];
#pragma unroll
    for (int q = 0; q < UT; ++q) { h[q] = 0.f; bn[q] = __ldg(bhn + j0 + q); }
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t a_row_off = (m >> 3) * 128 + (m & 7) * 16;

    // gi of step s2.  !kSplit: this thread fetches gate `ug` of its row (48 units, 192 B) into the staging area with one
    // bulk copy; kSplit: L2 prefetch of the thread's three 64-byte segments.
    auto prefetch_gi = [&](int s2) {
      if (kSplit) {
        if (s2 < len) {
          const int t2 = dir ? (len - 1 - s2) : s2;
          const float* gp = gi + ((int64_t)b * T + t2) * gi_ld + dir * G3 + j0;
#pragma unroll
          for (int g = 0; g < 3; ++g) asm volatile("prefetch.global.L2 [%0];\n" ::"l"(gp + g * H));
        }
      } else if (kGi == 2) {
        // fp16 rows, two staging buffers: step s2 lives in buffer s2 & 1 and is requested during step s2 - 2
        const int buf = s2 & 1;
        if (lane % 3 == ug && s2 < len) {
          const int t2 = dir ? (len - 1 - s2) : s2;
          const uint16_t* gp = reinterpret_cast<const uint16_t*>(gi) + ((int64_t)b * T + t2) * gi_ld + dir * G3 + (int)rank * NG;
          if (s2 >= 2) { tc::mbar_wait_bounded(&ctl.gi_rd[warp & 3], (s2 - 2) & 1); tc::fence_async_smem(); }
          tc::mbar_arrive_expect_tx(&ctl.gi_full[buf], NG * 2);
          tc::bulk_copy_g2s(tc::smem_u32(reinterpret_cast<uint8_t*>(gstage) + buf * GI16_BUF + m * GI16_ROW), gp, NG * 2, &ctl.gi_full[buf]);
        } else {
          tc::mbar_arrive(&ctl.gi_full[buf]);
        }
      } else if (kBlk) {
        // rows of this lane quarter are dealt out to its three warps: lane l of warp `ug` fetches row m when l % 3 == ug
        if (lane % 3 == ug && s2 < len) {
          const int t2 = dir ? (len - 1 - s2) : s2;
          const float* gp = gi + ((int64_t)b * T + t2) * gi_ld + dir * G3 + (int)rank * NG;
          // all three readers of the row are done with step s2-1.  Their reads were generic-proxy LDS, the refill is an
          // async-proxy write: without the proxy fence the copy can overtake them (seen as a rare stale-row race).
          if (s2 >= 1) { tc::mbar_wait_bounded(&ctl.gi_rd[warp & 3], (s2 - 1) & 1); tc::fence_async_smem(); }
          tc::mbar_arrive_expect_tx(&ctl.gi_full[0], NG * 4);
          tc::bulk_copy_g2s(tc::smem_u32(gstage + m * GROWB), gp, NG * 4, &ctl.gi_full[0]);
        } else {
          tc::mbar_arrive(&ctl.gi_full[0]);
        }
      } else if (s2 < len) {
        const int t2 = dir ? (len - 1 - s2) : s2;
        const float* gp = gi + ((int64_t)b * T + t2) * gi_ld + dir * G3 + ug * H + (int)rank * U;
        tc::mbar_arrive_expect_tx(&ctl.gi_full[0], U * 4);
        tc::bulk_copy_g2s(tc::smem_u32(gstage + (ug * BM + m) * GROW), gp, U * 4, &ctl.gi_full[0]);
      } else {
        tc::mbar_arrive(&ctl.gi_full[0]);
      }
    };
    prefetch_gi(0);
    if (kGi == 2 && maxlen > 1) prefetch_gi(1);
    for (int s = 0; s < maxlen; ++s) {
      const bool act = s < len;
      const int t = dir ? (len - 1 - s) : s;
      const int64_t bt = (int64_t)(act ? b : 0) * T + (act ? t : 0);
      float4 g4[12];                                             // gi: 3 gates x 16 units
      if (kSplit) {
        {
          const float* gp = act ? gi + bt * gi_ld + dir * G3 + j0 : nullptr;
#pragma unroll
          for (int g = 0; g < 3; ++g) {
            float4 q4[4];
            quad_load64(gp ? gp + g * H : nullptr, q4, lane);
#pragma unroll
            for (int v = 0; v < 4; ++v) g4[g * 4 + v] = q4[v];
          }
        }
        prefetch_gi(s + 1);
      } else if (kGi == 2) {
        tc::mbar_wait_bounded(&ctl.gi_full[s & 1], (s >> 1) & 1);
        if (act) {
          const uint8_t* rowp = reinterpret_cast<const uint8_t*>(gstage) + (s & 1) * GI16_BUF + m * GI16_ROW + ug * (UT * 2);
          uint4 raw[6];
#pragma unroll
          for (int g = 0; g < 3; ++g) {
            raw[2 * g] = *reinterpret_cast<const uint4*>(rowp + g * (U * 2));
            raw[2 * g + 1] = *reinterpret_cast<const uint4*>(rowp + g * (U * 2) + 16);
          }
#pragma unroll
          for (int i = 0; i < 6; ++i) {                            // 8 halfs -> two float4
            const float2 p0 = __half22float2(*reinterpret_cast<const __half2*>(&raw[i].x));
            const float2 p1 = __half22float2(*reinterpret_cast<const __half2*>(&raw[i].y));
            const float2 p2 = __half22float2(*reinterpret_cast<const __half2*>(&raw[i].z));
            const float2 p3 = __half22float2(*reinterpret_cast<const __half2*>(&raw[i].w));
            g4[2 * i] = make_float4(p0.x, p0.y, p1.x, p1.y);
            g4[2 * i + 1] = make_float4(p2.x, p2.y, p3.x, p3.y);
          }
        }
        if (act) {                                                 // reads performed (not merely issued) before the hand-back
#pragma unroll
          for (int v = 0; v < 12; ++v) asm volatile("" : "+f"(g4[v].x), "+f"(g4[v].y), "+f"(g4[v].z), "+f"(g4[v].w));
        }
        tc::mbar_arrive(&ctl.gi_rd[warp & 3]);
        if (s + 2 < maxlen) prefetch_gi(s + 2);
      } else {
        tc::mbar_wait_bounded(&ctl.gi_full[0], s & 1);
        if (act) {
#pragma unroll
          for (int g = 0; g < 3; ++g)
#pragma unroll
            for (int v = 0; v < 4; ++v)
              g4[g * 4 + v] = kBlk ? *reinterpret_cast<const float4*>(gstage + m * GROWB + g * U + ug * UT + v * 4)
                                   : *reinterpret_cast<const float4*>(gstage + (g * BM + m) * GROW + ug * UT + v * 4);
        }
        // The staging rows are refilled by OTHER threads once this thread has signalled below: make the loaded registers
        // a dependency of that signal, so the reads have been performed (not merely issued) when it is given.
        if (act) {
#pragma unroll
          for (int v = 0; v < 12; ++v) asm volatile("" : "+f"(g4[v].x), "+f"(g4[v].y), "+f"(g4[v].z), "+f"(g4[v].w));
        }
        if (kBlk) {
          tc::mbar_arrive(&ctl.gi_rd[warp & 3]);                   // non-blocking; only the lanes that refill a row wait on it
        } else {
          __syncwarp();
          asm volatile("bar.sync 1, %0;\n" ::"n"(EPI_THREADS) : "memory");
          tc::fence_async_smem();                                  // generic reads above -> async-proxy refill below
        }
        if (s + 1 < maxlen) prefetch_gi(s + 1);                    // a whole step of prefetch distance
      }
      if (s >= 1) {
        tc::mbar_wait_bounded(&ctl.mma_done, (s - 1) & 1);
        tc::fence_after_sync();
      }
      float sv[kSave ? 4 : 1][kSave ? UT : 1];                   // kSave: r, z, n, W_hn h + b_hn, written out after the hand-off
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float ar[8], az[8], an[8];
        if (s >= 1) {
          tc::tmem_ld8(trow + 0 * U + ug * UT + c * 8, ar);
          tc::tmem_ld8(trow + 1 * U + ug * UT + c * 8, az);
          tc::tmem_ld8(trow + 2 * U + ug * UT + c * 8, an);
          tc::tmem_ld_wait();
        } else {
#pragma unroll
          for (int q = 0; q < 8; ++q) ar[q] = az[q] = an[q] = 0.f;
        }
        if (act) {
          const float* gr = reinterpret_cast<const float*>(&g4[0 * 4 + c * 2]);
          const float* gz = reinterpret_cast<const float*>(&g4[1 * 4 + c * 2]);
          const float* gn = reinterpret_cast<const float*>(&g4[2 * 4 + c * 2]);
          float hn8[8], r8[8], z8[8], n8[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const int u = c * 8 + q;
            r8[q] = kSplit ? sigm(gr[q] + ar[q]) : sigm_hw(gr[q] + ar[q]);
            z8[q] = kSplit ? sigm(gz[q] + az[q]) : sigm_hw(gz[q] + az[q]);
            hn8[q] = an[q] + bn[u];
            n8[q] = kSplit ? tanh_fast(gn[q] + r8[q] * hn8[q]) : tanh_hw(gn[q] + r8[q] * hn8[q]);
            h[u] = (1.0f - z8[q]) * n8[q] + z8[q] * h[u];
          }
          if (kSave) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              sv[0][c * 8 + q] = r8[q]; sv[1][c * 8 + q] = z8[q]; sv[2][c * 8 + q] = n8[q]; sv[3][c * 8 + q] = hn8[q];
            }
          }
          // Before this CTA's slice is overwritten with h(s): every peer has finished MMA #s, i.e. it has consumed
          // h(s-1) -- which also means the outgoing bulk copies of h(s-1), whose SOURCE is this slice, have landed
          // (a copy still reading its source while the next step overwrote it was a rare lo-part race).
          if (c == 0 && s >= 1) tc::mbar_wait_bounded(&ctl.empty, (s - 1) & 1);
          // new h (8 units = one 16-byte core-matrix row) into this CTA's slice of the A operand
          const uint32_t off = (uint32_t)((rank * U + ug * UT + c * 8) >> 3) * A_LBO + a_row_off;
          float hh[8], ll[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) tc::split_f16(h[c * 8 + q], hh[q], ll[q]);
          *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(tc::pack_f16(hh[0], hh[1]), tc::pack_f16(hh[2], hh[3]),
                                                             tc::pack_f16(hh[4], hh[5]), tc::pack_f16(hh[6], hh[7]));
          if (kSplit)
            *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(tc::pack_f16(ll[0], ll[1]), tc::pack_f16(ll[2], ll[3]),
                                                               tc::pack_f16(ll[4], ll[5]), tc::pack_f16(ll[6], ll[7]));
        }
      }
      tc::fence_async_smem();                                    // slice visible to the async proxy (MMA, bulk copy)
      tc::fence_before_sync();                                   // TMEM reads ordered before the next step's MMA
      tc::mbar_arrive(&ctl.ready);                               // 384 arrivals: h(s) is complete in this CTA's slice
      // Global traffic goes AFTER the exchange has been kicked off: the proxy fence above waits for this thread's
      // outstanding memory operations, and a global store / cp.async in front of it put HBM latency on every step.
      quad_store64(act ? out + bt * (2 * H) + dir * H + j0 : nullptr, h, lane);
      if (kSave) {                                               // 16 stores that used to sit in front of the proxy fence
        float* gs = act ? gates + (bt * 2 + dir) * (4 * H) + j0 : nullptr;
#pragma unroll
        for (int g = 0; g < 4; ++g) quad_store64(gs ? gs + g * H : nullptr, sv[g], lane);
      }
    }
    // zeros past each clip's length (pad_packed_sequence semantics)
    if (b >= 0) {
      for (int t = len; t < T; ++t) {
        float* orow = out + ((int64_t)b * T + t) * (2 * H) + dir * H + j0;
#pragma unroll
        for (int v = 0; v < UT / 4; ++v) reinterpret_cast<float4*>(orow)[v] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  tc::fence_before_sync();
  tc::cluster_sync_all();                        // no CTA leaves while a peer may still touch its shared memory
  if (warp == EPI_THREADS / 32) tc::tmem_dealloc<TMEM_COLS>(tmem);
}

// gi_bias[d*576 + i] = b_ih_d[i] + (i < 384 ? b_hh_d[i] : 0)
__global__ void fold_bias_kernel(const float* __restrict__ bih_f, const float* __restrict__ bhh_f,
                                 const float* __restrict__ bih_r, const float* __restrict__ bhh_r, float* __restrict__ o, int blocked) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * G3) return;
  const int d = i / G3;
  int k = i % G3;
  if (blocked) k = ((k % NG) / U) * H + (k / NG) * U + k % U;        // [rank][gate][48] -> gate * 192 + rank * 48 + u
  const float* bi = d ? bih_r : bih_f;
  const float* bh = d ? bhh_r : bhh_f;
  o[i] = bi[k] + (k < 2 * H ? bh[k] : 0.f);
}

}  // namespace grutc
}  // namespace ssb

extern "C" int ssb_gru_fold_bias(const float* bih_f, const float* bhh_f, const float* bih_r, const float* bhh_r,
                                 float* out1152, int32_t gru_blocked, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(bih_f && bhh_f && bih_r && bhh_r && out1152, "null pointer");
  grutc::fold_bias_kernel<<<cdiv(2 * grutc::G3, 256), 256, 0, (cudaStream_t)stream>>>(bih_f, bhh_f, bih_r, bhh_r, out1152, gru_blocked);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// Tensor-core recurrence of one bidirectional layer.  gi [B,T,gi_ld] must already contain
// b_ih + (b_hr, b_hz, 0) (ssb_gru_fold_bias); bhh_* are the full b_hh vectors (the n part is read).
// order (nullable) = clip index per batch slot, e.g. sorted by length so that tiles are homogeneous.
// passes = 1 (fp16) or 3 (split-fp16).  Output / gates layouts as ssb_gru_layer_fwd.
// gi_layout: 0 = columns [dir][gate][192]; 1 = GRU-blocked [dir][rank][gate][48] as written through
// ssb_proj_prepack(.., gru_blocked = 1) + ssb_gru_fold_bias(.., gru_blocked = 1) (passes = 1, gates = NULL only);
// 2 = the same order stored as fp16 (ssb_proj_gemm out_half = 1), double-buffered staging.
extern "C" int ssb_gru_layer_fwd_tc(const void* gi_, int64_t gi_ld, const int64_t* lengths, const int32_t* order,
                                    int64_t B, int64_t T, int32_t hidden,
                                    const float* whh_f, const float* bhh_f, const float* whh_r, const float* bhh_r,
                                    float* out, float* gates, int32_t passes, int32_t gi_layout, void* stream) {
  using namespace ssb;
  using namespace ssb::grutc;
  const float* gi = reinterpret_cast<const float*>(gi_);     // gi_layout 2: fp16 elements, gi_ld counts them
  SSB_REQUIRE(gi && lengths && whh_f && bhh_f && whh_r && bhh_r && out, "null pointer");
  SSB_REQUIRE(hidden == H, "hidden must be 192");
  SSB_REQUIRE(gi_ld >= 2 * G3 && gi_ld % (gi_layout == 2 ? 8 : 4) == 0 && ((uintptr_t)gi & 15) == 0, "gi must be 16-byte aligned rows of >= 1152");
  SSB_REQUIRE(((uintptr_t)whh_f & 15) == 0 && ((uintptr_t)whh_r & 15) == 0 && ((uintptr_t)out & 15) == 0, "alignment");
  SSB_REQUIRE(passes == 1 || passes == 3, "passes must be 1 or 3");
  SSB_REQUIRE(gi_layout == 0 || ((gi_layout == 1 || gi_layout == 2) && passes == 1 && !gates), "GRU-blocked gi: single-pass inference only");
  if (B == 0) return SSB_OK;
  // Small batches: a 128-row tile that is only partly filled costs the tensor core nothing extra, but the per-step work of
  // the gate warps (gi reads, gate math, out / gates stores) is per REAL row.  So while all clusters still fit one wave
  // (148 SMs / 4 CTAs / 2 directions = 18 per direction) the clips are dealt out 64 or 32 per cluster: at the train
  // step's 128 clips per GPU that is 4 clusters per direction on 32 SMs instead of 1 + 1 on 8, each with one busy lane
  // quarter -- the serial step then costs little more than its MMA + exchange chain.
  int rows = BM;
  if (cdiv(B, 32) <= 18) rows = 32;
  else if (cdiv(B, 64) <= 18) rows = 64;
  dim3 grid(CL * cdiv(B, rows), 2), block(THREADS);
  cudaStream_t st = (cudaStream_t)stream;
#define SSB_LAUNCH_GRUTC(SPLIT, SAVE, BLK)                                                                          \
  do {                                                                                                               \
    SSB_CUDA(cudaFuncSetAttribute(gru_tc_fwd_kernel<SPLIT, SAVE, BLK>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                  (int)smem_bytes<SPLIT>()));                                                        \
    gru_tc_fwd_kernel<SPLIT, SAVE, BLK><<<grid, block, smem_bytes<SPLIT>(), st>>>(                                   \
        gi, gi_ld, whh_f, whh_r, bhh_f + 2 * H, bhh_r + 2 * H, lengths, order, (int)B, (int)T, out, gates, rows);    \
  } while (0)
  if (passes == 3)         { if (gates) SSB_LAUNCH_GRUTC(true, true, 0); else SSB_LAUNCH_GRUTC(true, false, 0); }
  else if (gi_layout == 2) { SSB_LAUNCH_GRUTC(false, false, 2); }
  else if (gi_layout == 1) { SSB_LAUNCH_GRUTC(false, false, 1); }
  else                     { if (gates) SSB_LAUNCH_GRUTC(false, true, 0); else SSB_LAUNCH_GRUTC(false, false, 0); }
#undef SSB_LAUNCH_GRUTC
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
