// K3a v2 -- GRU input projection for BOTH directions as one persistent tcgen05 GEMM (fp16 single-pass mode).
//
//   C[M, N] = [A1 | A2] (M x (K1+K2), fp32) . [W_fwd ; W_rev] (N x (K1+K2), fp32)^T + bias     N = 2 x 576 = 1152
//   (nn.GRU's x W_ih^T + b_ih of train_model_official.py:261-267 for the forward and reverse direction at once;
//   the concat of train:297 is two K segments of A)
//
// The output (M x 1152 fp32, 1.7 GB at 4096 clips) makes this an HBM-write-bound kernel; the design keeps
// everything else off the critical path:
//   * the weights are packed ONCE per call (ssb_proj_prepack) into fp16 blocks of 128 rows x 64 k in the
//     canonical no-swizzle K-major UMMA layout, so a weight block is ONE 16 KB bulk async copy (TMA, 1-D)
//     straight into its shared-memory stage -- no register staging, no conversion in the main loop;
//   * a persistent CTA owns a 128-row tile of A: it is read from HBM once, converted to fp16 once and stays
//     resident in shared memory while the CTA sweeps all nine 128-column tiles of N;
//   * warp roles: warps 0-7 epilogue (two per TMEM lane quarter), warp 8 weight producer, warp 9 MMA issuer;
//     4-stage weight ring and two 128-column TMEM accumulators, so the epilogue (TMEM -> +bias -> global)
//     of tile n overlaps the MMAs of tile n+1.
// K is laid out as [A1 padded to 8 | A2 padded to 8 | zeros to a multiple of 64] on both operands.
#include <stdlib.h>

#include "common.cuh"
#include "tc.cuh"

namespace ssb {
namespace gemm2 {

constexpr int BM = 128, BN = 128, KB = 64;
constexpr int CHUNK = 2048;                              // one k-chunk (8 k) of 128 rows: 16 groups x 128 B
constexpr int W_BLOCK = (KB / 8) * CHUNK;                // 16,384 B
constexpr int STAGES = 4, MAX_STAGES = 4;
constexpr int EPI_WARPS = 8, THREADS = (EPI_WARPS + 2) * 32;   // two epilogue warps per TMEM lane quarter
constexpr int TILE_LD = 36;                              // floats per row of an epilogue transpose tile (conflict-free 16-byte accesses)
constexpr int TMEM_COLS = 256;
constexpr int MAX_KP = 384, MAX_N = 1152;

struct Ctl {
  uint64_t full[MAX_STAGES], empty[MAX_STAGES], accfull[2], accempty[2];
  uint32_t tmem_base;
};

__host__ __device__ constexpr int round_up(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ constexpr int kp_of(int K1, int K2) { return round_up(round_up(K1, 8) + round_up(K2, 8), KB); }

// packed[(nt * nkb + kb) * W_BLOCK + (k/8) * CHUNK + (r/8) * 128 + (r%8) * 16 + (k%8) * 2]  for row nt*128 + r, column kb*64 + k
// split != 0: every 16 KB block is followed by the block of the fp16 residuals (w - fp16(w)), i.e. 32 KB per (nt, kb)
__global__ void prepack_kernel(const float* __restrict__ Wa, const float* __restrict__ Wb, int64_t ldw, int n_each,
                               int K1, int K2, int N, int gru_blocked, int split, uint8_t* __restrict__ packed) {
  const int K1p = round_up(K1, 8), Kp = kp_of(K1, K2), nkb = Kp / KB;
  const int items = round_up(N, BN) * (Kp / 8);          // one item = 8 consecutive k of one row
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < items; it += gridDim.x * blockDim.x) {
    const int n = it / (Kp / 8), c = it % (Kp / 8);
    float x[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int kk = c * 8 + j;
      int col = -1;
      if (kk < K1p) { if (kk < K1) col = kk; }
      else if (kk - K1p < K2) col = K1 + (kk - K1p);
      float v = 0.f;
      if (n < N && col >= 0) {
        int r = n % n_each;                                   // source row inside its direction's matrix
        if (gru_blocked) {                                    // packed row [rank][gate][u] <- W row gate * hid + rank * (hid/4) + u
          const int hid = n_each / 3, u4 = hid / 4;
          r = ((r % (3 * u4)) / u4) * hid + (r / (3 * u4)) * u4 + r % u4;
        }
        v = n < n_each ? Wa[(int64_t)r * ldw + col] : Wb[(int64_t)r * ldw + col];
      }
      x[j] = v;
    }
    const int nt = n / BN, r = n % BN, kb = c / (KB / 8), kc = c % (KB / 8);
    uint8_t* dst = packed + ((size_t)(nt * nkb + kb)) * (split ? 2 : 1) * W_BLOCK + kc * CHUNK + (r >> 3) * 128 + (r & 7) * 16;
    if (split) {
      float lo[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) tc::split_f16(x[j], x[j], lo[j]);
      *reinterpret_cast<uint4*>(dst + W_BLOCK) = make_uint4(tc::pack_f16(lo[0], lo[1]), tc::pack_f16(lo[2], lo[3]),
                                                            tc::pack_f16(lo[4], lo[5]), tc::pack_f16(lo[6], lo[7]));
    }
    *reinterpret_cast<uint4*>(dst) = make_uint4(tc::pack_f16(x[0], x[1]), tc::pack_f16(x[2], x[3]),
                                                tc::pack_f16(x[4], x[5]), tc::pack_f16(x[6], x[7]));
  }
}

struct Args {
  const float* A1; int64_t lda1; int K1;
  const float* A2; int64_t lda2; int K2;
  const uint8_t* Wp; const float* bias;
  float* C; int64_t ldc;                                   // out_half: C points at fp16 elements, ldc counts them
  int M, N, out_half;
  const int64_t* lengths; int T;
};

__global__ void __launch_bounds__(THREADS, 1) proj_gemm_kernel(Args a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K1p = round_up(a.K1, 8), Kp = kp_of(a.K1, a.K2), nkb = Kp / KB, nnt = (a.N + BN - 1) / BN;
  uint8_t* a_s = smem;                                       // Kp/8 chunks x 2048 B
  uint8_t* w_s = a_s + (Kp / 8) * CHUNK;                     // STAGES x 16 KB
  float* tile_s = reinterpret_cast<float*>(w_s + STAGES * W_BLOCK);        // EPI_WARPS x 32 x TILE_LD floats
  float* bias_s = tile_s + EPI_WARPS * 32 * TILE_LD;              // nnt * 128 floats
  uint8_t* live_s = reinterpret_cast<uint8_t*>(bias_s + nnt * BN);         // 128 bytes
  Ctl& ctl = *reinterpret_cast<Ctl*>(live_s + BM);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // known warp-uniform: role branches and MMA descriptors go uniform

  for (int i = tid; i < nnt * BN; i += THREADS) bias_s[i] = (a.bias && i < a.N) ? a.bias[i] : 0.f;
  if (warp == 0) tc::tmem_alloc<TMEM_COLS>(&ctl.tmem_base);
  if (tid == 32) {
    for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&ctl.full[s], 1); tc::mbar_init(&ctl.empty[s], 1); }
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&ctl.accfull[b], 1); tc::mbar_init(&ctl.accempty[b], EPI_WARPS); }
    tc::mbar_fence_init();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = __shfl_sync(0xffffffffu, ctl.tmem_base, 0);
  const uint32_t a_u = tc::smem_u32(a_s), w_u = tc::smem_u32(w_s);
  constexpr uint32_t idesc = tc::make_idesc_f16(BM, BN);
  uint32_t w_it = 0, acc_it = 0;                             // running weight-block / accumulator counters (ring phases)

  const int m_tiles = (a.M + BM - 1) / BM;
  for (int mt = blockIdx.x; mt < m_tiles; mt += gridDim.x) {
    const int m0 = mt * BM;
    // ---- tile liveness: a tile made only of padded frames (t >= lengths[b]) is never read downstream
    int live = 0;
    if (tid < BM) {
      const int m = m0 + tid;
      live = m < a.M && (!a.lengths || (m % a.T) < (int)a.lengths[m / a.T]);
      live_s[tid] = (uint8_t)live;
    }
    if (!__syncthreads_or(live)) continue;                  // also: every MMA of the previous tile has completed
    // ---- A tile: fp32 global -> fp16 canonical K-major, once per tile.  A warp covers 8 rows x 4 chunks per pass:
    // 128 B contiguous per row from global, conflict-free 16-byte stores.
    {
      const int items = BM * (Kp / 8);
      constexpr int NB = 4;                                  // items per batch: 8 x 16-byte loads in flight per thread
      for (int base = tid; base < items; base += NB * THREADS) {
        float4 v0[NB], v1[NB];
        uint32_t dst[NB];
#pragma unroll
        for (int u = 0; u < NB; ++u) {
          const int idx = base + u * THREADS;
          const int r8 = idx & 7, c4 = (idx >> 3) & 3, rest = idx >> 5, rg = rest & 15, cg = rest >> 4;
          const int row = rg * 8 + r8, chunk = cg * 4 + c4, kk = chunk * 8;
          v0[u] = make_float4(0.f, 0.f, 0.f, 0.f); v1[u] = v0[u];
          dst[u] = idx < items ? (uint32_t)(chunk * CHUNK + rg * 128 + r8 * 16) : 0xffffffffu;
          if (idx < items && m0 + row < a.M) {
            if (kk < K1p) {
              const float* g = a.A1 + (int64_t)(m0 + row) * a.lda1 + kk;
              if (kk < a.K1) v0[u] = __ldg(reinterpret_cast<const float4*>(g));
              if (kk + 4 < a.K1) v1[u] = __ldg(reinterpret_cast<const float4*>(g + 4));
            } else if (kk - K1p < a.K2) {
              const int k2 = kk - K1p;
              const float* g = a.A2 + (int64_t)(m0 + row) * a.lda2 + k2;
              v0[u] = __ldg(reinterpret_cast<const float4*>(g));
              if (k2 + 4 < a.K2) v1[u] = __ldg(reinterpret_cast<const float4*>(g + 4));
            }
          }
        }
#pragma unroll
        for (int u = 0; u < NB; ++u)
          if (dst[u] != 0xffffffffu)
            *reinterpret_cast<uint4*>(a_s + dst[u]) = make_uint4(tc::pack_f16(v0[u].x, v0[u].y), tc::pack_f16(v0[u].z, v0[u].w),
                                                                 tc::pack_f16(v1[u].x, v1[u].y), tc::pack_f16(v1[u].z, v1[u].w));
      }
    }
    tc::fence_async_smem();
    __syncthreads();

    if (warp == EPI_WARPS) {
      // =============================== weight producer ===============================
      if (lane == 0) {
        for (int nt = 0; nt < nnt; ++nt)
          for (int kb = 0; kb < nkb; ++kb, ++w_it) {
            const uint32_t s = w_it % STAGES, ph = (w_it / STAGES) & 1;
            tc::mbar_wait_bounded(&ctl.empty[s], ph ^ 1);
            tc::mbar_arrive_expect_tx(&ctl.full[s], W_BLOCK);
            tc::bulk_copy_g2s(w_u + s * W_BLOCK, a.Wp + (size_t)(nt * nkb + kb) * W_BLOCK, W_BLOCK, &ctl.full[s]);
          }
      } else {
        w_it += nnt * nkb;
      }
      w_it = __shfl_sync(0xffffffffu, w_it, 0);
    } else if (warp == EPI_WARPS + 1) {
      // =============================== MMA issuer ===============================
      if (tc::elect_one()) {
        for (int nt = 0; nt < nnt; ++nt, ++acc_it) {
          const uint32_t buf = acc_it & 1;
          tc::mbar_wait_bounded(&ctl.accempty[buf], ((acc_it >> 1) & 1) ^ 1);
          tc::fence_after_sync();
          for (int kb = 0; kb < nkb; ++kb, ++w_it) {
            const uint32_t s = w_it % STAGES, ph = (w_it / STAGES) & 1;
            tc::mbar_wait_bounded(&ctl.full[s], ph);
            tc::fence_after_sync();
#pragma unroll
            for (int j = 0; j < KB / 16; ++j) {
              const uint64_t da = tc::make_desc(a_u + (kb * (KB / 8) + 2 * j) * CHUNK, CHUNK, 128);
              const uint64_t db = tc::make_desc(w_u + s * W_BLOCK + 2 * j * CHUNK, CHUNK, 128);
              tc::mma_f16(tmem + buf * BN, da, db, idesc, (kb | j) != 0);
            }
            tc::mma_commit(&ctl.empty[s]);
          }
          tc::mma_commit(&ctl.accfull[buf]);
        }
      } else {
        w_it += nnt * nkb; acc_it += nnt;
      }
      w_it = __shfl_sync(0xffffffffu, w_it, 0);
      acc_it = __shfl_sync(0xffffffffu, acc_it, 0);
    } else {
      // =============================== epilogue warps ===============================
      // warp w reads TMEM lanes 32 (w % 4) ..; warps w and w + 4 split the four 32-column groups of a tile.
      // TMEM -> registers -> per-warp transpose tile (16-byte stores) -> + bias -> global: every store instruction
      // writes four full 128-byte row segments.
      float* tile = tile_s + warp * (32 * TILE_LD);
      const int q = warp & 3, half = warp >> 2;
      const int row0 = m0 + q * 32;
      const uint32_t tlane = tmem + ((uint32_t)(q * 32) << 16);
      const uint32_t live_mask = __ballot_sync(0xffffffffu, live_s[q * 32 + lane] != 0);
      const int rsub = lane >> 3, cch = lane & 7;          // read-back: 4 rows x 8 column chunks per instruction
      for (int nt = 0; nt < nnt; ++nt, ++acc_it) {
        const uint32_t buf = acc_it & 1;
        tc::mbar_wait_bounded(&ctl.accfull[buf], (acc_it >> 1) & 1);
        tc::fence_after_sync();
#pragma unroll 1
        for (int cgi = 0; cgi < 2; ++cgi) {
          const int cg = half * 2 + cgi;
          float v[32];
          tc::tmem_ld32(tlane + buf * BN + cg * 32, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<float4*>(tile + lane * TILE_LD + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          __syncwarp();
          if (a.out_half) {
            // fp16 output: 8 columns (16 bytes) per lane, 8 rows x 4 chunks per store instruction
            const int r8 = lane >> 2, c8 = lane & 3;
            const int n8 = nt * BN + cg * 32 + c8 * 8;
            if (n8 < a.N) {                                  // N is a multiple of 8
              const float4 b0 = *reinterpret_cast<const float4*>(bias_s + n8), b1 = *reinterpret_cast<const float4*>(bias_s + n8 + 4);
              uint16_t* c = reinterpret_cast<uint16_t*>(a.C) + (int64_t)row0 * a.ldc + n8;
#pragma unroll
              for (int it = 0; it < 4; ++it) {
                const int r = it * 8 + r8;
                const float4 o0 = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 8 * c8);
                const float4 o1 = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 8 * c8 + 4);
                if ((live_mask >> r) & 1u)
                  *reinterpret_cast<uint4*>(c + (int64_t)r * a.ldc) =
                      make_uint4(tc::pack_f16(o0.x + b0.x, o0.y + b0.y), tc::pack_f16(o0.z + b0.z, o0.w + b0.w),
                                 tc::pack_f16(o1.x + b1.x, o1.y + b1.y), tc::pack_f16(o1.z + b1.z, o1.w + b1.w));
              }
            }
            __syncwarp();
            continue;
          }
          const int n = nt * BN + cg * 32 + cch * 4;
          if (n < a.N) {                                     // N is a multiple of 4
            const float4 bv = *reinterpret_cast<const float4*>(bias_s + n);
            float* c = a.C + (int64_t)row0 * a.ldc + n;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int r = it * 4 + rsub;
              float4 o = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * cch);
              o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
              if ((live_mask >> r) & 1u) *reinterpret_cast<float4*>(c + (int64_t)r * a.ldc) = o;
            }
          }
          __syncwarp();
        }
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&ctl.accempty[buf]);
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<TMEM_COLS>(tmem);
}


// ---------------------------------------------------------------------------------------------------------------------
// Split mode (fp32-parity grade): C = A . W^T with A = A_hi + A_lo, W = W_hi + W_lo (fp16 each) and
// D += W_hi.A_hi + W_lo.A_hi + W_hi.A_lo.  The operand roles are SWAPPED with respect to the kernel above: the weights are
// the M = 128 operand (streamed: 32 KB hi|lo blocks through a 3-stage ring), the activations the N = 64 operand (a 64-row
// tile, hi and lo, resident: 96 KB at K = 384 -- a 128-row tile would not fit), so every MMA runs at the full M = 128
// rate and the accumulator comes out TRANSPOSED: TMEM lane = output column n, TMEM column = tile row m.  The epilogue
// therefore needs no shared-memory transpose: a warp's 32 lanes hold 32 consecutive n of one row, i.e. every store
// instruction writes one full 128-byte row segment of C.
//
// kALo = false ("weights split", the inference form): the activations are rounded to fp16 ONCE and D += W_hi.A + W_lo.A.
// Without the lo copy a 128-row tile fits (the same 96 KB), so each of the two MMAs per k-step has N = 128: per 128 rows
// 2 x 64 tensor-pipe cycles instead of 6 x 48 (an M = 128, K = 16 MMA costs about 32 + N/4 cycles), and the activation
// conversion is half the work.  Logits stay within 1e-4 of the CPU reference (bench.py's parity block); training keeps kALo.
//
// Tile height SBM and ring depth SSTAGES, measured on B200 (layer 0, 4096 clips): 64 rows with 2 / 3 / 4 stages of 32 KB:
// 0.94 / 0.84 / 0.77 ms; 96 or 128 rows (hi | lo as N = 192 / 256, 2 stages): 1.08 / 0.85 ms.  Taller tiles do not pay:
// the shared-memory pipe is the bound -- the MMAs read 128 B per cycle of operands (both K-major from shared memory) while
// the ring takes in another 70 B per cycle -- so the shared memory goes to the deepest ring that fits beside a 64-row tile
// (four stages; the bias is read from global memory to make room at K = 384).
template <bool kALo, int SBM, int SSTAGES>
__global__ void __launch_bounds__(THREADS, 1) proj_gemm_split_kernel(Args a) {
  constexpr int SCHUNK = SBM * 16;                         // one k-chunk (8 k) of one part of the tile
  constexpr int NROW = (kALo ? 2 : 1) * SBM;               // operand rows per k-chunk: SBM rows, or SBM hi rows | SBM lo rows
  constexpr int ACH = NROW * 16;                           // one k-chunk of the whole tile
  constexpr int ACC = NROW;                                // TMEM columns per accumulator
  constexpr int STMEM_COLS = 2 * ACC <= 256 ? 256 : 512;   // two accumulators
  constexpr int RG = SBM / 8;                              // 8-row groups per tile
  static_assert(SBM % 32 == 0 && NROW % 16 == 0 && NROW <= 256 && SSTAGES <= MAX_STAGES, "tile shape");
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K1p = round_up(a.K1, 8), Kp = kp_of(a.K1, a.K2), nkb = Kp / KB, nnt = (a.N + BN - 1) / BN;
  uint8_t* act_hi = smem;                                    // Kp/8 chunks x 1024 B
  uint8_t* w_s = act_hi + (Kp / 8) * ACH;                    // SSTAGES x (16 KB hi + 16 KB lo)
  uint8_t* live_s = w_s + SSTAGES * 2 * W_BLOCK;             // SBM bytes
  Ctl& ctl = *reinterpret_cast<Ctl*>(live_s + 128);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // known warp-uniform: role branches and MMA descriptors go uniform

  if (warp == 0) tc::tmem_alloc<STMEM_COLS>(&ctl.tmem_base);
  if (tid == 32) {
    for (int s = 0; s < SSTAGES; ++s) { tc::mbar_init(&ctl.full[s], 1); tc::mbar_init(&ctl.empty[s], 1); }
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&ctl.accfull[b], 1); tc::mbar_init(&ctl.accempty[b], EPI_WARPS); }
    tc::mbar_fence_init();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = __shfl_sync(0xffffffffu, ctl.tmem_base, 0);
  const uint32_t ah_u = tc::smem_u32(act_hi), w_u = tc::smem_u32(w_s);
  // M = 128 weight rows; N = the tile's operand rows (kALo: SBM hi rows | SBM lo rows side by side, so W_hi.[A_hi | A_lo]
  // is ONE MMA -- an M = 128, K = 16 MMA costs about 32 + N/4 cycles, the weight operand being fetched per instruction)
  constexpr uint32_t idesc = tc::make_idesc_f16(BN, NROW), idesc_lo = tc::make_idesc_f16(BN, SBM);
  uint32_t w_it = 0, acc_it = 0;

  const int m_tiles = (a.M + SBM - 1) / SBM;
  for (int mt = blockIdx.x; mt < m_tiles; mt += gridDim.x) {
    const int m0 = mt * SBM;
    int live = 0;
    if (tid < SBM) {
      const int m = m0 + tid;
      live = m < a.M && (!a.lengths || (m % a.T) < (int)a.lengths[m / a.T]);
      live_s[tid] = (uint8_t)live;
    }
    if (!__syncthreads_or(live)) continue;                  // also: every MMA of the previous tile has completed
    // ---- activation tile: fp32 global -> fp16 hi / lo, canonical K-major, once per tile (a warp covers 8 rows x 4
    // chunks per pass: 128 B contiguous per row from global, conflict-free 16-byte stores)
    {
      const int items = SBM * (Kp / 8);
      constexpr int NB = 4;
      for (int base = tid; base < items; base += NB * THREADS) {
        float4 v0[NB], v1[NB];
        uint32_t dst[NB];
#pragma unroll
        for (int u = 0; u < NB; ++u) {
          const int idx = base + u * THREADS;
          const int r8 = idx & 7, c4 = (idx >> 3) & 3, rest = idx >> 5, rg = rest % RG, cg = rest / RG;
          const int row = rg * 8 + r8, chunk = cg * 4 + c4, kk = chunk * 8;
          v0[u] = make_float4(0.f, 0.f, 0.f, 0.f); v1[u] = v0[u];
          dst[u] = idx < items ? (uint32_t)(chunk * ACH + rg * 128 + r8 * 16) : 0xffffffffu;
          if (idx < items && m0 + row < a.M) {
            if (kk < K1p) {
              const float* g = a.A1 + (int64_t)(m0 + row) * a.lda1 + kk;
              if (kk < a.K1) v0[u] = __ldg(reinterpret_cast<const float4*>(g));
              if (kk + 4 < a.K1) v1[u] = __ldg(reinterpret_cast<const float4*>(g + 4));
            } else if (kk - K1p < a.K2) {
              const int k2 = kk - K1p;
              const float* g = a.A2 + (int64_t)(m0 + row) * a.lda2 + k2;
              v0[u] = __ldg(reinterpret_cast<const float4*>(g));
              if (k2 + 4 < a.K2) v1[u] = __ldg(reinterpret_cast<const float4*>(g + 4));
            }
          }
        }
#pragma unroll
        for (int u = 0; u < NB; ++u)
          if (dst[u] != 0xffffffffu) {
            float x[8] = {v0[u].x, v0[u].y, v0[u].z, v0[u].w, v1[u].x, v1[u].y, v1[u].z, v1[u].w}, l[8];
            uint8_t* act_lo = act_hi + SCHUNK;               // kALo: the lo rows follow the hi rows inside every k-chunk
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (kALo) tc::split_f16(x[j], x[j], l[j]);     // otherwise pack_f16 rounds once
            *reinterpret_cast<uint4*>(act_hi + dst[u]) = make_uint4(tc::pack_f16(x[0], x[1]), tc::pack_f16(x[2], x[3]),
                                                                    tc::pack_f16(x[4], x[5]), tc::pack_f16(x[6], x[7]));
            if (kALo)
              *reinterpret_cast<uint4*>(act_lo + dst[u]) = make_uint4(tc::pack_f16(l[0], l[1]), tc::pack_f16(l[2], l[3]),
                                                                      tc::pack_f16(l[4], l[5]), tc::pack_f16(l[6], l[7]));
          }
      }
    }
    tc::fence_async_smem();
    __syncthreads();

    if (warp == EPI_WARPS) {
      // =============================== weight producer ===============================
      if (lane == 0) {
        for (int nt = 0; nt < nnt; ++nt)
          for (int kb = 0; kb < nkb; ++kb, ++w_it) {
            const uint32_t s = w_it % SSTAGES, ph = (w_it / SSTAGES) & 1;
            tc::mbar_wait_bounded(&ctl.empty[s], ph ^ 1);
            tc::mbar_arrive_expect_tx(&ctl.full[s], 2 * W_BLOCK);
            tc::bulk_copy_g2s(w_u + s * 2 * W_BLOCK, a.Wp + (size_t)(nt * nkb + kb) * 2 * W_BLOCK, 2 * W_BLOCK, &ctl.full[s]);
          }
      } else {
        w_it += nnt * nkb;
      }
      w_it = __shfl_sync(0xffffffffu, w_it, 0);
    } else if (warp == EPI_WARPS + 1) {
      // =============================== MMA issuer ===============================
      if (tc::elect_one()) {
        for (int nt = 0; nt < nnt; ++nt, ++acc_it) {
          const uint32_t buf = acc_it & 1;
          tc::mbar_wait_bounded(&ctl.accempty[buf], ((acc_it >> 1) & 1) ^ 1);
          tc::fence_after_sync();
          for (int kb = 0; kb < nkb; ++kb, ++w_it) {
            const uint32_t s = w_it % SSTAGES, ph = (w_it / SSTAGES) & 1;
            tc::mbar_wait_bounded(&ctl.full[s], ph);
            tc::fence_after_sync();
#pragma unroll
            for (int j = 0; j < KB / 16; ++j) {
              const uint32_t koff = (kb * (KB / 8) + 2 * j) * ACH;
              const uint64_t dwh = tc::make_desc(w_u + s * 2 * W_BLOCK + 2 * j * CHUNK, CHUNK, 128);
              const uint64_t dwl = tc::make_desc(w_u + s * 2 * W_BLOCK + W_BLOCK + 2 * j * CHUNK, CHUNK, 128);
              const uint64_t dact = tc::make_desc(ah_u + koff, ACH, 128);
              // columns [0, SBM): W_hi.A_hi + W_lo.A_hi; kALo: columns [SBM, 2 SBM): W_hi.A_lo (added in the epilogue)
              tc::mma_f16(tmem + buf * ACC, dwh, dact, idesc, (kb | j) != 0);
              tc::mma_f16(tmem + buf * ACC, dwl, dact, idesc_lo, true);
            }
            tc::mma_commit(&ctl.empty[s]);
          }
          tc::mma_commit(&ctl.accfull[buf]);
        }
      } else {
        w_it += nnt * nkb; acc_it += nnt;
      }
      w_it = __shfl_sync(0xffffffffu, w_it, 0);
      acc_it = __shfl_sync(0xffffffffu, acc_it, 0);
    } else {
      // =============================== epilogue warps ===============================
      // warp w reads TMEM lanes 32 (w % 4) .. (= output columns nt*128 + 32 (w % 4) + lane) and the 32 accumulator columns
      // (= tile rows) of every other 32-row block; one store instruction per row: 32 lanes x 4 B = one 128-byte segment of C
      constexpr int RB = SBM / 32, MYB = (RB + 1) / 2;         // row blocks of the tile; blocks per warp (half, half + 2, ..)
      const int q = warp & 3, half = warp >> 2;
      const uint32_t tlane = tmem + ((uint32_t)(q * 32) << 16);
      uint32_t live_mask[MYB];
#pragma unroll
      for (int k = 0; k < MYB; ++k) {
        live_mask[k] = 0;
        const int rb = half + 2 * k;
        if (rb < RB)
          for (int r = 0; r < 32; ++r) live_mask[k] |= (live_s[rb * 32 + r] != 0 ? 1u : 0u) << r;
      }
      for (int nt = 0; nt < nnt; ++nt, ++acc_it) {
        const uint32_t buf = acc_it & 1;
        tc::mbar_wait_bounded(&ctl.accfull[buf], (acc_it >> 1) & 1);
        tc::fence_after_sync();
        const int n = nt * BN + q * 32 + lane;
        const float bv = (a.bias && n < a.N) ? __ldg(a.bias + n) : 0.f;
#pragma unroll
        for (int k = 0; k < MYB; ++k) {
          const int rb = half + 2 * k;                         // warp-uniform
          float v[32];
          if (rb < RB) {
            tc::tmem_ld32(tlane + buf * ACC + rb * 32, v);
            if (kALo) {
              float w[32];
              tc::tmem_ld32(tlane + buf * ACC + SBM + rb * 32, w);
              tc::tmem_ld_wait();
#pragma unroll
              for (int r = 0; r < 32; ++r) v[r] += w[r];
            } else {
              tc::tmem_ld_wait();
            }
          }
          if (k == MYB - 1) {                                  // this warp's last read of the accumulator: it may be reused
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&ctl.accempty[buf]);
          }
          if (rb < RB && n < a.N) {
            float* c = a.C + (int64_t)(m0 + rb * 32) * a.ldc + n;
#pragma unroll
            for (int r = 0; r < 32; ++r)
              if ((live_mask[k] >> r) & 1u) c[(int64_t)r * a.ldc] = v[r] + bv;
          }
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<STMEM_COLS>(tmem);
}

inline size_t smem_bytes_split(int Kp, int nnt, int nrow, int stages) {     // nrow = operand rows per k-chunk (hi + lo)
  return (size_t)(Kp / 8) * nrow * 16 + (size_t)stages * 2 * W_BLOCK + 128 + sizeof(Ctl);
}

template <bool kALo, int SBM, int SSTAGES>
int launch_proj_split(const Args& a, int Kp, int nnt, cudaStream_t st) {
  const size_t smem = smem_bytes_split(Kp, nnt, (kALo ? 2 : 1) * SBM, SSTAGES);
  SSB_REQUIRE(smem <= 232448, "shared-memory budget");
  const int m_tiles = cdiv(a.M, SBM);
  auto kern = proj_gemm_split_kernel<kALo, SBM, SSTAGES>;
  SSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<m_tiles < kSMs ? m_tiles : kSMs, THREADS, smem, st>>>(a);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

inline size_t smem_bytes(int Kp, int nnt) {
  return (size_t)(Kp / 8) * CHUNK + STAGES * W_BLOCK + EPI_WARPS * 32 * TILE_LD * 4 + (size_t)nnt * BN * 4 + BM + sizeof(Ctl);
}

}  // namespace gemm2
}  // namespace ssb

// Bytes of the packed-weight workspace ssb_proj_prepack writes (caller-owned device memory).
extern "C" int64_t ssb_proj_packed_bytes(int64_t N, int64_t K1, int64_t K2) {
  using namespace ssb::gemm2;
  return (int64_t)round_up((int)N, BN) * kp_of((int)K1, (int)K2) * 2;
}
// same for the split (hi | lo) packing of ssb_proj_prepack_split
extern "C" int64_t ssb_proj_packed_bytes_split(int64_t N, int64_t K1, int64_t K2) { return 2 * ssb_proj_packed_bytes(N, K1, K2); }

// Packs [W_a ; W_b] (n_each rows each, row stride ldw, K1+K2 columns; W_b may be NULL with N = n_each) into the
// fp16 block layout ssb_proj_gemm streams.  A derived cache of the state-dict tensors: rebuild after they change.
// gru_blocked = 1 (n_each = 3 * hidden, hidden % 4 == 0): output column order becomes [dir][rank 0..3][gate r,z,n][hidden/4]
// instead of [dir][gate][hidden], so that the 3 x 48 values one GRU cluster CTA needs of a row are contiguous
// (ssb_gru_layer_fwd_tc gi_layout = 1); the bias passed to ssb_proj_gemm must use the same order (ssb_gru_fold_bias).
static int proj_prepack_impl(const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int64_t K1, int64_t K2,
                             int32_t gru_blocked, int32_t split, void* packed, void* stream) {
  using namespace ssb;
  using namespace ssb::gemm2;
  SSB_REQUIRE(W_a && packed, "null pointer");
  SSB_REQUIRE(K1 > 0 && K2 >= 0 && n_each > 0, "bad shape");
  SSB_REQUIRE(((uintptr_t)packed & 15) == 0, "packed must be 16-byte aligned");
  SSB_REQUIRE(!gru_blocked || n_each % 12 == 0, "GRU-blocked order needs n_each = 3 * hidden with hidden % 4 == 0");
  const int N = (int)(W_b ? 2 * n_each : n_each);
  SSB_REQUIRE(kp_of((int)K1, (int)K2) <= MAX_KP && N <= MAX_N, "K > 384 or N > 1152 not supported");
  const int items = round_up(N, BN) * (kp_of((int)K1, (int)K2) / 8);
  prepack_kernel<<<cdiv(items, 256), 256, 0, (cudaStream_t)stream>>>(W_a, W_b ? W_b : W_a, ldw, (int)n_each, (int)K1, (int)K2, N, gru_blocked,
                                                                   split, reinterpret_cast<uint8_t*>(packed));
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
extern "C" int ssb_proj_prepack(const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int64_t K1, int64_t K2,
                                int32_t gru_blocked, void* packed, void* stream) {
  return proj_prepack_impl(W_a, W_b, ldw, n_each, K1, K2, gru_blocked, 0, packed, stream);
}
// Split packing for ssb_proj_gemm_split: every 16 KB fp16 block is followed by the block of residuals w - fp16(w)
// (ssb_proj_packed_bytes_split bytes).
extern "C" int ssb_proj_prepack_split(const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int64_t K1, int64_t K2,
                                      int32_t gru_blocked, void* packed, void* stream) {
  return proj_prepack_impl(W_a, W_b, ldw, n_each, K1, K2, gru_blocked, 1, packed, stream);
}

// C[M, N] = [A1 | A2] . packed^T + bias (fp16 operands, fp32 accumulation in TMEM).  lengths/T (optional): rows of
// padded frames (m % T >= lengths[m / T]) are not written.  K1, K2, lda1, lda2 multiples of 4; 16-byte aligned bases.
// out_half = 1: C is an fp16 matrix (ldc in fp16 elements; N, ldc multiples of 8) -- the single-pass recurrence reads gi
// in fp16 (ssb_gru_layer_fwd_tc gi_layout = 2), which halves this kernel's dominant HBM stream.
extern "C" int ssb_proj_gemm(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                             const void* packed, const float* bias, void* C, int64_t ldc, int64_t M, int64_t N,
                             const int64_t* lengths, int64_t T, int32_t out_half, void* stream) {
  using namespace ssb;
  using namespace ssb::gemm2;
  SSB_REQUIRE(A1 && packed && C, "null pointer");
  SSB_REQUIRE(K1 > 0 && K1 % 4 == 0 && lda1 % 4 == 0, "K1/lda1 must be multiples of 4");
  SSB_REQUIRE(!A2 || (K2 > 0 && K2 % 4 == 0 && lda2 % 4 == 0), "K2/lda2 must be multiples of 4");
  SSB_REQUIRE(((uintptr_t)A1 & 15) == 0 && ((uintptr_t)A2 & 15) == 0 && ((uintptr_t)packed & 15) == 0, "operands must be 16-byte aligned");
  SSB_REQUIRE(!lengths || T > 0, "row mask needs T");
  SSB_REQUIRE(N % 4 == 0 && ldc % 4 == 0 && ((uintptr_t)C & 15) == 0, "N, ldc multiples of 4 and C 16-byte aligned");
  SSB_REQUIRE(!out_half || (N % 8 == 0 && ldc % 8 == 0), "fp16 output: N, ldc multiples of 8");
  if (!A2) K2 = 0;
  const int Kp = kp_of((int)K1, (int)K2);
  SSB_REQUIRE(Kp <= MAX_KP && N <= MAX_N && N > 0, "K > 384 or N > 1152 not supported");
  SSB_REQUIRE(M < (int64_t)1 << 31, "M too large");
  if (M == 0) return SSB_OK;
  Args a{A1, lda1, (int)K1, A2, lda2, (int)K2, reinterpret_cast<const uint8_t*>(packed), bias, reinterpret_cast<float*>(C), ldc, (int)M, (int)N, out_half ? 1 : 0, lengths, (int)T};
  const int nnt = cdiv(N, BN);
  const size_t smem = smem_bytes(Kp, nnt);
  const int m_tiles = cdiv(M, BM);
  SSB_CUDA(cudaFuncSetAttribute(proj_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  proj_gemm_kernel<<<m_tiles < kSMs ? m_tiles : kSMs, THREADS, smem, (cudaStream_t)stream>>>(a);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}

// C[M, N] = [A1 | A2] . W^T + bias at fp32-parity grade (fp16 hi/lo split of both operands, three products, fp32 accumulation
// in TMEM): the input projection of the split-mode recurrence for both directions in one persistent launch.  packed comes
// from ssb_proj_prepack_split.  Same alignment rules as ssb_proj_gemm; C is fp32.
static int proj_gemm_split_impl(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                                const void* packed, const float* bias, float* C, int64_t ldc, int64_t M, int64_t N,
                                const int64_t* lengths, int64_t T, bool act_lo, void* stream) {
  using namespace ssb;
  using namespace ssb::gemm2;
  SSB_REQUIRE(A1 && packed && C, "null pointer");
  SSB_REQUIRE(K1 > 0 && K1 % 4 == 0 && lda1 % 4 == 0, "K1/lda1 must be multiples of 4");
  SSB_REQUIRE(!A2 || (K2 > 0 && K2 % 4 == 0 && lda2 % 4 == 0), "K2/lda2 must be multiples of 4");
  SSB_REQUIRE(((uintptr_t)A1 & 15) == 0 && ((uintptr_t)A2 & 15) == 0 && ((uintptr_t)packed & 15) == 0, "operands must be 16-byte aligned");
  SSB_REQUIRE(!lengths || T > 0, "row mask needs T");
  SSB_REQUIRE(((uintptr_t)C & 3) == 0, "C must be 4-byte aligned");
  if (!A2) K2 = 0;
  const int Kp = kp_of((int)K1, (int)K2);
  SSB_REQUIRE(Kp <= MAX_KP && N <= MAX_N && N > 0, "K > 384 or N > 1152 not supported");
  SSB_REQUIRE(M < (int64_t)1 << 31, "M too large");
  if (M == 0) return SSB_OK;
  Args a{A1, lda1, (int)K1, A2, lda2, (int)K2, reinterpret_cast<const uint8_t*>(packed), bias, C, ldc, (int)M, (int)N, 0, lengths, (int)T};
  const int nnt = cdiv(N, BN);
  cudaStream_t st = (cudaStream_t)stream;
  if (!act_lo) return launch_proj_split<false, 128, 4>(a, Kp, nnt, st);
  // SSB_PROJ_STAGES = 2 | 3 forces a shallower weight ring (A/B runs; read once)
  static const int nst = [] { const char* e = getenv("SSB_PROJ_STAGES"); return e ? atoi(e) : 0; }();
  if (nst == 2) return launch_proj_split<true, 64, 2>(a, Kp, nnt, st);
  if (nst == 3) return launch_proj_split<true, 64, 3>(a, Kp, nnt, st);
  return launch_proj_split<true, 64, 4>(a, Kp, nnt, st);
}
extern "C" int ssb_proj_gemm_split(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                                   const void* packed, const float* bias, float* C, int64_t ldc, int64_t M, int64_t N,
                                   const int64_t* lengths, int64_t T, void* stream) {
  return proj_gemm_split_impl(A1, lda1, K1, A2, lda2, K2, packed, bias, C, ldc, M, N, lengths, T, true, stream);
}
// The inference form of the split projection: the same packed weights (ssb_proj_prepack_split), activations rounded to
// fp16 once, D += W_hi.A + W_lo.A on 128-row tiles.
extern "C" int ssb_proj_gemm_wsplit(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                                    const void* packed, const float* bias, float* C, int64_t ldc, int64_t M, int64_t N,
                                    const int64_t* lengths, int64_t T, void* stream) {
  return proj_gemm_split_impl(A1, lda1, K1, A2, lda2, K2, packed, bias, C, ldc, M, N, lengths, T, false, stream);
}

// ------------------------------------------------------------------------------------------------- no-allocation variant
// ssb_proj_gemm_static: ssb_proj_prepack + ssb_proj_gemm in one call whose packed-weight workspace is NOT caller- or
// cudaMalloc-owned: it lives in the file-scope device array below (kStaticPackedBytes = 884,736 bytes, the packing of the
// largest shape the kernel takes: N = 1152, K1p + K2p = 384; every GPU that loads the library holds it until exit).
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The call
// itself allocates nothing: no cudaMalloc / cudaFree / allocating library routine.
//   - The array is shared by every call on the device and keeps the last call's contents.  Nothing has to start at zero:
//     the pack kernel rewrites every byte the GEMM reads (zero padding included) before the GEMM, in stream order.
//   - Calls that overlap on the GPU (two streams, two host threads) race on the array; serialise them.
#include <mutex>
namespace ssb { namespace gemm2 {
constexpr size_t kStaticPackedBytes = (size_t)MAX_N * MAX_KP * 2;
static __device__ __align__(256) uint8_t g_static_packed[kStaticPackedBytes];

// Device address of g_static_packed on the current device.  In host code the array's name is only a handle, so the
// address is looked up once per device (with that device current) through cudaGetSymbolAddress and cached.
static int static_packed_ptr(uint8_t** out) {
  constexpr int kMaxDev = 64;
  static uint8_t* cache[kMaxDev] = {};
  static std::mutex mu;
  int dev = 0;
  SSB_CUDA(cudaGetDevice(&dev));
  SSB_REQUIRE(dev >= 0 && dev < kMaxDev, "device index out of range");
  std::lock_guard<std::mutex> g(mu);
  if (!cache[dev]) {
    void* p = nullptr;
    SSB_CUDA(cudaGetSymbolAddress(&p, g_static_packed));
    SSB_REQUIRE(p && ((uintptr_t)p & 255) == 0, "static workspace address");
    cache[dev] = reinterpret_cast<uint8_t*>(p);
  }
  *out = cache[dev];
  return SSB_OK;
}
}}  // namespace ssb::gemm2

extern "C" int64_t ssb_proj_static_max_bytes(void) { return (int64_t)ssb::gemm2::kStaticPackedBytes; }

extern "C" int ssb_proj_gemm_static(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                                    const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int32_t gru_blocked,
                                    const float* bias, void* C, int64_t ldc, int64_t M,
                                    const int64_t* lengths, int64_t T, int32_t out_half, void* stream) {
  using namespace ssb;
  using namespace ssb::gemm2;
  SSB_REQUIRE(A1 && W_a && C, "null pointer");
  SSB_REQUIRE(K1 > 0 && n_each > 0 && K2 >= 0, "bad shape");
  if (!A2) K2 = 0;
  const int64_t N = W_b ? 2 * n_each : n_each;
  SSB_REQUIRE(K1 <= MAX_KP && K2 <= MAX_KP && N <= MAX_N && ssb_proj_packed_bytes(N, K1, K2) <= (int64_t)kStaticPackedBytes,
              "ssb_proj_gemm_static: packed weights exceed the 884,736-byte static workspace (N <= 1152 and "
              "round_up(K1, 8) + round_up(K2, 8) <= 384); use ssb_proj_prepack + ssb_proj_gemm with a caller-owned workspace");
  uint8_t* packed = nullptr;
  if (int rc = static_packed_ptr(&packed)) return rc;
  if (int rc = proj_prepack_impl(W_a, W_b, ldw, n_each, K1, K2, gru_blocked, 0, packed, stream)) return rc;
  return ssb_proj_gemm(A1, lda1, K1, A2, lda2, K2, packed, bias, C, ldc, M, N, lengths, T, out_half, stream);
}
