// K2 v3 -- the single-pass (fp16) instance of the fused mouth-ROI CNN forward.  The split-mode and training instances
// moved to K2 v4 (roi_cnn_tc3.cu: TMA-staged frames, conv1 on exact raw pixels, standardisation in the epilogue); in
// single-pass mode with four 128-thread slots v4 measures 7.64 ms per 4096 clips against this kernel's 6.55 ms (its
// conv1 front end executes more instructions per frame and that mode is latency-bound per slot), so passes = 1 keeps the
// register-prefetched, standardised-plane formulation below until v4 catches up.  Only <kSplit = false, 4 slots, 128
// threads> is instantiated here.
//
// K2 v3 -- fused mouth-ROI CNN forward, all three convolutions on tcgen05, pooling folded into N (the v2
// formulation below), re-organised for latency hiding: in fp16 mode FOUR independent frame slots of 128 threads
// per CTA (one warp per TMEM lane quarter, one MMA-issuing thread per slot -- concurrent issuers reach ~40 cycles
// per small-N MMA per SM against ~70 for a single issuer, profiles/microbench), frames handed out dynamically by
// an atomic counter, the next frame's pixels prefetched into registers while the current one is in the tensor
// core, all per-thread addressing hoisted out of the frame loop, and conv2's tiles committed one by one so the
// epilogue of tile i overlaps the MMAs of tile i+1.
//
// Arithmetic contract: train_model_official.py:286-294 + TinyROICNN (209-229); live semantics with standardize=0.
//
// The two pooled convolutions are computed per POOLED pixel: the 2x2 window of conv outputs that a max-pool
// cell covers depends on one 4x4 input patch, so the four sub-positions become four groups of output columns
// of ONE GEMM (weights placed at their shifted positions inside the 4x4 patch, zeros elsewhere):
//   conv1+pool : M = pooled pixels, K = 16 (4x4 patch of the 1-channel frame),  N = 4 x 8  = 32
//   conv2+pool : M = pooled pixels, K = 128 (4x4 patch x 8 channels),          N = 4 x 16 = 64
//   conv3      : M = pixels,        K = 144 (9 taps x 16 channels),            N = 24 (padded 32)
// The epilogue of a pooled layer is then lane-local: max over the 4 column groups, + bias, ReLU -- no
// shuffles, no idle lanes.  conv2 and conv3 read their A operand straight from the activation planes through
// the shared-memory descriptor (no im2col copy): activations are fp16, pixel-major (8 channels = 16 bytes per
// pixel); conv2's input is kept in four (row parity, column parity) planes so that "pooled pixel m -> patch
// element (r,c)" is base(r,c) + 16*m, i.e. a legal canonical K-major operand with SBO = 128 and one k-chunk
// per patch element; a K=16 MMA takes two patch elements whose planes are LBO apart.  conv1's patches (32 B per
// pooled pixel) are gathered from the standardised fp16 frame, 3 tiles (384 pooled pixels) at a time.
//
// One CTA hosts kSlots independent "frame slots" of 256 threads (named barriers, own TMEM columns, own
// activation buffers) that share one copy of the weight operands; while one slot waits for its MMAs the
// other computes.  kSplit: operands are fp16 hi + lo; D += Ahi.Bhi + Ahi.Blo + Alo.Bhi (fp32-parity grade).
#include "common.cuh"
#include "tc.cuh"

namespace ssb {
namespace cnn3v3 {

constexpr int H0 = 48, W0 = 96, NPIX = H0 * W0;
constexpr int H1 = 24, W1 = 48, C1 = 8;                // pooled conv1 output
constexpr int H2 = 12, W2 = 24, C2 = 16;               // pooled conv2 output
constexpr int C3 = 24, C3P = 32, E = 32;
// Training: activations handed to the backward kernel instead of being recomputed there (one record per frame, the
// layout ssb_roi_cnn_bwd reads -- see roi_cnn.cuh: SAVED_*).
constexpr int SV_A1 = 0, SV_A2 = 36864, SV_IDX1 = 55296, SV_IDX2 = 57600, SV_MASK3 = 58752, SV_MEAN3 = 59904, SV_BYTES = 60000;
constexpr int PIX_LD = 100, PIX_ROWS = 50;             // standardised frame, fp16, interior at [1..48][1..96]
constexpr int PIX_BYTES = PIX_ROWS * PIX_LD * 2;       // 10,000
constexpr int P1_ROWS = 13, P1_COLS = 25;              // a1 parity plane: [13][25] pixels x 16 B
constexpr int P1_PLANE = P1_ROWS * P1_COLS * 16;       // 5,200
constexpr int A1_BYTES = 4 * P1_PLANE;                 // 20,800
constexpr int PITCH2 = 26, ROWS2 = 18;                 // a2: 2 channel planes x [18][26] pixels x 16 B (as in v1)
constexpr int PLANE2 = ROWS2 * PITCH2 * 16;            // 7,488
constexpr int A2_BYTES = 2 * PLANE2;                   // 14,976
constexpr int W1_CHUNK = 32 * 16, W2_CHUNK = 64 * 16;  // bytes per k-chunk of the weight operands
constexpr int W1_BYTES = 2 * W1_CHUNK, W2_BYTES = 16 * W2_CHUNK;   // 1,024 and 16,384 per part

template <bool kSplit>
struct Lay {
  static constexpr int P = kSplit ? 2 : 1;
  static constexpr int STRIP_TILES = kSplit ? 2 : 3;   // conv1 tiles (128 pooled pixels each) gathered per round
  static constexpr int PATCH_CHUNK = STRIP_TILES * 128 * 16, PATCH_BYTES = 2 * PATCH_CHUNK;
  static constexpr int W3_ROWS = kSplit ? 64 : 32;     // conv3 keeps hi|lo stacked along N (v1 scheme)
  static constexpr int W3_CHUNK = W3_ROWS * 16, W3_BYTES = 18 * W3_CHUNK;
  // per-slot region R0: {standardised frame, conv1 patches} during conv1, a2 afterwards
  static constexpr int R0_A = P * (PIX_BYTES + PATCH_BYTES), R0_B = P * A2_BYTES;
  static constexpr int R0_BYTES = ((R0_A > R0_B ? R0_A : R0_B) + 127) / 128 * 128;
  static constexpr int A1_ALLOC = (P * A1_BYTES + 1408 + 127) / 128 * 128;    // + over-read slack of the last conv2 tile (1,360 B)
  static constexpr int SLOT_BYTES = R0_BYTES + A1_ALLOC;
  static constexpr int OFF_PIX = 0, OFF_PATCH = P * PIX_BYTES;
};

struct SlotMisc {
  float part[9][C3P];                                  // conv3 channel sums per (tile, lane quarter): fixed-order reduction

  uint32_t red[16];
  uint64_t bar[3];
  long long next;                                      // next frame of this slot (dynamic hand-out)
};

struct Shared {                                        // one per CTA, after the slots and the weight operands
  alignas(16) float b1[C1];
  alignas(16) float b2[C2];
  alignas(16) float b3[C3P];
  float wfc[E][C3];
  float bfc[E];
  uint32_t tmem_base;
};

struct Weights {
  const float *w1, *b1, *w2, *b2, *w3, *b3, *wfc, *bfc;
};

__device__ __forceinline__ uint4 pack8(const float* v) {
  return make_uint4(tc::pack_f16(v[0], v[1]), tc::pack_f16(v[2], v[3]), tc::pack_f16(v[4], v[5]), tc::pack_f16(v[6], v[7]));
}
template <int ST>
__device__ __forceinline__ void slot_sync(int slot) { asm volatile("bar.sync %0, %1;\n" ::"r"(slot + 1), "n"(ST) : "memory"); }

// ---- weight operands (built once per CTA).  w1op / w2op: hi part, then (kSplit) lo part.
template <bool kSplit>
__device__ __forceinline__ void build_weights(uint8_t* w1op, uint8_t* w2op, uint8_t* w3op, const Weights& g, int tid, int nthreads) {
  using L = Lay<kSplit>;
  // conv1: row n = sub*8 + co, k = r*4 + c (patch position); w1[co][dy][dx] sits at r = i+dy, c = j+dx
  for (int idx = tid; idx < 32 * 2; idx += nthreads) {
    const int n = idx >> 1, kc = idx & 1, sub = n >> 3, co = n & 7, i = sub >> 1, j = sub & 1;
    float hi[8], lo[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int r = kc * 2 + (e >> 2), c = e & 3, dy = r - i, dx = c - j;
      const float w = (dy >= 0 && dy < 3 && dx >= 0 && dx < 3) ? g.w1[co * 9 + dy * 3 + dx] : 0.f;
      tc::split_f16(w, hi[e], lo[e]);
    }
    // split mode: rows [0,32) of a k-chunk = hi, rows [32,64) = lo, so that A_hi . [B_hi | B_lo] is ONE MMA with N = 64
    const uint32_t off = kc * (L::P * W1_CHUNK) + (n >> 3) * 128 + (n & 7) * 16;
    *reinterpret_cast<uint4*>(w1op + off) = pack8(hi);
    if (kSplit) *reinterpret_cast<uint4*>(w1op + W1_CHUNK + off) = pack8(lo);
  }
  // conv2: row n = sub*16 + co, k-chunk = patch element r*4 + c, 8 input channels per chunk
  for (int idx = tid; idx < 64 * 16; idx += nthreads) {
    const int n = idx >> 4, kc = idx & 15, sub = n >> 4, co = n & 15, i = sub >> 1, j = sub & 1;
    const int r = kc >> 2, c = kc & 3, dy = r - i, dx = c - j;
    const bool in = dy >= 0 && dy < 3 && dx >= 0 && dx < 3;
    float hi[8], lo[8];
#pragma unroll
    for (int ci = 0; ci < 8; ++ci) tc::split_f16(in ? g.w2[(co * C1 + ci) * 9 + dy * 3 + dx] : 0.f, hi[ci], lo[ci]);
    const uint32_t off = kc * (L::P * W2_CHUNK) + (n >> 3) * 128 + (n & 7) * 16;     // split: rows [0,64) hi, [64,128) lo
    *reinterpret_cast<uint4*>(w2op + off) = pack8(hi);
    if (kSplit) *reinterpret_cast<uint4*>(w2op + W2_CHUNK + off) = pack8(lo);
  }
  // conv3: chunk = tap*2 + channel plane; rows co (hi), co + 32 (lo, kSplit); other rows zero
  for (int idx = tid; idx < L::W3_ROWS * 18; idx += nthreads) {
    const int n = idx / 18, ch = idx % 18, tap = ch >> 1, pl = ch & 1;
    const int co = n & 31, part = n >> 5;
    float hi[8], lo[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) tc::split_f16(co < C3 ? g.w3[(co * C2 + pl * 8 + c) * 9 + tap] : 0.f, hi[c], lo[c]);
    const uint32_t off = ch * L::W3_CHUNK + (n >> 3) * 128 + (n & 7) * 16;
    *reinterpret_cast<uint4*>(w3op + off) = part ? pack8(lo) : pack8(hi);
  }
}

template <bool kSplit, int kSlots, int ST, bool kSave>
__global__ void __launch_bounds__(ST * kSlots, 1)
roi_cnn_v3_kernel(const uint8_t* __restrict__ R, const int64_t* __restrict__ lengths, int B, int T,
                   Weights gw, int standardize, float* __restrict__ out, int64_t out_stride,
                   unsigned int* __restrict__ counter, uint8_t* __restrict__ saved) {
  using L = Lay<kSplit>;
  constexpr int NWG = ST / 128;                        // warps per TMEM lane quarter
  constexpr int WPT = (NPIX / 4 + ST - 1) / ST;        // 4-pixel words per thread: 9 (ST=128) or 5 (ST=256)
  constexpr int TMEM_PER_SLOT = 512 / kSlots;
  constexpr bool kPipe2 = TMEM_PER_SLOT < 3 * (kSplit ? 128 : 64);   // conv2's three tiles do not fit: commit tile by tile
  static_assert(TMEM_PER_SLOT >= 2 * (kSplit ? 128 : 64), "a slot needs two conv2 tiles of TMEM columns");
  static_assert(kPipe2, "the non-pipelined conv2 path was removed");
  constexpr int GIT = L::STRIP_TILES * 256 / ST;       // gather items per thread in a full conv1 round
  constexpr int E1 = (L::STRIP_TILES + NWG - 1) / NWG; // conv1 epilogue tiles per thread in a full round
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* w1op = smem + kSlots * L::SLOT_BYTES;
  uint8_t* w2op = w1op + L::P * W1_BYTES;
  uint8_t* w3op = w2op + L::P * W2_BYTES;
  SlotMisc* smz = reinterpret_cast<SlotMisc*>(w3op + L::W3_BYTES);
  Shared& sh = *reinterpret_cast<Shared*>(smz + kSlots);
  // slot and warp index through a shuffle: known warp-uniform to the compiler (MMA descriptors stay in uniform registers)
  const int slot = __shfl_sync(0xffffffffu, (int)(threadIdx.x / ST), 0), tid = threadIdx.x % ST, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  uint8_t* r0 = smem + slot * L::SLOT_BYTES;           // pix / patches, later a2
  uint8_t* a1 = r0 + L::R0_BYTES;
  SlotMisc& mz = smz[slot];

  // ---- one-time setup (whole CTA)
  build_weights<kSplit>(w1op, w2op, w3op, gw, threadIdx.x, ST * kSlots);
  for (int i = threadIdx.x; i < E * C3; i += ST * kSlots) sh.wfc[i / C3][i % C3] = gw.wfc[i];
  if (threadIdx.x < C1) sh.b1[threadIdx.x] = gw.b1[threadIdx.x];
  if (threadIdx.x < C2) sh.b2[threadIdx.x] = gw.b2[threadIdx.x];
  if (threadIdx.x < C3P) sh.b3[threadIdx.x] = threadIdx.x < C3 ? gw.b3[threadIdx.x] : 0.f;
  if (threadIdx.x < E) sh.bfc[threadIdx.x] = gw.bfc[threadIdx.x];
  for (int i = tid; i < L::A1_ALLOC / 16; i += ST) reinterpret_cast<uint4*>(a1)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x < 32) tc::tmem_alloc<512>(&sh.tmem_base);
  if (tid == 32) { tc::mbar_init(&mz.bar[0], 1); tc::mbar_init(&mz.bar[1], 1); tc::mbar_init(&mz.bar[2], 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = __shfl_sync(0xffffffffu, sh.tmem_base, 0) + slot * TMEM_PER_SLOT;
  const uint32_t r0_u = tc::smem_u32(r0), a1_u = tc::smem_u32(a1);
  const uint32_t w1_u = tc::smem_u32(w1op), w2_u = tc::smem_u32(w2op), w3_u = tc::smem_u32(w3op);
  constexpr uint32_t idesc1 = tc::make_idesc_f16(128, 32), idesc2 = tc::make_idesc_f16(128, 64);
  constexpr uint32_t idesc1s = tc::make_idesc_f16(128, 64), idesc2s = tc::make_idesc_f16(128, 128);   // hi|lo stacked along N
  constexpr int C1W = kSplit ? 64 : 32, C2W = kSplit ? 128 : 64;        // TMEM columns per conv1 / conv2 tile
  constexpr uint32_t idesc3_hi = tc::make_idesc_f16(128, kSplit ? 64 : 32), idesc3_lo = tc::make_idesc_f16(128, 32);
  constexpr int COLS3 = kSplit ? 64 : 32;
  uint32_t ph0 = 0, ph1 = 0, ph2 = 0;
  const int q = warp & 3, wg = warp >> 2;
  const int row = q * 32 + lane;                       // this thread's row of every 128-row tile (= TMEM lane)
  const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);

  // ---- per-thread addressing, hoisted out of the frame loop
  // 4-pixel words of the frame per thread.  ST == 128: every warp owns 12 whole rows (288 words = 9 x 32), so the
  // left neighbour of a word is the previous lane (or lane 31 of the previous iteration) and each thread can
  // store ONE aligned 8-byte group [left neighbour's last pixel, a, b, c] -- the 1-pixel border shift of the
  // padded plane costs a shuffle instead of three misaligned stores.
  constexpr bool kRowWarp = (ST == 128);
  auto word_of = [&](int j) { return kRowWarp ? 288 * warp + 32 * j + lane : tid + j * ST; };
  uint32_t pixoff[WPT];                                // byte offset of word j's store in the padded frame plane
  uint32_t rowend = 0, rowstart = 0;                   // bit j: word j is the last / first of its row
#pragma unroll
  for (int j = 0; j < WPT; ++j) {
    const int word = word_of(j), y = word / (W0 / 4), x = (word % (W0 / 4)) * 4;
    pixoff[j] = ((y + 1) * PIX_LD + x + (kRowWarp ? 0 : 1)) * 2;
    if (word % (W0 / 4) == W0 / 4 - 1) rowend |= 1u << j;
    if (word % (W0 / 4) == 0) rowstart |= 1u << j;
  }
  uint32_t gsrc[GIT], gdst[GIT];                       // conv1 patch gather of round 0 (rounds advance by 8 pooled rows)
#pragma unroll
  for (int k = 0; k < GIT; ++k) {
    const int idx = tid + k * ST, kc = idx / (L::STRIP_TILES * 128), pxs = idx % (L::STRIP_TILES * 128);
    const int py = pxs / W1, px = pxs % W1;
    gsrc[k] = ((2 * py + 2 * kc) * PIX_LD + 2 * px) * 2;
    gdst[k] = L::OFF_PATCH + kc * L::PATCH_CHUNK + pxs * 16;
  }
  uint32_t a1off[E1];                                  // conv1 epilogue destination of round 0's tiles wg, wg+NWG, ..
#pragma unroll
  for (int k = 0; k < E1; ++k) {
    const int Pp = (wg + k * NWG) * 128 + row, rr = Pp / W1 + 1, cc = Pp % W1 + 1;
    a1off[k] = ((rr & 1) * 2 + (cc & 1)) * P1_PLANE + ((rr >> 1) * P1_COLS + (cc >> 1)) * 16;
  }
  int a2off[3];                                        // conv2 epilogue destination per tile (-1: virtual pixel)
#pragma unroll
  for (int tl = 0; tl < 3; ++tl) {
    const int vp = tl * 128 + row, py = vp / P1_COLS, px = vp % P1_COLS;
    a2off[tl] = (py < H2 && px < W2) ? ((py + 1) * PITCH2 + px + 1) * 16 : -1;
  }

  const long long frames = (long long)B * T;
  const long long handout_base = (long long)gridDim.x * kSlots;
  uint32_t wv[WPT];
  bool live = false;
  auto fetch = [&](long long fr) {                     // pixels of frame fr -> registers (loads stay in flight)
    live = fr < frames;
    if (live && lengths) live = (int)(fr % T) < (int)lengths[fr / T];
    if (live) {
#pragma unroll
      for (int j = 0; j < WPT; ++j) {
        const int word = word_of(j);
        wv[j] = word < NPIX / 4 ? __ldg(reinterpret_cast<const uint32_t*>(R + fr * NPIX) + word) : 0u;
      }
    }
  };
  long long f = (long long)blockIdx.x * kSlots + slot;
  fetch(f);
  while (f < frames) {
    unsigned int handout = 0;
    if (tid == 0) handout = atomicAdd(counter, 1u);    // this slot's NEXT frame; consumed after the first barrier
    float* o = out + f * out_stride;
    if (!live) {                                       // padded frame (t >= lengths[b]): zero embedding, no work
      if (tid < E) o[tid] = 0.f;
      if (tid == 0) mz.next = handout_base + handout;
      slot_sync<ST>(slot);
      f = mz.next;
      slot_sync<ST>(slot);
      fetch(f);
      continue;
    }
    // ================= P0: exact integer statistics, standardised fp16 frame (pixel (y,x) at [y+1][x+1])
    {
      uint32_t s1 = 0, s2 = 0;
#pragma unroll
      for (int j = 0; j < WPT; ++j) {
        s1 = __dp4a(wv[j], 0x01010101u, s1);
        s2 = __dp4a(wv[j], wv[j], s2);
      }
      s1 = __reduce_add_sync(0xffffffffu, s1);
      s2 = __reduce_add_sync(0xffffffffu, s2);
      if (lane == 0) { mz.red[warp] = s1; mz.red[8 + warp] = s2; }
      if (tid == 0) mz.next = handout_base + handout;
      // zero border of the frame plane(s): region R0 held a2 a moment ago
      if (tid < PIX_LD / 2) {
#pragma unroll
        for (int p = 0; p < L::P; ++p) {
          uint32_t* pw = reinterpret_cast<uint32_t*>(r0 + p * PIX_BYTES);
          pw[tid] = 0; pw[(PIX_ROWS - 1) * (PIX_LD / 2) + tid] = 0;                 // rows 0 and 49
          if (!kRowWarp) {
            uint16_t* ph = reinterpret_cast<uint16_t*>(r0 + p * PIX_BYTES);
            ph[tid * PIX_LD] = 0; ph[tid * PIX_LD + W0 + 1] = 0;                    // columns 0 and 97
          }
        }
      }
      slot_sync<ST>(slot);
      float mean_u = 0.f, inv_s = 1.0f / 255.0f;
      if (standardize) {
        uint32_t t1 = 0, t2 = 0;
#pragma unroll
        for (int i = 0; i < ST / 32; ++i) { t1 += mz.red[i]; t2 += mz.red[8 + i]; }
        // N*S2 - S1^2 is an exact 64-bit integer (< 2^41), so the cancellation happens before any rounding and the rest
        // is fp32: unbiased variance in raw pixel units, std clamped at 1e-6 * 255 (train_model_official.py:289-291)
        const unsigned long long d = (unsigned long long)NPIX * t2 - (unsigned long long)t1 * t1;
        const float var_u = (float)d * (1.0f / ((float)NPIX * (float)(NPIX - 1)));
        mean_u = (float)t1 * (1.0f / (float)NPIX);               // t1 < 2^24 and NPIX = 9 * 2^9: one rounding
        inv_s = 1.0f / fmaxf(sqrtf(var_u), 255.0e-6f);
      }
      const float nmean = -mean_u * inv_s;
      uint32_t carry_hi = 0, carry_lo = 0;             // lane 31's last pixel of the previous iteration (kRowWarp)
#pragma unroll
      for (int j = 0; j < WPT; ++j) {
        if (WPT * ST == NPIX / 4 || word_of(j) < NPIX / 4) {
          float hi[4], lo[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // byte k -> float through the 2^23 + b bit pattern (one PRMT, one exact FADD), then one FFMA
            const float xf = __uint_as_float(__byte_perm(wv[j], 0x4B000000u, 0x7440 + k)) - 8388608.0f;
            hi[k] = fmaf(xf, inv_s, nmean);
            if (kSplit) tc::split_f16(hi[k], hi[k], lo[k]);
          }
          uint8_t* d = r0 + pixoff[j];
          if (kRowWarp) {
            // [left neighbour's d | a] [b | c] as one 8-byte store; the row's last word also stores [d | 0]
#pragma unroll
            for (int p = 0; p < L::P; ++p) {
              const float* v = p ? lo : hi;
              uint32_t& carry = p ? carry_lo : carry_hi;
              const uint32_t ad = tc::pack_f16(v[0], v[3]);                       // a low half, d high half
              uint32_t left = __shfl_up_sync(0xffffffffu, ad, 1);
              if (lane == 0) left = carry;
              if ((rowstart >> j) & 1u) left = 0u;                                 // column 0 of the padded plane
              carry = __shfl_sync(0xffffffffu, ad, 31);
              *reinterpret_cast<uint2*>(d + p * PIX_BYTES) = make_uint2(__byte_perm(left, ad, 0x5432), tc::pack_f16(v[1], v[2]));
              if ((rowend >> j) & 1u) *reinterpret_cast<uint32_t*>(d + p * PIX_BYTES + 8) = ad >> 16;   // [d | column 97 = 0]
            }
          } else {
            const uint32_t ad = tc::pack_f16(hi[0], hi[3]);                        // column x+1 is odd: a | (b,c) | d
            *reinterpret_cast<uint16_t*>(d) = (uint16_t)(ad & 0xffffu);
            *reinterpret_cast<uint32_t*>(d + 2) = tc::pack_f16(hi[1], hi[2]);
            *reinterpret_cast<uint16_t*>(d + 6) = (uint16_t)(ad >> 16);
            if (kSplit) {
              const uint32_t al = tc::pack_f16(lo[0], lo[3]);
              *reinterpret_cast<uint16_t*>(d + PIX_BYTES) = (uint16_t)(al & 0xffffu);
              *reinterpret_cast<uint32_t*>(d + PIX_BYTES + 2) = tc::pack_f16(lo[1], lo[2]);
              *reinterpret_cast<uint16_t*>(d + PIX_BYTES + 6) = (uint16_t)(al >> 16);
            }
          }
        }
      }
    }
    const long long fnext = mz.next;
    fetch(fnext);                                      // next frame's pixels travel while this one is convolved
    slot_sync<ST>(slot);

    // ================= conv1 + ReLU + pool: rounds of STRIP_TILES tiles of 128 pooled pixels
#pragma unroll 1
    for (int tile0 = 0, rnd = 0; tile0 < 9; tile0 += L::STRIP_TILES, ++rnd) {
      const int ntiles = min(L::STRIP_TILES, 9 - tile0);
      if (9 % L::STRIP_TILES == 0) {                   // full rounds only: hoisted addressing (+16 frame rows per round)
        const uint32_t radd = rnd * (16 * PIX_LD * 2);
        // all loads first, then all stores: source and destination live in the same buffer, so the compiler would
        // otherwise keep every load behind the previous store and expose the LDS latency once per item
#pragma unroll
        for (int p = 0; p < L::P; ++p) {
          uint4 v[GIT];
#pragma unroll
          for (int k = 0; k < GIT; ++k) {
            const uint8_t* src = r0 + p * PIX_BYTES + gsrc[k] + radd;
            v[k].x = *reinterpret_cast<const uint32_t*>(src);
            v[k].y = *reinterpret_cast<const uint32_t*>(src + 4);
            v[k].z = *reinterpret_cast<const uint32_t*>(src + PIX_LD * 2);
            v[k].w = *reinterpret_cast<const uint32_t*>(src + PIX_LD * 2 + 4);
          }
#pragma unroll
          for (int k = 0; k < GIT; ++k) *reinterpret_cast<uint4*>(r0 + p * L::PATCH_BYTES + gdst[k]) = v[k];
        }
      } else {
        // gather the 4x4 patches: item = (k-chunk, pooled pixel); chunk kc holds patch rows 2kc, 2kc+1
        for (int idx = tid; idx < ntiles * 256; idx += ST) {
          const int kc = idx / (ntiles * 128), pxs = idx % (ntiles * 128);
          const int Pp = tile0 * 128 + pxs, py = Pp / W1, px = Pp % W1;
#pragma unroll
          for (int p = 0; p < L::P; ++p) {
            const uint8_t* src = r0 + p * PIX_BYTES + ((2 * py + 2 * kc) * PIX_LD + 2 * px) * 2;
            uint4 v;
            v.x = *reinterpret_cast<const uint32_t*>(src);
            v.y = *reinterpret_cast<const uint32_t*>(src + 4);
            v.z = *reinterpret_cast<const uint32_t*>(src + PIX_LD * 2);
            v.w = *reinterpret_cast<const uint32_t*>(src + PIX_LD * 2 + 4);
            *reinterpret_cast<uint4*>(r0 + L::OFF_PATCH + p * L::PATCH_BYTES + kc * L::PATCH_CHUNK + pxs * 16) = v;
          }
        }
      }
      tc::fence_async_smem();
      slot_sync<ST>(slot);
      if (warp == 0 && tc::elect_one()) {
        tc::fence_after_sync();
        for (int tl = 0; tl < ntiles; ++tl) {
          const uint64_t ah = tc::make_desc(r0_u + L::OFF_PATCH + tl * 2048, L::PATCH_CHUNK, 128);
          const uint64_t bh = tc::make_desc(w1_u, L::P * W1_CHUNK, 128);
          if (kSplit) {
            // columns [0,32) = A_hi.B_hi (+ A_lo.B_hi), columns [32,64) = A_hi.B_lo; the epilogue adds the halves
            const uint64_t al = tc::make_desc(r0_u + L::OFF_PATCH + L::PATCH_BYTES + tl * 2048, L::PATCH_CHUNK, 128);
            tc::mma_f16(tmem + tl * C1W, ah, bh, idesc1s, false);
            tc::mma_f16(tmem + tl * C1W, al, bh, idesc1, true);
          } else {
            tc::mma_f16(tmem + tl * C1W, ah, bh, idesc1, false);
          }
        }
        tc::mma_commit(&mz.bar[0]);
      }
      tc::mbar_wait_bounded(&mz.bar[0], ph0);
      ph0 ^= 1;
      tc::fence_after_sync();
#pragma unroll
      for (int k = 0; k < E1; ++k) {
        const int tl = wg + k * NWG;
        if (tl < ntiles) {
          float v[32];
          tc::tmem_ld32(tq + tl * C1W, v);
          if (kSplit) {
            float w[32];
            tc::tmem_ld32(tq + tl * C1W + 32, w);
            tc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += w[i];
          } else {
            tc::tmem_ld_wait();
          }
          float hi[8], lo[8], bb[8];
          // biases as 16-byte broadcast loads (a scalar LDS per use made them a tenth of the kernel's shared-memory wavefronts)
          *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(sh.b1);
          *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(sh.b1 + 4);
          if (kSave) {                                                    // fp32 activations + pool argmax for the backward pass
            float a[8];
            uint32_t am = 0;
#pragma unroll
            for (int co = 0; co < 8; ++co) {                              // first maximum in row-major order (ATen's choice)
              float m = v[co];
              uint32_t arg = 0;
              if (v[8 + co] > m) { m = v[8 + co]; arg = 1; }
              if (v[16 + co] > m) { m = v[16 + co]; arg = 2; }
              if (v[24 + co] > m) { m = v[24 + co]; arg = 3; }
              am |= arg << (2 * co);
              a[co] = fmaxf(m + bb[co], 0.f);
            }
            const int Pp = (tile0 + tl) * 128 + row;
            uint8_t* rec = saved + f * SV_BYTES;
            reinterpret_cast<float4*>(rec + SV_A1)[2 * Pp] = make_float4(a[0], a[1], a[2], a[3]);
            reinterpret_cast<float4*>(rec + SV_A1)[2 * Pp + 1] = make_float4(a[4], a[5], a[6], a[7]);
            reinterpret_cast<uint16_t*>(rec + SV_IDX1)[Pp] = (uint16_t)am;
          }
#pragma unroll
          for (int co = 0; co < 8; ++co) {
            const float m4 = fmaxf(fmaxf(v[co], v[8 + co]), fmaxf(v[16 + co], v[24 + co]));
            hi[co] = fmaxf(m4 + bb[co], 0.f);
            if (kSplit) tc::split_f16(hi[co], hi[co], lo[co]);          // single pass: pack8 rounds once
          }
          uint32_t off;
          if (9 % L::STRIP_TILES == 0) {
            off = a1off[k] + rnd * (4 * P1_COLS * 16);                  // 8 pooled rows further = 4 plane rows
          } else {
            const int Pp = (tile0 + tl) * 128 + row;                    // pooled pixel -> a1 padded (py+1, px+1)
            const int rr = Pp / W1 + 1, cc = Pp % W1 + 1;
            off = ((rr & 1) * 2 + (cc & 1)) * P1_PLANE + ((rr >> 1) * P1_COLS + (cc >> 1)) * 16;
          }
          *reinterpret_cast<uint4*>(a1 + off) = pack8(hi);
          if (kSplit) *reinterpret_cast<uint4*>(a1 + A1_BYTES + off) = pack8(lo);
        }
      }
      tc::fence_before_sync();
      slot_sync<ST>(slot);                              // patches and TMEM columns are reused by the next round
    }

    // ================= conv2 + ReLU + pool: 3 tiles of 128 virtual pooled pixels (v = py*25 + px)
    // region R0 becomes a2: clear it (its zero border is conv3's padding); the frame plane and patches are dead
    for (int i = tid; i < L::P * A2_BYTES / 16; i += ST) reinterpret_cast<uint4*>(r0)[i] = make_uint4(0, 0, 0, 0);
    tc::fence_async_smem();
    slot_sync<ST>(slot);
    auto conv2_issue = [&](int tl, uint32_t dcol) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          // patch elements (r, 2h) and (r, 2h+1): column-parity planes 0 and 1 of row-parity plane r&1
          const uint32_t a_off = (r & 1) * 2 * P1_PLANE + ((r >> 1) * P1_COLS + h) * 16 + tl * 2048;
          const uint32_t b_off = (r * 4 + 2 * h) * (L::P * W2_CHUNK);
          const uint64_t ah = tc::make_desc(a1_u + a_off, P1_PLANE, 128);
          const uint64_t bh = tc::make_desc(w2_u + b_off, L::P * W2_CHUNK, 128);
          if (kSplit) {
            // columns [0,64) = A_hi.B_hi + A_lo.B_hi, columns [64,128) = A_hi.B_lo
            const uint64_t al = tc::make_desc(a1_u + A1_BYTES + a_off, P1_PLANE, 128);
            tc::mma_f16(tmem + dcol, ah, bh, idesc2s, (r | h) != 0);
            tc::mma_f16(tmem + dcol, al, bh, idesc2, true);
          } else {
            tc::mma_f16(tmem + dcol, ah, bh, idesc2, (r | h) != 0);
          }
        }
    };
    auto conv2_epilogue = [&](int tl, uint32_t dcol) {
      if (kSplit) {
        // 256-thread slots: warpgroup wg owns output channels 8wg..8wg+7 (= channel plane wg of a2); per sub-position
        // it adds the A_hi.B_lo half (columns 64..) to the main half, then max / bias / ReLU / hi-lo split
        float m4[8];
        uint32_t am = 0;                                                 // pool argmax, 2 bits per channel (saved for training)
#pragma unroll
        for (int sub = 0; sub < 4; ++sub) {
          float a[8], bq[8];
          tc::tmem_ld8(tq + dcol + sub * 16 + wg * 8, a);
          tc::tmem_ld8(tq + dcol + 64 + sub * 16 + wg * 8, bq);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float x = a[i] + bq[i];
            if (!kSave) { m4[i] = sub == 0 ? x : fmaxf(m4[i], x); }
            else if (sub == 0) { m4[i] = x; }
            else if (x > m4[i]) { m4[i] = x; am = (am & ~(3u << (2 * i))) | ((uint32_t)sub << (2 * i)); }
          }
        }
        if (a2off[tl] >= 0) {
          float hi[8], lo[8];
#pragma unroll
          float bb[8];
          *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(sh.b2 + wg * 8);
          *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(sh.b2 + wg * 8 + 4);
#pragma unroll
          for (int i = 0; i < 8; ++i) tc::split_f16(fmaxf(m4[i] + bb[i], 0.f), hi[i], lo[i]);
          const uint32_t off = (uint32_t)a2off[tl] + wg * PLANE2;
          if (kSave) {
            const int vp = tl * 128 + row, pix = (vp / P1_COLS) * W2 + vp % P1_COLS;
            uint8_t* rec = saved + f * SV_BYTES;
            float4* d = reinterpret_cast<float4*>(rec + SV_A2) + pix * 4 + wg * 2;
            d[0] = make_float4(fmaxf(m4[0] + bb[0], 0.f), fmaxf(m4[1] + bb[1], 0.f), fmaxf(m4[2] + bb[2], 0.f), fmaxf(m4[3] + bb[3], 0.f));
            d[1] = make_float4(fmaxf(m4[4] + bb[4], 0.f), fmaxf(m4[5] + bb[5], 0.f), fmaxf(m4[6] + bb[6], 0.f), fmaxf(m4[7] + bb[7], 0.f));
            reinterpret_cast<uint16_t*>(rec + SV_IDX2)[pix * 2 + wg] = (uint16_t)am;
          }
          *reinterpret_cast<uint4*>(r0 + off) = pack8(hi);
          *reinterpret_cast<uint4*>(r0 + A2_BYTES + off) = pack8(lo);
        }
      } else {
        float v[32], w[32];
        tc::tmem_ld32(tq + dcol, v);
        tc::tmem_ld32(tq + dcol + 32, w);
        tc::tmem_ld_wait();
        if (a2off[tl] >= 0) {
          float hi[16], bb[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) *reinterpret_cast<float4*>(bb + 4 * i) = *reinterpret_cast<const float4*>(sh.b2 + 4 * i);
#pragma unroll
          for (int co = 0; co < 16; ++co) {
            const float m4 = fmaxf(fmaxf(v[co], v[16 + co]), fmaxf(w[co], w[16 + co]));
            hi[co] = fmaxf(m4 + bb[co], 0.f);
          }
          const uint32_t off = (uint32_t)a2off[tl];
          *reinterpret_cast<uint4*>(r0 + off) = pack8(hi);
          *reinterpret_cast<uint4*>(r0 + PLANE2 + off) = pack8(hi + 8);
        }
      }
    };
    {
      // two tiles of TMEM columns: tiles 0 and 1 in flight, tile 2 re-uses tile 0's columns once its epilogue is
      // done; every tile has its own commit, so an epilogue overlaps the next tile's MMAs
      if (warp == 0 && tc::elect_one()) {
        tc::fence_after_sync();
        conv2_issue(0, 0);
        tc::mma_commit(&mz.bar[0]);
        conv2_issue(1, C2W);
        tc::mma_commit(&mz.bar[1]);
      }
      tc::mbar_wait_bounded(&mz.bar[0], ph0); ph0 ^= 1;
      tc::fence_after_sync();
      conv2_epilogue(0, 0);
      tc::fence_before_sync();
      slot_sync<ST>(slot);
      if (warp == 0 && tc::elect_one()) {
        tc::fence_after_sync();
        conv2_issue(2, 0);
        tc::mma_commit(&mz.bar[0]);
      }
      tc::mbar_wait_bounded(&mz.bar[1], ph1); ph1 ^= 1;
      tc::fence_after_sync();
      conv2_epilogue(1, C2W);
      tc::mbar_wait_bounded(&mz.bar[0], ph0); ph0 ^= 1;
      tc::fence_after_sync();
      conv2_epilogue(2, 0);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    slot_sync<ST>(slot);

    // ================= conv3 + ReLU + mean: 3 tiles of 16 rows x 8 columns (12 rows real), 9 taps
    // every tile has its own commit (bar[0..2]): the channel sums of tile i are taken while tiles i+1.. are still in the
    // tensor core
    if (warp == 0 && tc::elect_one()) {
      tc::fence_after_sync();
#pragma unroll
      for (int tx = 0; tx < 3; ++tx) {
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int dy = tap / 3, dx = tap % 3;
          const uint32_t a_off = (dy * PITCH2 + tx * 8 + dx) * 16;
          const uint64_t db = tc::make_desc(w3_u + 2 * tap * L::W3_CHUNK, L::W3_CHUNK, 128);
          tc::mma_f16(tmem + tx * COLS3, tc::make_desc(r0_u + a_off, PLANE2, PITCH2 * 16), db, idesc3_hi, tap > 0);
          if (kSplit) tc::mma_f16(tmem + tx * COLS3, tc::make_desc(r0_u + A2_BYTES + a_off, PLANE2, PITCH2 * 16), db, idesc3_lo, true);
        }
        tc::mma_commit(&mz.bar[tx]);
      }
    }
    float acc[C3P];
#pragma unroll
    for (int c = 0; c < C3P; ++c) acc[c] = 0.f;
#pragma unroll
    for (int tx = 0; tx < 3; ++tx) {
      // every thread follows all three barriers (phase bookkeeping), only the tile's owners read its columns
      if (tx == 0) { tc::mbar_wait_bounded(&mz.bar[0], ph0); ph0 ^= 1; }
      if (tx == 1) { tc::mbar_wait_bounded(&mz.bar[1], ph1); ph1 ^= 1; }
      if (tx == 2) { tc::mbar_wait_bounded(&mz.bar[2], ph2); ph2 ^= 1; }
      tc::fence_after_sync();
      if (q < 3 && tx % NWG == wg) {
        float v[32];
        tc::tmem_ld32(tq + tx * COLS3, v);
        if (kSplit) {
          float w[32];
          tc::tmem_ld32(tq + tx * COLS3 + 32, w);
          tc::tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < C3; ++c) v[c] += w[c];
        } else {
          tc::tmem_ld_wait();
        }
        uint32_t on = 0;                                                 // relu'(conv3) per channel (saved for training)
#pragma unroll
        for (int c4 = 0; c4 < C3 / 4; ++c4) {
          const float4 bq = *reinterpret_cast<const float4*>(sh.b3 + 4 * c4);
          const float x0 = v[4 * c4 + 0] + bq.x, x1 = v[4 * c4 + 1] + bq.y, x2 = v[4 * c4 + 2] + bq.z, x3 = v[4 * c4 + 3] + bq.w;
          acc[4 * c4 + 0] += fmaxf(x0, 0.f);
          acc[4 * c4 + 1] += fmaxf(x1, 0.f);
          acc[4 * c4 + 2] += fmaxf(x2, 0.f);
          acc[4 * c4 + 3] += fmaxf(x3, 0.f);
          if (kSave) on |= ((x0 > 0.f ? 1u : 0u) | (x1 > 0.f ? 2u : 0u) | (x2 > 0.f ? 4u : 0u) | (x3 > 0.f ? 8u : 0u)) << (4 * c4);
        }
        if (kSave)                                                       // tile tx: image rows row/8, columns 8 tx + row%8
          reinterpret_cast<uint32_t*>(saved + f * SV_BYTES + SV_MASK3)[(row >> 3) * W2 + tx * 8 + (row & 7)] = on;
      }
    }
    if (q < 3) {
      // one transposing reduction per warp: lane c ends with channel c's sum over the warp's pixels
#pragma unroll
      for (int off = 16, n = 16; off >= 1; off >>= 1, n >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
          const float send = up ? acc[i] : acc[i + n], keep = up ? acc[i + n] : acc[i];
          acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      mz.part[wg * 3 + q][lane] = acc[0];
    }
    tc::fence_before_sync();
    slot_sync<ST>(slot);
    if (tid < E) {                                      // warp 0: deterministic channel means, then the 24 -> 32 fc
      float cs = 0.f;
#pragma unroll
      for (int i = 0; i < 3 * NWG; ++i) cs += mz.part[i][lane];
      cs *= 1.0f / (float)(H2 * W2);
      if (kSave && tid < C3) reinterpret_cast<float*>(saved + f * SV_BYTES + SV_MEAN3)[tid] = cs;
      float a = sh.bfc[tid];
#pragma unroll
      for (int c = 0; c < C3; ++c) a = fmaf(sh.wfc[tid][c], __shfl_sync(0xffffffffu, cs, c), a);
      o[tid] = a;
    }
    f = fnext;
    slot_sync<ST>(slot);                                // R0 and mz.sum are reused by the next frame
  }
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc<512>(sh.tmem_base);
}

template <bool kSplit, int kSlots>
constexpr size_t smem_total() {
  return kSlots * Lay<kSplit>::SLOT_BYTES + Lay<kSplit>::P * (W1_BYTES + W2_BYTES) + Lay<kSplit>::W3_BYTES +
         kSlots * sizeof(SlotMisc) + sizeof(Shared);
}

}  // namespace cnn3v3
}  // namespace ssb

// passes = 1 launch, called by ssb_roi_cnn_fwd_tc3 (roi_cnn_tc3.cu) after it has validated the arguments.
int ssb_roi_cnn_fwd_v3_(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T, const float* const* weights,
                        int32_t standardize, float* out, int64_t out_stride, void* stream) {
  using namespace ssb;
  using namespace ssb::cnn3v3;
  Weights gw{weights[0], weights[1], weights[2], weights[3], weights[4], weights[5], weights[6], weights[7]};
  const int64_t frames = B * T;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* counter = nullptr;
  if (int rc = ssb_scratch_counter_(st, &counter)) return rc;
  constexpr int kSlots = 4, ST = 128;
  constexpr size_t smem = smem_total<false, kSlots>();
  static_assert(smem <= 232448, "fp16-mode smem budget");
  const int64_t want = (frames + kSlots - 1) / kSlots;
  const int grid = (int)(want < kSMs ? want : kSMs);
  SSB_CUDA(cudaFuncSetAttribute(roi_cnn_v3_kernel<false, kSlots, ST, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  roi_cnn_v3_kernel<false, kSlots, ST, false><<<grid, ST * kSlots, smem, st>>>(R, lengths, (int)B, (int)T, gw, standardize, out, out_stride, counter, nullptr);
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
