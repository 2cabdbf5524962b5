// K2 v4 -- fused mouth-ROI CNN forward, all three convolutions on tcgen05, pooling folded into N.
//
// Arithmetic contract: train_model_official.py:286-294 + TinyROICNN (209-229); live semantics with standardize=0.
//
// Frame staging (r2): every frame slot owns a 4,608-byte shared-memory buffer that ONE `cp.async.bulk` (TMA, 1-D) fills
// with the raw u8 frame straight from HBM, completing on the slot's mbarrier; the copy of a slot's NEXT frame is issued
// as soon as the current frame's pixels have been consumed and travels under the rest of the frame (conv1 tail, conv2,
// conv3).  Frames are handed out by an atomic counter; padded frames (t >= lengths[b]) are skipped by the acquiring warp,
// which writes their zero embeddings itself.
//
// conv1 on EXACT inputs (r2): u8 pixels are exactly representable in fp16, so the A operand of conv1 is the RAW pixel
// patch (built from the staged bytes with two PRMTs and two HSUB2 per four pixels, no fp32 arithmetic) and the per-frame
// standardisation moves into the epilogue:  conv(w, (x - mu)/sigma) = (conv(w, x_raw) - mu * ws) / sigma, where ws is the
// sum of the weights over the taps that fall inside the frame (a 3 x 3 table of border classes per output channel; zero
// padding of the STANDARDISED frame means "mu" outside the raw one).  In split mode that removes the activation's lo part
// and one of the two conv1 MMAs; in both modes it removes the standardised plane, the patch gather from it and a
// synchronisation.  The fp32 statistics are exact integers (64-bit N*S2 - S1^2) as before.
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
// pooled pixel) are built two tiles (256 pooled pixels) at a time into a double buffer: the patches of round r+1 are
// written while the tensor core works on round r.
//
// One CTA hosts kSlots independent "frame slots" (named barriers, own TMEM columns, own activation buffers) that share
// one copy of the weight operands; while one slot waits for its MMAs the others compute.  fp16 mode: four 128-thread
// slots; kSplit: two 256-thread slots, operands are fp16 hi + lo and D += Ahi.Bhi + Ahi.Blo + Alo.Bhi (fp32-parity grade;
// conv1: x.Whi + x.Wlo with x exact).
#include <stdlib.h>

#include "common.cuh"
#include "tc.cuh"

namespace ssb {
namespace cnn3 {

constexpr int H0 = 48, W0 = 96, NPIX = H0 * W0;
constexpr int H1 = 24, W1 = 48, C1 = 8;                // pooled conv1 output
constexpr int H2 = 12, W2 = 24, C2 = 16;               // pooled conv2 output
constexpr int C3 = 24, C3P = 32, E = 32;
// Training: activations handed to the backward kernel instead of being recomputed there (one record per frame, the
// layout ssb_roi_cnn_bwd reads -- see roi_cnn.cuh: SAVED_*).
constexpr int SV_A1 = 0, SV_A2 = 36864, SV_IDX1 = 55296, SV_IDX2 = 57600, SV_MASK3 = 58752, SV_MEAN3 = 59904, SV_BYTES = 60000;
constexpr int RING_BYTES = NPIX;                       // raw u8 frame staged by one bulk copy
constexpr int C1_ROUNDS = 3;                           // conv1: 9 tiles in three rounds -- every round is a trip through the
                                                       // tensor pipe's queue behind the other slots' batches
constexpr int P1_ROWS = 13, P1_COLS = 25;              // a1 parity plane: [13][25] pixels x 16 B
constexpr int P1_PLANE = P1_ROWS * P1_COLS * 16;       // 5,200
constexpr int A1_BYTES = 4 * P1_PLANE;                 // 20,800
constexpr int PITCH2 = 26, ROWS2 = 18;                 // a2: 2 channel planes x [18][26] pixels x 16 B (as in v1)
constexpr int PLANE2 = ROWS2 * PITCH2 * 16;            // 7,488
constexpr int A2_BYTES = 2 * PLANE2;                   // 14,976
constexpr int W2_CHUNK = 64 * 16;                      // bytes per k-chunk of the conv2 weight operand
constexpr int W2_BYTES = 16 * W2_CHUNK;                // 16,384 per part

// kSplit: weights as fp16 hi + lo.  kALo: activations as hi + lo too (passes = 3; false = passes 2: activations rounded to
// fp16 once, products x.W_hi + x.W_lo).  kWide: 256-thread slots (two per CTA) with hi|lo stacked along N; narrow slots
// (128 threads, four per CTA, 128 TMEM columns each) issue the hi and lo products as two MMAs into the same columns.
template <bool kSplit, bool kALo = kSplit, bool kWide = kSplit>
struct Lay {
  static_assert(!kALo || (kSplit && kWide), "activation lo parts exist in the wide split mode only");
  static constexpr int TPR = kWide ? 4 : 3;            // conv1 tiles per round: 4 + 4 + 1 (wide) or 3 + 3 + 3 (narrow: smaller patch buffer)
  static constexpr int LAST_TILES = 9 - 2 * TPR;
  static constexpr int PATCH_PX = TPR * 128;           // pooled pixels per round
  static constexpr int PATCH_CHUNK = PATCH_PX * 16;    // one k-chunk (two patch rows) of a round
  static constexpr int PATCH_BUF = 2 * PATCH_CHUNK;    // 16,384 / 12,288 per round
  static constexpr int P = kSplit ? 2 : 1;             // weight parts
  static constexpr int PA = kALo ? 2 : 1;              // activation parts
  static constexpr int W3_ROWS = kSplit ? 64 : 32;     // conv3 keeps hi|lo stacked along N (v1 scheme)
  static constexpr int W3_CHUNK = W3_ROWS * 16, W3_BYTES = 18 * W3_CHUNK;
  static constexpr int W1_ROWS = kSplit ? 64 : 32;     // conv1 weights: rows [0,32) hi, [32,64) lo (split mode; the input is exact)
  static constexpr int W1_CHUNK = W1_ROWS * 16, W1_BYTES = 2 * W1_CHUNK;
  // conv1 double buffering (patches and TMEM columns of round r+1 while round r is in the tensor core / epilogue): split
  // mode only -- its slots have 256 TMEM columns and R0 is 30 KB anyway; the four fp16-mode slots have 128 columns each
  static constexpr int NBUF1 = kWide ? 2 : 1;
  // per-slot region R0: conv1 patch buffer(s) during conv1, a2 afterwards
  static constexpr int R0_A = NBUF1 * PATCH_BUF, R0_B = PA * A2_BYTES;
  static constexpr int R0_BYTES = ((R0_A > R0_B ? R0_A : R0_B) + 127) / 128 * 128;
  static constexpr int A1_ALLOC = (PA * A1_BYTES + 1408 + 127) / 128 * 128;    // + over-read slack of the last conv2 tile (1,360 B)
  static constexpr int SLOT_BYTES = RING_BYTES + R0_BYTES + A1_ALLOC;
};

struct SlotMisc {
  float part[9][C3P];                                  // conv3 channel sums per (tile, lane quarter): fixed-order reduction

  uint32_t red[16];
  uint64_t bar[3];
  uint64_t rbar;                                       // frame ring: the bulk copy's bytes complete here
  long long next;                                      // next frame of this slot (dynamic hand-out)
};

struct Shared {                                        // one per CTA, after the slots and the weight operands
  alignas(16) float b1[C1];
  alignas(16) float b2[C2];
  alignas(16) float b3[C3P];
  alignas(16) float ws1[9][C1];                        // conv1: sum of the effective weights over the in-frame taps, per border class
  uint32_t gtab[288];                                  // conv1 group order: [0,13) a1 cell offset of pixel 4g-1, bit 15 (py+1)&1, [16,21) py, [21,25) g
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
    // split mode: rows [0,32) of a k-chunk = hi, rows [32,64) = lo, so that x . [W_hi | W_lo] is ONE MMA with N = 64
    const uint32_t off = kc * L::W1_CHUNK + (n >> 3) * 128 + (n & 7) * 16;
    *reinterpret_cast<uint4*>(w1op + off) = pack8(hi);
    if (kSplit) *reinterpret_cast<uint4*>(w1op + 32 * 16 + off) = pack8(lo);
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

// conv1 row order.  The 1,152 pooled pixels of a frame are handled as 288 groups of four x-adjacent pixels (one thread
// builds a group's patches from one 8-byte + two 4-byte loads per image row).  Groups are ordered interior first
// (rows 1..22, columns 4..43), then the 44 groups that touch the left / right frame border, then the 24 of the first and
// last pooled row, so that the border-class arithmetic of the epilogue is confined to the last tiles; inside a 128-row
// tile, row m = e*32 + l holds pixel e of the tile's group l, which makes both the patch stores of the builder (fixed e,
// consecutive l) and the epilogue's TMEM lanes (warp = e, lane = l) contiguous.
__device__ __forceinline__ void group_pos(int og, int& py, int& g) {
  if (og < 220) { py = 1 + og / 10; g = 1 + og % 10; }
  else if (og < 264) { const int k = og - 220; py = 1 + (k >> 1); g = (k & 1) ? 11 : 0; }
  else { const int k = og - 264; py = k < 12 ? 0 : H1 - 1; g = k < 12 ? k : k - 12; }
}

// conv1's effective weight as the tensor core sees it: the fp16 value (single pass) or hi + lo (split mode)
template <bool kSplit>
__device__ __forceinline__ float eff_weight(float w) {
  float hi, lo;
  tc::split_f16(w, hi, lo);
  if (!kSplit) return hi;
  float lh, ll;
  tc::split_f16(lo, lh, ll);            // lo is rounded to fp16 once more when it is packed
  return hi + lh;
}

template <bool kSplit, bool kALo, int kSlots, int ST, bool kSave, bool kStack = kSplit && ST == 256>
__global__ void __launch_bounds__(ST * kSlots, 1)
roi_cnn_tc3_kernel(const uint8_t* __restrict__ R, const int64_t* __restrict__ lengths, int B, int T,
                   Weights gw, int standardize, float* __restrict__ out, int64_t out_stride,
                   unsigned int* __restrict__ counter, uint8_t* __restrict__ saved) {
  // kStack: the hi and lo weight parts sit side by side along N of ONE MMA (conv2: N = 128, conv3: N = 64) -- an MMA with
  // M = 128, K = 16 costs about 32 + N/4 cycles of the tensor pipe (the 4 KB A operand is fetched per instruction:
  // profiles/microbench/r2_tc_micro3.txt), so one N = 128 MMA (64) is cheaper than two N = 64 MMAs (96).  A narrow slot's
  // 128 TMEM columns then hold one conv2 tile (tiles run one after the other) or two conv3 tiles.
  constexpr bool kWide = ST == 256;
  static_assert(!kStack || kSplit, "only split weights are stacked");
  static_assert(ST == 128 || kStack, "256-thread slots exist in the stacked split mode only");
  static_assert(!kSave || kALo, "saved activations come from the full split mode");
  using L = Lay<kSplit, kALo, kWide>;
  constexpr int NWG = ST / 128;                        // warps per TMEM lane quarter
  constexpr int C1_TPR = L::TPR, C1_LAST = L::LAST_TILES, PATCH_CHUNK = L::PATCH_CHUNK, PATCH_BUF = L::PATCH_BUF;
  constexpr int TMEM_PER_SLOT = 512 / kSlots;
  constexpr int C1W = 32, C2W = kStack ? 128 : 64;     // TMEM columns per conv1 / conv2 tile
  constexpr int C2_FLY = TMEM_PER_SLOT / C2W;          // conv2 tiles in flight: 2 (tile 2 re-uses tile 0's columns) or 1
  static_assert(C2_FLY == 1 || C2_FLY == 2, "conv2 runs one or two tiles of TMEM columns, committed tile by tile");
  constexpr int E1 = C1_TPR / NWG;                     // conv1 epilogue tiles per thread and round: 4 (ST=128) or 2 (ST=256)
  constexpr int NB1 = L::NBUF1;
  static_assert(C1_TPR % NWG == 0 && ST >= 64, "conv1 round / slot shape");
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* w1op = smem + kSlots * L::SLOT_BYTES;
  uint8_t* w2op = w1op + L::W1_BYTES;
  uint8_t* w3op = w2op + L::P * W2_BYTES;
  SlotMisc* smz = reinterpret_cast<SlotMisc*>(w3op + L::W3_BYTES);
  Shared& sh = *reinterpret_cast<Shared*>(smz + kSlots);
  // slot and warp index through a shuffle: the compiler then knows they are warp-uniform (uniform registers and branches
  // for everything derived from them, in particular the MMA descriptors)
  const int slot = __shfl_sync(0xffffffffu, (int)(threadIdx.x / ST), 0), tid = threadIdx.x % ST, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  uint8_t* ring = smem + slot * L::SLOT_BYTES;         // raw u8 frame (bulk-copy destination)
  uint8_t* r0 = ring + RING_BYTES;                     // conv1 patches, later a2
  uint8_t* a1 = r0 + L::R0_BYTES;
  SlotMisc& mz = smz[slot];

  // ---- one-time setup (whole CTA)
  build_weights<kSplit>(w1op, w2op, w3op, gw, threadIdx.x, ST * kSlots);
  for (int i = threadIdx.x; i < E * C3; i += ST * kSlots) sh.wfc[i / C3][i % C3] = gw.wfc[i];
  if (threadIdx.x < C1) sh.b1[threadIdx.x] = gw.b1[threadIdx.x];
  if (threadIdx.x < C2) sh.b2[threadIdx.x] = gw.b2[threadIdx.x];
  if (threadIdx.x < C3P) sh.b3[threadIdx.x] = threadIdx.x < C3 ? gw.b3[threadIdx.x] : 0.f;
  if (threadIdx.x < E) sh.bfc[threadIdx.x] = gw.bfc[threadIdx.x];
  if (threadIdx.x >= 128 && threadIdx.x < 128 + 9 * C1) {
    // border-class table of conv1: class (rc, cc) in {first, interior, last} x {first, interior, last} row / column of the
    // frame; ws = sum of the effective weights over the taps whose input pixel lies inside the frame
    const int e = (threadIdx.x - 128) / C1, co = (threadIdx.x - 128) % C1, rc = e / 3, cc = e % 3;
    float acc = 0.f;
    for (int dy = 0; dy < 3; ++dy)
      for (int dx = 0; dx < 3; ++dx) {
        const bool in = !(rc == 0 && dy == 0) && !(rc == 2 && dy == 2) && !(cc == 0 && dx == 0) && !(cc == 2 && dx == 2);
        if (in) acc += eff_weight<kSplit>(gw.w1[co * 9 + dy * 3 + dx]);
      }
    sh.ws1[e][co] = acc;
  }
  for (int og = threadIdx.x; og < 288; og += ST * kSlots) {
    int py, g;
    group_pos(og, py, g);
    const int rr = py + 1;
    sh.gtab[og] = (uint32_t)(((rr >> 1) * P1_COLS + 2 * g) * 16) | ((uint32_t)(rr & 1) << 15) | ((uint32_t)py << 16) | ((uint32_t)g << 21);
  }
  for (int i = tid; i < L::A1_ALLOC / 16; i += ST) reinterpret_cast<uint4*>(a1)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x < 32) tc::tmem_alloc<512>(&sh.tmem_base);
  if (tid == 32) {
    tc::mbar_init(&mz.bar[0], 1); tc::mbar_init(&mz.bar[1], 1); tc::mbar_init(&mz.bar[2], 1); tc::mbar_init(&mz.rbar, 1);
    tc::mbar_fence_init();
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = __shfl_sync(0xffffffffu, sh.tmem_base, 0) + slot * TMEM_PER_SLOT;
  const uint32_t r0_u = tc::smem_u32(r0), a1_u = tc::smem_u32(a1), ring_u = tc::smem_u32(ring);
  const uint32_t w1_u = tc::smem_u32(w1op), w2_u = tc::smem_u32(w2op), w3_u = tc::smem_u32(w3op);
  constexpr uint32_t idesc1 = tc::make_idesc_f16(128, 32), idesc2 = tc::make_idesc_f16(128, 64);
  constexpr uint32_t idesc2s = tc::make_idesc_f16(128, 128);            // hi|lo stacked along N
  static_assert(NB1 * C1_TPR * C1W <= TMEM_PER_SLOT, "conv1 rounds do not fit the slot's TMEM columns");
  constexpr uint32_t idesc3_hi = tc::make_idesc_f16(128, kStack ? 64 : 32), idesc3_lo = tc::make_idesc_f16(128, 32);
  constexpr int COLS3 = kStack ? 64 : 32;
  constexpr int C3_FLY = TMEM_PER_SLOT / COLS3 >= 3 ? 3 : 2;           // conv3 tiles in flight
  static_assert(TMEM_PER_SLOT >= 2 * COLS3, "conv3 needs two tiles of TMEM columns");
  // base descriptors, built once: the issuing thread only adds compile-time offsets to them
  const uint64_t d_patch = tc::make_desc(r0_u, PATCH_CHUNK, 128), d_w1 = tc::make_desc(w1_u, L::W1_CHUNK, 128);
  const uint64_t d_a1 = tc::make_desc(a1_u, P1_PLANE, 128), d_w2 = tc::make_desc(w2_u, L::P * W2_CHUNK, 128);
  const uint64_t d_a2 = tc::make_desc(r0_u, PLANE2, PITCH2 * 16), d_w3 = tc::make_desc(w3_u, L::W3_CHUNK, 128);
  uint32_t ph0 = 0, ph1 = 0, ph2 = 0, rph = 0;
  const int q = warp & 3, wg = warp >> 2;
  const int row = q * 32 + lane;                       // this thread's row of every 128-row tile (= TMEM lane)
  const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);
  const uint32_t* ring32 = reinterpret_cast<const uint32_t*>(ring);

  int a2off[3];                                        // conv2 epilogue destination per tile (-1: virtual pixel)
#pragma unroll
  for (int tl = 0; tl < 3; ++tl) {
    const int vp = tl * 128 + row, py = vp / P1_COLS, px = vp % P1_COLS;
    a2off[tl] = (py < H2 && px < W2) ? ((py + 1) * PITCH2 + px + 1) * 16 : -1;
  }

  const long long frames = (long long)B * T;
  const long long handout_base = (long long)gridDim.x * kSlots;
  // ---- frame hand-out (warp 1 of the slot, all lanes): first live frame at or after `cand` in hand-out order; the
  // padded frames passed on the way get their zero embedding here (E = 32 = one lane per output)
  auto acquire = [&](long long cand) -> long long {
    while (cand < frames) {
      const bool lv = !lengths || (int)(cand % T) < (int)lengths[cand / T];
      if (lv) break;
      out[cand * out_stride + lane] = 0.f;
      unsigned int h = 0;
      if (lane == 0) h = atomicAdd(counter, 1u);
      cand = handout_base + __shfl_sync(0xffffffffu, h, 0);
    }
    return cand;
  };
  auto start_copy = [&](long long fr) {                // one TMA bulk copy: HBM frame -> this slot's ring buffer
    if (lane == 0 && fr < frames) {
      tc::mbar_arrive_expect_tx(&mz.rbar, RING_BYTES);
      tc::bulk_copy_g2s(ring_u, R + fr * NPIX, RING_BYTES, &mz.rbar);
    }
  };
  // ---- conv1 patches of round r (groups 64 r .. 64 r + 63): item = (k-chunk, group); a chunk holds two patch rows of
  // four raw pixels each, converted u8 -> fp16 exactly (0x6400 | b is 1024 + b; minus 1024).  A group's four pooled pixels
  // need image columns 8g-1 .. 8g+8 of each row: bytes 3 of word 2g-1, words 2g and 2g+1, byte 0 of word 2g+2.
  auto build = [&](int r) {
    const int ng = (r == C1_ROUNDS - 1) ? 32 * C1_LAST : 32 * C1_TPR;     // groups of this round
#pragma unroll
    for (int it0 = 0; it0 < 2 * 32 * C1_TPR; it0 += ST) {
      const int it = it0 + tid;
      if (it >= 2 * ng) break;
      const int kc = it >= ng ? 1 : 0, gl = it - kc * ng;
      const uint32_t gt = sh.gtab[r * 32 * C1_TPR + gl];
      const int py = (gt >> 16) & 31, g = gt >> 21;
      const int y0 = 2 * py - 1 + 2 * kc;              // patch rows 2kc, 2kc+1 = image rows y0, y0+1
      uint32_t o[4][4];
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const int y = y0 + rr;
        uint32_t wm = 0, wp = 0;
        uint2 w01 = make_uint2(0u, 0u);
        if ((unsigned)y < (unsigned)H0) {
          const uint32_t* rp = ring32 + y * (W0 / 4) + 2 * g;
          w01 = *reinterpret_cast<const uint2*>(rp);
          if (g > 0) wm = rp[-1];
          if (g < W1 / 4 - 1) wp = rp[2];
        }
        uint32_t b4[4];
        b4[0] = __funnelshift_r(wm, w01.x, 24);        // columns 8g-1 .. 8g+2
        b4[1] = __funnelshift_r(w01.x, w01.y, 8);      //         8g+1 .. 8g+4
        b4[2] = __funnelshift_r(w01.x, w01.y, 24);     //         8g+3 .. 8g+6
        b4[3] = __funnelshift_r(w01.y, wp, 8);         //         8g+5 .. 8g+8
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          uint32_t h01 = __byte_perm(b4[e], 0x64646464u, 0x5140), h23 = __byte_perm(b4[e], 0x64646464u, 0x5342);
          asm("sub.f16x2 %0, %0, %1;" : "+r"(h01) : "r"(0x64006400u));
          asm("sub.f16x2 %0, %0, %1;" : "+r"(h23) : "r"(0x64006400u));
          o[e][2 * rr] = h01; o[e][2 * rr + 1] = h23;
        }
      }
      uint8_t* pb = r0 + (r % NB1) * PATCH_BUF + kc * PATCH_CHUNK + ((gl >> 5) * 128 + (gl & 31)) * 16;
#pragma unroll
      for (int e = 0; e < 4; ++e) *reinterpret_cast<uint4*>(pb + e * 32 * 16) = make_uint4(o[e][0], o[e][1], o[e][2], o[e][3]);
    }
  };
  auto conv1_issue = [&](int r) {                      // tid 0: the round's tiles, K = 16, N = 32: x.W (fp16) or x.W_hi + x.W_lo
    const int nt = (r == C1_ROUNDS - 1) ? C1_LAST : C1_TPR;
    for (int tl = 0; tl < nt; ++tl) {
      const uint64_t da = tc::desc_add(d_patch, (r % NB1) * PATCH_BUF + tl * 2048);
      const uint32_t dcol = tmem + (r % NB1) * (C1_TPR * C1W) + tl * C1W;
      tc::mma_f16(dcol, da, d_w1, idesc1, false);
      if (kSplit) tc::mma_f16(dcol, da, tc::desc_add(d_w1, 32 * 16), idesc1, true);
    }
    tc::mma_commit(&mz.bar[r % NB1]);
  };

  long long f = frames;
  if (warp == 1) {
    f = acquire((long long)blockIdx.x * kSlots + slot);
    start_copy(f);
    if (lane == 0) mz.next = f;
  }
  slot_sync<ST>(slot);
  f = mz.next;
  while (f < frames) {
    unsigned int handout = 0;
    if (warp == 1 && lane == 0) handout = atomicAdd(counter, 1u);      // candidate for this slot's NEXT frame
    float* o = out + f * out_stride;
    tc::mbar_wait_bounded(&mz.rbar, rph);              // the frame's bytes have landed
    rph ^= 1;
    // ================= P0: exact integer statistics of the raw frame; first round of conv1 patches
    {
      uint32_t s1 = 0, s2 = 0;
#pragma unroll
      for (int i = 0; i < (NPIX / 4 + ST - 1) / ST; ++i) {
        const int w = i * ST + tid;
        const uint32_t v = (NPIX / 4) % ST == 0 || w < NPIX / 4 ? ring32[w] : 0u;
        s1 = __dp4a(v, 0x01010101u, s1);
        s2 = __dp4a(v, v, s2);
      }
      s1 = __reduce_add_sync(0xffffffffu, s1);
      s2 = __reduce_add_sync(0xffffffffu, s2);
      if (lane == 0) { mz.red[warp] = s1; mz.red[8 + warp] = s2; }
    }
    build(0);
    tc::fence_async_smem();
    slot_sync<ST>(slot);
    if (warp == 0 && tc::elect_one()) { tc::fence_after_sync(); conv1_issue(0); }
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
    const float nmean = -mean_u;
    float cint[8];                                     // interior pixels: relu(inv_s * max_sub(v) + cint), cint = b1 - mu/sigma * ws(interior)
    {
      float wsi[8], bb[8];
      *reinterpret_cast<float4*>(wsi) = *reinterpret_cast<const float4*>(sh.ws1[4]);
      *reinterpret_cast<float4*>(wsi + 4) = *reinterpret_cast<const float4*>(sh.ws1[4] + 4);
      *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(sh.b1);
      *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(sh.b1 + 4);
#pragma unroll
      for (int co = 0; co < 8; ++co) cint[co] = fmaf(nmean * inv_s, wsi[co], bb[co]);
    }

    // ================= conv1 + ReLU + pool: rounds of 4, 4 and 1 tiles.  Split mode: the patches of round r+1 are built
    // while the tensor core works on round r, TMEM columns and patch buffers alternate.  fp16 mode (one buffer): the
    // patches of round r+1 are built as soon as round r's MMAs have consumed theirs, before its epilogue.
    long long fnext = frames, cand = frames;
    int64_t cand_len = 0;
#pragma unroll 1
    for (int r = 0; r < C1_ROUNDS; ++r) {
      if (NB1 == 2 && r + 1 < C1_ROUNDS) build(r + 1);
      if (NB1 == 2 && (r & 1)) { tc::mbar_wait_bounded(&mz.bar[1], ph1); ph1 ^= 1; } else { tc::mbar_wait_bounded(&mz.bar[0], ph0); ph0 ^= 1; }
      tc::fence_after_sync();
      if (NB1 == 1 && r + 1 < C1_ROUNDS) build(r + 1);
      const int nt = (r == C1_ROUNDS - 1) ? C1_LAST : C1_TPR;
#pragma unroll
      for (int k = 0; k < E1; ++k) {
        const int tl = k * NWG + wg;
        if (tl < nt) {
          float v[32];
          tc::tmem_ld32(tq + (r % NB1) * (C1_TPR * C1W) + tl * C1W, v);
          tc::tmem_ld_wait();
          const uint32_t gt = sh.gtab[(r * C1_TPR + tl) * 32 + lane];
          const int py = (gt >> 16) & 31, g = gt >> 21;
          const int px = 4 * g + q, Pp = py * W1 + px;                    // this thread's pooled pixel (row m = q*32 + lane of the tile)
          const bool border = py == 0 || py == H1 - 1 || px == 0 || px == W1 - 1;
          float hi[8], lo[8];
          uint32_t am = 0;
          if (!__any_sync(0xffffffffu, border)) {
            // all four conv pixels of every lane are interior: one class, the maximum commutes with the affine map
#pragma unroll
            for (int co = 0; co < 8; ++co) {
              float m = v[co];
              if (!kSave) {
                m = fmaxf(fmaxf(m, v[8 + co]), fmaxf(v[16 + co], v[24 + co]));
              } else {                                                    // first maximum in row-major order (ATen's choice)
                uint32_t arg = 0;
                if (v[8 + co] > m) { m = v[8 + co]; arg = 1; }
                if (v[16 + co] > m) { m = v[16 + co]; arg = 2; }
                if (v[24 + co] > m) { m = v[24 + co]; arg = 3; }
                am |= arg << (2 * co);
              }
              hi[co] = fmaxf(fmaf(m, inv_s, cint[co]), 0.f);
            }
          } else {
            // border class of the four conv pixels (2py + i, 2px + j) of this pooled pixel
            const int rc0 = py == 0 ? 0 : 1, rc1 = py == H1 - 1 ? 2 : 1, cc0 = px == 0 ? 0 : 1, cc1 = px == W1 - 1 ? 2 : 1;
            float m[8];
#pragma unroll
            for (int sub = 0; sub < 4; ++sub) {
              const int e = ((sub >> 1) ? rc1 : rc0) * 3 + ((sub & 1) ? cc1 : cc0);
              float ws[8];
              *reinterpret_cast<float4*>(ws) = *reinterpret_cast<const float4*>(sh.ws1[e]);
              *reinterpret_cast<float4*>(ws + 4) = *reinterpret_cast<const float4*>(sh.ws1[e] + 4);
#pragma unroll
              for (int co = 0; co < 8; ++co) {
                const float x = fmaf(nmean, ws[co], v[sub * 8 + co]);    // conv(w, x_raw) - mu * (sum of in-frame weights)
                if (sub == 0) m[co] = x;
                else if (!kSave) m[co] = fmaxf(m[co], x);
                else if (x > m[co]) { m[co] = x; am = (am & ~(3u << (2 * co))) | ((uint32_t)sub << (2 * co)); }
              }
            }
            float bb[8];
            *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(sh.b1);
            *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(sh.b1 + 4);
#pragma unroll
            for (int co = 0; co < 8; ++co) hi[co] = fmaxf(fmaf(m[co], inv_s, bb[co]), 0.f);  // 1/sigma > 0 commutes with the max
          }
          if (kSave) {                                                    // fp32 activations + pool argmax for the backward pass
            uint8_t* rec = saved + f * SV_BYTES;
            reinterpret_cast<float4*>(rec + SV_A1)[2 * Pp] = make_float4(hi[0], hi[1], hi[2], hi[3]);
            reinterpret_cast<float4*>(rec + SV_A1)[2 * Pp + 1] = make_float4(hi[4], hi[5], hi[6], hi[7]);
            reinterpret_cast<uint16_t*>(rec + SV_IDX1)[Pp] = (uint16_t)am;
          }
          if (kALo) {
#pragma unroll
            for (int co = 0; co < 8; ++co) tc::split_f16(hi[co], hi[co], lo[co]);           // otherwise pack8 rounds once
          }
          // pooled pixel -> a1 padded parity planes: row parity from the table, column parity (q+1)&1, cell 2g + (q+1)/2
          const uint32_t off = (((gt >> 15) & 1) * 2 + ((q + 1) & 1)) * P1_PLANE + (gt & 0x7fffu) + ((q + 1) >> 1) * 16;
          *reinterpret_cast<uint4*>(a1 + off) = pack8(hi);
          if (kALo) *reinterpret_cast<uint4*>(a1 + A1_BYTES + off) = pack8(lo);
        }
      }
      tc::fence_before_sync();
      tc::fence_async_smem();
      slot_sync<ST>(slot);                              // patches of round r+1 complete; round r's TMEM columns and patch buffer free
      if (warp == 0 && r + 1 < C1_ROUNDS && tc::elect_one()) { tc::fence_after_sync(); conv1_issue(r + 1); }
      if (warp == 1) {
        if (r == 0) {
          // hand-out, split in two so that neither global round trip is waited for: the atomic was issued at the top of
          // the frame, the candidate's clip length is requested here and first looked at a round later
          cand = handout_base + __shfl_sync(0xffffffffu, handout, 0);
          cand_len = (lengths && cand < frames) ? lengths[cand / T] : (int64_t)T;
        }
        if (r == C1_ROUNDS - 2) {
          // the last patches are built: the ring buffer is free (every reader passed the barrier above, after its
          // fence.proxy.async) -- start the copy of this slot's next frame; it travels under the rest of this frame
          fnext = cand;
          if (cand < frames && (int64_t)(cand % T) >= cand_len) {        // a padded frame: zero embedding, look further
            out[cand * out_stride + lane] = 0.f;
            unsigned int h = 0;
            if (lane == 0) h = atomicAdd(counter, 1u);
            fnext = acquire(handout_base + __shfl_sync(0xffffffffu, h, 0));
          }
          start_copy(fnext);
          if (lane == 0) mz.next = fnext;
        }
      }
    }

    // ================= conv2 + ReLU + pool: 3 tiles of 128 virtual pooled pixels (v = py*25 + px)
    // region R0 becomes a2: zero its border cells (conv3's padding; the interior is overwritten by the epilogue below, the
    // rows past 13 only feed output rows that are never read); the patches are dead
    for (int i = tid; i < L::PA * 2 * 76; i += ST) {
      const int pl = i / 76, c = i % 76;               // plane (part, channel half), border cell
      const int cell = c < 26 ? c : c < 52 ? 13 * PITCH2 + (c - 26) : (1 + ((c - 52) >> 1)) * PITCH2 + ((c & 1) ? PITCH2 - 1 : 0);
      reinterpret_cast<uint4*>(r0 + pl * PLANE2)[cell] = make_uint4(0, 0, 0, 0);
    }
    tc::fence_async_smem();
    slot_sync<ST>(slot);
    fnext = mz.next;
    auto conv2_issue = [&](int tl, uint32_t dcol) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          // patch elements (r, 2h) and (r, 2h+1): column-parity planes 0 and 1 of row-parity plane r&1
          const uint32_t a_off = (r & 1) * 2 * P1_PLANE + ((r >> 1) * P1_COLS + h) * 16 + tl * 2048;
          const uint32_t b_off = (r * 4 + 2 * h) * (L::P * W2_CHUNK);
          const uint64_t ah = tc::desc_add(d_a1, a_off);
          const uint64_t bh = tc::desc_add(d_w2, b_off);
          if (kStack) {
            // columns [0,64) = A_hi.B_hi (+ A_lo.B_hi), columns [64,128) = A_hi.B_lo
            tc::mma_f16(tmem + dcol, ah, bh, idesc2s, (r | h) != 0);
            if (kALo) tc::mma_f16(tmem + dcol, tc::desc_add(d_a1, A1_BYTES + a_off), bh, idesc2, true);
          } else {
            tc::mma_f16(tmem + dcol, ah, bh, idesc2, (r | h) != 0);
            if (kSplit) tc::mma_f16(tmem + dcol, ah, tc::desc_add(bh, W2_CHUNK), idesc2, true);      // + A.B_lo, same columns
          }
        }
    };
    auto conv2_epilogue = [&](int tl, uint32_t dcol) {
      if (kStack && !kWide) {
        // 128-thread slot, stacked tile: columns [0,64) = A.B_hi, [64,128) = A.B_lo, 16 channels x 4 sub-positions each;
        // two sub-positions per pass keep the live registers at 64 + 16
        float m4[16];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float a[32], b[32];
          tc::tmem_ld32(tq + dcol + half * 32, a);
          tc::tmem_ld32(tq + dcol + 64 + half * 32, b);
          tc::tmem_ld_wait();
#pragma unroll
          for (int co = 0; co < 16; ++co) {
            const float x = fmaxf(a[co] + b[co], a[16 + co] + b[16 + co]);
            m4[co] = half ? fmaxf(m4[co], x) : x;
          }
        }
        if (a2off[tl] >= 0) {
          float hi[16], bb[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) *reinterpret_cast<float4*>(bb + 4 * i) = *reinterpret_cast<const float4*>(sh.b2 + 4 * i);
#pragma unroll
          for (int co = 0; co < 16; ++co) hi[co] = fmaxf(m4[co] + bb[co], 0.f);
          const uint32_t off = (uint32_t)a2off[tl];
          *reinterpret_cast<uint4*>(r0 + off) = pack8(hi);
          *reinterpret_cast<uint4*>(r0 + PLANE2 + off) = pack8(hi + 8);
        }
      } else if (kStack) {
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
          float bb[8];
          *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(sh.b2 + wg * 8);
          *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(sh.b2 + wg * 8 + 4);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            hi[i] = fmaxf(m4[i] + bb[i], 0.f);
            if (kALo) tc::split_f16(hi[i], hi[i], lo[i]);
          }
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
          if (kALo) *reinterpret_cast<uint4*>(r0 + A2_BYTES + off) = pack8(lo);
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
    if (C2_FLY == 2) {
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
    } else {
      // one stacked tile fills the slot's TMEM columns: the three tiles run one after the other (the other slots' MMAs
      // fill the tensor pipe meanwhile)
#pragma unroll
      for (int tl = 0; tl < 3; ++tl) {
        if (warp == 0 && tc::elect_one()) {
          tc::fence_after_sync();
          conv2_issue(tl, 0);
          tc::mma_commit(&mz.bar[0]);
        }
        tc::mbar_wait_bounded(&mz.bar[0], ph0); ph0 ^= 1;
        tc::fence_after_sync();
        conv2_epilogue(tl, 0);
        if (tl < 2) { tc::fence_before_sync(); slot_sync<ST>(slot); }
      }
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    slot_sync<ST>(slot);

    // ================= conv3 + ReLU + mean: 3 tiles of 16 rows x 8 columns (12 rows real), 9 taps
    // every tile has its own commit (bar[0..2]): the channel sums of tile i are taken while tiles i+1.. are still in the
    // tensor core
    auto conv3_issue = [&](int tx, uint32_t col) {
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int dy = tap / 3, dx = tap % 3;
        const uint32_t a_off = (dy * PITCH2 + tx * 8 + dx) * 16;
        const uint64_t db = tc::desc_add(d_w3, 2 * tap * L::W3_CHUNK);
        tc::mma_f16(tmem + col, tc::desc_add(d_a2, a_off), db, idesc3_hi, tap > 0);
        if (kALo) tc::mma_f16(tmem + col, tc::desc_add(d_a2, A2_BYTES + a_off), db, idesc3_lo, true);
        if (kSplit && !kStack) tc::mma_f16(tmem + col, tc::desc_add(d_a2, a_off), tc::desc_add(db, 32 * 16), idesc3_lo, true);
      }
      tc::mma_commit(&mz.bar[tx]);
    };
    if (warp == 0 && tc::elect_one()) {
      tc::fence_after_sync();
#pragma unroll
      for (int tx = 0; tx < C3_FLY; ++tx) conv3_issue(tx, tx * COLS3);
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
      const uint32_t col3 = (tx % C3_FLY) * COLS3;         // tile 2 re-uses tile 0's columns when only two fit
      if (q < 3 && tx % NWG == wg) {
        float v[32];
        tc::tmem_ld32(tq + col3, v);
        if (kStack) {
          float w[32];
          tc::tmem_ld32(tq + col3 + 32, w);
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
      if (C3_FLY == 2 && tx == 0) {                      // tile 0's columns have been read: tile 2 goes into them
        tc::fence_before_sync();
        slot_sync<ST>(slot);
        if (warp == 0 && tc::elect_one()) { tc::fence_after_sync(); conv3_issue(2, 0); }
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

template <bool kSplit, bool kALo, int kSlots>
constexpr size_t smem_total() {
  using L = Lay<kSplit, kALo, kSlots == 2>;
  return kSlots * L::SLOT_BYTES + L::W1_BYTES + L::P * W2_BYTES + L::W3_BYTES + kSlots * sizeof(SlotMisc) + sizeof(Shared);
}

}  // namespace cnn3
}  // namespace ssb

// K2 v3 (see the header comment).  Same arguments as ssb_roi_cnn_fwd_tc; passes = 1: four 128-thread frame
// slots per CTA; passes = 3 (split, fp32-parity grade): two 256-thread slots (shared memory bound).
// saved (nullable, passes = 3): B*T records of ssb_roi_cnn_saved_bytes() bytes that receive the pooled activations, pool
// argmax maps, conv3 ReLU mask and channel means of every valid frame, for ssb_roi_cnn_bwd.
extern "C" int64_t ssb_roi_cnn_saved_bytes(void) { return ssb::cnn3::SV_BYTES; }
int ssb_roi_cnn_fwd_v3_(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T, const float* const* weights,
                        int32_t standardize, float* out, int64_t out_stride, void* stream);
// passes = 1 runs K2 v3 unless SSB_K2_FP16_V4=1 (A/B switch for tools/bench_stage.py; read once)
static bool ssb_k2_single_pass_v4_() {
  static const bool v4 = [] { const char* e = getenv("SSB_K2_FP16_V4"); return e && e[0] == '1'; }();
  return v4;
}
// passes = 2 runs four 128-thread slots; SSB_K2_TC2_SLOTS=2 picks the two-slot x 256-thread instance (A/B switch, read once)
static int ssb_k2_tc2_slots_() {
  static const int n = [] { const char* e = getenv("SSB_K2_TC2_SLOTS"); return e && e[0] == '4' ? 4 : e && e[0] == '2' ? 2 : 0; }();
  return n;
}
namespace ssb { namespace cnn3 {
template <bool kSplit, bool kALo, int kSlots, int ST, bool kSave, bool kStack>
static int launch_k2(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T, const Weights& gw, int32_t standardize,
                     float* out, int64_t out_stride, unsigned int* counter, void* saved, cudaStream_t st) {
  constexpr size_t smem = smem_total<kSplit, kALo, kSlots>();
  static_assert(smem <= 232448, "shared-memory budget of one CTA");
  static_assert((kSlots == 2) == (ST == 256), "two wide or four narrow slots");
  const int64_t want = (B * T + kSlots - 1) / kSlots;
  const int grid = (int)(want < kSMs ? want : kSMs);
  auto kern = roi_cnn_tc3_kernel<kSplit, kALo, kSlots, ST, kSave, kStack>;
  SSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<grid, ST * kSlots, smem, st>>>(R, lengths, (int)B, (int)T, gw, standardize, out, out_stride, counter,
                                        reinterpret_cast<uint8_t*>(saved));
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
}}  // namespace ssb::cnn3
extern "C" int ssb_roi_cnn_fwd_tc3(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                                   int32_t roi_h, int32_t roi_w, const float* const* weights, int32_t standardize,
                                   float* out, int64_t out_stride, int32_t passes, void* saved, void* stream) {
  using namespace ssb;
  using namespace ssb::cnn3;
  SSB_REQUIRE(R && weights && out, "null pointer");
  SSB_REQUIRE(!saved || (passes == 3 && ((uintptr_t)saved & 15) == 0), "saved activations: split mode only, 16-byte aligned");
  SSB_REQUIRE(roi_h == H0 && roi_w == W0, "ROI must be 48x96 (train_model_official.py:38)");
  SSB_REQUIRE(out_stride >= E, "out_stride < 32");
  SSB_REQUIRE(passes >= 1 && passes <= 3, "passes must be 1, 2 or 3");
  for (int i = 0; i < 8; ++i) SSB_REQUIRE(weights[i], "null weight pointer");
  SSB_REQUIRE(((uintptr_t)R & 15) == 0, "R must be 16-byte aligned (frames are staged by 4,608-byte bulk copies)");
  if (B * T == 0) return SSB_OK;
  Weights gw{weights[0], weights[1], weights[2], weights[3], weights[4], weights[5], weights[6], weights[7]};
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* counter = nullptr;
  if (int rc = ssb_scratch_counter_(st, &counter)) return rc;
  if (passes == 3) {                                   // both operands split: fp32-parity grade, the training mode
    return saved ? launch_k2<true, true, 2, 256, true, true>(R, lengths, B, T, gw, standardize, out, out_stride, counter, saved, st)
                 : launch_k2<true, true, 2, 256, false, true>(R, lengths, B, T, gw, standardize, out, out_stride, counter, nullptr, st);
  }
  if (passes == 2) {                                   // split weights, activations rounded to fp16 once
    // four narrow slots (A/B: SSB_K2_TC2_SLOTS=2 two wide slots, =4 four narrow slots with hi / lo as separate MMAs)
    if (ssb_k2_tc2_slots_() == 2) return launch_k2<true, false, 2, 256, false, true>(R, lengths, B, T, gw, standardize, out, out_stride, counter, nullptr, st);
    if (ssb_k2_tc2_slots_() == 4) return launch_k2<true, false, 4, 128, false, false>(R, lengths, B, T, gw, standardize, out, out_stride, counter, nullptr, st);
    return launch_k2<true, false, 4, 128, false, true>(R, lengths, B, T, gw, standardize, out, out_stride, counter, nullptr, st);
  }
  if (!ssb_k2_single_pass_v4_()) return ssb_roi_cnn_fwd_v3_(R, lengths, B, T, weights, standardize, out, out_stride, stream);   // roi_cnn_v3.cu
  return launch_k2<false, false, 4, 128, false, false>(R, lengths, B, T, gw, standardize, out, out_stride, counter, nullptr, st);
}
