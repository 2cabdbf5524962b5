// Device-resident clip store: batch assembly by index (".npz ingestion").
//
// The reference assembles a batch on the host, one np.load + pad/trim per item and a torch.stack per batch
// (NPZWordDataset.__getitem__ train_model_official.py:134-172, collate_fn 174-204), which makes a B=1024 step
// input-bound.  Here the padded clips are ingested ONCE into HBM (a few hundred KB per clip: 10,000 clips = 4.8 GB of
// the 180 GB) and a batch is a gather of whole rows by the sampler's indices -- pure data movement at HBM speed.
#include "common.cuh"

namespace ssb {
namespace ds {

constexpr int THREADS = 256;
constexpr int SEG = THREADS * 16 * 4;      // bytes one CTA moves per segment: 4 x 16 B per thread, all loads first

// rows of row_bytes (multiple of 16, 16-byte aligned): grid (n, segments)
__global__ void __launch_bounds__(THREADS)
gather_rows16_kernel(const uint4* __restrict__ src, int64_t row_vec, const int64_t* __restrict__ index, uint4* __restrict__ dst) {
  const int64_t r = index[blockIdx.x];
  const uint4* s = src + r * row_vec;
  uint4* d = dst + (int64_t)blockIdx.x * row_vec;
  for (int64_t v0 = (int64_t)blockIdx.y * (SEG / 16); v0 < row_vec; v0 += (int64_t)gridDim.y * (SEG / 16)) {
    uint4 t[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t v = v0 + k * THREADS + threadIdx.x;
      if (v < row_vec) t[k] = __ldg(s + v);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t v = v0 + k * THREADS + threadIdx.x;
      if (v < row_vec) d[v] = t[k];
    }
  }
}

// small or odd-sized rows (lengths, labels): one thread per byte-unit of `unit` bytes
template <typename U>
__global__ void gather_rows_small_kernel(const U* __restrict__ src, int64_t row_units, const int64_t* __restrict__ index,
                                         int64_t n, U* __restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * row_units) return;
  const int64_t row = i / row_units, c = i - row * row_units;
  dst[i] = src[index[row] * row_units + c];
}

}  // namespace ds
}  // namespace ssb

// dst[i, :] = src[index[i], :] for i < n; rows are row_bytes long and contiguous.  index values must lie in
// [0, n_src) -- the host-side sampler guarantees it (silent_speech_b200/dataset.py validates before upload).
extern "C" int ssb_gather_rows(const void* src, int64_t row_bytes, const int64_t* index, int64_t n, void* dst, void* stream) {
  using namespace ssb;
  SSB_REQUIRE(row_bytes > 0 && n >= 0, "bad sizes");
  if (n == 0) return SSB_OK;
  SSB_REQUIRE(src && index && dst, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const uintptr_t al = (uintptr_t)src | (uintptr_t)dst | (uintptr_t)row_bytes;
  if ((al & 15) == 0 && row_bytes >= 4096) {
    const int64_t row_vec = row_bytes / 16;
    int64_t segs = (row_bytes + ds::SEG - 1) / ds::SEG;
    if (segs > 64) segs = 64;
    SSB_REQUIRE(n < ((int64_t)1 << 31), "too many rows");
    ds::gather_rows16_kernel<<<dim3((unsigned)n, (unsigned)segs), ds::THREADS, 0, st>>>((const uint4*)src, row_vec, index, (uint4*)dst);
  } else if ((al & 7) == 0) {
    const int64_t units = row_bytes / 8, total = n * units;
    ds::gather_rows_small_kernel<uint64_t><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const uint64_t*)src, units, index, n, (uint64_t*)dst);
  } else if ((al & 3) == 0) {
    const int64_t units = row_bytes / 4, total = n * units;
    ds::gather_rows_small_kernel<uint32_t><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const uint32_t*)src, units, index, n, (uint32_t*)dst);
  } else {
    const int64_t total = n * row_bytes;
    ds::gather_rows_small_kernel<uint8_t><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const uint8_t*)src, row_bytes, index, n, (uint8_t*)dst);
  }
  SSB_LAUNCH_CHECK();
  return SSB_OK;
}
