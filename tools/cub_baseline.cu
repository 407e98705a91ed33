// GPU stage baselines built from the LIBRARY calls the reference rasterizer makes (test / bench side only; never
// linked into libpdgs.so, never on the product path):
//   K2  cub::DeviceScan::InclusiveSum over tiles_touched[P]                       
//   K4  cub::DeviceRadixSort::SortPairs on (u64 key = tile << 32 | depth bits, u32 value = Gaussian id)
//       over key bits [0, 32 + bits(T))                                           
//   K3  duplicateWithKeys in the reference's shape (one thread per Gaussian in index order), see cubb_duplicate_with_keys
// The reference's own source for these call sites is absent from /root/reference (empty submodule); the calls are
// the published ones of graphdeco-inria/diff-gaussian-rasterization rasterizer_impl.cu [UPSTREAM-RECALL].  CUB is the
// one shipped with CUDA 12.9 in this image.  Plain C ABI so that tests/ and bench.py can drive it through ctypes on
// the very arrays the new pipeline produced.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

extern "C" {

size_t cubb_scan_temp_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceScan::InclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n);
  return bytes;
}

int cubb_inclusive_sum_u32(const uint32_t* in, uint32_t* out, int64_t n, void* temp, size_t temp_bytes, void* stream) {
  return (int)cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, (int)n, (cudaStream_t)stream);
}

size_t cubb_sort_u64_temp_bytes(int64_t n, int end_bit) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (int)n, 0, end_bit);
  return bytes;
}

// the reference's K4: keys_unsorted/values_unsorted -> keys/values, bits [0, end_bit)
int cubb_sort_pairs_u64(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out, int64_t n,
                        int end_bit, void* temp, size_t temp_bytes, void* stream) {
  return (int)cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, end_bit,
                                              (cudaStream_t)stream);
}

size_t cubb_sort_u32_temp_bytes(int64_t n, int end_bit) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (int)n, 0, end_bit);
  return bytes;
}

// 32-bit keys: what CUB needs for the two sorts of the new pipeline's decomposition (depth bits over P, tile id over R)
int cubb_sort_pairs_u32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out, int64_t n,
                        int end_bit, void* temp, size_t temp_bytes, void* stream) {
  return (int)cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, end_bit,
                                              (cudaStream_t)stream);
}

// The reference's K3 in the reference's shape (graphdeco-inria duplicateWithKeys + getRect
// [UPSTREAM-RECALL]; restated, not copied: the source is absent): one thread per Gaussian in INDEX order; a visible one
// (radii > 0) recomputes its tile rectangle from its pixel centre and radius, starts at the exclusive prefix of
// tiles_touched and writes, y outer / x inner, key = (tile << 32) | depth bits, value = its index.  It reads only K1
// outputs (xy, radii, depth bits) and the CUB scan of tiles_touched -- nothing the new pipeline's binning produced --
// so sorting its pairs with cub::DeviceRadixSort gives a point list that is independent of the code under test.
// error[0] counts Gaussians whose recomputed rectangle area differs from tiles_touched (must stay 0).
__global__ void cubb_duplicate_kernel(const float4* __restrict__ xy, const int32_t* __restrict__ radii,
                                      const uint32_t* __restrict__ depth_bits, const uint32_t* __restrict__ offsets,
                                      const uint32_t* __restrict__ tiles_touched, int64_t P, int grid_x, int grid_y,
                                      uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* error) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const int r = radii[idx];
  if (r <= 0) return;
  const float4 p = xy[idx];
  const int x0 = min(grid_x, max(0, (int)((p.x - r) / 16)));
  const int y0 = min(grid_y, max(0, (int)((p.y - r) / 16)));
  // upstream writes `p.x + max_radius + BLOCK_X - 1`: float additions left to right ((p.x + r) + 16) - 1, two roundings --
  // not the single `+ 15` of SURVEY A.4's shorthand, which differs in the last bit for a few Gaussians per million
  const int x1 = min(grid_x, max(0, (int)((p.x + r + 16 - 1) / 16)));
  const int y1 = min(grid_y, max(0, (int)((p.y + r + 16 - 1) / 16)));
  if ((uint32_t)((x1 - x0) * (y1 - y0)) != tiles_touched[idx]) { atomicAdd(error, 1u); return; }
  uint32_t off = idx == 0 ? 0u : offsets[idx - 1];
  const uint64_t d = depth_bits[idx];
  for (int y = y0; y < y1; ++y)
    for (int x = x0; x < x1; ++x) {
      keys[off] = ((uint64_t)(uint32_t)(y * grid_x + x) << 32) | d;
      vals[off] = (uint32_t)idx;
      ++off;
    }
}

int cubb_duplicate_with_keys(const float* xy4, const int32_t* radii, const uint32_t* depth_bits, const uint32_t* offsets,
                             const uint32_t* tiles_touched, int64_t P, int grid_x, int grid_y, uint64_t* keys, uint32_t* vals,
                             uint32_t* error, void* stream) {
  if (P <= 0) return 0;
  cubb_duplicate_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const float4*)xy4, radii, depth_bits, offsets, tiles_touched, P, grid_x, grid_y, keys, vals, error);
  return (int)cudaGetLastError();
}

}  // extern "C"
