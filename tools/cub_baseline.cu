// GPU stage baselines built from the LIBRARY calls the reference rasterizer makes (test / bench side only; never
// linked into libpdgs.so, never on the product path):
//   K2  cub::DeviceScan::InclusiveSum over tiles_touched[P]                       
//   K4  cub::DeviceRadixSort::SortPairs on (u64 key = tile << 32 | depth bits, u32 value = Gaussian id)
//       over key bits [0, 32 + bits(T))                                           
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

}  // extern "C"
