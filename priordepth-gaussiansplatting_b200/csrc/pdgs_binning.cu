// K2 scan, K3 duplicateWithKeys, K4 onesweep radix sort, K5 identifyTileRanges.
// Replaces cub::DeviceScan::InclusiveSum, duplicateWithKeys, cub::DeviceRadixSort::SortPairs and
// identifyTileRanges of the reference rasterizer (source absent).
//
// All kernels read the pair count R from device memory (Ctrl::num_rendered) and are guarded by
// the binning capacity, so the host can enqueue them before it knows R (no mid-forward bubble).
#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

// ------------------------------------------------------------------------------------------
// K2: single-pass inclusive scan of tiles_touched with decoupled look-back.
// state word: [63:62] flag (0 empty, 1 aggregate, 2 inclusive prefix), [61:0] value.
// ------------------------------------------------------------------------------------------
constexpr unsigned long long FLAG_AGG = 1ull << 62;
constexpr unsigned long long FLAG_INC = 2ull << 62;
constexpr unsigned long long VAL_MASK = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
  return *reinterpret_cast<const volatile unsigned long long*>(p);
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
  *reinterpret_cast<volatile unsigned long long*>(p) = v;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
  return *reinterpret_cast<const volatile uint32_t*>(p);
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }

__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(GeomView g, int64_t P) {
  __shared__ uint32_t s_tile;
  __shared__ unsigned long long s_warp[SCAN_THREADS / 32];
  __shared__ unsigned long long s_excl;
  if (threadIdx.x == 0) s_tile = atomicAdd(&g.ctrl->scan_ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  if (base + SCAN_ITEMS <= P) {
    const uint4* src = reinterpret_cast<const uint4*>(g.tiles_touched + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k) {
      const uint4 q = src[k];
      v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < P) ? g.tiles_touched[base + k] : 0u;
  }
  unsigned long long tsum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) tsum += v[k];
  // block exclusive scan of thread sums
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  unsigned long long woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    const unsigned long long c = s_warp[w];
    if (w < warp) woff += c;
    total += c;
  }
  if (threadIdx.x == 0) {
    unsigned long long excl = 0;
    if (tile == 0) {
      st_volatile_u64(&g.scan_state[0], FLAG_INC | total);
    } else {
      st_volatile_u64(&g.scan_state[tile], FLAG_AGG | total);
      int64_t t = (int64_t)tile - 1;
      while (true) {
        unsigned long long s;
        do { s = ld_volatile_u64(&g.scan_state[t]); } while ((s >> 62) == 0ull);
        excl += s & VAL_MASK;
        if ((s >> 62) == 2ull) break;
        --t;
      }
      st_volatile_u64(&g.scan_state[tile], FLAG_INC | (excl + total));
    }
    s_excl = excl;
    const int64_t ntiles = (P + SCAN_TILE - 1) / SCAN_TILE;
    if ((int64_t)tile == ntiles - 1) {
      const unsigned long long R = excl + total;
      g.ctrl->num_rendered = (uint32_t)(R > 0xffffffffull ? 0xffffffffull : R);
      g.ctrl->overflow = (R >= (unsigned long long)MAX_PAIRS) ? 1u : 0u;
    }
  }
  __syncthreads();
  unsigned long long run = s_excl + woff + (inc - tsum);
  uint32_t o[SCAN_ITEMS];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    run += v[k];
    o[k] = (uint32_t)run;
  }
  if (base + SCAN_ITEMS <= P) {
    uint4* dst = reinterpret_cast<uint4*>(g.point_offsets + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k) dst[k] = make_uint4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
      if (base + k < P) g.point_offsets[base + k] = o[k];
  }
}

// ------------------------------------------------------------------------------------------
// K3: emit (tile << 32 | depth bits, id) for every tile of every visible rect (y outer, x inner).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) duplicate_kernel(GeomView g, const int32_t* __restrict__ radii, int64_t P,
                                                        int grid_x, int grid_y, uint64_t* __restrict__ keys,
                                                        uint32_t* __restrict__ vals, int64_t capacity) {
  const uint32_t R = ld_volatile_u32(&g.ctrl->num_rendered);
  if ((int64_t)R > capacity || g.ctrl->overflow) return;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const int radius = radii[idx];
  if (radius <= 0) return;
  const float4 a = g.gA[idx];
  const float depth = g.gB[idx].w;
  int rx0, ry0, rx1, ry1;
  tile_rect(a.x, a.y, radius, grid_x, grid_y, rx0, ry0, rx1, ry1);
  uint32_t off = (idx == 0) ? 0u : g.point_offsets[idx - 1];
  const uint64_t dbits = (uint64_t)__float_as_uint(depth);
  for (int y = ry0; y < ry1; ++y) {
    for (int x = rx0; x < rx1; ++x) {
      keys[off] = ((uint64_t)(uint32_t)(y * grid_x + x) << 32) | dbits;
      vals[off] = (uint32_t)idx;
      ++off;
    }
  }
}

// ------------------------------------------------------------------------------------------
// K4: onesweep LSD radix sort, 8-bit digits.  One histogram kernel for all passes, then one
// chained-scan (decoupled look-back) kernel per pass.  look-back word: [31:30] flag, [29:0] count.
// ------------------------------------------------------------------------------------------
constexpr uint32_t LB_AGG = 1u << 30;
constexpr uint32_t LB_INC = 2u << 30;
constexpr uint32_t LB_MASK = (1u << 30) - 1;

__global__ void __launch_bounds__(256) sort_histogram_kernel(const uint64_t* __restrict__ keys, const uint32_t* n_dev,
                                                             int64_t capacity, int npass, int end_bit, uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_hist[MAX_PASSES * RADIX];
  const int64_t n = (int64_t)ld_volatile_u32(n_dev);
  if (n > capacity) return;
  for (int i = threadIdx.x; i < npass * RADIX; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t k = keys[i];
    for (int p = 0; p < npass; ++p) {
      const int nb = min(RADIX_BITS, end_bit - p * RADIX_BITS);  // the last pass may cover fewer than 8 bits
      atomicAdd(&s_hist[p * RADIX + (int)((k >> (p * RADIX_BITS)) & ((1u << nb) - 1u))], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < npass * RADIX; i += blockDim.x) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

__global__ void __launch_bounds__(SORT_THREADS) sort_onesweep_kernel(const uint64_t* __restrict__ keys_in,
                                                                    const uint32_t* __restrict__ vals_in,
                                                                    uint64_t* __restrict__ keys_out,
                                                                    uint32_t* __restrict__ vals_out, const uint32_t* n_dev,
                                                                    int64_t capacity, int pass, uint32_t digit_mask,
                                                                    const uint32_t* __restrict__ hist, uint32_t* ticket,
                                                                    uint32_t* lookback) {
  __shared__ uint64_t s_keys[SORT_TILE];
  __shared__ uint32_t s_vals[SORT_TILE];
  __shared__ uint32_t s_wcnt[SORT_THREADS / 32][RADIX];
  __shared__ uint32_t s_cta_base[RADIX];    // exclusive prefix of CTA digit counts (smem slot of the digit's first key)
  __shared__ uint32_t s_goff[RADIX];        // global offset of the digit's first key minus its smem slot
  __shared__ uint32_t s_scan[SORT_THREADS / 32];
  __shared__ uint32_t s_tile;

  const int64_t n = (int64_t)ld_volatile_u32(n_dev);
  if (n > capacity) return;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  for (int i = threadIdx.x; i < (SORT_THREADS / 32) * RADIX; i += SORT_THREADS) (&s_wcnt[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const int64_t tile_base = (int64_t)tile * SORT_TILE;
  if (tile_base >= n) return;
  const int valid = (int)((n - tile_base < SORT_TILE) ? (n - tile_base) : SORT_TILE);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int shift = pass * RADIX_BITS;

  // warp-striped load: warp w owns SORT_ITEMS*32 consecutive pairs, item i of lane l is seg + i*32 + l
  uint64_t key[SORT_ITEMS];
  uint32_t val[SORT_ITEMS];
  uint32_t rank[SORT_ITEMS];
  const int seg = warp * SORT_ITEMS * 32;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int local = seg + i * 32 + lane;
    if (local < valid) {
      key[i] = keys_in[tile_base + local];
      val[i] = vals_in[tile_base + local];
    } else {
      key[i] = ~0ull;  // padding: forced into the last bin below, excluded from the counts
      val[i] = 0u;
    }
  }
  // per-warp stable ranking with match_any
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const uint32_t d = (seg + i * 32 + lane < valid) ? ((uint32_t)(key[i] >> shift) & digit_mask) : (uint32_t)(RADIX - 1);
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    const uint32_t below = __popc(peers & ((1u << lane) - 1u));
    uint32_t pre = 0;
    if (lane == leader) {
      pre = s_wcnt[warp][d];
      s_wcnt[warp][d] = pre + __popc(peers);
    }
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank[i] = pre + below;
    __syncwarp();
  }
  __syncthreads();
  // thread d owns digit d: exclusive scan over warps, CTA count
  const int d = threadIdx.x;
  uint32_t cnt = 0;
#pragma unroll
  for (int w = 0; w < SORT_THREADS / 32; ++w) {
    const uint32_t c = s_wcnt[w][d];
    s_wcnt[w][d] = cnt;
    cnt += c;
  }
  const uint32_t n_invalid = (uint32_t)(SORT_TILE - valid);
  const uint32_t cnt_valid = cnt - ((d == RADIX - 1) ? n_invalid : 0u);
  // publish the aggregate as early as possible
  uint32_t* my_lb = lookback + (size_t)tile * RADIX + d;
  if (tile == 0) st_volatile_u32(my_lb, LB_INC | cnt_valid);
  else st_volatile_u32(my_lb, LB_AGG | cnt_valid);
  // global base of digit d = exclusive prefix over digits of the global histogram
  const uint32_t hcount = hist[pass * RADIX + d];
  uint32_t inc_c = cnt, inc_h = hcount;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, inc_c, o);
    const uint32_t b = __shfl_up_sync(0xffffffffu, inc_h, o);
    if (lane >= o) { inc_c += a; inc_h += b; }
  }
  __shared__ uint32_t s_scan_h[SORT_THREADS / 32];
  if (lane == 31) { s_scan[warp] = inc_c; s_scan_h[warp] = inc_h; }
  __syncthreads();
  uint32_t woff_c = 0, woff_h = 0;
#pragma unroll
  for (int w = 0; w < SORT_THREADS / 32; ++w) {
    if (w < warp) { woff_c += s_scan[w]; woff_h += s_scan_h[w]; }
  }
  const uint32_t cta_base = woff_c + inc_c - cnt;       // exclusive over digits (smem slot)
  const uint32_t gbase = woff_h + inc_h - hcount;       // exclusive over digits (global)
  // decoupled look-back for digit d
  uint32_t excl = 0;
  if (tile > 0) {
    int64_t t = (int64_t)tile - 1;
    while (true) {
      uint32_t s;
      do { s = ld_volatile_u32(lookback + (size_t)t * RADIX + d); } while ((s >> 30) == 0u);
      excl += s & LB_MASK;
      if ((s >> 30) == 2u) break;
      --t;
    }
    st_volatile_u32(my_lb, LB_INC | ((excl + cnt_valid) & LB_MASK));
  }
  s_cta_base[d] = cta_base;
  s_goff[d] = gbase + excl - cta_base;
  __syncthreads();
  // scatter into shared memory in sorted order
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const uint32_t dd = (seg + i * 32 + lane < valid) ? ((uint32_t)(key[i] >> shift) & digit_mask) : (uint32_t)(RADIX - 1);
    const uint32_t slot = s_cta_base[dd] + s_wcnt[warp][dd] + rank[i];
    s_keys[slot] = key[i];
    s_vals[slot] = val[i];
  }
  __syncthreads();
  // coalesced write-out; padding keys occupy slots >= valid and are dropped
  for (int j = threadIdx.x; j < valid; j += SORT_THREADS) {
    const uint64_t k = s_keys[j];
    const uint32_t dd = (uint32_t)(k >> shift) & digit_mask;
    const uint32_t dst = s_goff[dd] + (uint32_t)j;
    keys_out[dst] = k;
    vals_out[dst] = s_vals[j];
  }
}

// ------------------------------------------------------------------------------------------
// K5: per-tile [first, last) from the sorted keys.  ranges is zeroed beforehand.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ranges_kernel(const uint64_t* __restrict__ keys, const uint32_t* n_dev,
                                                     int64_t capacity, uint2* __restrict__ ranges) {
  const int64_t n = (int64_t)ld_volatile_u32(n_dev);
  if (n > capacity) return;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[t].x = 0;
    else {
      const uint32_t tp = (uint32_t)(keys[i - 1] >> 32);
      if (t != tp) {
        ranges[tp].y = (uint32_t)i;
        ranges[t].x = (uint32_t)i;
      }
    }
    if (i == n - 1) ranges[t].y = (uint32_t)n;
  }
}

}  // namespace

cudaError_t launch_scan(GeomView g, int64_t P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((P + SCAN_TILE - 1) / SCAN_TILE);
  scan_kernel<<<blocks, SCAN_THREADS, 0, s>>>(g, P);
  return cudaGetLastError();
}

cudaError_t launch_duplicate(const DeviceArgs& a, GeomView g, const int32_t* radii, uint64_t* keys, uint32_t* vals,
                             int64_t capacity, cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((a.P + 255) / 256);
  duplicate_kernel<<<blocks, 256, 0, s>>>(g, radii, a.P, a.grid_x, a.grid_y, keys, vals, capacity);
  return cudaGetLastError();
}

cudaError_t launch_sort(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, const uint32_t* n_dev,
                        int64_t capacity, int npass, int end_bit, void* scratch, cudaStream_t s) {
  if (capacity <= 0 || npass <= 0) return cudaSuccess;
  if (npass > MAX_PASSES) return cudaErrorInvalidValue;
  SortScratch sc = carve_sort_scratch(scratch, capacity);
  cudaError_t e = cudaMemsetAsync(scratch, 0, sc.zero_bytes, s);
  if (e != cudaSuccess) return e;
  uint64_t* kin = (npass % 2 == 0) ? keys_a : keys_b;
  uint32_t* vin = (npass % 2 == 0) ? vals_a : vals_b;
  uint64_t* kout = (npass % 2 == 0) ? keys_b : keys_a;
  uint32_t* vout = (npass % 2 == 0) ? vals_b : vals_a;
  int hist_blocks = (int)((capacity + 256 * 16 - 1) / (256 * 16));
  if (hist_blocks > 148 * 8) hist_blocks = 148 * 8;
  if (hist_blocks < 1) hist_blocks = 1;
  sort_histogram_kernel<<<hist_blocks, 256, 0, s>>>(kin, n_dev, capacity, npass, end_bit, sc.hist);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const unsigned tiles = (unsigned)sc.ntiles;
  for (int p = 0; p < npass; ++p) {
    const int nb = (end_bit - p * RADIX_BITS < RADIX_BITS) ? end_bit - p * RADIX_BITS : RADIX_BITS;
    sort_onesweep_kernel<<<tiles, SORT_THREADS, 0, s>>>(kin, vin, kout, vout, n_dev, capacity, p, (1u << nb) - 1u, sc.hist,
                                                         sc.tickets + p, sc.lookback + (size_t)p * sc.ntiles * RADIX);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    uint64_t* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  return cudaSuccess;
}

cudaError_t launch_ranges(const uint64_t* keys, const uint32_t* n_dev, int64_t capacity, uint2* ranges, int tiles,
                          cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), s);
  if (e != cudaSuccess) return e;
  if (capacity <= 0) return cudaSuccess;
  int blocks = (int)((capacity + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  ranges_kernel<<<blocks, 256, 0, s>>>(keys, n_dev, capacity, ranges);
  return cudaGetLastError();
}

}  // namespace pdgs
