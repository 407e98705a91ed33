// K2 scan, K3 duplicateWithKeys, K4 onesweep radix sorts, K5 identifyTileRanges.
// Replaces cub::DeviceScan::InclusiveSum, duplicateWithKeys, cub::DeviceRadixSort::SortPairs and
// identifyTileRanges of the reference rasterizer (source absent) with a
// pipeline that produces the identical sorted point list and tile ranges from less data movement.
//
// All kernels read the pair count R from device memory (Ctrl::num_rendered) and are guarded by
// the binning capacity, so the host can enqueue them before it knows R (no mid-forward bubble).
#include <cstdlib>

#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

// ------------------------------------------------------------------------------------------
// Binning pipeline ("depth first, then tile").  The reference sorts R 64-bit (tile | depth) keys
// with R 32-bit values in one radix sort.  The same sorted list comes out of two much smaller
// sorts, because an LSD radix sort is stable:
//   1. sort the P Gaussians by their 32 depth bits (4 passes over 8-byte pairs; culled Gaussians
//      carry key 0xffffffff and end up last), ties keep ascending index;
//   2. scan tiles_touched in that order and emit the (tile, index) pairs in that order;
//   3. stable-sort the R pairs by tile id only (bits(T) bits: 2 passes over 8-byte pairs).
// Pairs of one tile are then ordered by (depth bits, index): exactly the order a stable sort of the
// reference's keys over an index-ordered duplication produces.
// ------------------------------------------------------------------------------------------
constexpr unsigned long long FLAG_AGG = 1ull << 62;
constexpr unsigned long long FLAG_INC = 2ull << 62;
constexpr unsigned long long VAL_MASK = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// K2: single-pass inclusive scan (decoupled look-back) of tiles_touched taken in depth order:
// point_offsets[i] = sum_{k <= i} tiles_touched[order[k]].
// state word: [63:62] flag (0 empty, 1 aggregate, 2 inclusive prefix), [61:0] value.
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
    const uint4* src = reinterpret_cast<const uint4*>(g.order + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k) {
      const uint4 q = src[k];
      v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = __ldg(g.tiles_touched + v[k]);
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < P) ? __ldg(g.tiles_touched + g.order[base + k]) : 0u;
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
  if (warp == 0) {
    // warp-parallel decoupled look-back: lane l inspects tile (t - l); stop at the nearest inclusive prefix
    unsigned long long excl = 0;
    if (lane == 0) st_relaxed_u64(&g.scan_state[tile], (tile == 0 ? FLAG_INC : FLAG_AGG) | total);
    int64_t t = (int64_t)tile - 1;
    while (t >= 0) {
      const int64_t ti = t - lane;
      unsigned long long sv = FLAG_INC;  // tiles before the first one: inclusive prefix 0
      if (ti >= 0) {
        do { sv = ld_relaxed_u64(&g.scan_state[ti]); } while ((sv >> 62) == 0ull);
      }
      const unsigned incmask = __ballot_sync(0xffffffffu, (sv >> 62) == 2ull);
      const int first = incmask ? (__ffs(incmask) - 1) : 31;
      unsigned long long val = (lane <= first) ? (sv & VAL_MASK) : 0ull;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
      excl += val;
      if (incmask) break;
      t -= 32;
    }
    if (lane == 0) {
      if (tile != 0) st_relaxed_u64(&g.scan_state[tile], FLAG_INC | (excl + total));
      s_excl = excl;
      const int64_t ntiles = (P + SCAN_TILE - 1) / SCAN_TILE;
      if ((int64_t)tile == ntiles - 1) {
        const unsigned long long R = excl + total;
        g.ctrl->num_rendered = (uint32_t)(R > 0xffffffffull ? 0xffffffffull : R);
        g.ctrl->overflow = (R >= (unsigned long long)MAX_PAIRS) ? 1u : 0u;
      }
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
// K3: emit (tile, index) for every tile of every visible rect, Gaussians taken in depth order.
// Output-centric: a warp owns DUP_SEG consecutive sorted Gaussians, whose pairs form ONE contiguous
// output range; lane l writes output positions l, l+32, ... of that range and finds the owning Gaussian
// with a binary search over the segment's offsets in shared memory.  Every store is fully coalesced and
// a rect of thousands of tiles costs the same per pair as a rect of one.
// ------------------------------------------------------------------------------------------
constexpr int DUP_SEG = PDGS_DUP_SEG;  // sorted Gaussians per warp
constexpr int DUP_SEARCH = DUP_SEG <= 64 ? 6 : (DUP_SEG <= 128 ? 7 : 8);  // binary-search steps over a segment's offsets
static_assert(DUP_SEG == 64 || DUP_SEG == 128 || DUP_SEG == 256, "segment size");
constexpr int DUP_WARPS = 8;

__global__ void __launch_bounds__(DUP_WARPS * 32) duplicate_kernel(GeomView g, int64_t P, int grid_x,
                                                                   uint32_t* __restrict__ tkeys,
                                                                   uint32_t* __restrict__ vals, int64_t capacity) {
  __shared__ uint32_t s_off[DUP_WARPS][DUP_SEG + 1];
  __shared__ uint2 s_rect[DUP_WARPS][DUP_SEG];
  __shared__ uint32_t s_idx[DUP_WARPS][DUP_SEG];
  const uint32_t R = ld_relaxed_u32(&g.ctrl->num_rendered);
  if ((int64_t)R > capacity || g.ctrl->overflow) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t seg0 = ((int64_t)blockIdx.x * DUP_WARPS + warp) * DUP_SEG;
  if (seg0 >= P) return;
  const int n = (int)((P - seg0 < DUP_SEG) ? (P - seg0) : DUP_SEG);
  uint32_t* off = s_off[warp];
  if (lane == 0) off[0] = (seg0 == 0) ? 0u : g.point_offsets[seg0 - 1];
#pragma unroll
  for (int k = 0; k < DUP_SEG / 32; ++k) {
    const int j = k * 32 + lane;
    if (j < n) {
      const uint32_t idx = g.order[seg0 + j];
      off[j + 1] = g.point_offsets[seg0 + j];
      s_idx[warp][j] = idx;
      s_rect[warp][j] = __ldg(g.rect + idx);
    }
  }
  __syncwarp();
  const uint32_t first = off[0], total = off[n] - first;
  for (uint32_t e = lane; e < total; e += 32) {
    const uint32_t pos = first + e;
    // largest j in [0, n) with off[j] <= pos  (off is non-decreasing; empty Gaussians have off[j] == off[j+1])
    int lo = 0, hi = n;  // invariant: off[lo] <= pos < off[hi]
#pragma unroll
    for (int it = 0; it < DUP_SEARCH; ++it) {
      const int mid = (lo + hi) >> 1;
      if (hi - lo > 1) {
        if (off[mid] <= pos) lo = mid; else hi = mid;
      }
    }
    const uint2 rc = s_rect[warp][lo];
    const uint32_t rx0 = rc.x & 0xffffu, ry0 = rc.x >> 16, w = (rc.y & 0xffffu) - rx0;
    const uint32_t local = pos - off[lo];
    const uint32_t dy = local / w, dx = local - dy * w;
    tkeys[pos] = (ry0 + dy) * (uint32_t)grid_x + rx0 + dx;
    vals[pos] = s_idx[warp][lo];
  }
}

// K2 + K3 in one kernel (the steady-state path: a binning buffer is at hand).  A CTA takes DUP_TILE consecutive sorted
// Gaussians, eight warps of DUP_SEG each.  The pair count of a Gaussian is the area of its tile rect (culled
// Gaussians carry an empty rect), which the emit phase needs anyway, so tiles_touched is not gathered at all and the
// offsets never travel through HBM between two kernels: warp scan -> CTA scan -> decoupled look-back over the CTAs
// (tickets give the order, as in scan_kernel) -> emit.  point_offsets are still written (state view, resume path).
// Pairs that would land at or beyond `capacity` are not written; the host sees R > capacity and re-runs with a larger
// buffer through the resume path (duplicate_kernel above).
static_assert(DUP_TILE == DUP_WARPS * DUP_SEG, "fused scan + duplicate tile");

__global__ void __launch_bounds__(DUP_WARPS * 32) scan_duplicate_kernel(GeomView g, int64_t P, int grid_x,
                                                                        uint32_t* __restrict__ tkeys,
                                                                        uint32_t* __restrict__ vals, int64_t capacity) {
  __shared__ uint32_t s_off[DUP_WARPS][DUP_SEG + 1];
  __shared__ uint2 s_rect[DUP_WARPS][DUP_SEG];
  __shared__ uint32_t s_idx[DUP_WARPS][DUP_SEG];
  __shared__ uint32_t s_wsum[DUP_WARPS];
  __shared__ uint32_t s_tile;
  __shared__ unsigned long long s_excl;
  if (threadIdx.x == 0) s_tile = atomicAdd(&g.ctrl->scan_ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t seg0 = ((int64_t)tile * DUP_WARPS + warp) * DUP_SEG;
  const int n = (seg0 >= P) ? 0 : (int)((P - seg0 < DUP_SEG) ? (P - seg0) : DUP_SEG);
  uint32_t* off = s_off[warp];
  constexpr int K = DUP_SEG / 32;
  uint32_t idx[K];
  uint2 rc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) idx[k] = (k * 32 + lane < n) ? g.order[seg0 + k * 32 + lane] : 0u;
#pragma unroll
  for (int k = 0; k < K; ++k) rc[k] = (k * 32 + lane < n) ? __ldg(g.rect + idx[k]) : make_uint2(0u, 0u);
  uint32_t run = 0;  // pairs of the warp's Gaussians before block k
  if (lane == 0) off[0] = 0u;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int j = k * 32 + lane;
    uint32_t inc = ((rc[k].y & 0xffffu) - (rc[k].x & 0xffffu)) * ((rc[k].y >> 16) - (rc[k].x >> 16));
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += up;
    }
    off[j + 1] = run + inc;  // inclusive, relative to the warp's first pair
    s_idx[warp][j] = idx[k];
    s_rect[warp][j] = rc[k];
    run += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) s_wsum[warp] = run;
  __syncthreads();
  unsigned long long woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < DUP_WARPS; ++w) {
    const unsigned long long c = s_wsum[w];
    if (w < warp) woff += c;
    total += c;
  }
  if (warp == 0) {
    // warp-parallel decoupled look-back, as in scan_kernel
    unsigned long long excl = 0;
    if (lane == 0) st_relaxed_u64(&g.scan_state[tile], (tile == 0 ? FLAG_INC : FLAG_AGG) | total);
    int64_t t = (int64_t)tile - 1;
    while (t >= 0) {
      const int64_t ti = t - lane;
      unsigned long long sv = FLAG_INC;
      if (ti >= 0) {
        do { sv = ld_relaxed_u64(&g.scan_state[ti]); } while ((sv >> 62) == 0ull);
      }
      const unsigned incmask = __ballot_sync(0xffffffffu, (sv >> 62) == 2ull);
      const int first = incmask ? (__ffs(incmask) - 1) : 31;
      unsigned long long val = (lane <= first) ? (sv & VAL_MASK) : 0ull;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
      excl += val;
      if (incmask) break;
      t -= 32;
    }
    if (lane == 0) {
      if (tile != 0) st_relaxed_u64(&g.scan_state[tile], FLAG_INC | (excl + total));
      s_excl = excl;
      const int64_t ntiles = (P + DUP_TILE - 1) / DUP_TILE;
      if ((int64_t)tile == ntiles - 1) {
        const unsigned long long R = excl + total;
        g.ctrl->num_rendered = (uint32_t)(R > 0xffffffffull ? 0xffffffffull : R);
        g.ctrl->overflow = (R >= (unsigned long long)MAX_PAIRS) ? 1u : 0u;
      }
    }
  }
  __syncthreads();
  if (n == 0) return;
  const unsigned long long base = s_excl + woff;  // pairs before this warp's first
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int j = k * 32 + lane;
    if (j < n) g.point_offsets[seg0 + j] = (uint32_t)(base + off[j + 1]);
  }
  if (base >= (unsigned long long)capacity) return;
  uint32_t total_w = off[n];
  if (base + total_w > (unsigned long long)capacity) total_w = (uint32_t)((unsigned long long)capacity - base);
  const uint32_t first = (uint32_t)base;  // < capacity <= 2^30
  for (uint32_t e = lane; e < total_w; e += 32) {
    // largest j in [0, n) with off[j] <= e  (off is non-decreasing; empty Gaussians have off[j] == off[j+1])
    int lo = 0, hi = n;  // invariant: off[lo] <= e < off[hi]
#pragma unroll
    for (int it = 0; it < DUP_SEARCH; ++it) {
      const int mid = (lo + hi) >> 1;
      if (hi - lo > 1) {
        if (off[mid] <= e) lo = mid; else hi = mid;
      }
    }
    const uint2 r = s_rect[warp][lo];
    const uint32_t rx0 = r.x & 0xffffu, ry0 = r.x >> 16, w = (r.y & 0xffffu) - rx0;
    const uint32_t local = e - off[lo];
    const uint32_t dy = local / w, dx = local - dy * w;
    tkeys[first + e] = (ry0 + dy) * (uint32_t)grid_x + rx0 + dx;
    vals[first + e] = s_idx[warp][lo];
  }
}

// ------------------------------------------------------------------------------------------
// K4: onesweep LSD radix sort, 8-bit digits.  One histogram kernel for all passes, then one
// chained-scan (decoupled look-back) kernel per pass.  look-back word: [31:30] flag, [29:0] count.
// Keys are u32 (depth bits / tile ids) inside the rasterizer; the u64 instantiation serves the
// stand-alone pdgs_sort_pairs entry point.
// ------------------------------------------------------------------------------------------
#ifndef PDGS_SORT_CTAS
#define PDGS_SORT_CTAS 4  // resident CTAs per SM of the u32 onesweep pass (64 registers)
#endif
#ifndef PDGS_LB_WIDTH
#define PDGS_LB_WIDTH 4
#endif
constexpr int LB_WIDTH = PDGS_LB_WIDTH;
constexpr uint32_t LB_AGG = 1u << 30;
constexpr uint32_t LB_INC = 2u << 30;
constexpr uint32_t LB_MASK = (1u << 30) - 1;

template <typename KeyT>
__global__ void __launch_bounds__(256) sort_histogram_kernel(const KeyT* __restrict__ keys, const uint32_t* n_dev,
                                                             int64_t n_host, int64_t capacity, int npass, int end_bit,
                                                             uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_hist[MAX_PASSES * RADIX];
  const int64_t n = n_dev ? (int64_t)ld_relaxed_u32(n_dev) : n_host;
  if (n > capacity) return;
  for (int i = threadIdx.x; i < npass * RADIX; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t tail0 = 0;  // keys [tail0, n) go through the scalar loop below
  if (sizeof(KeyT) == 4 && (((uintptr_t)keys) & 15u) == 0) {
    // u32 keys: four 16-byte loads in flight per thread, and equal digits of neighbouring keys (the tiles of one
    // Gaussian's rect share their high digit, depth keys share their top byte) merged into one shared-memory atomic
    const int64_t nvec = n / 4;
    const uint4* k4 = reinterpret_cast<const uint4*>(keys);
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += 4 * stride) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < nvec) v[u] = k4[i0 + u * stride];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (i0 + u * stride >= nvec) break;
        const uint32_t kk[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
        for (int p = 0; p < npass; ++p) {
          const int nb = min(RADIX_BITS, end_bit - p * RADIX_BITS);
          const uint32_t msk = (1u << nb) - 1u;
          uint32_t d = (kk[0] >> (p * RADIX_BITS)) & msk, run = 1;
#pragma unroll
          for (int e = 1; e < 4; ++e) {
            const uint32_t de = (kk[e] >> (p * RADIX_BITS)) & msk;
            if (de == d) { ++run; } else { atomicAdd(&s_hist[p * RADIX + d], run); d = de; run = 1; }
          }
          atomicAdd(&s_hist[p * RADIX + d], run);
        }
      }
    }
    tail0 = nvec * 4;
  }
  for (int64_t i = tail0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const KeyT k = keys[i];
    for (int p = 0; p < npass; ++p) {
      const int nb = min(RADIX_BITS, end_bit - p * RADIX_BITS);  // the last pass may cover fewer than 8 bits
      atomicAdd(&s_hist[p * RADIX + (int)((uint32_t)(k >> (p * RADIX_BITS)) & ((1u << nb) - 1u))], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < npass * RADIX; i += blockDim.x) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

// Lanes of the warp holding the same digit.  Built from one ballot per digit bit: MATCH.ANY serialises over the
// distinct values in the warp (about 100 cycles per instruction per SM with random 8-bit digits, ncu r01h) and
// was the bottleneck of every pass.  Four instructions per bit (predicate, VOTE, SELP, LOP3): the sort passes are bound
// by the ALU pipe (ncu r02A: math-pipe throttle 1.6 stalls per issue, 170 instructions per key and pass), and the
// round-1 form spent about six per bit plus a run-time test of the digit width.  The digit is already masked to its
// width, so the bits above it are zero in every lane and their ballots change nothing: no width test is needed.
__device__ __forceinline__ unsigned warp_peers(uint32_t d, int /*nbits*/) {
  unsigned peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < RADIX_BITS; ++b) {
    unsigned m, x;
    asm("{\n\t.reg .pred p;\n\t.reg .u32 t;\n\t"
        "and.b32 t, %2, %3;\n\t"
        "setp.ne.u32 p, t, 0;\n\t"
        "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t"
        "selp.u32 %1, 0xffffffff, 0, p;\n\t}"
        : "=r"(m), "=r"(x)
        : "r"(d), "r"(1u << b));
    peers &= ~(m ^ x);  // lanes whose bit b equals mine
  }
  return peers;
}

template <typename KeyT, int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS, (sizeof(KeyT) == 4) ? (ITEMS <= 8 ? 6 : PDGS_SORT_CTAS) : 3) sort_onesweep_kernel(const KeyT* __restrict__ keys_in,
                                                                    const uint32_t* __restrict__ vals_in,
                                                                    KeyT* __restrict__ keys_out,
                                                                    uint32_t* __restrict__ vals_out, const uint32_t* n_dev,
                                                                    int64_t n_host, int64_t capacity, int shift,
                                                                    uint32_t digit_mask, const uint32_t* __restrict__ hist,
                                                                    uint32_t* ticket, uint32_t* lookback, int use_match) {
  constexpr int TILE_ITEMS = SORT_THREADS * ITEMS;
  constexpr int NW = SORT_THREADS / 32;
  __shared__ KeyT s_keys[TILE_ITEMS];
  __shared__ uint32_t s_vals[TILE_ITEMS];
  __shared__ uint32_t s_wcnt[NW][RADIX];
  __shared__ uint32_t s_cta_base[RADIX];  // exclusive prefix of CTA digit counts (smem slot of the digit's first key)
  __shared__ uint32_t s_goff[RADIX];      // global offset of the digit's first key minus its smem slot
  __shared__ uint32_t s_scan[NW], s_scan_h[NW];
  __shared__ uint32_t s_tile;

  const int64_t n = n_dev ? (int64_t)ld_relaxed_u32(n_dev) : n_host;
  if (n > capacity) return;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  for (int i = threadIdx.x; i < NW * RADIX; i += SORT_THREADS) (&s_wcnt[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const int64_t tile_base = (int64_t)tile * TILE_ITEMS;
  if (tile_base >= n) return;
  const int valid = (int)((n - tile_base < TILE_ITEMS) ? (n - tile_base) : TILE_ITEMS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nbits = __popc(digit_mask);

  // warp-striped load: warp w owns ITEMS*32 consecutive pairs, item i of lane l is seg + i*32 + l
  KeyT key[ITEMS];
  uint32_t rank[ITEMS];
  const int seg = warp * ITEMS * 32;
  const bool full = valid == TILE_ITEMS;  // every tile but the last: no per-item bounds test (the passes are ALU-bound)
  if (full) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) key[i] = keys_in[tile_base + seg + i * 32 + lane];
  } else {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int local = seg + i * 32 + lane;
      key[i] = (local < valid) ? keys_in[tile_base + local]
                               : (KeyT)~(KeyT)0;  // padding: digit = digit_mask, the last bin in use; ranked behind every real key of that
                                                  // bin (higher local index), so it lands in slots >= valid and is dropped
    }
  }
  // per-warp stable ranking: peers of every item first (independent MATCH ops), then one shared-memory atomic
  // per (item, distinct digit) issued back to back -- program order keeps equal digits of later items behind
  // earlier ones -- and only then the broadcasts that consume the atomics' results
  constexpr int CH = 8;  // items ranked together (bounds the live registers)
  static_assert(ITEMS % 4 == 0, "ITEMS must be a multiple of 4");
#pragma unroll
  for (int c = 0; c < ITEMS; c += CH) {
    unsigned peers[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (c + i < ITEMS) {
        const uint32_t d = (uint32_t)(key[c + i] >> shift) & digit_mask;
        // most significant pass: a warp holds few distinct digits (depth keys share their top byte, the 32
        // consecutive pairs of a warp come from a handful of Gaussians), where one MATCH.ANY beats nbits ballots
        peers[i] = use_match ? __match_any_sync(0xffffffffu, d) : warp_peers(d, nbits);
      }
    }
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (c + i < ITEMS) {
        const uint32_t d = (uint32_t)(key[c + i] >> shift) & digit_mask;
        rank[c + i] = 0;
        if (lane == __ffs(peers[i]) - 1) rank[c + i] = atomicAdd(&s_wcnt[warp][d], (uint32_t)__popc(peers[i]));
        __syncwarp();
      }
    }
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (c + i < ITEMS) {
        const uint32_t pre = __shfl_sync(0xffffffffu, rank[c + i], __ffs(peers[i]) - 1);
        rank[c + i] = pre + __popc(peers[i] & ((1u << lane) - 1u));
      }
    }
  }
  __syncthreads();
  // thread d owns digit d: exclusive scan over warps, CTA count
  const int d = threadIdx.x;
  uint32_t cnt = 0;
#pragma unroll
  for (int w = 0; w < NW; ++w) {
    const uint32_t c = s_wcnt[w][d];
    s_wcnt[w][d] = cnt;
    cnt += c;
  }
  const uint32_t n_invalid = (uint32_t)(TILE_ITEMS - valid);
  const uint32_t cnt_valid = cnt - (((uint32_t)d == digit_mask) ? n_invalid : 0u);
  // publish the aggregate as early as possible
  uint32_t* my_lb = lookback + (size_t)tile * RADIX + d;
  st_relaxed_u32(my_lb, (tile == 0 ? LB_INC : LB_AGG) | cnt_valid);
  // global base of digit d = exclusive prefix over digits of the global histogram
  const uint32_t hcount = hist[d];
  uint32_t inc_c = cnt, inc_h = hcount;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, inc_c, o);
    const uint32_t b = __shfl_up_sync(0xffffffffu, inc_h, o);
    if (lane >= o) { inc_c += a; inc_h += b; }
  }
  if (lane == 31) { s_scan[warp] = inc_c; s_scan_h[warp] = inc_h; }
  __syncthreads();
  uint32_t woff_c = 0, woff_h = 0;
#pragma unroll
  for (int w = 0; w < NW; ++w) {
    if (w < warp) { woff_c += s_scan[w]; woff_h += s_scan_h[w]; }
  }
  const uint32_t cta_base = woff_c + inc_c - cnt;  // exclusive over digits (smem slot)
  const uint32_t gbase = woff_h + inc_h - hcount;  // exclusive over digits (global)
  s_cta_base[d] = cta_base;
  __syncthreads();
  // scatter the keys into shared memory in sorted order, then fetch the values: neither needs anything from
  // other CTAs, so both run before the look-back (predecessors get time to publish, the value loads fly meanwhile)
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const uint32_t dd = (uint32_t)(key[i] >> shift) & digit_mask;
    const uint32_t slot = s_cta_base[dd] + s_wcnt[warp][dd] + rank[i];
    s_keys[slot] = key[i];
    rank[i] = slot;
  }
  uint32_t val[ITEMS];
  if (full && vals_in) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) val[i] = vals_in[tile_base + seg + i * 32 + lane];
  } else {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int local = seg + i * 32 + lane;
      val[i] = (local < valid) ? (vals_in ? vals_in[tile_base + local] : (uint32_t)(tile_base + local)) : 0u;
    }
  }
  // Decoupled look-back for digit d, LB_WIDTH predecessors per round trip.  All resident CTAs publish their aggregates
  // at about the same time, so a tile walks back over ~30 of them on average before it meets an inclusive prefix (ncu
  // r02B: 7.3 rounds of 4 per thread, 23 % of the pass's instructions at 17 per predecessor).  Fast path: when every word
  // of the round is already published, fold them with three predicated instructions each; the spin loop only runs for a
  // word that is still empty.
  uint32_t excl = 0;
  if (tile > 0) {
    // 32-bit tile arithmetic (ntiles < 2^22) and one base pointer per thread: the loop is instruction- and latency-bound
    const uint32_t* lb_d = lookback + d;
    int t = (int)tile - 1;
    bool found = false;
    while (!found) {
      uint32_t w[LB_WIDTH];
#pragma unroll
      for (int j = 0; j < LB_WIDTH; ++j) w[j] = (t >= j) ? ld_relaxed_u32(lb_d + (size_t)(uint32_t)(t - j) * RADIX) : LB_INC;
      uint32_t lowest = w[0];
#pragma unroll
      for (int j = 1; j < LB_WIDTH; ++j) lowest = min(lowest, w[j]);
      if (lowest < LB_AGG) {  // a word is not published yet: wait for the words in order
#pragma unroll
        for (int j = 0; j < LB_WIDTH; ++j) {
          if (!found) {
            uint32_t v = w[j];
            while ((v >> 30) == 0u) v = ld_relaxed_u32(lb_d + (size_t)(uint32_t)(t - j) * RADIX);
            excl += v & LB_MASK;
            found = (v >> 30) == 2u;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < LB_WIDTH; ++j) {
          if (!found) {
            excl += w[j] & LB_MASK;
            found = w[j] >= LB_INC;
          }
        }
      }
      t -= LB_WIDTH;
    }
    st_relaxed_u32(my_lb, LB_INC | ((excl + cnt_valid) & LB_MASK));
  }
  s_goff[d] = gbase + excl - cta_base;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) s_vals[rank[i]] = val[i];
  __syncthreads();
  // coalesced write-out; padding keys occupy slots >= valid and are dropped
  for (int j = threadIdx.x; j < valid; j += SORT_THREADS) {
    const KeyT k = s_keys[j];
    const uint32_t dd = (uint32_t)(k >> shift) & digit_mask;
    const uint32_t dst = s_goff[dd] + (uint32_t)j;
    keys_out[dst] = k;
    vals_out[dst] = s_vals[j];
  }
}

// ------------------------------------------------------------------------------------------
// K5: per-tile [first, last) from the sorted tile ids.  ranges is zeroed beforehand.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ranges_kernel(const uint32_t* __restrict__ tkeys, const uint32_t* n_dev,
                                                     int64_t capacity, uint2* __restrict__ ranges) {
  const int64_t n = (int64_t)ld_relaxed_u32(n_dev);
  if (n > capacity) return;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // four pairs per thread from one 16-byte load (tkeys is 128-byte aligned inside the binning buffer)
  const int64_t nvec = ((((uintptr_t)tkeys) & 15u) == 0) ? n / 4 : 0;
  const uint4* k4 = reinterpret_cast<const uint4*>(tkeys);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const uint4 q = k4[v];
    const int64_t i0 = 4 * v;
    const uint32_t t[5] = {i0 > 0 ? tkeys[i0 - 1] : q.x, q.x, q.y, q.z, q.w};
    if (i0 == 0) ranges[q.x].x = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (t[e + 1] != t[e]) {
        ranges[t[e]].y = (uint32_t)(i0 + e);
        ranges[t[e + 1]].x = (uint32_t)(i0 + e);
      }
    }
    if (i0 + 4 == n) ranges[q.w].y = (uint32_t)n;
  }
  for (int64_t i = 4 * nvec + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t t = tkeys[i];
    if (i == 0) ranges[t].x = 0;
    else {
      const uint32_t tp = tkeys[i - 1];
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

cudaError_t launch_duplicate(const DeviceArgs& a, GeomView g, uint32_t* tkeys, uint32_t* vals, int64_t capacity,
                             cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const int64_t per_block = DUP_WARPS * DUP_SEG;
  const unsigned blocks = (unsigned)((a.P + per_block - 1) / per_block);
  duplicate_kernel<<<blocks, DUP_WARPS * 32, 0, s>>>(g, a.P, a.grid_x, tkeys, vals, capacity);
  return cudaGetLastError();
}

cudaError_t launch_scan_duplicate(const DeviceArgs& a, GeomView g, uint32_t* tkeys, uint32_t* vals, int64_t capacity,
                                  cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((a.P + DUP_TILE - 1) / DUP_TILE);
  scan_duplicate_kernel<<<blocks, DUP_WARPS * 32, 0, s>>>(g, a.P, a.grid_x, tkeys, vals, capacity);
  return cudaGetLastError();
}

template <typename KeyT, int ITEMS>
static cudaError_t sort_impl(KeyT* keys_a, uint32_t* vals_a, KeyT* keys_b, uint32_t* vals_b, bool iota_vals,
                             const uint32_t* n_dev, int64_t n_host, int64_t capacity, int npass, int end_bit,
                             void* scratch, cudaStream_t s) {
  if (capacity <= 0 || npass <= 0) return cudaSuccess;
  if (npass > MAX_PASSES) return cudaErrorInvalidValue;
  // MATCH.ANY in the most significant pass pays only where a warp holds few distinct digits: the depth keys (iota_vals:
  // they share their top byte).  For the tile sort the rewritten ballots are faster (r02C: C4 0.166 -> 0.158 ms, C5
  // 4.95 -> 4.21 ms); PDGS_SORT_MATCH=0 / 2 forces it off / on everywhere.
  static const int match_env = [] { const char* e = getenv("PDGS_SORT_MATCH"); return e ? atoi(e) : 1; }();
  const bool match_last = match_env == 2 || (match_env == 1 && iota_vals);
  SortScratch sc = carve_sort_scratch(scratch, capacity, SORT_THREADS * ITEMS, npass);  // callers size for >= npass
  cudaError_t e = cudaMemsetAsync(scratch, 0, sc.zero_bytes, s);
  if (e != cudaSuccess) return e;
  KeyT* kin = (npass % 2 == 0) ? keys_a : keys_b;
  uint32_t* vin = (npass % 2 == 0) ? vals_a : vals_b;
  KeyT* kout = (npass % 2 == 0) ? keys_b : keys_a;
  uint32_t* vout = (npass % 2 == 0) ? vals_b : vals_a;
  // every block ends with npass * 256 global atomics on the same 256-entry histograms, so the block count, not the key
  // count, sets the cost of small inputs: PDGS_HIST_CTAS_PER_SM blocks per SM at most
  static const int hist_per_sm = [] { const char* e = getenv("PDGS_HIST_CTAS_PER_SM"); return e ? atoi(e) : 4; }();
  int hist_blocks = (int)((capacity + 256 * 16 - 1) / (256 * 16));
  if (hist_blocks > sm_count() * hist_per_sm) hist_blocks = sm_count() * hist_per_sm;
  if (hist_blocks < 1) hist_blocks = 1;
  sort_histogram_kernel<KeyT><<<hist_blocks, 256, 0, s>>>(kin, n_dev, n_host, capacity, npass, end_bit, sc.hist);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const unsigned tiles = (unsigned)sc.ntiles;
  for (int p = 0; p < npass; ++p) {
    const int nb = (end_bit - p * RADIX_BITS < RADIX_BITS) ? end_bit - p * RADIX_BITS : RADIX_BITS;
    sort_onesweep_kernel<KeyT, ITEMS><<<tiles, SORT_THREADS, 0, s>>>(
        kin, (p == 0 && iota_vals) ? nullptr : vin, kout, vout, n_dev, n_host, capacity, p * RADIX_BITS, (1u << nb) - 1u,
        sc.hist + p * RADIX, sc.tickets + p, sc.lookback + (size_t)p * sc.ntiles * RADIX,
        (match_last && npass >= 2 && p == npass - 1) ? 1 : 0);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    KeyT* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  return cudaSuccess;
}

cudaError_t launch_sort_u32(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, bool iota_vals,
                            const uint32_t* n_dev, int64_t n_host, int64_t capacity, int npass, int end_bit,
                            void* scratch, cudaStream_t s) {
  return sort_impl<uint32_t, SORT_ITEMS_U32>(keys_a, vals_a, keys_b, vals_b, iota_vals, n_dev, n_host, capacity, npass,
                                             end_bit, scratch, s);
}

cudaError_t launch_sort_u64(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, const uint32_t* n_dev,
                            int64_t capacity, int npass, int end_bit, void* scratch, cudaStream_t s) {
  return sort_impl<uint64_t, SORT_ITEMS_U64>(keys_a, vals_a, keys_b, vals_b, false, n_dev, 0, capacity, npass, end_bit,
                                             scratch, s);
}

cudaError_t launch_ranges(const uint32_t* tkeys, const uint32_t* n_dev, int64_t capacity, uint2* ranges, int tiles,
                          cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), s);
  if (e != cudaSuccess) return e;
  if (capacity <= 0) return cudaSuccess;
  int blocks = (int)((capacity + 255) / 256);
  if (blocks > sm_count() * 16) blocks = sm_count() * 16;
  ranges_kernel<<<blocks, 256, 0, s>>>(tkeys, n_dev, capacity, ranges);
  return cudaGetLastError();
}

}  // namespace pdgs
