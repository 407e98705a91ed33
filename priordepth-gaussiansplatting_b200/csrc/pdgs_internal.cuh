// Internal layout of the per-call state blobs and the launch functions shared
// between translation units of libpdgs.so.  The blobs play the role of the
// reference's geomBuffer / binningBuffer / imgBuffer but the
// layout is SoA float4 records sized for coalesced 16-byte gathers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "pdgs_math.cuh"

namespace pdgs {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int MAX_PASSES = 8;
constexpr int SORT_THREADS = 256;
#ifndef PDGS_SORT_ITEMS
#define PDGS_SORT_ITEMS 16
#endif
constexpr int SORT_ITEMS_U32 = PDGS_SORT_ITEMS;  // pairs per thread, 256 threads per CTA (u32 keys: depth bits, tile ids)
constexpr int SORT_ITEMS_U64 = 12;  // 3072 pairs per CTA (u64 keys: stand-alone pdgs_sort_pairs)
constexpr int DEPTH_SORT_PASSES = 4;
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 4096 counts per CTA
#ifndef PDGS_DUP_SEG
#define PDGS_DUP_SEG 128
#endif
constexpr int DUP_TILE = 8 * PDGS_DUP_SEG;  // sorted Gaussians per CTA of the fused scan + duplicate kernel (8 warps x DUP_SEG)
constexpr uint32_t MAX_PAIRS = 1u << 30;               // look-back words carry 30-bit counts

// Control block at the start of the geometry blob (device memory, 256 B).
struct Ctrl {
  uint32_t num_rendered;  // R = sum of tiles_touched
  uint32_t overflow;      // R >= 2^30
  uint32_t scan_ticket;
  uint32_t error;
  uint32_t sort_ticket[MAX_PASSES];
  uint32_t pad[52];
};
static_assert(sizeof(Ctrl) == 256, "Ctrl must be 256 bytes");

inline __host__ __device__ size_t align_up(size_t v, size_t a = 128) { return (v + a - 1) / a * a; }

// Number of SMs of the current device (grid sizing of the persistent / capped launches); asked once per device.
inline __host__ int sm_count() {
  static int cached[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;  // B200
    cached[dev] = n;
  }
  return cached[dev];
}

struct GeomView {
  Ctrl* ctrl;
  float4* gA;               // (x, y, conic.x, conic.y)
  float4* gB;               // (conic.z, opacity*h, 1/depth, depth)
  float4* gC;               // (r, g, b, clamped bits)
  uint32_t* tiles_touched;  // [P]
  uint2* rect;              // [P] tile rect (x0 | y0 << 16, x1 | y1 << 16); (0, 0) = culled
  uint32_t* depth_keys;     // [P] depth bits (0xffffffff = culled); sorted ascending after the depth sort
  uint32_t* order;          // [P] Gaussian indices in (depth bits, index) order
  uint32_t* depth_keys_b;   // [P] ping-pong buffers of the depth sort
  uint32_t* order_b;        // [P]
  uint32_t* point_offsets;  // [P] inclusive scan of tiles_touched taken in depth order
  unsigned long long* scan_state;  // [ceil(P/DUP_TILE)] (the stand-alone scan uses the first ceil(P/SCAN_TILE) words)
  void* sort_scratch;       // depth-sort histograms / tickets / look-back words
  size_t sort_scratch_bytes;
  size_t bytes;
};

struct SortScratch {
  uint32_t* hist;      // [MAX_PASSES][RADIX] global digit histograms
  uint32_t* tickets;   // [MAX_PASSES]
  uint32_t* lookback;  // [MAX_PASSES][ntiles][RADIX]
  size_t ntiles;
  size_t bytes;        // whole scratch
  size_t zero_bytes;   // prefix that must be zeroed before a sort
};

inline __host__ SortScratch carve_sort_scratch(void* base, int64_t capacity, int tile_items, int npass = MAX_PASSES) {
  SortScratch s;
  char* p = (char*)base;
  size_t off = 0;
  const size_t cap = (size_t)(capacity > 0 ? capacity : 1);
  s.ntiles = (cap + tile_items - 1) / tile_items;
  s.hist = (uint32_t*)(p + off); off += align_up(MAX_PASSES * RADIX * sizeof(uint32_t));
  s.tickets = (uint32_t*)(p + off); off += align_up(MAX_PASSES * sizeof(uint32_t));
  s.lookback = (uint32_t*)(p + off); off += align_up((size_t)npass * s.ntiles * RADIX * sizeof(uint32_t));
  s.bytes = off;
  s.zero_bytes = off;
  return s;
}

inline __host__ GeomView carve_geom(void* base, int64_t P) {
  GeomView v;
  char* p = (char*)base;
  size_t off = 0;
  const size_t Pn = (size_t)(P > 0 ? P : 1);
  v.ctrl = (Ctrl*)(p + off); off += align_up(sizeof(Ctrl));
  v.gA = (float4*)(p + off); off += align_up(Pn * sizeof(float4));
  v.gB = (float4*)(p + off); off += align_up(Pn * sizeof(float4));
  v.gC = (float4*)(p + off); off += align_up(Pn * sizeof(float4));
  v.tiles_touched = (uint32_t*)(p + off); off += align_up(Pn * sizeof(uint32_t));
  v.rect = (uint2*)(p + off); off += align_up(Pn * sizeof(uint2));
  v.depth_keys = (uint32_t*)(p + off); off += align_up(Pn * sizeof(uint32_t));
  v.order = (uint32_t*)(p + off); off += align_up(Pn * sizeof(uint32_t));
  v.depth_keys_b = (uint32_t*)(p + off); off += align_up(Pn * sizeof(uint32_t));
  v.order_b = (uint32_t*)(p + off); off += align_up(Pn * sizeof(uint32_t));
  v.point_offsets = (uint32_t*)(p + off); off += align_up(Pn * sizeof(uint32_t));
  const size_t nscan = (Pn + DUP_TILE - 1) / DUP_TILE;
  v.scan_state = (unsigned long long*)(p + off); off += align_up(nscan * sizeof(unsigned long long));
  v.sort_scratch = (void*)(p + off);
  v.sort_scratch_bytes = carve_sort_scratch(nullptr, P, SORT_THREADS * SORT_ITEMS_U32, DEPTH_SORT_PASSES).bytes;
  off += align_up(v.sort_scratch_bytes);
  v.bytes = off;
  return v;
}

struct ImageView {
  uint2* ranges;        // [tiles]
  float* final_T;       // [H*W]
  uint32_t* n_contrib;  // [H*W]
  uint32_t* n_tested;   // [H*W] list entries evaluated before the pixel stopped (roofline bookkeeping)
  uint32_t* n_accepted; // [H*W] entries blended into the pixel
  size_t bytes;
};

inline __host__ ImageView carve_image(void* base, int W, int H) {
  ImageView v;
  char* p = (char*)base;
  size_t off = 0;
  const size_t tiles = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  const size_t hw = (size_t)W * H;
  v.ranges = (uint2*)(p + off); off += align_up(tiles * sizeof(uint2));
  v.final_T = (float*)(p + off); off += align_up(hw * sizeof(float));
  v.n_contrib = (uint32_t*)(p + off); off += align_up(hw * sizeof(uint32_t));
  v.n_tested = (uint32_t*)(p + off); off += align_up(hw * sizeof(uint32_t));
  v.n_accepted = (uint32_t*)(p + off); off += align_up(hw * sizeof(uint32_t));
  v.bytes = off;
  return v;
}

struct BinningView {
  uint32_t* point_list;  // [cap] sorted Gaussian ids (buffer A values)
  uint32_t* tkeys_a;     // [cap] sorted tile ids (buffer A keys)
  uint32_t* tkeys_b;     // [cap]
  uint32_t* vals_b;      // [cap]
  uint8_t* pair_mask;    // [cap] per sorted pair: 8x4-pixel patches of its tile the Gaussian reaches (forward -> backward)
  void* sort_scratch;
  size_t sort_scratch_bytes;
  size_t bytes;
};

inline int tile_id_bits(int tiles) {
  int bits = 0;
  while ((1 << bits) <= tiles && bits < 31) ++bits;  // index of MSB + 1 (256 -> 9), as the reference trims its keys
  return bits;
}
inline int num_tile_sort_passes(int tiles) { return (tile_id_bits(tiles) + RADIX_BITS - 1) / RADIX_BITS; }

inline __host__ BinningView carve_binning(void* base, int64_t capacity) {
  BinningView v;
  char* p = (char*)base;
  size_t off = 0;
  const size_t cap = (size_t)(capacity > 0 ? capacity : 1);
  v.point_list = (uint32_t*)(p + off); off += align_up(cap * sizeof(uint32_t));
  v.tkeys_a = (uint32_t*)(p + off); off += align_up(cap * sizeof(uint32_t));
  v.tkeys_b = (uint32_t*)(p + off); off += align_up(cap * sizeof(uint32_t));
  v.vals_b = (uint32_t*)(p + off); off += align_up(cap * sizeof(uint32_t));
  v.pair_mask = (uint8_t*)(p + off); off += align_up(cap);
  SortScratch s = carve_sort_scratch(nullptr, capacity, SORT_THREADS * SORT_ITEMS_U32, 4);
  v.sort_scratch = (void*)(p + off);
  v.sort_scratch_bytes = s.bytes;
  off += align_up(s.bytes);
  v.bytes = off;
  return v;
}

// Per-Gaussian gradient accumulator written by the backward blend (12 floats, 3 x float4):
//   [0..3]  dmean2D.x, dmean2D.y, dconic.x, dconic.y
//   [4..7]  dconic.w, dopacity, dinvdepth, -
//   [8..11] dr, dg, db, -
constexpr int ACC_FLOATS = 12;

// ---- launchers (each enqueues on `stream`; returns cudaGetLastError()) ----
struct DeviceArgs {
  const float* means3D; const float* shs; const float* colors_precomp; const float* opacities;
  const float* scales; const float* rotations; const float* cov3D_precomp;
  const float* shs_dc; const float* shs_rest;  // split SH (features_dc / features_rest) instead of shs
  int raw;                                     // opacities / scales / rotations are pre-activation parameters
  const float* viewmatrix; const float* projmatrix; const float* campos; const float* bg;
  int64_t P; int M;
  float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
  int W, H, grid_x, grid_y, sh_degree, antialiasing;
};

cudaError_t launch_preprocess(const DeviceArgs& a, GeomView g, int32_t* radii, cudaStream_t s);
cudaError_t launch_scan(GeomView g, int64_t P, cudaStream_t s);
cudaError_t launch_scan_duplicate(const DeviceArgs& a, GeomView g, uint32_t* tkeys, uint32_t* vals, int64_t capacity,
                                  cudaStream_t s);
cudaError_t launch_duplicate(const DeviceArgs& a, GeomView g, uint32_t* tkeys, uint32_t* vals, int64_t capacity,
                             cudaStream_t s);
// Stable LSD radix sorts of (key, u32 value) pairs between buffers A and B.  The input must sit in A if npass is
// even, in B if odd; the result always lands in A.  n is read from *n_dev (or is n_host when n_dev is NULL) and
// must be <= capacity.  Key bits [0, end_bit) are significant (npass = ceil(end_bit / 8); the last pass masks its
// digit to the remaining bits).  iota_vals: the input values are the element indices and are not read.
cudaError_t launch_sort_u32(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, bool iota_vals,
                            const uint32_t* n_dev, int64_t n_host, int64_t capacity, int npass, int end_bit,
                            void* scratch, cudaStream_t s);
cudaError_t launch_sort_u64(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, const uint32_t* n_dev,
                            int64_t capacity, int npass, int end_bit, void* scratch, cudaStream_t s);
cudaError_t launch_ranges(const uint32_t* tkeys, const uint32_t* n_dev, int64_t capacity, uint2* ranges, int tiles,
                          cudaStream_t s);
// work_counters: also write ImageView::n_tested / n_accepted (roofline accounting; off on the product path)
cudaError_t launch_render_forward(const DeviceArgs& a, GeomView g, ImageView im, const uint32_t* point_list,
                                  uint8_t* pair_mask, int64_t capacity, float* out_color, float* out_invdepth,
                                  int work_counters, cudaStream_t s);
cudaError_t launch_render_backward(const DeviceArgs& a, GeomView g, ImageView im, const uint32_t* point_list,
                                   uint8_t* pair_mask, const float* dL_dcolor, const float* dL_dinvdepth, float* acc, cudaStream_t s);
struct GradPtrs {
  float* means3D; float* means2D; float* shs; float* colors_precomp; float* opacities; float* scales;
  float* rotations; float* cov3D_precomp; int accumulate;
  float* densify_grad_accum; float* densify_denom; int32_t* max_radii2D;
  float* color_grad_view;  // deferred SH backward: masked dL/drgb of this view instead of SH / direction gradients
  float* shs_dc; float* shs_rest;  // split SH gradients
};
cudaError_t launch_preprocess_backward(const DeviceArgs& a, GeomView g, const int32_t* radii, const float* acc,
                                       GradPtrs gp, cudaStream_t s);
// shs != NULL: packed [P,M,3] rows, gradient to grad_shs.  Otherwise split rows (shs_dc [P,3], shs_rest [P,3(M-1)]),
// gradients to grad_dc / grad_rest.
cudaError_t launch_sh_backward_views(const float* means3D, const float* shs, const float* shs_dc, const float* shs_rest,
                                     int64_t P, int M, int sh_degree, const float* campos_views, const float* color_grads,
                                     int n_views, float* grad_shs, float* grad_dc, float* grad_rest, float* grad_means3D,
                                     int accumulate, int64_t first, int64_t count, cudaStream_t s);
// alpha / clamp01: the loss sees clamp(color, 0, 1) * alpha (alpha: one [H*W] plane or NULL), see pdgs_loss.cu
cudaError_t launch_image_loss(const float* color, const float* target, const float* invdepth, const float* prior,
                              const float* mask, const float* alpha, int clamp01, int W, int H, float w_color,
                              float w_depth, float* dL_dcolor, float* dL_dinvdepth, float* loss, cudaStream_t s);
// SSIM (pdgs_ssim.cu).  sums: device double[2] zeroed by the caller (sum of the map, sum |img1 - img2|)
cudaError_t launch_ssim_forward(float C1, float C2, const float* img1, const float* img2, int planes, int W, int H,
                                float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sums,
                                int want_l1, const float* alpha, int clamp01, cudaStream_t s);
cudaError_t launch_ssim_finalize(const double* sums, double scale, float* out_ssim, float* out_l1, cudaStream_t s);
cudaError_t launch_ssim_backward(const float* img1, const float* img2, const float* dL_dmap, float dL_uniform,
                                 const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, int planes, int W,
                                 int H, float l1_weight, float* dL_dimg1, const double* fin_sums, double fin_scale,
                                 float* fin_ssim, float* fin_l1, const float* alpha, int clamp01, cudaStream_t s);
// Visibility-sparse Adam (pdgs_optim.cu).  One group = one parameter tensor [N, M] with its gradient and moments.
constexpr int ADAM_MAX_GROUPS = 8;
struct AdamGroup {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  double lr_host;  // caller's learning rate; launch_sparse_adam derives the fp32 step size `lr`
  float lr; int M; int vec4;
};
// visible [N] bytes, or (visible == NULL) radii [N] with visible = radii > 0.  step > 0: torch.optim.Adam bias
// correction for that step count; step == 0: none (the upstream sparse kernel's rule).
cudaError_t launch_sparse_adam(const AdamGroup* groups, int n_groups, const uint8_t* visible, const int32_t* radii,
                               int64_t N, double b1, double b2, double eps, int step, cudaStream_t s);
cudaError_t launch_mark_visible(const float* means3D, const float* viewmatrix, int64_t P, uint8_t* present,
                                cudaStream_t s);

}  // namespace pdgs
