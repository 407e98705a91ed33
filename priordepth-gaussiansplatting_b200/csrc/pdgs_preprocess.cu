// K1 preprocess (forward), fused K8+K9 preprocess backward, K10 mark_visible.
// Replaces preprocessCUDA / computeCov2DCUDA-backward / preprocessCUDA-backward / checkFrustum
// of the reference rasterizer (source absent).
#include "pdgs_async.cuh"
#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

constexpr int PRE_THREADS = 256;

// Accumulation into the per-Gaussian gradient rows (batch path) uses fire-and-forget reductions: the add happens in
// L2 and the thread never waits for the old value (a load-add-store would put one more DRAM round trip at the end of
// every thread of this latency-bound kernel).  Atomic, so several views' backward kernels may run concurrently on
// different streams and add into the same rows (ViewParallelRasterizer lanes).
__device__ __forceinline__ void red_add(float* p, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void load_camera(const DeviceArgs& a, Camera* cam) {
  // one CTA-wide cooperative load of the 35 camera floats into shared memory
  const int t = threadIdx.x;
  if (t < 16) cam->vm[t] = __ldg(a.viewmatrix + t);
  else if (t < 32) cam->pm[t - 16] = __ldg(a.projmatrix + (t - 16));
  else if (t < 35) cam->campos[t - 32] = a.campos ? __ldg(a.campos + (t - 32)) : 0.f;
  else if (t == 35) {
    cam->tanfovx = a.tanfovx; cam->tanfovy = a.tanfovy;
    cam->focal_x = a.focal_x; cam->focal_y = a.focal_y;
    cam->scale_modifier = a.scale_modifier;
    cam->W = a.W; cam->H = a.H; cam->grid_x = a.grid_x; cam->grid_y = a.grid_y;
    cam->sh_degree = a.sh_degree; cam->antialiasing = a.antialiasing;
  }
  __syncthreads();
}

struct ShGlobal {  // per-Gaussian SH block read through the read-only path
  const float* p;
  __device__ __forceinline__ float operator()(int i) const { return __ldg(p + i); }
};

struct ShRegs {  // SH block held in registers (48 floats, loaded as 12 x float4)
  float v[48];
  __device__ __forceinline__ float operator()(int i) const { return v[i]; }
};

// ---- fused activations (PdgsScene.raw_params; /root/reference/scene/gaussian_model.py:40-48, 103-131) ----
// scales = exp(_scaling), rotations = _rotation / max(|_rotation|, 1e-12) (torch.nn.functional.normalize),
// opacity = sigmoid(_opacity).  expf / sqrtf / division are the IEEE-rounded libdevice versions (the build does
// not use fast-math), the same functions torch's elementwise kernels call.
__device__ __forceinline__ void load_scale_rot(const DeviceArgs& a, int64_t idx, float sc[3], float q[4], float& qinv) {
  sc[0] = __ldg(a.scales + 3 * idx); sc[1] = __ldg(a.scales + 3 * idx + 1); sc[2] = __ldg(a.scales + 3 * idx + 2);
  const float4 q4 = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
  q[0] = q4.x; q[1] = q4.y; q[2] = q4.z; q[3] = q4.w;
  qinv = 1.0f;
  if (a.raw) {
    sc[0] = expf(sc[0]); sc[1] = expf(sc[1]); sc[2] = expf(sc[2]);
    const float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    qinv = 1.0f / fmaxf(n, 1e-12f);
    q[0] = q[0] * qinv; q[1] = q[1] * qinv; q[2] = q[2] * qinv; q[3] = q[3] * qinv;
  }
}
__device__ __forceinline__ float load_opacity(const DeviceArgs& a, int64_t idx) {
  const float x = __ldg(a.opacities + idx);
  return a.raw ? 1.0f / (1.0f + expf(-x)) : x;
}
// chain rule through the activations: sc, q are the ACTIVATED values, qinv = 1 / max(|_rotation|, eps)
__device__ __forceinline__ void raw_scale_rot_backward(const float sc[3], const float q[4], float qinv, float g_scale[3],
                                                       float g_rot[4]) {
  g_scale[0] *= sc[0]; g_scale[1] *= sc[1]; g_scale[2] *= sc[2];
  const float d = q[0] * g_rot[0] + q[1] * g_rot[1] + q[2] * g_rot[2] + q[3] * g_rot[3];
#pragma unroll
  for (int k = 0; k < 4; ++k) g_rot[k] = (g_rot[k] - q[k] * d) * qinv;
}
__device__ __forceinline__ float raw_opacity_backward(float o, float g) { return g * o * (1.0f - o); }

struct ShSplitGlobal {  // split SH row (features_dc + features_rest) through the read-only path
  const float* dc;
  const float* rest;
  __device__ __forceinline__ float operator()(int i) const { return i < 3 ? __ldg(dc + i) : __ldg(rest + (i - 3)); }
};

constexpr int SPLIT_REST = 45;                                   // floats per features_rest row at M == 16
constexpr int SPLIT_WARP_FLOATS = 32 * SPLIT_REST;               // a warp's 32 rows: 5760 contiguous bytes
constexpr int PRE_SPLIT_SMEM = (PRE_THREADS / 32) * SPLIT_WARP_FLOATS * (int)sizeof(float);

// A warp copies the features_rest rows of its 32 Gaussians (one contiguous, 16-byte aligned 5760-byte span) into
// its shared-memory slice with coalesced 16-byte loads; each lane then reads its own 45-float row with stride 45
// (odd, so conflict-free).  The 180-byte rows themselves are only 4-byte aligned, so per-thread vector loads are
// not possible on this layout.
// vmask: lanes (rows) whose Gaussian is visible; a 16-byte piece is fetched only if one of the (at most two) rows it
// overlaps is visible, so rows of culled Gaussians cost no DRAM traffic beyond the sectors they share with neighbours.
__device__ __forceinline__ void warp_stage_rest(const float* __restrict__ rest, int64_t warp_row0, int64_t P, float* slice,
                                                unsigned vmask) {
  const int lane = threadIdx.x & 31;
  const int64_t left = P - warp_row0;
  const float* src = rest + warp_row0 * SPLIT_REST;
  if (left >= 32) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(slice);
#pragma unroll
    for (int i = 0; i < 12; ++i) {
      const int j = lane + 32 * i;
      if (j < SPLIT_WARP_FLOATS / 4) {
        const int r0 = (4 * j) / SPLIT_REST, r1 = (4 * j + 3) / SPLIT_REST;
        if (((vmask >> r0) | (vmask >> r1)) & 1u) d4[j] = __ldg(s4 + j);
      }
    }
  } else {
    const int n = (int)(left > 0 ? left : 0) * SPLIT_REST;
    for (int j = lane; j < n; j += 32) slice[j] = __ldg(src + j);
  }
  __syncwarp();
}

#ifndef PDGS_PRE_CTAS
#define PDGS_PRE_CTAS 3
#endif
template <bool SPLIT>
__global__ void __launch_bounds__(PRE_THREADS, PDGS_PRE_CTAS) preprocess_kernel(DeviceArgs a, GeomView g, int32_t* __restrict__ radii) {
  __shared__ Camera cam;
  extern __shared__ __align__(16) float pre_stage[];  // SPLIT && M == 16: [warps][32 * 45]
  load_camera(a, &cam);
  const int64_t idx = (int64_t)blockIdx.x * PRE_THREADS + threadIdx.x;
  const bool in_range = idx < a.P;
  if (!SPLIT && !in_range) return;
  bool vis = false;
  GeomOut go;
  float mx = 0.f, my = 0.f, mz = 0.f;
  if (in_range) {
    mx = __ldg(a.means3D + 3 * idx); my = __ldg(a.means3D + 3 * idx + 1); mz = __ldg(a.means3D + 3 * idx + 2);
    float cov[6];
    if (a.cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; ++k) cov[k] = __ldg(a.cov3D_precomp + 6 * idx + k);
    } else {
      float sc[3], q[4], qinv;
      load_scale_rot(a, idx, sc, q, qinv);
      cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov);
    }
    vis = forward_geometry(cam, mx, my, mz, cov, go);
  }
  float* slice = nullptr;
  if (SPLIT && a.M == 16) {
    slice = pre_stage + (threadIdx.x >> 5) * SPLIT_WARP_FLOATS;
    const unsigned vmask = __ballot_sync(0xffffffffu, vis);
    if (vmask) warp_stage_rest(a.shs_rest, idx - (threadIdx.x & 31), a.P, slice, vmask);
  }
  if (!in_range) return;
  if (!vis) {
    radii[idx] = 0;
    g.tiles_touched[idx] = 0;
    g.rect[idx] = make_uint2(0u, 0u);
    g.depth_keys[idx] = 0xffffffffu;  // sorts behind every visible Gaussian (depth > 0.2 has the sign bit clear)
    return;
  }
  float rgb[3];
  unsigned clamped = 0;
  if (a.colors_precomp) {
    rgb[0] = __ldg(a.colors_precomp + 3 * idx);
    rgb[1] = __ldg(a.colors_precomp + 3 * idx + 1);
    rgb[2] = __ldg(a.colors_precomp + 3 * idx + 2);
  } else {
    const float dx = mx - cam.campos[0], dy = my - cam.campos[1], dz = mz - cam.campos[2];
    if (SPLIT) {
      if (a.M == 16) {
        ShRegs sh;
        sh.v[0] = __ldg(a.shs_dc + 3 * idx); sh.v[1] = __ldg(a.shs_dc + 3 * idx + 1); sh.v[2] = __ldg(a.shs_dc + 3 * idx + 2);
        const float* row = slice + (threadIdx.x & 31) * SPLIT_REST;
#pragma unroll
        for (int k = 0; k < SPLIT_REST; ++k) sh.v[3 + k] = row[k];
        clamped = sh_to_rgb(cam.sh_degree, dx, dy, dz, sh, rgb);
      } else {
        ShSplitGlobal sh{a.shs_dc + 3 * idx, a.shs_rest + (size_t)idx * 3 * (a.M - 1)};
        clamped = sh_to_rgb(cam.sh_degree, dx, dy, dz, sh, rgb);
      }
    } else if (a.M == 16) {
      ShRegs sh;
      const float4* src = reinterpret_cast<const float4*>(a.shs + (size_t)idx * 48);
#pragma unroll
      for (int k = 0; k < 12; ++k) {
        const float4 v = __ldg(src + k);
        sh.v[4 * k] = v.x; sh.v[4 * k + 1] = v.y; sh.v[4 * k + 2] = v.z; sh.v[4 * k + 3] = v.w;
      }
      clamped = sh_to_rgb(cam.sh_degree, dx, dy, dz, sh, rgb);
    } else {
      ShGlobal sh{a.shs + (size_t)idx * a.M * 3};
      clamped = sh_to_rgb(cam.sh_degree, dx, dy, dz, sh, rgb);
    }
  }
  const float opacity = load_opacity(a, idx);
  g.gA[idx] = make_float4(go.px, go.py, go.conA, go.conB);
  g.gB[idx] = make_float4(go.conC, opacity * go.h, PD_DIV(1.0f, go.depth), go.depth);
  g.gC[idx] = make_float4(rgb[0], rgb[1], rgb[2], __uint_as_float(clamped));
  radii[idx] = go.radius;
  g.depth_keys[idx] = __float_as_uint(go.depth);
  g.rect[idx] = make_uint2((uint32_t)go.rx0 | ((uint32_t)go.ry0 << 16), (uint32_t)go.rx1 | ((uint32_t)go.ry1 << 16));
  g.tiles_touched[idx] = (uint32_t)((go.rx1 - go.rx0) * (go.ry1 - go.ry0));
}

// K8+K9 for scenes with precomputed colours (no SH): one thread per Gaussian, plain loads and stores.
__global__ void __launch_bounds__(PRE_THREADS) preprocess_backward_kernel(DeviceArgs a, GeomView g,
                                                                        const int32_t* __restrict__ radii,
                                                                        const float* __restrict__ acc, GradPtrs gp) {
  __shared__ Camera cam;
  load_camera(a, &cam);
  const int64_t idx = (int64_t)blockIdx.x * PRE_THREADS + threadIdx.x;
  if (idx >= a.P) return;
  const bool vis = radii[idx] > 0;
  if (!vis) {
    if (!gp.accumulate) {  // dense zero rows for culled Gaussians
      if (gp.means3D) { gp.means3D[3 * idx] = 0.f; gp.means3D[3 * idx + 1] = 0.f; gp.means3D[3 * idx + 2] = 0.f; }
      if (gp.means2D) { gp.means2D[3 * idx] = 0.f; gp.means2D[3 * idx + 1] = 0.f; gp.means2D[3 * idx + 2] = 0.f; }
      if (gp.opacities) gp.opacities[idx] = 0.f;
      if (gp.colors_precomp) { gp.colors_precomp[3 * idx] = 0.f; gp.colors_precomp[3 * idx + 1] = 0.f; gp.colors_precomp[3 * idx + 2] = 0.f; }
      if (gp.scales) { gp.scales[3 * idx] = 0.f; gp.scales[3 * idx + 1] = 0.f; gp.scales[3 * idx + 2] = 0.f; }
      if (gp.rotations) reinterpret_cast<float4*>(gp.rotations)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gp.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; ++k) gp.cov3D_precomp[6 * idx + k] = 0.f;
      }
    }
    return;
  }
  const float4 a0 = __ldg(reinterpret_cast<const float4*>(acc) + 3 * idx);
  const float4 a1 = __ldg(reinterpret_cast<const float4*>(acc) + 3 * idx + 1);
  const float4 a2 = __ldg(reinterpret_cast<const float4*>(acc) + 3 * idx + 2);
  const float4 gA = g.gA[idx], gB = g.gB[idx], gC = g.gC[idx];
  BwdIn in;
  {
    // acc holds the raw pixel moments of z = o*G*dL/dalpha about the Gaussian's centre (pdgs_render.cu):
    //   a0 = (sum z dx, sum z dy, sum z dx^2, sum z dx dy), a1 = (sum z dy^2, sum z, sum w dL/dD, -), a2 = sum w dL/dC
    // with d = centre - pixel.  dL/dG = o*dL/dalpha, dG/dd = -G (A dx + B dy, C dy + B dx), dG/dconic = -0.5 G d d^T,
    // dL/do = sum G dL/dalpha = (sum z) / o.
    const float cA = gA.z, cB = gA.w, cC = gB.x, o = gB.y;
    in.g_mean2d_x = -(0.5f * (float)a.W) * (cA * a0.x + cB * a0.y);
    in.g_mean2d_y = -(0.5f * (float)a.H) * (cC * a0.y + cB * a0.x);
    in.g_conic_x = -0.5f * a0.z; in.g_conic_y = -0.5f * a0.w; in.g_conic_w = -0.5f * a1.x;
    in.g_opacity = (o > 0.f) ? a1.y / o : 0.f; in.g_invdepth = a1.z;
    in.g_rgb[0] = a2.x; in.g_rgb[1] = a2.y; in.g_rgb[2] = a2.z;
  }
  in.clamped = __float_as_uint(gC.w);

  const float mx = __ldg(a.means3D + 3 * idx), my = __ldg(a.means3D + 3 * idx + 1), mz = __ldg(a.means3D + 3 * idx + 2);
  float cov[6];
  float sc[3] = {0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f}, qinv = 1.0f;
  if (a.cov3D_precomp) {
#pragma unroll
    for (int k = 0; k < 6; ++k) cov[k] = __ldg(a.cov3D_precomp + 6 * idx + k);
  } else {
    load_scale_rot(a, idx, sc, q, qinv);
    cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov);
  }
  const float opacity = load_opacity(a, idx);
  float g_mean[3], g_op, g_cov[6];
  geometry_backward(cam, mx, my, mz, cov, opacity, in, g_mean, g_op, g_cov);
  if (a.raw) g_op = raw_opacity_backward(opacity, g_op);

  // colour (this kernel only serves precomputed colours; SH scenes take preprocess_backward_sh_kernel)
  if (gp.colors_precomp) {
    if (gp.accumulate) {
      red_add(gp.colors_precomp + 3 * idx, in.g_rgb[0]); red_add(gp.colors_precomp + 3 * idx + 1, in.g_rgb[1]);
      red_add(gp.colors_precomp + 3 * idx + 2, in.g_rgb[2]);
    } else {
      gp.colors_precomp[3 * idx] = in.g_rgb[0]; gp.colors_precomp[3 * idx + 1] = in.g_rgb[1]; gp.colors_precomp[3 * idx + 2] = in.g_rgb[2];
    }
  }
  // covariance
  if (a.cov3D_precomp) {
    if (gp.cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        if (gp.accumulate) red_add(gp.cov3D_precomp + 6 * idx + k, g_cov[k]); else gp.cov3D_precomp[6 * idx + k] = g_cov[k];
      }
    }
  } else {
    float g_scale[3], g_rot[4];
    cov3d_backward(sc, cam.scale_modifier, q, g_cov, g_scale, g_rot);
    if (a.raw) raw_scale_rot_backward(sc, q, qinv, g_scale, g_rot);
    if (gp.scales) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (gp.accumulate) red_add(gp.scales + 3 * idx + k, g_scale[k]); else gp.scales[3 * idx + k] = g_scale[k];
      }
    }
    if (gp.rotations) {
      if (gp.accumulate) red_add4(gp.rotations + 4 * idx, g_rot[0], g_rot[1], g_rot[2], g_rot[3]);
      else reinterpret_cast<float4*>(gp.rotations)[idx] = make_float4(g_rot[0], g_rot[1], g_rot[2], g_rot[3]);
    }
  }
  if (gp.means3D) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (gp.accumulate) red_add(gp.means3D + 3 * idx + k, g_mean[k]); else gp.means3D[3 * idx + k] = g_mean[k];
    }
  }
  if (gp.means2D) {
    if (gp.accumulate) { red_add(gp.means2D + 3 * idx, in.g_mean2d_x); red_add(gp.means2D + 3 * idx + 1, in.g_mean2d_y); }
    else { gp.means2D[3 * idx] = in.g_mean2d_x; gp.means2D[3 * idx + 1] = in.g_mean2d_y; gp.means2D[3 * idx + 2] = 0.f; }
  }
  if (gp.opacities) {
    if (gp.accumulate) red_add(gp.opacities + idx, g_op); else gp.opacities[idx] = g_op;
  }
  // densification statistics of this view (scene/gaussian_model.py:502-506, train.py:313-314)
  if (gp.densify_grad_accum)
    red_add(gp.densify_grad_accum + idx, sqrtf(in.g_mean2d_x * in.g_mean2d_x + in.g_mean2d_y * in.g_mean2d_y));
  if (gp.densify_denom) red_add(gp.densify_denom + idx, 1.0f);
  if (gp.max_radii2D) atomicMax(gp.max_radii2D + idx, radii[idx]);
}

// ------------------------------------------------------------------------------------------
// K8+K9 for the SH path, streamed through shared memory with bulk asynchronous copies.
// The generic kernel above is latency-bound (ncu r01f: 112 registers, 16 warps/SM, long-scoreboard
// stalls 16 cycles per issue, 43 % of DRAM peak): every thread gathers its own 192-byte SH row and
// read-modify-writes its own 192-byte gradient row with 16-byte accesses 192 bytes apart.  Here one
// elected thread moves the CTA's contiguous [256][3M] SH block and [256][12] accumulator block into
// shared memory with cp.async.bulk (completion on an mbarrier) and the CTA's SH-gradient block goes
// back with one cp.async.bulk store -- or, when gradients accumulate over views, one
// cp.reduce.async.bulk .add.f32, so the read half of the read-modify-write happens in L2 and never
// occupies the SM.  Two CTAs share an SM: one computes while the other's copies are in flight.
// ------------------------------------------------------------------------------------------
constexpr int PB_TILE = 256;

struct PbShared {  // the deferred-SH variant allocates only the part before sh_in
  float acc_in[PB_TILE * ACC_FLOATS];
  Camera cam;
  unsigned long long bar;
  float sh_in[PB_TILE * 48];
  float sh_out[PB_TILE * 48];
};

struct ShRow {  // SH block of one Gaussian in shared memory
  const float* p;
  __device__ __forceinline__ float operator()(int i) const { return p[i]; }
};
struct ShRowSplit {  // split SH row in shared memory: 3 DC floats + 3(M-1) rest floats
  const float* dc;
  const float* rest;
  __device__ __forceinline__ float operator()(int i) const { return i < 3 ? dc[i] : rest[i - 3]; }
};

// DEFER: the SH part of the backward is left to sh_backward_views_kernel (one pass per batch of views); this
// kernel then only emits the view's masked dL/drgb and never reads or writes an SH row.
// SPLIT: the scene holds features_dc [P,3] and features_rest [P,3(M-1)] instead of one [P,3M] array.  The staging
// regions sh_in / sh_out then hold the tile's rest block ([256][3(M-1)]) followed by its DC block ([256][3]) -- the
// same 256*3M floats -- each moved by its own bulk copy.

#ifndef PDGS_PB_DEFER_CTAS
#define PDGS_PB_DEFER_CTAS 3
#endif
template <bool DEFER, bool SPLIT>
__global__ void __launch_bounds__(PB_TILE, DEFER ? PDGS_PB_DEFER_CTAS : 2) preprocess_backward_sh_kernel(DeviceArgs a, GeomView g,
                                                                           const int32_t* __restrict__ radii,
                                                                           const float* __restrict__ acc, GradPtrs gp,
                                                                           int bulk_ok) {
  extern __shared__ __align__(128) unsigned char pb_raw[];
  PbShared& sm = *reinterpret_cast<PbShared*>(pb_raw);
  load_camera(a, &sm.cam);
  const Camera& cam = sm.cam;
  const int tid = threadIdx.x;
  const int M = a.M, row = 3 * M;  // floats per SH row
  const int rrow = 3 * (M - 1);    // SPLIT: floats per features_rest row
  float* const dc_in = sm.sh_in + PB_TILE * rrow;
  float* const dc_out = sm.sh_out + PB_TILE * rrow;
  if (tid == 0) mbar_init(reinterpret_cast<uint64_t*>(&sm.bar), 1);
  __syncthreads();
  uint32_t phase = 0;
  const int64_t ntiles = (a.P + PB_TILE - 1) / PB_TILE;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t t0 = tile * PB_TILE;
    const int n = (int)((a.P - t0 < PB_TILE) ? (a.P - t0) : PB_TILE);
    const bool bulk = bulk_ok && n == PB_TILE;  // full tile and 16-byte aligned arrays
    if (bulk) {
      if (tid == 0) {
        const uint32_t sh_bytes = DEFER ? 0u : (uint32_t)(PB_TILE * row * sizeof(float)), acc_bytes = PB_TILE * ACC_FLOATS * 4;
        mbar_expect_tx(reinterpret_cast<uint64_t*>(&sm.bar), sh_bytes + acc_bytes);
        if (!DEFER && !SPLIT) bulk_g2s(sm.sh_in, a.shs + (size_t)t0 * row, sh_bytes, reinterpret_cast<uint64_t*>(&sm.bar));
        if (!DEFER && SPLIT) {
          if (rrow > 0)
            bulk_g2s(sm.sh_in, a.shs_rest + (size_t)t0 * rrow, (uint32_t)(PB_TILE * rrow * sizeof(float)),
                     reinterpret_cast<uint64_t*>(&sm.bar));
          bulk_g2s(dc_in, a.shs_dc + (size_t)t0 * 3, (uint32_t)(PB_TILE * 3 * sizeof(float)), reinterpret_cast<uint64_t*>(&sm.bar));
        }
        bulk_g2s(sm.acc_in, acc + (size_t)t0 * ACC_FLOATS, acc_bytes, reinterpret_cast<uint64_t*>(&sm.bar));
      }
    } else {
      if (!DEFER && !SPLIT)
        for (int i = tid; i < n * row; i += PB_TILE) sm.sh_in[i] = __ldg(a.shs + (size_t)t0 * row + i);
      if (!DEFER && SPLIT) {
        for (int i = tid; i < n * rrow; i += PB_TILE) sm.sh_in[i] = __ldg(a.shs_rest + (size_t)t0 * rrow + i);
        for (int i = tid; i < n * 3; i += PB_TILE) dc_in[i] = __ldg(a.shs_dc + (size_t)t0 * 3 + i);
      }
      for (int i = tid; i < n * ACC_FLOATS; i += PB_TILE) sm.acc_in[i] = __ldg(acc + (size_t)t0 * ACC_FLOATS + i);
    }
    // per-thread inputs that are small or already coalesced come straight from global memory
    const int64_t idx = t0 + tid;
    const bool in_range = tid < n;
    const int radius = in_range ? radii[idx] : 0;
    const bool vis = radius > 0;
    float mx = 0.f, my = 0.f, mz = 0.f, opacity = 0.f;
    float cov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, sc[3] = {0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f}, qinv = 1.0f;
    float4 gA = make_float4(0.f, 0.f, 0.f, 0.f), gB = gA, gC = gA;
    if (vis) {
      mx = __ldg(a.means3D + 3 * idx); my = __ldg(a.means3D + 3 * idx + 1); mz = __ldg(a.means3D + 3 * idx + 2);
      if (a.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cov[k] = __ldg(a.cov3D_precomp + 6 * idx + k);
      } else {
        load_scale_rot(a, idx, sc, q, qinv);
      }
      opacity = load_opacity(a, idx);
      gA = g.gA[idx]; gB = g.gB[idx]; gC = g.gC[idx];
    }
    if (!DEFER && tid == 0) bulk_wait_read0();  // the previous tile's bulk store has finished reading sh_out
    __syncthreads();
    if (bulk) {
      mbar_wait(reinterpret_cast<uint64_t*>(&sm.bar), phase);
      phase ^= 1u;
    }
    float* out_row = DEFER ? nullptr : (SPLIT ? sm.sh_out + tid * rrow : sm.sh_out + tid * row);
    float* out_dc = dc_out + tid * 3;  // SPLIT only
    const bool want_sh = !DEFER && (SPLIT ? gp.shs_dc != nullptr : gp.shs != nullptr);
    if (!vis) {
      if (want_sh) {
        if (SPLIT) {
          for (int k = 0; k < rrow; ++k) out_row[k] = 0.f;
          out_dc[0] = 0.f; out_dc[1] = 0.f; out_dc[2] = 0.f;
        } else {
          for (int k = 0; k < row; ++k) out_row[k] = 0.f;
        }
      }
      if (DEFER && in_range) {
        gp.color_grad_view[3 * idx] = 0.f; gp.color_grad_view[3 * idx + 1] = 0.f; gp.color_grad_view[3 * idx + 2] = 0.f;
      }
      if (in_range && !gp.accumulate) {  // dense zero rows for culled Gaussians
        if (gp.means3D) { gp.means3D[3 * idx] = 0.f; gp.means3D[3 * idx + 1] = 0.f; gp.means3D[3 * idx + 2] = 0.f; }
        if (gp.means2D) { gp.means2D[3 * idx] = 0.f; gp.means2D[3 * idx + 1] = 0.f; gp.means2D[3 * idx + 2] = 0.f; }
        if (gp.opacities) gp.opacities[idx] = 0.f;
        if (gp.scales) { gp.scales[3 * idx] = 0.f; gp.scales[3 * idx + 1] = 0.f; gp.scales[3 * idx + 2] = 0.f; }
        if (gp.rotations) reinterpret_cast<float4*>(gp.rotations)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gp.cov3D_precomp) {
#pragma unroll
          for (int k = 0; k < 6; ++k) gp.cov3D_precomp[6 * idx + k] = 0.f;
        }
      }
    } else {
      const float4* accv = reinterpret_cast<const float4*>(sm.acc_in + tid * ACC_FLOATS);
      const float4 a0 = accv[0], a1 = accv[1], a2 = accv[2];
      BwdIn in;
      {
        // acc holds the raw pixel moments of z = o*G*dL/dalpha about the Gaussian's centre (pdgs_render.cu)
        const float cA = gA.z, cB = gA.w, cC = gB.x, o = gB.y;
        in.g_mean2d_x = -(0.5f * (float)a.W) * (cA * a0.x + cB * a0.y);
        in.g_mean2d_y = -(0.5f * (float)a.H) * (cC * a0.y + cB * a0.x);
        in.g_conic_x = -0.5f * a0.z; in.g_conic_y = -0.5f * a0.w; in.g_conic_w = -0.5f * a1.x;
        in.g_opacity = (o > 0.f) ? a1.y / o : 0.f; in.g_invdepth = a1.z;
        in.g_rgb[0] = a2.x; in.g_rgb[1] = a2.y; in.g_rgb[2] = a2.z;
      }
      in.clamped = __float_as_uint(gC.w);
      if (!a.cov3D_precomp) cov3d_from_scale_rot(sc, cam.scale_modifier, q, cov);
      float g_mean[3], g_op, g_cov[6];
      geometry_backward(cam, mx, my, mz, cov, opacity, in, g_mean, g_op, g_cov);
      if (a.raw) g_op = raw_opacity_backward(opacity, g_op);
      // colour: dL/dsh rows into shared memory, view-direction term into dL/dmean
      float grgb[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) grgb[c] = ((in.clamped >> c) & 1u) ? 0.f : in.g_rgb[c];
      if (DEFER) {
        gp.color_grad_view[3 * idx] = grgb[0]; gp.color_grad_view[3 * idx + 1] = grgb[1]; gp.color_grad_view[3 * idx + 2] = grgb[2];
      } else {
        const float dx = mx - cam.campos[0], dy = my - cam.campos[1], dz = mz - cam.campos[2];
        float gdir[3], basis[16];
        if (M == 16) {
          ShRegs sh;
          if (SPLIT) {
            sh.v[0] = dc_in[tid * 3]; sh.v[1] = dc_in[tid * 3 + 1]; sh.v[2] = dc_in[tid * 3 + 2];
            const float* src = sm.sh_in + tid * SPLIT_REST;  // odd stride: conflict-free scalar reads
#pragma unroll
            for (int k = 0; k < SPLIT_REST; ++k) sh.v[3 + k] = src[k];
          } else {
            const float4* src = reinterpret_cast<const float4*>(sm.sh_in + tid * 48);
#pragma unroll
            for (int k = 0; k < 12; ++k) {
              const float4 v = src[k];
              sh.v[4 * k] = v.x; sh.v[4 * k + 1] = v.y; sh.v[4 * k + 2] = v.z; sh.v[4 * k + 3] = v.w;
            }
          }
          sh_backward_dir(cam.sh_degree, dx, dy, dz, sh, grgb, gdir, basis);
        } else if (SPLIT) {
          ShRowSplit sh{dc_in + tid * 3, sm.sh_in + tid * rrow};
          sh_backward_dir(cam.sh_degree, dx, dy, dz, sh, grgb, gdir, basis);
        } else {
          ShRow sh{sm.sh_in + tid * row};
          sh_backward_dir(cam.sh_degree, dx, dy, dz, sh, grgb, gdir, basis);
        }
        g_mean[0] += gdir[0]; g_mean[1] += gdir[1]; g_mean[2] += gdir[2];
        if (want_sh) {
          const int ncoef = sh_num_coeffs(cam.sh_degree);
          if (M == 16) {
            float out[48];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
              const float bk = (k < ncoef) ? basis[k] : 0.f;
              out[3 * k] = bk * grgb[0]; out[3 * k + 1] = bk * grgb[1]; out[3 * k + 2] = bk * grgb[2];
            }
            if (SPLIT) {
              out_dc[0] = out[0]; out_dc[1] = out[1]; out_dc[2] = out[2];
#pragma unroll
              for (int k = 0; k < SPLIT_REST; ++k) out_row[k] = out[3 + k];
            } else {
              float4* dst = reinterpret_cast<float4*>(out_row);
#pragma unroll
              for (int k = 0; k < 12; ++k) dst[k] = make_float4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
            }
          } else {
            for (int k = 0; k < M; ++k) {
              const float bk = (k < ncoef) ? basis[k] : 0.f;
              float* o3 = SPLIT ? (k == 0 ? out_dc : out_row + 3 * (k - 1)) : out_row + 3 * k;
              o3[0] = bk * grgb[0]; o3[1] = bk * grgb[1]; o3[2] = bk * grgb[2];
            }
          }
        }
      }
      // covariance
      if (a.cov3D_precomp) {
        if (gp.cov3D_precomp) {
#pragma unroll
          for (int k = 0; k < 6; ++k) {
            if (gp.accumulate) red_add(gp.cov3D_precomp + 6 * idx + k, g_cov[k]); else gp.cov3D_precomp[6 * idx + k] = g_cov[k];
          }
        }
      } else {
        float g_scale[3], g_rot[4];
        cov3d_backward(sc, cam.scale_modifier, q, g_cov, g_scale, g_rot);
        if (a.raw) raw_scale_rot_backward(sc, q, qinv, g_scale, g_rot);
        if (gp.scales) {
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            if (gp.accumulate) red_add(gp.scales + 3 * idx + k, g_scale[k]); else gp.scales[3 * idx + k] = g_scale[k];
          }
        }
        if (gp.rotations) {
          if (gp.accumulate) red_add4(gp.rotations + 4 * idx, g_rot[0], g_rot[1], g_rot[2], g_rot[3]);
          else reinterpret_cast<float4*>(gp.rotations)[idx] = make_float4(g_rot[0], g_rot[1], g_rot[2], g_rot[3]);
        }
      }
      if (gp.means3D) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (gp.accumulate) red_add(gp.means3D + 3 * idx + k, g_mean[k]); else gp.means3D[3 * idx + k] = g_mean[k];
        }
      }
      if (gp.means2D) {
        if (gp.accumulate) { red_add(gp.means2D + 3 * idx, in.g_mean2d_x); red_add(gp.means2D + 3 * idx + 1, in.g_mean2d_y); }
        else { gp.means2D[3 * idx] = in.g_mean2d_x; gp.means2D[3 * idx + 1] = in.g_mean2d_y; gp.means2D[3 * idx + 2] = 0.f; }
      }
      if (gp.opacities) {
        if (gp.accumulate) red_add(gp.opacities + idx, g_op); else gp.opacities[idx] = g_op;
      }
      // densification statistics of this view (scene/gaussian_model.py:502-506, train.py:313-314)
      if (gp.densify_grad_accum)
        red_add(gp.densify_grad_accum + idx, sqrtf(in.g_mean2d_x * in.g_mean2d_x + in.g_mean2d_y * in.g_mean2d_y));
      if (gp.densify_denom) red_add(gp.densify_denom + idx, 1.0f);
      if (gp.max_radii2D) atomicMax(gp.max_radii2D + idx, radius);
    }
    if (want_sh) {
      fence_async_smem();  // generic-proxy writes of sh_out -> visible to the bulk store
      __syncthreads();
      if (SPLIT) {
        float* dst_rest = gp.shs_rest + (size_t)t0 * rrow;
        float* dst_dc = gp.shs_dc + (size_t)t0 * 3;
        if (bulk) {
          if (tid == 0) {
            const uint32_t rest_bytes = (uint32_t)(PB_TILE * rrow * sizeof(float)), dc_bytes = (uint32_t)(PB_TILE * 3 * sizeof(float));
            if (gp.accumulate) {
              if (rrow > 0) bulk_s2g_add_f32(dst_rest, sm.sh_out, rest_bytes);
              bulk_s2g_add_f32(dst_dc, dc_out, dc_bytes);
            } else {
              if (rrow > 0) bulk_s2g(dst_rest, sm.sh_out, rest_bytes);
              bulk_s2g(dst_dc, dc_out, dc_bytes);
            }
            bulk_commit();
          }
        } else {
          for (int i = tid; i < n * rrow; i += PB_TILE) {
            if (gp.accumulate) red_add(dst_rest + i, sm.sh_out[i]); else dst_rest[i] = sm.sh_out[i];
          }
          for (int i = tid; i < n * 3; i += PB_TILE) {
            if (gp.accumulate) red_add(dst_dc + i, dc_out[i]); else dst_dc[i] = dc_out[i];
          }
        }
      } else {
      float* dst = gp.shs + (size_t)t0 * row;
      if (bulk) {
        if (tid == 0) {
          const uint32_t sh_bytes = (uint32_t)(PB_TILE * row * sizeof(float));
          if (gp.accumulate) bulk_s2g_add_f32(dst, sm.sh_out, sh_bytes); else bulk_s2g(dst, sm.sh_out, sh_bytes);
          bulk_commit();
        }
      } else {
        for (int i = tid; i < n * row; i += PB_TILE) {
          if (gp.accumulate) red_add(dst + i, sm.sh_out[i]); else dst[i] = sm.sh_out[i];
        }
      }
      }
    } else {
      __syncthreads();  // sh_in / acc_in are free for the next tile's copies
    }
  }
  if (!DEFER && tid == 0) bulk_wait0();  // all bulk stores complete before the CTA's shared memory goes away
}

// ------------------------------------------------------------------------------------------
// SH backward of a whole batch of views (deferred from K9, see PdgsGrads.color_grad_view).
// One thread per Gaussian: the 3M SH coefficients are read ONCE and their gradient written ONCE per batch;
// per view only the 12-byte masked dL/drgb is read.  For each view that touched the Gaussian the thread
// rebuilds the basis for that camera's view direction, adds basis_k * dL/drgb_c to its 3M register
// accumulators and the direction term to dL/dmean.
// ------------------------------------------------------------------------------------------
constexpr int SHV_THREADS = 128;
constexpr int SHV_MAX_VIEWS = 256;
constexpr int SHV_VB = 8;          // views whose colour gradients are in flight per thread  // camera positions staged per launch (longer batches loop on the host)

template <int MT>  // MT = 16: fully unrolled register path; MT = 0: any M through global memory
__global__ void __launch_bounds__(SHV_THREADS) sh_backward_views_kernel(const float* __restrict__ means3D,
                                                                        const float* __restrict__ shs, int64_t P,
                                                                        int64_t cg_stride, int M, int sh_degree,
                                                                        const float* __restrict__ campos_views,
                                                                        const float* __restrict__ color_grads,
                                                                        int n_views, float* __restrict__ grad_shs,
                                                                        float* __restrict__ grad_means3D, int accumulate) {
  __shared__ float s_cam[SHV_MAX_VIEWS * 3];
  for (int i = threadIdx.x; i < n_views * 3; i += SHV_THREADS) s_cam[i] = __ldg(campos_views + i);
  __syncthreads();
  const int64_t idx = (int64_t)blockIdx.x * SHV_THREADS + threadIdx.x;
  if (idx >= P) return;
  const float mx = __ldg(means3D + 3 * idx), my = __ldg(means3D + 3 * idx + 1), mz = __ldg(means3D + 3 * idx + 2);
  const int ncoef = sh_num_coeffs(sh_degree);
  float gm[3] = {0.f, 0.f, 0.f};
  if (MT == 16) {
    ShRegs sh;
    const float4* src = reinterpret_cast<const float4*>(shs + (size_t)idx * 48);
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      const float4 v = __ldg(src + k);
      sh.v[4 * k] = v.x; sh.v[4 * k + 1] = v.y; sh.v[4 * k + 2] = v.z; sh.v[4 * k + 3] = v.w;
    }
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; ++k) acc[k] = 0.f;
    bool any = false;
    // the per-view colour gradients are fetched VB views ahead of their use: one dependent DRAM round trip per view
    // would otherwise serialise the loop (64 views = 64 round trips per thread)
    for (int v0 = 0; v0 < n_views; v0 += SHV_VB) {
      float gb[SHV_VB][3];
#pragma unroll
      for (int u = 0; u < SHV_VB; ++u) {
        const int v = (v0 + u < n_views) ? v0 + u : n_views - 1;
        const float* gsrc = color_grads + ((size_t)v * cg_stride + idx) * 3;
        gb[u][0] = __ldg(gsrc); gb[u][1] = __ldg(gsrc + 1); gb[u][2] = __ldg(gsrc + 2);
      }
#pragma unroll
      for (int u = 0; u < SHV_VB; ++u) {
        const int v = v0 + u;
        const float g[3] = {gb[u][0], gb[u][1], gb[u][2]};
        if (v >= n_views || (g[0] == 0.f && g[1] == 0.f && g[2] == 0.f)) continue;  // culled in this view / no gradient
        any = true;
        float gdir[3], basis[16];
        sh_backward_dir(sh_degree, mx - s_cam[3 * v], my - s_cam[3 * v + 1], mz - s_cam[3 * v + 2], sh, g, gdir, basis);
        gm[0] += gdir[0]; gm[1] += gdir[1]; gm[2] += gdir[2];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float bk = (k < ncoef) ? basis[k] : 0.f;
          acc[3 * k] = fmaf(bk, g[0], acc[3 * k]);
          acc[3 * k + 1] = fmaf(bk, g[1], acc[3 * k + 1]);
          acc[3 * k + 2] = fmaf(bk, g[2], acc[3 * k + 2]);
        }
      }
    }
    if (grad_shs && (any || !accumulate)) {
      float4* dst = reinterpret_cast<float4*>(grad_shs + (size_t)idx * 48);
#pragma unroll
      for (int k = 0; k < 12; ++k) {
        float4 o = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
        if (accumulate) { const float4 w = dst[k]; o.x += w.x; o.y += w.y; o.z += w.z; o.w += w.w; }
        dst[k] = o;
      }
    }
  } else {
    ShGlobal sh{shs + (size_t)idx * M * 3};
    float* dst = grad_shs ? grad_shs + (size_t)idx * M * 3 : nullptr;
    if (dst && !accumulate)
      for (int k = 0; k < 3 * M; ++k) dst[k] = 0.f;
    for (int v = 0; v < n_views; ++v) {
      const float* gsrc = color_grads + ((size_t)v * cg_stride + idx) * 3;
      const float g[3] = {__ldg(gsrc), __ldg(gsrc + 1), __ldg(gsrc + 2)};
      if (g[0] == 0.f && g[1] == 0.f && g[2] == 0.f) continue;
      float gdir[3], basis[16];
      sh_backward_dir(sh_degree, mx - s_cam[3 * v], my - s_cam[3 * v + 1], mz - s_cam[3 * v + 2], sh, g, gdir, basis);
      gm[0] += gdir[0]; gm[1] += gdir[1]; gm[2] += gdir[2];
      if (dst)
        for (int k = 0; k < ncoef; ++k) {
          dst[3 * k] += basis[k] * g[0]; dst[3 * k + 1] += basis[k] * g[1]; dst[3 * k + 2] += basis[k] * g[2];
        }
    }
  }
  if (grad_means3D && (gm[0] != 0.f || gm[1] != 0.f || gm[2] != 0.f)) {
    grad_means3D[3 * idx] += gm[0]; grad_means3D[3 * idx + 1] += gm[1]; grad_means3D[3 * idx + 2] += gm[2];
  }
}

// Split-SH variant (features_dc [P,3] + features_rest [P,3(M-1)]).  MT = 16: the CTA's 128 rest rows (one contiguous
// 23040-byte span) are staged in shared memory with coalesced loads, each thread works on its own 45-float row
// (odd stride, conflict-free), leaves the row's gradient in the same place and the CTA writes the span back coalesced.
template <int MT>
__global__ void __launch_bounds__(SHV_THREADS) sh_backward_views_split_kernel(
    const float* __restrict__ means3D, const float* __restrict__ shs_dc, const float* __restrict__ shs_rest, int64_t P,
    int64_t cg_stride, int M, int sh_degree, const float* __restrict__ campos_views, const float* __restrict__ color_grads, int n_views,
    float* __restrict__ grad_dc, float* __restrict__ grad_rest, float* __restrict__ grad_means3D, int accumulate) {
  __shared__ float s_cam[SHV_MAX_VIEWS * 3];
  __shared__ __align__(16) float s_rest[MT == 16 ? SHV_THREADS * SPLIT_REST : 4];
  for (int i = threadIdx.x; i < n_views * 3; i += SHV_THREADS) s_cam[i] = __ldg(campos_views + i);
  const int64_t row0 = (int64_t)blockIdx.x * SHV_THREADS;
  const int nrows = (int)((P - row0 < SHV_THREADS) ? (P - row0) : SHV_THREADS);
  if (MT == 16) {
    const float* src = shs_rest + row0 * SPLIT_REST;
    if (nrows == SHV_THREADS) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      float4* d4 = reinterpret_cast<float4*>(s_rest);
      for (int i = threadIdx.x; i < SHV_THREADS * SPLIT_REST / 4; i += SHV_THREADS) d4[i] = __ldg(s4 + i);
    } else {
      for (int i = threadIdx.x; i < nrows * SPLIT_REST; i += SHV_THREADS) s_rest[i] = __ldg(src + i);
    }
  }
  __syncthreads();
  const int64_t idx = row0 + threadIdx.x;
  const bool in_range = idx < P;
  const int ncoef = sh_num_coeffs(sh_degree);
  float gm[3] = {0.f, 0.f, 0.f};
  float mx = 0.f, my = 0.f, mz = 0.f;
  if (in_range) { mx = __ldg(means3D + 3 * idx); my = __ldg(means3D + 3 * idx + 1); mz = __ldg(means3D + 3 * idx + 2); }
  if (MT == 16) {
    float* row = s_rest + threadIdx.x * SPLIT_REST;
    if (in_range) {
      ShRegs sh;
      sh.v[0] = __ldg(shs_dc + 3 * idx); sh.v[1] = __ldg(shs_dc + 3 * idx + 1); sh.v[2] = __ldg(shs_dc + 3 * idx + 2);
#pragma unroll
      for (int k = 0; k < SPLIT_REST; ++k) sh.v[3 + k] = row[k];
      float acc[48];
#pragma unroll
      for (int k = 0; k < 48; ++k) acc[k] = 0.f;
      for (int v0 = 0; v0 < n_views; v0 += SHV_VB) {  // colour gradients fetched SHV_VB views ahead, as above
        float gb[SHV_VB][3];
#pragma unroll
        for (int u = 0; u < SHV_VB; ++u) {
          const int v = (v0 + u < n_views) ? v0 + u : n_views - 1;
          const float* gsrc = color_grads + ((size_t)v * cg_stride + idx) * 3;
          gb[u][0] = __ldg(gsrc); gb[u][1] = __ldg(gsrc + 1); gb[u][2] = __ldg(gsrc + 2);
        }
#pragma unroll
        for (int u = 0; u < SHV_VB; ++u) {
          const int v = v0 + u;
          const float g[3] = {gb[u][0], gb[u][1], gb[u][2]};
          if (v >= n_views || (g[0] == 0.f && g[1] == 0.f && g[2] == 0.f)) continue;
          float gdir[3], basis[16];
          sh_backward_dir(sh_degree, mx - s_cam[3 * v], my - s_cam[3 * v + 1], mz - s_cam[3 * v + 2], sh, g, gdir, basis);
          gm[0] += gdir[0]; gm[1] += gdir[1]; gm[2] += gdir[2];
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const float bk = (k < ncoef) ? basis[k] : 0.f;
            acc[3 * k] = fmaf(bk, g[0], acc[3 * k]);
            acc[3 * k + 1] = fmaf(bk, g[1], acc[3 * k + 1]);
            acc[3 * k + 2] = fmaf(bk, g[2], acc[3 * k + 2]);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < SPLIT_REST; ++k) row[k] = acc[3 + k];
      if (grad_dc) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          if (accumulate) grad_dc[3 * idx + c] += acc[c]; else grad_dc[3 * idx + c] = acc[c];
        }
      }
    }
    __syncthreads();
    if (grad_rest) {
      float* dst = grad_rest + row0 * SPLIT_REST;
      if (nrows == SHV_THREADS) {
        float4* d4 = reinterpret_cast<float4*>(dst);
        const float4* s4 = reinterpret_cast<const float4*>(s_rest);
        for (int i = threadIdx.x; i < SHV_THREADS * SPLIT_REST / 4; i += SHV_THREADS) {
          float4 o = s4[i];
          if (accumulate) { const float4 w = d4[i]; o.x += w.x; o.y += w.y; o.z += w.z; o.w += w.w; }
          d4[i] = o;
        }
      } else {
        for (int i = threadIdx.x; i < nrows * SPLIT_REST; i += SHV_THREADS) {
          if (accumulate) dst[i] += s_rest[i]; else dst[i] = s_rest[i];
        }
      }
    }
  } else if (in_range) {
    const int rrow = 3 * (M - 1);
    ShSplitGlobal sh{shs_dc + 3 * idx, shs_rest + (size_t)idx * rrow};
    float* ddc = grad_dc ? grad_dc + 3 * idx : nullptr;
    float* drest = grad_rest ? grad_rest + (size_t)idx * rrow : nullptr;
    if (!accumulate) {
      if (ddc) { ddc[0] = 0.f; ddc[1] = 0.f; ddc[2] = 0.f; }
      if (drest) for (int k = 0; k < rrow; ++k) drest[k] = 0.f;
    }
    for (int v = 0; v < n_views; ++v) {
      const float* gsrc = color_grads + ((size_t)v * cg_stride + idx) * 3;
      const float g[3] = {__ldg(gsrc), __ldg(gsrc + 1), __ldg(gsrc + 2)};
      if (g[0] == 0.f && g[1] == 0.f && g[2] == 0.f) continue;
      float gdir[3], basis[16];
      sh_backward_dir(sh_degree, mx - s_cam[3 * v], my - s_cam[3 * v + 1], mz - s_cam[3 * v + 2], sh, g, gdir, basis);
      gm[0] += gdir[0]; gm[1] += gdir[1]; gm[2] += gdir[2];
      if (ddc) { ddc[0] += basis[0] * g[0]; ddc[1] += basis[0] * g[1]; ddc[2] += basis[0] * g[2]; }
      if (drest)
        for (int k = 1; k < ncoef; ++k) {
          drest[3 * (k - 1)] += basis[k] * g[0]; drest[3 * (k - 1) + 1] += basis[k] * g[1]; drest[3 * (k - 1) + 2] += basis[k] * g[2];
        }
    }
  }
  if (in_range && grad_means3D && (gm[0] != 0.f || gm[1] != 0.f || gm[2] != 0.f)) {
    grad_means3D[3 * idx] += gm[0]; grad_means3D[3 * idx + 1] += gm[1]; grad_means3D[3 * idx + 2] += gm[2];
  }
}

__global__ void mark_visible_kernel(const float* __restrict__ means3D, const float* __restrict__ viewmatrix, int64_t P,
                                    uint8_t* __restrict__ present) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float vm[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) vm[k] = __ldg(viewmatrix + k);
  const float z = affine_row(vm, 2, __ldg(means3D + 3 * idx), __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2));
  present[idx] = (z > NEAR_Z) ? 1 : 0;
}

}  // namespace

cudaError_t launch_preprocess(const DeviceArgs& a, GeomView g, int32_t* radii, cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((a.P + PRE_THREADS - 1) / PRE_THREADS);
  if (a.shs_dc)
    preprocess_kernel<true><<<blocks, PRE_THREADS, a.M == 16 ? PRE_SPLIT_SMEM : 0, s>>>(a, g, radii);
  else
    preprocess_kernel<false><<<blocks, PRE_THREADS, 0, s>>>(a, g, radii);
  return cudaGetLastError();
}

cudaError_t launch_preprocess_backward(const DeviceArgs& a, GeomView g, const int32_t* radii, const float* acc,
                                       GradPtrs gp, cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  if (a.shs || a.shs_dc) {  // SH path: bulk-copy pipeline
    static bool configured[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || !configured[dev]) {
      e = cudaFuncSetAttribute(preprocess_backward_sh_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)sizeof(PbShared));
      if (e != cudaSuccess) return e;
      e = cudaFuncSetAttribute(preprocess_backward_sh_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)sizeof(PbShared));
      if (e != cudaSuccess) return e;
      if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const auto al16 = [](const void* p) { return (((uintptr_t)p) & 15u) == 0; };
    const int64_t ntiles = (a.P + PB_TILE - 1) / PB_TILE;
    if (gp.color_grad_view) {  // SH backward deferred to launch_sh_backward_views: only the accumulator block is staged
      const int bulk_ok = al16(acc);
      const unsigned blocks = (unsigned)(ntiles < PDGS_PB_DEFER_CTAS * sm_count() ? ntiles : PDGS_PB_DEFER_CTAS * sm_count());
      preprocess_backward_sh_kernel<true, false><<<blocks, PB_TILE, offsetof(PbShared, sh_in), s>>>(a, g, radii, acc, gp, bulk_ok);
      return cudaGetLastError();
    }
    const unsigned blocks = (unsigned)(ntiles < 2 * sm_count() ? ntiles : 2 * sm_count());  // persistent: two CTAs per SM
    if (a.shs_dc) {
      if (gp.shs_dc && a.M > 1 && !gp.shs_rest) return cudaErrorInvalidValue;
      const int bulk_ok = al16(a.shs_dc) && (!a.shs_rest || al16(a.shs_rest)) && al16(acc) && (!gp.shs_dc || al16(gp.shs_dc)) &&
                          (!gp.shs_rest || al16(gp.shs_rest));
      preprocess_backward_sh_kernel<false, true><<<blocks, PB_TILE, sizeof(PbShared), s>>>(a, g, radii, acc, gp, bulk_ok);
      return cudaGetLastError();
    }
    const int bulk_ok = al16(a.shs) && al16(acc) && (!gp.shs || al16(gp.shs));
    preprocess_backward_sh_kernel<false, false><<<blocks, PB_TILE, sizeof(PbShared), s>>>(a, g, radii, acc, gp, bulk_ok);
    return cudaGetLastError();
  }
  const unsigned blocks = (unsigned)((a.P + PRE_THREADS - 1) / PRE_THREADS);
  preprocess_backward_kernel<<<blocks, PRE_THREADS, 0, s>>>(a, g, radii, acc, gp);
  return cudaGetLastError();
}

// Gaussians [first, first + count) of a scene of P Gaussians (count <= 0: all of them).  color_grads is [n_views][P][3];
// every other array is indexed by Gaussian, so a range is the same kernel on offset pointers with the view stride P.
cudaError_t launch_sh_backward_views(const float* means3D, const float* shs, const float* shs_dc, const float* shs_rest,
                                     int64_t P, int M, int sh_degree, const float* campos_views, const float* color_grads,
                                     int n_views, float* grad_shs, float* grad_dc, float* grad_rest, float* grad_means3D,
                                     int accumulate, int64_t first, int64_t count, cudaStream_t s) {
  if (P <= 0 || n_views <= 0) return cudaSuccess;
  if (count <= 0) { first = 0; count = P; }
  if (first < 0 || first + count > P) return cudaErrorInvalidValue;
  const int rest = 3 * (M - 1);
  means3D += 3 * first;
  if (shs) shs += (size_t)first * 3 * M;
  if (shs_dc) shs_dc += 3 * first;
  if (shs_rest) shs_rest += (size_t)first * rest;
  if (grad_shs) grad_shs += (size_t)first * 3 * M;
  if (grad_dc) grad_dc += 3 * first;
  if (grad_rest) grad_rest += (size_t)first * rest;
  if (grad_means3D) grad_means3D += 3 * first;
  color_grads += 3 * first;
  const unsigned blocks = (unsigned)((count + SHV_THREADS - 1) / SHV_THREADS);
  for (int v0 = 0; v0 < n_views; v0 += SHV_MAX_VIEWS) {  // later chunks always add to what the first one wrote
    const int nv = (n_views - v0 < SHV_MAX_VIEWS) ? n_views - v0 : SHV_MAX_VIEWS;
    const int acc = (v0 == 0) ? accumulate : 1;
    const float* cg = color_grads + (size_t)v0 * P * 3;
    if (!shs && M == 16)
      sh_backward_views_split_kernel<16><<<blocks, SHV_THREADS, 0, s>>>(means3D, shs_dc, shs_rest, count, P, M, sh_degree,
                                                                        campos_views + 3 * v0, cg, nv, grad_dc, grad_rest,
                                                                        grad_means3D, acc);
    else if (!shs)
      sh_backward_views_split_kernel<0><<<blocks, SHV_THREADS, 0, s>>>(means3D, shs_dc, shs_rest, count, P, M, sh_degree,
                                                                       campos_views + 3 * v0, cg, nv, grad_dc, grad_rest,
                                                                       grad_means3D, acc);
    else if (M == 16)
      sh_backward_views_kernel<16><<<blocks, SHV_THREADS, 0, s>>>(means3D, shs, count, P, M, sh_degree, campos_views + 3 * v0,
                                                                  cg, nv, grad_shs, grad_means3D, acc);
    else
      sh_backward_views_kernel<0><<<blocks, SHV_THREADS, 0, s>>>(means3D, shs, count, P, M, sh_degree, campos_views + 3 * v0,
                                                                 cg, nv, grad_shs, grad_means3D, acc);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_mark_visible(const float* means3D, const float* viewmatrix, int64_t P, uint8_t* present,
                                cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((P + 255) / 256);
  mark_visible_kernel<<<blocks, 256, 0, s>>>(means3D, viewmatrix, P, present);
  return cudaGetLastError();
}

}  // namespace pdgs
