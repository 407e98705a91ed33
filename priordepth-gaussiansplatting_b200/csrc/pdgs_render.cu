// K6 forward alpha-blend (colour + inverse depth) and K7 backward blend.
// Replaces renderCUDA (forward/backward) of the reference rasterizer (source absent;
// SURVEY.md 2.3(b) K6/K7, A.7, A.8).
//
// One CTA per 16x16 tile, 8 warps, each warp owns an 8x4 pixel patch so that warps
// terminate early and skip non-overlapping Gaussians independently.  Batches of 256
// Gaussians are staged in shared memory as three float4 records.  The per-pair alpha is
// computed by ONE shared function in a fixed operation order so that forward and
// backward take bit-identical accept/reject decisions.
#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

constexpr int RT = 256;  // threads per tile CTA
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// power = -0.5*(A dx^2 + C dy^2) - B dx dy, fixed order (hA = -0.5A, hC = -0.5C, nB = -B are exact)
__device__ __forceinline__ float gauss_power(float dx, float dy, float hA, float nB, float hC) {
  const float u = __fmaf_rn(hA, dx, __fmul_rn(nB, dy));
  const float v = __fmul_rn(__fmul_rn(hC, dy), dy);
  return __fmaf_rn(dx, u, v);
}
__device__ __forceinline__ float gauss_G(float power) { return ex2_approx(__fmul_rn(power, LOG2E)); }

__device__ __forceinline__ void pixel_of_thread(int& lx, int& ly) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  lx = (warp & 1) * 8 + (lane & 7);
  ly = (warp >> 1) * 4 + (lane >> 3);
}

__global__ void __launch_bounds__(RT) render_forward_kernel(const float4* __restrict__ gA, const float4* __restrict__ gB,
                                                            const float4* __restrict__ gC,
                                                            const uint2* __restrict__ ranges,
                                                            const uint32_t* __restrict__ point_list, int W, int H,
                                                            const float* __restrict__ bg, float* __restrict__ out_color,
                                                            float* __restrict__ out_invdepth, float* __restrict__ final_T,
                                                            uint32_t* __restrict__ n_contrib,
                                                            uint32_t* __restrict__ n_tested,
                                                            uint32_t* __restrict__ n_accepted) {
  __shared__ float4 sA[RT];  // x, y, -0.5*conA, -conB
  __shared__ float4 sB[RT];  // -0.5*conC, opacity, 1/depth, -
  __shared__ float4 sC[RT];  // r, g, b
  int lx, ly;
  pixel_of_thread(lx, ly);
  const int px = blockIdx.x * TILE + lx, py = blockIdx.y * TILE + ly;
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t contributor = 0, last = 0, accepted = 0;
  bool done = !inside;
  for (uint32_t base = range.x; base < range.y; base += RT) {
    if (__syncthreads_count(done) == RT) break;
    const uint32_t i = base + threadIdx.x;
    if (i < range.y) {
      const uint32_t id = point_list[i];
      const float4 a = __ldg(gA + id), b = __ldg(gB + id);
      sA[threadIdx.x] = make_float4(a.x, a.y, -0.5f * a.z, -a.w);
      sB[threadIdx.x] = make_float4(-0.5f * b.x, b.y, b.z, 0.f);
      sC[threadIdx.x] = __ldg(gC + id);
    }
    __syncthreads();
    const int cnt = min((uint32_t)RT, range.y - base);
    for (int j = 0; !done && j < cnt; ++j) {
      ++contributor;
      const float4 a = sA[j];
      const float4 b = sB[j];
      const float dx = a.x - pxf, dy = a.y - pyf;
      const float power = gauss_power(dx, dy, a.z, a.w, b.x);
      if (power > 0.0f) continue;
      const float alpha = fminf(ALPHA_MAX, __fmul_rn(b.y, gauss_G(power)));
      if (alpha < ALPHA_MIN) continue;
      const float test_T = T * (1.0f - alpha);
      if (test_T < T_EPS) {
        done = true;
        continue;
      }
      const float4 c = sC[j];
      const float w = alpha * T;
      C0 = fmaf(c.x, w, C0);
      C1 = fmaf(c.y, w, C1);
      C2 = fmaf(c.z, w, C2);
      D = fmaf(b.z, w, D);
      T = test_T;
      last = contributor;
      ++accepted;
    }
  }
  if (inside) {
    const size_t pix = (size_t)py * W + px, hw = (size_t)W * H;
    final_T[pix] = T;
    n_contrib[pix] = last;
    n_tested[pix] = contributor;
    n_accepted[pix] = accepted;
    out_color[pix] = C0 + T * __ldg(bg + 0);
    out_color[hw + pix] = C1 + T * __ldg(bg + 1);
    out_color[2 * hw + pix] = C2 + T * __ldg(bg + 2);
    out_invdepth[pix] = D;
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(RT) render_backward_kernel(const float4* __restrict__ gA, const float4* __restrict__ gB,
                                                             const float4* __restrict__ gC,
                                                             const uint2* __restrict__ ranges,
                                                             const uint32_t* __restrict__ point_list, int W, int H,
                                                             const float* __restrict__ bg,
                                                             const float* __restrict__ final_T,
                                                             const uint32_t* __restrict__ n_contrib,
                                                             const float* __restrict__ dL_dcolor,
                                                             const float* __restrict__ dL_dinvdepth,
                                                             float* __restrict__ acc) {
  __shared__ float4 sA[RT];
  __shared__ float4 sB[RT];
  __shared__ float4 sC[RT];
  __shared__ uint32_t sId[RT];
  __shared__ uint32_t s_max[RT / 32];
  int lx, ly;
  pixel_of_thread(lx, ly);
  const int px = blockIdx.x * TILE + lx, py = blockIdx.y * TILE + ly;
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const size_t pix = (size_t)py * W + px, hw = (size_t)W * H;
  const float T_final = inside ? final_T[pix] : 0.f;
  const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;
  float gp0 = 0.f, gp1 = 0.f, gp2 = 0.f, gpd = 0.f;
  if (inside) {
    gp0 = dL_dcolor[pix];
    gp1 = dL_dcolor[hw + pix];
    gp2 = dL_dcolor[2 * hw + pix];
    if (dL_dinvdepth) gpd = dL_dinvdepth[pix];
  }
  const float bg_dot = __ldg(bg + 0) * gp0 + __ldg(bg + 1) * gp1 + __ldg(bg + 2) * gp2;
  // the tile only needs list positions below the largest n_contrib of its pixels
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t wmax = __reduce_max_sync(0xffffffffu, last_contributor);
  if (lane == 0) s_max[warp] = wmax;
  __syncthreads();
  uint32_t todo = 0;
#pragma unroll
  for (int w = 0; w < RT / 32; ++w) todo = max(todo, s_max[w]);
  todo = min(todo, range.y - range.x);

  float T = T_final;
  float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, ard = 0.f;          // accum_rec
  float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, ld = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * (float)W, ddely_dy = 0.5f * (float)H;

  for (uint32_t done = 0; done < todo; done += RT) {
    __syncthreads();
    const uint32_t progress = done + threadIdx.x;
    if (progress < todo) {
      const uint32_t id = point_list[range.x + (todo - 1 - progress)];
      const float4 a = __ldg(gA + id), b = __ldg(gB + id);
      sId[threadIdx.x] = id;
      sA[threadIdx.x] = make_float4(a.x, a.y, -0.5f * a.z, -a.w);
      sB[threadIdx.x] = make_float4(-0.5f * b.x, b.y, b.z, 0.f);
      sC[threadIdx.x] = __ldg(gC + id);
    }
    __syncthreads();
    const int cnt = min((uint32_t)RT, todo - done);
    for (int j = 0; j < cnt; ++j) {
      const uint32_t pos = todo - 1 - (done + j);  // position of this Gaussian in the tile list
      const float4 a = sA[j];
      const float4 b = sB[j];
      const float dx = a.x - pxf, dy = a.y - pyf;
      const float power = gauss_power(dx, dy, a.z, a.w, b.x);
      const float G = gauss_G(power);
      const float alpha = fminf(ALPHA_MAX, __fmul_rn(b.y, G));
      const bool active = (pos < last_contributor) && !(power > 0.0f) && !(alpha < ALPHA_MIN);
      if (!__any_sync(0xffffffffu, active)) continue;
      float g_mx = 0.f, g_my = 0.f, g_cx = 0.f, g_cy = 0.f, g_cw = 0.f, g_op = 0.f, g_id = 0.f, g_r = 0.f, g_g = 0.f,
            g_b = 0.f;
      if (active) {
        const float4 c = sC[j];
        T = __fdividef(T, 1.0f - alpha);
        const float w = alpha * T;
        float dL_dalpha;
        ar0 = last_alpha * lc0 + (1.f - last_alpha) * ar0; lc0 = c.x;
        ar1 = last_alpha * lc1 + (1.f - last_alpha) * ar1; lc1 = c.y;
        ar2 = last_alpha * lc2 + (1.f - last_alpha) * ar2; lc2 = c.z;
        ard = last_alpha * ld + (1.f - last_alpha) * ard; ld = b.z;
        dL_dalpha = (c.x - ar0) * gp0 + (c.y - ar1) * gp1 + (c.z - ar2) * gp2 + (b.z - ard) * gpd;
        g_r = w * gp0; g_g = w * gp1; g_b = w * gp2; g_id = w * gpd;
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
        const float dL_dG = b.y * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        // conic = (A, B, C) = (-2*a.z, -a.w, -2*b.x)
        const float A = -2.f * a.z, B = -a.w, Cc = -2.f * b.x;
        const float dG_ddelx = -gdx * A - gdy * B;
        const float dG_ddely = -gdy * Cc - gdx * B;
        g_mx = dL_dG * dG_ddelx * ddelx_dx;
        g_my = dL_dG * dG_ddely * ddely_dy;
        g_cx = -0.5f * gdx * dx * dL_dG;
        g_cy = -0.5f * gdx * dy * dL_dG;
        g_cw = -0.5f * gdy * dy * dL_dG;
        g_op = G * dL_dalpha;
      }
      g_mx = warp_sum(g_mx); g_my = warp_sum(g_my); g_cx = warp_sum(g_cx); g_cy = warp_sum(g_cy);
      g_cw = warp_sum(g_cw); g_op = warp_sum(g_op); g_id = warp_sum(g_id);
      g_r = warp_sum(g_r); g_g = warp_sum(g_g); g_b = warp_sum(g_b);
      if (lane == 0) {
        float* dst = acc + (size_t)sId[j] * ACC_FLOATS;
        red_add_v4(dst, g_mx, g_my, g_cx, g_cy);
        red_add_v4(dst + 4, g_cw, g_op, g_id, 0.f);
        red_add_v4(dst + 8, g_r, g_g, g_b, 0.f);
      }
    }
  }
}

}  // namespace

cudaError_t launch_render_forward(const DeviceArgs& a, GeomView g, ImageView im, const uint32_t* point_list,
                                  int64_t capacity, float* out_color, float* out_invdepth, cudaStream_t s) {
  (void)capacity;
  if (a.W <= 0 || a.H <= 0) return cudaSuccess;
  dim3 grid(a.grid_x, a.grid_y);
  render_forward_kernel<<<grid, RT, 0, s>>>(g.gA, g.gB, g.gC, im.ranges, point_list, a.W, a.H, a.bg, out_color,
                                            out_invdepth, im.final_T, im.n_contrib, im.n_tested, im.n_accepted);
  return cudaGetLastError();
}

cudaError_t launch_render_backward(const DeviceArgs& a, GeomView g, ImageView im, const uint32_t* point_list,
                                   const float* dL_dcolor, const float* dL_dinvdepth, float* acc, cudaStream_t s) {
  if (a.W <= 0 || a.H <= 0) return cudaSuccess;
  dim3 grid(a.grid_x, a.grid_y);
  render_backward_kernel<<<grid, RT, 0, s>>>(g.gA, g.gB, g.gC, im.ranges, point_list, a.W, a.H, a.bg, im.final_T,
                                             im.n_contrib, dL_dcolor, dL_dinvdepth, acc);
  return cudaGetLastError();
}

}  // namespace pdgs
