// Fused SSIM (11x11 Gaussian window, sigma 1.5, zero "same" padding) forward + backward, and the
// training loss built on it.  Replaces, behind the same map-level interface, the reference's
// fused-ssim module:
//   fusedssimCUDA           /root/reference/submodules/fused-ssim/ssim.cu:187-289
//   fusedssim_backwardCUDA  /root/reference/submodules/fused-ssim/ssim.cu:291-366
//   python wrapper          /root/reference/submodules/fused-ssim/fused_ssim/__init__.py:8-47
// whose CPU-runnable statement of the same maths is utils/loss_utils.py:64-89 (`_ssim`).
//
// B200 design: one CTA per 32x32 output tile of one plane, 256 threads.  Both separable passes are
// register-blocked (4 outputs per thread from a 14-value window) so that shared memory serves
// 128-bit row reads in the horizontal pass and conflict-free column reads in the vertical pass; the
// five moments (forward) / three weighted partial maps (backward) are convolved together from one
// staged halo instead of one pass per quantity, and the loss sums (sum of the SSIM map, sum |a-b|)
// are reduced in the same kernel, so the whole loss + cotangent costs two launches and 44 bytes per
// pixel-channel of HBM traffic.
#include <initializer_list>

#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

constexpr int ST = 32;        // tile edge (outputs)
constexpr int SHALO = 5;      // window radius
constexpr int SROWS = ST + 2 * SHALO;  // 42 staged rows
constexpr int SLEFT = 8;      // staged columns start at x = tx0 - 8 so that global rows are read as aligned float4
constexpr int SPITCH = 48;    // staged columns [tx0 - 8, tx0 + 40): 12 float4 per row
constexpr int SOFF = SLEFT - SHALO;  // window of output column x starts at staged column x + SOFF
constexpr int STHREADS = 256;

// normalised 1-D window exp(-(k-5)^2 / (2 * 1.5^2)), k = 0..10, rounded to fp32 (utils/loss_utils.py:46-48)
#define PDGS_WIN(k)                                                                                          \
  ((k) == 0 || (k) == 10 ? 0.001028380123898387f                                                             \
   : (k) == 1 || (k) == 9 ? 0.0075987582094967365f                                                           \
   : (k) == 2 || (k) == 8 ? 0.036000773310661316f                                                            \
   : (k) == 3 || (k) == 7 ? 0.10936068743467331f                                                             \
   : (k) == 4 || (k) == 6 ? 0.21300552785396576f                                                             \
                          : 0.26601171493530273f)

__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float block_sum256(float v, float* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = (lane < STHREADS / 32) ? s_red[lane] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;  // valid in thread 0
}

// Stage NQ planes' halo (SROWS x SPITCH, zero outside the image) around the tile at (tx0, ty0): dst[q][r][c] =
// src[q] * (mul ? mul : mul_uniform).  vec: the image rows can be read as aligned float4 (W % 4 == 0 and 16-byte
// aligned planes); all loads of a thread are issued before the first store.
// FIRST (loss path, src[0] = the rendered image): the multiplier applies to plane 0 only, after an optional clamp to
// [0, 1] -- the caller-side `rendered_image.clamp(0, 1)` and `image *= alpha_mask` of the reference
// (gaussian_renderer/__init__.py:121, train.py:148-150).
__device__ __forceinline__ float clamp01f(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

template <int NQ, bool FIRST = false>
__device__ __forceinline__ void stage_halo(float (*dst)[SROWS][SPITCH], const float* const (&src)[NQ],
                                           const float* __restrict__ mul, float mul_uniform, int W, int H, int tx0,
                                           int ty0, bool vec, bool clamp0 = false) {
  constexpr int ITEMS = SROWS * (SPITCH / 4);  // 504 float4 per plane
  constexpr int ROUNDS = (ITEMS + STHREADS - 1) / STHREADS;
  float4 v[ROUNDS][NQ];
#pragma unroll
  for (int rd = 0; rd < ROUNDS; ++rd) {
    const int it = threadIdx.x + rd * STHREADS;
    const int r = it / (SPITCH / 4), c4 = it - r * (SPITCH / 4);
    const int y = ty0 + r - SHALO, x0 = tx0 - SLEFT + 4 * c4;
    const bool row_ok = it < ITEMS && y >= 0 && y < H;
    float4 m = make_float4(mul_uniform, mul_uniform, mul_uniform, mul_uniform);
    if (vec) {
      const bool ok = row_ok && x0 >= 0 && x0 < W;  // W % 4 == 0: a float4 is entirely inside or outside
      const size_t p = (size_t)(ok ? y : 0) * W + (ok ? x0 : 0);
      if (mul && ok) m = __ldg(reinterpret_cast<const float4*>(mul + p));
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) t = __ldg(reinterpret_cast<const float4*>(src[q] + p));
        if (FIRST && q == 0 && clamp0) t = make_float4(clamp01f(t.x), clamp01f(t.y), clamp01f(t.z), clamp01f(t.w));
        if (FIRST && q != 0) v[rd][q] = t;
        else v[rd][q] = make_float4(t.x * m.x, t.y * m.y, t.z * m.z, t.w * m.w);
      }
    } else {
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        float e[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int x = x0 + i;
          e[i] = 0.f;
          if (row_ok && x >= 0 && x < W) {
            const size_t p = (size_t)y * W + x;
            float t = __ldg(src[q] + p);
            if (FIRST && q == 0 && clamp0) t = clamp01f(t);
            e[i] = (FIRST && q != 0) ? t : t * (mul ? __ldg(mul + p) : mul_uniform);
          }
        }
        v[rd][q] = make_float4(e[0], e[1], e[2], e[3]);
      }
    }
  }
#pragma unroll
  for (int rd = 0; rd < ROUNDS; ++rd) {
    const int it = threadIdx.x + rd * STHREADS;
    if (it < ITEMS) {
      const int r = it / (SPITCH / 4), c4 = it - r * (SPITCH / 4);
#pragma unroll
      for (int q = 0; q < NQ; ++q) *reinterpret_cast<float4*>(&dst[q][r][4 * c4]) = v[rd][q];
    }
  }
}

// 20 consecutive staged values starting at a 16-byte aligned column (the 4 outputs of a group need 17 of them)
__device__ __forceinline__ void load_row20(const float* row, float (&v)[20]) {
  const float4* r4 = reinterpret_cast<const float4*>(row);
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const float4 t = r4[k];
    v[4 * k] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
  }
}

// 11-tap filter of 4 adjacent outputs from a 14-value window starting at v[o]
template <int N>
__device__ __forceinline__ void taps4(const float (&v)[N], int o, float (&out)[4]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) s = fmaf(PDGS_WIN(k), v[o + j + k], s);
    out[j] = s;
  }
}

struct SsimFwdSmem {
  float in[2][SROWS][SPITCH];  // img1, img2 halo
  float h[5][SROWS][ST];       // horizontally filtered a, b, a^2, b^2, ab
  float red[STHREADS / 32];
};

// sums: device double[2] = (sum of the SSIM map, sum |a - b|), or NULL
__global__ void __launch_bounds__(STHREADS) ssim_forward_kernel(float C1, float C2, const float* __restrict__ img1,
                                                                const float* __restrict__ img2, int W, int H, int vec,
                                                                float* __restrict__ ssim_map,
                                                                float* __restrict__ dm_dmu1,
                                                                float* __restrict__ dm_dsigma1_sq,
                                                                float* __restrict__ dm_dsigma12,
                                                                double* __restrict__ sums, int want_l1,
                                                                const float* __restrict__ alpha, int clamp01) {
  extern __shared__ __align__(16) unsigned char ssim_smem[];
  SsimFwdSmem& sm = *reinterpret_cast<SsimFwdSmem*>(ssim_smem);
  const size_t plane = (size_t)blockIdx.z * W * H;
  const int tx0 = blockIdx.x * ST, ty0 = blockIdx.y * ST;
  {
    const float* const src[2] = {img1 + plane, img2 + plane};
    // loss path: img1 is staged as clamp(img1, 0, 1) * alpha (one alpha plane for all colour planes)
    if (alpha || clamp01) stage_halo<2, true>(sm.in, src, alpha, 1.f, W, H, tx0, ty0, vec != 0, clamp01 != 0);
    else stage_halo<2>(sm.in, src, nullptr, 1.f, W, H, tx0, ty0, vec != 0);
  }
  __syncthreads();
  // horizontal pass: item = (row r, group of 4 output columns); the products are formed once per staged value
  for (int it = threadIdx.x; it < SROWS * (ST / 4); it += STHREADS) {
    const int r = it / (ST / 4), xg = it - r * (ST / 4);
    float va[20], vb[20], vp[20], o[4];
    load_row20(&sm.in[0][r][4 * xg], va);
    load_row20(&sm.in[1][r][4 * xg], vb);
    taps4(va, SOFF, o);
    *reinterpret_cast<float4*>(&sm.h[0][r][4 * xg]) = make_float4(o[0], o[1], o[2], o[3]);
    taps4(vb, SOFF, o);
    *reinterpret_cast<float4*>(&sm.h[1][r][4 * xg]) = make_float4(o[0], o[1], o[2], o[3]);
#pragma unroll
    for (int k = SOFF; k < SOFF + 14; ++k) vp[k] = va[k] * vb[k];
    taps4(vp, SOFF, o);
    *reinterpret_cast<float4*>(&sm.h[4][r][4 * xg]) = make_float4(o[0], o[1], o[2], o[3]);
#pragma unroll
    for (int k = SOFF; k < SOFF + 14; ++k) { va[k] *= va[k]; vb[k] *= vb[k]; }
    taps4(va, SOFF, o);
    *reinterpret_cast<float4*>(&sm.h[2][r][4 * xg]) = make_float4(o[0], o[1], o[2], o[3]);
    taps4(vb, SOFF, o);
    *reinterpret_cast<float4*>(&sm.h[3][r][4 * xg]) = make_float4(o[0], o[1], o[2], o[3]);
  }
  __syncthreads();
  // vertical pass: thread = (column x, block of 4 output rows)
  const int x = threadIdx.x & 31, yb = threadIdx.x >> 5;
  float acc[5][4];
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    float col[14];
#pragma unroll
    for (int k = 0; k < 14; ++k) col[k] = sm.h[q][4 * yb + k][x];
    taps4(col, 0, acc[q]);
  }
  float part_m = 0.f, part_l1 = 0.f;
  const int px = tx0 + x;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int py = ty0 + 4 * yb + j;
    if (px < W && py < H) {
      const float mu1 = acc[0][j], mu2 = acc[1][j];
      const float s1 = acc[2][j] - mu1 * mu1, s2 = acc[3][j] - mu2 * mu2, s12 = acc[4][j] - mu1 * mu2;
      const float Cc = 2.f * mu1 * mu2 + C1, D = 2.f * s12 + C2;
      const float A = mu1 * mu1 + mu2 * mu2 + C1, B = s1 + s2 + C2;
      const float iA = rcp_fast(A), iB = rcp_fast(B);
      const float iAB = iA * iB;
      const float m = Cc * D * iAB;
      const size_t p = plane + (size_t)py * W + px;
      if (ssim_map) ssim_map[p] = m;
      if (dm_dmu1) {
        // total derivative w.r.t. mu1 (sigma1_sq = E[a^2] - mu1^2 and sigma12 = E[ab] - mu1 mu2 depend on it)
        dm_dmu1[p] = 2.f * iAB * (mu2 * (D - Cc) + mu1 * Cc * D * (iB - iA));
        dm_dsigma1_sq[p] = -m * iB;
        dm_dsigma12[p] = 2.f * Cc * iAB;
      }
      part_m += m;
      if (want_l1)
        part_l1 += fabsf(sm.in[0][4 * yb + j + SHALO][x + SLEFT] - sm.in[1][4 * yb + j + SHALO][x + SLEFT]);
    }
  }
  if (sums) {
    const float t = block_sum256(part_m, sm.red);
    if (threadIdx.x == 0) atomicAdd(sums, (double)t);
    if (want_l1) {
      const float u = block_sum256(part_l1, sm.red);
      if (threadIdx.x == 0) atomicAdd(sums + 1, (double)u);
    }
  }
}

// out[i] += scale * sums[i] (float accumulators of the C ABI); one thread
__global__ void ssim_finalize_kernel(const double* __restrict__ sums, double scale, float* out0, float* out1) {
  // atomic: several views (streams) may fold into the same accumulators concurrently
  if (out0) atomicAdd(out0, (float)(sums[0] * scale));
  if (out1) atomicAdd(out1, (float)(sums[1] * scale));
}

struct SsimBwdSmem {
  float p[3][SROWS][SPITCH];  // dL/dmap * (dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
  float h[3][SROWS][ST];
};

// finalize: when non-NULL, block (0,0,0) also folds the forward kernel's double sums into the float accumulators
__global__ void __launch_bounds__(STHREADS) ssim_backward_kernel(const float* __restrict__ img1,
                                                                 const float* __restrict__ img2,
                                                                 const float* __restrict__ dL_dmap, float dL_uniform,
                                                                 const float* __restrict__ dm_dmu1,
                                                                 const float* __restrict__ dm_dsigma1_sq,
                                                                 const float* __restrict__ dm_dsigma12, int W, int H,
                                                                 int vec, float l1_weight, float* __restrict__ dL_dimg1,
                                                                 const double* __restrict__ fin_sums, double fin_scale,
                                                                 float* fin_ssim, float* fin_l1,
                                                                 const float* __restrict__ alpha, int clamp01) {
  extern __shared__ __align__(16) unsigned char ssim_smem[];
  SsimBwdSmem& sm = *reinterpret_cast<SsimBwdSmem*>(ssim_smem);
  const size_t plane = (size_t)blockIdx.z * W * H;
  const int tx0 = blockIdx.x * ST, ty0 = blockIdx.y * ST;
  if (fin_sums && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
    atomicAdd(fin_ssim, (float)(fin_sums[0] * fin_scale));  // atomic: the accumulators are shared by concurrent views
    atomicAdd(fin_l1, (float)(fin_sums[1] * fin_scale));
  }
  {
    const float* const src[3] = {dm_dmu1 + plane, dm_dsigma1_sq + plane, dm_dsigma12 + plane};
    stage_halo<3>(sm.p, src, dL_dmap ? dL_dmap + plane : nullptr, dL_uniform, W, H, tx0, ty0, vec != 0);
  }
  __syncthreads();
  for (int it = threadIdx.x; it < 3 * SROWS * (ST / 4); it += STHREADS) {
    const int q = it / (SROWS * (ST / 4)), rem = it - q * (SROWS * (ST / 4));
    const int r = rem / (ST / 4), xg = rem - r * (ST / 4);
    float v[20], o[4];
    load_row20(&sm.p[q][r][4 * xg], v);
    taps4(v, SOFF, o);
    *reinterpret_cast<float4*>(&sm.h[q][r][4 * xg]) = make_float4(o[0], o[1], o[2], o[3]);
  }
  __syncthreads();
  const int x = threadIdx.x & 31, yb = threadIdx.x >> 5;
  float acc[3][4];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    float col[14];
#pragma unroll
    for (int k = 0; k < 14; ++k) col[k] = sm.h[q][4 * yb + k][x];
    taps4(col, 0, acc[q]);
  }
  const int px = tx0 + x;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int py = ty0 + 4 * yb + j;
    if (px < W && py < H) {
      const size_t p = plane + (size_t)py * W + px;
      const float raw = __ldg(img1 + p), b = __ldg(img2 + p);
      // loss path: the image the losses saw is a = clamp(raw, 0, 1) * alpha; its cotangent reaches raw through alpha
      // where 0 <= raw <= 1 and not at all elsewhere (torch.clamp's backward)
      const float am = alpha ? __ldg(alpha + (size_t)py * W + px) : 1.f;
      const float a = (clamp01 ? clamp01f(raw) : raw) * am;
      // dL/da(p) = sum_q w(q-p) dL/dm(q) [dm/dmu1(q) + 2 a(p) dm/dsigma1_sq(q) + b(p) dm/dsigma12(q)]
      float d = acc[0][j] + 2.f * a * acc[1][j] + b * acc[2][j];
      if (l1_weight != 0.f) {
        const float e = a - b;
        d += l1_weight * ((e > 0.f) ? 1.f : ((e < 0.f) ? -1.f : 0.f));
      }
      d *= am;
      if (clamp01 && !(raw >= 0.f && raw <= 1.f)) d = 0.f;
      dL_dimg1[p] = d;
    }
  }
}

cudaError_t opt_in_smem(const void* fn, size_t bytes, int* configured_device) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (*configured_device != dev) {
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return e;
    *configured_device = dev;
  }
  return cudaSuccess;
}

bool rows_vectorizable(int W, std::initializer_list<const void*> ptrs) {
  if (W % 4 != 0) return false;
  for (const void* p : ptrs)
    if (p && (((uintptr_t)p) & 15u) != 0) return false;
  return true;
}

}  // namespace

// sums: device double[2], zeroed by the caller on the same stream (sum of the map, sum |img1 - img2| when want_l1)
cudaError_t launch_ssim_forward(float C1, float C2, const float* img1, const float* img2, int planes, int W, int H,
                                float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sums,
                                int want_l1, const float* alpha, int clamp01, cudaStream_t s) {
  if (planes <= 0 || W <= 0 || H <= 0) return cudaSuccess;
  static thread_local int dev_f = -1;
  cudaError_t e = opt_in_smem((const void*)ssim_forward_kernel, sizeof(SsimFwdSmem), &dev_f);
  if (e != cudaSuccess) return e;
  const int vec = rows_vectorizable(W, {img1, img2, alpha}) && (((size_t)W * H) % 4 == 0);
  dim3 grid((W + ST - 1) / ST, (H + ST - 1) / ST, planes);
  ssim_forward_kernel<<<grid, STHREADS, sizeof(SsimFwdSmem), s>>>(C1, C2, img1, img2, W, H, vec, ssim_map, dm_dmu1,
                                                                  dm_dsigma1_sq, dm_dsigma12, sums, want_l1, alpha, clamp01);
  return cudaGetLastError();
}

cudaError_t launch_ssim_finalize(const double* sums, double scale, float* out_ssim, float* out_l1, cudaStream_t s) {
  ssim_finalize_kernel<<<1, 1, 0, s>>>(sums, scale, out_ssim, out_l1);
  return cudaGetLastError();
}

cudaError_t launch_ssim_backward(const float* img1, const float* img2, const float* dL_dmap, float dL_uniform,
                                 const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, int planes, int W,
                                 int H, float l1_weight, float* dL_dimg1, const double* fin_sums, double fin_scale,
                                 float* fin_ssim, float* fin_l1, const float* alpha, int clamp01, cudaStream_t s) {
  if (planes <= 0 || W <= 0 || H <= 0) return cudaSuccess;
  static thread_local int dev_b = -1;
  cudaError_t e = opt_in_smem((const void*)ssim_backward_kernel, sizeof(SsimBwdSmem), &dev_b);
  if (e != cudaSuccess) return e;
  const int vec = rows_vectorizable(W, {dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dL_dmap}) && (((size_t)W * H) % 4 == 0);
  dim3 grid((W + ST - 1) / ST, (H + ST - 1) / ST, planes);
  ssim_backward_kernel<<<grid, STHREADS, sizeof(SsimBwdSmem), s>>>(img1, img2, dL_dmap, dL_uniform, dm_dmu1, dm_dsigma1_sq,
                                                                   dm_dsigma12, W, H, vec, l1_weight, dL_dimg1, fin_sums,
                                                                   fin_scale, fin_ssim, fin_l1, alpha, clamp01);
  return cudaGetLastError();
}

}  // namespace pdgs
