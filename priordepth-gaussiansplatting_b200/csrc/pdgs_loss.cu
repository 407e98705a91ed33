// Image-space loss of one view and its cotangents in a single pass over the rendered planes.
// Replaces the caller-side elementwise torch kernels of the reference's training step:
//   Ll1 = l1_loss(image, gt_image)                              /root/reference/train.py:153, utils/loss_utils.py:40-41
//   Ll1depth_pure = |(invDepth - mono_invdepth) * depth_mask|.mean()   /root/reference/train.py:174
// and their autograd backward (sign / N), so that the cotangents the backward blend consumes are
// produced by one HBM pass (read 2 planes-sets, write 1) instead of ~8 elementwise launches.
#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

constexpr int LOSS_THREADS = 256;

__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

__device__ __forceinline__ float block_sum(float v, float* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = (lane < LOSS_THREADS / 32) ? s_red[lane] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;  // valid in thread 0
}

// The reference's caller clamps the rendered image and multiplies it by the camera's alpha mask before the losses
// (gaussian_renderer/__init__.py:121 `rendered_image.clamp(0, 1)`, train.py:148-150 `image *= alpha_mask`): with
// clamp01 / alpha the loss is taken on a = clamp(c, 0, 1) * alpha and the cotangent w.r.t. the raw colour c is
// dL/da * alpha where 0 <= c <= 1 and zero elsewhere (torch.clamp's backward).  alpha is one [H*W] plane shared by
// the three colour planes (NULL = ones).
__device__ __forceinline__ float loss_elem(float c, float t, float am, int clamp01, float gc, float& acc) {
  const bool pass = !clamp01 || (c >= 0.f && c <= 1.f);
  const float a = (clamp01 ? fminf(fmaxf(c, 0.f), 1.f) : c) * am;
  const float d = a - t;
  acc += fabsf(d);
  return pass ? gc * sgn(d) * am : 0.f;
}

// vec_ok: colour planes readable as aligned float4 (and H*W % 4 == 0 when alpha is given, so that a float4 never
// straddles two planes).
__global__ void __launch_bounds__(LOSS_THREADS) image_loss_kernel(const float* __restrict__ color,
                                                                  const float* __restrict__ target,
                                                                  const float* __restrict__ invdepth,
                                                                  const float* __restrict__ prior,
                                                                  const float* __restrict__ mask,
                                                                  const float* __restrict__ alpha, int clamp01,
                                                                  size_t n_color, size_t n_depth, float gc, float gd,
                                                                  float* __restrict__ dL_dcolor,
                                                                  float* __restrict__ dL_dinvdepth,
                                                                  float* __restrict__ loss, int vec_ok) {
  __shared__ float s_red[LOSS_THREADS / 32];
  const size_t stride = (size_t)gridDim.x * LOSS_THREADS;
  const size_t tid = (size_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
  float acc_c = 0.f, acc_d = 0.f;
  if (vec_ok) {
    const size_t n4 = n_color / 4, hw4 = n_depth / 4;
    const float4* c4 = reinterpret_cast<const float4*>(color);
    const float4* t4 = reinterpret_cast<const float4*>(target);
    const float4* a4 = reinterpret_cast<const float4*>(alpha);
    float4* g4 = reinterpret_cast<float4*>(dL_dcolor);
    for (size_t i = tid; i < n4; i += stride) {
      const float4 c = __ldg(c4 + i), t = __ldg(t4 + i);
      float4 m = make_float4(1.f, 1.f, 1.f, 1.f);
      if (alpha) m = __ldg(a4 + (i % hw4));
      float4 g;
      g.x = loss_elem(c.x, t.x, m.x, clamp01, gc, acc_c);
      g.y = loss_elem(c.y, t.y, m.y, clamp01, gc, acc_c);
      g.z = loss_elem(c.z, t.z, m.z, clamp01, gc, acc_c);
      g.w = loss_elem(c.w, t.w, m.w, clamp01, gc, acc_c);
      g4[i] = g;
    }
    for (size_t i = n4 * 4 + tid; i < n_color; i += stride)
      dL_dcolor[i] = loss_elem(color[i], target[i], alpha ? __ldg(alpha + (i % n_depth)) : 1.f, clamp01, gc, acc_c);
  } else {
    for (size_t i = tid; i < n_color; i += stride)
      dL_dcolor[i] = loss_elem(color[i], target[i], alpha ? __ldg(alpha + (i % n_depth)) : 1.f, clamp01, gc, acc_c);
  }
  if (prior) {
    for (size_t i = tid; i < n_depth; i += stride) {
      const float m = mask ? __ldg(mask + i) : 1.f;
      const float d = (__ldg(invdepth + i) - __ldg(prior + i)) * m;
      acc_d += fabsf(d);
      dL_dinvdepth[i] = gd * sgn(d) * m;
    }
  }
  const float sc = block_sum(acc_c, s_red);
  if (threadIdx.x == 0 && sc != 0.f) atomicAdd(loss + 0, sc / (float)n_color);
  if (prior) {
    const float sd = block_sum(acc_d, s_red);
    if (threadIdx.x == 0 && sd != 0.f) atomicAdd(loss + 1, sd / (float)n_depth);
  }
}

}  // namespace

cudaError_t launch_image_loss(const float* color, const float* target, const float* invdepth, const float* prior,
                              const float* mask, const float* alpha, int clamp01, int W, int H, float w_color,
                              float w_depth, float* dL_dcolor, float* dL_dinvdepth, float* loss, cudaStream_t s) {
  const size_t n_depth = (size_t)W * H, n_color = color ? 3 * n_depth : 0;  // color == NULL: depth-prior term only
  const float gc = n_color ? w_color / (float)n_color : 0.f, gd = w_depth / (float)n_depth;
  const int vec_ok = (((((uintptr_t)color) | ((uintptr_t)target) | ((uintptr_t)dL_dcolor) | ((uintptr_t)alpha)) & 15u) == 0 &&
                      (!alpha || n_depth % 4 == 0))
                         ? 1
                         : 0;
  size_t blocks = ((n_color ? n_color : n_depth) / 4 + LOSS_THREADS - 1) / LOSS_THREADS;
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  if (blocks < 1) blocks = 1;
  image_loss_kernel<<<(unsigned)blocks, LOSS_THREADS, 0, s>>>(color, target, invdepth, prior, mask, alpha, clamp01, n_color,
                                                              n_depth, gc, gd, dL_dcolor, dL_dinvdepth, loss, vec_ok);
  return cudaGetLastError();
}

}  // namespace pdgs
