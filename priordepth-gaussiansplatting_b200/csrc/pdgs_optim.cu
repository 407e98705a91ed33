// Visibility-sparse Adam: the optimiser step the reference takes when the rasterizer package
// exports SparseGaussianAdam (/root/reference/train.py:47-52, 332-334; scene/gaussian_model.py:25-28, 196-198):
// `optimizer.step(visible, N)` updates only the rows of Gaussians that were visible in the rendered view.  The
// kernel behind that class (`adamUpdate` of the accelerated rasterizer branch) is not in /root/reference; the update
// rule is restated in oracle/adam_oracle.py.
//
// HBM-bound elementwise work: per updated float 4 loads (param, grad, exp_avg, exp_avg_sq) and 3 stores.  All
// parameter groups of the model (xyz 3, f_dc 3, f_rest 45, opacity 1, scaling 3, rotation 4 floats per Gaussian) are
// handled by ONE launch; every thread owns one 16-byte chunk of a group's flat [N*M] array, looks up the visibility of
// the (at most two... four) rows its four floats belong to, and touches memory only when one of them is visible.
#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

constexpr int ADAM_THREADS = 256;

struct AdamLaunch {
  AdamGroup g[ADAM_MAX_GROUPS];
  unsigned block_end[ADAM_MAX_GROUPS];  // exclusive prefix of blocks per group
  int n_groups;
  const uint8_t* visible;  // [N] or NULL
  const int32_t* radii;    // [N] (visible = radii > 0) when `visible` is NULL
  int64_t N;
  float b1, b2, eps;
  float omb1, omb2;     // 1 - b1, 1 - b2 (see launch_sparse_adam for how they are rounded)
  float bc2_sqrt;       // sqrt(1 - b2^step); the 1 / (1 - b1^step) factor is folded into each group's step size
  int bias_correction;
};

__device__ __forceinline__ bool row_visible(const AdamLaunch& L, int64_t row) {
  return L.visible ? (__ldg(L.visible + row) != 0) : (__ldg(L.radii + row) > 0);
}

// One element, in a fixed operation order with every product and sum rounded separately (no contraction), so the
// result is reproducible bit for bit by the fp32 oracle.
__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, float lr, const AdamLaunch& L) {
  m = __fadd_rn(__fmul_rn(L.b1, m), __fmul_rn(L.omb1, g));
  v = __fadd_rn(__fmul_rn(L.b2, v), __fmul_rn(__fmul_rn(L.omb2, g), g));
  if (L.bias_correction) {
    // torch.optim.Adam: denom = sqrt(v) / sqrt(1 - b2^t) + eps, p -= (lr / (1 - b1^t)) * m / denom   (lr = step size here)
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), L.bc2_sqrt), L.eps);
    p = __fsub_rn(p, __fdiv_rn(__fmul_rn(lr, m), denom));
  } else {
    p = __fsub_rn(p, __fdiv_rn(__fmul_rn(lr, m), __fadd_rn(__fsqrt_rn(v), L.eps)));
  }
}

template <int MT>  // MT > 0: compile-time row length; MT == 0: runtime M
__device__ __forceinline__ void adam_group(const AdamLaunch& L, const AdamGroup& G, unsigned block_in_group) {
  const int M = MT > 0 ? MT : G.M;
  const int64_t total = L.N * (int64_t)M;
  const int64_t e0 = ((int64_t)block_in_group * ADAM_THREADS + threadIdx.x) * 4;
  if (e0 >= total) return;
  bool vis[4];
  bool any = false;
  int64_t prev_row = -1;
  bool prev_vis = false;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int64_t e = e0 + k;
    if (e < total) {
      const int64_t row = MT > 0 ? e / MT : e / M;
      if (row != prev_row) { prev_vis = row_visible(L, row); prev_row = row; }
      vis[k] = prev_vis;
    } else {
      vis[k] = false;
    }
    any |= vis[k];
  }
  if (!any) return;
  if (G.vec4 && e0 + 4 <= total) {
    float4 p = *reinterpret_cast<const float4*>(G.param + e0);
    const float4 g = __ldg(reinterpret_cast<const float4*>(G.grad + e0));
    float4 m = *reinterpret_cast<const float4*>(G.exp_avg + e0);
    float4 v = *reinterpret_cast<const float4*>(G.exp_avg_sq + e0);
    if (vis[0]) adam_elem(p.x, g.x, m.x, v.x, G.lr, L);
    if (vis[1]) adam_elem(p.y, g.y, m.y, v.y, G.lr, L);
    if (vis[2]) adam_elem(p.z, g.z, m.z, v.z, G.lr, L);
    if (vis[3]) adam_elem(p.w, g.w, m.w, v.w, G.lr, L);
    *reinterpret_cast<float4*>(G.param + e0) = p;
    *reinterpret_cast<float4*>(G.exp_avg + e0) = m;
    *reinterpret_cast<float4*>(G.exp_avg_sq + e0) = v;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (!vis[k]) continue;
      const int64_t e = e0 + k;
      float p = G.param[e], m = G.exp_avg[e], v = G.exp_avg_sq[e];
      adam_elem(p, __ldg(G.grad + e), m, v, G.lr, L);
      G.param[e] = p; G.exp_avg[e] = m; G.exp_avg_sq[e] = v;
    }
  }
}

__global__ void __launch_bounds__(ADAM_THREADS) sparse_adam_kernel(const __grid_constant__ AdamLaunch L) {
  int gi = 0;
  while (gi < L.n_groups - 1 && blockIdx.x >= L.block_end[gi]) ++gi;
  const unsigned first = gi == 0 ? 0u : L.block_end[gi - 1];
  const AdamGroup& G = L.g[gi];
  const unsigned b = blockIdx.x - first;
  switch (G.M) {
    case 1: adam_group<1>(L, G, b); break;
    case 3: adam_group<3>(L, G, b); break;
    case 4: adam_group<4>(L, G, b); break;
    case 45: adam_group<45>(L, G, b); break;
    case 48: adam_group<48>(L, G, b); break;
    default: adam_group<0>(L, G, b); break;
  }
}

}  // namespace

cudaError_t launch_sparse_adam(const AdamGroup* groups, int n_groups, const uint8_t* visible, const int32_t* radii,
                               int64_t N, double b1, double b2, double eps, int step, cudaStream_t s) {
  if (n_groups <= 0 || N <= 0) return cudaSuccess;
  if (n_groups > ADAM_MAX_GROUPS) return cudaErrorInvalidValue;
  AdamLaunch L;
  L.n_groups = n_groups; L.visible = visible; L.radii = radii; L.N = N;
  L.b1 = (float)b1; L.b2 = (float)b2; L.eps = (float)eps;
  L.bias_correction = step > 0 ? 1 : 0;
  L.bc2_sqrt = 1.0f;
  double bc1 = 1.0;
  if (step > 0) {  // the double-precision host scalars torch.optim.Adam computes, rounded to fp32 once
    L.omb1 = (float)(1.0 - b1); L.omb2 = (float)(1.0 - b2);
    bc1 = 1.0 - pow(b1, (double)step);
    L.bc2_sqrt = (float)sqrt(1.0 - pow(b2, (double)step));
  } else {         // the upstream sparse kernel subtracts in fp32: (1.0f - b1), (1.0f - b2)
    L.omb1 = 1.0f - L.b1; L.omb2 = 1.0f - L.b2;
  }
  unsigned blocks = 0;
  for (int i = 0; i < n_groups; ++i) {
    L.g[i] = groups[i];
    L.g[i].lr = (float)(groups[i].lr_host / bc1);
    const int64_t chunks = (N * (int64_t)groups[i].M + 3) / 4;
    const int64_t nb = (chunks + ADAM_THREADS - 1) / ADAM_THREADS;
    if (nb + blocks > 0x7fffffffLL) return cudaErrorInvalidValue;
    blocks += (unsigned)nb;
    L.block_end[i] = blocks;
    const uintptr_t al = (uintptr_t)groups[i].param | (uintptr_t)groups[i].grad | (uintptr_t)groups[i].exp_avg |
                         (uintptr_t)groups[i].exp_avg_sq;
    L.g[i].vec4 = (al & 15u) == 0 ? 1 : 0;
  }
  for (int i = n_groups; i < ADAM_MAX_GROUPS; ++i) { L.g[i] = AdamGroup{}; L.block_end[i] = blocks; }
  sparse_adam_kernel<<<blocks, ADAM_THREADS, 0, s>>>(L);
  return cudaGetLastError();
}

}  // namespace pdgs
