// Per-Gaussian math of the B200 rasterizer: forward geometry (projection, EWA
// covariance, radius, tile rect), SH colour, and the fused preprocess backward.
//
// Replaces the per-Gaussian device functions of the reference rasterizer
// (graphdeco-inria/diff-gaussian-rasterization @ 9c5c2028, branch dr_aa; source
// absent from /root/reference -- see SURVEY.md Appendix A.2-A.5, A.8).
//
// Everything that feeds an INTEGER result (radii, tile rects, depth sort keys,
// pixel centres) is written with explicit round-to-nearest intrinsics in one
// fixed ("canonical") operation order, documented in DESIGN.md, so that the
// fp32 bit-model under oracle/ can reproduce it bit for bit.  The functions are
// __host__ __device__ so the same source can be instantiated on the host; the
// product never executes the host instantiation.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define PD_HD __host__ __device__ __forceinline__

#if defined(__CUDA_ARCH__)
#define PD_MUL(a, b) __fmul_rn((a), (b))
#define PD_ADD(a, b) __fadd_rn((a), (b))
#define PD_SUB(a, b) __fsub_rn((a), (b))
#define PD_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define PD_DIV(a, b) __fdiv_rn((a), (b))
#define PD_SQRT(a) __fsqrt_rn((a))
#define PD_DMUL(a, b) __dmul_rn((a), (b))
#define PD_DADD(a, b) __dadd_rn((a), (b))
#else  // host build: compile with -ffp-contract=off
#define PD_MUL(a, b) ((float)((float)(a) * (float)(b)))
#define PD_ADD(a, b) ((float)((float)(a) + (float)(b)))
#define PD_SUB(a, b) ((float)((float)(a) - (float)(b)))
#define PD_FMA(a, b, c) fmaf((a), (b), (c))
#define PD_DIV(a, b) ((float)((float)(a) / (float)(b)))
#define PD_SQRT(a) sqrtf((a))
#define PD_DMUL(a, b) ((double)(a) * (double)(b))
#define PD_DADD(a, b) ((double)(a) + (double)(b))
#endif

namespace pdgs {

constexpr int TILE = 16;               // 16x16 pixel tiles (reference BLOCK_X/BLOCK_Y)
constexpr float NEAR_Z = 0.2f;         // A.2 near-plane cull
constexpr float DILATION = 0.3f;       // A.3 low-pass dilation
constexpr float AA_FLOOR = 0.000025f;  // A.3 antialiasing floor
constexpr float ALPHA_MIN = 1.0f / 255.0f;
constexpr float ALPHA_MAX = 0.99f;
constexpr float T_EPS = 0.0001f;

constexpr float SH_C0 = 0.28209479177387814f;
constexpr float SH_C1 = 0.4886025119029199f;
constexpr float SH_C2_0 = 1.0925484305920792f, SH_C2_1 = -1.0925484305920792f, SH_C2_2 = 0.31539156525252005f,
                SH_C2_3 = -1.0925484305920792f, SH_C2_4 = 0.5462742152960396f;
constexpr float SH_C3_0 = -0.5900435899266435f, SH_C3_1 = 2.890611442640554f, SH_C3_2 = -0.4570457994644658f,
                SH_C3_3 = 0.3731763325901154f, SH_C3_4 = -0.4570457994644658f, SH_C3_5 = 1.445305721320277f,
                SH_C3_6 = -0.5900435899266435f;

// Camera block, loaded once per CTA into shared memory.
struct Camera {
  float vm[16];  // world->view, flat index 4*c+r = element (r,c)
  float pm[16];  // full projection, same indexing
  float campos[3];
  float tanfovx, tanfovy, focal_x, focal_y;
  float scale_modifier;
  int W, H, grid_x, grid_y;
  int sh_degree, antialiasing;
};

// a0*b0 + a1*b1 + a2*b2 in the canonical contraction order.
PD_HD float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
  return PD_FMA(a2, b2, PD_FMA(a0, b0, PD_MUL(a1, b1)));
}
// m[r]*x + m[4+r]*y + m[8+r]*z + m[12+r]
PD_HD float affine_row(const float* m, int r, float x, float y, float z) {
  return PD_ADD(dot3c(m[r], x, m[4 + r], y, m[8 + r], z), m[12 + r]);
}

PD_HD float ndc2pix(float v, int S) {
  // reference uses double literals: ((v + 1.0) * S - 1.0) * 0.5, rounded once to fp32
  double d = PD_DMUL(PD_DADD(PD_DMUL(PD_DADD((double)v, 1.0), (double)S), -1.0), 0.5);
  return (float)d;
}

PD_HD int f2i_trunc_safe(float v) {
  v = fminf(fmaxf(v, -1.0e9f), 1.0e9f);
  return (int)v;
}

PD_HD void rotation_from_quat(const float q[4], float R[9]) {
  const float r = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = PD_SUB(1.0f, PD_MUL(2.0f, PD_FMA(y, y, PD_MUL(z, z))));
  R[1] = PD_MUL(2.0f, PD_FMA(x, y, -PD_MUL(r, z)));
  R[2] = PD_MUL(2.0f, PD_FMA(x, z, PD_MUL(r, y)));
  R[3] = PD_MUL(2.0f, PD_FMA(x, y, PD_MUL(r, z)));
  R[4] = PD_SUB(1.0f, PD_MUL(2.0f, PD_FMA(x, x, PD_MUL(z, z))));
  R[5] = PD_MUL(2.0f, PD_FMA(y, z, -PD_MUL(r, x)));
  R[6] = PD_MUL(2.0f, PD_FMA(x, z, -PD_MUL(r, y)));
  R[7] = PD_MUL(2.0f, PD_FMA(y, z, PD_MUL(r, x)));
  R[8] = PD_SUB(1.0f, PD_MUL(2.0f, PD_FMA(x, x, PD_MUL(y, y))));
}

// Sigma = (R S)(R S)^T packed [xx, xy, xz, yy, yz, zz]  (A.3)
PD_HD void cov3d_from_scale_rot(const float scale[3], float mod, const float q[4], float cov[6]) {
  float R[9];
  rotation_from_quat(q, R);
  const float s0 = PD_MUL(mod, scale[0]), s1 = PD_MUL(mod, scale[1]), s2 = PD_MUL(mod, scale[2]);
  float M[9];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    M[3 * i + 0] = PD_MUL(R[3 * i + 0], s0);
    M[3 * i + 1] = PD_MUL(R[3 * i + 1], s1);
    M[3 * i + 2] = PD_MUL(R[3 * i + 2], s2);
  }
  cov[0] = dot3c(M[0], M[0], M[1], M[1], M[2], M[2]);
  cov[1] = dot3c(M[0], M[3], M[1], M[4], M[2], M[5]);
  cov[2] = dot3c(M[0], M[6], M[1], M[7], M[2], M[8]);
  cov[3] = dot3c(M[3], M[3], M[4], M[4], M[5], M[5]);
  cov[4] = dot3c(M[3], M[6], M[4], M[7], M[5], M[8]);
  cov[5] = dot3c(M[6], M[6], M[7], M[7], M[8], M[8]);
}

// Intermediates of the EWA projection, shared by forward and backward.
struct Ewa {
  float tx, ty, tz;    // view-space position, lateral components clamped (A.3)
  float xmul, ymul;    // 0 where the 1.3*tanfov clamp is active
  float M0[3], M1[3];  // rows of J*Rw
  float v0[3], v1[3];  // Sigma*M0^T, Sigma*M1^T
  float a0, b, c0;     // cov2D before dilation
};

PD_HD void ewa_project(const Camera& cam, float pvx, float pvy, float pvz, const float cov[6], Ewa& e) {
  const float limx = PD_MUL(1.3f, cam.tanfovx), limy = PD_MUL(1.3f, cam.tanfovy);
  const float txtz = PD_DIV(pvx, pvz), tytz = PD_DIV(pvy, pvz);
  e.xmul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
  e.ymul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
  e.tx = PD_MUL(fminf(limx, fmaxf(-limx, txtz)), pvz);
  e.ty = PD_MUL(fminf(limy, fmaxf(-limy, tytz)), pvz);
  e.tz = pvz;
  const float tz2 = PD_MUL(pvz, pvz);
  const float J00 = PD_DIV(cam.focal_x, pvz);
  const float J02 = -PD_DIV(PD_MUL(cam.focal_x, e.tx), tz2);
  const float J11 = PD_DIV(cam.focal_y, pvz);
  const float J12 = -PD_DIV(PD_MUL(cam.focal_y, e.ty), tz2);
#pragma unroll
  for (int k = 0; k < 3; ++k) {  // Rw[r][k] = vm[4k + r]
    e.M0[k] = PD_FMA(J00, cam.vm[4 * k + 0], PD_MUL(J02, cam.vm[4 * k + 2]));
    e.M1[k] = PD_FMA(J11, cam.vm[4 * k + 1], PD_MUL(J12, cam.vm[4 * k + 2]));
  }
  const float S00 = cov[0], S01 = cov[1], S02 = cov[2], S11 = cov[3], S12 = cov[4], S22 = cov[5];
  e.v0[0] = dot3c(S00, e.M0[0], S01, e.M0[1], S02, e.M0[2]);
  e.v0[1] = dot3c(S01, e.M0[0], S11, e.M0[1], S12, e.M0[2]);
  e.v0[2] = dot3c(S02, e.M0[0], S12, e.M0[1], S22, e.M0[2]);
  e.v1[0] = dot3c(S00, e.M1[0], S01, e.M1[1], S02, e.M1[2]);
  e.v1[1] = dot3c(S01, e.M1[0], S11, e.M1[1], S12, e.M1[2]);
  e.v1[2] = dot3c(S02, e.M1[0], S12, e.M1[1], S22, e.M1[2]);
  e.a0 = dot3c(e.M0[0], e.v0[0], e.M0[1], e.v0[1], e.M0[2], e.v0[2]);
  e.b = dot3c(e.M0[0], e.v1[0], e.M0[1], e.v1[1], e.M0[2], e.v1[2]);
  e.c0 = dot3c(e.M1[0], e.v1[0], e.M1[1], e.v1[1], e.M1[2], e.v1[2]);
}

// Result of the forward geometry of one Gaussian.
struct GeomOut {
  int radius;              // 0 = culled
  int rx0, ry0, rx1, ry1;  // tile rect [min, max)
  float depth;             // view-space z (sort key bits)
  float px, py;            // pixel centre
  float conA, conB, conC;  // conic
  float h;                 // antialiasing opacity factor (1 when off)
};

PD_HD void tile_rect(float px, float py, int radius, int grid_x, int grid_y, int& rx0, int& ry0, int& rx1, int& ry1) {
  const float rf = (float)radius;
  const float tile = (float)TILE;
  rx0 = min(grid_x, max(0, f2i_trunc_safe(PD_DIV(PD_SUB(px, rf), tile))));
  ry0 = min(grid_y, max(0, f2i_trunc_safe(PD_DIV(PD_SUB(py, rf), tile))));
  rx1 = min(grid_x, max(0, f2i_trunc_safe(PD_DIV(PD_SUB(PD_ADD(PD_ADD(px, rf), tile), 1.0f), tile))));
  ry1 = min(grid_y, max(0, f2i_trunc_safe(PD_DIV(PD_SUB(PD_ADD(PD_ADD(py, rf), tile), 1.0f), tile))));
}

// Forward geometry (A.2-A.4). cov is the 3D covariance (already built).
PD_HD bool forward_geometry(const Camera& cam, float mx, float my, float mz, const float cov[6], GeomOut& g) {
  g.radius = 0;
  const float pvx = affine_row(cam.vm, 0, mx, my, mz);
  const float pvy = affine_row(cam.vm, 1, mx, my, mz);
  const float pvz = affine_row(cam.vm, 2, mx, my, mz);
  if (!(pvz > NEAR_Z)) return false;
  const float hx = affine_row(cam.pm, 0, mx, my, mz);
  const float hy = affine_row(cam.pm, 1, mx, my, mz);
  const float hw = affine_row(cam.pm, 3, mx, my, mz);
  const float pw = PD_DIV(1.0f, PD_ADD(hw, 0.0000001f));
  const float ndcx = PD_MUL(hx, pw), ndcy = PD_MUL(hy, pw);
  Ewa e;
  ewa_project(cam, pvx, pvy, pvz, cov, e);
  const float det0 = PD_FMA(e.a0, e.c0, -PD_MUL(e.b, e.b));
  const float a = PD_ADD(e.a0, DILATION), c = PD_ADD(e.c0, DILATION);
  const float det = PD_FMA(a, c, -PD_MUL(e.b, e.b));
  // det == 0 culls as upstream does; a non-finite covariance (NaN scale / rotation, exp() overflow of a raw scale) is
  // culled too: upstream's (int)NaN == 0 radius gives a zero-area rect there, here the same Gaussian must not reach
  // the binning with an inverted rect
  if (det == 0.0f || !(det == det)) return false;
  g.h = cam.antialiasing ? PD_SQRT(fmaxf(AA_FLOOR, PD_DIV(det0, det))) : 1.0f;
  const float det_inv = PD_DIV(1.0f, det);
  g.conA = PD_MUL(c, det_inv);
  g.conB = PD_MUL(-e.b, det_inv);
  g.conC = PD_MUL(a, det_inv);
  const float mid = PD_MUL(0.5f, PD_ADD(a, c));
  const float disc = PD_SQRT(fmaxf(0.1f, PD_FMA(mid, mid, -det)));
  const float lam = fmaxf(PD_ADD(mid, disc), PD_SUB(mid, disc));
  if (!(lam == lam)) return false;
  const int radius = f2i_trunc_safe(ceilf(PD_MUL(3.0f, PD_SQRT(lam))));
  if (radius <= 0) return false;
  g.px = ndc2pix(ndcx, cam.W);
  g.py = ndc2pix(ndcy, cam.H);
  tile_rect(g.px, g.py, radius, cam.grid_x, cam.grid_y, g.rx0, g.ry0, g.rx1, g.ry1);
  if (!(g.rx1 > g.rx0 && g.ry1 > g.ry0)) return false;  // empty (or, for a NaN centre, inverted) rect
  g.radius = radius;
  g.depth = pvz;
  return true;
}

// ---------------------------------------------------------------------------
// SH colour (A.5).  sh points at this Gaussian's [M][3] block (RGB innermost).
// Returns rgb and a 3-bit mask of channels clamped at 0.
// ---------------------------------------------------------------------------
struct ShBasis {
  float b[16];
};

PD_HD void sh_basis(int deg, float x, float y, float z, float b[16]) {
  b[0] = SH_C0;
  if (deg > 0) {
    b[1] = -SH_C1 * y;
    b[2] = SH_C1 * z;
    b[3] = -SH_C1 * x;
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = SH_C2_0 * xy;
      b[5] = SH_C2_1 * yz;
      b[6] = SH_C2_2 * (2.0f * zz - xx - yy);
      b[7] = SH_C2_3 * xz;
      b[8] = SH_C2_4 * (xx - yy);
      if (deg > 2) {
        b[9] = SH_C3_0 * y * (3.0f * xx - yy);
        b[10] = SH_C3_1 * xy * z;
        b[11] = SH_C3_2 * y * (4.0f * zz - xx - yy);
        b[12] = SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        b[13] = SH_C3_4 * x * (4.0f * zz - xx - yy);
        b[14] = SH_C3_5 * z * (xx - yy);
        b[15] = SH_C3_6 * x * (xx - 3.0f * yy);
      }
    }
  }
}

PD_HD int sh_num_coeffs(int deg) { return (deg + 1) * (deg + 1); }

template <typename ShLoad>
PD_HD unsigned sh_to_rgb(int deg, float dx, float dy, float dz, ShLoad sh, float rgb[3]) {
  const float inv = 1.0f / sqrtf(dx * dx + dy * dy + dz * dz);
  const float x = dx * inv, y = dy * inv, z = dz * inv;
  float b[16];
  sh_basis(deg, x, y, z, b);
  const int n = sh_num_coeffs(deg);
  float r = 0.f, g = 0.f, bl = 0.f;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    if (k < n) {
      r += b[k] * sh(3 * k + 0);
      g += b[k] * sh(3 * k + 1);
      bl += b[k] * sh(3 * k + 2);
    }
  }
  r += 0.5f;
  g += 0.5f;
  bl += 0.5f;
  unsigned clamped = (r < 0.f ? 1u : 0u) | (g < 0.f ? 2u : 0u) | (bl < 0.f ? 4u : 0u);
  rgb[0] = fmaxf(r, 0.f);
  rgb[1] = fmaxf(g, 0.f);
  rgb[2] = fmaxf(bl, 0.f);
  return clamped;
}

// ---------------------------------------------------------------------------
// Fused preprocess backward for one Gaussian (reference kernels K8
// computeCov2DCUDA-backward + K9 preprocessCUDA-backward).
// ---------------------------------------------------------------------------
struct BwdIn {
  float g_mean2d_x, g_mean2d_y;        // dL/dmean2D in NDC units (as returned to the caller)
  float g_conic_x, g_conic_y, g_conic_w;  // K7 convention: .y is per off-diagonal entry
  float g_opacity;                     // dL/d(conic_opacity.w)
  float g_rgb[3];
  float g_invdepth;
  unsigned clamped;
};

struct BwdOut {
  float g_mean[3];
  float g_opacity;
  float g_scale[3];
  float g_rot[4];
  float g_cov[6];
};

// dL/dsh[k][c] = basis_k * dL/drgb_c (masked); returns dL/d(unnormalised dir).
template <typename ShLoad>
PD_HD void sh_backward_dir(int deg, float dx, float dy, float dz, ShLoad sh, const float g[3], float gdir[3], float basis[16]) {
  const float len = sqrtf(dx * dx + dy * dy + dz * dz);
  const float inv = 1.0f / len;
  const float x = dx * inv, y = dy * inv, z = dz * inv;
  sh_basis(deg, x, y, z, basis);
  float ddx = 0.f, ddy = 0.f, ddz = 0.f;  // dL/d(normalised dir)
  if (deg > 0) {
    // s(k) = sum_c g[c] * sh[k][c]
#define PD_S(k) (g[0] * sh(3 * (k) + 0) + g[1] * sh(3 * (k) + 1) + g[2] * sh(3 * (k) + 2))
    const float s1 = PD_S(1), s2 = PD_S(2), s3 = PD_S(3);
    ddx = -SH_C1 * s3;
    ddy = -SH_C1 * s1;
    ddz = SH_C1 * s2;
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      const float s4 = PD_S(4), s5 = PD_S(5), s6 = PD_S(6), s7 = PD_S(7), s8 = PD_S(8);
      ddx += SH_C2_0 * y * s4 + SH_C2_2 * 2.f * -x * s6 + SH_C2_3 * z * s7 + SH_C2_4 * 2.f * x * s8;
      ddy += SH_C2_0 * x * s4 + SH_C2_1 * z * s5 + SH_C2_2 * 2.f * -y * s6 + SH_C2_4 * 2.f * -y * s8;
      ddz += SH_C2_1 * y * s5 + SH_C2_2 * 4.f * z * s6 + SH_C2_3 * x * s7;
      if (deg > 2) {
        const float s9 = PD_S(9), s10 = PD_S(10), s11 = PD_S(11), s12 = PD_S(12), s13 = PD_S(13), s14 = PD_S(14),
                    s15 = PD_S(15);
        ddx += SH_C3_0 * s9 * 6.f * xy + SH_C3_1 * s10 * yz + SH_C3_2 * s11 * -2.f * xy + SH_C3_3 * s12 * -6.f * xz +
               SH_C3_4 * s13 * (-3.f * xx + 4.f * zz - yy) + SH_C3_5 * s14 * 2.f * xz + SH_C3_6 * s15 * 3.f * (xx - yy);
        ddy += SH_C3_0 * s9 * 3.f * (xx - yy) + SH_C3_1 * s10 * xz + SH_C3_2 * s11 * (-3.f * yy + 4.f * zz - xx) +
               SH_C3_3 * s12 * -6.f * yz + SH_C3_4 * s13 * -2.f * xy + SH_C3_5 * s14 * -2.f * yz + SH_C3_6 * s15 * -6.f * xy;
        ddz += SH_C3_1 * s10 * xy + SH_C3_2 * s11 * 8.f * yz + SH_C3_3 * s12 * 3.f * (2.f * zz - xx - yy) +
               SH_C3_4 * s13 * 8.f * xz + SH_C3_5 * s14 * (xx - yy);
      }
    }
#undef PD_S
  }
  // through n = v/|v|:  dL/dv = (dL/dn - n (n . dL/dn)) / |v|
  const float nd = x * ddx + y * ddy + z * ddz;
  gdir[0] = (ddx - x * nd) * inv;
  gdir[1] = (ddy - y * nd) * inv;
  gdir[2] = (ddz - z * nd) * inv;
}

// Geometry part of the backward. cov = 3D covariance used in forward.
// Produces dL/dmean (geometry terms only), dL/dopacity and dL/dcov3D.
PD_HD void geometry_backward(const Camera& cam, float mx, float my, float mz, const float cov[6], float opacity,
                             const BwdIn& in, float g_mean[3], float& g_opacity, float g_cov[6]) {
  const float pvx = affine_row(cam.vm, 0, mx, my, mz);
  const float pvy = affine_row(cam.vm, 1, mx, my, mz);
  const float pvz = affine_row(cam.vm, 2, mx, my, mz);
  Ewa e;
  ewa_project(cam, pvx, pvy, pvz, cov, e);
  const float a = e.a0 + DILATION, c = e.c0 + DILATION, b = e.b;
  const float det0 = e.a0 * e.c0 - b * b;
  const float det = a * c - b * b;
  float ga = 0.f, gb = 0.f, gc = 0.f;
  g_opacity = in.g_opacity;
  if (cam.antialiasing) {
    const float ratio = det0 / det;
    const float h = sqrtf(fmaxf(AA_FLOOR, ratio));
    const float g_h = in.g_opacity * opacity;
    g_opacity = in.g_opacity * h;
    const float g_f = (ratio <= AA_FLOOR) ? 0.f : g_h / (2.f * h);
    const float w = DILATION, x = e.a0, y = e.c0, z = b;
    const float s = g_f / (det * det);
    ga += s * w * (w * y + y * y + z * z);
    gc += s * w * (w * x + x * x + z * z);
    gb += s * (-2.f * w * z * (w + x + y));
  }
  {
    const float inv2 = 1.f / (det * det);
    const float gx = in.g_conic_x, gy = in.g_conic_y, gz = in.g_conic_w;
    ga += inv2 * (-c * c * gx + 2.f * b * c * gy + (det - a * c) * gz);
    gc += inv2 * (-a * a * gz + 2.f * a * b * gy + (det - a * c) * gx);
    gb += inv2 * 2.f * (b * c * gx - (det + 2.f * b * b) * gy + a * b * gz);
  }
  const float* M0 = e.M0;
  const float* M1 = e.M1;
  g_cov[0] = M0[0] * M0[0] * ga + M0[0] * M1[0] * gb + M1[0] * M1[0] * gc;
  g_cov[3] = M0[1] * M0[1] * ga + M0[1] * M1[1] * gb + M1[1] * M1[1] * gc;
  g_cov[5] = M0[2] * M0[2] * ga + M0[2] * M1[2] * gb + M1[2] * M1[2] * gc;
  g_cov[1] = 2.f * M0[0] * M0[1] * ga + (M0[0] * M1[1] + M0[1] * M1[0]) * gb + 2.f * M1[0] * M1[1] * gc;
  g_cov[2] = 2.f * M0[0] * M0[2] * ga + (M0[0] * M1[2] + M0[2] * M1[0]) * gb + 2.f * M1[0] * M1[2] * gc;
  g_cov[4] = 2.f * M0[1] * M0[2] * ga + (M0[1] * M1[2] + M0[2] * M1[1]) * gb + 2.f * M1[1] * M1[2] * gc;
  float dM0[3], dM1[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    dM0[k] = 2.f * ga * e.v0[k] + gb * e.v1[k];
    dM1[k] = 2.f * gc * e.v1[k] + gb * e.v0[k];
  }
  float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    dJ00 += dM0[k] * cam.vm[4 * k + 0];
    dJ02 += dM0[k] * cam.vm[4 * k + 2];
    dJ11 += dM1[k] * cam.vm[4 * k + 1];
    dJ12 += dM1[k] * cam.vm[4 * k + 2];
  }
  const float tzi = 1.f / e.tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
  const float dtx = e.xmul * (-cam.focal_x * tz2) * dJ02;
  const float dty = e.ymul * (-cam.focal_y * tz2) * dJ12;
  float dtz = -cam.focal_x * tz2 * dJ00 - cam.focal_y * tz2 * dJ11 + 2.f * cam.focal_x * e.tx * tz3 * dJ02 +
              2.f * cam.focal_y * e.ty * tz3 * dJ12;
  dtz -= in.g_invdepth * tz2;  // invdepth = 1 / p_view.z
#pragma unroll
  for (int j = 0; j < 3; ++j) g_mean[j] = cam.vm[4 * j + 0] * dtx + cam.vm[4 * j + 1] * dty + cam.vm[4 * j + 2] * dtz;
  // screen-space mean gradient (NDC units) through the perspective divide
  const float hx = affine_row(cam.pm, 0, mx, my, mz);
  const float hy = affine_row(cam.pm, 1, mx, my, mz);
  const float hw = affine_row(cam.pm, 3, mx, my, mz);
  const float m_w = 1.f / (hw + 0.0000001f);
  const float mul1 = hx * m_w * m_w, mul2 = hy * m_w * m_w;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    g_mean[j] += (cam.pm[4 * j + 0] * m_w - cam.pm[4 * j + 3] * mul1) * in.g_mean2d_x +
                 (cam.pm[4 * j + 1] * m_w - cam.pm[4 * j + 3] * mul2) * in.g_mean2d_y;
  }
}

// dL/dcov3D -> dL/dscale, dL/dquaternion (quaternion used as given, A.3/A.8)
PD_HD void cov3d_backward(const float scale[3], float mod, const float q[4], const float g_cov[6], float g_scale[3],
                          float g_rot[4]) {
  float R[9];
  rotation_from_quat(q, R);
  const float s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
  float M[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) M[3 * i + k] = R[3 * i + k] * s[k];
  const float G[9] = {g_cov[0],        0.5f * g_cov[1], 0.5f * g_cov[2], 0.5f * g_cov[1], g_cov[3],
                      0.5f * g_cov[4], 0.5f * g_cov[2], 0.5f * g_cov[4], g_cov[5]};
  float dM[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k)
      dM[3 * i + k] = 2.f * (G[3 * i + 0] * M[0 + k] + G[3 * i + 1] * M[3 + k] + G[3 * i + 2] * M[6 + k]);
  float dR[9];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    g_scale[k] = mod * (R[0 + k] * dM[0 + k] + R[3 + k] * dM[3 + k] + R[6 + k] * dM[6 + k]);
#pragma unroll
    for (int i = 0; i < 3; ++i) dR[3 * i + k] = dM[3 * i + k] * s[k];
  }
  const float r = q[0], x = q[1], y = q[2], z = q[3];
  g_rot[0] = 2.f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
  g_rot[1] = 2.f * (y * dR[1] + z * dR[2] + y * dR[3] - 2.f * x * dR[4] - r * dR[5] + z * dR[6] + r * dR[7] - 2.f * x * dR[8]);
  g_rot[2] = 2.f * (-2.f * y * dR[0] + x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] + z * dR[7] - 2.f * y * dR[8]);
  g_rot[3] = 2.f * (-2.f * z * dR[0] - r * dR[1] + x * dR[2] + r * dR[3] - 2.f * z * dR[4] + y * dR[5] + x * dR[6] + y * dR[7]);
}

}  // namespace pdgs
