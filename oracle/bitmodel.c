/* fp32 bit-model of the rasterizer's INTEGER side: radii, tile rects, depth sort keys,
 * the (tile|depth) key list, its stable sort and the per-tile ranges.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/raster_oracle.py header).  PARITY UNPINNED: the
 * reference rasterizer source is absent from /root/reference; this restates SURVEY.md
 * Appendix A.2-A.4 and A.6 in IEEE fp32 with the operation order fixed in DESIGN.md
 * ("canonical fp32 order"): a*b + c*d + e*f is evaluated fma(e,f, fma(a,b, c*d)), the
 * affine rows add their constant last, ndc2Pix is done in double and rounded once.
 * Written independently of the CUDA source; compile with -ffp-contract=off so that
 * only the explicit fmaf() calls fuse.
 *
 * Call site being modelled: /root/reference/gaussian_renderer/__init__.py:104-112
 * (the rasterizer call); matrices follow /root/reference/scene/cameras.py:86-88
 * (flat index 4*c + r = element (r,c)).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BM_TILE 16

static float d3(float a0, float b0, float a1, float b1, float a2, float b2) {
  float t = a1 * b1;
  t = fmaf(a0, b0, t);
  return fmaf(a2, b2, t);
}

static float row(const float* m, int r, float x, float y, float z) {
  float t = d3(m[r], x, m[4 + r], y, m[8 + r], z);
  return t + m[12 + r];
}

static float pix_of_ndc(float v, int S) {
  double d = (((double)v + 1.0) * (double)S - 1.0) * 0.5;
  return (float)d;
}

static int trunc_i(float v) {
  if (!(v > -1.0e9f)) v = -1.0e9f; /* also maps NaN */
  if (v > 1.0e9f) v = 1.0e9f;
  return (int)v;
}

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static void quat_to_cov(const float* sc, float mod, const float* q, float* cov) {
  float r = q[0], x = q[1], y = q[2], z = q[3];
  float R[9];
  R[0] = 1.0f - 2.0f * fmaf(y, y, z * z);
  R[1] = 2.0f * fmaf(x, y, -(r * z));
  R[2] = 2.0f * fmaf(x, z, r * y);
  R[3] = 2.0f * fmaf(x, y, r * z);
  R[4] = 1.0f - 2.0f * fmaf(x, x, z * z);
  R[5] = 2.0f * fmaf(y, z, -(r * x));
  R[6] = 2.0f * fmaf(x, z, -(r * y));
  R[7] = 2.0f * fmaf(y, z, r * x);
  R[8] = 1.0f - 2.0f * fmaf(x, x, y * y);
  float s[3] = {mod * sc[0], mod * sc[1], mod * sc[2]};
  float M[9];
  for (int i = 0; i < 3; ++i)
    for (int k = 0; k < 3; ++k) M[3 * i + k] = R[3 * i + k] * s[k];
  cov[0] = d3(M[0], M[0], M[1], M[1], M[2], M[2]);
  cov[1] = d3(M[0], M[3], M[1], M[4], M[2], M[5]);
  cov[2] = d3(M[0], M[6], M[1], M[7], M[2], M[8]);
  cov[3] = d3(M[3], M[3], M[4], M[4], M[5], M[5]);
  cov[4] = d3(M[3], M[6], M[4], M[7], M[5], M[8]);
  cov[5] = d3(M[6], M[6], M[7], M[7], M[8], M[8]);
}

/* Forward geometry of P Gaussians.  Outputs (all length P unless noted):
 *   radii       int32   radius, 0 when culled
 *   radius_raw  int32   radius before the rect-area cull (diagnostic)
 *   rect        int32x4 tile rect x0,y0,x1,y1
 *   depth_bits  uint32  bit pattern of view-space z
 *   xy          f32x2   pixel centre
 *   tiles       uint32  tiles touched (0 when culled)
 * scales/rots may be NULL when cov3d is given.  Returns sum of tiles.                */
int64_t bm_preprocess(int64_t P, const float* means, const float* scales, float mod, const float* rots,
                      const float* cov3d, const float* vm, const float* pm, int W, int H, float tanfovx,
                      float tanfovy, int antialiasing, int32_t* radii, int32_t* radius_raw, int32_t* rect,
                      uint32_t* depth_bits, float* xy, uint32_t* tiles) {
  (void)antialiasing; /* h does not influence any integer output */
  const int gx = (W + BM_TILE - 1) / BM_TILE, gy = (H + BM_TILE - 1) / BM_TILE;
  const float fx = (float)W / (2.0f * tanfovx), fy = (float)H / (2.0f * tanfovy);
  const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
  int64_t total = 0;
  for (int64_t i = 0; i < P; ++i) {
    radii[i] = 0;
    radius_raw[i] = 0;
    tiles[i] = 0;
    depth_bits[i] = 0;
    xy[2 * i] = xy[2 * i + 1] = 0.f;
    rect[4 * i] = rect[4 * i + 1] = rect[4 * i + 2] = rect[4 * i + 3] = 0;
    const float x = means[3 * i], y = means[3 * i + 1], z = means[3 * i + 2];
    const float vx = row(vm, 0, x, y, z), vy = row(vm, 1, x, y, z), vz = row(vm, 2, x, y, z);
    if (!(vz > 0.2f)) continue;
    const float hx = row(pm, 0, x, y, z), hy = row(pm, 1, x, y, z), hw = row(pm, 3, x, y, z);
    const float pw = 1.0f / (hw + 0.0000001f);
    const float nx = hx * pw, ny = hy * pw;
    float cov[6];
    if (cov3d)
      memcpy(cov, cov3d + 6 * i, sizeof(cov));
    else
      quat_to_cov(scales + 3 * i, mod, rots + 4 * i, cov);
    float txtz = vx / vz, tytz = vy / vz;
    float cxx = fminf(limx, fmaxf(-limx, txtz)), cyy = fminf(limy, fmaxf(-limy, tytz));
    float tx = cxx * vz, ty = cyy * vz;
    float tz2 = vz * vz;
    float J00 = fx / vz, J02 = -((fx * tx) / tz2), J11 = fy / vz, J12 = -((fy * ty) / tz2);
    float M0[3], M1[3];
    for (int k = 0; k < 3; ++k) {
      M0[k] = fmaf(J00, vm[4 * k + 0], J02 * vm[4 * k + 2]);
      M1[k] = fmaf(J11, vm[4 * k + 1], J12 * vm[4 * k + 2]);
    }
    float v0[3], v1[3];
    v0[0] = d3(cov[0], M0[0], cov[1], M0[1], cov[2], M0[2]);
    v0[1] = d3(cov[1], M0[0], cov[3], M0[1], cov[4], M0[2]);
    v0[2] = d3(cov[2], M0[0], cov[4], M0[1], cov[5], M0[2]);
    v1[0] = d3(cov[0], M1[0], cov[1], M1[1], cov[2], M1[2]);
    v1[1] = d3(cov[1], M1[0], cov[3], M1[1], cov[4], M1[2]);
    v1[2] = d3(cov[2], M1[0], cov[4], M1[1], cov[5], M1[2]);
    float a0 = d3(M0[0], v0[0], M0[1], v0[1], M0[2], v0[2]);
    float b = d3(M0[0], v1[0], M0[1], v1[1], M0[2], v1[2]);
    float c0 = d3(M1[0], v1[0], M1[1], v1[1], M1[2], v1[2]);
    float a = a0 + 0.3f, c = c0 + 0.3f;
    float det = fmaf(a, c, -(b * b));
    if (det == 0.0f || !(det == det)) continue; /* non-finite covariance: culled (pdgs_math.cuh forward_geometry) */
    float mid = 0.5f * (a + c);
    float disc = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
    float lam = fmaxf(mid + disc, mid - disc);
    if (!(lam == lam)) continue;
    int rad = trunc_i(ceilf(3.0f * sqrtf(lam)));
    if (rad <= 0) continue;
    float px = pix_of_ndc(nx, W), py = pix_of_ndc(ny, H);
    float rf = (float)rad;
    int x0 = clampi(trunc_i((px - rf) / 16.0f), 0, gx);
    int y0 = clampi(trunc_i((py - rf) / 16.0f), 0, gy);
    int x1 = clampi(trunc_i((((px + rf) + 16.0f) - 1.0f) / 16.0f), 0, gx);
    int y1 = clampi(trunc_i((((py + rf) + 16.0f) - 1.0f) / 16.0f), 0, gy);
    radius_raw[i] = rad;
    rect[4 * i] = x0;
    rect[4 * i + 1] = y0;
    rect[4 * i + 2] = x1;
    rect[4 * i + 3] = y1;
    xy[2 * i] = px;
    xy[2 * i + 1] = py;
    memcpy(&depth_bits[i], &vz, 4);
    if (!(x1 > x0 && y1 > y0)) continue;
    int area = (x1 - x0) * (y1 - y0);
    radii[i] = rad;
    tiles[i] = (uint32_t)area;
    total += area;
  }
  return total;
}

typedef struct {
  uint64_t key;
  uint32_t val;
} bm_pair;

static int cmp_pair(const void* pa, const void* pb) {
  const bm_pair* a = (const bm_pair*)pa;
  const bm_pair* b = (const bm_pair*)pb;
  if (a->key != b->key) return a->key < b->key ? -1 : 1;
  if (a->val != b->val) return a->val < b->val ? -1 : 1; /* stable: emission order is ascending id */
  return 0;
}

/* A.6: emit (tile<<32 | depth_bits, id) for every tile of every visible rect (y outer, x inner,
 * ids ascending), sort stably by key, and derive ranges[T][2].  keys/vals have length R.   */
int bm_bin(int64_t P, const int32_t* radii, const int32_t* rect, const uint32_t* depth_bits, int gx, int gy,
           int64_t R, uint64_t* keys, uint32_t* vals, uint32_t* ranges) {
  bm_pair* prs = (bm_pair*)malloc(sizeof(bm_pair) * (size_t)(R > 0 ? R : 1));
  if (!prs) return -1;
  int64_t n = 0;
  for (int64_t i = 0; i < P; ++i) {
    if (radii[i] <= 0) continue;
    for (int y = rect[4 * i + 1]; y < rect[4 * i + 3]; ++y)
      for (int x = rect[4 * i]; x < rect[4 * i + 2]; ++x) {
        if (n >= R) {
          free(prs);
          return -2;
        }
        prs[n].key = ((uint64_t)(uint32_t)(y * gx + x) << 32) | depth_bits[i];
        prs[n].val = (uint32_t)i;
        ++n;
      }
  }
  if (n != R) {
    free(prs);
    return -3;
  }
  qsort(prs, (size_t)n, sizeof(bm_pair), cmp_pair);
  memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * (size_t)gy);
  for (int64_t k = 0; k < n; ++k) {
    keys[k] = prs[k].key;
    vals[k] = prs[k].val;
    uint32_t t = (uint32_t)(prs[k].key >> 32);
    if (k == 0 || (uint32_t)(prs[k - 1].key >> 32) != t) ranges[2 * t] = (uint32_t)k;
    if (k == n - 1 || (uint32_t)(prs[k + 1].key >> 32) != t) ranges[2 * t + 1] = (uint32_t)(k + 1);
  }
  free(prs);
  return 0;
}
