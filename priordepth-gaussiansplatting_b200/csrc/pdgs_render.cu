// K6 forward alpha-blend (colour + inverse depth) and K7 backward blend.
// Replaces renderCUDA (forward/backward) of the reference rasterizer (source absent;
// SURVEY.md 2.3(b) K6/K7, A.7, A.8).
//
// One CTA per 16x16 tile, 8 warps, each warp owns an 8x4 pixel patch.  Batches of 256 list
// entries are staged in shared memory as three float4 records.  Both kernels are
// instruction-issue bound (ncu r01b: issue active 87 %), so the design removes instructions:
//
//  * Hierarchical culling.  The thread that stages a Gaussian also tests its alpha >= 1/255
//    footprint (the ellipse 0.5 d^T Q d <= ln(255 o)) against the eight warp patches of the tile
//    (exact minimum of the quadratic over each 8x4 pixel box, with a conservative margin) and
//    stores an 8-bit patch mask.  Each warp ballots its bit and visits only the survivors.  A
//    culled Gaussian would have been rejected by every pixel of the patch, so images, n_contrib
//    and gradients are unchanged; list positions stay those of the un-culled list.  The forward
//    leaves each pair's mask in the binning buffer (pair_mask); the backward reads it back instead
//    of recomputing it and skips the record gathers of pairs that reach no patch (35 % on C4).
//    Computing the masks earlier, in duplicateWithKeys, was measured and is slower overall: here
//    the arithmetic fills issue slots that the staging phase leaves idle.
//  * Backward in two phases per warp.  Phase 1 (lane = pixel) walks the survivors back to front
//    and, per (pixel, Gaussian), produces just two scalars: w = alpha*T and z = o*G*dL/dalpha.  The
//    state carried between entries is T and one scalar (sum of w_j q_j of the entries behind), not
//    the reference's per-channel accum_rec.  (z, w) go to a per-warp [8 Gaussians][32 pixels]
//    shared-memory panel (rows padded to 280 bytes: conflict-free for lane = pixel and lane = row).
//    Phase 2 (lane = Gaussian x pixel row of the patch) turns a full panel into ten per-Gaussian
//    sums -- the moments sum z, z dx, z dy, z dx^2, z dx dy, z dy^2 about the Gaussian's own centre
//    (taken raw about the patch origin with compile-time pixel offsets, then shifted once) and
//    sum w*dL/dC_c, sum w*dL/dD -- then issues three red.global.add.v4.f32.  The moments are
//    additive over pixels and tiles; the fused preprocess backward (K8+K9) converts them to
//    dL/dmean2D, dL/dconic, dL/dopacity.  This replaces fifty shuffles per (warp, Gaussian).
//    Measured on C4 (B200): 8-row panels 0.86 ms vs 16-row panels 1.02 ms (smaller shared footprint
//    leaves more L1 for the gathers); fetching the next survivor's records early and 5-6 CTAs/SM were
//    both slower.  Shared memory per CTA decides the L1 size through the carveout: the same kernel padded
//    to 43 / 51 KB runs at 0.885 / 1.045 ms, so the panels are kept small (34.6 KB -> 164 KB carveout).
//    Double-buffered staging with one barrier per batch was also slower (forward 0.60 vs 0.55 ms).
//  * alpha = ex2(log2 o + log2e * power) with the scaled conic and log2 o staged once per entry, by ONE
//    shared function in a fixed operation order so that forward and backward take bit-identical
//    accept/reject decisions.
#include "pdgs_internal.cuh"

namespace pdgs {

namespace {

constexpr int RT = 256;  // threads per tile CTA
constexpr int NWARP = RT / 32;
constexpr float LOG2E = 1.4426950408889634f;
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Exponent of alpha in base 2: t = log2(o) + log2(e) * (-0.5*(A dx^2 + C dy^2) - B dx dy), fixed order.  The staged
// coefficients are pre-scaled (hA = -0.5 A log2e, nB = -B log2e, hC = -0.5 C log2e, L = log2 o), so
// alpha = ex2(t) costs no multiply by log2(e) or by the opacity, and "power > 0" is "t > L".
__device__ __forceinline__ float gauss_exponent(float dx, float dy, float hA, float nB, float hC, float L) {
  const float u = __fmaf_rn(hA, dx, __fmul_rn(nB, dy));
  const float v = __fmaf_rn(__fmul_rn(hC, dy), dy, L);
  return __fmaf_rn(dx, u, v);
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Mask of the warp patches (bit = warp index = row*2 + col; 2 columns of 8 px, 4 rows of 4 px) of the
// tile at (tx0, ty0) that the alpha >= 1/255 footprint of a Gaussian can reach.  Conservative: the
// threshold carries a margin far above the fp32 error of the per-pixel evaluation, and any
// non-finite / non-positive-definite input keeps every patch.
__device__ __forceinline__ uint32_t patch_mask(float gx, float gy, float A, float B, float C, float o, float log2o,
                                               float tx0, float ty0) {
  const float o255 = 255.0f * o;
  if (!(o255 >= 0.999f)) return (o255 == o255) ? 0u : 0xFFu;  // alpha < 1/255 everywhere
  if (!(A > 0.f) || !(C > 0.f)) return 0xFFu;
  // ln(255 o) = (log2 o + log2 255) ln 2, with the margin
  const float tau = fmaf((log2o + 7.994353436858858f) * 0.6931471805599453f, 1.002f, 0.02f);
  const float hA = 0.5f * A, hC = 0.5f * C;
  const float sx = -B * __fdividef(1.0f, A);  // argmin over dx on a row dy = const:  dx* = sx * dy
  const float sy = -B * __fdividef(1.0f, C);  // argmin over dy on a column dx = const
#ifndef PDGS_MASK_FAST
#define PDGS_MASK_FAST 1
#endif
#if PDGS_MASK_FAST
  // Per column box k: nearest x to the centre cx, and the x-only parts of the form at that x; per row box m likewise.
  // A box then costs two clamps and four FMAs.  The rounding margin is one absolute bound for the whole tile
  // (1e-5 of the largest magnitude the three terms can reach on it) instead of a relative one per box.
  float x0[2], cx[2], ak[2], bk[2], sk[2], y0[4], y1[4], cy[4], cm[4], dm[4], rm[4];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    x0[k] = (tx0 + 8.0f * k) - gx;  // pixel-centre box relative to the Gaussian
    cx[k] = fminf(fmaxf(0.f, x0[k]), x0[k] + 7.0f);
    bk[k] = B * cx[k];
    ak[k] = hA * cx[k] * cx[k];
    sk[k] = sy * cx[k];
  }
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    y0[m] = (ty0 + 4.0f * m) - gy;
    y1[m] = y0[m] + 3.0f;
    cy[m] = fminf(fmaxf(0.f, y0[m]), y1[m]);
    dm[m] = B * cy[m];
    cm[m] = hC * cy[m] * cy[m];
    rm[m] = sx * cy[m];
  }
  const float X = fmaxf(fabsf(x0[0]), fabsf(x0[1] + 7.0f)), Y = fmaxf(fabsf(y0[0]), fabsf(y1[3]));
  const float taum = fmaf(1e-5f, fmaf(hA * X, X, fmaf(fabsf(B) * X, Y, hC * Y * Y)), tau);
  uint32_t mask = 0;
#pragma unroll
  for (int m = 0; m < 4; ++m) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      // the box minimum of a convex quadratic centred outside the box lies on an edge facing the centre
      const float dyv = fminf(fmaxf(sk[k], y0[m]), y1[m]);
      const float q1 = fmaf(dyv, fmaf(hC, dyv, bk[k]), ak[k]);
      const float dxv = fminf(fmaxf(rm[m], x0[k]), x0[k] + 7.0f);
      const float q2 = fmaf(dxv, fmaf(hA, dxv, dm[m]), cm[m]);
      if (!(fminf(q1, q2) > taum)) mask |= 1u << (m * 2 + k);
    }
  }
#else
  float x0[2], cx[2], y0[4], cy[4];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    x0[k] = (tx0 + 8.0f * k) - gx;  // pixel-centre box relative to the Gaussian
    cx[k] = fminf(fmaxf(0.f, x0[k]), x0[k] + 7.0f);
  }
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    y0[m] = (ty0 + 4.0f * m) - gy;
    cy[m] = fminf(fmaxf(0.f, y0[m]), y0[m] + 3.0f);
  }
  uint32_t mask = 0;
#pragma unroll
  for (int m = 0; m < 4; ++m) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      // the box minimum of a convex quadratic centred outside the box lies on an edge facing the centre
      const float dyv = fminf(fmaxf(sy * cx[k], y0[m]), y0[m] + 3.0f);
      const float t1 = hA * cx[k] * cx[k], t2 = B * cx[k] * dyv, t3 = hC * dyv * dyv;
      const float q1 = (t1 + t2 + t3) - 1e-5f * (t1 + fabsf(t2) + t3);
      const float dxv = fminf(fmaxf(sx * cy[m], x0[k]), x0[k] + 7.0f);
      const float u1 = hA * dxv * dxv, u2 = B * dxv * cy[m], u3 = hC * cy[m] * cy[m];
      const float q2 = (u1 + u2 + u3) - 1e-5f * (u1 + fabsf(u2) + u3);
      if (!(fminf(q1, q2) > tau)) mask |= 1u << (m * 2 + k);
    }
  }
#endif
  return mask;
}

struct Stage {          // 49 bytes per staged list entry
  float4 a[RT];         // x, y, -0.5*conA*log2e, -conB*log2e
  float4 b[RT];         // -0.5*conC*log2e, log2(opacity), 1/depth, r
  float4 c[RT];         // g, b, Gaussian id (bits), -
  uint8_t mask[RT];     // warp patches the Gaussian can reach
};

// Stage list entry `src` (or nothing when !valid) into slot threadIdx.x.  The forward (REUSE = false) computes the
// patch mask and leaves it in pair_mask[src]; the backward (REUSE = true) reads it back instead of recomputing it and
// does not gather the records of an entry that reaches no patch (35 % of the pairs on C4).
template <bool REUSE>
__device__ __forceinline__ void stage_entry(Stage& st, bool valid, uint32_t src, const float4* __restrict__ gA,
                                            const float4* __restrict__ gB, const float4* __restrict__ gC,
                                            const uint32_t* __restrict__ point_list, uint8_t* __restrict__ pair_mask,
                                            float tx0, float ty0) {
  uint32_t m = 0;
  if (valid) {
    if (REUSE) m = pair_mask[src];
    if (!REUSE || m != 0) {
      const uint32_t id = __ldg(point_list + src);
      const float4 a = __ldg(gA + id), b = __ldg(gB + id), c = __ldg(gC + id);
      const float L = __log2f(b.y);
      st.a[threadIdx.x] = make_float4(a.x, a.y, (-0.5f * LOG2E) * a.z, -LOG2E * a.w);
      st.b[threadIdx.x] = make_float4((-0.5f * LOG2E) * b.x, L, b.z, c.x);
      st.c[threadIdx.x] = make_float4(c.y, c.z, __uint_as_float(id), 0.f);
      if (!REUSE) {
        m = patch_mask(a.x, a.y, a.z, a.w, b.x, b.y, L, tx0, ty0);
        pair_mask[src] = (uint8_t)m;
      }
    }
  }
  st.mask[threadIdx.x] = (uint8_t)m;
}

// Survivor loop indexing.  The survivor mask of 32 staged entries is bit-reversed, so its highest set bit `hb` is the
// first remaining list entry j = 31 - hb of the block.  Written as PTX so that the address of that entry's record is
// ONE multiply-add from the address of the block's last record (last - 16*hb) instead of the three integer
// instructions the compiler derives from j (it canonicalises every C++ spelling back to 31 - hb, + block base, * 16).
__device__ __forceinline__ uint32_t highest_bit(uint32_t m) {
  uint32_t r;
  asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(m));
  return r;
}
// mask of the hb low bits: and-ing it into the survivor mask clears its highest set bit hb in two instructions (BMSK +
// LOP3) instead of the three of `rm ^= 1u << hb` (MOV 1, SHF, LOP3); the survivor loops are issue-bound
__device__ __forceinline__ uint32_t low_bits(uint32_t hb) {
  uint32_t r;
  asm("bmsk.clamp.b32 %0, 0, %1;" : "=r"(r) : "r"(hb));
  return r;
}
__device__ __forceinline__ uint32_t record_addr(uint32_t last_addr, uint32_t hb) {
  uint32_t r;
  asm("mad.lo.u32 %0, %1, 0xfffffff0, %2;" : "=r"(r) : "r"(hb), "r"(last_addr));
  return r;
}
__device__ __forceinline__ float4 lds128_nv(uint32_t a) {  // not volatile: the forward loop has no shared-memory stores
  float4 v;
  asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}

#ifndef PDGS_FWD_CTAS
#define PDGS_FWD_CTAS 4
#endif
#ifndef PDGS_FWD_NOBRANCH
#define PDGS_FWD_NOBRANCH 1
#endif
// COUNT: also produce the per-pixel work counters n_tested / n_accepted (roofline FLOP accounting of bench.py; enabled
// per handle with pdgs_set_work_counters).  The product path runs COUNT = false: the counters cost two to three of the
// ~42 instructions of the survivor loop.
template <bool COUNT>
__global__ void __launch_bounds__(RT, PDGS_FWD_CTAS) render_forward_kernel(const float4* __restrict__ gA, const float4* __restrict__ gB,
                                                            const float4* __restrict__ gC,
                                                            const uint2* __restrict__ ranges,
                                                            const uint32_t* __restrict__ point_list,
                                                            uint8_t* __restrict__ pair_mask, int W, int H,
                                                            const float* __restrict__ bg, float* __restrict__ out_color,
                                                            float* __restrict__ out_invdepth, float* __restrict__ final_T,
                                                            uint32_t* __restrict__ n_contrib,
                                                            uint32_t* __restrict__ n_tested,
                                                            uint32_t* __restrict__ n_accepted) {
  __shared__ Stage st;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int px = blockIdx.x * TILE + (warp & 1) * 8 + (lane & 7);
  const int py = blockIdx.y * TILE + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const float tx0 = (float)(blockIdx.x * TILE), ty0 = (float)(blockIdx.y * TILE);
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const uint32_t wbit = 1u << warp;
  const uint32_t st_addr = (uint32_t)__cvta_generic_to_shared(&st.a[0]);
  constexpr uint32_t FWD_ST_B = (uint32_t)offsetof(Stage, b), FWD_ST_C = (uint32_t)offsetof(Stage, c);
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last = 0, accepted = 0, tested = range.y - range.x;
  bool done = !inside;
  for (uint32_t base = range.x; base < range.y; base += RT) {
    if (__syncthreads_count(done) == RT) break;
    stage_entry<false>(st, base + threadIdx.x < range.y, base + threadIdx.x, gA, gB, gC, point_list, pair_mask, tx0, ty0);
    __syncthreads();
    if (__all_sync(FULL, done)) continue;  // this warp's patch is finished; it only helps staging
    const int cnt = min((uint32_t)RT, range.y - base);
    const uint32_t pos0 = base - range.x;
    // shared address of the record of the last entry blended / of the entry that ended the pixel in THIS batch: one move
    // inside the loop, the list position is formed from it once per batch.  "None" is 0xffffffff, which no 16-byte
    // aligned record address can be (0 could: nothing forbids the stage from starting at shared address 0).
    constexpr uint32_t NO_REC = 0xffffffffu;
    uint32_t hit_rec = NO_REC, end_rec = NO_REC;
    for (int k = 0; k * 32 < cnt; ++k) {
      // survivors of this warp's patch, bit-reversed so that the leading one is the first list entry
      uint32_t rm = __brev(__ballot_sync(FULL, (st.mask[k * 32 + lane] & wbit) != 0));
      const uint32_t last_rec = st_addr + 16u * (uint32_t)(k * 32 + 31);
      while (rm) {
        const uint32_t hb = highest_bit(rm);
        rm &= low_bits(hb);
        const uint32_t rec = record_addr(last_rec, hb);  // entry j = k*32 + 31 - hb
        const float4 a = lds128_nv(rec);
        const float4 b = lds128_nv(rec + FWD_ST_B);
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float t = gauss_exponent(dx, dy, a.z, a.w, b.x, b.y);
        const float alpha = fminf(ALPHA_MAX, ex2_approx(t));
#if PDGS_FWD_NOBRANCH
        // no divergent branch: every lane evaluates the blend, the updates are selects
        const bool skip = done || t > b.y || alpha < ALPHA_MIN;
        const float test_T = T * (1.0f - alpha);
        const bool ends = !skip && test_T < T_EPS;
        const bool ok = !skip && !ends;
        float2 c;
        asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c.x), "=f"(c.y) : "r"(rec + FWD_ST_C));
        const float w = alpha * T;
        C0 = ok ? fmaf(b.w, w, C0) : C0;
        C1 = ok ? fmaf(c.x, w, C1) : C1;
        C2 = ok ? fmaf(c.y, w, C2) : C2;
        D = ok ? fmaf(b.z, w, D) : D;
        T = ok ? test_T : T;
        hit_rec = ok ? rec : hit_rec;
        done = done || ends;
        if (COUNT) {
          if (ends) end_rec = rec;
          if (ok) ++accepted;
        }
#else
        if (done || t > b.y || alpha < ALPHA_MIN) continue;
        const float test_T = T * (1.0f - alpha);
        if (test_T < T_EPS) {
          done = true;
          if (COUNT) end_rec = rec;
          continue;
        }
        float2 c;
        asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c.x), "=f"(c.y) : "r"(rec + FWD_ST_C));
        const float w = alpha * T;
        C0 = fmaf(b.w, w, C0);
        C1 = fmaf(c.x, w, C1);
        C2 = fmaf(c.y, w, C2);
        D = fmaf(b.z, w, D);
        T = test_T;
        hit_rec = rec;
        if (COUNT) ++accepted;
#endif
      }
      if (__all_sync(FULL, done)) break;
    }
    if (hit_rec != NO_REC) last = pos0 + ((hit_rec - st_addr) >> 4) + 1u;
    if (COUNT && end_rec != NO_REC) tested = pos0 + ((end_rec - st_addr) >> 4) + 1u;
  }
  if (inside) {
    const size_t pix = (size_t)py * W + px, hw = (size_t)W * H;
    final_T[pix] = T;
    n_contrib[pix] = last;
    if (COUNT) {
      n_tested[pix] = tested;      // list entries a pixel-serial walk evaluates before this pixel stops
      n_accepted[pix] = accepted;
    }
    out_color[pix] = C0 + T * __ldg(bg + 0);
    out_color[hw + pix] = C1 + T * __ldg(bg + 1);
    out_color[2 * hw + pix] = C2 + T * __ldg(bg + 2);
    out_invdepth[pix] = D;
  }
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Explicit 32-bit shared-memory addressing for the backward hot loop.  Through C++ references the compiler
// rebuilt the CTA's shared window base (S2R SR_CgaCtaId + LEA) on every list entry (ncu r01t: 3 instructions and
// the longest stall of the loop); an opaque base register plus ld/st.shared with immediate offsets avoids that.
__device__ __forceinline__ uint32_t smem_base(const void* p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p), r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ float2 lds64(uint32_t a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts64(uint32_t a, float x, float y) {
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t x) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(x) : "memory"); }

#ifndef PDGS_BWD_FB
#define PDGS_BWD_FB 8
#endif
// measured (profiles/README.md r02B, C4, blend_bwd ms): rm ^= 1 << hb 0.834, BMSK form 0.826; panel-full test on the
// row pointer instead of a slot counter 0.856 (one instruction less per survivor, yet slower), so the counter stays
#ifndef PDGS_BWD_BMSK
#define PDGS_BWD_BMSK 1
#endif
#ifndef PDGS_BWD_SLOTPTR
#define PDGS_BWD_SLOTPTR 0
#endif
// r02C (same probe): per-pixel block without a divergent branch 0.826 -> 0.809; list-position test as one signed
// compare against a per-32-block limit 0.826 -> 0.820; both 0.803
#ifndef PDGS_BWD_NOBRANCH
#define PDGS_BWD_NOBRANCH 1
#endif
#ifndef PDGS_BWD_THR
#define PDGS_BWD_THR 1
#endif
#ifndef PDGS_BWD_CTAS
#define PDGS_BWD_CTAS 5
#endif
constexpr int FB = PDGS_BWD_FB;   // Gaussians per phase-2 panel (8 or 16)
constexpr int FPARTS = 32 / FB;   // lanes per panel row in phase 2; each covers FB pixels = FB/8 pixel rows of the patch
static_assert(FB == 8 || FB == 16, "panel rows");

// Per-warp panel in dynamic shared memory (byte offsets from the warp's base):
//   FB rows of PROW bytes: [0, 256) (z, w) of the warp's 32 pixels, [256, 264) centre of the row's Gaussian,
//   [264, 268) its Gaussian id; then 32 x float4 (dL/dC rgb, dL/dD) of the warp's pixels.
// PROW = 280 makes both access patterns conflict-free: lane = pixel writes 8-byte columns of one row, lane = row
// reads one column (row stride 70 words: 6 g mod 32 is distinct and even for g = 0..15; with 8 rows the second
// lane group of a half-warp sits 16 words further, also distinct).
constexpr uint32_t PROW = 280, PR_GXY = 256, PR_POS = 264;
constexpr uint32_t WP_GPIX = FB * PROW, WP_BYTES = WP_GPIX + 32 * 16;
constexpr uint32_t ST_B = (uint32_t)offsetof(Stage, b), ST_C = (uint32_t)offsetof(Stage, c);
constexpr uint32_t BWD_PANELS = (uint32_t)((sizeof(Stage) + 15) / 16 * 16), BWD_SMAX = BWD_PANELS + NWARP * WP_BYTES,
                   BWD_SMEM = BWD_SMAX + NWARP * 4;
static_assert(PROW % 8 == 0 && WP_BYTES % 16 == 0, "panel alignment");

// Phase 2: reduce `count` panel rows to per-Gaussian sums and add them to the global accumulator.
// wp = shared-memory address of the warp's panel.  Lane = (row g, pixel part); each lane takes raw moments of
// z over its FB pixels about the patch origin (the pixel offsets are compile-time constants, so a pixel costs
// 1 + 2 [+ 2] FMAs instead of ten operations), then shifts them to the Gaussian's centre once.
__device__ __forceinline__ void flush_panel(uint32_t wp, int count, int lane, float X0f, float Y0f,
                                            float* __restrict__ acc) {
  __syncwarp();
  const int g = lane & (FB - 1), part = lane / FB;
  const uint32_t row = wp + PROW * g;
  uint32_t id = 0;
  if (part == 0 && g < count) id = lds32(row + PR_POS);
  float Sz = 0.f, Szx = 0.f, Szy = 0.f, Szxx = 0.f, Szxy = 0.f, Szyy = 0.f, Sw0 = 0.f, Sw1 = 0.f, Sw2 = 0.f, Swd = 0.f;
  if (g < count) {
    float m0 = 0.f, mx = 0.f, mxx = 0.f, my = 0.f, mxy = 0.f;  // raw moments: x = 0..7, y = 0..FB/8-1 (y^2 = y)
#pragma unroll
    for (int i = 0; i < FB; ++i) {
      const int p = part * FB + i;
      const float2 zw = lds64(row + 8 * p);
      const float4 gp = lds128(wp + WP_GPIX + 16 * p);
      const float x = (float)(i & 7);
      m0 += zw.x;
      if ((i & 7) != 0) {
        mx = fmaf(zw.x, x, mx);
        mxx = fmaf(zw.x, x * x, mxx);
      }
      if (i >= 8) {
        my += zw.x;
        if ((i & 7) != 0) mxy = fmaf(zw.x, x, mxy);
      }
      Sw0 = fmaf(zw.y, gp.x, Sw0);
      Sw1 = fmaf(zw.y, gp.y, Sw1);
      Sw2 = fmaf(zw.y, gp.z, Sw2);
      Swd = fmaf(zw.y, gp.w, Swd);
    }
    // d = centre - pixel = (bx - x, by - y)
    const float2 gxy = lds64(row + PR_GXY);
    const float bx = gxy.x - X0f;
    const float by = (gxy.y - Y0f) - (float)((FB / 8) * part);
    Sz = m0;
    Szx = fmaf(bx, m0, -mx);
    Szy = fmaf(by, m0, -my);
    Szxx = fmaf(bx, fmaf(bx, m0, -2.f * mx), mxx);
    Szyy = fmaf(by, fmaf(by, m0, -2.f * my), my);
    Szxy = fmaf(bx, Szy, fmaf(-by, mx, mxy));
  }
#pragma unroll
  for (int sh = 16; sh >= FB; sh >>= 1) {
    Sz += __shfl_xor_sync(FULL, Sz, sh);
    Szx += __shfl_xor_sync(FULL, Szx, sh);
    Szy += __shfl_xor_sync(FULL, Szy, sh);
    Szxx += __shfl_xor_sync(FULL, Szxx, sh);
    Szxy += __shfl_xor_sync(FULL, Szxy, sh);
    Szyy += __shfl_xor_sync(FULL, Szyy, sh);
    Sw0 += __shfl_xor_sync(FULL, Sw0, sh);
    Sw1 += __shfl_xor_sync(FULL, Sw1, sh);
    Sw2 += __shfl_xor_sync(FULL, Sw2, sh);
    Swd += __shfl_xor_sync(FULL, Swd, sh);
  }
  if (part == 0 && g < count) {
    float* dst = acc + (size_t)id * ACC_FLOATS;
    red_add_v4(dst, Szx, Szy, Szxx, Szxy);
    red_add_v4(dst + 4, Szyy, Sz, Swd, 0.f);
    red_add_v4(dst + 8, Sw0, Sw1, Sw2, 0.f);
  }
  __syncwarp();
}

__global__ void __launch_bounds__(RT, PDGS_BWD_CTAS) render_backward_kernel(const float4* __restrict__ gA, const float4* __restrict__ gB,
                                                                const float4* __restrict__ gC,
                                                                const uint2* __restrict__ ranges,
                                                                const uint32_t* __restrict__ point_list,
                                                                uint8_t* __restrict__ pair_mask, int W, int H,
                                                                const float* __restrict__ bg,
                                                                const float* __restrict__ final_T,
                                                                const uint32_t* __restrict__ n_contrib,
                                                                const float* __restrict__ dL_dcolor,
                                                                const float* __restrict__ dL_dinvdepth,
                                                                float* __restrict__ acc) {
  extern __shared__ __align__(16) unsigned char bwd_smem[];
  Stage& st = *reinterpret_cast<Stage*>(bwd_smem);
  uint32_t* s_max = reinterpret_cast<uint32_t*>(bwd_smem + BWD_SMAX);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sst = smem_base(bwd_smem), wp = sst + BWD_PANELS + WP_BYTES * warp;
  const int X0 = blockIdx.x * TILE + (warp & 1) * 8, Y0 = blockIdx.y * TILE + (warp >> 1) * 4;
  const int px = X0 + (lane & 7), py = Y0 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const float tx0 = (float)(blockIdx.x * TILE), ty0 = (float)(blockIdx.y * TILE);
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
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(wp + WP_GPIX + 16 * lane), "f"(gp0), "f"(gp1), "f"(gp2),
               "f"(gpd)
               : "memory");
  // the tile only needs list positions below the largest n_contrib of its pixels, each warp below its own
  const uint32_t wmax = __reduce_max_sync(FULL, last_contributor);
  if (lane == 0) s_max[warp] = wmax;
  __syncthreads();
  uint32_t todo = 0;
#pragma unroll
  for (int w = 0; w < NWARP; ++w) todo = max(todo, s_max[w]);
  todo = min(todo, range.y - range.x);
  const uint32_t wbit = 1u << warp;
  const float X0f = (float)X0, Y0f = (float)Y0;

  // Back to front.  With q_i = c_i . dL/dC + (1/z_i) dL/dD and w_i = alpha_i T_i, the reference's recurrence
  //   dL/dalpha_i = T_i (c_i - accum_rec_i) . dL/dC - T_final/(1 - alpha_i) (bg . dL/dC)
  // equals T_i q_i + (bgterm - sum_{j behind i} w_j q_j) / (1 - alpha_i), so the state carried from one entry to
  // the next is T and ONE scalar `behind` (four FMAs per pair instead of the per-channel accum_rec updates).
  float T = T_final;
  float behind = -T_final * (__ldg(bg + 0) * gp0 + __ldg(bg + 1) * gp1 + __ldg(bg + 2) * gp2);
#if !PDGS_BWD_SLOTPTR
  int slot = 0;
#endif
  uint32_t wrow = wp;  // shared address of the panel row being filled

  for (uint32_t done = 0; done < todo; done += RT) {
    __syncthreads();
    const uint32_t progress = done + threadIdx.x;
    stage_entry<true>(st, progress < todo, range.x + (todo - 1 - progress), gA, gB, gC, point_list, pair_mask, tx0, ty0);
    __syncthreads();
    const int cnt = min((uint32_t)RT, todo - done);
    const uint32_t top = todo - 1 - done;  // list position of staged entry 0 (entries go back to front)
    // staged entry j sits at list position top - j; this warp needs positions < wmax only, i.e. j >= jmin
    const int jmin = (int)top - (int)wmax + 1;
    if (jmin >= cnt) continue;
    for (int k = 0; k * 32 < cnt; ++k) {
      uint32_t mw = __ballot_sync(FULL, (st.mask[k * 32 + lane] & wbit) != 0);
      const int lo = jmin - k * 32;
      if (lo >= 32) continue;
      if (lo > 0) mw &= 0xffffffffu << lo;
      uint32_t rm = __brev(mw);  // leading one = first entry to visit
      const uint32_t last_rec = sst + 16u * (uint32_t)(k * 32 + 31);
      const uint32_t pos_last = top - (uint32_t)(k * 32 + 31);  // list position of the block's last entry (may wrap: only used + hb)
#if PDGS_BWD_THR
      // pos = pos_last + hb < last_contributor  <=>  (int)hb < (int)(last_contributor - pos_last): one compare per survivor
      const int hb_limit = (int)(last_contributor - pos_last);
#endif
      while (rm) {
        const uint32_t hb = highest_bit(rm);  // entry j = k*32 + 31 - hb
#if PDGS_BWD_BMSK
        rm &= low_bits(hb);
#else
        rm ^= 1u << hb;
#endif
        const uint32_t rec = record_addr(last_rec, hb);
        const float4 a = lds128(rec), b = lds128(rec + ST_B);
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float t = gauss_exponent(dx, dy, a.z, a.w, b.x, b.y);
        const float araw = ex2_approx(t);  // o * G
        const float alpha = fminf(ALPHA_MAX, araw);
#if PDGS_BWD_THR
        const bool active = ((int)hb < hb_limit) && !(t > b.y) && !(alpha < ALPHA_MIN);
#else
        const uint32_t pos = pos_last + hb;  // = top - j
        const bool active = (pos < last_contributor) && !(t > b.y) && !(alpha < ALPHA_MIN);
#endif
        if (__any_sync(FULL, active)) {
          float z = 0.f, w = 0.f;
          const float4 c = lds128(rec + ST_C);  // uniform address: one broadcast wavefront
          const uint32_t row_id = __float_as_uint(c.z);
#if PDGS_BWD_NOBRANCH
          {  // no divergent branch: every lane evaluates, inactive lanes keep their state and store zeros
            const float inv = rcp_approx(1.0f - alpha);  // 1 - alpha >= 0.01
            const float Tn = T * inv;
            const float wv = alpha * Tn;
            float q = b.w * gp0;
            q = fmaf(c.x, gp1, q);
            q = fmaf(c.y, gp2, q);
            q = fmaf(b.z, gpd, q);
            const float dL_dalpha = fmaf(Tn, q, behind * inv);
            const float bn = fmaf(-wv, q, behind);
            T = active ? Tn : T;
            behind = active ? bn : behind;
            w = active ? wv : 0.f;
            z = active ? araw * dL_dalpha : 0.f;
          }
#else
          if (active) {
            const float inv = rcp_approx(1.0f - alpha);  // 1 - alpha >= 0.01
            T *= inv;
            w = alpha * T;
            float q = b.w * gp0;
            q = fmaf(c.x, gp1, q);
            q = fmaf(c.y, gp2, q);
            q = fmaf(b.z, gpd, q);
            const float dL_dalpha = fmaf(T, q, behind * inv);
            behind = fmaf(-w, q, behind);
            z = araw * dL_dalpha;  // = o * G * dL/dalpha; K8/K9 takes the factor o into account
          }
#endif
          sts64(wrow + 8 * lane, z, w);
          sts64(wrow + PR_GXY, a.x, a.y);  // every lane stores the same words: no lane predicate, no conflict
          sts32(wrow + PR_POS, row_id);
          wrow += PROW;
#if PDGS_BWD_SLOTPTR
          if (wrow == wp + FB * PROW) {  // panel full
            flush_panel(wp, FB, lane, X0f, Y0f, acc);
            wrow = wp;
          }
#else
          if (++slot == FB) {
            flush_panel(wp, FB, lane, X0f, Y0f, acc);
            slot = 0;
            wrow = wp;
          }
#endif
        }
      }
    }
  }
#if PDGS_BWD_SLOTPTR
  if (wrow != wp) flush_panel(wp, (int)((wrow - wp) / PROW), lane, X0f, Y0f, acc);
#else
  if (slot > 0) flush_panel(wp, slot, lane, X0f, Y0f, acc);
#endif
}

}  // namespace

cudaError_t launch_render_forward(const DeviceArgs& a, GeomView g, ImageView im, const uint32_t* point_list,
                                  uint8_t* pair_mask, int64_t capacity, float* out_color, float* out_invdepth,
                                  int work_counters, cudaStream_t s) {
  (void)capacity;
  if (a.W <= 0 || a.H <= 0) return cudaSuccess;
  dim3 grid(a.grid_x, a.grid_y);
  if (work_counters)
    render_forward_kernel<true><<<grid, RT, 0, s>>>(g.gA, g.gB, g.gC, im.ranges, point_list, pair_mask, a.W, a.H, a.bg,
                                                    out_color, out_invdepth, im.final_T, im.n_contrib, im.n_tested,
                                                    im.n_accepted);
  else
    render_forward_kernel<false><<<grid, RT, 0, s>>>(g.gA, g.gB, g.gC, im.ranges, point_list, pair_mask, a.W, a.H, a.bg,
                                                     out_color, out_invdepth, im.final_T, im.n_contrib, im.n_tested,
                                                     im.n_accepted);
  return cudaGetLastError();
}

cudaError_t launch_render_backward(const DeviceArgs& a, GeomView g, ImageView im, const uint32_t* point_list,
                                   uint8_t* pair_mask, const float* dL_dcolor, const float* dL_dinvdepth, float* acc, cudaStream_t s) {
  if (a.W <= 0 || a.H <= 0) return cudaSuccess;
  // 49.3 KB of dynamic shared memory per CTA (above the 48 KB static limit): opt in once per device
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (configured_device != dev) {
    e = cudaFuncSetAttribute(render_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BWD_SMEM);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  dim3 grid(a.grid_x, a.grid_y);
  render_backward_kernel<<<grid, RT, BWD_SMEM, s>>>(g.gA, g.gB, g.gC, im.ranges, point_list, pair_mask, a.W, a.H, a.bg,
                                                    im.final_T, im.n_contrib, dL_dcolor, dL_dinvdepth, acc);
  return cudaGetLastError();
}

}  // namespace pdgs
