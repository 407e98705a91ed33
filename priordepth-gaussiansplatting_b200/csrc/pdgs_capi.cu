// C ABI of libpdgs.so (declared in include/pdgs.h) and the host-side orchestration that
// plays the role of the reference's rasterize_points.cu + rasterizer_impl.cu (source absent;
// SURVEY.md section 8(a) rows a5/a6, Appendix A.1).  Differences by design:
//   * no blocking device->host copy in the middle of forward: the binning/sort/blend kernels
//     are enqueued speculatively behind a device-side capacity guard and the host waits only
//     for the 8-byte pair count (event after the scan);
//   * everything runs on the caller's stream;
//   * no CPU fallback of any kind: without a CUDA device every entry point fails.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "../../include/pdgs.h"
#include "pdgs_internal.cuh"

using namespace pdgs;

struct StageRec { int stage; cudaEvent_t b, e; };

struct PdgsHandle {
  int device;
  unsigned long long* host_slot;  // pinned: [0] = num_rendered | overflow << 32
  double* dsums;                  // device double[8]: loss sums accumulated in double by the SSIM kernels
  cudaEvent_t ev;
  int timing;                     // per-stage events enabled
  int work_counters;              // forward blend also writes n_tested / n_accepted
  std::vector<cudaEvent_t> pool;  // idle timing events
  std::vector<StageRec> recs;     // recorded since the last pdgs_stage_times()
};

static cudaEvent_t take_event(PdgsHandle* h) {
  if (!h->pool.empty()) { cudaEvent_t e = h->pool.back(); h->pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

// Brackets one stage with a fresh event pair on the launching stream when timing is on; the pairs are
// resolved (and recycled) only in pdgs_stage_times(), so timing adds no host synchronisation.
struct StageTimer {
  PdgsHandle* h; cudaStream_t s; StageRec r;
  StageTimer(PdgsHandle* h_, int stage_, cudaStream_t s_) : h(h_), s(s_) {
    r.stage = stage_; r.b = r.e = nullptr;
    if (h->timing) { r.b = take_event(h); r.e = take_event(h); cudaEventRecord(r.b, s); }
  }
  ~StageTimer() {
    if (r.b) { cudaEventRecord(r.e, s); h->recs.push_back(r); }
  }
};

static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, const char* detail = "") {
  snprintf(g_err, sizeof(g_err), fmt, detail);
  return code;
}
#define PD_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t _e = (call);                                                 \
    if (_e != cudaSuccess) {                                                 \
      snprintf(g_err, sizeof(g_err), "%s failed: %s", #call, cudaGetErrorString(_e)); \
      return PDGS_E_CUDA;                                                    \
    }                                                                        \
  } while (0)

// ---- scratch for pdgs_sort_pairs_static ------------------------------------------------------------------------------
// Some hosts of this library may not allocate device memory while they run: a renderer embedded in an engine whose
// memory manager owns the whole device budget, a plugin called from inside somebody else's frame loop where a
// cudaMalloc (a device-wide synchronisation) would stall every other stream, or a process that handed all of its
// memory to a pool up front and has nothing left to ask the driver for.  For them the sort's working memory lives
// in the library image itself: one fixed array per GPU, resident from the moment the library is loaded on that GPU
// until the process exits, whether or not the entry point is ever called.
//
// The size follows carve_sort_scratch() for STATIC_SORT_MAX_PAIRS pairs in tiles of SORT_THREADS * SORT_ITEMS_U64:
// 8 digit histograms, 8 tickets, 8 passes of per-tile look-back words, and the 256-byte tail that holds the count.
constexpr int64_t STATIC_SORT_MAX_PAIRS = PDGS_SORT_STATIC_MAX_PAIRS;  // include/pdgs.h: 2^20
constexpr size_t STATIC_SORT_TILES = (STATIC_SORT_MAX_PAIRS + SORT_THREADS * SORT_ITEMS_U64 - 1) / (SORT_THREADS * SORT_ITEMS_U64);
constexpr size_t STATIC_SORT_BYTES = MAX_PASSES * RADIX * sizeof(uint32_t)                        // histograms
                                     + 128                                                         // tickets
                                     + MAX_PASSES * STATIC_SORT_TILES * RADIX * sizeof(uint32_t)   // look-back words
                                     + 256;                                                        // pair count
static __device__ __align__(256) unsigned char g_static_sort_scratch[STATIC_SORT_BYTES];  // ~2.7 MiB per GPU

// In host code `g_static_sort_scratch` is only a handle the runtime knows the array by; the address a kernel can use
// has to be asked for, per device, while that device is current.  Looked up once per device and kept.
constexpr int STATIC_MAX_DEVICES = 64;
static void* g_static_sort_ptr[STATIC_MAX_DEVICES] = {};
static std::mutex g_static_sort_mutex;

static int static_sort_scratch(void** out) {
  int dev = -1;
  PD_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) return fail(PDGS_E_INVALID, "pdgs_sort_pairs_static: device ordinal out of range");
  std::lock_guard<std::mutex> guard(g_static_sort_mutex);
  if (!g_static_sort_ptr[dev]) {
    void* p = nullptr;
    PD_CUDA(cudaGetSymbolAddress(&p, g_static_sort_scratch));
    if (!p) return fail(PDGS_E_CUDA, "pdgs_sort_pairs_static: cudaGetSymbolAddress returned NULL");
    g_static_sort_ptr[dev] = p;
  }
  *out = g_static_sort_ptr[dev];
  return PDGS_OK;
}

extern "C" {

const char* pdgs_last_error(void) { return g_err; }
const char* pdgs_version(void) { return "pdgs-b200 0.1 (sm_100a)"; }

int pdgs_create(PdgsHandle** out) {
  if (!out) return fail(PDGS_E_INVALID, "pdgs_create: out is NULL");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(PDGS_E_NO_DEVICE, "pdgs_create: no CUDA device (this library has no CPU path)");
  }
  PdgsHandle* h = new PdgsHandle();
  PD_CUDA(cudaGetDevice(&h->device));
  cudaDeviceProp prop;
  PD_CUDA(cudaGetDeviceProperties(&prop, h->device));
  if (prop.major != 10) {
    delete h;
    return fail(PDGS_E_NO_DEVICE, "pdgs_create: device is not sm_100 (kernels are built for sm_100a only)");
  }
  PD_CUDA(cudaHostAlloc((void**)&h->host_slot, 64, cudaHostAllocDefault));
  PD_CUDA(cudaEventCreateWithFlags(&h->ev, cudaEventDisableTiming));
  PD_CUDA(cudaMalloc((void**)&h->dsums, 64));
  h->timing = 0;
  h->work_counters = 0;
  *out = h;
  return PDGS_OK;
}

int pdgs_destroy(PdgsHandle* h) {
  if (!h) return PDGS_OK;
  cudaFreeHost(h->host_slot);
  cudaFree(h->dsums);
  cudaEventDestroy(h->ev);
  for (StageRec& r : h->recs) { cudaEventDestroy(r.b); cudaEventDestroy(r.e); }
  for (cudaEvent_t e : h->pool) cudaEventDestroy(e);
  delete h;
  return PDGS_OK;
}

int pdgs_set_stage_timing(PdgsHandle* h, int32_t enable) {
  if (!h) return fail(PDGS_E_INVALID, "pdgs_set_stage_timing: handle is NULL");
  h->timing = enable ? 1 : 0;
  for (StageRec& r : h->recs) { h->pool.push_back(r.b); h->pool.push_back(r.e); }
  h->recs.clear();
  return PDGS_OK;
}

int pdgs_set_work_counters(PdgsHandle* h, int32_t enable) {
  if (!h) return fail(PDGS_E_INVALID, "pdgs_set_work_counters: handle is NULL");
  h->work_counters = enable ? 1 : 0;
  return PDGS_OK;
}

int pdgs_stage_times(PdgsHandle* h, float* ms, int32_t* counts, int32_t n) {
  if (!h || !ms || n < PDGS_STAGE_COUNT) return fail(PDGS_E_INVALID, "pdgs_stage_times: bad argument");
  for (int i = 0; i < PDGS_STAGE_COUNT; ++i) { ms[i] = 0.f; if (counts) counts[i] = 0; }
  for (StageRec& r : h->recs) {
    float t = 0.f;
    PD_CUDA(cudaEventSynchronize(r.e));
    PD_CUDA(cudaEventElapsedTime(&t, r.b, r.e));
    ms[r.stage] += t;
    if (counts) counts[r.stage] += 1;
    h->pool.push_back(r.b); h->pool.push_back(r.e);
  }
  h->recs.clear();
  return PDGS_OK;
}

size_t pdgs_geom_bytes(int64_t P) { return carve_geom(nullptr, P).bytes; }
size_t pdgs_image_bytes(int32_t width, int32_t height) { return carve_image(nullptr, width, height).bytes; }
size_t pdgs_binning_bytes(int64_t capacity_pairs, int64_t P) {
  (void)P;
  return carve_binning(nullptr, capacity_pairs).bytes;
}
size_t pdgs_backward_scratch_bytes(int64_t P) { return align_up((size_t)(P > 0 ? P : 1) * ACC_FLOATS * sizeof(float)); }
size_t pdgs_sort_scratch_bytes(int64_t n) { return carve_sort_scratch(nullptr, n, SORT_THREADS * SORT_ITEMS_U64).bytes + 256; }

static bool aligned16(const void* p) { return (((uintptr_t)p) & 15u) == 0; }

static int make_args(const PdgsScene* sc, const PdgsCamera* cam, DeviceArgs& a) {
  if (!sc || !cam) return fail(PDGS_E_INVALID, "scene/camera is NULL");
  if (sc->P < 0) return fail(PDGS_E_INVALID, "P < 0");
  if (cam->image_width <= 0 || cam->image_height <= 0) return fail(PDGS_E_INVALID, "image size must be positive");
  // the reference traps on a culled point when prefiltered is set and its caller always passes False
  // (/root/reference/gaussian_renderer/__init__.py:47); refused here rather than silently ignored (include/pdgs.h)
  if (cam->prefiltered) return fail(PDGS_E_INVALID, "prefiltered=1 is not supported (the reference's caller always passes False)");
  if (sc->P > 0) {
    if (!sc->means3D || !sc->opacities) return fail(PDGS_E_INVALID, "means3D/opacities is NULL");
    const bool split = sc->shs_dc != nullptr;
    if ((int)(sc->shs != nullptr) + (int)split + (int)(sc->colors_precomp != nullptr) != 1)
      return fail(PDGS_E_INVALID, "exactly one of shs / (shs_dc, shs_rest) / colors_precomp must be given");
    if (!split && sc->shs_rest) return fail(PDGS_E_INVALID, "shs_rest given without shs_dc");
    if (split && sc->M > 1 && !sc->shs_rest) return fail(PDGS_E_INVALID, "shs_rest is NULL but M > 1");
    if (split && (!aligned16(sc->shs_dc) || (sc->shs_rest && !aligned16(sc->shs_rest))))
      return fail(PDGS_E_INVALID, "shs_dc / shs_rest must be 16-byte aligned");
    const bool sr = sc->scales != nullptr && sc->rotations != nullptr;
    if (sr == (sc->cov3D_precomp != nullptr) || ((sc->scales != nullptr) != (sc->rotations != nullptr)))
      return fail(PDGS_E_INVALID, "exactly one of (scales, rotations) / cov3D_precomp must be given");
    const bool any_sh = sc->shs || split;
    if (any_sh && (sc->M < 1 || sc->M > 16)) return fail(PDGS_E_INVALID, "SH coefficient count M must be in 1..16");
    if (any_sh && (cam->sh_degree < 0 || cam->sh_degree > 3 || (cam->sh_degree + 1) * (cam->sh_degree + 1) > sc->M))
      return fail(PDGS_E_INVALID, "sh_degree needs more coefficients than shs holds (or is outside 0..3)");
    if (sc->rotations && !aligned16(sc->rotations)) return fail(PDGS_E_INVALID, "rotations must be 16-byte aligned");
    if (sc->shs && sc->M == 16 && !aligned16(sc->shs)) return fail(PDGS_E_INVALID, "shs must be 16-byte aligned");
  }
  if (!cam->viewmatrix || !cam->projmatrix || !cam->bg) return fail(PDGS_E_INVALID, "viewmatrix/projmatrix/bg is NULL");
  if ((sc->shs || sc->shs_dc) && !cam->campos) return fail(PDGS_E_INVALID, "campos is NULL");
  a.shs_dc = sc->shs_dc; a.shs_rest = sc->shs_dc ? sc->shs_rest : nullptr; a.raw = sc->raw_params ? 1 : 0;
  a.means3D = sc->means3D; a.shs = sc->shs; a.colors_precomp = sc->colors_precomp; a.opacities = sc->opacities;
  a.scales = sc->scales; a.rotations = sc->rotations; a.cov3D_precomp = sc->cov3D_precomp;
  a.viewmatrix = cam->viewmatrix; a.projmatrix = cam->projmatrix; a.campos = cam->campos; a.bg = cam->bg;
  a.P = sc->P; a.M = (sc->shs || sc->shs_dc) ? sc->M : 0;
  a.tanfovx = cam->tanfovx; a.tanfovy = cam->tanfovy;
  a.W = cam->image_width; a.H = cam->image_height;
  a.focal_x = (float)a.W / (2.0f * cam->tanfovx);
  a.focal_y = (float)a.H / (2.0f * cam->tanfovy);
  a.scale_modifier = cam->scale_modifier;
  a.grid_x = (a.W + TILE - 1) / TILE; a.grid_y = (a.H + TILE - 1) / TILE;
  if (a.grid_x > 65535 || a.grid_y > 65535) return fail(PDGS_E_INVALID, "image larger than 65535 tiles per axis");
  a.sh_degree = cam->sh_degree; a.antialiasing = cam->antialiasing ? 1 : 0;
  return PDGS_OK;
}

#define PD_STEP(dbg, s, call)                                                                         \
  do {                                                                                                \
    PD_CUDA(call);                                                                                    \
    if (dbg) {                                                                                        \
      cudaError_t _e = cudaStreamSynchronize(s);                                                      \
      if (_e != cudaSuccess) {                                                                        \
        snprintf(g_err, sizeof(g_err), "debug: kernel behind %s failed: %s", #call, cudaGetErrorString(_e)); \
        return PDGS_E_CUDA;                                                                           \
      }                                                                                               \
    }                                                                                                 \
  } while (0)

// fused_scan: the scan of tiles_touched has not run yet; one kernel scans and emits the pairs, and the pair count is
// copied to the host behind it.  Otherwise (resume) point_offsets are valid and only the pairs are emitted.
static int run_binning_and_blend(PdgsHandle* h, const DeviceArgs& a, GeomView g, ImageView im, PdgsState* st,
                                 const int32_t* radii, const PdgsForwardOut* out, int debug, bool fused_scan, cudaStream_t s) {
  BinningView b = carve_binning(st->binning, st->capacity_pairs);
  if (b.bytes > st->binning_bytes) return fail(PDGS_E_INVALID, "binning buffer smaller than pdgs_binning_bytes(capacity)");
  const int tiles = a.grid_x * a.grid_y;
  const int tbits = tile_id_bits(tiles), npass = num_tile_sort_passes(tiles);
  uint32_t* kin = (npass % 2 == 0) ? b.tkeys_a : b.tkeys_b;
  uint32_t* vin = (npass % 2 == 0) ? b.point_list : b.vals_b;
  {
    StageTimer t(h, PDGS_STAGE_DUPLICATE, s);
    if (fused_scan) PD_STEP(debug, s, launch_scan_duplicate(a, g, kin, vin, st->capacity_pairs, s));
    else PD_STEP(debug, s, launch_duplicate(a, g, kin, vin, st->capacity_pairs, s));
  }
  if (fused_scan) {
    PD_CUDA(cudaMemcpyAsync(h->host_slot, g.ctrl, 8, cudaMemcpyDeviceToHost, s));
    PD_CUDA(cudaEventRecord(h->ev, s));
  }
  {
    StageTimer t(h, PDGS_STAGE_SORT, s);
    PD_STEP(debug, s, launch_sort_u32(b.tkeys_a, b.point_list, b.tkeys_b, b.vals_b, false, &g.ctrl->num_rendered, 0,
                                      st->capacity_pairs, npass, tbits, b.sort_scratch, s));
  }
  {
    StageTimer t(h, PDGS_STAGE_RANGES, s);
    PD_STEP(debug, s, launch_ranges(b.tkeys_a, &g.ctrl->num_rendered, st->capacity_pairs, im.ranges, tiles, s));
  }
  {
    StageTimer t(h, PDGS_STAGE_BLEND_FWD, s);
    PD_STEP(debug, s, launch_render_forward(a, g, im, b.point_list, b.pair_mask, st->capacity_pairs, out->color, out->invdepth,
                                            h->work_counters, s));
  }
  return PDGS_OK;
}

int pdgs_forward(PdgsHandle* h, const PdgsScene* scene, const PdgsCamera* cam, PdgsState* state,
                 const PdgsForwardOut* out, int32_t resume, int64_t* num_rendered, void* stream) {
  if (!h || !state || !out || !num_rendered) return fail(PDGS_E_INVALID, "pdgs_forward: NULL argument");
  DeviceArgs a;
  int rc = make_args(scene, cam, a);
  if (rc != PDGS_OK) return rc;
  if (!out->color || !out->invdepth || (a.P > 0 && !out->radii)) return fail(PDGS_E_INVALID, "output pointer is NULL");
  if (!state->geom || state->geom_bytes < pdgs_geom_bytes(a.P)) return fail(PDGS_E_INVALID, "geom buffer too small");
  if (!state->image || state->image_bytes < pdgs_image_bytes(a.W, a.H)) return fail(PDGS_E_INVALID, "image buffer too small");
  cudaStream_t s = (cudaStream_t)stream;
  GeomView g = carve_geom(state->geom, a.P);
  ImageView im = carve_image(state->image, a.W, a.H);
  const int debug = cam->debug;
  const bool have_binning = state->binning != nullptr && state->capacity_pairs > 0;

  if (!resume) {
    PD_CUDA(cudaMemsetAsync(g.ctrl, 0, sizeof(Ctrl), s));
    PD_CUDA(cudaMemsetAsync(g.scan_state, 0, ((size_t)(a.P > 0 ? a.P : 1) + DUP_TILE - 1) / DUP_TILE * 8, s));
    {
      StageTimer t(h, PDGS_STAGE_PREPROCESS, s);
      PD_STEP(debug, s, launch_preprocess(a, g, out->radii, s));
    }
    {
      StageTimer t(h, PDGS_STAGE_DEPTH_SORT, s);
      PD_STEP(debug, s, launch_sort_u32(g.depth_keys, g.order, g.depth_keys_b, g.order_b, true, nullptr, a.P, a.P,
                                        DEPTH_SORT_PASSES, 32, g.sort_scratch, s));
    }
    if (have_binning) {
      rc = run_binning_and_blend(h, a, g, im, state, out->radii, out, debug, true, s);
      if (rc != PDGS_OK) return rc;
    } else {  // sizing call: only the pair count is wanted
      {
        StageTimer t(h, PDGS_STAGE_SCAN, s);
        PD_STEP(debug, s, launch_scan(g, a.P, s));
      }
      PD_CUDA(cudaMemcpyAsync(h->host_slot, g.ctrl, 8, cudaMemcpyDeviceToHost, s));
      PD_CUDA(cudaEventRecord(h->ev, s));
    }
    PD_CUDA(cudaEventSynchronize(h->ev));
    const unsigned long long slot = h->host_slot[0];
    const uint32_t R = (uint32_t)(slot & 0xffffffffull);
    const uint32_t ovf = (uint32_t)(slot >> 32);
    *num_rendered = (int64_t)R;
    if (ovf) return fail(PDGS_E_OVERFLOW, "number of (tile, Gaussian) pairs >= 2^30");
    if (!have_binning || (int64_t)R > state->capacity_pairs) return PDGS_NEED_BINNING;
    return PDGS_OK;
  }
  // resume: geometry state is valid, R known; (re)run binning + blend with a big-enough buffer
  if (!have_binning) return fail(PDGS_E_INVALID, "resume without a binning buffer");
  if (*num_rendered > state->capacity_pairs) return fail(PDGS_E_INVALID, "resume: capacity < num_rendered");
  return run_binning_and_blend(h, a, g, im, state, out->radii, out, debug, false, s);
}

int pdgs_backward(PdgsHandle* h, const PdgsScene* scene, const PdgsCamera* cam, const PdgsState* state,
                  const int32_t* radii, int64_t num_rendered, const float* dL_dcolor, const float* dL_dinvdepth,
                  void* scratch, size_t scratch_bytes, const PdgsGrads* grads, void* stream) {
  if (!h || !state || !grads || !dL_dcolor) return fail(PDGS_E_INVALID, "pdgs_backward: NULL argument");
  DeviceArgs a;
  int rc = make_args(scene, cam, a);
  if (rc != PDGS_OK) return rc;
  if (a.P == 0) return PDGS_OK;
  if (!radii) return fail(PDGS_E_INVALID, "radii is NULL");
  if (!scratch || scratch_bytes < pdgs_backward_scratch_bytes(a.P)) return fail(PDGS_E_INVALID, "backward scratch too small");
  if (grads->rotations && !aligned16(grads->rotations)) return fail(PDGS_E_INVALID, "grad rotations must be 16-byte aligned");
  if (grads->shs && a.M == 16 && !aligned16(grads->shs)) return fail(PDGS_E_INVALID, "grad shs must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  GeomView g = carve_geom(state->geom, a.P);
  ImageView im = carve_image(state->image, a.W, a.H);
  const int debug = cam->debug;
  float* acc = (float*)scratch;
  {
    StageTimer t(h, PDGS_STAGE_BWD_ZERO, s);
    PD_CUDA(cudaMemsetAsync(acc, 0, (size_t)a.P * ACC_FLOATS * sizeof(float), s));
  }
  if (num_rendered > 0) {
    if (!state->binning) return fail(PDGS_E_INVALID, "binning buffer is NULL");
    BinningView b = carve_binning(state->binning, state->capacity_pairs);
    StageTimer t(h, PDGS_STAGE_BLEND_BWD, s);
    PD_STEP(debug, s, launch_render_backward(a, g, im, b.point_list, b.pair_mask, dL_dcolor, dL_dinvdepth, acc, s));
  }
  GradPtrs gp;
  gp.means3D = grads->means3D; gp.means2D = grads->means2D; gp.shs = grads->shs;
  gp.colors_precomp = grads->colors_precomp; gp.opacities = grads->opacities; gp.scales = grads->scales;
  gp.rotations = grads->rotations; gp.cov3D_precomp = grads->cov3D_precomp; gp.accumulate = grads->accumulate;
  gp.densify_grad_accum = grads->densify_grad_accum; gp.densify_denom = grads->densify_denom;
  gp.max_radii2D = grads->max_radii2D;
  gp.color_grad_view = (a.shs || a.shs_dc) ? grads->color_grad_view : nullptr;
  gp.shs_dc = a.shs_dc ? grads->shs_dc : nullptr; gp.shs_rest = a.shs_dc ? grads->shs_rest : nullptr;
  if (!a.shs) gp.shs = nullptr;
  if ((gp.shs_dc && !aligned16(gp.shs_dc)) || (gp.shs_rest && !aligned16(gp.shs_rest)))
    return fail(PDGS_E_INVALID, "grad shs_dc / shs_rest must be 16-byte aligned");
  {
    StageTimer t(h, PDGS_STAGE_PREPROCESS_BWD, s);
    PD_STEP(debug, s, launch_preprocess_backward(a, g, radii, acc, gp, s));
  }
  return PDGS_OK;
}

int pdgs_image_loss(PdgsHandle* h, const float* color, const float* target, const float* invdepth,
                    const float* prior, const float* mask, const float* alpha_mask, int32_t clamp01, int32_t width,
                    int32_t height, float w_color, float w_depth, float* dL_dcolor, float* dL_dinvdepth, float* loss,
                    void* stream) {
  if (!h || !color || !target || !dL_dcolor || !loss || width <= 0 || height <= 0)
    return fail(PDGS_E_INVALID, "pdgs_image_loss: bad argument");
  if (prior && (!invdepth || !dL_dinvdepth)) return fail(PDGS_E_INVALID, "pdgs_image_loss: prior given without invdepth / dL_dinvdepth");
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_LOSS, s);
  PD_CUDA(launch_image_loss(color, target, invdepth, prior, mask, alpha_mask, clamp01 ? 1 : 0, width, height, w_color,
                            w_depth, dL_dcolor, dL_dinvdepth, loss, s));
  return PDGS_OK;
}

size_t pdgs_ssim_scratch_bytes(int32_t width, int32_t height, int32_t planes) {
  if (width <= 0 || height <= 0 || planes <= 0) return 0;
  return 3 * align_up((size_t)planes * width * height * sizeof(float));
}

int pdgs_ssim_forward(PdgsHandle* h, float C1, float C2, const float* img1, const float* img2, int32_t planes,
                      int32_t width, int32_t height, float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq,
                      float* dm_dsigma12, float* ssim_sum, void* stream) {
  if (!h || !img1 || !img2 || planes < 0 || width <= 0 || height <= 0)
    return fail(PDGS_E_INVALID, "pdgs_ssim_forward: bad argument");
  const int nd = (dm_dmu1 != nullptr) + (dm_dsigma1_sq != nullptr) + (dm_dsigma12 != nullptr);
  if (nd != 0 && nd != 3) return fail(PDGS_E_INVALID, "pdgs_ssim_forward: give all three partial maps or none");
  if (planes > 65535) return fail(PDGS_E_INVALID, "pdgs_ssim_forward: more than 65535 planes");
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_LOSS, s);
  if (ssim_sum) PD_CUDA(cudaMemsetAsync(h->dsums, 0, 16, s));
  PD_CUDA(launch_ssim_forward(C1, C2, img1, img2, planes, width, height, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12,
                              ssim_sum ? h->dsums : nullptr, 0, nullptr, 0, s));
  if (ssim_sum && planes > 0) PD_CUDA(launch_ssim_finalize(h->dsums, 1.0, ssim_sum, nullptr, s));
  return PDGS_OK;
}

int pdgs_ssim_backward(PdgsHandle* h, const float* img1, const float* img2, const float* dL_dmap, float dL_dmap_uniform,
                       const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, int32_t planes,
                       int32_t width, int32_t height, float* dL_dimg1, void* stream) {
  if (!h || !img1 || !img2 || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1 || planes < 0 || width <= 0 ||
      height <= 0)
    return fail(PDGS_E_INVALID, "pdgs_ssim_backward: bad argument");
  if (planes > 65535) return fail(PDGS_E_INVALID, "pdgs_ssim_backward: more than 65535 planes");
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_LOSS, s);
  PD_CUDA(launch_ssim_backward(img1, img2, dL_dmap, dL_dmap_uniform, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, planes, width,
                               height, 0.f, dL_dimg1, nullptr, 0.0, nullptr, nullptr, nullptr, 0, s));
  return PDGS_OK;
}

int pdgs_image_loss_ssim(PdgsHandle* h, const float* color, const float* target, const float* invdepth,
                         const float* prior, const float* mask, const float* alpha_mask, int32_t clamp01, int32_t width,
                         int32_t height, float lambda_dssim, float w_depth, void* scratch, size_t scratch_bytes,
                         float* dL_dcolor, float* dL_dinvdepth, float* loss, void* stream) {
  if (!h || !color || !target || !dL_dcolor || !loss || width <= 0 || height <= 0)
    return fail(PDGS_E_INVALID, "pdgs_image_loss_ssim: bad argument");
  if (prior && (!invdepth || !dL_dinvdepth))
    return fail(PDGS_E_INVALID, "pdgs_image_loss_ssim: prior given without invdepth / dL_dinvdepth");
  if (!scratch || scratch_bytes < pdgs_ssim_scratch_bytes(width, height, 3))
    return fail(PDGS_E_INVALID, "pdgs_image_loss_ssim: scratch smaller than pdgs_ssim_scratch_bytes(width, height, 3)");
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_LOSS, s);
  const size_t n = (size_t)3 * width * height, seg = align_up(n * sizeof(float));
  float* d_mu = (float*)scratch;
  float* d_s1 = (float*)((char*)scratch + seg);
  float* d_s12 = (float*)((char*)scratch + 2 * seg);
  const float inv_n = 1.0f / (float)n;
  // partial maps into the scratch; sum of the SSIM map and sum |color - target| accumulated in double
  PD_CUDA(cudaMemsetAsync(h->dsums, 0, 16, s));
  const int cl = clamp01 ? 1 : 0;
  PD_CUDA(launch_ssim_forward(0.01f * 0.01f, 0.03f * 0.03f, color, target, 3, width, height, nullptr, d_mu, d_s1, d_s12,
                              h->dsums, 1, alpha_mask, cl, s));
  // d/dcolor [(1 - lambda) mean|c - t| + lambda (1 - mean ssim)]; its first block folds the sums into loss[2], loss[0]
  PD_CUDA(launch_ssim_backward(color, target, nullptr, -lambda_dssim * inv_n, d_mu, d_s1, d_s12, 3, width, height,
                               (1.0f - lambda_dssim) * inv_n, dL_dcolor, h->dsums, 1.0 / (double)n, loss + 2, loss + 0,
                               alpha_mask, cl, s));
  if (prior)
    PD_CUDA(launch_image_loss(nullptr, nullptr, invdepth, prior, mask, nullptr, 0, width, height, 0.f, w_depth, nullptr,
                              dL_dinvdepth, loss, s));
  return PDGS_OK;
}

int pdgs_sh_backward_views(PdgsHandle* h, const PdgsScene* scene, int32_t sh_degree, const float* campos_views,
                           const float* color_grads, int32_t n_views, float* grad_shs, float* grad_means3D,
                           int32_t accumulate, int64_t first, int64_t count, void* stream) {
  if (!h || !scene || !campos_views || !color_grads || n_views < 0)
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views: bad argument");
  if (scene->P <= 0 || n_views == 0) return PDGS_OK;
  if (!scene->means3D || !scene->shs || scene->M < 1 || scene->M > 16)
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views: means3D / shs missing or M outside 1..16");
  if (sh_degree < 0 || sh_degree > 3 || (sh_degree + 1) * (sh_degree + 1) > scene->M)
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views: sh_degree needs more coefficients than shs holds");
  if (scene->M == 16 && (!aligned16(scene->shs) || (grad_shs && !aligned16(grad_shs))))
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views: shs / grad_shs must be 16-byte aligned");
  if (count > 0 && (first < 0 || first + count > scene->P || first % 4 != 0))
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views: range outside [0, P) or `first` not a multiple of 4");
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_SH_BWD, s);
  PD_CUDA(launch_sh_backward_views(scene->means3D, scene->shs, nullptr, nullptr, scene->P, scene->M, sh_degree, campos_views,
                                   color_grads, n_views, grad_shs, nullptr, nullptr, grad_means3D, accumulate, first, count, s));
  return PDGS_OK;
}

int pdgs_sh_backward_views_split(PdgsHandle* h, const PdgsScene* scene, int32_t sh_degree, const float* campos_views,
                                 const float* color_grads, int32_t n_views, float* grad_dc, float* grad_rest,
                                 float* grad_means3D, int32_t accumulate, int64_t first, int64_t count, void* stream) {
  if (!h || !scene || !campos_views || !color_grads || n_views < 0)
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views_split: bad argument");
  if (scene->P <= 0 || n_views == 0) return PDGS_OK;
  if (!scene->means3D || !scene->shs_dc || scene->M < 1 || scene->M > 16 || (scene->M > 1 && !scene->shs_rest))
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views_split: means3D / shs_dc / shs_rest missing or M outside 1..16");
  if (sh_degree < 0 || sh_degree > 3 || (sh_degree + 1) * (sh_degree + 1) > scene->M)
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views_split: sh_degree needs more coefficients than the scene holds");
  if ((grad_dc != nullptr) != (grad_rest != nullptr || scene->M == 1) && scene->M > 1)
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views_split: grad_dc and grad_rest must be given together");
  if (!aligned16(scene->shs_dc) || (scene->shs_rest && !aligned16(scene->shs_rest)) || (grad_dc && !aligned16(grad_dc)) ||
      (grad_rest && !aligned16(grad_rest)))
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views_split: SH arrays must be 16-byte aligned");
  if (count > 0 && (first < 0 || first + count > scene->P || first % 256 != 0))
    return fail(PDGS_E_INVALID, "pdgs_sh_backward_views_split: range outside [0, P) or `first` not a multiple of 256");
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_SH_BWD, s);
  PD_CUDA(launch_sh_backward_views(scene->means3D, nullptr, scene->shs_dc, scene->shs_rest, scene->P, scene->M, sh_degree,
                                   campos_views, color_grads, n_views, nullptr, grad_dc, grad_rest, grad_means3D, accumulate,
                                   first, count, s));
  return PDGS_OK;
}

int pdgs_sparse_adam(PdgsHandle* h, const PdgsAdamGroup* groups, int32_t n_groups, const uint8_t* visible,
                     const int32_t* radii, int64_t N, double beta1, double beta2, double eps, int32_t step, void* stream) {
  if (!h || n_groups < 0 || N < 0 || step < 0) return fail(PDGS_E_INVALID, "pdgs_sparse_adam: bad argument");
  if (n_groups == 0 || N == 0) return PDGS_OK;
  if (!groups || n_groups > ADAM_MAX_GROUPS) return fail(PDGS_E_INVALID, "pdgs_sparse_adam: 1..8 groups");
  if (!visible && !radii) return fail(PDGS_E_INVALID, "pdgs_sparse_adam: visible or radii must be given");
  AdamGroup g[ADAM_MAX_GROUPS];
  for (int i = 0; i < n_groups; ++i) {
    if (!groups[i].param || !groups[i].grad || !groups[i].exp_avg || !groups[i].exp_avg_sq || groups[i].M < 1)
      return fail(PDGS_E_INVALID, "pdgs_sparse_adam: NULL tensor or M < 1 in a group");
    g[i].param = groups[i].param; g[i].grad = groups[i].grad; g[i].exp_avg = groups[i].exp_avg;
    g[i].exp_avg_sq = groups[i].exp_avg_sq; g[i].lr_host = groups[i].lr; g[i].lr = 0.f; g[i].M = groups[i].M; g[i].vec4 = 0;
  }
  cudaStream_t s = (cudaStream_t)stream;
  StageTimer t(h, PDGS_STAGE_ADAM, s);
  PD_CUDA(launch_sparse_adam(g, n_groups, visible, radii, N, beta1, beta2, eps, step, s));
  return PDGS_OK;
}

int pdgs_mark_visible(const float* means3D, const float* viewmatrix, const float* projmatrix, int64_t P,
                      uint8_t* present, void* stream) {
  (void)projmatrix;
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return fail(PDGS_E_INVALID, "pdgs_mark_visible: bad argument");
  PD_CUDA(launch_mark_visible(means3D, viewmatrix, P, present, (cudaStream_t)stream));
  return PDGS_OK;
}

int pdgs_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int64_t n,
                    int32_t end_bit, void* scratch, size_t scratch_bytes, void* stream) {
  if (n < 0 || end_bit < 1 || end_bit > 64) return fail(PDGS_E_INVALID, "pdgs_sort_pairs: bad n / end_bit");
  if (n == 0) return PDGS_OK;
  if ((uint64_t)n >= MAX_PAIRS) return fail(PDGS_E_OVERFLOW, "pdgs_sort_pairs: n >= 2^30");
  if (!keys_in || !vals_in || !keys_out || !vals_out || !scratch || scratch_bytes < pdgs_sort_scratch_bytes(n))
    return fail(PDGS_E_INVALID, "pdgs_sort_pairs: NULL buffer or scratch too small");
  cudaStream_t s = (cudaStream_t)stream;
  const int npass = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
  // the count lives in the last 256 bytes of the scratch
  SortScratch sc = carve_sort_scratch(scratch, n, SORT_THREADS * SORT_ITEMS_U64);
  uint32_t* cnt = (uint32_t*)((char*)scratch + sc.bytes);
  const uint32_t n32 = (uint32_t)n;
  PD_CUDA(cudaMemcpyAsync(cnt, &n32, 4, cudaMemcpyHostToDevice, s));
  if (npass % 2 == 0) {
    // input must sit in A (= out buffers) for an even number of passes: copy in -> out first
    PD_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, s));
    PD_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
  }
  PD_CUDA(launch_sort_u64(keys_out, vals_out, keys_in, vals_in, cnt, n, npass, end_bit, scratch, s));
  return PDGS_OK;
}

// pdgs_sort_pairs for callers that cannot allocate: same kernels, same result, but the working memory is the
// library's own fixed array (see g_static_sort_scratch above) instead of a caller buffer.  Nothing here, or below
// it, calls cudaMalloc / cudaFree or any routine that allocates.  The first call on a device looks up the array's
// device address (make that call before starting a stream capture); later calls only enqueue work.  All calls on
// a device share the one array: it is cleared at the start of every sort (the memset inside sort_impl), and two
// sorts that run on the GPU at the same time -- different streams, different host threads -- corrupt each other.
int pdgs_sort_pairs_static(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int64_t n,
                           int32_t end_bit, void* stream) {
  if (n < 0 || end_bit < 1 || end_bit > 64) return fail(PDGS_E_INVALID, "pdgs_sort_pairs_static: bad n / end_bit");
  if (n > STATIC_SORT_MAX_PAIRS)
    return fail(PDGS_E_INVALID, "pdgs_sort_pairs_static: n exceeds the built-in scratch (limit %s pairs); "
                                "use pdgs_sort_pairs with a caller-provided scratch", "1048576");
  if (n == 0) return PDGS_OK;
  if (!keys_in || !vals_in || !keys_out || !vals_out) return fail(PDGS_E_INVALID, "pdgs_sort_pairs_static: NULL buffer");
  void* scratch = nullptr;
  const int rc = static_sort_scratch(&scratch);
  if (rc != PDGS_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int npass = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
  SortScratch sc = carve_sort_scratch(scratch, n, SORT_THREADS * SORT_ITEMS_U64);
  if (sc.bytes + 256 > STATIC_SORT_BYTES) return fail(PDGS_E_INVALID, "pdgs_sort_pairs_static: scratch layout outgrew the array");
  uint32_t* cnt = (uint32_t*)((char*)scratch + sc.bytes);
  const uint32_t n32 = (uint32_t)n;
  PD_CUDA(cudaMemcpyAsync(cnt, &n32, 4, cudaMemcpyHostToDevice, s));
  if (npass % 2 == 0) {
    PD_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, s));
    PD_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
  }
  PD_CUDA(launch_sort_u64(keys_out, vals_out, keys_in, vals_in, cnt, n, npass, end_bit, scratch, s));
  return PDGS_OK;
}

int pdgs_state_view(const PdgsState* state, int64_t P, int32_t width, int32_t height, PdgsStateView* view) {
  if (!state || !view) return fail(PDGS_E_INVALID, "pdgs_state_view: NULL argument");
  GeomView g = carve_geom(state->geom, P);
  ImageView im = carve_image(state->image, width, height);
  memset(view, 0, sizeof(*view));
  if (state->binning) {
    BinningView b = carve_binning(state->binning, state->capacity_pairs);
    view->point_list = b.point_list;
    view->sorted_tiles = b.tkeys_a;
  }
  view->ranges = (const uint32_t*)im.ranges;
  view->final_T = im.final_T;
  view->n_contrib = im.n_contrib;
  view->n_tested = im.n_tested;
  view->n_accepted = im.n_accepted;
  view->tiles_touched = g.tiles_touched;
  view->point_offsets = g.point_offsets;
  view->depth_order = g.order;
  view->xy = (const float*)g.gA;
  view->conic_opacity = (const float*)g.gB;
  view->rgb = (const float*)g.gC;
  return PDGS_OK;
}

}  // extern "C"
