/* pdgs.h -- C ABI of the B200-native differentiable Gaussian-splat rasterizer (libpdgs.so).
 *
 * This is the drop-in boundary under the Python package `diff_gaussian_rasterization`.
 * Each entry point names the reference interface it replaces.  The reference's native
 * module is the pinned submodule graphdeco-inria/diff-gaussian-rasterization@9c5c2028
 * (branch dr_aa), whose source is ABSENT from /root/reference (.SUBMODULES.json: "missing");
 * the interface is therefore cited through its only caller,
 * /root/reference/gaussian_renderer/__init__.py:14,36-52,104-112, and through SURVEY.md
 * section 8(a)/(b) and Appendix B (upstream recall):
 *
 *   pdgs_forward       <- _C.rasterize_gaussians          (RasterizeGaussiansCUDA -> Rasterizer::forward)
 *   pdgs_backward      <- _C.rasterize_gaussians_backward (RasterizeGaussiansBackwardCUDA -> Rasterizer::backward)
 *   pdgs_mark_visible  <- _C.mark_visible                 (markVisible -> checkFrustum)
 *   pdgs_*_bytes       <- the reference's resize-callback "required<State>(n)" sizing of
 *                         geomBuffer / binningBuffer / imgBuffer
 *   pdgs_sort_pairs    <- the cub::DeviceRadixSort::SortPairs call inside Rasterizer::forward
 *                         (exposed so the sort stage can be tested and timed alone)
 *
 * Conventions: plain pointers and sizes, no torch types.  Every `const float*` / output
 * pointer is a DEVICE pointer on the current CUDA device unless marked "host".  All work is
 * enqueued on `stream`.  Matrices are the reference's transposed 4x4s (flat 4*c+r = (r,c),
 * /root/reference/scene/cameras.py:86-88).  Functions return 0 or a negative PDGS_E_* code;
 * pdgs_last_error() gives the message for the calling thread.
 */
#ifndef PDGS_H
#define PDGS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDGS_OK 0
#define PDGS_NEED_BINNING 1      /* not an error: binning capacity too small, call again with resume=1 */
#define PDGS_E_INVALID (-1)      /* bad argument (shape / null pointer / flag combination) */
#define PDGS_E_CUDA (-2)         /* a CUDA call or kernel failed; see pdgs_last_error() */
#define PDGS_E_OVERFLOW (-3)     /* number of (tile, Gaussian) pairs does not fit 2^30 */
#define PDGS_E_NO_DEVICE (-4)    /* no usable sm_100 device / kernels missing for this device */

typedef struct PdgsHandle PdgsHandle; /* per-thread/stream context: pinned counter slot, events */

/* Camera + flags: the fields of GaussianRasterizationSettings
 * (/root/reference/gaussian_renderer/__init__.py:36-50). */
typedef struct PdgsCamera {
  int32_t image_height, image_width;
  float tanfovx, tanfovy;
  const float* bg;          /* [3]  device */
  float scale_modifier;
  const float* viewmatrix;  /* [16] device */
  const float* projmatrix;  /* [16] device */
  int32_t sh_degree;        /* active degree 0..3 */
  const float* campos;      /* [3]  device */
  int32_t prefiltered;      /* must be 0: the reference traps on a culled point when set and always passes False
                             * (gaussian_renderer/__init__.py:47); pdgs_forward returns PDGS_E_INVALID for 1 */
  int32_t debug;            /* 1: synchronise + check after every kernel */
  int32_t antialiasing;
} PdgsCamera;

/* The Gaussians: arguments of GaussianRasterizer.forward
 * (/root/reference/gaussian_renderer/__init__.py:104-112).  Exactly one of shs /
 * colors_precomp and exactly one of (scales, rotations) / cov3D_precomp is non-NULL. */
typedef struct PdgsScene {
  int64_t P;                    /* number of Gaussians */
  int32_t M;                    /* SH coefficients per Gaussian in `shs` (1,4,9,16); 0 if shs NULL */
  const float* means3D;         /* [P,3] */
  const float* shs;             /* [P,M,3] */
  const float* colors_precomp;  /* [P,3] */
  const float* opacities;       /* [P] */
  const float* scales;          /* [P,3] */
  const float* rotations;       /* [P,4] */
  const float* cov3D_precomp;   /* [P,6] */
  /* Fused activations + split SH.  These let the caller hand over the optimiser's own
   * parameter tensors, so the per-iteration torch kernels of /root/reference/scene/gaussian_model.py:103-131
   * (exp / normalize / sigmoid) and the 192 B/Gaussian torch.cat of get_features (:115-119) disappear.
   *  - shs_dc + shs_rest replace `shs` (the `dc=`, `shs=` pair render() passes when separate_sh is set,
   *    /root/reference/gaussian_renderer/__init__.py:82-101): features_dc [P,1,3] and features_rest [P,M-1,3];
   *    M stays the total coefficient count.  Exactly one of shs / (shs_dc, shs_rest) / colors_precomp is given.
   *  - raw_params = 1: opacities / scales / rotations hold _opacity / _scaling / _rotation BEFORE activation; K1
   *    applies sigmoid / exp / x / max(|x|, 1e-12) (gaussian_model.py:40-48) and K9 returns gradients w.r.t. the
   *    raw tensors.  Ignored for cov3D_precomp (opacity is still activated). */
  const float* shs_dc;          /* [P,1,3] */
  const float* shs_rest;        /* [P,M-1,3] (may be NULL when M == 1) */
  int32_t raw_params;
} PdgsScene;

/* Per-call state, allocated by the caller with the sizes below (the reference's
 * geomBuffer / binningBuffer / imgBuffer).  It must stay alive until backward. */
typedef struct PdgsState {
  void* geom;    size_t geom_bytes;      /* pdgs_geom_bytes(P) */
  void* image;   size_t image_bytes;     /* pdgs_image_bytes(W,H) */
  void* binning; size_t binning_bytes;   /* pdgs_binning_bytes(capacity_pairs, P); may be NULL/0 */
  int64_t capacity_pairs;                /* pairs the binning buffer was sized for */
} PdgsState;

typedef struct PdgsForwardOut {
  float* color;      /* [3,H,W] */
  float* invdepth;   /* [1,H,W] */
  int32_t* radii;    /* [P] */
} PdgsForwardOut;

typedef struct PdgsGrads {       /* any pointer may be NULL = not wanted */
  float* means3D;        /* [P,3] */
  float* means2D;        /* [P,3] (.z = 0) */
  float* shs;            /* [P,M,3] */
  float* colors_precomp; /* [P,3] */
  float* opacities;      /* [P] */
  float* scales;         /* [P,3] */
  float* rotations;      /* [P,4] */
  float* cov3D_precomp;  /* [P,6] */
  int32_t accumulate;    /* 0: overwrite (dense, zeros for culled); 1: += into the buffers */
  /* Optional densification statistics of this view, updated for visible Gaussians (radii > 0) in the
   * same pass; they replace the caller-side torch ops at /root/reference/scene/gaussian_model.py:502-506
   * and /root/reference/train.py:313-314.  NULL = not wanted. */
  float* densify_grad_accum; /* [P] += || dL/dmeans2D.xy ||  (xyz_gradient_accum) */
  float* densify_denom;      /* [P] += 1                      (denom) */
  int32_t* max_radii2D;      /* [P]  = max(old, radii)        (max_radii2D) */
  /* Deferred SH backward (batch path).  When non-NULL (SH path only) backward writes this view's clamp-masked
   * dL/d(rgb) per Gaussian here (zeros for culled ones) and does NOT touch `shs` nor add the view-direction term
   * to `means3D`: both are produced for all views of a batch at once by pdgs_sh_backward_views(), which reads the
   * SH coefficients and writes their gradient once per batch instead of once per view. */
  float* color_grad_view;    /* [P,3] */
  /* Split SH gradients (scene given as shs_dc / shs_rest): replace `shs`. */
  float* shs_dc;             /* [P,1,3] */
  float* shs_rest;           /* [P,M-1,3] */
} PdgsGrads;

const char* pdgs_last_error(void);
const char* pdgs_version(void);

int pdgs_create(PdgsHandle** out);
int pdgs_destroy(PdgsHandle* h);

/* Per-stage device timing (no reference counterpart; the reference times whole iterations with
 * two CUDA events, /root/reference/train.py:87-88,121,186).  When enabled, every stage of every
 * forward/backward/loss call on this handle is bracketed by a CUDA event pair on the caller's
 * stream (no host synchronisation).  pdgs_stage_times() waits for the recorded pairs, returns per
 * stage the SUM of durations in ms and the number of pairs since the previous call, and resets. */
enum {
  PDGS_STAGE_PREPROCESS = 0, PDGS_STAGE_SCAN = 1, PDGS_STAGE_DUPLICATE = 2, PDGS_STAGE_SORT = 3,
  PDGS_STAGE_RANGES = 4, PDGS_STAGE_BLEND_FWD = 5, PDGS_STAGE_BWD_ZERO = 6, PDGS_STAGE_BLEND_BWD = 7,
  PDGS_STAGE_PREPROCESS_BWD = 8, PDGS_STAGE_LOSS = 9, PDGS_STAGE_DEPTH_SORT = 10, PDGS_STAGE_SH_BWD = 11,
  PDGS_STAGE_ADAM = 12, PDGS_STAGE_COUNT = 13
};
int pdgs_set_stage_timing(PdgsHandle* h, int32_t enable);
/* Work counters (no reference counterpart): when enabled, the forward blend of this handle also writes, per pixel, the
 * number of list entries a pixel-serial walk evaluates before the pixel stops (n_tested) and the number it blends
 * (n_accepted) -- the N_test / N_acc of SURVEY 8(d)'s algorithmic FLOP count, read through pdgs_state_view.  Off by
 * default: the counters cost instructions in the hottest loop and nothing of the path needs them. */
int pdgs_set_work_counters(PdgsHandle* h, int32_t enable);
int pdgs_stage_times(PdgsHandle* h, float* ms_sum, int32_t* counts, int32_t n);

size_t pdgs_geom_bytes(int64_t P);
size_t pdgs_image_bytes(int32_t width, int32_t height);
size_t pdgs_binning_bytes(int64_t capacity_pairs, int64_t P);
size_t pdgs_backward_scratch_bytes(int64_t P);

/* Forward.  resume=0: preprocess + depth sort, then (if state->capacity_pairs > 0) scan +
 * duplicateWithKeys (one kernel, timed as PDGS_STAGE_DUPLICATE), sort and blend are launched
 * speculatively behind a device-side guard -- no pair is written at or beyond the capacity --
 * and the host waits only for the pair count; without a binning buffer only the scan runs
 * (PDGS_STAGE_SCAN).  Returns PDGS_OK when the image is complete (enqueued), or
 * PDGS_NEED_BINNING with *num_rendered set when the binning buffer is absent/too small:
 * allocate pdgs_binning_bytes(>= *num_rendered, P) and call again with resume=1, which runs
 * only duplicateWithKeys -> sort -> ranges -> blend on the state already in `geom`. */
int pdgs_forward(PdgsHandle* h, const PdgsScene* scene, const PdgsCamera* cam, PdgsState* state,
                 const PdgsForwardOut* out, int32_t resume, int64_t* num_rendered, void* stream);

/* Backward.  dL_dinvdepth may be NULL (treated as zeros).  scratch: pdgs_backward_scratch_bytes(P). */
int pdgs_backward(PdgsHandle* h, const PdgsScene* scene, const PdgsCamera* cam, const PdgsState* state,
                  const int32_t* radii, int64_t num_rendered, const float* dL_dcolor, const float* dL_dinvdepth,
                  void* scratch, size_t scratch_bytes, const PdgsGrads* grads, void* stream);

/* SH part of the backward for a batch of views (see PdgsGrads.color_grad_view): for every Gaussian and every view v,
 *   grad_shs[k][c]  (+)= basis_k(dir_v) * color_grads[v][c],   dir_v = normalize(mean - campos_views[v])
 *   grad_means3D    +=  d(rgb)/d(mean) through dir_v           (the view-direction term of K9)
 * color_grads: [n_views][P][3] device, campos_views: [n_views][3] device.  accumulate = 0 overwrites grad_shs
 * (grad_means3D is always added to).  Replaces n_views read-modify-writes of the [P,M,3] SH gradient by one.
 * first / count select the Gaussians [first, first + count) (count <= 0: all; `first` a multiple of 4, for the split
 * variant of 256): the view-parallel step runs the pass in chunks along P and starts the all-reduce of a chunk's rows
 * while the next chunk is computed (SURVEY 8(e) "chunking along P"). */
int pdgs_sh_backward_views(PdgsHandle* h, const PdgsScene* scene, int32_t sh_degree, const float* campos_views,
                           const float* color_grads, int32_t n_views, float* grad_shs, float* grad_means3D,
                           int32_t accumulate, int64_t first, int64_t count, void* stream);
/* Same for a scene given as shs_dc / shs_rest: gradients go to grad_dc [P,1,3] and grad_rest [P,M-1,3]. */
int pdgs_sh_backward_views_split(PdgsHandle* h, const PdgsScene* scene, int32_t sh_degree, const float* campos_views,
                                 const float* color_grads, int32_t n_views, float* grad_dc, float* grad_rest,
                                 float* grad_means3D, int32_t accumulate, int64_t first, int64_t count, void* stream);

/* Image-space loss of one view and its cotangents in one pass (the caller-side torch ops at
 * /root/reference/train.py:153 `l1_loss` = utils/loss_utils.py:40-41 and :174 the depth-prior term):
 *   loss[0] += mean |color - target|                     over 3*H*W
 *   loss[1] += mean |(invdepth - prior) * mask|          over H*W      (skipped when prior is NULL)
 *   dL_dcolor    = w_color * sign(color - target) / (3*H*W)
 *   dL_dinvdepth = w_depth * sign((invdepth - prior) * mask) * mask / (H*W)
 * mask may be NULL (= ones).  loss is a 2-float DEVICE accumulator the caller zeroes.
 * clamp01 / alpha_mask restate what the reference's caller does to the rendered image BEFORE the losses
 * (/root/reference/gaussian_renderer/__init__.py:121 `rendered_image.clamp(0, 1)`; train.py:148-150
 * `image *= alpha_mask`): the loss is taken on a = clamp(color, 0, 1) * alpha_mask (alpha_mask: one [H*W] plane
 * shared by the three colour planes, NULL = ones) and dL_dcolor is the cotangent w.r.t. the RAW rasterizer output:
 * dL/da * alpha_mask where 0 <= color <= 1, zero elsewhere (torch.clamp's backward). */
int pdgs_image_loss(PdgsHandle* h, const float* color, const float* target, const float* invdepth,
                    const float* prior, const float* mask, const float* alpha_mask, int32_t clamp01, int32_t width,
                    int32_t height, float w_color, float w_depth, float* dL_dcolor, float* dL_dinvdepth, float* loss,
                    void* stream);

/* SSIM with the 11x11 Gaussian window (sigma 1.5, zero "same" padding) on planar fp32 images [planes][H][W]; the
 * map-level interface of the reference's fused-ssim module:
 *   fusedssim(C1, C2, img1, img2, train) -> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
 *       /root/reference/submodules/fused-ssim/ssim.cu:368-404, kernel :187-289, ext.cpp:5
 *   fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12) -> dL_dimg1
 *       /root/reference/submodules/fused-ssim/ssim.cu:406-444, kernel :291-366, ext.cpp:6
 * ssim_map may be NULL; the three partial maps are all given (train) or all NULL; ssim_sum (may be NULL) is a DEVICE
 * accumulator that receives += sum of the map.  dL_dmap NULL = the uniform value dL_dmap_uniform (a mean). */
int pdgs_ssim_forward(PdgsHandle* h, float C1, float C2, const float* img1, const float* img2, int32_t planes,
                      int32_t width, int32_t height, float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq,
                      float* dm_dsigma12, float* ssim_sum, void* stream);
int pdgs_ssim_backward(PdgsHandle* h, const float* img1, const float* img2, const float* dL_dmap, float dL_dmap_uniform,
                       const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, int32_t planes,
                       int32_t width, int32_t height, float* dL_dimg1, void* stream);
size_t pdgs_ssim_scratch_bytes(int32_t width, int32_t height, int32_t planes);

/* The reference's training loss and its cotangents in three launches (/root/reference/train.py:153-159, 174):
 *   L = (1 - lambda_dssim) * mean|color - target| + lambda_dssim * (1 - mean ssim(color, target))
 *       + w_depth * mean|(invdepth - prior) * mask|
 *   loss[0] += mean L1, loss[1] += mean masked depth L1 (prior != NULL), loss[2] += mean SSIM   (DEVICE float[3])
 *   dL_dcolor = dL/dcolor, dL_dinvdepth = dL/dinvdepth.  scratch: pdgs_ssim_scratch_bytes(width, height, 3) bytes.
 * clamp01 / alpha_mask: as in pdgs_image_loss -- with clamp01 = 1 this is the reference's training loss on the raw
 * rasterizer output (render() clamps, train.py masks, then l1_loss and fused_ssim see that image). */
int pdgs_image_loss_ssim(PdgsHandle* h, const float* color, const float* target, const float* invdepth,
                         const float* prior, const float* mask, const float* alpha_mask, int32_t clamp01, int32_t width,
                         int32_t height, float lambda_dssim, float w_depth, void* scratch, size_t scratch_bytes,
                         float* dL_dcolor, float* dL_dinvdepth, float* loss, void* stream);

/* Visibility-sparse Adam: the kernel under `SparseGaussianAdam.step(visibility, N)`, which the reference uses when the
 * rasterizer package exports that class (/root/reference/train.py:47-52, 332-334; scene/gaussian_model.py:25-28,
 * 196-198; the class itself lives in the accelerated rasterizer branch, absent from /root/reference).  For every
 * parameter group (tensor [N, M], row i = Gaussian i) and every row with visible[i] != 0 (or radii[i] > 0 when
 * `visible` is NULL):
 *   m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;  p -= lr * m / (sqrt(v) + eps)                      (step == 0)
 *   p -= (lr / (1 - b1^step)) * m / (sqrt(v) / sqrt(1 - b2^step) + eps)   (step > 0: torch.optim.Adam's correction)
 * Rows of invisible Gaussians are neither read nor written.  All groups (<= 8) are updated by one launch.  Arithmetic
 * is fp32; the scalars arrive as doubles so that step > 0 can round (1 - beta), lr / (1 - b1^step) and
 * sqrt(1 - b2^step) to fp32 once, as torch does, while step == 0 forms (1.0f - beta) in fp32 like the upstream kernel. */
typedef struct PdgsAdamGroup {
  float* param;        /* [N,M] updated in place */
  const float* grad;   /* [N,M] */
  float* exp_avg;      /* [N,M] */
  float* exp_avg_sq;   /* [N,M] */
  double lr;
  int32_t M;           /* floats per Gaussian in this tensor */
} PdgsAdamGroup;
int pdgs_sparse_adam(PdgsHandle* h, const PdgsAdamGroup* groups, int32_t n_groups, const uint8_t* visible,
                     const int32_t* radii, int64_t N, double beta1, double beta2, double eps, int32_t step, void* stream);

int pdgs_mark_visible(const float* means3D, const float* viewmatrix, const float* projmatrix, int64_t P,
                      uint8_t* present, void* stream);

/* Stand-alone stable LSD radix sort of (u64 key, u32 value) pairs over key bits [0, end_bit).
 * scratch: pdgs_sort_scratch_bytes(n).  Result in keys_out / vals_out.  keys_in/vals_in are clobbered. */
size_t pdgs_sort_scratch_bytes(int64_t n);
int pdgs_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int64_t n,
                    int32_t end_bit, void* scratch, size_t scratch_bytes, void* stream);

/* The same sort for callers that are not allowed to allocate device memory.  No counterpart in the reference:
 * upstream sizes CUB's temporary storage through its resize callbacks and lets torch allocate it.  Typical users are
 * hosts where all device memory belongs to somebody else -- an engine or inference server whose memory manager took
 * the whole budget at start-up, a plugin invoked inside another program's frame loop where a cudaMalloc would
 * stall every stream on the device, or an embedded deployment with a fixed memory map.
 *
 * The working memory is a fixed array inside libpdgs.so (about 2.7 MiB of device memory on every GPU the library is
 * loaded on, held until the process exits), large enough for PDGS_SORT_STATIC_MAX_PAIRS pairs; larger n is refused
 * with PDGS_E_INVALID and a message naming the limit.  The call allocates nothing, on host or device, and neither
 * does anything it calls.  Output is identical to pdgs_sort_pairs; keys_in / vals_in are clobbered the same way.
 *
 * Set-up: the first call on each device (made with that device current) looks up the array's device address; do that
 * first call before beginning a stream capture.  Sharing: every call on a device uses the same array.  Each sort
 * clears what it needs on entry, so nothing has to be reset between calls, but two calls whose GPU work overlaps
 * (different streams or host threads) race on it -- serialise them on one stream or with events. */
#define PDGS_SORT_STATIC_MAX_PAIRS (1 << 20)
int pdgs_sort_pairs_static(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int64_t n,
                           int32_t end_bit, void* stream);

/* Read-only views into the saved state, for parity tests (device pointers, valid until the
 * state buffers are freed). */
typedef struct PdgsStateView {
  const uint32_t* point_list;   /* [R] sorted Gaussian ids */
  const uint32_t* sorted_tiles; /* [R] tile id of each sorted pair; the reference's 64-bit key of pair i is
                                 *     (sorted_tiles[i] << 32) | bits(conic_opacity[point_list[i]][3]) */
  const uint32_t* ranges;       /* [T,2] */
  const float* final_T;         /* [H*W] */
  const uint32_t* n_contrib;    /* [H*W] */
  const uint32_t* n_tested;     /* [H*W] list entries evaluated by the forward blend (valid with pdgs_set_work_counters) */
  const uint32_t* n_accepted;   /* [H*W] list entries blended (valid with pdgs_set_work_counters) */
  const uint32_t* tiles_touched;/* [P] */
  const uint32_t* point_offsets;/* [P] inclusive scan of tiles_touched taken in depth order */
  const uint32_t* depth_order;  /* [P] Gaussian indices sorted by (depth bits, index); culled ones last */
  const float* xy;              /* [P,4] (x, y, conic.x, conic.y) */
  const float* conic_opacity;   /* [P,4] (conic.z, opacity, 1/depth, depth) */
  const float* rgb;             /* [P,4] (r, g, b, clamped-bits) */
} PdgsStateView;
int pdgs_state_view(const PdgsState* state, int64_t P, int32_t width, int32_t height, PdgsStateView* view);

#ifdef __cplusplus
}
#endif
#endif /* PDGS_H */
