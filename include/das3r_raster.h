/*
 * das3r_raster.h — C ABI of the B200-native (sm_100a) differentiable Gaussian-splat rasterizer and the 3-NN
 * initialiser that replace DAS3R's submodules/diff-gaussian-rasterization and submodules/simple-knn.
 *
 * Everything here is plain C: POD structs, raw DEVICE pointers, sizes, a CUDA stream passed as void*.
 * No torch types. The torch extension (das3r_b200/csrc/torch_ext.cpp) and the ctypes tests are the callers.
 *
 * Conventions (mirroring the reference library API, submodules/diff-gaussian-rasterization/
 * cuda_rasterizer/rasterizer.h:20-85):
 *   - a NULL input pointer means "this optional input is absent" (the reference passes empty tensors whose
 *     data pointer is null: diff_gaussian_rasterization/__init__.py:197-207 <-> forward.cu:205,241);
 *   - the caller owns all memory. The three growable state blobs are requested through allocator callbacks
 *     (rasterizer.h:31-33 uses std::function<char*(size_t)>), and handed back verbatim to the backward;
 *   - every function returns 0 (or a non-negative count) on success and a negative DAS3R_E* code on failure;
 *     das3r_last_error() then returns a human-readable message for the calling thread;
 *   - all work is enqueued on `stream` (cudaStream_t as void*; NULL = legacy default stream).
 */
#ifndef DAS3R_RASTER_H_
#define DAS3R_RASTER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DAS3R_ABI_VERSION 3

#define DAS3R_OK 0
#define DAS3R_EINVAL (-1)   /* bad argument (shape, null, size) */
#define DAS3R_ECUDA (-2)    /* a CUDA runtime call or kernel failed */
#define DAS3R_ENOMEM (-3)   /* an allocator callback returned NULL */
#define DAS3R_ELIMIT (-4)   /* size beyond a documented limit */

/* Allocator callback: return a DEVICE pointer to at least `bytes` bytes (256-byte aligned), or NULL.
 * Replaces std::function<char*(size_t)> of rasterizer.h:31-33 / resizeFunctional, rasterize_points.cu:27-33. */
typedef char* (*das3r_alloc_fn)(size_t bytes, void* ctx);

/* Per-call camera / configuration bundle. Field meaning follows GaussianRasterizationSettings
 * (diff_gaussian_rasterization/__init__.py:157-169) and the argument list of Rasterizer::forward
 * (rasterizer_impl.cu:198-221). Matrices are 16 floats in the reference's flat layout (column-major as the
 * kernels index them: auxiliary.h:58-77). */
typedef struct das3r_raster_params {
    int32_t P;              /* number of Gaussians */
    int32_t D;              /* active SH degree 0..3 */
    int32_t M;              /* SH coefficients per channel in `shs` rows ((max_degree+1)^2; 0 if shs absent) */
    int32_t width, height;  /* image size in pixels */
    float tan_fovx, tan_fovy;
    float scale_modifier;
    int32_t prefiltered;    /* nonzero: a culled point is a device-side trap (auxiliary.h:156-160) */
    int32_t debug;          /* nonzero: synchronise and check after every stage (CHECK_CUDA, auxiliary.h:166-173) */
    const float* background;  /* device [3] */
    const float* viewmatrix;  /* device [16] */
    const float* projmatrix;  /* device [16] */
    const float* campos;      /* device [3] */
} das3r_raster_params;

/* Per-Gaussian inputs (device, contiguous fp32). Exactly one of shs / colors_precomp and exactly one of
 * (scales, rotations) / cov3D_precomp must be non-NULL (the Python surface enforces it). */
typedef struct das3r_raster_inputs {
    const float* means3D;        /* [P,3] */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3] or NULL */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL */
    const float* rotations;      /* [P,4] or NULL (used un-normalised, forward.cu:127); 16-byte aligned */
    const float* cov3D_precomp;  /* [P,6] or NULL */
} das3r_raster_inputs;

/* Forward outputs (device, caller-allocated). out_color and radii need not be pre-zeroed. */
typedef struct das3r_raster_outputs {
    float* out_color;   /* [3,H,W] */
    int32_t* radii;     /* [P] */
    float* out_depth;   /* [H,W] or NULL: sum_i z_i alpha_i T_i (additive surface) */
    float* out_alpha;   /* [H,W] or NULL: 1 - final_T */
} das3r_raster_outputs;

/* The three state blobs produced by the forward and consumed by the backward. Their layout is private to this
 * library (see das3r_raster_state_layout for the test/debug view). */
typedef struct das3r_raster_state {
    char* geom;     size_t geom_bytes;
    char* binning;  size_t binning_bytes;
    char* image;    size_t image_bytes;
    int32_t num_rendered;
} das3r_raster_state;

/* Gradient outputs (device, caller-allocated, need NOT be pre-zeroed: every element is written).
 * Shapes follow RasterizeGaussiansBackwardCUDA, rasterize_points.cu:151-159. */
typedef struct das3r_raster_grads {
    float* dL_dmeans2D;    /* [P,3] (z is always 0: backward.cu:545-546 never writes it); may be NULL = not wanted */
    float* dL_dcolors;     /* [P,3]; may be NULL = not wanted */
    float* dL_dopacity;    /* [P] */
    float* dL_dmeans3D;    /* [P,3] */
    float* dL_dcov3D;      /* [P,6]; may be NULL = not wanted */
    float* dL_dsh;         /* [P,M,3] or NULL when M == 0 */
    float* dL_dscales;     /* [P,3] */
    float* dL_drotations;  /* [P,4]; 16-byte aligned */
    float* scratch;        /* device scratch, das3r_raster_backward_scratch_bytes(P) bytes, 16-byte aligned; contents
                              undefined */
    int32_t accumulate;    /* 0: every gradient element is overwritten (reference behaviour: fresh zero-filled
                              tensors). 1: gradients are ADDED to what the buffers hold and Gaussians that were not
                              rendered are left untouched - used to sum several views of a step without a second pass
                              (no reference counterpart; there autograd's AccumulateGrad does it in extra kernels). */
} das3r_raster_grads;

/* Library / ABI version (DAS3R_ABI_VERSION). No reference counterpart. */
int das3r_abi_version(void);

/* Message for the last failure on the calling thread ("" if none). Replaces the std::runtime_error thrown by
 * CHECK_CUDA (auxiliary.h:166-173) and AT_ERROR (rasterize_points.cu:58). */
const char* das3r_last_error(void);

/* Forward rasterization. Replaces CudaRasterizer::Rasterizer::forward (rasterizer.h:31-53,
 * rasterizer_impl.cu:198-336): preprocess -> tile binning -> per-tile depth sort -> tile ranges -> blend.
 * Returns num_rendered (>= 0) like the reference, or a negative error code. `state` receives the blob pointers
 * obtained from the allocators. Blocks the host once (4-byte read-back of num_rendered, as rasterizer_impl.cu:281). */
int das3r_raster_forward(const das3r_raster_params* params, const das3r_raster_inputs* in,
                         const das3r_raster_outputs* out,
                         das3r_alloc_fn geom_alloc, void* geom_ctx,
                         das3r_alloc_fn binning_alloc, void* binning_ctx,
                         das3r_alloc_fn image_alloc, void* image_ctx,
                         das3r_raster_state* state, void* stream);

/* The forward in two halves, for hosts that render several views of the same Gaussians per step (a data-parallel
 * training step, render.py's frame loop). No counterpart in the reference, whose Rasterizer::forward
 * (rasterizer_impl.cu:198-336) handles one view per call.
 *
 *   das3r_raster_forward_preprocess(n, views, stream)   first half for n >= 1 views at once (replaces preprocessCUDA,
 *                                                       forward.cu:155-256, and the prefix sum, rasterizer_impl.cu:277):
 *                                                       allocates every view's geom and image blobs through its
 *                                                       callbacks, projects the Gaussians, counts tile instances,
 *                                                       builds the tile ranges. Views that share their `shs` array
 *                                                       (degree 3, 16 coefficients), P and the image size are handled
 *                                                       by ONE kernel that reads every SH row once - two thirds of the
 *                                                       traffic of this stage; other views get the one-view kernel.
 *                                                       Every output is bit-identical to das3r_raster_forward's.
 *   das3r_raster_forward_render(view, stream_v)         second half for ONE view, on any stream ordered after the
 *                                                       first half: reads back num_rendered (blocks the host like
 *                                                       das3r_raster_forward), then binning, per-tile sort, rect lists,
 *                                                       blend. Returns num_rendered or a negative code.
 * das3r_raster_forward(...) == preprocess(1 view) + render. */
typedef struct das3r_raster_fwd_view {
    const das3r_raster_params* params;
    const das3r_raster_inputs* in;
    const das3r_raster_outputs* out;
    das3r_alloc_fn geom_alloc;    void* geom_ctx;
    das3r_alloc_fn binning_alloc; void* binning_ctx;
    das3r_alloc_fn image_alloc;   void* image_ctx;
    das3r_raster_state* state;    /* filled by the two halves */
} das3r_raster_fwd_view;
int das3r_raster_forward_preprocess(int32_t n_views, const das3r_raster_fwd_view* views, void* stream);
int das3r_raster_forward_render(const das3r_raster_fwd_view* view, void* stream);

/* Second half for n >= 1 views in ONE call (no reference counterpart; the reference renders one view per call,
 * rasterizer_impl.cu:277-336): a single host read-back fetches every view's num_rendered (the reference blocks once per
 * view, rasterizer_impl.cu:281), then duplicateWithKeys / sort / identifyTileRanges / renderCUDA run as ONE launch per
 * stage over all views (grid.y = view). Views that share P and the image size are batched, 16 at a time; the others are
 * grouped by size. Every output is bit-identical to das3r_raster_forward_render. num_rendered[v] receives view v's count.
 * Returns 0 or a negative code. */
int das3r_raster_forward_render_batch(int32_t n_views, const das3r_raster_fwd_view* views, int32_t* num_rendered,
                                      void* stream);

/* Backward. Replaces CudaRasterizer::Rasterizer::backward (rasterizer.h:55-84, rasterizer_impl.cu:340-434):
 * backward blend -> cov2D/projection/SH/cov3D gradients. `radii` is the forward's output. `dL_dpix` is [3,H,W]. */
int das3r_raster_backward(const das3r_raster_params* params, const das3r_raster_inputs* in,
                          const int32_t* radii, const das3r_raster_state* state,
                          const float* dL_dpix, const das3r_raster_grads* grads, void* stream);

/* The backward in two halves, so that a training step over several views reads what the views share only once.
 * It has no counterpart in the reference, whose Rasterizer::backward (rasterizer_impl.cu:340-434) handles one view per
 * call and leaves the sum over views to autograd.
 *
 *   das3r_raster_backward_blend(view, stream)      first half for ONE view: the backward blend (replaces renderCUDA,
 *                                                  backward.cu:399-557). Fills view->scratch
 *                                                  (das3r_raster_backward_scratch_bytes(P) bytes, one buffer per view)
 *                                                  with the screen-space gradient sums. Views are independent: issue
 *                                                  them on as many streams as you like.
 *   das3r_raster_backward_finish(n, views, grads, stream)
 *                                                  second half for n >= 1 views at once (computeCov2DCUDA,
 *                                                  preprocessCUDA, SH and cov3D backward: backward.cu:20-396):
 *                                                  grads = (grads->accumulate ? grads : 0) + sum over the views, i.e.
 *                                                  what n das3r_raster_backward calls with accumulate would leave,
 *                                                  summed in view order. All blend halves must have completed on
 *                                                  `stream` (or be ordered before it by events). grads->scratch is
 *                                                  not used. With n > 1, dL_dmeans2D / dL_dcolors / dL_dcov3D must be
 *                                                  NULL (per-view quantities).
 * When the views use the same `shs` array at degree 3 with scale / rotation inputs (DAS3R's default: the pose moves
 * means and rotations, not the appearance), the SH rows are read once and every gradient row is written once for the
 * whole step; otherwise the call runs one accumulating pass per view. All views must agree in P, D and M. */
typedef struct das3r_raster_view {
    const das3r_raster_params* params;
    const das3r_raster_inputs* in;
    const int32_t* radii;              /* the forward's output */
    const das3r_raster_state* state;   /* the forward's state blobs */
    const float* dL_dpix;              /* [3,H,W] (blend half only) */
    float* scratch;                    /* per-view, written by the blend half, read by the finish half */
    float* dL_dmeans3D;                /* optional [P,3]: this view's gradient w.r.t. ITS means3D goes here instead of
                                          into grads->dL_dmeans3D (DAS3R hands every view its own pose-transformed means
                                          and rotations, gaussian_renderer/__init__.py:83-92, so their gradients have to
                                          go back through that view's pose); NULL: summed into grads->dL_dmeans3D */
    float* dL_drotations;              /* optional [P,4], likewise for the rotations */
} das3r_raster_view;
int das3r_raster_backward_blend(const das3r_raster_view* view, void* stream);
int das3r_raster_backward_finish(int32_t n_views, const das3r_raster_view* views, const das3r_raster_grads* grads,
                                 void* stream);

/* The blend half for n >= 1 views in ONE launch (grid.y = view; views grouped by P and image size, 16 per launch).
 * Same results as n das3r_raster_backward_blend calls. No reference counterpart (renderCUDA, backward.cu:399-557, is
 * launched once per view). */
int das3r_raster_backward_blend_batch(int32_t n_views, const das3r_raster_view* views, void* stream);

/* das3r_raster_backward_finish with two extras (no reference counterpart):
 *   dsh_coeffs   coefficients per ROW of grads->dL_dsh: M (the reference's [P,M,3] layout, rasterize_points.cu:156; 0
 *                means M) down to (D+1)^2. At active degree D only the first (D+1)^2 coefficients of a gradient row can
 *                be non-zero (backward.cu:62-137); DAS3R trains at D = 0 for 3000 of 4000 iterations
 *                (train_gui.py:541-543), where a trimmed row is 12 bytes instead of 192 - less to write, and less to
 *                all-reduce in the data-parallel step. Trimmed rows need the fused path (views sharing 16-coefficient,
 *                16-byte aligned shs, scale / rotation inputs).
 *   first, count only Gaussians [first, first + count) are processed (count < 0: to the end): the caller can hand a
 *                finished slice of the gradient buffers to a collective while the next slice is computed. grads and the
 *                per-view outputs are still addressed by ABSOLUTE Gaussian index (row `first` is the first one written),
 *                unless grads_local != 0: then the pointers in `grads` address Gaussian `first` (a window of a
 *                slice-major gradient bucket); the per-view outputs stay absolute.
 * Alignment: rotations, dL_drotations (shared and per-view) and the accumulators must be 16-byte aligned (float4
 * accesses); violations are refused with DAS3R_EINVAL. */
int das3r_raster_backward_finish_ex(int32_t n_views, const das3r_raster_view* views, const das3r_raster_grads* grads,
                                    int32_t dsh_coeffs, int32_t first, int32_t count, int32_t grads_local,
                                    void* stream);

/* Backward for hosts that cannot allocate device memory. It has no counterpart in the reference.
 *
 * das3r_raster_backward needs `grads->scratch`, a P*48-byte device buffer the caller normally takes from its allocator
 * (the torch binding borrows it from torch's caching allocator). Some callers have no such source: a C or plugin host
 * that owns no device allocator at all, a process whose whole device was handed to another framework's memory pool at
 * start-up (so a cudaMalloc here would fail or carve up that pool), or a frame/step loop that may not risk the
 * device-wide stall a cudaMalloc/cudaFree can cause. For them this variant keeps the scratch inside the library, in a
 * fixed array in device memory that exists once per GPU from the moment the library is loaded there until the process
 * exits (DAS3R_STATIC_MAX_GAUSSIANS * 48 bytes = 6 MiB). `grads->scratch` is ignored; everything else - inputs, state
 * blobs, gradient outputs - is passed exactly as for das3r_raster_backward, and the results are the same.
 *
 *  - P above DAS3R_STATIC_MAX_GAUSSIANS is refused with DAS3R_ELIMIT before any work is enqueued.
 *  - Neither this call nor anything it runs calls cudaMalloc/cudaFree or an allocating library routine.
 *  - The first call on a device looks up the array's device address (a one-time set-up with that device current);
 *    make that first call before starting any stream capture.
 *  - All calls on a device share the one array and it keeps whatever the last call left in it; the part a call uses is
 *    cleared at its start, in stream order. Two calls whose GPU work can overlap (different streams, no dependency)
 *    would race on it: serialise them. */
#define DAS3R_STATIC_MAX_GAUSSIANS 131072
int das3r_raster_backward_static(const das3r_raster_params* params, const das3r_raster_inputs* in,
                                 const int32_t* radii, const das3r_raster_state* state,
                                 const float* dL_dpix, const das3r_raster_grads* grads, void* stream);

/* Bytes of device scratch the backward needs for P Gaussians (screen-space gradient accumulators; they replace
 * the reference's separate zero-filled dL_dconic/dL_dmean2D/dL_dcolor/dL_dopacity tensors, rasterize_points.cu:151-155). */
size_t das3r_raster_backward_scratch_bytes(int32_t P);

/* Frustum test. Replaces CudaRasterizer::Rasterizer::markVisible (rasterizer.h:24-29, rasterizer_impl.cu:141-153,
 * checkFrustum :54-66) with DAS3R's 0.001 near plane. present: device uint8/bool [P]. */
int das3r_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, void* stream);

/* Mean squared distance to the 3 nearest other points. Replaces SimpleKNN::knn (simple-knn/simple_knn.h,
 * simple_knn.cu:186-222) behind distCUDA2 (spatial.cu:15-25). points: device [P,3]; mean_dist2: device [P].
 * Scratch comes from the caller (the reference cudaMallocs internally, simple_knn.cu:189-217):
 * call das3r_knn_scratch_bytes(P) first. Does not block the host (the reference copies the bounding box back twice,
 * simple_knn.cu:197-201; here it stays on the device). */
size_t das3r_knn_scratch_bytes(int32_t P);
int das3r_knn_dist2(int32_t P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes,
                    void* stream);

/* Test/debug view of the private state layout: byte offsets of the arrays the parity tests compare with the
 * reference's GeometryState/BinningState/ImageState (rasterizer_impl.cu:155-194). No reference counterpart. */
typedef struct das3r_raster_state_layout {
    /* geom blob (P-sized) */
    size_t g_recA, g_recB, g_recC;   /* 48-byte records [P]: {x,y,h,-tau'} {conic a,b,c,opacity} {r,g,b,depth} */
    size_t g_tiles_touched;          /* uint32[P] */
    size_t g_clamped;                /* uint8[P], bit c = channel c clamped */
    size_t geom_bytes;
    /* image blob */
    size_t i_final_T, i_n_contrib;   /* float[N], uint32[N] */
    size_t i_ranges;                 /* uint2[T] */
    size_t i_tile_counts, i_cursors; /* uint32[T] each */
    size_t i_totals;                 /* uint32[4]: num_rendered, max tile count, 0, 0 */
    size_t image_bytes;
    /* binning blob (R-sized) */
    size_t b_unsorted;               /* uint64[R]: depth_bits<<32 | gaussian index, tile-bucketed */
    size_t b_point_list;             /* uint32[R] sorted */
    size_t b_keys_sorted;            /* uint64[R] tile<<32|depth_bits, written only when params.debug != 0 */
    size_t b_rect_off, b_rect_cnt;   /* uint32[T*8]: start / length of the list of each 8x4 pixel rectangle */
    size_t b_rect_list;              /* uint2[8*R] capacity: {tile-list position, pixel mask} per rectangle, in list order */
    size_t b_rect_gid;               /* uint32[8*R] capacity: Gaussian index of the same list entry */
    size_t binning_bytes;
} das3r_raster_state_layout;
int das3r_raster_state_layout_query(int32_t P, int32_t width, int32_t height, int32_t num_rendered,
                                    das3r_raster_state_layout* out);

/* Measurement hooks used by bench.py (no reference counterpart): when enabled, every stage of forward/backward is
 * bracketed by CUDA events on the launching stream; das3r_prof_collect waits for them and returns, per stage
 * (das3r_prof_stage_name(i), i < das3r_prof_num_stages()), the summed device time in ms and the number of launches.
 * das3r_launch_count() is the number of this library's kernels launched by the process so far. Not thread-safe. */
int das3r_prof_enable(int on);
int das3r_prof_num_stages(void);
const char* das3r_prof_stage_name(int i);
int das3r_prof_collect(double* ms_total, long long* calls, int reset);
long long das3r_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DAS3R_RASTER_H_ */
