// Launch-side declarations shared between forward.cu, backward.cu, knn.cu and api.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace das3r {

constexpr int TILE_X = 16;  // reference config.h:16-17 (BLOCK_X, BLOCK_Y)
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;

// Per-Gaussian record, 48 bytes, written by the forward preprocess; the blend kernels fetch the records of their list
// entries straight from this array (cp.async, 48 contiguous bytes per entry).
//   A = {x, y, h, -tau'}    pixel-space mean; h = conservative radius of the region where alpha >= 1/255 is possible
//                           (< 0: never), tau' = ln(255*opacity) with a margin: a pixel can contribute iff power >= -tau'
//   B = {conic a, b, c, opacity}   (reference: conic_opacity, forward.cu:253)
//   C = {r, g, b, depth}
struct __align__(16) GaussRec {
    float4 A, B, C;
};
static_assert(sizeof(GaussRec) == 48, "GaussRec must be 48 bytes");

// Screen-space gradient accumulator, 48 bytes per Gaussian (replaces the reference's four zero-filled tensors
// dL_dmean2D/dL_dconic/dL_dopacity/dL_dcolors, rasterize_points.cu:151-155).
//   f[0..1] dL_dmean2D.xy   f[2..4] dL_dconic (x, y, w)   f[5] dL_dopacity   f[6..8] dL_dcolor   f[9..11] pad
constexpr int ACC_STRIDE = 12;

struct PreFwdArgs {
    int P, D, M, W, H;
    unsigned gx, gy;
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_mod;
    int prefiltered;
    const float *means, *shs, *colors_precomp, *opac, *scales, *rots, *cov3D_precomp, *view, *proj, *campos;
    int* radii;
    GaussRec* rec;
    uint32_t* tiles_touched;
    uint8_t* clamped;
    uint32_t* tile_counts;
    int smem_hist;  // 1: per-block shared-memory histogram of tile counts, 0: direct global atomics
};

constexpr int MF_MAX_VIEWS = 16;
struct MultiFwdArgs {  // views that share shs (degree 3, M == 16) and the image size
    int n_views, T, smem_hist, pad_;
    PreFwdArgs v[MF_MAX_VIEWS];
};

struct PreBwdArgs {
    int P, D, M, W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_mod;
    const float *means, *shs, *scales, *rots, *cov3D_precomp, *view, *proj, *campos;
    const int* radii;
    const uint8_t* clamped;
    const float* acc;  // [P][ACC_STRIDE]
    float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales, *dL_drots;
    int accumulate;  // 1: outputs += gradient, invisible Gaussians untouched; 0: outputs = gradient (zeros if invisible)
};

// Multi-view preprocess backward (backward.cu): per-view inputs + what the views share.
constexpr int MV_MAX_VIEWS = 16;
struct MultiBwdView {
    const float *means, *scales, *rots, *view, *proj, *campos;
    const int* radii;
    const uint8_t* clamped;
    const float* acc;  // [P][ACC_STRIDE], filled by this view's backward blend
    float *dL_dmeans3D, *dL_drots;  // optional per-view outputs (see das3r_raster_view); NULL: summed into the shared ones
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_mod;
    int pad_;
};
struct MultiBwdArgs {
    int P, n_views;
    int accumulate_shared;  // the summed outputs: 1 = add to what the buffers hold (and leave unrendered Gaussians alone)
    int accumulate_view;    // the per-view outputs (MultiBwdView::dL_dmeans3D / dL_drots), likewise
    int D;                  // active SH degree 0..3 (rows of `shs` always have 16 coefficients)
    int dsh_coeffs;         // coefficients per row of dL_dsh: (D+1)^2 .. 16 (rows trimmed to the active bands or full)
    const float* shs;  // [P,16,3], shared by all views
    float *dL_dopacity, *dL_dmeans3D, *dL_dsh, *dL_dscales, *dL_drots;
    MultiBwdView v[MV_MAX_VIEWS];
};

// Batched binning + blend (forward.cu, backward.cu): the per-view pointers of up to RB_MAX_VIEWS views that share P and
// the image size; every stage is ONE launch with blockIdx.y = view.
constexpr int RB_MAX_VIEWS = 16;
struct ViewBin {
    // image blob
    uint32_t* tile_counts;
    uint2* ranges;
    uint32_t* cursors;
    uint32_t* totals;  // {num_rendered, largest tile, 0, 0}
    float* final_T;
    uint32_t* n_contrib;
    // geometry blob + the forward's radii
    const GaussRec* rec;
    const int* radii;
    const uint32_t* tiles_touched;
    // binning blob
    uint64_t* unsorted;
    uint32_t* point_list;
    uint32_t* rect_off;
    uint32_t* rect_cnt;
    uint2* rect_list;    // {tile-list position, pixel mask}
    uint32_t* rect_gid;  // Gaussian index of the same entry (the blends fetch its 48-byte record directly)
    uint64_t* keys_sorted;  // debug only
    // per-view inputs / outputs of the blends
    const float* bg;
    float* out_color;
    float* out_depth;
    float* out_alpha;
    const float* dL_dpix;
    float* acc;
};
struct BatchArgs {
    int n, P, W, H;
    unsigned gx, gy;
    int T, pad_;
    ViewBin v[RB_MAX_VIEWS];
};
static_assert(sizeof(BatchArgs) <= 4000, "BatchArgs must fit the 4 KB kernel parameter space");

// Fused DAS3R render front-end (frontend.cu, include/das3r_frontend.h)
constexpr int FE_MAX_VIEWS = 16;
struct FrontendFwdArgs {
    int P, n_views;
    int K;  // active SH coefficients (D+1)^2: only these are copied into the joint [P,16,3] rows
    int pad_;
    const float *xyz, *rotation, *opacity_raw, *scaling_raw, *features_dc, *features_rest, *conf_static, *poses;
    const int64_t* conf_index;
    const int64_t* filter_index;  // optional [P]: model row of output row i (render()'s boolean `filtering`, as indices)
    float *opacity, *scales, *shs;
    float* means3D[FE_MAX_VIEWS];
    float* rotations[FE_MAX_VIEWS];
};
struct FrontendBwdArgs {
    int P, n_views;
    int K;            // active SH coefficients
    int dsh_coeffs;   // coefficients per row of the incoming dL_dshs (K..16)
    int rest_coeffs;  // coefficients per row of the outgoing dL_dfeatures_rest (K-1 = trimmed, or 15)
    int row0, row_end;  // output rows [row0, row_end) are processed by this launch (row0 a multiple of 256)
    int out_row0;       // the parameter-gradient outputs and dL_dconf address this row (0, or row0 for bucket windows)
    const float *xyz, *rotation, *opacity_raw, *scaling_raw, *conf_static, *poses;
    const int64_t* conf_index;
    const int64_t* filter_index;
    const float *dL_dopacity, *dL_dscales, *dL_dshs;
    const float* dL_dmeans3D[FE_MAX_VIEWS];
    const float* dL_drotations[FE_MAX_VIEWS];
    float *dL_dxyz, *dL_drotation, *dL_dopacity_raw, *dL_dscaling_raw, *dL_dfeatures_dc, *dL_dfeatures_rest, *dL_dconf;
    float* partials;  // [blocks][n_views][16]
};
cudaError_t launch_frontend_fwd(const FrontendFwdArgs& a, cudaStream_t s);
size_t frontend_backward_scratch_bytes(int P, int n_views);
// dL_dposes != NULL: also reduce the per-block partial sums of ALL ceil(P/256) blocks to the pose gradients (the last
// range of a step)
cudaError_t launch_frontend_bwd(const FrontendBwdArgs& a, float* dL_dposes, cudaStream_t s);

// Fused photometric loss (loss.cu, include/das3r_loss.h)
struct LossArgs {
    int W, H;
    float lambda_dssim;
    const float *image, *gt, *static_mask;
};
cudaError_t loss_upload_window();  // once per device: the 11 window weights into constant memory
size_t loss_state_bytes(int W, int H);
cudaError_t launch_loss_fwd(const LossArgs& a, float* out, void* state, cudaStream_t s);
cudaError_t launch_loss_bwd(const LossArgs& a, const void* state, const float* grad_loss, float* dL_dimage,
                            float* dL_dstatic, float* dL_dgt, cudaStream_t s);

// Fused multi-tensor Adam (adam.cu, include/das3r_adam.h)
constexpr int ADAM_MAX_TENSORS = 16;
struct AdamTensor {
    float* param;
    const float* grad;
    float *exp_avg, *exp_avg_sq;
    int64_t n;
    float step_size;  // lr / (1 - beta1^t)
    int vec4;         // all four arrays 16-byte aligned
};
struct AdamArgs {
    int n_tensors;
    float b1w, beta2, b2w, bc2_sqrt, eps;  // 1-beta1, beta2, 1-beta2, sqrt(1-beta2^t), eps
    int block_start[ADAM_MAX_TENSORS + 1];
    AdamTensor t[ADAM_MAX_TENSORS];
};
int adam_chunk_elems();
cudaError_t launch_adam(const AdamArgs& a, int total_blocks, cudaStream_t s);

// forward.cu
cudaError_t launch_preprocess_fwd(const PreFwdArgs& a, bool stage_sh, cudaStream_t s);
cudaError_t launch_preprocess_fwd_multi(const MultiFwdArgs& m, cudaStream_t s);
// Binning and blend stages, batched over the views of `b` (one launch per stage, blockIdx.y = view).
cudaError_t launch_tile_scan(const BatchArgs& b, cudaStream_t s);
cudaError_t launch_scatter(const BatchArgs& b, cudaStream_t s);
// Per-tile sort + tile-ordered record streams. max_count = largest tile over all views. A view whose largest tile has
// more than tile_sort_capacity() entries needs temp[v] (tile_sort_fallback_bytes(R_v, T) bytes) for the library
// segmented sort that handles just those tiles; R[v] = that view's num_rendered.
int tile_sort_capacity();
size_t tile_sort_fallback_bytes(int R, int T);
cudaError_t launch_tile_sort(const BatchArgs& b, uint32_t max_count, const int* R, const uint32_t* view_max,
                             void* const* temp, const size_t* temp_bytes, cudaStream_t s);
cudaError_t launch_rect_lists(const BatchArgs& b, cudaStream_t s);
cudaError_t launch_blend_fwd(const BatchArgs& b, cudaStream_t s);
cudaError_t launch_mark_visible(int P, const float* means, const float* view, const float* proj, uint8_t* present,
                                cudaStream_t s);

// backward.cu
cudaError_t launch_blend_bwd(const BatchArgs& b, cudaStream_t s);
cudaError_t launch_preprocess_bwd(const PreBwdArgs& a, cudaStream_t s);
cudaError_t launch_preprocess_bwd_multi(const MultiBwdArgs& a, cudaStream_t s);

// One-wave launches must be spread evenly: the hardware block scheduler fills an SM up to its occupancy limit before
// moving on, so a grid smaller than one full wave leaves some SMs empty while others run at the limit (measured on
// blend_fwd: sm__cycles_active min 199k / max 717k). Returns the number of bytes of (unused) dynamic shared memory to
// request so that at most ceil(grid / #SMs) blocks fit on an SM; 0 when the grid spans several waves anyway.
size_t spread_pad_smem(const void* kernel, int grid, int block_threads, size_t static_smem);

// init.cu (include/das3r_init.h)
size_t pointmap_scratch_bytes(int64_t n_pixels);
cudaError_t launch_pointmap_select(const float* conf, const float* dyna, float log_thre, int64_t n, uint8_t* mask,
                                   float* conf_static, void* scratch, cudaStream_t s);
cudaError_t launch_pointmap_emit(int V, int H, int W, const float* K, const float* poses, const float* depth,
                                 const float* rgb, const uint8_t* mask, void* scratch, float* xyz, float* features_dc,
                                 float* features_rest, float* rotation, float* opacity, int64_t* conf_index,
                                 float opacity_raw, cudaStream_t s);
cudaError_t launch_pointmap_scales(int64_t P, const float* dist2, float* scaling, cudaStream_t s);

// comm.cu
cudaError_t launch_nvls_allreduce(float* mc, size_t n_f4, int rank, int world, int n_ctas, cudaStream_t s);

// knn.cu
size_t knn_scratch_bytes(int P);
cudaError_t knn_dist2(int P, const float* points, float* out, void* scratch, size_t scratch_bytes, cudaStream_t s);

}  // namespace das3r
