/* C ABI of the fused DAS3R render front-end: everything DAS3R's render() does to the
 * Gaussian parameters between the model and the rasterizer call, for all views of a step at once, and its backward.
 *
 * The reference does this with about ten separate torch kernels per view (clone, index, cat, matmul, unbind/stack,
 * sigmoid, exp, mul): reference gaussian_renderer/__init__.py:83-108 and :126,
 *   means3D   = (get_camera_from_tensor(pose) @ [xyz | 1]^T)^T[:, :3]        (:83-90, utils/pose_utils.py:57-85, :10-55)
 *   rotations = quadmultiply(pose[:4], _rotation)                            (:91, utils/pose_utils.py:86-104)
 *   opacity   = sigmoid(_opacity) * _conf_static.reshape(-1,1)[aggregated_mask]   (:95-97, scene/gaussian_model.py:194)
 *   scales    = exp(_scaling)                                                (:106, scene/gaussian_model.py:103-105)
 *   shs       = cat(_features_dc, _features_rest, dim=1)                     (:126, scene/gaussian_model.py:186-190)
 * Note the reference's asymmetry, kept here: the rotation matrix that moves the means is built from the NORMALISED
 * pose quaternion (quad2rotation divides by the norm), the quaternion product uses the RAW one.
 *
 * Same conventions as das3r_raster.h: extern "C", plain device pointers, caller owns all memory, work is enqueued on
 * `stream`, return 0 or a negative DAS3R_E* code with das3r_last_error(). No CPU path.
 */
#ifndef DAS3R_FRONTEND_H
#define DAS3R_FRONTEND_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DAS3R_FRONTEND_MAX_VIEWS 16 /* views per call */

typedef struct das3r_frontend_model {
    int32_t P;                   /* Gaussians */
    const float* xyz;            /* [P,3]   _xyz */
    const float* rotation;       /* [P,4]   _rotation (raw, un-normalised; r,x,y,z) */
    const float* opacity;        /* [P,1]   _opacity (pre-sigmoid) */
    const float* scaling;        /* [P,3]   _scaling (log) */
    const float* features_dc;    /* [P,1,3] _features_dc */
    const float* features_rest;  /* [P,15,3] _features_rest */
    const float* conf_static;    /* flat _conf_static (one value per pointmap pixel) */
    const int64_t* conf_index;   /* [P] position of Gaussian i in conf_static = nonzero(aggregated_mask); NULL: i.
                                    Precondition (not checked, it would cost a device round trip): every entry is a
                                    valid index into conf_static */
} das3r_frontend_model;

typedef struct das3r_frontend_outputs {
    float* opacity;    /* [P,1]    sigmoid(_opacity) * conf */
    float* scales;     /* [P,3]    exp(_scaling) */
    float* shs;        /* [P,16,3] {dc, rest} */
    float* means3D[DAS3R_FRONTEND_MAX_VIEWS];   /* per view [P,3] */
    float* rotations[DAS3R_FRONTEND_MAX_VIEWS]; /* per view [P,4] */
} das3r_frontend_outputs;

/* Gradients arriving from the rasterizer (das3r_raster_backward_finish: the three shared ones summed over the views,
 * means / rotations per view). */
typedef struct das3r_frontend_grads_in {
    const float* dL_dopacity;  /* [P,1] */
    const float* dL_dscales;   /* [P,3] */
    const float* dL_dshs;      /* [P,16,3] */
    const float* dL_dmeans3D[DAS3R_FRONTEND_MAX_VIEWS];   /* per view [P,3] */
    const float* dL_drotations[DAS3R_FRONTEND_MAX_VIEWS]; /* per view [P,4] */
} das3r_frontend_grads_in;

/* Gradients w.r.t. the model parameters and the poses; every element is written. */
typedef struct das3r_frontend_grads_out {
    float* dL_dxyz;            /* [P,3] */
    float* dL_drotation;       /* [P,4] */
    float* dL_dopacity;        /* [P,1] */
    float* dL_dscaling;        /* [P,3] */
    float* dL_dfeatures_dc;    /* [P,1,3] */
    float* dL_dfeatures_rest;  /* [P,15,3] */
    float* dL_dconf;           /* [P]: gradient w.r.t. conf_static[conf_index[i]] (the caller scatters it; indices are
                                  unique) */
    float* dL_dposes;          /* [n_views,7] */
    float* scratch;            /* das3r_frontend_backward_scratch_bytes(P, n_views) bytes */
} das3r_frontend_grads_out;

/* poses: [n_views,7] device array, {quaternion r,x,y,z ; translation} per view (camera_pose in render()). */
int das3r_frontend_forward(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                           const das3r_frontend_outputs* out, void* stream);

size_t das3r_frontend_backward_scratch_bytes(int32_t P, int32_t n_views);

/* Backward of das3r_frontend_forward: one pass over the Gaussians turns the per-view dL_dmeans3D / dL_drotations into
 * dL_dxyz / dL_drotation and per-block partial sums for the pose gradients, and applies the activation derivatives; a
 * second, tiny kernel reduces the partial sums (deterministically, no atomics) to dL_dposes. */
int das3r_frontend_backward(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                            const das3r_frontend_grads_in* gin, const das3r_frontend_grads_out* gout, void* stream);

/* Options of the _ex variants below (no reference counterpart beyond the lines cited). Zero-initialise, then set:
 *   sh_degree     active SH degree D (pc.active_sh_degree, gaussian_renderer/__init__.py:73): only the (D+1)^2 active
 *                 coefficients are copied into / out of the joint SH rows; the rasterizer never touches the others
 *                 (forward.cu:20-71). Default 3.
 *   filter_index  render()'s boolean `filtering` (gaussian_renderer/__init__.py:38-46,:85-86,:95,:107,:126) as the int64
 *                 device array of selected model rows, n_rows entries; NULL = all rows (n_rows = model->P). Outputs of
 *                 the forward have n_rows rows; the backward writes parameter gradients to the selected rows only (the
 *                 caller zero-fills the gradient buffers). conf_index / conf_static are indexed by OUTPUT row, as the
 *                 reference's `opacity * _conf_static[...]` product is.
 *   dsh_coeffs    backward: coefficients per row of gin->dL_dshs, (D+1)^2 .. 16 (0 = 16), matching
 *                 das3r_raster_backward_finish_ex.
 *   rest_coeffs   backward: coefficients per row of gout->dL_dfeatures_rest: 15 (0 = 15; inactive bands written as zeros)
 *                 or trimmed to the (D+1)^2 - 1 active ones for a gradient bucket that skips the zero bands (pass that
 *                 number, or -1; at degree 0 the rows are empty and the pointer may be NULL).
 *   first, count  backward: only output rows [first, first+count) (count < 0: to the end; first a multiple of 256).
 *   local_outputs backward: nonzero = the parameter-gradient pointers of gout (and dL_dconf) address row `first` (windows
 *                 of a slice-major gradient bucket) instead of row 0. Not with filter_index.
 *   finish_poses  backward: after this range, reduce the per-block partial sums of ALL rows to gout->dL_dposes (set on
 *                 the last range of a step; every range must have been processed by then). */
typedef struct das3r_frontend_options {
    int32_t sh_degree;
    int32_t n_rows;
    const int64_t* filter_index;
    int32_t dsh_coeffs, rest_coeffs;
    int32_t first, count;
    int32_t finish_poses;
    int32_t local_outputs;
} das3r_frontend_options;

int das3r_frontend_forward_ex(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                              const das3r_frontend_outputs* out, const das3r_frontend_options* opt, void* stream);
int das3r_frontend_backward_ex(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                               const das3r_frontend_grads_in* gin, const das3r_frontend_grads_out* gout,
                               const das3r_frontend_options* opt, void* stream);

#ifdef __cplusplus
}
#endif
#endif
