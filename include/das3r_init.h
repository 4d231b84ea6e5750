/* C ABI of the pointmap initialisation: what GaussianModel.create_from_cameras does between the
 * stacked per-view maps and the parameter tensors, reference scene/gaussian_model.py:573-659 -
 *   p3d   = depth_to_pts3d(K, poses, W, H, depth)            (:614, utils/pose_utils.py:572-583, :672-683, :609-659)
 *   mask  = confidence.reshape(-1) > log(conf_thre)          (:622-626)  -> aggregated_mask
 *   xyz   = p3d.reshape(-1,3)[mask] ; colour = rgb.permute(0,2,3,1).reshape(-1,3)[mask]           (:629-630)
 *   f_dc  = RGB2SH(colour) ; f_rest = 0                      (:634-637, utils/sh_utils.py:114-115)
 *   scale = log(sqrt(clamp_min(distCUDA2(xyz), 1e-7))) x3    (:641-642)
 *   rot   = (1,0,0,0) ; opacity = inverse_sigmoid(1/V)       (:643-647, utils/general_utils.py:18-19)
 *   conf_static = 1 - dyna_avg                               (:620, :659)
 * - as two passes over the V*H*W pixels plus the 3-NN search, instead of ~25 torch kernels with boolean-mask gathers.
 *
 * Same conventions as das3r_raster.h: extern "C", plain device pointers, caller owns all memory, work is enqueued on
 * `stream`, return >= 0 or a negative DAS3R_E* code with das3r_last_error(). No CPU path.
 */
#ifndef DAS3R_INIT_H
#define DAS3R_INIT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct das3r_pointmaps {
    int32_t V, H, W;
    const float* K;          /* [V,3,3] intrinsics; fx = K[0][0] (the reference asserts fx == fy), pp = K[0:2][2] */
    const float* poses;      /* [V,4,4] camera-to-world (camera.original_pose) */
    const float* depth;      /* [V,H,W] */
    const float* confidence; /* [V,H,W] */
    const float* dyna_avg;   /* [V,H,W] averaged dynamic mask */
    const float* rgb;        /* [V,3,H,W] */
    float log_conf_thre;     /* log(conf_thre), computed by the caller in fp32 */
} das3r_pointmaps;

/* Bytes of device scratch for das3r_pointmap_select / das3r_pointmap_init with n_pixels = V*H*W. */
size_t das3r_pointmap_scratch_bytes(int64_t n_pixels);

/* Pass 1: aggregated_mask[V*H*W] (uint8 0/1), conf_static[V*H*W] = 1 - dyna_avg, and the compaction offsets (kept in
 * scratch for pass 2). Returns the number of selected pixels P (blocks the host once on a 4-byte read-back: the caller
 * has to size the parameter tensors), or a negative code. */
int64_t das3r_pointmap_select(const das3r_pointmaps* in, uint8_t* aggregated_mask, float* conf_static, void* scratch,
                              size_t scratch_bytes, void* stream);

typedef struct das3r_init_outputs {
    float* xyz;            /* [P,3] */
    float* features_dc;    /* [P,1,3] */
    float* features_rest;  /* [P,15,3] (zeros) or NULL */
    float* scaling;        /* [P,3] */
    float* rotation;       /* [P,4] */
    float* opacity;        /* [P,1] */
    int64_t* conf_index;   /* [P] pixel index of Gaussian i = nonzero(aggregated_mask), or NULL */
    float* dist2;          /* [P] mean squared 3-NN distance before the clamp, or NULL */
} das3r_init_outputs;

/* Pass 2 (after das3r_pointmap_select on the same scratch and stream): unprojects the selected pixels, fills the
 * parameter tensors in pixel order, runs the 3-NN search (das3r_knn_dist2) on the new points and converts it to the log
 * scales. knn_scratch: das3r_knn_scratch_bytes(P) bytes. */
int das3r_pointmap_init(const das3r_pointmaps* in, const uint8_t* aggregated_mask, int64_t P, const das3r_init_outputs* out,
                        void* scratch, size_t scratch_bytes, void* knn_scratch, size_t knn_scratch_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif
