/* C ABI of the fused DAS3R photometric loss (loss part): what train_gui.py:565-577 computes
 * from the rendered image with about forty torch kernels (two multiplies by the static mask, abs, five depthwise 11x11
 * convolutions, a dozen element-wise ops, three reductions) and autograd - in two kernels forward and one backward.
 *
 *   image = image * static ; gt = gt * static                                  train_gui.py:567-569
 *   Ll1   = |image - gt|                                  (no reduction)        utils/loss_utils.py:17-21
 *   Lssim = ssim(image, gt, size_average=False)           (11x11 Gaussian window, sigma 1.5, zero padding, C1 = 0.01^2,
 *                                                          C2 = 0.03^2)         utils/loss_utils.py:26-67
 *   psnr  = mean_c 20 log10(1 / sqrt(mean((image - gt)^2)))                     utils/image_utils.py:17-19, train_gui.py:573
 *   loss  = mean((1 - lambda) * Ll1 + lambda * (1 - Lssim))                     train_gui.py:575-576
 *
 * The window is applied as two 1-D passes with the reference's float32 1-D weights (the reference multiplies them into
 * a 2-D window first: same numbers up to float rounding). Same conventions as das3r_raster.h: extern "C", device
 * pointers, caller owns all memory, work is enqueued on `stream`, 0 or a negative DAS3R_E* code, das3r_last_error().
 */
#ifndef DAS3R_LOSS_H
#define DAS3R_LOSS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct das3r_loss_inputs {
    int32_t width, height;
    float lambda_dssim;        /* opt.lambda_dssim (arguments/__init__.py:83: 0.2) */
    const float* image;        /* [3,H,W] rendered image */
    const float* gt;           /* [3,H,W] ground truth */
    const float* static_mask;  /* [H,W] _conf_static[uid], or NULL for all ones */
} das3r_loss_inputs;

/* Bytes of the state buffer the forward fills for the backward (15 derivative maps of H*W floats + block partials). */
size_t das3r_loss_state_bytes(int32_t width, int32_t height);

/* out[0] = loss, out[1] = psnr (device floats). Deterministic: block partial sums are added in a fixed order. */
int das3r_loss_forward(const das3r_loss_inputs* in, float* out, void* state, void* stream);

/* Gradient of `loss` times *grad_loss (a device float; NULL = 1) w.r.t. the image, and optionally w.r.t. the static
 * mask (it scales both image and gt) and the ground truth. Every element of the given outputs is written. */
int das3r_loss_backward(const das3r_loss_inputs* in, const void* state, const float* grad_loss, float* dL_dimage,
                        float* dL_dstatic, float* dL_dgt, void* stream);

#ifdef __cplusplus
}
#endif
#endif
