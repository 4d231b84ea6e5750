/* C ABI of the fused Adam step (optimizer part): DAS3R's
 *   self.optimizer = torch.optim.Adam(l, lr=0.0, eps=1e-15)          scene/gaussian_model.py:236-258 (seven groups,
 *   gaussians.optimizer.step()                                        one learning rate each), train_gui.py:580-586
 * for all parameter tensors of the optimizer in ONE launch: every element of parameter, gradient and the two moment
 * buffers is read once and parameter and moments are written once (28 bytes per element), instead of the five to six
 * passes per tensor of the library's default multi-tensor path.
 *
 * Semantics = torch.optim.Adam without weight decay / amsgrad / maximize (torch/optim/adam.py, _single_tensor_adam):
 *   m <- m + (1-b1) (g - m);  v <- b2 v + (1-b2) g g;  p <- p - (lr / (1-b1^t)) * m / (sqrt(v) / sqrt(1-b2^t) + eps)
 * Same conventions as das3r_raster.h (extern "C", device pointers, caller owns memory, stream, DAS3R_E* codes).
 */
#ifndef DAS3R_ADAM_H
#define DAS3R_ADAM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DAS3R_ADAM_MAX_TENSORS 16 /* per call */

typedef struct das3r_adam_tensor {
    float* param;        /* [n] updated in place */
    const float* grad;   /* [n] */
    float* exp_avg;      /* [n] first moment, updated in place */
    float* exp_avg_sq;   /* [n] second moment, updated in place */
    int64_t n;           /* elements */
    double lr;           /* the group's learning rate (a Python float in torch) */
} das3r_adam_tensor;

/* `step` is the 1-based step count AFTER this update (torch increments state['step'] before using it). */
/* betas and eps are doubles like the Python floats torch.optim.Adam holds: 1 - beta is formed in double before it is
 * rounded to float32 (1 - (float)0.999 differs from (float)0.001 by 1.3e-5 relative). */
int das3r_adam_step(int32_t n_tensors, const das3r_adam_tensor* tensors, double beta1, double beta2, double eps,
                    int64_t step, void* stream);

#ifdef __cplusplus
}
#endif
#endif
