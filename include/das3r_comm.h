/* C ABI of the multi-GPU piece of the hot path: the sum of the step's gradient bucket over the ranks.
 *
 * The reference is single-GPU (scripts/rendering_sintel.sh:3,57: GPU_ID=0) - there is no reference counterpart. In the
 * data-parallel plan every rank holds a replica of the Gaussians, rasterizes its own views and the per-Gaussian parameter
 * gradients are summed across ranks once per step. NCCL can do that sum (das3r_b200/sharding.py uses it by default and on
 * machines without NVSwitch multicast); on an NVSwitch box this kernel does it in the switch:
 *
 *   rank r owns every world-th share of the range; for its share it issues multimem.ld_reduce (the switch reads the
 *   element from all ranks' copies of the symmetric buffer and returns the sum) followed by multimem.st (the switch
 *   writes the sum into all ranks' copies). Each byte crosses a GPU's NVLink once in each direction.
 *
 * The caller provides the multicast mapping of a buffer that is symmetric across the ranks (same size and offset on
 * every rank; e.g. torch.distributed._symmetric_memory: handle.multicast_ptr) and brackets the call with a barrier over
 * the ranks on `stream` (all ranks' slices complete before, all ranks' stores landed after). Same conventions as
 * das3r_raster.h: plain pointers, work enqueued on `stream`, 0 or a negative DAS3R_E* code, das3r_last_error().
 */
#ifndef DAS3R_COMM_H
#define DAS3R_COMM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* In-place sum over `world` ranks of n_floats fp32 values starting at multicast_base + offset_floats (both multiples of 4
 * floats = 16 bytes). n_ctas thread blocks of 512 threads are used (0: a default that leaves most SMs to the gradient
 * kernels running next to it). */
int das3r_nvls_allreduce_sum(void* multicast_base, size_t offset_floats, size_t n_floats, int32_t rank, int32_t world,
                             int32_t n_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif
