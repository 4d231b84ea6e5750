// Fused multi-tensor Adam step for sm_100a (include/das3r_adam.h): one launch over all parameter tensors of DAS3R's
// optimizer (scene/gaussian_model.py:236-258, train_gui.py:580-586). HBM-bound: 16 bytes read + 12 bytes written per
// element, 128-bit accesses where the tensor allows it.
#include "common.cuh"
#include "kernels.cuh"

namespace das3r {

constexpr int AD_THREADS = 256;
constexpr int AD_PER_THREAD = 8;                          // elements per thread: two float4 per array in flight
constexpr int AD_CHUNK = AD_THREADS * AD_PER_THREAD;      // elements per block

__device__ __forceinline__ void adam_one(float& p, const float g, float& m, float& v, const float b1w, const float b2,
                                         const float b2w, const float step_size, const float bc2_sqrt, const float eps) {
    m = m + b1w * (g - m);               // exp_avg.lerp_(grad, 1 - beta1)
    v = v * b2 + b2w * g * g;            // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    const float denom = sqrtf(v) / bc2_sqrt + eps;
    p = p - step_size * (m / denom);     // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(AD_THREADS) adam_kernel(const __grid_constant__ AdamArgs a) {
    // which tensor does this block belong to (at most 16 entries: linear scan)
    int t = 0;
    while (t + 1 < a.n_tensors && (int)blockIdx.x >= a.block_start[t + 1]) t++;
    const AdamTensor& T = a.t[t];
    const int64_t base = (int64_t)(blockIdx.x - a.block_start[t]) * AD_CHUNK;
    const float step_size = T.step_size;
    if (T.vec4 && base + AD_CHUNK <= T.n) {
        float4* P = reinterpret_cast<float4*>(T.param + base);
        const float4* G = reinterpret_cast<const float4*>(T.grad + base);
        float4* M = reinterpret_cast<float4*>(T.exp_avg + base);
        float4* V = reinterpret_cast<float4*>(T.exp_avg_sq + base);
        float4 p[2], g[2], m[2], v[2];
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int i = threadIdx.x + k * AD_THREADS;
            p[k] = P[i]; g[k] = G[i]; m[k] = M[i]; v[k] = V[i];
        }
#pragma unroll
        for (int k = 0; k < 2; k++) {
            adam_one(p[k].x, g[k].x, m[k].x, v[k].x, a.b1w, a.beta2, a.b2w, step_size, a.bc2_sqrt, a.eps);
            adam_one(p[k].y, g[k].y, m[k].y, v[k].y, a.b1w, a.beta2, a.b2w, step_size, a.bc2_sqrt, a.eps);
            adam_one(p[k].z, g[k].z, m[k].z, v[k].z, a.b1w, a.beta2, a.b2w, step_size, a.bc2_sqrt, a.eps);
            adam_one(p[k].w, g[k].w, m[k].w, v[k].w, a.b1w, a.beta2, a.b2w, step_size, a.bc2_sqrt, a.eps);
            const int i = threadIdx.x + k * AD_THREADS;
            P[i] = p[k]; M[i] = m[k]; V[i] = v[k];
        }
    } else {
        for (int64_t i = base + threadIdx.x; i < min(base + (int64_t)AD_CHUNK, T.n); i += AD_THREADS) {
            float p = T.param[i], m = T.exp_avg[i], v = T.exp_avg_sq[i];
            adam_one(p, T.grad[i], m, v, a.b1w, a.beta2, a.b2w, step_size, a.bc2_sqrt, a.eps);
            T.param[i] = p; T.exp_avg[i] = m; T.exp_avg_sq[i] = v;
        }
    }
}

int adam_chunk_elems() { return AD_CHUNK; }

cudaError_t launch_adam(const AdamArgs& a, int total_blocks, cudaStream_t s) {
    if (total_blocks <= 0) return cudaSuccess;
    adam_kernel<<<total_blocks, AD_THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace das3r
