// In-switch all-reduce of a slice of the gradient bucket (include/das3r_comm.h): multimem.ld_reduce + multimem.st over
// the NVSwitch multicast mapping of a buffer that is symmetric across the ranks. No reference counterpart (the
// reference is single-GPU).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace das3r {

constexpr int AR_THREADS = 512;

constexpr int AR_UNROLL = 4;  // independent 16-byte reductions in flight per thread

__device__ __forceinline__ float4 mm_ld_reduce(const float* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(float* p, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

__global__ void __launch_bounds__(AR_THREADS) nvls_allreduce_kernel(float* __restrict__ mc, size_t n_f4, int rank,
                                                                    int world) {
    // this rank's share: one contiguous run of float4s; consecutive threads touch consecutive 16-byte elements, every
    // thread keeps AR_UNROLL reductions in flight before it stores them (the switch round trip is several microseconds)
    const size_t per = (n_f4 + (size_t)world - 1) / (size_t)world;
    const size_t begin = (size_t)rank * per;
    const size_t end = begin + per < n_f4 ? begin + per : n_f4;
    const size_t stride = (size_t)gridDim.x * AR_THREADS;
    for (size_t i0 = begin + (size_t)blockIdx.x * AR_THREADS + threadIdx.x; i0 < end; i0 += stride * AR_UNROLL) {
        float4 v[AR_UNROLL];
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) {
            const size_t i = i0 + (size_t)u * stride;
            if (i < end) v[u] = mm_ld_reduce(mc + 4 * i);
        }
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) {
            const size_t i = i0 + (size_t)u * stride;
            if (i < end) mm_st(mc + 4 * i, v[u]);
        }
    }
}

cudaError_t launch_nvls_allreduce(float* mc, size_t n_f4, int rank, int world, int n_ctas, cudaStream_t s) {
    if (n_f4 == 0) return cudaSuccess;
    nvls_allreduce_kernel<<<n_ctas, AR_THREADS, 0, s>>>(mc, n_f4, rank, world);
    return cudaGetLastError();
}

}  // namespace das3r
