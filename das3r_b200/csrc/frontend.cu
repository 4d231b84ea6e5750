// Fused DAS3R render front-end for sm_100a (see include/das3r_frontend.h): pose transform of means and rotations for
// all views of a step, opacity / scale activations, SH concatenation - and the backward, which reduces the per-view
// mean / rotation gradients to parameter gradients and pose gradients in one pass over the Gaussians.
// Reference: gaussian_renderer/__init__.py:83-108,:126; utils/pose_utils.py:10-104; scene/gaussian_model.py:103-113,
// :186-194. HBM-bound: every array is read or written once, the per-view arrays once per view.
#include "common.cuh"
#include "kernels.cuh"

namespace das3r {

constexpr int FE_THREADS = 256;
constexpr int FE_POSE_F = 16;  // per view in shared memory: R (9, row-major), raw quaternion (4), translation (3)
constexpr int FE_PART_F = 16;  // per view and block: sum g (3), sum g x^T (9, row-major), quaternion-product part (4)

// Pose (r,x,y,z,tx,ty,tz) -> rotation matrix of the NORMALISED quaternion (utils/pose_utils.py:36-55), raw quaternion,
// translation.
__device__ __forceinline__ void pose_to_smem(const float* __restrict__ pose, float* __restrict__ o) {
    const float q0 = pose[0], q1 = pose[1], q2 = pose[2], q3 = pose[3];
    const float norm = sqrtf(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);
    const float r = q0 / norm, x = q1 / norm, y = q2 / norm, z = q3 / norm;
    o[0] = 1 - 2 * (y * y + z * z);
    o[1] = 2 * (x * y - r * z);
    o[2] = 2 * (x * z + r * y);
    o[3] = 2 * (x * y + r * z);
    o[4] = 1 - 2 * (x * x + z * z);
    o[5] = 2 * (y * z - r * x);
    o[6] = 2 * (x * z - r * y);
    o[7] = 2 * (y * z + r * x);
    o[8] = 1 - 2 * (x * x + y * y);
    o[9] = q0; o[10] = q1; o[11] = q2; o[12] = q3;
    o[13] = pose[4]; o[14] = pose[5]; o[15] = pose[6];
}

// Block-cooperative copy between the split SH layout (dc [P,1,3], rest [P,15,3]) and the joint one ([P,16,3]): flat,
// coalesced on the split side, 15 independent loads in flight per thread before the first store (a load -> store loop
// keeps one 4-byte access per thread in flight and ran at a third of the bandwidth). TO_JOINT: dc/rest -> shs; else
// shs -> dc/rest.
template <bool TO_JOINT, typename TD, typename TR, typename TS>
__device__ __forceinline__ void sh_rows_copy(TD* dc, TR* rest, TS* shs, int base, int nvalid) {
    const int tid = threadIdx.x;
    {
        float v[3];
        const int total = nvalid * 3;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const int e = tid + k * FE_THREADS;
            const int g = e / 3, j = e - 3 * g;
            if (e < total) v[k] = TO_JOINT ? dc[(size_t)base * 3 + e] : shs[(size_t)(base + g) * 48 + j];
        }
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const int e = tid + k * FE_THREADS;
            const int g = e / 3, j = e - 3 * g;
            if (e < total) {
                if constexpr (TO_JOINT) shs[(size_t)(base + g) * 48 + j] = v[k];
                else dc[(size_t)base * 3 + e] = v[k];
            }
        }
    }
    const int total = nvalid * 45;
#pragma unroll 1
    for (int c = 0; c < 3; c++) {
        float v[15];
#pragma unroll
        for (int k = 0; k < 15; k++) {
            const int e = tid + (c * 15 + k) * FE_THREADS;
            const int g = e / 45, j = e - 45 * g;
            if (e < total) v[k] = TO_JOINT ? rest[(size_t)base * 45 + e] : shs[(size_t)(base + g) * 48 + 3 + j];
        }
#pragma unroll
        for (int k = 0; k < 15; k++) {
            const int e = tid + (c * 15 + k) * FE_THREADS;
            const int g = e / 45, j = e - 45 * g;
            if (e < total) {
                if constexpr (TO_JOINT) shs[(size_t)(base + g) * 48 + 3 + j] = v[k];
                else rest[(size_t)base * 45 + e] = v[k];
            }
        }
    }
}

__global__ void __launch_bounds__(FE_THREADS) frontend_fwd_kernel(const __grid_constant__ FrontendFwdArgs a) {
    __shared__ float s_pose[FE_MAX_VIEWS * FE_POSE_F];
    const int tid = threadIdx.x;
    const int base = blockIdx.x * FE_THREADS;
    const int idx = base + tid;
    const int nvalid = min(FE_THREADS, a.P - base);
    if (tid < a.n_views) pose_to_smem(a.poses + 7 * tid, s_pose + FE_POSE_F * tid);
    // all 16 coefficients, no row selection: block-cooperative flat copy; otherwise every thread copies the active
    // coefficients of its own row (the rasterizer never reads the bands above the active degree)
    const bool flat = a.K == 16 && a.filter_index == nullptr;
    if (flat) sh_rows_copy<true>(a.features_dc, a.features_rest, a.shs, base, nvalid);
    __syncthreads();
    if (idx >= a.P) return;
    // gaussian_renderer/__init__.py:85-86,:95,:107,:126: `[filtering]` selects model rows; output row idx = model row src
    const size_t src = a.filter_index ? (size_t)a.filter_index[idx] : (size_t)idx;
    if (!flat) {
        float* row = a.shs + (size_t)idx * 48;
#pragma unroll
        for (int j = 0; j < 3; j++) row[j] = a.features_dc[src * 3 + j];
        const int nrest = 3 * (a.K - 1);
        for (int j = 0; j < nrest; j++) row[3 + j] = a.features_rest[src * 45 + j];
    }

    // activations (scene/gaussian_model.py:103-105,:193-194; gaussian_renderer/__init__.py:95-97)
    const float o = a.opacity_raw[src];
    const float conf = a.conf_static[a.conf_index ? a.conf_index[idx] : (int64_t)idx];
    a.opacity[idx] = (1.0f / (1.0f + expf(-o))) * conf;
#pragma unroll
    for (int k = 0; k < 3; k++) a.scales[3 * idx + k] = expf(a.scaling_raw[3 * src + k]);

    const float X = a.xyz[3 * src], Y = a.xyz[3 * src + 1], Z = a.xyz[3 * src + 2];
    const float4 rq = __ldg(reinterpret_cast<const float4*>(a.rotation) + src);  // (w2, x2, y2, z2)
#pragma unroll 1
    for (int v = 0; v < a.n_views; v++) {
        const float* p = s_pose + FE_POSE_F * v;
        // gaussian_renderer/__init__.py:87-89: rows of [R | t] times (X, Y, Z, 1)
        float* m = a.means3D[v] + 3 * (size_t)idx;
        m[0] = p[0] * X + p[1] * Y + p[2] * Z + p[13];
        m[1] = p[3] * X + p[4] * Y + p[5] * Z + p[14];
        m[2] = p[6] * X + p[7] * Y + p[8] * Z + p[15];
        // utils/pose_utils.py:94-101 with q1 = raw pose quaternion, q2 = Gaussian rotation
        const float w1 = p[9], x1 = p[10], y1 = p[11], z1 = p[12];
        float4 r;
        r.x = w1 * rq.x - x1 * rq.y - y1 * rq.z - z1 * rq.w;
        r.y = w1 * rq.y + x1 * rq.x + y1 * rq.w - z1 * rq.z;
        r.z = w1 * rq.z - x1 * rq.w + y1 * rq.x + z1 * rq.y;
        r.w = w1 * rq.w + x1 * rq.z - y1 * rq.y + z1 * rq.x;
        reinterpret_cast<float4*>(a.rotations[v])[idx] = r;
    }
}

cudaError_t launch_frontend_fwd(const FrontendFwdArgs& a, cudaStream_t s) {
    const int blocks = (a.P + FE_THREADS - 1) / FE_THREADS;
    frontend_fwd_kernel<<<blocks, FE_THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

// Sum of `v` over the block, result valid in thread 0. `red` holds one float per warp.
__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x == 0)
#pragma unroll
        for (int w = 0; w < FE_THREADS / 32; w++) t += red[w];
    __syncthreads();
    return t;
}

// Sums 16 per-lane values over the warp with 16 shuffles instead of 80: at every butterfly step a lane hands half of
// its values to its partner and keeps the other half. Returns the complete sum of value `idx` (a function of the
// lane's bits 4..1; both lanes of a pair hold the same result).
__device__ __forceinline__ float warp_reduce16(const float (&v)[16], int lane, int& idx) {
    const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4, h1 = lane & 2;
    float a[8], b[4], c[2];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = (h4 ? v[k + 8] : v[k]) + __shfl_xor_sync(0xffffffffu, h4 ? v[k] : v[k + 8], 16);
#pragma unroll
    for (int k = 0; k < 4; k++) b[k] = (h3 ? a[k + 4] : a[k]) + __shfl_xor_sync(0xffffffffu, h3 ? a[k] : a[k + 4], 8);
#pragma unroll
    for (int k = 0; k < 2; k++) c[k] = (h2 ? b[k + 2] : b[k]) + __shfl_xor_sync(0xffffffffu, h2 ? b[k] : b[k + 2], 4);
    float d = (h1 ? c[1] : c[0]) + __shfl_xor_sync(0xffffffffu, h1 ? c[0] : c[1], 2);
    d += __shfl_xor_sync(0xffffffffu, d, 1);
    idx = (h4 ? 8 : 0) + (h3 ? 4 : 0) + (h2 ? 2 : 0) + (h1 ? 1 : 0);
    return d;
}

__global__ void __launch_bounds__(FE_THREADS) frontend_bwd_kernel(const __grid_constant__ FrontendBwdArgs a) {
    __shared__ float s_pose[FE_MAX_VIEWS * FE_POSE_F];
    __shared__ float s_warp[FE_MAX_VIEWS][FE_THREADS / 32][FE_PART_F];  // per view: every warp's 16 sums
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int base = a.row0 + blockIdx.x * FE_THREADS;
    const int idx = base + tid;
    const int nvalid = min(FE_THREADS, a.row_end - base);
    const bool in_range = idx < a.row_end;
    if (tid < a.n_views) pose_to_smem(a.poses + 7 * tid, s_pose + FE_POSE_F * tid);
    const bool flat = a.K == 16 && a.dsh_coeffs == 16 && a.rest_coeffs == 15 && a.filter_index == nullptr;
    if (flat)  // the split-layout outputs may be windows that start at row out_row0
        sh_rows_copy<false>(a.dL_dfeatures_dc - 3 * (size_t)a.out_row0, a.dL_dfeatures_rest - 45 * (size_t)a.out_row0,
                            a.dL_dshs, base, nvalid);
    __syncthreads();
    // output row idx of the forward = model row src (render()'s `[filtering]`); parameter gradients go to row src
    const size_t src = (in_range && a.filter_index) ? (size_t)a.filter_index[idx] : (size_t)idx;
    const size_t dst = src - (size_t)a.out_row0;  // row of the parameter-gradient outputs

    float X = 0.f, Y = 0.f, Z = 0.f;
    float4 rq = {0.f, 0.f, 0.f, 0.f};
    if (in_range) {
        if (!flat) {
            // dL_dshs rows have dsh_coeffs coefficients of which K are active; the rest rows written have rest_coeffs
            const float* g = a.dL_dshs + (size_t)idx * 3 * a.dsh_coeffs;
#pragma unroll
            for (int j = 0; j < 3; j++) a.dL_dfeatures_dc[dst * 3 + j] = g[j];
            const int nact = 3 * (a.K - 1), nrest = 3 * a.rest_coeffs;
            float* r = a.dL_dfeatures_rest + dst * (size_t)nrest;
            for (int j = 0; j < nrest; j++) r[j] = (j < nact) ? g[3 + j] : 0.f;
        }
        // activation derivatives
        const float o = a.opacity_raw[src];
        const float sg = 1.0f / (1.0f + expf(-o));
        const float conf = a.conf_static[a.conf_index ? a.conf_index[idx] : (int64_t)idx];
        const float go = a.dL_dopacity[idx];
        a.dL_dopacity_raw[dst] = go * conf * sg * (1.0f - sg);
        a.dL_dconf[idx - a.out_row0] = go * sg;
#pragma unroll
        for (int k = 0; k < 3; k++)
            a.dL_dscaling_raw[3 * dst + k] = a.dL_dscales[3 * idx + k] * expf(a.scaling_raw[3 * src + k]);
        X = a.xyz[3 * src]; Y = a.xyz[3 * src + 1]; Z = a.xyz[3 * src + 2];
        rq = __ldg(reinterpret_cast<const float4*>(a.rotation) + src);
    }

    float dX = 0.f, dY = 0.f, dZ = 0.f;
    float4 dr = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 2
    for (int v = 0; v < a.n_views; v++) {
        const float* p = s_pose + FE_POSE_F * v;
        float part[FE_PART_F];
#pragma unroll
        for (int k = 0; k < FE_PART_F; k++) part[k] = 0.f;
        if (in_range) {
            const float* gm = a.dL_dmeans3D[v] + 3 * (size_t)idx;
            const float g0 = gm[0], g1 = gm[1], g2 = gm[2];
            const float4 d = __ldg(reinterpret_cast<const float4*>(a.dL_drotations[v]) + idx);
            // means = R X + t
            dX += p[0] * g0 + p[3] * g1 + p[6] * g2;
            dY += p[1] * g0 + p[4] * g1 + p[7] * g2;
            dZ += p[2] * g0 + p[5] * g1 + p[8] * g2;
            part[0] = g0; part[1] = g1; part[2] = g2;
            part[3] = g0 * X; part[4] = g0 * Y; part[5] = g0 * Z;
            part[6] = g1 * X; part[7] = g1 * Y; part[8] = g1 * Z;
            part[9] = g2 * X; part[10] = g2 * Y; part[11] = g2 * Z;
            // rotations = q1 (x) q2: gradient w.r.t. q2 (the Gaussian) ...
            const float w1 = p[9], x1 = p[10], y1 = p[11], z1 = p[12];
            dr.x += d.x * w1 + d.y * x1 + d.z * y1 + d.w * z1;
            dr.y += -d.x * x1 + d.y * w1 + d.z * z1 - d.w * y1;
            dr.z += -d.x * y1 - d.y * z1 + d.z * w1 + d.w * x1;
            dr.w += -d.x * z1 + d.y * y1 - d.z * x1 + d.w * w1;
            // ... and w.r.t. q1 (the raw pose quaternion)
            part[12] = d.x * rq.x + d.y * rq.y + d.z * rq.z + d.w * rq.w;
            part[13] = -d.x * rq.y + d.y * rq.x - d.z * rq.w + d.w * rq.z;
            part[14] = -d.x * rq.z + d.y * rq.w + d.z * rq.x - d.w * rq.y;
            part[15] = -d.x * rq.w - d.y * rq.z + d.z * rq.y + d.w * rq.x;
        }
        // the warp's sums of the 16 partials of this view (the block sum follows after the loop: no barrier in here)
        int pi;
        const float ps = warp_reduce16(part, lane, pi);
        if ((lane & 1) == 0) s_warp[v][warp][pi] = ps;
    }
    if (in_range) {
        a.dL_dxyz[3 * dst] = dX; a.dL_dxyz[3 * dst + 1] = dY; a.dL_dxyz[3 * dst + 2] = dZ;
        reinterpret_cast<float4*>(a.dL_drotation)[dst] = dr;
    }
    // this block's partial sums -> scratch[block][view][16] (summed by frontend_pose_kernel)
    __syncthreads();
    for (int e = tid; e < a.n_views * FE_PART_F; e += FE_THREADS) {
        const int v = e / FE_PART_F, k = e - v * FE_PART_F;
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < FE_THREADS / 32; w++) t += s_warp[v][w][k];
        a.partials[(size_t)(a.row0 / FE_THREADS + blockIdx.x) * a.n_views * FE_PART_F + e] = t;
    }
}

// One block per view: sums the per-block partials in a fixed order and turns them into dL/d(pose) (7 values).
__global__ void __launch_bounds__(FE_THREADS) frontend_pose_kernel(const float* __restrict__ partials, int n_blocks,
                                                                  int n_views, const float* __restrict__ poses,
                                                                  float* __restrict__ dL_dposes) {
    __shared__ float red[FE_THREADS / 32][FE_PART_F];
    __shared__ float s_sum[FE_PART_F];
    const int v = blockIdx.x;
    // every thread sums its blocks' 16 partials (fixed assignment and order: run-to-run identical), then one shuffle
    // tree per value and a fixed-order sum over the warps
    float acc[FE_PART_F];
#pragma unroll
    for (int k = 0; k < FE_PART_F; k++) acc[k] = 0.f;
    for (int b = threadIdx.x; b < n_blocks; b += FE_THREADS) {
        const float4* p4 = reinterpret_cast<const float4*>(partials + ((size_t)b * n_views + v) * FE_PART_F);
#pragma unroll
        for (int q = 0; q < FE_PART_F / 4; q++) {
            const float4 t = p4[q];
            acc[4 * q] += t.x; acc[4 * q + 1] += t.y; acc[4 * q + 2] += t.z; acc[4 * q + 3] += t.w;
        }
    }
#pragma unroll
    for (int k = 0; k < FE_PART_F; k++) {
        float t = acc[k];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = t;
    }
    __syncthreads();
    if (threadIdx.x < FE_PART_F) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < FE_THREADS / 32; w++) t += red[w][threadIdx.x];
        s_sum[threadIdx.x] = t;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const float* pose = poses + 7 * v;
    const float q0 = pose[0], q1 = pose[1], q2 = pose[2], q3 = pose[3];
    const float norm = sqrtf(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);
    const float r = q0 / norm, x = q1 / norm, y = q2 / norm, z = q3 / norm;
    const float* M = s_sum + 3;  // M[3*a+b] = sum_i g_a X_b = dL/dR[a][b]
    // dL/d(normalised quaternion) = sum_ab M_ab dR_ab/dq (R as in utils/pose_utils.py:47-55)
    const float dr_ = 2 * (-z * M[1] + y * M[2] + z * M[3] - x * M[5] - y * M[6] + x * M[7]);
    const float dx_ = 2 * (y * M[1] + z * M[2] + y * M[3] - 2 * x * M[4] - r * M[5] + z * M[6] + r * M[7] - 2 * x * M[8]);
    const float dy_ = 2 * (-2 * y * M[0] + x * M[1] + r * M[2] + x * M[3] + z * M[5] - r * M[6] + z * M[7] - 2 * y * M[8]);
    const float dz_ = 2 * (-2 * z * M[0] - r * M[1] + x * M[2] + r * M[3] - 2 * z * M[4] + y * M[5] + x * M[6] + y * M[7]);
    // through q / |q|
    const float dot = r * dr_ + x * dx_ + y * dy_ + z * dz_;
    float* o = dL_dposes + 7 * v;
    o[0] = (dr_ - r * dot) / norm + s_sum[12];
    o[1] = (dx_ - x * dot) / norm + s_sum[13];
    o[2] = (dy_ - y * dot) / norm + s_sum[14];
    o[3] = (dz_ - z * dot) / norm + s_sum[15];
    o[4] = s_sum[0];
    o[5] = s_sum[1];
    o[6] = s_sum[2];
}

size_t frontend_backward_scratch_bytes(int P, int n_views) {
    const size_t blocks = (size_t)(P + FE_THREADS - 1) / FE_THREADS;
    return (blocks * (size_t)n_views * FE_PART_F * sizeof(float) + 255) / 256 * 256;
}

cudaError_t launch_frontend_bwd(const FrontendBwdArgs& a, float* dL_dposes, cudaStream_t s) {
    const int blocks = (a.row_end - a.row0 + FE_THREADS - 1) / FE_THREADS;
    if (blocks > 0) {
        frontend_bwd_kernel<<<blocks, FE_THREADS, 0, s>>>(a);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (dL_dposes) {
        const int all_blocks = (a.P + FE_THREADS - 1) / FE_THREADS;
        frontend_pose_kernel<<<a.n_views, FE_THREADS, 0, s>>>(a.partials, all_blocks, a.n_views, a.poses, dL_dposes);
    }
    return cudaGetLastError();
}

}  // namespace das3r
