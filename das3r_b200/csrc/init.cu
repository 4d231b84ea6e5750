// Pointmap initialisation for sm_100a (include/das3r_init.h): GaussianModel.create_from_cameras (reference
// scene/gaussian_model.py:573-659) as two passes over the pixels of all views + the 3-NN search. HBM-bound: every map
// is read once, every parameter tensor written once.
#include "common.cuh"
#include "kernels.cuh"

namespace das3r {

constexpr int PM_THREADS = 256;
constexpr int PM_PER_THREAD = 4;                      // consecutive pixels per thread
constexpr int PM_BLOCK_PIX = PM_THREADS * PM_PER_THREAD;

// Pass 1: mask + conf_static + number of selected pixels per block.
__global__ void __launch_bounds__(PM_THREADS) pointmap_select_kernel(const float* __restrict__ conf,
                                                                     const float* __restrict__ dyna, float log_thre,
                                                                     int64_t n, uint8_t* __restrict__ mask,
                                                                     float* __restrict__ conf_static,
                                                                     uint32_t* __restrict__ block_counts) {
    __shared__ uint32_t warp_cnt[PM_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * PM_BLOCK_PIX + (int64_t)threadIdx.x * PM_PER_THREAD;
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < PM_PER_THREAD; k++) {
        const int64_t i = base + k;
        if (i < n) {
            const bool m = conf[i] > log_thre;  // gaussian_model.py:624
            mask[i] = m ? 1 : 0;
            conf_static[i] = 1.0f - dyna[i];    // :620
            c += m ? 1u : 0u;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < PM_THREADS / 32; w++) t += warp_cnt[w];
        block_counts[blockIdx.x] = t;
    }
}

// Exclusive scan of the block counts in place (one block; a few thousand entries) + the total.
__global__ void __launch_bounds__(1024) pointmap_scan_kernel(uint32_t* __restrict__ counts, int n_blocks,
                                                             unsigned long long* __restrict__ total) {
    __shared__ unsigned long long warp_sum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (n_blocks + 1023) / 1024;
    const int beg = tid * per, end = min(n_blocks, beg + per);
    unsigned long long sum = 0;
    for (int i = beg; i < end; i++) sum += counts[i];
    unsigned long long incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = warp_sum[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        warp_sum[lane] = wi - w;
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    unsigned long long run = warp_sum[warp] + incl - sum;
    for (int i = beg; i < end; i++) {
        const uint32_t c = counts[i];
        counts[i] = (uint32_t)run;  // P < 2^31 is checked by the host
        run += c;
    }
}

struct PointmapEmitArgs {
    int V, H, W;
    int64_t n;
    const float *K, *poses, *depth, *rgb;
    const uint8_t* mask;
    const uint32_t* block_offsets;
    float *xyz, *features_dc, *features_rest, *rotation, *opacity;
    int64_t* conf_index;
    float opacity_raw;  // inverse_sigmoid(1/V)
};

// Pass 2: selected pixels, in pixel order, become Gaussians.
__global__ void __launch_bounds__(PM_THREADS) pointmap_emit_kernel(const PointmapEmitArgs a) {
    __shared__ uint32_t warp_off[PM_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * PM_BLOCK_PIX + (int64_t)threadIdx.x * PM_PER_THREAD;
    uint32_t m[PM_PER_THREAD];
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < PM_PER_THREAD; k++) {
        const int64_t i = base + k;
        m[k] = (i < a.n) ? a.mask[i] : 0u;
        c += m[k];
    }
    // exclusive position of this thread's first selected pixel inside the block (threads own consecutive pixels)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_off[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < PM_THREADS / 32; w++) wbase += (w < warp) ? warp_off[w] : 0u;
    size_t pos = (size_t)a.block_offsets[blockIdx.x] + wbase + incl - c;
    const int64_t HW = (int64_t)a.H * a.W;
#pragma unroll
    for (int k = 0; k < PM_PER_THREAD; k++) {
        if (!m[k]) continue;
        const int64_t i = base + k;
        const int v = (int)(i / HW);
        const int64_t pix = i - (int64_t)v * HW;
        const int py = (int)(pix / a.W), px = (int)(pix - (int64_t)py * a.W);
        // _fast_depthmap_to_pts3d (pose_utils.py:672-683): depth * (grid - pp) / focal, depth
        const float* Kv = a.K + 9 * v;
        const float focal = Kv[0], ppx = Kv[2], ppy = Kv[5];
        const float d = a.depth[i];
        const float cx = __fdiv_rn(__fmul_rn(d, (float)px - ppx), focal);
        const float cy = __fdiv_rn(__fmul_rn(d, (float)py - ppy), focal);
        // geotrf (pose_utils.py:634-637): R @ p + t with the upper 3x4 of the camera-to-world pose
        const float* T = a.poses + 16 * v;
        float* o = a.xyz + 3 * pos;
        o[0] = T[0] * cx + T[1] * cy + T[2] * d + T[3];
        o[1] = T[4] * cx + T[5] * cy + T[6] * d + T[7];
        o[2] = T[8] * cx + T[9] * cy + T[10] * d + T[11];
        // RGB2SH (sh_utils.py:114-115) of the pixel's colour; rgb is [V,3,H,W]
        const float* cv = a.rgb + (size_t)v * 3 * HW + pix;
        float* f = a.features_dc + 3 * pos;
        f[0] = __fdiv_rn(cv[0] - 0.5f, kSH0);
        f[1] = __fdiv_rn(cv[HW] - 0.5f, kSH0);
        f[2] = __fdiv_rn(cv[2 * HW] - 0.5f, kSH0);
        reinterpret_cast<float4*>(a.rotation)[pos] = make_float4(1.f, 0.f, 0.f, 0.f);
        a.opacity[pos] = a.opacity_raw;
        if (a.conf_index) a.conf_index[pos] = i;
        pos++;
    }
}

// scales = log(sqrt(clamp_min(dist2, 1e-7))) repeated x3 (gaussian_model.py:641-642)
__global__ void pointmap_scales_kernel(int64_t P, const float* __restrict__ dist2, float* __restrict__ scaling) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float s = logf(sqrtf(fmaxf(dist2[i], 0.0000001f)));
    scaling[3 * i] = s;
    scaling[3 * i + 1] = s;
    scaling[3 * i + 2] = s;
}

size_t pointmap_scratch_bytes(int64_t n_pixels) {
    const size_t blocks = (size_t)((n_pixels + PM_BLOCK_PIX - 1) / PM_BLOCK_PIX);
    return (blocks * 4 + 255) / 256 * 256 + 256;
}

cudaError_t launch_pointmap_select(const float* conf, const float* dyna, float log_thre, int64_t n, uint8_t* mask,
                                   float* conf_static, void* scratch, cudaStream_t s) {
    const int blocks = (int)((n + PM_BLOCK_PIX - 1) / PM_BLOCK_PIX);
    uint32_t* counts = reinterpret_cast<uint32_t*>(static_cast<char*>(scratch) + 256);
    unsigned long long* total = reinterpret_cast<unsigned long long*>(scratch);
    if (blocks > 0) pointmap_select_kernel<<<blocks, PM_THREADS, 0, s>>>(conf, dyna, log_thre, n, mask, conf_static, counts);
    pointmap_scan_kernel<<<1, 1024, 0, s>>>(counts, blocks, total);
    return cudaGetLastError();
}

cudaError_t launch_pointmap_emit(int V, int H, int W, const float* K, const float* poses, const float* depth,
                                 const float* rgb, const uint8_t* mask, void* scratch, float* xyz, float* features_dc,
                                 float* features_rest, float* rotation, float* opacity, int64_t* conf_index,
                                 float opacity_raw, cudaStream_t s) {
    PointmapEmitArgs a;
    a.V = V; a.H = H; a.W = W;
    a.n = (int64_t)V * H * W;
    a.K = K; a.poses = poses; a.depth = depth; a.rgb = rgb; a.mask = mask;
    a.block_offsets = reinterpret_cast<const uint32_t*>(static_cast<char*>(scratch) + 256);
    a.xyz = xyz; a.features_dc = features_dc; a.features_rest = features_rest; a.rotation = rotation;
    a.opacity = opacity; a.conf_index = conf_index; a.opacity_raw = opacity_raw;
    const int blocks = (int)((a.n + PM_BLOCK_PIX - 1) / PM_BLOCK_PIX);
    if (blocks > 0) pointmap_emit_kernel<<<blocks, PM_THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_pointmap_scales(int64_t P, const float* dist2, float* scaling, cudaStream_t s) {
    if (P > 0) pointmap_scales_kernel<<<(unsigned)((P + 255) / 256), 256, 0, s>>>(P, dist2, scaling);
    return cudaGetLastError();
}

}  // namespace das3r
