// Fused DAS3R photometric loss for sm_100a (include/das3r_loss.h): static-mask multiply, L1, SSIM (11x11 Gaussian window
// as two 1-D passes through shared memory), PSNR and the mean in one kernel + a tiny deterministic reduction; the
// backward convolves the five SSIM derivative maps the forward left behind and adds the L1 part.
// Reference: train_gui.py:565-577, utils/loss_utils.py:17-67, utils/image_utils.py:17-19.
//
// SSIM per pixel, with mu = w*x, e11 = w*(x1 x1), e22 = w*(x2 x2), e12 = w*(x1 x2) (w* = windowed sum):
//   A1 = 2 mu1 mu2 + C1, A2 = 2 (e12 - mu1 mu2) + C2, B1 = mu1^2 + mu2^2 + C1, B2 = (e11 - mu1^2) + (e22 - mu2^2) + C2,
//   S = A1 A2 / (B1 B2)                                                        (utils/loss_utils.py:49-62)
// and, because S is a function of the five windowed sums and the window is symmetric with zero padding,
//   dL/dx1 = w*(g dS/dmu1) + 2 x1 w*(g dS/de11) + x2 w*(g dS/de12),   g = dL/dS,   likewise for x2.
#include "common.cuh"
#include "kernels.cuh"
#include <math.h>

namespace das3r {

constexpr int LT = 16;            // output tile edge
constexpr int LR = 5;             // window radius (window_size 11)
constexpr int LTI = LT + 2 * LR;  // input tile edge: 26
constexpr int LOSS_PART_F = 4;    // per block: loss sum, squared-error sum of the three channels

__constant__ float c_win[2 * LR + 1];

// gaussian(11, 1.5) of utils/loss_utils.py:26-28 in float32: exp evaluated in double (Python floats), the list turned
// into a float32 tensor, divided by its float32 sum.
static void window_weights(float* w) {
    float s = 0.f;
    for (int x = 0; x < 2 * LR + 1; x++) {
        w[x] = (float)exp(-(double)((x - LR) * (x - LR)) / (2.0 * 1.5 * 1.5));
        s += w[x];
    }
    for (int x = 0; x < 2 * LR + 1; x++) w[x] = w[x] / s;
}

cudaError_t loss_upload_window() {
    float w[2 * LR + 1];
    window_weights(w);
    return cudaMemcpyToSymbol(c_win, w, sizeof(w));
}

// Separable window over five LTI x LTI input planes held in shared memory: horizontal pass into `h` (5 x LTI x LT),
// vertical pass returns the five windowed sums at this thread's pixel. All 256 threads must call it.
__device__ __forceinline__ void window5(const float (*in)[LTI][LTI], float (*h)[LTI][LT], float out[5]) {
    const int tid = threadIdx.y * LT + threadIdx.x;
    for (int e = tid; e < LTI * LT; e += LT * LT) {
        const int r = e / LT, c = e - r * LT;
        float s[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 2 * LR + 1; k++) {
            const float w = c_win[k];
#pragma unroll
            for (int q = 0; q < 5; q++) s[q] = fmaf(w, in[q][r][c + k], s[q]);
        }
#pragma unroll
        for (int q = 0; q < 5; q++) h[q][r][c] = s[q];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 5; q++) out[q] = 0.f;
#pragma unroll
    for (int k = 0; k < 2 * LR + 1; k++) {
        const float w = c_win[k];
#pragma unroll
        for (int q = 0; q < 5; q++) out[q] = fmaf(w, h[q][threadIdx.y + k][threadIdx.x], out[q]);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(LT * LT) loss_fwd_kernel(const LossArgs a, float* __restrict__ dmaps,
                                                           float* __restrict__ partials) {
    __shared__ float s_in[5][LTI][LTI];
    __shared__ float s_h[5][LTI][LT];
    __shared__ float s_red[LT * LT / 32][LOSS_PART_F];
    const int W = a.W, H = a.H;
    const size_t N = (size_t)W * H;
    const int x0 = blockIdx.x * LT - LR, y0 = blockIdx.y * LT - LR;
    const int px = blockIdx.x * LT + threadIdx.x, py = blockIdx.y * LT + threadIdx.y;
    const bool inside = px < W && py < H;
    const int tid = threadIdx.y * LT + threadIdx.x;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    float acc[LOSS_PART_F] = {0.f, 0.f, 0.f, 0.f};

    for (int ch = 0; ch < 3; ch++) {
        const float* img = a.image + (size_t)ch * N;
        const float* gt = a.gt + (size_t)ch * N;
        // masked inputs and their products, zero outside the image (conv2d padding, utils/loss_utils.py:46-54)
        for (int e = tid; e < LTI * LTI; e += LT * LT) {
            const int r = e / LTI, c = e - r * LTI;
            const int x = x0 + c, y = y0 + r;
            float v1 = 0.f, v2 = 0.f;
            if (x >= 0 && x < W && y >= 0 && y < H) {
                const size_t p = (size_t)y * W + x;
                const float s = a.static_mask ? a.static_mask[p] : 1.0f;
                v1 = img[p] * s;
                v2 = gt[p] * s;
            }
            s_in[0][r][c] = v1;
            s_in[1][r][c] = v2;
            s_in[2][r][c] = v1 * v1;
            s_in[3][r][c] = v2 * v2;
            s_in[4][r][c] = v1 * v2;
        }
        __syncthreads();
        const float x1 = s_in[0][threadIdx.y + LR][threadIdx.x + LR], x2 = s_in[1][threadIdx.y + LR][threadIdx.x + LR];
        float m[5];
        window5(s_in, s_h, m);
        if (inside) {
            const float mu1 = m[0], mu2 = m[1];
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
            const float sigma1_sq = m[2] - mu1_sq, sigma2_sq = m[3] - mu2_sq, sigma12 = m[4] - mu1_mu2;
            const float A1 = 2 * mu1_mu2 + C1, A2 = 2 * sigma12 + C2;
            const float B1 = mu1_sq + mu2_sq + C1, B2 = sigma1_sq + sigma2_sq + C2;
            const float S = (A1 * A2) / (B1 * B2);
            const float diff = x1 - x2;
            acc[0] += (1.0f - a.lambda_dssim) * fabsf(diff) + a.lambda_dssim * (1.0f - S);
            acc[1 + ch] = diff * diff;
            // derivative maps for the backward (header comment): dS/dmu1, dS/dmu2, dS/de11, dS/de22, dS/de12
            const float inv = 1.0f / (B1 * B2);
            const float common = 2.0f * (A2 - A1) * inv;
            const size_t p = (size_t)py * W + px;
            float* d = dmaps + (size_t)ch * 5 * N + p;
            d[0 * N] = mu2 * common - 2.0f * mu1 * S / B1 + 2.0f * mu1 * S / B2;
            d[1 * N] = mu1 * common - 2.0f * mu2 * S / B1 + 2.0f * mu2 * S / B2;
            d[2 * N] = -S / B2;
            d[3 * N] = -S / B2;
            d[4 * N] = 2.0f * A1 * inv;
        }
    }
    // block sums -> partials[block][4]
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int k = 0; k < LOSS_PART_F; k++) {
        float v = acc[k];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][k] = v;
    }
    __syncthreads();
    if (tid < LOSS_PART_F) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < LT * LT / 32; w++) v += s_red[w][tid];
        partials[(size_t)(blockIdx.y * gridDim.x + blockIdx.x) * LOSS_PART_F + tid] = v;
    }
}

// One block: adds the block partials in a fixed order (double accumulators), writes loss and psnr.
__global__ void __launch_bounds__(256) loss_finish_kernel(const float* __restrict__ partials, int n_blocks, int W, int H,
                                                          float* __restrict__ out) {
    __shared__ double s_red[8][LOSS_PART_F];
    __shared__ double s_sum[LOSS_PART_F];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = 0; k < LOSS_PART_F; k++) {
        double v = 0.0;
        for (int b = tid; b < n_blocks; b += 256) v += (double)partials[(size_t)b * LOSS_PART_F + k];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][k] = v;
    }
    __syncthreads();
    if (tid < LOSS_PART_F) {
        double v = 0.0;
        for (int w = 0; w < 8; w++) v += s_red[w][tid];
        s_sum[tid] = v;
    }
    __syncthreads();
    if (tid == 0) {
        const double N = (double)W * H;
        out[0] = (float)(s_sum[0] / (3.0 * N));
        float ps = 0.f;
        for (int c = 0; c < 3; c++) {
            const float mse = (float)(s_sum[1 + c] / N);
            ps += 20.0f * log10f(1.0f / sqrtf(mse));  // utils/image_utils.py:17-19
        }
        out[1] = ps / 3.0f;
    }
}

__global__ void __launch_bounds__(LT * LT) loss_bwd_kernel(const LossArgs a, const float* __restrict__ dmaps,
                                                           const float* __restrict__ grad_loss,
                                                           float* __restrict__ dL_dimage, float* __restrict__ dL_dstatic,
                                                           float* __restrict__ dL_dgt) {
    __shared__ float s_in[5][LTI][LTI];
    __shared__ float s_h[5][LTI][LT];
    const int W = a.W, H = a.H;
    const size_t N = (size_t)W * H;
    const int x0 = blockIdx.x * LT - LR, y0 = blockIdx.y * LT - LR;
    const int px = blockIdx.x * LT + threadIdx.x, py = blockIdx.y * LT + threadIdx.y;
    const bool inside = px < W && py < H;
    const int tid = threadIdx.y * LT + threadIdx.x;
    const float up = grad_loss ? *grad_loss : 1.0f;
    const float gS = -a.lambda_dssim * up / (3.0f * (float)N);          // d loss / d S
    const float gL = (1.0f - a.lambda_dssim) * up / (3.0f * (float)N);  // d loss / d |x1 - x2|
    const size_t p = (size_t)py * W + px;
    const float s = (inside && a.static_mask) ? a.static_mask[p] : 1.0f;
    float dstatic = 0.f;

    for (int ch = 0; ch < 3; ch++) {
        const float* dm = dmaps + (size_t)ch * 5 * N;
        for (int e = tid; e < LTI * LTI; e += LT * LT) {
            const int r = e / LTI, c = e - r * LTI;
            const int x = x0 + c, y = y0 + r;
            const bool in = x >= 0 && x < W && y >= 0 && y < H;
            const size_t q = (size_t)y * W + x;
#pragma unroll
            for (int k = 0; k < 5; k++) s_in[k][r][c] = in ? dm[(size_t)k * N + q] : 0.f;
        }
        __syncthreads();
        float m[5];
        window5(s_in, s_h, m);
        if (inside) {
            const float i1 = a.image[(size_t)ch * N + p], i2 = a.gt[(size_t)ch * N + p];
            const float x1 = i1 * s, x2 = i2 * s;
            const float diff = x1 - x2;
            const float sg = diff > 0.f ? 1.0f : (diff < 0.f ? -1.0f : 0.f);  // torch.abs backward: 0 at 0
            const float gx1 = gS * (m[0] + 2.0f * x1 * m[2] + x2 * m[4]) + gL * sg;
            const float gx2 = gS * (m[1] + 2.0f * x2 * m[3] + x1 * m[4]) - gL * sg;
            dL_dimage[(size_t)ch * N + p] = gx1 * s;
            if (dL_dgt) dL_dgt[(size_t)ch * N + p] = gx2 * s;
            dstatic += gx1 * i1 + gx2 * i2;
        }
    }
    if (inside && dL_dstatic) dL_dstatic[p] = dstatic;
}

size_t loss_state_bytes(int W, int H) {
    const size_t N = (size_t)W * H;
    const size_t blocks = (size_t)((W + LT - 1) / LT) * ((H + LT - 1) / LT);
    return (15 * N * sizeof(float) + 255) / 256 * 256 + (blocks * LOSS_PART_F * sizeof(float) + 255) / 256 * 256;
}

static float* loss_partials(void* state, int W, int H) {
    const size_t N = (size_t)W * H;
    return reinterpret_cast<float*>(static_cast<char*>(state) + (15 * N * sizeof(float) + 255) / 256 * 256);
}

cudaError_t launch_loss_fwd(const LossArgs& a, float* out, void* state, cudaStream_t s) {
    const dim3 grid((a.W + LT - 1) / LT, (a.H + LT - 1) / LT), block(LT, LT);
    float* partials = loss_partials(state, a.W, a.H);
    loss_fwd_kernel<<<grid, block, 0, s>>>(a, static_cast<float*>(state), partials);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    loss_finish_kernel<<<1, 256, 0, s>>>(partials, (int)(grid.x * grid.y), a.W, a.H, out);
    return cudaGetLastError();
}

cudaError_t launch_loss_bwd(const LossArgs& a, const void* state, const float* grad_loss, float* dL_dimage,
                            float* dL_dstatic, float* dL_dgt, cudaStream_t s) {
    const dim3 grid((a.W + LT - 1) / LT, (a.H + LT - 1) / LT), block(LT, LT);
    loss_bwd_kernel<<<grid, block, 0, s>>>(a, static_cast<const float*>(state), grad_loss, dL_dimage, dL_dstatic, dL_dgt);
    return cudaGetLastError();
}

}  // namespace das3r
