// Backward path of the sm_100a Gaussian-splat rasterizer:
//   blend_bwd       -> back-to-front replay per 16x16 tile (replaces renderCUDA, reference
//                      cuda_rasterizer/backward.cu:399-557); only pairs that really contribute reach the gradient
//                      math, and their 9 partials leave as two 128-bit + one 32-bit atomics (backward.cu:523-554 uses
//                      nine scalar ones per pair).
//   preprocess_bwd  -> computeCov2DCUDA (backward.cu:144-274) + preprocessCUDA (:346-396) + SH backward (:20-139) +
//                      cov3D backward (:278-341) fused into one pass that also writes the zeros the reference gets from
//                      nine torch::zeros fills (rasterize_points.cu:151-159).
#include "common.cuh"
#include "kernels.cuh"
#include "blend_common.cuh"
#include <stdlib.h>
#include <algorithm>

namespace das3r {

// blend_bwd (replaces renderCUDA, backward.cu:399-557): walks the rectangle's list back to front (see
// blend_common.cuh) from the warp's highest `n_contrib`. In the blend step every lane holds a DIFFERENT
// (pixel, Gaussian) pair that really contributes, evaluates backward.cu:490-554 for it and adds the 9 partial gradients
// to the Gaussian's 48-byte accumulator with two 128-bit and one 32-bit no-return atomics (REDG.E.ADD.F32x4) instead of
// the reference's nine scalar ones.
struct BwdPair {
    float4 col;  // {r, g, b, gaussian index}: fetched in prep so that apply starts without a shared-memory wait
    float2 d;
    float G, alpha, inv_1ma;
    bool skip;
};

__global__ void __launch_bounds__(32 * BLEND_WARPS, DAS3R_BWD_MIN_BLOCKS) blend_bwd_kernel(
    const __grid_constant__ BatchArgs b) {
    __shared__ __align__(16) BlendSmem<BWD_K> sm;
    const ViewBin& V = b.v[blockIdx.y];
    const uint2* __restrict__ ranges = V.ranges;
    const uint32_t* __restrict__ rect_off = V.rect_off;
    const uint32_t* __restrict__ rect_cnt = V.rect_cnt;
    const uint2* __restrict__ rect_list = V.rect_list;
    const uint32_t* __restrict__ rect_gid = V.rect_gid;
    const GaussRec* __restrict__ rec_by_id = V.rec;
    const float* __restrict__ bg = V.bg;
    const float* __restrict__ final_Ts = V.final_T;
    const uint32_t* __restrict__ n_contrib = V.n_contrib;
    const float* __restrict__ dL_dpixels = V.dL_dpix;
    float* __restrict__ acc_out = V.acc;
    const int W = b.W, H = b.H;
    const unsigned gx = b.gx;
    const int warp = threadIdx.x >> 5;
    const WarpRect wr = warp_rect(W, H, gx);
    const uint2 range = ranges[wr.tile];
    if (range.y == range.x) return;
    const int cnt = (int)rect_cnt[wr.tile * 8 + wr.rect];
    if (cnt == 0) return;
    const uint32_t off = rect_off[wr.tile * 8 + wr.rect];
    const int pix_id = W * wr.py + wr.px;
    const bool inside = wr.inside;
    const float2 pixf = {(float)wr.px, (float)wr.py};

    // backward.cu:431-458
    const float T_final = inside ? final_Ts[pix_id] : 0;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0;
    float accum_rec[3] = {0.f, 0.f, 0.f};
    float dL_dpixel[3] = {0.f, 0.f, 0.f};
    if (inside)
#pragma unroll
        for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * H * W + pix_id];
    float last_alpha = 0;
    float last_color[3] = {0.f, 0.f, 0.f};
    const float ddelx_dx = 0.5 * W;
    const float ddely_dy = 0.5 * H;
    float bg_dot_dpixel = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) bg_dot_dpixel += bg[i] * dL_dpixel[i];

    // Highest list position any pixel of this warp still needs: everything behind it is skipped.
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last_contributor);
    if (wmax == 0) return;

    walk_rect_list<true, BWD_K, BWD_NPAIR>(
        sm.ring[warp], sm.qmask[warp], rect_list + off, rect_gid + off, cnt, rec_by_id, wmax,
        [&]() { return last_contributor != 0u; },
        [&](const RingRec& r) {
            // backward.cu:490-500: everything that does not depend on the pixel's running state
            BwdPair x;
            const float4 A = r.A;
            const float4 con_o = r.B;
            x.d = make_float2(A.x - pixf.x, A.y - pixf.y);
            const float power = splat_power(con_o, x.d);
            x.G = expf(power);
            x.alpha = min(0.99f, __fmul_rn(con_o.w, x.G));
            // one IEEE reciprocal of (1 - alpha) serves both divisions of backward.cu:503,:534 (gradients are
            // compared at 1e-3; the forward's T and alpha decisions are not touched by this)
            x.inv_1ma = __frcp_rn(1.f - x.alpha);
            x.skip = __float_as_uint(A.z) >= last_contributor || x.alpha < 1.0f / 255.0f;
            x.col = r.C;
            return x;
        },
        [&](const RingRec& r, const BwdPair& x) {
            // backward.cu:501-554
            if (x.skip) return;  // behind this pixel's last contributor, or alpha < 1/255
            const float4 con_o = r.B;
            const float2 d = x.d;
            const float G = x.G, alpha = x.alpha, inv_1ma = x.inv_1ma;
            const float4 col4 = x.col;
            const float col[3] = {col4.x, col4.y, col4.z};
            T = T * inv_1ma;
            const float dchannel_dcolor = alpha * T;
            float dL_dalpha = 0.0f;
            float gcol[3];
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                const float c = col[ch];
                accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                last_color[ch] = c;
                const float dL_dchannel = dL_dpixel[ch];
                dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
                gcol[ch] = dchannel_dcolor * dL_dchannel;
            }
            dL_dalpha *= T;
            last_alpha = alpha;
            dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;
            const float dL_dG = con_o.w * dL_dalpha;
            const float gdx = G * d.x;
            const float gdy = G * d.y;
            const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
            const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
            float* dst = acc_out + (size_t)__float_as_uint(col4.w) * ACC_STRIDE;
            atomicAdd(reinterpret_cast<float4*>(dst),
                      make_float4(dL_dG * dG_ddelx * ddelx_dx, dL_dG * dG_ddely * ddely_dy, -0.5f * gdx * d.x * dL_dG,
                                  -0.5f * gdx * d.y * dL_dG));
            atomicAdd(reinterpret_cast<float4*>(dst) + 1,
                      make_float4(-0.5f * gdy * d.y * dL_dG, G * dL_dalpha, gcol[0], gcol[1]));
            atomicAdd(dst + 8, gcol[2]);
        });
}

cudaError_t launch_blend_bwd(const BatchArgs& b, cudaStream_t s) {
    const unsigned per_view = b.gx * b.gy * (8 / BLEND_WARPS);
    const size_t pad = spread_pad_smem((const void*)blend_bwd_kernel, (int)(per_view * b.n), 32 * BLEND_WARPS,
                                       sizeof(BlendSmem<BWD_K>));
    blend_bwd_kernel<<<dim3(per_view, b.n), 32 * BLEND_WARPS, pad, s>>>(b);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// preprocess_bwd
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float3 dnormvdv3(float3 v, float3 dv) {  // auxiliary.h:107-117
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward (backward.cu:20-139), restated per coefficient so that nothing but three running sums has to stay live:
//   dL_dsh[k] = B_k(dir) * dL_dRGB            (B_k = the k-th real SH basis function, as in the forward)
//   dL_ddir   = sum_k grad(B_k)(dir) * (sh[k] . dL_dRGB)
// which is the reference's dRGBdx/dRGBdy/dRGBdz contraction with the sums over channels and coefficients swapped.
// `get(k)` returns sh[k] as float3, `put(k, v)` stores dL_dsh[k]. Returns the view-direction part of dL_dmean.
template <typename Get, typename Put>
__device__ __forceinline__ float3 sh_backward(int deg, float3 pos, float3 campos, uint8_t clamped, float3 dL_dcolor,
                                              Get get, Put put) {
    float3 dir_orig = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    // PyTorch clamping rule: a clamped channel passes no gradient (backward.cu:31-34)
    const float3 g = {(clamped & 1) ? 0.f : dL_dcolor.x, (clamped & 2) ? 0.f : dL_dcolor.y,
                      (clamped & 4) ? 0.f : dL_dcolor.z};
    float3 dd = {0.f, 0.f, 0.f};
    auto term = [&](int k, float B, float bx, float by, float bz) {
        const float3 c = get(k);
        const float sdot = c.x * g.x + c.y * g.y + c.z * g.z;
        dd.x += bx * sdot;
        dd.y += by * sdot;
        dd.z += bz * sdot;
        put(k, make_float3(B * g.x, B * g.y, B * g.z));
    };
    put(0, make_float3(kSH0 * g.x, kSH0 * g.y, kSH0 * g.z));
    if (deg > 0) {
        term(1, -kSH1 * y, 0.f, -kSH1, 0.f);
        term(2, kSH1 * z, 0.f, 0.f, kSH1);
        term(3, -kSH1 * x, -kSH1, 0.f, 0.f);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            term(4, kSH2[0] * xy, kSH2[0] * y, kSH2[0] * x, 0.f);
            term(5, kSH2[1] * yz, 0.f, kSH2[1] * z, kSH2[1] * y);
            term(6, kSH2[2] * (2.f * zz - xx - yy), kSH2[2] * -2.f * x, kSH2[2] * -2.f * y, kSH2[2] * 4.f * z);
            term(7, kSH2[3] * xz, kSH2[3] * z, 0.f, kSH2[3] * x);
            term(8, kSH2[4] * (xx - yy), kSH2[4] * 2.f * x, kSH2[4] * -2.f * y, 0.f);
            if (deg > 2) {
                term(9, kSH3[0] * y * (3.f * xx - yy), kSH3[0] * 6.f * xy, kSH3[0] * 3.f * (xx - yy), 0.f);
                term(10, kSH3[1] * xy * z, kSH3[1] * yz, kSH3[1] * xz, kSH3[1] * xy);
                term(11, kSH3[2] * y * (4.f * zz - xx - yy), kSH3[2] * -2.f * xy,
                     kSH3[2] * (4.f * zz - xx - 3.f * yy), kSH3[2] * 8.f * yz);
                term(12, kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy), kSH3[3] * -6.f * xz, kSH3[3] * -6.f * yz,
                     kSH3[3] * 3.f * (2.f * zz - xx - yy));
                term(13, kSH3[4] * x * (4.f * zz - xx - yy), kSH3[4] * (4.f * zz - 3.f * xx - yy),
                     kSH3[4] * -2.f * xy, kSH3[4] * 8.f * xz);
                term(14, kSH3[5] * z * (xx - yy), kSH3[5] * 2.f * xz, kSH3[5] * -2.f * yz, kSH3[5] * (xx - yy));
                term(15, kSH3[6] * x * (xx - 3.f * yy), kSH3[6] * 3.f * (xx - yy), kSH3[6] * -6.f * xy, 0.f);
            }
        }
    }
    return dnormvdv3(dir_orig, dd);
}

// Gradients of one (view, Gaussian) pair with respect to the Gaussian's 3D mean, scale and rotation, given the
// screen-space gradients the backward blend accumulated: computeCov2DCUDA (backward.cu:144-274), the projection part of
// preprocessCUDA (:366-395) and computeCov3D (:278-341). Shared by the one-view and the multi-view kernel.
struct ViewConst {
    const float *view, *proj;
    float focal_x, focal_y, tan_fovx, tan_fovy, scale_mod;
};
struct GeomGrad {
    float3 dmean;   // without the SH (view direction) part
    float3 dscale;  // valid if has_sr
    float4 drot;    // valid if has_sr
    float dcov[6];  // dL/dcov3D
};
__device__ __forceinline__ void geom_backward(const ViewConst& vc, const float3 mean, const float* cov3D, const bool has_sr,
                                              const float3 sc, const float4 rq, const float dL_dmean2D_x,
                                              const float dL_dmean2D_y, const float3 dL_dconic, GeomGrad& g) {
    // ---- computeCov2DCUDA, backward.cu:162-273 ----
    const float h_x = vc.focal_x, h_y = vc.focal_y;
    Ewa e = ewa_project(mean, h_x, h_y, vc.tan_fovx, vc.tan_fovy, cov3D, vc.view);
    const float3 t = e.t;
    const float x_grad_mul = e.txtz < -e.limx || e.txtz > e.limx ? 0 : 1;
    const float y_grad_mul = e.tytz < -e.limy || e.tytz > e.limy ? 0 : 1;
    const Mat3& T = e.T;
    const Mat3& Vrk = e.Vrk;
    const float ca = e.a, cb = e.b, cc = e.c;
    float denom = ca * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dL_dcov[6];
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dL_dconic.x + 2 * cb * cc * dL_dconic.y + (denom - ca * cc) * dL_dconic.z);
        dL_dc = denom2inv * (-ca * ca * dL_dconic.z + 2 * ca * cb * dL_dconic.y + (denom - ca * cc) * dL_dconic.x);
        dL_db = denom2inv * 2 * (cb * cc * dL_dconic.x - (denom + 2 * cb * cb) * dL_dconic.y + ca * cb * dL_dconic.z);
        dL_dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
        dL_dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
        dL_dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
        dL_dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
                     2 * T.m[1][0] * T.m[1][1] * dL_dc;
        dL_dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
                     2 * T.m[1][0] * T.m[1][2] * dL_dc;
        dL_dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
                     2 * T.m[1][1] * T.m[1][2] * dL_dc;
    } else {
#pragma unroll
        for (int i = 0; i < 6; i++) dL_dcov[i] = 0;
    }
#pragma unroll
    for (int i = 0; i < 6; i++) g.dcov[i] = dL_dcov[i];

    float dL_dT00 = 2 * (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_da +
                    (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_db;
    float dL_dT01 = 2 * (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_da +
                    (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_db;
    float dL_dT02 = 2 * (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_da +
                    (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_db;
    float dL_dT10 = 2 * (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_dc +
                    (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_db;
    float dL_dT11 = 2 * (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_dc +
                    (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_db;
    float dL_dT12 = 2 * (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_dc +
                    (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_db;

    // W as built in ewa_project: W.m[c][r] = view[4*r + c]  (glm::mat3(view[0], view[4], view[8], ...))
    const float* vm = vc.view;
    const float W00 = vm[0], W01 = vm[4], W02 = vm[8];
    const float W10 = vm[1], W11 = vm[5], W12 = vm[9];
    const float W20 = vm[2], W21 = vm[6], W22 = vm[10];
    float dL_dJ00 = W00 * dL_dT00 + W01 * dL_dT01 + W02 * dL_dT02;
    float dL_dJ02 = W20 * dL_dT00 + W21 * dL_dT01 + W22 * dL_dT02;
    float dL_dJ11 = W10 * dL_dT10 + W11 * dL_dT11 + W12 * dL_dT12;
    float dL_dJ12 = W20 * dL_dT10 + W21 * dL_dT11 + W22 * dL_dT12;

    float tz = 1.f / t.z;
    float tz2 = tz * tz;
    float tz3 = tz2 * tz;
    float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                   (2 * h_y * t.y) * tz3 * dL_dJ12;
    float3 dL_dmean_acc = xform_vec_4x3_T({dL_dtx, dL_dty, dL_dtz}, vc.view);

    // ---- preprocessCUDA (backward.cu:366-395) ----
    const float* proj = vc.proj;
    float4 m_hom = xform_point_4x4(mean, proj);
    float m_w = 1.0f / (m_hom.w + 0.0000001f);
    float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    float3 dL_dmean;
    dL_dmean.x = (proj[0] * m_w - proj[3] * mul1) * dL_dmean2D_x + (proj[1] * m_w - proj[3] * mul2) * dL_dmean2D_y;
    dL_dmean.y = (proj[4] * m_w - proj[7] * mul1) * dL_dmean2D_x + (proj[5] * m_w - proj[7] * mul2) * dL_dmean2D_y;
    dL_dmean.z = (proj[8] * m_w - proj[11] * mul1) * dL_dmean2D_x + (proj[9] * m_w - proj[11] * mul2) * dL_dmean2D_y;
    dL_dmean_acc.x += dL_dmean.x;
    dL_dmean_acc.y += dL_dmean.y;
    dL_dmean_acc.z += dL_dmean.z;
    g.dmean = dL_dmean_acc;

    // ---- computeCov3D backward (backward.cu:278-341) ----
    if (has_sr) {
        float r = rq.x, x = rq.y, y = rq.z, z = rq.w;
        Mat3 R = mat3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                      2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                      2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
        Mat3 S = mat3(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
        float3 s = {vc.scale_mod * sc.x, vc.scale_mod * sc.y, vc.scale_mod * sc.z};
        S.m[0][0] = s.x;
        S.m[1][1] = s.y;
        S.m[2][2] = s.z;
        Mat3 M = mul(S, R);
        Mat3 dL_dSigma = mat3(dL_dcov[0], 0.5f * dL_dcov[1], 0.5f * dL_dcov[2], 0.5f * dL_dcov[1], dL_dcov[3],
                              0.5f * dL_dcov[4], 0.5f * dL_dcov[2], 0.5f * dL_dcov[4], dL_dcov[5]);
        // dL_dM = 2.0f * M * dL_dSigma  (glm: scalar*mat first, then the product)
        Mat3 M2;
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) M2.m[c][rr] = 2.0f * M.m[c][rr];
        Mat3 dL_dM = mul(M2, dL_dSigma);
        Mat3 Rt = transpose(R);
        Mat3 dL_dMt = transpose(dL_dM);
        float3 dsc;
        dsc.x = Rt.m[0][0] * dL_dMt.m[0][0] + Rt.m[0][1] * dL_dMt.m[0][1] + Rt.m[0][2] * dL_dMt.m[0][2];
        dsc.y = Rt.m[1][0] * dL_dMt.m[1][0] + Rt.m[1][1] * dL_dMt.m[1][1] + Rt.m[1][2] * dL_dMt.m[1][2];
        dsc.z = Rt.m[2][0] * dL_dMt.m[2][0] + Rt.m[2][1] * dL_dMt.m[2][1] + Rt.m[2][2] * dL_dMt.m[2][2];
        g.dscale = dsc;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            dL_dMt.m[0][k] *= s.x;
            dL_dMt.m[1][k] *= s.y;
            dL_dMt.m[2][k] *= s.z;
        }
        float4 dq;
        dq.x = 2 * z * (dL_dMt.m[0][1] - dL_dMt.m[1][0]) + 2 * y * (dL_dMt.m[2][0] - dL_dMt.m[0][2]) +
               2 * x * (dL_dMt.m[1][2] - dL_dMt.m[2][1]);
        dq.y = 2 * y * (dL_dMt.m[1][0] + dL_dMt.m[0][1]) + 2 * z * (dL_dMt.m[2][0] + dL_dMt.m[0][2]) +
               2 * r * (dL_dMt.m[1][2] - dL_dMt.m[2][1]) - 4 * x * (dL_dMt.m[2][2] + dL_dMt.m[1][1]);
        dq.z = 2 * x * (dL_dMt.m[1][0] + dL_dMt.m[0][1]) + 2 * r * (dL_dMt.m[2][0] - dL_dMt.m[0][2]) +
               2 * z * (dL_dMt.m[1][2] + dL_dMt.m[2][1]) - 4 * y * (dL_dMt.m[2][2] + dL_dMt.m[0][0]);
        dq.w = 2 * r * (dL_dMt.m[0][1] - dL_dMt.m[1][0]) + 2 * x * (dL_dMt.m[2][0] + dL_dMt.m[0][2]) +
               2 * y * (dL_dMt.m[1][2] + dL_dMt.m[2][1]) - 4 * z * (dL_dMt.m[1][1] + dL_dMt.m[0][0]);
        g.drot = dq;
    }
}

constexpr int PB_THREADS = 256;
constexpr int PB_ROW_F4 = 12;      // a degree-3 SH row: 48 floats
constexpr int PB_ROW_PAD_F4 = 13;  // padded pitch in shared memory (odd number of float4: conflict-free LDS/STS.128)

// STAGE (degree 3, M == 16): the block's 256 SH rows are brought into shared memory with coalesced 16-byte cp.async
// while the covariance / projection gradients are computed; each thread then turns ITS row into the dL_dsh row in
// registers, puts it back, and the block writes (or, in accumulate mode, adds) the 48 KB with coalesced 128-bit accesses.
template <bool STAGE>
__global__ void __launch_bounds__(PB_THREADS, 3) preprocess_bwd_kernel(const PreBwdArgs a) {
    extern __shared__ __align__(16) unsigned char pb_smem[];
    float4* stage = reinterpret_cast<float4*>(pb_smem);
    const int tid = threadIdx.x;
    const int base = blockIdx.x * PB_THREADS;
    const int idx = base + tid;
    const int nsh = 3 * a.M;  // floats per dL_dsh row
    const bool accum = a.accumulate != 0;
    const bool in_range = idx < a.P;
    const bool visible = in_range && (a.radii[idx] > 0);
    auto put1 = [accum](float* p, float v) { *p = accum ? *p + v : v; };
    auto put4 = [accum](float4* p, float4 v) {
        if (accum) {
            float4 o = *p;
            v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
        }
        *p = v;
    };

    if (STAGE) {
        const int nvalid = min(PB_THREADS, a.P - base);
        const float4* src = reinterpret_cast<const float4*>(a.shs) + (size_t)base * PB_ROW_F4;
        const int total = nvalid * PB_ROW_F4;
#pragma unroll
        for (int k = 0; k < PB_ROW_F4; k++) {
            int e = tid + k * PB_THREADS;
            if (e < total) {
                int g = e / PB_ROW_F4, j = e - g * PB_ROW_F4;
                cp_async_16(&stage[g * PB_ROW_PAD_F4 + j], src + e);
            }
        }
        cp_async_commit();
    }

    float3 mean = {0.f, 0.f, 0.f}, dL_dcol = {0.f, 0.f, 0.f}, dL_dmean_acc = {0.f, 0.f, 0.f};

    if (in_range && !visible && !accum) {
        // not rendered: every gradient is zero (the reference relies on torch::zeros for this, rasterize_points.cu:151-159)
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (a.dL_dmeans2D) a.dL_dmeans2D[3 * idx + k] = 0.f;
            if (a.dL_dcolors) a.dL_dcolors[3 * idx + k] = 0.f;
            a.dL_dmeans3D[3 * idx + k] = 0.f;
            a.dL_dscales[3 * idx + k] = 0.f;
        }
        a.dL_dopacity[idx] = 0.f;
        if (a.dL_dcov3D)
#pragma unroll
            for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * idx + k] = 0.f;
        reinterpret_cast<float4*>(a.dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.dL_dsh && !STAGE) {
            float* sh_out = a.dL_dsh + (size_t)idx * nsh;
            if (((a.M & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0)) {
                float4* o4 = reinterpret_cast<float4*>(sh_out);
                for (int j = 0; j < nsh / 4; j++) o4[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                for (int j = 0; j < nsh; j++) sh_out[j] = 0.f;
            }
        }
    }

    if (visible) {
        const float4* acc4 = reinterpret_cast<const float4*>(a.acc + (size_t)idx * ACC_STRIDE);
        const float4 q0 = acc4[0], q1 = acc4[1], q2 = acc4[2];
        const float dL_dmean2D_x = q0.x, dL_dmean2D_y = q0.y;
        const float3 dL_dconic = {q0.z, q0.w, q1.x};  // (x, y, w) of the reference's float4
        const float dL_dopac = q1.y;
        dL_dcol = {q1.z, q1.w, q2.x};

        if (a.dL_dmeans2D) {
            put1(&a.dL_dmeans2D[3 * idx + 0], dL_dmean2D_x);
            put1(&a.dL_dmeans2D[3 * idx + 1], dL_dmean2D_y);
            put1(&a.dL_dmeans2D[3 * idx + 2], 0.f);
        }
        if (a.dL_dcolors) {
            put1(&a.dL_dcolors[3 * idx + 0], dL_dcol.x);
            put1(&a.dL_dcolors[3 * idx + 1], dL_dcol.y);
            put1(&a.dL_dcolors[3 * idx + 2], dL_dcol.z);
        }
        put1(&a.dL_dopacity[idx], dL_dopac);

        mean = {a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]};
        float cov3D[6];
        float3 sc = {0.f, 0.f, 0.f};
        float4 rq = {0.f, 0.f, 0.f, 0.f};
        if (a.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) cov3D[k] = a.cov3D_precomp[6 * idx + k];
        } else {
            sc = {a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]};
            rq = __ldg(reinterpret_cast<const float4*>(a.rots) + idx);
            cov3d_from_scale_rot(sc, a.scale_mod, rq, cov3D);  // recomputed instead of stored (same code as forward)
        }

        ViewConst vc{a.view, a.proj, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, a.scale_mod};
        GeomGrad gg;
        geom_backward(vc, mean, cov3D, a.scales != nullptr, sc, rq, dL_dmean2D_x, dL_dmean2D_y, dL_dconic, gg);
        dL_dmean_acc = gg.dmean;
        if (a.dL_dcov3D)
#pragma unroll
            for (int i = 0; i < 6; i++) put1(&a.dL_dcov3D[6 * idx + i], gg.dcov[i]);
        if (a.scales != nullptr) {
            put1(&a.dL_dscales[3 * idx + 0], gg.dscale.x);
            put1(&a.dL_dscales[3 * idx + 1], gg.dscale.y);
            put1(&a.dL_dscales[3 * idx + 2], gg.dscale.z);
            put4(&reinterpret_cast<float4*>(a.dL_drots)[idx], gg.drot);
        } else if (!accum) {
            a.dL_dscales[3 * idx + 0] = 0.f;
            a.dL_dscales[3 * idx + 1] = 0.f;
            a.dL_dscales[3 * idx + 2] = 0.f;
            reinterpret_cast<float4*>(a.dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }

    // ---- SH backward (backward.cu:20-139) ----
    if (STAGE) {
        cp_async_wait_all();
        __syncthreads();  // every thread's copies have landed: the rows are complete
        float f[48];
        float4* myrow = &stage[tid * PB_ROW_PAD_F4];
        if (visible) {
#pragma unroll
            for (int j = 0; j < PB_ROW_F4; j++) {
                float4 v = myrow[j];
                f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w;
            }
            const float3 cam = {a.campos[0], a.campos[1], a.campos[2]};
            const float3 dm = sh_backward(
                3, mean, cam, a.clamped[idx], dL_dcol,
                [&](int k) { return make_float3(f[3 * k], f[3 * k + 1], f[3 * k + 2]); },
                [&](int k, float3 v) { f[3 * k] = v.x; f[3 * k + 1] = v.y; f[3 * k + 2] = v.z; });
            dL_dmean_acc.x += dm.x;
            dL_dmean_acc.y += dm.y;
            dL_dmean_acc.z += dm.z;
#pragma unroll
            for (int j = 0; j < PB_ROW_F4; j++)
                myrow[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
        } else {
#pragma unroll
            for (int j = 0; j < PB_ROW_F4; j++) myrow[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();
        const int nvalid = min(PB_THREADS, a.P - base);
        const int total = nvalid * PB_ROW_F4;
        float4* dst = reinterpret_cast<float4*>(a.dL_dsh) + (size_t)base * PB_ROW_F4;
#pragma unroll
        for (int k = 0; k < PB_ROW_F4; k++) {
            int e = tid + k * PB_THREADS;
            if (e < total) {
                int g = e / PB_ROW_F4, j = e - g * PB_ROW_F4;
                put4(&dst[e], stage[g * PB_ROW_PAD_F4 + j]);
            }
        }
    } else if (visible && a.shs != nullptr && a.dL_dsh != nullptr) {
        float* sh_out = a.dL_dsh + (size_t)idx * nsh;
        float f[48];
        const float* rowp = a.shs + (size_t)idx * nsh;
        const bool al = ((a.M & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
        load_sh_row(rowp, a.D, al, f);
        const int nw = 3 * (a.D + 1) * (a.D + 1);
        const float3 cam = {a.campos[0], a.campos[1], a.campos[2]};
        const float3 dm = sh_backward(
            a.D, mean, cam, a.clamped[idx], dL_dcol,
            [&](int k) { return make_float3(f[3 * k], f[3 * k + 1], f[3 * k + 2]); },
            [&](int k, float3 v) { f[3 * k] = v.x; f[3 * k + 1] = v.y; f[3 * k + 2] = v.z; });
        dL_dmean_acc.x += dm.x;
        dL_dmean_acc.y += dm.y;
        dL_dmean_acc.z += dm.z;
        if ((a.M == 16) && ((reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0)) {
            float4* o4 = reinterpret_cast<float4*>(sh_out);
#pragma unroll
            for (int j = 0; j < 12; j++) {
                float4 v;
                v.x = (4 * j + 0 < nw) ? f[4 * j + 0] : 0.f;
                v.y = (4 * j + 1 < nw) ? f[4 * j + 1] : 0.f;
                v.z = (4 * j + 2 < nw) ? f[4 * j + 2] : 0.f;
                v.w = (4 * j + 3 < nw) ? f[4 * j + 3] : 0.f;
                if (!accum || 4 * j < nw) put4(&o4[j], v);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 48; j++)
                if (j < nsh && (!accum || j < nw)) put1(&sh_out[j], (j < nw) ? f[j] : 0.f);
        }
    }

    if (visible) {
        put1(&a.dL_dmeans3D[3 * idx + 0], dL_dmean_acc.x);
        put1(&a.dL_dmeans3D[3 * idx + 1], dL_dmean_acc.y);
        put1(&a.dL_dmeans3D[3 * idx + 2], dL_dmean_acc.z);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// preprocess_bwd_multi: the same gradients for SEVERAL views of one step in one pass, summed over the views in
// registers / shared memory. What the views share - the SH rows - is read once, and every gradient row is written (or
// added to the caller's buffer) once instead of once per view; per view only the 48-byte accumulator row, mean,
// rotation, scale, radius and clamp bits are read. A Gaussian invisible in a view contributes nothing for it. Sums are
// formed in view order. Scale/rotation inputs, 16-coefficient SH rows, any active degree DEG: only the 3*(DEG+1)^2
// active floats of a row are read (staged through shared memory at degree 3, straight into registers below), and the
// dL_dsh rows the kernel writes have `dsh_coeffs` coefficients - 16 (the reference's layout; the inactive bands are
// written as zeros) or fewer, down to (DEG+1)^2 (rows trimmed to what can be non-zero: DAS3R trains at degree 0 for
// 3000 of its 4000 iterations, train_gui.py:541-543, where 180 of the 192 bytes of a full row are identically zero).
// ---------------------------------------------------------------------------------------------------------------
constexpr int MV_THREADS = 128;
constexpr int MV_OUT_PITCH = MV_THREADS + 1;  // [coefficient][thread] accumulators, odd pitch: conflict-free both ways

#ifndef DAS3R_MV_MIN_BLOCKS
#define DAS3R_MV_MIN_BLOCKS 4
#endif
template <int DEG>
__global__ void __launch_bounds__(MV_THREADS, DAS3R_MV_MIN_BLOCKS) preprocess_bwd_multi_kernel(const __grid_constant__ MultiBwdArgs a) {
    constexpr int NW = 3 * (DEG + 1) * (DEG + 1);  // active floats of an SH row
    constexpr bool STAGE = DEG == 3;
    extern __shared__ __align__(16) unsigned char mv_smem[];
    // the staged input rows are consumed into registers before the accumulators are first touched: one buffer serves both
    float4* in_rows = reinterpret_cast<float4*>(mv_smem);  // [MV_THREADS][13] (STAGE only)
    float* out_acc = reinterpret_cast<float*>(mv_smem);    // [NW][pitch]
    __shared__ uint8_t row_vis[MV_THREADS];
    const int tid = threadIdx.x;
    const int base = blockIdx.x * MV_THREADS;
    const int idx = base + tid;
    const bool in_range = idx < a.P;
    const bool accum = a.accumulate_shared != 0;
    const bool accum_view = a.accumulate_view != 0;
    const int nvalid = min(MV_THREADS, a.P - base);
    if (STAGE) {
        const float4* src = reinterpret_cast<const float4*>(a.shs) + (size_t)base * PB_ROW_F4;
        const int total = nvalid * PB_ROW_F4;
#pragma unroll
        for (int k = 0; k < PB_ROW_F4; k++) {
            const int e = tid + k * MV_THREADS;
            if (e < total) {
                const int g = e / PB_ROW_F4, j = e - g * PB_ROW_F4;
                cp_async_16(&in_rows[g * PB_ROW_PAD_F4 + j], src + e);
            }
        }
        cp_async_commit();
    }

    float3 s_mean = {0.f, 0.f, 0.f}, s_scale = {0.f, 0.f, 0.f};
    float4 s_rot = {0.f, 0.f, 0.f, 0.f};
    float s_opac = 0.f;
    bool any_vis = false;

    float f[48];
    if (STAGE) {
        cp_async_wait_all();
        __syncthreads();
        const float4* myrow = &in_rows[tid * PB_ROW_PAD_F4];
#pragma unroll
        for (int j = 0; j < PB_ROW_F4; j++) {
            const float4 v = myrow[j];
            f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w;
        }
        __syncthreads();  // every row has been read: the buffer becomes the accumulator array
    } else if (in_range) {
        load_sh_row(a.shs + (size_t)idx * 48, DEG, true, f);
    }
#pragma unroll
    for (int k = 0; k < NW; k++) out_acc[k * MV_OUT_PITCH + tid] = 0.f;

#pragma unroll 1
    for (int v = 0; v < a.n_views; v++) {
        const MultiBwdView& V = a.v[v];
        if (!in_range) continue;
        if (V.radii[idx] <= 0) {
            // not rendered in this view: a per-view output still has to be defined
            if (!accum_view) {
                if (V.dL_dmeans3D) {
                    V.dL_dmeans3D[3 * idx] = 0.f; V.dL_dmeans3D[3 * idx + 1] = 0.f; V.dL_dmeans3D[3 * idx + 2] = 0.f;
                }
                if (V.dL_drots) reinterpret_cast<float4*>(V.dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            continue;
        }
        any_vis = true;
        const float4* acc4 = reinterpret_cast<const float4*>(V.acc + (size_t)idx * ACC_STRIDE);
        const float4 q0 = acc4[0], q1 = acc4[1], q2 = acc4[2];
        const float3 mean = {V.means[3 * idx], V.means[3 * idx + 1], V.means[3 * idx + 2]};
        const float3 sc = {V.scales[3 * idx], V.scales[3 * idx + 1], V.scales[3 * idx + 2]};
        const float4 rq = __ldg(reinterpret_cast<const float4*>(V.rots) + idx);
        const uint8_t cl = V.clamped[idx];
        float cov3D[6];
        cov3d_from_scale_rot(sc, V.scale_mod, rq, cov3D);
        const ViewConst vc{V.view, V.proj, V.focal_x, V.focal_y, V.tan_fovx, V.tan_fovy, V.scale_mod};
        GeomGrad gg;
        geom_backward(vc, mean, cov3D, true, sc, rq, q0.x, q0.y, make_float3(q0.z, q0.w, q1.x), gg);
        const float3 cam = {V.campos[0], V.campos[1], V.campos[2]};
        const float3 dm = sh_backward(
            DEG, mean, cam, cl, make_float3(q1.z, q1.w, q2.x),
            [&](int k) { return make_float3(f[3 * k], f[3 * k + 1], f[3 * k + 2]); },
            [&](int k, float3 g) {
                out_acc[(3 * k + 0) * MV_OUT_PITCH + tid] += g.x;
                out_acc[(3 * k + 1) * MV_OUT_PITCH + tid] += g.y;
                out_acc[(3 * k + 2) * MV_OUT_PITCH + tid] += g.z;
            });
        const float3 gm = {gg.dmean.x + dm.x, gg.dmean.y + dm.y, gg.dmean.z + dm.z};
        if (V.dL_dmeans3D) {
            float* o = V.dL_dmeans3D + 3 * (size_t)idx;
            o[0] = accum_view ? o[0] + gm.x : gm.x;
            o[1] = accum_view ? o[1] + gm.y : gm.y;
            o[2] = accum_view ? o[2] + gm.z : gm.z;
        } else {
            s_mean.x += gm.x; s_mean.y += gm.y; s_mean.z += gm.z;
        }
        if (V.dL_drots) {
            float4* o = reinterpret_cast<float4*>(V.dL_drots) + idx;
            float4 r = gg.drot;
            if (accum_view) {
                const float4 p = *o;
                r.x += p.x; r.y += p.y; r.z += p.z; r.w += p.w;
            }
            *o = r;
        } else {
            s_rot.x += gg.drot.x; s_rot.y += gg.drot.y; s_rot.z += gg.drot.z; s_rot.w += gg.drot.w;
        }
        s_scale.x += gg.dscale.x; s_scale.y += gg.dscale.y; s_scale.z += gg.dscale.z;
        s_opac += q1.y;
    }

    row_vis[tid] = any_vis ? 1 : 0;
    if (in_range && (any_vis || !accum)) {
        auto put1 = [accum](float* p, float v) { *p = accum ? *p + v : v; };
        put1(&a.dL_dopacity[idx], s_opac);
        put1(&a.dL_dmeans3D[3 * idx + 0], s_mean.x);
        put1(&a.dL_dmeans3D[3 * idx + 1], s_mean.y);
        put1(&a.dL_dmeans3D[3 * idx + 2], s_mean.z);
        put1(&a.dL_dscales[3 * idx + 0], s_scale.x);
        put1(&a.dL_dscales[3 * idx + 1], s_scale.y);
        put1(&a.dL_dscales[3 * idx + 2], s_scale.z);
        float4* dr = reinterpret_cast<float4*>(a.dL_drots) + idx;
        if (accum) {
            const float4 o = *dr;
            s_rot.x += o.x; s_rot.y += o.y; s_rot.z += o.z; s_rot.w += o.w;
        }
        *dr = s_rot;
    }
    __syncthreads();
    // dL_dsh rows of the block, [nvalid][3 * dsh_coeffs] floats, contiguous in memory: coalesced write-out from the
    // transposed accumulators; floats beyond the active bands are zeros (overwrite mode) or left alone (accumulate)
    const int row_f = 3 * a.dsh_coeffs;
    float* dst_f = a.dL_dsh + (size_t)base * row_f;
    if ((row_f & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst_f) & 15) == 0)) {
        const int row_f4 = row_f >> 2;
        const int total = nvalid * row_f4;
        float4* dst = reinterpret_cast<float4*>(dst_f);
        for (int e = tid; e < total; e += MV_THREADS) {
            const int g = e / row_f4, j = e - g * row_f4;
            if (accum && (!row_vis[g] || 4 * j >= NW)) continue;  // nothing to add
            float4 o;
            o.x = (4 * j + 0 < NW) ? out_acc[(4 * j + 0) * MV_OUT_PITCH + g] : 0.f;
            o.y = (4 * j + 1 < NW) ? out_acc[(4 * j + 1) * MV_OUT_PITCH + g] : 0.f;
            o.z = (4 * j + 2 < NW) ? out_acc[(4 * j + 2) * MV_OUT_PITCH + g] : 0.f;
            o.w = (4 * j + 3 < NW) ? out_acc[(4 * j + 3) * MV_OUT_PITCH + g] : 0.f;
            if (accum) {
                const float4 p = dst[e];
                o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
            }
            dst[e] = o;
        }
    } else {
        const int total = nvalid * row_f;
        for (int e = tid; e < total; e += MV_THREADS) {
            const int g = e / row_f, j = e - g * row_f;
            if (accum && (!row_vis[g] || j >= NW)) continue;
            float o = (j < NW) ? out_acc[j * MV_OUT_PITCH + g] : 0.f;
            if (accum) o += dst_f[e];
            dst_f[e] = o;
        }
    }
}

cudaError_t launch_preprocess_bwd_multi(const MultiBwdArgs& a, cudaStream_t s) {
    const int blocks = (a.P + MV_THREADS - 1) / MV_THREADS;
    const int nw = 3 * (a.D + 1) * (a.D + 1);
    const size_t smem = std::max((a.D == 3 ? (size_t)MV_THREADS * PB_ROW_PAD_F4 * 16 : 0), (size_t)nw * MV_OUT_PITCH * sizeof(float));
    const void* fn = a.D == 0   ? (const void*)preprocess_bwd_multi_kernel<0>
                     : a.D == 1 ? (const void*)preprocess_bwd_multi_kernel<1>
                     : a.D == 2 ? (const void*)preprocess_bwd_multi_kernel<2>
                                : (const void*)preprocess_bwd_multi_kernel<3>;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    switch (a.D) {
        case 0: preprocess_bwd_multi_kernel<0><<<blocks, MV_THREADS, smem, s>>>(a); break;
        case 1: preprocess_bwd_multi_kernel<1><<<blocks, MV_THREADS, smem, s>>>(a); break;
        case 2: preprocess_bwd_multi_kernel<2><<<blocks, MV_THREADS, smem, s>>>(a); break;
        default: preprocess_bwd_multi_kernel<3><<<blocks, MV_THREADS, smem, s>>>(a); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_preprocess_bwd(const PreBwdArgs& a, cudaStream_t s) {
    const int blocks = (a.P + PB_THREADS - 1) / PB_THREADS;
    static const bool allow_stage = getenv("DAS3R_NO_SH_STAGE") == nullptr;
    const bool stage = allow_stage && a.shs != nullptr && a.dL_dsh != nullptr && a.D == 3 && a.M == 16 &&
                       ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0);
    if (stage) {
        const size_t smem = (size_t)PB_THREADS * PB_ROW_PAD_F4 * 16;
        cudaError_t e = cudaFuncSetAttribute(preprocess_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return e;
        preprocess_bwd_kernel<true><<<blocks, PB_THREADS, smem, s>>>(a);
    } else {
        preprocess_bwd_kernel<false><<<blocks, PB_THREADS, 0, s>>>(a);
    }
    return cudaGetLastError();
}

}  // namespace das3r
