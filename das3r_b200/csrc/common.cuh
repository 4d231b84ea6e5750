// Shared device helpers for the sm_100a rasterizer kernels.
//
// The small matrix type below follows glm's column-major convention and the exact operand order of glm 0.9.9.9's
// mat3 product (reference: third_party/glm/glm/detail/type_mat3x3.inl:486-518), because radii and tile rectangles
// are discontinuous functions of these fp32 values and must come out bit-identical to the reference
// (SURVEY.md §7 hard part 1). Do not "simplify" the expression shapes in this file.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "kernels.cuh"

namespace das3r {

constexpr float NEAR_Z = 0.001f;    // DAS3R README.md:41-44: auxiliary.h:154 patched 0.2f -> 0.001f

// SH basis constants (auxiliary.h:21-38).
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                      0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

struct Mat3 {
    float m[3][3];  // m[col][row]
};

// Column-major constructor, same argument order as glm::mat3(x0,y0,z0, x1,y1,z1, x2,y2,z2).
__device__ __forceinline__ Mat3 mat3(float x0, float y0, float z0, float x1, float y1, float z1, float x2, float y2,
                                     float z2) {
    Mat3 r;
    r.m[0][0] = x0; r.m[0][1] = y0; r.m[0][2] = z0;
    r.m[1][0] = x1; r.m[1][1] = y1; r.m[1][2] = z1;
    r.m[2][0] = x2; r.m[2][1] = y2; r.m[2][2] = z2;
    return r;
}

// R[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2]   (type_mat3x3.inl:509-517)
__device__ __forceinline__ Mat3 mul(const Mat3& A, const Mat3& B) {
    Mat3 R;
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
        for (int r = 0; r < 3; r++) {
            R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
        }
    }
    return R;
}

__device__ __forceinline__ Mat3 transpose(const Mat3& A) {
    Mat3 R;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) R.m[c][r] = A.m[r][c];
    return R;
}

// auxiliary.h:58-66
__device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* __restrict__ m) {
    float3 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
    };
    return t;
}

// auxiliary.h:68-77
__device__ __forceinline__ float4 xform_point_4x4(const float3& p, const float* __restrict__ m) {
    float4 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
    return t;
}

// auxiliary.h:89-97
__device__ __forceinline__ float3 xform_vec_4x3_T(const float3& p, const float* __restrict__ m) {
    float3 t = {
        m[0] * p.x + m[1] * p.y + m[2] * p.z,
        m[4] * p.x + m[5] * p.y + m[6] * p.z,
        m[8] * p.x + m[9] * p.y + m[10] * p.z,
    };
    return t;
}

// auxiliary.h:41-44 — evaluated in double because of the 1.0 / 0.5 literals, then rounded to float.
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// auxiliary.h:46-56
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rmin, uint2& rmax, unsigned gx,
                                          unsigned gy) {
    rmin = {min(gx, max((int)0, (int)((p.x - max_radius) / TILE_X))),
            min(gy, max((int)0, (int)((p.y - max_radius) / TILE_Y)))};
    rmax = {min(gx, max((int)0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X))),
            min(gy, max((int)0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)))};
}

// forward.cu:118-152 — scale/rotation -> 3D covariance (quaternion deliberately NOT normalised, :127).
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 scale, float mod, const float4 rot, float* cov3D) {
    Mat3 S = mat3(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
    S.m[0][0] = mod * scale.x;
    S.m[1][1] = mod * scale.y;
    S.m[2][2] = mod * scale.z;
    float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    // The products of the rotation matrix are written with explicit rounding/fusion intrinsics: in the reference
    // source they are plain `a*b +/- c*d` expressions whose FMA contraction is decided by ptxas, and the choice it made
    // for the reference's preprocessCUDA (sm_100 build, CUDA 12.9: x*z rounded and r*y fused, etc.) is pinned here so
    // that the tiny off-diagonal covariance entries - and through them conic, radii and tile rectangles - come out
    // bit-identical no matter how this kernel is scheduled.
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
    const float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
    const float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
    const float yy_zz = __fadd_rn(yy, zz), xx_zz = __fmaf_rn(x, x, zz), xx_yy = __fmaf_rn(x, x, yy);
    Mat3 R = mat3(__fsub_rn(1.f, __fadd_rn(yy_zz, yy_zz)), __fadd_rn(xy_m_rz, xy_m_rz), __fadd_rn(xz_p_ry, xz_p_ry),
                  __fadd_rn(xy_p_rz, xy_p_rz), __fsub_rn(1.f, __fadd_rn(xx_zz, xx_zz)), __fadd_rn(yz_m_rx, yz_m_rx),
                  __fadd_rn(xz_m_ry, xz_m_ry), __fadd_rn(yz_p_rx, yz_p_rx), __fsub_rn(1.f, __fadd_rn(xx_yy, xx_yy)));
    Mat3 M = mul(S, R);
    Mat3 Sigma = mul(transpose(M), M);
    cov3D[0] = Sigma.m[0][0];
    cov3D[1] = Sigma.m[0][1];
    cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1];
    cov3D[4] = Sigma.m[1][2];
    cov3D[5] = Sigma.m[2][2];
}

// Pieces of the EWA projection shared by forward (forward.cu:74-113) and backward (backward.cu:170-205).
struct Ewa {
    Mat3 T;       // W * J
    Mat3 Vrk;
    float3 t;     // clamped view-space mean
    float txtz, tytz, limx, limy;
    float a, b, c;  // cov2D[0][0] + 0.3, cov2D[0][1], cov2D[1][1] + 0.3
};

__device__ __forceinline__ Ewa ewa_project(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                           float tan_fovy, const float* cov3D, const float* __restrict__ view) {
    Ewa e;
    float3 t = xform_point_4x3(mean, view);
    e.limx = 1.3f * tan_fovx;
    e.limy = 1.3f * tan_fovy;
    e.txtz = t.x / t.z;
    e.tytz = t.y / t.z;
    t.x = min(e.limx, max(-e.limx, e.txtz)) * t.z;
    t.y = min(e.limy, max(-e.limy, e.tytz)) * t.z;
    e.t = t;
    Mat3 J = mat3(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z), 0.0f, focal_y / t.z,
                  -(focal_y * t.y) / (t.z * t.z), 0, 0, 0);
    Mat3 W = mat3(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    e.T = mul(W, J);
    e.Vrk = mat3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    Mat3 cov = mul(mul(transpose(e.T), transpose(e.Vrk)), e.T);
    e.a = cov.m[0][0] + 0.3f;
    e.b = cov.m[0][1];
    e.c = cov.m[1][1] + 0.3f;
    return e;
}

// Loads the first n floats (n = 3*(deg+1)^2) of an SH row into f[48]; 128-bit loads when the row is 16-byte aligned.
__device__ __forceinline__ void load_sh_row(const float* __restrict__ row, int deg, bool aligned, float* f) {
    if (aligned) {
        const float4* r4 = reinterpret_cast<const float4*>(row);
        if (deg == 0) {
            f[0] = __ldg(row); f[1] = __ldg(row + 1); f[2] = __ldg(row + 2);
        } else if (deg == 1) {
#pragma unroll
            for (int j = 0; j < 3; j++) {
                float4 v = __ldg(r4 + j);
                f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w;
            }
        } else if (deg == 2) {
#pragma unroll
            for (int j = 0; j < 6; j++) {
                float4 v = __ldg(r4 + j);
                f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w;
            }
            f[24] = __ldg(row + 24); f[25] = __ldg(row + 25); f[26] = __ldg(row + 26);
        } else {
#pragma unroll
            for (int j = 0; j < 12; j++) {
                float4 v = __ldg(r4 + j);
                f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w;
            }
        }
    } else {
        if (deg == 0) {
#pragma unroll
            for (int j = 0; j < 3; j++) f[j] = __ldg(row + j);
        } else if (deg == 1) {
#pragma unroll
            for (int j = 0; j < 12; j++) f[j] = __ldg(row + j);
        } else if (deg == 2) {
#pragma unroll
            for (int j = 0; j < 27; j++) f[j] = __ldg(row + j);
        } else {
#pragma unroll
            for (int j = 0; j < 48; j++) f[j] = __ldg(row + j);
        }
    }
}

// ---- small PTX wrappers -------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// cp.async (LDGSTS): 16 bytes global -> shared without a register round trip; completion by commit / wait groups.
// (The first blend kernel staged whole batches with TMA bulk copies + mbarrier; the warp-autonomous design gathers
// individual 48-byte records, which is what LDGSTS is for - DESIGN.md 3.3.)
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() {  // all but the N most recent groups have landed
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

}  // namespace das3r
