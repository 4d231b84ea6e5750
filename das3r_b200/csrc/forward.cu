// Forward path of the sm_100a Gaussian-splat rasterizer:
//   preprocess_fwd  -> per-Gaussian projection / EWA covariance / SH colour / tile rectangle  (replaces
//                      preprocessCUDA, reference cuda_rasterizer/forward.cu:155-256) + per-tile histogram
//   tile_scan       -> exclusive scan of the T tile counts = tile ranges (replaces the P-sized InclusiveSum,
//                      rasterizer_impl.cu:277, and identifyTileRanges, :116-138)
//   scatter         -> one (depth|index) entry per (Gaussian, tile) pair into its tile bucket (replaces
//                      duplicateWithKeys, rasterizer_impl.cu:70-111)
//   tile_sort       -> per-tile most-significant-digit bucket sort in shared memory, three size classes, bitonic
//                      network for skewed depths, library segmented sort only for tiles above 16384 entries (replaces
//                      the global 64-bit cub::DeviceRadixSort, rasterizer_impl.cu:303-308)
//   rect_lists      -> gathers the tile-ordered record streams through the sorted point list and builds, per 8x4-pixel
//                      rectangle, the list of {tile-list position, pixel mask} of the Gaussians that can
//                      reach alpha >= 1/255 there (no reference counterpart: there every thread of the tile looks at
//                      every Gaussian of the tile)
//   blend_fwd       -> front-to-back alpha blend, one warp per rectangle walking its list (replaces renderCUDA,
//                      forward.cu:261-374; see blend_common.cuh)
//
// The result of scatter+tile_sort is bit-identical to the reference's stable LSD radix sort of (tile<<32|depth)
// keys: equal keys keep emission order there, i.e. ascending Gaussian index, so the total order is
// (tile, depth bits, Gaussian index) and any correct sort on that triple reproduces point_list exactly.
#include "common.cuh"
#include "kernels.cuh"
#include "blend_common.cuh"
#include <cub/device/device_segmented_radix_sort.cuh>
#include <stdio.h>
#include <map>
#include <mutex>

namespace das3r {

// ---------------------------------------------------------------------------------------------------------------
// SH evaluation (forward.cu:20-71). f[] holds the row as 3*k+c floats; only the first 3*(deg+1)^2 are defined.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float3 sh_to_rgb(int deg, const float* f, float3 pos, float3 campos, uint8_t& clamped) {
    float3 dir = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x = dir.x / len;
    dir.y = dir.y / len;
    dir.z = dir.z / len;
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = __fmul_rn(kSH0, f[c]);  // rounded product, as in the reference's build: left to
                                                                  // the compiler, this multiply or the next one gets
                                                                  // fused into the subtraction below depending on the
                                                                  // surrounding kernel (seen in SASS: one-view vs
                                                                  // multi-view), a 1-ulp difference in the colour
    if (deg > 0) {
        float x = dir.x, y = dir.y, z = dir.z;
#pragma unroll
        for (int c = 0; c < 3; c++)
            res[c] = res[c] - kSH1 * y * f[3 + c] + kSH1 * z * f[6 + c] - kSH1 * x * f[9 + c];
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
            for (int c = 0; c < 3; c++)
                res[c] = res[c] + kSH2[0] * xy * f[12 + c] + kSH2[1] * yz * f[15 + c] +
                         kSH2[2] * (2.0f * zz - xx - yy) * f[18 + c] + kSH2[3] * xz * f[21 + c] +
                         kSH2[4] * (xx - yy) * f[24 + c];
            if (deg > 2) {
#pragma unroll
                for (int c = 0; c < 3; c++)
                    res[c] = res[c] + kSH3[0] * y * (3.0f * xx - yy) * f[27 + c] + kSH3[1] * xy * z * f[30 + c] +
                             kSH3[2] * y * (4.0f * zz - xx - yy) * f[33 + c] +
                             kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * f[36 + c] +
                             kSH3[4] * x * (4.0f * zz - xx - yy) * f[39 + c] + kSH3[5] * z * (xx - yy) * f[42 + c] +
                             kSH3[6] * x * (xx - 3.0f * yy) * f[45 + c];
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] += 0.5f;
    clamped = (uint8_t)((res[0] < 0 ? 1 : 0) | (res[1] < 0 ? 2 : 0) | (res[2] < 0 ? 4 : 0));
    return make_float3(res[0] < 0.0f ? 0.0f : res[0], res[1] < 0.0f ? 0.0f : res[1], res[2] < 0.0f ? 0.0f : res[2]);
}

// SH degrees 0..2 (the degrees DAS3R trains at) with every product and sum written as an explicit intrinsic: left to
// the compiler, expressions like xx - yy are contracted into different FMAs in the one-view and the multi-view kernel
// (seen at degree 2: a 1-ulp colour difference), and the two must agree bit for bit. Same formulas as sh_to_rgb
// (forward.cu:20-71); the result is within an ulp or two of the reference's build, far inside the 1e-5 bar. Degree 3
// keeps the plain expressions above (both kernels were verified equal there, and equal to the reference's build).
__device__ __forceinline__ float3 sh_to_rgb_low(int deg, const float* f, float3 pos, float3 campos, uint8_t& clamped) {
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = __fmul_rn(kSH0, f[c]);
    if (deg > 0) {
        const float dx = __fsub_rn(pos.x, campos.x), dy = __fsub_rn(pos.y, campos.y), dz = __fsub_rn(pos.z, campos.z);
        const float len = sqrtf(__fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx))));
        const float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);
        const float b1 = __fmul_rn(kSH1, y), b2 = __fmul_rn(kSH1, z), b3 = __fmul_rn(kSH1, x);
#pragma unroll
        for (int c = 0; c < 3; c++)
            res[c] = __fmaf_rn(-b3, f[9 + c], __fmaf_rn(b2, f[6 + c], __fmaf_rn(-b1, f[3 + c], res[c])));
        if (deg > 1) {
            const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
            const float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
            const float c0 = __fmul_rn(kSH2[0], xy), c1 = __fmul_rn(kSH2[1], yz);
            const float c2 = __fmul_rn(kSH2[2], __fsub_rn(__fsub_rn(__fadd_rn(zz, zz), xx), yy));
            const float c3 = __fmul_rn(kSH2[3], xz), c4 = __fmul_rn(kSH2[4], __fsub_rn(xx, yy));
#pragma unroll
            for (int c = 0; c < 3; c++)
                res[c] = __fmaf_rn(c4, f[24 + c], __fmaf_rn(c3, f[21 + c], __fmaf_rn(c2, f[18 + c],
                             __fmaf_rn(c1, f[15 + c], __fmaf_rn(c0, f[12 + c], res[c])))));
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = __fadd_rn(res[c], 0.5f);
    clamped = (uint8_t)((res[0] < 0 ? 1 : 0) | (res[1] < 0 ? 2 : 0) | (res[2] < 0 ? 4 : 0));
    return make_float3(res[0] < 0.0f ? 0.0f : res[0], res[1] < 0.0f ? 0.0f : res[1], res[2] < 0.0f ? 0.0f : res[2]);
}

constexpr int PRE_THREADS = 256;
constexpr int SH_ROW_F4 = 12;      // 48 floats = 12 float4 per degree-3 row
constexpr int SH_ROW_PAD_F4 = 13;  // padded row pitch in shared memory: 13 float4 (odd) -> conflict-free LDS.128

// ---------------------------------------------------------------------------------------------------------------
// preprocess_fwd. STAGE_SH: the block's 256 degree-3 SH rows (48 KB contiguous) are brought into shared memory with
// coalesced 16-byte cp.async while the geometry is computed, then read back row-wise without bank conflicts.
// ---------------------------------------------------------------------------------------------------------------
// What the geometry half of the preprocess leaves for the colour / record half.
struct PreGeom {
    bool vis;
    float my_radius;
    uint2 rmin, rmax;
    float2 point_image;
    float3 p_orig, conic;
    float depth, opacity, hrad, negtau;
};

// Geometry half of preprocessCUDA (forward.cu:182-237) for Gaussian idx of one view: near cull, projection, EWA
// covariance, conic, radius, tile rectangle, and the culling radius / threshold of the blend. Shared by the one-view and
// the multi-view kernel (same code, same rounding).
__device__ __forceinline__ PreGeom pre_geometry(const PreFwdArgs& a, const int idx) {
    bool vis = false;
    float my_radius = 0.f;
    uint2 rmin = {0, 0}, rmax = {0, 0};
    float2 point_image = {0.f, 0.f};
    float3 p_orig = {0.f, 0.f, 0.f};
    float3 conic = {0.f, 0.f, 0.f};
    float depth = 0.f, opacity = 0.f, hrad = -1.f, negtau = 1.f;

    if (idx < a.P) {
        p_orig = {a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]};
        // in_frustum (auxiliary.h:139-164) with the 0.001 near plane.
        float4 p_hom = xform_point_4x4(p_orig, a.proj);
        float p_w = 1.0f / (p_hom.w + 0.0000001f);
        float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
        float3 p_view = xform_point_4x3(p_orig, a.view);
        if (p_view.z <= NEAR_Z) {
            if (a.prefiltered) {
                printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                __trap();
            }
        } else {
            float cov3D[6];
            if (a.cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; k++) cov3D[k] = a.cov3D_precomp[6 * idx + k];
            } else {
                float3 sc = {a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]};
                float4 rq = __ldg(reinterpret_cast<const float4*>(a.rots) + idx);
                cov3d_from_scale_rot(sc, a.scale_mod, rq, cov3D);
            }
            Ewa e = ewa_project(p_orig, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, cov3D, a.view);
            float3 cov = {e.a, e.b, e.c};
            // forward.cu:218-237
            float det = (cov.x * cov.z - cov.y * cov.y);
            if (det != 0.0f) {
                float det_inv = 1.f / det;
                conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
                float mid = 0.5f * (cov.x + cov.z);
                float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
                float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
                my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
                point_image = {ndc_to_pix(p_proj.x, a.W), ndc_to_pix(p_proj.y, a.H)};
                tile_rect(point_image, (int)my_radius, rmin, rmax, a.gx, a.gy);
                if ((rmax.x - rmin.x) * (rmax.y - rmin.y) != 0) {
                    vis = true;
                    depth = p_view.z;
                    opacity = a.opac[idx];
                    // A pair can only contribute if alpha = opacity*exp(power) >= 1/255, i.e. power >= -tau with
                    // tau = ln(255*opacity). tau' = tau*1.0001 + 1e-4 (margin >> expf's error) is stored negated for the
                    // per-pixel test; h = radius of the bounding box of the ellipse {0.5*d^T Q d <= tau'} (computed in
                    // double from the ROUNDED conic the blend uses, plus margins) is the warp-level test. A pair skipped
                    // by either test is always one the reference skips by `alpha < 1/255` (forward.cu:340).
                    float o255 = 255.f * opacity;
                    if (o255 > 1.0f) {
                        double tau = (double)logf(o255) * 1.0001 + 1e-4;
                        negtau = -(float)(tau * 1.000001);
                        double qa = conic.x, qb = conic.y, qc = conic.z;
                        double qdet = qa * qc - qb * qb;
                        if (qdet > 0.0 && qa > 0.0 && qc > 0.0) {
                            double hx = sqrt(2.0 * tau * qc / qdet) * 1.0001 + 0.01;
                            double hy = sqrt(2.0 * tau * qa / qdet) * 1.0001 + 0.01;
                            hrad = (float)(fmax(hx, hy) * 1.000001);
                        } else {
                            hrad = 3.0e38f;  // degenerate conic: never cull at warp level
                            negtau = -3.0e38f;
                        }
                    }
                }
            }
        }
    }

    PreGeom g;
    g.vis = vis; g.my_radius = my_radius; g.rmin = rmin; g.rmax = rmax; g.point_image = point_image;
    g.p_orig = p_orig; g.conic = conic; g.depth = depth; g.opacity = opacity; g.hrad = hrad; g.negtau = negtau;
    return g;
}

// Colour / record half: radii and tile count, SH -> RGB (the row comes from shared memory when STAGE_SH), the 48-byte
// record, clamp bits, tile histogram (forward.cu:239-255).
// f_low: the first 27 floats of the Gaussian's SH row already in registers (degrees 0..2, multi-view kernel), or NULL.
// KNOWN_DEG >= 0: the degree as a compile-time constant (multi-view kernel: f_low then holds exactly its active floats).
template <bool STAGE_SH, int KNOWN_DEG = -1>
__device__ __forceinline__ void pre_finish(const PreFwdArgs& a, const int idx, const PreGeom& g,
                                           const float4* __restrict__ sh_row, uint32_t* __restrict__ hist,
                                           const float* f_low = nullptr) {
    if (idx < a.P) {
        a.radii[idx] = g.vis ? (int)g.my_radius : 0;
        a.tiles_touched[idx] = g.vis ? (g.rmax.y - g.rmin.y) * (g.rmax.x - g.rmin.x) : 0u;
    }
    if (g.vis) {
        float3 rgb;
        uint8_t cl = 0;
        if (a.colors_precomp == nullptr) {
            const float3 cam = {a.campos[0], a.campos[1], a.campos[2]};
            if constexpr (!STAGE_SH && KNOWN_DEG >= 0) {
                rgb = sh_to_rgb_low(KNOWN_DEG, f_low, g.p_orig, cam, cl);
            } else if (!STAGE_SH && a.D < 3) {
                float fl[27];
                const float* row = a.shs + (size_t)idx * 3 * a.M;
                const bool aligned = ((a.M & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
                load_sh_row(row, a.D, aligned, fl);
                rgb = sh_to_rgb_low(a.D, fl, g.p_orig, cam, cl);
            } else {
                float f[48];
                if (STAGE_SH) {
#pragma unroll
                    for (int j = 0; j < SH_ROW_F4; j++) {
                        float4 v = sh_row[j];
                        f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w;
                    }
                } else {
                    const float* row = a.shs + (size_t)idx * 3 * a.M;
                    bool aligned = ((a.M & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
                    load_sh_row(row, a.D, aligned, f);
                }
                rgb = sh_to_rgb(a.D, f, g.p_orig, cam, cl);
            }
        } else {
            rgb = {a.colors_precomp[3 * idx], a.colors_precomp[3 * idx + 1], a.colors_precomp[3 * idx + 2]};
        }
        GaussRec r;
        r.A = make_float4(g.point_image.x, g.point_image.y, g.hrad, g.negtau);
        r.B = make_float4(g.conic.x, g.conic.y, g.conic.z, g.opacity);
        r.C = make_float4(rgb.x, rgb.y, rgb.z, g.depth);
        a.rec[idx] = r;
        a.clamped[idx] = cl;
        // Tile histogram (this replaces the prefix sum over all P Gaussians of the reference).
        for (unsigned y = g.rmin.y; y < g.rmax.y; y++)
            for (unsigned x = g.rmin.x; x < g.rmax.x; x++) {
                if (a.smem_hist)
                    atomicAdd(&hist[y * a.gx + x], 1u);
                else
                    atomicAdd(&a.tile_counts[y * a.gx + x], 1u);
            }
    }
}

template <bool STAGE_SH>
__global__ void __launch_bounds__(PRE_THREADS) preprocess_fwd_kernel(const PreFwdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* sh_stage = reinterpret_cast<float4*>(smem_raw);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw + (STAGE_SH ? PRE_THREADS * SH_ROW_PAD_F4 * 16 : 0));

    const int tid = threadIdx.x;
    const int base = blockIdx.x * PRE_THREADS;
    const int idx = base + tid;
    const int T = (int)(a.gx * a.gy);

    if (a.smem_hist)
        for (int i = tid; i < T; i += PRE_THREADS) hist[i] = 0;

    if (STAGE_SH) {
        const int nvalid = min(PRE_THREADS, a.P - base);
        const float4* src = reinterpret_cast<const float4*>(a.shs) + (size_t)base * SH_ROW_F4;
        const int total = nvalid * SH_ROW_F4;
#pragma unroll
        for (int k = 0; k < SH_ROW_F4; k++) {
            int e = tid + k * PRE_THREADS;
            if (e < total) {
                int g = e / SH_ROW_F4, j = e - g * SH_ROW_F4;
                cp_async_16(&sh_stage[g * SH_ROW_PAD_F4 + j], src + e);
            }
        }
        cp_async_commit();
    }

    PreGeom g = pre_geometry(a, idx);

    if (STAGE_SH) cp_async_wait_all();
    __syncthreads();

    pre_finish<STAGE_SH>(a, idx, g, sh_stage + tid * SH_ROW_PAD_F4, hist);
    if (a.smem_hist) {
        __syncthreads();
        for (int i = tid; i < T; i += PRE_THREADS) {
            uint32_t v = hist[i];
            if (v) atomicAdd(&a.tile_counts[i], v);
        }
    }
}

// preprocess_fwd for SEVERAL views that share their SH rows. STAGE (degree 3, 16 coefficients): the block's 256 rows are
// staged once, every thread then runs the geometry and colour halves for each view with the row in shared memory - the
// 192 bytes of SH per Gaussian, two thirds of the kernel's traffic, are read once per step instead of once per view.
// !STAGE (degrees 0..2, the regime DAS3R trains in: train_gui.py:541-543): a thread reads the 12..108 active bytes of
// its row per view (first view from HBM, the others from L1/L2); what the views share here is the launch, the tile
// histograms and the Gaussian-major sweep. Same device functions as the one-view kernel, so every output is
// bit-identical to it (tests/test_gpu_multiview.py).
template <int DEG>
__global__ void __launch_bounds__(PRE_THREADS) preprocess_fwd_multi_kernel(const __grid_constant__ MultiFwdArgs m) {
    constexpr bool STAGE = DEG == 3;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* sh_stage = reinterpret_cast<float4*>(smem_raw);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw + (STAGE ? PRE_THREADS * SH_ROW_PAD_F4 * 16 : 0));
    const int tid = threadIdx.x;
    const int base = blockIdx.x * PRE_THREADS;
    const int idx = base + tid;
    const int P = m.v[0].P, T = m.T;
    if (m.smem_hist)
        for (int i = tid; i < T * m.n_views; i += PRE_THREADS) hist[i] = 0;
    if (STAGE) {
        const int nvalid = min(PRE_THREADS, P - base);
        const float4* src = reinterpret_cast<const float4*>(m.v[0].shs) + (size_t)base * SH_ROW_F4;
        const int total = nvalid * SH_ROW_F4;
#pragma unroll
        for (int k = 0; k < SH_ROW_F4; k++) {
            int e = tid + k * PRE_THREADS;
            if (e < total) {
                int g = e / SH_ROW_F4, j = e - g * SH_ROW_F4;
                cp_async_16(&sh_stage[g * SH_ROW_PAD_F4 + j], src + e);
            }
        }
        cp_async_commit();
        cp_async_wait_all();
    }
    __syncthreads();
    // degrees 0..2: the thread's active SH floats (3, 12 or 27) stay in registers for all views
    float fl[STAGE ? 1 : 3 * (DEG + 1) * (DEG + 1)];
    if (!STAGE && idx < P) {
        const PreFwdArgs& a0 = m.v[0];
        const bool aligned = ((a0.M & 3) == 0) && ((reinterpret_cast<uintptr_t>(a0.shs) & 15) == 0);
        load_sh_row(a0.shs + (size_t)idx * 3 * a0.M, DEG, aligned, fl);
    }
#pragma unroll 1
    for (int v = 0; v < m.n_views; v++) {
        const PreFwdArgs& a = m.v[v];
        const PreGeom g = pre_geometry(a, idx);
        pre_finish<STAGE, DEG>(a, idx, g, sh_stage + tid * SH_ROW_PAD_F4, hist + (size_t)v * T, STAGE ? nullptr : fl);
    }
    if (m.smem_hist) {
        __syncthreads();
        for (int i = tid; i < T * m.n_views; i += PRE_THREADS) {
            const uint32_t c = hist[i];
            if (c) atomicAdd(&m.v[i / T].tile_counts[i % T], c);
        }
    }
}

cudaError_t launch_preprocess_fwd_multi(const MultiFwdArgs& m_in, cudaStream_t s) {
    MultiFwdArgs m = m_in;
    const bool stage = m.v[0].D == 3 && m.v[0].M == 16;
    const size_t stage_bytes = stage ? (size_t)PRE_THREADS * SH_ROW_PAD_F4 * 16 : 0;
    const size_t hist_bytes = (size_t)m.T * m.n_views * 4;
    m.smem_hist = hist_bytes <= 96 * 1024 ? 1 : 0;
    for (int v = 0; v < m.n_views; v++) m.v[v].smem_hist = m.smem_hist;
    const size_t smem = stage_bytes + (m.smem_hist ? hist_bytes : 0);
    const int deg = stage ? 3 : m.v[0].D;
    const void* fn = deg == 0   ? (const void*)preprocess_fwd_multi_kernel<0>
                     : deg == 1 ? (const void*)preprocess_fwd_multi_kernel<1>
                     : deg == 2 ? (const void*)preprocess_fwd_multi_kernel<2>
                                : (const void*)preprocess_fwd_multi_kernel<3>;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int blocks = (m.v[0].P + PRE_THREADS - 1) / PRE_THREADS;
    switch (deg) {
        case 0: preprocess_fwd_multi_kernel<0><<<blocks, PRE_THREADS, smem, s>>>(m); break;
        case 1: preprocess_fwd_multi_kernel<1><<<blocks, PRE_THREADS, smem, s>>>(m); break;
        case 2: preprocess_fwd_multi_kernel<2><<<blocks, PRE_THREADS, smem, s>>>(m); break;
        default: preprocess_fwd_multi_kernel<3><<<blocks, PRE_THREADS, smem, s>>>(m); break;
    }
    return cudaGetLastError();
}

constexpr int SMEM_HIST_MAX_TILES = 8192;

cudaError_t launch_preprocess_fwd(const PreFwdArgs& a_in, bool stage_sh, cudaStream_t s) {
    PreFwdArgs a = a_in;
    const int T = (int)(a.gx * a.gy);
    a.smem_hist = (T <= SMEM_HIST_MAX_TILES) ? 1 : 0;
    const size_t hist_bytes = a.smem_hist ? (size_t)T * 4 : 0;
    const int blocks = (a.P + PRE_THREADS - 1) / PRE_THREADS;
    stage_sh = stage_sh && a.shs != nullptr && a.colors_precomp == nullptr && a.D == 3 && a.M == 16 &&
               ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
    if (stage_sh) {
        size_t smem = (size_t)PRE_THREADS * SH_ROW_PAD_F4 * 16 + hist_bytes;
        cudaError_t e =
            cudaFuncSetAttribute(preprocess_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        preprocess_fwd_kernel<true><<<blocks, PRE_THREADS, smem, s>>>(a);
    } else {
        preprocess_fwd_kernel<false><<<blocks, PRE_THREADS, hist_bytes, s>>>(a);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// tile_scan: one block. ranges[t] = (start, end) for non-empty tiles and (0,0) for empty ones, exactly what the
// reference's memset + identifyTileRanges leave behind (rasterizer_impl.cu:310-318).
// ---------------------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 1024;
__global__ void __launch_bounds__(SCAN_THREADS) tile_scan_kernel(const __grid_constant__ BatchArgs b) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t warp_max[32];
    const ViewBin& V = b.v[blockIdx.x];  // one block per view
    const uint32_t* __restrict__ counts = V.tile_counts;
    uint2* __restrict__ ranges = V.ranges;
    uint32_t* __restrict__ cursors = V.cursors;
    uint32_t* __restrict__ totals = V.totals;
    const int T = b.T;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (T + SCAN_THREADS - 1) / SCAN_THREADS;
    const int beg = tid * per, end = min(T, beg + per);
    uint32_t sum = 0, mx = 0;
    for (int i = beg; i < end; i++) {
        uint32_t c = counts[i];
        sum += c;
        mx = max(mx, c);
    }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 31) warp_sum[warp] = incl;
    if (lane == 0) warp_max[warp] = mx;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warp_sum[lane];
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        warp_sum[lane] = wi - w;  // exclusive
        uint32_t m = warp_max[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 31) totals[0] = wi;
        if (lane == 0) {
            totals[1] = m;
            totals[2] = 0;
            totals[3] = 0;
        }
    }
    __syncthreads();
    uint32_t run = warp_sum[warp] + incl - sum;
    for (int i = beg; i < end; i++) {
        uint32_t c = counts[i];
        ranges[i] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
        cursors[i] = run;
        run += c;
    }
}

cudaError_t launch_tile_scan(const BatchArgs& b, cudaStream_t s) {
    tile_scan_kernel<<<b.n, SCAN_THREADS, 0, s>>>(b);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// scatter: (Gaussian, tile) pairs into tile buckets. Entry = depth_bits << 32 | gaussian index. The slot order
// inside a bucket is arbitrary (atomic cursor); the tile sort makes the final order deterministic.
// ---------------------------------------------------------------------------------------------------------------
// SMEM: the block first counts its entries per tile in shared memory, reserves one contiguous run per (block, tile)
// with a single global atomic, then hands out slots with shared-memory atomics. Consecutive Gaussians of a pointmap
// land in the same few tiles, so this replaces ~R same-address global atomics by a handful per block.
template <bool SMEM>
__global__ void __launch_bounds__(256) scatter_kernel(const __grid_constant__ BatchArgs b) {
    extern __shared__ uint32_t s_u32[];
    const ViewBin& V = b.v[blockIdx.y];
    const int P = b.P;
    const unsigned gx = b.gx, gy = b.gy;
    const GaussRec* __restrict__ rec = V.rec;
    const int* __restrict__ radii = V.radii;
    const uint32_t* __restrict__ tiles_touched = V.tiles_touched;
    uint32_t* __restrict__ cursors = V.cursors;
    uint64_t* __restrict__ unsorted = V.unsorted;
    const int T = (int)(gx * gy);
    uint32_t* s_cnt = s_u32;
    uint32_t* s_base = s_u32 + T;
    const int tid = threadIdx.x;
    const int idx = blockIdx.x * blockDim.x + tid;
    if (SMEM) {
        for (int i = tid; i < T; i += 256) s_cnt[i] = 0;
        __syncthreads();
    }
    uint2 rmin = {0, 0}, rmax = {0, 0};
    uint64_t entry = 0;
    if (idx < P && tiles_touched[idx] != 0) {
        float4 A = rec[idx].A;
        float depth = rec[idx].C.w;
        tile_rect(make_float2(A.x, A.y), radii[idx], rmin, rmax, gx, gy);
        entry = ((uint64_t)__float_as_uint(depth) << 32) | (uint32_t)idx;
    }
    if (!SMEM) {
        for (unsigned y = rmin.y; y < rmax.y; y++)
            for (unsigned x = rmin.x; x < rmax.x; x++) {
                uint32_t slot = atomicAdd(&cursors[y * gx + x], 1u);
                unsorted[slot] = entry;
            }
        return;
    }
    for (unsigned y = rmin.y; y < rmax.y; y++)
        for (unsigned x = rmin.x; x < rmax.x; x++) atomicAdd(&s_cnt[y * gx + x], 1u);
    __syncthreads();
    for (int i = tid; i < T; i += 256) {
        uint32_t c = s_cnt[i];
        if (c) s_base[i] = atomicAdd(&cursors[i], c);
        s_cnt[i] = 0;
    }
    __syncthreads();
    for (unsigned y = rmin.y; y < rmax.y; y++)
        for (unsigned x = rmin.x; x < rmax.x; x++) {
            const unsigned t = y * gx + x;
            uint32_t slot = s_base[t] + atomicAdd(&s_cnt[t], 1u);
            unsorted[slot] = entry;
        }
}

cudaError_t launch_scatter(const BatchArgs& b, cudaStream_t s) {
    const size_t smem = (size_t)b.T * 8;
    const dim3 grid((b.P + 255) / 256, b.n);
    if (smem <= 48 * 1024)
        scatter_kernel<true><<<grid, 256, smem, s>>>(b);
    else
        scatter_kernel<false><<<grid, 256, 0, s>>>(b);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// tile_sort: one block of 512 threads per tile, bitonic network over N = 512*E keys held E per thread (blocked):
// strides < E in registers, E..16E by warp shuffles, >= 32E through shared memory.
// ---------------------------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 512;
constexpr int SORT_MAX_E = 32;

__device__ __forceinline__ void cmpx(uint64_t& a, uint64_t& b, bool up) {
    bool sw = (a > b) == up;
    uint64_t lo = sw ? b : a, hi = sw ? a : b;
    a = lo;
    b = hi;
}

template <int E>
__device__ __forceinline__ void bitonic_sort_tile(uint64_t* sm, const uint64_t* __restrict__ in, int n) {
    constexpr int N = SORT_THREADS * E;
    const int t = threadIdx.x;
    uint64_t v[E];
    // striped (coalesced) global read into shared, then blocked read-out
    for (int i = t; i < N; i += SORT_THREADS) sm[i] = (i < n) ? in[i] : ~0ull;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < E; r++) v[r] = sm[t * E + r];
    __syncthreads();

#pragma unroll 1
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll 1
        for (int j = k >> 1; j >= 32 * E; j >>= 1) {
            // cross-warp stride: exchange through shared memory, layout [r][t] (conflict-free both ways)
            const int pt = t ^ (j / E);
#pragma unroll
            for (int r = 0; r < E; r++) sm[r * SORT_THREADS + t] = v[r];
            __syncthreads();
            const bool up = (((t * E) & k) == 0);
            const bool keep_min = (((t * E) & j) == 0) == up;
#pragma unroll
            for (int r = 0; r < E; r++) {
                uint64_t o = sm[r * SORT_THREADS + pt];
                v[r] = keep_min ? min(v[r], o) : max(v[r], o);
            }
            __syncthreads();
        }
        {
            int j = min(k >> 1, 16 * E);
#pragma unroll 1
            for (; j >= E; j >>= 1) {
                const int mask = j / E;
                const bool up = (((t * E) & k) == 0) || (k == N);  // (i & N) is always 0 for i < N
                const bool keep_min = (((t * E) & j) == 0) == up;
#pragma unroll
                for (int r = 0; r < E; r++) {
                    uint64_t o = __shfl_xor_sync(0xffffffffu, v[r], mask);
                    v[r] = keep_min ? min(v[r], o) : max(v[r], o);
                }
            }
        }
        if (E > 1) {
#pragma unroll
            for (int j = E >> 1; j >= 1; j >>= 1) {
                if (j < k) {
#pragma unroll
                    for (int r = 0; r < E; r++) {
                        if ((r & j) == 0) {
                            const int i = t * E + r;
                            const bool up = ((i & k) == 0);
                            cmpx(v[r], v[r | j], up);
                        }
                    }
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < E; r++) sm[t * E + r] = v[r];
    __syncthreads();
}

// Per-tile sort, one block of 512 threads per tile. The keys of a tile are (depth bits << 32 | Gaussian index), all
// distinct, and the required order is plain ascending - which is exactly the order of the reference's stable LSD radix
// sort on tile|depth keys (ties in depth keep emission = index order). The common path is a most-significant-digit
// bucket sort entirely in shared memory:
//   1. every thread keeps its (striped, coalesced) keys in registers; a block-wide AND/OR finds the depth bits that
//      actually vary inside the tile, and the top LOG_NB of them become the digit (monotone in depth);
//   2. histogram, exclusive scan and scatter into 2^LOG_NB buckets with shared-memory atomics (order inside a bucket
//      is arbitrary at this point);
//   3. each bucket (a handful of keys) is put in order by one thread with an insertion sort on the FULL 64-bit key,
//      so equal depths come out in index order without any extra pass.
// If the depth distribution is so skewed that some bucket holds more than LEAF_MAX keys (e.g. thousands of Gaussians
// at exactly the same depth), the block falls back to the bitonic network over the whole tile.
// MAXE bounds the per-thread key count of the instantiation (registers / shared memory); MIN_N..512*MAXE is the size
// class this launch handles, blocks of other classes exit at once.
#ifndef DAS3R_SORT16_LOGNB
#define DAS3R_SORT16_LOGNB 12
#endif
constexpr int LEAF_MAX = 128;
constexpr int MED_CAP = 2048;  // buckets with more than 8 keys: at most 16384 / 9 of them

template <int MAXE, int MIN_N, int LOG_NB>
__global__ void __launch_bounds__(SORT_THREADS, (MAXE <= 8 ? 2 : 1))
    tile_sort_kernel(const __grid_constant__ BatchArgs b) {
    const ViewBin& V = b.v[blockIdx.y];
    const uint2* __restrict__ ranges = V.ranges;
    const uint64_t* __restrict__ unsorted = V.unsorted;
    uint32_t* __restrict__ point_list = V.point_list;
    uint64_t* __restrict__ keys_sorted = V.keys_sorted;
    constexpr int NB = 1 << LOG_NB;
    constexpr int PER = NB / SORT_THREADS;  // bucket counters per thread in the scan
    static_assert(NB % SORT_THREADS == 0 && PER >= 1, "bucket count must be a multiple of the block size");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* sm = reinterpret_cast<uint64_t*>(smem_raw);                                   // [512*MAXE] keys
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_raw + (size_t)SORT_THREADS * MAXE * 8);  // [NB] bucket sizes
    uint32_t* s_off = s_cnt + NB;                                                           // [NB] bucket starts
    uint32_t* s_cur = s_off + NB;                                                           // [NB] scatter cursors
    uint16_t* s_med = reinterpret_cast<uint16_t*>(s_cur + NB);                              // [MED_CAP] buckets > 8 keys
    __shared__ uint32_t s_red[2][SORT_THREADS / 32];
    __shared__ uint32_t s_warp_tot[SORT_THREADS / 32];
    __shared__ uint32_t s_flags[2];
    const uint32_t tile = blockIdx.x;
    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    if (n <= MIN_N || n > SORT_THREADS * MAXE) return;  // empty, or another launch's size class
    const uint64_t* in = unsorted + range.x;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

    // ---- 1. keys into registers, varying depth bits ----
    uint64_t k[MAXE];
    uint32_t v_and = 0xffffffffu, v_or = 0u;
#pragma unroll
    for (int e = 0; e < MAXE; e++) {
        const int i = t + e * SORT_THREADS;
        if (i < n) {
            k[e] = in[i];
            const uint32_t d = (uint32_t)(k[e] >> 32);
            v_and &= d;
            v_or |= d;
        }
    }
    for (int i = t; i < NB; i += SORT_THREADS) s_cnt[i] = 0;
    v_and = __reduce_and_sync(0xffffffffu, v_and);
    v_or = __reduce_or_sync(0xffffffffu, v_or);
    if (lane == 0) {
        s_red[0][warp] = v_and;
        s_red[1][warp] = v_or;
    }
    if (t == 0) {
        s_flags[0] = 0;
        s_flags[1] = 0;
    }
    __syncthreads();
    v_and = 0xffffffffu;
    v_or = 0u;
#pragma unroll
    for (int w = 0; w < SORT_THREADS / 32; w++) {
        v_and &= s_red[0][w];
        v_or |= s_red[1][w];
    }
    const uint32_t diff = v_and ^ v_or;
    const int top = diff ? (31 - __clz(diff)) : 0;
    const int shift = max(0, top - (LOG_NB - 1));

    // ---- 2. histogram ----
#pragma unroll
    for (int e = 0; e < MAXE; e++) {
        const int i = t + e * SORT_THREADS;
        if (i < n) atomicAdd(&s_cnt[((uint32_t)(k[e] >> 32) >> shift) & (NB - 1)], 1u);
    }
    __syncthreads();
    // exclusive scan of the NB counts (PER consecutive ones per thread) + largest bucket + list of medium buckets
    {
        uint32_t c[PER];
        uint32_t sum = 0, cmax = 0;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            c[j] = s_cnt[PER * t + j];
            sum += c[j];
            cmax = max(cmax, c[j]);
            if (c[j] > 8u) {
                const uint32_t slot = atomicAdd(&s_flags[1], 1u);
                if (slot < (uint32_t)MED_CAP) s_med[slot] = (uint16_t)(PER * t + j);
            }
        }
        if (cmax > (uint32_t)LEAF_MAX) atomicOr(&s_flags[0], 1u);
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp_tot[warp] = incl;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; w++) wbase += (w < warp) ? s_warp_tot[w] : 0u;
        uint32_t ex = wbase + incl - sum;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            s_off[PER * t + j] = ex;
            s_cur[PER * t + j] = ex;
            ex += c[j];
        }
    }
    __syncthreads();

    if (s_flags[0] != 0) {
        // skewed depths: bitonic network over the whole tile (reads the keys again from global memory)
        if (n <= SORT_THREADS * 1) bitonic_sort_tile<1>(sm, in, n);
        else if (MAXE >= 2 && n <= SORT_THREADS * 2) bitonic_sort_tile<(MAXE >= 2 ? 2 : 1)>(sm, in, n);
        else if (MAXE >= 4 && n <= SORT_THREADS * 4) bitonic_sort_tile<(MAXE >= 4 ? 4 : 1)>(sm, in, n);
        else if (MAXE >= 8 && n <= SORT_THREADS * 8) bitonic_sort_tile<(MAXE >= 8 ? 8 : 1)>(sm, in, n);
        else if (MAXE >= 16 && n <= SORT_THREADS * 16) bitonic_sort_tile<(MAXE >= 16 ? 16 : 1)>(sm, in, n);
        else bitonic_sort_tile<(MAXE >= 32 ? 32 : 1)>(sm, in, n);
    } else {
        // ---- scatter into buckets ----
#pragma unroll
        for (int e = 0; e < MAXE; e++) {
            const int i = t + e * SORT_THREADS;
            if (i < n) {
                const uint32_t pos = atomicAdd(&s_cur[((uint32_t)(k[e] >> 32) >> shift) & (NB - 1)], 1u);
                sm[pos] = k[e];
            }
        }
        __syncthreads();
        // ---- 3. leaf: every bucket is put in order on the FULL key ----
        // (a) small buckets (<= 8 keys): one thread each, keys in registers, fixed compare-exchange networks sized to
        //     the bucket (2: one exchange, 3..4: the 5-exchange network, 5..8: odd-even transposition)
        for (int b = t; b < NB; b += SORT_THREADS) {
            const uint32_t s0 = s_off[b], m = s_cnt[b];
            if (m == 2) {
                const uint64_t x = sm[s0], y = sm[s0 + 1];
                sm[s0] = min(x, y);
                sm[s0 + 1] = max(x, y);
            } else if (m > 2 && m <= 4) {
                uint64_t v0 = sm[s0], v1 = sm[s0 + 1], v2 = sm[s0 + 2], v3 = (m == 4) ? sm[s0 + 3] : ~0ull;
                cmpx(v0, v1, true);
                cmpx(v2, v3, true);
                cmpx(v0, v2, true);
                cmpx(v1, v3, true);
                cmpx(v1, v2, true);
                sm[s0] = v0;
                sm[s0 + 1] = v1;
                sm[s0 + 2] = v2;
                if (m == 4) sm[s0 + 3] = v3;
            } else if (m > 4 && m <= 8) {
                uint64_t v[8];
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = (j < (int)m) ? sm[s0 + j] : ~0ull;
#pragma unroll
                for (int round = 0; round < 8; round++) {
#pragma unroll
                    for (int j = (round & 1); j + 1 < 8; j += 2) {
                        const uint64_t lo = min(v[j], v[j + 1]), hi = max(v[j], v[j + 1]);
                        v[j] = lo;
                        v[j + 1] = hi;
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (j < (int)m) sm[s0 + j] = v[j];
            }
        }
        // (b) medium buckets (9..128 keys): one warp each, rank sort (a key's final slot = number of smaller keys)
        const int n_med = (int)min(s_flags[1], (uint32_t)MED_CAP);
        for (int q = warp; q < n_med; q += SORT_THREADS / 32) {
            const int b = s_med[q];
            const uint32_t s0 = s_off[b], m = s_cnt[b];
            if (m > 8 && m <= 32) {
                // the common medium bucket: one key per lane, one compare per key (ncu: with the four-slot loop below
                // for every bucket, half of the kernel's instructions were these compares)
                const uint64_t mine = ((uint32_t)lane < m) ? sm[s0 + lane] : ~0ull;
                uint32_t rank = 0;
                for (uint32_t j = 0; j < m; j++) rank += (sm[s0 + j] < mine) ? 1u : 0u;  // broadcast read
                __syncwarp();
                if ((uint32_t)lane < m) sm[s0 + rank] = mine;
            } else if (m > 32) {  // m <= LEAF_MAX is guaranteed by the flag test above
                uint64_t mine[LEAF_MAX / 32];
                uint32_t rank[LEAF_MAX / 32];
#pragma unroll
                for (int r = 0; r < LEAF_MAX / 32; r++) {
                    const uint32_t j = (uint32_t)lane + 32u * r;
                    mine[r] = (j < m) ? sm[s0 + j] : ~0ull;
                    rank[r] = 0;
                }
                for (uint32_t j = 0; j < m; j++) {
                    const uint64_t o = sm[s0 + j];  // broadcast read
#pragma unroll
                    for (int r = 0; r < LEAF_MAX / 32; r++) rank[r] += (o < mine[r]) ? 1u : 0u;
                }
                __syncwarp();
#pragma unroll
                for (int r = 0; r < LEAF_MAX / 32; r++)
                    if ((uint32_t)lane + 32u * r < m) sm[s0 + rank[r]] = mine[r];
            }
        }
        __syncthreads();
    }

    // The sorted keys leave as point_list (+ the debug key array); the tile-ordered record streams are produced by a
    // separate, fully parallel gather kernel so that no tile serialises long-latency gathers behind its sort.
    for (int i = threadIdx.x; i < n; i += SORT_THREADS) {
        const uint64_t key = sm[i];
        point_list[range.x + i] = (uint32_t)key;
        if (keys_sorted) keys_sorted[range.x + i] = ((uint64_t)tile << 32) | (key >> 32);
    }
}

int tile_sort_capacity() { return SORT_THREADS * SORT_MAX_E; }

template <int MAXE, int MIN_N, int LOG_NB>
static cudaError_t launch_tile_sort_variant(const BatchArgs& b, cudaStream_t s) {
    const size_t smem = (size_t)SORT_THREADS * MAXE * sizeof(uint64_t) + 3 * ((size_t)1 << LOG_NB) * sizeof(uint32_t) +
                        MED_CAP * sizeof(uint16_t);
    if (smem + 2048 > 48 * 1024) {  // + the kernel's static shared memory
        cudaError_t err = cudaFuncSetAttribute(tile_sort_kernel<MAXE, MIN_N, LOG_NB>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
    }
    tile_sort_kernel<MAXE, MIN_N, LOG_NB><<<dim3(b.T, b.n), SORT_THREADS, smem, s>>>(b);
    return cudaGetLastError();
}

// Two auxiliary streams per device for the crowded size classes of the tile sort (created on first use, never
// destroyed; cudaStreamNonBlocking, ordered against the caller's stream with per-call events).
namespace {
constexpr int kMaxDevices = 64;
struct SideStreams {
    cudaStream_t st[2] = {nullptr, nullptr};
    bool ready = false;
};
cudaError_t side_streams(cudaStream_t out[2]) {
    static std::mutex mu;
    static SideStreams per_dev[kMaxDevices];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    SideStreams& ss = per_dev[dev];
    if (!ss.ready) {
        for (int i = 0; i < 2; i++) {
            e = cudaStreamCreateWithFlags(&ss.st[i], cudaStreamNonBlocking);
            if (e != cudaSuccess) return e;
        }
        ss.ready = true;
    }
    out[0] = ss.st[0];
    out[1] = ss.st[1];
    return cudaSuccess;
}
}  // namespace

static cudaError_t sort_oversized_tiles(int R, int T, const uint2* ranges, uint64_t* unsorted, uint32_t* point_list,
                                        uint64_t* keys_sorted, void* temp, size_t temp_bytes, cudaStream_t s);

cudaError_t launch_tile_sort(const BatchArgs& b, uint32_t max_count, const int* R, const uint32_t* view_max,
                             void* const* temp, const size_t* temp_bytes, cudaStream_t s) {
    // Tiles are sorted by the cheapest variant that fits them (<= 4096, <= 8192, <= 16384 entries in shared memory;
    // beyond that a library segmented radix sort restricted to those tiles): every launch covers all tiles of all views
    // and a block whose tile belongs to another size class exits at once. The crowded classes (one block per SM, a few
    // dozen tiles) run on auxiliary streams next to the common one instead of in front of it.
    const bool need16 = max_count > SORT_THREADS * 8, need32 = max_count > SORT_THREADS * 16;
    const bool need_lib = max_count > (uint32_t)tile_sort_capacity();
    if (need_lib)
        for (int v = 0; v < b.n; v++)
            if (view_max[v] > (uint32_t)tile_sort_capacity() &&
                (temp[v] == nullptr || temp_bytes[v] < tile_sort_fallback_bytes(R[v], b.T)))
                return cudaErrorInvalidValue;
    cudaError_t e = cudaSuccess;
    cudaStream_t side[2] = {nullptr, nullptr};
    cudaEvent_t fork = nullptr, join[2] = {nullptr, nullptr};
    if (need16) {
        if ((e = side_streams(side)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&fork, cudaEventDisableTiming)) != cudaSuccess) return e;
        e = cudaEventRecord(fork, s);
        for (int c = 0; c < 2 && e == cudaSuccess; c++) {
            if (c == 1 && !need32) break;
            e = cudaStreamWaitEvent(side[c], fork, 0);
            if (e != cudaSuccess) break;
            e = (c == 0) ? launch_tile_sort_variant<16, SORT_THREADS * 8, DAS3R_SORT16_LOGNB>(b, side[c])
                         : launch_tile_sort_variant<32, SORT_THREADS * 16, 12>(b, side[c]);
            if (e != cudaSuccess) break;
            if (c == 1 && need_lib) {
                for (int v = 0; v < b.n && e == cudaSuccess; v++)
                    if (view_max[v] > (uint32_t)tile_sort_capacity())
                        e = sort_oversized_tiles(R[v], b.T, b.v[v].ranges, b.v[v].unsorted, b.v[v].point_list,
                                                 b.v[v].keys_sorted, temp[v], temp_bytes[v], side[c]);
                if (e != cudaSuccess) break;
            }
            if ((e = cudaEventCreateWithFlags(&join[c], cudaEventDisableTiming)) != cudaSuccess) break;
            e = cudaEventRecord(join[c], side[c]);
        }
    }
    if (e == cudaSuccess) e = launch_tile_sort_variant<8, 0, 10>(b, s);
    for (int c = 0; c < 2; c++)
        if (join[c]) {
            if (e == cudaSuccess) e = cudaStreamWaitEvent(s, join[c], 0);
            cudaEventDestroy(join[c]);  // released once the recorded work has completed
        }
    if (fork) cudaEventDestroy(fork);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();  // the tile-ordered record streams are written by rect_lists_kernel
}

// Tiles beyond the shared-memory capacity: library segmented radix sort of their (depth|index) entries (segment =
// tile; the other tiles are passed as empty segments), then the sorted entries become point_list / debug keys. The
// record streams are written by the common gather afterwards.
__global__ void oversized_to_list_kernel(int T, uint32_t capacity, const uint2* __restrict__ ranges,
                                         const uint64_t* __restrict__ sorted_entries,
                                         uint32_t* __restrict__ point_list, uint64_t* __restrict__ keys_sorted) {
    const uint32_t tile = blockIdx.x;  // one block per tile so the tile id is known without a search
    const uint2 range = ranges[tile];
    if (range.y - range.x <= capacity) return;
    for (uint32_t i = range.x + threadIdx.x; i < range.y; i += blockDim.x) {
        const uint64_t key = sorted_entries[i];
        point_list[i] = (uint32_t)key;
        if (keys_sorted) keys_sorted[i] = ((uint64_t)tile << 32) | (key >> 32);
    }
}

__global__ void oversized_offsets_kernel(int T, uint32_t capacity, const uint2* __restrict__ ranges,
                                         uint32_t* __restrict__ beg, uint32_t* __restrict__ end) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < T) {
        const uint2 r = ranges[i];
        const bool big = r.y - r.x > capacity;
        beg[i] = big ? r.x : 0u;
        end[i] = big ? r.y : 0u;
    }
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static size_t spread_pad_smem_uncached(const void* kernel, int grid, int block_threads, size_t static_smem);

size_t spread_pad_smem(const void* kernel, int grid, int block_threads, size_t static_smem) {
    // the answer only depends on (device, kernel, grid): memoise it, the queries cost several microseconds per launch
    struct Key {
        int dev;
        const void* k;
        int grid;
        bool operator<(const Key& o) const {
            return dev != o.dev ? dev < o.dev : (k != o.k ? k < o.k : grid < o.grid);
        }
    };
    static std::mutex mu;
    static std::map<Key, size_t> cache;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    std::lock_guard<std::mutex> lock(mu);
    Key key{dev, kernel, grid};
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    size_t v = spread_pad_smem_uncached(kernel, grid, block_threads, static_smem);
    cache[key] = v;
    return v;
}

static size_t spread_pad_smem_uncached(const void* kernel, int grid, int block_threads, size_t static_smem) {
    int dev = 0, sms = 0, smem_sm = 0, smem_optin = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (sms <= 0 || smem_sm <= 0) return 0;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, block_threads, 0) != cudaSuccess || occ <= 0)
        return 0;
    const int want = (grid + sms - 1) / sms;  // blocks per SM of an even spread
    if (want >= occ) return 0;                // more than one wave, or already exactly full
    // each block also pays 1 KB of driver-reserved shared memory
    size_t per_block = (size_t)smem_sm / (size_t)want;
    if (per_block <= static_smem + 1024) return 0;
    size_t dyn = per_block - static_smem - 1024;
    dyn = dyn / 256 * 256;
    if (dyn + static_smem > (size_t)smem_optin) dyn = (size_t)smem_optin - static_smem;
    // The attribute belongs to the kernel, not to this grid size: always raise it to the device limit, never to `dyn`
    // (a later, smaller request would otherwise lower it under a cached larger one).
    if (dyn > 48 * 1024 - static_smem)
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)((size_t)smem_optin - static_smem)) != cudaSuccess)
            return 0;
    return dyn;
}

size_t tile_sort_fallback_bytes(int R, int T) {
    size_t cub_bytes = 0;
    cub::DeviceSegmentedRadixSort::SortKeys((void*)nullptr, cub_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, R,
                                            T, (const uint32_t*)nullptr, (const uint32_t*)nullptr);
    return align_up((size_t)R * 8, 256) + 2 * align_up((size_t)T * 4, 256) + align_up(cub_bytes, 256);
}

static cudaError_t sort_oversized_tiles(int R, int T, const uint2* ranges, uint64_t* unsorted, uint32_t* point_list,
                                        uint64_t* keys_sorted, void* temp, size_t temp_bytes, cudaStream_t s) {
    char* p = static_cast<char*>(temp);
    uint64_t* sorted_entries = reinterpret_cast<uint64_t*>(p);
    p += align_up((size_t)R * 8, 256);
    uint32_t* beg = reinterpret_cast<uint32_t*>(p);
    p += align_up((size_t)T * 4, 256);
    uint32_t* end = reinterpret_cast<uint32_t*>(p);
    p += align_up((size_t)T * 4, 256);
    size_t cub_bytes = temp_bytes - (size_t)(p - static_cast<char*>(temp));
    const uint32_t cap = (uint32_t)tile_sort_capacity();
    oversized_offsets_kernel<<<(T + 255) / 256, 256, 0, s>>>(T, cap, ranges, beg, end);
    cudaError_t e = cub::DeviceSegmentedRadixSort::SortKeys(p, cub_bytes, unsorted, sorted_entries, R, T, beg, end, 0,
                                                            64, s);
    if (e != cudaSuccess) return e;
    oversized_to_list_kernel<<<T, 256, 0, s>>>(T, cap, ranges, sorted_entries, point_list, keys_sorted);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// rect_lists: for every tile and each of its eight 8x4 pixel rectangles, the ordered list of tile-list positions of
// the Gaussians whose conservative radius reaches the rectangle. One block per tile, warp w builds rectangle w's list
// by ballot compaction in a single pass (four chunks of 32 entries in flight per round); rectangle w of a tile with n
// entries owns the n slots starting at 8*range.x + w*n of the output array, so no counting pass or atomics are needed.
// ---------------------------------------------------------------------------------------------------------------
#ifndef DAS3R_RECT_MIN_BLOCKS
#define DAS3R_RECT_MIN_BLOCKS 4
#endif
constexpr int RECT_CHUNK = 1024;  // tile-list entries staged in shared memory at a time (4 per thread)

// The block stages a chunk of its tile's list in shared memory - Gaussian index and the 16-byte record head
// {x, y, h, -tau'}, fetched through the sorted point list - and its eight warps scan the chunk for their rectangles. No
// tile-ordered copy of the records is written any more: the lists carry the Gaussian index and the blends fetch the
// 48-byte records from the per-Gaussian array themselves (a gather either way, but 52 bytes per instance less to write
// and to read back; the separate gather kernel took as long as the whole sort).
__global__ void __launch_bounds__(256, DAS3R_RECT_MIN_BLOCKS) rect_lists_kernel(const __grid_constant__ BatchArgs b) {
    __shared__ uint32_t cand_s[8][256];   // per warp: circular buffer of chunk positions that reach the rectangle
    __shared__ float4 sA_s[RECT_CHUNK];   // {x, y, h, -tau'} of the chunk's entries
    __shared__ uint32_t sG_s[RECT_CHUNK];  // their Gaussian indices
    const ViewBin& V = b.v[blockIdx.y];
    const uint2* __restrict__ ranges = V.ranges;
    const uint32_t* __restrict__ point_list = V.point_list;
    const GaussRec* __restrict__ rec = V.rec;
    uint32_t* __restrict__ rect_off = V.rect_off;
    uint32_t* __restrict__ rect_cnt = V.rect_cnt;
    uint2* __restrict__ rect_list = V.rect_list;
    uint32_t* __restrict__ rect_gid = V.rect_gid;
    const unsigned gx = b.gx;
    const int W = b.W, H = b.H;
    const int tile = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t* gL = point_list + range.x;
    uint32_t* cand = cand_s[w];
    const uint32_t off = 8u * range.x + (uint32_t)w * (uint32_t)n;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int rx0 = tx * TILE_X + (w & 1) * 8, ry0 = ty * TILE_Y + (w >> 1) * 4;
    // pixels of the rectangle that exist in the image
    uint32_t pix_valid = 0;
    for (int r = 0; r < 4; r++)
        for (int j = 0; j < 8; j++)
            if (rx0 + j < W && ry0 + r < H) pix_valid |= 1u << (r * 8 + j);

    uint32_t head = 0, tail = 0, run = 0;
    int chunk0 = 0;
    // Pixel masks of `k` (<= 32) buffered candidates of the current chunk, one per lane; entries with a non-empty mask
    // are appended to the rectangle's list with their position in the tile list and their Gaussian index.
    auto emit = [&](int k) {
        uint32_t pos = 0, mask = 0, gid = 0;
        if (lane < k) {
            pos = cand[(head + lane) & 255u];
            gid = sG_s[pos];
            const float4 A = sA_s[pos];
            const float4 con_o = __ldg(reinterpret_cast<const float4*>(rec + gid) + 1);
            // splat_power() for the 32 pixels with the column / row factors hoisted (same operations, same rounding)
            float dx[8], ax[8], bx[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                dx[j] = A.x - (float)(rx0 + j);
                ax[j] = __fmul_rn(con_o.x, dx[j]);
                bx[j] = __fmul_rn(con_o.y, dx[j]);
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const float dy = A.y - (float)(ry0 + r);
                const float zz = __fmul_rn(__fmul_rn(con_o.z, dy), dy);
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float sq = __fmaf_rn(ax[j], dx[j], zz);
                    const float u = __fmul_rn(bx[j], dy);
                    const float power = __fmaf_rn(sq, -0.5f, -u);
                    if (!(power > 0.0f) && power >= A.w) mask |= 1u << (r * 8 + j);
                }
            }
            mask &= pix_valid;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, mask != 0u);
        if (mask) {
            const uint32_t slot = off + run + __popc(bal & lt_mask);
            rect_list[slot] = make_uint2((uint32_t)chunk0 + pos, mask);
            rect_gid[slot] = gid;
        }
        run += __popc(bal);
        head += (uint32_t)k;
    };

    for (chunk0 = 0; chunk0 < n; chunk0 += RECT_CHUNK) {
        const int cn = min(RECT_CHUNK, n - chunk0);
        // ---- stage: 4 entries per thread, all loads in flight before the first store ----
        {
            constexpr int PER = RECT_CHUNK / 256;
            uint32_t g[PER];
            float4 rA[PER];
#pragma unroll
            for (int u = 0; u < PER; u++) {
                const int i = threadIdx.x + u * 256;
                g[u] = (i < cn) ? gL[chunk0 + i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < PER; u++) {
                const int i = threadIdx.x + u * 256;
                if (i < cn) rA[u] = __ldg(reinterpret_cast<const float4*>(rec + g[u]));
            }
#pragma unroll
            for (int u = 0; u < PER; u++) {
                const int i = threadIdx.x + u * 256;
                if (i < cn) {
                    sA_s[i] = rA[u];
                    sG_s[i] = g[u];
                }
            }
        }
        __syncthreads();
        // ---- scan: every warp compacts the entries that reach its rectangle and evaluates their pixel masks ----
        for (int c0 = 0; c0 < cn; c0 += 128) {
            float4 a[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int e = c0 + 32 * u + lane;
                a[u] = (e < cn) ? sA_s[e] : make_float4(0.f, 0.f, -1.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool ov = reaches_rect(a[u], tx, ty, w);
                const unsigned m = __ballot_sync(0xffffffffu, ov);
                if (ov) cand[(tail + __popc(m & lt_mask)) & 255u] = (uint32_t)(c0 + 32 * u + lane);
                tail += __popc(m);
            }
            __syncwarp();
            while (tail - head >= 32u) emit(32);
            __syncwarp();
        }
        // candidates index the CURRENT chunk's shared-memory copy: drain them before it is replaced
        while (tail != head) emit((int)min(32u, tail - head));
        __syncthreads();
    }
    if (lane == 0) {
        rect_off[tile * 8 + w] = off;
        rect_cnt[tile * 8 + w] = run;
    }
}

cudaError_t launch_rect_lists(const BatchArgs& b, cudaStream_t s) {
    rect_lists_kernel<<<dim3(b.T, b.n), 256, 0, s>>>(b);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// blend_fwd (replaces renderCUDA, forward.cu:261-374): see blend_common.cuh for the walk. The blend step is exactly
// forward.cu:330-361 - exp, alpha clamp at 0.99, skip below 1/255, stop when T would drop under 1e-4, colour
// accumulation - with the reference's FMA contraction pinned, so the front-to-back recurrence per pixel, `final_T`,
// `n_contrib` (list positions) and the image are bit-identical; the early termination is per warp (ballot).
// ---------------------------------------------------------------------------------------------------------------
#ifdef DAS3R_BLEND_TIMING  // developer instrumentation (tools/variants.sh), never in the shipped library
__device__ unsigned long long g_blend_dbg[4 * 65536];
extern "C" int das3r_debug_blend_timing(unsigned long long* host_out, int n_words) {
    return (int)cudaMemcpyFromSymbol(host_out, g_blend_dbg, sizeof(unsigned long long) * n_words);
}
extern "C" int das3r_debug_walk_prof(unsigned long long* host_out, int n_words) {
    return (int)cudaMemcpyFromSymbol(host_out, g_walk_prof, sizeof(unsigned long long) * n_words);
}
__device__ __forceinline__ unsigned long long dbg_globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

struct FwdPair {
    float4 col;  // {r, g, b, depth (EXTRAS only)}
    float alpha;
    uint32_t pos1;  // tile-list position + 1
};

template <bool EXTRAS>
__global__ void __launch_bounds__(32 * BLEND_WARPS) blend_fwd_kernel(const __grid_constant__ BatchArgs b) {
    __shared__ __align__(16) BlendSmem<FWD_K> sm;
    const ViewBin& V = b.v[blockIdx.y];
    const uint2* __restrict__ ranges = V.ranges;
    const uint32_t* __restrict__ rect_off = V.rect_off;
    const uint32_t* __restrict__ rect_cnt = V.rect_cnt;
    const uint2* __restrict__ rect_list = V.rect_list;
    const uint32_t* __restrict__ rect_gid = V.rect_gid;
    const GaussRec* __restrict__ rec_by_id = V.rec;
    const float* __restrict__ bg = V.bg;
    float* __restrict__ final_T = V.final_T;
    uint32_t* __restrict__ n_contrib = V.n_contrib;
    float* __restrict__ out_color = V.out_color;
    float* __restrict__ out_depth = V.out_depth;
    float* __restrict__ out_alpha = V.out_alpha;
    const int W = b.W, H = b.H;
    const unsigned gx = b.gx;
    const int warp = threadIdx.x >> 5;
    const WarpRect wr = warp_rect(W, H, gx);
    const uint2 range = ranges[wr.tile];
    const int cnt = (range.y > range.x) ? (int)rect_cnt[wr.tile * 8 + wr.rect] : 0;
    const uint32_t off = cnt ? rect_off[wr.tile * 8 + wr.rect] : 0u;
    const float2 pixf = {(float)wr.px, (float)wr.py};
#ifdef DAS3R_BLEND_TIMING
    const unsigned long long dbg_t0 = dbg_globaltimer();
    const long long dbg_c0 = clock64();
#endif

    bool done = !wr.inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float Dz = 0.f;

    walk_rect_list<false, FWD_K, FWD_NPAIR>(
        sm.ring[warp], sm.qmask[warp], rect_list + off, rect_gid + off, cnt, rec_by_id, 0xffffffffu,
        [&]() { return !done; },
        [&](const RingRec& r) {
            FwdPair x;
            const float4 A = r.A;
            const float4 con_o = r.B;
            x.col = r.C;  // fetched with the rest: apply() then has no shared-memory latency in the pixel's chain
            if (EXTRAS) x.col.w = rec_by_id[__float_as_uint(x.col.w)].C.w;  // the record carries the index, not the depth
            x.pos1 = __float_as_uint(A.z) + 1u;
            const float power = splat_power(con_o, make_float2(A.x - pixf.x, A.y - pixf.y));
            // forward.cu:339-340 (power <= 0 is guaranteed by the pixel mask)
            x.alpha = min(0.99f, __fmul_rn(con_o.w, expf(power)));
            return x;
        },
        [&](const RingRec&, const FwdPair& x) {
            // forward.cu:341-361, written with selects instead of early returns (no divergent branches in the chain)
            const float alpha = x.alpha;
            const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
            const bool counts = !done && !(alpha < 1.0f / 255.0f);
            const bool stop = counts && (test_T < 0.0001f);
            const bool blend = counts && !stop;
            const float w = blend ? T : 0.f;  // C += T * (c * alpha) only where the pair contributes
            const float c0 = __fmaf_rn(w, __fmul_rn(x.col.x, alpha), C[0]);
            const float c1 = __fmaf_rn(w, __fmul_rn(x.col.y, alpha), C[1]);
            const float c2 = __fmaf_rn(w, __fmul_rn(x.col.z, alpha), C[2]);
            C[0] = blend ? c0 : C[0];
            C[1] = blend ? c1 : C[1];
            C[2] = blend ? c2 : C[2];
            if (EXTRAS) {
                const float dz = __fmaf_rn(w, __fmul_rn(x.col.w, alpha), Dz);
                Dz = blend ? dz : Dz;
            }
            T = blend ? test_T : T;
            last_contributor = blend ? x.pos1 : last_contributor;
            done = done || stop;
        });

#ifdef DAS3R_BLEND_TIMING
    {
        const unsigned wmaxc = __reduce_max_sync(0xffffffffu, last_contributor);
        if ((threadIdx.x & 31) == 0) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            unsigned long long* o = g_blend_dbg + 4 * (size_t)(blockIdx.x * BLEND_WARPS + warp);
            o[0] = dbg_t0;
            o[1] = dbg_globaltimer();
            o[2] = (unsigned long long)(clock64() - dbg_c0);
            o[3] = ((unsigned long long)smid << 48) | ((unsigned long long)wmaxc << 24) | (unsigned long long)cnt;
        }
    }
#endif
    if (wr.inside) {
        const int pix_id = W * wr.py + wr.px;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) out_color[ch * H * W + pix_id] = __fmaf_rn(T, bg[ch], C[ch]);
        if (EXTRAS) {
            if (out_depth) out_depth[pix_id] = Dz;
            if (out_alpha) out_alpha[pix_id] = 1.0f - T;
        }
    }
}

cudaError_t launch_blend_fwd(const BatchArgs& b, cudaStream_t s) {
    const unsigned per_view = b.gx * b.gy * (8 / BLEND_WARPS);
    bool extras = false;  // depth / alpha wanted by any view: all views take the EXTRAS kernel (it tests the pointers)
    for (int v = 0; v < b.n; v++) extras = extras || b.v[v].out_depth || b.v[v].out_alpha;
    const dim3 grid(per_view, b.n);
    if (extras) {
        const size_t pad = spread_pad_smem((const void*)blend_fwd_kernel<true>, (int)(per_view * b.n), 32 * BLEND_WARPS,
                                           sizeof(BlendSmem<FWD_K>));
        blend_fwd_kernel<true><<<grid, 32 * BLEND_WARPS, pad, s>>>(b);
    } else {
        size_t pad = spread_pad_smem((const void*)blend_fwd_kernel<false>, (int)(per_view * b.n), 32 * BLEND_WARPS,
                                     sizeof(BlendSmem<FWD_K>));
#ifdef DAS3R_BLEND_TIMING  // occupancy experiments
        if (const char* ex = getenv("DAS3R_BLEND_EXTRA_SMEM")) {
            pad = (size_t)atoi(ex);
            cudaFuncSetAttribute(blend_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad);
        }
#endif
        blend_fwd_kernel<false><<<grid, 32 * BLEND_WARPS, pad, s>>>(b);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// mark_visible (checkFrustum, rasterizer_impl.cu:54-66): present = view-space z > near.
// ---------------------------------------------------------------------------------------------------------------
__global__ void mark_visible_kernel(int P, const float* __restrict__ means, const float* __restrict__ view,
                                    const float* __restrict__ proj, uint8_t* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
    float3 p_view = xform_point_4x3(p, view);
    present[idx] = (p_view.z <= NEAR_Z) ? 0 : 1;
}

cudaError_t launch_mark_visible(int P, const float* means, const float* view, const float* proj, uint8_t* present,
                                cudaStream_t s) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means, view, proj, present);
    return cudaGetLastError();
}

}  // namespace das3r
