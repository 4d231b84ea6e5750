/*
 * raster_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, CPU restatement of the reference's algorithm for the hot path (DAS3R's vendored
 * diff-gaussian-rasterization + simple-knn), used only by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / `--impl reference-cpu` legs as the checker and as the timed CPU baseline. Nothing under
 * das3r_b200/ may import, link or call it; the product path has no CPU fallback.
 *
 * Parity pin: the reference ships no tests, golden vectors or fixtures for this path, so this port is
 * pinned against outputs of the reference CUDA code itself (oracle/_ref, built by oracle/build_ref.sh from
 * /root/reference and run on the GPU box; vectors committed under tests/golden/ with tests/golden/make_golden.py).
 * Integer outputs could still differ from the CUDA reference on a Gaussian that sits within an ulp of a rounding
 * threshold if a contraction was missed here; the golden tests state the tolerated count (0 on the committed vectors).
 *
 * Each function cites the reference lines it follows (paths relative to
 * /root/reference/submodules/diff-gaussian-rasterization/ unless noted).
 *
 * Build: gcc -O2 -fopenmp -mfma -ffp-contract=off -shared -fPIC raster_oracle.c -o _build/liboracle.so -lm
 * (contraction off: where the reference's sm_100 build fuses a multiply-add on the chain that decides the integer
 * outputs - transforms, covariance, determinant, radius - the fusion is spelled out with fmaf(); everything else is
 * evaluated unfused, which stays far inside the float tolerances)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define BLOCK_X 16 /* cuda_rasterizer/config.h:16-17 */
#define BLOCK_Y 16
#define NEAR_Z 0.001f /* DAS3R README.md:41-44 edit of cuda_rasterizer/auxiliary.h:154 */

static const float SH_C0 = 0.28209479177387814f; /* auxiliary.h:21-38 */
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                               0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f,  -0.5900435899266435f};

typedef struct { float x, y, z; } f3;
typedef struct { float m[3][3]; } mat3; /* m[col][row] like glm */

static mat3 mat3_make(float x0, float y0, float z0, float x1, float y1, float z1, float x2, float y2, float z2) {
    mat3 r = {{{x0, y0, z0}, {x1, y1, z1}, {x2, y2, z2}}};
    return r;
}
/* third_party/glm/glm/detail/type_mat3x3.inl:486-518 */
static mat3 mat3_mul(const mat3* A, const mat3* B) {
    mat3 R;
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++)
            /* a*b + c*d + e*f as nvcc contracts it: the middle product is rounded, the outer two are fused */
            R.m[c][r] = fmaf(A->m[2][r], B->m[c][2], fmaf(A->m[0][r], B->m[c][0], A->m[1][r] * B->m[c][1]));
    return R;
}
static mat3 mat3_T(const mat3* A) {
    mat3 R;
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++) R.m[c][r] = A->m[r][c];
    return R;
}
/* auxiliary.h:58-66 */
static float dot3p(float a0, float b0, float a1, float b1, float a2, float b2, float c) {
    /* a0*b0 + a1*b1 + a2*b2 + c as compiled for the reference: fma(a2,b2, fma(a0,b0, a1*b1)) + c */
    return fmaf(a2, b2, fmaf(a0, b0, a1 * b1)) + c;
}
static f3 xform43(f3 p, const float* m) {
    f3 t = {dot3p(m[0], p.x, m[4], p.y, m[8], p.z, m[12]), dot3p(m[1], p.x, m[5], p.y, m[9], p.z, m[13]),
            dot3p(m[2], p.x, m[6], p.y, m[10], p.z, m[14])};
    return t;
}
/* auxiliary.h:68-77 */
static void xform44(f3 p, const float* m, float o[4]) {
    o[0] = dot3p(m[0], p.x, m[4], p.y, m[8], p.z, m[12]);
    o[1] = dot3p(m[1], p.x, m[5], p.y, m[9], p.z, m[13]);
    o[2] = dot3p(m[2], p.x, m[6], p.y, m[10], p.z, m[14]);
    o[3] = dot3p(m[3], p.x, m[7], p.y, m[11], p.z, m[15]);
}
/* auxiliary.h:89-97 */
static f3 xformvec43T(f3 p, const float* m) {
    f3 t = {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
            m[8] * p.x + m[9] * p.y + m[10] * p.z};
    return t;
}
/* auxiliary.h:41-44 (double arithmetic because of the 1.0 / 0.5 literals) */
static float ndc2Pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:46-56 */
static void getRect(float px, float py, int max_radius, int gx, int gy, int rmin[2], int rmax[2]) {
    rmin[0] = imin(gx, imax(0, (int)((px - max_radius) / BLOCK_X)));
    rmin[1] = imin(gy, imax(0, (int)((py - max_radius) / BLOCK_Y)));
    rmax[0] = imin(gx, imax(0, (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X)));
    rmax[1] = imin(gy, imax(0, (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y)));
}

/* forward.cu:118-152 */
static void computeCov3D(f3 scale, float mod, const float rot[4], float* cov3D) {
    mat3 S = mat3_make(1, 0, 0, 0, 1, 0, 0, 0, 1);
    S.m[0][0] = mod * scale.x;
    S.m[1][1] = mod * scale.y;
    S.m[2][2] = mod * scale.z;
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3]; /* NOT normalised, forward.cu:127 */
    /* forward.cu:134-138. The `a*b +/- c*d` products are spelled with fmaf() the way ptxas contracted them in the
     * reference's sm_100 build (x*z, r*x, r*z, y*y, z*z rounded; the partner product fused), so the tiny off-diagonal
     * covariance entries match the CUDA reference bit for bit. */
    const float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    const float xy_m_rz = fmaf(x, y, -rz), xy_p_rz = fmaf(x, y, rz);
    const float xz_p_ry = fmaf(r, y, xz), xz_m_ry = fmaf(-r, y, xz);
    const float yz_m_rx = fmaf(y, z, -rx), yz_p_rx = fmaf(y, z, rx);
    const float yy_zz = yy + zz, xx_zz = fmaf(x, x, zz), xx_yy = fmaf(x, x, yy);
    mat3 R = mat3_make(1.f - (yy_zz + yy_zz), xy_m_rz + xy_m_rz, xz_p_ry + xz_p_ry,
                       xy_p_rz + xy_p_rz, 1.f - (xx_zz + xx_zz), yz_m_rx + yz_m_rx,
                       xz_m_ry + xz_m_ry, yz_p_rx + yz_p_rx, 1.f - (xx_yy + xx_yy));
    mat3 M = mat3_mul(&S, &R);
    mat3 Mt = mat3_T(&M);
    mat3 Sigma = mat3_mul(&Mt, &M);
    cov3D[0] = Sigma.m[0][0];
    cov3D[1] = Sigma.m[0][1];
    cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1];
    cov3D[4] = Sigma.m[1][2];
    cov3D[5] = Sigma.m[2][2];
}

typedef struct {
    mat3 T, Vrk;
    f3 t;
    float txtz, tytz, limx, limy, a, b, c;
} ewa_t;

/* forward.cu:74-113 and backward.cu:162-205 */
static ewa_t ewa(f3 mean, float fx, float fy, float tan_fovx, float tan_fovy, const float* cov3D, const float* vm) {
    ewa_t e;
    f3 t = xform43(mean, vm);
    e.limx = 1.3f * tan_fovx;
    e.limy = 1.3f * tan_fovy;
    e.txtz = t.x / t.z;
    e.tytz = t.y / t.z;
    t.x = fminf(e.limx, fmaxf(-e.limx, e.txtz)) * t.z;
    t.y = fminf(e.limy, fmaxf(-e.limy, e.tytz)) * t.z;
    e.t = t;
    mat3 J = mat3_make(fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z), 0.0f, fy / t.z, -(fy * t.y) / (t.z * t.z), 0, 0, 0);
    mat3 W = mat3_make(vm[0], vm[4], vm[8], vm[1], vm[5], vm[9], vm[2], vm[6], vm[10]);
    e.T = mat3_mul(&W, &J);
    e.Vrk = mat3_make(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    mat3 Tt = mat3_T(&e.T), Vt = mat3_T(&e.Vrk);
    mat3 tmp = mat3_mul(&Tt, &Vt);
    mat3 cov = mat3_mul(&tmp, &e.T);
    e.a = cov.m[0][0] + 0.3f;
    e.b = cov.m[0][1];
    e.c = cov.m[1][1] + 0.3f;
    return e;
}

/* forward.cu:20-71 */
static void colorFromSH(int deg, const float* sh /* [M][3] */, f3 pos, const float* campos, float rgb[3],
                        uint8_t clamped[3]) {
    f3 dir = {pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]};
    float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    float x = dir.x / len, y = dir.y / len, z = dir.z / len;
    for (int c = 0; c < 3; c++) {
        float res = SH_C0 * sh[c];
        if (deg > 0) {
            res = res - SH_C1 * y * sh[3 + c] + SH_C1 * z * sh[6 + c] - SH_C1 * x * sh[9 + c];
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + SH_C2[0] * xy * sh[12 + c] + SH_C2[1] * yz * sh[15 + c] +
                      SH_C2[2] * (2.0f * zz - xx - yy) * sh[18 + c] + SH_C2[3] * xz * sh[21 + c] +
                      SH_C2[4] * (xx - yy) * sh[24 + c];
                if (deg > 2) {
                    res = res + SH_C3[0] * y * (3.0f * xx - yy) * sh[27 + c] + SH_C3[1] * xy * z * sh[30 + c] +
                          SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[33 + c] +
                          SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + c] +
                          SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[39 + c] + SH_C3[5] * z * (xx - yy) * sh[42 + c] +
                          SH_C3[6] * x * (xx - 3.0f * yy) * sh[45 + c];
                }
            }
        }
        res += 0.5f;
        clamped[c] = res < 0;
        rgb[c] = res < 0.0f ? 0.0f : res;
    }
}

/* ------------------------------------------------------------------------------------------------------------------
 * Forward preprocess. forward.cu:155-256 (preprocessCUDA) + auxiliary.h:139-164 (in_frustum).
 * Outputs are P-sized arrays; rows of culled Gaussians keep radii = tiles_touched = 0 and are otherwise untouched.
 * ---------------------------------------------------------------------------------------------------------------- */
void oracle_preprocess(int P, int D, int M, const float* means, const float* scales, float scale_mod,
                       const float* rots, const float* opac, const float* shs, const float* cov3D_precomp,
                       const float* colors_precomp, const float* view, const float* proj, const float* campos, int W,
                       int H, float tan_fovx, float tan_fovy, int32_t* radii, float* xy /*[P][2]*/, float* depths,
                       float* cov3Ds /*[P][6]*/, float* rgb /*[P][3]*/, float* conic_opacity /*[P][4]*/,
                       uint32_t* tiles_touched, uint8_t* clamped /*[P][3]*/) {
    const float focal_y = H / (2.0f * tan_fovy); /* rasterizer_impl.cu:222-223 */
    const float focal_x = W / (2.0f * tan_fovx);
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        f3 p_orig = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
        float p_hom[4];
        xform44(p_orig, proj, p_hom);
        float p_w = 1.0f / (p_hom[3] + 0.0000001f);
        f3 p_proj = {p_hom[0] * p_w, p_hom[1] * p_w, p_hom[2] * p_w};
        f3 p_view = xform43(p_orig, view);
        if (p_view.z <= NEAR_Z) continue;
        const float* cov3D;
        if (cov3D_precomp) {
            cov3D = cov3D_precomp + 6 * idx;
        } else {
            f3 sc = {scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
            computeCov3D(sc, scale_mod, rots + 4 * idx, cov3Ds + 6 * idx);
            cov3D = cov3Ds + 6 * idx;
        }
        ewa_t e = ewa(p_orig, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, view);
        float cx = e.a, cy = e.b, cz = e.c;
        float det = fmaf(cx, cz, -(cy * cy)); /* FFMA(cov.x, cov.z, -(cov.y*cov.y)) in the reference build */
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float conic[3] = {cz * det_inv, -cy * det_inv, cx * det_inv};
        float mid = 0.5f * (cx + cz);
        float lambda1 = mid + sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        float lambda2 = mid - sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
        float pix = ndc2Pix(p_proj.x, W), piy = ndc2Pix(p_proj.y, H);
        int rmin[2], rmax[2];
        getRect(pix, piy, (int)my_radius, gx, gy, rmin, rmax);
        if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
        if (!colors_precomp) {
            colorFromSH(D, shs + (size_t)idx * M * 3, p_orig, campos, rgb + 3 * idx, clamped + 3 * idx);
        }
        depths[idx] = p_view.z;
        radii[idx] = (int32_t)my_radius;
        xy[2 * idx] = pix;
        xy[2 * idx + 1] = piy;
        conic_opacity[4 * idx + 0] = conic[0];
        conic_opacity[4 * idx + 1] = conic[1];
        conic_opacity[4 * idx + 2] = conic[2];
        conic_opacity[4 * idx + 3] = opac[idx];
        tiles_touched[idx] = (uint32_t)((rmax[1] - rmin[1]) * (rmax[0] - rmin[0]));
    }
}

/* rasterizer_impl.cu:35-50 */
static uint32_t getHigherMsb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step;
        else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* Inclusive scan of tiles_touched (rasterizer_impl.cu:277). Returns num_rendered. */
int64_t oracle_scan(int P, const uint32_t* tiles_touched, uint32_t* point_offsets) {
    uint32_t run = 0;
    for (int i = 0; i < P; i++) {
        run += tiles_touched[i];
        point_offsets[i] = run;
    }
    return P > 0 ? (int64_t)run : 0;
}

/* duplicateWithKeys (rasterizer_impl.cu:70-111) + stable LSD radix sort over bits [0, 32+bit) (:300-308, the CUB
 * call) + identifyTileRanges (:116-138) after the memset of ranges (:310). */
void oracle_bin_sort(int P, int R, int W, int H, const float* xy, const float* depths, const int32_t* radii,
                     const uint32_t* point_offsets, uint64_t* keys_unsorted, uint32_t* vals_unsorted,
                     uint64_t* keys_sorted, uint32_t* point_list, uint32_t* ranges /*[T][2]*/) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    for (int idx = 0; idx < P; idx++) {
        if (radii[idx] > 0) {
            uint32_t off = (idx == 0) ? 0 : point_offsets[idx - 1];
            int rmin[2], rmax[2];
            getRect(xy[2 * idx], xy[2 * idx + 1], radii[idx], gx, gy, rmin, rmax);
            for (int y = rmin[1]; y < rmax[1]; y++)
                for (int x = rmin[0]; x < rmax[0]; x++) {
                    uint64_t key = (uint64_t)(y * gx + x);
                    key <<= 32;
                    uint32_t dbits;
                    memcpy(&dbits, &depths[idx], 4);
                    key |= dbits;
                    keys_unsorted[off] = key;
                    vals_unsorted[off] = (uint32_t)idx;
                    off++;
                }
        }
    }
    const int end_bit = 32 + (int)getHigherMsb((uint32_t)(gx * gy));
    uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R > 0 ? R : 1));
    uint32_t* va = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R > 0 ? R : 1));
    uint64_t* kb = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R > 0 ? R : 1));
    uint32_t* vb = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R > 0 ? R : 1));
    memcpy(ka, keys_unsorted, sizeof(uint64_t) * (size_t)R);
    memcpy(va, vals_unsorted, sizeof(uint32_t) * (size_t)R);
    for (int shift = 0; shift < end_bit; shift += 8) {
        size_t cnt[257];
        memset(cnt, 0, sizeof(cnt));
        int nbits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint64_t mask = (1ull << nbits) - 1;
        for (int i = 0; i < R; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (int i = 0; i < R; i++) {
            size_t d = (ka[i] >> shift) & mask;
            kb[cnt[d]] = ka[i];
            vb[cnt[d]] = va[i];
            cnt[d]++;
        }
        uint64_t* tk = ka; ka = kb; kb = tk;
        uint32_t* tv = va; va = vb; vb = tv;
    }
    memcpy(keys_sorted, ka, sizeof(uint64_t) * (size_t)R);
    memcpy(point_list, va, sizeof(uint32_t) * (size_t)R);
    free(ka); free(va); free(kb); free(vb);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)(gx * gy));
    for (int idx = 0; idx < R; idx++) {
        uint32_t currtile = (uint32_t)(keys_sorted[idx] >> 32);
        if (idx == 0) ranges[2 * currtile] = 0;
        else {
            uint32_t prevtile = (uint32_t)(keys_sorted[idx - 1] >> 32);
            if (currtile != prevtile) {
                ranges[2 * prevtile + 1] = (uint32_t)idx;
                ranges[2 * currtile] = (uint32_t)idx;
            }
        }
        if (idx == R - 1) ranges[2 * currtile + 1] = (uint32_t)R;
    }
}

/* Forward blend, one pixel at a time. forward.cu:261-374 (renderCUDA): thresholds 0.99 / 1/255 / 1e-4 and the
 * `contributor` bookkeeping (list positions, :328,:361). features = rgb or colors_precomp (rasterizer_impl.cu:321).
 * out_depth/out_alpha (optional) are the additive outputs: sum z*alpha*T and 1 - final_T. */
void oracle_render_fwd(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* xy,
                       const float* features, const float* conic_opacity, const float* depths, const float* bg,
                       float* final_T, uint32_t* n_contrib, float* out_color, float* out_depth, float* out_alpha) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X;
#pragma omp parallel for schedule(dynamic, 64)
    for (int pix_id = 0; pix_id < W * H; pix_id++) {
        int px = pix_id % W, py = pix_id / W;
        int tile = (py / BLOCK_Y) * gx + (px / BLOCK_X);
        uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        float pixfx = (float)px, pixfy = (float)py;
        float T = 1.0f;
        uint32_t contributor = 0, last_contributor = 0;
        float C[3] = {0, 0, 0};
        float Dz = 0.f;
        for (uint32_t k = r0; k < r1; k++) {
            contributor++;
            uint32_t id = point_list[k];
            float dx = xy[2 * id] - pixfx, dy = xy[2 * id + 1] - pixfy;
            const float* co = conic_opacity + 4 * id;
            float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
            if (power > 0.0f) continue;
            float alpha = fminf(0.99f, co[3] * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) break; /* done = true */
            for (int ch = 0; ch < 3; ch++) C[ch] += features[3 * id + ch] * alpha * T;
            Dz += depths[id] * alpha * T;
            T = test_T;
            last_contributor = contributor;
        }
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        for (int ch = 0; ch < 3; ch++) out_color[ch * H * W + pix_id] = C[ch] + T * bg[ch];
        if (out_depth) out_depth[pix_id] = Dz;
        if (out_alpha) out_alpha[pix_id] = 1.0f - T;
    }
}

/* Backward blend. backward.cu:399-557. Per-Gaussian sums use atomics (order is nondeterministic there as well). */
void oracle_render_bwd(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* bg,
                       const float* xy, const float* conic_opacity, const float* colors, const float* final_Ts,
                       const uint32_t* n_contrib, const float* dL_dpixels, float* dL_dmean2D /*[P][3]*/,
                       float* dL_dconic2D /*[P][4]*/, float* dL_dopacity /*[P]*/, float* dL_dcolors /*[P][3]*/) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X;
    const float ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;
#pragma omp parallel for schedule(dynamic, 64)
    for (int pix_id = 0; pix_id < W * H; pix_id++) {
        int px = pix_id % W, py = pix_id / W;
        int tile = (py / BLOCK_Y) * gx + (px / BLOCK_X);
        uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        float pixfx = (float)px, pixfy = (float)py;
        const float T_final = final_Ts[pix_id];
        float T = T_final;
        const uint32_t last_contributor = n_contrib[pix_id];
        float accum_rec[3] = {0, 0, 0}, dL_dpixel[3], last_color[3] = {0, 0, 0};
        for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * H * W + pix_id];
        float last_alpha = 0;
        float bg_dot_dpixel = 0;
        for (int i = 0; i < 3; i++) bg_dot_dpixel += bg[i] * dL_dpixel[i];
        /* positions r1-1 ... r0; contributor counts down from toDo = r1-r0 (backward.cu:437,:486-488) */
        for (uint32_t pos = last_contributor; pos-- > 0;) {
            uint32_t id = point_list[r0 + pos];
            float dx = xy[2 * id] - pixfx, dy = xy[2 * id + 1] - pixfy;
            const float* co = conic_opacity + 4 * id;
            float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
            if (power > 0.0f) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, co[3] * G);
            if (alpha < 1.0f / 255.0f) continue;
            T = T / (1.f - alpha);
            const float dchannel_dcolor = alpha * T;
            float dL_dalpha = 0.0f;
            for (int ch = 0; ch < 3; ch++) {
                const float c = colors[3 * id + ch];
                accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                last_color[ch] = c;
                const float dL_dchannel = dL_dpixel[ch];
                dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
                float v = dchannel_dcolor * dL_dchannel;
#pragma omp atomic
                dL_dcolors[3 * id + ch] += v;
            }
            dL_dalpha *= T;
            last_alpha = alpha;
            dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
            const float dL_dG = co[3] * dL_dalpha;
            const float gdx = G * dx, gdy = G * dy;
            const float dG_ddelx = -gdx * co[0] - gdy * co[1];
            const float dG_ddely = -gdy * co[2] - gdx * co[1];
            float v0 = dL_dG * dG_ddelx * ddelx_dx, v1 = dL_dG * dG_ddely * ddely_dy;
            float c0 = -0.5f * gdx * dx * dL_dG, c1 = -0.5f * gdx * dy * dL_dG, c3 = -0.5f * gdy * dy * dL_dG;
            float vo = G * dL_dalpha;
#pragma omp atomic
            dL_dmean2D[3 * id + 0] += v0;
#pragma omp atomic
            dL_dmean2D[3 * id + 1] += v1;
#pragma omp atomic
            dL_dconic2D[4 * id + 0] += c0;
#pragma omp atomic
            dL_dconic2D[4 * id + 1] += c1;
#pragma omp atomic
            dL_dconic2D[4 * id + 3] += c3;
#pragma omp atomic
            dL_dopacity[id] += vo;
        }
    }
}

/* auxiliary.h:107-117 */
static f3 dnormvdv(f3 v, f3 dv) {
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    f3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

#define SHV(k, c) (sh[3 * (k) + (c)])

/* backward.cu:20-139 */
static f3 colorFromSH_bwd(int deg, const float* sh, f3 pos, const float* campos, const uint8_t* clamped,
                          const float* dL_dcolor, float* dL_dsh /* row of this Gaussian */) {
    f3 dir_orig = {pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]};
    float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    float g[3];
    for (int c = 0; c < 3; c++) g[c] = dL_dcolor[c] * (clamped[c] ? 0 : 1);
    float dx_[3] = {0, 0, 0}, dy_[3] = {0, 0, 0}, dz_[3] = {0, 0, 0};
    float w[16];
    int nw = (deg + 1) * (deg + 1);
    w[0] = SH_C0;
    if (deg > 0) {
        w[1] = -SH_C1 * y;
        w[2] = SH_C1 * z;
        w[3] = -SH_C1 * x;
        for (int c = 0; c < 3; c++) {
            dx_[c] = -SH_C1 * SHV(3, c);
            dy_[c] = -SH_C1 * SHV(1, c);
            dz_[c] = SH_C1 * SHV(2, c);
        }
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            w[4] = SH_C2[0] * xy;
            w[5] = SH_C2[1] * yz;
            w[6] = SH_C2[2] * (2.f * zz - xx - yy);
            w[7] = SH_C2[3] * xz;
            w[8] = SH_C2[4] * (xx - yy);
            for (int c = 0; c < 3; c++) {
                dx_[c] += SH_C2[0] * y * SHV(4, c) + SH_C2[2] * 2.f * -x * SHV(6, c) + SH_C2[3] * z * SHV(7, c) +
                          SH_C2[4] * 2.f * x * SHV(8, c);
                dy_[c] += SH_C2[0] * x * SHV(4, c) + SH_C2[1] * z * SHV(5, c) + SH_C2[2] * 2.f * -y * SHV(6, c) +
                          SH_C2[4] * 2.f * -y * SHV(8, c);
                dz_[c] += SH_C2[1] * y * SHV(5, c) + SH_C2[2] * 2.f * 2.f * z * SHV(6, c) + SH_C2[3] * x * SHV(7, c);
            }
            if (deg > 2) {
                w[9] = SH_C3[0] * y * (3.f * xx - yy);
                w[10] = SH_C3[1] * xy * z;
                w[11] = SH_C3[2] * y * (4.f * zz - xx - yy);
                w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                w[13] = SH_C3[4] * x * (4.f * zz - xx - yy);
                w[14] = SH_C3[5] * z * (xx - yy);
                w[15] = SH_C3[6] * x * (xx - 3.f * yy);
                for (int c = 0; c < 3; c++) {
                    dx_[c] += (SH_C3[0] * SHV(9, c) * 3.f * 2.f * xy + SH_C3[1] * SHV(10, c) * yz +
                               SH_C3[2] * SHV(11, c) * -2.f * xy + SH_C3[3] * SHV(12, c) * -3.f * 2.f * xz +
                               SH_C3[4] * SHV(13, c) * (-3.f * xx + 4.f * zz - yy) + SH_C3[5] * SHV(14, c) * 2.f * xz +
                               SH_C3[6] * SHV(15, c) * 3.f * (xx - yy));
                    dy_[c] += (SH_C3[0] * SHV(9, c) * 3.f * (xx - yy) + SH_C3[1] * SHV(10, c) * xz +
                               SH_C3[2] * SHV(11, c) * (-3.f * yy + 4.f * zz - xx) +
                               SH_C3[3] * SHV(12, c) * -3.f * 2.f * yz + SH_C3[4] * SHV(13, c) * -2.f * xy +
                               SH_C3[5] * SHV(14, c) * -2.f * yz + SH_C3[6] * SHV(15, c) * -3.f * 2.f * xy);
                    dz_[c] += (SH_C3[1] * SHV(10, c) * xy + SH_C3[2] * SHV(11, c) * 4.f * 2.f * yz +
                               SH_C3[3] * SHV(12, c) * 3.f * (2.f * zz - xx - yy) +
                               SH_C3[4] * SHV(13, c) * 4.f * 2.f * xz + SH_C3[5] * SHV(14, c) * (xx - yy));
                }
            }
        }
    }
    for (int k = 0; k < nw; k++)
        for (int c = 0; c < 3; c++) dL_dsh[3 * k + c] = w[k] * g[c];
    f3 dL_ddir = {dx_[0] * g[0] + dx_[1] * g[1] + dx_[2] * g[2], dy_[0] * g[0] + dy_[1] * g[1] + dy_[2] * g[2],
                  dz_[0] * g[0] + dz_[1] * g[1] + dz_[2] * g[2]};
    return dnormvdv(dir_orig, dL_ddir);
}

/* Backward preprocess: computeCov2DCUDA (backward.cu:144-274), preprocessCUDA (:346-396), computeCov3D (:278-341).
 * All gradient arrays must be zero-initialised by the caller (the reference uses torch::zeros, rasterize_points.cu:
 * 151-159). cov3Ds = geomState.cov3D or cov3D_precomp (rasterizer_impl.cu:402). */
void oracle_preprocess_bwd(int P, int D, int M, const float* means, const int32_t* radii, const float* shs,
                           const uint8_t* clamped, const float* scales, const float* rots, float scale_mod,
                           const float* cov3Ds, const float* view, const float* proj, int W, int H, float tan_fovx,
                           float tan_fovy, const float* campos, const float* dL_dmean2D /*[P][3]*/,
                           const float* dL_dconic /*[P][4]*/, const float* dL_dcolor /*[P][3]*/,
                           float* dL_dmeans /*[P][3]*/, float* dL_dcov3D /*[P][6]*/, float* dL_dsh /*[P][M][3]*/,
                           float* dL_dscale /*[P][3]*/, float* dL_drot /*[P][4]*/) {
    const float h_y = H / (2.0f * tan_fovy), h_x = W / (2.0f * tan_fovx);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (!(radii[idx] > 0)) continue;
        const float* cov3D = cov3Ds + 6 * idx;
        f3 mean = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
        f3 dLc = {dL_dconic[4 * idx], dL_dconic[4 * idx + 1], dL_dconic[4 * idx + 3]};
        ewa_t e = ewa(mean, h_x, h_y, tan_fovx, tan_fovy, cov3D, view);
        f3 t = e.t;
        const float x_grad_mul = e.txtz < -e.limx || e.txtz > e.limx ? 0 : 1;
        const float y_grad_mul = e.tytz < -e.limy || e.tytz > e.limy ? 0 : 1;
        const mat3 T = e.T, Vrk = e.Vrk;
        float a = e.a, b = e.b, c = e.c;
        float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcov = dL_dcov3D + 6 * idx;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dLc.x + 2 * b * c * dLc.y + (denom - a * c) * dLc.z);
            dL_dc = denom2inv * (-a * a * dLc.z + 2 * a * b * dLc.y + (denom - a * c) * dLc.x);
            dL_db = denom2inv * 2 * (b * c * dLc.x - (denom + 2 * b * b) * dLc.y + a * b * dLc.z);
            dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
            dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
            dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
            dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
                      2 * T.m[1][0] * T.m[1][1] * dL_dc;
            dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
                      2 * T.m[1][0] * T.m[1][2] * dL_dc;
            dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
                      2 * T.m[1][1] * T.m[1][2] * dL_dc;
        } else {
            for (int i = 0; i < 6; i++) dcov[i] = 0;
        }
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
        /* W[c][r] = view[4*r + c] */
        const float W00 = view[0], W01 = view[4], W02 = view[8], W10 = view[1], W11 = view[5], W12 = view[9],
                    W20 = view[2], W21 = view[6], W22 = view[10];
        float dL_dJ00 = W00 * dL_dT00 + W01 * dL_dT01 + W02 * dL_dT02;
        float dL_dJ02 = W20 * dL_dT00 + W21 * dL_dT01 + W22 * dL_dT02;
        float dL_dJ11 = W10 * dL_dT10 + W11 * dL_dT11 + W12 * dL_dT12;
        float dL_dJ12 = W20 * dL_dT10 + W21 * dL_dT11 + W22 * dL_dT12;
        float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
        float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                       (2 * h_y * t.y) * tz3 * dL_dJ12;
        f3 dtv = {dL_dtx, dL_dty, dL_dtz};
        f3 dm = xformvec43T(dtv, view);
        /* backward.cu:366-384 */
        float m_hom[4];
        xform44(mean, proj, m_hom);
        float m_w = 1.0f / (m_hom[3] + 0.0000001f);
        float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
        float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
        float gx_ = dL_dmean2D[3 * idx], gy_ = dL_dmean2D[3 * idx + 1];
        f3 d2;
        d2.x = (proj[0] * m_w - proj[3] * mul1) * gx_ + (proj[1] * m_w - proj[3] * mul2) * gy_;
        d2.y = (proj[4] * m_w - proj[7] * mul1) * gx_ + (proj[5] * m_w - proj[7] * mul2) * gy_;
        d2.z = (proj[8] * m_w - proj[11] * mul1) * gx_ + (proj[9] * m_w - proj[11] * mul2) * gy_;
        dm.x += d2.x; dm.y += d2.y; dm.z += d2.z;
        if (shs) {
            f3 ds = colorFromSH_bwd(D, shs + (size_t)idx * M * 3, mean, campos, clamped + 3 * idx, dL_dcolor + 3 * idx,
                                    dL_dsh + (size_t)idx * M * 3);
            dm.x += ds.x; dm.y += ds.y; dm.z += ds.z;
        }
        dL_dmeans[3 * idx] = dm.x;
        dL_dmeans[3 * idx + 1] = dm.y;
        dL_dmeans[3 * idx + 2] = dm.z;
        if (scales) { /* backward.cu:278-341 */
            const float* q = rots + 4 * idx;
            float r = q[0], x = q[1], y = q[2], z = q[3];
            mat3 R = mat3_make(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                               2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                               2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
            mat3 S = mat3_make(1, 0, 0, 0, 1, 0, 0, 0, 1);
            f3 s = {scale_mod * scales[3 * idx], scale_mod * scales[3 * idx + 1], scale_mod * scales[3 * idx + 2]};
            S.m[0][0] = s.x; S.m[1][1] = s.y; S.m[2][2] = s.z;
            mat3 Mm = mat3_mul(&S, &R);
            mat3 dSig = mat3_make(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                                  0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
            mat3 M2;
            for (int cc = 0; cc < 3; cc++)
                for (int rr = 0; rr < 3; rr++) M2.m[cc][rr] = 2.0f * Mm.m[cc][rr];
            mat3 dL_dM = mat3_mul(&M2, &dSig);
            mat3 Rt = mat3_T(&R);
            mat3 dMt = mat3_T(&dL_dM);
            dL_dscale[3 * idx + 0] = Rt.m[0][0] * dMt.m[0][0] + Rt.m[0][1] * dMt.m[0][1] + Rt.m[0][2] * dMt.m[0][2];
            dL_dscale[3 * idx + 1] = Rt.m[1][0] * dMt.m[1][0] + Rt.m[1][1] * dMt.m[1][1] + Rt.m[1][2] * dMt.m[1][2];
            dL_dscale[3 * idx + 2] = Rt.m[2][0] * dMt.m[2][0] + Rt.m[2][1] * dMt.m[2][1] + Rt.m[2][2] * dMt.m[2][2];
            for (int k = 0; k < 3; k++) {
                dMt.m[0][k] *= s.x;
                dMt.m[1][k] *= s.y;
                dMt.m[2][k] *= s.z;
            }
            float* dq = dL_drot + 4 * idx;
            dq[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) +
                    2 * x * (dMt.m[1][2] - dMt.m[2][1]);
            dq[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) +
                    2 * r * (dMt.m[1][2] - dMt.m[2][1]) - 4 * x * (dMt.m[2][2] + dMt.m[1][1]);
            dq[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) +
                    2 * z * (dMt.m[1][2] + dMt.m[2][1]) - 4 * y * (dMt.m[2][2] + dMt.m[0][0]);
            dq[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) +
                    2 * y * (dMt.m[1][2] + dMt.m[2][1]) - 4 * z * (dMt.m[1][1] + dMt.m[0][0]);
        }
    }
}

/* checkFrustum, rasterizer_impl.cu:54-66 with the 0.001 near plane. */
void oracle_mark_visible(int P, const float* means, const float* view, const float* proj, uint8_t* present) {
    (void)proj;
    for (int idx = 0; idx < P; idx++) {
        f3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
        f3 pv = xform43(p, view);
        present[idx] = pv.z <= NEAR_Z ? 0 : 1;
    }
}

/* simple-knn: exact mean of the 3 smallest squared distances to other indices (submodules/simple-knn/simple_knn.cu:
 * 132-184; only `i == idx` is skipped, duplicates count). Brute force O(P^2): small cases only. */
void oracle_knn(int P, const float* pts, float* out) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        for (int j = 0; j < P; j++) {
            if (j == i) continue;
            float dx = pts[3 * j] - pts[3 * i], dy = pts[3 * j + 1] - pts[3 * i + 1], dz = pts[3 * j + 2] - pts[3 * i + 2];
            float dist = dx * dx + dy * dy + dz * dz;
            for (int k = 0; k < 3; k++) {
                if (best[k] > dist) {
                    float t = best[k];
                    best[k] = dist;
                    dist = t;
                }
            }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
