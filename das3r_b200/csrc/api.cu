// C ABI of the rasterizer library (see include/das3r_raster.h). Host-side orchestration only: layout of the three
// state blobs, stage sequencing, error reporting. Replaces CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (reference cuda_rasterizer/rasterizer_impl.cu:141-434) and SimpleKNN::knn (simple-knn/simple_knn.cu:186-222).
#include "../../include/das3r_raster.h"
#include "../../include/das3r_frontend.h"
#include "../../include/das3r_loss.h"
#include "../../include/das3r_adam.h"
#include "../../include/das3r_comm.h"
#include "../../include/das3r_init.h"
#include <math.h>
#include "kernels.cuh"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <algorithm>
#include <atomic>
#include <mutex>

using namespace das3r;

namespace {

thread_local std::string g_err;

int fail(int code, const char* where, cudaError_t e) {
    char buf[512];
    snprintf(buf, sizeof(buf), "[das3r_raster] %s: %s", where, e == cudaSuccess ? "failed" : cudaGetErrorString(e));
    g_err = buf;
    return code;
}
int fail_msg(int code, const std::string& msg) {
    g_err = std::string("[das3r_raster] ") + msg;
    return code;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Cursor {
    size_t off = 0;
    size_t take(size_t bytes) {
        size_t o = off;
        off += align_up(bytes, 256);
        return o;
    }
};

void geom_layout(int P, das3r_raster_state_layout* L) {
    Cursor c;
    const size_t n = (size_t)P;
    L->g_recA = c.take(n * sizeof(GaussRec));  // AoS: A at +0, B at +16, C at +32 of each 48-byte record
    L->g_recB = L->g_recA + 16;
    L->g_recC = L->g_recA + 32;
    L->g_tiles_touched = c.take(n * 4);
    L->g_clamped = c.take(n);
    L->geom_bytes = c.off;
}

void image_layout(int W, int H, das3r_raster_state_layout* L) {
    Cursor c;
    const size_t N = (size_t)W * H;
    const size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    L->i_final_T = c.take(N * 4);
    L->i_n_contrib = c.take(N * 4);
    L->i_ranges = c.take(T * 8);
    L->i_tile_counts = c.take(T * 4);
    L->i_cursors = c.take(T * 4);
    L->i_totals = c.take(16);
    L->image_bytes = c.off;
}

// Tail of the binning blob used only by the large-tile fallback starts at b_fallback.
struct BinExtra {
    size_t b_fallback, fallback_bytes;
};

void binning_layout(int R, int T, bool with_keys, bool with_fallback, das3r_raster_state_layout* L, BinExtra* X) {
    Cursor c;
    const size_t n = (size_t)R;
    L->b_unsorted = c.take(n * 8);
    L->b_point_list = c.take(n * 4);
    L->b_rect_off = c.take((size_t)T * 8 * 4);   // uint32[T*8] start of each rectangle's list
    L->b_rect_cnt = c.take((size_t)T * 8 * 4);   // uint32[T*8] its length
    L->b_rect_list = c.take(n * 8 * 8);          // uint2 {position, pixel mask}, capacity 8 per instance (<= 8 rectangles)
    L->b_rect_gid = c.take(n * 8 * 4);           // uint32 Gaussian index of the same entry
    L->b_keys_sorted = with_keys ? c.take(n * 8) : (size_t)0;
    X->fallback_bytes = with_fallback ? tile_sort_fallback_bytes(R, T) : 0;
    X->b_fallback = with_fallback ? c.take(X->fallback_bytes) : 0;
    L->binning_bytes = c.off;
}

// ---- optional per-stage CUDA-event timing + kernel launch counter (bench.py's roofline leg) ------------------------
enum Stage {
    ST_PREPROCESS_FWD = 0, ST_TILE_SCAN, ST_SCATTER, ST_TILE_SORT, ST_BLEND_FWD, ST_BLEND_BWD, ST_PREPROCESS_BWD,
    ST_MEMSET, ST_RECT_LISTS, ST_FRONTEND_FWD, ST_FRONTEND_BWD, ST_COUNT
};
const char* const kStageNames[ST_COUNT] = {"preprocess_fwd", "tile_scan", "scatter", "tile_sort", "blend_fwd",
                                           "blend_bwd", "preprocess_bwd", "memset", "rect_lists", "frontend_fwd",
                                           "frontend_bwd"};
struct Pending {
    int stage;
    cudaEvent_t a, b;
};
struct Prof {
    bool enabled = false;
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> free_events;
    double ms[ST_COUNT] = {0};
    long long calls[ST_COUNT] = {0};
    cudaEvent_t get() {
        if (!free_events.empty()) {
            cudaEvent_t e = free_events.back();
            free_events.pop_back();
            return e;
        }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }
};
Prof g_prof;
std::atomic<long long> g_launches{0};

struct StageScope {
    int stage;
    cudaStream_t stream;
    cudaEvent_t a = nullptr;
    StageScope(int st, cudaStream_t s, int kernels) : stage(st), stream(s) {
        g_launches += kernels;
        if (g_prof.enabled) {
            a = g_prof.get();
            cudaEventRecord(a, stream);
        }
    }
    ~StageScope() {
        if (a) {
            cudaEvent_t b = g_prof.get();
            cudaEventRecord(b, stream);
            g_prof.pending.push_back({stage, a, b});
        }
    }
};

bool check_params(const das3r_raster_params* p) {
    return p && p->P >= 0 && p->width > 0 && p->height > 0 && p->D >= 0 && p->D <= 3 && p->M >= 0;
}

#define STAGE(call, name, stage_id, nkernels)                               \
    do {                                                                    \
        cudaError_t e_;                                                     \
        {                                                                   \
            StageScope scope_(stage_id, stream, nkernels);                  \
            e_ = (call);                                                    \
        }                                                                   \
        if (e_ != cudaSuccess) return fail(DAS3R_ECUDA, name, e_);          \
        if (params->debug) {                                                \
            e_ = cudaStreamSynchronize(stream);                             \
            if (e_ == cudaSuccess) e_ = cudaGetLastError();                 \
            if (e_ != cudaSuccess) return fail(DAS3R_ECUDA, name, e_);      \
        }                                                                   \
    } while (0)

}  // namespace

extern "C" {

int das3r_abi_version(void) { return DAS3R_ABI_VERSION; }

const char* das3r_last_error(void) { return g_err.c_str(); }

int das3r_prof_enable(int on) {
    g_prof.enabled = on != 0;
    return DAS3R_OK;
}

int das3r_prof_num_stages(void) { return ST_COUNT; }

const char* das3r_prof_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }

int das3r_prof_collect(double* ms_total, long long* calls, int reset) {
    for (auto& p : g_prof.pending) {
        float ms = 0.f;
        cudaError_t e = cudaEventSynchronize(p.b);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, p.a, p.b);
        if (e != cudaSuccess) return fail(DAS3R_ECUDA, "prof_collect", e);
        g_prof.ms[p.stage] += ms;
        g_prof.calls[p.stage] += 1;
        g_prof.free_events.push_back(p.a);
        g_prof.free_events.push_back(p.b);
    }
    g_prof.pending.clear();
    for (int i = 0; i < ST_COUNT; i++) {
        if (ms_total) ms_total[i] = g_prof.ms[i];
        if (calls) calls[i] = g_prof.calls[i];
        if (reset) {
            g_prof.ms[i] = 0;
            g_prof.calls[i] = 0;
        }
    }
    return DAS3R_OK;
}

long long das3r_launch_count(void) { return g_launches.load(); }

int das3r_raster_state_layout_query(int32_t P, int32_t width, int32_t height, int32_t num_rendered,
                                    das3r_raster_state_layout* out) {
    if (!out || P < 0 || width <= 0 || height <= 0 || num_rendered < 0) return fail_msg(DAS3R_EINVAL, "bad layout query");
    memset(out, 0, sizeof(*out));
    geom_layout(P, out);
    image_layout(width, height, out);
    BinExtra x;
    const int T = ((width + TILE_X - 1) / TILE_X) * ((height + TILE_Y - 1) / TILE_Y);
    binning_layout(num_rendered, T, true, false, out, &x);
    return DAS3R_OK;
}

// ---- forward in two halves ----------------------------------------------------------------------------------------
// Validation shared by every forward entry point. Returns DAS3R_OK, or 1 for the P == 0 early-out (image zeroed).
static int forward_check(const char* who, const das3r_raster_params* params, const das3r_raster_inputs* in,
                         const das3r_raster_outputs* out, das3r_raster_state* state, cudaStream_t stream) {
    if (!check_params(params) || !in || !out || !state)
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": null or invalid argument");
    if (!out->out_color || (params->P > 0 && !out->radii)) return fail_msg(DAS3R_EINVAL, std::string(who) + ": null output");
    const int P = params->P, W = params->width, H = params->height;
    if (P == 0) {
        // rasterize_points.cu:68,:81 - the image stays all-zero (no background) when there are no Gaussians
        memset(state, 0, sizeof(*state));
        cudaError_t e = cudaMemsetAsync(out->out_color, 0, (size_t)3 * W * H * 4, stream);
        if (e == cudaSuccess && out->out_depth) e = cudaMemsetAsync(out->out_depth, 0, (size_t)W * H * 4, stream);
        if (e == cudaSuccess && out->out_alpha) e = cudaMemsetAsync(out->out_alpha, 0, (size_t)W * H * 4, stream);
        if (e != cudaSuccess) return fail(DAS3R_ECUDA, "memset(out_color)", e);
        return 1;
    }
    if (!in->means3D || !in->opacities || !params->background || !params->viewmatrix || !params->projmatrix ||
        !params->campos)
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": means3D/opacities/background/matrices/campos must be non-null");
    if ((in->shs == nullptr) == (in->colors_precomp == nullptr))
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": provide exactly one of shs / colors_precomp");
    const bool has_sr = in->scales != nullptr && in->rotations != nullptr;
    if (has_sr == (in->cov3D_precomp != nullptr) || ((in->scales != nullptr) != (in->rotations != nullptr)))
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": provide exactly one of (scales, rotations) / cov3D_precomp");
    if (in->shs && (params->M < (params->D + 1) * (params->D + 1) || params->M > 16))
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": M must satisfy (D+1)^2 <= M <= 16");
    if (in->rotations && (reinterpret_cast<uintptr_t>(in->rotations) & 15) != 0)
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": rotations must be 16-byte aligned (they are read as float4)");
    return DAS3R_OK;
}

// Allocates the geom and image blobs of one view, clears its tile counters and fills the preprocess arguments.
static int forward_begin(const das3r_raster_params* params, const das3r_raster_inputs* in, const das3r_raster_outputs* out,
                         das3r_alloc_fn geom_alloc, void* geom_ctx, das3r_alloc_fn image_alloc, void* image_ctx,
                         das3r_raster_state* state, cudaStream_t stream, PreFwdArgs* a_out) {
    const int P = params->P, W = params->width, H = params->height;
    memset(state, 0, sizeof(*state));
    const unsigned gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int T = (int)(gx * gy);
    das3r_raster_state_layout L;
    memset(&L, 0, sizeof(L));
    geom_layout(P, &L);
    image_layout(W, H, &L);
    char* geom = geom_alloc(L.geom_bytes, geom_ctx);
    char* img = image_alloc(L.image_bytes, image_ctx);
    if (!geom || !img) return fail_msg(DAS3R_ENOMEM, "forward: allocator returned NULL");
    state->geom = geom;
    state->geom_bytes = L.geom_bytes;
    state->image = img;
    state->image_bytes = L.image_bytes;
    uint32_t* tile_counts = reinterpret_cast<uint32_t*>(img + L.i_tile_counts);
    STAGE(cudaMemsetAsync(tile_counts, 0, (size_t)T * 4, stream), "memset(tile_counts)", ST_MEMSET, 0);
    PreFwdArgs a;
    memset(&a, 0, sizeof(a));
    a.P = P; a.D = params->D; a.M = params->M; a.W = W; a.H = H;
    a.gx = gx; a.gy = gy;
    a.tan_fovx = params->tan_fovx; a.tan_fovy = params->tan_fovy;
    a.focal_y = H / (2.0f * params->tan_fovy);  // rasterizer_impl.cu:222-223
    a.focal_x = W / (2.0f * params->tan_fovx);
    a.scale_mod = params->scale_modifier;
    a.prefiltered = params->prefiltered;
    a.means = in->means3D; a.shs = in->shs; a.colors_precomp = in->colors_precomp; a.opac = in->opacities;
    a.scales = in->scales; a.rots = in->rotations; a.cov3D_precomp = in->cov3D_precomp;
    a.view = params->viewmatrix; a.proj = params->projmatrix; a.campos = params->campos;
    a.radii = out->radii;
    a.rec = reinterpret_cast<GaussRec*>(geom + L.g_recA);
    a.tiles_touched = reinterpret_cast<uint32_t*>(geom + L.g_tiles_touched);
    a.clamped = reinterpret_cast<uint8_t*>(geom + L.g_clamped);
    a.tile_counts = tile_counts;
    *a_out = a;
    return DAS3R_OK;
}

// Per-view pointers of the binning / blend kernels that live in the geom and image blobs.
static void bin_fill_state(ViewBin& vb, int P, int W, int H, char* geom, char* img, const int32_t* radii) {
    das3r_raster_state_layout L;
    memset(&L, 0, sizeof(L));
    geom_layout(P, &L);
    image_layout(W, H, &L);
    vb.tile_counts = reinterpret_cast<uint32_t*>(img + L.i_tile_counts);
    vb.ranges = reinterpret_cast<uint2*>(img + L.i_ranges);
    vb.cursors = reinterpret_cast<uint32_t*>(img + L.i_cursors);
    vb.totals = reinterpret_cast<uint32_t*>(img + L.i_totals);
    vb.final_T = reinterpret_cast<float*>(img + L.i_final_T);
    vb.n_contrib = reinterpret_cast<uint32_t*>(img + L.i_n_contrib);
    vb.rec = reinterpret_cast<const GaussRec*>(geom + L.g_recA);
    vb.tiles_touched = reinterpret_cast<const uint32_t*>(geom + L.g_tiles_touched);
    vb.radii = radii;
}

// ... and those in the binning blob.
static void bin_fill_binning(ViewBin& vb, const das3r_raster_state_layout& L, char* bin, bool debug) {
    vb.unsorted = reinterpret_cast<uint64_t*>(bin + L.b_unsorted);
    vb.point_list = reinterpret_cast<uint32_t*>(bin + L.b_point_list);
    vb.rect_off = reinterpret_cast<uint32_t*>(bin + L.b_rect_off);
    vb.rect_cnt = reinterpret_cast<uint32_t*>(bin + L.b_rect_cnt);
    vb.rect_list = reinterpret_cast<uint2*>(bin + L.b_rect_list);
    vb.rect_gid = reinterpret_cast<uint32_t*>(bin + L.b_rect_gid);
    vb.keys_sorted = debug ? reinterpret_cast<uint64_t*>(bin + L.b_keys_sorted) : nullptr;
}

static void batch_header(BatchArgs& b, int n, int P, int W, int H) {
    memset(&b, 0, sizeof(b));
    b.n = n; b.P = P; b.W = W; b.H = H;
    b.gx = (unsigned)((W + TILE_X - 1) / TILE_X);
    b.gy = (unsigned)((H + TILE_Y - 1) / TILE_Y);
    b.T = (int)(b.gx * b.gy);
}

// Tile ranges of n views that share P and the image size: one launch, one block per view.
static int forward_scan(int n, const das3r_raster_fwd_view* const* views, cudaStream_t stream) {
    const das3r_raster_params* params = views[0]->params;
    BatchArgs b;
    batch_header(b, n, params->P, params->width, params->height);
    for (int v = 0; v < n; v++)
        bin_fill_state(b.v[v], params->P, params->width, params->height, views[v]->state->geom, views[v]->state->image,
                       views[v]->out->radii);
    STAGE(launch_tile_scan(b, stream), "tile_scan", ST_TILE_SCAN, 1);
    return DAS3R_OK;
}

// Second half of the forward for n <= RB_MAX_VIEWS views that share P and the image size: ONE read-back of every
// view's num_rendered, then binning, per-tile sort, rect lists and the blend, each as one launch over all views.
// rendered[v] receives num_rendered of view v.
static int forward_finish(int n, const das3r_raster_fwd_view* const* views, int* rendered, cudaStream_t stream) {
    const das3r_raster_params* params = views[0]->params;
    const int P = params->P, W = params->width, H = params->height;
    BatchArgs b;
    batch_header(b, n, P, W, H);
    const int T = b.T;
    for (int v = 0; v < n; v++)
        bin_fill_state(b.v[v], P, W, H, views[v]->state->geom, views[v]->state->image, views[v]->out->radii);

    // The one host round trip of the forward (the reference has one per view, rasterizer_impl.cu:281). The counters of
    // all views land in a small page-locked buffer (one per host thread, allocated on first use and kept): copies into
    // pageable memory would each block the host on their own.
    static thread_local uint32_t(*pinned_totals)[4] = nullptr;
    if (pinned_totals == nullptr) {
        void* p = nullptr;
        if (cudaHostAlloc(&p, sizeof(uint32_t) * 4 * RB_MAX_VIEWS, cudaHostAllocDefault) == cudaSuccess)
            pinned_totals = static_cast<uint32_t(*)[4]>(p);
        else
            (void)cudaGetLastError();  // fall back to the stack buffer below
    }
    uint32_t stack_totals[RB_MAX_VIEWS][4];
    uint32_t(*host_totals)[4] = pinned_totals ? pinned_totals : stack_totals;
    cudaError_t e = cudaSuccess;
    for (int v = 0; v < n && e == cudaSuccess; v++)
        e = cudaMemcpyAsync(host_totals[v], b.v[v].totals, sizeof(uint32_t) * 4, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "read-back of num_rendered", e);

    int R[RB_MAX_VIEWS];
    uint32_t view_max[RB_MAX_VIEWS];
    void* temp[RB_MAX_VIEWS];
    size_t temp_bytes[RB_MAX_VIEWS];
    uint32_t max_count = 0;
    long long R_total = 0;
    bool any_fallback = false;
    for (int v = 0; v < n; v++) {
        if (host_totals[v][0] > 0x7fffffffu) return fail_msg(DAS3R_ELIMIT, "forward: more than 2^31-1 tile instances");
        R[v] = (int)host_totals[v][0];
        view_max[v] = host_totals[v][1];
        max_count = std::max(max_count, view_max[v]);
        R_total += R[v];
        das3r_raster_state* state = views[v]->state;
        state->num_rendered = R[v];
        rendered[v] = R[v];
        const bool fallback = view_max[v] > (uint32_t)tile_sort_capacity();
        any_fallback = any_fallback || fallback;
        das3r_raster_state_layout L;
        memset(&L, 0, sizeof(L));
        BinExtra X;
        binning_layout(R[v], T, views[v]->params->debug != 0, fallback, &L, &X);
        char* bin = views[v]->binning_alloc(L.binning_bytes, views[v]->binning_ctx);
        if (!bin && L.binning_bytes > 0) return fail_msg(DAS3R_ENOMEM, "forward: binning allocator returned NULL");
        state->binning = bin;
        state->binning_bytes = L.binning_bytes;
        bin_fill_binning(b.v[v], L, bin, views[v]->params->debug != 0);
        temp[v] = fallback ? bin + X.b_fallback : nullptr;
        temp_bytes[v] = X.fallback_bytes;
        b.v[v].bg = views[v]->params->background;
        b.v[v].out_color = views[v]->out->out_color;
        b.v[v].out_depth = views[v]->out->out_depth;
        b.v[v].out_alpha = views[v]->out->out_alpha;
    }
    if (R_total > 0) {
        STAGE(launch_scatter(b, stream), "scatter", ST_SCATTER, 1);
        STAGE(launch_tile_sort(b, max_count, R, view_max, temp, temp_bytes, stream), "tile_sort", ST_TILE_SORT,
              1 + (max_count > 4096 ? 1 : 0) + (max_count > 8192 ? 1 : 0) + (any_fallback ? 3 : 0));
        STAGE(launch_rect_lists(b, stream), "rect_lists", ST_RECT_LISTS, 1);
    }
    STAGE(launch_blend_fwd(b, stream), "blend_fwd", ST_BLEND_FWD, 1);
    return DAS3R_OK;
}

int das3r_raster_forward(const das3r_raster_params* params, const das3r_raster_inputs* in,
                         const das3r_raster_outputs* out, das3r_alloc_fn geom_alloc, void* geom_ctx,
                         das3r_alloc_fn binning_alloc, void* binning_ctx, das3r_alloc_fn image_alloc, void* image_ctx,
                         das3r_raster_state* state, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!geom_alloc || !binning_alloc || !image_alloc) return fail_msg(DAS3R_EINVAL, "forward: null allocator");
    int rc = forward_check("forward", params, in, out, state, stream);
    if (rc != DAS3R_OK) return rc < 0 ? rc : 0;
    PreFwdArgs a;
    rc = forward_begin(params, in, out, geom_alloc, geom_ctx, image_alloc, image_ctx, state, stream, &a);
    if (rc != DAS3R_OK) return rc;
    static const bool stage_sh = getenv("DAS3R_NO_SH_STAGE") == nullptr;
    STAGE(launch_preprocess_fwd(a, stage_sh, stream), "preprocess_fwd", ST_PREPROCESS_FWD, 1);
    das3r_raster_fwd_view fv;
    memset(&fv, 0, sizeof(fv));
    fv.params = params; fv.in = in; fv.out = out; fv.state = state;
    fv.binning_alloc = binning_alloc; fv.binning_ctx = binning_ctx;
    const das3r_raster_fwd_view* one = &fv;
    rc = forward_scan(1, &one, stream);
    if (rc != DAS3R_OK) return rc;
    int rendered = 0;
    rc = forward_finish(1, &one, &rendered, stream);
    return rc != DAS3R_OK ? rc : rendered;
}

int das3r_raster_forward_preprocess(int32_t n_views, const das3r_raster_fwd_view* views, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (n_views < 1 || !views) return fail_msg(DAS3R_EINVAL, "forward_preprocess: null or invalid argument");
    std::vector<PreFwdArgs> args((size_t)n_views);
    std::vector<char> live((size_t)n_views, 0);
    const das3r_raster_params* params = nullptr;  // for the STAGE macro
    for (int v = 0; v < n_views; v++) {
        const das3r_raster_fwd_view& V = views[v];
        if (!V.geom_alloc || !V.image_alloc) return fail_msg(DAS3R_EINVAL, "forward_preprocess: null allocator");
        int rc = forward_check("forward_preprocess", V.params, V.in, V.out, V.state, stream);
        if (rc < 0) return rc;
        if (rc == 1) continue;  // P == 0: nothing to preprocess
        rc = forward_begin(V.params, V.in, V.out, V.geom_alloc, V.geom_ctx, V.image_alloc, V.image_ctx, V.state, stream,
                           &args[v]);
        if (rc != DAS3R_OK) return rc;
        live[v] = 1;
        params = V.params;
    }
    if (!params) return DAS3R_OK;
    static const bool stage_sh = getenv("DAS3R_NO_SH_STAGE") == nullptr;
    // Views that can share one pass: same SH rows, same P and image size. Degree 3 stages the 16-coefficient rows
    // through shared memory (needs 16-byte alignment); lower degrees read their few active floats per thread.
    auto fusable = [&](int v) {
        const PreFwdArgs& a = args[v];
        if (!live[v] || !stage_sh || !a.shs || a.colors_precomp) return false;
        if (a.D == 3) return a.M == 16 && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
        return true;
    };
    std::vector<char> done((size_t)n_views, 0);
    for (int v0 = 0; v0 < n_views; v0++) {
        if (!live[v0] || done[v0]) continue;
        if (!fusable(v0)) {
            params = views[v0].params;
            STAGE(launch_preprocess_fwd(args[v0], stage_sh, stream), "preprocess_fwd", ST_PREPROCESS_FWD, 1);
            done[v0] = 1;
            continue;
        }
        MultiFwdArgs m;
        memset(&m, 0, sizeof(m));
        m.T = (int)(args[v0].gx * args[v0].gy);
        for (int v = v0; v < n_views && m.n_views < MF_MAX_VIEWS; v++) {
            if (done[v] || !fusable(v)) continue;
            const PreFwdArgs& a = args[v];
            if (a.shs != args[v0].shs || a.P != args[v0].P || a.W != args[v0].W || a.H != args[v0].H ||
                a.D != args[v0].D || a.M != args[v0].M)
                continue;
            m.v[m.n_views++] = a;
            done[v] = 1;
        }
        params = views[v0].params;
        if (m.n_views == 1)
            STAGE(launch_preprocess_fwd(m.v[0], stage_sh, stream), "preprocess_fwd", ST_PREPROCESS_FWD, 1);
        else
            STAGE(launch_preprocess_fwd_multi(m, stream), "preprocess_fwd_multi", ST_PREPROCESS_FWD, 1);
    }
    // tile ranges: one launch per group of views with the same P and image size
    std::fill(done.begin(), done.end(), 0);
    for (int v0 = 0; v0 < n_views; v0++) {
        if (!live[v0] || done[v0]) continue;
        const das3r_raster_fwd_view* grp[RB_MAX_VIEWS];
        int n = 0;
        for (int v = v0; v < n_views && n < RB_MAX_VIEWS; v++) {
            if (!live[v] || done[v]) continue;
            if (views[v].params->P != views[v0].params->P || views[v].params->width != views[v0].params->width ||
                views[v].params->height != views[v0].params->height)
                continue;
            grp[n++] = &views[v];
            done[v] = 1;
        }
        int rc = forward_scan(n, grp, stream);
        if (rc != DAS3R_OK) return rc;
    }
    return DAS3R_OK;
}

static int render_view_check(const das3r_raster_fwd_view* view) {
    if (!view || !check_params(view->params) || !view->out || !view->state || !view->binning_alloc)
        return fail_msg(DAS3R_EINVAL, "forward_render: null or invalid argument");
    if (view->params->P == 0) return 1;  // the preprocess half has zeroed the image
    if (!view->state->geom || !view->state->image || !view->out->out_color || !view->out->radii)
        return fail_msg(DAS3R_EINVAL, "forward_render: the view has not been through das3r_raster_forward_preprocess");
    return DAS3R_OK;
}

int das3r_raster_forward_render(const das3r_raster_fwd_view* view, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    int rc = render_view_check(view);
    if (rc != DAS3R_OK) return rc < 0 ? rc : 0;
    int rendered = 0;
    rc = forward_finish(1, &view, &rendered, stream);
    return rc != DAS3R_OK ? rc : rendered;
}

int das3r_raster_forward_render_batch(int32_t n_views, const das3r_raster_fwd_view* views, int32_t* num_rendered,
                                      void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (n_views < 1 || !views || !num_rendered)
        return fail_msg(DAS3R_EINVAL, "forward_render_batch: null or invalid argument");
    std::vector<char> todo((size_t)n_views, 0);
    for (int v = 0; v < n_views; v++) {
        num_rendered[v] = 0;
        int rc = render_view_check(&views[v]);
        if (rc < 0) return rc;
        todo[v] = rc == DAS3R_OK;
    }
    // groups of views with the same P and image size (normally: all of them) share their launches
    for (int v0 = 0; v0 < n_views; v0++) {
        if (!todo[v0]) continue;
        const das3r_raster_fwd_view* grp[RB_MAX_VIEWS];
        int idx[RB_MAX_VIEWS];
        int n = 0;
        for (int v = v0; v < n_views && n < RB_MAX_VIEWS; v++) {
            if (!todo[v]) continue;
            if (views[v].params->P != views[v0].params->P || views[v].params->width != views[v0].params->width ||
                views[v].params->height != views[v0].params->height || views[v].params->debug != views[v0].params->debug)
                continue;
            grp[n] = &views[v];
            idx[n++] = v;
            todo[v] = 0;
        }
        int rendered[RB_MAX_VIEWS];
        int rc = forward_finish(n, grp, rendered, stream);
        if (rc != DAS3R_OK) return rc;
        for (int i = 0; i < n; i++) num_rendered[idx[i]] = rendered[i];
    }
    return DAS3R_OK;
}

size_t das3r_raster_backward_scratch_bytes(int32_t P) {
    return align_up((size_t)(P > 0 ? P : 0) * ACC_STRIDE * sizeof(float), 256);
}

// The two halves of the backward, shared by das3r_raster_backward and the split / multi-view entry points.
// First half for n <= RB_MAX_VIEWS views that share P and the image size: every view's accumulators are cleared, then
// ONE launch runs the backward blend of all views.
struct BlendView {
    const das3r_raster_params* params;
    const das3r_raster_state* state;
    const float* dL_dpix;
    float* scratch;
};

static int backward_blend_impl(const char* who, int n, const BlendView* views, cudaStream_t stream) {
    const das3r_raster_params* params = views[0].params;  // the STAGE macro reads params->debug
    const int P = params->P, W = params->width, H = params->height;
    BatchArgs b;
    batch_header(b, n, P, W, H);
    const int T = b.T;
    long long R_total = 0;
    for (int v = 0; v < n; v++) {
        const BlendView& V = views[v];
        const das3r_raster_state* state = V.state;
        if (!V.dL_dpix || !state->geom || !state->image || !V.scratch)
            return fail_msg(DAS3R_EINVAL, std::string(who) + ": null buffer");
        if ((reinterpret_cast<uintptr_t>(V.scratch) & 15) != 0)
            return fail_msg(DAS3R_EINVAL, std::string(who) + ": the accumulator scratch must be 16-byte aligned");
        const int R = state->num_rendered;
        if (R < 0) return fail_msg(DAS3R_EINVAL, std::string(who) + ": negative num_rendered");
        if (R > 0 && !state->binning) return fail_msg(DAS3R_EINVAL, std::string(who) + ": binning blob is null");
        das3r_raster_state_layout L;
        memset(&L, 0, sizeof(L));
        geom_layout(P, &L);
        image_layout(W, H, &L);
        BinExtra X;
        binning_layout(R, T, V.params->debug != 0, false, &L, &X);
        if (state->geom_bytes < L.geom_bytes || state->image_bytes < L.image_bytes ||
            (R > 0 && state->binning_bytes < L.binning_bytes))
            return fail_msg(DAS3R_EINVAL, std::string(who) + ": state blobs are smaller than the forward's layout for "
                                                              "this num_rendered");
        bin_fill_state(b.v[v], P, W, H, state->geom, state->image, nullptr);
        if (R > 0) bin_fill_binning(b.v[v], L, state->binning, V.params->debug != 0);
        b.v[v].bg = V.params->background;
        b.v[v].dL_dpix = V.dL_dpix;
        b.v[v].acc = V.scratch;
        R_total += R;
        STAGE(cudaMemsetAsync(V.scratch, 0, (size_t)P * ACC_STRIDE * sizeof(float), stream), "memset(accumulators)",
              ST_MEMSET, 0);
    }
    if (R_total > 0) STAGE(launch_blend_bwd(b, stream), "blend_bwd", ST_BLEND_BWD, 1);
    return DAS3R_OK;
}

static void fill_prebwd_args(PreBwdArgs& a, const das3r_raster_params* params, const das3r_raster_inputs* in,
                             const int32_t* radii, const uint8_t* clamped, const float* acc,
                             const das3r_raster_grads* g, int accumulate) {
    memset(&a, 0, sizeof(a));
    const int W = params->width, H = params->height;
    a.P = params->P; a.D = params->D; a.M = in->shs ? params->M : 0; a.W = W; a.H = H;
    a.tan_fovx = params->tan_fovx; a.tan_fovy = params->tan_fovy;
    a.focal_y = H / (2.0f * params->tan_fovy);
    a.focal_x = W / (2.0f * params->tan_fovx);
    a.scale_mod = params->scale_modifier;
    a.means = in->means3D; a.shs = in->shs; a.scales = in->scales; a.rots = in->rotations;
    a.cov3D_precomp = in->cov3D_precomp;
    a.view = params->viewmatrix; a.proj = params->projmatrix; a.campos = params->campos;
    a.radii = radii; a.clamped = clamped; a.acc = acc;
    a.dL_dmeans2D = g->dL_dmeans2D; a.dL_dcolors = g->dL_dcolors; a.dL_dopacity = g->dL_dopacity;
    a.dL_dmeans3D = g->dL_dmeans3D; a.dL_dcov3D = g->dL_dcov3D; a.dL_dsh = in->shs ? g->dL_dsh : nullptr;
    a.dL_dscales = g->dL_dscales; a.dL_drots = g->dL_drotations;
    a.accumulate = accumulate;
}

// The kernels read rotations and write dL_drotations / the accumulators with 128-bit accesses.
static bool misaligned16(const void* p) { return p != nullptr && (reinterpret_cast<uintptr_t>(p) & 15) != 0; }

int das3r_raster_backward(const das3r_raster_params* params, const das3r_raster_inputs* in, const int32_t* radii,
                          const das3r_raster_state* state, const float* dL_dpix, const das3r_raster_grads* g,
                          void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!check_params(params) || !in || !state || !g) return fail_msg(DAS3R_EINVAL, "backward: null or invalid argument");
    const int P = params->P;
    if (P == 0) return DAS3R_OK;
    if (!radii || !g->dL_dopacity || !g->dL_dmeans3D || !g->dL_dscales || !g->dL_drotations)
        return fail_msg(DAS3R_EINVAL, "backward: null buffer");
    if (in->shs && !g->dL_dsh) return fail_msg(DAS3R_EINVAL, "backward: dL_dsh is null although shs is given");
    if (misaligned16(in->rotations) || misaligned16(g->dL_drotations) || misaligned16(g->scratch))
        return fail_msg(DAS3R_EINVAL, "backward: rotations, dL_drotations and scratch must be 16-byte aligned");
    BlendView bv{params, state, dL_dpix, g->scratch};
    int rc = backward_blend_impl("backward", 1, &bv, stream);
    if (rc != DAS3R_OK) return rc;
    das3r_raster_state_layout L;
    memset(&L, 0, sizeof(L));
    geom_layout(P, &L);
    const uint8_t* clamped = reinterpret_cast<const uint8_t*>(state->geom + L.g_clamped);
    PreBwdArgs a;
    fill_prebwd_args(a, params, in, radii, clamped, g->scratch, g, g->accumulate);
    STAGE(launch_preprocess_bwd(a, stream), "preprocess_bwd", ST_PREPROCESS_BWD, 1);
    return DAS3R_OK;
}

// ---- split / multi-view backward (see include/das3r_raster.h) ----------------------------------------------------
int das3r_raster_backward_blend(const das3r_raster_view* view, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!view || !check_params(view->params) || !view->state)
        return fail_msg(DAS3R_EINVAL, "backward_blend: null or invalid argument");
    if (view->params->P == 0) return DAS3R_OK;
    BlendView bv{view->params, view->state, view->dL_dpix, view->scratch};
    return backward_blend_impl("backward_blend", 1, &bv, stream);
}

int das3r_raster_backward_blend_batch(int32_t n_views, const das3r_raster_view* views, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (n_views < 1 || !views) return fail_msg(DAS3R_EINVAL, "backward_blend_batch: null or invalid argument");
    for (int v = 0; v < n_views; v++)
        if (!check_params(views[v].params) || !views[v].state)
            return fail_msg(DAS3R_EINVAL, "backward_blend_batch: null or invalid view");
    std::vector<char> todo((size_t)n_views, 1);
    for (int v0 = 0; v0 < n_views; v0++) {
        if (!todo[v0] || views[v0].params->P == 0) continue;
        BlendView grp[RB_MAX_VIEWS];
        int n = 0;
        for (int v = v0; v < n_views && n < RB_MAX_VIEWS; v++) {
            const das3r_raster_params* p = views[v].params;
            if (!todo[v] || p->P != views[v0].params->P || p->width != views[v0].params->width ||
                p->height != views[v0].params->height || p->debug != views[v0].params->debug)
                continue;
            grp[n++] = BlendView{p, views[v].state, views[v].dL_dpix, views[v].scratch};
            todo[v] = 0;
        }
        int rc = backward_blend_impl("backward_blend_batch", n, grp, stream);
        if (rc != DAS3R_OK) return rc;
    }
    return DAS3R_OK;
}

int das3r_raster_backward_finish_ex(int32_t n_views, const das3r_raster_view* views, const das3r_raster_grads* g,
                                    int32_t dsh_coeffs, int32_t first, int32_t count, int32_t grads_local,
                                    void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (n_views < 1 || !views || !g) return fail_msg(DAS3R_EINVAL, "backward_finish: null or invalid argument");
    for (int v = 0; v < n_views; v++) {
        const das3r_raster_view& V = views[v];
        if (!check_params(V.params) || !V.in || !V.state)
            return fail_msg(DAS3R_EINVAL, "backward_finish: null or invalid view");
        if (V.params->P != views[0].params->P || V.params->D != views[0].params->D || V.params->M != views[0].params->M)
            return fail_msg(DAS3R_EINVAL, "backward_finish: the views differ in P, D or M");
        if ((V.in->shs != nullptr) != (views[0].in->shs != nullptr))
            return fail_msg(DAS3R_EINVAL, "backward_finish: the views differ in their colour input");
    }
    const das3r_raster_params* params = views[0].params;  // for the STAGE macro
    const int P = params->P, D = params->D;
    const int M = views[0].in->shs ? params->M : 0;
    if (dsh_coeffs <= 0) dsh_coeffs = M;
    if (first < 0) return fail_msg(DAS3R_EINVAL, "backward_finish: negative first");
    if (count < 0) count = P - first;
    if (count < 0 || (long long)first + count > P)
        return fail_msg(DAS3R_EINVAL, "backward_finish: Gaussian range outside [0, P)");
    if (M > 0 && (dsh_coeffs < (D + 1) * (D + 1) || dsh_coeffs > M))
        return fail_msg(DAS3R_EINVAL, "backward_finish: dsh_coeffs must satisfy (D+1)^2 <= dsh_coeffs <= M");
    if (P == 0 || count == 0) return DAS3R_OK;
    for (int v = 0; v < n_views; v++)
        if (!views[v].radii || !views[v].scratch || !views[v].state->geom)
            return fail_msg(DAS3R_EINVAL, "backward_finish: null buffer in a view");
    if (!g->dL_dopacity || !g->dL_dmeans3D || !g->dL_dscales || !g->dL_drotations)
        return fail_msg(DAS3R_EINVAL, "backward_finish: null gradient buffer");
    if (views[0].in->shs && !g->dL_dsh)
        return fail_msg(DAS3R_EINVAL, "backward_finish: dL_dsh is null although shs is given");
    if (n_views > 1 && (g->dL_dmeans2D || g->dL_dcolors || g->dL_dcov3D))
        return fail_msg(DAS3R_EINVAL, "backward_finish: dL_dmeans2D / dL_dcolors / dL_dcov3D are per-view quantities; pass "
                                      "NULL for them when summing several views");
    das3r_raster_state_layout L;
    memset(&L, 0, sizeof(L));
    geom_layout(P, &L);
    for (int v = 0; v < n_views; v++) {
        if (views[v].state->geom_bytes < L.geom_bytes)
            return fail_msg(DAS3R_EINVAL, "backward_finish: geom blob is smaller than the forward's layout");
        if (misaligned16(views[v].in->rotations) || misaligned16(views[v].dL_drotations) || misaligned16(views[v].scratch))
            return fail_msg(DAS3R_EINVAL, "backward_finish: rotations, per-view dL_drotations and the accumulators must be "
                                          "16-byte aligned");
    }
    if (misaligned16(g->dL_drotations))
        return fail_msg(DAS3R_EINVAL, "backward_finish: dL_drotations must be 16-byte aligned");

    // One fused pass when the views share 16-coefficient SH rows and use scale / rotation inputs; otherwise one
    // accumulating pass per view, which gives the same sums (full rows only).
    const size_t o = (size_t)first;  // every pointer below is moved to the first Gaussian of the range
    const size_t og = grads_local ? 0 : o;  // ... except the summed outputs when they already address that Gaussian
    bool fused = M == 16 && views[0].in->shs != nullptr && !misaligned16(views[0].in->shs) &&
                 (n_views > 1 || dsh_coeffs != M);
    for (int v = 0; v < n_views && fused; v++) {
        const das3r_raster_inputs* in = views[v].in;
        fused = in->shs == views[0].in->shs && in->scales && in->rotations && !in->cov3D_precomp && !in->colors_precomp;
    }
    if (fused) {
        for (int v0 = 0; v0 < n_views; v0 += MV_MAX_VIEWS) {
            MultiBwdArgs a;
            memset(&a, 0, sizeof(a));
            a.P = count;
            a.n_views = std::min(MV_MAX_VIEWS, n_views - v0);
            a.accumulate_shared = (v0 == 0) ? g->accumulate : 1;
            a.accumulate_view = g->accumulate;
            a.D = D;
            a.dsh_coeffs = dsh_coeffs;
            a.shs = views[0].in->shs + o * 48;
            a.dL_dopacity = g->dL_dopacity + og; a.dL_dmeans3D = g->dL_dmeans3D + og * 3;
            a.dL_dsh = g->dL_dsh + og * 3 * dsh_coeffs;
            a.dL_dscales = g->dL_dscales + og * 3; a.dL_drots = g->dL_drotations + og * 4;
            for (int i = 0; i < a.n_views; i++) {
                const das3r_raster_view& V = views[v0 + i];
                MultiBwdView& m = a.v[i];
                m.means = V.in->means3D + o * 3; m.scales = V.in->scales + o * 3; m.rots = V.in->rotations + o * 4;
                m.view = V.params->viewmatrix; m.proj = V.params->projmatrix; m.campos = V.params->campos;
                m.radii = V.radii + o;
                m.clamped = reinterpret_cast<const uint8_t*>(V.state->geom + L.g_clamped) + o;
                m.acc = V.scratch + o * ACC_STRIDE;
                m.dL_dmeans3D = V.dL_dmeans3D ? V.dL_dmeans3D + o * 3 : nullptr;
                m.dL_drots = V.dL_drotations ? V.dL_drotations + o * 4 : nullptr;
                m.tan_fovx = V.params->tan_fovx; m.tan_fovy = V.params->tan_fovy;
                m.focal_x = V.params->width / (2.0f * V.params->tan_fovx);
                m.focal_y = V.params->height / (2.0f * V.params->tan_fovy);
                m.scale_mod = V.params->scale_modifier;
            }
            STAGE(launch_preprocess_bwd_multi(a, stream), "preprocess_bwd_multi", ST_PREPROCESS_BWD, 1);
        }
        return DAS3R_OK;
    }
    if (dsh_coeffs != M)
        return fail_msg(DAS3R_EINVAL, "backward_finish: dL_dsh rows trimmed to the active bands need the fused path (views "
                                      "sharing 16-coefficient, 16-byte aligned shs with scale / rotation inputs)");
    // The one-view kernel has one accumulate switch for all its outputs. A view with outputs of its own (they follow
    // grads->accumulate) and the summed outputs (which accumulate from the second view on) can only share a launch
    // if both agree; otherwise the view's own outputs are cleared first and the launch accumulates.
    for (int v = 0; v < n_views; v++) {
        const das3r_raster_view& V = views[v];
        das3r_raster_inputs in = *V.in;
        in.means3D += o * 3;
        if (in.shs) in.shs += o * 3 * M;
        if (in.scales) in.scales += o * 3;
        if (in.rotations) in.rotations += o * 4;
        if (in.cov3D_precomp) in.cov3D_precomp += o * 6;
        das3r_raster_grads gg = *g;
        gg.dL_dopacity += og; gg.dL_dmeans3D += og * 3; gg.dL_dscales += og * 3; gg.dL_drotations += og * 4;
        if (gg.dL_dsh) gg.dL_dsh += og * 3 * M;
        if (gg.dL_dmeans2D) gg.dL_dmeans2D += og * 3;
        if (gg.dL_dcolors) gg.dL_dcolors += og * 3;
        if (gg.dL_dcov3D) gg.dL_dcov3D += og * 6;
        das3r_raster_params pp = *V.params;
        pp.P = count;
        PreBwdArgs a;
        fill_prebwd_args(a, &pp, &in, V.radii + o, reinterpret_cast<const uint8_t*>(V.state->geom + L.g_clamped) + o,
                         V.scratch + o * ACC_STRIDE, &gg, v == 0 ? g->accumulate : 1);
        if (V.dL_dmeans3D || V.dL_drotations) {
            float* vm = V.dL_dmeans3D ? V.dL_dmeans3D + o * 3 : nullptr;
            float* vr = V.dL_drotations ? V.dL_drotations + o * 4 : nullptr;
            if (a.accumulate != g->accumulate) {  // the launch accumulates, the view's own outputs must be overwritten
                if (vm) STAGE(cudaMemsetAsync(vm, 0, (size_t)count * 3 * sizeof(float), stream), "memset(view dL_dmeans3D)", ST_MEMSET, 0);
                if (vr) STAGE(cudaMemsetAsync(vr, 0, (size_t)count * 4 * sizeof(float), stream), "memset(view dL_drotations)", ST_MEMSET, 0);
            }
            // the summed output this view does not feed still has to be defined
            if (!a.accumulate && vm)
                STAGE(cudaMemsetAsync(gg.dL_dmeans3D, 0, (size_t)count * 3 * sizeof(float), stream), "memset(dL_dmeans3D)", ST_MEMSET, 0);
            if (!a.accumulate && vr)
                STAGE(cudaMemsetAsync(gg.dL_drotations, 0, (size_t)count * 4 * sizeof(float), stream), "memset(dL_drotations)", ST_MEMSET, 0);
            if (vm) a.dL_dmeans3D = vm;
            if (vr) a.dL_drots = vr;
        }
        STAGE(launch_preprocess_bwd(a, stream), "preprocess_bwd", ST_PREPROCESS_BWD, 1);
    }
    return DAS3R_OK;
}

int das3r_raster_backward_finish(int32_t n_views, const das3r_raster_view* views, const das3r_raster_grads* g,
                                 void* stream_v) {
    return das3r_raster_backward_finish_ex(n_views, views, g, 0, 0, -1, 0, stream_v);
}

// ---- backward without any device allocation (see include/das3r_raster.h) ---------------------------------------
// The gradient accumulators live in this fixed array: one copy per GPU that loads the library, held until exit.
// In host code the name below is only a handle: kernels must be given the address cudaGetSymbolAddress returns.
static __device__ __align__(256) float g_static_accumulators[(size_t)DAS3R_STATIC_MAX_GAUSSIANS * ACC_STRIDE];

int das3r_raster_backward_static(const das3r_raster_params* params, const das3r_raster_inputs* in,
                                 const int32_t* radii, const das3r_raster_state* state, const float* dL_dpix,
                                 const das3r_raster_grads* grads, void* stream_v) {
    g_err.clear();
    if (!params || !grads) return fail_msg(DAS3R_EINVAL, "backward_static: null argument");
    if (params->P > DAS3R_STATIC_MAX_GAUSSIANS) {
        char buf[200];
        snprintf(buf, sizeof(buf),
                 "backward_static: %d Gaussians exceed the built-in scratch of %d (use das3r_raster_backward with "
                 "caller-provided scratch)",
                 params->P, DAS3R_STATIC_MAX_GAUSSIANS);
        return fail_msg(DAS3R_ELIMIT, buf);
    }
    // One-time set-up per device, with that device current: resolve the array's device address.
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static float* device_ptr[kMaxDevices] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "backward_static: cudaGetDevice", e);
    if (dev < 0 || dev >= kMaxDevices) return fail_msg(DAS3R_ELIMIT, "backward_static: device ordinal out of range");
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (device_ptr[dev] == nullptr) {
            float* ptr = nullptr;
            e = cudaGetSymbolAddress((void**)&ptr, g_static_accumulators);
            if (e != cudaSuccess || ptr == nullptr)
                return fail(DAS3R_ECUDA, "backward_static: cudaGetSymbolAddress", e);
            device_ptr[dev] = ptr;
        }
        scratch = device_ptr[dev];
    }
    das3r_raster_grads g = *grads;
    g.scratch = scratch;  // the regular backward clears the P*48 bytes it uses before accumulating into them
    return das3r_raster_backward(params, in, radii, state, dL_dpix, &g, stream_v);
}

int das3r_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, void* stream_v) {
    g_err.clear();
    if (P < 0) return fail_msg(DAS3R_EINVAL, "mark_visible: P < 0");
    if (P == 0) return DAS3R_OK;
    if (!means3D || !viewmatrix || !projmatrix || !present) return fail_msg(DAS3R_EINVAL, "mark_visible: null pointer");
    g_launches += 1;
    cudaError_t e =
        launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, static_cast<cudaStream_t>(stream_v));
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "mark_visible", e);
    return DAS3R_OK;
}

size_t das3r_knn_scratch_bytes(int32_t P) { return knn_scratch_bytes(P); }

int das3r_knn_dist2(int32_t P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes,
                    void* stream_v) {
    g_err.clear();
    if (P < 0) return fail_msg(DAS3R_EINVAL, "knn: P < 0");
    if (P == 0) return DAS3R_OK;
    if (!points || !mean_dist2 || !scratch) return fail_msg(DAS3R_EINVAL, "knn: null pointer");
    if (scratch_bytes < knn_scratch_bytes(P)) return fail_msg(DAS3R_EINVAL, "knn: scratch too small");
    g_launches += 7;  // init, bbox, morton, (library sort), gather, boxes, search
    cudaError_t e = knn_dist2(P, points, mean_dist2, scratch, scratch_bytes, static_cast<cudaStream_t>(stream_v));
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "knn", e);
    return DAS3R_OK;
}


// ---- fused DAS3R render front-end (include/das3r_frontend.h) ------------------------------------------------------
static bool frontend_model_ok(const das3r_frontend_model* m) {
    return m && m->P >= 0 &&
           (m->P == 0 || (m->xyz && m->rotation && m->opacity && m->scaling && m->features_dc && m->features_rest &&
                          m->conf_static && ((reinterpret_cast<uintptr_t>(m->rotation) & 15) == 0)));
}

static int frontend_rows(const das3r_frontend_model* model, const das3r_frontend_options* opt) {
    return (opt && opt->filter_index) ? opt->n_rows : model->P;
}

int das3r_frontend_forward_ex(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                              const das3r_frontend_outputs* out, const das3r_frontend_options* opt, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!frontend_model_ok(model) || !out || n_views < 1 || n_views > DAS3R_FRONTEND_MAX_VIEWS || !poses)
        return fail_msg(DAS3R_EINVAL, "frontend_forward: null or invalid argument (1..16 views, rotation 16-byte aligned)");
    const int D = opt ? opt->sh_degree : 3;
    if (D < 0 || D > 3) return fail_msg(DAS3R_EINVAL, "frontend_forward: sh_degree must be 0..3");
    const int rows = frontend_rows(model, opt);
    if (rows < 0) return fail_msg(DAS3R_EINVAL, "frontend_forward: negative n_rows");
    if (rows == 0) return DAS3R_OK;
    if (!out->opacity || !out->scales || !out->shs) return fail_msg(DAS3R_EINVAL, "frontend_forward: null output");
    FrontendFwdArgs a;
    memset(&a, 0, sizeof(a));
    a.P = rows; a.n_views = n_views; a.K = (D + 1) * (D + 1);
    a.xyz = model->xyz; a.rotation = model->rotation; a.opacity_raw = model->opacity; a.scaling_raw = model->scaling;
    a.features_dc = model->features_dc; a.features_rest = model->features_rest; a.conf_static = model->conf_static;
    a.conf_index = model->conf_index; a.poses = poses;
    a.filter_index = opt ? opt->filter_index : nullptr;
    a.opacity = out->opacity; a.scales = out->scales; a.shs = out->shs;
    for (int v = 0; v < n_views; v++) {
        if (!out->means3D[v] || !out->rotations[v] || (reinterpret_cast<uintptr_t>(out->rotations[v]) & 15))
            return fail_msg(DAS3R_EINVAL, "frontend_forward: null or misaligned per-view output");
        a.means3D[v] = out->means3D[v];
        a.rotations[v] = out->rotations[v];
    }
    cudaError_t e;
    {
        StageScope scope_(ST_FRONTEND_FWD, stream, 1);
        e = launch_frontend_fwd(a, stream);
    }
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "frontend_fwd", e);
    return DAS3R_OK;
}

int das3r_frontend_forward(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                           const das3r_frontend_outputs* out, void* stream_v) {
    return das3r_frontend_forward_ex(model, n_views, poses, out, nullptr, stream_v);
}

size_t das3r_frontend_backward_scratch_bytes(int32_t P, int32_t n_views) {
    return frontend_backward_scratch_bytes(P > 0 ? P : 0, n_views > 0 ? n_views : 0);
}

int das3r_frontend_backward_ex(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                               const das3r_frontend_grads_in* gin, const das3r_frontend_grads_out* gout,
                               const das3r_frontend_options* opt, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!frontend_model_ok(model) || !gin || !gout || n_views < 1 || n_views > DAS3R_FRONTEND_MAX_VIEWS || !poses)
        return fail_msg(DAS3R_EINVAL, "frontend_backward: null or invalid argument (1..16 views)");
    if (!gout->dL_dposes) return fail_msg(DAS3R_EINVAL, "frontend_backward: dL_dposes is null");
    const int D = opt ? opt->sh_degree : 3;
    if (D < 0 || D > 3) return fail_msg(DAS3R_EINVAL, "frontend_backward: sh_degree must be 0..3");
    const int K = (D + 1) * (D + 1);
    const int dsh = (opt && opt->dsh_coeffs > 0) ? opt->dsh_coeffs : 16;
    const int rest = (opt && opt->rest_coeffs > 0) ? opt->rest_coeffs : ((opt && opt->rest_coeffs < 0) ? K - 1 : 15);
    if (dsh < K || dsh > 16 || (rest != 15 && rest != K - 1))
        return fail_msg(DAS3R_EINVAL, "frontend_backward: dsh_coeffs must be (D+1)^2..16 and rest_coeffs 15 or (D+1)^2-1");
    const int rows = frontend_rows(model, opt);
    const int first = opt ? opt->first : 0;
    int count = opt ? opt->count : -1;
    const bool finish = opt ? opt->finish_poses != 0 : true;
    if (rows < 0 || first < 0 || first > rows || (first % 256) != 0)
        return fail_msg(DAS3R_EINVAL, "frontend_backward: bad row range (first must be a multiple of 256 within the rows)");
    if (count < 0 || first + count > rows) count = rows - first;
    if (rows == 0) {
        cudaError_t e = cudaMemsetAsync(gout->dL_dposes, 0, sizeof(float) * 7 * n_views, stream);
        if (e != cudaSuccess) return fail(DAS3R_ECUDA, "frontend_backward memset", e);
        return DAS3R_OK;
    }
    if (!gin->dL_dopacity || !gin->dL_dscales || !gin->dL_dshs || !gout->dL_dxyz || !gout->dL_drotation ||
        !gout->dL_dopacity || !gout->dL_dscaling || !gout->dL_dfeatures_dc || (!gout->dL_dfeatures_rest && rest > 0) ||
        !gout->dL_dconf || !gout->scratch || (reinterpret_cast<uintptr_t>(gout->dL_drotation) & 15))
        return fail_msg(DAS3R_EINVAL, "frontend_backward: null or misaligned buffer");
    FrontendBwdArgs a;
    memset(&a, 0, sizeof(a));
    a.P = rows; a.n_views = n_views; a.K = K; a.dsh_coeffs = dsh; a.rest_coeffs = rest;
    a.row0 = first; a.row_end = first + count;
    a.out_row0 = (opt && opt->local_outputs) ? first : 0;
    if (a.out_row0 && a.filter_index)
        return fail_msg(DAS3R_EINVAL, "frontend_backward: local_outputs cannot be combined with filter_index");
    a.xyz = model->xyz; a.rotation = model->rotation; a.opacity_raw = model->opacity; a.scaling_raw = model->scaling;
    a.conf_static = model->conf_static; a.conf_index = model->conf_index; a.poses = poses;
    a.filter_index = opt ? opt->filter_index : nullptr;
    a.dL_dopacity = gin->dL_dopacity; a.dL_dscales = gin->dL_dscales; a.dL_dshs = gin->dL_dshs;
    for (int v = 0; v < n_views; v++) {
        if (!gin->dL_dmeans3D[v] || !gin->dL_drotations[v] || (reinterpret_cast<uintptr_t>(gin->dL_drotations[v]) & 15))
            return fail_msg(DAS3R_EINVAL, "frontend_backward: null or misaligned per-view gradient");
        a.dL_dmeans3D[v] = gin->dL_dmeans3D[v];
        a.dL_drotations[v] = gin->dL_drotations[v];
    }
    a.dL_dxyz = gout->dL_dxyz; a.dL_drotation = gout->dL_drotation; a.dL_dopacity_raw = gout->dL_dopacity;
    a.dL_dscaling_raw = gout->dL_dscaling; a.dL_dfeatures_dc = gout->dL_dfeatures_dc;
    a.dL_dfeatures_rest = gout->dL_dfeatures_rest; a.dL_dconf = gout->dL_dconf;
    a.partials = gout->scratch;
    cudaError_t e;
    {
        StageScope scope_(ST_FRONTEND_BWD, stream, (count > 0 ? 1 : 0) + (finish ? 1 : 0));
        e = launch_frontend_bwd(a, finish ? gout->dL_dposes : nullptr, stream);
    }
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "frontend_bwd", e);
    return DAS3R_OK;
}

int das3r_frontend_backward(const das3r_frontend_model* model, int32_t n_views, const float* poses,
                            const das3r_frontend_grads_in* gin, const das3r_frontend_grads_out* gout, void* stream_v) {
    return das3r_frontend_backward_ex(model, n_views, poses, gin, gout, nullptr, stream_v);
}


// ---- fused photometric loss (include/das3r_loss.h) ----------------------------------------------------------------
static int loss_prepare(const das3r_loss_inputs* in, const char* who, LossArgs* a) {
    if (!in || in->width <= 0 || in->height <= 0 || !in->image || !in->gt || !(in->lambda_dssim >= 0.f && in->lambda_dssim <= 1.f))
        return fail_msg(DAS3R_EINVAL, std::string(who) + ": null or invalid argument");
    // the window weights live in constant memory: uploaded once per device
    static std::mutex mu;
    static bool uploaded[64] = {false};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) return fail(DAS3R_ECUDA, who, e);
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!uploaded[dev]) {
            e = loss_upload_window();
            if (e != cudaSuccess) return fail(DAS3R_ECUDA, "loss window upload", e);
            uploaded[dev] = true;
        }
    }
    a->W = in->width; a->H = in->height; a->lambda_dssim = in->lambda_dssim;
    a->image = in->image; a->gt = in->gt; a->static_mask = in->static_mask;
    return DAS3R_OK;
}

size_t das3r_loss_state_bytes(int32_t width, int32_t height) {
    return (width > 0 && height > 0) ? loss_state_bytes(width, height) : 0;
}

int das3r_loss_forward(const das3r_loss_inputs* in, float* out, void* state, void* stream_v) {
    g_err.clear();
    LossArgs a;
    int rc = loss_prepare(in, "loss_forward", &a);
    if (rc != DAS3R_OK) return rc;
    if (!out || !state) return fail_msg(DAS3R_EINVAL, "loss_forward: null output or state");
    cudaError_t e = launch_loss_fwd(a, out, state, static_cast<cudaStream_t>(stream_v));
    g_launches += 2;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "loss_fwd", e);
    return DAS3R_OK;
}

int das3r_loss_backward(const das3r_loss_inputs* in, const void* state, const float* grad_loss, float* dL_dimage,
                        float* dL_dstatic, float* dL_dgt, void* stream_v) {
    g_err.clear();
    LossArgs a;
    int rc = loss_prepare(in, "loss_backward", &a);
    if (rc != DAS3R_OK) return rc;
    if (!state || !dL_dimage) return fail_msg(DAS3R_EINVAL, "loss_backward: null state or dL_dimage");
    cudaError_t e = launch_loss_bwd(a, state, grad_loss, dL_dimage, dL_dstatic, dL_dgt, static_cast<cudaStream_t>(stream_v));
    g_launches += 1;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "loss_bwd", e);
    return DAS3R_OK;
}


// ---- pointmap initialisation (include/das3r_init.h) ---------------------------------------------------------------
static bool pointmaps_ok(const das3r_pointmaps* in) {
    return in && in->V > 0 && in->H > 0 && in->W > 0 && in->K && in->poses && in->depth && in->confidence &&
           in->dyna_avg && in->rgb;
}

size_t das3r_pointmap_scratch_bytes(int64_t n_pixels) { return pointmap_scratch_bytes(n_pixels > 0 ? n_pixels : 0); }

int64_t das3r_pointmap_select(const das3r_pointmaps* in, uint8_t* aggregated_mask, float* conf_static, void* scratch,
                              size_t scratch_bytes, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!pointmaps_ok(in) || !aggregated_mask || !conf_static || !scratch)
        return fail_msg(DAS3R_EINVAL, "pointmap_select: null or invalid argument");
    const int64_t n = (int64_t)in->V * in->H * in->W;
    if (scratch_bytes < pointmap_scratch_bytes(n)) return fail_msg(DAS3R_EINVAL, "pointmap_select: scratch too small");
    cudaError_t e = launch_pointmap_select(in->confidence, in->dyna_avg, in->log_conf_thre, n, aggregated_mask, conf_static,
                                           scratch, stream);
    g_launches += 2;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_select", e);
    unsigned long long total = 0;
    e = cudaMemcpyAsync(&total, scratch, sizeof(total), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_select read-back", e);
    if (total > 0x7fffffffull) return fail_msg(DAS3R_ELIMIT, "pointmap_select: more than 2^31-1 selected pixels");
    return (int64_t)total;
}

int das3r_pointmap_init(const das3r_pointmaps* in, const uint8_t* aggregated_mask, int64_t P, const das3r_init_outputs* out,
                        void* scratch, size_t scratch_bytes, void* knn_scratch, size_t knn_bytes, void* stream_v) {
    g_err.clear();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!pointmaps_ok(in) || !aggregated_mask || !out || !scratch || P < 0 || P > 0x7fffffffLL)
        return fail_msg(DAS3R_EINVAL, "pointmap_init: null or invalid argument");
    if (P == 0) return DAS3R_OK;
    if (!out->xyz || !out->features_dc || !out->scaling || !out->rotation || !out->opacity ||
        (reinterpret_cast<uintptr_t>(out->rotation) & 15))
        return fail_msg(DAS3R_EINVAL, "pointmap_init: null output (or rotation not 16-byte aligned)");
    const int64_t n = (int64_t)in->V * in->H * in->W;
    if (scratch_bytes < pointmap_scratch_bytes(n)) return fail_msg(DAS3R_EINVAL, "pointmap_init: scratch too small");
    if (!knn_scratch || knn_bytes < knn_scratch_bytes((int)P))
        return fail_msg(DAS3R_EINVAL, "pointmap_init: knn scratch too small (das3r_knn_scratch_bytes(P))");
    // inverse_sigmoid(1/V) as torch forms it in fp32: log(x / (1 - x)) with x = (float)(1/V)  (gaussian_model.py:646-647)
    const float x = (float)(1.0 / (double)in->V);
    const float opacity_raw = logf(x / (1.0f - x));
    cudaError_t e = cudaSuccess;
    if (out->features_rest) {  // features[:, 3:, 1:] = 0 (gaussian_model.py:637)
        e = cudaMemsetAsync(out->features_rest, 0, (size_t)P * 45 * sizeof(float), stream);
        if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_init memset", e);
    }
    e = launch_pointmap_emit(in->V, in->H, in->W, in->K, in->poses, in->depth, in->rgb, aggregated_mask,
                                         scratch, out->xyz, out->features_dc, out->features_rest, out->rotation,
                                         out->opacity, out->conf_index, opacity_raw, stream);
    g_launches += 1;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_emit", e);
    // the 3-NN distances go through out->dist2 when given, else through the scaling buffer's first P floats
    float* d2 = out->dist2 ? out->dist2 : out->scaling;
    g_launches += 7;
    e = knn_dist2((int)P, out->xyz, d2, knn_scratch, knn_bytes, stream);
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_init knn", e);
    if (!out->dist2) {
        // in place: scaling[3i..3i+2] are written from dist2 = scaling[i]; walk from the back would be needed - use the
        // knn scratch's head (free again after the search) as the staging copy instead
        e = cudaMemcpyAsync(knn_scratch, d2, (size_t)P * sizeof(float), cudaMemcpyDeviceToDevice, stream);
        if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_init copy", e);
        d2 = static_cast<float*>(knn_scratch);
    }
    e = launch_pointmap_scales(P, d2, out->scaling, stream);
    g_launches += 1;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "pointmap_scales", e);
    return DAS3R_OK;
}

// ---- in-switch all-reduce of a bucket slice (include/das3r_comm.h) ------------------------------------------------
int das3r_nvls_allreduce_sum(void* multicast_base, size_t offset_floats, size_t n_floats, int32_t rank, int32_t world,
                             int32_t n_ctas, void* stream_v) {
    g_err.clear();
    if (!multicast_base || world < 1 || rank < 0 || rank >= world)
        return fail_msg(DAS3R_EINVAL, "nvls_allreduce: null multicast pointer or bad rank / world");
    if ((offset_floats & 3) || (n_floats & 3) || (reinterpret_cast<uintptr_t>(multicast_base) & 15))
        return fail_msg(DAS3R_EINVAL, "nvls_allreduce: base, offset and length must be multiples of 16 bytes");
    if (n_floats == 0) return DAS3R_OK;
    if (n_ctas <= 0) n_ctas = 16;  // link-bound from 16 CTAs on (B200 HGX, 8 ranks); leaves the SMs to the gradient kernels
    cudaError_t e = launch_nvls_allreduce(static_cast<float*>(multicast_base) + offset_floats, n_floats / 4, rank, world,
                                          n_ctas, static_cast<cudaStream_t>(stream_v));
    g_launches += 1;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "nvls_allreduce", e);
    return DAS3R_OK;
}

// ---- fused Adam step (include/das3r_adam.h) -----------------------------------------------------------------------
int das3r_adam_step(int32_t n_tensors, const das3r_adam_tensor* tensors, double beta1, double beta2, double eps,
                    int64_t step, void* stream_v) {
    g_err.clear();
    if (n_tensors < 0 || n_tensors > DAS3R_ADAM_MAX_TENSORS || (n_tensors > 0 && !tensors) || step < 1 ||
        !(beta1 >= 0.0 && beta1 < 1.0) || !(beta2 >= 0.0 && beta2 < 1.0))
        return fail_msg(DAS3R_EINVAL, "adam_step: invalid argument (at most 16 tensors, step >= 1, betas in [0,1))");
    AdamArgs a;
    memset(&a, 0, sizeof(a));
    // bias corrections in double, like the Python floats of torch/optim/adam.py
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    a.b1w = (float)(1.0 - beta1);
    a.beta2 = (float)beta2;
    a.b2w = (float)(1.0 - beta2);
    a.bc2_sqrt = (float)sqrt(bc2);
    a.eps = (float)eps;
    const int chunk = adam_chunk_elems();
    int blocks = 0, k = 0;
    for (int i = 0; i < n_tensors; i++) {
        const das3r_adam_tensor& t = tensors[i];
        if (t.n < 0) return fail_msg(DAS3R_EINVAL, "adam_step: negative element count");
        if (t.n == 0) continue;
        if (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq) return fail_msg(DAS3R_EINVAL, "adam_step: null buffer");
        AdamTensor& o = a.t[k];
        o.param = t.param; o.grad = t.grad; o.exp_avg = t.exp_avg; o.exp_avg_sq = t.exp_avg_sq; o.n = t.n;
        o.step_size = (float)(t.lr / bc1);
        o.vec4 = (((reinterpret_cast<uintptr_t>(t.param) | reinterpret_cast<uintptr_t>(t.grad) |
                    reinterpret_cast<uintptr_t>(t.exp_avg) | reinterpret_cast<uintptr_t>(t.exp_avg_sq)) & 15) == 0) ? 1 : 0;
        a.block_start[k] = blocks;
        const int64_t nb = (t.n + chunk - 1) / chunk;
        if (nb > (int64_t)0x7fffffff - blocks) return fail_msg(DAS3R_ELIMIT, "adam_step: too many elements for one launch");
        blocks += (int)nb;
        k++;
    }
    a.block_start[k] = blocks;
    a.n_tensors = k;
    cudaError_t e = launch_adam(a, blocks, static_cast<cudaStream_t>(stream_v));
    if (blocks > 0) g_launches += 1;
    if (e != cudaSuccess) return fail(DAS3R_ECUDA, "adam", e);
    return DAS3R_OK;
}

}  // extern "C"
