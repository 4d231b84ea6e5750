// Thin PyTorch binding over the C ABI (include/das3r_raster.h). The exported functions have exactly the signatures of
// the reference's pybind module (submodules/diff-gaussian-rasterization/ext.cpp:15-18, rasterize_points.h:18-67) and
// of simple_knn._C.distCUDA2 (submodules/simple-knn/ext.cpp:15-16, spatial.h:14), so the reference's Python package
// code paths work unchanged on top of them. torch is used for memory and streams only.
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <tuple>
#include <string>
#include <vector>
#include "../../include/das3r_raster.h"
#include "../../include/das3r_frontend.h"
#include "../../include/das3r_loss.h"
#include "../../include/das3r_adam.h"
#include "../../include/das3r_comm.h"
#include "../../include/das3r_init.h"

namespace {

const float* fptr(const torch::Tensor& t) {
    // empty tensor (the reference passes torch.Tensor([]) for absent inputs) -> NULL
    return t.defined() && t.numel() > 0 ? t.data_ptr<float>() : nullptr;
}

torch::Tensor prep(const torch::Tensor& t, const char* name, const torch::Device& dev) {
    if (!t.defined() || t.numel() == 0) return t;
    TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, " must be float32");
    TORCH_CHECK(t.device() == dev, name, " must be on ", dev);
    return t.contiguous();
}

struct BlobAlloc {
    torch::Tensor t;
    torch::TensorOptions opts;
    static char* fn(size_t bytes, void* ctx) {
        auto* self = static_cast<BlobAlloc*>(ctx);
        self->t = torch::empty({(long long)bytes}, self->opts);
        return reinterpret_cast<char*>(self->t.data_ptr());
    }
};

void check(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string(what) + ": " + das3r_last_error());
}

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_impl(const torch::Tensor& background, const torch::Tensor& means3D_, const torch::Tensor& colors_,
               const torch::Tensor& opacity_, const torch::Tensor& scales_, const torch::Tensor& rotations_,
               const float scale_modifier, const torch::Tensor& cov3D_precomp_, const torch::Tensor& viewmatrix_,
               const torch::Tensor& projmatrix_, const float tan_fovx, const float tan_fovy, const int image_height,
               const int image_width, const torch::Tensor& sh_, const int degree, const torch::Tensor& campos_,
               const bool prefiltered, const bool debug, const bool extras) {
    if (means3D_.ndimension() != 2 || means3D_.size(1) != 3) {
        AT_ERROR("means3D must have dimensions (num_points, 3)");  // rasterize_points.cu:57-59
    }
    TORCH_CHECK(means3D_.is_cuda(), "means3D must be a CUDA tensor: this rasterizer has no CPU path");
    const auto dev = means3D_.device();
    c10::cuda::CUDAGuard guard(dev);
    const int P = means3D_.size(0), H = image_height, W = image_width;

    auto means3D = prep(means3D_, "means3D", dev);
    auto colors = prep(colors_, "colors_precomp", dev);
    auto opacity = prep(opacity_, "opacities", dev);
    auto scales = prep(scales_, "scales", dev);
    auto rotations = prep(rotations_, "rotations", dev);
    auto cov3D = prep(cov3D_precomp_, "cov3D_precomp", dev);
    auto view = prep(viewmatrix_, "viewmatrix", dev);
    auto proj = prep(projmatrix_, "projmatrix", dev);
    auto sh = prep(sh_, "shs", dev);
    auto campos = prep(campos_, "campos", dev);
    auto bg = prep(background, "bg", dev);

    auto float_opts = means3D.options().dtype(torch::kFloat32);
    torch::Tensor out_color = torch::empty({3, H, W}, float_opts);
    torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
    torch::Tensor depth, alpha;
    if (extras) {
        depth = torch::empty({H, W}, float_opts);
        alpha = torch::empty({H, W}, float_opts);
    }
    auto byte_opts = torch::TensorOptions().dtype(torch::kByte).device(dev);
    BlobAlloc geom{torch::empty({0}, byte_opts), byte_opts}, binning{torch::empty({0}, byte_opts), byte_opts},
        img{torch::empty({0}, byte_opts), byte_opts};

    das3r_raster_params p;
    memset(&p, 0, sizeof(p));
    p.P = P;
    p.D = degree;
    p.M = (sh.defined() && sh.numel() > 0) ? (int)sh.size(1) : 0;
    p.width = W;
    p.height = H;
    p.tan_fovx = tan_fovx;
    p.tan_fovy = tan_fovy;
    p.scale_modifier = scale_modifier;
    p.prefiltered = prefiltered;
    p.debug = debug;
    p.background = fptr(bg);
    p.viewmatrix = fptr(view);
    p.projmatrix = fptr(proj);
    p.campos = fptr(campos);
    das3r_raster_inputs in;
    in.means3D = fptr(means3D);
    in.shs = fptr(sh);
    in.colors_precomp = fptr(colors);
    in.opacities = fptr(opacity);
    in.scales = fptr(scales);
    in.rotations = fptr(rotations);
    in.cov3D_precomp = fptr(cov3D);
    das3r_raster_outputs out;
    out.out_color = out_color.data_ptr<float>();
    out.radii = P > 0 ? radii.data_ptr<int>() : nullptr;
    out.out_depth = extras ? depth.data_ptr<float>() : nullptr;
    out.out_alpha = extras ? alpha.data_ptr<float>() : nullptr;
    das3r_raster_state st;
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    int rendered = das3r_raster_forward(&p, &in, &out, BlobAlloc::fn, &geom, BlobAlloc::fn, &binning, BlobAlloc::fn,
                                        &img, &st, stream);
    check(rendered, "rasterize_gaussians");
    return std::make_tuple(rendered, out_color, radii, geom.t, binning.t, img.t, depth, alpha);
}

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor> RasterizeGaussiansCUDA(
    const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
    const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
    const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
    const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
    const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
    const bool prefiltered, const bool debug) {
    auto r = rasterize_impl(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                            viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                            prefiltered, debug, false);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r), std::get<3>(r), std::get<4>(r),
                           std::get<5>(r));
}

// Additive surface: same call, also returns depth [H,W] = sum z*alpha*T and alpha [H,W] = 1-final_T.
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansExCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                         const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                         const float scale_modifier, const torch::Tensor& cov3D_precomp,
                         const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                         const float tan_fovy, const int image_height, const int image_width, const torch::Tensor& sh,
                         const int degree, const torch::Tensor& campos, const bool prefiltered, const bool debug) {
    return rasterize_impl(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                          viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                          prefiltered, debug, true);
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor>
backward_impl(const torch::Tensor& background, const torch::Tensor& means3D_,
                               const torch::Tensor& radii_, const torch::Tensor& colors_, const torch::Tensor& scales_,
                               const torch::Tensor& rotations_, const float scale_modifier,
                               const torch::Tensor& cov3D_precomp_, const torch::Tensor& viewmatrix_,
                               const torch::Tensor& projmatrix_, const float tan_fovx, const float tan_fovy,
                               const torch::Tensor& dL_dout_color_, const torch::Tensor& sh_, const int degree,
                               const torch::Tensor& campos_, const torch::Tensor& geomBuffer, const int R,
                               const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                               const std::vector<torch::Tensor>* outs, const bool accumulate) {
    TORCH_CHECK(means3D_.is_cuda(), "means3D must be a CUDA tensor: this rasterizer has no CPU path");
    const auto dev = means3D_.device();
    c10::cuda::CUDAGuard guard(dev);
    const int P = means3D_.size(0);
    const int H = dL_dout_color_.size(1);  // rasterize_points.cu:142-143
    const int W = dL_dout_color_.size(2);

    auto means3D = prep(means3D_, "means3D", dev);
    auto colors = prep(colors_, "colors_precomp", dev);
    auto scales = prep(scales_, "scales", dev);
    auto rotations = prep(rotations_, "rotations", dev);
    auto cov3D = prep(cov3D_precomp_, "cov3D_precomp", dev);
    auto view = prep(viewmatrix_, "viewmatrix", dev);
    auto proj = prep(projmatrix_, "projmatrix", dev);
    auto sh = prep(sh_, "shs", dev);
    auto campos = prep(campos_, "campos", dev);
    auto bg = prep(background, "bg", dev);
    auto dL_dout = prep(dL_dout_color_, "dL_dout_color", dev);
    auto radii = radii_.contiguous();
    TORCH_CHECK(radii.scalar_type() == torch::kInt32, "radii must be int32");

    // rasterize_points.cu:145-149 takes M from sh.size(1) only when sh.size(0) != 0; an [0, M, 3] tensor (no Gaussians)
    // must still get an [0, M, 3] gradient or autograd rejects it, so the row length is read whenever sh is 3-D
    int M = 0;
    if (sh_.defined() && sh_.dim() == 3) M = sh_.size(1);

    auto o = means3D.options();
    // torch::empty, not zeros: the backward kernels write every element (incl. the zeros of invisible Gaussians).
    torch::Tensor dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations;
    if (outs) {
        // caller-provided gradient buffers, same order as the returned tuple
        TORCH_CHECK(outs->size() == 8, "expected 8 gradient tensors");
        const int64_t want[8] = {3LL * P, 3LL * P, (int64_t)P, 3LL * P, 6LL * P, 3LL * M * P, 3LL * P, 4LL * P};
        for (int i = 0; i < 8; i++) {
            const auto& t = (*outs)[i];
            if ((i == 0 || i == 1 || i == 4) && (!t.defined() || t.numel() == 0)) continue;  // optional outputs
            TORCH_CHECK(t.is_cuda() && t.device() == dev && t.scalar_type() == torch::kFloat32 && t.is_contiguous() &&
                            t.numel() == want[i],
                        "gradient buffer ", i, " must be a contiguous float32 CUDA tensor of ", want[i], " elements");
        }
        dL_dmeans2D = (*outs)[0]; dL_dcolors = (*outs)[1]; dL_dopacity = (*outs)[2]; dL_dmeans3D = (*outs)[3];
        dL_dcov3D = (*outs)[4]; dL_dsh = (*outs)[5]; dL_dscales = (*outs)[6]; dL_drotations = (*outs)[7];
    } else {
        dL_dmeans3D = torch::empty({P, 3}, o);
        dL_dmeans2D = torch::empty({P, 3}, o);
        dL_dcolors = torch::empty({P, 3}, o);
        dL_dopacity = torch::empty({P, 1}, o);
        dL_dcov3D = torch::empty({P, 6}, o);
        dL_dsh = torch::empty({P, M, 3}, o);
        dL_dscales = torch::empty({P, 3}, o);
        dL_drotations = torch::empty({P, 4}, o);
    }

    if (P != 0) {
        torch::Tensor scratch = torch::empty({(long long)das3r_raster_backward_scratch_bytes(P)},
                                             torch::TensorOptions().dtype(torch::kByte).device(dev));
        das3r_raster_params p;
        memset(&p, 0, sizeof(p));
        p.P = P;
        p.D = degree;
        p.M = M;
        p.width = W;
        p.height = H;
        p.tan_fovx = tan_fovx;
        p.tan_fovy = tan_fovy;
        p.scale_modifier = scale_modifier;
        p.debug = debug;
        p.background = fptr(bg);
        p.viewmatrix = fptr(view);
        p.projmatrix = fptr(proj);
        p.campos = fptr(campos);
        das3r_raster_inputs in;
        memset(&in, 0, sizeof(in));
        in.means3D = fptr(means3D);
        in.shs = fptr(sh);
        in.colors_precomp = fptr(colors);
        in.scales = fptr(scales);
        in.rotations = fptr(rotations);
        in.cov3D_precomp = fptr(cov3D);
        das3r_raster_state st;
        st.geom = reinterpret_cast<char*>(geomBuffer.data_ptr());
        st.geom_bytes = geomBuffer.numel();
        st.binning = binningBuffer.numel() > 0 ? reinterpret_cast<char*>(binningBuffer.data_ptr()) : nullptr;
        st.binning_bytes = binningBuffer.numel();
        st.image = reinterpret_cast<char*>(imageBuffer.data_ptr());
        st.image_bytes = imageBuffer.numel();
        st.num_rendered = R;
        das3r_raster_grads g;
        auto optp = [](torch::Tensor& t) { return t.defined() && t.numel() > 0 ? t.data_ptr<float>() : nullptr; };
        g.dL_dmeans2D = optp(dL_dmeans2D);
        g.dL_dcolors = optp(dL_dcolors);
        g.dL_dopacity = dL_dopacity.data_ptr<float>();
        g.dL_dmeans3D = dL_dmeans3D.data_ptr<float>();
        g.dL_dcov3D = optp(dL_dcov3D);
        g.dL_dsh = M > 0 ? dL_dsh.data_ptr<float>() : nullptr;
        g.dL_dscales = dL_dscales.data_ptr<float>();
        g.dL_drotations = dL_drotations.data_ptr<float>();
        g.scratch = reinterpret_cast<float*>(scratch.data_ptr());
        g.accumulate = accumulate ? 1 : 0;
        cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(das3r_raster_backward(&p, &in, radii.data_ptr<int>(), &st, dL_dout.data_ptr<float>(), &g, stream),
              "rasterize_gaussians_backward");
    }
    return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
                           dL_drotations);
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor>
RasterizeGaussiansBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D,
                               const torch::Tensor& radii, const torch::Tensor& colors, const torch::Tensor& scales,
                               const torch::Tensor& rotations, const float scale_modifier,
                               const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                               const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                               const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                               const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                               const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug) {
    return backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                         viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R,
                         binningBuffer, imageBuffer, debug, nullptr, false);
}

// Additive: same backward, but the 8 gradients go into caller-provided tensors (order of the returned tuple), either
// overwritten or accumulated (`accumulate`): sums the views of a step in place, e.g. into one flat all-reduce bucket.
void RasterizeGaussiansBackwardIntoCUDA(const torch::Tensor& background, const torch::Tensor& means3D,
                                        const torch::Tensor& radii, const torch::Tensor& colors,
                                        const torch::Tensor& scales, const torch::Tensor& rotations,
                                        const float scale_modifier, const torch::Tensor& cov3D_precomp,
                                        const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                                        const float tan_fovx, const float tan_fovy, const torch::Tensor& dL_dout_color,
                                        const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                                        const torch::Tensor& geomBuffer, const int R,
                                        const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
                                        const bool debug, const std::vector<torch::Tensor>& outs,
                                        const bool accumulate) {
    backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                  projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer,
                  imageBuffer, debug, &outs, accumulate);
}

torch::Tensor markVisible(torch::Tensor& means3D_, torch::Tensor& viewmatrix_, torch::Tensor& projmatrix_) {
    TORCH_CHECK(means3D_.is_cuda(), "means3D must be a CUDA tensor: this rasterizer has no CPU path");
    const auto dev = means3D_.device();
    c10::cuda::CUDAGuard guard(dev);
    const int P = means3D_.size(0);
    torch::Tensor present = torch::full({P}, false, means3D_.options().dtype(at::kBool));
    if (P != 0) {
        auto means3D = prep(means3D_, "means3D", dev);
        auto view = prep(viewmatrix_, "viewmatrix", dev);
        auto proj = prep(projmatrix_, "projmatrix", dev);
        cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(das3r_mark_visible(P, fptr(means3D), fptr(view), fptr(proj),
                                 reinterpret_cast<uint8_t*>(present.data_ptr<bool>()), stream),
              "mark_visible");
    }
    return present;
}

torch::Tensor distCUDA2(const torch::Tensor& points_) {
    TORCH_CHECK(points_.is_cuda(), "points must be a CUDA tensor: distCUDA2 has no CPU path");
    const auto dev = points_.device();
    c10::cuda::CUDAGuard guard(dev);
    const int P = points_.size(0);
    auto points = prep(points_, "points", dev);
    torch::Tensor means = torch::full({P}, 0.0, points_.options().dtype(torch::kFloat32));
    if (P != 0) {
        size_t bytes = das3r_knn_scratch_bytes(P);
        torch::Tensor scratch = torch::empty({(long long)bytes}, torch::TensorOptions().dtype(torch::kByte).device(dev));
        cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
        check(das3r_knn_dist2(P, fptr(points), means.data_ptr<float>(), scratch.data_ptr(), bytes, stream),
              "distCUDA2");
    }
    return means;
}

}  // namespace

// ---- split / multi-view forward (das3r_raster_forward_preprocess / _render; no reference counterpart) ------------
// (bg, means3D, rotations, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, campos) per view
using FwdView = std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, float, float, int,
                           int, torch::Tensor>;

// First half for all views: -> per view (color, radii, geomBuffer, imgBuffer). SH colours + scale/rotation inputs only.
static std::vector<std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>> ForwardPreprocessCUDA(
    const std::vector<FwdView>& views, const torch::Tensor& opacity_, const torch::Tensor& scales_,
    const torch::Tensor& sh_, const int degree, const float scale_modifier, const bool prefiltered, const bool debug) {
    TORCH_CHECK(!views.empty(), "no views");
    TORCH_CHECK(opacity_.is_cuda(), "opacities must be a CUDA tensor: this rasterizer has no CPU path");
    const auto dev = opacity_.device();
    c10::cuda::CUDAGuard guard(dev);
    auto opacity = prep(opacity_, "opacities", dev);
    auto scales = prep(scales_, "scales", dev);
    auto sh = prep(sh_, "shs", dev);
    TORCH_CHECK(sh_.dim() == 3, "shs must be [P,M,3]");
    const int P = opacity_.size(0), M = sh_.size(1);
    const size_t n = views.size();
    std::vector<das3r_raster_params> ps(n);
    std::vector<das3r_raster_inputs> ins(n);
    std::vector<das3r_raster_outputs> outs(n);
    std::vector<das3r_raster_state> sts(n);
    std::vector<das3r_raster_fwd_view> vs(n);
    std::vector<BlobAlloc> geomA(n), imgA(n);
    std::vector<torch::Tensor> keep, colors(n), radii(n);
    auto fo = opacity.options();
    for (size_t i = 0; i < n; i++) {
        const auto& V = views[i];
        TORCH_CHECK(std::get<1>(V).dim() == 2 && std::get<1>(V).size(0) == P && std::get<1>(V).size(1) == 3,
                    "means3D must have dimensions (num_points, 3) with one row per opacity");
        auto bg = prep(std::get<0>(V), "bg", dev);
        auto means3D = prep(std::get<1>(V), "means3D", dev);
        auto rotations = prep(std::get<2>(V), "rotations", dev);
        auto view = prep(std::get<3>(V), "viewmatrix", dev);
        auto proj = prep(std::get<4>(V), "projmatrix", dev);
        auto campos = prep(std::get<9>(V), "campos", dev);
        for (const auto& t : {bg, means3D, rotations, view, proj, campos}) keep.push_back(t);
        const int H = std::get<7>(V), W = std::get<8>(V);
        colors[i] = torch::empty({3, H, W}, fo);
        radii[i] = torch::empty({P}, fo.dtype(torch::kInt32));
        geomA[i].opts = torch::TensorOptions().dtype(torch::kByte).device(dev);
        imgA[i].opts = geomA[i].opts;
        memset(&ps[i], 0, sizeof(ps[i]));
        ps[i].P = P; ps[i].D = degree; ps[i].M = M; ps[i].width = W; ps[i].height = H;
        ps[i].tan_fovx = std::get<5>(V); ps[i].tan_fovy = std::get<6>(V);
        ps[i].scale_modifier = scale_modifier; ps[i].prefiltered = prefiltered; ps[i].debug = debug;
        ps[i].background = fptr(bg); ps[i].viewmatrix = fptr(view); ps[i].projmatrix = fptr(proj);
        ps[i].campos = fptr(campos);
        memset(&ins[i], 0, sizeof(ins[i]));
        ins[i].means3D = fptr(means3D); ins[i].shs = fptr(sh); ins[i].opacities = fptr(opacity);
        ins[i].scales = fptr(scales); ins[i].rotations = fptr(rotations);
        memset(&outs[i], 0, sizeof(outs[i]));
        outs[i].out_color = colors[i].data_ptr<float>();
        outs[i].radii = radii[i].data_ptr<int>();
        memset(&vs[i], 0, sizeof(vs[i]));
        vs[i].params = &ps[i]; vs[i].in = &ins[i]; vs[i].out = &outs[i]; vs[i].state = &sts[i];
        vs[i].geom_alloc = &BlobAlloc::fn; vs[i].geom_ctx = &geomA[i];
        vs[i].image_alloc = &BlobAlloc::fn; vs[i].image_ctx = &imgA[i];
    }
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    check(das3r_raster_forward_preprocess((int32_t)n, vs.data(), stream), "rasterize_gaussians_forward_preprocess");
    std::vector<std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>> res;
    auto bo = torch::TensorOptions().dtype(torch::kByte).device(dev);
    for (size_t i = 0; i < n; i++) {
        if (!geomA[i].t.defined()) geomA[i].t = torch::empty({0}, bo);
        if (!imgA[i].t.defined()) imgA[i].t = torch::empty({0}, bo);
        res.emplace_back(colors[i], radii[i], geomA[i].t, imgA[i].t);
    }
    return res;
}

// Second half for one view (current stream): -> (num_rendered, binningBuffer). color / radii are completed in place.
static std::tuple<int, torch::Tensor> ForwardRenderCUDA(const torch::Tensor& background, const torch::Tensor& color,
                                                        const torch::Tensor& radii, const torch::Tensor& geomBuffer,
                                                        const torch::Tensor& imgBuffer, const bool debug) {
    TORCH_CHECK(color.is_cuda() && color.dim() == 3, "color must be the [3,H,W] tensor forward_preprocess returned");
    const auto dev = color.device();
    c10::cuda::CUDAGuard guard(dev);
    auto bg = prep(background, "bg", dev);
    das3r_raster_params p;
    memset(&p, 0, sizeof(p));
    p.P = (int32_t)radii.size(0);
    p.height = (int32_t)color.size(1);
    p.width = (int32_t)color.size(2);
    p.debug = debug;
    p.background = fptr(bg);
    das3r_raster_outputs out;
    memset(&out, 0, sizeof(out));
    out.out_color = color.data_ptr<float>();
    out.radii = radii.data_ptr<int>();
    das3r_raster_state st;
    memset(&st, 0, sizeof(st));
    st.geom = geomBuffer.numel() ? reinterpret_cast<char*>(geomBuffer.data_ptr()) : nullptr;
    st.geom_bytes = geomBuffer.numel();
    st.image = imgBuffer.numel() ? reinterpret_cast<char*>(imgBuffer.data_ptr()) : nullptr;
    st.image_bytes = imgBuffer.numel();
    BlobAlloc binA;
    binA.opts = torch::TensorOptions().dtype(torch::kByte).device(dev);
    das3r_raster_fwd_view v;
    memset(&v, 0, sizeof(v));
    v.params = &p; v.out = &out; v.state = &st;
    v.binning_alloc = &BlobAlloc::fn; v.binning_ctx = &binA;
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    const int R = das3r_raster_forward_render(&v, stream);
    check(R, "rasterize_gaussians_forward_render");
    if (!binA.t.defined()) binA.t = torch::empty({0}, binA.opts);
    return std::make_tuple(R, binA.t);
}

// Second half for ALL views (current stream): views = (bg, color, radii, geomBuffer, imgBuffer) per view as returned by
// forward_preprocess; -> per view (num_rendered, binningBuffer). One host read-back, one launch per stage.
using RenderView = std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>;
static std::vector<std::tuple<int, torch::Tensor>> ForwardRenderBatchCUDA(const std::vector<RenderView>& views,
                                                                          const bool debug) {
    TORCH_CHECK(!views.empty(), "no views");
    const auto dev = std::get<1>(views[0]).device();
    TORCH_CHECK(dev.is_cuda(), "color must be a CUDA tensor: this rasterizer has no CPU path");
    c10::cuda::CUDAGuard guard(dev);
    const size_t n = views.size();
    std::vector<das3r_raster_params> ps(n);
    std::vector<das3r_raster_outputs> outs(n);
    std::vector<das3r_raster_state> sts(n);
    std::vector<das3r_raster_fwd_view> vs(n);
    std::vector<BlobAlloc> binA(n);
    std::vector<torch::Tensor> keep;
    for (size_t i = 0; i < n; i++) {
        const auto& color = std::get<1>(views[i]);
        const auto& radii = std::get<2>(views[i]);
        const auto& geom = std::get<3>(views[i]);
        const auto& img = std::get<4>(views[i]);
        TORCH_CHECK(color.is_cuda() && color.dim() == 3, "color must be the [3,H,W] tensor forward_preprocess returned");
        auto bg = prep(std::get<0>(views[i]), "bg", dev);
        keep.push_back(bg);
        memset(&ps[i], 0, sizeof(ps[i]));
        ps[i].P = (int32_t)radii.size(0);
        ps[i].height = (int32_t)color.size(1);
        ps[i].width = (int32_t)color.size(2);
        ps[i].debug = debug;
        ps[i].background = fptr(bg);
        memset(&outs[i], 0, sizeof(outs[i]));
        outs[i].out_color = color.data_ptr<float>();
        outs[i].radii = radii.numel() ? radii.data_ptr<int>() : nullptr;
        memset(&sts[i], 0, sizeof(sts[i]));
        sts[i].geom = geom.numel() ? reinterpret_cast<char*>(geom.data_ptr()) : nullptr;
        sts[i].geom_bytes = geom.numel();
        sts[i].image = img.numel() ? reinterpret_cast<char*>(img.data_ptr()) : nullptr;
        sts[i].image_bytes = img.numel();
        binA[i].opts = torch::TensorOptions().dtype(torch::kByte).device(dev);
        memset(&vs[i], 0, sizeof(vs[i]));
        vs[i].params = &ps[i]; vs[i].out = &outs[i]; vs[i].state = &sts[i];
        vs[i].binning_alloc = &BlobAlloc::fn; vs[i].binning_ctx = &binA[i];
    }
    std::vector<int32_t> R(n, 0);
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    check(das3r_raster_forward_render_batch((int32_t)n, vs.data(), R.data(), stream),
          "rasterize_gaussians_forward_render_batch");
    std::vector<std::tuple<int, torch::Tensor>> res;
    for (size_t i = 0; i < n; i++) {
        if (!binA[i].t.defined()) binA[i].t = torch::empty({0}, binA[i].opts);
        res.emplace_back((int)R[i], binA[i].t);
    }
    return res;
}

// ---- split / multi-view backward (das3r_raster_backward_blend / _finish; no reference counterpart) --------------
// First half for ALL views in one launch: views = (bg, dL_dout_color, geomBuffer, R, binningBuffer, imgBuffer);
// -> per view the [P,12] accumulator tensor the finish half reads.
using BlendViewT = std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, int, torch::Tensor, torch::Tensor>;
static std::vector<torch::Tensor> BackwardBlendBatchCUDA(const std::vector<BlendViewT>& views, const int P,
                                                         const bool debug) {
    TORCH_CHECK(!views.empty(), "no views");
    const auto dev = std::get<1>(views[0]).device();
    TORCH_CHECK(dev.is_cuda(), "dL_dout_color must be a CUDA tensor: this rasterizer has no CPU path");
    c10::cuda::CUDAGuard guard(dev);
    const size_t n = views.size();
    std::vector<das3r_raster_params> ps(n);
    std::vector<das3r_raster_state> sts(n);
    std::vector<das3r_raster_view> vs(n);
    std::vector<torch::Tensor> keep, accs;
    for (size_t i = 0; i < n; i++) {
        auto bg = prep(std::get<0>(views[i]), "bg", dev);
        auto dL = prep(std::get<1>(views[i]), "dL_dout_color", dev);
        TORCH_CHECK(dL.dim() == 3 && dL.size(0) == 3, "dL_dout_color must be [3,H,W]");
        keep.push_back(bg);
        keep.push_back(dL);
        const auto& geom = std::get<2>(views[i]);
        const auto& binning = std::get<4>(views[i]);
        const auto& img = std::get<5>(views[i]);
        accs.push_back(torch::empty({(long long)P, 12}, torch::TensorOptions().dtype(torch::kFloat32).device(dev)));
        memset(&ps[i], 0, sizeof(ps[i]));
        ps[i].P = P;
        ps[i].height = (int32_t)dL.size(1);
        ps[i].width = (int32_t)dL.size(2);
        ps[i].debug = debug;
        ps[i].background = fptr(bg);
        memset(&sts[i], 0, sizeof(sts[i]));
        sts[i].geom = geom.numel() ? reinterpret_cast<char*>(geom.data_ptr()) : nullptr;
        sts[i].geom_bytes = geom.numel();
        sts[i].binning = binning.numel() > 0 ? reinterpret_cast<char*>(binning.data_ptr()) : nullptr;
        sts[i].binning_bytes = binning.numel();
        sts[i].image = img.numel() ? reinterpret_cast<char*>(img.data_ptr()) : nullptr;
        sts[i].image_bytes = img.numel();
        sts[i].num_rendered = std::get<3>(views[i]);
        memset(&vs[i], 0, sizeof(vs[i]));
        vs[i].params = &ps[i];
        vs[i].state = &sts[i];
        vs[i].dL_dpix = dL.data_ptr<float>();
        vs[i].scratch = P > 0 ? accs.back().data_ptr<float>() : nullptr;
    }
    if (P == 0) return accs;
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    check(das3r_raster_backward_blend_batch((int32_t)n, vs.data(), stream), "rasterize_gaussians_backward_blend_batch");
    return accs;
}

// First half for one view: returns the [P,12] float32 accumulator tensor the finish half reads.
static torch::Tensor BackwardBlendCUDA(const torch::Tensor& background, const torch::Tensor& dL_dout_color_, const int P,
                                       const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                                       const torch::Tensor& imageBuffer, const bool debug) {
    TORCH_CHECK(dL_dout_color_.is_cuda(), "dL_dout_color must be a CUDA tensor: this rasterizer has no CPU path");
    const auto dev = dL_dout_color_.device();
    c10::cuda::CUDAGuard guard(dev);
    auto bg = prep(background, "bg", dev);
    auto dL_dout = prep(dL_dout_color_, "dL_dout_color", dev);
    torch::Tensor scratch = torch::empty({(long long)P, 12}, torch::TensorOptions().dtype(torch::kFloat32).device(dev));
    if (P == 0) return scratch;
    das3r_raster_params p;
    memset(&p, 0, sizeof(p));
    p.P = P;
    p.height = dL_dout_color_.size(1);
    p.width = dL_dout_color_.size(2);
    p.debug = debug;
    p.background = fptr(bg);
    das3r_raster_state st;
    st.geom = reinterpret_cast<char*>(geomBuffer.data_ptr());
    st.geom_bytes = geomBuffer.numel();
    st.binning = binningBuffer.numel() > 0 ? reinterpret_cast<char*>(binningBuffer.data_ptr()) : nullptr;
    st.binning_bytes = binningBuffer.numel();
    st.image = reinterpret_cast<char*>(imageBuffer.data_ptr());
    st.image_bytes = imageBuffer.numel();
    st.num_rendered = R;
    das3r_raster_view v;
    memset(&v, 0, sizeof(v));
    v.params = &p;
    v.state = &st;
    v.dL_dpix = dL_dout.data_ptr<float>();
    v.scratch = scratch.data_ptr<float>();
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    check(das3r_raster_backward_blend(&v, stream), "rasterize_gaussians_backward_blend");
    return scratch;
}

// (means3D, radii, scales, rotations, viewmatrix, projmatrix, tan_fovx, tan_fovy, campos, geomBuffer, accumulators,
//  image_height, image_width, dL_dmeans3D_view, dL_drotations_view) per view; the last two are optional outputs (empty
//  tensor: the view's gradient joins the summed dL_dmeans3D / dL_drotations)
using FinishView = std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
                              float, float, torch::Tensor, torch::Tensor, torch::Tensor, int, int, torch::Tensor,
                              torch::Tensor>;

// Second half for all views of a step: outs = {dL_dopacity [P,1], dL_dmeans3D [P,3], dL_dsh [P,M,3], dL_dscales [P,3],
// dL_drotations [P,4]} receive (accumulate: are increased by) the sum over the views.
// dsh_coeffs: coefficients per row of outs[2] (0 = M: the reference's [P,M,3]; (degree+1)^2 = rows trimmed to the active
// bands). first/count: only Gaussians [first, first+count) are processed (count < 0: to the end).
static void BackwardFinishExCUDA(const std::vector<FinishView>& views, const torch::Tensor& sh_, const int degree,
                                 const float scale_modifier, const std::vector<torch::Tensor>& outs,
                                 const bool accumulate, const bool debug, const int dsh_coeffs_, const int first,
                                 const int count) {
    TORCH_CHECK(!views.empty(), "no views");
    TORCH_CHECK(outs.size() == 5, "expected 5 gradient tensors (opacity, means3D, sh, scales, rotations)");
    const auto dev = std::get<0>(views[0]).device();
    TORCH_CHECK(dev.is_cuda(), "means3D must be a CUDA tensor: this rasterizer has no CPU path");
    c10::cuda::CUDAGuard guard(dev);
    const int P = std::get<0>(views[0]).size(0);
    int M = 0;
    if (sh_.defined() && sh_.dim() == 3) M = sh_.size(1);
    const int dsh_coeffs = dsh_coeffs_ > 0 ? dsh_coeffs_ : M;
    // outs either cover all P Gaussians or exactly the rows [first, first+count) (windows of a slice-major bucket)
    const int64_t rows_in_range = count < 0 ? (int64_t)P - first : (int64_t)count;
    const bool grads_local = rows_in_range != P && outs[0].numel() == rows_in_range;
    const int64_t Pw = grads_local ? rows_in_range : (int64_t)P;
    const int64_t want[5] = {Pw, 3LL * Pw, 3LL * dsh_coeffs * Pw, 3LL * Pw, 4LL * Pw};
    for (int i = 0; i < 5; i++)
        TORCH_CHECK(outs[i].is_cuda() && outs[i].device() == dev && outs[i].scalar_type() == torch::kFloat32 &&
                        outs[i].is_contiguous() && outs[i].numel() == want[i],
                    "gradient buffer ", i, " must be a contiguous float32 CUDA tensor of ", want[i], " elements");
    if (P == 0) return;
    auto sh = prep(sh_, "shs", dev);
    const size_t n = views.size();
    std::vector<das3r_raster_params> ps(n);
    std::vector<das3r_raster_inputs> ins(n);
    std::vector<das3r_raster_state> sts(n);
    std::vector<das3r_raster_view> vs(n);
    std::vector<torch::Tensor> keep;  // contiguous copies must outlive the launch
    for (size_t i = 0; i < n; i++) {
        const auto& V = views[i];
        TORCH_CHECK(std::get<0>(V).size(0) == P, "the views differ in the number of Gaussians");
        auto means3D = prep(std::get<0>(V), "means3D", dev);
        auto radii = std::get<1>(V).contiguous();
        TORCH_CHECK(radii.scalar_type() == torch::kInt32, "radii must be int32");
        auto scales = prep(std::get<2>(V), "scales", dev);
        auto rotations = prep(std::get<3>(V), "rotations", dev);
        auto view = prep(std::get<4>(V), "viewmatrix", dev);
        auto proj = prep(std::get<5>(V), "projmatrix", dev);
        auto campos = prep(std::get<8>(V), "campos", dev);
        const auto& geom = std::get<9>(V);
        const auto& acc = std::get<10>(V);
        TORCH_CHECK(acc.is_cuda() && acc.scalar_type() == torch::kFloat32 && acc.is_contiguous() &&
                        acc.numel() == 12LL * P,
                    "accumulators must be the tensor rasterize_gaussians_backward_blend returned");
        for (const auto& t : {means3D, radii, scales, rotations, view, proj, campos}) keep.push_back(t);
        memset(&ps[i], 0, sizeof(ps[i]));
        ps[i].P = P; ps[i].D = degree; ps[i].M = M;
        ps[i].height = std::get<11>(V); ps[i].width = std::get<12>(V);
        ps[i].tan_fovx = std::get<6>(V); ps[i].tan_fovy = std::get<7>(V);
        ps[i].scale_modifier = scale_modifier;
        ps[i].debug = debug;
        ps[i].viewmatrix = fptr(view); ps[i].projmatrix = fptr(proj); ps[i].campos = fptr(campos);
        memset(&ins[i], 0, sizeof(ins[i]));
        ins[i].means3D = fptr(means3D); ins[i].shs = fptr(sh); ins[i].scales = fptr(scales);
        ins[i].rotations = fptr(rotations);
        memset(&sts[i], 0, sizeof(sts[i]));
        sts[i].geom = reinterpret_cast<char*>(geom.data_ptr());
        sts[i].geom_bytes = geom.numel();
        memset(&vs[i], 0, sizeof(vs[i]));
        vs[i].params = &ps[i]; vs[i].in = &ins[i]; vs[i].radii = radii.data_ptr<int>(); vs[i].state = &sts[i];
        vs[i].scratch = acc.data_ptr<float>();
        const auto& gm = std::get<13>(V);
        const auto& gr = std::get<14>(V);
        if (gm.defined() && gm.numel() > 0) {
            TORCH_CHECK(gm.is_cuda() && gm.scalar_type() == torch::kFloat32 && gm.is_contiguous() && gm.numel() == 3LL * P,
                        "per-view dL_dmeans3D must be a contiguous float32 CUDA tensor of 3*P elements");
            vs[i].dL_dmeans3D = gm.data_ptr<float>();
        }
        if (gr.defined() && gr.numel() > 0) {
            TORCH_CHECK(gr.is_cuda() && gr.scalar_type() == torch::kFloat32 && gr.is_contiguous() && gr.numel() == 4LL * P,
                        "per-view dL_drotations must be a contiguous float32 CUDA tensor of 4*P elements");
            vs[i].dL_drotations = gr.data_ptr<float>();
        }
    }
    das3r_raster_grads g;
    memset(&g, 0, sizeof(g));
    g.dL_dopacity = outs[0].data_ptr<float>();
    g.dL_dmeans3D = outs[1].data_ptr<float>();
    g.dL_dsh = M > 0 ? outs[2].data_ptr<float>() : nullptr;
    g.dL_dscales = outs[3].data_ptr<float>();
    g.dL_drotations = outs[4].data_ptr<float>();
    g.accumulate = accumulate ? 1 : 0;
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    check(das3r_raster_backward_finish_ex((int32_t)n, vs.data(), &g, dsh_coeffs, first, count, grads_local ? 1 : 0, stream),
          "rasterize_gaussians_backward_finish");
}

static void BackwardFinishCUDA(const std::vector<FinishView>& views, const torch::Tensor& sh_, const int degree,
                               const float scale_modifier, const std::vector<torch::Tensor>& outs,
                               const bool accumulate, const bool debug) {
    BackwardFinishExCUDA(views, sh_, degree, scale_modifier, outs, accumulate, debug, 0, 0, -1);
}

// ---- fused DAS3R render front-end (include/das3r_frontend.h) ------------------------------------------------------
static void fill_model(das3r_frontend_model& m, const torch::Tensor& xyz, const torch::Tensor& rotation,
                       const torch::Tensor& opacity, const torch::Tensor& scaling, const torch::Tensor& f_dc,
                       const torch::Tensor& f_rest, const torch::Tensor& conf_static, const torch::Tensor& conf_index) {
    memset(&m, 0, sizeof(m));
    m.P = (int32_t)xyz.size(0);
    m.xyz = xyz.data_ptr<float>();
    m.rotation = rotation.data_ptr<float>();
    m.opacity = opacity.data_ptr<float>();
    m.scaling = scaling.data_ptr<float>();
    m.features_dc = f_dc.defined() && f_dc.numel() ? f_dc.data_ptr<float>() : nullptr;
    m.features_rest = f_rest.defined() && f_rest.numel() ? f_rest.data_ptr<float>() : nullptr;
    m.conf_static = conf_static.data_ptr<float>();
    m.conf_index = conf_index.defined() && conf_index.numel() ? conf_index.data_ptr<int64_t>() : nullptr;
}

static void check_model(const torch::Tensor& xyz, const torch::Tensor& rotation, const torch::Tensor& opacity,
                        const torch::Tensor& scaling, const torch::Tensor& conf_static, const torch::Tensor& conf_index,
                        const torch::Tensor& poses, int64_t rows = -1) {
    TORCH_CHECK(xyz.is_cuda(), "xyz must be a CUDA tensor: the front-end has no CPU path");
    const int64_t P = xyz.size(0);
    if (rows < 0) rows = P;  // output rows: fewer than P when render()'s `filtering` selects a subset
    auto ok = [&](const torch::Tensor& t, int64_t n) {
        return t.is_cuda() && t.device() == xyz.device() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() &&
               t.numel() == n;
    };
    TORCH_CHECK(ok(xyz, 3 * P) && ok(rotation, 4 * P) && ok(opacity, P) && ok(scaling, 3 * P),
                "xyz [P,3], rotation [P,4], opacity [P,1], scaling [P,3] must be contiguous float32 CUDA tensors");
    TORCH_CHECK(conf_static.is_cuda() && conf_static.scalar_type() == torch::kFloat32 && conf_static.is_contiguous(),
                "conf_static must be a contiguous float32 CUDA tensor");
    if (conf_index.defined() && conf_index.numel()) {
        TORCH_CHECK(conf_index.is_cuda() && conf_index.scalar_type() == torch::kInt64 && conf_index.is_contiguous() &&
                        conf_index.numel() == rows,
                    "conf_index must be a contiguous int64 CUDA tensor with one entry per (selected) Gaussian");
    } else {
        TORCH_CHECK(conf_static.numel() >= rows, "conf_static has fewer entries than there are (selected) Gaussians");
    }
    TORCH_CHECK(poses.is_cuda() && poses.scalar_type() == torch::kFloat32 && poses.is_contiguous() && poses.dim() == 2 &&
                    poses.size(1) == 7 && poses.size(0) >= 1 && poses.size(0) <= DAS3R_FRONTEND_MAX_VIEWS,
                "poses must be a contiguous float32 CUDA tensor [V,7] with 1 <= V <= 16");
}

// sh_degree: active SH degree (only its (D+1)^2 coefficients are copied into the joint rows); filter_index: optional
// int64 [rows] model rows selected by render()'s `filtering` (empty: all rows)
static std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, std::vector<torch::Tensor>, std::vector<torch::Tensor>>
FrontendForwardExCUDA(const torch::Tensor& xyz, const torch::Tensor& rotation, const torch::Tensor& opacity,
                      const torch::Tensor& scaling, const torch::Tensor& f_dc, const torch::Tensor& f_rest,
                      const torch::Tensor& conf_static, const torch::Tensor& conf_index, const torch::Tensor& poses,
                      const int sh_degree, const torch::Tensor& filter_index) {
    const bool filtered = filter_index.defined() && filter_index.numel() > 0;
    const int64_t Pm = xyz.size(0);
    const int64_t P = filtered ? filter_index.numel() : Pm;
    check_model(xyz, rotation, opacity, scaling, conf_static, conf_index, poses, P);
    if (filtered) {
        TORCH_CHECK(filter_index.is_cuda() && filter_index.scalar_type() == torch::kInt64 && filter_index.is_contiguous(),
                    "filter_index must be a contiguous int64 CUDA tensor");
    }
    TORCH_CHECK(f_dc.numel() == 3 * Pm && f_rest.numel() == 45 * Pm, "features_dc [P,1,3] / features_rest [P,15,3] sizes");
    TORCH_CHECK(f_dc.is_cuda() && f_dc.is_contiguous() && f_rest.is_cuda() && f_rest.is_contiguous() &&
                    f_dc.scalar_type() == torch::kFloat32 && f_rest.scalar_type() == torch::kFloat32,
                "features_dc [P,1,3] and features_rest [P,15,3] must be contiguous float32 CUDA tensors");
    c10::cuda::CUDAGuard guard(xyz.device());
    const int V = (int)poses.size(0);
    auto o = xyz.options();
    auto opac = torch::empty({P, 1}, o), scales = torch::empty({P, 3}, o), shs = torch::empty({P, 16, 3}, o);
    std::vector<torch::Tensor> means, rots;
    das3r_frontend_outputs out;
    memset(&out, 0, sizeof(out));
    out.opacity = opac.data_ptr<float>();
    out.scales = scales.data_ptr<float>();
    out.shs = shs.data_ptr<float>();
    for (int v = 0; v < V; v++) {
        means.push_back(torch::empty({P, 3}, o));
        rots.push_back(torch::empty({P, 4}, o));
        out.means3D[v] = means.back().data_ptr<float>();
        out.rotations[v] = rots.back().data_ptr<float>();
    }
    das3r_frontend_model m;
    fill_model(m, xyz, rotation, opacity, scaling, f_dc, f_rest, conf_static, conf_index);
    das3r_frontend_options opt;
    memset(&opt, 0, sizeof(opt));
    opt.sh_degree = sh_degree;
    opt.n_rows = (int32_t)P;
    opt.filter_index = filtered ? filter_index.data_ptr<int64_t>() : nullptr;
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(xyz.device().index()).stream();
    check(das3r_frontend_forward_ex(&m, V, poses.data_ptr<float>(), &out, &opt, stream), "frontend_forward");
    return std::make_tuple(opac, scales, shs, means, rots);
}

static std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, std::vector<torch::Tensor>, std::vector<torch::Tensor>>
FrontendForwardCUDA(const torch::Tensor& xyz, const torch::Tensor& rotation, const torch::Tensor& opacity,
                    const torch::Tensor& scaling, const torch::Tensor& f_dc, const torch::Tensor& f_rest,
                    const torch::Tensor& conf_static, const torch::Tensor& conf_index, const torch::Tensor& poses) {
    return FrontendForwardExCUDA(xyz, rotation, opacity, scaling, f_dc, f_rest, conf_static, conf_index, poses, 3,
                                 torch::Tensor());
}

// Backward of the front-end. outs: empty (the 8 gradient tensors are allocated: dxyz [P,3], drot [P,4], dop [P,1], dsc
// [P,3], ddc [P,1,3], drest [P,15,3], dconf [rows], dposes [V,7]) or those 8 tensors provided by the caller (e.g. windows
// of a gradient bucket; drest may be [P,(D+1)^2-1,3] = trimmed to the active bands; with filter_index the caller
// zero-fills them). scratch: empty (allocated) or a uint8 tensor of das3r_frontend_backward_scratch_bytes(rows, V) bytes
// that lives across the ranges of a step. first / count / finish_poses: see das3r_frontend_options.
static std::vector<torch::Tensor> FrontendBackwardExCUDA(
    const torch::Tensor& xyz, const torch::Tensor& rotation, const torch::Tensor& opacity, const torch::Tensor& scaling,
    const torch::Tensor& conf_static, const torch::Tensor& conf_index, const torch::Tensor& poses,
    const torch::Tensor& dL_dopacity, const torch::Tensor& dL_dscales, const torch::Tensor& dL_dshs,
    const std::vector<torch::Tensor>& dL_dmeans, const std::vector<torch::Tensor>& dL_drots, const int sh_degree,
    const torch::Tensor& filter_index, const std::vector<torch::Tensor>& outs_, const torch::Tensor& scratch_,
    const int first, const int count, const bool finish_poses) {
    const bool filtered = filter_index.defined() && filter_index.numel() > 0;
    const int64_t Pm = xyz.size(0);
    const int64_t P = filtered ? filter_index.numel() : Pm;
    check_model(xyz, rotation, opacity, scaling, conf_static, conf_index, poses, P);
    const int V = (int)poses.size(0);
    TORCH_CHECK(sh_degree >= 0 && sh_degree <= 3, "sh_degree must be 0..3");
    const int K = (sh_degree + 1) * (sh_degree + 1);
    TORCH_CHECK((int)dL_dmeans.size() == V && (int)dL_drots.size() == V, "one dL_dmeans3D / dL_drotations per view");
    if (filtered) {
        TORCH_CHECK(filter_index.is_cuda() && filter_index.scalar_type() == torch::kInt64 && filter_index.is_contiguous(),
                    "filter_index must be a contiguous int64 CUDA tensor");
    }
    c10::cuda::CUDAGuard guard(xyz.device());
    const auto dev = xyz.device();
    auto o = xyz.options();
    std::vector<torch::Tensor> keep;
    auto cg = [&](const torch::Tensor& t, int64_t n, const char* name) {
        TORCH_CHECK(t.is_cuda() && t.device() == dev && t.scalar_type() == torch::kFloat32 && t.numel() == n, name,
                    " must be a float32 CUDA tensor of ", n, " elements");
        keep.push_back(t.contiguous());
        return keep.back().data_ptr<float>();
    };
    TORCH_CHECK(dL_dshs.numel() % (3 * std::max<int64_t>(P, 1)) == 0, "dL_dshs must be [rows, coeffs, 3]");
    const int dsh = P > 0 ? (int)(dL_dshs.numel() / (3 * P)) : 16;
    TORCH_CHECK(dsh >= K && dsh <= 16, "dL_dshs rows must have (D+1)^2 .. 16 coefficients");
    das3r_frontend_grads_in gin;
    memset(&gin, 0, sizeof(gin));
    gin.dL_dopacity = cg(dL_dopacity, P, "dL_dopacity");
    gin.dL_dscales = cg(dL_dscales, 3 * P, "dL_dscales");
    gin.dL_dshs = cg(dL_dshs, 3LL * dsh * P, "dL_dshs");
    for (int v = 0; v < V; v++) {
        gin.dL_dmeans3D[v] = cg(dL_dmeans[v], 3 * P, "dL_dmeans3D");
        gin.dL_drotations[v] = cg(dL_drots[v], 4 * P, "dL_drotations");
    }
    std::vector<torch::Tensor> outs = outs_;
    if (outs.empty()) {
        auto mk = [&](std::initializer_list<int64_t> shape) {
            return filtered ? torch::zeros(shape, o) : torch::empty(shape, o);
        };
        outs = {mk({Pm, 3}), mk({Pm, 4}), mk({Pm, 1}), mk({Pm, 3}), mk({Pm, 1, 3}), mk({Pm, 15, 3}),
                torch::empty({P}, o), torch::empty({V, 7}, o)};
    }
    TORCH_CHECK(outs.size() == 8, "expected 8 gradient tensors");
    // outs either cover all model rows or exactly the rows [first, first+count) (windows of a slice-major bucket)
    const int64_t rows_in_range = std::min<int64_t>(count < 0 ? P - first : count, P - first);
    const bool local = !filtered && rows_in_range != Pm && outs[0].numel() == 3 * rows_in_range;
    const int64_t Pw = local ? rows_in_range : Pm;
    const int64_t rest_numel = outs[5].defined() ? outs[5].numel() : 0;
    TORCH_CHECK(rest_numel == 45 * Pw || rest_numel == 3LL * (K - 1) * Pw,
                "dL_dfeatures_rest must be [P,15,3] or trimmed to [P,(D+1)^2-1,3]");
    const int rest = rest_numel == 45 * Pw ? 15 : K - 1;
    const int64_t want[8] = {3 * Pw, 4 * Pw, Pw, 3 * Pw, 3 * Pw, rest_numel, local ? rows_in_range : P, 7LL * V};
    for (int i = 0; i < 8; i++) {
        if (i == 5 && rest_numel == 0) continue;
        TORCH_CHECK(outs[i].is_cuda() && outs[i].device() == dev && outs[i].scalar_type() == torch::kFloat32 &&
                        outs[i].is_contiguous() && outs[i].numel() == want[i],
                    "gradient tensor ", i, " must be a contiguous float32 CUDA tensor of ", want[i], " elements");
    }
    torch::Tensor scratch = scratch_;
    const size_t need = das3r_frontend_backward_scratch_bytes((int32_t)P, V);
    if (!scratch.defined() || scratch.numel() == 0)
        scratch = torch::empty({(long long)need}, torch::TensorOptions().dtype(torch::kByte).device(dev));
    TORCH_CHECK(scratch.is_cuda() && (size_t)scratch.numel() * scratch.element_size() >= need, "scratch too small");
    das3r_frontend_grads_out gout;
    memset(&gout, 0, sizeof(gout));
    gout.dL_dxyz = outs[0].data_ptr<float>();
    gout.dL_drotation = outs[1].data_ptr<float>();
    gout.dL_dopacity = outs[2].data_ptr<float>();
    gout.dL_dscaling = outs[3].data_ptr<float>();
    gout.dL_dfeatures_dc = outs[4].data_ptr<float>();
    gout.dL_dfeatures_rest = rest_numel ? outs[5].data_ptr<float>() : nullptr;
    gout.dL_dconf = outs[6].data_ptr<float>();
    gout.dL_dposes = outs[7].data_ptr<float>();
    gout.scratch = reinterpret_cast<float*>(scratch.data_ptr());
    das3r_frontend_model m;
    torch::Tensor none;
    fill_model(m, xyz, rotation, opacity, scaling, none, none, conf_static, conf_index);
    // the backward does not read the features, only their shapes matter; the C ABI wants non-null pointers
    m.features_dc = m.xyz;
    m.features_rest = m.xyz;
    das3r_frontend_options opt;
    memset(&opt, 0, sizeof(opt));
    opt.sh_degree = sh_degree;
    opt.n_rows = (int32_t)P;
    opt.filter_index = filtered ? filter_index.data_ptr<int64_t>() : nullptr;
    opt.dsh_coeffs = dsh;
    opt.rest_coeffs = rest == 15 ? 15 : -1;
    opt.first = first;
    opt.count = count;
    opt.finish_poses = finish_poses ? 1 : 0;
    opt.local_outputs = local ? 1 : 0;
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    check(das3r_frontend_backward_ex(&m, V, poses.data_ptr<float>(), &gin, &gout, &opt, stream), "frontend_backward");
    return outs;
}

static std::vector<torch::Tensor> FrontendBackwardCUDA(
    const torch::Tensor& xyz, const torch::Tensor& rotation, const torch::Tensor& opacity, const torch::Tensor& scaling,
    const torch::Tensor& conf_static, const torch::Tensor& conf_index, const torch::Tensor& poses,
    const torch::Tensor& dL_dopacity, const torch::Tensor& dL_dscales, const torch::Tensor& dL_dshs,
    const std::vector<torch::Tensor>& dL_dmeans, const std::vector<torch::Tensor>& dL_drots) {
    return FrontendBackwardExCUDA(xyz, rotation, opacity, scaling, conf_static, conf_index, poses, dL_dopacity, dL_dscales,
                                  dL_dshs, dL_dmeans, dL_drots, 3, torch::Tensor(), {}, torch::Tensor(), 0, -1, true);
}

// ---- fused photometric loss (include/das3r_loss.h) ----------------------------------------------------------------
static void loss_inputs(das3r_loss_inputs& in, const torch::Tensor& image, const torch::Tensor& gt,
                        const torch::Tensor& static_mask, float lambda_dssim) {
    TORCH_CHECK(image.is_cuda(), "image must be a CUDA tensor: the loss has no CPU path");
    TORCH_CHECK(image.dim() == 3 && image.size(0) == 3 && image.scalar_type() == torch::kFloat32 && image.is_contiguous(),
                "image must be a contiguous float32 [3,H,W] tensor");
    TORCH_CHECK(gt.is_cuda() && gt.device() == image.device() && gt.sizes() == image.sizes() &&
                    gt.scalar_type() == torch::kFloat32 && gt.is_contiguous(),
                "gt must be a contiguous float32 CUDA tensor of the image's shape");
    const bool has_static = static_mask.defined() && static_mask.numel() > 0;
    if (has_static)
        TORCH_CHECK(static_mask.is_cuda() && static_mask.device() == image.device() &&
                        static_mask.scalar_type() == torch::kFloat32 && static_mask.is_contiguous() &&
                        static_mask.numel() == image.size(1) * image.size(2),
                    "static mask must be a contiguous float32 CUDA tensor with H*W elements");
    memset(&in, 0, sizeof(in));
    in.height = (int32_t)image.size(1);
    in.width = (int32_t)image.size(2);
    in.lambda_dssim = lambda_dssim;
    in.image = image.data_ptr<float>();
    in.gt = gt.data_ptr<float>();
    in.static_mask = has_static ? static_mask.data_ptr<float>() : nullptr;
}

// -> (out [2] = {loss, psnr}, state)
static std::tuple<torch::Tensor, torch::Tensor> LossForwardCUDA(const torch::Tensor& image, const torch::Tensor& gt,
                                                                const torch::Tensor& static_mask, float lambda_dssim) {
    das3r_loss_inputs in;
    loss_inputs(in, image, gt, static_mask, lambda_dssim);
    c10::cuda::CUDAGuard guard(image.device());
    auto out = torch::empty({2}, image.options());
    auto state = torch::empty({(long long)das3r_loss_state_bytes(in.width, in.height)},
                              torch::TensorOptions().dtype(torch::kByte).device(image.device()));
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(image.device().index()).stream();
    check(das3r_loss_forward(&in, out.data_ptr<float>(), state.data_ptr(), stream), "loss_forward");
    return std::make_tuple(out, state);
}

// -> (dL_dimage, dL_dstatic or empty, dL_dgt or empty); grad_loss: 1-element float32 CUDA tensor or empty (= 1)
static std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> LossBackwardCUDA(
    const torch::Tensor& image, const torch::Tensor& gt, const torch::Tensor& static_mask, float lambda_dssim,
    const torch::Tensor& state, const torch::Tensor& grad_loss, bool want_static, bool want_gt) {
    das3r_loss_inputs in;
    loss_inputs(in, image, gt, static_mask, lambda_dssim);
    TORCH_CHECK(state.is_cuda() && (size_t)state.numel() >= das3r_loss_state_bytes(in.width, in.height),
                "state must be the tensor loss_forward returned");
    const float* gl = nullptr;
    torch::Tensor glc;
    if (grad_loss.defined() && grad_loss.numel() > 0) {
        TORCH_CHECK(grad_loss.is_cuda() && grad_loss.scalar_type() == torch::kFloat32 && grad_loss.numel() == 1,
                    "grad_loss must be a 1-element float32 CUDA tensor");
        glc = grad_loss.contiguous();
        gl = glc.data_ptr<float>();
    }
    c10::cuda::CUDAGuard guard(image.device());
    auto dimg = torch::empty_like(image);
    torch::Tensor dstatic, dgt;
    if (want_static) dstatic = torch::empty({image.size(1), image.size(2)}, image.options());
    if (want_gt) dgt = torch::empty_like(image);
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(image.device().index()).stream();
    check(das3r_loss_backward(&in, state.data_ptr(), gl, dimg.data_ptr<float>(),
                              want_static ? dstatic.data_ptr<float>() : nullptr, want_gt ? dgt.data_ptr<float>() : nullptr,
                              stream),
          "loss_backward");
    if (!want_static) dstatic = torch::empty({0}, image.options());
    if (!want_gt) dgt = torch::empty({0}, image.options());
    return std::make_tuple(dimg, dstatic, dgt);
}

// ---- fused Adam step (include/das3r_adam.h) -----------------------------------------------------------------------
// params / grads / exp_avgs / exp_avg_sqs: parallel lists of contiguous float32 CUDA tensors; lrs: one per tensor.
static void AdamStepCUDA(const std::vector<torch::Tensor>& params, const std::vector<torch::Tensor>& grads,
                         const std::vector<torch::Tensor>& exp_avgs, const std::vector<torch::Tensor>& exp_avg_sqs,
                         const std::vector<double>& lrs, double beta1, double beta2, double eps, int64_t step) {
    const size_t n = params.size();
    TORCH_CHECK(grads.size() == n && exp_avgs.size() == n && exp_avg_sqs.size() == n && lrs.size() == n,
                "params, grads, exp_avgs, exp_avg_sqs and lrs must have the same length");
    if (n == 0) return;
    TORCH_CHECK(params[0].is_cuda(), "parameters must be CUDA tensors: the fused Adam has no CPU path");
    const auto dev = params[0].device();
    c10::cuda::CUDAGuard guard(dev);
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    for (size_t i0 = 0; i0 < n; i0 += DAS3R_ADAM_MAX_TENSORS) {
        das3r_adam_tensor t[DAS3R_ADAM_MAX_TENSORS];
        const size_t m = std::min(n - i0, (size_t)DAS3R_ADAM_MAX_TENSORS);
        for (size_t j = 0; j < m; j++) {
            const size_t i = i0 + j;
            for (const torch::Tensor* x : {&params[i], &grads[i], &exp_avgs[i], &exp_avg_sqs[i]})
                TORCH_CHECK(x->is_cuda() && x->device() == dev && x->scalar_type() == torch::kFloat32 &&
                                x->is_contiguous() && x->numel() == params[i].numel(),
                            "tensor ", i, ": parameter, gradient and moments must be contiguous float32 CUDA tensors of "
                            "one size on one device");
            t[j].param = params[i].data_ptr<float>();
            t[j].grad = grads[i].data_ptr<float>();
            t[j].exp_avg = exp_avgs[i].data_ptr<float>();
            t[j].exp_avg_sq = exp_avg_sqs[i].data_ptr<float>();
            t[j].n = params[i].numel();
            t[j].lr = lrs[i];
        }
        check(das3r_adam_step((int32_t)m, t, beta1, beta2, eps, step, stream), "adam_step");
    }
}

// ---- pointmap initialisation (include/das3r_init.h) ---------------------------------------------------------------
// K [V,3,3], poses [V,4,4] (camera-to-world), depth / confidence / dyna_avg [V,H,W], rgb [V,3,H,W], log_conf_thre
// -> (xyz, features_dc, features_rest, scaling, rotation, opacity, conf_static [V,H,W], aggregated_mask [V*H*W] bool,
//     conf_index [P] int64, dist2 [P])
static std::vector<torch::Tensor> PointmapInitCUDA(const torch::Tensor& K_, const torch::Tensor& poses_,
                                                   const torch::Tensor& depth_, const torch::Tensor& conf_,
                                                   const torch::Tensor& dyna_, const torch::Tensor& rgb_,
                                                   const double log_conf_thre) {
    TORCH_CHECK(depth_.is_cuda(), "depth must be a CUDA tensor: the initialisation has no CPU path");
    const auto dev = depth_.device();
    c10::cuda::CUDAGuard guard(dev);
    TORCH_CHECK(depth_.dim() == 3, "depth must be [V,H,W]");
    const int64_t V = depth_.size(0), H = depth_.size(1), W = depth_.size(2);
    auto K = prep(K_, "K", dev), poses = prep(poses_, "poses", dev), depth = prep(depth_, "depth", dev);
    auto conf = prep(conf_, "confidence", dev), dyna = prep(dyna_, "dyna_avg", dev), rgb = prep(rgb_, "rgb", dev);
    TORCH_CHECK(K.numel() == 9 * V && poses.numel() == 16 * V && conf.numel() == V * H * W && dyna.numel() == V * H * W &&
                    rgb.numel() == 3 * V * H * W,
                "K [V,3,3], poses [V,4,4], confidence / dyna_avg [V,H,W], rgb [V,3,H,W]");
    das3r_pointmaps in;
    memset(&in, 0, sizeof(in));
    in.V = (int32_t)V; in.H = (int32_t)H; in.W = (int32_t)W;
    in.K = fptr(K); in.poses = fptr(poses); in.depth = fptr(depth); in.confidence = fptr(conf); in.dyna_avg = fptr(dyna);
    in.rgb = fptr(rgb);
    in.log_conf_thre = (float)log_conf_thre;
    const int64_t n = V * H * W;
    auto bo = torch::TensorOptions().dtype(torch::kByte).device(dev);
    auto fo = depth.options();
    auto mask = torch::empty({n}, torch::TensorOptions().dtype(torch::kBool).device(dev));
    auto conf_static = torch::empty({V, H, W}, fo);
    const size_t sbytes = das3r_pointmap_scratch_bytes(n);
    auto scratch = torch::empty({(long long)sbytes}, bo);
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    const int64_t P = das3r_pointmap_select(&in, reinterpret_cast<uint8_t*>(mask.data_ptr<bool>()),
                                            conf_static.data_ptr<float>(), scratch.data_ptr(), sbytes, stream);
    if (P < 0) throw std::runtime_error(std::string("pointmap_select: ") + das3r_last_error());
    auto xyz = torch::empty({P, 3}, fo), fdc = torch::empty({P, 1, 3}, fo), frest = torch::empty({P, 15, 3}, fo);
    auto scaling = torch::empty({P, 3}, fo), rotation = torch::empty({P, 4}, fo), opacity = torch::empty({P, 1}, fo);
    auto conf_index = torch::empty({P}, torch::TensorOptions().dtype(torch::kInt64).device(dev));
    auto dist2 = torch::empty({P}, fo);
    if (P > 0) {
        const size_t kbytes = das3r_knn_scratch_bytes((int32_t)P);
        auto kscratch = torch::empty({(long long)kbytes}, bo);
        das3r_init_outputs out;
        memset(&out, 0, sizeof(out));
        out.xyz = xyz.data_ptr<float>(); out.features_dc = fdc.data_ptr<float>(); out.features_rest = frest.data_ptr<float>();
        out.scaling = scaling.data_ptr<float>(); out.rotation = rotation.data_ptr<float>();
        out.opacity = opacity.data_ptr<float>(); out.conf_index = conf_index.data_ptr<int64_t>();
        out.dist2 = dist2.data_ptr<float>();
        check(das3r_pointmap_init(&in, reinterpret_cast<const uint8_t*>(mask.data_ptr<bool>()), P, &out, scratch.data_ptr(),
                                  sbytes, kscratch.data_ptr(), kbytes, stream),
              "pointmap_init");
    }
    return {xyz, fdc, frest, scaling, rotation, opacity, conf_static, mask, conf_index, dist2};
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.def("pointmap_init", &PointmapInitCUDA, py::call_guard<py::gil_scoped_release>());
    // The GIL is released around every call into the library: a host that drives several CUDA streams from several
    // threads (bench.py) must not serialise on the forward's read-back of num_rendered.
    m.def("rasterize_gaussians", &RasterizeGaussiansCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_ex", &RasterizeGaussiansExCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_backward_into", &RasterizeGaussiansBackwardIntoCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_forward_preprocess", &ForwardPreprocessCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_forward_render", &ForwardRenderCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_backward_blend", &BackwardBlendCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_backward_finish", &BackwardFinishCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_backward_finish_ex", &BackwardFinishExCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_forward_render_batch", &ForwardRenderBatchCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("rasterize_gaussians_backward_blend_batch", &BackwardBlendBatchCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("frontend_forward_ex", &FrontendForwardExCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("frontend_backward_ex", &FrontendBackwardExCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("frontend_forward", &FrontendForwardCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("frontend_backward", &FrontendBackwardCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("loss_forward", &LossForwardCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("loss_backward", &LossBackwardCUDA, py::call_guard<py::gil_scoped_release>());
    m.def("adam_step", &AdamStepCUDA);
    m.def("nvls_allreduce_sum", [](int64_t multicast_ptr, int64_t offset_floats, int64_t n_floats, int rank, int world,
                                   int n_ctas, int device_index) {
        c10::cuda::CUDAGuard guard((c10::DeviceIndex)device_index);
        cudaStream_t stream = c10::cuda::getCurrentCUDAStream(device_index).stream();
        check(das3r_nvls_allreduce_sum(reinterpret_cast<void*>(multicast_ptr), (size_t)offset_floats, (size_t)n_floats,
                                       rank, world, n_ctas, stream),
              "nvls_allreduce_sum");
    });
    m.def("prof_enable", [](bool on) { das3r_prof_enable(on ? 1 : 0); });
    m.def("prof_collect", [](bool reset) {
        int n = das3r_prof_num_stages();
        std::vector<double> ms(n);
        std::vector<long long> calls(n);
        if (das3r_prof_collect(ms.data(), calls.data(), reset ? 1 : 0) < 0) throw std::runtime_error(das3r_last_error());
        std::vector<std::tuple<std::string, double, long long>> out;
        for (int i = 0; i < n; i++) out.emplace_back(das3r_prof_stage_name(i), ms[i], calls[i]);
        return out;
    });
    m.def("launch_count", []() { return das3r_launch_count(); });
    m.def("mark_visible", &markVisible, py::call_guard<py::gil_scoped_release>());
    m.def("distCUDA2", &distCUDA2, py::call_guard<py::gil_scoped_release>());
    m.def("abi_version", []() { return das3r_abi_version(); });
}
