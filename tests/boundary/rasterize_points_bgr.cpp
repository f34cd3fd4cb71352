// The binding a maintainer of the reference would write: the three functions its pybind module
// exports (DGR/ext.cpp:15-18, signatures of DGR/rasterize_points.h:18-67) implemented over the C ABI
// of libbgr.so (include/bgr.h) instead of CudaRasterizer::Rasterizer. C++ only - no CUDA code here;
// ATen supplies tensors and the current stream, exactly as in DGR/rasterize_points.cu.
//
// Built by __graft_entry__.build() as the torch extension `bgr_boundary` and driven by the GPU
// parity tests next to the ctypes binding of the product (diff_gaussian_rasterization/_C.py).
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <stdexcept>
#include <string>
#include <tuple>

#include "bgr.h"

namespace {

// resizeFunctional (rasterize_points.cu:27-33): grow the byte tensor, hand back its address
void* grow(size_t n, void* user)
{
    auto* t = static_cast<torch::Tensor*>(user);
    t->resize_({(long long)n});
    return t->data_ptr();
}

// `.contiguous().data<float>()` of an empty tensor is nullptr: the reference's convention for an
// absent optional input (rasterize_points.cu:94-111)
const float* fptr(const torch::Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; }

torch::Tensor dense(const torch::Tensor& t) { return t.numel() ? t.contiguous() : t; }

void check(int rc)
{
    if (rc != BGR_OK) throw std::runtime_error(std::string("bgr error ") + std::to_string(rc) + ": " + bgr_last_error());
}

}  // namespace

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D,
                       const torch::Tensor& colors, const torch::Tensor& opacity,
                       const torch::Tensor& scales, const torch::Tensor& rotations,
                       const float scale_modifier, const torch::Tensor& cov3D_precomp,
                       const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                       const float tan_fovx, const float tan_fovy, const int image_height,
                       const int image_width, const torch::Tensor& sh, const int degree,
                       const torch::Tensor& campos, const bool prefiltered, const bool debug)
{
    if (means3D.ndimension() != 2 || means3D.size(1) != 3)
        AT_ERROR("means3D must have dimensions (num_points, 3)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0), H = image_height, W = image_width;
    auto float_opts = means3D.options().dtype(torch::kFloat32);
    auto out_color = torch::empty({3, H, W}, float_opts);          // every element is written
    auto radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
    auto bytes = means3D.options().dtype(torch::kByte);
    torch::Tensor geom = torch::empty({0}, bytes), binning = torch::empty({0}, bytes),
                  img = torch::empty({0}, bytes);
    int rendered = 0;
    if (P == 0) {
        out_color.zero_();                                         // rasterize_points.cu:81
        return std::make_tuple(rendered, out_color, radii, geom, binning, img);
    }
    const auto bg = dense(background), me = dense(means3D), sh_ = dense(sh), co = dense(colors),
               op = dense(opacity), sc = dense(scales), ro = dense(rotations), cv = dense(cov3D_precomp),
               vi = dense(viewmatrix), pr = dense(projmatrix), ca = dense(campos);
    BgrForwardArgs a{};
    a.P = P; a.D = degree; a.M = sh.size(0) != 0 ? (int)sh.size(1) : 0; a.width = W; a.height = H;
    a.background = fptr(bg); a.means3D = fptr(me); a.shs = fptr(sh_); a.colors_precomp = fptr(co);
    a.opacities = fptr(op); a.scales = fptr(sc); a.scale_modifier = scale_modifier;
    a.rotations = fptr(ro); a.cov3D_precomp = fptr(cv); a.viewmatrix = fptr(vi);
    a.projmatrix = fptr(pr); a.cam_pos = fptr(ca); a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.prefiltered = prefiltered; a.debug = debug;
    a.out_color = out_color.data_ptr<float>(); a.radii = radii.data_ptr<int>();
    check(bgr_forward(&a, grow, &geom, grow, &binning, grow, &img,
                      at::cuda::getCurrentCUDAStream().stream(), &rendered));
    return std::make_tuple(rendered, out_color, radii, geom, binning, img);
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor, torch::Tensor>
RasterizeGaussiansBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D,
                               const torch::Tensor& radii, const torch::Tensor& colors,
                               const torch::Tensor& scales, const torch::Tensor& rotations,
                               const float scale_modifier, const torch::Tensor& cov3D_precomp,
                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                               const float tan_fovx, const float tan_fovy,
                               const torch::Tensor& dL_dout_color, const torch::Tensor& sh,
                               const int degree, const torch::Tensor& campos,
                               const torch::Tensor& geomBuffer, const int R,
                               const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
                               const bool debug)
{
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    const int H = (int)dL_dout_color.size(1), W = (int)dL_dout_color.size(2);
    const int M = sh.size(0) != 0 ? (int)sh.size(1) : 0;
    auto o = means3D.options().dtype(torch::kFloat32);
    // torch::empty, not the reference's nine torch::zeros (rasterize_points.cu:151-159): bgr_backward
    // writes every element, culled rows as zeros
    auto dL_dmeans3D = torch::empty({P, 3}, o), dL_dmeans2D = torch::empty({P, 3}, o),
         dL_dcolors = torch::empty({P, 3}, o), dL_dopacity = torch::empty({P, 1}, o),
         dL_dcov3D = torch::empty({P, 6}, o), dL_dsh = torch::empty({P, M, 3}, o),
         dL_dscales = torch::empty({P, 3}, o), dL_drotations = torch::empty({P, 4}, o);
    if (P != 0) {
        auto scratch = torch::empty({(long long)bgr_backward_scratch_bytes(P)}, means3D.options().dtype(torch::kByte));
        const auto bg = dense(background), me = dense(means3D), sh_ = dense(sh), co = dense(colors),
                   sc = dense(scales), ro = dense(rotations), cv = dense(cov3D_precomp),
                   vi = dense(viewmatrix), pr = dense(projmatrix), ca = dense(campos),
                   dp = dense(dL_dout_color), ra = radii.contiguous();
        BgrBackwardArgs b{};
        b.P = P; b.D = degree; b.M = M; b.R = R; b.width = W; b.height = H;
        b.background = fptr(bg); b.means3D = fptr(me); b.shs = fptr(sh_); b.colors_precomp = fptr(co);
        b.scales = fptr(sc); b.scale_modifier = scale_modifier; b.rotations = fptr(ro);
        b.cov3D_precomp = fptr(cv); b.viewmatrix = fptr(vi); b.projmatrix = fptr(pr);
        b.cam_pos = fptr(ca); b.tan_fovx = tan_fovx; b.tan_fovy = tan_fovy;
        b.radii = ra.data_ptr<int>();
        b.geom_buffer = geomBuffer.numel() ? geomBuffer.data_ptr() : nullptr;
        b.binning_buffer = binningBuffer.numel() ? binningBuffer.data_ptr() : nullptr;
        b.image_buffer = imageBuffer.numel() ? imageBuffer.data_ptr() : nullptr;
        b.dL_dpix = fptr(dp); b.scratch = scratch.data_ptr();
        b.dL_dmean2D = dL_dmeans2D.data_ptr<float>(); b.dL_dcolor = dL_dcolors.data_ptr<float>();
        b.dL_dopacity = dL_dopacity.data_ptr<float>(); b.dL_dmean3D = dL_dmeans3D.data_ptr<float>();
        b.dL_dcov3D = dL_dcov3D.data_ptr<float>(); b.dL_dsh = M ? dL_dsh.data_ptr<float>() : nullptr;
        b.dL_dscale = dL_dscales.data_ptr<float>(); b.dL_drot = dL_drotations.data_ptr<float>();
        b.debug = debug;
        check(bgr_backward(&b, at::cuda::getCurrentCUDAStream().stream()));
    }
    // rasterize_points.cu:195
    return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
                           dL_dscales, dL_drotations);
}

torch::Tensor markVisible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix)
{
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    torch::Tensor present = torch::full({P}, false, means3D.options().dtype(at::kBool));
    if (P != 0) {
        const auto me = means3D.contiguous(), vi = viewmatrix.contiguous(), pr = projmatrix.contiguous();
        check(bgr_mark_visible(P, me.data_ptr<float>(), vi.data_ptr<float>(), pr.data_ptr<float>(),
                               reinterpret_cast<uint8_t*>(present.data_ptr<bool>()),
                               at::cuda::getCurrentCUDAStream().stream()));
    }
    return present;
}

// DGR/ext.cpp:15-18
PYBIND11_MODULE(TORCH_EXTENSION_NAME, m)
{
    m.def("rasterize_gaussians", &RasterizeGaussiansCUDA);
    m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA);
    m.def("mark_visible", &markVisible);
}
