// TEST INFRASTRUCTURE — not part of the product path.
//
// C-ABI shim around the *unmodified* reference rasterizer core
// (/root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/*.cu), compiled where
// the sources lie by oracle/Makefile into oracle/_ref/libref_rasterizer.so. The shim only
// forwards arguments to CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (rasterizer.h:20-85) and exposes the reference's own scratch layout functions
// (GeometryState/ImageState/BinningState::fromChunk, rasterizer_impl.cu:155-194) so the parity
// tests can read the reference's intermediate buffers. Nothing here is used by the product.
#include <cstdint>
#include <cstddef>
#include <functional>
#include <stdexcept>
#include <cuda_runtime.h>
#include "rasterizer_impl.h"  // reference header, in place (-I .../cuda_rasterizer)

typedef char* (*ref_alloc_fn)(size_t bytes, void* user);

extern "C" {

// Returns num_rendered (>= 0), or -1 when the reference threw (debug mode / non-RGB).
int ref_forward(ref_alloc_fn geom_alloc, void* geom_user,
                ref_alloc_fn binning_alloc, void* binning_user,
                ref_alloc_fn img_alloc, void* img_user,
                int P, int D, int M, const float* background, int width, int height,
                const float* means3D, const float* shs, const float* colors_precomp,
                const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy,
                int prefiltered, float* out_color, int* radii, int debug)
{
    try {
        return CudaRasterizer::Rasterizer::forward(
            [=](size_t n) { return geom_alloc(n, geom_user); },
            [=](size_t n) { return binning_alloc(n, binning_user); },
            [=](size_t n) { return img_alloc(n, img_user); },
            P, D, M, background, width, height, means3D, shs, colors_precomp, opacities, scales,
            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx,
            tan_fovy, prefiltered != 0, out_color, radii, debug != 0);
    } catch (const std::exception&) {
        return -1;
    }
}

int ref_backward(int P, int D, int M, int R, const float* background, int width, int height,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                 const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                 char* geom_buffer, char* binning_buffer, char* image_buffer, const float* dL_dpix,
                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                 float* dL_drot, int debug)
{
    try {
        CudaRasterizer::Rasterizer::backward(
            P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales,
            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
            tan_fovy, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_dmean2D,
            dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
            debug != 0);
        return 0;
    } catch (const std::exception&) {
        return -1;
    }
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, uint8_t* present)
{
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix,
                                            reinterpret_cast<bool*>(present));
}

// Byte offsets (relative to `base`) of every member of the reference's scratch structs, obtained
// by calling the reference's own fromChunk on the real base address.
// geom: depths, clamped, internal_radii, means2D, cov3D, conic_opacity, rgb, tiles_touched,
//       scanning_space, point_offsets, end
void ref_geom_layout(char* base, size_t P, size_t* off)
{
    char* chunk = base;
    auto g = CudaRasterizer::GeometryState::fromChunk(chunk, P);
    off[0] = (char*)g.depths - base;
    off[1] = (char*)g.clamped - base;
    off[2] = (char*)g.internal_radii - base;
    off[3] = (char*)g.means2D - base;
    off[4] = (char*)g.cov3D - base;
    off[5] = (char*)g.conic_opacity - base;
    off[6] = (char*)g.rgb - base;
    off[7] = (char*)g.tiles_touched - base;
    off[8] = (char*)g.scanning_space - base;
    off[9] = (char*)g.point_offsets - base;
    off[10] = chunk - base;
}

// img: accum_alpha, n_contrib, ranges, end
void ref_img_layout(char* base, size_t N, size_t* off)
{
    char* chunk = base;
    auto s = CudaRasterizer::ImageState::fromChunk(chunk, N);
    off[0] = (char*)s.accum_alpha - base;
    off[1] = (char*)s.n_contrib - base;
    off[2] = (char*)s.ranges - base;
    off[3] = chunk - base;
}

// binning: point_list, point_list_unsorted, point_list_keys, point_list_keys_unsorted,
//          list_sorting_space, end
void ref_binning_layout(char* base, size_t R, size_t* off)
{
    char* chunk = base;
    auto b = CudaRasterizer::BinningState::fromChunk(chunk, R);
    off[0] = (char*)b.point_list - base;
    off[1] = (char*)b.point_list_unsorted - base;
    off[2] = (char*)b.point_list_keys - base;
    off[3] = (char*)b.point_list_keys_unsorted - base;
    off[4] = (char*)b.list_sorting_space - base;
    off[5] = chunk - base;
}

}  // extern "C"
