"""Drop-in replacement for the reference's ``diff_gaussian_rasterization`` package
(reference: submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py).

Public surface, kept signature-for-signature:

* ``GaussianRasterizationSettings``  (reference __init__.py:157-169) — 12-field NamedTuple
* ``GaussianRasterizer``             (reference __init__.py:171-220) — ``forward`` / ``markVisible``
* ``rasterize_gaussians``            (reference __init__.py:21-42)
* ``_RasterizeGaussians``            (reference __init__.py:44-155)  — the autograd Function
* ``cpu_deep_copy_tuple``            (reference __init__.py:17-19)
* ``_C``                             — native module (here: ctypes over libbgr.so, sm_100a CUDA)

The autograd Function is produced by ``make_rasterize_function(backend)`` so that the parity
tests and the bench can run the very same Python glue over the reference's own CUDA code
(oracle/ref_backend.py) — the product always uses ``_C``.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

__all__ = [
    "GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
    "_RasterizeGaussians", "cpu_deep_copy_tuple", "make_rasterize_function",
    "RawGaussianRasterizer", "_RasterizeGaussiansRaw",
]


def cpu_deep_copy_tuple(input_tuple):
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _call_native(fn, args, debug, dump_name, what):
    """Debug mode keeps a CPU copy of the arguments and writes it out if the native call raises
    (reference __init__.py:83-90, 132-139)."""
    if not debug:
        return fn(*args)
    cpu_args = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(cpu_args, dump_name)
        print(f"\nAn error occured in {what}. Please forward {dump_name} for debugging.")
        raise


def make_rasterize_function(backend):
    """Builds the autograd Function over a backend exposing ``rasterize_gaussians`` and
    ``rasterize_gaussians_backward`` with the reference's ``_C`` signatures."""

    class _Rasterize(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                    cov3Ds_precomp, raster_settings):
            rs = raster_settings
            native_args = (
                rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered,
                rs.debug,
            )
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _call_native(
                backend.rasterize_gaussians, native_args, rs.debug, "snapshot_fw.dump", "forward")
            ctx.raster_settings = rs
            ctx.num_rendered = num_rendered
            ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii,
                                  sh, geomBuffer, binningBuffer, imgBuffer)
            return color, radii

        @staticmethod
        def backward(ctx, grad_out_color, _grad_radii):
            rs = ctx.raster_settings
            (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
             binningBuffer, imgBuffer) = ctx.saved_tensors
            native_args = (
                rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                grad_out_color, sh, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered,
                binningBuffer, imgBuffer, rs.debug,
            )
            (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
             grad_sh, grad_scales, grad_rotations) = _call_native(
                backend.rasterize_gaussians_backward, native_args, rs.debug, "snapshot_bw.dump",
                "backward")
            # forward-input order: means3D, means2D, sh, colors_precomp, opacities, scales,
            # rotations, cov3Ds_precomp, raster_settings
            return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities,
                    grad_scales, grad_rotations, grad_cov3Ds_precomp, None)

    return _Rasterize


_RasterizeGaussians = make_rasterize_function(_C)
_RasterizeGaussians.__name__ = "_RasterizeGaussians"
_RasterizeGaussians.__qualname__ = "_RasterizeGaussians"


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


def _absent() -> torch.Tensor:
    # empty CPU tensor == "not provided" for the native layer (reference __init__.py:197-207)
    return torch.Tensor([])


class GaussianRasterizer(nn.Module):
    _rasterize = staticmethod(rasterize_gaussians)
    _backend = _C

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return self._backend.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr = scales is not None or rotations is not None
        sr_complete = scales is not None and rotations is not None
        if (not sr_complete and cov3D_precomp is None) or (has_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or '
                            'precomputed 3D covariance!')
        shs = _absent() if shs is None else shs
        colors_precomp = _absent() if colors_precomp is None else colors_precomp
        scales = _absent() if scales is None else scales
        rotations = _absent() if rotations is None else rotations
        cov3D_precomp = _absent() if cov3D_precomp is None else cov3D_precomp
        return self._rasterize(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                               cov3D_precomp, self.raster_settings)


# ---- extended entry beside the drop-in one ----------------------------------
class _RasterizeGaussiansRaw(torch.autograd.Function):
    """The trainer's RAW leaves in, gradients of the raw leaves out: what
    gaussian_renderer/__init__.py:53-82 + scene/gaussian_model.py:95-115 compute with four
    elementwise kernels and a 192-byte-per-Gaussian SH concatenation in front of the rasterizer
    (and their backward behind it) happens inside the preprocess kernels."""

    @staticmethod
    def forward(ctx, xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, xyz, opacity_raw, scaling_raw, rotation_raw, rs.scale_modifier, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, features_dc,
                features_rest, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _call_native(
            _C.rasterize_gaussians_raw, args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(xyz, scaling_raw, rotation_raw, radii, features_dc, features_rest,
                              geomBuffer, binningBuffer, imgBuffer)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii):
        rs = ctx.raster_settings
        (xyz, scaling_raw, rotation_raw, radii, features_dc, features_rest, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (rs.bg, xyz, radii, scaling_raw, rotation_raw, rs.scale_modifier, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, features_dc, features_rest,
                rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        (g_means2D, g_opacity, g_xyz, g_dc, g_rest, g_scaling, g_rotation) = _call_native(
            _C.rasterize_gaussians_raw_backward, args, rs.debug, "snapshot_bw.dump", "backward")
        return g_xyz, g_means2D, g_dc, g_rest, g_opacity, g_scaling, g_rotation, None


class RawGaussianRasterizer(nn.Module):
    """``GaussianRasterizer`` for callers that hold the trainer's parameters: pass `_xyz`,
    `_features_dc`, `_features_rest`, `_opacity`, `_scaling`, `_rotation` as they are stored
    (scene/gaussian_model.py:47-52) instead of their activated views. Returns (color, radii) like
    the drop-in module; `means2D` receives the screen-space gradient the densifier reads."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw):
        return _RasterizeGaussiansRaw.apply(xyz, means2D, features_dc, features_rest, opacity_raw,
                                            scaling_raw, rotation_raw, self.raster_settings)
