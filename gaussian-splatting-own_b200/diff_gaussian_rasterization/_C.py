"""Native binding: the three functions the reference exposes from its pybind module
(``ext.cpp:15-18``), implemented over the C ABI of ``libbgr.so`` (``include/bgr.h``) with ctypes.

    rasterize_gaussians(...)           <- RasterizeGaussiansCUDA          rasterize_points.cu:35-115
    rasterize_gaussians_backward(...)  <- RasterizeGaussiansBackwardCUDA  rasterize_points.cu:117-196
    mark_visible(...)                  <- markVisible                     rasterize_points.cu:198-217

Same positional arguments, same return tuples, same "empty tensor means absent" convention.
PyTorch is used only for device memory and the current stream. There is NO CPU fallback: if the
shared library is missing or no CUDA device is present the import/call fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent / "libbgr.so"


class _ForwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("width", C.c_int), ("height", C.c_int),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("scale_modifier", C.c_float), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("cam_pos", C.c_void_p),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("prefiltered", C.c_int),
        ("debug", C.c_int), ("out_color", C.c_void_p), ("radii", C.c_void_p),
    ]


class _BackwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("R", C.c_int), ("width", C.c_int),
        ("height", C.c_int), ("background", C.c_void_p), ("means3D", C.c_void_p),
        ("shs", C.c_void_p), ("colors_precomp", C.c_void_p), ("scales", C.c_void_p),
        ("scale_modifier", C.c_float), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("cam_pos", C.c_void_p),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("radii", C.c_void_p),
        ("geom_buffer", C.c_void_p), ("binning_buffer", C.c_void_p), ("image_buffer", C.c_void_p),
        ("dL_dpix", C.c_void_p), ("scratch", C.c_void_p), ("dL_dmean2D", C.c_void_p),
        ("dL_dcolor", C.c_void_p), ("dL_dopacity", C.c_void_p), ("dL_dmean3D", C.c_void_p),
        ("dL_dcov3D", C.c_void_p), ("dL_dsh", C.c_void_p), ("dL_dscale", C.c_void_p),
        ("dL_drot", C.c_void_p), ("debug", C.c_int),
        ("sh_factor_multicast", C.c_void_p), ("campos_multicast", C.c_void_p),
        ("stages", C.c_int), ("gaussian_begin", C.c_int), ("gaussian_end", C.c_int),
    ]


BACKWARD_ACCUMULATE, BACKWARD_PREPROCESS = 1, 2


_ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_size_t, C.c_void_p)

_lib = None


def lib() -> C.CDLL:
    """Loads libbgr.so (once). Raises if it has not been built — never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise ImportError(
            f"{_LIB_PATH} not found: build it with `python gaussian-splatting-own_b200/build.py` "
            "(there is no CPU or PyTorch fallback for the rasterizer)")
    L = C.CDLL(str(_LIB_PATH))
    L.bgr_last_error.restype = C.c_char_p
    L.bgr_version.restype = C.c_char_p
    for name in ("bgr_geom_bytes", "bgr_binning_bytes", "bgr_backward_scratch_bytes"):
        getattr(L, name).restype = C.c_size_t
        getattr(L, name).argtypes = [C.c_int]
    L.bgr_image_bytes.restype = C.c_size_t
    L.bgr_image_bytes.argtypes = [C.c_int, C.c_int]
    L.bgr_forward.restype = C.c_int
    L.bgr_forward.argtypes = [C.POINTER(_ForwardArgs), _ALLOC_FN, C.c_void_p, _ALLOC_FN, C.c_void_p,
                              _ALLOC_FN, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
    L.bgr_backward.restype = C.c_int
    L.bgr_backward.argtypes = [C.POINTER(_BackwardArgs), C.c_void_p]
    L.bgr_forward_raw.restype = C.c_int
    L.bgr_forward_raw.argtypes = [C.POINTER(_ForwardArgs), C.c_void_p, _ALLOC_FN, C.c_void_p, _ALLOC_FN,
                                  C.c_void_p, _ALLOC_FN, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
    L.bgr_backward_raw.restype = C.c_int
    L.bgr_backward_raw.argtypes = [C.POINTER(_BackwardArgs), C.c_void_p, C.c_void_p, C.c_void_p]
    L.bgr_loss_scratch_bytes.restype = C.c_size_t
    L.bgr_loss_scratch_bytes.argtypes = [C.c_int, C.c_int]
    L.bgr_l1_ssim_loss.restype = C.c_int
    L.bgr_l1_ssim_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p]
    L.bgr_densification_stats.restype = C.c_int
    L.bgr_densification_stats.argtypes = [C.c_int] + [C.c_void_p] * 6
    L.bgr_adam_step.restype = C.c_int
    L.bgr_adam_step.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.POINTER(C.c_double), C.c_double,
                                C.c_double, C.c_double, C.c_int, C.c_void_p]
    L.bgr_mark_visible.restype = C.c_int
    L.bgr_mark_visible.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p]
    L.bgr_geom_layout.argtypes = [C.c_int, C.POINTER(C.c_size_t)]
    L.bgr_image_layout.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_size_t)]
    L.bgr_binning_layout.argtypes = [C.c_int, C.POINTER(C.c_size_t)]
    L.bgr_nvls_allreduce_f32.restype = C.c_int
    L.bgr_nvls_allreduce_f32.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int,
                                         C.c_void_p]
    L.bgr_sh_grad_from_factors.restype = C.c_int
    L.bgr_sh_grad_from_factors.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p]
    L.bgr_sh_grad_from_factors_range.restype = C.c_int
    L.bgr_sh_grad_from_factors_range.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                 C.c_void_p]
    L.bgr_nvls_allreduce_segments_f32.restype = C.c_int
    L.bgr_nvls_allreduce_segments_f32.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_size_t),
                                                  C.POINTER(C.c_size_t), C.c_int, C.c_int, C.c_int,
                                                  C.c_void_p]
    L.bgr_launch_count.restype = C.c_ulonglong
    L.bgr_launch_count.argtypes = []
    L.bgr_last_forward_info.restype = None
    L.bgr_last_forward_info.argtypes = [C.POINTER(C.c_int)]
    L.bgr_stage_timing.restype = None
    L.bgr_stage_timing.argtypes = [C.c_int]
    L.bgr_stage_count.restype = C.c_int
    L.bgr_stage_name.restype = C.c_char_p
    L.bgr_stage_name.argtypes = [C.c_int]
    L.bgr_stage_times.restype = C.c_int
    L.bgr_stage_times.argtypes = [C.POINTER(C.c_float), C.c_int]
    L.bgr_render_stats.restype = C.c_int
    L.bgr_render_stats.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.POINTER(C.c_ulonglong), C.c_void_p]
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(f"bgr error {rc}: {lib().bgr_last_error().decode()}")


def _ptr(t: torch.Tensor):
    """Device address of a contiguous float tensor, or None (NULL) for an empty one — the
    reference's convention for absent optional inputs (rasterize_points.cu:94-111)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t: torch.Tensor, dev) -> torch.Tensor:
    """Contiguous f32 tensor on the compute device. Empty tensors stay as they are."""
    if t.numel() == 0:
        return t
    if t.device != dev:
        raise RuntimeError(f"expected a tensor on {dev}, got {t.device}")
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected float32, got {t.dtype}")
    return t.contiguous()


def _stream(dev) -> int:
    # raw handle of the current stream (what torch.cuda.current_stream(dev).cuda_stream returns,
    # without building the Stream object: this runs two or three times per training step)
    if not isinstance(dev, torch.device):
        dev = torch.device(dev) if not isinstance(dev, int) else torch.device("cuda", dev)
    return torch._C._cuda_getCurrentRawStream(dev.index if dev.index is not None else torch.cuda.current_device())


class _on_device:
    """`with torch.cuda.device(dev)` that costs nothing when dev is already current."""

    __slots__ = ("ctx",)

    def __init__(self, dev):
        idx = dev.index
        self.ctx = None if idx is None or idx == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            return self.ctx.__exit__(*exc)
        return False


class _Grower:
    """The resize-callback the core calls to size a scratch buffer (resizeFunctional,
    rasterize_points.cu:27-33). The tensor it creates is handed back to Python.

    The ctypes callback closes over a plain list, not over `self`: a bound-method callback would
    form a reference cycle (self -> cb -> self) and keep hundreds of MB of scratch alive until the
    cyclic GC runs, forcing a fresh cudaMalloc on every call."""

    __slots__ = ("box", "cb")

    def __init__(self, dev):
        box = [None]

        def grow(nbytes, _user):
            try:
                box[0] = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
                return box[0].data_ptr()
            except Exception:  # noqa: BLE001  (reported as BGR_ERR_ALLOC by the library)
                return None

        self.box = box
        self.cb = _ALLOC_FN(grow)

    def take(self, dev):
        """The tensor the core asked for (an empty one if it never did); the grower forgets it."""
        t, self.box[0] = self.box[0], None
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=dev)


_grower_pool = threading.local()


def _take_growers(dev):
    """Three growers (geometry, binning, image) for one forward call. Building a ctypes callback
    costs several microseconds, so finished calls park theirs per thread and device."""
    pool = _grower_pool.__dict__.setdefault("free", {})
    g = pool.pop(dev, None)
    return g if g is not None else (_Grower(dev), _Grower(dev), _Grower(dev))


def _park_growers(dev, g):
    _grower_pool.__dict__.setdefault("free", {})[dev] = g


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                        image_width, sh, degree, campos, prefiltered, debug):
    """-> (num_rendered, out_color(3,H,W) f32, radii(P) i32, geomBuffer, binningBuffer, imgBuffer)"""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("the rasterizer is CUDA-only: means3D must be a CUDA tensor")
    L = lib()
    dev = means3D.device
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    growers = _take_growers(dev)
    geom, binning, img = growers
    with _on_device(dev):
        out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        if P == 0:
            out_color.zero_()
            return 0, out_color, radii, geom.take(dev), binning.take(dev), img.take(dev)
        M = int(sh.size(1)) if sh.size(0) != 0 else 0
        keep = [_f32c(t, dev) for t in (background, means3D, sh, colors, opacity, scales, rotations,
                                        cov3D_precomp, viewmatrix, projmatrix, campos)]
        bg_, means_, sh_, col_, op_, sc_, rot_, cov_, view_, proj_, cam_ = keep
        a = _ForwardArgs(
            P=P, D=int(degree), M=M, width=W, height=H, background=_ptr(bg_), means3D=_ptr(means_),
            shs=_ptr(sh_), colors_precomp=_ptr(col_), opacities=_ptr(op_), scales=_ptr(sc_),
            scale_modifier=float(scale_modifier), rotations=_ptr(rot_), cov3D_precomp=_ptr(cov_),
            viewmatrix=_ptr(view_), projmatrix=_ptr(proj_), cam_pos=_ptr(cam_),
            tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy), prefiltered=int(bool(prefiltered)),
            debug=int(bool(debug)), out_color=out_color.data_ptr(), radii=radii.data_ptr())
        rendered = C.c_int(0)
        rc = L.bgr_forward(C.byref(a), geom.cb, None, binning.cb, None, img.cb, None,
                           _stream(dev), C.byref(rendered))
        _check(rc)
    bufs = (geom.take(dev), binning.take(dev), img.take(dev))
    _park_growers(dev, growers)
    return (int(rendered.value), out_color, radii) + bufs


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations,
                                 scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                                 tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R,
                                 binningBuffer, imageBuffer, debug):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
    dL_drotations), shapes as rasterize_points.cu:151-159,195."""
    L = lib()
    dev = means3D.device
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    M = int(sh.size(1)) if sh.size(0) != 0 else 0
    with _on_device(dev):
        # One arena for all outputs: every element is written by the kernels, so no zero fill is
        # needed. Layout (floats per Gaussian):
        #   [means2D 3 | colors 3 | cov3D 6 || means3D 3 | opacity 1 | scales 3 | rotations 4 | sh 3M]
        # The part after || is what a trainer's leaves receive (59 floats at M = 16) and is
        # contiguous so that a data-parallel caller can sum it across ranks in place with one
        # collective — the kernels write straight into the send buffer. The small leaves come
        # first (dense exchange), rotations + sh last (row-sparse exchange, view_parallel.py).
        widths = (3, 3, 6, 3, 1, 3, 4, 3 * M)
        n_arena = P * sum(widths)
        arena = _grad_arena_provider(n_arena, dev) if _grad_arena_provider else None
        if arena is None:
            arena = torch.empty((n_arena,), dtype=torch.float32, device=dev)
        elif arena.numel() != n_arena or arena.dtype != torch.float32 or not arena.is_contiguous():
            raise RuntimeError("gradient arena provider returned a tensor of the wrong size/dtype")
        parts, o = [], 0
        for w in widths:
            parts.append(arena[o:o + P * w])
            o += P * w
        dL_dmeans2D = parts[0].view(P, 3)
        dL_dcolors = parts[1].view(P, 3)
        dL_dcov3D = parts[2].view(P, 6)
        dL_dmeans3D = parts[3].view(P, 3)
        dL_dopacity = parts[4].view(P, 1)
        dL_dscales = parts[5].view(P, 3)
        dL_drotations = parts[6].view(P, 4)
        dL_dsh = parts[7].view(P, M, 3)
        if P != 0:
            scratch = torch.empty((L.bgr_backward_scratch_bytes(P),), dtype=torch.uint8, device=dev)
            keep = [_f32c(t, dev) for t in (background, means3D, sh, colors, scales, rotations,
                                            cov3D_precomp, viewmatrix, projmatrix, campos,
                                            dL_dout_color)]
            bg_, means_, sh_, col_, sc_, rot_, cov_, view_, proj_, cam_, dpix_ = keep
            radii_ = radii.contiguous()
            a = _BackwardArgs(
                P=P, D=int(degree), M=M, R=int(R), width=W, height=H, background=_ptr(bg_),
                means3D=_ptr(means_), shs=_ptr(sh_), colors_precomp=_ptr(col_), scales=_ptr(sc_),
                scale_modifier=float(scale_modifier), rotations=_ptr(rot_),
                cov3D_precomp=_ptr(cov_), viewmatrix=_ptr(view_), projmatrix=_ptr(proj_),
                cam_pos=_ptr(cam_), tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy),
                radii=radii_.data_ptr(), geom_buffer=_ptr(geomBuffer),
                binning_buffer=_ptr(binningBuffer), image_buffer=_ptr(imageBuffer),
                dL_dpix=_ptr(dpix_), scratch=scratch.data_ptr(), dL_dmean2D=_ptr(dL_dmeans2D),
                dL_dcolor=_ptr(dL_dcolors), dL_dopacity=_ptr(dL_dopacity),
                dL_dmean3D=_ptr(dL_dmeans3D), dL_dcov3D=_ptr(dL_dcov3D), dL_dsh=_ptr(dL_dsh),
                dL_dscale=_ptr(dL_dscales), dL_drot=_ptr(dL_drotations), debug=int(bool(debug)))
            chunks = (_backward_ranges, lambda p0, p1: None) if _backward_ranges else None
            if _sh_factor_hook is not None and M > 0:
                # view-parallel path: multicast dL/dRGB instead of writing dL_dsh (see bgr.h)
                slots = _sh_factor_hook(P, M, int(degree), means_, dL_dsh)
                if slots is not None:
                    a.sh_factor_multicast, a.campos_multicast = slots[0], slots[1]
                    chunks = slots[2] if len(slots) > 2 else None
            if chunks is None:
                _check(L.bgr_backward(C.byref(a), _stream(dev)))
            else:
                # pipelined exchange: (bounds, after_chunk) — the hook's owner exchanges the
                # gradients of each finished Gaussian range while the next one is computed
                bounds, after_chunk = chunks
                a.stages = BACKWARD_ACCUMULATE
                _check(L.bgr_backward(C.byref(a), _stream(dev)))
                a.stages = BACKWARD_PREPROCESS
                for p0, p1 in zip(bounds[:-1], bounds[1:]):
                    a.gaussian_begin, a.gaussian_end = int(p0), int(p1)
                    _check(L.bgr_backward(C.byref(a), _stream(dev)))
                    after_chunk(int(p0), int(p1))
            # scratch was allocated on the stream the kernels run on: the caching allocator hands
            # it out again only to later work of that same stream, so no record_stream is needed
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
            dL_drotations)


# ---- extended entry: raw leaves in, raw-leaf gradients out ---------------
def rasterize_gaussians_raw(background, xyz, opacity_raw, scaling_raw, rotation_raw, scale_modifier,
                            viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width,
                            features_dc, features_rest, degree, campos, prefiltered, debug):
    """Forward over the trainer's RAW leaves (scene/gaussian_model.py:95-115): `_opacity` (P,1)
    logits, `_scaling` (P,3) log-scales, `_rotation` (P,4) unnormalised, `_features_dc` (P,1,3) and
    `_features_rest` (P,M-1,3); sigmoid / exp / normalize / cat happen inside the preprocess kernel
    (bgr_forward_raw). Same return tuple as rasterize_gaussians."""
    if xyz.dim() != 2 or xyz.size(1) != 3:
        raise RuntimeError("xyz must have dimensions (num_points, 3)")
    if not xyz.is_cuda:
        raise RuntimeError("the rasterizer is CUDA-only: xyz must be a CUDA tensor")
    L = lib()
    dev = xyz.device
    P, H, W = int(xyz.size(0)), int(image_height), int(image_width)
    growers = _take_growers(dev)
    geom, binning, img = growers
    with _on_device(dev):
        out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        if P == 0:
            out_color.zero_()
            return 0, out_color, radii, geom.take(dev), binning.take(dev), img.take(dev)
        if features_dc.dim() != 3 or features_dc.size(1) != 1 or features_rest.dim() != 3:
            raise RuntimeError("features_dc must be (P,1,3) and features_rest (P,M-1,3)")
        M = 1 + int(features_rest.size(1))
        keep = [_f32c(t, dev) for t in (background, xyz, features_dc, features_rest, opacity_raw,
                                        scaling_raw, rotation_raw, viewmatrix, projmatrix, campos)]
        bg_, xyz_, dc_, rest_, op_, sc_, rot_, view_, proj_, cam_ = keep
        a = _ForwardArgs(
            P=P, D=int(degree), M=M, width=W, height=H, background=_ptr(bg_), means3D=_ptr(xyz_),
            shs=_ptr(dc_), colors_precomp=None, opacities=_ptr(op_), scales=_ptr(sc_),
            scale_modifier=float(scale_modifier), rotations=_ptr(rot_), cov3D_precomp=None,
            viewmatrix=_ptr(view_), projmatrix=_ptr(proj_), cam_pos=_ptr(cam_),
            tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy), prefiltered=int(bool(prefiltered)),
            debug=int(bool(debug)), out_color=out_color.data_ptr(), radii=radii.data_ptr())
        rendered = C.c_int(0)
        rc = L.bgr_forward_raw(C.byref(a), _ptr(rest_), geom.cb, None, binning.cb, None, img.cb, None,
                               _stream(dev), C.byref(rendered))
        _check(rc)
    bufs = (geom.take(dev), binning.take(dev), img.take(dev))
    _park_growers(dev, growers)
    return (int(rendered.value), out_color, radii) + bufs


def rasterize_gaussians_raw_backward(background, xyz, radii, scaling_raw, rotation_raw, scale_modifier,
                                     viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                     features_dc, features_rest, degree, campos, geomBuffer, R,
                                     binningBuffer, imageBuffer, debug):
    """-> (dL_dmeans2D (P,3), dL_dopacity_raw (P,1), dL_dxyz (P,3), dL_dfeatures_dc (P,1,3),
    dL_dfeatures_rest (P,M-1,3), dL_dscaling_raw (P,3), dL_drotation_raw (P,4)): gradients of the
    raw leaves, the activations' derivatives applied inside the backward preprocess kernel."""
    L = lib()
    dev = xyz.device
    P = int(xyz.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    M = 1 + int(features_rest.size(1))
    with _on_device(dev):
        # one arena; [colors 3 | cov3D 6] are by-products nobody reads on this path
        widths = (3, 3, 6, 3, 1, 3, 4, 3, 3 * (M - 1))
        arena = torch.empty((P * sum(widths),), dtype=torch.float32, device=dev)
        parts, o = [], 0
        for w in widths:
            parts.append(arena[o:o + P * w])
            o += P * w
        dL_dmeans2D, dL_dcolors, dL_dcov3D = parts[0].view(P, 3), parts[1].view(P, 3), parts[2].view(P, 6)
        dL_dxyz, dL_dopacity = parts[3].view(P, 3), parts[4].view(P, 1)
        dL_dscaling, dL_drotation = parts[5].view(P, 3), parts[6].view(P, 4)
        dL_ddc, dL_drest = parts[7].view(P, 1, 3), parts[8].view(P, M - 1, 3)
        if P != 0:
            scratch = torch.empty((L.bgr_backward_scratch_bytes(P),), dtype=torch.uint8, device=dev)
            keep = [_f32c(t, dev) for t in (background, xyz, features_dc, features_rest, scaling_raw,
                                            rotation_raw, viewmatrix, projmatrix, campos, dL_dout_color)]
            bg_, xyz_, dc_, rest_, sc_, rot_, view_, proj_, cam_, dpix_ = keep
            radii_ = radii.contiguous()
            a = _BackwardArgs(
                P=P, D=int(degree), M=M, R=int(R), width=W, height=H, background=_ptr(bg_),
                means3D=_ptr(xyz_), shs=_ptr(dc_), colors_precomp=None, scales=_ptr(sc_),
                scale_modifier=float(scale_modifier), rotations=_ptr(rot_), cov3D_precomp=None,
                viewmatrix=_ptr(view_), projmatrix=_ptr(proj_), cam_pos=_ptr(cam_),
                tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy), radii=radii_.data_ptr(),
                geom_buffer=_ptr(geomBuffer), binning_buffer=_ptr(binningBuffer),
                image_buffer=_ptr(imageBuffer), dL_dpix=_ptr(dpix_), scratch=scratch.data_ptr(),
                dL_dmean2D=_ptr(dL_dmeans2D), dL_dcolor=_ptr(dL_dcolors), dL_dopacity=_ptr(dL_dopacity),
                dL_dmean3D=_ptr(dL_dxyz), dL_dcov3D=_ptr(dL_dcov3D), dL_dsh=_ptr(dL_ddc),
                dL_dscale=_ptr(dL_dscaling), dL_drot=_ptr(dL_drotation), debug=int(bool(debug)))
            _check(L.bgr_backward_raw(C.byref(a), _ptr(rest_), _ptr(dL_drest), _stream(dev)))
    return dL_dmeans2D, dL_dopacity, dL_dxyz, dL_ddc, dL_drest, dL_dscaling, dL_drotation


# ---- the step after the path ------------------------------------------------
def l1_ssim_loss(image, gt, lambda_dssim, need_grad=True):
    """-> (loss3, dL_dimage): loss3 = device tensor {loss, l1, ssim} of train.py:91-92's loss
    (utils/loss_utils.py:16-63), dL_dimage = d loss / d image (3,H,W) or None."""
    if not image.is_cuda or image.dim() != 3 or image.size(0) != 3 or gt.shape != image.shape:
        raise RuntimeError("l1_ssim_loss needs two CUDA tensors of shape (3, H, W)")
    L = lib()
    dev = image.device
    H, W = int(image.size(1)), int(image.size(2))
    with _on_device(dev):
        img_, gt_ = _f32c(image, dev), _f32c(gt, dev)
        scratch = torch.empty((L.bgr_loss_scratch_bytes(W, H),), dtype=torch.uint8, device=dev)
        loss3 = torch.empty((3,), dtype=torch.float32, device=dev)
        grad = torch.empty_like(img_) if need_grad else None
        _check(L.bgr_l1_ssim_loss(img_.data_ptr(), gt_.data_ptr(), W, H, float(lambda_dssim), scratch.data_ptr(),
                                  loss3.data_ptr(), grad.data_ptr() if need_grad else None, _stream(dev)))
    return loss3, grad


def densification_stats(radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom):
    """train.py:114-116 + scene/gaussian_model.py:405-407 for one view, in place, in one kernel."""
    P = int(radii.numel())
    dev = radii.device
    ok = radii.is_cuda and radii.dtype == torch.int32 and radii.is_contiguous() and \
        dL_dmeans2D.shape == (P, 3) and dL_dmeans2D.dtype == torch.float32 and dL_dmeans2D.is_contiguous()
    for t in (dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom):
        ok = ok and t.is_cuda and t.device == dev
    for t in (max_radii2D, xyz_gradient_accum, denom):
        ok = ok and t.numel() == P and t.dtype == torch.float32 and t.is_contiguous()
    if not ok:
        raise RuntimeError("densification_stats needs contiguous CUDA tensors: radii int32 [P], dL_dmeans2D f32 "
                           "[P,3], three f32 accumulators of P elements")
    with _on_device(dev):
        _check(lib().bgr_densification_stats(P, radii.data_ptr(), dL_dmeans2D.data_ptr(), max_radii2D.data_ptr(),
                                             xyz_gradient_accum.data_ptr(), denom.data_ptr(), _stream(dev)))


def adam_step(params, grads, exp_avgs, exp_avg_sqs, lrs, beta1, beta2, eps, step):
    """torch.optim.Adam's update of up to eight parameter groups (one contiguous f32 CUDA tensor each,
    one learning rate each) in one kernel launch; `step` is the 1-based update count."""
    n = len(params)
    if not (1 <= n <= 8) or not (len(grads) == len(exp_avgs) == len(exp_avg_sqs) == len(lrs) == n):
        raise RuntimeError("adam_step takes 1..8 groups with one gradient, two moments and one lr each")
    dev = params[0].device
    for group in (params, grads, exp_avgs, exp_avg_sqs):
        for t, p_ in zip(group, params):
            if not t.is_cuda or t.device != dev or t.dtype != torch.float32 or not t.is_contiguous() \
                    or t.numel() != p_.numel():
                raise RuntimeError("adam_step needs contiguous float32 CUDA tensors of matching sizes on one device")
    arr = lambda ts: (C.c_void_p * n)(*[t.data_ptr() for t in ts])  # noqa: E731
    numel = (C.c_size_t * n)(*[p_.numel() for p_ in params])
    lr = (C.c_double * n)(*[float(x) for x in lrs])
    with _on_device(dev):
        _check(lib().bgr_adam_step(n, arr(params), arr(grads), arr(exp_avgs), arr(exp_avg_sqs), numel, lr,
                                   float(beta1), float(beta2), float(eps), int(step), _stream(dev)))


def mark_visible(means3D, viewmatrix, projmatrix):
    """-> bool tensor (P,), True where the Gaussian is in front of the near plane."""
    if not means3D.is_cuda:
        raise RuntimeError("the rasterizer is CUDA-only: means3D must be a CUDA tensor")
    dev = means3D.device
    P = int(means3D.size(0))
    with _on_device(dev):
        present = torch.zeros((P,), dtype=torch.bool, device=dev)
        if P != 0:
            m, v, pj = _f32c(means3D, dev), _f32c(viewmatrix, dev), _f32c(projmatrix, dev)
            _check(lib().bgr_mark_visible(P, m.data_ptr(), v.data_ptr(), pj.data_ptr(),
                                          present.data_ptr(), _stream(dev)))
    return present


_grad_arena_provider = None


def set_grad_arena_provider(fn) -> None:
    """fn(numel, device) -> flat f32 tensor (or None) that rasterize_gaussians_backward uses for its
    outputs instead of a fresh allocation. The view-parallel path points it at a buffer in
    symmetric (peer-mapped, multicast-bound) memory so that the backward kernels write straight
    into the memory the NVLS all-reduce operates on. The previous call's gradients are
    overwritten by the next call."""
    global _grad_arena_provider
    _grad_arena_provider = fn


def nvls_allreduce(multicast_ptr: int, offset_floats: int, count_floats: int, rank: int, world: int,
                   device, max_blocks: int = 296) -> None:
    with torch.cuda.device(device):
        _check(lib().bgr_nvls_allreduce_f32(C.c_void_p(multicast_ptr), offset_floats, count_floats,
                                            rank, world, max_blocks, _stream(device)))


_sh_factor_hook = None
_backward_ranges = None


def set_backward_ranges(bounds) -> None:
    """bounds = [0, ..., P] (or None): run the backward as ACCUMULATE + one PREPROCESS call per
    Gaussian range (BgrBackwardArgs.stages). Results are identical to the single call; the
    view-parallel exchange uses the same split to overlap communication (view_parallel.py)."""
    global _backward_ranges
    _backward_ranges = list(bounds) if bounds else None


def set_sh_factor_hook(fn) -> None:
    """fn(P, M, degree, means3D, dL_dsh) -> (factor_slot_multicast_ptr, campos_slot_multicast_ptr) or
    None. When it returns pointers the backward multicasts the SH gradient factors instead of
    writing dL_dsh; the hook's owner later fills dL_dsh with sh_grad_from_factors()."""
    global _sh_factor_hook
    _sh_factor_hook = fn


def sh_grad_from_factors(P, D, M, world, means3D, campos_all_ptr: int, factors_all_ptr: int, dL_dsh,
                         device) -> None:
    with torch.cuda.device(device):
        _check(lib().bgr_sh_grad_from_factors(int(P), int(D), int(M), int(world),
                                              C.c_void_p(means3D.data_ptr()), C.c_void_p(campos_all_ptr),
                                              C.c_void_p(factors_all_ptr), C.c_void_p(dL_dsh.data_ptr()),
                                              _stream(device)))


def sh_grad_from_factors_range(P, begin, end, D, M, world, means3D, campos_all_ptr: int,
                               factors_all_ptr: int, dL_dsh, device) -> None:
    with torch.cuda.device(device):
        _check(lib().bgr_sh_grad_from_factors_range(
            int(P), int(begin), int(end), int(D), int(M), int(world), C.c_void_p(means3D.data_ptr()),
            C.c_void_p(campos_all_ptr), C.c_void_p(factors_all_ptr), C.c_void_p(dL_dsh.data_ptr()),
            _stream(device)))


def nvls_allreduce_segments(multicast_ptr: int, segments, rank: int, world: int, device,
                            max_blocks: int = 296) -> None:
    """segments: up to four (offset_floats, count_floats) pairs, all multiples of 4."""
    n = len(segments)
    offs = (C.c_size_t * max(n, 1))(*[int(o) for o, _ in segments])
    cnts = (C.c_size_t * max(n, 1))(*[int(c) for _, c in segments])
    with torch.cuda.device(device):
        _check(lib().bgr_nvls_allreduce_segments_f32(C.c_void_p(multicast_ptr), n, offs, cnts, rank, world,
                                                     max_blocks, _stream(device)))


def leaf_grad_arena(grads) -> torch.Tensor:
    """The flat f32 tensor that aliases (dL_dmeans3D, dL_dopacity, dL_dscales, dL_drotations,
    dL_dsh) of one rasterize_gaussians_backward result, in that (memory) order."""
    dL_dmeans3D, dL_dsh = grads[3], grads[5]
    arena = dL_dmeans3D._base
    P = dL_dmeans3D.size(0)
    start = 12 * P
    n = P * (3 + 3 * dL_dsh.size(1) + 1 + 3 + 4)
    flat = arena[start:start + n]
    assert flat.data_ptr() == dL_dmeans3D.data_ptr()
    return flat


# ---- diagnostics -------------------------------------------------------------------------------
def launch_count() -> int:
    return int(lib().bgr_launch_count())


def last_forward_info() -> dict:
    out = (C.c_int * 8)()
    lib().bgr_last_forward_info(out)
    return dict(zip(("global_sort_path", "nb_idx", "nb_depth", "_3", "max_tile_list", "smem_keys", "_6", "R"),
                    [int(x) for x in out]))


def stage_timing(enable: bool) -> None:
    lib().bgr_stage_timing(int(bool(enable)))


def stage_times() -> dict:
    L = lib()
    n = L.bgr_stage_count()
    buf = (C.c_float * n)()
    L.bgr_stage_times(buf, n)
    return {L.bgr_stage_name(i).decode(): float(buf[i]) for i in range(n) if buf[i] >= 0}


def render_stats(fwd_out, P, W, H) -> dict:
    R, _c, _r, geom, binning, img = fwd_out
    out = (C.c_ulonglong * 4)()
    dev = geom.device
    with _on_device(dev):
        _check(lib().bgr_render_stats(int(P), int(R), int(W), int(H), _ptr(geom), _ptr(binning),
                                      _ptr(img), out, _stream(dev)))
    return dict(zip(("staged", "examined", "exp_evaluated", "blended"), [int(x) for x in out]))


# ---- introspection helpers used by the parity tests ------------------------------------------
def geom_layout(P):
    o = (C.c_size_t * 2)()
    lib().bgr_geom_layout(int(P), o)
    return dict(zip(("rec", "tiles_touched"), [int(x) for x in o]))


def image_layout(W, H):
    o = (C.c_size_t * 4)()
    lib().bgr_image_layout(int(W), int(H), o)
    return dict(zip(("final_T", "n_contrib", "ranges", "control"), [int(x) for x in o]))


def binning_layout(R):
    o = (C.c_size_t * 3)()
    lib().bgr_binning_layout(int(R), o)
    return dict(zip(("point_list", "keys", "keys_alt"), [int(x) for x in o]))
