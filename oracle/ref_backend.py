"""TEST INFRASTRUCTURE — the reference's own CUDA rasterizer behind the `_C` call signatures.

Loads oracle/_ref/libref_rasterizer.so (the UNMODIFIED reference .cu files compiled for sm_100a by
oracle/Makefile + the forwarding shim oracle/ref_shim.cu) and re-states, in Python, exactly what
the reference's torch shim does around the core (rasterize_points.cu:35-217): the same output
allocations (`torch.full` / nine `torch.zeros`), the same three byte buffers grown through resize
callbacks, the same argument order and return tuples. Used as

  * the GPU oracle of the parity tests (tests/ -m gpu), including its intermediate buffers, and
  * the reference arm of bench.py (`--impl reference`).

Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent / "_ref" / "libref_rasterizer.so"
_ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_size_t, C.c_void_p)
_lib = None


def available() -> bool:
    return _LIB_PATH.exists()


def lib():
    global _lib
    if _lib is None:
        if not _LIB_PATH.exists():
            raise ImportError(f"{_LIB_PATH} missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(str(_LIB_PATH))
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.ref_forward.restype = ci
        L.ref_forward.argtypes = [_ALLOC_FN, vp, _ALLOC_FN, vp, _ALLOC_FN, vp, ci, ci, ci, vp, ci, ci,
                                  vp, vp, vp, vp, vp, cf, vp, vp, vp, vp, vp, cf, cf, ci, vp, vp, ci]
        L.ref_backward.restype = ci
        L.ref_backward.argtypes = [ci, ci, ci, ci, vp, ci, ci, vp, vp, vp, vp, cf, vp, vp, vp, vp, vp,
                                   cf, cf, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci]
        L.ref_mark_visible.restype = None
        L.ref_mark_visible.argtypes = [ci, vp, vp, vp, vp]
        for n in ("ref_geom_layout", "ref_img_layout", "ref_binning_layout"):
            getattr(L, n).restype = None
            getattr(L, n).argtypes = [vp, C.c_size_t, C.POINTER(C.c_size_t)]
        _lib = L
    return _lib


def _p(t):
    # `.contiguous().data<float>()` of an empty tensor is nullptr in the reference
    return None if t.numel() == 0 else t.data_ptr()


class _Resizer:
    """resizeFunctional (rasterize_points.cu:27-33): t.resize_({N}); return t.data_ptr().
    The callback closes over the tensor only (no self reference => no reference cycle, the
    buffers are released by refcount exactly as the reference's pybind module releases them)."""

    def __init__(self):
        t = torch.empty(0, dtype=torch.uint8, device="cuda")

        def f(n, _u):
            t.resize_(int(n))
            return t.data_ptr()

        self.t = t
        self.cb = _ALLOC_FN(f)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                        image_width, sh, degree, campos, prefiltered, debug):
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    L = lib()
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    out_color = torch.full((3, H, W), 0.0, dtype=torch.float32, device=means3D.device)
    radii = torch.full((P,), 0, dtype=torch.int32, device=means3D.device)
    g, b, i = _Resizer(), _Resizer(), _Resizer()
    rendered = 0
    if P != 0:
        M = sh.size(1) if sh.size(0) != 0 else 0
        ts = [t.contiguous() for t in (background, means3D, sh, colors, opacity, scales, rotations,
                                       cov3D_precomp, viewmatrix, projmatrix, campos)]
        bg_, me_, sh_, co_, op_, sc_, ro_, cv_, vi_, pr_, ca_ = ts
        rendered = L.ref_forward(g.cb, None, b.cb, None, i.cb, None, P, int(degree), M, _p(bg_), W, H,
                                 _p(me_), _p(sh_), _p(co_), _p(op_), _p(sc_), float(scale_modifier),
                                 _p(ro_), _p(cv_), _p(vi_), _p(pr_), _p(ca_), float(tan_fovx),
                                 float(tan_fovy), int(bool(prefiltered)), out_color.data_ptr(),
                                 radii.data_ptr(), int(bool(debug)))
        if rendered < 0:
            raise RuntimeError("reference rasterizer forward threw")
    return rendered, out_color, radii, g.t, b.t, i.t


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations,
                                 scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                                 tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R,
                                 binningBuffer, imageBuffer, debug):
    L = lib()
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.size(0) != 0 else 0
    o = dict(dtype=means3D.dtype, device=means3D.device)
    dL_dmeans3D = torch.zeros((P, 3), **o)
    dL_dmeans2D = torch.zeros((P, 3), **o)
    dL_dcolors = torch.zeros((P, 3), **o)
    dL_dconic = torch.zeros((P, 2, 2), **o)
    dL_dopacity = torch.zeros((P, 1), **o)
    dL_dcov3D = torch.zeros((P, 6), **o)
    dL_dsh = torch.zeros((P, M, 3), **o)
    dL_dscales = torch.zeros((P, 3), **o)
    dL_drotations = torch.zeros((P, 4), **o)
    if P != 0:
        ts = [t.contiguous() for t in (background, means3D, sh, colors, scales, rotations,
                                       cov3D_precomp, viewmatrix, projmatrix, campos, dL_dout_color)]
        bg_, me_, sh_, co_, sc_, ro_, cv_, vi_, pr_, ca_, dp_ = ts
        rc = L.ref_backward(P, int(degree), M, int(R), _p(bg_), W, H, _p(me_), _p(sh_), _p(co_),
                            _p(sc_), float(scale_modifier), _p(ro_), _p(cv_), _p(vi_), _p(pr_),
                            _p(ca_), float(tan_fovx), float(tan_fovy), radii.contiguous().data_ptr(),
                            geomBuffer.data_ptr(), binningBuffer.data_ptr(), imageBuffer.data_ptr(),
                            _p(dp_), dL_dmeans2D.data_ptr(), dL_dconic.data_ptr(),
                            dL_dopacity.data_ptr(), dL_dcolors.data_ptr(), dL_dmeans3D.data_ptr(),
                            dL_dcov3D.data_ptr(), _p(dL_dsh), dL_dscales.data_ptr(),
                            dL_drotations.data_ptr(), int(bool(debug)))
        if rc != 0:
            raise RuntimeError("reference rasterizer backward threw")
    rasterize_gaussians_backward.last_dL_dconic = dL_dconic
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
            dL_drotations)


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.size(0)
    present = torch.full((P,), False, dtype=torch.bool, device=means3D.device)
    if P != 0:
        lib().ref_mark_visible(P, means3D.contiguous().data_ptr(), viewmatrix.contiguous().data_ptr(),
                               projmatrix.contiguous().data_ptr(), present.data_ptr())
    return present


# ---- the reference's intermediate buffers, parsed with the reference's own layout code ---------
def _view(buf, off, count, dtype):
    nbytes = count * torch.empty((), dtype=dtype).element_size()
    return buf[off:off + nbytes].view(dtype)


def parse_geom(geomBuffer, P):
    off = (C.c_size_t * 11)()
    lib().ref_geom_layout(geomBuffer.data_ptr(), P, off)
    return {
        "depths": _view(geomBuffer, off[0], P, torch.float32),
        "clamped": _view(geomBuffer, off[1], 3 * P, torch.uint8).view(P, 3),
        "means2D": _view(geomBuffer, off[3], 2 * P, torch.float32).view(P, 2),
        "cov3D": _view(geomBuffer, off[4], 6 * P, torch.float32).view(P, 6),
        "conic_opacity": _view(geomBuffer, off[5], 4 * P, torch.float32).view(P, 4),
        "rgb": _view(geomBuffer, off[6], 3 * P, torch.float32).view(P, 3),
        "tiles_touched": _view(geomBuffer, off[7], P, torch.int32),
        "point_offsets": _view(geomBuffer, off[9], P, torch.int32),
    }


def parse_img(imgBuffer, W, H):
    N = W * H
    off = (C.c_size_t * 4)()
    lib().ref_img_layout(imgBuffer.data_ptr(), N, off)
    return {
        "final_T": _view(imgBuffer, off[0], N, torch.float32).view(H, W),
        "n_contrib": _view(imgBuffer, off[1], N, torch.int32).view(H, W),
        "ranges": _view(imgBuffer, off[2], 2 * N, torch.int32).view(N, 2),
    }


def parse_binning(binningBuffer, R):
    off = (C.c_size_t * 6)()
    lib().ref_binning_layout(binningBuffer.data_ptr(), R, off)
    return {
        "point_list": _view(binningBuffer, off[0], R, torch.int32),
        "point_list_keys": _view(binningBuffer, off[2], R, torch.int64),
    }
