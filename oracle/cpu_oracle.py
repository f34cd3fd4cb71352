"""TEST INFRASTRUCTURE — numpy front-end of the plain-C CPU restatement (oracle/rasterizer_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this module; the
product package never does. ``forward`` / ``backward`` mirror the data flow of
Rasterizer::forward / Rasterizer::backward (reference rasterizer_impl.cu:198-434) and return every
intermediate buffer by name so tests can compare stage by stage.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_libs = {}


def build() -> None:
    subprocess.check_call(["make", "-s", "-C", str(_HERE), "oracle"])


def _lib(dtype):
    name = "liboracle_f32.so" if dtype == np.float32 else "liboracle_f64.so"
    if name not in _libs:
        path = _HERE / name
        if not path.exists():
            build()
        L = C.CDLL(str(path))
        L.oracle_preprocess.restype = C.c_longlong
        L.oracle_render.restype = None
        L.oracle_render_backward.restype = None
        L.oracle_preprocess_backward.restype = None
        assert L.oracle_sizeof_real() == np.dtype(dtype).itemsize
        _libs[name] = L
    return _libs[name]


def _a(x, dtype):
    if x is None:
        return None
    return np.ascontiguousarray(np.asarray(x), dtype=dtype)


def _p(x):
    return None if x is None else x.ctypes.data_as(C.c_void_p)


def forward(scene: dict, cam, bg, sh_degree=3, scale_modifier=1.0, dtype=np.float32):
    """scene: means3D, opacities and (shs | colors_precomp) and ((scales, rotations) | cov3D_precomp).
    cam: object with image_height, image_width, tanfovx, tanfovy, viewmatrix, projmatrix, campos."""
    L = _lib(dtype)
    real = C.c_float if dtype == np.float32 else C.c_double
    means = _a(scene["means3D"], dtype)
    P = means.shape[0]
    W, H = int(cam.image_width), int(cam.image_height)
    shs = _a(scene.get("shs"), dtype)
    colors = _a(scene.get("colors_precomp"), dtype)
    M = shs.shape[1] if shs is not None else 0
    opac = _a(scene["opacities"], dtype).reshape(-1)
    scales = _a(scene.get("scales"), dtype)
    rots = _a(scene.get("rotations"), dtype)
    cov_pre = _a(scene.get("cov3D_precomp"), dtype)
    view = _a(cam.viewmatrix, dtype).reshape(-1)
    proj = _a(cam.projmatrix, dtype).reshape(-1)
    campos = _a(cam.campos, dtype).reshape(-1)
    bg = _a(bg, dtype).reshape(-1)

    out = {
        "radii": np.zeros(P, np.int32), "depths": np.zeros(P, dtype), "means2D": np.zeros((P, 2), dtype),
        "cov3D": np.zeros((P, 6), dtype), "conic_opacity": np.zeros((P, 4), dtype),
        "rgb": np.zeros((P, 3), dtype), "clamped": np.zeros((P, 3), np.uint8),
        "tiles_touched": np.zeros(P, np.uint32),
    }
    R = L.oracle_preprocess(
        P, int(sh_degree), M, W, H, _p(means), _p(shs), _p(colors), _p(opac), _p(scales),
        real(scale_modifier), _p(rots), _p(cov_pre), _p(view), _p(proj), _p(campos),
        real(cam.tanfovx), real(cam.tanfovy), _p(out["radii"]), _p(out["depths"]), _p(out["means2D"]),
        _p(out["cov3D"]), _p(out["conic_opacity"]), _p(out["rgb"]), _p(out["clamped"]),
        _p(out["tiles_touched"])) if P else 0
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["num_rendered"] = int(R)
    out["point_list_keys"] = np.zeros(R, np.uint64)
    out["point_list"] = np.zeros(R, np.uint32)
    out["ranges"] = np.zeros((tiles, 2), np.uint32)
    out["final_T"] = np.zeros((H, W), dtype)
    out["n_contrib"] = np.zeros((H, W), np.uint32)
    out["color"] = np.zeros((3, H, W), dtype)
    if P:  # P == 0: the reference skips the core and the image stays zero
        L.oracle_render(P, W, H, C.c_longlong(R), _p(out["radii"]), _p(out["depths"]),
                        _p(out["means2D"]), _p(out["conic_opacity"]), _p(out["rgb"]),
                        _p(out["tiles_touched"]), _p(bg), _p(out["point_list_keys"]),
                        _p(out["point_list"]), _p(out["ranges"]), _p(out["final_T"]),
                        _p(out["n_contrib"]), _p(out["color"]))
    out["_ctx"] = dict(means=means, shs=shs, colors=colors, scales=scales, rots=rots, cov_pre=cov_pre,
                       view=view, proj=proj, campos=campos, bg=bg, P=P, M=M, W=W, H=H,
                       D=int(sh_degree), mod=float(scale_modifier), tanx=float(cam.tanfovx),
                       tany=float(cam.tanfovy), dtype=dtype)
    return out


def backward(fwd: dict, dL_dpix):
    """Gradients for the forward result `fwd`, in the reference's shapes."""
    c = fwd["_ctx"]
    dtype = c["dtype"]
    L = _lib(dtype)
    real = C.c_float if dtype == np.float32 else C.c_double
    P, M, W, H = c["P"], c["M"], c["W"], c["H"]
    dpix = _a(dL_dpix, dtype)
    g = {
        "dL_dmeans2D": np.zeros((P, 3), dtype), "dL_dconic": np.zeros((P, 4), dtype),
        "dL_dopacity": np.zeros((P, 1), dtype), "dL_dcolors": np.zeros((P, 3), dtype),
        "dL_dmeans3D": np.zeros((P, 3), dtype), "dL_dcov3D": np.zeros((P, 6), dtype),
        "dL_dsh": np.zeros((P, M, 3), dtype), "dL_dscales": np.zeros((P, 3), dtype),
        "dL_drotations": np.zeros((P, 4), dtype),
    }
    if P == 0:
        return g
    colors = c["colors"] if c["colors"] is not None else fwd["rgb"]
    L.oracle_render_backward(W, H, _p(fwd["ranges"]), _p(fwd["point_list"]), _p(c["bg"]),
                             _p(fwd["means2D"]), _p(fwd["conic_opacity"]), _p(colors),
                             _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dpix),
                             _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]),
                             _p(g["dL_dcolors"]))
    cov3D = c["cov_pre"] if c["cov_pre"] is not None else fwd["cov3D"]
    L.oracle_preprocess_backward(
        P, c["D"], M, W, H, _p(c["means"]), _p(fwd["radii"]), _p(c["shs"]), _p(fwd["clamped"]),
        _p(c["scales"]), _p(c["rots"]), real(c["mod"]), _p(cov3D), _p(c["view"]), _p(c["proj"]),
        _p(c["campos"]), real(c["tanx"]), real(c["tany"]), _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]),
        _p(g["dL_dcolors"]), _p(g["dL_dmeans3D"]), _p(g["dL_dcov3D"]), _p(g["dL_dsh"]),
        _p(g["dL_dscales"]), _p(g["dL_drotations"]))
    return g
