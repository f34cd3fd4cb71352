"""Helpers shared by the tests: scene plumbing between torch / numpy / the three implementations."""
from __future__ import annotations

import numpy as np
import torch

import synthetic


def scene_np(scene):
    return {k: v.detach().cpu().numpy() for k, v in scene.items()}


def settings_for(cam, bg, sh_degree=3, scale_modifier=1.0, device="cuda", debug=False):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx,
        tanfovy=cam.tanfovy, bg=bg.to(device), scale_modifier=scale_modifier,
        viewmatrix=cam.viewmatrix.to(device), projmatrix=cam.projmatrix.to(device),
        sh_degree=sh_degree, campos=cam.campos.to(device), prefiltered=False, debug=debug)


def native_forward(backend, scene, cam, bg, sh_degree=3, scale_modifier=1.0, colors_precomp=None,
                   cov3D_precomp=None, prefiltered=False, debug=False):
    """Calls backend.rasterize_gaussians with the reference's 19 positional arguments."""
    e = torch.Tensor([])
    dev = "cuda"
    sh = e if colors_precomp is not None else scene["shs"].to(dev)
    col = colors_precomp.to(dev) if colors_precomp is not None else e
    sc = e if cov3D_precomp is not None else scene["scales"].to(dev)
    rot = e if cov3D_precomp is not None else scene["rotations"].to(dev)
    cov = cov3D_precomp.to(dev) if cov3D_precomp is not None else e
    args = (bg.to(dev), scene["means3D"].to(dev), col, scene["opacities"].to(dev), sc, rot,
            scale_modifier, cov, cam.viewmatrix.to(dev), cam.projmatrix.to(dev), cam.tanfovx,
            cam.tanfovy, cam.image_height, cam.image_width, sh, sh_degree, cam.campos.to(dev),
            bool(prefiltered), bool(debug))
    return args, backend.rasterize_gaussians(*args)


def native_backward(backend, fwd_args, fwd_out, dL_dpix):
    (bg, means3D, col, _op, sc, rot, mod, cov, view, proj, tx, ty, _h, _w, sh, deg, campos, _pf,
     dbg) = fwd_args
    R, _color, radii, geom, binning, img = fwd_out
    return backend.rasterize_gaussians_backward(bg, means3D, radii, col, sc, rot, mod, cov, view,
                                                proj, tx, ty, dL_dpix.to("cuda"), sh, deg, campos,
                                                geom, R, binning, img, dbg)


def bgr_intermediates(fwd_out, P, W, H):
    """Decodes this repo's scratch buffers (layout: include/bgr.h) into named arrays."""
    from diff_gaussian_rasterization import _C
    R, _c, _r, geom, binning, img = fwd_out
    gl, il, bl = _C.geom_layout(P), _C.image_layout(W, H), _C.binning_layout(R)
    rec = geom[gl["rec"]:gl["rec"] + P * 48].view(torch.float32).view(P, 12)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out = {
        "means2D": rec[:, 0:2], "conic_opacity": torch.stack([rec[:, 2], rec[:, 3], rec[:, 4], rec[:, 5]], 1),
        "rgb": rec[:, 6:9], "thr": rec[:, 9], "depths": rec[:, 10],
        "clampbits": rec[:, 11].contiguous().view(torch.int32),
        "tiles_touched": geom[gl["tiles_touched"]:gl["tiles_touched"] + 4 * P].view(torch.int32),
        "final_T": img[il["final_T"]:il["final_T"] + 4 * W * H].view(torch.float32).view(H, W),
        "n_contrib": img[il["n_contrib"]:il["n_contrib"] + 4 * W * H].view(torch.int32).view(H, W),
        "ranges": img[il["ranges"]:il["ranges"] + 8 * tiles].view(torch.int32).view(tiles, 2),
        "point_list": binning[bl["point_list"]:bl["point_list"] + 4 * R].view(torch.int32) if R else torch.zeros(0, dtype=torch.int32),
    }
    return out


def rel_abs_ok(a, b, rel, abs_):
    """|a-b| <= abs_ + rel*|b| elementwise -> fraction violating, max violation ratio."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    err = np.abs(a - b); tol = abs_ + rel * np.abs(b)
    bad = err > tol
    return bad.mean(), (err / tol).max() if err.size else 0.0
