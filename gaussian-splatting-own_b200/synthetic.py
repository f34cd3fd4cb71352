"""Seeded synthetic Gaussians + cameras for parity tests and the bench.

Everything is generated on the CPU with a seeded ``torch.Generator`` so that the scene is
bit-identical on every machine; callers move it to the GPU. The camera helpers restate the
reference's conventions (reference file:line):

* ``world_to_view``       -- utils/graphics_utils.py:38-49 (getWorld2View2, translate=0, scale=1)
* ``projection_matrix``   -- utils/graphics_utils.py:51-71 (getProjectionMatrix)
* ``make_camera``         -- scene/cameras.py:48-57 (znear/zfar, transposes, full projection,
                             camera centre) and gaussian_renderer/__init__.py:33-34 (tan fov)

They are pinned against the reference's own functions by tests/golden/camera_*.pt
(generated with tests/golden/make_golden_cpu.py, which imports /root/reference).
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple

import numpy as np
import torch


class Camera(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # (4,4) f32, world_view_transform (row-vector convention)
    projmatrix: torch.Tensor   # (4,4) f32, full_proj_transform
    campos: torch.Tensor       # (3,) f32


def world_to_view(R: np.ndarray, t: np.ndarray) -> np.ndarray:
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    Rt = np.linalg.inv(C2W)
    return np.float32(Rt)


def projection_matrix(znear: float, zfar: float, fovX: float, fovY: float) -> torch.Tensor:
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def _rot_x(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)


def _rot_y(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def make_camera(width: int, height: int, focal: float, view_seed: int = 0) -> Camera:
    """Fixed camera of §8(d); ``view_seed`` > 0 perturbs R, T (the 8 cameras of config c3)."""
    ry, rx = 0.30, -0.20
    T = np.array([0.1, -0.2, 4.0], dtype=np.float64)
    if view_seed:
        rng = np.random.RandomState(1000 + view_seed)
        ry += rng.uniform(-0.25, 0.25)
        rx += rng.uniform(-0.15, 0.15)
        T = T + rng.uniform(-0.3, 0.3, size=3)
    R = _rot_y(ry) @ _rot_x(rx)
    fovx = 2.0 * math.atan(width / (2.0 * focal))
    fovy = 2.0 * math.atan(height / (2.0 * focal))
    view = torch.tensor(world_to_view(R, T)).transpose(0, 1).contiguous()
    proj = projection_matrix(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = view.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return Camera(height, width, math.tan(fovx * 0.5), math.tan(fovy * 0.5), view, full, campos)


# focal length (px) per named resolution, §8(d)
RESOLUTIONS = {
    "800x800": (800, 800, 1111.0),
    "1080p": (1920, 1080, 1600.0),
    "4k": (3840, 2160, 3200.0),
}


def make_gaussians(P: int, seed: int = 0, sh_degree: int = 3,
                   log_scale_mean: float = math.log(0.005)) -> Dict[str, torch.Tensor]:
    """Post-activation Gaussian parameters exactly as the rasterizer receives them."""
    g = torch.Generator().manual_seed(seed)
    M = (sh_degree + 1) ** 2 if sh_degree >= 0 else 0
    u = torch.rand(P, 3, generator=g)
    lo = torch.tensor([-3.5, -3.5, -5.0])
    hi = torch.tensor([3.5, 3.5, 6.0])
    means3D = lo + u * (hi - lo)
    scales = torch.exp(log_scale_mean + 0.7 * torch.randn(P, 3, generator=g))
    q = torch.randn(P, 4, generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(2.0 * torch.randn(P, 1, generator=g))
    shs = torch.empty(P, 16, 3)
    shs[:, 0, :] = 0.8 * torch.randn(P, 3, generator=g)
    shs[:, 1:, :] = 0.15 * torch.randn(P, 15, 3, generator=g)
    shs = shs[:, :max(M, 1), :].contiguous() if M < 16 else shs
    return {
        "means3D": means3D.contiguous(),
        "scales": scales.contiguous(),
        "rotations": rotations.contiguous(),
        "opacities": opacities.contiguous(),
        "shs": shs.contiguous(),
    }


def make_upstream_grad(height: int, width: int, seed: int = 1) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return (1e-3 * torch.randn(3, height, width, generator=g)).contiguous()


def make_tiny_scene(P: int, width: int, height: int, seed: int = 0, focal: float | None = None,
                    scale_mean: float = 0.05, box: float = 1.5):
    """Small dense scene for oracle / finite-difference tests: larger splats, everything in view."""
    g = torch.Generator().manual_seed(seed)
    focal = focal if focal is not None else 0.9 * width
    cam = make_camera(width, height, focal)
    means3D = (torch.rand(P, 3, generator=g) * 2 - 1) * torch.tensor([box, box, 1.5])
    scales = torch.exp(math.log(scale_mean) + 0.5 * torch.randn(P, 3, generator=g))
    q = torch.randn(P, 4, generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(1.5 * torch.randn(P, 1, generator=g))
    shs = torch.empty(P, 16, 3)
    shs[:, 0, :] = 0.8 * torch.randn(P, 3, generator=g)
    shs[:, 1:, :] = 0.15 * torch.randn(P, 15, 3, generator=g)
    scene = {
        "means3D": means3D.contiguous(), "scales": scales.contiguous(),
        "rotations": rotations.contiguous(), "opacities": opacities.contiguous(),
        "shs": shs.contiguous(),
    }
    return scene, cam
