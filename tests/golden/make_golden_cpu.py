"""Generates tests/golden/sh_camera_golden.npz by IMPORTING the reference's own pure-torch helpers
from /root/reference (possible only in the build container, on the CPU):

  utils/sh_utils.py:57-112         eval_sh          -> SH colour of random directions, degrees 0..3
  utils/graphics_utils.py:38-71    getWorld2View2, getProjectionMatrix -> camera matrices

The committed .npz pins oracle/rasterizer_oracle.c (sh_rgb) and tests/synthetic.py
(rigid_world_to_camera, perspective, make_camera) against the reference.

    python tests/golden/make_golden_cpu.py
"""
import math
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.sh_utils import eval_sh  # noqa: E402
from utils.graphics_utils import getWorld2View2, getProjectionMatrix  # noqa: E402

out = {}
g = torch.Generator().manual_seed(123)
P = 257
means = torch.randn(P, 3, generator=g) * 2
campos = torch.tensor([0.3, -0.4, 1.7])
shs = torch.randn(P, 16, 3, generator=g) * 0.4
shs[:, 0] += 0.5
out["means3D"] = means.numpy()
out["campos"] = campos.numpy()
out["shs"] = shs.numpy()
dirs = means - campos
dirs = dirs / dirs.norm(dim=1, keepdim=True)
for deg in range(4):
    # gaussian_renderer/__init__.py:74-78: colour = clamp_min(eval_sh(deg, shs^T, dir) + 0.5, 0)
    sh_view = shs.transpose(1, 2)  # (P,3,16)
    rgb = eval_sh(deg, sh_view, dirs)
    out[f"rgb_raw_deg{deg}"] = (rgb + 0.5).numpy()
    out[f"rgb_deg{deg}"] = torch.clamp_min(rgb + 0.5, 0.0).numpy()

# cameras: the fixed §8(d) camera at the three resolutions (scene/cameras.py:48-57)
def rot_x(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)
def rot_y(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)
R = rot_y(0.30) @ rot_x(-0.20)
T = np.array([0.1, -0.2, 4.0])
for name, (W, H, f) in {"800x800": (800, 800, 1111.0), "1080p": (1920, 1080, 1600.0), "4k": (3840, 2160, 3200.0)}.items():
    fovx = 2 * math.atan(W / (2 * f)); fovy = 2 * math.atan(H / (2 * f))
    wvt = torch.tensor(getWorld2View2(R, T)).transpose(0, 1)
    proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).transpose(0, 1)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    center = wvt.inverse()[3, :3]
    out[f"view_{name}"] = wvt.numpy(); out[f"proj_{name}"] = full.numpy(); out[f"campos_{name}"] = center.numpy()
    out[f"tan_{name}"] = np.array([math.tan(fovx * 0.5), math.tan(fovy * 0.5)])
np.savez_compressed(Path(__file__).parent / "sh_camera_golden.npz", **out)
print("wrote", len(out), "arrays")
