"""Generates tests/golden/loss_golden.npz by IMPORTING the reference's own loss from
/root/reference (possible only in the build container, on the CPU):

  utils/loss_utils.py:16-17   l1_loss
  utils/loss_utils.py:36-63   ssim (create_window, _ssim)
  train.py:91-92              loss = (1 - lambda) * Ll1 + lambda * (1 - ssim)

and, for the optimizer, five steps of torch.optim.Adam(l, lr=0.0, eps=1e-15) over six small groups
with the reference's learning rates (scene/gaussian_model.py:149-160, arguments/__init__.py:74-81).
The committed .npz pins the torch restatement in tests/test_train_step.py (CPU) and the CUDA
kernels (bgr_l1_ssim_loss, bgr_adam_step; -m gpu).

    python tests/golden/make_golden_loss.py
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.loss_utils import l1_loss, ssim  # noqa: E402

out = {}
g = torch.Generator().manual_seed(7)
for name, (H, W) in {"a": (40, 56), "b": (33, 71)}.items():
    gt = torch.rand(3, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    img[:, :4, :5] = gt[:, :4, :5]                      # exact ties: sign(0) = 0 in the L1 gradient
    img.requires_grad_(True)
    lam = 0.2
    Ll1 = l1_loss(img, gt)
    s = ssim(img, gt)
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - s)
    loss.backward()
    out[f"{name}_image"] = img.detach().numpy()
    out[f"{name}_gt"] = gt.numpy()
    out[f"{name}_terms"] = np.array([loss.item(), Ll1.item(), s.item()], dtype=np.float64)
    out[f"{name}_grad"] = img.grad.numpy()

# Adam: the reference's six groups (sizes shrunk), its learning rates, eps 1e-15
shapes = {"xyz": (50, 3), "f_dc": (50, 1, 3), "f_rest": (50, 15, 3), "opacity": (50, 1), "scaling": (50, 3),
          "rotation": (50, 4)}
lrs = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
params = {k: torch.nn.Parameter(torch.randn(*s, generator=g)) for k, s in shapes.items()}
opt = torch.optim.Adam([{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
for k in shapes:
    out[f"adam_p0_{k}"] = params[k].detach().numpy().copy()
for step in range(5):
    for k in shapes:
        grad = torch.randn(*shapes[k], generator=g) * (10.0 ** -(step % 3))
        if k == "f_rest":
            grad[::2] = 0.0                               # culled Gaussians: zero rows
        params[k].grad = grad
        out[f"adam_g{step}_{k}"] = grad.numpy().copy()
    if step == 3:                                         # update_learning_rate (gaussian_model.py:162-168)
        opt.param_groups[0]["lr"] = 0.00008
    opt.step()
for k in shapes:
    out[f"adam_p5_{k}"] = params[k].detach().numpy().copy()
    out[f"adam_m5_{k}"] = opt.state[params[k]]["exp_avg"].numpy().copy()
    out[f"adam_v5_{k}"] = opt.state[params[k]]["exp_avg_sq"].numpy().copy()

dst = Path(__file__).resolve().parent / "loss_golden.npz"
np.savez_compressed(dst, **out)
print(dst, {k: v.shape for k, v in out.items() if not k.startswith("adam_g")})
