"""The step around the rasterizer: the trainer's loss and optimizer update,
each as a handful of libbgr launches instead of PyTorch's elementwise chains.

    loss = l1_ssim_loss(image, gt, lambda_dssim)       train.py:91-92, utils/loss_utils.py:16-63
    opt  = FusedAdam(param_groups, lr=0.0, eps=1e-15)  scene/gaussian_model.py:149-160

`FusedAdam` keeps `torch.optim.Adam`'s surface as the reference uses it: `param_groups` is the same
list of dicts (`params`, `lr`, `name`), so `update_learning_rate` (scene/gaussian_model.py:162-168)
works unchanged; `state[param]` holds `step`, `exp_avg`, `exp_avg_sq` under the same keys, which is
what the densification code reads and rewrites (`replace_tensor_to_optimizer`, `_prune_optimizer`,
`cat_tensors_to_optimizer`, :256-330). No CPU path: CUDA float32 parameters only.
"""
from __future__ import annotations

from collections import defaultdict

import torch

from diff_gaussian_rasterization import _C


class _L1SSIMLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        loss3, grad = _C.l1_ssim_loss(image, gt, lambda_dssim, need_grad=image.requires_grad)
        ctx.save_for_backward(grad)
        ctx.has_grad = grad is not None
        ctx.terms = loss3           # {loss, l1, ssim}, readable without another pass
        return loss3[0]

    @staticmethod
    def backward(ctx, grad_out):
        if not ctx.has_grad:
            return None, None, None
        (grad,) = ctx.saved_tensors
        return grad * grad_out, None, None


def l1_ssim_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> torch.Tensor:
    """(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)) of the reference's trainer,
    differentiable w.r.t. `image` (3,H,W)."""
    return _L1SSIMLoss.apply(image, gt, lambda_dssim)


def loss_terms(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> torch.Tensor:
    """Device tensor {loss, l1, ssim} without gradient (what train.py logs)."""
    return _C.l1_ssim_loss(image.detach(), gt, lambda_dssim, need_grad=False)[0]


@torch.no_grad()
def densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom) -> None:
    """The per-iteration bookkeeping of train.py:114-116, in place:
    `max_radii2D[vis] = max(max_radii2D[vis], radii[vis])` and `add_densification_stats`
    (scene/gaussian_model.py:405-407) with `vis = radii > 0`; `viewspace_grad` is the (P,3)
    gradient the rasterizer leaves on `means2D`."""
    _C.densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom)


class FusedAdam:
    """torch.optim.Adam(param_groups, lr, betas, eps) with one tensor per group, one launch per step."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        groups = list(params)
        if groups and not isinstance(groups[0], dict):
            groups = [{"params": groups}]
        self.defaults = {"lr": lr, "betas": tuple(betas), "eps": eps}
        self.param_groups = []
        for g in groups:
            g = dict(g)
            g["params"] = list(g["params"])
            g.setdefault("lr", lr)
            g.setdefault("betas", tuple(betas))
            g.setdefault("eps", eps)
            self.param_groups.append(g)
        self.state = defaultdict(dict)

    def zero_grad(self, set_to_none: bool = True):
        for g in self.param_groups:
            for p in g["params"]:
                if set_to_none:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.zero_()

    def _launches(self):
        """Parameters that have a gradient, chunked into launches of <= 8 with equal hyper-parameters
        and step counts."""
        work = []
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                step = int(st["step"]) + 1
                st["step"] = step
                work.append(((tuple(g["betas"]), float(g["eps"]), step), p, st, float(g["lr"])))
        launches = []
        for key, p, st, lr in work:
            if launches and launches[-1][0] == key and len(launches[-1][1]) < 8:
                launches[-1][1].append((p, st, lr))
            else:
                launches.append((key, [(p, st, lr)]))
        return launches

    @torch.no_grad()
    def step(self):
        for (betas, eps, step), items in self._launches():
            params = [p for p, _, _ in items]
            grads = [p.grad if p.grad.is_contiguous() else p.grad.contiguous() for p in params]
            _C.adam_step(params, grads, [st["exp_avg"] for _, st, _ in items],
                         [st["exp_avg_sq"] for _, st, _ in items], [lr for _, _, lr in items],
                         betas[0], betas[1], eps, step)

    def state_dict(self):
        index = {}
        packed = []
        for g in self.param_groups:
            ids = []
            for p in g["params"]:
                index.setdefault(id(p), len(index))
                ids.append(index[id(p)])
            packed.append({**{k: v for k, v in g.items() if k != "params"}, "params": ids})
        state = {}
        for g in self.param_groups:
            for p in g["params"]:
                if p in self.state and self.state[p]:
                    state[index[id(p)]] = dict(self.state[p])
        return {"state": state, "param_groups": packed}

    def load_state_dict(self, sd):
        flat = [p for g in self.param_groups for p in g["params"]]
        for g, saved in zip(self.param_groups, sd["param_groups"]):
            for k, v in saved.items():
                if k != "params":
                    g[k] = v
        self.state = defaultdict(dict)
        for i, st in sd["state"].items():
            self.state[flat[int(i)]] = dict(st)
