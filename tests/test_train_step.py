"""The step after the path: L1 + D-SSIM loss and the Adam update.

CPU part: the torch restatement below (the checker of the GPU part at full size) against golden
vectors made by importing the reference's own loss and torch.optim.Adam in the build container
(tests/golden/make_golden_loss.py), and the host logic of FusedAdam. GPU part (-m gpu): the CUDA
kernels (bgr_l1_ssim_loss, bgr_adam_step) against the golden vectors and against the restatement at
1080p."""
import math
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn.functional as F

GOLDEN = Path(__file__).resolve().parent / "golden" / "loss_golden.npz"
GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
LRS = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}


def torch_loss(img, gt, lam=0.2):
    """Restatement of train.py:91-92 over utils/loss_utils.py:16-63 -> (loss, l1, ssim)."""
    g = torch.tensor([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    window = g.mm(g.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().to(img)
    conv = lambda t: F.conv2d(t[None], window, padding=5, groups=3)[0]  # noqa: E731
    mu1, mu2 = conv(img), conv(gt)
    s1, s2, s12 = conv(img * img) - mu1 * mu1, conv(gt * gt) - mu2 * mu2, conv(img * gt) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))).mean()
    l1 = (img - gt).abs().mean()
    return (1.0 - lam) * l1 + lam * (1.0 - ssim), l1, ssim


def close(got, want):
    """Float32 agreement up to a few ulps of the operands: |got - want| <= 2e-6 |want| + 2e-7 max|want|
    (an element that is a small difference of large ones carries the rounding of the large ones)."""
    return bool(((got - want).abs() <= 2e-6 * want.abs() + 2e-7 * want.abs().max()).all())


def grad_close(got, want):
    """|got - want| <= 1e-4 |want| + 1e-5 max|want| elementwise (the gradient is ~1/(3HW) per pixel)."""
    tol = 1e-4 * want.abs() + 1e-5 * want.abs().max()
    return bool(((got - want).abs() <= tol).all())


@pytest.mark.parametrize("name", ["a", "b"])
def test_restatement_matches_the_reference_loss(name):
    z = np.load(GOLDEN)
    img = torch.tensor(z[f"{name}_image"], requires_grad=True)
    loss, l1, ssim = torch_loss(img, torch.tensor(z[f"{name}_gt"]))
    loss.backward()
    assert np.allclose([loss.item(), l1.item(), ssim.item()], z[f"{name}_terms"], rtol=1e-6, atol=1e-7)
    assert grad_close(img.grad, torch.tensor(z[f"{name}_grad"]))


def test_fused_adam_keeps_the_optimizer_surface_the_trainer_uses():
    """param_groups with names and per-group lr (gaussian_model.py:149-168), zero_grad, state keys."""
    import train_step
    params = {k: torch.nn.Parameter(torch.zeros(4, 3)) for k in GROUPS}
    opt = train_step.FusedAdam([{"params": [params[k]], "lr": LRS[k], "name": k} for k in GROUPS], lr=0.0, eps=1e-15)
    assert [g["name"] for g in opt.param_groups] == list(GROUPS)
    assert opt.param_groups[2]["lr"] == LRS["f_rest"] and opt.param_groups[0]["eps"] == 1e-15
    for g in opt.param_groups:                      # update_learning_rate's loop
        if g["name"] == "xyz":
            g["lr"] = 1e-5
    assert opt.param_groups[0]["lr"] == 1e-5
    params["xyz"].grad = torch.ones(4, 3)
    opt.zero_grad()
    assert params["xyz"].grad is None
    sd = opt.state_dict()
    assert [g["params"] for g in sd["param_groups"]] == [[i] for i in range(6)] and sd["state"] == {}
    params["xyz"].grad = torch.ones(4, 3)
    with pytest.raises(RuntimeError, match="CUDA"):  # no CPU path
        opt.step()
    import diff_gaussian_rasterization as d
    with pytest.raises(RuntimeError, match="CUDA"):
        d._C.l1_ssim_loss(torch.zeros(3, 8, 8), torch.zeros(3, 8, 8), 0.2)


# ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b"])
def test_cuda_loss_matches_the_reference_golden(name):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import train_step
    z = np.load(GOLDEN)
    img = torch.tensor(z[f"{name}_image"]).cuda().requires_grad_(True)
    gt = torch.tensor(z[f"{name}_gt"]).cuda()
    loss = train_step.l1_ssim_loss(img, gt, 0.2)
    loss.backward()
    terms = train_step.loss_terms(img, gt, 0.2).cpu().double().numpy()
    assert np.allclose(terms, z[f"{name}_terms"], rtol=2e-6, atol=1e-7)
    assert abs(loss.item() - z[f"{name}_terms"][0]) <= 2e-6 * abs(z[f"{name}_terms"][0]) + 1e-7
    assert grad_close(img.grad.cpu(), torch.tensor(z[f"{name}_grad"]))


@pytest.mark.gpu
@pytest.mark.parametrize("H,W", [(1080, 1920), (800, 800), (37, 53), (1, 1)])
def test_cuda_loss_matches_torch_at_size(H, W):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import train_step
    g = torch.Generator().manual_seed(H * 7 + W)
    gt = torch.rand(3, H, W, generator=g).cuda()
    img0 = (gt + 0.2 * torch.randn(3, H, W, generator=g).cuda()).clamp(0, 1)
    a = img0.clone().requires_grad_(True)
    b = img0.clone().requires_grad_(True)
    want, _, _ = torch_loss(a, gt, 0.2)
    want.backward()
    got = train_step.l1_ssim_loss(b, gt, 0.2)
    (3.0 * got).backward()                               # upstream factor passes through
    assert abs(got.item() - want.item()) <= 2e-6 * abs(want.item()) + 1e-7
    assert grad_close(b.grad / 3.0, a.grad)
    # deterministic: fixed-order sums
    again = train_step.loss_terms(img0, gt, 0.2)
    assert torch.equal(again, train_step.loss_terms(img0, gt, 0.2))


@pytest.mark.gpu
def test_cuda_adam_matches_torch_optim_adam_golden():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import train_step
    z = np.load(GOLDEN)
    params = {k: torch.nn.Parameter(torch.tensor(z[f"adam_p0_{k}"]).cuda()) for k in GROUPS}
    opt = train_step.FusedAdam([{"params": [params[k]], "lr": LRS[k], "name": k} for k in GROUPS], lr=0.0, eps=1e-15)
    for step in range(5):
        for k in GROUPS:
            params[k].grad = torch.tensor(z[f"adam_g{step}_{k}"]).cuda()
        if step == 3:
            opt.param_groups[0]["lr"] = 0.00008
        opt.step()
    for k in GROUPS:
        for got, key in ((params[k].detach(), "p5"), (opt.state[params[k]]["exp_avg"], "m5"),
                         (opt.state[params[k]]["exp_avg_sq"], "v5")):
            want = torch.tensor(z[f"adam_{key}_{k}"])
            assert close(got.cpu(), want), (k, key)
        assert opt.state[params[k]]["step"] == 5


@pytest.mark.gpu
def test_cuda_adam_large_groups_and_missing_gradients():
    """3M-row sized groups in one launch against torch.optim.Adam on the GPU; a group without a
    gradient is skipped as torch does."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import train_step
    P = 300_001
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3),
              "rotation": (P, 4)}
    g = torch.Generator(device="cuda").manual_seed(5)
    init = {k: torch.randn(*s, generator=g, device="cuda") for k, s in shapes.items()}
    pa = {k: torch.nn.Parameter(v.clone()) for k, v in init.items()}
    pb = {k: torch.nn.Parameter(v.clone()) for k, v in init.items()}
    # one group at a 4-byte offset: the kernel's 128-bit path must step aside for it
    shifted = torch.empty(init["scaling"].numel() + 1, device="cuda")
    shifted[1:] = init["scaling"].reshape(-1)
    pb["scaling"] = torch.nn.Parameter(shifted[1:].view(P, 3))
    assert pb["scaling"].data_ptr() % 16 != 0 and pb["scaling"].is_contiguous()
    oa = torch.optim.Adam([{"params": [pa[k]], "lr": LRS[k], "name": k} for k in GROUPS], lr=0.0, eps=1e-15)
    ob = train_step.FusedAdam([{"params": [pb[k]], "lr": LRS[k], "name": k} for k in GROUPS], lr=0.0, eps=1e-15)
    for step in range(3):
        for k in GROUPS:
            if k == "opacity" and step == 1:
                pa[k].grad = pb[k].grad = None
                continue
            gr = torch.randn(*shapes[k], generator=g, device="cuda") * 0.01
            pa[k].grad, pb[k].grad = gr, gr.clone()
        oa.step()
        ob.step()
    for k in GROUPS:
        assert close(pb[k].detach(), pa[k].detach()), k
    assert ob.state[pb["opacity"]]["step"] == 2 and ob.state[pb["xyz"]]["step"] == 3


@pytest.mark.gpu
def test_cuda_densification_stats_match_the_trainer_bookkeeping():
    """train.py:114-116 + scene/gaussian_model.py:405-407, restated with torch's masked ops."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import train_step
    P = 100_003
    g = torch.Generator(device="cuda").manual_seed(11)
    radii = (torch.randint(-2, 40, (P,), generator=g, device="cuda").clamp_min(0)).int()
    grad = torch.randn(P, 3, generator=g, device="cuda") * 1e-3
    mr = torch.rand(P, generator=g, device="cuda") * 30
    acc = torch.rand(P, 1, generator=g, device="cuda")
    den = torch.randint(0, 5, (P, 1), generator=g, device="cuda").float()
    w_mr, w_acc, w_den = mr.clone(), acc.clone(), den.clone()
    vis = radii > 0
    w_mr[vis] = torch.max(w_mr[vis], radii[vis])
    w_acc[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
    w_den[vis] += 1
    train_step.densification_stats(radii, grad, mr, acc, den)
    assert torch.equal(mr, w_mr) and torch.equal(den, w_den)
    assert torch.allclose(acc, w_acc, rtol=1e-6, atol=1e-9)
    assert 0 < int(vis.sum()) < P
    with pytest.raises(RuntimeError, match="densification_stats needs"):
        train_step.densification_stats(radii.long(), grad, mr, acc, den)
