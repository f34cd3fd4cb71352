"""GPU suite, part 5: the extended entry with the trainer's activations fused in (SURVEY.md §8 row
f1; bgr_forward_raw / bgr_backward_raw, RawGaussianRasterizer). Oracle: torch autograd over the
activations of scene/gaussian_model.py:95-115 (exp, normalize, sigmoid, cat) in front of the drop-in
path - the composition the reference's trainer runs (gaussian_renderer/__init__.py:53-82)."""
import pytest
import torch
import torch.nn.functional as F

import synthetic
import util
from test_gpu_full import make_case, violations
from test_gpu_parity import C_  # noqa: F401  (fixture)

pytestmark = pytest.mark.gpu

RAW_NAMES = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")


def raw_leaves(sc, M=16, seed=0):
    """The activated synthetic scene pulled back to the trainer's parameterisation
    (scene/gaussian_model.py:47-52); quaternions get non-unit lengths so that normalize matters."""
    g = torch.Generator().manual_seed(seed)
    P = sc["means3D"].shape[0]
    stretch = (0.5 + 1.5 * torch.rand(P, 1, generator=g)).cuda()
    o = sc["opacities"].clamp(1e-4, 1 - 1e-4)
    return {
        "xyz": sc["means3D"].clone(),
        "features_dc": sc["shs"][:, :1, :].contiguous(),
        "features_rest": sc["shs"][:, 1:M, :].contiguous(),
        "opacity": torch.log(o / (1 - o)),
        "scaling": torch.log(sc["scales"]),
        "rotation": sc["rotations"] * stretch,
    }


def run_composed(raw, cam, bg, deg, dpix):
    """torch autograd over the activations + the drop-in rasterizer."""
    import diff_gaussian_rasterization as d
    L = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    m2d = torch.zeros_like(L["xyz"], requires_grad=True)
    shs = torch.cat((L["features_dc"], L["features_rest"]), dim=1)
    img, radii = d.GaussianRasterizer(util.settings_for(cam, bg, sh_degree=deg))(
        means3D=L["xyz"], means2D=m2d, opacities=torch.sigmoid(L["opacity"]), shs=shs,
        scales=torch.exp(L["scaling"]), rotations=F.normalize(L["rotation"]))
    img.backward(dpix)
    return img.detach(), radii, {k: L[k].grad for k in RAW_NAMES}, m2d.grad


def run_fused(raw, cam, bg, deg, dpix):
    import diff_gaussian_rasterization as d
    L = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    m2d = torch.zeros_like(L["xyz"], requires_grad=True)
    img, radii = d.RawGaussianRasterizer(util.settings_for(cam, bg, sh_degree=deg))(
        L["xyz"], m2d, L["features_dc"], L["features_rest"], L["opacity"], L["scaling"], L["rotation"])
    img.backward(dpix)
    return img.detach(), radii, {k: L[k].grad for k in RAW_NAMES}, m2d.grad


def compare(P, res, M, deg, view_seed=0, seed=0, bg=(0.0, 0.0, 0.0)):
    sc, cam = make_case(P, res, view_seed=view_seed, seed=seed)
    raw = raw_leaves(sc, M, seed)
    bg = torch.tensor(bg)
    dpix = synthetic.make_upstream_grad(cam.image_height, cam.image_width).cuda()
    img_c, radii_c, g_c, m_c = run_composed(raw, cam, bg, deg, dpix)
    img_c2, _, g_c2, m_c2 = run_composed(raw, cam, bg, deg, dpix)        # run-to-run spread of the atomics
    img_f, radii_f, g_f, m_f = run_fused(raw, cam, bg, deg, dpix)
    # the activations are the same formulas; the quaternion norm may round differently from torch's
    # reduction kernel, which can move a radius (ceil of 3 sigma) on a handful of Gaussians
    differ = (radii_c != radii_f).float().mean().item()
    assert differ <= 1e-5, f"radii differ on {differ:.2e} of the Gaussians"
    assert (img_c - img_f).abs().max().item() <= (1e-5 if differ == 0 else 1e-3)
    assert 0 < int((radii_f > 0).sum()) < P
    for k in RAW_NAMES:
        assert g_f[k] is not None and g_f[k].shape == raw[k].shape, k
        assert torch.isfinite(g_f[k]).all(), k
        if g_c[k].numel() == 0:
            continue
        spread, _ = violations(g_c2[k], g_c[k])
        mine, _ = violations(g_f[k], g_c[k])
        assert mine <= max(3 * spread, 1e-4), f"{k}: {mine:.2e} outside the 1e-4/1e-6 bar (composed run-to-run {spread:.2e})"
        scale = g_c[k].abs().max().clamp_min(1e-30)
        ref_norm = ((g_c2[k] - g_c[k]).abs().max() / scale).item()
        assert ((g_f[k] - g_c[k]).abs().max() / scale).item() <= max(1e-3, 3 * ref_norm), k
    assert violations(m_f, m_c)[0] <= max(3 * violations(m_c2, m_c)[0], 1e-4)
    return raw


def test_raw_entry_matches_autograd_over_the_activations_degree3():
    compare(200_000, "800x800", M=16, deg=3, view_seed=1)


def test_raw_entry_1080p_with_background_and_ragged_size():
    compare(300_017, "1080p", M=16, deg=3, view_seed=3, seed=2, bg=(0.2, 0.5, 0.1))


def test_raw_entry_lower_active_degree_than_stored():
    # the trainer starts at active_sh_degree 0 with all 16 coefficients stored (train.py:69-70)
    compare(100_000, "800x800", M=16, deg=0, view_seed=2)
    compare(100_000, "800x800", M=16, deg=1, view_seed=2)


@pytest.mark.parametrize("M,deg", [(9, 2), (4, 1), (1, 0)])
def test_raw_entry_other_coefficient_counts(M, deg):
    compare(50_003, "800x800", M=M, deg=deg, view_seed=4, seed=5)


def test_raw_entry_rejects_what_it_cannot_take(C_):
    sc, cam = make_case(1000, "800x800")
    raw = raw_leaves(sc)
    bg = torch.zeros(3).cuda()
    common = (1.0, cam.viewmatrix.cuda(), cam.projmatrix.cuda(), cam.tanfovx, cam.tanfovy,
              cam.image_height, cam.image_width)
    with pytest.raises(RuntimeError, match="features_dc must be"):
        C_.rasterize_gaussians_raw(bg, raw["xyz"], raw["opacity"], raw["scaling"], raw["rotation"], *common,
                                   sc["shs"], raw["features_rest"], 3, cam.campos.cuda(), False, False)
    with pytest.raises(RuntimeError, match="CUDA-only"):
        C_.rasterize_gaussians_raw(bg, raw["xyz"].cpu(), raw["opacity"], raw["scaling"], raw["rotation"], *common,
                                   raw["features_dc"], raw["features_rest"], 3, cam.campos.cuda(), False, False)
    with pytest.raises(RuntimeError, match="SH degree does not fit"):
        C_.rasterize_gaussians_raw(bg, raw["xyz"], raw["opacity"], raw["scaling"], raw["rotation"], *common,
                                   raw["features_dc"], raw["features_rest"][:, :3].contiguous(), 3,
                                   cam.campos.cuda(), False, False)
