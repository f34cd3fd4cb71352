"""CPU suite: pins the oracle (oracle/rasterizer_oracle.c) before anything trusts it.

1. against outputs of the REFERENCE's own CUDA code (tests/golden/ref_small_*.npz, produced on a
   B200 by tests/golden/make_golden_gpu.py from oracle/_ref = the unmodified reference .cu files);
2. against the reference's pure-torch SH evaluation and camera helpers
   (tests/golden/sh_camera_golden.npz, produced by tests/golden/make_golden_cpu.py which imports
   /root/reference/utils);
3. its analytic backward against central finite differences of its own float64 forward.

Tolerances (BASELINE.json north_star): integer buffers bit-exact; image 1e-5 max-abs; gradients
1e-4 relative / 1e-6 absolute. The gradient bound is applied per element with a small allowance:
the reference accumulates with unordered float atomics and several backward terms cancel
catastrophically (denom - a*c), so two runs of the reference itself violate the element-wise bound
on ~1e-4 of the elements (measured on B200, DESIGN.md "Parity"). The golden upstream gradient is
O(1) (1e3 x the bench's), which scales the absolute term accordingly.
"""
from pathlib import Path

import numpy as np
import pytest

import synthetic
from oracle import cpu_oracle

GOLDEN = Path(__file__).parent / "golden"
GRAD_NAMES = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh",
              "dL_dscales", "dL_drotations")


def load_case(name):
    g = np.load(GOLDEN / f"ref_small_{name}.npz")
    P, W, H = int(g["P"]), int(g["W"]), int(g["H"])
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=int(g["seed"]))
    return g, {k: v.numpy() for k, v in scene.items()}, cam


def grad_close(mine, ref, rel=1e-4, abs_=1e-6, scale=1.0, max_bad_frac=5e-3, max_norm_err=2e-3):
    """element-wise |a-b| <= abs_*scale + rel*|b| except for a small fraction; and the worst error
    stays small against the tensor's magnitude."""
    mine = np.asarray(mine, np.float64).reshape(-1)
    ref = np.asarray(ref, np.float64).reshape(-1)
    if ref.size == 0:
        return True, "empty"
    err = np.abs(mine - ref)
    bad = err > abs_ * scale + rel * np.abs(ref)
    norm_err = err.max() / max(np.abs(ref).max(), 1e-30)
    ok = bad.mean() <= max_bad_frac and norm_err <= max_norm_err and np.isfinite(mine).all()
    return ok, f"bad {bad.sum()}/{bad.size} norm_err {norm_err:.2e}"


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_oracle_forward_matches_reference_cuda(name):
    g, scene, cam = load_case(name)
    o = cpu_oracle.forward(scene, cam, g["bg"], int(g["sh_degree"]), float(g["scale_modifier"]))
    vis = g["radii"] > 0
    assert o["num_rendered"] == int(g["num_rendered"])
    np.testing.assert_array_equal(o["radii"], g["radii"])
    np.testing.assert_array_equal(o["tiles_touched"], g["tiles_touched"].view(np.uint32))
    np.testing.assert_array_equal(o["point_list"], g["point_list"].view(np.uint32))
    np.testing.assert_array_equal(o["point_list_keys"], g["point_list_keys"].view(np.uint64))
    np.testing.assert_array_equal(o["ranges"], g["ranges"].view(np.uint32))
    np.testing.assert_array_equal(o["n_contrib"], g["n_contrib"].view(np.uint32))
    np.testing.assert_array_equal(o["clamped"][vis].astype(bool), g["clamped"][vis].astype(bool))
    np.testing.assert_array_equal(o["depths"][vis], g["depths"][vis])
    np.testing.assert_allclose(o["means2D"][vis], g["means2D"][vis], rtol=0, atol=1e-4)
    np.testing.assert_allclose(o["conic_opacity"][vis], g["conic_opacity"][vis], rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(o["rgb"][vis], g["rgb"][vis], rtol=0, atol=2e-6)
    np.testing.assert_allclose(o["cov3D"][vis], g["cov3D"][vis], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(o["final_T"], g["final_T"], rtol=0, atol=1e-5)
    assert np.abs(o["color"] - g["color"]).max() <= 1e-5


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_oracle_backward_matches_reference_cuda(name):
    g, scene, cam = load_case(name)
    o = cpu_oracle.forward(scene, cam, g["bg"], int(g["sh_degree"]), float(g["scale_modifier"]))
    gr = cpu_oracle.backward(o, g["dL_dpix"])
    scale = 1e3  # golden dL_dpix is 1e3 x the bench's 1e-3-sigma upstream gradient
    for k in GRAD_NAMES + ("dL_dconic",):
        ref = g[k].reshape(gr[k].shape)
        ok, msg = grad_close(gr[k], ref, scale=scale)
        assert ok, f"{k}: {msg}"


def test_oracle_sh_matches_reference_eval_sh():
    """utils/sh_utils.py:57-112 + gaussian_renderer/__init__.py:74-78 (imported when the fixture
    was generated) vs the oracle's SH stage, degrees 0..3."""
    g = np.load(GOLDEN / "sh_camera_golden.npz")
    means, campos, shs = g["means3D"], g["campos"], g["shs"]
    P = means.shape[0]
    cam = synthetic.make_camera(64, 64, 60.0)
    # a camera looking at everything: only the colour stage is compared, so place the points in
    # front of it by using a view matrix that is the identity with a z offset
    import torch
    view = torch.eye(4)
    view[3, 2] = 50.0
    cam = cam._replace(viewmatrix=view, projmatrix=view @ synthetic.perspective(
        0.01, 100.0, 1.0, 1.0).T, campos=torch.tensor(campos))
    scene = dict(means3D=means, shs=shs, opacities=np.full((P, 1), 0.5, np.float32),
                 scales=np.full((P, 3), 0.5, np.float32),
                 rotations=np.tile(np.array([1, 0, 0, 0], np.float32), (P, 1)))
    for deg in range(4):
        o = cpu_oracle.forward(scene, cam, np.zeros(3, np.float32), deg, 1.0)
        vis = o["radii"] > 0
        assert vis.sum() > P // 2
        np.testing.assert_allclose(o["rgb"][vis], g[f"rgb_deg{deg}"][vis], rtol=0, atol=2e-6)
        raw = g[f"rgb_raw_deg{deg}"]
        np.testing.assert_array_equal(o["clamped"][vis].astype(bool), (raw < 0)[vis])


@pytest.mark.parametrize("res", ["800x800", "1080p", "4k"])
def test_camera_helpers_match_reference(res):
    """scene/cameras.py:48-57 + utils/graphics_utils.py:38-71 vs synthetic.make_camera."""
    g = np.load(GOLDEN / "sh_camera_golden.npz")
    W, H, f = synthetic.RESOLUTIONS[res]
    cam = synthetic.make_camera(W, H, f)
    np.testing.assert_array_equal(cam.viewmatrix.numpy(), g[f"view_{res}"])
    np.testing.assert_array_equal(cam.projmatrix.numpy(), g[f"proj_{res}"])
    np.testing.assert_array_equal(cam.campos.numpy(), g[f"campos_{res}"])
    np.testing.assert_allclose([cam.tanfovx, cam.tanfovy], g[f"tan_{res}"], rtol=1e-15)


def test_oracle_backward_is_derivative_of_forward_f64():
    """Finite differences of the float64 forward vs the analytic backward. Inputs whose
    perturbation flips a discrete decision (alpha threshold, tile rect) are rare at eps=1e-6 and
    are excluded by a median-style criterion: 95% of the probed coordinates must agree to 1e-4."""
    P, W, H = 40, 48, 32
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=3, scale_mean=0.12)
    sc = {k: v.numpy().astype(np.float64) for k, v in scene.items()}
    bg = np.array([0.2, 0.5, 0.3])
    rng = np.random.RandomState(0)
    dpix = rng.randn(3, H, W)

    def loss(s):
        return float((cpu_oracle.forward(s, cam, bg, 3, 1.0, dtype=np.float64)["color"] * dpix).sum())

    o = cpu_oracle.forward(sc, cam, bg, 3, 1.0, dtype=np.float64)
    gr = cpu_oracle.backward(o, dpix)
    vis = np.nonzero(o["radii"] > 0)[0]
    assert len(vis) > 10
    checks = []
    # dL_dscales is w.r.t. mod*scale and dL_drotations w.r.t. the raw quaternion (A.11 quirks 9,10);
    # at scale_modifier 1 both are plain derivatives.
    for key, gname in (("means3D", "dL_dmeans3D"), ("opacities", "dL_dopacity"),
                       ("shs", "dL_dsh"), ("scales", "dL_dscales"), ("rotations", "dL_drotations")):
        for _ in range(12):
            i = int(rng.choice(vis))
            flat_idx = tuple([i] + [int(rng.randint(d)) for d in sc[key].shape[1:]])
            eps = 1e-6 * max(1.0, abs(sc[key][flat_idx]))
            sp = {k: v.copy() for k, v in sc.items()}
            sm = {k: v.copy() for k, v in sc.items()}
            sp[key][flat_idx] += eps
            sm[key][flat_idx] -= eps
            fd = (loss(sp) - loss(sm)) / (2 * eps)
            an = gr[gname][flat_idx]
            if key == "means3D":
                # the analytic mean gradient omits the path through the depth ordering (none) but
                # includes the screen-space path carried by dL_dmeans2D: already folded into
                # dL_dmeans3D by the preprocess backward.
                pass
            checks.append(abs(fd - an) <= 1e-4 * max(abs(fd), abs(an)) + 1e-7)
    assert np.mean(checks) >= 0.95, f"only {np.mean(checks):.2%} of finite-difference probes agree"


def test_oracle_edge_cases():
    cam = synthetic.make_camera(40, 24, 30.0)
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    # P == 0: image stays zero, not background (rasterize_points.cu:81)
    empty = dict(means3D=np.zeros((0, 3), np.float32), shs=np.zeros((0, 16, 3), np.float32),
                 opacities=np.zeros((0, 1), np.float32), scales=np.zeros((0, 3), np.float32),
                 rotations=np.zeros((0, 4), np.float32))
    o = cpu_oracle.forward(empty, cam, bg)
    assert o["num_rendered"] == 0 and not o["color"].any()
    # everything behind the camera: R == 0, image == background
    scene, _ = synthetic.make_tiny_scene(50, 40, 24, seed=1)
    sc = {k: v.numpy() for k, v in scene.items()}
    sc["means3D"][:, 2] = -50.0
    o = cpu_oracle.forward(sc, cam, bg)
    assert o["num_rendered"] == 0 and not o["radii"].any()
    np.testing.assert_array_equal(o["color"], np.broadcast_to(bg[:, None, None], (3, 24, 40)))
