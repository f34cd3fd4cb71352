"""GPU suite, part 2: the reference CUDA build side by side (when oracle/_ref travelled), the
full-size configurations of BASELINE.json through size-independent properties, the alternative
input modes, edge cases and the public autograd API."""
import math

import numpy as np
import pytest
import torch

import synthetic
import util
from test_gpu_parity import C_, to_cuda  # noqa: F401  (fixture)
from test_oracle_golden import GRAD_NAMES, grad_close

pytestmark = pytest.mark.gpu


def ref_backend_or_skip():
    from oracle import ref_backend
    if not ref_backend.available():
        pytest.skip("oracle/_ref/libref_rasterizer.so not present on this box")
    return ref_backend


def make_case(P, res, view_seed=0, seed=0):
    W, H, f = synthetic.RESOLUTIONS[res]
    return to_cuda(synthetic.make_gaussians(P, seed=seed)), synthetic.make_camera(W, H, f, view_seed)


def check_binning_properties(m, radii, R, P, W, H):
    """Holds for ANY correct implementation: keys rebuilt from (tile, depth) and sorted by torch
    (stable) must reproduce point_list and ranges exactly."""
    gx, gy = (W + 15) // 16, (H + 15) // 16
    tiles = m["tiles_touched"].long()
    assert int(tiles.sum()) == R
    assert torch.equal((tiles > 0), (radii > 0))
    vis = torch.nonzero(radii > 0).squeeze(1)
    xy, r = m["means2D"][vis], radii[vis].float()
    # tile rectangle (auxiliary.h:46-56) in float32, as specified
    x0 = ((xy[:, 0] - r) / 16).int().clamp(0, gx); y0 = ((xy[:, 1] - r) / 16).int().clamp(0, gy)
    x1 = ((((xy[:, 0] + r) + 16) - 1) / 16).int().clamp(0, gx)   # left to right, as specified
    y1 = ((((xy[:, 1] + r) + 16) - 1) / 16).int().clamp(0, gy)
    w, h = (x1 - x0).long(), (y1 - y0).long()
    assert torch.equal(w * h, tiles[vis])
    # expand instances: instance k of Gaussian i covers tile (y0 + k // w, x0 + k % w)
    rep = torch.repeat_interleave(torch.arange(len(vis), device="cuda"), w * h)
    first = torch.cumsum(w * h, 0) - w * h
    k = torch.arange(R, device="cuda") - first[rep]
    tile = (y0[rep].long() + k // w[rep]) * gx + x0[rep].long() + k % w[rep]
    depth_bits = m["depths"][vis].contiguous().view(torch.int32).long()[rep]
    keys = (tile << 32) | depth_bits
    order = torch.sort(keys, stable=True).indices
    assert torch.equal(vis[rep][order].int(), m["point_list"])
    st = tile[order]
    tid = torch.arange(gx * gy, device="cuda")
    lo = torch.searchsorted(st, tid, right=False); hi = torch.searchsorted(st, tid, right=True)
    empty = lo == hi
    exp = torch.stack([torch.where(empty, 0, lo), torch.where(empty, 0, hi)], 1).int()
    assert torch.equal(exp, m["ranges"])
    # a pixel cannot have blended more entries than its tile holds
    n = m["n_contrib"]
    ln = (m["ranges"][:, 1] - m["ranges"][:, 0]).view(gy, gx)
    per_pix = ln.repeat_interleave(16, 0)[:H].repeat_interleave(16, 1)[:, :W]
    assert bool((n <= per_pix).all())


def violations(a, b):
    """north_star's gradient bar, per element: |a - b| <= 1e-6 + 1e-4 |b|.
    -> (fraction of elements outside, largest error / tolerance)"""
    a_, b_ = a.double(), b.double()
    tol = 1e-6 + 1e-4 * b_.abs()
    err = (a_ - b_).abs()
    if err.numel() == 0:
        return 0.0, 0.0
    return (err > tol).double().mean().item(), (err / tol).max().item()


def compare_with_reference(C_, sc, cam, bg, deg=3, mod=1.0, bad_frac=1e-3, colors_precomp=None,
                           cov3D_precomp=None, prefiltered=False, backward=True):
    """Same inputs through the reference's own CUDA build (oracle/_ref) and through libbgr:
    integer buffers bit-equal, float intermediates bit-equal, image <= 1e-5, gradients at the
    north_star bar except for the fraction of elements on which two runs of the REFERENCE itself
    disagree (unordered float atomics; measured per call)."""
    ref = ref_backend_or_skip()
    P = sc["means3D"].shape[0]
    W, H = cam.image_width, cam.image_height
    kw = dict(colors_precomp=colors_precomp, cov3D_precomp=cov3D_precomp, prefiltered=prefiltered)
    a_r, o_r = util.native_forward(ref, sc, cam, bg, deg, mod, **kw)
    a_n, o_n = util.native_forward(C_, sc, cam, bg, deg, mod, **kw)
    assert o_r[0] == o_n[0]
    g = ref.parse_geom(o_r[3], P); im = ref.parse_img(o_r[5], W, H); b = ref.parse_binning(o_r[4], o_r[0])
    m = util.bgr_intermediates(o_n, P, W, H)
    vis = o_r[2] > 0
    assert torch.equal(o_r[2], o_n[2])
    assert torch.equal(g["tiles_touched"], m["tiles_touched"])
    assert torch.equal(b["point_list"], m["point_list"])
    assert torch.equal(im["ranges"][:m["ranges"].shape[0]], m["ranges"])
    assert torch.equal(im["n_contrib"], m["n_contrib"])
    # (the reference leaves geom.rgb unwritten when colours are precomputed)
    for k in ("depths", "means2D", "conic_opacity") + (("rgb",) if colors_precomp is None else ()):
        assert torch.equal(g[k][vis], m[k][vis]), k
    assert (o_r[1] - o_n[1]).abs().max().item() <= 1e-5
    assert (im["final_T"] - m["final_T"]).abs().max().item() <= 1e-5
    if not backward:
        return o_n
    dpix = synthetic.make_upstream_grad(H, W).cuda()
    g_r = util.native_backward(ref, a_r, o_r, dpix)
    g_r2 = util.native_backward(ref, a_r, o_r, dpix)   # the reference's own run-to-run spread
    g_n = util.native_backward(C_, a_n, o_n, dpix)
    for k, r, r2, n in zip(GRAD_NAMES, g_r, g_r2, g_n):
        assert r.shape == n.shape, k
        spread, _ = violations(r2, r)
        mine, _ = violations(n, r)
        assert torch.isfinite(n).all(), k
        # floor: 2e-5 of the elements, and never fewer than five (on a 20K-Gaussian scene 2e-5 of
        # dL_dcov3D is 2.4 elements, inside the run-to-run noise of the accumulation order)
        floor = max(2e-5, 5.0 / max(r.numel(), 1))
        assert mine <= max(3 * spread, floor), f"{k}: {mine:.2e} of elements off vs reference spread {spread:.2e}"
        # worst element against the tensor's magnitude: 1e-3, or 3x what two runs of the reference
        # differ by on their worst element (cancelling terms amplify the atomics' order noise)
        ref_norm_err = ((r2.double() - r.double()).abs().max() / r.double().abs().max().clamp_min(1e-30)).item() if r.numel() else 0.0
        ok, msg = grad_close(n.cpu().numpy(), r.cpu().numpy(), max_bad_frac=bad_frac,
                             max_norm_err=max(1e-3, 3 * ref_norm_err))
        assert ok, f"{k}: {msg} (reference run-to-run norm_err {ref_norm_err:.2e})"
    return o_n


def test_c1_100k_800x800_vs_reference_build(C_):
    sc, cam = make_case(100_000, "800x800")
    compare_with_reference(C_, sc, cam, torch.zeros(3))


def test_c1_random_background_degree2_vs_reference_build(C_):
    sc, cam = make_case(100_000, "800x800", view_seed=3, seed=5)
    compare_with_reference(C_, sc, cam, torch.tensor([0.3, 0.8, 0.5]), deg=2, mod=1.5)


def test_c2_3m_1080p_vs_reference_build(C_):
    sc, cam = make_case(3_000_000, "1080p")
    compare_with_reference(C_, sc, cam, torch.zeros(3))


@pytest.mark.parametrize("P,res,mod", [(3_000_000, "1080p", 1.0), (1_000_000, "4k", 2.0)])
def test_full_size_properties(C_, P, res, mod):
    sc, cam = make_case(P, res)
    W, H = cam.image_width, cam.image_height
    bg = torch.tensor([0.1, 0.2, 0.3])
    args, out = util.native_forward(C_, sc, cam, bg, 3, mod)
    R, color, radii, *_ = out
    m = util.bgr_intermediates(out, P, W, H)
    check_binning_properties(m, radii, R, P, W, H)
    assert torch.isfinite(color).all()
    T = m["final_T"]
    assert bool(((T >= 0) & (T <= 1)).all())
    # where nothing was blended the pixel is exactly background
    untouched = (m["n_contrib"] == 0)
    for ch in range(3):
        assert torch.equal(color[ch][untouched], torch.full_like(color[ch][untouched], bg[ch].item()) * T[untouched])
    # linearity of the backward in the upstream gradient, and determinism of the forward
    _, out2 = util.native_forward(C_, sc, cam, bg, 3, mod)
    assert torch.equal(out2[1], color) and torch.equal(out2[2], radii)
    dpix = synthetic.make_upstream_grad(H, W).cuda()
    g1 = util.native_backward(C_, args, out, dpix)
    g2 = util.native_backward(C_, args, out, 2.0 * dpix)
    for k, a, b in zip(GRAD_NAMES, g1, g2):
        ok, msg = grad_close(b.cpu().numpy(), 2.0 * a.cpu().numpy(), max_bad_frac=1e-3, max_norm_err=1e-3)
        assert ok, f"linearity {k}: {msg}"
    # culled Gaussians get exactly zero gradient in every output (the fused kernel writes them)
    dead = radii == 0
    for a in g1:
        assert not a[dead].any()


def test_precomputed_colour_and_covariance_modes_agree(C_):
    """gaussian_renderer/__init__.py:55-82: convert_SHs_python / compute_cov3D_python feed the
    same rasterizer through colors_precomp / cov3D_precomp."""
    P, W, H = 5000, 200, 120
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=31)
    sc = to_cuda(scene)
    bg = torch.tensor([0.2, 0.1, 0.0])
    _, base = util.native_forward(C_, sc, cam, bg, 3, 1.0)
    m = util.bgr_intermediates(base, P, W, H)
    vis = base[2] > 0
    rgb = torch.where(vis[:, None], m["rgb"], torch.zeros_like(m["rgb"])).contiguous()
    a1, o1 = util.native_forward(C_, sc, cam, bg, 3, 1.0, colors_precomp=rgb)
    assert torch.equal(o1[2], base[2]) and torch.equal(o1[1], base[1])
    # cov3D from scale/rotation in torch (scene/gaussian_model.py:27-31 semantics)
    q, s = sc["rotations"], sc["scales"]
    r, x, y, z = q.unbind(1)
    Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                      2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                      2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(P, 3, 3)
    L = Rm * s[:, None, :]
    Sig = L @ L.transpose(1, 2)
    cov = torch.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], 1).contiguous()
    a2, o2 = util.native_forward(C_, sc, cam, bg, 3, 1.0, cov3D_precomp=cov)
    assert (o2[2] != base[2]).float().mean().item() < 1e-3       # radii: ceil() of ~1e-7-different floats
    assert (o2[1] - base[1]).abs().max().item() < 2e-3
    # backward in the precomputed modes: dL_dcolors / dL_dcov3D are the live outputs
    dpix = synthetic.make_upstream_grad(H, W).cuda()
    gb = util.native_backward(C_, *_fw(C_, sc, cam, bg), dpix)
    g1 = util.native_backward(C_, a1, o1, dpix)
    assert torch.allclose(g1[1], gb[1], rtol=1e-4, atol=1e-7)     # dL_dcolors
    assert not g1[5].numel()                                      # no SH -> empty dL_dsh (P,0,3)
    g2 = util.native_backward(C_, a2, o2, dpix)
    assert not g2[6].any() and not g2[7].any()                    # no scale/rot gradient


def _fw(C_, sc, cam, bg):
    return util.native_forward(C_, sc, cam, bg, 3, 1.0)


def test_edge_cases(C_):
    cam = synthetic.make_camera(40, 24, 30.0)
    bg = torch.tensor([0.1, 0.2, 0.3])
    e = torch.Tensor([])
    # P == 0 -> zeros (not background), no buffers touched
    sc0 = dict(means3D=torch.zeros(0, 3), shs=torch.zeros(0, 16, 3), opacities=torch.zeros(0, 1),
               scales=torch.zeros(0, 3), rotations=torch.zeros(0, 4))
    a, o = util.native_forward(C_, to_cuda(sc0), cam, bg)
    assert o[0] == 0 and not o[1].any() and o[2].numel() == 0
    g = util.native_backward(C_, a, o, torch.ones(3, 24, 40))
    assert [t.shape for t in g] == [(0, 3), (0, 3), (0, 1), (0, 3), (0, 6), (0, 0, 3), (0, 3), (0, 4)]
    # everything behind the camera -> R == 0, image == background, gradients all zero
    scene, _ = synthetic.make_tiny_scene(300, 40, 24, seed=1)
    scene["means3D"][:, 2] = -50.0
    a, o = util.native_forward(C_, to_cuda(scene), cam, bg)
    assert o[0] == 0 and not o[2].any()
    assert torch.equal(o[1].cpu(), bg[:, None, None].expand(3, 24, 40))
    g = util.native_backward(C_, a, o, torch.ones(3, 24, 40))
    assert all(not t.any() for t in g)
    # one huge Gaussian covering every tile of a ragged image; opacity 1 -> alpha clamps at 0.99
    big = dict(means3D=torch.tensor([[0.0, 0.0, 0.0]]), shs=torch.ones(1, 16, 3) * 0.5,
               opacities=torch.ones(1, 1), scales=torch.ones(1, 3) * 500.0,
               rotations=torch.tensor([[1.0, 0, 0, 0]]))
    cam2 = synthetic.make_camera(50, 37, 40.0)
    a, o = util.native_forward(C_, to_cuda(big), cam2, bg)
    assert o[0] == 4 * 3 and int(o[2][0]) > 50
    m = util.bgr_intermediates(o, 1, 50, 37)
    assert bool((m["n_contrib"] == 1).all())
    assert torch.allclose(m["final_T"], torch.full_like(m["final_T"], 0.01), atol=1e-6)


def test_mark_visible(C_):
    sc, cam = make_case(200_000, "800x800")
    vis = C_.mark_visible(sc["means3D"], cam.viewmatrix.cuda(), cam.projmatrix.cuda())
    v = cam.viewmatrix.cuda()
    p = sc["means3D"]
    z = torch.addcmul(torch.addcmul(p[:, 1] * v[1, 2], p[:, 0], v[0, 2]), p[:, 2], v[2, 2]) + v[3, 2]
    assert vis.dtype == torch.bool and vis.shape == (200_000,)
    assert (vis != (z > 0.2)).float().mean().item() < 1e-5
    try:
        ref = ref_backend_or_skip()
    except pytest.skip.Exception:
        return
    assert torch.equal(vis, ref.mark_visible(p, v, cam.projmatrix.cuda()))


def test_public_autograd_api_and_debug_mode(C_):
    """The call a user makes: gaussian_renderer/__init__.py:85-93."""
    from diff_gaussian_rasterization import GaussianRasterizer
    P, W, H = 4000, 160, 96
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=41)
    sc = {k: v.cuda().requires_grad_(True) for k, v in scene.items()}
    bg = torch.tensor([0.0, 0.5, 1.0])
    outs = []
    for debug in (False, True):
        for t in sc.values():
            t.grad = None
        means2D = torch.zeros_like(sc["means3D"], requires_grad=True)
        means2D.retain_grad()
        rast = GaussianRasterizer(util.settings_for(cam, bg, debug=debug))
        color, radii = rast(means3D=sc["means3D"], means2D=means2D, shs=sc["shs"], colors_precomp=None,
                            opacities=sc["opacities"], scales=sc["scales"], rotations=sc["rotations"],
                            cov3D_precomp=None)
        assert color.shape == (3, H, W) and radii.dtype == torch.int32 and not radii.requires_grad
        (color * synthetic.make_upstream_grad(H, W).cuda()).sum().backward()
        assert means2D.grad.shape == (P, 3) and not means2D.grad[:, 2].any()
        assert means2D.grad[radii > 0].abs().sum() > 0
        outs.append((color.detach().clone(), {k: v.grad.clone() for k, v in sc.items()}))
        assert sc["opacities"].grad.shape == (P, 1) and sc["shs"].grad.shape == (P, 16, 3)
    assert torch.equal(outs[0][0], outs[1][0])
    for k in sc:
        assert torch.allclose(outs[0][1][k], outs[1][1][k], rtol=1e-3, atol=1e-7)
    # render under no_grad (render.py:38) returns plain tensors
    with torch.no_grad():
        c2, _ = rast(means3D=sc["means3D"], means2D=means2D, shs=sc["shs"], opacities=sc["opacities"],
                     scales=sc["scales"], rotations=sc["rotations"])
    assert torch.equal(c2, outs[0][0]) and not c2.requires_grad


def test_non_default_stream_and_second_device_context(C_):
    """All work is enqueued on torch's current stream."""
    P, W, H = 3000, 128, 64
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=51)
    sc = to_cuda(scene)
    bg = torch.zeros(3)
    _, base = util.native_forward(C_, sc, cam, bg)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        _, o = util.native_forward(C_, sc, cam, bg)
    s.synchronize()
    assert torch.equal(o[1], base[1])


def test_staged_backward_matches_single_call(C_):
    """BgrBackwardArgs.stages: ACCUMULATE once + PREPROCESS over Gaussian ranges must give what the
    single call gives (two executions differ only by the atomics' summation order)."""
    P = 50_001          # ragged: the last range is 1 Gaussian, bounds are not multiples of 4
    scene, cam = make_case(P, "800x800", view_seed=2)
    args, out = util.native_forward(C_, scene, cam, torch.zeros(3), 3, 1.0)
    dpix = synthetic.make_upstream_grad(cam.image_height, cam.image_width, seed=5)
    whole = [g.clone() for g in util.native_backward(C_, args, out, dpix)]
    C_.set_backward_ranges([0, 1, 4097, 20_000, 50_000, P])
    try:
        staged = util.native_backward(C_, args, out, dpix)
    finally:
        C_.set_backward_ranges(None)
    for name, a, b in zip(GRAD_NAMES, staged, whole):
        ok, msg = grad_close(a.cpu().numpy(), b.cpu().numpy(), scale=1e3)
        assert ok, f"{name}: {msg}"
    # rows of culled Gaussians are exact zeros in both
    culled = out[2] <= 0
    assert float(staged[3][culled].abs().max()) == 0.0 and float(staged[5][culled].abs().max()) == 0.0


@pytest.mark.parametrize("world,degree", [(1, 3), (3, 2), (8, 3)])
def test_sh_gradient_rebuild_from_factors(C_, world, degree):
    """View-parallel path, single-GPU part: dL_dsh = sum_v basis(dir_v) x dL/dRGB_v from `world`
    factor slots (local memory stands in for the multicast buffer) against an fp64 restatement of
    the reference's per-view SH backward (backward.cu:21-137), full range and a sub-range."""
    P = 70_003
    g = torch.Generator().manual_seed(world * 10 + degree)
    means = (torch.rand(P, 3, generator=g) * 10 - 5).cuda()
    campos = torch.zeros(world, 4)
    campos[:, :3] = torch.randn(world, 3, generator=g) * 8
    factors = torch.randn(world, P, 3, generator=g) * (torch.rand(world, P, 1, generator=g) < 0.4)
    campos, factors = campos.cuda(), factors.cuda()
    C0, C1 = 0.28209479177387814, 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
          -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
    want = torch.zeros(P, 16, 3, dtype=torch.float64, device="cuda")
    for v in range(world):
        d = means.double() - campos[v, :3].double()
        d = d / d.norm(dim=1, keepdim=True)
        x, y, z = d[:, 0], d[:, 1], d[:, 2]
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b = [torch.full_like(x, C0), -C1 * y, C1 * z, -C1 * x, C2[0] * xy, C2[1] * yz,
             C2[2] * (2 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy), C3[0] * y * (3 * xx - yy),
             C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy), C3[3] * z * (2 * zz - 3 * xx - 3 * yy),
             C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy), C3[6] * x * (xx - 3 * yy)]
        B = torch.stack(b, 1)[:, :(degree + 1) ** 2]
        want[:, :B.shape[1]] += B[:, :, None] * factors[v].double()[:, None, :]
    dsh = torch.full((P, 16, 3), float("nan"), device="cuda")
    C_.sh_grad_from_factors(P, degree, 16, world, means, campos.data_ptr(), factors.data_ptr(), dsh, "cuda")
    assert (dsh.double() - want).abs().max().item() < 1e-5      # fp32 products and sums, |x| <~ 10
    assert float(dsh[:, (degree + 1) ** 2:].abs().max()) == 0.0 if degree < 3 else True
    # sub-range: rows outside [begin, end) are left alone, rows inside are identical to the full call
    part = torch.full((P, 16, 3), 7.0, device="cuda")
    C_.sh_grad_from_factors_range(P, 1025, P - 30, degree, 16, world, means, campos.data_ptr(),
                                  factors.data_ptr(), part, "cuda")
    assert float(part[:1025].min()) == 7.0 and float(part[P - 30:].min()) == 7.0 and float(part[P - 30:].max()) == 7.0
    assert torch.equal(part[1025:P - 30], dsh[1025:P - 30])
