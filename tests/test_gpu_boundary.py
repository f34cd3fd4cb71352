"""GPU suite, part 4: parity THROUGH THE REFERENCE'S OWN BINDING SHAPE. tests/boundary/
rasterize_points_bgr.cpp is the file a maintainer of the reference would put in place of
rasterize_points.cu (rasterize_points.h:16-67, ext.cpp:15-18): the same three pybind functions with
torch::Tensor signatures, the same resize-lambda allocation of the geometry / binning / image
buffers, but calling include/bgr.h instead of CudaRasterizer::Rasterizer. It is compiled as a torch
extension (`bgr_boundary`, built in tree by __graft_entry__.build()) and linked against libbgr.so.
Here the parity cases run through that extension - no ctypes in the path - and the reference's
autograd class shape runs on top of it."""
import sys
from pathlib import Path

import pytest
import torch

import synthetic
import util
from test_gpu_full import compare_with_reference, make_case
from test_gpu_parity import C_  # noqa: F401  (fixture)
from test_gpu_round2 import _cov3d_torch

pytestmark = pytest.mark.gpu

BOUNDARY_DIR = Path(__file__).resolve().parent / "boundary"


@pytest.fixture(scope="module")
def ext():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    sys.path.insert(0, str(BOUNDARY_DIR))
    try:
        import bgr_boundary   # must have been built (tests/boundary/build.py); no fallback
    finally:
        sys.path.remove(str(BOUNDARY_DIR))
    return bgr_boundary


def test_extension_exports_the_reference_functions(ext):
    for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(ext, name))


def test_c1_through_the_extension_vs_reference_build(ext):
    sc, cam = make_case(100_000, "800x800")
    compare_with_reference(ext, sc, cam, torch.zeros(3))


def test_c2_3m_1080p_through_the_extension_vs_reference_build(ext):
    sc, cam = make_case(3_000_000, "1080p")
    compare_with_reference(ext, sc, cam, torch.zeros(3), bad_frac=2e-3)


def test_precomputed_inputs_through_the_extension(ext):
    sc, cam = make_case(200_000, "800x800", view_seed=6, seed=4)
    P = sc["means3D"].shape[0]
    g = torch.Generator().manual_seed(3)
    col = torch.rand(P, 3, generator=g)
    cov = _cov3d_torch(sc)
    compare_with_reference(ext, sc, cam, torch.tensor([0.1, 0.2, 0.3]), colors_precomp=col)
    compare_with_reference(ext, sc, cam, torch.zeros(3), cov3D_precomp=cov)


def test_extension_and_ctypes_binding_agree_bit_for_bit(ext, C_):
    """Two bindings, one library: everything the forward returns is bit-identical."""
    sc, cam = make_case(200_000, "800x800", view_seed=1)
    bg = torch.tensor([0.2, 0.4, 0.6])
    _, a = util.native_forward(ext, sc, cam, bg)
    _, b = util.native_forward(C_, sc, cam, bg)
    assert a[0] == b[0]
    assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
    P, W, H = sc["means3D"].shape[0], cam.image_width, cam.image_height
    ma, mb = util.bgr_intermediates(a, P, W, H), util.bgr_intermediates(b, P, W, H)
    for k in ("tiles_touched", "point_list", "ranges", "n_contrib", "final_T"):
        assert torch.equal(ma[k], mb[k]), k


def test_mark_visible_through_the_extension(ext, C_):
    sc, cam = make_case(50_000, "800x800", view_seed=4)
    a = ext.mark_visible(sc["means3D"], cam.viewmatrix.cuda(), cam.projmatrix.cuda())
    b = C_.mark_visible(sc["means3D"], cam.viewmatrix.cuda(), cam.projmatrix.cuda())
    assert a.dtype == torch.bool and torch.equal(a, b) and 0 < int(a.sum()) < a.numel()


def test_reference_autograd_shape_on_top_of_the_extension(ext, C_):
    """The Python layer (the reference's __init__.py shape: GaussianRasterizer -> autograd Function
    -> _C.rasterize_gaussians) with the extension in the place of _C: a full forward + backward
    gives the same image and gradients as with the ctypes binding."""
    import diff_gaussian_rasterization as d
    from test_gpu_full import violations
    sc, cam = make_case(80_000, "800x800", view_seed=5)
    bg = torch.tensor([0.0, 0.1, 0.0])
    dpix = synthetic.make_upstream_grad(cam.image_height, cam.image_width).cuda()
    fn = d.make_rasterize_function(ext)

    class OverExtension(d.GaussianRasterizer):
        _rasterize = staticmethod(lambda *a: fn.apply(*a))
        _backend = ext

    def run(cls):
        leaves = {k: v.clone().requires_grad_(True) for k, v in sc.items()}
        m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
        r = cls(util.settings_for(cam, bg))
        img, radii = r(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"],
                       scales=leaves["scales"], rotations=leaves["rotations"])
        img.backward(dpix)
        vis = r.markVisible(leaves["means3D"])
        return img.detach(), radii, {k: v.grad for k, v in leaves.items()}, m2d.grad, vis

    img_c, radii_c, g_c, m_c, vis_c = run(d.GaussianRasterizer)
    img_e, radii_e, g_e, m_e, vis_e = run(OverExtension)
    assert torch.equal(img_c, img_e) and torch.equal(radii_c, radii_e) and torch.equal(vis_c, vis_e)
    for k in g_c:
        assert g_e[k] is not None and g_e[k].shape == g_c[k].shape, k
        assert violations(g_e[k], g_c[k])[0] <= 1e-3, k   # same kernels; only the atomics' order differs
    assert violations(m_e, m_c)[0] <= 1e-3


def test_reference_pybind_extension_agrees_with_the_shim_arm():
    """The reference's own torch extension (oracle/build_ref_ext.py: its setup.py sources,
    unmodified) against oracle/_ref/libref_rasterizer.so (same .cu files behind ref_shim.cu), which
    every other comparison runs through: identical integer buffers and image, gradients within the
    reference's own run-to-run spread."""
    from oracle import build_ref_ext, ref_backend
    from test_gpu_full import violations
    stock = build_ref_ext.load()
    if stock is None or not ref_backend.available():
        pytest.skip("reference builds not present on this box")
    sc, cam = make_case(100_000, "800x800", view_seed=2)
    bg = torch.tensor([0.1, 0.2, 0.3])
    P, W, H = sc["means3D"].shape[0], cam.image_width, cam.image_height
    a_s, o_s = util.native_forward(stock, sc, cam, bg)
    a_r, o_r = util.native_forward(ref_backend, sc, cam, bg)
    assert o_s[0] == o_r[0] and torch.equal(o_s[1], o_r[1]) and torch.equal(o_s[2], o_r[2])
    gs, gr = ref_backend.parse_geom(o_s[3], P), ref_backend.parse_geom(o_r[3], P)
    vis = o_r[2] > 0
    for k in ("depths", "means2D", "conic_opacity", "rgb"):
        assert torch.equal(gs[k][vis], gr[k][vis]), k
    assert torch.equal(gs["tiles_touched"], gr["tiles_touched"])
    assert torch.equal(ref_backend.parse_binning(o_s[4], o_s[0])["point_list"],
                       ref_backend.parse_binning(o_r[4], o_r[0])["point_list"])
    i_s, i_r = ref_backend.parse_img(o_s[5], W, H), ref_backend.parse_img(o_r[5], W, H)
    assert torch.equal(i_s["n_contrib"], i_r["n_contrib"]) and torch.equal(i_s["final_T"], i_r["final_T"])
    dpix = synthetic.make_upstream_grad(H, W).cuda()
    g_s = util.native_backward(stock, a_s, o_s, dpix)
    g_r = util.native_backward(ref_backend, a_r, o_r, dpix)
    g_r2 = util.native_backward(ref_backend, a_r, o_r, dpix)
    for s_, r_, r2_ in zip(g_s, g_r, g_r2):
        assert s_.shape == r_.shape
        assert violations(s_, r_)[0] <= max(3 * violations(r2_, r_)[0], 2e-5)
