"""GPU suite, part 3 (round 2): the BASELINE.json configurations and product code paths that the
first round never reached, all against the reference's own CUDA build (oracle/_ref):

  * configs[4] 4K render stress: 2M Gaussians / scale_modifier 2.0, 5M / 0.5, 10M / degree 0
    (rasterizer_impl.cu:289-318 sorts 47 key bits there); the global-memory path of the per-tile
    sort forced on small and mid scenes; exact depth ties;
  * configs[3] 6M Gaussians at 1080p;
  * colors_precomp / cov3D_precomp input modes (null-pointer switches rasterizer_impl.cu:321,389,411),
    SH tensors with M = 4 / 9 coefficients and a misaligned SH base (the non-vectorised kernels);
  * prefiltered, the debug-mode error + snapshot path, the overflow guard;
  * gradients arbitrated by the float64 oracle: |ours - f64| against |reference - f64|.
"""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

import synthetic
import util
from test_gpu_full import compare_with_reference, make_case, ref_backend_or_skip, violations
from test_gpu_parity import C_, to_cuda  # noqa: F401  (fixture)
from test_oracle_golden import GRAD_NAMES

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(autouse=True)
def _needs_a_cuda_device():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


# ------------------------------------------------------------------------------------------------
# configs[4] and configs[3]
# ------------------------------------------------------------------------------------------------
def test_c4_4k_2m_scale_modifier_2_vs_reference_build(C_):
    sc, cam = make_case(2_000_000, "4k")
    compare_with_reference(C_, sc, cam, torch.zeros(3), deg=3, mod=2.0)
    info = C_.last_forward_info()
    assert info["nb_idx"] == 21 and info["global_sort_path"] == 0 and info["max_tile_list"] <= info["smem_keys"]


def test_c4_4k_10m_degree0_vs_reference_build(C_):
    """24 index bits + 32 depth bits per bucket key; 39M instances, lists of up to ~1 800."""
    sc, cam = make_case(10_000_000, "4k")
    compare_with_reference(C_, sc, cam, torch.tensor([0.2, 0.4, 0.6]), deg=0, mod=1.0)
    info = C_.last_forward_info()
    assert info["nb_idx"] == 24 and info["nb_depth"] == 32 and info["R"] > 30_000_000, info


def test_c4_4k_5m_scale_modifier_half_degree1(C_):
    sc, cam = make_case(5_000_000, "4k", view_seed=2)
    compare_with_reference(C_, sc, cam, torch.zeros(3), deg=1, mod=0.5)


def test_c3_6m_1080p_vs_reference_build(C_):
    sc, cam = make_case(6_000_000, "1080p", view_seed=5)
    compare_with_reference(C_, sc, cam, torch.zeros(3))


def test_global_memory_sort_path_on_small_scenes():
    """BGR_TILE_SORT_SMEM_KEYS=96 (read per call) shrinks the shared-memory capacity of the per-tile
    sort so that every longer list takes the global-memory path; the reference-fixture and oracle
    parity tests are re-run that way in a child process."""
    env = dict(os.environ, BGR_TILE_SORT_SMEM_KEYS="96", BGR_EXPECT_GLOBAL_SORT="1")
    r = subprocess.run([sys.executable, "-m", "pytest", str(ROOT / "tests" / "test_gpu_parity.py"), "-x", "-q",
                        "-m", "gpu", "-k", "golden or oracle", "-p", "no:cacheprovider"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "failed" not in r.stdout


def test_global_memory_sort_path_mid_scene_vs_reference_build(C_, monkeypatch):
    monkeypatch.setenv("BGR_TILE_SORT_SMEM_KEYS", "128")
    sc, cam = make_case(1_000_000, "1080p", view_seed=1)
    compare_with_reference(C_, sc, cam, torch.zeros(3))
    info = C_.last_forward_info()
    assert info["global_sort_path"] == 1 and info["max_tile_list"] > 128


def test_exact_depth_ties_keep_index_order_vs_reference_build(C_):
    """Gaussians with bit-identical view depth in one tile: the reference's stable sort leaves
    them in emission (= index) order (rasterizer_impl.cu:98-108, 303-308). Here they land in their
    bucket in atomic order, so the per-tile sort has to notice the tie and sort the index digits."""
    P = 60_000
    sc, cam = make_case(P, "800x800", view_seed=2, seed=12)
    g = torch.Generator().manual_seed(5)
    pick = torch.randint(0, 300, (P,), generator=g).cuda()          # 300 distinct centres, ~200 copies each
    sc["means3D"] = sc["means3D"][:300][pick].contiguous()
    sc["scales"] = (sc["scales"] * 3.0).contiguous()
    # (hundreds of coincident splats per pixel: the reference's own atomics are noisy here, so the
    # per-element gradient allowance is the run-to-run one only; this test is about the order)
    o_n = compare_with_reference(C_, sc, cam, torch.tensor([0.1, 0.1, 0.1]), bad_frac=1e-2)
    m = util.bgr_intermediates(o_n, P, cam.image_width, cam.image_height)
    # there really are adjacent list entries with equal depth
    pl = m["point_list"].long()
    d = m["depths"][pl]
    assert int((d[1:] == d[:-1]).sum()) > 1000


# ------------------------------------------------------------------------------------------------
# input modes
# ------------------------------------------------------------------------------------------------
def _cov3d_torch(sc, mod=1.0):
    """scene/gaussian_model.py:27-31 (build_covariance_from_scaling_rotation + strip_symmetric)."""
    q, s = sc["rotations"], sc["scales"] * mod
    r, x, y, z = q.unbind(1)
    Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                      2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                      2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    L = Rm * s[:, None, :]
    Sig = L @ L.transpose(1, 2)
    return torch.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], 1).contiguous()


def test_colors_precomp_mode_vs_reference_build(C_):
    sc, cam = make_case(200_000, "800x800", view_seed=4, seed=3)
    g = torch.Generator().manual_seed(9)
    colors = torch.rand(sc["means3D"].shape[0], 3, generator=g).cuda()
    compare_with_reference(C_, sc, cam, torch.tensor([0.1, 0.0, 0.3]), colors_precomp=colors)


def test_cov3d_precomp_mode_vs_reference_build(C_):
    sc, cam = make_case(200_000, "800x800", view_seed=6, seed=4)
    cov = _cov3d_torch(sc)
    compare_with_reference(C_, sc, cam, torch.zeros(3), cov3D_precomp=cov)


def test_both_precomp_modes_vs_reference_build(C_):
    sc, cam = make_case(150_000, "1080p", view_seed=7, seed=6)
    g = torch.Generator().manual_seed(10)
    colors = torch.rand(sc["means3D"].shape[0], 3, generator=g).cuda()
    compare_with_reference(C_, sc, cam, torch.tensor([0.5, 0.5, 0.5]), colors_precomp=colors,
                           cov3D_precomp=_cov3d_torch(sc))


@pytest.mark.parametrize("M,deg", [(4, 1), (9, 2), (9, 1), (1, 0), (25, 3)])
def test_sh_tensors_with_other_coefficient_counts_vs_reference_build(C_, M, deg):
    """M != 16 runs the non-vectorised SH loads/stores of both preprocess kernels."""
    sc, cam = make_case(120_000, "800x800", view_seed=M, seed=M)
    shs = sc["shs"]
    if M <= 16:
        sc["shs"] = shs[:, :M, :].contiguous()
    else:
        sc["shs"] = torch.cat([shs, 0.05 * torch.ones(shs.shape[0], M - 16, 3, device="cuda")], 1).contiguous()
    compare_with_reference(C_, sc, cam, torch.zeros(3), deg=deg)


def test_misaligned_sh_base_vs_reference_build(C_):
    """A (P,16,3) view starting 4 bytes into its storage: contiguous, but not 16-byte aligned."""
    sc, cam = make_case(100_000, "800x800", view_seed=9, seed=8)
    P = sc["means3D"].shape[0]
    store = torch.empty(P * 48 + 1, device="cuda")
    store[1:].copy_(sc["shs"].reshape(-1))
    sc["shs"] = store[1:].view(P, 16, 3)
    assert sc["shs"].is_contiguous() and sc["shs"].data_ptr() % 16 == 4
    rot = torch.empty(P * 4 + 1, device="cuda")
    rot[1:].copy_(sc["rotations"].reshape(-1))
    sc["rotations"] = rot[1:].view(P, 4)
    compare_with_reference(C_, sc, cam, torch.zeros(3))


def test_prefiltered_flag_on_an_all_visible_scene(C_):
    """prefiltered=True only arms the device trap for culled points (auxiliary.h:156-160); with
    every Gaussian in front of the near plane the results are those of prefiltered=False."""
    P, W, H = 20_000, 320, 200
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=61)
    sc = to_cuda(scene)
    assert bool(C_.mark_visible(sc["means3D"], cam.viewmatrix.cuda(), cam.projmatrix.cuda()).all())
    bg = torch.tensor([0.3, 0.2, 0.1])
    _, base = util.native_forward(C_, sc, cam, bg)
    _, pre = util.native_forward(C_, sc, cam, bg, prefiltered=True)
    assert pre[0] == base[0] and torch.equal(pre[1], base[1]) and torch.equal(pre[2], base[2])
    compare_with_reference(C_, sc, cam, bg, prefiltered=True)


_TRAP_CHILD = r"""
import sys, os
sys.path[:0] = [{root!r}, {root!r} + "/gaussian-splatting-own_b200", {root!r} + "/tests"]
import torch, synthetic, util
from diff_gaussian_rasterization import GaussianRasterizer
P, W, H = 3000, 96, 64
scene, cam = synthetic.make_tiny_scene(P, W, H, seed=71)
scene["means3D"][::7, 2] = -40.0            # behind the camera: culled => the prefiltered trap fires
sc = {{k: v.cuda() for k, v in scene.items()}}
rs = util.settings_for(cam, torch.zeros(3), debug=True)._replace(prefiltered=True)
try:
    GaussianRasterizer(rs)(means3D=sc["means3D"], means2D=torch.zeros_like(sc["means3D"]), shs=sc["shs"],
                           opacities=sc["opacities"], scales=sc["scales"], rotations=sc["rotations"])
except RuntimeError as e:
    print("RAISED:", e, flush=True)
    sys.stderr.flush()
    os._exit(42)          # the CUDA context is gone: skip torch's teardown
os._exit(0)
"""


def test_debug_mode_reports_the_failing_stage_and_writes_the_snapshot(tmp_path):
    """reference __init__.py:83-90 + auxiliary.h:166-173: with debug=True a kernel fault surfaces
    as a RuntimeError right after the stage that caused it, and the CPU copy of the arguments is
    saved to snapshot_fw.dump. The fault is the `prefiltered` trap (auxiliary.h:156-160); it
    poisons the CUDA context, hence the child process."""
    r = subprocess.run([sys.executable, "-c", _TRAP_CHILD.format(root=str(ROOT))], cwd=tmp_path,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 42, r.stdout[-1500:] + r.stderr[-1500:]
    assert "RAISED:" in r.stdout and "preprocess" in r.stdout
    assert "An error occured in forward" in r.stdout
    assert "[CUDA ERROR]" in r.stderr
    dump = tmp_path / "snapshot_fw.dump"
    assert dump.exists()
    args = torch.load(dump, weights_only=False)
    assert len(args) == 19 and args[1].shape == (3000, 3) and args[17] is True and args[18] is True
    assert all(not t.is_cuda for t in args if isinstance(t, torch.Tensor))


@pytest.mark.parametrize("P", [40_000, 140_000])
def test_overflow_guard_counts_in_64_bits(C_, P):
    """P giant Gaussians x 32 400 tiles at 4K: 1.3e9 >= 2^30 instances, and at P = 140 000 the
    true total 4.5e9 wraps a 32-bit counter to 2.4e8. Both must be refused before any binning
    buffer is sized (the reference overflows its int silently, rasterizer_impl.cu:280-283)."""
    W, H, f = synthetic.RESOLUTIONS["4k"]
    cam = synthetic.make_camera(W, H, f)
    sc = dict(means3D=torch.zeros(P, 3), shs=torch.full((P, 16, 3), 0.2), opacities=torch.full((P, 1), 0.5),
              scales=torch.full((P, 3), 800.0), rotations=torch.tensor([[1.0, 0, 0, 0]]).repeat(P, 1))
    with pytest.raises(RuntimeError, match=r"2\^30"):
        util.native_forward(C_, to_cuda(sc), cam, torch.zeros(3))


# ------------------------------------------------------------------------------------------------
# gradients arbitrated in float64
# ------------------------------------------------------------------------------------------------
def _f64_backward_of_reference_forward(ref, sc, cam, bg, deg, mod, o_r, dpix):
    """The plain-C oracle's float64 backward (oracle/rasterizer_oracle.c, backward.cu:144-557
    restated) applied to the float32 forward state the REFERENCE build produced on this GPU: same
    lists, same per-pixel final_T / n_contrib, same conics. What it returns is what exact
    arithmetic makes of the very inputs both float32 backward passes start from."""
    from oracle import cpu_oracle
    P = sc["means3D"].shape[0]
    W, H = cam.image_width, cam.image_height
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    g = ref.parse_geom(o_r[3], P); im = ref.parse_img(o_r[5], W, H); b = ref.parse_binning(o_r[4], o_r[0])
    f8 = lambda t: np.ascontiguousarray(t.detach().cpu().numpy().astype(np.float64))  # noqa: E731
    vis = (o_r[2] > 0).cpu().numpy()
    zero_culled = lambda a: np.where(vis.reshape((-1,) + (1,) * (a.ndim - 1)), a, 0)  # noqa: E731
    fwd = {
        "radii": o_r[2].cpu().numpy().astype(np.int32),
        "ranges": np.ascontiguousarray(im["ranges"][:tiles].cpu().numpy().view(np.uint32)),
        "point_list": np.ascontiguousarray(b["point_list"].cpu().numpy().view(np.uint32)),
        "means2D": zero_culled(f8(g["means2D"])), "conic_opacity": zero_culled(f8(g["conic_opacity"])),
        "rgb": zero_culled(f8(g["rgb"])), "cov3D": zero_culled(f8(g["cov3D"])),
        "clamped": np.ascontiguousarray(zero_culled(g["clamped"].cpu().numpy().astype(np.uint8))),
        "final_T": f8(im["final_T"]), "n_contrib": np.ascontiguousarray(im["n_contrib"].cpu().numpy().view(np.uint32)),
    }
    fwd["_ctx"] = dict(means=f8(sc["means3D"]), shs=f8(sc["shs"]), colors=None, scales=f8(sc["scales"]),
                       rots=f8(sc["rotations"]), cov_pre=None, view=f8(cam.viewmatrix).reshape(-1),
                       proj=f8(cam.projmatrix).reshape(-1), campos=f8(cam.campos).reshape(-1), bg=f8(bg).reshape(-1),
                       P=P, M=sc["shs"].shape[1], W=W, H=H, D=deg, mod=float(mod), tanx=float(cam.tanfovx),
                       tany=float(cam.tanfovy), dtype=np.float64)
    return cpu_oracle.backward(fwd, dpix.cpu().numpy().astype(np.float64))


@pytest.mark.parametrize("P,W,H,seed,deg,mod", [
    (20_000, 400, 304, 81, 3, 1.0),       # dense tiny-scene splats (long lists, deep blending)
    (12_000, 333, 210, 82, 2, 1.6),       # ragged image, elongated splats
])
def test_gradients_arbitrated_by_the_float64_oracle(C_, P, W, H, seed, deg, mod):
    """north_star: gradients within 1e-4 relative / 1e-6 absolute. Two float32 implementations
    that accumulate in different orders cannot be compared with each other at that bar on
    cancelling terms, so both are measured against float64: per tensor, our worst error and our
    count of elements outside the bar may not exceed 1.5 x the reference's own (+ a floor of two
    float32 ulps of the tensor's magnitude / a handful of elements)."""
    ref = ref_backend_or_skip()
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=seed)
    sc = to_cuda(scene)
    bg = torch.tensor([0.2, 0.1, 0.3])
    a_r, o_r = util.native_forward(ref, sc, cam, bg, deg, mod)
    a_n, o_n = util.native_forward(C_, sc, cam, bg, deg, mod)
    assert o_r[0] == o_n[0] and torch.equal(o_r[2], o_n[2])
    dpix = synthetic.make_upstream_grad(H, W, seed=seed).cuda()
    g_r = util.native_backward(ref, a_r, o_r, dpix)
    g_n = util.native_backward(C_, a_n, o_n, dpix)
    g64 = _f64_backward_of_reference_forward(ref, sc, cam, bg, deg, mod, o_r, dpix)
    report = []
    for k, r, n in zip(GRAD_NAMES, g_r, g_n):
        want = torch.tensor(g64[k].reshape(tuple(r.shape)), device="cuda")
        e_r = (r.double() - want).abs()
        e_n = (n.double() - want).abs()
        mag = want.abs().max().item()
        floor = 2.4e-7 * mag
        bad_r, worst_r = violations(r, want)
        bad_n, worst_n = violations(n, want)
        report.append(f"{k}: max err ours {e_n.max().item():.3e} ref {e_r.max().item():.3e} (|g|max {mag:.3e}); "
                      f"outside bar ours {bad_n:.2e} ref {bad_r:.2e}; worst/tol ours {worst_n:.1f} ref {worst_r:.1f}")
        assert e_n.max().item() <= 1.5 * e_r.max().item() + floor, report[-1]
        assert bad_n <= 1.5 * bad_r + 8.0 / max(want.numel(), 1), report[-1]
        # rms error: the typical element, not only the worst one
        assert e_n.pow(2).mean().sqrt().item() <= 1.5 * e_r.pow(2).mean().sqrt().item() + floor, report[-1]
    print("\n".join(report))
