"""GPU suite (pytest -m gpu): the sm_100a CUDA path, called through the C ABI of libbgr.so, against

  * the committed golden fixtures produced by the reference's own CUDA code (tests/golden/),
  * the CPU oracle (oracle/rasterizer_oracle.c) on fresh seeded scenes,
  * the reference CUDA build itself when oracle/_ref/libref_rasterizer.so travelled to the box,
  * size-independent properties at BASELINE.json's full sizes (3M @ 1080p, 4K stress).

Bars (north_star): radii, tiles_touched, point_list, ranges, n_contrib bit-exact; image <= 1e-5
max-abs; gradients 1e-4 relative / 1e-6 absolute per element, with the allowance explained in
test_oracle_golden.py (two runs of the reference differ by more than that on ~2e-4 of the
dL_dcov3D elements; measured numbers in DESIGN.md).
"""
import os
from pathlib import Path

import numpy as np
import pytest
import torch

import synthetic
import util
from test_oracle_golden import GRAD_NAMES, grad_close, load_case

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"


@pytest.fixture(scope="module")
def C_():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from diff_gaussian_rasterization import _C
    _C.lib()  # must exist: there is no fallback
    return _C


def to_cuda(scene):
    return {k: torch.as_tensor(v).cuda() for k, v in scene.items()}


# ------------------------------------------------------------------------------------------------
# golden fixtures of the reference's CUDA code
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_against_reference_golden_fixture(C_, name):
    g, scene, cam = load_case(name)
    P, W, H = int(g["P"]), int(g["W"]), int(g["H"])
    sc = to_cuda(scene)
    bg = torch.tensor(g["bg"])
    args, out = util.native_forward(C_, sc, cam, bg, int(g["sh_degree"]), float(g["scale_modifier"]))
    R, color, radii, *_ = out
    if os.environ.get("BGR_EXPECT_GLOBAL_SORT"):   # child run of test_global_memory_sort_path_on_small_scenes
        info = C_.last_forward_info()
        assert info["smem_keys"] == 96 and info["global_sort_path"] == int(info["max_tile_list"] > 96)
    m = util.bgr_intermediates(out, P, W, H)
    vis = torch.tensor(g["radii"] > 0, device="cuda")
    assert R == int(g["num_rendered"])
    for mine, key in ((radii, "radii"), (m["tiles_touched"], "tiles_touched"),
                      (m["point_list"], "point_list"), (m["ranges"], "ranges"),
                      (m["n_contrib"], "n_contrib")):
        assert torch.equal(mine.cpu(), torch.tensor(g[key]).view(mine.dtype).reshape(mine.shape)), key
    # float intermediates are bit-identical too (FMA contraction pinned to the reference build)
    for key in ("depths", "means2D", "conic_opacity", "rgb"):
        ref = torch.tensor(g[key], device="cuda")
        assert torch.equal(m[key][vis], ref[vis]), key
    assert (color.cpu() - torch.tensor(g["color"])).abs().max().item() <= 1e-5
    assert (m["final_T"].cpu() - torch.tensor(g["final_T"])).abs().max().item() <= 1e-5

    grads = util.native_backward(C_, args, out, torch.tensor(g["dL_dpix"]))
    for k, t in zip(GRAD_NAMES, grads):
        ok, msg = grad_close(t.cpu().numpy(), g[k].reshape(t.shape), scale=1e3)
        assert ok, f"{k}: {msg}"


# ------------------------------------------------------------------------------------------------
# CPU oracle on fresh scenes (different seeds / sizes / modes from the fixtures)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("P,W,H,seed,deg,mod,bg", [
    (800, 128, 96, 21, 3, 1.0, (0.0, 0.0, 0.0)),
    (2500, 161, 83, 22, 1, 1.7, (0.9, 0.1, 0.4)),      # ragged size, large splats
    (64, 16, 16, 23, 3, 1.0, (1.0, 1.0, 1.0)),         # a single tile
    (1, 64, 48, 24, 0, 1.0, (0.2, 0.2, 0.2)),          # a single Gaussian
])
def test_against_cpu_oracle(C_, P, W, H, seed, deg, mod, bg):
    from oracle import cpu_oracle
    scene, cam = synthetic.make_tiny_scene(P, W, H, seed=seed)
    if P == 1:
        scene["means3D"][:] = torch.tensor([0.05, -0.02, 0.3])
    bg_t = torch.tensor(bg)
    o = cpu_oracle.forward(util.scene_np(scene), cam, np.array(bg, np.float32), deg, mod)
    args, out = util.native_forward(C_, to_cuda(scene), cam, bg_t, deg, mod)
    R, color, radii, *_ = out
    m = util.bgr_intermediates(out, P, W, H)
    assert R == o["num_rendered"]
    assert np.array_equal(radii.cpu().numpy(), o["radii"])
    assert np.array_equal(m["tiles_touched"].cpu().numpy().view(np.uint32), o["tiles_touched"])
    assert np.array_equal(m["point_list"].cpu().numpy().view(np.uint32), o["point_list"])
    assert np.array_equal(m["ranges"].cpu().numpy().view(np.uint32), o["ranges"])
    nc = m["n_contrib"].cpu().numpy().view(np.uint32)
    # the CPU oracle does not fuse multiply-adds like the GPU: a pixel whose alpha or T sits
    # within an ulp of a threshold may differ. Allow 1e-4 of the pixels; the bit-exact bar is
    # carried by the reference fixtures and the reference build.
    assert (nc != o["n_contrib"]).mean() <= 1e-4
    # Image: 1e-5 where the splats are well conditioned. With scale_modifier 1.7 the elongated
    # splats have det = a*c - b*b cancelling by ~1e3, which amplifies the oracle's un-fused
    # rounding of a, b, c into ~1e-4 relative conic differences (the CUDA path is bit-identical to
    # the reference build there, see test_gpu_full.py) -> looser bound for that case only.
    img_tol = 1e-5 if mod == 1.0 else 1e-3
    assert np.abs(color.cpu().numpy() - o["color"]).max() <= img_tol + 1e-2 * (nc != o["n_contrib"]).any()

    dpix = synthetic.make_upstream_grad(H, W, seed=seed) * 1e3
    grads = util.native_backward(C_, args, out, dpix)
    og = cpu_oracle.backward(o, dpix.numpy())
    for k, t in zip(GRAD_NAMES, grads):
        ok, msg = grad_close(t.cpu().numpy(), og[k].reshape(t.shape), scale=1e3, max_bad_frac=1e-2,
                             max_norm_err=5e-3)
        assert ok, f"{k}: {msg}"
