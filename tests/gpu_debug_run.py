"""Developer script (not a pytest): runs every stage of libbgr against the reference CUDA build and
torch primitives on a GPU box and prints a detailed report. Used through gpurun while developing.

    python tests/gpu_debug_run.py [small|mid|full ...]
"""
import sys
import time
import traceback
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "gaussian-splatting-own_b200"))
sys.path.insert(0, str(ROOT / "tests"))
import synthetic  # noqa: E402
import util  # noqa: E402
from diff_gaussian_rasterization import _C  # noqa: E402
from oracle import ref_backend  # noqa: E402


def section(name):
    print("\n" + "=" * 20, name, "=" * 20, flush=True)


def check_scan():
    section("scan")
    for n in (1, 7, 2048, 2049, 100_000, 3_000_001):
        x = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda")
        y = _C.inclusive_scan_u32(x)
        ref = torch.cumsum(x.long(), 0).int()
        print(n, "ok" if torch.equal(y, ref) else f"MISMATCH first at {int((y != ref).nonzero()[0])}")


def check_sort():
    section("radix sort")
    g = torch.Generator(device="cuda").manual_seed(0)
    cases = ((1, 0, 8, False), (1000, 0, 64, False), (4096, 3, 20, True), (4097, 22, 61, False),
             (300_000, 0, 45, True), (3_500_000, 22, 61, False), (3_500_000, 0, 39, True),
             (5_000_000, 10, 10, False))
    for n, lo, hi, pairs in cases:
        for ties in (False, True):
            keys = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
            if ties and hi - lo > 6:   # only 64 distinct values in the sorted field -> many ties
                keys = keys & ~(((1 << (hi - lo - 6)) - 1) << lo)
            vals = torch.arange(n, dtype=torch.int32, device="cuda") if pairs else None
            ks, vs = _C.radix_sort(keys, vals, lo, hi)
            if hi - lo == 64:
                field = keys ^ (-2**63)  # unsigned order
            else:
                field = (keys >> lo) & ((1 << (hi - lo)) - 1)
            order = torch.sort(field, stable=True).indices if hi > lo else torch.arange(n, device="cuda")
            ok = torch.equal(ks, keys[order]) and (vs is None or torch.equal(vs, vals[order]))
            print(f"n={n} bits=[{lo},{hi}) pairs={pairs} ties={ties}: {'ok' if ok else 'MISMATCH'}")
            if not ok:
                bad = (ks != keys[order]).nonzero()
                print("   first bad index", int(bad[0]) if len(bad) else None, "count", len(bad))


def compare(scene, cam, bg, deg, mod, label, time_it=False, dpix=None):
    section(label)
    P = scene["means3D"].shape[0]
    W, H = cam.image_width, cam.image_height
    a_ref, o_ref = util.native_forward(ref_backend, scene, cam, bg, deg, mod)
    a_new, o_new = util.native_forward(_C, scene, cam, bg, deg, mod)
    torch.cuda.synchronize()
    Rr, Rn = o_ref[0], o_new[0]
    print("P", P, "R ref/new", Rr, Rn)
    g = ref_backend.parse_geom(o_ref[3], P)
    im = ref_backend.parse_img(o_ref[5], W, H)
    b = ref_backend.parse_binning(o_ref[4], Rr)
    mine = util.bgr_intermediates(o_new, P, W, H)
    vis = o_ref[2] > 0
    print("visible", int(vis.sum()), "mean tiles/visible", Rr / max(int(vis.sum()), 1))
    print("radii equal:", torch.equal(o_ref[2], o_new[2]), "mismatches", int((o_ref[2] != o_new[2]).sum()))
    print("tiles_touched equal:", torch.equal(g["tiles_touched"], mine["tiles_touched"]),
          int((g["tiles_touched"] != mine["tiles_touched"]).sum()))
    for name in ("depths", "means2D", "conic_opacity", "rgb"):
        r = g[name][vis]; m = mine[name][vis]
        neq = (r.view(torch.int32) != m.contiguous().view(torch.int32))
        print(f"{name}: bit-mismatches {int(neq.sum())} / {neq.numel()}  max abs diff {float((r - m).abs().max()) if r.numel() else 0:.3e}")
        if int(neq.sum()) and neq.dim() == 2:
            print("    per-column mismatches:", neq.sum(0).tolist())
    rc = g["clamped"][vis].bool()
    mc = torch.stack([(mine["clampbits"][vis] >> k) & 1 for k in range(3)], 1).bool()
    print("clamped equal:", torch.equal(rc, mc))
    if Rr == Rn:
        print("point_list equal:", torch.equal(b["point_list"], mine["point_list"]),
              int((b["point_list"] != mine["point_list"]).sum()))
    tiles = mine["ranges"].shape[0]
    print("ranges equal:", torch.equal(im["ranges"][:tiles], mine["ranges"]))
    print("n_contrib equal:", torch.equal(im["n_contrib"], mine["n_contrib"]),
          int((im["n_contrib"] != mine["n_contrib"]).sum()), "sum", int(im["n_contrib"].long().sum()))
    print("final_T max abs diff", float((im["final_T"] - mine["final_T"]).abs().max()))
    print("image max abs diff", float((o_ref[1] - o_new[1]).abs().max()))

    if dpix is None:
        dpix = synthetic.make_upstream_grad(H, W)
    g_ref = util.native_backward(ref_backend, a_ref, o_ref, dpix)
    g_new = util.native_backward(_C, a_new, o_new, dpix)
    g_ref2 = util.native_backward(ref_backend, a_ref, o_ref, dpix)
    torch.cuda.synchronize()
    for n_, r, m in zip(("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"), g_ref, g_ref2):
        err = (r.double() - m.double()).abs(); tol = 1e-6 + 1e-4 * r.double().abs()
        print(f"  ref-vs-ref {n_}: max err {err.max().item() if r.numel() else 0:.3e} viol {int((err > tol).sum())} max err/tol {(err / tol).max().item() if r.numel() else 0:.2f}")
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations")
    for n_, r, m in zip(names, g_ref, g_new):
        r64, m64 = r.double(), m.double()
        err = (r64 - m64).abs()
        scale = r64.abs().max().item() if r.numel() else 0.0
        tol = 1e-6 + 1e-4 * r64.abs()
        bad = (err > tol)
        # tolerance relative to the per-tensor magnitude as a second view
        print(f"{n_}: max|ref| {scale:.3e} max err {err.max().item() if r.numel() else 0:.3e} "
              f"viol(1e-4rel+1e-6abs) {int(bad.sum())}/{bad.numel()} max err/tol {(err / tol).max().item() if r.numel() else 0:.2f} "
              f"nan {int(torch.isnan(m).sum())}")

    if time_it:
        def t(fn, n=10):
            for _ in range(3): fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            for _ in range(n): fn()
            e1.record(); torch.cuda.synchronize()
            return e0.elapsed_time(e1) / n
        print("forward ms  ref %.3f  new %.3f" % (t(lambda: util.native_forward(ref_backend, scene_d, cam, bg, deg, mod)),
                                                  t(lambda: util.native_forward(_C, scene_d, cam, bg, deg, mod))))
        print("backward ms ref %.3f  new %.3f" % (t(lambda: util.native_backward(ref_backend, a_ref, o_ref, dpix_d)),
                                                  t(lambda: util.native_backward(_C, a_new, o_new, dpix_d))))


scene_d = None
dpix_d = None


def run_case(P, res, label, deg=3, mod=1.0, bg=(0.0, 0.0, 0.0), time_it=False, tiny=None):
    global scene_d, dpix_d
    try:
        if tiny:
            scene, cam = synthetic.make_tiny_scene(P, *tiny, seed=5)
        else:
            W, H, f = synthetic.RESOLUTIONS[res]
            cam = synthetic.make_camera(W, H, f)
            scene = synthetic.make_gaussians(P)
        scene_d = {k: v.cuda() for k, v in scene.items()}
        dpix_d = synthetic.make_upstream_grad(cam.image_height, cam.image_width).cuda()
        compare(scene_d, cam, torch.tensor(bg), deg, mod, label, time_it, dpix_d)
    except Exception:
        traceback.print_exc()


if __name__ == "__main__":
    what = sys.argv[1:] or ["prims", "small"]
    print(torch.cuda.get_device_name(0))
    if "prims" in what:
        try: check_scan()
        except Exception: traceback.print_exc()
        try: check_sort()
        except Exception: traceback.print_exc()
    if "small" in what:
        run_case(500, None, "tiny 500 @ 96x64", tiny=(96, 64))
        run_case(3000, None, "tiny 3000 @ 130x70 bg", tiny=(130, 70), bg=(0.3, 0.7, 0.2), deg=2, mod=1.2)
    if "mid" in what:
        run_case(100_000, "800x800", "c1: 100K @ 800x800", time_it=True)
    if "full" in what:
        run_case(3_000_000, "1080p", "c2: 3M @ 1080p", time_it=True)
