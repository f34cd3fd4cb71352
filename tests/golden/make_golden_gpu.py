"""Runs the REFERENCE's own CUDA rasterizer (oracle/_ref, unmodified reference .cu files built for
sm_100a) on small seeded scenes on a B200 and stores inputs-by-seed + outputs as fixtures:

    tests/golden/ref_small_*.npz

Run on the GPU box:  gpurun -- 'python tests/golden/make_golden_gpu.py gpurun_out/golden'
then copy gpurun_out/golden/*.npz into tests/golden/ and commit. The not-gpu tests pin the CPU
oracle (oracle/rasterizer_oracle.c) against these files.
"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "gaussian-splatting-own_b200"))
sys.path.insert(0, str(ROOT / "tests"))
import synthetic  # noqa: E402
from oracle import ref_backend  # noqa: E402
import util  # noqa: E402

CASES = {
    # name: (P, W, H, seed, sh_degree, scale_modifier, bg)
    "a": (600, 96, 64, 11, 3, 1.0, (0.0, 0.0, 0.0)),
    "b": (1500, 112, 80, 12, 2, 1.3, (0.3, 0.6, 0.1)),
    "c": (300, 50, 37, 13, 0, 0.7, (1.0, 1.0, 1.0)),   # ragged image size, degree 0
}


def main(outdir):
    outdir = Path(outdir); outdir.mkdir(parents=True, exist_ok=True)
    for name, (P, W, H, seed, deg, mod, bg) in CASES.items():
        scene, cam = synthetic.make_tiny_scene(P, W, H, seed=seed)
        bg_t = torch.tensor(bg, dtype=torch.float32)
        args, out = util.native_forward(ref_backend, scene, cam, bg_t, sh_degree=deg, scale_modifier=mod)
        R, color, radii, geom, binning, img = out
        dpix = synthetic.make_upstream_grad(H, W, seed=seed + 100) * 1e3
        grads = util.native_backward(ref_backend, args, out, dpix)
        g = ref_backend.parse_geom(geom, P); b = ref_backend.parse_binning(binning, R); i = ref_backend.parse_img(img, W, H)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        vis = (radii > 0)
        rec = dict(P=P, W=W, H=H, seed=seed, sh_degree=deg, scale_modifier=mod, bg=np.array(bg, np.float32),
                   num_rendered=R, color=color.cpu().numpy(), radii=radii.cpu().numpy(),
                   tiles_touched=g["tiles_touched"].cpu().numpy(), depths=(g["depths"] * vis).cpu().numpy(),
                   means2D=(g["means2D"] * vis[:, None]).cpu().numpy(),
                   conic_opacity=(g["conic_opacity"] * vis[:, None]).cpu().numpy(),
                   rgb=(g["rgb"] * vis[:, None]).cpu().numpy(), cov3D=g["cov3D"].cpu().numpy(),
                   clamped=(g["clamped"] * vis[:, None]).cpu().numpy(),
                   point_list=b["point_list"].cpu().numpy(), point_list_keys=b["point_list_keys"].cpu().numpy(),
                   ranges=i["ranges"][:tiles].cpu().numpy(), final_T=i["final_T"].cpu().numpy(),
                   n_contrib=i["n_contrib"].cpu().numpy(), dL_dpix=dpix.numpy())
        for k, t in zip(("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh",
                         "dL_dscales", "dL_drotations"), grads):
            rec[k] = t.cpu().numpy()
        rec["dL_dconic"] = ref_backend.rasterize_gaussians_backward.last_dL_dconic.cpu().numpy()
        np.savez_compressed(outdir / f"ref_small_{name}.npz", **rec)
        print(name, "R", R, "visible", int(vis.sum()), "mean n_contrib", float(i["n_contrib"].float().mean()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
