"""Single-GPU micro-benchmark + check of the view-parallel dL_dsh rebuild kernel
(bgr_sh_grad_from_factors): `world` factor slots in local memory stand in for the multicast
buffer. Not a pytest; run on a GPU box:  python tests/micro_sh_rebuild.py [P] [world]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "gaussian-splatting-own_b200"):
    sys.path.insert(0, str(p))
from diff_gaussian_rasterization import _C  # noqa: E402


def reference(means, campos, factors, D):
    """fp64 restatement of sum_v basis(dir_v) x factor_v (reference backward.cu:21-137 per view)."""
    C0, C1 = 0.28209479177387814, 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
          -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
    P = means.shape[0]
    out = torch.zeros(P, 16, 3, dtype=torch.float64, device=means.device)
    for v in range(campos.shape[0]):
        d = means.double() - campos[v, :3].double()
        d = d / d.norm(dim=1, keepdim=True)
        x, y, z = d[:, 0], d[:, 1], d[:, 2]
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b = [torch.full_like(x, C0), -C1 * y, C1 * z, -C1 * x, C2[0] * xy, C2[1] * yz,
             C2[2] * (2 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy), C3[0] * y * (3 * xx - yy),
             C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy), C3[3] * z * (2 * zz - 3 * xx - 3 * yy),
             C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy), C3[6] * x * (xx - 3 * yy)]
        B = torch.stack(b, 1)[:, :(D + 1) ** 2]
        out[:, :B.shape[1]] += B[:, :, None] * factors[v].double()[:, None, :]
    return out


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
    world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(0)
    means = (torch.rand(P, 3, generator=g) * 10 - 5).to(dev)
    campos = torch.zeros(world, 4)
    campos[:, :3] = torch.randn(world, 3, generator=g) * 8
    campos = campos.to(dev)
    factors = torch.randn(world, P, 3, generator=g)
    factors *= (torch.rand(world, P, 1, generator=g) < 0.37)     # visibility of the bench scene
    factors = factors.to(dev)
    dsh = torch.empty(P, 16, 3, device=dev)
    for D in (3, 1):
        dsh.fill_(float("nan"))
        _C.sh_grad_from_factors(P, D, 16, world, means, campos.data_ptr(), factors.data_ptr(), dsh, dev)
        n = min(P, 200_000)
        want = reference(means[:n], campos, factors[:, :n], D)
        err = (dsh[:n].double() - want).abs().max().item()
        tail = reference(means[-1000:], campos, factors[:, -1000:], D)
        err = max(err, (dsh[-1000:].double() - tail).abs().max().item())
        print(f"D={D} max abs err vs fp64 {err:.3e} finite {bool(torch.isfinite(dsh).all())}")
        assert err < 1e-4 and torch.isfinite(dsh).all()
    # range variant must leave rows outside the range alone
    dsh.fill_(7.0)
    _C.sh_grad_from_factors_range(P, 1024, P - 1028, 3, 16, world, means, campos.data_ptr(),
                                  factors.data_ptr(), dsh, dev)
    assert float(dsh[:1024].min()) == 7.0 and float(dsh[P - 1028:].min()) == 7.0 and float(dsh[P - 1028:].max()) == 7.0
    assert not bool((dsh[1024:P - 1028] == 7.0).all())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        _C.sh_grad_from_factors(P, 3, 16, world, means, campos.data_ptr(), factors.data_ptr(), dsh, dev)
    e0.record()
    for _ in range(20):
        _C.sh_grad_from_factors(P, 3, 16, world, means, campos.data_ptr(), factors.data_ptr(), dsh, dev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    gb = (world * P * 12 + P * 12 + P * 192) / 1e9
    print(f"sh_grad_from_factors P={P} world={world}: {ms:.4f} ms, {gb / ms * 1e3:.0f} GB/s algorithmic")


if __name__ == "__main__":
    main()
