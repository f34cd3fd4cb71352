"""Multi-GPU check of the view-parallel exchange (run under torchrun on >= 2 B200s; driven by
tests/test_gpu_dp.py, which skips below two devices). Every rank renders its own camera. The leaf
gradients summed by libbgr's exchange (SH-gradient factor multicast fused into the backward kernel
+ in-switch sum, pipelined over Gaussian ranges; or the dense in-switch sum) must

  * equal the float64 sum of the per-rank gradients at the path's own bar (1e-4 relative / 1e-6
    absolute per element), except on the fraction of elements where two executions of the SAME
    backward on one rank already disagree (unordered float atomics; measured here, per tensor), and
  * be bit-identical on all ranks (checked through the peer mapping).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/dp_check.py \
        [P] [factor|dense] [chunks]
"""
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "gaussian-splatting-own_b200", ROOT / "tests"):
    sys.path.insert(0, str(p))
import synthetic  # noqa: E402
import view_parallel as vp  # noqa: E402
from diff_gaussian_rasterization import _C  # noqa: E402, F401


def outside_bar(a, want64):
    err = (a.double() - want64).abs()
    tol = 1e-6 + 1e-4 * want64.abs()
    return (err > tol).double().mean().item(), err.max().item()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
    mode = sys.argv[2] if len(sys.argv) > 2 else "factor"
    chunks = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    W, H, f = synthetic.RESOLUTIONS["800x800"]
    scene = {k: v.to(dev) for k, v in synthetic.make_gaussians(P).items()}
    cam = synthetic.make_camera(W, H, f, view_seed=rank)
    dpix = synthetic.make_upstream_grad(H, W, seed=rank + 1).to(dev)
    bg = torch.zeros(3, device=dev)
    ok = True
    t0 = time.time()

    # this rank's own gradients, twice (private buffers, no exchange): the float64 sum over ranks
    # of the first run is the expectation, the second run measures the run-to-run spread
    solo = vp.ViewParallelRasterizer(exchange="none")
    g1 = [g.clone() for g in solo.step(scene, cam, bg, lambda c: dpix).grads]
    g2 = [g.clone() for g in solo.step(scene, cam, bg, lambda c: dpix).grads]
    want, spread = {}, {}
    for name, i in vp._GRAD_INDEX.items():
        w = g1[i].double()
        dist.all_reduce(w)
        want[name] = w
        s = torch.tensor([outside_bar(g2[i], g1[i].double())[0]], device=dev, dtype=torch.float64)
        dist.all_reduce(s)                       # sum of the ranks' spreads bounds the sum's spread
        spread[name] = float(s)
    torch.cuda.synchronize()

    ex = vp.NvlsExchange(P, 16, dev, mode=mode, chunks=chunks)
    r = vp.ViewParallelRasterizer(exchange="nvls")
    r._nvls = ex
    peer = (rank + 1) % world
    for it in range(3):   # several times: the arena and the factor slots are reused
        res = r.step(scene, cam, bg, lambda c: dpix)
        torch.cuda.synchronize()
        # bit-identical replicas: compare my arena with the peer's through its mapped buffer
        mine = ex.buf[12 * P:ex.numel]
        theirs = ex.hdl.get_buffer(peer, (ex.numel,), torch.float32, 0)[12 * P:]
        ident = torch.equal(mine, theirs)
        ok = ok and ident
        for name, i in vp._GRAD_INDEX.items():
            got = res.grads[i]
            frac, worst = outside_bar(got, want[name])
            mag = want[name].abs().max().item()
            allowed = max(3.0 * spread[name], 2e-5)
            good = frac <= allowed and worst <= 1e-3 * mag and bool(torch.isfinite(got).all())
            ok = ok and good
            sys.stdout.write(
                f"[rank {rank}/{world} {time.time() - t0:5.1f}s] mode={ex.mode} chunks={ex.chunks} it={it} {name}: "
                f"outside 1e-4/1e-6 bar {frac:.2e} (same-rank run-to-run {spread[name]:.2e}, allowed {allowed:.2e}) "
                f"max err {worst:.3e} of |g|max {mag:.3e} identical_to_peer {ident} {'ok' if good else 'FAIL'}\n")
            sys.stdout.flush()
        ex.hdl.barrier(channel=0)   # nobody overwrites its arena while a peer still compares
        torch.cuda.synchronize()
    # one write per line: eight ranks share the launcher's pipe
    sys.stdout.write(f"DP_CHECK rank {rank} world {world} {'PASS' if ok else 'FAIL'}\n")
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
