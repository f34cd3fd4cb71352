"""Multi-GPU check of the view-parallel exchange (run under torchrun on >= 2 B200s; not a pytest —
the round-end GPU suite has one GPU). Every rank renders its own camera; the leaf gradients summed
by libbgr's exchange kernels (factor multicast / row-sparse / dense) must equal the NCCL all-reduce of the same
per-rank gradients, and be bit-identical on all ranks.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/dp_check.py
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
from diff_gaussian_rasterization import _C  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
    W, H, f = synthetic.RESOLUTIONS["800x800"]
    scene = {k: v.to(dev) for k, v in synthetic.make_gaussians(P).items()}
    cam = synthetic.make_camera(W, H, f, view_seed=rank)
    dpix = synthetic.make_upstream_grad(H, W, seed=rank + 1).to(dev)
    bg = torch.zeros(3, device=dev)
    ok = True
    t0 = time.time()

    # reference result: plain NCCL on private gradients
    r0 = vp.ViewParallelRasterizer(exchange="nccl")
    base = r0.step(scene, cam, bg, lambda c: dpix)
    want = [g.clone() for g in base.grads]
    torch.cuda.synchronize()

    mode = sys.argv[2] if len(sys.argv) > 2 else "factor"
    chunks = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    ex = vp.NvlsExchange(P, 16, dev, mode=mode, chunks=chunks)
    r = vp.ViewParallelRasterizer(exchange="nvls")
    r._nvls = ex
    peer = (rank + 1) % world
    for sparse in ((mode,) if mode == "factor" else (False, True)):
        ex.sparse = sparse
        for it in range(2):   # twice: the arena is reused
            res = r.step(scene, cam, bg, lambda c: dpix)
            torch.cuda.synchronize()
            # bit-identical replicas: compare my arena with the peer's through its mapped buffer
            mine = ex.buf[12 * P:ex.numel]
            theirs = ex.hdl.get_buffer(peer, (ex.numel,), torch.float32, 0)[12 * P:]
            ident = torch.equal(mine, theirs)
            ok = ok and ident
            for name, i in vp._GRAD_INDEX.items():
                got, exp = res.grads[i], want[i]
                # two executions of the backward differ at the atomics' rounding level, so this is
                # a tolerance comparison; the exactness claim is "identical across ranks".
                err = (got.double() - exp.double()).abs()
                tol = 5e-4 * exp.abs().max().item() + 1e-3 * exp.double().abs()   # atomics-order noise of two runs
                bad = int((err > tol).sum())
                ok = ok and bad == 0
                print(f"[rank {rank} {time.time() - t0:5.1f}s] sparse={sparse} it={it} {name}: max err "
                      f"{err.max().item():.3e} bad {bad} identical_to_peer {ident}", flush=True)
            ex.hdl.barrier(channel=0)   # nobody overwrites its arena while a peer still compares
            torch.cuda.synchronize()
    print(f"DP_CHECK rank {rank}", "PASS" if ok else "FAIL", flush=True)
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
