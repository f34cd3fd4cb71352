"""CPU suite for the N > 1 path: world_size-2 gloo runs of the view-parallel step with a stub
backend that has libbgr's call signatures and arena layout (the rasterizer itself is CUDA-only).
Checks the host logic: view dealing, zero-copy aliasing of the leaf gradients, the in-place sum,
the packed fallback used for the reference backend, averaging, async handles, statistics."""
import os
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "gaussian-splatting-own_b200"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

import synthetic  # noqa: E402
import view_parallel as vp  # noqa: E402


class StubBackend:
    """Deterministic stand-in: gradients depend on the camera and the upstream gradient, laid out
    either in one arena (libbgr) or as separate tensors (reference shim)."""

    def __init__(self, arena: bool):
        self.arena = arena

    def rasterize_gaussians(self, bg, means3D, colors, opacity, scales, rotations, mod, cov, view,
                            proj, tx, ty, H, W, sh, deg, campos, prefiltered, debug):
        P = means3D.size(0)
        color = torch.full((3, H, W), float(campos.sum())) + bg[:, None, None]
        radii = (means3D[:, 2] > 0).int()
        return P // 2, color, radii, torch.zeros(8, dtype=torch.uint8), torch.zeros(8, dtype=torch.uint8), \
            torch.zeros(8, dtype=torch.uint8)

    def rasterize_gaussians_backward(self, bg, means3D, radii, colors, scales, rotations, mod, cov,
                                     view, proj, tx, ty, dL, sh, deg, campos, geom, R, binning, img, debug):
        P, M = means3D.size(0), sh.size(1)
        widths = (3, 3, 6, 3, 1, 3, 4, 3 * M)
        s = float(dL.sum()) + float(campos[0])
        if self.arena:
            a = torch.empty(P * sum(widths))
            parts, o = [], 0
            for w in widths:
                parts.append(a[o:o + P * w]); o += P * w
        else:
            parts = [torch.empty(P * w) for w in widths]
        for k, p in enumerate(parts):
            p.copy_(torch.arange(p.numel(), dtype=torch.float32) * 1e-3 + s * (k + 1))
        m2, col, cov3, m3, op, sc, rot, dsh = parts
        return (m2.view(P, 3), col.view(P, 3), op.view(P, 1), m3.view(P, 3), cov3.view(P, 6),
                dsh.view(P, M, 3), sc.view(P, 3), rot.view(P, 4))


def _worker(rank, world, port, arena, average, async_op, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene = synthetic.make_gaussians(50, seed=0)
        cams = [synthetic.make_camera(32, 16, 20.0, view_seed=v) for v in range(5)]
        cam = cams[vp.view_for_rank(3, world, rank, len(cams))]
        r = vp.ViewParallelRasterizer(StubBackend(arena), average=average)
        res = r.step(scene, cam, torch.zeros(3), lambda c: torch.ones_like(c) * (rank + 1), async_op=async_op)
        if res.work is not None:
            res.work.wait()
        stats = vp.visibility_statistics(res.radii, res.grads[0])
        out[rank] = ([g.clone() for g in res.grads], [s.clone() for s in stats])
    finally:
        dist.destroy_process_group()


def _run(arena, average=False, async_op=False, world=2):
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + (os.getpid() + hash((arena, average, async_op))) % 2000
    mp.spawn(_worker, args=(world, port, arena, average, async_op, out), nprocs=world, join=True)
    return out


def _single(rank, world, arena, average):
    """What each rank would have produced alone (no process group)."""
    scene = synthetic.make_gaussians(50, seed=0)
    cams = [synthetic.make_camera(32, 16, 20.0, view_seed=v) for v in range(5)]
    cam = cams[vp.view_for_rank(3, world, rank, len(cams))]
    r = vp.ViewParallelRasterizer(StubBackend(arena))
    scale = 1.0 / world if average else 1.0
    return r.step(scene, cam, torch.zeros(3), lambda c: torch.ones_like(c) * (rank + 1) * scale).grads


@pytest.mark.parametrize("arena,average,async_op", [(True, False, False), (False, False, False),
                                                    (True, True, True)])
def test_two_ranks_sum_leaf_gradients(arena, average, async_op):
    out = _run(arena, average, async_op)
    solo = [_single(r, 2, arena, average) for r in range(2)]
    leaf = set(vp._GRAD_INDEX.values())
    for rank in range(2):
        grads, _ = out[rank]
        for i, g in enumerate(grads):
            if i in leaf:      # summed over ranks, identical on both
                torch.testing.assert_close(g, solo[0][i] + solo[1][i], rtol=1e-6, atol=1e-6)
            else:              # view-local quantities are left alone
                torch.testing.assert_close(g, solo[rank][i])
    # replicas stay bit-identical after the exchange
    for i in leaf:
        assert torch.equal(out[0][0][i], out[1][0][i])
    # statistics: both ranks see the batch totals
    assert torch.equal(out[0][1][1], out[1][1][1]) and float(out[0][1][1].max()) == 2.0


def test_flat_view_is_zero_copy_for_the_arena_layout():
    g = StubBackend(True).rasterize_gaussians_backward(
        None, torch.zeros(7, 3), None, None, None, None, 1.0, None, None, None, 0, 0, torch.ones(1),
        torch.zeros(7, 16, 3), 3, torch.zeros(3), None, 0, None, None, False)
    flat, aliased = vp.flat_leaf_gradients(g)
    assert aliased and flat.numel() == 7 * 59 and flat.data_ptr() == g[3].data_ptr()
    flat.mul_(2.0)   # in place on the arena == in place on every leaf gradient
    assert float(g[5][-1, -1, -1]) == pytest.approx(float(flat[-1]))
    g2 = StubBackend(False).rasterize_gaussians_backward(
        None, torch.zeros(7, 3), None, None, None, None, 1.0, None, None, None, 0, 0, torch.ones(1),
        torch.zeros(7, 16, 3), 3, torch.zeros(3), None, 0, None, None, False)
    flat2, aliased2 = vp.flat_leaf_gradients(g2)
    assert not aliased2 and flat2.numel() == 7 * 59
    flat2.mul_(2.0)
    vp.scatter_flat(flat2, g2)
    torch.testing.assert_close(torch.cat([g2[i].reshape(-1) for i in (3, 2, 6, 7, 5)]), flat2)


def test_real_backend_arena_layout_matches():
    """The slicing done by _C.rasterize_gaussians_backward (checked here without a GPU by
    re-deriving its widths) must keep the five leaves adjacent in LEAF_ORDER."""
    src = (ROOT / "gaussian-splatting-own_b200" / "diff_gaussian_rasterization" / "_C.py").read_text()
    assert "widths = (3, 3, 6, 3, 1, 3, 4, 3 * M)" in src
    assert vp.LEAF_ORDER == ("means3D", "opacities", "scales", "rotations", "shs")


def test_view_dealing_covers_every_view_once_per_epoch():
    world, views = 4, 12
    seen = [vp.view_for_rank(s, world, r, views) for s in range(views // world) for r in range(world)]
    assert sorted(seen) == list(range(views))
    assert vp.view_for_rank(5, 8, 3, 7) == (5 * 8 + 3) % 7
    with pytest.raises(ValueError):
        vp.view_for_rank(0, 2, 0, 0)


def test_chunk_bounds_for_the_pipelined_exchange():
    for P, chunks in [(3_000_000, 2), (3_000_000, 4), (400_000, 2), (5000, 4), (1024, 4), (4, 2), (1_000_004, 3)]:
        b = vp.chunk_bounds(P, chunks)
        assert b[0] == 0 and b[-1] == P and len(b) - 1 <= chunks
        assert all(x < y for x, y in zip(b[:-1], b[1:]))
        assert all(x % 1024 == 0 for x in b[:-1])          # interior bounds: whole preprocess blocks
        sizes = [y - x for x, y in zip(b[:-1], b[1:])]
        assert max(sizes) - min(sizes[:-1] or sizes) <= 0 or len(sizes) == 1 or sizes[:-1].count(sizes[0]) == len(sizes) - 1
    assert vp.chunk_bounds(0, 4) == [0, 0]
    assert vp.chunk_bounds(3_000_000, 1) == [0, 3_000_000]
    assert vp.chunk_bounds(3_000_000, 2) == [0, 1_500_160, 3_000_000]
