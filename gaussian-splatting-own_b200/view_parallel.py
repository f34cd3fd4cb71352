"""View-batch data parallelism around the rasterizer.

The reference is single-GPU and renders one view per step (train.py:76-86). The path shards
naturally over views: every rank holds the full, replicated Gaussian set, renders a different
camera, and the per-Gaussian gradients of the trainer's leaves are summed across ranks before the
optimizer — ONE exchange step, no other data-path collective.

What is specific to this implementation:

* the backward kernels write the five leaf gradients (means3D 3, opacity 1, scales 3, rotations 4,
  sh 3M floats per Gaussian = 59 at M = 16) into one contiguous arena
  (diff_gaussian_rasterization/_C.py: rasterize_gaussians_backward), so the exchange runs IN PLACE
  on the memory the kernels produced — no pack / unpack copies;
* exchange "nccl": one NCCL all-reduce of the arena (708 MB at 3M Gaussians) behind the backward,
  `async_op` optional. The baseline, and what other backends (reference build, stubs) get;
* exchange "nvls" (`NvlsExchange`, libbgr's own kernels over NVLink multicast): the backward
  multicasts the 3-float SH-gradient factor per Gaussian while it computes, every rank rebuilds
  dL_dsh locally, the other 11 floats are summed in the switch, and the whole exchange is
  pipelined under the backward preprocess in Gaussian ranges (see the class docstring; DESIGN.md
  §6 has the measurements: 0.21 / 0.23 / 0.33 ms per step at 2 / 4 / 8 GPUs instead of 1.3–1.9 ms).

`torch.distributed` (NCCL on GPUs, gloo in the CPU tests) and `torch.distributed._symmetric_memory`
(allocation, rendezvous, barrier) are plumbing only.
"""
from __future__ import annotations

from typing import Callable, NamedTuple, Optional, Sequence

import torch
import torch.distributed as dist

LEAF_ORDER = ("means3D", "opacities", "scales", "rotations", "shs")   # memory order in the arena
# index of each leaf's gradient in the 8-tuple returned by rasterize_gaussians_backward
_GRAD_INDEX = {"means3D": 3, "shs": 5, "opacities": 2, "scales": 6, "rotations": 7}


def chunk_bounds(P: int, chunks: int, align: int = 1024) -> list:
    """[0, b1, ..., P]: at most `chunks` consecutive Gaussian ranges of equal size rounded up to
    `align` (the multicast path needs bounds that are multiples of 4; 1024 keeps the ranges whole
    blocks of the backward preprocess). The last range takes the remainder."""
    if P <= 0:
        return [0, 0]
    chunks = max(1, int(chunks))
    step = max(align, -(-(-(-P // chunks)) // align) * align)
    return list(range(0, P, step)) + [P]


def view_for_rank(step: int, world: int, rank: int, num_views: int) -> int:
    """Camera index rendered by `rank` at `step`: views are dealt round-robin, `world` per step,
    wrapping around the view list (the DP analogue of train.py:76-79's one random view per step)."""
    if num_views <= 0:
        raise ValueError("no views")
    return (step * world + rank) % num_views


def flat_leaf_gradients(grads: Sequence[torch.Tensor]) -> tuple[torch.Tensor, bool]:
    """Returns (flat, aliased). If the five leaf gradients already sit back to back in one
    allocation (the layout libbgr's backward produces) `flat` is a zero-copy view over them and
    `aliased` is True; otherwise they are packed into a new buffer (reference backend, stubs)."""
    parts = [grads[_GRAD_INDEX[k]] for k in LEAF_ORDER]
    base = parts[0]
    contiguous = all(p.is_contiguous() for p in parts)
    if contiguous and base._base is not None:
        ptr = base.data_ptr()
        ok = True
        for p in parts:
            ok = ok and p.data_ptr() == ptr and p._base is base._base
            ptr += p.numel() * p.element_size()
        if ok:
            root = base._base
            start = (base.data_ptr() - root.data_ptr()) // root.element_size()
            n = sum(p.numel() for p in parts)
            return root.view(-1)[start:start + n], True
    return torch.cat([p.reshape(-1) for p in parts]), False


def scatter_flat(flat: torch.Tensor, grads: Sequence[torch.Tensor]) -> None:
    """Inverse of the packing branch of flat_leaf_gradients."""
    o = 0
    for k in LEAF_ORDER:
        g = grads[_GRAD_INDEX[k]]
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()


class NvlsExchange:
    """The gradient arena in symmetric (peer-mapped, multicast-bound) memory and the exchange over
    it, done by libbgr's own kernels (csrc/nvls_allreduce.cu, csrc/preprocess_bwd.cu).
    Construction is collective; keep ONE instance per process for a given (P, M).

    mode "factor" (default): dL/dsh[k] = basis_k(dir) * dL/dRGB is rank one per view, so the
        backward kernel does not write the 3M floats of dL/dsh at all: it multicasts the 3 floats
        of dL/dRGB per Gaussian (multimem.st, replicated by the NVSwitch into every GPU's copy of
        slot [rank]) while it computes — the all-gather is fused into the producing kernel. After
        one barrier every rank rebuilds sum_v basis(dir_v) x dL/dRGB_v locally (deterministic,
        identical everywhere) and only the remaining 11 floats/Gaussian are summed in the switch.
        Fabric traffic at 3M Gaussians, 8 GPUs: 36 MB out + 252 MB in (overlapped with the
        backward) + a 132 MB in-switch sum, instead of a 708 MB all-reduce.
    mode "dense": the whole 59-float block in the switch (fallback when P % 4 != 0 or M == 0)."""

    NON_LEAF_FLOATS = 12   # means2D 3 + colors 3 + cov3D 6 precede the leaf block in the arena
    SMALL_FLOATS = 11      # ... + rotations 4: everything except sh

    def __init__(self, P: int, M: int, device, group: Optional[dist.ProcessGroup] = None,
                 max_blocks: Optional[int] = None, mode: str = "factor", chunks: int = 2):
        import torch.distributed._symmetric_memory as symm_mem
        from diff_gaussian_rasterization import _C
        self._C = _C
        self.P, self.M, self.device = P, M, torch.device(device)
        self.leaf_floats = self.SMALL_FLOATS + 3 * M
        self.numel = P * (self.NON_LEAF_FLOATS + self.leaf_floats)
        self._max_blocks_arg = max_blocks
        grp = group if group is not None else dist.group.WORLD
        world = dist.get_world_size(grp)
        up4 = lambda n: (n + 3) // 4 * 4  # noqa: E731
        self.campos_off = up4(self.numel)
        self.factors_off = self.campos_off + 4 * world
        aligned = P % 4 == 0
        if mode not in ("factor", "dense"):
            raise ValueError("mode must be 'factor' or 'dense'")
        if mode == "factor" and not (aligned and M > 0):
            mode = "dense"
        self.mode = mode
        # CTAs of the in-switch sum. Mode "factor" runs it BESIDE the HBM-bound dL_dsh rebuild and
        # the next range's preprocess: its multimem requests are slow (fabric round trips) and, with
        # a CTA on every SM, starve the other kernels' loads (measured: rebuild 0.14 -> 0.48 ms).
        # 32 CTAs keep the fabric as busy and leave 116 SMs alone. Alone (dense) it gets 296.
        self.max_blocks = self._max_blocks_arg if self._max_blocks_arg is not None else \
            (32 if mode == "factor" else 296)
        total = self.factors_off + (world * P * 3 if mode == "factor" else 0)
        self.buf = symm_mem.empty(total, dtype=torch.float32, device=self.device)
        self.hdl = symm_mem.rendezvous(self.buf, grp)
        if not self.hdl.multicast_ptr:
            raise RuntimeError("NVLink multicast (NVLS) is not available for this group")
        self.rank, self.world = self.hdl.rank, self.hdl.world_size
        self._pending = None     # (D, means3D, dL_dsh) of the backward whose factors are in flight
        self._side = self._side2 = None
        # mode "factor": the backward preprocess runs in `chunks` Gaussian ranges (bounds are
        # multiples of 1024) and each finished range is exchanged on a second stream — barrier,
        # local dL_dsh rebuild, in-switch sum of its 11 floats/Gaussian — under the next range.
        self.chunks = max(1, int(chunks))
        self.bounds = chunk_bounds(P, self.chunks)
        self._piped = False
        self.profile = False     # diagnostics: CUDA events around the pieces of all_reduce()
        self._marks = []
        self._side_marks = []
        self._pipe_marks = []    # profile, pipelined mode: per range (ready, barrier, sum, rebuild0, rebuild1)
        self._pipe_calls = []    # ... per call: (marks of its ranges, end event)
        self._active = 0

    # The hooks that redirect the backward's outputs into the symmetric arena are installed only
    # while this exchange is entered (`with exchange: backward(...)`): a backward issued outside
    # the block — another rasterizer of the same size, gradient accumulation over several views —
    # gets private buffers and a complete dL_dsh as usual.
    def __enter__(self):
        if self._active == 0:
            self._C.set_grad_arena_provider(self._provide)
            if self.mode == "factor":
                self._C.set_sh_factor_hook(self._factor_slots)
        self._active += 1
        return self

    def __exit__(self, *exc):
        self._active -= 1
        if self._active == 0:
            self._C.set_grad_arena_provider(None)
            self._C.set_sh_factor_hook(None)
        return False

    # ---- hooks called by _C.rasterize_gaussians_backward -------------------------------------
    def _provide(self, numel, dev):
        return self.buf[:self.numel] if (numel == self.numel and torch.device(dev) == self.buf.device) else None

    def _factor_slots(self, P, M, degree, means3D, dL_dsh):
        if P != self.P or M != self.M or means3D.device != self.buf.device:
            return None
        self._pending = (degree, means3D, dL_dsh)
        mc = self.hdl.multicast_ptr
        slots = (mc + 4 * (self.factors_off + self.rank * 3 * P), mc + 4 * (self.campos_off + 4 * self.rank))
        if self.chunks <= 1 or len(self.bounds) <= 2:
            return slots
        self._piped = True
        return slots + ((self.bounds, self._after_chunk),)

    def _after_chunk(self, p0: int, p1: int) -> None:
        """Called right after the backward preprocess of Gaussians [p0, p1) was queued on the current
        stream: exchange that range on the side stream while the next range is computed."""
        degree, means3D, dL_dsh = self._pending
        P, n = self.P, p1 - p0
        cur = torch.cuda.current_stream(self.device)
        self._make_streams()
        ev = (lambda: torch.cuda.Event(enable_timing=True)) if self.profile else (lambda: None)
        e_ready, e_bar, e_sum, e_rb0, e_rb1 = ev(), ev(), ev(), ev(), ev()
        if self.profile:
            e_ready.record(cur)
        self._side.wait_stream(cur)
        leaf0 = P * self.NON_LEAF_FLOATS
        base = self.buf.data_ptr()
        with torch.cuda.stream(self._side):
            self.hdl.barrier(channel=0)   # every rank has produced (and multicast) this range
            ready = torch.cuda.Event()
            ready.record()
            if self.profile:
                e_bar.record()
            # the fabric-bound kernel first (one CTA per SM, high-priority stream): the HBM-bound
            # rebuild and the next range's preprocess fill the rest of the machine around it
            segs = [(leaf0 + 3 * p0, 3 * n), (leaf0 + 3 * P + p0, n), (leaf0 + 4 * P + 3 * p0, 3 * n),
                    (leaf0 + 7 * P + 4 * p0, 4 * n)]   # means3D | opacity | scales | rotations rows
            self._C.nvls_allreduce_segments(self.hdl.multicast_ptr, segs, self.rank, self.world,
                                            self.device, self.max_blocks)
            if self.profile:
                e_sum.record()
        self._side2.wait_event(ready)
        with torch.cuda.stream(self._side2):
            if self.profile:
                e_rb0.record()
            self._C.sh_grad_from_factors_range(P, p0, p1, degree, self.M, self.world, means3D,
                                               base + 4 * self.campos_off, base + 4 * self.factors_off,
                                               dL_dsh, self.device)
            if self.profile:
                e_rb1.record()
                self._pipe_marks.append((e_ready, e_bar, e_sum, e_rb0, e_rb1))

    def _make_streams(self):
        if self._side is None:
            self._side = torch.cuda.Stream(self.device, priority=-1)   # in-switch sums
            self._side2 = torch.cuda.Stream(self.device)                # local dL_dsh rebuild

    # ---- the exchange ------------------------------------------------------------------------
    def all_reduce(self) -> None:
        """Sum of the leaf block across ranks, in place, on the current stream."""
        P, mc = self.P, self.hdl.multicast_ptr
        leaf0 = P * self.NON_LEAF_FLOATS
        red = lambda off, cnt: self._C.nvls_allreduce(mc, off, cnt, self.rank, self.world, self.device,  # noqa: E731
                                                      self.max_blocks)
        if self._piped:
            # every range was exchanged on the side stream as it was produced
            cur = torch.cuda.current_stream(self.device)
            cur.wait_stream(self._side)
            cur.wait_stream(self._side2)
            self._piped = False
            self._pending = None
            e_join = None
            if self.profile:
                e_join = torch.cuda.Event(enable_timing=True)
                e_join.record()
            self.hdl.barrier(channel=0)    # sums broadcast everywhere; factor slots free for reuse
            if self.profile and self._pipe_marks:
                e_end = torch.cuda.Event(enable_timing=True)
                e_end.record()
                self._pipe_calls.append((self._pipe_marks, e_join, e_end))
                self._pipe_marks = []
            return
        mark = self._mark
        mark("start")
        self.hdl.barrier(channel=0)        # every rank's backward (stores, multicasts) is complete
        mark("barrier_in")
        if self.mode == "factor" and self._pending is not None:
            degree, means3D, dL_dsh = self._pending
            base = self.buf.data_ptr()
            # The local rebuild of dL_dsh (HBM-bound) runs beside the in-switch sum (NVLink-bound).
            # The sum is launched FIRST: its one CTA per SM is resident at once and the rebuild's
            # 23k CTAs fill in around it; the other way round the block scheduler hands out the
            # rebuild's CTAs first and the two kernels run back to back (measured).
            cur = torch.cuda.current_stream(self.device)
            self._make_streams()
            ready = torch.cuda.Event()
            ready.record()
            red(leaf0, P * self.SMALL_FLOATS)
            self._side2.wait_event(ready)
            with torch.cuda.stream(self._side2):
                rb0 = rb1 = None
                if self.profile:
                    rb0 = torch.cuda.Event(enable_timing=True)
                    rb0.record()
                self._C.sh_grad_from_factors(P, degree, self.M, self.world, means3D,
                                             base + 4 * self.campos_off, base + 4 * self.factors_off,
                                             dL_dsh, self.device)
                if self.profile:
                    rb1 = torch.cuda.Event(enable_timing=True)
                    rb1.record()
                    self._side_marks.append((self._marks[-1][1], rb0, rb1))
            mark("in_switch_sum")
            cur.wait_stream(self._side2)
            mark("sh_rebuild_tail")
            self._pending = None
        else:
            red(leaf0, P * self.leaf_floats)
        self.hdl.barrier(channel=0)        # sums broadcast everywhere; factor slots free for reuse
        mark("barrier_out")

    def _mark(self, name):
        if self.profile:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._marks.append((name, e))

    def profile_report(self):
        """Mean milliseconds per piece over the profiled calls (synchronises)."""
        torch.cuda.synchronize(self.device)
        if self._pipe_calls:
            # pipelined mode. Per Gaussian range j: wait_j = from "its backward preprocess is done on
            # the main stream" to "every rank passed the barrier" (includes waiting for the slowest
            # rank's range), sum_j = the in-switch sum, rebuild_j = the local dL_dsh rebuild (beside
            # the sum). tail = from the LAST range's preprocess being done to the end of the
            # exchange: the part of the exchange that nothing overlaps.
            acc, calls = {}, len(self._pipe_calls)
            for marks, e_join, e_end in self._pipe_calls:
                for j, (e_ready, e_bar, e_sum, e_rb0, e_rb1) in enumerate(marks):
                    for k, v in ((f"range{j}_barrier_wait", e_ready.elapsed_time(e_bar)),
                                 (f"range{j}_in_switch_sum", e_bar.elapsed_time(e_sum)),
                                 (f"range{j}_sh_rebuild", e_rb0.elapsed_time(e_rb1))):
                        acc[k] = acc.get(k, 0.0) + v
                last_ready = marks[-1][0]
                acc["tail_after_last_preprocess"] = acc.get("tail_after_last_preprocess", 0.0) + last_ready.elapsed_time(e_end)
                acc["barrier_out"] = acc.get("barrier_out", 0.0) + e_join.elapsed_time(e_end)
            self._pipe_calls = []
            return {k: round(v / calls, 4) for k, v in acc.items()}
        acc, n = {}, 0
        prev = None
        for name, e in self._marks:
            if name == "start":
                n += 1
            else:
                acc[name] = acc.get(name, 0.0) + prev.elapsed_time(e)
            prev = e
        self._marks = []
        out = {k: round(v / max(n, 1), 4) for k, v in acc.items()}
        if self._side_marks:   # the rebuild on its own stream, relative to the end of barrier_in
            m = len(self._side_marks)
            out["rebuild_start_after_barrier"] = round(sum(b.elapsed_time(r0) for b, r0, _ in self._side_marks) / m, 4)
            out["rebuild_duration"] = round(sum(r0.elapsed_time(r1) for _, r0, r1 in self._side_marks) / m, 4)
            self._side_marks = []
        return out

    def close(self):
        self._active = 0
        self._C.set_grad_arena_provider(None)
        self._C.set_sh_factor_hook(None)


class StepResult(NamedTuple):
    color: torch.Tensor
    radii: torch.Tensor
    num_rendered: int
    grads: tuple           # the backend's 8-tuple; leaf entries hold the SUM over ranks
    work: Optional[object]  # async handle of the all-reduce (None when finished or world == 1)


class ViewParallelRasterizer:
    """forward + backward of one view per rank and the gradient exchange.

    backend: module/object with the reference's `_C` call signatures (defaults to libbgr).
    loss_grad: callable(color) -> dL/dcolor (the trainer's loss backward; a stub in the bench).
    """

    def __init__(self, backend=None, group: Optional[dist.ProcessGroup] = None, average: bool = False,
                 exchange: str = "nccl"):
        if backend is None:
            from diff_gaussian_rasterization import _C as backend  # noqa: N811
        self.backend = backend
        self.group = group
        self.average = average
        if exchange not in ("nccl", "nvls", "none"):
            raise ValueError("exchange must be 'nccl', 'nvls' or 'none' (per-rank gradients, no sum)")
        self.exchange = exchange
        self._nvls: Optional[NvlsExchange] = None

    @property
    def world(self) -> int:
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def step(self, scene: dict, cam, bg: torch.Tensor, loss_grad: Callable[[torch.Tensor], torch.Tensor],
             sh_degree: int = 3, scale_modifier: float = 1.0, async_op: bool = False) -> StepResult:
        e = torch.Tensor([])
        dev = scene["means3D"].device
        view, proj, campos = (cam.viewmatrix.to(dev), cam.projmatrix.to(dev), cam.campos.to(dev))
        fwd = self.backend.rasterize_gaussians(
            bg, scene["means3D"], e, scene["opacities"], scene["scales"], scene["rotations"],
            scale_modifier, e, view, proj, cam.tanfovx, cam.tanfovy, cam.image_height,
            cam.image_width, scene["shs"], sh_degree, campos, False, False)
        R, color, radii, geom, binning, img = fwd
        if self.exchange == "nvls" and self.world > 1 and self._nvls is None:
            self._nvls = NvlsExchange(scene["means3D"].size(0), scene["shs"].size(1), dev, self.group)
        dL_dpix = loss_grad(color)
        if self.average and self.world > 1:
            dL_dpix = dL_dpix / self.world     # fold the 1/G of a mean into the upstream gradient
        import contextlib
        with (self._nvls if self._nvls is not None else contextlib.nullcontext()):
            grads = self.backend.rasterize_gaussians_backward(
                bg, scene["means3D"], radii, e, scene["scales"], scene["rotations"], scale_modifier, e,
                view, proj, cam.tanfovx, cam.tanfovy, dL_dpix, scene["shs"], sh_degree, campos, geom, R,
                binning, img, False)
        work = None
        if self.world > 1 and self._nvls is not None:
            self._nvls.all_reduce()
        elif self.world > 1 and self.exchange != "none":
            flat, aliased = flat_leaf_gradients(grads)
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
            if not aliased:
                if work is not None:
                    work.wait()
                    work = None
                scatter_flat(flat, grads)
        return StepResult(color, radii, int(R), tuple(grads), work)


def visibility_statistics(radii: torch.Tensor, dL_dmeans2D: torch.Tensor,
                          group: Optional[dist.ProcessGroup] = None) -> tuple[torch.Tensor, torch.Tensor]:
    """Densification statistics of the view batch (scene/gaussian_model.py:405-407 per view):
    sum over ranks of ||dL/dmean2D[:, :2]|| for visible Gaussians and of the visibility count."""
    vis = radii > 0
    norm = torch.where(vis, dL_dmeans2D[:, :2].norm(dim=-1), torch.zeros((), device=radii.device))
    stats = torch.stack([norm, vis.to(norm.dtype)])
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, group=group)
    return stats[0], stats[1]
