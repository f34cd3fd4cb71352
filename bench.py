#!/usr/bin/env python
"""Benchmark of the rasterizer hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|reference-ext]

One "step" = forward + backward of ONE 1920x1080 view over 3M synthetic SH-degree-3 Gaussians
(BASELINE.json configs[2]; seeded scene of SURVEY.md §8(d)). With N > 1 ranks (torchrun) every
rank renders its own camera over the replicated Gaussians and the 59 floats/Gaussian of leaf
gradients are summed with ONE NCCL all-reduce issued in place on the buffer the backward kernels
wrote (weak scaling: per-GPU work fixed).

  value : views/s, inputs resident in HBM, native calls (C ABI), CUDA-event timed, max over ranks
  e2e   : views/s through the public API (GaussianRasterizer + autograd) including per step the
          H2D copy of the camera block and the (3,H,W) loss weights from pinned host memory and the
          D2H read of the loss
  --impl reference : the UNMODIFIED reference CUDA kernels (oracle/_ref/libref_rasterizer.so,
          compiled from /root/reference by oracle/Makefile) behind the same Python glue and the
          same harness. The reference has no CPU rasterizer, so the comparison
          column named by north_star is its CUDA extension on the same B200; the CPU column is
          `cpu_baseline` (the plain-C oracle port, one view of the same workload).
  --impl reference-ext : the reference's own torch extension (its setup.py sources compiled
          unmodified by oracle/build_ref_ext.py) - the stock pybind path, to show it agrees with
          the shim arm above.

Between timed iterations nothing is flushed explicitly: the inputs (708 MB of Gaussians) are
far larger than the 126 MB L2, which is stated in config.l2.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
for p in (ROOT, ROOT / "gaussian-splatting-own_b200", ROOT / "tests"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

import numpy as np  # noqa: E402
import torch  # noqa: E402

GRAD_FLOATS = 59  # means3D 3 + sh 48 + opacity 1 + scales 3 + rotations 4


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=("ours", "reference", "reference-ext"), default="ours")
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--res", choices=("800x800", "1080p", "4k"), default="1080p")
    ap.add_argument("--mode", choices=("step", "sweep", "f1", "train"), default="step",
                    help="step: the train-step bench (default). sweep: BASELINE.json configs[4], the "
                         "forward-only 4K grid over Gaussians x SH degree x scale_modifier, both arms "
                         "per cell, written to --out. f1: the trainer's activation glue + drop-in path "
                         "against the raw-leaf entry with the activations fused (SURVEY.md 8 row f1). "
                         "train: one whole training iteration (activations, rasterizer, L1 + D-SSIM loss, Adam) "
                         "composed as the reference's trainer does vs the fused entries (rows f1 + f2)")
    ap.add_argument("--sweep-gaussians", default="1,2,5,10", help="millions of Gaussians per cell")
    ap.add_argument("--sweep-degrees", default="0,3")
    ap.add_argument("--sweep-mods", default="0.5,1,2")
    ap.add_argument("--out", default="gpurun_out/sweep_4k.json")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--chunks", type=int, default=2,
                    help="N > 1, --exchange nvls: Gaussian ranges the backward preprocess is pipelined "
                         "against the exchange in (1 = exchange after the backward)")
    ap.add_argument("--nvls-blocks", type=int, default=0,
                    help="CTAs of the in-switch sum kernel (0 = the exchange mode's default)")
    ap.add_argument("--exchange", choices=("nvls", "nvls-dense", "nccl"), default="nvls",
                    help="N > 1, ours: in-switch multimem kernel of libbgr (default) or ncclAllReduce")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region (B200_PROFILING.md recipe)
# --------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in open(self.path).read().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "power_w_max": float(max(power)), "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------
# workload
# --------------------------------------------------------------------------------------------
def load_backend(impl):
    if impl == "ours":
        from diff_gaussian_rasterization import _C
        _C.lib()  # fail loudly if libbgr.so is missing: there is no fallback
        return _C
    if impl == "reference-ext":
        # the reference's own torch extension (its setup.py sources, pybind module), built by
        # oracle/build_ref_ext.py: same three functions, same signatures
        from oracle import build_ref_ext
        return build_ref_ext.load()
    from oracle import ref_backend
    if not ref_backend.available():
        return None
    ref_backend.lib()
    return ref_backend


class Workload:
    def __init__(self, args, rank, device):
        import synthetic
        self.W, self.H, focal = synthetic.RESOLUTIONS[args.res]
        self.P = args.gaussians
        self.cam = synthetic.make_camera(self.W, self.H, focal, view_seed=rank)
        self.scene_cpu = synthetic.make_gaussians(self.P, seed=0)
        self.scene = {k: v.to(device) for k, v in self.scene_cpu.items()}
        self.bg = torch.zeros(3, device=device)
        self.dpix_cpu = synthetic.make_upstream_grad(self.H, self.W, seed=1)
        self.dpix = self.dpix_cpu.to(device)
        self.device = device
        c = self.cam
        self.view, self.proj, self.campos = (c.viewmatrix.to(device), c.projmatrix.to(device),
                                             c.campos.to(device))
        self.empty = torch.Tensor([])

    def fwd_args(self):
        s, c, e = self.scene, self.cam, self.empty
        return (self.bg, s["means3D"], e, s["opacities"], s["scales"], s["rotations"], 1.0, e,
                self.view, self.proj, c.tanfovx, c.tanfovy, self.H, self.W, s["shs"], 3, self.campos,
                False, False)

    def bwd_args(self, out):
        s, c, e = self.scene, self.cam, self.empty
        R, _color, radii, geom, binning, img = out
        return (self.bg, s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, self.view,
                self.proj, c.tanfovx, c.tanfovy, self.dpix, s["shs"], 3, self.campos, geom, R,
                binning, img, False)


def make_native_step(backend, wl, world, impl, exchange="nccl", chunks=2, nvls_blocks=0):
    """forward + backward through the backend's `_C`-signature calls; N > 1 adds the all-reduce."""
    import torch.distributed as dist
    import view_parallel as vp
    nvls = None
    make_native_step.fallback = None
    if world > 1 and impl == "ours" and exchange.startswith("nvls"):
        try:
            nvls = vp.NvlsExchange(wl.P, 16, wl.device,
                                   mode={"nvls": "factor", "nvls-dense": "dense"}[exchange],
                                   chunks=chunks, max_blocks=nvls_blocks or None)
            ok = torch.ones(1, device=wl.device)
        except Exception as e:  # no NVLink multicast / symmetric memory on this box
            make_native_step.fallback = f"{type(e).__name__}: {e}"[:200]
            ok = torch.zeros(1, device=wl.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)      # every rank takes the same path
        if float(ok) == 0.0:
            if nvls is not None:
                nvls.close()
            nvls = None
            make_native_step.fallback = make_native_step.fallback or "a peer rank could not set up NVLS"
            print(f"[bench] NVLS exchange unavailable ({make_native_step.fallback}); "
                  "using the NCCL all-reduce of the gradient arena", file=sys.stderr)
    make_native_step.nvls = nvls

    import contextlib

    def step():
        out = backend.rasterize_gaussians(*wl.fwd_args())
        with (nvls if nvls is not None else contextlib.nullcontext()):   # outputs -> symmetric arena
            grads = backend.rasterize_gaussians_backward(*wl.bwd_args(out))
        if world > 1:
            if nvls is not None:
                nvls.all_reduce()                                     # in-switch sum, own kernel
            elif impl == "ours":
                flat, aliased = vp.flat_leaf_gradients(grads)         # zero-copy view of the arena
                assert aliased
                dist.all_reduce(flat)                                 # in place, one collective
            else:                                                     # packed: one collective too
                flat, _ = vp.flat_leaf_gradients(grads)
                dist.all_reduce(flat)
                vp.scatter_flat(flat, grads)
        return out, grads
    return step


def make_e2e_step(backend, wl, world, impl):
    """The call a user makes (gaussian_renderer/__init__.py:85-93) + loss stub, with the per-step
    host inputs (camera block, loss weights) copied from pinned memory and the loss read back."""
    import torch.distributed as dist
    import diff_gaussian_rasterization as dgr
    if impl == "ours":
        Rasterizer = dgr.GaussianRasterizer
    else:
        fn = dgr.make_rasterize_function(backend)

        class Rasterizer(dgr.GaussianRasterizer):
            _rasterize = staticmethod(lambda *a: fn.apply(*a))
            _backend = backend
    dev = wl.device
    c = wl.cam
    cam_host = torch.cat([c.viewmatrix.reshape(-1), c.projmatrix.reshape(-1), c.campos.reshape(-1),
                          torch.zeros(3)]).pin_memory()
    w_host = wl.dpix_cpu.pin_memory()
    cam_dev = torch.empty_like(cam_host, device=dev)
    w_dev = torch.empty_like(w_host, device=dev)
    loss_host = torch.zeros((), dtype=torch.float32).pin_memory()
    leaves = {k: v.clone().requires_grad_(True) for k, v in wl.scene.items()}
    h2d = cam_host.numel() * 4 + w_host.numel() * 4
    # The loss weights are first needed after the forward: their copy runs on a second stream
    # under the forward kernels (both arms use this harness). The step ends with a
    # synchronize, so the next step's copy cannot overtake this step's reads.
    copy_stream = torch.cuda.Stream(device=dev)
    w_ready = torch.cuda.Event()

    nvls = getattr(make_native_step, "nvls", None) if world > 1 else None

    def step():
        cam_dev.copy_(cam_host, non_blocking=True)
        # queued first: with eight ranks sharing the host's memory system the 24.9 MB copy takes
        # longer than alone, and it needs the whole forward to hide under
        with torch.cuda.stream(copy_stream):
            w_dev.copy_(w_host, non_blocking=True)
            w_ready.record(copy_stream)
        for t in leaves.values():
            t.grad = None
        settings = dgr.GaussianRasterizationSettings(
            image_height=wl.H, image_width=wl.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy,
            bg=cam_dev[35:38], scale_modifier=1.0, viewmatrix=cam_dev[0:16].view(4, 4),
            projmatrix=cam_dev[16:32].view(4, 4), sh_degree=3, campos=cam_dev[32:35],
            prefiltered=False, debug=False)
        means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, radii_ = Rasterizer(settings)(
            means3D=leaves["means3D"], means2D=means2D, shs=leaves["shs"], colors_precomp=None,
            opacities=leaves["opacities"], scales=leaves["scales"], rotations=leaves["rotations"],
            cov3D_precomp=None)
        torch.cuda.current_stream().wait_event(w_ready)
        loss = (color * w_dev).sum()
        if nvls is not None:
            with nvls:                 # the backward writes into the symmetric arena
                loss.backward()
            g0 = leaves["means3D"].grad
            if g0.data_ptr() != nvls.buf.data_ptr() + 4 * 12 * wl.P:
                # autograd copied instead of handing out views of the arena: the in-place exchange
                # would not reach the leaves' .grad. Never time a numerically wrong step.
                raise RuntimeError("e2e: leaf gradients do not alias the exchange arena")
            nvls.all_reduce()          # in place, in the switch
        else:
            loss.backward()
            if world > 1:              # one collective over the flattened leaf gradients
                flat = torch.cat([t.grad.reshape(-1) for t in leaves.values()])
                dist.all_reduce(flat)
                o = 0
                for t in leaves.values():
                    t.grad.copy_(flat[o:o + t.grad.numel()].view_as(t.grad))
                    o += t.grad.numel()
        loss_host.copy_(loss.detach(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(loss_host)
    return step, h2d, 4


def timed(step, steps, warmup, world, device):
    """W warm-up steps, then exactly K steps between barrier + synchronize on both sides; CUDA
    events on the launching (current) stream; max over ranks."""
    import torch.distributed as dist
    for _ in range(warmup):
        step()
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    ms = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


# --------------------------------------------------------------------------------------------
# roofline (SURVEY.md §8(d) algorithmic bytes / flops per unit; DESIGN.md "Measurement")
# --------------------------------------------------------------------------------------------
def peaks():
    path = ROOT / "MEASURED_PEAKS.json"
    if path.exists():
        d = json.loads(path.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)", 1965.0


NCU_RAW = "profiles/r2_g_ncu_full_raw.csv"
_STAGE_KERNEL = {"render_backward": "render_backward_kernel", "render_forward": "render_forward_kernel",
                 "preprocess_backward": "preprocess_backward_kernel", "preprocess": "preprocess_kernel",
                 "scatter_keys": "scatter_keys_kernel", "tile_sort": "tile_sort_kernel"}


def ncu_traffic(stage):
    """DRAM bytes (read + write) of one launch of the stage's kernel, from the committed
    `ncu --set full` capture of this same command; None when the capture is not there."""
    import csv
    path = ROOT / NCU_RAW
    name = _STAGE_KERNEL.get(stage)
    if name is None or not path.exists():
        return None
    try:
        rows = list(csv.reader(open(path)))
        h, units = rows[0], rows[1]
        k, r, w = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for row in rows[2:]:
            if name in row[k]:
                return int(float(row[r]) * scale[units[r]] + float(row[w]) * scale[units[w]])
    except (ValueError, KeyError, IndexError):
        return None
    return None


def roofline_report(backend, wl, steps, device):
    """Per-stage durations from CUDA events recorded on the launch stream inside libbgr (mean over
    `steps` extra steps run right after the timed region, same process and inputs), combined with
    the algorithmic bytes / flops of each stage."""
    P, W, H = wl.P, wl.W, wl.H
    backend.stage_timing(True)
    acc, n = {}, 0
    out = None
    for _ in range(steps):
        out = backend.rasterize_gaussians(*wl.fwd_args())
        backend.rasterize_gaussians_backward(*wl.bwd_args(out))
        torch.cuda.synchronize(device)
        for k, v in backend.stage_times().items():
            acc[k] = acc.get(k, 0.0) + v
        n += 1
    backend.stage_timing(False)
    ms = {k: v / n for k, v in acc.items()}
    R = int(out[0])
    V = int((out[2] > 0).sum().item())
    C = P - V
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    st = backend.render_stats(out, P, W, H)
    il = backend.image_layout(W, H)
    n_contrib = out[5][il["n_contrib"]:il["n_contrib"] + 4 * W * H].view(torch.int32)
    n_pair = int(n_contrib.long().sum().item())
    hbm_peak, peak_src, sm_mhz = peaks()
    n_sm = torch.cuda.get_device_properties(device).multi_processor_count
    fp32_peak = n_sm * 128 * 2 * sm_mhz * 1e6 / 1e12  # TFLOP/s at max SM clock
    bytes_ = {
        "preprocess": 20 * P + 291 * V,
        "tile_scan": 4 * (tiles + 2 * int(tiles ** 0.5) + 1) + 12 * tiles,   # coverage grid in, ranges + cursors out
        "scatter_keys": 20 * V + 8 * R,                                      # rect inputs in, one 8 B key per instance out
        "tile_sort": 12 * R,                                                 # keys in, sorted Gaussian ids out
        "clear_accumulators": 48 * P,
        "preprocess_backward": 555 * V + 284 * C,
    }
    flops = {
        # 25 flop per pair whose exponential is evaluated (11 for pairs rejected at the power
        # test) + 6 per blended pair; backward: the same re-evaluation + 75 per blended pair
        "render_forward": 11 * (st["examined"] - st["exp_evaluated"]) + 25 * st["exp_evaluated"] + 6 * st["blended"],
        "render_backward": 11 * (st["examined"] - st["exp_evaluated"]) + 25 * st["exp_evaluated"] + 75 * st["blended"],
    }
    stages = []
    for name, t in ms.items():
        if name in bytes_:
            a = bytes_[name] / (t * 1e-3) / 1e9
            stages.append({"kernel": name, "ms": round(t, 4), "bound": "hbm", "achieved": round(a, 1),
                           "peak": hbm_peak, "unit": "GB/s", "frac": round(a / hbm_peak, 4),
                           "algorithmic_bytes": bytes_[name]})
        else:
            a = flops[name] / (t * 1e-3) / 1e12
            stages.append({"kernel": name, "ms": round(t, 4), "bound": "fp32", "achieved": round(a, 2),
                           "peak": round(fp32_peak, 1), "unit": "TFLOP/s", "frac": round(a / fp32_peak, 4),
                           "algorithmic_flops": flops[name]})
    top = max(stages, key=lambda s: s["ms"])
    roof = dict(top)
    roof["traffic"] = ncu_traffic(top["kernel"]) if (P, W, H) == (3_000_000, 1920, 1080) else None
    if roof["traffic"] is not None:
        roof["traffic_source"] = f"{NCU_RAW} (dram__bytes_read.sum + dram__bytes_write.sum, one launch, same workload)"
    roof["peak_source"] = peak_src if top["bound"] == "hbm" else \
        f"{n_sm} SMs x 128 FMA lanes x 2 x {sm_mhz:.0f} MHz (no tensor-core work on this path)"
    # the dominant HBM-bound stage, so that the line always carries one bandwidth roofline too
    hbm = [s for s in stages if s["bound"] == "hbm"]
    workload_stats = {"P": P, "visible": V, "num_rendered": R, "n_pair_sum_n_contrib": n_pair,
                      "mean_tile_list": round(R / tiles, 1), **st}
    hbm_top = dict(max(hbm, key=lambda s: s["ms"])) if hbm else None
    if hbm_top is not None and (P, W, H) == (3_000_000, 1920, 1080):
        hbm_top["traffic"] = ncu_traffic(hbm_top["kernel"])
    return roof, stages, hbm_top, workload_stats, peak_src


def dp_breakdown(backend, wl, impl, world, device, ms_per_step):
    """N > 1 diagnostic, outside the timed region: every rank times its own view WITHOUT the
    exchange (private gradient buffers, plain backward kernel). The step time is bounded below by
    the slowest view; what is left is the exchange (barriers, multicast, in-switch sum, rebuild)."""
    import torch.distributed as dist
    if impl == "ours":
        backend.set_grad_arena_provider(None)
        backend.set_sh_factor_hook(None)
    n = 20
    for _ in range(3):
        out = backend.rasterize_gaussians(*wl.fwd_args())
        backend.rasterize_gaussians_backward(*wl.bwd_args(out))
    torch.cuda.synchronize(device)
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = backend.rasterize_gaussians(*wl.fwd_args())
        backend.rasterize_gaussians_backward(*wl.bwd_args(out))
    e1.record()
    torch.cuda.synchronize(device)
    mine = torch.tensor([e0.elapsed_time(e1) / n], device=device)
    every = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    local = [round(float(t), 4) for t in every]
    return {"local_ms_per_rank": local, "slowest_view_ms": max(local),
            "exchange_ms": round(ms_per_step - max(local), 4),
            "note": "local = this rank's view, forward + backward without any exchange"}


def cpu_baseline(wl):
    """One view (forward + backward) of the SAME workload on the host cores with the plain-C
    oracle port (the reference has no CPU rasterizer)."""
    from oracle import cpu_oracle
    sc = {k: v.numpy() for k, v in wl.scene_cpu.items()}
    t0 = time.perf_counter()
    o = cpu_oracle.forward(sc, wl.cam, np.zeros(3, np.float32), 3, 1.0)
    t1 = time.perf_counter()
    cpu_oracle.backward(o, wl.dpix_cpu.numpy())
    t2 = time.perf_counter()
    return {"value": round(1.0 / (t2 - t0), 5), "unit": "views/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"1 view fwd+bwd of the full workload ({wl.P} Gaussians, {wl.W}x{wl.H}); "
                      f"forward {t1 - t0:.2f} s, backward {t2 - t1:.2f} s (OpenMP over Gaussians / "
                      f"tiles on all cores; the per-Gaussian gradient sums are a serial pass)"}


# --------------------------------------------------------------------------------------------
# f1: the step in front of the path (activations + SH cat) composed in torch vs fused in the kernels
# --------------------------------------------------------------------------------------------
def run_f1(args, device):
    """One view of the bench scene through the public autograd API, three ways, CUDA-event timed:
    composed = the trainer's glue (scene/gaussian_model.py:95-115 via gaussian_renderer/__init__.py:
    53-82: sigmoid, exp, normalize, cat) + GaussianRasterizer, gradients to the raw leaves by torch
    autograd; fused = RawGaussianRasterizer (bgr_forward_raw / bgr_backward_raw); glue_only = the
    activations and their backward without any rasterizer between them."""
    import torch.nn.functional as F
    import diff_gaussian_rasterization as dgr
    load_backend("ours")
    wl = Workload(args, 0, device)
    s, c = wl.scene, wl.cam
    o = s["opacities"].clamp(1e-4, 1 - 1e-4)
    raw = {"xyz": s["means3D"], "features_dc": s["shs"][:, :1, :].contiguous(),
           "features_rest": s["shs"][:, 1:, :].contiguous(), "opacity": torch.log(o / (1 - o)),
           "scaling": torch.log(s["scales"]), "rotation": s["rotations"] * 1.25}
    leaves = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    settings = dgr.GaussianRasterizationSettings(
        image_height=wl.H, image_width=wl.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=wl.bg,
        scale_modifier=1.0, viewmatrix=wl.view, projmatrix=wl.proj, sh_degree=3, campos=wl.campos,
        prefiltered=False, debug=False)
    drop_in, fused_mod = dgr.GaussianRasterizer(settings), dgr.RawGaussianRasterizer(settings)

    def clear():
        for t in leaves.values():
            t.grad = None

    def activate():
        L = leaves
        return (torch.cat((L["features_dc"], L["features_rest"]), dim=1), torch.sigmoid(L["opacity"]),
                torch.exp(L["scaling"]), F.normalize(L["rotation"]))

    def composed():
        clear()
        shs, opac, scales, rots = activate()
        m2d = torch.zeros_like(leaves["xyz"], requires_grad=True)
        img, _ = drop_in(means3D=leaves["xyz"], means2D=m2d, opacities=opac, shs=shs, scales=scales,
                         rotations=rots)
        img.backward(wl.dpix)

    def fused():
        clear()
        L = leaves
        m2d = torch.zeros_like(L["xyz"], requires_grad=True)
        img, _ = fused_mod(L["xyz"], m2d, L["features_dc"], L["features_rest"], L["opacity"], L["scaling"],
                           L["rotation"])
        img.backward(wl.dpix)

    g_act = None

    def glue_only():
        nonlocal g_act
        clear()
        acts = activate()
        if g_act is None:
            g_act = [torch.ones_like(a) for a in acts]
        torch.autograd.backward(acts, g_act)

    out = {"mode": "f1", "metric": "ms per view (forward + backward to the raw leaves)",
           "config": {"workload": f"{args.gaussians} Gaussians SH3, {args.res}, one view, inputs resident"},
           "steps": args.steps, "warmup": args.warmup}
    for name, fn in (("composed_ms", composed), ("fused_ms", fused), ("glue_only_ms", glue_only)):
        out[name] = round(timed(fn, args.steps, args.warmup, 1, device) / args.steps, 4)
    out["saving_ms"] = round(out["composed_ms"] - out["fused_ms"], 4)
    out["speedup"] = round(out["composed_ms"] / out["fused_ms"], 3)
    print(json.dumps(out))


# --------------------------------------------------------------------------------------------
# f1 + f2: a whole training iteration, composed as the reference's trainer does vs fused
# --------------------------------------------------------------------------------------------
def run_train(args, device):
    """One training iteration of train.py:79-128 without densification, CUDA-event timed, inputs
    resident: activations (scene/gaussian_model.py:95-115) -> rasterizer -> loss = 0.8 L1 + 0.2
    (1 - SSIM) (utils/loss_utils.py) -> backward -> Adam over the six groups (gaussian_model.py:
    149-160) -> zero_grad. Three ways: the reference's rasterizer with torch loss and
    torch.optim.Adam ("reference"), this library's drop-in rasterizer in the same composition
    ("drop_in"), and the fused entries RawGaussianRasterizer + l1_ssim_loss + FusedAdam ("fused")."""
    import math
    import torch.nn.functional as F
    import diff_gaussian_rasterization as dgr
    import train_step
    load_backend("ours")
    ref = load_backend("reference")
    wl = Workload(args, 0, device)
    s, c = wl.scene, wl.cam
    o = s["opacities"].clamp(1e-4, 1 - 1e-4)
    raw0 = {"xyz": s["means3D"], "f_dc": s["shs"][:, :1, :].contiguous(), "f_rest": s["shs"][:, 1:, :].contiguous(),
            "opacity": torch.log(o / (1 - o)), "scaling": torch.log(s["scales"]), "rotation": s["rotations"] * 1.25}
    lrs = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005,
           "rotation": 0.001}
    gt = (torch.rand(3, wl.H, wl.W, generator=torch.Generator().manual_seed(3))).to(device)
    settings = dgr.GaussianRasterizationSettings(
        image_height=wl.H, image_width=wl.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=wl.bg,
        scale_modifier=1.0, viewmatrix=wl.view, projmatrix=wl.proj, sh_degree=3, campos=wl.campos,
        prefiltered=False, debug=False)
    g1 = torch.tensor([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g1 = (g1 / g1.sum()).unsqueeze(1)
    window = g1.mm(g1.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().to(device)

    def torch_loss(img):          # utils/loss_utils.py:16-63, train.py:91-92
        conv = lambda t: F.conv2d(t[None], window, padding=5, groups=3)[0]  # noqa: E731
        mu1, mu2 = conv(img), conv(gt)
        s1, s2, s12 = conv(img * img) - mu1 * mu1, conv(gt * gt) - mu2 * mu2, conv(img * gt) - mu1 * mu2
        ssim = (((2 * mu1 * mu2 + 0.0001) * (2 * s12 + 0.0009)) /
                ((mu1 * mu1 + mu2 * mu2 + 0.0001) * (s1 + s2 + 0.0009))).mean()
        return 0.8 * (img - gt).abs().mean() + 0.2 * (1.0 - ssim)

    def make(kind):
        P_ = {k: torch.nn.Parameter(v.clone()) for k, v in raw0.items()}
        groups = [{"params": [P_[k]], "lr": lrs[k], "name": k} for k in raw0]
        if kind == "fused":
            opt = train_step.FusedAdam(groups, lr=0.0, eps=1e-15)
            rast = dgr.RawGaussianRasterizer(settings)

            def step():
                m2d = torch.zeros_like(P_["xyz"], requires_grad=True)
                img, _ = rast(P_["xyz"], m2d, P_["f_dc"], P_["f_rest"], P_["opacity"], P_["scaling"], P_["rotation"])
                train_step.l1_ssim_loss(img, gt, 0.2).backward()
                opt.step()
                opt.zero_grad(set_to_none=True)
            return step
        opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        if kind == "reference":
            fn = dgr.make_rasterize_function(ref)

            class Rast(dgr.GaussianRasterizer):
                _rasterize = staticmethod(lambda *a: fn.apply(*a))
                _backend = ref
        else:
            Rast = dgr.GaussianRasterizer
        rast = Rast(settings)

        def step():
            m2d = torch.zeros_like(P_["xyz"], requires_grad=True)
            img, _ = rast(means3D=P_["xyz"], means2D=m2d, opacities=torch.sigmoid(P_["opacity"]),
                          shs=torch.cat((P_["f_dc"], P_["f_rest"]), dim=1), scales=torch.exp(P_["scaling"]),
                          rotations=F.normalize(P_["rotation"]))
            torch_loss(img).backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
        return step

    out = {"mode": "train", "metric": "ms per training iteration (activations + rasterizer + L1/D-SSIM loss + Adam)",
           "config": {"workload": f"{args.gaussians} Gaussians SH3, {args.res}, one view per iteration, inputs resident"},
           "steps": args.steps, "warmup": args.warmup}
    kinds = (["reference"] if ref is not None else []) + ["drop_in", "fused"]
    for kind in kinds:
        step = make(kind)
        out[f"{kind}_ms"] = round(timed(step, args.steps, args.warmup, 1, device) / args.steps, 4)
        del step
        torch.cuda.empty_cache()
    if "reference_ms" in out:
        out["speedup_vs_reference"] = round(out["reference_ms"] / out["fused_ms"], 3)
    out["speedup_vs_drop_in"] = round(out["drop_in_ms"] / out["fused_ms"], 3)

    # the two f2 pieces on their own: loss forward + backward on a fixed image, one optimizer step
    img0 = (gt + 0.1 * torch.randn_like(gt)).clamp(0, 1)

    def loss_fused():
        x = img0.detach().requires_grad_(True)
        train_step.l1_ssim_loss(x, gt, 0.2).backward()

    def loss_torch():
        x = img0.detach().requires_grad_(True)
        torch_loss(x).backward()

    def adam_pair(fused):
        P_ = {k: torch.nn.Parameter(v.clone()) for k, v in raw0.items()}
        for t in P_.values():
            t.grad = torch.randn_like(t) * 1e-3
        groups = [{"params": [P_[k]], "lr": lrs[k], "name": k} for k in raw0]
        opt = train_step.FusedAdam(groups, lr=0.0, eps=1e-15) if fused else torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        return opt.step
    pieces = {}
    for name, fn in (("loss_fused_ms", loss_fused), ("loss_torch_ms", loss_torch),
                     ("adam_fused_ms", adam_pair(True)), ("adam_torch_ms", adam_pair(False))):
        pieces[name] = round(timed(fn, args.steps, args.warmup, 1, device) / args.steps, 4)
        torch.cuda.empty_cache()
    out["pieces"] = pieces
    print(json.dumps(out))


# --------------------------------------------------------------------------------------------
# configs[4]: render-only sweep at 3840x2160 (tile-list length and sort-size stress)
# --------------------------------------------------------------------------------------------
def run_sweep(args, device):
    """Forward only, both arms per cell on the same inputs: P x SH degree x scale_modifier at 4K.
    Per cell: visible Gaussians, tile instances R, mean / max tile list, the key format libbgr
    chose, per-stage ms of libbgr (CUDA events inside the library), ms per forward of each arm
    (CUDA events, 3 warm-up + 8 timed), and the parity of the two arms' outputs (integer buffers
    bit-equal, image max-abs difference)."""
    import synthetic
    import util
    ours = load_backend("ours")
    ref = load_backend("reference")
    W, H, focal = synthetic.RESOLUTIONS[args.res if args.res != "1080p" else "4k"]
    cam = synthetic.make_camera(W, H, focal)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    bg = torch.zeros(3, device=device)
    cells = []
    for pm in [float(x) for x in args.sweep_gaussians.split(",")]:
        P = int(round(pm * 1e6))
        scene = {k: v.to(device) for k, v in synthetic.make_gaussians(P, seed=0).items()}
        for deg in [int(x) for x in args.sweep_degrees.split(",")]:
            for mod in [float(x) for x in args.sweep_mods.split(",")]:
                def fwd(backend):
                    return util.native_forward(backend, scene, cam, bg, deg, mod)[1]

                def time_arm(backend, n_warm=3, n=8):
                    for _ in range(n_warm):
                        out = fwd(backend)
                    torch.cuda.synchronize(device)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(n):
                        out = fwd(backend)
                    e1.record()
                    torch.cuda.synchronize(device)
                    return e0.elapsed_time(e1) / n, out

                ms_ours, out = time_arm(ours)
                ours.stage_timing(True)
                acc = {}
                for _ in range(3):
                    fwd(ours)
                    torch.cuda.synchronize(device)
                    for k, v in ours.stage_times().items():
                        acc[k] = acc.get(k, 0.0) + v / 3
                ours.stage_timing(False)
                info = ours.last_forward_info()
                R = int(out[0])
                m = util.bgr_intermediates(out, P, W, H)
                lens = (m["ranges"][:, 1] - m["ranges"][:, 0]).long()
                st = ours.render_stats(out, P, W, H)
                cell = {"gaussians": P, "sh_degree": deg, "scale_modifier": mod,
                        "visible": int((out[2] > 0).sum()), "num_rendered": R,
                        "mean_tile_list": round(R / tiles, 1), "max_tile_list": int(lens.max()),
                        "sum_n_contrib": int(m["n_contrib"].long().sum()),
                        "binning": {"index_bits": info["nb_idx"], "depth_bits": info["nb_depth"],
                                    "smem_sort_capacity": info["smem_keys"],
                                    "global_sort_path": info["global_sort_path"]},
                        "ours_ms": round(ms_ours, 4), "ours_stage_ms": {k: round(v, 4) for k, v in acc.items()},
                        "pairs_examined": st["examined"], "pairs_blended": st["blended"]}
                if ref is not None:
                    ms_ref, o_r = time_arm(ref)
                    im = ref.parse_img(o_r[5], W, H)
                    b = ref.parse_binning(o_r[4], o_r[0])
                    cell["reference_ms"] = round(ms_ref, 4)
                    cell["speedup"] = round(ms_ref / ms_ours, 3)
                    cell["parity"] = {
                        "num_rendered": int(o_r[0]) == R,
                        "radii": bool(torch.equal(o_r[2], out[2])),
                        "point_list": bool(torch.equal(b["point_list"], m["point_list"])),
                        "ranges": bool(torch.equal(im["ranges"][:tiles], m["ranges"])),
                        "n_contrib": bool(torch.equal(im["n_contrib"], m["n_contrib"])),
                        "image_max_abs_diff": float((o_r[1] - out[1]).abs().max()),
                    }
                    del o_r, im, b
                cells.append(cell)
                print(json.dumps(cell), file=sys.stderr, flush=True)
                del out, m
        del scene
        torch.cuda.empty_cache()
    doc = {"config": f"BASELINE.json configs[4]: render-only sweep at {W}x{H}, seeded scene of SURVEY.md 8d, "
                     "forward only, inputs resident", "gpu": torch.cuda.get_device_name(device),
           "reference": "oracle/_ref (the reference's unmodified .cu files, sm_100a)" if ref is not None else None,
           "cells": cells}
    out_path = Path(args.out)
    out_path.parent.mkdir(parents=True, exist_ok=True)
    out_path.write_text(json.dumps(doc, indent=1))
    ok = all(all(v is True or (k == "image_max_abs_diff" and v <= 1e-5) for k, v in c.get("parity", {}).items())
             for c in cells)
    print(json.dumps({"mode": "sweep", "cells": len(cells), "parity_all_cells": ok, "out": str(out_path),
                      "speedup_min": min((c.get("speedup", 0) for c in cells), default=None),
                      "speedup_max": max((c.get("speedup", 0) for c in cells), default=None)}))


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        print(f"--gpus {args.gpus} needs one rank per GPU: launch with torch.distributed.run "
              f"--nproc-per-node {args.gpus}", file=sys.stderr)
        sys.exit(2)
    backend = load_backend(args.impl)
    if backend is None:
        if rank == 0:
            print(json.dumps({"impl": args.impl, "unavailable":
                              "oracle/_ref/libref_rasterizer.so (or ref_dgr_C, for reference-ext) missing: build with "
                              "`make -C oracle ref` / `python oracle/build_ref_ext.py` where /root/reference exists"}))
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    if args.mode == "sweep":
        if world > 1:
            raise SystemExit("--mode sweep is a single-GPU run")
        run_sweep(args, device)
        return
    if args.mode == "f1":
        if world > 1 or args.impl != "ours":
            raise SystemExit("--mode f1 is a single-GPU run of this library")
        run_f1(args, device)
        return
    if args.mode == "train":
        if world > 1 or args.impl != "ours":
            raise SystemExit("--mode train is a single-GPU run (it times all arms itself)")
        run_train(args, device)
        return

    wl = Workload(args, rank, device)
    native = make_native_step(backend, wl, world, args.impl, args.exchange, args.chunks, args.nvls_blocks)

    launches0 = backend.launch_count() if args.impl == "ours" else 0
    native()  # first call: lazy init (pinned control block, func attributes)
    torch.cuda.synchronize(device)
    per_step_launches = (backend.launch_count() - launches0) if args.impl == "ours" else None

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total = timed(native, args.steps, args.warmup, world, device)
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms_total / args.steps
    value = world * 1000.0 / ms_per_step

    e2e = None
    if not args.skip_e2e:
        e2e_step, h2d, d2h = make_e2e_step(backend, wl, world, args.impl)
        ms_e2e = timed(e2e_step, args.steps, args.warmup, world, device) / args.steps
        e2e = {"value": round(world * 1000.0 / ms_e2e, 3), "unit": "views/s", "ms_per_step": round(ms_e2e, 4),
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "api": "GaussianRasterizer(settings)(...) + loss.backward(); Gaussians stay resident "
                      "(as in train.py), camera block + (3,H,W) loss weights copied from pinned memory"}

    line = {
        "metric": "train views/s (rasterizer fwd+bwd per view, 1080p, SH3)", "value": round(value, 3),
        "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": args.impl,
        "config": {"workload": f"{wl.P} Gaussians SH3 @ {wl.W}x{wl.H}, one camera per GPU "
                               f"(BASELINE.json configs[2]; seeded scene of SURVEY.md 8d)",
                   "gaussians": wl.P, "width": wl.W, "height": wl.H, "sh_degree": 3,
                   "parallelism": f"view-batch dp{world}" + (f", all-reduce of 59 floats/Gaussian ({args.exchange if args.impl == 'ours' else 'nccl'})" if world > 1 else ""),
                   "l2": "no flush: per-step inputs (708 MB at 3M) exceed the 126 MB L2"},
        "e2e": e2e, "clocks": clocks,
    }
    if args.impl == "ours":
        line["gpu_launches"] = per_step_launches * args.steps
        line["gpu_launches_per_step"] = per_step_launches
        if world == 1:
            roof, stages, hbm_top, stats, _ = roofline_report(backend, wl, max(3, min(args.steps, 10)), device)
            line["roofline"] = roof
            line["roofline_hbm"] = hbm_top
            line["roofline_stages"] = stages
            line["workload_stats"] = stats
        if world == 1 and not args.skip_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
    else:
        line["reference"] = (("the reference's own torch extension (the five sources of its setup.py, unmodified, "
                              "pybind module built by oracle/build_ref_ext.py for sm_100a) on the same GPU; "
                              if args.impl == "reference-ext" else
                              "unmodified reference CUDA kernels (oracle/_ref, nvcc sm_100a) on the same GPU; ") +
                             "the reference has no CPU implementation of this path")
        line["cpu_baseline"] = {"value": line["value"], "unit": "views/s", "cores": 0, "kind": "reference",
                                "sample": "not a CPU run: reference CUDA extension, same workload and harness"}
    if world > 1:
        if getattr(make_native_step, "fallback", None):
            line["config"]["parallelism"] += f" [fell back to nccl: {make_native_step.fallback}]"
        nvls = getattr(make_native_step, "nvls", None)
        pieces = None
        if nvls is not None:      # exchange pieces, outside the timed region
            nvls.profile = True
            for _ in range(20):
                native()
            pieces = nvls.profile_report()
            nvls.profile = False
        line["dp"] = dp_breakdown(backend, wl, args.impl, world, device, ms_per_step)
        line["dp"]["exchange_pieces_ms_rank0"] = pieces
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
