"""Where is the GPU idle inside the end-to-end step? Runs bench.py's e2e step under the torch
profiler (CUPTI kernel timestamps), and for one steady-state step prints every device activity in
start order with the idle gap before it. Not a pytest; run on a GPU box:

    python tests/e2e_gap_profile.py [--gaussians N]
"""
import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "gaussian-splatting-own_b200", ROOT / "tests"):
    sys.path.insert(0, str(p))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--res", default="1080p")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    backend = bench.load_backend("ours")
    wl = bench.Workload(a, 0, dev)
    step, _, _ = bench.make_e2e_step(backend, wl, 1, "ours")
    for _ in range(10):
        step()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(6):
            step()
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # split into steps at the D2H copy of the loss (last activity of a step)
    steps, cur = [], []
    for e in evs:
        cur.append(e)
        if "DtoH" in e.name and cur:
            # the control-block read-back in the forward is also a DtoH copy: a step ends at the
            # second one
            if sum("DtoH" in x.name for x in cur) == 2:
                steps.append(cur)
                cur = []
    if len(steps) < 4:
        print("could not split the trace into steps", len(steps))
        return
    st = steps[3]
    prev_step_end = max(e.time_range.end for e in steps[2])
    t_first = st[0].time_range.start
    print(f"step: {len(st)} device activities; idle between the previous step's last activity and this "
          f"step's first: {t_first - prev_step_end:.1f} us")
    busy_end = t_first
    idle_total = 0.0
    compute = [e for e in st if "HtoD" not in e.name]     # the H2D copies run on their own stream
    for e in compute:
        gap = e.time_range.start - busy_end
        if gap > 0:
            idle_total += gap
        dur = e.time_range.end - e.time_range.start
        print(f"  +{e.time_range.start - t_first:9.1f} us  gap {max(gap, 0):7.1f}  dur {dur:8.1f}  {e.name[:90]}")
        busy_end = max(busy_end, e.time_range.end)
    span = busy_end - t_first
    print(f"step span (first to last device activity) {span:.1f} us, idle inside {idle_total:.1f} us, "
          f"idle before the step {t_first - prev_step_end:.1f} us")
    for e in st:
        if "HtoD" in e.name:
            print(f"  H2D copy at +{e.time_range.start - t_first:.1f} us, {e.time_range.end - e.time_range.start:.1f} us")


if __name__ == "__main__":
    main()
