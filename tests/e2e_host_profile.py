"""Host-side profile of the end-to-end step (where does the Python time go between the
synchronisation points?). Not a pytest; run on a GPU box: python tests/e2e_host_profile.py"""
import cProfile
import pstats
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "gaussian-splatting-own_b200"):
    sys.path.insert(0, str(p))
import bench  # noqa: E402


def main():
    sys.argv = ["bench.py"]
    args = bench.parse_args() if hasattr(bench, "parse_args") else None
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    backend = bench.load_backend("ours")

    class A:
        res = "1080p"; gaussians = 3_000_000
    wl = bench.Workload(A, 0, dev)
    step, _, _ = bench.make_e2e_step(backend, wl, 1, "ours")
    for _ in range(5):
        step()
    t0 = time.perf_counter()
    for _ in range(50):
        step()
    dt = (time.perf_counter() - t0) / 50
    print(f"wall per step {dt * 1e3:.3f} ms")
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(20):
            step()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=60))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(50):
        step()
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(22)


if __name__ == "__main__":
    main()
