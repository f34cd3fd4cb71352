"""Host-side cost of one end-to-end step (bench.py's e2e step: public autograd API, pinned H2D
copies, loss read-back) measured where the GPU work is negligible: a 2 000-Gaussian scene, so that
wall time per step ~ Python + launch overhead. Prints the wall time and a cProfile listing. Not a
pytest; run on a GPU box:

    python tests/host_overhead_probe.py [--impl ours|reference]
"""
import argparse
import cProfile
import pstats
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "gaussian-splatting-own_b200", ROOT / "tests"):
    sys.path.insert(0, str(p))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=2000)
    ap.add_argument("--res", default="800x800")
    ap.add_argument("--impl", default="ours")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    backend = bench.load_backend(a.impl)
    wl = bench.Workload(a, 0, dev)
    step, _, _ = bench.make_e2e_step(backend, wl, 1, a.impl)
    for _ in range(50):
        step()
    torch.cuda.synchronize()
    n = 500
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    print(f"{a.impl}: wall per step {1e6 * (time.perf_counter() - t0) / n:.1f} us at {a.gaussians} Gaussians, {a.res}")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(200):
        step()
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(28)
    st.sort_stats("cumulative").print_stats(22)


if __name__ == "__main__":
    main()
