"""GPU suite, multi-device part: the view-parallel gradient exchange on every GPU of the box
(skips below two devices). Runs tests/dp_check.py under torchrun: the in-switch sum / factor
multicast result against the float64 sum of the per-rank gradients at the path's own bar, and
bit-identical replicas, in the pipelined factor mode the bench uses and in the dense mode."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _run(world, P, mode, chunks, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "dp_check.py"),
           str(P), mode, str(chunks)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    tail = r.stdout[-3000:] + r.stderr[-1500:]
    assert r.returncode == 0, tail
    for rank in range(world):
        assert f"DP_CHECK rank {rank} world {world} PASS" in r.stdout, tail
    return r.stdout


@pytest.mark.parametrize("mode,chunks,P", [("factor", 4, 400_000), ("factor", 1, 100_004), ("dense", 1, 100_003)])
def test_view_parallel_exchange_against_fp64_sum(mode, chunks, P):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs at least two CUDA devices")
    world = min(8, torch.cuda.device_count())
    out = _run(world, P, mode, chunks, 29610 + chunks)
    print(out[-2500:])
