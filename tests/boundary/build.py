"""Builds tests/boundary/rasterize_points_bgr.cpp (the reference-side binding over include/bgr.h) as
the torch extension `bgr_boundary`, in tree, linked against the product's libbgr.so.

    python tests/boundary/build.py
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parents[1]
LIBDIR = ROOT / "gaussian-splatting-own_b200" / "diff_gaussian_rasterization"
SUFFIX = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
OUT = HERE / ("bgr_boundary" + SUFFIX)
SRC = HERE / "rasterize_points_bgr.cpp"


def build(force: bool = False) -> Path:
    deps = [SRC, ROOT / "include" / "bgr.h"]
    if not force and OUT.exists() and all(OUT.stat().st_mtime >= d.stat().st_mtime for d in deps):
        return OUT
    import torch
    from torch.utils import cpp_extension as ce
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{cuda_home}/include", f"-I{ROOT / 'include'}",
                                                    f"-I{sysconfig.get_paths()['include']}"]
    torch_lib = Path(torch.__file__).parent / "lib"
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=bgr_boundary",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *inc, str(SRC), "-o", str(OUT),
           f"-L{LIBDIR}", "-l:libbgr.so", f"-Wl,-rpath,$ORIGIN/../../gaussian-splatting-own_b200/diff_gaussian_rasterization",
           f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda",
           f"-Wl,-rpath,{torch_lib}"]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
