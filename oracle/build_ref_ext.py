"""TEST INFRASTRUCTURE - builds a checker, never the product.

Compiles the reference's OWN torch extension (the five sources of DGR/setup.py:21-28: the three
cuda_rasterizer/*.cu, rasterize_points.cu and ext.cpp) from where they lie under /root/reference,
unmodified, for sm_100a, into oracle/_ref/ref_dgr_C<EXT_SUFFIX> (pybind module name `ref_dgr_C`, so
it imports next to the product in one process). Flags are what torch's BuildExtension gives that
setup.py (no fast-math, default -fmad) plus `-include cstdint`, which gcc 13 needs for
rasterizer_impl.h. Nothing is copied out of the reference tree; only the built .so lands in
oracle/_ref/ (git-ignored, shipped to the GPU box by gpurun).

    python oracle/build_ref_ext.py        # no-op when the reference tree is absent or the .so is newer

`bench.py --impl reference-ext` and tests/test_gpu_boundary.py use it to show that the stock pybind
path (rasterize_points.cu: torch::zeros outputs, resize lambdas) agrees with the shim arm
(oracle/ref_shim.cu) that the other comparisons run through.
"""
from __future__ import annotations

import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path("/root/reference/submodules/diff-gaussian-rasterization")
OUTDIR = HERE / "_ref"
NAME = "ref_dgr_C"
OUT = OUTDIR / (NAME + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))
SOURCES = ["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu", "cuda_rasterizer/backward.cu",
           "rasterize_points.cu", "ext.cpp"]
NVCC = "/usr/local/cuda/bin/nvcc"


def build(force: bool = False):
    srcs = [REF / s for s in SOURCES]
    if not all(s.exists() for s in srcs):
        return OUT if OUT.exists() else None
    if not force and OUT.exists() and all(OUT.stat().st_mtime >= s.stat().st_mtime for s in srcs):
        return OUT
    import torch
    from torch.utils import cpp_extension as ce
    OUTDIR.mkdir(exist_ok=True)
    objdir = OUTDIR / "ref_ext_obj"
    objdir.mkdir(exist_ok=True)
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{REF / 'third_party' / 'glm'}"]
    defs = [f"-DTORCH_EXTENSION_NAME={NAME}", "-DTORCH_API_INCLUDE_EXTENSION_H",
            f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    # the -D__CUDA_NO_HALF... set is what BuildExtension adds to every nvcc command
    nv = [NVCC, "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-include", "cstdint",
          "-Xcompiler", "-fPIC", "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
          "-D__CUDA_NO_BFLOAT16_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__", "--expt-relaxed-constexpr",
          "-Xcudafe", "--diag_suppress=177", *defs, *inc]
    cxx = ["g++", "-O3", "-std=c++17", "-fPIC", "-include", "cstdint", *defs, *inc]

    def compile_one(src: Path):
        obj = objdir / (src.name + ".o")
        cmd = (nv if src.suffix == ".cu" else cxx) + ["-c", str(src), "-o", str(obj)]
        subprocess.check_call(cmd)
        return obj

    with ThreadPoolExecutor(len(srcs)) as ex:
        objs = list(ex.map(compile_one, srcs))
    torch_lib = Path(torch.__file__).parent / "lib"
    subprocess.check_call(["g++", "-shared", *map(str, objs), "-o", str(OUT), f"-L{torch_lib}", "-lc10", "-ltorch",
                           "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-L/usr/local/cuda/lib64",
                           "-lcudart", f"-Wl,-rpath,{torch_lib}"])
    for o in objs:
        o.unlink()
    objdir.rmdir()
    return OUT


def load():
    """-> the reference's pybind module, or None when it was not built."""
    if not OUT.exists():
        return None
    import torch  # noqa: F401  (libtorch must be loaded first)
    sys.path.insert(0, str(OUTDIR))
    try:
        import ref_dgr_C
    finally:
        sys.path.remove(str(OUTDIR))
    return ref_dgr_C


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
