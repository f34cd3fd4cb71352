"""CPU suite: the C-ABI library loads and exports exactly what include/bgr.h declares, the pure
host entry points behave, and the Python drop-in surface mirrors the reference's
(diff_gaussian_rasterization/__init__.py:157-220). No compute call is made here (no GPU)."""
import ctypes as C
import inspect
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]
HEADER = ROOT / "include" / "bgr.h"


@pytest.fixture(scope="module")
def lib():
    from diff_gaussian_rasterization import _C
    return _C.lib()


def declared_symbols():
    text = HEADER.read_text()
    return sorted(set(re.findall(r"BGR_API\s+[\w\s\*]+?\b(bgr_\w+)\s*\(", text)))


def test_header_declares_the_reference_entry_points():
    syms = declared_symbols()
    for s in ("bgr_forward", "bgr_backward", "bgr_mark_visible", "bgr_last_error"):
        assert s in syms
    assert len(syms) >= 15


def test_library_exports_every_declared_symbol(lib):
    for s in declared_symbols():
        assert hasattr(lib, s), f"libbgr.so does not export {s}"


def test_host_only_entry_points(lib):
    assert b"sm_100a" in lib.bgr_version()
    assert lib.bgr_geom_bytes(0) >= 0
    g1, g2 = lib.bgr_geom_bytes(1000), lib.bgr_geom_bytes(2000)
    assert g2 > g1 >= 1000 * 52          # 48 B record + tiles_touched per Gaussian
    # final_T + n_contrib per pixel; ranges + cursor per tile; control block; coverage grid
    assert lib.bgr_image_bytes(1920, 1080) >= 1920 * 1080 * 8 + 8160 * 12 + 32 + 121 * 69 * 4
    assert lib.bgr_binning_bytes(1 << 20) >= (1 << 20) * 20   # ids + two key arrays (upper bound)
    assert lib.bgr_backward_scratch_bytes(1000) >= 1000 * 48
    off = (C.c_size_t * 2)()
    lib.bgr_geom_layout(1000, off)
    assert list(off) == sorted(off) and off[0] == 0 and all(o % 256 == 0 for o in off)
    off4 = (C.c_size_t * 4)()
    lib.bgr_image_layout(800, 800, off4)
    assert off4[1] - off4[0] >= 800 * 800 * 4 and list(off4) == sorted(off4)
    off3 = (C.c_size_t * 3)()
    lib.bgr_binning_layout(12345, off3)
    assert list(off3) == sorted(off3) and off3[0] == 0


def test_invalid_arguments_are_rejected_without_a_gpu(lib):
    from diff_gaussian_rasterization import _C
    n = C.c_int(0)
    cb = _C._ALLOC_FN(lambda nbytes, user: None)
    assert lib.bgr_forward(None, cb, None, cb, None, cb, None, None, C.byref(n)) == -1
    assert b"null" in lib.bgr_last_error()
    assert lib.bgr_backward(None, None) == -1
    assert lib.bgr_forward_raw(None, None, cb, None, cb, None, cb, None, None, C.byref(n)) == -1
    assert lib.bgr_backward_raw(None, None, None, None) == -1
    assert lib.bgr_mark_visible(-1, None, None, None, None, None) == -1


def test_python_surface_matches_reference_names():
    import diff_gaussian_rasterization as d
    for name in ("GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
                 "_RasterizeGaussians", "cpu_deep_copy_tuple", "_C"):
        assert hasattr(d, name)
    assert d.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    sig = inspect.signature(d.GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs",
                                    "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    # the three native functions with the reference's positional arity (ext.cpp:15-18)
    assert len(inspect.signature(d._C.rasterize_gaussians).parameters) == 19
    assert len(inspect.signature(d._C.rasterize_gaussians_backward).parameters) == 21
    assert len(inspect.signature(d._C.mark_visible).parameters) == 3
    assert issubclass(d._RasterizeGaussians, torch.autograd.Function)


def _settings(d):
    return d.GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4),
                                           torch.eye(4), 3, torch.zeros(3), False, False)


def test_input_mode_validation_messages():
    """Same exceptions, same (misspelt) text as the reference (__init__.py:191-195)."""
    import diff_gaussian_rasterization as d
    r = d.GaussianRasterizer(_settings(d))
    m, o = torch.zeros(4, 3), torch.zeros(4, 1)
    sh, col = torch.zeros(4, 16, 3), torch.zeros(4, 3)
    sc, rot, cov = torch.ones(4, 3), torch.zeros(4, 4), torch.zeros(4, 6)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, o, shs=sh, colors_precomp=col, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, o, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, o, shs=sh, scales=sc, rotations=rot, cov3D_precomp=cov)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, o, shs=sh, scales=sc)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, o, shs=sh)


def test_no_cpu_fallback():
    """CPU tensors must fail loudly: the product has no CPU or eager-PyTorch path."""
    import diff_gaussian_rasterization as d
    r = d.GaussianRasterizer(_settings(d))
    m, o = torch.zeros(4, 3), torch.zeros(4, 1)
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r(m, m, o, shs=torch.zeros(4, 16, 3), scales=torch.ones(4, 3), rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r.markVisible(m)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        d._C.rasterize_gaussians(torch.zeros(3), torch.zeros(4, 2), *([torch.Tensor([])] * 17))


def test_product_never_imports_the_oracle():
    """The product path must not import, load or link anything under oracle/."""
    pkg = ROOT / "gaussian-splatting-own_b200"
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|liboracle|libref_rasterizer|cpu_oracle|"
                     r"#include\s+\"[^\"]*oracle", re.M)
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        assert not pat.search(f.read_text()), f"{f} reaches into oracle/"


def test_bench_reads_roofline_traffic_from_the_committed_ncu_capture():
    """bench.py's `roofline.traffic` comes from profiles/*_ncu_full_raw.csv (one launch of the
    dominant kernel, DRAM read + write bytes); the file and the kernel rows must be there."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", ROOT / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert (ROOT / bench.NCU_RAW).exists()
    bwd = bench.ncu_traffic("render_backward")
    fwd = bench.ncu_traffic("render_forward")
    pre = bench.ncu_traffic("preprocess_backward")
    assert isinstance(bwd, int) and 2e8 < bwd < 6e8          # ~357 MB at 3M Gaussians / 1080p
    assert isinstance(fwd, int) and 1e8 < fwd < 4e8
    assert isinstance(pre, int) and 1e9 < pre < 2e9          # ~1.41 GB: HBM-bound stage
    assert bench.ncu_traffic("no_such_stage") is None


def test_ctypes_structs_match_the_c_header_field_by_field(tmp_path):
    """The argument blocks cross the boundary by layout, not by name: compile include/bgr.h as plain
    C and compare sizeof / offsetof of every field with the ctypes mirrors in _C.py."""
    import shutil
    import subprocess
    from diff_gaussian_rasterization import _C
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    pairs = (("BgrForwardArgs", _C._ForwardArgs), ("BgrBackwardArgs", _C._BackwardArgs))
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "bgr.h"', 'int main(void) {']
    for cname, ct in pairs:
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in ct._fields_:
            lines.append(f'  printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-I", str(ROOT / "include"), str(src), "-o", str(exe)],
                   check=True, capture_output=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True,
                                                 text=True).stdout.splitlines())
    for cname, ct in pairs:
        assert int(got[cname]) == C.sizeof(ct), cname
        for fname, _ in ct._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(ct, fname).offset, f"{cname}.{fname}"


def test_boundary_extension_builds_and_exposes_the_reference_signatures():
    """tests/boundary/rasterize_points_bgr.cpp (the reference's pybind functions over include/bgr.h)
    compiles as a torch extension and carries the reference's arities (ext.cpp:15-18)."""
    import importlib.util
    import sys
    spec = importlib.util.spec_from_file_location("boundary_build", ROOT / "tests" / "boundary" / "build.py")
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    out = b.build()
    assert out.exists()
    sys.path.insert(0, str(out.parent))
    try:
        import bgr_boundary
    finally:
        sys.path.remove(str(out.parent))
    doc = bgr_boundary.rasterize_gaussians.__doc__
    assert doc.count("arg") == 19 and "Tuple[int" in doc.replace("tuple[int", "Tuple[int")
    assert bgr_boundary.rasterize_gaussians_backward.__doc__.count("arg") == 21
    assert bgr_boundary.mark_visible.__doc__.count("arg") == 3
    # no GPU here: the binding must refuse CPU tensors through the library's own error, not crash
    with pytest.raises(RuntimeError):
        bgr_boundary.rasterize_gaussians(torch.zeros(3), torch.zeros(4, 2), *([torch.Tensor([])] * 2),
                                         torch.Tensor([]), torch.Tensor([]), 1.0, torch.Tensor([]),
                                         torch.eye(4), torch.eye(4), 1.0, 1.0, 16, 16, torch.Tensor([]), 3,
                                         torch.zeros(3), False, False)


def test_raw_leaf_entry_surface():
    """The extended entry beside the drop-in one: names, arities, and the same
    loud failure on CPU tensors."""
    import diff_gaussian_rasterization as d
    assert issubclass(d._RasterizeGaussiansRaw, torch.autograd.Function)
    sig = inspect.signature(d.RawGaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "xyz", "means2D", "features_dc", "features_rest", "opacity_raw",
                                    "scaling_raw", "rotation_raw"]
    assert len(inspect.signature(d._C.rasterize_gaussians_raw).parameters) == 18
    assert len(inspect.signature(d._C.rasterize_gaussians_raw_backward).parameters) == 20
    r = d.RawGaussianRasterizer(_settings(d))
    z = torch.zeros
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r(z(4, 3), z(4, 3), z(4, 1, 3), z(4, 15, 3), z(4, 1), z(4, 3), z(4, 4))
    with pytest.raises(RuntimeError, match="num_points, 3"):
        d._C.rasterize_gaussians_raw(z(3), z(4, 2), *([torch.Tensor([])] * 16))
