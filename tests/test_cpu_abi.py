"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/gsr_b200.h declares
(and nothing torch-typed), argument validation works without a GPU, and the Python operator surface keeps the
reference's names, argument order and error behaviour.  No kernel is launched here."""
import ctypes as C
import inspect
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gsr_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"GSR_API\s+[\w\s\*]+?\b(gsr_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    names = _declared_symbols()
    for must in ("gsr_forward", "gsr_backward", "gsr_mark_visible", "gsr_amr_forward", "gsr_geom_bytes",
                 "gsr_binning_bytes", "gsr_image_bytes", "gsr_sort_pairs", "gsr_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    out = subprocess.run(["nm", "-D", "--defined-only", _capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.strip()}
    missing = [s for s in _declared_symbols() if s not in exported]
    assert not missing, f"declared in gsr_b200.h but not exported: {missing}"
    # every declared symbol is also bound (with argtypes) by the ctypes layer
    unbound = [s for s in _declared_symbols() if s not in _capi.SYMBOLS]
    assert not unbound, f"declared but not bound in _capi.SYMBOLS: {unbound}"
    # the C ABI is plain C with no torch/ATen dependency
    needed = subprocess.run(["readelf", "-d", _capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "libtorch" not in needed and "libc10" not in needed


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "gsr_b200.h"\nint main(void){ gsr_scene s; gsr_camera c; gsr_grads g; (void)s;(void)c;(void)g; return GSR_OK; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o",
                    str(tmp_path / "t.o")], check=True)


def test_version_and_size_queries():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    assert b"sm_100a" in lib.gsr_version()
    for n, want in [(2500, 12), (8160, 13), (32400, 15), (2040, 11)]:       # getHigherMsb, SURVEY T10
        assert lib.gsr_higher_msb(n) == want
    g1, g2 = lib.gsr_geom_bytes(1000), lib.gsr_geom_bytes(2000)
    assert 0 < g1 < g2
    assert lib.gsr_binning_bytes(0, 0, 1920, 1080, 0) > 0
    assert lib.gsr_binning_bytes(10_000, 3_000, 1920, 1080, 0) >= 4 * 10_000 + 16 * 3_000
    base, amr = lib.gsr_image_bytes(1920, 1080, 0), lib.gsr_image_bytes(1920, 1080, 1)
    assert base >= 8 * 1920 * 1080 and amr >= 8 * 1920 * 1080
    assert lib.gsr_sort_temp_bytes(1) > 0


def test_argument_validation_needs_no_gpu():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    from gaussian_splatting_with_eye_tracking_b200._capi import GsrCamera, GsrScene, lib
    n = C.c_int(0)
    rc = lib.gsr_forward(None, None, _capi.ALLOC_FN(lambda *a: None), None, None, None, C.byref(n), 0, None)
    assert rc == _capi.GSR_ERR_INVALID and b"null" in lib.gsr_last_error()
    sc = GsrScene(P=10, sh_degree=3, sh_coeffs=4)         # degree 3 needs 16 coefficients
    cam = GsrCamera(width=64, height=64, tan_fovx=1.0, tan_fovy=1.0)
    rc = lib.gsr_forward(C.byref(sc), C.byref(cam), _capi.ALLOC_FN(lambda *a: None), None, None, None, C.byref(n), 0, None)
    assert rc == _capi.GSR_ERR_INVALID
    with pytest.raises(_capi.GsrError):
        _capi.check(rc, "gsr_forward")
    assert lib.gsr_mark_visible(-1, None, None, None, None, None) == _capi.GSR_ERR_INVALID
    assert lib.gsr_sort_pairs(None, None, None, None, None, None, 5, 0, None, 0, None) == _capi.GSR_ERR_INVALID


def test_dropin_packages_export_reference_names():
    import diff_gaussian_rasterization as R
    import diff_gaussian_rasterization_amr as A
    for m in (R, A):
        for name in ("GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians", "_C"):
            assert hasattr(m, name), (m.__name__, name)
        for fn in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
            assert callable(getattr(m._C, fn))
    # settings: the reference's 12 fields in order (diff_gaussian_rasterization/__init__.py:157-169)
    assert R.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    # `_C` signatures: 19 / 21 / 25 positional arguments (rasterize_points.h:18-62, AMR :19-53)
    def positional(fn):      # keyword-only-by-convention extensions (gaze=, band=) carry defaults and come last
        return [n for n, q in inspect.signature(fn).parameters.items() if q.default is inspect.Parameter.empty]
    assert len(positional(R._C.rasterize_gaussians)) == 19
    assert len(positional(R._C.rasterize_gaussians_backward)) == 21
    assert len(positional(A._C.rasterize_gaussians)) == 25
    amr_params = list(inspect.signature(A._C.rasterize_gaussians).parameters)
    assert amr_params[18:25] == ["foveaStep", "out_color_precomp", "geomBuffer_precomp", "binningBuffer_precomp",
                                 "imageBuffer_precomp", "interpolate_image", "debug"]
    # module forward keyword surface
    assert list(inspect.signature(R.GaussianRasterizer.forward).parameters)[1:] == [
        "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    assert list(inspect.signature(A.GaussianRasterizer.forward).parameters)[9:] == [
        "foveaStep", "out_color_precomp", "geomBuffer_precomp", "binningBuffer_precomp", "imageBuffer_precomp", "interpolate_image"]
    # autograd arity: 9 inputs baseline, 15 inputs AMR (gaussian_renderer_amr/__init__.py:183-207)
    assert len(inspect.signature(R._RasterizeGaussians.forward).parameters) == 1 + 9
    assert len(inspect.signature(A._RasterizeGaussians.forward).parameters) == 1 + 15


def _settings(mod):
    return mod.GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                             torch.zeros(3), False, False)


def test_optional_input_checks_raise_like_the_reference():
    import diff_gaussian_rasterization as R
    import diff_gaussian_rasterization_amr as A
    m = torch.zeros(4, 3)
    for mod in (R, A):
        r = mod.GaussianRasterizer(_settings(mod))
        with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
            r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
        with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
            r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=m, scales=m,
              rotations=torch.zeros(4, 4))
        with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
            r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), colors_precomp=m)
        with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
            r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4),
              cov3D_precomp=torch.zeros(4, 6))


def test_host_shim_rejects_bad_inputs_before_any_launch():
    from gaussian_splatting_with_eye_tracking_b200 import _host
    e = torch.empty(0)
    bad = torch.zeros(4, 2)
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        _host.rasterize_gaussians(torch.zeros(3), bad, e, torch.zeros(4, 1), e, e, 1.0, e, torch.eye(4), torch.eye(4), 1.0,
                                  1.0, 8, 8, e, 0, torch.zeros(3), False, False)
    # CPU tensors are refused: there is no CPU fallback in the product path
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        _host.rasterize_gaussians(torch.zeros(3), torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), torch.zeros(4, 3),
                                  torch.zeros(4, 4), 1.0, e, torch.eye(4), torch.eye(4), 1.0, 1.0, 8, 8, e, 0, torch.zeros(3),
                                  False, False)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gaussian_splatting_with_eye_tracking_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "gsr_oracle" not in txt, f
    for f in ("diff_gaussian_rasterization/__init__.py", "diff_gaussian_rasterization_amr/__init__.py"):
        assert "oracle" not in open(os.path.join(ROOT, f)).read()
