"""Drop-in for the reference package `diff_gaussian_rasterization_amr`
(submodules/diff-gaussian-rasterization-amr/diff_gaussian_rasterization_amr/__init__.py): 32x32 binning tiles,
per-tile AMR level, 5-step progressive ("fovea") rendering.  `_RasterizeGaussians.apply` takes the reference's 15
positional inputs and returns its 5 outputs (gaussian_renderer_amr/__init__.py:183-207)."""
from gaussian_splatting_with_eye_tracking_b200 import _host as _host_mod
from gaussian_splatting_with_eye_tracking_b200.rasterizer import GaussianRasterizationSettings  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import GaussianRasterizerAMR as GaussianRasterizer  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import RasterizeGaussiansAMR as _RasterizeGaussians  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import cpu_deep_copy_tuple  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import rasterize_gaussians_amr as rasterize_gaussians  # noqa: F401

_C = _host_mod.C_amr
