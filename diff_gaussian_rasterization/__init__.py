"""Drop-in for the reference package `diff_gaussian_rasterization`
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py): same exported names
(`GaussianRasterizationSettings`, `GaussianRasterizer`, `rasterize_gaussians`, `_RasterizeGaussians`, `_C`),
backed by the sm_100a library instead of the pybind/CUB extension."""
from gaussian_splatting_with_eye_tracking_b200 import _host as _host_mod
from gaussian_splatting_with_eye_tracking_b200.rasterizer import GaussianRasterizationSettings  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import GaussianRasterizerBaseline as GaussianRasterizer  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import RasterizeGaussiansBaseline as _RasterizeGaussians  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import cpu_deep_copy_tuple  # noqa: F401
from gaussian_splatting_with_eye_tracking_b200.rasterizer import rasterize_gaussians_baseline as rasterize_gaussians  # noqa: F401

_C = _host_mod.C_baseline
