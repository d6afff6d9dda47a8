"""B200-native differentiable Gaussian-splatting rasterizer (hot path of
XinShuo-ph/gaussian_splatting_with_eye_tracking), behind the reference's own operator API.

Layout
------
csrc/            hand-written sm_100a CUDA kernels + the extern "C" ABI (include/gsr_b200.h)
_capi.py         ctypes binding of that ABI (loads libgsr_b200.so; raises if it cannot)
_host.py         PyTorch host shim with the reference's `_C` call signatures
rasterizer.py    GaussianRasterizationSettings / GaussianRasterizer / autograd functions (baseline + AMR)
scenes.py        deterministic synthetic Gaussian clouds and cameras
dist.py          data-parallel views (NCCL / NVLink peer-memory gradient exchange) and screen-tile-band split
loss_utils.py    l1_loss / ssim / l1_ssim_loss of the reference's utils/loss_utils.py as one kernel per direction
optim.py         FusedAdam: torch.optim.Adam's state layout, every parameter group in one launch
scene_io.py      the reference's point-cloud PLY and cameras.json, plyfile-free
model.py         GaussianModel with the reference class's interface (densify / prune / reset bookkeeping, schedules)

The drop-in packages `diff_gaussian_rasterization` and `diff_gaussian_rasterization_amr` at the repository root
re-export this API under the reference's import names.
"""
from .rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizerAMR,
    GaussianRasterizerBaseline,
    RasterizeGaussiansAMR,
    RasterizeGaussiansBaseline,
    RasterizeGaussiansRaw,
    rasterize_gaussians_amr,
    rasterize_gaussians_baseline,
    rasterize_gaussians_raw,
    render_raw,
)

__all__ = [
    "GaussianRasterizationSettings", "GaussianRasterizerBaseline", "GaussianRasterizerAMR",
    "RasterizeGaussiansBaseline", "RasterizeGaussiansAMR", "rasterize_gaussians_baseline", "rasterize_gaussians_amr",
]
