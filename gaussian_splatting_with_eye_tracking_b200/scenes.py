"""Deterministic synthetic Gaussian clouds, cameras and gaze trajectories.

No trained scene ships with the reference and there is no network, so every benchmark and parity input is
generated here from fixed seeds with a CPU ``torch.Generator`` (then moved to the GPU by the caller).  Tensors
follow the rasterizer's input contract: means [P,3], SH [P,16,3], opacity [P,1] (activated),
scales [P,3] (activated), rotations [P,4] unit quaternions (r,x,y,z).

Camera matrices follow the reference convention (scene/cameras.py:54-57, utils/graphics_utils.py:38-71):
``world_view_transform`` = W2C^T, ``full_proj_transform`` = (P @ W2C)^T, i.e. stored transposed so that the
kernels read them column-major.
"""
from __future__ import annotations

import math
from typing import NamedTuple

import torch

SH_C0 = 0.28209479177387814


class SynthCamera(NamedTuple):
    image_width: int
    image_height: int
    FoVx: float
    FoVy: float
    world_view_transform: torch.Tensor   # [4,4]
    full_proj_transform: torch.Tensor    # [4,4]
    camera_center: torch.Tensor          # [3]

    @property
    def tanfovx(self):
        return math.tan(self.FoVx * 0.5)

    @property
    def tanfovy(self):
        return math.tan(self.FoVy * 0.5)

    def to(self, device):
        return self._replace(world_view_transform=self.world_view_transform.to(device),
                             full_proj_transform=self.full_proj_transform.to(device),
                             camera_center=self.camera_center.to(device))


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:51-71 (getProjectionMatrix), float32."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4, dtype=torch.float32)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_camera(eye, target, width, height, fovx, fovy=None, znear=0.01, zfar=100.0, up=(0.0, 1.0, 0.0)) -> SynthCamera:
    """COLMAP-style camera (+z forward, +y down) looking from ``eye`` at ``target``."""
    if fovy is None:
        fovy = 2.0 * math.atan(math.tan(fovx / 2) * height / width)
    eye_t = torch.tensor(eye, dtype=torch.float64)
    f = torch.tensor(target, dtype=torch.float64) - eye_t
    f = f / f.norm()
    upv = torch.tensor(up, dtype=torch.float64)
    r = torch.linalg.cross(f, -upv)          # +x right when +y is down
    r = r / r.norm()
    d = torch.linalg.cross(f, r)             # +y down
    R = torch.stack([r, d, f])               # rows: camera axes in world coordinates
    W2C = torch.eye(4, dtype=torch.float64)
    W2C[:3, :3] = R
    W2C[:3, 3] = -R @ eye_t
    wvt = W2C.to(torch.float32).t().contiguous()
    proj = projection_matrix(znear, zfar, fovx, fovy).t().contiguous()
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return SynthCamera(int(width), int(height), float(fovx), float(fovy), wvt, full, center)


def _quats(g, n):
    q = torch.randn(n, 4, generator=g)
    return q / q.norm(dim=1, keepdim=True)


def _sh(g, n):
    sh = torch.zeros(n, 16, 3)
    sh[:, 0, :] = torch.randn(n, 3, generator=g) * 0.5 / SH_C0
    sh[:, 1:, :] = torch.randn(n, 15, 3, generator=g) * 0.05
    return sh


def _ball(g, n, radius):
    v = torch.randn(n, 3, generator=g)
    v = v / v.norm(dim=1, keepdim=True)
    return v * radius * torch.rand(n, 1, generator=g).pow(1.0 / 3.0)


def make_object_scene(P=100_000, seed=1001, log_scale_mean=math.log(0.012)):
    """S-obj (config C2): NeRF-synthetic-scale object, unit-ish ball of anisotropic Gaussians."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    means = _ball(g, P, 1.3)
    scales = torch.exp(torch.randn(P, 3, generator=g) * 0.6 + log_scale_mean)
    rots = _quats(g, P)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0)
    return dict(means3D=means.contiguous(), scales=scales.contiguous(), rotations=rots.contiguous(),
                opacities=opac.contiguous(), shs=_sh(g, P).contiguous(), sh_degree=3, bg=torch.ones(3))


def object_cameras(n=8, width=800, height=800, fov=0.6911, distance=4.0):
    cams = []
    for i in range(n):
        a = 2 * math.pi * i / n
        eye = (distance * math.sin(a), 0.0, -distance * math.cos(a))
        cams.append(look_at_camera(eye, (0.0, 0.0, 0.0), width, height, fov, fov))
    return cams


# frozen so that N/P is about 8 pairs per Gaussian at 1080p with 16x16 tiles; tuned once, see DESIGN.md
UNBOUNDED_LOG_SCALE_MEAN = math.log(0.013)   # N/P = 5.9 .. 10.6 over the 8 ring cameras, mean 8.0 (3 M, 1080p)


def make_unbounded_scene(P=3_000_000, seed=2002, log_scale_mean=UNBOUNDED_LOG_SCALE_MEAN):
    """S-scene (configs C3/C4/C5): 60 % foreground blob, 30 % far shell, 10 % dense cluster (hot tiles)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    n_fg = int(P * 0.6)
    n_bg = int(P * 0.3)
    n_cl = P - n_fg - n_bg
    fg = torch.randn(n_fg, 3, generator=g) * torch.tensor([3.0, 1.2, 3.0])
    d = torch.randn(n_bg, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    rad = 6.0 + 24.0 * torch.rand(n_bg, 1, generator=g)
    bgp = d * rad
    cl = torch.randn(n_cl, 3, generator=g) * 0.4 + torch.tensor([1.0, 0.0, 2.0])
    means = torch.cat([fg, bgp, cl], 0)
    log_s = torch.randn(P, 3, generator=g) * 0.7 + log_scale_mean
    log_s[n_fg:n_fg + n_bg] += torch.log(rad / 6.0)
    perm = torch.randperm(P, generator=g)        # the reference's point order is not spatially sorted either
    means, log_s = means[perm], log_s[perm]
    scales = torch.exp(log_s)
    rots = _quats(g, P)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.5 - 0.5)
    return dict(means3D=means.contiguous(), scales=scales.contiguous(), rotations=rots.contiguous(),
                opacities=opac.contiguous(), shs=_sh(g, P).contiguous(), sh_degree=3, bg=torch.zeros(3))


def unbounded_cameras(n=8, width=1920, height=1080, fovx=1.20, radius=4.5, y=0.3):
    cams = []
    for i in range(n):
        a = 2 * math.pi * i / n
        eye = (radius * math.sin(a), y, -radius * math.cos(a))
        cams.append(look_at_camera(eye, (0.0, 0.0, 0.0), width, height, fovx))
    return cams


def gaze_trajectory(n_frames=900, width=3840, height=2160, rate_hz=90.0, seed=4004):
    """Synthetic 'recorded' gaze (none ships with the reference): smooth pursuit + a saccade every 0.4 s."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    out = []
    sacc_every = int(round(0.4 * rate_hz))
    transit = max(1, int(round(0.020 * rate_hz)))
    off = torch.zeros(2)
    src = torch.zeros(2)
    dst = torch.zeros(2)
    t_sacc = -10**9
    for i in range(n_frames):
        t = i / rate_hz
        base = torch.tensor([width * (0.5 + 0.35 * math.sin(2 * math.pi * 0.13 * t)),
                             height * (0.5 + 0.30 * math.sin(2 * math.pi * 0.21 * t + 1.0))])
        if i > 0 and i % sacc_every == 0:
            target = torch.rand(2, generator=g) * 0.8 + 0.1
            src = off.clone()
            dst = torch.tensor([target[0] * width, target[1] * height]) - base
            t_sacc = i
        k = min(1.0, (i - t_sacc) / transit) if t_sacc >= 0 else 1.0
        off = src + (dst - src) * k if t_sacc >= 0 else off
        p = base + off
        out.append((float(min(max(p[0], 0.0), width - 1.0)), float(min(max(p[1], 0.0), height - 1.0))))
        off = off * 0.97           # pursuit pulls the gaze back to the moving target between saccades
    return out


def fovea_radii(width):
    """gaussian_renderer_amr/__init__.py:103-106: W/2, W/4, W/8, W/16."""
    return (width / 2.0, width / 4.0, width / 8.0, width / 16.0)


def raster_settings_for(cam: SynthCamera, bg, sh_degree=3, scale_modifier=1.0, debug=False):
    from .rasterizer import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg,
        scale_modifier=scale_modifier, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
        sh_degree=sh_degree, campos=cam.camera_center, prefiltered=False, debug=debug)
