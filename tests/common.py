"""Shared helpers of the parity tests: seeded cases, runners for (ours via the C-ABI | CPU oracle | reference CUDA)
and the comparison rules from BASELINE.md section 5 (bit-exact integers, 1e-4 image, 1e-3 relative gradients)."""
from __future__ import annotations

import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from gaussian_splatting_with_eye_tracking_b200 import scenes  # noqa: E402
from oracle import oracle as orc  # noqa: E402

IMG_TOL = 1e-4          # max-abs per channel (north_star)
GRAD_RTOL = 1e-3        # relative (north_star), plus an absolute floor of FLOOR * ||g||_inf:
GRAD_FLOOR_REF = 1e-5   #   vs the reference CUDA rasterizer (same expf, float atomics in a different order; measured <= 9e-7)
GRAD_FLOOR_ORACLE = 3e-4  # vs the CPU oracle: libm expf vs MUFU.EX2 flips single alpha/T threshold decisions; the REFERENCE
#                           CUDA rasterizer itself sits at the same distance from the oracle (measured 9.1e-5, tools/gradstats.py)


# ---------------------------------------------------------------------------------------------------------------
# cases
# ---------------------------------------------------------------------------------------------------------------
def _cloud(P, seed, spread=(1.2, 1.0, 1.2), log_scale=math.log(0.05), sigma=0.5, opacity_max=1.0):
    g = torch.Generator().manual_seed(seed)
    means = (torch.rand(P, 3, generator=g) - 0.5) * 2 * torch.tensor(spread)
    scales = torch.exp(torch.randn(P, 3, generator=g) * sigma + log_scale)
    rots = torch.randn(P, 4, generator=g)
    rots = rots / rots.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0).clamp_max(opacity_max)
    shs = torch.zeros(P, 16, 3)
    shs[:, 0] = torch.randn(P, 3, generator=g) * 0.5 / scenes.SH_C0
    shs[:, 1:] = torch.randn(P, 15, 3, generator=g) * 0.08
    return dict(means3D=means, scales=scales, rotations=rots, opacities=opac, shs=shs, sh_degree=3,
                bg=torch.tensor([0.1, 0.4, 0.7]))


def make_case(name):
    """Returns (scene dict of CPU tensors, SynthCamera)."""
    if name == "tiny":
        return _cloud(64, 11, log_scale=math.log(0.12)), scenes.look_at_camera((0.2, 0.1, -3.0), (0, 0, 0), 64, 48, 0.8)
    if name == "small":
        return _cloud(4000, 12), scenes.look_at_camera((1.0, 0.4, -3.2), (0, 0, 0), 160, 120, 0.9)
    if name == "ragged":          # W, H not multiples of 16 or 32
        return _cloud(3000, 13), scenes.look_at_camera((-0.8, 0.2, -3.0), (0, 0, 0), 203, 117, 0.9)
    if name == "inside":          # camera inside the cloud: near-plane culling, huge splats, clamped t.x/t.y
        return _cloud(3000, 14, spread=(3.0, 3.0, 3.0), log_scale=math.log(0.15)), scenes.look_at_camera(
            (0.1, 0.0, -0.2), (0, 0, 1.0), 128, 96, 1.2)
    if name == "dense":           # long per-tile lists (> 256 entries -> several batches) and early termination
        return _cloud(12000, 15, spread=(0.5, 0.5, 0.5), log_scale=math.log(0.06)), scenes.look_at_camera(
            (0.0, 0.0, -2.5), (0, 0, 0), 96, 96, 0.6)
    if name in ("deg0", "deg1", "deg2"):
        sc = _cloud(1500, 16 + int(name[-1]))
        sc["sh_degree"] = int(name[-1])
        return sc, scenes.look_at_camera((0.5, -0.3, -3.0), (0, 0, 0), 112, 80, 0.9)
    if name == "precomp":         # colors_precomp + cov3D_precomp instead of SH / scale+rotation
        sc = _cloud(1500, 20)
        L = torch.randn(1500, 3, 3, generator=torch.Generator().manual_seed(21)) * 0.05
        S = L @ L.transpose(1, 2) + 1e-4 * torch.eye(3)
        sc["cov3D_precomp"] = torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).contiguous()
        sc["colors_precomp"] = torch.rand(1500, 3, generator=torch.Generator().manual_seed(22))
        sc["scales"] = sc["rotations"] = sc["shs"] = None
        return sc, scenes.look_at_camera((0.5, 0.3, -3.0), (0, 0, 0), 112, 80, 0.9)
    if name == "offscreen":       # everything behind the camera: N == 0
        return _cloud(500, 23), scenes.look_at_camera((0.0, 0.0, -3.0), (0, 0, -6.0), 64, 64, 0.8)
    if name == "obj":             # scaled-down config C2
        sc = scenes.make_object_scene(P=20000, log_scale_mean=math.log(0.02))
        return sc, scenes.object_cameras(8, 256, 256)[3]
    raise KeyError(name)


ALL_CASES = ["tiny", "small", "ragged", "inside", "dense", "deg0", "deg1", "deg2", "precomp", "offscreen"]


def upstream_gradient(cam, seed=3003):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(3, cam.image_height, cam.image_width, generator=g) * 2 - 1


def _np(t):
    return None if t is None else t.detach().cpu().numpy()


def to_oracle(sc, cam):
    s = orc.Scene(_np(sc["means3D"]), _np(sc["opacities"]), _np(sc.get("scales")), _np(sc.get("rotations")),
                  _np(sc.get("shs")), _np(sc.get("colors_precomp")), _np(sc.get("cov3D_precomp")), sc["sh_degree"])
    c = orc.Camera(cam.image_width, cam.image_height, cam.tanfovx, cam.tanfovy, _np(cam.world_view_transform),
                   _np(cam.full_proj_transform), _np(cam.camera_center), _np(sc["bg"]))
    return s, c


# ---------------------------------------------------------------------------------------------------------------
# runners
# ---------------------------------------------------------------------------------------------------------------
def _absent(dev):
    return torch.empty(0, device=dev)


def cuda_inputs(sc, cam, dev="cuda"):
    d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    for k in ("scales", "rotations", "shs", "colors_precomp", "cov3D_precomp"):
        if d.get(k) is None:
            d[k] = _absent(dev)
    return d, cam.to(dev)


def fwd_args(d, cam):
    return (d["bg"], d["means3D"], d["colors_precomp"], d["opacities"], d["scales"], d["rotations"], 1.0,
            d["cov3D_precomp"], cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy,
            cam.image_height, cam.image_width, d["shs"], d["sh_degree"], cam.camera_center, False)


def run_ours(sc, cam, dL_dpix=None, dev="cuda"):
    """Forward (+ backward) through the `_C`-shaped host shim, i.e. through the C-ABI."""
    from gaussian_splatting_with_eye_tracking_b200 import _host, views
    d, camd = cuda_inputs(sc, cam, dev)
    N, color, radii, geom, binning, img = _host.rasterize_gaussians(*fwd_args(d, camd), False)
    P = d["means3D"].shape[0]
    out = dict(N=N, out_color=color, radii=radii)
    if P:
        out.update(views.geom_view(geom, P))
        out["radii"] = radii
        out.update(views.binning_view(binning, N))
        out.update(views.image_view(img, cam.image_width, cam.image_height, False))
        out["tiles"] = views.tiles_from_ranges(out["ranges"], N)
        out["keys"] = views.sorted_keys(out["tiles"], out["point_list"], out["depths"])   # reference's (tile << 32 | depth) keys
    if dL_dpix is not None:
        g = _host.rasterize_gaussians_backward(
            d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0, d["cov3D_precomp"],
            camd.world_view_transform, camd.full_proj_transform, cam.tanfovx, cam.tanfovy, dL_dpix.to(dev), d["shs"],
            d["sh_degree"], camd.camera_center, geom, N, binning, img, False)
        names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"]
        out["grads"] = dict(zip(names, g))
    torch.cuda.synchronize()
    return out


def run_reference(sc, cam, dL_dpix=None, dev="cuda"):
    """Same, through the unmodified reference CUDA rasterizer (oracle/_ref)."""
    from oracle import ref_cuda
    ref = ref_cuda.RefRasterizer(False)
    d, camd = cuda_inputs(sc, cam, dev)
    a = fwd_args(d, camd)
    N, color, radii, geom, binning, img = ref.forward(*a[:12], cam.image_height, cam.image_width, d["shs"], d["sh_degree"],
                                                      camd.camera_center)
    P = d["means3D"].shape[0]
    out = dict(N=N, out_color=color, radii=radii)
    out.update(ref.parse(geom, binning, img, P, N, cam.image_width, cam.image_height))
    if dL_dpix is not None:
        g = ref.backward(d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0,
                         d["cov3D_precomp"], camd.world_view_transform, camd.full_proj_transform, cam.tanfovx, cam.tanfovy,
                         dL_dpix.to(dev), d["shs"], d["sh_degree"], camd.camera_center, geom, N, binning, img)
        out["grads"] = g
    torch.cuda.synchronize()
    return out


def run_oracle(sc, cam, dL_dpix=None):
    s, c = to_oracle(sc, cam)
    st = orc.forward(s, c)
    out = dict(N=st.num_rendered, out_color=st.out_color, radii=st.radii, depths=st.depths, means2D=st.means2D,
               cov3D=st.cov3D, conic_opacity=st.conic_opacity, rgb=st.rgb, tiles_touched=st.tiles_touched,
               clamped=st.clamped, keys=st.keys.view(np.int64), point_list=st.point_list.view(np.int32),
               ranges=st.ranges.view(np.int32), n_contrib=st.n_contrib.view(np.int32), accum_alpha=st.final_T)
    if dL_dpix is not None:
        out["grads"] = orc.backward(s, c, st, _np(dL_dpix))
    return out


# ---------------------------------------------------------------------------------------------------------------
# comparison rules
# ---------------------------------------------------------------------------------------------------------------
def A(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


def assert_ints_equal(a, b, what):
    a, b = A(a), A(b)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    bad = np.nonzero(a.reshape(-1) != b.reshape(-1))[0]
    assert bad.size == 0, f"{what}: {bad.size} of {a.size} differ (first at {bad[:5]}: {a.reshape(-1)[bad[:5]]} vs {b.reshape(-1)[bad[:5]]})"


def assert_bits_equal(a, b, what, mask=None):
    """float arrays compared as raw bits (+0 == -0 tolerated) on the rows selected by mask."""
    a, b = A(a), A(b)
    if mask is not None:
        a, b = a[A(mask)], b[A(mask)]
    bad = np.nonzero((a != b) & ~(np.isnan(a) & np.isnan(b)))
    assert bad[0].size == 0, f"{what}: {bad[0].size} of {a.size} values differ bitwise; max abs diff {np.abs(a - b).max()}"


def grad_close(a, b, rtol=GRAD_RTOL, floor_factor=GRAD_FLOOR_ORACLE):
    a, b = A(a).astype(np.float64), A(b).astype(np.float64)
    floor = floor_factor * max(np.abs(a).max(initial=0.0), np.abs(b).max(initial=0.0)) + 1e-12
    err = np.abs(a - b) - rtol * np.maximum(np.abs(a), np.abs(b)) - floor
    return float(err.max(initial=-1.0)), float(np.abs(a - b).max(initial=0.0))


def assert_grads_close(ga, gb, names=None, rtol=GRAD_RTOL, floor_factor=GRAD_FLOOR_ORACLE):
    for k in (names or ["dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations", "dL_dcolors"]):
        if k not in ga or k not in gb:
            continue
        a, b = A(ga[k]), A(gb[k])
        if a.size == 0 and b.size == 0:
            continue
        viol, mx = grad_close(a.reshape(b.shape), b, rtol, floor_factor)
        assert viol <= 0, f"{k}: gradient mismatch beyond {rtol} relative (max abs diff {mx}, |g|max {np.abs(b).max()})"


# ---------------------------------------------------------------------------------------------------------------
# AMR / foveated runners
# ---------------------------------------------------------------------------------------------------------------
def amr_args(d, cam, step, precomp, gb, bb, ib, interp):
    return fwd_args(d, cam) + (step, precomp, gb, bb, ib, interp, False)


def run_ours_amr(sc, cam, stepped, interpolate, gaze=None, dev="cuda"):
    """single pass (foveaStep=-2) or the 5-step protocol of gaussian_renderer_amr/__init__.py:132-608."""
    from gaussian_splatting_with_eye_tracking_b200 import _host, views
    d, camd = cuda_inputs(sc, cam, dev)
    e = torch.empty(0, device=dev)
    eb = torch.empty(0, dtype=torch.uint8, device=dev)
    P = d["means3D"].shape[0]
    out = {}
    if not stepped:
        N, color, radii, gb, bb, ib = _host.rasterize_gaussians_amr(*amr_args(d, camd, -2, e, eb, eb, eb, interpolate), gaze=gaze)
        out["images"] = [color]
    else:
        N, color0, radii, gb, bb, ib = _host.rasterize_gaussians_amr(*amr_args(d, camd, 0, e, eb, eb, eb, False), gaze=gaze)
        acc = color0.clone()
        imgs = [color0]
        for k in (1, 2, 3, 4):
            _, ck, _, gb, bb, ib = _host.rasterize_gaussians_amr(
                *amr_args(d, camd, k, acc, gb, bb, ib, interpolate if k == 4 else False), gaze=gaze)
            imgs.append(ck)
            acc = acc + ck
        out["images"] = imgs
        color = acc
    out.update(N=N, out_color=color, radii=radii)
    out.update(views.geom_view(gb, P))
    out["radii"] = radii
    out.update(views.binning_view(bb, N))
    out.update(views.image_view(ib, cam.image_width, cam.image_height, True))
    out["tiles"] = views.tiles_from_ranges(out["ranges"], N)
    out["keys"] = views.sorted_keys(out["tiles"], out["point_list"], out["depths"])
    torch.cuda.synchronize()
    return out


def run_oracle_amr(sc, cam, stepped, interpolate, gaze=None):
    s, c = to_oracle(sc, cam)
    st = orc.amr_build(s, c)
    if not stepped:
        img = orc.amr_render(c, st, -2, None, interpolate, gaze)
        imgs = [img]
    else:
        acc = np.zeros((3, c.H, c.W), np.float32)
        imgs = [acc.copy()]
        for k in (1, 2, 3, 4):
            ck = orc.amr_render(c, st, k, acc, interpolate if k == 4 else False, gaze)
            imgs.append(ck)
            acc = acc + ck
        img = acc
    return dict(N=st.num_rendered, out_color=img, images=imgs, radii=st.radii, keys=st.keys.view(np.int64),
                point_list=st.point_list.view(np.int32), ranges=st.ranges.view(np.int32),
                n_intersections=st.n_intersections.view(np.int32), levels=st.levels.view(np.int32),
                levels_last=st.levels_last.view(np.int32), levels_current=st.levels_current.view(np.int32),
                thresholds=st.thresholds.view(np.int32), tiles_touched=st.tiles_touched, n_contrib=st.n_contrib.view(np.int32))
