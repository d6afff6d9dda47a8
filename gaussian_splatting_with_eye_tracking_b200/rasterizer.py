"""Operator API of the rasterizer: settings tuple, autograd functions and nn.Module wrappers.

Keeps the public surface of the reference's two Python packages so that ``gaussian_renderer/__init__.py`` and
``gaussian_renderer_amr/__init__.py`` run unchanged on top of it:

* baseline: ``diff_gaussian_rasterization/__init__.py`` (settings :157-169, module :171-220, autograd :44-155)
* AMR:      ``diff_gaussian_rasterization_amr/__init__.py`` (15-input ``.apply`` -> 5 outputs, :72-179; module :252-366)

Both flavours are generated from one implementation here; only the argument marshalling differs.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from . import _host


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def cpu_deep_copy_tuple(input_tuple):
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


def _call_with_snapshot(fn, args, debug, dump_name, what):
    """debug=True keeps the reference's crash-dump behaviour (diff_gaussian_rasterization/__init__.py:83-90)."""
    if not debug:
        return fn(*args)
    cpu_args = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(cpu_args, dump_name)
        print(f"\nAn error occured in {what}. Writing {dump_name} for debugging.\n")
        raise


def _forward_args(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp):
    return (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree,
            rs.campos, rs.prefiltered)


def _backward_impl(ctx, grad_out_color, C_mod):
    rs = ctx.raster_settings
    (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
     imgBuffer) = ctx.saved_tensors
    args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos, geomBuffer,
            ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
    (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
     grad_rotations) = _call_with_snapshot(C_mod.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
    return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
            grad_cov3Ds_precomp)


# ----------------------------------------------------------------------------------------------------------------
# baseline flavour
# ----------------------------------------------------------------------------------------------------------------
class RasterizeGaussiansBaseline(torch.autograd.Function):
    _C = _host.C_baseline

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
        rs = raster_settings
        args = _forward_args(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp) + (rs.debug,)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _call_with_snapshot(
            RasterizeGaussiansBaseline._C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii=None):
        return _backward_impl(ctx, grad_out_color, RasterizeGaussiansBaseline._C) + (None,)


def rasterize_gaussians_baseline(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                                 raster_settings):
    return RasterizeGaussiansBaseline.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                            cov3Ds_precomp, raster_settings)


# ----------------------------------------------------------------------------------------------------------------
# raw-parameter entry (extension): the model's RAW tensors go straight to the kernels
# ----------------------------------------------------------------------------------------------------------------
class RasterizeGaussiansRaw(torch.autograd.Function):
    """Rasterize from the raw parameters of scene/gaussian_model.py -- ``_opacity`` (logits), ``_scaling`` (log-scales),
    ``_rotation`` (un-normalised quaternions), ``_features_dc`` [P,1,3] and ``_features_rest`` [P,M-1,3] -- without the activated
    copies the reference's caller materialises every frame (gaussian_model.py:95-115: sigmoid, exp, normalize, torch.cat =
    236 B/Gaussian read and written again before the rasterizer starts).  The kernels activate on load; the backward returns
    the gradients of the activated quantities and they are chained through the activations here (11 floats per Gaussian)."""

    @staticmethod
    def forward(ctx, means3D, means2D, features_dc, features_rest, raw_opacity, raw_scaling, raw_rotation, raster_settings):
        rs = raster_settings
        e = means3D.new_empty(0)
        args = (rs.bg, means3D, e, raw_opacity, raw_scaling, raw_rotation, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                rs.tanfovy, rs.image_height, rs.image_width, features_dc, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _host.rasterize_gaussians(*args, sh_rest=features_rest)
        ctx.raster_settings, ctx.num_rendered = rs, num_rendered
        ctx.save_for_backward(means3D, features_dc, features_rest, raw_opacity, raw_scaling, raw_rotation, radii, geomBuffer, binningBuffer, imgBuffer)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii=None):
        rs = ctx.raster_settings
        means3D, f_dc, f_rest, raw_op, raw_sc, raw_rot, radii, geomBuffer, binningBuffer, imgBuffer = ctx.saved_tensors
        e = means3D.new_empty(0)
        (g_means2D, _g_col, g_op, g_means3D, _g_cov, g_sh, g_sc, g_rot) = _host.rasterize_gaussians_backward(
            rs.bg, means3D, radii, e, raw_sc, raw_rot, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
            grad_out_color, f_dc, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug, sh_rest=f_rest)
        sig = torch.sigmoid(raw_op)
        g_raw_op = g_op.view_as(raw_op) * sig * (1.0 - sig)
        g_raw_sc = g_sc * torch.exp(raw_sc)
        n = raw_rot.norm(dim=1, keepdim=True).clamp_min(1e-12)
        qh = raw_rot / n
        g_raw_rot = (g_rot - qh * (qh * g_rot).sum(dim=1, keepdim=True)) / n
        return g_means3D, g_means2D, g_sh[:, :1], g_sh[:, 1:], g_raw_op, g_raw_sc, g_raw_rot, None


def rasterize_gaussians_raw(means3D, means2D, features_dc, features_rest, raw_opacity, raw_scaling, raw_rotation, raster_settings):
    return RasterizeGaussiansRaw.apply(means3D, means2D, features_dc, features_rest, raw_opacity, raw_scaling, raw_rotation, raster_settings)


def render_raw(viewpoint_camera, pc, bg_color, scaling_modifier=1.0, debug=False):
    """``gaussian_renderer.render`` (gaussian_renderer/__init__.py:18-113) on the raw-parameter entry: same result dictionary, no
    activated parameter copies.  ``pc`` is a GaussianModel (``_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation``)."""
    import math
    screenspace_points = torch.zeros_like(pc._xyz, requires_grad=True)
    rs = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False, debug=debug)
    image, radii = rasterize_gaussians_raw(pc._xyz, screenspace_points, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling,
                                           pc._rotation, rs)
    return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii}


def _check_optional_inputs(shs, colors_precomp, scales, rotations, cov3D_precomp):
    # same two checks, same messages as the reference (diff_gaussian_rasterization/__init__.py:191-195)
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    has_sr = scales is not None or rotations is not None
    if ((scales is None or rotations is None) and cov3D_precomp is None) or (has_sr and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')


def _absent():
    return torch.Tensor([])


class _RasterizerBase(nn.Module):
    _C = None

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return self._C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)


class GaussianRasterizerBaseline(_RasterizerBase):
    _C = _host.C_baseline

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        _check_optional_inputs(shs, colors_precomp, scales, rotations, cov3D_precomp)
        opt = [x if x is not None else _absent() for x in (shs, colors_precomp, scales, rotations, cov3D_precomp)]
        shs, colors_precomp, scales, rotations, cov3D_precomp = opt
        return rasterize_gaussians_baseline(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                            cov3D_precomp, self.raster_settings)


# ----------------------------------------------------------------------------------------------------------------
# AMR / foveated flavour (forward only -- the reference's AMR backward cannot run)
# ----------------------------------------------------------------------------------------------------------------
class RasterizeGaussiansAMR(torch.autograd.Function):
    _C = _host.C_amr
    # gaze-driven level cap (extension, off by default): set to (cx, cy, (r1, r2, r3, r4)) in pixels
    gaze = None

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, foveaStep,
                out_color_precomp, geomBuffer_precomp, binningBuffer_precomp, imageBuffer_precomp, interpolate_image,
                raster_settings):
        rs = raster_settings
        args = _forward_args(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp) + (
            foveaStep, out_color_precomp, geomBuffer_precomp, binningBuffer_precomp, imageBuffer_precomp,
            interpolate_image, rs.debug)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _host.rasterize_gaussians_amr(
            *args, gaze=RasterizeGaussiansAMR.gaze)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        return color, radii, geomBuffer, binningBuffer, imgBuffer

    @staticmethod
    def backward(ctx, *grads):
        raise NotImplementedError(
            "the AMR/foveated rasterizer is forward-only (the reference's AMR backward is non-functional); "
            "use diff_gaussian_rasterization for training")


def rasterize_gaussians_amr(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, foveaStep,
                            out_color_precomp, geomBuffer_precomp, binningBuffer_precomp, imageBuffer_precomp,
                            interpolate_image, raster_settings):
    return RasterizeGaussiansAMR.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                                       foveaStep, out_color_precomp, geomBuffer_precomp, binningBuffer_precomp,
                                       imageBuffer_precomp, interpolate_image, raster_settings)


class GaussianRasterizerAMR(_RasterizerBase):
    _C = _host.C_amr

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, foveaStep=int(0), out_color_precomp=None, geomBuffer_precomp=None,
                binningBuffer_precomp=None, imageBuffer_precomp=None, interpolate_image=True):
        _check_optional_inputs(shs, colors_precomp, scales, rotations, cov3D_precomp)
        opt = [x if x is not None else _absent() for x in (shs, colors_precomp, scales, rotations, cov3D_precomp)]
        shs, colors_precomp, scales, rotations, cov3D_precomp = opt
        if out_color_precomp is None:
            out_color_precomp = _absent()
        bufs = [b if b is not None else _absent().to(torch.uint8)
                for b in (geomBuffer_precomp, binningBuffer_precomp, imageBuffer_precomp)]
        return rasterize_gaussians_amr(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                       foveaStep, out_color_precomp, bufs[0], bufs[1], bufs[2], interpolate_image,
                                       self.raster_settings)
