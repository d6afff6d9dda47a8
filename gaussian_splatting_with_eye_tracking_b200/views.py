"""Typed tensor views of the three opaque work buffers (gsr_view_* in include/gsr_b200.h).

The byte tensors returned by ``_C.rasterize_gaussians`` are opaque to callers, exactly like the reference's
geomBuffer / binningBuffer / imgBuffer.  Tests, the tile-band driver and debugging use these zero-copy views; the
analogue in the reference is ParseBuffers of the amr-debug fork (diff-gaussian-rasterization-amr-debug/rasterize_points.cu:37-62).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _capi
from ._capi import GsrBinningView, GsrGeomView, GsrImageView, lib


def _slice(buf, ptr, nbytes, dtype):
    if not ptr:
        return None
    off = int(ptr) - buf.data_ptr()
    assert 0 <= off and off + nbytes <= buf.numel(), "view outside its buffer"
    return buf[off:off + nbytes].view(dtype)


def geom_view(geom: torch.Tensor, P: int) -> dict:
    v = GsrGeomView()
    _capi.check(lib.gsr_view_geom(C.c_void_p(geom.data_ptr()), P, C.byref(v)), "gsr_view_geom")
    nblk = (P + 255) // 256
    return dict(
        depths=_slice(geom, v.depths, 4 * P, torch.float32),
        clamped=_slice(geom, v.clamped, 3 * P, torch.uint8).view(P, 3),
        radii=_slice(geom, v.radii, 4 * P, torch.int32),
        means2D=_slice(geom, v.means2D, 8 * P, torch.float32).view(P, 2),
        cov3D=_slice(geom, v.cov3D, 24 * P, torch.float32).view(P, 6),
        conic_opacity=_slice(geom, v.conic_opacity, 16 * P, torch.float32).view(P, 4),
        rgb=_slice(geom, v.rgb, 12 * P, torch.float32).view(P, 3),
        tiles_touched=_slice(geom, v.tiles_touched, 4 * P, torch.int32),
        block_offsets=_slice(geom, v.block_offsets, 4 * (nblk + 1), torch.int32),
        depth_order=_slice(geom, v.depth_order, 4 * P, torch.int32),
        num_rendered=_slice(geom, v.num_rendered, 4, torch.int32),
    )


def binning_view(binning: torch.Tensor, N: int) -> dict:
    v = GsrBinningView()
    _capi.check(lib.gsr_view_binning(C.c_void_p(binning.data_ptr()), N, C.byref(v)), "gsr_view_binning")
    # tile ids are never materialised per pair (run-length binning): see tiles_from_ranges
    return dict(point_list=_slice(binning, v.point_list, 4 * N, torch.int32))


def tiles_from_ranges(ranges: torch.Tensor, N: int) -> torch.Tensor:
    """Tile id of every entry of the sorted list, from the tile ranges: lists are stored back to back in tile order."""
    counts = (ranges[:, 1] - ranges[:, 0]).long()
    tiles = torch.repeat_interleave(torch.arange(ranges.shape[0], device=ranges.device), counts)
    assert tiles.numel() == N, f"tile ranges cover {tiles.numel()} entries, expected {N}"
    return tiles.to(torch.int32)


def sorted_keys(tiles: torch.Tensor, point_list: torch.Tensor, depths: torch.Tensor) -> torch.Tensor:
    """The reference's sorted 64-bit keys (tile << 32 | depth bits, rasterizer_impl.cu:102-105) recomposed from the
    hierarchical sort's outputs, for bit-exact comparison."""
    d = depths.view(torch.int32)[point_list.long()].long() & 0xFFFFFFFF
    return (tiles.long() << 32) | d


def image_view(image: torch.Tensor, W: int, H: int, amr: bool) -> dict:
    v = GsrImageView()
    mode = _capi.GSR_MODE_AMR if amr else _capi.GSR_MODE_BASELINE
    _capi.check(lib.gsr_view_image(C.c_void_p(image.data_ptr()), W, H, mode, C.byref(v)), "gsr_view_image")
    nt = v.tiles_x * v.tiles_y
    out = dict(tiles_x=v.tiles_x, tiles_y=v.tiles_y,
               accum_alpha=_slice(image, v.accum_alpha, 4 * W * H, torch.float32),
               n_contrib=_slice(image, v.n_contrib, 4 * W * H, torch.int32),
               ranges=_slice(image, v.ranges, 8 * nt, torch.int32).view(nt, 2))
    if amr:
        out.update(n_intersections=_slice(image, v.n_intersections, 4 * nt, torch.int32),
                   levels=_slice(image, v.levels, 4 * nt, torch.int32),
                   levels_last=_slice(image, v.levels_last, 4 * nt, torch.int32),
                   levels_current=_slice(image, v.levels_current, 4 * nt, torch.int32),
                   thresholds=_slice(image, v.thresholds, 12, torch.int32))
    return out


def sort_pairs(keys: torch.Tensor, vals: torch.Tensor, end_bit: int):
    """Stand-alone stable radix sort of (int64 key, int32 value) pairs on bits [0, end_bit) (gsr_sort_pairs)."""
    assert keys.is_cuda and vals.is_cuda and keys.dtype == torch.int64 and vals.dtype == torch.int32
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    kt, vt = torch.empty_like(keys), torch.empty_like(vals)
    tb = int(lib.gsr_sort_temp_bytes(n))
    temp = torch.empty(tb, dtype=torch.uint8, device=keys.device)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    with torch.cuda.device(keys.device):
        rc = lib.gsr_sort_pairs(p(keys), p(ko), p(vals), p(vo), p(kt), p(vt), n, int(end_bit), p(temp), tb,
                                C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _capi.check(rc, "gsr_sort_pairs")
    return ko, vo
