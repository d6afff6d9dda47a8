"""PyTorch host side of the `_C` operator surface.

Mirrors, argument for argument, what the reference's pybind module exposes:

* ``rasterize_gaussians``          <- RasterizeGaussiansCUDA          (rasterize_points.cu:35-115, 19 args -> 6-tuple)
* ``rasterize_gaussians_backward`` <- RasterizeGaussiansBackwardCUDA  (rasterize_points.cu:117-196, 21 args -> 8-tuple)
* ``mark_visible``                 <- markVisible                     (rasterize_points.cu:198-217)
* ``rasterize_gaussians_amr``      <- AMR RasterizeGaussiansCUDA      (diff-gaussian-rasterization-amr/rasterize_points.cu:65-202, 25 args)

PyTorch is plumbing only: it owns device memory (outputs and the three growable work buffers, handed to the
library through an allocation callback -- the C analogue of ``resizeFunctional``), supplies the current stream,
and wraps results as tensors.  All compute is in libgsr_b200.so.
"""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import torch

from . import _capi
from ._capi import GsrCamera, GsrGaze, GsrGrads, GsrScene, lib


def _dev_tensor(t: torch.Tensor, name: str, dtype=torch.float32):
    """Absent input (numel == 0, reference: empty CPU tensor -> null data pointer) -> None; else contiguous CUDA."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    return t.contiguous()


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class _Arena:
    """Receives the library's allocation requests and keeps the byte tensors (geom / binning / image).

    One module-level C callback serves every call: the `user` pointer carries a ticket into `_LIVE`, so no per-call
    ctypes thunk is created and -- more importantly -- no reference cycle (arena -> thunk -> bound method -> arena)
    keeps the work buffers alive until Python's cyclic GC happens to run (measured: 35 GB of reserved memory and 2.6
    cudaMalloc per frame with the cyclic version)."""

    _LIVE = {}
    _next_ticket = 1

    def __init__(self, device):
        self.device = device
        self.bufs = {}
        self.ticket = None

    def __enter__(self):
        self.ticket = _Arena._next_ticket
        _Arena._next_ticket += 1
        _Arena._LIVE[self.ticket] = self
        return self

    def __exit__(self, *exc):
        _Arena._LIVE.pop(self.ticket, None)
        return False

    @property
    def user(self):
        return C.c_void_p(self.ticket)

    @staticmethod
    def _bucket(nbytes: int) -> int:
        """Round large requests up to a quarter-octave grid (<= 25 % slack): the binning buffer's size follows the
        per-view pair count, and a handful of distinct sizes keeps torch's caching allocator in a steady state."""
        if nbytes <= (1 << 20):
            return nbytes
        step = 1 << (nbytes.bit_length() - 3)
        return (nbytes + step - 1) // step * step

    def alloc(self, which, nbytes):
        try:
            t = torch.empty(self._bucket(int(nbytes)), dtype=torch.uint8, device=self.device)
        except Exception:  # out of memory etc.: report through the status code
            return None
        self.bufs[int(which)] = t
        return t.data_ptr()

    def get(self, which):
        t = self.bufs.get(which)
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=self.device)


def _alloc_trampoline(user, which, nbytes):
    arena = _Arena._LIVE.get(int(user or 0))
    return arena.alloc(which, nbytes) if arena is not None else None


_ALLOC_CB = _capi.ALLOC_FN(_alloc_trampoline)


def _scene_camera(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                  projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, band=None, sh_rest=None):
    """``sh_rest`` is not None: raw-parameter entry (gsr_scene.raw_params) -- opacity / scales / rotations are the RAW tensors of
    scene/gaussian_model.py (logit, log-scale, un-normalised quaternion), ``sh`` = features_dc [P,1,3], ``sh_rest`` =
    features_rest [P,M-1,3]; the kernels apply the activations on load."""
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P = int(means3D.size(0))
    keep = SimpleNamespace()
    keep.means3D = _dev_tensor(means3D, "means3D") if P else None
    keep.colors = _dev_tensor(colors, "colors_precomp")
    keep.opacity = _dev_tensor(opacity, "opacities")
    keep.scales = _dev_tensor(scales, "scales")
    keep.rotations = _dev_tensor(rotations, "rotations")
    keep.cov3D = _dev_tensor(cov3D_precomp, "cov3D_precomp")
    keep.sh = _dev_tensor(sh, "sh")
    keep.bg = _dev_tensor(background, "bg")
    keep.view = _dev_tensor(viewmatrix, "viewmatrix")
    keep.proj = _dev_tensor(projmatrix, "projmatrix")
    keep.campos = _dev_tensor(campos, "campos")
    M = int(sh.size(1)) if (sh is not None and sh.dim() >= 2 and sh.size(0) != 0) else 0
    keep.sh_rest = None
    raw = sh_rest is not None
    if raw:
        if M != 1:
            raise RuntimeError("raw-parameter entry: sh must be features_dc [P,1,3]")
        keep.sh_rest = _dev_tensor(sh_rest, "features_rest")
        M = 1 + (int(sh_rest.size(1)) if keep.sh_rest is not None else 0)
    sc = GsrScene(P=P, sh_degree=int(degree), sh_coeffs=M, means3D=_p(keep.means3D), shs=_p(keep.sh),
                  colors_precomp=_p(keep.colors), opacities=_p(keep.opacity), scales=_p(keep.scales),
                  rotations=_p(keep.rotations), cov3D_precomp=_p(keep.cov3D), scale_modifier=float(scale_modifier),
                  raw_params=1 if raw else 0, shs_rest=_p(keep.sh_rest))
    cam = GsrCamera(width=int(image_width), height=int(image_height), tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy),
                    background=_p(keep.bg), viewmatrix=_p(keep.view), projmatrix=_p(keep.proj), cam_pos=_p(keep.campos),
                    prefiltered=1 if prefiltered else 0, band_row0=int(band[0]) if band else 0,
                    band_row1=int(band[1]) if band else 0)
    return sc, cam, keep, P, M


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ----------------------------------------------------------------------------------------------------------------
# baseline
# ----------------------------------------------------------------------------------------------------------------
def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug, band=None, sh_rest=None):
    """``sh_rest`` (extension): raw-parameter entry, see ``_scene_camera``.
    ``band`` = (row0, row1): bin and blend only tile rows [row0, row1) (extension for multi-GPU tile bands;
    rows of out_color outside the band are left unwritten).  None = the reference's whole-frame behaviour."""
    sc, cam, keep, P, _ = _scene_camera(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                                        image_width, sh, degree, campos, prefiltered, band, sh_rest)
    dev = means3D.device if means3D.is_cuda else torch.device("cuda")
    with torch.cuda.device(dev):
        out_color = torch.empty((3, int(image_height), int(image_width)), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        rendered = C.c_int(0)
        with _Arena(dev) as arena:
            rc = lib.gsr_forward(C.byref(sc), C.byref(cam), _ALLOC_CB, arena.user, _p(out_color), _p(radii) if P else None,
                                 C.byref(rendered), 1 if debug else 0, _stream())
        _capi.check(rc, "gsr_forward")
    return (int(rendered.value), out_color, radii, arena.get(_capi.GSR_BUF_GEOM), arena.get(_capi.GSR_BUF_BINNING),
            arena.get(_capi.GSR_BUF_IMAGE))


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug, band=None, compact_sh=False,
                                 between_phases=None, preprocess_chunks=1, after_chunk=None, dense_out=None, push=None, sh_rest=None):
    """``compact_sh=True`` (extension, view-parallel training): slot 5 of the returned tuple is the [P,3] colour
    gradient after the SH clamp mask instead of the [P,M,3] SH gradient, which is then not written at all;
    ``sh_grad_from_views`` rebuilds the SH gradient summed over views from those.
    ``between_phases(dL_dcolors_clamped)`` (with compact_sh): called after the blend-stage backward has been queued,
    when that tensor is final in stream order -- the hook where a view-parallel caller starts its all-gather so that
    it overlaps the preprocess-stage backward (gsr_backward_phase).
    ``preprocess_chunks`` > 1 (with between_phases): the preprocess stage runs as that many launches over consecutive
    ranges of Gaussians (gsr_backward_preprocess_range) and ``after_chunk(first, count, (dL_dmeans3D, dL_dopacity,
    dL_dscales, dL_drotations))`` is called after each with views of that range's rows, final in stream order -- where
    the caller starts reducing them while the next range is still being computed.
    ``dense_out``: caller-owned float32 buffer of at least 11 * ceil4(P) elements that receives the four dense gradients
    (means3D, opacity, scales, rotations, back to back) instead of a fresh allocation -- a view-parallel caller passes
    peer-mapped memory here so that dist.PeerAllReduce sums them in place over NVLink.
    ``push`` = (peer_bases_dev_ptr, staging_offset_bytes, rank, world, rows_per_rank, epoch)
    (= dist.PeerAllReduce.push_args): fused exchange (gsr_grads.push_*) -- the preprocess stage stores the 11 dense floats of
    every Gaussian into the OWNER rank's staging area (self-validating 128-byte lines) instead of writing the four dense tensors;
    dist.PeerAllReduce.owner_sum, queued beside it, fills them (in ``dense_out``) on every rank."""
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    sc, cam, keep, P, M = _scene_camera(background, means3D, colors, torch.empty(0), scales, rotations, scale_modifier,
                                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree,
                                        campos, False, band, sh_rest)
    dev = means3D.device
    f32 = dict(dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        # the four "dense" parameter gradients are carved out of one allocation, back to back (16-byte aligned each for
        # P % 4 == 0): a view-parallel caller all-reduces them as ONE in-place NCCL call (dist.GradientAllReduce
        # recognises adjacent tensors) instead of packing them into a bucket and back
        P4 = (P + 3) // 4 * 4
        if dense_out is None:
            flat = torch.empty((11 * P4,), **f32)
        else:
            if (not dense_out.is_cuda or dense_out.dtype != torch.float32 or not dense_out.is_contiguous()
                    or dense_out.numel() < 11 * P4 or dense_out.data_ptr() % 16):
                raise RuntimeError("dense_out must be a contiguous, 16-byte aligned float32 CUDA buffer of >= 11*ceil4(P) elements")
            flat = dense_out.view(-1)[:11 * P4]
        dL_dmeans3D = flat[0:3 * P].view(P, 3)
        dL_dopacity = flat[3 * P4:3 * P4 + P].view(P, 1)
        dL_dscales = flat[4 * P4:4 * P4 + 3 * P].view(P, 3)
        dL_drotations = flat[7 * P4:7 * P4 + 4 * P].view(P, 4)
        dL_dmeans2D = torch.empty((P, 3), **f32)
        dL_dcolors = torch.empty((P, 3), **f32)
        dL_dconic = torch.empty((P, 2, 2), **f32)
        dL_dcov3D = torch.empty((P, 6), **f32)
        compact = bool(compact_sh) and M > 0
        dL_dsh = torch.empty((P, 3), **f32) if compact else torch.empty((P, M, 3), **f32)
        if P != 0:
            scratch = torch.empty((P, 12), **f32)
            dpix = _dev_tensor(dL_dout_color, "dL_dout_color")
            radii_t = _dev_tensor(radii, "radii", torch.int32)
            grads = GsrGrads(dL_dmeans2D=_p(dL_dmeans2D), dL_dcolors=_p(dL_dcolors), dL_dopacity=_p(dL_dopacity),
                             dL_dmeans3D=_p(dL_dmeans3D), dL_dcov3D=_p(dL_dcov3D),
                             dL_dsh=_p(dL_dsh) if (M and not compact) else None,
                             dL_dscales=_p(dL_dscales), dL_drotations=_p(dL_drotations), dL_dconic=_p(dL_dconic),
                             scratch=_p(scratch), dL_dcolors_clamped=_p(dL_dsh) if compact else None)
            if push is not None:
                grads.push_bases = C.c_void_p(int(push[0])); grads.push_offset = int(push[1]); grads.push_rank = int(push[2])
                grads.push_world = int(push[3]); grads.push_rows_per_rank = int(push[4])
                grads.push_epoch = int(push[5]) & 0xffffffff
            bufs = (_p(geomBuffer.contiguous()), _p(binningBuffer.contiguous()), _p(imageBuffer.contiguous()))
            if compact and between_phases is not None:
                rc = lib.gsr_backward_phase(C.byref(sc), C.byref(cam), _p(radii_t), *bufs, int(R), _p(dpix), C.byref(grads),
                                            _capi.GSR_BWD_BLEND, 1 if debug else 0, _stream())
                _capi.check(rc, "gsr_backward_phase(blend)")
                between_phases(dL_dsh)
                nch = max(1, min(int(preprocess_chunks), (P + 255) // 256))
                if nch == 1:
                    rc = lib.gsr_backward_phase(C.byref(sc), C.byref(cam), _p(radii_t), *bufs, int(R), _p(dpix), C.byref(grads),
                                                _capi.GSR_BWD_PREPROCESS, 1 if debug else 0, _stream())
                    _capi.check(rc, "gsr_backward_phase(preprocess)")
                    if after_chunk is not None:
                        after_chunk(0, P, (dL_dmeans3D, dL_dopacity, dL_dscales, dL_drotations))
                else:
                    per = ((P + nch - 1) // nch + 255) // 256 * 256
                    for first in range(0, P, per):
                        count = min(per, P - first)
                        rc = lib.gsr_backward_preprocess_range(C.byref(sc), C.byref(cam), _p(radii_t), *bufs, int(R), C.byref(grads),
                                                               first, count, 1 if debug else 0, _stream())
                        _capi.check(rc, "gsr_backward_preprocess_range")
                        if after_chunk is not None:
                            after_chunk(first, count, tuple(t[first:first + count] for t in (dL_dmeans3D, dL_dopacity, dL_dscales, dL_drotations)))
            else:
                rc = lib.gsr_backward(C.byref(sc), C.byref(cam), _p(radii_t), *bufs, int(R), _p(dpix),
                                      C.byref(grads), 1 if debug else 0, _stream())
                _capi.check(rc, "gsr_backward")
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)


def sh_grad_from_views(means3D, cam_positions, dL_dcolors_clamped, degree, M):
    """SH gradient summed over views from the per-view clamped colour gradients (gsr_sh_grad_from_views).
    cam_positions [K,3], dL_dcolors_clamped [K,P,3] -> dL_dsh [P,M,3]."""
    P, K = int(means3D.size(0)), int(cam_positions.size(0))
    m = _dev_tensor(means3D, "means3D")
    cams = _dev_tensor(cam_positions, "cam_positions")
    g = _dev_tensor(dL_dcolors_clamped, "dL_dcolors_clamped")
    if tuple(g.shape) != (K, P, 3):
        raise RuntimeError(f"dL_dcolors_clamped must have dimensions ({K}, {P}, 3)")
    with torch.cuda.device(m.device):
        out = torch.empty((P, int(M), 3), dtype=torch.float32, device=m.device)
        rc = lib.gsr_sh_grad_from_views(P, int(degree), int(M), _p(m), K, _p(cams), _p(g), _p(out), _stream())
        _capi.check(rc, "gsr_sh_grad_from_views")
    return out


def last_counts(geomBuffer, P):
    """(pairs, runs, visible Gaussians) of the forward this thread ran last on ``geomBuffer`` (gsr_last_counts): host
    integers the forward already holds -- no device round trip."""
    n, r, v = C.c_longlong(0), C.c_longlong(0), C.c_longlong(0)
    rc = lib.gsr_last_counts(_p(geomBuffer), int(P), C.byref(n), C.byref(r), C.byref(v))
    _capi.check(rc, "gsr_last_counts")
    return int(n.value), int(r.value), int(v.value)


def blend_stats(geomBuffer, binningBuffer, imageBuffer, P, image_height, image_width):
    """(pair_tests_forward, hits, pair_tests_backward) of a baseline frame just built by rasterize_gaussians: the
    algorithmic work of the blend stage, counted by one extra kernel (gsr_blend_stats).  Measurement only."""
    dev = geomBuffer.device
    with torch.cuda.device(dev):
        tot = torch.empty(3, dtype=torch.int64, device=dev)
        rc = lib.gsr_blend_stats(int(P), int(image_width), int(image_height), _p(geomBuffer), _p(binningBuffer), _p(imageBuffer),
                                 _p(tot), _stream())
        _capi.check(rc, "gsr_blend_stats")
    return tuple(int(x) for x in tot.tolist())


def pack_meta_words(P):
    return int(lib.gsr_pack_meta_words(int(P)))


def pack_visible_rows(rows, radii, cap, out_rows=None, out_meta=None):
    """rows [P,3] float32, radii [P] int32 -> (packed [cap,3], meta [pack_meta_words(P)] int32): the rows of the
    Gaussians with radii > 0 in index order plus the mask / prefix words that locate them (gsr_pack_visible_rows)."""
    P = int(rows.size(0))
    r = _dev_tensor(rows, "rows")
    rad = radii if (radii.dtype == torch.int32 and radii.is_contiguous()) else radii.to(torch.int32).contiguous()
    if not rad.is_cuda or tuple(r.shape) != (P, 3) or int(rad.numel()) != P:
        raise RuntimeError("pack_visible_rows: rows must be [P,3] float32 and radii [P] on the GPU")
    cap = int(cap)
    with torch.cuda.device(r.device):
        packed = out_rows if out_rows is not None else torch.empty((cap, 3), dtype=torch.float32, device=r.device)
        meta = out_meta if out_meta is not None else torch.empty((pack_meta_words(P),), dtype=torch.int32, device=r.device)
        if packed.numel() < 3 * cap or meta.numel() < pack_meta_words(P) or not packed.is_contiguous() or not meta.is_contiguous():
            raise RuntimeError("pack_visible_rows: output buffers too small")
        tb = int(lib.gsr_pack_temp_bytes(P))
        temp = torch.empty((max(tb, 4),), dtype=torch.uint8, device=r.device)
        rc = lib.gsr_pack_visible_rows(P, _p(rad), _p(r), cap, _p(packed), _p(meta), _p(temp), tb, _stream())
        _capi.check(rc, "gsr_pack_visible_rows")
    return packed, meta


def sh_grad_from_packed_views(means3D, cam_positions, meta, rows, degree, M, out=None):
    """As sh_grad_from_views with every view's colour gradient packed by pack_visible_rows:
    cam_positions [K,3], meta [K,L] int32 (L >= pack_meta_words(P)), rows [K,cap,3] -> dL_dsh [P,M,3]
    (``out``: a caller-owned contiguous float32 [P,M,3] tensor to write instead of a fresh allocation)."""
    P, K = int(means3D.size(0)), int(cam_positions.size(0))
    m = _dev_tensor(means3D, "means3D")
    cams = _dev_tensor(cam_positions, "cam_positions")
    g = _dev_tensor(rows, "rows")
    if meta.dtype != torch.int32 or not meta.is_cuda or not meta.is_contiguous() or meta.dim() != 2 or int(meta.size(0)) != K:
        raise RuntimeError("meta must be a contiguous int32 [K, L] GPU tensor")
    if g.dim() != 3 or int(g.size(0)) != K or int(g.size(2)) != 3:
        raise RuntimeError(f"rows must have dimensions ({K}, cap, 3)")
    with torch.cuda.device(m.device):
        if out is None:
            out = torch.empty((P, int(M), 3), dtype=torch.float32, device=m.device)
        elif tuple(out.shape) != (P, int(M), 3) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != m.device:
            raise RuntimeError(f"out must be a contiguous float32 ({P}, {int(M)}, 3) tensor on the same device")
        rc = lib.gsr_sh_grad_from_packed_views(P, int(degree), int(M), _p(m), K, _p(cams), _p(meta), int(meta.size(1)),
                                               _p(g), int(g.size(1)), _p(out), _stream())
        _capi.check(rc, "gsr_sh_grad_from_packed_views")
    return out


def mark_visible(means3D, viewmatrix, projmatrix):
    P = int(means3D.size(0))
    present = torch.empty((P,), dtype=torch.bool, device=means3D.device)
    if P != 0:
        m = _dev_tensor(means3D, "means3D")
        v = _dev_tensor(viewmatrix, "viewmatrix")
        pr = _dev_tensor(projmatrix, "projmatrix")
        with torch.cuda.device(means3D.device):
            rc = lib.gsr_mark_visible(P, _p(m), _p(v), _p(pr), _p(present), _stream())
            _capi.check(rc, "gsr_mark_visible")
    return present


# ----------------------------------------------------------------------------------------------------------------
# AMR / foveated
# ----------------------------------------------------------------------------------------------------------------
def rasterize_gaussians_amr(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                            viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                            prefiltered, foveaStep, out_color_precomp, geomBuffer_precomp, binningBuffer_precomp,
                            imageBuffer_precomp, interpolate_image, debug, gaze=None, band=None):
    """``gaze`` = (cx, cy, (r1, r2, r3, r4)) enables the gaze-driven level cap (extension; None = reference).
    ``band`` = (row0, row1) restricts binning and blending to 32-px tile rows [row0, row1) (extension); rows of out_color
    outside the band are left unwritten (only the band's rows are cleared and rendered)."""
    step = int(foveaStep)
    if step >= 1 and means3D.is_cuda and means3D.dim() == 2:
        # progressive rounds 1..4 read nothing of the scene but its size (everything was built by step 0): marshal only what the
        # library validates -- five calls per frame make the host side of this protocol the bottleneck
        P = int(means3D.size(0))
        keep = SimpleNamespace(means3D=means3D, bg=_dev_tensor(background, "bg"), view=_dev_tensor(viewmatrix, "viewmatrix"),
                               proj=_dev_tensor(projmatrix, "projmatrix"), campos=_dev_tensor(campos, "campos"), sh=sh if sh.numel() else None,
                               colors=colors if colors.numel() else None)
        M = int(sh.size(1)) if (sh is not None and sh.dim() >= 2 and sh.size(0) != 0) else 0
        dummy = _p(means3D)
        sc = GsrScene(P=P, sh_degree=int(degree), sh_coeffs=M, means3D=dummy, shs=_p(keep.sh), colors_precomp=_p(keep.colors), opacities=dummy,
                      scales=dummy, rotations=dummy, cov3D_precomp=None, scale_modifier=float(scale_modifier), raw_params=0, shs_rest=None)
        cam = GsrCamera(width=int(image_width), height=int(image_height), tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy),
                        background=_p(keep.bg), viewmatrix=_p(keep.view), projmatrix=_p(keep.proj), cam_pos=_p(keep.campos),
                        prefiltered=1 if prefiltered else 0, band_row0=int(band[0]) if band else 0, band_row1=int(band[1]) if band else 0)
    else:
        sc, cam, keep, P, _ = _scene_camera(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                            cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                                            image_width, sh, degree, campos, prefiltered, band)
    dev = means3D.device if means3D.is_cuda else torch.device("cuda")
    gz = _gaze_struct(gaze)
    with torch.cuda.device(dev):
        out_color = torch.empty((3, int(image_height), int(image_width)), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        if step >= 1:
            radii.zero_()          # steps >= 1 do not preprocess; the reference returns its zero-filled radii
        rendered = C.c_int(0)
        precomp = _dev_tensor(out_color_precomp, "out_color_precomp")
        gp = geomBuffer_precomp if (geomBuffer_precomp is not None and geomBuffer_precomp.numel()) else None
        bp = binningBuffer_precomp if (binningBuffer_precomp is not None and binningBuffer_precomp.numel()) else None
        ip = imageBuffer_precomp if (imageBuffer_precomp is not None and imageBuffer_precomp.numel()) else None
        with _Arena(dev) as arena:
            rc = lib.gsr_amr_forward(C.byref(sc), C.byref(cam), step, _p(precomp), _p(gp), _p(bp), _p(ip),
                                     1 if interpolate_image else 0, C.byref(gz) if gz is not None else None, _ALLOC_CB,
                                     arena.user, _p(out_color), _p(radii) if (P and step < 1) else None, C.byref(rendered),
                                     1 if debug else 0, _stream())
        _capi.check(rc, "gsr_amr_forward")
    if step > 0:   # the returned buffers ARE the passed-in ones (A/rasterize_points.cu:182-191)
        return int(rendered.value), out_color, radii, geomBuffer_precomp, binningBuffer_precomp, imageBuffer_precomp
    return (int(rendered.value), out_color, radii, arena.get(_capi.GSR_BUF_GEOM), arena.get(_capi.GSR_BUF_BINNING),
            arena.get(_capi.GSR_BUF_IMAGE))


def _gaze_struct(gaze):
    if gaze is None:
        return None
    return GsrGaze(cx=float(gaze[0]), cy=float(gaze[1]), radii=(C.c_float * 4)(*[float(r) for r in gaze[2]]))


def amr_relevel(imageBuffer, image_height, image_width):
    """Recompute percentile thresholds and tile levels from the per-tile pair counts stored in ``imageBuffer``
    (after they were summed over tile-band ranks in place).  gsr_amr_relevel."""
    with torch.cuda.device(imageBuffer.device):
        rc = lib.gsr_amr_relevel(_p(imageBuffer), int(image_width), int(image_height), _stream())
    _capi.check(rc, "gsr_amr_relevel")


def amr_render_prebuilt(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                        projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                        geomBuffer, binningBuffer, imageBuffer, interpolate_image, debug=False, gaze=None, band=None):
    """Single-pass AMR render (all lattice rounds) from buffers built by ``rasterize_gaussians_amr(foveaStep=0)``.
    gsr_amr_render_prebuilt; used by the tile-band driver after the global re-levelling."""
    sc, cam, keep, P, _ = _scene_camera(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                                        image_width, sh, degree, campos, prefiltered, band)
    dev = means3D.device
    gz = _gaze_struct(gaze)
    with torch.cuda.device(dev):
        out_color = torch.empty((3, int(image_height), int(image_width)), dtype=torch.float32, device=dev)
        rc = lib.gsr_amr_render_prebuilt(C.byref(sc), C.byref(cam), _p(geomBuffer), _p(binningBuffer), _p(imageBuffer),
                                         1 if interpolate_image else 0, C.byref(gz) if gz is not None else None,
                                         _p(out_color), 1 if debug else 0, _stream())
        _capi.check(rc, "gsr_amr_render_prebuilt")
    return out_color


# `_C`-shaped namespaces, one per reference extension module
C_baseline = SimpleNamespace(rasterize_gaussians=rasterize_gaussians,
                             rasterize_gaussians_backward=rasterize_gaussians_backward,
                             mark_visible=mark_visible)
C_amr = SimpleNamespace(rasterize_gaussians=rasterize_gaussians_amr,
                        rasterize_gaussians_backward=rasterize_gaussians_backward,
                        mark_visible=mark_visible)
