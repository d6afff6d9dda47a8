"""Driver of the UNMODIFIED reference CUDA rasterizer built into oracle/_ref/ (see oracle/Makefile, ref_shim.cu).

TEST / BASELINE INFRASTRUCTURE: used by the GPU parity tests, oracle/make_golden.py and bench.py's reference arm.
It needs a GPU and the prebuilt oracle/_ref/*.so (built in the container that has /root/reference; the .so files
travel to the GPU box with the repo snapshot).  Nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIBS = {False: os.path.join(HERE, "_ref", "libref_raster.so"), True: os.path.join(HERE, "_ref", "libref_raster_amr.so")}
ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_int, C.c_size_t)


def available(amr: bool = False) -> bool:
    return os.path.exists(LIBS[amr])


class _GeomView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("depths", "clamped", "internal_radii", "means2D", "cov3D", "conic_opacity",
                                          "rgb", "point_offsets", "tiles_touched")]


class _BinView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("keys_unsorted", "keys", "list_unsorted", "list")]


class _ImgView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("ranges", "n_contrib", "accum_alpha", "n_intersections", "levels",
                                          "levels_last", "levels_current")]


def _p(t):
    return None if (t is None or t.numel() == 0) else C.c_void_p(t.data_ptr())


def _slice(buf: torch.Tensor, ptr, nbytes: int, dtype):
    """View `nbytes` of the byte tensor `buf` starting at raw device address `ptr` as `dtype`."""
    off = int(ptr) - buf.data_ptr()
    assert 0 <= off and off + nbytes <= buf.numel(), "reference view outside its buffer"
    return buf[off:off + nbytes].view(dtype)


class RefRasterizer:
    def __init__(self, amr: bool = False):
        if not available(amr):
            raise RuntimeError(f"{LIBS[amr]} is missing: run `make -C oracle ref` where /root/reference exists")
        self.amr = amr
        self.lib = C.CDLL(LIBS[amr])
        self.lib.ref_last_error.restype = C.c_char_p
        self.lib.ref_forward.restype = C.c_int
        self.lib.ref_backward.restype = C.c_int

    # -- forward: mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-115 / AMR :65-202): allocates like the binding
    def forward(self, bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered=False, fovea_step=-1,
                out_color_precomp=None, geom_pre=None, bin_pre=None, img_pre=None, interpolate=False, debug=False):
        dev = means3D.device
        P = means3D.shape[0]
        out_color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        bufs = {}

        def alloc(_u, which, n):
            t = torch.empty(int(n), dtype=torch.uint8, device=dev)
            bufs[int(which)] = t
            return t.data_ptr()

        cb = ALLOC_FN(alloc)
        M = sh.shape[1] if (sh is not None and sh.numel() and sh.dim() >= 2) else 0
        rc = self.lib.ref_forward(
            C.c_int(P), C.c_int(degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H), _p(means3D), _p(sh), _p(colors),
            _p(opacity), _p(scales), C.c_float(scale_modifier), _p(rotations), _p(cov3D_precomp), _p(viewmatrix),
            _p(projmatrix), _p(campos), C.c_float(tan_fovx), C.c_float(tan_fovy), C.c_int(1 if prefiltered else 0),
            C.c_int(fovea_step), _p(out_color_precomp), _p(geom_pre), _p(bin_pre), _p(img_pre), _p(out_color), _p(radii),
            C.c_int(1 if interpolate else 0), C.c_int(1 if debug else 0), cb, None)
        if rc < 0:
            raise RuntimeError(f"reference forward failed ({rc}): {self.lib.ref_last_error().decode()}")
        if self.amr and fovea_step > 0:
            return rc, out_color, radii, geom_pre, bin_pre, img_pre
        return rc, out_color, radii, bufs.get(0), bufs.get(1), bufs.get(2)

    def backward(self, bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                 tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geom, R, binning, image, debug=False):
        dev = means3D.device
        P = means3D.shape[0]
        H, W = dL_dout_color.shape[1], dL_dout_color.shape[2]
        M = sh.shape[1] if (sh is not None and sh.numel() and sh.dim() >= 2) else 0
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)  # noqa: E731  (rasterize_points.cu:151-159)
        g = dict(dL_dmeans3D=z(P, 3), dL_dmeans2D=z(P, 3), dL_dcolors=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1),
                 dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
        rc = self.lib.ref_backward(
            C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(R), _p(bg), C.c_int(W), C.c_int(H), _p(means3D), _p(sh),
            _p(colors), _p(scales), C.c_float(scale_modifier), _p(rotations), _p(cov3D_precomp), _p(viewmatrix),
            _p(projmatrix), _p(campos), C.c_float(tan_fovx), C.c_float(tan_fovy), _p(radii), _p(geom), _p(binning),
            _p(image), _p(dL_dout_color.contiguous()), _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]),
            _p(g["dL_dcolors"]), _p(g["dL_dmeans3D"]), _p(g["dL_dcov3D"]), _p(g["dL_dsh"]) if M else _p(g["dL_dcolors"]),
            _p(g["dL_dscales"]), _p(g["dL_drotations"]), C.c_int(1 if debug else 0))
        if rc != 0:
            raise RuntimeError(f"reference backward failed: {self.lib.ref_last_error().decode()}")
        return g

    # -- decoded intermediate buffers (the reference's own fromChunk layout)
    def parse(self, geom, binning, image, P, N, W, H):
        gv, bv, iv = _GeomView(), _BinView(), _ImgView()
        self.lib.ref_view_geom(C.c_void_p(geom.data_ptr()), C.c_int(P), C.byref(gv))
        self.lib.ref_view_binning(C.c_void_p(binning.data_ptr()), C.c_int(N), C.byref(bv))
        self.lib.ref_view_image(C.c_void_p(image.data_ptr()), C.c_int(W), C.c_int(H), C.byref(iv))
        block = 32 if self.amr else 16
        nt = ((W + block - 1) // block) * ((H + block - 1) // block)
        out = dict(
            depths=_slice(geom, gv.depths, 4 * P, torch.float32),
            clamped=_slice(geom, gv.clamped, 3 * P, torch.uint8).view(P, 3),
            means2D=_slice(geom, gv.means2D, 8 * P, torch.float32).view(P, 2),
            cov3D=_slice(geom, gv.cov3D, 24 * P, torch.float32).view(P, 6),
            conic_opacity=_slice(geom, gv.conic_opacity, 16 * P, torch.float32).view(P, 4),
            rgb=_slice(geom, gv.rgb, 12 * P, torch.float32).view(P, 3),
            point_offsets=_slice(geom, gv.point_offsets, 4 * P, torch.int32),
            tiles_touched=_slice(geom, gv.tiles_touched, 4 * P, torch.int32),
            keys=_slice(binning, bv.keys, 8 * N, torch.int64),
            keys_unsorted=_slice(binning, bv.keys_unsorted, 8 * N, torch.int64),
            point_list=_slice(binning, bv.list, 4 * N, torch.int32),
            ranges=_slice(image, iv.ranges, 8 * nt, torch.int32).view(nt, 2),
            n_contrib=_slice(image, iv.n_contrib, 4 * W * H, torch.int32),
            accum_alpha=_slice(image, iv.accum_alpha, 4 * W * H, torch.float32),
        )
        if self.amr:
            out.update(
                n_intersections=_slice(image, iv.n_intersections, 4 * nt, torch.int32),
                levels=_slice(image, iv.levels, 4 * nt, torch.int32),
                levels_last=_slice(image, iv.levels_last, 4 * nt, torch.int32),
                levels_current=_slice(image, iv.levels_current, 4 * nt, torch.int32))
        return out
