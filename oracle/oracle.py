"""ctypes/numpy driver of the CPU oracle (oracle/gsr_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module; nothing under
gaussian_splatting_with_eye_tracking_b200/ does.  It orchestrates the restated stages in the reference's order
(CudaRasterizer::Rasterizer::forward, cuda_rasterizer/rasterizer_impl.cu:198-336; backward :340-434; AMR fork
diff-gaussian-rasterization-amr/cuda_rasterizer/rasterizer_impl.cu:296-694) on numpy arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "gsr_oracle.c")
LIB = os.path.join(HERE, "_build", "libgsr_oracle.so")

GCC_FLAGS = ["-O2", "-std=c11", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared",
             "-fvisibility=hidden", "-Wno-misleading-indentation"]


def build(force: bool = False) -> str:
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    cmd = ["gcc", *GCC_FLAGS, "-o", LIB + ".tmp", SRC, "-lm"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed building the oracle:\n" + res.stderr)
    os.replace(LIB + ".tmp", LIB)
    return LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.orc_scan.restype = C.c_longlong
        _lib.orc_higher_msb.restype = C.c_uint32
        _lib.orc_higher_msb.argtypes = [C.c_uint32]
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(lib().orc_num_threads())


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def higher_msb(n: int) -> int:
    return int(lib().orc_higher_msb(C.c_uint32(n)))


class Scene:
    """Inputs of the rasterizer as numpy arrays (same tensor contract as the `_C` call)."""

    def __init__(self, means3D, opacities, scales=None, rotations=None, shs=None, colors_precomp=None,
                 cov3D_precomp=None, sh_degree=0, scale_modifier=1.0):
        self.means3D = _f(means3D)
        self.opacities = _f(opacities).reshape(-1)
        self.scales = _f(scales)
        self.rotations = _f(rotations)
        self.shs = _f(shs)
        self.colors_precomp = _f(colors_precomp)
        self.cov3D_precomp = _f(cov3D_precomp)
        self.sh_degree = int(sh_degree)
        self.scale_modifier = float(scale_modifier)
        self.P = int(self.means3D.shape[0])
        self.M = int(self.shs.shape[1]) if self.shs is not None else 0


class Camera:
    def __init__(self, width, height, tanfovx, tanfovy, viewmatrix, projmatrix, campos, bg):
        self.W, self.H = int(width), int(height)
        self.tanfovx, self.tanfovy = float(tanfovx), float(tanfovy)
        self.view = _f(viewmatrix).reshape(16)      # as stored by the Python side (transposed) = column-major for the kernels
        self.proj = _f(projmatrix).reshape(16)
        self.campos = _f(campos).reshape(3)
        self.bg = _f(bg).reshape(3)


class ForwardState:
    """Everything the reference keeps in its geom / binning / image buffers, as named arrays."""


def preprocess(sc: Scene, cam: Camera, block: int) -> ForwardState:
    L = lib()
    P = sc.P
    st = ForwardState()
    st.block = block
    st.gx, st.gy = (cam.W + block - 1) // block, (cam.H + block - 1) // block
    st.radii = np.zeros(P, np.int32)
    st.means2D = np.zeros((P, 2), np.float32)
    st.depths = np.zeros(P, np.float32)
    st.cov3D = np.zeros((P, 6), np.float32)
    st.rgb = np.zeros((P, 3), np.float32)
    st.conic_opacity = np.zeros((P, 4), np.float32)
    st.tiles_touched = np.zeros(P, np.uint32)
    st.clamped = np.zeros((P, 3), np.uint8)
    L.orc_preprocess(C.c_int(P), C.c_int(sc.sh_degree), C.c_int(sc.M), _ptr(sc.means3D), _ptr(sc.scales),
                     C.c_float(sc.scale_modifier), _ptr(sc.rotations), _ptr(sc.opacities), _ptr(sc.shs),
                     _ptr(sc.cov3D_precomp), _ptr(sc.colors_precomp), _ptr(cam.view), _ptr(cam.proj), _ptr(cam.campos),
                     C.c_int(cam.W), C.c_int(cam.H), C.c_float(cam.tanfovx), C.c_float(cam.tanfovy), C.c_int(block),
                     _ptr(st.radii), _ptr(st.means2D), _ptr(st.depths), _ptr(st.cov3D), _ptr(st.rgb),
                     _ptr(st.conic_opacity), _ptr(st.tiles_touched), _ptr(st.clamped))
    return st


def bin_tiles(st: ForwardState) -> None:
    """scan -> duplicateWithKeys -> SortPairs -> identifyTileRanges."""
    L = lib()
    P = st.radii.shape[0]
    st.point_offsets = np.zeros(P, np.uint32)
    N = int(L.orc_scan(C.c_int(P), _ptr(st.tiles_touched), _ptr(st.point_offsets)))
    st.num_rendered = N
    st.keys_unsorted = np.zeros(N, np.uint64)
    st.vals_unsorted = np.zeros(N, np.uint32)
    L.orc_duplicate(C.c_int(P), _ptr(st.means2D), _ptr(st.depths), _ptr(st.point_offsets), _ptr(st.radii),
                    C.c_int(st.block), C.c_int(st.gx), C.c_int(st.gy), _ptr(st.keys_unsorted), _ptr(st.vals_unsorted))
    st.end_bit = 32 + higher_msb(st.gx * st.gy)
    st.keys = np.zeros(N, np.uint64)
    st.point_list = np.zeros(N, np.uint32)
    L.orc_sort(C.c_longlong(N), _ptr(st.keys_unsorted), _ptr(st.vals_unsorted), _ptr(st.keys), _ptr(st.point_list),
               C.c_int(st.end_bit))
    st.ranges = np.zeros((st.gx * st.gy, 2), np.uint32)
    L.orc_ranges(C.c_longlong(N), _ptr(st.keys), _ptr(st.ranges), C.c_int(st.gx * st.gy))


def sort_pairs(keys, vals, end_bit):
    keys = np.ascontiguousarray(keys, np.uint64)
    vals = np.ascontiguousarray(vals, np.uint32)
    ko, vo = np.zeros_like(keys), np.zeros_like(vals)
    lib().orc_sort(C.c_longlong(keys.shape[0]), _ptr(keys), _ptr(vals), _ptr(ko), _ptr(vo), C.c_int(end_bit))
    return ko, vo


def forward(sc: Scene, cam: Camera) -> ForwardState:
    """Baseline rasterizer forward (16x16 tiles)."""
    st = preprocess(sc, cam, 16)
    bin_tiles(st)
    st.colors = sc.colors_precomp if sc.colors_precomp is not None else st.rgb
    st.final_T = np.zeros(cam.W * cam.H, np.float32)
    st.n_contrib = np.zeros(cam.W * cam.H, np.uint32)
    st.out_color = np.zeros((3, cam.H, cam.W), np.float32)
    lib().orc_blend_forward(C.c_int(cam.W), C.c_int(cam.H), C.c_int(16), _ptr(st.ranges), _ptr(st.point_list),
                            _ptr(st.means2D), _ptr(st.colors), _ptr(st.conic_opacity), _ptr(cam.bg), _ptr(st.final_T),
                            _ptr(st.n_contrib), _ptr(st.out_color))
    return st


def backward(sc: Scene, cam: Camera, st: ForwardState, dL_dpix) -> dict:
    """Rasterizer::backward: blend backward, then cov2D / preprocess backward.  Returns the 8 reference outputs."""
    L = lib()
    P, M = sc.P, sc.M
    dL_dpix = _f(dL_dpix)
    g = dict(dL_dmeans2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
             dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolors=np.zeros((P, 3), np.float32),
             dL_dmeans3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
             dL_dsh=np.zeros((P, M, 3), np.float32), dL_dscales=np.zeros((P, 3), np.float32),
             dL_drotations=np.zeros((P, 4), np.float32))
    L.orc_blend_backward(C.c_int(P), C.c_int(cam.W), C.c_int(cam.H), C.c_int(16), _ptr(st.ranges), _ptr(st.point_list),
                         _ptr(cam.bg), _ptr(st.means2D), _ptr(st.conic_opacity), _ptr(st.colors), _ptr(st.final_T),
                         _ptr(st.n_contrib), _ptr(dL_dpix), _ptr(g["dL_dmeans2D"]), _ptr(g["dL_dconic"]),
                         _ptr(g["dL_dopacity"]), _ptr(g["dL_dcolors"]))
    cov3D = sc.cov3D_precomp if sc.cov3D_precomp is not None else st.cov3D
    shs = sc.shs if sc.colors_precomp is None else None
    scales = sc.scales if sc.cov3D_precomp is None else None
    L.orc_preprocess_backward(C.c_int(P), C.c_int(sc.sh_degree), C.c_int(M), _ptr(sc.means3D), _ptr(st.radii), _ptr(shs),
                              _ptr(st.clamped), _ptr(scales), _ptr(sc.rotations), C.c_float(sc.scale_modifier),
                              _ptr(cov3D), _ptr(cam.view), _ptr(cam.proj), C.c_int(cam.W), C.c_int(cam.H),
                              C.c_float(cam.tanfovx), C.c_float(cam.tanfovy), _ptr(cam.campos), _ptr(g["dL_dmeans2D"]),
                              _ptr(g["dL_dconic"]), _ptr(g["dL_dmeans3D"]), _ptr(g["dL_dcolors"]), _ptr(g["dL_dcov3D"]),
                              _ptr(g["dL_dsh"]), _ptr(g["dL_dscales"]), _ptr(g["dL_drotations"]))
    return g


# ---------------------------------------------------------------------------------------------------------------
# AMR / foveated
# ---------------------------------------------------------------------------------------------------------------
def amr_build(sc: Scene, cam: Camera) -> ForwardState:
    """foveaStep == 0: preprocess + binning on 32x32 tiles + percentile levels; no blend."""
    st = preprocess(sc, cam, 32)
    bin_tiles(st)
    st.colors = sc.colors_precomp if sc.colors_precomp is not None else st.rgb
    nt = st.gx * st.gy
    st.n_intersections = np.zeros(nt, np.uint32)
    st.levels = np.zeros(nt, np.uint32)
    st.levels_last = np.zeros(nt, np.uint32)
    st.levels_current = np.zeros(nt, np.uint32)
    st.thresholds = np.zeros(3, np.uint32)
    lib().orc_amr_levels(C.c_int(nt), _ptr(st.ranges), _ptr(st.n_intersections), _ptr(st.levels), _ptr(st.thresholds))
    st.final_T = np.zeros(cam.W * cam.H, np.float32)
    st.n_contrib = np.zeros(cam.W * cam.H, np.uint32)
    return st


def _gaze_arr(gaze):
    if gaze is None:
        return None
    cx, cy, radii = gaze
    return np.asarray([cx, cy, *radii], np.float32)


def amr_render(cam: Camera, st: ForwardState, fovea_step: int, out_color_precomp=None, interpolate=False, gaze=None):
    """One call of the AMR forward for fovea_step != 0 on buffers built by amr_build (steps must run in order)."""
    L = lib()
    nt = st.gx * st.gy
    gz = _gaze_arr(gaze)
    L.orc_fovea_levels(C.c_int(fovea_step), _ptr(st.levels_last), _ptr(st.levels_current), _ptr(st.levels), C.c_int(nt),
                       C.c_int(st.gx), C.c_int(32), _ptr(gz))
    out = np.zeros((3, cam.H, cam.W), np.float32)
    lv = st.levels_current if fovea_step >= 1 else st.levels
    L.orc_blend_amr(C.c_int(cam.W), C.c_int(cam.H), _ptr(st.ranges), _ptr(lv), _ptr(st.levels_last), C.c_int(fovea_step),
                    _ptr(st.point_list), _ptr(st.means2D), _ptr(st.colors), _ptr(st.conic_opacity), _ptr(cam.bg),
                    _ptr(st.final_T), _ptr(st.n_contrib), _ptr(out))
    if interpolate:
        pre = _f(out_color_precomp) if out_color_precomp is not None and np.size(out_color_precomp) else None
        L.orc_interpolate(C.c_int(cam.W), C.c_int(cam.H), _ptr(lv), _ptr(st.levels_last), C.c_int(fovea_step),
                          _ptr(st.final_T), _ptr(st.n_contrib), _ptr(out), _ptr(pre))
    return out
