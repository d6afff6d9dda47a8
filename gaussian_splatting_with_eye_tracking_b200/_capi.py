"""ctypes binding of include/gsr_b200.h.

This module is the only place that touches the shared library.  It fails loudly: if libgsr_b200.so is missing and
cannot be built, the first call into it raises -- there is no CPU or PyTorch fallback anywhere in the product path.
(The library is mapped on first use, not at import, so that a process which never calls the product -- bench.py's
reference arm -- does not carry it.)
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)

# status codes (gsr_b200.h)
GSR_OK, GSR_ERR_INVALID, GSR_ERR_CUDA, GSR_ERR_ALLOC, GSR_ERR_WORKSPACE, GSR_ERR_SORT = range(6)
GSR_BWD_ALL, GSR_BWD_BLEND, GSR_BWD_PREPROCESS = 0, 1, 2
GSR_BUF_GEOM, GSR_BUF_BINNING, GSR_BUF_IMAGE = range(3)
GSR_MODE_BASELINE, GSR_MODE_AMR = 0, 1

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_int, C.c_size_t)


class GsrScene(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("sh_degree", C.c_int), ("sh_coeffs", C.c_int),
        ("means3D", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p),
        ("scales", C.c_void_p), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("scale_modifier", C.c_float), ("raw_params", C.c_int), ("shs_rest", C.c_void_p),
    ]


class GsrCamera(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
        ("background", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("cam_pos", C.c_void_p),
        ("prefiltered", C.c_int), ("band_row0", C.c_int), ("band_row1", C.c_int),
    ]


class GsrGaze(C.Structure):
    _fields_ = [("cx", C.c_float), ("cy", C.c_float), ("radii", C.c_float * 4)]


class GsrGrads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
        "dL_drotations", "dL_dconic", "scratch", "dL_dcolors_clamped")] + [
        ("push_bases", C.c_void_p), ("push_offset", C.c_size_t), ("push_rank", C.c_int), ("push_world", C.c_int),
        ("push_rows_per_rank", C.c_int), ("push_epoch", C.c_uint)]


class GsrAdamTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_longlong), ("lr", C.c_double), ("step", C.c_int)]


class GsrGeomView(C.Structure):
    _fields_ = [("P", C.c_int)] + [(n, C.c_void_p) for n in (
        "depths", "clamped", "radii", "means2D", "cov3D", "conic_opacity", "rgb", "tiles_touched", "block_offsets",
        "depth_order", "num_rendered")]


class GsrBinningView(C.Structure):
    _fields_ = [("N", C.c_longlong)] + [(n, C.c_void_p) for n in (
        "keys_unsorted", "values_unsorted", "tiles_sorted", "point_list")]


class GsrImageView(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("width", "height", "mode", "tiles_x", "tiles_y")] + [(n, C.c_void_p) for n in (
        "accum_alpha", "n_contrib", "ranges", "n_intersections", "levels", "levels_last", "levels_current",
        "thresholds")]


# every symbol include/gsr_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "gsr_geom_bytes": (C.c_size_t, [C.c_int]),
    "gsr_binning_bytes": (C.c_size_t, [C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int]),
    "gsr_image_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "gsr_forward": (C.c_int, [C.POINTER(GsrScene), C.POINTER(GsrCamera), ALLOC_FN, C.c_void_p, C.c_void_p, C.c_void_p,
                              c_int_p, C.c_int, C.c_void_p]),
    "gsr_backward": (C.c_int, [C.POINTER(GsrScene), C.POINTER(GsrCamera), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_int, C.c_void_p, C.POINTER(GsrGrads), C.c_int, C.c_void_p]),
    "gsr_backward_phase": (C.c_int, [C.POINTER(GsrScene), C.POINTER(GsrCamera), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_void_p, C.POINTER(GsrGrads), C.c_int, C.c_int, C.c_void_p]),
    "gsr_backward_preprocess_range": (C.c_int, [C.POINTER(GsrScene), C.POINTER(GsrCamera), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_int, C.POINTER(GsrGrads), C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "gsr_mark_visible": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_sh_grad_from_views": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_pack_meta_words": (C.c_size_t, [C.c_int]),
    "gsr_pack_temp_bytes": (C.c_size_t, [C.c_int]),
    "gsr_pack_visible_rows": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "gsr_sh_grad_from_packed_views": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                                C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "gsr_last_counts": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "gsr_amr_forward": (C.c_int, [C.POINTER(GsrScene), C.POINTER(GsrCamera), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_int, C.POINTER(GsrGaze), ALLOC_FN, C.c_void_p, C.c_void_p, C.c_void_p,
                                  c_int_p, C.c_int, C.c_void_p]),
    "gsr_amr_relevel": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "gsr_amr_render_prebuilt": (C.c_int, [C.POINTER(GsrScene), C.POINTER(GsrCamera), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                          C.POINTER(GsrGaze), C.c_void_p, C.c_int, C.c_void_p]),
    "gsr_blend_stats": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_view_geom": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(GsrGeomView)]),
    "gsr_view_binning": (C.c_int, [C.c_void_p, C.c_longlong, C.POINTER(GsrBinningView)]),
    "gsr_view_image": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(GsrImageView)]),
    "gsr_sort_temp_bytes": (C.c_size_t, [C.c_longlong]),
    "gsr_sort_pairs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong,
                                 C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "gsr_higher_msb": (C.c_uint, [C.c_uint]),
    "gsr_l1_ssim_partials": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "gsr_l1_ssim_forward": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_l1_ssim_backward": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_adam_step": (C.c_int, [C.POINTER(GsrAdamTensor), C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p]),
    "gsr_pack_rgb8": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "gsr_peer_flag_bytes": (C.c_size_t, [C.c_int]),
    "gsr_peer_barrier": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_uint, C.c_void_p]),
    "gsr_peer_owner_sum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_size_t,
                                     C.c_int, C.c_int, C.c_uint, C.c_int, C.c_void_p]),
    "gsr_peer_error": (C.c_int, []),
    "gsr_pack_visible_rows_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_pack_noalloc_max_gaussians": (C.c_int, []),
    "gsr_peer_allreduce_sum": (C.c_int, [C.c_void_p, C.c_size_t, C.c_longlong, C.c_size_t, C.c_int, C.c_int, C.c_uint, C.c_void_p]),
    "gsr_peer_allreduce_sum_mc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_longlong, C.c_size_t, C.c_int, C.c_int, C.c_uint, C.c_void_p]),
    "gsr_launch_count": (C.c_ulonglong, []),
    "gsr_set_option": (C.c_int, [C.c_int, C.c_int]),
    "gsr_profile_enable": (None, [C.c_int]),
    "gsr_profile_read": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "gsr_last_error": (C.c_char_p, []),
    "gsr_version": (C.c_char_p, []),
}


def _load() -> C.CDLL:
    path = _build.LIB_PATH
    if not os.path.exists(path):
        # first use in a fresh checkout: compile in-tree (needs nvcc); raises if that is impossible
        _build.build()
    elif _build.is_stale():
        try:                       # sources newer than the library: rebuild when a compiler is at hand
            _build.build()
        except RuntimeError:
            pass                   # no nvcc here: the prebuilt library is used; a header/library mismatch raises below
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


class _LazyLib:
    """The shared library, dlopen'ed on FIRST USE rather than at import: a process that only imports the package for
    its host-side helpers (scenes.py in bench.py's reference arm) never maps libgsr_b200.so.  The first attribute access
    loads (or builds) the library and raises if that is impossible -- there is still no fallback of any kind."""
    _real = None

    def _get(self) -> C.CDLL:
        if _LazyLib._real is None:
            _LazyLib._real = _load()
        return _LazyLib._real

    def __getattr__(self, name):
        return getattr(self._get(), name)


lib = _LazyLib()
LIB_PATH = _build.LIB_PATH


def loaded() -> bool:
    """True once libgsr_b200.so has been mapped into this process."""
    return _LazyLib._real is not None


class GsrError(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != GSR_OK:
        msg = lib.gsr_last_error()
        raise GsrError(f"{what} failed (status {rc}): {msg.decode() if msg else ''}")


PROF_SLOTS = ["preprocess", "scan", "duplicate", "sort", "ranges", "blend", "amr_levels", "amr_blend", "bwd_blend",
              "bwd_preprocess", "depth_sort", "expand", "sort_pass0", "bin_scatter", "bin_sort"] + [f"sort_pass{i}" for i in range(3, 8)]


def profile_enable(on: bool) -> None:
    lib.gsr_profile_enable(1 if on else 0)


def profile_read() -> dict:
    """{stage: (total_ms, count)} of the intervals recorded since the last read (synchronises on them)."""
    n = len(PROF_SLOTS)
    ms = (C.c_float * n)()
    cnt = (C.c_int * n)()
    check(lib.gsr_profile_read(ms, cnt), "gsr_profile_read")
    return {PROF_SLOTS[i]: (float(ms[i]), int(cnt[i])) for i in range(n) if cnt[i]}


def launch_count() -> int:
    return int(lib.gsr_launch_count())
