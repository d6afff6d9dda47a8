"""The bucketed depth sort (csrc/gsr_depthsort.cu) and its onesweep fallback against the reference's
cub::DeviceRadixSort::SortPairs order (rasterizer_impl.cu:303-317): sorted keys, point_list and ranges bit-identical.

* forced fallback (GSR_OPT_FORCE_SORT_FALLBACK): same outputs as the bucket path and as the reference
* natural overflow: thousands of Gaussians with IDENTICAL depth bits (a plane facing the camera) -- one histogram bin
  holds them all, the range buckets cannot split it, the library must take the fallback by itself; ties resolve by
  ascending Gaussian index exactly like the reference's stable sort
* narrow depth range far from the camera (few distinct float exponents) and a single Gaussian
"""
import math

import numpy as np
import pytest
import torch

import common
from common import A
from oracle import ref_cuda

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not ref_cuda.available(False), reason="oracle/_ref/libref_raster.so not built")

GSR_OPT_FORCE_SORT_FALLBACK = 3


def _force(on):
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    _capi.check(_capi.lib.gsr_set_option(GSR_OPT_FORCE_SORT_FALLBACK, 1 if on else 0), "gsr_set_option")


def _same_lists(ours, ref):
    assert ours["N"] == ref["N"]
    common.assert_ints_equal(ours["radii"], ref["radii"], "radii")
    if ref["N"]:
        assert torch.equal(ours["keys"].view(torch.int64), ref["keys"].view(torch.int64)), "sorted (tile, depth) keys differ"
        assert torch.equal(ours["point_list"].view(torch.int32), ref["point_list"].view(torch.int32)), "sorted Gaussian indices differ"
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "tile ranges")
    common.assert_ints_equal(ours["n_contrib"], ref["n_contrib"], "n_contrib")
    common.assert_bits_equal(ours["out_color"], ref["out_color"], "image (bitwise)")


@needs_ref
@pytest.mark.parametrize("case", ["tiny", "small", "dense", "inside", "offscreen", "obj"])
def test_forced_fallback_equals_bucket_path_and_reference(case, cuda_device):
    sc, cam = common.make_case(case)
    ref = common.run_reference(sc, cam)
    a = common.run_ours(sc, cam)
    try:
        _force(True)
        b = common.run_ours(sc, cam)
    finally:
        _force(False)
    _same_lists(a, ref)
    _same_lists(b, ref)


def _plane_scene(P, seed, z=3.0, jitter=0.0):
    """Gaussians on a plane facing the camera: view-space depth identical (jitter = 0) for all of them."""
    g = torch.Generator().manual_seed(seed)
    sc = common._cloud(P, seed, log_scale=math.log(0.03))
    xy = (torch.rand(P, 2, generator=g) - 0.5) * 3.0
    zz = torch.full((P, 1), z) + (torch.rand(P, 1, generator=g) - 0.5) * jitter
    sc["means3D"] = torch.cat([xy, zz], 1).contiguous()
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    cam = scenes.look_at_camera((0.0, 0.0, 0.0), (0.0, 0.0, 1.0), 192, 128, 1.0)
    return sc, cam


@needs_ref
def test_equal_depths_overflow_takes_the_fallback_by_itself(cuda_device):
    sc, cam = _plane_scene(30000, 41)
    ref = common.run_reference(sc, cam)
    d = A(ref["depths"])[A(ref["radii"]) > 0]
    assert d.size > 8000 and np.unique(d.view(np.uint32)).size < 8, "the case must put > one bucket of keys on a few depth values"
    ours = common.run_ours(sc, cam)
    _same_lists(ours, ref)


@needs_ref
@pytest.mark.parametrize("P,z,jitter", [(20000, 120.0, 0.5), (9000, 3.0, 1e-4), (1, 3.0, 0.0), (5000, 0.25, 0.05)])
def test_narrow_depth_ranges(P, z, jitter, cuda_device):
    sc, cam = _plane_scene(P, 43 + P % 7, z=z, jitter=jitter)
    ref = common.run_reference(sc, cam)
    ours = common.run_ours(sc, cam)
    _same_lists(ours, ref)


@needs_ref
def test_amr_lists_through_the_fallback(cuda_device):
    sc, cam = common.make_case("dense")
    a = common.run_ours_amr(sc, cam, False, False)
    try:
        _force(True)
        b = common.run_ours_amr(sc, cam, False, False)
    finally:
        _force(False)
    assert torch.equal(a["point_list"], b["point_list"])
    common.assert_ints_equal(a["ranges"], b["ranges"], "ranges")
    common.assert_ints_equal(a["levels"], b["levels"], "levels")
    common.assert_bits_equal(a["out_color"], b["out_color"], "AMR image")


def test_sort_watchdog_surfaces_as_gsr_err_sort(cuda_device):
    """a tripped look-back watchdog used to be written to a header word nobody read.  The last binning kernel now
    posts it to the host mailbox; the next call of the thread returns GSR_ERR_SORT (5), a debug forward reports it itself."""
    from gaussian_splatting_with_eye_tracking_b200 import _capi, _host
    sc, cam = common.make_case("small")
    d, camd = common.cuda_inputs(sc, cam)
    args = common.fwd_args(d, camd)
    _capi.check(_capi.lib.gsr_set_option(4, 1), "gsr_set_option")           # GSR_OPT_FAKE_SORT_ERROR
    try:
        _host.rasterize_gaussians(*args, False)                            # queues the frame; the flag arrives with it
        torch.cuda.synchronize()
        with pytest.raises(_capi.GsrError, match="status 5"):
            _host.rasterize_gaussians(*args, False)
        with pytest.raises(_capi.GsrError, match="status 5"):              # debug: synchronous, reported by the frame itself
            _host.rasterize_gaussians(*args, True)
    finally:
        _capi.check(_capi.lib.gsr_set_option(4, 0), "gsr_set_option")
        torch.cuda.synchronize()
    _host.rasterize_gaussians(*args, False)                                # healthy again
    torch.cuda.synchronize()
    _host.rasterize_gaussians(*args, False)
