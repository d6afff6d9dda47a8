"""GPU parity against the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built from /root/reference sources for
sm_100a).  This is the strongest pin: same inputs, same GPU, reference's own kernels."""
import numpy as np
import pytest
import torch

import common
from common import A
from oracle import ref_cuda

pytestmark = pytest.mark.gpu

needs_ref = pytest.mark.skipif(not ref_cuda.available(False), reason="oracle/_ref/libref_raster.so not built")
needs_ref_amr = pytest.mark.skipif(not ref_cuda.available(True), reason="oracle/_ref/libref_raster_amr.so not built")


@needs_ref
@pytest.mark.parametrize("case", common.ALL_CASES + ["obj"])
def test_forward_vs_reference_cuda(case, cuda_device):
    sc, cam = common.make_case(case)
    ours = common.run_ours(sc, cam)
    ref = common.run_reference(sc, cam)
    vis = A(ref["radii"]) > 0
    common.assert_ints_equal(ours["radii"], ref["radii"], "radii")
    assert ours["N"] == ref["N"]
    common.assert_ints_equal(ours["tiles_touched"], ref["tiles_touched"], "tiles_touched")
    common.assert_bits_equal(ours["depths"], ref["depths"], "depths", vis)
    common.assert_bits_equal(ours["means2D"], ref["means2D"], "means2D", vis)
    if ref["N"]:
        common.assert_ints_equal(ours["keys"], ref["keys"], "sorted tile/depth keys")
        common.assert_ints_equal(ours["point_list"], ref["point_list"], "sorted Gaussian indices")
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "tile ranges")
    common.assert_bits_equal(ours["conic_opacity"], ref["conic_opacity"], "conic_opacity", vis)
    if sc.get("colors_precomp") is None:
        common.assert_bits_equal(ours["rgb"], ref["rgb"], "rgb", vis)
    common.assert_ints_equal(ours["n_contrib"], ref["n_contrib"], "n_contrib")
    diff = np.abs(A(ours["out_color"]) - A(ref["out_color"])).max()
    assert diff <= common.IMG_TOL, f"image max-abs vs reference CUDA: {diff}"
    # same expf, same operation order -> expected to be bit-identical
    common.assert_bits_equal(ours["out_color"], ref["out_color"], "image (bitwise)")
    common.assert_bits_equal(ours["accum_alpha"], ref["accum_alpha"], "final_T (bitwise)")


@needs_ref
@pytest.mark.parametrize("case", ["tiny", "small", "ragged", "inside", "dense", "deg1", "precomp", "obj"])
def test_backward_vs_reference_cuda(case, cuda_device):
    sc, cam = common.make_case(case)
    dpix = common.upstream_gradient(cam)
    ours = common.run_ours(sc, cam, dpix)
    ref = common.run_reference(sc, cam, dpix)
    common.assert_grads_close(ours["grads"], ref["grads"], floor_factor=common.GRAD_FLOOR_REF)


def _ref_amr(sc, cam, stepped, interpolate):
    ref = ref_cuda.RefRasterizer(True)
    d, camd = common.cuda_inputs(sc, cam)
    a = common.fwd_args(d, camd)
    H, W = cam.image_height, cam.image_width
    e = torch.empty(0, device="cuda")
    P = d["means3D"].shape[0]

    def call(step, precomp, gb, bb, ib, interp):
        return ref.forward(*a[:12], H, W, d["shs"], d["sh_degree"], camd.camera_center, False, step, precomp, gb, bb, ib, interp)

    if not stepped:
        N, color, radii, gb, bb, ib = call(-2, e, None, None, None, interpolate)
        imgs = [color]
    else:
        N, c0, radii, gb, bb, ib = call(0, e, None, None, None, False)
        acc = c0.clone()
        imgs = [c0]
        for k in (1, 2, 3, 4):
            _, ck, _, gb, bb, ib = call(k, acc, gb, bb, ib, interpolate if k == 4 else False)
            imgs.append(ck)
            acc = acc + ck
        color = acc
    out = dict(N=N, out_color=color, images=imgs, radii=radii)
    out.update(ref.parse(gb, bb, ib, P, N, W, H))
    torch.cuda.synchronize()
    return out


def _ref_amr_or_optout(sc, cam, stepped, interpolate):
    """The AMR pin must not vanish silently: when oracle/_ref/libref_raster_amr.so exists but cannot run, the test FAILS
    unless GSR_ALLOW_AMR_REF_SKIP=1 says the box is known not to address registered host memory."""
    import os
    try:
        return _ref_amr(sc, cam, stepped, interpolate)
    except RuntimeError as ex:
        if os.environ.get("GSR_ALLOW_AMR_REF_SKIP") == "1":
            pytest.skip(str(ex))
        raise


@needs_ref_amr
@pytest.mark.parametrize("case", ["small", "ragged", "dense"])
@pytest.mark.parametrize("stepped", [False, True])
def test_amr_vs_reference_cuda(case, stepped, cuda_device):
    sc, cam = common.make_case(case)
    ref = _ref_amr_or_optout(sc, cam, stepped, False)
    ours = common.run_ours_amr(sc, cam, stepped, False)
    common.assert_ints_equal(ours["radii"], ref["radii"], "radii")
    assert ours["N"] == ref["N"]
    common.assert_ints_equal(ours["keys"], ref["keys"], "sorted keys")
    common.assert_ints_equal(ours["point_list"], ref["point_list"], "sorted indices")
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "ranges")
    common.assert_ints_equal(ours["n_intersections"], ref["n_intersections"], "n_intersections")
    common.assert_ints_equal(ours["levels"], ref["levels"], "AMR tile levels")
    common.assert_ints_equal(ours["levels_current"], ref["levels_current"], "levels_current")
    common.assert_ints_equal(ours["levels_last"], ref["levels_last"], "levels_last")
    for k, (x, y) in enumerate(zip(ours["images"], ref["images"])):
        dmax = np.abs(A(x) - A(y)).max()
        assert dmax <= common.IMG_TOL, f"image {k}: max-abs {dmax}"


@needs_ref_amr
def test_amr_interpolated_vs_reference_cuda(cuda_device):
    sc, cam = common.make_case("small")
    ref = _ref_amr_or_optout(sc, cam, False, True)
    ours = common.run_ours_amr(sc, cam, False, True)
    dmax = np.abs(A(ours["out_color"]) - A(ref["out_color"])).max()
    assert dmax <= common.IMG_TOL, f"interpolated AMR image max-abs {dmax}"


# ---------------------------------------------------------------------------------------------------------------
# BASELINE.json sizes against the LIVE reference (not invariants): C2 = 100 k Gaussians, 800x800, forward + backward;
# C3 = 3 M Gaussians, 1920x1080, forward on two ring cameras + backward on one; AMR levels / lists at C3.
# This is where the exact-preserving tile cull of the blend kernels (cannot_touch / reject_below, gsr_blend.cu) decides
# ~30 M pairs per frame: a flipped n_contrib or a single differing pixel bit fails here.
# Reference code matched: cuda_rasterizer/forward.cu:261-374 (blend), rasterizer_impl.cu:198-336 (lists), backward.cu.
# ---------------------------------------------------------------------------------------------------------------
def _full_scene(name, cam_index):
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    if name == "C2":
        return scenes.make_object_scene(P=100_000), scenes.object_cameras(8, 800, 800)[cam_index]
    if name == "C3":
        return scenes.make_unbounded_scene(P=3_000_000, seed=2002), scenes.unbounded_cameras(8, 1920, 1080)[cam_index]
    raise KeyError(name)


def _assert_forward_identical(ours, ref, sc):
    vis = A(ref["radii"]) > 0
    common.assert_ints_equal(ours["radii"], ref["radii"], "radii")
    assert ours["N"] == ref["N"]
    common.assert_ints_equal(ours["tiles_touched"], ref["tiles_touched"], "tiles_touched")
    common.assert_bits_equal(ours["depths"], ref["depths"], "depths", vis)
    common.assert_bits_equal(ours["means2D"], ref["means2D"], "means2D", vis)
    common.assert_bits_equal(ours["conic_opacity"], ref["conic_opacity"], "conic_opacity", vis)
    common.assert_bits_equal(ours["rgb"], ref["rgb"], "rgb", vis)
    assert torch.equal(ours["point_list"].view(torch.int32), ref["point_list"].view(torch.int32)), "sorted Gaussian indices differ"
    assert torch.equal(ours["keys"].view(torch.int64), ref["keys"].view(torch.int64)), "sorted (tile, depth) keys differ"
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "tile ranges")
    common.assert_ints_equal(ours["n_contrib"], ref["n_contrib"], "n_contrib")
    common.assert_bits_equal(ours["accum_alpha"], ref["accum_alpha"], "final_T (bitwise)")
    common.assert_bits_equal(ours["out_color"], ref["out_color"], "image (bitwise)")


@needs_ref
@pytest.mark.parametrize("name,cam_index", [("C2", 1), ("C2", 6), ("C3", 2), ("C3", 5)])
def test_full_size_forward_vs_reference_cuda(name, cam_index, cuda_device):
    sc, cam = _full_scene(name, cam_index)
    ours = common.run_ours(sc, cam)
    ref = common.run_reference(sc, cam)
    _assert_forward_identical(ours, ref, sc)


@needs_ref
@pytest.mark.parametrize("name,cam_index", [("C2", 1), ("C3", 2)])
def test_full_size_backward_vs_reference_cuda(name, cam_index, cuda_device):
    sc, cam = _full_scene(name, cam_index)
    dpix = common.upstream_gradient(cam)
    ours = common.run_ours(sc, cam, dpix)["grads"]
    ours = {k: v.clone() for k, v in ours.items()}
    torch.cuda.empty_cache()
    ref = {k: v.clone() for k, v in common.run_reference(sc, cam, dpix)["grads"].items()}
    torch.cuda.empty_cache()
    ref2 = common.run_reference(sc, cam, dpix)["grads"]
    # The reference accumulates every Gaussian's gradient with float atomicAdd in scheduling order
    # (cuda_rasterizer/backward.cu:523-554): at this size its OWN two runs differ by more than the small-scene floor of
    # 1e-5 |g|max.  Bar: 1e-3 relative + max(1e-5 |g|max, 4 x the reference's own run-to-run spread of that tensor).
    for k in ["dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations", "dL_dcolors"]:
        a, b, b2 = A(ours[k]).astype(np.float64), A(ref[k]).astype(np.float64), A(ref2[k]).astype(np.float64)
        a = a.reshape(b.shape)
        gmax = max(np.abs(a).max(initial=0.0), np.abs(b).max(initial=0.0))
        spread = np.abs(b - b2).max(initial=0.0)
        floor = max(common.GRAD_FLOOR_REF * gmax, 4.0 * spread) + 1e-12
        err = np.abs(a - b) - common.GRAD_RTOL * np.maximum(np.abs(a), np.abs(b)) - floor
        assert err.max(initial=-1.0) <= 0, (f"{k}: max abs diff {np.abs(a - b).max()} with |g|max {gmax}, reference run-to-run spread {spread}")
        # and in aggregate: relative L2 distance far inside the 1e-3 bar
        denom = np.linalg.norm(b.ravel())
        if denom > 0:
            assert np.linalg.norm((a - b).ravel()) / denom <= 1e-4, f"{k}: relative L2 {np.linalg.norm((a - b).ravel()) / denom}"


@needs_ref_amr
def test_full_size_amr_vs_reference_cuda(cuda_device):
    sc, cam = _full_scene("C3", 2)
    ref = _ref_amr_or_optout(sc, cam, False, False)
    ours = common.run_ours_amr(sc, cam, False, False)
    assert ours["N"] == ref["N"]
    assert torch.equal(ours["point_list"].view(torch.int32), ref["point_list"].view(torch.int32)), "AMR sorted indices differ"
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "ranges")
    common.assert_ints_equal(ours["n_intersections"], ref["n_intersections"], "n_intersections")
    common.assert_ints_equal(ours["levels"], ref["levels"], "AMR tile levels")
    dmax = np.abs(A(ours["out_color"]) - A(ref["out_color"])).max()
    assert dmax <= common.IMG_TOL, f"AMR single-pass image at C3: max-abs {dmax}"
    common.assert_bits_equal(ours["out_color"], ref["out_color"], "AMR image (bitwise)")
