"""GPU parity: the sm_100a path, called through the C-ABI, against the CPU oracle on the same seeded inputs.

Bar: integer/index stages bit-exact; image max-abs <= 1e-4; gradients <= 1e-3 relative."""
import numpy as np
import pytest
import torch

import common
from common import A

pytestmark = pytest.mark.gpu


def _vis(o):
    return A(o["radii"]) > 0


@pytest.mark.parametrize("case", common.ALL_CASES)
def test_forward_matches_oracle(case, cuda_device):
    sc, cam = common.make_case(case)
    ours = common.run_ours(sc, cam)
    ref = common.run_oracle(sc, cam)
    # --- integer stages: bit-exact
    common.assert_ints_equal(ours["radii"], ref["radii"], "radii")
    assert ours["N"] == ref["N"], f"num_rendered {ours['N']} vs {ref['N']}"
    vis = _vis(ref)
    common.assert_ints_equal(A(ours["tiles_touched"]), ref["tiles_touched"].view(np.int32), "tiles_touched")
    common.assert_bits_equal(ours["depths"], ref["depths"], "depth bits", vis)
    common.assert_bits_equal(ours["means2D"], ref["means2D"], "means2D", vis)
    if ref["N"]:
        common.assert_ints_equal(ours["keys"], ref["keys"], "sorted tile/depth keys")
        common.assert_ints_equal(ours["point_list"], ref["point_list"], "sorted Gaussian indices")
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "tile ranges")
    # --- per-Gaussian floats: pinned arithmetic -> bitwise
    common.assert_bits_equal(ours["conic_opacity"], ref["conic_opacity"], "conic_opacity", vis)
    if sc.get("colors_precomp") is None:
        common.assert_bits_equal(ours["rgb"], ref["rgb"], "rgb", vis)
        common.assert_ints_equal(A(ours["clamped"])[vis], ref["clamped"][vis], "clamped flags")
    if sc.get("cov3D_precomp") is None:
        common.assert_bits_equal(ours["cov3D"], ref["cov3D"], "cov3D", vis)
    # --- image: CUDA expf vs libm expf differ by ulps -> tolerance
    # The alpha < 1/255 and T < 1e-4 tests are discontinuities: a 1-ulp difference between CUDA's expf (MUFU.EX2) and
    # libm's expf can flip one (pixel, Gaussian) decision, worth up to ~1/255 in colour.  Against the CPU
    # oracle a handful of such pixels is tolerated and bounded; against the reference CUDA rasterizer
    # (test_gpu_reference.py) the image is required to match bit for bit.
    diff = np.abs(A(ours["out_color"]) - ref["out_color"]).max(axis=0)
    nc_diff = int((A(ours["n_contrib"]) != ref["n_contrib"]).sum())
    flips = int((diff > common.IMG_TOL).sum())
    budget = max(2, diff.size // 5000)
    assert flips <= budget, f"{flips} pixels exceed {common.IMG_TOL} (budget {budget}); max-abs {diff.max()}"
    assert diff.max() <= 6e-3, f"image max-abs {diff.max()} is more than one threshold flip"
    assert nc_diff <= budget, f"n_contrib differs on {nc_diff} pixels"


@pytest.mark.parametrize("case", ["tiny", "small", "ragged", "inside", "dense", "deg1", "precomp"])
def test_backward_matches_oracle(case, cuda_device):
    sc, cam = common.make_case(case)
    dpix = common.upstream_gradient(cam)
    ours = common.run_ours(sc, cam, dpix)
    ref = common.run_oracle(sc, cam, dpix)
    common.assert_grads_close(ours["grads"], ref["grads"])


def test_empty_scene(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case("tiny")
    for k in ("means3D", "scales", "rotations", "opacities", "shs"):
        sc[k] = sc[k][:0]
    d, camd = common.cuda_inputs(sc, cam)
    N, color, radii, geom, binning, img = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
    assert N == 0 and radii.numel() == 0
    assert float(color.abs().max()) == 0.0        # reference: nothing launched, out_color keeps its zero fill


def test_bad_means_shape_raises(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case("tiny")
    d, camd = common.cuda_inputs(sc, cam)
    d["means3D"] = d["means3D"][:, :2].contiguous()
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        _host.rasterize_gaussians(*common.fwd_args(d, camd), False)


def test_mark_visible(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    from oracle import oracle as orc
    import ctypes as C
    sc, cam = common.make_case("inside")
    d, camd = common.cuda_inputs(sc, cam)
    got = A(_host.mark_visible(d["means3D"], camd.world_view_transform, camd.full_proj_transform))
    m = np.ascontiguousarray(A(sc["means3D"]), np.float32)
    v = np.ascontiguousarray(A(cam.world_view_transform), np.float32)
    want = np.zeros(m.shape[0], np.uint8)
    orc.lib().orc_mark_visible(C.c_int(m.shape[0]), m.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p),
                               want.ctypes.data_as(C.c_void_p))
    assert 0 < want.sum() < want.size
    common.assert_ints_equal(got.astype(np.uint8), want, "mark_visible")


@pytest.mark.parametrize("n,end_bit", [(1, 40), (31, 44), (4608, 45), (4609, 45), (100_003, 47), (1_000_000, 45), (5000, 8), (70_000, 64)])
def test_radix_sort_matches_stable_sort(n, end_bit, cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import views
    g = torch.Generator().manual_seed(n)
    hi = torch.randint(0, 1 << 13, (n,), generator=g, dtype=torch.int64)
    lo = torch.randint(0, 1 << 20, (n,), generator=g, dtype=torch.int64) << 10     # many duplicate low parts -> stability matters
    keys = ((hi << 32) | lo)
    if end_bit == 64:
        keys = keys | (torch.randint(0, 2, (n,), generator=g, dtype=torch.int64) << 63)
    vals = torch.arange(n, dtype=torch.int32)
    ko, vo = views.sort_pairs(keys.cuda(), vals.cuda(), end_bit)
    mask = (1 << end_bit) - 1
    mk = (keys.numpy().view(np.uint64) & np.uint64(mask))
    order = np.argsort(mk, kind="stable")
    common.assert_ints_equal(vo, vals.numpy()[order], "sorted values (stable)")
    common.assert_ints_equal(ko, keys.numpy()[order], "sorted keys")
