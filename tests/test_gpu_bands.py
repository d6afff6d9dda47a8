"""Screen-tile bands on one GPU: the ranks of a band-parallel frame are emulated one after the other
and the stitched result must equal the whole-frame render bit for bit (tiles are independent after binning)."""
import numpy as np
import pytest
import torch

import common
from common import A

pytestmark = pytest.mark.gpu


def _bands(tiles_y, world):
    from gaussian_splatting_with_eye_tracking_b200 import dist as gdist
    return gdist.balanced_bands([1.0 + (r % 3) for r in range(tiles_y)], world)


@pytest.mark.parametrize("case,world", [("small", 2), ("ragged", 3), ("dense", 4)])
def test_baseline_bands_stitch_to_full_frame(case, world, cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host, views
    sc, cam = common.make_case(case)
    d, camd = common.cuda_inputs(sc, cam)
    H, W = cam.image_height, cam.image_width
    args = common.fwd_args(d, camd)
    Nf, full, radii_f, gf, bf, imf = _host.rasterize_gaussians(*args, False)
    full_ranges = A(views.image_view(imf, W, H, False)["ranges"])
    full_pl = A(views.binning_view(bf, Nf)["point_list"])
    dpix = common.upstream_gradient(cam).cuda()
    gfull = _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii_f, d["colors_precomp"], d["scales"], d["rotations"], 1.0,
                                               d["cov3D_precomp"], camd.world_view_transform, camd.full_proj_transform, cam.tanfovx,
                                               cam.tanfovy, dpix, d["shs"], d["sh_degree"], camd.camera_center, gf, Nf, bf, imf, False)
    tiles_y, tiles_x = (H + 15) // 16, (W + 15) // 16
    bands = _bands(tiles_y, world)
    stitched = torch.full_like(full, float("nan"))
    n_sum, radii_max = 0, torch.zeros_like(radii_f)
    gsum = [torch.zeros_like(t) for t in gfull]
    for (r0, r1) in bands:
        N, img, radii, gb, bb, ib = _host.rasterize_gaussians(*args, False, band=(r0, r1))
        n_sum += N
        radii_max = torch.maximum(radii_max, radii)
        y0, y1 = min(r0 * 16, H), min(r1 * 16, H)
        stitched[:, y0:y1] = img[:, y0:y1]
        rg = A(views.image_view(ib, W, H, False)["ranges"]).reshape(tiles_y, tiles_x, 2)
        pl = A(views.binning_view(bb, N)["point_list"])
        fr = full_ranges.reshape(tiles_y, tiles_x, 2)
        cnt = rg[..., 1] - rg[..., 0]
        assert (cnt[:r0] == 0).all() and (cnt[r1:] == 0).all()                       # nothing binned outside the band
        common.assert_ints_equal(cnt[r0:r1], (fr[..., 1] - fr[..., 0])[r0:r1], "per-tile pair counts inside the band")
        for ty in range(r0, r1):                                                      # same depth-ordered list per tile
            for tx in (0, tiles_x // 2, tiles_x - 1):
                a0, a1 = rg[ty, tx]; b0, b1 = fr[ty, tx]
                assert np.array_equal(pl[a0:a1], full_pl[b0:b1])
        g = _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0,
                                               d["cov3D_precomp"], camd.world_view_transform, camd.full_proj_transform, cam.tanfovx,
                                               cam.tanfovy, dpix, d["shs"], d["sh_degree"], camd.camera_center, gb, N, bb, ib, False,
                                               band=(r0, r1))
        for acc, t in zip(gsum, g):
            acc += t
    assert n_sum == Nf
    common.assert_ints_equal(radii_max, radii_f, "max over bands of radii")
    common.assert_bits_equal(stitched, full, "stitched bands vs whole frame")
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"]
    common.assert_grads_close(dict(zip(names, gsum)), dict(zip(names, gfull)), floor_factor=common.GRAD_FLOOR_REF)


@pytest.mark.parametrize("case,world", [("small", 2), ("dense", 3)])
@pytest.mark.parametrize("interpolate", [False, True])
def test_amr_bands_match_single_gpu(case, world, interpolate, cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host, views
    sc, cam = common.make_case(case)
    d, camd = common.cuda_inputs(sc, cam)
    H, W = cam.image_height, cam.image_width
    args = common.fwd_args(d, camd)
    e = torch.empty(0, device="cuda"); eb = torch.empty(0, dtype=torch.uint8, device="cuda")
    single = common.run_ours_amr(sc, cam, False, interpolate)
    tiles_y = (H + 31) // 32
    bands = _bands(tiles_y, world)
    built = []
    for band in bands:
        N, _, radii, gb, bb, ib = _host.rasterize_gaussians_amr(*args, 0, e, eb, eb, eb, False, False, band=band)
        built.append((gb, bb, ib, views.image_view(ib, W, H, True)))
    total = sum(b[3]["n_intersections"].clone() for b in built)          # what the NCCL all-reduce produces
    common.assert_ints_equal(total, single["n_intersections"], "summed band counts vs single GPU")
    stitched = torch.zeros(3, H, W, device="cuda")
    for band, (gb, bb, ib, iv) in zip(bands, built):
        iv["n_intersections"].copy_(total)
        _host.amr_relevel(ib, H, W)
        common.assert_ints_equal(iv["levels"], single["levels"], "levels after global re-levelling")
        common.assert_ints_equal(iv["thresholds"], single["thresholds"], "thresholds")
        img = _host.amr_render_prebuilt(*args, gb, bb, ib, interpolate, band=band)
        y0, y1 = min(band[0] * 32, H), min(band[1] * 32, H)
        # rows outside the band are not this call's to write (left untouched, as in the 16-px path): only the band is read
        stitched[:, y0:y1] = img[:, y0:y1]
    common.assert_bits_equal(stitched, single["out_color"], "stitched AMR bands vs single GPU")


def test_bad_band_is_rejected(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case("tiny")
    d, camd = common.cuda_inputs(sc, cam)
    for band in [(2, 1), (0, 99), (3, 0)]:
        with pytest.raises(RuntimeError, match="band"):
            _host.rasterize_gaussians(*common.fwd_args(d, camd), False, band=band)
