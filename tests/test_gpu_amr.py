"""GPU parity of the AMR / foveated forward (32x32 tiles, percentile levels, 5-step protocol) against the oracle."""
import numpy as np
import pytest

import common
from common import A

pytestmark = pytest.mark.gpu

AMR_CASES = ["small", "ragged", "dense", "inside"]


def _check_common(ours, ref):
    common.assert_ints_equal(ours["radii"], ref["radii"], "radii (32-px tiles)")
    assert ours["N"] == ref["N"]
    common.assert_ints_equal(ours["keys"], ref["keys"], "sorted keys")
    common.assert_ints_equal(ours["point_list"], ref["point_list"], "sorted indices")
    common.assert_ints_equal(ours["ranges"], ref["ranges"], "tile ranges")
    common.assert_ints_equal(ours["n_intersections"], ref["n_intersections"], "n_intersections")
    common.assert_ints_equal(ours["thresholds"], ref["thresholds"], "percentile thresholds")
    common.assert_ints_equal(ours["levels"], ref["levels"], "AMR tile levels")


@pytest.mark.parametrize("case", AMR_CASES)
@pytest.mark.parametrize("interpolate", [False, True])
def test_single_pass(case, interpolate, cuda_device):
    sc, cam = common.make_case(case)
    ours = common.run_ours_amr(sc, cam, False, interpolate)
    ref = common.run_oracle_amr(sc, cam, False, interpolate)
    _check_common(ours, ref)
    common.assert_ints_equal(ours["levels_current"], ref["levels_current"], "levels_current")
    common.assert_ints_equal(ours["levels_last"], ref["levels_last"], "levels_last")
    diff = np.abs(A(ours["out_color"]) - ref["out_color"]).max()
    assert diff <= common.IMG_TOL, f"AMR image max-abs {diff}"
    lv = ref["levels"]
    assert lv.min() >= 1 and lv.max() <= 4


@pytest.mark.parametrize("case", ["small", "ragged"])
def test_five_step_protocol(case, cuda_device):
    sc, cam = common.make_case(case)
    ours = common.run_ours_amr(sc, cam, True, False)
    ref = common.run_oracle_amr(sc, cam, True, False)
    _check_common(ours, ref)
    common.assert_ints_equal(ours["levels_current"], ref["levels_current"], "levels_current after step 4")
    common.assert_ints_equal(ours["levels_last"], ref["levels_last"], "levels_last after step 4")
    for k, (a, b) in enumerate(zip(ours["images"], ref["images"])):
        d = np.abs(A(a) - b).max()
        assert d <= common.IMG_TOL, f"step {k}: max-abs {d}"
    assert float(A(ours["images"][0]).max()) == 0.0      # step 0 builds only
    # invariant: sum of the steps == single pass without interpolation, bit for bit
    single = common.run_ours_amr(sc, cam, False, False)
    common.assert_bits_equal(ours["out_color"], single["out_color"], "sum(steps) vs single pass")


def test_step4_interpolation(cuda_device):
    sc, cam = common.make_case("small")
    ours = common.run_ours_amr(sc, cam, True, True)
    ref = common.run_oracle_amr(sc, cam, True, True)
    for k, (a, b) in enumerate(zip(ours["images"], ref["images"])):
        d = np.abs(A(a) - b).max()
        assert d <= common.IMG_TOL, f"step {k}: max-abs {d}"


def test_gaze_cap_extension(cuda_device):
    """Extension A7 (no reference code: parity unpinned, oracle only).  Infinite radii == reference behaviour."""
    sc, cam = common.make_case("small")
    W, H = cam.image_width, cam.image_height
    inf = (1e9, 1e9, 1e9, 1e9)
    base = common.run_ours_amr(sc, cam, False, False)
    same = common.run_ours_amr(sc, cam, False, False, gaze=(W / 2, H / 2, inf))
    common.assert_bits_equal(base["out_color"], same["out_color"], "gaze with infinite radii")
    gz = (0.3 * W, 0.6 * H, (W / 2, W / 4, W / 8, W / 16))
    ours = common.run_ours_amr(sc, cam, False, True, gaze=gz)
    ref = common.run_oracle_amr(sc, cam, False, True, gaze=gz)
    common.assert_ints_equal(ours["levels"], ref["levels"], "gaze-capped levels")
    assert np.abs(A(ours["out_color"]) - ref["out_color"]).max() <= common.IMG_TOL
    assert (ref["levels"] < common.run_oracle_amr(sc, cam, False, False)["levels"]).any()
    st_o = common.run_ours_amr(sc, cam, True, False, gaze=gz)
    st_r = common.run_oracle_amr(sc, cam, True, False, gaze=gz)
    assert np.abs(A(st_o["out_color"]) - st_r["out_color"]).max() <= common.IMG_TOL
