"""View-parallel gradient exchange: the SH gradient leaves the backward as its rank-1 factor
(the clamped colour gradient, 3 floats per Gaussian) and gsr_sh_grad_from_views rebuilds the summed [P,M,3] tensor.
Checked against the dense SH gradient of the ordinary backward, through the C ABI."""
import pytest
import torch

import common

pytestmark = pytest.mark.gpu


def _close(a, b, what):
    """1e-3 relative with the float-atomics floor of the parity tests (blend-stage sums reorder between runs)."""
    viol, mx = common.grad_close(common.A(a).reshape(common.A(b).shape), common.A(b), common.GRAD_RTOL, common.GRAD_FLOOR_REF)
    assert viol <= 0, f"{what}: mismatch beyond 1e-3 relative (max abs diff {mx})"


def _backward(sc, cam, dpix, compact):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    d, camd = common.cuda_inputs(sc, cam)
    N, color, radii, geom, binning, img = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
    g = _host.rasterize_gaussians_backward(
        d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0, d["cov3D_precomp"],
        camd.world_view_transform, camd.full_proj_transform, cam.tanfovx, cam.tanfovy, dpix.cuda(), d["shs"],
        d["sh_degree"], camd.camera_center, geom, N, binning, img, False, compact_sh=compact)
    return d, camd, g


@pytest.mark.parametrize("case", ["small", "deg1", "inside"])
def test_compact_backward_plus_rebuild_equals_dense_sh_gradient(case):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case(case)
    dpix = common.upstream_gradient(cam)
    d, camd, dense = _backward(sc, cam, dpix, False)
    _, _, comp = _backward(sc, cam, dpix, True)
    P, M = d["shs"].shape[0], d["shs"].shape[1]
    assert tuple(comp[5].shape) == (P, 3) and tuple(dense[5].shape) == (P, M, 3)
    # every other gradient is untouched by the mode (blend atomics reorder: same tolerance as the parity tests)
    for i in (0, 1, 2, 3, 4, 6, 7):
        _close(comp[i], dense[i], f"grad[{i}]")
    rebuilt = _host.sh_grad_from_views(d["means3D"], camd.camera_center.reshape(1, 3), comp[5].reshape(1, P, 3), d["sh_degree"], M)
    torch.cuda.synchronize()
    _close(rebuilt, dense[5], "dL_dsh rebuilt from one view")
    # rows of culled Gaussians are exactly zero in both
    culled = (comp[5].abs().sum(1) == 0)
    assert bool((rebuilt[culled] == 0).all())


def test_rebuild_sums_views_in_order():
    from gaussian_splatting_with_eye_tracking_b200 import _host, scenes
    sc, cam0 = common.make_case("small")
    cam1 = scenes.look_at_camera((-1.2, 0.5, -3.0), (0, 0, 0), cam0.image_width, cam0.image_height, 0.9)
    gs, cams, dense = [], [], None
    for cam in (cam0, cam1):
        dpix = common.upstream_gradient(cam)
        d, camd, g_dense = _backward(sc, cam, dpix, False)
        _, _, g_comp = _backward(sc, cam, dpix, True)
        gs.append(g_comp[5]); cams.append(camd.camera_center.reshape(3))
        dense = g_dense[5] if dense is None else dense + g_dense[5]
    P, M = d["shs"].shape[0], d["shs"].shape[1]
    both = _host.sh_grad_from_views(d["means3D"], torch.stack(cams), torch.stack(gs), d["sh_degree"], M)
    _close(both, dense, "dL_dsh summed over two views")
    # bit-identical on repeat (fixed summation order): what makes the result equal on every rank
    again = _host.sh_grad_from_views(d["means3D"], torch.stack(cams), torch.stack(gs), d["sh_degree"], M)
    assert torch.equal(both, again)


def test_two_phase_backward_matches_single_call_and_hook_sees_final_tensor():
    """gsr_backward_phase: BLEND then PREPROCESS == one gsr_backward; the clamped colour gradient is final (in stream
    order) when the between-phases hook runs, which is where the view-parallel all-gather starts."""
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case("small")
    dpix = common.upstream_gradient(cam)
    d, camd, one = _backward(sc, cam, dpix, True)
    N, color, radii, geom, binning, img = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
    seen = {}

    def hook(drgb):
        seen["at_hook"] = drgb.clone()          # stream-ordered copy at the hook point
    two = _host.rasterize_gaussians_backward(
        d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0, d["cov3D_precomp"],
        camd.world_view_transform, camd.full_proj_transform, cam.tanfovx, cam.tanfovy, dpix.cuda(), d["shs"],
        d["sh_degree"], camd.camera_center, geom, N, binning, img, False, compact_sh=True, between_phases=hook)
    torch.cuda.synchronize()
    assert torch.equal(seen["at_hook"], two[5])          # not rewritten by the second phase
    for i in range(8):
        _close(two[i], one[i], f"grad[{i}] two-phase vs single call")


def test_last_counts_match_the_forward_outputs():
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case("small")
    d, camd = common.cuda_inputs(sc, cam)
    N, color, radii, geom, binning, img = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
    n, r, v = _host.last_counts(geom, d["means3D"].shape[0])
    assert n == N and v == int((radii > 0).sum()) and 0 < v <= r <= n
    with pytest.raises(RuntimeError):
        _host.last_counts(binning, d["means3D"].shape[0])          # not the geometry buffer of the last forward


@pytest.mark.parametrize("P", [1, 31, 32, 1000, 1025, 70001])
def test_pack_visible_rows_layout(P):
    """mask word g/32 bit g%32, prefix = set bits before the word, rows in index order (bit-exact copies)."""
    from gaussian_splatting_with_eye_tracking_b200 import _host
    gen = torch.Generator().manual_seed(P)
    radii = ((torch.rand(P, generator=gen) < 0.4).to(torch.int32) * 7).cuda()
    rows = torch.randn(P, 3, generator=gen).cuda()
    V = int((radii > 0).sum())
    for cap in (V, V + 13, max(0, V - 5)):
        packed, meta = _host.pack_visible_rows(rows, radii, cap)
        torch.cuda.synchronize()
        W32 = (P + 31) // 32
        assert meta.numel() == 2 * W32 == _host.pack_meta_words(P)
        bits = torch.zeros(W32 * 32, dtype=torch.int64); bits[:P] = (radii > 0).cpu().to(torch.int64)
        words = (bits.view(W32, 32) << torch.arange(32)).sum(1)
        counts = bits.view(W32, 32).sum(1)
        assert torch.equal(meta[:W32].cpu().to(torch.int64) & 0xffffffff, words)
        assert torch.equal(meta[W32:].cpu().to(torch.int64), torch.cumsum(counts, 0) - counts)
        want = rows[radii > 0][:cap]
        assert torch.equal(packed[: want.shape[0]], want)


def test_packed_rebuild_is_bit_identical_to_dense_rebuild():
    from gaussian_splatting_with_eye_tracking_b200 import _host, scenes
    sc, cam0 = common.make_case("small")
    cam1 = scenes.look_at_camera((-1.2, 0.5, -3.0), (0, 0, 0), cam0.image_width, cam0.image_height, 0.9)
    gs, cams, radiis = [], [], []
    for cam in (cam0, cam1):
        dpix = common.upstream_gradient(cam)
        d, camd = common.cuda_inputs(sc, cam)
        N, color, radii, geom, binning, img = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
        g = _host.rasterize_gaussians_backward(
            d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0, d["cov3D_precomp"],
            camd.world_view_transform, camd.full_proj_transform, cam.tanfovx, cam.tanfovy, dpix.cuda(), d["shs"],
            d["sh_degree"], camd.camera_center, geom, N, binning, img, False, compact_sh=True)
        assert _host.last_counts(geom, radii.numel())[2] == int((radii > 0).sum())
        gs.append(g[5]); cams.append(camd.camera_center.reshape(3)); radiis.append(radii)
    P, M = d["shs"].shape[0], d["shs"].shape[1]
    dense = _host.sh_grad_from_views(d["means3D"], torch.stack(cams), torch.stack(gs), d["sh_degree"], M)
    cap = (max(int((r > 0).sum()) for r in radiis) + 255) // 256 * 256
    cap = min(cap, P)
    L = _host.pack_meta_words(P)
    rows_all = torch.zeros(2, cap, 3, device="cuda")
    meta_all = torch.zeros(2, L + 4, dtype=torch.int32, device="cuda")          # strided meta, as the exchange lays it out
    for v in range(2):
        _host.pack_visible_rows(gs[v], radiis[v], cap, rows_all[v], meta_all[v, :L])
    packed = _host.sh_grad_from_packed_views(d["means3D"], torch.stack(cams), meta_all, rows_all, d["sh_degree"], M)
    torch.cuda.synchronize()
    assert torch.equal(packed, dense)
    assert float(dense.abs().max()) > 0


@pytest.mark.parametrize("chunks", [2, 3, 7])
def test_range_split_preprocess_backward_matches_single_launch(chunks):
    """gsr_backward_preprocess_range: consecutive ranges cover [0, P) once; every range's rows are final at its hook."""
    from gaussian_splatting_with_eye_tracking_b200 import _host
    sc, cam = common.make_case("small")
    dpix = common.upstream_gradient(cam)
    d, camd, one = _backward(sc, cam, dpix, True)
    N, color, radii, geom, binning, img = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
    P = d["means3D"].shape[0]
    seen, snaps = [], []

    def after(first, count, rows):
        seen.append((first, count))
        assert [tuple(r.shape) for r in rows] == [(count, 3), (count, 1), (count, 3), (count, 4)]
        assert all(r.is_contiguous() for r in rows)
        snaps.append([r.clone() for r in rows])
    split = _host.rasterize_gaussians_backward(
        d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0, d["cov3D_precomp"],
        camd.world_view_transform, camd.full_proj_transform, cam.tanfovx, cam.tanfovy, dpix.cuda(), d["shs"],
        d["sh_degree"], camd.camera_center, geom, N, binning, img, False, compact_sh=True,
        between_phases=lambda drgb: None, preprocess_chunks=chunks, after_chunk=after)
    torch.cuda.synchronize()
    assert seen[0][0] == 0 and sum(c for _, c in seen) == P and all(f % 256 == 0 for f, _ in seen)
    assert all(seen[i][0] + seen[i][1] == seen[i + 1][0] for i in range(len(seen) - 1))
    for (first, count), snap in zip(seen, snaps):          # rows were final when the hook ran
        for r, full in zip(snap, (split[3], split[2], split[6], split[7])):
            assert torch.equal(r, full[first:first + count])
    for i in range(8):
        _close(split[i], one[i], f"grad[{i}] range-split vs single launch")
