"""The formats either side of the path feed it unchanged: a cloud written as the reference's PLY and read
back (raw parameters -> activations) renders to the image of the original tensors; a camera taken through cameras.json does
too; and the pieces of a training iteration (rasterizer autograd API + loss_utils.l1_ssim_loss + optim.FusedAdam) fit a target
image like the plain-torch chain (utils/loss_utils.py formulas via F.conv2d + torch.optim.Adam) on the same rasterizer."""
import numpy as np
import pytest
import torch

import common

pytestmark = pytest.mark.gpu


def _render(d, cam, dev, means2D=None):
    import diff_gaussian_rasterization as dgr
    rs = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"].to(dev), 1.0,
                                           cam.world_view_transform.to(dev), cam.full_proj_transform.to(dev), d["sh_degree"],
                                           cam.camera_center.to(dev), False, False)
    return dgr.GaussianRasterizer(rs)(means3D=d["means3D"], means2D=means2D, opacities=d["opacities"], shs=d["shs"],
                                      scales=d["scales"], rotations=d["rotations"])


def test_ply_and_cameras_json_round_trip_renders_the_same_image(cuda_device, tmp_path):
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    sc, cam = common.make_case("small")
    d = {k: (v.to(cuda_device) if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    with torch.no_grad():
        img0, radii0 = _render(d, cam, cuda_device)
        op = sc["opacities"].clamp(1e-6, 1 - 1e-6)
        path = str(tmp_path / "point_cloud" / "iteration_1" / "point_cloud.ply")
        scene_io.save_ply(path, sc["means3D"], sc["shs"][:, :1], sc["shs"][:, 1:], torch.log(op / (1 - op)), torch.log(sc["scales"]),
                          sc["rotations"] * 1.7)                                   # rotations are stored unnormalised
        a = scene_io.activate(scene_io.load_ply(path, device=cuda_device))
        assert a["means3D"].is_cuda and torch.equal(a["means3D"].cpu(), sc["means3D"]) and torch.equal(a["shs"].cpu(), sc["shs"])
        a["bg"] = sc["bg"]
        img1, radii1 = _render(a, cam, cuda_device)
        assert float((radii0 != radii1).float().mean()) < 5e-3                   # log/exp, logit/sigmoid round trips move a few radii by one
        assert float((img0 - img1).abs().max()) < 5e-3 and float((img0 - img1).abs().mean()) < 1e-5
        # the camera through cameras.json (camera centre + camera-to-world rotation + focal lengths)
        W2C = cam.world_view_transform.t().double().numpy()
        entry = scene_io.camera_to_json(0, W2C[:3, :3].T, W2C[:3, 3], cam.FoVx, cam.FoVy, cam.image_width, cam.image_height, "view0")
        jp = str(tmp_path / "cameras.json")
        scene_io.save_cameras_json(jp, [entry])
        cam2 = scene_io.cameras_from_json(jp)[0]
        np.testing.assert_allclose(cam2.full_proj_transform.numpy(), cam.full_proj_transform.numpy(), atol=2e-5)
        img2, _ = _render(d, cam2, cuda_device)
        assert float((img0 - img2).abs().mean()) < 1e-4


def _torch_loss(x, y, lam, win):
    import torch.nn.functional as F
    conv = lambda t: F.conv2d(t[None], win, padding=5, groups=3)[0]  # noqa: E731
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    ssim = (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))).mean()
    return (1.0 - lam) * (x - y).abs().mean() + lam * (1.0 - ssim)


def test_training_iterations_fit_a_target_like_the_plain_torch_chain(cuda_device):
    """train.py:83-121 in miniature: 40 iterations on one view; both chains start from the same perturbed scene."""
    import math
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils, optim
    sc, cam = common.make_case("small")
    d = {k: (v.to(cuda_device) if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    with torch.no_grad():
        target, _ = _render(d, cam, cuda_device)
    g = torch.Generator().manual_seed(99)
    start = {"means3D": sc["means3D"] + 0.01 * torch.randn(sc["means3D"].shape, generator=g),
             "shs": sc["shs"] + 0.2 * torch.randn(sc["shs"].shape, generator=g)}
    w1 = torch.tensor([math.exp(-(x - 5) ** 2 / 4.5) for x in range(11)])
    w1 = w1 / w1.sum()
    win = (w1[:, None] @ w1[None, :]).expand(3, 1, 11, 11).contiguous().to(cuda_device)
    curves = {}
    for chain in ("fused", "torch"):
        m3 = torch.nn.Parameter(start["means3D"].clone().to(cuda_device))
        sh = torch.nn.Parameter(start["shs"].clone().to(cuda_device))
        groups = [{"params": [m3], "lr": 1e-4, "name": "xyz"}, {"params": [sh], "lr": 1e-2, "name": "f"}]
        opt = (optim.FusedAdam if chain == "fused" else torch.optim.Adam)(groups, lr=0.0, eps=1e-15)
        losses = []
        for it in range(40):
            cur = dict(d, means3D=m3, shs=sh)
            m2d = torch.zeros_like(m3, requires_grad=True)          # screenspace_points (gaussian_renderer/__init__.py:26-30)
            img, _ = _render(cur, cam, cuda_device, means2D=m2d)
            loss = loss_utils.l1_ssim_loss(img, target, 0.2) if chain == "fused" else _torch_loss(img, target, 0.2, win)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
            losses.append(float(loss.detach()))
        curves[chain] = losses
    for chain, ls in curves.items():
        assert all(math.isfinite(v) for v in ls), chain
        assert ls[-1] < 0.6 * ls[0], (chain, ls[0], ls[-1])                      # the fit makes progress
    assert abs(curves["fused"][0] - curves["torch"][0]) < 1e-5                    # same loss at the common start
    assert abs(curves["fused"][-1] - curves["torch"][-1]) < 0.05 * curves["torch"][0]   # and the same trajectory, up to float atomics
