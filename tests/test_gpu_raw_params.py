"""Raw-parameter entry: rendering from the model's raw tensors equals the reference's path -- the getters of
scene/gaussian_model.py:95-115 (sigmoid, exp, normalize, cat) followed by the standard rasterizer -- and its gradients equal
autograd through those getters."""
import numpy as np
import pytest
import torch

import common
from common import A

pytestmark = pytest.mark.gpu


def _model(P, seed, dev):
    from gaussian_splatting_with_eye_tracking_b200 import model
    g = torch.Generator().manual_seed(seed)
    pc = model.GaussianModel(3)
    raw = {"xyz": (torch.rand(P, 3, generator=g) - 0.5) * 2.4, "features_dc": torch.randn(P, 1, 3, generator=g) * 1.5,
           "features_rest": torch.randn(P, 15, 3, generator=g) * 0.08, "opacity": torch.randn(P, 1, generator=g) * 2.0,
           "scaling": torch.randn(P, 3, generator=g) * 0.5 + np.log(0.05), "rotation": torch.randn(P, 4, generator=g) * 3.0}
    pc._adopt({k: v.to(dev) for k, v in raw.items()})
    return pc


@pytest.mark.parametrize("degree", [0, 2, 3])
def test_raw_entry_equals_activated_path(degree, cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import rasterizer, scenes
    dev = cuda_device
    pc = _model(5000, 91 + degree, dev)
    pc.active_sh_degree = degree
    cam = scenes.look_at_camera((0.9, 0.4, -3.1), (0, 0, 0), 208, 144, 0.9).to(dev)
    bg = torch.tensor([0.1, 0.4, 0.7], device=dev)
    dpix = common.upstream_gradient(cam).to(dev)

    # reference path: activated copies through the getters, standard entry
    import math
    rs = rasterizer.GaussianRasterizationSettings(cam.image_height, cam.image_width, math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), bg, 1.0,
                                                  cam.world_view_transform, cam.full_proj_transform, degree, cam.camera_center, False, False)
    m2d = torch.zeros_like(pc._xyz, requires_grad=True)
    img_a, radii_a = rasterizer.GaussianRasterizerBaseline(rs)(means3D=pc.get_xyz, means2D=m2d, opacities=pc.get_opacity, shs=pc.get_features,
                                                                scales=pc.get_scaling, rotations=pc.get_rotation)
    (img_a * dpix).sum().backward()
    ga = {k: v.grad.clone() for k, v in pc._p.items()}
    ga["means2D"] = m2d.grad.clone()
    for v in pc._p.values():
        v.grad = None

    out = rasterizer.render_raw(cam, pc, bg)
    (out["render"] * dpix).sum().backward()
    gb = {k: v.grad.clone() for k, v in pc._p.items()}
    gb["means2D"] = out["viewspace_points"].grad.clone()

    # activations: sigmoid / exp are torch's expressions bit for bit; the quaternion norm may round differently in the last place,
    # which can move a radius by one -> the image bar is the north_star's 1e-4, radii may differ on a vanishing fraction
    d = np.abs(A(out["render"]) - A(img_a)).max()
    assert d <= common.IMG_TOL, f"image max-abs {d}"
    mism = int((out["radii"] != radii_a).sum())
    assert mism <= max(2, pc._xyz.shape[0] // 2000), f"{mism} radii differ"
    assert torch.equal(out["visibility_filter"], out["radii"] > 0)
    for k in ga:
        viol, mx = common.grad_close(gb[k], ga[k], common.GRAD_RTOL, 1e-4)
        assert viol <= 0, f"{k}: max abs diff {mx}"


def test_raw_entry_argument_checks(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host, _capi
    dev = cuda_device
    pc = _model(64, 5, dev)
    sc, cam = common.make_case("tiny")
    d, camd = common.cuda_inputs(sc, cam, dev)
    e = torch.empty(0, device=dev)
    args = (d["bg"], pc._xyz.detach(), e, pc._opacity.detach(), pc._scaling.detach(), pc._rotation.detach(), 1.0, e, camd.world_view_transform,
            camd.full_proj_transform, cam.tanfovx, cam.tanfovy, cam.image_height, cam.image_width)
    with pytest.raises(RuntimeError):          # sh must be features_dc [P,1,3]
        _host.rasterize_gaussians(*args, pc.get_features.detach(), 3, camd.camera_center, False, False, sh_rest=pc._features_rest.detach())
    N, img, *_ = _host.rasterize_gaussians(*args, pc._features_dc.detach(), 3, camd.camera_center, False, False, sh_rest=pc._features_rest.detach())
    assert bool(torch.isfinite(img).all())
