"""GPU parity of the training-step surround, through the C ABI via the reference-named wrappers:
the fused L1+SSIM loss against (i) outputs of the reference's own utils/loss_utils.py (tests/golden/loss_*.npz), (ii) the
numpy oracle, (iii) a plain PyTorch fp32/fp64 restatement run live (F.conv2d, autograd) at sizes up to 1080p; and the
multi-tensor Adam against torch.optim.Adam.  Floating point: values within 2e-6 absolute (float32 sums of ~1e4..1e6 terms),
gradients within 1e-4 of the largest entry -- written below."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import loss_oracle as lo  # noqa: E402

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(ROOT, "tests", "golden")
VAL_TOL = 2e-6
GRAD_TOL = 1e-4       # x max|grad|


def torch_ref_loss(x, y, lam):
    """utils/loss_utils.py:17-63 + train.py:87-88 restated with plain torch ops (any dtype/device)."""
    w1 = torch.tensor(lo.gaussian_window(), dtype=torch.float32)
    w2 = (w1[:, None] @ w1[None, :]).to(x.dtype).to(x.device)
    C = x.shape[0]
    win = w2.expand(C, 1, 11, 11).contiguous()
    conv = lambda a: F.conv2d(a[None], win, padding=5, groups=C)[0]
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu1 * mu2 + c1) * (2 * s12 + c2)) / ((mu1 * mu1 + mu2 * mu2 + c1) * (s1 + s2 + c2))).mean()
    l1 = (x - y).abs().mean()
    return (1.0 - lam) * l1 + lam * (1.0 - ssim), l1, ssim


@pytest.mark.parametrize("name", ["small", "tiny", "edge32", "mono"])
def test_fused_loss_matches_the_reference_outputs(cuda_device, name):
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils
    z = np.load(os.path.join(GOLDEN, f"loss_{name}.npz"))
    lam = float(z["lambda_dssim"])
    x = torch.from_numpy(z["image"]).to(cuda_device).requires_grad_(True)
    y = torch.from_numpy(z["gt"]).to(cuda_device)
    loss, terms = loss_utils.l1_ssim_loss(x, y, lam, return_terms=True)
    loss.backward()
    t = terms.cpu().numpy()
    assert abs(float(loss.detach()) - float(z["loss_f64"])) < VAL_TOL and abs(t[0] - float(loss.detach())) == 0
    assert abs(t[1] - float(z["l1_f64"])) < VAL_TOL and abs(t[2] - float(z["ssim_f64"])) < VAL_TOL
    g, want = x.grad.cpu().numpy().astype(np.float64), z["grad_f64"]
    assert np.abs(g - want).max() < GRAD_TOL * np.abs(want).max()
    # the reference's own float32 run is no closer to its float64 run than that
    assert np.abs(g - z["grad_f32"]).max() < GRAD_TOL * np.abs(want).max()
    # the separate entry points (loss_utils.l1_loss / ssim) and their gradients
    x2 = x.detach().clone().requires_grad_(True)
    s = loss_utils.ssim(x2, y)
    s.backward()
    assert abs(float(s) - float(z["ssim_f64"])) < VAL_TOL
    assert np.abs(x2.grad.cpu().numpy() - z["ssim_grad_f64"]).max() < GRAD_TOL * np.abs(z["ssim_grad_f64"]).max()
    x3 = x.detach().clone().requires_grad_(True)
    l = loss_utils.l1_loss(x3, y)
    (3.0 * l).backward()                                     # a non-unit upstream gradient
    assert abs(float(l) - float(z["l1_f64"])) < VAL_TOL
    want_l1 = 3.0 * np.sign(z["image"].astype(np.float64) - z["gt"]) / z["image"].size
    assert np.abs(x3.grad.cpu().numpy() - want_l1).max() < 1e-12 + 1e-6 * np.abs(want_l1).max()
    # oracle == kernel within the same bars (keeps the three descriptions tied together)
    assert abs(lo.loss(z["image"], z["gt"], lam) - float(loss)) < VAL_TOL


@pytest.mark.parametrize("shape", [(3, 1080, 1920), (3, 800, 800), (3, 33, 65), (3, 5, 3), (2, 3, 100, 52)])
def test_fused_loss_against_live_torch(cuda_device, shape):
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils
    g = torch.Generator().manual_seed(sum(shape))
    gt = torch.rand(shape, generator=g)
    img = (gt + 0.2 * torch.randn(shape, generator=g)).clamp(0, 1)
    x = img.to(cuda_device).requires_grad_(True)
    y = gt.to(cuda_device)
    loss, terms = loss_utils.l1_ssim_loss(x, y, 0.2, return_terms=True)
    (2.0 * loss).backward()
    xr = img.to(cuda_device).double().reshape(-1, *shape[-2:]).requires_grad_(True)
    lr, l1r, sr = torch_ref_loss(xr, y.double().reshape(-1, *shape[-2:]), 0.2)
    (2.0 * lr).backward()
    assert abs(float(loss) - float(lr)) < VAL_TOL
    assert abs(float(terms[1]) - float(l1r)) < VAL_TOL and abs(float(terms[2]) - float(sr)) < VAL_TOL
    want = xr.grad.reshape(shape)
    assert x.grad.shape == x.shape
    assert float((x.grad.double() - want).abs().max()) < GRAD_TOL * float(want.abs().max())
    # deterministic: same bits on a second run (fixed-order reduction, no atomics)
    x_b = img.to(cuda_device).requires_grad_(True)
    loss_b = loss_utils.l1_ssim_loss(x_b, y, 0.2)
    (2.0 * loss_b).backward()
    assert torch.equal(loss_b, loss) and torch.equal(x_b.grad, x.grad)
    # no gradient wanted -> same value, no maps written
    with torch.no_grad():
        # (the map-less template instance may contract its FMAs differently: equal to rounding, not to the bit)
        assert abs(float(loss_utils.l1_ssim_loss(img.to(cuda_device), y, 0.2)) - float(loss.detach())) < 1e-6


def test_identical_images(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils
    x = torch.rand(3, 70, 90, device=cuda_device).requires_grad_(True)
    loss, terms = loss_utils.l1_ssim_loss(x, x.detach().clone(), 0.2, return_terms=True)
    loss.backward()
    assert float(terms[1]) == 0.0 and abs(float(terms[2]) - 1.0) < 1e-6 and abs(float(loss)) < 1e-6
    assert float(x.grad.abs().max()) < 1e-9


def _groups(dev, P, g):
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}   # arguments/__init__.py
    params = {k: torch.randn(s, generator=g) for k, s in shapes.items()}
    return shapes, lrs, params


@pytest.mark.parametrize("P", [1, 1000, 4099, 70001])
def test_fused_adam_matches_torch_adam(cuda_device, P):
    from gaussian_splatting_with_eye_tracking_b200 import _capi, optim
    g = torch.Generator().manual_seed(P)
    shapes, lrs, init = _groups(cuda_device, P, g)
    mk = lambda: {k: torch.nn.Parameter(v.clone().to(cuda_device)) for k, v in init.items()}
    pa, pb = mk(), mk()
    ref = torch.optim.Adam([{"params": [pa[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)   # gaussian_model.py:151-160
    ours = optim.FusedAdam([{"params": [pb[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    for step in range(1, 6):
        n0 = _capi.launch_count()
        for k, s in shapes.items():
            gr = torch.randn(s, generator=g) * 10.0 ** (-(step % 3))
            gr[::3] = 0.0                              # Gaussians not visible in this view
            pa[k].grad = gr.to(cuda_device); pb[k].grad = gr.to(cuda_device)
        if step == 3:                                  # the learning-rate schedule edits param_groups in place (gaussian_model.py:162-168)
            for o in (ref, ours):
                for grp in o.param_groups:
                    if grp["name"] == "xyz":
                        grp["lr"] = 3.3e-5
        ref.step(); ours.step()
        assert _capi.launch_count() - n0 == 1          # all six groups in one launch
        for k in shapes:
            for a, b in ((pa[k].data, pb[k].data), (ref.state[pa[k]]["exp_avg"], ours.state[pb[k]]["exp_avg"]),
                         (ref.state[pa[k]]["exp_avg_sq"], ours.state[pb[k]]["exp_avg_sq"])):
                assert float((a - b).abs().max()) <= 2e-6 * float(a.abs().max()) + 1e-30, (k, step)
    assert int(ours.state[pb["xyz"]]["step"]) == 5


def test_fused_adam_survives_densification_style_state_surgery(cuda_device):
    """scene/gaussian_model.py:335-352 (cat_tensors_to_optimizer): the state entry is re-keyed to a longer parameter."""
    from gaussian_splatting_with_eye_tracking_b200 import optim
    p = torch.nn.Parameter(torch.randn(10, 3, device=cuda_device))
    opt = optim.FusedAdam([{"params": [p], "lr": 1e-2, "name": "xyz"}], lr=0.0, eps=1e-15)
    p.grad = torch.randn_like(p); opt.step()
    ext = torch.randn(4, 3, device=cuda_device)
    group = opt.param_groups[0]
    st = opt.state.get(group["params"][0])
    st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)), dim=0)
    st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
    del opt.state[group["params"][0]]
    group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0].data, ext), dim=0).requires_grad_(True))
    opt.state[group["params"][0]] = st
    q = group["params"][0]
    before = q.data.clone()
    q.grad = torch.ones_like(q); opt.step()
    assert int(st["step"]) == 2 and not torch.equal(q.data, before) and torch.isfinite(q.data).all()
