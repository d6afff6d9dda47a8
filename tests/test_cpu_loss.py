"""CPU checks of the training-step surround: the numpy oracle of the L1+SSIM loss and of Adam is pinned
against outputs of the reference's own utils/loss_utils.py (tests/golden/loss_*.npz) and against torch.optim.Adam, and the
product wrappers refuse CPU tensors (no fallback).  No kernel is launched here."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import loss_oracle as lo  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
CASES = ["small", "tiny", "edge32", "mono"]


@pytest.mark.parametrize("name", CASES)
def test_loss_oracle_matches_the_reference_loss_utils(name):
    z = np.load(os.path.join(GOLDEN, f"loss_{name}.npz"))
    x, y, lam = z["image"], z["gt"], float(z["lambda_dssim"])
    # float64 reference outputs: the restatement agrees to rounding
    assert abs(lo.l1(x, y) - float(z["l1_f64"])) < 1e-12
    assert abs(lo.ssim(x, y) - float(z["ssim_f64"])) < 1e-10
    assert abs(lo.loss(x, y, lam) - float(z["loss_f64"])) < 1e-10
    g = lo.loss_grad(x, y, 1.0 - lam, -lam)
    np.testing.assert_allclose(g, z["grad_f64"], rtol=1e-7, atol=1e-12 * np.abs(z["grad_f64"]).max() + 1e-15)
    gs = lo.loss_grad(x, y, 0.0, 1.0)
    np.testing.assert_allclose(gs, z["ssim_grad_f64"], rtol=1e-7, atol=1e-15)
    # and the float32 run of the reference sits within float32 noise of it (the bar the GPU kernel is held to)
    assert abs(lo.loss(x, y, lam) - float(z["loss_f32"])) < 2e-6
    scale = np.abs(z["grad_f64"]).max()
    assert np.abs(z["grad_f32"] - g).max() < 1e-4 * scale


def test_window_is_the_reference_window():
    w = lo.gaussian_window()
    assert w.shape == (11,) and abs(w.sum() - 1.0) < 1e-6 and np.allclose(w, w[::-1])
    assert abs(w[5] - 0.26601172) < 1e-6          # gaussian(11, 1.5)[5] as float32


def test_adam_oracle_matches_torch_adam():
    g = torch.Generator().manual_seed(5)
    shapes, lrs = [(37, 3), (37, 1, 3), (37, 15, 3), (37, 1), (37, 3), (37, 4)], [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3]
    params = [torch.nn.Parameter(torch.randn(s, generator=g)) for s in shapes]
    opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(params, lrs)], lr=0.0, eps=1e-15)   # gaussian_model.py:151-160
    mine = [(p.detach().numpy().copy(), np.zeros(s, np.float32), np.zeros(s, np.float32)) for p, s in zip(params, shapes)]
    for step in range(1, 5):
        grads = [torch.randn(s, generator=g) * (10.0 ** (-step)) for s in shapes]
        grads[3][::2] = 0.0                         # Gaussians invisible in this view: zero gradient, moments still decay
        for p, gr in zip(params, grads):
            p.grad = gr.clone()
        opt.step()
        mine = [lo.adam_step(p_, gr.numpy(), m_, v_, step, lr) for (p_, m_, v_), gr, lr in zip(mine, grads, lrs)]
        for p, (p_, m_, v_) in zip(params, mine):
            st = opt.state[p]
            # float32 rounding of (g - m) where the two nearly cancel: absolute floor relative to the largest entry
            for mine_, ref_ in ((p_, p.detach().numpy()), (m_, st["exp_avg"].numpy()), (v_, st["exp_avg_sq"].numpy())):
                np.testing.assert_allclose(mine_, ref_, rtol=2e-6, atol=1e-6 * np.abs(ref_).max())


def test_product_wrappers_refuse_cpu_tensors():
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils, optim
    x, y = torch.rand(3, 16, 16), torch.rand(3, 16, 16)
    for fn in (loss_utils.l1_loss, loss_utils.ssim, loss_utils.l1_ssim_loss):
        with pytest.raises(RuntimeError, match="CUDA"):
            fn(x, y)
    with pytest.raises(NotImplementedError):
        loss_utils.ssim(x, y, window_size=7)
    p = torch.nn.Parameter(torch.zeros(4))
    p.grad = torch.ones(4)
    opt = optim.FusedAdam([{"params": [p], "lr": 1e-3, "name": "xyz"}], lr=0.0, eps=1e-15)
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.step()
    assert opt.param_groups[0]["name"] == "xyz"     # extra group keys survive (gaussian_model.py:163-168 looks groups up by name)
    with pytest.raises(NotImplementedError):
        optim.FusedAdam([p], amsgrad=True)


def test_adam_argument_validation_without_a_gpu():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    assert _capi.lib.gsr_adam_step(None, 0, 0.9, 0.999, 1e-15, None) == _capi.GSR_OK       # nothing to do, nothing launched
    assert _capi.lib.gsr_adam_step(None, 2, 0.9, 0.999, 1e-15, None) == _capi.GSR_ERR_INVALID
    assert _capi.lib.gsr_l1_ssim_partials(3, 1080, 1920) == 3 * 34 * 60
    assert _capi.lib.gsr_l1_ssim_forward(3, 0, 5, None, None, 1.0, 0.0, 0.0, None, None, None, None) == _capi.GSR_ERR_INVALID
