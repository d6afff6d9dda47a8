"""Makes tests/golden/loss_*.npz by running the reference's OWN utils/loss_utils.py (pure torch, CPU) in the build
container:  python tests/golden/make_loss_golden.py   (needs /root/reference; the fixtures are committed)."""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

CASES = {"small": (3, 45, 70, 11), "tiny": (3, 7, 9, 12), "edge32": (3, 64, 33, 13), "mono": (1, 40, 40, 14)}
LAMBDA = 0.2   # arguments/__init__.py: lambda_dssim

for name, (C, H, W, seed) in CASES.items():
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    img = (gt + 0.25 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
    img[:, : H // 3] = gt[:, : H // 3]                      # a region with x == y exactly (sign(0) = 0, ssim = 1)
    out = {"image": img.numpy().copy(), "gt": gt.numpy()}
    for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        x = img.to(dt).clone().requires_grad_(True)
        y = gt.to(dt)
        Ll1 = ref.l1_loss(x, y)
        s = ref.ssim(x, y)
        loss = (1.0 - LAMBDA) * Ll1 + LAMBDA * (1.0 - s)      # train.py:87-88
        loss.backward()
        out[f"l1_{tag}"] = Ll1.detach().numpy(); out[f"ssim_{tag}"] = s.detach().numpy()
        out[f"loss_{tag}"] = loss.detach().numpy(); out[f"grad_{tag}"] = x.grad.numpy()
        x2 = img.to(dt).clone().requires_grad_(True)
        ref.ssim(x2, y).backward()
        out[f"ssim_grad_{tag}"] = x2.grad.numpy()
    np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), lambda_dssim=LAMBDA, **out)
    print(name, float(out["loss_f32"]), float(out["ssim_f64"]))
