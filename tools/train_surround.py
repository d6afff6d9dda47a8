"""Device time of the training-step surround kernels (fused L1+SSIM forward/backward, multi-tensor Adam) next to the plain
PyTorch chain the reference runs (utils/loss_utils.py via F.conv2d + autograd; torch.optim.Adam), C3 sizes.
    python tools/train_surround.py [--gaussians 3000000] [--width 1920 --height 1080]
Prints one JSON object; bench.py embeds the same measurement as extra.train_surround."""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _time(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); b.synchronize()
    return a.elapsed_time(b) / iters


def torch_loss(x, y, lam, win):
    C = x.shape[0]
    conv = lambda a: F.conv2d(a[None], win, padding=5, groups=C)[0]
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    ssim = (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))).mean()
    return (1.0 - lam) * (x - y).abs().mean() + lam * (1.0 - ssim)


def measure(P=3_000_000, W=1920, H=1080, dev=None, hbm_gbs=None):
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils, optim
    dev = dev or torch.device("cuda:0")
    g = torch.Generator().manual_seed(7)
    gt = torch.rand(3, H, W, generator=g).to(dev)
    img = (gt + 0.1 * torch.randn(3, H, W, generator=g).to(dev)).clamp(0, 1)
    out = {"image": [3, H, W], "gaussians": P}
    # ---- loss
    x = img.clone().requires_grad_(True)

    def ours():
        x.grad = None
        loss_utils.l1_ssim_loss(x, gt, 0.2).backward()
    w1 = torch.exp(-(torch.arange(11.0) - 5) ** 2 / 4.5)            # gaussian(11, 1.5), utils/loss_utils.py:23-25
    w1 = (w1 / w1.sum()).to(dev)
    win = (w1[:, None] @ w1[None, :]).expand(3, 1, 11, 11).contiguous()

    def ref():
        x.grad = None
        torch_loss(x, gt, 0.2, win).backward()
    with torch.no_grad():
        t_fwd = _time(lambda: loss_utils.l1_ssim_loss(img, gt, 0.2))
    t_ours, t_ref = _time(ours), _time(ref)
    px = 3 * H * W
    # algorithmic bytes: fwd reads image+gt (8 B/px/channel) writes 3 maps (12); bwd reads 3 maps + image + gt (20) writes grad (4)
    out["l1_ssim"] = {"fwd_only_ms": round(t_fwd, 4), "fwd_bwd_ms": round(t_ours, 4), "torch_fwd_bwd_ms": round(t_ref, 4),
                      "speedup": round(t_ref / t_ours, 2), "alg_bytes_fwd_bwd": 44 * px,
                      "alg_gbs": round(44 * px / (t_ours * 1e-3) / 1e9, 1)}
    # ---- Adam over the six groups (59 floats per Gaussian)
    shapes = [(P, 3), (P, 1, 3), (P, 15, 3), (P, 1), (P, 3), (P, 4)]
    lrs = [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3]

    def mk():
        ps = [torch.nn.Parameter(torch.randn(s, device=dev)) for s in shapes]
        for p in ps:
            p.grad = torch.randn_like(p) * 1e-3
        return ps
    pa = mk()
    ours_opt = optim.FusedAdam([{"params": [p], "lr": lr} for p, lr in zip(pa, lrs)], lr=0.0, eps=1e-15)
    t_a = _time(ours_opt.step)
    del ours_opt, pa
    pb = mk()
    ref_opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(pb, lrs)], lr=0.0, eps=1e-15)
    t_b = _time(ref_opt.step)
    del ref_opt, pb
    nb = 28 * 59 * P
    out["adam"] = {"ms": round(t_a, 4), "torch_adam_ms": round(t_b, 4), "speedup": round(t_b / t_a, 2), "alg_bytes": nb,
                   "alg_gbs": round(nb / (t_a * 1e-3) / 1e9, 1)}
    if hbm_gbs:
        out["adam"]["frac_of_hbm_peak"] = round(out["adam"]["alg_gbs"] / hbm_gbs, 3)
        out["l1_ssim"]["frac_of_hbm_peak"] = round(out["l1_ssim"]["alg_gbs"] / hbm_gbs, 3)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    a = ap.parse_args()
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    print(json.dumps(measure(a.gaussians, a.width, a.height, hbm_gbs=peak)))
