"""One launch each of the training-surround kernels at C3 sizes for `ncu` (loss forward with maps, loss backward, Adam over
the six groups), bracketed by cudaProfilerStart/Stop:
    ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/surround python tools/profile_surround.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import loss_utils, optim  # noqa: E402

dev = torch.device("cuda:0")
P, H, W = 3_000_000, 1080, 1920
g = torch.Generator().manual_seed(7)
gt = torch.rand(3, H, W, generator=g).to(dev)
x = (gt + 0.1 * torch.randn(3, H, W, generator=g).to(dev)).clamp(0, 1).requires_grad_(True)
shapes = [(P, 3), (P, 16, 3), (P, 1), (P, 3), (P, 4)]
ps = [torch.nn.Parameter(torch.randn(s, device=dev)) for s in shapes]
opt = optim.FusedAdam([{"params": [p], "lr": 1e-3} for p in ps], lr=0.0, eps=1e-15)


def step():
    x.grad = None
    loss_utils.l1_ssim_loss(x, gt, 0.2).backward()
    for p in ps:
        p.grad = torch.full_like(p, 1e-3)
    opt.step()


step(); step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok")
