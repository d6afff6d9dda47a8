"""Multi-tensor Adam for the Gaussian parameter groups.

``FusedAdam`` is a ``torch.optim.Optimizer`` with the state layout of ``torch.optim.Adam`` -- ``state[p]`` holds ``"step"``,
``"exp_avg"``, ``"exp_avg_sq"`` -- because the reference's densification code edits exactly those entries in place
(``scene/gaussian_model.py:284-303`` replace_tensor_to_optimizer, ``:305-320`` _prune_optimizer, ``:335-352``
cat_tensors_to_optimizer) and builds the optimizer as ``torch.optim.Adam(l, lr=0.0, eps=1e-15)`` with one group per
parameter (``:149-160``).  ``step()`` updates every tensor of every group in ONE kernel launch (``gsr_adam_step``,
csrc/gsr_train.cu) instead of torch's per-tensor / foreach chains; 28 B of HBM traffic per parameter float.

No CPU path: parameters must be contiguous float32 CUDA tensors, otherwise ``step()`` raises.
"""
from __future__ import annotations

import ctypes as C

import torch

from ._capi import GsrAdamTensor, check, lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("weight_decay / amsgrad are not used by the reference (gaussian_model.py:160)")
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=0, amsgrad=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # one launch per (device, betas, eps) combination -- the reference has a single one
        batches = {}
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad
                if not p.is_cuda or not g.is_cuda:
                    raise RuntimeError("FusedAdam: parameters and gradients must be CUDA tensors (no CPU fallback)")
                if p.dtype != torch.float32 or g.dtype != torch.float32:
                    raise RuntimeError("FusedAdam: float32 only")
                if g.is_sparse:
                    raise RuntimeError("FusedAdam: sparse gradients are not supported")
                if not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous")
                g = g.contiguous()
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if m.shape != p.shape or v.shape != p.shape or not m.is_contiguous() or not v.is_contiguous():
                    raise RuntimeError("FusedAdam: optimizer state does not match its parameter")
                st["step"] += 1
                key = (p.device, float(b1), float(b2), float(group["eps"]))
                batches.setdefault(key, []).append((p, g, m, v, float(group["lr"]), int(st["step"])))
        for (device, b1, b2, eps), items in batches.items():
            arr = (GsrAdamTensor * len(items))()
            for i, (p, g, m, v, lr, step) in enumerate(items):
                arr[i].param = p.data_ptr(); arr[i].grad = g.data_ptr()
                arr[i].exp_avg = m.data_ptr(); arr[i].exp_avg_sq = v.data_ptr()
                arr[i].n = p.numel(); arr[i].lr = lr; arr[i].step = step
            with torch.cuda.device(device):
                check(lib.gsr_adam_step(arr, len(items), b1, b2, eps, C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                      "gsr_adam_step")
            del items  # the tensors outlive the launch through self.state / the parameters themselves
        return loss
