"""The training loss either side of the rasterizer, same names as the reference's ``utils/loss_utils.py``.

* ``l1_loss(network_output, gt)``                      <- utils/loss_utils.py:17-18
* ``ssim(img1, img2, window_size=11, size_average=True)`` <- utils/loss_utils.py:39-63
* ``l1_ssim_loss(image, gt, lambda_dssim)``            <- train.py:87-88, both terms from ONE kernel per direction:
  ``(1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt))``

All three run ``gsr_l1_ssim_forward`` / ``gsr_l1_ssim_backward`` of libgsr_b200.so (csrc/gsr_train.cu); there is no
PyTorch or CPU fallback -- CPU tensors raise.  Gradients flow to the first argument only (the reference never
differentiates the ground-truth image).
"""
from __future__ import annotations

import ctypes as C

import torch

from ._capi import check, lib


def _as_chw(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    if t.dim() == 4:                       # [B,C,H,W] -> planes; the window never crosses a plane, so B*C channels
        t = t.reshape(-1, t.shape[-2], t.shape[-1])
    if t.dim() != 3:
        raise RuntimeError(f"{name} must be [C,H,W] or [B,C,H,W], got {tuple(t.shape)}")
    return t.contiguous()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _L1SSIM(torch.autograd.Function):
    """out[0] = w_l1 * L1 + w_ssim * SSIM + w_const (device scalar; no host synchronisation in either direction)."""

    @staticmethod
    def forward(ctx, image, gt, w_l1, w_ssim, w_const):
        shape = image.shape
        x = _as_chw(image, "image")
        y = _as_chw(gt.detach(), "gt")
        if x.shape != y.shape:
            raise RuntimeError(f"image {tuple(x.shape)} and gt {tuple(y.shape)} differ in shape")
        Cn, H, W = x.shape
        need_grad = image.requires_grad
        with torch.cuda.device(x.device):
            maps = torch.empty((3, Cn, H, W), dtype=torch.float32, device=x.device) if (need_grad and w_ssim != 0.0) else None
            partials = torch.empty(2 * lib.gsr_l1_ssim_partials(Cn, H, W), dtype=torch.float32, device=x.device)
            out = torch.empty(3, dtype=torch.float32, device=x.device)
            check(lib.gsr_l1_ssim_forward(Cn, H, W, x.data_ptr(), y.data_ptr(), w_l1, w_ssim, w_const,
                                          None if maps is None else maps.data_ptr(), partials.data_ptr(), out.data_ptr(),
                                          _stream()), "gsr_l1_ssim_forward")
        ctx.save_for_backward(x, y, maps if maps is not None else torch.empty(0, device=x.device))
        ctx.w = (float(w_l1), float(w_ssim))
        ctx.shape = shape
        ctx.mark_non_differentiable(out)   # the 3-vector is auxiliary; gradient flows through the scalar view below
        return out[0].clone(), out

    @staticmethod
    def backward(ctx, dloss, _daux):
        x, y, maps = ctx.saved_tensors
        Cn, H, W = x.shape
        w_l1, w_ssim = ctx.w
        dloss = dloss.contiguous().to(torch.float32)
        with torch.cuda.device(x.device):
            dimg = torch.empty_like(x)
            check(lib.gsr_l1_ssim_backward(Cn, H, W, x.data_ptr(), y.data_ptr(), w_l1, w_ssim,
                                           maps.data_ptr() if maps.numel() else None, dloss.data_ptr(), dimg.data_ptr(),
                                           _stream()), "gsr_l1_ssim_backward")
        return dimg.view(ctx.shape), None, None, None, None


def l1_ssim_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2, return_terms: bool = False):
    """train.py:87-88 fused.  ``return_terms=True`` also returns the device 3-vector {loss, L1, SSIM} (train.py logs Ll1)."""
    lam = float(lambda_dssim)
    loss, terms = _L1SSIM.apply(image, gt, 1.0 - lam, -lam, lam)
    return (loss, terms) if return_terms else loss


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    return _L1SSIM.apply(network_output, gt, 1.0, 0.0, 0.0)[0]


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    if window_size != 11:
        raise NotImplementedError("the fused kernel is built for the reference's window_size=11 (loss_utils.py:39)")
    if not size_average:
        raise NotImplementedError("size_average=False is not used on the training path (train.py:88)")
    return _L1SSIM.apply(img1, img2, 0.0, 1.0, 0.0)[0]
