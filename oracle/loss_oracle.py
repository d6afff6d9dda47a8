"""CPU restatement of the training loss and optimizer either side of the rasterizer -- TEST INFRASTRUCTURE ONLY
(imported by tests/, never by the product package).

* ``gaussian_window``  <- utils/loss_utils.py:23-25 (gaussian(11, 1.5), normalised)
* ``ssim_map``/``ssim`` <- utils/loss_utils.py:39-63 (five zero-padded 11x11 window means, C1 = 0.01^2, C2 = 0.03^2)
* ``l1``               <- utils/loss_utils.py:17-18
* ``loss``             <- train.py:87-88
* ``loss_grad``        <- the analytic gradient of ``loss`` w.r.t. the image (what autograd derives from the above)
* ``adam_step``        <- torch.optim.Adam as the reference configures it (scene/gaussian_model.py:160: eps=1e-15,
                          betas (0.9, 0.999), no weight decay, no amsgrad); torch 2.11 (third-party, in this image)

Pinned: tests/test_cpu_loss.py checks these against outputs of the reference's own utils/loss_utils.py
(tests/golden/loss_*.npz, made by tests/golden/make_loss_golden.py in the build container) and against
torch.optim.Adam run live on CPU.  float64 numpy throughout unless ``dtype`` says otherwise.
"""
import numpy as np


def gaussian_window(window_size=11, sigma=1.5):
    g = np.array([np.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    g = g.astype(np.float32)              # torch.Tensor([...]) stores float32 before normalising
    # torch's float32 sum of these 11 values equals the correctly rounded double sum (a sequential float32 sum is one ulp
    # lower); reproduced here so the window is bit-identical to create_window's (checked in tests/test_cpu_loss.py)
    return (g / np.float32(g.astype(np.float64).sum())).astype(np.float64)


def window_2d(w):
    """create_window (loss_utils.py:27-31): _1D_window.mm(_1D_window.t()) in float32 -- each entry one rounded product,
    so the reference's window is NOT exactly separable (relative 6e-8); the oracle keeps it, the GPU kernel runs the
    separable form (difference far below float32 noise of the sums, bounded in tests/test_gpu_loss.py)."""
    w32 = w.astype(np.float32)
    return (w32[:, None] * w32[None, :]).astype(np.float64)


def _window_mean(a, w):
    """Zero-padded window mean of a [C,H,W] array = conv2d(padding=5, groups=C) with the reference's 2-D window."""
    h = len(w) // 2
    C, H, W = a.shape
    p = np.zeros((C, H + 2 * h, W + 2 * h), dtype=a.dtype)
    p[:, h:h + H, h:h + W] = a
    w2 = window_2d(w)
    out = np.zeros_like(a)
    for i in range(len(w)):
        for j in range(len(w)):
            out = out + w2[i, j] * p[:, i:i + H, j:j + W]
    return out


def _moments(x, y, w):
    mu1, mu2 = _window_mean(x, w), _window_mean(y, w)
    e11, e22, e12 = _window_mean(x * x, w), _window_mean(y * y, w), _window_mean(x * y, w)
    return mu1, mu2, e11, e22, e12


C1 = 0.01 ** 2
C2 = 0.03 ** 2


def ssim_map(x, y):
    w = gaussian_window()
    mu1, mu2, e11, e22, e12 = _moments(x, y, w)
    s1, s2, s12 = e11 - mu1 * mu1, e22 - mu2 * mu2, e12 - mu1 * mu2
    return ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))


def ssim(x, y):
    return float(ssim_map(np.asarray(x, np.float64), np.asarray(y, np.float64)).mean())


def l1(x, y):
    return float(np.abs(np.asarray(x, np.float64) - np.asarray(y, np.float64)).mean())


def loss(x, y, lambda_dssim=0.2):
    return (1.0 - lambda_dssim) * l1(x, y) + lambda_dssim * (1.0 - ssim(x, y))


def loss_grad(x, y, w_l1, w_ssim):
    """d/dx of  w_l1 * L1(x, y) + w_ssim * SSIM(x, y)."""
    x = np.asarray(x, np.float64); y = np.asarray(y, np.float64)
    w = gaussian_window()
    mu1, mu2, e11, e22, e12 = _moments(x, y, w)
    s1, s2, s12 = e11 - mu1 * mu1, e22 - mu2 * mu2, e12 - mu1 * mu2
    A, B = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    Cc, D = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    val = A * B / (Cc * D)
    d_mu1 = 2 * mu2 * (B - A) / (Cc * D) - val * 2 * mu1 / Cc + val * 2 * mu1 / D
    d_e11 = -val / D
    d_e12 = 2 * A / (Cc * D)
    n = x.size
    g = _window_mean(d_mu1, w) + 2 * x * _window_mean(d_e11, w) + y * _window_mean(d_e12, w)
    return w_ssim * g / n + w_l1 * np.sign(x - y) / n


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-15, dtype=np.float32):
    """One torch.optim.Adam update (step = 1-based count after the update).  Returns new (p, m, v)."""
    p, g, m, v = (np.asarray(a, dtype) for a in (p, g, m, v))
    one = dtype(1.0)
    m = m + (g - m) * dtype(1.0 - beta1)
    v = v * dtype(beta2) + dtype(1.0 - beta2) * g * g
    bc1, bc2 = 1.0 - beta1 ** step, 1.0 - beta2 ** step
    denom = np.sqrt(v) / dtype(bc2 ** 0.5) + dtype(eps)
    p = p - dtype(lr / bc1) * (m / denom)
    return p.astype(dtype), m.astype(dtype), v.astype(dtype) * one
