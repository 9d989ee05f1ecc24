"""Per-pixel loss pieces used by the residual (reference utils/loss_utils.py:39-122): |a-b| and the SSIM map
(11x11 Gaussian window, sigma 1.5, C1 = 0.01^2, C2 = 0.03^2), written with separable convolutions in torch so
that autograd and forward-AD both work.  The fused CUDA SSIM of the reference (submodules/fused-ssim) is out of
scope: the LM driver disables SSIM (train_jvp.py:45-46,212)."""
import math

import torch
import torch.nn.functional as F


def l1_loss_per_pixel(network_output, gt):
    return torch.abs(network_output - gt)


def _window(size, sigma, like):
    g = torch.tensor([math.exp(-(x - size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(size)],
                     dtype=like.dtype, device=like.device)
    return g / g.sum()


def _blur(x, w1d):
    C = x.shape[-3]
    k = w1d.numel()
    wx = w1d.view(1, 1, 1, k).expand(C, 1, 1, k)
    wy = w1d.view(1, 1, k, 1).expand(C, 1, k, 1)
    x = F.conv2d(x, wx, padding=(0, k // 2), groups=C)
    return F.conv2d(x, wy, padding=(k // 2, 0), groups=C)


def ssim_per_pixel(img1, img2, window_size=11):
    """SSIM map with the shape of the inputs ([B,3,H,W] or [3,H,W])."""
    squeeze = img1.dim() == 3
    if squeeze:
        img1, img2 = img1[None], img2[None]
    w = _window(window_size, 1.5, img1)
    mu1, mu2 = _blur(img1, w), _blur(img2, w)
    mu1_sq, mu2_sq, mu12 = mu1 * mu1, mu2 * mu2, mu1 * mu2
    s1 = _blur(img1 * img1, w) - mu1_sq
    s2 = _blur(img2 * img2, w) - mu2_sq
    s12 = _blur(img1 * img2, w) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    m = ((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))
    return m[0] if squeeze else m
