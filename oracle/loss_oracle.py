"""CPU oracle of the residual side of the LM path: the per-pixel loss blocks and their derivatives.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; nothing under gaussian-splatting-lm_b200/ (the product) does.

Restates, in numpy float64 (the window itself is rounded to float32 exactly as the reference builds it):
  gaussian / create_window          /root/reference/utils/loss_utils.py:49-57
  _ssim_per_pixel                   /root/reference/utils/loss_utils.py:98-117   (F.conv2d, 11x11 window, zero padding 5)
  l1_loss_per_pixel                 /root/reference/utils/loss_utils.py:43-44
  compute_batch_loss_block          /root/reference/solver/batch_training_loss.py:10-31
and the forward-mode / reverse-mode derivatives PyTorch derives from them (torch.abs has derivative sign(x), 0 at 0).

Parity status: PINNED.  tests/golden/loss_fixture.npz holds outputs of the reference's own Python functions
(imported from /root/reference in the build container by tests/golden/make_loss_golden.py, CPU torch fp32):
both residual blocks, a forward-AD tangent and an autograd cotangent pull-back;
tests/test_loss_oracle.py checks this restatement against them.
"""
import math

import numpy as np

C1 = 0.01 ** 2
C2 = 0.03 ** 2
EPS = 1e-6


def gaussian_window_1d(size=11, sigma=1.5):
    """loss_utils.py:49-51: float32 tensor of exp values, normalised in float32 (torch's sum of 11 floats is the
    correctly rounded one, which a float64 accumulation reproduces bit for bit)."""
    g = np.array([math.exp(-(x - size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(size)], dtype=np.float32)
    return g / np.float32(g.astype(np.float64).sum())


def window_2d(size=11):
    """loss_utils.py:53-57: outer product of the 1-D window, in float32."""
    w = gaussian_window_1d(size)
    return (w[:, None] * w[None, :]).astype(np.float32)


def blur(x, w2d=None):
    """F.conv2d(x, window, padding=size//2, groups=channel) on [..., H, W] (zero padding; cross-correlation with a
    symmetric window)."""
    w = (window_2d() if w2d is None else w2d).astype(np.float64)
    k = w.shape[0]
    r = k // 2
    x = np.asarray(x, dtype=np.float64)
    H, W = x.shape[-2:]
    pad = np.zeros(x.shape[:-2] + (H + 2 * r, W + 2 * r))
    pad[..., r:r + H, r:r + W] = x
    out = np.zeros_like(x)
    for i in range(k):
        for j in range(k):
            out += w[i, j] * pad[..., i:i + H, j:j + W]
    return out


def _stats(x, y):
    mu1, mu2 = blur(x), blur(y)
    s1 = blur(x * x) - mu1 * mu1
    s2 = blur(y * y) - mu2 * mu2
    s12 = blur(x * y) - mu1 * mu2
    return mu1, mu2, s1, s2, s12


def ssim_per_pixel(x, y):
    """loss_utils.py:98-117."""
    mu1, mu2, s1, s2, s12 = _stats(x, y)
    return ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))


def _ssim_partials(x, y):
    """SSIM map m and its partials with respect to the three x-dependent window statistics
    mu1 = G*x, E11 = G*(x x), E12 = G*(x y) (sigma1^2 = E11 - mu1^2, sigma12 = E12 - mu1 mu2)."""
    mu1, mu2, s1, s2, s12 = _stats(x, y)
    A1, A2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    B1, B2 = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    m = A1 * A2 / (B1 * B2)
    dm_dmu1 = 2 * mu2 * (A2 - A1) / (B1 * B2) - 2 * mu1 * m * (1 / B1 - 1 / B2)
    dm_dE11 = -m / B2
    dm_dE12 = 2 * A1 / (B1 * B2)
    return m, dm_dmu1, dm_dE11, dm_dE12


def loss_block(images, alpha_masks, gt, alphas, betas, disable_ssim=False):
    """compute_batch_loss_block (batch_training_loss.py:10-31): returns (Ll1_per_pixel, ssim_loss_per_pixel)."""
    x = np.asarray(images, dtype=np.float64)
    if alpha_masks is not None:
        x = x * np.asarray(alpha_masks, dtype=np.float64)
    y = np.asarray(gt, dtype=np.float64)
    if disable_ssim:
        r = x - y
        return r, r
    a = np.asarray(alphas, dtype=np.float64).reshape(-1, 1, 1, 1)
    b = np.asarray(betas, dtype=np.float64).reshape(-1, 1, 1, 1)
    m = ssim_per_pixel(x, y)
    return a * np.sqrt(np.abs(x - y) + EPS), b * np.sqrt(np.abs(1.0 - m) + EPS)


def _linearisation(images, alpha_masks, gt, alphas, betas):
    x = np.asarray(images, dtype=np.float64)
    mask = None if alpha_masks is None else np.asarray(alpha_masks, dtype=np.float64)
    if mask is not None:
        x = x * mask
    y = np.asarray(gt, dtype=np.float64)
    a = np.asarray(alphas, dtype=np.float64).reshape(-1, 1, 1, 1)
    b = np.asarray(betas, dtype=np.float64).reshape(-1, 1, 1, 1)
    m, pa, pb, pc = _ssim_partials(x, y)
    d1 = a * np.sign(x - y) / (2 * np.sqrt(np.abs(x - y) + EPS))
    d2 = -b * np.sign(1.0 - m) / (2 * np.sqrt(np.abs(1.0 - m) + EPS))
    return x, y, mask, d1, d2, pa, pb, pc


def loss_block_jvp(images, alpha_masks, gt, alphas, betas, t_images):
    """Tangents of the two blocks for an image tangent t_images (what forward AD through
    compute_batch_loss_block returns)."""
    x, y, mask, d1, d2, pa, pb, pc = _linearisation(images, alpha_masks, gt, alphas, betas)
    dx = np.asarray(t_images, dtype=np.float64)
    if mask is not None:
        dx = dx * mask
    dm = pa * blur(dx) + pb * blur(2 * x * dx) + pc * blur(y * dx)
    return d1 * dx, d2 * dm


def loss_block_vjp(images, alpha_masks, gt, alphas, betas, u1, u2):
    """Pull-back of cotangents (u1, u2) on the two blocks to the images (what autograd returns)."""
    x, y, mask, d1, d2, pa, pb, pc = _linearisation(images, alpha_masks, gt, alphas, betas)
    w = d2 * np.asarray(u2, dtype=np.float64)
    gx = d1 * np.asarray(u1, dtype=np.float64) + blur(pa * w) + 2 * x * blur(pb * w) + y * blur(pc * w)
    if mask is not None:
        gx = gx * mask
    return gx
