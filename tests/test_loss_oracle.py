"""CPU: pins oracle/loss_oracle.py on outputs of the reference's own Python loss functions
(tests/golden/loss_fixture.npz, produced by tests/golden/make_loss_golden.py from utils/loss_utils.py:91-117 and
solver/batch_training_loss.py:10-31), and checks its self-consistency (finite differences, adjoint identity)."""
import os

import numpy as np
import pytest

import loss_oracle as lo
from helpers import rel_l2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(ROOT, "tests", "golden", "loss_fixture.npz"))


def _case(fx, name):
    g = lambda k: fx[f"{name}_{k}"] if f"{name}_{k}" in fx.files else None
    return g


def test_ssim_map_matches_reference(fx):
    m = lo.ssim_per_pixel(fx["ssim_x"], fx["ssim_y"])
    assert np.abs(m - fx["ssim_map"]).max() < 2e-5           # the reference map is fp32 with sigma^2 = E - mu^2 cancellation


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_loss_blocks_match_reference(fx, name):
    g = _case(fx, name)
    r1, r2 = lo.loss_block(g("img"), g("mask"), g("gt"), g("alphas"), g("betas"), bool(g("disable_ssim")))
    assert rel_l2(r1, g("r1")) < 1e-6
    assert rel_l2(r2, g("r2")) < 5e-6                        # measured 3e-7 .. 6e-7 (the reference is fp32)


@pytest.mark.parametrize("name", ["a", "b"])
def test_loss_derivatives_match_reference_autograd(fx, name):
    g = _case(fx, name)
    t1, t2 = lo.loss_block_jvp(g("img"), g("mask"), g("gt"), g("alphas"), g("betas"), g("t_img"))
    gx = lo.loss_block_vjp(g("img"), g("mask"), g("gt"), g("alphas"), g("betas"), g("u1"), g("u2"))
    assert rel_l2(t1, g("t1")) < 1e-6                        # measured 5e-8
    assert rel_l2(t2, g("t2")) < 2e-5                        # measured 1.4e-6
    assert rel_l2(gx, g("gx")) < 5e-6                        # measured 1.2e-7


def test_oracle_jvp_is_the_derivative_and_vjp_its_adjoint():
    rng = np.random.default_rng(3)
    B, H, W = 2, 19, 23
    img, gt = rng.random((B, 3, H, W)), rng.random((B, 3, H, W))
    mask = (rng.random((B, 1, H, W)) > 0.2).astype(np.float64)
    al, be = np.array([0.3, 0.4]), np.array([0.2, 0.1])
    t = rng.standard_normal((B, 3, H, W))
    h = 1e-6
    p = lo.loss_block(img + h * t, mask, gt, al, be)
    m = lo.loss_block(img - h * t, mask, gt, al, be)
    t1, t2 = lo.loss_block_jvp(img, mask, gt, al, be, t)
    assert rel_l2(t1, (p[0] - m[0]) / (2 * h)) < 1e-6
    assert rel_l2(t2, (p[1] - m[1]) / (2 * h)) < 1e-6
    u1, u2 = rng.standard_normal((B, 3, H, W)), rng.standard_normal((B, 3, H, W))
    gx = lo.loss_block_vjp(img, mask, gt, al, be, u1, u2)
    lhs = (t1 * u1).sum() + (t2 * u2).sum()
    assert abs(lhs - (gx * t).sum()) < 1e-10 * max(1.0, abs(lhs))
