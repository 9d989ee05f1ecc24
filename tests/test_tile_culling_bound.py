"""CPU check of the bound behind exact tile culling (csrc/binning.cu, tile_can_blend): a numpy float32 restatement of
the test is run on random (Gaussian, tile) pairs and every pair it would cull is verified pixel by pixel against the
reference's blend test (forward.cu:354-364: power = -0.5 (A dx^2 + C dy^2) - B dx dy; skip if power > 0 or
o exp(power) < 1/255), evaluated in float32 like the kernel.  It must never cull a tile with a blendable pixel, and it
should cull most tiles that have none (tightness, so the test cannot pass by never culling)."""
import numpy as np

f32 = np.float32


def tile_can_blend(x, y, A, B, C, o, tx, ty, W, H):
    """float32 restatement of tile_can_blend (same formulas; numpy rounds every operation separately)."""
    tau = np.log(f32(255.0) * o).astype(f32) if o > 0 else f32(-np.inf)
    px0, px1 = f32(tx * 16), f32(min(tx * 16 + 15, W - 1))
    py0, py1 = f32(ty * 16), f32(min(ty * 16 + 15, H - 1))
    dx_lo, dx_hi, dy_lo, dy_hi = x - px1, x - px0, y - py1, y - py0
    cx = min(max(f32(0), dx_lo), dx_hi)
    cy = min(max(f32(0), dy_lo), dy_hi)
    if not (A > 0 and C > 0):
        return True

    def parabola_min(K, L, lo, hi):
        t = min(max(-L / K, lo), hi)
        return t * (f32(0.5) * K * t + L)

    qmin = f32(0)
    if cx != 0 or cy != 0:
        qmin = f32(np.inf)
        if cx != 0:
            qmin = parabola_min(C, B * cx, dy_lo, dy_hi) + f32(0.5) * A * cx * cx
        if cy != 0:
            qmin = min(qmin, parabola_min(A, B * cy, dx_lo, dx_hi) + f32(0.5) * C * cy * cy)
    DX, DY = max(abs(dx_lo), abs(dx_hi)), max(abs(dy_lo), abs(dy_hi))
    smax = abs(B) * DX * DY + f32(0.5) * (A * DX * DX + C * DY * DY)
    slack = f32(4e-6) * smax + f32(1e-5) * abs(tau) + f32(1e-3)
    return not (qmin - slack > tau)


def any_pixel_blends(x, y, A, B, C, o, tx, ty, W, H):
    px = np.arange(tx * 16, min(tx * 16 + 16, W), dtype=np.float32)[None, :]
    py = np.arange(ty * 16, min(ty * 16 + 16, H), dtype=np.float32)[:, None]
    dx, dy = x - px, y - py
    power = f32(-0.5) * (A * dx * dx + C * dy * dy) - B * dx * dy
    alpha = np.minimum(f32(0.99), o * np.exp(power, dtype=np.float32))
    return bool(((power <= 0) & (alpha >= f32(1.0 / 255.0))).any())


def test_bound_never_culls_a_blendable_tile_and_is_tight():
    rng = np.random.default_rng(12)
    W, H = 640, 360
    culled = kept_empty = blendable = 0
    for _ in range(4000):
        # covariance of a projected Gaussian: rotated ellipse with the 0.3 px dilation of forward.cu:218-222
        s1, s2 = np.exp(rng.uniform(-1.5, 4.0, 2))
        th = rng.uniform(0, np.pi)
        R = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
        cov = R @ np.diag([s1 * s1, s2 * s2]) @ R.T + 0.3 * np.eye(2)
        det = cov[0, 0] * cov[1, 1] - cov[0, 1] ** 2
        A, B, C = f32(cov[1, 1] / det), f32(-cov[0, 1] / det), f32(cov[0, 0] / det)
        o = f32(rng.choice([rng.uniform(0.001, 0.02), rng.uniform(0.02, 1.0)]))
        x, y = f32(rng.uniform(-40, W + 40)), f32(rng.uniform(-40, H + 40))
        lam = 0.5 * (cov[0, 0] + cov[1, 1]) + np.sqrt(max(0.1, (0.5 * (cov[0, 0] + cov[1, 1])) ** 2 - det))
        rad = int(np.ceil(3.0 * np.sqrt(lam)))                     # forward.cu:240-243
        gx, gy = (W + 15) // 16, (H + 15) // 16
        x0, x1 = min(gx, max(0, int((x - rad) / 16))), min(gx, max(0, int((x + rad + 15) / 16)))
        y0, y1 = min(gy, max(0, int((y - rad) / 16))), min(gy, max(0, int((y + rad + 15) / 16)))
        tiles = [(tx, ty) for ty in range(y0, y1) for tx in range(x0, x1)]
        if len(tiles) > 12:
            tiles = [tiles[i] for i in rng.choice(len(tiles), 12, replace=False)]
        for tx, ty in tiles:
            keep = tile_can_blend(x, y, A, B, C, o, tx, ty, W, H)
            hit = any_pixel_blends(x, y, A, B, C, o, tx, ty, W, H)
            assert keep or not hit, (x, y, A, B, C, o, tx, ty)
            blendable += hit
            culled += not keep
            kept_empty += keep and not hit
    assert blendable > 2000 and culled > 2000
    # tightness: of the tiles no pixel can blend, the bound (continuous box minimum + slack) keeps few
    assert kept_empty < 0.25 * (culled + kept_empty), (culled, kept_empty)
