"""Exact tile culling (GslmCamera::tile_culling = 1, csrc/binning.cu): dropping the (tile, Gaussian) instances that
provably stay below alpha = 1/255 at every pixel of the tile must change NOTHING the reference's forward pass defines
-- image, inverse depth, radii, final_T -- nor the compact linearisation cache the derivative kernels read (so every
gradient / tangent / J^T J v is unchanged too); only num_rendered and the sorted lists get shorter.  Compared bit for
bit against mode 0, whose binning is pinned on the reference's fixture and kernels (test_gpu_parity.py,
test_gpu_vs_reference.py)."""
import numpy as np
import pytest
import torch

import helpers as h
import scene_synth as ss

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def C():
    from diff_gaussian_rasterization import _C
    return _C


KEYS = ("final_T", "cstart", "n_contrib_c", "records", "masks", "cid", "g_list", "ranges", "point_list")


def _forward(C, gd, ct, aa, culling, P, W, H):
    """forward pass + a snapshot of its state (the sort scratch is ONE buffer shared by all forward calls: the views
    into it must be copied before the next call)"""
    C.TILE_CULLING = culling
    try:
        fwd = h.c_forward(C, gd, ct, antialiasing=aa, keep_sort_scratch=True)
        v = C.buffer_view(P, W, H, fwd[0], fwd[3], fwd[4], fwd[5], C.LAST_SORT_SCRATCH)
        return fwd, {k: (v[k].clone() if k in KEYS else v[k]) for k in KEYS + ("num_cached",)}
    finally:
        C.TILE_CULLING = False


CASES = [  # P, degree, antialiasing, scale multiplier, W, H, opacity transform
    (10000, 3, False, 1.0, 128, 128, None),
    (10000, 1, True, 1.0, 200, 120, None),
    (3000, 0, False, 4.0, 257, 131, None),                       # large splats, ragged image size
    (20000, 2, False, 0.6, 320, 240, lambda o: o * 0.05),        # faint: most of the 3-sigma square is below 1/255
    (5000, 2, True, 2.0, 160, 96, lambda o: np.maximum(o, 0.97)),  # nearly opaque: early termination everywhere
]


@pytest.mark.parametrize("P,deg,aa,smul,W,H,opf", CASES)
def test_culling_changes_nothing_but_the_list_lengths(C, P, deg, aa, smul, W, H, opf):
    g = ss.make_gaussians(P, sh_degree=deg, seed=40 + deg, scale_mult=smul)
    # anisotropy: stretch one axis of every third Gaussian by 8x (thin, rotated ellipses are the hard case for a
    # box-vs-ellipse test)
    g["scales"][::3, 0] *= 8.0
    if opf is not None:
        g["opacities"] = opf(g["opacities"]).astype(np.float32)
    cam = ss.hemisphere_rig(4, W, H)[1]
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=(0.1, 0.2, 0.3))
    f0, v0 = _forward(C, gd, ct, aa, False, P, W, H)
    f1, v1 = _forward(C, gd, ct, aa, True, P, W, H)
    R0, R1 = int(f0[0]), int(f1[0])
    assert 0 < R1 < R0
    assert torch.equal(f0[1], f1[1]) and torch.equal(f0[6], f1[6]) and torch.equal(f0[2], f1[2])
    assert v0["num_cached"] == v1["num_cached"] > 0
    for k in ("final_T", "cstart", "n_contrib_c", "records", "masks", "cid", "g_list"):
        assert torch.equal(v0[k], v1[k]), k
    # the culled lists are subsequences of the reference's lists, tile by tile, and keep every blended instance
    r0, r1 = v0["ranges"].cpu().numpy().astype(np.int64), v1["ranges"].cpu().numpy().astype(np.int64)
    p0, p1 = v0["point_list"].cpu().numpy(), v1["point_list"].cpu().numpy()
    cid, cstart = v1["cid"].cpu().numpy(), v1["cstart"].cpu().numpy().astype(np.int64)
    assert ((r1[:, 1] - r1[:, 0]) <= (r0[:, 1] - r0[:, 0])).all()
    for t in np.linspace(0, len(r0) - 1, 12).astype(int):
        full, kept = p0[r0[t, 0]:r0[t, 1]].tolist(), p1[r1[t, 0]:r1[t, 1]].tolist()
        it = iter(full)
        assert all(any(x == y for y in it) for x in kept)
        it = iter(kept)
        assert all(any(x == y for y in it) for x in cid[cstart[t]:cstart[t + 1]].tolist())
    # gradients come from the cache: identical bits
    u = torch.randn(3, H, W, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    g0 = h.c_backward(C, gd, ct, f0, u, antialiasing=aa)
    g1 = h.c_backward(C, gd, ct, f1, u, antialiasing=aa)
    assert all((a is None and b is None) or torch.equal(a, b) for a, b in zip(g0, g1))
    print(f"instances {R0} -> {R1} ({R1 / R0:.2f}), cached {v1['num_cached']}")


def test_culling_at_full_size_is_bit_identical(C):
    """1M Gaussians at 1920x1080 (BASELINE config C's view size): same image bits, same cache."""
    g, cams = ss.make_config("C", sh_degree=3, views=2)
    gd, ct = h.to_dev(g), h.cam_tensors(cams[1])
    f0, v0 = _forward(C, gd, ct, False, False, 1_000_000, 1920, 1080)
    f1, v1 = _forward(C, gd, ct, False, True, 1_000_000, 1920, 1080)
    assert int(f1[0]) < int(f0[0]) and f0[0].cached == f1[0].cached
    assert torch.equal(f0[1], f1[1]) and torch.equal(f0[6], f1[6])
    for k in ("final_T", "cstart", "n_contrib_c", "records", "masks", "cid", "g_list"):
        assert torch.equal(v0[k], v1[k]), k
    print(f"instances {int(f0[0])} -> {int(f1[0])}")
