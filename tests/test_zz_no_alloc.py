"""GPU test of gslm_jvp_cached_noalloc, the J v entry point for hosts that cannot hand the library device scratch.

Who that is: a C / Go / Java embedder with no device allocator of its own, a process whose tensor framework already
owns all of HBM (cudaMalloc would fail), or a viewer that must keep cudaMalloc's device-wide synchronisation out of
its frame loop.  For them the 48-byte-per-Gaussian tangent records live in a 12 MiB array inside the library instead
of in caller memory.

What is checked, through the C ABI (diff_gaussian_rasterization._C -> libgslm_b200.so):
  * the result equals gslm_jvp_cached's (caller-provided scratch) on the same linearised view -- bit for bit, which
    is also what tests/test_gpu_parity.py asks of the cached entry point, and a fortiori within the 1e-3 JVP
    tolerance of BASELINE.json;
  * on calls in a row that share the built-in array: a large scene, then a different, smaller scene (stale records of
    the first call are still in the array), then the large scene again with another tangent;
  * more Gaussians than GSLM_NOALLOC_MAX_GAUSSIANS are refused with an error that names the limit;
  * a warmed-up call does not shrink the free device memory.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a problem in this additive entry
point must not hide the parity tests.  It runs alone (fresh process, first call does the address look-up) as well as
after the other tests."""
import numpy as np
import pytest
import torch

import scene_synth as ss
import helpers as h

pytestmark = pytest.mark.gpu

DER_TOL = 1e-3     # JVPs (north_star)


@pytest.fixture(scope="module")
def C():
    from diff_gaussian_rasterization import _C
    _C.lib()
    return _C


def _linearised(C, name, deg, tangent_seed, view=1):
    if name == "small":
        g = ss.make_gaussians(1500, sh_degree=deg, seed=21, scale_mult=0.8, extent=1.0)
        cam = ss.hemisphere_rig(3, 150, 100)[view]
    else:
        g, cams = ss.make_config(name, sh_degree=deg, views=view + 1)
        cam = cams[view]
    c = ss.camera_dict(cam, bg=(0.1, 0.2, 0.3))
    t = ss.random_tangent(g, seed=tangent_seed)
    gd, td, ct = h.to_dev(g), h.to_dev(t), h.cam_tensors(cam, bg=c["bg"])
    return gd, td, ct, h.c_forward(C, gd, ct)


def _both(C, gd, td, ct, fwd):
    """(J v through caller scratch, J v through the built-in array), colour and inverse-depth tangents."""
    R, color, radii, geom, binning, img, invd = fwd
    args = (h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, geom, R, binning, img)
    ref_c, ref_d = C.jvp_cached(*args, want_invdepth=True)
    # outputs are the caller's; poison them so that an entry point that wrote nothing cannot pass
    out_c = torch.full((3, ct["H"], ct["W"]), float("nan"), device="cuda")
    out_d = torch.full((1, ct["H"], ct["W"]), float("nan"), device="cuda")
    C.jvp_cached_noalloc(*args, out_c, out_d)
    torch.cuda.synchronize()
    return ref_c, ref_d, out_c, out_d


def _assert_same(ref_c, ref_d, out_c, out_d, what):
    assert torch.isfinite(out_c).all() and torch.isfinite(out_d).all(), what
    assert float(ref_c.abs().max()) > 0, what                     # a non-trivial tangent image
    assert h.rel_l2(out_c.cpu().numpy(), ref_c.cpu().numpy()) < DER_TOL, what
    assert h.rel_l2(out_d.cpu().numpy(), ref_d.cpu().numpy()) < DER_TOL, what
    assert torch.equal(out_c, ref_c) and torch.equal(out_d, ref_d), what   # same kernels, same records: same bits


def test_noalloc_jvp_matches_caller_scratch_on_calls_in_a_row(C):
    big = _linearised(C, "B", 3, tangent_seed=9)
    small = _linearised(C, "small", 1, tangent_seed=5)
    assert big[0]["means3D"].shape[0] <= C.NOALLOC_MAX_GAUSSIANS
    _assert_same(*_both(C, *big), "first call (large scene)")
    _assert_same(*_both(C, *small), "second call (small scene over the large scene's stale records)")
    gd, _, ct, fwd = big
    g_host = {k: (v.cpu().numpy() if torch.is_tensor(v) else v) for k, v in gd.items()}
    td2 = h.to_dev(ss.random_tangent(g_host, seed=31))
    _assert_same(*_both(C, gd, td2, ct, fwd), "third call (large scene, new tangent)")


def test_noalloc_jvp_takes_no_device_memory(C):
    gd, td, ct, fwd = _linearised(C, "small", 2, tangent_seed=3)
    R, color, radii, geom, binning, img, invd = fwd
    args = (h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, geom, R, binning, img)
    out_c = torch.empty((3, ct["H"], ct["W"]), device="cuda")
    C.jvp_cached_noalloc(*args, out_c)            # warm-up: address look-up, module load
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        C.jvp_cached_noalloc(*args, out_c)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after >= free_before


def test_noalloc_jvp_refuses_more_gaussians_than_the_built_in_array_holds(C):
    P = C.NOALLOC_MAX_GAUSSIANS + 1
    g = ss.make_gaussians(P, sh_degree=0, seed=4)
    cam = ss.hemisphere_rig(3, 64, 48)[0]
    c = ss.camera_dict(cam, bg=(0.0, 0.0, 0.0))
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=c["bg"])
    td = h.to_dev(ss.random_tangent(g, seed=1))
    dummy = torch.zeros(256, dtype=torch.uint8, device="cuda")    # never read: the size check comes first
    radii = torch.zeros(P, dtype=torch.int32, device="cuda")
    out_c = torch.zeros((3, ct["H"], ct["W"]), device="cuda")
    with pytest.raises(RuntimeError) as e:
        C.jvp_cached_noalloc(h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, dummy, C.NumRendered(0, 0),
                             dummy, dummy, out_c)
    assert str(C.NOALLOC_MAX_GAUSSIANS) in str(e.value)
    assert float(out_c.abs().max()) == 0.0                        # nothing was launched
