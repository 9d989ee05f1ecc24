"""CPU: the C-ABI library loads and exports every symbol include/gslm_b200.h declares; size queries (host-only
functions) behave; the Python surface has the reference's names and raises the reference's exceptions.
No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gaussian-splatting-lm_b200", "libgslm_b200.so")
HEADER = os.path.join(ROOT, "include", "gslm_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gslm_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import __graft_entry__
        __graft_entry__.build()
    return ctypes.CDLL(LIB)


def test_every_declared_symbol_is_exported(lib):
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gslm_b200.h but not exported"


def test_binding_lists_the_same_symbols():
    from diff_gaussian_rasterization import _C
    assert sorted(_C.EXPORTS) == _declared_symbols()


def test_ctypes_mirrors_match_the_compiled_struct_layouts(lib):
    """The ctypes Structures of the binding against sizeof / offsetof as the library was compiled."""
    from diff_gaussian_rasterization import _C
    lib.gslm_sizeof.restype = ctypes.c_size_t
    lib.gslm_sizeof.argtypes = [ctypes.c_char_p]
    lib.gslm_camera_offsetof.restype = ctypes.c_size_t
    lib.gslm_camera_offsetof.argtypes = [ctypes.c_char_p]
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    names = re.findall(r"typedef struct (Gslm[A-Za-z]+)", src)
    assert len(names) >= 9
    for n in names:
        assert lib.gslm_sizeof(n.encode()) == ctypes.sizeof(getattr(_C, n)) > 0, n
    assert lib.gslm_sizeof(b"NoSuchStruct") == 0
    for f, _ in _C.GslmCamera._fields_:
        assert lib.gslm_camera_offsetof(f.encode()) == getattr(_C.GslmCamera, f).offset, f
    assert lib.gslm_abi_version() >= 2


def test_size_queries(lib):
    for n in ("gslm_geom_bytes", "gslm_image_bytes", "gslm_binning_bytes", "gslm_backward_scratch_bytes",
              "gslm_jvp_scratch_bytes", "gslm_jtjv_scratch_bytes", "gslm_binning_temp_bytes"):
        if hasattr(lib, n):
            getattr(lib, n).restype = ctypes.c_size_t
    assert lib.gslm_abi_version() >= 1
    g0, g1 = lib.gslm_geom_bytes(ctypes.c_int32(0)), lib.gslm_geom_bytes(ctypes.c_int32(1000))
    assert g1 > g0 and g1 % 128 == 0
    # per-pixel state: final_T + n_contrib + n_contrib_c = 12 B / pixel (+ per-tile ranges / compact offsets)
    i = lib.gslm_image_bytes(ctypes.c_int32(1920), ctypes.c_int32(1080))
    assert 12 * 1920 * 1080 <= i <= 13 * 1920 * 1080
    b0, b1 = lib.gslm_binning_bytes(ctypes.c_int64(0)), lib.gslm_binning_bytes(ctypes.c_int64(1 << 20))
    assert b1 > b0 and b1 >= (1 << 20) * (48 + 4 + 4)
    assert lib.gslm_backward_scratch_bytes(ctypes.c_int64(1000)) >= 1000 * 40


def test_python_surface_names_and_errors():
    import diff_gaussian_rasterization as dgr
    fields = dgr.GaussianRasterizationSettings._fields
    # DGR/diff_gaussian_rasterization/__init__.py:287-300
    assert fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
                      "projmatrix", "sh_degree", "campos", "prefiltered", "debug", "antialiasing")
    rs = dgr.GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                           torch.zeros(3), False, False, False)
    r = dgr.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), shs=None, colors_precomp=None, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 6))
    # no CPU path: host tensors are refused loudly
    with pytest.raises(RuntimeError, match="CUDA"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))


def test_utils_tangent_helpers():
    import torch.autograd.forward_ad as fwAD
    from diff_gaussian_rasterization.utils import get_tangent, has_tangent
    x = torch.ones(3)
    assert not has_tangent(x) and not has_tangent(1.0)
    assert get_tangent(2.0) == 0.0 and torch.equal(get_tangent(x), torch.zeros(3))
    with fwAD.dual_level():
        d = fwAD.make_dual(x, 2 * x)
        assert has_tangent(d) and has_tangent([x, d])
        assert torch.equal(get_tangent(d), 2 * x)


def test_argument_validation_without_a_gpu(lib):
    """Entry points reject malformed arguments before touching the device: checked here on the CPU-only box."""
    lib.gslm_last_error.restype = ctypes.c_char_p
    lib.gslm_multimem_allreduce_f32.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int32, ctypes.c_int32,
                                                ctypes.c_void_p]
    assert lib.gslm_multimem_allreduce_f32(None, 1024, 0, 2, None) != 0            # no multicast mapping
    assert b"multimem" in lib.gslm_last_error()
    assert lib.gslm_multimem_allreduce_f32(ctypes.c_void_p(4096), 1023, 0, 2, None) != 0   # n not a multiple of 4
    assert lib.gslm_multimem_allreduce_f32(ctypes.c_void_p(4096), 1024, 2, 2, None) != 0   # rank outside the world
    assert lib.gslm_multimem_allreduce_f32(ctypes.c_void_p(4100), 1024, 0, 2, None) != 0   # misaligned buffer
    lib.gslm_forward_batch.argtypes = [ctypes.c_void_p] + [ctypes.c_int32] + [ctypes.c_void_p] * 10
    assert lib.gslm_forward_batch(None, 1, None, None, None, None, None, None, None, None, None, None) != 0
    lib.gslm_densification_stats.argtypes = [ctypes.c_int32] + [ctypes.c_void_p] * 6
    assert lib.gslm_densification_stats(10, None, None, None, None, None, None) != 0
    assert lib.gslm_densification_stats(0, None, None, None, None, None, None) == 0      # empty model: nothing to do
    assert lib.gslm_abi_version() >= 3
