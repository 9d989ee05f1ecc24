"""Forward-AD helpers with the names and behaviour of DGR/diff_gaussian_rasterization/utils.py."""
import torch
import torch.autograd.forward_ad as fwAD


def has_tangent(x):
    """True iff x (a tensor, or a list of them) carries a forward-mode tangent; floats never do."""
    if isinstance(x, (float, int)):
        return False
    if isinstance(x, (list, tuple)):
        return any(has_tangent(e) for e in x)
    return fwAD.unpack_dual(x).tangent is not None


def get_tangent(x):
    """Tangent of x: 0.0 for floats, zeros_like for tensors without one, element-wise for lists."""
    if isinstance(x, (float, int)):
        return 0.0
    if isinstance(x, (list, tuple)):
        return [get_tangent(e) for e in x]
    if isinstance(x, torch.Tensor):
        t = fwAD.unpack_dual(x).tangent
        return t if t is not None else torch.zeros_like(x)
    raise ValueError(f"Unsupported type for tangent extraction: {type(x)}")


def primal(x):
    """Strip a forward-mode tangent (tensors or lists of tensors)."""
    if isinstance(x, (list, tuple)):
        return [primal(e) for e in x]
    if isinstance(x, torch.Tensor):
        return fwAD.unpack_dual(x).primal
    return x
