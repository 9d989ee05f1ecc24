"""Parameter-space generalized vectors (reference solver/gaussian_model_state.py): the seven tensors
xyz[P,3], features_dc[P,1,3], features_rest[P,M-1,3], scaling[P,3], rotation[P,4], opacity[P,1], exposure[C,3,4]
with + - * dot, per-group damping and masks.

B200 difference: a state created here owns ONE packed fp32 buffer (`flat`) and the seven attributes are views
into it, so every vector operation is a single kernel over 59 P floats (the reference launches 7+ per op) and
the buffer doubles as the NCCL payload of the per-iteration allreduce.  States built from seven arbitrary
tensors (the reference constructor) still work; they are packed lazily on first arithmetic."""
import torch

FIELDS = ("xyz_grad", "features_dc_grad", "features_rest_grad", "scaling_grad", "rotation_grad", "opacity_grad",
          "exposure_grad")
_DAMPS = ("xyz_damp", "features_dc_damp", "features_rest_damp", "scaling_damp", "rotation_damp", "opacity_damp",
          "exposure_damp")
_MASKS = ("mask_xyz", "mask_features_dc", "mask_features_rest", "mask_scaling", "mask_rotation", "mask_opacity",
          "mask_exposure")
_MODEL = ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity", "_exposure")


class GaussianModelDampMatrix:
    """Per-parameter-group diagonal damping."""

    def __init__(self, xyz_damp, features_dc_damp, features_rest_damp, scaling_damp, rotation_damp, opacity_damp,
                 exposure_damp):
        for n, v in zip(_DAMPS, (xyz_damp, features_dc_damp, features_rest_damp, scaling_damp, rotation_damp,
                                 opacity_damp, exposure_damp)):
            setattr(self, n, v)

    def values(self):
        return tuple(getattr(self, n) for n in _DAMPS)

    def __neg__(self):
        return GaussianModelDampMatrix(*(-v for v in self.values()))


class GaussianModelParamGroupMask:
    """True = that parameter group is frozen (zeroed in every state built with this mask)."""

    def __init__(self, mask_xyz=False, mask_features_dc=False, mask_features_rest=False, mask_scaling=False,
                 mask_rotation=False, mask_opacity=False, mask_exposure=False):
        for n, v in zip(_MASKS, (mask_xyz, mask_features_dc, mask_features_rest, mask_scaling, mask_rotation,
                                 mask_opacity, mask_exposure)):
            setattr(self, n, v)

    def values(self):
        return tuple(getattr(self, n) for n in _MASKS)


class GaussianModelSplatMask:
    """Per-splat mask: rows selected by mask_out_filter are zeroed."""

    def __init__(self, mask_out_filter):
        self.mask_out_filter = mask_out_filter


def _pack(tensors):
    """One flat fp32 buffer holding `tensors` back to back + the views into it."""
    sizes = [t.numel() for t in tensors]
    flat = torch.empty(sum(sizes), dtype=tensors[0].dtype, device=tensors[0].device)
    views, off = [], 0
    for t, n in zip(tensors, sizes):
        v = flat[off:off + n].view(t.shape)
        v.copy_(t)
        views.append(v)
        off += n
    return flat, views


class GaussianModelState:
    """Updates to (or gradients of) the Gaussian parameters as one generalized vector."""

    def __init__(self, xyz_grad, features_dc_grad, features_rest_grad, scaling_grad, rotation_grad, opacity_grad,
                 exposure_grad, param_mask=None, splat_mask=None, _flat=None):
        vals = (xyz_grad, features_dc_grad, features_rest_grad, scaling_grad, rotation_grad, opacity_grad,
                exposure_grad)
        for n, v in zip(FIELDS, vals):
            setattr(self, n, v)
        self.flat = _flat
        # masks zero groups / rows in place, in the constructor only (reference :64-87)
        if param_mask is not None:
            assert isinstance(param_mask, GaussianModelParamGroupMask), \
                "param_mask must be an instance of GaussianModelParamGroupMask"
            for n, m in zip(FIELDS, param_mask.values()):
                if m:
                    getattr(self, n).zero_()
        if splat_mask is not None:
            for n in FIELDS[:6]:
                getattr(self, n)[splat_mask.mask_out_filter] = 0
            # exposure is per camera, not per splat; the reference indexes it with the splat filter as well,
            # which only works when the shapes happen to allow it -- left untouched here

    # ---- construction ----
    @classmethod
    def _from_shapes(cls, like, fill=None, zero=True):
        sizes = [t.numel() for t in like]
        flat = (torch.zeros if zero else torch.empty)(sum(sizes), dtype=torch.float32, device=like[0].device)
        views, off = [], 0
        for t, n in zip(like, sizes):
            views.append(flat[off:off + n].view(t.shape))
            off += n
        return cls(*views, _flat=flat)

    @classmethod
    def from_gaussians(cls, gaussians, param_mask=None, splat_mask=None, zero=True):
        """A zero state shaped like the model's parameters.  zero=False: only the exposure block is cleared -- for
        callers that overwrite the six Gaussian blocks right away (saves a 59 P-float memset per product)."""
        s = cls._from_shapes([getattr(gaussians, n) for n in _MODEL], zero=zero)
        if not zero:
            s.exposure_grad.zero_()
        return cls(*s.tensors(), param_mask=param_mask, splat_mask=splat_mask, _flat=s.flat)

    @classmethod
    def from_gaussians_grad(cls, gaussians, param_mask=None, splat_mask=None):
        params = [getattr(gaussians, n) for n in _MODEL]
        grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in params]
        flat, views = _pack(grads)
        return cls(*views, param_mask=param_mask, splat_mask=splat_mask, _flat=flat)

    @classmethod
    def from_flat(cls, flat, like):
        """Views of `flat` shaped like the seven tensors of `like` (a state or a list of tensors)."""
        shapes = [t.shape for t in (like.tensors() if isinstance(like, GaussianModelState) else like)]
        views, off = [], 0
        for sh in shapes:
            n = 1
            for d in sh:
                n *= d
            views.append(flat[off:off + n].view(sh))
            off += n
        return cls(*views, _flat=flat)

    def tensors(self):
        return tuple(getattr(self, n) for n in FIELDS)

    def packed(self):
        """The flat buffer (packs once if the state was built from loose tensors)."""
        if self.flat is None:
            self.flat, views = _pack(self.tensors())
            for n, v in zip(FIELDS, views):
                setattr(self, n, v)
        return self.flat

    # ---- reference API ----
    @property
    def length(self):
        return sum(t.numel() for t in self.tensors())

    def as_1d_tensor(self):
        return self.packed().clone()

    def load_1d_tensor(self, T):
        assert T.shape[0] == self.length, "Input tensor must match the length of the model state"
        new = GaussianModelState.from_flat(T, self)
        self.__init__(*new.tensors(), _flat=T)

    def index_to_desc(self, index):
        assert index < self.length, "Index out of bounds for GaussianModelState"
        for n in FIELDS:
            t = getattr(self, n)
            if index < t.numel():
                coords = []
                for dim in reversed(t.shape):
                    coords.append(index % dim)
                    index //= dim
                return n, tuple(reversed(coords))
            index -= t.numel()

    def _like(self, flat):
        return GaussianModelState.from_flat(flat, self)

    def __add__(self, other):
        return self._like(self.packed() + other.packed())

    def __sub__(self, other):
        return self._like(self.packed() - other.packed())

    def __mul__(self, other):
        if isinstance(other, (int, float)):
            return self._like(self.packed() * other)
        if isinstance(other, GaussianModelDampMatrix):
            out = self._like(self.packed().clone())
            for t, d in zip(out.tensors(), other.values()):
                t.mul_(d)
            return out
        raise TypeError(f"Can only multiply by scalar values, not {type(other)}")

    def __rmul__(self, other):
        return self.__mul__(other)

    def dot_tensor(self, other, damp=1.0):
        """<self, other>_damp as a 0-dim device tensor (no host synchronisation)."""
        if isinstance(damp, (int, float)):
            return torch.dot(self.packed(), other.packed()) * damp
        if isinstance(damp, GaussianModelDampMatrix):
            s = 0.0
            for a, b, d in zip(self.tensors(), other.tensors(), damp.values()):
                if a.numel():
                    s = s + d * torch.dot(a.reshape(-1), b.reshape(-1))
            return s
        raise TypeError(f"damp must be a scalar or GaussianModelDampMatrix, not {type(damp)}")

    def dot(self, other, damp=1.0):
        s = self.dot_tensor(other, damp)
        return s.item() if isinstance(s, torch.Tensor) else float(s)
