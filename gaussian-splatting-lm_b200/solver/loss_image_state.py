"""Residual-space generalized vectors (reference solver/loss_image_state.py): per batch of views three image
stacks -- the L1-type block, the SSIM-type block and the depth block -- with + - * dot, forward-AD unpacking and
the vector-Jacobian `backward`."""
import torch
import torch.autograd.forward_ad as fwAD

_BLOCKS = ("Ll1_per_pixel", "ssim_loss_per_pixel", "Ll1depth_per_pixel")


def _sumsq(t):
    return torch.linalg.vector_norm(t.flatten(), ord=2) ** 2


class BatchLossImageState:
    """Losses of several images; same-type losses of all images are contiguous ([B,3,H,W] stacks)."""

    def __init__(self, Ll1_per_pixel, ssim_loss_per_pixel, Ll1depth_per_pixel, sizes_list, has_depth):
        assert Ll1depth_per_pixel is not None, "Ll1depth_per_pixel must not be None"
        self.Ll1_per_pixel = Ll1_per_pixel
        self.ssim_loss_per_pixel = ssim_loss_per_pixel
        self.Ll1depth_per_pixel = Ll1depth_per_pixel
        self.Ll1_scalar = _sumsq(Ll1_per_pixel)
        # with disable_ssim both blocks are the SAME tensor (solver/batch_training_loss.py:15-17): reuse the norm
        self.ssim_loss_scalar = self.Ll1_scalar if ssim_loss_per_pixel is Ll1_per_pixel else _sumsq(ssim_loss_per_pixel)
        self.Ll1depth_scalar = _sumsq(Ll1depth_per_pixel)
        self.loss_scalar = self.Ll1_scalar + self.ssim_loss_scalar + self.Ll1depth_scalar
        self.sizes_list = sizes_list
        self.has_depth = has_depth

    def blocks(self):
        return tuple(getattr(self, n) for n in _BLOCKS)

    def _new(self, blocks):
        return BatchLossImageState(*blocks, self.sizes_list, self.has_depth)

    def check_invariant(self):
        for name, t in zip(_BLOCKS, self.blocks()):
            if name == "Ll1depth_per_pixel" and not self.has_depth:
                continue
            c = t.clone()
            for i, (H, W) in enumerate(self.sizes_list):
                c[i, :, :H, :W] = 0
            assert torch.all(c == 0), f"{name} has values outside valid image regions"

    @property
    def length(self):
        n = sum(2 * 3 * H * W for H, W in self.sizes_list)
        if self.has_depth:
            n += sum(H * W for H, W in self.sizes_list)
        return n

    def index_to_desc(self, index):
        assert index < self.length, "Index out of bounds for BatchLossImageState"
        for name, t in zip(_BLOCKS, self.blocks()):
            if index < t.numel():
                coords = []
                for dim in reversed(t.shape):
                    coords.append(index % dim)
                    index //= dim
                return name, tuple(reversed(coords))
            index -= t.numel()

    def unpack_dual(self):
        prim, tang = [], []
        for t in self.blocks():
            p, d = fwAD.unpack_dual(t)
            prim.append(p)
            tang.append(d if d is not None else torch.zeros_like(p))
        if self.ssim_loss_per_pixel is self.Ll1_per_pixel:   # keep the aliasing
            prim[1], tang[1] = prim[0], tang[0]
        return self._new(prim), self._new(tang)

    def backward(self, v, retain_graph=False):
        assert isinstance(v, BatchLossImageState), "v must be an instance of BatchLossImageState"
        if self.ssim_loss_per_pixel is self.Ll1_per_pixel:
            # aliased blocks: the reference back-propagates the same tensor twice (J^T of [I; I]); by linearity
            # that is ONE backward pass with the summed cotangent
            pairs = [(self.Ll1_per_pixel, v.Ll1_per_pixel + v.ssim_loss_per_pixel),
                     (self.Ll1depth_per_pixel, v.Ll1depth_per_pixel)]
        else:
            pairs = list(zip(self.blocks(), v.blocks()))
        for mine, theirs in pairs:
            if mine.requires_grad:
                mine.backward(theirs, retain_graph=retain_graph)

    def zero_like(self):
        return self._new([torch.zeros_like(t) for t in self.blocks()])

    def _binary(self, other, op):
        a = op(self.Ll1_per_pixel, other.Ll1_per_pixel)
        aliased = self.ssim_loss_per_pixel is self.Ll1_per_pixel and other.ssim_loss_per_pixel is other.Ll1_per_pixel
        b = a if aliased else op(self.ssim_loss_per_pixel, other.ssim_loss_per_pixel)
        return self._new((a, b, op(self.Ll1depth_per_pixel, other.Ll1depth_per_pixel)))

    def __add__(self, other):
        return self._binary(other, torch.add)

    def __sub__(self, other):
        return self._binary(other, torch.sub)

    def __mul__(self, other):
        if not isinstance(other, (int, float)):
            raise TypeError("Can only multiply by scalar values")
        a = self.Ll1_per_pixel * other
        b = a if self.ssim_loss_per_pixel is self.Ll1_per_pixel else self.ssim_loss_per_pixel * other
        return self._new((a, b, self.Ll1depth_per_pixel * other))

    def __rmul__(self, other):
        return self.__mul__(other)

    def dot_tensor(self, other, damp=1.0):
        if not isinstance(damp, (int, float)):
            raise TypeError("Damping factor must be a scalar")
        s = torch.sum(self.Ll1_per_pixel * other.Ll1_per_pixel)
        if self.ssim_loss_per_pixel is self.Ll1_per_pixel and other.ssim_loss_per_pixel is other.Ll1_per_pixel:
            s = 2.0 * s
        else:
            s = s + torch.sum(self.ssim_loss_per_pixel * other.ssim_loss_per_pixel)
        return (s + torch.sum(self.Ll1depth_per_pixel * other.Ll1depth_per_pixel)) * damp

    def dot(self, other, damp=1.0):
        return self.dot_tensor(other, damp).item()


class MultiBatchLossImageState:
    """Several batch losses (the mini-batches of LinearSolverFunctions) as one generalized vector."""

    def __init__(self, batch_losses):
        self.batch_losses = batch_losses
        self.Ll1_scalar = sum(b.Ll1_scalar for b in batch_losses)
        self.ssim_loss_scalar = sum(b.ssim_loss_scalar for b in batch_losses)
        self.Ll1depth_scalar = sum(b.Ll1depth_scalar for b in batch_losses)
        self.loss_scalar = self.Ll1_scalar + self.ssim_loss_scalar + self.Ll1depth_scalar

    def check_invariant(self):
        for b in self.batch_losses:
            b.check_invariant()

    @property
    def length(self):
        return sum(b.length for b in self.batch_losses)

    def unpack_dual(self):
        duals = [b.unpack_dual() for b in self.batch_losses]
        return MultiBatchLossImageState([d[0] for d in duals]), MultiBatchLossImageState([d[1] for d in duals])

    def backward(self, v, retain_graph=False):
        assert isinstance(v, MultiBatchLossImageState), "v must be an instance of MultiBatchLossImageState"
        for b, vb in zip(self.batch_losses, v.batch_losses):
            b.backward(vb, retain_graph=retain_graph)

    def zero_like(self):
        return MultiBatchLossImageState([b.zero_like() for b in self.batch_losses])

    def __add__(self, other):
        return MultiBatchLossImageState([a + b for a, b in zip(self.batch_losses, other.batch_losses)])

    def __sub__(self, other):
        return MultiBatchLossImageState([a - b for a, b in zip(self.batch_losses, other.batch_losses)])

    def __mul__(self, other):
        return MultiBatchLossImageState([b * other for b in self.batch_losses])

    def __rmul__(self, other):
        return self.__mul__(other)

    def dot(self, other, damp=1.0):
        return sum(a.dot(b, damp) for a, b in zip(self.batch_losses, other.batch_losses))
