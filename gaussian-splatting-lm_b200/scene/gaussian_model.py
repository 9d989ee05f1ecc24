"""GaussianModel: raw parameters + activations + make_dual / zero_grad / update_step
(scene/gaussian_model.py:42-50,71-139,192-224 of the reference).  Everything else the reference class does
(point-cloud init, densify/prune, PLY IO, Adam) is out of scope of the LM hot path."""
from contextlib import contextmanager

import numpy as np
import torch
import torch.autograd.forward_ad as fwAD
from torch import nn

_PARAMS = ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity", "_exposure")
_STATE_FIELDS = ("xyz_grad", "features_dc_grad", "features_rest_grad", "scaling_grad", "rotation_grad",
                 "opacity_grad", "exposure_grad")


def inverse_sigmoid(x):
    return torch.log(x / (1 - x))


class GaussianModel:
    def setup_functions(self):
        self.scaling_activation = torch.exp
        self.scaling_inverse_activation = torch.log
        self.opacity_activation = torch.sigmoid
        self.inverse_opacity_activation = inverse_sigmoid
        self.rotation_activation = torch.nn.functional.normalize

    def __init__(self, sh_degree, optimizer_type="default"):
        self.active_sh_degree = 0
        self.optimizer_type = optimizer_type
        self.max_sh_degree = sh_degree
        for n in _PARAMS:
            setattr(self, n, torch.empty(0))
        self.pretrained_exposures = None
        self.exposure_mapping = {}
        self.setup_functions()

    @classmethod
    def from_activated(cls, g, device="cuda", num_cameras=1, requires_grad=True):
        """Build a model whose ACTIVATED values equal the dict of scene_synth.make_gaussians
        (means3D, scales, rotations, opacities, shs, sh_degree)."""
        def T(a):
            return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32, device=device)
        m = cls(int(g["sh_degree"]))
        m.active_sh_degree = int(g["sh_degree"])
        shs = T(g["shs"])
        raw = {"_xyz": T(g["means3D"]), "_features_dc": shs[:, :1].contiguous(),
               "_features_rest": shs[:, 1:].contiguous(), "_scaling": torch.log(T(g["scales"])),
               "_rotation": T(g["rotations"]), "_opacity": inverse_sigmoid(T(g["opacities"]).reshape(-1, 1)),
               "_exposure": torch.eye(3, 4, device=device)[None].repeat(num_cameras, 1, 1)}
        for n, v in raw.items():
            setattr(m, n, nn.Parameter(v.requires_grad_(requires_grad)) if requires_grad else v)
        return m

    # ---- forward-AD plumbing (reference :71-139) ----
    @contextmanager
    def make_dual(self, v):
        orig = {n: getattr(self, n) for n in _PARAMS}
        try:
            for n, f in zip(_PARAMS, _STATE_FIELDS):
                setattr(self, n, fwAD.make_dual(orig[n], getattr(v, f)))
            yield
        finally:
            for n in _PARAMS:
                setattr(self, n, orig[n])

    def zero_grad(self):
        for n in _PARAMS:
            getattr(self, n).grad = None

    def update_step(self, s):
        for n, f in zip(_PARAMS, _STATE_FIELDS):
            getattr(self, n).data += getattr(s, f)

    # ---- activations (reference :192-224) ----
    @property
    def get_scaling(self):
        return self.scaling_activation(self._scaling)

    @property
    def get_rotation(self):
        return self.rotation_activation(self._rotation)

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_features_dc(self):
        return self._features_dc

    @property
    def get_features_rest(self):
        return self._features_rest

    @property
    def get_opacity(self):
        return self.opacity_activation(self._opacity)

    @property
    def get_exposure(self):
        return self._exposure

    def get_exposure_from_name(self, image_name):
        if self.pretrained_exposures is None:
            return self._exposure[self.exposure_mapping[image_name]]
        return self.pretrained_exposures[image_name]

    def get_covariance(self, scaling_modifier=1):
        """strip_symmetric(L L^T), L = R(q) diag(s * modifier) (reference :42-46, utils/general_utils.py:78-128)."""
        s = self.get_scaling * scaling_modifier
        q = self.get_rotation        # already normalised; build_rotation normalises again (idempotent)
        q = q / q.norm(dim=1, keepdim=True)
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                         2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                         2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
        L = R * s[:, None, :]
        S = L @ L.transpose(1, 2)
        return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], dim=1)
