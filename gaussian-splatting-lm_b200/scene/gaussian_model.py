"""GaussianModel: raw parameters + activations + make_dual / zero_grad / update_step
(scene/gaussian_model.py:42-50,71-139,192-224 of the reference).  Everything else the reference class does
(point-cloud init, densify/prune, Adam) is out of scope of the LM hot path.  save_ply / load_ply read and write the
reference's checkpoint format (scene/gaussian_model.py:315-404: one binary little-endian PLY `vertex` element with
float properties x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_*, SH stored channel-major) with numpy only --
the reference goes through the `plyfile` package -- so scenes trained by the reference load here and vice versa."""
import os
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
        # bumped by every method that changes parameter VALUES (update_step, load_ply): `.data +=` leaves the
        # tensors' own version counters untouched, and LinearSolverFunctions' cached linearisation must notice
        self._param_version = 0
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
        m._param_version += 1
        return m

    # ---- checkpoint format (reference :315-404) ----
    def construct_list_of_attributes(self):
        names = ["x", "y", "z", "nx", "ny", "nz"]
        names += [f"f_dc_{i}" for i in range(self._features_dc.shape[1] * self._features_dc.shape[2])]
        names += [f"f_rest_{i}" for i in range(self._features_rest.shape[1] * self._features_rest.shape[2])]
        names.append("opacity")
        names += [f"scale_{i}" for i in range(self._scaling.shape[1])]
        names += [f"rot_{i}" for i in range(self._rotation.shape[1])]
        return names

    def save_ply(self, path):
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        cpu = lambda t: t.detach().cpu().numpy().astype(np.float32)
        xyz = cpu(self._xyz)
        cols = np.concatenate((xyz, np.zeros_like(xyz),
                               cpu(self._features_dc.detach().transpose(1, 2).flatten(start_dim=1)),
                               cpu(self._features_rest.detach().transpose(1, 2).flatten(start_dim=1)),
                               cpu(self._opacity), cpu(self._scaling), cpu(self._rotation)), axis=1)
        names = self.construct_list_of_attributes()
        assert cols.shape[1] == len(names)
        header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % cols.shape[0]
        header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
        with open(path, "wb") as f:
            f.write(header.encode("ascii"))
            f.write(np.ascontiguousarray(cols, dtype="<f4").tobytes())

    @staticmethod
    def _read_ply_vertices(path):
        """{property name: float64 column} of the first element of a PLY file (binary little/big endian or ascii)."""
        types = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
                 "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
                 "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}
        with open(path, "rb") as f:
            if f.readline().strip() != b"ply":
                raise ValueError(f"{path}: not a PLY file")
            fmt, count, props, in_first, seen = None, 0, [], False, 0
            while True:
                line = f.readline()
                if not line:
                    raise ValueError(f"{path}: truncated PLY header")
                tok = line.decode("ascii").split()
                if not tok or tok[0] == "comment":
                    continue
                if tok[0] == "format":
                    fmt = tok[1]
                elif tok[0] == "element":
                    seen += 1
                    in_first = seen == 1
                    if in_first:
                        count = int(tok[2])
                elif tok[0] == "property" and in_first:
                    if tok[1] == "list":
                        raise ValueError(f"{path}: list properties are not supported in the vertex element")
                    props.append((tok[2], types[tok[1]]))
                elif tok[0] == "end_header":
                    break
            if fmt == "ascii":
                data = np.loadtxt(f, max_rows=count, ndmin=2)
                return {n: data[:, i].astype(np.float64) for i, (n, _) in enumerate(props)}
            order = "<" if fmt == "binary_little_endian" else ">"
            rec = np.dtype([(n, order + t) for n, t in props])
            data = np.frombuffer(f.read(rec.itemsize * count), dtype=rec, count=count)
            return {n: data[n].astype(np.float64) for n, _ in props}

    def load_ply(self, path, use_train_test_exp=False, device="cuda", requires_grad=True):
        v = self._read_ply_vertices(path)
        by_index = lambda prefix: sorted((k for k in v if k.startswith(prefix)), key=lambda k: int(k.split("_")[-1]))
        xyz = np.stack((v["x"], v["y"], v["z"]), axis=1)
        P = xyz.shape[0]
        f_dc = np.stack((v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]), axis=1).reshape(P, 3, 1)
        rest_names = by_index("f_rest_")
        assert len(rest_names) == 3 * (self.max_sh_degree + 1) ** 2 - 3
        f_rest = np.stack([v[k] for k in rest_names], axis=1).reshape(P, 3, (self.max_sh_degree + 1) ** 2 - 1) \
            if rest_names else np.zeros((P, 3, 0))
        scales = np.stack([v[k] for k in by_index("scale_")], axis=1)
        rots = np.stack([v[k] for k in by_index("rot")], axis=1)
        T = lambda a: torch.tensor(a, dtype=torch.float, device=device)
        raw = {"_xyz": T(xyz), "_features_dc": T(f_dc).transpose(1, 2).contiguous(),
               "_features_rest": T(f_rest).transpose(1, 2).contiguous(), "_opacity": T(v["opacity"][:, None]),
               "_scaling": T(scales), "_rotation": T(rots)}
        for n, t in raw.items():
            setattr(self, n, nn.Parameter(t.requires_grad_(True)) if requires_grad else t)
        if self._exposure.numel() == 0:
            self._exposure = torch.eye(3, 4, device=device)[None].clone()
        self.active_sh_degree = self.max_sh_degree
        self._param_version += 1

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
        self._param_version += 1

    def param_signature(self):
        """Changes whenever a parameter tensor is replaced, resized, modified in place through autograd-visible ops,
        or stepped by update_step / load_ply: what a cached linearisation is valid for."""
        return (self._param_version, int(self.active_sh_degree)) + tuple(
            (getattr(self, n).data_ptr(), getattr(self, n)._version, tuple(getattr(self, n).shape)) for n in _PARAMS[:6])

    # ---- densification bookkeeping (reference scene/gaussian_model.py:561-563, train_jvp.py:311-316) ----
    def init_densification_stats(self):
        """max_radii2D [P], xyz_gradient_accum [P,1], denom [P,1] as training_setup creates them (reference :168-170)."""
        P, dev = self._xyz.shape[0], self._xyz.device
        self.max_radii2D = torch.zeros(P, device=dev)
        self.xyz_gradient_accum = torch.zeros((P, 1), device=dev)
        self.denom = torch.zeros((P, 1), device=dev)

    def add_densification_stats(self, viewspace_point_tensor, update_filter):
        """The reference's method: accumulates ||viewspace grad[:, :2]|| and the visit count on `update_filter`."""
        grad = viewspace_point_tensor.grad if getattr(viewspace_point_tensor, "grad", None) is not None \
            else viewspace_point_tensor
        self.xyz_gradient_accum[update_filter] += torch.norm(grad[update_filter, :2], dim=-1, keepdim=True)
        self.denom[update_filter] += 1

    def update_densification_stats(self, radii, viewspace_grad):
        """train_jvp.py:313-316 in ONE kernel on CUDA (gslm_densification_stats): max_radii2D, xyz_gradient_accum and
        denom updated on the visible set radii > 0.  radii: [P] (max over the batch's views); viewspace_grad: [P,3]
        (the batch's dL_dmeans2D, e.g. LinearSolverFunctions.viewspace_grad) or a tensor carrying it in .grad."""
        if not hasattr(self, "max_radii2D") or self.max_radii2D.shape[0] != self._xyz.shape[0]:
            self.init_densification_stats()
        g = viewspace_grad.grad if getattr(viewspace_grad, "grad", None) is not None else viewspace_grad
        if radii.is_cuda:
            from diff_gaussian_rasterization import _C
            _C.densification_stats(radii, g.detach(), self.max_radii2D, self.xyz_gradient_accum.view(-1),
                                   self.denom.view(-1))
            return
        vis = radii > 0
        self.max_radii2D[vis] = torch.max(self.max_radii2D[vis], radii[vis].to(self.max_radii2D.dtype))
        self.add_densification_stats(g, vis)

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
