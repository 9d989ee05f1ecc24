"""Synthetic Gaussian clouds and camera rigs -- the benchmark / test workload generator.

Everything is generated on the CPU from a seeded numpy Generator so that the same scene is
reproduced on every rank and every GPU count.  Camera matrices follow the reference's
conventions exactly:
  * world->view and projection as utils/graphics_utils.py:38-71 (getWorld2View2 / getProjectionMatrix),
  * stored TRANSPOSED (row-vector form) with full_proj = wv @ proj and camera_center = inverse(wv)[3,:3]
    as scene/cameras.py:86-89.
"""
import math
from types import SimpleNamespace

import numpy as np


def make_gaussians(P, sh_degree=3, seed=0, extent=1.0, scale_mult=1.0, shell_fraction=0.0):
    """Random Gaussian cloud in *activated* form (what the rasterizer consumes):
    means3D [P,3], scales [P,3] (>0), rotations [P,4] (unit), opacities [P,1] in (0,1), shs [P,M,3].

    means ~ U(-extent, extent)^3; scale = exp(N(log s0, 0.5^2)) with s0 = 0.5 * 2 / P^(1/3) (keeps the
    mean tiles/Gaussian O(10) at every P); rotation = normalize(N(0,1)^4); opacity = sigmoid(N(0,1.5^2));
    SH: DC ~ N(0,1), rest ~ N(0,0.1^2).  shell_fraction > 0 puts that share of the points on a far
    shell r in [8,25] with 6x scales ("360-scale" config D)."""
    rng = np.random.default_rng(seed)
    M = (sh_degree + 1) ** 2
    means = rng.uniform(-extent, extent, size=(P, 3))
    s0 = 0.5 * 2.0 / (P ** (1.0 / 3.0)) * scale_mult
    scales = np.exp(rng.normal(math.log(s0), 0.5, size=(P, 3)))
    n_shell = int(P * shell_fraction)
    if n_shell:
        d = rng.normal(size=(n_shell, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        r = rng.uniform(8.0, 25.0, size=(n_shell, 1))
        means[:n_shell] = d * r
        scales[:n_shell] *= 6.0
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = 1.0 / (1.0 + np.exp(-rng.normal(0.0, 1.5, size=(P, 1))))
    shs = np.empty((P, M, 3))
    shs[:, 0] = rng.normal(0.0, 1.0, size=(P, 3))
    if M > 1:
        shs[:, 1:] = rng.normal(0.0, 0.1, size=(P, M - 1, 3))
    f = np.float32
    return {"means3D": means.astype(f), "scales": scales.astype(f), "rotations": q.astype(f),
            "opacities": opac.astype(f), "shs": shs.astype(f), "sh_degree": int(sh_degree)}


def world2view(R, t):
    """utils/graphics_utils.py:38-51 with translate = 0, scale = 1."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.T
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    return np.float32(Rt)


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:53-71."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    Pm = np.zeros((4, 4), np.float32)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def look_at_camera(eye, target, W, H, fovx, up=(0.0, 0.0, 1.0), znear=0.01, zfar=100.0):
    """A camera at `eye` looking at `target` (COLMAP convention: +z forward, +y down), as a namespace with the
    fields of scene/cameras.py: image_width/height, FoVx/FoVy, world_view_transform, full_proj_transform,
    camera_center (numpy float32; matrices transposed like the reference's)."""
    eye = np.asarray(eye, np.float64)
    fwd = np.asarray(target, np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    upv = np.asarray(up, np.float64)
    right = np.cross(fwd, upv)
    if np.linalg.norm(right) < 1e-6:
        right = np.cross(fwd, np.array([0.0, 1.0, 0.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w_R = np.stack([right, down, fwd], axis=1)      # columns = camera axes in world coordinates
    w2c_R = c2w_R.T
    t = -w2c_R @ eye
    # the reference stores R = c2w rotation and T = w2c translation (scene/dataset_readers.py convention)
    wv = world2view(c2w_R, t).T                        # transposed, row-vector form
    fovy = 2.0 * math.atan(math.tan(fovx / 2) * H / W)
    proj = projection_matrix(znear, zfar, fovx, fovy).T
    full = (wv.astype(np.float32) @ proj.astype(np.float32)).astype(np.float32)
    center = np.linalg.inv(wv.astype(np.float64))[3, :3].astype(np.float32)
    return SimpleNamespace(image_width=int(W), image_height=int(H), FoVx=float(fovx), FoVy=float(fovy),
                           znear=znear, zfar=zfar, world_view_transform=np.ascontiguousarray(wv, np.float32),
                           projection_matrix=np.ascontiguousarray(proj, np.float32),
                           full_proj_transform=np.ascontiguousarray(full), camera_center=center)


def hemisphere_rig(V, W, H, radius=4.0, fovx=0.6911):
    """Blender-like rig (configs A, B, C): V cameras on a Fibonacci upper hemisphere looking at the origin."""
    cams = []
    golden = math.pi * (3.0 - math.sqrt(5.0))
    for i in range(V):
        z = (i + 0.5) / V * 0.9 + 0.05            # height in (0.05, 0.95): upper hemisphere, away from the pole
        r = math.sqrt(max(0.0, 1.0 - z * z))
        th = golden * i
        eye = radius * np.array([r * math.cos(th), r * math.sin(th), z])
        cams.append(look_at_camera(eye, (0.0, 0.0, 0.0), W, H, fovx))
    return cams


def ring_rig(V, W, H, radius=4.0, fovx=0.6911):
    """360-like rig (config D): ring of radius 4, elevation 10-30 degrees, looking at the origin."""
    cams = []
    for i in range(V):
        th = 2.0 * math.pi * i / V
        el = math.radians(10.0 + 20.0 * ((i * 7) % V) / max(V - 1, 1))
        eye = radius * np.array([math.cos(el) * math.cos(th), math.cos(el) * math.sin(th), math.sin(el)])
        cams.append(look_at_camera(eye, (0.0, 0.0, 0.0), W, H, fovx))
    return cams


CONFIGS = {
    # name: (P, V, W, H, rig, shell_fraction, extent)   SURVEY.md 8: sizes A-D
    "A": (10_000, 4, 128, 128, "hemisphere", 0.0, 1.0),
    "B": (100_000, 16, 800, 800, "hemisphere", 0.0, 1.0),
    "C": (1_000_000, 64, 1920, 1080, "hemisphere", 0.0, 1.0),
    "D": (3_000_000, 256, 1600, 1060, "ring", 1.0 / 3.0, 3.0),
}


def make_config(name, sh_degree=3, views=None, P=None):
    Pn, V, W, H, rig, shell, extent = CONFIGS[name]
    Pn = P or Pn
    V = views or V
    idx = "ABCD".index(name)
    g = make_gaussians(Pn, sh_degree=sh_degree, seed=1000 + idx, extent=extent, shell_fraction=shell)
    cams = hemisphere_rig(V, W, H) if rig == "hemisphere" else ring_rig(V, W, H)
    return g, cams


def random_tangent(g, seed, mask_xyz=False):
    """v ~ N(0,1) per parameter on the activated inputs (same keys as make_gaussians)."""
    rng = np.random.default_rng(seed)
    t = {k: rng.normal(size=g[k].shape).astype(np.float32) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    if mask_xyz:
        t["means3D"][:] = 0.0
    return t


def camera_dict(cam, bg=(0.0, 0.0, 0.0), antialiasing=False):
    """Oracle-style camera dict from a camera namespace."""
    return {"W": cam.image_width, "H": cam.image_height, "tan_fovx": math.tan(cam.FoVx * 0.5),
            "tan_fovy": math.tan(cam.FoVy * 0.5), "viewmatrix": cam.world_view_transform.reshape(16),
            "projmatrix": cam.full_proj_transform.reshape(16), "campos": cam.camera_center,
            "bg": np.asarray(bg, np.float32), "antialiasing": bool(antialiasing)}


def torch_camera(cam, device="cuda", original_image=None, alpha_mask=None, uid=0):
    """Camera namespace with torch tensors on `device` and the attributes the solver reads from the reference's
    scene.cameras.Camera: image sizes, FoV, the three matrices/centre, original_image [3,H,W], alpha_mask, depth flags."""
    import torch
    H, W = cam.image_height, cam.image_width
    if original_image is None:
        original_image = torch.zeros(3, H, W, device=device)      # gt = 0
    return SimpleNamespace(
        image_width=W, image_height=H, FoVx=cam.FoVx, FoVy=cam.FoVy, znear=cam.znear, zfar=cam.zfar, uid=uid,
        image_name=f"synth_{uid:04d}",
        world_view_transform=torch.from_numpy(cam.world_view_transform).to(device),
        projection_matrix=torch.from_numpy(cam.projection_matrix).to(device),
        full_proj_transform=torch.from_numpy(cam.full_proj_transform).to(device),
        camera_center=torch.from_numpy(cam.camera_center).to(device),
        original_image=original_image, alpha_mask=alpha_mask, depth_reliable=False, invdepthmap=None)
