"""A tiny differentiable stand-in for batch_training_loss (CPU, torch only) used to test the host-side solver
plumbing without the rasterizer: each 'camera' renders a fixed random linear map of the ACTIVATED parameters
followed by tanh, and the residual is image - gt, aliased into both blocks like disable_ssim=True."""
from types import SimpleNamespace

import numpy as np
import torch

import scene_synth as ss
from scene.gaussian_model import GaussianModel
from solver.loss_image_state import BatchLossImageState


def make_toy(P=24, n_cams=4, H=6, W=5, seed=0):
    g = ss.make_gaussians(P, sh_degree=1, seed=seed)
    m = GaussianModel.from_activated(g, device="cpu", num_cameras=n_cams)
    gen = torch.Generator().manual_seed(seed + 1)
    D = P * (3 + 12 + 3 + 4 + 1)
    cams = []
    for i in range(n_cams):
        cams.append(SimpleNamespace(image_height=H, image_width=W, uid=i,
                                    A=torch.randn(3 * H * W, D, generator=gen) / np.sqrt(D),
                                    original_image=torch.rand(3, H, W, generator=gen), alpha_mask=None))

    def loss_func(gaussians, viewpoint_cams, batch_stats=None):
        feat = torch.cat([gaussians.get_xyz.reshape(-1), gaussians.get_features.reshape(-1),
                          gaussians.get_scaling.reshape(-1), gaussians.get_rotation.reshape(-1),
                          gaussians.get_opacity.reshape(-1)])
        imgs = torch.stack([torch.tanh(c.A @ feat).view(3, H, W) for c in viewpoint_cams])
        gt = torch.stack([c.original_image for c in viewpoint_cams])
        r = imgs - gt
        return BatchLossImageState(r, r, torch.zeros(0, requires_grad=True), [(H, W)] * len(viewpoint_cams), False)

    return m, cams, loss_func
