"""LinearSolverFunctions: the callables CG / CGLS consume (reference solver/solver_functions.py:15-138) --
matvec = J v, matvec_T = J^T u, dot, saxpy, evaluate_loss, get_initial_solution -- plus the additive
normal_matvec = J^T J v (+ D v) that plugs into conjugate_gradient(matvec=...).

Two execution modes with identical results:

generic  (any loss_func): exactly the reference's composition -- forward-mode AD through loss_func for J v,
         loss.backward for J^T u; the rasterizer kernels run underneath via autograd.Function.{jvp,backward}.
cached   (loss_func is a functools.partial of solver.batch_training_loss.batch_training_loss; disable_ssim=True is
         the configuration the LM driver uses, train_jvp.py:212): the linearisation point is
         fixed for the whole CG solve, so evaluate_loss() keeps, per view, the sorted instance records, tile
         ranges, final_T / n_contrib and the pixel weights; every later matvec / matvec_T / normal_matvec then
         costs only the tangent-preprocess, one tile kernel and the per-Gaussian adjoint kernel per view: no
         preprocess, no scan / sort, no host synchronisation.
         normal_matvec uses the fused tangent->adjoint tile kernel, so J v is never written to HBM.
         With disable_ssim=False (L1 + SSIM blocks) the residual Jacobian is no longer diagonal in image space
         (11x11 windows), so J^T J v = J_r^T (L^T L) J_r v runs as tangent tile kernel -> loss-block JVP -> loss-block
         VJP (csrc/loss_blocks.cu, one kernel each) -> adjoint tile kernel, still on the cached linearisation.

Multi-GPU (torch.distributed initialised, one process per GPU): views are sharded over ranks, parameters and
parameter-space vectors are replicated, and each normal_matvec / matvec_T ends with ONE allreduce(sum) of the
packed 59 P-float result; residual-space dots allreduce one scalar.
"""
import functools
import math

import torch
import torch.autograd.forward_ad as fwAD

from diff_gaussian_rasterization import _C
from solver.batch_training_loss import batch_training_loss, gather_targets
from solver.gaussian_model_state import (GaussianModelDampMatrix, GaussianModelParamGroupMask,
                                         GaussianModelSplatMask, GaussianModelState)
from solver.loss_image_state import BatchLossImageState, MultiBatchLossImageState


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


_PEER_CACHE = {}        # device index -> _PeerAccumulator or None (decided once per device, on all ranks alike)


class _PeerAccumulator:
    """The rank's activated-space accumulator (59 P floats) as a SYMMETRIC allocation with an NVSwitch multicast
    mapping (torch.distributed._symmetric_memory supplies the allocation, the handle exchange and the cross-rank
    barrier; the reduction itself is this library's multimem kernel, csrc/peer_reduce.cu).
    ONE buffer per device: `view(n)` hands out its first n floats and re-allocates (collectively: every rank sees
    the same n, P being replicated) only when n outgrows the capacity or falls below a quarter of it, so densification /
    pruning / a new scene do not pile up rendezvoused buffers."""

    HEADROOM = 1.25         # capacity over the requested size: the Gaussian count drifts between LM steps

    def __init__(self, n, dev, dist):
        self.dev, self.dist = dev, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.t = self.hdl = None
        self.cap = self.mc = 0
        self._allocate(n)

    def _allocate(self, n):
        import torch.distributed._symmetric_memory as sm
        self.t = self.hdl = None                    # drop the old buffer before asking for the new one
        cap = (int(n * self.HEADROOM) + 3) // 4 * 4
        self.t = sm.empty(cap, dtype=torch.float32, device=self.dev)
        self.t.zero_()
        self.hdl = sm.rendezvous(self.t, self.dist.group.WORLD.group_name)
        self.mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        if not self.mc:
            raise RuntimeError("no multicast mapping (NVLS) for the symmetric accumulator")
        self.cap = cap

    def view(self, n):
        """The first n floats of the buffer.  Views handed out earlier stay valid (several LinearSolverFunctions of
        different sizes may share the buffer) until a re-allocation; allreduce() refuses a view that outlived one."""
        if n > self.cap or (n + 3) // 4 * 4 < self.cap // 4:
            self._allocate(n)
        return self.t[:n]

    def allreduce(self, flat):
        """Sum `flat` (a tensor from view()) over the ranks, in place, on the current stream: barrier (all partials
        written) -> multimem reduce + broadcast of this rank's slice -> barrier (all slices stored)."""
        if flat.data_ptr() != self.t.data_ptr() or flat.numel() > self.cap:
            raise RuntimeError("gslm_b200: accumulator view is stale (the symmetric buffer was re-allocated for "
                               "another size since); linearize again before the next product")
        n_pad = (flat.numel() + 3) // 4 * 4         # <= cap; the floats past the view are summed too and never read
        self.hdl.barrier(channel=0)
        _C.multimem_allreduce(self.mc, n_pad, self.rank, self.world)
        self.hdl.barrier(channel=0)


def _peer_accumulator(n, dev, dist):
    """The device's symmetric accumulator, or None when peer memory is disabled (GSLM_PEER_REDUCE=0) or unavailable on
    this box (no NVSwitch multicast): the product then ends in an NCCL allreduce instead."""
    import os
    # measured (profiles/r2l_peer_reduce.txt, 236 MB): 8 GPUs 0.57 ms against NCCL's 0.64; 2 GPUs 0.60 against 0.47 --
    # with two ranks the switch has nothing to save (each link carries half the payload either way), so the multimem
    # path is taken from 3 ranks up unless GSLM_PEER_REDUCE forces it on (1) or off (0)
    mode = os.environ.get("GSLM_PEER_REDUCE", "auto")
    if mode == "0" or dev.type != "cuda" or (mode == "auto" and dist.get_world_size() < 3):
        return None
    key = dev.index
    if key not in _PEER_CACHE:
        acc, why = None, ""
        try:
            acc = _PeerAccumulator(n, dev, dist)
        except Exception as e:      # noqa: BLE001
            why = f"{type(e).__name__}: {e}"
        # the ranks must agree: one rank on NCCL while the others wait in the symmetric-memory barrier would hang
        ok = torch.tensor([1 if acc is not None else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            import warnings
            warnings.warn("gslm_b200: peer-memory reduction unavailable on at least one rank"
                          + (f" (here: {why})" if why else "") + "; using NCCL")
            acc = None
        _PEER_CACHE[key] = acc
    return _PEER_CACHE[key]


class _ViewCache:
    """Linearisation state of one view (everything a CG iteration needs, all device-resident)."""
    __slots__ = ("cam", "H", "W", "R", "radii", "geom", "binning", "img", "weight", "mask", "residual", "cam_args")


def balanced_view_assignment(costs, world):
    """Greedy (longest-processing-time-first) assignment of views to ranks on a per-view cost (the
    instance count R), with at most ceil(V / world) views per rank so that memory and launch counts stay even.
    Deterministic: ties go to the lower view index / lower rank.  Returns world ascending index lists."""
    V = len(costs)
    cap = (V + world - 1) // world
    load, owned = [0] * world, [[] for _ in range(world)]
    for i in sorted(range(V), key=lambda i: (-int(costs[i]), i)):
        r = min((r for r in range(world) if len(owned[r]) < cap), key=lambda r: (load[r], r))
        owned[r].append(i)
        load[r] += int(costs[i])
    return [sorted(o) for o in owned]


class LinearSolverFunctions:

    def __init__(self, loss_func, gaussians, viewpoint_cams, batch_size=-1, param_mask=None, splat_mask=None,
                 cached=None, shard_views=True, tile_culling=True, balance_views=True):
        self.loss_func = loss_func
        # exact tile culling in the cached mode's own forward passes (include/gslm_b200.h, GslmCamera::tile_culling):
        # residuals, linearisation cache and every product are bit-identical with and without it
        self.tile_culling = bool(tile_culling)
        self.gaussians = gaussians
        self.all_cams = list(viewpoint_cams)
        d = _dist() if shard_views else None
        self.dist = d
        # data-parallel axis: views.  Default: rank r owns views r, r + world, ... (round robin); in cached mode the
        # assignment is re-balanced on the per-view instance counts below (_balance_views)
        self.view_ids = list(range(d.get_rank(), len(self.all_cams), d.get_world_size())) if d else \
            list(range(len(self.all_cams)))
        self.viewpoint_cams = [self.all_cams[i] for i in self.view_ids]
        self.B = len(self.viewpoint_cams)
        self.batch_size = max(self.B, 1) if batch_size < 0 else batch_size
        self.batch_stats = {}
        self.param_mask = param_mask
        self.splat_mask = splat_mask
        legal = (isinstance(loss_func, functools.partial) and loss_func.func is batch_training_loss
                 and not loss_func.keywords.get("train_test_exp", False)
                 and not getattr(loss_func.keywords.get("pipe"), "compute_cov3D_python", False)
                 and not getattr(loss_func.keywords.get("pipe"), "convert_SHs_python", False))
        if cached and not legal:
            raise ValueError("cached=True needs loss_func = partial(batch_training_loss, ...) without exposure "
                             "training and without the Python SH / covariance paths")
        self.cached = legal if cached is None else bool(cached)
        self.ssim = legal and not bool(loss_func.keywords.get("disable_ssim", False))
        self._views = None
        self._scratch = None
        self._plan = None
        if d and self.cached and balance_views and len(self.all_cams) > d.get_world_size():
            self._balance_views(batch_size)

    def _balance_views(self, batch_size):
        """Views -> ranks by greedy balancing on the per-view instance count R.  Every rank counts its
        round-robin share (gslm_count_instances: preprocess + culling + scan, no rendering), ONE all_gather makes the
        full list known everywhere, and the deterministic LPT assignment is computed identically on every rank."""
        d = self.dist
        world, rank = d.get_world_size(), d.get_rank()
        V = len(self.all_cams)
        pipe, bg = self._kw("pipe"), self._kw("bg")
        a = self._activated(with_layer=False)
        e = torch.empty(0, device=a["means3D"].device)
        views = [(c.world_view_transform, c.full_proj_transform, math.tan(c.FoVx * 0.5), math.tan(c.FoVy * 0.5),
                  int(c.image_height), int(c.image_width), c.camera_center) for c in self.viewpoint_cams]
        mine = _C.count_instances(bg, a["means3D"], e, a["opacities"], a["scales"], a["rotations"], 1.0, e, views,
                                  a["shs"], a["sh_degree"], pipe.antialiasing, tile_culling=self.tile_culling)
        per = (V + world - 1) // world
        t = torch.zeros(per, dtype=torch.int64, device=a["means3D"].device)
        t[:len(mine)] = torch.tensor(mine, dtype=torch.int64)
        allc = torch.empty(per * world, dtype=torch.int64, device=t.device)
        d.all_gather_into_tensor(allc, t)
        allc = allc.view(world, per).cpu()
        costs = [0] * V
        for r in range(world):
            for k, i in enumerate(range(r, V, world)):
                costs[i] = int(allc[r, k])
        self.view_costs = costs
        self.view_ids = balanced_view_assignment(costs, world)[rank]
        self.viewpoint_cams = [self.all_cams[i] for i in self.view_ids]
        self.B = len(self.viewpoint_cams)
        self.batch_size = max(self.B, 1) if batch_size < 0 else batch_size

    # ------------------------------------------------------------------ reference API
    def get_initial_solution(self):
        return GaussianModelState.from_gaussians(self.gaussians, param_mask=self.param_mask,
                                                 splat_mask=self.splat_mask)

    def _batches(self):
        for start in range(0, self.B, self.batch_size):
            yield start, min(start + self.batch_size, self.B)

    def evaluate_loss(self, with_batch_stats=False, with_grad=False):
        """r(theta) at the current parameters; in cached mode this is also the linearisation pass."""
        if self.cached and not with_grad and not with_batch_stats:
            self.linearize()
            return self.loss
        batch_stats = {} if with_batch_stats else None
        with torch.enable_grad() if with_grad else torch.no_grad():
            losses = []
            for start, end in self._batches():
                loss = self.loss_func(gaussians=self.gaussians, viewpoint_cams=self.viewpoint_cams[start:end],
                                      batch_stats=batch_stats)
                losses.append(loss)
                if with_batch_stats:
                    for k, v in batch_stats.items():
                        if start == 0:
                            self.batch_stats[k] = [v]
                        else:
                            self.batch_stats[k].append(v)
            self.loss = MultiBatchLossImageState(losses)
        return self.loss

    def _scalar(self, t):
        if self.dist:
            # a rank that owns no view holds the Python int 0 here: it must still enter the collective
            t = torch.as_tensor(t, dtype=torch.float32, device=self.gaussians._xyz.device).clone()
            self.dist.all_reduce(t)
        return t

    @property
    def loss_scalar(self):
        return self._scalar(self.loss.loss_scalar)

    @property
    def Ll1_scalar(self):
        return self._scalar(self.loss.Ll1_scalar)

    @property
    def depth_loss_scalar(self):
        return self._scalar(self.loss.Ll1depth_scalar)

    @property
    def get_batch_stats(self):
        return self.batch_stats

    def matvec(self, v):
        assert isinstance(v, GaussianModelState), "v must be an instance of GaussianModelState"
        if self.cached:
            return self._matvec_cached(v)
        with torch.no_grad(), fwAD.dual_level(), self.gaussians.make_dual(v):
            tangents = []
            for start, end in self._batches():
                loss_dual = self.loss_func(gaussians=self.gaussians, viewpoint_cams=self.viewpoint_cams[start:end])
                tangents.append(loss_dual.unpack_dual()[1])
            return MultiBatchLossImageState(tangents)

    def matvec_T(self, vs):
        assert isinstance(vs, MultiBatchLossImageState), "vs must be an instance of MultiBatchLossImageState"
        if self.cached:
            return self._matvec_T_cached(vs)
        with torch.enable_grad():
            self.gaussians.zero_grad()
            for idx, (start, end) in enumerate(self._batches()):
                loss_batch = self.loss_func(gaussians=self.gaussians, viewpoint_cams=self.viewpoint_cams[start:end])
                loss_batch.backward(vs.batch_losses[idx], retain_graph=True)
        g = self.gaussians
        for name in ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity"):
            grad = getattr(g, name).grad
            assert grad is None or not torch.isnan(grad).any(), f"NaN detected in gaussians.{name}.grad"
        out = GaussianModelState.from_gaussians_grad(g, param_mask=None, splat_mask=None)
        if self.dist:
            self.dist.all_reduce(out.packed())
        return GaussianModelState(*out.tensors(), param_mask=self.param_mask, splat_mask=self.splat_mask,
                                  _flat=out.flat)

    def dot(self, v1, v2, damp=1.0):
        s = v1.dot(v2, damp)
        if self.dist and isinstance(v1, (MultiBatchLossImageState, BatchLossImageState)):
            t = torch.tensor([s], dtype=torch.float64, device=self.gaussians._xyz.device)
            self.dist.all_reduce(t)
            s = t.item()
        return s

    def saxpy(self, a, x, y):
        return a * x + y

    # ------------------------------------------------------------------ host staging of parameter-space vectors
    # (additive; bench.py's end-to-end leg).  Parameter-space vectors are replicated over ranks, so with N ranks each
    # rank moves only 1/N of the bytes over its own PCIe link and NVLink does the rest.
    def _shard(self, n):
        d = self.dist
        world, rank = (d.get_world_size(), d.get_rank()) if d else (1, 0)
        per = (n + world - 1) // world
        return world, rank, per

    def state_from_host(self, host_flat, state):
        """state.packed() <- host_flat (pinned float32 [n], the same on every rank): each rank copies its 1/N
        slice host->device, then ONE all_gather over NVLink.  Returns the bytes this rank copied."""
        flat = state.packed()
        n = flat.numel()
        world, rank, per = self._shard(n)
        if world == 1:
            flat.copy_(host_flat, non_blocking=True)
            return 4 * n
        if getattr(self, "_stage", None) is None or self._stage.numel() != per * world:
            self._stage = torch.empty(per * world, dtype=torch.float32, device=flat.device)
        lo, hi = rank * per, min((rank + 1) * per, n)
        mine = self._stage[rank * per:(rank + 1) * per]
        if hi > lo:
            mine[:hi - lo].copy_(host_flat[lo:hi], non_blocking=True)
        self.dist.all_gather_into_tensor(self._stage, mine)
        flat.copy_(self._stage[:n])
        return 4 * max(hi - lo, 0)

    def state_to_host(self, state, host_flat):
        """host_flat[lo:hi] <- this rank's 1/N slice of the (replicated) state; returns the bytes copied."""
        flat = state.packed()
        n = flat.numel()
        world, rank, per = self._shard(n)
        lo, hi = rank * per, min((rank + 1) * per, n)
        if hi > lo:
            host_flat[lo:hi].copy_(flat[lo:hi], non_blocking=True)
        return 4 * max(hi - lo, 0)

    def host_pipeline(self, depth=2):
        """Double-buffered host staging for a stream of INDEPENDENT products (block of probe vectors / right-hand
        sides): see HostPipeline."""
        return HostPipeline(self, depth)

    # ------------------------------------------------------------------ cached mode
    def _kw(self, name, default=None):
        return self.loss_func.keywords.get(name, default)

    def _activated(self, with_layer=True):
        """Kernel inputs at the linearisation point + what the activation JVP / VJP need."""
        g = self.gaussians
        with torch.no_grad():
            raw_rot = g._rotation.detach()
            norm = raw_rot.norm(dim=1, keepdim=True).clamp_min(1e-12)   # F.normalize eps
            act = {"means3D": g._xyz.detach().contiguous(), "opacities": torch.sigmoid(g._opacity.detach()),
                   "scales": torch.exp(g._scaling.detach()), "rotations": raw_rot / norm, "rot_norm": norm,
                   "shs": torch.cat((g._features_dc.detach(), g._features_rest.detach()), dim=1).contiguous(),
                   "sh_degree": int(g.active_sh_degree)}
            if not with_layer:
                return act
            # the differentiated activation layer (csrc/param_space.cu) + its persistent tangent buffers
            act["layer"] = _C.Activations(act["scales"], act["rotations"], act["rot_norm"], act["opacities"],
                                          act["shs"].shape[1])
            act["tangent"] = tuple(torch.empty_like(act[k]) for k in ("shs", "scales", "rotations", "opacities"))
        return act

    def _check_depth_branch(self):
        """The depth residual block is not implemented by the reference (solver/batch_training_loss.py:88-93 raises);
        the cached path must not silently drop it."""
        w = self._kw("depth_l1_weight")
        it = self._kw("iteration", 0)
        if w is not None and w(it) > 0 and any(getattr(c, "depth_reliable", False) for c in self.viewpoint_cams):
            raise NotImplementedError("Ll1depth_pure_per_pixel is not implemented in this version.")

    def _param_signature(self):
        g = self.gaussians
        if hasattr(g, "param_signature"):
            return g.param_signature()
        return tuple((getattr(g, n).data_ptr(), getattr(g, n)._version) for n in
                     ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity"))

    def _gauss_args(self):
        a = self._act
        e = self._empty
        return (a["means3D"], e, a["opacities"], a["scales"], a["rotations"], 1.0, e, a["shs"], a["sh_degree"])

    LINEARIZE_CHUNK = 16   # views per gslm_forward_batch call

    def _targets(self, start, end, like):
        """gather_targets of the batch [start, end), cached: targets and alpha masks are constants of the object."""
        if not hasattr(self, "_target_cache"):
            self._target_cache = {}
        key = (start, end, tuple(like.shape))
        if key not in self._target_cache:
            self._target_cache[key] = gather_targets(self.viewpoint_cams[start:end], like)
        return self._target_cache[key]

    def linearize(self):
        """Forward-render every local view once; keep per-view caches, residuals and pixel weights.
        The views go through gslm_forward_batch in chunks of LINEARIZE_CHUNK (one host synchronisation per chunk plus
        one overlapped read-back per view) and the clamp mask / alpha mask / residual / pixel weight of every view
        come out of the library's fused epilogue: no per-view torch kernels."""
        self._check_depth_branch()
        pipe, bg = self._kw("pipe"), self._kw("bg")
        dev = self.gaussians._xyz.device
        self._empty = torch.empty(0, device=dev)
        self._act = a = self._activated()
        self._lin_signature = self._param_signature()
        self._views = []
        self._plan = None
        self._eval_cache = {}          # loss_value()'s stacked targets are re-gathered after every linearisation
        self._loss_batches = []
        batches, specs, view_args = [], [], []
        for start, end in self._batches():
            cams = self.viewpoint_cams[start:end]
            Hm, Wm = max(c.image_height for c in cams), max(c.image_width for c in cams)
            uniform = all((c.image_height, c.image_width) == (Hm, Wm) for c in cams)
            new = torch.empty if uniform else torch.zeros       # padding must read as zero
            resid = new((len(cams), 3, Hm, Wm), dtype=torch.float32, device=dev)
            gt, masks = self._targets(start, end, resid)
            # ssim mode: `resid` collects the clamped renders, cmask = d clamp / d colour (fused into the loss kernels)
            cmask = new(resid.shape, dtype=torch.float32, device=dev) if self.ssim else None
            batches.append((cams, resid, gt, masks, cmask))
            for i, cam in enumerate(cams):
                H, W = int(cam.image_height), int(cam.image_width)
                vc = _ViewCache()
                vc.cam, vc.H, vc.W = cam, H, W
                tfx, tfy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
                vc.cam_args = (bg, cam.world_view_transform, cam.full_proj_transform, tfx, tfy, H, W,
                               cam.camera_center, pipe.antialiasing, pipe.debug)
                view_args.append((cam.world_view_transform, cam.full_proj_transform, tfx, tfy, H, W, cam.camera_center))
                if self.ssim:
                    # the alpha mask is applied inside the loss kernels
                    vc.mask, vc.weight = None, None
                    specs.append(_C.ResidualSpec(image=resid[i, :, :H, :W], mask=cmask[i, :, :H, :W],
                                                 apply_alpha_mask=False, subtract_gt=False))
                else:
                    # d clamp(x,0,1)/dx as torch defines it (inclusive bounds) times the alpha mask; two aliased
                    # residual blocks => multiplicity 2 in J^T J (batch_training_loss.py:15-17)
                    vc.mask = torch.empty((3, H, W), dtype=torch.float32, device=dev)
                    vc.weight = torch.empty((3, H, W), dtype=torch.float32, device=dev)
                    specs.append(_C.ResidualSpec(image=resid[i, :, :H, :W], mask=vc.mask, weight=vc.weight,
                                                 gt=gt[i, :, :H, :W],
                                                 alpha_mask=masks[i, 0, :H, :W] if masks is not None else None,
                                                 weight_scale=2.0, apply_alpha_mask=True, subtract_gt=True))
                self._views.append(vc)
        for c0 in range(0, len(self._views), self.LINEARIZE_CHUNK):
            c1 = min(c0 + self.LINEARIZE_CHUNK, len(self._views))
            outs = _C.forward_batch(bg, a["means3D"], self._empty, a["opacities"], a["scales"], a["rotations"], 1.0,
                                    self._empty, view_args[c0:c1], a["shs"], a["sh_degree"], pipe.antialiasing,
                                    pipe.debug, residual=specs[c0:c1], tile_culling=self.tile_culling)
            for vc, (R, _color, radii, geom, binning, img, _) in zip(self._views[c0:c1], outs):
                vc.R, vc.radii, vc.geom, vc.binning, vc.img = R, radii, geom, binning, img
        losses = []
        for cams, resid, gt, masks, cmask in batches:
            depth_block = torch.zeros((0,), dtype=torch.float32, device=dev)
            sizes = [(c.image_height, c.image_width) for c in cams]
            if self.ssim:
                lam = self._kw("opt").lambda_dssim
                al = torch.tensor([math.sqrt((1.0 - lam) / (3 * h * w)) for h, w in sizes], device=dev)
                be = torch.tensor([math.sqrt(lam / (3 * h * w)) for h, w in sizes], device=dev)
                lb = _C.LossBatch(resid, masks, gt, al, be, tangent_masks=cmask)
                r1, r2 = lb.blocks(want_lin=True)
                self._loss_batches.append(lb)
                losses.append(BatchLossImageState(r1, r2, depth_block, sizes, False))
            else:
                losses.append(BatchLossImageState(resid, resid, depth_block, sizes, False))
        self.loss = MultiBatchLossImageState(losses)
        return self.loss

    def _render_chunks(self):
        """(start, (end of batch 1, end of batch 2, ...)) of the view chunks loss_value() renders with one library
        call: consecutive batches are merged while the chunk stays within RENDER_CHUNK views of ONE image size (the
        residual of a padded batch depends on the padding, so batches of mixed sizes stay on their own)."""
        chunk = None
        for b0, b1 in self._batches():
            sizes = {(int(c.image_height), int(c.image_width)) for c in self.viewpoint_cams[b0:b1]}
            if chunk is not None and len(sizes) == 1 and sizes == chunk[2] and b1 - chunk[0] <= self.RENDER_CHUNK:
                chunk[1].append(b1)
                continue
            if chunk is not None:
                yield chunk[0], tuple(chunk[1])
            chunk = [b0, [b1], sizes]
        if chunk is not None:
            yield chunk[0], tuple(chunk[1])

    RENDER_CHUNK = 16      # views per gslm_render_batch call in loss_value()

    def loss_value(self):
        """||r(theta)||^2 at the current parameters as a 0-dim device tensor, WITHOUT building a linearisation:
        render-only forward passes (gslm_render) and the residual reduced on the fly.  What the back-tracking line
        search needs (train_jvp.py:262-279 calls evaluate_loss + loss_scalar seven times per LM step).  The stacked
        targets / alpha masks of the views are gathered on the first call and reused until the next linearize():
        they are constants of a line search."""
        if not self.cached:
            self.evaluate_loss()
            return self.loss_scalar
        self._check_depth_branch()
        pipe, bg = self._kw("pipe"), self._kw("bg")
        dev = self.gaussians._xyz.device
        empty = torch.empty(0, device=dev)
        a = self._activated(with_layer=False)
        total = torch.zeros((), dtype=torch.float32, device=dev)
        if not hasattr(self, "_eval_cache"):
            self._eval_cache = {}       # targets and the render buffer of each chunk: constant over a line search
        for start, bounds in self._render_chunks():
            cams = self.viewpoint_cams[start:bounds[-1]]
            sizes = [(int(c.image_height), int(c.image_width)) for c in cams]
            Hm, Wm = max(h for h, _ in sizes), max(w for _, w in sizes)
            uniform = all(sz == (Hm, Wm) for sz in sizes)
            if start not in self._eval_cache:
                images = torch.zeros((len(cams), 3, Hm, Wm), dtype=torch.float32, device=dev)
                # targets are gathered batch by batch: a view without an alpha mask is zeroed when another view of
                # its BATCH has one (solver/batch_training_loss.py:61-67 of the reference), so the masks of a chunk
                # of several batches are assembled from the per-batch results
                parts = [gather_targets(self.viewpoint_cams[b0:b1], images[b0 - start:b1 - start])
                         for b0, b1 in zip((start,) + bounds[:-1], bounds)]
                gt = torch.cat([p[0] for p in parts])
                masks = None
                if any(p[1] is not None for p in parts):
                    masks = torch.cat([p[1] if p[1] is not None else torch.ones_like(p[0][:, :1]) for p in parts])
                self._eval_cache[start] = (images, gt, masks)
            images, gt, masks = self._eval_cache[start]
            views = [(c.world_view_transform, c.full_proj_transform, math.tan(c.FoVx * 0.5), math.tan(c.FoVy * 0.5),
                      h, w, c.camera_center) for c, (h, w) in zip(cams, sizes)]
            # one library call and one host synchronisation per chunk of views (gslm_render_batch)
            _, colors = _C.render_batch(bg, a["means3D"], empty, a["opacities"], a["scales"], a["rotations"], 1.0,
                                        empty, views, a["shs"], a["sh_degree"], pipe.antialiasing, pipe.debug,
                                        out=[images[i] for i in range(len(cams))] if uniform else None,
                                        tile_culling=self.tile_culling)
            if not uniform:
                for i, (h, w) in enumerate(sizes):
                    images[i, :, :h, :w].copy_(colors[i])
            images.clamp_(0, 1)
            if self.ssim:
                lam = self._kw("opt").lambda_dssim
                al = torch.tensor([math.sqrt((1.0 - lam) / (3 * h * w)) for h, w in sizes], device=dev)
                be = torch.tensor([math.sqrt(lam / (3 * h * w)) for h, w in sizes], device=dev)
                r1, r2 = _C.LossBatch(images, masks, gt, al, be).blocks(want_lin=False)
                total = total + torch.linalg.vector_norm(r1) ** 2 + torch.linalg.vector_norm(r2) ** 2
            else:
                r = (images * masks if masks is not None else images) - gt
                total = total + 2.0 * torch.linalg.vector_norm(r) ** 2     # two aliased blocks
        return self._scalar(total)

    def _tangent_inputs(self, v):
        """Activation JVP: raw-parameter tangent -> tangents of the kernel inputs (scene/gaussian_model.py:192-220),
        one kernel into buffers that live as long as the linearisation."""
        a = self._act
        t_shs, t_scales, t_rot, t_op = a["tangent"]
        a["layer"].jvp(v, t_shs, t_scales, t_rot, t_op)
        return (v.xyz_grad.contiguous(), None, t_op, t_scales, t_rot, None, t_shs)

    def _act_accumulators(self):
        """Packed activated-space gradient buffer [means3D | shs | scales | rotations | opacity] (59 P floats:
        the allreduce payload) and the 8-tuple of sinks the kernels write into."""
        a = self._act
        P, M = a["means3D"].shape[0], a["shs"].shape[1]
        # uninitialised: the batched adjoint calls run in overwrite mode.  Multi-GPU: the accumulator is the rank's
        # copy of a symmetric, multicast-mapped allocation, summed over ranks in place (csrc/peer_reduce.cu)
        n = P * (3 + 3 * M + 3 + 4 + 1)
        dev = a["means3D"].device
        self._peer = _peer_accumulator(n, dev, self.dist) if self.dist else None
        flat = self._peer.view(n) if self._peer is not None else torch.empty(n, dtype=torch.float32, device=dev)
        off = 0
        views = {}
        for name, shape in (("means3D", (P, 3)), ("shs", (P, M, 3)), ("scales", (P, 3)), ("rotations", (P, 4)),
                            ("opacities", (P, 1))):
            n = 1
            for d in shape:
                n *= d
            views[name] = flat[off:off + n].view(shape)
            off += n
        sinks = (None, None, views["opacities"], views["means3D"], None, views["shs"], views["scales"],
                 views["rotations"])
        self._flat_acc = flat
        return flat, views, sinks

    # Gaussian-range slices of the last adjoint launch when a collective follows (multi-GPU).  1 = the collective
    # starts when the whole adjoint stage is done.  Measured on 2 and 8 x B200 (profiles/r2g_*): cutting the launch
    # into 4 ranges so that NCCL reduces slice k while slice k + 1 is produced LOSES (adjoint stage 1.47 -> 2.34 ms
    # per step at 32 views/rank): the stage is latency-bound and a quarter-size launch fills 41 % of the SMs.
    N_SLICES = 1

    def _tail(self, views, v=None, damp=None):
        """The end of every parameter-space product: (allreduce over ranks) -> activation VJP (+ damping), run per
        Gaussian-range slice as the library's adjoint stage completes it (`_ProductTail`)."""
        return _ProductTail(self, views, v, damp)

    def _to_state(self, views, v=None, damp=None):
        """Activation VJP (+ damping) of a complete activated-space gradient (this rank's: no collective)."""
        t = _ProductTail(self, views, v, damp, local=True)
        t(0, 0, self._act["means3D"].shape[0])
        return t.finish()

    def _require_linearized(self):
        """The reference's matvec / matvec_T render at the CURRENT parameters on every call
        (solver/solver_functions.py:75-113); a cached linearisation is therefore only reused while the parameters
        are the ones it was taken at -- after update_step / load_ply / any in-place change it is rebuilt."""
        if self._views is None or self._lin_signature != self._param_signature():
            self.linearize()

    def _get_plan(self):
        if self._plan is None:
            self._plan = _C.JtjvBatchPlan([(vc.cam_args, vc.radii, vc.geom, vc.R, vc.binning, vc.img, vc.weight)
                                           for vc in self._views], self._act["means3D"].device)
        return self._plan

    def _image_tangents(self, gargs, targs):
        """J_r v for every local view in ONE library call (16 views per kernel launch): per batch of views a padded
        [n,3,Hmax,Wmax] tensor holding mask * (J v)."""
        dev = self._act["means3D"].device
        tens, dests, staged, k = [], [], [], 0
        for start, end in self._batches():
            cams = self.viewpoint_cams[start:end]
            Hm, Wm = max(c.image_height for c in cams), max(c.image_width for c in cams)
            uniform = all((c.image_height, c.image_width) == (Hm, Wm) for c in cams)
            t = (torch.empty if uniform else torch.zeros)((len(cams), 3, Hm, Wm), dtype=torch.float32, device=dev)
            for i in range(len(cams)):
                vc = self._views[k + i]
                if (vc.H, vc.W) == (Hm, Wm):
                    dests.append(t[i])                       # contiguous: the kernel writes in place
                else:
                    buf = torch.empty((3, vc.H, vc.W), dtype=torch.float32, device=dev)
                    dests.append(buf)
                    staged.append((t, i, buf, vc))
            tens.append(t)
            k += len(cams)
        _C.jvp_batch(gargs, targs, self._get_plan(), dests)
        for t, i, buf, vc in staged:
            t[i, :, :vc.H, :vc.W] = buf
        if self.ssim:                     # the clamp mask is applied inside the loss-block kernels (tangent_masks)
            return tens
        k = 0
        for t in tens:
            for i in range(t.shape[0]):
                vc = self._views[k + i]
                t[i, :, :vc.H, :vc.W].mul_(vc.mask)
            k += t.shape[0]
        return tens

    def _pull_back(self, cots, sinks, tail):
        """sinks = sum over local views of J_r^T (mask * cot): ONE library call; cots = per-batch padded tensors."""
        per_view, k = [], 0
        for cot in cots:
            for i in range(cot.shape[0]):
                vc = self._views[k + i]
                c = cot[i, :, :vc.H, :vc.W]
                per_view.append(c.contiguous() if self.ssim else c * vc.mask)     # ssim: already masked by the VJP kernel
            k += cot.shape[0]
        _C.backward_batch(self._gauss_args(), self._get_plan(), per_view, sinks, accumulate=False,
                          n_slices=tail.n_slices, on_slice=tail)

    def _matvec_cached(self, v):
        self._require_linearized()
        tens = self._image_tangents(self._gauss_args(), self._tangent_inputs(v))
        out = []
        for idx, ((start, end), t) in enumerate(zip(self._batches(), tens)):
            cams = self.viewpoint_cams[start:end]
            depth_block = torch.zeros((0,), dtype=torch.float32, device=t.device)
            t1, t2 = self._loss_batches[idx].jvp(t) if self.ssim else (t, t)
            out.append(BatchLossImageState(t1, t2, depth_block, [(c.image_height, c.image_width) for c in cams], False))
        return MultiBatchLossImageState(out)

    # densification bookkeeping (train_jvp.py:311-316): set track_viewspace = True and every cached matvec_T also
    # leaves the batch's viewspace gradient dL_dmeans2D [P,3] in self.viewspace_grad (summed over all views and
    # ranks); max_radii() gives the per-Gaussian maximum screen radius over all views.  Feed both to
    # GaussianModel.update_densification_stats.
    track_viewspace = False
    viewspace_grad = None

    def max_radii(self):
        self._require_linearized()
        P = self._act["means3D"].shape[0]
        r = torch.zeros(P, dtype=torch.int32, device=self._act["means3D"].device)
        for vc in self._views:
            torch.maximum(r, vc.radii, out=r)
        if self.dist:
            self.dist.all_reduce(r, op=self.dist.ReduceOp.MAX)
        return r

    def _matvec_T_cached(self, vs):
        self._require_linearized()
        flat, views, sinks = self._act_accumulators()
        if self.track_viewspace:
            self.viewspace_grad = torch.empty((self._act["means3D"].shape[0], 3), dtype=torch.float32, device=flat.device)
            sinks = (self.viewspace_grad,) + sinks[1:]
        cots = []
        for idx, _ in enumerate(self._batches()):
            u = vs.batch_losses[idx]
            if self.ssim:
                cots.append(self._loss_batches[idx].vjp(u.Ll1_per_pixel, u.ssim_loss_per_pixel))
            else:
                cots.append(u.Ll1_per_pixel + u.ssim_loss_per_pixel)     # J^T of the two aliased blocks
        tail = self._tail(views)
        self._pull_back(cots, sinks, tail)
        if self.dist and self.track_viewspace:
            self.dist.all_reduce(self.viewspace_grad)
        return tail.finish()

    def _normal_matvec_ssim(self, gargs, targs, sinks, quad, tail):
        """J_r^T (L^T L) J_r v: image tangents from the batched tangent tile kernel, the loss-block JVP and VJP
        kernels in image space, then the batched adjoint tile kernel."""
        cots = []
        for lb, t in zip(self._loss_batches, self._image_tangents(gargs, targs)):
            t1, t2 = lb.jvp(t)
            if quad is not None:
                quad += (t1 * t1).sum() + (t2 * t2).sum()
            cots.append(lb.vjp(t1, t2))
        self._pull_back(cots, sinks, tail)

    def normal_matvec(self, v, damp=None, return_quad=False):
        """J^T J v (+ damp * v): the Gauss-Newton product CG needs, one fused tile kernel per view.
        damp: None, a float (lambda: adds lambda v) or a GaussianModelDampMatrix (adds D v).
        return_quad=True also returns v^T J^T J v as a 0-dim device tensor (CG's p^T A p without a second pass)."""
        assert isinstance(v, GaussianModelState), "v must be an instance of GaussianModelState"
        if not self.cached:
            out = self.matvec_T(self.matvec(v))
            if damp is not None:
                out = out + v * damp
            return (out, None) if return_quad else out
        self._require_linearized()
        targs = self._tangent_inputs(v)
        gargs = self._gauss_args()
        flat, views, sinks = self._act_accumulators()
        quad = torch.zeros(1, dtype=torch.float32, device=flat.device) if return_quad else None
        tail = self._tail(views, v, damp)
        if self.ssim:
            self._normal_matvec_ssim(gargs, targs, sinks, quad, tail)
        else:
            _C.jtjv_batch(gargs, targs, self._get_plan(), sinks, out_vJJv=quad, accumulate=False,
                          n_slices=tail.n_slices, on_slice=tail)
        if self.dist and quad is not None:
            self.dist.all_reduce(quad)
        out = tail.finish()
        return (out, quad) if return_quad else out


class _ProductTail:
    """Callable handed to the library as the per-slice callback of gslm_jtjv_batch_sliced / gslm_backward_batch_sliced:
    the end of every parameter-space product.

    Single GPU: one call for [0, P): the activation VJP (+ damping) on the caller's stream.
    Multi-GPU, N_SLICES == 1 (default): one call for [0, P) on the caller's stream: the sum over ranks -- in place on
    the symmetric accumulator through this library's multimem kernel (csrc/peer_reduce.cu) from 3 ranks up, one NCCL
    allreduce otherwise -- then the activation VJP.
    Multi-GPU, N_SLICES > 1 (kept for measurement, slower on B200: DESIGN.md 5): the library cuts the last adjoint
    launch into Gaussian ranges; after each launch this object records an event on the compute stream and, on a
    communication stream, all-reduces the five row blocks of the range as ONE grouped NCCL call and runs the
    activation VJP of the range.  finish() joins the streams and returns the masked raw-space state."""

    _GROUPS = ("means3D", "shs", "scales", "rotations", "opacities")

    def __init__(self, F, views, v, damp, local=False):
        self.F, self.views, self.v, self.damp = F, views, v, damp
        self.out = GaussianModelState.from_gaussians(F.gaussians, zero=False)
        self.dist = None if local else F.dist
        self.n_slices = F.N_SLICES if self.dist else 1
        self.d7 = None
        if damp is not None:
            self.d7 = damp.values() if isinstance(damp, GaussianModelDampMatrix) else (float(damp),) * 7
        if self.dist and self.n_slices > 1:
            dev = self.out.flat.device
            if getattr(F, "_comm_stream", None) is None:
                F._comm_stream = torch.cuda.Stream(dev)
            self.comm = F._comm_stream
            self.cur = torch.cuda.current_stream(dev)
            P = F._act["means3D"].shape[0]
            self.rows = [views[n].view(P, -1) for n in self._GROUPS]

    def _vjp(self, g_lo, g_hi):
        vw = self.views
        grads = (vw["means3D"], vw["shs"], vw["scales"], vw["rotations"], vw["opacities"])
        layer = self.F._act["layer"]
        if self.d7 is None:
            layer.vjp(*grads, self.out, g_lo=g_lo, g_hi=g_hi)
        else:
            layer.vjp(*grads, self.out, self.v, self.d7[:6], g_lo=g_lo, g_hi=g_hi)

    def __call__(self, k, g_lo, g_hi):
        if not self.dist:
            self._vjp(g_lo, g_hi)
            return
        import torch.distributed as dist
        if self.n_slices == 1:
            # the whole vector at once, on the caller's stream: the multimem kernel over NVSwitch peer memory, or one
            # NCCL allreduce where no multicast mapping exists
            peer = getattr(self.F, "_peer", None)
            if peer is not None:
                peer.allreduce(self.flat_all())
            else:
                dist.all_reduce(self.flat_all())
            self._vjp(g_lo, g_hi)
            return
        ev = torch.cuda.Event()
        ev.record(self.cur)
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(ev)
            if g_hi > g_lo:
                with dist._coalescing_manager():          # one grouped NCCL launch for the five row blocks
                    for r in self.rows:
                        dist.all_reduce(r[g_lo:g_hi])
            self._vjp(g_lo, g_hi)

    def flat_all(self):
        """The packed 59 P-float accumulator the five group views alias."""
        return self.F._flat_acc

    def finish(self):
        F = self.F
        if self.dist and self.n_slices > 1:
            self.cur.wait_stream(self.comm)
            for t in self.views.values():       # produced on `cur`, consumed on `comm`
                t.record_stream(self.comm)
            self.out.flat.record_stream(self.comm)
        out = self.out
        if self.d7 is not None and out.exposure_grad.numel():
            out.exposure_grad.add_(self.v.exposure_grad, alpha=float(self.d7[6]))
        return GaussianModelState(*out.tensors(), param_mask=F.param_mask, splat_mask=F.splat_mask, _flat=out.flat)


class HostPipeline:
    """Streams host-resident parameter-space vectors through F.normal_matvec with the PCIe copies off the critical
    path: the host->device copy of vector k + 1 (copy-in stream) and the device->host copy of result k - 1 (copy-out
    stream) overlap the kernels of product k (the caller's stream).  Every product still pays its own H2D and D2H --
    they are just not serialised with the compute, as bench.py's strictly sequential leg is (`e2e.serial`).
    Valid for INDEPENDENT products (a block of probe vectors or right-hand sides); a CG whose vectors live on the host
    cannot overlap (v_{k+1} depends on result k) -- keep those vectors on the device (solver/lm.py).

    Multi-GPU: parameter-space vectors are replicated, so every rank moves only 1/N of each vector over its own PCIe
    link; one all_gather over NVLink (on a dedicated process group, so it does not queue behind the products'
    allreduce) completes the input, and each rank returns its 1/N slice of the result.

        pipe = F.host_pipeline()
        for v_host, out_host in work:        # pinned float32 [n] tensors
            pipe.submit(v_host, out_host)
        pipe.flush()                          # all results have landed in their out_host buffers
    """

    def __init__(self, F, depth=2):
        self.F = F
        self.depth = max(int(depth), 1)
        dev = F.gaussians._xyz.device
        self.dev = dev
        self.s_in, self.s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        self.states = [GaussianModelState.from_gaussians(F.gaussians) for _ in range(self.depth)]
        self.n = self.states[0].packed().numel()
        d = F.dist
        self.world, self.rank = (d.get_world_size(), d.get_rank()) if d else (1, 0)
        self.per = (self.n + self.world - 1) // self.world
        self.group = d.new_group(backend="nccl") if (d and dev.type == "cuda") else None
        self.stage = [torch.empty(self.per * self.world, dtype=torch.float32, device=dev) for _ in range(self.depth)] \
            if self.world > 1 else None
        self.ev_in = [torch.cuda.Event() for _ in range(self.depth)]
        self.ev_free = [None] * self.depth          # compute of the product that last read slot i has finished
        self.ev_out = []                             # pending result copies
        self.k = 0
        self.bytes_in = self.bytes_out = 0

    def submit(self, host_in, host_out):
        F, slot = self.F, self.k % self.depth
        self.k += 1
        cur = torch.cuda.current_stream(self.dev)
        flat = self.states[slot].packed()
        lo, hi = self.rank * self.per, min((self.rank + 1) * self.per, self.n)
        with torch.cuda.stream(self.s_in):
            if self.ev_free[slot] is not None:
                self.s_in.wait_event(self.ev_free[slot])
            if self.world == 1:
                flat.copy_(host_in, non_blocking=True)
            else:
                st = self.stage[slot]
                mine = st[self.rank * self.per:(self.rank + 1) * self.per]
                if hi > lo:
                    mine[:hi - lo].copy_(host_in[lo:hi], non_blocking=True)
                F.dist.all_gather_into_tensor(st, mine, group=self.group)
                flat.copy_(st[:self.n])
            self.ev_in[slot].record(self.s_in)
        self.bytes_in = 4 * max(hi - lo, 0)
        cur.wait_event(self.ev_in[slot])
        out = F.normal_matvec(self.states[slot])
        done = torch.cuda.Event()
        done.record(cur)
        self.ev_free[slot] = done
        res = out.packed()
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(done)
            if hi > lo:
                host_out[lo:hi].copy_(res[lo:hi], non_blocking=True)
            res.record_stream(self.s_out)
            e = torch.cuda.Event()
            e.record(self.s_out)
        self.ev_out.append(e)
        self.ev_out = self.ev_out[-2 * self.depth:]
        self.bytes_out = 4 * max(hi - lo, 0)

    def flush(self):
        """Make the caller's stream wait for every pending copy (so CUDA events on it bracket the whole pipeline) and
        block the host until the results are in host memory."""
        cur = torch.cuda.current_stream(self.dev)
        for e in self.ev_out:
            cur.wait_event(e)
        cur.wait_stream(self.s_in)
        for e in self.ev_out:
            e.synchronize()
        self.ev_out = []
