"""Host-side mirror of the reference's GaussianModel (src/scene/gaussian_model.{h,cpp}) and of
the loop body the reference's train() is missing (src/utils/train_utils.cpp:128-145).

Same member names and semantics as the reference: raw parameters xyz_/features_dc_/
features_rest_/scaling_ (log)/rotation_ (w,x,y,z)/opacity_ (logit), statistics max_radii2D_/
xyz_gradient_accum_/denom_, getters applying exp / normalise / sigmoid / cat, setup() with six
Adam groups, update_learning_rate(), oneup_SH_degree().  Differences, all following the
*intent* recorded in SURVEY.md §5.1: xyz_gradient_accum is (N,1) (defect 1), the xyz schedule
decays over position_lr_max_steps (defect 2), spatial_lr_scale defaults to 1 (defect 3), and
the six libtorch Adam instances are ONE fused CUDA launch (gs_adam_densify_step) with six
segments and six step counters.  torch is used for device memory only.
"""
from __future__ import annotations

from dataclasses import dataclass

import ctypes as C

import numpy as np
import torch

from .camera import get_expon_lr_func
from .rasterizer import GaussianRasterizationSettings, Rasterizer


@dataclass
class OptimizationParams:
    """Defaults of the reference's OptimizationParams (src/arguments/params.h:52-68)."""
    iterations: int = 30_000
    position_lr_init: float = 0.00016
    position_lr_final: float = 0.0000016
    position_lr_delay_mult: float = 0.01
    position_lr_max_steps: int = 30_000
    feature_lr: float = 0.0025
    opacity_lr: float = 0.05
    scaling_lr: float = 0.005
    rotation_lr: float = 0.001
    percent_dense: float = 0.01
    lambda_dssim: float = 0.2
    densification_interval: int = 100
    opacity_reset_interval: int = 3000
    densify_from_iter: int = 500
    densify_until_iter: int = 15_000
    densify_grad_threshold: float = 0.0002
    random_background: bool = False


@dataclass
class PipelineParams:
    """src/arguments/params.h:93-97"""
    convert_SHs_python: bool = False
    compute_cov3D_python: bool = False
    debug: bool = False


class GaussianModel:
    GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")  # order of setup(), gaussian_model.cpp:323-345

    def __init__(self, sh_degree: int, device: int = 0):
        self.max_sh_degree = int(sh_degree)
        self.active_sh_degree = 0
        self.device = torch.device("cuda", device)
        self.spatial_lr_scale = 1.0
        self.percent_dense = 0.01
        e = lambda *s: torch.empty(*s, device=self.device)
        self.xyz_ = e(0, 3); self.features_dc_ = e(0, 1, 3); self.features_rest_ = e(0, 0, 3)
        self.scaling_ = e(0, 3); self.rotation_ = e(0, 4); self.opacity_ = e(0, 1)
        self.max_radii2D_ = torch.empty(0, dtype=torch.int32, device=self.device)
        self.xyz_gradient_accum_ = e(0, 1); self.denom_ = e(0, 1)
        self._opt_ready = False

    # ---- construction from a synthetic scene (replaces create_from_pcd / Scene loading) ----
    def load_scene(self, scene, active_sh_degree=None):
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(self.device)
        self.xyz_ = d(scene.xyz); self.features_dc_ = d(scene.features_dc); self.features_rest_ = d(scene.features_rest)
        self.scaling_ = d(scene.log_scales); self.rotation_ = d(scene.raw_rotations); self.opacity_ = d(scene.opacity_logits)
        self.max_sh_degree = scene.max_sh_degree
        self.active_sh_degree = self.max_sh_degree if active_sh_degree is None else int(active_sh_degree)
        N = self.xyz_.shape[0]
        self.max_radii2D_ = torch.zeros(N, dtype=torch.int32, device=self.device)
        return self

    @property
    def N(self):
        return int(self.xyz_.shape[0])

    # ---- getters (src/scene/gaussian_model.cpp:270-304) ----
    def get_xyz(self): return self.xyz_
    def get_scaling(self): return torch.exp(self.scaling_)
    def get_rotation(self): return torch.nn.functional.normalize(self.rotation_, dim=1)
    def get_opacity(self): return torch.sigmoid(self.opacity_)
    def get_features(self): return torch.cat((self.features_dc_, self.features_rest_), dim=1)

    def oneup_SH_degree(self):
        """src/scene/gaussian_model.cpp:307-313"""
        if self.active_sh_degree < self.max_sh_degree:
            self.active_sh_degree += 1

    # ---- optimiser (src/scene/gaussian_model.cpp:316-361) ----
    def params(self):
        return [self.xyz_, self.features_dc_, self.features_rest_, self.opacity_, self.scaling_, self.rotation_]

    def setup(self, opt: OptimizationParams):
        N = self.N
        self.percent_dense = opt.percent_dense
        self.xyz_gradient_accum_ = torch.zeros((N, 1), device=self.device)
        self.denom_ = torch.zeros((N, 1), device=self.device)
        self.widths = [int(p.numel() // max(N, 1)) for p in self.params()]
        self.lr = [opt.position_lr_init * self.spatial_lr_scale, opt.feature_lr, opt.feature_lr / 20.0,
                   opt.opacity_lr, opt.scaling_lr, opt.rotation_lr]
        self.exp_avg = [torch.zeros_like(p) for p in self.params()]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params()]
        self.grads = [torch.zeros_like(p) for p in self.params()]
        self.grad_means2D = torch.zeros((N, 2), device=self.device)
        self.step_count = [0] * 6
        self.xyz_scheduler_args = get_expon_lr_func(opt.position_lr_init * self.spatial_lr_scale,
                                                    opt.position_lr_final * self.spatial_lr_scale, 0,
                                                    opt.position_lr_delay_mult, opt.position_lr_max_steps)
        self._opt_ready = True

    def update_learning_rate(self, iteration: int) -> float:
        """src/scene/gaussian_model.cpp:355-361"""
        self.lr[0] = self.xyz_scheduler_args(iteration)
        return self.lr[0]

    # ---- densification (published 3DGS algorithm; oracle/orc_densify.c is the specification) ----
    def densify_and_prune(self, rast: Rasterizer, max_grad: float, min_opacity: float, extent: float,
                          max_screen_size: float = 0.0, seed: int = 0):
        """Clone / split / prune in one classification + one scatter pass (csrc/densify.cu), carrying the Adam
        moments along; the statistics restart from zero, as after the published step.  Returns the counts
        [kept originals, kept clones, kept split parents, pruned]."""
        N = self.N
        counts = rast.densify_plan(self.xyz_gradient_accum_, self.denom_, self.opacity_, self.scaling_, max_grad,
                                   min_opacity, extent, self.percent_dense, max_screen_size)
        n_new = counts[0] + counts[1] + 2 * counts[2]
        src = self.params()
        shapes = [(n_new,) + tuple(p.shape[1:]) for p in src]
        new = lambda: [torch.empty(s, dtype=torch.float32, device=self.device) for s in shapes]
        dst_p, dst_m, dst_v = new(), new(), new()
        rast.densify_apply(N, self.widths, (src, self.exp_avg, self.exp_avg_sq), (dst_p, dst_m, dst_v), seed)
        (self.xyz_, self.features_dc_, self.features_rest_, self.opacity_, self.scaling_, self.rotation_) = dst_p
        self.exp_avg, self.exp_avg_sq = dst_m, dst_v
        self.grads = [torch.zeros_like(p) for p in dst_p]
        self.grad_means2D = torch.zeros((n_new, 2), device=self.device)
        self.xyz_gradient_accum_ = torch.zeros((n_new, 1), device=self.device)
        self.denom_ = torch.zeros((n_new, 1), device=self.device)
        self.max_radii2D_ = torch.zeros(n_new, dtype=torch.int32, device=self.device)
        return counts

    def reset_opacity(self, rast: Rasterizer):
        """opacity <- logit(min(sigmoid(opacity), 0.01)), fresh Adam state for the group (opacity_reset_interval_)."""
        rast.reset_opacity(self.opacity_, self.exp_avg[3], self.exp_avg_sq[3])

    # ---- checkpoint (src/scene/gaussian_model.cpp:228-262; files: checkpoint.py) ----
    def capture(self, tensors_path, opt_path_for_xyz=None, opt_path_for_f_dc=None, opt_path_for_f_rest=None,
                opt_path_for_opacity=None, opt_path_for_scaling=None, opt_path_for_rotation=None, fmt="libtorch",
                moments=None):
        """GaussianModel::capture: the eleven tensors + one Adam archive per parameter group, in the reference's own
        torch::save format by default."""
        from . import checkpoint
        checkpoint.capture(self, tensors_path, opt_path_for_xyz, opt_path_for_f_dc, opt_path_for_f_rest,
                           opt_path_for_opacity, opt_path_for_scaling, opt_path_for_rotation, fmt=fmt, moments=moments)

    def restore(self, params: OptimizationParams, tensors_path, opt_path_for_xyz=None, opt_path_for_f_dc=None,
                opt_path_for_f_rest=None, opt_path_for_opacity=None, opt_path_for_scaling=None,
                opt_path_for_rotation=None):
        """GaussianModel::restore (gaussian_model.cpp:247-262): setup(params), then every file that exists."""
        from . import checkpoint
        checkpoint.restore(self, tensors_path, opt_path_for_xyz, opt_path_for_f_dc, opt_path_for_f_rest,
                           opt_path_for_opacity, opt_path_for_scaling, opt_path_for_rotation, opt_params=params)

    def grad_sinks(self):
        g = self.grads
        return dict(means3D=g[0], sh_dc=g[1], sh_rest=g[2], opacities=g[3], scales=g[4], rotations=g[5],
                    means2D=self.grad_means2D)


class NativeTrainer:
    """ctypes handle on the NATIVE training loop (csrc/trainer.cu, gs_trainer_*): the model, the optimiser state and
    the pipelined loop live in C++; Python only hands over the initial parameters, the cameras and the pinned
    ground-truth images, and receives the losses.  Same schedules as Trainer / the reference's OptimizationParams."""

    def __init__(self, scene, opt: OptimizationParams | None = None, device: int = 0, bg=(0.0, 0.0, 0.0),
                 spatial_lr_scale: float = 1.0, lambda_dssim: float | None = None, seed: int = 0,
                 active_sh_degree: int | None = None):
        from . import _lib
        self._L = _lib
        o = opt or OptimizationParams()
        cfg = _lib.GsTrainConfig(
            o.position_lr_init, o.position_lr_final, o.position_lr_delay_mult, o.position_lr_max_steps, o.feature_lr,
            o.opacity_lr, o.scaling_lr, o.rotation_lr, o.percent_dense,
            o.lambda_dssim if lambda_dssim is None else lambda_dssim, o.densification_interval,
            o.opacity_reset_interval, o.densify_from_iter, o.densify_until_iter, o.densify_grad_threshold,
            spatial_lr_scale, (C.c_float * 3)(*bg), seed,
            scene.max_sh_degree if active_sh_degree is None else active_sh_degree,   # (as GaussianModel.load_scene)
            int(o.random_background))
        f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        self._init = [f32(scene.xyz), f32(scene.features_dc), f32(scene.features_rest), f32(scene.opacity_logits),
                      f32(scene.log_scales), f32(scene.raw_rotations)]
        self.K = 1 + self._init[2].shape[1]
        self._t = C.c_void_p()
        ptr = lambda a: a.ctypes.data_as(C.c_void_p) if a.size else None
        rc = _lib.lib.gs_trainer_create(C.byref(self._t), device, C.byref(cfg), int(self._init[0].shape[0]), self.K,
                                        *[ptr(a) for a in self._init])
        if rc != 0:
            raise RuntimeError("gs_trainer_create: " + (_lib.lib.gs_last_error(None) or b"").decode())
        self._keep = []

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what}: " + (self._L.lib.gs_trainer_last_error(self._t) or b"").decode())

    def add_view(self, cam, gt_host: torch.Tensor):
        """cam: camera.Camera; gt_host: (3,H,W) float32 host tensor (pinned: asynchronous copies); kept alive here."""
        assert gt_host.dtype == torch.float32 and not gt_host.is_cuda and gt_host.is_contiguous()
        self._keep.append(gt_host)
        self._check(self._L.lib.gs_trainer_add_view(self._t, C.byref(cam._st), C.c_void_p(gt_host.data_ptr())), "add_view")

    def run(self, n: int, cyclic: bool = False, densify_extent: float = 0.0):
        losses = np.zeros(n, np.float32)
        self._check(self._L.lib.gs_trainer_run(self._t, int(n), losses.ctypes.data_as(C.c_void_p), int(cyclic),
                                               float(densify_extent)), "gs_trainer_run")
        return losses.tolist()

    def save(self, path: str):
        """gs_trainer_save: the live training state into one flat-container file (checkpoint.native_to_libtorch
        converts it into the reference's torch::save archives)."""
        self._check(self._L.lib.gs_trainer_save(self._t, str(path).encode()), "gs_trainer_save")

    def load(self, path: str):
        self._check(self._L.lib.gs_trainer_load(self._t, str(path).encode()), "gs_trainer_load")

    def set_pair_capacity(self, pairs: int):
        """Fixed (tile, Gaussian) pair capacity (default: probed by the first run with one forward per view)."""
        self._check(self._L.lib.gs_trainer_set_pair_capacity(self._t, int(pairs)), "set_pair_capacity")

    @property
    def N(self):
        return int(self._L.lib.gs_trainer_num_gaussians(self._t))

    @property
    def iteration(self):
        return int(self._L.lib.gs_trainer_iteration(self._t))

    @property
    def retries(self):
        return int(self._L.lib.gs_trainer_retries(self._t))

    def params(self):
        N = self.N
        shapes = [(N, 3), (N, 1, 3), (N, self.K - 1, 3), (N, 1), (N, 3), (N, 4)]
        out = [np.zeros(s, np.float32) for s in shapes]
        self._check(self._L.lib.gs_trainer_get_params(self._t, *[a.ctypes.data_as(C.c_void_p) if a.size else None
                                                                   for a in out]), "get_params")
        return out

    def close(self):
        if getattr(self, "_t", None):
            self._L.lib.gs_trainer_destroy(self._t)
            self._t = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ViewSampler:
    """Camera sampling WITHOUT replacement, as the training loop the reference transliterates does it: a stack of
    all training views, one popped at a random position per iteration, refilled when empty
    (loop skeleton: src/utils/train_utils.cpp:128-145).  Seeded, so that every rank of a data-parallel job can
    draw the same permutation and take its own slice of it."""

    def __init__(self, n_views: int, seed: int = 0):
        self.n, self.rng, self.stack = int(n_views), np.random.Generator(np.random.Philox(seed)), []

    def __call__(self) -> int:
        if not self.stack:
            self.stack = list(range(self.n))
        return self.stack.pop(int(self.rng.integers(0, len(self.stack))))


def render(viewpoint_camera, pc: GaussianModel, pipe: PipelineParams, bg_color, rast: Rasterizer,
           scaling_modifier: float = 1.0, fused_activations: bool = True, with_visibility: bool = True,
           cache_settings: bool = True):
    """The missing render(viewpoint_cam, gaussians, pipe, background) call.  Returns
    dict(render=(3,H,W), radii=(N) int32, visibility_filter=radii>0).
    fused_activations=True feeds the RAW parameters (activations fused in the kernels);
    False reproduces the getter chain of the reference (exp / normalise / sigmoid / cat)."""
    key = (tuple(bg_color), pc.active_sh_degree, pc.max_sh_degree, scaling_modifier, pipe.debug)
    cache = viewpoint_camera.__dict__.setdefault("_raster_settings", {})
    st = cache.get(key) if cache_settings else None
    if st is None:   # raster settings of a camera are fixed (src/scene/camera.cpp:56-71): build them once
        st = GaussianRasterizationSettings.from_camera(
            viewpoint_camera, bg_color, pc.active_sh_degree, pc.max_sh_degree, scaling_modifier, pipe.debug)
        if cache_settings:      # (a random background changes every iteration: not cached)
            cache[key] = st
    if fused_activations:
        color, radii = rast.forward(st, pc.xyz_, pc.scaling_, pc.rotation_, pc.opacity_, sh_dc=pc.features_dc_,
                                    sh_rest=pc.features_rest_, raw_params=True)
    else:
        color, radii = rast.forward(st, pc.get_xyz(), pc.get_scaling(), pc.get_rotation(), pc.get_opacity(),
                                    shs=pc.get_features())
    out = dict(render=color, radii=radii)
    if with_visibility:   # one more elementwise kernel; the fused Adam/stats step derives it from radii itself
        out["visibility_filter"] = radii > 0
    return out


class Trainer:
    """One training iteration = the loop body missing at src/utils/train_utils.cpp:138-145:
    render -> loss -> backward -> densification stats -> optimizer step, all on the caller's
    stream; only the ground-truth image comes from the host and only the loss goes back."""

    def __init__(self, gaussians: GaussianModel, opt: OptimizationParams, pipe: PipelineParams | None = None,
                 bg=(0.0, 0.0, 0.0), lambda_dssim: float | None = None):
        self.g, self.opt, self.pipe = gaussians, opt, pipe or PipelineParams()
        self.bg = tuple(bg)
        self.rast = Rasterizer(gaussians.device.index or 0)
        if not gaussians._opt_ready:
            gaussians.setup(opt)
        self.lambda_dssim = opt.lambda_dssim if lambda_dssim is None else lambda_dssim
        self.iteration = 0
        # active SH degree as a FUNCTION of the iteration number (one more every 1000 iterations,
        # src/utils/train_utils.cpp:133-136 / gaussian_model.cpp:307-313), so that repeating an iteration after a
        # capacity overflow cannot raise it twice
        self._sh_degree_at_0 = gaussians.active_sh_degree
        self.scale_capacity_on_densify = True
        self.bg_seed = 0                # random_background: colour = f(bg_seed, iteration)
        self._cur_bg = self.bg
        self._gt_dev = None
        self._loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
        # capacity mode (no host round-trip inside the forward): enabled by reserve_pairs()
        self.pair_capacity = 0
        self.grad_sync = None           # multi-GPU: callable run between backward and the optimiser step (all-reduce)
        self._copy_stream = None
        self._gt_ready = None
        self._guard_tensor = None
        self.skip_guard = None          # device float the Adam step honours (all-reduced slab tail when multi-GPU)
        self.retries = 0
        self.fused_step = True          # single GPU: backward and Adam in one call (gs_backward_adam_step)
        self.on_resize = None           # callable run after a densification step changed N (e.g. rebuild a gradient slab)
        self.peer = None                # multi-GPU: parallel.PeerShardedAdam (exchange fused into K4b / K5)
        self.sharded = None             # multi-GPU: parallel.ShardedAdam (reduce-scatter / sharded Adam / all-gather)

    @staticmethod
    def _mix32(x: int) -> int:
        x &= 0xFFFFFFFF
        x ^= x >> 16
        x = (x * 0x7feb352d) & 0xFFFFFFFF
        x ^= x >> 15
        x = (x * 0x846ca68b) & 0xFFFFFFFF
        x ^= x >> 16
        return x

    def background_for(self, iteration: int):
        """The background colour of iteration `iteration`: the fixed one, or with opt.random_background
        (random_background_, src/arguments/params.h:68) a fresh uniform colour per iteration -- a function of (seed,
        iteration), the same counter-based generator as the native trainer's."""
        if not self.opt.random_background:
            return self.bg
        m = self._mix32
        return tuple((m(self.bg_seed ^ m(iteration * 3 + ch + 0x51ed270b)) >> 8) / 16777216.0 for ch in range(3))

    def shard_optimizer(self, rank: int, world: int, group=None):
        """Shard the Adam state over the ranks (parallel.ShardedAdam).  Call before reserve_pairs()."""
        from . import parallel
        adam = lambda segs, guard: self.rast.adam_step_flat(segs, skip_guard=guard)
        self.sharded = parallel.ShardedAdam(self.g, rank, world, adam, group=group)
        return self.sharded

    def shard_optimizer_peer(self, rank: int, world: int, group=None, pull: bool = False):
        """Fused multi-GPU step over peer memory (parallel.PeerShardedAdam).  Call before reserve_pairs().
        pull=True: the layout for train_batch_device (several views per rank and step, one exchange)."""
        from . import parallel
        self.peer = parallel.PeerShardedAdam(self.g, self.rast, rank, world, group=group, pull=pull)
        return self.peer

    def reserve_pairs(self, pairs: int, guard: torch.Tensor | None = None):
        """Fix the (tile, Gaussian) pair capacity so that forwards run without synchronising the host.  An
        iteration whose forward overflows is skipped on the device (Adam guard) and repeated by train_step
        with twice the capacity.  guard: caller-owned device float raised on overflow (the gradient slab's
        guard element when the gradients are all-reduced); default: the context's own flag."""
        self.pair_capacity = int(pairs)
        self.rast.set_duplicate_capacity(self.pair_capacity)
        self._guard_tensor = guard
        if guard is not None:
            self.rast.set_overflow_mirror(guard)
            self.skip_guard = guard.data_ptr()
        else:
            self.skip_guard = self.rast.overflow_guard_ptr()

    def forward_backward(self, cam, gt_dev: torch.Tensor, accumulate: bool = False, gt_host=None,
                         fused_optimizer: bool = False):
        """render + loss + backward for one view with the gradients left in the model's sinks.
        gt_host: pinned host image to copy into gt_dev first.  The image is not needed before the loss, so the
        copy is issued on a side stream AFTER the render has been enqueued and runs underneath it; the compute
        stream only waits for it in front of the loss kernel."""
        if self._guard_tensor is not None and not accumulate:
            self._guard_tensor.zero_()   # caller-owned guard: lowered at the start of every iteration
        out = render(cam, self.g, self.pipe, self._cur_bg, self.rast, with_visibility=False,
                     cache_settings=not self.opt.random_background)
        if gt_host is not None:
            main = torch.cuda.current_stream(self.g.device)
            if self._copy_stream is None:
                self._copy_stream = torch.cuda.Stream(device=self.g.device)
                self._gt_ready = torch.cuda.Event()
                self._gt_free = torch.cuda.Event()
                self._gt_free.record(main)
            self._copy_stream.wait_event(self._gt_free)   # the previous iteration's loss kernel has read gt_dev
            with torch.cuda.stream(self._copy_stream):
                gt_dev.copy_(gt_host, non_blocking=True)
                self._gt_ready.record(self._copy_stream)
            main.wait_event(self._gt_ready)
        loss, dL = self.rast.loss(out["render"], gt_dev, self.lambda_dssim)
        if gt_host is not None:
            self._gt_free.record(torch.cuda.current_stream(self.g.device))
        if fused_optimizer:
            g = self.g
            g.step_count = [s + 1 for s in g.step_count]
            self.rast.backward_adam_step(dL, g.exp_avg, g.exp_avg_sq, g.lr, g.step_count, max_radii2D=g.max_radii2D_,
                                         xyz_grad_accum=g.xyz_gradient_accum_, denom=g.denom_,
                                         skip_guard=self.skip_guard if self.pair_capacity else None)
        else:
            self.rast.backward(dL, grads=self.g.grad_sinks(), accumulate=accumulate)
        return loss, out

    def optimizer_step(self, radii):
        g = self.g
        g.step_count = [s + 1 for s in g.step_count]
        if self.peer is not None:
            self.peer.step(g.lr, g.step_count, radii=radii, dL_dmeans2D=g.grad_means2D, max_radii2D=g.max_radii2D_,
                           xyz_grad_accum=g.xyz_gradient_accum_, denom=g.denom_, use_guard=bool(self.pair_capacity))
            return
        if self.sharded is not None:
            self.sharded.step(g.lr, g.step_count, use_guard=bool(self.pair_capacity))
            none6 = [None] * 6    # statistics only (per-rank partial sums, merged when consumed)
            self.rast.adam_densify_step(g.N, none6, none6, none6, none6, g.widths, g.lr, g.step_count, radii=radii,
                                        dL_dmeans2D=g.grad_means2D, max_radii2D=g.max_radii2D_,
                                        xyz_grad_accum=g.xyz_gradient_accum_, denom=g.denom_,
                                        skip_guard=self.skip_guard if self.pair_capacity else None)
            return
        self.rast.adam_densify_step(g.N, g.params(), g.grads, g.exp_avg, g.exp_avg_sq, g.widths, g.lr, g.step_count,
                                    radii=radii, dL_dmeans2D=g.grad_means2D, max_radii2D=g.max_radii2D_,
                                    xyz_grad_accum=g.xyz_gradient_accum_, denom=g.denom_,
                                    skip_guard=self.skip_guard if self.pair_capacity else None)

    # ---- checkpoint of the LIVE training state (the start_checkpoint hook of src/utils/train_utils.cpp:110-113) ----
    def capture(self, folder: str, fmt: str = "libtorch"):
        """Write the seven files of a checkpoint of the model being trained into `folder`, named *_<iteration>
        (search_for_max_iteration finds the newest).  Drains nothing: call between iterations.  In peer mode the
        sharded Adam moments are gathered first (every rank gets the full state; rank 0 writes)."""
        from . import checkpoint
        torch.cuda.synchronize(self.g.device)
        moments = None
        if self.peer is not None:
            moments = self.peer.gather_moments()
            import torch.distributed as dist
            if dist.is_initialized() and dist.get_rank(self.peer.group) != 0:
                return None
        assert self.sharded is None, "checkpointing the NCCL-sharded optimiser is not wired (use the peer mode)"
        paths = checkpoint.checkpoint_paths(folder, self.iteration)
        self.g.capture(*paths, fmt=fmt, moments=moments)
        return paths

    def restore(self, folder: str, iteration: int | None = None):
        """Load the checkpoint of `iteration` (default: the newest in `folder`) into the live model: parameters,
        statistics, Adam moments and step counters, learning rates, the iteration counter and the SH-degree
        schedule.  Returns the iteration restored."""
        from . import checkpoint
        assert self.sharded is None, "restoring into the NCCL-sharded optimiser is not wired (use the peer mode)"
        if iteration is None:
            iteration = checkpoint.search_for_max_iteration(folder)
            if iteration < 0:
                raise FileNotFoundError(f"no checkpoint in {folder}")
        paths = checkpoint.checkpoint_paths(folder, iteration)
        torch.cuda.synchronize(self.g.device)
        g = self.g
        old = self.peer
        if old is not None:            # the parameters are views of the peer buffer: release it, re-shard afterwards
            old.detach_model()
            old.close()
            self.peer = None
        g.restore(self.opt, *paths)
        self.iteration = int(iteration)
        self._sh_degree_at_0 = g.active_sh_degree - self.iteration // 1000
        if old is not None:
            from . import parallel
            self.peer = parallel.PeerShardedAdam(g, self.rast, old.rank, old.world, group=old.group,
                                                 moments=(g.exp_avg, g.exp_avg_sq))
            if self.pair_capacity:
                self.reserve_pairs(self.pair_capacity, guard=self.peer.guard)
        elif self.on_resize is not None:
            self.on_resize()
        return self.iteration

    def densification_due(self, iteration: int) -> bool:
        o = self.opt
        return (iteration < o.densify_until_iter and iteration > o.densify_from_iter
                and iteration % o.densification_interval == 0) or \
               (iteration < o.densify_until_iter and iteration % o.opacity_reset_interval == 0)

    def densification_step(self, iteration: int, extent: float, white_background: bool = False):
        """The block of the published loop that follows the optimiser bookkeeping of iteration `iteration`
        (schedule parameters: src/arguments/params.h:61-67).  Changes N: the caller must have drained any
        pipelined iterations.  Data parallel (all-reduce or peer mode): the per-rank statistics are merged first, so
        every rank takes the same decisions with the same seed; the peer mode re-shards the optimiser state over
        new peer buffers."""
        o, g = self.opt, self.g
        assert self.sharded is None, "densification with the NCCL-sharded optimiser is not wired (use the peer mode)"
        done = None
        if iteration < o.densify_until_iter:
            if iteration > o.densify_from_iter and iteration % o.densification_interval == 0:
                size_threshold = 20.0 if iteration > o.opacity_reset_interval else 0.0
                if self.peer is not None or self.grad_sync is not None:
                    # data parallel: every rank accumulated the statistics of its own views -> merge, then every
                    # rank takes the same decisions (same statistics, same seed) on its replica
                    from . import parallel
                    parallel.merge_densify_stats(g.max_radii2D_, g.xyz_gradient_accum_, g.denom_)
                if self.peer is not None:
                    # the step needs (and produces) the FULL optimiser state: gather the moment shards, release the
                    # peer buffers (N changes), densify, shard again over new buffers
                    old = self.peer
                    m_full, v_full = old.gather_moments()
                    old.detach_model()
                    old.close()
                    g.exp_avg, g.exp_avg_sq = m_full, v_full
                n_before = g.N
                done = g.densify_and_prune(self.rast, o.densify_grad_threshold, 0.005, extent, size_threshold,
                                           seed=iteration)
                if self.pair_capacity and self.scale_capacity_on_densify and g.N > n_before:
                    # the (tile, Gaussian) pair count grows with N: keep the fixed capacity ahead of it
                    self.pair_capacity = int(self.pair_capacity * (g.N / max(n_before, 1)) * 1.1) + 1024
                    self.rast.set_duplicate_capacity(self.pair_capacity)
                if self.peer is not None:
                    from . import parallel
                    self.peer = parallel.PeerShardedAdam(g, self.rast, old.rank, old.world, group=old.group,
                                                         moments=(g.exp_avg, g.exp_avg_sq))
                    if self.pair_capacity:
                        self.reserve_pairs(self.pair_capacity, guard=self.peer.guard)
                elif self.on_resize is not None:
                    self.on_resize()
            if iteration % o.opacity_reset_interval == 0 or (white_background and iteration == o.densify_from_iter):
                if self.peer is not None:
                    # every rank resets its whole replica (same values everywhere); the moments are sharded: the
                    # kernel gets throw-away full-size moments and the owned shard is cleared here
                    scratch = torch.empty_like(g.opacity_)
                    self.rast.reset_opacity(g.opacity_, scratch, scratch)
                    g.exp_avg[3].zero_()
                    g.exp_avg_sq[3].zero_()
                    torch.cuda.synchronize(g.opacity_.device)
                    self.peer.barrier()
                else:
                    g.reset_opacity(self.rast)
        return done

    def check_overflow(self) -> bool:
        """After the step's stream synchronisation: did the last forward overflow the fixed capacity?  If so the
        device skipped the optimiser step; roll the host-side counters back and double the capacity."""
        if not self.pair_capacity:
            return False
        D, overflow = self.rast.forward_status()
        if overflow:
            self.g.step_count = [s - 1 for s in self.g.step_count]
            self.iteration -= 1
            self.retries += 1
            self.pair_capacity = max(2 * self.pair_capacity, int(1.25 * D))
            self.rast.set_duplicate_capacity(self.pair_capacity)
        return overflow

    def apply_schedules(self, iteration: int):
        """Position learning rate and active SH degree of iteration `iteration` (idempotent)."""
        self.g.update_learning_rate(iteration)
        self.g.active_sh_degree = min(self.g.max_sh_degree, self._sh_degree_at_0 + iteration // 1000)
        self._cur_bg = self.background_for(iteration)

    def train_step_device(self, cam, gt_dev: torch.Tensor, gt_host=None, iteration: int | None = None):
        """Device-resident step: ground truth already in HBM; returns the loss as a device tensor.
        iteration: run the step AS iteration number `iteration` (a repeated iteration after a capacity overflow keeps
        its own number and with it its learning rate and SH degree); default: the next one."""
        if iteration is None:
            self.iteration += 1
            iteration = self.iteration
        self.apply_schedules(iteration)
        if self.fused_step and self.grad_sync is None and self.peer is None and self.sharded is None:
            # single GPU: K4b and K5 fused (gs_backward_adam_step) -- the gradient tensors never reach HBM
            loss, out = self.forward_backward(cam, gt_dev, gt_host=gt_host, fused_optimizer=True)
            return loss
        loss, out = self.forward_backward(cam, gt_dev, gt_host=gt_host)
        if self.grad_sync is not None:
            self.grad_sync()
        self.optimizer_step(out["radii"])
        return loss

    def train_batch_device(self, cams, gts_dev, gts_host=None, exchange_events=None):
        """ONE optimiser step on a batch of views (north_star's multi-view step): render + loss +
        backward per view with the parameter gradients ACCUMULATED over the views (the gradient of the summed
        per-view losses), the densification statistics taken view by view (xyz_gradient_accum += |dL/dmean2D| and
        denom += 1 are per-view sums, src/scene/gaussian_model.h:18-20), then ONE gradient exchange and ONE Adam.
        Data parallel: every rank passes ITS views of the batch; exchange = the trainer's mode (all-reduce of the
        gradient slab, or the pull layout of the peer optimiser: shard_optimizer_peer(pull=True)).
        Needs a caller-owned overflow guard (reserve_pairs(pairs, guard=tensor)): the context's own flag is reset by
        every forward, the caller's stays raised for the rest of the step.  exchange_events: two CUDA events recorded
        around the exchange + Adam part (benchmarks report it separately).  Returns the mean loss (device tensor)."""
        assert self.pair_capacity and self._guard_tensor is not None, "train_batch_device: reserve_pairs(pairs, guard=...)"
        assert self.sharded is None and (self.peer is None or self.peer.pull), "batch steps: all-reduce or pull layout"
        g = self.g
        self.iteration += 1
        self.apply_schedules(self.iteration)
        none6 = [None] * 6
        total = None
        for v, cam in enumerate(cams):
            loss, out = self.forward_backward(cam, gts_dev[v], accumulate=v > 0,
                                              gt_host=None if gts_host is None else gts_host[v])
            self.rast.adam_densify_step(g.N, none6, none6, none6, none6, g.widths, g.lr, [1] * 6, radii=out["radii"],
                                        dL_dmeans2D=g.grad_means2D, max_radii2D=g.max_radii2D_,
                                        xyz_grad_accum=g.xyz_gradient_accum_, denom=g.denom_, skip_guard=self.skip_guard)
            total = loss if total is None else total + loss
        if exchange_events is not None:
            exchange_events[0].record()
        if self.grad_sync is not None:
            self.grad_sync()
        self.optimizer_step(None)
        if exchange_events is not None:
            exchange_events[1].record()
        return total / len(cams)

    def train_step(self, cam, gt_host: torch.Tensor) -> float:
        """End-to-end step with HOST buffers: pinned ground-truth image H2D (overlapped with the render, see
        forward_backward), loss D2H."""
        dev = self.g.device
        if self._gt_dev is None or self._gt_dev.shape != gt_host.shape:
            self._gt_dev = torch.empty(gt_host.shape, dtype=torch.float32, device=dev)
        while True:
            loss = self.train_step_device(cam, self._gt_dev, gt_host=gt_host)
            self._loss_host.copy_(loss, non_blocking=True)
            torch.cuda.current_stream(self.g.device).synchronize()
            if not self.check_overflow():   # capacity mode only: an overflowed iteration was skipped on the device
                return float(self._loss_host[0])

    def train(self, cams, gts_host, n_steps: int, first: int = 0, depth: int = 2, on_loss=None,
              densify_extent: float | None = None, sampler=None):
        """The training loop of the reference (src/utils/train_utils.cpp:120-150: for each iteration pick a view,
        render, loss, backward, optimiser step, log) with HOST ground-truth images, as a software pipeline: the
        host enqueues iteration i while the device is still working on iteration i - depth + 1, so the device
        never waits for the host.  Every iteration still copies its pinned ground truth host->device and its loss
        device->host; the host just reads loss i after it has enqueued iteration i + depth - 1.
        cams / gts_host: sequences cycled by iteration index (first + i), or drawn by `sampler` (a ViewSampler:
        without replacement, as the published loop does).  Returns the list of per-iteration
        losses (floats, in order); on_loss(i, loss) is called as each arrives.

        Fixed pair capacity (reserve_pairs) is required -- the synchronous forward would stall the pipeline.  An
        iteration whose forward overflows the capacity is skipped on the device (Adam guard: parameters, moments and
        densification statistics untouched).  The host learns of it `depth - 1` iterations late and recovers AT ONCE:
        it drains the pipeline, takes back the step counters of the void iterations, grows the capacity and runs
        those iterations again under their ORIGINAL iteration numbers (learning rate, SH degree) before it enqueues
        anything new -- so a capacity outgrown in the middle of a long run (densification) costs a few repeated
        iterations, never the rest of the schedule.  (Iterations that were already in flight behind a void one and
        did fit have run with a bias-correction counter one ahead.)

        densify_extent (the scene extent): run the densification schedule of the published loop
        (densification_step) -- the pipeline is drained first, since N changes, and the pair capacity follows N.
        Unlike the published loop, whose freshly concatenated parameters carry no gradient in that iteration, the
        iteration's optimiser step has already been applied when the model is densified."""
        assert self.pair_capacity, "Trainer.train needs reserve_pairs() (capacity mode)"
        dev = self.g.device
        shape = gts_host[0].shape
        if self._gt_dev is None or self._gt_dev.shape != shape:
            self._gt_dev = torch.empty(shape, dtype=torch.float32, device=dev)
        ring = max(2, int(depth)) + 1
        if getattr(self, "_ring", None) is None or self._ring[0].shape[0] != ring:
            self._ring = (torch.zeros(ring, 2, dtype=torch.float32).pin_memory(),
                          [torch.cuda.Event() for _ in range(ring)])
        ctx_guard = self.rast.wrap(self.rast.overflow_guard_ptr(), 1)   # the context's own flag (stable address)
        slots, events = self._ring
        stream = torch.cuda.current_stream(dev)
        losses, pending, void = [None] * n_steps, [], []

        def retire():
            j, slot, it_no = pending.pop(0)
            events[slot].synchronize()
            if slots[slot, 1] != 0.0:      # iteration j overflowed: skipped on the device
                void.append((j, it_no))
                return
            losses[j] = float(slots[slot, 0])
            if on_loss is not None:
                on_loss(j, losses[j])

        drawn = {}

        def enqueue(j, slot, it_no):
            if j not in drawn:      # (a repeated iteration after an overflow keeps its view)
                drawn[j] = sampler() if sampler is not None else (first + j) % len(cams)
            v = drawn[j]
            loss = self.train_step_device(cams[v], self._gt_dev, gt_host=gts_host[v], iteration=it_no)
            slots[slot, 0:1].copy_(loss.reshape(1), non_blocking=True)
            # (looked up per iteration: a densification step in peer mode installs a new caller-owned guard)
            slots[slot, 1:2].copy_(self._guard_tensor if self._guard_tensor is not None else ctx_guard, non_blocking=True)
            events[slot].record(stream)
            pending.append((j, slot, it_no))

        def recover():
            """Drain, then repeat every void iteration (one at a time, growing the capacity until it fits)."""
            while pending:
                retire()
            while void:
                batch = list(void)
                del void[:]
                D, _ = self.rast.forward_status()
                self.pair_capacity = max(2 * self.pair_capacity, int(1.25 * D))
                self.rast.set_duplicate_capacity(self.pair_capacity)
                self.retries += len(batch)
                for j, it_no in batch:
                    self.g.step_count = [s - 1 for s in self.g.step_count]   # the void attempt's increment
                    enqueue(j, 0, it_no)
                    retire()

        for i in range(n_steps):
            if len(pending) >= depth:
                retire()
            if void:
                recover()
            self.iteration += 1
            enqueue(i, i % ring, self.iteration)
            if densify_extent is not None and self.densification_due(self.iteration):
                recover()
                self.densification_step(self.iteration, densify_extent)
        recover()
        return losses
