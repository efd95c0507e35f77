#!/usr/bin/env python
"""bench.py — the headline benchmark of BASELINE.json on B200.

A "step" is ONE training iteration of the reference's (missing) loop body on one camera view
per rank: render (K1-K3) -> L1 + D-SSIM loss -> backward (K4a, K4b) -> fused Adam + densification stats
(K5), on the synthetic 1.6M-Gaussian / 1246x825 / SH3 scene (BASELINE.json configs[2], the
configuration the metric is quoted on).  metric = views/s (= train it/s at N=1).  With N>1 each
rank renders its own camera (camera sharding, replicated Gaussians); by default (--dp-mode peer) the
gradient exchange is fused into the kernels on either side over NVLink peer memory (K4b stores into
the owner ranks' slots, the sharded Adam kernel sums them and stores the new parameters into every
replica); --dp-mode allreduce is the library baseline (one NCCL all-reduce + replicated Adam).
The headline runs with WARM Adam state (every moment non-zero: the kernels' exact skip of
never-touched vectors cannot fire); `best_case` reports the cold-state step beside it.

Printed JSON keys follow the driver contract; see DESIGN.md "Measurement" for the definitions
of roofline / cpu_baseline / e2e / gpu_launches.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
from ctypes import c_int as C_int

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "train_views_per_s"
UNIT = "views/s"
DEFAULT_CONFIG = "cfg2_1p6m_1246x825_sh3"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=DEFAULT_CONFIG)
    ap.add_argument("--views", type=int, default=8, help="distinct camera views cycled per rank")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dp-mode", default="peer", choices=["peer", "sharded", "allreduce"],
                    help="N>1: peer = exchange fused into K4b/K5 over NVLink peer memory (sharded Adam); sharded = NCCL "
                         "reduce-scatter + sharded Adam + all-gather; allreduce = NCCL all-reduce + replicated Adam")
    ap.add_argument("--no-fused", action="store_true", help="N=1: separate K4b and K5 launches instead of the fused kernel")
    ap.add_argument("--e2e-api", default="loop", choices=["loop", "step", "native"],
                    help="e2e leg: Trainer.train (pipelined Python loop, default), one synchronous Trainer.train_step per "
                         "step, or the native C++ loop gs_trainer_run (N=1 only; wall-clock timed)")
    ap.add_argument("--N", type=int, default=None, help="override the Gaussian count (debugging only)")
    ap.add_argument("--adam-state", default="warm", choices=["warm", "cold"],
                    help="warm (default, the headline): every Adam moment non-zero before the timed region, as in a "
                         "scene whose Gaussians have all received a gradient at some point -- the kernels' exact skip "
                         "of never-touched vectors cannot fire (adam_idle_frac 0).  cold: moments start at zero, the "
                         "Gaussians this synthetic ball hides from every view stay idle and are skipped (best case)")
    ap.add_argument("--variant", default="default", choices=["default", "stress"],
                    help="stress: opacity logits shifted by -2 (long blend lists, little early termination)")
    ap.add_argument("--sustain-seconds", type=float, default=2.0,
                    help="length of the continuous run behind the `sustained` key (0: skip)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the secondary measurements (best_case, stress, sustained, north-star batch)")
    ap.add_argument("--batch-views", type=int, default=0,
                    help="> 0: north_star's multi-view step (BASELINE.json configs[4]): ONE optimiser step per batch of this "
                         "many views, split V/g per GPU, gradients accumulated locally over a rank's views, ONE exchange "
                         "(strong scaling: the batch is the same at every GPU count).  Default config for it: cfg4")
    ap.add_argument("--mode", default="train", choices=["train", "render", "raster"],
                    help="train: the headline train step (default).  render: forward-only render throughput sweep over "
                         "the config's cameras (BASELINE.json configs[3]: Mpix/s with the sort / blend breakdown).  raster: "
                         "one forward + backward of the render path, GPU beside the CPU oracle (BASELINE.json configs[0])")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# ---- clocks sampler (B200_PROFILING.md recipe) -------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            c = [x.strip() for x in r.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            try:
                pw.append(float(c[3]))
            except ValueError:
                pass
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "sm_min_mhz": min(sm) if sm else None,
                "power_w_max": max(pw) if pw else None}


# ---- algorithmic bytes ---------------------------------------------------
def stage_bytes(N, N_vis, D, D_eff, P, T, K, active_floats=None):
    """active_floats: parameter floats whose Adam vector is NOT at the exact fixed point g = m = v = 0 (counted on the
    device after the timed loop); only those are written back (moments + parameter, 12 B).  None: all of them."""
    w = 11 + 3 * K
    if active_floats is None:
        active_floats = w * N
    tile_bits = max(1, int(np.ceil(np.log2(max(T, 2)))))
    passes = int(np.ceil(tile_bits / 8))
    kb = 2 if T <= 65536 else 4
    return {
        "preprocess": (44 + 12 * K) * N + 48 * N_vis + 8 * (N - N_vis) + 16 * N,   # + 16-B rectangle/alive-mask record
        # two-level sort: argsort of N depth keys, 4 passes of (hist 4 + read 8 + write 8) B
        "depth_sort": 4 * 20 * N - 8 * N,
        # scan along the depth order: order id (4) + 16-B rectangle record, read twice; (tile, id) pairs written once
        "scan_duplicate": 2 * 20 * N + (kb + 4) * D,
        "sort": passes * (kb + 2 * (kb + 4)) * D,
        "ranges": 8 * T,
        # deferred SH -> RGB of the split-phase multi-GPU step (gs_set_color_gate): xyz + dc + rest + radius read,
        # colour + clamp mask written; K1 then reads 12 K bytes per Gaussian less (subtracted where it is used)
        "sh_color": (16 + 12 * K) * N + 13 * N_vis,
        "blend_fwd": 40 * D_eff + 20 * P,
        "blend_bwd": 40 * D_eff + 20 * P + 36 * N_vis,
        "preprocess_bwd": (44 + 12 * K) * N_vis + 36 * N_vis + 4 * N + (44 + 12 * K) * N,
        # p, g, m, v read for every float; m, v, p written for the active ones only
        "adam_stats": 16 * w * N + 12 * active_floats + 36 * N_vis,
        # K4b + K5 fused (single GPU): record 48 + parameters 4w + moments 8w in for every Gaussian; moments 8w +
        # parameters 4w out for the ACTIVE vectors only; radii / clamp mask / statistics ~ 24 B  (w = 11 + 3K)
        "preprocess_bwd_adam": (48 + 12 * w + 24) * N + 12 * active_floats,
        # (1-lambda) L1 + lambda D-SSIM: fwd reads image + target, writes 3 partial maps; bwd reads them + both images,
        # writes dL/dimage (12 B per pixel each)
        "loss": 12 * P * (2 + 3 + 5 + 1),
    }


def measured_issue_pct():
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            return json.load(f).get("issue_pct", {})
    except (OSError, ValueError):
        return {}


def measured_traffic(stage):
    """DRAM bytes per launch of the stage's kernel from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            v = json.load(f).get(stage)
        return float(v) if v else None
    except (OSError, ValueError):
        return None


# ---- CPU leg: the oracle port timed on the host cores ------------------------------------------
class OracleCamera:
    """A camera built on the ORACLE's restatement of the reference conventions (orc_camera_tensors,
    src/scene/camera.cpp:44-45,66-71), so that the CPU reference arm never loads libgsplat_b200.so.  Same fields
    as 3d_gaussian_splatting_b200.camera.Camera (tests/test_conventions_golden.py pins both to the same vectors)."""

    def __init__(self, lib, eye, target, up, focal_px, H, W, uid=0):
        import ctypes as C
        eye = np.asarray(eye, np.float64); target = np.asarray(target, np.float64); up = np.asarray(up, np.float64)
        fwd = target - eye
        fwd /= np.linalg.norm(fwd)
        right = np.cross(fwd, up)
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        Rw2c = np.stack([right, down, fwd], axis=0)
        t = -Rw2c @ eye
        lib.orc_focal2fov.restype = C.c_double
        lib.orc_focal2fov.argtypes = [C.c_double, C.c_double]
        fovx, fovy = lib.orc_focal2fov(float(focal_px), float(W)), lib.orc_focal2fov(float(focal_px), float(H))
        d = lambda a: np.ascontiguousarray(a, np.float64)
        R, tt, tr = d(Rw2c.T), d(t), d(np.zeros(3))
        wv, pj, fp, cp = np.zeros(16), np.zeros(16), np.zeros(16), np.zeros(3)
        P = lambda a: a.ctypes.data_as(C.c_void_p)
        lib.orc_camera_tensors(P(R), P(tt), P(tr), C.c_double(1.0), C.c_double(0.01), C.c_double(100.0),
                               C.c_double(fovx), C.c_double(fovy), P(wv), P(pj), P(fp), P(cp))
        self.uid, self.image_height, self.image_width = uid, int(H), int(W)
        self.FoVx, self.FoVy = fovx, fovy
        self.world_view_transform = wv.astype(np.float32).reshape(4, 4)
        self.full_proj_transform = fp.astype(np.float32).reshape(4, 4)
        self.camera_center = cp.astype(np.float32)
        self.tanfovx, self.tanfovy = float(np.float32(np.tan(fovx * 0.5))), float(np.float32(np.tan(fovy * 0.5)))


def load_synthetic_standalone():
    """synthetic.py loaded from its file, WITHOUT the package __init__ (which binds libgsplat_b200.so): scene
    generation is pure numpy; the cameras come from OracleCamera."""
    import importlib.util
    path = os.path.join(ROOT, "3d_gaussian_splatting_b200", "synthetic.py")
    spec = importlib.util.spec_from_file_location("_gs_synthetic_standalone", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod      # (dataclasses resolves the module of a class through sys.modules)
    spec.loader.exec_module(mod)
    return mod


class CpuStep:
    """One full train step (render, loss, backward, Adam + stats) of the CPU oracle on the same scene, stated on the
    RAW parameters exactly as the GPU path runs it (orc_activate_raw: specified exp / sigmoid, quaternion normalised
    once inside the covariance construction; activation Jacobians of src/scene/gaussian_model.cpp:54-59 applied to the
    oracle's gradients; xyz learning rate from the reference's schedule, gaussian_model.cpp:355-361).
    oracle/ is the checker; this is the one place bench.py executes it (cpu_baseline / --impl reference), and
    tests/test_gpu_raw_path.py compares Trainer.train_step with it."""

    LR = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 0.05, 5e-3, 1e-3]   # src/arguments/params.h:52-60

    def __init__(self, scene, cams, deg, views):
        from oracle.oracle_lib import Oracle, RasterSettings
        self.o = Oracle()
        self.RS = RasterSettings
        self.scene, self.cams, self.deg = scene, cams, deg
        self.raw = [scene.xyz.copy(), scene.features_dc.copy(), scene.features_rest.copy(),
                    scene.opacity_logits.copy(), scene.log_scales.copy(), scene.raw_rotations.copy()]
        self.m = [np.zeros_like(p) for p in self.raw]
        self.v = [np.zeros_like(p) for p in self.raw]
        self.lr = list(self.LR)
        N = scene.N
        self.maxr = np.zeros(N, np.int32); self.acc = np.zeros(N, np.float32); self.den = np.zeros(N, np.float32)
        self.step = 0
        self.views = views
        # all host cores, whatever OMP_NUM_THREADS the launcher exported (torchrun sets it to 1)
        want = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.o.lib.orc_set_threads.restype = C_int
        self.cores = int(self.o.lib.orc_set_threads(C_int(want)))

    def settings(self, cam):
        return self.RS(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, (0.0, 0.0, 0.0), 1.0,
                       cam.world_view_transform.reshape(-1), cam.full_proj_transform.reshape(-1), cam.camera_center,
                       self.deg, self.deg)

    def forward_backward(self, cam, dL=None, gt=None):
        """(loss or None, gradients w.r.t. the six raw parameter groups, pre, g) for one view."""
        o = self.o
        s = self.settings(cam)
        xyz, dc, rest, opl, ls, rq = self.raw
        sc, rot, op = o.activate_raw(ls, rq, opl)
        shs = np.ascontiguousarray(np.concatenate([dc, rest], axis=1))
        pre, b, fwd = o.forward(s, xyz, sc, rot, op, shs)
        loss = None
        if dL is None:
            if gt is None:
                gt = np.zeros_like(fwd["out_color"])
            loss, dL, _, _ = o.photometric_loss(fwd["out_color"], gt, LAMBDA_DSSIM)
        g = o.backward(s, xyz, sc, rot, shs, pre, b, fwd, dL)
        g_ls = (g["dL_dscales"] * sc).astype(np.float32)
        g_op = (g["dL_dopacities"].reshape(-1, 1) * (op * (1 - op))).astype(np.float32)
        grads = [g["dL_dmeans3D"], np.ascontiguousarray(g["dL_dshs"][:, :1]), np.ascontiguousarray(g["dL_dshs"][:, 1:]),
                 g_op, g_ls, np.ascontiguousarray(g["dL_drotations"], dtype=np.float32)]
        return loss, grads, pre, g

    def run(self, gt=None):
        o = self.o
        cam = self.cams[self.step % self.views]
        self.step += 1
        # update_learning_rate(iteration): position_lr_init -> position_lr_final over position_lr_max_steps
        self.lr[0] = float(o.lib.orc_expon_lr(1.6e-4, 1.6e-6, 0, 0.01, 30_000, self.step))
        loss, grads, pre, g = self.forward_backward(cam, gt=gt)
        for k in range(6):
            if self.raw[k].size:
                o.adam_step(self.raw[k], np.ascontiguousarray(grads[k], dtype=np.float32), self.m[k], self.v[k],
                            self.lr[k], 0.9, 0.999, 1e-8, self.step)
        o.densify_stats(pre["radii"], g["dL_dmeans2D"], self.maxr, self.acc, self.den)
        return loss


def cpu_scene(args, n_cameras):
    """(scene, cameras, degree, CpuStep-ready) built without touching the product library."""
    from oracle.oracle_lib import Oracle
    lib = Oracle().lib
    syn = load_synthetic_standalone()
    factory = lambda eye, target, up, focal, H, W, uid=0: OracleCamera(lib, eye, target, up, focal, H, W, uid)
    return syn.make_config(args.config, n_cameras=n_cameras, N=args.N, look_at_camera=factory)


def run_reference(args, rank, world):
    """--impl reference: the reference has no implementation of this path, so the CPU
    arm is the oracle port on all host cores, one full step of the same workload per step.  Nothing of the product
    is loaded here (no libgsplat_b200.so): scene generation is numpy, cameras come from the oracle's conventions."""
    if rank != 0:
        return
    scene, cams, deg = cpu_scene(args, max(args.views, 1))
    cpu = CpuStep(scene, cams, deg, args.views)
    steps, warm = max(1, min(args.steps, 6)), max(0, min(args.warmup, 1))
    for _ in range(warm):
        cpu.run()
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu.run()
    dt = (time.perf_counter() - t0) / steps
    val = 1.0 / dt
    cam = cams[0]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.config, "N": scene.N, "W": cam.image_width, "H": cam.image_height,
                   "sh_degree": deg, "views_per_rank_per_step": 1, "loss": LOSS_NAME,
                   "note": "reference repo has no render path; CPU arm = this repo's oracle port (OpenMP), built for "
                           "accuracy (float64 accumulators, serial per-pair gradient reduction), not for speed"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cpu.cores, "kind": "port",
                         "sample": f"{steps} full train steps of the same workload (steps capped at 6)"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


# ---- GPU arm -----------------------------------------------------------------------------------
# depth keys (+ histograms), preprocess, depth sort 4 one-sweep passes, dup block sums + scan + duplicate, tile sort 2 passes, tile
# ranges, tile order, blend_forward, ssim forward + backward + finalize, blend_backward, preprocess_backward, adam_stats
OWN_KERNELS_PER_STEP = 1 + 1 + 4 + 3 + 2 + 2 + 1 + 3 + 1 + 1 + 1   # separate K4b and K5; the fused single-GPU step: one less
LAMBDA_DSSIM = 0.2   # lambda_dssim_ default (src/arguments/params.h:62)
LOSS_NAME = "0.8*L1 + 0.2*(1-SSIM)"
LIB_KERNELS_PER_STEP = 0   # no library kernels on the path (the radix sort and scans are this repo's own)


def warm_adam_state(model):
    """Every Adam moment non-zero (m = 1e-8, v = 1e-14: the size a gradient of ~1e-7 leaves behind): the state of a
    scene in which every Gaussian has received a gradient at some point.  exp_avg_sq decays by 0.999 per step, so in a
    real training run a vector that has been touched once stays active for the rest of the run."""
    for t in model.exp_avg:
        if t is not None and t.numel():
            t.fill_(1e-8)
    for t in model.exp_avg_sq:
        if t is not None and t.numel():
            t.fill_(1e-14)


class Setup:
    pass


def setup_trainer(args, scene, my_cams, deg, rank, world, local_rank, adam_state, dp_mode):
    """Model + trainer + ground-truth images (device and pinned host) + fixed pair capacity for one workload."""
    import torch
    gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
    par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
    dev = torch.device("cuda", local_rank)
    u = Setup()
    u.dp_mode = dp_mode
    model = gm.GaussianModel(deg, local_rank).load_scene(scene)
    opt = gm.OptimizationParams()
    trainer = gm.Trainer(model, opt, bg=(0.0, 0.0, 0.0), lambda_dssim=LAMBDA_DSSIM)
    trainer.fused_step = not args.no_fused
    u.model, u.opt, u.trainer = model, opt, trainer
    u.sharded, u.gslab = None, None
    u.grad_mb = 4e-6 * sum(p.numel() for p in model.params())
    guard = None
    if world > 1 and u.dp_mode == "peer":
        # K4b stores gradient rows into the owner's slots over NVLink; K5 sums them, updates the owned shard and
        # stores the parameters into every replica; two one-float NCCL all-reduces order the ranks
        try:
            u.sharded = trainer.shard_optimizer_peer(rank, world)
            guard = u.sharded.guard
        except RuntimeError as e:       # raised on every rank alike (PeerShardedAdam): fall back to the NCCL mode
            if rank == 0:
                print(f"[bench] {e}; falling back to --dp-mode sharded", file=sys.stderr)
            u.dp_mode = "sharded"
    if world > 1 and u.dp_mode == "peer":
        pass
    elif world > 1 and u.dp_mode == "sharded":
        # gradients reduce-scattered, Adam on the 1/world slice this rank owns, parameters all-gathered
        u.sharded = trainer.shard_optimizer(rank, world)
        guard = u.sharded.guard
    else:
        # one contiguous gradient slab (a single all-reduce per step when world > 1)
        u.gslab = par.GradientSlab(model.params())
        guard = u.gslab.guard
        model.grads = u.gslab.views
    if adam_state == "warm":
        if world > 1 and u.dp_mode == "sharded":
            u.sharded.exp_avg.fill_(1e-8)
            u.sharded.exp_avg_sq.fill_(1e-14)
        else:
            warm_adam_state(model)

    # ground truth = render of the scene with SH perturbed by N(0, 0.05^2)  
    rng = np.random.Generator(np.random.Philox(99))
    gt_model = gm.GaussianModel(deg, local_rank).load_scene(scene)
    gt_model.features_dc_ += torch.from_numpy((0.05 * rng.standard_normal(scene.features_dc.shape)).astype(np.float32)).to(dev)
    u.gts_dev, u.gts_host = [], []
    max_pairs = 0
    for cam in my_cams:
        out = gm.render(cam, gt_model, gm.PipelineParams(), (0.0, 0.0, 0.0), trainer.rast)
        u.gts_dev.append(out["render"].clone())
        u.gts_host.append(out["render"].cpu().pin_memory())
        max_pairs = max(max_pairs, trainer.rast.forward_status()[0])   # (synchronised by the .cpu() above)
    del gt_model
    torch.cuda.synchronize()
    # fixed (tile, Gaussian) pair capacity: the forward then never waits for the host (include/gsplat_b200.h);
    # an overflowing iteration would be skipped on the device and is reported below
    trainer.reserve_pairs(int(1.3 * max_pairs) + 1024, guard=guard if world > 1 else None)
    if world > 1 and u.sharded is None:
        trainer.grad_sync = u.gslab.all_reduce     # ONE NCCL all-reduce of the gradient slab per iteration
    return u


def adam_activity(model, sharded=None):
    """(idle fraction, active floats): share of parameter floats whose Adam vector is at the exact fixed point
    (second moment still 0: never received a gradient).  Counted on the device."""
    if model.exp_avg_sq is None:
        ts = [sharded.exp_avg_sq] if sharded is not None and hasattr(sharded, "exp_avg_sq") else []
    else:
        ts = [t for t in model.exp_avg_sq if t is not None and t.numel()]
    if not ts:
        return None, None
    nz = sum(int((t != 0).sum().item()) for t in ts)
    tot = sum(t.numel() for t in ts)
    return round(1.0 - nz / max(1, tot), 4), nz


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
    gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
    par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    scene, cams, deg = syn.make_config(args.config, n_cameras=max(args.views * world, 1), N=args.N, variant=args.variant)
    my_cams = par.shard_cameras(cams, rank, world)[:args.views]
    H, W = my_cams[0].image_height, my_cams[0].image_width
    P, T, K, N = H * W, ((H + 15) // 16) * ((W + 15) // 16), (deg + 1) ** 2, scene.N

    u = setup_trainer(args, scene, my_cams, deg, rank, world, local_rank, args.adam_state, args.dp_mode)
    args.dp_mode = u.dp_mode
    model, opt, trainer, sharded, gslab = u.model, u.opt, u.trainer, u.sharded, u.gslab
    gts_dev, gts_host, grad_mb = u.gts_dev, u.gts_host, u.grad_mb
    g = model

    def step_device(i):
        return trainer.train_step_device(my_cams[i % args.views], gts_dev[i % args.views])

    def step_e2e(i):
        # the public call: pinned host image in, loss (python float) out
        return trainer.train_step(my_cams[i % args.views], gts_host[i % args.views])

    host_ms = [0.0]   # host time the last timed() loop needed to ENQUEUE a step (the queue is empty when it starts)

    def timed(fn, steps, warmup, sampler=None):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h0 = time.perf_counter()
        for i in range(steps):
            fn(warmup + i)
        host_ms[0] = (time.perf_counter() - h0) * 1e3 / max(steps, 1)
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), clocks

    warm = max(args.warmup, 3)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    total_ms, clocks = timed(step_device, args.steps, warm, sampler)
    host_enqueue_ms = host_ms[0]
    peer_phases = None
    if world > 1 and args.dp_mode == "peer" and getattr(sharded, "split", False):
        # where the split-phase exchange sits inside a pipelined step: a second, instrumented pass (events only)
        sharded.profile_phases = True
        trainer.rast.set_profiling(True)
        timed(step_device, args.steps, 1)
        sharded.profile_phases = False
        peer_phases = sharded.phase_ms()
        # the library's per-stage events of the LAST pipelined iteration: stage durations as they are while the SH
        # phase of the previous step is still running beside them
        peer_phases["stages_while_overlapped"] = {k: round(v, 4) for k, v in trainer.rast.stage_ms().items()}
        trainer.rast.set_profiling(False)
    assert not trainer.rast.forward_status()[1], "pair capacity overflow inside the timed region"
    ms_per_step = total_ms / args.steps
    value = world * 1.0 / (ms_per_step * 1e-3)
    # end to end through the public training loop (Trainer.train, the reference's train() loop shape): pinned HOST
    # ground truth copied in and the loss copied out EVERY iteration; the loop is a depth-2 software pipeline, so
    # the host reads loss i while the device already works on iteration i + 1.  One call = args.steps iterations.
    e2e_losses = []

    def loop_e2e(first, n):
        e2e_losses[:] = trainer.train(my_cams[:args.views], gts_host[:args.views], n, first=first)

    if args.e2e_api == "native" and world == 1:
        # the loop in C++ (csrc/trainer.cu): its own model copy, its own streams; timed by the host clock around the
        # call, which returns only after the last iteration's loss has arrived
        nt = gm.NativeTrainer(scene, opt, device=local_rank, lambda_dssim=LAMBDA_DSSIM)
        for cam, gt in zip(my_cams[:args.views], gts_host[:args.views]):
            nt.add_view(cam, gt)
        nt.run(warm, cyclic=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e2e_losses[:] = nt.run(args.steps, cyclic=True)
        e2e_ms = (time.perf_counter() - t0) * 1e3
        assert all(np.isfinite(l) for l in e2e_losses) and nt.retries == 0
        nt.close()
    elif args.e2e_api in ("loop", "native"):
        # timed() runs fn(0) as its one warm-up call (`warm` iterations), then fn(1) timed (args.steps iterations)
        e2e_ms, _ = timed(lambda i: loop_e2e(0, warm) if i == 0 else loop_e2e(warm, args.steps), 1, 1)
        assert len(e2e_losses) == args.steps and all(np.isfinite(l) for l in e2e_losses), "e2e loop lost a loss"
    else:
        e2e_ms, _ = timed(step_e2e, args.steps, warm)
    e2e_value = world * 1.0 / (e2e_ms / args.steps * 1e-3)

    # ---- per-stage device times (CUDA events recorded by the library around each stage) ----
    def profile_stages(tr, gts):
        """Median stage times of 6 steady iterations + workload counters of the last view."""
        m = tr.g
        tr.rast.set_profiling(True)
        acc = {}
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        extra = {"loss": [], "adam_stats": []}
        fused = world == 1 and tr.fused_step
        for i in range(8):
            cam, gt = my_cams[i % args.views], gts[i % args.views]
            out = gm.render(cam, m, tr.pipe, tr.bg, tr.rast)
            ev[0].record(); loss, dL = tr.rast.loss(out["render"], gt, LAMBDA_DSSIM); ev[1].record()
            if fused:     # K4b and K5 are one kernel: its time is the library's stage 7
                m.step_count = [s + 1 for s in m.step_count]
                tr.rast.backward_adam_step(dL, m.exp_avg, m.exp_avg_sq, m.lr, m.step_count,
                                           max_radii2D=m.max_radii2D_, xyz_grad_accum=m.xyz_gradient_accum_,
                                           denom=m.denom_, skip_guard=tr.skip_guard)
            else:
                tr.rast.backward(dL, grads=m.grad_sinks())
                if tr.grad_sync is not None:      # (all-reduce mode: the replicas must keep receiving the SUMMED gradient;
                    tr.grad_sync()                #  its time lands in neither the backward's nor the Adam stage)
            ev[2].record()
            if not fused:
                tr.optimizer_step(out["radii"])
            ev[3].record()
            torch.cuda.synchronize()
            if i >= 2:
                for k, v in tr.rast.stage_ms().items():
                    acc.setdefault(k, []).append(v)
                extra["loss"].append(ev[0].elapsed_time(ev[1]))
                if not fused:
                    extra["adam_stats"].append(ev[2].elapsed_time(ev[3]))
        tr.rast.set_profiling(False)
        st = {k: float(np.median(v)) for k, v in acc.items()}
        if st.get("sh_color", 0.0) <= 0.0:      # (only the split-phase multi-GPU step defers the colour evaluation)
            st.pop("sh_color", None)
        st.update({k: float(np.median(v)) for k, v in extra.items() if v})
        if fused:
            st["preprocess_bwd_adam"] = st.pop("preprocess_bwd")
        n_vis = int((out["radii"] > 0).sum().item())
        d_pairs = int(tr.rast.debug_get("NUM_RENDERED")[0])
        ncon = torch.from_numpy(tr.rast.debug_get("N_CONTRIB").astype(np.int64)).reshape(H, W)
        pad = torch.zeros(((H + 15) // 16 * 16, (W + 15) // 16 * 16), dtype=torch.int64)
        pad[:H, :W] = ncon
        d_eff = int(pad.reshape(pad.shape[0] // 16, 16, pad.shape[1] // 16, 16).amax(dim=(1, 3)).sum().item())
        return st, n_vis, d_pairs, d_eff, int(ncon.sum().item()), fused

    stage_ms, N_vis, D, D_eff, interactions, fused = profile_stages(trainer, gts_dev)
    adam_idle_frac, active_floats = adam_activity(g, sharded)

    # ---- sustained: one continuous run of >= sustain_seconds (the 20-step region above lasts ~30 ms; this one says
    # what the clocks do when the step runs for seconds).  The headline model TRAINS during such a run: at the
    # reference's learning rates 1500 Adam steps grow the synthetic scene's splats, D rises past the fixed pair
    # capacity and every later step is voided by the overflow guard (Trainer.train would grow the capacity; this
    # bare loop does not) -- that run measures a drifting workload, not the clocks.  So the figure reported is taken
    # with every learning rate held at 0: the same kernels, instructions and HBM traffic per step (the Adam kernel
    # still reads and writes every moment and parameter; p <- p - 0 * m/(sqrt(v) + eps)), the scene and with it D /
    # D_eff stay the headline's.  The live-training run is kept beside it with its final D and capacity status. ----
    sustained = None
    if args.sustain_seconds > 0 and not args.no_extras:
        n_sus = max(args.steps, int(args.sustain_seconds * 1e3 / ms_per_step) + 1)

        def sustained_run(frozen):
            marks, host_t = [], [0.0, 0.0]
            saved_lr, saved_sched = list(g.lr), g.xyz_scheduler_args
            if frozen:
                g.lr = [0.0] * len(g.lr)
                g.xyz_scheduler_args = lambda it: 0.0

            def step_marked(i):
                if i == 0:
                    host_t[0] = time.perf_counter()
                if i % 100 == 0:
                    e = torch.cuda.Event(enable_timing=True)
                    e.record()
                    marks.append(e)
                step_device(i)
                host_t[1] = time.perf_counter()

            sus_ms, sus_clocks = timed(step_marked, n_sus, 0, ClockSampler(local_rank) if rank == 0 else None)
            g.lr, g.xyz_scheduler_args = saved_lr, saved_sched
            d_end, ovf = trainer.rast.forward_status()
            per100 = [round(a.elapsed_time(b) / 100.0, 4) for a, b in zip(marks, marks[1:])]
            return {"ms_per_step": sus_ms / n_sus, "value": world * 1e3 / (sus_ms / n_sus), "unit": UNIT,
                    "steps": n_sus, "seconds": round(sus_ms * 1e-3, 3), "clocks": sus_clocks,
                    # device ms per step over each block of 100 steps, and the HOST time the Python loop needed to
                    # enqueue a step (when it reaches the device time the loop, not the GPU, sets the pace)
                    "ms_per_step_by_100": per100,
                    "host_enqueue_ms_per_step": round((host_t[1] - host_t[0]) * 1e3 / n_sus, 4),
                    "D_end": int(d_end), "capacity_ok": not ovf}

        sustained = sustained_run(True)
        sustained["workload"] = ("frozen: all learning rates 0 during this run (identical kernels and traffic per step; the "
                                 "scene, D and D_eff stay the headline's)")

    # ---- secondary workloads (single GPU): the same step with (a) cold Adam state -- the synthetic ball hides ~70 %
    # of its Gaussians from every view, their vectors stay at the exact fixed point and are skipped: the best case;
    # (b) the stress variant of the scene (faint Gaussians: long blend lists, little early termination) ----
    def secondary(adam_state, variant):
        sc2 = scene if variant == args.variant else syn.make_config(args.config, n_cameras=max(args.views, 1), N=args.N,
                                                                    variant=variant)[0]
        u2 = setup_trainer(args, sc2, my_cams, deg, rank, world, local_rank, adam_state, args.dp_mode)
        t2 = u2.trainer
        ms2, _ = timed(lambda i: t2.train_step_device(my_cams[i % args.views], u2.gts_dev[i % args.views]),
                       args.steps, warm)
        ok = not t2.rast.forward_status()[1]
        st2, nv2, d2, de2, ia2, _ = profile_stages(t2, u2.gts_dev)
        idle2, act2 = adam_activity(u2.model, u2.sharded)
        by2 = stage_bytes(N, nv2, d2, de2, P, T, K, act2)
        dom2 = "preprocess_bwd_adam" if "preprocess_bwd_adam" in st2 else max(st2, key=lambda k: st2[k])
        pk = peaks()[0]
        res = {"ms_per_step": ms2 / args.steps, "value": 1e3 / (ms2 / args.steps), "unit": UNIT,
               "adam_idle_frac": idle2, "D": d2, "D_eff": de2, "pixel_gaussian_interactions": ia2,
               "capacity_ok": ok,
               "stages_ms": {k: round(v, 4) for k, v in st2.items()},
               "roofline_frac": {dom2: round(by2[dom2] / (st2[dom2] * 1e-3) / 1e9 / pk, 4)}}
        t2.rast.close()
        del u2, t2
        torch.cuda.empty_cache()
        return res

    best_case = stress = None
    if world == 1 and not args.no_extras:
        if args.adam_state == "warm":
            best_case = secondary("cold", args.variant)
        if args.variant == "default":
            stress = secondary(args.adam_state, "stress")

    # ---- north_star's multi-view step (BASELINE.json configs[4]): 6 M Gaussians, ONE optimiser step per 64-view
    # batch split over the GPUs (strong scaling), gradients accumulated per rank, one exchange per step ----
    north = None
    if not args.no_extras and args.variant == "default":
        try:
            north = measure_batch(args, rank, world, local_rank, "cfg4_6m_1246x825_sh3", 64, 2, 1, args.dp_mode)
        except Exception as e:      # (reported, never fatal for the headline)
            north = {"error": f"{type(e).__name__}: {e}"}

    by = stage_bytes(N, N_vis, D, D_eff, P, T, K, active_floats if world == 1 else None)
    if "sh_color" in stage_ms:
        by["preprocess"] -= 12 * K * N
    if world > 1 and args.dp_mode == "peer" and active_floats is not None:
        # adam_peer_kernel, local HBM bytes of one rank: the `world` gradient slots + m, v, p of its R-row shard read;
        # m, v, p of the active vectors written; plus what the other ranks' kernels store into this replica
        wfl, R = 11 + 3 * K, sharded.R
        share = 1.0
        if getattr(sharded, "split", False):
            # split-phase step: the stage timed on the caller's stream is the GEOMETRY phase (11 of the 11 + 3 K floats
            # of a Gaussian); the SH phase runs beside the next forward and is reported in peer_phases_ms
            share = 11.0 / wfl
            wfl = 11
        act = active_floats * share
        by["adam_stats"] = int((4 * world + 12) * wfl * R + 12 * act + 4 * act * (world - 1) + 36 * N_vis)
    elif world > 1 and args.dp_mode == "sharded":
        wfl = 11 + 3 * K
        by["adam_stats"] = 28 * wfl * N // world + 36 * N_vis
    peak, peak_kind = peaks()
    # the roofline object is for the dominant HBM-bound stage (the blend kernels are bound by issue slots and dependency
    # latency by construction -- their share of the HBM peak says nothing; `stages` carries it for completeness)
    hbm_bound = [k for k in stage_ms if k in ("preprocess_bwd_adam", "adam_stats", "preprocess_bwd", "preprocess",
                                              "sh_color", "loss")]
    dom = max(hbm_bound or stage_ms, key=lambda k: stage_ms[k])
    ach = by[dom] / (stage_ms[dom] * 1e-3) / 1e9
    by = {k: v for k, v in by.items() if k in stage_ms}
    step_bytes = sum(by.values())
    stages = {k: {"ms": round(stage_ms[k], 4), "alg_GB": round(by[k] / 1e9, 4),
                  "GBps": round(by[k] / (stage_ms[k] * 1e-3) / 1e9, 1) if stage_ms[k] > 0 else None,
                  "frac_hbm": round(by[k] / (stage_ms[k] * 1e-3) / 1e9 / peak, 4) if stage_ms[k] > 0 else None}
              for k in stage_ms}
    for k, v in measured_issue_pct().items():      # issue-slot utilisation from the committed ncu capture
        if k in stages:
            stages[k]["issue_pct_ncu"] = v
    # (the fused single-GPU step has no separate K4b: its K4b+K5 kernel is counted whole)
    # the depth argsort runs beside K1 on the context's second stream (capi.cu): the two stages overlap, the critical
    # path holds the longer one (GS_NO_SORT_OVERLAP=1 serialises them again)
    sort_overlapped = os.environ.get("GS_NO_SORT_OVERLAP", "0") != "1"
    head_ms = (max(stage_ms.get("preprocess", 0.0), stage_ms.get("depth_sort", 0.0)) if sort_overlapped
               else stage_ms.get("preprocess", 0.0) + stage_ms.get("depth_sort", 0.0))
    fwd_bwd_ms = head_ms + sum(stage_ms.get(k, 0.0) for k in ("scan_duplicate", "sort", "ranges", "blend_fwd",
                                                              "blend_bwd", "preprocess_bwd", "preprocess_bwd_adam"))
    for k in ("preprocess", "depth_sort"):
        if sort_overlapped and k in stages:
            stages[k]["concurrent_with"] = "depth_sort" if k == "preprocess" else "preprocess"

    # ---- the rasterizer alone (BASELINE.json "fwd+bwd raster ms"): gs_forward + gs_backward, no loss, no optimiser;
    # dL/dimage = seeded U(-1,1)/P, loss-independent.  Single GPU only (the multi-GPU modes route
    # the backward's stores to the owner ranks).
    raster = None
    if world == 1:
        rs = np.random.Generator(np.random.Philox(4321))
        dL_seed = torch.from_numpy(((2.0 * rs.random((3, H, W)) - 1.0) / P).astype(np.float32)).to(dev)
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(12)]
        for i, e in enumerate(evs):
            cam = my_cams[i % args.views]
            e[0].record()
            gm.render(cam, g, trainer.pipe, trainer.bg, trainer.rast, with_visibility=False)
            e[1].record()
            trainer.rast.backward(dL_seed, grads=g.grad_sinks())
            e[2].record()
        torch.cuda.synchronize()
        f_ms = float(np.median([e[0].elapsed_time(e[1]) for e in evs[3:]]))
        b_ms = float(np.median([e[1].elapsed_time(e[2]) for e in evs[3:]]))
        fb_ms = float(np.median([e[0].elapsed_time(e[2]) for e in evs[3:]]))
        rb = sum(by.get(k, 0) for k in ("preprocess", "depth_sort", "scan_duplicate", "sort", "ranges", "blend_fwd",
                                        "blend_bwd")) + stage_bytes(N, N_vis, D, D_eff, P, T, K)["preprocess_bwd"]
        raster = {"fwd_ms": round(f_ms, 4), "bwd_ms": round(b_ms, 4), "fwd_bwd_ms": round(fb_ms, 4),
                  "alg_GB": round(rb / 1e9, 4), "frac_hbm": round(rb / (fb_ms * 1e-3) / 1e9 / peak, 4),
                  "what": "gs_forward + gs_backward (K1-K3, K4a, K4b with the six gradient tensors written to HBM), "
                          "dL/dimage = seeded U(-1,1)/P, median of 9 after 3 warm-ups, CUDA events"}

    # (last use of the headline model on one GPU: the live-training companion of `sustained` lets it drift)
    if sustained is not None and world == 1:
        live = sustained_run(False)
        sustained["live_training"] = {k: live[k] for k in ("ms_per_step", "ms_per_step_by_100", "D_end", "capacity_ok")}
        sustained["live_training"]["note"] = ("the model trains at the reference's learning rates: the workload drifts (D "
                                              "grows); capacity_ok false = D passed the fixed pair capacity and the later "
                                              "steps were voided by the guard")

    # ---- N > 1: the replicas must be bit-identical after the timed loops (every rank holds the same model) ----
    replicas_identical = last_loss = barrier_timeouts = None
    if world > 1:
        torch.cuda.synchronize()
        cs = torch.zeros(1, dtype=torch.int64, device=dev)
        for t in g.params():
            if t.numel():
                cs += t.contiguous().view(torch.int32).to(torch.int64).sum()
        allcs = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(allcs, cs)
        replicas_identical = all(int(c.item()) == int(allcs[0].item()) for c in allcs)
        last_loss = float(e2e_losses[-1]) if e2e_losses else None
        if args.dp_mode == "peer":
            barrier_timeouts = sharded.barrier_timeouts()
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = CpuStep(scene, cams, deg, args.views)
        gt_np = gts_host[0].numpy()
        cpu.run(gt_np)  # warm-up (page faults, OpenMP pool)
        t0 = time.perf_counter()
        nrep = 2
        for _ in range(nrep):
            cpu.run(gt_np)
        dt = (time.perf_counter() - t0) / nrep
        cpu_baseline = {"value": 1.0 / dt, "unit": UNIT, "cores": cpu.cores, "kind": "port",
                        "sample": f"{nrep} full train steps of the same workload after 1 warm-up ({dt:.2f} s/step)"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.config, "N": N, "W": W, "H": H, "sh_degree": deg, "tiles": T,
                       "views_per_rank_per_step": 1, "loss": LOSS_NAME, "variant": args.variant,
                       "adam_state": ("warm: every Adam moment non-zero before the timed region, so no vector is at the "
                                      "kernels' exact skip point (steady state of a trained scene)" if args.adam_state == "warm"
                                      else "cold: moments start at zero; never-touched vectors are skipped"),
                       "N_vis": N_vis, "D": D, "D_eff": D_eff,
                       "pixel_gaussian_interactions": interactions,
                       "l2_policy": "inputs larger than L2 (params+Adam state 1.13 GB per step vs 126 MB L2)",
                       "parallelism": (f"camera-sharded dp{world}, {grad_mb:.0f} MB of gradients per rank: "
                                       + {"peer": "K4b stores them into the owner ranks' slots over NVLink, K5 sums the "
                                                  "slots, updates the owned 1/%d of the rows and stores the parameters "
                                                  "into every replica (2 flag barriers per step: gs_peer_barrier, one 32-thread "
                                                  "kernel each, no collective)" % world,
                                          "sharded": "NCCL reduce-scatter -> Adam on the owned 1/%d slice -> NCCL "
                                                     "all-gather" % world,
                                          "allreduce": "one NCCL all-reduce, replicated Adam"}[args.dp_mode])
                       if world > 1 else "single GPU"},
            # share of parameter floats whose Adam moments are still exactly 0 (never received a gradient): the Adam
            # kernels skip those vectors (an exact fixed point of the update), so K5's time depends on it
            "adam_idle_frac": adam_idle_frac,
            # host time per step the Python loop spent enqueuing the timed region (below ms_per_step: the device sets
            # the pace; at or above it: the host does)
            "host_enqueue_ms_per_step": round(host_enqueue_ms, 4),
            # the same step over >= sustain_seconds of continuous running, with its own clock record
            "sustained": sustained,
            # the same step with cold Adam state (this scene hides ~70 % of its Gaussians from every view: their
            # vectors are skipped) and on the stress variant of the scene (long blend lists)
            "best_case": best_case, "stress": stress,
            # north_star's 64-view batch on the 6 M-Gaussian scene: views/s, exchange time reported separately
            "north_star_batch": north,
            "replicas_identical": replicas_identical, "last_loss": last_loss, "barrier_timeouts": barrier_timeouts,
            # split-phase exchange (N > 1, peer mode): device ms from the end of barrier 1 to the end of each part
            "peer_phases_ms": peer_phases,
            "raster": raster,
            "raster_fwd_bwd_ms": fwd_bwd_ms, "raster_fwd_bwd_includes_adam": bool(fused), "train_it_per_s_per_gpu": 1e3 / ms_per_step,
            "stages": stages,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak,
                         # (the committed ncu captures are single-GPU: the peer kernels of N > 1 have none)
                         "traffic": measured_traffic(dom) if world == 1 or dom not in ("adam_stats", "preprocess_bwd") else None,
                         "peak_kind": peak_kind,
                         "step_frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak},
            "cpu_baseline": cpu_baseline,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": int(gts_host[0].numel() * 4 + 172),
                    "d2h_bytes_per_step": 4 if args.e2e_api == "step" else 8,
                    "api": {"loop": "Trainer.train: depth-2 pipelined loop, ground truth H2D + loss/overflow flag D2H every "
                                    "iteration",
                            "native": "gs_trainer_run (C++ loop over the C ABI): depth-2 pipeline, ground truth H2D + "
                                      "loss/overflow flag D2H every iteration; host-clock timed",
                            "step": "Trainer.train_step: one synchronous call per step"}[
                                args.e2e_api if world == 1 or args.e2e_api != "native" else "loop"]},
            "pair_capacity": trainer.pair_capacity, "capacity_retries": trainer.retries,
            # own kernels only (NCCL's are not counted).  Peer mode: the routed K4b is ONE launch for all owners; two
            # flag-barrier kernels per step; the split-phase step adds the colour kernel, a second Adam launch and a
            # third barrier kernel
            "gpu_launches": (OWN_KERNELS_PER_STEP - (1 if world == 1 and trainer.fused_step else 0)
                             + ((2 if getattr(sharded, "flag_barriers", False) else 0)
                                + (3 if getattr(sharded, "split", False) else 0)
                                if world > 1 and args.dp_mode == "peer" else 0)
                             + (1 if world > 1 and args.dp_mode == "sharded" else 0)) * args.steps,
            "library_launches": LIB_KERNELS_PER_STEP * args.steps,
            "clocks": clocks,
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure_batch(args, rank, world, local_rank, config, V, steps, warm, dp_mode):
    """north_star's multi-view step on `config`: V views per optimiser step, V/world per rank.  Returns the fields of
    the result (every rank must call; values are maxima over ranks)."""
    import torch
    import torch.distributed as dist
    syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
    gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
    par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
    dev = torch.device("cuda", local_rank)
    assert V % world == 0, "--batch-views must be a multiple of the GPU count"
    scene, cams, deg = syn.make_config(config, n_cameras=V, N=args.N)
    my_cams = par.shard_cameras(cams, rank, world)          # V / world views
    H, W = my_cams[0].image_height, my_cams[0].image_width
    model = gm.GaussianModel(deg, local_rank).load_scene(scene)
    trainer = gm.Trainer(model, gm.OptimizationParams(), bg=(0.0, 0.0, 0.0), lambda_dssim=LAMBDA_DSSIM)
    mode = dp_mode if world > 1 else "single"
    gslab = None
    if mode == "peer":
        try:
            peer = trainer.shard_optimizer_peer(rank, world, pull=True)
            guard = peer.guard
        except RuntimeError as e:
            if rank == 0:
                print(f"[bench] {e}; batch mode falls back to --dp-mode allreduce", file=sys.stderr)
            mode = "allreduce"
    if mode != "peer":
        gslab = par.GradientSlab(model.params())
        model.grads = gslab.views
        guard = gslab.guard
        if world > 1:
            trainer.grad_sync = gslab.all_reduce
    warm_adam_state(model)
    rng = np.random.Generator(np.random.Philox(99))
    gt_model = gm.GaussianModel(deg, local_rank).load_scene(scene)
    gt_model.features_dc_ += torch.from_numpy((0.05 * rng.standard_normal(scene.features_dc.shape)).astype(np.float32)).to(dev)
    gts_dev, gts_host, max_pairs = [], [], 0
    for cam in my_cams:
        out = gm.render(cam, gt_model, gm.PipelineParams(), (0.0, 0.0, 0.0), trainer.rast)
        gts_dev.append(out["render"].clone())
        gts_host.append(out["render"].cpu().pin_memory())
        max_pairs = max(max_pairs, trainer.rast.forward_status()[0])
    del gt_model
    torch.cuda.synchronize()
    trainer.reserve_pairs(int(1.3 * max_pairs) + 1024, guard=guard)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ex_ms = []

    def step(timed):
        loss = trainer.train_batch_device(my_cams, gts_dev, exchange_events=ev if timed else None)
        return loss

    def run(n, host=False):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for _ in range(n):
            if host:      # end to end: every view's ground truth comes from pinned host memory, the loss goes back
                last = trainer.train_batch_device(my_cams, [trainer._gt_dev] * len(my_cams), gts_host=gts_host)
                trainer._loss_host.copy_(last, non_blocking=True)
            else:
                last = step(True)
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / n, last

    trainer._gt_dev = torch.empty_like(gts_dev[0])
    for _ in range(warm):
        step(False)
    ms_step, last = run(steps)
    torch.cuda.synchronize()
    ex = torch.tensor([ev[0].elapsed_time(ev[1])], device=dev, dtype=torch.float64)      # (the last step's)
    if world > 1:
        dist.all_reduce(ex, op=dist.ReduceOp.MAX)
    e2e_ms, last_e2e = run(max(1, steps), host=True)
    torch.cuda.synchronize()
    ok = float(guard.item()) == 0.0
    cs = torch.zeros(1, dtype=torch.int64, device=dev)
    for t in model.params():
        if t.numel():
            cs += t.contiguous().view(torch.int32).to(torch.int64).sum()
    identical = None
    if world > 1:
        allcs = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(allcs, cs)
        identical = all(int(c.item()) == int(allcs[0].item()) for c in allcs)
    res = {"workload": config, "N": scene.N, "W": W, "H": H, "sh_degree": deg, "batch_views": V,
           "views_per_rank": len(my_cams), "steps": steps, "warmup": warm,
           "ms_per_step": ms_step, "views_per_s": V / (ms_step * 1e-3),
           "exchange_ms": float(ex.item()),
           "exchange": {"single": "none (one GPU): Adam + statistics only",
                        "peer": "pull over NVLink peer memory: the owner's Adam kernel reads its rows of every rank's "
                                "gradient slab, updates its 1/%d of the model and stores the parameters into every "
                                "replica (2 flag barriers per step: gs_peer_barrier)" % world,
                        "allreduce": "one NCCL all-reduce of the gradient slab, replicated Adam"}[mode],
           "exchange_bytes_per_rank": (0 if world == 1 else int(2 * (world - 1) / world * 4 * sum(p.numel() for p in model.params()))),
           "e2e_ms_per_step": e2e_ms, "e2e_views_per_s": V / (e2e_ms * 1e-3),
           "h2d_bytes_per_step": int(len(my_cams) * gts_host[0].numel() * 4), "d2h_bytes_per_step": 4,
           "capacity_ok": ok, "replicas_identical": identical, "last_loss": float(last.item()),
           "adam_state": "warm", "scaling": "strong"}
    if mode == "peer":
        res["barrier_timeouts"] = trainer.peer.barrier_timeouts()
        trainer.peer.detach_model()
        trainer.peer.close()
    trainer.rast.close()
    del trainer, model, gts_dev, gts_host
    torch.cuda.empty_cache()
    return res


def run_batch(args, rank, world, local_rank):
    """--batch-views V: the multi-view step as the headline of its own line (strong scaling over the GPU count)."""
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    config = args.config if args.config != DEFAULT_CONFIG else "cfg4_6m_1246x825_sh3"
    warm = max(args.warmup, 3)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    r = measure_batch(args, rank, world, local_rank, config, args.batch_views, args.steps, warm, args.dp_mode)
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": r["views_per_s"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {k: r[k] for k in ("workload", "N", "W", "H", "sh_degree", "batch_views", "views_per_rank", "adam_state",
                                         "exchange")},
            "exchange_ms": r["exchange_ms"], "exchange_bytes_per_rank": r["exchange_bytes_per_rank"],
            "e2e": {"value": r["e2e_views_per_s"], "unit": UNIT, "ms_per_step": r["e2e_ms_per_step"],
                    "h2d_bytes_per_step": r["h2d_bytes_per_step"], "d2h_bytes_per_step": r["d2h_bytes_per_step"],
                    "api": "Trainer.train_batch_device with pinned host ground truth per view, loss D2H per step"},
            "capacity_ok": r["capacity_ok"], "replicas_identical": r["replicas_identical"], "last_loss": r["last_loss"],
            "clocks": clocks}), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_raster(args, local_rank):
    """--mode raster: ONE forward + backward of the render path (BASELINE.json configs[0]: 10 k Gaussians, one 256x256
    camera, SH degree 0 -- the mandated CPU measurement), the CUDA path beside the CPU oracle on the host cores, same
    seeded inputs, dL/dimage = seeded U(-1,1)/P; medians of >= 5 repetitions after warm-up."""
    import torch
    pkg = importlib.import_module("3d_gaussian_splatting_b200")
    syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
    torch.cuda.set_device(local_rank)
    scene, cams, deg = syn.make_config(args.config, n_cameras=1, N=args.N)
    cam = cams[0]
    H, W = cam.image_height, cam.image_width
    P = H * W
    rs = np.random.Generator(np.random.Philox(4321))
    dL = ((2.0 * rs.random((3, H, W)) - 1.0) / P).astype(np.float32)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()
    Pm = [d(scene.xyz), d(scene.features_dc), d(scene.features_rest), d(scene.opacity_logits), d(scene.log_scales),
          d(scene.raw_rotations)]
    rast = pkg.Rasterizer(local_rank)
    sg = pkg.GaussianRasterizationSettings.from_camera(cam, (0.0, 0.0, 0.0), deg, deg)
    dLd = d(dL)
    reps = max(args.steps, 9)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(reps + 3)]
    for e in evs:
        e[0].record()
        rast.forward(sg, Pm[0], Pm[4], Pm[5], Pm[3], sh_dc=Pm[1], sh_rest=Pm[2] if Pm[2].numel() else None, raw_params=True)
        e[1].record()
        rast.backward(dLd)
        e[2].record()
    torch.cuda.synchronize()
    gf = float(np.median([e[0].elapsed_time(e[1]) for e in evs[3:]]))
    gb = float(np.median([e[1].elapsed_time(e[2]) for e in evs[3:]]))
    gfb = float(np.median([e[0].elapsed_time(e[2]) for e in evs[3:]]))
    D = int(rast.debug_get("NUM_RENDERED")[0])
    cpu = CpuStep(scene, cams, deg, 1)
    cf, cb = [], []
    for i in range(1 + max(5, min(args.steps, 9))):
        t0 = time.perf_counter()
        s = cpu.settings(cam)
        xyz, dc, rest, opl, ls, rq = cpu.raw
        sc, rot, op = cpu.o.activate_raw(ls, rq, opl)
        shs = np.ascontiguousarray(np.concatenate([dc, rest], axis=1))
        pre, b, fwd = cpu.o.forward(s, xyz, sc, rot, op, shs)
        t1 = time.perf_counter()
        cpu.o.backward(s, xyz, sc, rot, shs, pre, b, fwd, dL)
        t2 = time.perf_counter()
        if i:
            cf.append((t1 - t0) * 1e3); cb.append((t2 - t1) * 1e3)
    print(json.dumps({
        "metric": "raster_fwd_bwd_ms", "value": gfb, "unit": "ms", "n_gpus": 1, "steps": reps, "warmup": 3,
        "higher_is_better": False, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.config, "mode": "one forward + backward of the render path", "N": scene.N, "W": W,
                   "H": H, "sh_degree": deg, "D": D},
        "gpu": {"fwd_ms": round(gf, 4), "bwd_ms": round(gb, 4), "fwd_bwd_ms": round(gfb, 4),
                "what": "gs_forward + gs_backward through the C ABI (raw parameters), CUDA events, median"},
        "cpu_baseline": {"fwd_ms": round(float(np.median(cf)), 3), "bwd_ms": round(float(np.median(cb)), 3),
                         "value": round(float(np.median(cf) + np.median(cb)), 3), "unit": "ms", "cores": cpu.cores,
                         "kind": "port", "sample": f"median of {len(cf)} forward + backward passes of the same inputs "
                                                   "(oracle port, OpenMP; built for accuracy, not speed)"},
    }), flush=True)


def run_render_sweep(args, local_rank):
    """--mode render: forward-only render throughput over the config's cameras (BASELINE.json configs[3]: 4 M
    Gaussians at 1920x1080), Mpix/s with the per-stage breakdown.  One JSON line; not the headline metric."""
    import torch
    syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
    gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
    torch.cuda.set_device(local_rank)
    scene, cams, deg = syn.make_config(args.config, N=args.N)
    cams = cams[:max(1, args.views)]
    H, W = cams[0].image_height, cams[0].image_width
    P, T, K, N = H * W, ((H + 15) // 16) * ((W + 15) // 16), (deg + 1) ** 2, scene.N
    model = gm.GaussianModel(deg, local_rank).load_scene(scene)
    rast = importlib.import_module("3d_gaussian_splatting_b200.rasterizer").Rasterizer(local_rank)
    pipe, bg = gm.PipelineParams(), (0.0, 0.0, 0.0)
    max_pairs, per_view = 0, []
    for cam in cams:      # synchronous mode: exact pair counts per view
        out = gm.render(cam, model, pipe, bg, rast, with_visibility=False)
        torch.cuda.synchronize()
        Dv = rast.forward_status()[0]
        max_pairs = max(max_pairs, Dv)
        per_view.append(Dv)
    rast.set_duplicate_capacity(int(1.2 * max_pairs) + 1024)      # from here on no host round-trip
    warm = max(args.warmup, 3)
    for i in range(warm):
        gm.render(cams[i % len(cams)], model, pipe, bg, rast, with_visibility=False)
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    n = max(args.steps, len(cams))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        gm.render(cams[i % len(cams)], model, pipe, bg, rast, with_visibility=False)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    assert not rast.forward_status()[1], "pair capacity overflow inside the timed region"
    ms = e0.elapsed_time(e1) / n
    rast.set_profiling(True)
    acc = {}
    for i in range(len(cams)):
        out = gm.render(cams[i], model, pipe, bg, rast, with_visibility=False)
        torch.cuda.synchronize()
        for k, v in rast.stage_ms().items():
            if v > 0:
                acc.setdefault(k, []).append(v)
    rast.set_profiling(False)
    stage_ms = {k: float(np.median(v)) for k, v in acc.items()}
    N_vis = int((out["radii"] > 0).sum().item())
    D = int(rast.debug_get("NUM_RENDERED")[0])
    ncon = torch.from_numpy(rast.debug_get("N_CONTRIB").astype(np.int64)).reshape(H, W)
    pad = torch.zeros(((H + 15) // 16 * 16, (W + 15) // 16 * 16), dtype=torch.int64)
    pad[:H, :W] = ncon
    D_eff = int(pad.reshape(pad.shape[0] // 16, 16, pad.shape[1] // 16, 16).amax(dim=(1, 3)).sum().item())
    by = stage_bytes(N, N_vis, D, D_eff, P, T, K)
    peak, peak_kind = peaks()
    fwd_bytes = sum(by[k] for k in ("preprocess", "depth_sort", "scan_duplicate", "sort", "ranges", "blend_fwd"))
    binning = sum(stage_ms.get(k, 0.0) for k in ("depth_sort", "scan_duplicate", "sort", "ranges"))
    print(json.dumps({
        "metric": "render_mpix_per_s", "value": P / (ms * 1e-3) / 1e6, "unit": "Mpix/s", "n_gpus": 1, "steps": n,
        "warmup": warm, "ms_per_step": ms, "frames_per_s": 1e3 / ms, "higher_is_better": True, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": args.config, "mode": "forward-only render", "N": N, "W": W, "H": H, "sh_degree": deg,
                   "tiles": T, "views": len(cams), "N_vis": N_vis, "D_last_view": D, "D_max": max_pairs,
                   "D_mean": float(np.mean(per_view)), "D_eff_last_view": D_eff,
                   "pixel_gaussian_interactions": int(ncon.sum().item())},
        "breakdown_ms": {"preprocess": round(stage_ms.get("preprocess", 0.0), 4), "binning_total": round(binning, 4),
                         "depth_sort": round(stage_ms.get("depth_sort", 0.0), 4),
                         "scan_duplicate": round(stage_ms.get("scan_duplicate", 0.0), 4),
                         "tile_sort": round(stage_ms.get("sort", 0.0), 4), "ranges": round(stage_ms.get("ranges", 0.0), 4),
                         "blend_fwd": round(stage_ms.get("blend_fwd", 0.0), 4),
                         "note": "depth_sort runs beside preprocess on a second stream (not additive)"},
        "roofline": {"bound": "hbm", "kernel": "forward (K1 + K2 + K3)", "achieved": fwd_bytes / (ms * 1e-3) / 1e9,
                     "peak": peak, "unit": "GB/s", "frac": fwd_bytes / (ms * 1e-3) / 1e9 / peak, "traffic": None,
                     "peak_kind": peak_kind},
        "clocks": clocks,
    }), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.batch_views > 0:
        run_batch(args, rank, world, local_rank)
    elif args.mode == "render":
        if rank == 0:
            run_render_sweep(args, local_rank)
    elif args.mode == "raster":
        if rank == 0:
            run_raster(args, local_rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
