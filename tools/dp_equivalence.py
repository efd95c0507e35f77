"""2+ ranks: the three data-parallel modes must train the same model.  Runs a few steps of a small config in
each mode from identical replicas and prints the max-abs parameter difference against the all-reduce mode
(world 2: the rank-ordered slot sum equals the all-reduce sum, so the peer mode is expected to be bit-identical
up to the non-deterministic atomics of the backward blend)."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world, lrk = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lrk)
dev = torch.device("cuda", lrk)
dist.init_process_group("nccl", device_id=dev)
syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
scene, cams, deg = syn.make_config("cfg1_100k_800_sh3", n_cameras=2 * world)
mine = par.shard_cameras(cams, rank, world)
results = {}
gts = None
for mode in ("allreduce", "sharded", "peer"):
    model = gm.GaussianModel(deg, lrk).load_scene(scene)
    tr = gm.Trainer(model, gm.OptimizationParams(), lambda_dssim=0.2)
    if gts is None:
        gt_model = gm.GaussianModel(deg, lrk).load_scene(scene)
        gt_model.features_dc_ += 0.05
        gts = [gm.render(c, gt_model, gm.PipelineParams(), (0., 0., 0.), tr.rast)["render"].clone() for c in mine]
    if mode == "peer":
        sh = tr.shard_optimizer_peer(rank, world); guard = sh.guard
    elif mode == "sharded":
        sh = tr.shard_optimizer(rank, world); guard = sh.guard
    else:
        slab = par.GradientSlab(model.params()); model.grads = slab.views; guard = slab.guard
        tr.grad_sync = slab.all_reduce
    tr.reserve_pairs(4_000_000, guard=guard)
    losses = []
    for it in range(6):
        losses.append(float(tr.train_step_device(mine[it % len(mine)], gts[it % len(mine)])))
    torch.cuda.synchronize()
    dist.barrier()
    flat = torch.cat([p.reshape(-1) for p in model.params()]).clone()
    # all ranks must hold the same replica
    other = flat.clone()
    dist.broadcast(other, src=0)
    results[mode] = (flat, float((flat - other).abs().max()), losses)
    if mode == "peer":
        tr.peer.close()
    del tr, model
ref = results["allreduce"][0]
for mode, (flat, rep_diff, losses) in results.items():
    d = (flat - ref).abs().max().item()
    print(f"rank {rank} {mode:9s}: max|p - p_allreduce| = {d:.3e}, replica divergence vs rank 0 = {rep_diff:.1e}, "
          f"losses {['%.5f' % l for l in losses]}", flush=True)
dist.destroy_process_group()
