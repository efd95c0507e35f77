"""World-size-2 CPU (gloo) test of the camera-sharded data-parallel host logic:
camera sharding, the contiguous gradient slab + single all-reduce, statistics merge."""
import importlib
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
    N = 1001
    widths = [3, 3, 45, 1, 3, 4]
    params = [torch.zeros(N, w) for w in widths]
    slab = par.GradientSlab(params)
    assert all(o % 4 == 0 for o in slab.offsets) and slab.flat.numel() >= 59 * N
    cams = list(range(13))
    mine = par.shard_cameras(cams, rank, world)
    # "gradient of view c" = c+1 everywhere in group (c % 6); accumulate local views, then all-reduce
    for c in mine:
        slab.views[c % 6] += float(c + 1)
    slab.all_reduce()
    want = [sum(c + 1 for c in cams if c % 6 == k) for k in range(6)]
    ok = all(torch.all(v == w).item() for v, w in zip(slab.views, want))
    maxr = torch.tensor([rank + 1, 5 - rank], dtype=torch.int32)
    acc = torch.tensor([1.0 + rank, 2.0]); den = torch.tensor([1.0, float(rank)])
    par.merge_densify_stats(maxr, acc, den)
    ok &= maxr.tolist() == [2, 5] and acc.tolist() == [3.0, 4.0] and den.tolist() == [2.0, 1.0]
    q.put((rank, mine, bool(ok)))
    dist.destroy_process_group()


def test_camera_sharding_and_gradient_allreduce_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert all(ok for _, _, ok in res)
    shards = [m for _, m, _ in res]
    assert sorted(shards[0] + shards[1]) == list(range(13)) and abs(len(shards[0]) - len(shards[1])) <= 1


def test_shard_cameras_partition():
    par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
    for world in (1, 2, 4, 8):
        allv = []
        for r in range(world):
            allv += par.shard_cameras(list(range(64)), r, world)
        assert sorted(allv) == list(range(64))


class _StubModel:
    """The six parameter groups of GaussianModel::setup (src/scene/gaussian_model.cpp:323-345) on CPU tensors."""

    def __init__(self, N, seed):
        g = torch.Generator().manual_seed(seed)
        w = [(3,), (1, 3), (15, 3), (1,), (3,), (4,)]
        (self.xyz_, self.features_dc_, self.features_rest_, self.opacity_, self.scaling_,
         self.rotation_) = [torch.randn((N,) + s, generator=g) for s in w]

    def params(self):
        return [self.xyz_, self.features_dc_, self.features_rest_, self.opacity_, self.scaling_, self.rotation_]


def _torch_adam_segments(segs, guard, b1=0.9, b2=0.999, eps=1e-8):
    if guard is not None and float(guard[0]) != 0.0:
        return
    for p, g, m, v, lr, step in segs:
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        denom = (v.sqrt() / (1 - b2 ** step) ** 0.5).add_(eps)
        p.addcdiv_(m, denom, value=-lr / (1 - b1 ** step))


def _sharded_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    par = importlib.import_module("3d_gaussian_splatting_b200.parallel")
    N = 777
    model = _StubModel(N, seed=5)                       # identical replicas
    dense = [p.clone().requires_grad_(True) for p in model.params()]
    lr = [1.6e-4, 2.5e-3, 1.25e-4, 5e-2, 5e-3, 1e-3]
    ref = torch.optim.Adam([{"params": [p], "lr": l} for p, l in zip(dense, lr)], eps=1e-8)
    sh = par.ShardedAdam(model, rank, world, _torch_adam_segments, align=64)
    ok = sh.total % (world * 64) == 0 and sum(b - a for _, a, b in sh.pieces) <= sh.shard_len
    ok &= all(p.data_ptr() == sh.param_slab[o:].data_ptr() for p, o in zip(model.params(), sh.offsets))
    for it in range(1, 4):
        tot = [torch.zeros_like(p) for p in dense]
        for r in range(world):                           # every rank can rebuild every rank's gradient
            g = torch.Generator().manual_seed(100 * it + r)
            for k, p in enumerate(dense):
                gr = torch.randn(p.shape, generator=g)
                tot[k] += gr
                if r == rank:
                    model.grads[k].copy_(gr)
        for p, t in zip(dense, tot):
            p.grad = t
        ref.step()
        sh.step(lr, [it] * 6)
        for p, d in zip(model.params(), dense):
            ok &= torch.allclose(p, d.detach(), rtol=1e-5, atol=1e-7)
    # a raised guard on ONE rank makes every rank skip the step
    before = sh.param_slab.clone()
    sh.guard.fill_(1.0 if rank == 1 else 0.0)
    sh.step(lr, [4] * 6, use_guard=True)
    ok &= torch.equal(before, sh.param_slab)
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_sharded_adam_world2_matches_dense_adam():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sharded_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]


def test_view_sampler_without_replacement():
    gm = importlib.import_module("3d_gaussian_splatting_b200.gaussian_model")
    s = gm.ViewSampler(7, seed=3)
    for _ in range(3):                      # every epoch is a permutation of all views
        assert sorted(s() for _ in range(7)) == list(range(7))
    a, b = gm.ViewSampler(50, seed=1), gm.ViewSampler(50, seed=1)
    seq = [a() for _ in range(120)]
    assert seq == [b() for _ in range(120)] and seq[:50] != list(range(50))
