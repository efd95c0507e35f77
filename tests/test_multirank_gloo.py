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
