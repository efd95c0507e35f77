"""2-rank probe: can one process map another GPU's buffer (CUDA IPC through the C ABI) and store into it?
Run: python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/peer_probe.py"""
import ctypes as C
import importlib
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3d_gaussian_splatting_b200")
L = importlib.import_module("3d_gaussian_splatting_b200._lib")
rast = pkg.Rasterizer(lr)
lib, ctx = L.lib, rast._ctx
n = 64 << 20
try:
    ptr = C.c_void_p()
    h = C.create_string_buffer(64)
    L.check(ctx, lib.gs_peer_alloc(ctx, n, C.byref(ptr), h))
    handles = [None] * world
    dist.all_gather_object(handles, bytes(h.raw))
    peers = {}
    for r in range(world):
        if r != rank:
            q = C.c_void_p()
            L.check(ctx, lib.gs_peer_open(ctx, handles[r], C.byref(q)))
            peers[r] = q.value
    rt = C.CDLL("libcudart.so.12")
    src = torch.full((n // 4,), float(rank + 1), device=dev)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.current_stream().cuda_stream
    for r, q in peers.items():
        for it in range(3):
            if it == 2: e0.record()
            rt.cudaMemcpyAsync(C.c_void_p(q), C.c_void_p(src.data_ptr()), C.c_size_t(n), 3, C.c_void_p(st))
        e1.record()
    torch.cuda.synchronize(); dist.barrier()
    back = torch.empty(4, device=dev)
    rt.cudaMemcpy(C.c_void_p(back.data_ptr()), ptr, C.c_size_t(16), 3)
    print(f"[ipc] rank {rank}: own buffer now holds {back.tolist()} (peer wrote), push {n / 1e6 / e0.elapsed_time(e1):.1f} GB/s", flush=True)
except Exception as e:
    print(f"[ipc] rank {rank} FAILED: {e!r}", flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(n // 4, dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(t, dist.group.WORLD)
    peer = hdl.get_buffer((rank + 1) % world, (n // 4,), torch.float32)
    peer.fill_(float(rank + 1))
    torch.cuda.synchronize(); dist.barrier()
    print(f"[symm] rank {rank}: own buffer {t[:2].tolist()}", flush=True)
except Exception as e:
    print(f"[symm] rank {rank} FAILED: {e!r}", flush=True)
dist.destroy_process_group()
