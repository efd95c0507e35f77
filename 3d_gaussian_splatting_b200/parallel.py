"""Camera-sharded data parallelism.  The reference is single-process, single-device
(no NCCL/MPI call site anywhere); this is the B200-native extension: one process per GPU,
Gaussians (parameters + Adam moments) replicated, a batch of camera views split across ranks,
per-Gaussian gradients summed with ONE all-reduce over a contiguous slab, then the identical
fused Adam step on every rank.  The densification statistics are additive (sum / max), so they
are merged only when they are consumed (every densification_interval iterations).

Device-agnostic on purpose: the same code runs under gloo on CPU tensors in the tests.
"""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.distributed as dist


def shard_cameras(cameras: Sequence, rank: int, world: int) -> List:
    """Round-robin camera assignment: rank r renders views r, r+world, r+2*world, ...
    Every view is rendered by exactly one rank; shard sizes differ by at most one."""
    return [cameras[i] for i in range(rank, len(cameras), world)]


class GradientSlab:
    """One contiguous fp32 buffer holding the gradients of all parameter groups (59 floats per
    Gaussian at SH3), each group 16-byte aligned, exposed as views shaped like the parameters."""

    def __init__(self, params: Sequence[torch.Tensor]):
        self.sizes = [int(p.numel()) for p in params]
        self.offsets, off = [], 0
        for s in self.sizes:
            self.offsets.append(off)
            off += (s + 3) // 4 * 4
        dev = params[0].device if len(params) else "cpu"
        # 4 trailing floats; [0] is the step guard: a rank whose forward overflowed its fixed pair capacity raises it
        # to 1, the all-reduce spreads it, and the fused Adam step of EVERY rank then skips the iteration
        self.guard_offset = off
        off += 4
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.guard = self.flat[self.guard_offset:self.guard_offset + 1]
        self.views = [self.flat[o:o + s].view_as(p) for o, s, p in zip(self.offsets, self.sizes, params)]

    def zero_(self):
        self.flat.zero_()

    def all_reduce(self, group=None):
        """Sum over ranks (gradient of the summed per-view losses).  In place, one collective."""
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        return self.flat


def merge_densify_stats(max_radii2D: torch.Tensor, xyz_gradient_accum: torch.Tensor, denom: torch.Tensor,
                        group=None):
    """Combine the per-rank statistics (src/scene/gaussian_model.h:18-20) accumulated on disjoint
    camera shards: sums for xyz_gradient_accum/denom, max for max_radii2D.  After the call the
    buffers hold the GLOBAL statistics on every rank (call once, right before they are consumed,
    then reset them as the densification step does)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(xyz_gradient_accum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)
    return max_radii2D, xyz_gradient_accum, denom
