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


class ShardedAdam:
    """Optimiser state sharded across the ranks (ZeRO-1 shape, but over ONE flat slab of 59 floats per Gaussian):

        gradients   : reduce-scatter   -> every rank holds the SUM over ranks of its 1/world slice
        Adam        : each rank updates the slice of the parameter slab it owns (moments exist only there)
        parameters  : all-gather (in place) -> every rank holds the updated replica again

    Same bytes on the wire as the all-reduce it replaces, but the Adam step -- the HBM-bound 16 B/float kernel that
    is the largest single stage of a step -- shrinks by 1/world, and so does the moment memory.
    The model's six parameter tensors (and gradient sinks) are re-bound as views of two flat slabs, each group
    16-byte aligned, the total padded to world * align floats so that every slice boundary is kernel-friendly.

    adam_fn(segments, skip_guard): segments = [(p, g, m, v, lr, step), ...] 1-D views of equal length -- the
    rasterizer's adam_step_flat on the GPU; a plain callable in the gloo tests.
    """

    def __init__(self, model, rank: int, world: int, adam_fn, group=None, align: int = 1024):
        params = model.params()
        self.model, self.rank, self.world, self.group, self.adam_fn = model, rank, world, group, adam_fn
        self.sizes = [int(p.numel()) for p in params]
        self.offsets, off = [], 0
        for s in self.sizes:
            self.offsets.append(off)
            off += (s + 3) // 4 * 4
        q = world * align
        self.total = (off + q - 1) // q * q
        self.shard_len = self.total // world
        self.lo, self.hi = rank * self.shard_len, (rank + 1) * self.shard_len
        dev = params[0].device
        self.param_slab = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.grad_slab = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.grad_shard = torch.zeros(self.shard_len, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(self.shard_len, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(self.shard_len, dtype=torch.float32, device=dev)
        self.guard = torch.zeros(1, dtype=torch.float32, device=dev)
        views = []
        for o, s, p in zip(self.offsets, self.sizes, params):
            v = self.param_slab[o:o + s].view_as(p)
            v.copy_(p)
            views.append(v)
        (model.xyz_, model.features_dc_, model.features_rest_, model.opacity_, model.scaling_, model.rotation_) = views
        model.grads = [self.grad_slab[o:o + s].view_as(p) for o, s, p in zip(self.offsets, self.sizes, params)]
        model.exp_avg = model.exp_avg_sq = None     # the dense moments are not used (and not kept) in this mode
        # the pieces of the owned slice, cut at the parameter-group boundaries: (group, start, stop) in slab floats
        self.pieces = [(k, max(o, self.lo), min(o + s, self.hi)) for k, (o, s) in enumerate(zip(self.offsets, self.sizes))
                       if max(o, self.lo) < min(o + s, self.hi)]
        b = dist.get_backend(group) if dist.is_initialized() else "none"
        self._native = (b == "nccl")   # gloo has neither reduce_scatter_tensor nor in-place all_gather_into_tensor

    def segments(self, lr, step):
        lo = self.lo
        return [(self.param_slab[a:b], self.grad_shard[a - lo:b - lo], self.exp_avg[a - lo:b - lo],
                 self.exp_avg_sq[a - lo:b - lo], lr[k], step[k]) for k, a, b in self.pieces]

    def step(self, lr, step, use_guard: bool = False):
        """lr / step: per-group learning rates and (already incremented) step counts.  The gradient slab is
        consumed (left zeroed for the next backward when the backward accumulates; the plain backward overwrites)."""
        multi = dist.is_initialized() and self.world > 1
        if multi and use_guard:    # a rank whose forward overflowed its pair capacity makes EVERY rank skip the step
            dist.all_reduce(self.guard, op=dist.ReduceOp.MAX, group=self.group)
        if not multi:
            self.grad_shard.copy_(self.grad_slab)
        elif self._native:
            dist.reduce_scatter_tensor(self.grad_shard, self.grad_slab, op=dist.ReduceOp.SUM, group=self.group)
        else:
            dist.all_reduce(self.grad_slab, op=dist.ReduceOp.SUM, group=self.group)
            self.grad_shard.copy_(self.grad_slab[self.lo:self.hi])
        self.adam_fn(self.segments(lr, step), self.guard if use_guard else None)
        if not multi:
            return
        if self._native:
            dist.all_gather_into_tensor(self.param_slab, self.param_slab[self.lo:self.hi], group=self.group)
        else:
            parts = [torch.empty(self.shard_len, dtype=torch.float32, device=self.param_slab.device)
                     for _ in range(self.world)]
            dist.all_gather(parts, self.param_slab[self.lo:self.hi].clone(), group=self.group)
            self.param_slab.copy_(torch.cat(parts))


class PeerShardedAdam:
    """The multi-GPU step with reduce-scatter, optimiser and all-gather FUSED into the kernels on either side
    (include/gsplat_b200.h, "the multi-GPU step over peer memory"):

        backward (K4b) : stores the gradient rows of shard q straight into rank q's slot buffer over NVLink
        barrier 1      : gs_peer_barrier (flag kernel; carries the MAX of the overflow guards)
        Adam (K5)      : sums the world slots of the owned rows, updates them, stores the new parameters into
                         every rank's replica over NVLink -- in TWO phases (split-phase step): the geometry groups
                         (xyz, opacity, scaling, rotation: 44 of a Gaussian's 236 B) + statistics on the caller's
                         stream, then barrier 2a and the next forward starts; the two SH groups (192 B) on a second
                         stream, then barrier 2b on a second flag block and an event -- the colour gate of the next
                         forward (gs_set_color_gate), which evaluates SH -> RGB only after K1, the depth argsort, the
                         duplicate stage and the tile sort.  The all-gather of 81 % of the parameter bytes thereby
                         overlaps the front of the next iteration.  Used from 4 ranks (GS_PEER_SPLIT=sm|ce|off
                         overrides; GS_PEER_NO_SPLIT=1: one phase, one barrier 2).

    Per rank and step the wire carries (world-1)/world of the gradient and parameter bytes once each -- the same
    as reduce-scatter + all-gather -- but no byte is staged, re-read or re-written in HBM for the exchange, and
    Adam touches 1/world of the model.  Every rank allocates ONE peer-mapped buffer holding its parameter replica
    and its slot buffer; the 64-byte IPC handles travel through torch.distributed (all_gather_object).

    rast: the rank's Rasterizer.  exchange(handle) -> list of all ranks' handles (default: all_gather_object).
    """

    def __init__(self, model, rast, rank: int, world: int, group=None, moments=None, pull: bool = False):
        """moments: optional (exp_avg[6], exp_avg_sq[6]) FULL (N, C) arrays to start the owned shards from (after a
        densification step, which needs and produces the full optimiser state); default: zeros.

        pull=True -- the layout for steps that render SEVERAL views per rank (north_star's 64-view batch): every rank
        accumulates the gradients of all its views into ONE local full-size slab (plain local reduce-adds, no
        routing), and after a barrier the owner's Adam kernel reads rows [r*R, (r+1)*R) of EVERY rank's slab straight
        over NVLink (peer loads), sums them in rank order, updates, and stores the parameters into every replica:
        one exchange per step, however many views went into it.  The buffer is the same size as in the push layout
        (world slots of R rows == one slab of N rows)."""
        self.model, self.rast, self.rank, self.world, self.group = model, rast, rank, world, group
        self.pull = bool(pull)
        params = model.params()
        N = self.N = int(params[0].shape[0])
        self.widths = [int(p.numel() // max(N, 1)) for p in params]
        R = self.R = ((N + world - 1) // world + 127) // 128 * 128
        # layout of a rank's buffer (floats): six (N, C) parameter arrays, then world slots of six (R, C) arrays
        self.p_off, off = [], 0
        for c in self.widths:
            self.p_off.append(off)
            off += (N * c + 3) // 4 * 4
        self.s_off = []
        if self.pull:
            # ONE gradient slab of world * R rows per group (rows past N are never written nor read)
            self.g_off = []
            for c in self.widths:
                self.g_off.append(off)
                off += world * R * c
        else:
            for s in range(world):
                row = []
                for c in self.widths:
                    row.append(off)
                    off += R * c
                self.s_off.append(row)
        # the flag block of the kernel barriers (gs_peer_barrier): 32 words, zeroed with the buffer
        off = (off + 31) // 32 * 32
        self.f_off = off
        off += 32
        self.f_off2 = off       # second flag block: the barrier of the SH phase, which runs on its own stream
        off += 32
        self.total = off
        self.epoch = 0
        self.epoch2 = 0
        import os
        self.flag_barriers = os.environ.get("GS_PEER_NCCL_BARRIER", "0") != "1"
        # split-phase step (see the class docstring): needs the flag barriers (a collective on a second stream would
        # have to be ordered against NCCL's own stream use) and a model with SH coefficients
        self.split = (world > 1 and self.flag_barriers and os.environ.get("GS_PEER_NO_SPLIT", "0") != "1"
                      and self.widths[1] + self.widths[2] > 0)
        # GS_PEER_SPLIT: "sm" -- a second Adam kernel on the second stream stores the SH rows into every replica, its
        # grid capped at 3 x SMs / ranks (GS_PEER_BESIDE_FORWARD); "ce" -- the Adam kernel updates the SH rows in
        # the own replica only and COPY ENGINES distribute them (gs_peer_copy); "off" -- one phase.  Measured (B200,
        # cfg2, one view per rank per step, ms/step; profiles/r02k_split_phase.md): 8 GPUs sm 1.56 / ce 1.83 / off
        # 1.71; 4 GPUs sm 1.49 / ce 1.52 / off 1.57; 2 GPUs sm 1.46 / off 1.42 (half the model's SH rows per rank is
        # HBM work, not NVLink time, and the two-kernel forward costs 45 us).  Default: sm from 4 ranks, else off.
        # (pull layout -- several views per rank and step: the exchange is a small share of the step and the split
        # measured within noise of one phase, 25.2 against 24.7 ms at 8 GPUs: off unless asked for)
        default_mode = "sm" if (world >= 4 and not self.pull) else "off"
        self.split_mode = os.environ.get("GS_PEER_SPLIT", default_mode) if self.split else "off"
        if self.split_mode == "off":
            self.split = False
        assert self.split_mode in ("off", "ce", "sm"), self.split_mode
        self.base, handle = rast.peer_alloc(self.total * 4)
        handles = [None] * world
        if world > 1:
            dist.all_gather_object(handles, handle, group=group)
        else:
            handles[0] = handle
        # map the peers' buffers; every rank must succeed or every rank gives up (same exception everywhere, so
        # the caller can fall back to an NCCL mode without ranks diverging)
        self.bases, failure = [], None
        for r in range(world):
            try:
                self.bases.append(self.base if r == rank else rast.peer_open(handles[r]))
            except Exception as e:      # e.g. CUDA IPC not permitted between these processes
                failure = failure or f"rank {rank} cannot map rank {r}: {e}"
                self.bases.append(0)
        if world > 1:
            ok = torch.tensor([0.0 if failure else 1.0], device=params[0].device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if ok.item() == 0.0:
                for r, b in enumerate(self.bases):
                    if r != rank and b:
                        rast.peer_close(b)
                rast.peer_free(self.base)
                raise RuntimeError("peer memory unavailable: " + (failure or "another rank failed to map a peer"))
        self.flat = rast.wrap(self.base, self.total)
        views = []
        for o, c, p in zip(self.p_off, self.widths, params):
            v = self.flat[o:o + N * c].view_as(p)
            v.copy_(p)
            views.append(v)
        (model.xyz_, model.features_dc_, model.features_rest_, model.opacity_, model.scaling_, model.rotation_) = views
        if self.pull:   # the model's gradient sinks ARE the local slab
            model.grads = [self.flat[o:o + N * c].view_as(p) for o, c, p in zip(self.g_off, self.widths, params)]
        else:
            model.grads = [None] * 6
        lo, hi = min(rank * R, N), min((rank + 1) * R, N)
        shard = []
        for full in (moments if moments is not None else (None, None)):
            ts = [torch.zeros(R * c, dtype=torch.float32, device=self.flat.device) for c in self.widths]
            if full is not None:
                for t, f, c in zip(ts, full, self.widths):
                    if c > 0 and hi > lo:
                        t[:(hi - lo) * c].copy_(f.reshape(N, c)[lo:hi].reshape(-1))
            shard.append(ts)
        model.exp_avg, model.exp_avg_sq = shard
        self.guard = torch.zeros(1, dtype=torch.float32, device=self.flat.device)
        self.sync = torch.zeros(1, dtype=torch.float32, device=self.flat.device)
        # params[r][k]: rank r's replica; slots[s][k]: MY slot for source s; routing[q][k]: my slot INSIDE rank q
        self.flag_ptrs = [self.bases[r] + 4 * self.f_off for r in range(world)]
        self.flag_ptrs2 = [self.bases[r] + 4 * self.f_off2 for r in range(world)]
        self.geo_widths = [self.widths[0], 0, 0, self.widths[3], self.widths[4], self.widths[5]]
        self.sh_widths = [0, self.widths[1], self.widths[2], 0, 0, 0]
        self.side = self.ev_fork = self.ev_sh = None
        self.profile_phases, self.phase_events = False, []
        if self.split:
            dev = self.flat.device
            self.copy_streams = [torch.cuda.Stream(device=dev) for _ in range(max(1, world - 1))]
            self.ev_copy = [torch.cuda.Event() for _ in range(max(1, world - 1))]
            self.ev_shadam = torch.cuda.Event()
            self.side = self.copy_streams[0]
            self.ev_fork = torch.cuda.Event()
            self.ev_sh = torch.cuda.Event()
            self.ev_sh.record(torch.cuda.current_stream(dev))     # (creates the handle; nothing to wait for yet)
            rast.set_color_gate(self.ev_sh)
        self.param_ptrs = [[self.bases[r] + 4 * self.p_off[k] for k in range(6)] for r in range(world)]
        if self.pull:
            # source s of my shard = rows [rank*R, ...) of rank s's slab (peer-mapped for s != rank)
            self.slot_ptrs = [[self.bases[s] + 4 * (self.g_off[k] + rank * R * self.widths[k]) for k in range(6)]
                              for s in range(world)]
        else:
            self.slot_ptrs = [[self.base + 4 * self.s_off[s][k] for k in range(6)] for s in range(world)]
            routing = [[self.bases[q] + 4 * self.s_off[rank][k] for k in range(6)] for q in range(world)]
            rast.set_gradient_routing(world, rank, R, routing)
        self.barrier()   # every replica initialised and every mapping in place before anyone stores remotely

    def join(self):
        """Make the current stream wait for the SH phase of the last step (a no-op without the split).  The forward
        does this itself through the colour gate; anything ELSE that reads or writes the SH coefficients, the SH
        moments or the gradient slots on the current stream must call this first."""
        if self.ev_sh is not None:
            torch.cuda.current_stream(self.flat.device).wait_event(self.ev_sh)

    def barrier(self, with_guard: bool = False, nccl: bool = False, second: bool = False):
        """Stream-ordered barrier across the ranks.  Default: gs_peer_barrier -- one 32-thread kernel per rank that
        stores an epoch flag into every rank's flag block over NVLink and waits for the others' (a few microseconds,
        against ~25 for a one-float NCCL all-reduce).  with_guard: the overflow guards ride along, self.guard becomes
        their maximum over the ranks.  nccl=True (or GS_PEER_NCCL_BARRIER=1): the collective instead."""
        if not (dist.is_initialized() and self.world > 1):
            return
        if nccl or not self.flag_barriers:
            dist.all_reduce(self.guard if with_guard else self.sync, op=dist.ReduceOp.MAX, group=self.group)
            return
        if second:          # the SH phase's own flag block and epoch sequence (it runs on its own stream)
            self.epoch2 += 1
            self.rast.peer_barrier(self.world, self.rank, self.flag_ptrs2, self.epoch2)
            return
        self.epoch += 1
        g = self.guard if with_guard else None
        self.rast.peer_barrier(self.world, self.rank, self.flag_ptrs, self.epoch, guard_in=g, guard_out=g)

    def phase_ms(self):
        """Median device times of the split-phase step's parts over the steps taken with profile_phases = True, from the
        end of barrier 1: SH copies (or SH Adam kernel) done on the second stream, + its barrier 2b, Adam kernel of the
        caller's stream done, + its barrier 2a."""
        torch.cuda.synchronize(self.flat.device)
        if not self.phase_events:
            return None
        import statistics
        rows = [[e[0].elapsed_time(e[k]) for k in (1, 2, 3, 4)] for e in self.phase_events]
        self.phase_events = []
        names = ("sh_phase_done", "sh_barrier_done", "adam_done", "geo_barrier_done")
        return {n: round(statistics.median(r[i] for r in rows), 4) for i, n in enumerate(names)}

    def barrier_timeouts(self) -> int:
        """Barriers of this rank that gave up waiting for a peer (~2 s each); 0 in a healthy run.  Synchronises."""
        torch.cuda.synchronize(self.flat.device)
        return int(self.flat[self.f_off + 16:self.f_off + 17].view(torch.int32).item()) + \
            int(self.flat[self.f_off2 + 16:self.f_off2 + 17].view(torch.int32).item())

    def slot(self, src: int, k: int) -> torch.Tensor:
        """The gradient rows source rank `src` stored for my shard, group k: (R, widths[k]) (push layout)."""
        assert not self.pull
        o = self.s_off[src][k]
        return self.flat[o:o + self.R * self.widths[k]].view(self.R, self.widths[k])

    def step(self, lr, step, radii=None, dL_dmeans2D=None, max_radii2D=None, xyz_grad_accum=None, denom=None,
             use_guard: bool = False):
        m = self.model
        # barrier 1: every rank's backward has landed in the slots; carries the overflow guard along
        self.barrier(with_guard=True)
        guard = self.guard if use_guard else None
        if not self.split:
            self.rast.adam_step_peer(self.world, self.rank, self.N, self.R, self.param_ptrs, self.slot_ptrs, m.exp_avg,
                                     m.exp_avg_sq, self.widths, lr, step, radii=radii, dL_dmeans2D=dL_dmeans2D,
                                     max_radii2D=max_radii2D, xyz_grad_accum=xyz_grad_accum, denom=denom,
                                     skip_guard=guard, sparse_rest=not self.pull)
            self.barrier()   # barrier 2: every replica complete before the next forward reads it
            return
        main = torch.cuda.current_stream(self.flat.device)
        prof = self.phase_events if self.profile_phases else None
        if prof is not None:      # (measurement only: CUDA events around the two phases of this step)
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            prof.append(evs)
            evs[0].record(main)
        if self.split_mode == "ce":
            # Geometry phase on the caller's stream, stored into every replica by the kernel: what K1 and the binning
            # stages of the next forward read -- the critical path, so nothing bulky shares NVLink with it for long.
            self.ev_fork.record(main)
            self.rast.adam_step_peer(self.world, self.rank, self.N, self.R, self.param_ptrs, self.slot_ptrs, m.exp_avg,
                                     m.exp_avg_sq, self.geo_widths, lr, step, radii=radii, dL_dmeans2D=dL_dmeans2D,
                                     max_radii2D=max_radii2D, xyz_grad_accum=xyz_grad_accum, denom=denom,
                                     skip_guard=guard, sparse_rest=not self.pull)
            if prof is not None:
                evs[3].record(main)
            self.barrier()   # barrier 2a: xyz / opacity / scaling / rotation of every replica complete
            if prof is not None:
                evs[4].record(main)
            # SH phase beside it: the SH groups are updated in the OWN replica only (GS_PEER_SH_LOCAL, second stream),
            # then the copy engines carry the owned rows to the other replicas -- one stream per peer, several engines
            # busy at a time, no SM involved -- beside the next forward; then that phase's own barrier (2b) and the
            # event the next colour kernel waits for
            self.side.wait_event(self.ev_fork)
            with torch.cuda.stream(self.side):
                self.rast.adam_step_peer(self.world, self.rank, self.N, self.R, self.param_ptrs, self.slot_ptrs,
                                         m.exp_avg, m.exp_avg_sq, self.sh_widths, lr, step, skip_guard=guard,
                                         sparse_rest=not self.pull, sh_local=True)
                self.ev_shadam.record(self.side)
            rows = max(0, min(self.R, self.N - self.rank * self.R))
            for d in range(1, self.world):
                r = (self.rank + d) % self.world
                st = self.copy_streams[d - 1]
                if d > 1:
                    st.wait_event(self.ev_shadam)
                with torch.cuda.stream(st):
                    for k in (2, 1):
                        w = self.widths[k]
                        if w and rows:
                            off = 4 * self.rank * self.R * w
                            self.rast.peer_copy(self.param_ptrs[r][k] + off, self.param_ptrs[self.rank][k] + off,
                                                4 * rows * w)
                    if d > 1:
                        self.ev_copy[d - 1].record(st)
            with torch.cuda.stream(self.side):          # (self.side is copy_streams[0])
                for d in range(2, self.world):
                    self.side.wait_event(self.ev_copy[d - 1])
                if prof is not None:
                    evs[1].record(self.side)
                self.barrier(second=True)
                self.ev_sh.record(self.side)
                if prof is not None:
                    evs[2].record(self.side)
            return
        # "sm": SH phase as a second Adam kernel on the second stream: f_dc / f_rest (81 % of the bytes), then its own
        # barrier (2b) and the event the next forward's colour kernel waits for.  (The previous step's SH phase is
        # complete: this step's forward waited for its event before the colour kernel.)
        self.ev_fork.record(main)
        self.side.wait_event(self.ev_fork)
        with torch.cuda.stream(self.side):
            self.rast.adam_step_peer(self.world, self.rank, self.N, self.R, self.param_ptrs, self.slot_ptrs, m.exp_avg,
                                     m.exp_avg_sq, self.sh_widths, lr, step, skip_guard=guard,
                                     sparse_rest=not self.pull, beside_forward=True)
            if prof is not None:
                evs[1].record(self.side)
            self.barrier(second=True)
            self.ev_sh.record(self.side)
            if prof is not None:
                evs[2].record(self.side)
        # geometry phase on the caller's stream: what K1 and the binning stages of the next forward read
        self.rast.adam_step_peer(self.world, self.rank, self.N, self.R, self.param_ptrs, self.slot_ptrs, m.exp_avg,
                                 m.exp_avg_sq, self.geo_widths, lr, step, radii=radii, dL_dmeans2D=dL_dmeans2D,
                                 max_radii2D=max_radii2D, xyz_grad_accum=xyz_grad_accum, denom=denom,
                                 skip_guard=guard, sparse_rest=not self.pull)
        if prof is not None:
            evs[3].record(main)
        self.barrier()   # barrier 2a: xyz / opacity / scaling / rotation of every replica complete
        if prof is not None:
            evs[4].record(main)

    def gather_moments(self):
        """The FULL (N, C) Adam moments on every rank (all-gather of the shards): what a densification step or a
        checkpoint needs.  Rare (once per densification interval), so plain NCCL."""
        self.join()
        out = []
        for shard in (self.model.exp_avg, self.model.exp_avg_sq):
            full = []
            for t, c, p in zip(shard, self.widths, self.model.params()):
                if c == 0:
                    full.append(torch.zeros_like(p))
                    continue
                buf = torch.empty(self.world * self.R * c, dtype=torch.float32, device=t.device)
                if dist.is_initialized() and self.world > 1:
                    dist.all_gather_into_tensor(buf, t, group=self.group)
                else:
                    buf.copy_(t)
                full.append(buf[:self.N * c].view_as(p).clone())
            out.append(full)
        return out[0], out[1]

    def detach_model(self):
        """Give the model private copies of its parameters (they are views of the peer buffer) so that the buffer
        can be released."""
        self.join()
        m = self.model
        (m.xyz_, m.features_dc_, m.features_rest_, m.opacity_, m.scaling_, m.rotation_) = [p.clone() for p in m.params()]

    def close(self):
        torch.cuda.synchronize(self.flat.device)
        self.barrier(nccl=True)               # nobody is still storing into a peer's buffer
        if self.split:
            self.rast.set_color_gate(None)
        if not self.pull:
            self.rast.set_gradient_routing(0)
        for r, b in enumerate(self.bases):
            if r != self.rank:
                self.rast.peer_close(b)
        torch.cuda.synchronize(self.flat.device)
        self.barrier(nccl=True)               # every mapping is gone before any owner frees (the flag blocks too)
        torch.cuda.synchronize(self.flat.device)
        self.flat = None
        self.rast.peer_free(self.base)
