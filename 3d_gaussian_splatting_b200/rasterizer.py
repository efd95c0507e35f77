"""The render call the reference trainer is missing (src/utils/train_utils.cpp:138-145),
as a Python host-side mirror over the C ABI.

`GaussianRasterizationSettings` carries the raster settings north_star lists; `Rasterizer` owns
one gs_ctx (one per process/device) and exposes forward / backward on torch CUDA tensors
(PyTorch is used for device memory and streams only).  `render()` is the reference-shaped
entry point: render(viewpoint_camera, gaussians, pipe, bg_color).  There is no fallback path:
everything below calls libgsplat_b200.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import GsGaussians, GsGradients, GsSettings, check, lib


@dataclass
class GaussianRasterizationSettings:
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: tuple
    scale_modifier: float
    viewmatrix: np.ndarray      # flat world_view_transform_ (src/scene/camera.cpp:66-68)
    projmatrix: np.ndarray      # flat full_proj_transform_  (src/scene/camera.cpp:70)
    campos: np.ndarray          # camera_center_             (src/scene/camera.cpp:71)
    sh_degree: int
    max_sh_degree: int = 3
    prefiltered: bool = False
    debug: bool = False         # PipelineParams::debug_ (src/arguments/params.h:97): sync + error + NaN/inf checks
    published_clamp_grad: bool = False   # GS_PUBLISHED_CLAMP_GRAD: the widely used derivative of the 1.3 tanfov guard

    def to_c(self) -> GsSettings:
        cached = self.__dict__.get("_c")
        if cached is not None:      # settings objects are immutable in practice; the ctypes struct is built once
            return cached
        st = GsSettings()
        st.H, st.W = int(self.image_height), int(self.image_width)
        st.tanfovx, st.tanfovy = float(self.tanfovx), float(self.tanfovy)
        vm = np.asarray(self.viewmatrix, dtype=np.float32).reshape(-1)
        pm = np.asarray(self.projmatrix, dtype=np.float32).reshape(-1)
        cp = np.asarray(self.campos, dtype=np.float32).reshape(-1)
        for k in range(3):
            st.bg[k] = float(self.bg[k])
            st.campos[k] = float(cp[k])
        for k in range(16):
            st.viewmatrix[k] = float(vm[k])
            st.projmatrix[k] = float(pm[k])
        st.scale_modifier = float(self.scale_modifier)
        st.sh_degree, st.max_sh_degree = int(self.sh_degree), int(self.max_sh_degree)
        st.prefiltered = int(self.prefiltered)
        st.debug = (_lib.GS_DEBUG_CHECKS if self.debug else 0) | (_lib.GS_PUBLISHED_CLAMP_GRAD if self.published_clamp_grad else 0)
        self.__dict__["_c"] = st
        return st

    @staticmethod
    def from_camera(cam, bg=(0.0, 0.0, 0.0), sh_degree=3, max_sh_degree=3, scale_modifier=1.0, debug=False,
                    published_clamp_grad=False):
        return GaussianRasterizationSettings(
            cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, tuple(bg), scale_modifier,
            cam.world_view_transform.reshape(-1), cam.full_proj_transform.reshape(-1), cam.camera_center,
            sh_degree, max_sh_degree, False, debug, published_clamp_grad)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_cuda:
        raise TypeError(f"{name} must be a float32 CUDA tensor")
    return t if t.is_contiguous() else t.contiguous()


@dataclass
class _Packed:
    g: GsGaussians
    keep: list = field(default_factory=list)  # tensors that must outlive the call
    K: int = 1
    split_sh: bool = False


class Rasterizer:
    """One rasterizer context on one device (ABI: one context per host thread and device)."""

    def __init__(self, device: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("libgsplat_b200 needs a CUDA (sm_100) device; there is no CPU fallback")
        self.device = torch.device("cuda", device)
        self._ctx = C.c_void_p()
        check(None, lib.gs_create(C.byref(self._ctx), int(device)))

    def close(self):
        if getattr(self, "_ctx", None):
            lib.gs_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- input packing ----------------------------------------------------------------------
    def _pack(self, means3D, scales, rotations, opacities, shs=None, sh_dc=None, sh_rest=None,
              colors_precomp=None, cov3D_precomp=None, raw_params=False) -> _Packed:
        g = GsGaussians()
        keep = []
        def reg(t, name):
            t = _f32c(t, name)
            keep.append(t)
            return t
        means3D = reg(means3D, "means3D")
        g.means3D = means3D.data_ptr()
        opacities = reg(opacities, "opacities")
        g.opacities = opacities.data_ptr()
        if cov3D_precomp is not None:
            g.cov3D_precomp = reg(cov3D_precomp, "cov3D_precomp").data_ptr()
        else:
            g.scales = reg(scales, "scales").data_ptr()
            g.rotations = reg(rotations, "rotations").data_ptr()
        K, split = 1, False
        if colors_precomp is not None:
            g.colors_precomp = reg(colors_precomp, "colors_precomp").data_ptr()
        elif shs is not None:           # get_features() layout (N,K,3)
            shs = reg(shs, "shs")
            K = shs.shape[1]
            g.sh_dc = shs.data_ptr()
            g.sh_dc_stride = 3 * K
            g.sh_rest = shs.data_ptr() + 12 if K > 1 else None
            g.sh_rest_stride = 3 * K
        else:                            # features_dc_ (N,1,3) + features_rest_ (N,K-1,3), no cat
            sh_dc = reg(sh_dc, "sh_dc")
            g.sh_dc = sh_dc.data_ptr()
            g.sh_dc_stride = 3
            K = 1
            if sh_rest is not None and sh_rest.numel() > 0:
                sh_rest = reg(sh_rest, "sh_rest")
                K = 1 + sh_rest.shape[1]
                g.sh_rest = sh_rest.data_ptr()
                g.sh_rest_stride = 3 * (K - 1)
            split = True
        g.raw_params = int(bool(raw_params))
        return _Packed(g, keep, K, split)

    # ---- forward ------------------------------------------------------------------------------
    def forward(self, settings: GaussianRasterizationSettings, means3D, scales=None, rotations=None,
                opacities=None, shs=None, sh_dc=None, sh_rest=None, colors_precomp=None, cov3D_precomp=None,
                raw_params=False, out_color=None, radii=None):
        """Returns (color (3,H,W) float32, radii (N) int32).  Keeps the forward state in the context."""
        N = int(means3D.shape[0])
        pk = self._pack(means3D, scales, rotations, opacities, shs, sh_dc, sh_rest, colors_precomp,
                        cov3D_precomp, raw_params)
        st = settings.to_c()
        if out_color is None:
            out_color = torch.empty((3, st.H, st.W), dtype=torch.float32, device=self.device)
        if radii is None:
            radii = torch.empty((N,), dtype=torch.int32, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_forward(self._ctx, C.byref(st), N, C.byref(pk.g), _ptr(out_color), _ptr(radii),
                                        C.c_void_p(stream)))
        self._last = (st, N, pk)
        return out_color, radii

    # ---- backward -----------------------------------------------------------------------------
    def backward(self, dL_dout_color: torch.Tensor, grads: Optional[dict] = None, accumulate: bool = False):
        """Gradients w.r.t. every Gaussian parameter of the last forward.  `grads` may supply
        pre-allocated sinks (same keys as the returned dict)."""
        st, N, pk = self._last
        dev = self.device
        K = pk.K
        grads = dict(grads or {})
        def sink(name, shape):
            if name not in grads:
                grads[name] = (torch.zeros if accumulate else torch.empty)(shape, dtype=torch.float32, device=dev)
            return grads[name]
        gg = GsGradients()
        if getattr(self, "_routing", False):
            # gs_set_gradient_routing: the six parameter gradients go to the owner ranks' slot buffers
            gg.dL_dmeans2D = sink("means2D", (N, 2)).data_ptr()
            dL = _f32c(dL_dout_color, "dL_dout_color")
            stream = torch.cuda.current_stream(dev).cuda_stream
            check(self._ctx, lib.gs_backward(self._ctx, C.byref(st), N, C.byref(pk.g), _ptr(dL), C.byref(gg),
                                             C.c_void_p(stream)))
            return grads
        gg.dL_dmeans3D = sink("means3D", (N, 3)).data_ptr()
        gg.dL_dopacities = sink("opacities", (N, 1)).data_ptr()
        gg.dL_dmeans2D = sink("means2D", (N, 2)).data_ptr()
        if pk.g.cov3D_precomp:
            gg.dL_dcov3D = sink("cov3D_precomp", (N, 6)).data_ptr()
        else:
            gg.dL_dscales = sink("scales", (N, 3)).data_ptr()
            gg.dL_drotations = sink("rotations", (N, 4)).data_ptr()
        if pk.g.colors_precomp:
            gg.dL_dcolors = sink("colors_precomp", (N, 3)).data_ptr()
        elif pk.split_sh:
            gg.dL_dsh_dc = sink("sh_dc", (N, 1, 3)).data_ptr()
            gg.sh_dc_stride = 3
            if K > 1:
                gg.dL_dsh_rest = sink("sh_rest", (N, K - 1, 3)).data_ptr()
                gg.sh_rest_stride = 3 * (K - 1)
        else:
            t = sink("shs", (N, K, 3))
            gg.dL_dsh_dc = t.data_ptr()
            gg.sh_dc_stride = 3 * K
            if K > 1:
                gg.dL_dsh_rest = t.data_ptr() + 12
                gg.sh_rest_stride = 3 * K
        gg.accumulate = int(accumulate)
        dL = _f32c(dL_dout_color, "dL_dout_color")
        stream = torch.cuda.current_stream(dev).cuda_stream
        check(self._ctx, lib.gs_backward(self._ctx, C.byref(st), N, C.byref(pk.g), _ptr(dL), C.byref(gg),
                                         C.c_void_p(stream)))
        return grads

    def backward_adam_step(self, dL_dout_color, exp_avg, exp_avg_sq, lr, step, beta1=0.9, beta2=0.999, eps=1e-8,
                           dL_dmeans2D=None, max_radii2D=None, xyz_grad_accum=None, denom=None, skip_guard=None):
        """Backward of the last forward with the Adam update fused into its per-Gaussian kernel
        (gs_backward_adam_step): the tensors given to forward() -- means3D, sh_dc, sh_rest, opacities, scales,
        rotations -- are the parameters and are updated in place; exp_avg / exp_avg_sq in the order of
        GaussianModel::setup (xyz, f_dc, f_rest, opacity, scaling, rotation)."""
        st, N, pk = self._last
        if isinstance(skip_guard, torch.Tensor):
            skip_guard = skip_guard.data_ptr()
        VP6 = C.c_void_p * 6
        arr = lambda ts: VP6(*[None if t is None or t.numel() == 0 else t.data_ptr() for t in ts])
        dL = _f32c(dL_dout_color, "dL_dout_color")
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_backward_adam_step(
            self._ctx, C.byref(st), N, C.byref(pk.g), _ptr(dL), _ptr(dL_dmeans2D), arr(exp_avg), arr(exp_avg_sq),
            (C.c_double * 6)(*[float(x) for x in lr]), (C.c_longlong * 6)(*[int(x) for x in step]), float(beta1),
            float(beta2), float(eps), _ptr(max_radii2D), _ptr(xyz_grad_accum), _ptr(denom),
            C.c_void_p(skip_guard) if skip_guard else None, C.c_void_p(stream)))

    # ---- running without the host round-trip ---------------------------------------------------
    def set_duplicate_capacity(self, pairs: int):
        """pairs > 0: fixed (tile, Gaussian) pair capacity, gs_forward never waits for the host; 0: synchronous."""
        check(self._ctx, lib.gs_set_duplicate_capacity(self._ctx, int(pairs)))

    def set_overflow_mirror(self, t: Optional[torch.Tensor]):
        """A caller-owned device float (e.g. the tail element of the all-reduced gradient slab) that the forward
        raises to 1.0 when it overflows the fixed capacity."""
        check(self._ctx, lib.gs_set_overflow_mirror(self._ctx, None if t is None else C.c_void_p(t.data_ptr())))

    def overflow_guard_ptr(self) -> int:
        return int(lib.gs_overflow_guard(self._ctx) or 0)

    def forward_status(self):
        """(D, overflow) of the last forward; the caller must have synchronised the stream it ran on."""
        D, ov = C.c_ulonglong(0), C.c_int(0)
        check(self._ctx, lib.gs_get_forward_status(self._ctx, C.byref(D), C.byref(ov)))
        return int(D.value), bool(ov.value)

    # ---- optimiser / loss ---------------------------------------------------------------------
    def adam_densify_step(self, N, params, grads, exp_avg, exp_avg_sq, widths, lr, step, beta1=0.9, beta2=0.999,
                          eps=1e-8, radii=None, dL_dmeans2D=None, max_radii2D=None, xyz_grad_accum=None,
                          denom=None, zero_grads=False, skip_guard=None):
        """Fused Adam over the six parameter groups of GaussianModel::setup
        (src/scene/gaussian_model.cpp:323-345) + densification statistics.  skip_guard: device float
        (tensor or raw pointer); non-zero at execution time turns the call into a no-op."""
        if isinstance(skip_guard, torch.Tensor):
            skip_guard = skip_guard.data_ptr()
        VP6 = C.c_void_p * 6
        arr = lambda ts: VP6(*[None if t is None else t.data_ptr() for t in ts])
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_adam_densify_step(
            self._ctx, int(N), arr(params), arr(grads), arr(exp_avg), arr(exp_avg_sq), (C.c_int * 6)(*widths),
            (C.c_double * 6)(*[float(x) for x in lr]), (C.c_longlong * 6)(*[int(s) for s in step]), float(beta1),
            float(beta2), float(eps), _ptr(radii), _ptr(dL_dmeans2D), _ptr(max_radii2D), _ptr(xyz_grad_accum),
            _ptr(denom), int(zero_grads), C.c_void_p(skip_guard) if skip_guard else None, C.c_void_p(stream)))

    def adam_step_flat(self, segments, beta1=0.9, beta2=0.999, eps=1e-8, zero_grads=False, skip_guard=None):
        """Adam over up to six contiguous float ranges: segments = [(p, g, m, v, lr, step), ...] with 1-D views of
        equal length (the slice of a flat parameter slab a rank owns, cut at the parameter-group boundaries)."""
        assert len(segments) <= 6
        VP6 = C.c_void_p * 6
        pad = lambda xs, fill: list(xs) + [fill] * (6 - len(xs))
        ptrs = lambda k: VP6(*pad([s[k].data_ptr() for s in segments], None))
        if isinstance(skip_guard, torch.Tensor):
            skip_guard = skip_guard.data_ptr()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_adam_step_flat(
            self._ctx, len(segments), ptrs(0), ptrs(1), ptrs(2), ptrs(3),
            (C.c_ulonglong * 6)(*pad([int(s[0].numel()) for s in segments], 0)),
            (C.c_double * 6)(*pad([float(s[4]) for s in segments], 0.0)),
            (C.c_longlong * 6)(*pad([int(s[5]) for s in segments], 1)), float(beta1), float(beta2), float(eps),
            int(zero_grads), C.c_void_p(skip_guard) if skip_guard else None, C.c_void_p(stream)))

    # ---- densify / prune / opacity reset ------------------------------------------------------
    def densify_plan(self, xyz_grad_accum, denom, opacities, scales, grad_threshold, min_opacity, extent,
                     percent_dense=0.01, max_screen_size=0.0):
        """Classify every Gaussian (gs_densify_plan).  Returns [kept originals, kept clones, kept split parents,
        pruned]; the new N is the first two plus twice the third.  Synchronises the current stream."""
        N = int(opacities.shape[0])
        counts = (C.c_int * 4)()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_densify_plan(self._ctx, N, _ptr(xyz_grad_accum), _ptr(denom), _ptr(opacities),
                                             _ptr(scales), float(grad_threshold), float(min_opacity), float(extent),
                                             float(percent_dense), float(max_screen_size), counts, C.c_void_p(stream)))
        return list(counts)

    def densify_apply(self, N, widths, src, dst, seed=0):
        """src / dst: (params, exp_avg, exp_avg_sq), each six tensors; moves rows and moments (gs_densify_apply)."""
        VP6 = C.c_void_p * 6
        arr = lambda ts: VP6(*[None if t is None or t.numel() == 0 else t.data_ptr() for t in ts])
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_densify_apply(self._ctx, int(N), (C.c_int * 6)(*widths), arr(src[0]), arr(src[1]),
                                              arr(src[2]), arr(dst[0]), arr(dst[1]), arr(dst[2]), int(seed),
                                              C.c_void_p(stream)))

    def reset_opacity(self, opacities, exp_avg, exp_avg_sq):
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_reset_opacity(self._ctx, int(opacities.shape[0]), _ptr(opacities), _ptr(exp_avg),
                                              _ptr(exp_avg_sq), C.c_void_p(stream)))

    # ---- multi-GPU step over peer memory --------------------------------------------------------
    def peer_alloc(self, nbytes: int):
        """Zero-filled device buffer other processes may map: (device pointer, 64-byte IPC handle)."""
        ptr, h = C.c_void_p(), C.create_string_buffer(64)
        check(self._ctx, lib.gs_peer_alloc(self._ctx, int(nbytes), C.byref(ptr), h))
        return int(ptr.value), bytes(h.raw)

    def peer_open(self, handle: bytes) -> int:
        ptr = C.c_void_p()
        check(self._ctx, lib.gs_peer_open(self._ctx, handle, C.byref(ptr)))
        return int(ptr.value)

    def peer_close(self, ptr: int):
        check(self._ctx, lib.gs_peer_close(self._ctx, C.c_void_p(ptr)))

    def peer_free(self, ptr: int):
        check(self._ctx, lib.gs_peer_free(self._ctx, C.c_void_p(ptr)))

    def peer_barrier(self, nranks: int, rank: int, flag_blocks, epoch: int, guard_in=None, guard_out=None):
        """gs_peer_barrier: flag-based barrier across the ranks of the peer-memory step (one 32-thread kernel)."""
        tab = (C.c_void_p * nranks)(*[int(p) for p in flag_blocks])
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_peer_barrier(self._ctx, nranks, rank, tab, int(epoch) & 0xFFFFFFFF, _ptr(guard_in),
                                             _ptr(guard_out), C.c_void_p(stream)))

    def peer_copy(self, dst_ptr: int, src_ptr: int, nbytes: int):
        """gs_peer_copy: one contiguous copy-engine copy on the current stream (either side may be peer-mapped)."""
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_peer_copy(self._ctx, C.c_void_p(int(dst_ptr)), C.c_void_p(int(src_ptr)), int(nbytes),
                                          C.c_void_p(stream)))

    def set_color_gate(self, event=None):
        """gs_set_color_gate: event = a torch.cuda.Event that has been recorded at least once (its handle exists
        from then on); later forwards run K1 without the SH -> RGB evaluation and evaluate the colours after the
        event's most recent record, just ahead of the blend.  None: back to the one-kernel K1."""
        self._gate_event = event      # keep the event object (and its handle) alive
        check(self._ctx, lib.gs_set_color_gate(self._ctx, C.c_void_p(event.cuda_event) if event is not None else None))

    def wrap(self, ptr: int, nfloats: int) -> torch.Tensor:
        """A float32 torch view of raw device memory (no copy, no ownership)."""
        class _Raw:
            pass
        r = _Raw()
        r.__cuda_array_interface__ = dict(shape=(int(nfloats),), typestr="<f4", data=(int(ptr), False), version=2)
        t = torch.as_tensor(r, device=self.device)
        assert t.data_ptr() == ptr
        return t

    def set_gradient_routing(self, nranks: int, rank: int = 0, rows_per_shard: int = 0, table=None):
        """table[q][k]: device pointer of this rank's slot for group k inside rank q's slot buffer; nranks=0: off."""
        if not nranks:
            check(self._ctx, lib.gs_set_gradient_routing(self._ctx, 0, 0, 0, None))
            self._routing = False
            return
        flat = (C.c_void_p * (nranks * 6))(*[int(table[q][k]) for q in range(nranks) for k in range(6)])
        check(self._ctx, lib.gs_set_gradient_routing(self._ctx, nranks, rank, rows_per_shard, flat))
        self._routing = True

    def adam_step_peer(self, nranks, rank, N, rows_per_shard, params, slots, exp_avg, exp_avg_sq, widths, lr, step,
                       beta1=0.9, beta2=0.999, eps=1e-8, radii=None, dL_dmeans2D=None, max_radii2D=None,
                       xyz_grad_accum=None, denom=None, skip_guard=None, sparse_rest=True, beside_forward=False,
                       sh_local=False):
        """params[r][k] / slots[s][k]: device pointers (ints); exp_avg / exp_avg_sq: six shard-local tensors.
        sparse_rest=False: the pull layout (GS_PEER_PULL): slots point into the ranks' full gradient slabs."""
        if isinstance(skip_guard, torch.Tensor):
            skip_guard = skip_guard.data_ptr()
        def tab(t):
            # the pointer tables of the peer optimiser are the same lists every step: build the ctypes array once
            cache = self.__dict__.setdefault("_peer_tab_cache", {})
            hit = cache.get(id(t))
            if hit is None or hit[0] is not t:
                hit = cache[id(t)] = (t, (C.c_void_p * (nranks * 6))(*[int(t[r][k]) for r in range(nranks) for k in range(6)]))
            return hit[1]
        VP6 = C.c_void_p * 6
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_adam_step_peer(
            self._ctx, nranks, rank, int(N), int(rows_per_shard), tab(params), tab(slots),
            VP6(*[t.data_ptr() for t in exp_avg]), VP6(*[t.data_ptr() for t in exp_avg_sq]), (C.c_int * 6)(*widths),
            (C.c_double * 6)(*[float(x) for x in lr]), (C.c_longlong * 6)(*[int(x) for x in step]), float(beta1),
            float(beta2), float(eps), _ptr(radii), _ptr(dL_dmeans2D), _ptr(max_radii2D), _ptr(xyz_grad_accum),
            _ptr(denom), C.c_void_p(skip_guard) if skip_guard else None, (0 if sparse_rest else 1) | (2 if beside_forward else 0) | (4 if sh_local else 0),
            C.c_void_p(stream)))

    def loss(self, image, gt, lambda_dssim=0.0, want_grad=True):
        """(1-lambda)*L1 + lambda*(1-SSIM); returns (loss 1-element device tensor, dL/dimage)."""
        H, W = int(image.shape[1]), int(image.shape[2])
        loss = torch.empty((1,), dtype=torch.float32, device=self.device)
        d = torch.empty_like(image) if want_grad else None
        stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._ctx, lib.gs_loss(self._ctx, H, W, _ptr(_f32c(image, "image")), _ptr(_f32c(gt, "gt")),
                                     float(lambda_dssim), _ptr(loss), _ptr(d), C.c_void_p(stream)))
        return loss, d

    # ---- parity harness / profiling -------------------------------------------------------------
    _DBG_DTYPES = dict(MEANS2D=np.float32, DEPTHS=np.float32, CONIC_OPACITY=np.float32, RGB=np.float32,
                       CLAMPED=np.uint8, TILES_TOUCHED=np.uint32, DEPTH_ORDER=np.uint32, KEYS_SORTED=np.uint64,
                       POINT_LIST=np.uint32, RANGES=np.uint32, FINAL_T=np.float32, N_CONTRIB=np.uint32,
                       DL_DMEAN2D=np.float32, DL_DCONIC=np.float32, DL_DOPACITY=np.float32, DL_DCOLOR=np.float32,
                       NUM_RENDERED=np.uint64)

    def debug_get(self, name: str) -> np.ndarray:
        what = _lib.DBG[name]
        n = int(lib.gs_debug_size(self._ctx, what))
        if n < 0:
            check(self._ctx, n)
        out = np.empty(n // np.dtype(self._DBG_DTYPES[name]).itemsize, dtype=self._DBG_DTYPES[name])
        check(self._ctx, lib.gs_debug_get(self._ctx, what, out.ctypes.data_as(C.c_void_p), n))
        return out

    def set_profiling(self, enabled: bool):
        check(self._ctx, lib.gs_set_profiling(self._ctx, int(enabled)))

    def stage_ms(self) -> dict:
        ms = (C.c_float * _lib.GS_NUM_STAGES)()
        check(self._ctx, lib.gs_get_stage_ms(self._ctx, ms))
        return dict(zip(_lib.STAGE_NAMES, [float(x) for x in ms]))
