"""ctypes front-end of the CPU oracle — TEST INFRASTRUCTURE, never imported by the product.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module.  It loads oracle/_build/liboracle.so (fp32) or liboracle_f64.so (fp64,
finite-difference validation) and exposes the stage functions of orc_*.c on numpy arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")


def build(force: bool = False) -> None:
    """Compile the oracle with the committed Makefile (gcc only, a few seconds)."""
    libs = [os.path.join(_BUILD, n) for n in ("liboracle.so", "liboracle_f64.so")]
    if not force and all(os.path.exists(p) for p in libs):
        srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
        if min(os.path.getmtime(p) for p in libs) >= max(os.path.getmtime(s) for s in srcs):
            return
    subprocess.run(["make", "-C", _HERE, "-s"], check=True)


def _settings_type(real):
    class Settings(C.Structure):
        _fields_ = [
            ("H", C.c_int), ("W", C.c_int),
            ("tanfovx", real), ("tanfovy", real),
            ("bg", real * 3), ("scale_modifier", real),
            ("viewmatrix", real * 16), ("projmatrix", real * 16), ("campos", real * 3),
            ("sh_degree", C.c_int), ("max_sh_degree", C.c_int),
            ("prefiltered", C.c_int), ("debug", C.c_int),
        ]
    return Settings


@dataclass
class RasterSettings:
    """Raster settings of the missing render() call (see oracle/orc_common.h for the mapping
    of every field to the reference object that supplies it)."""
    H: int
    W: int
    tanfovx: float
    tanfovy: float
    bg: tuple
    scale_modifier: float
    viewmatrix: np.ndarray   # 16 floats, flat world_view_transform_ (column-major V)
    projmatrix: np.ndarray   # 16 floats, flat full_proj_transform_ (column-major P*V)
    campos: np.ndarray       # 3 floats
    sh_degree: int
    max_sh_degree: int
    prefiltered: int = 0
    debug: int = 0


class Oracle:
    def __init__(self, f64: bool = False):
        build()
        self.f64 = f64
        self.real = C.c_double if f64 else C.c_float
        self.dt = np.float64 if f64 else np.float32
        self.lib = C.CDLL(os.path.join(_BUILD, "liboracle_f64.so" if f64 else "liboracle.so"))
        self.Settings = _settings_type(self.real)
        self.lib.orc_scan_tiles.restype = C.c_uint64
        self.lib.orc_l1_loss.restype = C.c_double
        self.lib.orc_photometric_loss.restype = C.c_double
        self.lib.orc_focal2fov.restype = C.c_double
        self.lib.orc_expon_lr.restype = C.c_float
        self.lib.orc_expon_lr.argtypes = [C.c_float, C.c_float, C.c_int, C.c_float, C.c_int, C.c_int]
        self.lib.orc_focal2fov.argtypes = [C.c_double, C.c_double]

    def set_tile_cull(self, enabled: bool):
        """Exact tile culling on/off (orc_common.h); on by default -- off only to prove it is conservative."""
        self.lib.orc_set_tile_cull(C.c_int(1 if enabled else 0))

    # ---- helpers -------------------------------------------------------------------
    def arr(self, a, dtype=None):
        return np.ascontiguousarray(a, dtype=self.dt if dtype is None else dtype)

    @staticmethod
    def p(a):
        return None if a is None else a.ctypes.data_as(C.c_void_p)

    def settings(self, s: RasterSettings):
        st = self.Settings()
        st.H, st.W = int(s.H), int(s.W)
        st.tanfovx, st.tanfovy = float(s.tanfovx), float(s.tanfovy)
        for i in range(3):
            st.bg[i] = float(s.bg[i])
            st.campos[i] = float(s.campos[i])
        st.scale_modifier = float(s.scale_modifier)
        vm = np.asarray(s.viewmatrix, dtype=np.float64).reshape(-1)
        pm = np.asarray(s.projmatrix, dtype=np.float64).reshape(-1)
        for i in range(16):
            st.viewmatrix[i] = float(vm[i])
            st.projmatrix[i] = float(pm[i])
        st.sh_degree, st.max_sh_degree = int(s.sh_degree), int(s.max_sh_degree)
        st.prefiltered, st.debug = int(s.prefiltered), int(s.debug)
        return st

    @staticmethod
    def _split_sh(shs):
        """(N,K,3) contiguous -> (dc ptr view, dc stride, rest view, rest stride)."""
        N, K, _ = shs.shape
        flat = shs.reshape(N, K * 3)
        return flat, 3 * K, (flat[:, 3:] if K > 1 else None), 3 * K

    def activate_raw(self, log_scales, raw_rotations, opacity_logits):
        """The activations as the RAW-parameter path of the kernels evaluates them (orc_activate_raw): specified exp /
        sigmoid; the quaternion stays un-normalised (the covariance construction normalises it, once)."""
        ls = self.arr(log_scales); rq = self.arr(raw_rotations); ol = self.arr(opacity_logits).reshape(-1)
        N = ls.shape[0]
        sc = np.zeros((N, 3), self.dt); rot = np.zeros((N, 4), self.dt); op = np.zeros(N, self.dt)
        self.lib.orc_activate_raw(C.c_int(N), self.p(ls), self.p(rq), self.p(ol), self.p(sc), self.p(rot), self.p(op))
        return sc, rot, op.reshape(-1, 1)

    # ---- stages --------------------------------------------------------------------
    def preprocess(self, s, means3D, scales, rotations, opacities, shs, colors_precomp=None, cov3D_precomp=None):
        st = self.settings(s)
        means3D = self.arr(means3D); N = means3D.shape[0]
        scales = self.arr(scales); rotations = self.arr(rotations); opacities = self.arr(opacities).reshape(-1)
        shs = self.arr(shs)
        K = shs.shape[1]
        cp = None if colors_precomp is None else self.arr(colors_precomp)
        cv = None if cov3D_precomp is None else self.arr(cov3D_precomp)
        out = dict(
            means2D=np.zeros((N, 2), self.dt), depths=np.zeros(N, self.dt), cov3D=np.zeros((N, 6), self.dt),
            conic_opacity=np.zeros((N, 4), self.dt), rgb=np.zeros((N, 3), self.dt),
            clamped=np.zeros((N, 3), np.uint8), radii=np.zeros(N, np.int32), tiles_touched=np.zeros(N, np.uint32))
        rest_ptr = C.c_void_p(shs.ctypes.data + 3 * shs.itemsize) if K > 1 else None
        self.lib.orc_preprocess(C.byref(st), C.c_int(N), self.p(means3D), self.p(scales), self.p(rotations),
                                self.p(opacities), self.p(shs), C.c_int(3 * K), rest_ptr, C.c_int(3 * K),
                                self.p(cp), self.p(cv), self.p(out["means2D"]), self.p(out["depths"]),
                                self.p(out["cov3D"]), self.p(out["conic_opacity"]), self.p(out["rgb"]),
                                self.p(out["clamped"]), self.p(out["radii"]), self.p(out["tiles_touched"]))
        return out

    def bin(self, s, pre):
        N = pre["radii"].shape[0]
        gx, gy = (s.W + 15) // 16, (s.H + 15) // 16
        offsets = np.zeros(N, np.uint32)
        D = int(self.lib.orc_scan_tiles(C.c_int(N), self.p(pre["tiles_touched"]), self.p(offsets)))
        keys = np.zeros(max(D, 1), np.uint64); values = np.zeros(max(D, 1), np.uint32)
        self.lib.orc_duplicate_with_keys(C.c_int(s.H), C.c_int(s.W), C.c_int(N), self.p(pre["means2D"]),
                                         self.p(pre["depths"]), self.p(pre["radii"]), self.p(pre["conic_opacity"]),
                                         self.p(offsets), self.p(keys), self.p(values))
        keys_unsorted = keys.copy(); values_unsorted = values.copy()
        self.lib.orc_sort_pairs(C.c_uint64(D), self.p(keys), self.p(values))
        ranges = np.zeros((gx * gy, 2), np.uint32)
        self.lib.orc_identify_tile_ranges(C.c_uint64(D), self.p(keys), C.c_int(gx * gy), self.p(ranges))
        return dict(D=D, offsets=offsets, keys=keys[:D], values=values[:D], ranges=ranges,
                    keys_unsorted=keys_unsorted[:D], values_unsorted=values_unsorted[:D])

    def blend_forward(self, s, pre, b):
        st = self.settings(s)
        H, W = s.H, s.W
        out_color = np.zeros((3, H, W), self.dt); final_T = np.zeros((H, W), self.dt)
        n_contrib = np.zeros((H, W), np.uint32); stats = np.zeros(3, np.uint64)
        vals = np.ascontiguousarray(b["values"]) if b["D"] > 0 else np.zeros(1, np.uint32)
        self.lib.orc_blend_forward(C.byref(st), self.p(b["ranges"]), self.p(vals), self.p(pre["means2D"]),
                                   self.p(pre["conic_opacity"]), self.p(pre["rgb"]), self.p(out_color),
                                   self.p(final_T), self.p(n_contrib), self.p(stats))
        return dict(out_color=out_color, final_T=final_T, n_contrib=n_contrib,
                    D_eff=int(stats[0]), sum_contrib=int(stats[1]), interactions=int(stats[2]))

    def blend_backward(self, s, pre, b, fwd, dL_dpix):
        st = self.settings(s)
        N = pre["radii"].shape[0]
        dL_dpix = self.arr(dL_dpix)
        g = dict(dL_dmean2D=np.zeros((N, 2), self.dt), dL_dconic=np.zeros((N, 3), self.dt),
                 dL_dopacity=np.zeros(N, self.dt), dL_dcolor=np.zeros((N, 3), self.dt))
        vals = np.ascontiguousarray(b["values"]) if b["D"] > 0 else np.zeros(1, np.uint32)
        self.lib.orc_blend_backward(C.byref(st), C.c_int(N), self.p(b["ranges"]), self.p(vals),
                                    self.p(pre["means2D"]), self.p(pre["conic_opacity"]), self.p(pre["rgb"]),
                                    self.p(fwd["final_T"]), self.p(fwd["n_contrib"]), self.p(dL_dpix),
                                    self.p(g["dL_dmean2D"]), self.p(g["dL_dconic"]), self.p(g["dL_dopacity"]),
                                    self.p(g["dL_dcolor"]))
        return g

    def preprocess_backward(self, s, means3D, scales, rotations, shs, pre, g, cov3D_precomp=None,
                            colors_are_precomp=False):
        st = self.settings(s)
        means3D = self.arr(means3D); N = means3D.shape[0]
        scales = self.arr(scales); rotations = self.arr(rotations); shs = self.arr(shs)
        K = shs.shape[1]
        cv = None if cov3D_precomp is None else self.arr(cov3D_precomp)
        out = dict(dL_dmeans3D=np.zeros((N, 3), self.dt), dL_dscales=np.zeros((N, 3), self.dt),
                   dL_drotations=np.zeros((N, 4), self.dt), dL_dshs=np.zeros((N, K, 3), self.dt),
                   dL_dcov3D=np.zeros((N, 6), self.dt))
        rest_ptr = C.c_void_p(shs.ctypes.data + 3 * shs.itemsize) if K > 1 else None
        grest_ptr = C.c_void_p(out["dL_dshs"].ctypes.data + 3 * shs.itemsize) if K > 1 else None
        self.lib.orc_preprocess_backward(
            C.byref(st), C.c_int(N), self.p(means3D), self.p(scales), self.p(rotations), self.p(shs),
            C.c_int(3 * K), rest_ptr, C.c_int(3 * K), self.p(cv), C.c_int(int(colors_are_precomp)),
            self.p(pre["radii"]), self.p(pre["clamped"]), self.p(g["dL_dmean2D"]), self.p(g["dL_dconic"]),
            self.p(g["dL_dcolor"]), self.p(out["dL_dmeans3D"]), self.p(out["dL_dscales"]),
            self.p(out["dL_drotations"]), self.p(out["dL_dshs"]), C.c_int(3 * K), grest_ptr, C.c_int(3 * K),
            self.p(out["dL_dcov3D"]))
        out["dL_dopacities"] = g["dL_dopacity"].copy()
        out["dL_dmeans2D"] = g["dL_dmean2D"].copy()
        return out

    # ---- composed render path (what the reference's render() + backward would do) ----
    def forward(self, s, means3D, scales, rotations, opacities, shs, **kw):
        pre = self.preprocess(s, means3D, scales, rotations, opacities, shs, **kw)
        b = self.bin(s, pre)
        fwd = self.blend_forward(s, pre, b)
        return pre, b, fwd

    def backward(self, s, means3D, scales, rotations, shs, pre, b, fwd, dL_dpix, **kw):
        g = self.blend_backward(s, pre, b, fwd, dL_dpix)
        return self.preprocess_backward(s, means3D, scales, rotations, shs, pre, g, **kw)

    # ---- optimiser / loss ------------------------------------------------------------
    def adam_step(self, p, g, m, v, lr, beta1, beta2, eps, step):
        assert p.dtype == np.float32 and p.flags.c_contiguous
        self.lib.orc_adam_step(C.c_size_t(p.size), self.p(p), self.p(g), self.p(m), self.p(v), C.c_double(lr),
                               C.c_double(beta1), C.c_double(beta2), C.c_double(eps), C.c_longlong(step))

    def densify_stats(self, radii, dL_dmean2D, max_radii2D, accum, denom):
        self.lib.orc_densify_stats(C.c_int(radii.shape[0]), self.p(radii), self.p(dL_dmean2D),
                                   self.p(max_radii2D), self.p(accum), self.p(denom))

    # ---- densify / prune / opacity reset (orc_densify.c) -------------------------------------
    class DensifyParams(C.Structure):
        _fields_ = [("grad_threshold", C.c_float), ("min_opacity", C.c_float), ("extent", C.c_float),
                    ("percent_dense", C.c_float), ("max_screen_size", C.c_float), ("seed", C.c_ulonglong)]

    def split_normals(self, seed, parent, copy):
        z = (C.c_float * 3)()
        self.lib.orc_split_normals(C.c_ulonglong(seed), C.c_ulonglong(parent), C.c_int(copy), z)
        return np.array(z[:], np.float32)

    def densify_and_prune(self, params, exp_avg, exp_avg_sq, xyz_grad_accum, denom, max_radii2D, grad_threshold,
                          min_opacity, extent, percent_dense=0.01, max_screen_size=0.0, seed=0):
        """params / exp_avg / exp_avg_sq: six float32 arrays (xyz, f_dc, f_rest, opacity, scaling, rotation), N rows.
        Returns (new_params, new_exp_avg, new_exp_avg_sq, counts[4])."""
        N = params[0].shape[0]
        widths = [int(p.size // max(N, 1)) for p in params]
        f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        P, M, V = [f32(a) for a in params], [f32(a) for a in exp_avg], [f32(a) for a in exp_avg_sq]
        cap = 2 * N + 1
        out = [[np.zeros((cap, max(w, 1)), np.float32) for w in widths] for _ in range(3)]
        FP6 = C.POINTER(C.c_float) * 6
        ptr = lambda arrs: FP6(*[a.ctypes.data_as(C.POINTER(C.c_float)) for a in arrs])
        prm = self.DensifyParams(grad_threshold, min_opacity, extent, percent_dense, max_screen_size, seed)
        counts = (C.c_int * 4)()
        self.lib.orc_densify_and_prune.restype = C.c_int
        n = self.lib.orc_densify_and_prune(
            C.c_int(N), (C.c_int * 6)(*widths), ptr(P), ptr(M), ptr(V),
            f32(xyz_grad_accum).ctypes.data_as(C.POINTER(C.c_float)), f32(denom).ctypes.data_as(C.POINTER(C.c_float)),
            np.ascontiguousarray(max_radii2D, dtype=np.int32).ctypes.data_as(C.POINTER(C.c_int)), C.byref(prm),
            ptr(out[0]), ptr(out[1]), ptr(out[2]), counts)
        shp = lambda a, src: a.reshape(-1)[:n * (src.size // max(N, 1))].reshape((n,) + src.shape[1:])
        res = [[shp(o, src).copy() for o, src in zip(out[j], params)] for j in range(3)]
        return res[0], res[1], res[2], list(counts)

    def reset_opacity(self, opacity, exp_avg, exp_avg_sq):
        for a in (opacity, exp_avg, exp_avg_sq):
            assert a.dtype == np.float32 and a.flags.c_contiguous
        fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
        self.lib.orc_reset_opacity(C.c_int(opacity.size), fp(opacity), fp(exp_avg), fp(exp_avg_sq))

    def photometric_loss(self, img, gt, lambda_dssim=0.2, want_grad=True):
        """(1-lambda) L1 + lambda (1 - SSIM) (orc_ssim.c); returns (loss, dL/dimg or None, l1, ssim)."""
        img = self.arr(img); gt = self.arr(gt)
        _, H, W = img.shape
        d = np.zeros_like(img) if want_grad else None
        l1, ss = C.c_double(0), C.c_double(0)
        loss = self.lib.orc_photometric_loss(C.c_int(H), C.c_int(W), self.p(img), self.p(gt), C.c_double(lambda_dssim),
                                             self.p(d), C.byref(l1), C.byref(ss))
        return float(loss), d, float(l1.value), float(ss.value)

    def l1_loss(self, img, gt):
        img = self.arr(img); gt = self.arr(gt)
        d = np.zeros_like(img)
        loss = self.lib.orc_l1_loss(C.c_size_t(img.size), self.p(img), self.p(gt), self.p(d))
        return float(loss), d
