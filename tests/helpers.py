"""Shared helpers of the parity tests: build identical inputs for the CPU oracle (numpy) and the
CUDA path (torch tensors through the C ABI)."""
import importlib

import numpy as np

from oracle.oracle_lib import RasterSettings

syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
cammod = importlib.import_module("3d_gaussian_splatting_b200.camera")
rastmod = importlib.import_module("3d_gaussian_splatting_b200.rasterizer")


def oracle_settings(cam, deg, maxdeg, bg=(0.0, 0.0, 0.0), scale_modifier=1.0):
    return RasterSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, tuple(bg), scale_modifier,
                          cam.world_view_transform.reshape(-1), cam.full_proj_transform.reshape(-1),
                          cam.camera_center, deg, maxdeg)


def gpu_settings(cam, deg, maxdeg, bg=(0.0, 0.0, 0.0), scale_modifier=1.0, debug=False):
    return rastmod.GaussianRasterizationSettings.from_camera(cam, bg, deg, maxdeg, scale_modifier, debug)


def small_scene(N=2000, H=96, W=128, deg=3, seed=3, scale_boost=0.0, focal=120.0, eye=(0.4, 0.3, 3.5)):
    scene = syn.make_scene(N, deg, seed=seed)
    scene.log_scales += np.float32(scale_boost)
    cam = cammod.look_at_camera(eye, (0, 0, 0), (0, 1, 0), focal, H, W)
    return scene, cam


def rel_err(a, b):
    """max |a-b| relative to the largest magnitude of the reference array b (per-array scale)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def elem_err(a, b, floor=1e-3):
    """Per-element relative error with an absolute floor: max |a-b| / max(|b|, floor * max|b|).  An element 1000x
    smaller than the largest must still be right to 1e-3 of ITSELF; below the floor the bar is 1e-3 * floor * max."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    if b.size == 0:
        return 0.0
    scale = np.maximum(np.abs(b), floor * max(np.abs(b).max(), 1e-300))
    return float((np.abs(a - b) / scale).max())


def frac_elem_beyond(a, b, tol=1e-3, floor=1e-3):
    """Share of the elements whose floored per-element relative error (elem_err's measure) exceeds tol."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    if b.size == 0:
        return 0.0
    scale = np.maximum(np.abs(b), floor * max(np.abs(b).max(), 1e-300))
    return float((np.abs(a - b) / scale > tol).mean())


def min_row_cosine(a, b, floor=1e-3):
    """Smallest cosine similarity between matching per-Gaussian rows (rows whose reference norm is above floor * the
    largest row norm; the others are covered by elem_err's absolute floor).  1.0 when no row qualifies."""
    a = np.asarray(a, np.float64).reshape(len(a), -1); b = np.asarray(b, np.float64).reshape(len(b), -1)
    if a.shape[1] < 2 or b.size == 0:
        return 1.0
    nb = np.linalg.norm(b, axis=1); na = np.linalg.norm(a, axis=1)
    keep = nb > floor * max(nb.max(), 1e-300)
    if not keep.any():
        return 1.0
    cos = (a[keep] * b[keep]).sum(1) / np.maximum(na[keep] * nb[keep], 1e-300)
    return float(cos.min())


def raw_reference(oracle, scene, cam, deg, maxdeg, bg=(0.0, 0.0, 0.0), dL=None):
    """The oracle's statement of the RAW-parameter path (the one the training step and bench.py run): specified
    exp / sigmoid (orc_activate_raw), un-normalised quaternion into the covariance construction, and -- for the
    backward -- the activation Jacobians of the getters (src/scene/gaussian_model.cpp:54-59, :270-298) applied to
    the oracle's gradients, so that every gradient is with respect to the STORED parameter."""
    sc, rq, op = oracle.activate_raw(scene.log_scales, scene.raw_rotations, scene.opacity_logits)
    shs = np.ascontiguousarray(np.concatenate([scene.features_dc, scene.features_rest], axis=1)[:, :(maxdeg + 1) ** 2])
    so = oracle_settings(cam, deg, maxdeg, bg)
    pre, b, fwd = oracle.forward(so, scene.xyz, sc, rq, op, shs)
    out = dict(pre=pre, b=b, fwd=fwd, so=so, shs=shs, act=(sc, rq, op))
    if dL is not None:
        g = oracle.backward(so, scene.xyz, sc, rq, shs, pre, b, fwd, dL)
        g = dict(g)
        g["dL_dscales"] = (g["dL_dscales"] * sc).astype(np.float32)                       # d exp(s)/ds = exp(s)
        g["dL_dopacities"] = (g["dL_dopacities"].reshape(-1, 1) * (op * (1 - op))).astype(np.float32)  # d sigmoid
        # dL_drotations is already with respect to the raw quaternion (the oracle differentiates the normalisation)
        out["g_ref"] = g
    return out
