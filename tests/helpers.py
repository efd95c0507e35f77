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
