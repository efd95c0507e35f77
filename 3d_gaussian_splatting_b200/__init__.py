"""3d_gaussian_splatting_b200 — B200-native (sm_100a) differentiable Gaussian-splat rasterizer.

Drop-in for the render/rasterize call missing from seiya-kumada/3d_gaussian_splatting
(src/utils/train_utils.cpp:138-145).  Product code = csrc/ (hand-written CUDA + C ABI) and this
thin host layer.  The package name starts with a digit, so import it with
importlib.import_module("3d_gaussian_splatting_b200").
"""
from . import _lib  # noqa: F401  (raises ImportError when libgsplat_b200.so is not built)
from .camera import Camera, focal2fov, get_expon_lr_func, get_projection_matrix, get_world2view_2, look_at_camera
from .rasterizer import GaussianRasterizationSettings, Rasterizer

__all__ = ["Camera", "focal2fov", "get_expon_lr_func", "get_projection_matrix", "get_world2view_2",
           "look_at_camera", "GaussianRasterizationSettings", "Rasterizer"]
