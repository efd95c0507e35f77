"""Host-side camera conventions — thin Python mirror of the reference's
src/utils/graphics_utils.{h,cpp} and src/scene/camera.{h,cpp}, implemented ONCE in C++
(csrc/host_conventions.cpp) and reached through the C ABI.  Same names and argument meaning as
the reference functions.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib


def _d(n, vals):
    return (C.c_double * n)(*[float(v) for v in np.asarray(vals, dtype=np.float64).reshape(-1)])


def focal2fov(focal: float, pixels: float) -> float:
    """src/utils/graphics_utils.cpp:4-7"""
    return float(_lib.lib.gs_focal2fov(float(focal), float(pixels)))


def get_world2view_2(R, t, translate=(0.0, 0.0, 0.0), scale: float = 1.0) -> np.ndarray:
    """src/utils/graphics_utils.cpp:32-43 — returns the row-major 4x4 world->view matrix (float64)."""
    out = (C.c_double * 16)()
    _lib.lib.gs_get_world2view2(_d(9, R), _d(3, t), _d(3, translate), float(scale), out)
    return np.array(out, dtype=np.float64).reshape(4, 4)


def get_projection_matrix(znear: float, zfar: float, fovX: float, fovY: float) -> np.ndarray:
    """src/utils/graphics_utils.cpp:46-72"""
    out = (C.c_double * 16)()
    _lib.lib.gs_get_projection_matrix(float(znear), float(zfar), float(fovX), float(fovY), out)
    return np.array(out, dtype=np.float64).reshape(4, 4)


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """src/utils/general_utils.cpp:112-142 (same parameter order as general_utils.h:8-13)."""
    def fn(step: int) -> float:
        return float(_lib.lib.gs_expon_lr(float(lr_init), float(lr_final), int(lr_delay_steps),
                                          float(lr_delay_mult), int(max_steps), int(step)))
    return fn


class Camera:
    """Mirror of the reference Camera (src/scene/camera.h:5-43): R is the camera-to-world rotation,
    T the world-to-camera translation (W2V[:3,:3] = R^T, W2V[:3,3] = T), FoVx/FoVy in radians.
    The derived tensors follow the constructor at src/scene/camera.cpp:44-45,66-71 and are
    exposed as the fp32 fields of a gs_settings struct (fp64 -> fp32 once, at the boundary)."""

    def __init__(self, R, T, FoVx: float, FoVy: float, image_height: int, image_width: int,
                 trans=(0.0, 0.0, 0.0), scale: float = 1.0, uid: int = 0, image_name: str = ""):
        self.uid, self.image_name = uid, image_name
        self.R = np.asarray(R, dtype=np.float64).reshape(3, 3)
        self.T = np.asarray(T, dtype=np.float64).reshape(3)
        self.FoVx, self.FoVy = float(FoVx), float(FoVy)
        self.image_height, self.image_width = int(image_height), int(image_width)
        self.zfar, self.znear = 100.0, 0.01
        st = _lib.GsSettings()
        _lib.lib.gs_camera_settings(_d(9, self.R), _d(3, self.T), _d(3, trans), float(scale), self.FoVx, self.FoVy,
                                    self.image_height, self.image_width, C.byref(st))
        self._st = st
        self.world_view_transform = np.array(st.viewmatrix, dtype=np.float32).reshape(4, 4)
        self.full_proj_transform = np.array(st.projmatrix, dtype=np.float32).reshape(4, 4)
        self.camera_center = np.array(st.campos, dtype=np.float32)
        self.tanfovx, self.tanfovy = float(st.tanfovx), float(st.tanfovy)


def look_at_camera(eye, target, up, focal_px: float, H: int, W: int, uid: int = 0) -> Camera:
    """Synthetic camera: position `eye`, looking at `target`; +z forward, +x right, +y down
    (the projection of graphics_utils.cpp:46-72 has z_sign=+1)."""
    eye = np.asarray(eye, np.float64); target = np.asarray(target, np.float64); up = np.asarray(up, np.float64)
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, up)       # image x
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)     # image y
    Rw2c = np.stack([right, down, fwd], axis=0)
    t = -Rw2c @ eye
    return Camera(Rw2c.T, t, focal2fov(focal_px, W), focal2fov(focal_px, H), H, W, uid=uid)
