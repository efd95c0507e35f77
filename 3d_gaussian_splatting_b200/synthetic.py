"""Deterministic synthetic scenes and cameras for the BASELINE.json configs.

Replaces the reference's Scene / dataset readers (out of scope: they need real datasets that do
not exist on the box).  numpy Philox (counter-based) streams with fixed seeds, so every process
and rank generates identical bits.  Parameters are produced RAW, in the storage convention of
GaussianModel::CoreParams (src/scene/gaussian_model.h:9-40): log-scales, un-normalised
quaternions (w,x,y,z), opacity logits, features_dc (N,1,3) + features_rest (N,K-1,3).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

# (the camera module is imported lazily: it binds libgsplat_b200.so, and bench.py's CPU reference arm loads this
#  file on its own -- scene generation is pure numpy -- with a camera factory built on the oracle's conventions)

# name -> (N, W, H, focal_px, sh_degree, n_cameras)   [BASELINE.json configs 0..4]
CONFIGS = {
    "cfg0_10k_256_sh0": (10_000, 256, 256, 280.0, 0, 1),
    "cfg1_100k_800_sh3": (100_000, 800, 800, 1111.1, 3, 100),
    "cfg2_1p6m_1246x825_sh3": (1_600_000, 1246, 825, 1000.0, 3, 64),
    "cfg3_4m_1080p_sh3": (4_000_000, 1920, 1080, 1600.0, 3, 16),
    "cfg4_6m_1246x825_sh3": (6_000_000, 1246, 825, 1000.0, 3, 64),
}


@dataclass
class SyntheticScene:
    xyz: np.ndarray            # (N,3)
    log_scales: np.ndarray     # (N,3)   scaling_ (log space)
    raw_rotations: np.ndarray  # (N,4)   rotation_ (w,x,y,z), not normalised
    opacity_logits: np.ndarray # (N,1)   opacity_ (logit)
    features_dc: np.ndarray    # (N,1,3)
    features_rest: np.ndarray  # (N,K-1,3)
    max_sh_degree: int

    @property
    def N(self):
        return self.xyz.shape[0]

    # activations of the getters, src/scene/gaussian_model.cpp:270-298 (float32 numpy)
    def activated(self):
        scales = np.exp(self.log_scales).astype(np.float32)
        q = self.raw_rotations
        rot = (q / np.maximum(np.linalg.norm(q, axis=1, keepdims=True), 1e-12)).astype(np.float32)
        opac = (1.0 / (1.0 + np.exp(-self.opacity_logits.astype(np.float32)))).astype(np.float32)
        shs = np.concatenate([self.features_dc, self.features_rest], axis=1).astype(np.float32)
        return scales, rot, opac, shs


def make_scene(N: int, max_sh_degree: int = 3, seed: int = 1234, radius: float = 1.0) -> SyntheticScene:
    rng = np.random.Generator(np.random.Philox(seed))
    K = (max_sh_degree + 1) ** 2
    # positions: uniform in the ball of radius R
    d = rng.standard_normal((N, 3)).astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = radius * rng.random((N, 1)) ** (1.0 / 3.0)
    xyz = (d * r).astype(np.float32)
    s0 = 0.9 * radius * N ** (-1.0 / 3.0)
    log_scales = (math.log(s0) + 0.7 * rng.standard_normal((N, 3))).astype(np.float32)
    raw_rot = rng.standard_normal((N, 4)).astype(np.float32)
    opacity_logits = (-1.0 + 1.5 * rng.standard_normal((N, 1))).astype(np.float32)
    f_dc = (0.5 * rng.standard_normal((N, 1, 3))).astype(np.float32)
    f_rest = (0.1 * rng.standard_normal((N, K - 1, 3))).astype(np.float32)
    return SyntheticScene(xyz, log_scales, raw_rot, opacity_logits, f_dc, f_rest, max_sh_degree)


def orbit_cameras(n: int, W: int, H: int, focal_px: float, radius: float = 4.0, elevation_deg: float = 30.0,
                  look_at_camera=None):
    """n cameras on a sphere of radius 4R looking at the origin, up = +y, elevation swinging +-30 deg.
    look_at_camera: camera factory (eye, target, up, focal_px, H, W, uid=) -- default: camera.look_at_camera."""
    if look_at_camera is None:
        from .camera import look_at_camera
    cams = []
    for k in range(n):
        az = 2.0 * math.pi * k / max(n, 1)
        el = math.radians(elevation_deg) * math.sin(2.0 * math.pi * k / max(n, 1) * 3.0 + 0.5)
        eye = radius * np.array([math.cos(el) * math.sin(az), math.sin(el), math.cos(el) * math.cos(az)])
        cams.append(look_at_camera(eye, (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), focal_px, H, W, uid=k))
    return cams


def make_config(name: str, n_cameras: int | None = None, N: int | None = None, look_at_camera=None,
                variant: str = "default"):
    """(scene, cameras, sh_degree) for one of the BASELINE.json configs; seed = 1234 + config index.
    variant "stress": the same scene with the opacity logits shifted by -2 (N(-3, 1.5^2): faint Gaussians, long
    blend lists, little early termination) -- the second workload bench.py reports beside the default one."""
    n, W, H, focal, deg, ncam = CONFIGS[name]
    idx = list(CONFIGS).index(name)
    scene = make_scene(N if N is not None else n, max_sh_degree=deg, seed=1234 + idx)
    if variant == "stress":
        scene.opacity_logits -= np.float32(2.0)
    elif variant != "default":
        raise ValueError(f"unknown workload variant {variant!r}")
    cams = orbit_cameras(n_cameras if n_cameras is not None else ncam, W, H, focal, look_at_camera=look_at_camera)
    return scene, cams, deg
