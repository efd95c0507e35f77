"""ctypes binding of libgsplat_b200.so — the C ABI declared in include/gsplat_b200.h.

The library is built in-tree (3d_gaussian_splatting_b200/build.py).  There is no Python or CPU
fallback: if the shared library is missing, importing this module raises, and every compute
entry point returns GS_ECUDA when no sm_100 GPU is present.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgsplat_b200.so")

GS_OK, GS_EINVAL, GS_ENOMEM, GS_ECUDA, GS_EOVERFLOW, GS_ESTATE, GS_ENUMERIC = 0, -1, -2, -3, -4, -5, -6
GS_DEBUG_CHECKS, GS_PUBLISHED_CLAMP_GRAD = 1, 2
GS_NUM_STAGES = 9
STAGE_NAMES = ("preprocess", "depth_sort", "scan_duplicate", "sort", "ranges", "blend_fwd", "blend_bwd", "preprocess_bwd", "sh_color")

DBG = dict(MEANS2D=0, DEPTHS=1, CONIC_OPACITY=2, RGB=3, CLAMPED=4, TILES_TOUCHED=5, DEPTH_ORDER=6,
           KEYS_SORTED=7, POINT_LIST=8, RANGES=9, FINAL_T=10, N_CONTRIB=11, DL_DMEAN2D=12, DL_DCONIC=13,
           DL_DOPACITY=14, DL_DCOLOR=15, NUM_RENDERED=16)


class GsSettings(C.Structure):
    _fields_ = [
        ("H", C.c_int), ("W", C.c_int),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float),
        ("bg", C.c_float * 3), ("scale_modifier", C.c_float),
        ("viewmatrix", C.c_float * 16), ("projmatrix", C.c_float * 16), ("campos", C.c_float * 3),
        ("sh_degree", C.c_int), ("max_sh_degree", C.c_int),
        ("prefiltered", C.c_int), ("debug", C.c_int),
    ]


class GsGaussians(C.Structure):
    _fields_ = [
        ("means3D", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p), ("opacities", C.c_void_p),
        ("sh_dc", C.c_void_p), ("sh_rest", C.c_void_p),
        ("sh_dc_stride", C.c_int), ("sh_rest_stride", C.c_int),
        ("colors_precomp", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("raw_params", C.c_int),
    ]


class GsGradients(C.Structure):
    _fields_ = [
        ("dL_dmeans3D", C.c_void_p), ("dL_dscales", C.c_void_p), ("dL_drotations", C.c_void_p),
        ("dL_dopacities", C.c_void_p), ("dL_dsh_dc", C.c_void_p), ("dL_dsh_rest", C.c_void_p),
        ("sh_dc_stride", C.c_int), ("sh_rest_stride", C.c_int),
        ("dL_dmeans2D", C.c_void_p), ("dL_dcolors", C.c_void_p), ("dL_dcov3D", C.c_void_p),
        ("accumulate", C.c_int),
    ]


class GsTrainConfig(C.Structure):
    """gs_train_config (OptimizationParams of src/arguments/params.h:52-68 + spatial_lr_scale, bg, seed)"""
    _fields_ = [
        ("position_lr_init", C.c_double), ("position_lr_final", C.c_double), ("position_lr_delay_mult", C.c_double),
        ("position_lr_max_steps", C.c_int),
        ("feature_lr", C.c_double), ("opacity_lr", C.c_double), ("scaling_lr", C.c_double), ("rotation_lr", C.c_double),
        ("percent_dense", C.c_double), ("lambda_dssim", C.c_double),
        ("densification_interval", C.c_int), ("opacity_reset_interval", C.c_int), ("densify_from_iter", C.c_int),
        ("densify_until_iter", C.c_int),
        ("densify_grad_threshold", C.c_double), ("spatial_lr_scale", C.c_double),
        ("bg", C.c_float * 3), ("seed", C.c_uint), ("active_sh_degree", C.c_int), ("random_background", C.c_int),
    ]


# every symbol include/gsplat_b200.h declares (tests check the library exports all of them)
EXPORTS = (
    "gs_version", "gs_create", "gs_destroy", "gs_last_error", "gs_forward", "gs_backward",
    "gs_adam_densify_step", "gs_loss", "gs_loss_static", "gs_debug_size", "gs_debug_get", "gs_set_profiling", "gs_get_stage_ms",
    "gs_focal2fov", "gs_get_world2view2", "gs_get_projection_matrix", "gs_camera_settings", "gs_expon_lr",
    "gs_adam_step_flat", "gs_trainer_create", "gs_trainer_destroy", "gs_trainer_last_error", "gs_trainer_add_view", "gs_trainer_run",
    "gs_trainer_num_gaussians", "gs_trainer_iteration", "gs_trainer_retries", "gs_trainer_get_params",
    "gs_trainer_set_pair_capacity", "gs_trainer_save", "gs_trainer_load", "gs_peer_barrier", "gs_set_color_gate", "gs_peer_copy",
    "gs_backward_adam_step", "gs_densify_plan", "gs_densify_apply", "gs_reset_opacity", "gs_debug_reuse_blend_gradients", "gs_set_gradient_routing", "gs_adam_step_peer", "gs_peer_alloc", "gs_peer_open", "gs_peer_close", "gs_peer_free", "gs_set_duplicate_capacity", "gs_set_overflow_mirror", "gs_overflow_guard", "gs_get_forward_status",
)


class GsError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libgsplat_b200 error {code}: {msg}")
        self.code = code


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python 3d_gaussian_splatting_b200/build.py` "
            "(or __graft_entry__.build()).  There is no CPU/Python fallback for the render path.")
    lib = C.CDLL(LIB_PATH)
    vp, i, f, d = C.c_void_p, C.c_int, C.c_float, C.c_double
    lib.gs_version.restype = C.c_char_p
    lib.gs_last_error.restype = C.c_char_p
    lib.gs_last_error.argtypes = [vp]
    lib.gs_create.argtypes = [C.POINTER(vp), i]
    lib.gs_destroy.argtypes = [vp]
    lib.gs_destroy.restype = None
    lib.gs_forward.argtypes = [vp, C.POINTER(GsSettings), i, C.POINTER(GsGaussians), vp, vp, vp]
    lib.gs_backward.argtypes = [vp, C.POINTER(GsSettings), i, C.POINTER(GsGaussians), vp, C.POINTER(GsGradients), vp]
    lib.gs_adam_densify_step.argtypes = [vp, i, vp * 6, vp * 6, vp * 6, vp * 6, i * 6, d * 6, C.c_longlong * 6,
                                         d, d, d, vp, vp, vp, vp, vp, i, vp, vp]
    lib.gs_adam_step_flat.argtypes = [vp, i, vp * 6, vp * 6, vp * 6, vp * 6, C.c_ulonglong * 6, d * 6, C.c_longlong * 6,
                                      d, d, d, i, vp, vp]
    lib.gs_backward_adam_step.argtypes = [vp, C.POINTER(GsSettings), i, C.POINTER(GsGaussians), vp, vp, vp * 6, vp * 6,
                                          d * 6, C.c_longlong * 6, d, d, d, vp, vp, vp, vp, vp]
    lib.gs_debug_reuse_blend_gradients.argtypes = [vp, i]
    lib.gs_trainer_create.argtypes = [C.POINTER(vp), i, C.POINTER(GsTrainConfig), i, i, vp, vp, vp, vp, vp, vp]
    lib.gs_trainer_destroy.argtypes = [vp]
    lib.gs_trainer_destroy.restype = None
    lib.gs_trainer_last_error.argtypes = [vp]
    lib.gs_trainer_last_error.restype = C.c_char_p
    lib.gs_trainer_add_view.argtypes = [vp, C.POINTER(GsSettings), vp]
    lib.gs_trainer_run.argtypes = [vp, i, vp, i, C.c_float]
    lib.gs_trainer_num_gaussians.argtypes = [vp]
    lib.gs_trainer_iteration.argtypes = [vp]
    lib.gs_trainer_retries.argtypes = [vp]
    lib.gs_trainer_set_pair_capacity.argtypes = [vp, C.c_ulonglong]
    lib.gs_trainer_save.argtypes = [vp, C.c_char_p]
    lib.gs_trainer_load.argtypes = [vp, C.c_char_p]
    lib.gs_trainer_get_params.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    f = C.c_float
    lib.gs_densify_plan.argtypes = [vp, i, vp, vp, vp, vp, f, f, f, f, f, i * 4, vp]
    lib.gs_densify_apply.argtypes = [vp, i, i * 6, vp * 6, vp * 6, vp * 6, vp * 6, vp * 6, vp * 6, C.c_ulonglong, vp]
    lib.gs_reset_opacity.argtypes = [vp, i, vp, vp, vp, vp]
    lib.gs_set_gradient_routing.argtypes = [vp, i, i, i, C.POINTER(vp)]
    lib.gs_adam_step_peer.argtypes = [vp, i, i, i, i, C.POINTER(vp), C.POINTER(vp), vp * 6, vp * 6, i * 6, d * 6,
                                      C.c_longlong * 6, d, d, d, vp, vp, vp, vp, vp, vp, i, vp]
    lib.gs_peer_barrier.argtypes = [vp, i, i, C.POINTER(vp), C.c_uint, vp, vp, vp]
    lib.gs_peer_alloc.argtypes = [vp, C.c_ulonglong, C.POINTER(vp), C.c_char_p]
    lib.gs_peer_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    lib.gs_peer_close.argtypes = [vp, vp]
    lib.gs_peer_free.argtypes = [vp, vp]
    lib.gs_set_duplicate_capacity.argtypes = [vp, C.c_ulonglong]
    lib.gs_set_overflow_mirror.argtypes = [vp, vp]
    lib.gs_overflow_guard.argtypes = [vp]
    lib.gs_overflow_guard.restype = vp
    lib.gs_get_forward_status.argtypes = [vp, C.POINTER(C.c_ulonglong), C.POINTER(i)]
    lib.gs_loss.argtypes = [vp, i, i, vp, vp, f, vp, vp, vp]
    lib.gs_set_color_gate.argtypes = [vp, vp]
    lib.gs_peer_copy.argtypes = [vp, vp, vp, C.c_ulonglong, vp]
    lib.gs_loss_static.argtypes = [i, i, i, vp, vp, f, vp, vp, vp]
    lib.gs_debug_size.argtypes = [vp, i]
    lib.gs_debug_size.restype = C.c_longlong
    lib.gs_debug_get.argtypes = [vp, i, vp, C.c_size_t]
    lib.gs_set_profiling.argtypes = [vp, i]
    lib.gs_get_stage_ms.argtypes = [vp, f * GS_NUM_STAGES]
    lib.gs_focal2fov.argtypes = [d, d]
    lib.gs_focal2fov.restype = d
    lib.gs_get_world2view2.argtypes = [d * 9, d * 3, d * 3, d, d * 16]
    lib.gs_get_world2view2.restype = None
    lib.gs_get_projection_matrix.argtypes = [d, d, d, d, d * 16]
    lib.gs_get_projection_matrix.restype = None
    lib.gs_camera_settings.argtypes = [d * 9, d * 3, d * 3, d, d, d, i, i, C.POINTER(GsSettings)]
    lib.gs_camera_settings.restype = None
    lib.gs_expon_lr.argtypes = [f, f, i, f, i, i]
    lib.gs_expon_lr.restype = f
    return lib


lib = _load()


def check(ctx, rc: int) -> None:
    if rc != GS_OK:
        msg = lib.gs_last_error(ctx)
        raise GsError(rc, msg.decode() if msg else "")
