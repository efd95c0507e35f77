"""Real-data ingestion next to the camera objects: COLMAP binary model readers and the NeRF++
normalisation that yields the scene extent / spatial_lr_scale, mirroring the reference's
src/scene/colmap_loader.cpp and src/scene/dataset_readers.cpp (same names, same argument meaning, same errors).
Pure host code: numpy + struct.  Image decoding is left to the caller (the reference uses cv::imread, which this
image does not ship): CameraInfo carries the image path, and `load_image` may be passed to attach pixels.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np

from .camera import Camera, focal2fov, get_world2view_2

# src/scene/colmap_loader.cpp:196-208
CAMERA_MODELS = {0: ("SIMPLE_PINHOLE", 3), 1: ("PINHOLE", 4), 2: ("SIMPLE_RADIAL", 4), 3: ("RADIAL", 5),
                 4: ("OPENCV", 8), 5: ("OPENCV_FISHEYE", 8), 6: ("FULL_OPENCV", 12), 7: ("FOV", 5),
                 8: ("SIMPLE_RADIAL_FISHEYE", 4), 9: ("RADIAL_FISHEYE", 5), 10: ("THIN_PRISM_FISHEYE", 12)}


@dataclass
class Image:
    """colmap::Image (src/scene/colmap_loader.cpp:10-27)"""
    id_: int
    qvec_: np.ndarray          # (4) w, x, y, z
    tvec_: np.ndarray          # (3)
    camera_id_: int
    name_: str
    xys_: np.ndarray           # (P, 2)
    point3D_ids_: np.ndarray   # (P) int64


@dataclass
class ColmapCamera:
    """colmap::Camera (src/scene/colmap_loader.cpp:29-40)"""
    id_: int
    model_: str
    width_: int
    height_: int
    params_: np.ndarray


def _read(f, fmt: str):
    size = struct.calcsize("<" + fmt)
    buf = f.read(size)
    if len(buf) != size:
        raise RuntimeError("unexpected end of COLMAP model file")
    return struct.unpack("<" + fmt, buf)


def read_extrinsics_binary(path_to_model_file: str) -> Dict[int, Image]:
    """src/scene/colmap_loader.cpp:120-171 (images.bin)"""
    try:
        f = open(path_to_model_file, "rb")
    except OSError:
        raise RuntimeError("Unable to open file: " + path_to_model_file)
    images: Dict[int, Image] = {}
    with f:
        (num_reg_images,) = _read(f, "Q")
        for _ in range(num_reg_images):
            props = _read(f, "idddddddi")
            image_id, qvec, tvec, camera_id = props[0], np.array(props[1:5]), np.array(props[5:8]), props[8]
            name = bytearray()
            while True:
                c = f.read(1)
                if not c:
                    raise RuntimeError("unexpected end of COLMAP model file")
                if c == b"\x00":
                    break
                name += c
            (num_points2D,) = _read(f, "Q")
            raw = f.read(24 * num_points2D)
            if len(raw) != 24 * num_points2D:
                raise RuntimeError("unexpected end of COLMAP model file")
            rec = np.frombuffer(raw, dtype=np.dtype([("x", "<f8"), ("y", "<f8"), ("id", "<i8")]))
            images[image_id] = Image(image_id, qvec, tvec, camera_id, name.decode("utf-8"),
                                     np.stack([rec["x"], rec["y"]], axis=1) if num_points2D else np.zeros((0, 2)),
                                     rec["id"].copy())
    return images


def read_intrinsics_binary(path_to_model_file: str) -> Dict[int, ColmapCamera]:
    """src/scene/colmap_loader.cpp:222-250 (cameras.bin)"""
    try:
        f = open(path_to_model_file, "rb")
    except OSError:
        raise RuntimeError("Unable to open file: " + path_to_model_file)
    cameras: Dict[int, ColmapCamera] = {}
    with f:
        (num_cameras,) = _read(f, "Q")
        for _ in range(num_cameras):
            camera_id, model_id, width, height = _read(f, "iiQQ")
            if model_id not in CAMERA_MODELS:          # (std::unordered_map::at throws in the reference)
                raise RuntimeError(f"unknown COLMAP camera model id {model_id}")
            model, nparams = CAMERA_MODELS[model_id]
            params = np.array(_read(f, "d" * nparams))
            cameras[camera_id] = ColmapCamera(camera_id, model, width, height, params)
    return cameras


def qvec2rotmat(qvec) -> np.ndarray:
    """src/scene/colmap_loader.cpp:265-280 (no normalisation, exactly as there)"""
    q = np.asarray(qvec, dtype=np.float64)
    return np.array([
        [1 - 2 * q[2] * q[2] - 2 * q[3] * q[3], 2 * q[1] * q[2] - 2 * q[0] * q[3], 2 * q[3] * q[1] + 2 * q[0] * q[2]],
        [2 * q[1] * q[2] + 2 * q[0] * q[3], 1 - 2 * q[1] * q[1] - 2 * q[3] * q[3], 2 * q[2] * q[3] - 2 * q[0] * q[1]],
        [2 * q[3] * q[1] - 2 * q[0] * q[2], 2 * q[2] * q[3] + 2 * q[0] * q[1], 1 - 2 * q[1] * q[1] - 2 * q[2] * q[2]]])


@dataclass
class CameraInfo:
    """src/scene/dataset_readers.cpp:13-35"""
    uid_: int
    R_: np.ndarray
    T_: np.ndarray
    FovY_: float
    FovX_: float
    image_: Optional[np.ndarray]
    image_path_: str
    image_name_: str
    width_: int
    height_: int

    def to_camera(self) -> Camera:
        return Camera(self.R_, self.T_, self.FovX_, self.FovY_, self.height_, self.width_, uid=self.uid_,
                      image_name=self.image_name_)


def read_colmap_cameras(cam_extrinsics: Dict[int, Image], cam_intrinsics: Dict[int, ColmapCamera], images_folder: str,
                        load_image: Optional[Callable[[str], np.ndarray]] = None) -> List[CameraInfo]:
    """src/scene/dataset_readers.cpp:38-93: R is the TRANSPOSED COLMAP rotation (camera-to-world), T the COLMAP
    translation; only undistorted models (PINHOLE / SIMPLE_PINHOLE)."""
    cam_infos = []
    for extr in cam_extrinsics.values():
        intr = cam_intrinsics[extr.camera_id_]
        height, width, uid = intr.height_, intr.width_, intr.id_
        R = qvec2rotmat(extr.qvec_).T
        T = np.array(extr.tvec_)
        if intr.model_ == "SIMPLE_PINHOLE":
            fx = intr.params_[0]
            FovY, FovX = focal2fov(fx, height), focal2fov(fx, width)
        elif intr.model_ == "PINHOLE":
            fx, fy = intr.params_[0], intr.params_[1]
            FovY, FovX = focal2fov(fy, height), focal2fov(fx, width)
        else:
            raise RuntimeError("Colmap camera model not handled: only undistorted datasets (PINHOLE or SIMPLE_PINHOLE "
                               "cameras) supported!")
        image_path = os.path.join(images_folder, os.path.basename(extr.name_))
        image_name = os.path.splitext(os.path.basename(image_path))[0]
        image = load_image(image_path) if load_image is not None else None
        cam_infos.append(CameraInfo(uid, R, T, FovY, FovX, image, image_path, image_name, width, height))
    return cam_infos


def get_center_and_diag(cam_centers) -> Tuple[np.ndarray, float]:
    """src/scene/dataset_readers.cpp:96-110 ("diag" is the largest distance from the mean centre)"""
    c = np.asarray(cam_centers, dtype=np.float64).reshape(-1, 3)
    center = c.sum(axis=0) / float(len(c))
    return center, float(np.linalg.norm(c - center, axis=1).max()) if len(c) else 0.0


def get_nerfpp_norm(train_cam_infos: List[CameraInfo]) -> dict:
    """src/scene/dataset_readers.cpp:111-127: {"translate": -centre, "radius": 1.1 * diag}; radius is the
    cameras_extent that scales the position learning rate (spatial_lr_scale) and the densification thresholds."""
    centers = []
    for cam in train_cam_infos:
        C2W = np.linalg.inv(get_world2view_2(cam.R_, cam.T_))
        centers.append(C2W[:3, 3])
    center, diagonal = get_center_and_diag(centers)
    return {"translate": -center, "radius": diagonal * 1.1}


@dataclass
class SceneInfo:
    train_cameras: List[CameraInfo]
    test_cameras: List[CameraInfo]
    nerf_normalization: dict = field(default_factory=dict)


def read_colmap_scene_info(path: str, images: str = "", eval: bool = False, llffhold: int = 8,
                           load_image: Optional[Callable[[str], np.ndarray]] = None) -> SceneInfo:
    """src/scene/dataset_readers.cpp:140-198: binary model under <path>/sparse/0, cameras sorted by image name,
    every llffhold-th one held out when eval; the text model is "Not implemented", as in the reference."""
    try:
        cam_extrinsics = read_extrinsics_binary(os.path.join(path, "sparse", "0", "images.bin"))
        cam_intrinsics = read_intrinsics_binary(os.path.join(path, "sparse", "0", "cameras.bin"))
    except RuntimeError:
        raise RuntimeError("Not implemented")
    reading_dir = images if images else "images"
    cam_infos = sorted(read_colmap_cameras(cam_extrinsics, cam_intrinsics, os.path.join(path, reading_dir), load_image),
                       key=lambda c: c.image_name_)
    if eval:
        train = [c for i, c in enumerate(cam_infos) if i % llffhold != 0]
        test = [c for i, c in enumerate(cam_infos) if i % llffhold == 0]
    else:
        train, test = cam_infos, []
    return SceneInfo(train, test, get_nerfpp_norm(train))
