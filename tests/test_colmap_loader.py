"""COLMAP ingestion (3d_gaussian_splatting_b200/colmap.py) against the reference's own unit-test values
(src/scene/colmap_loader.cpp:297-326 qvec2rotmat; src/scene/dataset_readers.cpp:232-255 get_center_and_diag) and
against model files written here in COLMAP's binary layout (the reference's fixtures are private datasets)."""
import importlib
import math
import struct

import numpy as np
import pytest

cm = importlib.import_module("3d_gaussian_splatting_b200.colmap")


def test_qvec2rotmat_reference_values():
    r = cm.qvec2rotmat([1.0, 2.0, 3.0, 4.0])          # the reference's test vector (not normalised)
    np.testing.assert_array_equal(r, [[-49, 4, 22], [20, -39, 20], [10, 28, -25]])
    q = np.array([0.5, -0.5, 0.5, 0.5])
    R = cm.qvec2rotmat(q)
    np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-15)


def test_get_center_and_diag_reference_values():
    center, diag = cm.get_center_and_diag([[i, i, i] for i in range(10)])
    np.testing.assert_array_equal(center, [4.5, 4.5, 4.5])
    assert diag == pytest.approx(math.sqrt(3.0 * 4.5 * 4.5), rel=1e-6)


def write_model(root, n_images=11, model_id=1):
    sp = root / "sparse" / "0"
    sp.mkdir(parents=True)
    rng = np.random.default_rng(0)
    with open(sp / "cameras.bin", "wb") as f:
        f.write(struct.pack("<Q", 1))
        params = ([400.0, 410.0, 256.0, 192.0] + [0.0] * 8)[:cm.CAMERA_MODELS[model_id][1]]
        f.write(struct.pack("<iiQQ", 1, model_id, 512, 384) + struct.pack(f"<{len(params)}d", *params))
    poses = []
    with open(sp / "images.bin", "wb") as f:
        f.write(struct.pack("<Q", n_images))
        for i in range(n_images, 0, -1):                      # written in reverse: the reader must sort by name
            q = rng.standard_normal(4); q /= np.linalg.norm(q)
            t = rng.standard_normal(3) * 2
            npts = 792 if i == 1 else int(rng.integers(0, 5))
            f.write(struct.pack("<idddddddi", i, *q, *t, 1) + f"sub/Image_{i:06d}.jpg".encode() + b"\x00")
            f.write(struct.pack("<Q", npts))
            for _ in range(npts):
                f.write(struct.pack("<ddq", *rng.random(2), int(rng.integers(-1, 1000))))
            poses.append((i, q, t))
    return poses


def test_read_binary_model_and_scene_info(tmp_path):
    poses = write_model(tmp_path)
    ext = cm.read_extrinsics_binary(str(tmp_path / "sparse" / "0" / "images.bin"))
    intr = cm.read_intrinsics_binary(str(tmp_path / "sparse" / "0" / "cameras.bin"))
    assert len(ext) == 11 and len(intr) == 1
    im = ext[1]                                               # the shape of the reference's assertions on image 1
    assert im.id_ == 1 and im.camera_id_ == 1 and im.name_ == "sub/Image_000001.jpg"
    assert im.xys_.shape == (792, 2) and im.point3D_ids_.shape == (792,)
    cam = intr[1]
    assert cam.model_ == "PINHOLE" and cam.width_ == 512 and cam.height_ == 384 and len(cam.params_) == 4
    for i, q, t in poses:
        np.testing.assert_array_equal(ext[i].qvec_, q); np.testing.assert_array_equal(ext[i].tvec_, t)
    info = cm.read_colmap_scene_info(str(tmp_path), eval=True, llffhold=4)
    names = [c.image_name_ for c in info.train_cameras + info.test_cameras]
    assert sorted(names) == [f"Image_{i:06d}" for i in range(1, 12)]
    assert [c.image_name_ for c in info.test_cameras] == ["Image_000001", "Image_000005", "Image_000009"]
    c0 = info.test_cameras[0]
    np.testing.assert_allclose(c0.R_, cm.qvec2rotmat(ext[1].qvec_).T)
    assert c0.FovX_ == pytest.approx(2 * math.atan(512 / (2 * 400.0))) and c0.FovY_ == pytest.approx(2 * math.atan(384 / (2 * 410.0)))
    assert c0.image_path_.endswith("images/Image_000001.jpg") and c0.image_ is None
    # NeRF++ normalisation: camera centres are -R_colmap^T t
    centers = np.array([-(cm.qvec2rotmat(ext[int(c.image_name_[-6:])].qvec_).T @ ext[int(c.image_name_[-6:])].tvec_)
                        for c in info.train_cameras])
    center = centers.mean(axis=0)
    np.testing.assert_allclose(info.nerf_normalization["translate"], -center, atol=1e-12)
    assert info.nerf_normalization["radius"] == pytest.approx(1.1 * np.linalg.norm(centers - center, axis=1).max())
    assert c0.to_camera().image_width == 512                  # feeds the Camera the rasterizer settings come from


def test_errors_match_the_reference(tmp_path):
    with pytest.raises(RuntimeError, match="Unable to open file"):
        cm.read_extrinsics_binary(str(tmp_path / "missing.bin"))
    with pytest.raises(RuntimeError, match="Not implemented"):
        cm.read_colmap_scene_info(str(tmp_path))               # no binary model: the text path is a stub there too
    write_model(tmp_path, n_images=2, model_id=4)               # OPENCV: a distorted model
    with pytest.raises(RuntimeError, match="only undistorted datasets"):
        cm.read_colmap_scene_info(str(tmp_path))
    (tmp_path / "t").mkdir()
    write_model(tmp_path / "t", n_images=3)
    blob = (tmp_path / "t" / "sparse" / "0" / "images.bin").read_bytes()
    (tmp_path / "t" / "sparse" / "0" / "images.bin").write_bytes(blob[:-5])
    with pytest.raises(RuntimeError):
        cm.read_extrinsics_binary(str(tmp_path / "t" / "sparse" / "0" / "images.bin"))
