"""The reference's own known-answer tests, re-expressed as tests of
(a) the CPU oracle's convention functions and (b) the product's host-side C++ mirror
(csrc/host_conventions.cpp, reached through the C ABI).  These vectors are the only place the
reference pins anything on the render path: quaternion order, Sigma packing, activations,
world->view / projection layout, LR schedule, COLMAP qvec convention."""
import ctypes as C
import math

import numpy as np
import pytest

f32 = np.float32


def _rot(oracle, q):
    q = np.ascontiguousarray(q, dtype=f32)
    R = np.zeros((q.shape[0], 3, 3), f32)
    oracle.lib.orc_build_rotation(q.shape[0], oracle.p(q), oracle.p(R))
    return R


def test_build_rotation_golden(oracle):
    """src/utils/general_utils.cpp:147-187: q=(.5,.5,.5,.5) and the un-normalised (.25,.25,.25,.25)
    both give R=[[0,0,1],[1,0,0],[0,1,0]] (build_rotation normalises internally)."""
    R = _rot(oracle, [[0.5, 0.5, 0.5, 0.5], [0.25, 0.25, 0.25, 0.25]])
    want = np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0]], f32)
    np.testing.assert_allclose(R[0], want, atol=1e-6)
    np.testing.assert_allclose(R[1], want, atol=1e-6)


def test_build_scaling_rotation_golden(oracle):
    """src/utils/general_utils.cpp:189-241: L = R * diag(s) for s = .5 and .25."""
    q = np.array([[0.5, 0.5, 0.5, 0.5], [0.25, 0.25, 0.25, 0.25]], f32)
    s = np.array([[0.5, 0.5, 0.5], [0.25, 0.25, 0.25]], f32)
    L = np.zeros((2, 3, 3), f32)
    oracle.lib.orc_build_scaling_rotation(2, oracle.p(s), oracle.p(q), oracle.p(L))
    base = np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0]], f32)
    np.testing.assert_allclose(L[0], 0.5 * base, atol=1e-6)
    np.testing.assert_allclose(L[1], 0.25 * base, atol=1e-6)


def test_strip_lowerdiag_order_golden(oracle):
    """src/utils/general_utils.cpp:243-292: packing order [00,01,02,11,12,22]."""
    M = np.arange(18, dtype=f32).reshape(2, 3, 3)
    out = np.zeros((2, 6), f32)
    oracle.lib.orc_strip_symmetric(2, oracle.p(M), oracle.p(out))
    np.testing.assert_array_equal(out[0], [0, 1, 2, 4, 5, 8])
    np.testing.assert_array_equal(out[1], [9, 10, 11, 13, 14, 17])


def test_covariance_golden(oracle):
    """src/scene/gaussian_model.cpp:409-453 (s=.5 -> diag .25, s=.25 -> diag .0625, off-diag 0) and
    :552-562 (s=e^0, q=ones -> identity packed [1,0,0,1,0,1])."""
    q = np.array([[0.5, 0.5, 0.5, 0.5], [0.25, 0.25, 0.25, 0.25], [1, 1, 1, 1]], f32)
    s = np.array([[0.5] * 3, [0.25] * 3, [1.0] * 3], f32)
    cov = np.zeros((3, 6), f32)
    oracle.lib.orc_covariance(3, oracle.p(s), oracle.real(1.0), oracle.p(q), oracle.p(cov))
    np.testing.assert_allclose(cov[0], [0.25, 0, 0, 0.25, 0, 0.25], atol=1e-6)
    np.testing.assert_allclose(cov[1], [0.0625, 0, 0, 0.0625, 0, 0.0625], atol=1e-6)
    np.testing.assert_allclose(cov[2], [1, 0, 0, 1, 0, 1], atol=1e-6)


def test_activations_golden(oracle):
    """src/scene/gaussian_model.cpp:496-550: exp(0)=1, normalize(ones4)=.5, sigmoid(0)=.5."""
    ls = np.zeros((1, 3), f32); rq = np.ones((1, 4), f32); ol = np.zeros((1,), f32)
    s = np.zeros((1, 3), f32); r = np.zeros((1, 4), f32); o = np.zeros((1,), f32)
    oracle.lib.orc_activate(1, oracle.p(ls), oracle.p(rq), oracle.p(ol), oracle.p(s), oracle.p(r), oracle.p(o))
    np.testing.assert_allclose(s, 1.0); np.testing.assert_allclose(r, 0.5); np.testing.assert_allclose(o, 0.5)


def _w2v(oracle, R, t):
    out = np.zeros(16, np.float64)
    oracle.lib.orc_get_world2view(oracle.p(np.asarray(R, np.float64)), oracle.p(np.asarray(t, np.float64)), oracle.p(out))
    return out.reshape(4, 4)


def test_world2view_golden(oracle, pkg):
    """src/utils/graphics_utils.cpp:81-97: Rt[:3,:3] = R^T, Rt[:3,3] = t; :100-117: translating by
    (1,1,1) with identity R and t=0 gives Rt(1,3) = -1."""
    R = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1]], np.float64); t = np.array([1.0, 2.0, 3.0])
    Rt = _w2v(oracle, R, t)
    np.testing.assert_allclose(Rt[:3, :3], R.T); np.testing.assert_allclose(Rt[:3, 3], t); assert Rt[3, 3] == 1
    out = np.zeros(16, np.float64)
    I = np.eye(3); z = np.zeros(3); one = np.ones(3)
    oracle.lib.orc_get_world2view_2(oracle.p(I), oracle.p(z), oracle.p(one), C.c_double(1.0), oracle.p(out))
    assert out.reshape(4, 4)[1, 3] == pytest.approx(-1.0)
    # product host mirror
    np.testing.assert_allclose(pkg.get_world2view_2(I, z, one, 1.0), out.reshape(4, 4), atol=1e-15)
    np.testing.assert_allclose(pkg.get_world2view_2(R, t), Rt, atol=1e-15)


def test_projection_matrix_golden(oracle, pkg):
    """src/utils/graphics_utils.cpp:120-135: znear 1, zfar 10, fov pi/2 -> P00=P11=1, P02=P12=0,
    P32=1, P22=10/9, P23=-10/9."""
    for P in (pkg.get_projection_matrix(1.0, 10.0, math.pi / 2, math.pi / 2),):
        assert P[0, 0] == pytest.approx(1.0) and P[1, 1] == pytest.approx(1.0)
        assert P[0, 2] == 0 and P[1, 2] == 0 and P[3, 2] == 1.0
        assert P[2, 2] == pytest.approx(10.0 / 9.0) and P[2, 3] == pytest.approx(-10.0 / 9.0)
    out = np.zeros(16, np.float64)
    oracle.lib.orc_get_projection_matrix(C.c_double(1.0), C.c_double(10.0), C.c_double(math.pi / 2),
                                         C.c_double(math.pi / 2), oracle.p(out))
    np.testing.assert_allclose(out.reshape(4, 4), P, atol=1e-15)


def test_expon_lr_golden(oracle, pkg):
    """src/utils/general_utils.cpp:294-331: lr(-1)=0; lr(0)=0.05 with delay_mult .5 (delay active),
    0.1 without delay; and src/scene/gaussian_model.cpp:689-708: initial xyz lr 1.6e-4."""
    lib = oracle.lib
    assert lib.orc_expon_lr(0.1, 0.01, 10, 0.5, 100, -1) == 0.0
    assert lib.orc_expon_lr(0.1, 0.01, 10, 0.5, 100, 0) == pytest.approx(0.05, rel=1e-6)
    assert lib.orc_expon_lr(0.1, 0.01, 0, 0.5, 100, 0) == pytest.approx(0.1, rel=1e-6)
    assert lib.orc_expon_lr(1.6e-4, 1.6e-6, 0, 0.01, 30000, 0) == pytest.approx(1.6e-4, rel=1e-6)
    assert lib.orc_expon_lr(1.6e-4, 1.6e-6, 0, 0.01, 30000, 30000) == pytest.approx(1.6e-6, rel=1e-5)
    fn = pkg.get_expon_lr_func(0.1, 0.01, 10, 0.5, 100)
    for step in (-1, 0, 3, 10, 50, 100, 1000):
        assert fn(step) == lib.orc_expon_lr(0.1, 0.01, 10, 0.5, 100, step)


def test_qvec2rotmat_golden(oracle):
    """src/scene/colmap_loader.cpp:313-327: qvec (1,2,3,4), not normalised."""
    R = np.zeros(9, np.float64)
    oracle.lib.orc_qvec2rotmat(oracle.p(np.array([1.0, 2, 3, 4])), oracle.p(R))
    np.testing.assert_array_equal(R.reshape(3, 3), [[-49, 4, 22], [20, -39, 20], [10, 28, -25]])


def test_focal2fov(oracle, pkg):
    """src/utils/graphics_utils.cpp:4-7"""
    assert pkg.focal2fov(500.0, 1000) == pytest.approx(math.pi / 2)
    assert oracle.lib.orc_focal2fov(500.0, 1000.0) == pytest.approx(math.pi / 2)


def test_camera_tensors_match_oracle(oracle, pkg):
    """Camera ctor (src/scene/camera.cpp:44-45,66-71): product mirror vs oracle restatement, and the
    layout facts the kernels rely on (flat buffer = column-major V; campos = -R^T-rotated t)."""
    rng = np.random.default_rng(0)
    A = rng.standard_normal((3, 3)); Q, _ = np.linalg.qr(A)
    if np.linalg.det(Q) < 0:
        Q[:, 0] *= -1
    t = rng.standard_normal(3)
    fx, fy = 0.9, 0.7
    cam = pkg.Camera(Q, t, fx, fy, 480, 640)
    wv = np.zeros(16); pj = np.zeros(16); fp = np.zeros(16); cp = np.zeros(3)
    oracle.lib.orc_camera_tensors(oracle.p(Q.copy()), oracle.p(t), oracle.p(np.zeros(3)), C.c_double(1.0),
                                  C.c_double(0.01), C.c_double(100.0), C.c_double(fx), C.c_double(fy),
                                  oracle.p(wv), oracle.p(pj), oracle.p(fp), oracle.p(cp))
    np.testing.assert_allclose(cam.world_view_transform.reshape(-1), wv, rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(cam.full_proj_transform.reshape(-1), fp, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(cam.camera_center, cp, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(cam.camera_center, -Q @ t, rtol=1e-5, atol=1e-6)   # C = -R t (R is cam-to-world)
    V = cam.world_view_transform.T  # row-major V
    np.testing.assert_allclose(V[:3, :3], Q.T, atol=1e-6); np.testing.assert_allclose(V[:3, 3], t, atol=1e-6)
    assert cam.tanfovx == pytest.approx(math.tan(fx / 2), rel=1e-6)


def test_reference_arm_cameras_match_product_cameras(oracle):
    """bench.py --impl reference builds its cameras on the oracle's conventions (so that the CPU arm never loads
    libgsplat_b200.so); they must be the product's cameras (to rounding of the float64 intermediate)."""
    import importlib
    import sys
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        import bench
    finally:
        sys.argv = argv
    syn = importlib.import_module("3d_gaussian_splatting_b200.synthetic")
    a = syn.orbit_cameras(5, 1246, 825, 1000.0)
    b = syn.orbit_cameras(5, 1246, 825, 1000.0, look_at_camera=lambda e, t, u, f, H, W, uid=0:
                          bench.OracleCamera(oracle.lib, e, t, u, f, H, W, uid))
    for x, y in zip(a, b):
        np.testing.assert_allclose(x.world_view_transform, y.world_view_transform, rtol=0, atol=1e-7)
        np.testing.assert_allclose(x.full_proj_transform, y.full_proj_transform, rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(x.camera_center, y.camera_center, rtol=1e-6, atol=1e-7)
        assert x.tanfovx == y.tanfovx and x.tanfovy == y.tanfovy
        assert (x.image_height, x.image_width) == (y.image_height, y.image_width)
