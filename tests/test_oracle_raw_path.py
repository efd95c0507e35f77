"""CPU: the oracle's statement of the RAW-parameter path (helpers.raw_reference -- what the GPU tests compare the
training path with) against the activated-input path with the getters' Jacobians applied by hand
(src/scene/gaussian_model.cpp:54-59, :270-298), and the specified exp against libm."""
import numpy as np

from helpers import oracle_settings, raw_reference, rel_err, small_scene


def test_specified_exp_and_sigmoid(oracle):
    x = np.linspace(-30, 30, 200001).astype(np.float32)
    sc, rot, op = oracle.activate_raw(np.stack([x, x, x], 1), np.ones((x.size, 4), np.float32), x)
    ref = np.exp(x.astype(np.float64))
    assert np.abs(sc[:, 0] / ref - 1).max() < 2e-6          # argument reduction in fp32: |x| ulp(x log2 e) ln 2
    near = np.abs(x) < 12                                   # the range log-scales and logits live in
    assert np.abs(sc[near, 0] / ref[near] - 1).max() < 1.2e-6
    assert np.abs(op[:, 0] - 1 / (1 + np.exp(-x.astype(np.float64)))).max() < 2e-7
    # whole-range behaviour: flush below 2^-126, +inf above 2^128, rotation passed through
    e, _, o = oracle.activate_raw(np.float32([[-100, 0, 95]]), np.float32([[3, 0, 4, 0]]), np.float32([[-200.0]]))
    assert e[0, 0] == 0 and e[0, 1] == 1 and np.isinf(e[0, 2]) and o[0, 0] == 0
    assert np.array_equal(rot[0], [1, 1, 1, 1])


def test_raw_reference_equals_activated_path_with_manual_jacobians(oracle):
    scene, cam = small_scene(N=1500, H=64, W=96, deg=2, seed=12, scale_boost=0.5)
    rng = np.random.default_rng(0)
    dL = ((rng.random((3, 64, 96)) * 2 - 1) / (64 * 96)).astype(np.float32)
    r = raw_reference(oracle, scene, cam, 2, 2, dL=dL)
    sc, rot, op, shs = scene.activated()
    so = oracle_settings(cam, 2, 2)
    pre, b, fwd = oracle.forward(so, scene.xyz, sc, rot, op, shs)
    g = oracle.backward(so, scene.xyz, sc, rot, shs, pre, b, fwd, dL)
    assert np.abs(fwd["out_color"] - r["fwd"]["out_color"]).max() < 2e-4     # activations differ by ulps
    qn = np.linalg.norm(scene.raw_rotations, axis=1, keepdims=True)
    g_rq = (g["dL_drotations"] - rot * (rot * g["dL_drotations"]).sum(1, keepdims=True)) / qn
    gr = r["g_ref"]
    same = pre["radii"] == r["pre"]["radii"]
    assert same.mean() > 0.995
    assert rel_err(gr["dL_drotations"][same], g_rq[same]) < 5e-3
    assert rel_err(gr["dL_dscales"][same], (g["dL_dscales"] * sc)[same]) < 5e-3
    assert rel_err(gr["dL_dopacities"][same], (g["dL_dopacities"].reshape(-1, 1) * op * (1 - op))[same]) < 5e-3
    assert rel_err(gr["dL_dmeans3D"][same], g["dL_dmeans3D"][same]) < 5e-3
