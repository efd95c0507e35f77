"""The densify / prune / opacity-reset oracle (oracle/orc_densify.c) against an independent numpy restatement
of the published 3DGS algorithm (densify_and_clone -> densify_and_split -> prune_points; reset_opacity), phase by
phase with explicit concatenations and masks as the published code does it."""
import numpy as np
import pytest

from oracle.oracle_lib import Oracle

WIDTHS = [(3,), (1, 3), (8, 3), (1,), (3,), (4,)]


def make_state(N, seed):
    rng = np.random.default_rng(seed)
    P = [rng.standard_normal((N,) + w).astype(np.float32) for w in WIDTHS]
    P[4] = (rng.uniform(-6.0, -1.0, (N, 3))).astype(np.float32)          # log scales
    P[3] = rng.uniform(-7.0, 3.0, (N, 1)).astype(np.float32)             # opacity logits (some below 0.005)
    M = [(rng.standard_normal(p.shape) * 1e-3).astype(np.float32) for p in P]
    V = [(rng.random(p.shape) * 1e-6).astype(np.float32) for p in P]
    denom = rng.integers(0, 4, N).astype(np.float32)                      # zeros: never visible -> NaN -> 0
    accum = (rng.random(N) * 6e-4 * np.maximum(denom, 1)).astype(np.float32)
    accum[denom == 0] = 0
    maxr = rng.integers(0, 40, N).astype(np.int32)
    return P, M, V, accum, denom, maxr


def numpy_densify(o, P, M, V, accum, denom, thr, min_op, extent, pd, max_screen, seed):
    N = P[0].shape[0]
    with np.errstate(invalid="ignore", divide="ignore"):
        grads = accum / denom
    grads[np.isnan(grads)] = 0
    scale_max = lambda S: np.exp(S).max(axis=1)
    # clone
    sel = (grads >= thr) & (scale_max(P[4]) <= pd * extent)
    P1 = [np.concatenate([p, p[sel]]) for p in P]
    M1 = [np.concatenate([m, np.zeros_like(m[sel])]) for m in M]
    V1 = [np.concatenate([v, np.zeros_like(v[sel])]) for v in V]
    N1 = P1[0].shape[0]
    # split
    pg = np.zeros(N1, np.float32); pg[:N] = grads
    sel2 = (pg >= thr) & (scale_max(P1[4]) > pd * extent)
    idx = np.nonzero(sel2)[0]
    rep = np.concatenate([idx, idx])
    copy = np.concatenate([np.zeros_like(idx), np.ones_like(idx)])
    z = np.stack([o.split_normals(seed, int(i), int(c)) for i, c in zip(rep, copy)]) if len(rep) else np.zeros((0, 3), np.float32)
    sc = np.exp(P1[4][rep]).astype(np.float32)
    smp = (z * sc).astype(np.float32)
    q = P1[5][rep]
    q = q / np.linalg.norm(q, axis=1, keepdims=True)
    r, x, y, zz = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([1 - 2 * (y * y + zz * zz), 2 * (x * y - r * zz), 2 * (x * zz + r * y),
                  2 * (x * y + r * zz), 1 - 2 * (x * x + zz * zz), 2 * (y * zz - r * x),
                  2 * (x * zz - r * y), 2 * (y * zz + r * x), 1 - 2 * (x * x + y * y)], axis=1).reshape(-1, 3, 3)
    new_xyz = np.einsum("nij,nj->ni", R, smp) + P1[0][rep]
    new_scl = np.log(sc / (0.8 * 2))
    newP = [p[rep].copy() for p in P1]
    newP[0], newP[4] = new_xyz.astype(np.float32), new_scl.astype(np.float32)
    P2 = [np.concatenate([p, n]) for p, n in zip(P1, newP)]
    M2 = [np.concatenate([m, np.zeros_like(m[rep])]) for m in M1]
    V2 = [np.concatenate([v, np.zeros_like(v[rep])]) for v in V1]
    prune = np.concatenate([sel2, np.zeros(len(rep), bool)])
    keep = ~prune
    P3, M3, V3 = [p[keep] for p in P2], [m[keep] for m in M2], [v[keep] for v in V2]
    # prune (statistics were zeroed by the postfix: the screen-size test never fires)
    mask = (1 / (1 + np.exp(-P3[3].reshape(-1)))) < min_op
    if max_screen > 0:
        mask |= scale_max(P3[4]) > 0.1 * extent
    keep = ~mask
    return [p[keep] for p in P3], [m[keep] for m in M3], [v[keep] for v in V3]


@pytest.mark.parametrize("N,seed,max_screen", [(500, 1, 0.0), (3000, 2, 20.0), (64, 3, 20.0)])
def test_densify_matches_numpy_restatement(N, seed, max_screen):
    o = Oracle()
    P, M, V, accum, denom, maxr = make_state(N, seed)
    extent, pd, thr, min_op = 3.0, 0.01, 2e-4, 0.005
    if max_screen > 0:
        P[4][::17] += 6.0      # a few Gaussians larger than 0.1 * extent
    got_P, got_M, got_V, counts = o.densify_and_prune(P, M, V, accum, denom, maxr, thr, min_op, extent, pd, max_screen, seed=42)
    ref_P, ref_M, ref_V = numpy_densify(o, P, M, V, accum, denom, thr, min_op, extent, pd, max_screen, 42)
    assert got_P[0].shape[0] == ref_P[0].shape[0] == counts[0] + counts[1] + 2 * counts[2]
    assert counts[1] > 0 and counts[2] > 0 and counts[3] > 0, counts     # every branch exercised
    for k in range(6):
        tol = dict(rtol=2e-6, atol=2e-6) if k in (0, 4) else dict(rtol=0, atol=0)   # only split xyz / scaling are computed
        np.testing.assert_allclose(got_P[k], ref_P[k], **tol)
        np.testing.assert_array_equal(got_M[k], ref_M[k])
        np.testing.assert_array_equal(got_V[k], ref_V[k])


def test_split_normals_are_standard_normal_and_order_free():
    o = Oracle()
    z = np.stack([o.split_normals(7, i, c) for i in range(4000) for c in (0, 1)])
    assert abs(z.mean()) < 0.03 and abs(z.std() - 1) < 0.03
    assert abs(np.corrcoef(z[:, 0], z[:, 1])[0, 1]) < 0.05 and abs(np.corrcoef(z[::2, 0], z[1::2, 0])[0, 1]) < 0.05
    np.testing.assert_array_equal(o.split_normals(7, 123, 1), o.split_normals(7, 123, 1))
    assert not np.array_equal(o.split_normals(7, 123, 1), o.split_normals(8, 123, 1))


def test_reset_opacity():
    o = Oracle()
    op = np.array([-8.0, -4.6, 0.0, 5.0], np.float32)
    m = np.ones(4, np.float32); v = np.ones(4, np.float32)
    want = np.log(np.minimum(1 / (1 + np.exp(-op.astype(np.float64))), 0.01) / (1 - np.minimum(1 / (1 + np.exp(-op.astype(np.float64))), 0.01)))
    o.reset_opacity(op, m, v)
    np.testing.assert_allclose(op, want, rtol=1e-5)
    assert not m.any() and not v.any()
