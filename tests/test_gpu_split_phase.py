"""The split-phase multi-GPU step (parallel.PeerShardedAdam, include/gsplat_b200.h: gs_set_color_gate,
GS_PEER_BESIDE_FORWARD) on ONE device: (a) a forward with the colour gate -- K1 without SH -> RGB, colours filled in
by sh_color_kernel ahead of the blend -- equals the one-kernel forward bit for bit, image, intermediates and every
gradient; (b) gs_adam_step_peer called once per phase with the other phase's groups masked out (widths 0), the SH
phase with a capped grid, equals the single call bit for bit."""
import importlib

import numpy as np
import pytest

from helpers import gpu_settings, small_scene

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("N,deg,maxdeg", [(3000, 2, 2), (2048, 3, 3), (1000, 0, 3), (640, 1, 1), (100, 3, 3)])
def test_colour_gate_forward_backward_bit_identical(rasterizer, N, deg, maxdeg):
    """N = 3000 / 1000 / 100: a ragged tail (generic kernels) behind the TMA-staged blocks; N = 100: generic only."""
    scene, cam = small_scene(N=N, H=96, W=128, deg=maxdeg, seed=17 + N, scale_boost=0.5)
    sg = gpu_settings(cam, deg, maxdeg, bg=(0.1, 0.2, 0.3))
    K = (maxdeg + 1) ** 2
    P = [dev(np.ascontiguousarray(a, dtype=np.float32)) for a in
         (scene.xyz, scene.features_dc, scene.features_rest[:, :K - 1], scene.opacity_logits, scene.log_scales,
          scene.raw_rotations)]
    rng = np.random.default_rng(5)
    dL = dev(((rng.random((3, 96, 128)) * 2 - 1) / (96 * 128)).astype(np.float32))
    def run():
        color, radii = rasterizer.forward(sg, P[0], P[4], P[5], P[3], sh_dc=P[1], sh_rest=P[2] if K > 1 else None,
                                          raw_params=True)
        color, radii = color.clone(), radii.clone()
        inter = {k: rasterizer.debug_get(k).copy() for k in
                 ("RGB", "CLAMPED", "CONIC_OPACITY", "MEANS2D", "POINT_LIST", "RANGES", "N_CONTRIB", "FINAL_T")}
        # (the blend backward's atomics are order-dependent: gradients are compared to atomic noise; K4b's inputs that
        #  the gate touches -- colours and clamp masks -- are compared exactly through the debug arrays)
        g = {k: v.clone() for k, v in rasterizer.backward(dL).items()}
        return color, radii, inter, g

    base = run()
    ev = torch.cuda.Event()
    ev.record()
    rasterizer.set_color_gate(ev)
    try:
        gated = run()
    finally:
        rasterizer.set_color_gate(None)
    again = run()
    for other in (gated, again):
        assert torch.equal(base[0], other[0]) and torch.equal(base[1], other[1])
        for k in base[2]:
            np.testing.assert_array_equal(base[2][k], other[2][k], err_msg=k)
        for k in base[3]:
            a, b = base[3][k], other[3][k]
            scale = a.abs().max().item() + 1e-30
            assert (a - b).abs().max().item() <= 1e-3 * scale, k      # (atomic order of the blend backward; the bar of the parity tests)
    assert (base[2]["RGB"] != 0).any() and base[1].gt(0).any()


def test_colour_gate_waits_for_the_event(rasterizer):
    """The colour kernel runs after the gate event: SH coefficients written on a SECOND stream, behind a long-running
    kernel there, reach the image; without the wait the forward would blend the stale (zero) coefficients."""
    N = 2048
    scene, cam = small_scene(N=N, H=96, W=128, deg=1, seed=23, scale_boost=0.5)
    sg = gpu_settings(cam, 1, 1)
    P = [dev(np.ascontiguousarray(a, dtype=np.float32)) for a in
         (scene.xyz, scene.features_dc, scene.features_rest[:, :3], scene.opacity_logits, scene.log_scales,
          scene.raw_rotations)]
    fw = lambda dc, rest: rasterizer.forward(sg, P[0], P[4], P[5], P[3], sh_dc=dc, sh_rest=rest, raw_params=True)[0].clone()
    want = fw(P[1], P[2])
    dc, rest = torch.zeros_like(P[1]), torch.zeros_like(P[2])
    stale = fw(dc, rest)
    assert not torch.equal(want, stale)
    side, ev = torch.cuda.Stream(), torch.cuda.Event()
    ev.record()
    rasterizer.set_color_gate(ev)
    try:
        torch.cuda.synchronize()
        with torch.cuda.stream(side):
            big = torch.empty(1 << 28, device="cuda")
            for _ in range(20):
                big.fill_(1.0)                   # ~20 x 1 GiB of writes: several milliseconds on the side stream
            dc.copy_(P[1])
            rest.copy_(P[2])
            ev.record(side)
        got = fw(dc, rest)                       # enqueued at once on the main stream; K1 / binning run meanwhile
        torch.cuda.synchronize()
    finally:
        rasterizer.set_color_gate(None)
    assert torch.equal(got, want)


def test_split_phase_peer_adam_equals_single_call(pkg, monkeypatch):
    """Three virtual ranks on one device: geometry groups + statistics in one call, SH groups in a second call with the
    capped grid (GS_PEER_BESIDE_FORWARD; capped at 3 blocks here so that the grid-stride loop really loops) == all six
    groups in one call, bit for bit, replicas, moments, slots."""
    monkeypatch.setenv("GS_PEER_ADAM_GRID", "3")      # read when the context is created
    rasterizer = pkg.Rasterizer(0)
    try:
        _split_phase_adam(rasterizer)
    finally:
        rasterizer.close()


def _split_phase_adam(rasterizer):
    G, N = 3, 3000
    R = ((N + G - 1) // G + 127) // 128 * 128
    widths = [3, 3, 24, 1, 3, 4]
    geo = [3, 0, 0, 1, 3, 4]
    sh = [0, 3, 24, 0, 0, 0]
    lr = [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3]
    rng = np.random.default_rng(11)
    P0 = [dev(rng.standard_normal((N, w)).astype(np.float32)) for w in widths]
    grads = [[(rng.standard_normal((N, w)) * 1e-3).astype(np.float32) for w in widths] for _ in range(G)]
    for s_ in range(G):
        grads[s_][2][rng.random(N) < 0.6] = 0.0       # sparse f_rest rows, as the routed backward leaves them
    radii = dev((rng.random(N) < 0.7).astype(np.int32) * 5)
    g2 = dev(rng.standard_normal((N, 2)).astype(np.float32))

    def world(split):
        reps = [[p.clone() for p in P0] for _ in range(G)]
        m1 = [[torch.zeros(R * w, device="cuda") for w in widths] for _ in range(G)]
        m2 = [[torch.zeros(R * w, device="cuda") for w in widths] for _ in range(G)]
        stats = [(torch.zeros(N, dtype=torch.int32, device="cuda"), torch.zeros(N, device="cuda"),
                  torch.zeros(N, device="cuda")) for _ in range(G)]
        slots_out = []
        for step in (1, 2, 3):
            for r in range(G):
                lo, hi = r * R, min((r + 1) * R, N)
                sl = [[torch.zeros(R, w, device="cuda") for w in widths] for _ in range(G)]
                for s_ in range(G):
                    for k in range(6):
                        sl[s_][k][:hi - lo] = dev(grads[s_][k][lo:hi]) * step
                pp = [[t.data_ptr() for t in rep] for rep in reps]
                sp = [[t.data_ptr() for t in row] for row in sl]
                st = dict(radii=radii, dL_dmeans2D=g2, max_radii2D=stats[r][0], xyz_grad_accum=stats[r][1],
                          denom=stats[r][2])
                if split == "ce":
                    # SH groups into the own replica only (GS_PEER_SH_LOCAL), then copy-engine copies of the owned rows
                    rasterizer.adam_step_peer(G, r, N, R, pp, sp, m1[r], m2[r], widths, lr, [step] * 6, sh_local=True, **st)
                    for k in (1, 2):
                        off = 4 * lo * widths[k]
                        for q in range(G):
                            if q != r:
                                rasterizer.peer_copy(pp[q][k] + off, pp[r][k] + off, 4 * (hi - lo) * widths[k])
                elif split:
                    rasterizer.adam_step_peer(G, r, N, R, pp, sp, m1[r], m2[r], sh, lr, [step] * 6, beside_forward=True)
                    rasterizer.adam_step_peer(G, r, N, R, pp, sp, m1[r], m2[r], geo, lr, [step] * 6, **st)
                else:
                    rasterizer.adam_step_peer(G, r, N, R, pp, sp, m1[r], m2[r], widths, lr, [step] * 6, **st)
                slots_out.append(sl)
        torch.cuda.synchronize()
        return reps, m1, m2, stats, slots_out

    a, b = world(False), world(True)
    _same_world(a, b)
    _same_world(a, world("ce"))
    assert not all(torch.equal(x, y) for x, y in zip(a[0][0], P0))      # the steps did something


def _same_world(a, b):
    for ra, rb in zip(a[0], b[0]):
        assert all(torch.equal(x, y) for x, y in zip(ra, rb))
    for q in (1, 2):
        for ra, rb in zip(a[q], b[q]):
            assert all(torch.equal(x, y) for x, y in zip(ra, rb))
    for sa, sb in zip(a[3], b[3]):
        assert all(torch.equal(x, y) for x, y in zip(sa, sb))
    for sa, sb in zip(a[4], b[4]):
        for rowa, rowb in zip(sa, sb):
            assert all(torch.equal(x, y) for x, y in zip(rowa, rowb))


def test_tiny_first_moments_fused_and_peer_kernels(rasterizer):
    """First moments below 2^-100 (Gaussians unseen for ~480 iterations and more; they end up stuck at the smallest
    denormal) have a packed path of their own in the fused K4b+K5 kernel and in the peer Adam kernel
    (GS_ADAM_FAST_T, csrc/gs_common.cuh).  Both must give the bits of gs_adam_densify_step, whose kernel keeps such
    vectors on the scalar IEEE path (itself pinned to the oracle by test_adam_packed_fast_path_is_ieee).  Whole row
    blocks are tiny so that whole warps take the path; large second moments make some quotients denormal (the
    scalar fallback inside the path), zero gradients exercise the pure decay."""
    L = importlib.import_module("3d_gaussian_splatting_b200._lib")
    from test_gpu_parity import seeded_dL
    N, deg = 3000, 2
    # (the camera sits inside the cloud: the Gaussians behind it are culled and receive no gradient -- pure decay)
    scene, cam = small_scene(N=N, H=96, W=128, deg=deg, seed=43, scale_boost=0.5, eye=(0.1, 0.0, 0.2))
    K = (deg + 1) ** 2
    widths = [3, 3, 3 * (K - 1), 1, 3, 4]
    sg = gpu_settings(cam, deg, deg)
    P = [dev(np.ascontiguousarray(a, dtype=np.float32)) for a in
         (scene.xyz, scene.features_dc, scene.features_rest[:, :K - 1], scene.opacity_logits, scene.log_scales,
          scene.raw_rotations)]
    lr = [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3]
    rng = np.random.default_rng(29)

    def moments():
        M, V = [], []
        for p in P:
            shp = tuple(p.shape)
            m = (rng.standard_normal(shp) * 1e-4).astype(np.float32)
            v = (rng.random(shp) * 1e-8 + 1e-16).astype(np.float32)
            tiny = ((10.0 ** rng.uniform(-45, -31, shp)) * rng.choice([-1.0, 1.0], shp)).astype(np.float32)
            m[:2000] = tiny[:2000]
            m[:2000:5] = np.float32(1e-45)
            v[:1000] = (10.0 ** rng.uniform(-18, -13, shp)).astype(np.float32)[:1000]
            v[1000:1500] = (10.0 ** rng.uniform(-6, 2, shp)).astype(np.float32)[1000:1500]      # denormal quotients
            M.append(dev(m)); V.append(dev(v))
        return M, V

    # ---- fused K4b+K5 against gs_backward + gs_adam_densify_step on the same blend-gradient records ----
    M, V = moments()
    _, radii = rasterizer.forward(sg, P[0], P[4], P[5], P[3], sh_dc=P[1], sh_rest=P[2], raw_params=True)
    dL = dev(seeded_dL(96, 128, 5))
    G = rasterizer.backward(dL)
    Gk = [G["means3D"], G["sh_dc"], G["sh_rest"], G["opacities"], G["scales"], G["rotations"]]
    assert sum(int((g[:2000] == 0).all(dim=-1).sum()) for g in Gk) > 0          # pure-decay rows among the tiny ones
    A, MA, VA = [p.clone() for p in P], [m.clone() for m in M], [v.clone() for v in V]
    rasterizer.adam_densify_step(N, A, Gk, MA, VA, widths, lr, [7] * 6)
    PF, MF, VF = [p.clone() for p in P], [m.clone() for m in M], [v.clone() for v in V]
    L.check(rasterizer._ctx, L.lib.gs_debug_reuse_blend_gradients(rasterizer._ctx, 1))
    saved = [p.clone() for p in P]
    try:
        rasterizer.backward_adam_step(dL, MF, VF, lr, [7] * 6)       # updates P in place
    finally:
        L.check(rasterizer._ctx, L.lib.gs_debug_reuse_blend_gradients(rasterizer._ctx, 0))
    torch.cuda.synchronize()
    for k in range(6):
        assert torch.equal(P[k], A[k]), ("fused param", k)
        assert torch.equal(MF[k], MA[k]) and torch.equal(VF[k], VA[k]), ("fused moments", k)
        assert not torch.equal(P[k], saved[k])
        P[k].copy_(saved[k])
    # ---- peer Adam kernel, one rank: slot 0 holds the gradients ----
    R = (N + 127) // 128 * 128
    M, V = moments()
    A, MA, VA = [p.clone() for p in P], [m.clone() for m in M], [v.clone() for v in V]
    grads = [dev((rng.standard_normal(tuple(p.shape)) * 1e-3 * (rng.random(tuple(p.shape)) < 0.5)).astype(np.float32)) for p in P]
    rasterizer.adam_densify_step(N, A, grads, MA, VA, widths, lr, [9] * 6)
    slots = [torch.zeros(R * w, device="cuda") for w in widths]
    mp = [torch.zeros(R * w, device="cuda") for w in widths]
    vp = [torch.zeros(R * w, device="cuda") for w in widths]
    for k, w in enumerate(widths):
        slots[k][:N * w] = grads[k].reshape(-1)
        mp[k][:N * w] = M[k].reshape(-1)
        vp[k][:N * w] = V[k].reshape(-1)
    rep = [p.clone() for p in P]
    rasterizer.adam_step_peer(1, 0, N, R, [[t.data_ptr() for t in rep]], [[t.data_ptr() for t in slots]], mp, vp, widths,
                              lr, [9] * 6, sparse_rest=False)
    torch.cuda.synchronize()
    for k, w in enumerate(widths):
        assert torch.equal(rep[k], A[k]), ("peer param", k)
        assert torch.equal(mp[k][:N * w], MA[k].reshape(-1)) and torch.equal(vp[k][:N * w], VA[k].reshape(-1)), ("peer moments", k)
