"""K5 oracle pinned to torch.optim.Adam on CPU (the reference's optimizers are libtorch Adam with
defaults: src/scene/gaussian_model.cpp:323-345), and the densification statistics / L1 loss."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")


def test_adam_matches_torch(oracle):
    rng = np.random.default_rng(0)
    n = 4099
    p0 = rng.standard_normal(n).astype(np.float32)
    for lr in (1.6e-4, 0.05):
        p = p0.copy(); m = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
        tp = torch.nn.Parameter(torch.from_numpy(p0.copy()))
        opt = torch.optim.Adam([tp], lr=lr)            # defaults beta=(0.9,0.999), eps=1e-8, as libtorch
        gmax = np.zeros(n, np.float32)
        for step in range(1, 8):
            g = (rng.standard_normal(n) * 10 ** rng.uniform(-5, 0, n)).astype(np.float32)
            tp.grad = torch.from_numpy(g.copy())
            opt.step()
            oracle.adam_step(p, g, m, v, lr, 0.9, 0.999, 1e-8, step)
            st = opt.state[tp]
            # torch's Python Adam forms exp_avg with lerp_ (m + (g-m)*(1-b1)); libtorch C++ (the reference)
            # uses mul_/add_ (m*b1 + g*(1-b1)): identical up to one rounding of the larger operand
            ref_m = st["exp_avg"].numpy()
            gmax = np.maximum(gmax, np.abs(g))
            assert np.all(np.abs(m - ref_m) <= 2e-6 * np.abs(ref_m) + 3e-7 * gmax)
            np.testing.assert_allclose(v, st["exp_avg_sq"].numpy(), rtol=2e-6, atol=1e-20)
            np.testing.assert_allclose(p, tp.detach().numpy(), rtol=0, atol=4e-7 * max(1.0, lr * 100))


def test_densify_stats(oracle):
    radii = np.array([0, 3, 5, 0, 2], np.int32)
    g2 = np.array([[3, 4], [3, 4], [0, 0], [1, 1], [6, 8]], np.float32)
    maxr = np.array([1, 4, 1, 9, 0], np.int32); acc = np.ones(5, np.float32); den = np.zeros(5, np.float32)
    oracle.densify_stats(radii, g2, maxr, acc, den)
    np.testing.assert_array_equal(maxr, [1, 4, 5, 9, 2])
    np.testing.assert_allclose(acc, [1, 6, 1, 1, 11])
    np.testing.assert_array_equal(den, [0, 1, 1, 0, 1])


def test_l1_loss(oracle):
    a = np.array([[[0.5, 0.2]], [[0.1, 0.9]], [[0.3, 0.3]]], np.float32)
    b = np.array([[[0.25, 0.2]], [[0.6, 0.4]], [[0.3, 0.0]]], np.float32)
    loss, d = oracle.l1_loss(a, b)
    assert loss == pytest.approx(np.abs(a - b).mean(), rel=1e-6)
    np.testing.assert_allclose(d, np.sign(a - b) / a.size)


def _torch_ssim_loss(img, gt, lam):
    """The published trainer's loss: 11x11 Gaussian window (sigma 1.5), zero-padded conv2d per channel."""
    import torch.nn.functional as F
    g = torch.tensor([np.exp(-(i - 5) ** 2 / (2 * 1.5 ** 2)) for i in range(11)], dtype=torch.float64)
    g = g / g.sum()
    w = (g[:, None] * g[None, :]).expand(3, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t[None], w, padding=5, groups=3)[0]
    mu1, mu2 = conv(img), conv(gt)
    s11, s22, s12 = conv(img * img) - mu1 * mu1, conv(gt * gt) - mu2 * mu2, conv(img * gt) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s11 + s22 + C2))
    return (1 - lam) * (img - gt).abs().mean() + lam * (1 - ssim.mean())


@pytest.mark.parametrize("H,W,lam", [(37, 53, 0.2), (16, 16, 1.0), (9, 70, 0.5)])
def test_photometric_loss_matches_torch(oracle, H, W, lam):
    """D-SSIM + L1 (lambda_dssim_, src/arguments/params.h:62) against a float64 torch conv2d/autograd restatement."""
    rng = np.random.default_rng(H)
    a = rng.random((3, H, W)).astype(np.float32); b = np.clip(a + 0.2 * rng.standard_normal((3, H, W)), 0, 1).astype(np.float32)
    loss, d, l1, ss = oracle.photometric_loss(a, b, lam)
    ta = torch.tensor(a, dtype=torch.float64, requires_grad=True); tb = torch.tensor(b, dtype=torch.float64)
    ref = _torch_ssim_loss(ta, tb, lam)
    ref.backward()
    assert loss == pytest.approx(ref.item(), rel=1e-6)
    np.testing.assert_allclose(d, ta.grad.numpy(), rtol=1e-4, atol=1e-9)
    assert l1 == pytest.approx(np.abs(a.astype(np.float64) - b).mean(), rel=1e-6) and 0 < ss < 1
