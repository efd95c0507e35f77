"""gs_loss_static: the allocation-free variant of gs_loss (SSIM workspace in a file-scope device array,
include/gsplat_b200.h).  It has no counterpart in the reference; it was added
for callers that may not call cudaMalloc.  Checked through the C ABI against the existing entry point gs_loss, which
tests/test_gpu_parity.py pins to the oracle.  This file is collected last on purpose."""
import ctypes as C
import importlib

import numpy as np
import pytest

MAX_PIXELS = 262144  # GS_LOSS_STATIC_MAX_PIXELS


def _images(H, W, seed):
    rng = np.random.default_rng(seed)
    a = rng.random((3, H, W)).astype(np.float32)
    b = np.clip(a + 0.2 * rng.standard_normal((3, H, W)), 0, 1).astype(np.float32)
    return a, b


def test_static_loss_rejects_sizes_beyond_the_array(pkg):
    """The size check comes before any CUDA call, so it is testable without a GPU."""
    L = pkg._lib
    dummy = C.c_void_p(256)
    for H, W in [(513, 512), (1, MAX_PIXELS + 1), (825, 1246)]:
        rc = L.lib.gs_loss_static(0, H, W, dummy, dummy, 0.2, dummy, dummy, None)
        assert rc == L.GS_EINVAL
        msg = L.lib.gs_last_error(None)
        assert b"GS_LOSS_STATIC_MAX_PIXELS" in msg and b"gs_loss" in msg
    assert L.lib.gs_loss_static(0, 0, 8, dummy, dummy, 0.2, dummy, dummy, None) == L.GS_EINVAL
    assert L.lib.gs_loss_static(0, 8, 8, dummy, dummy, 1.5, dummy, dummy, None) == L.GS_EINVAL
    assert L.lib.gs_loss_static(-1, 8, 8, dummy, dummy, 0.2, dummy, dummy, None) == L.GS_EINVAL


@pytest.mark.gpu
def test_static_loss_matches_gs_loss_on_consecutive_calls(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    L = pkg._lib
    rast = pkg.Rasterizer(0)
    try:
        stream = torch.cuda.current_stream(0).cuda_stream
        # three calls in a row on the shared array: a large image first, then smaller ones that find the first call's
        # maps and block partials still in the array, the last at exactly the maximum size
        cases = [(384, 500, 0.2), (75, 201, 0.2), (37, 53, 1.0), (512, 512, 0.2)]
        inputs = []
        for k, (H, W, lam) in enumerate(cases):
            a, b = _images(H, W, 100 + k)
            img, gt = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
            ref_loss, ref_d = rast.loss(img, gt, lam)          # existing entry point (gs_loss)
            out_loss = torch.full((1,), float("nan"), device="cuda")
            out_d = torch.full_like(img, float("nan"))
            inputs.append((H, W, lam, img, gt, ref_loss, ref_d, out_loss, out_d))
        torch.cuda.synchronize()

        def call(H, W, lam, img, gt, out_loss, out_d):
            rc = L.lib.gs_loss_static(0, H, W, C.c_void_p(img.data_ptr()), C.c_void_p(gt.data_ptr()), lam,
                                      C.c_void_p(out_loss.data_ptr()), C.c_void_p(out_d.data_ptr()), C.c_void_p(stream))
            assert rc == L.GS_OK, L.lib.gs_last_error(None)

        # first call on this device resolves the array's address (set-up); after it, device memory must not move
        H, W, lam, img, gt, _, _, out_loss, out_d = inputs[0]
        call(H, W, lam, img, gt, out_loss, out_d)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(0)
        for rep in range(2):
            for (H, W, lam, img, gt, ref_loss, ref_d, out_loss, out_d) in inputs:
                out_loss.fill_(float("nan")); out_d.fill_(float("nan"))
                call(H, W, lam, img, gt, out_loss, out_d)      # back to back on one stream, no sync in between
            torch.cuda.synchronize()
            for (H, W, lam, img, gt, ref_loss, ref_d, out_loss, out_d) in inputs:
                # same kernels, same launch shape: the tolerance of test_photometric_loss_dssim (1e-5 / 1e-4) is met
                # with zero difference
                assert abs(out_loss.item() - ref_loss.item()) <= 1e-5 * abs(ref_loss.item())
                err = (out_d - ref_d).abs().max().item()
                assert err <= 1e-4 * ref_d.abs().max().item()
                assert torch.equal(out_loss, ref_loss) and torch.equal(out_d, ref_d)
        free_after, _ = torch.cuda.mem_get_info(0)
        assert free_after == free_before, "gs_loss_static changed the device's free memory"

        # loss only (no gradient wanted), and lambda = 0 through the same kernels (gs_loss uses its pure-L1 kernel there)
        H, W, lam, img, gt, ref_loss, _, out_loss, _ = inputs[1]
        out_loss.fill_(float("nan"))
        rc = L.lib.gs_loss_static(0, H, W, C.c_void_p(img.data_ptr()), C.c_void_p(gt.data_ptr()), lam,
                                  C.c_void_p(out_loss.data_ptr()), None, C.c_void_p(stream))
        assert rc == L.GS_OK
        torch.cuda.synchronize()
        assert torch.equal(out_loss, ref_loss)
        l1_ref, d_ref = rast.loss(img, gt, 0.0)
        out_d = torch.empty_like(img)
        call(H, W, 0.0, img, gt, out_loss, out_d)
        torch.cuda.synchronize()
        assert abs(out_loss.item() - l1_ref.item()) <= 1e-6
        assert (out_d - d_ref).abs().max().item() <= 1e-4 * d_ref.abs().max().item()

        # beyond the maximum: rejected with a clear error, outputs untouched
        big = torch.zeros(3, 513, 512, device="cuda")
        out_loss.fill_(7.0)
        rc = L.lib.gs_loss_static(0, 513, 512, C.c_void_p(big.data_ptr()), C.c_void_p(big.data_ptr()), 0.2,
                                  C.c_void_p(out_loss.data_ptr()), None, C.c_void_p(stream))
        assert rc == L.GS_EINVAL and b"exceeds the static workspace" in L.lib.gs_last_error(None)
        torch.cuda.synchronize()
        assert out_loss.item() == 7.0
    finally:
        rast.close()
