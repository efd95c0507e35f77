// Test extension: drives include/gsplat_b200_torch.h exactly as the reference's training loop would
// (render -> loss -> backward with libtorch autograd) and hands the results to Python for comparison with the
// ctypes path.  Built in-tree by tests/shim/build_shim.py.
#include <torch/extension.h>

#include "gsplat_b200_torch.h"

namespace {
std::shared_ptr<gsplat_b200::Context> g_ctx;

std::vector<torch::Tensor> render_loss_backward(int64_t H, int64_t W, double FoVx, double FoVy, torch::Tensor view,
                                                torch::Tensor proj, torch::Tensor campos, torch::Tensor bg, int64_t deg,
                                                int64_t maxdeg, torch::Tensor xyz, torch::Tensor f_dc, torch::Tensor f_rest,
                                                torch::Tensor opacity, torch::Tensor scaling, torch::Tensor rotation,
                                                torch::Tensor gt, double lambda_dssim) {
  pybind11::gil_scoped_release no_gil;  // the autograd engine must not be entered with the GIL held
  if (!g_ctx) g_ctx = std::make_shared<gsplat_b200::Context>(xyz.get_device());
  std::vector<torch::Tensor> params = {xyz, f_dc, f_rest, opacity, scaling, rotation};
  for (auto& p : params) p.set_requires_grad(true);
  auto st = gsplat_b200::make_settings((int)H, (int)W, FoVx, FoVy, view, proj, campos, bg, (int)deg, (int)maxdeg);
  // the "screenspace points" of the published API: a zero tensor whose .grad receives dL/dmean2D
  auto means2D = torch::zeros({xyz.size(0), 2}, xyz.options()).set_requires_grad(true);
  auto out = gsplat_b200::Rasterize::apply(g_ctx, st, xyz, f_dc, f_rest, opacity, scaling, rotation, means2D);
  auto loss = gsplat_b200::photometric_loss(out[0], gt, lambda_dssim, g_ctx);
  loss.backward();
  return {out[0].detach(), out[1], loss.detach(), xyz.grad(), f_dc.grad(), f_rest.grad(), opacity.grad(), scaling.grad(),
          rotation.grad(), means2D.grad()};
}

// Two renders on ONE context before the first one's backward: the context keeps the state of its last forward only,
// so the stale backward must throw; an evaluation render whose backward never runs must not leak its handle (the
// use count of the context returns to 1 once the graph is gone).  Returns {threw, use_count_after}.
std::vector<int64_t> stale_backward_and_leak_check(int64_t H, int64_t W, double FoVx, double FoVy, torch::Tensor view,
                                                   torch::Tensor proj, torch::Tensor campos, torch::Tensor bg, int64_t deg,
                                                   int64_t maxdeg, torch::Tensor xyz, torch::Tensor f_dc, torch::Tensor f_rest,
                                                   torch::Tensor opacity, torch::Tensor scaling, torch::Tensor rotation) {
  pybind11::gil_scoped_release no_gil;
  auto ctx = std::make_shared<gsplat_b200::Context>(xyz.get_device());
  xyz.set_requires_grad(true);
  auto st = gsplat_b200::make_settings((int)H, (int)W, FoVx, FoVy, view, proj, campos, bg, (int)deg, (int)maxdeg);
  int64_t threw = 0;
  {
    auto a = gsplat_b200::Rasterize::apply(ctx, st, xyz, f_dc, f_rest, opacity, scaling, rotation);
    auto b = gsplat_b200::Rasterize::apply(ctx, st, xyz, f_dc, f_rest, opacity, scaling, rotation);
    try {
      a[0].sum().backward();
    } catch (const std::exception&) {
      threw = 1;
    }
    b[0].sum().backward();                    // the latest render differentiates fine
    auto c = gsplat_b200::Rasterize::apply(ctx, st, xyz, f_dc, f_rest, opacity, scaling, rotation);   // never differentiated
  }
  return {threw, (int64_t)ctx.use_count()};
}
}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("render_loss_backward", &render_loss_backward);
  m.def("stale_backward_and_leak_check", &stale_backward_and_leak_check);
}
