// gsplat_b200_torch.h — libtorch shim over the C ABI: lets the reference's own objects
// (torch::Tensor members of GaussianModel, src/scene/gaussian_model.h:10-25; Camera, src/scene/camera.h:5-43) drive
// the kernels through autograd, so that the commented-out loop body at src/utils/train_utils.cpp:138-144
//     auto render_pkg = render(viewpoint_cam, gaussians, pipe, background);
//     auto loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt));  loss.backward();
// works as written, with the six torch::optim::Adam objects of GaussianModel::setup (gaussian_model.cpp:323-345)
// left in place.  Header-only; needs libtorch (the reference already links it) and libgsplat_b200.so.
// This is the COMPATIBLE path: gradients are materialised as torch tensors for torch's optimisers.  The fast path
// (fused backward + Adam, no gradient tensors) is gs_backward_adam_step / gs_adam_densify_step in gsplat_b200.h.
#pragma once

#include <c10/cuda/CUDAStream.h>
#include <torch/torch.h>

#include <memory>
#include <stdexcept>
#include <string>

#include "gsplat_b200.h"

namespace gsplat_b200 {

// One context per (thread, device), destroyed with the handle.
struct Context {
  gs_ctx* ctx = nullptr;
  // stamped by every Rasterize::forward: the gs_ctx keeps the state of its LAST forward only, so a backward whose
  // stamp is stale (another render ran on this context in between) must fail loudly instead of differentiating the
  // wrong view
  uint64_t generation = 0;
  explicit Context(int device = 0) {
    if (gs_create(&ctx, device) != GS_OK) throw std::runtime_error(std::string("gs_create: ") + gs_last_error(nullptr));
  }
  ~Context() { gs_destroy(ctx); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
};

inline void check(gs_ctx* ctx, int rc, const char* what) {
  if (rc != GS_OK) throw std::runtime_error(std::string(what) + ": " + gs_last_error(ctx));
}

inline const float* fptr(const torch::Tensor& t) { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; }

// Raster settings from the reference's Camera fields (world_view_transform_, full_proj_transform_,
// camera_center_ are the transposed / flattened tensors built at src/scene/camera.cpp:66-71; FoVx_/FoVy_ :24-25).
inline gs_settings make_settings(int H, int W, double FoVx, double FoVy, const torch::Tensor& world_view_transform,
                                 const torch::Tensor& full_proj_transform, const torch::Tensor& camera_center,
                                 const torch::Tensor& bg_color, int sh_degree, int max_sh_degree,
                                 float scaling_modifier = 1.0f, bool debug = false) {
  gs_settings st{};
  st.H = H;
  st.W = W;
  st.tanfovx = (float)std::tan(FoVx * 0.5);
  st.tanfovy = (float)std::tan(FoVy * 0.5);
  auto vm = world_view_transform.to(torch::kCPU, torch::kFloat32).contiguous();
  auto pm = full_proj_transform.to(torch::kCPU, torch::kFloat32).contiguous();
  auto cc = camera_center.to(torch::kCPU, torch::kFloat32).contiguous();
  auto bg = bg_color.to(torch::kCPU, torch::kFloat32).contiguous();
  for (int i = 0; i < 16; ++i) {
    st.viewmatrix[i] = vm.data_ptr<float>()[i];
    st.projmatrix[i] = pm.data_ptr<float>()[i];
  }
  for (int i = 0; i < 3; ++i) {
    st.campos[i] = cc.data_ptr<float>()[i];
    st.bg[i] = bg.data_ptr<float>()[i];
  }
  st.scale_modifier = scaling_modifier;
  st.sh_degree = sh_degree;
  st.max_sh_degree = max_sh_degree;
  st.prefiltered = 0;
  st.debug = debug ? 1 : 0;
  return st;
}

// render(): forward(ctx, settings, xyz_, features_dc_, features_rest_, opacity_, scaling_, rotation_) ->
// {image (3,H,W), radii (N) int32}.  The six inputs are the RAW parameter tensors (the activations of
// get_scaling / get_rotation / get_opacity / get_features, gaussian_model.cpp:270-304, are fused in the kernels),
// so loss.backward() deposits .grad on exactly the tensors the reference's optimisers own.
// means2D (optional): an (N,2) zero tensor with requires_grad, the "screenspace points" of the published API --
// its .grad receives dL/dmean2D, whose per-Gaussian norm feeds xyz_gradient_accum_ (src/scene/gaussian_model.h:19).
class Rasterize : public torch::autograd::Function<Rasterize> {
 public:
  static torch::autograd::variable_list forward(torch::autograd::AutogradContext* actx, std::shared_ptr<Context> gs,
                                                gs_settings st, torch::Tensor xyz, torch::Tensor features_dc,
                                                torch::Tensor features_rest, torch::Tensor opacity,
                                                torch::Tensor scaling, torch::Tensor rotation,
                                                torch::Tensor means2D = torch::Tensor()) {
    TORCH_CHECK(xyz.is_cuda() && xyz.scalar_type() == torch::kFloat32, "parameters must be float32 CUDA tensors");
    xyz = xyz.contiguous();
    features_dc = features_dc.contiguous();
    features_rest = features_rest.contiguous();
    opacity = opacity.contiguous();
    scaling = scaling.contiguous();
    rotation = rotation.contiguous();
    const int N = (int)xyz.size(0);
    const int K = 1 + (features_rest.dim() == 3 ? (int)features_rest.size(1) : 0);
    gs_gaussians in{};
    in.means3D = fptr(xyz);
    in.scales = fptr(scaling);
    in.rotations = fptr(rotation);
    in.opacities = fptr(opacity);
    in.sh_dc = fptr(features_dc);
    in.sh_dc_stride = 3;
    in.sh_rest = K > 1 ? fptr(features_rest) : nullptr;
    in.sh_rest_stride = 3 * (K - 1);
    in.raw_params = 1;
    auto image = torch::empty({3, st.H, st.W}, xyz.options());
    auto radii = torch::empty({N}, xyz.options().dtype(torch::kInt32));
    void* stream = c10::cuda::getCurrentCUDAStream().stream();
    check(gs->ctx, gs_forward(gs->ctx, &st, N, &in, image.data_ptr<float>(), radii.data_ptr<int>(), stream), "gs_forward");
    actx->save_for_backward({xyz, features_dc, features_rest, opacity, scaling, rotation});
    actx->saved_data["K"] = (int64_t)K;
    // The context and the settings travel as opaque blobs (autograd only stores IValues).  The context handle lives
    // in a one-byte tensor whose DELETER releases it: freed with the autograd node whether or not backward ever
    // runs (an evaluation render with requires_grad), and never twice (retain_graph, a second backward).
    auto* handle = new std::shared_ptr<Context>(gs);
    actx->saved_data["gs"] = torch::from_blob(handle, {1}, [handle](void*) { delete handle; }, torch::kUInt8);
    actx->saved_data["generation"] = (int64_t)(++gs->generation);
    actx->saved_data["has_means2D"] = means2D.defined();
    actx->saved_data["st"] = torch::from_blob(&st, {(int64_t)sizeof(gs_settings)}, torch::kUInt8).clone();
    actx->mark_non_differentiable({radii});
    return {image, radii};
  }

  static torch::autograd::variable_list backward(torch::autograd::AutogradContext* actx,
                                                 torch::autograd::variable_list grad_out) {
    auto saved = actx->get_saved_variables();
    auto xyz = saved[0], features_dc = saved[1], features_rest = saved[2], opacity = saved[3], scaling = saved[4],
         rotation = saved[5];
    const std::shared_ptr<Context>& gs =
        *reinterpret_cast<std::shared_ptr<Context>*>(actx->saved_data["gs"].toTensor().data_ptr());
    TORCH_CHECK((int64_t)gs->generation == actx->saved_data["generation"].toInt(),
                "gsplat_b200::Rasterize: another render ran on this context between this forward and its backward; "
                "the context keeps the state of its last forward only (use one Context per in-flight render)");
    gs_ctx* ctx = gs->ctx;
    gs_settings st;
    std::memcpy(&st, actx->saved_data["st"].toTensor().data_ptr<uint8_t>(), sizeof(gs_settings));
    const int K = (int)actx->saved_data["K"].toInt();
    const int N = (int)xyz.size(0);
    gs_gaussians in{};
    in.means3D = fptr(xyz);
    in.scales = fptr(scaling);
    in.rotations = fptr(rotation);
    in.opacities = fptr(opacity);
    in.sh_dc = fptr(features_dc);
    in.sh_dc_stride = 3;
    in.sh_rest = K > 1 ? fptr(features_rest) : nullptr;
    in.sh_rest_stride = 3 * (K - 1);
    in.raw_params = 1;
    auto dL = grad_out[0].contiguous();
    auto g_xyz = torch::empty_like(xyz), g_dc = torch::empty_like(features_dc), g_rest = torch::empty_like(features_rest);
    auto g_op = torch::empty_like(opacity), g_scl = torch::empty_like(scaling), g_rot = torch::empty_like(rotation);
    auto g_m2d = torch::empty({N, 2}, xyz.options());
    gs_gradients g{};
    g.dL_dmeans3D = g_xyz.data_ptr<float>();
    g.dL_dscales = g_scl.data_ptr<float>();
    g.dL_drotations = g_rot.data_ptr<float>();
    g.dL_dopacities = g_op.data_ptr<float>();
    g.dL_dsh_dc = g_dc.data_ptr<float>();
    g.sh_dc_stride = 3;
    g.dL_dsh_rest = K > 1 ? g_rest.data_ptr<float>() : nullptr;
    g.sh_rest_stride = 3 * (K - 1);
    g.dL_dmeans2D = g_m2d.data_ptr<float>();
    void* stream = c10::cuda::getCurrentCUDAStream().stream();
    check(ctx, gs_backward(ctx, &st, N, &in, dL.data_ptr<float>(), &g, stream), "gs_backward");
    // inputs of forward(): gs, st, the six parameter tensors, means2D
    return {torch::Tensor(), torch::Tensor(), g_xyz, g_dc, g_rest, g_op, g_scl, g_rot,
            actx->saved_data["has_means2D"].toBool() ? g_m2d : torch::Tensor()};
  }
};

// The loss line of the loop: returns the scalar loss tensor and dL/dimage is produced by autograd as usual.
// (For the fused path call gs_loss directly: it returns dL/dimage in the same pass.)
inline torch::Tensor photometric_loss(const torch::Tensor& image, const torch::Tensor& gt, double lambda_dssim,
                                      std::shared_ptr<Context> gs) {
  class Loss : public torch::autograd::Function<Loss> {
   public:
    static torch::Tensor forward(torch::autograd::AutogradContext* actx, std::shared_ptr<Context> gs, torch::Tensor image,
                                 torch::Tensor gt, double lambda) {
      image = image.contiguous();
      gt = gt.contiguous();
      auto loss = torch::empty({1}, image.options());
      auto dL = torch::empty_like(image);
      void* stream = c10::cuda::getCurrentCUDAStream().stream();
      check(gs->ctx,
            gs_loss(gs->ctx, (int)image.size(1), (int)image.size(2), image.data_ptr<float>(), gt.data_ptr<float>(),
                    (float)lambda, loss.data_ptr<float>(), dL.data_ptr<float>(), stream),
            "gs_loss");
      actx->save_for_backward({dL});
      return loss.squeeze();
    }
    static torch::autograd::variable_list backward(torch::autograd::AutogradContext* actx,
                                                   torch::autograd::variable_list grad_out) {
      auto dL = actx->get_saved_variables()[0];
      return {torch::Tensor(), dL * grad_out[0], torch::Tensor(), torch::Tensor()};
    }
  };
  return Loss::apply(std::move(gs), image, gt, lambda_dssim);
}

}  // namespace gsplat_b200
