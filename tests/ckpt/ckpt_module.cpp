// Test-only extension: the reference's own serialisation calls (src/scene/gaussian_model.cpp:85-131, :165-201) --
// torch::save / torch::load of a std::vector<torch::Tensor> and of a torch::optim::Adam -- so that the archives
// 3d_gaussian_splatting_b200/checkpoint.py writes and reads are checked against libtorch itself.
#include <torch/extension.h>
#include <torch/torch.h>

#include <memory>
#include <string>
#include <vector>

static void save_tensors(const std::string& path, std::vector<torch::Tensor> tensors) { torch::save(tensors, path); }

static std::vector<torch::Tensor> load_tensors(const std::string& path) {
  std::vector<torch::Tensor> tensors;
  torch::load(tensors, path);
  return tensors;
}

// one Adam over one parameter, as GaussianModel::setup builds it (gaussian_model.cpp:323-345): lr only, defaults else
static std::unique_ptr<torch::optim::Adam> make_adam(torch::Tensor param, double lr) {
  std::vector<torch::Tensor> params = {param};
  return std::make_unique<torch::optim::Adam>(params, torch::optim::AdamOptions(lr));
}

// writes an Adam whose state for `param` is (step, exp_avg, exp_avg_sq)
static void save_adam(const std::string& path, torch::Tensor param, double lr, int64_t step, torch::Tensor exp_avg,
                      torch::Tensor exp_avg_sq) {
  auto opt = make_adam(param, lr);
  auto st = std::make_unique<torch::optim::AdamParamState>();
  st->step(step);
  st->exp_avg(exp_avg);
  st->exp_avg_sq(exp_avg_sq);
  opt->state()[param.unsafeGetTensorImpl()] = std::move(st);
  torch::save(*opt, path);
}

// loads an Adam archive into a freshly set-up optimiser (what CoreParams::restore does) and returns
// {step, lr, exp_avg, exp_avg_sq}
static std::vector<torch::Tensor> load_adam(const std::string& path, torch::Tensor param, double lr) {
  auto opt = make_adam(param, lr);
  torch::load(*opt, path);
  auto it = opt->state().find(param.unsafeGetTensorImpl());
  TORCH_CHECK(it != opt->state().end(), "no state for the parameter after torch::load");
  auto& st = static_cast<torch::optim::AdamParamState&>(*it->second);
  const double got_lr = static_cast<torch::optim::AdamOptions&>(opt->param_groups()[0].options()).lr();
  return {torch::tensor(st.step()), torch::tensor(got_lr, torch::kFloat64), st.exp_avg(), st.exp_avg_sq()};
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("save_tensors", &save_tensors);
  m.def("load_tensors", &load_tensors);
  m.def("save_adam", &save_adam);
  m.def("load_adam", &load_adam);
}
