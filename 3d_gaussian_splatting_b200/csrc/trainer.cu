// trainer.cu — the training loop the reference leaves as a stub (src/utils/train_utils.cpp:97-147: `train`
// builds the model and the background, then `for (iteration = first_iter; ...)` with the body commented out at
// :138-144), as NATIVE host code over this library's own entry points: no Python, no libtorch.
//
// One gs_trainer owns the model (six parameter slabs, twelve Adam-moment slabs, the three statistics), a gs_ctx,
// a copy stream and a ring of pinned result slots.  gs_trainer_run is a depth-2 software pipeline: iteration i is
// enqueued (render -> H2D of its ground truth underneath -> L1 + D-SSIM loss -> backward fused with Adam and the
// statistics -> D2H of loss and overflow flag) while the device still works on iteration i - 1; the host only ever
// waits on the event of the iteration two back.  Schedules follow the reference's parameters
// (src/arguments/params.h:52-68): exponential position learning rate (general_utils.cpp:112-142), one more SH
// degree every 1000 iterations (gaussian_model.cpp:307-313), densification every densification_interval_ after
// densify_from_iter_ until densify_until_iter_, opacity reset every opacity_reset_interval_.
// Views are drawn without replacement from a stack refilled when empty (the published loop's viewpoint_stack), by
// a specified counter-based generator; GS_SAMPLE_CYCLIC walks them in order instead.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <string>
#include <vector>

#include "gs_common.cuh"

namespace {

struct Slab {
  float* p = nullptr;
  size_t floats = 0;
};

struct View {
  gs_settings st;
  const float* gt_host;  // (3,H,W), caller-owned, must outlive the trainer (pinned memory makes the copy async)
};

inline uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

}  // namespace

struct gs_trainer {
  gs_ctx* ctx = nullptr;
  int device = 0;
  gs_train_config cfg;
  int N = 0, K = 1, widths[6] = {3, 3, 0, 1, 3, 4};
  float* P[6] = {};
  float* M[6] = {};
  float* V[6] = {};
  int* max_radii2D = nullptr;
  float* accum = nullptr;
  float* denom = nullptr;
  int* radii = nullptr;
  std::vector<View> views;
  size_t max_pixels = 0;
  float* image = nullptr;
  float* dL = nullptr;
  float* gt_dev = nullptr;
  float* loss_dev = nullptr;
  cudaStream_t stream = nullptr, copy_stream = nullptr;
  cudaEvent_t gt_ready = nullptr, gt_free = nullptr;
  static const int RING = 4;
  float* slots = nullptr;  // pinned: RING x {loss, overflow flag}
  cudaEvent_t done[RING] = {};
  int iteration = 0, active_sh_degree = 0;
  int sh_degree_at_0 = 0;  // the active SH degree is a function of the iteration number: sh_degree_at_0 + iteration / 1000
  long long step_count[6] = {0, 0, 0, 0, 0, 0};
  unsigned long long pair_capacity = 0;
  int retries = 0;
  std::vector<int> stack;  // views not yet drawn in this pass
  uint32_t draws = 0;
};

#define TR_CUDA(t, expr)                                                                             \
  do {                                                                                               \
    cudaError_t _e = (expr);                                                                         \
    if (_e != cudaSuccess) {                                                                         \
      gs_set_error((t)->ctx, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return GS_ECUDA;                                                                               \
    }                                                                                                \
  } while (0)
#define TR_CHECK(expr)          \
  do {                          \
    int _rc = (expr);           \
    if (_rc != GS_OK) return _rc; \
  } while (0)

static void trainer_free_model(gs_trainer* t) {
  for (int k = 0; k < 6; ++k) {
    cudaFree(t->P[k]);
    cudaFree(t->M[k]);
    cudaFree(t->V[k]);
    t->P[k] = t->M[k] = t->V[k] = nullptr;
  }
  cudaFree(t->max_radii2D);
  cudaFree(t->accum);
  cudaFree(t->denom);
  cudaFree(t->radii);
  t->max_radii2D = nullptr;
  t->accum = t->denom = nullptr;
  t->radii = nullptr;
}

static int trainer_alloc_stats(gs_trainer* t, int N) {
  cudaFree(t->max_radii2D);
  cudaFree(t->accum);
  cudaFree(t->denom);
  cudaFree(t->radii);
  const size_t n = (size_t)(N > 0 ? N : 1);
  TR_CUDA(t, cudaMalloc(&t->max_radii2D, n * 4));
  TR_CUDA(t, cudaMalloc(&t->accum, n * 4));
  TR_CUDA(t, cudaMalloc(&t->denom, n * 4));
  TR_CUDA(t, cudaMalloc(&t->radii, n * 4));
  TR_CUDA(t, cudaMemsetAsync(t->max_radii2D, 0, n * 4, t->stream));
  TR_CUDA(t, cudaMemsetAsync(t->accum, 0, n * 4, t->stream));
  TR_CUDA(t, cudaMemsetAsync(t->denom, 0, n * 4, t->stream));
  return GS_OK;
}

extern "C" {

int gs_trainer_create(gs_trainer** out, int device, const gs_train_config* cfg, int N, int sh_coeffs,
                      const float* xyz, const float* f_dc, const float* f_rest, const float* opacity,
                      const float* scaling, const float* rotation) {
  if (!out || !cfg || N < 0 || sh_coeffs < 1 || !xyz || !f_dc || !opacity || !scaling || !rotation ||
      (sh_coeffs > 1 && !f_rest)) {
    gs_set_error(nullptr, "gs_trainer_create: bad argument");
    return GS_EINVAL;
  }
  gs_trainer* t = new (std::nothrow) gs_trainer();
  if (!t) return GS_ENOMEM;
  t->device = device;
  t->cfg = *cfg;
  int rc = gs_create(&t->ctx, device);
  if (rc != GS_OK) {
    delete t;
    return rc;
  }
  t->N = N;
  t->K = sh_coeffs;
  t->widths[2] = 3 * (sh_coeffs - 1);
  {
    const int maxdeg = (int)lround(sqrt((double)sh_coeffs)) - 1;
    t->active_sh_degree = cfg->active_sh_degree < 0 ? 0 : (cfg->active_sh_degree > maxdeg ? maxdeg : cfg->active_sh_degree);
    t->sh_degree_at_0 = t->active_sh_degree;
  }
  const float* src[6] = {xyz, f_dc, f_rest, opacity, scaling, rotation};
  auto fail = [&](int code) {
    gs_trainer_destroy(t);
    return code;
  };
  if (cudaSetDevice(device) != cudaSuccess) return fail(GS_ECUDA);
  if (cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(GS_ECUDA);
  if (cudaStreamCreateWithFlags(&t->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return fail(GS_ECUDA);
  cudaEventCreateWithFlags(&t->gt_ready, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&t->gt_free, cudaEventDisableTiming);
  for (int i = 0; i < gs_trainer::RING; ++i) cudaEventCreateWithFlags(&t->done[i], cudaEventDisableTiming);
  if (cudaMallocHost(&t->slots, sizeof(float) * 2 * gs_trainer::RING) != cudaSuccess) return fail(GS_ECUDA);
  if (cudaMalloc(&t->loss_dev, 16) != cudaSuccess) return fail(GS_ECUDA);
  for (int k = 0; k < 6; ++k) {
    const size_t bytes = std::max<size_t>((size_t)N * t->widths[k] * 4, 16);
    if (cudaMalloc(&t->P[k], bytes) != cudaSuccess || cudaMalloc(&t->M[k], bytes) != cudaSuccess ||
        cudaMalloc(&t->V[k], bytes) != cudaSuccess)
      return fail(GS_ENOMEM);
    cudaMemsetAsync(t->M[k], 0, bytes, t->stream);
    cudaMemsetAsync(t->V[k], 0, bytes, t->stream);
    if (t->widths[k] > 0 && N > 0)
      if (cudaMemcpyAsync(t->P[k], src[k], (size_t)N * t->widths[k] * 4, cudaMemcpyDefault, t->stream) != cudaSuccess)
        return fail(GS_ECUDA);
  }
  if (trainer_alloc_stats(t, N) != GS_OK) return fail(GS_ECUDA);
  if (cudaStreamSynchronize(t->stream) != cudaSuccess) return fail(GS_ECUDA);
  cudaEventRecord(t->gt_free, t->stream);
  *out = t;
  return GS_OK;
}

void gs_trainer_destroy(gs_trainer* t) {
  if (!t) return;
  cudaSetDevice(t->device);
  if (t->stream) cudaStreamSynchronize(t->stream);
  trainer_free_model(t);
  cudaFree(t->image);
  cudaFree(t->dL);
  cudaFree(t->gt_dev);
  cudaFree(t->loss_dev);
  if (t->slots) cudaFreeHost(t->slots);
  if (t->gt_ready) cudaEventDestroy(t->gt_ready);
  if (t->gt_free) cudaEventDestroy(t->gt_free);
  for (int i = 0; i < gs_trainer::RING; ++i)
    if (t->done[i]) cudaEventDestroy(t->done[i]);
  if (t->copy_stream) cudaStreamDestroy(t->copy_stream);
  if (t->stream) cudaStreamDestroy(t->stream);
  gs_destroy(t->ctx);
  delete t;
}

const char* gs_trainer_last_error(const gs_trainer* t) { return gs_last_error(t ? t->ctx : nullptr); }

int gs_trainer_add_view(gs_trainer* t, const gs_settings* camera, const float* gt_host) {
  if (!t || !camera || !gt_host || camera->H <= 0 || camera->W <= 0) {
    gs_set_error(t ? t->ctx : nullptr, "gs_trainer_add_view: bad argument");
    return GS_EINVAL;
  }
  View v;
  v.st = *camera;
  v.gt_host = gt_host;
  const int maxdeg = (int)lround(sqrt((double)t->K)) - 1;
  v.st.max_sh_degree = maxdeg;
  for (int c = 0; c < 3; ++c) v.st.bg[c] = t->cfg.bg[c];
  v.st.scale_modifier = 1.0f;
  v.st.prefiltered = 0;
  v.st.debug = 0;
  t->views.push_back(v);
  const size_t px = (size_t)camera->H * camera->W;
  if (px > t->max_pixels) {
    TR_CUDA(t, cudaSetDevice(t->device));
    TR_CUDA(t, cudaStreamSynchronize(t->stream));
    cudaFree(t->image);
    cudaFree(t->dL);
    cudaFree(t->gt_dev);
    TR_CUDA(t, cudaMalloc(&t->image, px * 12));
    TR_CUDA(t, cudaMalloc(&t->dL, px * 12));
    TR_CUDA(t, cudaMalloc(&t->gt_dev, px * 12));
    t->max_pixels = px;
  }
  return GS_OK;
}

int gs_trainer_num_gaussians(const gs_trainer* t) { return t ? t->N : -1; }
int gs_trainer_iteration(const gs_trainer* t) { return t ? t->iteration : -1; }
int gs_trainer_retries(const gs_trainer* t) { return t ? t->retries : -1; }

int gs_trainer_set_pair_capacity(gs_trainer* t, unsigned long long pairs) {
  if (!t || pairs == 0) return GS_EINVAL;
  t->pair_capacity = pairs;
  return gs_set_duplicate_capacity(t->ctx, pairs);
}

// ---- checkpoint of the live training state (the start_checkpoint hook of src/utils/train_utils.cpp:110-113) ----
// One file in the flat container of 3d_gaussian_splatting_b200/checkpoint.py (libgsplat_b200.so does not link
// libtorch; checkpoint.native_to_libtorch converts it into the reference's torch::save archives):
//   magic "GSB200CK" | u32 version 1 | u32 n_entries | entry*;  entry := u8 dtype (0 f32, 1 i32, 2 i64) | u8 ndim |
//   u16 0 | u64 dims[ndim] | raw bytes.
// Entries: [0] i64[8] {iteration, active_sh_degree, sh_degree_at_0, N, K, draws, pair_capacity, 0}; [1] i64[6] Adam
// step counters; [2..7] parameters, [8..13] exp_avg, [14..19] exp_avg_sq (group order of GaussianModel::setup);
// [20] max_radii2D i32; [21] xyz_gradient_accum; [22] denom; [23] i32 views not yet drawn in this pass.
namespace {
struct FlatWriter {
  FILE* f;
  bool ok;
  void entry(int dtype, const std::vector<unsigned long long>& dims, const void* data, size_t bytes) {
    const unsigned char head[4] = {(unsigned char)dtype, (unsigned char)dims.size(), 0, 0};
    ok = ok && fwrite(head, 1, 4, f) == 4;
    for (unsigned long long d : dims) ok = ok && fwrite(&d, 8, 1, f) == 1;
    if (bytes) ok = ok && fwrite(data, 1, bytes, f) == bytes;
  }
};
}  // namespace

int gs_trainer_save(gs_trainer* t, const char* path) {
  if (!t || !path) return GS_EINVAL;
  TR_CUDA(t, cudaSetDevice(t->device));
  TR_CUDA(t, cudaStreamSynchronize(t->stream));
  const std::string tmp = std::string(path) + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) {
    gs_set_error(t->ctx, "gs_trainer_save: cannot open %s", tmp.c_str());
    return GS_EINVAL;
  }
  FlatWriter w{f, true};
  const unsigned int ver = 1, n_entries = 24;
  w.ok = fwrite("GSB200CK", 1, 8, f) == 8 && fwrite(&ver, 4, 1, f) == 1 && fwrite(&n_entries, 4, 1, f) == 1;
  const long long head[8] = {t->iteration, t->active_sh_degree, t->sh_degree_at_0, t->N, t->K, (long long)t->draws,
                             (long long)t->pair_capacity, 0};
  w.entry(2, {8}, head, sizeof(head));
  w.entry(2, {6}, t->step_count, sizeof(t->step_count));
  std::vector<char> host;
  auto dev_entry = [&](int dtype, const void* dptr, unsigned long long rows, unsigned long long cols) -> int {
    const size_t bytes = (size_t)rows * cols * 4;
    host.resize(bytes ? bytes : 1);
    if (bytes) TR_CUDA(t, cudaMemcpy(host.data(), dptr, bytes, cudaMemcpyDeviceToHost));
    w.entry(dtype, {rows, cols}, host.data(), bytes);
    return GS_OK;
  };
  float* const* slabs[3] = {t->P, t->M, t->V};
  for (int a = 0; a < 3; ++a)
    for (int k = 0; k < 6; ++k) {
      const int rc = dev_entry(0, slabs[a][k], (unsigned long long)t->N, (unsigned long long)t->widths[k]);
      if (rc != GS_OK) {
        fclose(f);
        remove(tmp.c_str());
        return rc;
      }
    }
  int rc = dev_entry(1, t->max_radii2D, (unsigned long long)t->N, 1);
  if (rc == GS_OK) rc = dev_entry(0, t->accum, (unsigned long long)t->N, 1);
  if (rc == GS_OK) rc = dev_entry(0, t->denom, (unsigned long long)t->N, 1);
  if (rc != GS_OK) {
    fclose(f);
    remove(tmp.c_str());
    return rc;
  }
  w.entry(1, {(unsigned long long)t->stack.size()}, t->stack.data(), t->stack.size() * 4);
  const bool closed = fclose(f) == 0;
  if (!w.ok || !closed || rename(tmp.c_str(), path) != 0) {  // a crash never leaves a truncated file under the final name
    remove(tmp.c_str());
    gs_set_error(t->ctx, "gs_trainer_save: writing %s failed", path);
    return GS_EINVAL;
  }
  return GS_OK;
}

int gs_trainer_load(gs_trainer* t, const char* path) {
  if (!t || !path) return GS_EINVAL;
  TR_CUDA(t, cudaSetDevice(t->device));
  TR_CUDA(t, cudaStreamSynchronize(t->stream));
  FILE* f = fopen(path, "rb");
  if (!f) {
    gs_set_error(t->ctx, "gs_trainer_load: cannot open %s", path);
    return GS_EINVAL;
  }
  struct Entry {
    int dtype;
    std::vector<unsigned long long> dims;
    std::vector<char> data;
  };
  std::vector<Entry> e;
  bool ok = true;
  char magic[8];
  unsigned int ver = 0, n = 0;
  ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, "GSB200CK", 8) == 0 && fread(&ver, 4, 1, f) == 1 && ver == 1 &&
       fread(&n, 4, 1, f) == 1 && n == 24;
  for (unsigned int i = 0; ok && i < n; ++i) {
    unsigned char head[4];
    ok = fread(head, 1, 4, f) == 4 && head[0] <= 3 && head[1] <= 4;
    if (!ok) break;
    Entry en;
    en.dtype = head[0];
    size_t count = 1;
    for (int d = 0; ok && d < head[1]; ++d) {
      unsigned long long v = 0;
      ok = fread(&v, 8, 1, f) == 1 && v < (1ull << 40);
      en.dims.push_back(v);
      count *= (size_t)v;
    }
    const size_t bytes = count * ((en.dtype == 2 || en.dtype == 3) ? 8 : 4);
    if (ok && bytes > ((size_t)1 << 40)) ok = false;
    if (ok) {
      en.data.resize(bytes ? bytes : 1);
      ok = bytes == 0 || fread(en.data.data(), 1, bytes, f) == bytes;
    }
    e.push_back(std::move(en));
  }
  fclose(f);
  if (!ok || e.size() != 24 || e[0].data.size() < 64 || e[1].data.size() < 48) {
    gs_set_error(t->ctx, "gs_trainer_load: %s is not a gs_trainer checkpoint (or is truncated)", path);
    return GS_EINVAL;
  }
  const long long* head = (const long long*)e[0].data.data();
  const int N = (int)head[3], K = (int)head[4];
  if (K != t->K || N < 0) {
    gs_set_error(t->ctx, "gs_trainer_load: checkpoint has %d SH coefficients, the trainer %d", K, t->K);
    return GS_EINVAL;
  }
  for (int a = 0; a < 3; ++a)
    for (int k = 0; k < 6; ++k)
      if (e[2 + 6 * a + k].data.size() < (size_t)N * t->widths[k] * 4) {
        gs_set_error(t->ctx, "gs_trainer_load: entry %d is too short", 2 + 6 * a + k);
        return GS_EINVAL;
      }
  if (e[20].data.size() < (size_t)N * 4 || e[21].data.size() < (size_t)N * 4 || e[22].data.size() < (size_t)N * 4) {
    gs_set_error(t->ctx, "gs_trainer_load: statistics entries are too short");
    return GS_EINVAL;
  }
  float** slabs[3] = {t->P, t->M, t->V};
  for (int a = 0; a < 3; ++a)
    for (int k = 0; k < 6; ++k) {
      const size_t bytes = (size_t)N * t->widths[k] * 4;
      cudaFree(slabs[a][k]);
      slabs[a][k] = nullptr;
      TR_CUDA(t, cudaMalloc(&slabs[a][k], bytes > 16 ? bytes : 16));
      if (bytes) TR_CUDA(t, cudaMemcpy(slabs[a][k], e[2 + 6 * a + k].data.data(), bytes, cudaMemcpyHostToDevice));
    }
  t->N = N;
  TR_CHECK(trainer_alloc_stats(t, N));
  TR_CUDA(t, cudaStreamSynchronize(t->stream));
  if (N > 0) {
    TR_CUDA(t, cudaMemcpy(t->max_radii2D, e[20].data.data(), (size_t)N * 4, cudaMemcpyHostToDevice));
    TR_CUDA(t, cudaMemcpy(t->accum, e[21].data.data(), (size_t)N * 4, cudaMemcpyHostToDevice));
    TR_CUDA(t, cudaMemcpy(t->denom, e[22].data.data(), (size_t)N * 4, cudaMemcpyHostToDevice));
  }
  t->iteration = (int)head[0];
  t->active_sh_degree = (int)head[1];
  t->sh_degree_at_0 = (int)head[2];
  t->draws = (uint32_t)head[5];
  if (head[6] > 0 && (unsigned long long)head[6] > t->pair_capacity) {
    t->pair_capacity = (unsigned long long)head[6];
    TR_CHECK(gs_set_duplicate_capacity(t->ctx, t->pair_capacity));
  }
  memcpy(t->step_count, e[1].data.data(), sizeof(t->step_count));
  const int* st = (const int*)e[23].data.data();
  const size_t ns = e[23].dims.empty() ? 0 : (size_t)e[23].dims[0];
  t->stack.assign(st, st + ns);
  for (int v : t->stack)
    if (v < 0 || v >= (int)t->views.size()) {
      t->stack.clear();  // the views of this trainer differ from the checkpoint's: start a fresh pass
      break;
    }
  return GS_OK;
}

int gs_trainer_get_params(gs_trainer* t, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling,
                          float* rotation) {
  if (!t) return GS_EINVAL;
  float* dst[6] = {xyz, f_dc, f_rest, opacity, scaling, rotation};
  TR_CUDA(t, cudaSetDevice(t->device));
  for (int k = 0; k < 6; ++k)
    if (dst[k] && t->widths[k] > 0 && t->N > 0)
      TR_CUDA(t, cudaMemcpyAsync(dst[k], t->P[k], (size_t)t->N * t->widths[k] * 4, cudaMemcpyDefault, t->stream));
  TR_CUDA(t, cudaStreamSynchronize(t->stream));
  return GS_OK;
}

}  // extern "C"

namespace {

gs_gaussians model_view(const gs_trainer* t) {
  gs_gaussians in;
  memset(&in, 0, sizeof(in));
  in.means3D = t->P[0];
  in.sh_dc = t->P[1];
  in.sh_dc_stride = 3;
  in.sh_rest = t->K > 1 ? t->P[2] : nullptr;
  in.sh_rest_stride = 3 * (t->K - 1);
  in.opacities = t->P[3];
  in.scales = t->P[4];
  in.rotations = t->P[5];
  in.raw_params = 1;
  return in;
}

int next_view(gs_trainer* t, int sampling) {
  const int nv = (int)t->views.size();
  if (sampling == GS_SAMPLE_CYCLIC) return (int)(t->draws++ % (uint32_t)nv);
  if (t->stack.empty())
    for (int i = 0; i < nv; ++i) t->stack.push_back(i);
  const uint32_t h = mix32((uint32_t)t->cfg.seed ^ mix32(t->draws++ + 0x9e3779b9u));
  const size_t pos = (size_t)(h % (uint32_t)t->stack.size());
  const int v = t->stack[pos];
  t->stack.erase(t->stack.begin() + (long)pos);
  return v;
}

// Iteration number `iter_no`, enqueued on t->stream; its loss and overflow flag land in pinned slot `slot`.  Learning
// rate and SH degree are functions of iter_no, so an iteration repeated after a capacity overflow runs under the
// schedule of its own number (one more SH degree every 1000 iterations: src/utils/train_utils.cpp:133-136).
int enqueue_iteration(gs_trainer* t, int iter_no, int view, int slot) {
  const gs_train_config& c = t->cfg;
  {
    const int maxdeg = (int)lround(sqrt((double)t->K)) - 1;
    const int deg = t->sh_degree_at_0 + iter_no / 1000;
    t->active_sh_degree = deg > maxdeg ? maxdeg : deg;
  }
  const View& v = t->views[(size_t)view];
  gs_settings st = v.st;
  st.sh_degree = t->active_sh_degree;
  if (c.random_background) {
    // random_background_ (src/arguments/params.h:68): a function of (seed, iteration), so a repeated iteration draws
    // the same colour; the same generator as Trainer.background_for in gaussian_model.py
    for (int ch = 0; ch < 3; ++ch)
      st.bg[ch] = (float)(mix32((uint32_t)c.seed ^ mix32((uint32_t)iter_no * 3u + (uint32_t)ch + 0x51ed270bu)) >> 8) * (1.0f / 16777216.0f);
  }
  const gs_gaussians in = model_view(t);
  TR_CHECK(gs_forward(t->ctx, &st, t->N, &in, t->image, t->radii, t->stream));
  // ground truth: host -> device on the copy stream, underneath the render that was just enqueued
  const size_t bytes = (size_t)st.H * st.W * 12;
  TR_CUDA(t, cudaStreamWaitEvent(t->copy_stream, t->gt_free, 0));
  TR_CUDA(t, cudaMemcpyAsync(t->gt_dev, v.gt_host, bytes, cudaMemcpyHostToDevice, t->copy_stream));
  TR_CUDA(t, cudaEventRecord(t->gt_ready, t->copy_stream));
  TR_CUDA(t, cudaStreamWaitEvent(t->stream, t->gt_ready, 0));
  TR_CHECK(gs_loss(t->ctx, st.H, st.W, t->image, t->gt_dev, (float)c.lambda_dssim, t->loss_dev, t->dL, t->stream));
  TR_CUDA(t, cudaEventRecord(t->gt_free, t->stream));
  double lr[6] = {0, c.feature_lr, c.feature_lr / 20.0, c.opacity_lr, c.scaling_lr, c.rotation_lr};
  lr[0] = (double)gs_expon_lr((float)(c.position_lr_init * c.spatial_lr_scale), (float)(c.position_lr_final * c.spatial_lr_scale),
                              0, (float)c.position_lr_delay_mult, c.position_lr_max_steps, iter_no);
  for (int k = 0; k < 6; ++k) t->step_count[k] += 1;
  TR_CHECK(gs_backward_adam_step(t->ctx, &st, t->N, &in, t->dL, nullptr, t->M, t->V, lr, t->step_count, 0.9, 0.999, 1e-8,
                                 t->max_radii2D, t->accum, t->denom, gs_overflow_guard(t->ctx), t->stream));
  TR_CUDA(t, cudaMemcpyAsync(t->slots + 2 * slot, t->loss_dev, 4, cudaMemcpyDeviceToHost, t->stream));
  TR_CUDA(t, cudaMemcpyAsync(t->slots + 2 * slot + 1, gs_overflow_guard(t->ctx), 4, cudaMemcpyDeviceToHost, t->stream));
  TR_CUDA(t, cudaEventRecord(t->done[slot], t->stream));
  return GS_OK;
}

int densify_if_due(gs_trainer* t, float extent) {
  const gs_train_config& c = t->cfg;
  const int it = t->iteration;
  if (extent <= 0.0f || it >= c.densify_until_iter) return GS_OK;
  const bool densify = it > c.densify_from_iter && c.densification_interval > 0 && it % c.densification_interval == 0;
  const bool reset = c.opacity_reset_interval > 0 && it % c.opacity_reset_interval == 0;
  if (!densify && !reset) return GS_OK;
  if (densify) {
    int counts[4];
    const float size_threshold = it > c.opacity_reset_interval ? 20.0f : 0.0f;
    TR_CHECK(gs_densify_plan(t->ctx, t->N, t->accum, t->denom, t->P[3], t->P[4], (float)c.densify_grad_threshold, 0.005f,
                             extent, (float)c.percent_dense, size_threshold, counts, t->stream));
    const int n_new = counts[0] + counts[1] + 2 * counts[2];
    float *nP[6] = {}, *nM[6] = {}, *nV[6] = {};
    for (int k = 0; k < 6; ++k) {
      const size_t bytes = std::max<size_t>((size_t)n_new * t->widths[k] * 4, 16);
      TR_CUDA(t, cudaMalloc(&nP[k], bytes));
      TR_CUDA(t, cudaMalloc(&nM[k], bytes));
      TR_CUDA(t, cudaMalloc(&nV[k], bytes));
    }
    TR_CHECK(gs_densify_apply(t->ctx, t->N, t->widths, t->P, t->M, t->V, nP, nM, nV, (unsigned long long)it, t->stream));
    TR_CUDA(t, cudaStreamSynchronize(t->stream));
    for (int k = 0; k < 6; ++k) {
      cudaFree(t->P[k]);
      cudaFree(t->M[k]);
      cudaFree(t->V[k]);
      t->P[k] = nP[k];
      t->M[k] = nM[k];
      t->V[k] = nV[k];
    }
    if (t->pair_capacity > 0 && n_new > t->N) {
      // the (tile, Gaussian) pair count grows with N: keep the fixed capacity ahead of it
      t->pair_capacity = (unsigned long long)((double)t->pair_capacity * ((double)n_new / (double)(t->N > 0 ? t->N : 1)) * 1.1) + 1024;
      TR_CHECK(gs_set_duplicate_capacity(t->ctx, t->pair_capacity));
    }
    t->N = n_new;
    TR_CHECK(trainer_alloc_stats(t, n_new));
  }
  if (reset) TR_CHECK(gs_reset_opacity(t->ctx, t->N, t->P[3], t->M[3], t->V[3], t->stream));
  return GS_OK;
}

}  // namespace

extern "C" int gs_trainer_run(gs_trainer* t, int n_iterations, float* losses, int sampling, float densify_extent) {
  if (!t || n_iterations < 0 || (n_iterations > 0 && !losses) || t->views.empty()) {
    gs_set_error(t ? t->ctx : nullptr, "gs_trainer_run: bad argument (add at least one view first)");
    return GS_EINVAL;
  }
  TR_CUDA(t, cudaSetDevice(t->device));
  // fixed pair capacity so that no forward waits for the host: sized once from one synchronous forward per view
  if (t->pair_capacity == 0) {
    unsigned long long worst = 0;
    const gs_gaussians in = model_view(t);
    for (const View& v : t->views) {
      gs_settings st = v.st;
      st.sh_degree = t->active_sh_degree;
      TR_CHECK(gs_forward(t->ctx, &st, t->N, &in, t->image, t->radii, t->stream));
      TR_CUDA(t, cudaStreamSynchronize(t->stream));
      unsigned long long D = 0;
      int ov = 0;
      TR_CHECK(gs_get_forward_status(t->ctx, &D, &ov));
      worst = std::max(worst, D);
    }
    t->pair_capacity = (unsigned long long)(1.3 * (double)worst) + 1024;
    TR_CHECK(gs_set_duplicate_capacity(t->ctx, t->pair_capacity));
  }
  struct Pending {
    int j, view, slot, iter_no;
  };
  std::vector<Pending> pending, redo;
  const int depth = 2;
  auto retire = [&]() -> int {
    const Pending p = pending.front();
    pending.erase(pending.begin());
    TR_CUDA(t, cudaEventSynchronize(t->done[p.slot]));
    if (t->slots[2 * p.slot + 1] != 0.0f) redo.push_back(p);  // overflowed: skipped on the device
    else losses[p.j] = t->slots[2 * p.slot];
    return GS_OK;
  };
  // An overflowed iteration was void on the device (parameters, moments, statistics untouched).  Recover AT ONCE:
  // drain, take back the void attempts' step counters, grow the capacity, repeat those iterations under their own
  // iteration numbers -- before anything new is enqueued, so an outgrown capacity never voids the rest of a run.
  auto recover = [&]() -> int {
    while (!pending.empty()) TR_CHECK(retire());
    while (!redo.empty()) {
      std::vector<Pending> batch;
      batch.swap(redo);
      unsigned long long D = 0;
      int ov = 0;
      TR_CHECK(gs_get_forward_status(t->ctx, &D, &ov));
      t->pair_capacity = std::max(2 * t->pair_capacity, (unsigned long long)(1.25 * (double)D));
      TR_CHECK(gs_set_duplicate_capacity(t->ctx, t->pair_capacity));
      t->retries += (int)batch.size();
      for (const Pending& p : batch) {
        for (int k = 0; k < 6; ++k) t->step_count[k] -= 1;
        TR_CHECK(enqueue_iteration(t, p.iter_no, p.view, 0));
        pending.push_back({p.j, p.view, 0, p.iter_no});
        TR_CHECK(retire());
      }
    }
    return GS_OK;
  };
  for (int j = 0; j < n_iterations; ++j) {
    if ((int)pending.size() >= depth) TR_CHECK(retire());
    if (!redo.empty()) TR_CHECK(recover());
    const int view = next_view(t, sampling);
    const int slot = j % gs_trainer::RING;
    t->iteration += 1;
    TR_CHECK(enqueue_iteration(t, t->iteration, view, slot));
    pending.push_back({j, view, slot, t->iteration});
    if (densify_extent > 0.0f) {
      const gs_train_config& c = t->cfg;
      const int it = t->iteration;
      const bool due = it < c.densify_until_iter &&
                       ((it > c.densify_from_iter && c.densification_interval > 0 && it % c.densification_interval == 0) ||
                        (c.opacity_reset_interval > 0 && it % c.opacity_reset_interval == 0));
      if (due) {
        TR_CHECK(recover());
        TR_CHECK(densify_if_due(t, densify_extent));
      }
    }
  }
  TR_CHECK(recover());
  return GS_OK;
}
