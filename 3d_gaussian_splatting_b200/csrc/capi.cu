// capi.cu — the C ABI of libgsplat_b200.so (include/gsplat_b200.h): context, workspace
// management and stage orchestration.  No torch types, no exceptions across the boundary, no
// CPU fallback: compute entry points fail with GS_ECUDA when no sm_100 device is usable.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "gs_common.cuh"

// ---- launchers implemented in the other translation units ----
cudaError_t gs_launch_preprocess(const gs_settings&, const gs_gaussians&, int, GsGeom*, float*, int*, uint32_t*,
                                 uint4*, uint8_t*, int*, int defer_color, cudaStream_t);
cudaError_t gs_launch_sh_color(const gs_settings&, const gs_gaussians&, int, const int* radii, GsGeom*, uint8_t* clamped,
                               cudaStream_t);
cudaError_t gs_launch_depth_sort(int, const uint32_t*, uint32_t*, uint32_t*, uint32_t*, uint32_t*, const uint32_t*,
                                 unsigned long long*, uint32_t*, uint32_t, int, int, uint32_t, cudaStream_t);
size_t gs_onesweep_status_words(size_t n, int items);
int gs_dup_num_blocks(int N);
cudaError_t gs_launch_dup_scan(int, const uint32_t*, const uint32_t*, uint32_t*, unsigned long long*, uint32_t*,
                               unsigned long long, uint32_t*, float*, float*, uint32_t*, cudaStream_t);
cudaError_t gs_launch_duplicate(int, const uint32_t*, const uint4*, const uint32_t*, void*, int, uint32_t*, int,
                                uint32_t, uint32_t*, uint32_t*, cudaStream_t);
cudaError_t gs_launch_tile_sort(const uint32_t*, uint32_t, int, int, void*, void*, uint32_t*, uint32_t*, const uint32_t*,
                                unsigned long long*, uint32_t*, uint32_t, int, int, uint32_t, int*, cudaStream_t);
cudaError_t gs_launch_ranges(const uint32_t*, const void*, int, uint2*, int, cudaStream_t);
cudaError_t gs_launch_depth_keys(const gs_settings&, const float*, int, uint32_t*, uint32_t*, cudaStream_t);
cudaError_t gs_launch_order_tiles(int, const uint2*, uint32_t*, cudaStream_t);
cudaError_t gs_launch_blend_forward(const GsDims&, const uint32_t*, const uint2*, const uint32_t*, const GsGeom*,
                                    const float[3], float*, float4*, uint32_t*, float4*, uint2*, uint32_t*, uint32_t*,
                                    bool, cudaStream_t);
cudaError_t gs_launch_blend_backward(const GsDims&, const uint2*, const uint32_t*, const GsGeom*, const float[3],
                                     const float4*, const uint32_t*, const float*, const float4*, const uint2*,
                                     const uint32_t*, uint32_t*, const uint32_t*, float4*, int, bool, cudaStream_t);
cudaError_t gs_launch_preprocess_backward(const gs_settings&, const gs_gaussians&, const gs_gradients&, int, int,
                                          const int*, const uint8_t*, const float4*, cudaStream_t);
cudaError_t gs_launch_preprocess_backward_adam(const gs_settings&, const gs_gaussians&, float*, int, const int*,
                                               const uint8_t*, const float4*, float* const[6], float* const[6],
                                               const double[6], const long long[6], double, double, double, int*,
                                               float*, float*, const float*, int*, cudaStream_t);
cudaError_t gs_launch_densify_plan(int, const float*, const float*, const float*, const float*, float, float, float,
                                   float, float, uint8_t*, uint32_t*, uint32_t*, uint32_t*, cudaStream_t);
cudaError_t gs_launch_densify_apply(int, const uint8_t*, const uint32_t*, const uint32_t*, const int[6], float* const[6],
                                    float* const[6], float* const[6], float* const[6], float* const[6],
                                    float* const[6], unsigned long long, cudaStream_t);
cudaError_t gs_launch_reset_opacity(int, float*, float*, float*, cudaStream_t);
cudaError_t gs_launch_preprocess_backward_routed(const gs_settings&, const gs_gaussians&, const gs_gradients&, int, int,
                                                 int, int, float* const[], const int*, const uint8_t*, const float4*,
                                                 cudaStream_t);
cudaError_t gs_launch_adam_peer(int, int, int, int, float* const[], const float* const[], float* const[6],
                                float* const[6], const int[6], const double[6], const long long[6], double, double,
                                double, const int*, const float*, int*, float*, float*, const float*, int, int max_blocks, int local_mask,
                                cudaStream_t);
cudaError_t gs_launch_adam(int, float* const[6], float* const[6], float* const[6], float* const[6], const int[6],
                           const double[6], const long long[6], double, double, double, const int*, const float*, int*,
                           float*, float*, int, const float*, cudaStream_t);
cudaError_t gs_launch_adam_counts(int, int, float* const[], float* const[], float* const[], float* const[],
                                  const size_t[], const double[], const long long[], double, double, double,
                                  const int*, const float*, int*, float*, float*, int, const float*, cudaStream_t);
cudaError_t gs_launch_l1_loss(size_t, const float*, const float*, float, float*, double*, float*, cudaStream_t);
cudaError_t gs_launch_count_nonfinite(const float*, size_t, uint32_t*, cudaStream_t);
cudaError_t gs_launch_peer_barrier(int, int, uint32_t* const[], uint32_t, const float*, float*, cudaStream_t);
size_t gs_ssim_workspace_bytes(int H, int W);
cudaError_t gs_launch_photometric_loss(int, int, const float*, const float*, float, float*, double*, float*, float*,
                                       cudaStream_t);

// ---- context ----------------------------------------------------------------------------------
struct DevBuf {
  void* ptr = nullptr;
  size_t cap = 0;
};

struct gs_ctx {
  int device = 0;
  std::string err;
  // per-Gaussian
  DevBuf geom, depths, radii, tiles, rects, clamped, grad_geom;
  DevBuf sort_k0, sort_k1, sort_v0, sort_v1, dup_sums;  // depth argsort ping-pong (order = sort_v1), scan partials
  DevBuf depth_keys;                                    // sort keys written ahead of K1 (depth_key_kernel)
  // the depth argsort runs beside K1 on this stream (fork after the counters are cleared, join before the scan)
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool overlap_depth_sort = true;
  // colour gate (gs_set_color_gate): K1 leaves the SH -> RGB evaluation to a kernel that runs just ahead of the
  // forward blend, after this event -- the multi-GPU step records it once the peers have delivered the new SH
  // coefficients, so that their all-gather overlaps K1 and the binning stages
  cudaEvent_t color_gate = nullptr;
  cudaEvent_t ev_k1 = nullptr, ev_color = nullptr;  // K1 done (the colour kernel needs its radii) / colours complete
  cudaStream_t color_stream = nullptr;              // default priority (the sort's stream is high priority: a 12 500-block
                                                    // kernel there would push the duplicate stage aside, not run beside it)
  int peer_side_grid = 0;  // GS_PEER_ADAM_GRID: grid cap of a gs_adam_step_peer call flagged GS_PEER_BESIDE_FORWARD (0: automatic)
  // per-duplicate: tile ids (u16, u32 beyond 65536 tiles) and Gaussian ids, ping-pong
  DevBuf keys_a, keys_b, vals_a, vals_b;
  DevBuf radix_counts;
  // one-sweep radix passes (radix_sort.cuh): 16 ticket counters + the digit histograms of the depth argsort (4 x 256)
  // and of the tile sort (4 x 256), cleared at the start of every forward; per-tile status words (tagged, never
  // cleared after the allocation); the next unused tag
  DevBuf sort_meta, sort_status;
  uint32_t sort_tag = 8;
  // tuning (environment, read at gs_create): pairs per thread of a tile (8 / 16), persistent grids, poll back-off
  int depth_sort_items = 16, tile_sort_items = 16, depth_sort_grid = 592;  // (8 / 8: 1.379 ms per step against 1.346, profiles/r02e)
  uint32_t sort_sleep_ns = 0;
  int radix_grid = 296;
  int key_bytes = 2;
  const void* sorted_tiles = nullptr;    // tile ids of the sorted list
  const uint32_t* point_list = nullptr;  // Gaussian ids of the sorted list
  // per-tile / per-pixel
  DevBuf ranges, final_state, n_contrib, tile_order, tile_last;
  // segmented backward: per-segment pixel checkpoints, (tile, segment) work list, {work_count, work_cursor}
  DevBuf ckpt, work_items, counters;
  int num_sms = 148;
  DevBuf loss_acc, ssim_work, tail_grads;
  DevBuf dz_codes, dz_counts;  // densification plan: per-Gaussian outcome codes, per-block counts -> offsets + totals
  int dz_N = -1;               // N of the pending plan (gs_densify_plan -> gs_densify_apply)
  uint32_t* d_host_pinned = nullptr;  // readback: bytes 0..7 D (u64), 8..11 overflow flag (written by the scan kernel)
  uint32_t* d_host_mapped = nullptr;  // device address of the same buffer
  cudaEvent_t ev_zero = nullptr;      // grad_geom cleared on the side stream during the forward
  bool grad_geom_zeroed = false;
  bool work_cursor_zeroed = false;    // counters[1] still zero from the forward's clear
  uint64_t dup_cap_fixed = 0;         // > 0: caller-fixed pair capacity, no host synchronisation in gs_forward
  float* overflow_mirror = nullptr;   // optional caller-owned device float raised to 1.0 on overflow
  bool D_pending = false;             // D of the last forward not yet read back
  bool reuse_blend_gradients = false; // test hook: gs_backward_adam_step consumes the records of the last gs_backward
  // multi-GPU gradient routing (gs_set_gradient_routing): rows [q*R, (q+1)*R) of the six parameter gradients go
  // to route_table[q*6 + k] (the slot this rank owns inside rank q's buffer; q == route_rank: local memory)
  int route_shards = 0, route_rows = 0, route_rank = 0;
  float* route_table[8 * 6] = {};
  // state of the last forward
  int N = -1, H = 0, W = 0, ntiles = 0;
  uint64_t D = 0;
  bool have_forward = false, have_backward = false;
  // profiling
  bool profiling = false;
  cudaEvent_t ev[2 * GS_NUM_STAGES] = {};
  bool ev_valid[GS_NUM_STAGES] = {};
};

static std::mutex g_err_mu;
static std::string g_last_error;

void gs_set_error(gs_ctx* ctx, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_last_error = buf;
}

static int ensure(gs_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return GS_OK;
  if (b.ptr) cudaFree(b.ptr);
  b.ptr = nullptr;
  b.cap = 0;
  size_t want = bytes + bytes / 4 + 256;
  cudaError_t e = cudaMalloc(&b.ptr, want);
  if (e != cudaSuccess) {
    cudaGetLastError();
    e = cudaMalloc(&b.ptr, bytes);  // retry without slack
    want = bytes;
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    gs_set_error(ctx, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    return GS_ENOMEM;
  }
  b.cap = want;
  return GS_OK;
}
#define GS_ENSURE(ctx, buf, bytes)                   \
  do {                                               \
    int _r = ensure(ctx, buf, bytes);                \
    if (_r != GS_OK) return _r;                      \
  } while (0)

struct StageTimer {
  gs_ctx* c;
  int stage;
  cudaStream_t s;
  StageTimer(gs_ctx* ctx, int st, cudaStream_t stream) : c(ctx), stage(st), s(stream) {
    if (c->profiling) cudaEventRecord(c->ev[2 * stage], s);
  }
  ~StageTimer() {
    if (c->profiling) {
      cudaEventRecord(c->ev[2 * stage + 1], s);
      c->ev_valid[stage] = true;
    }
  }
};

extern "C" {

const char* gs_version(void) { return "gsplat_b200 0.1 (sm_100a)"; }

const char* gs_last_error(const gs_ctx* ctx) {
  if (ctx) return ctx->err.c_str();
  std::lock_guard<std::mutex> lk(g_err_mu);
  static thread_local std::string copy;
  copy = g_last_error;
  return copy.c_str();
}

int gs_create(gs_ctx** out, int device) {
  if (!out) return GS_EINVAL;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    gs_set_error(nullptr, "no CUDA device available (%s); libgsplat_b200 has no CPU fallback",
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    return GS_ECUDA;
  }
  if (device < 0 || device >= count) {
    gs_set_error(nullptr, "device %d out of range [0,%d)", device, count);
    return GS_EINVAL;
  }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess || prop.major != 10) {
    gs_set_error(nullptr, "device %d is sm_%d%d; this library contains sm_100a code only", device,
                 e == cudaSuccess ? prop.major : 0, e == cudaSuccess ? prop.minor : 0);
    return GS_ECUDA;
  }
  gs_ctx* c = new (std::nothrow) gs_ctx();
  if (!c) return GS_ENOMEM;
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  c->radix_grid = 4 * c->num_sms < 608 ? 4 * c->num_sms : 608;  // RS_MAX_GRID
  {
    const char* env = getenv("GS_RADIX_GRID");  // tuning switch: persistent grid of the tile sort
    if (env && atoi(env) >= 1 && atoi(env) <= 1024) c->radix_grid = atoi(env);
  }
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaHostAlloc((void**)&c->d_host_pinned, 64, cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer((void**)&c->d_host_mapped, c->d_host_pinned, 0) != cudaSuccess) {
    gs_set_error(nullptr, "cudaSetDevice/cudaMallocHost failed: %s", cudaGetErrorString(cudaGetLastError()));
    delete c;
    return GS_ECUDA;
  }
  for (int i = 0; i < 2 * GS_NUM_STAGES; ++i) cudaEventCreate(&c->ev[i]);
  // highest priority: the sort's twelve small launches must not queue behind K1's 12 500 blocks
  int prio_lo = 0, prio_hi = 0;
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  if (cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_zero, cudaEventDisableTiming) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->color_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_k1, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_color, cudaEventDisableTiming) != cudaSuccess) {
    gs_set_error(nullptr, "cudaStreamCreate/cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
    gs_destroy(c);
    return GS_ECUDA;
  }
  {
    const char* e1 = getenv("GS_DEPTH_SORT_ITEMS");
    const char* e2 = getenv("GS_TILE_SORT_ITEMS");
    const char* e3 = getenv("GS_DEPTH_SORT_GRID");
    const char* e4 = getenv("GS_SORT_SLEEP_NS");
    if (e1 && atoi(e1) == 8) c->depth_sort_items = 8;
    if (e2 && atoi(e2) == 8) c->tile_sort_items = 8;
    c->depth_sort_grid = c->radix_grid;
    if (e3 && atoi(e3) >= 1 && atoi(e3) <= 1024) c->depth_sort_grid = atoi(e3);
    if (e4 && atoi(e4) >= 0 && atoi(e4) <= 10000) c->sort_sleep_ns = (uint32_t)atoi(e4);
  }
  c->peer_side_grid = 0;  // 0: from the rank count (gs_peer_side_grid)
  {
    const char* env = getenv("GS_PEER_ADAM_GRID");  // tuning switch: blocks of an Adam call that runs beside a forward
    if (env && atoi(env) >= 1) c->peer_side_grid = atoi(env);
  }
  {
    const char* env = getenv("GS_NO_SORT_OVERLAP");  // measurement switch: run the depth argsort after K1 on the caller's stream
    if (env && env[0] == '1') c->overlap_depth_sort = false;
  }
  *out = c;
  return GS_OK;
}

void gs_destroy(gs_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  DevBuf* bufs[] = {&c->geom,   &c->depths, &c->radii,  &c->tiles,   &c->rects, &c->clamped,   &c->grad_geom, &c->sort_k0,
                    &c->sort_k1, &c->sort_v0, &c->sort_v1, &c->dup_sums, &c->radix_counts, &c->depth_keys, &c->sort_meta, &c->sort_status,
                    &c->keys_a, &c->keys_b, &c->vals_a, &c->vals_b,  &c->ranges,    &c->final_state,  &c->n_contrib,
                    &c->tile_order, &c->tile_last, &c->ckpt, &c->work_items, &c->counters, &c->loss_acc, &c->ssim_work,
                    &c->tail_grads, &c->dz_codes,
                    &c->dz_counts};
  for (DevBuf* b : bufs)
    if (b->ptr) cudaFree(b->ptr);
  if (c->d_host_pinned) cudaFreeHost(c->d_host_pinned);
  if (c->side) cudaStreamDestroy(c->side);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->ev_zero) cudaEventDestroy(c->ev_zero);
  if (c->color_stream) cudaStreamDestroy(c->color_stream);
  if (c->ev_k1) cudaEventDestroy(c->ev_k1);
  if (c->ev_color) cudaEventDestroy(c->ev_color);
  for (int i = 0; i < 2 * GS_NUM_STAGES; ++i)
    if (c->ev[i]) cudaEventDestroy(c->ev[i]);
  delete c;
}

int gs_set_profiling(gs_ctx* c, int enabled) {
  if (!c) return GS_EINVAL;
  c->profiling = enabled != 0;
  for (int i = 0; i < GS_NUM_STAGES; ++i) c->ev_valid[i] = false;
  return GS_OK;
}

int gs_get_stage_ms(gs_ctx* c, float ms[GS_NUM_STAGES]) {
  if (!c || !ms) return GS_EINVAL;
  cudaSetDevice(c->device);
  for (int i = 0; i < GS_NUM_STAGES; ++i) {
    ms[i] = 0.f;
    if (c->ev_valid[i]) {
      cudaEventSynchronize(c->ev[2 * i + 1]);
      cudaEventElapsedTime(&ms[i], c->ev[2 * i], c->ev[2 * i + 1]);
    }
  }
  return GS_OK;
}

static int validate(gs_ctx* c, const gs_settings* st, int N, const gs_gaussians* in) {
  if (!c) return GS_EINVAL;
  if (!st || !in || N < 0) {
    gs_set_error(c, "null settings/inputs or negative N");
    return GS_EINVAL;
  }
  if (st->H <= 0 || st->W <= 0 || !(st->tanfovx > 0.f) || !(st->tanfovy > 0.f)) {
    gs_set_error(c, "invalid image size %dx%d or tanfov", st->W, st->H);
    return GS_EINVAL;
  }
  if (st->max_sh_degree < 0 || st->max_sh_degree > 3 || st->sh_degree < 0 || st->sh_degree > st->max_sh_degree) {
    gs_set_error(c, "sh_degree %d / max_sh_degree %d out of range (0..3)", st->sh_degree, st->max_sh_degree);
    return GS_EINVAL;
  }
  if (N > 0) {
    if (!in->means3D || !in->opacities) {
      gs_set_error(c, "means3D and opacities are required");
      return GS_EINVAL;
    }
    if (!in->cov3D_precomp && (!in->scales || !in->rotations)) {
      gs_set_error(c, "scales and rotations are required unless cov3D_precomp is given");
      return GS_EINVAL;
    }
    if (!in->colors_precomp) {
      if (!in->sh_dc || in->sh_dc_stride < 3) {
        gs_set_error(c, "sh_dc (stride >= 3) is required unless colors_precomp is given");
        return GS_EINVAL;
      }
      const int need = 3 * ((st->sh_degree + 1) * (st->sh_degree + 1) - 1);
      if (need > 0 && (!in->sh_rest || in->sh_rest_stride < need)) {
        gs_set_error(c, "sh_rest with stride >= %d floats is required for sh_degree %d", need, st->sh_degree);
        return GS_EINVAL;
      }
    }
    if (in->rotations && (((uintptr_t)in->rotations) & 15)) {
      gs_set_error(c, "rotations must be 16-byte aligned");
      return GS_EINVAL;
    }
  }
  return GS_OK;
}

int gs_forward(gs_ctx* c, const gs_settings* st, int N, const gs_gaussians* in, float* out_color, int* radii,
               void* stream_) {
  int rc = validate(c, st, N, in);
  if (rc != GS_OK) return rc;
  if (!out_color) {
    gs_set_error(c, "out_color is required");
    return GS_EINVAL;
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  const GsDims d = gs_dims(*st);
  const size_t P = (size_t)d.H * d.W;
  const size_t Nn = (size_t)(N > 0 ? N : 1);
  c->have_forward = false;
  c->have_backward = false;

  GS_ENSURE(c, c->geom, Nn * sizeof(GsGeom));
  GS_ENSURE(c, c->depths, Nn * 4);
  GS_ENSURE(c, c->radii, Nn * 4);
  GS_ENSURE(c, c->tiles, Nn * 4);
  GS_ENSURE(c, c->rects, Nn * 16);
  GS_ENSURE(c, c->clamped, Nn);
  GS_ENSURE(c, c->depth_keys, Nn * 4);
  GS_ENSURE(c, c->grad_geom, Nn * 48);
  c->grad_geom_zeroed = false;
  GS_ENSURE(c, c->sort_k0, Nn * 4);
  GS_ENSURE(c, c->sort_k1, Nn * 4);
  GS_ENSURE(c, c->sort_v0, Nn * 4);
  GS_ENSURE(c, c->sort_v1, Nn * 4);
  GS_ENSURE(c, c->dup_sums, (size_t)(gs_dup_num_blocks(N > 0 ? N : 1) + 1) * 4);
  GS_ENSURE(c, c->sort_meta, 64 + 8 * 256 * 4);
  GS_ENSURE(c, c->ranges, (size_t)d.ntiles * sizeof(uint2));
  GS_ENSURE(c, c->final_state, P * 16);
  GS_ENSURE(c, c->tile_order, (size_t)d.ntiles * 4);
  GS_ENSURE(c, c->tile_last, (size_t)d.ntiles * 8);  // running maximum + arrival counter per tile
  GS_ENSURE(c, c->counters, 64);
  GS_ENSURE(c, c->n_contrib, P * 4);
  // device counters: [0] backward work items, [1] backward cursor, bytes 8..15 D (u64), [4] D (u32, clamped to the
  // capacity), [5] overflow flag, [6] overflow guard (float)
  uint32_t* cnt = (uint32_t*)c->counters.ptr;
  unsigned long long* D64_dev = (unsigned long long*)(cnt + 2);
  uint32_t* D_dev = cnt + 4;
  uint32_t* overflow_dev = cnt + 5;
  float* guard_dev = (float*)(cnt + 6);  // 1.0f when this forward overflowed its pair capacity
  c->D_pending = false;
  c->key_bytes = d.ntiles <= 65536 ? 2 : 4;
  int tile_bits = 1;
  while ((1u << tile_bits) < (uint32_t)d.ntiles) ++tile_bits;

  uint64_t D = 0;
  bool tile_last_zeroed = false;
  // K1 stores the radii into the caller's array as well (a copy operation less on the stream) when it is 16-byte aligned
  int* radii_user = (radii && (((uintptr_t)radii) & 15) == 0) ? radii : nullptr;
  GS_CUDA_CHECK(c, cudaMemsetAsync(c->counters.ptr, 0, 64, stream));
  c->work_cursor_zeroed = true;
  // one-sweep sorts: tickets and digit histograms start at zero; eight fresh tags (depth passes, tile passes)
  GS_CUDA_CHECK(c, cudaMemsetAsync(c->sort_meta.ptr, 0, 64 + 8 * 256 * 4, stream));
  uint32_t* sort_tickets = (uint32_t*)c->sort_meta.ptr;
  uint32_t* depth_hist = sort_tickets + 16;
  uint32_t* tile_hist = depth_hist + 4 * 256;
  const uint32_t tag0 = c->sort_tag;
  c->sort_tag = c->sort_tag + 8 >= (1u << 30) ? 8u : c->sort_tag + 8;
  auto ensure_status = [&](size_t words) -> int {
    if (words * 8 <= c->sort_status.cap) return GS_OK;
    const int r = ensure(c, c->sort_status, words * 8);
    if (r != GS_OK) return r;
    // a status word is valid only if it carries the tag of the launch reading it: fresh memory must not by accident
    if (cudaMemsetAsync(c->sort_status.ptr, 0, c->sort_status.cap, stream) != cudaSuccess) return GS_ECUDA;
    return GS_OK;
  };
  {
    const int r = ensure_status(gs_onesweep_status_words(Nn, 8));
    if (r != GS_OK) return r;
  }
  const int defer_color = (c->color_gate != nullptr && !in->colors_precomp) ? 1 : 0;
  bool color_on_side = false;
  if (!defer_color) c->ev_valid[8] = false;  // (no colour kernel in this forward: no stale stage time)
  if (N > 0) {
    // The depth argsort reads only the view depth, so it is forked onto the context's second stream and runs
    // beside K1 (the latency-bound N-sized sort fills the issue slots K1 leaves); the scan joins the two.
    const bool overlap = c->overlap_depth_sort;
    cudaStream_t sort_stream = overlap ? c->side : stream;
    if (overlap) {
      GS_CUDA_CHECK(c, cudaEventRecord(c->ev_fork, stream));
      GS_CUDA_CHECK(c, cudaStreamWaitEvent(c->side, c->ev_fork, 0));
      // K3's per-tile running maximum + arrival counters, cleared off the critical path (joined with the sort)
      GS_CUDA_CHECK(c, cudaMemsetAsync(c->tile_last.ptr, 0, (size_t)d.ntiles * 8, c->side));
      tile_last_zeroed = true;
    } else {
      StageTimer t(c, 0, stream);
      GS_CUDA_CHECK(c, gs_launch_preprocess(*st, *in, N, (GsGeom*)c->geom.ptr, (float*)c->depths.ptr,
                                            (int*)c->radii.ptr, (uint32_t*)c->tiles.ptr, (uint4*)c->rects.ptr,
                                            (uint8_t*)c->clamped.ptr, radii_user, defer_color, stream));
    }
    {
      StageTimer t(c, 1, sort_stream);
      GS_CUDA_CHECK(c, gs_launch_depth_keys(*st, in->means3D, N, (uint32_t*)c->depth_keys.ptr, depth_hist, sort_stream));
      GS_CUDA_CHECK(c, gs_launch_depth_sort(N, (const uint32_t*)c->depth_keys.ptr, (uint32_t*)c->sort_k0.ptr,
                                            (uint32_t*)c->sort_k1.ptr, (uint32_t*)c->sort_v0.ptr,
                                            (uint32_t*)c->sort_v1.ptr, depth_hist,
                                            (unsigned long long*)c->sort_status.ptr, sort_tickets, tag0,
                                            c->depth_sort_grid, c->depth_sort_items, c->sort_sleep_ns, sort_stream));
    }
    if (overlap) {
      GS_CUDA_CHECK(c, cudaEventRecord(c->ev_join, c->side));
      // the backward's accumulation records are cleared here, off the critical path (consumed by gs_backward*)
      GS_CUDA_CHECK(c, cudaMemsetAsync(c->grad_geom.ptr, 0, (size_t)N * 48, c->side));
      GS_CUDA_CHECK(c, cudaEventRecord(c->ev_zero, c->side));
      c->grad_geom_zeroed = true;
      {
        StageTimer t(c, 0, stream);
        GS_CUDA_CHECK(c, gs_launch_preprocess(*st, *in, N, (GsGeom*)c->geom.ptr, (float*)c->depths.ptr,
                                              (int*)c->radii.ptr, (uint32_t*)c->tiles.ptr, (uint4*)c->rects.ptr,
                                              (uint8_t*)c->clamped.ptr, radii_user, defer_color, stream));
      }
      if (defer_color) {
        // deferred SH -> RGB on a stream of its own, beside the duplicate stage and the tile sort (HBM-bound next to
        // latency-bound work): after K1 (its radii), after the gate (the caller recorded it BEFORE this call)
        GS_CUDA_CHECK(c, cudaEventRecord(c->ev_k1, stream));
        GS_CUDA_CHECK(c, cudaStreamWaitEvent(c->color_stream, c->ev_k1, 0));
        GS_CUDA_CHECK(c, cudaStreamWaitEvent(c->color_stream, c->color_gate, 0));
        {
          StageTimer t(c, 8, c->color_stream);
          GS_CUDA_CHECK(c, gs_launch_sh_color(*st, *in, N, (const int*)c->radii.ptr, (GsGeom*)c->geom.ptr,
                                              (uint8_t*)c->clamped.ptr, c->color_stream));
        }
        GS_CUDA_CHECK(c, cudaEventRecord(c->ev_color, c->color_stream));
        color_on_side = true;
      }
      GS_CUDA_CHECK(c, cudaStreamWaitEvent(stream, c->ev_join, 0));
    }
    {
      StageTimer t(c, 2, stream);
      const bool async_mode = c->dup_cap_fixed > 0;
      const unsigned long long cap = async_mode ? c->dup_cap_fixed : 0xFFFFFFF0ull;
      GS_CUDA_CHECK(c, gs_launch_dup_scan(N, (const uint32_t*)c->sort_v1.ptr, (const uint32_t*)c->tiles.ptr,
                                          (uint32_t*)c->dup_sums.ptr, D64_dev, D_dev, cap, overflow_dev, guard_dev,
                                          c->overflow_mirror, c->d_host_mapped, stream));
      if (async_mode) {
        // caller-fixed capacity: nothing to wait for; an overflow poisons the step (guard) and is reported by
        // gs_get_forward_status once the caller has synchronised
        D = cap;
        c->D_pending = true;
      } else {
        // the one host synchronisation of the forward: number of (Gaussian, tile) duplicates, to size the lists
        GS_CUDA_CHECK(c, cudaStreamSynchronize(stream));
        D = *(const unsigned long long*)c->d_host_pinned;
        if (D > 0xFFFFFFF0ull) {
          gs_set_error(c, "duplicate list too long for 32-bit indices (%llu pairs)", (unsigned long long)D);
          return GS_EOVERFLOW;
        }
      }
      const size_t Dn = (size_t)(D > 0 ? D : 1);
      GS_ENSURE(c, c->keys_a, Dn * c->key_bytes);
      GS_ENSURE(c, c->keys_b, Dn * c->key_bytes);
      GS_ENSURE(c, c->vals_a, Dn * 4);
      GS_ENSURE(c, c->vals_b, Dn * 4);
      {
        const int r = ensure_status(gs_onesweep_status_words(Dn, 8));  // (the depth argsort has joined)
        if (r != GS_OK) return r;
      }
      if (D > 0)
        GS_CUDA_CHECK(c, gs_launch_duplicate(N, (const uint32_t*)c->sort_v1.ptr, (const uint4*)c->rects.ptr,
                                             (const uint32_t*)c->dup_sums.ptr, c->keys_a.ptr, c->key_bytes,
                                             (uint32_t*)c->vals_a.ptr, d.gx, (uint32_t)D, overflow_dev, tile_hist,
                                             stream));
    }
    {
      StageTimer t(c, 3, stream);
      int in_b = 0;
      GS_CUDA_CHECK(c, gs_launch_tile_sort(D_dev, (uint32_t)D, tile_bits, c->key_bytes, c->keys_a.ptr, c->keys_b.ptr,
                                           (uint32_t*)c->vals_a.ptr, (uint32_t*)c->vals_b.ptr, tile_hist,
                                           (unsigned long long*)c->sort_status.ptr, sort_tickets + 8, tag0 + 4,
                                           c->radix_grid, c->tile_sort_items, c->sort_sleep_ns, &in_b, stream));
      c->sorted_tiles = in_b ? c->keys_b.ptr : c->keys_a.ptr;
      c->point_list = (const uint32_t*)(in_b ? c->vals_b.ptr : c->vals_a.ptr);
    }
  } else {
    GS_ENSURE(c, c->keys_a, 8);
    GS_ENSURE(c, c->vals_a, 4);
    *(unsigned long long*)c->d_host_pinned = 0ull;
    c->d_host_pinned[2] = 0u;
    c->sorted_tiles = c->keys_a.ptr;
    c->point_list = (const uint32_t*)c->vals_a.ptr;
  }
  {
    StageTimer t(c, 4, stream);
    GS_CUDA_CHECK(c, gs_launch_ranges(D_dev, c->sorted_tiles, c->key_bytes, (uint2*)c->ranges.ptr, d.ntiles, stream));
    GS_CUDA_CHECK(c, gs_launch_order_tiles(d.ntiles, (const uint2*)c->ranges.ptr, (uint32_t*)c->tile_order.ptr,
                                           stream));
  }
  // segment bookkeeping of the backward: at most D/GS_SEG + ntiles segments over all tiles
  const size_t max_segs = (size_t)(D >> GS_SEG_SHIFT) + (size_t)d.ntiles;
  GS_ENSURE(c, c->work_items, max_segs * sizeof(uint2));
  GS_ENSURE(c, c->ckpt, max_segs * GS_TILE_PIX * sizeof(float4));
  if (color_on_side) {
    GS_CUDA_CHECK(c, cudaStreamWaitEvent(stream, c->ev_color, 0));
  } else if (defer_color && N > 0) {
    // (GS_NO_SORT_OVERLAP: no second stream) wait for the gate, then fill colours / clamp masks in
    GS_CUDA_CHECK(c, cudaStreamWaitEvent(stream, c->color_gate, 0));
    StageTimer t(c, 8, stream);
    GS_CUDA_CHECK(c, gs_launch_sh_color(*st, *in, N, (const int*)c->radii.ptr, (GsGeom*)c->geom.ptr,
                                        (uint8_t*)c->clamped.ptr, stream));
  }
  {
    StageTimer t(c, 5, stream);
    GS_CUDA_CHECK(c, gs_launch_blend_forward(d, (const uint32_t*)c->tile_order.ptr, (const uint2*)c->ranges.ptr,
                                             c->point_list, (const GsGeom*)c->geom.ptr, st->bg,
                                             out_color, (float4*)c->final_state.ptr, (uint32_t*)c->n_contrib.ptr,
                                             (float4*)c->ckpt.ptr, (uint2*)c->work_items.ptr,
                                             (uint32_t*)c->counters.ptr, (uint32_t*)c->tile_last.ptr, tile_last_zeroed, stream));
  }
  if (radii && !radii_user && N > 0)
    GS_CUDA_CHECK(c, cudaMemcpyAsync(radii, c->radii.ptr, (size_t)N * 4, cudaMemcpyDeviceToDevice, stream));
  if (st->debug & GS_DEBUG_CHECKS) {
    GS_CUDA_CHECK(c, cudaStreamSynchronize(stream));
    GS_CUDA_CHECK(c, cudaGetLastError());
  }
  c->N = N;
  c->H = d.H;
  c->W = d.W;
  c->ntiles = d.ntiles;
  c->D = D;
  c->have_forward = true;
  return GS_OK;
}

// debug: count NaN / inf in a device array; GS_ENUMERIC with a message naming it when there are any
static int scan_nonfinite(gs_ctx* c, const float* p, size_t n, const char* name, cudaStream_t stream) {
  if (!p || n == 0) return GS_OK;
  GS_ENSURE(c, c->loss_acc, 64);
  uint32_t* cnt = (uint32_t*)c->loss_acc.ptr + 8;
  GS_CUDA_CHECK(c, cudaMemsetAsync(cnt, 0, 4, stream));
  GS_CUDA_CHECK(c, gs_launch_count_nonfinite(p, n, cnt, stream));
  uint32_t h = 0;
  GS_CUDA_CHECK(c, cudaMemcpyAsync(&h, cnt, 4, cudaMemcpyDeviceToHost, stream));
  GS_CUDA_CHECK(c, cudaStreamSynchronize(stream));
  if (h) {
    gs_set_error(c, "debug: %u non-finite values (NaN / inf) in %s", h, name);
    return GS_ENUMERIC;
  }
  return GS_OK;
}

int gs_backward(gs_ctx* c, const gs_settings* st, int N, const gs_gaussians* in, const float* dL_dout_color,
                const gs_gradients* grads, void* stream_) {
  int rc = validate(c, st, N, in);
  if (rc != GS_OK) return rc;
  if (!dL_dout_color || !grads) {
    gs_set_error(c, "dL_dout_color and grads are required");
    return GS_EINVAL;
  }
  if (!c->have_forward || c->N != N || c->H != st->H || c->W != st->W) {
    gs_set_error(c, "gs_backward needs the matching gs_forward (same N, H, W) to have run on this context");
    return GS_ESTATE;
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  if (N == 0) return GS_OK;
  const GsDims d = gs_dims(*st);
  GS_ENSURE(c, c->grad_geom, (size_t)N * 48);
  if (!(c->reuse_blend_gradients && c->have_backward)) {
    StageTimer t(c, 6, stream);
    if (c->grad_geom_zeroed) {
      GS_CUDA_CHECK(c, cudaStreamWaitEvent(stream, c->ev_zero, 0));
      c->grad_geom_zeroed = false;
    } else {
      GS_CUDA_CHECK(c, cudaMemsetAsync(c->grad_geom.ptr, 0, (size_t)N * 48, stream));
    }
    const bool cursor_fresh = c->work_cursor_zeroed;  // cleared with the counters by the forward, consumed once
    c->work_cursor_zeroed = false;
    GS_CUDA_CHECK(c, gs_launch_blend_backward(d, (const uint2*)c->ranges.ptr, c->point_list,
                                              (const GsGeom*)c->geom.ptr, st->bg, (const float4*)c->final_state.ptr,
                                              (const uint32_t*)c->n_contrib.ptr, dL_dout_color,
                                              (const float4*)c->ckpt.ptr, (const uint2*)c->work_items.ptr,
                                              (const uint32_t*)c->counters.ptr, (uint32_t*)c->counters.ptr + 1,
                                              (const uint32_t*)c->tile_last.ptr, (float4*)c->grad_geom.ptr,
                                              c->num_sms, cursor_fresh, stream));
  }
  {
    StageTimer t(c, 7, stream);
    if (c->route_shards == 0) {
      GS_CUDA_CHECK(c, gs_launch_preprocess_backward(*st, *in, *grads, 0, N, (const int*)c->radii.ptr,
                                                     (const uint8_t*)c->clamped.ptr,
                                                     (const float4*)c->grad_geom.ptr, stream));
    } else {
      // one launch for all owner ranks: consecutive blocks store to different owners' slots (PbRoute)
      if (grads->accumulate || in->colors_precomp || in->cov3D_precomp || !in->sh_rest) {
        gs_set_error(c, "gradient routing needs the raw split-SH parameter layout and accumulate == 0");
        return GS_EINVAL;
      }
      GS_CUDA_CHECK(c, gs_launch_preprocess_backward_routed(*st, *in, *grads, N, c->route_shards, c->route_rank,
                                                            c->route_rows, c->route_table, (const int*)c->radii.ptr,
                                                            (const uint8_t*)c->clamped.ptr,
                                                            (const float4*)c->grad_geom.ptr, stream));
    }
  }
  if (st->debug & GS_DEBUG_CHECKS) {
    GS_CUDA_CHECK(c, cudaStreamSynchronize(stream));
    GS_CUDA_CHECK(c, cudaGetLastError());
    // anomaly scan (the role of --detect_anomaly, src/train.cpp:39): the blend-gradient records and every sink
    const int kmax = (st->max_sh_degree + 1) * (st->max_sh_degree + 1);
    const struct { const float* p; size_t n; const char* name; } sinks[] = {
        {(const float*)c->grad_geom.ptr, (size_t)N * 12, "blend gradients (dL/dmean2D, conic, opacity, colour)"},
        {c->route_shards ? nullptr : grads->dL_dmeans3D, (size_t)N * 3, "dL_dmeans3D"},
        {c->route_shards ? nullptr : grads->dL_dscales, (size_t)N * 3, "dL_dscales"},
        {c->route_shards ? nullptr : grads->dL_drotations, (size_t)N * 4, "dL_drotations"},
        {c->route_shards ? nullptr : grads->dL_dopacities, (size_t)N, "dL_dopacities"},
        {grads->dL_dmeans2D, (size_t)N * 2, "dL_dmeans2D"},
        {grads->dL_dcov3D, (size_t)N * 6, "dL_dcov3D"},
        {grads->dL_dcolors, (size_t)N * 3, "dL_dcolors"},
        {(!c->route_shards && grads->sh_dc_stride == 3) ? grads->dL_dsh_dc : nullptr, (size_t)N * 3, "dL_dsh_dc"},
        {(!c->route_shards && kmax > 1 && grads->sh_rest_stride == 3 * (kmax - 1)) ? grads->dL_dsh_rest : nullptr,
         (size_t)N * 3 * (kmax - 1), "dL_dsh_rest"},
        {(!c->route_shards && grads->dL_dsh_dc && grads->sh_dc_stride == 3 * kmax && grads->sh_dc_stride != 3) ? grads->dL_dsh_dc : nullptr,
         (size_t)N * 3 * kmax, "dL_dshs"},
    };
    for (const auto& s : sinks) {
      const int r = scan_nonfinite(c, s.p, s.n, s.name, stream);
      if (r != GS_OK) return r;
    }
  }
  c->have_backward = true;
  return GS_OK;
}

int gs_backward_adam_step(gs_ctx* c, const gs_settings* st, int N, const gs_gaussians* in, const float* dL_dout_color,
                          float* dL_dmeans2D, float* const exp_avg[6], float* const exp_avg_sq[6], const double lr[6],
                          const long long step[6], double beta1, double beta2, double eps, int* max_radii2D,
                          float* xyz_grad_accum, float* denom, const float* skip_guard, void* stream_) {
  int rc = validate(c, st, N, in);
  if (rc != GS_OK) return rc;
  if (!dL_dout_color || !exp_avg || !exp_avg_sq || !lr || !step) {
    gs_set_error(c, "gs_backward_adam_step: null argument");
    return GS_EINVAL;
  }
  if (in->colors_precomp || in->cov3D_precomp || !in->sh_dc || in->sh_dc_stride != 3 || !in->scales || !in->rotations) {
    gs_set_error(c, "gs_backward_adam_step needs the trainable layout: means3D, scales, rotations, opacities, "
                    "sh_dc (N,1,3) and sh_rest (N,K-1,3) as separate dense tensors");
    return GS_EINVAL;
  }
  const int wrest = 3 * ((st->max_sh_degree + 1) * (st->max_sh_degree + 1) - 1);
  if (wrest > 0 && (!in->sh_rest || in->sh_rest_stride != wrest)) {
    gs_set_error(c, "gs_backward_adam_step: sh_rest must be dense (stride %d)", wrest);
    return GS_EINVAL;
  }
  for (int k = 0; k < 6; ++k) {
    if (step[k] < 1) {
      gs_set_error(c, "segment %d: step must be >= 1 (libtorch increments before use)", k);
      return GS_EINVAL;
    }
    if (!(k == 2 && wrest == 0) && (!exp_avg[k] || !exp_avg_sq[k])) {
      gs_set_error(c, "segment %d: exp_avg / exp_avg_sq missing", k);
      return GS_EINVAL;
    }
  }
  if (!c->have_forward || c->N != N || c->H != st->H || c->W != st->W) {
    gs_set_error(c, "gs_backward_adam_step needs the matching gs_forward (same N, H, W) to have run on this context");
    return GS_ESTATE;
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  if (N == 0) return GS_OK;
  const GsDims d = gs_dims(*st);
  GS_ENSURE(c, c->grad_geom, (size_t)N * 48);
  if (!(c->reuse_blend_gradients && c->have_backward)) {
    StageTimer t(c, 6, stream);
    if (c->grad_geom_zeroed) {
      GS_CUDA_CHECK(c, cudaStreamWaitEvent(stream, c->ev_zero, 0));
      c->grad_geom_zeroed = false;
    } else {
      GS_CUDA_CHECK(c, cudaMemsetAsync(c->grad_geom.ptr, 0, (size_t)N * 48, stream));
    }
    const bool cursor_fresh = c->work_cursor_zeroed;  // cleared with the counters by the forward, consumed once
    c->work_cursor_zeroed = false;
    GS_CUDA_CHECK(c, gs_launch_blend_backward(d, (const uint2*)c->ranges.ptr, c->point_list,
                                              (const GsGeom*)c->geom.ptr, st->bg, (const float4*)c->final_state.ptr,
                                              (const uint32_t*)c->n_contrib.ptr, dL_dout_color,
                                              (const float4*)c->ckpt.ptr, (const uint2*)c->work_items.ptr,
                                              (const uint32_t*)c->counters.ptr, (uint32_t*)c->counters.ptr + 1,
                                              (const uint32_t*)c->tile_last.ptr, (float4*)c->grad_geom.ptr,
                                              c->num_sms, cursor_fresh, stream));
  }
  {
    StageTimer t(c, 7, stream);
    int covered = 0;
    GS_CUDA_CHECK(c, gs_launch_preprocess_backward_adam(*st, *in, dL_dmeans2D, N, (const int*)c->radii.ptr,
                                                        (const uint8_t*)c->clamped.ptr,
                                                        (const float4*)c->grad_geom.ptr, exp_avg, exp_avg_sq, lr, step,
                                                        beta1, beta2, eps, max_radii2D, xyz_grad_accum, denom,
                                                        skip_guard, &covered, stream));
    if (covered < N) {
      // ragged tail (< 128 rows) or a layout the fused kernel does not take: separate K4b into scratch gradients,
      // then K5 on those rows.  Scratch rows are addressed with the global row index (re-based pointers).
      const long long b = covered, rows = (long long)N - covered;
      GS_ENSURE(c, c->tail_grads, (size_t)rows * (17 + wrest) * 4 + 7 * 16);
      float* base = (float*)c->tail_grads.ptr;
      auto take = [&](long long floats) {
        float* p = base;
        base += (floats + 3) / 4 * 4;
        return p;
      };
      float* gk[6] = {take(rows * 3), take(rows * 3), take(rows * (wrest > 0 ? wrest : 1)), take(rows),
                      take(rows * 3), take(rows * 4)};
      float* g2 = take(rows * 2);
      gs_gradients g;
      memset(&g, 0, sizeof(g));
      g.dL_dmeans3D = gk[0] - 3 * b;
      g.dL_dsh_dc = gk[1] - 3 * b;
      g.dL_dsh_rest = wrest > 0 ? gk[2] - (long long)wrest * b : nullptr;
      g.dL_dopacities = gk[3] - b;
      g.dL_dscales = gk[4] - 3 * b;
      g.dL_drotations = gk[5] - 4 * b;
      g.dL_dmeans2D = (dL_dmeans2D ? dL_dmeans2D : g2 - 2 * b);
      g.sh_dc_stride = 3;
      g.sh_rest_stride = wrest;
      GS_CUDA_CHECK(c, gs_launch_preprocess_backward(*st, *in, g, (int)b, N, (const int*)c->radii.ptr,
                                                     (const uint8_t*)c->clamped.ptr, (const float4*)c->grad_geom.ptr,
                                                     stream));
      const int widths[6] = {3, 3, wrest, 1, 3, 4};
      float* pk[6] = {const_cast<float*>(in->means3D) + 3 * b, const_cast<float*>(in->sh_dc) + 3 * b,
                      wrest > 0 ? const_cast<float*>(in->sh_rest) + (long long)wrest * b : nullptr,
                      const_cast<float*>(in->opacities) + b, const_cast<float*>(in->scales) + 3 * b,
                      const_cast<float*>(in->rotations) + 4 * b};
      float* mk[6];
      float* vk[6];
      size_t cnt[6];
      for (int k = 0; k < 6; ++k) {
        mk[k] = exp_avg[k] ? exp_avg[k] + (long long)widths[k] * b : nullptr;
        vk[k] = exp_avg_sq[k] ? exp_avg_sq[k] + (long long)widths[k] * b : nullptr;
        cnt[k] = pk[k] ? (size_t)rows * widths[k] : 0;
      }
      const bool stats = max_radii2D && xyz_grad_accum && denom;
      GS_CUDA_CHECK(c, gs_launch_adam_counts((int)rows, 6, pk, gk, mk, vk, cnt, lr, step, beta1, beta2, eps,
                                             stats ? (const int*)c->radii.ptr + b : nullptr,
                                             stats ? g.dL_dmeans2D + 2 * b : nullptr,
                                             stats ? max_radii2D + b : nullptr, stats ? xyz_grad_accum + b : nullptr,
                                             stats ? denom + b : nullptr, 0, skip_guard, stream));
    }
  }
  if (st->debug & GS_DEBUG_CHECKS) {
    GS_CUDA_CHECK(c, cudaStreamSynchronize(stream));
    GS_CUDA_CHECK(c, cudaGetLastError());
    // the parameter gradients never reach HBM here: scan what feeds them, and the updated parameters
    int r = scan_nonfinite(c, (const float*)c->grad_geom.ptr, (size_t)N * 12,
                           "blend gradients (dL/dmean2D, conic, opacity, colour)", stream);
    if (r == GS_OK) r = scan_nonfinite(c, in->means3D, (size_t)N * 3, "means3D after the fused Adam step", stream);
    if (r == GS_OK) r = scan_nonfinite(c, in->scales, (size_t)N * 3, "scales after the fused Adam step", stream);
    if (r == GS_OK) r = scan_nonfinite(c, in->rotations, (size_t)N * 4, "rotations after the fused Adam step", stream);
    if (r == GS_OK) r = scan_nonfinite(c, in->opacities, (size_t)N, "opacities after the fused Adam step", stream);
    if (r != GS_OK) return r;
  }
  c->have_backward = true;
  return GS_OK;
}

// ---- densify / prune / opacity reset (densify.cu) ----
int gs_densify_plan(gs_ctx* c, int N, const float* xyz_grad_accum, const float* denom, const float* opacities,
                    const float* scales, float grad_threshold, float min_opacity, float extent, float percent_dense,
                    float max_screen_size, int counts[4], void* stream_) {
  if (!c || N < 0 || !counts || (N > 0 && (!xyz_grad_accum || !denom || !opacities || !scales))) {
    gs_set_error(c, "gs_densify_plan: null argument");
    return GS_EINVAL;
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  const int nb = (N + 255) / 256;
  GS_ENSURE(c, c->dz_codes, (size_t)N + 16);
  GS_ENSURE(c, c->dz_counts, ((size_t)nb * 4 + 8) * 4);
  uint32_t* totals = (uint32_t*)c->dz_counts.ptr + (size_t)nb * 4;
  GS_CUDA_CHECK(c, gs_launch_densify_plan(N, xyz_grad_accum, denom, opacities, scales, grad_threshold, min_opacity,
                                          extent, percent_dense, max_screen_size, (uint8_t*)c->dz_codes.ptr,
                                          (uint32_t*)c->dz_counts.ptr, totals, nullptr, stream));
  uint32_t h[4];
  GS_CUDA_CHECK(c, cudaMemcpyAsync(h, totals, 16, cudaMemcpyDeviceToHost, stream));
  GS_CUDA_CHECK(c, cudaStreamSynchronize(stream));
  const long long n_new = (long long)h[0] + h[1] + 2ll * h[2];
  if (n_new > 0x7fffffffll) {
    gs_set_error(c, "gs_densify_plan: %lld Gaussians after the step exceed the int32 index range", n_new);
    return GS_EINVAL;
  }
  counts[0] = (int)h[0];
  counts[1] = (int)h[1];
  counts[2] = (int)h[2];
  counts[3] = (int)((long long)h[3] - n_new);
  c->dz_N = N;
  return GS_OK;
}

int gs_densify_apply(gs_ctx* c, int N, const int widths[6], float* const src_params[6], float* const src_exp_avg[6],
                     float* const src_exp_avg_sq[6], float* const dst_params[6], float* const dst_exp_avg[6],
                     float* const dst_exp_avg_sq[6], unsigned long long seed, void* stream_) {
  if (!c || !widths || !src_params || !src_exp_avg || !src_exp_avg_sq || !dst_params || !dst_exp_avg || !dst_exp_avg_sq) {
    gs_set_error(c, "gs_densify_apply: null argument");
    return GS_EINVAL;
  }
  if (c->dz_N != N) {
    gs_set_error(c, "gs_densify_apply needs the gs_densify_plan of the same N to have run on this context");
    return GS_ESTATE;
  }
  if (widths[0] != 3 || widths[3] != 1 || widths[4] != 3 || widths[5] != 4) {
    gs_set_error(c, "gs_densify_apply: group widths must be xyz 3, f_dc, f_rest, opacity 1, scaling 3, rotation 4");
    return GS_EINVAL;
  }
  for (int k = 0; k < 6; ++k)
    if (widths[k] > 0 && N > 0 && (!src_params[k] || !src_exp_avg[k] || !src_exp_avg_sq[k] || !dst_params[k] ||
                                   !dst_exp_avg[k] || !dst_exp_avg_sq[k])) {
      gs_set_error(c, "gs_densify_apply: group %d has a null slab", k);
      return GS_EINVAL;
    }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  const int nb = (N + 255) / 256;
  const uint32_t* totals = (const uint32_t*)c->dz_counts.ptr + (size_t)nb * 4;
  GS_CUDA_CHECK(c, gs_launch_densify_apply(N, (const uint8_t*)c->dz_codes.ptr, (const uint32_t*)c->dz_counts.ptr, totals,
                                           widths, src_params, src_exp_avg, src_exp_avg_sq, dst_params, dst_exp_avg,
                                           dst_exp_avg_sq, seed, (cudaStream_t)stream_));
  c->dz_N = -1;
  c->have_forward = false;  // N changed: the forward state no longer matches the model
  return GS_OK;
}

int gs_reset_opacity(gs_ctx* c, int N, float* opacities, float* exp_avg, float* exp_avg_sq, void* stream_) {
  if (!c || N < 0 || (N > 0 && (!opacities || !exp_avg || !exp_avg_sq))) {
    gs_set_error(c, "gs_reset_opacity: null argument");
    return GS_EINVAL;
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, gs_launch_reset_opacity(N, opacities, exp_avg, exp_avg_sq, (cudaStream_t)stream_));
  return GS_OK;
}

int gs_debug_reuse_blend_gradients(gs_ctx* c, int on) {
  if (!c) return GS_EINVAL;
  c->reuse_blend_gradients = on != 0;
  return GS_OK;
}

int gs_adam_densify_step(gs_ctx* c, int N, float* const params[6], float* const grads[6], float* const exp_avg[6],
                         float* const exp_avg_sq[6], const int widths[6], const double lr[6], const long long step[6],
                         double beta1, double beta2, double eps, const int* radii, const float* dL_dmeans2D,
                         int* max_radii2D, float* xyz_grad_accum, float* denom, int zero_grads,
                         const float* skip_guard, void* stream_) {
  if (!c || N < 0 || !params || !grads || !exp_avg || !exp_avg_sq || !widths || !lr || !step) {
    gs_set_error(c, "gs_adam_densify_step: null argument");
    return GS_EINVAL;
  }
  for (int s = 0; s < 6; ++s) {
    if (widths[s] > 0 && params[s] && (!grads[s] || !exp_avg[s] || !exp_avg_sq[s])) {
      gs_set_error(c, "segment %d: grads/exp_avg/exp_avg_sq missing", s);
      return GS_EINVAL;
    }
    if (step[s] < 1) {
      gs_set_error(c, "segment %d: step must be >= 1 (libtorch increments before use)", s);
      return GS_EINVAL;
    }
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, gs_launch_adam(N, params, grads, exp_avg, exp_avg_sq, widths, lr, step, beta1, beta2, eps, radii,
                                  dL_dmeans2D, max_radii2D, xyz_grad_accum, denom, zero_grads, skip_guard,
                                  (cudaStream_t)stream_));
  return GS_OK;
}

int gs_adam_step_flat(gs_ctx* c, int nseg, float* const params[], float* const grads[], float* const exp_avg[],
                      float* const exp_avg_sq[], const unsigned long long counts[], const double lr[],
                      const long long step[], double beta1, double beta2, double eps, int zero_grads,
                      const float* skip_guard, void* stream_) {
  if (!c || nseg < 0 || nseg > 6 || (nseg > 0 && (!params || !grads || !exp_avg || !exp_avg_sq || !counts || !lr || !step))) {
    gs_set_error(c, "gs_adam_step_flat: bad argument (at most 6 segments)");
    return GS_EINVAL;
  }
  size_t cnt[6] = {0, 0, 0, 0, 0, 0};
  for (int s = 0; s < nseg; ++s) {
    cnt[s] = (size_t)counts[s];
    if (cnt[s] > 0 && (!params[s] || !grads[s] || !exp_avg[s] || !exp_avg_sq[s])) {
      gs_set_error(c, "segment %d: null pointer", s);
      return GS_EINVAL;
    }
    if (step[s] < 1) {
      gs_set_error(c, "segment %d: step must be >= 1 (libtorch increments before use)", s);
      return GS_EINVAL;
    }
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, gs_launch_adam_counts(0, nseg, params, grads, exp_avg, exp_avg_sq, cnt, lr, step, beta1, beta2, eps,
                                         nullptr, nullptr, nullptr, nullptr, nullptr, zero_grads, skip_guard,
                                         (cudaStream_t)stream_));
  return GS_OK;
}

int gs_set_gradient_routing(gs_ctx* c, int nranks, int rank, int rows_per_shard, float* const table[]) {
  if (!c) return GS_EINVAL;
  if (nranks == 0) {
    c->route_shards = 0;
    return GS_OK;
  }
  if (nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks || rows_per_shard <= 0 || rows_per_shard % 128 || !table) {
    gs_set_error(c, "gs_set_gradient_routing: 1..8 ranks, rows_per_shard a positive multiple of 128");
    return GS_EINVAL;
  }
  for (int i = 0; i < nranks * 6; ++i) {
    if (!table[i] || ((uintptr_t)table[i] & 15)) {
      gs_set_error(c, "gs_set_gradient_routing: slot %d null or not 16-byte aligned", i);
      return GS_EINVAL;
    }
    c->route_table[i] = table[i];
  }
  c->route_shards = nranks;
  c->route_rank = rank;
  c->route_rows = rows_per_shard;
  return GS_OK;
}

// Grid of an Adam call that runs beside a forward (GS_PEER_BESIDE_FORWARD).  The kernel is bound by NVLink, and the
// more stores it keeps queued the longer the loads of the kernels beside it wait; too few blocks and the phase
// itself is late.  The rows per rank go with 1 / nranks, so does the cap: 3 * SMs / nranks (8 GPUs: 55, measured
// 40-64 blocks 1.56 ms/step, 24: 1.74, 148: 1.63, 296: 1.71; 4 GPUs: 111, measured 64-128 blocks 1.49-1.50, 49:
// 1.58, 296: 1.52 -- profiles/r02k_split_phase.md).  GS_PEER_ADAM_GRID overrides.
static int gs_peer_side_grid(const gs_ctx* c, int nranks) {
  if (c->peer_side_grid > 0) return c->peer_side_grid;
  const int g = 3 * c->num_sms / (nranks > 0 ? nranks : 1);
  return g > 16 ? g : 16;
}

int gs_adam_step_peer(gs_ctx* c, int nranks, int rank, int N, int rows_per_shard, float* const params[],
                      const float* const slots[], float* const exp_avg[6], float* const exp_avg_sq[6],
                      const int widths[6], const double lr[6], const long long step[6], double beta1, double beta2,
                      double eps, const int* radii, const float* dL_dmeans2D, int* max_radii2D,
                      float* xyz_grad_accum, float* denom, const float* skip_guard, int flags, void* stream_) {
  if (!c || nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks || N < 0 || rows_per_shard <= 0 ||
      rows_per_shard % 128 || !params || !slots || !exp_avg || !exp_avg_sq || !widths || !lr || !step) {
    gs_set_error(c, "gs_adam_step_peer: bad argument (1..8 ranks, rows_per_shard a positive multiple of 128)");
    return GS_EINVAL;
  }
  if ((long long)rows_per_shard * nranks < N) {
    gs_set_error(c, "gs_adam_step_peer: %d shards of %d rows do not cover N = %d", nranks, rows_per_shard, N);
    return GS_EINVAL;
  }
  for (int k = 0; k < 6; ++k) {
    if (step[k] < 1) {
      gs_set_error(c, "segment %d: step must be >= 1 (libtorch increments before use)", k);
      return GS_EINVAL;
    }
    if (widths[k] <= 0) continue;
    bool ok = exp_avg[k] && exp_avg_sq[k] && !((uintptr_t)exp_avg[k] & 15) && !((uintptr_t)exp_avg_sq[k] & 15);
    for (int r = 0; r < nranks; ++r)
      ok = ok && params[r * 6 + k] && slots[r * 6 + k] && !((uintptr_t)params[r * 6 + k] & 15) &&
           !((uintptr_t)slots[r * 6 + k] & 15);
    if (!ok) {
      gs_set_error(c, "segment %d: null or unaligned pointer", k);
      return GS_EINVAL;
    }
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, gs_launch_adam_peer(nranks, rank, N, rows_per_shard, params, slots, exp_avg, exp_avg_sq, widths, lr,
                                       step, beta1, beta2, eps, radii, dL_dmeans2D, max_radii2D, xyz_grad_accum,
                                       denom, skip_guard, (flags & GS_PEER_PULL) ? 0 : 1,
                                       (flags & GS_PEER_BESIDE_FORWARD) ? gs_peer_side_grid(c, nranks) : 0,
                                       (flags & GS_PEER_SH_LOCAL) ? 0x6 : 0, (cudaStream_t)stream_));
  return GS_OK;
}

int gs_peer_copy(gs_ctx* c, void* dst, const void* src, unsigned long long bytes, void* stream_) {
  if (!c || (bytes && (!dst || !src))) {
    gs_set_error(c, "gs_peer_copy: bad argument");
    return GS_EINVAL;
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  if (bytes) GS_CUDA_CHECK(c, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
  return GS_OK;
}

int gs_peer_barrier(gs_ctx* c, int nranks, int rank, void* const flag_blocks[], unsigned int epoch, const float* guard_in,
                    float* guard_out, void* stream_) {
  if (!c || nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks || !flag_blocks) {
    gs_set_error(c, "gs_peer_barrier: bad argument (1..8 ranks)");
    return GS_EINVAL;
  }
  uint32_t* f[8] = {};
  for (int r = 0; r < nranks; ++r) {
    if (!flag_blocks[r] || ((uintptr_t)flag_blocks[r] & 3)) {
      gs_set_error(c, "gs_peer_barrier: flag block %d null or unaligned", r);
      return GS_EINVAL;
    }
    f[r] = (uint32_t*)flag_blocks[r];
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, gs_launch_peer_barrier(nranks, rank, f, epoch, guard_in, guard_out, (cudaStream_t)stream_));
  return GS_OK;
}

// ---- peer memory (one process per GPU, NVLink): plain cudaMalloc + CUDA IPC handles ----
int gs_peer_alloc(gs_ctx* c, unsigned long long bytes, void** ptr, unsigned char handle[64]) {
  if (!c || !ptr || !handle) {
    gs_set_error(c, "gs_peer_alloc: null argument");
    return GS_EINVAL;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  void* p = nullptr;
  GS_CUDA_CHECK(c, cudaMalloc(&p, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    gs_set_error(c, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
    return GS_ECUDA;
  }
  GS_CUDA_CHECK(c, cudaMemset(p, 0, (size_t)bytes));
  memcpy(handle, &h, 64);
  *ptr = p;
  return GS_OK;
}

int gs_peer_open(gs_ctx* c, const unsigned char handle[64], void** ptr) {
  if (!c || !ptr || !handle) {
    gs_set_error(c, "gs_peer_open: null argument");
    return GS_EINVAL;
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  GS_CUDA_CHECK(c, cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return GS_OK;
}

int gs_peer_close(gs_ctx* c, void* ptr) {
  if (!c) return GS_EINVAL;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, cudaIpcCloseMemHandle(ptr));
  return GS_OK;
}

int gs_peer_free(gs_ctx* c, void* ptr) {
  if (!c) return GS_EINVAL;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, cudaFree(ptr));
  return GS_OK;
}

int gs_set_duplicate_capacity(gs_ctx* c, unsigned long long pairs) {
  if (!c) return GS_EINVAL;
  if (pairs > 0xFFFFFFF0ull) {
    gs_set_error(c, "capacity of %llu pairs exceeds 32-bit indexing", pairs);
    return GS_EINVAL;
  }
  c->dup_cap_fixed = pairs;
  return GS_OK;
}

int gs_set_color_gate(gs_ctx* c, void* cuda_event) {
  if (!c) return GS_EINVAL;
  c->color_gate = (cudaEvent_t)cuda_event;
  return GS_OK;
}

int gs_set_overflow_mirror(gs_ctx* c, float* device_float) {
  if (!c) return GS_EINVAL;
  c->overflow_mirror = device_float;
  return GS_OK;
}

const float* gs_overflow_guard(gs_ctx* c) {
  if (!c) return nullptr;
  cudaSetDevice(c->device);
  if (ensure(c, c->counters, 64) != GS_OK) return nullptr;
  return (const float*)((uint32_t*)c->counters.ptr + 6);
}

int gs_get_forward_status(gs_ctx* c, unsigned long long* num_duplicates, int* overflow) {
  if (!c) return GS_EINVAL;
  if (!c->have_forward) {
    gs_set_error(c, "gs_get_forward_status: no forward has run on this context");
    return GS_ESTATE;
  }
  // the values were copied to pinned memory on the forward's stream: the caller has synchronised it
  const unsigned long long D = *(const unsigned long long*)c->d_host_pinned;
  if (num_duplicates) *num_duplicates = D;
  if (overflow) *overflow = c->N > 0 ? (int)c->d_host_pinned[2] : 0;
  return GS_OK;
}

int gs_loss(gs_ctx* c, int H, int W, const float* image, const float* gt, float lambda_dssim, float* loss_out,
            float* dL_dimage, void* stream_) {
  if (!c || H <= 0 || W <= 0 || !image || !gt || !loss_out) {
    gs_set_error(c, "gs_loss: bad argument");
    return GS_EINVAL;
  }
  if (!(lambda_dssim >= 0.0f && lambda_dssim <= 1.0f)) {
    gs_set_error(c, "gs_loss: lambda_dssim %g outside [0,1]", (double)lambda_dssim);
    return GS_EINVAL;
  }
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_ENSURE(c, c->loss_acc, 64);
  if (lambda_dssim != 0.0f) {
    GS_ENSURE(c, c->ssim_work, gs_ssim_workspace_bytes(H, W));
    GS_CUDA_CHECK(c, gs_launch_photometric_loss(H, W, image, gt, lambda_dssim, (float*)c->ssim_work.ptr,
                                                (double*)c->loss_acc.ptr, loss_out, dL_dimage,
                                                (cudaStream_t)stream_));
    return GS_OK;
  }
  GS_CUDA_CHECK(c, gs_launch_l1_loss((size_t)3 * H * W, image, gt, 1.0f, dL_dimage, (double*)c->loss_acc.ptr,
                                     loss_out, (cudaStream_t)stream_));
  return GS_OK;
}

// ---- parity harness ---------------------------------------------------------------------------
static long long dbg_size(gs_ctx* c, int what) {
  if (!c->have_forward) return GS_ESTATE;
  if (c->D_pending) {  // capacity mode: the forward did not wait for D
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    const unsigned long long D = *(const unsigned long long*)c->d_host_pinned;
    c->D = D < c->dup_cap_fixed ? D : c->dup_cap_fixed;
    c->D_pending = false;
  }
  const long long N = c->N, D = (long long)c->D, T = c->ntiles, P = (long long)c->H * c->W;
  switch (what) {
    case GS_DBG_MEANS2D: return N * 8;
    case GS_DBG_DEPTHS: return N * 4;
    case GS_DBG_CONIC_OPACITY: return N * 16;
    case GS_DBG_RGB: return N * 12;
    case GS_DBG_CLAMPED: return N;
    case GS_DBG_TILES_TOUCHED: return N * 4;
    case GS_DBG_DEPTH_ORDER: return N * 4;
    case GS_DBG_KEYS_SORTED: return D * 8;
    case GS_DBG_POINT_LIST: return D * 4;
    case GS_DBG_RANGES: return T * 8;
    case GS_DBG_FINAL_T: return P * 4;
    case GS_DBG_N_CONTRIB: return P * 4;
    case GS_DBG_DL_DMEAN2D: return c->have_backward ? N * 8 : GS_ESTATE;
    case GS_DBG_DL_DCONIC: return c->have_backward ? N * 12 : GS_ESTATE;
    case GS_DBG_DL_DOPACITY: return c->have_backward ? N * 4 : GS_ESTATE;
    case GS_DBG_DL_DCOLOR: return c->have_backward ? N * 12 : GS_ESTATE;
    case GS_DBG_NUM_RENDERED: return 8;
    default: return GS_EINVAL;
  }
}

long long gs_debug_size(gs_ctx* c, int what) {
  if (!c) return GS_EINVAL;
  return dbg_size(c, what);
}

int gs_debug_get(gs_ctx* c, int what, void* host_dst, size_t bytes) {
  if (!c || (!host_dst && bytes)) return GS_EINVAL;
  const long long need = dbg_size(c, what);
  if (need < 0) {
    gs_set_error(c, "gs_debug_get(%d): not available in the current state", what);
    return (int)need;
  }
  if ((size_t)need != bytes) {
    gs_set_error(c, "gs_debug_get(%d): expected %lld bytes, got %zu", what, need, bytes);
    return GS_EINVAL;
  }
  if (bytes == 0) return GS_OK;
  GS_CUDA_CHECK(c, cudaSetDevice(c->device));
  GS_CUDA_CHECK(c, cudaDeviceSynchronize());
  const size_t N = (size_t)c->N;
  const void* src = nullptr;
  switch (what) {
    case GS_DBG_DEPTHS: src = c->depths.ptr; break;
    case GS_DBG_CLAMPED: src = c->clamped.ptr; break;
    case GS_DBG_TILES_TOUCHED: src = c->tiles.ptr; break;
    case GS_DBG_DEPTH_ORDER: src = c->sort_v1.ptr; break;
    case GS_DBG_KEYS_SORTED: {
      // the 64-bit keys (tile << 32 | depth bits) are implied by the two-level sort: rebuild them on the host
      const size_t D = (size_t)c->D;
      std::vector<uint32_t> ids(D), dep(N);
      std::vector<uint8_t> tk(D * (size_t)c->key_bytes);
      GS_CUDA_CHECK(c, cudaMemcpy(ids.data(), c->point_list, D * 4, cudaMemcpyDeviceToHost));
      GS_CUDA_CHECK(c, cudaMemcpy(tk.data(), c->sorted_tiles, tk.size(), cudaMemcpyDeviceToHost));
      GS_CUDA_CHECK(c, cudaMemcpy(dep.data(), c->depths.ptr, N * 4, cudaMemcpyDeviceToHost));
      uint64_t* o = (uint64_t*)host_dst;
      for (size_t i = 0; i < D; ++i) {
        const uint32_t tile = c->key_bytes == 2 ? ((const uint16_t*)tk.data())[i] : ((const uint32_t*)tk.data())[i];
        o[i] = ((uint64_t)tile << 32) | dep[ids[i]];
      }
      return GS_OK;
    }
    case GS_DBG_POINT_LIST: src = c->point_list; break;
    case GS_DBG_RANGES: src = c->ranges.ptr; break;
    case GS_DBG_FINAL_T: {  // .x of the per-pixel final state {T, C.r, C.g, C.b}
      const size_t P = (size_t)c->H * c->W;
      std::vector<float> fs(P * 4);
      GS_CUDA_CHECK(c, cudaMemcpy(fs.data(), c->final_state.ptr, P * 16, cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < P; ++i) ((float*)host_dst)[i] = fs[4 * i];
      return GS_OK;
    }
    case GS_DBG_N_CONTRIB: src = c->n_contrib.ptr; break;
    case GS_DBG_NUM_RENDERED: {
      const uint64_t D = c->D;
      memcpy(host_dst, &D, 8);
      return GS_OK;
    }
    default: break;
  }
  if (src) {
    GS_CUDA_CHECK(c, cudaMemcpy(host_dst, src, bytes, cudaMemcpyDeviceToHost));
    return GS_OK;
  }
  // records that are packed on the device: copy the AoS and unpack on the host
  const bool is_grad = what >= GS_DBG_DL_DMEAN2D && what <= GS_DBG_DL_DCOLOR;
  std::vector<float> rec(N * 12);
  GS_CUDA_CHECK(c, cudaMemcpy(rec.data(), is_grad ? c->grad_geom.ptr : c->geom.ptr, N * 48, cudaMemcpyDeviceToHost));
  float* o = (float*)host_dst;
  for (size_t i = 0; i < N; ++i) {
    const float* r = rec.data() + 12 * i;
    switch (what) {
      case GS_DBG_MEANS2D: o[2 * i] = r[0]; o[2 * i + 1] = r[1]; break;
      case GS_DBG_CONIC_OPACITY:
        o[4 * i] = 2.0f * r[2]; o[4 * i + 1] = r[3]; o[4 * i + 2] = 2.0f * r[4]; o[4 * i + 3] = r[5];
        break;
      case GS_DBG_RGB: o[3 * i] = r[6]; o[3 * i + 1] = r[7]; o[3 * i + 2] = r[8]; break;
      case GS_DBG_DL_DMEAN2D: o[2 * i] = r[0]; o[2 * i + 1] = r[1]; break;
      case GS_DBG_DL_DCONIC: o[3 * i] = r[2]; o[3 * i + 1] = r[3]; o[3 * i + 2] = r[4]; break;
      case GS_DBG_DL_DOPACITY: o[i] = r[5]; break;
      case GS_DBG_DL_DCOLOR: o[3 * i] = r[6]; o[3 * i + 1] = r[7]; o[3 * i + 2] = r[8]; break;
      default: return GS_EINVAL;
    }
  }
  return GS_OK;
}

}  // extern "C"
