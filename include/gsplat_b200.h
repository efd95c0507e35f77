/*
 * gsplat_b200.h — C ABI of libgsplat_b200.so, the B200-native (sm_100a) differentiable
 * Gaussian-splat rasterizer that fills the render/rasterize call the reference trainer
 * (seiya-kumada/3d_gaussian_splatting) is missing at src/utils/train_utils.cpp:138-145.
 *
 * The reference has NO FFI / plugin interface for this path: the boundary is
 * the call surface BASELINE.json's north_star spells out, and every argument is backed by a
 * reference object, cited next to each declaration below.  Plain pointers and sizes only;
 * no C++/torch types; no exceptions cross this boundary.  All device pointers are owned by
 * the caller (in the reference they would be the torch::Tensor storages of
 * GaussianModel::CoreParams and Camera); the context owns every intermediate workspace.
 *
 * Threading: one context per (host thread, device); calls on a context are not re-entrant.
 * All GPU work is enqueued on the `stream` argument (a cudaStream_t passed as void*).
 * There is NO CPU fallback: every compute entry point returns GS_ECUDA when no sm_100 device
 * is usable.
 */
#ifndef GSPLAT_B200_H
#define GSPLAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_API __attribute__((visibility("default")))

/* Error codes.  The reference signals errors with C++ exceptions (std::runtime_error at
 * src/arguments/params.cpp:148, src/scene/colmap_loader.cpp:125) and assert(false)
 * (src/scene/scene.cpp:51); the ABI maps them to integers + gs_last_error(). */
enum {
  GS_OK = 0,
  GS_EINVAL = -1,    /* bad argument (null pointer, N < 0, sh degree out of range ...) */
  GS_ENOMEM = -2,    /* device allocation failed */
  GS_ECUDA = -3,     /* CUDA runtime / launch error, or no usable device */
  GS_EOVERFLOW = -4, /* duplicate (Gaussian,tile) list exceeds the 32-bit index space */
  GS_ESTATE = -5,    /* backward without a matching forward, etc. */
  GS_ENUMERIC = -6   /* debug: a gradient is NaN or infinite (the role of --detect_anomaly, src/train.cpp:39) */
};

typedef struct gs_ctx gs_ctx;

/*
 * Raster settings = the arguments of the missing render(viewpoint_cam, gaussians, pipe, bg):
 *   H, W            Camera::image_height_/image_width_      src/scene/camera.cpp:57-58
 *   tanfovx/y       tan(FoVx_/2), tan(FoVy_/2)              src/scene/camera.cpp:38-39
 *   bg              background                              src/utils/train_utils.cpp:115-117
 *   scale_modifier  get_covariance(scaling_modifier)        src/scene/gaussian_model.h:90
 *   viewmatrix      world_view_transform_ (flat, fp64->fp32) src/scene/camera.cpp:66-68
 *   projmatrix      full_proj_transform_  (flat, fp64->fp32) src/scene/camera.cpp:70
 *   campos          camera_center_                           src/scene/camera.cpp:71
 *   sh_degree       CoreParams::active_sh_degree_ ; max_sh_degree = max_sh_degree_
 *                                                            src/scene/gaussian_model.h:11,43
 *   debug           PipelineParams::debug_                   src/arguments/params.h:97
 * Because the reference stores the transposed matrices, the flat buffers read as
 * column-major V and P*V: element [4*c + r] is M[r][c].
 */
typedef struct {
  int H, W;
  float tanfovx, tanfovy;
  float bg[3];
  float scale_modifier;
  float viewmatrix[16];
  float projmatrix[16];
  float campos[3];
  int sh_degree, max_sh_degree;
  int prefiltered, debug; /* debug: bit flags, below */
} gs_settings;
/* gs_settings.debug bits.  GS_DEBUG_CHECKS = PipelineParams::debug_ (src/arguments/params.h:97): every compute call
 * synchronises and reports CUDA errors at once, and the backward scans its gradients for NaN / inf (GS_ENUMERIC: the
 * role of --detect_anomaly, src/train.cpp:39).  GS_PUBLISHED_CLAMP_GRAD: differentiate the 1.3 * tanfov guard of the
 * EWA projection the way the widely used implementation does (factor 2 in dJ/dt.z for a clamped component too)
 * instead of the true derivative of the clamped forward; gradients differ only for Gaussians outside the guard. */
#define GS_DEBUG_CHECKS 1
#define GS_PUBLISHED_CLAMP_GRAD 2

/*
 * Per-Gaussian inputs (device pointers) = the GaussianModel getters
 * (src/scene/gaussian_model.cpp:270-298): get_xyz (N,3), get_scaling (N,3), get_rotation (N,4)
 * order (w,x,y,z), get_opacity (N,1), get_features (N,K,3).  SH is addressed as two strided
 * views so that both get_features()' concatenated tensor (sh_dc = shs, sh_rest = shs + 3,
 * both strides 3*K) and the separate features_dc_/features_rest_ parameters
 * (strides 3 and 3*(K-1); saves the torch::cat copy at gaussian_model.cpp:289-292) work.
 * colors_precomp / cov3D_precomp are the PipelineParams::convert_SHs_python_ /
 * compute_cov3D_python_ branches (src/arguments/params.h:95-96); cov3D layout is
 * strip_symmetric's [xx,xy,xz,yy,yz,zz] (src/utils/general_utils.cpp:49-62).
 * raw_params != 0: scales/rotations/opacities point at the RAW parameters scaling_ (log),
 * rotation_ (un-normalised) and opacity_ (logit) of CoreParams (gaussian_model.h:13-17) and
 * the activations exp / normalise / sigmoid (gaussian_model.cpp:54-59) are fused into the
 * kernels, forward and backward (gradients are then w.r.t. the raw parameters).
 */
typedef struct {
  const float *means3D;
  const float *scales;
  const float *rotations;
  const float *opacities;
  const float *sh_dc;
  const float *sh_rest;
  int sh_dc_stride;   /* floats between consecutive Gaussians in sh_dc   */
  int sh_rest_stride; /* floats between consecutive Gaussians in sh_rest */
  const float *colors_precomp; /* nullable, (N,3) */
  const float *cov3D_precomp;  /* nullable, (N,6) */
  int raw_params;
} gs_gaussians;

/* Gradient sinks (device pointers, same layouts as the inputs; any may be NULL to skip).
 * dL_dmeans2D is (N,2): the screen-space gradient whose norm feeds xyz_gradient_accum_
 * (src/scene/gaussian_model.h:19).  accumulate != 0 adds into the buffers instead of
 * overwriting them (gradient accumulation over several views) -- except dL_dmeans2D, which is a
 * per-VIEW quantity (the statistics take its norm view by view) and is always overwritten. */
typedef struct {
  float *dL_dmeans3D;
  float *dL_dscales;
  float *dL_drotations;
  float *dL_dopacities;
  float *dL_dsh_dc;
  float *dL_dsh_rest;
  int sh_dc_stride;
  int sh_rest_stride;
  float *dL_dmeans2D;
  float *dL_dcolors; /* nullable; filled when colors_precomp was given */
  float *dL_dcov3D;  /* nullable; filled when cov3D_precomp was given */
  int accumulate;
} gs_gradients;

/* ---- lifecycle ------------------------------------------------------------------------- */
GS_API const char *gs_version(void);
GS_API int gs_create(gs_ctx **out, int device);
GS_API void gs_destroy(gs_ctx *ctx);
GS_API const char *gs_last_error(const gs_ctx *ctx); /* ctx may be NULL: last global error */

/* ---- the render call (forward) -----------------------------------------------------------
 * out_color (3,H,W) CHW like Camera::original_image_ (src/scene/camera.cpp:56-58);
 * radii (N) int32, visibility == radii > 0 (consumer: max_radii2D_, gaussian_model.h:18).
 * The forward state (geometry, sorted lists, final_T, n_contrib) stays in the context until
 * the next gs_forward; gs_backward consumes it. */
GS_API int gs_forward(gs_ctx *ctx, const gs_settings *settings, int N, const gs_gaussians *in,
                      float *out_color, int *radii, void *stream);

/* ---- backward: gradients to every Gaussian parameter ------------------------------------ */
GS_API int gs_backward(gs_ctx *ctx, const gs_settings *settings, int N, const gs_gaussians *in,
                       const float *dL_dout_color /* (3,H,W) */, const gs_gradients *grads, void *stream);

/* ---- fused Adam + densification statistics ------------------------------------------------
 * The reference builds six torch::optim::Adam with only `lr` set
 * (src/scene/gaussian_model.cpp:323-345: xyz, f_dc, f_rest, opacity, scaling, rotation) =>
 * libtorch defaults beta=(0.9,0.999), eps=1e-8, no weight decay, no amsgrad, six independent
 * step counters (hyper-parameters are double, as in torch::optim::AdamOptions).  One launch updates all six segments (segment s has N*widths[s] floats).
 * Statistics on Gaussians with radii > 0 (fields at src/scene/gaussian_model.h:18-20):
 *   max_radii2D = max(max_radii2D, radii); xyz_grad_accum += |dL_dmeans2D.xy|; denom += 1.
 * radii / dL_dmeans2D / the three stat buffers may be NULL to skip the statistics.
 * zero_grads != 0 re-zeroes the gradient buffers in the same pass.
 * skip_guard (nullable): device float; when it is non-zero at execution time the whole call is a
 * no-op (parameters, moments, statistics and gradients untouched).  Pass gs_overflow_guard(ctx) --
 * or, with several GPUs, the all-reduced copy of it -- so that a forward that overflowed its fixed
 * pair capacity (gs_set_duplicate_capacity) cannot corrupt the model; the caller then grows the
 * capacity and repeats the iteration. */
GS_API int gs_adam_densify_step(gs_ctx *ctx, int N, float *const params[6], float *const grads[6],
                                float *const exp_avg[6], float *const exp_avg_sq[6], const int widths[6],
                                const double lr[6], const long long step[6], double beta1, double beta2,
                                double eps, const int *radii, const float *dL_dmeans2D, int *max_radii2D,
                                float *xyz_grad_accum, float *denom, int zero_grads, const float *skip_guard,
                                void *stream);

/* ---- backward + optimiser in one call (the single-GPU training step) -------------------------
 * gs_backward followed by gs_adam_densify_step on the same model, with the per-Gaussian half of the backward
 * and the Adam update FUSED into one kernel: a block turns its 128 gradient rows into parameter updates while
 * they are still in shared memory, so the six gradient tensors are never written to or read from HBM
 * (1464 B instead of 2172 B per Gaussian at SH degree 3, one launch less).  Bit-identical to the two calls.
 * `in` must be the trainable layout (raw or activated; sh_dc (N,1,3) + sh_rest (N,K-1,3) dense, no precomputed
 * colours / covariances); its six arrays ARE the parameters and are updated in place, in the group order of
 * GaussianModel::setup (src/scene/gaussian_model.cpp:323-345).  dL_dmeans2D: optional (N,2) sink.  Statistics
 * and skip_guard as in gs_adam_densify_step. */
GS_API int gs_backward_adam_step(gs_ctx *ctx, const gs_settings *settings, int N, const gs_gaussians *in,
                                 const float *dL_dout_color, float *dL_dmeans2D, float *const exp_avg[6],
                                 float *const exp_avg_sq[6], const double lr[6], const long long step[6],
                                 double beta1, double beta2, double eps, int *max_radii2D, float *xyz_grad_accum,
                                 float *denom, const float *skip_guard, void *stream);

/* ---- densify / prune / opacity reset with optimiser-state surgery ------------------------------
 * The consumers of the statistics gs_adam_densify_step accumulates (max_radii2D_, xyz_gradient_accum_, denom_,
 * src/scene/gaussian_model.h:18-20; percent_dense_ :44) under the schedule the reference parses
 * (src/arguments/params.h:61-67: densification_interval_, opacity_reset_interval_, densify_from_iter_,
 * densify_until_iter_, densify_grad_threshold_).  The reference declares the state, not the step; the algorithm
 * is the published 3DGS densify_and_prune (clone small high-gradient Gaussians, split large ones into two
 * samples shrunk by 1.6, prune split parents / faint / over-large ones), specified in oracle/orc_densify.c.
 *
 * Two calls, because the caller owns the slabs and N changes:
 *   gs_densify_plan   classifies every Gaussian (raw opacity logits and log-scales in) and returns
 *                     counts = {kept originals, kept clones, kept split parents, pruned};
 *                     the new N is counts[0] + counts[1] + 2*counts[2].  Synchronises the stream (12 bytes back).
 *   gs_densify_apply  moves parameters AND Adam moments from the old slabs to caller-allocated new ones of the
 *                     new N, in the published order [originals | clones | first children | second children];
 *                     new Gaussians get zero moments.  Group order and widths as gs_adam_densify_step.
 *                     seed: the counter-based generator of the split samples (oracle/orc_densify.c).
 * The three statistics buffers are all-zero after the published step: the caller re-allocates them zeroed.
 * gs_reset_opacity: opacity <- logit(min(sigmoid(opacity), 0.01)), fresh Adam state for the opacity group. */
GS_API int gs_densify_plan(gs_ctx *ctx, int N, const float *xyz_grad_accum, const float *denom,
                           const float *opacities, const float *scales, float grad_threshold, float min_opacity,
                           float extent, float percent_dense, float max_screen_size, int counts[4], void *stream);
GS_API int gs_densify_apply(gs_ctx *ctx, int N, const int widths[6], float *const src_params[6],
                            float *const src_exp_avg[6], float *const src_exp_avg_sq[6], float *const dst_params[6],
                            float *const dst_exp_avg[6], float *const dst_exp_avg_sq[6], unsigned long long seed,
                            void *stream);
GS_API int gs_reset_opacity(gs_ctx *ctx, int N, float *opacities, float *exp_avg, float *exp_avg_sq, void *stream);

/* Test hook: with on != 0, gs_backward and gs_backward_adam_step skip the blend backward and consume the
 * per-Gaussian gradient records the previous backward left in the context -- the blend backward sums with fp32 atomics, so this is what
 * makes "fused == separate" checkable bit for bit. */
GS_API int gs_debug_reuse_blend_gradients(gs_ctx *ctx, int on);

/* The same Adam update over up to six arbitrary contiguous float ranges (counts[s] floats each): the form a
 * rank uses when the optimiser state is SHARDED across GPUs -- gradients reduce-scattered, each rank updating
 * the slice of the flat parameter slab it owns (a slice intersects at most the six parameter groups), the
 * updated parameters all-gathered (3d_gaussian_splatting_b200/parallel.py ShardedAdam).  No statistics. */
GS_API int gs_adam_step_flat(gs_ctx *ctx, int nseg, float *const params[], float *const grads[],
                             float *const exp_avg[], float *const exp_avg_sq[], const unsigned long long counts[],
                             const double lr[], const long long step[], double beta1, double beta2, double eps,
                             int zero_grads, const float *skip_guard, void *stream);

/* ---- peer memory for the multi-GPU step (one process per GPU on one NVLink / NVSwitch node) ----
 * gs_peer_alloc: device buffer (zero-filled) that other processes may map; handle = the 64-byte CUDA IPC handle
 * to hand to the peers (any transport -- the Python host uses torch.distributed).  gs_peer_open maps a peer's
 * buffer into this process (peer access enabled on demand); kernels then load / store through the returned
 * pointer over NVLink.  The reference has no multi-device path: B200-native extension. */
GS_API int gs_peer_alloc(gs_ctx *ctx, unsigned long long bytes, void **ptr, unsigned char handle[64]);
GS_API int gs_peer_open(gs_ctx *ctx, const unsigned char handle[64], void **ptr);
GS_API int gs_peer_close(gs_ctx *ctx, void *ptr);
GS_API int gs_peer_free(gs_ctx *ctx, void *ptr);

/* ---- the multi-GPU step over peer memory: reduce-scatter, optimiser and all-gather fused into K4b and K5 ----
 * Rank r of nranks (<= 8) OWNS the Gaussians [r*R, (r+1)*R), R = rows_per_shard (a multiple of 128), and holds
 * the Adam moments of those rows only.  Every rank keeps a full parameter replica and a slot buffer with one
 * slot per source rank; both are peer-mapped (gs_peer_alloc / gs_peer_open).
 *
 * gs_set_gradient_routing: from now on gs_backward stores the six parameter gradients of rows [q*R, (q+1)*R)
 * through table[q*6 + k] -- this rank's slot inside rank q's buffer (group k in the order xyz, f_dc, f_rest,
 * opacity, scaling, rotation; (R, widths[k]) floats, 16-byte aligned) -- instead of the gs_gradients sinks
 * (dL_dmeans2D stays local).  The stores cross NVLink from inside the backward kernel; nothing is staged.
 * nranks == 0 switches routing off.
 *
 * gs_adam_step_peer: sums the nranks slots of the owned rows in rank order, applies the Adam step of
 * gs_adam_densify_step to them and stores the new parameters into the replica of every rank: params[r*6 + k] is
 * rank r's (N, widths[k]) array (r == rank: local).  slots[s*6 + k]: the local slot of source rank s.
 * exp_avg / exp_avg_sq: (R, widths[k]) shard-local moments.  Statistics as in gs_adam_densify_step (per rank).
 * flags: 0 -- the PUSH layout above (remote f_rest slots are written sparsely by the routed backward and re-zeroed
 * by this call).  GS_PEER_PULL -- the layout for steps of several views per rank: no routing; every rank accumulates
 * all its views into one LOCAL (nranks*R, widths[k]) gradient slab, and slots[s*6 + k] points at rows [rank*R, ..)
 * of rank s's slab (peer-mapped for s != rank), which this call reads over NVLink: ONE exchange per step.  Nothing
 * is re-zeroed (the next step's first backward overwrites).
 *
 * Ordering is the caller's: every rank's gs_backward must have completed before any rank's gs_adam_step_peer
 * runs, and every gs_adam_step_peer before the next gs_forward -- two stream-ordered barriers per step (the
 * Python host uses a one-float NCCL all-reduce for each: the overflow guard and the loss). */
GS_API int gs_set_gradient_routing(gs_ctx *ctx, int nranks, int rank, int rows_per_shard, float *const table[]);
GS_API int gs_adam_step_peer(gs_ctx *ctx, int nranks, int rank, int N, int rows_per_shard, float *const params[],
                             const float *const slots[], float *const exp_avg[6], float *const exp_avg_sq[6],
                             const int widths[6], const double lr[6], const long long step[6], double beta1,
                             double beta2, double eps, const int *radii, const float *dL_dmeans2D,
                             int *max_radii2D, float *xyz_grad_accum, float *denom, const float *skip_guard,
                             int flags, void *stream);
#define GS_PEER_PULL 1
/* GS_PEER_BESIDE_FORWARD: the call runs on a second stream beside the next gs_forward (the SH groups of the
 * split-phase step, see gs_set_color_gate): it is bound by NVLink, not by the SMs, so its grid is capped at
 * 3 x SM count / nranks (grid-stride loop; GS_PEER_ADAM_GRID overrides).  Measured on 8 GPUs: the more remote stores the
 * kernel keeps queued, the longer the loads of K1 and the sorts beside it wait (40-64 blocks: 1.56 ms/step, 296:
 * 1.71, none of the split: 1.71). */
#define GS_PEER_BESIDE_FORWARD 2
/* GS_PEER_SH_LOCAL: the two SH groups (f_dc, f_rest) are updated in the rank's OWN replica only; the caller
 * distributes the owned rows to the other replicas with gs_peer_copy -- copy engines over NVLink, no SM involved --
 * on a second stream, beside the next gs_forward (see gs_set_color_gate).  The copy engines leave the kernels
 * beside them alone, but their all-gather is slower than the kernel's stores (8 GPUs: 336 GB/s per rank against
 * 560 GB/s; 1.83 against 1.56 ms/step, profiles/r02k_split_phase.md), so this is not the default. */
#define GS_PEER_SH_LOCAL 4
/* One contiguous device-to-device copy (cudaMemcpyAsync: a copy engine; dst or src may be peer-mapped). */
GS_API int gs_peer_copy(gs_ctx *ctx, void *dst, const void *src, unsigned long long bytes, void *stream);
/* The two barriers of the peer-memory step as ONE small kernel each instead of a collective: every rank owns a
 * 128-byte flag block inside its peer-mapped buffer (zeroed by gs_peer_alloc); flag_blocks[r] is rank r's block as
 * mapped here.  The call stores (guard_in, epoch) into this rank's slot of every block and waits until every slot of
 * its own block has reached `epoch` -- so it returns (in stream order) only after every rank's preceding kernels on
 * the same barrier have completed and their peer stores are visible.  epoch: a number that grows by one per barrier
 * (the same sequence on every rank).  guard_out receives the maximum of the ranks' guard_in (the overflow guard of
 * the step; both may be NULL).  A peer that does not arrive within ~2 s is counted in word 16 of the block and the
 * kernel returns instead of hanging. */
GS_API int gs_peer_barrier(gs_ctx *ctx, int nranks, int rank, void *const flag_blocks[], unsigned int epoch,
                           const float *guard_in, float *guard_out, void *stream);

/* Colour gate (multi-GPU step; no counterpart in the single-device reference).  With a non-NULL cudaEvent_t set,
 * gs_forward runs K1 WITHOUT the SH -> RGB evaluation (it reads 44 of a Gaussian's 236 B), bins and sorts, then makes
 * its stream wait for the event and evaluates colours + clamp masks (bit-identical to the undeferred K1) just ahead
 * of the forward blend.  The caller records the event -- BEFORE calling gs_forward, on whatever stream -- once the SH
 * coefficients of its replica are complete, e.g. after a gs_adam_step_peer over the two SH groups + gs_peer_barrier
 * on a second stream: the all-gather of 81 % of the parameter bytes then overlaps K1 and the binning stages of the
 * next iteration (parallel.PeerShardedAdam).  xyz / scaling / rotation / opacity must be complete when gs_forward is
 * called.  NULL restores the one-kernel K1. */
GS_API int gs_set_color_gate(gs_ctx *ctx, void *cuda_event);
#define GS_PEER_FLAG_BYTES 128

/* ---- running without the host round-trip ----------------------------------------------------
 * By default gs_forward synchronises the stream once, to size the (tile, Gaussian) pair lists from
 * the data-dependent pair count D.  gs_set_duplicate_capacity(ctx, pairs > 0) fixes the capacity
 * instead: gs_forward then enqueues everything without waiting (the kernels read D from device
 * memory).  If D exceeds the capacity the lists are truncated, the forward's outputs are invalid,
 * and the overflow is signalled (a) on the device through the float at gs_overflow_guard(ctx) (and
 * the optional caller-owned mirror, e.g. one element of an all-reduced gradient slab) which
 * gs_adam_densify_step honours, and (b) on the host through gs_get_forward_status once the caller
 * has synchronised the stream.  pairs == 0 restores the synchronous behaviour. */
GS_API int gs_set_duplicate_capacity(gs_ctx *ctx, unsigned long long pairs);
GS_API int gs_set_overflow_mirror(gs_ctx *ctx, float *device_float);
GS_API const float *gs_overflow_guard(gs_ctx *ctx);
GS_API int gs_get_forward_status(gs_ctx *ctx, unsigned long long *num_duplicates, int *overflow);

/* ---- photometric loss ----------------------------------------------------------------------
 * loss = (1 - lambda_dssim) * L1 + lambda_dssim * (1 - SSIM)   (lambda_dssim_ = 0.2 at
 * src/arguments/params.h:62; the loop body that would call it is missing at
 * src/utils/train_utils.cpp:138-145).  lambda_dssim == 0 => pure L1.
 * image, gt, dL_dimage are (3,H,W) device buffers; loss_out is ONE device float. */
GS_API int gs_loss(gs_ctx *ctx, int H, int W, const float *image, const float *gt, float lambda_dssim,
                   float *loss_out, float *dL_dimage, void *stream);

/* gs_loss_static: gs_loss for callers that may not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and result as gs_loss minus the context (`device` = CUDA device ordinal of the buffers); the SSIM
 * workspace is a file-scope device array inside the library (9 * GS_LOSS_STATIC_MAX_PIXELS floats + block partials
 * = 9.25 MiB, held by every GPU that loads the library until the process exits), so the call makes no cudaMalloc /
 * cudaFree and needs no gs_ctx.  H * W > GS_LOSS_STATIC_MAX_PIXELS (or more than 10 922 tiles of 32x32) returns
 * GS_EINVAL with a message in gs_last_error(NULL).  lambda_dssim = 0 runs the same kernels (gs_loss switches to a
 * pure-L1 kernel there; equal to rounding).  All calls on a device share the array: calls that overlap on the GPU
 * (different streams or host threads) race; nothing in it needs clearing between calls.  The array's device address
 * is resolved once per device on the first call there (cudaGetSymbolAddress): make that call outside a capture. */
#define GS_LOSS_STATIC_MAX_PIXELS 262144 /* e.g. 512 x 512 */
GS_API int gs_loss_static(int device, int H, int W, const float *image, const float *gt, float lambda_dssim,
                          float *loss_out, float *dL_dimage, void *stream);

/* ---- parity harness: read intermediates of the last gs_forward / gs_backward --------------
 * Copies (synchronously) the named workspace array to host memory.  `bytes` must equal the
 * array's size; gs_debug_size returns it (or a negative error). */
enum {
  GS_DBG_MEANS2D = 0,       /* (N,2) f32 */
  GS_DBG_DEPTHS = 1,        /* (N) f32 */
  GS_DBG_CONIC_OPACITY = 2, /* (N,4) f32 */
  GS_DBG_RGB = 3,           /* (N,3) f32 */
  GS_DBG_CLAMPED = 4,       /* (N) u8, bit c set when channel c was clamped at 0 */
  GS_DBG_TILES_TOUCHED = 5, /* (N) u32 */
  GS_DBG_DEPTH_ORDER = 6,   /* (N) u32 Gaussian ids in (depth bits, id) order: the first level of the sort.  Keys are written
                               ahead of K1, so only the relative order of VISIBLE Gaussians (radii > 0) is specified */
  GS_DBG_KEYS_SORTED = 7,   /* (D) u64 tile << 32 | depth bits, rebuilt from the sorted (tile, id) pairs */
  GS_DBG_POINT_LIST = 8,    /* (D) u32 */
  GS_DBG_RANGES = 9,        /* (T,2) u32 */
  GS_DBG_FINAL_T = 10,      /* (H,W) f32 */
  GS_DBG_N_CONTRIB = 11,    /* (H,W) u32 */
  GS_DBG_DL_DMEAN2D = 12,   /* (N,2) f32, after gs_backward */
  GS_DBG_DL_DCONIC = 13,    /* (N,3) f32 d/d(A,B,C), after gs_backward */
  GS_DBG_DL_DOPACITY = 14,  /* (N) f32 w.r.t. the ACTIVATED opacity, after gs_backward */
  GS_DBG_DL_DCOLOR = 15,    /* (N,3) f32, after gs_backward */
  GS_DBG_NUM_RENDERED = 16  /* 1 x u64: D */
};
GS_API long long gs_debug_size(gs_ctx *ctx, int what);
GS_API int gs_debug_get(gs_ctx *ctx, int what, void *host_dst, size_t bytes);

/* Per-stage device time of the last forward/backward in milliseconds (CUDA events recorded
 * only when gs_set_profiling(ctx,1)); order: preprocess, depth sort, scan+duplicate, tile sort,
 * ranges, blend, blend_bwd, preprocess_bwd, deferred colour (gs_set_color_gate; 0 otherwise).  The role of the CUDA-event timers the reference says it lacks
 * (src/utils/train_utils.cpp:119-121). */
#define GS_NUM_STAGES 9
GS_API int gs_set_profiling(gs_ctx *ctx, int enabled);
GS_API int gs_get_stage_ms(gs_ctx *ctx, float ms[GS_NUM_STAGES]);

/* ---- the training loop, native ------------------------------------------------------------------
 * The loop the reference leaves as a stub (src/utils/train_utils.cpp:97-147, body commented out at :138-144),
 * over the entry points above: no Python, no libtorch.  The trainer OWNS the model (parameters, Adam moments,
 * statistics) on `device`; gs_trainer_run is a depth-2 software pipeline (host one iteration behind the device,
 * ground truth host->device and loss device->host every iteration), with the reference's schedules
 * (src/arguments/params.h:52-68) and, when densify_extent > 0, densification / opacity reset.
 * gs_train_config mirrors OptimizationParams (same names without the trailing underscore). */
typedef struct gs_train_config {
  double position_lr_init, position_lr_final, position_lr_delay_mult;
  int position_lr_max_steps;
  double feature_lr, opacity_lr, scaling_lr, rotation_lr;
  double percent_dense, lambda_dssim;
  int densification_interval, opacity_reset_interval, densify_from_iter, densify_until_iter;
  double densify_grad_threshold;
  double spatial_lr_scale; /* GaussianModel::spatial_lr_scale_ (the NeRF++ radius) */
  float bg[3];             /* background colour (white_background_ ? 1 : 0) */
  unsigned int seed;       /* view sampling */
  int active_sh_degree;    /* SH degree at the first iteration (the reference starts at 0 and adds one every 1000) */
  int random_background;   /* random_background_ (src/arguments/params.h:68): a fresh uniform background colour every
                              iteration (counter-based, from `seed` and the iteration number) instead of bg */
} gs_train_config;
typedef struct gs_trainer gs_trainer;
#define GS_SAMPLE_SHUFFLE 0 /* without replacement: a stack of all views, one popped at random, refilled when empty */
#define GS_SAMPLE_CYCLIC 1  /* views in the order they were added */

/* Parameters in the layout of GaussianModel (xyz (N,3), features_dc (N,1,3), features_rest (N,K-1,3), opacity
 * (N,1), scaling (N,3), rotation (N,4); raw, i.e. before the activations); host or device pointers; copied. */
GS_API int gs_trainer_create(gs_trainer **out, int device, const gs_train_config *cfg, int N, int sh_coeffs,
                             const float *xyz, const float *features_dc, const float *features_rest,
                             const float *opacity, const float *scaling, const float *rotation);
GS_API void gs_trainer_destroy(gs_trainer *t);
GS_API const char *gs_trainer_last_error(const gs_trainer *t);
/* camera: H, W, tanfov, matrices, campos (gs_camera_settings fills them); gt_host: (3,H,W) floats, caller-owned
 * and alive as long as the trainer (pinned memory makes the per-iteration copy asynchronous). */
GS_API int gs_trainer_add_view(gs_trainer *t, const gs_settings *camera, const float *gt_host);
/* n_iterations more iterations; losses[n_iterations] (host) receives every iteration's loss in order. */
GS_API int gs_trainer_run(gs_trainer *t, int n_iterations, float *losses, int sampling, float densify_extent);
GS_API int gs_trainer_num_gaussians(const gs_trainer *t);
GS_API int gs_trainer_iteration(const gs_trainer *t);
GS_API int gs_trainer_retries(const gs_trainer *t); /* iterations repeated after a pair-capacity overflow */
/* Fix the (tile, Gaussian) pair capacity of the trainer's forwards instead of letting the first gs_trainer_run probe
 * it with one synchronous forward per view.  Too small a value is not an error: an overflowing iteration is void on
 * the device, and the loop repeats it at once under its own iteration number with a grown capacity (and after a
 * densification step the capacity follows N). */
GS_API int gs_trainer_set_pair_capacity(gs_trainer *t, unsigned long long pairs);
/* Checkpoint of the LIVE training state -- parameters, Adam moments and step counters, densification statistics,
 * iteration, SH-degree schedule, pair capacity, the sampler's position -- the start_checkpoint hook the reference's
 * train() leaves commented out (src/utils/train_utils.cpp:110-113; what it would save:
 * GaussianModel::capture, src/scene/gaussian_model.cpp:228-262).  One file in the dependency-free flat container of
 * 3d_gaussian_splatting_b200/checkpoint.py (this library does not link libtorch); checkpoint.native_to_libtorch turns
 * it into the seven torch::save archives CoreParams::restore reads (gaussian_model.cpp:152-202).  Call between
 * gs_trainer_run calls.  gs_trainer_load replaces the model (N may differ) and resumes exactly where save left. */
GS_API int gs_trainer_save(gs_trainer *t, const char *path);
GS_API int gs_trainer_load(gs_trainer *t, const char *path);
GS_API int gs_trainer_get_params(gs_trainer *t, float *xyz, float *features_dc, float *features_rest,
                                 float *opacity, float *scaling, float *rotation); /* any may be NULL */

/* ---- host-side conventions (CPU, no device needed) -----------------------------------------
 * C restatements of the reference's math helpers so that a C/C++ caller can build gs_settings
 * exactly as Camera's constructor would. */
GS_API double gs_focal2fov(double focal, double pixels);                  /* graphics_utils.cpp:4-7 */
GS_API void gs_get_world2view2(const double R[9], const double t[3], const double translate[3], double scale,
                               double Rt_out[16]);                         /* graphics_utils.cpp:32-43 */
GS_API void gs_get_projection_matrix(double znear, double zfar, double fovX, double fovY,
                                     double P_out[16]);                    /* graphics_utils.cpp:46-72 */
/* Camera ctor tensors (src/scene/camera.cpp:44-45,66-71) -> a filled gs_settings
 * (H, W, tanfov, matrices, campos; bg/sh/scale_modifier left to the caller). */
GS_API void gs_camera_settings(const double R[9], const double t[3], const double translate[3], double scale,
                               double fovX, double fovY, int H, int W, gs_settings *out);
GS_API float gs_expon_lr(float lr_init, float lr_final, int lr_delay_steps, float lr_delay_mult, int max_steps,
                         int step);                                        /* general_utils.cpp:112-142 */

#ifdef __cplusplus
}
#endif
#endif /* GSPLAT_B200_H */
