// optim.cu — K5: fused Adam over the six parameter groups of
// GaussianModel::setup (src/scene/gaussian_model.cpp:323-345) + densification statistics
// (fields at src/scene/gaussian_model.h:18-20), and the L1 photometric loss.
// Adam follows libtorch's Adam::step with its defaults (only lr is set by the reference), in the
// arithmetic of libtorch's CUDA elementwise kernels (the reference keeps its tensors on kCUDA):
//   m = fma(1-b1, g, m*b1);  v = fma((1-b2)*g, g, v*b2);
//   denom = sqrt(v) * (1/sqrt(1-b2^t)) + eps;  p = fma(-lr/(1-b1^t), m/denom, p)
// with IEEE sqrt and division -- the same specified numerics as oracle/orc_optim.c, so the
// result is bit-identical to the oracle.
// Purely HBM-bound: 28 B per parameter float (+4 B when the gradient is re-zeroed); streaming
// 16-byte accesses (evict-first: nothing here is reused within the step), 8 independent vector
// loads per thread in flight at <= 64 registers so four 256-thread blocks stay resident per SM.
#include "gs_common.cuh"

#define AD_THREADS 256
#define AD_VEC_PER_THREAD 2
#define AD_FLOATS_PER_BLOCK (AD_THREADS * AD_VEC_PER_THREAD * 4)

struct AdamSeg {
  float* p;
  float* g;
  float* m;
  float* v;
  size_t n;         // floats in this segment
  int block_start;  // first block of this segment
  int vec_ok;       // all four pointers 16-byte aligned
  float neg_step, inv_bc2_sqrt;
};
struct AdamArgs {
  AdamSeg seg[6];
  int stats_block_start, total_blocks;
  float b1, b2, omb1, omb2, eps;
  int zero_grads;
  const float* skip_guard;  // device float; non-zero => the step is invalid (pair-capacity overflow), do nothing
  int N;
  const int* radii;
  const float* dL_dmeans2D;
  int* max_radii2D;
  float* xyz_grad_accum;
  float* denom;
};

__global__ void __launch_bounds__(AD_THREADS, 4) adam_stats_kernel(const __grid_constant__ AdamArgs a) {
  if (a.skip_guard && __ldg(a.skip_guard) != 0.0f) return;
  const int b = blockIdx.x;
  if (b >= a.stats_block_start) {
    // ---- densification statistics, one thread per Gaussian ----
    const int i = (b - a.stats_block_start) * AD_THREADS + threadIdx.x;
    if (i < a.N) {
      const int r = a.radii[i];
      if (r > 0) {
        if (r > a.max_radii2D[i]) a.max_radii2D[i] = r;
        const float2 g = reinterpret_cast<const float2*>(a.dL_dmeans2D)[i];
        a.xyz_grad_accum[i] += __fsqrt_rn(__fadd_rn(__fmul_rn(g.x, g.x), __fmul_rn(g.y, g.y)));
        a.denom[i] += 1.0f;
      }
    }
    return;
  }
  int si = 0;
#pragma unroll
  for (int k = 1; k < 6; ++k)
    if (b >= a.seg[k].block_start) si = k;
  const AdamSeg& s = a.seg[si];
  const float b1 = a.b1, b2 = a.b2, omb1 = a.omb1, omb2 = a.omb2, eps = a.eps;
  const float neg_step = s.neg_step, inv_bc2_sqrt = s.inv_bc2_sqrt;
  const size_t base = (size_t)(b - s.block_start) * AD_FLOATS_PER_BLOCK;
  if (s.vec_ok && base + AD_FLOATS_PER_BLOCK <= s.n) {
    float4 p[AD_VEC_PER_THREAD], g[AD_VEC_PER_THREAD], m[AD_VEC_PER_THREAD], v[AD_VEC_PER_THREAD];
    const size_t v0 = base / 4 + threadIdx.x;
    float4* __restrict__ pp = reinterpret_cast<float4*>(s.p) + v0;
    float4* __restrict__ gp = reinterpret_cast<float4*>(s.g) + v0;
    float4* __restrict__ mp = reinterpret_cast<float4*>(s.m) + v0;
    float4* __restrict__ vp = reinterpret_cast<float4*>(s.v) + v0;
#pragma unroll
    for (int k = 0; k < AD_VEC_PER_THREAD; ++k) {
      g[k] = __ldcs(gp + k * AD_THREADS);
      m[k] = __ldcs(mp + k * AD_THREADS);
      v[k] = __ldcs(vp + k * AD_THREADS);
      p[k] = __ldcs(pp + k * AD_THREADS);
    }
    // packed fast path when the whole warp's vectors are inside its range, scalar path otherwise (gs_common.cuh)
    const GsAdamConst kc = {b1, b2, omb1, omb2, eps};
    GsAdamQuad q[AD_VEC_PER_THREAD];
    bool idle[AD_VEC_PER_THREAD];
    bool any_active = false;
    int mode = GS_ADAM_FAST;
#pragma unroll
    for (int k = 0; k < AD_VEC_PER_THREAD; ++k) {
      idle[k] = gs_adam_idle(g[k], m[k], v[k]);  // exact fixed point (gs_common.cuh): nothing to write
      const int ok = gs_adam_quad_moments(g[k], m[k], v[k], kc, q[k]);
      if (!idle[k]) mode = min(mode, ok);
      any_active = any_active || !idle[k];
    }
    if (!__any_sync(0xffffffffu, any_active)) return;
    if (__all_sync(0xffffffffu, mode == GS_ADAM_FAST)) {
#pragma unroll
      for (int k = 0; k < AD_VEC_PER_THREAD; ++k) gs_adam_quad_step<false>(p[k], m[k], v[k], q[k], eps, neg_step, inv_bc2_sqrt);
    } else if (__all_sync(0xffffffffu, mode >= GS_ADAM_FAST_Z)) {
#pragma unroll
      for (int k = 0; k < AD_VEC_PER_THREAD; ++k) gs_adam_quad_step<true>(p[k], m[k], v[k], q[k], eps, neg_step, inv_bc2_sqrt);
    } else {
      // (tiny first moments, GS_ADAM_FAST_T, stay on the scalar path in THIS kernel: with eight vectors per thread the
      //  extra branch costs 10 registers and spills at the 64 that four blocks per SM allow -- 0.44 -> 0.50 ms for
      //  every state, measured; the fused kernel and the peer kernel, one or two vectors in flight, have the path)
#pragma unroll
      for (int k = 0; k < AD_VEC_PER_THREAD; ++k)
        if (!idle[k]) gs_adam_quad_scalar(p[k], g[k], m[k], v[k], kc, neg_step, inv_bc2_sqrt);
    }
#pragma unroll
    for (int k = 0; k < AD_VEC_PER_THREAD; ++k) {
      if (idle[k]) continue;
      __stcs(mp + k * AD_THREADS, m[k]);
      __stcs(vp + k * AD_THREADS, v[k]);
      __stcs(pp + k * AD_THREADS, p[k]);
      if (a.zero_grads) __stcs(gp + k * AD_THREADS, make_float4(0.f, 0.f, 0.f, 0.f));
    }
  } else {
    // ragged tail / unaligned segment: scalar, still coalesced
    for (size_t idx = base + threadIdx.x; idx < s.n && idx < base + AD_FLOATS_PER_BLOCK; idx += AD_THREADS) {
      float p = s.p[idx], m = s.m[idx], v = s.v[idx];
      gs_adam_one(p, s.g[idx], m, v, b1, b2, omb1, omb2, eps, neg_step, inv_bc2_sqrt);
      s.p[idx] = p;
      s.m[idx] = m;
      s.v[idx] = v;
      if (a.zero_grads) s.g[idx] = 0.f;
    }
  }
}

// counts[s]: floats in segment s (0 = absent)
cudaError_t gs_launch_adam_counts(int N, int nseg, float* const params[], float* const grads[], float* const exp_avg[],
                                  float* const exp_avg_sq[], const size_t counts[], const double lr[],
                                  const long long step[], double beta1, double beta2, double eps, const int* radii,
                                  const float* dL_dmeans2D, int* max_radii2D, float* xyz_grad_accum, float* denom,
                                  int zero_grads, const float* skip_guard, cudaStream_t stream) {
  AdamArgs a;
  int blocks = 0;
  for (int s = 0; s < 6; ++s) {
    AdamSeg& g = a.seg[s];
    const bool on = s < nseg && params[s] && counts[s] > 0;
    g.p = on ? params[s] : nullptr;
    g.g = on ? grads[s] : nullptr;
    g.m = on ? exp_avg[s] : nullptr;
    g.v = on ? exp_avg_sq[s] : nullptr;
    g.n = on ? counts[s] : 0;
    g.block_start = blocks;
    g.vec_ok = ((((uintptr_t)g.p | (uintptr_t)g.g | (uintptr_t)g.m | (uintptr_t)g.v) & 15) == 0);
    // libtorch evaluates the bias corrections in double and applies them as fp32 scalars
    const long long st = on ? step[s] : 1;
    const double bc1 = 1.0 - pow(beta1, (double)st);
    const double bc2 = 1.0 - pow(beta2, (double)st);
    g.neg_step = (float)(-((on ? lr[s] : 0.0) / bc1));
    g.inv_bc2_sqrt = 1.0f / (float)sqrt(bc2);
    blocks += (int)((g.n + AD_FLOATS_PER_BLOCK - 1) / AD_FLOATS_PER_BLOCK);
  }
  a.stats_block_start = blocks;
  const bool stats = radii && dL_dmeans2D && max_radii2D && xyz_grad_accum && denom;
  if (stats) blocks += (N + AD_THREADS - 1) / AD_THREADS;
  a.total_blocks = blocks;
  a.b1 = (float)beta1;
  a.b2 = (float)beta2;
  a.omb1 = (float)(1.0 - beta1);
  a.omb2 = (float)(1.0 - beta2);
  a.eps = (float)eps;
  a.zero_grads = zero_grads;
  a.skip_guard = skip_guard;
  a.N = N;
  a.radii = radii;
  a.dL_dmeans2D = dL_dmeans2D;
  a.max_radii2D = max_radii2D;
  a.xyz_grad_accum = xyz_grad_accum;
  a.denom = denom;
  if (blocks == 0) return cudaSuccess;
  adam_stats_kernel<<<blocks, AD_THREADS, 0, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t gs_launch_adam(int N, float* const params[6], float* const grads[6], float* const exp_avg[6],
                           float* const exp_avg_sq[6], const int widths[6], const double lr[6],
                           const long long step[6], double beta1, double beta2, double eps, const int* radii,
                           const float* dL_dmeans2D, int* max_radii2D, float* xyz_grad_accum, float* denom,
                           int zero_grads, const float* skip_guard, cudaStream_t stream) {
  size_t counts[6];
  for (int s = 0; s < 6; ++s) counts[s] = (params[s] && widths[s] > 0) ? (size_t)N * widths[s] : 0;
  return gs_launch_adam_counts(N, 6, params, grads, exp_avg, exp_avg_sq, counts, lr, step, beta1, beta2, eps, radii,
                               dL_dmeans2D, max_radii2D, xyz_grad_accum, denom, zero_grads, skip_guard, stream);
}

// ---- sharded step over peer memory (multi-GPU, one process per GPU) ----
// Rank r owns the Gaussians [r*R, (r+1)*R).  The backward of every rank has stored its gradient rows for this
// shard straight into this rank's slot buffers (K4b writes through NVLink, preprocess_bwd.cu); this kernel sums
// the G slots in RANK ORDER (deterministic; for G = 2 the same value as any all-reduce), applies Adam to the
// owned rows (moments exist only here) and stores the new parameters into the replica of EVERY rank -- its own
// through HBM, the others through NVLink -- so reduce-scatter, optimiser and all-gather are one pass over the
// shard with no staging copy.
#define ADP_MAX_RANKS 8
struct AdamPeerSeg {
  float* p;                         // own replica, at the shard's first row
  float* m;
  float* v;
  const float* src[ADP_MAX_RANKS];  // gradient slots (local memory), one per source rank
  float* dst[ADP_MAX_RANKS - 1];    // the other ranks' replicas (peer memory), at the shard's first row
  size_t n;
  int block_start;
  float neg_step, inv_bc2_sqrt;
};
struct AdamPeerArgs {
  AdamPeerSeg seg[6];
  int stats_block_start, total_blocks;
  float b1, b2, omb1, omb2, eps;
  const float* skip_guard;
  int rank;        // the slot this rank's own backward wrote (local, dense)
  int sparse_seg;  // segment whose REMOTE slots are written sparsely and must be re-zeroed after use (-1: none)
  int local_mask;  // bit k: segment k is updated in the OWN replica only (the caller distributes it with copy engines)
  int N;
  const int* radii;
  const float* dL_dmeans2D;
  int* max_radii2D;
  float* xyz_grad_accum;
  float* denom;
};

template <int G>
__global__ void __launch_bounds__(AD_THREADS, 2) adam_peer_kernel(const __grid_constant__ AdamPeerArgs a) {
  // an overflowed forward voids the step -- but the SPARSE slots (see below) must still be cleared
  const bool skip = a.skip_guard && __ldg(a.skip_guard) != 0.0f;
  // grid-stride over the work blocks: the launch may be capped (GS_PEER_ADAM_GRID) so that a call that runs beside the
  // next forward on a second stream -- bound by NVLink, not by the SMs -- leaves most warp slots to the forward
  for (int b = blockIdx.x; b < a.total_blocks; b += gridDim.x) {
  if (b >= a.stats_block_start) {
    if (skip) return;
    const int i = (b - a.stats_block_start) * AD_THREADS + threadIdx.x;
    if (i < a.N) {
      const int r = a.radii[i];
      if (r > 0) {
        if (r > a.max_radii2D[i]) a.max_radii2D[i] = r;
        const float2 g = reinterpret_cast<const float2*>(a.dL_dmeans2D)[i];
        a.xyz_grad_accum[i] += __fsqrt_rn(__fadd_rn(__fmul_rn(g.x, g.x), __fmul_rn(g.y, g.y)));
        a.denom[i] += 1.0f;
      }
    }
    continue;
  }
  int si = 0;
#pragma unroll
  for (int k = 1; k < 6; ++k)
    if (b >= a.seg[k].block_start) si = k;
  const AdamPeerSeg& s = a.seg[si];
  // The f_rest slots are written SPARSELY by the peers' backward (only rows with a non-zero gradient cross
  // NVLink, preprocess_bwd.cu): a row that is not written must read as zero, so whatever this kernel consumes
  // from a remote f_rest slot it puts back to zero (local stores, non-zero vectors only).
  const bool sparse = si == a.sparse_seg;
  const bool remote = !((a.local_mask >> si) & 1);
  if (skip && !sparse) continue;
  const float b1 = a.b1, b2 = a.b2, omb1 = a.omb1, omb2 = a.omb2, eps = a.eps;
  const float neg_step = s.neg_step, inv_bc2_sqrt = s.inv_bc2_sqrt;
  constexpr int FPB = AD_THREADS * 4;  // one float4 per thread: G + 3 vector loads in flight
  const size_t base = (size_t)(b - s.block_start) * FPB;
  if (base + FPB <= s.n) {
    const size_t v0 = base / 4 + threadIdx.x;
    float4 gs[G];
#pragma unroll
    for (int r = 0; r < G; ++r) gs[r] = __ldcs(reinterpret_cast<const float4*>(s.src[r]) + v0);
    float4 m = make_float4(0.f, 0.f, 0.f, 0.f), v = m, p = m;
    if (!skip) {
      m = __ldcs(reinterpret_cast<const float4*>(s.m) + v0);
      v = __ldcs(reinterpret_cast<const float4*>(s.v) + v0);
      p = __ldcs(reinterpret_cast<const float4*>(s.p) + v0);
    }
    if (sparse) {
#pragma unroll
      for (int r = 0; r < G; ++r)
        if (r != a.rank && ((__float_as_uint(gs[r].x) | __float_as_uint(gs[r].y) | __float_as_uint(gs[r].z) |
                             __float_as_uint(gs[r].w)) & 0x7fffffffu))
          reinterpret_cast<float4*>(const_cast<float*>(s.src[r]))[v0] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float4 g = gs[0];
#pragma unroll
    for (int r = 1; r < G; ++r) {
      g.x = __fadd_rn(g.x, gs[r].x);
      g.y = __fadd_rn(g.y, gs[r].y);
      g.z = __fadd_rn(g.z, gs[r].z);
      g.w = __fadd_rn(g.w, gs[r].w);
    }
    // exact fixed point: every replica already holds this p, nothing crosses NVLink
    const bool idle = skip || gs_adam_idle(g, m, v);
    if (!__any_sync(0xffffffffu, !idle)) continue;
    {
      const GsAdamConst kc = {b1, b2, omb1, omb2, eps};
      GsAdamQuad q;
      const int ok = idle ? GS_ADAM_FAST : gs_adam_quad_moments(g, m, v, kc, q);
      if (__all_sync(0xffffffffu, ok == GS_ADAM_FAST)) gs_adam_quad_step<false>(p, m, v, q, eps, neg_step, inv_bc2_sqrt);
      else if (__all_sync(0xffffffffu, ok >= GS_ADAM_FAST_Z)) gs_adam_quad_step<true>(p, m, v, q, eps, neg_step, inv_bc2_sqrt);
      else if (__all_sync(0xffffffffu, ok >= GS_ADAM_FAST_T)) gs_adam_quad_step_tiny(p, g, m, v, q, kc, neg_step, inv_bc2_sqrt, !idle);
      else if (!idle) gs_adam_quad_scalar(p, g, m, v, kc, neg_step, inv_bc2_sqrt);
    }
    if (idle) continue;
    if (remote) {
#pragma unroll
      for (int r = 0; r < G - 1; ++r) reinterpret_cast<float4*>(s.dst[r])[v0] = p;  // NVLink stores first
    }
    __stcs(reinterpret_cast<float4*>(s.m) + v0, m);
    __stcs(reinterpret_cast<float4*>(s.v) + v0, v);
    reinterpret_cast<float4*>(s.p)[v0] = p;   // read again by the next forward: normal caching
  } else {
    for (size_t idx = base + threadIdx.x; idx < s.n && idx < base + FPB; idx += AD_THREADS) {
      float g = 0.f;
#pragma unroll
      for (int r = 0; r < G; ++r) {
        const float gr = s.src[r][idx];
        g = r == 0 ? gr : __fadd_rn(g, gr);
        if (sparse && r != a.rank && gr != 0.0f) const_cast<float*>(s.src[r])[idx] = 0.0f;
      }
      if (skip) continue;
      float p = s.p[idx], m = s.m[idx], v = s.v[idx];
      gs_adam_one(p, g, m, v, b1, b2, omb1, omb2, eps, neg_step, inv_bc2_sqrt);
      if (remote) {
#pragma unroll
        for (int r = 0; r < G - 1; ++r) s.dst[r][idx] = p;
      }
      s.p[idx] = p;
      s.m[idx] = m;
      s.v[idx] = v;
    }
  }
  }
}

// params[r*6+k]: rank r's (N, widths[k]) array (r == rank: local, else peer-mapped); slots[s*6+k]: the local slot
// for source rank s (R x widths[k]); exp_avg / exp_avg_sq: shard-local (R x widths[k]).
cudaError_t gs_launch_adam_peer(int nranks, int rank, int N, int R, float* const params[],
                                const float* const slots[], float* const exp_avg[6], float* const exp_avg_sq[6],
                                const int widths[6], const double lr[6], const long long step[6], double beta1,
                                double beta2, double eps, const int* radii, const float* dL_dmeans2D,
                                int* max_radii2D, float* xyz_grad_accum, float* denom, const float* skip_guard,
                                int sparse_rest, int max_blocks, int local_mask, cudaStream_t stream) {
  AdamPeerArgs a;
  a.local_mask = local_mask;
  const long long row0 = (long long)rank * R;
  const long long rows = row0 >= N ? 0 : ((long long)N - row0 < R ? (long long)N - row0 : (long long)R);
  int blocks = 0;
  for (int k = 0; k < 6; ++k) {
    AdamPeerSeg& g = a.seg[k];
    const bool on = widths[k] > 0 && rows > 0 && params[rank * 6 + k];
    const size_t off = (size_t)row0 * widths[k];
    g.p = on ? params[rank * 6 + k] + off : nullptr;
    g.m = on ? exp_avg[k] : nullptr;
    g.v = on ? exp_avg_sq[k] : nullptr;
    for (int s = 0; s < ADP_MAX_RANKS; ++s) g.src[s] = (on && s < nranks) ? slots[s * 6 + k] : nullptr;
    int d = 0;
    for (int r = 0; r < nranks; ++r)
      if (r != rank) g.dst[d++] = on ? params[r * 6 + k] + off : nullptr;
    for (; d < ADP_MAX_RANKS - 1; ++d) g.dst[d] = nullptr;
    g.n = on ? (size_t)rows * widths[k] : 0;
    g.block_start = blocks;
    const long long st = on ? step[k] : 1;
    const double bc1 = 1.0 - pow(beta1, (double)st);
    const double bc2 = 1.0 - pow(beta2, (double)st);
    g.neg_step = (float)(-((on ? lr[k] : 0.0) / bc1));
    g.inv_bc2_sqrt = 1.0f / (float)sqrt(bc2);
    blocks += (int)((g.n + AD_THREADS * 4 - 1) / (AD_THREADS * 4));
  }
  a.stats_block_start = blocks;
  const bool stats = radii && dL_dmeans2D && max_radii2D && xyz_grad_accum && denom;
  if (stats) blocks += (N + AD_THREADS - 1) / AD_THREADS;
  a.total_blocks = blocks;
  a.b1 = (float)beta1;
  a.b2 = (float)beta2;
  a.omb1 = (float)(1.0 - beta1);
  a.omb2 = (float)(1.0 - beta2);
  a.eps = (float)eps;
  a.skip_guard = skip_guard;
  a.rank = rank;
  a.sparse_seg = sparse_rest ? 2 : -1;  // f_rest: 76% of the gradient bytes at SH degree 3 (push layout only)
  a.N = N;
  a.radii = radii;
  a.dL_dmeans2D = dL_dmeans2D;
  a.max_radii2D = max_radii2D;
  a.xyz_grad_accum = xyz_grad_accum;
  a.denom = denom;
  if (blocks == 0) return cudaSuccess;
  if (max_blocks > 0 && blocks > max_blocks) blocks = max_blocks;
  switch (nranks) {
#define ADP_CASE(G) case G: adam_peer_kernel<G><<<blocks, AD_THREADS, 0, stream>>>(a); break;
    ADP_CASE(1) ADP_CASE(2) ADP_CASE(3) ADP_CASE(4) ADP_CASE(5) ADP_CASE(6) ADP_CASE(7) ADP_CASE(8)
#undef ADP_CASE
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// ---- L1 loss --------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
l1_loss_kernel(size_t n, const float* __restrict__ img, const float* __restrict__ gt, float scale,
               float* __restrict__ dL_dimg, double* __restrict__ acc) {
  float local = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float d = img[i] - gt[i];
    local += fabsf(d);
    if (dL_dimg) dL_dimg[i] = d > 0.f ? scale : (d < 0.f ? -scale : 0.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  __shared__ float s_part[8];
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += s_part[k];
    atomicAdd(acc, (double)t);
  }
}
__global__ void loss_finalize_kernel(const double* acc, double inv_n, float* loss_out) {
  *loss_out = (float)(acc[0] * inv_n);
}

cudaError_t gs_launch_l1_loss(size_t n, const float* img, const float* gt, float weight, float* dL_dimg, double* acc,
                              float* loss_out, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(acc, 0, sizeof(double), stream);
  if (e != cudaSuccess) return e;
  const int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  l1_loss_kernel<<<blocks > 0 ? blocks : 1, 256, 0, stream>>>(n, img, gt, weight / (float)n, dL_dimg, acc);
  loss_finalize_kernel<<<1, 1, 0, stream>>>(acc, (double)weight / (double)n, loss_out);
  return cudaGetLastError();
}

// ---- debug: NaN / inf count (gs_settings.debug & GS_DEBUG_CHECKS; the role of --detect_anomaly) ----
__global__ void __launch_bounds__(256) count_nonfinite_kernel(const float* __restrict__ p, size_t n, uint32_t* __restrict__ count) {
  uint32_t bad = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    bad += ((__float_as_uint(p[i]) & 0x7f800000u) == 0x7f800000u) ? 1u : 0u;
  bad = __reduce_add_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(count, bad);
}
cudaError_t gs_launch_count_nonfinite(const float* p, size_t n, uint32_t* count, cudaStream_t stream) {
  const size_t want = (n + 255) / 256;
  count_nonfinite_kernel<<<(int)(want < 1184 ? (want ? want : 1) : 1184), 256, 0, stream>>>(p, n, count);
  return cudaGetLastError();
}

// ---- barrier across the ranks of the peer-memory step, in one small kernel (no NCCL, no host) ----
// Every rank owns a flag block {uint32 epoch[8]; float guard[8]; uint32 timeouts} inside its peer-mapped buffer.
// Rank r signals by storing (its guard, then `epoch` with release semantics at system scope) into slot r of EVERY
// rank's block, then waits until every slot of its OWN block has reached `epoch` (acquire loads).  Stream order
// puts the call after the kernel whose peer stores it publishes (K4b's rows in the owners' slots, the Adam kernel's
// parameters in the replicas): a kernel boundary completes those stores, the fence orders them before the flag.
// The guards ride along: guard_out = max over the ranks' guard_in (a pair-capacity overflow on ANY rank voids the
// step on EVERY rank).  Epochs only grow, and no rank can be more than one barrier ahead of another, so `>=` is the
// test.  A rank that does not arrive within ~2 s (a crashed peer) is reported in the block's timeout counter and the
// kernel returns instead of hanging the GPU.
struct PeerBarrierArgs {
  uint32_t* flags[ADP_MAX_RANKS];  // flag block of every rank (own: local memory)
  int G, rank;
  uint32_t epoch;
  const float* guard_in;
  float* guard_out;
};
__global__ void __launch_bounds__(32) peer_barrier_kernel(const __grid_constant__ PeerBarrierArgs a) {
  const int r = threadIdx.x;
  float g = 0.0f;
  if (r < a.G) {
    const float mine = a.guard_in ? *a.guard_in : 0.0f;
    __threadfence_system();
    float* gslot = reinterpret_cast<float*>(a.flags[r] + 8) + a.rank;
    uint32_t* eslot = a.flags[r] + a.rank;
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(gslot), "f"(mine) : "memory");
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(eslot), "r"(a.epoch) : "memory");
    const uint32_t* wait = a.flags[a.rank] + r;
    const long long t0 = clock64();
    uint32_t seen;
    while (true) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(wait) : "memory");
      if ((int32_t)(seen - a.epoch) >= 0) break;
      if (clock64() - t0 > 4000000000ll) {  // ~2 s
        atomicAdd(a.flags[a.rank] + 16, 1u);
        break;
      }
      __nanosleep(100);
    }
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(g) : "l"(reinterpret_cast<const float*>(a.flags[a.rank] + 8) + r) : "memory");
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) g = fmaxf(g, __shfl_xor_sync(0xffffffffu, g, o));
  if (r == 0 && a.guard_out) *a.guard_out = g;
}
cudaError_t gs_launch_peer_barrier(int G, int rank, uint32_t* const flags[], uint32_t epoch, const float* guard_in,
                                   float* guard_out, cudaStream_t stream) {
  PeerBarrierArgs a;
  for (int r = 0; r < ADP_MAX_RANKS; ++r) a.flags[r] = r < G ? flags[r] : nullptr;
  a.G = G;
  a.rank = rank;
  a.epoch = epoch;
  a.guard_in = guard_in;
  a.guard_out = guard_out;
  peer_barrier_kernel<<<1, 32, 0, stream>>>(a);
  return cudaGetLastError();
}
