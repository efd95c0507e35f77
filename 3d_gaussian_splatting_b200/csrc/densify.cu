// densify.cu — densify / prune / opacity reset with optimiser-state surgery: the consumers of
// the statistics K5 accumulates (max_radii2D_, xyz_gradient_accum_, denom_, src/scene/gaussian_model.h:18-20)
// under the reference's schedule parameters (src/arguments/params.h:61-67).  The reference declares the state but
// not the step; the algorithm is the published 3DGS one, specified phase by phase in oracle/orc_densify.c.
//
// The published code runs three passes over growing tensors (clone: cat; split: cat + mask; prune: mask), each
// re-allocating all six parameter tensors and rebuilding the optimiser state.  Here the outcome of all three is
// decided per Gaussian in ONE classification pass (every test of the three phases depends only on the Gaussian's
// own values), three stream compactions place the survivors in the published order
//     [kept originals | kept clones | kept first children | kept second children],
// and ONE scatter pass moves parameters and Adam moments (new Gaussians start with zero moments) from the old
// slabs to the new ones, coalesced on the read side.  Between the two passes the host reads the three totals
// (12 bytes) to size the new slabs -- the only synchronisation, once per densification interval.
#include <math.h>

#include "gs_common.cuh"

#define DZ_THREADS 256

struct gs_densify_params_dev {
  float grad_threshold, min_opacity, extent, percent_dense, max_screen_size;
  unsigned long long seed;
};

__device__ __forceinline__ uint32_t dz_mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

// the specified split-sample generator of oracle/orc_densify.c (counter-based: parent index, copy, pair)
__device__ __forceinline__ void dz_split_normals(unsigned long long seed, unsigned long long parent, int copy, float z[3]) {
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const unsigned long long c = (parent * 2ull + (unsigned)copy) * 2ull + (unsigned)j;
    const uint32_t h1 = dz_mix32((uint32_t)seed ^ dz_mix32((uint32_t)c ^ dz_mix32((uint32_t)(c >> 32) + (uint32_t)(seed >> 32))));
    const uint32_t h2 = dz_mix32(h1 ^ 0x85ebca6bu);
    const float u1 = (float)((h1 >> 8) + 1u) * 0x1p-24f;
    const float u2 = (float)(h2 >> 8) * 0x1p-24f;
    const float rad = sqrtf(-2.0f * logf(u1));
    const float ang = 6.283185307179586f * u2;
    if (j == 0) {
      z[0] = rad * cosf(ang);
      z[1] = rad * sinf(ang);
    } else {
      z[2] = rad * cosf(ang);
    }
  }
}

#define DZ_KEEP_ORIG 1u
#define DZ_KEEP_CLONE 2u
#define DZ_KEEP_CHILDREN 4u
// (the fourth per-block counter is the number of CANDIDATES -- originals that are not split parents, clones and
//  children -- so that candidates - kept = pruned by the opacity / size tests, the figure the oracle reports)

__device__ __forceinline__ float dz_max3(float a, float b, float c) { return fmaxf(a, fmaxf(b, c)); }

// ---- pass 1: classify + per-block counts ---------------------------------------------------------
__global__ void __launch_bounds__(DZ_THREADS)
densify_classify_kernel(const int N, const float* __restrict__ accum, const float* __restrict__ denom,
                        const float* __restrict__ opacity, const float* __restrict__ scaling,
                        const gs_densify_params_dev prm, uint8_t* __restrict__ codes, uint32_t* __restrict__ block_counts) {
  const int i = blockIdx.x * DZ_THREADS + threadIdx.x;
  uint32_t code = 0, candidates = 0;
  if (i < N) {
    float g = accum[i] / denom[i];
    if (isnan(g)) g = 0.0f;
    const float s0 = scaling[3 * (size_t)i], s1 = scaling[3 * (size_t)i + 1], s2 = scaling[3 * (size_t)i + 2];
    const float e0 = expf(s0), e1 = expf(s1), e2 = expf(s2);
    const float smax = dz_max3(e0, e1, e2);
    const float limit = prm.percent_dense * prm.extent;
    const bool hot = g >= prm.grad_threshold;
    const bool clone = hot && smax <= limit;
    const bool split = hot && smax > limit;
    const bool faint = 1.0f / (1.0f + expf(-opacity[i])) < prm.min_opacity;
    const bool use_size = prm.max_screen_size > 0.0f;  // (the screen-size half never fires: oracle quirk 1)
    const bool big = use_size && smax > 0.1f * prm.extent;
    const bool drop = faint || big;
    if (!split && !drop) code |= DZ_KEEP_ORIG;
    if (clone && !drop) code |= DZ_KEEP_CLONE;
    if (split) {
      // the children carry log(exp(s) / 1.6); the size test reads exp() of that, as the published code does
      const float c0 = expf(logf(e0 / (0.8f * 2.0f))), c1 = expf(logf(e1 / (0.8f * 2.0f))), c2 = expf(logf(e2 / (0.8f * 2.0f)));
      const bool cbig = use_size && dz_max3(c0, c1, c2) > 0.1f * prm.extent;
      if (!faint && !cbig) code |= DZ_KEEP_CHILDREN;
    }
    candidates = split ? 2u : (clone ? 2u : 1u);
    codes[i] = (uint8_t)code;
  }
  // block totals of the three keep flags (+ pruned, for the report)
  __shared__ uint32_t s_cnt[4];
  if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0;
  __syncthreads();
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    const uint32_t m = __ballot_sync(0xffffffffu, (code >> b) & 1u);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&s_cnt[b], __popc(m));
  }
  const uint32_t cand = __reduce_add_sync(0xffffffffu, candidates);
  if ((threadIdx.x & 31) == 0 && cand) atomicAdd(&s_cnt[3], cand);
  __syncthreads();
  if (threadIdx.x < 4) block_counts[4 * blockIdx.x + threadIdx.x] = s_cnt[threadIdx.x];
}

// ---- pass 1b: exclusive scan of the block counts (one block), totals to device + pinned host ------
__global__ void __launch_bounds__(1024)
densify_scan_kernel(const int nb, uint32_t* __restrict__ block_counts, uint32_t* __restrict__ totals, uint32_t* host_mirror) {
  __shared__ uint32_t s_part[32][4];
  __shared__ uint32_t s_carry[4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 4) s_carry[tid] = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int b = base + tid;
    uint32_t v[4], inc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[k] = b < nb ? block_counts[4 * b + k] : 0u;
      inc[k] = v[k];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc[k], o);
        if (lane >= o) inc[k] += t;
      }
      if (lane == 31) s_part[warp][k] = inc[k];
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t w = s_part[lane][k], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
          if (lane >= o) wi += t;
        }
        s_part[lane][k] = wi - w;  // exclusive over warps
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t excl = s_carry[k] + s_part[warp][k] + inc[k] - v[k];
      if (b < nb) block_counts[4 * b + k] = excl;
    }
    __syncthreads();
    if (tid == 1023) {
#pragma unroll
      for (int k = 0; k < 4; ++k) s_carry[k] += s_part[31][k] + inc[k];
    }
    __syncthreads();
  }
  if (tid < 4) {
    totals[tid] = s_carry[tid];
    if (host_mirror) host_mirror[tid] = s_carry[tid];
  }
}

// ---- pass 2: scatter parameters + moments to the new slabs ----------------------------------------
struct DzSlabs {
  const float* sp[6];
  const float* sm[6];
  const float* sv[6];
  float* dp[6];
  float* dm[6];
  float* dv[6];
  int w[6];
};

__global__ void __launch_bounds__(DZ_THREADS)
densify_scatter_kernel(const int N, const uint8_t* __restrict__ codes, const uint32_t* __restrict__ block_offsets,
                       const uint32_t* __restrict__ totals, const __grid_constant__ DzSlabs sl,
                       const unsigned long long seed) {
  __shared__ int s_dst[3][DZ_THREADS];  // destination row of the original / the clone / the first child, or -1
  __shared__ uint32_t s_warp[8][3];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row0 = blockIdx.x * DZ_THREADS;
  const int i = row0 + tid;
  const uint32_t code = i < N ? codes[i] : 0u;
  const uint32_t S0 = totals[0], S1 = totals[1], S2 = totals[2];
  uint32_t rank[3];
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    const uint32_t m = __ballot_sync(0xffffffffu, (code >> b) & 1u);
    rank[b] = __popc(m & ((1u << lane) - 1u));
    if (lane == 0) s_warp[warp][b] = __popc(m);
  }
  __syncthreads();
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    uint32_t before = block_offsets[4 * blockIdx.x + b];
    for (int w = 0; w < warp; ++w) before += s_warp[w][b];
    const uint32_t seg = b == 0 ? 0u : (b == 1 ? S0 : S0 + S1);
    s_dst[b][tid] = ((code >> b) & 1u) ? (int)(seg + before + rank[b]) : -1;
  }
  __syncthreads();
  const int rows = min(DZ_THREADS, N - row0);
  // every group: coalesced read of the block's rows, up to four destination rows per source row
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const int w = sl.w[k];
    if (w == 0) continue;
    const float* __restrict__ sp = sl.sp[k] + (size_t)row0 * w;
    const float* __restrict__ sm = sl.sm[k] + (size_t)row0 * w;
    const float* __restrict__ sv = sl.sv[k] + (size_t)row0 * w;
    const bool child_computed = (k == 0 || k == 4);  // xyz and scaling of the children are written below
    int r = tid / w, c = tid % w;
    const int dr = DZ_THREADS / w, dc = DZ_THREADS % w;
    for (int f = tid; f < rows * w; f += DZ_THREADS) {
      const int d0 = s_dst[0][r], d1 = s_dst[1][r], d2 = s_dst[2][r];
      if (d0 >= 0 || d1 >= 0 || d2 >= 0) {
        const float pv = sp[f];
        if (d0 >= 0) {
          sl.dp[k][(size_t)d0 * w + c] = pv;
          sl.dm[k][(size_t)d0 * w + c] = sm[f];
          sl.dv[k][(size_t)d0 * w + c] = sv[f];
        }
        if (d1 >= 0) {
          sl.dp[k][(size_t)d1 * w + c] = pv;
          sl.dm[k][(size_t)d1 * w + c] = 0.0f;
          sl.dv[k][(size_t)d1 * w + c] = 0.0f;
        }
        if (d2 >= 0) {
          const size_t a = (size_t)d2 * w + c, b2 = (size_t)(d2 + (int)S2) * w + c;
          if (!child_computed) {
            sl.dp[k][a] = pv;
            sl.dp[k][b2] = pv;
          }
          sl.dm[k][a] = 0.0f;
          sl.dv[k][a] = 0.0f;
          sl.dm[k][b2] = 0.0f;
          sl.dv[k][b2] = 0.0f;
        }
      }
      r += dr;
      c += dc;
      if (c >= w) { c -= w; ++r; }
    }
  }
  // split children: position sampled from the parent's own Gaussian, scale shrunk by 1.6
  const int d2 = s_dst[2][tid];
  if (i < N && d2 >= 0) {
    const float* q = sl.sp[5] + 4 * (size_t)i;
    const float* s = sl.sp[4] + 3 * (size_t)i;
    const float* x = sl.sp[0] + 3 * (size_t)i;
    const float sc[3] = {expf(s[0]), expf(s[1]), expf(s[2])};
    const float nq = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const float qr = q[0] / nq, qx = q[1] / nq, qy = q[2] / nq, qz = q[3] / nq;
    const float R[9] = {1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qr * qz), 2 * (qx * qz + qr * qy),
                        2 * (qx * qy + qr * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qr * qx),
                        2 * (qx * qz - qr * qy), 2 * (qy * qz + qr * qx), 1 - 2 * (qx * qx + qy * qy)};
#pragma unroll
    for (int rep = 0; rep < 2; ++rep) {
      float z[3];
      dz_split_normals(seed, (unsigned long long)i, rep, z);
      const float smp[3] = {z[0] * sc[0], z[1] * sc[1], z[2] * sc[2]};
      const size_t d = (size_t)(d2 + rep * (int)S2);
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        sl.dp[0][3 * d + a] = R[3 * a] * smp[0] + R[3 * a + 1] * smp[1] + R[3 * a + 2] * smp[2] + x[a];
        sl.dp[4][3 * d + a] = logf(sc[a] / (0.8f * 2.0f));
      }
    }
  }
}

__global__ void __launch_bounds__(DZ_THREADS)
reset_opacity_kernel(const int N, float* __restrict__ opacity, float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq) {
  const int i = blockIdx.x * DZ_THREADS + threadIdx.x;
  if (i >= N) return;
  const float x = fminf(1.0f / (1.0f + expf(-opacity[i])), 0.01f);
  opacity[i] = logf(x / (1.0f - x));
  exp_avg[i] = 0.0f;
  exp_avg_sq[i] = 0.0f;
}

// ---- launchers ------------------------------------------------------------------------------------
cudaError_t gs_launch_densify_plan(int N, const float* accum, const float* denom, const float* opacity,
                                   const float* scaling, float grad_threshold, float min_opacity, float extent,
                                   float percent_dense, float max_screen_size, uint8_t* codes, uint32_t* block_counts,
                                   uint32_t* totals, uint32_t* host_mirror, cudaStream_t stream) {
  gs_densify_params_dev prm;
  prm.grad_threshold = grad_threshold;
  prm.min_opacity = min_opacity;
  prm.extent = extent;
  prm.percent_dense = percent_dense;
  prm.max_screen_size = max_screen_size;
  prm.seed = 0;
  const int nb = (N + DZ_THREADS - 1) / DZ_THREADS;
  if (nb > 0) densify_classify_kernel<<<nb, DZ_THREADS, 0, stream>>>(N, accum, denom, opacity, scaling, prm, codes, block_counts);
  densify_scan_kernel<<<1, 1024, 0, stream>>>(nb, block_counts, totals, host_mirror);
  return cudaGetLastError();
}

cudaError_t gs_launch_densify_apply(int N, const uint8_t* codes, const uint32_t* block_offsets, const uint32_t* totals,
                                    const int widths[6], float* const src_p[6], float* const src_m[6],
                                    float* const src_v[6], float* const dst_p[6], float* const dst_m[6],
                                    float* const dst_v[6], unsigned long long seed, cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  DzSlabs sl;
  for (int k = 0; k < 6; ++k) {
    sl.w[k] = widths[k];
    sl.sp[k] = src_p[k];
    sl.sm[k] = src_m[k];
    sl.sv[k] = src_v[k];
    sl.dp[k] = dst_p[k];
    sl.dm[k] = dst_m[k];
    sl.dv[k] = dst_v[k];
  }
  const int nb = (N + DZ_THREADS - 1) / DZ_THREADS;
  densify_scatter_kernel<<<nb, DZ_THREADS, 0, stream>>>(N, codes, block_offsets, totals, sl, seed);
  return cudaGetLastError();
}

cudaError_t gs_launch_reset_opacity(int N, float* opacity, float* exp_avg, float* exp_avg_sq, cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  reset_opacity_kernel<<<(N + DZ_THREADS - 1) / DZ_THREADS, DZ_THREADS, 0, stream>>>(N, opacity, exp_avg, exp_avg_sq);
  return cudaGetLastError();
}
