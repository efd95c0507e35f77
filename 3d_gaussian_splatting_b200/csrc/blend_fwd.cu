// blend_fwd.cu — K3: forward alpha compositing, one 256-thread block per
// 16x16 tile.  Each warp owns an 8x4 pixel sub-block (tight footprint => whole-warp rejects
// are common for small splats).  Gaussian records (48 B AoS) are gathered in sorted order into
// shared memory 256 at a time and broadcast-read by all pixels.
// Skip decisions (power > 0, alpha < 1/255) use the specified numerics in gs_common.cuh and
// are bit-identical to the oracle's; colour accumulation uses FMAs (smooth differences only).
#include "gs_common.cuh"

#define BF_THREADS 256

__global__ void __launch_bounds__(BF_THREADS)
blend_forward_kernel(GsDims d, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                     const GsGeom* __restrict__ geom, float bg0, float bg1, float bg2, float* __restrict__ out_color,
                     float* __restrict__ final_T, uint32_t* __restrict__ n_contrib) {
  __shared__ float4 s_q0[BF_THREADS];
  __shared__ float4 s_q1[BF_THREADS];
  __shared__ float s_b[BF_THREADS];

  const int tile = blockIdx.x;
  const int tx = tile % d.gx, ty = tile / d.gx;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int px = tx * GS_TILE + (warp & 1) * 8 + (lane & 7);
  const int py = ty * GS_TILE + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < d.W && py < d.H;
  const float pixx = (float)px, pixy = (float)py;

  const uint2 range = ranges[tile];
  int todo = (int)(range.y - range.x);
  const int rounds = (todo + BF_THREADS - 1) / BF_THREADS;

  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
  uint32_t contributor = 0, last_contributor = 0;

  for (int r = 0; r < rounds; ++r, todo -= BF_THREADS) {
    if (__syncthreads_count(done) == BF_THREADS) break;
    const uint32_t idx = range.x + r * BF_THREADS + tid;
    if (idx < range.y) {
      const uint32_t id = __ldg(point_list + idx);
      const float4* gp = reinterpret_cast<const float4*>(geom + id);
      const float4 q0 = __ldg(gp), q1 = __ldg(gp + 1);
      const float b = __ldg(reinterpret_cast<const float*>(gp + 2));
      s_q0[tid] = q0;
      s_q1[tid] = q1;
      s_b[tid] = b;
    }
    __syncthreads();
    const int n = min(BF_THREADS, todo);
    for (int j = 0; !done && j < n; ++j) {
      ++contributor;
      const float4 q0 = s_q0[j];
      const float4 q1 = s_q1[j];
      const float dx = q0.x - pixx, dy = q0.y - pixy;
      const float power = gs_power(q0.z, q0.w, q1.x, dx, dy);
      if (power > 0.0f) continue;
      float G;
      const float alpha = gs_alpha(q1.y, power, G);
      if (alpha < GS_ALPHA_MIN) continue;
      const float test_T = T * (1.0f - alpha);
      if (test_T < GS_T_MIN) {
        done = true;
        continue;
      }
      const float w = alpha * T;
      C0 = fmaf(q1.z, w, C0);
      C1 = fmaf(q1.w, w, C1);
      C2 = fmaf(s_b[j], w, C2);
      T = test_T;
      last_contributor = contributor;
    }
  }
  if (inside) {
    const size_t pix = (size_t)py * d.W + px;
    const size_t plane = (size_t)d.H * d.W;
    final_T[pix] = T;
    n_contrib[pix] = last_contributor;
    out_color[pix] = fmaf(T, bg0, C0);
    out_color[plane + pix] = fmaf(T, bg1, C1);
    out_color[2 * plane + pix] = fmaf(T, bg2, C2);
  }
}

cudaError_t gs_launch_blend_forward(const GsDims& d, const uint2* ranges, const uint32_t* point_list,
                                    const GsGeom* geom, const float bg[3], float* out_color, float* final_T,
                                    uint32_t* n_contrib, cudaStream_t stream) {
  if (d.ntiles == 0) return cudaSuccess;
  blend_forward_kernel<<<d.ntiles, BF_THREADS, 0, stream>>>(d, ranges, point_list, geom, bg[0], bg[1], bg[2],
                                                           out_color, final_T, n_contrib);
  return cudaGetLastError();
}
