// blend_bwd.cu — K4a: backward alpha compositing.  Same tiling as K3; each
// pixel walks its tile list back-to-front from its last contributor, recomputing alpha and
// transmittance.  Per-Gaussian partial gradients are reduced across the warp with shuffles and
// leave the SM as three 16-byte vector reductions (red.global.add.v4.f32) per (warp, Gaussian).
//
// Gradient record per Gaussian (48 B, pre-zeroed):
//   g0 = { dL/dmean2D.x, dL/dmean2D.y, dL/dA, dL/dB }   (mean2D gradient in NDC units: x0.5W/0.5H)
//   g1 = { dL/dC, dL/dopacity, dL/dr, dL/dg }
//   g2 = { dL/db, -, -, - }
#include "gs_common.cuh"

#define BB_THREADS 256

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(BB_THREADS)
blend_backward_kernel(GsDims d, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      const GsGeom* __restrict__ geom, float bg0, float bg1, float bg2,
                      const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpixels, float4* __restrict__ grad_geom) {
  __shared__ float4 s_q0[BB_THREADS];
  __shared__ float4 s_q1[BB_THREADS];
  __shared__ float s_b[BB_THREADS];
  __shared__ uint32_t s_id[BB_THREADS];

  const int tile = blockIdx.x;
  const int tx = tile % d.gx, ty = tile / d.gx;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int px = tx * GS_TILE + (warp & 1) * 8 + (lane & 7);
  const int py = ty * GS_TILE + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < d.W && py < d.H;
  const float pixx = (float)px, pixy = (float)py;
  const size_t pix = (size_t)py * d.W + px;
  const size_t plane = (size_t)d.H * d.W;

  const uint2 range = ranges[tile];
  const int total = (int)(range.y - range.x);
  const int rounds = (total + BB_THREADS - 1) / BB_THREADS;

  const float T_final = inside ? final_T[pix] : 0.f;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix] : 0;
  float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f;
  if (inside) {
    dLp0 = dL_dpixels[pix];
    dLp1 = dL_dpixels[plane + pix];
    dLp2 = dL_dpixels[2 * plane + pix];
  }
  const float bg_dot = bg0 * dLp0 + bg1 * dLp1 + bg2 * dLp2;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * (float)d.W, ddely_dy = 0.5f * (float)d.H;

  // the whole warp can start at the largest last_contributor among its lanes
  int warp_last = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));

  int contributor = total;  // index (0-based) of the entry about to be processed is contributor-1
  int todo = total;
  for (int r = 0; r < rounds; ++r, todo -= BB_THREADS) {
    __syncthreads();
    const int k = r * BB_THREADS + tid;  // position counted from the back
    if (k < total) {
      const uint32_t id = __ldg(point_list + range.y - 1 - k);
      const float4* gp = reinterpret_cast<const float4*>(geom + id);
      s_id[tid] = id;
      s_q0[tid] = __ldg(gp);
      s_q1[tid] = __ldg(gp + 1);
      s_b[tid] = __ldg(reinterpret_cast<const float*>(gp + 2));
    }
    __syncthreads();
    const int n = min(BB_THREADS, todo);
    for (int j = 0; j < n; ++j) {
      --contributor;
      if (contributor >= warp_last) continue;  // warp-uniform
      const float4 q0 = s_q0[j];
      const float4 q1 = s_q1[j];
      const float dx = q0.x - pixx, dy = q0.y - pixy;
      const float power = gs_power(q0.z, q0.w, q1.x, dx, dy);
      float G = 0.f;
      float alpha = 0.f;
      bool valid = (contributor < last_contributor) && !(power > 0.0f);
      if (valid) {
        alpha = gs_alpha(q1.y, power, G);
        valid = !(alpha < GS_ALPHA_MIN);
      }
      if (__ballot_sync(0xffffffffu, valid) == 0u) continue;

      float v_mx = 0.f, v_my = 0.f, v_A = 0.f, v_B = 0.f, v_C = 0.f, v_o = 0.f, v_r = 0.f, v_g = 0.f, v_b = 0.f;
      if (valid) {
        const float cb = s_b[j];
        T = T / (1.0f - alpha);
        const float dchannel_dcolor = alpha * T;
        acc0 = last_alpha * lc0 + (1.0f - last_alpha) * acc0;
        acc1 = last_alpha * lc1 + (1.0f - last_alpha) * acc1;
        acc2 = last_alpha * lc2 + (1.0f - last_alpha) * acc2;
        lc0 = q1.z;
        lc1 = q1.w;
        lc2 = cb;
        float dL_dalpha = (q1.z - acc0) * dLp0 + (q1.w - acc1) * dLp1 + (cb - acc2) * dLp2;
        v_r = dchannel_dcolor * dLp0;
        v_g = dchannel_dcolor * dLp1;
        v_b = dchannel_dcolor * dLp2;
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final / (1.0f - alpha)) * bg_dot;
        const float dL_dG = q1.y * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        const float A = 2.0f * q0.z, B = q0.w, C = 2.0f * q1.x;
        const float dG_ddelx = -gdx * A - gdy * B;
        const float dG_ddely = -gdy * C - gdx * B;
        v_mx = dL_dG * dG_ddelx * ddelx_dx;
        v_my = dL_dG * dG_ddely * ddely_dy;
        v_A = -0.5f * gdx * dx * dL_dG;
        v_B = -gdx * dy * dL_dG;
        v_C = -0.5f * gdy * dy * dL_dG;
        v_o = G * dL_dalpha;
      }
      v_mx = warp_sum(v_mx);
      v_my = warp_sum(v_my);
      v_A = warp_sum(v_A);
      v_B = warp_sum(v_B);
      v_C = warp_sum(v_C);
      v_o = warp_sum(v_o);
      v_r = warp_sum(v_r);
      v_g = warp_sum(v_g);
      v_b = warp_sum(v_b);
      if (lane < 3) {
        float4* dst = grad_geom + 3 * (size_t)s_id[j] + lane;
        float4 v;
        if (lane == 0) v = make_float4(v_mx, v_my, v_A, v_B);
        else if (lane == 1) v = make_float4(v_C, v_o, v_r, v_g);
        else v = make_float4(v_b, 0.f, 0.f, 0.f);
        atomicAdd(dst, v);
      }
    }
  }
}

cudaError_t gs_launch_blend_backward(const GsDims& d, const uint2* ranges, const uint32_t* point_list,
                                     const GsGeom* geom, const float bg[3], const float* final_T,
                                     const uint32_t* n_contrib, const float* dL_dpixels, float4* grad_geom,
                                     cudaStream_t stream) {
  if (d.ntiles == 0) return cudaSuccess;
  blend_backward_kernel<<<d.ntiles, BB_THREADS, 0, stream>>>(d, ranges, point_list, geom, bg[0], bg[1], bg[2],
                                                            final_T, n_contrib, dL_dpixels, grad_geom);
  return cudaGetLastError();
}
