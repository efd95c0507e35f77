// ssim.cu — photometric loss of the train step: (1 - lambda) * L1 + lambda * (1 - SSIM),
// lambda = lambda_dssim_ (src/arguments/params.h:62); the loop body that would call it is missing in the
// reference (src/utils/train_utils.cpp:138-145).  SSIM: per channel, 11x11 Gaussian window (sigma 1.5),
// zero-padded "same" convolution, C1 = 0.01^2, C2 = 0.03^2 -- the oracle's definition (oracle/orc_ssim.c).
//
// Two fused kernels over 32x16 pixel tiles of one channel (grid.z = 3):
//   ssim_forward_kernel   stages the 42x26 halo of image and target in shared memory, runs the separable
//                         window over the five products (x, y, xx, yy, xy) -- rows, then columns -- evaluates the
//                         SSIM map and its partial derivatives w.r.t. the windowed means (dmu1, de11, de12;
//                         12 B per value, the only intermediate that goes to HBM) and reduces the L1 and SSIM sums.
//   ssim_backward_kernel  runs the same window over the three partial maps (the adjoint of a symmetric zero-padded
//                         window is itself) and combines  dL/dx = w*dmu1 + 2x w*de11 + y w*de12  with the L1 sign.
#include "gs_common.cuh"

#define SS_TX 32
#define SS_TY 16
#define SS_R 5
#define SS_HX (SS_TX + 2 * SS_R)  // 42
#define SS_HY (SS_TY + 2 * SS_R)  // 26
#define SS_THREADS (SS_TX * SS_TY)

struct SsimWindow {
  float g[2 * SS_R + 1];
};

__global__ void __launch_bounds__(SS_THREADS)
ssim_forward_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, SsimWindow win,
                    float* __restrict__ dmu, float* __restrict__ de11, float* __restrict__ de12,
                    double* __restrict__ acc /* [0] sum |x-y|, [1] sum SSIM */) {
  __shared__ float s_x[SS_HY][SS_HX + 1];
  __shared__ float s_y[SS_HY][SS_HX + 1];
  __shared__ float s_h[5][SS_HY][SS_TX];
  __shared__ float s_red[2][SS_THREADS / 32];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_TX + tx;
  const int x0 = blockIdx.x * SS_TX, y0 = blockIdx.y * SS_TY;
  const size_t plane = (size_t)H * W;
  const float* xp = img + blockIdx.z * plane;
  const float* yp = gt + blockIdx.z * plane;

  for (int k = tid; k < SS_HY * SS_HX; k += SS_THREADS) {
    const int r = k / SS_HX, c = k - r * SS_HX;
    const int gy = y0 + r - SS_R, gx = x0 + c - SS_R;
    const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
    s_x[r][c] = in ? __ldg(xp + (size_t)gy * W + gx) : 0.f;
    s_y[r][c] = in ? __ldg(yp + (size_t)gy * W + gx) : 0.f;
  }
  __syncthreads();
  // rows: 26 x 32 outputs, five products each
  for (int k = tid; k < SS_HY * SS_TX; k += SS_THREADS) {
    const int r = k / SS_TX, c = k - r * SS_TX;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
    for (int t = 0; t <= 2 * SS_R; ++t) {
      const float w = win.g[t], xv = s_x[r][c + t], yv = s_y[r][c + t];
      a0 = fmaf(w, xv, a0);
      a1 = fmaf(w, yv, a1);
      a2 = fmaf(w, xv * xv, a2);
      a3 = fmaf(w, yv * yv, a3);
      a4 = fmaf(w, xv * yv, a4);
    }
    s_h[0][r][c] = a0;
    s_h[1][r][c] = a1;
    s_h[2][r][c] = a2;
    s_h[3][r][c] = a3;
    s_h[4][r][c] = a4;
  }
  __syncthreads();
  // columns + map
  float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
  for (int t = 0; t <= 2 * SS_R; ++t) {
    const float w = win.g[t];
    m1 = fmaf(w, s_h[0][ty + t][tx], m1);
    m2 = fmaf(w, s_h[1][ty + t][tx], m2);
    e11 = fmaf(w, s_h[2][ty + t][tx], e11);
    e22 = fmaf(w, s_h[3][ty + t][tx], e22);
    e12 = fmaf(w, s_h[4][ty + t][tx], e12);
  }
  const int gx = x0 + tx, gy = y0 + ty;
  float l1 = 0.f, ss = 0.f;
  if (gx < W && gy < H) {
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const float s11 = e11 - m1 * m1, s22 = e22 - m2 * m2, s12 = e12 - m1 * m2;
    const float A1 = 2.f * m1 * m2 + C1, A2 = 2.f * s12 + C2;
    const float B1 = m1 * m1 + m2 * m2 + C1, B2 = s11 + s22 + C2;
    const float inv = 1.0f / (B1 * B2);
    const float S = A1 * A2 * inv;
    const float dA1 = A2 * inv, dA2 = A1 * inv, dB1 = -S / B1, dB2 = -S / B2;
    const size_t o = blockIdx.z * plane + (size_t)gy * W + gx;
    dmu[o] = 2.f * (dA1 * m2 + dB1 * m1 - dB2 * m1 - dA2 * m2);
    de11[o] = dB2;
    de12[o] = 2.f * dA2;
    ss = S;
    l1 = fabsf(s_x[ty + SS_R][tx + SS_R] - s_y[ty + SS_R][tx + SS_R]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if ((tid & 31) == 0) {
    s_red[0][tid >> 5] = l1;
    s_red[1][tid >> 5] = ss;
  }
  __syncthreads();
  if (tid < 2) {
    float t = 0.f;
    for (int k = 0; k < SS_THREADS / 32; ++k) t += s_red[tid][k];
    atomicAdd(acc + tid, (double)t);
  }
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_backward_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, SsimWindow win,
                     const float* __restrict__ dmu, const float* __restrict__ de11, const float* __restrict__ de12,
                     float k_ssim /* -lambda / n */, float k_l1 /* (1 - lambda) / n */, float* __restrict__ dL_dimg) {
  __shared__ float s_in[3][SS_HY][SS_HX + 1];
  __shared__ float s_h[3][SS_HY][SS_TX];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_TX + tx;
  const int x0 = blockIdx.x * SS_TX, y0 = blockIdx.y * SS_TY;
  const size_t plane = (size_t)H * W, base = blockIdx.z * plane;
  for (int k = tid; k < SS_HY * SS_HX; k += SS_THREADS) {
    const int r = k / SS_HX, c = k - r * SS_HX;
    const int gy = y0 + r - SS_R, gx = x0 + c - SS_R;
    const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
    const size_t o = base + (size_t)gy * W + gx;
    s_in[0][r][c] = in ? __ldg(dmu + o) : 0.f;
    s_in[1][r][c] = in ? __ldg(de11 + o) : 0.f;
    s_in[2][r][c] = in ? __ldg(de12 + o) : 0.f;
  }
  __syncthreads();
  for (int k = tid; k < SS_HY * SS_TX; k += SS_THREADS) {
    const int r = k / SS_TX, c = k - r * SS_TX;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int t = 0; t <= 2 * SS_R; ++t) {
      const float w = win.g[t];
      a0 = fmaf(w, s_in[0][r][c + t], a0);
      a1 = fmaf(w, s_in[1][r][c + t], a1);
      a2 = fmaf(w, s_in[2][r][c + t], a2);
    }
    s_h[0][r][c] = a0;
    s_h[1][r][c] = a1;
    s_h[2][r][c] = a2;
  }
  __syncthreads();
  const int gx = x0 + tx, gy = y0 + ty;
  if (gx >= W || gy >= H) return;
  float c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
  for (int t = 0; t <= 2 * SS_R; ++t) {
    const float w = win.g[t];
    c0 = fmaf(w, s_h[0][ty + t][tx], c0);
    c1 = fmaf(w, s_h[1][ty + t][tx], c1);
    c2 = fmaf(w, s_h[2][ty + t][tx], c2);
  }
  const size_t o = base + (size_t)gy * W + gx;
  const float x = __ldg(img + o), y = __ldg(gt + o);
  const float dS = c0 + 2.f * x * c1 + y * c2;
  const float d = x - y;
  const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
  dL_dimg[o] = fmaf(k_ssim, dS, k_l1 * sg);
}

__global__ void photometric_finalize_kernel(const double* acc, double inv_n, double lambda, float* loss_out) {
  *loss_out = (float)((1.0 - lambda) * acc[0] * inv_n + lambda * (1.0 - acc[1] * inv_n));
}

size_t gs_ssim_workspace_bytes(int H, int W) { return (size_t)9 * H * W * sizeof(float); }

cudaError_t gs_launch_photometric_loss(int H, int W, const float* img, const float* gt, float lambda, float* work,
                                       double* acc, float* loss_out, float* dL_dimg, cudaStream_t stream) {
  SsimWindow win;
  double g[2 * SS_R + 1], sum = 0;
  for (int i = 0; i <= 2 * SS_R; ++i) {
    const double d = (double)(i - SS_R);
    g[i] = exp(-(d * d) / (2.0 * 1.5 * 1.5));
    sum += g[i];
  }
  for (int i = 0; i <= 2 * SS_R; ++i) win.g[i] = (float)(g[i] / sum);
  const size_t n = (size_t)3 * H * W;
  float* dmu = work;
  float* de11 = work + n;
  float* de12 = work + 2 * n;
  cudaError_t e = cudaMemsetAsync(acc, 0, 2 * sizeof(double), stream);
  if (e != cudaSuccess) return e;
  const dim3 grid((W + SS_TX - 1) / SS_TX, (H + SS_TY - 1) / SS_TY, 3), block(SS_TX, SS_TY);
  ssim_forward_kernel<<<grid, block, 0, stream>>>(H, W, img, gt, win, dmu, de11, de12, acc);
  if (dL_dimg)
    ssim_backward_kernel<<<grid, block, 0, stream>>>(H, W, img, gt, win, dmu, de11, de12, -lambda / (float)n,
                                                     (1.0f - lambda) / (float)n, dL_dimg);
  photometric_finalize_kernel<<<1, 1, 0, stream>>>(acc, 1.0 / (double)n, (double)lambda, loss_out);
  return cudaGetLastError();
}
