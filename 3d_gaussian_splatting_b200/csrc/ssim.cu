// ssim.cu — photometric loss of the train step: (1 - lambda) * L1 + lambda * (1 - SSIM),
// lambda = lambda_dssim_ (src/arguments/params.h:62); the loop body that would call it is missing in the
// reference (src/utils/train_utils.cpp:138-145).  SSIM: per channel, 11x11 Gaussian window (sigma 1.5),
// zero-padded "same" convolution, C1 = 0.01^2, C2 = 0.03^2 -- the oracle's definition (oracle/orc_ssim.c).
//
// Two fused kernels over 32x32 pixel tiles of one channel (grid.z = 3), 256 threads each producing 4 outputs per
// pass so every staged value (and, in the forward, its three products) is reused from registers by 4 windows:
//   ssim_forward_kernel   stages the 42x42 halo of image and target in shared memory, runs the separable
//                         window over the five products (x, y, xx, yy, xy) -- rows, then columns -- evaluates the
//                         SSIM map and its partial derivatives w.r.t. the windowed means (dmu1, de11, de12;
//                         12 B per value, the only intermediate that goes to HBM) and reduces the L1 and SSIM sums.
//   ssim_backward_kernel  runs the same window over the three partial maps (the adjoint of a symmetric zero-padded
//                         window is itself) and combines  dL/dx = w*dmu1 + 2x w*de11 + y w*de12  with the L1 sign.
#include "gs_common.cuh"

#include <mutex>

#define SS_TX 32
#define SS_TY 32
#define SS_R 5
#define SS_HX (SS_TX + 2 * SS_R)  // 42
#define SS_HY (SS_TY + 2 * SS_R)  // 42
#define SS_THREADS 256            // every thread produces 4 outputs of each pass (register-blocked window)
#define SS_V 4
#define SS_PITCH (SS_HX + 1)      // 43: with 4-wide column groups the 32 lanes of a warp hit 32 different banks

struct SsimWindow {
  float g[2 * SS_R + 1];
};

// Column pass for one thread: 4 consecutive rows of column `col` from NP row-filtered planes.
template <int NP>
__device__ __forceinline__ void ssim_columns(const float (*s_h)[SS_HY][SS_TX], const SsimWindow& win, int row0, int col,
                                             float out[NP][SS_V]) {
#pragma unroll
  for (int q = 0; q < NP; ++q) {
    float v[SS_V + 2 * SS_R];
#pragma unroll
    for (int j = 0; j < SS_V + 2 * SS_R; ++j) v[j] = s_h[q][row0 + j][col];
#pragma unroll
    for (int u = 0; u < SS_V; ++u) {
      float a = 0.f;
#pragma unroll
      for (int t = 0; t <= 2 * SS_R; ++t) a = fmaf(win.g[t], v[u + t], a);
      out[q][u] = a;
    }
  }
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_forward_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, SsimWindow win,
                    float* __restrict__ dmu, float* __restrict__ de11, float* __restrict__ de12,
                    float2* __restrict__ partial /* per block: {sum |x-y|, sum SSIM} */) {
  __shared__ float s_x[SS_HY][SS_PITCH];
  __shared__ float s_y[SS_HY][SS_PITCH];
  __shared__ float s_h[5][SS_HY][SS_TX];
  __shared__ float s_red[2][SS_THREADS / 32];
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * SS_TX, y0 = blockIdx.y * SS_TY;
  const size_t plane = (size_t)H * W;
  const float* xp = img + blockIdx.z * plane;
  const float* yp = gt + blockIdx.z * plane;

  // halo staging: all loads of a thread are issued before the first store (one memory latency per block, not nine)
  constexpr int SS_STAGE = (SS_HY * SS_HX + SS_THREADS - 1) / SS_THREADS;
  {
    float vx[SS_STAGE], vy[SS_STAGE];
#pragma unroll
    for (int i = 0; i < SS_STAGE; ++i) {
      const int k = tid + i * SS_THREADS;
      const int r = k / SS_HX, c = k - r * SS_HX;
      const int gy = y0 + r - SS_R, gx = x0 + c - SS_R;
      const bool in = k < SS_HY * SS_HX && gy >= 0 && gy < H && gx >= 0 && gx < W;
      vx[i] = in ? __ldg(xp + (size_t)gy * W + gx) : 0.f;
      vy[i] = in ? __ldg(yp + (size_t)gy * W + gx) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < SS_STAGE; ++i) {
      const int k = tid + i * SS_THREADS;
      const int r = k / SS_HX, c = k - r * SS_HX;
      if (k < SS_HY * SS_HX) {
        s_x[r][c] = vx[i];
        s_y[r][c] = vy[i];
      }
    }
  }
  __syncthreads();
  // rows: 42 rows x 8 groups of 4 adjacent outputs; the 14 staged values (and their products) are reused by the 4
  for (int k = tid; k < SS_HY * (SS_TX / SS_V); k += SS_THREADS) {
    const int r = k / (SS_TX / SS_V), c = (k - r * (SS_TX / SS_V)) * SS_V;
    float xv[SS_V + 2 * SS_R], yv[SS_V + 2 * SS_R];
#pragma unroll
    for (int j = 0; j < SS_V + 2 * SS_R; ++j) {
      xv[j] = s_x[r][c + j];
      yv[j] = s_y[r][c + j];
    }
    float acc[5][SS_V];
#pragma unroll
    for (int u = 0; u < SS_V; ++u) acc[0][u] = acc[1][u] = acc[2][u] = acc[3][u] = acc[4][u] = 0.f;
#pragma unroll
    for (int j = 0; j < SS_V + 2 * SS_R; ++j) {
      const float xx = xv[j] * xv[j], yy = yv[j] * yv[j], xy = xv[j] * yv[j];
#pragma unroll
      for (int u = 0; u < SS_V; ++u) {
        const int t = j - u;
        if (t >= 0 && t <= 2 * SS_R) {
          const float w = win.g[t];
          acc[0][u] = fmaf(w, xv[j], acc[0][u]);
          acc[1][u] = fmaf(w, yv[j], acc[1][u]);
          acc[2][u] = fmaf(w, xx, acc[2][u]);
          acc[3][u] = fmaf(w, yy, acc[3][u]);
          acc[4][u] = fmaf(w, xy, acc[4][u]);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int u = 0; u < SS_V; ++u) s_h[q][r][c + u] = acc[q][u];
  }
  __syncthreads();
  // columns + map: thread = column tx, rows 4*rg .. 4*rg+3
  const int tx = tid & 31, row0 = (tid >> 5) * SS_V;
  float m[5][SS_V];
  ssim_columns<5>(s_h, win, row0, tx, m);
  const int gx = x0 + tx;
  float l1 = 0.f, ss = 0.f;
#pragma unroll
  for (int u = 0; u < SS_V; ++u) {
    const int gy = y0 + row0 + u;
    if (gx < W && gy < H) {
      const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
      const float m1 = m[0][u], m2 = m[1][u];
      const float s11 = m[2][u] - m1 * m1, s22 = m[3][u] - m2 * m2, s12 = m[4][u] - m1 * m2;
      const float A1 = 2.f * m1 * m2 + C1, A2 = 2.f * s12 + C2;
      const float B1 = m1 * m1 + m2 * m2 + C1, B2 = s11 + s22 + C2;
      const float inv = 1.0f / (B1 * B2);
      const float S = A1 * A2 * inv;
      const float dA1 = A2 * inv, dA2 = A1 * inv, dB1 = -S / B1, dB2 = -S / B2;
      const size_t o = blockIdx.z * plane + (size_t)gy * W + gx;
      dmu[o] = 2.f * (dA1 * m2 + dB1 * m1 - dB2 * m1 - dA2 * m2);
      de11[o] = dB2;
      de12[o] = 2.f * dA2;
      ss += S;
      l1 += fabsf(s_x[row0 + u + SS_R][tx + SS_R] - s_y[row0 + u + SS_R][tx + SS_R]);
    }
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
  if (tid == 0) {
    float t0 = 0.f, t1 = 0.f;
    for (int k = 0; k < SS_THREADS / 32; ++k) {
      t0 += s_red[0][k];
      t1 += s_red[1][k];
    }
    // one partial per block, summed in double by the finalize kernel: deterministic, no same-address atomics
    partial[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = make_float2(t0, t1);
  }
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_backward_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, SsimWindow win,
                     const float* __restrict__ dmu, const float* __restrict__ de11, const float* __restrict__ de12,
                     float k_ssim /* -lambda / n */, float k_l1 /* (1 - lambda) / n */, float* __restrict__ dL_dimg) {
  __shared__ float s_in[3][SS_HY][SS_PITCH];
  __shared__ float s_h[3][SS_HY][SS_TX];
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * SS_TX, y0 = blockIdx.y * SS_TY;
  const size_t plane = (size_t)H * W, base = blockIdx.z * plane;
  constexpr int SS_STAGE = (SS_HY * SS_HX + SS_THREADS - 1) / SS_THREADS;
  {
    float v0[SS_STAGE], v1[SS_STAGE], v2[SS_STAGE];
#pragma unroll
    for (int i = 0; i < SS_STAGE; ++i) {
      const int k = tid + i * SS_THREADS;
      const int r = k / SS_HX, c = k - r * SS_HX;
      const int gy = y0 + r - SS_R, gx = x0 + c - SS_R;
      const bool in = k < SS_HY * SS_HX && gy >= 0 && gy < H && gx >= 0 && gx < W;
      const size_t o = base + (size_t)gy * W + gx;
      v0[i] = in ? __ldg(dmu + o) : 0.f;
      v1[i] = in ? __ldg(de11 + o) : 0.f;
      v2[i] = in ? __ldg(de12 + o) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < SS_STAGE; ++i) {
      const int k = tid + i * SS_THREADS;
      const int r = k / SS_HX, c = k - r * SS_HX;
      if (k < SS_HY * SS_HX) {
        s_in[0][r][c] = v0[i];
        s_in[1][r][c] = v1[i];
        s_in[2][r][c] = v2[i];
      }
    }
  }
  __syncthreads();
  for (int k = tid; k < SS_HY * (SS_TX / SS_V); k += SS_THREADS) {
    const int r = k / (SS_TX / SS_V), c = (k - r * (SS_TX / SS_V)) * SS_V;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      float v[SS_V + 2 * SS_R];
#pragma unroll
      for (int j = 0; j < SS_V + 2 * SS_R; ++j) v[j] = s_in[q][r][c + j];
#pragma unroll
      for (int u = 0; u < SS_V; ++u) {
        float a = 0.f;
#pragma unroll
        for (int t = 0; t <= 2 * SS_R; ++t) a = fmaf(win.g[t], v[u + t], a);
        s_h[q][r][c + u] = a;
      }
    }
  }
  __syncthreads();
  const int tx = tid & 31, row0 = (tid >> 5) * SS_V;
  float cv[3][SS_V];
  ssim_columns<3>(s_h, win, row0, tx, cv);
  const int gx = x0 + tx;
  if (gx >= W) return;
#pragma unroll
  for (int u = 0; u < SS_V; ++u) {
    const int gy = y0 + row0 + u;
    if (gy < H) {
      const size_t o = base + (size_t)gy * W + gx;
      const float x = __ldg(img + o), y = __ldg(gt + o);
      const float dS = cv[0][u] + 2.f * x * cv[1][u] + y * cv[2][u];
      const float d = x - y;
      const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
      dL_dimg[o] = fmaf(k_ssim, dS, k_l1 * sg);
    }
  }
}

__global__ void __launch_bounds__(1024)
photometric_finalize_kernel(const float2* __restrict__ partial, int nblocks, double inv_n, double lambda,
                            float* __restrict__ loss_out) {
  __shared__ double s_a[32], s_b[32];
  double a = 0, b = 0;
  for (int i = threadIdx.x; i < nblocks; i += 1024) {
    const float2 p = partial[i];
    a += (double)p.x;
    b += (double)p.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  if ((threadIdx.x & 31) == 0) {
    s_a[threadIdx.x >> 5] = a;
    s_b[threadIdx.x >> 5] = b;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0, tb = 0;
    for (int k = 0; k < 32; ++k) {
      ta += s_a[k];
      tb += s_b[k];
    }
    *loss_out = (float)((1.0 - lambda) * ta * inv_n + lambda * (1.0 - tb * inv_n));
  }
}

static int ssim_blocks(int H, int W) { return ((W + SS_TX - 1) / SS_TX) * ((H + SS_TY - 1) / SS_TY) * 3; }
size_t gs_ssim_workspace_bytes(int H, int W) {
  return (((size_t)9 * H * W + 3) & ~(size_t)3) * sizeof(float) + (size_t)ssim_blocks(H, W) * sizeof(float2);
}

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
  (void)acc;
  float2* partial = reinterpret_cast<float2*>(work + ((3 * n + 3) & ~(size_t)3));
  const dim3 grid((W + SS_TX - 1) / SS_TX, (H + SS_TY - 1) / SS_TY, 3), block(SS_THREADS);
  ssim_forward_kernel<<<grid, block, 0, stream>>>(H, W, img, gt, win, dmu, de11, de12, partial);
  if (dL_dimg)
    ssim_backward_kernel<<<grid, block, 0, stream>>>(H, W, img, gt, win, dmu, de11, de12, -lambda / (float)n,
                                                     (1.0f - lambda) / (float)n, dL_dimg);
  photometric_finalize_kernel<<<1, 1024, 0, stream>>>(partial, ssim_blocks(H, W), 1.0 / (double)n, (double)lambda, loss_out);
  return cudaGetLastError();
}

// ---- allocation-free variant of the photometric loss ------------------------------------------------------------
// gs_loss_static: the same three kernels as gs_loss, with the SSIM workspace (three partial-derivative maps of
// 3*H*W floats and one float2 per block) in a file-scope device array instead of a cudaMalloc'ed context buffer, for
// callers that may not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// The array exists once per device that loads the library and is shared by every
// call: calls that overlap on the GPU (two streams, two host threads) race on it.  Nothing in it has to start at
// zero: every word a call reads (the maps inside the image, one partial per launched block) is written earlier in
// the same call, so the contents left by the previous call are never observed.
#define GS_LOSS_STATIC_MAX_BLOCKS 32768
static __device__ __align__(256) float g_loss_static_scratch[(size_t)9 * GS_LOSS_STATIC_MAX_PIXELS + 2 * GS_LOSS_STATIC_MAX_BLOCKS];

#define GS_LOSS_STATIC_MAX_DEVICES 64
static float* s_loss_static_ptr[GS_LOSS_STATIC_MAX_DEVICES];  // device address of the array, per device ordinal
static std::mutex s_loss_static_mu;

// In host code `g_loss_static_scratch` is only a handle for the runtime, not a device address: resolve it once per
// device, with that device current.  (The first resolution on a device may make the runtime load this module there;
// a caller that captures graphs makes one call outside the capture first.)
static cudaError_t loss_static_scratch(int device, float** out) {
  std::lock_guard<std::mutex> lk(s_loss_static_mu);
  if (!s_loss_static_ptr[device]) {
    float* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_loss_static_scratch);
    if (e != cudaSuccess) return e;
    s_loss_static_ptr[device] = p;
  }
  *out = s_loss_static_ptr[device];
  return cudaSuccess;
}

extern "C" GS_API int gs_loss_static(int device, int H, int W, const float* image, const float* gt, float lambda_dssim,
                                     float* loss_out, float* dL_dimage, void* stream) {
  if (device < 0 || device >= GS_LOSS_STATIC_MAX_DEVICES || H <= 0 || W <= 0 || !image || !gt || !loss_out) {
    gs_set_error(nullptr, "gs_loss_static: bad argument");
    return GS_EINVAL;
  }
  if (!(lambda_dssim >= 0.0f && lambda_dssim <= 1.0f)) {
    gs_set_error(nullptr, "gs_loss_static: lambda_dssim %g outside [0,1]", (double)lambda_dssim);
    return GS_EINVAL;
  }
  if ((long long)H * W > GS_LOSS_STATIC_MAX_PIXELS || ssim_blocks(H, W) > GS_LOSS_STATIC_MAX_BLOCKS) {
    gs_set_error(nullptr,
                 "gs_loss_static: %d x %d = %lld pixels exceeds the static workspace (GS_LOSS_STATIC_MAX_PIXELS = %d, "
                 "at most %d blocks of 32x32); use gs_loss",
                 H, W, (long long)H * W, GS_LOSS_STATIC_MAX_PIXELS, GS_LOSS_STATIC_MAX_BLOCKS / 3);
    return GS_EINVAL;
  }
  GS_CUDA_CHECK((gs_ctx*)nullptr, cudaSetDevice(device));
  float* work = nullptr;
  GS_CUDA_CHECK((gs_ctx*)nullptr, loss_static_scratch(device, &work));
  // same layout inside the array as gs_ssim_workspace_bytes() describes for the context's buffer
  GS_CUDA_CHECK((gs_ctx*)nullptr, gs_launch_photometric_loss(H, W, image, gt, lambda_dssim, work, nullptr, loss_out,
                                                             dL_dimage, (cudaStream_t)stream));
  return GS_OK;
}
