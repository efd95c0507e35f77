/*
 * oracle/orc_ssim.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Photometric loss of the train step the reference's loop is missing (src/utils/train_utils.cpp:138-145):
 *     loss = (1 - lambda) * L1(img, gt) + lambda * (1 - SSIM(img, gt)),   lambda = lambda_dssim_ = 0.2
 * (src/arguments/params.h:62).  SSIM follows the published 3DGS trainer the reference transliterates: per
 * channel, 11x11 Gaussian window (sigma 1.5, normalised, separable), zero padding ("same" convolution),
 * C1 = 0.01^2, C2 = 0.03^2, mean over all 3*H*W values.  Pinned against a torch (conv2d + autograd)
 * restatement in tests/test_oracle_optim.py.
 *
 * Backward: with mu1 = w*x, mu2 = w*y, e11 = w*(x*x), e22 = w*(y*y), e12 = w*(x*y) (w* = windowed mean)
 * and the per-pixel map S(mu1, mu2, e11, e22, e12), the gradient w.r.t. the rendered image x is
 *     dL/dx = w*(dS/dmu1) + 2 x . w*(dS/de11) + y . w*(dS/de12)          (the window is symmetric),
 * scaled by -lambda / (3 H W); the L1 part adds (1 - lambda) * sign(x - y) / (3 H W).
 */
#include "orc_common.h"

#define SSIM_R 5
#define SSIM_W 11

static void ssim_window(double g[SSIM_W]) {
  double sum = 0;
  for (int i = 0; i < SSIM_W; ++i) {
    const double d = (double)(i - SSIM_R);
    g[i] = exp(-(d * d) / (2.0 * 1.5 * 1.5));
    sum += g[i];
  }
  for (int i = 0; i < SSIM_W; ++i) g[i] /= sum;
}

/* separable "same" convolution with zero padding of one H x W plane (double accumulation) */
static void conv_plane(int H, int W, const double *in, const double g[SSIM_W], double *tmp, double *out) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {
      double acc = 0;
      for (int k = -SSIM_R; k <= SSIM_R; ++k) {
        const int xx = x + k;
        if (xx >= 0 && xx < W) acc += g[k + SSIM_R] * in[(size_t)y * W + xx];
      }
      tmp[(size_t)y * W + x] = acc;
    }
#pragma omp parallel for schedule(static)
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {
      double acc = 0;
      for (int k = -SSIM_R; k <= SSIM_R; ++k) {
        const int yy = y + k;
        if (yy >= 0 && yy < H) acc += g[k + SSIM_R] * tmp[(size_t)yy * W + x];
      }
      out[(size_t)y * W + x] = acc;
    }
}

int orc_ssim_available(void) { return 1; }

/* Returns the loss; writes dL/dimg (3,H,W) when dL_dimg != NULL; *l1_out / *ssim_out (nullable) receive the
 * two terms (mean |x-y| and mean SSIM). */
double orc_photometric_loss(int H, int W, const real *img, const real *gt, double lambda, real *dL_dimg,
                            double *l1_out, double *ssim_out) {
  const size_t P = (size_t)H * W, n = 3 * P;
  const double C1 = 0.01 * 0.01, C2 = 0.03 * 0.03;
  double g[SSIM_W];
  ssim_window(g);
  double *buf = (double *)malloc(sizeof(double) * P * 12);
  double *x = buf, *y = buf + P, *t = buf + 2 * P, *tmp = buf + 3 * P;
  double *mu1 = buf + 4 * P, *mu2 = buf + 5 * P, *e11 = buf + 6 * P, *e22 = buf + 7 * P, *e12 = buf + 8 * P;
  double *dm = buf + 9 * P, *d11 = buf + 10 * P, *d12 = buf + 11 * P;
  double l1 = 0, ssim = 0;
  for (int c = 0; c < 3; ++c) {
    const real *xi = img + c * P, *yi = gt + c * P;
    double l1c = 0;
#pragma omp parallel for schedule(static) reduction(+ : l1c)
    for (size_t i = 0; i < P; ++i) {
      x[i] = (double)xi[i];
      y[i] = (double)yi[i];
      l1c += fabs(x[i] - y[i]);
    }
    l1 += l1c;
    conv_plane(H, W, x, g, tmp, mu1);
    conv_plane(H, W, y, g, tmp, mu2);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < P; ++i) t[i] = x[i] * x[i];
    conv_plane(H, W, t, g, tmp, e11);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < P; ++i) t[i] = y[i] * y[i];
    conv_plane(H, W, t, g, tmp, e22);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < P; ++i) t[i] = x[i] * y[i];
    conv_plane(H, W, t, g, tmp, e12);
    double sc = 0;
#pragma omp parallel for schedule(static) reduction(+ : sc)
    for (size_t i = 0; i < P; ++i) {
      const double m1 = mu1[i], m2 = mu2[i];
      const double s11 = e11[i] - m1 * m1, s22 = e22[i] - m2 * m2, s12 = e12[i] - m1 * m2;
      const double A1 = 2 * m1 * m2 + C1, A2 = 2 * s12 + C2;
      const double B1 = m1 * m1 + m2 * m2 + C1, B2 = s11 + s22 + C2;
      const double S = (A1 * A2) / (B1 * B2);
      sc += S;
      /* partials of S w.r.t. the five windowed means, then chained to mu1 / e11 / e12 */
      const double dS_dA1 = A2 / (B1 * B2), dS_dA2 = A1 / (B1 * B2);
      const double dS_dB1 = -S / B1, dS_dB2 = -S / B2;
      const double dS_ds11 = dS_dB2, dS_ds12 = 2 * dS_dA2;
      /* s11 = e11 - m1^2, s12 = e12 - m1 m2 */
      dm[i] = dS_dA1 * 2 * m2 + dS_dB1 * 2 * m1 + dS_ds11 * (-2 * m1) + dS_ds12 * (-m2);
      d11[i] = dS_ds11;
      d12[i] = dS_ds12;
    }
    ssim += sc;
    if (dL_dimg) {
      conv_plane(H, W, dm, g, tmp, mu1);   /* reuse mu1, e11, e12 as outputs */
      conv_plane(H, W, d11, g, tmp, e11);
      conv_plane(H, W, d12, g, tmp, e12);
      real *go = dL_dimg + c * P;
      const double ks = -lambda / (double)n, kl = (1.0 - lambda) / (double)n;
#pragma omp parallel for schedule(static)
      for (size_t i = 0; i < P; ++i) {
        const double dS = mu1[i] + 2 * x[i] * e11[i] + y[i] * e12[i];
        const double d = x[i] - y[i];
        const double sg = d > 0 ? 1.0 : (d < 0 ? -1.0 : 0.0);
        go[i] = (real)(ks * dS + kl * sg);
      }
    }
  }
  free(buf);
  l1 /= (double)n;
  ssim /= (double)n;
  if (l1_out) *l1_out = l1;
  if (ssim_out) *ssim_out = ssim;
  return (1.0 - lambda) * l1 + lambda * (1.0 - ssim);
}
