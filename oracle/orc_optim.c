/*
 * oracle/orc_optim.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Stage K5: Adam update and densification statistics, plus the L1 photometric
 * loss used by the train step.
 *   Adam: the reference builds six torch::optim::Adam with only `lr` set
 *         (src/scene/gaussian_model.cpp:323-345) => libtorch defaults beta=(0.9,0.999),
 *         eps=1e-8, no weight decay, no amsgrad.  The update order restated here is
 *         libtorch's Adam::step: m.mul_(b1).add_(g, 1-b1); v.mul_(b2).addcmul_(g, g, 1-b2);
 *         denom = (sqrt(v) / sqrt(1-b2^t)).add_(eps); p.addcdiv_(m, denom, -lr/(1-b1^t)), with
 *         the bias corrections evaluated in double and applied as fp32 scalars.
 *         SPECIFIED NUMERICS (shared by definition with csrc/optim.cu): the reference keeps its
 *         tensors on torch::kCUDA (src/scene/gaussian_model.cpp:319-320), where libtorch's
 *         elementwise kernels (a) contract `a + alpha*b` into one FMA, (b) evaluate addcmul as
 *         fma(alpha*b, c, a), (c) divide by a host scalar by multiplying with its fp32 reciprocal
 *         and (d) evaluate addcdiv as fma(alpha, b/c, a) with an IEEE division.  Each line below is
 *         one such kernel; sqrt and the division are correctly rounded.
 *         Pinned against torch.optim.Adam on CPU in tests/test_oracle_optim.py (<= 2 ulp of p).
 *   Stats: fields max_radii2D_, xyz_gradient_accum_, denom_ (src/scene/gaussian_model.h:18-20).
 */
#include "orc_common.h"

void orc_adam_step(size_t n, float *p, const float *g, float *m, float *v, double lr, double beta1,
                   double beta2, double eps, long long step) {
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  const float b1 = (float)beta1, b2 = (float)beta2;
  const float omb1 = (float)(1.0 - beta1), omb2 = (float)(1.0 - beta2);
  const float inv_bc2_sqrt = 1.0f / (float)sqrt(bc2);
  const float neg_step = (float)(-(lr / bc1));
  const float epsf = (float)eps;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n; ++i) {
    const float gi = g[i];
    const float mi = fmaf(omb1, gi, m[i] * b1);
    const float vi = fmaf(omb2 * gi, gi, v[i] * b2);
    const float denom = sqrtf(vi) * inv_bc2_sqrt + epsf;
    m[i] = mi;
    v[i] = vi;
    p[i] = fmaf(neg_step, mi / denom, p[i]);
  }
}

void orc_densify_stats(int N, const int *radii, const float *dL_dmean2D, int *max_radii2D, float *xyz_grad_accum,
                       float *denom) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < N; ++i) {
    if (radii[i] > 0) {
      if (radii[i] > max_radii2D[i]) max_radii2D[i] = radii[i];
      const float gx = dL_dmean2D[2 * i], gy = dL_dmean2D[2 * i + 1];
      xyz_grad_accum[i] += sqrtf(gx * gx + gy * gy);
      denom[i] += 1.0f;
    }
  }
}

/* L1 loss = mean |img - gt| over 3*H*W; dL/dimg = sign(img - gt) / (3*H*W).
 * (lambda_dssim_ of src/arguments/params.h:62 weights the D-SSIM term, a "next" row.) */
double orc_l1_loss(size_t n, const real *img, const real *gt, real *dL_dimg) {
  double acc = 0;
  const real inv = RC(1) / RC(n);
#pragma omp parallel for schedule(static) reduction(+ : acc)
  for (size_t i = 0; i < n; ++i) {
    const real d = img[i] - gt[i];
    acc += (double)R_FABS(d);
    if (dL_dimg) dL_dimg[i] = d > 0 ? inv : (d < 0 ? -inv : RC(0));
  }
  return acc / (double)n;
}

/* Thread control for the timed CPU leg (bench.py): launchers such as torchrun export OMP_NUM_THREADS=1, which
 * would silently turn the "all host cores" baseline into a single-threaded one. */
#ifdef _OPENMP
#include <omp.h>
int orc_set_threads(int n) {
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
}
#else
int orc_set_threads(int n) { (void)n; return 1; }
#endif
