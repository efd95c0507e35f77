/*
 * oracle/orc_blend.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Stages K3 / K4a: per-pixel front-to-back alpha compositing over each
 * 16x16 tile's depth-sorted list, and its reverse-order backward.  No reference
 * counterpart; image layout is CHW to match Camera::original_image_
 * (src/scene/camera.cpp:56-58).  Constants (†) in orc_common.h.
 *
 * Gradient convention note: like the published algorithm, the min(0.99, .) alpha clamp is
 * NOT treated as a gradient gate (d alpha / d G = opacity even when clamped).
 * dL/dconic.y is the true partial derivative w.r.t. the off-diagonal conic entry B
 * (power = -0.5*(A dx^2 + C dy^2) - B dx dy).
 * Backward sums are accumulated in double, per (tile, list entry), then reduced per
 * Gaussian in list order: deterministic, and more accurate than the fp32 atomics on the GPU.
 */
#include "orc_common.h"

/* K3.  out_color (3,H,W); final_T (H,W); n_contrib (H,W) = index (1-based, within the tile
 * list) of the last Gaussian that contributed.  stats[0] = D_eff (sum over tiles of the
 * largest n_contrib... see below), stats[1] = sum over pixels of n_contrib,
 * stats[2] = number of (pixel, Gaussian) pairs that passed both skip tests. */
void orc_blend_forward(const orc_settings *st, const uint32_t *ranges, const uint32_t *point_list,
                       const real *means2D, const real *conic_opacity, const real *rgb, real *out_color,
                       real *final_T, uint32_t *n_contrib, uint64_t *stats) {
  const int H = st->H, W = st->W;
  const int gx = orc_grid_x(W), gy = orc_grid_y(H);
  uint64_t d_eff = 0, sum_contrib = 0, sum_blend = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : d_eff, sum_contrib, sum_blend)
  for (int tile = 0; tile < gx * gy; ++tile) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    uint32_t tile_max_scanned = 0;
    for (int ly = 0; ly < ORC_TILE; ++ly)
      for (int lx = 0; lx < ORC_TILE; ++lx) {
        const int px = tx * ORC_TILE + lx, py = ty * ORC_TILE + ly;
        if (px >= W || py >= H) continue;
        const real pixx = RC(px), pixy = RC(py);
        real T = RC(1), C0 = 0, C1 = 0, C2 = 0;
        uint32_t contributor = 0, last = 0, scanned = 0;
        for (uint32_t j = r0; j < r1; ++j) {
          const uint32_t g = point_list[j];
          ++contributor;
          scanned = contributor;
          const real dx = means2D[2 * g] - pixx, dy = means2D[2 * g + 1] - pixy;
          const real *co = conic_opacity + 4 * g;
          const real power = orc_power(RC(0.5) * co[0], co[1], RC(0.5) * co[2], dx, dy);
          if (power > RC(0)) continue;
          const real alpha = R_FMIN(ORC_ALPHA_MAX, co[3] * orc_exp(power));
          if (alpha < ORC_ALPHA_MIN) continue;
          const real test_T = T * (RC(1) - alpha);
          if (test_T < ORC_T_MIN) break;
          const real w = alpha * T;
          C0 += rgb[3 * g] * w;
          C1 += rgb[3 * g + 1] * w;
          C2 += rgb[3 * g + 2] * w;
          T = test_T;
          last = contributor;
          ++sum_blend;
        }
        const size_t pix = (size_t)py * W + px;
        final_T[pix] = T;
        n_contrib[pix] = last;
        out_color[0 * (size_t)H * W + pix] = C0 + T * st->bg[0];
        out_color[1 * (size_t)H * W + pix] = C1 + T * st->bg[1];
        out_color[2 * (size_t)H * W + pix] = C2 + T * st->bg[2];
        sum_contrib += last;
        if (scanned > tile_max_scanned) tile_max_scanned = scanned;
      }
    d_eff += tile_max_scanned; /* entries this tile had to stage before every pixel was done */
  }
  if (stats) {
    stats[0] = d_eff;
    stats[1] = sum_contrib;
    stats[2] = sum_blend;
  }
}

/* K4a.  dL_dpixels (3,H,W).  Outputs (overwritten): dL_dmean2D (N,2) [pixel-space gradient
 * scaled by 0.5*W / 0.5*H, i.e. w.r.t. NDC — the quantity the densification statistic
 * xyz_gradient_accum_ (src/scene/gaussian_model.h:19) accumulates the norm of],
 * dL_dconic (N,3) = d/d(A,B,C), dL_dopacity (N), dL_dcolor (N,3). */
void orc_blend_backward(const orc_settings *st, int N, const uint32_t *ranges, const uint32_t *point_list,
                        const real *means2D, const real *conic_opacity, const real *rgb, const real *final_T,
                        const uint32_t *n_contrib, const real *dL_dpixels, real *dL_dmean2D, real *dL_dconic,
                        real *dL_dopacity, real *dL_dcolor) {
  const int H = st->H, W = st->W;
  const int gx = orc_grid_x(W), gy = orc_grid_y(H);
  const int ntiles = gx * gy;
  /* total list length = max range end */
  uint64_t total = 0;
  for (int t = 0; t < ntiles; ++t)
    if (ranges[2 * t + 1] > total) total = ranges[2 * t + 1];
  double *part = (double *)calloc((size_t)total * 9 + 1, sizeof(double));
  const real ddelx_dx = RC(0.5) * RC(W), ddely_dy = RC(0.5) * RC(H);
#pragma omp parallel for schedule(dynamic, 4)
  for (int tile = 0; tile < ntiles; ++tile) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    (void)r1;
    for (int ly = 0; ly < ORC_TILE; ++ly)
      for (int lx = 0; lx < ORC_TILE; ++lx) {
        const int px = tx * ORC_TILE + lx, py = ty * ORC_TILE + ly;
        if (px >= W || py >= H) continue;
        const size_t pix = (size_t)py * W + px;
        const real pixx = RC(px), pixy = RC(py);
        const real T_final = final_T[pix];
        real T = T_final;
        const uint32_t last = n_contrib[pix];
        const real dLp[3] = {dL_dpixels[pix], dL_dpixels[(size_t)H * W + pix], dL_dpixels[2 * (size_t)H * W + pix]};
        const real bg_dot = st->bg[0] * dLp[0] + st->bg[1] * dLp[1] + st->bg[2] * dLp[2];
        real accum[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0;
        for (uint32_t k = last; k >= 1; --k) {
          const uint32_t j = r0 + k - 1;
          const uint32_t g = point_list[j];
          const real dx = means2D[2 * g] - pixx, dy = means2D[2 * g + 1] - pixy;
          const real *co = conic_opacity + 4 * g;
          const real power = orc_power(RC(0.5) * co[0], co[1], RC(0.5) * co[2], dx, dy);
          if (power > RC(0)) continue;
          const real G = orc_exp(power);
          const real alpha = R_FMIN(ORC_ALPHA_MAX, co[3] * G);
          if (alpha < ORC_ALPHA_MIN) continue;
          T = T / (RC(1) - alpha);
          const real dchannel_dcolor = alpha * T;
          real dL_dalpha = 0;
          double *pp = part + (size_t)j * 9;
          for (int ch = 0; ch < 3; ++ch) {
            const real c = rgb[3 * g + ch];
            accum[ch] = last_alpha * last_color[ch] + (RC(1) - last_alpha) * accum[ch];
            last_color[ch] = c;
            dL_dalpha += (c - accum[ch]) * dLp[ch];
            pp[6 + ch] += (double)(dchannel_dcolor * dLp[ch]);
          }
          dL_dalpha *= T;
          last_alpha = alpha;
          dL_dalpha += (-T_final / (RC(1) - alpha)) * bg_dot;
          const real dL_dG = co[3] * dL_dalpha;
          const real gdx = G * dx, gdy = G * dy;
          const real dG_ddelx = -gdx * co[0] - gdy * co[1];
          const real dG_ddely = -gdy * co[2] - gdx * co[1];
          pp[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
          pp[1] += (double)(dL_dG * dG_ddely * ddely_dy);
          pp[2] += (double)(RC(-0.5) * gdx * dx * dL_dG);
          pp[3] += (double)(-gdx * dy * dL_dG);
          pp[4] += (double)(RC(-0.5) * gdy * dy * dL_dG);
          pp[5] += (double)(G * dL_dalpha);
        }
      }
  }
  /* deterministic per-Gaussian reduction in list order */
  double *acc = (double *)calloc((size_t)N * 9 + 1, sizeof(double));
  for (uint64_t j = 0; j < total; ++j) {
    const uint32_t g = point_list[j];
    for (int c = 0; c < 9; ++c) acc[(size_t)g * 9 + c] += part[j * 9 + c];
  }
  for (int g = 0; g < N; ++g) {
    const double *a = acc + (size_t)g * 9;
    dL_dmean2D[2 * g] = (real)a[0];
    dL_dmean2D[2 * g + 1] = (real)a[1];
    dL_dconic[3 * g] = (real)a[2];
    dL_dconic[3 * g + 1] = (real)a[3];
    dL_dconic[3 * g + 2] = (real)a[4];
    dL_dopacity[g] = (real)a[5];
    dL_dcolor[3 * g] = (real)a[6];
    dL_dcolor[3 * g + 1] = (real)a[7];
    dL_dcolor[3 * g + 2] = (real)a[8];
  }
  free(part);
  free(acc);
}
