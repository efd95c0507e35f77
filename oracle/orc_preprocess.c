/*
 * oracle/orc_preprocess.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Stage K1: per-Gaussian frustum cull, EWA projection, conic, radius,
 * tile rectangle, SH -> RGB.  The reference has no counterpart (the call is missing at
 * src/utils/train_utils.cpp:138-145); cov3D follows src/scene/gaussian_model.cpp:18-28 and
 * src/utils/general_utils.cpp:12-99, the matrix layout follows src/scene/camera.cpp:66-71.
 * Every fp operation is written as a separately rounded step in a fixed order; the CUDA
 * kernel (csrc/preprocess.cu, compiled with --fmad=false) uses the same order, which is
 * what makes radii / tiles_touched / keys bit-exact.
 */
#include "orc_common.h"

void orc_cov3d_one(const real *s, real mod, const real *q, real *cov6);

/* column-major 4x4 (flat reference buffer) applied to a point */
static inline void xform4x3(const real *m, const real *p, real *o) {
  o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
  o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
  o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}
static inline void xform4x4(const real *m, const real *p, real *o) {
  o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
  o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
  o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
  o[3] = m[3] * p[0] + m[7] * p[1] + m[11] * p[2] + m[15];
}

/* EWA: 2D covariance (a,b,c) of the projected Gaussian, before dilation. */
void orc_cov2d_one(const real *t_view, real focal_x, real focal_y, real tanfovx, real tanfovy,
                   const real *cov6, const real *V, real *abc) {
  real limx = ORC_FOV_GUARD * tanfovx, limy = ORC_FOV_GUARD * tanfovy;
  real txtz = t_view[0] / t_view[2], tytz = t_view[1] / t_view[2];
  real tx = R_FMIN(limx, R_FMAX(-limx, txtz)) * t_view[2];
  real ty = R_FMIN(limy, R_FMAX(-limy, tytz)) * t_view[2];
  real tz = t_view[2];
  /* J = [[fx/tz, 0, -fx*tx/tz^2], [0, fy/tz, -fy*ty/tz^2]] */
  real J00 = focal_x / tz, J02 = -(focal_x * tx) / (tz * tz);
  real J11 = focal_y / tz, J12 = -(focal_y * ty) / (tz * tz);
  /* W = rotation part of V (world -> view), W[r][c] = V[4*c + r]; T = J * W (2x3) */
  real T00 = J00 * V[0] + J02 * V[2];
  real T01 = J00 * V[4] + J02 * V[6];
  real T02 = J00 * V[8] + J02 * V[10];
  real T10 = J11 * V[1] + J12 * V[2];
  real T11 = J11 * V[5] + J12 * V[6];
  real T12 = J11 * V[9] + J12 * V[10];
  /* cov2D = T * Sigma * T^T ; first U = Sigma * T^T rows */
  real c0 = cov6[0], c1 = cov6[1], c2 = cov6[2], c3 = cov6[3], c4 = cov6[4], c5 = cov6[5];
  real U00 = c0 * T00 + c1 * T01 + c2 * T02; /* (Sigma * T0)_x */
  real U01 = c1 * T00 + c3 * T01 + c4 * T02;
  real U02 = c2 * T00 + c4 * T01 + c5 * T02;
  real U10 = c0 * T10 + c1 * T11 + c2 * T12;
  real U11 = c1 * T10 + c3 * T11 + c4 * T12;
  real U12 = c2 * T10 + c4 * T11 + c5 * T12;
  abc[0] = T00 * U00 + T01 * U01 + T02 * U02;
  abc[1] = T00 * U10 + T01 * U11 + T02 * U12;
  abc[2] = T10 * U10 + T11 * U11 + T12 * U12;
}

/* SH -> RGB for one Gaussian; sh_dc points at 3 floats, sh_rest at (K-1)*3 floats. */
void orc_sh_to_rgb_one(int deg, const real *pos, const real *campos, const real *sh_dc, const real *sh_rest,
                       real *rgb, unsigned char *clamped) {
  real dx = pos[0] - campos[0], dy = pos[1] - campos[1], dz = pos[2] - campos[2];
  real len = R_SQRT(dx * dx + dy * dy + dz * dz);
  real x = dx / len, y = dy / len, z = dz / len;
  for (int c = 0; c < 3; ++c) {
    real res = ORC_SH_C0 * sh_dc[c];
    if (deg > 0) {
      const real *s = sh_rest; /* s[3*(k-1)+c] is coefficient k */
      res = res - ORC_SH_C1 * y * s[0 + c] + ORC_SH_C1 * z * s[3 + c] - ORC_SH_C1 * x * s[6 + c];
      if (deg > 1) {
        real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        res = res + RC(ORC_SH_C2[0]) * xy * s[9 + c] + RC(ORC_SH_C2[1]) * yz * s[12 + c] +
              RC(ORC_SH_C2[2]) * (RC(2) * zz - xx - yy) * s[15 + c] + RC(ORC_SH_C2[3]) * xz * s[18 + c] +
              RC(ORC_SH_C2[4]) * (xx - yy) * s[21 + c];
        if (deg > 2) {
          res = res + RC(ORC_SH_C3[0]) * y * (RC(3) * xx - yy) * s[24 + c] +
                RC(ORC_SH_C3[1]) * xy * z * s[27 + c] +
                RC(ORC_SH_C3[2]) * y * (RC(4) * zz - xx - yy) * s[30 + c] +
                RC(ORC_SH_C3[3]) * z * (RC(2) * zz - RC(3) * xx - RC(3) * yy) * s[33 + c] +
                RC(ORC_SH_C3[4]) * x * (RC(4) * zz - xx - yy) * s[36 + c] +
                RC(ORC_SH_C3[5]) * z * (xx - yy) * s[39 + c] +
                RC(ORC_SH_C3[6]) * x * (xx - RC(3) * yy) * s[42 + c];
        }
      }
    }
    res = res + RC(0.5);
    clamped[c] = (unsigned char)(res < RC(0));
    rgb[c] = R_FMAX(res, RC(0));
  }
}

/*
 * K1 forward.  Inputs follow the GaussianModel getters (src/scene/gaussian_model.cpp:270-298):
 * means3D (N,3), scales (N,3) activated, rotations (N,4) wxyz, opacities (N), SH split into
 * sh_dc (row stride dc_stride floats) and sh_rest (row stride rest_stride floats) so both the
 * concatenated get_features() layout (N,K,3) and the separate features_dc_/features_rest_
 * tensors are addressable without a copy.  colors_precomp / cov3D_precomp are the
 * PipelineParams::convert_SHs_python_/compute_cov3D_python_ branches (src/arguments/params.h:95-96).
 * Outputs: means2D (N,2), depths (N), cov3D (N,6), conic_opacity (N,4), rgb (N,3),
 * clamped (N,3) u8, radii (N) i32, tiles_touched (N) u32 (tiles of the 3-sigma rectangle that
 * survive the exact tile cull of orc_common.h).  Culled Gaussians get radii = 0,
 * tiles_touched = 0 and zeros elsewhere.
 */
int orc_tile_cull_enabled = 1;
void orc_set_tile_cull(int enabled) { orc_tile_cull_enabled = enabled; }

void orc_preprocess(const orc_settings *st, int N, const real *means3D, const real *scales,
                    const real *rotations, const real *opacities, const real *sh_dc, int dc_stride,
                    const real *sh_rest, int rest_stride, const real *colors_precomp,
                    const real *cov3D_precomp, real *means2D, real *depths, real *cov3D, real *conic_opacity,
                    real *rgb, unsigned char *clamped, int *radii, uint32_t *tiles_touched) {
  const int gx = orc_grid_x(st->W), gy = orc_grid_y(st->H);
  const real focal_x = RC(st->W) / (RC(2) * st->tanfovx);
  const real focal_y = RC(st->H) / (RC(2) * st->tanfovy);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < N; ++i) {
    radii[i] = 0;
    tiles_touched[i] = 0;
    means2D[2 * i] = means2D[2 * i + 1] = 0;
    depths[i] = 0;
    for (int c = 0; c < 6; ++c) cov3D[6 * i + c] = 0;
    for (int c = 0; c < 4; ++c) conic_opacity[4 * i + c] = 0;
    for (int c = 0; c < 3; ++c) { rgb[3 * i + c] = 0; clamped[3 * i + c] = 0; }

    const real *p = means3D + 3 * i;
    real pv[3];
    xform4x3(st->viewmatrix, p, pv);
    if (!(pv[2] > ORC_NEAR_Z)) continue; /* near cull (also rejects NaN) */

    real ph[4];
    xform4x4(st->projmatrix, p, ph);
    real pw = RC(1) / (ph[3] + ORC_W_EPS);
    real ndc_x = ph[0] * pw, ndc_y = ph[1] * pw;

    real c6[6];
    if (cov3D_precomp) {
      for (int c = 0; c < 6; ++c) c6[c] = cov3D_precomp[6 * i + c];
    } else {
      orc_cov3d_one(scales + 3 * i, st->scale_modifier, rotations + 4 * i, c6);
    }

    real abc[3];
    orc_cov2d_one(pv, focal_x, focal_y, st->tanfovx, st->tanfovy, c6, st->viewmatrix, abc);
    real a = abc[0] + ORC_DILATION, b = abc[1], c = abc[2] + ORC_DILATION;
    real det = a * c - b * b;
    if (!(det > RC(0))) continue;
    real det_inv = RC(1) / det;
    real con_x = c * det_inv, con_y = -b * det_inv, con_z = a * det_inv;

    real mid = RC(0.5) * (a + c);
    real disc = R_SQRT(R_FMAX(ORC_LAMBDA_FLOOR, mid * mid - det));
    real lambda1 = mid + disc, lambda2 = mid - disc;
    real rad_f = R_CEIL(RC(3) * R_SQRT(R_FMAX(lambda1, lambda2)));
    if (!(rad_f < RC(1e9))) continue; /* overflow / NaN guard */
    int rad = (int)rad_f;

    real px = ((ndc_x + RC(1)) * RC(st->W) - RC(1)) * RC(0.5);
    real py = ((ndc_y + RC(1)) * RC(st->H) - RC(1)) * RC(0.5);
    if (!(R_FABS(px) < RC(1e9)) || !(R_FABS(py) < RC(1e9))) continue;

    int x0 = (int)((px - RC(rad)) / RC(ORC_TILE));
    int y0 = (int)((py - RC(rad)) / RC(ORC_TILE));
    int x1 = (int)((px + RC(rad) + RC(ORC_TILE - 1)) / RC(ORC_TILE));
    int y1 = (int)((py + RC(rad) + RC(ORC_TILE - 1)) / RC(ORC_TILE));
    x0 = x0 < 0 ? 0 : (x0 > gx ? gx : x0);
    x1 = x1 < 0 ? 0 : (x1 > gx ? gx : x1);
    y0 = y0 < 0 ? 0 : (y0 > gy ? gy : y0);
    y1 = y1 < 0 ? 0 : (y1 > gy ? gy : y1);
    int area = (x1 - x0) * (y1 - y0);
    if (area <= 0) continue;
    /* exact tile culling (orc_common.h): tiles_touched counts the tiles that can receive alpha >= 1/255 */
    int alive = area;
    if (orc_tile_cull_enabled && area <= ORC_CULL_MAX_TILES) {
      const orc_cull cs = orc_cull_setup(RC(0.5) * con_x, con_y, RC(0.5) * con_z, opacities[i]);
      alive = 0;
      for (int ty = y0; ty < y1; ++ty)
        for (int tx = x0; tx < x1; ++tx) alive += orc_tile_alive(&cs, px, py, tx, ty);
    }

    if (colors_precomp) {
      for (int ch = 0; ch < 3; ++ch) rgb[3 * i + ch] = colors_precomp[3 * i + ch];
    } else {
      orc_sh_to_rgb_one(st->sh_degree, p, st->campos, sh_dc + (size_t)dc_stride * i,
                        sh_rest ? sh_rest + (size_t)rest_stride * i : NULL, rgb + 3 * i, clamped + 3 * i);
    }
    for (int c2 = 0; c2 < 6; ++c2) cov3D[6 * i + c2] = c6[c2];
    depths[i] = pv[2];
    radii[i] = rad;
    means2D[2 * i] = px;
    means2D[2 * i + 1] = py;
    conic_opacity[4 * i + 0] = con_x;
    conic_opacity[4 * i + 1] = con_y;
    conic_opacity[4 * i + 2] = con_z;
    conic_opacity[4 * i + 3] = opacities[i];
    tiles_touched[i] = (uint32_t)alive;
  }
}
