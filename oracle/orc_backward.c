/*
 * oracle/orc_backward.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Stage K4b: backward of the per-Gaussian preprocess — conic -> cov2D ->
 * (cov3D, view-space mean) ; mean2D -> mean3D ; SH backward ; cov3D -> (scale, quaternion).
 * No reference counterpart.  The forward it differentiates is
 * orc_preprocess.c; the cov3D parametrisation is the reference's
 * (src/utils/general_utils.cpp:12-99, src/scene/gaussian_model.cpp:18-28), including the
 * quaternion normalisation inside build_rotation (general_utils.cpp:14-19), whose Jacobian
 * is therefore part of dL/drotation.
 *
 * Deviation from the widely used implementation, chosen because finite differences are this
 * oracle's independent check: when x/z (y/z) is clamped by the 1.3*tanfov guard, the forward
 * is J02 = -fx*lim/tz, so dJ02/dtz carries a factor 1 (not 2) and dJ02/dtx = 0.  Setting bit 1 of
 * settings.debug selects the widely used form (factor 2 always) -- the compatibility switch of the product.
 */
#include "orc_common.h"

void orc_preprocess_backward(const orc_settings *st, int N, const real *means3D, const real *scales,
                             const real *rotations, const real *sh_dc, int dc_stride, const real *sh_rest,
                             int rest_stride, const real *cov3D_precomp, int colors_are_precomp,
                             const int *radii, const unsigned char *clamped, const real *dL_dmean2D,
                             const real *dL_dconic, const real *dL_dcolor, real *dL_dmeans3D, real *dL_dscales,
                             real *dL_drotations, real *dL_dsh_dc, int gdc_stride, real *dL_dsh_rest,
                             int grest_stride, real *dL_dcov3D_out) {
  const real focal_x = RC(st->W) / (RC(2) * st->tanfovx);
  const real focal_y = RC(st->H) / (RC(2) * st->tanfovy);
  const real *V = st->viewmatrix, *M = st->projmatrix;
  const int K = (st->max_sh_degree + 1) * (st->max_sh_degree + 1);
  const int deg = st->sh_degree;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < N; ++i) {
    real gm[3] = {0, 0, 0}, gs[3] = {0, 0, 0}, gq[4] = {0, 0, 0, 0}, gcov[6] = {0, 0, 0, 0, 0, 0};
    real gsh[16][3];
    for (int k = 0; k < 16; ++k) gsh[k][0] = gsh[k][1] = gsh[k][2] = 0;
    if (radii[i] > 0) {
      const real *p = means3D + 3 * i;
      /* ---------- recompute forward quantities ---------- */
      real t[3];
      t[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
      t[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
      t[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
      real c6[6];
      real qn[4], qlen = 1, R[9], sm[3];
      if (cov3D_precomp) {
        for (int c = 0; c < 6; ++c) c6[c] = cov3D_precomp[6 * i + c];
      } else {
        const real *q = rotations + 4 * i;
        qlen = R_SQRT(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
        for (int c = 0; c < 4; ++c) qn[c] = q[c] / qlen;
        const real r = qn[0], x = qn[1], y = qn[2], z = qn[3];
        R[0] = RC(1) - RC(2) * (y * y + z * z); R[1] = RC(2) * (x * y - r * z); R[2] = RC(2) * (x * z + r * y);
        R[3] = RC(2) * (x * y + r * z); R[4] = RC(1) - RC(2) * (x * x + z * z); R[5] = RC(2) * (y * z - r * x);
        R[6] = RC(2) * (x * z - r * y); R[7] = RC(2) * (y * z + r * x); R[8] = RC(1) - RC(2) * (x * x + y * y);
        for (int c = 0; c < 3; ++c) sm[c] = st->scale_modifier * scales[3 * i + c];
        real L[9];
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) L[3 * a + b] = R[3 * a + b] * sm[b];
        c6[0] = L[0] * L[0] + L[1] * L[1] + L[2] * L[2];
        c6[1] = L[0] * L[3] + L[1] * L[4] + L[2] * L[5];
        c6[2] = L[0] * L[6] + L[1] * L[7] + L[2] * L[8];
        c6[3] = L[3] * L[3] + L[4] * L[4] + L[5] * L[5];
        c6[4] = L[3] * L[6] + L[4] * L[7] + L[5] * L[8];
        c6[5] = L[6] * L[6] + L[7] * L[7] + L[8] * L[8];
      }
      const real limx = ORC_FOV_GUARD * st->tanfovx, limy = ORC_FOV_GUARD * st->tanfovy;
      const real txtz = t[0] / t[2], tytz = t[1] / t[2];
      const int x_free = !(txtz < -limx || txtz > limx), y_free = !(tytz < -limy || tytz > limy);
      const real tx = R_FMIN(limx, R_FMAX(-limx, txtz)) * t[2];
      const real ty = R_FMIN(limy, R_FMAX(-limy, tytz)) * t[2];
      const real tz = t[2];
      const real J00 = focal_x / tz, J02 = -(focal_x * tx) / (tz * tz);
      const real J11 = focal_y / tz, J12 = -(focal_y * ty) / (tz * tz);
      const real W0[3] = {V[0], V[4], V[8]}, W1[3] = {V[1], V[5], V[9]}, W2[3] = {V[2], V[6], V[10]};
      real T0[3], T1[3];
      for (int c = 0; c < 3; ++c) {
        T0[c] = J00 * W0[c] + J02 * W2[c];
        T1[c] = J11 * W1[c] + J12 * W2[c];
      }
      const real S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
      real U0[3], U1[3];
      for (int c = 0; c < 3; ++c) {
        U0[c] = S[c][0] * T0[0] + S[c][1] * T0[1] + S[c][2] * T0[2];
        U1[c] = S[c][0] * T1[0] + S[c][1] * T1[1] + S[c][2] * T1[2];
      }
      const real a = T0[0] * U0[0] + T0[1] * U0[1] + T0[2] * U0[2] + ORC_DILATION;
      const real b = T0[0] * U1[0] + T0[1] * U1[1] + T0[2] * U1[2];
      const real c = T1[0] * U1[0] + T1[1] * U1[1] + T1[2] * U1[2] + ORC_DILATION;
      const real det = a * c - b * b;
      /* ---------- conic -> cov2D ---------- */
      const real gA = dL_dconic[3 * i], gB = dL_dconic[3 * i + 1], gC = dL_dconic[3 * i + 2];
      const real d2 = RC(1) / (det * det);
      const real ga = d2 * (-c * c * gA + b * c * gB - b * b * gC);
      const real gb = d2 * (RC(2) * b * c * gA - (a * c + b * b) * gB + RC(2) * a * b * gC);
      const real gc = d2 * (-b * b * gA + a * b * gB - a * a * gC);
      /* ---------- cov2D -> cov3D (packed) ---------- */
      gcov[0] = T0[0] * T0[0] * ga + T0[0] * T1[0] * gb + T1[0] * T1[0] * gc;
      gcov[3] = T0[1] * T0[1] * ga + T0[1] * T1[1] * gb + T1[1] * T1[1] * gc;
      gcov[5] = T0[2] * T0[2] * ga + T0[2] * T1[2] * gb + T1[2] * T1[2] * gc;
      gcov[1] = RC(2) * T0[0] * T0[1] * ga + (T0[0] * T1[1] + T0[1] * T1[0]) * gb + RC(2) * T1[0] * T1[1] * gc;
      gcov[2] = RC(2) * T0[0] * T0[2] * ga + (T0[0] * T1[2] + T0[2] * T1[0]) * gb + RC(2) * T1[0] * T1[2] * gc;
      gcov[4] = RC(2) * T0[1] * T0[2] * ga + (T0[1] * T1[2] + T0[2] * T1[1]) * gb + RC(2) * T1[1] * T1[2] * gc;
      /* ---------- cov2D -> T -> J -> view-space mean ---------- */
      real gT0[3], gT1[3];
      for (int k = 0; k < 3; ++k) {
        gT0[k] = RC(2) * ga * U0[k] + gb * U1[k];
        gT1[k] = gb * U0[k] + RC(2) * gc * U1[k];
      }
      const real gJ00 = gT0[0] * W0[0] + gT0[1] * W0[1] + gT0[2] * W0[2];
      const real gJ02 = gT0[0] * W2[0] + gT0[1] * W2[1] + gT0[2] * W2[2];
      const real gJ11 = gT1[0] * W1[0] + gT1[1] * W1[1] + gT1[2] * W1[2];
      const real gJ12 = gT1[0] * W2[0] + gT1[1] * W2[1] + gT1[2] * W2[2];
      const real tz2 = RC(1) / (tz * tz), tz3 = tz2 / tz;
      const real gtx = x_free ? -focal_x * tz2 * gJ02 : RC(0);
      const real gty = y_free ? -focal_y * tz2 * gJ12 : RC(0);
      /* debug bit 1 (GS_PUBLISHED_CLAMP_GRAD of include/gsplat_b200.h): the widely used implementation's factor 2
       * for a clamped component as well (its dJ02/dtz treats the clamped t.x as independent of t.z) */
      const int published = (st->debug & 2) != 0;
      const real kx = (x_free || published) ? RC(2) : RC(1), ky = (y_free || published) ? RC(2) : RC(1);
      const real gtz = -focal_x * tz2 * gJ00 - focal_y * tz2 * gJ11 + kx * focal_x * tx * tz3 * gJ02 +
                       ky * focal_y * ty * tz3 * gJ12;
      gm[0] = W0[0] * gtx + W1[0] * gty + W2[0] * gtz;
      gm[1] = W0[1] * gtx + W1[1] * gty + W2[1] * gtz;
      gm[2] = W0[2] * gtx + W1[2] * gty + W2[2] * gtz;
      /* ---------- mean2D -> mean3D through the projection ---------- */
      real ph[4];
      for (int r = 0; r < 4; ++r) ph[r] = M[r] * p[0] + M[4 + r] * p[1] + M[8 + r] * p[2] + M[12 + r];
      const real pw = RC(1) / (ph[3] + ORC_W_EPS);
      const real g2x = dL_dmean2D[2 * i], g2y = dL_dmean2D[2 * i + 1];
      for (int j = 0; j < 3; ++j) {
        const real dndcx = M[4 * j + 0] * pw - M[4 * j + 3] * ph[0] * pw * pw;
        const real dndcy = M[4 * j + 1] * pw - M[4 * j + 3] * ph[1] * pw * pw;
        gm[j] += g2x * dndcx + g2y * dndcy;
      }
      /* ---------- SH backward ---------- */
      if (!colors_are_precomp) {
        real gr[3];
        for (int ch = 0; ch < 3; ++ch) gr[ch] = clamped[3 * i + ch] ? RC(0) : dL_dcolor[3 * i + ch];
        const real *dc = sh_dc + (size_t)dc_stride * i;
        const real *s = sh_rest ? sh_rest + (size_t)rest_stride * i : NULL;
        (void)dc;
        real dvx = p[0] - st->campos[0], dvy = p[1] - st->campos[1], dvz = p[2] - st->campos[2];
        const real len = R_SQRT(dvx * dvx + dvy * dvy + dvz * dvz);
        const real x = dvx / len, y = dvy / len, z = dvz / len;
        real basis[16];
        for (int k = 0; k < 16; ++k) basis[k] = 0;
        real ddx = 0, ddy = 0, ddz = 0; /* dL/d(normalised dir) */
        basis[0] = ORC_SH_C0;
        if (deg > 0) {
          basis[1] = -ORC_SH_C1 * y; basis[2] = ORC_SH_C1 * z; basis[3] = -ORC_SH_C1 * x;
          for (int ch = 0; ch < 3; ++ch) {
            ddx += -ORC_SH_C1 * s[6 + ch] * gr[ch];
            ddy += -ORC_SH_C1 * s[0 + ch] * gr[ch];
            ddz += ORC_SH_C1 * s[3 + ch] * gr[ch];
          }
          if (deg > 1) {
            const real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            const real c20 = RC(ORC_SH_C2[0]), c21 = RC(ORC_SH_C2[1]), c22 = RC(ORC_SH_C2[2]),
                       c23 = RC(ORC_SH_C2[3]), c24 = RC(ORC_SH_C2[4]);
            basis[4] = c20 * xy; basis[5] = c21 * yz; basis[6] = c22 * (RC(2) * zz - xx - yy);
            basis[7] = c23 * xz; basis[8] = c24 * (xx - yy);
            for (int ch = 0; ch < 3; ++ch) {
              const real s4 = s[9 + ch], s5 = s[12 + ch], s6 = s[15 + ch], s7 = s[18 + ch], s8 = s[21 + ch];
              ddx += (c20 * y * s4 + c22 * RC(-2) * x * s6 + c23 * z * s7 + c24 * RC(2) * x * s8) * gr[ch];
              ddy += (c20 * x * s4 + c21 * z * s5 + c22 * RC(-2) * y * s6 + c24 * RC(-2) * y * s8) * gr[ch];
              ddz += (c21 * y * s5 + c22 * RC(4) * z * s6 + c23 * x * s7) * gr[ch];
            }
            if (deg > 2) {
              const real c30 = RC(ORC_SH_C3[0]), c31 = RC(ORC_SH_C3[1]), c32 = RC(ORC_SH_C3[2]),
                         c33 = RC(ORC_SH_C3[3]), c34 = RC(ORC_SH_C3[4]), c35 = RC(ORC_SH_C3[5]),
                         c36 = RC(ORC_SH_C3[6]);
              basis[9] = c30 * y * (RC(3) * xx - yy);
              basis[10] = c31 * xy * z;
              basis[11] = c32 * y * (RC(4) * zz - xx - yy);
              basis[12] = c33 * z * (RC(2) * zz - RC(3) * xx - RC(3) * yy);
              basis[13] = c34 * x * (RC(4) * zz - xx - yy);
              basis[14] = c35 * z * (xx - yy);
              basis[15] = c36 * x * (xx - RC(3) * yy);
              for (int ch = 0; ch < 3; ++ch) {
                const real s9 = s[24 + ch], s10 = s[27 + ch], s11 = s[30 + ch], s12 = s[33 + ch],
                           s13 = s[36 + ch], s14 = s[39 + ch], s15 = s[42 + ch];
                ddx += (c30 * RC(6) * xy * s9 + c31 * yz * s10 + c32 * RC(-2) * xy * s11 +
                        c33 * RC(-6) * xz * s12 + c34 * (RC(4) * zz - RC(3) * xx - yy) * s13 +
                        c35 * RC(2) * xz * s14 + c36 * (RC(3) * xx - RC(3) * yy) * s15) * gr[ch];
                ddy += (c30 * (RC(3) * xx - RC(3) * yy) * s9 + c31 * xz * s10 +
                        c32 * (RC(4) * zz - xx - RC(3) * yy) * s11 + c33 * RC(-6) * yz * s12 +
                        c34 * RC(-2) * xy * s13 + c35 * RC(-2) * yz * s14 + c36 * RC(-6) * xy * s15) * gr[ch];
                ddz += (c31 * xy * s10 + c32 * RC(8) * yz * s11 +
                        c33 * (RC(6) * zz - RC(3) * xx - RC(3) * yy) * s12 + c34 * RC(8) * xz * s13 +
                        c35 * (xx - yy) * s14) * gr[ch];
              }
            }
          }
        }
        const int nact = (deg + 1) * (deg + 1);
        for (int k = 0; k < nact; ++k)
          for (int ch = 0; ch < 3; ++ch) gsh[k][ch] = basis[k] * gr[ch];
        /* through the normalisation of the view direction */
        const real dot = x * ddx + y * ddy + z * ddz;
        gm[0] += (ddx - x * dot) / len;
        gm[1] += (ddy - y * dot) / len;
        gm[2] += (ddz - z * dot) / len;
      }
      /* ---------- cov3D -> scale, quaternion ---------- */
      if (!cov3D_precomp) {
        /* full symmetric gradient, dL/dL = 2 * GS * L */
        const real GS[3][3] = {{gcov[0], RC(0.5) * gcov[1], RC(0.5) * gcov[2]},
                               {RC(0.5) * gcov[1], gcov[3], RC(0.5) * gcov[4]},
                               {RC(0.5) * gcov[2], RC(0.5) * gcov[4], gcov[5]}};
        real gL[9], gR[9];
        for (int a2 = 0; a2 < 3; ++a2)
          for (int b2 = 0; b2 < 3; ++b2) {
            real acc = 0;
            for (int k = 0; k < 3; ++k) acc += GS[a2][k] * (R[3 * k + b2] * sm[b2]);
            gL[3 * a2 + b2] = RC(2) * acc;
          }
        for (int b2 = 0; b2 < 3; ++b2) {
          real acc = 0;
          for (int a2 = 0; a2 < 3; ++a2) acc += gL[3 * a2 + b2] * R[3 * a2 + b2];
          gs[b2] = st->scale_modifier * acc;
        }
        for (int a2 = 0; a2 < 3; ++a2)
          for (int b2 = 0; b2 < 3; ++b2) gR[3 * a2 + b2] = gL[3 * a2 + b2] * sm[b2];
        const real r = qn[0], x = qn[1], y = qn[2], z = qn[3];
        real gn[4];
        gn[0] = RC(2) * (-z * gR[1] + y * gR[2] + z * gR[3] - x * gR[5] - y * gR[6] + x * gR[7]);
        gn[1] = RC(2) * (y * gR[1] + z * gR[2] + y * gR[3] - RC(2) * x * gR[4] - r * gR[5] + z * gR[6] +
                         r * gR[7] - RC(2) * x * gR[8]);
        gn[2] = RC(2) * (-RC(2) * y * gR[0] + x * gR[1] + r * gR[2] + x * gR[3] + z * gR[5] - r * gR[6] +
                         z * gR[7] - RC(2) * y * gR[8]);
        gn[3] = RC(2) * (-RC(2) * z * gR[0] - r * gR[1] + x * gR[2] + r * gR[3] - RC(2) * z * gR[4] +
                         y * gR[5] + x * gR[6] + y * gR[7]);
        const real dotq = qn[0] * gn[0] + qn[1] * gn[1] + qn[2] * gn[2] + qn[3] * gn[3];
        for (int k = 0; k < 4; ++k) gq[k] = (gn[k] - qn[k] * dotq) / qlen;
      }
    }
    for (int c = 0; c < 3; ++c) dL_dmeans3D[3 * i + c] = gm[c];
    if (dL_dscales)
      for (int c = 0; c < 3; ++c) dL_dscales[3 * i + c] = gs[c];
    if (dL_drotations)
      for (int c = 0; c < 4; ++c) dL_drotations[4 * i + c] = gq[c];
    if (dL_dcov3D_out)
      for (int c = 0; c < 6; ++c) dL_dcov3D_out[6 * i + c] = gcov[c];
    if (dL_dsh_dc) {
      for (int ch = 0; ch < 3; ++ch) dL_dsh_dc[(size_t)gdc_stride * i + ch] = gsh[0][ch];
      if (dL_dsh_rest)
        for (int k = 1; k < K; ++k)
          for (int ch = 0; ch < 3; ++ch) dL_dsh_rest[(size_t)grest_stride * i + 3 * (k - 1) + ch] = gsh[k][ch];
    }
  }
}
