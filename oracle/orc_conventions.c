/*
 * oracle/orc_conventions.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Restatement of the conventions the reference DOES fix; each function cites the
 * reference lines it follows and is pinned by the reference's own known-answer tests.
 */
#include "orc_common.h"

/* quaternion (r,x,y,z) -> row-major 3x3 R; normalises q itself.
 * Follows src/utils/general_utils.cpp:12-40. */
void orc_build_rotation(int N, const real *q_in, real *R_out) {
  for (int i = 0; i < N; ++i) {
    const real *q = q_in + 4 * i;
    real *R = R_out + 9 * i;
    real norm = R_SQRT(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    real r = q[0] / norm, x = q[1] / norm, y = q[2] / norm, z = q[3] / norm;
    R[0] = RC(1) - RC(2) * (y * y + z * z);
    R[1] = RC(2) * (x * y - r * z);
    R[2] = RC(2) * (x * z + r * y);
    R[3] = RC(2) * (x * y + r * z);
    R[4] = RC(1) - RC(2) * (x * x + z * z);
    R[5] = RC(2) * (y * z - r * x);
    R[6] = RC(2) * (x * z - r * y);
    R[7] = RC(2) * (y * z + r * x);
    R[8] = RC(1) - RC(2) * (x * x + y * y);
  }
}

/* L = R * diag(s).  Follows src/utils/general_utils.cpp:88-99. */
void orc_build_scaling_rotation(int N, const real *s, const real *q, real *L_out) {
  for (int i = 0; i < N; ++i) {
    real R[9];
    orc_build_rotation(1, q + 4 * i, R);
    real *L = L_out + 9 * i;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) L[3 * r + c] = R[3 * r + c] * s[3 * i + c];
  }
}

/* (N,3,3) -> (N,6) in the order [00,01,02,11,12,22].
 * Follows src/utils/general_utils.cpp:49-62 (strip_lowerdiag) / :73 (strip_symmetric). */
void orc_strip_symmetric(int N, const real *M, real *out) {
  for (int i = 0; i < N; ++i) {
    const real *m = M + 9 * i;
    real *o = out + 6 * i;
    o[0] = m[0]; o[1] = m[1]; o[2] = m[2]; o[3] = m[4]; o[4] = m[5]; o[5] = m[8];
  }
}

/* Sigma = L L^T with L = R diag(mod*s), packed (N,6).
 * Follows src/scene/gaussian_model.cpp:18-28 (build_covariance_from_scaling_rotation)
 * as called by get_covariance (:301-304).  The scalar per-Gaussian routine used by
 * orc_preprocess is orc_cov3d_one below; this is the batched form the reference tests pin. */
void orc_cov3d_one(const real *s, real mod, const real *q, real *cov6) {
  real norm = R_SQRT(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  real r = q[0] / norm, x = q[1] / norm, y = q[2] / norm, z = q[3] / norm;
  real R00 = RC(1) - RC(2) * (y * y + z * z);
  real R01 = RC(2) * (x * y - r * z);
  real R02 = RC(2) * (x * z + r * y);
  real R10 = RC(2) * (x * y + r * z);
  real R11 = RC(1) - RC(2) * (x * x + z * z);
  real R12 = RC(2) * (y * z - r * x);
  real R20 = RC(2) * (x * z - r * y);
  real R21 = RC(2) * (y * z + r * x);
  real R22 = RC(1) - RC(2) * (x * x + y * y);
  real s0 = mod * s[0], s1 = mod * s[1], s2 = mod * s[2];
  real L00 = R00 * s0, L01 = R01 * s1, L02 = R02 * s2;
  real L10 = R10 * s0, L11 = R11 * s1, L12 = R12 * s2;
  real L20 = R20 * s0, L21 = R21 * s1, L22 = R22 * s2;
  cov6[0] = L00 * L00 + L01 * L01 + L02 * L02;
  cov6[1] = L00 * L10 + L01 * L11 + L02 * L12;
  cov6[2] = L00 * L20 + L01 * L21 + L02 * L22;
  cov6[3] = L10 * L10 + L11 * L11 + L12 * L12;
  cov6[4] = L10 * L20 + L11 * L21 + L12 * L22;
  cov6[5] = L20 * L20 + L21 * L21 + L22 * L22;
}

void orc_covariance(int N, const real *scales, real mod, const real *q, real *cov6) {
  for (int i = 0; i < N; ++i) orc_cov3d_one(scales + 3 * i, mod, q + 4 * i, cov6 + 6 * i);
}

/* Activations of GaussianModel.  Follows src/scene/gaussian_model.cpp:54-59 and the
 * getters :270-298: scaling = exp, rotation = L2 normalise, opacity = sigmoid. */
void orc_activate(int N, const real *log_scales, const real *raw_rot, const real *logit_opacity,
                  real *scales, real *rot, real *opacity) {
  for (int i = 0; i < N; ++i) {
    for (int c = 0; c < 3; ++c) scales[3 * i + c] = R_EXP(log_scales[3 * i + c]);
    const real *q = raw_rot + 4 * i;
    real n = R_SQRT(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    n = R_FMAX(n, RC(1e-12)); /* torch normalize eps */
    for (int c = 0; c < 4; ++c) rot[4 * i + c] = q[c] / n;
    opacity[i] = RC(1) / (RC(1) + R_EXP(-logit_opacity[i]));
  }
}

/* The same activations as the RAW-parameter path of the kernels evaluates them (raw_params = 1): scaling and
 * opacity through the specified exp (orc_common.h orc_exp_any / orc_sigmoid); the quaternion is handed on
 * un-normalised, because the covariance construction normalises it itself (src/utils/general_utils.cpp:14-19),
 * exactly once -- as K1 does. */
void orc_activate_raw(int N, const real *log_scales, const real *raw_rot, const real *logit_opacity,
                      real *scales, real *rot, real *opacity) {
  for (int i = 0; i < N; ++i) {
    for (int c = 0; c < 3; ++c) scales[3 * i + c] = orc_exp_any(log_scales[3 * i + c]);
    for (int c = 0; c < 4; ++c) rot[4 * i + c] = raw_rot[4 * i + c];
    opacity[i] = orc_sigmoid(logit_opacity[i]);
  }
}

/* ---- camera conventions (float64 on the host, as in the reference) ---- */

/* focal2fov.  Follows src/utils/graphics_utils.cpp:4-7. */
double orc_focal2fov(double focal, double pixels) { return 2.0 * atan(pixels / (2.0 * focal)); }

static void mat4_inverse_rigid_general(const double *m, double *inv) {
  /* general 4x4 inverse by cofactors (row-major) */
  double a[16];
  memcpy(a, m, sizeof(a));
  double t[16];
  t[0] = a[5] * a[10] * a[15] - a[5] * a[11] * a[14] - a[9] * a[6] * a[15] + a[9] * a[7] * a[14] + a[13] * a[6] * a[11] - a[13] * a[7] * a[10];
  t[4] = -a[4] * a[10] * a[15] + a[4] * a[11] * a[14] + a[8] * a[6] * a[15] - a[8] * a[7] * a[14] - a[12] * a[6] * a[11] + a[12] * a[7] * a[10];
  t[8] = a[4] * a[9] * a[15] - a[4] * a[11] * a[13] - a[8] * a[5] * a[15] + a[8] * a[7] * a[13] + a[12] * a[5] * a[11] - a[12] * a[7] * a[9];
  t[12] = -a[4] * a[9] * a[14] + a[4] * a[10] * a[13] + a[8] * a[5] * a[14] - a[8] * a[6] * a[13] - a[12] * a[5] * a[10] + a[12] * a[6] * a[9];
  t[1] = -a[1] * a[10] * a[15] + a[1] * a[11] * a[14] + a[9] * a[2] * a[15] - a[9] * a[3] * a[14] - a[13] * a[2] * a[11] + a[13] * a[3] * a[10];
  t[5] = a[0] * a[10] * a[15] - a[0] * a[11] * a[14] - a[8] * a[2] * a[15] + a[8] * a[3] * a[14] + a[12] * a[2] * a[11] - a[12] * a[3] * a[10];
  t[9] = -a[0] * a[9] * a[15] + a[0] * a[11] * a[13] + a[8] * a[1] * a[15] - a[8] * a[3] * a[13] - a[12] * a[1] * a[11] + a[12] * a[3] * a[9];
  t[13] = a[0] * a[9] * a[14] - a[0] * a[10] * a[13] - a[8] * a[1] * a[14] + a[8] * a[2] * a[13] + a[12] * a[1] * a[10] - a[12] * a[2] * a[9];
  t[2] = a[1] * a[6] * a[15] - a[1] * a[7] * a[14] - a[5] * a[2] * a[15] + a[5] * a[3] * a[14] + a[13] * a[2] * a[7] - a[13] * a[3] * a[6];
  t[6] = -a[0] * a[6] * a[15] + a[0] * a[7] * a[14] + a[4] * a[2] * a[15] - a[4] * a[3] * a[14] - a[12] * a[2] * a[7] + a[12] * a[3] * a[6];
  t[10] = a[0] * a[5] * a[15] - a[0] * a[7] * a[13] - a[4] * a[1] * a[15] + a[4] * a[3] * a[13] + a[12] * a[1] * a[7] - a[12] * a[3] * a[5];
  t[14] = -a[0] * a[5] * a[14] + a[0] * a[6] * a[13] + a[4] * a[1] * a[14] - a[4] * a[2] * a[13] - a[12] * a[1] * a[6] + a[12] * a[2] * a[5];
  t[3] = -a[1] * a[6] * a[11] + a[1] * a[7] * a[10] + a[5] * a[2] * a[11] - a[5] * a[3] * a[10] - a[9] * a[2] * a[7] + a[9] * a[3] * a[6];
  t[7] = a[0] * a[6] * a[11] - a[0] * a[7] * a[10] - a[4] * a[2] * a[11] + a[4] * a[3] * a[10] + a[8] * a[2] * a[7] - a[8] * a[3] * a[6];
  t[11] = -a[0] * a[5] * a[11] + a[0] * a[7] * a[9] + a[4] * a[1] * a[11] - a[4] * a[3] * a[9] - a[8] * a[1] * a[7] + a[8] * a[3] * a[5];
  t[15] = a[0] * a[5] * a[10] - a[0] * a[6] * a[9] - a[4] * a[1] * a[10] + a[4] * a[2] * a[9] + a[8] * a[1] * a[6] - a[8] * a[2] * a[5];
  double det = a[0] * t[0] + a[1] * t[4] + a[2] * t[8] + a[3] * t[12];
  for (int i = 0; i < 16; ++i) inv[i] = t[i] / det;
}

/* Rt[:3,:3] = R^T, Rt[:3,3] = t, Rt[3,3] = 1 (row-major 4x4).
 * Follows src/utils/graphics_utils.cpp:10-29. */
void orc_get_world2view(const double *R33, const double *t3, double *Rt) {
  memset(Rt, 0, 16 * sizeof(double));
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Rt[4 * i + j] = R33[3 * j + i];
  for (int i = 0; i < 3; ++i) Rt[4 * i + 3] = t3[i];
  Rt[15] = 1.0;
}

/* Recentred/scaled variant.  Follows src/utils/graphics_utils.cpp:32-43. */
void orc_get_world2view_2(const double *R33, const double *t3, const double *translate3, double scale,
                          double *Rt) {
  double W2V[16], C2W[16];
  orc_get_world2view(R33, t3, W2V);
  mat4_inverse_rigid_general(W2V, C2W);
  for (int i = 0; i < 3; ++i) C2W[4 * i + 3] = (C2W[4 * i + 3] + translate3[i]) * scale;
  mat4_inverse_rigid_general(C2W, Rt);
}

/* Row-major 4x4 projection.  Follows src/utils/graphics_utils.cpp:46-72. */
void orc_get_projection_matrix(double znear, double zfar, double fovX, double fovY, double *P) {
  double tanHalfFovY = tan(fovY / 2), tanHalfFovX = tan(fovX / 2);
  double top = tanHalfFovY * znear, bottom = -top;
  double right = tanHalfFovX * znear, left = -right;
  memset(P, 0, 16 * sizeof(double));
  double z_sign = 1.0;
  P[0] = 2.0 * znear / (right - left);
  P[5] = 2.0 * znear / (top - bottom);
  P[2] = (right + left) / (right - left);
  P[6] = (top + bottom) / (top - bottom);
  P[14] = z_sign;
  P[10] = z_sign * zfar / (zfar - znear);
  P[11] = -(zfar * znear) / (zfar - znear);
}

/* Camera tensors as the Camera ctor stores them.  Follows src/scene/camera.cpp:66-71:
 *   world_view_transform = W2V^T, projection_matrix = P^T,
 *   full_proj_transform  = W2V^T * P^T  (= (P*W2V)^T),
 *   camera_center        = row 3 of inverse(world_view_transform).
 * Outputs are the FLAT buffers of those row-major tensors (so element [4*c+r] of the flat
 * buffer is V[r][c]: a column-major view of V and of P*V).  znear .01 / zfar 100 are the
 * ctor's constants (camera.cpp:44-45) and are passed in by the caller. */
void orc_camera_tensors(const double *R33, const double *t3, const double *translate3, double scale,
                        double znear, double zfar, double fovX, double fovY, double *world_view_T,
                        double *proj_T, double *full_proj_T, double *campos3) {
  double V[16], P[16];
  orc_get_world2view_2(R33, t3, translate3, scale, V);
  orc_get_projection_matrix(znear, zfar, fovX, fovY, P);
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) {
      world_view_T[4 * r + c] = V[4 * c + r];
      proj_T[4 * r + c] = P[4 * c + r];
    }
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += world_view_T[4 * r + k] * proj_T[4 * k + c];
      full_proj_T[4 * r + c] = s;
    }
  double inv[16];
  mat4_inverse_rigid_general(world_view_T, inv);
  campos3[0] = inv[12]; campos3[1] = inv[13]; campos3[2] = inv[14];
}

/* Exponential LR schedule.  Follows src/utils/general_utils.cpp:112-142 (float arithmetic,
 * double only inside sin as in the reference's std::sin(0.5 * M_PI * ...)). */
float orc_expon_lr(float lr_init, float lr_final, int lr_delay_steps, float lr_delay_mult, int max_steps,
                   int step) {
  if (step < 0 || (lr_init == 0.0f && lr_final == 0.0f)) return 0.0f;
  float delay_rate;
  if (lr_delay_steps > 0) {
    float c = (float)step / lr_delay_steps;
    c = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
    delay_rate = (float)(lr_delay_mult + (1 - lr_delay_mult) * sin(0.5 * M_PI * c));
  } else {
    delay_rate = 1.0f;
  }
  float t = (float)step / max_steps;
  t = t < 0.0f ? 0.0f : (t > 1.0f ? 1.0f : t);
  float log_lerp = expf(logf(lr_init) * (1 - t) + logf(lr_final) * t);
  return delay_rate * log_lerp;
}

/* qvec (w,x,y,z, NOT normalised) -> R, the COLMAP convention.
 * Follows src/scene/colmap_loader.cpp:265-279 (pinned by its integer known-answer test). */
void orc_qvec2rotmat(const double *q, double *R) {
  R[0] = 1 - 2 * q[2] * q[2] - 2 * q[3] * q[3];
  R[1] = 2 * q[1] * q[2] - 2 * q[0] * q[3];
  R[2] = 2 * q[3] * q[1] + 2 * q[0] * q[2];
  R[3] = 2 * q[1] * q[2] + 2 * q[0] * q[3];
  R[4] = 1 - 2 * q[1] * q[1] - 2 * q[3] * q[3];
  R[5] = 2 * q[2] * q[3] - 2 * q[0] * q[1];
  R[6] = 2 * q[3] * q[1] - 2 * q[0] * q[2];
  R[7] = 2 * q[2] * q[3] + 2 * q[0] * q[1];
  R[8] = 1 - 2 * q[1] * q[1] - 2 * q[2] * q[2];
}
