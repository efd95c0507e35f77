// host_conventions.cpp — host-side (CPU, float64) mirror of the reference's camera / schedule
// helpers, exported through the C ABI so a C or C++ trainer can fill gs_settings exactly as
// the reference's Camera constructor would.  Each function cites the reference lines it
// restates.  No CUDA here.
#include <cmath>
#include <cstring>

#include "../../include/gsplat_b200.h"

namespace {

struct Mat4 {
  double m[4][4];
};

Mat4 identity() {
  Mat4 r{};
  for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0;
  return r;
}

// Gauss-Jordan with partial pivoting (the reference uses cv::Matx::inv / torch::inverse).
Mat4 inverse(const Mat4& a) {
  double aug[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      aug[i][j] = a.m[i][j];
      aug[i][4 + j] = (i == j) ? 1.0 : 0.0;
    }
  for (int col = 0; col < 4; ++col) {
    int piv = col;
    for (int r = col + 1; r < 4; ++r)
      if (std::fabs(aug[r][col]) > std::fabs(aug[piv][col])) piv = r;
    if (piv != col)
      for (int j = 0; j < 8; ++j) std::swap(aug[piv][j], aug[col][j]);
    const double d = aug[col][col];
    for (int j = 0; j < 8; ++j) aug[col][j] /= d;
    for (int r = 0; r < 4; ++r) {
      if (r == col) continue;
      const double f = aug[r][col];
      if (f != 0.0)
        for (int j = 0; j < 8; ++j) aug[r][j] -= f * aug[col][j];
    }
  }
  Mat4 out{};
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) out.m[i][j] = aug[i][4 + j];
  return out;
}

// Rt[:3,:3] = R^T, Rt[:3,3] = t — src/utils/graphics_utils.cpp:10-29
Mat4 world2view(const double R[9], const double t[3]) {
  Mat4 Rt{};
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) Rt.m[i][j] = R[3 * j + i];
    Rt.m[i][3] = t[i];
  }
  Rt.m[3][3] = 1.0;
  return Rt;
}

// recentre + rescale the camera centre — src/utils/graphics_utils.cpp:32-43
Mat4 world2view2(const double R[9], const double t[3], const double translate[3], double scale) {
  Mat4 c2w = inverse(world2view(R, t));
  for (int i = 0; i < 3; ++i) c2w.m[i][3] = (c2w.m[i][3] + translate[i]) * scale;
  return inverse(c2w);
}

// src/utils/graphics_utils.cpp:46-72 (z_sign = +1)
Mat4 projection(double znear, double zfar, double fovX, double fovY) {
  const double ty = std::tan(fovY / 2), tx = std::tan(fovX / 2);
  const double top = ty * znear, bottom = -top, right = tx * znear, left = -right;
  Mat4 P{};
  P.m[0][0] = 2.0 * znear / (right - left);
  P.m[1][1] = 2.0 * znear / (top - bottom);
  P.m[0][2] = (right + left) / (right - left);
  P.m[1][2] = (top + bottom) / (top - bottom);
  P.m[3][2] = 1.0;
  P.m[2][2] = zfar / (zfar - znear);
  P.m[2][3] = -(zfar * znear) / (zfar - znear);
  return P;
}

Mat4 transpose(const Mat4& a) {
  Mat4 r{};
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i];
  return r;
}

Mat4 matmul(const Mat4& a, const Mat4& b) {
  Mat4 r{};
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += a.m[i][k] * b.m[k][j];
      r.m[i][j] = s;
    }
  return r;
}

}  // namespace

extern "C" {

double gs_focal2fov(double focal, double pixels) { return 2.0 * std::atan(pixels / (2.0 * focal)); }

void gs_get_world2view2(const double R[9], const double t[3], const double translate[3], double scale,
                        double Rt_out[16]) {
  const Mat4 v = world2view2(R, t, translate, scale);
  std::memcpy(Rt_out, v.m, sizeof(v.m));
}

void gs_get_projection_matrix(double znear, double zfar, double fovX, double fovY, double P_out[16]) {
  const Mat4 p = projection(znear, zfar, fovX, fovY);
  std::memcpy(P_out, p.m, sizeof(p.m));
}

// Camera ctor: src/scene/camera.cpp:44-45 (znear .01, zfar 100), :66-71 (transposed tensors,
// full_proj = world_view^T-tensor bmm projection^T-tensor, camera_center = inverse(wvt)[3,:3]).
void gs_camera_settings(const double R[9], const double t[3], const double translate[3], double scale, double fovX,
                        double fovY, int H, int W, gs_settings* out) {
  const double znear = 0.01, zfar = 100.0;
  const Mat4 wvt = transpose(world2view2(R, t, translate, scale));
  const Mat4 prj = transpose(projection(znear, zfar, fovX, fovY));
  const Mat4 full = matmul(wvt, prj);
  const Mat4 inv = inverse(wvt);
  out->H = H;
  out->W = W;
  out->tanfovx = (float)std::tan(fovX * 0.5);
  out->tanfovy = (float)std::tan(fovY * 0.5);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      out->viewmatrix[4 * i + j] = (float)wvt.m[i][j];
      out->projmatrix[4 * i + j] = (float)full.m[i][j];
    }
  for (int k = 0; k < 3; ++k) out->campos[k] = (float)inv.m[3][k];
}

// src/utils/general_utils.cpp:112-142
float gs_expon_lr(float lr_init, float lr_final, int lr_delay_steps, float lr_delay_mult, int max_steps, int step) {
  if (step < 0 || (lr_init == 0.0f && lr_final == 0.0f)) return 0.0f;
  float delay_rate = 1.0f;
  if (lr_delay_steps > 0) {
    float frac = static_cast<float>(step) / lr_delay_steps;
    frac = std::fmin(std::fmax(frac, 0.0f), 1.0f);
    delay_rate = static_cast<float>(lr_delay_mult + (1 - lr_delay_mult) * std::sin(0.5 * M_PI * frac));
  }
  float u = static_cast<float>(step) / max_steps;
  u = std::fmin(std::fmax(u, 0.0f), 1.0f);
  const float log_lerp = std::exp(std::log(lr_init) * (1 - u) + std::log(lr_final) * u);
  return delay_rate * log_lerp;
}

}  // extern "C"
