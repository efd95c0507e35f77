// preprocess_bwd.cu — K4b: backward of the per-Gaussian preprocess.
// conic -> cov2D -> (cov3D, view-space mean); mean2D -> mean3D; SH backward; cov3D ->
// (scale, quaternion) through the reference's parametrisation (q normalised inside,
// src/utils/general_utils.cpp:14-19; L = R diag(s), Sigma = L L^T, gaussian_model.cpp:18-28).
// One thread per Gaussian.  Two kernels share the math (pb_compute):
//  * preprocess_backward_bulk_kernel — the measured path: every per-Gaussian input slice of a
//    128-row block (xyz, scales, rotations, opacities, the 48-B blend-gradient records, the 180-B
//    SH rows) arrives by TMA bulk copy on one mbarrier while the threads already work, the
//    gradients are written in place into the same shared-memory slices and leave as bulk stores
//    (or bulk fp32 reduce-adds when accumulating across views).  No strided LDG/STG, no index math.
//  * preprocess_backward_kernel — generic fallback (precomputed cov3D / colours, strided or
//    unaligned tensors, ragged tail): SH rows staged through shared memory with coalesced loops.
// With raw_params the activation Jacobians (exp, sigmoid, normalise) are applied here too.
#include <stdlib.h>

#include "gs_common.cuh"
#include "tma_bulk.cuh"

#define PB_THREADS 256

struct PreBwdArgs {
  gs_gaussians in;
  gs_gradients out;
  int N;
  const int* radii;
  const uint8_t* clamped;
  const float4* grad_geom;
  float focal_x, focal_y;
  int rest_width_act;  // 3*((deg+1)^2-1)
  int rest_width_max;  // 3*((maxdeg+1)^2-1)
  int pitch;           // odd, >= rest_width_max
};

struct PbOut {
  float gm[3], gs[3], gq[4], gcov[6], gdc[3], gcol[3], gr[3], basis[15];
  float g2x, g2y, gop;
};

__device__ __forceinline__ void pb_zero(PbOut& o) {
#pragma unroll
  for (int k = 0; k < 3; ++k) o.gm[k] = o.gs[k] = o.gdc[k] = o.gcol[k] = o.gr[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) o.gq[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 6; ++k) o.gcov[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 15; ++k) o.basis[k] = 0.f;
  o.g2x = o.g2y = o.gop = 0.f;
}

// Backward of one visible Gaussian.  p: mean; q4/scl/opac_in: rotation, scale, opacity as stored by the
// caller (raw when a.in.raw_params); cov_pre: precomputed cov3D row or nullptr; G0/G1/G2x: the blend
// gradient record; cm: colour clamp mask; srow: this Gaussian's staged SH "rest" row.
__device__ __forceinline__ void pb_compute(const gs_settings& st, const PreBwdArgs& a, const float p[3], const float4 q4,
                                           const float scl[3], const float opac_in, const float* cov_pre,
                                           const float4 G0, const float4 G1, const float G2x, const uint32_t cm,
                                           const float* srow, const bool use_sh, PbOut& o) {
  float* gm = o.gm;
  float* gs = o.gs;
  float* gq = o.gq;
  float* gcov = o.gcov;
  float* gdc = o.gdc;
  float* gcol = o.gcol;
  float* gr = o.gr;
  float* basis = o.basis;
  {
    o.g2x = G0.x;
    o.g2y = G0.y;
    const float g2x = G0.x, g2y = G0.y;
    const float gA = G0.z, gB = G0.w, gC = G1.x;
    float gop = G1.y;
    gcol[0] = G1.z;
    gcol[1] = G1.w;
    gcol[2] = G2x;
    const float* V = st.viewmatrix;
    const float* M = st.projmatrix;
    float t[3];
    t[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
    t[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
    t[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
    // ---- recompute cov3D ----
    float c6[6], qn[4] = {1.f, 0.f, 0.f, 0.f}, qlen = 1.f, R[9], sm[3] = {0.f, 0.f, 0.f}, sact[3] = {0.f, 0.f, 0.f};
    const bool own_cov = cov_pre == nullptr;
    if (!own_cov) {
#pragma unroll
      for (int c = 0; c < 6; ++c) c6[c] = cov_pre[c];
    } else {
      const float q[4] = {q4.x, q4.y, q4.z, q4.w};
      qlen = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
      for (int c = 0; c < 4; ++c) qn[c] = q[c] / qlen;
      const float r = qn[0], x = qn[1], y = qn[2], z = qn[3];
      R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
      R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
      R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float sv = scl[c];
        if (a.in.raw_params) sv = gs_exp_any(sv);
        sact[c] = sv;
        sm[c] = st.scale_modifier * sv;
      }
      float L[9];
#pragma unroll
      for (int r2 = 0; r2 < 3; ++r2)
#pragma unroll
        for (int c2 = 0; c2 < 3; ++c2) L[3 * r2 + c2] = R[3 * r2 + c2] * sm[c2];
      c6[0] = L[0] * L[0] + L[1] * L[1] + L[2] * L[2];
      c6[1] = L[0] * L[3] + L[1] * L[4] + L[2] * L[5];
      c6[2] = L[0] * L[6] + L[1] * L[7] + L[2] * L[8];
      c6[3] = L[3] * L[3] + L[4] * L[4] + L[5] * L[5];
      c6[4] = L[3] * L[6] + L[4] * L[7] + L[5] * L[8];
      c6[5] = L[6] * L[6] + L[7] * L[7] + L[8] * L[8];
    }
    // ---- recompute EWA ----
    const float limx = GS_FOV_GUARD * st.tanfovx, limy = GS_FOV_GUARD * st.tanfovy;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    const bool x_free = !(txtz < -limx || txtz > limx), y_free = !(tytz < -limy || tytz > limy);
    const float tx = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    const float ty = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    const float tz = t[2];
    const float fx = a.focal_x, fy = a.focal_y;
    const float J00 = fx / tz, J02 = -(fx * tx) / (tz * tz);
    const float J11 = fy / tz, J12 = -(fy * ty) / (tz * tz);
    const float W0[3] = {V[0], V[4], V[8]}, W1[3] = {V[1], V[5], V[9]}, W2[3] = {V[2], V[6], V[10]};
    float T0[3], T1[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      T0[c] = J00 * W0[c] + J02 * W2[c];
      T1[c] = J11 * W1[c] + J12 * W2[c];
    }
    const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    float U0[3], U1[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      U0[c] = S[c][0] * T0[0] + S[c][1] * T0[1] + S[c][2] * T0[2];
      U1[c] = S[c][0] * T1[0] + S[c][1] * T1[1] + S[c][2] * T1[2];
    }
    const float ca = T0[0] * U0[0] + T0[1] * U0[1] + T0[2] * U0[2] + GS_DILATION;
    const float cb = T0[0] * U1[0] + T0[1] * U1[1] + T0[2] * U1[2];
    const float cc = T1[0] * U1[0] + T1[1] * U1[1] + T1[2] * U1[2] + GS_DILATION;
    const float det = ca * cc - cb * cb;
    // ---- conic -> cov2D ----
    const float d2 = 1.0f / (det * det);
    const float ga = d2 * (-cc * cc * gA + cb * cc * gB - cb * cb * gC);
    const float gb = d2 * (2.f * cb * cc * gA - (ca * cc + cb * cb) * gB + 2.f * ca * cb * gC);
    const float gc = d2 * (-cb * cb * gA + ca * cb * gB - ca * ca * gC);
    // ---- cov2D -> cov3D ----
    gcov[0] = T0[0] * T0[0] * ga + T0[0] * T1[0] * gb + T1[0] * T1[0] * gc;
    gcov[3] = T0[1] * T0[1] * ga + T0[1] * T1[1] * gb + T1[1] * T1[1] * gc;
    gcov[5] = T0[2] * T0[2] * ga + T0[2] * T1[2] * gb + T1[2] * T1[2] * gc;
    gcov[1] = 2.f * T0[0] * T0[1] * ga + (T0[0] * T1[1] + T0[1] * T1[0]) * gb + 2.f * T1[0] * T1[1] * gc;
    gcov[2] = 2.f * T0[0] * T0[2] * ga + (T0[0] * T1[2] + T0[2] * T1[0]) * gb + 2.f * T1[0] * T1[2] * gc;
    gcov[4] = 2.f * T0[1] * T0[2] * ga + (T0[1] * T1[2] + T0[2] * T1[1]) * gb + 2.f * T1[1] * T1[2] * gc;
    // ---- cov2D -> T -> J -> t ----
    float gT0[3], gT1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      gT0[k] = 2.f * ga * U0[k] + gb * U1[k];
      gT1[k] = gb * U0[k] + 2.f * gc * U1[k];
    }
    const float gJ00 = gT0[0] * W0[0] + gT0[1] * W0[1] + gT0[2] * W0[2];
    const float gJ02 = gT0[0] * W2[0] + gT0[1] * W2[1] + gT0[2] * W2[2];
    const float gJ11 = gT1[0] * W1[0] + gT1[1] * W1[1] + gT1[2] * W1[2];
    const float gJ12 = gT1[0] * W2[0] + gT1[1] * W2[1] + gT1[2] * W2[2];
    const float tz2 = 1.0f / (tz * tz), tz3 = tz2 / tz;
    const float gtx = x_free ? -fx * tz2 * gJ02 : 0.f;
    const float gty = y_free ? -fy * tz2 * gJ12 : 0.f;
    // GS_PUBLISHED_CLAMP_GRAD (settings.debug bit 1): the widely used implementation's factor 2 for a clamped
    // component too; default: the true derivative of the clamped forward (factor 1), which finite differences confirm
    const bool published = (st.debug & GS_PUBLISHED_CLAMP_GRAD) != 0;
    const float kx = (x_free || published) ? 2.f : 1.f, ky = (y_free || published) ? 2.f : 1.f;
    const float gtz = -fx * tz2 * gJ00 - fy * tz2 * gJ11 + kx * fx * tx * tz3 * gJ02 + ky * fy * ty * tz3 * gJ12;
#pragma unroll
    for (int c = 0; c < 3; ++c) gm[c] = W0[c] * gtx + W1[c] * gty + W2[c] * gtz;
    // ---- mean2D -> mean3D ----
    float ph[4];
#pragma unroll
    for (int r2 = 0; r2 < 4; ++r2) ph[r2] = M[r2] * p[0] + M[4 + r2] * p[1] + M[8 + r2] * p[2] + M[12 + r2];
    const float pw = 1.0f / (ph[3] + GS_W_EPS);
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const float dndcx = M[4 * j + 0] * pw - M[4 * j + 3] * ph[0] * pw * pw;
      const float dndcy = M[4 * j + 1] * pw - M[4 * j + 3] * ph[1] * pw * pw;
      gm[j] += g2x * dndcx + g2y * dndcy;
    }
    // ---- SH backward (row staged in smem is overwritten by its own gradient) ----
    if (use_sh) {
      const int deg = st.sh_degree;
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) gr[ch] = ((cm >> ch) & 1u) ? 0.f : gcol[ch];
      const float dvx = p[0] - st.campos[0], dvy = p[1] - st.campos[1], dvz = p[2] - st.campos[2];
      const float len = sqrtf(dvx * dvx + dvy * dvy + dvz * dvz);
      const float x = dvx / len, y = dvy / len, z = dvz / len;
      float ddx = 0.f, ddy = 0.f, ddz = 0.f;
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) gdc[ch] = GS_SH_C0 * gr[ch];
      if (deg > 0) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        basis[0] = -GS_SH_C1 * y; basis[1] = GS_SH_C1 * z; basis[2] = -GS_SH_C1 * x;
        if (deg > 1) {
          basis[3] = GS_SH_C2_0 * xy; basis[4] = GS_SH_C2_1 * yz; basis[5] = GS_SH_C2_2 * (2.f * zz - xx - yy);
          basis[6] = GS_SH_C2_3 * xz; basis[7] = GS_SH_C2_4 * (xx - yy);
        }
        if (deg > 2) {
          basis[8] = GS_SH_C3_0 * y * (3.f * xx - yy);
          basis[9] = GS_SH_C3_1 * xy * z;
          basis[10] = GS_SH_C3_2 * y * (4.f * zz - xx - yy);
          basis[11] = GS_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
          basis[12] = GS_SH_C3_4 * x * (4.f * zz - xx - yy);
          basis[13] = GS_SH_C3_5 * z * (xx - yy);
          basis[14] = GS_SH_C3_6 * x * (xx - 3.f * yy);
        }
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
          const float g = gr[ch];
          const float s1 = srow[0 + ch], s2 = srow[3 + ch], s3 = srow[6 + ch];
          float ax = -GS_SH_C1 * s3, ay = -GS_SH_C1 * s1, az = GS_SH_C1 * s2;
          if (deg > 1) {
            const float s4 = srow[9 + ch], s5 = srow[12 + ch], s6 = srow[15 + ch], s7 = srow[18 + ch],
                        s8 = srow[21 + ch];
            ax += GS_SH_C2_0 * y * s4 + GS_SH_C2_2 * -2.f * x * s6 + GS_SH_C2_3 * z * s7 + GS_SH_C2_4 * 2.f * x * s8;
            ay += GS_SH_C2_0 * x * s4 + GS_SH_C2_1 * z * s5 + GS_SH_C2_2 * -2.f * y * s6 + GS_SH_C2_4 * -2.f * y * s8;
            az += GS_SH_C2_1 * y * s5 + GS_SH_C2_2 * 4.f * z * s6 + GS_SH_C2_3 * x * s7;
            if (deg > 2) {
              const float s9 = srow[24 + ch], s10 = srow[27 + ch], s11 = srow[30 + ch], s12 = srow[33 + ch],
                          s13 = srow[36 + ch], s14 = srow[39 + ch], s15 = srow[42 + ch];
              ax += GS_SH_C3_0 * 6.f * xy * s9 + GS_SH_C3_1 * yz * s10 + GS_SH_C3_2 * -2.f * xy * s11 +
                    GS_SH_C3_3 * -6.f * xz * s12 + GS_SH_C3_4 * (4.f * zz - 3.f * xx - yy) * s13 +
                    GS_SH_C3_5 * 2.f * xz * s14 + GS_SH_C3_6 * (3.f * xx - 3.f * yy) * s15;
              ay += GS_SH_C3_0 * (3.f * xx - 3.f * yy) * s9 + GS_SH_C3_1 * xz * s10 +
                    GS_SH_C3_2 * (4.f * zz - xx - 3.f * yy) * s11 + GS_SH_C3_3 * -6.f * yz * s12 +
                    GS_SH_C3_4 * -2.f * xy * s13 + GS_SH_C3_5 * -2.f * yz * s14 + GS_SH_C3_6 * -6.f * xy * s15;
              az += GS_SH_C3_1 * xy * s10 + GS_SH_C3_2 * 8.f * yz * s11 +
                    GS_SH_C3_3 * (6.f * zz - 3.f * xx - 3.f * yy) * s12 + GS_SH_C3_4 * 8.f * xz * s13 +
                    GS_SH_C3_5 * (xx - yy) * s14;
            }
          }
          ddx += ax * g;
          ddy += ay * g;
          ddz += az * g;
        }
        const float dot = x * ddx + y * ddy + z * ddz;
        gm[0] += (ddx - x * dot) / len;
        gm[1] += (ddy - y * dot) / len;
        gm[2] += (ddz - z * dot) / len;
      }
    }
    // ---- cov3D -> scale, quaternion ----
    if (own_cov) {
      const float GS[3][3] = {{gcov[0], 0.5f * gcov[1], 0.5f * gcov[2]},
                              {0.5f * gcov[1], gcov[3], 0.5f * gcov[4]},
                              {0.5f * gcov[2], 0.5f * gcov[4], gcov[5]}};
      float gL[9], gR[9];
#pragma unroll
      for (int a2 = 0; a2 < 3; ++a2)
#pragma unroll
        for (int b2 = 0; b2 < 3; ++b2) {
          float acc = 0.f;
#pragma unroll
          for (int k = 0; k < 3; ++k) acc += GS[a2][k] * (R[3 * k + b2] * sm[b2]);
          gL[3 * a2 + b2] = 2.f * acc;
        }
#pragma unroll
      for (int b2 = 0; b2 < 3; ++b2) {
        float acc = 0.f;
#pragma unroll
        for (int a2 = 0; a2 < 3; ++a2) acc += gL[3 * a2 + b2] * R[3 * a2 + b2];
        gs[b2] = st.scale_modifier * acc;
        if (a.in.raw_params) gs[b2] *= sact[b2];  // d exp(raw)/d raw = exp(raw)
      }
#pragma unroll
      for (int a2 = 0; a2 < 3; ++a2)
#pragma unroll
        for (int b2 = 0; b2 < 3; ++b2) gR[3 * a2 + b2] = gL[3 * a2 + b2] * sm[b2];
      const float r = qn[0], x = qn[1], y = qn[2], z = qn[3];
      float gn[4];
      gn[0] = 2.f * (-z * gR[1] + y * gR[2] + z * gR[3] - x * gR[5] - y * gR[6] + x * gR[7]);
      gn[1] = 2.f * (y * gR[1] + z * gR[2] + y * gR[3] - 2.f * x * gR[4] - r * gR[5] + z * gR[6] + r * gR[7] -
                     2.f * x * gR[8]);
      gn[2] = 2.f * (-2.f * y * gR[0] + x * gR[1] + r * gR[2] + x * gR[3] + z * gR[5] - r * gR[6] + z * gR[7] -
                     2.f * y * gR[8]);
      gn[3] = 2.f * (-2.f * z * gR[0] - r * gR[1] + x * gR[2] + r * gR[3] - 2.f * z * gR[4] + y * gR[5] +
                     x * gR[6] + y * gR[7]);
      const float dotq = qn[0] * gn[0] + qn[1] * gn[1] + qn[2] * gn[2] + qn[3] * gn[3];
#pragma unroll
      for (int k = 0; k < 4; ++k) gq[k] = (gn[k] - qn[k] * dotq) / qlen;
    }
    if (a.in.raw_params) {
      const float sg = gs_sigmoid(opac_in);
      gop *= sg * (1.0f - sg);
    }
    o.gop = gop;
  }
}


__global__ void __launch_bounds__(PB_THREADS)
preprocess_backward_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ PreBwdArgs a, const int first) {
  extern __shared__ float s_sh[];
  const int tid = threadIdx.x;
  const int row0 = first + blockIdx.x * PB_THREADS;
  const int i = row0 + tid;
  const int rows = min(PB_THREADS, a.N - row0);
  const bool use_sh = a.in.colors_precomp == nullptr;

  if (use_sh && a.rest_width_act > 0) {
    const int stride = a.in.sh_rest_stride;
    const float* base = a.in.sh_rest + (size_t)row0 * stride;
    const int total = (rows - 1) * stride + a.rest_width_act;
    int row = tid / stride, col = tid % stride;
    const int drow = PB_THREADS / stride, dcol = PB_THREADS % stride;
    for (int f = tid; f < total; f += PB_THREADS) {
      if (col < a.rest_width_act) s_sh[row * a.pitch + col] = __ldg(base + f);
      row += drow;
      col += dcol;
      if (col >= stride) { col -= stride; ++row; }
    }
  }
  __syncthreads();

  PbOut o;
  pb_zero(o);
  float* srow = s_sh + tid * a.pitch;
  bool visible = false;
  if (i < a.N) visible = a.radii[i] > 0;
  if (visible) {
    const float4 G0 = a.grad_geom[3 * (size_t)i], G1 = a.grad_geom[3 * (size_t)i + 1];
    const float G2x = a.grad_geom[3 * (size_t)i + 2].x;
    const float p[3] = {a.in.means3D[3 * i], a.in.means3D[3 * i + 1], a.in.means3D[3 * i + 2]};
    const float* cov_pre = a.in.cov3D_precomp ? a.in.cov3D_precomp + 6 * (size_t)i : nullptr;
    float4 q4 = make_float4(1.f, 0.f, 0.f, 0.f);
    float scl[3] = {0.f, 0.f, 0.f};
    if (!cov_pre) {
      q4 = *reinterpret_cast<const float4*>(a.in.rotations + 4 * (size_t)i);
      scl[0] = a.in.scales[3 * i];
      scl[1] = a.in.scales[3 * i + 1];
      scl[2] = a.in.scales[3 * i + 2];
    }
    const float opac_in = a.in.raw_params ? a.in.opacities[i] : 0.f;
    const uint32_t cm = use_sh ? a.clamped[i] : 0u;
    pb_compute(st, a, p, q4, scl, opac_in, cov_pre, G0, G1, G2x, cm, srow, use_sh, o);
  }
  const float* gm = o.gm;
  const float* gs = o.gs;
  const float* gq = o.gq;
  const float* gcov = o.gcov;
  const float* gdc = o.gdc;
  const float* gcol = o.gcol;
  const float* gr = o.gr;
  const float* basis = o.basis;
  const float g2x = o.g2x, g2y = o.g2y, gop = o.gop;

  // ---- outputs ----
  const bool acc = a.out.accumulate != 0;
  if (i < a.N) {
    if (a.out.dL_dmeans3D) {
      float* o = a.out.dL_dmeans3D + 3 * (size_t)i;
#pragma unroll
      for (int c = 0; c < 3; ++c) o[c] = acc ? o[c] + gm[c] : gm[c];
    }
    if (a.out.dL_dscales) {
      float* o = a.out.dL_dscales + 3 * (size_t)i;
#pragma unroll
      for (int c = 0; c < 3; ++c) o[c] = acc ? o[c] + gs[c] : gs[c];
    }
    if (a.out.dL_drotations) {
      float4* o = reinterpret_cast<float4*>(a.out.dL_drotations + 4 * (size_t)i);
      float4 v = make_float4(gq[0], gq[1], gq[2], gq[3]);
      if (acc) { const float4 u = *o; v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w; }
      *o = v;
    }
    if (a.out.dL_dopacities) {
      float* o = a.out.dL_dopacities + i;
      *o = acc ? *o + gop : gop;
    }
    if (a.out.dL_dmeans2D) {  // a per-VIEW quantity (densification statistics): never accumulated
      float2* o = reinterpret_cast<float2*>(a.out.dL_dmeans2D + 2 * (size_t)i);
      *o = make_float2(g2x, g2y);
    }
    if (a.out.dL_dcov3D) {
      float* o = a.out.dL_dcov3D + 6 * (size_t)i;
#pragma unroll
      for (int c = 0; c < 6; ++c) o[c] = acc ? o[c] + gcov[c] : gcov[c];
    }
    if (a.out.dL_dcolors) {
      float* o = a.out.dL_dcolors + 3 * (size_t)i;
#pragma unroll
      for (int c = 0; c < 3; ++c) o[c] = acc ? o[c] + gcol[c] : gcol[c];
    }
    if (use_sh && a.out.dL_dsh_dc) {
      float* o = a.out.dL_dsh_dc + (size_t)i * a.out.sh_dc_stride;
#pragma unroll
      for (int c = 0; c < 3; ++c) o[c] = acc ? o[c] + gdc[c] : gdc[c];
    }
    // this thread's smem row <- SH gradient of the "rest" coefficients (zeros when inactive)
    if (use_sh && a.out.dL_dsh_rest && a.rest_width_max > 0) {
      const int kmax = a.rest_width_max / 3;
#pragma unroll
      for (int k = 0; k < 15; ++k)
        if (k < kmax) {
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) srow[3 * k + ch] = basis[k] * gr[ch];
        }
    }
  }
  if (!(use_sh && a.out.dL_dsh_rest && a.rest_width_max > 0)) return;
  __syncthreads();
  {
    const int stride = a.out.sh_rest_stride;
    float* base = a.out.dL_dsh_rest + (size_t)row0 * stride;
    const int width = acc ? a.rest_width_act : a.rest_width_max;  // accumulate: inactive rows add 0
    if (width == 0) return;
    const int total = (rows - 1) * stride + width;
    int row = tid / stride, col = tid % stride;
    const int drow = PB_THREADS / stride, dcol = PB_THREADS % stride;
    for (int f = tid; f < total; f += PB_THREADS) {
      if (col < width) {
        const float v = s_sh[row * a.pitch + col];
        base[f] = acc ? base[f] + v : v;
      }
      row += drow;
      col += dcol;
      if (col >= stride) { col -= stride; ++row; }
    }
  }
}

// ---- TMA-staged path ---------------------------------------------------------------------------
#define PBB_ROWS 128
// shared-memory slices of one block (floats); every slice offset and size is a multiple of 16 B
#define PBB_OFF_SH 0                               // rows x 45 (SH rest in, dSH rest out)
#define PBB_OFF_XYZ (PBB_OFF_SH + PBB_ROWS * 45)   // rows x 3  (means in, dmeans3D out)
#define PBB_OFF_SCL (PBB_OFF_XYZ + PBB_ROWS * 3)   // rows x 3
#define PBB_OFF_ROT (PBB_OFF_SCL + PBB_ROWS * 3)   // rows x 4
#define PBB_OFF_OP (PBB_OFF_ROT + PBB_ROWS * 4)    // rows x 1
#define PBB_OFF_GG (PBB_OFF_OP + PBB_ROWS)         // rows x 12 (blend gradient records in)
#define PBB_OFF_M2D (PBB_OFF_GG + PBB_ROWS * 12)   // rows x 2  (dmeans2D out)
#define PBB_OFF_DC (PBB_OFF_M2D + PBB_ROWS * 2)    // rows x 3  (dSH dc out)
#define PBB_FLOATS (PBB_OFF_DC + PBB_ROWS * 3)
#define PBB_SMEM_BYTES (PBB_FLOATS * 4 + 16)

// Multi-GPU gradient routing (capi.cu, gs_set_gradient_routing): rows [q*R, (q+1)*R) of the six parameter
// gradients belong to rank q and are stored through vb[q][k] -- the "virtual base" of this rank's slot inside
// rank q's buffer, re-based so that the global row index addresses it (slot + (i - q*R)*C == vb + i*C).
// Consecutive blocks serve DIFFERENT owners (block b -> owner (b % G + rank + 1) % G, chunk b / G), so all G-1
// NVLink destinations and the local HBM are written concurrently for the whole kernel, and no GPU's ingress is
// hit by every sender at once.
struct PbRoute {
  int G;  // 0: no routing (sinks of PreBwdArgs)
  int rank, R;
  int nblk[8];    // full 128-row blocks among rank q's valid rows
  float* vb[8][6];  // order: xyz, f_dc, f_rest, opacity, scaling, rotation
};

__global__ void __launch_bounds__(PBB_ROWS)
preprocess_backward_bulk_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ PreBwdArgs a,
                                const int first, const __grid_constant__ PbRoute rt) {
  extern __shared__ __align__(128) float s_buf[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_buf + PBB_FLOATS);
  const int tid = threadIdx.x;
  size_t row0 = (size_t)first + (size_t)blockIdx.x * PBB_ROWS;
  float* o_xyz = a.out.dL_dmeans3D;
  float* o_dc = a.out.dL_dsh_dc;
  float* o_sh = a.out.dL_dsh_rest;
  float* o_op = a.out.dL_dopacities;
  float* o_scl = a.out.dL_dscales;
  float* o_rot = a.out.dL_drotations;
  bool sparse_sh = false;  // remote f_rest slot: only rows with a non-zero gradient are sent (see below)
  if (rt.G > 0) {
    const int q = ((int)(blockIdx.x % rt.G) + rt.rank + 1) % rt.G;
    sparse_sh = q != rt.rank;
    const int c = (int)(blockIdx.x / rt.G);
    if (c >= rt.nblk[q]) return;  // (whole block: nothing has touched the barrier yet)
    row0 = (size_t)q * rt.R + (size_t)c * PBB_ROWS;
    o_xyz = rt.vb[q][0];
    o_dc = rt.vb[q][1];
    o_sh = rt.vb[q][2];
    o_op = rt.vb[q][3];
    o_scl = rt.vb[q][4];
    o_rot = rt.vb[q][5];
  }
  const size_t i = row0 + tid;
  const int W = a.rest_width_max;  // == sh_rest_stride (dense rows), 0..45
  const bool load_sh = a.rest_width_act > 0;

  if (tid == 0) {
    tma::mbar_init(bar, 1);
    tma::fence_mbar_init();
    const uint32_t sh_bytes = load_sh ? (uint32_t)(PBB_ROWS * W * 4) : 0u;
    const uint32_t op_bytes = a.in.raw_params ? PBB_ROWS * 4u : 0u;
    tma::mbar_arrive_expect_tx(bar, sh_bytes + op_bytes + PBB_ROWS * (12u + 12u + 16u + 48u));
    tma::load(s_buf + PBB_OFF_GG, a.grad_geom + 3 * row0, PBB_ROWS * 48u, bar);
    tma::load(s_buf + PBB_OFF_XYZ, a.in.means3D + 3 * row0, PBB_ROWS * 12u, bar);
    tma::load(s_buf + PBB_OFF_SCL, a.in.scales + 3 * row0, PBB_ROWS * 12u, bar);
    tma::load(s_buf + PBB_OFF_ROT, a.in.rotations + 4 * row0, PBB_ROWS * 16u, bar);
    if (op_bytes) tma::load(s_buf + PBB_OFF_OP, a.in.opacities + row0, op_bytes, bar);
    if (sh_bytes) tma::load(s_buf + PBB_OFF_SH, a.in.sh_rest + row0 * W, sh_bytes, bar);
  }
  const bool visible = a.radii[i] > 0;
  const uint32_t cm = a.clamped[i];
  __syncthreads();  // barrier initialised before anyone polls it
  tma::mbar_wait(bar, 0);

  PbOut o;
  pb_zero(o);
  float* srow = s_buf + PBB_OFF_SH + tid * W;
  if (visible) {
    const float4* gg = reinterpret_cast<const float4*>(s_buf + PBB_OFF_GG) + 3 * tid;
    const float4 G0 = gg[0], G1 = gg[1];
    const float G2x = s_buf[PBB_OFF_GG + 12 * tid + 8];
    const float p[3] = {s_buf[PBB_OFF_XYZ + 3 * tid], s_buf[PBB_OFF_XYZ + 3 * tid + 1], s_buf[PBB_OFF_XYZ + 3 * tid + 2]};
    const float scl[3] = {s_buf[PBB_OFF_SCL + 3 * tid], s_buf[PBB_OFF_SCL + 3 * tid + 1], s_buf[PBB_OFF_SCL + 3 * tid + 2]};
    const float4 q4 = reinterpret_cast<const float4*>(s_buf + PBB_OFF_ROT)[tid];
    const float opac_in = a.in.raw_params ? s_buf[PBB_OFF_OP + tid] : 0.f;
    pb_compute(st, a, p, q4, scl, opac_in, nullptr, G0, G1, G2x, cm, srow, true, o);
  }
  __shared__ uint32_t s_live[PBB_ROWS / 32];
  {
    // dL/dSH = basis x dL/dcolour: a Gaussian whose colour gradient is exactly zero (culled, or behind the
    // early-termination front of every pixel it covers) has an all-zero f_rest gradient row
    const bool live = visible && (o.gr[0] != 0.0f || o.gr[1] != 0.0f || o.gr[2] != 0.0f);
    const uint32_t lw = __ballot_sync(0xffffffffu, live);
    if ((tid & 31) == 0) s_live[tid >> 5] = lw;
  }
  // gradients overwrite the input slices in place (each thread touches only its own rows)
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    s_buf[PBB_OFF_XYZ + 3 * tid + c] = o.gm[c];
    s_buf[PBB_OFF_SCL + 3 * tid + c] = o.gs[c];
    s_buf[PBB_OFF_DC + 3 * tid + c] = o.gdc[c];
  }
  reinterpret_cast<float4*>(s_buf + PBB_OFF_ROT)[tid] = make_float4(o.gq[0], o.gq[1], o.gq[2], o.gq[3]);
  s_buf[PBB_OFF_OP + tid] = o.gop;
  reinterpret_cast<float2*>(s_buf + PBB_OFF_M2D)[tid] = make_float2(o.g2x, o.g2y);
  {
    const int kmax = W / 3;
#pragma unroll
    for (int k = 0; k < 15; ++k)
      if (k < kmax) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) srow[3 * k + ch] = o.basis[k] * o.gr[ch];
      }
  }
  tma::fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) {
    const bool acc = a.out.accumulate != 0;
#define PBB_OUT(gptr, soff, bytes)                                          \
  do {                                                                      \
    if (gptr) {                                                             \
      if (acc) tma::reduce_add_f32((gptr), s_buf + (soff), (bytes));        \
      else tma::store((gptr), s_buf + (soff), (bytes));                     \
    }                                                                       \
  } while (0)
    if (W > 0 && !sparse_sh) PBB_OUT(o_sh ? o_sh + row0 * W : nullptr, PBB_OFF_SH, (uint32_t)(PBB_ROWS * W * 4));
    PBB_OUT(o_xyz ? o_xyz + 3 * row0 : nullptr, PBB_OFF_XYZ, PBB_ROWS * 12u);
    PBB_OUT(o_scl ? o_scl + 3 * row0 : nullptr, PBB_OFF_SCL, PBB_ROWS * 12u);
    PBB_OUT(o_rot ? o_rot + 4 * row0 : nullptr, PBB_OFF_ROT, PBB_ROWS * 16u);
    PBB_OUT(o_op ? o_op + row0 : nullptr, PBB_OFF_OP, PBB_ROWS * 4u);
    if (a.out.dL_dmeans2D) tma::store(a.out.dL_dmeans2D + 2 * row0, s_buf + PBB_OFF_M2D, PBB_ROWS * 8u);  // per view: never accumulated
    PBB_OUT(o_dc ? o_dc + 3 * row0 : nullptr, PBB_OFF_DC, PBB_ROWS * 12u);
#undef PBB_OUT
    tma::commit_group();
  }
  if (sparse_sh && W > 0 && o_sh) {
    // SPARSE push of the f_rest rows (76% of the gradient bytes): a row crosses NVLink only if it is non-zero;
    // the owner re-zeroes what it consumes (adam_peer_kernel), so an unwritten row reads as zero.  One warp per
    // row, 4-byte lanes: a 180-B row is two store instructions.
    const int warp = tid >> 5, lane = tid & 31;
    for (int r = warp; r < PBB_ROWS; r += PBB_ROWS / 32) {
      if (!((s_live[r >> 5] >> (r & 31)) & 1u)) continue;
      float* dst = o_sh + (row0 + r) * W;
      const float* src = s_buf + PBB_OFF_SH + r * W;
      for (int c = lane; c < W; c += 32) dst[c] = src[c];
    }
  }
  if (tid == 0) tma::wait_group_read0();  // shared memory must stay alive until the copies have read it
}

// ---- K4b + K5 fused (single-GPU training step) -------------------------------------------------
// The gradient rows a block has just produced sit in shared memory; instead of storing them (236 B/row) for
// the optimiser to read back (another 236 B/row), the block applies the Adam update of optim.cu to its own
// 128 rows right away: moments from HBM, parameters from L2 (the block's own TMA loads fetched them a moment
// ago), gradients from shared memory.  Per row: 48 B record + 236 B parameters in; 472 B moments in and out;
// 236 B parameters out = 1464 B instead of 2172 B for the two separate kernels, and one launch less.  The
// densification statistics of K5 are folded in as well (radius, |dL/dmean2D|), so dL_dmeans2D need not be
// stored either (optional sink).  Bit-identical to K4b followed by K5: same pb_compute, same gs_adam_one.
struct FusedAdamArgs {
  float* p[6];  // group order of GaussianModel::setup: xyz, f_dc, f_rest, opacity, scaling, rotation
  float* m[6];
  float* v[6];
  float neg_step[6], inv_bc2_sqrt[6];
  float b1, b2, omb1, omb2, eps;
  const float* skip_guard;
  int l2_prefetch;   // pull the block's moments into L2 at block start (cp.async.bulk.prefetch.L2)
  int* max_radii2D;  // the three statistics: all or none
  float* xyz_grad_accum;
  float* denom;
};


// The block's 128 x (11 + 3K) gradient floats as chunks of 32 float4 (512 B), every chunk inside ONE group because
// each group's slice is 128 rows x C floats = C chunks: chunk index -> (group, offset).  A warp walks chunks
// warp, warp + 4, ... in rounds of PBF_UNROLL, DOUBLE-BUFFERED: the moments and parameters of the next round are
// requested (3 x PBF_UNROLL independent 16-byte loads per thread) before the current round is computed, so the L2
// latency of the Adam phase's loads -- 35 % of the kernel's stall samples when every round waited for its own loads,
// profiles/r02a -- is spent under arithmetic.  Only the group and one 32-bit vector index are kept per chunk.
// the 14 chunks behind the W f_rest chunks: xyz 3, f_dc 3, opacity 1, scaling 3, rotation 4
__constant__ int c_pbf_group[14] = {0, 0, 0, 1, 1, 1, 3, 4, 4, 4, 5, 5, 5, 5};
__constant__ int c_pbf_sub[14] = {0, 1, 2, 0, 1, 2, 0, 0, 1, 2, 0, 1, 2, 3};
__constant__ int c_pbf_width[14] = {3, 3, 3, 3, 3, 3, 1, 3, 3, 3, 4, 4, 4, 4};
__constant__ int c_pbf_soff[14] = {PBB_OFF_XYZ, PBB_OFF_XYZ, PBB_OFF_XYZ, PBB_OFF_DC, PBB_OFF_DC, PBB_OFF_DC, PBB_OFF_OP,
                                   PBB_OFF_SCL, PBB_OFF_SCL, PBB_OFF_SCL, PBB_OFF_ROT, PBB_OFF_ROT, PBB_OFF_ROT, PBB_OFF_ROT};

template <int PBF_UNROLL>
struct PbfRound {
  float4 p[PBF_UNROLL], m[PBF_UNROLL], v[PBF_UNROLL];
  uint32_t vec[PBF_UNROLL];   // float4 index of this lane's vector inside the group's arrays
  uint32_t gvec[PBF_UNROLL];  // float4 index of its gradient inside the block's shared memory
  int k[PBF_UNROLL];          // group (-1: no such chunk)
};

template <int PBF_UNROLL>
__device__ __forceinline__ void pbf_request(PbfRound<PBF_UNROLL>& R, const FusedAdamArgs& f, const size_t row0, const int W,
                                            const int nchunks, const int c0, const int lane) {
#pragma unroll
  for (int u = 0; u < PBF_UNROLL; ++u) {
    const int c = c0 + 4 * u;
    R.k[u] = -1;
    if (c < nchunks) {
      const bool rest = c < W;
      const int t = rest ? 0 : c - W;
      const int k = rest ? 2 : c_pbf_group[t];
      const int sub = rest ? c : c_pbf_sub[t];
      const int cw = rest ? W : c_pbf_width[t];
      const int soff = rest ? PBB_OFF_SH : c_pbf_soff[t];
      R.k[u] = k;
      R.gvec[u] = (uint32_t)(soff / 4 + sub * 32 + lane);                    // (every slice offset is a multiple of 4)
      R.vec[u] = (uint32_t)((row0 * cw) / 4 + (size_t)sub * 32 + lane);      // (row0 is a multiple of 128)
      R.m[u] = __ldcs(reinterpret_cast<const float4*>(f.m[k]) + R.vec[u]);
      R.v[u] = __ldcs(reinterpret_cast<const float4*>(f.v[k]) + R.vec[u]);
      R.p[u] = reinterpret_cast<const float4*>(f.p[k])[R.vec[u]];
    }
  }
}

template <int PBF_UNROLL>
__device__ __forceinline__ void pbf_update(PbfRound<PBF_UNROLL>& R, const float* s_buf, const FusedAdamArgs& f, const GsAdamConst& kc) {
  float4 g[PBF_UNROLL];
  GsAdamQuad q[PBF_UNROLL];
  bool idle[PBF_UNROLL];
  bool any_active = false;
  int mode = GS_ADAM_FAST;  // the weakest path any of this lane's live vectors needs
#pragma unroll
  for (int u = 0; u < PBF_UNROLL; ++u) {
    idle[u] = true;
    if (R.k[u] >= 0) {
      g[u] = reinterpret_cast<const float4*>(s_buf)[R.gvec[u]];
      idle[u] = gs_adam_idle(g[u], R.m[u], R.v[u]);  // exact fixed point (gs_common.cuh): nothing to write
      const int ok = gs_adam_quad_moments(g[u], R.m[u], R.v[u], kc, q[u]);
      if (!idle[u]) mode = min(mode, ok);
      any_active = any_active || !idle[u];
    }
  }
  if (!__any_sync(0xffffffffu, any_active)) return;
  // packed paths (gs_common.cuh): the round's pairs in one basic block -- independent dependency chains
  if (__all_sync(0xffffffffu, mode == GS_ADAM_FAST)) {
#pragma unroll
    for (int u = 0; u < PBF_UNROLL; ++u)
      if (R.k[u] >= 0)
        gs_adam_quad_step<false>(R.p[u], R.m[u], R.v[u], q[u], kc.eps, f.neg_step[R.k[u]], f.inv_bc2_sqrt[R.k[u]]);
  } else if (__all_sync(0xffffffffu, mode >= GS_ADAM_FAST_Z)) {
#pragma unroll
    for (int u = 0; u < PBF_UNROLL; ++u)
      if (R.k[u] >= 0)
        gs_adam_quad_step<true>(R.p[u], R.m[u], R.v[u], q[u], kc.eps, f.neg_step[R.k[u]], f.inv_bc2_sqrt[R.k[u]]);
  } else if (__all_sync(0xffffffffu, mode >= GS_ADAM_FAST_T)) {
    // tiny first moments (Gaussians unseen for ~480 iterations and more): still packed, see gs_common.cuh
#pragma unroll
    for (int u = 0; u < PBF_UNROLL; ++u) {
      const int k = R.k[u] >= 0 ? R.k[u] : 0;
      gs_adam_quad_step_tiny(R.p[u], g[u], R.m[u], R.v[u], q[u], kc, f.neg_step[k], f.inv_bc2_sqrt[k],
                             R.k[u] >= 0 && !idle[u]);
    }
  } else {
#pragma unroll
    for (int u = 0; u < PBF_UNROLL; ++u)
      if (R.k[u] >= 0 && !idle[u])
        gs_adam_quad_scalar(R.p[u], g[u], R.m[u], R.v[u], kc, f.neg_step[R.k[u]], f.inv_bc2_sqrt[R.k[u]]);
  }
#pragma unroll
  for (int u = 0; u < PBF_UNROLL; ++u)
    if (R.k[u] >= 0 && !idle[u]) {
      const int k = R.k[u];
      __stcs(reinterpret_cast<float4*>(f.m[k]) + R.vec[u], R.m[u]);
      __stcs(reinterpret_cast<float4*>(f.v[k]) + R.vec[u], R.v[u]);
      reinterpret_cast<float4*>(f.p[k])[R.vec[u]] = R.p[u];  // read again by the next forward
    }
}

template <int PBF_UNROLL, int MIN_BLOCKS, int DEPTH>
__global__ void __launch_bounds__(PBB_ROWS, MIN_BLOCKS)
preprocess_backward_adam_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ PreBwdArgs a,
                                const __grid_constant__ FusedAdamArgs f) {
  if (f.skip_guard && __ldg(f.skip_guard) != 0.0f) return;  // overflowed forward: the whole step is void
  extern __shared__ __align__(128) float s_buf[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_buf + PBB_FLOATS);
  const int tid = threadIdx.x;
  const size_t row0 = (size_t)blockIdx.x * PBB_ROWS;
  const size_t i = row0 + tid;
  const int W = a.rest_width_max;
  const bool load_sh = a.rest_width_act > 0;

  if (tid == 0) {
    tma::mbar_init(bar, 1);
    tma::fence_mbar_init();
    const uint32_t sh_bytes = load_sh ? (uint32_t)(PBB_ROWS * W * 4) : 0u;
    const uint32_t op_bytes = a.in.raw_params ? PBB_ROWS * 4u : 0u;
    tma::mbar_arrive_expect_tx(bar, sh_bytes + op_bytes + PBB_ROWS * (12u + 12u + 16u + 48u));
    tma::load(s_buf + PBB_OFF_GG, a.grad_geom + 3 * row0, PBB_ROWS * 48u, bar);
    tma::load(s_buf + PBB_OFF_XYZ, a.in.means3D + 3 * row0, PBB_ROWS * 12u, bar);
    tma::load(s_buf + PBB_OFF_SCL, a.in.scales + 3 * row0, PBB_ROWS * 12u, bar);
    tma::load(s_buf + PBB_OFF_ROT, a.in.rotations + 4 * row0, PBB_ROWS * 16u, bar);
    if (op_bytes) tma::load(s_buf + PBB_OFF_OP, a.in.opacities + row0, op_bytes, bar);
    if (sh_bytes) tma::load(s_buf + PBB_OFF_SH, a.in.sh_rest + row0 * W, sh_bytes, bar);
    // the Adam phase below reads the moments (and the one parameter slice not staged above) with plain loads:
    // pull them into L2 now, underneath the gradient math
    const int cw[6] = {3, 3, W, 1, 3, 4};
    if (f.l2_prefetch) {
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (cw[k] > 0) {
          tma::prefetch_l2(f.m[k] + row0 * cw[k], PBB_ROWS * 4u * cw[k]);
          tma::prefetch_l2(f.v[k] + row0 * cw[k], PBB_ROWS * 4u * cw[k]);
        }
      tma::prefetch_l2(f.p[1] + 3 * row0, PBB_ROWS * 12u);
      if (!load_sh && W > 0) tma::prefetch_l2(f.p[2] + row0 * W, (uint32_t)(PBB_ROWS * W * 4));
      if (!a.in.raw_params) tma::prefetch_l2(f.p[3] + row0, PBB_ROWS * 4u);
    }
  }
  const int radius = a.radii[i];
  const bool visible = radius > 0;
  const uint32_t cm = a.clamped[i];
  // the three statistics are read-modify-write: request them now, consume them after the gradient math
  int stat_maxr = 0;
  float stat_acc = 0.f, stat_den = 0.f;
  if (f.max_radii2D && visible) {
    stat_maxr = f.max_radii2D[i];
    stat_acc = f.xyz_grad_accum[i];
    stat_den = f.denom[i];
  }
  __syncthreads();
  tma::mbar_wait(bar, 0);

  PbOut o;
  pb_zero(o);
  float* srow = s_buf + PBB_OFF_SH + tid * W;
  if (visible) {
    const float4* gg = reinterpret_cast<const float4*>(s_buf + PBB_OFF_GG) + 3 * tid;
    const float4 G0 = gg[0], G1 = gg[1];
    const float G2x = s_buf[PBB_OFF_GG + 12 * tid + 8];
    const float p[3] = {s_buf[PBB_OFF_XYZ + 3 * tid], s_buf[PBB_OFF_XYZ + 3 * tid + 1], s_buf[PBB_OFF_XYZ + 3 * tid + 2]};
    const float scl[3] = {s_buf[PBB_OFF_SCL + 3 * tid], s_buf[PBB_OFF_SCL + 3 * tid + 1], s_buf[PBB_OFF_SCL + 3 * tid + 2]};
    const float4 q4 = reinterpret_cast<const float4*>(s_buf + PBB_OFF_ROT)[tid];
    const float opac_in = a.in.raw_params ? s_buf[PBB_OFF_OP + tid] : 0.f;
    pb_compute(st, a, p, q4, scl, opac_in, nullptr, G0, G1, G2x, cm, srow, true, o);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    s_buf[PBB_OFF_XYZ + 3 * tid + c] = o.gm[c];
    s_buf[PBB_OFF_SCL + 3 * tid + c] = o.gs[c];
    s_buf[PBB_OFF_DC + 3 * tid + c] = o.gdc[c];
  }
  reinterpret_cast<float4*>(s_buf + PBB_OFF_ROT)[tid] = make_float4(o.gq[0], o.gq[1], o.gq[2], o.gq[3]);
  s_buf[PBB_OFF_OP + tid] = o.gop;
  {
    const int kmax = W / 3;
#pragma unroll
    for (int k = 0; k < 15; ++k)
      if (k < kmax) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) srow[3 * k + ch] = o.basis[k] * o.gr[ch];
      }
  }
  // first Adam round's moments / parameters: requested before the block's gradients are complete in shared memory
  const int nchunks = W + 14;
  const int warp = tid >> 5, lane = tid & 31;
  constexpr int STRIDE = 4 * PBF_UNROLL;  // chunks one round of this warp covers
  PbfRound<PBF_UNROLL> R[DEPTH];         // ring of rounds in flight: DEPTH - 1 rounds are requested ahead of the one computed
#pragma unroll
  for (int d = 0; d < DEPTH - 1; ++d) pbf_request(R[d], f, row0, W, nchunks, warp + d * STRIDE, lane);
  // densification statistics of this row (K5's second half) and the optional dL_dmeans2D sink
  if (f.max_radii2D && visible) {
    if (radius > stat_maxr) f.max_radii2D[i] = radius;
    f.xyz_grad_accum[i] = __fadd_rn(stat_acc, __fsqrt_rn(__fadd_rn(__fmul_rn(o.g2x, o.g2x), __fmul_rn(o.g2y, o.g2y))));
    f.denom[i] = __fadd_rn(stat_den, 1.0f);
  }
  if (a.out.dL_dmeans2D) reinterpret_cast<float2*>(a.out.dL_dmeans2D)[i] = make_float2(o.g2x, o.g2y);
  __syncthreads();
  // ---- Adam over the block's rows (chunked and double-buffered, see pbf_chunk) ----
  const GsAdamConst kc = {f.b1, f.b2, f.omb1, f.omb2, f.eps};
  for (int c0 = warp; c0 < nchunks; c0 += DEPTH * STRIDE) {
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) {
      pbf_request(R[(d + DEPTH - 1) % DEPTH], f, row0, W, nchunks, c0 + (d + DEPTH - 1) * STRIDE, lane);
      pbf_update(R[d], s_buf, f, kc);
    }
  }
}

static bool aligned16(const void* p) { return (((uintptr_t)p) & 15) == 0; }

// Rows [row_begin, row_end) only (the whole model: 0, N).  The multi-GPU step launches one range per owner rank
// with the parameter-gradient sinks pointing into that rank's slot buffer (capi.cu, gradient routing).
cudaError_t gs_launch_preprocess_backward(const gs_settings& st, const gs_gaussians& in, const gs_gradients& out,
                                          int row_begin, int row_end, const int* radii, const uint8_t* clamped,
                                          const float4* grad_geom, cudaStream_t stream) {
  if (row_end <= row_begin) return cudaSuccess;
  PreBwdArgs a;
  a.in = in;
  a.out = out;
  a.N = row_end;
  a.radii = radii;
  a.clamped = clamped;
  a.grad_geom = grad_geom;
  a.focal_x = (float)st.W / (2.0f * st.tanfovx);
  a.focal_y = (float)st.H / (2.0f * st.tanfovy);
  const int kact = (st.sh_degree + 1) * (st.sh_degree + 1);
  const int kmax = (st.max_sh_degree + 1) * (st.max_sh_degree + 1);
  a.rest_width_act = in.colors_precomp ? 0 : 3 * (kact - 1);
  a.rest_width_max = in.colors_precomp ? 0 : 3 * (kmax - 1);
  a.pitch = a.rest_width_max | 1;

  // TMA-staged path: dense, 16-byte aligned reference tensors (the GaussianModel layout) and the SH path
  const int W = a.rest_width_max;
  const bool dense = !in.colors_precomp && !in.cov3D_precomp && !out.dL_dcov3D && !out.dL_dcolors &&
                     (W == 0 || (in.sh_rest_stride == W && (!out.dL_dsh_rest || out.sh_rest_stride == W))) &&
                     (!out.dL_dsh_dc || out.sh_dc_stride == 3) && (W * 4 * PBB_ROWS) % 16 == 0 &&
                     aligned16(in.means3D) && aligned16(in.scales) && aligned16(in.rotations) &&
                     aligned16(in.opacities) && (W == 0 || aligned16(in.sh_rest)) && aligned16(out.dL_dmeans3D) &&
                     aligned16(out.dL_dscales) && aligned16(out.dL_drotations) && aligned16(out.dL_dopacities) &&
                     aligned16(out.dL_dmeans2D) && aligned16(out.dL_dsh_dc) && aligned16(out.dL_dsh_rest) &&
                     aligned16(grad_geom);
  int first = row_begin;
  const int N = row_end - row_begin;
  if (dense && N >= PBB_ROWS && row_begin % 4 == 0) {
    static bool attr_set = false;
    if (!attr_set) {
      cudaError_t e = cudaFuncSetAttribute(preprocess_backward_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           PBB_SMEM_BYTES);
      if (e != cudaSuccess) return e;
      attr_set = true;
    }
    const int blocks = N / PBB_ROWS;
    PbRoute none;
    memset(&none, 0, sizeof(none));
    preprocess_backward_bulk_kernel<<<blocks, PBB_ROWS, PBB_SMEM_BYTES, stream>>>(st, a, first, none);
    first += blocks * PBB_ROWS;
  }
  if (first < row_end) {
    const size_t smem = a.rest_width_max > 0 ? (size_t)PB_THREADS * a.pitch * sizeof(float) : 0;
    const int blocks = (row_end - first + PB_THREADS - 1) / PB_THREADS;
    preprocess_backward_kernel<<<blocks, PB_THREADS, smem, stream>>>(st, a, first);
  }
  return cudaGetLastError();
}


// Fused K4b + K5 over rows [0, floor(N/128)*128); returns the number of rows it covered in *covered (0 when the
// layout does not qualify -- the caller then runs the separate kernels).  Parameters are in.{means3D, sh_dc,
// sh_rest, opacities, scales, rotations}, updated in place.
cudaError_t gs_launch_preprocess_backward_adam(const gs_settings& st, const gs_gaussians& in, float* dL_dmeans2D, int N,
                                               const int* radii, const uint8_t* clamped, const float4* grad_geom,
                                               float* const exp_avg[6], float* const exp_avg_sq[6], const double lr[6],
                                               const long long step[6], double beta1, double beta2, double eps,
                                               int* max_radii2D, float* xyz_grad_accum, float* denom,
                                               const float* skip_guard, int* covered, cudaStream_t stream) {
  *covered = 0;
  if (N < PBB_ROWS) return cudaSuccess;
  PreBwdArgs a;
  a.in = in;
  memset(&a.out, 0, sizeof(a.out));
  a.out.dL_dmeans2D = dL_dmeans2D;
  a.N = N;
  a.radii = radii;
  a.clamped = clamped;
  a.grad_geom = grad_geom;
  a.focal_x = (float)st.W / (2.0f * st.tanfovx);
  a.focal_y = (float)st.H / (2.0f * st.tanfovy);
  const int kact = (st.sh_degree + 1) * (st.sh_degree + 1);
  const int kmax = (st.max_sh_degree + 1) * (st.max_sh_degree + 1);
  a.rest_width_act = 3 * (kact - 1);
  a.rest_width_max = 3 * (kmax - 1);
  a.pitch = a.rest_width_max | 1;
  const int W = a.rest_width_max;
  float* params[6] = {const_cast<float*>(in.means3D), const_cast<float*>(in.sh_dc), const_cast<float*>(in.sh_rest),
                      const_cast<float*>(in.opacities), const_cast<float*>(in.scales), const_cast<float*>(in.rotations)};
  bool ok = !in.colors_precomp && !in.cov3D_precomp && in.sh_dc && in.sh_dc_stride == 3 &&
            (W == 0 || (in.sh_rest && in.sh_rest_stride == W)) && (W * 4 * PBB_ROWS) % 16 == 0 &&
            aligned16(grad_geom) && aligned16(dL_dmeans2D);
  FusedAdamArgs f;
  for (int k = 0; k < 6; ++k) {
    const bool on = !(k == 2 && W == 0);
    f.p[k] = on ? params[k] : nullptr;
    f.m[k] = on ? exp_avg[k] : nullptr;
    f.v[k] = on ? exp_avg_sq[k] : nullptr;
    if (on) ok = ok && f.p[k] && f.m[k] && f.v[k] && aligned16(f.p[k]) && aligned16(f.m[k]) && aligned16(f.v[k]);
    const double bc1 = 1.0 - pow(beta1, (double)step[k]);
    const double bc2 = 1.0 - pow(beta2, (double)step[k]);
    f.neg_step[k] = (float)(-(lr[k] / bc1));
    f.inv_bc2_sqrt[k] = 1.0f / (float)sqrt(bc2);
  }
  if (!ok) return cudaSuccess;
  f.b1 = (float)beta1;
  f.b2 = (float)beta2;
  f.omb1 = (float)(1.0 - beta1);
  f.omb2 = (float)(1.0 - beta2);
  f.eps = (float)eps;
  f.skip_guard = skip_guard;
  const bool stats = max_radii2D && xyz_grad_accum && denom;
  f.max_radii2D = stats ? max_radii2D : nullptr;
  f.xyz_grad_accum = stats ? xyz_grad_accum : nullptr;
  f.denom = stats ? denom : nullptr;
  // tuning switch (measurement only): GS_FUSED_VARIANT = chunks per Adam round x resident blocks per SM
  static int variant = -1;
  if (variant < 0) {
    const char* env = getenv("GS_FUSED_VARIANT");
    variant = env ? atoi(env) : 0;
    if (variant < 0 || variant > 7) variant = 0;
  }
  static int no_pf = -1;
  if (no_pf < 0) {
    const char* env = getenv("GS_FUSED_NO_L2_PREFETCH");
    no_pf = env && env[0] == '1';
  }
  f.l2_prefetch = !no_pf;
  const int blocks = N / PBB_ROWS;
#define PBF_LAUNCH(U, B, D)                                                                                        \
  do {                                                                                                             \
    cudaError_t e = cudaFuncSetAttribute(preprocess_backward_adam_kernel<U, B, D>,                                 \
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, PBB_SMEM_BYTES);             \
    if (e != cudaSuccess) return e;                                                                                \
    preprocess_backward_adam_kernel<U, B, D><<<blocks, PBB_ROWS, PBB_SMEM_BYTES, stream>>>(st, a, f);              \
  } while (0)
  switch (variant) {
    case 2: PBF_LAUNCH(2, 4, 2); break;
    case 3: PBF_LAUNCH(1, 6, 2); break;
    case 6: PBF_LAUNCH(1, 5, 3); break;
    default: PBF_LAUNCH(1, 5, 2); break;
  }
#undef PBF_LAUNCH
  *covered = blocks * PBB_ROWS;
  return cudaGetLastError();
}


// The routed K4b of the multi-GPU step: ONE launch covers all owners (interleaved, see PbRoute); table[q*6 + k] is
// this rank's slot for group k inside rank q's buffer.  The ragged tail (N % 128 rows, all in one shard) goes
// through the generic kernel.  `out` supplies only the local dL_dmeans2D sink.
cudaError_t gs_launch_preprocess_backward_routed(const gs_settings& st, const gs_gaussians& in, const gs_gradients& out,
                                                 int N, int G, int rank, int R, float* const table[], const int* radii,
                                                 const uint8_t* clamped, const float4* grad_geom, cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  const int kmax = (st.max_sh_degree + 1) * (st.max_sh_degree + 1);
  const int W = 3 * (kmax - 1);
  const int cw[6] = {3, 3, W, 1, 3, 4};
  PbRoute rt;
  memset(&rt, 0, sizeof(rt));
  rt.G = G;
  rt.rank = rank;
  rt.R = R;
  int max_blk = 0;
  for (int q = 0; q < G; ++q) {
    const long long b = (long long)q * R;
    const long long rows = b >= N ? 0 : ((long long)N - b < R ? (long long)N - b : (long long)R);
    rt.nblk[q] = (int)(rows / PBB_ROWS);
    if (rt.nblk[q] > max_blk) max_blk = rt.nblk[q];
    for (int k = 0; k < 6; ++k) rt.vb[q][k] = (k == 2 && W == 0) ? nullptr : table[q * 6 + k] - b * cw[k];
  }
  PreBwdArgs a;
  a.in = in;
  a.out = out;
  a.out.dL_dmeans3D = a.out.dL_dsh_dc = a.out.dL_dsh_rest = a.out.dL_dopacities = a.out.dL_dscales = a.out.dL_drotations = nullptr;
  a.out.dL_dcov3D = a.out.dL_dcolors = nullptr;
  a.out.accumulate = 0;
  a.N = N;
  a.radii = radii;
  a.clamped = clamped;
  a.grad_geom = grad_geom;
  a.focal_x = (float)st.W / (2.0f * st.tanfovx);
  a.focal_y = (float)st.H / (2.0f * st.tanfovy);
  const int kact = (st.sh_degree + 1) * (st.sh_degree + 1);
  a.rest_width_act = 3 * (kact - 1);
  a.rest_width_max = W;
  a.pitch = W | 1;
  const bool dense = !in.colors_precomp && !in.cov3D_precomp && (W == 0 || in.sh_rest_stride == W) && in.sh_dc_stride == 3 &&
                     (W * 4 * PBB_ROWS) % 16 == 0 && aligned16(in.means3D) && aligned16(in.scales) &&
                     aligned16(in.rotations) && aligned16(in.opacities) && (W == 0 || aligned16(in.sh_rest)) &&
                     aligned16(out.dL_dmeans2D) && aligned16(grad_geom);
  if (!dense) return cudaErrorInvalidValue;
  if (max_blk > 0) {
    cudaError_t e = cudaFuncSetAttribute(preprocess_backward_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         PBB_SMEM_BYTES);
    if (e != cudaSuccess) return e;
    preprocess_backward_bulk_kernel<<<G * max_blk, PBB_ROWS, PBB_SMEM_BYTES, stream>>>(st, a, 0, rt);
  }
  // tails: rows of shard q beyond its full blocks
  for (int q = 0; q < G; ++q) {
    const long long b = (long long)q * R;
    const long long rows = b >= N ? 0 : ((long long)N - b < R ? (long long)N - b : (long long)R);
    const long long t0 = b + (long long)rt.nblk[q] * PBB_ROWS, t1 = b + rows;
    if (t1 <= t0) continue;
    gs_gradients g = out;
    g.dL_dmeans3D = rt.vb[q][0];
    g.dL_dsh_dc = rt.vb[q][1];
    g.dL_dsh_rest = rt.vb[q][2];
    g.dL_dopacities = rt.vb[q][3];
    g.dL_dscales = rt.vb[q][4];
    g.dL_drotations = rt.vb[q][5];
    g.dL_dcov3D = g.dL_dcolors = nullptr;
    g.sh_dc_stride = 3;
    g.sh_rest_stride = W;
    g.accumulate = 0;
    PreBwdArgs at = a;
    at.out = g;
    at.N = (int)t1;
    const size_t smem = W > 0 ? (size_t)PB_THREADS * at.pitch * sizeof(float) : 0;
    const int blocks = (int)((t1 - t0 + PB_THREADS - 1) / PB_THREADS);
    preprocess_backward_kernel<<<blocks, PB_THREADS, smem, stream>>>(st, at, (int)t0);
  }
  return cudaGetLastError();
}
