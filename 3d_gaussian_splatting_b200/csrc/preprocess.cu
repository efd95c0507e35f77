// preprocess.cu — K1: per-Gaussian frustum cull, EWA projection, conic, radius, tile
// rectangle and SH -> RGB.  One thread per Gaussian.
//
// This translation unit is compiled with --fmad=false: every fp32 operation is rounded
// individually and written in the same order as the CPU oracle, so that radii,
// tiles_touched, depths (-> sort keys) and in practice all fp outputs are bit-identical.
// cov3D follows the reference's parametrisation: q=(r,x,y,z) normalised inside
// (src/utils/general_utils.cpp:12-40), L = R*diag(mod*s) (:88-99), Sigma = L*L^T packed
// [xx,xy,xz,yy,yz,zz] (src/scene/gaussian_model.cpp:18-28).
//
// Memory, measured path (preprocess_bulk_kernel): the reference's row-major (N,C) tensors are
// consumed as they are -- each 128-row block pulls its xyz / scale / rotation / opacity / SH-dc /
// SH-rest slices into shared memory with TMA bulk copies on one mbarrier (the AoS -> thread-row
// transposition happens there; 45- and 3-float row pitches are odd, hence bank-conflict free),
// and the 48-B geometry records, depths, radii, tile counts and clamp masks leave as bulk stores.
// Generic fallback (preprocess_kernel: strided / unaligned tensors, precomputed cov3D or colours,
// ragged tail): direct loads, SH rows staged through shared memory with coalesced loops.
#include "gs_common.cuh"
#include "radix_sort.cuh"
#include "tma_bulk.cuh"

#define PP_THREADS 256

struct PreprocessArgs {
  gs_gaussians in;
  int N;
  GsGeom* geom;
  float* depths;
  int* radii;
  int* radii_user;  // optional second copy straight into the caller's radii array (16-byte aligned), or nullptr
  uint32_t* tiles_touched;
  uint4* rects;  // {x0 | y0 << 16, x1 | y1 << 16, alive-tile mask lo, hi}; zeros when culled (mask unused when area > 64)
  uint8_t* clamped;
  float focal_x, focal_y;
  int gx, gy;
  int rest_width;  // active SH floats per Gaussian beyond dc: 3*((deg+1)^2-1)
  int rest_pitch;  // smem row pitch (odd)
  int defer_color; // 1: leave colour and clamp mask at zero, the SH rows are not read (sh_color_kernel fills them in later)
};

__device__ __forceinline__ void cov3d_from_scale_rot(const float s[3], float mod, const float q[4], float c6[6]) {
  float norm = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  float r = q[0] / norm, x = q[1] / norm, y = q[2] / norm, z = q[3] / norm;
  float R00 = 1.0f - 2.0f * (y * y + z * z);
  float R01 = 2.0f * (x * y - r * z);
  float R02 = 2.0f * (x * z + r * y);
  float R10 = 2.0f * (x * y + r * z);
  float R11 = 1.0f - 2.0f * (x * x + z * z);
  float R12 = 2.0f * (y * z - r * x);
  float R20 = 2.0f * (x * z - r * y);
  float R21 = 2.0f * (y * z + r * x);
  float R22 = 1.0f - 2.0f * (x * x + y * y);
  float s0 = mod * s[0], s1 = mod * s[1], s2 = mod * s[2];
  float L00 = R00 * s0, L01 = R01 * s1, L02 = R02 * s2;
  float L10 = R10 * s0, L11 = R11 * s1, L12 = R12 * s2;
  float L20 = R20 * s0, L21 = R21 * s1, L22 = R22 * s2;
  c6[0] = L00 * L00 + L01 * L01 + L02 * L02;
  c6[1] = L00 * L10 + L01 * L11 + L02 * L12;
  c6[2] = L00 * L20 + L01 * L21 + L02 * L22;
  c6[3] = L10 * L10 + L11 * L11 + L12 * L12;
  c6[4] = L10 * L20 + L11 * L21 + L12 * L22;
  c6[5] = L20 * L20 + L21 * L21 + L22 * L22;
}

__device__ __forceinline__ void cov2d_ewa(const float t[3], float fx, float fy, float tanfovx, float tanfovy,
                                          const float c6[6], const float* V, float abc[3]) {
  float limx = GS_FOV_GUARD * tanfovx, limy = GS_FOV_GUARD * tanfovy;
  float txtz = t[0] / t[2], tytz = t[1] / t[2];
  float tx = fminf(limx, fmaxf(-limx, txtz)) * t[2];
  float ty = fminf(limy, fmaxf(-limy, tytz)) * t[2];
  float tz = t[2];
  float J00 = fx / tz, J02 = -(fx * tx) / (tz * tz);
  float J11 = fy / tz, J12 = -(fy * ty) / (tz * tz);
  float T00 = J00 * V[0] + J02 * V[2];
  float T01 = J00 * V[4] + J02 * V[6];
  float T02 = J00 * V[8] + J02 * V[10];
  float T10 = J11 * V[1] + J12 * V[2];
  float T11 = J11 * V[5] + J12 * V[6];
  float T12 = J11 * V[9] + J12 * V[10];
  float c0 = c6[0], c1 = c6[1], c2 = c6[2], c3 = c6[3], c4 = c6[4], c5 = c6[5];
  float U00 = c0 * T00 + c1 * T01 + c2 * T02;
  float U01 = c1 * T00 + c3 * T01 + c4 * T02;
  float U02 = c2 * T00 + c4 * T01 + c5 * T02;
  float U10 = c0 * T10 + c1 * T11 + c2 * T12;
  float U11 = c1 * T10 + c3 * T11 + c4 * T12;
  float U12 = c2 * T10 + c4 * T11 + c5 * T12;
  abc[0] = T00 * U00 + T01 * U01 + T02 * U02;
  abc[1] = T00 * U10 + T01 * U11 + T02 * U12;
  abc[2] = T10 * U10 + T11 * U11 + T12 * U12;
}

// s points at this Gaussian's staged "rest" row: s[3*(k-1)+c] is coefficient k, channel c.
__device__ __forceinline__ void sh_to_rgb(int deg, const float p[3], const float campos[3], const float dc[3],
                                          const float* s, float rgb[3], uint32_t& clamp_mask) {
  float dx = p[0] - campos[0], dy = p[1] - campos[1], dz = p[2] - campos[2];
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float x = dx / len, y = dy / len, z = dz / len;
  clamp_mask = 0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float res = GS_SH_C0 * dc[c];
    if (deg > 0) {
      res = res - GS_SH_C1 * y * s[0 + c] + GS_SH_C1 * z * s[3 + c] - GS_SH_C1 * x * s[6 + c];
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        res = res + GS_SH_C2_0 * xy * s[9 + c] + GS_SH_C2_1 * yz * s[12 + c] +
              GS_SH_C2_2 * (2.0f * zz - xx - yy) * s[15 + c] + GS_SH_C2_3 * xz * s[18 + c] +
              GS_SH_C2_4 * (xx - yy) * s[21 + c];
        if (deg > 2) {
          res = res + GS_SH_C3_0 * y * (3.0f * xx - yy) * s[24 + c] + GS_SH_C3_1 * xy * z * s[27 + c] +
                GS_SH_C3_2 * y * (4.0f * zz - xx - yy) * s[30 + c] +
                GS_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * s[33 + c] +
                GS_SH_C3_4 * x * (4.0f * zz - xx - yy) * s[36 + c] + GS_SH_C3_5 * z * (xx - yy) * s[39 + c] +
                GS_SH_C3_6 * x * (xx - 3.0f * yy) * s[42 + c];
        }
      }
    }
    res = res + 0.5f;
    if (res < 0.0f) clamp_mask |= (1u << c);
    rgb[c] = fmaxf(res, 0.0f);
  }
}

struct PpOut {
  GsGeom g;
  float depth;
  int radius;
  uint32_t tiles, cmask;
  uint4 rect;
  GsCull cs;  // cull parameters of a visible Gaussian (deferred, warp-cooperative cull of the bulk kernel)
};

// One Gaussian.  p / scl / q4 / opac_in as stored by the caller (raw when a.in.raw_params); cov_pre, col_pre:
// precomputed cov3D / colour rows or nullptr; dcp: the three SH dc floats; srow: the staged SH "rest" row.
// kInlineCull: evaluate the exact tile cull here, one thread looping over its rectangle (generic kernel); otherwise
// tiles / rect mask are left as "all tiles" and the caller culls cooperatively (pp_warp_cull).
template <bool kInlineCull>
__device__ __forceinline__ PpOut pp_compute(const gs_settings& st, const PreprocessArgs& a, const float p[3],
                                            const float scl[3], const float4 q4, const float opac_in,
                                            const float* cov_pre, const float* col_pre, const float* dcp,
                                            const float* srow) {
  GsGeom g;
  g.q0 = make_float4(0.f, 0.f, 0.f, 0.f);
  g.q1 = g.q0;
  g.q2 = g.q0;
  int rad_out = 0;
  uint32_t tiles = 0, cmask = 0;
  uint4 rect = make_uint4(0u, 0u, 0u, 0u);
  float depth = 0.f;
  GsCull cs_out;
  cs_out.hA = cs_out.B = cs_out.hC = cs_out.rA = cs_out.rC = cs_out.thr = 0.f;

  const float* V = st.viewmatrix;
  const float* M = st.projmatrix;
  float pv[3];
  pv[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
  pv[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
  pv[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
  bool ok = pv[2] > GS_NEAR_Z;
  if (ok) {
    float ph[4];
    ph[0] = M[0] * p[0] + M[4] * p[1] + M[8] * p[2] + M[12];
    ph[1] = M[1] * p[0] + M[5] * p[1] + M[9] * p[2] + M[13];
    ph[2] = M[2] * p[0] + M[6] * p[1] + M[10] * p[2] + M[14];
    ph[3] = M[3] * p[0] + M[7] * p[1] + M[11] * p[2] + M[15];
    float pw = 1.0f / (ph[3] + GS_W_EPS);
    float ndc_x = ph[0] * pw, ndc_y = ph[1] * pw;

    float c6[6];
    if (cov_pre) {
#pragma unroll
      for (int c = 0; c < 6; ++c) c6[c] = cov_pre[c];
    } else {
      float s[3] = {scl[0], scl[1], scl[2]};
      float q[4] = {q4.x, q4.y, q4.z, q4.w};
      if (a.in.raw_params) {
        s[0] = gs_exp_any(s[0]);
        s[1] = gs_exp_any(s[1]);
        s[2] = gs_exp_any(s[2]);
      }
      cov3d_from_scale_rot(s, st.scale_modifier, q, c6);
    }
    float abc[3];
    cov2d_ewa(pv, a.focal_x, a.focal_y, st.tanfovx, st.tanfovy, c6, V, abc);
    float ca = abc[0] + GS_DILATION, cb = abc[1], cc = abc[2] + GS_DILATION;
    float det = ca * cc - cb * cb;
    ok = det > 0.0f;
    if (ok) {
      float det_inv = 1.0f / det;
      float con_x = cc * det_inv, con_y = -cb * det_inv, con_z = ca * det_inv;
      float mid = 0.5f * (ca + cc);
      float disc = sqrtf(fmaxf(GS_LAMBDA_FLOOR, mid * mid - det));
      float lambda1 = mid + disc, lambda2 = mid - disc;
      float rad_f = ceilf(3.0f * sqrtf(fmaxf(lambda1, lambda2)));
      float px = ((ndc_x + 1.0f) * (float)st.W - 1.0f) * 0.5f;
      float py = ((ndc_y + 1.0f) * (float)st.H - 1.0f) * 0.5f;
      ok = (rad_f < 1e9f) && (fabsf(px) < 1e9f) && (fabsf(py) < 1e9f);
      if (ok) {
        int rad = (int)rad_f;
        int x0 = (int)((px - (float)rad) / (float)GS_TILE);
        int y0 = (int)((py - (float)rad) / (float)GS_TILE);
        int x1 = (int)((px + (float)rad + (float)(GS_TILE - 1)) / (float)GS_TILE);
        int y1 = (int)((py + (float)rad + (float)(GS_TILE - 1)) / (float)GS_TILE);
        x0 = min(a.gx, max(0, x0));
        x1 = min(a.gx, max(0, x1));
        y0 = min(a.gy, max(0, y0));
        y1 = min(a.gy, max(0, y1));
        int area = (x1 - x0) * (y1 - y0);
        if (area > 0) {
          float rgb[3];
          if (col_pre) {
            rgb[0] = col_pre[0];
            rgb[1] = col_pre[1];
            rgb[2] = col_pre[2];
          } else if (a.defer_color) {
            rgb[0] = rgb[1] = rgb[2] = 0.f;  // sh_color_kernel: same sh_to_rgb, same translation unit, same bits
          } else {
            const float dc[3] = {dcp[0], dcp[1], dcp[2]};
            sh_to_rgb(st.sh_degree, p, st.campos, dc, srow, rgb, cmask);
          }
          float opac = opac_in;
          if (a.in.raw_params) opac = gs_sigmoid(opac);
          // exact tile culling (gs_common.cuh): which tiles of the rectangle can receive alpha >= 1/255
          uint32_t mlo = 0xFFFFFFFFu, mhi = 0xFFFFFFFFu;
          const GsCull cs = gs_cull_setup(0.5f * con_x, con_y, 0.5f * con_z, opac);
          cs_out = cs;
          if (kInlineCull && area <= GS_CULL_MAX_TILES) {
            mlo = mhi = 0u;
            int bit = 0;
            for (int ty = y0; ty < y1; ++ty)
              for (int tx = x0; tx < x1; ++tx, ++bit)
                if (gs_tile_alive(cs, px, py, tx, ty)) {
                  if (bit < 32) mlo |= 1u << bit;
                  else mhi |= 1u << (bit - 32);
                }
            area = __popc(mlo) + __popc(mhi);
          }
          g.q0 = make_float4(px, py, 0.5f * con_x, con_y);
          g.q1 = make_float4(0.5f * con_z, opac, rgb[0], rgb[1]);
          g.q2 = make_float4(rgb[2], pv[2], cs.thr, 0.f);
          rad_out = rad;
          tiles = (uint32_t)area;
          rect = make_uint4((uint32_t)x0 | ((uint32_t)y0 << 16), (uint32_t)x1 | ((uint32_t)y1 << 16),
                            mlo, mhi);
          depth = pv[2];
        }
      }
    }
  }
  PpOut o;
  o.g = g;
  o.depth = depth;
  o.radius = rad_out;
  o.tiles = tiles;
  o.cmask = cmask;
  o.rect = rect;
  o.cs = cs_out;
  return o;
}

// Exact tile cull of one warp's Gaussians, load-balanced: the candidate (Gaussian, tile) pairs of the 32 lanes are
// flattened (exclusive scan of the rectangle areas), lane l tests candidates l, l+32, ...: it evaluates gs_tile_alive
// with the owner's parameters and ORs the result into the owner's 64-bit mask.  Same decisions as the per-thread loop
// (same function, same inputs), without the divergence of 32 different rectangle sizes.
// Owner of a candidate, without a search: the lanes that have candidates store their records compacted (slot = rank
// among those lanes); per batch of 32 candidates one REDUX.OR collects the bit mask of the offsets that start inside
// the batch, and the owner's slot is (owners started before the batch) + popc(starts at or below my position) - 1.
// (The first version's 5-step binary search in shared memory was 20% of K1's stall samples.)
// sw: this warp's scratch (32 slots x PP_CULL_WORDS words).
#define PP_CULL_WORDS 15  // 14 used; odd pitch: conflict-free owner look-ups
__device__ __forceinline__ void pp_warp_cull(PpOut& o, uint32_t* sw) {
  const int lane = threadIdx.x & 31;
  const int x0 = (int)(o.rect.x & 0xFFFFu), y0 = (int)(o.rect.x >> 16);
  const int w = (int)(o.rect.y & 0xFFFFu) - x0, h = (int)(o.rect.y >> 16) - y0;
  const int area = w * h;
  const bool cand = area > 0 && area <= GS_CULL_MAX_TILES;
  const int cnt = cand ? area : 0;
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  const int off = incl - cnt;
  const uint32_t owners = __ballot_sync(0xffffffffu, cand);
  uint32_t* me = sw + __popc(owners & ((1u << lane) - 1u)) * PP_CULL_WORDS;
  if (cand) {
    me[0] = (uint32_t)off;
    me[1] = __float_as_uint(o.g.q0.x);
    me[2] = __float_as_uint(o.g.q0.y);
    me[3] = __float_as_uint(o.cs.hA);
    me[4] = __float_as_uint(o.cs.B);
    me[5] = __float_as_uint(o.cs.hC);
    me[6] = __float_as_uint(o.cs.rA);
    me[7] = __float_as_uint(o.cs.rC);
    me[8] = __float_as_uint(o.cs.thr);
    me[9] = (uint32_t)x0 | ((uint32_t)y0 << 16);
    me[10] = (uint32_t)w;
    me[11] = __float_as_uint(1.0f / (float)w);
    me[12] = 0u;
    me[13] = 0u;
  }
  __syncwarp();
  int started = 0;  // owners whose first candidate lies before the current batch
  for (int cb = 0; cb < total; cb += 32) {
    const uint32_t rel = (uint32_t)(off - cb);  // wraps for offsets below cb
    const uint32_t starts = __reduce_or_sync(0xffffffffu, (cand && rel < 32u) ? 1u << rel : 0u);
    const int c = cb + lane;
    if (c < total) {
      uint32_t* ow = sw + (started + __popc(starts & (0xFFFFFFFFu >> (31 - lane))) - 1) * PP_CULL_WORDS;
      const int k = c - (int)ow[0];
      const int ww = (int)ow[10];
      const int row = __float2int_rd(((float)k + 0.5f) * __uint_as_float(ow[11]));  // exact for k, ww <= 64
      const int col = k - row * ww;
      GsCull cs;
      cs.hA = __uint_as_float(ow[3]);
      cs.B = __uint_as_float(ow[4]);
      cs.hC = __uint_as_float(ow[5]);
      cs.rA = __uint_as_float(ow[6]);
      cs.rC = __uint_as_float(ow[7]);
      cs.thr = __uint_as_float(ow[8]);
      const int tx = (int)(ow[9] & 0xFFFFu) + col, ty = (int)(ow[9] >> 16) + row;
      if (gs_tile_alive(cs, __uint_as_float(ow[1]), __uint_as_float(ow[2]), tx, ty)) atomicOr(&ow[12 + (k >> 5)], 1u << (k & 31));
    }
    started += __popc(starts);
  }
  __syncwarp();
  if (cand) {
    o.rect.z = me[12];
    o.rect.w = me[13];
    o.tiles = (uint32_t)(__popc(me[12]) + __popc(me[13]));
  }
  __syncwarp();
}

__global__ void __launch_bounds__(PP_THREADS)
preprocess_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ PreprocessArgs a, const int first) {
  extern __shared__ float s_sh[];
  const int tid = threadIdx.x;
  const int row0 = first + blockIdx.x * PP_THREADS;
  const int i = row0 + tid;
  const int rows = min(PP_THREADS, a.N - row0);

  // ---- stage the SH "rest" rows of this block: coalesced global reads, odd-pitch smem ----
  if (a.rest_width > 0 && a.in.colors_precomp == nullptr && !a.defer_color) {
    const int stride = a.in.sh_rest_stride;
    const float* base = a.in.sh_rest + (size_t)row0 * stride;
    const int total = (rows - 1) * stride + a.rest_width;  // never read past the last row
    int row = tid / stride, col = tid % stride;
    const int drow = PP_THREADS / stride, dcol = PP_THREADS % stride;
    for (int f = tid; f < total; f += PP_THREADS) {
      if (col < a.rest_width) s_sh[row * a.rest_pitch + col] = __ldg(base + f);
      row += drow;
      col += dcol;
      if (col >= stride) { col -= stride; ++row; }
    }
  }
  __syncthreads();
  if (i >= a.N) return;

  const float p[3] = {a.in.means3D[3 * i], a.in.means3D[3 * i + 1], a.in.means3D[3 * i + 2]};
  const float* cov_pre = a.in.cov3D_precomp ? a.in.cov3D_precomp + 6 * (size_t)i : nullptr;
  const float* col_pre = a.in.colors_precomp ? a.in.colors_precomp + 3 * (size_t)i : nullptr;
  float scl[3] = {0.f, 0.f, 0.f};
  float4 q4 = make_float4(1.f, 0.f, 0.f, 0.f);
  if (!cov_pre) {
    scl[0] = a.in.scales[3 * i];
    scl[1] = a.in.scales[3 * i + 1];
    scl[2] = a.in.scales[3 * i + 2];
    q4 = *reinterpret_cast<const float4*>(a.in.rotations + 4 * (size_t)i);
  }
  const float* dcp = col_pre ? nullptr : a.in.sh_dc + (size_t)i * a.in.sh_dc_stride;
  const PpOut o = pp_compute<true>(st, a, p, scl, q4, a.in.opacities[i], cov_pre, col_pre, dcp, s_sh + tid * a.rest_pitch);
  a.geom[i] = o.g;
  a.depths[i] = o.depth;
  a.radii[i] = o.radius;
  if (a.radii_user) a.radii_user[i] = o.radius;
  a.tiles_touched[i] = o.tiles;
  a.rects[i] = o.rect;
  a.clamped[i] = (uint8_t)o.cmask;
}

// ---- TMA-staged path ---------------------------------------------------------------------------
#define PPB_ROWS 128
#define PPB_OFF_SH 0                               // rows x 45 floats
#define PPB_OFF_XYZ (PPB_OFF_SH + PPB_ROWS * 45)   // rows x 3
#define PPB_OFF_SCL (PPB_OFF_XYZ + PPB_ROWS * 3)   // rows x 3
#define PPB_OFF_ROT (PPB_OFF_SCL + PPB_ROWS * 3)   // rows x 4
#define PPB_OFF_OP (PPB_OFF_ROT + PPB_ROWS * 4)    // rows x 1
#define PPB_OFF_DC (PPB_OFF_OP + PPB_ROWS)         // rows x 3
#define PPB_OFF_GEOM (PPB_OFF_DC + PPB_ROWS * 3)   // rows x 12 (out)
#define PPB_OFF_DEPTH (PPB_OFF_GEOM + PPB_ROWS * 12)
#define PPB_OFF_RADII (PPB_OFF_DEPTH + PPB_ROWS)
#define PPB_OFF_TILES (PPB_OFF_RADII + PPB_ROWS)
#define PPB_OFF_RECT (PPB_OFF_TILES + PPB_ROWS)    // rows x 4
#define PPB_OFF_CLAMP (PPB_OFF_RECT + PPB_ROWS * 4) // rows x 1 byte
#define PPB_FLOATS (PPB_OFF_CLAMP + PPB_ROWS / 4)
#define PPB_SMEM_BYTES (PPB_FLOATS * 4 + 16)

__global__ void __launch_bounds__(PPB_ROWS)
preprocess_bulk_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ PreprocessArgs a) {
  extern __shared__ __align__(128) float s_buf[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_buf + PPB_FLOATS);
  const int tid = threadIdx.x;
  const size_t row0 = (size_t)blockIdx.x * PPB_ROWS;
  const int W = a.in.sh_rest_stride;  // dense rows (0 when the model has no higher-order SH)

  if (tid == 0) {
    tma::mbar_init(bar, 1);
    tma::fence_mbar_init();
    const uint32_t sh_bytes = (a.rest_width > 0 && !a.defer_color) ? (uint32_t)(PPB_ROWS * W * 4) : 0u;
    const uint32_t dc_bytes = a.defer_color ? 0u : PPB_ROWS * 12u;
    tma::mbar_arrive_expect_tx(bar, sh_bytes + dc_bytes + PPB_ROWS * (12u + 12u + 16u + 4u));
    tma::load(s_buf + PPB_OFF_XYZ, a.in.means3D + 3 * row0, PPB_ROWS * 12u, bar);
    tma::load(s_buf + PPB_OFF_SCL, a.in.scales + 3 * row0, PPB_ROWS * 12u, bar);
    tma::load(s_buf + PPB_OFF_ROT, a.in.rotations + 4 * row0, PPB_ROWS * 16u, bar);
    tma::load(s_buf + PPB_OFF_OP, a.in.opacities + row0, PPB_ROWS * 4u, bar);
    if (dc_bytes) tma::load(s_buf + PPB_OFF_DC, a.in.sh_dc + 3 * row0, dc_bytes, bar);
    if (sh_bytes) tma::load(s_buf + PPB_OFF_SH, a.in.sh_rest + row0 * W, sh_bytes, bar);
  }
  __syncthreads();
  tma::mbar_wait(bar, 0);

  const float p[3] = {s_buf[PPB_OFF_XYZ + 3 * tid], s_buf[PPB_OFF_XYZ + 3 * tid + 1], s_buf[PPB_OFF_XYZ + 3 * tid + 2]};
  const float scl[3] = {s_buf[PPB_OFF_SCL + 3 * tid], s_buf[PPB_OFF_SCL + 3 * tid + 1], s_buf[PPB_OFF_SCL + 3 * tid + 2]};
  const float4 q4 = reinterpret_cast<const float4*>(s_buf + PPB_OFF_ROT)[tid];
  PpOut o = pp_compute<false>(st, a, p, scl, q4, s_buf[PPB_OFF_OP + tid], nullptr, nullptr, s_buf + PPB_OFF_DC + 3 * tid,
                              s_buf + PPB_OFF_SH + tid * W);
  // the (not yet written) output slices serve as per-warp scratch of the cooperative tile cull
  static_assert(4 * 32 * PP_CULL_WORDS <= PPB_OFF_CLAMP - PPB_OFF_GEOM, "cull scratch does not fit the output slices");
  pp_warp_cull(o, reinterpret_cast<uint32_t*>(s_buf + PPB_OFF_GEOM) + (tid >> 5) * 32 * PP_CULL_WORDS);
  __syncthreads();
  float4* sg = reinterpret_cast<float4*>(s_buf + PPB_OFF_GEOM) + 3 * tid;
  sg[0] = o.g.q0;
  sg[1] = o.g.q1;
  sg[2] = o.g.q2;
  s_buf[PPB_OFF_DEPTH + tid] = o.depth;
  reinterpret_cast<int*>(s_buf)[PPB_OFF_RADII + tid] = o.radius;
  reinterpret_cast<uint32_t*>(s_buf)[PPB_OFF_TILES + tid] = o.tiles;
  reinterpret_cast<uint4*>(s_buf + PPB_OFF_RECT)[tid] = o.rect;
  reinterpret_cast<uint8_t*>(s_buf + PPB_OFF_CLAMP)[tid] = (uint8_t)o.cmask;
  tma::fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) {
    tma::store(a.geom + row0, s_buf + PPB_OFF_GEOM, PPB_ROWS * 48u);
    tma::store(a.depths + row0, s_buf + PPB_OFF_DEPTH, PPB_ROWS * 4u);
    tma::store(a.radii + row0, s_buf + PPB_OFF_RADII, PPB_ROWS * 4u);
    if (a.radii_user) tma::store(a.radii_user + row0, s_buf + PPB_OFF_RADII, PPB_ROWS * 4u);
    tma::store(a.tiles_touched + row0, s_buf + PPB_OFF_TILES, PPB_ROWS * 4u);
    tma::store(a.rects + row0, s_buf + PPB_OFF_RECT, PPB_ROWS * 16u);
    tma::store(a.clamped + row0, s_buf + PPB_OFF_CLAMP, PPB_ROWS);
    tma::commit_group();
    tma::wait_group_read0();
  }
}


// ---- depth keys ahead of K1 ---------------------------------------------------------------------
// The depth argsort needs nothing but the view depth, 12 of the 236 B a Gaussian occupies, so it does not have to
// wait for K1: this kernel evaluates pv[2] with the expression pp_compute uses (same translation unit, no
// contraction: the same bits) and the sort runs beside K1 on the context's second stream (capi.cu).  A Gaussian
// behind the near plane gets key 0 like in K1; one that K1 culls later (degenerate covariance, empty rectangle)
// keeps its depth here -- it emits no (tile, Gaussian) pair, so its place in the order is immaterial.
// The kernel also takes the digit histograms of all four radix passes of the depth argsort (a digit histogram does
// not depend on the order of the keys, so the one-sweep passes need no histogram launches of their own,
// radix_sort.cuh): per-block shared-memory counters, then one global atomic per non-empty bin.  The two high bytes
// of a depth are nearly constant over a scene, so a warp whose 32 keys agree on a digit adds 32 with one atomic.
#define DK_ITEMS 8
__global__ void __launch_bounds__(256)
depth_key_kernel(const __grid_constant__ gs_settings st, const float* __restrict__ means3D, int N,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ hist /* [4][256] */) {
  __shared__ uint32_t s_hist[4][256];
  const int tid = threadIdx.x, lane = tid & 31;
#pragma unroll
  for (int d = 0; d < 4; ++d) s_hist[d][tid] = 0;
  __syncthreads();
  const float* V = st.viewmatrix;
  const int base = blockIdx.x * (256 * DK_ITEMS);
#pragma unroll
  for (int k = 0; k < DK_ITEMS; ++k) {
    const int i = base + k * 256 + tid;
    const bool valid = i < N;
    uint32_t key = 0;
    if (valid) {
      const float p[3] = {__ldg(means3D + 3 * (size_t)i), __ldg(means3D + 3 * (size_t)i + 1), __ldg(means3D + 3 * (size_t)i + 2)};
      const float z = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
      key = z > GS_NEAR_Z ? __float_as_uint(z) : 0u;
      keys[i] = key;
    }
    if (valid) {
      atomicAdd(&s_hist[0][key & 0xFFu], 1u);
      atomicAdd(&s_hist[1][(key >> 8) & 0xFFu], 1u);
    }
    warp_run_hist(s_hist[2], (key >> 16) & 0xFFu, valid, lane);  // (i increases with the lane: valid lanes are a prefix)
    warp_run_hist(s_hist[3], key >> 24, valid, lane);
  }
  __syncthreads();
#pragma unroll
  for (int d = 0; d < 4; ++d) {
    const uint32_t c = s_hist[d][tid];
    if (c) atomicAdd(hist + d * 256 + tid, c);
  }
}

cudaError_t gs_launch_depth_keys(const gs_settings& st, const float* means3D, int N, uint32_t* keys, uint32_t* hist,
                                 cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  depth_key_kernel<<<(N + 256 * DK_ITEMS - 1) / (256 * DK_ITEMS), 256, 0, stream>>>(st, means3D, N, keys, hist);
  return cudaGetLastError();
}

static bool pp_aligned16(const void* p) { return (((uintptr_t)p) & 15) == 0; }

// ---- deferred colour (multi-GPU step, capi.cu: gs_set_color_gate) -------------------------------------------------
// With a colour gate the forward runs K1 WITHOUT the SH -> RGB evaluation (defer_color: 44 of a Gaussian's 236 B are
// read), bins and sorts, and only then -- after the gate event, i.e. once the peers' optimiser kernels have delivered
// the new SH coefficients into this replica over NVLink -- evaluates the colours with this kernel, just ahead of the
// forward blend.  The all-gather of the SH coefficients (81 % of the parameter bytes at degree 3) thereby overlaps
// K1, the depth argsort, the duplicate stage and the tile sort of the next iteration.  Same sh_to_rgb, same
// translation unit (no contraction): colours and clamp masks are bit-identical to the undeferred K1.
// A Gaussian gets a colour iff K1 gave it a radius (K1 evaluates the colour under exactly that condition).
#define SHC_OFF_SH 0                              // rows x 45 floats
#define SHC_OFF_XYZ (SHC_OFF_SH + PPB_ROWS * 45)  // rows x 3
#define SHC_OFF_DC (SHC_OFF_XYZ + PPB_ROWS * 3)   // rows x 3
#define SHC_FLOATS (SHC_OFF_DC + PPB_ROWS * 3)
#define SHC_SMEM_BYTES (SHC_FLOATS * 4 + 16)
struct ShColorArgs {
  gs_gaussians in;
  int N, first;
  const int* radii;
  GsGeom* geom;
  uint8_t* clamped;
  int rest_width;
};

__device__ __forceinline__ void shc_store(const ShColorArgs& a, size_t i, const float rgb[3], uint32_t cmask) {
  float* rec = reinterpret_cast<float*>(a.geom + i);  // {x,y,hA,B} {hC,opacity,r,g} {b,depth,thr,-}
  *reinterpret_cast<float2*>(rec + 6) = make_float2(rgb[0], rgb[1]);
  rec[8] = rgb[2];
  a.clamped[i] = (uint8_t)cmask;
}

__global__ void __launch_bounds__(PPB_ROWS)
sh_color_bulk_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ ShColorArgs a) {
  extern __shared__ __align__(128) float s_buf[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_buf + SHC_FLOATS);
  const int tid = threadIdx.x;
  const size_t row0 = (size_t)blockIdx.x * PPB_ROWS;
  const int W = a.in.sh_rest_stride;
  // (the rows are requested before the radii are known: one memory latency per block, as in K1, which cannot know
  //  the visibility before it has read a row either)
  if (tid == 0) {
    tma::mbar_init(bar, 1);
    tma::fence_mbar_init();
    const uint32_t sh_bytes = a.rest_width > 0 ? (uint32_t)(PPB_ROWS * W * 4) : 0u;
    tma::mbar_arrive_expect_tx(bar, sh_bytes + PPB_ROWS * (12u + 12u));
    tma::load(s_buf + SHC_OFF_XYZ, a.in.means3D + 3 * row0, PPB_ROWS * 12u, bar);
    tma::load(s_buf + SHC_OFF_DC, a.in.sh_dc + 3 * row0, PPB_ROWS * 12u, bar);
    if (sh_bytes) tma::load(s_buf + SHC_OFF_SH, a.in.sh_rest + row0 * W, sh_bytes, bar);
  }
  const int radius = a.radii[row0 + tid];
  __syncthreads();
  tma::mbar_wait(bar, 0);
  if (radius > 0) {
    const float p[3] = {s_buf[SHC_OFF_XYZ + 3 * tid], s_buf[SHC_OFF_XYZ + 3 * tid + 1], s_buf[SHC_OFF_XYZ + 3 * tid + 2]};
    const float dc[3] = {s_buf[SHC_OFF_DC + 3 * tid], s_buf[SHC_OFF_DC + 3 * tid + 1], s_buf[SHC_OFF_DC + 3 * tid + 2]};
    float rgb[3];
    uint32_t cmask = 0;
    sh_to_rgb(st.sh_degree, p, st.campos, dc, s_buf + SHC_OFF_SH + tid * W, rgb, cmask);
    shc_store(a, row0 + tid, rgb, cmask);
  }
}

// generic rows (ragged tail, strided or unaligned tensors): one thread per Gaussian, SH row read in place
__global__ void __launch_bounds__(PPB_ROWS)
sh_color_kernel(const __grid_constant__ gs_settings st, const __grid_constant__ ShColorArgs a) {
  const size_t i = (size_t)a.first + (size_t)blockIdx.x * PPB_ROWS + threadIdx.x;
  if (i >= (size_t)a.N || a.radii[i] <= 0) return;
  const float p[3] = {a.in.means3D[3 * i], a.in.means3D[3 * i + 1], a.in.means3D[3 * i + 2]};
  const float* dcp = a.in.sh_dc + i * a.in.sh_dc_stride;
  const float dc[3] = {dcp[0], dcp[1], dcp[2]};
  float srow[45];
#pragma unroll
  for (int k = 0; k < 45; ++k) srow[k] = k < a.rest_width ? a.in.sh_rest[i * a.in.sh_rest_stride + k] : 0.f;
  float rgb[3];
  uint32_t cmask = 0;
  sh_to_rgb(st.sh_degree, p, st.campos, dc, srow, rgb, cmask);
  shc_store(a, i, rgb, cmask);
}

static bool pp_dense_layout(const gs_settings& st, const gs_gaussians& in, int rest_width) {
  const int kmax = (st.max_sh_degree + 1) * (st.max_sh_degree + 1);
  const int Wmax = 3 * (kmax - 1);
  return !in.colors_precomp && !in.cov3D_precomp && in.sh_dc_stride == 3 &&
         (Wmax == 0 || in.sh_rest_stride == Wmax) && (rest_width == 0 || in.sh_rest_stride <= 45) &&
         pp_aligned16(in.means3D) && pp_aligned16(in.scales) && pp_aligned16(in.rotations) &&
         pp_aligned16(in.opacities) && pp_aligned16(in.sh_dc) && (rest_width == 0 || pp_aligned16(in.sh_rest));
}

cudaError_t gs_launch_sh_color(const gs_settings& st, const gs_gaussians& in, int N, const int* radii, GsGeom* geom,
                               uint8_t* clamped, cudaStream_t stream) {
  if (N == 0 || in.colors_precomp) return cudaSuccess;
  ShColorArgs a;
  a.in = in;
  a.N = N;
  a.first = 0;
  a.radii = radii;
  a.geom = geom;
  a.clamped = clamped;
  const int kact = (st.sh_degree + 1) * (st.sh_degree + 1);
  a.rest_width = 3 * (kact - 1);
  if (pp_dense_layout(st, in, a.rest_width) && N >= PPB_ROWS) {
    static bool attr_set = false;
    if (!attr_set) {
      cudaError_t e =
          cudaFuncSetAttribute(sh_color_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SHC_SMEM_BYTES);
      if (e != cudaSuccess) return e;
      attr_set = true;
    }
    const int blocks = N / PPB_ROWS;
    sh_color_bulk_kernel<<<blocks, PPB_ROWS, SHC_SMEM_BYTES, stream>>>(st, a);
    a.first = blocks * PPB_ROWS;
  }
  if (a.first < N) sh_color_kernel<<<(N - a.first + PPB_ROWS - 1) / PPB_ROWS, PPB_ROWS, 0, stream>>>(st, a);
  return cudaGetLastError();
}

// defer_color: K1 without the SH -> RGB evaluation (colours and clamp masks stay zero until gs_launch_sh_color)
cudaError_t gs_launch_preprocess(const gs_settings& st, const gs_gaussians& in, int N, GsGeom* geom, float* depths,
                                 int* radii, uint32_t* tiles_touched, uint4* rects, uint8_t* clamped,
                                 int* radii_user, int defer_color, cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  PreprocessArgs a;
  a.in = in;
  a.defer_color = (defer_color && !in.colors_precomp) ? 1 : 0;
  a.N = N;
  a.geom = geom;
  a.depths = depths;
  a.radii = radii;
  a.radii_user = radii_user;
  a.tiles_touched = tiles_touched;
  a.rects = rects;
  a.clamped = clamped;
  a.focal_x = (float)st.W / (2.0f * st.tanfovx);
  a.focal_y = (float)st.H / (2.0f * st.tanfovy);
  a.gx = (st.W + GS_TILE - 1) / GS_TILE;
  a.gy = (st.H + GS_TILE - 1) / GS_TILE;
  const int kact = (st.sh_degree + 1) * (st.sh_degree + 1);
  a.rest_width = in.colors_precomp ? 0 : 3 * (kact - 1);
  a.rest_pitch = a.rest_width | 1;
  size_t smem = (size_t)PP_THREADS * a.rest_pitch * sizeof(float);
  if (a.rest_width == 0) smem = 0;
  const bool dense = pp_dense_layout(st, in, a.rest_width);
  int first = 0;
  if (dense && N >= PPB_ROWS) {
    static bool attr_set = false;
    if (!attr_set) {
      cudaError_t e =
          cudaFuncSetAttribute(preprocess_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PPB_SMEM_BYTES);
      if (e != cudaSuccess) return e;
      attr_set = true;
    }
    const int blocks = N / PPB_ROWS;
    preprocess_bulk_kernel<<<blocks, PPB_ROWS, PPB_SMEM_BYTES, stream>>>(st, a);
    first = blocks * PPB_ROWS;
  }
  if (first < N) {
    const int blocks = (N - first + PP_THREADS - 1) / PP_THREADS;
    preprocess_kernel<<<blocks, PP_THREADS, smem, stream>>>(st, a, first);
  }
  return cudaGetLastError();
}
