/*
 * oracle/orc_densify.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Densify / prune / opacity reset with optimiser-state surgery: the consumers of the statistics the
 * reference accumulates (max_radii2D_, xyz_gradient_accum_, denom_, src/scene/gaussian_model.h:18-20;
 * percent_dense_ :44) under the schedule parameters it parses (src/arguments/params.h:61-67:
 * densification_interval_, opacity_reset_interval_, densify_from_iter_, densify_until_iter_,
 * densify_grad_threshold_).  The reference declares the state and the parameters but not the step itself
 * (SURVEY.md section 8f-2), so this restates the PUBLISHED algorithm of the 3D Gaussian Splatting trainer the
 * reference transliterates (Kerbl et al. 2023, gaussian_model.py: densify_and_prune / densify_and_clone /
 * densify_and_split / prune_points / reset_opacity), phase by phase and literally -- including two of its
 * quirks: (1) the statistics are zeroed by the clone phase's postfix, so the screen-size pruning test sees
 * max_radii2D == 0 and never fires; (2) the split children are appended as [all first copies, all second
 * copies].  The CUDA kernels do the same in one compaction pass; this file is their specification.
 *
 * Parity unpinned against a run of the reference (it has no such code); pinned against a numpy restatement of
 * the published algorithm in tests/test_oracle_densify.py.
 *
 * SPECIFIED RNG for the split samples (the published code calls torch.normal): counter-based, so the result
 * does not depend on the processing order.  For parent index i (numbering BEFORE the step), copy r in {0,1}
 * and pair j in {0,1}:  c = (i*2 + r)*2 + j;  h1 = mix(seed_lo ^ mix(c_lo ^ mix(c_hi + seed_hi)));
 * h2 = mix(h1 ^ 0x85ebca6b);  u1 = ((h1 >> 8) + 1) * 2^-24 in (0,1];  u2 = (h2 >> 8) * 2^-24 in [0,1);
 * z_a = sqrt(-2 ln u1) cos(2 pi u2), z_b = sqrt(-2 ln u1) sin(2 pi u2);  sample = (z_a(j=0), z_b(j=0), z_a(j=1)).
 * mix = the "lowbias32" integer finaliser.  (ln / cos / sin are the platform's: float parity to ~1e-6.)
 */
#include "orc_common.h"

static inline uint32_t orc_mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

void orc_split_normals(unsigned long long seed, unsigned long long parent, int copy, float z[3]) {
  for (int j = 0; j < 2; ++j) {
    const unsigned long long c = (parent * 2ull + (unsigned)copy) * 2ull + (unsigned)j;
    const uint32_t h1 = orc_mix32((uint32_t)seed ^ orc_mix32((uint32_t)c ^ orc_mix32((uint32_t)(c >> 32) + (uint32_t)(seed >> 32))));
    const uint32_t h2 = orc_mix32(h1 ^ 0x85ebca6bu);
    const float u1 = (float)((h1 >> 8) + 1u) * 0x1p-24f;
    const float u2 = (float)(h2 >> 8) * 0x1p-24f;
    const float rad = sqrtf(-2.0f * logf(u1));
    const float ang = 6.283185307179586f * u2;
    if (j == 0) {
      z[0] = rad * cosf(ang);
      z[1] = rad * sinf(ang);
    } else {
      z[2] = rad * cosf(ang);
    }
  }
}

typedef struct {
  float grad_threshold;  /* densify_grad_threshold_ (params.h:67) */
  float min_opacity;     /* 0.005 in the published schedule */
  float extent;          /* scene extent (cameras_extent) */
  float percent_dense;   /* percent_dense_ (gaussian_model.h:44) */
  float max_screen_size; /* 0 = none (the published loop passes 20 after the first opacity reset) */
  unsigned long long seed;
} orc_densify_params;

static float max_scale(const float *s) {
  const float a = expf(s[0]), b = expf(s[1]), c = expf(s[2]);
  return fmaxf(a, fmaxf(b, c));
}

static float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

/* rows: growing arrays of one Gaussian's 6 parameter groups + 12 moment groups */
typedef struct {
  int n, cap, w[6];
  float *p[6], *m[6], *v[6];
} dz_store;

static void dz_init(dz_store *s, int cap, const int w[6]) {
  s->n = 0;
  s->cap = cap;
  for (int k = 0; k < 6; ++k) {
    s->w[k] = w[k];
    const size_t bytes = sizeof(float) * (size_t)cap * (size_t)(w[k] > 0 ? w[k] : 1);
    s->p[k] = (float *)malloc(bytes);
    s->m[k] = (float *)malloc(bytes);
    s->v[k] = (float *)malloc(bytes);
  }
}
static void dz_free(dz_store *s) {
  for (int k = 0; k < 6; ++k) {
    free(s->p[k]);
    free(s->m[k]);
    free(s->v[k]);
  }
}
/* append row `src` of store a to store b; zero_moments: new Gaussians start with fresh Adam state */
static int dz_append(dz_store *b, const dz_store *a, int src, int zero_moments) {
  const int d = b->n++;
  for (int k = 0; k < 6; ++k) {
    const int w = b->w[k];
    memcpy(b->p[k] + (size_t)d * w, a->p[k] + (size_t)src * w, sizeof(float) * w);
    if (zero_moments) {
      memset(b->m[k] + (size_t)d * w, 0, sizeof(float) * w);
      memset(b->v[k] + (size_t)d * w, 0, sizeof(float) * w);
    } else {
      memcpy(b->m[k] + (size_t)d * w, a->m[k] + (size_t)src * w, sizeof(float) * w);
      memcpy(b->v[k] + (size_t)d * w, a->v[k] + (size_t)src * w, sizeof(float) * w);
    }
  }
  return d;
}

/* Group order everywhere: xyz, f_dc, f_rest, opacity, scaling, rotation (GaussianModel::setup,
 * src/scene/gaussian_model.cpp:323-345).  params / exp_avg / exp_avg_sq: N rows in; out_*: capacity 2N rows.
 * Returns the new N; counts[4] = {kept originals, kept clones, kept split parents (each contributes 2
 * children), pruned}.  The statistics are all zero afterwards (the caller re-allocates them). */
int orc_densify_and_prune(int N, const int widths[6], float *const params[6], float *const exp_avg[6],
                          float *const exp_avg_sq[6], const float *xyz_grad_accum, const float *denom,
                          const int *max_radii2D, const orc_densify_params *prm, float *const out_params[6],
                          float *const out_exp_avg[6], float *const out_exp_avg_sq[6], int counts[4]) {
  (void)max_radii2D; /* zeroed by the clone phase's postfix before the screen-size test reads it (quirk 1) */
  dz_store in, cur;
  in.n = N;
  in.cap = N;
  for (int k = 0; k < 6; ++k) {
    in.w[k] = widths[k];
    in.p[k] = params[k];
    in.m[k] = exp_avg[k];
    in.v[k] = exp_avg_sq[k];
  }
  float *grads = (float *)malloc(sizeof(float) * (size_t)(N > 0 ? N : 1));
  for (int i = 0; i < N; ++i) {
    const float g = xyz_grad_accum[i] / denom[i];
    grads[i] = isnan(g) ? 0.0f : g;
  }
  const float dense_limit = prm->percent_dense * prm->extent;
  /* ---- phase 1: densify_and_clone ---- */
  dz_init(&cur, 4 * N + 4, widths);
  for (int i = 0; i < N; ++i) dz_append(&cur, &in, i, 0);
  for (int i = 0; i < N; ++i)
    if (grads[i] >= prm->grad_threshold && max_scale(in.p[4] + 3 * (size_t)i) <= dense_limit) dz_append(&cur, &in, i, 1);
  const int N1 = cur.n;
  /* ---- phase 2: densify_and_split (N = 2 children) ---- */
  unsigned char *sel = (unsigned char *)calloc((size_t)N1 + 1, 1);
  int ns = 0;
  for (int i = 0; i < N1; ++i) {
    const float pg = i < N ? grads[i] : 0.0f;
    sel[i] = pg >= prm->grad_threshold && max_scale(cur.p[4] + 3 * (size_t)i) > dense_limit;
    ns += sel[i];
  }
  for (int r = 0; r < 2; ++r)
    for (int i = 0; i < N1; ++i) {
      if (!sel[i]) continue;
      const int d = dz_append(&cur, &cur, i, 1);
      const float *q = cur.p[5] + 4 * (size_t)i;
      const float *s = cur.p[4] + 3 * (size_t)i;
      const float *x = cur.p[0] + 3 * (size_t)i;
      float z[3];
      orc_split_normals(prm->seed, (unsigned long long)i, r, z);
      const float sc[3] = {expf(s[0]), expf(s[1]), expf(s[2])};
      const float smp[3] = {z[0] * sc[0], z[1] * sc[1], z[2] * sc[2]};
      /* build_rotation of the published code: normalise q = (r, x, y, z), then the usual matrix */
      const float nq = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
      const float qr = q[0] / nq, qx = q[1] / nq, qy = q[2] / nq, qz = q[3] / nq;
      const float R[9] = {1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qr * qz), 2 * (qx * qz + qr * qy),
                          2 * (qx * qy + qr * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qr * qx),
                          2 * (qx * qz - qr * qy), 2 * (qy * qz + qr * qx), 1 - 2 * (qx * qx + qy * qy)};
      float *nx = cur.p[0] + 3 * (size_t)d, *nsl = cur.p[4] + 3 * (size_t)d;
      for (int a = 0; a < 3; ++a) {
        nx[a] = R[3 * a] * smp[0] + R[3 * a + 1] * smp[1] + R[3 * a + 2] * smp[2] + x[a];
        nsl[a] = logf(sc[a] / (0.8f * 2.0f));
      }
    }
  const int N2 = cur.n;
  /* ---- phase 3: prune (split parents; low opacity; over-large in world space) ---- */
  int kept[3] = {0, 0, 0}, pruned = 0, n_out = 0;
  for (int i = 0; i < N2; ++i) {
    int drop = i < N1 && sel[i];
    const int parent_drop = drop;
    if (!drop) {
      drop = sigmoidf_(cur.p[3][i]) < prm->min_opacity;
      if (prm->max_screen_size > 0.0f) drop = drop || max_scale(cur.p[4] + 3 * (size_t)i) > 0.1f * prm->extent;
    }
    if (drop) {
      pruned += !parent_drop;
      continue;
    }
    for (int k = 0; k < 6; ++k) {
      const int w = widths[k];
      memcpy(out_params[k] + (size_t)n_out * w, cur.p[k] + (size_t)i * w, sizeof(float) * w);
      memcpy(out_exp_avg[k] + (size_t)n_out * w, cur.m[k] + (size_t)i * w, sizeof(float) * w);
      memcpy(out_exp_avg_sq[k] + (size_t)n_out * w, cur.v[k] + (size_t)i * w, sizeof(float) * w);
    }
    kept[i < N ? 0 : (i < N1 ? 1 : 2)]++;
    ++n_out;
  }
  if (counts) {
    counts[0] = kept[0];
    counts[1] = kept[1];
    counts[2] = kept[2] / 2;
    counts[3] = pruned;
  }
  free(sel);
  free(grads);
  dz_free(&cur);
  return n_out;
}

/* reset_opacity of the published trainer: opacity <- inverse_sigmoid(min(sigmoid(opacity), 0.01)) and fresh
 * Adam state for the opacity group (opacity_reset_interval_, src/arguments/params.h:64). */
void orc_reset_opacity(int N, float *opacity, float *exp_avg, float *exp_avg_sq) {
  for (int i = 0; i < N; ++i) {
    const float x = fminf(sigmoidf_(opacity[i]), 0.01f);
    opacity[i] = logf(x / (1.0f - x));
    exp_avg[i] = 0.0f;
    exp_avg_sq[i] = 0.0f;
  }
}
