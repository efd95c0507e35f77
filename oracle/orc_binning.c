/*
 * oracle/orc_binning.c — TEST INFRASTRUCTURE (see orc_common.h header).
 * Stage K2: inclusive scan of tiles_touched,
 * duplicateWithKeys (key = tile_id << 32 | fp32 bits of view depth, value = Gaussian id),
 * STABLE sort by key, identifyTileRanges.  No reference counterpart.
 * All outputs are integers and must match the GPU bit for bit.
 */
#include "orc_common.h"

static inline uint32_t f32_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}

/* tile rectangle of one Gaussian: same expression as orc_preprocess */
static inline void tile_rect(real px, real py, int rad, int gx, int gy, int *x0, int *y0, int *x1, int *y1) {
  int a0 = (int)((px - RC(rad)) / RC(ORC_TILE));
  int b0 = (int)((py - RC(rad)) / RC(ORC_TILE));
  int a1 = (int)((px + RC(rad) + RC(ORC_TILE - 1)) / RC(ORC_TILE));
  int b1 = (int)((py + RC(rad) + RC(ORC_TILE - 1)) / RC(ORC_TILE));
  *x0 = a0 < 0 ? 0 : (a0 > gx ? gx : a0);
  *x1 = a1 < 0 ? 0 : (a1 > gx ? gx : a1);
  *y0 = b0 < 0 ? 0 : (b0 > gy ? gy : b0);
  *y1 = b1 < 0 ? 0 : (b1 > gy ? gy : b1);
}

/* inclusive prefix sum; returns the total D (number of (Gaussian, tile) duplicates) */
uint64_t orc_scan_tiles(int N, const uint32_t *tiles_touched, uint32_t *point_offsets) {
  uint64_t acc = 0;
  for (int i = 0; i < N; ++i) {
    acc += tiles_touched[i];
    point_offsets[i] = (uint32_t)acc;
  }
  return acc;
}

/* duplicateWithKeys: Gaussian i writes its keys at [offsets[i-1], offsets[i]) in row-major
 * tile order (y outer, x inner). */
void orc_duplicate_with_keys(int H, int W, int N, const real *means2D, const real *depths, const int *radii,
                             const real *conic_opacity, const uint32_t *point_offsets, uint64_t *keys,
                             uint32_t *values) {
  const int gx = orc_grid_x(W), gy = orc_grid_y(H);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < N; ++i) {
    if (radii[i] <= 0) continue;
    uint32_t off = (i == 0) ? 0u : point_offsets[i - 1];
    int x0, y0, x1, y1;
    tile_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, &x0, &y0, &x1, &y1);
    uint32_t dbits = f32_bits((float)depths[i]);
    const int cull = orc_tile_cull_enabled && (x1 - x0) * (y1 - y0) <= ORC_CULL_MAX_TILES;
    const real *co = conic_opacity + 4 * (size_t)i;
    const orc_cull cs = orc_cull_setup(RC(0.5) * co[0], co[1], RC(0.5) * co[2], co[3]);
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        if (cull && !orc_tile_alive(&cs, means2D[2 * i], means2D[2 * i + 1], x, y)) continue;
        uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
        key = (key << 32) | dbits;
        keys[off] = key;
        values[off] = (uint32_t)i;
        ++off;
      }
  }
}

/* stable LSD radix sort of (key,value) pairs, 8-bit digits, skipping constant digits */
void orc_sort_pairs(uint64_t D, uint64_t *keys, uint32_t *values) {
  if (D == 0) return;
  uint64_t *k2 = (uint64_t *)malloc(D * sizeof(uint64_t));
  uint32_t *v2 = (uint32_t *)malloc(D * sizeof(uint32_t));
  uint64_t *ka = keys, *kb = k2;
  uint32_t *va = values, *vb = v2;
  for (int pass = 0; pass < 8; ++pass) {
    const int shift = 8 * pass;
    size_t hist[256];
    memset(hist, 0, sizeof(hist));
    for (uint64_t i = 0; i < D; ++i) hist[(ka[i] >> shift) & 0xFF]++;
    int constant = 0;
    for (int d = 0; d < 256; ++d)
      if (hist[d] == D) constant = 1;
    if (constant) continue;
    size_t sum = 0;
    for (int d = 0; d < 256; ++d) {
      size_t c = hist[d];
      hist[d] = sum;
      sum += c;
    }
    for (uint64_t i = 0; i < D; ++i) {
      size_t dst = hist[(ka[i] >> shift) & 0xFF]++;
      kb[dst] = ka[i];
      vb[dst] = va[i];
    }
    uint64_t *tk = ka; ka = kb; kb = tk;
    uint32_t *tv = va; va = vb; vb = tv;
  }
  if (ka != keys) {
    memcpy(keys, ka, D * sizeof(uint64_t));
    memcpy(values, va, D * sizeof(uint32_t));
  }
  free(k2);
  free(v2);
}

/* identifyTileRanges: ranges[2*t] = first index of tile t, ranges[2*t+1] = one past last.
 * Tiles without entries keep (0,0). */
void orc_identify_tile_ranges(uint64_t D, const uint64_t *keys, int num_tiles, uint32_t *ranges) {
  memset(ranges, 0, (size_t)num_tiles * 2 * sizeof(uint32_t));
  for (uint64_t i = 0; i < D; ++i) {
    uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
      ranges[2 * cur] = 0;
    } else {
      uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
      if (cur != prev) {
        ranges[2 * prev + 1] = (uint32_t)i;
        ranges[2 * cur] = (uint32_t)i;
      }
    }
    if (i == D - 1) ranges[2 * cur + 1] = (uint32_t)D;
  }
}
