// blend_fwd.cu — K3: forward alpha compositing.
//
// A 16x16 tile is eight 8x4-pixel boxes; ONE WARP owns a box (one pixel per lane) and walks the tile's list on its
// own, 32 entries at a time, until its 32 pixels have terminated -- there is no block barrier (a block is only a
// scheduling unit: BF_WARPS boxes of one tile, default one).  Tiles are launched heaviest-first (tile_order,
// binning.cu) because per-tile work is extremely skewed.  The 48-B records of the next chunk travel global -> shared
// as cp.async copies while the current chunk is walked.  Warp-level culling: the loading lane tests its entry against
// the warp's box (the conservative alpha-reach test of the exact tile cull, gs_common.cuh); one ballot gives the
// entries the warp walks.  Entries are evaluated four at a time (power / alpha of four entries are independent) and
// the short transmittance chain behind them is branch-free.
//
// For the segmented backward (blend_bwd.cu) the kernel also leaves, per tile:
//   * a checkpoint (T, C.rgb) of every live pixel at each GS_SEG-entry boundary of the tile list,
//   * the work items (tile, segment) for segment < ceil(max_pixel(n_contrib) / GS_SEG).
// Skip decisions (power > 0, alpha < 1/255) use the specified numerics in gs_common.cuh and are
// bit-identical to the oracle's; colour accumulation uses FMAs (smooth differences only).
#include <stdlib.h>

#include "gs_common.cuh"

// A tile is 8 boxes of 8x4 pixels, one warp each.  The warps are autonomous, so a block is only a scheduling unit:
// BF_WARPS boxes of one tile per block (small blocks free their registers and shared memory as soon as their boxes
// have terminated instead of waiting for the slowest box of the tile).
#define BF_WARPS_DEFAULT 1

// 48-byte record, global -> shared, asynchronously and without staging registers (LDGSTS); .ca keeps the line in L1
// for the other seven warps of the tile
__device__ __forceinline__ void bf_copy_record(GsGeom* dst, const GsGeom* src) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s), "l"(src) : "memory");
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s + 16u), "l"(reinterpret_cast<const char*>(src) + 16) : "memory");
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s + 32u), "l"(reinterpret_cast<const char*>(src) + 32) : "memory");
}
__device__ __forceinline__ void bf_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bf_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

template <int BF_WARPS>
__global__ void __launch_bounds__(32 * BF_WARPS, 32 / BF_WARPS)
blend_forward_kernel(GsDims d, const uint32_t* __restrict__ tile_order, const uint2* __restrict__ ranges,
                     const uint32_t* __restrict__ point_list, const GsGeom* __restrict__ geom, float bg0, float bg1,
                     float bg2, float* __restrict__ out_color, float4* __restrict__ final_state,
                     uint32_t* __restrict__ n_contrib, float4* __restrict__ ckpt, uint2* __restrict__ work_items,
                     uint32_t* __restrict__ work_count, uint32_t* __restrict__ tile_last /* [ntiles] max, then [ntiles] arrivals; zeroed */) {
  // per-warp slices, double-buffered: the 32 records of the chunk the warp is walking and of the next one
  constexpr int BF_BLOCKS_PER_TILE = 8 / BF_WARPS;
  __shared__ GsGeom s_g_all[BF_WARPS][2][33];  // slot 32 of each buffer: the null record (opacity 0: never contributes)

  const int tile = (int)tile_order[blockIdx.x / BF_BLOCKS_PER_TILE];
  const int tx = tile % d.gx, ty = tile / d.gx;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int box = (int)(blockIdx.x % BF_BLOCKS_PER_TILE) * BF_WARPS + warp;  // 0..7: column box & 1, row box >> 1
  const int lx = (box & 1) * 8 + (lane & 7), ly = (box >> 1) * 4 + (lane >> 3);
  const int px = tx * GS_TILE + lx, py = ty * GS_TILE + ly;
  const bool inside = px < d.W && py < d.H;
  const float pixx = (float)px, pixy = (float)py;

  const uint2 range = ranges[tile];
  const int total = (int)(range.y - range.x);
  float4* my_ckpt = ckpt + ((size_t)(range.x >> GS_SEG_SHIFT) + (size_t)tile) * GS_TILE_PIX + (ly * GS_TILE + lx);
  // the warp's 8x4 box of pixel centres
  const float bx = (float)(tx * GS_TILE + (box & 1) * 8), by = (float)(ty * GS_TILE + (box >> 1) * 4);
  const uint32_t* plist = point_list + range.x + lane;

  if (lane < 2) {
    GsGeom z;
    z.q0 = z.q1 = z.q2 = make_float4(0.f, 0.f, 0.f, 0.f);
    s_g_all[warp][lane][32] = z;
  }
  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
  uint32_t last_contributor = 0;

  // Every warp walks the tile list on its own, 32 entries at a time, and stops when its 32 pixels have terminated:
  // no block barrier inside the loop (the first version staged 256 entries per block between two barriers, and 30%
  // of its issue slots were warps waiting there for the slowest of the eight, profiles/r01k).  The records of the
  // next chunk are in flight (cp.async, global -> shared without registers) while the current one is walked, the
  // ids one chunk further ahead; the eight warps of a tile gather the same records, so most gathers hit L1.
  uint32_t id_next = 0;
  if (lane < total) {
    const uint32_t id0 = __ldg(plist);
    bf_copy_record(&s_g_all[warp][0][lane], geom + id0);
  }
  bf_commit();
  if (32 + lane < total) id_next = __ldg(plist + 32);
  for (int pos = 0, buf = 0; pos < total; pos += 32, buf ^= 1) {
    if (__all_sync(0xffffffffu, done)) break;
    // segment boundary: checkpoint the state in front of entry pos
    if ((pos & (GS_SEG - 1)) == 0 && pos > 0 && !done)
      my_ckpt[(size_t)(pos >> GS_SEG_SHIFT) * GS_TILE_PIX] = make_float4(T, C0, C1, C2);
    __syncwarp();  // every lane has finished reading the buffer the next copies overwrite
    if (pos + 32 + lane < total) bf_copy_record(&s_g_all[warp][buf ^ 1][lane], geom + id_next);
    bf_commit();
    if (pos + 64 + lane < total) id_next = __ldg(plist + pos + 64);
    bf_wait_1();  // this chunk's records have landed (the next chunk's may still be in flight)
    __syncwarp();
    const GsGeom* s_g = s_g_all[warp][buf];
    // ---- the loading lane decides whether its entry can reach alpha >= 1/255 anywhere in the warp's box
    //      (gs_box_alive: the same conservative test as the tile cull, so skipping a dead (warp, entry) pair
    //      never changes a pixel) ----
    bool alive = false;
    if (pos + lane < total) {
      const float4 q0 = s_g[lane].q0;
      const float hC = s_g[lane].q1.x;
      GsCull c;
      c.hA = q0.z;
      c.B = q0.w;
      c.hC = hC;
      c.rA = __frcp_rn(2.0f * q0.z);
      c.rC = __frcp_rn(2.0f * hC);
      c.thr = s_g[lane].q2.z;
      alive = gs_box_alive(c, q0.x, q0.y, bx, bx + 7.0f, by, by + 3.0f);
    }
    uint32_t word = __ballot_sync(0xffffffffu, alive);
    // ---- walk the alive entries in list order ----
    int last_j = -1;  // last entry of this chunk that contributed to my pixel
    while (word) {
      // up to four alive entries at a time: their power / alpha are independent, only T is sequential
      int jj[4];
      float alpha[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        jj[u] = word ? __ffs(word) - 1 : 32;  // 32: the null record (alpha = 0)
        word &= word - 1;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = jj[u];
        const float4 q0 = s_g[j].q0;
        const float hC = s_g[j].q1.x, op = s_g[j].q1.y;
        const float dx = q0.x - pixx, dy = q0.y - pixy;
        const float power = gs_power(q0.z, q0.w, hC, dx, dy);
        float G;
        const float a = gs_alpha(op, power, G);
        alpha[u] = (!(power > 0.0f) && !(a < GS_ALPHA_MIN)) ? a : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        // branch-free: an entry that does not contribute runs the same FMAs with weight 0 (state unchanged bit for bit)
        const int j = jj[u];
        const bool cand = alpha[u] != 0.0f && !done;
        const float test_T = T * (1.0f - alpha[u]);
        const bool term = cand && test_T < GS_T_MIN;
        done |= term;
        const bool app = cand && !term;
        const float w = app ? alpha[u] * T : 0.0f;
        C0 = fmaf(s_g[j].q1.z, w, C0);
        C1 = fmaf(s_g[j].q1.w, w, C1);
        C2 = fmaf(s_g[j].q2.x, w, C2);
        T = app ? test_T : T;
        last_j = app ? j : last_j;
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
    if (last_j >= 0) last_contributor = (uint32_t)(pos + last_j + 1);
  }
  asm volatile("cp.async.wait_all;" ::: "memory");  // no copy may outlive the warp
  if (inside) {
    const size_t pix = (size_t)py * d.W + px;
    const size_t plane = (size_t)d.H * d.W;
    final_state[pix] = make_float4(T, C0, C1, C2);  // out = C + T * bg; the backward rebuilds it from here
    n_contrib[pix] = last_contributor;
    out_color[pix] = fmaf(T, bg0, C0);
    out_color[plane + pix] = fmaf(T, bg1, C1);
    out_color[2 * plane + pix] = fmaf(T, bg2, C2);
  }
  // ---- work items of the segmented backward: (tile, s) for s < ceil(tile_last / GS_SEG), written by the last of
  //      the tile's eight boxes to arrive (tile_last is a running maximum, the second half of the array counts
  //      arrivals; both are zeroed before the launch) ----
  uint32_t wl = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wl = max(wl, __shfl_xor_sync(0xffffffffu, wl, o));
  uint32_t arrived = 0;
  if (lane == 0) {
    if (wl > 0) atomicMax(tile_last + tile, wl);
    __threadfence();
    arrived = atomicAdd(tile_last + d.ntiles + tile, 1u);
  }
  arrived = __shfl_sync(0xffffffffu, arrived, 0);
  if (arrived == 7u) {
    uint32_t tl = 0;
    if (lane == 0) {
      __threadfence();
      tl = atomicMax(tile_last + tile, 0u);
    }
    tl = __shfl_sync(0xffffffffu, tl, 0);
    const uint32_t nseg = (tl + GS_SEG - 1) >> GS_SEG_SHIFT;
    uint32_t pos = 0;
    if (lane == 0 && nseg > 0) pos = atomicAdd(work_count, nseg);
    pos = __shfl_sync(0xffffffffu, pos, 0);
    for (uint32_t s = lane; s < nseg; s += 32) work_items[pos + s] = make_uint2((uint32_t)tile, s);
  }
}

cudaError_t gs_launch_blend_forward(const GsDims& d, const uint32_t* tile_order, const uint2* ranges,
                                    const uint32_t* point_list, const GsGeom* geom, const float bg[3],
                                    float* out_color, float4* final_state, uint32_t* n_contrib, float4* ckpt,
                                    uint2* work_items, uint32_t* work_count, uint32_t* tile_last,
                                    bool tile_last_zeroed, cudaStream_t stream) {
  if (d.ntiles == 0) return cudaSuccess;
  if (!tile_last_zeroed) {  // (the forward clears it on its second stream, off the critical path, when it forks one)
    cudaError_t e = cudaMemsetAsync(tile_last, 0, (size_t)d.ntiles * 8, stream);
    if (e != cudaSuccess) return e;
  }
  static int warps = 0;
  if (warps == 0) {
    const char* env = getenv("GS_BF_WARPS");  // tuning switch: boxes of a tile per block (1, 2, 4 or 8)
    warps = env ? atoi(env) : BF_WARPS_DEFAULT;
    if (warps != 1 && warps != 2 && warps != 4 && warps != 8) warps = BF_WARPS_DEFAULT;
  }
#define BF_LAUNCH(WARPS)                                                                                              \
  blend_forward_kernel<WARPS><<<d.ntiles * (8 / WARPS), 32 * WARPS, 0, stream>>>(                                     \
      d, tile_order, ranges, point_list, geom, bg[0], bg[1], bg[2], out_color, final_state, n_contrib, ckpt, work_items, \
      work_count, tile_last)
  switch (warps) {
    case 1: BF_LAUNCH(1); break;
    case 2: BF_LAUNCH(2); break;
    case 4: BF_LAUNCH(4); break;
    default: BF_LAUNCH(8); break;
  }
#undef BF_LAUNCH
  return cudaGetLastError();
}
