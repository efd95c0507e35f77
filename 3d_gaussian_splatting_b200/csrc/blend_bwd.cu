// blend_bwd.cu — K4a: backward alpha compositing, segmented.
//
// Per-tile work is extremely skewed (mean ~300 processed list entries, max several thousand), and a
// pixel's walk through its tile list is sequential, so "one block per tile" is bound by the heaviest
// tile.  Instead the forward leaves a per-pixel checkpoint (T, C.rgb) at every GS_SEG-entry boundary
// and a list of (tile, segment) work items; here every work item is independent: a block restarts
// from the checkpoint and replays its GS_SEG entries FRONT-TO-BACK with exactly the forward's
// arithmetic (so T and the prefix colour C are the forward's bits), and gets the colour behind the
// current Gaussian from the final pixel colour:  Cb = out - C_prefix_after  (includes bg * T_final).
//     dL/dalpha_i = sum_ch dL/dC_ch * (c_ch * T_i - Cb_ch / (1 - alpha_i))
// Persistent warps pull work items (a 16x8 half of a tile x one segment) from an atomic cursor; every thread owns
// a 2x2 pixel quad, so a segment needs only two cross-lane reductions per Gaussian: per-thread partial sums of
// the 9 gradient components are reduced with a halving butterfly (8 values in 4+2+1+1+1 shuffles,
// the 9th in 5) and leave the SM as ONE predicated red.global.add.f32 (9 lanes, 9 consecutive floats).
//
// Gradient record per Gaussian (48 B, pre-zeroed):
//   { dL/dmean2D.x, dL/dmean2D.y, dL/dA, dL/dB } { dL/dC, dL/dopacity, dL/dr, dL/dg } { dL/db, -, -, - }
// (mean2D gradient in NDC units, x0.5W / x0.5H).  Skip decisions use the specified numerics of
// gs_common.cuh and are identical to the forward's and the oracle's.
#include "gs_common.cuh"

#define BB_THREADS 64
#define BB_WARPS (BB_THREADS / 32)
#define BB_PER_LANE (GS_SEG / 32)

// Every WARP is autonomous: it pulls its own work item -- (tile, segment, upper or lower 16x8 half of the tile) --
// stages the segment's records into its own shared-memory slice and never meets a block barrier.  (The first
// version paired the two halves of a tile in one 64-thread block: 17% of its issue slots were spent at the barrier
// waiting for the slower half, profiles/r01k.)
__global__ void __launch_bounds__(BB_THREADS, 10)
blend_backward_kernel(GsDims d, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      const GsGeom* __restrict__ geom, float bg0, float bg1, float bg2,
                      const float4* __restrict__ final_state, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                      const float4* __restrict__ ckpt, const uint2* __restrict__ work_items,
                      const uint32_t* __restrict__ work_count, uint32_t* __restrict__ work_cursor,
                      const uint32_t* __restrict__ tile_last, float* __restrict__ grad_geom) {
  __shared__ float4 s_q0_all[BB_WARPS][GS_SEG];
  __shared__ float4 s_q1_all[BB_WARPS][GS_SEG];
  __shared__ float s_b_all[BB_WARPS][GS_SEG];
  __shared__ uint32_t s_id_all[BB_WARPS][GS_SEG];

  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float4* const s_q0 = s_q0_all[wib];
  float4* const s_q1 = s_q1_all[wib];
  float* const s_b = s_b_all[wib];
  uint32_t* const s_id = s_id_all[wib];
  const size_t plane = (size_t)d.H * d.W;
  const float ddelx_dx = 0.5f * (float)d.W, ddely_dy = 0.5f * (float)d.H;
  const uint32_t nwork = 2u * *work_count;  // two halves per (tile, segment)

  for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(work_cursor, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= nwork) break;
    const uint2 item = work_items[w >> 1];
    const int half = (int)(w & 1u);
    const int tile = (int)item.x;
    const int seg = (int)item.y;
    const int k0 = seg << GS_SEG_SHIFT;                                  // first list entry of the segment
    const int n = min(GS_SEG, (int)tile_last[tile] - k0);                // entries in it
    const uint2 range = ranges[tile];
    const int tx = tile % d.gx, ty = tile / d.gx;
    const int qx = 2 * (lane & 7), qy = 2 * (lane >> 3) + 8 * half;
    const int x0 = tx * GS_TILE + qx, y0 = ty * GS_TILE + qy;

    // per-pixel state
    float pixx[4], pixy[4], T[4], C0[4], C1[4], C2[4], O0[4], O1[4], O2[4], dL0[4], dL1[4], dL2[4];
    int last[4];
    const float4* ck = ckpt + ((size_t)(range.x >> GS_SEG_SHIFT) + (size_t)tile + (size_t)seg) * GS_TILE_PIX;
    int my_last = 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int px = x0 + (p & 1), py = y0 + (p >> 1);
      const bool inside = px < d.W && py < d.H;
      pixx[p] = (float)px;
      pixy[p] = (float)py;
      const size_t pix = (size_t)py * d.W + px;
      last[p] = inside ? (int)n_contrib[pix] : 0;
      const bool active = last[p] > k0;  // this pixel still had contributors in this segment
      dL0[p] = active ? dL_dpixels[pix] : 0.f;
      dL1[p] = active ? dL_dpixels[plane + pix] : 0.f;
      dL2[p] = active ? dL_dpixels[2 * plane + pix] : 0.f;
      const float4 fs = active ? final_state[pix] : make_float4(0.f, 0.f, 0.f, 0.f);
      O0[p] = fmaf(fs.x, bg0, fs.y);  // final pixel colour incl. background, as the forward wrote it
      O1[p] = fmaf(fs.x, bg1, fs.z);
      O2[p] = fmaf(fs.x, bg2, fs.w);
      float4 st = make_float4(1.f, 0.f, 0.f, 0.f);
      if (active && seg > 0) st = ck[(qy + (p >> 1)) * GS_TILE + qx + (p & 1)];
      T[p] = st.x;
      C0[p] = st.y;
      C1[p] = st.z;
      C2[p] = st.w;
      if (!active) last[p] = 0;
      my_last = max(my_last, last[p]);
    }
    int warp_last = my_last;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
    const int jend = min(n, warp_last - k0);  // this warp only needs entries k < warp_last
    if (jend <= 0) continue;

    // gather the records this warp needs (coalesced ids, L2-resident 48-B records); the loading lane also decides
    // whether its entry can reach alpha >= 1/255 anywhere in the warp's 16x8 pixel box (gs_box_alive, the
    // conservative test of the exact tile cull): the warp then walks only those
    uint32_t words[BB_PER_LANE];
    __syncwarp();  // every lane has finished reading the previous item's records
#pragma unroll
    for (int u = 0; u < BB_PER_LANE; ++u) {
      const int j = lane + 32 * u;
      bool alive = false;
      if (j < jend) {
        const uint32_t id = __ldg(point_list + range.x + (uint32_t)(k0 + j));
        const float4* gp = reinterpret_cast<const float4*>(geom + id);
        const float4 q0 = __ldg(gp), q1 = __ldg(gp + 1), q2 = __ldg(gp + 2);
        s_id[j] = id;
        s_q0[j] = q0;
        s_q1[j] = q1;
        s_b[j] = q2.x;
        GsCull c;
        c.hA = q0.z;
        c.B = q0.w;
        c.hC = q1.x;
        c.rA = __frcp_rn(2.0f * q0.z);
        c.rC = __frcp_rn(2.0f * q1.x);
        c.thr = q2.z;
        const float bx = (float)(tx * GS_TILE), by = (float)(ty * GS_TILE + 8 * half);
        alive = gs_box_alive(c, q0.x, q0.y, bx, bx + 15.0f, by, by + 7.0f);
      }
      words[u] = __ballot_sync(0xffffffffu, alive);
    }
    __syncwarp();

#pragma unroll
    for (int wd = 0; wd < BB_PER_LANE; ++wd) {
     uint32_t word = words[wd];
     while (word) {
      const int j = (wd << 5) + __ffs(word) - 1;
      word &= word - 1;
      const int k = k0 + j;
      const float4 q0 = s_q0[j];
      const float4 q1 = s_q1[j];
      const float cb = s_b[j];
      float a_mx = 0.f, a_my = 0.f, a_A = 0.f, a_B = 0.f, a_C = 0.f, a_o = 0.f, a_r = 0.f, a_g = 0.f, a_b = 0.f;
      bool any = false;
      // Branch-free over the four pixels (a skipped pixel runs the same arithmetic with alpha = 0, G = 0, which leaves
      // its state and every sum unchanged bit for bit): four independent dependency chains the scheduler can
      // interleave, instead of eight divergent branches per entry.
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const float dx = q0.x - pixx[p], dy = q0.y - pixy[p];
        const float power = gs_power(q0.z, q0.w, q1.x, dx, dy);
        float Gf;
        const float af = gs_alpha(q1.y, power, Gf);
        const bool valid = k < last[p] && !(power > 0.0f) && !(af < GS_ALPHA_MIN);
        any |= valid;
        const float alpha = valid ? af : 0.0f;
        const float G = valid ? Gf : 0.0f;
        const float one_m = 1.0f - alpha;
        const float Ti = T[p];
        const float wgt = alpha * Ti;
        C0[p] = fmaf(q1.z, wgt, C0[p]);   // same FMAs as the forward: prefix colour incl. this Gaussian
        C1[p] = fmaf(q1.w, wgt, C1[p]);
        C2[p] = fmaf(cb, wgt, C2[p]);
        T[p] = Ti * one_m;
        const float inv = __fdividef(1.0f, one_m);
        a_r = fmaf(wgt, dL0[p], a_r);
        a_g = fmaf(wgt, dL1[p], a_g);
        a_b = fmaf(wgt, dL2[p], a_b);
        float dL_dalpha = (q1.z * Ti - (O0[p] - C0[p]) * inv) * dL0[p];
        dL_dalpha = fmaf(q1.w * Ti - (O1[p] - C1[p]) * inv, dL1[p], dL_dalpha);
        dL_dalpha = fmaf(cb * Ti - (O2[p] - C2[p]) * inv, dL2[p], dL_dalpha);
        a_o = fmaf(G, dL_dalpha, a_o);
        const float wg = q1.y * dL_dalpha * G;  // dL/dG * G
        const float wdx = wg * dx, wdy = wg * dy;
        a_mx += wdx;               // sum w dx
        a_my += wdy;               // sum w dy
        a_A = fmaf(wdx, dx, a_A);  // sum w dx^2
        a_B = fmaf(wdx, dy, a_B);  // sum w dx dy
        a_C = fmaf(wdy, dy, a_C);  // sum w dy^2
      }
      if (__ballot_sync(0xffffffffu, any) == 0u) continue;
      // moments -> gradient components (A = 2*hA, C = 2*hC)
      const float A = 2.0f * q0.z, B = q0.w, C = 2.0f * q1.x;
      float v[8];
      v[0] = -(A * a_mx + B * a_my) * ddelx_dx;  // dL/dmean2D.x
      v[1] = -(C * a_my + B * a_mx) * ddely_dy;  // dL/dmean2D.y
      v[2] = -0.5f * a_A;                        // dL/dA
      v[3] = -a_B;                               // dL/dB
      v[4] = -0.5f * a_C;                        // dL/dC
      v[5] = a_o;                                // dL/dopacity
      v[6] = a_r;
      v[7] = a_g;
      float v8 = a_b;
      // halving butterfly: 8 values -> each quad of lanes ends up owning the total of one value
      const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4;
      float w4[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float send = up16 ? v[i] : v[i + 4];
        const float keep = up16 ? v[i + 4] : v[i];
        w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
      }
      float w2[2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const float send = up8 ? w4[i] : w4[i + 2];
        const float keep = up8 ? w4[i + 2] : w4[i];
        w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
      }
      float w1;
      {
        const float send = up4 ? w2[0] : w2[1];
        const float keep = up4 ? w2[1] : w2[0];
        w1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
      }
      w1 += __shfl_xor_sync(0xffffffffu, w1, 2);
      w1 += __shfl_xor_sync(0xffffffffu, w1, 1);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v8 += __shfl_xor_sync(0xffffffffu, v8, o);
      // lane L (L % 4 == 0) owns value index 4*bit4 + 2*bit3 + bit2; lane 1 carries the ninth
      const int idx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
      float* dst = grad_geom + 12 * (size_t)s_id[j];
      const bool ninth = lane == 1;
      if ((lane & 3) == 0 || ninth) atomicAdd(dst + (ninth ? 8 : idx), ninth ? v8 : w1);
     }
    }
  }
}

cudaError_t gs_launch_blend_backward(const GsDims& d, const uint2* ranges, const uint32_t* point_list,
                                     const GsGeom* geom, const float bg[3], const float4* final_state,
                                     const uint32_t* n_contrib,
                                     const float* dL_dpixels, const float4* ckpt, const uint2* work_items,
                                     const uint32_t* work_count, uint32_t* work_cursor, const uint32_t* tile_last,
                                     float4* grad_geom, int num_sms, cudaStream_t stream) {
  if (d.ntiles == 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(work_cursor, 0, sizeof(uint32_t), stream);
  if (e != cudaSuccess) return e;
  const int blocks = num_sms * 10;  // persistent: 95 regs x 64 threads -> 10 blocks per SM
  blend_backward_kernel<<<blocks, BB_THREADS, 0, stream>>>(d, ranges, point_list, geom, bg[0], bg[1], bg[2],
                                                          final_state, n_contrib,
                                                          dL_dpixels, ckpt, work_items, work_count, work_cursor,
                                                          tile_last, reinterpret_cast<float*>(grad_geom));
  return cudaGetLastError();
}
