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
// Persistent warps pull work items (a 16x4 pixel box of a tile x one segment) from an atomic cursor; every thread owns
// two x-adjacent pixels (BB_QR = 1; 2x2 quads and 16x8 boxes with BB_QR = 2: 96 registers and 20 warps/SM against 61
// and 32, measured 0.328 against 0.302 ms), so a segment needs four cross-lane reductions per Gaussian: per-thread partial sums of
// the 9 gradient components are reduced with a halving butterfly (8 values in 4+2+1+1+1 shuffles,
// the 9th in 5) and leave the SM as ONE predicated red.global.add.f32 (9 lanes, 9 consecutive floats).
//
// Gradient record per Gaussian (48 B, pre-zeroed):
//   { dL/dmean2D.x, dL/dmean2D.y, dL/dA, dL/dB } { dL/dC, dL/dopacity, dL/dr, dL/dg } { dL/db, -, -, - }
// (mean2D gradient in NDC units, x0.5W / x0.5H).  Skip decisions use the specified numerics of
// gs_common.cuh and are identical to the forward's and the oracle's.
#include "gs_common.cuh"

#define BB_THREADS 32
#define BB_WARPS (BB_THREADS / 32)
#define BB_PER_LANE (GS_SEG / 32)
// pixel rows per lane: a lane owns BB_QR rows of two x-adjacent pixels, a warp a 16 x (4 BB_QR) box, a tile has
// BB_PARTS such boxes
#ifndef BB_QR
#define BB_QR 1
#endif
#define BB_NP (2 * BB_QR)
#define BB_PARTS (4 / BB_QR)
#define BB_BLOCKS_PER_SM ((BB_QR == 1 ? 32 : 20) / BB_WARPS)

#ifndef BB_PAIR
#define BB_PAIR 1
#endif
// One entry: halving butterfly, 8 values -> each quad of lanes ends up owning the total of one value (4+2+1+1+1
// shuffles), the ninth in 5 more; ONE predicated red.global.add.f32 (9 lanes, 9 consecutive floats).
__device__ __forceinline__ void bb_reduce_one(const float v[8], float v8, uint32_t id, int lane, float* __restrict__ grad_geom) {
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
  float* dst = grad_geom + 12 * (size_t)id;
  const bool ninth = lane == 1;
  if ((lane & 3) == 0 || ninth) atomicAdd(dst + (ninth ? 8 : idx), ninth ? v8 : w1);
}
// Two entries at once: the half-warps swap entries (9 shuffles), then each half reduces its entry's 8 values by
// halving (4+2+1+1) and the ninth plainly (4): 21 shuffles per pair; one predicated red for both (18 lanes).
__device__ __forceinline__ void bb_reduce_pair(const float a[8], float a8, uint32_t ida, const float b[8], float b8, uint32_t idb,
                                               int lane, float* __restrict__ grad_geom) {
  const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4, up2 = lane & 2;
  float x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float send = up16 ? a[i] : b[i];
    const float keep = up16 ? b[i] : a[i];
    x[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  float x8;
  {
    const float send = up16 ? a8 : b8;
    const float keep = up16 ? b8 : a8;
    x8 = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  float w4[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float send = up8 ? x[i] : x[i + 4];
    const float keep = up8 ? x[i + 4] : x[i];
    w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  float w2[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float send = up4 ? w4[i] : w4[i + 2];
    const float keep = up4 ? w4[i + 2] : w4[i];
    w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  float w1;
  {
    const float send = up2 ? w2[0] : w2[1];
    const float keep = up2 ? w2[1] : w2[0];
    w1 = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  w1 += __shfl_xor_sync(0xffffffffu, w1, 1);
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) x8 += __shfl_xor_sync(0xffffffffu, x8, o);
  // even lane L of a half-warp owns value index 4*bit3 + 2*bit2 + bit1 of its half's entry; lanes 1 / 17 the ninth
  const int idx = ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
  float* dst = grad_geom + 12 * (size_t)(up16 ? idb : ida);
  const bool ninth = (lane & 15) == 1;
  if ((lane & 1) == 0 || ninth) atomicAdd(dst + (ninth ? 8 : idx), ninth ? x8 : w1);
}

// Every WARP is autonomous: it pulls its own work item -- (tile, segment, one of the tile's BB_PARTS pixel boxes) --
// stages the segment's records into its own shared-memory slice and never meets a block barrier.  (The first
// version paired the two halves of a tile in one 64-thread block: 17% of its issue slots were spent at the barrier
// waiting for the slower half, profiles/r01k.)
__global__ void __launch_bounds__(BB_THREADS, BB_BLOCKS_PER_SM)
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
  const uint32_t nwork = (uint32_t)BB_PARTS * *work_count;  // BB_PARTS boxes per (tile, segment)

  for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(work_cursor, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= nwork) break;
    const uint2 item = work_items[w / BB_PARTS];
    const int part = (int)(w % BB_PARTS);
    const int tile = (int)item.x;
    const int seg = (int)item.y;
    const int k0 = seg << GS_SEG_SHIFT;                                  // first list entry of the segment
    const int n = min(GS_SEG, (int)tile_last[tile] - k0);                // entries in it
    const uint2 range = ranges[tile];
    const int tx = tile % d.gx, ty = tile / d.gx;
    const int qx = 2 * (lane & 7), qy = BB_QR * (lane >> 3) + 4 * BB_QR * part;
    const int x0 = tx * GS_TILE + qx, y0 = ty * GS_TILE + qy;

    // per-pixel state
    float pixx[BB_NP], pixy[BB_NP], T[BB_NP], C0[BB_NP], C1[BB_NP], C2[BB_NP], O0[BB_NP], O1[BB_NP], O2[BB_NP], dL0[BB_NP], dL1[BB_NP],
        dL2[BB_NP];
    int last[BB_NP];
    const float4* ck = ckpt + ((size_t)(range.x >> GS_SEG_SHIFT) + (size_t)tile + (size_t)seg) * GS_TILE_PIX;
    int my_last = 0;
#pragma unroll
    for (int p = 0; p < BB_NP; ++p) {
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
    // the state of the two x-adjacent pixels of each quad row as packed pairs (FFMA2 / FMUL2 / FADD2 inner loop)
    gs_f2 npx2[BB_QR], npy2[BB_QR], T2[BB_QR], C0_2[BB_QR], C1_2[BB_QR], C2_2[BB_QR], O0_2[BB_QR], O1_2[BB_QR], O2_2[BB_QR], dL0_2[BB_QR],
        dL1_2[BB_QR], dL2_2[BB_QR];
#pragma unroll
    for (int h = 0; h < BB_QR; ++h) {
      npx2[h] = gs_f2_pack(-pixx[2 * h], -pixx[2 * h + 1]);
      npy2[h] = gs_f2_pack(-pixy[2 * h], -pixy[2 * h + 1]);
      T2[h] = gs_f2_pack(T[2 * h], T[2 * h + 1]);
      C0_2[h] = gs_f2_pack(C0[2 * h], C0[2 * h + 1]);
      C1_2[h] = gs_f2_pack(C1[2 * h], C1[2 * h + 1]);
      C2_2[h] = gs_f2_pack(C2[2 * h], C2[2 * h + 1]);
      O0_2[h] = gs_f2_pack(O0[2 * h], O0[2 * h + 1]);
      O1_2[h] = gs_f2_pack(O1[2 * h], O1[2 * h + 1]);
      O2_2[h] = gs_f2_pack(O2[2 * h], O2[2 * h + 1]);
      dL0_2[h] = gs_f2_pack(dL0[2 * h], dL0[2 * h + 1]);
      dL1_2[h] = gs_f2_pack(dL1[2 * h], dL1[2 * h + 1]);
      dL2_2[h] = gs_f2_pack(dL2[2 * h], dL2[2 * h + 1]);
    }
    const gs_f2 kNeg1 = gs_f2_pack(-1.0f, -1.0f), kOne = gs_f2_pack(1.0f, 1.0f);
    const gs_f2 kNegLog2e = gs_f2_pack(-1.44269504088896340736f, -1.44269504088896340736f);

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
        const float bx = (float)(tx * GS_TILE), by = (float)(ty * GS_TILE + 4 * BB_QR * part);
        alive = gs_box_alive(c, q0.x, q0.y, bx, bx + 15.0f, by, by + (float)(4 * BB_QR - 1));
      }
      words[u] = __ballot_sync(0xffffffffu, alive);
    }
    __syncwarp();

#if BB_PAIR
    float pend[8], pend8 = 0.f;
    uint32_t pend_id = 0;
    bool have_pend = false;
#endif
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
      bool any = false;
      // Branch-free and packed: the two x-adjacent pixels of a quad row share every FFMA2 / FMUL2 / FADD2 (each half
      // is the scalar IEEE operation, so T and the prefix colour stay the forward's bits); a skipped pixel runs the
      // same arithmetic with alpha = 0, G = 0, which leaves its state and every sum unchanged.
      const gs_f2 mx2 = gs_f2_pack(q0.x, q0.x), my2 = gs_f2_pack(q0.y, q0.y);
      const gs_f2 hA2 = gs_f2_pack(q0.z, q0.z), B2 = gs_f2_pack(q0.w, q0.w), hC2 = gs_f2_pack(q1.x, q1.x);
      const gs_f2 op2 = gs_f2_pack(q1.y, q1.y);
      const gs_f2 cr2 = gs_f2_pack(q1.z, q1.z), cg2 = gs_f2_pack(q1.w, q1.w), cb2 = gs_f2_pack(cb, cb);
      gs_f2 s_mx, s_my, s_A, s_B, s_C, s_o, s_r, s_g, s_b2;  // initialised by the first pair (h == 0)
#pragma unroll
      for (int h = 0; h < BB_QR; ++h) {
        const gs_f2 dx2 = gs_f2_add(mx2, npx2[h]), dy2 = gs_f2_add(my2, npy2[h]);
        // gs_power: q = fma(dx, fma(hA, dx, B * dy), (hC * dy) * dy), power = -q
        const gs_f2 u2 = gs_f2_fma(hA2, dx2, gs_f2_mul(B2, dy2));
        const gs_f2 qq2 = gs_f2_fma(dx2, u2, gs_f2_mul(gs_f2_mul(hC2, dy2), dy2));
        const gs_f2 t2 = gs_f2_mul(qq2, kNegLog2e);  // power * log2(e): (-q) * c and q * (-c) are the same bits
        float qq[2], tt[2], al[2], Gm[2];
        gs_f2_unpack(qq2, qq[0], qq[1]);
        gs_f2_unpack(t2, tt[0], tt[1]);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          float G;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(G) : "f"(tt[e]));
          float a = q1.y * G;
          if (fabsf(a - GS_ALPHA_MIN) < 8e-8f) {  // guard band of gs_alpha: the skip decision is the oracle's
            G = gs_exp_ref(-qq[e]);
            a = __fmul_rn(q1.y, G);
          }
          a = fminf(GS_ALPHA_MAX, a);
          const bool valid = k < last[2 * h + e] && !(qq[e] < 0.0f) && !(a < GS_ALPHA_MIN);  // power > 0 <=> q < 0
          any |= valid;
          al[e] = valid ? a : 0.0f;
          Gm[e] = valid ? G : 0.0f;
        }
        const gs_f2 alpha2 = gs_f2_pack(al[0], al[1]), G2 = gs_f2_pack(Gm[0], Gm[1]);
        const gs_f2 one_m2 = gs_f2_fma(alpha2, kNeg1, kOne);  // 1 - alpha, one rounding
        const gs_f2 Ti2 = T2[h];
        const gs_f2 wgt2 = gs_f2_mul(alpha2, Ti2);
        C0_2[h] = gs_f2_fma(cr2, wgt2, C0_2[h]);  // same FMAs as the forward: prefix colour incl. this Gaussian
        C1_2[h] = gs_f2_fma(cg2, wgt2, C1_2[h]);
        C2_2[h] = gs_f2_fma(cb2, wgt2, C2_2[h]);
        T2[h] = gs_f2_mul(Ti2, one_m2);
        float om[2];
        gs_f2_unpack(one_m2, om[0], om[1]);
        float iv[2];  // 1 / (1 - alpha), 1 - alpha in [0.01, 1]: one MUFU.RCP each (gradient value only, no decision)
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iv[0]) : "f"(om[0]));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iv[1]) : "f"(om[1]));
        const gs_f2 inv2 = gs_f2_pack(iv[0], iv[1]);
        s_r = h == 0 ? gs_f2_mul(wgt2, dL0_2[h]) : gs_f2_fma(wgt2, dL0_2[h], s_r);
        s_g = h == 0 ? gs_f2_mul(wgt2, dL1_2[h]) : gs_f2_fma(wgt2, dL1_2[h], s_g);
        s_b2 = h == 0 ? gs_f2_mul(wgt2, dL2_2[h]) : gs_f2_fma(wgt2, dL2_2[h], s_b2);
        // dL/dalpha = sum_ch (c_ch * Ti - (O_ch - C_ch) / (1 - alpha)) * dL_ch
        gs_f2 f0 = gs_f2_fma(gs_f2_mul(gs_f2_fma(C0_2[h], kNeg1, O0_2[h]), inv2), kNeg1, gs_f2_mul(cr2, Ti2));
        gs_f2 f1 = gs_f2_fma(gs_f2_mul(gs_f2_fma(C1_2[h], kNeg1, O1_2[h]), inv2), kNeg1, gs_f2_mul(cg2, Ti2));
        gs_f2 f2 = gs_f2_fma(gs_f2_mul(gs_f2_fma(C2_2[h], kNeg1, O2_2[h]), inv2), kNeg1, gs_f2_mul(cb2, Ti2));
        gs_f2 dLa2 = gs_f2_mul(f0, dL0_2[h]);
        dLa2 = gs_f2_fma(f1, dL1_2[h], dLa2);
        dLa2 = gs_f2_fma(f2, dL2_2[h], dLa2);
        s_o = h == 0 ? gs_f2_mul(G2, dLa2) : gs_f2_fma(G2, dLa2, s_o);
        const gs_f2 wg2 = gs_f2_mul(gs_f2_mul(op2, dLa2), G2);  // dL/dG * G
        const gs_f2 wdx2 = gs_f2_mul(wg2, dx2), wdy2 = gs_f2_mul(wg2, dy2);
        s_mx = h == 0 ? wdx2 : gs_f2_add(s_mx, wdx2);                              // sum w dx
        s_my = h == 0 ? wdy2 : gs_f2_add(s_my, wdy2);                              // sum w dy
        s_A = h == 0 ? gs_f2_mul(wdx2, dx2) : gs_f2_fma(wdx2, dx2, s_A);           // sum w dx^2
        s_B = h == 0 ? gs_f2_mul(wdx2, dy2) : gs_f2_fma(wdx2, dy2, s_B);           // sum w dx dy
        s_C = h == 0 ? gs_f2_mul(wdy2, dy2) : gs_f2_fma(wdy2, dy2, s_C);           // sum w dy^2
      }
      if (__ballot_sync(0xffffffffu, any) == 0u) continue;
      float a_mx, a_my, a_A, a_B, a_C, a_o, a_r, a_g, a_b;
      {
        float lo, hi;
        gs_f2_unpack(s_mx, lo, hi); a_mx = lo + hi;
        gs_f2_unpack(s_my, lo, hi); a_my = lo + hi;
        gs_f2_unpack(s_A, lo, hi); a_A = lo + hi;
        gs_f2_unpack(s_B, lo, hi); a_B = lo + hi;
        gs_f2_unpack(s_C, lo, hi); a_C = lo + hi;
        gs_f2_unpack(s_o, lo, hi); a_o = lo + hi;
        gs_f2_unpack(s_r, lo, hi); a_r = lo + hi;
        gs_f2_unpack(s_g, lo, hi); a_g = lo + hi;
        gs_f2_unpack(s_b2, lo, hi); a_b = lo + hi;
      }
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
#if BB_PAIR
      // two entries per butterfly (21 shuffles per pair instead of 2 x 14): the first of a pair waits in registers
      if (__ballot_sync(0xffffffffu, have_pend) == 0u) {  // (warp-uniform by construction; the vote tells the compiler)
#pragma unroll
        for (int i = 0; i < 8; ++i) pend[i] = v[i];
        pend8 = v8;
        pend_id = s_id[j];
        have_pend = true;
        continue;
      }
      have_pend = false;
      bb_reduce_pair(pend, pend8, pend_id, v, v8, s_id[j], lane, grad_geom);
#else
      bb_reduce_one(v, v8, s_id[j], lane, grad_geom);
#endif
     }
    }
#if BB_PAIR
    if (__ballot_sync(0xffffffffu, have_pend) != 0u) {
      have_pend = false;
      bb_reduce_one(pend, pend8, pend_id, lane, grad_geom);
    }
#endif
  }
}

cudaError_t gs_launch_blend_backward(const GsDims& d, const uint2* ranges, const uint32_t* point_list,
                                     const GsGeom* geom, const float bg[3], const float4* final_state,
                                     const uint32_t* n_contrib,
                                     const float* dL_dpixels, const float4* ckpt, const uint2* work_items,
                                     const uint32_t* work_count, uint32_t* work_cursor, const uint32_t* tile_last,
                                     float4* grad_geom, int num_sms, bool cursor_zeroed, cudaStream_t stream) {
  if (d.ntiles == 0) return cudaSuccess;
  if (!cursor_zeroed) {  // (the forward's counter clear covers the first backward after it)
    cudaError_t e = cudaMemsetAsync(work_cursor, 0, sizeof(uint32_t), stream);
    if (e != cudaSuccess) return e;
  }
  const int blocks = num_sms * BB_BLOCKS_PER_SM;  // persistent: 96 regs x 64 threads -> 10 blocks per SM
  blend_backward_kernel<<<blocks, BB_THREADS, 0, stream>>>(d, ranges, point_list, geom, bg[0], bg[1], bg[2],
                                                          final_state, n_contrib,
                                                          dL_dpixels, ckpt, work_items, work_count, work_cursor,
                                                          tile_last, reinterpret_cast<float*>(grad_geom));
  return cudaGetLastError();
}
