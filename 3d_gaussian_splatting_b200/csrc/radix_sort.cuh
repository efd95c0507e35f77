// radix_sort.cuh — this library's own stable LSD radix-sort pass (8-bit digits) for (key, value)
// pairs, used twice by the binning stage (binning.cu):
//   * Gaussians by view depth   (u32 keys = fp32 bits of the depth, 4 passes over N pairs)
//   * duplicates by tile id     (u16/u32 keys, ceil(log2(T)/8) passes over D pairs)
// One pass = three launches on a fixed persistent grid (so the element count may live in device
// memory and no host round-trip is needed):
//   radix_hist_kernel    per-block digit histograms of the block's contiguous chunk -> counts[256][B]
//   radix_scan_kernel    exclusive scan of every digit's row of counts (one block per digit)
//   radix_scatter_kernel re-reads the chunk tile by tile (4096 pairs), ranks every pair stably
//                        (ballot-built peer masks + per-warp counters, then a prefix over warps), reorders the
//                        tile by digit in shared memory and writes digit-contiguous runs, so global
//                        stores coalesce.
// Stability is what makes the two-level scheme exact: pairs that compare equal keep their input order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_BINS 256
#define RS_MAX_GRID 1024  // persistent grid size limit (the scan kernel has one thread per chunk)
// pairs per thread and tile: 16 (4096-pair tiles) for the multi-million pair lists; 8 (2048-pair tiles, twice the
// blocks, half the serial ranking rounds per block) for the N-sized depth argsort, which is latency-bound

struct RadixChunk {
  uint32_t begin, end;  // element range of this block
};

__device__ __forceinline__ RadixChunk radix_chunk(uint32_t n, uint32_t tile) {
  const uint32_t tiles = (n + tile - 1) / tile;
  const uint32_t tpb = (tiles + gridDim.x - 1) / gridDim.x;
  RadixChunk c;
  const uint64_t b = (uint64_t)blockIdx.x * tpb * tile;
  const uint64_t e = b + (uint64_t)tpb * tile;
  c.begin = (uint32_t)(b < n ? b : n);
  c.end = (uint32_t)(e < n ? e : n);
  return c;
}

// Per-block digit histogram of the block's chunk.  16-byte loads (8 u16 / 4 u32 keys per lane), per-warp
// private shared-memory histograms, and run-length aggregation inside each lane's vector: tile ids of one
// Gaussian's rectangle and the high bytes of depths come in runs, which would otherwise serialise the
// shared-memory atomics on a single bin.  (match.any is avoided on purpose: measured at ~64 SM cycles
// per instruction on sm_100, it alone cost 60% of the SM's time in the first version of this kernel.)
template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const KeyT* __restrict__ keys, uint32_t n_host, const uint32_t* __restrict__ n_dev, int shift,
                  uint32_t tile, uint32_t* __restrict__ counts) {
  __shared__ uint32_t s_hist[RS_WARPS][RS_BINS];
  constexpr int VEC = 16 / sizeof(KeyT);
  const uint32_t n = n_dev ? *n_dev : n_host;
  const RadixChunk c = radix_chunk(n, tile);
  const int warp = threadIdx.x >> 5;
  for (int k = threadIdx.x; k < RS_WARPS * RS_BINS; k += RS_THREADS) (&s_hist[0][0])[k] = 0;
  __syncthreads();
  uint32_t* h = s_hist[warp];
  const uint32_t nvec = (c.end - c.begin) / VEC;  // c.begin is a multiple of the tile size: 16-byte aligned
  const uint4* vp = reinterpret_cast<const uint4*>(keys + c.begin);
  for (uint32_t v = threadIdx.x; v < nvec; v += RS_THREADS) {
    const uint4 q = __ldg(vp + v);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    uint32_t prev = 0xFFFFFFFFu, run = 0;
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      const uint32_t key = sizeof(KeyT) == 2 ? ((w[k >> 1] >> (16 * (k & 1))) & 0xFFFFu) : w[k & 3];
      const uint32_t d = (key >> shift) & 0xFFu;
      if (d != prev) {
        if (run) atomicAdd(&h[prev], run);
        prev = d;
        run = 0;
      }
      ++run;
    }
    atomicAdd(&h[prev], run);
  }
  for (uint32_t i = c.begin + nvec * VEC + threadIdx.x; i < c.end; i += RS_THREADS)
    atomicAdd(&h[((uint32_t)keys[i] >> shift) & 0xFFu], 1u);
  __syncthreads();
  uint32_t tot = 0;
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) tot += s_hist[w][threadIdx.x];
  counts[threadIdx.x * gridDim.x + blockIdx.x] = tot;
}

// Block d scans row d of counts[256][B] in place (exclusive) and leaves the row total in
// counts[256*B + d]; the scatter kernel scans those 256 totals itself.
#define RS_SCAN_THREADS RS_MAX_GRID
__global__ void __launch_bounds__(RS_SCAN_THREADS) radix_scan_kernel(uint32_t* __restrict__ counts, uint32_t B) {
  __shared__ uint32_t s_w[RS_SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t* row = counts + (size_t)blockIdx.x * B;
  const uint32_t v = threadIdx.x < B ? row[threadIdx.x] : 0u;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  uint32_t off = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < RS_SCAN_THREADS / 32; ++w) {
    const uint32_t t = s_w[w];
    off += (w < warp) ? t : 0u;
    tot += t;
  }
  if (threadIdx.x < B) row[threadIdx.x] = off + incl - v;
  if (threadIdx.x == 0) counts[(size_t)RS_BINS * B + blockIdx.x] = tot;
}

// vals_in == nullptr: the value of element i is i (first pass of an argsort)
template <typename KeyT, int RS_ITEMS>
__global__ void __launch_bounds__(RS_THREADS, 4)
radix_scatter_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ vals_out, uint32_t n_host, const uint32_t* __restrict__ n_dev, int shift,
                     const uint32_t* __restrict__ counts) {
  constexpr uint32_t RS_TILE = RS_THREADS * RS_ITEMS;
  __shared__ KeyT s_keys[RS_TILE];
  __shared__ uint32_t s_vals[RS_TILE];
  __shared__ uint32_t s_whist[RS_WARPS][RS_BINS + 1];
  __shared__ uint32_t s_run[RS_BINS];    // next global slot of every digit for this block
  __shared__ uint32_t s_gbase[RS_BINS];  // global slot - position in the reordered tile (mod 2^32)
  __shared__ uint32_t s_start[RS_BINS];  // first position of every digit in the reordered tile
  __shared__ uint32_t s_wsum[RS_WARPS];

  const uint32_t n = n_dev ? *n_dev : n_host;
  const RadixChunk c = radix_chunk(n, RS_TILE);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  {  // first global slot of (digit tid, this block) = all smaller digits + this digit in earlier blocks
    const uint32_t rowtot = counts[RS_BINS * gridDim.x + tid];
    uint32_t incl = rowtot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) off += (w < warp) ? s_wsum[w] : 0u;
    s_run[tid] = counts[tid * gridDim.x + blockIdx.x] + off + incl - rowtot;
    __syncthreads();
  }

  for (uint32_t t0 = c.begin; t0 < c.end; t0 += RS_TILE) {
    const uint32_t tile_n = min(RS_TILE, c.end - t0);
    for (int k = tid; k < RS_WARPS * (RS_BINS + 1); k += RS_THREADS) (&s_whist[0][0])[k] = 0;
    __syncthreads();

    // ---- load (warp w owns items [w*512, w*512+512) of the tile, 32 consecutive per round) + rank ----
    KeyT key[RS_ITEMS];
    uint32_t val[RS_ITEMS];
    uint16_t rank[RS_ITEMS];
    const uint32_t wbase = warp * (32 * RS_ITEMS) + lane;
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
      const uint32_t li = wbase + j * 32;
      const bool valid = li < tile_n;
      key[j] = valid ? keys_in[t0 + li] : (KeyT)0;
      val[j] = valid ? (vals_in ? vals_in[t0 + li] : t0 + li) : 0u;
    }
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
      const bool valid = wbase + j * 32 < tile_n;
      const uint32_t d = valid ? (((uint32_t)key[j] >> shift) & 0xFFu) : 256u;
      // lanes holding the same digit, from 9 ballots (8 digit bits + validity)
      uint32_t peers = __ballot_sync(0xffffffffu, valid);
      peers = valid ? peers : ~peers;
#pragma unroll
      for (int bit = 0; bit < 8; ++bit) {
        const bool one = (d >> bit) & 1u;
        const uint32_t bal = __ballot_sync(0xffffffffu, one);
        peers &= one ? bal : ~bal;
      }
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if (lane == leader) {
        old = s_whist[warp][d];
        s_whist[warp][d] = old + (uint32_t)__popc(peers);
      }
      old = __shfl_sync(0xffffffffu, old, leader);
      rank[j] = (uint16_t)(old + (uint32_t)__popc(peers & lt_mask));
      __syncwarp();
    }
    __syncthreads();

    // ---- per digit: exclusive prefix over warps, tile total, exclusive scan over digits ----
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
      const uint32_t t = s_whist[w][tid];
      s_whist[w][tid] = total;
      total += t;
    }
    uint32_t incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) woff += (w < warp) ? s_wsum[w] : 0u;
    const uint32_t start = woff + incl - total;
    s_start[tid] = start;
    const uint32_t run = s_run[tid];
    s_gbase[tid] = run - start;
    s_run[tid] = run + total;
    __syncthreads();

    // ---- reorder the tile by digit in shared memory ----
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
      if (wbase + j * 32 < tile_n) {
        const uint32_t d = ((uint32_t)key[j] >> shift) & 0xFFu;
        const uint32_t pos = s_start[d] + s_whist[warp][d] + rank[j];
        s_keys[pos] = key[j];
        s_vals[pos] = val[j];
      }
    }
    __syncthreads();

    // ---- digit-contiguous runs go out with coalesced stores ----
    for (uint32_t i = tid; i < tile_n; i += RS_THREADS) {
      const KeyT k = s_keys[i];
      const uint32_t d = ((uint32_t)k >> shift) & 0xFFu;
      const uint32_t dst = s_gbase[d] + i;
      if (keys_out) keys_out[dst] = k;
      vals_out[dst] = s_vals[i];
    }
    __syncthreads();
  }
}

// One full pass.  `counts` must hold 256 * (grid + 1) uint32.  keys_out may be null on the last pass when the
// sorted keys are not needed.
template <typename KeyT, int ITEMS>
static cudaError_t radix_pass(const KeyT* keys_in, KeyT* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                              uint32_t n_host, const uint32_t* n_dev, int shift, uint32_t* counts, int grid,
                              cudaStream_t stream) {
  radix_hist_kernel<KeyT><<<grid, RS_THREADS, 0, stream>>>(keys_in, n_host, n_dev, shift, RS_THREADS * ITEMS, counts);
  radix_scan_kernel<<<RS_BINS, RS_SCAN_THREADS, 0, stream>>>(counts, (uint32_t)grid);
  radix_scatter_kernel<KeyT, ITEMS><<<grid, RS_THREADS, 0, stream>>>(keys_in, keys_out, vals_in, vals_out, n_host, n_dev,
                                                                     shift, counts);
  return cudaGetLastError();
}
