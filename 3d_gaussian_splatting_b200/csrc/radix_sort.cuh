// radix_sort.cuh — this library's own stable LSD radix sort (8-bit digits) for (key, value) pairs, used twice by the
// binning stage (binning.cu):
//   * Gaussians by view depth   (u32 keys = fp32 bits of the depth, 4 passes over N pairs)
//   * duplicates by tile id     (u16/u32 keys, ceil(log2(T)/8) passes over D pairs)
// ONE launch per pass ("one-sweep": chained scan with decoupled look-back, below); the digit histograms of all
// passes are taken by the kernels that write the keys.  The element count may live in device memory (fixed-capacity
// mode: no host round-trip).  A tile of 2048 / 4096 pairs is ranked stably (ballot-built peer masks + per-warp
// counters, then a prefix over warps -- match.any is avoided on purpose: measured at ~64 SM cycles per instruction on
// sm_100), reordered by digit in shared memory and written as digit-contiguous runs, so global stores coalesce.
// Stability is what makes the two-level scheme exact: pairs that compare equal keep their input order.
// (Round 1 ran three launches per pass -- per-block histograms, a scan kernel, the scatter: 13 launches for the depth
//  argsort, 6 for the tile sort, each bound by launch latency on data that fits in L2.)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_BINS 256
#define RS_MAX_GRID 1024  // persistent grid size limit

// ---- one-sweep pass: chained scan with decoupled look-back (no separate histogram / scan launches) -------------
// The per-pass digit histograms of ALL passes are produced by the kernels that WRITE the keys (depth_key_kernel for
// the depth argsort, duplicate_with_keys_kernel for the tile sort): a digit histogram does not depend on the order of
// the keys, so it can be taken before the first pass.  One pass is then ONE launch: a block takes the next 2048- or
// 4096-pair tile by ticket (so that every tile it will wait for has been started before it), ranks the tile stably as
// above, publishes its per-digit counts, looks back over the preceding tiles' published counts until it meets one
// that carries an inclusive prefix, publishes its own inclusive prefix, and scatters.  A status word is 64 bits --
// {tag << 2 | flag, value} -- written and read with single 8-byte accesses, and the tag (a fresh number per launch)
// means the status array never has to be cleared.
#define OS_FLAG_AGGREGATE 1u
#define OS_FLAG_INCLUSIVE 2u
__device__ __forceinline__ unsigned long long os_load(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void os_store(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Histogram update of one warp-wide batch of digits with run aggregation: a lane whose digit differs from its left
// neighbour's heads a run and adds the run's length with ONE shared-memory atomic.  Digits that come in runs (the
// high bytes of depths, the high byte of a rectangle's tile ids) would otherwise serialise 32 atomics on one bin.
// The valid lanes must form a prefix of the warp; every lane of the warp calls.
__device__ __forceinline__ void warp_run_hist(uint32_t* h, uint32_t dig, bool valid, int lane) {
  const uint32_t prev = __shfl_up_sync(0xffffffffu, dig, 1);
  const bool head = valid && (lane == 0 || prev != dig);
  const uint32_t heads = __ballot_sync(0xffffffffu, head);
  const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
  if (head) {
    const uint32_t after = (heads | ~vmask) & ~((2u << lane) - 1u);  // next head, or first invalid lane, above this one
    const int end = after ? __ffs(after) - 1 : 32;
    atomicAdd(&h[dig], (uint32_t)(end - lane));
  }
}

template <typename KeyT, int RS_ITEMS>
__global__ void __launch_bounds__(RS_THREADS, 4)
onesweep_scatter_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out, const uint32_t* __restrict__ vals_in,
                        uint32_t* __restrict__ vals_out, uint32_t n_host, const uint32_t* __restrict__ n_dev, int shift,
                        const uint32_t* __restrict__ digit_hist, unsigned long long* __restrict__ status,
                        uint32_t* __restrict__ ticket, uint32_t tag, uint32_t sleep_ns) {
  constexpr uint32_t RS_TILE = RS_THREADS * RS_ITEMS;
  __shared__ KeyT s_keys[RS_TILE];
  __shared__ uint32_t s_vals[RS_TILE];
  __shared__ uint32_t s_whist[RS_WARPS][RS_BINS + 1];
  __shared__ uint32_t s_dstart[RS_BINS];  // first global slot of every digit (exclusive scan of the pass's histogram)
  __shared__ uint32_t s_gbase[RS_BINS];   // global slot - position in the reordered tile (mod 2^32)
  __shared__ uint32_t s_start[RS_BINS];   // first position of every digit in the reordered tile
  __shared__ uint32_t s_wsum[RS_WARPS];
  __shared__ uint32_t s_tile;

  const uint32_t n = n_dev ? *n_dev : n_host;
  const uint32_t ntiles = (n + RS_TILE - 1) / RS_TILE;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t cnt0 = digit_hist[tid];
  if (__syncthreads_or(cnt0 == n && n > 0)) {
    // every key carries the same digit (the top byte of the depths of a scene a few octaves deep, the high byte of
    // the tile ids of a small image): a stable pass over it is the identity -- copy the tiles through
    while (true) {
      if (tid == 0) s_tile = atomicAdd(ticket, 1u);
      __syncthreads();
      const uint32_t tile = s_tile;
      __syncthreads();
      if (tile >= ntiles) return;
      const uint32_t t0 = tile * RS_TILE, tile_n = min(RS_TILE, n - t0);
      for (uint32_t i = tid; i < tile_n; i += RS_THREADS) {
        if (keys_out) keys_out[t0 + i] = keys_in[t0 + i];
        vals_out[t0 + i] = vals_in ? vals_in[t0 + i] : t0 + i;
      }
    }
  }
  {  // exclusive scan over the digits of this pass's histogram
    const uint32_t cnt = cnt0;
    uint32_t incl = cnt;
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
    s_dstart[tid] = off + incl - cnt;
    __syncthreads();
  }

  while (true) {
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int k = tid; k < RS_WARPS * (RS_BINS + 1); k += RS_THREADS) (&s_whist[0][0])[k] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= ntiles) return;
    const uint32_t t0 = tile * RS_TILE;
    const uint32_t tile_n = min(RS_TILE, n - t0);

    // ---- load (warp w owns items [w*32*ITEMS, (w+1)*32*ITEMS) of the tile, 32 consecutive per round) + rank ----
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

    // ---- per digit (thread = digit): prefix over warps, tile total; publish, look back, publish ----
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
      const uint32_t t = s_whist[w][tid];
      s_whist[w][tid] = total;
      total += t;
    }
    unsigned long long* my = status + (size_t)tile * RS_BINS + tid;
    const unsigned long long hi = (unsigned long long)tag << 34;
    uint32_t excl = 0;
    if (tile == 0) {
      os_store(my, hi | ((unsigned long long)OS_FLAG_INCLUSIVE << 32) | total);
    } else {
      os_store(my, hi | ((unsigned long long)OS_FLAG_AGGREGATE << 32) | total);
      // Look back over the predecessors, nearest first, EIGHT status words per round (independent loads: one L2
      // latency per round instead of one per tile -- tiles that start together see ~sqrt(2 * tiles) aggregate-only
      // predecessors before the first inclusive one, and the serial walk was most of a pass, profiles/r02g).
      const unsigned long long* look = my - RS_BINS;
      int remaining = (int)tile;  // predecessors not yet examined (tile 0 always carries an inclusive prefix)
      bool done = false;
      while (!done) {
        unsigned long long sv[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) sv[k] = k < remaining ? os_load(look - (size_t)k * RS_BINS) : 0ull;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (!done && k < remaining) {
            while ((sv[k] >> 34) != (unsigned long long)tag) {  // not published yet in this pass
              if (sleep_ns) __nanosleep(sleep_ns);
              sv[k] = os_load(look - (size_t)k * RS_BINS);
            }
            excl += (uint32_t)sv[k];
            done = ((uint32_t)(sv[k] >> 32) & 3u) == OS_FLAG_INCLUSIVE;
          }
        }
        look -= 8 * RS_BINS;
        remaining -= 8;
      }
      os_store(my, hi | ((unsigned long long)OS_FLAG_INCLUSIVE << 32) | (unsigned long long)(excl + total));
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
    s_gbase[tid] = s_dstart[tid] + excl - start;
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

// One one-sweep pass = one launch.  digit_hist: the 256 counts of this pass's digit over all n keys; status: 256
// 64-bit words per tile of RS_THREADS * ITEMS pairs (never cleared, see above; zeroed once when allocated);
// ticket: one zeroed counter per launch; tag: a number no earlier launch on this status array has used.
template <typename KeyT, int ITEMS>
static cudaError_t onesweep_pass(const KeyT* keys_in, KeyT* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                                 uint32_t n_host, const uint32_t* n_dev, uint32_t n_max, int shift,
                                 const uint32_t* digit_hist, unsigned long long* status, uint32_t* ticket, uint32_t tag,
                                 int max_grid, uint32_t sleep_ns, cudaStream_t stream) {
  const uint32_t tile = RS_THREADS * ITEMS;
  long long grid = ((long long)n_max + tile - 1) / tile;
  if (grid > max_grid) grid = max_grid;
  if (grid < 1) grid = 1;
  onesweep_scatter_kernel<KeyT, ITEMS><<<(int)grid, RS_THREADS, 0, stream>>>(keys_in, keys_out, vals_in, vals_out, n_host,
                                                                            n_dev, shift, digit_hist, status, ticket, tag, sleep_ns);
  return cudaGetLastError();
}
