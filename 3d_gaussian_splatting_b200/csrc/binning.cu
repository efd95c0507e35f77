// binning.cu — K2: from per-Gaussian tile rectangles to per-tile,
// depth-ordered point lists.  The sorted list is, bit for bit, the one a stable sort of the 64-bit
// keys (tile_id << 32 | fp32 bits of the view depth) over the duplicates emitted in Gaussian order
// produces (the oracle's definition) -- but the 64-bit keys are never materialised:
//
//   1. depth sort    argsort of the N Gaussians by depth bits (stable LSD radix, 4 passes over N pairs)
//   2. duplicate     the Gaussians emit their (tile id, Gaussian id) pairs IN DEPTH ORDER
//                    (offsets = exclusive scan of the rectangle areas along that order; K1 leaves an 8-B
//                    rectangle record per Gaussian so this stage gathers 8 B per Gaussian, nothing else)
//   3. tile sort     stable LSD radix sort of the D pairs by tile id only: ceil(log2(T) / 8) passes of
//                    6 B pairs (u16 tile ids) instead of 6 passes of 12 B pairs
//   4. ranges        per-tile binary search in the sorted tile ids
//
// Because both sorts are stable, the order inside a tile is (depth, Gaussian id) ascending = the
// oracle's.  Every element count that depends on the data (D) is read from device memory by the
// kernels, which run on fixed persistent grids: the stage needs no host round-trip by itself.
// HBM traffic per duplicate: 6 (emit) + passes * (2 + 6 + 6) B  vs  12 + 8 + 6 * 24 + 8 B for the
// one-level 64-bit sort.
#include "gs_common.cuh"
#include "radix_sort.cuh"

// ---- 1. depth sort ----------------------------------------------------------------------------
// hist: the four digit histograms taken by depth_key_kernel; status / tickets / tag: one-sweep bookkeeping
// (radix_sort.cuh).  Four launches; the result (Gaussian ids in depth order) lands in v1.
cudaError_t gs_launch_depth_sort(int N, const uint32_t* depth_bits, uint32_t* k0, uint32_t* k1, uint32_t* v0,
                                 uint32_t* v1, const uint32_t* hist, unsigned long long* status, uint32_t* tickets,
                                 uint32_t tag, int max_grid, int items, uint32_t sleep_ns, cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  const uint32_t n = (uint32_t)N;
#define GS_DEPTH_PASSES(IT)                                                                                              \
  do {                                                                                                                   \
    onesweep_pass<uint32_t, IT>(depth_bits, k0, nullptr, v0, n, nullptr, n, 0, hist, status, tickets + 0, tag + 0,      \
                                max_grid, sleep_ns, stream);                                                             \
    onesweep_pass<uint32_t, IT>(k0, k1, v0, v1, n, nullptr, n, 8, hist + 256, status, tickets + 1, tag + 1, max_grid,    \
                                sleep_ns, stream);                                                                       \
    onesweep_pass<uint32_t, IT>(k1, k0, v1, v0, n, nullptr, n, 16, hist + 512, status, tickets + 2, tag + 2, max_grid,   \
                                sleep_ns, stream);                                                                       \
    return onesweep_pass<uint32_t, IT>(k0, nullptr, v0, v1, n, nullptr, n, 24, hist + 768, status, tickets + 3, tag + 3, \
                                       max_grid, sleep_ns, stream);                                                      \
  } while (0)
  if (items == 16) GS_DEPTH_PASSES(16);
  GS_DEPTH_PASSES(8);
#undef GS_DEPTH_PASSES
}

// ---- 2. scan of tiles_touched along the depth order + duplicate ---------------------------------
#define DUP_THREADS 256
#define DUP_ITEMS 4
#define DUP_BLOCK (DUP_THREADS * DUP_ITEMS)

__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* s_warp, uint32_t& block_total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t off = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < DUP_THREADS / 32; ++w) {
    const uint32_t t = s_warp[w];
    off += (w < warp) ? t : 0u;
    tot += t;
  }
  block_total = tot;
  return off + incl - v;
}

// per-Gaussian record written by K1: {x0 | y0 << 16, x1 | y1 << 16, alive mask lo, hi}; the 64-bit mask (bit =
// row-major index inside the rectangle) is only meaningful for rectangles of <= 64 tiles, larger ones keep all
__device__ __forceinline__ uint32_t rect_area(uint4 r) {
  return ((r.y & 0xFFFFu) - (r.x & 0xFFFFu)) * ((r.y >> 16) - (r.x >> 16));
}
__device__ __forceinline__ uint32_t rect_alive(uint4 r) {
  const uint32_t area = rect_area(r);
  return area > GS_CULL_MAX_TILES ? area : (uint32_t)(__popc(r.z) + __popc(r.w));
}
__global__ void __launch_bounds__(DUP_THREADS)
dup_block_sums_kernel(int N, const uint32_t* __restrict__ order, const uint32_t* __restrict__ tiles_touched,
                      uint32_t* __restrict__ block_sums) {
  __shared__ uint32_t s_warp[DUP_THREADS / 32];
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < DUP_ITEMS; ++k) {
    const int r = blockIdx.x * DUP_BLOCK + k * DUP_THREADS + threadIdx.x;
    if (r < N) sum += __ldg(tiles_touched + order[r]);  // = rect_alive(rects[.]): K1 wrote both; 4-B gathers, 6.4 MB array
  }
  uint32_t tot;
  block_exclusive_scan_256(sum, s_warp, tot);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// one block: exclusive scan of the block sums in place; the grand total D goes to *total (64-bit, so
// a wrapped 32-bit scan is detectable) and, clamped to the capacity of the pair buffers, to *total32 for
// the kernels that index with it.  D > cap raises the overflow flag and the float guards (the step is
// then invalid: the Adam kernel skips its update when it sees the guard, see optim.cu).
__global__ void __launch_bounds__(1024)
dup_scan_sums_kernel(int nb, uint32_t* __restrict__ block_sums, unsigned long long* __restrict__ total,
                     uint32_t* __restrict__ total32, unsigned long long cap, uint32_t* __restrict__ overflow,
                     float* __restrict__ guard, float* __restrict__ mirror, uint32_t* __restrict__ host_status) {
  __shared__ unsigned long long s_warp[32];
  const int per = (nb + 1023) / 1024;
  const int lo = min(nb, (int)threadIdx.x * per), hi = min(nb, lo + per);
  unsigned long long sum = 0;
  for (int i = lo; i < hi; ++i) sum += block_sums[i];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned long long off = 0, tot = 0;
  for (int w = 0; w < 32; ++w) {
    const unsigned long long t = s_warp[w];
    off += (w < warp) ? t : 0ull;
    tot += t;
  }
  unsigned long long run = off + incl - sum;
  for (int i = lo; i < hi; ++i) {
    const uint32_t t = block_sums[i];
    block_sums[i] = (uint32_t)run;
    run += t;
  }
  if (threadIdx.x == 0) {
    *total = tot;
    *total32 = (uint32_t)(tot > cap ? cap : tot);
    if (tot > cap) {
      *overflow = 1u;
      *guard = 1.0f;
      if (mirror) *mirror = 1.0f;
    }
    // status words straight into the caller's pinned (device-mapped) host buffer: D (u64), overflow flag --
    // no copy operation on the stream
    if (host_status) {
      *reinterpret_cast<unsigned long long*>(host_status) = tot;
      host_status[2] = tot > cap ? 1u : 0u;
    }
  }
}

// index (inside the rectangle) of the k-th alive tile: binary descent over popcounts (the generic __fns costs more)
__device__ __forceinline__ uint32_t kth_set_bit64(uint32_t lo, uint32_t hi, uint32_t k) {
  const uint32_t c = (uint32_t)__popc(lo);
  uint32_t word = lo, pos = 0;
  if (k >= c) { k -= c; word = hi; pos = 32; }
  uint32_t t = (uint32_t)__popc(word & 0xFFFFu);
  if (k >= t) { k -= t; word >>= 16; pos += 16; }
  t = (uint32_t)__popc(word & 0xFFu);
  if (k >= t) { k -= t; word >>= 8; pos += 8; }
  t = (uint32_t)__popc(word & 0xFu);
  if (k >= t) { k -= t; word >>= 4; pos += 4; }
  t = (uint32_t)__popc(word & 0x3u);
  if (k >= t) { k -= t; word >>= 2; pos += 2; }
  if (k >= (word & 1u)) pos += 1;
  return pos;
}

// A block expands 1024 consecutive ranks of the depth order.  Thread t first scans the tile counts of
// ranks 4t .. 4t+3.  The block's output slots [0, total) are then produced in rounds of DUP_ROUND: every rank
// with tiles marks its first slot with (rank + 1), an inclusive max-scan over the marks (8 slots per thread, one
// 16-byte shared load / store each) turns them into the owner of every slot, and the slots are dealt round-robin
// to the threads: slot j reads its owner, derives its tile from the owner's rectangle and alive-tile mask (exact
// tile culling, gs_common.cuh), so both output arrays are written fully coalesced whatever the rectangle sizes are
// (no per-thread loops over big splats, and no per-slot search: the first version's 10-step binary search per slot
// was 2/3 of this kernel's instructions).  cap: capacity of the pair buffers (pairs past it are dropped and
// *overflow is raised -- only possible when the caller fixed the capacity in advance).
#define DUP_SLOTS 8
#define DUP_ROUND (DUP_THREADS * DUP_SLOTS)
template <typename KeyT>
__global__ void __launch_bounds__(DUP_THREADS)
duplicate_with_keys_kernel(int N, const uint32_t* __restrict__ order, const uint4* __restrict__ rects,
                           const uint32_t* __restrict__ block_base, KeyT* __restrict__ tile_ids,
                           uint32_t* __restrict__ values, int gx, uint32_t cap, uint32_t* __restrict__ overflow,
                           uint32_t* __restrict__ hist /* [sizeof(KeyT)][256]: digit histograms of the tile ids */) {
  // the tile sort's one-sweep passes take their digit histograms from here (radix_sort.cuh): the pairs are counted
  // by the kernel that writes them
  constexpr int NDIG = (int)sizeof(KeyT);
  __shared__ uint32_t s_hist[NDIG][256];
  __shared__ uint32_t s_warp[DUP_THREADS / 32];
  __shared__ uint32_t s_wmax[DUP_THREADS / 32];
  __shared__ uint32_t s_off[DUP_BLOCK];
  __shared__ uint32_t s_id[DUP_BLOCK];
  __shared__ uint4 s_rect[DUP_BLOCK];
  __shared__ __align__(16) uint16_t s_owner[DUP_ROUND];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int d = 0; d < NDIG; ++d) s_hist[d][tid] = 0;  // (ordered before the first use by the scan's barrier)
  const int r0 = blockIdx.x * DUP_BLOCK + tid * DUP_ITEMS;
  uint32_t cnt[DUP_ITEMS], myoff[DUP_ITEMS];
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < DUP_ITEMS; ++k) {
    uint32_t id = 0;
    uint4 rc = make_uint4(0u, 0u, 0u, 0u);
    if (r0 + k < N) {
      id = order[r0 + k];
      rc = __ldg(rects + id);
    }
    s_id[tid * DUP_ITEMS + k] = id;
    s_rect[tid * DUP_ITEMS + k] = rc;
    cnt[k] = rect_alive(rc);
    sum += cnt[k];
  }
  uint32_t total;
  uint32_t off = block_exclusive_scan_256(sum, s_warp, total);
#pragma unroll
  for (int k = 0; k < DUP_ITEMS; ++k) {
    s_off[tid * DUP_ITEMS + k] = off;
    myoff[k] = off;
    off += cnt[k];
  }
  const uint32_t base = block_base[blockIdx.x];
  uint32_t limit = total;
  if ((unsigned long long)base + total > cap) {
    limit = base < cap ? cap - base : 0u;  // (the scan kernel has raised the overflow flag)
  }
  uint32_t carry = 0;  // owner mark of the last slot of the previous round
  for (uint32_t lo = 0; lo < limit; lo += DUP_ROUND) {
    reinterpret_cast<uint4*>(s_owner)[tid] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();  // (also orders s_off / s_id / s_rect before the first round's reads)
#pragma unroll
    for (int k = 0; k < DUP_ITEMS; ++k) {
      const uint32_t rel = myoff[k] - lo;  // wraps for offsets below lo
      if (cnt[k] > 0 && rel < DUP_ROUND) s_owner[rel] = (uint16_t)(tid * DUP_ITEMS + k + 1);
    }
    __syncthreads();
    // inclusive max-scan of the marks (they increase with the slot): owner of a slot = last mark at or before it
    const uint4 m = reinterpret_cast<const uint4*>(s_owner)[tid];
    uint32_t e[DUP_SLOTS] = {m.x & 0xFFFFu, m.x >> 16, m.y & 0xFFFFu, m.y >> 16,
                             m.z & 0xFFFFu, m.z >> 16, m.w & 0xFFFFu, m.w >> 16};
#pragma unroll
    for (int i = 1; i < DUP_SLOTS; ++i) e[i] = max(e[i], e[i - 1]);
    uint32_t incl = e[DUP_SLOTS - 1];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl = max(incl, t);
    }
    uint32_t pre = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) pre = 0;
    if (lane == 31) s_wmax[warp] = incl;
    __syncthreads();
    uint32_t tot = carry;
    pre = max(pre, carry);
#pragma unroll
    for (int w = 0; w < DUP_THREADS / 32; ++w) {
      const uint32_t t = s_wmax[w];
      if (w < warp) pre = max(pre, t);
      tot = max(tot, t);
    }
    carry = tot;
#pragma unroll
    for (int i = 0; i < DUP_SLOTS; ++i) e[i] = max(e[i], pre);
    reinterpret_cast<uint4*>(s_owner)[tid] =
        make_uint4(e[0] | (e[1] << 16), e[2] | (e[3] << 16), e[4] | (e[5] << 16), e[6] | (e[7] << 16));
    __syncthreads();
#pragma unroll
    for (int i = 0; i < DUP_SLOTS; ++i) {
      const uint32_t jr = (uint32_t)tid + (uint32_t)i * DUP_THREADS;
      const uint32_t j = lo + jr;
      const bool valid = j < limit;
      uint32_t tile = 0;
      if (valid) {
        const uint32_t owner = (uint32_t)s_owner[jr] - 1u;
        const uint4 rc = s_rect[owner];
        uint32_t k = j - s_off[owner];
        const uint32_t x0 = rc.x & 0xFFFFu, y0 = rc.x >> 16, w = (rc.y & 0xFFFFu) - x0;
        uint32_t row;
        if (rect_area(rc) > GS_CULL_MAX_TILES) {
          row = k / w;
        } else {
          k = kth_set_bit64(rc.z, rc.w, k);
          // k < 64, w <= 64: (k + 0.5) / w is at least 1/128 away from an integer, far beyond the division's error
          row = __float2uint_rz(__fdividef((float)k + 0.5f, (float)w));
        }
        const uint32_t col = k - row * w;
        tile = (y0 + row) * (uint32_t)gx + x0 + col;
        tile_ids[base + j] = (KeyT)tile;
        values[base + j] = s_id[owner];
      }
      // digit histograms (slots, hence valid lanes, increase with the lane; a rectangle row is a run of high bytes)
      if (valid) atomicAdd(&s_hist[0][tile & 0xFFu], 1u);
#pragma unroll
      for (int d = 1; d < NDIG; ++d) warp_run_hist(s_hist[d], (tile >> (8 * d)) & 0xFFu, valid, lane);
    }
    __syncthreads();  // the next round clears s_owner
  }
  __syncthreads();
#pragma unroll
  for (int d = 0; d < NDIG; ++d) {
    const uint32_t c = s_hist[d][tid];
    if (c) atomicAdd(hist + d * 256 + tid, c);
  }
}

int gs_dup_num_blocks(int N) { return (N + DUP_BLOCK - 1) / DUP_BLOCK; }

cudaError_t gs_launch_dup_scan(int N, const uint32_t* order, const uint32_t* tiles_touched, uint32_t* block_sums,
                               unsigned long long* total, uint32_t* total32, unsigned long long cap,
                               uint32_t* overflow, float* guard, float* mirror, uint32_t* host_status,
                               cudaStream_t stream) {
  if (N == 0) return cudaMemsetAsync(total, 0, 8, stream);
  const int nb = gs_dup_num_blocks(N);
  dup_block_sums_kernel<<<nb, DUP_THREADS, 0, stream>>>(N, order, tiles_touched, block_sums);
  dup_scan_sums_kernel<<<1, 1024, 0, stream>>>(nb, block_sums, total, total32, cap, overflow, guard, mirror, host_status);
  return cudaGetLastError();
}

cudaError_t gs_launch_duplicate(int N, const uint32_t* order, const uint4* rects, const uint32_t* block_base,
                                void* tile_ids, int key_bytes, uint32_t* values, int gx, uint32_t cap,
                                uint32_t* overflow, uint32_t* hist, cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  const int nb = gs_dup_num_blocks(N);
  if (key_bytes == 2)
    duplicate_with_keys_kernel<uint16_t><<<nb, DUP_THREADS, 0, stream>>>(N, order, rects, block_base, (uint16_t*)tile_ids,
                                                                        values, gx, cap, overflow, hist);
  else
    duplicate_with_keys_kernel<uint32_t><<<nb, DUP_THREADS, 0, stream>>>(N, order, rects, block_base, (uint32_t*)tile_ids,
                                                                        values, gx, cap, overflow, hist);
  return cudaGetLastError();
}

// ---- 3. tile sort -------------------------------------------------------------------------------
// Sorts the D pairs by tile id.  Buffers a = input; the result is in `a` after an even number of
// passes and in `b` after an odd number (returned through *result_in_b).
cudaError_t gs_launch_tile_sort(const uint32_t* D_dev, uint32_t D_max, int tile_bits, int key_bytes, void* keys_a,
                                void* keys_b, uint32_t* vals_a, uint32_t* vals_b, const uint32_t* hist,
                                unsigned long long* status, uint32_t* tickets, uint32_t tag, int max_grid, int items,
                                uint32_t sleep_ns, int* result_in_b, cudaStream_t stream) {
  const int passes = (tile_bits + 7) / 8;
  bool in_a = true;
  for (int p = 0; p < passes; ++p) {
    uint32_t* vi = in_a ? vals_a : vals_b;
    uint32_t* vo = in_a ? vals_b : vals_a;
#define GS_TILE_PASS(KeyT, IT)                                                                                         \
  onesweep_pass<KeyT, IT>((const KeyT*)(in_a ? keys_a : keys_b), (KeyT*)(in_a ? keys_b : keys_a), vi, vo, 0u, D_dev,   \
                          D_max, 8 * p, hist + 256 * p, status, tickets + p, tag + (uint32_t)p, max_grid, sleep_ns, stream)
    if (key_bytes == 2) {
      if (items == 16) GS_TILE_PASS(uint16_t, 16);
      else GS_TILE_PASS(uint16_t, 8);
    } else {
      if (items == 16) GS_TILE_PASS(uint32_t, 16);
      else GS_TILE_PASS(uint32_t, 8);
    }
#undef GS_TILE_PASS
    in_a = !in_a;
  }
  *result_in_b = in_a ? 0 : 1;
  return cudaGetLastError();
}

// 64-bit status words the one-sweep passes need for a list of up to n pairs in tiles of 256 * items pairs
size_t gs_onesweep_status_words(size_t n, int items) {
  const size_t tile = (size_t)RS_THREADS * items;
  return ((n + tile - 1) / tile + 1) * RS_BINS;
}

// ---- 4. tile ranges -----------------------------------------------------------------------------
// One WARP per tile boundary: a 33-ary search (32 probes per round, one ballot) finds the first list position of tile
// t in ~log33(D) = 5 dependent load rounds instead of the ~23 of a per-thread binary search (the kernel is pure
// latency: 13 -> 4 us).  A block of 32 warps resolves 31 tiles: the range of a tile is the pair of its own and its
// right neighbour's first position.
template <typename KeyT>
__device__ __forceinline__ uint32_t warp_lower_bound_tile(const KeyT* __restrict__ keys, uint32_t n, uint32_t t, int lane) {
  uint32_t lo = 0, hi = n;  // the answer lies in [lo, hi]
  while (hi - lo > 32u) {
    const unsigned long long span = hi - lo;
    const uint32_t p = lo + (uint32_t)((span * (unsigned long long)(lane + 1)) / 33ull);  // lo < p < hi, increasing in lane
    const uint32_t c = (uint32_t)__popc(__ballot_sync(0xffffffffu, (uint32_t)keys[p] < t));  // probes below t: a prefix
    const uint32_t p_below = lo + (uint32_t)((span * (unsigned long long)c) / 33ull);        // probe c - 1 (c >= 1)
    const uint32_t p_above = lo + (uint32_t)((span * (unsigned long long)(c + 1)) / 33ull);  // probe c     (c <= 31)
    if (c < 32u) hi = p_above;
    if (c > 0u) lo = p_below + 1u;
  }
  const bool below = lo + (uint32_t)lane < hi && (uint32_t)keys[lo + lane] < t;
  return lo + (uint32_t)__popc(__ballot_sync(0xffffffffu, below));
}
template <typename KeyT>
__global__ void __launch_bounds__(1024)
identify_tile_ranges_kernel(const uint32_t* __restrict__ D_dev, const KeyT* __restrict__ keys, int ntiles,
                            uint2* __restrict__ ranges) {
  __shared__ uint32_t s_lb[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * 31 + warp;
  const uint32_t n = *D_dev;
  const uint32_t lb = t <= ntiles ? warp_lower_bound_tile(keys, n, (uint32_t)t, lane) : n;
  if (lane == 0) s_lb[warp] = lb;
  __syncthreads();
  if (threadIdx.x >= 31) return;
  const int tt = blockIdx.x * 31 + (int)threadIdx.x;
  if (tt >= ntiles) return;
  const uint32_t a = s_lb[threadIdx.x], b = s_lb[threadIdx.x + 1];
  ranges[tt] = a == b ? make_uint2(0u, 0u) : make_uint2(a, b);  // empty tiles read (0,0), as in the oracle
}

cudaError_t gs_launch_ranges(const uint32_t* D_dev, const void* keys, int key_bytes, uint2* ranges, int ntiles,
                             cudaStream_t stream) {
  if (ntiles == 0) return cudaSuccess;
  const int blocks = (ntiles + 30) / 31;
  if (key_bytes == 2)
    identify_tile_ranges_kernel<uint16_t><<<blocks, 1024, 0, stream>>>(D_dev, (const uint16_t*)keys, ntiles, ranges);
  else
    identify_tile_ranges_kernel<uint32_t><<<blocks, 1024, 0, stream>>>(D_dev, (const uint32_t*)keys, ntiles, ranges);
  return cudaGetLastError();
}

// ---- tile scheduling order: longest lists first --------------------------------------------------
// Per-tile work is extremely skewed (mean ~300 processed entries, max several thousand), and the
// blend kernels run one block per tile, so launching the heavy tiles first (longest-processing-time
// order) removes the tail.  Counting sort by a 7-bit log-scale bucket of the list length; one block.
#define ORDER_BUCKETS 128
__device__ __forceinline__ int length_bucket(uint32_t len) {
  if (len < 4) return (int)len;
  const int msb = 31 - __clz(len);                      // >= 2
  const int sub = (int)((len >> (msb - 2)) & 3u);       // next two bits
  return min(ORDER_BUCKETS - 1, 4 * (msb - 1) + sub);   // monotone in len
}
__global__ void __launch_bounds__(1024)
order_tiles_kernel(int ntiles, const uint2* __restrict__ ranges, uint32_t* __restrict__ tile_order) {
  __shared__ uint32_t s_count[ORDER_BUCKETS];
  __shared__ uint32_t s_start[ORDER_BUCKETS];
  for (int b = threadIdx.x; b < ORDER_BUCKETS; b += blockDim.x) s_count[b] = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < ntiles; t += blockDim.x) {
    const uint2 r = ranges[t];
    atomicAdd(&s_count[length_bucket(r.y - r.x)], 1u);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t acc = 0;
    for (int b = ORDER_BUCKETS - 1; b >= 0; --b) {  // descending: heaviest bucket first
      s_start[b] = acc;
      acc += s_count[b];
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < ntiles; t += blockDim.x) {
    const uint2 r = ranges[t];
    const uint32_t pos = atomicAdd(&s_start[length_bucket(r.y - r.x)], 1u);
    tile_order[pos] = (uint32_t)t;
  }
}

cudaError_t gs_launch_order_tiles(int ntiles, const uint2* ranges, uint32_t* tile_order, cudaStream_t stream) {
  if (ntiles == 0) return cudaSuccess;
  order_tiles_kernel<<<1, 1024, 0, stream>>>(ntiles, ranges, tile_order);
  return cudaGetLastError();
}
