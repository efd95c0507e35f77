// binning.cu — K2: inclusive scan of tiles_touched,
// duplicateWithKeys (key = tile_id << 32 | fp32 bits of the view depth, value = Gaussian id),
// stable radix sort restricted to the significant bits, identifyTileRanges.
// All outputs are integers and bit-exact against the oracle.
#include <cub/cub.cuh>

#include "gs_common.cuh"

// ---- K2a: duplicateWithKeys ------------------------------------------------------------------
// One thread per Gaussian for small rectangles; rectangles wider than a warp's worth of tiles are
// emitted cooperatively by the whole warp so one huge splat does not serialise a thread.
__global__ void __launch_bounds__(256)
duplicate_with_keys_kernel(int N, const GsGeom* __restrict__ geom, const float* __restrict__ depths,
                           const int* __restrict__ radii, const uint32_t* __restrict__ offsets,
                           uint64_t* __restrict__ keys, uint32_t* __restrict__ values, int gx, int gy) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  int rad = 0;
  uint32_t off = 0, dbits = 0;
  int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
  if (i < N) {
    rad = radii[i];
    if (rad > 0) {
      off = (i == 0) ? 0u : offsets[i - 1];
      const float4 q0 = geom[i].q0;
      const float px = q0.x, py = q0.y;
      x0 = min(gx, max(0, (int)__fdiv_rn(__fsub_rn(px, (float)rad), (float)GS_TILE)));
      y0 = min(gy, max(0, (int)__fdiv_rn(__fsub_rn(py, (float)rad), (float)GS_TILE)));
      x1 = min(gx, max(0, (int)__fdiv_rn(__fadd_rn(__fadd_rn(px, (float)rad), (float)(GS_TILE - 1)), (float)GS_TILE)));
      y1 = min(gy, max(0, (int)__fdiv_rn(__fadd_rn(__fadd_rn(py, (float)rad), (float)(GS_TILE - 1)), (float)GS_TILE)));
      dbits = __float_as_uint(depths[i]);
    }
  }
  const int area = (x1 - x0) * (y1 - y0);
  const bool big = area > 64;
  if (rad > 0 && !big) {
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
        values[off] = (uint32_t)i;
        ++off;
      }
  }
  // warp-cooperative emission of the big rectangles
  unsigned big_mask = __ballot_sync(0xffffffffu, big);
  while (big_mask) {
    const int src = __ffs(big_mask) - 1;
    big_mask &= big_mask - 1;
    const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
    const int bx1 = __shfl_sync(0xffffffffu, x1, src), by1 = __shfl_sync(0xffffffffu, y1, src);
    const uint32_t boff = __shfl_sync(0xffffffffu, off, src), bd = __shfl_sync(0xffffffffu, dbits, src);
    const int bi = __shfl_sync(0xffffffffu, i, src);
    const int w = bx1 - bx0, total = w * (by1 - by0);
    for (int k = lane; k < total; k += 32) {
      const int y = by0 + k / w, x = bx0 + k % w;
      keys[boff + k] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | bd;
      values[boff + k] = (uint32_t)bi;
    }
  }
}

// ---- K2c: identifyTileRanges -------------------------------------------------------------------
__global__ void __launch_bounds__(256)
identify_tile_ranges_kernel(uint32_t D, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= D) return;
  const uint32_t cur = (uint32_t)(keys[i] >> 32);
  if (i == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = i;
      ranges[cur].x = i;
    }
  }
  if (i == D - 1) ranges[cur].y = D;
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

// ---- host-side launchers ---------------------------------------------------------------------
size_t gs_scan_temp_bytes(int N) {
  size_t bytes = 0;
  cub::DeviceScan::InclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, N);
  return bytes;
}
cudaError_t gs_launch_scan(void* temp, size_t temp_bytes, const uint32_t* tiles_touched, uint32_t* offsets, int N,
                           cudaStream_t stream) {
  return cub::DeviceScan::InclusiveSum(temp, temp_bytes, tiles_touched, offsets, N, stream);
}

cudaError_t gs_launch_duplicate(int N, const GsGeom* geom, const float* depths, const int* radii,
                                const uint32_t* offsets, uint64_t* keys, uint32_t* values, int gx, int gy,
                                cudaStream_t stream) {
  if (N == 0) return cudaSuccess;
  duplicate_with_keys_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, geom, depths, radii, offsets, keys, values,
                                                                  gx, gy);
  return cudaGetLastError();
}

size_t gs_sort_temp_bytes(uint32_t D, int end_bit) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)D, 0, end_bit);
  return bytes;
}
cudaError_t gs_launch_sort(void* temp, size_t temp_bytes, const uint64_t* keys_in, uint64_t* keys_out,
                           const uint32_t* vals_in, uint32_t* vals_out, uint32_t D, int end_bit,
                           cudaStream_t stream) {
  if (D == 0) return cudaSuccess;
  return cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, (int)D, 0, end_bit,
                                         stream);
}

cudaError_t gs_launch_ranges(uint32_t D, const uint64_t* keys, uint2* ranges, int ntiles, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, (size_t)ntiles * sizeof(uint2), stream);
  if (e != cudaSuccess || D == 0) return e;
  identify_tile_ranges_kernel<<<(D + 255) / 256, 256, 0, stream>>>(D, keys, ranges);
  return cudaGetLastError();
}
