// binning.cu -- instance emission, tile-digit histograms and tile ranges (K4, K6 of SURVEY 2.2).
//
// The reference emits (tile<<32 | depth_bits, id) pairs in Gaussian-id order (duplicateWithKeys,
// rasterizer_impl.cu:70-111), sorts them on 32+log2(tiles) bits (:303-311) and finds tile boundaries
// (identifyTileRanges, :116-138).  A stable sort on (tile, depth) with ties in emission order is the total
// order (tile, depth bits, Gaussian id), so the same point_list is obtained with far less traffic by
//   1. sorting the P Gaussians ONCE by (depth bits, id)          -- P-level, 4 radix passes (radix_sort.cu);
//   2. gathering their tile rectangles into that order + chunk sums (here), scan (preprocess.cu);
//   3. emitting the instances of the Gaussians in that order     -- here; only the tile id is written as key,
//      and the digit histograms of the tile sort are accumulated on the fly;
//   4. a stable 1-2 pass radix sort on the tile id alone         -- R-level, 16 B per instance and pass
//      instead of 6 passes x 24 B;
//   5. reading the tile boundaries off the sorted tile ids       -- here.
// The 64-bit keys of the reference are never materialised on this path (gsr_debug_sorted_keys rebuilds
// them for the bit-exact parity tests).  Every R-level kernel takes the instance count from DEVICE memory
// (total) bounded by the caller's capacity, so the forward can be enqueued without a host round trip.
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int DUP_THREADS = 256;   // chunk size of chunk_offsets (num_pre_blocks)
constexpr uint32_t COOP_THRESHOLD = 32;

GSR_DEVINL uint32_t rect_count(uint2 r) {  // r.x = x0 | x1 << 16, r.y = y0 | y1 << 16
  return ((r.x >> 16) - (r.x & 0xffffu)) * ((r.y >> 16) - (r.y & 0xffffu));
}

// Tile rectangles into depth order + instance count of every 256-Gaussian chunk of the depth-ordered list.
__global__ void __launch_bounds__(256)
gather_rects_kernel(int P, const uint32_t* __restrict__ order, const uint2* __restrict__ rect,
                    uint2* __restrict__ rect_sorted, uint32_t* __restrict__ chunk_sums,
                    uint32_t* __restrict__ chunk_sums_c /* bin-level instance counts (two-level binning) or NULL */) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_w[8], s_c[8];
  const int s = blockIdx.x * 256 + threadIdx.x;
  uint32_t v = 0, vc = 0;
  if (s < P) {
    const uint2 r = rect[order[s]];
    rect_sorted[s] = r;
    v = rect_count(r);
    vc = rect_count(ts_coarse_rect(r));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v += __shfl_xor_sync(0xffffffffu, v, o);
    vc += __shfl_xor_sync(0xffffffffu, vc, o);
  }
  if ((threadIdx.x & 31) == 0) { s_w[threadIdx.x >> 5] = v; s_c[threadIdx.x >> 5] = vc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0, tc = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { t += s_w[w]; tc += s_c[w]; }
    chunk_sums[blockIdx.x] = t;
    if (chunk_sums_c) chunk_sums_c[blockIdx.x] = tc;
  }
}

// One thread per Gaussian of the depth-ordered list; tiles of a Gaussian in row-major order like the
// reference loop (rasterizer_impl.cu:98-109).  Gaussians covering many tiles are emitted by their whole warp.
// hist[pass][256] accumulates the digit counts of the tile sort (shift1 = bits of digit 0; n_passes 1 or 2).
__global__ void __launch_bounds__(DUP_THREADS)
emit_instances_kernel(int P, const uint32_t* __restrict__ order, const uint2* __restrict__ rect_sorted,
                      const uint32_t* __restrict__ chunk_offsets, int tiles_x, long long cap, int n_passes,
                      uint32_t mask0, int shift1, uint32_t mask1, uint32_t* __restrict__ tkeys,
                      uint32_t* __restrict__ vals, uint32_t* __restrict__ hist) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_warp[DUP_THREADS / 32];
  __shared__ uint32_t s_hist[2 * SORT_RADIX];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int i = tid; i < 2 * SORT_RADIX; i += DUP_THREADS) s_hist[i] = 0;
  const int s = blockIdx.x * DUP_THREADS + tid;
  uint2 r = make_uint2(0, 0);
  uint32_t idx = 0;
  if (s < P) { r = rect_sorted[s]; idx = order[s]; }
  const uint32_t x0 = r.x & 0xffffu, x1 = r.x >> 16, y0 = r.y & 0xffffu, y1 = r.y >> 16;
  const uint32_t n = (x1 - x0) * (y1 - y0);

  uint32_t inc = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  uint32_t woff = 0;
#pragma unroll
  for (int ww = 0; ww < DUP_THREADS / 32; ++ww) woff += (ww < w) ? s_warp[ww] : 0u;
  const long long off = (long long)chunk_offsets[blockIdx.x] + woff + inc - n;

  auto put = [&](long long o, uint32_t tile, uint32_t id) {
    if (o < cap) {  // over capacity: the host re-runs with a larger workspace (GSR_ECAPACITY)
      tkeys[o] = tile;
      vals[o] = id;
      atomicAdd(&s_hist[tile & mask0], 1u);
      if (n_passes > 1) atomicAdd(&s_hist[SORT_RADIX + ((tile >> shift1) & mask1)], 1u);
    }
  };
  if (n > 0 && n <= COOP_THRESHOLD) {
    long long o = off;
    for (uint32_t y = y0; y < y1; ++y)
      for (uint32_t x = x0; x < x1; ++x) put(o++, y * tiles_x + x, idx);
  }
  uint32_t big = __ballot_sync(0xffffffffu, n > COOP_THRESHOLD);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t bn = __shfl_sync(0xffffffffu, n, src);
    const long long boff = __shfl_sync(0xffffffffu, off, src);
    const uint32_t bx0 = __shfl_sync(0xffffffffu, x0, src);
    const uint32_t by0 = __shfl_sync(0xffffffffu, y0, src);
    const uint32_t bx1 = __shfl_sync(0xffffffffu, x1, src);
    const uint32_t bidx = __shfl_sync(0xffffffffu, idx, src);
    const uint32_t bw = bx1 - bx0;
    for (uint32_t k = lane; k < bn; k += 32) put(boff + k, (by0 + k / bw) * tiles_x + (bx0 + k % bw), bidx);
  }
  __syncthreads();
  for (int i = tid; i < n_passes * SORT_RADIX; i += DUP_THREADS)
    if (s_hist[i]) atomicAdd(&hist[i], s_hist[i]);
}

// identifyTileRanges (rasterizer_impl.cu:116-138) on the sorted tile ids; `ranges` is zeroed beforehand so
// empty tiles read (0,0) exactly like the reference's cudaMemset (:313). Four instances per thread.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(long long cap, const unsigned long long* __restrict__ n_dev,
                   const uint32_t* __restrict__ tkeys_sorted, uint2* __restrict__ ranges) {
  pdl_launch_dependents();
  pdl_wait();
  long long L = cap;
  if (n_dev) L = min((long long)*n_dev, cap);
  const long long i0 = ((long long)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i0 >= L) return;
  uint32_t k[4];
  if (i0 + 4 <= L) {
    const uint4 v = *reinterpret_cast<const uint4*>(tkeys_sorted + i0);
    k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) k[j] = (i0 + j < L) ? tkeys_sorted[i0 + j] : 0u;
  }
  uint32_t prev = i0 > 0 ? tkeys_sorted[i0 - 1] : 0u;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const long long i = i0 + j;
    if (i < L) {
      const uint32_t cur = k[j];
      if (i == 0) ranges[cur].x = 0;
      else if (cur != prev) { ranges[prev].y = (uint32_t)i; ranges[cur].x = (uint32_t)i; }
      if (i == L - 1) ranges[cur].y = (uint32_t)L;
      prev = cur;
    }
  }
}

// Test hook: rebuild the reference's sorted 64-bit keys = (tile << 32) | depth bits.
__global__ void __launch_bounds__(256)
rebuild_keys_kernel(long long L, const uint32_t* __restrict__ tkeys_sorted, const uint32_t* __restrict__ point_list,
                    const float4* __restrict__ rec, uint64_t* __restrict__ keys) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= L) return;
  keys[i] = ((uint64_t)tkeys_sorted[i] << 32) | __float_as_uint(rec[3 * (size_t)point_list[i]].z);
}

cudaError_t launch_gather_rects(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g) {
  count_launch(); launch_k(gather_rects_kernel, num_pre_blocks(P), 256, 0, st, P, order, g.rect, g.rect_sorted, g.chunk_offsets,
           g.chunk_offsets_c);
  return cudaGetLastError();
}

cudaError_t launch_emit_instances(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g, int tiles_x,
                                  long long cap, const SortPlan& plan, uint32_t* tkeys, uint32_t* vals,
                                  uint32_t* hist) {
  const uint32_t mask0 = (1u << plan.bits[0]) - 1u;
  const int shift1 = plan.n_passes > 1 ? plan.shift[1] : 0;
  const uint32_t mask1 = plan.n_passes > 1 ? (1u << plan.bits[1]) - 1u : 0u;
  count_launch(); launch_k(emit_instances_kernel, num_pre_blocks(P), DUP_THREADS, 0, st, 
      P, order, g.rect_sorted, g.chunk_offsets, tiles_x, cap, plan.n_passes, mask0, shift1, mask1, tkeys, vals, hist);
  return launch_digit_offsets(st, hist, plan.n_passes);
}

cudaError_t launch_tile_ranges(cudaStream_t st, long long cap, const unsigned long long* n_dev,
                               const uint32_t* tkeys_sorted, uint2* ranges, size_t n_tiles) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, sizeof(uint2) * n_tiles, st);
  if (e != cudaSuccess || cap <= 0) return e;
  count_launch(); launch_k(tile_ranges_kernel, (unsigned)((cap + 1023) / 1024), 256, 0, st, cap, n_dev, tkeys_sorted, ranges);
  return cudaGetLastError();
}

cudaError_t launch_rebuild_keys(cudaStream_t st, long long L, const uint32_t* tkeys_sorted,
                                const uint32_t* point_list, const float4* rec, uint64_t* keys) {
  if (L <= 0) return cudaSuccess;
  count_launch(); launch_k(rebuild_keys_kernel, (unsigned)((L + 255) / 256), 256, 0, st, L, tkeys_sorted, point_list, rec, keys);
  return cudaGetLastError();
}

}  // namespace gsr
