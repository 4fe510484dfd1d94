// binning.cu -- instance emission, tile-digit histograms and tile ranges (K4, K6 of SURVEY 2.2).
//
// The reference emits (tile<<32 | depth_bits, id) pairs in Gaussian-id order (duplicateWithKeys,
// rasterizer_impl.cu:70-111), sorts them on 32+log2(tiles) bits (:303-311) and finds tile boundaries
// (identifyTileRanges, :116-138).  A stable sort on (tile, depth) with ties in emission order is the total
// order (tile, depth bits, Gaussian id), so the same point_list is obtained with far less traffic by
//   1. sorting the P Gaussians ONCE by (depth bits, id)          -- P-level, 4 radix passes (radix_sort.cu);
//   2. emitting the instances of the Gaussians in that order     -- here; only the tile id is written as key;
//   3. a stable 1-2 pass radix sort on the tile id alone         -- R-level, 16 B per instance and pass
//      instead of 6 passes x 24 B;
//   4. reading the tile boundaries off the sorted tile ids       -- here.
// The 64-bit keys of the reference are never materialised on this path (gsr_debug_sorted_keys rebuilds
// them for the bit-exact parity tests).
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int DUP_THREADS = 256;   // chunk size of chunk_offsets (num_pre_blocks)
constexpr uint32_t COOP_THRESHOLD = 32;

// getRect (auxiliary.h:46-56), same operation order as preprocess.cu::tile_rect
GSR_DEVINL void rect_from_rec(float px, float py, int radius, int tiles_x, int tiles_y, uint2& rmin, uint2& rmax) {
  const float r = (float)radius;
  int x0 = (int)fmul(fsub(px, r), 0.0625f);
  int y0 = (int)fmul(fsub(py, r), 0.0625f);
  int x1 = (int)fmul(fadd(fadd(fadd(px, r), 16.0f), -1.0f), 0.0625f);
  int y1 = (int)fmul(fadd(fadd(fadd(py, r), 16.0f), -1.0f), 0.0625f);
  rmin.x = min((unsigned)tiles_x, (unsigned)max(0, x0));
  rmin.y = min((unsigned)tiles_y, (unsigned)max(0, y0));
  rmax.x = min((unsigned)tiles_x, (unsigned)max(0, x1));
  rmax.y = min((unsigned)tiles_y, (unsigned)max(0, y1));
}

// One thread per Gaussian of the depth-ordered list; tiles of a Gaussian in row-major order like the
// reference loop (rasterizer_impl.cu:98-109).  Gaussians covering many tiles are emitted by their whole warp.
__global__ void __launch_bounds__(DUP_THREADS)
emit_instances_kernel(int P, const uint32_t* __restrict__ order, const float4* __restrict__ rec_a,
                      const int* __restrict__ radii, const uint32_t* __restrict__ tiles_touched,
                      const uint32_t* __restrict__ chunk_offsets, int tiles_x, int tiles_y,
                      uint32_t* __restrict__ tkeys, uint32_t* __restrict__ vals) {
  __shared__ uint32_t s_warp[DUP_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int s = blockIdx.x * DUP_THREADS + tid;
  const uint32_t idx = s < P ? order[s] : 0u;
  const uint32_t n = s < P ? tiles_touched[idx] : 0u;

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
  const uint32_t off = chunk_offsets[blockIdx.x] + woff + inc - n;

  uint2 rmin = make_uint2(0, 0), rmax = make_uint2(0, 0);
  if (n > 0) {
    const float4 a = rec_a[idx];
    rect_from_rec(a.x, a.y, radii[idx], tiles_x, tiles_y, rmin, rmax);
  }
  if (n > 0 && n <= COOP_THRESHOLD) {
    uint32_t o = off;
    for (unsigned y = rmin.y; y < rmax.y; ++y)
      for (unsigned x = rmin.x; x < rmax.x; ++x) {
        tkeys[o] = y * tiles_x + x;
        vals[o] = idx;
        ++o;
      }
  }
  uint32_t big = __ballot_sync(0xffffffffu, n > COOP_THRESHOLD);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t bn = __shfl_sync(0xffffffffu, n, src);
    const uint32_t boff = __shfl_sync(0xffffffffu, off, src);
    const uint32_t bx0 = __shfl_sync(0xffffffffu, rmin.x, src);
    const uint32_t by0 = __shfl_sync(0xffffffffu, rmin.y, src);
    const uint32_t bx1 = __shfl_sync(0xffffffffu, rmax.x, src);
    const uint32_t bidx = __shfl_sync(0xffffffffu, idx, src);
    const uint32_t bw = bx1 - bx0;
    for (uint32_t k = lane; k < bn; k += 32) {
      tkeys[boff + k] = (by0 + k / bw) * tiles_x + (bx0 + k % bw);
      vals[boff + k] = bidx;
    }
  }
}

// Histograms of the two tile-id digits over the emitted instances. Consecutive instances belong to the same
// Gaussian, so the high digit is nearly warp-uniform: it is aggregated with match.any before the shared atomic.
__global__ void __launch_bounds__(256)
tile_digit_hist_kernel(const uint32_t* __restrict__ tkeys, long long n, int n_passes, uint32_t* __restrict__ hist) {
  __shared__ uint32_t s[2 * 256];
  for (int i = threadIdx.x; i < 512; i += 256) s[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * 256;
  const long long n_round = (n + 31) / 32 * 32;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n_round; i += stride) {
    const bool ok = i < n;
    const uint32_t k = ok ? tkeys[i] : 0xFFFFFFFFu;
    if (ok) atomicAdd(&s[k & 255u], 1u);
    if (n_passes > 1) {
      const uint32_t d1 = ok ? ((k >> 8) & 255u) : 0x100u;
      const uint32_t peers = __match_any_sync(0xffffffffu, d1);
      if (ok && (__ffs(peers) - 1) == lane) atomicAdd(&s[256 + d1], (uint32_t)__popc(peers));
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_passes * 256; i += 256)
    if (s[i]) atomicAdd(&hist[i], s[i]);
}

__global__ void tile_digit_offsets_kernel(uint32_t* hist, int n_passes) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (w >= n_passes) return;
  uint32_t run = 0;
  for (int c = 0; c < 8; ++c) {
    uint32_t v = hist[w * 256 + c * 32 + lane];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    hist[w * 256 + c * 32 + lane] = run + inc - v;
    run += __shfl_sync(0xffffffffu, inc, 31);
  }
}

// identifyTileRanges (rasterizer_impl.cu:116-138) on the sorted tile ids; `ranges` is zeroed beforehand so
// empty tiles read (0,0) exactly like the reference's cudaMemset (:313).
__global__ void __launch_bounds__(256)
tile_ranges_kernel(long long L, const uint32_t* __restrict__ tkeys_sorted, uint2* __restrict__ ranges) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= L) return;
  const uint32_t cur = tkeys_sorted[i];
  if (i == 0) ranges[cur].x = 0;
  else {
    const uint32_t prev = tkeys_sorted[i - 1];
    if (cur != prev) { ranges[prev].y = (uint32_t)i; ranges[cur].x = (uint32_t)i; }
  }
  if (i == L - 1) ranges[cur].y = (uint32_t)L;
}

// Test hook: rebuild the reference's sorted 64-bit keys = (tile << 32) | depth bits.
__global__ void __launch_bounds__(256)
rebuild_keys_kernel(long long L, const uint32_t* __restrict__ tkeys_sorted, const uint32_t* __restrict__ point_list,
                    const float4* __restrict__ rec_a, uint64_t* __restrict__ keys) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= L) return;
  keys[i] = ((uint64_t)tkeys_sorted[i] << 32) | __float_as_uint(rec_a[point_list[i]].z);
}

cudaError_t launch_emit_instances(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g,
                                  const int* radii, int tiles_x, int tiles_y, uint32_t* tkeys, uint32_t* vals) {
  count_launch(); emit_instances_kernel<<<num_pre_blocks(P), DUP_THREADS, 0, st>>>(P, order, g.rec_a, radii, g.tiles_touched,
                                                                  g.chunk_offsets, tiles_x, tiles_y, tkeys, vals);
  return cudaGetLastError();
}

cudaError_t launch_tile_digit_hist(cudaStream_t st, const uint32_t* tkeys, long long n, int n_passes,
                                   uint32_t* hist) {
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)min((long long)148 * 8, (n + 255) / 256);
  count_launch(); tile_digit_hist_kernel<<<blocks, 256, 0, st>>>(tkeys, n, n_passes, hist);
  count_launch(); tile_digit_offsets_kernel<<<1, 64, 0, st>>>(hist, n_passes);
  return cudaGetLastError();
}

cudaError_t launch_tile_ranges(cudaStream_t st, long long L, const uint32_t* tkeys_sorted, uint2* ranges,
                               size_t n_tiles) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, sizeof(uint2) * n_tiles, st);
  if (e != cudaSuccess || L <= 0) return e;
  count_launch(); tile_ranges_kernel<<<(unsigned)((L + 255) / 256), 256, 0, st>>>(L, tkeys_sorted, ranges);
  return cudaGetLastError();
}

cudaError_t launch_rebuild_keys(cudaStream_t st, long long L, const uint32_t* tkeys_sorted,
                                const uint32_t* point_list, const float4* rec_a, uint64_t* keys) {
  if (L <= 0) return cudaSuccess;
  count_launch(); rebuild_keys_kernel<<<(unsigned)((L + 255) / 256), 256, 0, st>>>(L, tkeys_sorted, point_list, rec_a, keys);
  return cudaGetLastError();
}

}  // namespace gsr
