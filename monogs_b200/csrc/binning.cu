// binning.cu -- instance key emission (K4 of SURVEY 2.2).
//
// Replaces duplicateWithKeys (rasterizer_impl.cu:70-111): for every visible
// Gaussian and every tile of its bounding rect, in row-major tile order,
//   key = (tile_id << 32) | float_bits(view depth),  value = Gaussian index.
// The write offset of a Gaussian is the exclusive prefix sum of tiles_touched
// (block offsets come from finish_scan_kernel, the in-block part is scanned
// here), so the unsorted order equals the reference's and the stable sort that
// follows yields a bit-identical point_list.
// Gaussians covering many tiles are emitted cooperatively by their whole warp
// instead of serially by one thread (the reference's load imbalance).
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int DUP_THREADS = 256;   // must equal PRE_THREADS: block_sums are per 256 Gaussians
constexpr uint32_t COOP_THRESHOLD = 32;

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

__global__ void __launch_bounds__(DUP_THREADS)
emit_keys_kernel(int P, const float4* __restrict__ rec_a, const int* __restrict__ radii,
                 const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ block_offsets,
                 int tiles_x, int tiles_y, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  __shared__ uint32_t s_warp[DUP_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int idx = blockIdx.x * DUP_THREADS + tid;
  const uint32_t n = idx < P ? tiles_touched[idx] : 0u;

  // exclusive scan of n inside the block
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
  uint32_t off = block_offsets[blockIdx.x] + woff + inc - n;

  uint2 rmin = make_uint2(0, 0), rmax = make_uint2(0, 0);
  uint32_t dbits = 0;
  if (n > 0) {
    const float4 a = rec_a[idx];
    rect_from_rec(a.x, a.y, radii[idx], tiles_x, tiles_y, rmin, rmax);
    dbits = __float_as_uint(a.z);
  }

  // small footprints: one thread writes its own instances
  if (n > 0 && n <= COOP_THRESHOLD) {
    uint32_t o = off;
    for (unsigned y = rmin.y; y < rmax.y; ++y)
      for (unsigned x = rmin.x; x < rmax.x; ++x) {
        keys[o] = ((uint64_t)(y * tiles_x + x) << 32) | dbits;
        vals[o] = (uint32_t)idx;
        ++o;
      }
  }
  // large footprints: the warp emits them together, one Gaussian at a time
  uint32_t big = __ballot_sync(0xffffffffu, n > COOP_THRESHOLD);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t bn = __shfl_sync(0xffffffffu, n, src);
    const uint32_t boff = __shfl_sync(0xffffffffu, off, src);
    const uint32_t bx0 = __shfl_sync(0xffffffffu, rmin.x, src);
    const uint32_t by0 = __shfl_sync(0xffffffffu, rmin.y, src);
    const uint32_t bx1 = __shfl_sync(0xffffffffu, rmax.x, src);
    const uint32_t bd = __shfl_sync(0xffffffffu, dbits, src);
    const uint32_t bidx = (uint32_t)(blockIdx.x * DUP_THREADS + (w << 5) + src);
    const uint32_t bw = bx1 - bx0;
    for (uint32_t k = lane; k < bn; k += 32) {
      const uint32_t y = by0 + k / bw, x = bx0 + k % bw;
      keys[boff + k] = ((uint64_t)(y * tiles_x + x) << 32) | bd;
      vals[boff + k] = bidx;
    }
  }
}

cudaError_t launch_emit_keys(cudaStream_t st, int P, const GeomViews& g, const int* radii, int tiles_x,
                             int tiles_y, uint64_t* keys, uint32_t* vals) {
  emit_keys_kernel<<<num_pre_blocks(P), DUP_THREADS, 0, st>>>(P, g.rec_a, radii, g.tiles_touched, g.block_sums,
                                                             tiles_x, tiles_y, keys, vals);
  return cudaGetLastError();
}

}  // namespace gsr
