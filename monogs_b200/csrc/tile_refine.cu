// tile_refine.cu -- level 2 of the two-level instance binning (level 1: tile_scatter.cu with COARSE = true).
//
// Level 1 leaves, for every bin of 4x4 tiles, the list of the Gaussians whose tile rectangle meets the bin, in
// (depth, id) order, each with a 16-bit mask of the bin's tiles it covers.  The reference's sorted instance list
// (duplicateWithKeys + SortPairs + identifyTileRanges, rasterizer_impl.cu:70-138,303-321) is, tile by tile, the
// sub-list of its bin's list whose mask has the tile's bit -- in the same order.  Every bin list is cut into segments
// of TS2_SEG entries; per segment one CTA
//   refine_count_kernel    counts the entries of each of the 16 tiles                       (2 B/entry)
//   refine_scan_kernel     per bin: exclusive prefix of those counts over the bin's segments, tile totals; the last
//                          CTA scans the totals of ALL tiles into list starts and writes the ranges
//   refine_scatter_kernel  ranks the entries of every tile with one warp ballot per (32 entries, tile) -- a vote ranks
//                          all the entries of a tile at once (~6 at the headline scene), where the single-level
//                          scatter spends a 13-vote match per instance -- and writes the ids to their final slots.
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int RF_THREADS = 256;
constexpr int RF_WARPS = RF_THREADS / 32;
constexpr int RF_ITEMS = TS2_SEG / RF_THREADS;  // entries per thread, warp-striped (a warp owns 32 * RF_ITEMS consecutive entries)
static_assert(TS2_SEG % RF_THREADS == 0 && TS2_BIN_TILES == 16, "16 tiles per bin: lanes 0..15 own one tile each");

struct RfSeg { uint32_t bin, first, n; };  // entries [first, first + n) of the bin lists

GSR_DEVINL RfSeg rf_segment(uint32_t seg, const uint32_t* __restrict__ seg_bin, const uint32_t* __restrict__ seg_prefix,
                            const uint32_t* __restrict__ cstart, const uint32_t* __restrict__ ctotals, long long cap) {
  RfSeg s;
  s.bin = seg_bin[seg];
  const uint32_t q = seg - seg_prefix[s.bin], st = cstart[s.bin], tot = ctotals[s.bin];
  const long long first = (long long)st + (long long)q * TS2_SEG;
  long long end = min((long long)st + tot, first + TS2_SEG);
  end = min(end, cap);  // an undersized list (GSR_ECAPACITY): entries beyond it were never written
  s.first = (uint32_t)min(first, cap);
  s.n = end > first ? (uint32_t)(end - first) : 0u;
  return s;
}

// Lane t < 16 returns how many of the warp's 32 * RF_ITEMS masks have bit t.
GSR_DEVINL uint32_t rf_warp_counts(const uint32_t (&m)[RF_ITEMS], int lane) {
  uint32_t cnt = 0;
#pragma unroll
  for (int i = 0; i < RF_ITEMS; ++i) {
#pragma unroll
    for (int t = 0; t < TS2_BIN_TILES; ++t) {
      const uint32_t b = __ballot_sync(0xffffffffu, (m[i] >> t) & 1u);
      if (lane == t) cnt += __popc(b);
    }
  }
  return cnt;
}

__global__ void __launch_bounds__(RF_THREADS)
refine_count_kernel(const uint16_t* __restrict__ masks, const uint32_t* __restrict__ seg_bin,
                    const uint32_t* __restrict__ seg_prefix, const uint32_t* __restrict__ cstart,
                    const uint32_t* __restrict__ ctotals, int n_bins, uint32_t max_segs, long long cap,
                    uint32_t* __restrict__ seg_counts, uint32_t* __restrict__ warp_counts /*[segs][8 warps][16]*/) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_cnt[RF_WARPS][TS2_BIN_TILES];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t n_segs = min(seg_prefix[n_bins], max_segs);
  for (uint32_t seg = blockIdx.x; seg < n_segs; seg += gridDim.x) {
    const RfSeg sg = rf_segment(seg, seg_bin, seg_prefix, cstart, ctotals, cap);
    uint32_t m[RF_ITEMS];
#pragma unroll
    for (int i = 0; i < RF_ITEMS; ++i) {
      const uint32_t e = (uint32_t)(w * 32 * RF_ITEMS + i * 32 + lane);
      m[i] = e < sg.n ? (uint32_t)masks[sg.first + e] : 0u;
    }
    const uint32_t cnt = rf_warp_counts(m, lane);
    if (lane < TS2_BIN_TILES) {
      s_cnt[w][lane] = cnt;
      warp_counts[((size_t)seg * RF_WARPS + w) * TS2_BIN_TILES + lane] = cnt;  // the scatter kernel's warp offsets
    }
    __syncthreads();
    if (tid < TS2_BIN_TILES) {
      uint32_t t = 0;
#pragma unroll
      for (int ww = 0; ww < RF_WARPS; ++ww) t += s_cnt[ww][tid];
      seg_counts[(size_t)seg * TS2_BIN_TILES + tid] = t;
    }
    __syncthreads();
  }
}

constexpr int RS2_THREADS = 1024;
__global__ void __launch_bounds__(RS2_THREADS)
refine_scan_kernel(int n_bins, int bins_x, int tiles_x, int tiles_y, const uint32_t* __restrict__ seg_prefix,
                   uint32_t max_segs, uint32_t* __restrict__ seg_counts, uint32_t* __restrict__ totals,
                   uint32_t* __restrict__ ticket, long long cap, uint32_t* __restrict__ start,
                   uint2* __restrict__ ranges) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_warp[33];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int bin = blockIdx.x * (RS2_THREADS / 32) + w;
  if (bin < n_bins && lane < TS2_BIN_TILES) {
    const uint32_t s0 = min(seg_prefix[bin], max_segs), s1 = min(seg_prefix[bin + 1], max_segs);
    uint32_t run = 0;
    for (uint32_t sb = s0; sb < s1; sb += 8) {  // the loads of a batch issued together, then the stores
      uint32_t v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = sb + j < s1 ? seg_counts[(size_t)(sb + j) * TS2_BIN_TILES + lane] : 0u;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (sb + j < s1) seg_counts[(size_t)(sb + j) * TS2_BIN_TILES + lane] = run;
        run += v[j];
      }
    }
    const int x = ((bin % bins_x) << TS2_CSHIFT) + (lane & (TS2_CB - 1));
    const int y = ((bin / bins_x) << TS2_CSHIFT) + (lane >> TS2_CSHIFT);
    if (x < tiles_x && y < tiles_y) totals[y * tiles_x + x] = run;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (tid == 0) *ticket = 0u;  // the render half of a forward may be enqueued again on the same plan
  // ---- last CTA: list start of every tile = exclusive scan of the totals in tile order; ranges ----
  const int n_tiles = tiles_x * tiles_y;
  const int per = (n_tiles + RS2_THREADS - 1) / RS2_THREADS;
  const int t0 = min(n_tiles, tid * per), t1 = min(n_tiles, t0 + per);
  uint32_t mine = 0;
  for (int t = t0; t < t1; ++t) mine += *reinterpret_cast<volatile uint32_t*>(&totals[t]);
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  if (w == 0) {
    const uint32_t x = s_warp[lane];
    uint32_t xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, xi, o);
      if (lane >= o) xi += v;
    }
    s_warp[lane] = xi - x;
  }
  __syncthreads();
  uint32_t run = s_warp[w] + inc - mine;
  for (int t = t0; t < t1; ++t) {
    const uint32_t n = *reinterpret_cast<volatile uint32_t*>(&totals[t]);
    start[t] = run;
    // clipped to the list capacity; (0,0) for empty tiles like the reference's memset (rasterizer_impl.cu:313)
    const uint32_t lo = (uint32_t)min((long long)run, cap), hi = (uint32_t)min((long long)run + n, cap);
    ranges[t] = hi > lo ? make_uint2(lo, hi) : make_uint2(0u, 0u);
    run += n;
  }
}

__global__ void __launch_bounds__(RF_THREADS, 4)
refine_scatter_kernel(const uint32_t* __restrict__ ids, const uint16_t* __restrict__ masks,
                      const uint32_t* __restrict__ seg_bin, const uint32_t* __restrict__ seg_prefix,
                      const uint32_t* __restrict__ cstart, const uint32_t* __restrict__ ctotals, int n_bins,
                      int bins_x, int tiles_x, int tiles_y, uint32_t max_segs,
                      const uint32_t* __restrict__ seg_counts /*exclusive over the segments of a bin*/,
                      const uint32_t* __restrict__ warp_counts, const uint32_t* __restrict__ start, long long cap,
                      uint32_t* __restrict__ point_list) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_cnt[RF_WARPS * TS2_BIN_TILES];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t cap32 = (uint32_t)min(cap, (long long)0xffffffffll);
  const uint32_t n_segs = min(seg_prefix[n_bins], max_segs);
  for (uint32_t seg = blockIdx.x; seg < n_segs; seg += gridDim.x) {
    const RfSeg sg = rf_segment(seg, seg_bin, seg_prefix, cstart, ctotals, cap);
    if (tid < RF_WARPS * TS2_BIN_TILES) s_cnt[tid] = warp_counts[(size_t)seg * RF_WARPS * TS2_BIN_TILES + tid];
    uint32_t m[RF_ITEMS], id[RF_ITEMS];
#pragma unroll
    for (int i = 0; i < RF_ITEMS; ++i) {
      const uint32_t e = (uint32_t)(w * 32 * RF_ITEMS + i * 32 + lane);
      m[i] = e < sg.n ? (uint32_t)masks[sg.first + e] : 0u;
    }
#pragma unroll
    for (int i = 0; i < RF_ITEMS; ++i) {
      const uint32_t e = (uint32_t)(w * 32 * RF_ITEMS + i * 32 + lane);
      id[i] = e < sg.n ? ids[sg.first + e] : 0u;
    }
    // lane t < 16: first list slot of tile t for this warp's entries
    uint32_t mine = 0;
    if (lane < TS2_BIN_TILES) {
      const int x = ((sg.bin % bins_x) << TS2_CSHIFT) + (lane & (TS2_CB - 1));
      const int y = ((sg.bin / bins_x) << TS2_CSHIFT) + (lane >> TS2_CSHIFT);
      if (x < tiles_x && y < tiles_y) mine = start[y * tiles_x + x] + seg_counts[(size_t)seg * TS2_BIN_TILES + lane];
    }
    __syncthreads();
    if (lane < TS2_BIN_TILES) {
#pragma unroll
      for (int ww = 0; ww < RF_WARPS; ++ww) mine += ww < w ? s_cnt[ww * TS2_BIN_TILES + lane] : 0u;
    }
    uint32_t base[TS2_BIN_TILES];  // warp-uniform: next free slot of every tile
#pragma unroll
    for (int t = 0; t < TS2_BIN_TILES; ++t) base[t] = __shfl_sync(0xffffffffu, mine, t);
#pragma unroll
    for (int i = 0; i < RF_ITEMS; ++i) {
#pragma unroll
      for (int t = 0; t < TS2_BIN_TILES; ++t) {
        const bool in = (m[i] >> t) & 1u;
        const uint32_t b = __ballot_sync(0xffffffffu, in);
        const uint32_t pos = base[t] + __popc(b & lt);
        if (in && pos < cap32) point_list[pos] = id[i];
        base[t] += __popc(b);
      }
    }
    __syncthreads();
  }
}

cudaError_t launch_refine(cudaStream_t st, const GeomViews& g, const RefineViews& rv, int width, int height,
                          long long cap, uint2* ranges, uint32_t* point_list) {
  const int tiles_x = (width + TILE_X - 1) / TILE_X, tiles_y = (height + TILE_Y - 1) / TILE_Y;
  const int bins_x = (tiles_x + TS2_CB - 1) / TS2_CB, n_bins = (int)num_bins(width, height);
  const unsigned max_segs = (unsigned)rv.max_segs;
  const unsigned grid = max_segs < 148u * 8u ? max_segs : 148u * 8u;
  count_launch();
  launch_k(refine_count_kernel, grid, RF_THREADS, 0, st, rv.masks, rv.seg_bin, g.ts_seg_prefix, g.ts_cstart,
           g.ts_ctotals, n_bins, max_segs, cap, rv.seg_counts, rv.warp_counts);
  count_launch();
  launch_k(refine_scan_kernel, (n_bins + RS2_THREADS / 32 - 1) / (RS2_THREADS / 32), RS2_THREADS, 0, st, n_bins, bins_x,
           tiles_x, tiles_y, g.ts_seg_prefix, max_segs, rv.seg_counts, g.ts_totals, g.refine_ticket, cap, g.ts_start,
           ranges);
  count_launch();
  launch_k(refine_scatter_kernel, grid, RF_THREADS, 0, st, rv.ids, rv.masks, rv.seg_bin, g.ts_seg_prefix, g.ts_cstart,
           g.ts_ctotals, n_bins, bins_x, tiles_x, tiles_y, max_segs, rv.seg_counts, rv.warp_counts, g.ts_start, cap,
           point_list);
  return cudaGetLastError();
}

}  // namespace gsr
