// tile_scatter.cu -- instance binning WITHOUT an instance-level sort (replaces K4 + K5 + K6 of SURVEY 2.2:
// duplicateWithKeys, cub::DeviceRadixSort::SortPairs, identifyTileRanges; rasterizer_impl.cu:70-138,303-321)
// for images of up to TS_MAX_TILES tiles; larger images use the radix path of binning.cu.
//
// The Gaussians are already in (depth, id) order, so the reference's sorted instance list is: for every tile, the
// Gaussians whose rectangle contains the tile, in that order.  With the depth-ordered instance sequence cut into nb
// contiguous chunks of equal instance count, the final position of an instance of tile t in chunk b is
//     start[t] + sum_{b' < b} count[b'][t] + (number of earlier instances of t inside chunk b)
// which needs an nb x tiles count matrix instead of key/value passes over the R instances:
//   tile_count_kernel    each CTA walks its chunk and histograms the tile ids in shared memory   (reads 8 B/Gaussian)
//   tile_offsets_kernel  column scans of the matrix + tile totals                                (a few MB, L2)
//   tile_scatter_kernel  each CTA scans the totals into tile starts (CTA 0 also writes the ranges), walks its chunk
//                        again, in order, ranks the instances of every piece stably per tile (match.any + per-warp
//                        tile counters, like one radix-sort CTA step but over the whole tile id) and writes the
//                        Gaussian ids straight to their final slots                               (4 B/instance)
// Instance keys are never written or read; traffic is ~4 B per instance instead of ~44 B.
//
// Two-level binning (COARSE = true; the default path for images of up to TS_MAX_TILES bins = 65 536 tiles): the same
// three kernels run on BINS of 4x4 tiles -- a Gaussian of the headline scene covers 4.9 tiles but only 1.7 bins, and
// with 80 instead of 1200 destinations the per-piece bookkeeping and the rank's vote count shrink -- and write, per
// bin-level instance, the Gaussian id and a 16-bit mask of the bin's tiles its rectangle covers.  tile_refine.cu then
// splits every bin list into its 16 tile lists with warp ballots (one vote ranks ~6 instances of one tile).
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int TS_THREADS = 256;
constexpr int TS_WARPS = TS_THREADS / 32;
constexpr int TS_GPT = 4;                          // Gaussians per thread in a staged group
constexpr int TS_GROUP = TS_THREADS * TS_GPT;      // 1024 Gaussians = 4 entries of chunk_offsets
constexpr int TS_ITEMS = TS_ITEMS_PER_THREAD;      // instances per thread per piece
constexpr int TS_PIECE = TS_THREADS * TS_ITEMS;    // 2048 instances ranked at a time
#ifndef TS_MATCH_BATCH
#define TS_MATCH_BATCH 4
#endif
static_assert(TS_GROUP % 256 == 0, "a staged window is whole chunk_offsets chunks (256 Gaussians each)");

// CTA b owns the instances [k_lo, k_hi) of the depth-ordered instance sequence (an even split of `total`, so one
// chunk of screen-filling Gaussians cannot unbalance the grid); c0 = the 256-Gaussian chunk that contains k_lo.
GSR_DEVINL void ts_block_range(const uint32_t* __restrict__ chunk_offsets, int n_chunks, unsigned long long total,
                               int b, int nb, unsigned long long& k_lo, unsigned long long& k_hi, int& c0) {
  k_lo = total * (unsigned long long)b / nb;
  k_hi = total * (unsigned long long)(b + 1) / nb;
  // last chunk whose first instance index is <= k_lo, by a block-wide 256-ary search (two rounds of one load each
  // for 1 M Gaussians, where a per-thread binary search is a chain of 12 dependent loads from L2)
  int lo = 0, n = n_chunks;  // the answer lies in [lo, lo + n); chunk_offsets[lo] <= k_lo
  while (n > 1) {
    const int step = (n + TS_THREADS - 1) / TS_THREADS;
    const int idx = lo + (int)threadIdx.x * step;
    const int cnt = __syncthreads_count(idx < lo + n && (unsigned long long)chunk_offsets[idx] <= k_lo);
    const int nlo = lo + (cnt > 0 ? cnt - 1 : 0) * step;
    n = min(step, lo + n - nlo);
    lo = nlo;
  }
  c0 = lo;
}

// Last chunk >= c whose first instance index is <= k (the chunk that contains instance k). Block-wide: every thread
// tests one chunk per round, so the usual 0..4 chunk step costs one coalesced load and one barrier.
GSR_DEVINL int ts_advance_chunk(const uint32_t* __restrict__ chunk_offsets, int n_chunks, int c,
                                unsigned long long k) {
  for (;;) {
    const int idx = c + 1 + (int)threadIdx.x;
    const int cnt = __syncthreads_count(idx < n_chunks && (unsigned long long)chunk_offsets[idx] <= k);
    c += cnt;
    if (cnt < TS_THREADS) return c;
  }
}

struct TsGroup {          // one group of depth-consecutive Gaussians staged in shared memory
  uint32_t prefix[TS_GROUP + 1];  // exclusive instance prefix inside the group
  uint2 rect[TS_GROUP];
  uint32_t id[TS_GROUP];
  uint32_t warp_sum[TS_WARPS];
};

GSR_DEVINL uint32_t ts_rect_count(uint2 r) { return ((r.x >> 16) - (r.x & 0xffffu)) * ((r.y >> 16) - (r.y & 0xffffu)); }

// Loads the window of TS_GROUP Gaussians that starts at chunk c (rectangles in depth order) and builds its instance
// prefix. Returns the window's instance count.
template <bool WITH_IDS, bool COARSE>
GSR_DEVINL uint32_t ts_load_group(TsGroup& sg, int c, int P, const uint2* __restrict__ rect_sorted,
                                  const uint32_t* __restrict__ order) {
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int s0 = c * 256 + tid * TS_GPT;
  uint2 r[TS_GPT];
  uint32_t id[TS_GPT];
  if (s0 + TS_GPT <= P) {
    const uint4 a = *reinterpret_cast<const uint4*>(rect_sorted + s0);
    const uint4 b = *reinterpret_cast<const uint4*>(rect_sorted + s0 + 2);
    r[0] = make_uint2(a.x, a.y); r[1] = make_uint2(a.z, a.w); r[2] = make_uint2(b.x, b.y); r[3] = make_uint2(b.z, b.w);
    if (WITH_IDS) {
      const uint4 o = *reinterpret_cast<const uint4*>(order + s0);
      id[0] = o.x; id[1] = o.y; id[2] = o.z; id[3] = o.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < TS_GPT; ++j) {
      r[j] = make_uint2(0u, 0u); id[j] = 0u;
      if (s0 + j < P) { r[j] = rect_sorted[s0 + j]; if (WITH_IDS) id[j] = order[s0 + j]; }
    }
  }
  uint32_t n[TS_GPT], mine = 0;
#pragma unroll
  for (int j = 0; j < TS_GPT; ++j) { n[j] = ts_rect_count(COARSE ? ts_coarse_rect(r[j]) : r[j]); mine += n[j]; }
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();  // previous group fully consumed
  if (lane == 31) sg.warp_sum[w] = inc;
#pragma unroll
  for (int j = 0; j < TS_GPT; ++j) {
    sg.rect[tid * TS_GPT + j] = r[j];
    if (WITH_IDS) sg.id[tid * TS_GPT + j] = id[j];
  }
  __syncthreads();
  uint32_t run = inc - mine;
#pragma unroll
  for (int ww = 0; ww < TS_WARPS; ++ww) run += (ww < w) ? sg.warp_sum[ww] : 0u;
#pragma unroll
  for (int j = 0; j < TS_GPT; ++j) { sg.prefix[tid * TS_GPT + j] = run; run += n[j]; }
  if (tid == TS_THREADS - 1) sg.prefix[TS_GROUP] = run;
  __syncthreads();
  return sg.prefix[TS_GROUP];
}

// Walks a run of consecutive instances of the staged group: one binary search for the first instance, then
// x/y/slot increments (row-major tile order inside a rectangle like rasterizer_impl.cu:98-109).
template <bool COARSE>
struct TsWalker {
  uint32_t slot, x, y, x0, x1, y1;
  uint32_t fx0, fx1, fy0, fy1;  // COARSE: the tile rectangle (x, y, ... then walk its bins)
  GSR_DEVINL void load_rect(const TsGroup& sg) {
    uint2 r = sg.rect[slot];
    if (COARSE) {
      fx0 = r.x & 0xffffu; fx1 = r.x >> 16; fy0 = r.y & 0xffffu; fy1 = r.y >> 16;
      r = ts_coarse_rect(r);
    }
    x0 = r.x & 0xffffu; x1 = r.x >> 16; y = r.y & 0xffffu; y1 = r.y >> 16; x = x0;
  }
  GSR_DEVINL void seek(const TsGroup& sg, uint32_t k) {  // k < instance count of the group
    uint32_t l = 0;  // last slot with prefix[slot] <= k; prefix[0] = 0
#pragma unroll
    for (int step = TS_GROUP / 2; step > 0; step >>= 1)
      if (sg.prefix[l + step] <= k) l += step;
    slot = l;
    load_rect(sg);
    const uint32_t j = k - sg.prefix[l], wdt = x1 - x0, row = j / wdt;
    y += row; x += j - row * wdt;
  }
  GSR_DEVINL uint32_t tile(int tiles_x) const { return y * (uint32_t)tiles_x + x; }
  // COARSE: which of the 16 tiles of bin (x, y) the tile rectangle covers, bit 4 * dy + dx
  GSR_DEVINL uint32_t mask() const {
    const uint32_t bx = x << TS2_CSHIFT, by = y << TS2_CSHIFT;
    const uint32_t lx0 = max(fx0, bx) - bx, lx1 = min(fx1, bx + TS2_CB) - bx;
    const uint32_t ly0 = max(fy0, by) - by, ly1 = min(fy1, by + TS2_CB) - by;
    const uint32_t row = (1u << lx1) - (1u << lx0);
    const uint32_t rows = ((1u << (4u * ly1)) - (1u << (4u * ly0))) & 0x1111u;
    return row * rows;
  }
  GSR_DEVINL void next(const TsGroup& sg, uint32_t k_next) {  // k_next = index of the instance stepped to
    if (++x == x1) {
      x = x0;
      if (++y == y1) {  // rectangle exhausted: next Gaussian that touches a tile
        ++slot;
        load_rect(sg);  // slot < TS_GROUP: instance k_next exists
        if (x1 == x0 || y1 == y) seek(sg, k_next);  // culled Gaussians in between: search instead of stepping
      }
    }
  }
};

template <bool COARSE>
__global__ void __launch_bounds__(TS_THREADS)
tile_count_kernel(int P, const uint2* __restrict__ rect_sorted, const uint32_t* __restrict__ chunk_offsets,
                  const unsigned long long* __restrict__ total_dev, int tiles_x, int n_tiles,
                  uint32_t* __restrict__ counts /*[gridDim.x][n_tiles]*/) {
  pdl_launch_dependents();
  extern __shared__ __align__(16) unsigned char ts_smem[];
  TsGroup& sg = *reinterpret_cast<TsGroup*>(ts_smem);
  uint32_t* hist = reinterpret_cast<uint32_t*>(ts_smem + sizeof(TsGroup));
  const int tid = threadIdx.x;
  for (int t = tid; t < n_tiles; t += TS_THREADS) hist[t] = 0u;
  pdl_wait();
  const int n_chunks = num_pre_blocks(P);
  unsigned long long k_lo, k_hi;
  int c;
  ts_block_range(chunk_offsets, n_chunks, *total_dev, blockIdx.x, gridDim.x, k_lo, k_hi, c);
  for (unsigned long long kk = k_lo; kk < k_hi;) {
    c = ts_advance_chunk(chunk_offsets, n_chunks, c, kk);
    const unsigned long long off = chunk_offsets[c];
    const uint32_t n = ts_load_group<false, COARSE>(sg, c, P, rect_sorted, nullptr);
    const uint32_t b = (uint32_t)(kk - off);
    const uint32_t e = (uint32_t)min((unsigned long long)n, k_hi - off);
    if (e <= b) break;  // cannot happen (chunk c holds instance kk); never spin on inconsistent inputs
    kk = off + e;
    {  // thread t walks the contiguous run [k, k_end) of the window's instances [b, e)
      const uint32_t per = (e - b + TS_THREADS - 1) / TS_THREADS;
      uint32_t k = b + tid * per;
      const uint32_t k_end = min(e, k + per);
      if (k < k_end) {
        TsWalker<COARSE> wk;
        wk.seek(sg, k);
        for (;;) {
          atomicAdd(&hist[wk.tile(tiles_x)], 1u);
          if (++k == k_end) break;
          wk.next(sg, k);
        }
      }
    }
  }
  __syncthreads();
  for (int t = tid; t < n_tiles; t += TS_THREADS) counts[(size_t)blockIdx.x * n_tiles + t] = hist[t];
}

// One CTA per strip of 32 tiles: counts[b][t] -> exclusive prefix over b (in place), totals[t] = column sum.
// Warp w owns a contiguous band of rows; lane = column, so every row access is one 128-byte line.
constexpr int TS_OFF_WARPS = 32;
__global__ void __launch_bounds__(TS_OFF_WARPS * 32)
tile_offsets_kernel(int nb, int n_tiles, uint32_t* __restrict__ counts, uint32_t* __restrict__ totals) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t part[TS_OFF_WARPS][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int t = blockIdx.x * 32 + lane;
  const int rows_per_warp = (nb + TS_OFF_WARPS - 1) / TS_OFF_WARPS;
  const int r0 = w * rows_per_warp, r1 = min(nb, r0 + rows_per_warp);
  uint32_t sum = 0u;
  if (t < n_tiles) {
#pragma unroll 8
    for (int b = r0; b < r1; ++b) sum += counts[(size_t)b * n_tiles + t];
  }
  part[w][lane] = sum;
  __syncthreads();
  uint32_t run = 0u, all = 0u;
#pragma unroll
  for (int ww = 0; ww < TS_OFF_WARPS; ++ww) {
    const uint32_t p = part[ww][lane];
    run += ww < w ? p : 0u;
    all += p;
  }
  if (t < n_tiles) {
    for (int b0 = r0; b0 < r1; b0 += 8) {  // loads of a batch issued together, then the stores
      uint32_t c[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) c[j] = b0 + j < r1 ? counts[(size_t)(b0 + j) * n_tiles + t] : 0u;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (b0 + j < r1) counts[(size_t)(b0 + j) * n_tiles + t] = run;
        run += c[j];
      }
    }
    if (w == 0) totals[t] = all;
  }
}

constexpr int TS_STAGE_WORDS = TS_PIECE + TS_PIECE / 32;
static_assert(sizeof(TsGroup) + TS_STAGE_WORDS * 4 <= TS_GROUP_SMEM, "ts_blocks() budgets TS_GROUP_SMEM bytes");
static_assert(TS_GROUP <= 1024 && TS_MAX_TILES <= (1 << 12), "stage word = tile << 10 | slot; 13-bit peer masks");

// a[0..n) (shared memory) -> its exclusive prefix, in place; returns the total. Block-wide (TS_THREADS threads).
GSR_DEVINL uint32_t ts_block_scan(uint32_t* a, int n, uint32_t* warp_sum) {
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int per = (n + TS_THREADS - 1) / TS_THREADS;
  const int t0 = min(n, tid * per), t1 = min(n, t0 + per);
  uint32_t mine = 0u;
  for (int t = t0; t < t1; ++t) mine += a[t];
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  __syncthreads();  // warp_sum free
  if (lane == 31) warp_sum[w] = inc;
  __syncthreads();
  uint32_t run = inc - mine, all = 0u;
#pragma unroll
  for (int ww = 0; ww < TS_WARPS; ++ww) { run += (ww < w) ? warp_sum[ww] : 0u; all += warp_sum[ww]; }
  for (int t = t0; t < t1; ++t) {
    const uint32_t c = a[t];
    a[t] = run;
    run += c;
  }
  __syncthreads();
  return all;
}

// COARSE = false: destinations are tiles, `list` = point_list (Gaussian ids), CTA 0 writes starts and ranges.
// COARSE = true : destinations are bins (tiles_x = bins per row, n_tiles = bins), `list` = the bin lists' ids and
//                 `masks` their tile masks; CTA 0 writes the bin starts and the segment table of level 2
//                 (seg_prefix[bin] = first segment of the bin, seg_bin[segment] = its bin).
template <bool COARSE>
__global__ void __launch_bounds__(TS_THREADS, TS_MAX_OCC)
tile_scatter_kernel(int P, const uint2* __restrict__ rect_sorted, const uint32_t* __restrict__ order,
                    const uint32_t* __restrict__ chunk_offsets, const unsigned long long* __restrict__ total_dev,
                    int tiles_x, int n_tiles, const uint32_t* __restrict__ counts /*exclusive over chunks*/,
                    const uint32_t* __restrict__ totals, long long cap, uint32_t* __restrict__ start_out,
                    uint2* __restrict__ ranges, uint32_t* __restrict__ list, uint16_t* __restrict__ masks,
                    uint32_t* __restrict__ seg_prefix, uint32_t* __restrict__ seg_bin, uint32_t max_segs) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) unsigned char ts_smem[];
  TsGroup& sg = *reinterpret_cast<TsGroup*>(ts_smem);
  uint32_t* cur = reinterpret_cast<uint32_t*>(ts_smem + sizeof(TsGroup));   // next free slot of every tile
  uint32_t* pbase = cur + n_tiles;                                          // the same, at the start of the piece
  uint32_t* stage = pbase + n_tiles;                                        // piece: tile << 10 | slot, padded
  uint16_t* whist = reinterpret_cast<uint16_t*>(stage + TS_STAGE_WORDS);    // [TS_WARPS][n_tiles]
  uint16_t* mstage = whist + (size_t)TS_WARPS * n_tiles + 2;                // COARSE: tile mask of every staged instance
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;

  // ---- tile starts = exclusive scan of the tile totals (every CTA; CTA 0 publishes starts and ranges) ----
  for (int t = tid; t < n_tiles; t += TS_THREADS) { const uint32_t c = totals[t]; pbase[t] = c; cur[t] = c; }
  __syncthreads();
  ts_block_scan(cur, n_tiles, sg.warp_sum);
  if (blockIdx.x == 0) {
    for (int t = tid; t < n_tiles; t += TS_THREADS) {
      const uint32_t st = cur[t], n = pbase[t];
      start_out[t] = st;
      if (!COARSE) {
        // clipped to the list capacity: an undersized list (GSR_ECAPACITY) must not send the renderer out of
        // bounds; (0,0) for empty tiles like the reference's memset (rasterizer_impl.cu:313)
        const uint32_t lo = (uint32_t)min((long long)st, cap), hi = (uint32_t)min((long long)st + n, cap);
        ranges[t] = hi > lo ? make_uint2(lo, hi) : make_uint2(0u, 0u);
      }
    }
    if (COARSE) {
      uint32_t* nseg = reinterpret_cast<uint32_t*>(whist);  // [n_tiles] (whist holds 4 words per tile)
      for (int t = tid; t < n_tiles; t += TS_THREADS) nseg[t] = (pbase[t] + TS2_SEG - 1) / TS2_SEG;
      __syncthreads();
      const uint32_t all = ts_block_scan(nseg, n_tiles, sg.warp_sum);
      for (int t = tid; t < n_tiles; t += TS_THREADS) {
        const uint32_t first = nseg[t], n = (pbase[t] + TS2_SEG - 1) / TS2_SEG;
        seg_prefix[t] = first;
        for (uint32_t q = 0; q < n; ++q)
          if (first + q < max_segs) seg_bin[first + q] = (uint32_t)t;
      }
      if (tid == 0) seg_prefix[n_tiles] = min(all, max_segs);
      __syncthreads();
    }
  }
  for (int t = tid; t < n_tiles; t += TS_THREADS) cur[t] += counts[(size_t)blockIdx.x * n_tiles + t];

  const int n_chunks = num_pre_blocks(P);
  unsigned long long k_lo, k_hi;
  int c;
  ts_block_range(chunk_offsets, n_chunks, *total_dev, blockIdx.x, gridDim.x, k_lo, k_hi, c);
  uint16_t* my_hist = whist + (size_t)w * n_tiles;
  const int n_words = TS_WARPS * n_tiles / 2 + 1;

  // Pieces of up to TS_PIECE consecutive instances; the staged window of Gaussians slides with them (it starts at
  // the chunk that contains the piece's first instance), so pieces are full wherever TS_GROUP Gaussians have
  // TS_PIECE instances.
  int loaded = -1;
  uint32_t n_all = 0;
  for (unsigned long long kk = k_lo; kk < k_hi;) {
    c = ts_advance_chunk(chunk_offsets, n_chunks, c, kk);
    if (c != loaded) { n_all = ts_load_group<true, COARSE>(sg, c, P, rect_sorted, order); loaded = c; }
    const unsigned long long off = chunk_offsets[c];
    const uint32_t piece = (uint32_t)(kk - off);
    const uint32_t n = (uint32_t)min((unsigned long long)n_all, k_hi - off);
    {
      if (n <= piece) break;  // cannot happen (chunk c holds instance kk); never spin on inconsistent inputs
      const uint32_t piece_n = min((uint32_t)TS_PIECE, n - piece);
      kk += piece_n;
      for (int i = tid; i < n_words; i += TS_THREADS) reinterpret_cast<uint32_t*>(whist)[i] = 0u;
      // warp w ranks the contiguous instances [wbase, wbase + 32 * rounds) of the piece
      const int rounds = (int)((piece_n + TS_THREADS - 1) / TS_THREADS);
      const uint32_t wbase = (uint32_t)(w * rounds * 32);

      // ---- (tile, slot) of every instance: thread t walks the run [t * rounds, (t+1) * rounds) of the piece and
      //      stages it in instance order (one pad word per 32 keeps both access patterns conflict-free) ----
      {
        uint32_t local = (uint32_t)(tid * rounds);
        const uint32_t local_end = min(piece_n, local + (uint32_t)rounds);
        if (local < local_end) {
          TsWalker<COARSE> wk;
          wk.seek(sg, piece + local);
          for (;;) {
            stage[local + (local >> 5)] = wk.tile(tiles_x) << 10 | wk.slot;
            if (COARSE) mstage[local] = (uint16_t)wk.mask();
            if (++local == local_end) break;
            wk.next(sg, piece + local);
          }
        }
      }
      __syncthreads();
      uint32_t tile[TS_ITEMS], sr[TS_ITEMS];  // sr = slot << 16 | rank
#pragma unroll
      for (int i = 0; i < TS_ITEMS; ++i) {
        const uint32_t local = wbase + i * 32 + lane;
        tile[i] = 0xFFFFFFFFu; sr[i] = 0u;
        if (i < rounds && local < piece_n) {
          const uint32_t v = stage[local + (local >> 5)];
          tile[i] = v >> 10;
          sr[i] = (v & 1023u) << 16;
        }
      }

      // ---- stable rank per tile inside the warp's instances (match.any issued TS_MATCH_BATCH rounds ahead: the
      //      counter updates of one round hide the match latency of the next) ----
#pragma unroll
      for (int i0 = 0; i0 < TS_ITEMS; i0 += TS_MATCH_BATCH) {
        uint32_t peers[TS_MATCH_BATCH];
#pragma unroll
        for (int j = 0; j < TS_MATCH_BATCH; ++j) {
          const int i = i0 + j;
          peers[j] = 0u;
          // 13 bits: tile ids < TS_MAX_TILES = 2^12, bit 12 of the invalid marker 0xFFFFFFFF separates idle lanes
          if (i < rounds && wbase + i * 32 < piece_n)  // warp-uniform branch
            peers[j] = (COARSE && n_tiles <= 128) ? warp_peers<8>(tile[i]) : warp_peers<13>(tile[i]);
        }
#pragma unroll
        for (int j = 0; j < TS_MATCH_BATCH; ++j) {
          const int i = i0 + j;
          if (i < rounds && wbase + i * 32 < piece_n) {  // warp-uniform
            const uint32_t below = __popc(peers[j] & ((1u << lane) - 1u));
            uint32_t prev = 0;
            if (tile[i] != 0xFFFFFFFFu && below == 0) {
              prev = my_hist[tile[i]];
              my_hist[tile[i]] = (uint16_t)(prev + __popc(peers[j]));
            }
            prev = __shfl_sync(0xffffffffu, prev, __ffs(peers[j]) - 1);
            sr[i] |= prev + below;
            __syncwarp();
          }
        }
      }
      __syncthreads();

      // ---- per tile: exclusive prefix over the warps, advance the tile's cursor ----
      for (int t = tid; t < n_tiles; t += TS_THREADS) {
        uint32_t run = 0;
#pragma unroll
        for (int ww = 0; ww < TS_WARPS; ++ww) {
          const uint32_t c = whist[(size_t)ww * n_tiles + t];
          whist[(size_t)ww * n_tiles + t] = (uint16_t)run;
          run += c;
        }
        const uint32_t c0 = cur[t];
        pbase[t] = c0;
        cur[t] = c0 + run;
      }
      __syncthreads();

      // ---- final slots ----
#pragma unroll
      for (int i = 0; i < TS_ITEMS; ++i) {
        if (tile[i] != 0xFFFFFFFFu) {
          const long long pos = (long long)pbase[tile[i]] + my_hist[tile[i]] + (sr[i] & 0xffffu);
          if (pos < cap) {
            list[pos] = sg.id[sr[i] >> 16];
            if (COARSE) masks[pos] = mstage[wbase + i * 32 + lane];
          }
        }
      }
      __syncthreads();
    }
  }
}

// Test hook: the sorted tile id of every list position, from the tile starts (last tile with start <= i).
__global__ void __launch_bounds__(256)
rebuild_keys_from_starts_kernel(long long L, int n_tiles, const uint32_t* __restrict__ start,
                                const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
                                uint64_t* __restrict__ keys) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= L) return;
  int l = 0, h = n_tiles - 1;
  while (l < h) {
    const int m = (l + h + 1) >> 1;
    if ((long long)start[m] <= i) l = m; else h = m - 1;
  }
  keys[i] = ((uint64_t)(uint32_t)l << 32) | __float_as_uint(rec[3 * (size_t)point_list[i]].z);
}

static size_t ts_scatter_smem(int n_tiles, bool coarse) {
  return sizeof(TsGroup) + TS_STAGE_WORDS * 4 + (size_t)n_tiles * TS_TILE_SMEM + 8 + (coarse ? TS_PIECE * 2 : 0);
}
static size_t ts_count_smem(int n_tiles) { return sizeof(TsGroup) + (size_t)n_tiles * 4; }

static cudaError_t ts_set_attributes() {
  static bool attr = false;
  if (attr) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(tile_count_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)ts_count_smem(TS_MAX_TILES));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(tile_count_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)ts_count_smem(TS_MAX_TILES));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(tile_scatter_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)ts_scatter_smem(TS_MAX_TILES, false));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(tile_scatter_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)ts_scatter_smem(TS_MAX_TILES, true));
  if (e != cudaSuccess) return e;
  attr = true;
  return cudaSuccess;
}

cudaError_t launch_tile_count(cudaStream_t st, int P, const GeomViews& g, int tiles_x, int n_tiles) {
  cudaError_t e = ts_set_attributes();
  if (e != cudaSuccess) return e;
  count_launch();
  launch_k(tile_count_kernel<false>, ts_blocks(n_tiles), TS_THREADS, ts_count_smem(n_tiles), st,
      P, g.rect_sorted, g.chunk_offsets, g.total, tiles_x, n_tiles, g.ts_counts);
  return cudaGetLastError();
}

cudaError_t launch_tile_offsets_scatter(cudaStream_t st, int P, const GeomViews& g, int tiles_x, int n_tiles,
                                        long long cap, uint2* ranges, uint32_t* point_list) {
  const int nb = ts_blocks(n_tiles);
  count_launch();
  launch_k(tile_offsets_kernel, (n_tiles + 31) / 32, TS_OFF_WARPS * 32, 0, st, nb, n_tiles, g.ts_counts, g.ts_totals);
  count_launch();
  launch_k(tile_scatter_kernel<false>, nb, TS_THREADS, ts_scatter_smem(n_tiles, false), st,
      P, g.rect_sorted, g.order_a, g.chunk_offsets, g.total, tiles_x, n_tiles, g.ts_counts, g.ts_totals, cap,
      g.ts_start, ranges, point_list, nullptr, nullptr, nullptr, 0u);
  return cudaGetLastError();
}

// Level 1 of the two-level binning: the same kernels on bins of TS2_CB x TS2_CB tiles.
cudaError_t launch_bin_count(cudaStream_t st, int P, const GeomViews& g, int bins_x, int n_bins) {
  cudaError_t e = ts_set_attributes();
  if (e != cudaSuccess) return e;
  count_launch();
  launch_k(tile_count_kernel<true>, ts_blocks_coarse(n_bins), TS_THREADS, ts_count_smem(n_bins), st,
      P, g.rect_sorted, g.chunk_offsets_c, g.total_c, bins_x, n_bins, g.ts_counts);
  return cudaGetLastError();
}

cudaError_t launch_bin_offsets_scatter(cudaStream_t st, int P, const GeomViews& g, int bins_x, int n_bins,
                                       long long cap, const RefineViews& rv) {
  const int nb = ts_blocks_coarse(n_bins);
  count_launch();
  launch_k(tile_offsets_kernel, (n_bins + 31) / 32, TS_OFF_WARPS * 32, 0, st, nb, n_bins, g.ts_counts, g.ts_ctotals);
  count_launch();
  launch_k(tile_scatter_kernel<true>, nb, TS_THREADS, ts_scatter_smem(n_bins, true), st,
      P, g.rect_sorted, g.order_a, g.chunk_offsets_c, g.total_c, bins_x, n_bins, g.ts_counts, g.ts_ctotals, cap,
      g.ts_cstart, nullptr, rv.ids, rv.masks, g.ts_seg_prefix, rv.seg_bin, (uint32_t)rv.max_segs);
  return cudaGetLastError();
}

cudaError_t launch_rebuild_keys_from_starts(cudaStream_t st, long long L, int n_tiles, const GeomViews& g,
                                            const uint32_t* point_list, uint64_t* keys) {
  if (L <= 0) return cudaSuccess;
  count_launch();
  launch_k(rebuild_keys_from_starts_kernel, (unsigned)((L + 255) / 256), 256, 0, st, L, n_tiles, g.ts_start, point_list,
                                                                               g.rec, keys);
  return cudaGetLastError();
}

}  // namespace gsr
