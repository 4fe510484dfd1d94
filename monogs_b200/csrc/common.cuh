// common.cuh -- shared definitions for the sm_100a rasterizer kernels.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>
#include "../../include/gsr_b200.h"

namespace gsr {

// number of kernels launched by this library since load (gsr_profile_kernel_launches; bench.py's gpu_launches)
unsigned long long count_launch(unsigned long long add = 1);

constexpr int TILE_X = 16;   // binning tile: must equal the reference's BLOCK_X/BLOCK_Y
constexpr int TILE_Y = 16;   // (config.h:15-17) or radii->tiles->keys->ranges would differ
constexpr int TILE_PIX = TILE_X * TILE_Y;

// ---------------------------------------------------------------------------
// Workspace carving. Every sub-array is 256-byte aligned.
// ---------------------------------------------------------------------------
__host__ __device__ inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Per-Gaussian geometry state ("GeometryState" of rasterizer_impl.h:33-48), re-laid out as ONE packed 48-byte
// record per Gaussian = 3 consecutive float4, so a renderer's gather of an instance touches two 32-byte sectors:
//   rec[3i + 0] = {mean2D.x, mean2D.y, depth(view z), tau}   tau = cull threshold ln(255 o) + margins (preprocess.cu)
//   rec[3i + 1] = {conic.x, conic.y, conic.z, opacity}
//   rec[3i + 2] = {r, g, b, 0}
struct GeomViews {
  float4* rec;               // [3P]
  uint32_t* tiles_touched;   // [P]
  uint2* rect;               // [P] tile rectangle x0 | x1 << 16, y0 | y1 << 16 (empty when culled)
  uint2* rect_sorted;        // [P] the same in (depth, id) order
  uint8_t* clamped;          // [3P] SH clamp flags (forward.cu:69-71)
  // depth-order machinery (forward only)
  uint32_t* dkeys_a;         // [P] float bits of view depth (0xFFFFFFFF for culled Gaussians)
  uint32_t* dkeys_b;         // [P]
  uint32_t* order_a;         // [P] Gaussian ids; after the 4-pass sort: ids in (depth, id) order
  uint32_t* order_b;         // [P]
  uint32_t* ts_counts;       // [TS_BLOCKS][tiles] per-chunk tile counts -> exclusive prefixes (tile_scatter.cu)
  uint32_t* ts_totals;       // [tiles] instances of every tile
  uint32_t* ts_start;        // [tiles] first list position of every tile
  uint32_t* chunk_offsets;   // [ceil(P/256)+1] exclusive instance offsets of 256-Gaussian chunks, depth order
  // two-level binning (tile_scatter.cu level 1 on 4x4-tile bins, tile_refine.cu level 2)
  uint32_t* chunk_offsets_c; // [ceil(P/256)+1] the same for bin-level instances
  uint32_t* ts_ctotals;      // [bins + 1] bin-level instances of every bin
  uint32_t* ts_cstart;       // [bins + 1] first bin-list position of every bin
  uint32_t* ts_seg_prefix;   // [bins + 1] first segment (TS2_SEG bin-list entries) of every bin; [bins] = segment count
  unsigned long long* total_c; // [1] bin-level instance count
  uint32_t* refine_ticket;   // [1] (zeroed region)
  // --- small region zeroed at the start of every forward ---
  uint32_t* hist;            // [4*256] depth digit histograms -> exclusive offsets
  uint32_t* tickets;         // [4]
  uint32_t* lookback;        // [4 * sort_blocks(P) * 256]
  unsigned long long* total; // [1] num_rendered
  int* error_flag;           // [1]
  size_t zero_begin, zero_end;
};

__host__ __device__ inline int num_pre_blocks(int P) { return (P + 255) / 256; }
__host__ __device__ inline size_t num_tiles(int W, int H) {
  return (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
}
constexpr int SORT_RADIX_BITS = 8;   // widest digit of a pass
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_TILE_MIN = 256 * 8;             // smallest tile a sort pass uses (look-back words are sized for it)
#ifndef GSR_SORT_SMALL_N_LOG2
#define GSR_SORT_SMALL_N_LOG2 21
#endif
constexpr long long SORT_SMALL_N = 1ll << GSR_SORT_SMALL_N_LOG2;      // up to here the 2048-key tile is used (more CTAs in flight)
constexpr int SORT_MAX_PASSES = 8;
__host__ __device__ inline size_t sort_blocks(int64_t n) { return (size_t)((n + SORT_TILE_MIN - 1) / SORT_TILE_MIN); }

// Digit layout of an LSD sort on key bits [0, total_bits): ceil(bits/8) passes of (nearly) EQUAL width, so that
// e.g. an 11-bit tile id is sorted as 6+5 bits (two passes of long runs) rather than 8+3.
struct SortPlan {
  int n_passes;
  int shift[SORT_MAX_PASSES];
  int bits[SORT_MAX_PASSES];
};
inline SortPlan make_sort_plan(int total_bits) {
  SortPlan pl{};
  if (total_bits < 1) total_bits = 1;
  pl.n_passes = (total_bits + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
  const int base = total_bits / pl.n_passes, rem = total_bits % pl.n_passes;
  int sh = 0;
  for (int p = 0; p < pl.n_passes; ++p) {
    pl.bits[p] = base + (p < rem ? 1 : 0);
    pl.shift[p] = sh;
    sh += pl.bits[p];
  }
  return pl;
}

constexpr int TS_MAX_TILES = 4096;   // above this the per-tile shared-memory state no longer fits: radix path
// Two-level binning: level 1 scatters into bins of TS2_CB x TS2_CB tiles, level 2 splits every bin list into its tiles.
constexpr int TS2_CSHIFT = 2;
constexpr int TS2_CB = 1 << TS2_CSHIFT;
constexpr int TS2_BIN_TILES = TS2_CB * TS2_CB;         // 16 tiles per bin = one bit each of a u16 mask
constexpr int TS2_SEG = 2048;                          // bin-list entries one level-2 CTA handles
__host__ __device__ inline size_t num_bins(int W, int H) {
  const size_t tx = (W + TILE_X - 1) / TILE_X, ty = (H + TILE_Y - 1) / TILE_Y;
  return ((tx + TS2_CB - 1) / TS2_CB) * ((ty + TS2_CB - 1) / TS2_CB);
}
#ifndef TS_ITEMS_PER_THREAD
#define TS_ITEMS_PER_THREAD 8          // instances per thread and piece of the scatter kernel
#endif
#ifndef TS_MAX_OCC
#define TS_MAX_OCC 4                   // CTAs per SM the scatter kernel's registers allow
#endif
constexpr int TS_GROUP_SMEM = 16512 + (256 * TS_ITEMS_PER_THREAD / 32 * 33) * 4;  // staged Gaussians + staged piece
constexpr int TS_TILE_SMEM = 24;     // per tile: cursor + piece base (u32) + 8 per-warp u16 counters
// CTAs of the sort-free binning = rows of its count matrix: as many per SM as the scatter kernel's shared memory
// allows (at most TS_MAX_OCC: its register budget), on the 148 SMs of a B200.
__host__ __device__ inline int ts_blocks(size_t n_tiles) {
  const size_t smem = (size_t)TS_GROUP_SMEM + n_tiles * TS_TILE_SMEM + 1024;
  size_t occ = (size_t)232448 / smem;
  occ = occ > TS_MAX_OCC ? TS_MAX_OCC : (occ < 1 ? 1 : occ);
  return 148 * (int)occ;
}
#ifndef GSR_TS_COARSE_OCC
#define GSR_TS_COARSE_OCC TS_MAX_OCC
#endif
__host__ __device__ inline int ts_blocks_coarse(size_t n_bins) {  // level 1 of the two-level binning (+ staged masks)
  const size_t smem = (size_t)TS_GROUP_SMEM + 256 * TS_ITEMS_PER_THREAD * 2 + n_bins * TS_TILE_SMEM + 1024;
  size_t occ = (size_t)232448 / smem;
  occ = occ > GSR_TS_COARSE_OCC ? GSR_TS_COARSE_OCC : (occ < 1 ? 1 : occ);
  return 148 * (int)occ;
}
__host__ __device__ inline size_t carve_geom(char* base, int P, int W, int H, GeomViews* v) {
  const size_t ts_tiles = num_tiles(W, H) <= (size_t)TS_MAX_TILES ? num_tiles(W, H) : 0;
  const size_t ts_bins = num_bins(W, H) <= (size_t)TS_MAX_TILES ? num_bins(W, H) : 0;
  const size_t fine_tiles = ts_bins ? num_tiles(W, H) : 0;
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  char* a = take(sizeof(float4) * 3 * (size_t)P);
  char* t = take(sizeof(uint32_t) * (size_t)P);
  char* rc = take(sizeof(uint2) * (size_t)P); char* rs = take(sizeof(uint2) * (size_t)P);
  char* cl = take((size_t)3 * P);
  char* ka = take(4 * (size_t)P); char* kb = take(4 * (size_t)P);
  char* oa = take(4 * (size_t)P); char* ob = take(4 * (size_t)P);
  char* co = take(sizeof(uint32_t) * (size_t)(num_pre_blocks(P) + 1));
  const size_t tc_words = (size_t)ts_blocks(ts_tiles) * ts_tiles, tc_words_c = (size_t)ts_blocks_coarse(ts_bins) * ts_bins;
  char* tc = take(sizeof(uint32_t) * (tc_words > tc_words_c ? tc_words : tc_words_c));
  char* tto = take(sizeof(uint32_t) * (fine_tiles + 1));
  char* tst = take(sizeof(uint32_t) * (fine_tiles + 1));
  char* coc = take(sizeof(uint32_t) * (size_t)(num_pre_blocks(P) + 1));
  char* cto = take(sizeof(uint32_t) * (ts_bins + 1));
  char* cst = take(sizeof(uint32_t) * (ts_bins + 1));
  char* csp = take(sizeof(uint32_t) * (ts_bins + 1));
  size_t zb = off;
  char* h = take(sizeof(uint32_t) * 4 * 256);
  char* tk = take(sizeof(uint32_t) * 4);
  char* lb = take(sizeof(uint32_t) * 4 * sort_blocks(P > 0 ? P : 1) * SORT_RADIX);
  char* tot = take(256);
  size_t ze = off;
  if (v) {
    v->rec = (float4*)a;
    v->tiles_touched = (uint32_t*)t; v->clamped = (uint8_t*)cl;
    v->rect = (uint2*)rc; v->rect_sorted = (uint2*)rs;
    v->dkeys_a = (uint32_t*)ka; v->dkeys_b = (uint32_t*)kb; v->order_a = (uint32_t*)oa; v->order_b = (uint32_t*)ob;
    v->chunk_offsets = (uint32_t*)co;
    v->ts_counts = (uint32_t*)tc; v->ts_totals = (uint32_t*)tto; v->ts_start = (uint32_t*)tst;
    v->hist = (uint32_t*)h; v->tickets = (uint32_t*)tk; v->lookback = (uint32_t*)lb;
    v->total = (unsigned long long*)tot; v->error_flag = (int*)(tot ? tot + 64 : nullptr);
    v->total_c = (unsigned long long*)(tot ? tot + 128 : nullptr);
    v->refine_ticket = (uint32_t*)(tot ? tot + 192 : nullptr);
    v->chunk_offsets_c = (uint32_t*)coc; v->ts_ctotals = (uint32_t*)cto; v->ts_cstart = (uint32_t*)cst;
    v->ts_seg_prefix = (uint32_t*)csp;
    v->zero_begin = zb; v->zero_end = ze;
  }
  return off;
}

// Per-pixel + per-tile state ("ImageState", rasterizer_impl.h:50-57).
struct ImageViews {
  float* final_T;        // [H*W]
  uint32_t* n_contrib;   // [H*W]
  uint2* ranges;         // [tiles]
  uint32_t* tile_work;   // [tiles] largest n_contrib of the tile (written by the forward renderer)
  uint32_t* tile_order;  // [tiles] tiles by descending tile_work: the backward renderer's CTA -> tile map
  uint32_t* tile_order_fwd;  // [tiles] the same from the PREVIOUS forward's tile_work: the forward renderer's map
};
__host__ __device__ inline size_t carve_image(char* base, int W, int H, ImageViews* v) {
  size_t off = 0;
  size_t N = (size_t)W * H;
  size_t tiles = num_tiles(W, H);
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  char* a = take(4 * N); char* b = take(4 * N); char* c = take(8 * tiles);
  char* d = take(4 * tiles); char* e = take(4 * tiles); char* f = take(4 * tiles);
  if (v) {
    v->final_T = (float*)a; v->n_contrib = (uint32_t*)b; v->ranges = (uint2*)c;
    v->tile_work = (uint32_t*)d; v->tile_order = (uint32_t*)e; v->tile_order_fwd = (uint32_t*)f;
  }
  return off;
}

// Instance binning workspace ("BinningState", rasterizer_impl.h:59-69, minus the final point_list which the
// caller owns).  Instances are emitted in DEPTH order, so only the tile id (<= 16 bits) is left to sort on.
struct BinningViews {
  uint32_t* tkeys_a;     // [R] tile ids, emission (= depth) order
  uint32_t* tkeys_b;     // [R]
  uint32_t* vals_a;      // [R] Gaussian ids
  uint32_t* vals_b;      // [R]
  uint64_t* keys64_a;    // generic u64 sorter (gsr_sort_pairs_u64 test hook) aliases of the same memory
  uint64_t* keys64_b;
  uint32_t* hist;        // [8 * 256] digit histograms -> exclusive offsets
  uint32_t* tickets;     // [8]
  uint32_t* lookback;    // [8 * sort_blocks(R) * 256]
  size_t zero_begin;     // byte range [zero_begin, zero_end(n_passes)) must be zeroed before a sort
  size_t lookback_off;   // byte offset of `lookback`
  size_t lookback_pass_bytes;
  size_t zero_end(int n_passes) const { return lookback_off + (size_t)n_passes * lookback_pass_bytes; }
};
__host__ __device__ inline size_t carve_binning(char* base, int64_t R, BinningViews* v) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  size_t n = (size_t)(R > 0 ? R : 1);
  char* ka = take(8 * n); char* kb = take(8 * n); char* va = take(4 * n); char* vb = take(4 * n);
  size_t zb = off;
  char* h = take(4 * (size_t)SORT_MAX_PASSES * SORT_RADIX);
  char* t = take(4 * (size_t)SORT_MAX_PASSES);
  char* lb = take(4 * (size_t)SORT_MAX_PASSES * sort_blocks(n) * SORT_RADIX);
  if (v) {
    v->tkeys_a = (uint32_t*)ka; v->tkeys_b = (uint32_t*)kb; v->vals_a = (uint32_t*)va; v->vals_b = (uint32_t*)vb;
    v->keys64_a = (uint64_t*)ka; v->keys64_b = (uint64_t*)kb;
    v->hist = (uint32_t*)h; v->tickets = (uint32_t*)t; v->lookback = (uint32_t*)lb;
    v->zero_begin = zb;
    v->lookback_off = (size_t)(lb - base);
    v->lookback_pass_bytes = 4 * sort_blocks(n) * SORT_RADIX;
  }
  return off;
}

// Two-level binning workspace (aliases the same caller buffer as BinningViews; never used together): the bin lists
// written by level 1 and the per-segment tile counts of level 2.  Sized for `R` = the list capacity (the number of
// bin-level instances never exceeds the number of tile-level instances).
struct RefineViews {
  uint32_t* ids;         // [R] Gaussian ids, bin-major, (depth, id) order inside a bin
  uint16_t* masks;       // [R] which of the bin's 16 tiles the Gaussian's rectangle covers (bit 4*dy + dx)
  uint32_t* seg_bin;     // [max_segs] bin of every segment
  uint32_t* seg_counts;  // [max_segs][16] per-segment tile counts -> exclusive prefixes over the segments of a bin
  uint32_t* warp_counts; // [max_segs][8][16] the same per warp (256 entries) of the segment
  size_t max_segs;
};
__host__ __device__ inline size_t refine_max_segs(int64_t R, size_t n_bins) {
  return (size_t)((R > 0 ? R : 0) / TS2_SEG) + n_bins + 1;
}
__host__ __device__ inline size_t carve_refine(char* base, int64_t R, size_t n_bins, RefineViews* v) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  const size_t n = (size_t)(R > 0 ? R : 1), segs = refine_max_segs(R, n_bins);
  char* a = take(4 * n); char* b = take(2 * n); char* c = take(4 * segs); char* d = take(4 * segs * TS2_BIN_TILES);
  char* e = take(4 * segs * TS2_BIN_TILES * (TS2_SEG / 256));
  if (v) {
    v->ids = (uint32_t*)a; v->masks = (uint16_t*)b; v->seg_bin = (uint32_t*)c; v->seg_counts = (uint32_t*)d;
    v->warp_counts = (uint32_t*)e;
    v->max_segs = segs;
  }
  return off;
}

// Packed per-Gaussian partial gradients accumulated by the backward renderer
// (one 48-byte record so the 10 atomics of a warp hit 1-2 sectors):
//   [0]=dL/dmean2D.x [1]=dL/dmean2D.y [2]=dL/dconic.x [3]=dL/dconic.y
//   [4]=dL/dconic.w  [5]=dL/dopacity  [6..8]=dL/dcolor rgb [9]=dL/ddepth
constexpr int GRAD_STRIDE = 12;
struct BackwardViews {
  float* packed;        // [P * GRAD_STRIDE]
  double* tau_partial;  // [nblocks * 6]
  uint16_t* touched_list;   // [nblocks * 256] chunk-local indices of the Gaussians with a non-zero record, ascending
  uint32_t* touched_count;  // [nblocks]
  uint8_t* touched_flag;    // [P] 1 = the backward renderer added something to the Gaussian's record
  size_t zero_bytes;        // leading bytes (packed + tau_partial + touched_flag) that must be zero before the backward renderer runs
};
__host__ __device__ inline size_t carve_backward(char* base, int P, BackwardViews* v) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  char* a = take(sizeof(float) * GRAD_STRIDE * (size_t)P);
  char* b = take(sizeof(double) * 6 * (size_t)num_pre_blocks(P));
  char* tf = take((size_t)P);
  const size_t zb = off;
  char* c = take(sizeof(uint16_t) * 256 * (size_t)num_pre_blocks(P));
  char* d = take(sizeof(uint32_t) * (size_t)num_pre_blocks(P));
  if (v) {
    v->packed = (float*)a; v->tau_partial = (double*)b; v->touched_list = (uint16_t*)c; v->touched_count = (uint32_t*)d;
    v->touched_flag = (uint8_t*)tf;
    v->zero_bytes = zb;
  }
  return off;
}

// ---------------------------------------------------------------------------
// Explicitly rounded fp32 arithmetic.  nvcc never re-contracts these
// intrinsics, so the quantities that feed integer outputs (radii, tile rects,
// depth key bits, the n_touched threshold) follow exactly the operation
// sequence the reference kernels execute (its FMA contraction was read from
// the PTX nvcc 12.9 generates for forward.cu; see DESIGN.md "Bit-exactness").
// ---------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float frcp(float a) { return __frcp_rn(a); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }

// a0*b0 + a1*b1 + a2*b2 the way nvcc contracts a left-associated sum of
// products: fma(a2,b2, fma(a0,b0, a1*b1)).
__device__ __forceinline__ float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
  return ffma(a2, b2, ffma(a0, b0, fmul(a1, b1)));
}
// matrix[r] * x + matrix[r+4] * y + matrix[r+8] * z + matrix[r+12]  (auxiliary.h:58-77)
__device__ __forceinline__ float affine_row(const float* m, int r, float x, float y, float z) {
  return fadd(m[r + 12], dot3c(x, m[r], y, m[r + 4], z, m[r + 8]));
}

// power = -0.5f*(con.x*dx*dx + con.z*dy*dy) - con.y*dx*dy  (forward.cu:345), in the operation order of the
// reference's SASS: q = fma(dx, dx*cx, dy*(dy*cz)); power = fma(q, -0.5, -(dy*(dx*cy))).
__device__ __forceinline__ float gauss_power(float dx, float dy, float cx, float cy, float cz) {
  const float q = ffma(dx, fmul(dx, cx), fmul(dy, fmul(dy, cz)));
  return ffma(q, -0.5f, -fmul(dy, fmul(dx, cy)));
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

#define GSR_DEVINL __device__ __forceinline__

// Lanes of the warp that hold the same value in the low NBITS bits (all 32 lanes must call). One ballot per bit:
// constant cost, where match.any takes time proportional to the number of DISTINCT values in the warp -- high for
// the nearly all-distinct tile ids ranked by tile_scatter.cu (-8 us there; GSR_MATCH_ANY=1 selects the match.any
// instruction). The 8-bit digits of the onesweep passes keep match.any: ballots were 2x slower in that kernel.
#ifndef GSR_MATCH_ANY
#define GSR_MATCH_ANY 0
#endif
template <int NBITS>
GSR_DEVINL uint32_t warp_peers(uint32_t v) {
#if GSR_MATCH_ANY
  return __match_any_sync(0xffffffffu, v);
#else
  uint32_t peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < NBITS; ++b) {
    const bool bit = (v >> b) & 1u;
    const uint32_t vote = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? vote : ~vote;
  }
  return peers;
#endif
}

// Programmatic dependent launch: every kernel of the library lets its successor in the stream start launching at
// once (pdl_launch_dependents) and waits for its predecessor's results (pdl_wait) before it touches global memory,
// so launch latency, CTA scheduling and the shared-memory prologue overlap the predecessor's tail.  Every kernel
// must execute pdl_wait on every path: completion of a grid then implies completion of all earlier grids.
// Tiles by descending replay length (counting sort on tile_work / 4) by ONE CTA of TILE_ORDER_THREADS threads: the
// body of tile_order_kernel (render_bwd.cu), also run by an extra CTA of chunk_scan_kernel for the forward renderer.
constexpr int TILE_ORDER_THREADS = 1024;
GSR_DEVINL void tile_order_body(int n_tiles, const uint32_t* __restrict__ tile_work, uint32_t* __restrict__ tile_order) {
  __shared__ uint32_t s_bin[TILE_ORDER_THREADS];
  __shared__ uint32_t s_warp[TILE_ORDER_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  s_bin[tid] = 0;
  __syncthreads();
  for (int t = tid; t < n_tiles; t += TILE_ORDER_THREADS)
    atomicAdd(&s_bin[TILE_ORDER_THREADS - 1 - min(tile_work[t] >> 2, (uint32_t)TILE_ORDER_THREADS - 1)], 1u);
  __syncthreads();
  // exclusive scan of the 1024 bins (bin 0 = heaviest)
  const uint32_t c = s_bin[tid];
  uint32_t inc = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += v;
    }
    s_warp[lane] = w;
  }
  __syncthreads();
  s_bin[tid] = inc - c + (warp ? s_warp[warp - 1] : 0u);
  __syncthreads();
  for (int t = tid; t < n_tiles; t += TILE_ORDER_THREADS) {
    const uint32_t pos = atomicAdd(&s_bin[TILE_ORDER_THREADS - 1 - min(tile_work[t] >> 2, (uint32_t)TILE_ORDER_THREADS - 1)], 1u);
    tile_order[pos] = (uint32_t)t;
  }
}

// Tile rectangle (x0 | x1 << 16, y0 | y1 << 16, exclusive ends) -> rectangle of TS2_CB x TS2_CB-tile bins.
GSR_DEVINL uint2 ts_coarse_rect(uint2 r) {
  const uint32_t x0 = r.x & 0xffffu, x1 = r.x >> 16, y0 = r.y & 0xffffu, y1 = r.y >> 16;
  if (x1 <= x0 || y1 <= y0) return make_uint2(0u, 0u);
  return make_uint2((x0 >> TS2_CSHIFT) | (((x1 + TS2_CB - 1) >> TS2_CSHIFT) << 16),
                    (y0 >> TS2_CSHIFT) | (((y1 + TS2_CB - 1) >> TS2_CSHIFT) << 16));
}
GSR_DEVINL void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
GSR_DEVINL void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // api.cu: environment GSR_PDL=0 turns the launch attribute off (A/B measurements)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                            Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// Fused L1 loss (loss.cu): what the backward tile renderer needs to form dL/dpixel on the fly.
struct LossView {
  int kind;  // GSR_LOSS_NONE: use the dL_dpix images instead
  const float* color; const float* depth; const float* opacity;   // rendered images of the forward
  const float* gt_rgb; const float* gt_depth;                      // targets ([3,H,W], [1,H,W] or NULL)
  const uint8_t* mask; const uint8_t* grad_mask;                   // [H,W] 0/1 or NULL
  const float* exposure_a; const float* exposure_b;                // device scalars or NULL
  const float* out;                                                // out[8] of launch_l1_loss (device)
  const float* upstream;                                           // dL/dloss (device scalar) or NULL = 1
};

#ifdef __CUDACC__
// dL/d(colour rgb, depth) of pixel `pix` for the fused L1 losses (the per-pixel half of loss.cu's reduction):
//   dL/dI_c = up * k_rgb * sign(e^a I_c + b - gt_c) on the masked pixels, dL/dD = up * k_d * sign(D - gtD) on the
//   valid-depth pixels (k_rgb = out[3] has e^a folded in, k_d = out[4]).
// Used by the backward tile renderer's prologue and by gsr_l1_loss_grad, so both form the very same numbers.
__device__ __forceinline__ void loss_pixel_grad(const LossView& loss, size_t pix, size_t HW, float& dpr, float& dpg,
                                                float& dpb, float& dpd) {
  dpr = dpg = dpb = dpd = 0.0f;
  const float up = loss.upstream ? *loss.upstream : 1.0f;
  const float ea = loss.exposure_a ? expf(*loss.exposure_a) : 1.0f;
  const float eb = loss.exposure_b ? *loss.exposure_b : 0.0f;
  const bool om = loss.kind == GSR_LOSS_TRACKING ? (loss.opacity[pix] > 0.99f) : true;
  bool m = loss.mask ? loss.mask[pix] != 0 : true;
  if (loss.kind == GSR_LOSS_TRACKING) m = m && om && (loss.grad_mask ? loss.grad_mask[pix] != 0 : true);
  auto sgn = [](float d) { return d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); };
  if (m) {
    const float k = up * loss.out[3];
    dpr = k * sgn(ea * loss.color[pix] + eb - loss.gt_rgb[pix]);
    dpg = k * sgn(ea * loss.color[HW + pix] + eb - loss.gt_rgb[HW + pix]);
    dpb = k * sgn(ea * loss.color[2 * HW + pix] + eb - loss.gt_rgb[2 * HW + pix]);
  }
  if (loss.gt_depth) {
    const float gd = loss.gt_depth[pix];
    if (gd > 0.f && om) dpd = up * loss.out[4] * sgn(loss.depth[pix] - gd);
  }
}
#endif

// Camera constants passed by value (kernel parameter -> constant bank).
struct CameraParams {
  float view[16];      // W2C^T, flat index 4*col+row  
  float proj[16];      // (P*W2C)^T
  float campos[3];
  float tan_fovx, tan_fovy;
  float focal_x, focal_y;
  int W, H;
  int tiles_x, tiles_y;
};

}  // namespace gsr
