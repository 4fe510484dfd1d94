// radix_sort.cu -- stable LSD radix sort of (key, u32 value) pairs, one kernel per pass with decoupled
// look-back ("onesweep"), hand-written for sm_100a.  Replaces cub::DeviceRadixSort::SortPairs at
// rasterizer_impl.cu:306-311 (u64 tile|depth keys) and simple_knn.cu:210-213 (u32 Morton keys).
//
// A pass sorts on `bits` (1..8) key bits starting at `shift`; the caller chooses the digit widths (SortPlan), e.g.
// 4 x 8 for the 32 depth bits and two BALANCED digits (6+5 at 1200 tiles) for the tile id, so that every pass
// scatters into few, long runs.  Per pass, per tile of 256 x ITEMS keys (warp-striped: a warp owns 32*ITEMS
// consecutive keys, so stability follows from lane/item order):
//   1. rank keys inside the warp with match.any on the digit + per-warp digit counters in shared memory;
//   2. thread d combines the 8 warp counters of digit d, publishes the tile's count for d and walks back over the
//      preceding tiles' published counts (flagged 32-bit words: bits 31..30 = state, 29..0 = count), EIGHT
//      predecessors per round trip -- the inputs here are small (1-20 M keys = 1-3 waves of CTAs), so most
//      predecessors are still running and a one-at-a-time walk would serialise on L2 latency;
//   3. keys and values are scattered into shared memory in tile-sorted order and written out with one coalesced
//      run per digit.
// The element count may live in device memory (n_dev): the forward pass launches the R-level sort before the host
// knows num_rendered; CTAs beyond the actual count exit at once.
// The digit histograms needed up front are NOT computed from the keys on the rasterizer path: preprocess derives the
// depth digits per Gaussian and the instance emitter derives the tile digits (see preprocess.cu / binning.cu).
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
#ifndef GSR_RS_LOOKBACK_BATCH
#define GSR_RS_LOOKBACK_BATCH 8
#endif
constexpr int RS_LOOKBACK_BATCH = GSR_RS_LOOKBACK_BATCH;
#ifndef GSR_RS_MATCH_BATCH
#define GSR_RS_MATCH_BATCH 8
#endif
constexpr int RS_MATCH_BATCH = GSR_RS_MATCH_BATCH;
constexpr uint32_t FLAG_LOCAL = 1u << 30;
constexpr uint32_t FLAG_INCL = 2u << 30;
constexpr uint32_t FLAG_MASK = 3u << 30;
constexpr uint32_t VALUE_MASK = ~FLAG_MASK;

template <typename KeyT, int ITEMS>
struct RsSmem {
  KeyT keys[RS_THREADS * ITEMS];
  uint32_t vals[RS_THREADS * ITEMS];
  uint32_t warp_hist[RS_WARPS][SORT_RADIX];
  uint32_t bin_start[SORT_RADIX];    // exclusive digit prefix inside the tile
  uint32_t global_base[SORT_RADIX];  // global index of tile-position 0 of each digit (mod 2^32)
  uint32_t warp_tot[RS_WARPS];
  uint32_t block_id;
};

GSR_DEVINL uint32_t ld_volatile_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
GSR_DEVINL void st_volatile_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <typename KeyT, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS, ITEMS == 16 ? 3 : 4)
onesweep_pass_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, long long n_cap,
                     const unsigned long long* __restrict__ n_dev, int shift, uint32_t dmask,
                     const uint32_t* __restrict__ digit_offsets /*[256] exclusive*/,
                     uint32_t* __restrict__ ticket, uint32_t* __restrict__ lookback /*[nblocks][256]*/) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int TILE = RS_THREADS * ITEMS;
  extern __shared__ __align__(16) unsigned char rs_smem_raw[];
  RsSmem<KeyT, ITEMS>& sm = *reinterpret_cast<RsSmem<KeyT, ITEMS>*>(rs_smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;

  if (tid == 0) sm.block_id = atomicAdd(ticket, 1u);
#pragma unroll
  for (int i = tid; i < RS_WARPS * SORT_RADIX; i += RS_THREADS) (&sm.warp_hist[0][0])[i] = 0;
  __syncthreads();
  const uint32_t bid = sm.block_id;
  long long n = n_cap;
  if (n_dev) n = min((long long)*n_dev, n_cap);
  const long long tile_base = (long long)bid * TILE;
  if (tile_base >= n) return;  // launched for the capacity, not needed for the actual count
  const int tile_n = (int)min((long long)TILE, n - tile_base);

  // ---- load (warp-striped) ----
  KeyT key[ITEMS];
  uint32_t val[ITEMS];
  uint32_t rank[ITEMS];
  const int wbase = w * 32 * ITEMS;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int t = wbase + i * 32 + lane;
    key[i] = (t < tile_n) ? keys_in[tile_base + t] : ~KeyT(0);
  }
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int t = wbase + i * 32 + lane;
    val[i] = (t < tile_n) ? vals_in[tile_base + t] : 0u;
  }

  // ---- rank inside the warp (padding keys take the largest digit and sort last); the match.any of a batch of
  //      items are issued back to back so that their latencies overlap, then the counters are updated in order ----
  constexpr int MB = ITEMS < RS_MATCH_BATCH ? ITEMS : RS_MATCH_BATCH;
#pragma unroll
  for (int i0 = 0; i0 < ITEMS; i0 += MB) {
    uint32_t dg[MB], peers[MB];
#pragma unroll
    for (int j = 0; j < MB; ++j) {
      const int t = wbase + (i0 + j) * 32 + lane;
      dg[j] = (t < tile_n) ? (uint32_t)((key[i0 + j] >> shift) & dmask) : dmask;
      peers[j] = __match_any_sync(0xffffffffu, dg[j]);  // (ballot-based peer masks measured 2x slower here)
    }
#pragma unroll
    for (int j = 0; j < MB; ++j) {
      const uint32_t below = __popc(peers[j] & ((1u << lane) - 1u));
      uint32_t prev = 0;
      if (below == 0) {
        prev = sm.warp_hist[w][dg[j]];
        sm.warp_hist[w][dg[j]] = prev + __popc(peers[j]);
      }
      prev = __shfl_sync(0xffffffffu, prev, __ffs(peers[j]) - 1);
      rank[i0 + j] = prev + below;
      __syncwarp();
    }
  }
  __syncthreads();

  // ---- per-digit combine + look-back (thread d owns digit d; threads above the radix idle along) ----
  {
    const uint32_t d = (uint32_t)tid;
    const bool owner = d <= dmask;
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < RS_WARPS; ++ww) {
      uint32_t c = sm.warp_hist[ww][d];
      sm.warp_hist[ww][d] = run;
      run += c;
    }
    uint32_t count = run;
    if (d == dmask) count -= (uint32_t)(TILE - tile_n);  // padding keys were ranked as the largest digit
    uint32_t* my_status = lookback + (size_t)bid * SORT_RADIX + d;
    if (owner) st_volatile_u32(my_status, (bid == 0 ? FLAG_INCL : FLAG_LOCAL) | count);

    // exclusive scan of `run` over digits (tile-local bin starts; padding included, it sorts last)
    uint32_t inc = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) sm.warp_tot[w] = inc;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int ww = 0; ww < RS_WARPS; ++ww) woff += (ww < w) ? sm.warp_tot[ww] : 0u;
    const uint32_t bstart = woff + inc - run;
    sm.bin_start[d] = bstart;

    uint32_t excl = 0;
    if (owner && bid > 0) {
      long long p = (long long)bid - 1;
      bool done = false;
      while (!done) {
        uint32_t s[RS_LOOKBACK_BATCH];
#pragma unroll
        for (int i = 0; i < RS_LOOKBACK_BATCH; ++i) {
          const long long q = p - i;
          s[i] = q >= 0 ? ld_volatile_u32(lookback + (size_t)q * SORT_RADIX + d) : FLAG_INCL;
        }
        int consumed = 0;
#pragma unroll
        for (int i = 0; i < RS_LOOKBACK_BATCH; ++i) {
          const uint32_t f = s[i] & FLAG_MASK;
          if (!done && consumed == i && f != 0) {  // stop at the first predecessor that has not published yet
            excl += s[i] & VALUE_MASK;
            consumed = i + 1;
            done = f == FLAG_INCL;
          }
        }
        p -= consumed;
      }
      st_volatile_u32(my_status, FLAG_INCL | ((excl + count) & VALUE_MASK));
    }
    if (owner) sm.global_base[d] = digit_offsets[d] + excl - bstart;
  }
  __syncthreads();

  // ---- scatter into tile-sorted order in shared memory ----
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int t = wbase + i * 32 + lane;
    const uint32_t d = (t < tile_n) ? (uint32_t)((key[i] >> shift) & dmask) : dmask;
    const uint32_t pos = sm.bin_start[d] + sm.warp_hist[w][d] + rank[i];
    sm.keys[pos] = key[i];
    sm.vals[pos] = val[i];
  }
  __syncthreads();

  // ---- coalesced write-out (padding occupies positions >= tile_n) ----
#pragma unroll 4
  for (int p = tid; p < tile_n; p += RS_THREADS) {
    const KeyT k = sm.keys[p];
    const uint32_t d = (uint32_t)((k >> shift) & dmask);
    const uint32_t g = sm.global_base[d] + (uint32_t)p;
    keys_out[g] = k;
    vals_out[g] = sm.vals[p];
  }
}

// Generic digit histogram over the keys (test hook / kNN path).
template <typename KeyT>
__global__ void __launch_bounds__(256)
digit_hist_kernel(const KeyT* __restrict__ keys, long long n, SortPlan plan, uint32_t* __restrict__ hist) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s[SORT_MAX_PASSES * SORT_RADIX];
  for (int i = threadIdx.x; i < plan.n_passes * SORT_RADIX; i += 256) s[i] = 0;
  __syncthreads();
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    KeyT k = keys[i];
    for (int p = 0; p < plan.n_passes; ++p)
      atomicAdd(&s[p * SORT_RADIX + (uint32_t)((k >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < plan.n_passes * SORT_RADIX; i += 256)
    if (s[i]) atomicAdd(&hist[i], s[i]);
}

// Exclusive prefix over the 256 bins of each pass' digit histogram (one warp per pass).
__global__ void digit_offsets_kernel(uint32_t* hist, int n_passes) {
  pdl_launch_dependents();
  pdl_wait();
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

cudaError_t launch_digit_offsets(cudaStream_t st, uint32_t* hist, int n_passes) {
  count_launch(); launch_k(digit_offsets_kernel, 1, 32 * n_passes, 0, st, hist, n_passes);
  return cudaGetLastError();
}

template <typename KeyT, int ITEMS>
static cudaError_t launch_pass(cudaStream_t st, const KeyT* kin, KeyT* kout, const uint32_t* vin, uint32_t* vout,
                               long long n_cap, const unsigned long long* n_dev, int shift, int bits,
                               const uint32_t* digit_offsets, uint32_t* ticket, uint32_t* lookback) {
  static bool attr_set = false;
  const size_t smem = sizeof(RsSmem<KeyT, ITEMS>);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(onesweep_pass_kernel<KeyT, ITEMS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const unsigned nblocks = (unsigned)((n_cap + RS_THREADS * ITEMS - 1) / (RS_THREADS * ITEMS));
  count_launch();
  launch_k(onesweep_pass_kernel<KeyT, ITEMS>, nblocks, RS_THREADS, smem, st, kin, kout, vin, vout, n_cap, n_dev, shift,
                                                                      (1u << bits) - 1u, digit_offsets, ticket,
                                                                      lookback);
  return cudaGetLastError();
}

template <typename KeyT>
static cudaError_t sort_passes_impl(cudaStream_t st, long long n_cap, const unsigned long long* n_dev, KeyT* keys_a,
                                    KeyT* keys_b, uint32_t* vals_a, uint32_t* vals_b, const SortPlan& plan,
                                    const uint32_t* digit_offsets, uint32_t* tickets, uint32_t* lookback,
                                    uint32_t* final_vals) {
  if (n_cap <= 0) return cudaSuccess;
  const size_t lb_stride = sort_blocks(n_cap) * SORT_RADIX;  // words per pass, sized for the small tile
  const bool small = n_cap <= SORT_SMALL_N;
  KeyT* kin = keys_a; KeyT* kout = keys_b;
  uint32_t* vin = vals_a; uint32_t* vout = vals_b;
  for (int p = 0; p < plan.n_passes; ++p) {
    uint32_t* vdst = (p == plan.n_passes - 1 && final_vals) ? final_vals : vout;
    cudaError_t e = small
        ? launch_pass<KeyT, 8>(st, kin, kout, vin, vdst, n_cap, n_dev, plan.shift[p], plan.bits[p],
                               digit_offsets + p * SORT_RADIX, tickets + p, lookback + (size_t)p * lb_stride)
        : launch_pass<KeyT, 16>(st, kin, kout, vin, vdst, n_cap, n_dev, plan.shift[p], plan.bits[p],
                                digit_offsets + p * SORT_RADIX, tickets + p, lookback + (size_t)p * lb_stride);
    if (e != cudaSuccess) return e;
    KeyT* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  return cudaSuccess;
}

cudaError_t sort_pairs_u64_prehist(cudaStream_t st, long long n_cap, const unsigned long long* n_dev,
                                   uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                                   const SortPlan& plan, const uint32_t* digit_offsets, uint32_t* tickets,
                                   uint32_t* lookback, uint32_t* final_vals) {
  return sort_passes_impl<uint64_t>(st, n_cap, n_dev, keys_a, keys_b, vals_a, vals_b, plan, digit_offsets, tickets,
                                    lookback, final_vals);
}

cudaError_t sort_pairs_u32_prehist(cudaStream_t st, long long n_cap, const unsigned long long* n_dev,
                                   uint32_t* keys_a, uint32_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                                   const SortPlan& plan, const uint32_t* digit_offsets, uint32_t* tickets,
                                   uint32_t* lookback, uint32_t* final_vals) {
  return sort_passes_impl<uint32_t>(st, n_cap, n_dev, keys_a, keys_b, vals_a, vals_b, plan, digit_offsets, tickets,
                                    lookback, final_vals);
}

cudaError_t launch_digit_hist_u64(cudaStream_t st, const uint64_t* keys, long long n, const SortPlan& plan,
                                  uint32_t* hist) {
  if (n <= 0) return cudaSuccess;
  int blocks = (int)min((long long)148 * 8, (n + 255) / 256);
  count_launch(); launch_k(digit_hist_kernel<uint64_t>, blocks, 256, 0, st, keys, n, plan, hist);
  return launch_digit_offsets(st, hist, plan.n_passes);
}

cudaError_t launch_digit_hist_u32(cudaStream_t st, const uint32_t* keys, long long n, const SortPlan& plan,
                                  uint32_t* hist) {
  if (n <= 0) return cudaSuccess;
  int blocks = (int)min((long long)148 * 8, (n + 255) / 256);
  count_launch(); launch_k(digit_hist_kernel<uint32_t>, blocks, 256, 0, st, keys, n, plan, hist);
  return launch_digit_offsets(st, hist, plan.n_passes);
}

}  // namespace gsr
