// render_fwd.cu -- forward tile renderer (K7 of SURVEY 2.2).
//
// Replaces renderCUDA<3> (forward.cu:263-392).  One 256-thread block per
// 16x16 tile, but each WARP owns an 8x4 pixel sub-rectangle:
//   * the tile's instance list is staged in batches of 256 through shared
//     memory with cp.async (LDGSTS) gathers of three 16-byte records per
//     Gaussian, double buffered so the gather of batch b+1 overlaps the
//     blending of batch b;
//   * before blending, every lane tests ONE staged Gaussian against the warp's
//     pixel rectangle (conservative reach of alpha >= 1/255, precomputed per
//     Gaussian) and a ballot turns that into a 32-bit survivor mask -- the
//     per-pixel loop then only visits Gaussians that can contribute to this
//     warp, which skips exactly the iterations the reference `continue`s;
//   * a warp whose 32 pixels are all saturated stops blending, the block
//     stops fetching when all 8 warps are done.
// The blended quantities follow the reference's fp32 operation order exactly
// (explicit fmul/ffma), so final_T, n_contrib and the integer n_touched are
// bit-identical and colour/depth/opacity agree to the last bit.
#include <cstdlib>
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

#ifdef GSR_CTA_TRACE
// Debug build only (tools/time_kernels.py --trace): per-CTA start/end globaltimer and SM id.
__device__ unsigned long long g_trace_render_fwd[3 * 32768];
GSR_DEVINL unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
GSR_DEVINL unsigned trace_smid() { unsigned s; asm volatile("mov.u32 %0, %smid;" : "=r"(s)); return s; }
#endif

// A CTA renders RF_WARPS of the 8 warp rectangles of a tile (8: the whole 16x16 tile, 4: a 16x8 half, 2: a 16x4
// quarter); the sub-tile index is folded into blockIdx.y.  Smaller CTAs re-stage the tile's list (cheap: ~15
// instructions per record and thread) but shorten the tail of the launch and stop at their own pixels' saturation.
#ifndef GSR_RENDER_WARPS
#define GSR_RENDER_WARPS 8
#endif
#ifndef GSR_RENDER_BATCH
#define GSR_RENDER_BATCH 256
#endif
constexpr int RF_WARPS = GSR_RENDER_WARPS;
constexpr int RF_SUB = 8 / RF_WARPS;          // CTAs per tile
constexpr int RF_THREADS = 32 * RF_WARPS;
constexpr int RF_BATCH = GSR_RENDER_BATCH;
constexpr int RF_PER_THREAD = RF_BATCH / RF_THREADS;
static_assert(RF_WARPS == 8 || RF_WARPS == 4 || RF_WARPS == 2, "RF_WARPS");
static_assert(RF_BATCH % RF_THREADS == 0 && RF_BATCH % 32 == 0, "RF_BATCH");

struct RfStage {
  float4 a[RF_BATCH];
  float4 b[RF_BATCH];
  float4 c[RF_BATCH];
  uint32_t id[RF_BATCH];
};

GSR_DEVINL void cp_async16(void* smem_dst, const void* gmem_src) {
  uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
GSR_DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
GSR_DEVINL void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

#ifndef GSR_RF_MIN_CTAS
#define GSR_RF_MIN_CTAS 1
#endif
__global__ void __launch_bounds__(RF_THREADS, GSR_RF_MIN_CTAS)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                  int tiles_x, const float4* __restrict__ rec_a, const float4* __restrict__ rec_b,
                  const float4* __restrict__ rec_c, const float* __restrict__ bg, float* __restrict__ final_T,
                  uint32_t* __restrict__ n_contrib, float* __restrict__ out_color, float* __restrict__ out_depth,
                  float* __restrict__ out_opacity, int* __restrict__ n_touched, uint32_t* __restrict__ tile_work,
                  const uint32_t* __restrict__ tile_order) {
  pdl_launch_dependents();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_enter = trace_now();
#endif
  pdl_wait();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_start = trace_now();
#endif
  __shared__ __align__(16) RfStage stage[2];

  const int tid = threadIdx.x, lane = tid & 31;
  int tile_y = blockIdx.y / RF_SUB, tile_x = blockIdx.x;
  if (RF_SUB == 1 && tile_order) {  // heaviest tiles (of the previous forward on this workspace) first
    const uint32_t t = tile_order[blockIdx.y * gridDim.x + blockIdx.x];
    tile_y = (int)(t / (uint32_t)tiles_x);
    tile_x = (int)(t - (uint32_t)tile_y * (uint32_t)tiles_x);
  }
  const int w = (blockIdx.y % RF_SUB) * RF_WARPS + (tid >> 5);   // warp rectangle within the tile
  const int tile = tile_y * tiles_x + tile_x;
  // warp sub-rectangle: 2 columns x 4 rows of 8x4-pixel rectangles
  const int wx0 = tile_x * TILE_X + (w & 1) * 8;
  const int wy0 = tile_y * TILE_Y + (w >> 1) * 4;
  const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
  const bool inside = pxi < W && pyi < H;
  const float pixx = (float)pxi, pixy = (float)pyi;
  const float rx0 = (float)wx0, rx1 = (float)(wx0 + 7), ry0 = (float)wy0, ry1 = (float)(wy0 + 3);

  const uint2 range = ranges[tile];
  const int total = (int)(range.y - range.x);
  const int n_batches = (total + RF_BATCH - 1) / RF_BATCH;

  float T = 1.0f, Cr = 0.f, Cg = 0.f, Cb = 0.f, D = 0.f;
  uint32_t last_contributor = 0;
  bool done = !inside;

  // prologue: gather batch 0, prefetch the ids of batch 1
  uint32_t id_next[RF_PER_THREAD];
#pragma unroll
  for (int k = 0; k < RF_PER_THREAD; ++k) id_next[k] = 0;
  if (n_batches > 0) {
#pragma unroll
    for (int k = 0; k < RF_PER_THREAD; ++k) {
      const int t = tid + k * RF_THREADS;
      if (t < total) {
        const uint32_t id = point_list[range.x + t];
        stage[0].id[t] = id;
        cp_async16(&stage[0].a[t], rec_a + id);
        cp_async16(&stage[0].b[t], rec_b + id);
        cp_async16(&stage[0].c[t], rec_c + id);
      }
    }
    cp_async_commit();
#pragma unroll
    for (int k = 0; k < RF_PER_THREAD; ++k) {
      const int t = tid + k * RF_THREADS;
      if (RF_BATCH + t < total) id_next[k] = point_list[range.x + RF_BATCH + t];
    }
  }

  for (int b = 0; b < n_batches; ++b) {
    const int cur = b & 1;
    // issue the gather of batch b+1 into the other buffer (free since the barrier ending batch b-1)
    if (b + 1 < n_batches) {
      RfStage& nx = stage[cur ^ 1];
#pragma unroll
      for (int k = 0; k < RF_PER_THREAD; ++k) {
        const int t = tid + k * RF_THREADS;
        if ((b + 1) * RF_BATCH + t < total) {
          nx.id[t] = id_next[k];
          cp_async16(&nx.a[t], rec_a + id_next[k]);
          cp_async16(&nx.b[t], rec_b + id_next[k]);
          cp_async16(&nx.c[t], rec_c + id_next[k]);
        }
      }
      cp_async_commit();
#pragma unroll
      for (int k = 0; k < RF_PER_THREAD; ++k) {
        const int e2 = (b + 2) * RF_BATCH + tid + k * RF_THREADS;
        if (e2 < total) id_next[k] = point_list[range.x + e2];
      }
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();

    const RfStage& sg = stage[cur];
    const int count = min(RF_BATCH, total - b * RF_BATCH);
    const bool warp_live = !__all_sync(0xffffffffu, done);
    if (warp_live) {
      for (int g0 = 0; g0 < count; g0 += 32) {
        // lane-parallel conservative cull of 32 staged Gaussians against this warp's rectangle
        const int jj = g0 + lane;
        bool hit = false;
        if (jj < count) {
          const float4 a = sg.a[jj];
          const float ey = sg.c[jj].w;
          hit = (a.x + a.w >= rx0) && (a.x - a.w <= rx1) && (a.y + ey >= ry0) && (a.y - ey <= ry1);
        }
        uint32_t mask = __ballot_sync(0xffffffffu, hit);
        while (mask) {
          const int j = g0 + __ffs(mask) - 1;
          mask &= mask - 1;
          const float4 a = sg.a[j];
          const float4 co = sg.b[j];
          const float dx = fsub(a.x, pixx), dy = fsub(a.y, pixy);
          const float power = gauss_power(dx, dy, co.x, co.y, co.z);
          const float alpha = fminf(fmul(co.w, expf(power)), 0.99f);
          const float test_T = fmul(T, fsub(1.0f, alpha));
          bool act = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
          if (act && test_T < 0.0001f) { done = true; act = false; }
          if (act) {
            const float4 c = sg.c[j];
            Cr = ffma(T, fmul(alpha, c.x), Cr);
            Cg = ffma(T, fmul(alpha, c.y), Cg);
            Cb = ffma(T, fmul(alpha, c.z), Cb);
            D = ffma(T, fmul(alpha, a.z), D);
            T = test_T;
            last_contributor = (uint32_t)(b * RF_BATCH + j + 1);
          }
          // n_touched: pixels where this Gaussian is seen with T(1-alpha) > 0.5 (forward.cu:369-371)
          const uint32_t tv = __ballot_sync(0xffffffffu, act && test_T > 0.5f);
          if (tv != 0 && lane == 0) atomicAdd(&n_touched[sg.id[j]], __popc(tv));
          if (__all_sync(0xffffffffu, done)) { mask = 0; g0 = count; }
        }
      }
    }
    // everyone must be done reading `cur` before batch b+2 is gathered into it
    if (__syncthreads_and(done)) break;
  }
  cp_async_wait<0>();

  if (inside) {
    const size_t pix = (size_t)pyi * W + pxi;
    const size_t HW = (size_t)H * W;
    final_T[pix] = T;
    n_contrib[pix] = last_contributor;
    out_color[pix] = ffma(T, bg[0], Cr);
    out_color[HW + pix] = ffma(T, bg[1], Cg);
    out_color[2 * HW + pix] = ffma(T, bg[2], Cb);
    out_depth[pix] = D;
    out_opacity[pix] = fsub(1.0f, T);
  }
  // the tile's replay length (its largest n_contrib) = the backward renderer's work for this tile: it starts its
  // heaviest tiles first (tile_order_kernel, render_bwd.cu)
  if (RF_SUB == 1 && tile_work) {
    __shared__ uint32_t s_work[RF_WARPS];
    const uint32_t wm = __reduce_max_sync(0xffffffffu, inside ? last_contributor : 0u);
    if (lane == 0) s_work[tid >> 5] = wm;
    __syncthreads();
    if (tid == 0) {
      uint32_t m = 0;
#pragma unroll
      for (int i = 0; i < RF_WARPS; ++i) m = max(m, s_work[i]);
      tile_work[tile] = m;
    }
  }
#ifdef GSR_CTA_TRACE
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned b = blockIdx.y * gridDim.x + blockIdx.x;
    if (b < 32768) {
      g_trace_render_fwd[3 * b] = t_start; g_trace_render_fwd[3 * b + 1] = trace_now();
      g_trace_render_fwd[3 * b + 2] = ((unsigned long long)(t_start - t_enter) << 32) | trace_smid();
    }
  }
#endif
}

cudaError_t launch_render_fwd(cudaStream_t st, int W, int H, const ImageViews& im, const uint32_t* point_list,
                              const GeomViews& g, const float* bg, float* out_color, float* out_depth,
                              float* out_opacity, int* n_touched) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  dim3 grid(tiles_x, tiles_y * RF_SUB);
  // The forward renderer starts the tiles that were heaviest in the PREVIOUS forward on this image workspace first
  // (tracking / mapping iterate over nearly the same view; a tile's work is only known after it ran). A fresh
  // workspace holds arbitrary values: any content sorts into a valid permutation, so only the schedule depends on
  // it, never a result. GSR_FWD_TILE_ORDER=0 restores the linear CTA -> tile map (A/B measurements).
  static const bool fwd_order = [] { const char* e = getenv("GSR_FWD_TILE_ORDER"); return !(e && e[0] == '0'); }();
  const uint32_t* order = nullptr;
  if (RF_SUB == 1 && fwd_order && tiles_x * tiles_y > 2 * 148) {
    cudaError_t e = launch_tile_order(st, tiles_x * tiles_y, im.tile_work, im.tile_order_fwd);
    if (e != cudaSuccess) return e;
    order = im.tile_order_fwd;
  }
  count_launch(); launch_k(render_fwd_kernel, grid, RF_THREADS, 0, st, im.ranges, point_list, W, H, tiles_x, g.rec_a, g.rec_b, g.rec_c,
                                                 bg, im.final_T, im.n_contrib, out_color, out_depth, out_opacity,
                                                 n_touched, im.tile_work, order);
  return cudaGetLastError();
}

}  // namespace gsr

#ifdef GSR_CTA_TRACE
extern "C" int gsr_debug_cta_trace_render_fwd(unsigned long long* host_out, int n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, gsr::g_trace_render_fwd, sizeof(unsigned long long) * 3 * (size_t)n_ctas);
}
#endif
