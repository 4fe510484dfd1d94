// render_fwd.cu -- forward tile renderer (K7 of SURVEY 2.2).
//
// Replaces renderCUDA<3> (forward.cu:263-392).  One 256-thread CTA per 16x16 tile; warp = 8x4 pixels, lane = pixel
// (render_common.cuh):
//   * the tile's instance list is staged in batches of 256 through shared memory with cp.async (LDGSTS) gathers of
//     one packed 48-byte record per Gaussian, double buffered so the gather of batch b+1 overlaps the blending of
//     batch b;
//   * per chunk of 128 staged Gaussians every lane tests ONE Gaussian against the warp's pixel rectangle (exact
//     conservative test on the conic, render_common.cuh) and a ballot compacts the survivors into the warp's queue
//     in shared memory -- the blending loop then only visits Gaussians that can contribute to this warp, which
//     skips exactly (pixel, Gaussian) pairs the reference `continue`s on;
//   * a warp whose 32 pixels are all saturated stops blending, the CTA stops fetching when all 8 warps are done.
// The blended quantities follow the reference's fp32 operation order exactly (explicit fmul/ffma), so final_T,
// n_contrib and the integer n_touched are bit-identical and colour/depth/opacity agree to the last bit.
#include <cstdlib>
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace gsr {

#ifdef GSR_CTA_TRACE
// Debug build only (tools/time_kernels.py --trace): per-CTA start/end globaltimer and SM id.
__device__ unsigned long long g_trace_render_fwd[3 * 32768];
GSR_DEVINL unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
GSR_DEVINL unsigned trace_smid() { unsigned s; asm volatile("mov.u32 %0, %smid;" : "=r"(s)); return s; }
#endif

#ifndef GSR_RF_CHUNK
#define GSR_RF_CHUNK 128
#endif
constexpr int RF_CHUNK = GSR_RF_CHUNK;   // staged Gaussians queued at a time (early exit wastes at most one chunk's tests)
static_assert(RF_CHUNK % 32 == 0 && RENDER_BATCH % RF_CHUNK == 0 && RF_CHUNK <= 256, "RF_CHUNK");

#ifndef GSR_RF_MIN_CTAS
#define GSR_RF_MIN_CTAS 4
#endif
#ifndef GSR_RF_UNROLL
#define GSR_RF_UNROLL 4
#endif
constexpr int RF_UNROLL = GSR_RF_UNROLL;   // queued Gaussians blended per loop round
__global__ void __launch_bounds__(RENDER_THREADS, GSR_RF_MIN_CTAS)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                  int tiles_x, const float4* __restrict__ rec, const float* __restrict__ bg,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                  float* __restrict__ out_depth, float* __restrict__ out_opacity, int* __restrict__ n_touched,
                  uint32_t* __restrict__ tile_work, const uint32_t* __restrict__ tile_order) {
  pdl_launch_dependents();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_enter = trace_now();
#endif
  pdl_wait();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_start = trace_now();
#endif
  __shared__ __align__(16) RenderStage stage[2];
  __shared__ uint8_t s_queue[RENDER_WARPS][RF_CHUNK];

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  int tile_y = blockIdx.y, tile_x = blockIdx.x;
  if (tile_order) {  // heaviest tiles (of the previous forward on this workspace) first
    const uint32_t t = tile_order[blockIdx.y * gridDim.x + blockIdx.x];
    tile_y = (int)(t / (uint32_t)tiles_x);
    tile_x = (int)(t - (uint32_t)tile_y * (uint32_t)tiles_x);
  }
  const int tile = tile_y * tiles_x + tile_x;
  // warp rectangle: 2 columns x 4 rows of 8x4-pixel rectangles
  const int wx0 = tile_x * TILE_X + (w & 1) * WARP_RECT_W;
  const int wy0 = tile_y * TILE_Y + (w >> 1) * WARP_RECT_H;
  const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
  const bool inside = pxi < W && pyi < H;
  const float pixx = (float)pxi, pixy = (float)pyi;
  const float rx0 = (float)wx0, rx1 = (float)(wx0 + WARP_RECT_W - 1), ry0 = (float)wy0, ry1 = (float)(wy0 + WARP_RECT_H - 1);
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint8_t* const my_queue = s_queue[w];

  const uint2 range = ranges[tile];
  const int total = (int)(range.y - range.x);
  const int n_batches = (total + RENDER_BATCH - 1) / RENDER_BATCH;

  float T = 1.0f, Cr = 0.f, Cg = 0.f, Cb = 0.f, D = 0.f;
  uint32_t last_contributor = 0;
  bool done = !inside;
  bool touch_live = true;   // warp-uniform: some pixel of the warp can still satisfy T(1-alpha) > 0.5

  // prologue: gather batch 0, prefetch the id of batch 1
  uint32_t id_next = 0;
  if (n_batches > 0) {
    if (tid < total) stage_record(stage[0], tid, point_list[range.x + tid], rec);
    cp_async_commit();
    if (RENDER_BATCH + tid < total) id_next = point_list[range.x + RENDER_BATCH + tid];
  }

  for (int b = 0; b < n_batches; ++b) {
    const int cur = b & 1;
    // issue the gather of batch b+1 into the other buffer (free since the barrier ending batch b-1)
    if (b + 1 < n_batches) {
      if ((b + 1) * RENDER_BATCH + tid < total) stage_record(stage[cur ^ 1], tid, id_next, rec);
      cp_async_commit();
      const int e2 = (b + 2) * RENDER_BATCH + tid;
      if (e2 < total) id_next = point_list[range.x + e2];
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();

    const RenderStage& sg = stage[cur];
    const int count = min(RENDER_BATCH, total - b * RENDER_BATCH);
    for (int c0 = 0; c0 < count; c0 += RF_CHUNK) {
      if (__all_sync(0xffffffffu, done)) break;
      // ---- queue the chunk's Gaussians that can reach this warp's rectangle (lane-parallel conservative cull) ----
      int qn = 0;
      const int c1 = min(c0 + RF_CHUNK, count);
      for (int g0 = c0; g0 < c1; g0 += 32) {
        const int jj = g0 + lane;
        const bool valid = jj < c1;
        const float4 q0 = sg.rec[3 * (valid ? jj : c0)];
        const float4 q1 = sg.rec[3 * (valid ? jj : c0) + 1];
        const bool hit = valid && cull_hit(cull_gauss(q0, q1), rx0, rx1, ry0, ry1);
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (hit) my_queue[qn + __popc(m & lt_mask)] = (uint8_t)jj;   // batch index < RENDER_BATCH = 256
        qn += __popc(m);
      }
      __syncwarp();
      // ---- blend the queued Gaussians in list order (prefetching Gaussian i+1's record one iteration ahead was
      //      measured SLOWER here: 0.094 vs 0.081 ms; it pays in the backward kernel) ----
      // One queued Gaussian: alpha for this lane's pixel (independent of the pixel's running state) ...
      auto pair_alpha = [&](int j, float& alpha, bool& ok) {
        const float4 a = sg.rec[3 * j];
        const float4 co = sg.rec[3 * j + 1];
        const float dx = fsub(a.x, pixx), dy = fsub(a.y, pixy);
        const float power = gauss_power(dx, dy, co.x, co.y, co.z);
        alpha = fminf(fmul(co.w, expf(power)), 0.99f);
        ok = !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
      };
      // ... and its blend into the running state, in list order (forward.cu:352-371)
      auto blend = [&](int j, float alpha, bool ok) {
        const float test_T = fmul(T, fsub(1.0f, alpha));
        bool act = !done && ok;
        if (act && test_T < 0.0001f) { done = true; act = false; }
        if (act) {
          const float4 a = sg.rec[3 * j];
          const float4 c = sg.rec[3 * j + 2];
          Cr = ffma(T, fmul(alpha, c.x), Cr);
          Cg = ffma(T, fmul(alpha, c.y), Cg);
          Cb = ffma(T, fmul(alpha, c.z), Cb);
          D = ffma(T, fmul(alpha, a.z), D);
          T = test_T;
          last_contributor = (uint32_t)(b * RENDER_BATCH + j + 1);
        }
        // n_touched: pixels where this Gaussian is seen with T(1-alpha) > 0.5 (forward.cu:369-371). T only falls, so
        // once no pixel of the warp has T > 0.5 the vote can never fire again and is skipped for the rest of the tile.
        if (touch_live) {
          const uint32_t tv = __ballot_sync(0xffffffffu, act && test_T > 0.5f);
          if (tv != 0) {   // (a pixel that fired has T > 0.5: still live, no second vote needed)
            if (lane == 0) atomicAdd(&n_touched[sg.id[j]], __popc(tv));
          } else {
            touch_live = __any_sync(0xffffffffu, !done && T > 0.5f);
          }
        }
      };
      // RF_UNROLL queued Gaussians per round: their alphas are independent (several exp chains in flight), only the
      // short T / colour updates are ordered; the all-done vote and the loop bookkeeping are paid once per round
      int i = 0;
      for (; i + RF_UNROLL <= qn; i += RF_UNROLL) {
        int j[RF_UNROLL];
        float al[RF_UNROLL];
        bool ok[RF_UNROLL];
#pragma unroll
        for (int u = 0; u < RF_UNROLL; ++u) { j[u] = (int)my_queue[i + u]; pair_alpha(j[u], al[u], ok[u]); }
#pragma unroll
        for (int u = 0; u < RF_UNROLL; ++u) blend(j[u], al[u], ok[u]);
        if (__all_sync(0xffffffffu, done)) break;
      }
      if (i + RF_UNROLL > qn) {   // (not taken after the break above: there i + RF_UNROLL <= qn)
        for (; i < qn; ++i) {
          if (__all_sync(0xffffffffu, done)) break;
          const int j0 = (int)my_queue[i];
          float al0;
          bool ok0;
          pair_alpha(j0, al0, ok0);
          blend(j0, al0, ok0);
        }
      }
      __syncwarp();   // the queue is rewritten by the next chunk
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
  if (tile_work) {
    __shared__ uint32_t s_work[RENDER_WARPS];
    const uint32_t wm = __reduce_max_sync(0xffffffffu, inside ? last_contributor : 0u);
    if (lane == 0) s_work[w] = wm;
    __syncthreads();
    if (tid == 0) {
      uint32_t m = 0;
#pragma unroll
      for (int i = 0; i < RENDER_WARPS; ++i) m = max(m, s_work[i]);
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
                              float* out_opacity, int* n_touched, bool tile_order_ready) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  dim3 grid(tiles_x, tiles_y);
  // The forward renderer starts the tiles that were heaviest in the PREVIOUS forward on this image workspace first
  // (tracking / mapping iterate over nearly the same view; a tile's work is only known after it ran). A fresh
  // workspace holds arbitrary values: any content sorts into a valid permutation, so only the schedule depends on
  // it, never a result. GSR_FWD_TILE_ORDER=0 restores the linear CTA -> tile map (A/B measurements).
  static const bool fwd_order = [] { const char* e = getenv("GSR_FWD_TILE_ORDER"); return !(e && e[0] == '0'); }();
  const uint32_t* order = nullptr;
  if (fwd_order && tiles_x * tiles_y > 2 * 148) {
    if (!tile_order_ready) {   // (gsr_forward / gsr_forward_enqueue: already sorted by chunk_scan_kernel's extra CTA)
      cudaError_t e = launch_tile_order(st, tiles_x * tiles_y, im.tile_work, im.tile_order_fwd);
      if (e != cudaSuccess) return e;
    }
    order = im.tile_order_fwd;
  }
  count_launch(); launch_k(render_fwd_kernel, grid, RENDER_THREADS, 0, st, im.ranges, point_list, W, H, tiles_x, g.rec,
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
