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
// PPL = pixels per lane: 1 (warp = 8x4 pixels, 8 warps per tile) or 2 (warp = 8x8 pixels, 4 warps; a lane's two pixels
// sit 4 rows apart).  With two pixels per lane the record fetch, the queue walk, the loop bookkeeping, the x part of
// the quadratic form and the votes of a (warp, Gaussian) pair are paid once per 64 pixels; pairs that reach only one
// half of the 8x8 rectangle cost more.  Measured: neutral at 640x480 (headline 0.068 ms either way, C2 +1 %),
// render_fwd -8 % at 1200x680 (C3 / C4) and -10 % at 1920x1080 (C5) -> chosen by image size in launch_render_fwd.
template <int PPL>
__global__ void __launch_bounds__(32 * RENDER_WARPS / PPL, PPL == 1 ? GSR_RF_MIN_CTAS : 6)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                  int tiles_x, const float4* __restrict__ rec, const float* __restrict__ bg,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                  float* __restrict__ out_depth, float* __restrict__ out_opacity, int* __restrict__ n_touched,
                  uint32_t* __restrict__ tile_work, const uint32_t* __restrict__ tile_order) {
  static_assert(PPL == 1 || PPL == 2, "PPL");
  constexpr int RF_PPL = PPL;
  constexpr int RF_WARPS = RENDER_WARPS / PPL;
  constexpr int RF_THREADS = 32 * RF_WARPS;
  constexpr int RF_SLOTS = RENDER_BATCH / RF_THREADS;   // staged records per thread and batch
  pdl_launch_dependents();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_enter = trace_now();
#endif
  pdl_wait();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_start = trace_now();
#endif
  __shared__ __align__(16) RenderStage stage[2];
  __shared__ uint8_t s_queue[RF_WARPS][RF_CHUNK];

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  int tile_y = blockIdx.y, tile_x = blockIdx.x;
  if (tile_order) {  // heaviest tiles (of the previous forward on this workspace) first
    const uint32_t t = tile_order[blockIdx.y * gridDim.x + blockIdx.x];
    tile_y = (int)(t / (uint32_t)tiles_x);
    tile_x = (int)(t - (uint32_t)tile_y * (uint32_t)tiles_x);
  }
  const int tile = tile_y * tiles_x + tile_x;
  // warp rectangle: 2 columns x (4 / RF_PPL) rows of 8 x (4 RF_PPL)-pixel rectangles; a lane's pixels sit 4 rows apart
  constexpr int RECT_H = WARP_RECT_H * RF_PPL;
  const int wx0 = tile_x * TILE_X + (w & 1) * WARP_RECT_W;
  const int wy0 = tile_y * TILE_Y + (w >> 1) * RECT_H;
  const int pxi = wx0 + (lane & 7);
  int pyi[RF_PPL];
  bool inside[RF_PPL];
  float pixy[RF_PPL];
#pragma unroll
  for (int p = 0; p < RF_PPL; ++p) {
    pyi[p] = wy0 + (lane >> 3) + WARP_RECT_H * p;
    inside[p] = pxi < W && pyi[p] < H;
    pixy[p] = (float)pyi[p];
  }
  const float pixx = (float)pxi;
  const float rx0 = (float)wx0, rx1 = (float)(wx0 + WARP_RECT_W - 1), ry0 = (float)wy0, ry1 = (float)(wy0 + RECT_H - 1);
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint8_t* const my_queue = s_queue[w];

  const uint2 range = ranges[tile];
  const int total = (int)(range.y - range.x);
  const int n_batches = (total + RENDER_BATCH - 1) / RENDER_BATCH;

  float T[RF_PPL], Cr[RF_PPL], Cg[RF_PPL], Cb[RF_PPL], D[RF_PPL];
  uint32_t last_contributor[RF_PPL];
  bool done[RF_PPL];
#pragma unroll
  for (int p = 0; p < RF_PPL; ++p) {
    T[p] = 1.0f; Cr[p] = 0.f; Cg[p] = 0.f; Cb[p] = 0.f; D[p] = 0.f;
    last_contributor[p] = 0; done[p] = !inside[p];
  }
  auto all_done = [&]() {
    bool d = done[0];
#pragma unroll
    for (int p = 1; p < RF_PPL; ++p) d = d && done[p];
    return d;
  };
  bool touch_live = true;   // warp-uniform: some pixel of the warp can still satisfy T(1-alpha) > 0.5

  // prologue: gather batch 0, prefetch the ids of batch 1 (RF_SLOTS staged records per thread)
  uint32_t id_next[RF_SLOTS];
#pragma unroll
  for (int q = 0; q < RF_SLOTS; ++q) id_next[q] = 0;
  if (n_batches > 0) {
#pragma unroll
    for (int q = 0; q < RF_SLOTS; ++q) {
      const int t = tid + q * RF_THREADS;
      if (t < total) stage_record(stage[0], t, point_list[range.x + t], rec);
    }
    cp_async_commit();
#pragma unroll
    for (int q = 0; q < RF_SLOTS; ++q) {
      const int t = tid + q * RF_THREADS;
      if (RENDER_BATCH + t < total) id_next[q] = point_list[range.x + RENDER_BATCH + t];
    }
  }

  for (int b = 0; b < n_batches; ++b) {
    const int cur = b & 1;
    // issue the gather of batch b+1 into the other buffer (free since the barrier ending batch b-1)
    if (b + 1 < n_batches) {
#pragma unroll
      for (int q = 0; q < RF_SLOTS; ++q) {
        const int t = tid + q * RF_THREADS;
        if ((b + 1) * RENDER_BATCH + t < total) stage_record(stage[cur ^ 1], t, id_next[q], rec);
      }
      cp_async_commit();
#pragma unroll
      for (int q = 0; q < RF_SLOTS; ++q) {
        const int e2 = (b + 2) * RENDER_BATCH + tid + q * RF_THREADS;
        if (e2 < total) id_next[q] = point_list[range.x + e2];
      }
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();

    const RenderStage& sg = stage[cur];
    const int count = min(RENDER_BATCH, total - b * RENDER_BATCH);
    for (int c0 = 0; c0 < count; c0 += RF_CHUNK) {
      if (__all_sync(0xffffffffu, all_done())) break;
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
      // One queued Gaussian: alpha for this lane's pixels (independent of the pixels' running state) ...
      auto pair_alpha = [&](int j, float (&alpha)[RF_PPL], bool (&ok)[RF_PPL]) {
        const float4 a = sg.rec[3 * j];
        const float4 co = sg.rec[3 * j + 1];
        const float dx = fsub(a.x, pixx);
#pragma unroll
        for (int p = 0; p < RF_PPL; ++p) {
          const float dy = fsub(a.y, pixy[p]);
          const float power = gauss_power(dx, dy, co.x, co.y, co.z);
          alpha[p] = fminf(fmul(co.w, expf(power)), 0.99f);
          ok[p] = !(power > 0.0f) && !(alpha[p] < 1.0f / 255.0f);
        }
      };
      // ... and its blend into the running state, in list order (forward.cu:352-371)
      auto blend = [&](int j, const float (&alpha)[RF_PPL], const bool (&ok)[RF_PPL]) {
        bool act[RF_PPL], fire[RF_PPL];
        bool any_act = false;
#pragma unroll
        for (int p = 0; p < RF_PPL; ++p) {
          const float test_T = fmul(T[p], fsub(1.0f, alpha[p]));
          act[p] = !done[p] && ok[p];
          if (act[p] && test_T < 0.0001f) { done[p] = true; act[p] = false; }
          fire[p] = act[p] && test_T > 0.5f;
          any_act = any_act || act[p];
        }
        if (any_act) {
          const float4 a = sg.rec[3 * j];
          const float4 c = sg.rec[3 * j + 2];
#pragma unroll
          for (int p = 0; p < RF_PPL; ++p) {
            if (act[p]) {
              Cr[p] = ffma(T[p], fmul(alpha[p], c.x), Cr[p]);
              Cg[p] = ffma(T[p], fmul(alpha[p], c.y), Cg[p]);
              Cb[p] = ffma(T[p], fmul(alpha[p], c.z), Cb[p]);
              D[p] = ffma(T[p], fmul(alpha[p], a.z), D[p]);
              T[p] = fmul(T[p], fsub(1.0f, alpha[p]));
              last_contributor[p] = (uint32_t)(b * RENDER_BATCH + j + 1);
            }
          }
        }
        // n_touched: pixels where this Gaussian is seen with T(1-alpha) > 0.5 (forward.cu:369-371). T only falls, so
        // once no pixel of the warp has T > 0.5 the vote can never fire again and is skipped for the rest of the tile.
        if (touch_live) {
          uint32_t n_fire = 0;
#pragma unroll
          for (int p = 0; p < RF_PPL; ++p) n_fire += __popc(__ballot_sync(0xffffffffu, fire[p]));
          if (n_fire != 0) {   // (a pixel that fired has T > 0.5: still live, no second vote needed)
            if (lane == 0) atomicAdd(&n_touched[sg.id[j]], n_fire);
          } else {
            bool live = false;
#pragma unroll
            for (int p = 0; p < RF_PPL; ++p) live = live || (!done[p] && T[p] > 0.5f);
            touch_live = __any_sync(0xffffffffu, live);
          }
        }
      };
      // RF_UNROLL queued Gaussians per round: their alphas are independent (several exp chains in flight), only the
      // short T / colour updates are ordered; the all-done vote and the loop bookkeeping are paid once per round
      constexpr int U = RF_UNROLL / RF_PPL > 0 ? RF_UNROLL / RF_PPL : 1;
      int i = 0;
      for (; i + U <= qn; i += U) {
        int j[U];
        float al[U][RF_PPL];
        bool ok[U][RF_PPL];
#pragma unroll
        for (int u = 0; u < U; ++u) { j[u] = (int)my_queue[i + u]; pair_alpha(j[u], al[u], ok[u]); }
#pragma unroll
        for (int u = 0; u < U; ++u) blend(j[u], al[u], ok[u]);
        if (__all_sync(0xffffffffu, all_done())) break;
      }
      if (i + U > qn) {   // (not taken after the break above: there i + U <= qn)
        for (; i < qn; ++i) {
          if (__all_sync(0xffffffffu, all_done())) break;
          const int j0 = (int)my_queue[i];
          float al0[RF_PPL];
          bool ok0[RF_PPL];
          pair_alpha(j0, al0, ok0);
          blend(j0, al0, ok0);
        }
      }
      __syncwarp();   // the queue is rewritten by the next chunk
    }
    // everyone must be done reading `cur` before batch b+2 is gathered into it
    if (__syncthreads_and(all_done())) break;
  }
  cp_async_wait<0>();

  uint32_t my_work = 0;
#pragma unroll
  for (int p = 0; p < RF_PPL; ++p) {
    if (inside[p]) {
      const size_t pix = (size_t)pyi[p] * W + pxi;
      const size_t HW = (size_t)H * W;
      final_T[pix] = T[p];
      n_contrib[pix] = last_contributor[p];
      out_color[pix] = ffma(T[p], bg[0], Cr[p]);
      out_color[HW + pix] = ffma(T[p], bg[1], Cg[p]);
      out_color[2 * HW + pix] = ffma(T[p], bg[2], Cb[p]);
      out_depth[pix] = D[p];
      out_opacity[pix] = fsub(1.0f, T[p]);
      my_work = max(my_work, last_contributor[p]);
    }
  }
  // the tile's replay length (its largest n_contrib) = the backward renderer's work for this tile: it starts its
  // heaviest tiles first (tile_order_kernel, render_bwd.cu)
  if (tile_work) {
    __shared__ uint32_t s_work[RF_WARPS];
    const uint32_t wm = __reduce_max_sync(0xffffffffu, my_work);
    if (lane == 0) s_work[w] = wm;
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
  // two pixels per lane above 640x480-class images (see the kernel's header); GSR_RF_PPL=1|2 forces one (A/B runs)
  static const int forced_ppl = [] { const char* e = getenv("GSR_RF_PPL"); return e ? atoi(e) : 0; }();
  const bool two = forced_ppl ? forced_ppl == 2 : tiles_x * tiles_y > 2048;
  count_launch();
  if (two)
    launch_k(render_fwd_kernel<2>, grid, 32 * RENDER_WARPS / 2, 0, st, im.ranges, point_list, W, H, tiles_x, g.rec, bg,
             im.final_T, im.n_contrib, out_color, out_depth, out_opacity, n_touched, im.tile_work, order);
  else
    launch_k(render_fwd_kernel<1>, grid, 32 * RENDER_WARPS, 0, st, im.ranges, point_list, W, H, tiles_x, g.rec, bg,
             im.final_T, im.n_contrib, out_color, out_depth, out_opacity, n_touched, im.tile_work, order);
  return cudaGetLastError();
}

}  // namespace gsr

#ifdef GSR_CTA_TRACE
extern "C" int gsr_debug_cta_trace_render_fwd(unsigned long long* host_out, int n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, gsr::g_trace_render_fwd, sizeof(unsigned long long) * 3 * (size_t)n_ctas);
}
#endif
