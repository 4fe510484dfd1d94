// render_bwd.cu -- backward tile renderer (K8 of SURVEY 2.2).
//
// Replaces renderCUDA<3> of backward.cu:563-787 and its block-wide
// shared-memory tree reduction (render_cuda_reduce_sum, backward.cu:548-559).
// Same tile/warp decomposition and cp.async staging as the forward kernel,
// traversed back to front.  Differences from the reference's schedule (the
// math per (pixel, Gaussian) pair is the same):
//   * the traversal starts at the tile's largest n_contrib instead of the end
//     of the list (entries behind every pixel's last contributor are no-ops);
//   * no block barrier inside the per-Gaussian loop: each warp culls, replays
//     and reduces on its own;
//   * the ten partial gradients of a (warp, Gaussian) pair are reduced with a
//     12-shuffle "transpose" butterfly that leaves each total in a distinct
//     lane, and those ten lanes issue ONE warp-wide red.global.add.f32 into a
//     packed 48-byte per-Gaussian record (1-2 sectors per instruction);
//   * warps in which no pixel is touched by the Gaussian do nothing.
#include <cstdlib>
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

#ifdef GSR_CTA_TRACE
// Debug build only (tools/time_kernels.py --trace): per-CTA start/end globaltimer and SM id.
__device__ unsigned long long g_trace_render_bwd[3 * 32768];
GSR_DEVINL unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
GSR_DEVINL unsigned trace_smid() { unsigned s; asm volatile("mov.u32 %0, %smid;" : "=r"(s)); return s; }
#endif

// CTA = RB_WARPS of the tile's 8 warp rectangles (see render_fwd.cu); the back-to-front walk then starts at the
// largest n_contrib of the CTA's OWN pixels.
#ifndef GSR_RENDER_WARPS
#define GSR_RENDER_WARPS 8
#endif
#ifndef GSR_RENDER_BATCH
#define GSR_RENDER_BATCH 256
#endif
constexpr int RB_WARPS = GSR_RENDER_WARPS;
constexpr int RB_SUB = 8 / RB_WARPS;
constexpr int RB_THREADS = 32 * RB_WARPS;
constexpr int RB_BATCH = GSR_RENDER_BATCH;
constexpr int RB_PER_THREAD = RB_BATCH / RB_THREADS;
static_assert(RB_WARPS == 8 || RB_WARPS == 4 || RB_WARPS == 2, "RB_WARPS");
static_assert(RB_BATCH % RB_THREADS == 0 && RB_BATCH % 32 == 0, "RB_BATCH");

struct RbStage {
  float4 a[RB_BATCH];
  float4 b[RB_BATCH];
  float4 c[RB_BATCH];
  uint32_t id[RB_BATCH];
};

GSR_DEVINL void rb_cp_async16(void* smem_dst, const void* gmem_src) {
  uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
GSR_DEVINL void rb_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
GSR_DEVINL void rb_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Reduce 10 per-lane values over the warp with 12 shuffles. On return the
// total of value k sits in v0 of the lanes whose bits 4..1 select k:
//   k = 5*b4 + (b3 ? 3 + b2 : (b2 ? 2 : b1))   (see owner_slot()).
GSR_DEVINL float reduce10(float v[10], int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  float w[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const float keep = b4 ? v[i + 5] : v[i];
    const float send = b4 ? v[i] : v[i + 5];
    w[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  float u[3];
  {  // 5 -> (0,1,2) | (3,4,-)
    const float s0 = b3 ? w[0] : w[3], k0 = b3 ? w[3] : w[0];
    const float s1 = b3 ? w[1] : w[4], k1 = b3 ? w[4] : w[1];
    const float s2 = b3 ? w[2] : 0.0f, k2 = b3 ? 0.0f : w[2];
    u[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 8);
    u[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 8);
    u[2] = k2 + __shfl_xor_sync(0xffffffffu, s2, 8);
  }
  float t[2];
  {  // 3 -> (0,1) | (2,-)
    const float s0 = b2 ? u[0] : u[2], k0 = b2 ? u[2] : u[0];
    const float s1 = b2 ? u[1] : 0.0f, k1 = b2 ? 0.0f : u[1];
    t[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 4);
    t[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 4);
  }
  float r;
  {  // 2 -> 0 | 1
    const float s0 = b1 ? t[0] : t[1], k0 = b1 ? t[1] : t[0];
    r = k0 + __shfl_xor_sync(0xffffffffu, s0, 2);
  }
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;
}

// Which of the 10 values a lane ends up owning (-1: none / duplicate).
GSR_DEVINL int owner_slot(int lane) {
  if (lane & 1) return -1;
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  int k5;
  if (!b3) {             // group (0,1,2)
    if (!b2) k5 = b1 ? 1 : 0;   // (0,1)
    else { if (b1) return -1; k5 = 2; }
  } else {               // group (3,4,-)
    if (!b2) k5 = b1 ? 4 : 3;
    else return -1;
  }
  return (b4 ? 5 : 0) + k5;
}

// Tiles by descending replay length (counting sort on tile_work / 4, one CTA): the hardware hands CTAs out in
// linear order, so with this map the long tiles run first and the kernel's tail consists of the shortest ones
// (measured on the headline scene: the backward renderer's last CTA started at 54 % of the kernel's span and ran
// for up to half of it). Order inside a bin is arbitrary -- it only affects scheduling.
constexpr int TO_THREADS = 1024;
// zero_ptr / zero_n16: optional buffer (the backward's partial-gradient scratch) cleared by the OTHER CTAs of the
// launch while CTA 0 sorts -- one launch instead of a memset node (which also breaks the PDL chain) + a 1-CTA kernel.
__global__ void __launch_bounds__(TO_THREADS)
tile_order_kernel(int n_tiles, const uint32_t* __restrict__ tile_work, uint32_t* __restrict__ tile_order,
                  float4* __restrict__ zero_ptr, size_t zero_n16) {
  pdl_launch_dependents();
  pdl_wait();
  if (blockIdx.x > 0) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (size_t i = (size_t)(blockIdx.x - 1) * TO_THREADS + threadIdx.x; i < zero_n16;
         i += (size_t)(gridDim.x - 1) * TO_THREADS)
      zero_ptr[i] = z;
    return;
  }
  __shared__ uint32_t s_bin[TO_THREADS];
  __shared__ uint32_t s_warp[TO_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  s_bin[tid] = 0;
  __syncthreads();
  for (int t = tid; t < n_tiles; t += TO_THREADS)
    atomicAdd(&s_bin[TO_THREADS - 1 - min(tile_work[t] >> 2, (uint32_t)TO_THREADS - 1)], 1u);
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
  for (int t = tid; t < n_tiles; t += TO_THREADS) {
    const uint32_t pos = atomicAdd(&s_bin[TO_THREADS - 1 - min(tile_work[t] >> 2, (uint32_t)TO_THREADS - 1)], 1u);
    tile_order[pos] = (uint32_t)t;
  }
}

#ifndef GSR_RB_MIN_CTAS
#define GSR_RB_MIN_CTAS 1
#endif
__global__ void __launch_bounds__(RB_THREADS, GSR_RB_MIN_CTAS)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                  int tiles_x, const float* __restrict__ bg, const float4* __restrict__ rec_a,
                  const float4* __restrict__ rec_b, const float4* __restrict__ rec_c,
                  const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_depth, LossView loss,
                  float* __restrict__ packed /*[P][GRAD_STRIDE]*/, const uint32_t* __restrict__ tile_order) {
  pdl_launch_dependents();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_enter = trace_now();
#endif
  pdl_wait();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_start = trace_now();
#endif
  __shared__ __align__(16) RbStage stage[2];
  __shared__ uint32_t s_max[RB_THREADS / 32];

  const int tid = threadIdx.x, lane = tid & 31, wl = tid >> 5;
  int tile_y = blockIdx.y / RB_SUB, tile_x = blockIdx.x;
  if (RB_SUB == 1 && tile_order) {  // heaviest tiles first (tile_order_kernel)
    const uint32_t t = tile_order[blockIdx.y * gridDim.x + blockIdx.x];
    tile_y = (int)(t / (uint32_t)tiles_x);
    tile_x = (int)(t - (uint32_t)tile_y * (uint32_t)tiles_x);
  }
  const int w = (blockIdx.y % RB_SUB) * RB_WARPS + wl;   // warp rectangle within the tile
  const int tile = tile_y * tiles_x + tile_x;
  const int wx0 = tile_x * TILE_X + (w & 1) * 8;
  const int wy0 = tile_y * TILE_Y + (w >> 1) * 4;
  const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
  const bool inside = pxi < W && pyi < H;
  const float pixx = (float)pxi, pixy = (float)pyi;
  const float rx0 = (float)wx0, rx1 = (float)(wx0 + 7), ry0 = (float)wy0, ry1 = (float)(wy0 + 3);
  const size_t pix = (size_t)pyi * W + pxi;
  const size_t HW = (size_t)H * W;

  const uint2 range = ranges[tile];
  const float T_final = inside ? final_T[pix] : 0.0f;
  const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;
  float dpr = 0.f, dpg = 0.f, dpb = 0.f, dpd = 0.f;
  if (inside && loss.kind == GSR_LOSS_NONE) {
    if (dL_dpix) { dpr = dL_dpix[pix]; dpg = dL_dpix[HW + pix]; dpb = dL_dpix[2 * HW + pix]; }
    if (dL_dpix_depth) dpd = dL_dpix_depth[pix];
  } else if (inside) {
    loss_pixel_grad(loss, pix, HW, dpr, dpg, dpb, dpd);  // fused L1 loss gradient (loss.cu)
  }
  const float bg_dot_dpixel = bg[0] * dpr + bg[1] * dpg + bg[2] * dpb;

  // tile-wide start of the back-to-front walk
  uint32_t wm = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wm = max(wm, __shfl_xor_sync(0xffffffffu, wm, o));
  if (lane == 0) s_max[wl] = wm;
  __syncthreads();
  uint32_t tile_last = 0;
#pragma unroll
  for (int i = 0; i < RB_THREADS / 32; ++i) tile_last = max(tile_last, s_max[i]);
  const int total = (int)tile_last;  // list positions [0, total) are replayed, from the back
  const int n_batches = (total + RB_BATCH - 1) / RB_BATCH;

  float T = T_final;
  float acc_r = 0.f, acc_g = 0.f, acc_b = 0.f, acc_d = 0.f;
  float last_alpha = 0.f, last_r = 0.f, last_g = 0.f, last_b = 0.f, last_d = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
  const int slot = owner_slot(lane);

  // batch k covers list positions pos = total-1 - (k*RB_BATCH + t), t = 0..255 (descending)
  uint32_t id_next[RB_PER_THREAD];
#pragma unroll
  for (int k = 0; k < RB_PER_THREAD; ++k) id_next[k] = 0;
  if (n_batches > 0) {
#pragma unroll
    for (int k = 0; k < RB_PER_THREAD; ++k) {
      const int t = tid + k * RB_THREADS;
      if (t < total) {
        const uint32_t id = point_list[range.x + (total - 1 - t)];
        stage[0].id[t] = id;
        rb_cp_async16(&stage[0].a[t], rec_a + id);
        rb_cp_async16(&stage[0].b[t], rec_b + id);
        rb_cp_async16(&stage[0].c[t], rec_c + id);
      }
    }
    rb_commit();
#pragma unroll
    for (int k = 0; k < RB_PER_THREAD; ++k) {
      const int t = tid + k * RB_THREADS;
      if (RB_BATCH + t < total) id_next[k] = point_list[range.x + (total - 1 - RB_BATCH - t)];
    }
  }

  for (int b = 0; b < n_batches; ++b) {
    const int cur = b & 1;
    if (b + 1 < n_batches) {
      RbStage& nx = stage[cur ^ 1];
#pragma unroll
      for (int k = 0; k < RB_PER_THREAD; ++k) {
        const int t = tid + k * RB_THREADS;
        if ((b + 1) * RB_BATCH + t < total) {
          nx.id[t] = id_next[k];
          rb_cp_async16(&nx.a[t], rec_a + id_next[k]);
          rb_cp_async16(&nx.b[t], rec_b + id_next[k]);
          rb_cp_async16(&nx.c[t], rec_c + id_next[k]);
        }
      }
      rb_commit();
#pragma unroll
      for (int k = 0; k < RB_PER_THREAD; ++k) {
        const int e2 = (b + 2) * RB_BATCH + tid + k * RB_THREADS;
        if (e2 < total) id_next[k] = point_list[range.x + (total - 1 - e2)];
      }
      rb_wait<1>();
    } else {
      rb_wait<0>();
    }
    __syncthreads();

    const RbStage& sg = stage[cur];
    const int count = min(RB_BATCH, total - b * RB_BATCH);
    for (int g0 = 0; g0 < count; g0 += 32) {
      const int jj = g0 + lane;
      bool hit = false;
      if (jj < count) {
        const float4 a = sg.a[jj];
        const float ey = sg.c[jj].w;
        // entries behind this WARP's last contributor are no-ops for all of its pixels (the walk starts at the
        // tile's largest n_contrib, which is typically 10-20 % beyond a warp's own)
        const uint32_t posj = (uint32_t)(total - 1 - (b * RB_BATCH + jj));
        hit = (posj < wm) && (a.x + a.w >= rx0) && (a.x - a.w <= rx1) && (a.y + ey >= ry0) && (a.y - ey <= ry1);
      }
      uint32_t mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int j = g0 + __ffs(mask) - 1;
        mask &= mask - 1;
        const uint32_t pos = (uint32_t)(total - 1 - (b * RB_BATCH + j));  // 0-based list position
        const float4 a = sg.a[j];
        const float4 co = sg.b[j];
        const float dx = fsub(a.x, pixx), dy = fsub(a.y, pixy);
        const float power = gauss_power(dx, dy, co.x, co.y, co.z);
        const float G = expf(power);
        const float alpha = fminf(fmul(co.w, G), 0.99f);
        // skip rules of backward.cu:676-689
        const bool skip = !inside || pos >= last_contributor || power > 0.0f || alpha < 1.0f / 255.0f;
        if (__ballot_sync(0xffffffffu, !skip) == 0) continue;

        float v[10];
#pragma unroll
        for (int i = 0; i < 10; ++i) v[i] = 0.0f;
        if (!skip) {
          const float4 c = sg.c[j];
          // 1 / (1 - alpha): ONE reciprocal (SFU estimate + a Newton step: <= 1 ulp on [0.01, 1], no special cases), shared by the
          // transmittance recovery and the background term; the reference's two IEEE divisions
          // (backward.cu:700,742) cost ~20 instructions per (warp, Gaussian) here. Gradients are compared at 1e-4.
          const float one_m_alpha = 1.0f - alpha;
          float inv_1ma;
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_1ma) : "f"(one_m_alpha));
          inv_1ma = fmaf(inv_1ma, fmaf(-one_m_alpha, inv_1ma, 1.0f), inv_1ma);  // one Newton step
          T = T * inv_1ma;
          const float dchannel_dcolor = alpha * T;
          float dL_dalpha = 0.0f;
          acc_r = last_alpha * last_r + (1.f - last_alpha) * acc_r; last_r = c.x;
          dL_dalpha += (c.x - acc_r) * dpr;
          acc_g = last_alpha * last_g + (1.f - last_alpha) * acc_g; last_g = c.y;
          dL_dalpha += (c.y - acc_g) * dpg;
          acc_b = last_alpha * last_b + (1.f - last_alpha) * acc_b; last_b = c.z;
          dL_dalpha += (c.z - acc_b) * dpb;
          acc_d = last_alpha * last_d + (1.f - last_alpha) * acc_d; last_d = a.z;
          dL_dalpha += (a.z - acc_d) * dpd;
          dL_dalpha *= T;
          last_alpha = alpha;
          dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;

          const float dL_dG = co.w * dL_dalpha;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * co.x - gdy * co.y;
          const float dG_ddely = -gdy * co.z - gdx * co.y;
          v[0] = dL_dG * dG_ddelx * ddelx_dx;
          v[1] = dL_dG * dG_ddely * ddely_dy;
          v[2] = -0.5f * gdx * dx * dL_dG;
          v[3] = -0.5f * gdx * dy * dL_dG;
          v[4] = -0.5f * gdy * dy * dL_dG;
          v[5] = G * dL_dalpha;
          v[6] = dchannel_dcolor * dpr;
          v[7] = dchannel_dcolor * dpg;
          v[8] = dchannel_dcolor * dpb;
          v[9] = dchannel_dcolor * dpd;
        }
        const float r = reduce10(v, lane);
        if (slot >= 0) atomicAdd(packed + (size_t)sg.id[j] * GRAD_STRIDE + slot, r);
      }
    }
    __syncthreads();
  }
  rb_wait<0>();
#ifdef GSR_CTA_TRACE
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned b = blockIdx.y * gridDim.x + blockIdx.x;
    if (b < 32768) {
      g_trace_render_bwd[3 * b] = t_start; g_trace_render_bwd[3 * b + 1] = trace_now();
      g_trace_render_bwd[3 * b + 2] = ((unsigned long long)(t_start - t_enter) << 32) | trace_smid();
    }
  }
#endif
}

cudaError_t launch_tile_order(cudaStream_t st, int n_tiles, const uint32_t* tile_work, uint32_t* tile_order) {
  count_launch(); launch_k(tile_order_kernel, 1, TO_THREADS, 0, st, n_tiles, tile_work, tile_order, (float4*)nullptr,
                           (size_t)0);
  return cudaGetLastError();
}

static bool tile_order_enabled() {  // GSR_TILE_ORDER=0: linear CTA -> tile map (A/B measurements)
  static const bool on = [] { const char* e = getenv("GSR_TILE_ORDER"); return !(e && e[0] == '0'); }();
  return on;
}

cudaError_t launch_render_bwd(cudaStream_t st, int W, int H, const ImageViews& im, const uint32_t* point_list,
                              const GeomViews& g, const float* bg, const float* dL_dpix,
                              const float* dL_dpix_depth, const LossView& loss, float* packed, size_t zero_bytes) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  dim3 grid(tiles_x, tiles_y * RB_SUB);
  const uint32_t* order = nullptr;
  if (RB_SUB == 1 && tile_order_enabled() && tiles_x * tiles_y > 2 * 148) {
    const bool z = zero_bytes > 0 && zero_bytes % 16 == 0;
    count_launch(); launch_k(tile_order_kernel, z ? 1 + 2 * 148 : 1, TO_THREADS, 0, st, tiles_x * tiles_y,
                             (const uint32_t*)im.tile_work, im.tile_order, reinterpret_cast<float4*>(packed),
                             z ? zero_bytes / 16 : (size_t)0);
    if (z) zero_bytes = 0;
    order = im.tile_order;
  }
  if (zero_bytes > 0) {
    cudaError_t e = cudaMemsetAsync(packed, 0, zero_bytes, st);
    if (e != cudaSuccess) return e;
  }
  count_launch(); launch_k(render_bwd_kernel, grid, RB_THREADS, 0, st, im.ranges, point_list, W, H, tiles_x, bg, g.rec_a, g.rec_b,
                                                 g.rec_c, im.final_T, im.n_contrib, dL_dpix, dL_dpix_depth, loss, packed, order);
  return cudaGetLastError();
}

}  // namespace gsr

#ifdef GSR_CTA_TRACE
extern "C" int gsr_debug_cta_trace_render_bwd(unsigned long long* host_out, int n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, gsr::g_trace_render_bwd, sizeof(unsigned long long) * 3 * (size_t)n_ctas);
}
#endif
