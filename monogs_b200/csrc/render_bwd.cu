// render_bwd.cu -- backward tile renderer (K8 of SURVEY 2.2).
//
// Replaces renderCUDA<3> of backward.cu:563-787 and its block-wide shared-memory tree reduction
// (render_cuda_reduce_sum, backward.cu:548-559).  Same tile / warp decomposition, packed-record staging and
// per-warp queue as the forward kernel (render_common.cuh), traversed back to front.  The math per
// (pixel, Gaussian) pair is the reference's; the schedule is not:
//   * the traversal starts at the tile's largest n_contrib instead of the end of the list, and a warp only queues
//     entries in front of ITS OWN last contributor (entries behind it are no-ops for all of its pixels);
//   * no block barrier inside the per-Gaussian loop: each warp culls, replays and reduces on its own;
//   * TWO-PHASE TRANSPOSED REDUCTION.  Every gradient a pixel contributes to a Gaussian is a fixed polynomial in
//     the pixel offset times one of TWO per-pair scalars: S = G dL/dalpha (-> mean2D, conic, opacity) and A = alpha T
//     (-> colour, depth; times the pixel's constant dL/dpixel).  Phase 1 (lane = pixel) replays 16 queued Gaussians
//     and parks S and A in shared memory; phase 2 (lane = Gaussian x pixel half) re-reads them transposed and sums
//     its Gaussian's moments sum S, S dx, S dy, S dx^2, S dx dy, S dy^2 and sum A dL/dpixel over the warp's 32 pixels
//     in REGISTERS -- no cross-lane reduction of ten values per pair (round 1: a 12-shuffle butterfly + ~40 selects
//     and adds per pair, a third of the kernel's instructions), only one 16-lane exchange per 16 Gaussians -- and
//     adds them to the packed 48-byte gradient record with ten red.global.add.f32 per Gaussian;
//   * warps in which no pixel is touched by the Gaussian never see it (exact cull).
#include <cstdlib>
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace gsr {

#ifdef GSR_CTA_TRACE
// Debug build only (tools/time_kernels.py --trace): per-CTA start/end globaltimer and SM id.
__device__ unsigned long long g_trace_render_bwd[3 * 32768];
GSR_DEVINL unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
GSR_DEVINL unsigned trace_smid() { unsigned s; asm volatile("mov.u32 %0, %smid;" : "=r"(s)); return s; }
#endif

#ifndef GSR_RB_UNROLL
#define GSR_RB_UNROLL 8
#endif
constexpr int RB_UNROLL = GSR_RB_UNROLL;   // unroll of the phase-2 pixel loop (register budget: 3 CTAs per SM)
#ifndef GSR_RB_P1
#define GSR_RB_P1 2
#endif
constexpr int RB_P1 = GSR_RB_P1;          // queued Gaussians replayed per phase-1 loop round
constexpr int RB_GROUP = 16;              // queued Gaussians per transposed-reduction round
constexpr int RB_PITCH = 33;              // floats per Gaussian row of the S / A tiles: 32 pixels + 1 (conflict-free both ways)

// Tiles by descending replay length (counting sort on tile_work / 4, one CTA): the hardware hands CTAs out in
// linear order, so with this map the long tiles run first and the kernel's tail consists of the shortest ones
// (measured on the headline scene: the backward renderer's last CTA started at 54 % of the kernel's span and ran
// for up to half of it). Order inside a bin is arbitrary -- it only affects scheduling.
constexpr int TO_THREADS = TILE_ORDER_THREADS;
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
  tile_order_body(n_tiles, tile_work, tile_order);
}

// Dynamic shared memory of the backward renderer (65 KB: above the 48 KB static limit; 3 CTAs per SM fit).
#ifndef GSR_RB_BATCH
#define GSR_RB_BATCH 256
#endif
#ifndef GSR_RB_TILE_ACC
#define GSR_RB_TILE_ACC 1
#endif
constexpr int GRAD_ACC = 10;              // values accumulated per Gaussian (the first 10 of the packed record)
constexpr int RB_BATCH = GSR_RB_BATCH;    // list entries staged per round (<= one per thread; queue indices are u8)
static_assert(RB_BATCH <= RENDER_THREADS && RB_BATCH <= 256 && RB_BATCH % 32 == 0, "RB_BATCH");
using RbStage = RenderStageT<RB_BATCH>;
struct RbShared {
  RbStage stage[2];
  float4 dp[RENDER_WARPS][32];                 // dL/d(r, g, b, depth) of every warp's pixels
  float S[RENDER_WARPS][RB_GROUP * RB_PITCH];  // [Gaussian of the round][pixel]
  float A[RENDER_WARPS][RB_GROUP * RB_PITCH];
  uint8_t queue[RENDER_WARPS][RB_BATCH];
#if GSR_RB_TILE_ACC
  // the tile's partial gradients of the staged batch: the 8 warps add their (warp, Gaussian) sums here (shared-memory
  // atomics) and ONE global atomic per value and (tile, Gaussian) goes out when the batch is done -- the reference's
  // count (one per tile after its block reduction, backward.cu:773-786): an eighth of the global atomics and, with
  // them, of the fp32 accumulation-order noise on cancellation-dominated gradients (run-to-run spread of dL/drot on
  // the anisotropic layered 2 M scene 3-10e-5 -> the reference's 1-2e-5).  (max_contrib is only used before the
  // first batch and shares the space.)
  float acc[RB_BATCH][GRAD_ACC];
#else
  uint32_t max_contrib_[RENDER_WARPS];
#endif
};

#ifndef GSR_RB_MIN_CTAS
#define GSR_RB_MIN_CTAS 3
#endif
// 3 CTAs per SM: 3 x (dynamic + 1 KB reserved) <= 228 KB
static_assert(GSR_RB_MIN_CTAS != 3 || sizeof(RbShared) <= 76800, "RbShared no longer fits three times into an SM");
__global__ void __launch_bounds__(RENDER_THREADS, GSR_RB_MIN_CTAS)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                  int tiles_x, const float* __restrict__ bg, const float4* __restrict__ rec,
                  const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_depth, LossView loss,
                  float* __restrict__ packed /*[P][GRAD_STRIDE]*/, uint8_t* __restrict__ touched_flag /*[P]*/,
                  const uint32_t* __restrict__ tile_order) {
  pdl_launch_dependents();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_enter = trace_now();
#endif
  pdl_wait();
#ifdef GSR_CTA_TRACE
  const unsigned long long t_start = trace_now();
#endif
  extern __shared__ __align__(16) unsigned char rb_smem_raw[];
  RbShared& sh = *reinterpret_cast<RbShared*>(rb_smem_raw);
  RbStage* const stage = sh.stage;
#if GSR_RB_TILE_ACC
  uint32_t* const s_max = reinterpret_cast<uint32_t*>(&sh.acc[0][0]);
#else
  uint32_t* const s_max = sh.max_contrib_;
#endif

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  int tile_y = blockIdx.y, tile_x = blockIdx.x;
  if (tile_order) {  // heaviest tiles first (tile_order_kernel)
    const uint32_t t = tile_order[blockIdx.y * gridDim.x + blockIdx.x];
    tile_y = (int)(t / (uint32_t)tiles_x);
    tile_x = (int)(t - (uint32_t)tile_y * (uint32_t)tiles_x);
  }
  const int tile = tile_y * tiles_x + tile_x;
  const int wx0 = tile_x * TILE_X + (w & 1) * WARP_RECT_W;
  const int wy0 = tile_y * TILE_Y + (w >> 1) * WARP_RECT_H;
  const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
  const bool inside = pxi < W && pyi < H;
  const float pixx = (float)pxi, pixy = (float)pyi;
  const float rx0 = (float)wx0, rx1 = (float)(wx0 + WARP_RECT_W - 1), ry0 = (float)wy0, ry1 = (float)(wy0 + WARP_RECT_H - 1);
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint8_t* const my_queue = sh.queue[w];
  float* const my_S = sh.S[w];
  float* const my_A = sh.A[w];
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
  sh.dp[w][lane] = make_float4(dpr, dpg, dpb, dpd);

  // last contributor of the warp (bounds its queue) and of the tile (start of the walk)
  const uint32_t wm = __reduce_max_sync(0xffffffffu, last_contributor);
  if (lane == 0) s_max[w] = wm;
  __syncthreads();
  uint32_t tile_last = 0;
#pragma unroll
  for (int i = 0; i < RENDER_WARPS; ++i) tile_last = max(tile_last, s_max[i]);
#if GSR_RB_TILE_ACC
  __syncthreads();   // s_max read by everyone before the accumulators (same space) are cleared
  if (tid < RB_BATCH) {
#pragma unroll
    for (int c = 0; c < GRAD_ACC; ++c) sh.acc[tid][c] = 0.0f;
  }
  // (ordered before the first accumulation by the barrier that follows the first batch's gather)
#endif
  const int total = (int)tile_last;  // list positions [0, total) are replayed, from the back
  const int n_batches = (total + RB_BATCH - 1) / RB_BATCH;

  float T = T_final;
  float acc_w = 0.f, last_w = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
  // phase-2 role of this lane: Gaussian g2 of the round, pixels [16 h2, 16 h2 + 16) of the warp's rectangle
  const int g2 = lane & (RB_GROUP - 1), h2 = lane >> 4;
  const float py2 = (float)(wy0 + 2 * h2);

  // batch k covers list positions pos = total-1 - (k*RB_BATCH + t), t = 0..255 (descending)
  uint32_t id_next = 0;
  if (n_batches > 0) {
    if (tid < RB_BATCH && tid < total) stage_record(stage[0], tid, point_list[range.x + (total - 1 - tid)], rec);
    cp_async_commit();
    if (tid < RB_BATCH && RB_BATCH + tid < total) id_next = point_list[range.x + (total - 1 - RB_BATCH - tid)];
  }

  for (int b = 0; b < n_batches; ++b) {
    const int cur = b & 1;
    if (b + 1 < n_batches) {
      if (tid < RB_BATCH && (b + 1) * RB_BATCH + tid < total) stage_record(stage[cur ^ 1], tid, id_next, rec);
      cp_async_commit();
      const int e2 = (b + 2) * RB_BATCH + tid;
      if (tid < RB_BATCH && e2 < total) id_next = point_list[range.x + (total - 1 - e2)];
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();

    const RbStage& sg = stage[cur];
    const int count = min(RB_BATCH, total - b * RB_BATCH);
    // ---- queue the batch's Gaussians that can reach this warp's rectangle and lie in front of its last
    //      contributor (the walk starts at the TILE's largest n_contrib) ----
    int qn = 0;
    if ((uint32_t)(total - b * RB_BATCH - count) < wm) {   // else: the whole batch lies behind this warp's pixels
      for (int g0 = 0; g0 < count; g0 += 32) {
        const int jj = g0 + lane;
        const bool valid = jj < count;
        const float4 q0 = sg.rec[3 * (valid ? jj : 0)];
        const float4 q1 = sg.rec[3 * (valid ? jj : 0) + 1];
        const uint32_t posj = (uint32_t)(total - 1 - (b * RB_BATCH + jj));
        const bool hit = valid && posj < wm && cull_hit(cull_gauss(q0, q1), rx0, rx1, ry0, ry1);
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (hit) my_queue[qn + __popc(m & lt_mask)] = (uint8_t)jj;
        qn += __popc(m);
      }
    }
    __syncwarp();

    for (int i0 = 0; i0 < qn; i0 += RB_GROUP) {
      const int gn = min(RB_GROUP, qn - i0);
      // ---- phase 1 (lane = pixel): replay gn Gaussians, park S = G dL/dalpha and A = alpha T per (Gaussian, pixel).
      //      RB_P1 queued Gaussians per round: their records, exp chains and colour dot products are independent
      //      (issued together), only the short transmittance / accumulator recurrences are ordered.
      auto p1_front = [&](int q, float& G, float& alpha, float& wk, bool& skip) {
        const int j = (int)my_queue[q];
        const float4 a = sg.rec[3 * j], co = sg.rec[3 * j + 1], c = sg.rec[3 * j + 2];
        const uint32_t pos = (uint32_t)(total - 1 - (b * RB_BATCH + j));  // 0-based list position
        const float dx = fsub(a.x, pixx), dy = fsub(a.y, pixy);
        const float power = gauss_power(dx, dy, co.x, co.y, co.z);
        G = expf(power);
        alpha = fminf(fmul(co.w, G), 0.99f);
        // skip rules of backward.cu:676-689
        skip = !inside || pos >= last_contributor || power > 0.0f || alpha < 1.0f / 255.0f;
        // w = colour . dL/dpixel (+ depth): dL/dalpha = T sum_ch (c_ch - accum_rec_ch) dL/dpixel_ch
        // (backward.cu:707-731); the per-channel recurrence accum_rec = last_alpha last_c + (1 - last_alpha) accum_rec
        // is linear, so it is carried for this ONE scalar instead of the four channels.
        wk = fmaf(c.x, dpr, fmaf(c.y, dpg, fmaf(c.z, dpb, a.z * dpd)));
      };
      auto p1_back = [&](int k, float G, float alpha, float wk, bool skip) {
        float S = 0.0f, A = 0.0f;
        if (!skip) {
          // 1 / (1 - alpha): ONE reciprocal (SFU estimate + a Newton step: <= 1 ulp on [0.01, 1], no special cases),
          // shared by the transmittance recovery and the background term; the reference's two IEEE divisions
          // (backward.cu:700,742) cost ~20 instructions per pair here. Gradients are compared at 1e-4.
          const float one_m_alpha = 1.0f - alpha;
          float inv_1ma;
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_1ma) : "f"(one_m_alpha));
          inv_1ma = fmaf(inv_1ma, fmaf(-one_m_alpha, inv_1ma, 1.0f), inv_1ma);  // one Newton step
          T = T * inv_1ma;
          A = alpha * T;                                   // dchannel_dcolor (backward.cu:702)
          acc_w = fmaf(last_alpha, last_w - acc_w, acc_w);
          last_w = wk;
          last_alpha = alpha;
          const float dL_dalpha = fmaf(wk - acc_w, T, (-T_final * inv_1ma) * bg_dot_dpixel);
          S = G * dL_dalpha;
        }
        my_S[k * RB_PITCH + lane] = S;
        my_A[k * RB_PITCH + lane] = A;
      };
      int k = 0;
      for (; k + RB_P1 <= gn; k += RB_P1) {
        float G[RB_P1], al[RB_P1], wk[RB_P1];
        bool sk[RB_P1];
#pragma unroll
        for (int u = 0; u < RB_P1; ++u) p1_front(i0 + k + u, G[u], al[u], wk[u], sk[u]);
#pragma unroll
        for (int u = 0; u < RB_P1; ++u) p1_back(k + u, G[u], al[u], wk[u], sk[u]);
      }
      for (; k < gn; ++k) {
        float G, al, wk;
        bool sk;
        p1_front(i0 + k, G, al, wk, sk);
        p1_back(k, G, al, wk, sk);
      }
      __syncwarp();
      // ---- phase 2 (lane = Gaussian g2 x pixel half h2): moments of S and A dL/dpixel over 16 pixels in registers
      {
        const bool have = g2 < gn;
        const int jg = (int)my_queue[i0 + (have ? g2 : 0)];
        const float4 ag = sg.rec[3 * jg];
        const float4 cog = sg.rec[3 * jg + 1];
        const float* rowS = my_S + g2 * RB_PITCH + 16 * h2;
        const float* rowA = my_A + g2 * RB_PITCH + 16 * h2;
        const float4* dp = sh.dp[w] + 16 * h2;
        float m0 = 0.f, mx = 0.f, my = 0.f, mxx = 0.f, mxy = 0.f, myy = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, cd = 0.f;
        if (have) {
#pragma unroll RB_UNROLL
          for (int k = 0; k < 16; ++k) {
            const float S = rowS[k], A = rowA[k];
            const float4 d = dp[k];
            // pixel k of this half: x = wx0 + (k & 7), y = wy0 + 2 h2 + (k >> 3); same fp32 offsets as phase 1
            const float dx = fsub(ag.x, (float)(wx0 + (k & 7))), dy = fsub(ag.y, py2 + (float)(k >> 3));
            const float sx = S * dx, sy = S * dy;
            m0 += S; mx += sx; my += sy;
            mxx = fmaf(sx, dx, mxx); mxy = fmaf(sx, dy, mxy); myy = fmaf(sy, dy, myy);
            cr = fmaf(A, d.x, cr); cg = fmaf(A, d.y, cg); cb = fmaf(A, d.z, cb); cd = fmaf(A, d.w, cd);
          }
        }
        // the two pixel halves of a Gaussian sit 16 lanes apart
        m0 += __shfl_xor_sync(0xffffffffu, m0, 16); mx += __shfl_xor_sync(0xffffffffu, mx, 16);
        my += __shfl_xor_sync(0xffffffffu, my, 16); mxx += __shfl_xor_sync(0xffffffffu, mxx, 16);
        mxy += __shfl_xor_sync(0xffffffffu, mxy, 16); myy += __shfl_xor_sync(0xffffffffu, myy, 16);
        cr += __shfl_xor_sync(0xffffffffu, cr, 16); cg += __shfl_xor_sync(0xffffffffu, cg, 16);
        cb += __shfl_xor_sync(0xffffffffu, cb, 16); cd += __shfl_xor_sync(0xffffffffu, cd, 16);
        if (have && h2 == 0) {
          // dL/dG = o dL/dalpha; dG/ddelx = -G (dx cx + dy cy), dG/ddely = -G (dy cz + dx cy)   (backward.cu:746-771)
          const float o = cog.w;
#if GSR_RB_TILE_ACC
          float* dst = sh.acc[jg];
#else
          float* dst = packed + (size_t)sg.id[jg] * GRAD_STRIDE;
          touched_flag[sg.id[jg]] = 1;
#endif
          atomicAdd(dst + 0, -o * (cog.x * mx + cog.y * my) * ddelx_dx);
          atomicAdd(dst + 1, -o * (cog.z * my + cog.y * mx) * ddely_dy);
          atomicAdd(dst + 2, -0.5f * o * mxx);
          atomicAdd(dst + 3, -0.5f * o * mxy);
          atomicAdd(dst + 4, -0.5f * o * myy);
          atomicAdd(dst + 5, m0);
          atomicAdd(dst + 6, cr);
          atomicAdd(dst + 7, cg);
          atomicAdd(dst + 8, cb);
          atomicAdd(dst + 9, cd);
        }
      }
      __syncwarp();   // the S / A tiles are rewritten by the next round
    }
    __syncthreads();
#if GSR_RB_TILE_ACC
    // flush the batch: thread t owns staged entry t (it also re-gathers slot t next, so sg.id[t] is still valid here);
    // the next batch's accumulation starts behind the barrier at the top of the loop
    if (tid < count) {
      float v[GRAD_ACC];
      bool any = false;
#pragma unroll
      for (int c = 0; c < GRAD_ACC; ++c) { v[c] = sh.acc[tid][c]; any |= v[c] != 0.0f; }
      if (any) {
        const uint32_t id = sg.id[tid];
        float* dst = packed + (size_t)id * GRAD_STRIDE;
#pragma unroll
        for (int c = 0; c < GRAD_ACC; ++c) { atomicAdd(dst + c, v[c]); sh.acc[tid][c] = 0.0f; }
        touched_flag[id] = 1;   // the per-Gaussian backward reads this byte instead of the 48-byte record to classify
      }
    }
#endif
  }
  cp_async_wait<0>();
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
                              const float* dL_dpix_depth, const LossView& loss, float* packed, uint8_t* touched_flag,
                              size_t zero_bytes) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  dim3 grid(tiles_x, tiles_y);
  const uint32_t* order = nullptr;
  if (tile_order_enabled() && tiles_x * tiles_y > 2 * 148) {
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
  static const cudaError_t smem_attr = cudaFuncSetAttribute(render_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                            (int)sizeof(RbShared));
  if (smem_attr != cudaSuccess) return smem_attr;
  count_launch(); launch_k(render_bwd_kernel, grid, RENDER_THREADS, sizeof(RbShared), st, im.ranges, point_list, W, H, tiles_x, bg, g.rec,
                                                 im.final_T, im.n_contrib, dL_dpix, dL_dpix_depth, loss, packed, touched_flag, order);
  return cudaGetLastError();
}

}  // namespace gsr

#ifdef GSR_CTA_TRACE
extern "C" int gsr_debug_cta_trace_render_bwd(unsigned long long* host_out, int n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, gsr::g_trace_render_bwd, sizeof(unsigned long long) * 3 * (size_t)n_ctas);
}
#endif
