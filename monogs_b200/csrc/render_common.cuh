// render_common.cuh -- what the forward and backward tile renderers share: the packed per-Gaussian record, its
// cp.async staging and the conservative (pixel rectangle, Gaussian) cull.
//
// Pixel layout.  CTA = one 16x16 tile (the reference's BLOCK_X x BLOCK_Y, config.h:15-17), warp w = an 8x4 pixel
// rectangle, lane = one pixel.  Every warp keeps its OWN queue of the staged Gaussians that can reach its pixels and
// walks it independently of the other warps of the tile.
//
// Cull.  A (rectangle, Gaussian) pair is dropped only when EVERY pixel of the rectangle fails the reference's
// own test `alpha = min(0.99, o * exp(power)) < 1/255 -> continue` (forward.cu:352-358, backward.cu:683-689), i.e.
// when the minimum over the rectangle of q = -power = (cx dx^2 + cz dy^2)/2 + cy dx dy exceeds ln(255 o).  q is a
// convex quadratic (the conic of a visible Gaussian is positive definite), so its minimum over a rectangle that
// does not contain the centre lies on a face towards the centre and is found in closed form; the comparison carries
// margins for every rounding involved (tau from preprocess: +0.1 % + 1e-3; here: 2e-6 x the largest term
// magnitude), and a conic that is not positive definite is never culled.  Skipping such pairs cannot change a
// result bit: the reference `continue`s on exactly those pixels.  Measured on the headline scene: 25 % of the pairs
// a bounding-box test lets through have no pixel above 1/255; this test removes 97 % of them
// (profiles/r02_render_kernels_ncu_full.txt).
//
// Measured and rejected (profiles/r02_subwarp_queue_variants.txt): splitting every warp into 2 or 4 sub-warps with
// their own queues over 4x4 / 4x2 pixel sub-rectangles (one warp instruction = 2-4 different Gaussians).  It raises
// lane efficiency for splats of a few pixels, but the consumed front of the headline scene's tile lists is
// dominated by splats that cover whole tiles, so the iteration count fell by only 10 % while every iteration grew
// (per-lane instead of warp-uniform addressing, 4x the cull tests): forward 0.096 vs 0.080 ms, backward 0.204 vs
// 0.168 ms.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int RENDER_WARPS = 8;                 // warps per CTA = 8x4 rectangles per tile
constexpr int RENDER_THREADS = 32 * RENDER_WARPS;
#ifndef GSR_RENDER_BATCH
#define GSR_RENDER_BATCH 256
#endif
constexpr int RENDER_BATCH = GSR_RENDER_BATCH;  // list entries staged per round (one per thread)
static_assert(RENDER_BATCH == RENDER_THREADS, "one staged record per thread");
constexpr int WARP_RECT_W = 8, WARP_RECT_H = 4;

// One staged batch of B list entries: the 48-byte records (3 x float4, see GeomViews::rec) + the Gaussian ids.
template <int B>
struct RenderStageT {
  float4 rec[3 * B];
  uint32_t id[B];
};
using RenderStage = RenderStageT<RENDER_BATCH>;

GSR_DEVINL void cp_async16(void* smem_dst, const void* gmem_src) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
GSR_DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
GSR_DEVINL void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Gather record `id` into slot t of a stage: three 16-byte LDGSTS from ONE 48-byte location (two 32-byte sectors).
template <int B>
GSR_DEVINL void stage_record(RenderStageT<B>& st, int t, uint32_t id, const float4* __restrict__ rec) {
  st.id[t] = id;
  const float4* src = rec + 3 * (size_t)id;
  cp_async16(&st.rec[3 * t + 0], src + 0);
  cp_async16(&st.rec[3 * t + 1], src + 1);
  cp_async16(&st.rec[3 * t + 2], src + 2);
}

// Per-Gaussian part of the cull.
struct CullGauss {
  float gx, gy, cx, cy, cz, tau;
  float ncy_rcx, ncy_rcz;   // -cy/cx, -cy/cz
  bool pd;                  // conic positive definite (else: never culled)
};
GSR_DEVINL CullGauss cull_gauss(const float4 q0, const float4 q1) {
  CullGauss c;
  c.gx = q0.x; c.gy = q0.y; c.tau = q0.w;
  c.cx = q1.x; c.cy = q1.y; c.cz = q1.z;
  c.pd = (c.cx > 0.0f) && (c.cz > 0.0f) && (c.cx * c.cz > c.cy * c.cy);
  float rcx, rcz;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcx) : "f"(c.cx));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcz) : "f"(c.cz));
  c.ncy_rcx = -c.cy * rcx;
  c.ncy_rcz = -c.cy * rcz;
  return c;
}
// true = some pixel of [rx0, rx1] x [ry0, ry1] (pixel-centre coordinates, inclusive) may pass the alpha test.
GSR_DEVINL bool cull_hit(const CullGauss& c, float rx0, float rx1, float ry0, float ry1) {
  // dx = gx - pixel.x ranges over [X0, X1], dy likewise (forward.cu:343-345 uses xy - pixf)
  const float X0 = c.gx - rx1, X1 = c.gx - rx0, Y0 = c.gy - ry1, Y1 = c.gy - ry0;
  const float dxc = fminf(fmaxf(0.0f, X0), X1), dyc = fminf(fmaxf(0.0f, Y0), Y1);   // the point nearest the centre
  // minimum of q on the face dx = dxc (if the centre is outside in x) and on the face dy = dyc (outside in y)
  const float dy1 = fminf(fmaxf(c.ncy_rcz * dxc, Y0), Y1);
  const float dx2 = fminf(fmaxf(c.ncy_rcx * dyc, X0), X1);
  const float qa = 0.5f * (c.cx * dxc * dxc + c.cz * dy1 * dy1) + c.cy * dxc * dy1;
  const float qb = 0.5f * (c.cx * dx2 * dx2 + c.cz * dyc * dyc) + c.cy * dx2 * dyc;
  const float inf = __int_as_float(0x7f800000);
  float q = fminf(dxc != 0.0f ? qa : inf, dyc != 0.0f ? qb : inf);
  if (dxc == 0.0f && dyc == 0.0f) q = 0.0f;                                        // centre inside the rectangle
  // rounding of the reference's fp32 `power` (and of q here) scales with the largest term magnitude in the rectangle
  const float DX = fmaxf(fabsf(X0), fabsf(X1)), DY = fmaxf(fabsf(Y0), fabsf(Y1));
  const float mag = 0.5f * (c.cx * DX * DX + c.cz * DY * DY) + fabsf(c.cy) * DX * DY;
  return !c.pd || !(q > c.tau + 2e-6f * mag);                                      // NaN anywhere -> hit
}

}  // namespace gsr
