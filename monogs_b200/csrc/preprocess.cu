// preprocess.cu -- forward per-Gaussian stage (K1, K2, K3 of SURVEY 2.2).
//
// Replaces preprocessCUDA (forward.cu:157-258), checkFrustum
// (rasterizer_impl.cu:54-66) and the cub::DeviceScan::InclusiveSum call
// (rasterizer_impl.cu:280).  One thread per Gaussian; inputs are staged
// through shared memory so every global load is a fully coalesced 128-bit
// access, outputs are three float4 records (SoA of float4).
//
// Besides the reference's outputs this kernel derives, at per-GAUSSIAN cost,
// everything the instance-level stages need:
//   * depth-digit histograms for the first four radix passes (weighted by the
//     number of tiles the Gaussian touches),
//   * a difference array over tile ids (+1 at the start of each covered tile
//     row, -1 after its end) whose prefix sum is the per-tile instance count:
//     that gives the tile ranges (identifyTileRanges, rasterizer_impl.cu:116-138)
//     and the tile-digit histograms without ever reading the instance keys.
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

// SH basis constants (auxiliary.h:22-39).
__device__ const float kC0 = 0.28209479177387814f;
__device__ const float kC1 = 0.4886025119029199f;
__device__ const float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

// SH -> RGB (forward.cu:22-73). Returns unclamped+0.5 colour; caller clamps.
GSR_DEVINL float3 sh_to_rgb(int deg, int M, const float* __restrict__ sh /* [M][3] of this Gaussian */,
                            float3 pos, float3 campos) {
  float3 dir = make_float3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
  // glm: dir / length(dir) is a true division; keep it one.
  float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
  float x = dir.x / len, y = dir.y / len, z = dir.z / len;
  float3 r;
  auto S = [&](int k, int c) { return sh[3 * k + c]; };
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float v = kC0 * S(0, c);
    if (deg > 0) {
      v = v - kC1 * y * S(1, c) + kC1 * z * S(2, c) - kC1 * x * S(3, c);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        v = v + kC2[0] * xy * S(4, c) + kC2[1] * yz * S(5, c) + kC2[2] * (2.0f * zz - xx - yy) * S(6, c) +
            kC2[3] * xz * S(7, c) + kC2[4] * (xx - yy) * S(8, c);
        if (deg > 2) {
          v = v + kC3[0] * y * (3.0f * xx - yy) * S(9, c) + kC3[1] * xy * z * S(10, c) +
              kC3[2] * y * (4.0f * zz - xx - yy) * S(11, c) +
              kC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12, c) +
              kC3[4] * x * (4.0f * zz - xx - yy) * S(13, c) + kC3[5] * z * (xx - yy) * S(14, c) +
              kC3[6] * x * (xx - 3.0f * yy) * S(15, c);
        }
      }
    }
    v += 0.5f;
    (&r.x)[c] = v;
  }
  return r;
}

// 3D covariance from scale + quaternion, operation-for-operation what nvcc
// emits for computeCov3D (forward.cu:120-154).  q = (r,x,y,z), NOT re-normalised.
GSR_DEVINL void cov3d_from_scale_rot(float3 s_in, float mod, float4 q, float* cov) {
  const float sx = fmul(mod, s_in.x), sy = fmul(mod, s_in.y), sz = fmul(mod, s_in.z);
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  const float yy = fmul(y, y), zz = fmul(z, z);
  const float rz = fmul(r, z), xz = fmul(x, z), rx = fmul(r, x);
  const float yy_zz = fadd(yy, zz);
  const float xx_zz = ffma(x, x, zz);
  const float xx_yy = ffma(x, x, yy);
  auto twice = [](float v) { return fadd(v, v); };
  const float R00 = fsub(1.0f, twice(yy_zz));
  // ptxas fuses one product of each off-diagonal sum into an FFMA (read from the SASS of the
  // reference build, not its PTX: plain mul.f32/add.f32 may still be contracted by ptxas).
  const float a01 = twice(ffma(x, y, -rz));
  const float a02 = twice(ffma(r, y, xz));
  const float a10 = twice(ffma(x, y, rz));
  const float R11 = fsub(1.0f, twice(xx_zz));
  const float a12 = twice(ffma(y, z, -rx));
  const float a20 = twice(ffma(-r, y, xz));
  const float a21 = twice(ffma(y, z, rx));
  const float R22 = fsub(1.0f, twice(xx_yy));
  // M = S * R (glm column-major); the products with the zero entries of S
  // vanish for finite inputs.
  const float m0 = fmul(sx, R00), m1 = fmul(sy, a01), m2 = fmul(sz, a02);
  const float m3 = fmul(sx, a10), m4 = fmul(sy, R11), m5 = fmul(sz, a12);
  const float m6 = fmul(sx, a20), m7 = fmul(sy, a21), m8 = fmul(sz, R22);
  cov[0] = ffma(m2, m2, ffma(m0, m0, fmul(m1, m1)));
  cov[1] = ffma(m5, m2, ffma(m3, m0, fmul(m4, m1)));
  cov[2] = ffma(m8, m2, ffma(m6, m0, fmul(m7, m1)));
  cov[3] = ffma(m5, m5, ffma(m3, m3, fmul(m4, m4)));
  cov[4] = ffma(m8, m5, ffma(m6, m3, fmul(m7, m4)));
  cov[5] = ffma(m8, m8, ffma(m6, m6, fmul(m7, m7)));
}

// EWA projection of the 3D covariance (computeCov2D, forward.cu:76-115),
// same operation order as the reference's generated code. v = viewmatrix.
GSR_DEVINL float3 cov2d_project(float tx, float ty, float tz, float focal_x, float focal_y, float tan_fovx,
                                float tan_fovy, const float* cov, const float* v) {
  const float limx = fmul(tan_fovx, 1.3f), limy = fmul(tan_fovy, 1.3f);
  const float txtz = fdiv(tx, tz), tytz = fdiv(ty, tz);
  const float cx = fminf(limx, fmaxf(-limx, txtz));
  const float cy = fminf(limy, fmaxf(-limy, tytz));
  const float ntz = -tz;
  const float tz2 = fmul(tz, tz);
  const float J00 = fdiv(focal_x, tz);
  const float J02 = fdiv(fmul(focal_x, fmul(cx, ntz)), tz2);
  const float J11 = fdiv(focal_y, tz);
  const float J12 = fdiv(fmul(focal_y, fmul(cy, ntz)), tz2);
  const float T00 = ffma(v[2], J02, fmul(v[0], J00));
  const float T01 = ffma(v[6], J02, fmul(v[4], J00));
  const float T02 = ffma(J02, v[10], fmul(v[8], J00));
  const float T10 = ffma(v[2], J12, fmul(J11, v[1]));
  const float T11 = ffma(v[6], J12, fmul(J11, v[5]));
  const float T12 = ffma(J12, v[10], fmul(J11, v[9]));
  const float A0 = ffma(T02, cov[2], ffma(T00, cov[0], fmul(T01, cov[1])));
  const float B0 = ffma(T12, cov[2], ffma(T10, cov[0], fmul(T11, cov[1])));
  const float A1 = ffma(T02, cov[4], ffma(T00, cov[1], fmul(T01, cov[3])));
  const float B1 = ffma(T12, cov[4], ffma(T10, cov[1], fmul(T11, cov[3])));
  const float A2 = ffma(T02, cov[5], ffma(T00, cov[2], fmul(T01, cov[4])));
  const float B2 = ffma(T12, cov[5], ffma(T10, cov[2], fmul(T11, cov[4])));
  const float c00 = ffma(T02, A2, ffma(T00, A0, fmul(T01, A1)));
  const float c01 = ffma(T02, B2, ffma(T00, B0, fmul(T01, B1)));
  const float c11 = ffma(T12, B2, ffma(T10, B0, fmul(T11, B1)));
  return make_float3(fadd(c00, 0.3f), c01, fadd(c11, 0.3f));
}

// ndc2Pix (auxiliary.h:41-44): evaluated in double because of the 1.0 / 0.5
// literals; nvcc contracts (v+1)*S-1 into one double fma.
GSR_DEVINL float ndc_to_pix(float v, int S) {
  double t = __fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0);
  return (float)__dmul_rn(t, 0.5);
}

// getRect (auxiliary.h:46-56).
GSR_DEVINL void tile_rect(float px, float py, int radius, int tiles_x, int tiles_y, uint2& rmin, uint2& rmax) {
  const float r = (float)radius;
  const float inv = 0.0625f;  // division by BLOCK_X = 16 is exact as a multiply
  int x0 = (int)fmul(fsub(px, r), inv);
  int y0 = (int)fmul(fsub(py, r), inv);
  int x1 = (int)fmul(fadd(fadd(fadd(px, r), 16.0f), -1.0f), inv);
  int y1 = (int)fmul(fadd(fadd(fadd(py, r), 16.0f), -1.0f), inv);
  rmin.x = min((unsigned)tiles_x, (unsigned)max(0, x0));
  rmin.y = min((unsigned)tiles_y, (unsigned)max(0, y0));
  rmax.x = min((unsigned)tiles_x, (unsigned)max(0, x1));
  rmax.y = min((unsigned)tiles_y, (unsigned)max(0, y1));
}

constexpr int PRE_THREADS = 256;

// Coalesced 128-bit load of the block's PRE_THREADS * 3 consecutive floats ([P,3] AoS input) -- the first 192
// threads fetch one float4 each; the value is parked in a register so that all staging loads of the kernel are in
// flight together, and stored to shared memory afterwards (stage_store3).
GSR_DEVINL float4 stage_load3(const float* __restrict__ g, long long base_elem, long long total_elem) {
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  const int t = threadIdx.x;
  if (t < PRE_THREADS * 3 / 4) {
    const long long e = base_elem + 4ll * t;  // base_elem is a multiple of 768 floats -> 16-byte aligned
    if (e + 4 <= total_elem) v = __ldg(reinterpret_cast<const float4*>(g + e));
    else {
      if (e < total_elem) v.x = __ldg(g + e);
      if (e + 1 < total_elem) v.y = __ldg(g + e + 1);
      if (e + 2 < total_elem) v.z = __ldg(g + e + 2);
    }
  }
  return v;
}
GSR_DEVINL void stage_store3(float* s, float4 v) {
  if (threadIdx.x < PRE_THREADS * 3 / 4) reinterpret_cast<float4*>(s)[threadIdx.x] = v;
}

// One 256-Gaussian chunk's global input loads of a thread, parked in registers: the NEXT chunk's loads are in flight
// while the current chunk is computed.
struct PreLoads {
  float4 mean, scale, rot;
  float opac;
  float3 col;
};
GSR_DEVINL PreLoads pre_issue_loads(int chunk, int P, const float* __restrict__ means3D,
                                    const float* __restrict__ colors_precomp, const float* __restrict__ opacities,
                                    const float* __restrict__ scales, const float* __restrict__ rotations,
                                    bool has_cov3D) {
  PreLoads L;
  const long long blk_base = (long long)chunk * PRE_THREADS;
  const int idx = (int)(blk_base + threadIdx.x);
  L.mean = stage_load3(means3D, blk_base * 3, (long long)P * 3);
  L.scale = make_float4(0.f, 0.f, 0.f, 0.f);
  L.rot = make_float4(1.f, 0.f, 0.f, 0.f);
  if (!has_cov3D) {
    L.scale = stage_load3(scales, blk_base * 3, (long long)P * 3);
    if (idx < P) L.rot = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
  }
  L.opac = idx < P ? __ldg(opacities + idx) : 0.f;
  L.col = make_float3(0.f, 0.f, 0.f);
  if (colors_precomp != nullptr && idx < P)
    L.col = make_float3(__ldg(colors_precomp + (size_t)idx * 3), __ldg(colors_precomp + (size_t)idx * 3 + 1),
                        __ldg(colors_precomp + (size_t)idx * 3 + 2));
  return L;
}

constexpr int PRE_BLOCKS_PER_SM = 4;  // persistent grid: 148 SMs x 4 CTAs walk the chunks

__global__ void __launch_bounds__(PRE_THREADS, PRE_BLOCKS_PER_SM)
preprocess_fwd_kernel(int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ shs,
                      const float* __restrict__ colors_precomp, const float* __restrict__ opacities,
                      const float* __restrict__ scales, float scale_modifier,
                      const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                      const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                      const float* __restrict__ campos, int W, int H, float tan_fovx, float tan_fovy,
                      float focal_x, float focal_y, int tiles_x, int tiles_y, int prefiltered,
                      int* __restrict__ radii, GeomViews g) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float s_mean[PRE_THREADS * 3];
  __shared__ __align__(16) float s_scale[PRE_THREADS * 3];
  __shared__ float s_cam[36];
  __shared__ uint32_t s_hist[4 * 256];

  const int tid = threadIdx.x;
  const int n_chunks = num_pre_blocks(P);

  if (tid < 16) s_cam[tid] = viewmatrix[tid];
  else if (tid < 32) s_cam[tid] = projmatrix[tid - 16];
  else if (tid < 35) s_cam[tid] = campos ? campos[tid - 32] : 0.0f;
  for (int i = tid; i < 4 * 256; i += PRE_THREADS) s_hist[i] = 0;
  PreLoads cur = pre_issue_loads(blockIdx.x, P, means3D, colors_precomp, opacities, scales, rotations,
                                 cov3D_precomp != nullptr);

  for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
  const int idx = chunk * PRE_THREADS + tid;
  stage_store3(s_mean, cur.mean);
  stage_store3(s_scale, cur.scale);
  const float4 my_rot = cur.rot;
  const float my_opac = cur.opac;
  const float3 my_col = cur.col;
  if (chunk + (int)gridDim.x < n_chunks)  // the next chunk's loads fly during this chunk's arithmetic
    cur = pre_issue_loads(chunk + gridDim.x, P, means3D, colors_precomp, opacities, scales, rotations,
                          cov3D_precomp != nullptr);
  __syncthreads();

  uint32_t my_tiles = 0;
  uint2 my_rect = make_uint2(0u, 0u);  // x0 | x1 << 16, y0 | y1 << 16 (empty for culled Gaussians)
  uint32_t my_dkey = 0u;
  if (idx < P) {
    const float* view = s_cam;
    const float* proj = s_cam + 16;
    const float px = s_mean[3 * tid], py = s_mean[3 * tid + 1], pz = s_mean[3 * tid + 2];
    int my_radius_i = 0;
    float4 ra = make_float4(0.f, 0.f, 0.f, -1e30f), rb = make_float4(0.f, 0.f, 0.f, 0.f),
           rc = make_float4(0.f, 0.f, 0.f, 0.f);

    // near culling (in_frustum, auxiliary.h:139-164): p_view.z <= 0.2 -> invisible
    const float depth = affine_row(view, 2, px, py, pz);
    bool alive = !(depth <= 0.2f);  // NaN depth stays alive, like the reference test
    if (!alive && prefiltered) atomicExch(g.error_flag, 1);

    if (alive) {
      const float hx = affine_row(proj, 0, px, py, pz);
      const float hy = affine_row(proj, 1, px, py, pz);
      const float hw = affine_row(proj, 3, px, py, pz);
      const float p_w = frcp(fadd(hw, 0.0000001f));
      const float projx = fmul(hx, p_w), projy = fmul(hy, p_w);

      float cov[6];
      if (cov3D_precomp != nullptr) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cov[k] = __ldg(cov3D_precomp + (size_t)idx * 6 + k);
      } else {
        cov3d_from_scale_rot(make_float3(s_scale[3 * tid], s_scale[3 * tid + 1], s_scale[3 * tid + 2]),
                             scale_modifier, my_rot, cov);
      }
      const float tx = affine_row(view, 0, px, py, pz);
      const float ty = affine_row(view, 1, px, py, pz);
      const float3 c2 = cov2d_project(tx, ty, depth, focal_x, focal_y, tan_fovx, tan_fovy, cov, view);

      const float det = ffma(c2.x, c2.z, -fmul(c2.y, c2.y));  // FFMA(a, c, -(b*b)) in the reference SASS
      if (det == 0.0f) alive = false;
      if (alive) {
        const float det_inv = frcp(det);
        const float conx = fmul(c2.z, det_inv), cony = fmul(det_inv, -c2.y), conz = fmul(c2.x, det_inv);
        const float mid = fmul(fadd(c2.x, c2.z), 0.5f);
        const float sq = fsqrt(fmaxf(ffma(mid, mid, -det), 0.1f));
        const float lam = fmaxf(fadd(mid, sq), fsub(mid, sq));
        const float rad_f = ceilf(fmul(fsqrt(lam), 3.0f));
        my_radius_i = (int)rad_f;
        const float pix_x = ndc_to_pix(projx, W), pix_y = ndc_to_pix(projy, H);
        uint2 rmin, rmax;
        tile_rect(pix_x, pix_y, my_radius_i, tiles_x, tiles_y, rmin, rmax);
        my_tiles = (rmax.y - rmin.y) * (rmax.x - rmin.x);
        if (my_tiles == 0) {
          alive = false;
          my_radius_i = 0;
        } else {
          my_rect = make_uint2(rmin.x | (rmax.x << 16), rmin.y | (rmax.y << 16));
          const float opac = my_opac;
          // Cull threshold of the renderers (render_common.cuh): a pixel can pass the reference's alpha >= 1/255
          // test only where q = -power <= ln(255 o). Margins: 0.1 % + 1e-3 cover expf / fmul rounding of alpha and
          // logf here; the renderers add a term for the rounding of `power` itself. NaN opacity -> NaN -> never culled.
          const float lg = logf(255.0f * opac);
          const float tau = lg + 1e-3f * fabsf(lg) + 1e-3f;
          float3 rgb;
          if (colors_precomp != nullptr) {
            rgb = my_col;
          } else {
            rgb = sh_to_rgb(D, M, shs + (size_t)idx * M * 3, make_float3(px, py, pz),
                            make_float3(s_cam[32], s_cam[33], s_cam[34]));
            g.clamped[3 * (size_t)idx + 0] = rgb.x < 0.f;
            g.clamped[3 * (size_t)idx + 1] = rgb.y < 0.f;
            g.clamped[3 * (size_t)idx + 2] = rgb.z < 0.f;
            rgb.x = fmaxf(rgb.x, 0.f); rgb.y = fmaxf(rgb.y, 0.f); rgb.z = fmaxf(rgb.z, 0.f);
          }
          ra = make_float4(pix_x, pix_y, depth, tau);
          rb = make_float4(conx, cony, conz, opac);
          rc = make_float4(rgb.x, rgb.y, rgb.z, 0.0f);

        }
      }
    }
    radii[idx] = my_radius_i;
    g.tiles_touched[idx] = my_tiles;
    g.rect[idx] = my_rect;
    g.rec[3 * (size_t)idx + 0] = ra;
    g.rec[3 * (size_t)idx + 1] = rb;
    g.rec[3 * (size_t)idx + 2] = rc;
    // depth sort key: positive floats order like their bit patterns; culled Gaussians go last
    const uint32_t dkey = my_tiles ? __float_as_uint(ra.z) : 0xFFFFFFFFu;
    g.dkeys_a[idx] = dkey;
    my_dkey = dkey;
    g.order_a[idx] = (uint32_t)idx;
    atomicAdd(&s_hist[dkey & 255u], 1u);
    atomicAdd(&s_hist[256 + ((dkey >> 8) & 255u)], 1u);
  }
  {  // the two high depth digits take few distinct values per warp: one shared atomic per distinct value
    const uint32_t hi = idx < P ? (my_dkey >> 16) : 0x10000u;
    const uint32_t peers = __match_any_sync(0xffffffffu, hi);
    if (idx < P && (__ffs(peers) - 1) == (tid & 31)) {
      const uint32_t c = __popc(peers);
      atomicAdd(&s_hist[512 + (hi & 255u)], c);
      atomicAdd(&s_hist[768 + (hi >> 8)], c);
    }
  }
  __syncthreads();  // s_mean / s_scale are rewritten by the next chunk; last chunk: s_hist complete
  }  // chunk loop
  for (int i = tid; i < 4 * 256; i += PRE_THREADS) {
    uint32_t c = s_hist[i];
    if (c) atomicAdd(&g.hist[i], c);
  }
}

constexpr int SCAN_THREADS = 1024;
static_assert(SCAN_THREADS == TILE_ORDER_THREADS, "chunk_scan_kernel's extra CTA runs tile_order_body");

GSR_DEVINL uint32_t block_exclusive_scan_1024(uint32_t v, uint32_t* s_warp, uint32_t& block_total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t x = s_warp[lane];
    uint32_t xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, xi, o);
      if (lane >= o) xi += t;
    }
    s_warp[lane] = xi - x;
    if (lane == 31) s_warp[32] = xi;
  }
  __syncthreads();
  block_total = s_warp[32];
  return s_warp[w] + inc - v;
}

// chunk sums -> exclusive chunk offsets (in place) + total = num_rendered (cub InclusiveSum, rasterizer_impl.cu:280).
// One block; 4 consecutive chunks per thread and round (one 16-byte load), so 1M Gaussians are a single round.
__global__ void __launch_bounds__(SCAN_THREADS)
chunk_scan_kernel(int n_chunks, uint32_t* __restrict__ chunk_sums, unsigned long long* __restrict__ total,
                  uint32_t* __restrict__ chunk_sums_c, unsigned long long* __restrict__ total_c, int n_tiles,
                  const uint32_t* __restrict__ tile_work, uint32_t* __restrict__ tile_order) {
  pdl_launch_dependents();
  pdl_wait();
  if (blockIdx.x == 2) {   // optional CTA 2: the forward renderer's tile order (independent work, saves a launch)
    tile_order_body(n_tiles, tile_work, tile_order);
    return;
  }
  if (blockIdx.x == 1) { chunk_sums = chunk_sums_c; total = total_c; }  // CTA 1: the bin-level counts
  __shared__ uint32_t s_warp[33];
  unsigned long long carry = 0;
  for (int base = 0; base < n_chunks; base += SCAN_THREADS * 4) {
    const int i0 = base + threadIdx.x * 4;
    uint32_t v[4] = {0u, 0u, 0u, 0u};
    if (i0 + 4 <= n_chunks) {
      const uint4 x = *reinterpret_cast<const uint4*>(chunk_sums + i0);
      v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) if (i0 + j < n_chunks) v[j] = chunk_sums[i0 + j];
    }
    uint32_t tot;
    const uint32_t ex = block_exclusive_scan_1024(v[0] + v[1] + v[2] + v[3], s_warp, tot);
    unsigned long long run = carry + ex;
    if (i0 + 4 <= n_chunks) {
      uint4 o;
      o.x = (uint32_t)run; o.y = (uint32_t)(run + v[0]); o.z = (uint32_t)(run + v[0] + v[1]);
      o.w = (uint32_t)(run + v[0] + v[1] + v[2]);
      *reinterpret_cast<uint4*>(chunk_sums + i0) = o;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (i0 + j < n_chunks) chunk_sums[i0 + j] = (uint32_t)run;
        run += v[j];
      }
    }
    carry += tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) { *total = carry; chunk_sums[n_chunks] = (uint32_t)carry; }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                                    uint8_t* __restrict__ present) {
  pdl_launch_dependents();
  pdl_wait();
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float x = means3D[3 * (size_t)idx], y = means3D[3 * (size_t)idx + 1], z = means3D[3 * (size_t)idx + 2];
  present[idx] = !(affine_row(viewmatrix, 2, x, y, z) <= 0.2f);
}

// Camera matrices of one render() call (gsr_camera_setup): 16 threads, thread t = element (row i, column j).
__global__ void camera_setup_kernel(const float* __restrict__ R, int rs, int cs, const float* __restrict__ T, int ts,
                                    const float* __restrict__ proj_T, float* __restrict__ viewmatrix,
                                    float* __restrict__ projmatrix, float* __restrict__ campos) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float v[16];
  const int t = threadIdx.x, i = t >> 2, j = t & 3;
  // viewmatrix = W2C^T with W2C = [R T; 0 0 0 1]: v[i][j] = W2C[j][i]
  float e;
  if (i < 3) e = j < 3 ? R[j * rs + i * cs] : 0.0f;
  else e = j < 3 ? T[j * ts] : 1.0f;
  v[t] = e;
  viewmatrix[t] = e;
  __syncthreads();
  // projmatrix = viewmatrix x proj_T, k = 0..3 in order, one rounding per step
  float acc = __fmul_rn(v[i * 4], proj_T[j]);
#pragma unroll
  for (int k = 1; k < 4; ++k) acc = __fmaf_rn(v[i * 4 + k], proj_T[k * 4 + j], acc);
  projmatrix[t] = acc;
  if (t < 3) {
    // campos = -R^-1 T by the adjugate (R^-1 = adj(R) / det R)
    const float a = R[0], b = R[cs], c = R[2 * cs], d = R[rs], f = R[rs + cs], g = R[rs + 2 * cs], h = R[2 * rs],
                k2 = R[2 * rs + cs], m = R[2 * rs + 2 * cs];
    const float det = a * (f * m - g * k2) - b * (d * m - g * h) + c * (d * k2 - f * h);
    float r0, r1, r2;  // row t of adj(R)
    if (t == 0) { r0 = f * m - g * k2; r1 = c * k2 - b * m; r2 = b * g - c * f; }
    else if (t == 1) { r0 = g * h - d * m; r1 = a * m - c * h; r2 = c * d - a * g; }
    else { r0 = d * k2 - f * h; r1 = b * h - a * k2; r2 = a * f - b * d; }
    campos[t] = -(r0 * T[0] + r1 * T[ts] + r2 * T[2 * ts]) / det;
  }
}

// ---------------------------------------------------------------------------
cudaError_t launch_preprocess_fwd(cudaStream_t st, int P, int D, int M, const float* means3D, const float* shs,
                                  const float* colors_precomp, const float* opacities, const float* scales,
                                  float scale_modifier, const float* rotations, const float* cov3D_precomp,
                                  const float* viewmatrix, const float* projmatrix, const float* campos, int W,
                                  int H, float tan_fovx, float tan_fovy, int prefiltered, int* radii,
                                  const GeomViews& g) {
  const float focal_y = H / (2.0f * tan_fovy);
  const float focal_x = W / (2.0f * tan_fovx);
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  const int grid = num_pre_blocks(P) < 148 * PRE_BLOCKS_PER_SM ? num_pre_blocks(P) : 148 * PRE_BLOCKS_PER_SM;
  count_launch(); launch_k(preprocess_fwd_kernel, grid, PRE_THREADS, 0, st, 
      P, D, M, means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp,
      viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy, focal_x, focal_y, tiles_x, tiles_y, prefiltered,
      radii, g);
  return launch_digit_offsets(st, g.hist, 4);
}

// offsets of the depth-ordered chunks + total; `order` = Gaussian ids in depth order
cudaError_t launch_chunk_scan(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g, int n_tiles,
                              const uint32_t* tile_work, uint32_t* tile_order_fwd) {
  const int nc = num_pre_blocks(P);
  cudaError_t e = launch_gather_rects(st, P, order, g);
  if (e != cudaSuccess) return e;
  count_launch(); launch_k(chunk_scan_kernel, tile_order_fwd ? 3 : 2, SCAN_THREADS, 0, st, nc, g.chunk_offsets, g.total, g.chunk_offsets_c,
           g.total_c, n_tiles, tile_work, tile_order_fwd);
  return cudaGetLastError();
}

cudaError_t launch_camera_setup(cudaStream_t st, const float* R, int rs, int cs, const float* T, int ts,
                                const float* proj_T, float* viewmatrix, float* projmatrix, float* campos) {
  count_launch(); launch_k(camera_setup_kernel, 1, 16, 0, st, R, rs, cs, T, ts, proj_T, viewmatrix, projmatrix, campos);
  return cudaGetLastError();
}

cudaError_t launch_mark_visible(cudaStream_t st, int P, const float* means3D, const float* viewmatrix,
                                uint8_t* present) {
  count_launch(); launch_k(mark_visible_kernel, (P + 255) / 256, 256, 0, st, P, means3D, viewmatrix, present);
  return cudaGetLastError();
}

}  // namespace gsr
