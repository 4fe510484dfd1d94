// preprocess_bwd.cu -- backward per-Gaussian stage (K9 + K10 of SURVEY 2.2, fused).
//
// Replaces computeCov2DCUDA (backward.cu:150-346) and preprocessCUDA<3>
// (backward.cu:418-539) with ONE kernel, one thread per Gaussian:
//   conic grad -> cov2D grad -> cov3D grad (-> scale / quaternion grad),
//   mean grad through the covariance Jacobian, the projection and the depth,
//   SH colour grad (when shs are given), and the SE(3) pose gradient
//   dL/dtau = [rho | theta] evaluated at tau = 0, including the reference's
//   quirks: the clamped t inside skew(t) (backward.cu:179-180,278), the
//   x/y_grad_mul factors (:182-183), the 1e-7 epsilons (:210,450), the
//   denom2inv branch (:212) and proj_raw elements [0],[5],[11] only (:469-484).
// The per-Gaussian dL_dtau is optionally stored ([P,6], the reference's raw
// output) and always reduced in-kernel: fp64 block partials, then a final
// deterministic reduction -- this replaces the [P,6] round trip + torch.sum
// of diff_gaussian_rasterization/__init__.py:152.
// The input gradients come from the packed 48-byte records written by the
// backward renderer; every output array is fully written here (zeros for
// invisible Gaussians), so the caller never has to pre-zero anything.
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int PB_THREADS = 256;

__device__ const float bC0 = 0.28209479177387814f;
__device__ const float bC1 = 0.4886025119029199f;
__device__ const float bC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__device__ const float bC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

struct V3 { float x, y, z; };
GSR_DEVINL V3 v3(float x, float y, float z) { return V3{x, y, z}; }
GSR_DEVINL float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GSR_DEVINL V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GSR_DEVINL V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
GSR_DEVINL V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }

// Normalisation backward (dnormvdv, auxiliary.h:107-117).
GSR_DEVINL V3 dnorm(V3 v, V3 dv) {
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
  return v3(((sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv,
            (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv,
            (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * inv);
}

// SH colour backward (computeColorFromSH, backward.cu:21-145). Writes dL_dsh
// of this Gaussian, returns dL/dmean through the view direction.
GSR_DEVINL V3 sh_backward(int deg, int M, const float* __restrict__ sh, const uint8_t* __restrict__ clamped,
                          V3 pos, V3 campos, V3 dL_dRGB, float* __restrict__ dL_dsh) {
  V3 dir_orig = v3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
  const float len = sqrtf(dot(dir_orig, dir_orig));
  const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  dL_dRGB.x *= clamped[0] ? 0.f : 1.f;
  dL_dRGB.y *= clamped[1] ? 0.f : 1.f;
  dL_dRGB.z *= clamped[2] ? 0.f : 1.f;
  auto S = [&](int k) { return v3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]); };
  auto put = [&](int k, float wgt) {
    dL_dsh[3 * k] = wgt * dL_dRGB.x; dL_dsh[3 * k + 1] = wgt * dL_dRGB.y; dL_dsh[3 * k + 2] = wgt * dL_dRGB.z;
  };
  V3 dx = v3(0, 0, 0), dy = v3(0, 0, 0), dz = v3(0, 0, 0);
  put(0, bC0);
  if (deg > 0) {
    put(1, -bC1 * y); put(2, bC1 * z); put(3, -bC1 * x);
    dx = (-bC1) * S(3); dy = (-bC1) * S(1); dz = bC1 * S(2);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      put(4, bC2[0] * xy); put(5, bC2[1] * yz); put(6, bC2[2] * (2.f * zz - xx - yy));
      put(7, bC2[3] * xz); put(8, bC2[4] * (xx - yy));
      dx = dx + (bC2[0] * y) * S(4) + (bC2[2] * 2.f * -x) * S(6) + (bC2[3] * z) * S(7) + (bC2[4] * 2.f * x) * S(8);
      dy = dy + (bC2[0] * x) * S(4) + (bC2[1] * z) * S(5) + (bC2[2] * 2.f * -y) * S(6) + (bC2[4] * 2.f * -y) * S(8);
      dz = dz + (bC2[1] * y) * S(5) + (bC2[2] * 2.f * 2.f * z) * S(6) + (bC2[3] * x) * S(7);
      if (deg > 2) {
        put(9, bC3[0] * y * (3.f * xx - yy)); put(10, bC3[1] * xy * z);
        put(11, bC3[2] * y * (4.f * zz - xx - yy)); put(12, bC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
        put(13, bC3[4] * x * (4.f * zz - xx - yy)); put(14, bC3[5] * z * (xx - yy));
        put(15, bC3[6] * x * (xx - 3.f * yy));
        dx = dx + (bC3[0] * 3.f * 2.f * xy) * S(9) + (bC3[1] * yz) * S(10) + (bC3[2] * -2.f * xy) * S(11) +
             (bC3[3] * -3.f * 2.f * xz) * S(12) + (bC3[4] * (-3.f * xx + 4.f * zz - yy)) * S(13) +
             (bC3[5] * 2.f * xz) * S(14) + (bC3[6] * 3.f * (xx - yy)) * S(15);
        dy = dy + (bC3[0] * 3.f * (xx - yy)) * S(9) + (bC3[1] * xz) * S(10) +
             (bC3[2] * (-3.f * yy + 4.f * zz - xx)) * S(11) + (bC3[3] * -3.f * 2.f * yz) * S(12) +
             (bC3[4] * -2.f * xy) * S(13) + (bC3[5] * -2.f * yz) * S(14) + (bC3[6] * -3.f * 2.f * xy) * S(15);
        dz = dz + (bC3[1] * xy) * S(10) + (bC3[2] * 4.f * 2.f * yz) * S(11) +
             (bC3[3] * 3.f * (2.f * zz - xx - yy)) * S(12) + (bC3[4] * 4.f * 2.f * xz) * S(13) +
             (bC3[5] * (xx - yy)) * S(14);
      }
    }
  }
  for (int k = (deg + 1) * (deg + 1); k < M; ++k) put(k, 0.0f);
  const V3 dL_ddir = v3(dot(dx, dL_dRGB), dot(dy, dL_dRGB), dot(dz, dL_dRGB));
  return dnorm(dir_orig, dL_ddir);
}

#ifndef PB_BLOCKS_PER_SM
#define PB_BLOCKS_PER_SM 2
#endif
// Loads of one Gaussian, in two groups. EVERY Gaussian streams its radius and its packed 48-byte gradient record
// (52 B); only a TOUCHED Gaussian -- visible and with a non-zero record -- goes on to read its mean / scale /
// rotation (44 B) and to run the ~500-instruction gradient chain. Everything the kernel outputs is linear in the ten
// packed partials, so an all-zero record means all-zero outputs: on a map where most Gaussians are occluded or
// off-screen for a given keyframe (headline scene: 3.5 % touched) the touched ones are compacted to the front
// threads of the CTA and seven of its eight warps skip the arithmetic.
struct PbLoads {
  float4 rot;
  float m[3], s[3];
};
GSR_DEVINL PbLoads pb_issue_loads(int idx, const float* __restrict__ means3D, const float* __restrict__ scales,
                                  const float* __restrict__ rotations, bool has_cov3D) {
  PbLoads L;
  L.rot = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 3; ++k) { L.s[k] = 0.f; L.m[k] = __ldg(means3D + 3 * (size_t)idx + k); }
  if (!has_cov3D) {
#pragma unroll
    for (int k = 0; k < 3; ++k) L.s[k] = __ldg(scales + 3 * (size_t)idx + k);
    L.rot = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
  }
  return L;
}

// Kernel 1 of 2 (streaming, high occupancy): every Gaussian reads its radius and its packed record (52 B), is
// classified, gets its marks written and -- when untouched -- its all-zero outputs; the touched ones of every
// 256-Gaussian chunk are listed (chunk-local index, ascending) for kernel 2.
__global__ void __launch_bounds__(PB_THREADS)
preprocess_bwd_classify_kernel(int P, int M, const int* __restrict__ radii, const float* __restrict__ shs,
                               const uint8_t* __restrict__ touched_flag, float* __restrict__ dL_dmeans2D,
                               float* __restrict__ dL_dcolors, float* __restrict__ dL_dopacity,
                               float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D,
                               float* __restrict__ dL_dsh, float* __restrict__ dL_dscales,
                               float* __restrict__ dL_drots, float* __restrict__ dL_dtau_pg, int p_begin,
                               float* __restrict__ rows, int accumulate, uint32_t* __restrict__ touched_words,
                               uint8_t* __restrict__ vis_u8, uint16_t* __restrict__ touched_list,
                               uint32_t* __restrict__ touched_count) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_wcnt[PB_THREADS / 32];
  const int tid = threadIdx.x;
  const bool rows_touched_only = (accumulate & 2) != 0;   // rows pre-zeroed by the caller: only touched rows are written
  const bool rows_add = (accumulate & 1) != 0;
  const int n_chunks = (P - p_begin + PB_THREADS - 1) / PB_THREADS;
  for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
  const int chunk_base = p_begin + chunk * PB_THREADS;  // P = end of this launch's Gaussian range
  int n_touched = 0;
  {  // ---- phase A (thread = Gaussian of the chunk): classify, marks, zero outputs of the untouched ----
    const int idx = chunk_base + tid;
    const bool in_range = idx < P;
    int my_radius = 0;
    uint8_t flag = 0;
    if (in_range) {
      my_radius = __ldg(radii + idx);
      flag = __ldg(touched_flag + idx);   // set by the backward renderer for every record it added to (5 B per Gaussian
                                          // here instead of the 48-byte record)
    }
    const bool visible = in_range && my_radius > 0;
    const bool nonzero = visible && flag != 0;
    // with SH colours the zero-gradient Gaussians still get their dL_dsh written by the full path
    const bool touched = nonzero || (visible && shs && dL_dsh);
    const unsigned bal = __ballot_sync(0xffffffffu, touched);
    if (touched_words) {
      // marks for the sparse window exchange (rows_exchange.cu): bit i of word idx/32 = this rank holds a non-zero
      // gradient row for Gaussian idx; vis_u8 = number of this rank's keyframes that see the Gaussian.
      const unsigned nz = __ballot_sync(0xffffffffu, nonzero);
      if ((tid & 31) == 0 && in_range) {
        uint32_t* wp = touched_words + (idx >> 5);
        *wp = rows_add ? (*wp | nz) : nz;
      }
      if (in_range) vis_u8[idx] = (uint8_t)((rows_add ? vis_u8[idx] : 0) + (visible ? 1 : 0));
    }
    if ((tid & 31) == 0) s_wcnt[tid >> 5] = __popc(bal);
    __syncthreads();
    int before = 0;
#pragma unroll
    for (int w2 = 0; w2 < PB_THREADS / 32; ++w2) {
      const int c = (int)s_wcnt[w2];
      before += (w2 < (tid >> 5)) ? c : 0;
      n_touched += c;
    }
    if (touched) touched_list[(size_t)chunk_base + before + __popc(bal & ((1u << (tid & 31)) - 1u))] = (uint16_t)tid;
    if (in_range && !touched) {
      // all-zero gradients (and, for the window rows, the visibility flag)
      const size_t i3 = 3 * (size_t)idx;
      if (dL_dmeans2D) { dL_dmeans2D[i3] = 0.f; dL_dmeans2D[i3 + 1] = 0.f; dL_dmeans2D[i3 + 2] = 0.f; }
      if (dL_dcolors) { dL_dcolors[i3] = 0.f; dL_dcolors[i3 + 1] = 0.f; dL_dcolors[i3 + 2] = 0.f; }
      if (dL_dopacity) dL_dopacity[idx] = 0.f;
      if (dL_dmeans3D) { dL_dmeans3D[i3] = 0.f; dL_dmeans3D[i3 + 1] = 0.f; dL_dmeans3D[i3 + 2] = 0.f; }
      if (dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; ++k) dL_dcov3D[6 * (size_t)idx + k] = 0.f;
      }
      if (dL_dsh) for (int k = 0; k < M * 3; ++k) dL_dsh[(size_t)idx * M * 3 + k] = 0.f;
      if (dL_dscales) { dL_dscales[i3] = 0.f; dL_dscales[i3 + 1] = 0.f; dL_dscales[i3 + 2] = 0.f; }
      if (dL_drots) *reinterpret_cast<float4*>(dL_drots + 4 * (size_t)idx) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (dL_dtau_pg) {
#pragma unroll
        for (int k = 0; k < 6; ++k) dL_dtau_pg[6 * (size_t)idx + k] = 0.f;
      }
      if (rows && !rows_touched_only) {
        float4* r4 = reinterpret_cast<float4*>(rows + 16 * (size_t)idx);
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rows_add) {
          if (visible) rows[16 * (size_t)idx + 15] += 1.f;
        } else {
          r4[0] = z; r4[1] = z; r4[2] = z; r4[3] = make_float4(0.f, 0.f, 0.f, visible ? 1.f : 0.f);
        }
      }
    }
    __syncthreads();
  }
  if (tid == 0) touched_count[chunk_base / PB_THREADS] = (uint32_t)n_touched;
  }  // chunk loop
}

__global__ void __launch_bounds__(PB_THREADS, PB_BLOCKS_PER_SM)
preprocess_bwd_kernel(int P, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                      const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                      const float* __restrict__ scales, const float* __restrict__ rotations, float scale_modifier,
                      const float* __restrict__ cov3D_precomp, const float* __restrict__ viewmatrix,
                      const float* __restrict__ projmatrix, const float* __restrict__ projmatrix_raw,
                      const float* __restrict__ campos, float h_x, float h_y, float tan_fovx, float tan_fovy,
                      const float* __restrict__ packed, float* __restrict__ dL_dmeans2D,
                      float* __restrict__ dL_dcolors, float* __restrict__ dL_dopacity,
                      float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
                      float* __restrict__ dL_dscales, float* __restrict__ dL_drots,
                      float* __restrict__ dL_dtau_pg, double* __restrict__ tau_partial, int p_begin,
                      float* __restrict__ rows /*[P,16] window rows or NULL*/, int accumulate,
                      const uint16_t* __restrict__ touched_list, const uint32_t* __restrict__ touched_count) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_cam[40];
  __shared__ double s_red[PB_THREADS / 32][6];
  const int tid = threadIdx.x;
  if (tid < 16) s_cam[tid] = viewmatrix[tid];
  else if (tid < 32) s_cam[tid] = projmatrix[tid - 16];
  else if (tid == 32) s_cam[32] = projmatrix_raw[0];
  else if (tid == 33) s_cam[33] = projmatrix_raw[5];
  else if (tid == 34) s_cam[34] = projmatrix_raw[11];
  else if (tid < 38) s_cam[tid] = campos ? campos[tid - 35] : 0.f;
  __syncthreads();
  const float* v = s_cam;
  const float* p = s_cam + 16;

  // Kernel 2 of 2: the gradient chain for the touched Gaussians only. Persistent grid; WARP g of the grid walks the
  // chunks g, g + n_warps, ... and its lanes take the chunk's listed Gaussians 32 at a time; every thread keeps ONE
  // fp64 pose-gradient accumulator (fixed chunk -> warp and slot -> lane assignment: the sums are deterministic).
  const bool rows_add = (accumulate & 1) != 0;
  double tacc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  const int n_chunks = (P - p_begin + PB_THREADS - 1) / PB_THREADS;
  const int lane = tid & 31;
  const int n_warps = gridDim.x * (PB_THREADS / 32);
  for (int chunk = blockIdx.x * (PB_THREADS / 32) + (tid >> 5); chunk < n_chunks; chunk += n_warps) {
  const int chunk_base = p_begin + chunk * PB_THREADS;  // P = end of this launch's Gaussian range
  const int n_touched = (int)touched_count[chunk_base / PB_THREADS];
  for (int k0 = 0; k0 < n_touched; k0 += 32) {
  // ---- phase B (thread k < n_touched = k-th touched Gaussian of the chunk): the gradient chain ----
  float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (k0 + lane < n_touched) {
  const int idx = chunk_base + (int)touched_list[(size_t)chunk_base + k0 + lane];
  float g2x = 0.f, g2y = 0.f, gop = 0.f, gcr = 0.f, gcg = 0.f, gcb = 0.f;
  V3 dmean = v3(0, 0, 0);
  float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dscale[3] = {0.f, 0.f, 0.f};
  float drot[4] = {0.f, 0.f, 0.f, 0.f};
  // the record was read by this CTA a moment ago (L1 / L2 hit); the 44 B of parameters are read here only
  const float4 pk0 = __ldg(reinterpret_cast<const float4*>(packed + (size_t)idx * GRAD_STRIDE));
  const float4 pk1 = __ldg(reinterpret_cast<const float4*>(packed + (size_t)idx * GRAD_STRIDE + 4));
  const float4 pk2 = __ldg(reinterpret_cast<const float4*>(packed + (size_t)idx * GRAD_STRIDE + 8));
  const PbLoads L = pb_issue_loads(idx, means3D, scales, rotations, cov3D_precomp != nullptr);
  const float4 ld_rot = L.rot;
  const float ld_m[3] = {L.m[0], L.m[1], L.m[2]}, ld_s[3] = {L.s[0], L.s[1], L.s[2]};
  {
    g2x = pk0.x; g2y = pk0.y;
    const float dcx = pk0.z, dcy = pk0.w, dcz = pk1.x;
    gop = pk1.y; gcr = pk1.z; gcg = pk1.w; gcb = pk2.x;
    const float gdepth = pk2.y;

    const V3 m = v3(ld_m[0], ld_m[1], ld_m[2]);

    // ---- 3D covariance (recomputed; it is not stored by the forward) ----
    float cov[6];
    float Rg[3][3];  // glm-style [col][row] rotation, and scale s
    float s[3] = {0.f, 0.f, 0.f};
    float qr = 0.f, qx = 0.f, qy = 0.f, qz = 0.f;
    if (cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; ++k) cov[k] = cov3D_precomp[(size_t)idx * 6 + k];
    } else {
      s[0] = scale_modifier * ld_s[0];
      s[1] = scale_modifier * ld_s[1];
      s[2] = scale_modifier * ld_s[2];
      qr = ld_rot.x; qx = ld_rot.y; qy = ld_rot.z; qz = ld_rot.w;
      Rg[0][0] = 1.f - 2.f * (qy * qy + qz * qz); Rg[0][1] = 2.f * (qx * qy - qr * qz); Rg[0][2] = 2.f * (qx * qz + qr * qy);
      Rg[1][0] = 2.f * (qx * qy + qr * qz); Rg[1][1] = 1.f - 2.f * (qx * qx + qz * qz); Rg[1][2] = 2.f * (qy * qz - qr * qx);
      Rg[2][0] = 2.f * (qx * qz - qr * qy); Rg[2][1] = 2.f * (qy * qz + qr * qx); Rg[2][2] = 1.f - 2.f * (qx * qx + qy * qy);
      // M = S*R : Mg[i][j] = s[j] * Rg[i][j];  Sigma = M^T M : Sigma[i][j] = sum_k Mg[j][k]*Mg[i][k]... (symmetric)
      float Mg[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Mg[i][j] = s[j] * Rg[i][j];
      // glm: (transpose(M)*M)[i][j] = sum_k Mt[k][j] * M[i][k] = sum_k M[j][k] * M[i][k]
      auto sig = [&](int i, int j) { return Mg[j][0] * Mg[i][0] + Mg[j][1] * Mg[i][1] + Mg[j][2] * Mg[i][2]; };
      cov[0] = sig(0, 0); cov[1] = sig(0, 1); cov[2] = sig(0, 2); cov[3] = sig(1, 1); cov[4] = sig(1, 2); cov[5] = sig(2, 2);
    }

    // ---- cov2D backward (backward.cu:171-343) ----
    V3 t = v3(v[0] * m.x + v[4] * m.y + v[8] * m.z + v[12], v[1] * m.x + v[5] * m.y + v[9] * m.z + v[13],
              v[2] * m.x + v[6] * m.y + v[10] * m.z + v[14]);
    const V3 pC = t;  // unclamped camera-space point
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float J00 = h_x / t.z, J11 = h_y / t.z;
    const float J02 = -(h_x * t.x) / (t.z * t.z), J12 = -(h_y * t.y) / (t.z * t.z);
    // T rows (= rows of J'R), R[i][j] = v[4j+i]
    const float T0[3] = {v[0] * J00 + v[2] * J02, v[4] * J00 + v[6] * J02, v[8] * J00 + v[10] * J02};
    const float T1[3] = {v[1] * J11 + v[2] * J12, v[5] * J11 + v[6] * J12, v[9] * J11 + v[10] * J12};
    const float V[3][3] = {{cov[0], cov[1], cov[2]}, {cov[1], cov[3], cov[4]}, {cov[2], cov[4], cov[5]}};
    float r0[3], r1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      r0[k] = T0[0] * V[k][0] + T0[1] * V[k][1] + T0[2] * V[k][2];
      r1[k] = T1[0] * V[k][0] + T1[1] * V[k][1] + T1[2] * V[k][2];
    }
    const float a = (T0[0] * r0[0] + T0[1] * r0[1] + T0[2] * r0[2]) + 0.3f;
    const float b = T0[0] * r1[0] + T0[1] * r1[1] + T0[2] * r1[2];
    const float c = (T1[0] * r1[0] + T1[1] * r1[1] + T1[2] * r1[2]) + 0.3f;
    const float denom = a * c - b * b;
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    if (denom2inv != 0.f) {
      dL_da = denom2inv * (-c * c * dcx + 2.f * b * c * dcy + (denom - a * c) * dcz);
      dL_dc = denom2inv * (-a * a * dcz + 2.f * a * b * dcy + (denom - a * c) * dcx);
      dL_db = denom2inv * 2.f * (b * c * dcx - (denom + 2.f * b * b) * dcy + a * b * dcz);
      dcov[0] = T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc;
      dcov[3] = T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc;
      dcov[5] = T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc;
      dcov[1] = 2.f * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2.f * T1[0] * T1[1] * dL_dc;
      dcov[2] = 2.f * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2.f * T1[0] * T1[2] * dL_dc;
      dcov[4] = 2.f * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2.f * T1[1] * T1[2] * dL_dc;
    }
    float dT0[3], dT1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      dT0[k] = 2.f * r0[k] * dL_da + r1[k] * dL_db;
      dT1[k] = 2.f * r1[k] * dL_dc + r0[k] * dL_db;
    }
    const float dJ00 = v[0] * dT0[0] + v[4] * dT0[1] + v[8] * dT0[2];
    const float dJ02 = v[2] * dT0[0] + v[6] * dT0[1] + v[10] * dT0[2];
    const float dJ11 = v[1] * dT1[0] + v[5] * dT1[1] + v[9] * dT1[2];
    const float dJ12 = v[2] * dT1[0] + v[6] * dT1[1] + v[10] * dT1[2];
    const float itz = 1.f / t.z, itz2 = itz * itz, itz3 = itz2 * itz;
    const float dtx = x_grad_mul * -h_x * itz2 * dJ02;
    const float dty = y_grad_mul * -h_y * itz2 * dJ12;
    const float dtz = -h_x * itz2 * dJ00 - h_y * itz2 * dJ11 + (2.f * h_x * t.x) * itz3 * dJ02 +
                      (2.f * h_y * t.y) * itz3 * dJ12;
    // pose: d p_C / d rho = I, d p_C / d theta = -skew(t) with the CLAMPED t
    tau[0] += dtx; tau[1] += dty; tau[2] += dtz;
    tau[3] += -dty * t.z + dtz * t.y;
    tau[4] += dtx * t.z - dtz * t.x;
    tau[5] += -dtx * t.y + dty * t.x;
    // mean through t = W2C * mean (ASSIGN in the reference, backward.cu:297)
    dmean = v3(v[0] * dtx + v[1] * dty + v[2] * dtz, v[4] * dtx + v[5] * dty + v[6] * dtz,
               v[8] * dtx + v[9] * dty + v[10] * dtz);
    // rotation part of the pose gradient through W (backward.cu:299-343)
    {
      const V3 dWc1 = v3(J00 * dT0[0], J11 * dT1[0], J02 * dT0[0] + J12 * dT1[0]);
      const V3 dWc2 = v3(J00 * dT0[1], J11 * dT1[1], J02 * dT0[1] + J12 * dT1[1]);
      const V3 dWc3 = v3(J00 * dT0[2], J11 * dT1[2], J02 * dT0[2] + J12 * dT1[2]);
      const V3 c1 = v3(v[0], v[1], v[2]), c2 = v3(v[4], v[5], v[6]), c3 = v3(v[8], v[9], v[10]);
      // dot(dWc, (-skew(c)).col_i) summed over the three columns == sum_k (c_k x dWc_k)
      const V3 th = cross(c1, dWc1) + cross(c2, dWc2) + cross(c3, dWc3);
      tau[3] += th.x; tau[4] += th.y; tau[5] += th.z;
    }

    // ---- projection of the mean (backward.cu:446-512) ----
    const float mhx = p[0] * m.x + p[4] * m.y + p[8] * m.z + p[12];
    const float mhy = p[1] * m.x + p[5] * m.y + p[9] * m.z + p[13];
    const float mhw = p[3] * m.x + p[7] * m.y + p[11] * m.z + p[15];
    const float m_w = 1.0f / (mhw + 0.0000001f);
    const float mul1 = mhx * m_w * m_w, mul2 = mhy * m_w * m_w;
    dmean.x += (p[0] * m_w - p[3] * mul1) * g2x + (p[1] * m_w - p[3] * mul2) * g2y;
    dmean.y += (p[4] * m_w - p[7] * mul1) * g2x + (p[5] * m_w - p[7] * mul2) * g2y;
    dmean.z += (p[8] * m_w - p[11] * mul1) * g2x + (p[9] * m_w - p[11] * mul2) * g2y;
    {
      const float al = m_w, be = -mhx * m_w * m_w, ga = -mhy * m_w * m_w;
      const float pa = s_cam[32], pb = s_cam[33], pe = s_cam[34];
      const V3 d1 = v3(al * pa, 0.f, be * pe), d2 = v3(0.f, al * pb, ga * pe);
      const V3 d1t = cross(pC, d1), d2t = cross(pC, d2);  // (-skew(p_C))^T d = p_C x d
      tau[0] += g2x * d1.x + g2y * d2.x;
      tau[1] += g2x * d1.y + g2y * d2.y;
      tau[2] += g2x * d1.z + g2y * d2.z;
      tau[3] += g2x * d1t.x + g2y * d2t.x;
      tau[4] += g2x * d1t.y + g2y * d2t.y;
      tau[5] += g2x * d1t.z + g2y * d2t.z;
    }
    // ---- depth = p_C.z (backward.cu:514-528) ----
    dmean.x += gdepth * v[2]; dmean.y += gdepth * v[6]; dmean.z += gdepth * v[10];
    tau[2] += gdepth;
    tau[3] += gdepth * pC.y;
    tau[4] += -gdepth * pC.x;

    // ---- SH colours ----
    if (shs && dL_dsh) {
      const V3 dm = sh_backward(D, M, shs + (size_t)idx * M * 3, clamped + 3 * (size_t)idx, m,
                                v3(s_cam[35], s_cam[36], s_cam[37]), v3(gcr, gcg, gcb), dL_dsh + (size_t)idx * M * 3);
      dmean = dmean + dm;
      tau[0] -= dm.x; tau[1] -= dm.y; tau[2] -= dm.z;
    }

    // ---- scale / rotation (computeCov3D backward, backward.cu:350-413) ----
    if (!cov3D_precomp) {
      float Mg[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Mg[i][j] = s[j] * Rg[i][j];
      const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                              {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                              {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
      // dL_dM = 2 * M * dL_dSigma (glm product: (A*B)[i][j] = sum_k A[k][j] * B[i][k])
      float dM[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) dM[i][j] = 2.f * (Mg[0][j] * dS[i][0] + Mg[1][j] * dS[i][1] + Mg[2][j] * dS[i][2]);
      // dL_dMt[a][b] = dM[b][a];  Rt[a][b] = Rg[b][a]
      dscale[0] = Rg[0][0] * dM[0][0] + Rg[1][0] * dM[1][0] + Rg[2][0] * dM[2][0];
      dscale[1] = Rg[0][1] * dM[0][1] + Rg[1][1] * dM[1][1] + Rg[2][1] * dM[2][1];
      dscale[2] = Rg[0][2] * dM[0][2] + Rg[1][2] * dM[1][2] + Rg[2][2] * dM[2][2];
      float Mt[3][3];
#pragma unroll
      for (int a2 = 0; a2 < 3; ++a2)
#pragma unroll
        for (int b2 = 0; b2 < 3; ++b2) Mt[a2][b2] = dM[b2][a2] * s[a2];
      drot[0] = 2.f * qz * (Mt[0][1] - Mt[1][0]) + 2.f * qy * (Mt[2][0] - Mt[0][2]) + 2.f * qx * (Mt[1][2] - Mt[2][1]);
      drot[1] = 2.f * qy * (Mt[1][0] + Mt[0][1]) + 2.f * qz * (Mt[2][0] + Mt[0][2]) + 2.f * qr * (Mt[1][2] - Mt[2][1]) -
                4.f * qx * (Mt[2][2] + Mt[1][1]);
      drot[2] = 2.f * qx * (Mt[1][0] + Mt[0][1]) + 2.f * qr * (Mt[2][0] - Mt[0][2]) + 2.f * qz * (Mt[1][2] + Mt[2][1]) -
                4.f * qy * (Mt[2][2] + Mt[0][0]);
      drot[3] = 2.f * qr * (Mt[0][1] - Mt[1][0]) + 2.f * qx * (Mt[2][0] + Mt[0][2]) + 2.f * qy * (Mt[1][2] + Mt[2][1]) -
                4.f * qz * (Mt[1][1] + Mt[0][0]);
      // the reference's dL_dscale is w.r.t. the UNmodified scale only through s = mod*scale in M;
      // it stores dot(Rt, dL_dMt) without the modifier (backward.cu:394-397) -- same here.
    }
  }
  {
    const size_t i3 = 3 * (size_t)idx;
    if (dL_dmeans2D) { dL_dmeans2D[i3] = g2x; dL_dmeans2D[i3 + 1] = g2y; dL_dmeans2D[i3 + 2] = 0.f; }
    if (dL_dcolors) { dL_dcolors[i3] = gcr; dL_dcolors[i3 + 1] = gcg; dL_dcolors[i3 + 2] = gcb; }
    if (dL_dopacity) dL_dopacity[idx] = gop;
    if (dL_dmeans3D) { dL_dmeans3D[i3] = dmean.x; dL_dmeans3D[i3 + 1] = dmean.y; dL_dmeans3D[i3 + 2] = dmean.z; }
    if (dL_dcov3D) {
#pragma unroll
      for (int k = 0; k < 6; ++k) dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
    }
    if (dL_dscales) { dL_dscales[i3] = dscale[0]; dL_dscales[i3 + 1] = dscale[1]; dL_dscales[i3 + 2] = dscale[2]; }
    if (dL_drots) *reinterpret_cast<float4*>(dL_drots + 4 * (size_t)idx) = make_float4(drot[0], drot[1], drot[2], drot[3]);
    if (dL_dtau_pg) {
#pragma unroll
      for (int k = 0; k < 6; ++k) dL_dtau_pg[6 * (size_t)idx + k] = tau[k];
    }
    if (rows) {
      // mapping-window row (monogs_b200/window.py): xyz 3 | rgb 3 | opacity 1 | scale 3 | rotation 4 |
      // ||dL/dmean2D[:2]|| | visible -- one 64-byte row per Gaussian, so a Gaussian range is one contiguous slice
      float4* r4 = reinterpret_cast<float4*>(rows + 16 * (size_t)idx);
      float4 o0 = make_float4(dmean.x, dmean.y, dmean.z, gcr);
      float4 o1 = make_float4(gcg, gcb, gop, dscale[0]);
      float4 o2 = make_float4(dscale[1], dscale[2], drot[0], drot[1]);
      float4 o3 = make_float4(drot[2], drot[3], sqrtf(g2x * g2x + g2y * g2y), 1.f);
      if (rows_add) {
        const float4 a0 = r4[0], a1 = r4[1], a2 = r4[2], a3 = r4[3];
        o0.x += a0.x; o0.y += a0.y; o0.z += a0.z; o0.w += a0.w;
        o1.x += a1.x; o1.y += a1.y; o1.z += a1.z; o1.w += a1.w;
        o2.x += a2.x; o2.y += a2.y; o2.z += a2.z; o2.w += a2.w;
        o3.x += a3.x; o3.y += a3.y; o3.z += a3.z; o3.w += a3.w;
      }
      r4[0] = o0; r4[1] = o1; r4[2] = o2; r4[3] = o3;
    }
  }
  }  // touched
#pragma unroll
  for (int k = 0; k < 6; ++k) tacc[k] += (double)tau[k];
  }  // 32 listed Gaussians
  }  // chunk loop

  // ---- block reduction of dL/dtau in fp64 ----
  double td[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    td[k] = tacc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) td[k] += __shfl_xor_sync(0xffffffffu, td[k], o);
  }
  if ((tid & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) s_red[tid >> 5][k] = td[k];
  }
  __syncthreads();
  if (tid < 6) {
    double sacc = 0.0;
#pragma unroll
    for (int w2 = 0; w2 < PB_THREADS / 32; ++w2) sacc += s_red[w2][tid];
    tau_partial[((size_t)(p_begin / PB_THREADS) + blockIdx.x) * 6 + tid] = sacc;
  }
}

// Deterministic final reduction of the block partials: one block, fixed order.
__global__ void __launch_bounds__(256) tau_final_kernel(int n_blocks, const double* __restrict__ partial,
                                                        float* __restrict__ dL_dtau) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ double s[8][6];
  const int tid = threadIdx.x;
  double acc[6] = {0, 0, 0, 0, 0, 0};
  for (int b = tid; b < n_blocks; b += 256)
#pragma unroll
    for (int k = 0; k < 6; ++k) acc[k] += partial[(size_t)b * 6 + k];
#pragma unroll
  for (int k = 0; k < 6; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
  if ((tid & 31) == 0)
#pragma unroll
    for (int k = 0; k < 6; ++k) s[tid >> 5][k] = acc[k];
  __syncthreads();
  if (tid < 6) {
    double t = 0;
    for (int w = 0; w < 8; ++w) t += s[w][tid];
    dL_dtau[tid] = (float)t;
  }
}

cudaError_t launch_preprocess_bwd(cudaStream_t st, int P, int D, int M, const float* means3D, const int* radii,
                                  const float* shs, const uint8_t* clamped, const float* scales,
                                  const float* rotations, float scale_modifier, const float* cov3D_precomp,
                                  const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw,
                                  const float* campos, int W, int H, float tan_fovx, float tan_fovy,
                                  const BackwardViews& bw, float* dL_dmeans2D, float* dL_dcolors,
                                  float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                                  float* dL_dscales, float* dL_drots, float* dL_dtau_pg, float* dL_dtau,
                                  int p_begin, int p_end, float* rows, int accumulate, uint32_t* touched_words,
                                  uint8_t* vis_u8) {
  const float focal_y = H / (2.0f * tan_fovy);
  const float focal_x = W / (2.0f * tan_fovx);
  // p_begin is a multiple of 256 (checked by the caller); persistent grid of 148 SMs x PB_BLOCKS_PER_SM CTAs. Partial slot of CTA b
  // = p_begin / 256 + b: distinct over the ranges of one backward; slots no CTA writes stay zero (memset with packed)
  const int n_chunks = num_pre_blocks(p_end - p_begin);
  int nb = (n_chunks + PB_THREADS / 32 - 1) / (PB_THREADS / 32);   // kernel 2: one warp per chunk and round
  if (nb > 148 * PB_BLOCKS_PER_SM) nb = 148 * PB_BLOCKS_PER_SM;
  if (n_chunks > 0) {
    const int nb1 = n_chunks < 148 * 8 ? n_chunks : 148 * 8;
    count_launch();
    launch_k(preprocess_bwd_classify_kernel, nb1, PB_THREADS, 0, st, p_end, M, radii, shs, bw.touched_flag, dL_dmeans2D,
             dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drots, dL_dtau_pg, p_begin, rows,
             accumulate, touched_words, vis_u8, bw.touched_list, bw.touched_count);
    count_launch();
    launch_k(preprocess_bwd_kernel, nb, PB_THREADS, 0, st, p_end, D, M, means3D, radii, shs, clamped, scales, rotations,
                                                     scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                                                     projmatrix_raw, campos, focal_x, focal_y, tan_fovx, tan_fovy,
                                                     bw.packed, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D,
                                                     dL_dcov3D, dL_dsh, dL_dscales, dL_drots, dL_dtau_pg,
                                                     bw.tau_partial, p_begin, rows, accumulate, bw.touched_list,
                                                     bw.touched_count);
  }
  // the pose gradient is complete once the LAST range has run (ranges are issued in ascending order)
  if (dL_dtau && p_end == P) {
    // slots above the last range's are never written (ranges ascend), the others are written or still zero
    count_launch(); launch_k(tau_final_kernel, 1, 256, 0, st, p_begin / PB_THREADS + nb, bw.tau_partial, dL_dtau);
  }
  return cudaGetLastError();
}

}  // namespace gsr
