// stats.cu -- the per-Gaussian bookkeeping on the OUTPUT side of the rasterizer, fused:
//   densify_stats_kernel   Mapper's per-keyframe statistics update (utils/slam_mapper.py:452-460,
//                          gaussian_splatting/scene/gaussian_model.py:888-892): for the visible Gaussians
//                          max_radii_2d = max(max_radii_2d, radii); xyz_gradient_accum += ||dL/dmean2D.xy||; denom += 1
//                          -- one pass instead of 2 boolean-mask gathers (each a nonzero() + host sync), a norm, 3
//                          scatter-updates per keyframe.
//   visibility_prune_kernel the window's observation count and prune mask (slam_mapper.py:399-440): nr_obs = number of
//                          keyframes with n_touched > 0, to_prune = nr_obs <= coviz && kf_idx >= min_kf -- on the
//                          device, where the reference moves every visibility vector to the CPU (`nr_obs`, `kf_idx`
//                          live there, gaussian_model.py:51-52,639-640).
// Both are streaming passes: 4-byte reads/writes per Gaussian per array, bound by HBM.
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ grad2d /*[P,3]*/,
                     float* __restrict__ max_radii_2d, float* __restrict__ grad_accum, float* __restrict__ denom) {
  pdl_launch_dependents();
  pdl_wait();
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
    const int r = radii[i];
    if (r > 0) {
      if (max_radii_2d) max_radii_2d[i] = fmaxf(max_radii_2d[i], (float)r);
      if (grad2d) {
        const float gx = grad2d[3 * (size_t)i], gy = grad2d[3 * (size_t)i + 1];
        grad_accum[i] += sqrtf(fadd(fmul(gx, gx), fmul(gy, gy)));
        denom[i] += 1.0f;
      }
    }
  }
}

struct TouchedPtrs { const int* p[VIS_MAX_KEYFRAMES]; };

__global__ void __launch_bounds__(256)
visibility_prune_kernel(int P, int K, TouchedPtrs touched, const int* __restrict__ kf_idx, int min_kf, int coviz,
                        int* __restrict__ nr_obs, uint8_t* __restrict__ to_prune) {
  pdl_launch_dependents();
  pdl_wait();
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
    int n = 0;
#pragma unroll 4
    for (int k = 0; k < K; ++k) n += touched.p[k][i] > 0;
    if (nr_obs) nr_obs[i] = n;
    if (to_prune) to_prune[i] = (n <= coviz && kf_idx[i] >= min_kf) ? 1 : 0;
  }
}

static int stats_grid(int P) {
  const int need = (P + 255) / 256;
  return need < 148 * 8 ? need : 148 * 8;
}

cudaError_t launch_densify_stats(cudaStream_t st, int P, const int* radii, const float* grad2d, float* max_radii_2d,
                                 float* grad_accum, float* denom) {
  count_launch(); launch_k(densify_stats_kernel, stats_grid(P), 256, 0, st, P, radii, grad2d, max_radii_2d, grad_accum, denom);
  return cudaGetLastError();
}

cudaError_t launch_visibility_prune(cudaStream_t st, int P, int K, const int* const* touched, const int* kf_idx,
                                    int min_kf, int coviz, int* nr_obs, uint8_t* to_prune) {
  TouchedPtrs tp{};
  for (int k = 0; k < K; ++k) tp.p[k] = touched[k];
  count_launch(); launch_k(visibility_prune_kernel, stats_grid(P), 256, 0, st, P, K, tp, kf_idx, min_kf, coviz, nr_obs, to_prune);
  return cudaGetLastError();
}

}  // namespace gsr
