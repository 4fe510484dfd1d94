// oracle/ref_shim.cu -- TEST INFRASTRUCTURE, not product code.
//
// A C-ABI wrapper (our own code) around the UNMODIFIED reference rasterizer
// and simple-knn, whose sources are compiled where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/libref_dgr.so.
// Entry points wrapped:
//   CudaRasterizer::Rasterizer::forward   (DGR/cuda_rasterizer/rasterizer.h:33-60)
//   CudaRasterizer::Rasterizer::backward  (DGR/cuda_rasterizer/rasterizer.h:62-92)
//   CudaRasterizer::Rasterizer::markVisible (rasterizer.h:26-31)
//   SimpleKNN::knn                        (SKNN/simple_knn.h)
// All pointers are DEVICE pointers. The three state buffers the reference
// obtains through std::function<char*(size_t)> are owned by this shim
// (grow-only cudaMalloc), so torch is not needed to build or run it.
// Only tests/, __graft_entry__.smoke() and bench.py --impl reference may load it.
#include <cstdint>
#include <cstdio>
#include <functional>
#include <cuda_runtime.h>
#include "cuda_rasterizer/rasterizer_impl.h"
#include "simple_knn.h"

namespace {
struct Buf {
  char* p = nullptr;
  size_t cap = 0;
  char* get(size_t n) {
    if (n > cap) {
      if (p) cudaFree(p);
      size_t want = n + n / 4 + 4096;
      if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; cap = 0; return nullptr; }
      cap = want;
    }
    return p;
  }
};
// One forward's saved state. The reference keeps it in three torch byte tensors saved on the autograd
// context (diff_gaussian_rasterization/__init__.py:104); here a small ring of slots plays that role so that
// several forwards can be alive before their backwards run (a mapping window sums the keyframes' losses and
// calls backward once, utils/slam_mapper.py:273-322,394).
struct State {
  Buf geom, bin, img;
  int P = 0, R = 0, N = 0;
};
constexpr int kSlots = 16;
State g_state[kSlots];
int g_slot = 0;
#define g_geom (g_state[g_slot].geom)
#define g_bin (g_state[g_slot].bin)
#define g_img (g_state[g_slot].img)
#define g_P (g_state[g_slot].P)
#define g_R (g_state[g_slot].R)
#define g_N (g_state[g_slot].N)
}  // namespace

extern "C" {

// Selects the state slot used by the following ref_forward / ref_backward / ref_last_state calls.
int ref_set_slot(int slot) {
  if (slot < 0 || slot >= kSlots) return -1;
  g_slot = slot;
  return kSlots;
}

int ref_forward(int P, int D, int M, const float* bg, int W, int H,
                const float* means3D, const float* shs, const float* colors,
                const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos,
                float tan_fovx, float tan_fovy, int prefiltered,
                float* out_color, float* out_depth, float* out_opacity,
                int* radii, int* n_touched, int debug) {
  std::function<char*(size_t)> fg = [](size_t n) { return g_geom.get(n); };
  std::function<char*(size_t)> fb = [](size_t n) { return g_bin.get(n); };
  std::function<char*(size_t)> fi = [](size_t n) { return g_img.get(n); };
  int R = CudaRasterizer::Rasterizer::forward(
      fg, fb, fi, P, D, M, bg, W, H, means3D, shs, colors, opacities, scales,
      scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
      tan_fovx, tan_fovy, prefiltered != 0, out_color, out_depth, out_opacity,
      radii, n_touched, debug != 0);
  g_P = P; g_R = R; g_N = W * H;
  return R;
}

void ref_backward(int P, int D, int M, int R, const float* bg, int W, int H,
                  const float* means3D, const float* shs, const float* colors,
                  const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const float* viewmatrix,
                  const float* projmatrix, const float* projmatrix_raw,
                  const float* campos, float tan_fovx, float tan_fovy,
                  const int* radii, const float* dL_dpix, const float* dL_dpix_depth,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                  float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D,
                  float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  float* dL_dtau, int debug) {
  CudaRasterizer::Rasterizer::backward(
      P, D, M, R, bg, W, H, means3D, shs, colors, scales, scale_modifier, rotations,
      cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, campos, tan_fovx,
      tan_fovy, radii, g_geom.p, g_bin.p, g_img.p, dL_dpix, dL_dpix_depth,
      dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth, dL_dmean3D,
      dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dtau, debug != 0);
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix,
                      unsigned char* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix,
                                          reinterpret_cast<bool*>(present));
}

// Internal state of the last ref_forward call (device pointers into the
// shim-owned buffers), for bit-exact checks of keys / lists / ranges.
void ref_last_state(void** depths, void** means2D, void** cov3D, void** conic_opacity,
                    void** tiles_touched, void** point_offsets, void** point_list_keys,
                    void** point_list, void** ranges, void** n_contrib,
                    void** accum_alpha) {
  char* c = g_geom.p;
  CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(c, g_P);
  c = g_bin.p;
  CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(c, g_R);
  c = g_img.p;
  CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(c, g_N);
  *depths = g.depths; *means2D = g.means2D; *cov3D = g.cov3D;
  *conic_opacity = g.conic_opacity; *tiles_touched = g.tiles_touched;
  *point_offsets = g.point_offsets; *point_list_keys = b.point_list_keys;
  *point_list = b.point_list; *ranges = im.ranges; *n_contrib = im.n_contrib;
  *accum_alpha = im.accum_alpha;
}

void ref_knn(int P, float* points, float* mean_dists) {
  SimpleKNN::knn(P, reinterpret_cast<float3*>(points), mean_dists);
}

int ref_sync() { return (int)cudaDeviceSynchronize(); }

}  // extern "C"
