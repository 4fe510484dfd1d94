// insert.cu -- Gaussian insertion from an RGB-D keyframe: the device side of
// GaussianModel.create_viewpoint_pcd (gaussian_splatting/scene/gaussian_model.py:121-319), the caller of
// distCUDA2.  The reference builds it from ~40 framework ops (two sort-based torch.median, meshgrid/stack,
// boolean-mask indexing x4, a batched 3x3 matmul, cat, ...); here it is
//   select_kth_kernel x4      lower median by 8-bit MSD radix select (torch.median semantics, :177,:218)
//   insert_count/compact      densification mask (:191-226) + stable compaction in the reference's (W,H) order
//   insert_points_kernel      gather kept pixels, exposure + clamp (:135-139), unprojection K^-1, c2w (:262-279)
//   insert_finish_kernel      log-scale from the kNN distance (:294-306), identity rotation, opacity logit(0.5)
// with distCUDA2 (knn.cu) in between.  The random subsample (:247-259) stays torch.randperm on the host RNG, like
// the reference, so that a seeded run selects the same pixels.
#include <mutex>
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int SEL_THREADS = 256;
constexpr int SEL_ITEMS = 8;  // values per thread per CTA

// ws layout (u32 words): [0] prefix key, [1] remaining rank k, [2..5] tickets, [8 .. 8+4*256) histograms
GSR_DEVINL uint32_t f32_sortable(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
GSR_DEVINL float sortable_f32(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void __launch_bounds__(SEL_THREADS)
select_kth_kernel(long long n, const float* __restrict__ a, const float* __restrict__ b, int pass, uint32_t k0,
                  uint32_t* __restrict__ ws, float* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_hist[256];
  __shared__ bool s_last;
  const int tid = threadIdx.x;
  s_hist[tid] = 0;
  __syncthreads();
  const int shift = 24 - 8 * pass;
  const uint32_t prefix = pass == 0 ? 0u : ws[0];
  const uint32_t himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
  const long long base = (long long)blockIdx.x * SEL_THREADS * SEL_ITEMS;
#pragma unroll
  for (int i = 0; i < SEL_ITEMS; ++i) {
    const long long j = base + (long long)i * SEL_THREADS + tid;
    if (j < n) {
      const float v = b ? fabsf(a[j] - b[j]) : a[j];
      const uint32_t k = f32_sortable(v);
      if ((k & himask) == prefix) atomicAdd(&s_hist[(k >> shift) & 0xffu], 1u);
    }
  }
  __syncthreads();
  uint32_t* hist = ws + 8 + 256 * pass;
  if (s_hist[tid]) atomicAdd(&hist[tid], s_hist[tid]);
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = atomicAdd(&ws[2 + pass], 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last CTA of the pass: find the bin that holds rank k (ascending order)
  s_hist[tid] = *reinterpret_cast<volatile uint32_t*>(&hist[tid]);
  __syncthreads();
  if (tid == 0) {
    uint32_t k = pass == 0 ? k0 : ws[1], bin = 0;
    for (; bin < 255; ++bin) {
      if (k < s_hist[bin]) break;
      k -= s_hist[bin];
    }
    const uint32_t np = prefix | (bin << shift);
    ws[0] = np;
    ws[1] = k;
    if (pass == 3) *out = sortable_f32(np);
  }
}

// Flat index i of the reference's (W,H)-ordered flattening (gaussian_model.py:180-186): i = x*H + y.
GSR_DEVINL bool insert_keep(int i, int W, int H, const float* __restrict__ gt_depth,
                            const float* __restrict__ render_depth, const float* __restrict__ render_opacity,
                            bool init, float median_err) {
  const int x = i / H, y = i - x * H;
  const size_t p = (size_t)y * W + x;
  const float d = gt_depth[p];
  bool m = d >= 1e-3f;                                   // :191
  if (!init) {
    const float rd = render_depth[p];
    const bool low_opacity = render_opacity ? render_opacity[p] < 0.5f : true;            // :197-200
    const bool c1 = d < rd;                                                                // :207
    const bool c2 = fabsf(d - rd) > 50.0f * median_err;                                    // :218-220
    m = m && (low_opacity || (c1 && c2));                                                  // :223-226
  }
  return m;
}

constexpr int INS_THREADS = 1024;

__global__ void __launch_bounds__(INS_THREADS)
insert_count_kernel(int W, int H, const float* __restrict__ gt_depth, const float* __restrict__ render_depth,
                    const float* __restrict__ render_opacity, int init, const float* __restrict__ median_err,
                    uint32_t* __restrict__ block_counts) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * INS_THREADS + threadIdx.x;
  const bool keep = i < W * H && insert_keep(i, W, H, gt_depth, render_depth, render_opacity, init != 0,
                                             init ? 0.f : *median_err);
  const int c = __syncthreads_count(keep);
  if (threadIdx.x == 0) block_counts[blockIdx.x] = (uint32_t)c;
}

__global__ void __launch_bounds__(INS_THREADS)
insert_compact_kernel(int W, int H, const float* __restrict__ gt_depth, const float* __restrict__ render_depth,
                      const float* __restrict__ render_opacity, int init, const float* __restrict__ median_err,
                      const uint32_t* __restrict__ block_counts, int n_blocks, int32_t* __restrict__ sel,
                      uint32_t* __restrict__ total) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_warp[INS_THREADS / 32];
  __shared__ uint32_t s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // exclusive offset of this block = sum of the earlier blocks' counts (n_blocks <= a few thousand)
  uint32_t part = 0, all = 0;
  for (int b = tid; b < n_blocks; b += INS_THREADS) {
    const uint32_t c = block_counts[b];
    all += c;
    if (b < (int)blockIdx.x) part += c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    part += __shfl_xor_sync(0xffffffffu, part, o);
    all += __shfl_xor_sync(0xffffffffu, all, o);
  }
  if (lane == 0) s_warp[warp] = part;
  __syncthreads();
  if (tid == 0) {
    uint32_t t = 0;
    for (int w = 0; w < INS_THREADS / 32; ++w) t += s_warp[w];
    s_base = t;
  }
  __syncthreads();
  if (blockIdx.x == 0) {  // block 0 also publishes the total
    __shared__ uint32_t s_all[INS_THREADS / 32];
    if (lane == 0) s_all[warp] = all;
    __syncthreads();
    if (tid == 0) {
      uint32_t t = 0;
      for (int w = 0; w < INS_THREADS / 32; ++w) t += s_all[w];
      *total = t;
    }
  }
  const int i = blockIdx.x * INS_THREADS + tid;
  const bool keep = i < W * H && insert_keep(i, W, H, gt_depth, render_depth, render_opacity, init != 0,
                                             init ? 0.f : *median_err);
  const uint32_t bal = __ballot_sync(0xffffffffu, keep);
  __syncthreads();
  if (lane == 0) s_warp[warp] = __popc(bal);
  __syncthreads();
  uint32_t off = s_base;
  for (int w = 0; w < warp; ++w) off += s_warp[w];
  if (keep) sel[off + __popc(bal & ((1u << lane) - 1u))] = i;
}

__global__ void __launch_bounds__(256)
insert_points_kernel(int K, const int32_t* __restrict__ sel, const long long* __restrict__ pick, int W, int H,
                     const float* __restrict__ rgb, const float* __restrict__ exposure_a,
                     const float* __restrict__ exposure_b, const float* __restrict__ gt_depth,
                     const int32_t* __restrict__ segmentation, const float* __restrict__ Rw2c,
                     const float* __restrict__ Tw2c, float fx, float fy, float cx, float cy,
                     float* __restrict__ xyz, float* __restrict__ feat, int32_t* __restrict__ ids) {
  pdl_launch_dependents();
  pdl_wait();
  // K^-1 of the upper-triangular intrinsics in closed form and c2w = [R | T]^-1 by the adjugate (general R, like the
  // reference's torch.linalg.inv, :163-165) -- once per CTA, instead of two cuSOLVER calls with host syncs
  __shared__ float K_inv[9];
  __shared__ float c2w[12];
  if (threadIdx.x == 0) {
    K_inv[0] = 1.0f / fx; K_inv[1] = 0.f; K_inv[2] = -cx / fx;
    K_inv[3] = 0.f; K_inv[4] = 1.0f / fy; K_inv[5] = -cy / fy;
    K_inv[6] = 0.f; K_inv[7] = 0.f; K_inv[8] = 1.f;
    const float a = Rw2c[0], b = Rw2c[1], c = Rw2c[2], d = Rw2c[3], e = Rw2c[4], f = Rw2c[5], g = Rw2c[6],
                h = Rw2c[7], i = Rw2c[8];
    const float A = e * i - f * h, B = -(d * i - f * g), Cc = d * h - e * g;
    const float inv_det = 1.0f / (a * A + b * B + c * Cc);
    float inv[9] = {A * inv_det, -(b * i - c * h) * inv_det, (b * f - c * e) * inv_det,
                    B * inv_det, (a * i - c * g) * inv_det, -(a * f - c * d) * inv_det,
                    Cc * inv_det, -(a * h - b * g) * inv_det, (a * e - b * d) * inv_det};
    for (int r = 0; r < 3; ++r) {
      c2w[4 * r] = inv[3 * r]; c2w[4 * r + 1] = inv[3 * r + 1]; c2w[4 * r + 2] = inv[3 * r + 2];
      c2w[4 * r + 3] = -(inv[3 * r] * Tw2c[0] + inv[3 * r + 1] * Tw2c[1] + inv[3 * r + 2] * Tw2c[2]);
    }
  }
  __syncthreads();
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= K) return;
  const int i = sel[pick ? pick[j] : j];
  const int x = i / H, y = i - x * H;
  const size_t p = (size_t)y * W + x, HW = (size_t)H * W;
  float r = rgb[p], g = rgb[HW + p], b = rgb[2 * HW + p];
  if (exposure_a) {  // not init: clamp(exp(a) * rgb + b, 0, 1)   (:135-139)
    const float ea = expf(*exposure_a), eb = *exposure_b;
    r = fminf(fmaxf(ea * r + eb, 0.f), 1.f);
    g = fminf(fmaxf(ea * g + eb, 0.f), 1.f);
    b = fminf(fmaxf(ea * b + eb, 0.f), 1.f);
  }
  feat[3 * (size_t)j] = r; feat[3 * (size_t)j + 1] = g; feat[3 * (size_t)j + 2] = b;
  if (ids) ids[j] = segmentation[p];
  const float d = gt_depth[p];
  const float u = (float)x + 0.5f, v = (float)y + 0.5f;  // pixel centres (:238-239)
  float pc[3];
#pragma unroll
  for (int k = 0; k < 3; ++k)
    pc[k] = fmaf(K_inv[3 * k + 2], 1.0f, fmaf(K_inv[3 * k + 1], v, K_inv[3 * k] * u)) * d;  // :262-272
#pragma unroll
  for (int k = 0; k < 3; ++k)
    xyz[3 * (size_t)j + k] =
        fmaf(c2w[4 * k + 2], pc[2], fmaf(c2w[4 * k + 1], pc[1], c2w[4 * k] * pc[0])) + c2w[4 * k + 3];  // :275-277
}

__global__ void __launch_bounds__(256)
insert_finish_kernel(int K, const float* __restrict__ dist2, const float* __restrict__ median_depth,
                     int scale_cols, float* __restrict__ scales, float* __restrict__ rots,
                     float* __restrict__ opacities) {
  pdl_launch_dependents();
  pdl_wait();
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= K) return;
  const float point_size = fminf(0.05f, 0.01f * *median_depth);                    // :169-177
  const float s = logf(sqrtf(fmaxf(dist2[j], 0.0000001f) * point_size));           // :294-303
  for (int k = 0; k < scale_cols; ++k) scales[(size_t)j * scale_cols + k] = s;      // :304-306
  *reinterpret_cast<float4*>(rots + 4 * (size_t)j) = make_float4(1.f, 0.f, 0.f, 0.f);  // :308-311
  opacities[j] = 0.0f;  // inverse_sigmoid(0.5) = log(0.5 / 0.5)                     (:313-318)
}

cudaError_t launch_select_kth(cudaStream_t st, long long n, const float* a, const float* b, long long k,
                              float* out, uint32_t* ws) {
  cudaError_t e = cudaMemsetAsync(ws, 0, INSERT_SELECT_WS_BYTES, st);
  if (e != cudaSuccess) return e;
  const int nb = (int)((n + SEL_THREADS * SEL_ITEMS - 1) / (SEL_THREADS * SEL_ITEMS));
  for (int pass = 0; pass < 4; ++pass) {
    count_launch();
    launch_k(select_kth_kernel, nb, SEL_THREADS, 0, st, n, a, b, pass, (uint32_t)k, ws, out);
  }
  return cudaGetLastError();
}

cudaError_t launch_insert_select(cudaStream_t st, int W, int H, const float* gt_depth, const float* render_depth,
                                 const float* render_opacity, int init, const float* median_err, int32_t* sel,
                                 uint32_t* total, uint32_t* block_counts) {
  const int nb = (W * H + INS_THREADS - 1) / INS_THREADS;
  count_launch();
  launch_k(insert_count_kernel, nb, INS_THREADS, 0, st, W, H, gt_depth, render_depth, render_opacity, init,
           median_err, block_counts);
  count_launch();
  launch_k(insert_compact_kernel, nb, INS_THREADS, 0, st, W, H, gt_depth, render_depth, render_opacity, init,
           median_err, (const uint32_t*)block_counts, nb, sel, total);
  return cudaGetLastError();
}

cudaError_t launch_insert_points(cudaStream_t st, int K, const int32_t* sel, const long long* pick, int W, int H,
                                 const float* rgb, const float* exposure_a, const float* exposure_b,
                                 const float* gt_depth, const int32_t* segmentation, const float* R,
                                 const float* T, float fx, float fy, float cx, float cy, float* xyz, float* feat,
                                 int32_t* ids) {
  count_launch();
  launch_k(insert_points_kernel, (K + 255) / 256, 256, 0, st, K, sel, pick, W, H, rgb, exposure_a, exposure_b,
           gt_depth, segmentation, R, T, fx, fy, cx, cy, xyz, feat, ids);
  return cudaGetLastError();
}

cudaError_t launch_insert_finish(cudaStream_t st, int K, const float* dist2, const float* median_depth,
                                 int scale_cols, float* scales, float* rots, float* opacities) {
  count_launch();
  launch_k(insert_finish_kernel, (K + 255) / 256, 256, 0, st, K, dist2, median_depth, scale_cols, scales, rots,
           opacities);
  return cudaGetLastError();
}

// ---- gsr_insert_select_static: the same two kernels with their scratch in a file-scope device array ----------
// For callers that may not allocate device memory (stream capture, serving loops, harnesses that forbid cudaMalloc).
// The block counts of insert_count_kernel live in g_insert_static_counts (16 KB, held by every GPU that loads the
// library until the process exits) instead of a caller workspace.  Nothing in it has to start at zero: every word the
// compaction kernel reads ([0, n_blocks)) was written by the count kernel of the SAME call.  All calls on a device
// share the array, so two calls that overlap on the GPU (different streams) race; calls on one stream are ordered.
constexpr int INSERT_STATIC_MAX_BLOCKS = GSR_INSERT_STATIC_MAX_PIXELS / INS_THREADS;
static_assert(GSR_INSERT_STATIC_MAX_PIXELS % INS_THREADS == 0, "whole blocks");
static __device__ __align__(256) uint32_t g_insert_static_counts[INSERT_STATIC_MAX_BLOCKS];

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mu;
uint32_t* g_static_ptr[STATIC_MAX_DEVICES] = {};

// The device address of g_insert_static_counts on the CURRENT device (the name itself is only a host-side handle);
// looked up once per device.
cudaError_t insert_static_scratch(uint32_t** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[dev]) {
    uint32_t* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_insert_static_counts);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return cudaSuccess;
}
}  // namespace

}  // namespace gsr

extern "C" int gsr_insert_select_static(void* stream, int width, int height, const float* gt_depth,
                                        const float* render_depth, const float* render_opacity, int init,
                                        const float* median_err, int32_t* selected, uint32_t* count) {
  if (width <= 0 || height <= 0 || !gt_depth || !selected || !count) return GSR_EINVAL;
  if (!init && (!render_depth || !median_err)) return GSR_EINVAL;
  if ((long long)width * height > (long long)GSR_INSERT_STATIC_MAX_PIXELS) return GSR_ESTATICLIMIT;
  uint32_t* counts = nullptr;
  cudaError_t e = gsr::insert_static_scratch(&counts);
  if (e != cudaSuccess) return (int)e;
  e = gsr::launch_insert_select((cudaStream_t)stream, width, height, gt_depth, render_depth, render_opacity, init,
                                median_err, selected, count, counts);
  return e == cudaSuccess ? GSR_OK : (int)e;
}
