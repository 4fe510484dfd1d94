// knn.cu -- distCUDA2: mean squared distance to the 3 nearest other points
// (K11-K14 of SURVEY 2.2).
//
// Replaces SimpleKNN::knn (simple-knn/simple_knn.cu:185-221).  Like the
// reference this is an EXACT 3-NN (Morton order is only used to prune), so the
// result does not depend on the traversal and is bit-identical as long as the
// squared distance is evaluated with the reference's operation order
//   d2 = fma(dz,dz, fma(dx,dx, dy*dy))       (simple_knn.cu:134-135 as compiled)
// and the mean as ((b0+b1)+b2)/3 (simple_knn.cu:182).
// B200 design: points are gathered ONCE into Morton order as float4 (the
// reference chases points[indices[i]] for every candidate), boxes form a
// 3-level hierarchy (32 / 1024 / 32768 points; the reference tests every
// 1024-point box for every query), and a warp = 32 Morton-consecutive queries
// walks the hierarchy together: a box is opened when ANY lane needs it, so all
// loads are warp-uniform broadcasts and there is no divergence.  No host
// synchronisation, no allocation (the reference has 2 blocking copies, a
// cudaMalloc/cudaFree and 5 thrust vectors per call).
#include <cfloat>
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int KNN_LEAF = 32;
constexpr int KNN_L2 = 32 * KNN_LEAF;   // 1024 points
constexpr int KNN_L3 = 32 * KNN_L2;     // 32768 points

struct KnnBox { float4 lo; float4 hi; };  // xyz used

struct KnnViews {
  uint32_t* codes_a; uint32_t* codes_b; uint32_t* idx_a; uint32_t* idx_b;
  uint32_t* hist; uint32_t* tickets; uint32_t* lookback; uint32_t* minmax;  // 6 sortable-uint encoded floats
  float4* sorted; KnnBox* leaf; KnnBox* l2; KnnBox* l3;
  size_t zero_begin, zero_end;
};

static size_t carve_knn(char* base, int P, KnnViews* v) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  const size_t n = (size_t)(P > 0 ? P : 1);
  char* ca = take(4 * n); char* cb = take(4 * n); char* ia = take(4 * n); char* ib = take(4 * n);
  size_t zb = off;
  char* h = take(4 * 4 * 256); char* t = take(4 * 4);
  char* lb = take(4 * 4 * sort_blocks((int64_t)n) * SORT_RADIX);
  size_t ze = off;
  char* mm = take(64);
  char* so = take(16 * n);
  char* lf = take(sizeof(KnnBox) * ((n + KNN_LEAF - 1) / KNN_LEAF));
  char* b2 = take(sizeof(KnnBox) * ((n + KNN_L2 - 1) / KNN_L2));
  char* b3 = take(sizeof(KnnBox) * ((n + KNN_L3 - 1) / KNN_L3));
  if (v) {
    v->codes_a = (uint32_t*)ca; v->codes_b = (uint32_t*)cb; v->idx_a = (uint32_t*)ia; v->idx_b = (uint32_t*)ib;
    v->hist = (uint32_t*)h; v->tickets = (uint32_t*)t; v->lookback = (uint32_t*)lb; v->minmax = (uint32_t*)mm;
    v->sorted = (float4*)so; v->leaf = (KnnBox*)lf; v->l2 = (KnnBox*)b2; v->l3 = (KnnBox*)b3;
    v->zero_begin = zb; v->zero_end = ze;
  }
  return off;
}
size_t knn_workspace_bytes(int P) { return carve_knn(nullptr, P, nullptr); }

GSR_DEVINL uint32_t f2sortable(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
GSR_DEVINL float sortable2f(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void knn_init_minmax(uint32_t* mm) {
  if (threadIdx.x < 3) mm[threadIdx.x] = 0xffffffffu;
  else if (threadIdx.x < 6) mm[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256) knn_minmax_kernel(int P, const float* __restrict__ pts, uint32_t* mm) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float c = pts[3 * (size_t)i + k];
      lo[k] = fminf(lo[k], c); hi[k] = fmaxf(hi[k], c);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
      hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
    }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      atomicMin(&mm[k], f2sortable(lo[k]));
      atomicMax(&mm[3 + k], f2sortable(hi[k]));
    }
  }
}

GSR_DEVINL uint32_t spread10(uint32_t x) {  // prepMorton, simple_knn.cu:45-52
  x = (x | (x << 16)) & 0x030000FF;
  x = (x | (x << 8)) & 0x0300F00F;
  x = (x | (x << 4)) & 0x030C30C3;
  x = (x | (x << 2)) & 0x09249249;
  return x;
}

__global__ void __launch_bounds__(256)
knn_morton_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ mm,
                  uint32_t* __restrict__ codes, uint32_t* __restrict__ idx) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  uint32_t q[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float lo = sortable2f(mm[k]), hi = sortable2f(mm[3 + k]);
    const float ext = hi - lo;
    float u = ext > 0.f ? (pts[3 * (size_t)i + k] - lo) / ext : 0.f;
    u = fminf(fmaxf(u, 0.f), 1.f);
    q[k] = (uint32_t)(u * 1023.0f);
  }
  codes[i] = spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2);
  idx[i] = (uint32_t)i;
}

GSR_DEVINL KnnBox box_merge_warp(float3 lo, float3 hi) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo.x = fminf(lo.x, __shfl_xor_sync(0xffffffffu, lo.x, o));
    lo.y = fminf(lo.y, __shfl_xor_sync(0xffffffffu, lo.y, o));
    lo.z = fminf(lo.z, __shfl_xor_sync(0xffffffffu, lo.z, o));
    hi.x = fmaxf(hi.x, __shfl_xor_sync(0xffffffffu, hi.x, o));
    hi.y = fmaxf(hi.y, __shfl_xor_sync(0xffffffffu, hi.y, o));
    hi.z = fmaxf(hi.z, __shfl_xor_sync(0xffffffffu, hi.z, o));
  }
  KnnBox b;
  b.lo = make_float4(lo.x, lo.y, lo.z, 0.f);
  b.hi = make_float4(hi.x, hi.y, hi.z, 0.f);
  return b;
}

// One 1024-thread block = one level-2 box = 32 leaves; gathers points into Morton order.
__global__ void __launch_bounds__(1024)
knn_gather_boxes_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                        float4* __restrict__ sorted, KnnBox* __restrict__ leaf, KnnBox* __restrict__ l2) {
  __shared__ KnnBox s_leaf[32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int i = blockIdx.x * KNN_L2 + tid;
  float3 lo = make_float3(FLT_MAX, FLT_MAX, FLT_MAX), hi = make_float3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
  if (i < P) {
    const uint32_t o = order[i];
    const float x = pts[3 * (size_t)o], y = pts[3 * (size_t)o + 1], z = pts[3 * (size_t)o + 2];
    sorted[i] = make_float4(x, y, z, __uint_as_float(o));
    lo = make_float3(x, y, z); hi = lo;
  }
  KnnBox b = box_merge_warp(lo, hi);
  if (lane == 0) {
    s_leaf[w] = b;
    if (blockIdx.x * KNN_L2 + w * 32 < P) leaf[blockIdx.x * 32 + w] = b;
  }
  __syncthreads();
  if (w == 0) {
    KnnBox m = s_leaf[lane];
    KnnBox t = box_merge_warp(make_float3(m.lo.x, m.lo.y, m.lo.z), make_float3(m.hi.x, m.hi.y, m.hi.z));
    if (lane == 0) l2[blockIdx.x] = t;
  }
}

__global__ void knn_l3_kernel(int n_l2, const KnnBox* __restrict__ l2, KnnBox* __restrict__ l3) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int n_l3 = (n_l2 + 31) / 32;
  if (b >= n_l3) return;
  const int j = b * 32 + lane;
  float3 lo = make_float3(FLT_MAX, FLT_MAX, FLT_MAX), hi = make_float3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
  if (j < n_l2) {
    KnnBox m = l2[j];
    lo = make_float3(m.lo.x, m.lo.y, m.lo.z); hi = make_float3(m.hi.x, m.hi.y, m.hi.z);
  }
  KnnBox t = box_merge_warp(lo, hi);
  if (lane == 0) l3[b] = t;
}

// distBoxPoint (simple_knn.cu:119-129), same operation order as the point distance
// so that fp32(box distance) <= fp32(distance to any point inside) holds exactly.
GSR_DEVINL float box_dist2(const KnnBox& b, float px, float py, float pz) {
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (px < b.lo.x || px > b.hi.x) dx = fminf(fabsf(fsub(px, b.lo.x)), fabsf(fsub(px, b.hi.x)));
  if (py < b.lo.y || py > b.hi.y) dy = fminf(fabsf(fsub(py, b.lo.y)), fabsf(fsub(py, b.hi.y)));
  if (pz < b.lo.z || pz > b.hi.z) dz = fminf(fabsf(fsub(pz, b.lo.z)), fabsf(fsub(pz, b.hi.z)));
  return ffma(dz, dz, ffma(dx, dx, fmul(dy, dy)));
}

GSR_DEVINL void kbest3(float d, float& b0, float& b1, float& b2) {  // updateKBest<3>, simple_knn.cu:131-145
  if (b0 > d) { float t = b0; b0 = d; d = t; }
  if (b1 > d) { float t = b1; b1 = d; d = t; }
  if (b2 > d) { b2 = d; }
}

GSR_DEVINL float pt_dist2(float4 c, float px, float py, float pz) {
  const float dx = fsub(c.x, px), dy = fsub(c.y, py), dz = fsub(c.z, pz);
  return ffma(dz, dz, ffma(dx, dx, fmul(dy, dy)));
}

__global__ void __launch_bounds__(256)
knn_query_kernel(int P, const float4* __restrict__ sorted, const KnnBox* __restrict__ leaf,
                 const KnnBox* __restrict__ l2, const KnnBox* __restrict__ l3, float* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  const bool valid = i < P;
  const int n_leaf = (P + KNN_LEAF - 1) / KNN_LEAF, n_l2 = (P + KNN_L2 - 1) / KNN_L2, n_l3 = (P + KNN_L3 - 1) / KNN_L3;
  float4 me = valid ? sorted[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  const float px = me.x, py = me.y, pz = me.z;

  // seed bound from the +-3 Morton neighbours (simple_knn.cu:156-163)
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  if (valid) {
    for (int j = max(0, i - 3); j <= min(P - 1, i + 3); ++j) {
      if (j == i) continue;
      kbest3(pt_dist2(sorted[j], px, py, pz), b0, b1, b2);
    }
  }
  const float reject = b2;
  b0 = b1 = b2 = FLT_MAX;

  // warp-synchronous descent: open a box if any lane still needs it
  for (int c3 = 0; c3 < n_l3; ++c3) {
    const KnnBox B3 = l3[c3];
    const float d3 = box_dist2(B3, px, py, pz);
    if (!__any_sync(0xffffffffu, valid && !(d3 > reject || d3 > b2))) continue;
    const int e2 = min(n_l2, (c3 + 1) * 32);
    for (int c2 = c3 * 32; c2 < e2; ++c2) {
      const KnnBox B2 = l2[c2];
      const float d2 = box_dist2(B2, px, py, pz);
      if (!__any_sync(0xffffffffu, valid && !(d2 > reject || d2 > b2))) continue;
      const int e1 = min(n_leaf, (c2 + 1) * 32);
      for (int c1 = c2 * 32; c1 < e1; ++c1) {
        const KnnBox B1 = leaf[c1];
        const float d1 = box_dist2(B1, px, py, pz);
        if (!__any_sync(0xffffffffu, valid && !(d1 > reject || d1 > b2))) continue;
        const int e0 = min(P, (c1 + 1) * KNN_LEAF);
        for (int j = c1 * KNN_LEAF; j < e0; ++j) {
          const float d = pt_dist2(sorted[j], px, py, pz);
          if (j != i) kbest3(d, b0, b1, b2);
        }
      }
    }
  }
  if (valid) out[__float_as_uint(me.w)] = fdiv(fadd(fadd(b0, b1), b2), 3.0f);
}

cudaError_t launch_knn(cudaStream_t st, int P, const float* points, float* mean_dist2, void* ws) {
  if (P <= 0) return cudaSuccess;
  KnnViews v;
  carve_knn((char*)ws, P, &v);
  cudaError_t e = cudaMemsetAsync((char*)ws + v.zero_begin, 0, v.zero_end - v.zero_begin, st);
  if (e != cudaSuccess) return e;
  count_launch(); knn_init_minmax<<<1, 32, 0, st>>>(v.minmax);
  count_launch(); knn_minmax_kernel<<<min(148 * 4, (P + 255) / 256), 256, 0, st>>>(P, points, v.minmax);
  count_launch(); knn_morton_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, points, v.minmax, v.codes_a, v.idx_a);
  const SortPlan plan = make_sort_plan(30);  // 30-bit Morton codes: 4 passes (8+8+7+7)
  e = launch_digit_hist_u32(st, v.codes_a, P, plan, v.hist);
  if (e != cudaSuccess) return e;
  e = sort_pairs_u32_prehist(st, P, nullptr, v.codes_a, v.codes_b, v.idx_a, v.idx_b, plan, v.hist, v.tickets,
                             v.lookback, nullptr);
  if (e != cudaSuccess) return e;
  // 4 passes: result is back in the "a" buffers
  const int n_l2 = (P + KNN_L2 - 1) / KNN_L2;
  count_launch(); knn_gather_boxes_kernel<<<n_l2, 1024, 0, st>>>(P, points, v.idx_a, v.sorted, v.leaf, v.l2);
  const int n_l3 = (n_l2 + 31) / 32;
  count_launch(); knn_l3_kernel<<<(n_l3 + 7) / 8, 256, 0, st>>>(n_l2, v.l2, v.l3);
  count_launch(); knn_query_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, v.sorted, v.leaf, v.l2, v.l3, mean_dist2);
  return cudaGetLastError();
}

}  // namespace gsr
