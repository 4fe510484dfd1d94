// surgery.cu -- map growth and pruning next to the rasterizer (second half): the tensor surgery
// GaussianModel does on its parameters and Adam state around every keyframe insertion and prune
// (gaussian_splatting/scene/gaussian_model.py:642-778: cat_tensors_to_optimizer, _prune_optimizer,
// densification_postfix, prune_points).
//
// The reference re-creates every tensor: 15 torch.cat (parameter, exp_avg, exp_avg_sq of 5 groups) that each copy
// all P old rows to append K, and on a prune ~19 boolean-mask indexings, each a nonzero + gather + host sync.  Here
// the rows live in capacity-backed buffers (monogs_b200/optimizer_surgery.py):
//   rows_append_kernel    ONE launch writes the K new rows of all groups and zero-fills their moment rows: O(K);
//   keep_count / keep_scan / rows_compact_kernel   ONE stable compaction of all row arrays by the keep mask into
//                         the alternate buffers: the mask is scanned once, every surviving row is moved once, the
//                         host reads the new row count once.
// Pure data movement: results are bit-identical to the reference's (tests/test_gpu_surgery.py, golden vectors made
// by the reference functions themselves).
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int SG_THREADS = 256;

struct AppendArgs {
  int n;
  float* param[SURGERY_MAX_GROUPS];
  float* exp_avg[SURGERY_MAX_GROUPS];      // may be NULL (no optimizer state yet / not an optimizer group)
  float* exp_avg_sq[SURGERY_MAX_GROUPS];
  const float* rows[SURGERY_MAX_GROUPS];   // [K, cols]
  int cols[SURGERY_MAX_GROUPS];
};

// blockIdx.y = group; threads over the K * cols new elements of that group.
__global__ void __launch_bounds__(SG_THREADS)
rows_append_kernel(AppendArgs a, long long P_old, long long K) {
  pdl_launch_dependents();
  pdl_wait();
  const int g = blockIdx.y;
  const long long n = K * a.cols[g];
  const long long base = P_old * a.cols[g];
  for (long long i = (long long)blockIdx.x * SG_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * SG_THREADS) {
    a.param[g][base + i] = a.rows[g][i];
    if (a.exp_avg[g]) a.exp_avg[g][base + i] = 0.0f;
    if (a.exp_avg_sq[g]) a.exp_avg_sq[g][base + i] = 0.0f;
  }
}

// keep[P] (0/1 bytes) -> number of kept rows per 256-row chunk
__global__ void __launch_bounds__(SG_THREADS)
keep_count_kernel(long long P, const uint8_t* __restrict__ keep, uint32_t* __restrict__ chunk_sums) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = (long long)blockIdx.x * SG_THREADS + threadIdx.x;
  const int c = __syncthreads_count(i < P && keep[i] != 0);
  if (threadIdx.x == 0) chunk_sums[blockIdx.x] = (uint32_t)c;
}

// exclusive scan of the chunk sums in place (one CTA, any length) + total
__global__ void __launch_bounds__(1024)
keep_scan_kernel(int n_chunks, uint32_t* __restrict__ chunk_sums, uint32_t* __restrict__ total) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < n_chunks; base += 1024) {
    const int i = base + tid;
    const uint32_t v = i < n_chunks ? chunk_sums[i] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    if (w == 0) {
      uint32_t x = s_warp[lane], xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, xi, o);
        if (lane >= o) xi += t;
      }
      s_warp[lane] = xi - x;
    }
    __syncthreads();
    const uint32_t carry = s_carry;
    if (i < n_chunks) chunk_sums[i] = carry + s_warp[w] + inc - v;
    __syncthreads();
    if (tid == 1023) s_carry = carry + s_warp[w] + inc;
    __syncthreads();
  }
  if (tid == 0) *total = s_carry;
}

struct CompactArgs {
  int n;
  const uint32_t* src[SURGERY_MAX_ARRAYS];   // rows as 32-bit words
  uint32_t* dst[SURGERY_MAX_ARRAYS];
  int words[SURGERY_MAX_ARRAYS];             // 32-bit words per row
};

// thread = source row; kept rows move to chunk offset + rank inside the chunk (stable)
__global__ void __launch_bounds__(SG_THREADS)
rows_compact_kernel(CompactArgs a, long long P, const uint8_t* __restrict__ keep,
                    const uint32_t* __restrict__ chunk_offsets) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ uint32_t s_w[SG_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const long long i = (long long)blockIdx.x * SG_THREADS + tid;
  const bool k = i < P && keep[i] != 0;
  const uint32_t bal = __ballot_sync(0xffffffffu, k);
  if (lane == 0) s_w[w] = __popc(bal);
  __syncthreads();
  uint32_t before = 0;
#pragma unroll
  for (int ww = 0; ww < SG_THREADS / 32; ++ww) before += ww < w ? s_w[ww] : 0u;
  if (!k) return;
  const long long d = (long long)chunk_offsets[blockIdx.x] + before + __popc(bal & ((1u << lane) - 1u));
  for (int j = 0; j < a.n; ++j) {
    const int nw = a.words[j];
    const uint32_t* s = a.src[j] + i * nw;
    uint32_t* t = a.dst[j] + d * nw;
    for (int q = 0; q < nw; ++q) t[q] = s[q];
  }
}

cudaError_t launch_rows_append(cudaStream_t st, int n_groups, float* const* param, float* const* exp_avg,
                               float* const* exp_avg_sq, const float* const* rows, const int* cols, long long P_old,
                               long long K) {
  AppendArgs a{};
  a.n = n_groups;
  long long most = 0;
  for (int g = 0; g < n_groups; ++g) {
    a.param[g] = param[g]; a.exp_avg[g] = exp_avg ? exp_avg[g] : nullptr;
    a.exp_avg_sq[g] = exp_avg_sq ? exp_avg_sq[g] : nullptr; a.rows[g] = rows[g]; a.cols[g] = cols[g];
    most = most > K * cols[g] ? most : K * cols[g];
  }
  long long bx = (most + SG_THREADS - 1) / SG_THREADS;
  if (bx > 148 * 8) bx = 148 * 8;
  count_launch();
  launch_k(rows_append_kernel, dim3((unsigned)bx, (unsigned)n_groups), SG_THREADS, 0, st, a, P_old, K);
  return cudaGetLastError();
}

size_t surgery_ws_bytes(long long P) { return align_up(4 * (size_t)((P + SG_THREADS - 1) / SG_THREADS + 1)) + 256; }

cudaError_t launch_rows_compact(cudaStream_t st, int n_arrays, const void* const* src, void* const* dst,
                                const int* row_bytes, long long P, const uint8_t* keep, uint32_t* new_count,
                                void* ws) {
  CompactArgs a{};
  a.n = n_arrays;
  for (int j = 0; j < n_arrays; ++j) {
    a.src[j] = (const uint32_t*)src[j]; a.dst[j] = (uint32_t*)dst[j]; a.words[j] = row_bytes[j] / 4;
  }
  const int n_chunks = (int)((P + SG_THREADS - 1) / SG_THREADS);
  uint32_t* chunk = (uint32_t*)ws;
  count_launch(); launch_k(keep_count_kernel, n_chunks, SG_THREADS, 0, st, P, keep, chunk);
  count_launch(); launch_k(keep_scan_kernel, 1, 1024, 0, st, n_chunks, chunk, new_count);
  count_launch(); launch_k(rows_compact_kernel, n_chunks, SG_THREADS, 0, st, a, P, keep, (const uint32_t*)chunk);
  return cudaGetLastError();
}

}  // namespace gsr
