// rows_exchange.cu -- the ONE exchange step of the sharded mapping window, as our own kernel over
// NVLink / NVSwitch peer memory instead of a library allreduce.
//
// What is exchanged: the [P,16] fp32 window rows (include/gsr_b200.h) every rank's per-Gaussian backward wrote
// into ITS copy of a symmetric buffer, summed over the ranks.  The reference has no counterpart (it renders the
// window's keyframes one after the other on one GPU and lets autograd form the sums, utils/slam_mapper.py:273-322,394).
//
// Why not a dense allreduce: a keyframe's gradient only reaches the Gaussians that were actually composited in
// front of the saturation depth -- a few 100 k of the 1 M rows at the headline scene; all other rows are exactly
// zero on every rank.  preprocess_bwd (rows mode) therefore also writes a bitmap of its non-zero rows, and this
// kernel moves only rows that are non-zero on SOME rank:
//   phase 0  every rank multicasts its bitmap (P/8 bytes) and its per-Gaussian visibility counts (P bytes, u8) into
//            slot [rank] of every peer (multimem.st);
//   barrier  per-CTA cross-rank barrier on flag words in the symmetric buffer (release/acquire at system scope,
//            monotonic epochs, bounded spin -> status flag instead of a hang);
//   phase 1  vis_count[i] = sum over ranks of the visibility counts (local, every rank computes all of it);
//            rank r owns Gaussians [r*share, (r+1)*share): for every row of its share whose OR-ed bit is set,
//            four lanes pull the row reduced IN THE SWITCH (multimem.ld_reduce.add.v4.f32, 4 x 16 B) and multicast
//            the sum back to every rank (multimem.st);
//   barrier  every rank's rows are final when the kernel completes.
// Rows nobody touched keep their local (all-zero gradient) content.  Column 15 of a row (the dense `visible`
// flag) is NOT meaningful after this exchange: the visibility sum is returned in vis_count[P] instead.
//
// CTA c of rank A only consumes bitmap / visibility granules that CTA c of every rank pushed, so the per-CTA
// barrier (CTA c of all ranks) is sufficient inside the kernel; kernel completion on a rank implies that every
// CTA of every peer has passed the final barrier, i.e. all writes into this rank's rows have been issued and
// fenced.
#include "common.cuh"
#include "kernels.cuh"

namespace gsr {

constexpr int XCHG_THREADS = 1024;

XchgLayout xchg_layout(int P, int world) {
  XchgLayout L;
  L.n_gran = (P + 127) / 128;
  L.words_pad = L.n_gran * 4;
  L.p_pad = L.n_gran * 128;
  L.rows = 0;
  L.bitmap = align_up((size_t)64 * (size_t)(P > 0 ? P : 1));
  L.vis = L.bitmap + align_up((size_t)world * L.words_pad * 4);
  L.flags = L.vis + align_up((size_t)world * L.p_pad);
  L.total = L.flags + align_up((size_t)XCHG_MAX_CTAS * XCHG_MAX_WORLD * 4);
  return L;
}

GSR_DEVINL float4 mm_ld_reduce_add(const void* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
GSR_DEVINL void mm_st(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
GSR_DEVINL void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
GSR_DEVINL uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
GSR_DEVINL unsigned long long xchg_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// CTA c of every rank meets here. flags (in every rank's buffer): u32[XCHG_MAX_CTAS][XCHG_MAX_WORLD]; rank q's
// arrival at epoch e is the value e in word [c][q] of EVERY rank's flags.
GSR_DEVINL void xchg_barrier(char* local, const unsigned long long* peers, size_t flags_off, int rank, int world,
                             uint32_t epoch, int* status) {
  // Release pattern with cumulativity (PTX memory model): the CTA barrier orders every thread's earlier writes
  // (incl. the relaxed multimem stores) before the flag threads, whose fence.sys + st.release.sys then publish them
  // at system scope -- ONE fence per flag thread instead of one per thread of the CTA (round 1: 1024 x 148 fences
  // per barrier, 13-21 us per barrier at 2 GPUs).
  __syncthreads();
  const int t = threadIdx.x;
  if (t < world) {
    uint32_t* theirs = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(peers[t]) + flags_off) +
                       (size_t)blockIdx.x * XCHG_MAX_WORLD + rank;
    __threadfence_system();
    st_release_sys(theirs, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(local + flags_off) +
                           (size_t)blockIdx.x * XCHG_MAX_WORLD + t;
    const unsigned long long t0 = xchg_now_ns();
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (xchg_now_ns() - t0 > 20000000000ull) {  // 20 s: a peer is gone; report instead of hanging the GPU
        if (status) atomicExch(status, 1);
        break;
      }
    }
  }
  __syncthreads();
}

// Phase timestamps of the LAST exchange on this device (CTA 0, thread 0; globaltimer ns): kernel start, phase 0 done,
// barrier 1 passed, phase 1 done, barrier 2 passed. Read back with gsr_debug_exchange_trace (tools/xchg_check.py).
__device__ unsigned long long g_xchg_trace[5];

__global__ void __launch_bounds__(XCHG_THREADS, 1)
rows_exchange_kernel(int P, int rank, int world, char* __restrict__ local, char* __restrict__ mc,
                     const unsigned long long* __restrict__ peers, XchgLayout L, uint32_t epoch,
                     float* __restrict__ vis_count, int* __restrict__ status) {
  pdl_launch_dependents();
  pdl_wait();
  const bool tracer = blockIdx.x == 0 && threadIdx.x == 0;
  if (tracer) g_xchg_trace[0] = xchg_now_ns();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nC = gridDim.x, c = blockIdx.x;
  const int n_gran = L.n_gran;
  const int share = (n_gran + world - 1) / world;  // granules per rank

  // ---- phase 0: multicast my bitmap granules and visibility granules into slot [rank] of every peer ----
  {
    const size_t slot = L.bitmap + (size_t)rank * L.words_pad * 4;
    // bitmap granule g (16 B) of share r = g / share is consumed by CTA (g % share) % nC of rank r
    const int per_share = c < share ? (share - c + nC - 1) / nC : 0;
    for (int i = tid; i < per_share * world; i += XCHG_THREADS) {
      const int r = i / per_share, k = i - r * per_share;
      const int g = r * share + c + k * nC;
      if (g < n_gran) {
        const float4 v = *reinterpret_cast<const float4*>(local + slot + (size_t)g * 16);
        mm_st(mc + slot + (size_t)g * 16, v);
      }
    }
    // visibility granule g (128 B = 8 pieces) is consumed by CTA g % nC of every rank
    const size_t vslot = L.vis + (size_t)rank * L.p_pad;
    const int my_gran = c < n_gran ? (n_gran - c + nC - 1) / nC : 0;
    for (int i = tid; i < my_gran * 8; i += XCHG_THREADS) {
      const int g = c + (i >> 3) * nC;
      const size_t off = vslot + (size_t)g * 128 + (size_t)(i & 7) * 16;
      mm_st(mc + off, *reinterpret_cast<const float4*>(local + off));
    }
  }
  if (tracer) g_xchg_trace[1] = xchg_now_ns();
  xchg_barrier(local, peers, L.flags, rank, world, epoch, status);
  if (tracer) g_xchg_trace[2] = xchg_now_ns();

  // ---- phase 1a: visibility sum (every rank, all Gaussians) ----
  if (vis_count) {
    const int my_gran = c < n_gran ? (n_gran - c + nC - 1) / nC : 0;
    for (int i = tid; i < my_gran * 8; i += XCHG_THREADS) {
      const int g = c + (i >> 3) * nC;
      const size_t e0 = (size_t)g * 128 + (size_t)(i & 7) * 16;  // first Gaussian of this 16-entry piece
      unsigned acc[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) acc[k] = 0;
      for (int q = 0; q < world; ++q) {
        const uint4 u = __ldcv(reinterpret_cast<const uint4*>(local + L.vis + (size_t)q * L.p_pad + e0));
        const unsigned wv[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] += (wv[k >> 2] >> (8 * (k & 3))) & 0xffu;
      }
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (e0 + k < (size_t)P) vis_count[e0 + k] = (float)acc[k];
    }
  }

  // ---- phase 1b: rows of my share that are non-zero on some rank: in-switch sum, multicast back ----
  {
    const int per_share = c < share ? (share - c + nC - 1) / nC : 0;  // my granules: gl = c + k*nC
    const int sub = lane >> 2, part = lane & 3;
    for (int task = warp; task < per_share * 4; task += XCHG_THREADS / 32) {
      const int g = rank * share + c + (task >> 2) * nC;
      if (g >= n_gran) continue;
      const int word = g * 4 + (task & 3);
      uint32_t m = 0;
      for (int q = lane; q < world; q += 32)
        m |= __ldcv(reinterpret_cast<const uint32_t*>(local + L.bitmap + ((size_t)q * L.words_pad + word) * 4));
      m = __reduce_or_sync(0xffffffffu, m);
      const int n = __popc(m);
      const size_t row0 = (size_t)word * 32;
      float4 v[4];
      size_t off[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = i * 8 + sub;
        off[i] = ~(size_t)0;
        if (k < n) {
          const int bit = __fns(m, 0, k + 1);
          off[i] = (row0 + bit) * 64 + (size_t)part * 16;
          v[i] = mm_ld_reduce_add(mc + L.rows + off[i]);
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (off[i] != ~(size_t)0) mm_st(mc + L.rows + off[i], v[i]);
    }
  }
  if (tracer) g_xchg_trace[3] = xchg_now_ns();
  xchg_barrier(local, peers, L.flags, rank, world, epoch + 1, status);
  if (tracer) g_xchg_trace[4] = xchg_now_ns();
}

cudaError_t launch_rows_exchange(cudaStream_t st, int P, int capacity, int rank, int world, int n_ctas,
                                 void* local_buf, void* multicast_buf, const unsigned long long* peer_bufs,
                                 uint32_t epoch, float* vis_count, int* status) {
  // offsets and slot strides follow the buffer's CAPACITY (so they never move while the map grows inside it);
  // the loops cover the live P rows only
  XchgLayout L = xchg_layout(capacity, world);
  L.n_gran = (P + 127) / 128;
  count_launch();
  launch_k(rows_exchange_kernel, n_ctas, XCHG_THREADS, 0, st, P, rank, world, (char*)local_buf,
           (char*)multicast_buf, peer_bufs, L, epoch, vis_count, status);
  return cudaGetLastError();
}

cudaError_t read_exchange_trace(unsigned long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_xchg_trace, sizeof(g_xchg_trace));
}

}  // namespace gsr
