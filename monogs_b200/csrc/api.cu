// api.cu -- the C ABI declared in include/gsr_b200.h.
// Host-side orchestration only: argument checks, workspace carving, kernel
// launches on the caller's stream.  See the header for which reference
// interface each entry point replaces.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include "common.cuh"
#include "kernels.cuh"

#include <atomic>
using namespace gsr;

namespace gsr {
static std::atomic<unsigned long long> g_launches{0};
unsigned long long count_launch(unsigned long long add) { return g_launches.fetch_add(add) + add; }
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("GSR_PDL");
    return !(e && strcmp(e, "0") == 0);
  }();
  return on;
}
}  // namespace gsr

namespace {

// Number of bits needed to represent n (getHigherMsb, rasterizer_impl.cu:35-50).
int higher_msb(uint32_t n) {
  int b = 0;
  while (n >> b) ++b;
  return b;
}
struct HostSlot {
  unsigned long long total;
  int error_flag;
};
// The forward's one host read (num_rendered + prefilter flag) goes through a pinned slot + event taken from a
// pool, one per forward IN FLIGHT and per device: concurrent forwards (several streams, threads or devices of one
// process) neither share a slot nor serialise on a lock while the host waits. The mutex only guards the free list.
struct ReadSlot {
  HostSlot* pinned = nullptr;
  cudaEvent_t ev = nullptr;
  int device = -1;
};
std::mutex g_mu;
std::vector<ReadSlot> g_free_slots;
int acquire_read_slot(ReadSlot* out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    for (size_t i = 0; i < g_free_slots.size(); ++i)
      if (g_free_slots[i].device == dev) {
        *out = g_free_slots[i];
        g_free_slots[i] = g_free_slots.back();
        g_free_slots.pop_back();
        return GSR_OK;
      }
  }
  ReadSlot s;
  s.device = dev;
  e = cudaHostAlloc((void**)&s.pinned, sizeof(HostSlot), cudaHostAllocDefault);
  if (e != cudaSuccess) return (int)e;
  e = cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming);  // bound to the CURRENT device = the stream's
  if (e != cudaSuccess) { cudaFreeHost(s.pinned); return (int)e; }
  *out = s;
  return GSR_OK;
}
struct ReadSlotGuard {  // returns the slot to the pool on every exit path
  ReadSlot s;
  ~ReadSlotGuard() {
    if (!s.pinned) return;
    std::lock_guard<std::mutex> lk(g_mu);
    g_free_slots.push_back(s);
  }
};

int check(cudaError_t e) { return e == cudaSuccess ? GSR_OK : (int)e; }
int debug_sync(cudaStream_t st, int debug) {
  if (!debug) return GSR_OK;
  cudaError_t e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  return check(e);
}
// ---- optional per-stage CUDA-event timing (gsr_profile_*) --------------------------------------------------
// When enabled, every kernel group of a call is bracketed by two events on the caller's stream; nothing is
// synchronised until gsr_profile_collect. Used by bench.py to time the dominant kernel inside the timed region.
struct StageSpan { int stage; cudaEvent_t a, b; };
std::mutex g_prof_mu;
bool g_prof_on = false;
unsigned g_prof_mask = 0;   // stages to time (bit = GSR_STAGE_*); 0 = all
std::vector<StageSpan> g_spans;
std::vector<cudaEvent_t> g_free_events;
cudaEvent_t prof_event() {
  cudaEvent_t e = nullptr;
  if (!g_free_events.empty()) { e = g_free_events.back(); g_free_events.pop_back(); }
  else cudaEventCreate(&e);
  return e;
}
struct StageTimer {
  cudaStream_t st; int idx = -1;
  StageTimer(cudaStream_t s, int stage) : st(s) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!g_prof_on || (g_prof_mask && !(g_prof_mask >> stage & 1u))) return;
    StageSpan sp{stage, prof_event(), prof_event()};
    cudaEventRecord(sp.a, st);
    g_spans.push_back(sp);
    idx = (int)g_spans.size() - 1;
  }
  ~StageTimer() {
    if (idx < 0) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (idx < (int)g_spans.size()) cudaEventRecord(g_spans[idx].b, st);
  }
};

#define GSR_TRY(expr)                 \
  do {                                \
    int _rc = (expr);                 \
    if (_rc != GSR_OK) return _rc;    \
  } while (0)

}  // namespace

extern "C" {

const char* gsr_version(void) { return "gsr_b200 0.1 (sm_100a)"; }

size_t gsr_geom_bytes(int P, int width, int height) { return carve_geom(nullptr, P, width, height, nullptr); }
size_t gsr_image_bytes(int width, int height) { return carve_image(nullptr, width, height, nullptr); }
size_t gsr_point_list_bytes(int64_t R) { return align_up(4 * (size_t)(R > 0 ? R : 1)); }
size_t gsr_binning_bytes(int64_t R) {  // the larger of the radix-sort and the two-level layouts (any image size)
  const size_t a = carve_binning(nullptr, R, nullptr), b = carve_refine(nullptr, R, (size_t)TS_MAX_TILES, nullptr);
  return a > b ? a : b;
}
size_t gsr_backward_scratch_bytes(int P) { return carve_backward(nullptr, P, nullptr); }
size_t gsr_knn_bytes(int P) { return knn_workspace_bytes(P); }

// ---- forward building blocks (enqueue only; no host synchronisation) ---------------------------------------
static int check_forward_args(int P, int sh_degree, int sh_coeffs, int width, int height, const float* means3D,
                              const float* shs, const float* colors_precomp, const float* opacities,
                              const float* scales, const float* rotations, const float* cov3D_precomp,
                              const float* viewmatrix, const float* projmatrix, const float* campos,
                              const int* radii, const void* geom_ws) {
  if (!means3D || !opacities || !viewmatrix || !projmatrix || !radii || !geom_ws) return GSR_EINVAL;
  if (!shs && !colors_precomp) return GSR_ENOCOLOR;       // rasterizer_impl.cu:245-248
  if (shs && !colors_precomp && (!campos || sh_coeffs < (sh_degree + 1) * (sh_degree + 1) || sh_degree > 3))
    return GSR_EINVAL;
  if (!cov3D_precomp && (!scales || !rotations)) return GSR_ENOCOV;
  if (num_tiles(width, height) > 65536) return GSR_EINVAL;  // tile id must fit two 8-bit digits
  if ((width + TILE_X - 1) / TILE_X > 65535 || (height + TILE_Y - 1) / TILE_Y > 65535) return GSR_EINVAL;
  return GSR_OK;
}

// preprocess + depth sort of the Gaussians + rectangles into depth order + scan -> g.total (device)
static int enqueue_plan(cudaStream_t st, int P, int sh_degree, int sh_coeffs, int width, int height,
                        const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
                        const float* scales, float scale_modifier, const float* rotations,
                        const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                        const float* campos, float tan_fovx, float tan_fovy, int prefiltered, int* radii,
                        char* geom_ws, const GeomViews& g, int debug, const ImageViews* im = nullptr) {
  // im (the single-call forwards): the forward renderer's tile order is sorted by an extra CTA of the chunk scan --
  // one launch less than a separate sort kernel in front of the renderer
  GSR_TRY(check(cudaMemsetAsync(geom_ws + g.zero_begin, 0, g.zero_end - g.zero_begin, st)));
  {
    StageTimer tm(st, GSR_STAGE_PREPROCESS);
    GSR_TRY(check(launch_preprocess_fwd(st, P, sh_degree, sh_coeffs, means3D, colors_precomp ? nullptr : shs,
                                        colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp,
                                        viewmatrix, projmatrix, campos, width, height, tan_fovx, tan_fovy,
                                        prefiltered, radii, g)));
  }
  GSR_TRY(debug_sync(st, debug));
  {  // Gaussians into (depth bits, id) order: 4 stable passes, result back in dkeys_a / order_a
    StageTimer tm(st, GSR_STAGE_DEPTH_SORT);
    const SortPlan plan = make_sort_plan(32);
    GSR_TRY(check(sort_pairs_u32_prehist(st, P, nullptr, g.dkeys_a, g.dkeys_b, g.order_a, g.order_b, plan, g.hist,
                                         g.tickets, g.lookback, nullptr)));
  }
  GSR_TRY(debug_sync(st, debug));
  {
    StageTimer tm(st, GSR_STAGE_SCAN);
    GSR_TRY(check(launch_chunk_scan(st, P, g.order_a, g, (int)num_tiles(width, height), im ? im->tile_work : nullptr,
                                    im ? im->tile_order_fwd : nullptr)));
  }
  return GSR_OK;
}

static SortPlan tile_sort_plan(int W, int H) { return make_sort_plan(higher_msb((uint32_t)num_tiles(W, H))); }

// instance emit + tile sort + ranges + render. `cap` bounds the instances written; the actual count is read
// from n_dev (device) when given, else it IS cap.
// Binning implementation: 2 = two-level sort-free binning (tile_scatter.cu on 4x4-tile bins + tile_refine.cu; the
// default for images of up to TS_MAX_TILES bins), 1 = single-level sort-free binning (<= TS_MAX_TILES tiles),
// 0 = instance radix sort (any size).  GSR_BINNING=radix / single / twolevel forces a mode for every image size that admits it
// (A/B measurements); gsr_debug_set_binning overrides it at run time (tests compare the implementations in one
// process).
enum { BIN_RADIX = 0, BIN_SINGLE = 1, BIN_TWO_LEVEL = 2 };
static std::atomic<int> g_binning_mode{-1};  // -1: not decided yet, 0: automatic, 1: radix, 2: single-level, 3: two-level
// `cap` = the list capacity of the call (num_rendered, or the caller's bound on it): the two-level path has six short
// launches, the single-level one three, so small lists (tracking-sized maps) stay on the single level where the
// image admits it (measured: C2 100 k Gaussians 38 vs 52 us, headline 96 vs 91 us, C4 2 M Gaussians 282 vs 166 us).
constexpr long long TWO_LEVEL_MIN_INSTANCES = 2ll << 20;
static int binning_impl(int width, int height, long long cap = -1) {
  int mode = g_binning_mode.load();
  if (mode < 0) {
    const char* e = getenv("GSR_BINNING");
    mode = (e && strcmp(e, "radix") == 0) ? 1 : (e && strcmp(e, "single") == 0) ? 2
         : (e && strcmp(e, "twolevel") == 0) ? 3 : 0;
    g_binning_mode.store(mode);
  }
  if (mode == 1) return BIN_RADIX;
  const bool single_ok = num_tiles(width, height) <= (size_t)TS_MAX_TILES;
  if (mode == 2 && single_ok) return BIN_SINGLE;
  if (mode == 0 && single_ok && cap >= 0 && cap < TWO_LEVEL_MIN_INSTANCES) return BIN_SINGLE;
  return num_bins(width, height) <= (size_t)TS_MAX_TILES ? BIN_TWO_LEVEL : BIN_RADIX;
}
static bool binning_ws_needed(int width, int height, long long cap = -1) {
  return binning_impl(width, height, cap) != BIN_SINGLE;
}

static int enqueue_render(cudaStream_t st, int P, int width, int height, const float* background,
                          const GeomViews& g, const ImageViews& im, char* binning_ws, uint32_t* point_list,
                          long long cap, const unsigned long long* n_dev, float* out_color, float* out_depth,
                          float* out_opacity, int* n_touched, int debug, bool tile_order_ready = false) {
  const int tiles_x = (width + TILE_X - 1) / TILE_X;
  const size_t n_tiles = num_tiles(width, height);
  const int impl = binning_impl(width, height, cap);
  if (impl == BIN_TWO_LEVEL) {
    // level 1 (tile_scatter.cu on bins of 4x4 tiles) + level 2 (tile_refine.cu)
    const int bins_x = (tiles_x + TS2_CB - 1) / TS2_CB, n_bins = (int)num_bins(width, height);
    RefineViews rv;
    carve_refine(binning_ws, cap, (size_t)n_bins, &rv);
    {
      StageTimer tm(st, GSR_STAGE_EMIT);
      GSR_TRY(check(launch_bin_count(st, P, g, bins_x, n_bins)));
      GSR_TRY(check(launch_bin_offsets_scatter(st, P, g, bins_x, n_bins, cap, rv)));
    }
    GSR_TRY(debug_sync(st, debug));
    {
      StageTimer tm(st, GSR_STAGE_TILE_SORT);
      GSR_TRY(check(launch_refine(st, g, rv, width, height, cap, im.ranges, point_list)));
    }
    GSR_TRY(debug_sync(st, debug));
  } else if (impl == BIN_SINGLE) {
    // sort-free binning (tile_scatter.cu): count matrix -> column scans -> in-order scatter
    {
      StageTimer tm(st, GSR_STAGE_EMIT);
      GSR_TRY(check(launch_tile_count(st, P, g, tiles_x, (int)n_tiles)));
    }
    GSR_TRY(debug_sync(st, debug));
    {
      StageTimer tm(st, GSR_STAGE_TILE_SORT);
      GSR_TRY(check(launch_tile_offsets_scatter(st, P, g, tiles_x, (int)n_tiles, cap, im.ranges, point_list)));
    }
    GSR_TRY(debug_sync(st, debug));
  } else if (cap > 0) {
    BinningViews bv;
    carve_binning(binning_ws, cap, &bv);
    const SortPlan plan = tile_sort_plan(width, height);
    GSR_TRY(check(cudaMemsetAsync(binning_ws + bv.zero_begin, 0, bv.zero_end(plan.n_passes) - bv.zero_begin, st)));
    {
      StageTimer tm(st, GSR_STAGE_EMIT);
      GSR_TRY(check(launch_emit_instances(st, P, g.order_a, g, tiles_x, cap, plan, bv.tkeys_a, bv.vals_a, bv.hist)));
    }
    GSR_TRY(debug_sync(st, debug));
    {
      StageTimer tm(st, GSR_STAGE_TILE_SORT);
      GSR_TRY(check(sort_pairs_u32_prehist(st, cap, n_dev, bv.tkeys_a, bv.tkeys_b, bv.vals_a, bv.vals_b, plan,
                                           bv.hist, bv.tickets, bv.lookback, point_list)));
    }
    GSR_TRY(debug_sync(st, debug));
    {
      StageTimer tm(st, GSR_STAGE_RANGES);
      GSR_TRY(check(launch_tile_ranges(st, cap, n_dev, (plan.n_passes & 1) ? bv.tkeys_b : bv.tkeys_a, im.ranges,
                                       num_tiles(width, height))));
    }
  } else {
    GSR_TRY(check(launch_tile_ranges(st, 0, nullptr, nullptr, im.ranges, num_tiles(width, height))));
  }
  {
    StageTimer tm(st, GSR_STAGE_RENDER_FWD);
    GSR_TRY(check(launch_render_fwd(st, width, height, im, point_list, g, background, out_color, out_depth,
                                    out_opacity, n_touched, tile_order_ready)));
  }
  return debug_sync(st, debug);
}

// num_rendered + prefilter flag -> the slot's pinned words; `finish` waits for the two copies on the slot's event.
static int enqueue_total_read(cudaStream_t st, const GeomViews& g, const ReadSlot& rs) {
  GSR_TRY(check(cudaMemcpyAsync(&rs.pinned->total, g.total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st)));
  GSR_TRY(check(cudaMemcpyAsync(&rs.pinned->error_flag, g.error_flag, sizeof(int), cudaMemcpyDeviceToHost, st)));
  return check(cudaEventRecord(rs.ev, st));
}
static int finish_total_read(const ReadSlot& rs, int64_t* num_rendered) {
  GSR_TRY(check(cudaEventSynchronize(rs.ev)));
  const HostSlot slot = *rs.pinned;
  if (slot.error_flag) return GSR_EPREFILTER;
  if (slot.total >= (1ull << 30)) return GSR_EOVERFLOW;
  *num_rendered = (int64_t)slot.total;
  return GSR_OK;
}

int gsr_forward_plan(void* stream, int P, int sh_degree, int sh_coeffs, int width, int height,
                     const float* means3D, const float* shs, const float* colors_precomp,
                     const float* opacities, const float* scales, float scale_modifier,
                     const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                     const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                     int prefiltered, int* radii, void* geom_ws, int64_t* num_rendered, int debug) {
  if (!num_rendered) return GSR_EINVAL;
  *num_rendered = 0;
  if (P < 0 || width <= 0 || height <= 0) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  GSR_TRY(check_forward_args(P, sh_degree, sh_coeffs, width, height, means3D, shs, colors_precomp, opacities, scales,
                             rotations, cov3D_precomp, viewmatrix, projmatrix, campos, radii, geom_ws));
  cudaStream_t st = (cudaStream_t)stream;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  GSR_TRY(enqueue_plan(st, P, sh_degree, sh_coeffs, width, height, means3D, shs, colors_precomp, opacities, scales,
                       scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                       prefiltered, radii, (char*)geom_ws, g, debug));
  // the one host read of the forward (rasterizer_impl.cu:284)
  ReadSlotGuard rs;
  GSR_TRY(acquire_read_slot(&rs.s));
  GSR_TRY(enqueue_total_read(st, g, rs.s));
  return finish_total_read(rs.s, num_rendered);
}

int gsr_forward_render(void* stream, int P, int width, int height, const float* background, const int* radii,
                       const void* geom_ws, void* image_ws, void* binning_ws, uint32_t* point_list,
                       int64_t num_rendered, float* out_color, float* out_depth, float* out_opacity,
                       int* n_touched, int debug) {
  if (P < 0 || width <= 0 || height <= 0 || num_rendered < 0) return GSR_EINVAL;
  if (P == 0) return GSR_OK;  // reference returns its zero-filled outputs (rasterize_points.cu:85)
  if (!background || !radii || !geom_ws || !image_ws || !out_color || !out_depth || !out_opacity || !n_touched)
    return GSR_EINVAL;
  if (num_rendered > 0 && (!point_list || (!binning_ws && binning_ws_needed(width, height, num_rendered)))) return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  ImageViews im;
  carve_image((char*)image_ws, width, height, &im);
  return enqueue_render(st, P, width, height, background, g, im, (char*)binning_ws, point_list, num_rendered,
                        nullptr, out_color, out_depth, out_opacity, n_touched, debug);
}

int gsr_forward_enqueue(void* stream, int P, int sh_degree, int sh_coeffs, int width, int height,
                        const float* background, const float* means3D, const float* shs,
                        const float* colors_precomp, const float* opacities, const float* scales,
                        float scale_modifier, const float* rotations, const float* cov3D_precomp,
                        const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                        float tan_fovy, int prefiltered, int* radii, void* geom_ws, void* image_ws, void* binning_ws,
                        uint32_t* point_list, int64_t capacity, float* out_color, float* out_depth,
                        float* out_opacity, int* n_touched, uint64_t* total_out) {
  if (P <= 0 || width <= 0 || height <= 0 || capacity <= 0 || capacity >= (1ll << 30)) return GSR_EINVAL;
  GSR_TRY(check_forward_args(P, sh_degree, sh_coeffs, width, height, means3D, shs, colors_precomp, opacities, scales,
                             rotations, cov3D_precomp, viewmatrix, projmatrix, campos, radii, geom_ws));
  if (!background || !image_ws || !out_color || !out_depth || !out_opacity || !n_touched || !point_list)
    return GSR_EINVAL;
  if (!binning_ws && binning_ws_needed(width, height, capacity)) return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  ImageViews im;
  carve_image((char*)image_ws, width, height, &im);
  GSR_TRY(enqueue_plan(st, P, sh_degree, sh_coeffs, width, height, means3D, shs, colors_precomp, opacities, scales,
                       scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                       prefiltered, radii, (char*)geom_ws, g, 0, &im));
  if (total_out)
    GSR_TRY(check(cudaMemcpyAsync(total_out, g.total, sizeof(unsigned long long), cudaMemcpyDefault, st)));
  return enqueue_render(st, P, width, height, background, g, im, (char*)binning_ws, point_list, capacity, g.total,
                        out_color, out_depth, out_opacity, n_touched, 0, true);
}

int gsr_forward(void* stream, int P, int sh_degree, int sh_coeffs, int width, int height, const float* background,
                const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
                const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                float tan_fovy, int prefiltered, int* radii, void* geom_ws, void* image_ws, void* binning_ws,
                uint32_t* point_list, int64_t capacity, float* out_color, float* out_depth, float* out_opacity,
                int* n_touched, int64_t* num_rendered, int debug) {
  if (!num_rendered) return GSR_EINVAL;
  *num_rendered = 0;
  if (P < 0 || width <= 0 || height <= 0 || capacity < 0 || capacity >= (1ll << 30)) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  GSR_TRY(check_forward_args(P, sh_degree, sh_coeffs, width, height, means3D, shs, colors_precomp, opacities, scales,
                             rotations, cov3D_precomp, viewmatrix, projmatrix, campos, radii, geom_ws));
  if (!background || !image_ws || !out_color || !out_depth || !out_opacity || !n_touched) return GSR_EINVAL;
  if (capacity > 0 && (!point_list || (!binning_ws && binning_ws_needed(width, height, capacity)))) return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  ImageViews im;
  carve_image((char*)image_ws, width, height, &im);
  GSR_TRY(enqueue_plan(st, P, sh_degree, sh_coeffs, width, height, means3D, shs, colors_precomp, opacities, scales,
                       scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                       prefiltered, radii, (char*)geom_ws, g, debug, &im));
  ReadSlotGuard rs;
  GSR_TRY(acquire_read_slot(&rs.s));
  GSR_TRY(enqueue_total_read(st, g, rs.s));
  // everything downstream is enqueued BEFORE the host waits for num_rendered: the GPU never idles on the read
  GSR_TRY(enqueue_render(st, P, width, height, background, g, im, (char*)binning_ws, point_list, capacity, g.total,
                         out_color, out_depth, out_opacity, n_touched, debug, true));
  GSR_TRY(finish_total_read(rs.s, num_rendered));
  return *num_rendered > capacity ? GSR_ECAPACITY : GSR_OK;
}

static int backward_impl(void* stream, int P, int sh_degree, int sh_coeffs, int width, int height,
                         const float* background, const float* means3D, const float* shs,
                         const float* colors_precomp, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* projmatrix_raw, const float* campos, float tan_fovx,
                         float tan_fovy, const int* radii, const void* geom_ws, const void* image_ws,
                         const uint32_t* point_list, int64_t num_rendered, const float* dL_dout_color,
                         const float* dL_dout_depth, const LossView& loss, void* scratch, float* dL_dmeans2D,
                         float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                         float* dL_dsh, float* dL_dscales, float* dL_drotations, float* dL_dtau_per_gaussian,
                         float* dL_dtau, int debug) {
  if (P < 0 || width <= 0 || height <= 0 || num_rendered < 0) return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (P == 0) {
    if (dL_dtau) GSR_TRY(check(cudaMemsetAsync(dL_dtau, 0, 6 * sizeof(float), st)));
    return GSR_OK;
  }
  if (!background || !means3D || !viewmatrix || !projmatrix || !projmatrix_raw || !radii || !geom_ws ||
      !image_ws || !scratch)
    return GSR_EINVAL;
  if (!cov3D_precomp && (!scales || !rotations)) return GSR_ENOCOV;
  if (num_rendered > 0 && !point_list) return GSR_EINVAL;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  ImageViews im;
  carve_image((char*)image_ws, width, height, &im);
  BackwardViews bw;
  carve_backward((char*)scratch, P, &bw);
  // the partial-gradient scratch is zeroed by the backward renderer's prologue kernel (alongside its tile ordering)
  const size_t zero_bytes = bw.zero_bytes;
  if (num_rendered > 0) {
    StageTimer tm(st, GSR_STAGE_RENDER_BWD);
    GSR_TRY(check(launch_render_bwd(st, width, height, im, point_list, g, background, dL_dout_color,
                                    dL_dout_depth, loss, bw.packed, bw.touched_flag, zero_bytes)));
    GSR_TRY(debug_sync(st, debug));
  } else {
    GSR_TRY(check(cudaMemsetAsync(bw.packed, 0, zero_bytes, st)));
  }
  const float* shs_used = colors_precomp ? nullptr : shs;
  StageTimer tm_pb(st, GSR_STAGE_PREPROCESS_BWD);
  GSR_TRY(check(launch_preprocess_bwd(st, P, sh_degree, sh_coeffs, means3D, radii, shs_used, g.clamped, scales,
                                      rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                                      projmatrix_raw, campos, width, height, tan_fovx, tan_fovy, bw, dL_dmeans2D,
                                      dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
                                      dL_drotations, dL_dtau_per_gaussian, dL_dtau, 0, P, nullptr, 0)));
  return debug_sync(st, debug);
}

static LossView no_loss() {
  LossView l{};
  l.kind = GSR_LOSS_NONE;
  return l;
}

int gsr_backward(void* stream, int P, int sh_degree, int sh_coeffs, int width, int height,
                 const float* background, const float* means3D, const float* shs, const float* colors_precomp,
                 const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                 const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw,
                 const float* campos, float tan_fovx, float tan_fovy, const int* radii, const void* geom_ws,
                 const void* image_ws, const uint32_t* point_list, int64_t num_rendered,
                 const float* dL_dout_color, const float* dL_dout_depth, void* scratch, float* dL_dmeans2D,
                 float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscales, float* dL_drotations, float* dL_dtau_per_gaussian, float* dL_dtau,
                 int debug) {
  return backward_impl(stream, P, sh_degree, sh_coeffs, width, height, background, means3D, shs, colors_precomp,
                       scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw,
                       campos, tan_fovx, tan_fovy, radii, geom_ws, image_ws, point_list, num_rendered,
                       dL_dout_color, dL_dout_depth, no_loss(), scratch, dL_dmeans2D, dL_dcolors, dL_dopacity,
                       dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, dL_dtau_per_gaussian, dL_dtau,
                       debug);
}

int gsr_l1_loss_forward(void* stream, int kind, int width, int height, const float* color, const float* depth,
                        const float* opacity, const float* gt_rgb, const float* gt_depth, const uint8_t* mask,
                        const uint8_t* grad_mask, const float* exposure_a, const float* exposure_b,
                        float lambda_rgb, void* scratch, float* out) {
  if ((kind != GSR_LOSS_MAPPING && kind != GSR_LOSS_TRACKING) || width <= 0 || height <= 0) return GSR_EINVAL;
  if (!color || !gt_rgb || !scratch || !out) return GSR_EINVAL;
  if (gt_depth && !depth) return GSR_EINVAL;
  if (kind == GSR_LOSS_TRACKING && !opacity) return GSR_EINVAL;
  return check(launch_l1_loss((cudaStream_t)stream, kind, width, height, color, depth, opacity, gt_rgb, gt_depth,
                              mask, grad_mask, exposure_a, exposure_b, lambda_rgb, (double*)scratch, out));
}

static int to_loss_view(const gsr_l1_loss* loss, LossView* lv) {
  if (!loss || (loss->kind != GSR_LOSS_MAPPING && loss->kind != GSR_LOSS_TRACKING)) return GSR_EINVAL;
  if (!loss->color || !loss->gt_rgb || !loss->out || (loss->gt_depth && !loss->depth)) return GSR_EINVAL;
  if (loss->kind == GSR_LOSS_TRACKING && !loss->opacity) return GSR_EINVAL;
  lv->kind = loss->kind;
  lv->color = loss->color; lv->depth = loss->depth; lv->opacity = loss->opacity;
  lv->gt_rgb = loss->gt_rgb; lv->gt_depth = loss->gt_depth;
  lv->mask = loss->mask; lv->grad_mask = loss->grad_mask;
  lv->exposure_a = loss->exposure_a; lv->exposure_b = loss->exposure_b;
  lv->out = loss->out; lv->upstream = loss->upstream;
  return GSR_OK;
}

int gsr_l1_loss_grad(void* stream, int width, int height, const gsr_l1_loss* loss, float* dL_dcolor,
                     float* dL_ddepth) {
  if (width <= 0 || height <= 0 || (!dL_dcolor && !dL_ddepth)) return GSR_EINVAL;
  LossView lv{};
  GSR_TRY(to_loss_view(loss, &lv));
  return check(launch_l1_loss_grad((cudaStream_t)stream, width, height, lv, dL_dcolor, dL_ddepth));
}

int gsr_backward_l1(void* stream, int P, int sh_degree, int sh_coeffs, int width, int height,
                    const float* background, const float* means3D, const float* shs, const float* colors_precomp,
                    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                    const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw,
                    const float* campos, float tan_fovx, float tan_fovy, const int* radii, const void* geom_ws,
                    const void* image_ws, const uint32_t* point_list, int64_t num_rendered,
                    const gsr_l1_loss* loss, void* scratch, float* dL_dmeans2D, float* dL_dcolors,
                    float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                    float* dL_drotations, float* dL_dtau_per_gaussian, float* dL_dtau, int debug) {
  LossView lv{};
  GSR_TRY(to_loss_view(loss, &lv));
  return backward_impl(stream, P, sh_degree, sh_coeffs, width, height, background, means3D, shs, colors_precomp,
                       scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw,
                       campos, tan_fovx, tan_fovy, radii, geom_ws, image_ws, point_list, num_rendered, nullptr,
                       nullptr, lv, scratch, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
                       dL_dscales, dL_drotations, dL_dtau_per_gaussian, dL_dtau, debug);
}

static int backward_render_impl(void* stream, int P, int width, int height, const float* background,
                                const void* geom_ws, const void* image_ws, const uint32_t* point_list,
                                int64_t num_rendered, const float* dL_dout_color, const float* dL_dout_depth,
                                const LossView& loss, void* scratch, int debug);

int gsr_backward_render(void* stream, int P, int width, int height, const float* background, const void* geom_ws,
                        const void* image_ws, const uint32_t* point_list, int64_t num_rendered,
                        const float* dL_dout_color, const float* dL_dout_depth, void* scratch, int debug) {
  return backward_render_impl(stream, P, width, height, background, geom_ws, image_ws, point_list, num_rendered,
                              dL_dout_color, dL_dout_depth, no_loss(), scratch, debug);
}

int gsr_backward_render_l1(void* stream, int P, int width, int height, const float* background, const void* geom_ws,
                           const void* image_ws, const uint32_t* point_list, int64_t num_rendered,
                           const gsr_l1_loss* loss, void* scratch, int debug) {
  LossView lv{};
  GSR_TRY(to_loss_view(loss, &lv));
  return backward_render_impl(stream, P, width, height, background, geom_ws, image_ws, point_list, num_rendered,
                              nullptr, nullptr, lv, scratch, debug);
}

static int backward_render_impl(void* stream, int P, int width, int height, const float* background,
                                const void* geom_ws, const void* image_ws, const uint32_t* point_list,
                                int64_t num_rendered, const float* dL_dout_color, const float* dL_dout_depth,
                                const LossView& loss, void* scratch, int debug) {
  if (P < 0 || width <= 0 || height <= 0 || num_rendered < 0) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  if (!background || !geom_ws || !image_ws || !scratch) return GSR_EINVAL;
  if (num_rendered > 0 && !point_list) return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  ImageViews im;
  carve_image((char*)image_ws, width, height, &im);
  BackwardViews bw;
  carve_backward((char*)scratch, P, &bw);
  const size_t zero_bytes = bw.zero_bytes;
  if (num_rendered > 0) {
    StageTimer tm(st, GSR_STAGE_RENDER_BWD);
    GSR_TRY(check(launch_render_bwd(st, width, height, im, point_list, g, background, dL_dout_color,
                                    dL_dout_depth, loss, bw.packed, bw.touched_flag, zero_bytes)));
  } else {
    GSR_TRY(check(cudaMemsetAsync(bw.packed, 0, zero_bytes, st)));
  }
  return debug_sync(st, debug);
}

int gsr_backward_rows(void* stream, int P, int p_begin, int p_end, int width, int height, const float* means3D,
                      const float* scales, float scale_modifier, const float* rotations, const float* viewmatrix,
                      const float* projmatrix, const float* projmatrix_raw, float tan_fovx, float tan_fovy,
                      const int* radii, const void* geom_ws, void* scratch, float* rows, int accumulate,
                      float* dL_dmeans2D, float* dL_dtau, int debug) {
  return gsr_backward_rows_marked(stream, P, p_begin, p_end, width, height, means3D, scales, scale_modifier,
                                  rotations, viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy, radii,
                                  geom_ws, scratch, rows, accumulate, dL_dmeans2D, dL_dtau, nullptr, nullptr, debug);
}

size_t gsr_rows_exchange_bytes(int P, int world) {
  if (P < 0 || world < 1 || world > XCHG_MAX_WORLD) return 0;
  return xchg_layout(P, world).total;
}

int gsr_rows_exchange_layout(int P, int world, size_t* offsets, int* words_per_rank, int* vis_bytes_per_rank) {
  if (P < 0 || world < 1 || world > XCHG_MAX_WORLD || !offsets) return GSR_EINVAL;
  const XchgLayout L = xchg_layout(P, world);
  offsets[0] = L.rows; offsets[1] = L.bitmap; offsets[2] = L.vis; offsets[3] = L.flags;
  if (words_per_rank) *words_per_rank = L.words_pad;
  if (vis_bytes_per_rank) *vis_bytes_per_rank = L.p_pad;
  return GSR_OK;
}

int gsr_rows_exchange(void* stream, int P, int capacity, int rank, int world, int n_ctas, void* local_buf,
                      void* multicast_buf, const uint64_t* peer_bufs, uint32_t epoch, float* vis_count, int* status) {
  if (P < 0 || capacity < P || world < 1 || world > XCHG_MAX_WORLD || rank < 0 || rank >= world) return GSR_EINVAL;
  if (n_ctas < 1 || n_ctas > XCHG_MAX_CTAS || !local_buf || !multicast_buf || !peer_bufs) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  return check(launch_rows_exchange((cudaStream_t)stream, P, capacity, rank, world, n_ctas, local_buf, multicast_buf,
                                    reinterpret_cast<const unsigned long long*>(peer_bufs), epoch, vis_count,
                                    status));
}

int gsr_debug_exchange_trace(uint64_t* host_out) {
  if (!host_out) return GSR_EINVAL;
  return check(read_exchange_trace(reinterpret_cast<unsigned long long*>(host_out)));
}

int gsr_backward_rows_marked(void* stream, int P, int p_begin, int p_end, int width, int height,
                             const float* means3D, const float* scales, float scale_modifier,
                             const float* rotations, const float* viewmatrix, const float* projmatrix,
                             const float* projmatrix_raw, float tan_fovx, float tan_fovy, const int* radii,
                             const void* geom_ws, void* scratch, float* rows, int accumulate, float* dL_dmeans2D,
                             float* dL_dtau, uint32_t* touched_words, uint8_t* vis_u8, int debug) {
  if ((touched_words == nullptr) != (vis_u8 == nullptr)) return GSR_EINVAL;
  if (P < 0 || width <= 0 || height <= 0 || p_begin < 0 || p_end < p_begin || p_end > P) return GSR_EINVAL;
  if (p_begin % 256 != 0) return GSR_EINVAL;  // ranges start on a 256-Gaussian boundary (tau block partials)
  cudaStream_t st = (cudaStream_t)stream;
  if (P == 0) {
    if (dL_dtau) GSR_TRY(check(cudaMemsetAsync(dL_dtau, 0, 6 * sizeof(float), st)));
    return GSR_OK;
  }
  if (!means3D || !scales || !rotations || !viewmatrix || !projmatrix || !projmatrix_raw || !radii || !geom_ws ||
      !scratch || !rows)
    return GSR_EINVAL;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  BackwardViews bw;
  carve_backward((char*)scratch, P, &bw);
  StageTimer tm_pb(st, GSR_STAGE_PREPROCESS_BWD);
  GSR_TRY(check(launch_preprocess_bwd(st, P, 0, 0, means3D, radii, nullptr, g.clamped, scales, rotations,
                                      scale_modifier, nullptr, viewmatrix, projmatrix, projmatrix_raw, nullptr, width,
                                      height, tan_fovx, tan_fovy, bw, dL_dmeans2D, nullptr, nullptr, nullptr, nullptr,
                                      nullptr, nullptr, nullptr, nullptr, dL_dtau, p_begin, p_end, rows, accumulate,
                                      touched_words, vis_u8)));
  return debug_sync(st, debug);
}

size_t gsr_insert_ws_bytes(int width, int height) {
  if (width <= 0 || height <= 0) return 0;
  return INSERT_SELECT_WS_BYTES + align_up((size_t)((width * (size_t)height + 1023) / 1024) * 4);
}

int gsr_select_kth_f32(void* stream, int64_t n, const float* a, const float* b, int64_t k, float* out, void* ws) {
  if (n <= 0 || n >= (int64_t)1 << 32 || k < 0 || k >= n || !a || !out || !ws) return GSR_EINVAL;
  return check(launch_select_kth((cudaStream_t)stream, n, a, b, k, out, (uint32_t*)ws));
}

int gsr_insert_select(void* stream, int width, int height, const float* gt_depth, const float* render_depth,
                      const float* render_opacity, int init, const float* median_err, int32_t* selected,
                      uint32_t* count, void* ws) {
  if (width <= 0 || height <= 0 || !gt_depth || !selected || !count || !ws) return GSR_EINVAL;
  if (!init && (!render_depth || !median_err)) return GSR_EINVAL;
  return check(launch_insert_select((cudaStream_t)stream, width, height, gt_depth, render_depth, render_opacity,
                                    init, median_err, selected, count,
                                    (uint32_t*)((char*)ws + INSERT_SELECT_WS_BYTES)));
}

int gsr_insert_points(void* stream, int K, const int32_t* selected, const int64_t* pick, int width, int height,
                      const float* rgb, const float* exposure_a, const float* exposure_b, const float* gt_depth,
                      const int32_t* segmentation, const float* R, const float* T, float fx, float fy, float cx,
                      float cy, float* xyz, float* features, int32_t* ids) {
  if (K < 0 || width <= 0 || height <= 0) return GSR_EINVAL;
  if (K == 0) return GSR_OK;
  if (!selected || !rgb || !gt_depth || !R || !T || !xyz || !features || fx == 0.f || fy == 0.f) return GSR_EINVAL;
  if ((exposure_a == nullptr) != (exposure_b == nullptr) || (ids && !segmentation)) return GSR_EINVAL;
  return check(launch_insert_points((cudaStream_t)stream, K, selected, reinterpret_cast<const long long*>(pick),
                                    width, height, rgb, exposure_a, exposure_b, gt_depth, segmentation, R, T, fx,
                                    fy, cx, cy, xyz, features, ids));
}

int gsr_insert_finish(void* stream, int K, const float* dist2, const float* median_depth, int scale_cols,
                      float* scales, float* rotations, float* opacities) {
  if (K < 0 || (scale_cols != 1 && scale_cols != 3)) return GSR_EINVAL;
  if (K == 0) return GSR_OK;
  if (!dist2 || !median_depth || !scales || !rotations || !opacities) return GSR_EINVAL;
  return check(launch_insert_finish((cudaStream_t)stream, K, dist2, median_depth, scale_cols, scales, rotations,
                                    opacities));
}

int gsr_mark_visible(void* stream, int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present) {
  (void)projmatrix;  // the reference computes p_proj but only tests p_view.z (auxiliary.h:149-154)
  if (P < 0) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  if (!means3D || !viewmatrix || !present) return GSR_EINVAL;
  return check(launch_mark_visible((cudaStream_t)stream, P, means3D, viewmatrix, present));
}

int gsr_camera_setup(void* stream, const float* R, int r_stride_row, int r_stride_col, const float* T,
                     int t_stride, const float* proj_T, float* viewmatrix, float* projmatrix, float* campos) {
  if (!R || !T || !proj_T || !viewmatrix || !projmatrix || !campos) return GSR_EINVAL;
  return check(launch_camera_setup((cudaStream_t)stream, R, r_stride_row, r_stride_col, T, t_stride, proj_T,
                                   viewmatrix, projmatrix, campos));
}

int gsr_densify_stats(void* stream, int P, const int* radii, const float* dL_dmeans2D, float* max_radii_2d,
                      float* xyz_gradient_accum, float* denom) {
  if (P < 0) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  if (!radii || (dL_dmeans2D && (!xyz_gradient_accum || !denom))) return GSR_EINVAL;
  return check(launch_densify_stats((cudaStream_t)stream, P, radii, dL_dmeans2D, max_radii_2d, xyz_gradient_accum,
                                    denom));
}

int gsr_visibility_prune(void* stream, int P, int K, const int* const* n_touched, const int* kf_idx, int min_kf,
                         int coviz, int* nr_obs, uint8_t* to_prune) {
  if (P < 0 || K < 0 || K > VIS_MAX_KEYFRAMES) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  if ((K > 0 && !n_touched) || (to_prune && !kf_idx)) return GSR_EINVAL;
  for (int k = 0; k < K; ++k)
    if (!n_touched[k]) return GSR_EINVAL;
  return check(launch_visibility_prune((cudaStream_t)stream, P, K, n_touched, kf_idx, min_kf, coviz, nr_obs,
                                       to_prune));
}

size_t gsr_surgery_ws_bytes(int64_t P) { return P > 0 ? surgery_ws_bytes(P) : 256; }

int gsr_rows_append(void* stream, int n_groups, float* const* param, float* const* exp_avg, float* const* exp_avg_sq,
                    const float* const* new_rows, const int* cols, int64_t P_old, int64_t K) {
  if (n_groups < 1 || n_groups > SURGERY_MAX_GROUPS || P_old < 0 || K < 0 || !param || !new_rows || !cols)
    return GSR_EINVAL;
  if (K == 0) return GSR_OK;
  for (int g = 0; g < n_groups; ++g)
    if (!param[g] || !new_rows[g] || cols[g] < 1) return GSR_EINVAL;
  return check(launch_rows_append((cudaStream_t)stream, n_groups, param, exp_avg, exp_avg_sq, new_rows, cols, P_old, K));
}

int gsr_rows_compact(void* stream, int n_arrays, const void* const* src, void* const* dst, const int* row_bytes,
                     int64_t P, const uint8_t* keep, uint32_t* new_count, void* ws) {
  if (n_arrays < 0 || n_arrays > SURGERY_MAX_ARRAYS || P < 0 || P >= ((int64_t)1 << 32) || !new_count || !ws)
    return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (P == 0) return check(cudaMemsetAsync(new_count, 0, sizeof(uint32_t), st));
  if (!keep || (n_arrays > 0 && (!src || !dst || !row_bytes))) return GSR_EINVAL;
  for (int j = 0; j < n_arrays; ++j)
    if (!src[j] || !dst[j] || src[j] == dst[j] || row_bytes[j] < 4 || row_bytes[j] % 4 != 0) return GSR_EINVAL;
  return check(launch_rows_compact(st, n_arrays, src, dst, row_bytes, P, keep, new_count, ws));
}

int gsr_knn_dist2(void* stream, int P, const float* points, float* mean_dist2, void* knn_ws) {
  if (P < 0) return GSR_EINVAL;
  if (P == 0) return GSR_OK;
  if (!points || !mean_dist2 || !knn_ws) return GSR_EINVAL;
  return check(launch_knn((cudaStream_t)stream, P, points, mean_dist2, knn_ws));
}

int gsr_sort_pairs_u64(void* stream, int64_t n, const uint64_t* keys_in, const uint32_t* values_in,
                       uint64_t* keys_out, uint32_t* values_out, int end_bit, void* ws) {
  if (n < 0 || end_bit <= 0 || end_bit > 64 || n >= (1ll << 30)) return GSR_EINVAL;
  if (n == 0) return GSR_OK;
  if (!keys_in || !values_in || !keys_out || !values_out || !ws) return GSR_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  BinningViews bv;
  carve_binning((char*)ws, n, &bv);
  const SortPlan plan = make_sort_plan(end_bit);
  const int n_passes = plan.n_passes;
  GSR_TRY(check(cudaMemsetAsync((char*)ws + bv.zero_begin, 0, bv.zero_end(n_passes) - bv.zero_begin, st)));
  GSR_TRY(check(cudaMemcpyAsync(bv.keys64_a, keys_in, 8 * (size_t)n, cudaMemcpyDeviceToDevice, st)));
  GSR_TRY(check(cudaMemcpyAsync(bv.vals_a, values_in, 4 * (size_t)n, cudaMemcpyDeviceToDevice, st)));
  GSR_TRY(check(launch_digit_hist_u64(st, bv.keys64_a, n, plan, bv.hist)));
  GSR_TRY(check(sort_pairs_u64_prehist(st, n, nullptr, bv.keys64_a, bv.keys64_b, bv.vals_a, bv.vals_b, plan, bv.hist,
                                       bv.tickets, bv.lookback, values_out)));
  const uint64_t* sorted = (n_passes & 1) ? bv.keys64_b : bv.keys64_a;
  return check(cudaMemcpyAsync(keys_out, sorted, 8 * (size_t)n, cudaMemcpyDeviceToDevice, st));
}

int gsr_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = (on & 1) != 0;
  g_prof_mask = (unsigned)on >> 8;   // optional stage selection: on = 1 | (1 << (8 + GSR_STAGE_x)) | ...
  return GSR_OK;
}

unsigned long long gsr_profile_kernel_launches(void) { return count_launch(0); }

int gsr_profile_collect(double* ms_sum, int* launches) {
  if (!ms_sum || !launches) return GSR_EINVAL;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (int i = 0; i < GSR_NUM_STAGES; ++i) { ms_sum[i] = 0.0; launches[i] = 0; }
  int rc = GSR_OK;
  for (auto& sp : g_spans) {
    float ms = 0.f;
    cudaError_t e = cudaEventSynchronize(sp.b);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, sp.a, sp.b);
    if (e != cudaSuccess) rc = (int)e;
    else if (sp.stage >= 0 && sp.stage < GSR_NUM_STAGES) { ms_sum[sp.stage] += ms; launches[sp.stage] += 1; }
    g_free_events.push_back(sp.a);
    g_free_events.push_back(sp.b);
  }
  g_spans.clear();
  return rc;
}

int gsr_debug_geom_views(const void* geom_ws, int P, int width, int height, const float** rec_a,
                         const float** rec_b, const float** rec_c, const uint32_t** tiles_touched) {
  if (!geom_ws) return GSR_EINVAL;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  // ONE packed record per Gaussian: element i of each view lives at view + 12 * i floats
  if (rec_a) *rec_a = (const float*)g.rec;
  if (rec_b) *rec_b = (const float*)(g.rec + 1);
  if (rec_c) *rec_c = (const float*)(g.rec + 2);
  if (tiles_touched) *tiles_touched = g.tiles_touched;
  return GSR_OK;
}

int gsr_debug_image_views(const void* image_ws, int width, int height, const float** final_T,
                          const uint32_t** n_contrib, const uint32_t** ranges) {
  if (!image_ws) return GSR_EINVAL;
  ImageViews im;
  carve_image((char*)image_ws, width, height, &im);
  if (final_T) *final_T = im.final_T;
  if (n_contrib) *n_contrib = im.n_contrib;
  if (ranges) *ranges = (const uint32_t*)im.ranges;
  return GSR_OK;
}

int gsr_binning_needed(int width, int height) {
  if (width <= 0 || height <= 0) return 0;
  return binning_ws_needed(width, height) ? 1 : 0;
}

int gsr_debug_set_binning(int mode) {
  if (mode < 0 || mode > 3) return GSR_EINVAL;
  g_binning_mode.store(mode);
  return GSR_OK;
}

int gsr_debug_sorted_keys(void* stream, const void* geom_ws, const void* binning_ws, const uint32_t* point_list,
                          int P, int64_t num_rendered, int width, int height, uint64_t* keys_sorted) {
  if (!geom_ws || !keys_sorted || !point_list) return GSR_EINVAL;
  GeomViews g;
  carve_geom((char*)geom_ws, P, width, height, &g);
  if (binning_impl(width, height) != BIN_RADIX)
    return check(launch_rebuild_keys_from_starts((cudaStream_t)stream, num_rendered, (int)num_tiles(width, height), g,
                                                 point_list, keys_sorted));
  if (!binning_ws) return GSR_EINVAL;
  BinningViews bv;
  carve_binning((char*)binning_ws, num_rendered, &bv);
  const uint32_t* tk = (tile_sort_plan(width, height).n_passes & 1) ? bv.tkeys_b : bv.tkeys_a;
  return check(launch_rebuild_keys((cudaStream_t)stream, num_rendered, tk, point_list, g.rec, keys_sorted));
}

}  // extern "C"
