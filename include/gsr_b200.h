/*
 * gsr_b200.h -- C ABI of the B200-native Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary for the one hot path of s-esposito/MonoGS:
 * the differentiable rasterizer with SE(3) pose gradients and simple-knn's
 * distCUDA2.  Every entry point below replaces one interface of the
 * reference (paths relative to /root/reference/submodules):
 *
 *   gsr_forward_plan + gsr_forward_render
 *        <- CudaRasterizer::Rasterizer::forward
 *           (diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:33-60,
 *            rasterizer_impl.cu:198-344), bound to Python by
 *            RasterizeGaussiansCUDA (rasterize_points.cu:35-122, ext.cpp:16)
 *   gsr_backward
 *        <- CudaRasterizer::Rasterizer::backward (rasterizer.h:62-92,
 *           rasterizer_impl.cu:348-455), bound by
 *           RasterizeGaussiansBackwardCUDA (rasterize_points.cu:124-211)
 *   gsr_mark_visible
 *        <- CudaRasterizer::Rasterizer::markVisible (rasterizer.h:26-31,
 *           rasterizer_impl.cu:141-153), bound by markVisible
 *           (rasterize_points.cu:213-232)
 *   gsr_knn_dist2
 *        <- SimpleKNN::knn (simple-knn/simple_knn.cu:185-221), bound by
 *           distCUDA2 (simple-knn/spatial.cu:15-26)
 *   gsr_sort_pairs_u64 / gsr_debug_*  -- test hooks for the bit-exact
 *           parity of keys, lists and ranges (rasterizer_impl.cu:70-138,306-311)
 *
 * The reference's std::function<char*(size_t)> resize callbacks
 * (rasterizer.h:34-36) become caller-allocated workspaces whose sizes are
 * returned by the gsr_*_bytes functions; the forward is split in two calls
 * around the one unavoidable host read of num_rendered
 * (rasterizer_impl.cu:284).
 *
 * Conventions: all data pointers are DEVICE pointers (fp32 / int32 row-major,
 * exactly the tensors the reference binding receives); a NULL pointer means
 * "absent" like the reference's empty tensors (forward.cu:207,243);
 * `stream` is a cudaStream_t passed as void*; every function returns 0 on
 * success, a negative GSR_E* code for bad arguments, or a positive
 * cudaError_t value.  No torch types appear here.
 */
#ifndef GSR_B200_H_
#define GSR_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_OK 0
#define GSR_EINVAL (-1)     /* bad argument combination / size            */
#define GSR_ENOCOLOR (-2)   /* neither shs nor colors_precomp              */
#define GSR_ENOCOV (-3)     /* neither scales+rotations nor cov3D_precomp  */
#define GSR_EPREFILTER (-4) /* prefiltered=1 but a point was culled        */
#define GSR_EOVERFLOW (-5)  /* num_rendered does not fit the index type    */
#define GSR_ECAPACITY (-6)  /* gsr_forward: num_rendered > capacity; re-run gsr_forward_render with *num_rendered */
#define GSR_ESTATICLIMIT (-7) /* a *_static entry point: the size exceeds the fixed scratch built into the library */

/* Library / build identification ("sm_100a", version). */
const char* gsr_version(void);

/* ---- workspace sizes (bytes). All workspaces need 256-byte alignment. -- */
/* Per-Gaussian state written by forward and re-read by backward. */
size_t gsr_geom_bytes(int P, int width, int height);
/* Per-pixel state (final T, n_contrib) + per-tile ranges and work counters; saved for backward. Re-using the same
 * image workspace for consecutive forwards of nearly the same view (tracking / mapping iterations) lets the forward
 * renderer schedule its heaviest tiles first from the previous call's per-tile work; whatever the workspace holds,
 * only the CTA schedule depends on it, never a result (a fresh, uninitialised workspace is fine). */
size_t gsr_image_bytes(int width, int height);
/* Sorted instance list (Gaussian ids in tile|depth order); saved for backward. */
size_t gsr_point_list_bytes(int64_t num_rendered);
/* Temporary key/value ping-pong buffers + radix-sort scratch; forward only. */
size_t gsr_binning_bytes(int64_t num_rendered);
/* Scratch for gsr_backward (packed per-Gaussian partial gradients, tau partials). */
size_t gsr_backward_scratch_bytes(int P);
/* Scratch for gsr_knn_dist2. */
size_t gsr_knn_bytes(int P);

/* ---- forward, phase 1: preprocess + tile-count scan ---------------------
 * Projects every Gaussian, fills radii[P] (int32) and the geometry state,
 * sorts the Gaussians by (view depth, id) and returns num_rendered = sum of
 * tiles touched through *num_rendered (host int64).  This call synchronises `stream` once (the reference does a
 * blocking cudaMemcpy at the same place). */
int gsr_forward_plan(void* stream, int P, int sh_degree, int sh_coeffs /*M*/,
                     int width, int height,
                     const float* means3D, const float* shs,
                     const float* colors_precomp, const float* opacities,
                     const float* scales, float scale_modifier,
                     const float* rotations, const float* cov3D_precomp,
                     const float* viewmatrix, const float* projmatrix,
                     const float* campos, float tan_fovx, float tan_fovy,
                     int prefiltered, int* radii, void* geom_ws,
                     int64_t* num_rendered, int debug);

/* ---- forward, phase 2: key emit + sort + ranges + render ----------------
 * out_color[3,H,W], out_depth[1,H,W], out_opacity[1,H,W] are fully written;
 * n_touched[P] (int32) must be zero on entry and is atomically incremented. */
int gsr_forward_render(void* stream, int P, int width, int height,
                       const float* background, const int* radii,
                       const void* geom_ws, void* image_ws, void* binning_ws,
                       uint32_t* point_list, int64_t num_rendered,
                       float* out_color, float* out_depth, float* out_opacity,
                       int* n_touched, int debug);

/* ---- forward, single call without a GPU bubble ---------------------------
 * Same work as gsr_forward_plan + gsr_forward_render, but the instance-level kernels are enqueued for a
 * caller-chosen `capacity` (binning_ws / point_list sized with gsr_binning_bytes(capacity) /
 * gsr_point_list_bytes(capacity)) and read the actual count from device memory, so the host wait for
 * num_rendered (rasterizer_impl.cu:284) overlaps the rest of the forward. Returns GSR_ECAPACITY when
 * *num_rendered > capacity: the per-Gaussian state in geom_ws is valid, the images are not -- allocate for
 * *num_rendered and call gsr_forward_render. n_touched must be zero on entry (and re-zeroed before that retry). */
int gsr_forward(void* stream, int P, int sh_degree, int sh_coeffs /*M*/, int width, int height,
                const float* background, const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities, const float* scales,
                float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos,
                float tan_fovx, float tan_fovy, int prefiltered, int* radii, void* geom_ws,
                void* image_ws, void* binning_ws, uint32_t* point_list, int64_t capacity,
                float* out_color, float* out_depth, float* out_opacity, int* n_touched,
                int64_t* num_rendered, int debug);

/* ---- forward without any host interaction (CUDA-graph capturable) --------------------------------------------
 * gsr_forward minus the host read of num_rendered: every kernel is enqueued for `capacity` instances and the call
 * returns at once; nothing in it synchronises, allocates or touches library-global host state, so it may be issued
 * while `stream` is being captured into a CUDA graph (monogs_b200/tracking.py captures a whole tracking iteration).
 * total_out (optional; pinned-host or device memory, 8 bytes) receives num_rendered through a stream-ordered copy:
 * when it exceeds `capacity` the images of that launch are invalid (lists were clipped, nothing went out of
 * bounds) and the caller re-runs with larger buffers. n_touched must be zero on entry. */
int gsr_forward_enqueue(void* stream, int P, int sh_degree, int sh_coeffs /*M*/, int width, int height,
                        const float* background, const float* means3D, const float* shs,
                        const float* colors_precomp, const float* opacities, const float* scales,
                        float scale_modifier, const float* rotations, const float* cov3D_precomp,
                        const float* viewmatrix, const float* projmatrix, const float* campos,
                        float tan_fovx, float tan_fovy, int prefiltered, int* radii, void* geom_ws,
                        void* image_ws, void* binning_ws, uint32_t* point_list, int64_t capacity,
                        float* out_color, float* out_depth, float* out_opacity, int* n_touched,
                        uint64_t* total_out);

/* ---- backward -----------------------------------------------------------
 * Outputs (all optional except where noted; NULL = not wanted):
 *   dL_dmeans2D[P,3], dL_dcolors[P,3], dL_dopacity[P,1], dL_dmeans3D[P,3],
 *   dL_dcov3D[P,6], dL_dsh[P,M,3], dL_dscales[P,3], dL_drotations[P,4],
 *   dL_dtau_per_gaussian[P,6] (the reference's raw output), and
 *   dL_dtau[6] = sum over Gaussians (rho = [0:3], theta = [3:6]), reduced
 *   in-kernel (replaces torch.sum in diff_gaussian_rasterization/__init__.py:152).
 * Every output given is fully written (no pre-zeroing needed).
 * dL_dout_color[3,H,W] / dL_dout_depth[1,H,W] may be NULL (treated as zero). */
int gsr_backward(void* stream, int P, int sh_degree, int sh_coeffs,
                 int width, int height, const float* background,
                 const float* means3D, const float* shs,
                 const float* colors_precomp, const float* scales,
                 float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* projmatrix_raw,
                 const float* campos, float tan_fovx, float tan_fovy,
                 const int* radii, const void* geom_ws, const void* image_ws,
                 const uint32_t* point_list, int64_t num_rendered,
                 const float* dL_dout_color, const float* dL_dout_depth,
                 void* scratch,
                 float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                 float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscales, float* dL_drotations,
                 float* dL_dtau_per_gaussian, float* dL_dtau, int debug);

/* ---- fused L1 photometric + depth loss (north star (d); the callers' losses, utils/slam_utils.py:58-146) ----
 * gsr_l1_loss_forward reduces the rendered images against the targets: out[8] (device) =
 *   { loss, l1_rgb, l1_depth, k_rgb, k_depth, dL/d exposure_a, dL/d exposure_b, mean opacity }.
 *   GSR_LOSS_MAPPING : lambda_rgb * mean_{mask,3ch}|e^a I + b - gt| + (1 - lambda_rgb) * mean_{gtD>0}|D - gtD|
 *                      (get_loss_mapping :101-146; lambda_rgb is its `lambda_depth` = 0.9)
 *   GSR_LOSS_TRACKING: 0.5 * mean(opacity) * mean_{3HW}(m|e^a I + b - gt|) + mean_{gtD>0, opacity>0.99}|D - gtD|,
 *                      m = mask * grad_mask * (opacity > 0.99)   (get_loss_tracking :58-98)
 *   mask / grad_mask: uint8 0/1 [H,W] or NULL (= all ones); exposure_a/b: device scalars or NULL (a = b = 0);
 *   gt_depth may be NULL (no depth term). scratch: 64 bytes of device memory.
 * gsr_backward_l1 is gsr_backward with dL/dpixel formed inside the backward tile renderer from the same
 * images (no gradient images are materialised): loss->out is the out[8] above, loss->upstream the device scalar
 * dL/dloss (NULL = 1). */
#define GSR_LOSS_NONE 0
#define GSR_LOSS_MAPPING 1
#define GSR_LOSS_TRACKING 2
typedef struct gsr_l1_loss {
  int kind;
  const float* color;      /* [3,H,W] rendered (forward output) */
  const float* depth;      /* [1,H,W] */
  const float* opacity;    /* [1,H,W] */
  const float* gt_rgb;     /* [3,H,W] */
  const float* gt_depth;   /* [1,H,W] or NULL */
  const uint8_t* mask;     /* [H,W] or NULL */
  const uint8_t* grad_mask;/* [H,W] or NULL (tracking) */
  const float* exposure_a; /* device scalar or NULL */
  const float* exposure_b; /* device scalar or NULL */
  const float* out;        /* out[8] of gsr_l1_loss_forward */
  const float* upstream;   /* device scalar dL/dloss or NULL */
} gsr_l1_loss;
int gsr_l1_loss_forward(void* stream, int kind, int width, int height, const float* color,
                        const float* depth, const float* opacity, const float* gt_rgb,
                        const float* gt_depth, const uint8_t* mask, const uint8_t* grad_mask,
                        const float* exposure_a, const float* exposure_b, float lambda_rgb,
                        void* scratch, float* out);
/* The gradient images gsr_backward_l1 forms on the fly, written out: dL_dcolor[3,H,W] and dL_ddepth[1,H,W]
 * (either may be NULL) = upstream * d loss / d(color, depth) -- what autograd hands the rasterizer's backward after
 * get_loss_mapping / get_loss_tracking (utils/slam_utils.py:58-146). Same device code as the fused path. */
int gsr_l1_loss_grad(void* stream, int width, int height, const gsr_l1_loss* loss, float* dL_dcolor,
                     float* dL_ddepth);
int gsr_backward_l1(void* stream, int P, int sh_degree, int sh_coeffs,
                    int width, int height, const float* background,
                    const float* means3D, const float* shs,
                    const float* colors_precomp, const float* scales,
                    float scale_modifier, const float* rotations,
                    const float* cov3D_precomp, const float* viewmatrix,
                    const float* projmatrix, const float* projmatrix_raw,
                    const float* campos, float tan_fovx, float tan_fovy,
                    const int* radii, const void* geom_ws, const void* image_ws,
                    const uint32_t* point_list, int64_t num_rendered,
                    const gsr_l1_loss* loss, void* scratch,
                    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                    float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                    float* dL_dscales, float* dL_drotations,
                    float* dL_dtau_per_gaussian, float* dL_dtau, int debug);

/* ---- backward for the sharded mapping window (monogs_b200/window.py; no reference counterpart: it forms
 * the sums loss.backward() makes over the window's keyframes, utils/slam_mapper.py:273-322,394) ------------
 * gsr_backward_render: the per-pixel half (zeroes the partial-gradient scratch, then the backward tile renderer).
 * gsr_backward_rows:    the per-Gaussian half on the Gaussian range [p_begin, p_end) (p_begin a multiple of
 *   256), written -- or ADDED when accumulate != 0 -- into 64-byte window rows
 *   rows[P][16] = { dL/dxyz 3 | dL/drgb 3 | dL/dopacity 1 | dL/dscale 3 | dL/drotation 4 |
 *                   ||dL/dmean2D[:2]|| (densification statistic, gaussian_model.py:888-892) | visible (radii>0) }.
 *   A Gaussian range is one contiguous slice of `rows`, so the caller can hand finished ranges to an
 *   allreduce while later ranges are still being computed. Ranges must be issued in ascending order; the
 *   call with p_end == P also writes dL_dtau[6] (rho | theta). dL_dmeans2D[P,3] is optional.
 *   Precomputed colours + scales/rotations only (what MonoGS's mapper renders with).
 *   `accumulate` is a bit set: GSR_ROWS_ADD (1) adds to the rows instead of overwriting them;
 *   GSR_ROWS_TOUCHED_ONLY (2): the caller zero-filled `rows` (or they hold earlier keyframes' sums) and only the rows
 *   of Gaussians whose gradient is non-zero are written / added -- every other row is left untouched, and column 15
 *   (the dense visibility flag) is not maintained (the sparse exchange carries visibility in vis_u8). */
#define GSR_ROWS_ADD 1
#define GSR_ROWS_TOUCHED_ONLY 2
int gsr_backward_render(void* stream, int P, int width, int height, const float* background,
                        const void* geom_ws, const void* image_ws, const uint32_t* point_list,
                        int64_t num_rendered, const float* dL_dout_color, const float* dL_dout_depth,
                        void* scratch, int debug);
/* gsr_backward_render with dL/dpixel formed in the tile renderer from a gsr_l1_loss descriptor (see gsr_backward_l1). */
int gsr_backward_render_l1(void* stream, int P, int width, int height, const float* background,
                           const void* geom_ws, const void* image_ws, const uint32_t* point_list,
                           int64_t num_rendered, const gsr_l1_loss* loss, void* scratch, int debug);
int gsr_backward_rows(void* stream, int P, int p_begin, int p_end, int width, int height,
                      const float* means3D, const float* scales, float scale_modifier,
                      const float* rotations, const float* viewmatrix, const float* projmatrix,
                      const float* projmatrix_raw, float tan_fovx, float tan_fovy, const int* radii,
                      const void* geom_ws, void* scratch, float* rows, int accumulate,
                      float* dL_dmeans2D, float* dL_dtau, int debug);

/* ---- sparse in-switch exchange of the window rows (csrc/rows_exchange.cu; no reference counterpart) ---------
 * One symmetric buffer per rank (same size and layout on every rank, mapped for peer access AND bound to an
 * NVLS multicast object by the caller, e.g. torch.distributed._symmetric_memory), zero-filled once:
 *   [ rows f32[P][16] | touched bitmaps u32[world][words_per_rank] | visibility counts u8[world][vis_bytes_per_rank]
 *     | barrier flags ]                       (offsets: gsr_rows_exchange_layout; size: gsr_rows_exchange_bytes)
 * gsr_backward_rows_marked = gsr_backward_rows that also writes (ORs / adds when accumulate != 0) this rank's
 *   marks into slot [rank] of its own buffer: touched_words = local + offsets[1] + 4*rank*words_per_rank (bit i of
 *   word i/32 = row i is non-zero here), vis_u8 = local + offsets[2] + rank*vis_bytes_per_rank (keyframes of this
 *   rank that see Gaussian i). Both NULL = plain gsr_backward_rows.
 * The buffer is laid out for a CAPACITY of rows (the `P` argument of gsr_rows_exchange_bytes / _layout): offsets and
 *   slot strides depend on the capacity only, so a map that grows inside it (keyframe insertion, densification)
 *   keeps its buffer, its marks pointers and its barrier flags; gsr_rows_exchange takes the live P <= capacity.
 * gsr_rows_exchange: ONE kernel, enqueued on every rank with the same P / capacity / world / n_ctas / epoch: rows that are
 *   non-zero on some rank are summed in the NVSwitch (multimem.ld_reduce) and multicast back (multimem.st); all
 *   other rows are left alone (zero gradients everywhere). vis_count[P] (local, optional) = sum over the ranks of
 *   the visibility counts; column 15 of the rows is not meaningful afterwards. peer_bufs = DEVICE array of
 *   world pointers to every rank's buffer in this process's address space (peer_bufs[rank] == local_buf);
 *   multicast_buf = the multicast address of the same buffer. epoch must advance by 2 per call, starting at 1,
 *   identically on all ranks. *status (device int, optional) is set to 1 if a peer did not arrive within 20 s
 *   (the kernel then returns instead of hanging). n_ctas <= 256, at most one CTA per SM (they spin on peers). */
size_t gsr_rows_exchange_bytes(int P, int world);
int gsr_rows_exchange_layout(int P, int world, size_t* offsets /*[4]: rows, bitmaps, visibility, flags*/,
                             int* words_per_rank, int* vis_bytes_per_rank);
int gsr_backward_rows_marked(void* stream, int P, int p_begin, int p_end, int width, int height,
                             const float* means3D, const float* scales, float scale_modifier,
                             const float* rotations, const float* viewmatrix, const float* projmatrix,
                             const float* projmatrix_raw, float tan_fovx, float tan_fovy, const int* radii,
                             const void* geom_ws, void* scratch, float* rows, int accumulate, float* dL_dmeans2D,
                             float* dL_dtau, uint32_t* touched_words, uint8_t* vis_u8, int debug);
int gsr_rows_exchange(void* stream, int P, int capacity, int rank, int world, int n_ctas, void* local_buf,
                      void* multicast_buf, const uint64_t* peer_bufs, uint32_t epoch, float* vis_count, int* status);
/* Measurement hook: globaltimer (ns) at kernel start / after the marks multicast / after barrier 1 / after the row
 * sums / after barrier 2 of the LAST gsr_rows_exchange on the current device (host_out[5]; synchronise first). */
int gsr_debug_exchange_trace(uint64_t* host_out);

/* ---- Gaussian insertion from an RGB-D keyframe: the device side of
 * GaussianModel.create_viewpoint_pcd (gaussian_splatting/scene/gaussian_model.py:121-319), the caller of distCUDA2.
 * Images are the reference's row-major [H,W] / [3,H,W] tensors; "flat index" i = x*H + y is the reference's
 * (W,H)-ordered flattening (:180-186), which fixes the order of the selected pixels.
 *  gsr_select_kth_f32  out[0] = the element of ascending rank k of v[i] = b ? |a[i]-b[i]| : a[i]   (k = (n-1)/2 is
 *                      torch.median's lower median, :177,:218); 4 radix-select launches, no sort. ws >= 8192 bytes.
 *  gsr_insert_select   densification mask (:191-226: depth >= 1e-3, and unless init: opacity < 0.5 or (depth in front
 *                      of the rendered depth and |error| > 50 * *median_err)) + stable compaction:
 *                      selected[0..*count) = flat indices in ascending order. ws = gsr_insert_ws_bytes(W, H).
 *  gsr_insert_points   for j < K: i = selected[pick ? pick[j] : j] (pick = the host's randperm subsample, int64,
 *                      :247-259): features[j] = rgb (clamp(exp(a) rgb + b, 0, 1) when exposure_a/b given, :135-139),
 *                      ids[j] = segmentation (optional), xyz[j] = c2w * (K^-1 * (x+.5, y+.5, 1) * depth)  (:262-279),
 *                      K = [[fx,0,cx],[0,fy,cy],[0,0,1]] and c2w = [R | T]^-1 inverted in the kernel (the reference
 *                      calls torch.linalg.inv, :163-165). R: row-major 3x3, T: 3 floats (device, world-to-camera).
 *  gsr_insert_finish   scales[j][0..scale_cols) = log(sqrt(max(dist2[j], 1e-7) * min(0.05, 0.01 * *median_depth)))
 *                      (:169-177,:294-306), rotations[j] = (1,0,0,0), opacities[j] = logit(0.5) = 0  (:308-318). */
size_t gsr_insert_ws_bytes(int width, int height);
int gsr_select_kth_f32(void* stream, int64_t n, const float* a, const float* b, int64_t k, float* out, void* ws);
int gsr_insert_select(void* stream, int width, int height, const float* gt_depth, const float* render_depth,
                      const float* render_opacity, int init, const float* median_err, int32_t* selected,
                      uint32_t* count, void* ws);
/* gsr_insert_select WITHOUT a workspace argument, for callers that cannot allocate device memory (CUDA graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same kernels and same results as gsr_insert_select; the per-block counts live in
 * a 16 KB `static __device__` array inside the library (held by every GPU that loads it until the process exits),
 * which bounds the image: width * height <= GSR_INSERT_STATIC_MAX_PIXELS, else GSR_ESTATICLIMIT (nothing launched).
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  The device current at the
 * call must be the stream's.  All calls on a device share the array: calls on one stream are ordered, calls that
 * overlap on the GPU (different streams, no dependency) race and give undefined `selected` / `count`. */
#define GSR_INSERT_STATIC_MAX_PIXELS (4 * 1024 * 1024)
int gsr_insert_select_static(void* stream, int width, int height, const float* gt_depth, const float* render_depth,
                             const float* render_opacity, int init, const float* median_err, int32_t* selected,
                             uint32_t* count);
int gsr_insert_points(void* stream, int K, const int32_t* selected, const int64_t* pick, int width, int height,
                      const float* rgb, const float* exposure_a, const float* exposure_b, const float* gt_depth,
                      const int32_t* segmentation, const float* R, const float* T, float fx, float fy, float cx,
                      float cy, float* xyz, float* features, int32_t* ids);
int gsr_insert_finish(void* stream, int K, const float* dist2, const float* median_depth, int scale_cols,
                      float* scales, float* rotations, float* opacities);

/* ---- map growth / pruning: the tensor surgery of GaussianModel (gaussian_splatting/scene/gaussian_model.py:642-778:
 * cat_tensors_to_optimizer, _prune_optimizer, densification_postfix, prune_points) on capacity-backed row arrays
 * (monogs_b200/optimizer_surgery.py is the drop-in host side).  All pointer tables are HOST arrays of device pointers.
 *  gsr_rows_append   for every group g < n_groups (<= 8): param[g][P_old .. P_old+K) = new_rows[g] ([K, cols[g]] fp32)
 *                    and, where given, exp_avg[g] / exp_avg_sq[g] rows [P_old, P_old+K) = 0 (the zeros_like extension
 *                    of :711-718) -- ONE launch, O(K) (the reference's torch.cat copies all P_old rows 15 times).
 *                    The arrays must have room for P_old + K rows.
 *  gsr_rows_compact  stable compaction by keep[P] (uint8 0/1, = ~mask of prune_points :679-695) of n_arrays (<= 32) row
 *                    arrays at once: row i of src[j] (row_bytes[j] bytes, a multiple of 4) moves to row
 *                    #{i' < i : keep[i']} of dst[j] (dst != src). *new_count (device u32) = number of kept rows.
 *                    ws: gsr_surgery_ws_bytes(P) bytes of device scratch. */
size_t gsr_surgery_ws_bytes(int64_t P);
int gsr_rows_append(void* stream, int n_groups, float* const* param, float* const* exp_avg, float* const* exp_avg_sq,
                    const float* const* new_rows, const int* cols, int64_t P_old, int64_t K);
int gsr_rows_compact(void* stream, int n_arrays, const void* const* src, void* const* dst, const int* row_bytes,
                     int64_t P, const uint8_t* keep, uint32_t* new_count, void* ws);

/* present[P] (uint8 0/1) = p_view.z > 0.2  (auxiliary.h:139-164). */
int gsr_mark_visible(void* stream, int P, const float* means3D,
                     const float* viewmatrix, const float* projmatrix,
                     uint8_t* present);

/* Per-call camera matrices of the reference's render() wrapper in ONE launch, no host synchronisation
 * (gaussian_splatting/gaussian_renderer/__init__.py:63-84, utils/camera_utils.py:171-178,224-231: a 4x4
 * assembly, a general 4x4 inverse and a bmm per call there):
 *   viewmatrix[16]  = getWorld2View(R, T)^T          (row-major; element [4*c + r] = W2C[r][c])
 *   projmatrix[16]  = viewmatrix x proj_T            (proj_T = projection matrix already transposed)
 *   campos[3]       = row 3 of viewmatrix^-1 = -R^-1 T (general 3x3 inverse, R need not be orthonormal)
 * R and T are read with element strides (r_stride_row, r_stride_col, t_stride) so that views of a 4x4 pose work. */
int gsr_camera_setup(void* stream, const float* R, int r_stride_row, int r_stride_col, const float* T,
                     int t_stride, const float* proj_T, float* viewmatrix, float* projmatrix, float* campos);

/* Mapper statistics of ONE keyframe, fused (utils/slam_mapper.py:452-460, gaussian_model.py:888-892): for every
 * Gaussian with radii > 0: max_radii_2d = max(max_radii_2d, radii); xyz_gradient_accum += ||dL_dmeans2D[:, :2]||;
 * denom += 1. max_radii_2d may be NULL; dL_dmeans2D NULL skips the gradient statistics. All arrays on the device. */
int gsr_densify_stats(void* stream, int P, const int* radii, const float* dL_dmeans2D /*[P,3]*/,
                      float* max_radii_2d /*[P]*/, float* xyz_gradient_accum /*[P]*/, float* denom /*[P]*/);
/* Window observation count and prune mask (utils/slam_mapper.py:399-440): nr_obs[i] = #{k < K : n_touched[k][i] > 0};
 * to_prune[i] = nr_obs[i] <= coviz && kf_idx[i] >= min_kf. n_touched = HOST array of K (<= 16) device pointers (the
 * n_touched outputs of the window's renders); kf_idx device i32[P]; either output may be NULL. */
int gsr_visibility_prune(void* stream, int P, int K, const int* const* n_touched, const int* kf_idx, int min_kf,
                         int coviz, int* nr_obs, uint8_t* to_prune);

/* mean_dist2[P] = mean squared distance to the 3 nearest other points. */
int gsr_knn_dist2(void* stream, int P, const float* points, float* mean_dist2,
                  void* knn_ws);

/* ---- measurement hook ------------------------------------------------------
 * Per-stage device timing with CUDA events recorded on the CALLER's stream around each kernel group
 * (bench.py's roofline leg). gsr_profile_enable(1) starts recording every stage, gsr_profile_enable(1 | mask << 8)
 * only the stages whose bit (1 << GSR_STAGE_x) is set in mask -- an event record between two kernels takes away their
 * programmatic-dependent-launch overlap, so timing all nine stages costs the step ~3 %; gsr_profile_enable(0) stops.
 * gsr_profile_collect waits for the recorded events, writes ms_sum[GSR_NUM_STAGES] (total device milliseconds per
 * stage since the last collect) and launches[GSR_NUM_STAGES] (how many times each stage ran) and clears the record. */
#define GSR_STAGE_PREPROCESS 0      /* preprocess_fwd_kernel (+ digit offsets)            */
#define GSR_STAGE_DEPTH_SORT 1      /* 4 onesweep passes over P (depth bits, id)           */
#define GSR_STAGE_SCAN 2            /* chunk sums + scan -> num_rendered                   */
#define GSR_STAGE_EMIT 3            /* binning, first half: bin lists (two-level) / tile counts / instance emission */
#define GSR_STAGE_TILE_SORT 4       /* binning, second half: refine / offsets + scatter / onesweep passes over R   */
#define GSR_STAGE_RANGES 5          /* tile_ranges_kernel                                  */
#define GSR_STAGE_RENDER_FWD 6      /* render_fwd_kernel                                   */
#define GSR_STAGE_RENDER_BWD 7      /* render_bwd_kernel                                   */
#define GSR_STAGE_PREPROCESS_BWD 8  /* preprocess_bwd_kernel + tau reduction               */
#define GSR_NUM_STAGES 9
int gsr_profile_enable(int on);
int gsr_profile_collect(double* ms_sum, int* launches);
/* Kernels launched by this library since it was loaded (always counted). */
unsigned long long gsr_profile_kernel_launches(void);

/* ---- test / parity hooks --------------------------------------------------*/
/* Stable LSD radix sort of (u64 key, u32 value) pairs on key bits
 * [0, end_bit) -- the operation cub::DeviceRadixSort::SortPairs performs at
 * rasterizer_impl.cu:306-311. `ws` needs gsr_binning_bytes(n) bytes. */
int gsr_sort_pairs_u64(void* stream, int64_t n, const uint64_t* keys_in,
                       const uint32_t* values_in, uint64_t* keys_out,
                       uint32_t* values_out, int end_bit, void* ws);

/* Pointers into the workspaces, for bit-exact checks against the reference's
 * GeometryState / BinningState / ImageState (rasterizer_impl.h:33-62).
 * geom record layout: see DESIGN.md "Data layout in HBM". */
/* The three views point INTO one packed 48-byte record per Gaussian: element i of a view is the float4 at
 * view + 12 * i floats (rec_a: x, y, depth, cull threshold; rec_b: conic xyz, opacity; rec_c: rgb, 0). */
int gsr_debug_geom_views(const void* geom_ws, int P, int width, int height, const float** rec_a,
                         const float** rec_b, const float** rec_c, const uint32_t** tiles_touched);
int gsr_debug_image_views(const void* image_ws, int width, int height,
                          const float** final_T, const uint32_t** n_contrib,
                          const uint32_t** ranges /*uint2[tiles]*/);
/* 1 when a forward at this image size needs the binning workspace (gsr_binning_bytes): the two-level sort-free
 * binning (bin lists) and the instance radix sort (keys) do; 0 when binning_ws may be NULL (single-level sort-free
 * binning, a test mode). */
int gsr_binning_needed(int width, int height);
/* Test hook: 0 = automatic choice of the binning implementation (single-level sort-free binning for lists of fewer
 * than 2 Mi instances on images of <= 4096 tiles, else two-level sort-free binning for images of up to 4096 bins of
 * 4x4 tiles = 65 536 tiles, instance radix sort above), 1 = radix sort at every size, 2 = single-level and
 * 3 = two-level wherever the image admits them. All produce the identical point_list / ranges. A forward and the
 * gsr_debug_sorted_keys call that inspects it must run under the same mode. */
int gsr_debug_set_binning(int mode);
/* Rebuilds the reference's sorted 64-bit keys (tile << 32 | depth bits, rasterizer_impl.cu:102-104) of the
 * LAST gsr_forward_render call that used these workspaces into keys_sorted[num_rendered] (device). */
int gsr_debug_sorted_keys(void* stream, const void* geom_ws, const void* binning_ws,
                          const uint32_t* point_list, int P, int64_t num_rendered, int width, int height,
                          uint64_t* keys_sorted);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H_ */
