// kernels.cuh -- host-side launchers of the sm_100a kernels (one per .cu file).
#pragma once
#include "common.cuh"

namespace gsr {

// preprocess.cu
cudaError_t launch_preprocess_fwd(cudaStream_t st, int P, int D, int M, const float* means3D, const float* shs,
                                  const float* colors_precomp, const float* opacities, const float* scales,
                                  float scale_modifier, const float* rotations, const float* cov3D_precomp,
                                  const float* viewmatrix, const float* projmatrix, const float* campos, int W,
                                  int H, float tan_fovx, float tan_fovy, int prefiltered, int* radii,
                                  const GeomViews& g);
// tile_order_fwd != NULL: an extra CTA also sorts the tiles by the previous forward's tile_work (the forward renderer's map)
cudaError_t launch_chunk_scan(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g, int n_tiles,
                              const uint32_t* tile_work, uint32_t* tile_order_fwd);
cudaError_t launch_mark_visible(cudaStream_t st, int P, const float* means3D, const float* viewmatrix,
                                uint8_t* present);
cudaError_t launch_camera_setup(cudaStream_t st, const float* R, int rs, int cs, const float* T, int ts,
                                const float* proj_T,
                                float* viewmatrix, float* projmatrix, float* campos);

// binning.cu
cudaError_t launch_gather_rects(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g);
cudaError_t launch_emit_instances(cudaStream_t st, int P, const uint32_t* order, const GeomViews& g, int tiles_x,
                                  long long cap, const SortPlan& plan, uint32_t* tkeys, uint32_t* vals,
                                  uint32_t* hist);
cudaError_t launch_tile_ranges(cudaStream_t st, long long cap, const unsigned long long* n_dev,
                               const uint32_t* tkeys_sorted, uint2* ranges, size_t n_tiles);
cudaError_t launch_rebuild_keys(cudaStream_t st, long long L, const uint32_t* tkeys_sorted,
                                const uint32_t* point_list, const float4* rec, uint64_t* keys);

// tile_scatter.cu  (sort-free binning for images of <= TS_MAX_TILES tiles)
cudaError_t launch_tile_count(cudaStream_t st, int P, const GeomViews& g, int tiles_x, int n_tiles);
cudaError_t launch_tile_offsets_scatter(cudaStream_t st, int P, const GeomViews& g, int tiles_x, int n_tiles,
                                        long long cap, uint2* ranges, uint32_t* point_list);
// two-level binning: level 1 on bins of TS2_CB x TS2_CB tiles (tile_scatter.cu), level 2 = tile_refine.cu
cudaError_t launch_bin_count(cudaStream_t st, int P, const GeomViews& g, int bins_x, int n_bins);
cudaError_t launch_bin_offsets_scatter(cudaStream_t st, int P, const GeomViews& g, int bins_x, int n_bins,
                                       long long cap, const RefineViews& rv);
cudaError_t launch_refine(cudaStream_t st, const GeomViews& g, const RefineViews& rv, int width, int height,
                          long long cap, uint2* ranges, uint32_t* point_list);
cudaError_t launch_rebuild_keys_from_starts(cudaStream_t st, long long L, int n_tiles, const GeomViews& g,
                                            const uint32_t* point_list, uint64_t* keys);

// radix_sort.cu  (n_cap = element count or capacity; n_dev = optional device-resident count <= n_cap)
cudaError_t sort_pairs_u64_prehist(cudaStream_t st, long long n_cap, const unsigned long long* n_dev,
                                   uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                                   const SortPlan& plan, const uint32_t* digit_offsets, uint32_t* tickets,
                                   uint32_t* lookback, uint32_t* final_vals);
cudaError_t sort_pairs_u32_prehist(cudaStream_t st, long long n_cap, const unsigned long long* n_dev,
                                   uint32_t* keys_a, uint32_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                                   const SortPlan& plan, const uint32_t* digit_offsets, uint32_t* tickets,
                                   uint32_t* lookback, uint32_t* final_vals);
cudaError_t launch_digit_hist_u64(cudaStream_t st, const uint64_t* keys, long long n, const SortPlan& plan,
                                  uint32_t* hist);
cudaError_t launch_digit_hist_u32(cudaStream_t st, const uint32_t* keys, long long n, const SortPlan& plan,
                                  uint32_t* hist);
cudaError_t launch_digit_offsets(cudaStream_t st, uint32_t* hist, int n_passes);

// render_fwd.cu
cudaError_t launch_render_fwd(cudaStream_t st, int W, int H, const ImageViews& im, const uint32_t* point_list,
                              const GeomViews& g, const float* bg, float* out_color, float* out_depth,
                              float* out_opacity, int* n_touched, bool tile_order_ready = false);

// render_bwd.cu
// tile_order[0..n_tiles) = tiles by descending tile_work (any content of tile_work gives a valid permutation)
cudaError_t launch_tile_order(cudaStream_t st, int n_tiles, const uint32_t* tile_work, uint32_t* tile_order);
cudaError_t launch_render_bwd(cudaStream_t st, int W, int H, const ImageViews& im, const uint32_t* point_list,
                              const GeomViews& g, const float* bg, const float* dL_dpix,
                              const float* dL_dpix_depth, const LossView& loss, float* packed,
                              uint8_t* touched_flag /* [P], inside the zeroed bytes: set for every Gaussian added to */,
                              size_t zero_bytes /* from `packed` on, zeroed before the renderer runs (multiple of 16) */);

// loss.cu
cudaError_t launch_l1_loss(cudaStream_t st, int kind, int W, int H, const float* color, const float* depth,
                           const float* opacity, const float* gt_rgb, const float* gt_depth, const uint8_t* mask,
                           const uint8_t* grad_mask, const float* exposure_a, const float* exposure_b,
                           float lambda_rgb, double* acc, float* out);

cudaError_t launch_l1_loss_grad(cudaStream_t st, int W, int H, const LossView& loss, float* dL_dcolor,
                                float* dL_ddepth);

// preprocess_bwd.cu
cudaError_t launch_preprocess_bwd(cudaStream_t st, int P, int D, int M, const float* means3D, const int* radii,
                                  const float* shs, const uint8_t* clamped, const float* scales,
                                  const float* rotations, float scale_modifier, const float* cov3D_precomp,
                                  const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw,
                                  const float* campos, int W, int H, float tan_fovx, float tan_fovy,
                                  const BackwardViews& bw, float* dL_dmeans2D, float* dL_dcolors,
                                  float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                                  float* dL_dscales, float* dL_drots, float* dL_dtau_pg, float* dL_dtau,
                                  int p_begin, int p_end, float* rows, int accumulate,
                                  uint32_t* touched_words = nullptr, uint8_t* vis_u8 = nullptr);

// rows_exchange.cu  (sparse in-switch exchange of the mapping-window rows)
struct XchgLayout {
  size_t rows, bitmap, vis, flags, total;  // byte offsets into the symmetric buffer, and its size
  int n_gran;                              // 128-Gaussian granules = ceil(P / 128)
  int words_pad;                           // bitmap words per rank (4 per granule)
  int p_pad;                               // visibility bytes per rank (128 per granule)
};
constexpr int XCHG_MAX_CTAS = 256;
constexpr int XCHG_MAX_WORLD = 32;
XchgLayout xchg_layout(int P, int world);
cudaError_t launch_rows_exchange(cudaStream_t st, int P, int capacity, int rank, int world, int n_ctas,
                                 void* local_buf, void* multicast_buf, const unsigned long long* peer_bufs,
                                 uint32_t epoch, float* vis_count, int* status);
cudaError_t read_exchange_trace(unsigned long long* host_out /*[5]*/);

// surgery.cu  (append rows to / compact the rows of the parameter + Adam-state arrays)
constexpr int SURGERY_MAX_GROUPS = 8;
constexpr int SURGERY_MAX_ARRAYS = 32;
cudaError_t launch_rows_append(cudaStream_t st, int n_groups, float* const* param, float* const* exp_avg,
                               float* const* exp_avg_sq, const float* const* rows, const int* cols, long long P_old,
                               long long K);
size_t surgery_ws_bytes(long long P);
cudaError_t launch_rows_compact(cudaStream_t st, int n_arrays, const void* const* src, void* const* dst,
                                const int* row_bytes, long long P, const uint8_t* keep, uint32_t* new_count, void* ws);

// stats.cu
constexpr int VIS_MAX_KEYFRAMES = 16;  // keyframes per window the visibility kernel takes in one launch
cudaError_t launch_densify_stats(cudaStream_t st, int P, const int* radii, const float* grad2d, float* max_radii_2d,
                                 float* grad_accum, float* denom);
cudaError_t launch_visibility_prune(cudaStream_t st, int P, int K, const int* const* touched, const int* kf_idx,
                                    int min_kf, int coviz, int* nr_obs, uint8_t* to_prune);

// insert.cu  (Gaussian insertion from an RGB-D keyframe, gaussian_model.py:121-319)
constexpr size_t INSERT_SELECT_WS_BYTES = 8192;  // radix-select state: prefix, rank, tickets, 4 x 256 histogram
cudaError_t launch_select_kth(cudaStream_t st, long long n, const float* a, const float* b, long long k,
                              float* out, uint32_t* ws);
cudaError_t launch_insert_select(cudaStream_t st, int W, int H, const float* gt_depth, const float* render_depth,
                                 const float* render_opacity, int init, const float* median_err, int32_t* sel,
                                 uint32_t* total, uint32_t* block_counts);
cudaError_t launch_insert_points(cudaStream_t st, int K, const int32_t* sel, const long long* pick, int W, int H,
                                 const float* rgb, const float* exposure_a, const float* exposure_b,
                                 const float* gt_depth, const int32_t* segmentation, const float* R,
                                 const float* T, float fx, float fy, float cx, float cy, float* xyz, float* feat,
                                 int32_t* ids);
cudaError_t launch_insert_finish(cudaStream_t st, int K, const float* dist2, const float* median_depth,
                                 int scale_cols, float* scales, float* rots, float* opacities);

// knn.cu
size_t knn_workspace_bytes(int P);
cudaError_t launch_knn(cudaStream_t st, int P, const float* points, float* mean_dist2, void* ws);

}  // namespace gsr
