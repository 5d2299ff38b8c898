/* splat_b200 -- C ABI of the B200-native differentiable Gaussian-splat render.
 *
 * This is the drop-in boundary for the render path SplatFormer reaches through the
 * gsplat v0.1.11 Python API.  The reference binds this path as a torch C++ extension
 * (`gsplat.cuda._C`, loaded by `import gsplat` at /root/reference/utils/gs_utils.py:3);
 * gsplat v0.1.11 itself is a pip-from-git dependency (/root/reference/README.md:27) that
 * is not on disk, so each entry point below cites the reference CALL SITE it serves and
 * the upstream binding name it stands in for.
 *
 * Conventions (all entry points)
 *  - plain C: raw DEVICE pointers, sizes and scalars only; no torch types.
 *  - asynchronous on `stream` (a cudaStream_t passed as void*); no host synchronisation,
 *    no allocation, no ownership transfer, no global mutable state inside the library.
 *    Scratch memory is supplied by the caller (`*_temp_bytes` tells how much).
 *  - return value: 0 on success, otherwise a cudaError_t value (see sb_error_string);
 *    negative values are argument errors detected on the host (SB_ERR_*).
 *  - all float tensors are binary32, row-major, contiguous; ids/bins are int32, keys int64.
 *  - every call is batched over `num_views` views that share the Gaussians and the image
 *    size; the gsplat-compatible Python API calls with num_views = 1.  Per-view arrays are
 *    laid out view-major: (V, N, ...), (V, H, W, ...), tile bins (V*T, 2).
 *  - cameras: `viewmats` is (V, 3, 4) row-major world->camera (first 12 floats of a 3x4 or
 *    4x4 matrix, stride `viewmat_stride` floats between views); `intrins` is (V, 4) =
 *    fx, fy, cx, cy on the DEVICE, or NULL to use the scalar fx, fy, cx, cy arguments for
 *    every view (what gsplat.project_gaussians receives at gs_utils.py:88-91).
 */
#ifndef SPLAT_B200_H_
#define SPLAT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB_ABI_VERSION 3
#define SB_ERR_BAD_ARG (-1)
#define SB_ERR_TEMP_TOO_SMALL (-2)
#define SB_ERR_STATIC_SCRATCH_EXCEEDED (-3) /* sb_cumsum_i32_static: count > SB_CUMSUM_STATIC_MAX_COUNT */

typedef void* sb_stream_t; /* cudaStream_t */

#if defined(__GNUC__)
#define SB_API __attribute__((visibility("default")))
#else
#define SB_API
#endif

SB_API int sb_abi_version(void);
SB_API const char* sb_error_string(int code);

/* ---- a3 / a13: gsplat.spherical_harmonics (gs_utils.py:78) ---------------------------
 * upstream bindings: _C.compute_sh_forward / _C.compute_sh_backward.
 * dirs (V,N,3), coeffs (N,K,3) shared by all views, colors (V,N,3).
 * `degree` is the storage degree (K = (degree+1)^2), `degrees_to_use` <= degree.
 * backward: v_coeffs (N,K,3) = sum over views of Y_k(dir_v) * v_colors[v]; bases above
 * degrees_to_use receive 0; no gradient to dirs. */
SB_API int sb_sh_fwd(int num_views, int n, int degree, int degrees_to_use, const float* dirs,
              const float* coeffs, float* colors, sb_stream_t stream);
SB_API int sb_sh_bwd(int num_views, int n, int degree, int degrees_to_use, const float* dirs,
              const float* v_colors, float* v_coeffs, sb_stream_t stream);

/* ---- a4 / a12: gsplat.project_gaussians (gs_utils.py:82-95) ---------------------------
 * upstream bindings: _C.project_gaussians_forward / _backward.
 * Outputs (V,N,..): xys(2) depths radii(int) conics(3) compensation num_tiles_hit(int)
 * cov3d(6); compensation and cov3d may be NULL (cov3d is written for the Gaussians in front
 * of the view's near plane, also when they are culled later).
 * Culled Gaussians get radii = num_tiles_hit = 0 and zeros elsewhere. */
SB_API int sb_project_fwd(int num_views, int n, const float* means, const float* scales, float glob_scale,
                   const float* quats, const float* viewmats, int viewmat_stride,
                   const float* intrins, float fx, float fy, float cx, float cy, int img_h,
                   int img_w, int block_width, float clip_thresh, float* xys, float* depths,
                   int* radii, float* conics, float* compensation, int* num_tiles_hit,
                   float* cov3d, sb_stream_t stream);
/* v_* inputs are (V,N,..); v_depths, v_compensation and v_cov3d may be NULL.
 * Outputs v_means (N,3), v_scales (N,3), v_quats (N,4) are the sums over views (assigned,
 * not accumulated).  v_viewmats (V,3,4) may be NULL; if given it must be zero-filled by
 * the caller and receives the pose gradient. */
SB_API int sb_project_bwd(int num_views, int n, const float* means, const float* scales, float glob_scale,
                   const float* quats, const float* viewmats, int viewmat_stride,
                   const float* intrins, float fx, float fy, float cx, float cy, int img_h,
                   int img_w, int block_width, float clip_thresh, const int* radii,
                   const float* v_xys, const float* v_depths, const float* v_conics,
                   const float* v_compensation, const float* v_cov3d, float* v_means,
                   float* v_scales, float* v_quats, float* v_viewmats, sb_stream_t stream);

/* ---- a5: gsplat.utils.compute_cumulative_intersects ------------------------------------
 * inclusive int32 prefix sum over the flattened (V*N) tile counts (single-pass scan with
 * decoupled look-back, csrc/radix_sort.cu; `temp` is its look-back state). */
SB_API size_t sb_cumsum_temp_bytes(int count);
SB_API int sb_cumsum_i32(int count, const int* in, int* out, void* temp, size_t temp_bytes,
                  sb_stream_t stream);
/* The same prefix sum for callers that may not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The look-back state lives in a file-scope
 * `static __device__` array of the library (64 + 8 * SB_CUMSUM_STATIC_MAX_COUNT / 4096 =
 * 32 832 bytes, held by every GPU that loads the library until the process exits) -- the one
 * exception to "no global mutable state" above.  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine); the caller still passes `in` and `out`.
 *  - count > SB_CUMSUM_STATIC_MAX_COUNT returns SB_ERR_STATIC_SCRATCH_EXCEEDED before any
 *    CUDA call; count < 0 or a NULL pointer returns SB_ERR_BAD_ARG; count == 0 is a no-op.
 *  - the array is shared by every call on a device and keeps the previous call's contents:
 *    each call clears, on `stream`, the counter and the status words it is going to use.
 *    Calls that overlap on the GPU (streams with no ordering between them) RACE on it:
 *    serialise them.
 *  - the array's device address is looked up once per device (cudaGetSymbolAddress with that
 *    device current): make the first call on a device outside a stream capture.
 * Output identical to sb_cumsum_i32 (integer arithmetic, bit-exact). */
#define SB_CUMSUM_STATIC_MAX_COUNT (1 << 24)
SB_API int sb_cumsum_i32_static(int count, const int* in, int* out, sb_stream_t stream);

/* ---- a6: gsplat.utils.map_gaussian_to_intersects (_C.map_gaussian_to_intersects) -------
 * key = ((view*T + tile) << 32) | float32 bits of depth; gaussian_ids are view-local.
 * xys_stride: floats between consecutive pixel centres (2 for a dense (V*N,2) array, 12 to read
 * them straight out of the packed raster record); must be even (8-byte aligned float2 loads). */
SB_API int sb_map_gaussian_to_intersects(int num_views, int n, const float* xys, int xys_stride,
                                  const float* depths,
                                  const int* radii, const int* cum_tiles_hit, int tiles_x,
                                  int tiles_y, int block_width, int64_t* isect_ids,
                                  int* gaussian_ids, sb_stream_t stream);

/* Fused path, global-sort branch: emission in DEPTH ORDER.  sb_depth_order_keys writes, per (view,
 * Gaussian), key = view << 32 | depth bits and value = the flat index; after a stable sort of those
 * V*N pairs (sb_sort_pairs, end_bit = 32 + bits(V)) the values are the emission `order`.
 * sb_gather_i32 reorders the tile counts, sb_cumsum_i32 turns them into offsets, and
 * sb_map_gaussian_to_intersects_ordered emits slot j = Gaussian order[j].  The emitted keys are then
 * already ascending in depth within every tile, so the big sort only has to run over the TILE bits
 * [32, end_bit) (sb_sort_pairs_range, stable): 2 radix passes instead of 6.  Per-tile result:
 * ascending depth bits, ties by ascending Gaussian id -- identical to sorting the full 64-bit key. */
SB_API int sb_depth_order_keys(int num_views, int n, const float* depths, const int* radii, int64_t* keys,
                        int* index, sb_stream_t stream);
SB_API int sb_gather_i32(int count, const int* src, const int* index, int* dst, sb_stream_t stream);
SB_API int sb_map_gaussian_to_intersects_ordered(int num_views, int n, const float* xys, int xys_stride,
                                          const float* depths, const int* radii, const int* order,
                                          const int* cum_tiles_hit_ordered, int tiles_x, int tiles_y,
                                          int block_width, int64_t* isect_ids, int* gaussian_ids,
                                          sb_stream_t stream);

/* Compact variant of the same branch (what render.py uses): with depth-ordered emission the key only has
 * to carry the tile, so it is a uint32 (view*T + tile): 8 bytes per pair through the sort instead of 12. */
SB_API int sb_map_gaussian_to_tiles_ordered(int num_views, int n, const float* xys, int xys_stride,
                                     const float* depths, const int* radii, const int* order,
                                     const int* cum_tiles_hit_ordered, int tiles_x, int tiles_y,
                                     int block_width, uint32_t* tile_ids, int* gaussian_ids,
                                     sb_stream_t stream);
/* sb_sort_tiles_global: stable sort of the (tile index, Gaussian id) pairs on the tile index
 * (num_tiles_total = V*T distinct values), keeping ONLY the sorted ids (the key is implied by the bin); the
 * last sweep counts the pairs per tile, and a single-pass scan + one elementwise kernel inside the same call
 * turn the counts into tile_bins (num_tiles_total, 2), (0,0) for empty tiles.
 * Temp: sb_sort_tiles_global_temp_bytes(count, num_tiles_total). */
SB_API size_t sb_sort_tiles_global_temp_bytes(int count, int num_tiles_total);
SB_API int sb_sort_tiles_global(int count, int num_tiles_total, const uint32_t* tile_ids, const int* gaussian_ids,
                         int* gaussian_ids_sorted, int* tile_bins, const uint32_t* digit_hist, void* temp,
                         size_t temp_bytes, sb_stream_t stream);
SB_API size_t sb_sort32_temp_bytes(int count);
SB_API int sb_sort_pairs32(int count, int end_bit, const uint32_t* keys_in, const int* vals_in,
                    uint32_t* keys_out, int* vals_out, void* temp, size_t temp_bytes, sb_stream_t stream);
SB_API int sb_tile_bin_edges32(int count, const uint32_t* tile_ids_sorted, int num_tiles_total, int* tile_bins,
                        sb_stream_t stream);

/* ---- a7: the 64-bit (tile|depth) key sort (upstream: torch.sort + torch.gather) --------
 * stable LSD radix sort of (key, value) pairs over key bits [0, end_bit): hand-written one-sweep passes
 * (8-bit digits, decoupled look-back, csrc/radix_sort.cu), count < 2^30. */
SB_API size_t sb_sort_temp_bytes(int num_intersects);
SB_API int sb_sort_pairs_range(int count, int begin_bit, int end_bit, const int64_t* keys_in,
                        const int* vals_in, int64_t* keys_out, int* vals_out, void* temp,
                        size_t temp_bytes, sb_stream_t stream);
SB_API int sb_sort_pairs(int num_intersects, int end_bit, const int64_t* keys_in, const int* vals_in,
                  int64_t* keys_out, int* vals_out, void* temp, size_t temp_bytes,
                  sb_stream_t stream);

/* ---- a8: gsplat.utils.get_tile_bin_edges (_C.get_tile_bin_edges) ------------------------
 * tile_bins (num_tiles_total, 2) int32 must be zero-filled by the caller. */
SB_API int sb_tile_bin_edges(int num_intersects, const int64_t* isect_ids_sorted, int num_tiles_total,
                      int* tile_bins, sb_stream_t stream);

/* ---- a10 / a11: gsplat.rasterize_gaussians (gs_utils.py:96-109) -------------------------
 * upstream bindings: _C.rasterize_forward / _C.rasterize_backward (3 channels).
 * The kernels read one packed 48-byte record per (view, Gaussian):
 *   [x, y, hx, hy | A, B, C, log2(opacity) | r, g, b, opacity]
 *   hx, hy: conservative half extents of the region where alpha >= 1/255 (used only to skip
 *   work); A = -0.5*log2(e)*conic.a, B = -log2(e)*conic.b, C = -0.5*log2(e)*conic.c, so that
 *   log2(alpha) = dx*(A*dx + B*dy) + C*dy*dy + log2(opacity).  Built from the API tensors by
 *   sb_pack_splats.  Gradients come back in a packed 48-byte record of MOMENTS that the caller
 *   zero-fills: [M1x, M1y, M2xx, M2xy | M2yy, M0, v_r, v_g | v_b, 0, 0, 0] with w = dL/dsigma per
 *   (pixel, Gaussian) pair, M0 = sum w, M1x = sum w*dx, ... M2yy = sum w*dy*dy.
 *   sb_unpack_splat_grads turns them into the API gradients: v_xy = (a*M1x + b*M1y,
 *   b*M1x + c*M1y), v_conic = (M2xx/2, M2xy, M2yy/2), v_opacity = -M0/opacity.
 * view_has_isects ((V,) int32 flags, may be NULL): when given, a view whose flag is 0 (not a
 * single intersection) gets out_img = background and final_T = 0 (alpha 1) -- the behaviour of
 * upstream's `num_intersects < 1` early return, applied per view on the device.
 * out_alpha (may be NULL) receives 1 - final_T.  clamp_rgb_max1 != 0 stores min(rgb, 1) -- the
 * wrapper's final clamp (gs_utils.py:111) folded into the kernel; the backward then takes that
 * stored image as out_img with clamped_rgb_max1 != 0 and passes no gradient through channels
 * that reached 1.
 * Fused render loss (the step right after the path, train.py:275,278): with gt (V,H,W,3) given
 * (float32 in [0,1], or uint8 with gt_is_u8 != 0, read as value / 255),
 * the forward also writes loss_partial (V, T, 2) = per-tile sums of |rgb - gt| and (rgb - gt)^2
 * (deterministic; the caller sums them), and the backward derives the image gradient itself:
 * v_out = v_loss[view] * sign(out_img - gt) (+ v_out_img if that is given too), so neither the
 * loss nor its gradient image ever exists as a tensor.  v_loss is a (V,) device array. */
SB_API int sb_pack_splats(int count, const float* xys, const float* conics, const float* colors,
                   const float* opacity, float* packed, sb_stream_t stream);
SB_API int sb_rasterize_fwd(int num_views, int n, const float* packed, const int* gaussian_ids_sorted,
                     const int* tile_bins, const float* background, int img_h, int img_w,
                     int block_width, const int* view_has_isects, float* out_img, float* final_ts,
                     int* final_idx, float* out_alpha, int clamp_rgb_max1, const void* gt,
                     int gt_is_u8, float* loss_partial, sb_stream_t stream);
SB_API int sb_rasterize_bwd(int num_views, int n, const float* packed, const int* gaussian_ids_sorted,
                     const int* tile_bins, const float* background, int img_h, int img_w,
                     int block_width, const float* final_ts, const int* final_idx,
                     const float* v_out_img, const float* v_out_alpha, const float* out_img,
                     int clamped_rgb_max1, const void* gt, int gt_is_u8, const float* v_loss,
                     float* v_packed, sb_stream_t stream);
SB_API int sb_unpack_splat_grads(int count, const float* v_packed, const float* conics,
                          const float* opacity, float* v_xys, float* v_conics, float* v_colors,
                          float* v_opacity, sb_stream_t stream);

/* ---- gsplat.utils.compute_cov2d_bounds [upstream gsplat/utils.py -> _C.compute_cov2d_bounds] ----
 * (SURVEY.md 8b re-export list, 8f-4; never reached by SplatFormer.)  cov2d (count,3) = upper
 * triangle (a, b, c) of the 2x2 screen covariance -> conics (count,3) = its inverse (c, -b, a)/det
 * and radii (count,) int32 = ceil(3 sqrt(lambda_max)), discriminant floored at 0.1 as in the
 * projection (A.2).  det == 0: conic 0, radius 0. */
SB_API int sb_cov2d_bounds(int count, const float* cov2d, float* conics, int* radii, sb_stream_t stream);

/* ---- a5-a9 for the fused path: fixed-capacity tile segments + per-tile depth sort ---------
 * Same per-tile order as sb_map_gaussian_to_intersects + sb_sort_pairs + sb_tile_bin_edges
 * (ascending depth bits, ties by ascending Gaussian id = the stable sort's emission order) without
 * the scan, the host read of the intersection count and the 64-bit global radix sort.
 * sb_views_project_fwd, given tile_counts (V*T, zero-filled) and tile_keys (V*T*tile_capacity),
 * drops (depth bits << 32 | gaussian id) of every intersection into slot atomicAdd(count) of its
 * tile's segment [tile*capacity, (tile+1)*capacity).  sb_sort_tiles then sorts every segment and
 * writes gaussian_ids_sorted (same V*T*capacity indexing), tile_bins (V*T,2) = segment ranges
 * ((0,0) for empty tiles), view_has_isects (V, zero-filled by the caller) and totals (2 ints,
 * zero-filled): [0] = number of intersections, [1] = flags: bit 0 if a tile held more than tile_capacity
 * Gaussians (its list is then truncated: the caller must redo the binning with the chain above), bit 1 if
 * the rect-list pool ran out (see below).
 * tile_capacity: power of two in [64, sb_tile_sort_capacity()], V*T*capacity < 2^31. */
SB_API int sb_tile_sort_capacity(void);
/* The same binning for Gaussians that were projected by an earlier sb_project_fwd call (what
 * gsplat.rasterize_gaussians receives, gs_utils.py:96-109): xys / depths / radii (V,N,..) -> slot claims in
 * tile_counts (V*T, zero-filled) and keys in tile_keys (V*T*tile_capacity), then sb_sort_tiles as above. */
SB_API int sb_bin_to_tile_segments(int num_views, int n, const float* xys, int xys_stride, const float* depths,
                            const int* radii, int tiles_x, int tiles_y, int block_width, int* tile_counts,
                            uint64_t* tile_keys, int tile_capacity, sb_stream_t stream);
SB_API int sb_sort_tiles(int num_views, int tiles_per_view, int tile_capacity, const int* tile_counts,
                  const uint64_t* tile_keys, int* gaussian_ids_sorted, int* tile_bins,
                  int* view_has_isects, int* totals, int n, int tiles_x, const float* packed,
                  int* rect_ids, int* rect_meta, int* pool_head, int pool_capacity, sb_stream_t stream);

/* ---- a10 / a11 for the fused path: rect-list rasteriser (block width 16 only) -------------
 * A 16x16 tile is eight 8x4-pixel rectangles (rect r: columns 8*(r&1).., rows 4*(r>>1)..).  Given
 * rect_ids (a pool of pool_capacity ints, plus 32 ints of slack), rect_meta (V*T*8, 2) and pool_head (one
 * int, zero-filled), sb_sort_tiles additionally writes, for every rectangle, the depth-ordered ids of the
 * Gaussians whose {alpha >= 1/255} region reaches it (exact ellipse / rectangle test on the packed records,
 * conservatively inflated) into the pool: rect_meta[rect] = (start, count), the list is
 * rect_ids[start .. start + count); the eight rectangles of a tile without any Gaussian get (-1, 0), which lets
 * sb_rasterize_rect_fwd write such a tile in one go (its loss sums then sit in rectangle 0's entry of
 * loss_partial, zeros in the other seven).  A tile claims room for its eight lists with one atomic on *pool_head;
 * if the pool runs out, bit 1 (value 2) of totals[1] is set, the tile's lists are left empty and *pool_head
 * keeps counting, so that it reports the capacity the scene needs (gaussian_ids_sorted may be NULL when only
 * the rect lists are wanted).  The two kernels below give each rectangle to one warp,
 * which walks its own list with no block-level synchronisation: same compositing order, same images
 * and gradients as sb_rasterize_fwd / _bwd (the dropped pairs are ones those kernels skip).
 * final_idx here is the position in the rectangle's list of the last composited Gaussian (-1: none);
 * it is only meaningful to sb_rasterize_rect_bwd.  The forward is a persistent kernel (resident CTAs x
 * SMs) whose warps pull rectangle numbers from *work_counter: one int that must be 0 on entry and is
 * left at an unspecified value.  loss_partial here is (V, T, 8, 2): one (sum |rgb - gt|, sum (rgb - gt)^2)
 * pair per rectangle.  All other arguments as sb_rasterize_fwd / _bwd. */
SB_API int sb_rasterize_rect_fwd(int num_views, int n, const float* packed, const int* rect_ids,
                          const int* rect_meta, const float* background,
                          int img_h, int img_w, const int* view_has_isects, float* out_img,
                          float* final_ts, int* final_idx, float* out_alpha, int clamp_rgb_max1,
                          const void* gt, int gt_is_u8, float* loss_partial, int* work_counter,
                          sb_stream_t stream);
SB_API int sb_rasterize_rect_bwd(int num_views, int n, const float* packed, const int* rect_ids,
                          const int* rect_meta, const float* background,
                          int img_h, int img_w, const float* final_ts, const int* final_idx,
                          const float* v_out_img, const float* v_out_alpha, const float* out_img,
                          int clamped_rgb_max1, const void* gt, int gt_is_u8, const float* v_loss,
                          float* v_packed, sb_stream_t stream);

/* ---- a1 / a2: fused multi-view projection + SH colour ----------------------------------
 * The view loop of rasterize_gaussians_to_multiimgs (gs_utils.py:20-27) with the per-view
 * colour path of gs_utils.py:64-79 folded in; one launch for all V views of a scene.
 *   sh_degree = -1: color_src is (N,3) view-independent rgb (degree-0 path, gs_utils.py:65)
 *   sh_degree 0..4: colour = max(SH(mean - campos_v) + 0.5, 0); the coefficients are either one
 *                   (N,K,3) array (color_src, color_rest = NULL) or SplatFormer's own split
 *                   features_dc (N,3) = color_src and features_rest (N,K-1,3) = color_rest
 *                   (gs_utils.py:59-60), which removes the torch.cat the reference does per view.
 * flags (OR of SB_FLAG_*) let the kernels apply the wrapper's activations themselves, so the raw
 * SplatFormer parameters go in and the gradients w.r.t. the raw parameters come out:
 *   SB_FLAG_RAW_SCALES     scales are log-scales, scale = exp(raw)              (gs_utils.py:45)
 *   SB_FLAG_RAW_OPACITIES  opacities are logits, opacity = sigmoid(raw)         (gs_utils.py:54)
 *   SB_FLAG_COLOR_SIGMOID  sh_degree = -1 only: colour = sigmoid(color_src)     (gs_utils.py:65)
 *   SB_FLAG_C2W_OPENGL     viewmats holds OpenGL camera-to-world matrices; the kernel forms
 *                          [R^T | -R^T T] with R <- R diag(1,-1,-1) (gs_utils.py:32-42) and takes
 *                          the camera centre from the matrix (campos may be NULL)
 *   SB_FLAG_INTRINS_SHARED intrins is (1,4): one fx fy cx cy for all views, as SplatFormer's
 *                          cameras are (dataset/GS.py:392-394 holds them as 0-dim tensors)
 *   SB_FLAG_RAW_QUATS      quats are raw: the kernel does the wrapper's q / |q| and replaces a
 *                          quaternion that does not come out unit (zero, NaN, infinite) by
 *                          (0,0,0,1), which then receives no gradient          (gs_utils.py:46-51)
 * quats may be un-normalised in every mode (the projection normalises -- a second time with
 * SB_FLAG_RAW_QUATS, exactly as gsplat does with the wrapper's output -- and the backward
 * differentiates through the normalisations).
 * stats (backward, may be NULL): two ints the caller zero-fills; the kernel adds the number of
 * (view, Gaussian) records it scanned to [0] and the number with a non-zero gradient (the ones whose
 * VJP it evaluated) to [1].
 * Writes the packed raster record (V,N,12) for drawn Gaussians, depths/radii/num_tiles_hit
 * (V,N) for all.  The backward consumes the packed gradient record of sb_rasterize_bwd and
 * writes the per-Gaussian sums over views: v_means (N,3) v_scales (N,3) v_quats (N,4)
 * v_opacities (N,) v_color_src / v_color_rest (same shapes as the inputs).
 * tile_counts / tile_keys / tile_capacity (may be NULL / NULL / 0): binning for sb_sort_tiles, see
 * above. */
#define SB_FLAG_RAW_SCALES 1
#define SB_FLAG_RAW_OPACITIES 2
#define SB_FLAG_COLOR_SIGMOID 4
#define SB_FLAG_C2W_OPENGL 8
#define SB_FLAG_INTRINS_SHARED 16
#define SB_FLAG_RAW_QUATS 32
SB_API int sb_views_project_fwd(int num_views, int n, const float* means, const float* scales,
                         float glob_scale, const float* quats, const float* opacities,
                         const float* color_src, const float* color_rest, int sh_degree, int flags,
                         const float* viewmats, int viewmat_stride, const float* intrins, float fx,
                         float fy, float cx, float cy, const float* campos, int img_h, int img_w,
                         int block_width, float clip_thresh, float* packed, float* depths,
                         int* radii, int* num_tiles_hit, int* tile_counts, uint64_t* tile_keys,
                         int tile_capacity, sb_stream_t stream);
SB_API int sb_views_project_bwd(int num_views, int n, const float* means, const float* scales,
                         float glob_scale, const float* quats, int sh_degree, int flags,
                         const float* viewmats, int viewmat_stride, const float* intrins, float fx,
                         float fy, float cx, float cy, const float* campos, int img_h, int img_w,
                         int block_width, float clip_thresh, const float* packed,
                         const float* v_packed, float* v_means, float* v_scales, float* v_quats,
                         float* v_opacities, float* v_color_src, float* v_color_rest, int* stats,
                         sb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SPLAT_B200_H_ */
