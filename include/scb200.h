/*
 * scb200.h — C ABI of the B200-native Gaussian-splatting hot path (libscb200.so).
 *
 * Every entry point replaces one call the reference makes into a third-party CUDA
 * extension (citations are to /root/reference):
 *
 *   sc_projection_fwd/bwd      gsplat.fully_fused_projection   street_gaussian/models/street_gaussian_renderer.py:219-232
 *   sc_isect_count/emit        gsplat.isect_tiles              street_gaussian_renderer.py:243-252
 *   sc_exclusive_scan_i32      (torch.cumsum inside isect_tiles)
 *   sc_radix_sort_pairs        (cub::DeviceRadixSort inside isect_tiles)
 *   sc_isect_offset_encode     gsplat.isect_offset_encode      street_gaussian_renderer.py:253
 *   sc_sh_fwd/bwd              gsplat.spherical_harmonics      street_gaussian_renderer.py:256-260
 *   sc_rasterize_fwd/bwd       gsplat.rasterize_to_pixels      street_gaussian_renderer.py:267-280
 *                              (+ max_hit for diff_point_rasterization.PointRasterizer,
 *                               data_processor/utils/render_utils.py:150-177)
 *   sc_knn3_mean_dist2         simple_knn._C.distCUDA2         submodules/simple-knn/spatial.cu:16-25,
 *                                                              simple_knn.cu:186-222
 *   sc_isect_prepare/emit_sorted   the same isect_tiles(sort=True) as two calls around its one host read-back
 *   sc_bin_prepare/finish          isect_tiles(sort=True) + isect_offset_encode for ONE camera and a tile grid of
 *                                  at most 128 x 128 (street_gaussian_renderer.py:243-253): the fast path
 *   sc_*_opacity, sc_sh_*_view     the kernels above with the renderer's glue folded in
 *                                  (street_gaussian_renderer.py:233-238, :256-266)
 *   sc_photometric_fwd/bwd     loss_utils.l1_loss + ssim       street_gaussian/utils/loss_utils.py:21-38, :96-131,
 *                                                              combined as in train.py:183-186
 *   sc_compact_rows            prune_optimizer / prune_points  street_gaussian/models/gaussian_model.py:363-382, :416-431
 *   sc_densification_stats     set_max_radii2D + add_densification_stats
 *                                                              street_gaussian/models/street_gaussian_model.py:486-519
 *   sc_assemble_fwd/bwd        StreetGaussianModel.get_xyz / get_rotation  street_gaussian/models/street_gaussian_model.py:296-364
 *   sc_assemble_features_fwd/bwd  StreetGaussianModel.get_features (Fourier DC of the actors)  street_gaussian/models/street_gaussian_model.py:366-386
 *   sc_assemble_act_fwd/bwd    StreetGaussianModel.get_scaling / get_opacity  street_gaussian/models/street_gaussian_model.py:273-294, :389-407
 *   sc_densify_classify/move   densify_and_prune (clone + split + prune)  street_gaussian/models/gaussian_model.py:433-551
 *   sc_bin_finish_dev, sc_rasterize_fwd_depth_dev   the second half of isect_tiles(sort=True) and the compositing forward
 *                                  WITHOUT the host read-back of n_isects (street_gaussian_renderer.py:243-280): counts stay on
 *                                  the device, outputs sized by a capacity — a camera pass that can be captured in a CUDA graph
 *   sc_adam_step               torch.optim.Adam.step() as set up by GaussianModel.training_setup   gaussian_model.py:287-305
 *   sc_ply_pack/unpack         GaussianModel.make_ply / load_ply row layout   gaussian_model.py:82-156
 *   sc_lidar_gather/pack       make_lidar_ply / transform_lidar_ply / the uint8 packing of render_condition
 *                              street_gaussian/pointcloud_processor/base_processor.py:32-63, waymo_processor.py:179-242
 *   sc_rasterize_limit_warps, sc_exclusive_scan_i32_noalloc   no counterpart in the reference (scheduling knob; allocation-free scan)
 *
 * Conventions: all pointers are DEVICE pointers unless named host_*; row-major contiguous
 * fp32 / int32 / int64 arrays; `stream` is a cudaStream_t passed as void*; nothing here
 * synchronises the stream; every function returns 0 on success or a negative sc_status
 * (sc_last_error() gives a thread-local message).  No torch types cross this boundary.
 */
#ifndef SCB200_H
#define SCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum sc_status {
    SC_OK = 0,
    SC_ERR_BAD_ARG = -1,
    SC_ERR_CUDA = -2,
    SC_ERR_WORKSPACE = -3,
    SC_ERR_UNSUPPORTED = -4
};

const char* sc_last_error(void);
int sc_version(void);
/* number of kernel launches issued through this library since load (bench.py "gpu_launches") */
int64_t sc_launch_count(void);

/* ---- A1: fused projection ------------------------------------------------------------ */
/* means[N,3] quats[N,4](wxyz) scales[N,3] viewmats[C,4,4] Ks[C,3,3] ->
 * radii[C,N] i32, means2d[C,N,2], depths[C,N], conics[C,N,3], compensations[C,N] (nullable).
 * Culled Gaussians: radius 0, float outputs 0. */
int sc_projection_fwd(int64_t N, int C, const float* means, const float* quats, const float* scales,
                      const float* viewmats, const float* Ks, int width, int height, float eps2d,
                      float near_plane, float far_plane, float radius_clip, int32_t* radii,
                      float* means2d, float* depths, float* conics, float* compensations,
                      void* stream);
/* v_* inputs [C,N,...] (v_compensations nullable) -> v_means[N,3] v_quats[N,4] v_scales[N,3]
 * (each nullable; summed over cameras).  accumulate == 0: outputs are WRITTEN (zeros for
 * Gaussians no camera sees); accumulate != 0 (gradient-arena mode of the multi-view step):
 * outputs are += and Gaussians no camera sees are not touched. */
int sc_projection_bwd(int64_t N, int C, const float* means, const float* quats, const float* scales,
                      const float* viewmats, const float* Ks, int width, int height, float eps2d,
                      const int32_t* radii, const float* v_means2d, const float* v_depths,
                      const float* v_conics, const float* v_compensations, float* v_means,
                      float* v_quats, float* v_scales, int accumulate, void* stream);

/* ---- A5: spherical harmonics ----------------------------------------------------------- */
/* dirs[M,3] (un-normalised), coeffs[M,K,3], masks[M] u8 (nullable) -> colors[M,3]. */
int sc_sh_fwd(int64_t M, int K, int degree, const float* dirs, const float* coeffs,
              const uint8_t* masks, float* colors, void* stream);
/* v_colors[M,3] -> v_coeffs[M,K,3], v_dirs[M,3] (nullable).  accumulate == 0: fully written,
 * zeros where masked; accumulate != 0: += on unmasked entries only (gradient-arena mode). */
int sc_sh_bwd(int64_t M, int K, int degree, const float* dirs, const float* coeffs,
              const uint8_t* masks, const float* v_colors, float* v_coeffs, float* v_dirs,
              int accumulate, void* stream);

/* "View mode" of the same kernels, used by the fused renderer path (rows A2, A5, A6 of
 * SURVEY.md section 8 folded into their producers; K = 4 or 16):
 *   sc_projection_fwd_opacity: additionally opacities_out[C,N] = opacities[N] * compensation
 *     (plain opacity when compensations == NULL) and, both nullable, the renderer's result-packing
 *     outputs (row A11, street_gaussian_renderer.py:293-294): visibility[C,N] u8 = radii > 0 and
 *     radii_norm[C,N] = float(radii) * (1 / max(height, width));
 *   sc_sh_fwd_view: dirs = means - campos (campos: DEVICE float[3]), mask = radii > 0,
 *     colors4[M,4] = (clamp_min(colour + 0.5, 0), depths[M]);
 *   sc_sh_bwd_view: upstream v_colors4[M,4], clamp mask from the saved colors4_fwd; v_means
 *     receives the direction gradient;
 *   sc_projection_bwd_opacity: v_depths read with a stride (4: the depth channel of v_colors4),
 *     v_opacities_in = gradient w.r.t. opacities_out, v_opacities[N] receives its share. */
int sc_projection_fwd_opacity(int64_t N, int C, const float* means, const float* quats, const float* scales,
                              const float* opacities, const float* viewmats, const float* Ks, int width,
                              int height, float eps2d, float near_plane, float far_plane, float radius_clip,
                              int32_t* radii, float* means2d, float* depths, float* conics,
                              float* compensations, float* opacities_out, uint8_t* visibility,
                              float* radii_norm, void* stream);
int sc_projection_bwd_opacity(int64_t N, int C, const float* means, const float* quats, const float* scales,
                              const float* opacities, const float* viewmats, const float* Ks, int width,
                              int height, float eps2d, const int32_t* radii, const float* v_means2d,
                              const float* v_depths, int v_depths_stride, const float* v_conics,
                              const float* v_opacities_in, int use_compensations, float* v_means,
                              float* v_quats, float* v_scales, float* v_opacities, int accumulate,
                              void* stream);
int sc_sh_fwd_view(int64_t M, int K, int degree, const float* means, const float* campos,
                   const float* coeffs, const int32_t* radii, const float* depths, float* colors4,
                   void* stream);
int sc_sh_bwd_view(int64_t M, int K, int degree, const float* means, const float* campos,
                   const float* coeffs, const int32_t* radii, const float* colors4_fwd,
                   const float* v_colors4, float* v_coeffs, float* v_means, int accumulate, void* stream);

/* ---- A3/A4: tile intersection ---------------------------------------------------------- */
/* tiles_per_gauss[i] = size of the tile AABB of entry i (0 when radii[i] == 0).  Optional
 * (both or neither; need depths): depth_keys[i] = raw depth bits (0xffffffff when the entry
 * emits nothing), order_init[i] = i — the input of the depth pre-sort below. */
int sc_isect_count(int64_t CN, const float* means2d, const int32_t* radii, const float* depths,
                   int tile_size, int tile_width, int tile_height, int32_t* tiles_per_gauss,
                   uint32_t* depth_keys, uint32_t* order_init, void* stream);
/* counts_in_order[j] = tiles_per_gauss[order[j]] */
int sc_isect_gather_counts(int64_t CN, const uint32_t* order, const int32_t* tiles_per_gauss,
                           int32_t* counts_in_order, void* stream);
/* out[i] = sum_{j<i} in[j] (int64 accumulate); total (nullable) receives the grand total.
 * workspace >= sc_scan_workspace_bytes(n). */
int64_t sc_scan_workspace_bytes(int64_t n);
int sc_exclusive_scan_i32(int64_t n, const int32_t* in, int64_t* out, int64_t* total,
                          void* workspace, int64_t workspace_bytes, void* stream);
/* Allocation-free variant of sc_exclusive_scan_i32.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernels, same results; the block-sum scratch is a file-scope
 * `static __device__` array inside the library (64 KiB + 256 B, held by every GPU that loads the library until
 * the process exits) instead of a caller workspace.  Calls no cudaMalloc, cudaFree or other allocating routine.
 * Limit: n <= sc_exclusive_scan_i32_noalloc_max_n() = 2^24 elements; larger n returns SC_ERR_WORKSPACE with
 * a message naming the limit.  Every call on a device shares the one array (it keeps the previous call's
 * block sums; nothing needs clearing, each call overwrites what it reads): calls that OVERLAP on the same
 * GPU (different streams, no ordering between them) race — order them on one stream or with events. */
int64_t sc_exclusive_scan_i32_noalloc_max_n(void);
int sc_exclusive_scan_i32_noalloc(int64_t n, const int32_t* in, int64_t* out, int64_t* total, void* stream);
/* emits the row-major tile list of entry order[j] at cum_tiles[j] for j = 0..C*N-1
 * (order == NULL: identity, i.e. camera-major then Gaussian order, gsplat's unsorted output):
 * isect_ids = cam << (32+tile_n_bits) | tile << 32 | bits(depth); flatten_ids = cam*N + g.
 * With `order` = the stable depth order of the entries, a stable sort of the emitted pairs on
 * the (camera,tile) bits [32, 32+tile_n_bits+cam_bits) ALONE yields exactly the list a stable
 * sort on the full key yields (ties fall back to entry index in both) — 2 sort passes over
 * the intersection list instead of 6. */
int sc_isect_emit(int64_t N, int C, const float* means2d, const int32_t* radii, const float* depths,
                  const int64_t* cum_tiles, const uint32_t* order, int tile_size, int tile_width,
                  int tile_height, int tile_n_bits, int64_t n_isects, int64_t* isect_ids,
                  int32_t* flatten_ids, void* stream);
/* stable LSD radix sort of (key, u32 value) pairs on key bits [begin_bit, end_bit), n < 2^30;
 * u64-key and u32-key variants.  Both buffer pairs are ping-pong storage; RETURNS 0 if the
 * sorted pairs ended in keys_out/vals_out, 1 if they ended in keys_in/vals_in (even pass
 * count), <0 on error.  workspace >= sc_sort_workspace_bytes(n). */
int64_t sc_sort_workspace_bytes(int64_t n);
int sc_radix_sort_pairs(int64_t n, uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out,
                        uint32_t* vals_out, int begin_bit, int end_bit, void* workspace,
                        int64_t workspace_bytes, void* stream);
int sc_radix_sort_pairs_u32(int64_t n, uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out,
                            uint32_t* vals_out, int begin_bit, int end_bit, void* workspace,
                            int64_t workspace_bytes, void* stream);
/* Fused binning pipeline (what isect_tiles(sort=True) runs): two calls around the one host
 * read-back of the counts.
 *   sc_isect_prepare: tiles_per_gauss[CN]; order-preserving compaction of the emitting entries,
 *     their stable depth order, the exclusive scan of their tile counts — all left in
 *     `workspace` (>= sc_isect_workspace_bytes, 256-byte aligned), whose first 16 bytes hold
 *     int64 {n_visible, n_isects} once the stream has drained.
 *   sc_isect_emit_sorted: emits the (key, id) pairs in depth order (building the sort
 *     histograms on the fly) and stably sorts them on the (camera, tile) bits; the sorted list
 *     ends in isect_ids / flatten_ids; tmp_* are ping-pong buffers of the same sizes;
 *     sort_workspace >= sc_sort_workspace_bytes(n_isects). */
int64_t sc_isect_workspace_bytes(int64_t CN, int tile_width, int tile_height);
int sc_isect_prepare(int64_t CN, const float* means2d, const int32_t* radii, const float* depths,
                     int tile_size, int tile_width, int tile_height, int32_t* tiles_per_gauss,
                     void* workspace, int64_t workspace_bytes, void* stream);
int sc_isect_emit_sorted(int64_t N, int C, const float* means2d, const int32_t* radii, const float* depths,
                         int tile_size, int tile_width, int tile_height, int tile_n_bits, int cam_bits,
                         int64_t n_visible, int64_t n_isects, const void* prepare_workspace,
                         int64_t* isect_ids, int32_t* flatten_ids, int64_t* tmp_ids, int32_t* tmp_flatten,
                         void* sort_workspace, int64_t sort_workspace_bytes, void* stream);
int sc_isect_offset_encode(int64_t n_isects, const int64_t* isect_ids, int C, int tile_width,
                           int tile_height, int tile_n_bits, int32_t* offsets, void* stream);
/* Fast binning path: ONE camera (C == 1) and a tile grid of at most 128 x 128 tiles — every
 * configuration the reference renders (isect_tiles + isect_offset_encode,
 * street_gaussian_renderer.py:243-253).  Same results bit for bit as the generic pipeline above.
 *   sc_bin_supported: 1 when (C, tile grid) qualifies.
 *   sc_bin_prepare: like sc_isect_prepare (same workspace; its first 24 bytes hold int64 {n_visible,
 *     n_isects, n_items} once the stream has drained, n_items = sum of the tile-box widths), and
 *     additionally the tile boxes of the visible Gaussians in depth order and the histograms of
 *     their tile columns / rows.  ONE cooperative kernel (grid-wide barriers, one CTA per SM).
 *   sc_bin_finish: emits the intersections in depth order and sorts them by tile column INSIDE the
 *     emit kernel, sorts by tile row, writes flatten_ids[n_isects], isect_ids[n_isects] (nullable:
 *     skipped when the caller only composites) and offsets[tile_height * tile_width].
 *     sort_workspace >= sc_bin_sort_workspace_bytes(n_isects), 256-byte aligned. */
int sc_bin_supported(int C, int tile_width, int tile_height);
/* leave n_sms SMs out of the cooperative grid of sc_bin_prepare (process-wide): a concurrent kernel that holds
 * SMs — an NCCL collective on a side stream — would otherwise keep the cooperative launch waiting until it ends */
int sc_bin_reserve_sms(int n_sms);
int sc_bin_prepare(int64_t N, const float* means2d, const int32_t* radii, const float* depths, int tile_size,
                   int tile_width, int tile_height, int32_t* tiles_per_gauss, void* workspace,
                   int64_t workspace_bytes, void* stream);
int64_t sc_bin_sort_workspace_bytes(int64_t n_isects);
/* Sync-free form of sc_bin_finish (same reference call, isect_tiles + isect_offset_encode): the counts
 * {n_visible, n_isects, n_items} are NOT read back, they stay in the first 24 bytes of the prepare workspace;
 * flatten_ids / isect_ids hold `capacity` entries and sort_workspace >= sc_bin_sort_workspace_bytes(capacity).
 * A list longer than the capacity is DROPPED (all offsets 0, the view renders empty; no out-of-bounds access):
 * compare the counts with the capacity once the stream has passed and re-run with a larger one.  With it a camera pass has no host
 * dependency and can be captured in a CUDA graph (street_crafter_b200.pipeline). */
int sc_bin_finish_dev(int64_t N, const float* depths, int tile_width, int tile_height, int64_t capacity,
                      const void* prepare_workspace, int32_t* flatten_ids, int64_t* isect_ids, int32_t* offsets,
                      void* sort_workspace, int64_t sort_workspace_bytes, void* stream);
/* byte offset, inside the prepare workspace, of 16 uint64 %globaltimer stamps that CTA 0 of the cooperative
 * kernel leaves at its phase boundaries (measurement only: tools/bin_phases.py) */
int64_t sc_bin_phase_stamps_offset(int64_t N, int tile_width, int tile_height);
int sc_bin_finish(int64_t N, const float* depths, int tile_width, int tile_height, int64_t n_visible,
                  int64_t n_isects, int64_t n_items, const void* prepare_workspace, int32_t* flatten_ids, int64_t* isect_ids,
                  int32_t* offsets, void* sort_workspace, int64_t sort_workspace_bytes, void* stream);

/* ---- A7/A8: compositing ---------------------------------------------------------------- */
/* CH in 1..5; backgrounds[C,CH] nullable; max_hit 0 = unlimited.
 * render_colors[C,H,W,CH], render_alphas[C,H,W]; saved for backward (nullable):
 * last_ids[C,H,W] i32 (index into flatten_ids of the last contributing splat, -1 if none),
 * final_T[C,H,W] (exact final transmittance; 1-render_alpha loses bits near saturation).
 * stats (nullable, measurement only): two u64 counters ADDED to: [0] visited (pixel,splat)
 * pairs, [1] contributing pairs — the algorithmic-FLOP basis of the roofline. */
int sc_rasterize_fwd(int C, int64_t N, int64_t n_isects, int CH, int width, int height,
                     int tile_size, const float* means2d, const float* conics, const float* colors,
                     const float* opacities, const float* backgrounds, const int32_t* isect_offsets,
                     const int32_t* flatten_ids, int max_hit, float* render_colors,
                     float* render_alphas, int32_t* last_ids, float* final_T, uint64_t* stats,
                     void* stream);
/* sc_rasterize_fwd for the RGB + depth layout (CH = 4, no background, no max_hit) that also writes the
 * renderer's depth map depth_norm[C,H,W] = render_colors[..., 3] / max(render_alphas, 1e-10)
 * (row A11, street_gaussian_renderer.py:283; IEEE division). */
int sc_rasterize_fwd_depth(int C, int64_t N, int64_t n_isects, int width, int height, int tile_size,
                           const float* means2d, const float* conics, const float* colors4,
                           const float* opacities, const int32_t* isect_offsets, const int32_t* flatten_ids,
                           float* render_colors, float* render_alphas, float* depth_norm, int32_t* last_ids,
                           float* final_T, uint64_t* stats, void* stream);
/* the same behind sc_bin_finish_dev: flatten_ids holds `capacity` entries, the list length is read from the device
 * (n_isects_dev = the second int64 of the prepare workspace; 0 when it exceeds the capacity).  (sc_rasterize_bwd needs
 * no such form: it walks a tile's list from last_ids down to isect_offsets and never looks at the list's end;
 * pass the capacity as its n_isects.) */
int sc_rasterize_fwd_depth_dev(int C, int64_t N, int64_t capacity, const int64_t* n_isects_dev, int width, int height,
                               int tile_size, const float* means2d, const float* conics, const float* colors4,
                               const float* opacities, const int32_t* isect_offsets, const int32_t* flatten_ids,
                               float* render_colors, float* render_alphas, float* depth_norm, int32_t* last_ids,
                               float* final_T, uint64_t* stats, void* stream);

/* Resident compositing warps per SM for all later sc_rasterize_* launches (0 = as many as fit: 16).  A smaller
 * number leaves registers to the HBM-bound kernels a caller runs beside the compositing on another stream
 * (street_crafter_b200.pipeline); no counterpart in the reference (a scheduling knob of this library). */
int sc_rasterize_limit_warps(int warps_per_sm);

/* gradient buffers must be ZEROED by the caller; v_means2d_abs nullable; v_render_alphas nullable
 * (= no gradient flows into the alpha map). */
int sc_rasterize_bwd(int C, int64_t N, int64_t n_isects, int CH, int width, int height,
                     int tile_size, const float* means2d, const float* conics, const float* colors,
                     const float* opacities, const float* backgrounds, const int32_t* isect_offsets,
                     const int32_t* flatten_ids, const float* final_T,
                     const int32_t* last_ids, const float* v_render_colors,
                     const float* v_render_alphas, float* v_means2d, float* v_means2d_abs,
                     float* v_conics, float* v_colors, float* v_opacities, void* stream);

/* ---- N3: fused photometric loss --------------------------------------- */
/* loss = (1 - lambda_dssim) * lambda_l1 * L1 + lambda_dssim * (1 - SSIM)
 * (street_gaussian/utils/loss_utils.py:21-38, :84-131 combined as at train.py:183-186; no LPIPS).
 * image element (c, y, x), c < 3, lives at image[c * channel_stride + (y * W + x) * pixel_stride]
 * (the compositor's [H,W,4] output: strides 1 / 4; a contiguous [3,H,W] tensor: H*W / 1);
 * gt[3,H,W] contiguous; mask[H,W] u8 nullable (reference: l1 over masked pixels, SSIM on inputs
 * zeroed outside the mask).  out3 = {loss, L1, SSIM} (device floats).  The backward reads the
 * upstream scalar from DEVICE memory (nullable = 1) and WRITES d loss / d image for the three
 * channels with the given output strides, plus zeros for `zero_extra_channels` further channels
 * (3, 4, ... at the same strides: the depth channel of the compositor's interleaved output).
 * workspace >= sc_photometric_workspace_bytes, kept between the two calls. */
int64_t sc_photometric_workspace_bytes(int H, int W);
int sc_photometric_fwd(int H, int W, const float* image, int64_t channel_stride, int64_t pixel_stride,
                       const float* gt, const uint8_t* mask, float lambda_dssim, float lambda_l1,
                       void* workspace, int64_t workspace_bytes, float* out3, void* stream);
int sc_photometric_bwd(int H, int W, const float* image, int64_t channel_stride, int64_t pixel_stride,
                       const float* gt, const uint8_t* mask, float lambda_dssim, float lambda_l1,
                       const void* workspace, const float* upstream, float* v_image,
                       int64_t v_channel_stride, int64_t v_pixel_stride, int zero_extra_channels,
                       void* stream);

/* N3, second half: the per-iteration densification statistics in one pass
 * (street_gaussian/models/street_gaussian_model.py:486-497 set_max_radii2D, :503-519
 * add_densification_stats with absgrad, called at train.py:283-284).  For every Gaussian with
 * visibility != 0:  xyz_gradient_accum[i,0] += |absgrad_i * 0.5 * (W,H)|,
 * xyz_gradient_accum[i,1] += |grad_i * 0.5 * (W,H)|, denom[i] += 1,
 * max_radii2D[i] = max(max_radii2D[i], radii[i]).  All arrays fp32, visibility u8 / bool. */
int sc_densification_stats(int64_t N, const float* means2d_absgrad /*[N,2]*/, const float* means2d_grad /*[N,2]*/,
                           const uint8_t* visibility /*[N]*/, const float* radii /*[N]*/, int width, int height,
                           float* xyz_gradient_accum /*[N,2]*/, float* denom /*[N]*/, float* max_radii2D /*[N]*/,
                           void* stream);

/* N4 (prune / clone / split half): order-preserving compaction of up to SC_COMPACT_MAX_ARRAYS
 * row-major fp32 arrays that share one keep mask — parameters, Adam exp_avg / exp_avg_sq and the
 * densification statistics of every attribute in ONE launch
 * (street_gaussian/models/gaussian_model.py:363-382 prune_optimizer, :416-431 prune_points; the same
 * gather serves the `[selected_pts_mask]` selections of densify_and_clone / densify_and_split).
 * keep[N] u8 / bool; dest_index[N+1] i64 = exclusive prefix sum of keep with the total in entry N
 * (sc_exclusive_scan_i32 with `total` = out + N);
 * src / dst / row_floats are HOST arrays of n_arrays entries (device pointers inside);
 * dst[a] must hold sum(keep) rows of row_floats[a] floats. */
#define SC_COMPACT_MAX_ARRAYS 32
int sc_compact_rows(int64_t N, const uint8_t* keep, const int64_t* dest_index, int n_arrays,
                    const float* const* host_src, float* const* host_dst, const int* host_row_floats,
                    void* stream);

/* N4 (clone / split half): the whole densification cycle of
 * street_gaussian/models/gaussian_model.py:521-551 densify_and_prune (densify_and_clone :493-519,
 * densify_and_split :452-491, cat_optimizer :384-408, prune_points :416-431) as classify -> scan -> move.
 * sc_densify_classify writes four 0/1 segments flags[4N] = [stay | clone | child | split-selected]
 * (stay: kept original; clone: surviving copy; child: split row whose children survive the final prune) and
 * the number of clone-SELECTED rows (for the reference's scalar_dict) to *n_clone_selected (device).
 * Thresholds are the float32 values the reference compares against: clone_max_scale = percent_dense * extent,
 * split_min_scale = float32(percent_dense) * float32(extent), big_ws_scale = extent * percent_big_ws (used when
 * prune_big != 0, i.e. max_screen_size is set), split_shrink = 0.8 * n_copies; grad_column picks
 * xyz_gradient_accum[:, 0] (the base model) or [:, 1] (`densify_grad_abs_bkgd`, gaussian_model_bkgd.py:102-105).
 * The caller runs sc_exclusive_scan_i32 over the 4N flags (dest_index[4N+1]) and reads the segment totals
 * (n_stay, n_clone, n_child, n_split); sc_densify_move (two launches: the source-row index of the appended rows
 * into index_workspace, then the move) writes, for every array of the table, the output rows
 * [stay | clones | children of copy 0 | children of copy 1 ...] (n_stay + n_clone + n_copies * n_child rows):
 * kind PARAM rows are copied everywhere, MOMENT rows are copied to `stay` rows and zero elsewhere, the XYZ
 * array gets R(rotation) (noise * exp(scaling)) + xyz for children, the SCALING array log(exp(s) / split_shrink).
 * noise[n_copies * n_split, 3] is the standard-normal draw of torch.normal (row c * n_split + k belongs to copy
 * c of the k-th split-selected row).  Table arrays are HOST arrays as in sc_compact_rows. */
#define SC_DENSIFY_PARAM 0
#define SC_DENSIFY_MOMENT 1
#define SC_DENSIFY_XYZ 2
#define SC_DENSIFY_SCALING 3
int sc_densify_classify(int64_t N, const float* xyz_gradient_accum /*[N,2]*/, const float* denom /*[N]*/,
                        const float* scaling /*[N,3]*/, const float* opacity /*[N]*/, float max_grad,
                        float clone_max_scale, float split_min_scale, float min_opacity, int prune_big,
                        float big_ws_scale, float split_shrink, int grad_column /*0 | 1*/, int32_t* flags /*[4N]*/,
                        int32_t* n_clone_selected /*[1]*/, void* stream);
int sc_densify_move(int64_t N, const int32_t* flags /*[4N]*/, const int64_t* dest_index /*[4N+1]*/,
                    int64_t n_stay, int64_t n_clone, int64_t n_child, int64_t n_split, int n_copies,
                    const float* noise, const float* scaling, const float* rotation, float split_shrink,
                    int32_t* index_workspace /*[n_clone + 2 n_child]*/, int n_arrays,
                    const float* const* host_src, float* const* host_dst, const int* host_row_floats,
                    const int* host_kinds, void* stream);

/* N2 (geometric half): scene-graph assembly of the projection kernel's `means` / `quats` inputs —
 * StreetGaussianModel.get_xyz / get_rotation (street_gaussian/models/street_gaussian_model.py:296-364) with
 * quaternion_to_matrix / quaternion_raw_multiply (street_gaussian/utils/general_utils.py:125-146, :245-264) —
 * over the concatenation of up to SC_ASSEMBLE_MAX_SEGMENTS models per call.  Segment s covers output rows
 * [seg_start[s], seg_start[s+1]) (global row numbers: a scene with more segments is assembled by several calls
 * on the same outputs).  pose_rot[s] == NULL: plain segment (background, sky): xyz copied, rotation
 * F.normalize'd.  Otherwise an actor under the rigid pose (pose_rot[s][4] RAW wxyz, pose_trans[s][3], device
 * pointers): xyz_w = R(pose_rot / |pose_rot|) x' + t, rot_w = normalize(pose_rot (x) q), where rows with
 * flip[s][k] != 0 mirror x along flip_axis and pre-multiply q by flip_quat (host float[4]) — the training-time
 * symmetry prior (:313-318, :347-350).  All table arrays are HOST arrays of device pointers.
 * sc_assemble_bwd writes the gradients of every segment's own arrays (v_xyz[s][n_s,3], v_rot[s][n_s,4]; NULL =
 * not wanted) and ACCUMULATES the tracked-pose gradients into v_pose_rot[s][4] / v_pose_trans[s][3] (NULL = pose
 * not optimised; the caller zeroes them).
 * host_shell (HOST float[n_segments][4] = centre xyz, 2 R; NULL or 2 R = 0: none): the sky model's shell
 * projection of its positions, GaussianModelSky.get_xyz (street_gaussian/models/gaussian_model_sky.py:68-76):
 * d = xyz - c, ratio = |d| / (2 R), xyz_w = c + d / ratio where ratio < 1 — plain segments only. */
#define SC_ASSEMBLE_MAX_SEGMENTS 32
int sc_assemble_fwd(int n_segments, const int64_t* host_seg_start /*[n_segments+1]*/, const float* const* host_xyz,
                    const float* const* host_rot, const float* const* host_pose_rot,
                    const float* const* host_pose_trans, const uint8_t* const* host_flip, int flip_axis,
                    const float* host_flip_quat /*[4]*/, const float* host_shell /*[n_segments][4], nullable*/,
                    float* out_xyz /*[N,3]*/, float* out_rot /*[N,4]*/, void* stream);
int sc_assemble_bwd(int n_segments, const int64_t* host_seg_start, const float* const* host_xyz,
                    const float* const* host_rot, const float* const* host_pose_rot,
                    const float* const* host_pose_trans, const uint8_t* const* host_flip, int flip_axis,
                    const float* host_flip_quat, const float* host_shell, const float* v_out_xyz /*[N,3]*/,
                    const float* v_out_rot /*[N,4]*/,
                    float* const* host_v_xyz, float* const* host_v_rot, float* const* host_v_pose_rot,
                    float* const* host_v_pose_trans, void* stream);

/* N2, activations of the concatenated set: StreetGaussianModel.get_scaling (street_gaussian_model.py:273-294) =
 * cat of exp(_scaling) per model (clamped to clamp_max[s] for the sky model, gaussian_model_sky.py:63-66; pass
 * +inf otherwise) and get_opacity (:389-407) = cat of sigmoid(_opacity).  Same segment convention as
 * sc_assemble_fwd; host_clamp_max is a HOST float[n_segments].  The backward recomputes the activations from
 * the raw parameters and writes every segment's own gradient arrays (NULL = not wanted). */
int sc_assemble_act_fwd(int n_segments, const int64_t* host_seg_start, const float* const* host_scaling /*[n_s,3]*/,
                        const float* const* host_opacity /*[n_s]*/, const float* host_clamp_max,
                        float* out_scaling /*[N,3]*/, float* out_opacity /*[N]*/, void* stream);
int sc_assemble_act_bwd(int n_segments, const int64_t* host_seg_start, const float* const* host_scaling,
                        const float* const* host_opacity, const float* host_clamp_max,
                        const float* v_out_scaling /*[N,3]*/, const float* v_out_opacity /*[N]*/,
                        float* const* host_v_scaling, float* const* host_v_opacity, void* stream);

/* N2, features of the concatenated set: StreetGaussianModel.get_features (street_gaussian_model.py:366-386) —
 * cat(f_dc, f_rest) per model (gaussian_model.py:238-241); for actors the DC colour is the inverse Fourier sum
 * over n_fourier[s] coefficients, sum_c f_dc[:, c, :] * idft_base[s][c] (get_features_fourier,
 * gaussian_model_actor.py:67-76; idft_base = IDFT(time, C) of street_gaussian/utils/sh_utils.py:120-130, evaluated by
 * the host).  f_dc[s] is [n_s, n_fourier[s], 3], f_rest[s] [n_s, K-1, 3], out_features [N, K, 3];
 * host_idft_base is a HOST float[n_segments][SC_ASSEMBLE_MAX_FOURIER] (plain segments: n_fourier 1, base 1). */
#define SC_ASSEMBLE_MAX_FOURIER 8
int sc_assemble_features_fwd(int n_segments, const int64_t* host_seg_start, int K, const float* const* host_f_dc,
                             const float* const* host_f_rest, const int* host_n_fourier, const float* host_idft_base,
                             float* out_features, void* stream);
int sc_assemble_features_bwd(int n_segments, const int64_t* host_seg_start, int K, const int* host_n_fourier,
                             const float* host_idft_base, const float* v_out_features, float* const* host_v_f_dc,
                             float* const* host_v_f_rest, void* stream);

/* ---- A10: distCUDA2 -------------------------------------------------------------------- */
/* points[P,3] -> mean squared distance to the 3 nearest other points [P] (+inf if P < 4,
 * as the reference's FLT_MAX placeholders overflow).  Never synchronises (scene bounds stay
 * on the device, unlike the reference's two blocking cudaMemcpy). */
int64_t sc_knn_workspace_bytes(int64_t P);
int sc_knn3_mean_dist2(int64_t P, const float* points, float* mean_dist2, void* workspace,
                       int64_t workspace_bytes, void* stream);

/* ---- measurement helper ----------------------------------------------------------------- */
/* FP32 FMA / SFU(ex2) issue-rate probes used by bench.py as the roofline denominator of the
 * compositing kernels: every thread runs `iters` rounds of 8 independent dependent chains.
 * flops = grid*block*iters*16 (fma) ; ex2 ops = grid*block*iters*8.  sink[grid*block] out. */
int sc_probe_fp32(int kind /*0 fma, 1 ex2*/, int grid, int block, int iters, float* sink, void* stream);

/* ---- N1: LiDAR condition pipeline (street_gaussian/pointcloud_processor/waymo_processor.py:179-242) ---- */
/* sc_lidar_gather: the point set of ONE condition frame from device-resident clouds.  segs_dev: n_segs (<= 512)
 * records of sc_lidar_segment_bytes() bytes {int64 dst0, int64 count, const void* src, int32 is_f64,
 * int32 has_pose, double pose[12]}: `count` rows of 6 values (xyz rgb; float32 world-space background clouds,
 * float64 actor clouds posed by the 3x4 float64 object->world matrix exactly as base_processor.py:58-63) land
 * at [dst0, dst0 + count).  Writes means[n,3], rgb[n,3] and the isotropic splat scales[n,3]: `scale`, or with
 * use_ndc_scale scale * z / fx * 0.5 * min(H, W) (data_processor/utils/render_utils.py:116-122; w2c_row2 =
 * third row of the float32 world->camera matrix).
 * sc_lidar_pack: compositor output (rgb[H*W,3], alpha[H*W]) -> uint8 [H*W,4] = (x * 255).astype(uint8) of the
 * reference's PNG pair (waymo_processor.py:237-242) and / or float [4,H*W] of the quantised values. */
int64_t sc_lidar_segment_bytes(void);
int sc_lidar_gather(int64_t n_points, int n_segs, const void* segs_dev, const float* w2c_row2, float scale, float fx,
                    int height, int width, int use_ndc_scale, float* means, float* rgb, float* scales, void* stream);
int sc_lidar_pack(int64_t n_pix, const float* rgb, const float* alpha, uint8_t* out_u8, float* out_f32, void* stream);

/* ---- N4 / 8(e): optimiser step and checkpoint rows on the flat parameter arena ---------------------------- */
/* sc_adam_step: torch.optim.Adam(lr per group, eps) as GaussianModel.training_setup configures it
 * (street_gaussian/models/gaussian_model.py:293-305) over ONE flat fp32 buffer that holds every parameter
 * group: host_group_start[n_groups+1] ascending element offsets (multiples of 4), host_lr[n_groups]; elements
 * [lo, hi) (multiples of 4) are updated — the slice one rank owns after the gradient reduce-scatter.
 * params / exp_avg / exp_avg_sq are updated in place; `step` is the 1-based step count of the bias
 * correction; beta1 / beta2 / eps are the Python doubles torch forms 1 - beta from.  n_groups <= 64. */
int sc_adam_step(int n_groups, const int64_t* host_group_start, const float* host_lr, int64_t lo, int64_t hi,
                 float* params, const float* grads, float* exp_avg, float* exp_avg_sq, double beta1, double beta2,
                 double eps, int64_t step, void* stream);
/* Row layout of GaussianModel.make_ply / load_ply (gaussian_model.py:82-99, :107-156, :328-342):
 * x y z nx ny nz | f_dc (3) | f_rest (3 (K-1), channel-major) | opacity | scale (3) | rot (4) | semantic (S).
 * xyz[N,3] f_dc[N,1,3] f_rest[N,K-1,3] opacity[N,1] scaling[N,3] rotation[N,4] semantic[N,S] <-> rows[N,width]. */
int sc_ply_row_width(int K, int S);
int sc_ply_pack(int64_t N, int K, int S, const float* xyz, const float* f_dc, const float* f_rest, const float* opacity,
                const float* scaling, const float* rotation, const float* semantic, float* rows, void* stream);
int sc_ply_unpack(int64_t N, int K, int S, const float* rows, float* xyz, float* f_dc, float* f_rest, float* opacity,
                  float* scaling, float* rotation, float* semantic, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SCB200_H */
