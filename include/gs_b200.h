/*
 * gs_b200.h -- C ABI of the B200 (sm_100a) Gaussian-splatting rasteriser.
 *
 * One entry point per stage of the reference operator
 * GaussianPointCloudRasterisation (reference file
 * taichi_3d_gaussian_splatting/GaussianPointCloudRasterisation.py, "R.py" below).
 * Every pointer is a raw CUDA device pointer unless it says "host"; sizes are
 * element counts; `stream` is a cudaStream_t passed as void*.  No torch types.
 * All functions return 0 on success, a negative GS_E_* code for bad arguments or
 * capacity problems, or a positive cudaError_t from the launch.  The caller owns
 * every buffer (the Python host allocates them through torch's caching allocator).
 *
 * Arithmetic: the forward entry points (inverse_pose .. blend_fwd) follow the
 * "literal fp32" profile documented in DESIGN.md (no FMA contraction, IEEE / and
 * sqrt, gs_expf polynomial) and are bit-exact with oracle/gs_oracle.c; the
 * backward entry points use fast fp32 (FMA, ex2.approx) and are compared at
 * 1e-4 relative.
 */
#ifndef GS_B200_H
#define GS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_OK 0
#define GS_E_INVALID (-1)   /* bad argument (null pointer, W or H not a multiple of 16, ...) */
#define GS_E_CAPACITY (-2)  /* a count exceeds what the int32 index contract allows */
#define GS_E_STATIC_LIMIT (-3) /* a *_static entry point was asked for more than its fixed scratch holds */

#define GS_TILE 16          /* R.py:27-28 */
#define GS_BOUNDARY_TILES 3 /* R.py:26 */
#define GS_FEATURES 56      /* R.py:208-236 */
#define GS_RECORD_FLOATS 12 /* packed per-splat record: u v a b c rescale alpha r g b z s; s = log2(rescale
                             * alpha 255 / (1 + band)), the backward blend's decision band (see gs_blend_bwd) */
#define GS_GRAD_SLOTS 12    /* per-splat reduced gradients, see gs_blend_bwd */

/* ABI / build identification.  Returns e.g. 100 for the first ABI revision. */
int gs_abi_version(void);
/* Human-readable text for a status returned by any entry point (static storage). */
const char *gs_status_string(int status);

/* Bytes of scratch `workspace` the forward needs for n points / m visible points / k keys. */
int64_t gs_project_workspace_bytes(int64_t n);
int64_t gs_sort_workspace_bytes(int64_t k);        /* gs_sort_pairs and gs_sort_pairs32 */
int64_t gs_point_sort_workspace_bytes(int64_t m);  /* gs_sort_points_by_depth */
int64_t gs_emit_workspace_bytes(int64_t m);        /* gs_emit_tile_keys */

/*
 * Pose inverse.  Replaces inverse_SE3_qt_torch (reference utils.py:426-432, used R.py:845).
 * q_pc_cam [n_obj,4] xyzw, t_pc_cam [n_obj,3] -> q_cam_pc, t_cam_pc.
 */
int gs_inverse_pose(const float *q_pc_cam, const float *t_pc_cam, int n_obj, float *q_cam_pc,
                    float *t_cam_pc, void *stream);

/*
 * Cull + ordered compaction + projection + SH colour + tile count + offset scan in one pass.
 * Replaces filter_point_in_camera (R.py:31-78), the mask compaction (R.py:861-865),
 * generate_point_attributes_in_camera_plane (R.py:239-315, including the in-place
 * quaternion normalisation of `features`), generate_num_overlap_tiles (R.py:106-128)
 * and the cumsum/cat (R.py:913-922).
 *
 * Outputs are in ascending point-id order (decoupled look-back scan), each array
 * sized for n entries; only the first M are written.
 *   What the path itself reads (required): ids [n] i32, point_in_camera [n,3], records [n,12] (the
 *   packed per-splat record both blend kernels stage, see GS_RECORD_FLOATS), bbox [n,4] u16
 *   (tu0,tu1,tv0,tv1), depth_key [n] i32, totals [2] i64 = {M, K}.
 *   The reference-named per-point buffers (R.py:873-884), each NULLABLE (NULL = not written):
 *   point_uv [n,2], conic [n,4], alpha [n], color [n,3], radii [n], num_tiles [n] i32 -- the operator
 *   asks for point_uv / num_tiles only when a backward hook will show them, tests ask for all --
 *   and tile_offsets [n] i64 (exclusive scan of num_tiles), which only gs_emit_keys reads.  With
 *   tile_offsets the look-back word packs 27 bits of points and 35 bits of tiles: GS_E_CAPACITY if
 *   n >= 2^27 or n x (tiles per image) >= 2^35 (a carry could otherwise corrupt the point count).
 *   slot_of_point [n] i32 (NULLABLE): for EVERY point its position in the visible list, or -1 -- the
 *   inverse of ids, read by gs_point_bwd_batch.
 * features: 16-byte aligned (rows are fetched with cp.async.bulk); GS_E_INVALID otherwise.
 * workspace: gs_project_workspace_bytes(n) bytes, contents irrelevant on entry.
 */
int gs_project_cull_sh(const float *xyz, float *features, const int8_t *invalid_mask,
                       const int32_t *object_id, int64_t n, const float *intrinsics /*3x3*/,
                       const float *q_cam_pc, const float *t_cam_pc, int n_obj,
                       float near_plane, float far_plane, int width, int height,
                       float depth_to_sort_key_scale,
                       int32_t *ids, float *point_uv, float *point_in_camera, float *conic,
                       float *alpha, float *color, float *radii, int32_t *num_tiles,
                       int64_t *tile_offsets, float *records, uint16_t *bbox,
                       int32_t *depth_key, int32_t *slot_of_point, int64_t *totals, void *workspace,
                       void *stream);

/*
 * gs_project_cull_sh without a caller-allocated workspace, for callers that cannot allocate (code under
 * CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE
 * REFERENCE.  Same arguments, meaning and results as
 * gs_project_cull_sh minus `workspace`: the look-back scratch lives in a file-scope device array of the
 * library (8 bytes per 128 points, 1 MiB; every GPU that loads the library holds it until exit),
 * whose device address is resolved once per device.  Calls no cudaMalloc / cudaFree / allocating
 * routine.  n > GS_PROJECT_STATIC_MAX_POINTS returns GS_E_STATIC_LIMIT.  All calls on a device share the
 * array; each call clears what it uses first (in stream order), so calls in sequence are independent,
 * but calls that OVERLAP on the GPU (two streams without a dependency) race on it -- serialise them.
 */
#define GS_PROJECT_STATIC_MAX_POINTS (1ll << 24)
int gs_project_cull_sh_static(const float *xyz, float *features, const int8_t *invalid_mask,
                              const int32_t *object_id, int64_t n, const float *intrinsics /*3x3*/,
                              const float *q_cam_pc, const float *t_cam_pc, int n_obj,
                              float near_plane, float far_plane, int width, int height,
                              float depth_to_sort_key_scale,
                              int32_t *ids, float *point_uv, float *point_in_camera, float *conic,
                              float *alpha, float *color, float *radii, int32_t *num_tiles,
                              int64_t *tile_offsets, float *records, uint16_t *bbox,
                              int32_t *depth_key, int32_t *slot_of_point, int64_t *totals,
                              void *stream);

/*
 * Key/payload emission.  Replaces generate_point_sort_key_by_num_overlap_tiles (R.py:131-172).
 * keys [k] i64 = depth_key + (tile_id << 32), payload [k] i32 = offset in the visible list.
 */
int gs_emit_keys(const uint16_t *bbox, const int32_t *depth_key, const int64_t *tile_offsets,
                 int64_t m, int64_t k, int width, int64_t *keys, int32_t *payload, void *stream);

/*
 * Stable ascending sort of (key, payload) pairs.  Replaces Tensor.sort() + gather (R.py:947-950).
 * Only the significant bits are sorted: `depth_bits` low bits and `tile_bits` bits from bit 32
 * (pass depth_bits = 32, tile_bits = 32 for a full signed 64-bit sort).  keys_alt/payload_alt
 * are same-sized ping-pong buffers.  *result_in_alt (host int) tells where the result is.
 */
int gs_sort_pairs(int64_t *keys, int32_t *payload, int64_t *keys_alt, int32_t *payload_alt,
                  int64_t k, int depth_bits, int tile_bits, void *workspace,
                  int *result_in_alt /*host*/, void *stream);

/*
 * Depth-first key pipeline (the default product path when depth keys are non-negative, i.e.
 * near_plane >= 0).  Together the next three calls replace generate_point_sort_key_by_num_overlap_tiles
 * (R.py:131-172), Tensor.sort() + gather (R.py:947-950) and find_tile_start_and_end (R.py:175-193):
 * the resulting payload order and tile ranges equal those of a stable sort of the reference's
 * 64-bit keys (ties in emission order), at a third of the memory traffic.
 *
 * gs_sort_points_by_depth: order [m] i32 = row indices of the visible list sorted by depth_key
 *   (stable; depth_key >= 0 must fit `depth_bits` <= 31 bits).  If m_dev is not null, m is only a
 *   capacity (buffers and workspace sized for it) and the count is read on the device from *m_dev
 *   -- the host can enqueue the sort before it has read the totals of gs_project_cull_sh back.
 * gs_emit_tile_keys: walks `order`, and for every tile of every point's bounding box writes
 *   tile_ids [k] u32 and payload [k] i32 (= row index), contiguous per point, points in `order`.
 *   digit_hist (nullable) [4 * 256] u32: if given it is zeroed and receives, per 8-bit digit p of the
 *   low key_bits bits, how many emitted ids have each digit value (bins p * 256 + value) -- the
 *   histogram gs_sort_pairs32 would otherwise compute with one more pass over the ids.
 * gs_sort_pairs32: stable ascending sort of (u32 key, i32 payload) by the low `key_bits` bits;
 *   ping-pong buffers and *result_in_alt as gs_sort_pairs.  digit_hist (nullable): the digit
 *   histogram of the keys in the layout above, if the caller already has it.
 * gs_tile_ranges32: as gs_tile_ranges, on sorted tile ids.
 * gs_sorted_keys: rebuilds the reference's sorted 64-bit keys (depth_key[payload] + (tile << 32),
 *   R.py:159-160) for inspection and the parity tests; not on the product path.
 */
int gs_sort_points_by_depth(const int32_t *depth_key, int64_t m, const int64_t *m_dev, int depth_bits,
                            int32_t *order, void *workspace, void *stream);
int gs_emit_tile_keys(const int32_t *order, const uint16_t *bbox, int64_t m, int64_t k, int width,
                      uint32_t *tile_ids, int32_t *payload, void *workspace, uint32_t *digit_hist,
                      int key_bits, void *stream);
int gs_sort_pairs32(uint32_t *keys, int32_t *payload, uint32_t *keys_alt, int32_t *payload_alt,
                    int64_t k, int key_bits, void *workspace, const uint32_t *digit_hist,
                    int *result_in_alt /*host*/, void *stream);
int gs_tile_ranges32(const uint32_t *sorted_tile_ids, int64_t k, int32_t *tile_start,
                     int32_t *tile_end, int num_tiles, void *stream);
int gs_sorted_keys(const uint32_t *sorted_tile_ids, const int32_t *sorted_payload,
                   const int32_t *depth_key, int64_t k, int64_t *keys, void *stream);

/*
 * Tile-binned key pipeline (the default product path for images up to 4096 x 4096 and depth keys
 * that fit 31 bits).  After gs_sort_points_by_depth, ONE call replaces
 * generate_point_sort_key_by_num_overlap_tiles (R.py:131-172), Tensor.sort() + gather
 * (R.py:947-950) and find_tile_start_and_end (R.py:175-193): the depth-ordered points are split,
 * stably, first over the tile columns and then over the tile rows they cover, so that no (key,
 * payload) pair per list entry is ever written or sorted.  payload [k] i32, tile_start / tile_end
 * [tiles] i32 are bit-identical to those of the pair-sort chain above (and so to a stable sort of the
 * reference's 64-bit keys); empty tiles read (0, 0) as in the reference.
 *   entries: scratch, gs_bin_entries_capacity(m, k, width) x 8 bytes; workspace:
 *   gs_bin_workspace_bytes.  GS_E_CAPACITY if the image has more than 256 tile columns or rows
 *   or k >= 2^30 (callers then use the pair-sort chain).
 */
int64_t gs_bin_workspace_bytes(int64_t m, int64_t k, int width, int height);
int64_t gs_bin_entries_capacity(int64_t m, int64_t k, int width);
int gs_bin_tiles(const int32_t *order, const uint16_t *bbox, int64_t m, int64_t k, int width, int height,
                 uint64_t *entries, int32_t *payload, int32_t *tile_start, int32_t *tile_end,
                 void *workspace, void *stream);

/*
 * Tile ranges.  Replaces find_tile_start_and_end (R.py:175-193).  start/end [num_tiles] i32
 * are zeroed by this call (R.py:954-957) and then filled.
 */
int gs_tile_ranges(const int64_t *sorted_keys, int64_t k, int32_t *tile_start, int32_t *tile_end,
                   int num_tiles, void *stream);

/*
 * Forward blend.  Replaces gaussian_point_rasterisation (R.py:318-485).
 * image [H,W,3], depth [H,W], acc_alpha [H,W], last_effective [H,W] i32, valid_count [H,W] i32.
 * Every pixel is written (empty tiles give zeros).  With rgb_only != 0 only image is written.
 * What a following gs_blend_bwd needs from the forward (both nullable, both or neither):
 *   cell_masks [4, mask_stride] u8, mask_stride >= k: for every list entry the forward visits, which
 *     8x4-pixel cells of each 8x8 quarter of its tile the splat can reach (plane q = 2 * column + row
 *     half; bit 0 = upper cell, bit 1 = lower); entries the forward never reached (every pixel of the
 *     quarter already saturated) are left unwritten and are not read;
 *   quarter_flags [tiles, 4] u8: 1 if the pixels of that 8x8 quarter reach a list entry with an
 *     irregular conic (not positive definite) or a non-finite field, else 0 (every byte is written).
 */
int gs_blend_fwd(int height, int width, const int32_t *tile_start, const int32_t *tile_end,
                 const int32_t *sorted_payload, const float *records, float *image, float *depth,
                 float *acc_alpha, int32_t *last_effective, int32_t *valid_count, int rgb_only,
                 uint8_t *cell_masks, int64_t mask_stride, uint8_t *quarter_flags, void *stream);

/*
 * Backward blend.  Replaces loop 1 of gaussian_point_rasterisation_backward (R.py:531-705).
 * splat_grads [m, 12] f32 must be zero on entry; per visible point it receives
 *   0:duv.u 1:duv.v 2:dcov00 3:dcov01 4:dcov11 5:dr 6:dg 7:db 8:dlogit 9:|duv| 10:#pixels(as float) 11:0
 * map_on_image [H,W,2] gets the per-pixel sum of |duv| (R.py:700-704).  It may be null: the
 * magnitudes (the map and slot 9) are observable only through the reference's backward hook, and
 * a caller without one skips them (slot 9 then stays 0).
 * cell_masks, quarter_flags: REQUIRED, what gs_blend_fwd wrote for the same tile lists (the backward
 * never redoes the culling arithmetic).  scratch: gs_blend_bwd_scratch_bytes(H, W) bytes, contents
 * irrelevant on entry (one flag byte per pixel: which pixels the literal pass re-does).
 * records: as written by gs_project_cull_sh, INCLUDING the 12th float s of every record: with
 *   rho = |b| / sqrt(a c), tau = max(2 ln(255 rescale alpha), 0),
 *   band = min(3e-6 + 6e-7 tau (1 + rho) / (1 - rho), 0.02),   s = log2(255 rescale alpha / (1 + band))
 * (0 for records that are irregular or can never reach alpha = 1/255).  band bounds the relative difference
 * between this kernel's fast alpha and the reference's fp32 alpha near the threshold; pairs inside it
 * are re-decided with the reference's arithmetic.
 * Which (pixel, entry) pairs take part is decided as the reference decides it (alpha >= 1/255 on its
 * fp32 backward arithmetic, R.py:626-634): slot 10 equals the reference's in_camera_num_affected_pixels.
 */
int64_t gs_blend_bwd_scratch_bytes(int height, int width);
int gs_blend_bwd(int height, int width, const int32_t *tile_start, const int32_t *tile_end,
                 const int32_t *sorted_payload, const float *records, const float *image_grad,
                 const float *acc_alpha, const int32_t *last_effective, float *splat_grads,
                 float *map_on_image, const uint8_t *cell_masks, int64_t mask_stride,
                 const uint8_t *quarter_flags, void *scratch, void *stream);

/*
 * Per-point chain rule + SH band mask + factor scaling.  Replaces loop 2 of the backward kernel
 * (R.py:707-772), _clear_grad_by_color_max_sh_band (R.py:1167-1182) and the factor multiplies
 * (R.py:1105-1125).  With accumulate == 0, grad_xyz [n,3] and grad_features [n,56] must be zero on
 * entry and the rows of the m visible points are stored; with accumulate != 0 those rows are added
 * to what the buffers hold (several views summed into one gradient buffer, no zero fill per view).
 * Optional hook outputs (may be null), all [m,...] in list order:
 * hook_grad_xyz [m,3], hook_grad_features [m,56], hook_grad_viewspace [m,2], hook_magnitude [m],
 * hook_num_affected [m] i32 (R.py:1127-1142).
 */
int gs_point_bwd(const float *xyz, const float *features, const int32_t *object_id,
                 const float *intrinsics, const float *q_cam_pc, const float *t_cam_pc,
                 const float *t_pc_cam, int n_obj, const int32_t *ids, int64_t m,
                 const float *point_in_camera, const float *splat_grads, int color_max_sh_band,
                 float q_factor, float s_factor, float alpha_factor, float color_factor,
                 float high_order_color_factor, float *grad_xyz, float *grad_features,
                 float *hook_grad_xyz, float *hook_grad_features, float *hook_grad_viewspace,
                 float *hook_magnitude, int32_t *hook_num_affected, int accumulate, void *stream);

/*
 * The per-point chain rule of a whole batch of views in one pass over the points (the view-sharded
 * step; not in the reference, which renders one view per step).  Equivalent to
 * gs_point_bwd with accumulate != 0 called once per view on zeroed buffers, but every point's feature
 * row is read once, what depends on the view is summed in registers, and the gradient rows of ALL points
 * of the range [first_point, first_point + num_points) are STORED once (zeros for points no view sees):
 * no zero fill, no read-modify-write.  Calls over disjoint ranges are independent, so a step can be cut
 * into chunks that are exchanged (all-reduced) while the next chunk is computed.
 * views_device: DEVICE array of num_views gs_view_t (16-byte aligned), one per view of the batch, all
 * of whose pointers are device pointers into that view's forward / backward-blend state.
 * Hook outputs are not produced: a batch with a backward hook uses gs_point_bwd per view.
 */
typedef struct gs_view_t {
    const int32_t *slot_of_point;  /* [n]   gs_project_cull_sh: position in the visible list or -1 */
    const float *point_in_camera;  /* [m,3] gs_project_cull_sh */
    const float *splat_grads;      /* [m,12] gs_blend_bwd */
    const float *intrinsics;       /* [3,3] */
    const float *q_cam_pc;         /* [n_obj,4] gs_inverse_pose */
    const float *t_pc_cam;         /* [n_obj,3] the INPUT translation (R.py:731-732,749) */
    int64_t reserved[2];           /* pads the record to 64 bytes */
} gs_view_t;
int gs_point_bwd_batch(const float *xyz, const float *features, const int32_t *object_id,
                       int64_t first_point, int64_t num_points, const gs_view_t *views_device,
                       int num_views, int color_max_sh_band, float q_factor, float s_factor,
                       float alpha_factor, float color_factor, float high_order_color_factor,
                       float *grad_xyz, float *grad_features, void *stream);

/*
 * gs_point_bwd_batch for ONE view over all n points, the view's record passed from the HOST (it travels as
 * a kernel argument: no device table).  Replaces, for a plain backward without a hook, the zero fill of the
 * N x 3 / N x 56 gradient tensors (R.py:1051-1058) + gs_point_bwd: every row is stored once, zeros for the
 * points the view does not see.
 */
int gs_point_bwd_all_points(const float *xyz, const float *features, const int32_t *object_id, int64_t n,
                            const gs_view_t *view /* host */, int color_max_sh_band, float q_factor,
                            float s_factor, float alpha_factor, float color_factor,
                            float high_order_color_factor, float *grad_xyz, float *grad_features, void *stream);

/*
 * gs_point_bwd_batch over ALL n points, fused with the step's one exchange: the sum of the N x 59
 * gradient over the ranks of a view-sharded step (not in the reference, which is
 * single-GPU), done in the switch over NVLink multicast while the pass is still running.  One process
 * per GPU; every rank calls it once per step with the same n, chunk count and epoch.
 *   The gradient buffer is symmetric memory of 59 n floats rounded up to a multiple of 4, plus 4: feature
 *   rows first ([n,56]), then positions ([n,3]); `local_base` is this rank's mapping, `multicast_base` the
 *   multicast mapping of the same allocation (16-byte aligned).  `flags[r]` is rank r's flag array
 *   (gs_exchange_flag_words(chunks, world) u32, symmetric, zero before the first call) as mapped into
 *   THIS process; `scratch` is local device memory of gs_exchange_scratch_bytes(chunks).  `epoch` must be
 *   > 0 and strictly increasing from call to call on a flag array (nothing is reset between steps).
 *   chunks <= number of 128-point blocks (4 measured best).  reducer_ctas: CTAs that run the reduction
 *   (0 = the measured optimum: 72 for 2 ranks, 48 up to 4, 32 beyond).
 * On return of the KERNEL (stream order) local_base holds the summed gradient on every rank.  A rank that
 * waits ~10 s for a peer gives up, records a non-zero word at scratch[1] and finishes without hanging.
 * How the host gets the mappings is its business (torch.distributed._symmetric_memory in this package).
 */
typedef struct gs_exchange_t {
    void *multicast_base;
    float *local_base;
    void *flags[16];
    int32_t rank, world, chunks, reducer_ctas;
    uint32_t epoch;
    uint32_t reserved;
    void *scratch;
} gs_exchange_t;
int64_t gs_exchange_flag_words(int chunks, int world);
int64_t gs_exchange_scratch_bytes(int chunks);
int gs_point_bwd_batch_exchange(const float *xyz, const float *features, const int32_t *object_id,
                                int64_t n, const gs_view_t *views_device, int num_views,
                                int color_max_sh_band, float q_factor, float s_factor, float alpha_factor,
                                float color_factor, float high_order_color_factor,
                                const gs_exchange_t *exchange /* host */, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Loss front-end: what runs between the rasteriser's image and the
 * cotangent handed back to gs_blend_bwd.  Replaces, per iteration, GaussianPointTrainer.py:172-182
 * (`torch.clamp(image, 0, 1)`, `permute(2, 0, 1)`, `loss.backward()`) and LossFunction.py:20-52
 * (`L = (1 - lambda) L1 + lambda (1 - ssim(...))`, scale regulariser), including
 * pytorch_msssim.ssim(X, Y, data_range=1, size_average=True): 11-tap window, sigma 1.5, valid
 * filtering, K1 = 0.01, K2 = 0.03.
 *
 * `image` and `gt` are 3-channel fp32 images addressed by element strides (h, w, c), so both the
 * rasteriser's H x W x 3 output and a permuted 3 x H x W view are read in place.  The clamp to
 * [0, 1] is applied on load (a no-op for an image that is already clamped).  H, W >= 11.
 * --------------------------------------------------------------------------------------------- */
int64_t gs_loss_workspace_bytes(int height, int width); /* `workspace` of gs_loss_fwd */
int64_t gs_loss_maps_bytes(int height, int width);      /* `maps`: 9 (H-10)(W-10) floats */

/*
 * Forward.  scalars [6] f32 receives
 *   0: L   1: L1   2: 1 - SSIM   3: regulariser value (0 if absent)   4: valid points   5: SSIM
 * (LossFunction.forward returns (0, 1, 2)).  `maps` receives the three SSIM derivative maps
 * gs_loss_bwd consumes.  With features != null the scale regulariser (LossFunction.py:40-52:
 * mean over points with invalid_mask == 0 of ||exp(features[:, 4:7])||_2) times reg_weight is
 * added to L; features is [n, 56] f32, invalid_mask [n] i8.
 */
int gs_loss_fwd(const float *image, int64_t image_stride_h, int64_t image_stride_w,
                int64_t image_stride_c, const float *gt, int64_t gt_stride_h, int64_t gt_stride_w,
                int64_t gt_stride_c, int height, int width, float lambda_value,
                const float *features, const int8_t *invalid_mask, int64_t n, float reg_weight,
                float *maps, void *workspace, float *scalars, void *stream);

/*
 * Backward of the image terms: grad_image [H, W, 3] f32 (contiguous, the rasteriser's layout)
 * receives upstream * dL/d(image), zero where the clamp is active.  `upstream` is a device scalar
 * (the cotangent of L) or null for 1.
 */
int gs_loss_bwd(const float *image, int64_t image_stride_h, int64_t image_stride_w,
                int64_t image_stride_c, const float *gt, int64_t gt_stride_h, int64_t gt_stride_w,
                int64_t gt_stride_c, int height, int width, float lambda_value, const float *maps,
                const float *upstream, float *grad_image, void *stream);

/*
 * Backward of the scale regulariser into columns 4:7 of grad_features [n, 56] (stored, or added
 * with accumulate != 0; other columns are not touched).  `scalars` is gs_loss_fwd's output (slot
 * 4 is read).
 */
int gs_scale_reg_bwd(const float *features, const int8_t *invalid_mask, int64_t n, float reg_weight,
                     const float *scalars, const float *upstream, float *grad_features,
                     int accumulate, void *stream);

/* ---------------------------------------------------------------------------------------------
 * After the backward.
 * --------------------------------------------------------------------------------------------- */

/*
 * Densification statistics.  Replaces the body of GaussianPointAdaptiveController.update
 * (GaussianPointAdaptiveController.py:130-146): for each of the m visible points (hook payload, list
 * order; ids distinct) with id = ids[i]
 *   acc_num_in_camera[id] += 1;  acc_num_pixels[id] += num_affected[i];
 *   acc_view_grad[id] += magnitude[i];
 *   acc_view_grad_avg[id] += magnitude[i] / num_affected[i]   (NaN counted as 0);
 *   acc_pos_grad[id, 0:3] += grad_xyz[i, 0:3];  acc_pos_grad_norm[id] += ||grad_xyz[i]||_2.
 * The accumulators are the controller's [n] / [n,3] tensors (AC.py:113-125).
 */
int gs_controller_accumulate(const int32_t *ids, int64_t m, const int32_t *num_affected,
                             const float *magnitude, const float *grad_xyz,
                             int32_t *acc_num_in_camera, int32_t *acc_num_pixels,
                             float *acc_view_grad, float *acc_view_grad_avg, float *acc_pos_grad,
                             float *acc_pos_grad_norm, void *stream);

/*
 * One Adam update of `count` fp32 elements, in place.  Replaces torch.optim.Adam.step for the
 * trainer's two optimisers (GaussianPointTrainer.py:131-134,183-184; amsgrad off, no weight decay):
 *   m += (1 - beta1) (g - m);  v = beta2 v + (1 - beta2) g^2;
 *   p -= lr / (1 - beta1^step) * m / (sqrt(v) / sqrt(1 - beta2^step) + eps).
 * `step` is the 1-based update count.  `grad` is read in place (e.g. a slice of the all-reduced
 * N x 59 gradient buffer).  row_mask (nullable): [count / row_len] i8; elements of rows whose
 * byte is non-zero are left untouched (not reference behaviour; null = update everything).
 */
int gs_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, int64_t count,
                 float lr, float beta1, float beta2, float eps, int64_t step, const int8_t *row_mask,
                 int row_len, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GS_B200_H */
