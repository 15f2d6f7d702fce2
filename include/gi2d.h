/*
 * gi2d.h -- C ABI of libgi2d.so: the B200-native (sm_100a) GaussianImage / gsplat 2D render path.
 *
 * Drop-in boundary.  The reference binds its native code with pybind11 (gsplat/gsplat/cuda/csrc/ext.cpp:4-24);
 * each function below replaces one of those bindings (or a torch op the reference's Python layer calls between
 * them).  Plain pointers and sizes only -- no torch types.  All pointers are DEVICE pointers unless stated
 * otherwise; inputs are borrowed, outputs are caller-allocated and FULLY WRITTEN by the call (no pre-zeroing
 * needed -- the reference's torch::zeros fills are folded into the kernels).  Every call is asynchronous on
 * `stream` (the reference launches on the legacy default stream, quirk Q9) and returns
 *   0                       success (launch accepted)
 *   GI2D_ERR_INVALID   (1)  invalid argument
 *   GI2D_ERR_WORKSPACE (2)  workspace too small
 *   GI2D_ERR_STATIC_LIMIT (3)  size beyond what the library's static scratch holds (only gi2d_scan_tiles_static)
 *   1000 + cudaError_t      CUDA launch error
 * Tiles are 16x16 pixels (reference config.h:1-4; helpers.cuh:36-41 hard-codes 16, quirk Q3):
 * TX = ceil(W/16), TY = ceil(H/16).  Layouts are the reference's: (N,2)/(N,3) fp32 row-major, images (H,W,3) fp32,
 * tile_bins (rows,2) int32, keys int64 = (tile << 32) | depth_bits.
 *
 * The reference-side binding a maintainer would add (ctypes) is shown in INTEGRATION.md.
 */
#ifndef GI2D_H
#define GI2D_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st *cudaStream_t;
#endif

#define GI2D_ABI_VERSION 2
#define GI2D_ERR_INVALID 1
#define GI2D_ERR_WORKSPACE 2
#ifndef GI2D_ERR_STATIC_LIMIT
#define GI2D_ERR_STATIC_LIMIT 3 /* problem larger than the library's static scratch (gi2d_scan_tiles_static) */
#endif
#define GI2D_ERR_CUDA_BASE 1000

int gi2d_abi_version(void);
/* Static description of a status code returned by any function below. */
const char *gi2d_status_string(int status);

/* Helper for callers that replay a captured CUDA graph of a whole iteration (no counterpart in the reference, whose trainer
 * runs eagerly): replaces every small 1-D MEMSET node of `graph` (a cudaGraph_t; height 1, element size 1/2/4, at most
 * max_bytes) by a kernel node that writes the same bytes, with the node's dependencies and dependents.  Memset nodes between
 * kernel nodes -- e.g. the semaphore clear in front of ATen's reduction kernels, which F.mse_loss records -- cost several
 * microseconds of idle time on either side in a replay; a fill kernel does not.  The graph must not have been instantiated
 * yet (torch: CUDAGraph(keep_graph=True), raw_cuda_graph(), then instantiate()).  *n_replaced = nodes replaced. */
int gi2d_graph_fold_memsets(void *graph, size_t max_bytes, int *n_replaced);

/* ---- projection ------------------------------------------------------------------------------------------- */

/* Replaces project_gaussians_2d_forward (ext.cpp:16 -> bindings.cu:781-839 -> foward2d.cu:12-69).
 * means2d (N,2) in (-1,1), L (N,3) = (l11,l21,l22).  Outputs: xys (N,2), depths (N) == 0, radii (N) int32,
 * conics (N,3), num_tiles_hit (N) int32.  `clip_thresh` of the reference is accepted-and-ignored there (Q11),
 * so it is not part of this signature. */
int gi2d_project_cholesky_fwd(int N, const float *means2d, const float *L, int H, int W, int TX, int TY, float *xys,
                              float *depths, int *radii, float *conics, int *num_tiles_hit, cudaStream_t stream);

/* Codec decode prologue fused into the Cholesky projection.  Replaces, for decompress_wo_ec / decompress
 * (gaussianimage_cholesky.py:161-166, :214-218), the torch ops
 *     means = tanh(xyz.float());  cholesky = codes * scale + beta (quantize.py:86-87) + cholesky_bound;
 *     colors = sum_k codebook_k[index[:, k]] (quantize.py:146-150)
 * and the project_gaussians_2d call that follows.  xyz_half (N,2) fp16; codes (N,3) uint8 or fp32 (codes_are_f32);
 * scale3 / beta3 / bound3: 3 fp32 each (DEVICE); index (N,n_quantizers) int32 or int64 (index_is_i64), values in
 * [0, codebook_size) (clamped into range, where torch would raise); codebooks (n_quantizers, codebook_size, 3) fp32.
 * Outputs as gi2d_project_cholesky_fwd plus colors (N,3).  xys / conics / colors must be 16-byte aligned. */
int gi2d_decode_project_cholesky_fwd(int N, const void *xyz_half, const void *codes, int codes_are_f32, const float *scale3,
                                     const float *beta3, const float *bound3, const void *index, int index_is_i64,
                                     int n_quantizers, int codebook_size, const float *codebooks, int H, int W, int TX, int TY,
                                     float *xys, float *depths, int *radii, float *conics, int *num_tiles_hit, float *colors,
                                     cudaStream_t stream);

/* Replaces project_gaussians_2d_scale_rot_forward (ext.cpp:18 -> bindings.cu:841-899 -> foward2d.cu:71-123).
 * scales2d (N,2), rotation (N,1) radians. */
int gi2d_project_scale_rot_fwd(int N, const float *means2d, const float *scales2d, const float *rotation, int H, int W,
                               int TX, int TY, float *xys, float *depths, int *radii, float *conics,
                               int *num_tiles_hit, cudaStream_t stream);

/* Either projection (mode 0: p_a = L (N,3), p_b unused; mode 1: p_a = scales2d (N,2), p_b = rotation (N,1)) that ALSO
 * publishes sum(num_tiles_hit) -- the number compute_cumulative_intersects reads back with `.item()`
 * (gsplat/gsplat/utils.py:116-118, quirk Q12) -- as a by-product of the kernel: to *total_dev (device) and / or to
 * *total_host, MAPPED PINNED HOST memory that the caller presets to a negative value and polls.  By the time the caller
 * enters rasterize_gaussians_sum the word has normally arrived, so the binning workspace can be sized exactly without a
 * stream synchronisation and without an overflow / retry protocol.  Either pointer may be NULL. */
int gi2d_project_fwd_total(int mode, int N, const float *means2d, const float *p_a, const float *p_b, int H, int W, int TX, int TY,
                           float *xys, float *depths, int *radii, float *conics, int *num_tiles_hit, int *total_dev,
                           int *total_host, cudaStream_t stream);

/* Replaces project_gaussians_2d_backward (ext.cpp:17 -> bindings.cu:902-949 -> backward2d.cu:8-51), including the
 * reference's doubled off-diagonal term (Q5).  v_cov2d (N,3) may be NULL (the Python layer discards it).
 * Outputs v_mean2d (N,2), v_L (N,3); rows with radii <= 0 are zero. */
int gi2d_project_cholesky_bwd(int N, const float *L, int H, int W, const int *radii, const float *conics,
                              const float *v_xy, const float *v_conic, float *v_cov2d, float *v_mean2d, float *v_L,
                              cudaStream_t stream);

/* Replaces project_gaussians_2d_scale_rot_backward (ext.cpp:19 -> bindings.cu:951-1004 -> backward2d.cu:53-109).
 * Outputs v_mean2d (N,2), v_scale (N,2), v_rot (N,1). */
int gi2d_project_scale_rot_bwd(int N, const float *scales2d, const float *rotation, int H, int W, const int *radii,
                               const float *conics, const float *v_xy, const float *v_conic, float *v_cov2d,
                               float *v_mean2d, float *v_scale, float *v_rot, cudaStream_t stream);

/* The two projection backward passes reading v_xy / v_conic directly from the (N,12) gradient rows that
 * gi2d_raster_sum_bwd_packed writes (columns 0-1 and 2-4), i.e. without materialising (N,2)/(N,3) tensors. */
int gi2d_project_cholesky_bwd_packed(int N, const float *L, int H, int W, const int *radii, const float *conics,
                                     const float *grads12, float *v_mean2d, float *v_L, cudaStream_t stream);
int gi2d_project_scale_rot_bwd_packed(int N, const float *scales2d, const float *rotation, int H, int W, const int *radii,
                                      const float *conics, const float *grads12, float *v_mean2d, float *v_scale,
                                      float *v_rot, cudaStream_t stream);

/* Replaces compute_cov2d_bounds (ext.cpp:12 -> bindings.cu:21-61).  cov2d (N,3) upper triangle;
 * conics (N,3), radii (N,1) FLOAT (as in the reference). */
int gi2d_cov2d_bounds(int N, const float *cov2d, float *conics, float *radii, cudaStream_t stream);

/* ---- binning, staged and reference-typed (the public gsplat.utils functions) ---------------------------------- */

/* Replaces torch.cumsum(int32) of compute_cumulative_intersects (gsplat/gsplat/utils.py:116-118).
 * cum_tiles_hit (N) int32 inclusive scan; the total is copied (4 bytes, async on `stream`) to *total_out, which may
 * be a device pointer or a pinned host pointer (the caller synchronises before reading it, like `.item()` does). */
size_t gi2d_scan_workspace_bytes(int N);
int gi2d_scan_tiles(int N, const int *num_tiles_hit, int *cum_tiles_hit, int *total_out, void *ws, size_t ws_bytes,
                    cudaStream_t stream);

/* gi2d_scan_tiles WITHOUT a workspace argument and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernel, same results as gi2d_scan_tiles; the scratch is a file-scope device array
 * of GI2D_SCAN_STATIC_BYTES bytes inside the library, held by every GPU that loads it until the process exits.
 * Limit: N <= GI2D_SCAN_STATIC_MAX_N, beyond it GI2D_ERR_STATIC_LIMIT is returned and nothing is launched.  The call makes
 * no cudaMalloc / cudaFree and calls no allocating library routine (one async memset of the scratch, one kernel, one
 * 4-byte async copy when total_out is given).  ALL calls on a device share the one array: calls that overlap on the GPU
 * (different streams, or graphs replayed concurrently) race -- serialise them, or use gi2d_scan_tiles with own workspaces. */
#define GI2D_SCAN_STATIC_MAX_N (1 << 22)
#define GI2D_SCAN_STATIC_BYTES 24832
int gi2d_scan_tiles_static(int N, const int *num_tiles_hit, int *cum_tiles_hit, int *total_out, cudaStream_t stream);

/* Replaces map_gaussian_to_intersects (ext.cpp:13 -> bindings.cu:274-313 -> forward.cu:100-136).
 * isect_ids (n_isect) int64, gaussian_ids (n_isect) int32. */
int gi2d_emit_isects(int N, const float *xys, const float *depths, const int *radii, const int *cum_tiles_hit, int TX,
                     int TY, int n_isect, long long *isect_ids, int *gaussian_ids, cudaStream_t stream);

/* Replaces torch.sort + torch.gather of bin_and_sort_gaussians (utils.py:164-165): stable ascending sort of the
 * int64 keys on bits [begin_bit, end_bit) carrying the int32 ids.  begin=0,end=64 is the full signed sort; in the 2D
 * path all depth bits are zero (Q2) and begin=32,end=32+ceil(log2(TX*TY)) gives the same result in 2-3 passes. */
size_t gi2d_sort_workspace_bytes(int n, int begin_bit, int end_bit);
int gi2d_sort_isects(int n, const long long *ids_in, const int *gids_in, long long *ids_out, int *gids_out,
                     int begin_bit, int end_bit, void *ws, size_t ws_bytes, cudaStream_t stream);

/* Replaces get_tile_bin_edges (ext.cpp:14 -> bindings.cu:315-330 -> forward.cu:141-163).  tile_bins (n_bins,2);
 * tile ids >= n_bins are skipped instead of written out of bounds (the reference's Q4 hazard). */
int gi2d_tile_bins(int n, const long long *isect_ids_sorted, int n_bins, int *tile_bins, cudaStream_t stream);

/* ---- binning, fused and sync-free (what rasterize_gaussians_sum runs) ------------------------------------------ */

/* cumsum + map + sort + bin edges (utils.py:116-167) in one asynchronous chain with NO device->host read (the
 * reference's `.item()`, Q12): the intersection count stays on the device, `capacity` bounds it.
 *   gids_sorted (capacity) int32,
 *   tile_bins (TX*TY,2) int32: [lo,hi) per tile; empty tiles hold (INT_MAX, 0) until gi2d_raster_sum_fwd_packed
 *     rewrites them to the reference's (0, 0) -- consumers must treat lo >= hi as empty,
 *   packed (capacity*12 floats): per sorted intersection {x,y,opacity,conic.x | conic.y,conic.z,r,g | b,gid,ex,ey}
 *     (ex, ey: half-extents in pixels of the region where the alpha test can pass -- +inf = unbounded, -1 = never; the
 *     rasterizers cull with them)
 *   status_out (2) int32: [0] = n_isect, [1] = 1 if n_isect > capacity (outputs truncated; retry with more).
 * Requires all depths == 0 (true for both 2D projections). n_isect < 2^30. */
size_t gi2d_bin_fused_workspace_bytes(int N, int TX, int TY, int capacity);
/* How many kernels gi2d_bin_fused launches for this shape on the current device: 1 for images of at most 4096 tiles
 * whose grid is co-resident (one cooperative counting-sort kernel), else 1 + one onesweep pass per 8 tile-id bits. */
int gi2d_bin_fused_kernel_count(int N, int TX, int TY);
/* The records of an existing sorted list rebuilt from other per-Gaussian attributes (conics / colours / opacities): what
 * a caller that rasterizes the SAME geometry again with other colours needs instead of a second binning pass
 * (gaussianimage_cholesky.py:81,89,95 calls rasterize_gaussians_sum three times per forward on one projection, Q13).
 * gids_sorted / status are outputs of gi2d_bin_fused for that geometry; writes min(status[0], capacity) records. */
int gi2d_repack_records(int capacity, const int *gids_sorted, const int *status, const float *xys, const float *conics,
                        const float *colors, const float *opacities, float *packed, cudaStream_t stream);
int gi2d_bin_fused(int N, int TX, int TY, const float *xys, const int *radii, const int *num_tiles_hit,
                   const float *conics, const float *colors, const float *opacities, int capacity, void *ws,
                   size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed, int *status_out,
                   cudaStream_t stream);

/* Row-sharded variant (render-only decode of very large frames, one rank per band of tile rows): only intersections
 * with tile rows [ty_begin, ty_end) are counted, emitted and sorted; tile ids and tile_bins stay global (TX*TY rows). */
int gi2d_bin_fused_rows(int N, int TX, int TY, int ty_begin, int ty_end, const float *xys, const int *radii,
                        const int *num_tiles_hit, const float *conics, const float *colors, const float *opacities,
                        int capacity, void *ws, size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed,
                        int *status_out, cudaStream_t stream);

/* Same, and additionally publishes {n_isect, overflow} to `status_host` -- MAPPED PINNED HOST memory (may be NULL) -- as
 * soon as the first kernel of the chain knows them: [1] = overflow, then (after a system-scope fence) [0] = n_isect.
 * A caller that presets status_host[0] = -1 can poll it instead of synchronising the stream; this is how
 * rasterize_gaussians_sum replaces the reference's mid-pipeline `.item()` (gsplat/gsplat/utils.py:117, quirk Q12)
 * without ever making the GPU wait for the host. */
int gi2d_bin_fused_rows_notify(int N, int TX, int TY, int ty_begin, int ty_end, const float *xys, const int *radii,
                               const int *num_tiles_hit, const float *conics, const float *colors, const float *opacities,
                               int capacity, void *ws, size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed,
                               int *status_out, int *status_host, cudaStream_t stream);

/* gi2d_bin_fused_rows_notify with optional extras, as one argument block:
 *   status_sticky (2 int32, device, may be NULL): [0] = max over all launches of n_isect (atomicMax), [1] |= overflow.
 *     Only the host clears it -- a caller replaying a captured graph polls it now and then instead of after every launch.
 *   zero_grads12 ((N,12) fp32, 16-byte aligned, may be NULL): zero-filled by the chain (it walks every Gaussian anyway):
 *     the gradient rows gi2d_raster_sum_bwd_packed_acc accumulates into, without a memset in front of the backward pass. */
typedef struct gi2d_bin_args {
    int N, TX, TY, ty_begin, ty_end, capacity;
    const float *xys;
    const int *radii, *num_tiles_hit;
    const float *conics, *colors, *opacities;
    void *ws;
    size_t ws_bytes;
    int *gids_sorted, *tile_bins;
    float *packed;
    int *status_out, *status_host, *status_sticky;
    float *zero_grads12;
} gi2d_bin_args;
int gi2d_bin_fused_ex(const gi2d_bin_args *args, cudaStream_t stream);
/* gi2d_repack_records that also zero-fills zero_grads12 ((N,12), may be NULL) like gi2d_bin_fused_ex does. */
int gi2d_repack_records_ex(int capacity, const int *gids_sorted, const int *status, const float *xys, const float *conics,
                           const float *colors, const float *opacities, float *packed, int N, float *zero_grads12,
                           cudaStream_t stream);

/* ---- accumulated-sum rasterizer --------------------------------------------------------------------------------- */

/* Replaces rasterize_sum_forward (ext.cpp:22 -> bindings.cu:400-469 -> forward.cu:512-627).
 * tile_bins must have >= TX*TY rows.  out_img (H,W,3); final_Ts (H,W) == 1 and final_idx (H,W) int32 may both be
 * NULL.  `background` of the reference is unused by its kernel (Q6) and not part of this signature. */
int gi2d_raster_sum_fwd(int TX, int TY, int W, int H, const int *gids_sorted, const int *tile_bins, const float *xys,
                        const float *conics, const float *colors, const float *opacities, float *out_img,
                        float *final_Ts, int *final_idx, cudaStream_t stream);

/* Same over the packed records of gi2d_bin_fused (bulk-TMA staging).  If d_status != NULL and d_status[0] == 0 the
 * image is filled with background[0..2], which is what rasterize_sum.py:121-129 returns for zero intersections.
 * Also rewrites the (INT_MAX, 0) markers of empty tiles in tile_bins to (0, 0). */
int gi2d_raster_sum_fwd_packed(int TX, int TY, int W, int H, int *tile_bins, const float *packed,
                               const int *d_status, const float *background, float *out_img, float *final_Ts,
                               int *final_idx, cudaStream_t stream);

/* Row-sharded variant: rasterizes only the tiles of rows [ty_begin, ty_end) into the same full-size out_img. */
int gi2d_raster_sum_fwd_packed_rows(int TX, int TY, int W, int H, int ty_begin, int ty_end, int *tile_bins,
                                    const float *packed, const int *d_status, const float *background, float *out_img,
                                    float *final_Ts, int *final_idx, cudaStream_t stream);

/* Row-sharded variant with the all-gather FUSED into the rasterizer: `out_img_mc` is the NVLS multicast address of a
 * symmetric frame buffer (same layout as out_img on every rank; e.g. torch symmetric memory's `multicast_ptr`); every
 * pixel store is a `multimem.st`, which the NVSwitch replicates into the buffer of every rank of the multicast group
 * (including this one).  out_img (this rank's own replica) is then only used for alignment checks.  The caller
 * synchronises the ranks (a barrier after the kernel) before anybody reads the frame.  NULL = plain local stores. */
int gi2d_raster_sum_fwd_packed_rows_mc(int TX, int TY, int W, int H, int ty_begin, int ty_end, int *tile_bins,
                                       const float *packed, const int *d_status, const float *background, float *out_img,
                                       float *out_img_mc, float *final_Ts, int *final_idx, cudaStream_t stream);

/* Row-sharded variant with the all-gather fused into the rasterizer by PEER STORES over NVLink: every full-tile row goes to
 * out_img and to the same offset of each of the n_peers (<= 7) frame buffers of the other ranks, `peer_imgs` being a HOST
 * array of peer-mapped DEVICE addresses (e.g. torch symmetric memory's buffer_ptrs).  W % 4 == 0 and 16-byte aligned
 * buffers are required when n_peers > 0; the image must consist of whole tiles (ragged tiles are stored locally only). */
int gi2d_raster_sum_fwd_packed_rows_p2p(int TX, int TY, int W, int H, int ty_begin, int ty_end, int *tile_bins, const float *packed,
                                        float *out_img, float *const *peer_imgs, int n_peers, cudaStream_t stream);

/* Replaces rasterize_sum_backward (ext.cpp:23 -> bindings.cu:706-779 -> backward.cu:696-862).
 * v_out (H,W,3).  Outputs v_xy (N,2), v_conic (N,3), v_colors (N,3), v_opacity (N,1) are zeroed by the call.
 * final_idx is accepted for parity and may be NULL.  fp32 atomic accumulation (order-nondeterministic, as in the
 * reference). */
int gi2d_raster_sum_bwd(int N, int TX, int TY, int W, int H, const int *gids_sorted, const int *tile_bins,
                        const float *xys, const float *conics, const float *colors, const float *opacities,
                        const int *final_idx, const float *v_out, float *v_xy, float *v_conic, float *v_colors,
                        float *v_opacity, cudaStream_t stream);

/* Same over the packed records; `grads12` is one 16-byte-aligned buffer of (N,12) floats, zeroed by the call, one
 * 48-byte row per Gaussian: {v_xy.x, v_xy.y, v_conic.x, v_conic.y | v_conic.z, v_colors.rgb | v_opacity, 0, 0, 0}. */
int gi2d_raster_sum_bwd_packed(int N, int TX, int TY, int W, int H, const int *tile_bins, const float *packed,
                               const float *v_out, float *grads12, cudaStream_t stream);

/* Same WITHOUT the zero fill (grads12 was zeroed by gi2d_bin_fused_ex / gi2d_repack_records_ex); v_scale: optional DEVICE
 * scalar multiplied into v_out (NULL = 1). */
int gi2d_raster_sum_bwd_packed_acc(int N, int TX, int TY, int W, int H, const int *tile_bins, const float *packed,
                                   const float *v_out, const float *v_scale, float *grads12, cudaStream_t stream);

/* ---- loss epilogue --------------------------------------------------------------------------------- */

/* gi2d_raster_sum_fwd_packed that ends in the training loss the reference's callers compute next: image =
 * clamp(out_img, 0, 1) in the target's layout (gaussianimage_cholesky.py:84-85), loss = F.mse_loss(image, gt)
 * (utils.py:20-25).  gt: (3,H,W) planes when gt_is_chw else (H,W,3).  Writes *loss (deterministic), v_out (H,W,3) =
 * d loss / d out_img (zero where the clamp is active), and optionally the clamped image (H,W,3; NULL to skip).
 * ws: gi2d_raster_mse_workspace_bytes() bytes whose first word is ZERO before the first call (the kernel resets it). */
size_t gi2d_raster_mse_workspace_bytes(void);
int gi2d_raster_sum_fwd_packed_mse(int TX, int TY, int W, int H, int *tile_bins, const float *packed, const int *d_status,
                                   const float *background, const float *gt, int gt_is_chw, int clamp01, float *v_out,
                                   float *image, float *loss, void *ws, size_t ws_bytes, cudaStream_t stream);
/* gi2d_raster_sum_bwd_packed with v_out multiplied by the DEVICE scalar *v_scale (the upstream gradient of the loss). */
int gi2d_raster_sum_bwd_packed_scaled(int N, int TX, int TY, int W, int H, const int *tile_bins, const float *packed,
                                      const float *v_out, const float *v_scale, float *grads12, cudaStream_t stream);

/* ---- optimizer step -------------------------------------------------------------------------- */

/* The Adan update of optimizer.py:232-279 (`_single_tensor_adan`; optimizer.py:296-362 is the same arithmetic as ~25
 * torch._foreach launches) over all tensors of a parameter group in ONE launch (one per 16 tensors).  This is what the
 * reference's own plug-in seam expects from the external module `fused_adan` (optimizer.py:365-449: adan_multi_tensor /
 * adan_single_tensor; the CUDA extension of github.com/sail-sg/Adan, neither vendored nor pinned by the reference).
 * numel / params / ... are HOST arrays of n_tensors entries (element counts; fp32 DEVICE pointers), all updated in place.
 * The scalars are the ones the reference computes in Python doubles (optimizer.py:169-171, 210-214).  grads are scaled
 * by clip_global_grad_norm in place when it is not 1 (as optimizer.py:240 does).  update_neg_pre_grad != 0 also stores
 * neg_pre_grad = -grad (optimizer.py:268, :394-395), which the reference's fused wrapper does with two more launches. */
int gi2d_adan_multi_tensor(int n_tensors, const long long *numel, void *const *params, void *const *grads,
                           void *const *exp_avgs, void *const *exp_avg_sqs, void *const *exp_avg_diffs,
                           void *const *neg_pre_grads, double beta1, double beta2, double beta3, double bias_correction1,
                           double bias_correction2, double bias_correction3_sqrt, double lr, double weight_decay, double eps,
                           int no_prox, double clip_global_grad_norm, int update_neg_pre_grad, cudaStream_t stream);

/* CUDA-graph-capturable form: hyper-parameters and the step counter live in a DEVICE block of gi2d_adan_state_bytes()
 * bytes (fill a host copy with gi2d_adan_state_init and upload it; re-upload when lr changes).  The kernel derives the
 * bias corrections of step+1 in double and the last block to finish advances the counter, so replaying the same graph
 * performs successive optimizer steps.  At most 16 tensors per call. */
size_t gi2d_adan_state_bytes(void);
int gi2d_adan_state_init(void *state_host, double beta1, double beta2, double beta3, double lr, double weight_decay, double eps,
                         double clip_global_grad_norm, int step);
int gi2d_adan_multi_tensor_graph(int n_tensors, const long long *numel, void *const *params, void *const *grads,
                                 void *const *exp_avgs, void *const *exp_avg_sqs, void *const *exp_avg_diffs,
                                 void *const *neg_pre_grads, void *state_dev, int no_prox, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GI2D_H */
