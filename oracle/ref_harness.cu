/*
 * ref_harness.cu -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Thin extern "C" launchers around the REFERENCE's own CUDA kernels, so that the GPU parity tests and
 * the reference-CUDA timing script can run the reference's arithmetic on the GPU box without torch's
 * extension builder.  The kernels themselves are NOT in this repository: oracle/Makefile compiles
 * foward2d.cu, backward2d.cu, forward.cu, backward.cu straight from
 * /root/reference/gsplat/gsplat/cuda/csrc/ (where they lie) and links them with this file into
 * oracle/_ref/libgsplat_ref_{jit,fast}.so (git-ignored, shipped to the GPU box by gpurun).
 *
 * Launch shapes and argument casts restate the reference's host wrappers:
 *   bindings.cu:781-839  project_gaussians_2d_forward_tensor      (grid ceil(N/256) x 256)
 *   bindings.cu:841-899  project_gaussians_2d_scale_rot_forward_tensor
 *   bindings.cu:902-949  project_gaussians_2d_backward_tensor
 *   bindings.cu:951-1004 project_gaussians_2d_scale_rot_backward_tensor
 *   bindings.cu:274-313  map_gaussian_to_intersects_tensor
 *   bindings.cu:315-330  get_tile_bin_edges_tensor
 *   bindings.cu:400-469  rasterize_forward_sum_tensor               (grid tiles, block 16x16)
 *   bindings.cu:706-779  rasterize_backward_sum_tensor
 * All kernels go to the legacy default stream like the reference (quirk Q9); callers synchronise.
 * Outputs must be zero-filled by the caller (the reference uses torch::zeros).  `background_host` is a HOST pointer to
 * 3 floats (the kernels take the colour by value; they never use it, quirk Q6).
 */
#include "forward.cuh"
#include "backward.cuh"
#include "forward2d.cuh"
#include "backward2d.cuh"
#include "config.h"

#define REF_API extern "C" __attribute__((visibility("default")))

static inline int last_err() { return (int)cudaGetLastError(); }

REF_API int ref_project_cholesky_fwd(int N, const float *means2d, const float *L, int H, int W, int TX, int TY,
                                     float clip, float *xys, float *depths, int *radii, float *conics,
                                     int *num_tiles_hit) {
    dim3 img(W, H, 1), tb(TX, TY, 1);
    project_gaussians_2d_forward_kernel<<<(N + N_THREADS - 1) / N_THREADS, N_THREADS>>>(
        N, (const float2 *)means2d, (const float3 *)L, img, tb, clip, (float2 *)xys, depths, radii, (float3 *)conics,
        num_tiles_hit);
    return last_err();
}

REF_API int ref_project_scale_rot_fwd(int N, const float *means2d, const float *scales, const float *rot, int H, int W,
                                      int TX, int TY, float clip, float *xys, float *depths, int *radii, float *conics,
                                      int *num_tiles_hit) {
    dim3 img(W, H, 1), tb(TX, TY, 1);
    project_gaussians_2d_scale_rot_forward_kernel<<<(N + N_THREADS - 1) / N_THREADS, N_THREADS>>>(
        N, (const float2 *)means2d, (const float2 *)scales, rot, img, tb, clip, (float2 *)xys, depths, radii,
        (float3 *)conics, num_tiles_hit);
    return last_err();
}

REF_API int ref_project_cholesky_bwd(int N, const float *means2d, const float *L, int H, int W, const int *radii,
                                     const float *conics, const float *v_xy, const float *v_depth,
                                     const float *v_conic, float *v_cov2d, float *v_mean2d, float *v_L) {
    dim3 img(W, H, 1);
    project_gaussians_2d_backward_kernel<<<(N + N_THREADS - 1) / N_THREADS, N_THREADS>>>(
        N, (const float2 *)means2d, (const float3 *)L, img, radii, (const float3 *)conics, (const float2 *)v_xy,
        v_depth, (const float3 *)v_conic, (float3 *)v_cov2d, (float2 *)v_mean2d, (float3 *)v_L);
    return last_err();
}

REF_API int ref_project_scale_rot_bwd(int N, const float *means2d, const float *scales, const float *rot, int H, int W,
                                      const int *radii, const float *conics, const float *v_xy, const float *v_depth,
                                      const float *v_conic, float *v_cov2d, float *v_mean2d, float *v_scale,
                                      float *v_rot) {
    dim3 img(W, H, 1);
    project_gaussians_2d_scale_rot_backward_kernel<<<(N + N_THREADS - 1) / N_THREADS, N_THREADS>>>(
        N, (const float2 *)means2d, (const float2 *)scales, rot, img, radii, (const float3 *)conics,
        (const float2 *)v_xy, v_depth, (const float3 *)v_conic, (float3 *)v_cov2d, (float2 *)v_mean2d,
        (float2 *)v_scale, v_rot);
    return last_err();
}

REF_API int ref_map_gaussian_to_intersects(int N, const float *xys, const float *depths, const int *radii,
                                           const int *cum_tiles_hit, int TX, int TY, long long *isect_ids,
                                           int *gaussian_ids) {
    dim3 tb(TX, TY, 1);
    map_gaussian_to_intersects<<<(N + N_THREADS - 1) / N_THREADS, N_THREADS>>>(
        N, (const float2 *)xys, depths, radii, cum_tiles_hit, tb, (int64_t *)isect_ids, gaussian_ids);
    return last_err();
}

REF_API int ref_get_tile_bin_edges(int n, const long long *isect_ids_sorted, int *tile_bins) {
    get_tile_bin_edges<<<(n + N_THREADS - 1) / N_THREADS, N_THREADS>>>(n, (const int64_t *)isect_ids_sorted,
                                                                        (int2 *)tile_bins);
    return last_err();
}

REF_API int ref_rasterize_sum_fwd(int TX, int TY, int W, int H, const int *gids_sorted, const int *tile_bins,
                                  const float *xys, const float *conics, const float *colors, const float *opac,
                                  const float *background_host, float *out_img, float *final_Ts, int *final_idx) {
    dim3 tb(TX, TY, 1), block(BLOCK_X, BLOCK_Y, 1), img(W, H, 1);
    rasterize_forward_sum<<<tb, block>>>(tb, img, gids_sorted, (const int2 *)tile_bins, (const float2 *)xys,
                                         (const float3 *)conics, (const float3 *)colors, opac, final_Ts, final_idx,
                                         (float3 *)out_img, *(const float3 *)background_host);
    return last_err();
}

REF_API int ref_rasterize_sum_bwd(int TX, int TY, int W, int H, const int *gids_sorted, const int *tile_bins,
                                  const float *xys, const float *conics, const float *colors, const float *opac,
                                  const float *background_host, const float *final_Ts, const int *final_idx,
                                  const float *v_out, const float *v_out_alpha, float *v_xy, float *v_conic,
                                  float *v_colors, float *v_opacity) {
    dim3 tb(TX, TY, 1), block(BLOCK_X, BLOCK_Y, 1), img(W, H, 1);
    rasterize_backward_sum_kernel<<<tb, block>>>(
        tb, img, gids_sorted, (const int2 *)tile_bins, (const float2 *)xys, (const float3 *)conics,
        (const float3 *)colors, opac, *(const float3 *)background_host, final_Ts, final_idx, (const float3 *)v_out,
        v_out_alpha, (float2 *)v_xy, (float3 *)v_conic, (float3 *)v_colors, v_opacity);
    return last_err();
}

REF_API int ref_sync(void) { return (int)cudaDeviceSynchronize(); }
