/*
 * gi2d_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, fp32, IEEE round-to-nearest) of the reference's gsplat 2D render
 * path: projection -> tile binning/sort -> accumulated-sum rasterization, forward and backward.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's shared library.  The product (gaussianimage_b200/) never links or calls it.
 *
 * Every function cites the reference lines it follows (paths relative to
 * /root/reference/gsplat/gsplat/cuda/csrc/ unless stated otherwise).
 *
 * Floating-point contraction is PINNED in source: this file must be compiled with
 * -ffp-contract=off; every fused multiply-add below is an explicit fmaf() and mirrors the FFMA
 * placement that `nvcc 12.9 -O3 -arch=sm_100a` (the reference's JIT build, cuda/_backend.py:39-40)
 * produces for the reference kernels (read off `cuobjdump -sass`).
 * rcp.rn / sqrt.rn in that build are correctly rounded, so 1.0f/x and sqrtf() reproduce them
 * bit for bit; the only operations a CPU cannot reproduce bit-wise are MUFU.EX2 (`__expf`) and
 * libdevice sinf/cosf -- handled with the `ambig` outputs below.
 *
 * Parity pinning status: PINNED.  compute_cov2d_bounds / map_gaussian_to_intersects / get_tile_bin_edges
 * against the reference's own _torch_impl.py (tests/golden/ref_torch_impl.npz, made by
 * tests/golden/make_golden.py in the build container).  Projection (both parametrisations),
 * rasterize-sum forward/backward and projection backward -- which the reference neither tests nor
 * implements on the CPU -- against OUTPUTS OF THE REFERENCE'S OWN CUDA KERNELS, compiled in place from
 * /root/reference (oracle/_ref) and run on a B200: tests/golden/ref_cuda_kernels.npz, made by
 * tests/golden/make_golden_refcuda.py, checked by tests/test_oracle_vs_refcuda_golden_cpu.py
 * (integers, tile lists, bin edges and Cholesky projection floats bit-exact; image 1e-5; gradients
 * 1e-4).  See DESIGN.md section 6.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GI2D_BLOCK 16      /* config.h:1-2 BLOCK_X/BLOCK_Y (hard-coded in helpers.cuh:36-41, quirk Q3) */
#define GI2D_BATCH 256     /* config.h:3 BLOCK_SIZE: only the first 256 sorted intersections count (Q1) */

/* (int) cast of the CUDA build = cvt.rzi.s32.f32: truncates, saturates, NaN -> 0. */
static inline int32_t f2i_rz(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int32_t)v;
}
/* F2I.CEIL of the CUDA build ((int)ceil(x) fused): saturating, NaN -> 0. */
static inline int32_t f2i_ceil(float v) { return f2i_rz(ceilf(v)); }
/* CUDA fmaxf/fminf: if one operand is NaN the other is returned. */
static inline float cu_fmaxf(float a, float b) { return fmaxf(a, b); }
static inline float cu_fminf(float a, float b) { return fminf(a, b); }
static inline int32_t imin(int32_t a, int32_t b) { return a < b ? a : b; }
static inline int32_t imax(int32_t a, int32_t b) { return a > b ? a : b; }

/* helpers.cuh:45-68 compute_cov2d_bounds.  SASS contraction of the JIT build:
 * det = fma(a, c, -(b*b)); disc = fma(bh, bh, -det). Returns 0 when det == 0. */
static inline int cov2d_bounds(float a, float b, float c, float conic[3], float *radius_f) {
    float bb = b * b;
    float det = fmaf(a, c, -bb);
    if (det == 0.f) return 0;
    float inv = 1.0f / det;                    /* rcp.rn */
    conic[0] = c * inv;
    conic[1] = b * (-inv);
    conic[2] = a * inv;
    float bh = (a + c) * 0.5f;
    float disc = cu_fmaxf(fmaf(bh, bh, -det), 0.1f);
    float sq = sqrtf(disc);                    /* sqrt.rn */
    float v1 = bh + sq, v2 = bh - sq;
    *radius_f = 3.f * sqrtf(cu_fmaxf(v1, v2)); /* caller applies ceil */
    return 1;
}

/* helpers.cuh:11-43 get_bbox + get_tile_bbox with /16 folded to *0.0625 and fused (SASS):
 * tmin = clamp((int)fma(c, 1/16, -(r/16)), 0, T); tmax = clamp((int)(fma(c, 1/16, r/16) + 1), 0, T). */
static inline void tile_bbox(float cx, float cy, float radius, int TX, int TY, int32_t tmin[2], int32_t tmax[2]) {
    float r16 = radius * 0.0625f;
    tmin[0] = imin(imax(0, f2i_rz(fmaf(cx, 0.0625f, -r16))), TX);
    tmax[0] = imin(imax(0, f2i_rz(fmaf(cx, 0.0625f, r16) + 1.0f)), TX);
    tmin[1] = imin(imax(0, f2i_rz(fmaf(cy, 0.0625f, -r16))), TY);
    tmax[1] = imin(imax(0, f2i_rz(fmaf(cy, 0.0625f, r16) + 1.0f)), TY);
}

/* Shared tail of both projection kernels: foward2d.cu:50-67 / :105-122. */
static inline void project_tail(int i, float cx, float cy, float a, float b, float c, int TX, int TY,
                                float *xys, float *depths, int32_t *radii, float *conics, int32_t *num_tiles_hit) {
    float conic[3], rf;
    radii[i] = 0;
    num_tiles_hit[i] = 0;
    if (!cov2d_bounds(a, b, c, conic, &rf)) return;    /* outputs stay zero (bindings.cu:808-817, Q11) */
    float radius = ceilf(rf);
    conics[3 * i + 0] = conic[0]; conics[3 * i + 1] = conic[1]; conics[3 * i + 2] = conic[2];
    xys[2 * i + 0] = cx; xys[2 * i + 1] = cy;
    radii[i] = f2i_ceil(rf);                             /* written before the area test (Q11) */
    int32_t tmin[2], tmax[2];
    tile_bbox(cx, cy, radius, TX, TY, tmin, tmax);
    int32_t area = (int32_t)((uint32_t)(tmax[0] - tmin[0]) * (uint32_t)(tmax[1] - tmin[1]));
    if (area <= 0) return;
    num_tiles_hit[i] = area;
    depths[i] = 0.f;
}

/* foward2d.cu:12-69 project_gaussians_2d_forward_kernel (Cholesky). All outputs must be
 * zero-initialised by the caller like bindings.cu:808-817 does. */
void gi2d_oracle_project_cholesky_fwd(int N, const float *means2d, const float *L, int H, int W, int TX, int TY,
                                      float *xys, float *depths, int32_t *radii, float *conics, int32_t *num_tiles_hit) {
    float hw = (float)(uint32_t)W * 0.5f, hh = (float)(uint32_t)H * 0.5f;
    for (int i = 0; i < N; ++i) {
        float cx = fmaf(hw, means2d[2 * i + 0], hw);    /* :37-38, fused by nvcc */
        float cy = fmaf(means2d[2 * i + 1], hh, hh);
        float l11 = L[3 * i + 0], l21 = L[3 * i + 1], l22 = L[3 * i + 2];
        float a = l11 * l11;                             /* :47 */
        float b = l11 * l21;
        float c = fmaf(l21, l21, l22 * l22);
        project_tail(i, cx, cy, a, b, c, TX, TY, xys, depths, radii, conics, num_tiles_hit);
    }
}

/* foward2d.cu:71-123 project_gaussians_2d_scale_rot_forward_kernel.
 * R = rotmat2d(theta) (helpers.cuh:274-285; glm is column-major, R[col][row]: R[0]=(cos,-sin),
 * R[1]=(sin,cos)), S = diag(sx, sy), M = R*S, cov = M*transpose(M).  glm's mat2*mat2
 * (third_party/glm/glm/detail/type_mat2x2.inl) evaluates out[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1].
 * SASS of the JIT build keeps the multiplications by the literal zeros of S and contracts each sum
 * into one FFMA, e.g. M[0][0] = fma(cos, sx, 0*sin); for finite inputs that is round(cos*sx)
 * exactly, so the four entries are p = cos*sx, u = -(sin*sx) (column 0) and q = sin*sy, v = cos*sy
 * (column 1).  The products of M*M^T are contracted as
 *   cov.x = tmp[0][0] = fma(p, p, q*q)
 *   cov.y = tmp[0][1] = fma(p, u, q*v)
 *   cov.z = tmp[1][1] = fma(u, u, v*v)
 * sinf/cosf are libdevice on the GPU and libm here: not bit-reproducible (1-2 ulp), so tile lists
 * of this variant are compared exactly GPU-vs-GPU (oracle/_ref), and against this file exactly
 * OUTSIDE the set of Gaussians whose radius, evaluated in float64 (gi2d_oracle.py:
 * scale_rot_radius_shadow), sits within 2e-5 of a ceil step -- the only place where 1-2 ulp of
 * sin/cos can change an integer output. */
void gi2d_oracle_project_scale_rot_fwd(int N, const float *means2d, const float *scales, const float *rot, int H, int W,
                                       int TX, int TY, float *xys, float *depths, int32_t *radii, float *conics,
                                       int32_t *num_tiles_hit) {
    float hw = (float)(uint32_t)W * 0.5f, hh = (float)(uint32_t)H * 0.5f;
    for (int i = 0; i < N; ++i) {
        float cx = fmaf(hw, means2d[2 * i + 0], hw);
        float cy = fmaf(means2d[2 * i + 1], hh, hh);
        float cs = cosf(rot[i]), sn = sinf(rot[i]);
        float sx = scales[2 * i + 0], sy = scales[2 * i + 1];
        float p = cs * sx, u = -(sn * sx), q = sn * sy, v = cs * sy;
        float a = fmaf(p, p, q * q);
        float b = fmaf(p, u, q * v);
        float c = fmaf(u, u, v * v);
        project_tail(i, cx, cy, a, b, c, TX, TY, xys, depths, radii, conics, num_tiles_hit);
    }
}

/* bindings.cu:21-61 compute_cov2d_bounds_kernel / utils.py:77-96: conic (N,3), radius (N,1) float. */
void gi2d_oracle_cov2d_bounds(int N, const float *cov2d, float *conics, float *radii) {
    for (int i = 0; i < N; ++i) {
        float conic[3], rf;
        if (!cov2d_bounds(cov2d[3 * i], cov2d[3 * i + 1], cov2d[3 * i + 2], conic, &rf)) continue;
        conics[3 * i] = conic[0]; conics[3 * i + 1] = conic[1]; conics[3 * i + 2] = conic[2];
        radii[i] = ceilf(rf);
    }
}

/* utils.py:116-118 compute_cumulative_intersects: inclusive int32 scan, returns the total. */
int32_t gi2d_oracle_cumsum(int N, const int32_t *num_tiles_hit, int32_t *cum) {
    int32_t acc = 0;
    for (int i = 0; i < N; ++i) { acc += num_tiles_hit[i]; cum[i] = acc; }
    return acc;
}

/* forward.cu:100-136 map_gaussian_to_intersects: key = (tile << 32) | depth_bits, value = gaussian id. */
void gi2d_oracle_map_intersects(int N, const float *xys, const float *depths, const int32_t *radii, const int32_t *cum,
                                int TX, int TY, int64_t *isect_ids, int32_t *gaussian_ids) {
    for (int i = 0; i < N; ++i) {
        if (radii[i] <= 0) continue;
        int32_t tmin[2], tmax[2];
        tile_bbox(xys[2 * i], xys[2 * i + 1], (float)radii[i], TX, TY, tmin, tmax);
        int32_t cur = (i == 0) ? 0 : cum[i - 1];
        int32_t dbits; memcpy(&dbits, &depths[i], 4);
        int64_t depth_id = (int64_t)dbits;               /* sign-extends, as `(int64_t)*(int32_t*)&depth` does */
        for (int32_t ty = tmin[1]; ty < tmax[1]; ++ty)
            for (int32_t tx = tmin[0]; tx < tmax[0]; ++tx) {
                int64_t tile = (int64_t)(ty * TX + tx);
                isect_ids[cur] = (tile << 32) | depth_id;
                gaussian_ids[cur] = i;
                ++cur;
            }
    }
}

/* utils.py:164-165 torch.sort(int64 keys) + gather.  Signed ascending order, ties keep emission
 * order (stable; the CUDA radix sort torch dispatches to is stable -- quirk Q2). LSD radix, 8x8 bit. */
void gi2d_oracle_sort_intersects(int n, const int64_t *ids, const int32_t *gids, int64_t *ids_sorted,
                                 int32_t *gids_sorted) {
    if (n <= 0) return;
    uint64_t *ka = malloc(sizeof(uint64_t) * n), *kb = malloc(sizeof(uint64_t) * n);
    int32_t *va = malloc(sizeof(int32_t) * n), *vb = malloc(sizeof(int32_t) * n);
    for (int i = 0; i < n; ++i) { ka[i] = (uint64_t)ids[i] ^ 0x8000000000000000ull; va[i] = gids[i]; }
    for (int pass = 0; pass < 8; ++pass) {
        size_t count[257]; memset(count, 0, sizeof count);
        int shift = pass * 8;
        for (int i = 0; i < n; ++i) count[((ka[i] >> shift) & 0xff) + 1]++;
        for (int d = 0; d < 256; ++d) count[d + 1] += count[d];
        for (int i = 0; i < n; ++i) { size_t p = count[(ka[i] >> shift) & 0xff]++; kb[p] = ka[i]; vb[p] = va[i]; }
        uint64_t *tk = ka; ka = kb; kb = tk; int32_t *tv = va; va = vb; vb = tv;
    }
    for (int i = 0; i < n; ++i) { ids_sorted[i] = (int64_t)(ka[i] ^ 0x8000000000000000ull); gids_sorted[i] = va[i]; }
    free(ka); free(kb); free(va); free(vb);
}

/* forward.cu:141-163 get_tile_bin_edges. tile_bins has n_bins rows (zero-initialised by the caller);
 * the reference allocates num_intersects rows and indexes by tile id (Q4) -- rows >= n_bins are
 * skipped here instead of written out of bounds. */
void gi2d_oracle_tile_bin_edges(int n, const int64_t *ids_sorted, int n_bins, int32_t *tile_bins) {
    for (int i = 0; i < n; ++i) {
        int32_t cur = (int32_t)(ids_sorted[i] >> 32);
        if (i == 0 && cur >= 0 && cur < n_bins) tile_bins[2 * cur] = 0;
        if (i == n - 1 && cur >= 0 && cur < n_bins) tile_bins[2 * cur + 1] = n;
        if (i == 0) continue;
        int32_t prev = (int32_t)(ids_sorted[i - 1] >> 32);
        if (prev != cur) {
            if (prev >= 0 && prev < n_bins) tile_bins[2 * prev + 1] = i;
            if (cur >= 0 && cur < n_bins) tile_bins[2 * cur] = i;
        }
    }
}

/* One (pixel, gaussian) evaluation, forward.cu:595-603 / backward.cu:797-807.  FFMA placement of
 * the JIT build: q = fma(dx, cx*dx, dy*(cz*dy)); sigma = fma(dy, cy*dx, 0.5*q).
 * `__expf(-sigma)` = ex2.approx(sigma * -log2e) (one FMUL by 0xBFB8AA3B, then MUFU.EX2). */
static inline float eval_sigma(float x, float y, float cx, float cy, float cz, float px, float py, float *dx_o, float *dy_o) {
    float dx = x - px, dy = y - py;
    float q = fmaf(dx, cx * dx, dy * (cz * dy));
    float h = q * 0.5f;
    *dx_o = dx; *dy_o = dy;
    return fmaf(dy, cy * dx, h);
}
static inline float cu_expf_neg(float sigma) { return exp2f(sigma * -1.4426950216293334961f); }

#define GI2D_ALPHA_MIN (1.f / 255.f)
#define GI2D_AMBIG_REL 4e-6f   /* MUFU.EX2 is accurate to ~2 ulp; pairs this close to 1/255 may flip */

/* forward.cu:512-627 rasterize_forward_sum (Q1, Q6).  out_img (H,W,3), final_Ts (H,W) == 1,
 * final_idx (H,W) int32.  `ambig` (H,W) uint8, may be NULL: set to 1 for a pixel when some pair's
 * alpha lies within GI2D_AMBIG_REL of the 1/255 threshold (a GPU ex2.approx may decide it differently). */
void gi2d_oracle_raster_sum_fwd(int TX, int TY, int W, int H, const int32_t *gids_sorted, const int32_t *tile_bins,
                                const float *xys, const float *conics, const float *colors, const float *opac,
                                float *out_img, float *final_Ts, int32_t *final_idx, uint8_t *ambig) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < TX * TY; ++t) {
        int ty = t / TX, tx = t % TX;
        int lo = tile_bins[2 * t], hi = tile_bins[2 * t + 1];
        int end = hi < lo + GI2D_BATCH ? hi : lo + GI2D_BATCH;      /* Q1: `done = true` at forward.cu:613 */
        for (int i = ty * GI2D_BLOCK; i < (ty + 1) * GI2D_BLOCK && i < H; ++i)
            for (int j = tx * GI2D_BLOCK; j < (tx + 1) * GI2D_BLOCK && j < W; ++j) {
                float px = (float)j, py = (float)i;
                float r = 0.f, g = 0.f, b = 0.f;
                int32_t cur = 0; uint8_t amb = 0;
                for (int idx = lo; idx < end; ++idx) {
                    int gid = gids_sorted[idx];
                    float dx, dy;
                    float sigma = eval_sigma(xys[2 * gid], xys[2 * gid + 1], conics[3 * gid], conics[3 * gid + 1],
                                             conics[3 * gid + 2], px, py, &dx, &dy);
                    float raw = opac[gid] * cu_expf_neg(sigma);
                    float alpha = cu_fminf(raw, 1.f);
                    if (fabsf(alpha - GI2D_ALPHA_MIN) <= GI2D_AMBIG_REL * GI2D_ALPHA_MIN && !(sigma < 0.f)) amb = 1;
                    if (sigma < 0.f || alpha < GI2D_ALPHA_MIN) continue;
                    r = fmaf(alpha, colors[3 * gid + 0], r);          /* forward.cu:607-609, contracted */
                    g = fmaf(alpha, colors[3 * gid + 1], g);
                    b = fmaf(alpha, colors[3 * gid + 2], b);
                    cur = idx;
                }
                size_t pix = (size_t)i * W + j;
                out_img[3 * pix] = r; out_img[3 * pix + 1] = g; out_img[3 * pix + 2] = b;
                final_Ts[pix] = 1.f;
                final_idx[pix] = cur;
                if (ambig) ambig[pix] = amb;
            }
    }
}

/* backward.cu:696-862 rasterize_backward_sum_kernel (Q7).  The kernel visits sorted indices
 * idx <= final_idx[pix] that pass the same validity test; because final_idx is the last valid index
 * among the first 256 of the tile, that is exactly the forward's contributing set.  Per-pair terms
 * are fp32 as in the kernel; the sums over pixels (warp reduce + atomicAdd in the reference, order
 * nondeterministic) are accumulated in double here.  Outputs (zero-initialised by the caller):
 * v_xy (N,2), v_conic (N,3), v_colors (N,3), v_opacity (N,1). */
void gi2d_oracle_raster_sum_bwd(int TX, int TY, int W, int H, const int32_t *gids_sorted, const int32_t *tile_bins,
                                const float *xys, const float *conics, const float *colors, const float *opac,
                                const int32_t *final_idx, const float *v_out, int N, float *v_xy, float *v_conic,
                                float *v_colors, float *v_opacity) {
    int nthr = 1;
#ifdef _OPENMP
    nthr = omp_get_max_threads();
    if (nthr > 16) nthr = 16;                 /* bounds the private accumulators: nthr * N * 72 bytes */
#endif
    double *acc_all = calloc((size_t)nthr * N * 9, sizeof(double));
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthr)
    for (int t = 0; t < TX * TY; ++t) {
        int tid = 0;
#ifdef _OPENMP
        tid = omp_get_thread_num();
#endif
        double *acc = acc_all + (size_t)tid * N * 9;
        int ty = t / TX, tx = t % TX;
        int lo = tile_bins[2 * t], hi = tile_bins[2 * t + 1];
        for (int i = ty * GI2D_BLOCK; i < (ty + 1) * GI2D_BLOCK && i < H; ++i)
            for (int j = tx * GI2D_BLOCK; j < (tx + 1) * GI2D_BLOCK && j < W; ++j) {
                size_t pix = (size_t)i * W + j;
                float px = (float)j, py = (float)i;
                float vo0 = v_out[3 * pix], vo1 = v_out[3 * pix + 1], vo2 = v_out[3 * pix + 2];
                int32_t bin_final = final_idx[pix];
                for (int idx = lo; idx < hi && idx <= bin_final; ++idx) {       /* backward.cu:783-787 */
                    int gid = gids_sorted[idx];
                    float cx = conics[3 * gid], cy = conics[3 * gid + 1], cz = conics[3 * gid + 2];
                    float dx, dy;
                    float sigma = eval_sigma(xys[2 * gid], xys[2 * gid + 1], cx, cy, cz, px, py, &dx, &dy);
                    float vis = cu_expf_neg(sigma);
                    float op = opac[gid];
                    float alpha = cu_fminf(op * vis, 1.f);
                    if (sigma < 0.f || alpha < GI2D_ALPHA_MIN) continue;
                    double *a = acc + (size_t)gid * 9;
                    a[0] += alpha * vo0; a[1] += alpha * vo1; a[2] += alpha * vo2;          /* :814 */
                    float v_alpha = colors[3 * gid] * vo0 + colors[3 * gid + 1] * vo1 + colors[3 * gid + 2] * vo2;
                    float v_sigma = -op * vis * v_alpha;                                     /* :822 */
                    a[3] += 0.5f * v_sigma * dx * dx;                                        /* :823-825 */
                    a[4] += 0.5f * v_sigma * dx * dy;
                    a[5] += 0.5f * v_sigma * dy * dy;
                    a[6] += v_sigma * (cx * dx + cy * dy);                                   /* :826-827 */
                    a[7] += v_sigma * (cy * dx + cz * dy);
                    a[8] += vis * v_alpha;                                                   /* :828 */
                }
            }
    }
    for (int g = 0; g < N; ++g) {
        double a[9] = {0};
        for (int k = 0; k < nthr; ++k) for (int c = 0; c < 9; ++c) a[c] += acc_all[((size_t)k * N + g) * 9 + c];
        v_colors[3 * g] = (float)a[0]; v_colors[3 * g + 1] = (float)a[1]; v_colors[3 * g + 2] = (float)a[2];
        v_conic[3 * g] = (float)a[3]; v_conic[3 * g + 1] = (float)a[4]; v_conic[3 * g + 2] = (float)a[5];
        v_xy[2 * g] = (float)a[6]; v_xy[2 * g + 1] = (float)a[7];
        v_opacity[g] = (float)a[8];
    }
    free(acc_all);
}

/* helpers.cuh:71-82 cov2d_to_conic_vjp: V = -X G X with X = conic, G = v_conic (both symmetric),
 * v_cov2d = (V00, V10 + V01, V11). */
static inline void conic_vjp(const float *conic, const float *v_conic, float out[3]) {
    float x0 = conic[0], x1 = conic[1], x2 = conic[2];
    float g0 = v_conic[0], g1 = v_conic[1], g2 = v_conic[2];
    /* A = X G */
    float a00 = x0 * g0 + x1 * g1, a01 = x0 * g1 + x1 * g2;
    float a10 = x1 * g0 + x2 * g1, a11 = x1 * g1 + x2 * g2;
    /* V = -(A X) */
    float v00 = -(a00 * x0 + a01 * x1), v01 = -(a00 * x1 + a01 * x2);
    float v10 = -(a10 * x0 + a11 * x1), v11 = -(a10 * x1 + a11 * x2);
    out[0] = v00; out[1] = v10 + v01; out[2] = v11;
}

/* backward2d.cu:8-51 project_gaussians_2d_backward_kernel (Q5: the off-diagonal term is doubled).
 * Outputs zero-initialised by the caller; rows with radii <= 0 stay zero. v_cov2d may be NULL. */
void gi2d_oracle_project_cholesky_bwd(int N, const float *L, int H, int W, const int32_t *radii, const float *conics,
                                      const float *v_xy, const float *v_conic, float *v_cov2d, float *v_mean2d,
                                      float *v_L) {
    float hw = 0.5f * (float)(uint32_t)W, hh = 0.5f * (float)(uint32_t)H;
    for (int i = 0; i < N; ++i) {
        if (radii[i] <= 0) continue;
        float G[3];
        conic_vjp(conics + 3 * i, v_conic + 3 * i, G);
        if (v_cov2d) { v_cov2d[3 * i] = G[0]; v_cov2d[3 * i + 1] = G[1]; v_cov2d[3 * i + 2] = G[2]; }
        float l11 = L[3 * i], l21 = L[3 * i + 1], l22 = L[3 * i + 2];
        v_L[3 * i + 0] = 2 * l11 * G[0] + 2 * G[1] * l21;       /* :39-41 */
        v_L[3 * i + 1] = 2 * l11 * G[1] + 2 * l21 * G[2];
        v_L[3 * i + 2] = 2 * l22 * G[2];
        v_mean2d[2 * i + 0] = v_xy[2 * i + 0] * hw;             /* :48-49 */
        v_mean2d[2 * i + 1] = v_xy[2 * i + 1] * hh;
    }
}

/* backward2d.cu:53-109 project_gaussians_2d_scale_rot_backward_kernel (Q5).
 * glm is column-major: mat2(a,b,c,d) has columns (a,b),(c,d); m[c][r].  Written out by hand:
 * R = [[cs, sn],[-sn, cs]] as (row,col) i.e. R[0]=(cs,-sn) is column 0.
 * We keep everything as explicit (row,col) 2x2 arrays m[r][c] with the glm semantics
 * glm_m[c][r] == m[r][c]. */
typedef struct { float m[2][2]; } m2;   /* m[row][col] */
static inline m2 m2mul(m2 a, m2 b) {
    m2 o;
    for (int r = 0; r < 2; ++r) for (int c = 0; c < 2; ++c) o.m[r][c] = a.m[r][0] * b.m[0][c] + a.m[r][1] * b.m[1][c];
    return o;
}
static inline m2 m2T(m2 a) { m2 o = {{{a.m[0][0], a.m[1][0]}, {a.m[0][1], a.m[1][1]}}}; return o; }
static inline m2 m2add(m2 a, m2 b) {
    m2 o; for (int r = 0; r < 2; ++r) for (int c = 0; c < 2; ++c) o.m[r][c] = a.m[r][c] + b.m[r][c]; return o;
}
void gi2d_oracle_project_scale_rot_bwd(int N, const float *scales, const float *rot, int H, int W, const int32_t *radii,
                                       const float *conics, const float *v_xy, const float *v_conic, float *v_cov2d,
                                       float *v_mean2d, float *v_scale, float *v_rot) {
    float hw = 0.5f * (float)(uint32_t)W, hh = 0.5f * (float)(uint32_t)H;
    for (int i = 0; i < N; ++i) {
        if (radii[i] <= 0) continue;
        float G[3];
        conic_vjp(conics + 3 * i, v_conic + 3 * i, G);
        if (v_cov2d) { v_cov2d[3 * i] = G[0]; v_cov2d[3 * i + 1] = G[1]; v_cov2d[3 * i + 2] = G[2]; }
        float cs = cosf(rot[i]), sn = sinf(rot[i]);
        float sx = scales[2 * i], sy = scales[2 * i + 1];
        /* helpers.cuh:274-298: glm R[0][0]=cs,R[0][1]=-sn,R[1][0]=sn,R[1][1]=cs => (row,col): */
        m2 R = {{{cs, sn}, {-sn, cs}}};
        m2 Rg = {{{-sn, cs}, {-cs, -sn}}};
        m2 S = {{{sx, 0.f}, {0.f, sy}}};
        m2 M = m2mul(R, S);
        m2 theta_g = m2add(m2mul(m2mul(Rg, S), m2T(M)), m2mul(m2mul(M, m2T(S)), m2T(Rg)));   /* :88 */
        m2 Sxg = {{{2.f * sx, 0.f}, {0.f, 0.f}}}, Syg = {{{0.f, 0.f}, {0.f, 2.f * sy}}};
        m2 sig_x = m2mul(m2mul(R, Sxg), m2T(R)), sig_y = m2mul(m2mul(R, Syg), m2T(R));        /* :95-96 */
        /* glm X[0][1] is (col 0,row 1) == m[1][0] here */
        v_scale[2 * i + 0] = G[0] * sig_x.m[0][0] + 2 * G[1] * sig_x.m[1][0] + G[2] * sig_x.m[1][1];   /* :102-104 */
        v_scale[2 * i + 1] = G[0] * sig_y.m[0][0] + 2 * G[1] * sig_y.m[1][0] + G[2] * sig_y.m[1][1];
        v_rot[i] = G[0] * theta_g.m[0][0] + 2 * G[1] * theta_g.m[1][0] + G[2] * theta_g.m[1][1];
        v_mean2d[2 * i + 0] = v_xy[2 * i + 0] * hw;
        v_mean2d[2 * i + 1] = v_xy[2 * i + 1] * hh;
    }
}

/* Number of (pixel, gaussian) evaluations E of one forward pass: sum over tiles of
 * in-image pixels x min(count, 256).  Used by bench.py to turn time into FLOP/s. */
int64_t gi2d_oracle_count_evals(int TX, int TY, int W, int H, const int32_t *tile_bins) {
    int64_t e = 0;
    for (int t = 0; t < TX * TY; ++t) {
        int ty = t / TX, tx = t % TX;
        int cnt = tile_bins[2 * t + 1] - tile_bins[2 * t];
        if (cnt > GI2D_BATCH) cnt = GI2D_BATCH;
        if (cnt < 0) cnt = 0;
        int ph = imin(GI2D_BLOCK, H - ty * GI2D_BLOCK), pw = imin(GI2D_BLOCK, W - tx * GI2D_BLOCK);
        e += (int64_t)ph * pw * cnt;
    }
    return e;
}
