// gi2d_common.cuh -- shared device helpers for the gi2d (GaussianImage 2D render path) kernels, sm_100a.
//
// Arithmetic policy: the tile lists and the alpha>=1/255 decisions are discontinuous functions of the
// floating-point results, so every operation that feeds them is PINNED with round-to-nearest
// intrinsics (__fmul_rn/__fadd_rn/__fmaf_rn are never re-contracted by nvcc or ptxas).  The placement of
// the fused multiply-adds restates what `nvcc 12.9 -O3` (the reference's JIT build,
// gsplat/gsplat/cuda/_backend.py:39-40) generates for the reference kernels on sm_100 -- read off
// `cuobjdump -sass`, see DESIGN.md section 4 -- so integer outputs match that build bit for bit.
// This file must never be compiled with --use_fast_math.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GI2D_TILE 16          // reference config.h:1-2 (BLOCK_X/BLOCK_Y); helpers.cuh:36-41 hard-codes it (Q3)
#define GI2D_TILE_PIX 256
#define GI2D_BATCH 256        // reference config.h:3; only the first 256 sorted intersections of a tile count (Q1)
#define GI2D_REC_FLOATS 12    // packed per-intersection record: 3 x float4 = 48 B

#define GI2D_OK 0
#define GI2D_ERR_INVALID 1
#define GI2D_ERR_WORKSPACE 2
#ifndef GI2D_ERR_STATIC_LIMIT
#define GI2D_ERR_STATIC_LIMIT 3   // size beyond the static scratch of gi2d_scan_tiles_static (include/gi2d.h)
#endif
#define GI2D_ERR_CUDA_BASE 1000   // 1000 + cudaError_t

namespace gi2d {

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// reference helpers.cuh:45-68 compute_cov2d_bounds.  Returns false when det == 0.
// r3 = 3*sqrt(lambda_max) BEFORE the ceil (callers need both ceil-as-float and ceil-as-int).
__device__ __forceinline__ bool cov2d_bounds(float a, float b, float c, float3 &conic, float &r3) {
    const float det = ffma(a, c, -fmul(b, b));
    if (det == 0.f) return false;
    const float inv = __frcp_rn(det);
    conic.x = fmul(c, inv);
    conic.y = fmul(b, -inv);
    conic.z = fmul(a, inv);
    const float bh = fmul(fadd(a, c), 0.5f);
    const float disc = fmaxf(ffma(bh, bh, -det), 0.1f);
    const float sq = __fsqrt_rn(disc);
    const float v1 = fadd(bh, sq), v2 = fsub(bh, sq);
    r3 = fmul(__fsqrt_rn(fmaxf(v1, v2)), 3.f);
    return true;
}

// reference helpers.cuh:11-43 get_bbox/get_tile_bbox: tile-space bbox, min inclusive, max exclusive,
// clamped to [0,TX]x[0,TY]; the division by 16 is the exact multiplication by 0.0625 fused as in the JIT build.
__device__ __forceinline__ void tile_bbox(float cx, float cy, float radius, int TX, int TY, int &x0, int &x1, int &y0,
                                          int &y1) {
    const float r16 = fmul(radius, 0.0625f);
    x0 = min(max(0, __float2int_rz(ffma(cx, 0.0625f, -r16))), TX);
    x1 = min(max(0, __float2int_rz(fadd(ffma(cx, 0.0625f, r16), 1.0f))), TX);
    y0 = min(max(0, __float2int_rz(ffma(cy, 0.0625f, -r16))), TY);
    y1 = min(max(0, __float2int_rz(fadd(ffma(cy, 0.0625f, r16), 1.0f))), TY);
}

// One (pixel, gaussian) evaluation, reference forward.cu:595-600 / backward.cu:797-803:
// sigma = 0.5*(cx*dx*dx + cz*dy*dy) + cy*dx*dy with the JIT build's contraction.
__device__ __forceinline__ float eval_sigma(float dx, float dy, float cx, float cy, float cz) {
    const float q = ffma(dx, fmul(cx, dx), fmul(dy, fmul(cz, dy)));
    return ffma(dy, fmul(cy, dx), fmul(q, 0.5f));
}
// `__expf(-sigma)` of the reference = one FMUL by -log2(e) (0xBFB8AA3B) + MUFU.EX2.
__device__ __forceinline__ float exp_neg(float sigma) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmul(sigma, -1.4426950216293334961f)));
    return r;
}
#define GI2D_ALPHA_MIN (1.f / 255.f)

// ---- TMA 1-D bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier helpers --------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// global -> shared::cta bulk copy of `bytes` (multiple of 16, both addresses 16-B aligned), completion on `bar`.
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Conservative pixel-space extent of the region where alpha = min(1, opac*exp(-sigma)) can reach 1/255
// (the reference's per-pair test, forward.cu:602-604).  Pairs outside it are provably rejected by the
// reference, so skipping them cannot change any output bit.  Returns false when no bound is available
// (degenerate conic): the caller must then treat the whole tile as covered.  ex/ey < 0 means "never visible".
__device__ __forceinline__ bool visible_extent(float cx, float cy, float cz, float opac, float &ex, float &ey) {
    const float D = cx * cz - cy * cy;
    // |rho| close to 1 (or a non positive-definite conic) makes fp32 sigma cancellation-prone: no culling.
    if (!(cx > 0.f) || !(cz > 0.f) || !(D > 2e-3f * cx * cz) || !(opac > 0.f) || !(opac < 3.0e38f)) return false;
    const float smax = __logf(255.f * opac) * 1.002f + 0.01f;   // sigma <= ln(255*opac), padded
    if (!(smax > 0.f)) { ex = ey = -1.f; return true; }
    // approximate reciprocal / square root (MUFU, a few ulp): the bound is padded by 0.1 % + 0.01 px, three orders more
    const float k = __fdividef(2.f * smax, D);
    const float ax = k * cz, ay = k * cx;
    ex = ax * rsqrtf(ax) * 1.001f + 0.01f;
    ey = ay * rsqrtf(ay) * 1.001f + 0.01f;
    return (ex == ex) && (ey == ey);
}

// The extent as it is CARRIED IN A PACKED RECORD (third float4: {b, gid, ex, ey}), computed once where the record is
// packed instead of once per tile and rasterizer pass:  (+inf, +inf) = no bound (whole tile), (-1, -1) = never visible.
__device__ __forceinline__ float2 record_extent(float cx, float cy, float cz, float opac) {
    float ex, ey;
    if (!visible_extent(cx, cy, cz, opac, ex, ey)) return make_float2(__int_as_float(0x7f800000), __int_as_float(0x7f800000));
    return make_float2(ex, ey);
}

// Library-owned pairs of device words {count, generation} for self-cleaning device-side protocols (grid barrier of the
// one-kernel binning, last-block ticket of the projection total): zeroed once per device, left zero by every user.  Every
// launch gets its OWN pair: eager launches cycle through a ring (a pair is long idle again when the ring wraps), launches
// recorded into a CUDA graph -- which replay at any later time, on any stream -- get a pair that is never handed out again.
// nullptr when no pair can be provided (pool not yet created and the stream is capturing, or the capture pairs are spent).
unsigned int *sync_slot(cudaStream_t stream);

}  // namespace gi2d
