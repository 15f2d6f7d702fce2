// project.cu -- 2D projection kernels (forward/backward, Cholesky and scale-rotation parametrisations).
//
// Replaces the reference's project_gaussians_2d_forward_kernel / ..._scale_rot_forward_kernel
// (gsplat/gsplat/cuda/csrc/foward2d.cu:12-123), the two backward kernels (backward2d.cu:8-109) and
// compute_cov2d_bounds_kernel (bindings.cu:21-39) together with their torch::zeros fills
// (bindings.cu:808-817, 870-876, 918-928, 968-981): every kernel here writes EVERY output element, so no
// memset precedes it.  HBM-bound: 52 B/Gaussian forward, 68 B/Gaussian backward.
//
// Layout: the API keeps the reference's array-of-structs tensors ((N,2),(N,3) fp32).  A 256-Gaussian block moves its
// 3-float rows as 192 coalesced 16-byte vectors through shared memory (rows start at 256*12 B = 3072 B multiples),
// so global traffic is full 128-B lines in both directions; the per-thread reads of shared memory are stride-3 words,
// which is bank-conflict free.
#include <cuda_fp16.h>

#include "gi2d_common.cuh"

namespace gi2d {

constexpr int PB = 256;  // Gaussians per block

// cooperative (N,3) row staging: global -> shared
template <bool VEC>
__device__ __forceinline__ void stage_in3(float *s, const float *g, int base, int nb) {
    const int tid = threadIdx.x;
    if (VEC && nb == PB) {
        if (tid < PB * 3 / 4) reinterpret_cast<float4 *>(s)[tid] = __ldg(reinterpret_cast<const float4 *>(g + (size_t)base * 3) + tid);
    } else {
        for (int k = tid; k < nb * 3; k += PB) s[k] = __ldg(g + (size_t)base * 3 + k);
    }
}
template <bool VEC>
__device__ __forceinline__ void stage_out3(const float *s, float *g, int base, int nb) {
    const int tid = threadIdx.x;
    if (VEC && nb == PB) {
        if (tid < PB * 3 / 4) reinterpret_cast<float4 *>(g + (size_t)base * 3)[tid] = reinterpret_cast<const float4 *>(s)[tid];
    } else {
        for (int k = tid; k < nb * 3; k += PB) g[(size_t)base * 3 + k] = s[k];
    }
}

// Shared tail of both forward kernels (foward2d.cu:50-67 / :105-122): conic, radius, tile extent, tile count.
__device__ __forceinline__ void project_tail(float cx, float cy, float a, float b, float c, int TX, int TY, float2 &xy,
                                             int &radius, float3 &conic, int &tiles) {
    float r3;
    xy = make_float2(0.f, 0.f);
    conic = make_float3(0.f, 0.f, 0.f);
    radius = 0;
    tiles = 0;
    if (!cov2d_bounds(a, b, c, conic, r3)) {   // det == 0: the reference leaves its zero-filled outputs (Q11)
        conic = make_float3(0.f, 0.f, 0.f);
        return;
    }
    xy = make_float2(cx, cy);
    radius = __float2int_ru(r3);               // (int)ceil(.) -- written before the area test (Q11)
    int x0, x1, y0, y1;
    tile_bbox(cx, cy, ceilf(r3), TX, TY, x0, x1, y0, y1);
    const int area = (int)((unsigned)(x1 - x0) * (unsigned)(y1 - y0));
    if (area > 0) tiles = area;
}

// MODE 0: Cholesky (l11,l21,l22) rows in p3; MODE 1: scale-rotation, scales (N,2) in p2, rotation (N,) in p1.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(PB) project_fwd_kernel(int N, const float *__restrict__ means2d,
                                                         const float *__restrict__ p3, const float *__restrict__ p2,
                                                         const float *__restrict__ p1, float hw, float hh, int TX,
                                                         int TY, float *__restrict__ xys, float *__restrict__ depths,
                                                         int *__restrict__ radii, float *__restrict__ conics,
                                                         int *__restrict__ num_tiles_hit, unsigned int *total_slot = nullptr,
                                                         int *total_dev = nullptr, volatile int *total_host = nullptr) {
    __shared__ __align__(16) float s3[PB * 3];
    const int base = blockIdx.x * PB, tid = threadIdx.x, i = base + tid;
    const int nb = min(PB, N - base);
    if (MODE == 0) {
        stage_in3<VEC>(s3, p3, base, nb);
        __syncthreads();
    }
    float2 xy = make_float2(0.f, 0.f);
    float3 conic = make_float3(0.f, 0.f, 0.f);
    int radius = 0, tiles = 0;
    if (i < N) {
        float mx, my;
        if (VEC) { const float2 m = __ldg(reinterpret_cast<const float2 *>(means2d) + i); mx = m.x; my = m.y; }
        else { mx = __ldg(means2d + 2 * (size_t)i); my = __ldg(means2d + 2 * (size_t)i + 1); }
        const float cx = ffma(hw, mx, hw);          // foward2d.cu:37-38: 0.5*W*mx + 0.5*W, one FFMA in the JIT build
        const float cy = ffma(my, hh, hh);
        float a, b, c;
        if (MODE == 0) {
            const float l11 = s3[3 * tid], l21 = s3[3 * tid + 1], l22 = s3[3 * tid + 2];
            a = fmul(l11, l11);                     // foward2d.cu:47
            b = fmul(l11, l21);
            c = ffma(l21, l21, fmul(l22, l22));
        } else {
            float sx, sy;
            if (VEC) { const float2 s = __ldg(reinterpret_cast<const float2 *>(p2) + i); sx = s.x; sy = s.y; }
            else { sx = __ldg(p2 + 2 * (size_t)i); sy = __ldg(p2 + 2 * (size_t)i + 1); }
            const float th = __ldg(p1 + i);
            const float cs = cosf(th), sn = sinf(th);   // helpers.cuh:274-285 (libdevice, same code as the JIT build)
            // M = R*S (glm column-major), cov = M*M^T; FFMA placement of the JIT build, see oracle/gi2d_oracle.c
            const float p = fmul(cs, sx), u = -fmul(sn, sx), q = fmul(sn, sy), v = fmul(cs, sy);
            a = ffma(p, p, fmul(q, q));
            b = ffma(p, u, fmul(q, v));
            c = ffma(u, u, fmul(v, v));
        }
        project_tail(cx, cy, a, b, c, TX, TY, xy, radius, conic, tiles);
        if (VEC) reinterpret_cast<float2 *>(xys)[i] = xy;
        else { xys[2 * (size_t)i] = xy.x; xys[2 * (size_t)i + 1] = xy.y; }
        depths[i] = 0.f;                            // foward2d.cu:67 (and the zero fill otherwise)
        radii[i] = radius;
        num_tiles_hit[i] = tiles;
    }
    __syncthreads();                                // s3 is reused for the conic rows
    s3[3 * tid] = conic.x; s3[3 * tid + 1] = conic.y; s3[3 * tid + 2] = conic.z;
    __syncthreads();
    stage_out3<VEC>(s3, conics, base, nb);
    // The intersection total (what compute_cumulative_intersects' `.item()` reads, utils.py:116-118) as a by-product: block
    // sums meet in a library-owned word, the last block to arrive publishes the total -- to device memory and/or to a MAPPED
    // PINNED HOST word the host can poll without a stream synchronisation -- and leaves the pair of words zero again.
    if (total_slot) {
        __shared__ int s_sum[PB / 32];
        const int w = __reduce_add_sync(0xffffffffu, tiles);
        if ((tid & 31) == 0) s_sum[tid >> 5] = w;
        __syncthreads();
        if (tid == 0) {
            int b = 0;
#pragma unroll
            for (int k = 0; k < PB / 32; ++k) b += s_sum[k];
            // ONE 64-bit atomic per block on the pair of words: low word += block sum, high word += 1 (arrivals)
            unsigned long long *pair = reinterpret_cast<unsigned long long *>(total_slot);
            const unsigned long long old = atomicAdd(pair, (1ull << 32) | (unsigned long long)(unsigned int)b);
            if ((unsigned int)(old >> 32) == gridDim.x - 1) {             // last block: total complete, leave the pair zero
                const int total = (int)((unsigned int)old + (unsigned int)b);
                atomicExch(pair, 0ull);
                if (total_dev) *total_dev = total;
                if (total_host) *total_host = total;        // the word stands alone: nothing else has to be visible with it
            }
        }
    }
}

// reference helpers.cuh:71-82 cov2d_to_conic_vjp: V = -X G X, v_cov2d = (V00, V10+V01, V11).
__device__ __forceinline__ float3 conic_vjp(float3 X, float3 G) {
    const float a00 = X.x * G.x + X.y * G.y, a01 = X.x * G.y + X.y * G.z;
    const float a10 = X.y * G.x + X.z * G.y, a11 = X.y * G.y + X.z * G.z;
    const float v00 = -(a00 * X.x + a01 * X.y), v01 = -(a00 * X.y + a01 * X.z);
    const float v10 = -(a10 * X.x + a11 * X.y), v11 = -(a10 * X.y + a11 * X.z);
    return make_float3(v00, v10 + v01, v11);
}

// Backward of both parametrisations (backward2d.cu:8-109), including the reference's doubled off-diagonal term (Q5).
// MODE 0: out3 = v_L (N,3).  MODE 1: out2 = v_scale (N,2), out1 = v_rot (N,).  v_cov2d may be null.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(PB) project_bwd_kernel(int N, const float *__restrict__ p3, const float *__restrict__ p2,
                                                         const float *__restrict__ p1, float hw, float hh,
                                                         const int *__restrict__ radii, const float *__restrict__ conics,
                                                         const float *__restrict__ v_xy, const float *__restrict__ v_conic,
                                                         float *__restrict__ v_cov2d, float *__restrict__ v_mean2d,
                                                         float *__restrict__ out3, float *__restrict__ out2,
                                                         float *__restrict__ out1, const float4 *__restrict__ g12) {
    __shared__ __align__(16) float sA[PB * 3];   // conics, then v_cov2d
    __shared__ __align__(16) float sB[PB * 3];   // v_conic, then v_L
    __shared__ __align__(16) float sC[PB * 3];   // L
    const int base = blockIdx.x * PB, tid = threadIdx.x, i = base + tid;
    const int nb = min(PB, N - base);
    stage_in3<VEC>(sA, conics, base, nb);
    if (!g12) stage_in3<VEC>(sB, v_conic, base, nb);
    if (MODE == 0) stage_in3<VEC>(sC, p3, base, nb);
    __syncthreads();
    float3 G = make_float3(0.f, 0.f, 0.f), o3 = make_float3(0.f, 0.f, 0.f);
    if (i < N) {
        float2 vm = make_float2(0.f, 0.f), o2 = make_float2(0.f, 0.f);
        float o1 = 0.f;
        if (__ldg(radii + i) > 0) {
            float vx, vy;
            float3 vc;
            if (g12) {      // (N,12) rows of gi2d_raster_sum_bwd_packed: {v_xy.x, v_xy.y, v_conic.x, v_conic.y | v_conic.z, ...}
                const float4 a = __ldg(g12 + 3 * (size_t)i);
                vx = a.x; vy = a.y;
                vc = make_float3(a.z, a.w, __ldg(reinterpret_cast<const float *>(g12) + 12 * (size_t)i + 4));
            } else {
                vc = make_float3(sB[3 * tid], sB[3 * tid + 1], sB[3 * tid + 2]);
                if (VEC) { const float2 v = __ldg(reinterpret_cast<const float2 *>(v_xy) + i); vx = v.x; vy = v.y; }
                else { vx = __ldg(v_xy + 2 * (size_t)i); vy = __ldg(v_xy + 2 * (size_t)i + 1); }
            }
            G = conic_vjp(make_float3(sA[3 * tid], sA[3 * tid + 1], sA[3 * tid + 2]), vc);
            vm = make_float2(vx * hw, vy * hh);                          // backward2d.cu:48-49
            if (MODE == 0) {
                const float l11 = sC[3 * tid], l21 = sC[3 * tid + 1], l22 = sC[3 * tid + 2];
                o3.x = 2.f * l11 * G.x + 2.f * G.y * l21;                // backward2d.cu:39-41
                o3.y = 2.f * l11 * G.y + 2.f * l21 * G.z;
                o3.z = 2.f * l22 * G.z;
            } else {
                float sx, sy;
                if (VEC) { const float2 s = __ldg(reinterpret_cast<const float2 *>(p2) + i); sx = s.x; sy = s.y; }
                else { sx = __ldg(p2 + 2 * (size_t)i); sy = __ldg(p2 + 2 * (size_t)i + 1); }
                const float th = __ldg(p1 + i);
                const float cs = cosf(th), sn = sinf(th);
                // (row,col) matrices: R = [[cs, sn],[-sn, cs]], Rg = dR/dtheta = [[-sn, cs],[-cs, -sn]], S = diag(sx,sy)
                // Sigma = R S S R^T.  d/dsx = R diag(2sx,0) R^T, d/dsy = R diag(0,2sy) R^T,
                // d/dtheta = Rg S (RS)^T + (RS) S Rg^T   (backward2d.cu:84-96)
                const float sxx = sx * sx, syy = sy * sy;
                // sigma_x_g = 2sx * r0 r0^T with r0 = column 0 of R = (cs, -sn)
                const float X00 = 2.f * sx * cs * cs, X10 = -2.f * sx * cs * sn, X11 = 2.f * sx * sn * sn;
                // sigma_y_g = 2sy * r1 r1^T with r1 = column 1 of R = (sn, cs)
                const float Y00 = 2.f * sy * sn * sn, Y10 = 2.f * sy * sn * cs, Y11 = 2.f * sy * cs * cs;
                // theta_g = A + A^T with A = Rg S^2 R^T
                const float A00 = -sn * sxx * cs + cs * syy * sn;        // row0(Rg S^2) . row0(R)
                const float A01 = sn * sxx * sn + cs * syy * cs;         // row0(Rg S^2) . row1(R)
                const float A10 = -cs * sxx * cs - sn * syy * sn;        // row1(Rg S^2) . row0(R)
                const float A11 = cs * sxx * sn - sn * syy * cs;
                const float T00 = 2.f * A00, T10 = A10 + A01, T11 = 2.f * A11;
                o2.x = G.x * X00 + 2.f * G.y * X10 + G.z * X11;          // backward2d.cu:102-104 (Q5: 2*G12)
                o2.y = G.x * Y00 + 2.f * G.y * Y10 + G.z * Y11;
                o1 = G.x * T00 + 2.f * G.y * T10 + G.z * T11;
            }
        }
        if (VEC) reinterpret_cast<float2 *>(v_mean2d)[i] = vm;
        else { v_mean2d[2 * (size_t)i] = vm.x; v_mean2d[2 * (size_t)i + 1] = vm.y; }
        if (MODE == 1) {
            if (VEC) reinterpret_cast<float2 *>(out2)[i] = o2;
            else { out2[2 * (size_t)i] = o2.x; out2[2 * (size_t)i + 1] = o2.y; }
            out1[i] = o1;
        }
    }
    __syncthreads();
    sA[3 * tid] = G.x; sA[3 * tid + 1] = G.y; sA[3 * tid + 2] = G.z;
    if (MODE == 0) { sB[3 * tid] = o3.x; sB[3 * tid + 1] = o3.y; sB[3 * tid + 2] = o3.z; }
    __syncthreads();
    if (v_cov2d) stage_out3<VEC>(sA, v_cov2d, base, nb);
    if (MODE == 0) stage_out3<VEC>(sB, out3, base, nb);
}

// reference bindings.cu:21-39 compute_cov2d_bounds_kernel: conic (N,3), radius (N,1) as float; rows with det==0 -> 0.
__global__ void __launch_bounds__(PB) cov2d_bounds_kernel(int N, const float *__restrict__ cov2d, float *__restrict__ conics,
                                                          float *__restrict__ radii) {
    const int i = blockIdx.x * PB + threadIdx.x;
    if (i >= N) return;
    float3 conic;
    float r3;
    const bool ok = cov2d_bounds(__ldg(cov2d + 3 * (size_t)i), __ldg(cov2d + 3 * (size_t)i + 1), __ldg(cov2d + 3 * (size_t)i + 2), conic, r3);
    conics[3 * (size_t)i] = ok ? conic.x : 0.f;
    conics[3 * (size_t)i + 1] = ok ? conic.y : 0.f;
    conics[3 * (size_t)i + 2] = ok ? conic.z : 0.f;
    radii[i] = ok ? ceilf(r3) : 0.f;
}

static inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline int cuda_status() {
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? GI2D_OK : GI2D_ERR_CUDA_BASE + (int)e;
}

// ---------------------------------------------------------------------------------------------------------------
// Codec decode prologue fused into the Cholesky projection: what decompress_wo_ec does with torch ops
// before every decode render (gaussianimage_cholesky.py:161-166) --
//     means    = tanh(xyz.float())                                   fp16 positions
//     cholesky = codes * scale + beta + cholesky_bound               UniformQuantizer.decompress (quantize.py:86-87)
//     colors   = sum_k codebook_k[index[:, k]]                       VectorQuantizer.decompress (quantize.py:146-150)
// -- and project_gaussians_2d, in one pass: 4+3(+12)+8K B in, 32+12 B out per Gaussian instead of seven kernels that
// write and re-read means / cholesky / colours.  Roundings follow the torch ops one by one (mul, add, add; 0 + a + b).
template <bool CODE_F32, bool IDX64>
__global__ void __launch_bounds__(PB) decode_project_kernel(int N, const __half2 *__restrict__ xyz, const void *__restrict__ codes,
                                                            const float *__restrict__ scale3, const float *__restrict__ beta3,
                                                            const float *__restrict__ bound3, const void *__restrict__ index,
                                                            int K, int C, const float *__restrict__ codebooks, float hw, float hh,
                                                            int TX, int TY, float *__restrict__ xys, float *__restrict__ depths,
                                                            int *__restrict__ radii, float *__restrict__ conics,
                                                            int *__restrict__ num_tiles_hit, float *__restrict__ colors) {
    extern __shared__ __align__(16) float s_cb[];      // K*C*3 codebook floats (when they fit), then 3*PB staging floats
    const int ncb = K * C * 3;
    float *s3 = s_cb + ((ncb + 3) & ~3);
    const int base = blockIdx.x * PB, tid = threadIdx.x, i = base + tid;
    const int nb = min(PB, N - base);
    for (int k = tid; k < ncb; k += PB) s_cb[k] = __ldg(codebooks + k);
    __syncthreads();
    float2 xy = make_float2(0.f, 0.f);
    float3 conic = make_float3(0.f, 0.f, 0.f), col = make_float3(0.f, 0.f, 0.f);
    int radius = 0, tiles = 0;
    if (i < N) {
        const float2 z = __half22float2(xyz[i]);
        const float mx = tanhf(z.x), my = tanhf(z.y);            // gaussianimage_cholesky.py:163 (libdevice tanhf, as torch)
        float q0, q1, q2;
        if (CODE_F32) {
            const float *c = reinterpret_cast<const float *>(codes) + 3 * (size_t)i;
            q0 = __ldg(c); q1 = __ldg(c + 1); q2 = __ldg(c + 2);
        } else {
            const uint8_t *c = reinterpret_cast<const uint8_t *>(codes) + 3 * (size_t)i;
            q0 = (float)__ldg(c); q1 = (float)__ldg(c + 1); q2 = (float)__ldg(c + 2);
        }
        const float3 scale = make_float3(__ldg(scale3), __ldg(scale3 + 1), __ldg(scale3 + 2));
        const float3 beta = make_float3(__ldg(beta3), __ldg(beta3 + 1), __ldg(beta3 + 2));
        const float3 bound = make_float3(__ldg(bound3), __ldg(bound3 + 1), __ldg(bound3 + 2));
        const float l11 = __fadd_rn(__fadd_rn(__fmul_rn(q0, scale.x), beta.x), bound.x);    // quantize.py:87, then :165
        const float l21 = __fadd_rn(__fadd_rn(__fmul_rn(q1, scale.y), beta.y), bound.y);
        const float l22 = __fadd_rn(__fadd_rn(__fmul_rn(q2, scale.z), beta.z), bound.z);
        for (int k = 0; k < K; ++k) {                            // quantize.py:147-149: recon = 0; recon += embed[0, index[:, k]]
            long long e = IDX64 ? __ldg(reinterpret_cast<const long long *>(index) + (size_t)i * K + k)
                                : (long long)__ldg(reinterpret_cast<const int *>(index) + (size_t)i * K + k);
            e = e < 0 ? 0 : (e >= C ? C - 1 : e);               // torch would raise; never read outside the codebook
            const float *row = s_cb + ((size_t)k * C + (size_t)e) * 3;
            col.x = __fadd_rn(col.x, row[0]); col.y = __fadd_rn(col.y, row[1]); col.z = __fadd_rn(col.z, row[2]);
        }
        const float cx = ffma(hw, mx, hw);
        const float cy = ffma(my, hh, hh);
        const float a = fmul(l11, l11), b = fmul(l11, l21), c = ffma(l21, l21, fmul(l22, l22));
        project_tail(cx, cy, a, b, c, TX, TY, xy, radius, conic, tiles);
        reinterpret_cast<float2 *>(xys)[i] = xy;
        depths[i] = 0.f;
        radii[i] = radius;
        num_tiles_hit[i] = tiles;
    }
    s3[3 * tid] = conic.x; s3[3 * tid + 1] = conic.y; s3[3 * tid + 2] = conic.z;
    __syncthreads();
    stage_out3<true>(s3, conics, base, nb);
    __syncthreads();
    s3[3 * tid] = col.x; s3[3 * tid + 1] = col.y; s3[3 * tid + 2] = col.z;
    __syncthreads();
    stage_out3<true>(s3, colors, base, nb);
}

}  // namespace gi2d

using namespace gi2d;

extern "C" {

int gi2d_project_cholesky_fwd(int N, const float *means2d, const float *L, int H, int W, int TX, int TY, float *xys,
                              float *depths, int *radii, float *conics, int *num_tiles_hit, cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    const float hw = (float)(unsigned)W * 0.5f, hh = (float)(unsigned)H * 0.5f;
    const int grid = (N + PB - 1) / PB;
    const bool vec = aligned16(means2d) && aligned16(L) && aligned16(xys) && aligned16(conics);
    if (vec)
        project_fwd_kernel<0, true><<<grid, PB, 0, stream>>>(N, means2d, L, nullptr, nullptr, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit);
    else
        project_fwd_kernel<0, false><<<grid, PB, 0, stream>>>(N, means2d, L, nullptr, nullptr, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit);
    return cuda_status();
}

int gi2d_project_scale_rot_fwd(int N, const float *means2d, const float *scales2d, const float *rotation, int H, int W,
                               int TX, int TY, float *xys, float *depths, int *radii, float *conics, int *num_tiles_hit,
                               cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    const float hw = (float)(unsigned)W * 0.5f, hh = (float)(unsigned)H * 0.5f;
    const int grid = (N + PB - 1) / PB;
    const bool vec = aligned16(means2d) && aligned16(scales2d) && aligned16(xys) && aligned16(conics);
    if (vec)
        project_fwd_kernel<1, true><<<grid, PB, 0, stream>>>(N, means2d, nullptr, scales2d, rotation, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit);
    else
        project_fwd_kernel<1, false><<<grid, PB, 0, stream>>>(N, means2d, nullptr, scales2d, rotation, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit);
    return cuda_status();
}

// Either projection, additionally publishing the total of num_tiles_hit (see the kernel): total_dev (device int) and / or
// total_host (mapped pinned host int, preset by the caller to a negative value and polled) may be NULL.
int gi2d_project_fwd_total(int mode, int N, const float *means2d, const float *p_a, const float *p_b, int H, int W, int TX, int TY,
                           float *xys, float *depths, int *radii, float *conics, int *num_tiles_hit, int *total_dev,
                           int *total_host, cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0 || (mode != 0 && mode != 1)) return GI2D_ERR_INVALID;
    if (N == 0) {
        if (total_dev) cudaMemsetAsync(total_dev, 0, sizeof(int), stream);
        if (total_host) *reinterpret_cast<volatile int *>(total_host) = 0;
        return cuda_status();
    }
    unsigned int *slot = (total_dev || total_host) ? sync_slot(stream) : nullptr;
    if ((total_dev || total_host) && !slot) return GI2D_ERR_INVALID;
    const float hw = (float)(unsigned)W * 0.5f, hh = (float)(unsigned)H * 0.5f;
    const int grid = (N + PB - 1) / PB;
    const bool vec = aligned16(means2d) && aligned16(p_a) && aligned16(xys) && aligned16(conics);
    if (mode == 0) {
        if (vec) project_fwd_kernel<0, true><<<grid, PB, 0, stream>>>(N, means2d, p_a, nullptr, nullptr, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit, slot, total_dev, total_host);
        else project_fwd_kernel<0, false><<<grid, PB, 0, stream>>>(N, means2d, p_a, nullptr, nullptr, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit, slot, total_dev, total_host);
    } else {
        if (vec) project_fwd_kernel<1, true><<<grid, PB, 0, stream>>>(N, means2d, nullptr, p_a, p_b, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit, slot, total_dev, total_host);
        else project_fwd_kernel<1, false><<<grid, PB, 0, stream>>>(N, means2d, nullptr, p_a, p_b, hw, hh, TX, TY, xys, depths, radii, conics, num_tiles_hit, slot, total_dev, total_host);
    }
    return cuda_status();
}

int gi2d_project_cholesky_bwd(int N, const float *L, int H, int W, const int *radii, const float *conics,
                              const float *v_xy, const float *v_conic, float *v_cov2d, float *v_mean2d, float *v_L,
                              cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    const float hw = 0.5f * (float)(unsigned)W, hh = 0.5f * (float)(unsigned)H;
    const int grid = (N + PB - 1) / PB;
    const bool vec = aligned16(L) && aligned16(conics) && aligned16(v_xy) && aligned16(v_conic) && aligned16(v_mean2d) &&
                     aligned16(v_L) && (v_cov2d == nullptr || aligned16(v_cov2d));
    if (vec)
        project_bwd_kernel<0, true><<<grid, PB, 0, stream>>>(N, L, nullptr, nullptr, hw, hh, radii, conics, v_xy, v_conic, v_cov2d, v_mean2d, v_L, nullptr, nullptr, nullptr);
    else
        project_bwd_kernel<0, false><<<grid, PB, 0, stream>>>(N, L, nullptr, nullptr, hw, hh, radii, conics, v_xy, v_conic, v_cov2d, v_mean2d, v_L, nullptr, nullptr, nullptr);
    return cuda_status();
}

int gi2d_project_scale_rot_bwd(int N, const float *scales2d, const float *rotation, int H, int W, const int *radii,
                               const float *conics, const float *v_xy, const float *v_conic, float *v_cov2d,
                               float *v_mean2d, float *v_scale, float *v_rot, cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    const float hw = 0.5f * (float)(unsigned)W, hh = 0.5f * (float)(unsigned)H;
    const int grid = (N + PB - 1) / PB;
    const bool vec = aligned16(scales2d) && aligned16(conics) && aligned16(v_xy) && aligned16(v_conic) &&
                     aligned16(v_mean2d) && aligned16(v_scale) && (v_cov2d == nullptr || aligned16(v_cov2d));
    if (vec)
        project_bwd_kernel<1, true><<<grid, PB, 0, stream>>>(N, nullptr, scales2d, rotation, hw, hh, radii, conics, v_xy, v_conic, v_cov2d, v_mean2d, nullptr, v_scale, v_rot, nullptr);
    else
        project_bwd_kernel<1, false><<<grid, PB, 0, stream>>>(N, nullptr, scales2d, rotation, hw, hh, radii, conics, v_xy, v_conic, v_cov2d, v_mean2d, nullptr, v_scale, v_rot, nullptr);
    return cuda_status();
}

// The same two backward passes reading v_xy / v_conic straight from the (N,12) gradient rows written by
// gi2d_raster_sum_bwd_packed (no intermediate (N,2)/(N,3) tensors).
int gi2d_project_cholesky_bwd_packed(int N, const float *L, int H, int W, const int *radii, const float *conics,
                                     const float *grads12, float *v_mean2d, float *v_L, cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0 || !aligned16(grads12)) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    const float hw = 0.5f * (float)(unsigned)W, hh = 0.5f * (float)(unsigned)H;
    const int grid = (N + PB - 1) / PB;
    const float4 *g = reinterpret_cast<const float4 *>(grads12);
    if (aligned16(L) && aligned16(conics) && aligned16(v_mean2d) && aligned16(v_L))
        project_bwd_kernel<0, true><<<grid, PB, 0, stream>>>(N, L, nullptr, nullptr, hw, hh, radii, conics, nullptr, nullptr, nullptr, v_mean2d, v_L, nullptr, nullptr, g);
    else
        project_bwd_kernel<0, false><<<grid, PB, 0, stream>>>(N, L, nullptr, nullptr, hw, hh, radii, conics, nullptr, nullptr, nullptr, v_mean2d, v_L, nullptr, nullptr, g);
    return cuda_status();
}

int gi2d_project_scale_rot_bwd_packed(int N, const float *scales2d, const float *rotation, int H, int W, const int *radii,
                                      const float *conics, const float *grads12, float *v_mean2d, float *v_scale,
                                      float *v_rot, cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0 || !aligned16(grads12)) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    const float hw = 0.5f * (float)(unsigned)W, hh = 0.5f * (float)(unsigned)H;
    const int grid = (N + PB - 1) / PB;
    const float4 *g = reinterpret_cast<const float4 *>(grads12);
    if (aligned16(scales2d) && aligned16(conics) && aligned16(v_mean2d) && aligned16(v_scale))
        project_bwd_kernel<1, true><<<grid, PB, 0, stream>>>(N, nullptr, scales2d, rotation, hw, hh, radii, conics, nullptr, nullptr, nullptr, v_mean2d, nullptr, v_scale, v_rot, g);
    else
        project_bwd_kernel<1, false><<<grid, PB, 0, stream>>>(N, nullptr, scales2d, rotation, hw, hh, radii, conics, nullptr, nullptr, nullptr, v_mean2d, nullptr, v_scale, v_rot, g);
    return cuda_status();
}

int gi2d_decode_project_cholesky_fwd(int N, const void *xyz_half, const void *codes, int codes_are_f32, const float *scale3,
                                     const float *beta3, const float *bound3, const void *index, int index_is_i64,
                                     int n_quantizers, int codebook_size, const float *codebooks, int H, int W, int TX, int TY,
                                     float *xys, float *depths, int *radii, float *conics, int *num_tiles_hit, float *colors,
                                     cudaStream_t stream) {
    if (N < 0 || H <= 0 || W <= 0 || n_quantizers < 0 || codebook_size <= 0 || !scale3 || !beta3 || !bound3) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    if (!xyz_half || !codes || (n_quantizers && (!index || !codebooks))) return GI2D_ERR_INVALID;
    if (!aligned16(xys) || !aligned16(conics) || !aligned16(colors) || (reinterpret_cast<uintptr_t>(xyz_half) & 3)) return GI2D_ERR_INVALID;
    const size_t ncb = (size_t)n_quantizers * codebook_size * 3;
    const size_t smem = (((ncb + 3) & ~(size_t)3) + 3 * PB) * sizeof(float);
    if (smem > 200 * 1024) return GI2D_ERR_INVALID;           // codebooks of the codec are 2 x 8 x 3 floats
    const float hw = (float)(unsigned)W * 0.5f, hh = (float)(unsigned)H * 0.5f;
    const int grid = (N + PB - 1) / PB;
#define GI2D_DECODE(F, I)                                                                                                 \
    {                                                                                                                     \
        if (smem > 48 * 1024) cudaFuncSetAttribute(decode_project_kernel<F, I>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        decode_project_kernel<F, I><<<grid, PB, smem, stream>>>(N, reinterpret_cast<const __half2 *>(xyz_half), codes, scale3, beta3, bound3, index, \
                                                                n_quantizers, codebook_size, codebooks, hw, hh, TX, TY, xys, depths, \
                                                                radii, conics, num_tiles_hit, colors);                   \
    }
    if (codes_are_f32) { if (index_is_i64) GI2D_DECODE(true, true) else GI2D_DECODE(true, false) }
    else { if (index_is_i64) GI2D_DECODE(false, true) else GI2D_DECODE(false, false) }
#undef GI2D_DECODE
    return cuda_status();
}

int gi2d_cov2d_bounds(int N, const float *cov2d, float *conics, float *radii, cudaStream_t stream) {
    if (N < 0) return GI2D_ERR_INVALID;
    if (N == 0) return GI2D_OK;
    cov2d_bounds_kernel<<<(N + PB - 1) / PB, PB, 0, stream>>>(N, cov2d, conics, radii);
    return cuda_status();
}

}  // extern "C"
