// adan.cu -- the Adan update of the reference's optimizer (optimizer.py:232-279 `_single_tensor_adan`, the arithmetic
// definition; optimizer.py:296-362 is the same sequence as ~25 torch._foreach launches) as ONE kernel over all the
// tensors of a parameter group.  It fills the reference's own plug-in seam: Adan(fused=True) imports an external
// module `fused_adan` (optimizer.py:385,427,453; the CUDA extension of github.com/sail-sg/Adan, which the reference
// neither vendors nor pins) and calls adan_multi_tensor / adan_single_tensor.
//
// HBM-bound elementwise work: 6 arrays read, 5 (6 when the gradient is clipped) written, 44 B per element; float4
// accesses, a block owns 1024 consecutive elements of one tensor, the (tensor, chunk) of a block comes from a table
// passed by value.  Two ways to supply the scalars:
//   host   the ten floats the reference passes (bias corrections computed by the host in double, optimizer.py:169-171);
//   device hyper-parameters and the step counter live in device memory, bias corrections are computed in double on
//          the device and the last block to finish advances the counter: the step can sit inside a CUDA graph.
#include <cstdint>
#include <algorithm>
#include <cuda_runtime.h>

#include "../../include/gi2d.h"

namespace gi2d {

constexpr int ADAN_MAX_TENSORS = 16;
constexpr int ADAN_THREADS = 256;
constexpr int ADAN_CHUNK = ADAN_THREADS * 4;      // elements per block: one float4 per thread and array (C2: 547 blocks)

struct AdanTable {
    float *p[ADAN_MAX_TENSORS], *g[ADAN_MAX_TENSORS], *m[ADAN_MAX_TENSORS], *v[ADAN_MAX_TENSORS], *d[ADAN_MAX_TENSORS],
        *n[ADAN_MAX_TENSORS];
    long long numel[ADAN_MAX_TENSORS];
    int blk_end[ADAN_MAX_TENSORS];                // exclusive prefix end of the blocks of tensor i
    int count;
};

struct AdanScalars {
    float beta1, beta2, beta3, one_m_beta1, one_m_beta2, one_m_beta3;
    float bc3_sqrt, eps, neg_step, neg_step_diff, decay;     // decay = 1 - lr*wd (no_prox) or 1 + lr*wd
    float clip;
    int no_prox, write_neg;
};

// Everything the reference computes on the host in Python doubles (optimizer.py:169-171, :210-212, :262-263) and then
// hands to fp32 kernels as scalars.
__host__ __device__ inline AdanScalars adan_scalars(double beta1, double beta2, double beta3, double bc1, double bc2,
                                                    double bc3_sqrt, double lr, double wd, double eps, int no_prox,
                                                    double clip, int write_neg) {
    AdanScalars s;
    s.beta1 = (float)beta1; s.beta2 = (float)beta2; s.beta3 = (float)beta3;
    s.one_m_beta1 = (float)(1.0 - beta1); s.one_m_beta2 = (float)(1.0 - beta2); s.one_m_beta3 = (float)(1.0 - beta3);
    s.bc3_sqrt = (float)bc3_sqrt; s.eps = (float)eps;
    s.neg_step = (float)(-(lr / bc1));
    s.neg_step_diff = (float)(-(lr * beta2 / bc2));
    s.decay = (float)(no_prox ? 1.0 - lr * wd : 1.0 + lr * wd);
    s.clip = (float)clip;
    s.no_prox = no_prox; s.write_neg = write_neg;
    return s;
}

// One element, in the reference's order of operations (optimizer.py:240-279).
__device__ __forceinline__ void adan_element(float &p, float &g, float &m, float &v, float &d, float &n, const AdanScalars &s) {
    // Roundings placed where torch's CUDA kernels place them: every in-place op rounds once; add_(x, alpha) and
    // addcdiv_/addcmul_ are a multiply feeding one fused multiply-add.
    g = __fmul_rn(g, s.clip);                                                     // :240  grad.mul_(clip)
    n = __fadd_rn(n, g);                                                          // :244  neg_pre_grad + grad = grad difference
    m = __fmaf_rn(g, s.one_m_beta1, __fmul_rn(m, s.beta1));                       // :246  m_t
    d = __fmaf_rn(n, s.one_m_beta2, __fmul_rn(d, s.beta2));                       // :247  diff_t
    n = __fadd_rn(__fmul_rn(n, s.beta2), g);                                      // :250
    v = __fmaf_rn(__fmul_rn(s.one_m_beta3, n), n, __fmul_rn(v, s.beta3));         // :251  n_t
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), s.bc3_sqrt), s.eps);   // :255
    if (s.no_prox) {
        p = __fmul_rn(p, s.decay);                                                // :260
        p = __fmaf_rn(s.neg_step, __fdiv_rn(m, denom), p);                        // :261
        p = __fmaf_rn(s.neg_step_diff, __fdiv_rn(d, denom), p);
    } else {
        p = __fmaf_rn(s.neg_step, __fdiv_rn(m, denom), p);                        // :264
        p = __fmaf_rn(s.neg_step_diff, __fdiv_rn(d, denom), p);
        p = __fdiv_rn(p, s.decay);                                                // :266
    }
    if (s.write_neg) n = -g;                          // :268  neg_pre_grad = -grad (for the next step)
}

// device-resident hyper-parameters: doubles {beta1, beta2, beta3, lr, weight_decay, eps, clip}, ints {step, ticket}
struct AdanDeviceState {
    double beta1, beta2, beta3, lr, weight_decay, eps, clip;
    int step, ticket;
};

template <bool DEV>
__global__ void __launch_bounds__(ADAN_THREADS) adan_kernel(AdanTable t, AdanScalars hs, AdanDeviceState *st, int no_prox) {
    __shared__ AdanScalars s_s;
    AdanScalars s;
    if (DEV) {
        if (threadIdx.x == 0) {
            const int step = st->step + 1;            // every block reads the counter before the last one advances it
            const double b1 = st->beta1, b2 = st->beta2, b3 = st->beta3;
            s_s = adan_scalars(b1, b2, b3, 1.0 - pow(b1, (double)step), 1.0 - pow(b2, (double)step),
                               sqrt(1.0 - pow(b3, (double)step)), st->lr, st->weight_decay, st->eps, no_prox, st->clip, 1);
        }
        __syncthreads();
        s = s_s;
    } else {
        s = hs;
    }
    int ti = 0;
    while (ti < t.count - 1 && (int)blockIdx.x >= t.blk_end[ti]) ++ti;
    const int blk0 = ti ? t.blk_end[ti - 1] : 0;
    const long long numel = t.numel[ti];
    const long long base = (long long)(blockIdx.x - blk0) * ADAN_CHUNK;
    float *P = t.p[ti], *G = t.g[ti], *M = t.m[ti], *V = t.v[ti], *D = t.d[ti], *Nn = t.n[ti];
    const bool write_g = s.clip != 1.f;
    const bool vec = ((((uintptr_t)P | (uintptr_t)G | (uintptr_t)M | (uintptr_t)V | (uintptr_t)D | (uintptr_t)Nn) & 15) == 0);
#pragma unroll
    for (int it = 0; it < ADAN_CHUNK / (4 * ADAN_THREADS); ++it) {
        const long long i = base + 4ll * (it * ADAN_THREADS + threadIdx.x);
        if (i >= numel) break;
        if (vec && i + 4 <= numel) {
            float4 p = *reinterpret_cast<float4 *>(P + i), g = *reinterpret_cast<const float4 *>(G + i);
            float4 m = *reinterpret_cast<float4 *>(M + i), v = *reinterpret_cast<float4 *>(V + i);
            float4 d = *reinterpret_cast<float4 *>(D + i), n = *reinterpret_cast<float4 *>(Nn + i);
            adan_element(p.x, g.x, m.x, v.x, d.x, n.x, s);
            adan_element(p.y, g.y, m.y, v.y, d.y, n.y, s);
            adan_element(p.z, g.z, m.z, v.z, d.z, n.z, s);
            adan_element(p.w, g.w, m.w, v.w, d.w, n.w, s);
            *reinterpret_cast<float4 *>(P + i) = p; *reinterpret_cast<float4 *>(M + i) = m;
            *reinterpret_cast<float4 *>(V + i) = v; *reinterpret_cast<float4 *>(D + i) = d;
            *reinterpret_cast<float4 *>(Nn + i) = n;
            if (write_g) *reinterpret_cast<float4 *>(G + i) = g;
        } else {
            for (long long j = i; j < min(i + 4, numel); ++j) {
                float p = P[j], g = G[j], m = M[j], v = V[j], d = D[j], n = Nn[j];
                adan_element(p, g, m, v, d, n, s);
                P[j] = p; M[j] = m; V[j] = v; D[j] = d; Nn[j] = n;
                if (write_g) G[j] = g;
            }
        }
    }
    if (DEV) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&st->ticket, 1) == (int)gridDim.x - 1) {     // last block of the launch: advance the step
                st->ticket = 0;
                st->step = st->step + 1;
                __threadfence();
            }
        }
    }
}

static int adan_launch(int n_tensors, const long long *numel, void *const *params, void *const *grads, void *const *exp_avgs,
                       void *const *exp_avg_sqs, void *const *exp_avg_diffs, void *const *neg_pre_grads,
                       const AdanScalars *hs, AdanDeviceState *st, int no_prox, cudaStream_t stream) {
    if (n_tensors < 0 || (n_tensors > 0 && (!numel || !params || !grads || !exp_avgs || !exp_avg_sqs || !exp_avg_diffs || !neg_pre_grads)))
        return GI2D_ERR_INVALID;
    if (st && n_tensors > ADAN_MAX_TENSORS) return GI2D_ERR_INVALID;     // the device step counter advances once per launch
    for (int first = 0; first < n_tensors; first += ADAN_MAX_TENSORS) {
        AdanTable t;
        t.count = 0;
        int blocks = 0;
        for (int i = first; i < std::min(n_tensors, first + ADAN_MAX_TENSORS); ++i) {
            if (numel[i] < 0) return GI2D_ERR_INVALID;
            if (numel[i] == 0) continue;
            if (!params[i] || !grads[i] || !exp_avgs[i] || !exp_avg_sqs[i] || !exp_avg_diffs[i] || !neg_pre_grads[i]) return GI2D_ERR_INVALID;
            const int k = t.count++;
            t.p[k] = (float *)params[i]; t.g[k] = (float *)grads[i]; t.m[k] = (float *)exp_avgs[i];
            t.v[k] = (float *)exp_avg_sqs[i]; t.d[k] = (float *)exp_avg_diffs[i]; t.n[k] = (float *)neg_pre_grads[i];
            t.numel[k] = numel[i];
            const long long nb = (numel[i] + ADAN_CHUNK - 1) / ADAN_CHUNK;
            if (nb + blocks > 0x7fffffffll) return GI2D_ERR_INVALID;
            blocks += (int)nb;
            t.blk_end[k] = blocks;
        }
        if (!t.count) continue;
        if (st) adan_kernel<true><<<blocks, ADAN_THREADS, 0, stream>>>(t, AdanScalars{}, st, no_prox);
        else adan_kernel<false><<<blocks, ADAN_THREADS, 0, stream>>>(t, *hs, nullptr, no_prox);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return GI2D_ERR_CUDA_BASE + (int)e;
    }
    return 0;
}

}  // namespace gi2d

using namespace gi2d;

extern "C" {

int gi2d_adan_multi_tensor(int n_tensors, const long long *numel, void *const *params, void *const *grads,
                           void *const *exp_avgs, void *const *exp_avg_sqs, void *const *exp_avg_diffs,
                           void *const *neg_pre_grads, double beta1, double beta2, double beta3, double bias_correction1,
                           double bias_correction2, double bias_correction3_sqrt, double lr, double weight_decay, double eps,
                           int no_prox, double clip_global_grad_norm, int update_neg_pre_grad, cudaStream_t stream) {
    if (!(bias_correction1 != 0.0) || !(bias_correction2 != 0.0) || !(bias_correction3_sqrt != 0.0)) return GI2D_ERR_INVALID;
    const AdanScalars s = adan_scalars(beta1, beta2, beta3, bias_correction1, bias_correction2, bias_correction3_sqrt, lr,
                                       weight_decay, eps, no_prox ? 1 : 0, clip_global_grad_norm, update_neg_pre_grad ? 1 : 0);
    return adan_launch(n_tensors, numel, params, grads, exp_avgs, exp_avg_sqs, exp_avg_diffs, neg_pre_grads, &s, nullptr,
                       no_prox ? 1 : 0, stream);
}

size_t gi2d_adan_state_bytes(void) { return sizeof(AdanDeviceState); }

int gi2d_adan_state_init(void *state_host, double beta1, double beta2, double beta3, double lr, double weight_decay, double eps,
                         double clip_global_grad_norm, int step) {
    if (!state_host || step < 0) return GI2D_ERR_INVALID;
    AdanDeviceState *s = (AdanDeviceState *)state_host;
    s->beta1 = beta1; s->beta2 = beta2; s->beta3 = beta3; s->lr = lr; s->weight_decay = weight_decay; s->eps = eps;
    s->clip = clip_global_grad_norm; s->step = step; s->ticket = 0;
    return 0;
}

int gi2d_adan_multi_tensor_graph(int n_tensors, const long long *numel, void *const *params, void *const *grads,
                                 void *const *exp_avgs, void *const *exp_avg_sqs, void *const *exp_avg_diffs,
                                 void *const *neg_pre_grads, void *state_dev, int no_prox, cudaStream_t stream) {
    if (!state_dev) return GI2D_ERR_INVALID;
    return adan_launch(n_tensors, numel, params, grads, exp_avgs, exp_avg_sqs, exp_avg_diffs, neg_pre_grads, nullptr,
                       (AdanDeviceState *)state_dev, no_prox ? 1 : 0, stream);
}

}  // extern "C"
