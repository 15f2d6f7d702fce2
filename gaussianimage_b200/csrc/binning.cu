// binning.cu -- tile-intersection counting, prefix sum, emission, onesweep radix sort and tile bin edges.
//
// Replaces, for the 2D path, the reference's torch.cumsum + .item() (gsplat/gsplat/utils.py:116-118),
// map_gaussian_to_intersects (csrc/forward.cu:100-136), torch.sort + torch.gather on 64-bit keys
// (utils.py:164-165) and get_tile_bin_edges (forward.cu:141-163), including their torch::zeros fills
// (bindings.cu:289-296, 319-321).
//
// Two shapes of the same kernels:
//   * staged, reference-typed entry points (gi2d_scan_tiles / gi2d_emit_isects / gi2d_sort_isects / gi2d_tile_bins):
//     int64 keys `(tile << 32) | depth_bits`, int32 values, sizes known on the host -- what the public
//     gsplat.utils functions return.
//   * the fused, sync-free chain used by rasterize_gaussians_sum (gi2d_bin_fused): compact uint32 tile keys
//     (depths are identically 0 in the 2D path, quirk Q2, so only ceil(log2(TX*TY)) key bits are sorted), the
//     intersection count stays on the device, grids are sized from a host-side capacity, and the last kernel also
//     packs a 48-byte per-intersection record so that the rasterizers can stage a tile's batch with one bulk TMA copy.
//
// The sort is an LSD onesweep radix sort (8-bit digits): one pass = one kernel that ranks a block's keys
// (warp match_any multisplit, stable), resolves the block's global digit offsets by decoupled look-back over the
// preceding blocks' digit counts, and scatters.  Stability + ascending-gid emission order gives the per-tile
// ascending-gid order the reference gets from cub (quirk Q2).  All of this is HBM/L2-bound integer work.
#include "gi2d_common.cuh"
#include "gi2d.h"
#include <cooperative_groups.h>
#include <mutex>
#include <stdlib.h>

namespace gi2d {

constexpr int SB = 256;           // threads per block everywhere in this file
constexpr int RADIX = 256;
constexpr uint32_t FLAG_AGG = 1u << 30, FLAG_INCL = 2u << 30, VAL_MASK = (1u << 30) - 1;
constexpr int MAX_PASSES = 8;

// control block (uint32 words), zeroed by one memset per call
enum { CTL_N = 0, CTL_OVERFLOW = 1, CTL_SCAN_TICKET = 2, CTL_SORT_TICKET = 3 /* ..10 */, CTL_HIST = 16 };
constexpr int CTL_WORDS = CTL_HIST + MAX_PASSES * RADIX;

struct PassPlan {
    int npass;
    int shift[MAX_PASSES];
    int bits[MAX_PASSES];
};

__device__ __forceinline__ uint32_t ld_volatile(const uint32_t *p) { return *reinterpret_cast<const volatile uint32_t *>(p); }
__device__ __forceinline__ void st_volatile(uint32_t *p, uint32_t v) { *reinterpret_cast<volatile uint32_t *>(p) = v; }

__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// Block-wide inclusive scan of one int per thread (NT threads). Returns the inclusive value; *block_total = sum.
template <int NT>
__device__ __forceinline__ int block_incl_scan(int v, int *s_warp /*NT/32*/, int *block_total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = warp_incl_scan(v, lane);
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = lane < NT / 32 ? s_warp[lane] : 0;
        w = warp_incl_scan(w, lane);
        if (lane < NT / 32) s_warp[lane] = w;
    }
    __syncthreads();
    if (warp > 0) incl += s_warp[warp - 1];
    *block_total = s_warp[NT / 32 - 1];
    return incl;
}

// Decoupled look-back (single-pass chained scan): executed by warp 0 of block `vbid`; returns the sum of
// the totals of blocks [0, vbid).  state[] words are FLAG|value (30-bit values), zero = not yet published.
__device__ __forceinline__ int lookback_exclusive(uint32_t *state, int vbid, int block_total) {
    const int lane = threadIdx.x & 31;
    if (lane == 0) st_volatile(state + vbid, (vbid == 0 ? FLAG_INCL : FLAG_AGG) | (uint32_t)block_total);
    if (vbid == 0) return 0;
    int excl = 0;
    int look = vbid - 1;
    while (true) {
        const int idx = look - lane;
        uint32_t w = idx >= 0 ? ld_volatile(state + idx) : FLAG_INCL;
        while (__any_sync(0xffffffffu, (w >> 30) == 0)) w = idx >= 0 ? ld_volatile(state + idx) : FLAG_INCL;
        const unsigned incl_mask = __ballot_sync(0xffffffffu, (w & FLAG_INCL) != 0);
        const int val = (int)(w & VAL_MASK);
        if (incl_mask) {
            const int first = __ffs(incl_mask) - 1;
            excl += warp_sum(lane <= first ? val : 0);
            break;
        }
        excl += warp_sum(val);
        look -= 32;
    }
    if (lane == 0) st_volatile(state + vbid, FLAG_INCL | (uint32_t)(excl + block_total));
    return excl;
}

// ---------------------------------------------------------------------------------------------------------------
// Emission (reference forward.cu:100-136), warp-cooperative: the 32 Gaussians of a warp own one contiguous slot range;
// lanes walk that range together so stores are coalesced and a Gaussian covering thousands of tiles is spread
// over the warp.  Slot k of Gaussian g is tile (y0 + k / w, x0 + k % w): tile-row-major like the reference's loop.
// Slots beyond the recomputed bbox area (inconsistent num_tiles_hit) keep the reference's zero fill.
// FMT 0: uint32 tile key (+ per-pass digit histogram in shared memory). FMT 1: int64 (tile<<32)|depth_bits.
template <int FMT>
__device__ __forceinline__ void warp_emit(int cnt, int off, int x0, int w, int y0, int area, int gid, int dbits,
                                          int TX, int capacity, void *keys_out, int *vals_out, uint32_t *s_hist,
                                          const PassPlan &plan) {
    const int lane = threadIdx.x & 31;
    const int incl = warp_incl_scan(cnt, lane);
    const int excl = incl - cnt;
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    const int base = __shfl_sync(0xffffffffu, off, 0);
    for (int k0 = 0; k0 < total; k0 += 32) {
        const int k = k0 + lane;
        const bool active = k < total;
        int lo = 0, hi = 31;
#pragma unroll
        for (int it = 0; it < 5; ++it) {      // smallest lane whose inclusive count exceeds k
            const int mid = (lo + hi) >> 1;
            const int v = __shfl_sync(0xffffffffu, incl, mid);
            if (v > k) hi = mid; else lo = mid + 1;
        }
        const int g = min(lo, 31);
        const int m = k - __shfl_sync(0xffffffffu, excl, g);
        const int ga = __shfl_sync(0xffffffffu, area, g), gw = __shfl_sync(0xffffffffu, w, g);
        const int gx0 = __shfl_sync(0xffffffffu, x0, g), gy0 = __shfl_sync(0xffffffffu, y0, g);
        const int ggid = __shfl_sync(0xffffffffu, gid, g), gdb = __shfl_sync(0xffffffffu, dbits, g);
        if (!active) continue;
        int tile = 0, id = 0, db = 0;
        if (m < ga) {
            const int row = m / gw;
            tile = (gy0 + row) * TX + gx0 + (m - row * gw);
            id = ggid;
            db = gdb;
        }
        const long long pos = (long long)base + k;
        if (pos < capacity) {
            if (FMT == 0) {
                reinterpret_cast<uint32_t *>(keys_out)[pos] = (uint32_t)tile;
                for (int p = 0; p < plan.npass; ++p)
                    atomicAdd(&s_hist[p * RADIX + (((uint32_t)tile >> plan.shift[p]) & ((1u << plan.bits[p]) - 1))], 1u);
            } else {
                reinterpret_cast<long long *>(keys_out)[pos] = ((long long)tile << 32) | (long long)db;   // sign-extends like forward.cu:125
            }
            vals_out[pos] = id;
        }
    }
}

// Tile bbox of Gaussian i, optionally clipped to the tile-row window [wy0, wy1) of a row-sharded render.
__device__ __forceinline__ void gaussian_bbox(const float *__restrict__ xys, const int *__restrict__ radii, int i, int TX,
                                              int TY, int &x0, int &w, int &y0, int &area, int wy0 = 0, int wy1 = 0x7fffffff) {
    x0 = w = y0 = area = 0;
    const int r = __ldg(radii + i);
    if (r <= 0) return;                                   // forward.cu:113-114
    int x1, y1;
    tile_bbox(__ldg(xys + 2 * (size_t)i), __ldg(xys + 2 * (size_t)i + 1), (float)r, TX, TY, x0, x1, y0, y1);
    y0 = max(y0, wy0);
    y1 = min(y1, wy1);
    w = x1 - x0;
    area = max(0, w) * max(0, y1 - y0);
    if (w <= 0) { w = 1; area = 0; }
}

// Optional extras of the fused chain (gi2d_bin_fused_ex): a STICKY status pair that only the host clears -- [0] = largest
// intersection count seen, [1] |= overflow -- so that a caller replaying a captured graph can poll now and then instead of
// after every launch; and the (N,12) gradient rows of the backward pass zero-filled by the kernel that walks every Gaussian
// anyway (instead of a memset node in front of the backward rasterizer).
struct FusedExtras {
    int *sticky;
    float4 *zero_g12;
};
__device__ __forceinline__ void publish_sticky(int *sticky, int total, int over) {
    if (!sticky) return;
    atomicMax(sticky, total);
    if (over) atomicOr(sticky + 1, 1);
}
__device__ __forceinline__ void zero_grad_row(float4 *zero_g12, int i, int N) {
    if (zero_g12 && i < N) {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        zero_g12[3 * (size_t)i] = z; zero_g12[3 * (size_t)i + 1] = z; zero_g12[3 * (size_t)i + 2] = z;
    }
}

// Inclusive int32 scan of num_tiles_hit (utils.py:116) -> cum (may be null) and the total in ctl[CTL_N].
// FUSED additionally emits the compact (tile, gid) pairs and the digit histograms, presets tile_bins to
// (INT_MAX, 0) for the atomicMin/Max of the last sort pass and writes status_out = {n_isect, overflow}.
// One Gaussian per thread, SCAN_T threads per block: 70k Gaussians are 69 blocks, so the look-back chain is short.
template <bool FUSED, int SCAN_T>
__global__ void __launch_bounds__(SCAN_T) scan_kernel(int N, const int *__restrict__ num_tiles_hit, int *__restrict__ cum,
                                                      uint32_t *ctl, uint32_t *scan_state, const float *__restrict__ xys,
                                                      const int *__restrict__ radii, int TX, int TY, int capacity,
                                                      uint32_t *keys_out, int *vals_out, PassPlan plan,
                                                      int2 *tile_bins_init, int n_bins, int *status_out, volatile int *status_host, int use_ticket, int dbg, int wy0, int wy1,
                                                      FusedExtras ex = FusedExtras{nullptr, nullptr}) {
    __shared__ int s_warp[SCAN_T / 32];
    __shared__ int s_bcast[2];
    __shared__ uint32_t s_hist[FUSED ? MAX_PASSES * RADIX : 1];
    const int tid = threadIdx.x;
    if (FUSED) {
        for (int k = blockIdx.x * SCAN_T + tid; k < n_bins; k += gridDim.x * SCAN_T) tile_bins_init[k] = make_int2(0x7fffffff, 0);
        for (int k = tid; k < plan.npass * RADIX; k += SCAN_T) s_hist[k] = 0;
    }
    if (use_ticket && tid == 0) s_bcast[0] = (int)atomicAdd(ctl + CTL_SCAN_TICKET, 1u);   // look-back order = start order
    __syncthreads();
    const int vbid = use_ticket ? s_bcast[0] : (int)blockIdx.x;
    const int i = vbid * SCAN_T + tid;
    int cnt = i < N ? __ldg(num_tiles_hit + i) : 0;
    if (FUSED) zero_grad_row(ex.zero_g12, i, N);
    int x0 = 0, w = 1, y0 = 0, area = 0;
    const bool windowed = FUSED && (wy0 > 0 || wy1 < TY);
    if (windowed && cnt > 0) {              // row-sharded render: only the tile rows [wy0, wy1) of this rank count
        gaussian_bbox(xys, radii, i, TX, TY, x0, w, y0, area, wy0, wy1);
        cnt = area;
    }
    int block_total;
    const int incl = block_incl_scan<SCAN_T>(cnt, s_warp, &block_total);
    if (tid < 32) {
        const int e = lookback_exclusive(scan_state, vbid, block_total);
        if (tid == 0) s_bcast[1] = e;
    }
    __syncthreads();
    const int block_excl = s_bcast[1];
    if (i < N && cum) cum[i] = block_excl + incl;
    if (vbid == (int)gridDim.x - 1 && tid == 0) {
        const int total = block_excl + block_total;
        ctl[CTL_N] = (uint32_t)total;
        if (FUSED) {
            const int over = total > capacity ? 1 : 0;
            ctl[CTL_OVERFLOW] = (uint32_t)over;
            if (status_out) { status_out[0] = total; status_out[1] = over; }
            publish_sticky(ex.sticky, total, over);
            if (status_host) {          // mapped pinned host memory: the host polls [0] (preset to -1) instead of synchronising
                status_host[1] = over;
                __threadfence_system();
                status_host[0] = total;
            }
        }
    }
    if (FUSED && !(dbg & 1)) {
        if (!windowed && i < N && cnt > 0) gaussian_bbox(xys, radii, i, TX, TY, x0, w, y0, area);
        warp_emit<0>(cnt, block_excl + incl - cnt, x0, w, y0, area, i, 0, TX, capacity, keys_out, vals_out, s_hist, plan);
        __syncthreads();
        if (!(dbg & 2))
        for (int k = tid; k < plan.npass * RADIX; k += SCAN_T) {
            const uint32_t c = s_hist[k];
            if (c) atomicAdd(ctl + CTL_HIST + k, c);
        }
    }
}

// map_gaussian_to_intersects with reference-typed outputs (int64 keys, int32 ids), offsets from cum_tiles_hit.
__global__ void __launch_bounds__(SB) emit64_kernel(int N, const float *__restrict__ xys, const float *__restrict__ depths,
                                                    const int *__restrict__ radii, const int *__restrict__ cum, int TX,
                                                    int TY, int n_isect, long long *isect_ids, int *gaussian_ids) {
    const int i = blockIdx.x * SB + threadIdx.x;
    int cnt = 0, off = 0, x0 = 0, w = 1, y0 = 0, area = 0, db = 0;
    if (i < N) {
        const int hi = __ldg(cum + i);
        off = i == 0 ? 0 : __ldg(cum + i - 1);
        cnt = max(0, hi - off);
        gaussian_bbox(xys, radii, i, TX, TY, x0, w, y0, area);
        db = __float_as_int(__ldg(depths + i));
    }
    PassPlan none;
    none.npass = 0;
    warp_emit<1>(cnt, off, x0, w, y0, area, i, db, TX, n_isect, isect_ids, gaussian_ids, nullptr, none);
}

// ---------------------------------------------------------------------------------------------------------------
// Histogram of all passes' digits at once (staged path; the fused path gets it from the emit kernel).
template <typename KeyT>
__global__ void __launch_bounds__(SB) hist_kernel(int n, const KeyT *__restrict__ keys, uint32_t *ctl, PassPlan plan, KeyT flip) {
    __shared__ uint32_t s_hist[MAX_PASSES * RADIX];
    for (int k = threadIdx.x; k < plan.npass * RADIX; k += SB) s_hist[k] = 0;
    __syncthreads();
    for (int i = blockIdx.x * SB + threadIdx.x; i < n; i += gridDim.x * SB) {
        const KeyT key = keys[i] ^ flip;
        for (int p = 0; p < plan.npass; ++p)
            atomicAdd(&s_hist[p * RADIX + (int)((key >> plan.shift[p]) & (KeyT)((1u << plan.bits[p]) - 1))], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < plan.npass * RADIX; k += SB) {
        const uint32_t c = s_hist[k];
        if (c) atomicAdd(ctl + CTL_HIST + k, c);
    }
}

// Exclusive scan of one value per thread over the SB threads of the block (s_tmp: SB/32 words; contains barriers).
__device__ __forceinline__ uint32_t block_excl_scan_u32(uint32_t v, uint32_t *s_tmp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t incl = (uint32_t)warp_incl_scan((int)v, lane);
    __syncthreads();                       // s_tmp may still be read from a previous call
    if (lane == 31) s_tmp[warp] = incl;
    __syncthreads();
    uint32_t wprefix = 0;
#pragma unroll
    for (int w = 0; w < SB / 32; ++w) wprefix += (w < warp) ? s_tmp[w] : 0;
    return wprefix + incl - v;
}

struct PackSrc {          // sources of the packed records written by the LAST pass of the fused chain
    const float *xys, *conics, *colors, *opac;
    float4 *packed;
    int *tile_bins;       // (n_bins,2), pre-set to (INT_MAX, 0) by the scan kernel
    int n_bins;
};

// One onesweep pass.  Block tile = SB*ITEMS keys, warp-striped so that (item, lane) order == memory order.
//   1. rank the keys of the tile by digit (warp match_any multisplit, stable);
//   2. thread d owns digit d: block count, decoupled look-back over the preceding blocks (8 loads in flight per
//      step), global digit base from the pass histogram;
//   3. stage (key, value) in shared memory in block-sorted order, then write them out with consecutive threads on
//      consecutive addresses of each digit run.
// LAST (fused chain, uint32 tile keys, most significant digit): the sorted keys themselves are not written; instead
// the per-tile bin edges are derived from neighbouring keys (atomicMin/Max only matter at block/digit seams) and the
// 48-byte packed record of each intersection is gathered and written at its sorted position.
template <typename KeyT, int ITEMS, bool LAST>
__global__ void __launch_bounds__(SB) onesweep_kernel(const uint32_t *__restrict__ d_n, int n_host, int capacity,
                                                      const KeyT *__restrict__ keys_in, const int *__restrict__ vals_in,
                                                      KeyT *__restrict__ keys_out, int *__restrict__ vals_out,
                                                      uint32_t *ctl, uint32_t *lookback /* [blocks][RADIX] */, int pass,
                                                      int shift, int bits, KeyT flip, int use_ticket, PackSrc ps) {
    constexpr int WARPS = SB / 32, TILE = SB * ITEMS;
    __shared__ uint32_t s_cnt[WARPS][RADIX];     // per-warp digit counts, then per-warp exclusive offsets
    __shared__ uint32_t s_off[RADIX];            // block-exclusive offset of each digit's run in the staged order
    __shared__ uint32_t s_base[RADIX];           // global position of staged index 0 of each digit's run, minus s_off
    __shared__ uint32_t s_tmp[WARPS];
    __shared__ KeyT s_keys[TILE];
    __shared__ int s_vals[TILE];
    __shared__ int s_vbid;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = d_n ? min((int)*d_n, capacity) : n_host;
    if (use_ticket && tid == 0) s_vbid = (int)atomicAdd(ctl + CTL_SORT_TICKET + pass, 1u);
#pragma unroll
    for (int w = 0; w < WARPS; ++w) s_cnt[w][tid] = 0;
    __syncthreads();
    const int vbid = use_ticket ? s_vbid : (int)blockIdx.x;
    const int tile_start = vbid * TILE;
    if (tile_start >= n) return;                 // whole block past the end: nobody looks back at it
    const int nvalid = min(TILE, n - tile_start);
    const uint32_t mask = (1u << bits) - 1;

    KeyT key[ITEMS];
    int val[ITEMS];
    uint32_t rank[ITEMS];
    const int wofs = warp * 32 * ITEMS + lane;   // index inside the tile of this thread's item 0
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
        const int li = wofs + it * 32;
        const bool ok = li < nvalid;
        key[it] = ok ? keys_in[tile_start + li] : (KeyT)0;
        val[it] = ok ? vals_in[tile_start + li] : 0;
    }
    const unsigned lt_mask = (1u << lane) - 1;
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
        const bool ok = wofs + it * 32 < nvalid;
        const uint32_t digit = ok ? (uint32_t)(((key[it] ^ flip) >> shift) & (KeyT)mask) : RADIX;   // RADIX = "no key"
        const unsigned peers = __match_any_sync(0xffffffffu, digit);
        const uint32_t before = ok ? s_cnt[warp][digit] : 0;
        __syncwarp();
        if (ok && (peers & lt_mask) == 0) s_cnt[warp][digit] = before + __popc(peers);   // lowest lane of the group
        __syncwarp();
        rank[it] = before + __popc(peers & lt_mask);
    }
    __syncthreads();
    {   // thread d owns digit d
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            const uint32_t c = s_cnt[w][tid];
            s_cnt[w][tid] = run;
            run += c;
        }
        uint32_t *my = lookback + (size_t)vbid * RADIX + tid;
        st_volatile(my, (vbid == 0 ? FLAG_INCL : FLAG_AGG) | run);
        const uint32_t h = ctl[CTL_HIST + pass * RADIX + tid];
        const uint32_t off = block_excl_scan_u32(run, s_tmp);      // run of digit d starts here in the staged order
        const uint32_t gbase = block_excl_scan_u32(h, s_tmp);      // exclusive scan of the pass histogram
        uint32_t excl = 0;
        int pb = vbid - 1;
        while (pb >= 0) {
            constexpr int LB = 16;                  // predecessor words in flight per step
            uint32_t w[LB];
#pragma unroll
            for (int j = 0; j < LB; ++j) w[j] = (pb - j >= 0) ? ld_volatile(lookback + (size_t)(pb - j) * RADIX + tid) : FLAG_INCL;
            int j = 0;
            bool done = false;
#pragma unroll
            for (; j < LB; ++j) {
                if ((w[j] >> 30) == 0) break;      // not published yet: re-read from here
                excl += w[j] & VAL_MASK;
                if (w[j] & FLAG_INCL) { done = true; break; }
            }
            if (done) break;
            pb -= j;
        }
        if (vbid > 0) st_volatile(my, FLAG_INCL | (excl + run));
        s_off[tid] = off;
        s_base[tid] = gbase + excl - off;
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
        if (wofs + it * 32 < nvalid) {
            const uint32_t digit = (uint32_t)(((key[it] ^ flip) >> shift) & (KeyT)mask);
            const uint32_t li = s_off[digit] + s_cnt[warp][digit] + rank[it];
            s_keys[li] = key[it];
            s_vals[li] = val[it];
        }
    }
    __syncthreads();
    if (!LAST) {
        for (int li = tid; li < nvalid; li += SB) {
            const KeyT k = s_keys[li];
            const uint32_t digit = (uint32_t)(((k ^ flip) >> shift) & (KeyT)mask);
            const uint32_t pos = s_base[digit] + li;
            keys_out[pos] = k;
            vals_out[pos] = s_vals[li];
        }
    } else {
        for (int li = tid; li < nvalid; li += SB) {      // bin edges: atomics only matter at block / digit seams
            const KeyT k = s_keys[li];
            const uint32_t digit = (uint32_t)(((k ^ flip) >> shift) & (KeyT)mask);
            const uint32_t pos = s_base[digit] + li;
            const int cur = (int)k;
            if (cur < ps.n_bins) {
                const bool first = li == 0 || (uint32_t)(((s_keys[li - 1] ^ flip) >> shift) & (KeyT)mask) != digit;
                const bool last = li == nvalid - 1 || (uint32_t)(((s_keys[li + 1] ^ flip) >> shift) & (KeyT)mask) != digit;
                if (first || (int)s_keys[li - 1] != cur) atomicMin(ps.tile_bins + 2 * cur, (int)pos);
                if (last || (int)s_keys[li + 1] != cur) atomicMax(ps.tile_bins + 2 * cur + 1, (int)pos + 1);
            }
        }
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {             // packed records: all gathers of the thread in flight together
            const int li = tid + it * SB;
            if (li < nvalid) {
                const KeyT k = s_keys[li];
                const int v = s_vals[li];
                const uint32_t digit = (uint32_t)(((k ^ flip) >> shift) & (KeyT)mask);
                const uint32_t pos = s_base[digit] + li;
                vals_out[pos] = v;
                const size_t g = (size_t)v;
                const float x = __ldg(ps.xys + 2 * g), y = __ldg(ps.xys + 2 * g + 1);
                const float cx = __ldg(ps.conics + 3 * g), cy = __ldg(ps.conics + 3 * g + 1), cz = __ldg(ps.conics + 3 * g + 2);
                const float r = __ldg(ps.colors + 3 * g), gg = __ldg(ps.colors + 3 * g + 1), b = __ldg(ps.colors + 3 * g + 2);
                const float o = __ldg(ps.opac + g);
                __stcs(ps.packed + 3 * (size_t)pos + 0, make_float4(x, y, o, cx));
                __stcs(ps.packed + 3 * (size_t)pos + 1, make_float4(cy, cz, r, gg));
                const float2 e = record_extent(cx, cy, cz, o);
                __stcs(ps.packed + 3 * (size_t)pos + 2, make_float4(b, __int_as_float(v), e.x, e.y));
            }
        }
    }
}

// get_tile_bin_edges (forward.cu:141-163) on sorted keys; tile = key >> TSHIFT; rows >= n_bins are skipped (Q4).
template <typename KeyT, int TSHIFT>
__global__ void __launch_bounds__(SB) tile_bins_kernel(int n, const KeyT *__restrict__ keys_sorted, int n_bins, int *tile_bins) {
    for (int i = blockIdx.x * SB + threadIdx.x; i < n; i += gridDim.x * SB) {
        const int cur = (int)(keys_sorted[i] >> TSHIFT);
        const bool cur_ok = cur >= 0 && cur < n_bins;
        if (i == 0 && cur_ok) tile_bins[2 * cur] = 0;
        if (i == n - 1 && cur_ok) tile_bins[2 * cur + 1] = n;
        if (i == 0) continue;
        const int prev = (int)(keys_sorted[i - 1] >> TSHIFT);
        if (prev != cur) {
            if (prev >= 0 && prev < n_bins) tile_bins[2 * prev + 1] = i;
            if (cur_ok) tile_bins[2 * cur] = i;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// One-kernel binning for images of at most 4096 tiles (every GaussianImage fitting configuration: 768x512 has 1536
// tiles, 1020x678 has 2752): counting sort by tile, stable in Gaussian id, in ONE cooperative launch whose blocks meet
// at two grid-wide barriers.  Replaces memset + scan + two onesweep passes; needs no zero-initialised memory.
//   phase 1  a block owns NT consecutive Gaussians, a warp 32 of them; the warp walks its (Gaussian, tile) pairs in
//            emission order (forward.cu:115-134: Gaussian-major, tile-row-major) and counts them per tile in its own
//            row of byte counters (a Gaussian hits a tile at most once, so a count is <= 32);
//   phase 2  per tile: the counts of the warps become exclusive prefixes (bytes inside a group of 8 warps, 16-bit
//            across the groups), the block total goes to row `block` of a global (blocks x T) matrix;  grid barrier;
//   phase 3a every warp of the grid takes 4 tile columns of that matrix and scans them over the blocks, in place
//            (O(blocks x T) traffic in total);  grid barrier;
//   phase 3b per block: exclusive scan of the tile totals in tile order = first slot of every tile; plus the block's
//            row of the matrix = first slot of this block inside every tile; block 0 writes tile_bins and status;
//   phase 4  the warp walks its pairs again in the same order and writes gid and the 48-byte packed record at
//            slot = tile base + preceding blocks + preceding warps + rank inside the warp (match_any).
// Stability in (block, warp, emission order) = ascending Gaussian id per tile, the order the reference gets (Q2).
// Grid barriers cost microseconds only while there is at most one block per SM, so the block shape NT (256/512/1024)
// is picked as the smallest that keeps the grid within the SM count.
constexpr int SMALL_T = 1024;                 // largest block shape: threads per block = Gaussians per block

struct SmallArgs {
    int N, TX, TY, T, capacity;
    const float *xys;
    const int *radii, *nth;
    const float *conics, *colors, *opac;
    uint32_t *mat;            // (gridDim.x, Tpad): block totals per tile, then (in place) the totals of the preceding blocks
    uint32_t *tot;            // (Tpad): totals over all blocks
    uint32_t *ctl;            // zeroed control block: [CTL_SCAN_TICKET] is the grid-barrier counter
    int *gids_sorted, *tile_bins;
    float4 *packed;
    int *status_out;
    volatile int *status_host;
    int dbg;                  // experiments: stop after phase 1 / 2 / 3 (bits 8 / 16 / 32)
    unsigned int *bar;        // {arrival count, generation}: library-owned, zero once, left zero-count by every barrier; null = cg grid.sync
    FusedExtras ex;
};

// Grid-wide barrier of a co-resident (cooperatively launched) grid on two words of ordinary device memory: sense reversal,
// the last block to arrive resets the count and advances the generation.  Self-cleaning, so the words are zeroed once per
// (device, stream) and never again -- no memset node in front of the kernel.
__device__ __forceinline__ void grid_barrier(unsigned int *bar, unsigned int nb) {
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int g, old, v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(bar + 1) : "memory");   // cannot advance before this block arrives
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(bar) : "memory");
        if (old == nb - 1) {
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(bar), "r"(0u) : "memory");
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar + 1) : "memory");
        } else {
            do {
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar + 1) : "memory");
            } while (v == g);
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
        }
    }
    __syncthreads();
}

// Calls f(tile, gid_or_0, owner_lane, active) for every (Gaussian, tile) pair of the warp, 32 pairs per round, in
// emission order.  All 32 lanes call f in every round (f may use warp-wide primitives).
template <typename F>
__device__ __forceinline__ void warp_pairs(int cnt, int x0, int w, int y0, int TX, int gid, F f) {
    const int lane = threadIdx.x & 31;
    const int incl = warp_incl_scan(cnt, lane);
    const int excl = incl - cnt;
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    for (int k0 = 0; k0 < total; k0 += 32) {
        const int k = k0 + lane;
        int lo = 0, hi = 31;
#pragma unroll
        for (int it = 0; it < 5; ++it) {      // smallest lane whose inclusive count exceeds k
            const int mid = (lo + hi) >> 1;
            const int v = __shfl_sync(0xffffffffu, incl, mid);
            if (v > k) hi = mid; else lo = mid + 1;
        }
        const int g = min(lo, 31);
        const int m = k - __shfl_sync(0xffffffffu, excl, g);
        const int gw = __shfl_sync(0xffffffffu, w, g);
        const int gx0 = __shfl_sync(0xffffffffu, x0, g), gy0 = __shfl_sync(0xffffffffu, y0, g);
        const int ggid = __shfl_sync(0xffffffffu, gid, g);
        const int row = m / gw;
        const int tile = (gy0 + row) * TX + gx0 + (m - row * gw);
        f(tile, ggid, g, k < total);
    }
}

template <int TBITS, int NT>
__global__ void __launch_bounds__(NT, 1) bin_small_kernel(SmallArgs a) {
    constexpr int TP = 1 << TBITS;                         // padded tile count
    constexpr int SMALL_WARPS = NT / 32, NGRP = SMALL_WARPS / 8;
    extern __shared__ __align__(16) unsigned char s_raw[];
    unsigned char (*s_cnt)[TP] = reinterpret_cast<unsigned char (*)[TP]>(s_raw);                       // [32][TP]
    unsigned short (*s_grp)[TP] = reinterpret_cast<unsigned short (*)[TP]>(s_raw + SMALL_WARPS * TP);  // [NGRP][TP]
    int *s_base = reinterpret_cast<int *>(s_raw + SMALL_WARPS * TP + NGRP * TP * 2);                      // [TP]
    __shared__ int s_warp[NT / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = gridDim.x, b = blockIdx.x;

    {   // zero the byte counters (16 bytes per store)
        uint4 *z = reinterpret_cast<uint4 *>(s_raw);
        for (int k = tid; k < SMALL_WARPS * TP / 16; k += NT) z[k] = make_uint4(0, 0, 0, 0);
    }
    const int i = b * NT + tid;
    int cnt = 0, x0 = 0, w = 1, y0 = 0, area = 0;
    // every load of the kernel's input is issued here, independent of the others; the record fields are only consumed
    // in phase 4, their latency hides behind phases 1-3
    float rx = 0.f, ry = 0.f, ro = 0.f, rcx = 0.f, rcy = 0.f, rcz = 0.f, rr = 0.f, rg = 0.f, rb = 0.f;
    zero_grad_row(a.ex.zero_g12, i, a.N);
    if (i < a.N) {
        const size_t g = (size_t)i;
        const int nth = __ldg(a.nth + i), rad = __ldg(a.radii + i);
        rx = __ldg(a.xys + 2 * g); ry = __ldg(a.xys + 2 * g + 1);
        rcx = __ldg(a.conics + 3 * g); rcy = __ldg(a.conics + 3 * g + 1); rcz = __ldg(a.conics + 3 * g + 2);
        rr = __ldg(a.colors + 3 * g); rg = __ldg(a.colors + 3 * g + 1); rb = __ldg(a.colors + 3 * g + 2);
        ro = __ldg(a.opac + g);
        if (nth > 0 && rad > 0) {                          // forward.cu:113-114
            int x1, y1;
            tile_bbox(rx, ry, (float)rad, a.TX, a.TY, x0, x1, y0, y1);
            w = x1 - x0;
            area = max(0, w) * max(0, y1 - y0);
            if (w <= 0) { w = 1; area = 0; }
            cnt = min(nth, area);                          // the projections always give num_tiles_hit == area
        }
    }
    __syncthreads();
    const unsigned lt = (1u << lane) - 1;
    // ---- phase 1: per-warp, per-tile counts.  The first four rounds (128 pairs of the warp) also remember what phase 4
    // needs -- tile, owner lane and the pair's rank among the warp's pairs of that tile -- so that phase 4 does not walk
    // and match the pairs a second time: one word per lane and round, bit 31 = valid.
    uint32_t memo0 = 0, memo1 = 0, memo2 = 0, memo3 = 0;
    {
        int round = 0;
        warp_pairs(cnt, x0, w, y0, a.TX, i, [&](int tile, int, int owner, bool active) {
            const unsigned peers = __match_any_sync(0xffffffffu, active ? tile : -1 - lane);
            const int rk = __popc(peers & lt);
            const int before = active ? (int)s_cnt[warp][tile] : 0;
            __syncwarp();
            if (active && rk == 0) s_cnt[warp][tile] = (unsigned char)(before + __popc(peers));
            __syncwarp();
            const uint32_t v = active ? (0x80000000u | (uint32_t)tile | ((uint32_t)owner << 12) | ((uint32_t)(before + rk) << 17)) : 0u;
            memo0 = round == 0 ? v : memo0; memo1 = round == 1 ? v : memo1;
            memo2 = round == 2 ? v : memo2; memo3 = round == 3 ? v : memo3;
            ++round;
        });
    }
    __syncthreads();
    if (a.dbg & 8) return;
    // ---- phase 2: prefixes over the warps of the block, block totals to the global matrix.  Four adjacent tiles per
    // thread: their byte counters are one 32-bit word per warp row, and the running sums stay packed (a warp adds at most
    // 32 to a tile, so seven rows stay below 256 per byte); the group totals are widened to 16-bit pairs.
    {
        uint32_t *cnt32 = reinterpret_cast<uint32_t *>(s_raw);                         // [SMALL_WARPS][TP / 4]
        for (int wd = tid; wd < TP / 4; wd += NT) {
            uint32_t tot_lo = 0, tot_hi = 0;                                           // tiles {0,1} and {2,3} of the word
#pragma unroll
            for (int g = 0; g < NGRP; ++g) {
                reinterpret_cast<uint2 *>(&s_grp[g][0])[wd] = make_uint2(tot_lo, tot_hi);
                uint32_t run = 0, c = 0;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    c = cnt32[(8 * g + q) * (TP / 4) + wd];
                    cnt32[(8 * g + q) * (TP / 4) + wd] = run;                          // exclusive inside the group: <= 7 * 32
                    if (q < 7) run += c;
                }
                tot_lo += __byte_perm(run, 0, 0x4140) + __byte_perm(c, 0, 0x4140);
                tot_hi += __byte_perm(run, 0, 0x4342) + __byte_perm(c, 0, 0x4342);
            }
            reinterpret_cast<uint4 *>(a.mat + (size_t)b * TP)[wd] = make_uint4(tot_lo & 0xffffu, tot_lo >> 16, tot_hi & 0xffffu, tot_hi >> 16);
        }
    }
    if (a.dbg & 16) return;
    // grid-wide barrier (cooperative launch: all blocks are co-resident, the barrier state belongs to the driver, so
    // this path needs no zero-initialised memory at all)
    if (!a.bar || (a.dbg & 128)) __threadfence();       // own barrier: its release (after the block barrier) is cumulative over the block's writes
    if (a.bar) grid_barrier(a.bar, gridDim.x); else cooperative_groups::this_grid().sync();
    // ---- phase 3a: exclusive scan over the BLOCKS of every tile column, in place.  A warp takes FOUR adjacent columns
    // (16 bytes of every row) and splits the rows into eight contiguous groups, one per quarter-warp row of lanes: each
    // lane sums its rows serially (all loads of a batch in flight at once), three shuffles turn the group sums into group
    // offsets, and the lane writes its rows' exclusive prefixes.  Items are dealt round-robin over the blocks first so
    // that every SM takes part.  The first batch of rows stays in registers; later batches (large grids) are re-read from L2.
    {
        constexpr int RB3 = NT == 1024 ? 16 : 19;            // 8 x 19 rows cover a one-block-per-SM grid (148 SMs) in one batch
        const int gwarp = warp * nb + b, nwarps = nb * SMALL_WARPS;
        const int c4 = lane & 3, rgrp = lane >> 2;
        const int R = (nb + 7) >> 3;
        const int r_lo = rgrp * R, r_hi = min(nb, r_lo + R);
        for (int item = gwarp; item < TP / 4; item += nwarps) {
            uint32_t *col = a.mat + item * 4 + c4;
            uint32_t v[RB3], sum = 0;
#pragma unroll
            for (int u = 0; u < RB3; ++u) v[u] = (r_lo + u < r_hi) ? __ldcg(col + (size_t)(r_lo + u) * TP) : 0u;
#pragma unroll
            for (int u = 0; u < RB3; ++u) sum += v[u];
            for (int r0 = r_lo + RB3; r0 < r_hi; r0 += RB3) {
                uint32_t t[RB3];
#pragma unroll
                for (int u = 0; u < RB3; ++u) t[u] = (r0 + u < r_hi) ? __ldcg(col + (size_t)(r0 + u) * TP) : 0u;
#pragma unroll
                for (int u = 0; u < RB3; ++u) sum += t[u];
            }
            uint32_t incl = sum;
#pragma unroll
            for (int d = 4; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const uint32_t total = __shfl_sync(0xffffffffu, incl, 28 + c4);
            uint32_t run = incl - sum;
#pragma unroll
            for (int u = 0; u < RB3; ++u) {
                if (r_lo + u < r_hi) col[(size_t)(r_lo + u) * TP] = run;
                run += v[u];
            }
            for (int r0 = r_lo + RB3; r0 < r_hi; r0 += RB3) {
                uint32_t t[RB3];
#pragma unroll
                for (int u = 0; u < RB3; ++u) t[u] = (r0 + u < r_hi) ? __ldcg(col + (size_t)(r0 + u) * TP) : 0u;
#pragma unroll
                for (int u = 0; u < RB3; ++u) {
                    if (r0 + u < r_hi) col[(size_t)(r0 + u) * TP] = run;
                    run += t[u];
                }
            }
            if (rgrp == 0) a.tot[item * 4 + c4] = total;
        }
    }
    if (!a.bar || (a.dbg & 128)) __threadfence();
    if (a.bar) grid_barrier(a.bar, gridDim.x); else cooperative_groups::this_grid().sync();
    if (a.dbg & 64) return;
    // ---- phase 3b: first slot of every tile = exclusive scan of the totals in tile order + the preceding blocks
    constexpr int PER = TP / NT;                               // consecutive tiles per thread
    int tot_t[PER], pre_t[PER], mine = 0;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
        tot_t[q] = (int)__ldcg(a.tot + tid * PER + q);
        pre_t[q] = (int)__ldcg(a.mat + (size_t)b * TP + tid * PER + q);
    }
#pragma unroll
    for (int q = 0; q < PER; ++q) mine += tot_t[q];
    int block_total;
    const int incl = block_incl_scan<NT>(mine, s_warp, &block_total);
    {
        int base = incl - mine;
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const int t = tid * PER + q;
            s_base[t] = base + pre_t[q];
            // on overflow (n_isect > capacity, reported through status) the ranges are clipped to what was written
            if (b == 0 && t < a.T)
                reinterpret_cast<int2 *>(a.tile_bins)[t] = tot_t[q] > 0 ? make_int2(min(base, a.capacity), min(base + tot_t[q], a.capacity)) : make_int2(0, 0);
            base += tot_t[q];
        }
    }
    if (b == 0 && tid == 0) {
        const int over = block_total > a.capacity ? 1 : 0;
        a.ctl[CTL_N] = (uint32_t)block_total;
        a.ctl[CTL_OVERFLOW] = (uint32_t)over;
        if (a.status_out) { a.status_out[0] = block_total; a.status_out[1] = over; }
        publish_sticky(a.ex.sticky, block_total, over);
        if (a.status_host) {
            a.status_host[1] = over;
            __threadfence_system();
            a.status_host[0] = block_total;
        }
    }
    __syncthreads();
    if (a.dbg & 32) return;
    // ---- phase 4: scatter gid + packed record to the final slot
    const int grp = warp >> 3;
    const float2 rext = record_extent(rcx, rcy, rcz, ro);      // once per Gaussian; every intersection's record carries it
    auto emit = [&](int pos, int gid, int owner, bool active) {
        const float x = __shfl_sync(0xffffffffu, rx, owner), y = __shfl_sync(0xffffffffu, ry, owner), o = __shfl_sync(0xffffffffu, ro, owner);
        const float cx = __shfl_sync(0xffffffffu, rcx, owner), cy = __shfl_sync(0xffffffffu, rcy, owner), cz = __shfl_sync(0xffffffffu, rcz, owner);
        const float r = __shfl_sync(0xffffffffu, rr, owner), gg = __shfl_sync(0xffffffffu, rg, owner), bl = __shfl_sync(0xffffffffu, rb, owner);
        const float ex = __shfl_sync(0xffffffffu, rext.x, owner), ey = __shfl_sync(0xffffffffu, rext.y, owner);
        if (active && pos < a.capacity) {
            a.gids_sorted[pos] = gid;
            __stcs(a.packed + 3 * (size_t)pos + 0, make_float4(x, y, o, cx));
            __stcs(a.packed + 3 * (size_t)pos + 1, make_float4(cy, cz, r, gg));
            __stcs(a.packed + 3 * (size_t)pos + 2, make_float4(bl, __int_as_float(gid), ex, ey));
        }
    };
    const int warp_total = __reduce_add_sync(0xffffffffu, cnt);
    if (warp_total <= 128) {               // every pair of the warp is in the memo words (warp-uniform branch)
#pragma unroll
        for (int round = 0; round < 4; ++round) {
            if (round * 32 >= warp_total) break;
            const uint32_t v = round == 0 ? memo0 : round == 1 ? memo1 : round == 2 ? memo2 : memo3;
            const bool active = (v >> 31) != 0;
            const int tile = (int)(v & 4095u), owner = (int)((v >> 12) & 31u), rank = (int)((v >> 17) & 63u);
            const int pos = active ? s_base[tile] + (int)s_grp[grp][tile] + (int)s_cnt[warp][tile] + rank : 0;
            emit(pos, i - lane + owner, owner, active);
        }
        return;
    }
    warp_pairs(cnt, x0, w, y0, a.TX, i, [&](int tile, int gid, int owner, bool active) {
        const unsigned peers = __match_any_sync(0xffffffffu, active ? tile : -1 - lane);
        int pos = 0;
        if (active) {
            const int before = s_cnt[warp][tile];
            pos = s_base[tile] + (int)s_grp[grp][tile] + before + __popc(peers & lt);
        }
        __syncwarp();
        if (active && (peers & lt) == 0) s_cnt[warp][tile] += (unsigned char)__popc(peers);
        __syncwarp();
        emit(pos, gid, owner, active);
    });
}

// ---------------------------------------------------------------------------------------------------------------
static inline int cuda_status() {
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? GI2D_OK : GI2D_ERR_CUDA_BASE + (int)e;
}
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline int ceil_log2(long long v) { int b = 0; while ((1ll << b) < v) ++b; return b; }

static PassPlan make_plan(int lo_bit, int hi_bit) {   // sort key bits [lo_bit, hi_bit)
    PassPlan p;
    p.npass = 0;
    for (int s = lo_bit; s < hi_bit && p.npass < MAX_PASSES; s += 8) {
        p.shift[p.npass] = s;
        p.bits[p.npass] = (hi_bit - s) < 8 ? (hi_bit - s) : 8;
        ++p.npass;
    }
    return p;
}

constexpr int SORT_ITEMS64 = 8, SORT_TILE64 = SB * SORT_ITEMS64;    // int64 keys (staged path): 2048 per block

// Launch shapes of the fused chain.  Everything here is latency-bound at GaussianImage sizes (the whole chain moves a
// few MB), so the shapes trade look-back chain length against per-block serial work; defaults were picked from
// measurements on B200 (profiles/), and can be overridden for experiments through the environment.
struct Tuning { int scan_t, pass_items, last_items; };
static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}
static const Tuning &tuning() {
    static Tuning t = {env_int("GI2D_SCAN_T", 1024), env_int("GI2D_PASS_ITEMS", 4), env_int("GI2D_LAST_ITEMS", 4)};
    return t;
}
constexpr int MIN_SCAN_T = 256, MIN_SORT_TILE = SB * 4;

// Look-back order may be blockIdx order when every block of the grid is resident at once (no ticket atomics).
template <typename K>
static bool all_resident(K kernel, int threads, int grid) {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    static const void *keys[64];
    static int vals[64], used = 0;          // tiny cache keyed by kernel address (the query costs microseconds)
    const void *key = reinterpret_cast<const void *>(kernel);
    int per_sm = -1;
    for (int i = 0; i < used; ++i)
        if (keys[i] == key) per_sm = vals[i];
    if (per_sm < 0) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess) per_sm = 0;
        if (used < 64) { keys[used] = key; vals[used] = per_sm; ++used; }
    }
    return grid <= per_sm * sms;
}

template <typename K>
static bool all_resident_smem(K kernel, int threads, size_t smem, int grid) {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    static const void *keys[4];
    static int vals[4], used = 0;
    const void *key = reinterpret_cast<const void *>(kernel);
    int per_sm = -1;
    for (int i = 0; i < used; ++i)
        if (keys[i] == key) per_sm = vals[i];
    if (per_sm < 0) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess) per_sm = 0;
        if (used < 4) { keys[used] = key; vals[used] = per_sm; ++used; }
        cudaGetLastError();
    }
    return grid <= per_sm * sms;
}

struct FusedLayout {          // carve-up of the caller's workspace for gi2d_bin_fused
    size_t ctl, tile_state, tile_cur, zero_scatter, scan_state, lookback, zero_bytes, keys_a, keys_b, vals_b, mat, rows, total;
    int scan_blocks, sort_blocks, npass;
};
static FusedLayout fused_layout(int N, int T, int capacity) {
    FusedLayout L;
    L.npass = make_plan(0, ceil_log2(T > 1 ? T : 2)).npass;
    L.scan_blocks = (N + MIN_SCAN_T - 1) / MIN_SCAN_T;                 // sized for the smallest block shapes
    L.sort_blocks = (capacity + MIN_SORT_TILE - 1) / MIN_SORT_TILE;
    size_t o = 0;
    L.ctl = o; o += align_up(sizeof(uint32_t) * CTL_WORDS, 256);
    L.tile_state = o; o += align_up(sizeof(uint32_t) * (size_t)((T + 1023) / 1024 + 1), 256);   // scatter path: look-back words
    L.tile_cur = o; o += align_up(sizeof(uint32_t) * (size_t)T, 256);                            // scatter path: per-tile counters
    L.zero_scatter = o;
    L.scan_state = o; o += align_up(sizeof(uint32_t) * (size_t)L.scan_blocks, 256);
    L.lookback = o; o += align_up(sizeof(uint32_t) * (size_t)L.npass * L.sort_blocks * RADIX, 256);
    L.zero_bytes = o;
    L.keys_a = o; o += align_up(sizeof(uint32_t) * (size_t)capacity, 256);
    L.keys_b = o; o += align_up(sizeof(uint32_t) * (size_t)capacity, 256);
    L.vals_b = o; o += align_up(sizeof(int) * (size_t)capacity, 256);
    L.mat = o;                                                         // one-kernel path (T <= 4096): (blocks + 1, Tpad) uint32
    if (T <= 4096) o += align_up(sizeof(uint32_t) * ((size_t)((N + 255) / 256) + 1) * (T <= 2048 ? 2048 : 4096), 256);
    L.rows = o;                                                        // scatter path: one 48-byte row per Gaussian
    if (T > 4096 || env_int("GI2D_SCATTER_BIN", 1) == 2) o += align_up(48 * (size_t)N, 256);
    L.total = o;
    return L;
}


// Chooses the one-kernel path's kernel for (N, T) on the current device, or nullptr when the shape must take the
// scan + onesweep chain (more than 4096 tiles, or a grid that cannot be co-resident).
struct SmallPlan { const void *fn; int nt, tp; size_t smem; };
static SmallPlan small_plan(int N, int T) {
    SmallPlan p{nullptr, 0, 0, 0};
    static const int small_ok = env_int("GI2D_SMALL_BIN", 1);
    static const int small_nt_env = env_int("GI2D_SMALL_NT", 0);
    if (!small_ok || T > 4096) return p;
    static int n_sms = 0;
    if (!n_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    // smallest block shape that keeps the grid within one block per SM (grid barriers are cheap only then)
    const int small_nt = small_nt_env ? small_nt_env : (N <= 256 * n_sms ? 256 : N <= 512 * n_sms ? 512 : 1024);
    const int tp = T <= 2048 ? 2048 : 4096;
    // prefer that shape; fall back to 1024 Gaussians per block when the grid would not be co-resident
    for (int attempt = 0; attempt < 2 && !p.fn; ++attempt) {
        const int nt = attempt == 0 ? (small_nt == 512 ? 512 : small_nt == 1024 ? 1024 : 256) : 1024;
        const int grid = (N + nt - 1) / nt;
        const size_t smem = (size_t)(nt / 32) * tp + (size_t)(nt / 256) * tp * 2 + (size_t)tp * 4;
#define GI2D_TRY(TB, NTT)                                                                                                  \
    if (tp == (1 << TB) && nt == NTT) {                                                                                    \
        static bool attr = (cudaFuncSetAttribute(bin_small_kernel<TB, NTT>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                                 (NTT / 32) * (1 << TB) + (NTT / 256) * (1 << TB) * 2 + (1 << TB) * 4), true); \
        (void)attr;                                                                                                        \
        if (all_resident_smem(bin_small_kernel<TB, NTT>, NTT, smem, grid))                                                 \
            p = SmallPlan{reinterpret_cast<const void *>(bin_small_kernel<TB, NTT>), NTT, tp, smem};                       \
    }
        GI2D_TRY(11, 256) GI2D_TRY(11, 512) GI2D_TRY(11, 1024) GI2D_TRY(12, 256) GI2D_TRY(12, 512) GI2D_TRY(12, 1024)
#undef GI2D_TRY
    }
    return p;
}

// ---------------------------------------------------------------------------------------------------------------
// Scatter path for images of MORE than 4096 tiles (the 8K decode: 129 600 tiles, 19 intersections per tile on average).
// A radix sort of the (tile, gid) pairs moves every pair once per 8 tile-id bits and is latency-bound at these sizes
// (C5: scan 61 us + 3 passes = 253 us).  Tiles are small, so bin by tile with atomics instead and repair the order
// inside each tile:
//   K1 tile_count    every (Gaussian, tile) pair adds 1 to its tile's counter (RED.ADD, no return value);
//   K2 tile_scan     exclusive scan of the counters in place (single pass, decoupled look-back) = first slot of every
//                    tile; the total goes to status / the mapped host word;
//   K3 tile_scatter  every pair takes slot = atomicAdd(cursor[tile], 1) and stores its gid there -- any order inside
//                    the tile; afterwards cursor[tile] is the END of the tile;
//   K4 tile_sort_pack  one warp per tile sorts the tile's gids ascending (a Gaussian hits a tile at most once, so the
//                    order is total and equals the reference's stable sort, Q2): <= 32 by shuffle ranking, <= 512 by a
//                    bitonic network in shared memory, longer in place in global memory; then writes tile_bins and the
//                    48-byte packed records.
// Deterministic output although the scatter is not.  Slots >= capacity are never written (overflow is reported).
struct ScatterArgs {
    int N, TX, TY, T, capacity, wy0, wy1;
    const float *xys;
    const int *radii, *nth;
    const float *conics, *colors, *opac;
    uint32_t *cur;            // (T): counters -> exclusive starts -> ends
    float4 *rows;             // (N,3): one 48-byte record per GAUSSIAN, written by the count pass, gathered by the sort pass
    uint32_t *scan_state, *ctl;
    int *gids_sorted, *tile_bins;
    float4 *packed;
    int *status_out;
    volatile int *status_host;
    FusedExtras ex;
};

__device__ __forceinline__ void scatter_rect(const ScatterArgs &a, int i, int &cnt, int &x0, int &w, int &y0) {
    cnt = 0; x0 = 0; w = 1; y0 = 0;
    if (i >= a.N) return;
    const int nth = __ldg(a.nth + i);
    if (nth <= 0) return;
    int area;
    gaussian_bbox(a.xys, a.radii, i, a.TX, a.TY, x0, w, y0, area, a.wy0, a.wy1);
    const bool windowed = a.wy0 > 0 || a.wy1 < a.TY;
    cnt = windowed ? area : min(nth, area);        // the projections always give num_tiles_hit == area
}

template <bool SCATTER>
__global__ void __launch_bounds__(256) tile_count_scatter_kernel(ScatterArgs a) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    int cnt, x0, w, y0;
    scatter_rect(a, i, cnt, x0, w, y0);
    if (!SCATTER) zero_grad_row(a.ex.zero_g12, i, a.N);
    if (!SCATTER && i < a.N && cnt > 0) {
        // The record of this Gaussian as ONE 48-byte row (coalesced reads here): the sort pass then gathers 3 x 16 bytes
        // per intersection instead of nine scalars from four arrays (five 32-byte sectors for 36 useful bytes).
        const size_t q = (size_t)i;
        const float x = __ldg(a.xys + 2 * q), y = __ldg(a.xys + 2 * q + 1);
        const float cx = __ldg(a.conics + 3 * q), cy = __ldg(a.conics + 3 * q + 1), cz = __ldg(a.conics + 3 * q + 2);
        const float r = __ldg(a.colors + 3 * q), gg = __ldg(a.colors + 3 * q + 1), b = __ldg(a.colors + 3 * q + 2);
        const float o = __ldg(a.opac + q);
        a.rows[3 * q + 0] = make_float4(x, y, o, cx);
        a.rows[3 * q + 1] = make_float4(cy, cz, r, gg);
        const float2 e = record_extent(cx, cy, cz, o);
        a.rows[3 * q + 2] = make_float4(b, __int_as_float(i), e.x, e.y);
    }
    warp_pairs(cnt, x0, w, y0, a.TX, i, [&](int tile, int gid, int, bool active) {
        if (!active) return;
        if (!SCATTER) {
            atomicAdd(a.cur + tile, 1u);
        } else {
            const uint32_t pos = atomicAdd(a.cur + tile, 1u);
            if (pos < (uint32_t)a.capacity) a.gids_sorted[pos] = gid;
        }
    });
}

template <int NT>
__global__ void __launch_bounds__(NT) tile_scan_kernel(ScatterArgs a, int use_ticket) {
    __shared__ int s_warp[NT / 32];
    __shared__ int s_bcast[2];
    const int tid = threadIdx.x;
    if (use_ticket && tid == 0) s_bcast[0] = (int)atomicAdd(a.ctl + CTL_SCAN_TICKET, 1u);
    __syncthreads();
    const int vbid = use_ticket ? s_bcast[0] : (int)blockIdx.x;
    const int t = vbid * NT + tid;
    const int c = t < a.T ? (int)a.cur[t] : 0;
    int block_total;
    const int incl = block_incl_scan<NT>(c, s_warp, &block_total);
    if (tid < 32) {
        const int e = lookback_exclusive(a.scan_state, vbid, block_total);
        if (tid == 0) s_bcast[1] = e;
    }
    __syncthreads();
    const int excl = s_bcast[1] + incl - c;
    if (t < a.T) a.cur[t] = (uint32_t)excl;
    if (vbid == (int)gridDim.x - 1 && tid == 0) {
        const int total = s_bcast[1] + block_total;
        const int over = total > a.capacity ? 1 : 0;
        a.ctl[CTL_N] = (uint32_t)total;
        a.ctl[CTL_OVERFLOW] = (uint32_t)over;
        if (a.status_out) { a.status_out[0] = total; a.status_out[1] = over; }
        publish_sticky(a.ex.sticky, total, over);
        if (a.status_host) {
            a.status_host[1] = over;
            __threadfence_system();
            a.status_host[0] = total;
        }
    }
}

// Ascending in-place sort of buf[0, L) by one warp: bitonic network in its all-merges-ascending form (first step of a
// merge pairs i with i ^ (k - 1)), so virtual +inf padding beyond L never has to move.  buf may be shared or global.
__device__ __forceinline__ void warp_bitonic(int *buf, int L) {
    const int lane = threadIdx.x & 31;
    int P = 64;
    while (P < L) P <<= 1;
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const bool flip = j == (k >> 1);
            for (int i = lane; i < P; i += 32) {
                const int l = flip ? (i ^ (k - 1)) : (i ^ j);
                if (l > i && l < L) {
                    const int x = buf[i], y = buf[l];
                    if (x > y) { buf[i] = y; buf[l] = x; }
                }
            }
            __syncwarp();
        }
    }
}

constexpr int SORT_SMEM = 512;     // gids per warp sorted in shared memory

constexpr int SORT_TPW = 2;         // tiles per warp, handled in lockstep: twice the loads in flight per warp (the kernel is latency-bound)

__global__ void __launch_bounds__(256) tile_sort_pack_kernel(ScatterArgs a) {
    __shared__ int s_buf[8][SORT_SMEM];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile0 = (blockIdx.x * 8 + warp) * SORT_TPW;
    if (tile0 >= a.T) return;
    const int cap = a.capacity;
    int lo[SORT_TPW], L[SORT_TPW];
    {   // ranges of the warp's consecutive tiles: cur[] holds the END of every tile after the scatter pass
        uint32_t e[SORT_TPW + 1];
        e[0] = tile0 ? a.cur[tile0 - 1] : 0u;
#pragma unroll
        for (int q = 0; q < SORT_TPW; ++q) e[q + 1] = tile0 + q < a.T ? a.cur[tile0 + q] : e[q];
#pragma unroll
        for (int q = 0; q < SORT_TPW; ++q) {
            lo[q] = min((int)min(e[q], (uint32_t)0x7fffffff), cap);
            L[q] = min((int)min(e[q + 1], (uint32_t)0x7fffffff), cap) - lo[q];
            if (lane == 0 && tile0 + q < a.T)
                reinterpret_cast<int2 *>(a.tile_bins)[tile0 + q] = L[q] > 0 ? make_int2(lo[q], lo[q] + L[q]) : make_int2(0, 0);
        }
    }
    auto put_record = [&](int base, int slot, int v) {
        const float4 *row = a.rows + 3 * (size_t)v;
        const float4 r0 = __ldg(row), r1 = __ldg(row + 1), r2 = __ldg(row + 2);
        float4 *dst = a.packed + 3 * (size_t)(base + slot);
        __stcs(dst + 0, r0);
        __stcs(dst + 1, r1);
        __stcs(dst + 2, r2);
    };
    bool all_small = true;
#pragma unroll
    for (int q = 0; q < SORT_TPW; ++q) all_small = all_small && L[q] <= 32;
    if (all_small) {                                // one gid per lane and tile: its rank is its slot, nothing is re-read
        int v[SORT_TPW], rank[SORT_TPW], Lmax = 0;
#pragma unroll
        for (int q = 0; q < SORT_TPW; ++q) {
            v[q] = lane < L[q] ? a.gids_sorted[lo[q] + lane] : 0x7fffffff;       // the loads of all tiles are in flight together
            rank[q] = 0;
            Lmax = max(Lmax, L[q]);
        }
        for (int k = 0; k < Lmax; ++k) {
#pragma unroll
            for (int q = 0; q < SORT_TPW; ++q) rank[q] += __shfl_sync(0xffffffffu, v[q], k) < v[q];    // (+inf padding never counts)
        }
        float4 r[SORT_TPW][3];
#pragma unroll
        for (int q = 0; q < SORT_TPW; ++q)
            if (lane < L[q]) {
                const float4 *row = a.rows + 3 * (size_t)v[q];
                r[q][0] = __ldg(row); r[q][1] = __ldg(row + 1); r[q][2] = __ldg(row + 2);
            }
#pragma unroll
        for (int q = 0; q < SORT_TPW; ++q)
            if (lane < L[q]) {
                a.gids_sorted[lo[q] + rank[q]] = v[q];
                float4 *dst = a.packed + 3 * (size_t)(lo[q] + rank[q]);
                __stcs(dst + 0, r[q][0]); __stcs(dst + 1, r[q][1]); __stcs(dst + 2, r[q][2]);
            }
        return;
    }
    for (int q = 0; q < SORT_TPW; ++q) {            // a longer list among them: one tile after the other
        if (L[q] <= 0) continue;
        int *g = a.gids_sorted + lo[q];
        if (L[q] <= 32) {
            const int v = lane < L[q] ? g[lane] : 0x7fffffff;
            int rank = 0;
            for (int k = 0; k < L[q]; ++k) rank += __shfl_sync(0xffffffffu, v, k) < v;
            __syncwarp();
            if (lane < L[q]) {
                g[rank] = v;
                put_record(lo[q], rank, v);
            }
            continue;
        }
        const int *src = g;
        if (L[q] <= SORT_SMEM) {
            for (int k = lane; k < L[q]; k += 32) s_buf[warp][k] = g[k];
            __syncwarp();
            warp_bitonic(s_buf[warp], L[q]);
            for (int k = lane; k < L[q]; k += 32) g[k] = s_buf[warp][k];
            src = s_buf[warp];
        } else {
            __syncwarp();
            warp_bitonic(g, L[q]);                   // rare: a tile with more than 512 intersections
        }
        __syncwarp();
        for (int k = lane; k < L[q]; k += 32) put_record(lo[q], k, src[k]);
        __syncwarp();                                // s_buf is reused by the next tile of this warp
    }
}

// Records of an EXISTING sorted list with other per-Gaussian attributes: a caller that renders the same geometry several
// times with different colours (gaussianimage_cholesky.py:81,89,95 -- three rasterize_gaussians_sum calls per forward on
// the same xys / radii / num_tiles_hit, quirk Q13) re-uses gids_sorted and tile_bins and only rebuilds the 48-byte records.
__global__ void __launch_bounds__(256) repack_kernel(int capacity, const int *__restrict__ gids_sorted, const int *__restrict__ status,
                                                      const float *__restrict__ xys, const float *__restrict__ conics,
                                                      const float *__restrict__ colors, const float *__restrict__ opac,
                                                      float4 *__restrict__ packed, int N = 0, float4 *zero_g12 = nullptr) {
    for (int i = blockIdx.x * 256 + threadIdx.x; i < N; i += gridDim.x * 256) zero_grad_row(zero_g12, i, N);
    const int n = min(__ldg(status), capacity);
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        const int v = __ldg(gids_sorted + i);
        const size_t g = (size_t)v;
        const float x = __ldg(xys + 2 * g), y = __ldg(xys + 2 * g + 1);
        const float cx = __ldg(conics + 3 * g), cy = __ldg(conics + 3 * g + 1), cz = __ldg(conics + 3 * g + 2);
        const float r = __ldg(colors + 3 * g), gg = __ldg(colors + 3 * g + 1), b = __ldg(colors + 3 * g + 2);
        const float o = __ldg(opac + g);
        __stcs(packed + 3 * (size_t)i + 0, make_float4(x, y, o, cx));
        __stcs(packed + 3 * (size_t)i + 1, make_float4(cy, cz, r, gg));
        const float2 e = record_extent(cx, cy, cz, o);
        __stcs(packed + 3 * (size_t)i + 2, make_float4(b, __int_as_float(v), e.x, e.y));
    }
}

}  // namespace gi2d

using namespace gi2d;

extern "C" {

// ---- staged, reference-typed entry points ------------------------------------------------------------------------

size_t gi2d_scan_workspace_bytes(int N) {
    return align_up(sizeof(uint32_t) * CTL_WORDS, 256) + align_up(sizeof(uint32_t) * (size_t)((N + 1023) / 1024), 256);
}

// compute_cumulative_intersects (utils.py:99-118): cum (N,) int32 inclusive; the total is written to *total_out
// (device or pinned-host-mapped pointer) by a 4-byte async copy on `stream`.
int gi2d_scan_tiles(int N, const int *num_tiles_hit, int *cum_tiles_hit, int *total_out, void *ws, size_t ws_bytes,
                    cudaStream_t stream) {
    if (N <= 0) return GI2D_ERR_INVALID;
    const size_t need = gi2d_scan_workspace_bytes(N);
    if (ws_bytes < need) return GI2D_ERR_WORKSPACE;
    uint32_t *ctl = reinterpret_cast<uint32_t *>(ws);
    uint32_t *state = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(ws) + align_up(sizeof(uint32_t) * CTL_WORDS, 256));
    cudaMemsetAsync(ws, 0, need, stream);
    PassPlan none;
    none.npass = 0;
    const int grid = (N + 1023) / 1024;
    const int ticket = all_resident(scan_kernel<false, 1024>, 1024, grid) ? 0 : 1;
    scan_kernel<false, 1024><<<grid, 1024, 0, stream>>>(N, num_tiles_hit, cum_tiles_hit, ctl, state, nullptr, nullptr, 0, 0, 0, nullptr,
                                                   nullptr, none, nullptr, 0, nullptr, nullptr, ticket, 0, 0, 0x7fffffff);
    if (total_out) cudaMemcpyAsync(total_out, ctl + CTL_N, sizeof(int), cudaMemcpyDefault, stream);
    return cuda_status();
}

// map_gaussian_to_intersects (utils.py:12-50): every one of the n_isect slots is written.
int gi2d_emit_isects(int N, const float *xys, const float *depths, const int *radii, const int *cum_tiles_hit, int TX,
                     int TY, int n_isect, long long *isect_ids, int *gaussian_ids, cudaStream_t stream) {
    if (N <= 0 || n_isect < 0) return GI2D_ERR_INVALID;
    if (n_isect == 0) return GI2D_OK;
    // slots not covered by any Gaussian's range (cum inconsistent with n_isect) keep the reference's zero fill
    cudaMemsetAsync(isect_ids, 0, sizeof(long long) * (size_t)n_isect, stream);
    cudaMemsetAsync(gaussian_ids, 0, sizeof(int) * (size_t)n_isect, stream);
    emit64_kernel<<<(N + SB - 1) / SB, SB, 0, stream>>>(N, xys, depths, radii, cum_tiles_hit, TX, TY, n_isect, isect_ids, gaussian_ids);
    return cuda_status();
}

size_t gi2d_sort_workspace_bytes(int n, int begin_bit, int end_bit) {
    const PassPlan plan = make_plan(begin_bit, end_bit);
    const int blocks = (n + SORT_TILE64 - 1) / SORT_TILE64;
    size_t o = align_up(sizeof(uint32_t) * CTL_WORDS, 256);
    o += align_up(sizeof(uint32_t) * (size_t)plan.npass * blocks * RADIX, 256);
    o += align_up(sizeof(long long) * (size_t)n, 256) + align_up(sizeof(int) * (size_t)n, 256);
    return o;
}

// torch.sort(isect_ids) + gather (utils.py:164-165): stable ascending SIGNED int64 sort of key bits
// [begin_bit, end_bit) (0..64 sorts everything; 32..32+ceil(log2(tiles)) suffices when all depths are 0).
int gi2d_sort_isects(int n, const long long *ids_in, const int *gids_in, long long *ids_out, int *gids_out,
                     int begin_bit, int end_bit, void *ws, size_t ws_bytes, cudaStream_t stream) {
    if (n < 0 || begin_bit < 0 || end_bit > 64 || begin_bit >= end_bit) return GI2D_ERR_INVALID;
    if (n == 0) return GI2D_OK;
    if (ws_bytes < gi2d_sort_workspace_bytes(n, begin_bit, end_bit)) return GI2D_ERR_WORKSPACE;
    const PassPlan plan = make_plan(begin_bit, end_bit);
    const int blocks = (n + SORT_TILE64 - 1) / SORT_TILE64;
    char *base = reinterpret_cast<char *>(ws);
    size_t o = 0;
    uint32_t *ctl = reinterpret_cast<uint32_t *>(base + o); o += align_up(sizeof(uint32_t) * CTL_WORDS, 256);
    uint32_t *lookback = reinterpret_cast<uint32_t *>(base + o); o += align_up(sizeof(uint32_t) * (size_t)plan.npass * blocks * RADIX, 256);
    const size_t zero_bytes = o;
    unsigned long long *tmp_k = reinterpret_cast<unsigned long long *>(base + o); o += align_up(sizeof(long long) * (size_t)n, 256);
    int *tmp_v = reinterpret_cast<int *>(base + o);
    cudaMemsetAsync(ws, 0, zero_bytes, stream);
    const unsigned long long flip = end_bit == 64 ? 0x8000000000000000ull : 0ull;   // signed order
    const unsigned long long *kin = reinterpret_cast<const unsigned long long *>(ids_in);
    unsigned long long *kout = reinterpret_cast<unsigned long long *>(ids_out);
    hist_kernel<unsigned long long><<<min(blocks, 1024), SB, 0, stream>>>(n, kin, ctl, plan, flip);
    const int ticket = all_resident(onesweep_kernel<unsigned long long, SORT_ITEMS64, false>, SB, blocks) ? 0 : 1;
    PackSrc none{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0};
    // ping-pong so that the last pass lands in (ids_out, gids_out)
    const unsigned long long *src_k = kin;
    const int *src_v = gids_in;
    for (int p = 0; p < plan.npass; ++p) {
        const bool to_out = ((plan.npass - 1 - p) % 2) == 0;
        unsigned long long *dst_k = to_out ? kout : tmp_k;
        int *dst_v = to_out ? gids_out : tmp_v;
        onesweep_kernel<unsigned long long, SORT_ITEMS64, false><<<blocks, SB, 0, stream>>>(
            nullptr, n, n, src_k, src_v, dst_k, dst_v, ctl, lookback + (size_t)p * blocks * RADIX, p, plan.shift[p], plan.bits[p], flip,
            ticket, none);
        src_k = dst_k;
        src_v = dst_v;
    }
    return cuda_status();
}

// get_tile_bin_edges (utils.py:53-74): tile_bins (n_bins,2) int32, fully written (zero for empty tiles).
int gi2d_tile_bins(int n, const long long *isect_ids_sorted, int n_bins, int *tile_bins, cudaStream_t stream) {
    if (n < 0 || n_bins < 0) return GI2D_ERR_INVALID;
    if (n_bins > 0) cudaMemsetAsync(tile_bins, 0, sizeof(int) * 2 * (size_t)n_bins, stream);
    if (n == 0) return cuda_status();
    const int blocks = min((n + SB - 1) / SB, 4096);
    tile_bins_kernel<long long, 32><<<blocks, SB, 0, stream>>>(n, isect_ids_sorted, n_bins, tile_bins);
    return cuda_status();
}

// ---- fused, sync-free chain ---------------------------------------------------------------------------------------

// Barrier words of the one-kernel path: every launch gets its own library-owned pair (gi2d_common.cuh: sync_slot), so two
// grids of this kernel running at the same moment -- two streams, two graphs replayed side by side -- never share a barrier.
// GI2D_OWN_BARRIER=0, or no pair available, falls back to cooperative-groups' grid.sync().
static unsigned int *small_barrier(cudaStream_t stream) {
    static const int own = env_int("GI2D_OWN_BARRIER", 1);
    return own ? sync_slot(stream) : nullptr;
}

// GI2D_SCATTER_BIN: 0 = never, 1 (default) = images of more than 4096 tiles, 2 = whenever the one-kernel path does not apply
static bool use_scatter(int T) {
    static const int mode = env_int("GI2D_SCATTER_BIN", 1);
    return mode == 2 || (mode == 1 && T > 4096);
}

// Number of kernels gi2d_bin_fused launches for this shape on the current device (1 = the one-kernel path; otherwise the
// scan kernel + one onesweep pass per 8 tile-id bits); the zero-fill of the chain's control block is a memset, not a kernel.
int gi2d_bin_fused_kernel_count(int N, int TX, int TY) {
    if (N <= 0 || TX <= 0 || TY <= 0) return 0;
    const int T = TX * TY;
    if (small_plan(N, T).fn) return 1;
    if (use_scatter(T)) return 4;
    return 1 + make_plan(0, ceil_log2(T > 1 ? T : 2)).npass;
}

size_t gi2d_bin_fused_workspace_bytes(int N, int TX, int TY, int capacity) {
    return fused_layout(N, TX * TY, capacity).total;
}

// scan + emit + sort + bin edges + record packing with the intersection count left on the device, restricted to the
// tile rows [ty_begin, ty_end) (a row-sharded render bins only its own rows; tile ids stay global):
//   memset(control) -> scan_kernel<FUSED> -> onesweep pass(es) -> last onesweep pass (+ bin edges + packed records)
//   status_out[0] = n_isect (may exceed capacity), status_out[1] = 1 when n_isect > capacity (results truncated).
//   gids_sorted (capacity,) int32, packed (capacity*3) float4,
//   tile_bins (TX*TY,2) int32: [lo,hi) per tile; EMPTY tiles hold (INT_MAX, 0) until gi2d_raster_sum_fwd_packed
//   rewrites them to the reference's (0, 0).
// Precondition (2D path): depths are all zero, so the key is the tile id alone.
static int bin_fused_impl(int N, int TX, int TY, int ty_begin, int ty_end, const float *xys, const int *radii,
                          const int *num_tiles_hit, const float *conics, const float *colors, const float *opacities,
                          int capacity, void *ws, size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed,
                          int *status_out, int *status_host, FusedExtras ex, cudaStream_t stream) {
    if (N <= 0 || TX <= 0 || TY <= 0 || capacity <= 0 || ty_begin < 0 || ty_end > TY || ty_begin >= ty_end) return GI2D_ERR_INVALID;
    const int T = TX * TY;
    const FusedLayout L = fused_layout(N, T, capacity);
    if (ws_bytes < L.total) return GI2D_ERR_WORKSPACE;
    char *base = reinterpret_cast<char *>(ws);
    uint32_t *ctl = reinterpret_cast<uint32_t *>(base + L.ctl);
    uint32_t *scan_state = reinterpret_cast<uint32_t *>(base + L.scan_state);
    uint32_t *lookback = reinterpret_cast<uint32_t *>(base + L.lookback);
    uint32_t *keys_a = reinterpret_cast<uint32_t *>(base + L.keys_a);
    uint32_t *keys_b = reinterpret_cast<uint32_t *>(base + L.keys_b);
    int *vals_b = reinterpret_cast<int *>(base + L.vals_b);
    const PassPlan plan = make_plan(0, ceil_log2(T > 1 ? T : 2));

    // images of at most 4096 tiles, all blocks co-resident: ONE cooperative kernel (counting sort by tile)
    static const int dbg = env_int("GI2D_DBG", 0);
    if (ty_begin == 0 && ty_end == TY) {
        const SmallPlan sp = small_plan(N, T);
        if (sp.fn) {
            SmallArgs a{N, TX, TY, T, capacity, xys, radii, num_tiles_hit, conics, colors, opacities,
                        reinterpret_cast<uint32_t *>(base + L.mat), reinterpret_cast<uint32_t *>(base + L.mat) + (size_t)((N + 255) / 256) * sp.tp,
                        ctl, gids_sorted, tile_bins, reinterpret_cast<float4 *>(packed), status_out, status_host, dbg,
                        small_barrier(stream), ex};
            void *args[] = {&a};
            const cudaError_t e = cudaLaunchCooperativeKernel(sp.fn, dim3((N + sp.nt - 1) / sp.nt), dim3(sp.nt), args, sp.smem, stream);
            return e == cudaSuccess ? GI2D_OK : GI2D_ERR_CUDA_BASE + (int)e;
        }
    }

    if (use_scatter(T)) {          // count -> scan -> scatter -> per-tile sort + records
        cudaMemsetAsync(ws, 0, L.zero_scatter, stream);
        ScatterArgs a{N, TX, TY, T, capacity, ty_begin, ty_end, xys, radii, num_tiles_hit, conics, colors, opacities,
                      reinterpret_cast<uint32_t *>(base + L.tile_cur), reinterpret_cast<float4 *>(base + L.rows),
                      reinterpret_cast<uint32_t *>(base + L.tile_state), ctl,
                      gids_sorted, tile_bins, reinterpret_cast<float4 *>(packed), status_out, status_host, ex};
        const int gblocks = (N + 255) / 256, sblocks = (T + 1023) / 1024;
        tile_count_scatter_kernel<false><<<gblocks, 256, 0, stream>>>(a);
        tile_scan_kernel<1024><<<sblocks, 1024, 0, stream>>>(a, all_resident(tile_scan_kernel<1024>, 1024, sblocks) ? 0 : 1);
        tile_count_scatter_kernel<true><<<gblocks, 256, 0, stream>>>(a);
        tile_sort_pack_kernel<<<(T + 8 * SORT_TPW - 1) / (8 * SORT_TPW), 256, 0, stream>>>(a);
        return cuda_status();
    }

    cudaMemsetAsync(ws, 0, L.zero_bytes, stream);
    // values ping-pong between gids_sorted and vals_b; start so that the last pass lands in gids_sorted
    int *v0 = (plan.npass % 2 == 0) ? gids_sorted : vals_b;
    uint32_t *k0 = keys_a;
    const Tuning &tn = tuning();
    int2 *bins2 = reinterpret_cast<int2 *>(tile_bins);
#define GI2D_SCAN(NT)                                                                                                      \
    {                                                                                                                      \
        const int grid = (N + NT - 1) / NT;                                                                                \
        const int ticket = all_resident(scan_kernel<true, NT>, NT, grid) ? 0 : 1;                                          \
        scan_kernel<true, NT><<<grid, NT, 0, stream>>>(N, num_tiles_hit, nullptr, ctl, scan_state, xys, radii, TX, TY,     \
                                                       capacity, k0, v0, plan, bins2, T, status_out, status_host, ticket, dbg,        \
                                                       ty_begin, ty_end, ex);                                             \
    }
    if (tn.scan_t == 256) GI2D_SCAN(256) else if (tn.scan_t == 512) GI2D_SCAN(512) else GI2D_SCAN(1024)
#undef GI2D_SCAN
    if (dbg & 4) return cuda_status();      // experiments: scan kernel only
    PackSrc none{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0};
    PackSrc ps{xys, conics, colors, opacities, reinterpret_cast<float4 *>(packed), tile_bins, T};
    const uint32_t *src_k = k0;
    const int *src_v = v0;
    for (int p = 0; p < plan.npass; ++p) {
        uint32_t *dst_k = (src_k == keys_a) ? keys_b : keys_a;
        int *dst_v = (src_v == gids_sorted) ? vals_b : gids_sorted;
        uint32_t *lb = lookback + (size_t)p * L.sort_blocks * RADIX;
        const bool last = p == plan.npass - 1;
#define GI2D_SWEEP(ITEMS, LASTF, PS)                                                                                       \
    {                                                                                                                      \
        const int grid = (capacity + SB * ITEMS - 1) / (SB * ITEMS);                                                       \
        const int ticket = all_resident(onesweep_kernel<uint32_t, ITEMS, LASTF>, SB, grid) ? 0 : 1;                        \
        onesweep_kernel<uint32_t, ITEMS, LASTF><<<grid, SB, 0, stream>>>(ctl + CTL_N, 0, capacity, src_k, src_v, dst_k,     \
                                                                        dst_v, ctl, lb, p, plan.shift[p], plan.bits[p], 0u, \
                                                                        ticket, PS);                                       \
    }
        const int items = last ? tn.last_items : tn.pass_items;
        if (last) {
            if (items == 4) GI2D_SWEEP(4, true, ps) else if (items == 8) GI2D_SWEEP(8, true, ps) else GI2D_SWEEP(16, true, ps)
        } else {
            if (items == 4) GI2D_SWEEP(4, false, none) else if (items == 8) GI2D_SWEEP(8, false, none) else GI2D_SWEEP(16, false, none)
        }
#undef GI2D_SWEEP
        src_k = dst_k;
        src_v = dst_v;
    }
    return cuda_status();
}

int gi2d_bin_fused_rows_notify(int N, int TX, int TY, int ty_begin, int ty_end, const float *xys, const int *radii,
                               const int *num_tiles_hit, const float *conics, const float *colors, const float *opacities,
                               int capacity, void *ws, size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed,
                               int *status_out, int *status_host, cudaStream_t stream) {
    return bin_fused_impl(N, TX, TY, ty_begin, ty_end, xys, radii, num_tiles_hit, conics, colors, opacities, capacity, ws, ws_bytes,
                          gids_sorted, tile_bins, packed, status_out, status_host, FusedExtras{nullptr, nullptr}, stream);
}

// gi2d_bin_fused_rows_notify with the optional extras of struct gi2d_bin_args (include/gi2d.h): a sticky status pair and the
// zero fill of the backward pass's gradient rows.
int gi2d_bin_fused_ex(const gi2d_bin_args *a, cudaStream_t stream) {
    if (!a) return GI2D_ERR_INVALID;
    if (a->zero_grads12 && (reinterpret_cast<uintptr_t>(a->zero_grads12) & 15)) return GI2D_ERR_INVALID;
    return bin_fused_impl(a->N, a->TX, a->TY, a->ty_begin, a->ty_end, a->xys, a->radii, a->num_tiles_hit, a->conics, a->colors,
                          a->opacities, a->capacity, a->ws, a->ws_bytes, a->gids_sorted, a->tile_bins, a->packed, a->status_out,
                          a->status_host, FusedExtras{a->status_sticky, reinterpret_cast<float4 *>(a->zero_grads12)}, stream);
}

int gi2d_bin_fused_rows(int N, int TX, int TY, int ty_begin, int ty_end, const float *xys, const int *radii,
                        const int *num_tiles_hit, const float *conics, const float *colors, const float *opacities,
                        int capacity, void *ws, size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed,
                        int *status_out, cudaStream_t stream) {
    return gi2d_bin_fused_rows_notify(N, TX, TY, ty_begin, ty_end, xys, radii, num_tiles_hit, conics, colors, opacities, capacity, ws,
                                      ws_bytes, gids_sorted, tile_bins, packed, status_out, nullptr, stream);
}

int gi2d_bin_fused(int N, int TX, int TY, const float *xys, const int *radii, const int *num_tiles_hit,
                   const float *conics, const float *colors, const float *opacities, int capacity, void *ws,
                   size_t ws_bytes, int *gids_sorted, int *tile_bins, float *packed, int *status_out,
                   cudaStream_t stream) {
    if (TY <= 0) return GI2D_ERR_INVALID;
    return gi2d_bin_fused_rows(N, TX, TY, 0, TY, xys, radii, num_tiles_hit, conics, colors, opacities, capacity, ws, ws_bytes,
                               gids_sorted, tile_bins, packed, status_out, stream);
}

int gi2d_repack_records_ex(int capacity, const int *gids_sorted, const int *status, const float *xys, const float *conics,
                           const float *colors, const float *opacities, float *packed, int N, float *zero_grads12,
                           cudaStream_t stream) {
    if (capacity < 0 || !gids_sorted || !status || !xys || !conics || !colors || !opacities || !packed) return GI2D_ERR_INVALID;
    if (zero_grads12 && (N <= 0 || (reinterpret_cast<uintptr_t>(zero_grads12) & 15))) return GI2D_ERR_INVALID;
    if (capacity == 0 && !zero_grads12) return GI2D_OK;
    const int work = std::max(capacity, zero_grads12 ? N : 0);
    const int grid = std::min((work + 255) / 256, 8 * 148);
    repack_kernel<<<grid, 256, 0, stream>>>(capacity, gids_sorted, status, xys, conics, colors, opacities,
                                            reinterpret_cast<float4 *>(packed), zero_grads12 ? N : 0,
                                            reinterpret_cast<float4 *>(zero_grads12));
    return cuda_status();
}

int gi2d_repack_records(int capacity, const int *gids_sorted, const int *status, const float *xys, const float *conics,
                        const float *colors, const float *opacities, float *packed, cudaStream_t stream) {
    return gi2d_repack_records_ex(capacity, gids_sorted, status, xys, conics, colors, opacities, packed, 0, nullptr, stream);
}

// ---- allocation-free variant of gi2d_scan_tiles (no counterpart in the reference; see include/gi2d.h) -----------------
// The scan's scratch (control block + one look-back word per block of 1024 Gaussians) lives in a file-scope device array,
// resident on every GPU that loads the library until the process exits: GI2D_SCAN_STATIC_BYTES = 24832 bytes, enough for
// GI2D_SCAN_STATIC_MAX_N = 2^22 Gaussians.  In host code `g_scan_static` is only a handle: the device address is looked up
// once per device with cudaGetSymbolAddress (that device current) and cached.  All calls on a device share the array.
constexpr size_t SCAN_STATIC_CTL_BYTES = (sizeof(uint32_t) * CTL_WORDS + 255) / 256 * 256;
constexpr size_t SCAN_STATIC_BYTES = SCAN_STATIC_CTL_BYTES + sizeof(uint32_t) * (GI2D_SCAN_STATIC_MAX_N / 1024);
static_assert(SCAN_STATIC_BYTES == GI2D_SCAN_STATIC_BYTES, "include/gi2d.h states the size of the static scratch");
static __device__ __align__(256) unsigned char g_scan_static[SCAN_STATIC_BYTES];

static int scan_static_address(void **out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static void *addr[MAX_DEV] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return GI2D_ERR_CUDA_BASE + (int)e;
    if (dev < 0 || dev >= MAX_DEV) return GI2D_ERR_INVALID;
    std::lock_guard<std::mutex> g(mu);
    if (!addr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_scan_static);      // the current device's copy of the array
        if (e != cudaSuccess) return GI2D_ERR_CUDA_BASE + (int)e;
        if (!p) return GI2D_ERR_INVALID;
        addr[dev] = p;
    }
    *out = addr[dev];
    return GI2D_OK;
}

int gi2d_scan_tiles_static(int N, const int *num_tiles_hit, int *cum_tiles_hit, int *total_out, cudaStream_t stream) {
    if (N <= 0 || !num_tiles_hit || !cum_tiles_hit) return GI2D_ERR_INVALID;
    if (N > GI2D_SCAN_STATIC_MAX_N) return GI2D_ERR_STATIC_LIMIT;
    void *ws = nullptr;
    const int rc = scan_static_address(&ws);
    if (rc != GI2D_OK) return rc;
    const size_t need = gi2d_scan_workspace_bytes(N);      // <= SCAN_STATIC_BYTES for every accepted N
    uint32_t *ctl = reinterpret_cast<uint32_t *>(ws);
    uint32_t *state = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(ws) + SCAN_STATIC_CTL_BYTES);
    cudaMemsetAsync(ws, 0, need, stream);                  // the array keeps the previous call's flags: everything read starts at zero
    PassPlan none;
    none.npass = 0;
    const int grid = (N + 1023) / 1024;
    // look-back in ticket order: correct whether or not all blocks are co-resident
    scan_kernel<false, 1024><<<grid, 1024, 0, stream>>>(N, num_tiles_hit, cum_tiles_hit, ctl, state, nullptr, nullptr, 0, 0, 0, nullptr,
                                                   nullptr, none, nullptr, 0, nullptr, nullptr, 1, 0, 0, 0x7fffffff);
    if (total_out) cudaMemcpyAsync(total_out, ctl + CTL_N, sizeof(int), cudaMemcpyDefault, stream);
    return cuda_status();
}

}  // extern "C"
