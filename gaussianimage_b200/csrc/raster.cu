// raster.cu -- accumulated-sum rasterizer, forward and backward.
//
// Replaces the reference's rasterize_forward_sum (gsplat/gsplat/cuda/csrc/forward.cu:512-627) and
// rasterize_backward_sum_kernel (backward.cu:696-862) with their torch::zeros fills (bindings.cu:443-451,
// 752-756).  The reference runs one CTA of 256 pixel-threads per 16x16 tile with two block barriers per batch; here
// ONE WARP owns a tile, warps are independent pipelines (persistent grid, no block-wide barrier at all), and:
//   * the tile's batch (<= 256 intersections, quirk Q1) is a contiguous span of 48-byte packed records
//     (binning.cu) brought into shared memory in chunks of 64 by bulk TMA copies (cp.async.bulk + mbarrier expect-tx)
//     into a per-warp ring, one chunk in flight while the previous one is evaluated -- no per-thread index-indirect
//     gathers, no colour loads from global in the inner loop;
//   * every pair the reference would reject with `alpha < 1/255` for geometric reasons is culled before it is
//     evaluated, using a conservative per-Gaussian pixel extent (gi2d_common.cuh: visible_extent).  Culling only
//     removes pairs the reference rejects, so no output bit changes;
//   * forward: a lane owns a 4x2 pixel block and walks, in ascending order, only the records whose rectangle touches
//     that block (per-lane bit masks built with 12 ballots per 32 records) -- the summation order per pixel is the
//     reference's, the image is bit-identical;
//   * backward: unit-parallel over the records' visible rectangles -- a lane owns a contiguous piece of the flattened
//     (record, row, 4-pixel group) list, partial gradients are combined per record through shared memory with plain
//     stores, and one set of (vector) reductions per (tile, Gaussian) goes to global memory; no shuffle reductions.
//     (The sum rasterizer has no transmittance chain, so the order over pixels is free.)
// Per-pair arithmetic up to the two comparisons is the reference's instruction sequence (gi2d_common.cuh).
// FP32-pipe / issue bound; no tensor cores (not a dense contraction).
#include "gi2d_common.cuh"
#include <type_traits>

namespace gi2d {

// -DGI2D_TRACE (tools/trace_raster.py builds its own copy of the library with it; never the product build): every warp of the
// rasterizers writes {globaltimer at entry, clock at entry, clock after its first chunk landed, clock at exit, SM id, records}
// to a buffer, to see launch ramp, load latency and tail separately.
#ifdef GI2D_TRACE
__device__ unsigned long long *g_trace = nullptr;
#define GI2D_TRACE_DECL unsigned long long tr_[6] = {0, 0, 0, 0, 0, 0}
#define GI2D_TRACE_ENTRY do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_[0])); tr_[1] = clock64(); unsigned sm_; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_)); tr_[4] = sm_; } while (0)
#define GI2D_TRACE_FIRST do { if (!tr_[2]) tr_[2] = clock64(); } while (0)
#define GI2D_TRACE_WORK(n) do { tr_[5] += (n); } while (0)
#define GI2D_TRACE_EXIT(widx) do { tr_[3] = clock64(); if (g_trace && (threadIdx.x & 31) == 0) for (int k_ = 0; k_ < 6; ++k_) g_trace[6 * (size_t)(widx) + k_] = tr_[k_]; } while (0)
#else
#define GI2D_TRACE_DECL
#define GI2D_TRACE_ENTRY
#define GI2D_TRACE_FIRST
#define GI2D_TRACE_WORK(n)
#define GI2D_TRACE_EXIT(widx)
#endif

struct GatherSrc {        // index-indirect sources of the reference-typed entry points
    const int *gids_sorted;
    const float *xys, *conics, *colors, *opac;
};

// Local (tile-relative, inclusive) pixel rectangle in which the record can pass the alpha test.
// Returns false for "never visible in this tile".
__device__ __forceinline__ bool record_rect(const float4 &A, const float4 &C, int tx0, int ty0, int &lx0, int &lx1, int &ly0,
                                            int &ly1) {
    const float ex = C.z, ey = C.w;                       // carried in the record (record_extent): +inf = whole tile
    lx0 = 0; lx1 = GI2D_TILE - 1; ly0 = 0; ly1 = GI2D_TILE - 1;
    if (ex < 0.f) return false;
    // pixel centres sit at integer coordinates (px = (float)j, forward.cu:537-538)
    const float fx0 = ceilf(A.x - ex) - (float)tx0, fx1 = floorf(A.x + ex) - (float)tx0;
    const float fy0 = ceilf(A.y - ey) - (float)ty0, fy1 = floorf(A.y + ey) - (float)ty0;
    if (!(fx0 <= fx1) || !(fy0 <= fy1)) return false;
    if (fx1 < 0.f || fy1 < 0.f || fx0 > (float)(GI2D_TILE - 1) || fy0 > (float)(GI2D_TILE - 1)) return false;
    lx0 = (int)fmaxf(fx0, 0.f); lx1 = (int)fminf(fx1, (float)(GI2D_TILE - 1));
    ly0 = (int)fmaxf(fy0, 0.f); ly1 = (int)fminf(fy1, (float)(GI2D_TILE - 1));
    return true;
}

// NVLS multicast stores (PTX multimem.st): one store, replicated by the switch to every device of the multicast group.
__device__ __forceinline__ void multimem_st_v4(float *mc, float4 v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_v2(float *mc, float2 v) {
    asm volatile("multimem.st.weak.global.v2.f32 [%0], {%1, %2};" ::"l"(mc), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ void multimem_st(float *mc, float v) {
    asm volatile("multimem.st.weak.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

// Per-warp chunk pipeline shared by both rasterizers.  A warp walks the tiles  first, first + stride, ...  of its range; the
// records of a tile (<= 256, quirk Q1) arrive in chunks of PIPE_CHUNK by bulk TMA copies (cp.async.bulk + mbarrier
// expect-tx, SASS UBLKCP) into a ring of PIPE_NST stages that belongs to the warp alone.  The PRODUCER cursor (pk, p_off)
// runs PIPE_NST chunks ahead of the consumer -- across tile boundaries, skipping empty tiles -- so a copy always has a whole
// chunk of evaluation to land; tile ranges are fetched one tile ahead of either cursor.  Everything here is warp-uniform;
// lane 0 issues the copies.  No block-wide barrier is involved.
constexpr int PIPE_CHUNK = 64;                    // records per chunk (two per lane)
constexpr int PIPE_NST = 2;                       // chunks in flight per warp

struct ChunkPipe {
    const int2 *bins2;
    const float4 *packed;
    float4 *ring;
    uint64_t *bars;
    // Work ITEMS: a tile is 1 << ishift items, dealt to different warps.  Either every item of a tile takes all of the
    // tile's chunks (forward: the items are the two halves of the tile's pixels) or the items
    // split the chunk sequence between them (backward: item j takes chunks j, j + items, ...; gradients add up in global
    // memory anyway).  The item stride of a warp is a multiple of the items per tile, so j (jidx) is a constant of the warp.
    int tile_begin, n_items, stride, lane, ishift;
    int sshift, jidx;                             // chunk subsequences per tile = 1 << sshift (0 unless the items split the chunks); this warp's
    int stage_f4;                                 // float4 per ring stage (>= 3 * PIPE_CHUNK; the forward appends a zero record)
    int pk, p_off, p_cs;
    int2 p_rg, p_nx;
    uint32_t seq;                                 // chunks consumed so far

    // (lo, cnt <= 256) of the tile of item q; empty tiles may still hold the (INT_MAX, 0) preset of gi2d_bin_fused
    __device__ __forceinline__ int2 load_range(int q) const {
        const int2 rg = __ldcg(bins2 + tile_begin + (q >> ishift));
        const int c = rg.y - rg.x;
        return make_int2(rg.x, c > 0 && rg.x != 0x7fffffff ? min(c, GI2D_BATCH) : 0);     // Q1: first 256 only
    }
    // Records per chunk of a tile with n records.  Forward: always PIPE_CHUNK.  Backward (items split the chunks): the tile's
    // records are cut into EQUAL chunks, their number a multiple of the items per tile -- with chunks of 64 a tile of 111
    // records gave one warp 64 records and its partner 47, and since partners sit in CTAs of opposite parity, the SMs' shares
    // of the records differed by +-25 % (tools/trace_raster.py).
    __device__ __forceinline__ int chunk_size(int n) const {
        if (sshift == 0) return PIPE_CHUNK;
        int k = (n + PIPE_CHUNK - 1) / PIPE_CHUNK;
        const int mask = (1 << sshift) - 1;
        if (n > 32) k = (k + mask) & ~mask;
        return k <= 1 ? n : k == 2 ? (n + 1) >> 1 : k == 3 ? ((n + 2) * 171) >> 9 : k == 4 ? (n + 3) >> 2 : (n + k - 1) / k;
    }
    __device__ __forceinline__ void produce(int st) {
        while (pk < n_items && p_off >= p_rg.y) {          // item exhausted (or empty): move on
            pk += stride;
            p_rg = p_nx;
            p_cs = chunk_size(p_rg.y);
            p_off = jidx * p_cs;
            p_nx = (pk + stride < n_items) ? load_range(pk + stride) : make_int2(0, 0);
        }
        if (pk >= n_items) return;
        const uint32_t bytes = (uint32_t)min(p_cs, p_rg.y - p_off) * 48u;
        if (lane == 0) {
            mbar_expect_tx(&bars[st], bytes);
            tma_load_1d(ring + (size_t)st * stage_f4, packed + 3 * (size_t)(p_rg.x + p_off), bytes, &bars[st]);
        }
        p_off += p_cs << sshift;
    }
    __device__ __forceinline__ void init(const int *tile_bins, const float4 *packed_, float4 *ring_, uint64_t *bars_, int tile_begin_,
                                         int n_tiles, int ishift_, bool split_chunks, int first, int stride_, int lane_, bool tma,
                                         int stage_f4_ = PIPE_CHUNK * 3) {
        stage_f4 = stage_f4_;
        bins2 = reinterpret_cast<const int2 *>(tile_bins);
        packed = packed_; ring = ring_; bars = bars_;
        tile_begin = tile_begin_; ishift = ishift_; n_items = n_tiles << ishift_; stride = stride_; lane = lane_;
        sshift = split_chunks ? ishift_ : 0;
        jidx = split_chunks ? (first & ((1 << ishift_) - 1)) : 0;
        pk = first; seq = 0;
        p_rg = p_nx = make_int2(0, 0);
        p_cs = PIPE_CHUNK; p_off = 0;
        if (!tma) return;
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < PIPE_NST; ++s) mbar_init(&bars[s], 1);
            fence_mbar_init();
        }
        __syncwarp();
        if (pk < n_items) p_rg = load_range(pk);
        if (pk + stride < n_items) p_nx = load_range(pk + stride);
        p_cs = chunk_size(p_rg.y);
        p_off = jidx * p_cs;
#pragma unroll
        for (int s = 0; s < PIPE_NST; ++s) produce(s);
    }
    // stage of the chunk about to be consumed; waits until its copy has landed
    __device__ __forceinline__ float4 *acquire() {
        const int st = (int)(seq % PIPE_NST);
        mbar_wait(&bars[st], (seq / PIPE_NST) & 1u);
        return ring + (size_t)st * stage_f4;
    }
    // every lane has left the stage (the caller has just executed __syncwarp): refill it with the chunk PIPE_NST ahead
    __device__ __forceinline__ void release() {
        produce((int)(seq % PIPE_NST));
        ++seq;
    }
};

// Index-indirect gather of records [first, first + n) into a stage (the reference-typed entry points, no packed records).
__device__ __forceinline__ void gather_chunk(float4 *rec, const GatherSrc &src, int first, int n, int lane) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int t = 32 * h + lane;
        if (t < n) {
            const int g = __ldg(src.gids_sorted + first + t);
            const float x = __ldg(src.xys + 2 * (size_t)g), y = __ldg(src.xys + 2 * (size_t)g + 1);
            const float cx = __ldg(src.conics + 3 * (size_t)g), cy = __ldg(src.conics + 3 * (size_t)g + 1),
                        cz = __ldg(src.conics + 3 * (size_t)g + 2);
            const float r = __ldg(src.colors + 3 * (size_t)g), gg = __ldg(src.colors + 3 * (size_t)g + 1),
                        b = __ldg(src.colors + 3 * (size_t)g + 2);
            const float o = __ldg(src.opac + g);
            rec[3 * t + 0] = make_float4(x, y, o, cx);
            rec[3 * t + 1] = make_float4(cy, cz, r, gg);
            const float2 e = record_extent(cx, cy, cz, o);
            rec[3 * t + 2] = make_float4(b, __int_as_float(g), e.x, e.y);
        }
    }
    __syncwarp();
}

// Forward.  WARP-AUTONOMOUS, persistent: every warp of the grid is an independent tile pipeline -- no block-wide barrier
// anywhere in the kernel.  Tiles are dealt round-robin to CTAs first (tile k -> CTA k % grid, warp slot (k / grid) % 8), so
// every SM gets the same number of tiles whatever the tile count.
//   * lane <-> a 4x2 PIXEL BLOCK of the 16x16 tile (bx = lane & 3, by = lane >> 2): 8 pixels, 24 accumulators in registers;
//     the column terms (dx, cx*dx, cy*dx) are shared by the two rows and the row terms (dy, dy*(cz*dy)) by the four columns,
//     so a pair costs 13 FP32-pipe instructions instead of the 18 of a one-pixel thread -- and they are the SAME
//     operations in the same order as the reference's (gi2d_common.cuh: eval_sigma), only hoisted;
//   * the tile's records arrive in CHUNKS of 64 by bulk TMA copies into a per-warp ring of FW_NST stages; a chunk is in
//     flight while the previous one is evaluated, across tile boundaries (the producer cursor runs ahead of the consumer);
//   * per chunk, lane t derives the visible rectangle of record t and 32+t; 4 + 8 ballots over the column / row block bits
//     give EVERY lane the 64-bit mask of the records that touch ITS block (the rectangle is a product of a column range
//     and a row range), and the lane walks only those, in ascending order.  The per-pixel summation order is therefore
//     the reference's (ascending sorted index, forward.cu:566-614) and the image stays bit-identical, while the culling
//     granularity drops from 8x4 (59 pixel slots per record at 768x512 / 70k) to 4x2 (34; 21 are inside the rectangle).
//     Lanes run a common trip count (the warp maximum of the mask populations); a lane without work idles predicated off.
// Loss epilogue: with LOSS the kernel does not write the image; per pixel it forms what the caller would
// compute next --  image = clamp(out_img, 0, 1)  (gaussianimage_cholesky.py:84),  F.mse_loss(image, gt)  (utils.py:20-25) --
// and writes the gradient of that loss w.r.t. out_img,  v_out = 2/(3HW) (image - gt) [0 <= out_img <= 1],  straight into the
// buffer the backward rasterizer reads.  The squared error is summed per thread over the warp's tiles, reduced once per
// CTA, and the last CTA to finish adds the per-CTA partials in index order: a deterministic loss, no atomics on floats.
// Frame buffers of the OTHER ranks of a row-sharded render (peer-mapped device addresses, e.g. torch symmetric memory's
// buffer_ptrs): the epilogue repeats every full-tile row store to each of them over NVLink (plain st.global.v4).
struct PeerOut {
    float *p[7];
    int n;
};

struct LossArgs {
    const float *gt;          // target, (3,H,W) planes when gt_chw else (H,W,3)
    int gt_chw, clamp;
    int gt_vec;               // gt may be read with 16-byte loads (aligned base, W % 4 == 0)
    float inv_numel;          // 1 / (3 H W)
    float *v_out;             // (H,W,3): d loss / d out_img
    float *image;             // optional (H,W,3): the clamped image
    float *partials;          // (gridDim.x)
    unsigned int *ticket;     // zero before the launch; reset by the kernel
    float *loss;              // (1)
};

constexpr int FW_WARPS = 8;                       // warps per CTA = independent pipelines per CTA
constexpr int FW_CHUNK = PIPE_CHUNK;
constexpr int FW_THREADS = FW_WARPS * 32;
constexpr int FW_STAGE_F4 = (PIPE_CHUNK + 1) * 3;  // a stage = 64 records + one all-zero record (what a lane without work evaluates)
constexpr size_t FW_RING_BYTES = (size_t)FW_WARPS * PIPE_NST * FW_STAGE_F4 * 16;
constexpr size_t FW_OUT_BYTES = (size_t)FW_WARPS * GI2D_TILE_PIX * 3 * 4;        // one output tile per warp (epilogue staging)
constexpr size_t FW_SMEM = FW_RING_BYTES + (size_t)FW_WARPS * PIPE_NST * 8;
constexpr size_t FW_SMEM_EXCH = FW_SMEM + FW_OUT_BYTES;

// SPLIT = 1: a warp renders a whole tile, a lane a 4x2 pixel block (images with many tiles: 8K decode).
// SPLIT = 2: a warp renders HALF a tile (8 pixel rows), a lane a 2x2 block; the two halves of a tile go to different
//            warps (different CTAs), each streaming the tile's records itself -- twice the warps for the fitting images
//            (768x512 has only 1536 tiles for 148 SMs), the same instruction count (finer culling pays for the second
//            set of masks).
// EXCH: the frame buffers of other ranks are written too (NVLS multicast address or peer-mapped buffers): the output tile is
//       staged in shared memory and leaves as whole 192-byte rows.
template <bool TMA, bool AUX, bool LOSS, int SPLIT, bool EXCH = false>
__global__ void __launch_bounds__(FW_THREADS, 3) raster_fwd_kernel(int TX, int W, int H, int tile_begin, int n_tiles, int *tile_bins,
                                                                const float4 *__restrict__ packed, GatherSrc src,
                                                                const int *__restrict__ d_status,
                                                                const float *__restrict__ background, float *__restrict__ out_img,
                                                                float *__restrict__ final_Ts, int *__restrict__ final_idx,
                                                                float *out_mc, int vec_ok, LossArgs la = LossArgs{},
                                                                PeerOut peers = PeerOut{}) {
    GI2D_TRACE_DECL;
    GI2D_TRACE_ENTRY;
    constexpr int ISHIFT = SPLIT == 2 ? 1 : 0;
    constexpr int BW = 4 / SPLIT;                         // pixels per lane and row
    constexpr int XSH = SPLIT == 2 ? 1 : 2;               // log2(BW)
    constexpr int NBX = GI2D_TILE / BW, NBY = 8 / SPLIT;  // lane blocks across / down the warp's part of the tile (NBX * NBY = 32)
    constexpr int ROWS = GI2D_TILE / SPLIT;               // pixel rows of the warp's part
    extern __shared__ __align__(128) unsigned char fw_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = gridDim.x, stride = G * FW_WARPS;
    const int bxm = lane % NBX, bym = lane / NBX;         // this lane's pixel block
    const bool none_at_all = d_status && background && __ldg(d_status) == 0;   // rasterize_sum.py:121-129
    const int first = warp * G + blockIdx.x, n_items = n_tiles << ISHIFT;
    ChunkPipe pipe;
    float4 *const my_ring = reinterpret_cast<float4 *>(fw_smem) + (size_t)warp * PIPE_NST * FW_STAGE_F4;
    if (lane < 3 * PIPE_NST) my_ring[(lane / 3) * FW_STAGE_F4 + 3 * PIPE_CHUNK + lane % 3] = make_float4(0.f, 0.f, 0.f, 0.f);   // the zero records
    __syncwarp();
    pipe.init(tile_bins, packed, my_ring,
              reinterpret_cast<uint64_t *>(fw_smem + FW_RING_BYTES + (EXCH ? FW_OUT_BYTES : 0)) + warp * PIPE_NST, tile_begin, n_tiles,
              ISHIFT, false, first, stride, lane, TMA, FW_STAGE_F4);
    float4 *const s_out4 = reinterpret_cast<float4 *>(fw_smem + FW_RING_BYTES) + (size_t)warp * (GI2D_TILE_PIX * 3 / 4);   // EXCH only
    const int part = first & (SPLIT - 1);                 // which half of its tiles this warp renders (constant: stride is even)

    float sq_sum = 0.f;                                    // LOSS: this thread's share of the squared error
    int q = first;
    int2 c_nx = q < n_items ? pipe.load_range(q) : make_int2(0, 0);
    for (; q < n_items; q += stride) {
        const int2 rg = c_nx;
        if (q + stride < n_items) c_nx = pipe.load_range(q + stride);
        const int tile = tile_begin + (q >> ISHIFT);
        const int ty = tile / TX, tx = tile - ty * TX;
        const int tx0 = tx * GI2D_TILE, ty0 = ty * GI2D_TILE;
        const int lo = rg.x, cnt = rg.y;
        if (TMA && cnt == 0 && lane == 0 && part == 0) reinterpret_cast<int2 *>(tile_bins)[tile] = make_int2(0, 0);   // canonical (0, 0) like the reference's zero fill
        const int row0 = ROWS * part;                      // first tile row of this warp's part
        const float pxf = (float)(tx0 + BW * bxm), pyf = (float)(ty0 + row0 + 2 * bym);

        float acc[2][BW][3];
        int cur[2][BW];
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int c = 0; c < BW; ++c) {
                acc[r][c][0] = acc[r][c][1] = acc[r][c][2] = 0.f;
                cur[r][c] = 0;
            }
        float T_final = 1.f;                               // T is never updated in the sum rasterizer (Q6)

        for (int off = 0; off < cnt; off += FW_CHUNK) {
            const int n = min(FW_CHUNK, cnt - off);
            float4 *rec;
            if (TMA) {
                rec = pipe.acquire();
            } else {                                       // reference-typed entry: index-indirect gather by the warp itself
                rec = pipe.ring;
                gather_chunk(rec, src, lo + off, n, lane);
            }
            GI2D_TRACE_FIRST;
            GI2D_TRACE_WORK(n);
            // which records of the chunk touch this lane's pixel block: 64-bit mask from 2 x (NBX + NBY = 12) ballots
            uint32_t m[2] = {0u, 0u};
            bool big = false;                              // a record that needs the careful walk: opacity not <= 1 (min(1, .) of forward.cu:601) ...
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h == 1 && n <= 32) break;              // sparse tiles (8K decode: ~19 records per tile): one round of ballots
                const int t = 32 * h + lane;
                uint32_t colbits = 0, rowbits = 0;
                if (t < n) {
                    int lx0, lx1, ly0, ly1;
                    const float4 A = rec[3 * t];
                    const float4 Bq = rec[3 * t + 1];
                    // ... or whose fields could make sigma NaN (non-finite, or large enough to overflow): min(1, NaN) = 1 there
                    // ... or that carries no extent (+inf: visible_extent's guard D > 2e-3 cx cz, cx, cz > 0 failed).  With the guard,
                    // sigma >= 1e-3 * (0.5 cx dx^2 + 0.5 cz dy^2), three orders above its fp32 rounding error: `sigma < 0` is
                    // impossible and the fast walk does not test it.
                    const float4 Cq = rec[3 * t + 2];
                    big = big || !(A.z <= 1.f) || !(fabsf(A.x) < 1e6f) || !(fabsf(A.y) < 1e6f) || !(fabsf(A.w) < 1e18f) ||
                          !(fabsf(Bq.x) < 1e18f) || !(fabsf(Bq.y) < 1e18f) || !(Cq.z < 3.0e38f);
                    if (record_rect(A, Cq, tx0, ty0, lx0, lx1, ly0, ly1)) {
                        const int r0 = max(ly0, row0) - row0, r1 = min(ly1, row0 + ROWS - 1) - row0;
                        if (r0 <= r1) {
                            colbits = (2u << (lx1 >> XSH)) - (1u << (lx0 >> XSH));
                            rowbits = (2u << (r1 >> 1)) - (1u << (r0 >> 1));
                        }
                    }
                }
                uint32_t myC = 0, myR = 0;
#pragma unroll
                for (int j = 0; j < NBX; ++j) {
                    const uint32_t b = __ballot_sync(0xffffffffu, (colbits >> j) & 1u);
                    myC = bxm == j ? b : myC;
                }
#pragma unroll
                for (int j = 0; j < NBY; ++j) {
                    const uint32_t b = __ballot_sync(0xffffffffu, (rowbits >> j) & 1u);
                    myR = bym == j ? b : myR;
                }
                m[h] = myC & myR;
            }
            const int n_it = __reduce_max_sync(0xffffffffu, __popc(m[0]) + __popc(m[1]));
            const bool any_big = __any_sync(0xffffffffu, big);
            const int idx0 = lo + off;
            // The walk.  A lane whose mask has run out evaluates the stage's all-zero record (sigma = 0, alpha = 0: rejected by
            // the alpha test like any invisible pair), so the body needs no "lane active" predicate.  The halved conic terms
            // are exact (a factor 0.5 commutes with every rounding of eval_sigma), so  ffma(dy, w, ffma(dx, hu, ht3))  is
            // bit for bit  ffma(dy, w, 0.5 * ffma(dx, u, t3)).  With every opacity of the chunk <= 1 (always, in GaussianImage),
            // opac * exp(-sigma) <= opac <= 1 for every pair that passes sigma >= 0, and the min(1, .) is dropped.
            auto walk = [&](auto clamp_tag) {
                constexpr bool CLAMP1 = decltype(clamp_tag)::value;
                uint32_t m0 = m[0], m1 = m[1];
                int base = 0;
                for (int it = 0; it < n_it; ++it) {
                    const bool z0 = m0 == 0u;
                    m0 = z0 ? m1 : m0;
                    m1 = z0 ? 0u : m1;
                    base = z0 ? 32 : base;
                    const int t = m0 != 0u ? __ffs((int)m0) - 1 + base : PIPE_CHUNK;
                    m0 &= m0 - 1u;
                    const float4 A = rec[3 * t], B = rec[3 * t + 1];
                    const float cb = rec[3 * t + 2].x;
                    const float hcx = fmul(A.w, 0.5f), hcz = fmul(B.y, 0.5f);
                    float dx[BW], hu[BW], w[BW];
#pragma unroll
                    for (int c = 0; c < BW; ++c) {
                        dx[c] = fsub(A.x, pxf + (float)c);
                        hu[c] = fmul(hcx, dx[c]);               // 0.5 * (cx * dx)
                        w[c] = fmul(B.x, dx[c]);                // cy * dx
                    }
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        const float dy = fsub(A.y, pyf + (float)r);
                        const float ht3 = fmul(dy, fmul(hcz, dy));                         // 0.5 * (dy * (cz * dy))
#pragma unroll
                        for (int c = 0; c < BW; ++c) {
                            const float sigma = ffma(dy, w[c], ffma(dx[c], hu[c], ht3));    // eval_sigma, hoisted
                            const float ov = fmul(A.z, exp_neg(sigma));
                            const float alpha = CLAMP1 ? fminf(ov, 1.f) : ov;
                            if ((!CLAMP1 || !(sigma < 0.f)) && !(alpha < GI2D_ALPHA_MIN)) { // forward.cu:602-604
                                acc[r][c][0] = ffma(alpha, B.z, acc[r][c][0]);             // forward.cu:607-609
                                acc[r][c][1] = ffma(alpha, B.w, acc[r][c][1]);
                                acc[r][c][2] = ffma(alpha, cb, acc[r][c][2]);
                                if (AUX) cur[r][c] = idx0 + t;                              // forward.cu:611
                            }
                        }
                    }
                }
            };
            if (any_big) walk(std::true_type{}); else walk(std::false_type{});
            __syncwarp();                                  // every lane has left the stage
            if (TMA) pipe.release();                       // refill it with the chunk PIPE_NST ahead
        }
        if (cnt == 0 && none_at_all) {                     // no intersections at all: ones * background, final_Ts = zeros
            const float b0 = __ldg(background), b1 = __ldg(background + 1), b2 = __ldg(background + 2);
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int c = 0; c < BW; ++c) { acc[r][c][0] = b0; acc[r][c][1] = b1; acc[r][c][2] = b2; }
            T_final = 0.f;
        }
        // ---- epilogue: the lane owns BW consecutive pixels of two rows = two runs of 12 * BW contiguous bytes.  With `out_mc`
        // (the NVLS multicast alias of a symmetric frame buffer) every store is a `multimem.st`, which the NVSwitch replicates
        // into the frame buffer of EVERY rank -- the all-gather of a row-sharded render happens inside the rasterizer.
        float *const dst_img = LOSS ? la.v_out : out_img;
        const bool vec_tile = vec_ok && tx0 + GI2D_TILE <= W && ty0 + GI2D_TILE <= H;     // uniform over the warp
        constexpr bool staged = EXCH;
        // LOSS: the target's values for the lane's BW pixels of a row are loaded together BEFORE their first use (as 16- / 8-byte
        // loads on full tiles): element by element between the arithmetic they were a chain of 24 dependent L2 round trips at
        // the end of every tile -- 4.5 us per tile, and at the fitting sizes a warp's tile IS the kernel's critical path.
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int i = ty0 + row0 + 2 * bym + r, j0 = tx0 + BW * bxm;
            const size_t pix0 = (size_t)i * W + j0;
            float gtv[LOSS ? 3 * BW : 1];
            if (LOSS) {
                const size_t plane = (size_t)H * W;
                if (vec_tile && la.gt_vec) {
                    if (la.gt_chw) {
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch) {
                            if (BW == 4) {
                                const float4 g = __ldg(reinterpret_cast<const float4 *>(la.gt + ch * plane + pix0));
                                gtv[ch] = g.x; gtv[3 + ch] = g.y; gtv[(6 + ch) % (3 * BW)] = g.z; gtv[(9 + ch) % (3 * BW)] = g.w;
                            } else {
                                const float2 g = __ldg(reinterpret_cast<const float2 *>(la.gt + ch * plane + pix0));
                                gtv[ch] = g.x; gtv[3 + ch] = g.y;
                            }
                        }
                    } else {
                        if (BW == 4) {
#pragma unroll
                            for (int k4 = 0; k4 < 3; ++k4) {
                                const float4 g = __ldg(reinterpret_cast<const float4 *>(la.gt + 3 * pix0) + k4);
                                gtv[(4 * k4) % (3 * BW)] = g.x; gtv[(4 * k4 + 1) % (3 * BW)] = g.y;
                                gtv[(4 * k4 + 2) % (3 * BW)] = g.z; gtv[(4 * k4 + 3) % (3 * BW)] = g.w;
                            }
                        } else {
#pragma unroll
                            for (int k2 = 0; k2 < 3; ++k2) {
                                const float2 g = __ldg(reinterpret_cast<const float2 *>(la.gt + 3 * pix0) + k2);
                                gtv[2 * k2] = g.x; gtv[2 * k2 + 1] = g.y;
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < BW; ++c)
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch) {
                            const bool inside = i < H && j0 + c < W;
                            const float *ptr = la.gt_chw ? la.gt + ch * plane + pix0 + c : la.gt + 3 * (pix0 + c) + ch;
                            gtv[3 * c + ch] = inside ? __ldg(ptr) : 0.f;
                        }
                }
            }
            float v[3 * BW];
#pragma unroll
            for (int c = 0; c < BW; ++c) { v[3 * c] = acc[r][c][0]; v[3 * c + 1] = acc[r][c][1]; v[3 * c + 2] = acc[r][c][2]; }
            if (LOSS) {
                const float sc = 2.f * la.inv_numel;
                float cl[3 * BW];
#pragma unroll
                for (int c = 0; c < BW; ++c) {
                    const bool inside = i < H && j0 + c < W;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        const float raw = v[3 * c + ch];
                        cl[3 * c + ch] = la.clamp ? fminf(fmaxf(raw, 0.f), 1.f) : raw;
                        const float d = inside ? cl[3 * c + ch] - gtv[3 * c + ch] : 0.f;
                        sq_sum += d * d;
                        // torch.clamp passes the gradient where min <= x <= max
                        v[3 * c + ch] = (!la.clamp || (raw >= 0.f && raw <= 1.f)) ? sc * d : 0.f;
                    }
                }
                if (la.image) {                             // the clamped image (optional output)
                    if (vec_tile && la.gt_vec && BW == 4) {
#pragma unroll
                        for (int k4 = 0; k4 < 3; ++k4)
                            *reinterpret_cast<float4 *>(la.image + 3 * pix0 + 4 * k4) = make_float4(cl[4 * k4], cl[4 * k4 + 1], cl[4 * k4 + 2], cl[4 * k4 + 3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < BW; ++c)
                            if (i < H && j0 + c < W) {
#pragma unroll
                                for (int ch = 0; ch < 3; ++ch) la.image[3 * (pix0 + c) + ch] = cl[3 * c + ch];
                            }
                    }
                }
            }
            if (vec_tile && staged) {
                // frame buffers of other ranks are written too (multicast / peer stores): stage the lane's run into the warp's
                // output tile (rows of 48 floats); the tile then leaves as whole 192-byte rows, 12 consecutive lanes on 12
                // consecutive float4 -- full-size NVLink write packets instead of scattered 16-byte ones
                float *const so = reinterpret_cast<float *>(s_out4) + (row0 + 2 * bym + r) * (GI2D_TILE * 3) + 3 * BW * bxm;
                if (BW == 4) {
#pragma unroll
                    for (int k4 = 0; k4 < 3; ++k4)
                        reinterpret_cast<float4 *>(so)[k4] = make_float4(v[4 * k4], v[4 * k4 + 1], v[4 * k4 + 2], v[4 * k4 + 3]);
                } else {
#pragma unroll
                    for (int k2 = 0; k2 < 3; ++k2) reinterpret_cast<float2 *>(so)[k2] = make_float2(v[2 * k2], v[2 * k2 + 1]);
                }
            } else if (vec_tile) {                         // local image only: the lane's 12 * BW contiguous bytes, straight from registers
                if (BW == 4) {
#pragma unroll
                    for (int k4 = 0; k4 < 3; ++k4)
                        *reinterpret_cast<float4 *>(dst_img + 3 * pix0 + 4 * k4) = make_float4(v[4 * k4], v[4 * k4 + 1], v[4 * k4 + 2], v[4 * k4 + 3]);
                } else {
#pragma unroll
                    for (int k2 = 0; k2 < 3; ++k2)
                        *reinterpret_cast<float2 *>(dst_img + 3 * pix0 + 2 * k2) = make_float2(v[2 * k2], v[2 * k2 + 1]);
                }
            } else if (i < H) {
#pragma unroll
                for (int c = 0; c < BW; ++c)
                    if (j0 + c < W) {
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch) {
                            if (out_mc) multimem_st(out_mc + 3 * (pix0 + c) + ch, v[3 * c + ch]);
                            else dst_img[3 * (pix0 + c) + ch] = v[3 * c + ch];
                        }
                    }
            }
            if (AUX && i < H) {
#pragma unroll
                for (int c = 0; c < BW; ++c)
                    if (j0 + c < W) {
                        final_Ts[pix0 + c] = T_final;
                        final_idx[pix0 + c] = cur[r][c];
                    }
            }
        }
        if (vec_tile && staged) {
            __syncwarp();
            constexpr int F4_ROW = GI2D_TILE * 3 / 4, F4_PART = ROWS * F4_ROW;            // 12 float4 per row; 192 / 96 per part
#pragma unroll
            for (int jj = 0; jj < F4_PART / 32; ++jj) {
                const int idx = jj * 32 + lane, rr = idx / F4_ROW, c4 = idx - rr * F4_ROW;
                const float4 o4 = s_out4[row0 * F4_ROW + idx];
                const size_t off = ((size_t)(ty0 + row0 + rr) * W + tx0) * 3 + 4 * c4;
                if (out_mc) multimem_st_v4(out_mc + off, o4);
                else *reinterpret_cast<float4 *>(dst_img + off) = o4;
                for (int pk = 0; pk < peers.n; ++pk) __stcs(reinterpret_cast<float4 *>(peers.p[pk] + off), o4);
            }
            __syncwarp();                                  // the staging tile is free for the next tile
        }
    }
    GI2D_TRACE_EXIT(blockIdx.x * FW_WARPS + warp);
    if (LOSS) {
        __shared__ float s_red[FW_WARPS];
        __shared__ bool s_last;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) sq_sum += __shfl_xor_sync(0xffffffffu, sq_sum, d);
        if (lane == 0) s_red[warp] = sq_sum;
        __syncthreads();
        if (tid == 0) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < FW_WARPS; ++w) t += s_red[w];
            la.partials[blockIdx.x] = t;
            __threadfence();
            s_last = atomicAdd(la.ticket, 1u) == gridDim.x - 1;
        }
        __syncthreads();
        if (s_last && warp == 0) {                        // last CTA: fixed-order sum of the partials
            __threadfence();
            float t = 0.f;
            for (int c = lane; c < (int)gridDim.x; c += 32) t += *reinterpret_cast<volatile float *>(la.partials + c);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
            if (lane == 0) {
                *la.loss = t * la.inv_numel;
                *la.ticket = 0;
            }
        }
    }
}

// Backward.  WARP-AUTONOMOUS like the forward: a warp takes a tile's records in chunks of 64 through the same TMA ring, and
// no block-wide barrier exists.  Work unit = one ALIGNED 4x2 PIXEL BLOCK touched by one staged record's visible
// rectangle (column terms shared by the two rows, row terms by the four columns, as in the forward).  The units of all
// records of a chunk form one flattened list (record-major, row-major inside a record); lane l owns the CONTIGUOUS piece
// [l*R, l*R+R) of it (R = ceil(units/32)), so every lane has the same number of eight-pixel evaluations whatever the
// records' footprints are, and a lane's consecutive units mostly belong to one record: its partial sums stay in registers
// and are flushed only when the piece crosses into the next record -- with nine PLAIN shared-memory stores into slot
// (lane + record): along the flattened list the (lane, record) pairs form a monotone staircase, so that sum is unique per
// flush (shared-memory float atomics are CAS loops on sm_100: `ATOMS.CAST.SPIN`).  No shuffle reductions, no shared
// atomics; after a __syncwarp the lane that staged record t adds up the slots of the lanes [pref[t]/R, (pref[t+1]-1)/R]
// whose pieces intersect it, turns the sums into the nine gradients and issues ONE set of global reductions per
// (tile chunk, Gaussian): two 16-byte `red.global.add.v4.f32` + one scalar in the packed layout.
//
// Sums kept per record (m = vis * v_alpha for a contributing pixel, 0 otherwise; dx, dy as in backward.cu:797-799):
//   S_m = sum m, S_x = sum m dx, S_y = sum m dy, S_xx = sum m dx dx, S_xy = sum m dx dy, S_yy = sum m dy dy, and
//   sum alpha * v_out per channel.  With v_sigma = -opac * m (backward.cu:822) the reference's per-pixel terms
//   (:823-828) factor into  v_conic = -0.5 opac (S_xx, S_xy, S_yy),  v_xy = -opac (cx S_x + cy S_y, cy S_x + cz S_y),
//   v_opacity = S_m  -- the same fp32 sums re-associated, well inside the 1e-4 gradient tolerance.
__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

struct GradOut {            // VEC12: g12 = (N,12) rows {v_xy.x, v_xy.y, v_conic.xyz, v_colors.rgb, v_opacity, 0, 0, 0}
    float *g12;
    float *v_xy, *v_conic, *v_colors, *v_opacity;
};

constexpr int BW_WARPS = 8;
constexpr int BW_THREADS = BW_WARPS * 32;
struct __align__(16) BwdWarpSmem {
    float4 ring[PIPE_NST][PIPE_CHUNK * 3];        // staged records
    float4 v4[GI2D_TILE_PIX * 3 / 4];             // upstream gradient of the tile, interleaved like the image: [16 rows][16 px][rgb]
    uint16_t pref[PIPE_CHUNK + 2];                // exclusive prefix of the unit counts; [PIPE_CHUNK] = total
    uint8_t rect[PIPE_CHUNK];                     // bx0 | nbx<<2 | by0<<5: the 4x2 pixel blocks a record's rectangle touches
    uint64_t bars[PIPE_NST];
};
constexpr size_t BW_SMEM = sizeof(BwdWarpSmem) * BW_WARPS;

// ishift: a tile is 1 << ishift work items (warps); item j of a tile takes its chunks j, j + (1 << ishift), ... -- the per-record
// sums of different chunks are independent and the gradients meet in global memory, so the split needs no communication;
// it doubles / quadruples the warps in flight for the fitting images, whose 1536-2752 tiles do not fill 148 SMs.
template <bool TMA, bool VEC12>
__global__ void __launch_bounds__(BW_THREADS, 3) raster_bwd_kernel(int TX, int W, int H, int n_tiles, int ishift, const int *__restrict__ tile_bins,
                                                                const float4 *__restrict__ packed, GatherSrc src,
                                                                const float *__restrict__ v_out, GradOut out, int vec_ok,
                                                                const float *__restrict__ v_scale = nullptr) {
    GI2D_TRACE_DECL;
    GI2D_TRACE_ENTRY;
    extern __shared__ __align__(128) unsigned char bw_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    BwdWarpSmem &S = reinterpret_cast<BwdWarpSmem *>(bw_smem)[warp];
    const int G = gridDim.x, stride = G * BW_WARPS;
    ChunkPipe pipe;
    const int first = warp * G + blockIdx.x, n_items = n_tiles << ishift;
    pipe.init(tile_bins, packed, &S.ring[0][0], S.bars, 0, n_tiles, ishift, true, first, stride, lane, TMA);
    const float vsc = v_scale ? __ldg(v_scale) : 1.f;     // upstream gradient of a scalar loss whose d/d out_img is already in v_out

    int q = first;
    int2 c_nx = q < n_items ? pipe.load_range(q) : make_int2(0, 0);
    for (; q < n_items; q += stride) {
        const int2 rg = c_nx;
        if (q + stride < n_items) c_nx = pipe.load_range(q + stride);
        const int lo = rg.x, cnt = rg.y;
        const int cs = pipe.chunk_size(cnt);               // records per chunk of this tile (equal chunks)
        if (cnt <= pipe.jidx * cs) continue;               // nothing for this item in the tile
        const int k = q >> ishift;
        const int ty = k / TX, tx = k - ty * TX;
        const int tx0 = tx * GI2D_TILE, ty0 = ty * GI2D_TILE;
        // the tile of upstream gradients in the image's own interleaved layout (a straight copy of 16 rows x 12 float4), zero
        // outside the image (such pixels then contribute exact zeros)
        if (vec_ok && tx0 + GI2D_TILE <= W && ty0 + GI2D_TILE <= H) {
#pragma unroll
            for (int q = 0; q < 6; ++q) {
                const int f = q * 32 + lane, row = f / 12, c4 = f - row * 12;
                float4 g4 = __ldg(reinterpret_cast<const float4 *>(v_out + ((size_t)(ty0 + row) * W + tx0) * 3) + c4);
                g4.x *= vsc; g4.y *= vsc; g4.z *= vsc; g4.w *= vsc;
                S.v4[f] = g4;
            }
        } else {
            float *const sv = reinterpret_cast<float *>(S.v4);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int p = q * 32 + lane, j = tx0 + (p & 15), i = ty0 + (p >> 4);
                float vr = 0.f, vg = 0.f, vb = 0.f;
                if (i < H && j < W) {
                    const size_t pix = (size_t)i * W + j;
                    vr = __ldg(v_out + 3 * pix) * vsc; vg = __ldg(v_out + 3 * pix + 1) * vsc; vb = __ldg(v_out + 3 * pix + 2) * vsc;
                }
                sv[3 * p] = vr; sv[3 * p + 1] = vg; sv[3 * p + 2] = vb;
            }
        }
        // (visibility of S.v4 to the other lanes: the __syncwarp after the prefix stores below)

        for (int off = pipe.jidx * cs; off < cnt; off += cs << pipe.sshift) {
            const int n = min(cs, cnt - off);
            float4 *rec;
            if (TMA) {
                rec = pipe.acquire();
            } else {
                rec = pipe.ring;
                gather_chunk(rec, src, lo + off, n, lane);
            }
            GI2D_TRACE_FIRST;
            GI2D_TRACE_WORK(n);
            // per record: visible rectangle -> the aligned 4x2 pixel blocks it touches = units; warp-wide exclusive scan of the counts
            int units[2], ubeg[2];
            int run = 0;
            bool big = false;                              // a record whose alpha needs the min(1, .) of backward.cu:805 (see the forward)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int t = 32 * h + lane;
                int lx0 = 0, lx1 = 0, ly0 = 0, ly1 = 0, un = 0;
                uint32_t rect = 0;
                if (t < n) {
                    const float4 A = rec[3 * t], Bq = rec[3 * t + 1];
                    big = big || !(A.z <= 1.f) || !(fabsf(A.x) < 1e6f) || !(fabsf(A.y) < 1e6f) || !(fabsf(A.w) < 1e18f) ||
                          !(fabsf(Bq.x) < 1e18f) || !(fabsf(Bq.y) < 1e18f) || !(rec[3 * t + 2].z < 3.0e38f);
                }
                if (t < n && record_rect(rec[3 * t], rec[3 * t + 2], tx0, ty0, lx0, lx1, ly0, ly1)) {
                    const int bx0 = lx0 >> 2, nbx = (lx1 >> 2) - bx0 + 1, by0 = ly0 >> 1, nby = (ly1 >> 1) - by0 + 1;
                    un = nbx * nby;
                    rect = (uint32_t)bx0 | ((uint32_t)nbx << 2) | ((uint32_t)by0 << 5);
                }
                S.rect[t] = (uint8_t)rect;
                int incl = un;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                units[h] = un;
                ubeg[h] = run + incl - un;
                S.pref[t] = (uint16_t)ubeg[h];
                run += __shfl_sync(0xffffffffu, incl, 31);
            }
            const int total = run;
            if (lane == 0) S.pref[PIPE_CHUNK] = (uint16_t)total;
            const bool any_big = __any_sync(0xffffffffu, big);
            __syncwarp();
            if (total > 0) {
                // ---- main loop: this lane's contiguous piece of the flattened unit list
                const int R = (total + 31) >> 5;
                int u = lane * R;
                const int uend = min(total, u + R);
                if (u < uend) {
                    // record containing unit u: the largest t with pref[t] <= u (records without units share their successor's prefix)
                    int t = 0;
#pragma unroll
                    for (int step = PIPE_CHUNK / 2; step > 0; step >>= 1)
                        if ((int)S.pref[t + step] <= u) t += step;
                    float S_m = 0.f, S_x = 0.f, S_y = 0.f, S_xx = 0.f, S_xy = 0.f, S_yy = 0.f, a_r = 0.f, a_g = 0.f, a_b = 0.f;
                    // current record (registers) and its block range
                    float x, y, opac, hcx, cy, hcz, cr, cg, cb;      // hcx, hcz: HALF the conic's diagonal (0.5 cx dx dx + 0.5 cz dy dy: exact)
                    int bx0, nbx, by0, base, rec_end, rmul;       // rmul: kk / nbx == (kk * rmul) >> 7 for kk < 128 (nbx in 1..4)
                    size_t gid;
                    auto load_record = [&]() {
                        const float4 A = rec[3 * t], B = rec[3 * t + 1];
                        const float2 C = *reinterpret_cast<const float2 *>(&rec[3 * t + 2]);
                        cb = C.x; gid = (size_t)__float_as_int(C.y);
                        x = A.x; y = A.y; opac = A.z; hcx = fmul(A.w, 0.5f); cy = B.x; hcz = fmul(B.y, 0.5f); cr = B.z; cg = B.w;
                        const uint32_t rect = S.rect[t];
                        bx0 = rect & 3; nbx = (rect >> 2) & 7; by0 = rect >> 5;
                        rmul = nbx == 3 ? 43 : 128 >> (nbx >> 1);                       // 128, 64, 43, 32
                        base = (int)S.pref[t];
                        rec_end = (int)S.pref[t + 1];
                    };
                    // One flat loop over the piece; crossing into the next record is a short divergent branch (flush, find, load),
                    // the eight-pixel body stays converged across the warp.  CLAMP1 = false (every opacity of the chunk <= 1, so
                    // alpha = opac * vis needs no min): the colour sums run over vis instead of alpha and take the factor opac in
                    // the flush -- one select per pixel (vis or 0) instead of two.
                    auto piece = [&](auto clamp_tag) {
                        constexpr bool CLAMP1 = decltype(clamp_tag)::value;
                        // this lane's share of the record's gradients -> ONE set of global reductions (linear in the sums, so
                        // the shares of the lanes that split a record simply add up in memory like those of other tiles do)
                        auto flush = [&]() {
                            const float nop = -opac, hop = -0.5f * opac;              // v_sigma = -opac * m ; the conic terms carry 0.5 (Q7)
                            const float cx = hcx + hcx, cz = hcz + hcz;
                            const float g_x = nop * fmaf(cx, S_x, cy * S_y), g_y = nop * fmaf(cy, S_x, cz * S_y);   // backward.cu:826-827
                            const float g_cxx = hop * S_xx, g_cxy = hop * S_xy, g_cyy = hop * S_yy;                // :823-825
                            const float csc = CLAMP1 ? 1.f : opac;
                            const float o_r = csc * a_r, o_g = csc * a_g, o_b = csc * a_b;
                            if (VEC12) {
                                float *row = out.g12 + 12 * gid;
                                red_add_v4(row, g_x, g_y, g_cxx, g_cxy);
                                red_add_v4(row + 4, g_cyy, o_r, o_g, o_b);
                                atomicAdd(row + 8, S_m);
                            } else {
                                atomicAdd(out.v_xy + 2 * gid, g_x); atomicAdd(out.v_xy + 2 * gid + 1, g_y);
                                atomicAdd(out.v_conic + 3 * gid, g_cxx); atomicAdd(out.v_conic + 3 * gid + 1, g_cxy); atomicAdd(out.v_conic + 3 * gid + 2, g_cyy);
                                atomicAdd(out.v_colors + 3 * gid, o_r); atomicAdd(out.v_colors + 3 * gid + 1, o_g); atomicAdd(out.v_colors + 3 * gid + 2, o_b);
                                atomicAdd(out.v_opacity + gid, S_m);
                            }
                        };
                        for (; u < uend; ++u) {
                            if (u >= rec_end) {
                                flush();
                                S_m = S_x = S_y = S_xx = S_xy = S_yy = a_r = a_g = a_b = 0.f;
                                do { ++t; } while ((int)S.pref[t + 1] <= u);        // next record that has units
                                load_record();
                            }
                            const int kk = u - base;
                            const int rowb = (kk * rmul) >> 7;
                            const int bx = bx0 + kk - rowb * nbx, ly0 = 2 * (by0 + rowb);
                            const float fx0 = (float)(tx0 + 4 * bx);
                            float dx[4], hu[4], cw[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                dx[j] = fsub(x, fx0 + (float)j);
                                hu[j] = fmul(hcx, dx[j]);
                                cw[j] = fmul(cy, dx[j]);
                            }
                            // Pixels of the block outside the record's rectangle need no test of their own: the rectangle is a
                            // conservative bound of where the alpha test can pass, so that test rejects them (as in the forward).
#pragma unroll
                            for (int r = 0; r < 2; ++r) {
                                const int ly = ly0 + r;
                                const float dy = fsub(y, (float)(ty0 + ly));
                                const float ht3 = fmul(dy, fmul(hcz, dy));
                                const float4 e0 = S.v4[ly * 12 + bx * 3], e1 = S.v4[ly * 12 + bx * 3 + 1], e2 = S.v4[ly * 12 + bx * 3 + 2];
                                const float vr[4] = {e0.x, e0.w, e1.z, e2.y}, vg[4] = {e0.y, e1.x, e1.w, e2.z}, vb[4] = {e0.z, e1.y, e2.x, e2.w};
                                float u_m = 0.f, u_x = 0.f;          // this row's sum m and sum m dx (dy is constant over the row)
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    const float sigma = ffma(dy, cw[j], ffma(dx[j], hu[j], ht3));
                                    const float vis = exp_neg(sigma);
                                    const float ov = fmul(opac, vis);
                                    const float raw = CLAMP1 ? fminf(ov, 1.f) : ov;
                                    const bool valid = (!CLAMP1 || !(sigma < 0.f)) && !(raw < GI2D_ALPHA_MIN);                  // backward.cu:806-808
                                    const float v_alpha = fmaf(cr, vr[j], fmaf(cg, vg[j], cb * vb[j]));                          // :818-820
                                    float m;
                                    if (CLAMP1) {
                                        const float alpha = valid ? raw : 0.f;
                                        a_r = fmaf(alpha, vr[j], a_r); a_g = fmaf(alpha, vg[j], a_g); a_b = fmaf(alpha, vb[j], a_b);   // backward.cu:814
                                        m = valid ? vis * v_alpha : 0.f;
                                    } else {
                                        const float vis0 = valid ? vis : 0.f;
                                        a_r = fmaf(vis0, vr[j], a_r); a_g = fmaf(vis0, vg[j], a_g); a_b = fmaf(vis0, vb[j], a_b);
                                        m = vis0 * v_alpha;
                                    }
                                    const float mdx = m * dx[j];
                                    u_m += m;
                                    u_x += mdx;
                                    S_xx = fmaf(mdx, dx[j], S_xx);
                                }
                                S_m += u_m; S_x += u_x;
                                const float mdy = u_m * dy;
                                S_y += mdy;
                                S_yy = fmaf(mdy, dy, S_yy);
                                S_xy = fmaf(u_x, dy, S_xy);
                            }
                        }
                        flush();
                    };
                    load_record();
                    if (any_big) piece(std::true_type{}); else piece(std::false_type{});
                }
            }
            __syncwarp();                                  // every lane has left the stage, rect / pref are free again
            if (TMA) pipe.release();
        }
    }
    GI2D_TRACE_EXIT(blockIdx.x * BW_WARPS + warp);
}

static inline int cuda_status() {
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? GI2D_OK : GI2D_ERR_CUDA_BASE + (int)e;
}

// float4 image stores need 16-byte aligned rows: W % 4 == 0 and an aligned base (nullptr counts as aligned).
static inline int vec_ok(int W, const void *base) { return (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0); }

// Persistent grids: as many CTAs as are co-resident (SMs x occupancy), never more than there are tiles (tiles are dealt to
// CTAs first, then to the warps of a CTA, so a grid larger than tiles / 8 only leaves warp slots empty).
template <typename K>
static int persistent_grid(K kernel, int threads, size_t smem, int n_tiles, const char *env_cap) {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    static const void *keys[16];
    static int vals[16], used = 0;
    const void *key = reinterpret_cast<const void *>(kernel);
    int per_sm = 0;
    for (int i = 0; i < used; ++i)
        if (keys[i] == key) per_sm = vals[i];
    if (!per_sm) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
        const char *v = getenv(env_cap);                  // experiments: GI2D_FWD_CTAS_PER_SM / GI2D_BWD_CTAS_PER_SM
        if (v && atoi(v) > 0 && per_sm > atoi(v)) per_sm = atoi(v);
        if (used < 16) { keys[used] = key; vals[used] = per_sm; ++used; }
        cudaGetLastError();
    }
    const int g = sms * per_sm;
    return n_tiles < g ? (n_tiles < 1 ? 1 : n_tiles) : g;
}
template <typename K>
static int fwd_grid(K kernel, int n_items) { return persistent_grid(kernel, FW_THREADS, FW_SMEM, n_items, "GI2D_FWD_CTAS_PER_SM"); }
// Warps per tile of the forward rasterizer.  One: two half-tile warps double the warps in flight but rebuild the masks twice
// (+21 % instructions) and were not faster on any configuration (B200, CUPTI: 768x512 / 70k 15.3 vs 15.9 us, 1020x678 / 100k
// 19.8 vs 21.8 us, 8K / 1M 206 vs 276 us); GI2D_FWD_SPLIT=2 keeps the variant reachable for experiments.
static int fwd_split(int n_tiles) {
    static const int forced = getenv("GI2D_FWD_SPLIT") ? atoi(getenv("GI2D_FWD_SPLIT")) : 0;
    (void)n_tiles;
    return forced == 2 ? 2 : 1;
}
struct FwdArgs {
    int TX, W, H, tile_begin, n_tiles;
    int *tile_bins;
    const float4 *packed;
    GatherSrc src;
    const int *d_status;
    const float *background;
    float *out_img, *final_Ts;
    int *final_idx;
    float *out_mc;
    int vec_ok;
    LossArgs la;
    PeerOut peers;
};
template <bool TMA, bool AUX, bool LOSS>
static int fwd_grid_for(int n_tiles) {
    return fwd_split(n_tiles) == 2 ? fwd_grid(raster_fwd_kernel<TMA, AUX, LOSS, 2>, 2 * n_tiles) : fwd_grid(raster_fwd_kernel<TMA, AUX, LOSS, 1>, n_tiles);
}
// the row-sharded renderer's variant with exchange stores (one warp per tile)
static void launch_fwd_exch(const FwdArgs &a, cudaStream_t stream) {
    auto k = raster_fwd_kernel<true, false, false, 1, true>;
    const int grid = persistent_grid(k, FW_THREADS, FW_SMEM_EXCH, a.n_tiles, "GI2D_FWD_CTAS_PER_SM");
    k<<<grid, FW_THREADS, FW_SMEM_EXCH, stream>>>(a.TX, a.W, a.H, a.tile_begin, a.n_tiles, a.tile_bins, a.packed, a.src, a.d_status, a.background,
                                                    a.out_img, a.final_Ts, a.final_idx, a.out_mc, a.vec_ok, a.la, a.peers);
}
template <bool TMA, bool AUX, bool LOSS>
static void launch_fwd(const FwdArgs &a, cudaStream_t stream) {
    const int grid = fwd_grid_for<TMA, AUX, LOSS>(a.n_tiles);
    if (fwd_split(a.n_tiles) == 2)
        raster_fwd_kernel<TMA, AUX, LOSS, 2><<<grid, FW_THREADS, FW_SMEM, stream>>>(a.TX, a.W, a.H, a.tile_begin, a.n_tiles, a.tile_bins, a.packed, a.src, a.d_status,
                                                                                   a.background, a.out_img, a.final_Ts, a.final_idx, a.out_mc, a.vec_ok, a.la, a.peers);
    else
        raster_fwd_kernel<TMA, AUX, LOSS, 1><<<grid, FW_THREADS, FW_SMEM, stream>>>(a.TX, a.W, a.H, a.tile_begin, a.n_tiles, a.tile_bins, a.packed, a.src, a.d_status,
                                                                                   a.background, a.out_img, a.final_Ts, a.final_idx, a.out_mc, a.vec_ok, a.la, a.peers);
}
template <typename K>
static int bwd_grid(K kernel, int n_items) { return persistent_grid(kernel, BW_THREADS, BW_SMEM, n_items, "GI2D_BWD_CTAS_PER_SM"); }
// log2 of the warps per tile of the backward rasterizer (they split the tile's chunks of 64 records): 2 warps per tile when
// the image has so few tiles that one warp per tile leaves most warp slots of the 148 SMs empty AND a tile has more than
// one chunk to split (the projections of this path give ~2.5 intersections per Gaussian).  Measured on B200
// (CUPTI kernel time, L2 flushed): 768x512 / 70k (1536 tiles) 23.5 -> 21.1 us; 1020x678 / 100k (2752 tiles) 29.3 -> 32.4 us;
// 768x512 / 10k 9.2 -> 11.8 us; four warps per tile are slower everywhere.
static int bwd_ishift(int N, int n_tiles) {
    static const int forced = getenv("GI2D_BWD_SPLIT") ? atoi(getenv("GI2D_BWD_SPLIT")) : 0;
    if (forced == 1) return 0;
    if (forced == 2) return 1;
    if (forced == 4) return 2;
    const double per_tile = 2.5 * (double)N / (double)n_tiles;
    return (n_tiles <= 2048 && per_tile > 64.0) ? 1 : 0;
}

}  // namespace gi2d

using namespace gi2d;

extern "C" {

// rasterize_sum_forward (reference ext.cpp:22 / bindings.cu:400-469), reference-typed: index-indirect sources.
// out_img (H,W,3), final_Ts (H,W), final_idx (H,W) are fully written.
int gi2d_raster_sum_fwd(int TX, int TY, int W, int H, const int *gids_sorted, const int *tile_bins, const float *xys,
                        const float *conics, const float *colors, const float *opacities, float *out_img,
                        float *final_Ts, int *final_idx, cudaStream_t stream) {
    if (TX <= 0 || TY <= 0 || W <= 0 || H <= 0) return GI2D_ERR_INVALID;
    GatherSrc src{gids_sorted, xys, conics, colors, opacities};
    const FwdArgs a{TX, W, H, 0, TX * TY, const_cast<int *>(tile_bins), nullptr, src, nullptr, nullptr, out_img, final_Ts, final_idx, nullptr,
                    vec_ok(W, out_img), LossArgs{}, PeerOut{}};
    if (final_Ts && final_idx) launch_fwd<false, true, false>(a, stream);
    else launch_fwd<false, false, false>(a, stream);
    return cuda_status();
}

// Forward over the packed records of gi2d_bin_fused (bulk-TMA staging).  d_status[0] == 0 (no intersections
// anywhere) makes the kernel write `background` instead, as rasterize_sum.py:121-129 does.
int gi2d_raster_sum_fwd_packed_rows_mc(int TX, int TY, int W, int H, int ty_begin, int ty_end, int *tile_bins,
                                       const float *packed, const int *d_status, const float *background, float *out_img,
                                       float *out_img_mc, float *final_Ts, int *final_idx, cudaStream_t stream) {
    if (TX <= 0 || TY <= 0 || W <= 0 || H <= 0 || ty_begin < 0 || ty_end > TY || ty_begin >= ty_end) return GI2D_ERR_INVALID;
    const int vok = vec_ok(W, out_img) && vec_ok(W, out_img_mc);
    GatherSrc src{nullptr, nullptr, nullptr, nullptr, nullptr};
    const float4 *p = reinterpret_cast<const float4 *>(packed);
    const int nt = (ty_end - ty_begin) * TX, first = ty_begin * TX;
    const FwdArgs a{TX, W, H, first, nt, tile_bins, p, src, d_status, background, out_img, final_Ts, final_idx, out_img_mc, vok, LossArgs{}, PeerOut{}};
    if (out_img_mc) {
        if (final_Ts || final_idx) return GI2D_ERR_INVALID;           // the multicast variant writes the image only
        launch_fwd_exch(a, stream);
    } else if (final_Ts && final_idx) launch_fwd<true, true, false>(a, stream);
    else launch_fwd<true, false, false>(a, stream);
    return cuda_status();
}

// Row-sharded variant with the all-gather fused into the rasterizer by PEER STORES: every full-tile row is stored to out_img
// and to the same offset of each of the n_peers (<= 7) frame buffers in `peer_imgs` (HOST array of peer-mapped DEVICE
// addresses).  Requires W % 4 == 0, 16-byte aligned buffers and an image of whole tiles (ragged tiles are stored locally only).
int gi2d_raster_sum_fwd_packed_rows_p2p(int TX, int TY, int W, int H, int ty_begin, int ty_end, int *tile_bins, const float *packed,
                                        float *out_img, float *const *peer_imgs, int n_peers, cudaStream_t stream) {
    if (TX <= 0 || TY <= 0 || W <= 0 || H <= 0 || ty_begin < 0 || ty_end > TY || ty_begin >= ty_end || n_peers < 0 || n_peers > 7 ||
        (n_peers && !peer_imgs))
        return GI2D_ERR_INVALID;
    int vok = vec_ok(W, out_img);
    PeerOut po{};
    po.n = n_peers;
    for (int k = 0; k < n_peers; ++k) {
        po.p[k] = peer_imgs[k];
        vok = vok && vec_ok(W, peer_imgs[k]);
    }
    if (n_peers && !vok) return GI2D_ERR_INVALID;
    GatherSrc src{nullptr, nullptr, nullptr, nullptr, nullptr};
    const FwdArgs a{TX, W, H, ty_begin * TX, (ty_end - ty_begin) * TX, tile_bins, reinterpret_cast<const float4 *>(packed), src, nullptr, nullptr,
                    out_img, nullptr, nullptr, nullptr, vok, LossArgs{}, po};
    if (n_peers) launch_fwd_exch(a, stream);
    else launch_fwd<true, false, false>(a, stream);
    return cuda_status();
}

int gi2d_raster_sum_fwd_packed_rows(int TX, int TY, int W, int H, int ty_begin, int ty_end, int *tile_bins,
                                    const float *packed, const int *d_status, const float *background, float *out_img,
                                    float *final_Ts, int *final_idx, cudaStream_t stream) {
    return gi2d_raster_sum_fwd_packed_rows_mc(TX, TY, W, H, ty_begin, ty_end, tile_bins, packed, d_status, background, out_img,
                                              nullptr, final_Ts, final_idx, stream);
}

int gi2d_raster_sum_fwd_packed(int TX, int TY, int W, int H, int *tile_bins, const float *packed,
                               const int *d_status, const float *background, float *out_img, float *final_Ts,
                               int *final_idx, cudaStream_t stream) {
    return gi2d_raster_sum_fwd_packed_rows(TX, TY, W, H, 0, TY, tile_bins, packed, d_status, background, out_img, final_Ts,
                                           final_idx, stream);
}

// rasterize_sum_backward (reference ext.cpp:23 / bindings.cu:706-779).  The four outputs are zeroed here.
// final_idx is accepted for signature parity and not needed: the contributing set is recomputed (see raster.cu header).
int gi2d_raster_sum_bwd(int N, int TX, int TY, int W, int H, const int *gids_sorted, const int *tile_bins,
                        const float *xys, const float *conics, const float *colors, const float *opacities,
                        const int *final_idx, const float *v_out, float *v_xy, float *v_conic, float *v_colors,
                        float *v_opacity, cudaStream_t stream) {
    (void)final_idx;
    if (N <= 0 || TX <= 0 || TY <= 0) return GI2D_ERR_INVALID;
    cudaMemsetAsync(v_xy, 0, sizeof(float) * 2 * (size_t)N, stream);
    cudaMemsetAsync(v_conic, 0, sizeof(float) * 3 * (size_t)N, stream);
    cudaMemsetAsync(v_colors, 0, sizeof(float) * 3 * (size_t)N, stream);
    cudaMemsetAsync(v_opacity, 0, sizeof(float) * (size_t)N, stream);
    GatherSrc src{gids_sorted, xys, conics, colors, opacities};
    GradOut out{nullptr, v_xy, v_conic, v_colors, v_opacity};
    const int ish = bwd_ishift(N, TX * TY);
    raster_bwd_kernel<false, false><<<bwd_grid(raster_bwd_kernel<false, false>, (TX * TY) << ish), BW_THREADS, BW_SMEM, stream>>>(TX, W, H, TX * TY, ish, tile_bins, nullptr, src, v_out, out, vec_ok(W, v_out));
    return cuda_status();
}

// Backward over the packed records.  `grads12` is ONE buffer of (N,12) floats, zero-filled here, one 48-byte row
// per Gaussian: {v_xy.x, v_xy.y, v_conic.x, v_conic.y | v_conic.z, v_colors.r, .g, .b | v_opacity, 0, 0, 0}
// (16-byte aligned rows so that the kernel can use vector reductions).
int gi2d_raster_sum_bwd_packed(int N, int TX, int TY, int W, int H, const int *tile_bins, const float *packed,
                               const float *v_out, float *grads12, cudaStream_t stream) {
    if (N <= 0 || TX <= 0 || TY <= 0) return GI2D_ERR_INVALID;
    if (reinterpret_cast<uintptr_t>(grads12) & 15) return GI2D_ERR_INVALID;
    cudaMemsetAsync(grads12, 0, sizeof(float) * 12 * (size_t)N, stream);
    GatherSrc src{nullptr, nullptr, nullptr, nullptr, nullptr};
    GradOut out{grads12, nullptr, nullptr, nullptr, nullptr};
    const int ish = bwd_ishift(N, TX * TY);
    raster_bwd_kernel<true, true><<<bwd_grid(raster_bwd_kernel<true, true>, (TX * TY) << ish), BW_THREADS, BW_SMEM, stream>>>(TX, W, H, TX * TY, ish, tile_bins, reinterpret_cast<const float4 *>(packed), src, v_out, out, vec_ok(W, v_out));
    return cuda_status();
}

// gi2d_raster_sum_bwd_packed[_scaled] WITHOUT the zero fill: `grads12` was zeroed by the binning chain of the forward pass
// (gi2d_bin_fused_ex / gi2d_repack_records_ex: zero_grads12), so no memset node sits in front of the kernel.  v_scale may be NULL.
int gi2d_raster_sum_bwd_packed_acc(int N, int TX, int TY, int W, int H, const int *tile_bins, const float *packed,
                                   const float *v_out, const float *v_scale, float *grads12, cudaStream_t stream) {
    if (N <= 0 || TX <= 0 || TY <= 0) return GI2D_ERR_INVALID;
    if (reinterpret_cast<uintptr_t>(grads12) & 15) return GI2D_ERR_INVALID;
    GatherSrc src{nullptr, nullptr, nullptr, nullptr, nullptr};
    GradOut out{grads12, nullptr, nullptr, nullptr, nullptr};
    const int ish = bwd_ishift(N, TX * TY);
    raster_bwd_kernel<true, true><<<bwd_grid(raster_bwd_kernel<true, true>, (TX * TY) << ish), BW_THREADS, BW_SMEM, stream>>>(
        TX, W, H, TX * TY, ish, tile_bins, reinterpret_cast<const float4 *>(packed), src, v_out, out, vec_ok(W, v_out), v_scale);
    return cuda_status();
}

// Loss epilogue: forward over the packed records that ends in the L2 loss the reference's training step
// computes from the image (gaussianimage_cholesky.py:84-85 clamp + layout change, utils.py:20-25 F.mse_loss), and writes
// d loss / d out_img into v_out for gi2d_raster_sum_bwd_packed_scaled.  The image itself is optional.
int gi2d_raster_sum_fwd_packed_mse(int TX, int TY, int W, int H, int *tile_bins, const float *packed, const int *d_status,
                                   const float *background, const float *gt, int gt_is_chw, int clamp01, float *v_out,
                                   float *image, float *loss, void *ws, size_t ws_bytes, cudaStream_t stream) {
    if (TX <= 0 || TY <= 0 || W <= 0 || H <= 0 || !gt || !v_out || !loss || !ws) return GI2D_ERR_INVALID;
    const int nt = TX * TY;
    const int grid = fwd_grid_for<true, false, true>(nt);
    if (ws_bytes < sizeof(float) * (size_t)grid + 16) return GI2D_ERR_WORKSPACE;
    LossArgs la{gt, gt_is_chw ? 1 : 0, clamp01 ? 1 : 0, vec_ok(W, gt) && vec_ok(W, image), 1.f / (3.f * (float)H * (float)W), v_out, image,
                reinterpret_cast<float *>(reinterpret_cast<char *>(ws) + 16), reinterpret_cast<unsigned int *>(ws), loss};
    GatherSrc src{nullptr, nullptr, nullptr, nullptr, nullptr};
    const FwdArgs a{TX, W, H, 0, nt, tile_bins, reinterpret_cast<const float4 *>(packed), src, d_status, background, nullptr, nullptr, nullptr,
                    nullptr, vec_ok(W, v_out), la, PeerOut{}};
    launch_fwd<true, false, true>(a, stream);
    return cuda_status();
}

// Workspace of gi2d_raster_sum_fwd_packed_mse: a ticket word (must be ZERO before the first call; the kernel leaves it
// zero) and one partial sum per resident CTA.
size_t gi2d_raster_mse_workspace_bytes(void) {
    const int g1 = fwd_grid(raster_fwd_kernel<true, false, true, 1>, 1 << 30), g2 = fwd_grid(raster_fwd_kernel<true, false, true, 2>, 1 << 30);
    return 16 + sizeof(float) * (size_t)(g1 > g2 ? g1 : g2);
}

// gi2d_raster_sum_bwd_packed with v_out multiplied by the device scalar *v_scale (the upstream gradient of the loss).
int gi2d_raster_sum_bwd_packed_scaled(int N, int TX, int TY, int W, int H, const int *tile_bins, const float *packed,
                                      const float *v_out, const float *v_scale, float *grads12, cudaStream_t stream) {
    if (N <= 0 || TX <= 0 || TY <= 0) return GI2D_ERR_INVALID;
    if (reinterpret_cast<uintptr_t>(grads12) & 15) return GI2D_ERR_INVALID;
    cudaMemsetAsync(grads12, 0, sizeof(float) * 12 * (size_t)N, stream);
    GatherSrc src{nullptr, nullptr, nullptr, nullptr, nullptr};
    GradOut out{grads12, nullptr, nullptr, nullptr, nullptr};
    const int ish = bwd_ishift(N, TX * TY);
    raster_bwd_kernel<true, true><<<bwd_grid(raster_bwd_kernel<true, true>, (TX * TY) << ish), BW_THREADS, BW_SMEM, stream>>>(TX, W, H, TX * TY, ish, tile_bins, reinterpret_cast<const float4 *>(packed), src, v_out, out, vec_ok(W, v_out), v_scale);
    return cuda_status();
}

}  // extern "C"

#ifdef GI2D_TRACE
extern "C" int gi2d_debug_set_trace(void *buf) {
    return (int)cudaMemcpyToSymbol(gi2d::g_trace, &buf, sizeof(buf));
}
#endif
