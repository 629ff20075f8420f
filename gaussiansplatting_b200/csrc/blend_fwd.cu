// blend_fwd.cu — K5 tile ranges (buildTileRanges, sort.metal:553-589) and K6 the per-tile front-to-back
// alpha blend (tiledForward, tiled_shaders.metal:307-385; graded fp32 mode, SURVEY note H).
//
// K6: one CTA per 16x16 tile, one thread per pixel, each warp an 8x4 pixel patch.  The tile's sorted list is
// staged through shared memory 256 entries at a time (one 48-B record gather per thread, then broadcast
// LDS.128 in the inner loop).  FP32-issue bound; all arithmetic is the pinned sequence of gs_common.cuh, so the
// image, lastContribIdx and T_final are bit-identical to the oracle.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

// lb(t) = first index whose key's tile is >= t  (the reference's binary search, sort.metal:568-577)
__device__ __forceinline__ uint32_t tile_lower_bound(const uint64_t* __restrict__ keys, uint32_t n, uint32_t t) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if ((uint32_t)(keys[mid] >> 32) < t) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// start = lb(t); count = run length of tile t = lb(t+1) - lb(t)  (empty tiles keep start = insertion point)
__global__ void tile_ranges_kernel(const uint64_t* __restrict__ keys, uint32_t n, uint32_t num_tiles, uint2* __restrict__ ranges) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_tiles) return;
    const uint32_t a = tile_lower_bound(keys, n, t);
    const uint32_t b = tile_lower_bound(keys, n, t + 1);
    ranges[t] = make_uint2(a, b - a);
}

// K5 for sorted u32 tile ids: thread j owns the boundary between entries j-1 and j; the first entry of a run writes
// `start` for its tile and for every empty tile skipped since the previous run (buildTileRanges gives an empty tile
// the insertion point as start, sort.metal:568-577).  counts = start[t+1] - start[t] in a second tiny pass.
__global__ void __launch_bounds__(256) tile_starts_kernel(const uint32_t* __restrict__ tiles, uint32_t max_pairs,
                                                          const uint32_t* __restrict__ d_n, uint32_t num_tiles,
                                                          uint2* __restrict__ ranges) {
    const uint32_t n = min(*d_n, max_pairs);
    const uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (n == 0) { if (t0 < num_tiles) ranges[t0] = make_uint2(0u, 0u); return; }
    const uint32_t j0 = t0 * 4u;                              // four consecutive entries per thread
    if (j0 >= n) return;
    int64_t tp = j0 ? (int64_t)min(tiles[j0 - 1], num_tiles - 1u) : -1;
#pragma unroll
    for (uint32_t q = 0; q < 4u; q++) {
        const uint32_t j = j0 + q;
        if (j >= n) break;
        const uint32_t t = min(tiles[j], num_tiles - 1u);
        for (int64_t tt = tp + 1; tt <= (int64_t)t; tt++) ranges[tt].x = j;
        if (j == n - 1u) for (uint32_t tt = t + 1u; tt < num_tiles; tt++) ranges[tt].x = n;
        tp = t;
    }
}
__global__ void tile_counts_kernel(uint32_t max_pairs, const uint32_t* __restrict__ d_n, uint32_t num_tiles, uint2* __restrict__ ranges) {
    const uint32_t n = min(*d_n, max_pairs);
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_tiles || n == 0) return;
    const uint32_t a = ranges[t].x, b = (t + 1u < num_tiles) ? ranges[t + 1u].x : n;
    ranges[t].y = b - a;
}
__global__ void rebuild_keys64_kernel(const uint32_t* __restrict__ tiles, const uint32_t* __restrict__ vals,
                                      const uint32_t* __restrict__ dkey, uint32_t n, uint64_t* __restrict__ keys) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) keys[j] = ((uint64_t)tiles[j] << 32) | (uint64_t)dkey[vals[j] & GS_ID_MASK];
}

#ifndef GS_FWD_PPT
#define GS_FWD_PPT 2
#endif
constexpr int FB = 256;   // raw list entries staged per round

// K6.  PPT pixels per thread, NT = 256/PPT threads per tile.  Each round a CTA reads FB consecutive entries of the
// tile's sorted list, drops those flagged GS_CULL_BIT (pairs.cu: the splat cannot reach any pixel of this tile),
// compacts the survivors in list order into shared memory (one 48-B record gather each) and every thread walks them.
template <int PPT, bool COUNT>
__global__ void __launch_bounds__(256 / PPT) blend_forward_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ svals,
                                                            const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                            float4* __restrict__ image, uint32_t* __restrict__ last_idx,
                                                            float* __restrict__ tfinal, uint32_t* __restrict__ tile_flags,
                                                            unsigned long long* __restrict__ counters) {
    constexpr int NT = 256 / PPT, NW = NT / 32, K = FB / NT;
    __shared__ float4 s_r0[FB], s_r1[FB];
    __shared__ float s_b[FB];
    __shared__ uint32_t s_idx[FB];
    __shared__ uint32_t s_w[NW];
    __shared__ uint32_t s_mask[NW][FB / 32];          // per warp region: which staged entries can reach it at all
    __shared__ uint32_t s_degen;
    const uint32_t tile = blockIdx.x;
    const uint32_t tX = tile % ntx, tY = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint2 range = ranges[tile];                                // :320-323
    if (threadIdx.x == 0) s_degen = 0;

    uint32_t px[PPT], py[PPT];
    float pxf[PPT], pyf[PPT], Cr[PPT], Cg[PPT], Cb[PPT], T[PPT], Tp[PPT];
    uint32_t last[PPT];
    bool done[PPT];
    int ndone = 0;
    uint32_t nC = 0;   // COUNT only: contributing evaluations (roofline accounting)
#pragma unroll
    for (int j = 0; j < PPT; j++) {
        uint32_t x, y;
        gs::tile_pixel<PPT>(warp, lane, j, x, y);
        px[j] = tX * 16u + x; py[j] = tY * 16u + y;
        pxf[j] = (float)px[j] + 0.5f; pyf[j] = (float)py[j] + 0.5f;   // :328
        Cr[j] = 0.f; Cg[j] = 0.f; Cb[j] = 0.f; T[j] = 1.0f; Tp[j] = 1.0f;
        last[j] = 0xFFFFFFFFu;
        done[j] = !(px[j] < W && py[j] < H);
        ndone += done[j] ? 1 : 0;
    }
    for (uint32_t base = 0; base < range.y; base += FB) {
        if (__syncthreads_and(ndone == PPT)) break;
        // ---- stage: drop culled pairs, keep list order
        const uint32_t cntraw = min((uint32_t)FB, range.y - base);
        uint32_t v[K], nl = 0;
#pragma unroll
        for (int q = 0; q < K; q++) {
            const uint32_t j = threadIdx.x * K + q;
            v[q] = (j < cntraw) ? svals[range.x + base + j] : GS_CULL_BIT;
            nl += (v[q] >> 31) ^ 1u;
        }
        uint32_t total;
        uint32_t slot = gs::cta_excl_scan<NW>(nl, s_w, &total);
#pragma unroll
        for (int q = 0; q < K; q++) {
            if (v[q] >> 31) continue;
            const uint32_t g = v[q];
            const float4 r0 = rec[3 * (size_t)g], r2 = rec[3 * (size_t)g + 2];
            float4 r1 = rec[3 * (size_t)g + 1];
            if (__float_as_uint(r2.w) & 1u) { r1.y = 0.f; s_degen = 1u; }   // :350-351 conic-magnitude skip (forward only)
            s_r0[slot] = r0; s_r1[slot] = r1; s_b[slot] = r2.x; s_idx[slot] = base + threadIdx.x * K + q;
            slot++;
        }
        __syncthreads();
        // ---- per-region footprint masks: lane tests one staged entry against every warp's pixel rectangle
        for (uint32_t s0 = warp * 32u; s0 < total; s0 += NT) {
            const uint32_t sl = s0 + lane;
            const bool valid = sl < total;
            const gs::CullRec cr = gs::cull_rec(s_r0[valid ? sl : 0u], s_r1[valid ? sl : 0u]);
#pragma unroll
            for (int w = 0; w < NW; w++) {
                uint32_t rx, ry; float ex, ey;
                gs::warp_region<PPT>((uint32_t)w, rx, ry, ex, ey);
                const bool keep = valid && !gs::box_cannot_contribute(cr, (float)(tX * 16u + rx) + 0.5f, (float)(tY * 16u + ry) + 0.5f, ex, ey);
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
                if (lane == 0) s_mask[w][s0 >> 5] = m;
            }
        }
        __syncthreads();
        if (ndone < PPT) {
            const uint32_t nwords = (total + 31u) >> 5;
            for (uint32_t wd = 0; wd < nwords; wd++) {               // :334, entries that can reach this warp's pixels only
                uint32_t m = s_mask[warp][wd];
                while (m) {
                    const uint32_t k = wd * 32u + (uint32_t)__ffs(m) - 1u;
                    m &= m - 1u;
                    const float4 a = s_r0[k];
                    const float4 b = s_r1[k];
#pragma unroll
                    for (int j = 0; j < PPT; j++) {
                        if (done[j]) continue;
                        const float dx = gs::fsub(pxf[j], a.x), dy = gs::fsub(pyf[j], a.y);   // :347
                        const float power = gs::blend_power(a.z, a.w, b.x, dx, dy);
                        if (!(power <= 0.0f) || power < -4.5f) continue;  // :359
                        const float G = gs::exp_blend(power);             // :362
                        const float alpha = fminf(gs::fmul(b.y, G), GS_ALPHA_MAX);   // :363
                        if (alpha < GS_ALPHA_MIN) continue;               // :366
                        const float w = gs::fmul(alpha, T[j]);            // :369
                        Cr[j] = gs::ffma(b.z, w, Cr[j]); Cg[j] = gs::ffma(b.w, w, Cg[j]); Cb[j] = gs::ffma(s_b[k], w, Cb[j]);
                        Tp[j] = T[j];
                        if (COUNT) nC++;
                        T[j] = gs::fmul(T[j], gs::fsub(1.0f, alpha));     // :370
                        last[j] = range.x + s_idx[k];                     // :372
                        if (!(T[j] > GS_T_EPS)) { done[j] = true; ndone++; }   // :334 loop condition
                    }
                    if (ndone == PPT) { wd = nwords; break; }
                }
            }
        }
    }
    if (COUNT) {
        // E = list entries the reference's loop visits: up to and including the entry that ended the walk, else all
        unsigned long long e = 0, c = nC;
#pragma unroll
        for (int j = 0; j < PPT; j++)
            if (px[j] < W && py[j] < H) e += (!(T[j] > GS_T_EPS)) ? (last[j] - range.x + 1u) : range.y;
        for (int d = 16; d >= 1; d >>= 1) { e += __shfl_xor_sync(0xFFFFFFFFu, e, d); c += __shfl_xor_sync(0xFFFFFFFFu, c, d); }
        if (lane == 0) { atomicAdd(&counters[0], e); atomicAdd(&counters[1], c); }
        return;
    }
#pragma unroll
    for (int j = 0; j < PPT; j++) {
        if (px[j] < W && py[j] < H) {
            const size_t pix = (size_t)py[j] * W + px[j];
            image[pix] = make_float4(gs::fadd(Cr[j], T[j]), gs::fadd(Cg[j], T[j]), gs::fadd(Cb[j], T[j]), 1.0f);   // :377-384 white background
            last_idx[pix] = last[j];                                  // :382
            // T_final as tiledBackward's first pass (:430-460) would recompute it: the entry that drives T below
            // 1e-4 is NOT applied there (break before apply), every earlier one is.
            tfinal[pix] = (T[j] < GS_T_EPS) ? Tp[j] : T[j];
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) tile_flags[tile] = s_degen;
}

}  // namespace

static int g_fwd_ppt = GS_FWD_PPT;
void gs_tune_fwd_ppt(int v) { if (v == 1 || v == 2 || v == 4) g_fwd_ppt = v; }

void gs_launch_tile_ranges(const uint64_t* sorted_keys, uint32_t num_pairs, uint32_t num_tiles, uint32_t* /*lb*/,
                           uint2* ranges, cudaStream_t st) {
    if (num_tiles == 0) return;
    tile_ranges_kernel<<<(num_tiles + 127) / 128, 128, 0, st>>>(sorted_keys, num_pairs, num_tiles, ranges);
}

void gs_launch_tile_ranges32(const uint32_t* sorted_tiles, uint32_t max_pairs, const uint32_t* d_n, uint32_t num_tiles,
                             uint2* ranges, cudaStream_t st) {
    if (num_tiles == 0) return;
    const uint32_t work = (max_pairs + 3u) / 4u > num_tiles ? (max_pairs + 3u) / 4u : num_tiles;
    tile_starts_kernel<<<(work + 255) / 256, 256, 0, st>>>(sorted_tiles, max_pairs, d_n, num_tiles, ranges);
    tile_counts_kernel<<<(num_tiles + 255) / 256, 256, 0, st>>>(max_pairs, d_n, num_tiles, ranges);
}
void gs_launch_rebuild_keys64(const uint32_t* sorted_tiles, const uint32_t* sorted_vals, const uint32_t* dkey, uint32_t n,
                              uint64_t* keys, cudaStream_t st) {
    if (n) rebuild_keys64_kernel<<<(n + 255) / 256, 256, 0, st>>>(sorted_tiles, sorted_vals, dkey, n, keys);
}

void gs_launch_blend_forward(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                             uint32_t ntx, uint32_t nty, float4* image, uint32_t* last_idx, float* tfinal, cudaStream_t st) {
    if (ntx * nty == 0) return;
#define GS_FWD_LAUNCH(PPT)                                                                                            \
    blend_forward_kernel<PPT, false><<<ntx * nty, 256 / PPT, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, image, last_idx, \
                                                                       tfinal, O.tile_flags, nullptr)
    if (g_fwd_ppt == 4) GS_FWD_LAUNCH(4);
    else if (g_fwd_ppt == 2) GS_FWD_LAUNCH(2);
    else GS_FWD_LAUNCH(1);
}

// counts E (entries visited) and C (contributing evaluations) of the forward walk; counters[2] zeroed by the caller
void gs_launch_blend_count(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                           uint32_t ntx, uint32_t nty, unsigned long long* counters, cudaStream_t st) {
    if (ntx * nty == 0) return;
    blend_forward_kernel<1, true><<<ntx * nty, 256, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, nullptr, nullptr, nullptr,
                                                             O.tile_flags, counters);
}
