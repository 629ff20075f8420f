// blend_fwd.cu — K5 tile ranges (buildTileRanges, sort.metal:553-589) and K6 the per-tile front-to-back
// alpha blend (tiledForward, tiled_shaders.metal:307-385; graded fp32 mode, SURVEY note H).
//
// K6: one CTA of 128 threads per 16x16 tile, two pixels per thread held as one packed fp32 pair (FFMA2), each warp an 8x8
// pixel region.  Two kernels share the visit arithmetic:
//   blend_forward_kernel        the four warps stage the tile's sorted list through shared memory 128 entries at a time (one
//                               48-B record gather per thread, footprint bits per warp region, broadcast LDS.128 in the walk);
//                               the default of a render-only frame
//   blend_forward_warp_kernel   every warp walks the list on its own, 32 entries at a time, and stops with its own 64 pixels;
//                               the default of a training frame, where it also records the visit words the backward walks on
// Issue-bound (FP32 pipe close behind); all arithmetic is the pinned sequence of gs_common.cuh, so the image, lastContribIdx
// and T_final are bit-identical to the oracle from either kernel.
#include "gs_common.cuh"
#include "gs_kernels.h"

#ifndef GS_FWD_MINB
#define GS_FWD_MINB 10     // minimum resident CTAs per SM asked of ptxas (48 registers; measured 243 -> 235 us at C2)
#endif

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
    const uint32_t n = min(*d_n, max_pairs);                  // device-side pair count: the grid is fixed, threads stride
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (n == 0) { for (; t0 < num_tiles; t0 += stride) ranges[t0] = make_uint2(0u, 0u); return; }
    for (; (uint64_t)t0 * 4u < n; t0 += stride) {
        const uint32_t j0 = t0 * 4u;                          // four consecutive entries per thread
        int64_t tp = j0 ? (int64_t)min(tiles[j0 - 1], num_tiles - 1u) : -1;
        uint32_t tt4[4] = {0u, 0u, 0u, 0u};
        if (j0 + 4u <= n) { const uint4 q = *reinterpret_cast<const uint4*>(tiles + j0); tt4[0] = q.x; tt4[1] = q.y; tt4[2] = q.z; tt4[3] = q.w; }
        else for (uint32_t q = 0; q < 4u && j0 + q < n; q++) tt4[q] = tiles[j0 + q];
#pragma unroll
        for (uint32_t q = 0; q < 4u; q++) {
            const uint32_t j = j0 + q;
            if (j >= n) break;
            const uint32_t t = min(tt4[q], num_tiles - 1u);
            for (int64_t tt = tp + 1; tt <= (int64_t)t; tt++) ranges[tt].x = j;
            if (j == n - 1u) for (uint32_t tt = t + 1u; tt < num_tiles; tt++) ranges[tt].x = n;
            tp = t;
        }
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
#ifndef GS_FWD_FB
#define GS_FWD_FB 128
#endif
constexpr int FB = GS_FWD_FB;   // raw list entries staged per round

// K6.  PPT (2 or 4) pixels per thread, NT = 256/PPT threads per tile.  Each round a CTA reads FB consecutive entries of
// the tile's sorted list, drops those flagged GS_CULL_BIT (pairs.cu: the splat cannot reach any pixel of this tile),
// compacts the survivors in list order into shared memory (one 48-B record gather each), marks per warp the survivors
// that can reach the warp's own pixel rectangle, and every warp walks its marked entries.
// The walk is branch-free and keeps the thread's pixels two to a packed fp32 pair (FFMA2): a pixel that does not take an
// entry runs the same instructions with alpha forced to 0, which leaves C and T bit-unchanged (x*0 + C, T*(1-0)); a
// finished pixel has NaN coordinates, so every test on it fails.  Per half the operation sequence is the pinned one.
// launch order of the blend CTAs (gs::tile_order_cta); the binned pipeline computes it inside bin_fill instead
__global__ void __launch_bounds__(1024) tile_order_kernel(const uint2* __restrict__ ranges, uint32_t nt, uint32_t* __restrict__ order) {
    gs::tile_order_cta<1024>(ranges, nt, order);
}

struct FwdEnt { float4 r0, r1; float b; uint32_t idx; uint32_t pad0, pad1; };   // 48 B: one base address, three LDS
// vis (optional, PPT 2 only; four planes of `vstride` words, one per warp: word (w, entry) at vis[w * vstride + entry], so that
// a warp's words of consecutive entries are contiguous): for every STAGED list entry four 64-bit words, one per warp, "which of the warp's 64 pixel slots
// took the entry" (two ballots: slot j = 0 / 1 of every lane; zero = the warp never took it), at vis[entry * 4 + warp].  A warp
// keeps its word in shared memory; the CTA writes a round's words out once, 32 contiguous bytes per entry.  The backward
// walk takes its per-slot decisions from these words instead of re-deriving them (blend_bwd.cu).  Entries behind the round in
// which the whole tile finished are never staged and get no words: no pixel's lastContribIdx lies beyond that round.
template <int PPT, bool COUNT, bool VIS>
__global__ void __launch_bounds__(256 / PPT, GS_FWD_MINB) blend_forward_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ svals,
                                                            const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                            float4* __restrict__ image, uint32_t* __restrict__ last_idx,
                                                            float* __restrict__ tfinal, uint32_t* __restrict__ tile_flags,
                                                            const uint32_t* __restrict__ order,
                                                            unsigned long long* __restrict__ counters, uint2* __restrict__ vis, size_t vstride) {
    constexpr int NT = 256 / PPT, NW = NT / 32, K = FB / NT, NP = PPT / 2;
    __shared__ FwdEnt s_e[FB];
    __shared__ uint32_t s_w[NW];
    __shared__ unsigned char s_reach[FB];             // per staged entry: which warp regions it can reach at all (bit w)
    static_assert(!VIS || (PPT == 2 && !COUNT), "visit words are recorded by the two-pixel blend only");
    __shared__ uint2 s_vis[VIS ? FB : 1][NW];   // this round's visit words (zero = the warp did not take the entry)
    __shared__ uint32_t s_degen;
    const uint32_t tile = order ? order[blockIdx.x] : blockIdx.x;
    const uint32_t tX = tile % ntx, tY = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint2 range = ranges[tile];                                // :320-323
    if (threadIdx.x == 0) s_degen = 0;
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(s_e);
    // this warp's column of s_vis as a warp-uniform address (the broadcast tells the compiler so: no per-thread register, no
    // special-register read inside the walk)
    const uint32_t vbase = (uint32_t)__cvta_generic_to_shared(s_vis) + __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0) * (uint32_t)sizeof(uint2);

    uint32_t px[PPT], py[PPT], last[PPT];
    float Tp[PPT];
    gs::f32x2 X2[NP], Y2[NP], T2[NP], Cr2[NP], Cg2[NP], Cb2[NP];
    uint32_t nC = 0;   // COUNT only: contributing evaluations (roofline accounting)
    bool live = false;
#pragma unroll
    for (int p = 0; p < NP; p++) {
        float xf[2], yf[2], t0[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int j = 2 * p + h;
            uint32_t x, y;
            gs::tile_pixel<PPT>(warp, lane, j, x, y);
            px[j] = tX * 16u + x; py[j] = tY * 16u + y;
            const bool in = px[j] < W && py[j] < H;
            xf[h] = (float)px[j] + 0.5f; yf[h] = (float)py[j] + 0.5f;   // :328
            t0[h] = in ? 1.0f : 0.0f;                                   // a pixel outside the image never walks (T = 0 <= eps)
            Tp[j] = 1.0f; last[j] = 0xFFFFFFFFu;
            live |= in;
        }
        X2[p] = gs::pk2(xf[0], xf[1]); Y2[p] = gs::pk2(yf[0], yf[1]);
        T2[p] = gs::pk2(t0[0], t0[1]); Cr2[p] = gs::bc2(0.f); Cg2[p] = gs::bc2(0.f); Cb2[p] = gs::bc2(0.f);
    }
    bool wlive = __any_sync(0xFFFFFFFFu, live);      // warp-uniform: some pixel of this warp still walks
    constexpr bool rec_vis = VIS;   // a kernel of its own: the render-only frame keeps the walk without the two votes and the store
    uint32_t prev_total = 0;                          // staged entries of the previous round, whose visit words are still in s_vis
    for (uint32_t base = 0; base < range.y; base += FB) {
        const bool all_done = __syncthreads_and(!wlive);
        if (rec_vis) {                                // the previous round's words: 32 contiguous bytes per entry, one thread each
            for (uint32_t k = threadIdx.x; k < prev_total; k += NT) {
#pragma unroll
                for (int w = 0; w < NW; w++) vis[(size_t)w * vstride + s_e[k].idx] = s_vis[k][w];
            }
            prev_total = 0;
        }
        if (all_done) break;
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
            s_e[slot].r0 = r0; s_e[slot].r1 = r1; s_e[slot].b = r2.x; s_e[slot].idx = range.x + base + threadIdx.x * K + q;
            // footprint against every warp's pixel rectangle, by the thread that staged the entry: bit w = it can reach warp w
            const gs::CullRec cr = gs::cull_rec(r0, r1);
            uint32_t reach = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) {
                uint32_t rx, ry; float ex, ey;
                gs::warp_region<PPT>((uint32_t)w, rx, ry, ex, ey);
                if (!gs::box_cannot_contribute(cr, (float)(tX * 16u + rx) + 0.5f, (float)(tY * 16u + ry) + 0.5f, ex, ey)) reach |= 1u << w;
            }
            s_reach[slot] = (unsigned char)reach;
            if (rec_vis) {
#pragma unroll
                for (int w = 0; w < NW; w++) s_vis[slot][w] = make_uint2(0u, 0u);
            }
            slot++;
        }
        __syncthreads();
        prev_total = total;
        if (wlive) {
            const uint32_t nwords = (total + 31u) >> 5;
            for (uint32_t wd = 0; wd < nwords && wlive; wd++) {      // :334, entries that can reach this warp's pixels only
                const uint32_t sl = wd * 32u + lane;
                uint32_t m = __ballot_sync(0xFFFFFFFFu, sl < total && ((s_reach[sl] >> warp) & 1u));
                uint32_t wbase = ebase + wd * 32u * (uint32_t)sizeof(FwdEnt);
                // opaque to the compiler: otherwise it re-derives the shared window's base from %cluster_ctaid inside the visit loop
                // (five uniform-datapath instructions per visit, which still take issue slots)
                asm volatile("" : "+r"(wbase));
                while (m) {
                    const uint32_t kb = (uint32_t)__ffs(m) - 1u;
                    m &= m - 1u;
                    const uint32_t addr = wbase + kb * (uint32_t)sizeof(FwdEnt);
                    const float4 a = gs::lds128(addr);
                    const float4 b = gs::lds128(addr + 16u);
                    const float2 c = gs::lds64(addr + 32u);
                    const uint32_t idx = __float_as_uint(c.y);
                    const gs::f32x2 sx = gs::bc2(a.x), sy = gs::bc2(a.y), Ah = gs::bc2(a.z), Bn = gs::bc2(a.w), Ch = gs::bc2(b.x),
                                    op = gs::bc2(b.y), cr = gs::bc2(b.z), cg = gs::bc2(b.w), cb = gs::bc2(c.x);
                    bool alive = false;
                    bool took0 = false, took1 = false;
#pragma unroll
                    for (int p = 0; p < NP; p++) {
                        const gs::f32x2 dx = gs::fsub2(X2[p], sx), dy = gs::fsub2(Y2[p], sy);   // :347
                        const gs::f32x2 pw2 = gs::blend_power2(Ah, Bn, Ch, dx, dy);
                        const gs::f32x2 G2 = gs::exp_blend2(pw2);                             // :362
                        const gs::f32x2 a2 = gs::fmul2(op, G2);                               // :363
                        float pw[2], al[2], Tn[2], To[2];
                        gs::up2(pw2, pw[0], pw[1]); gs::up2(a2, al[0], al[1]); gs::up2(T2[p], To[0], To[1]);
                        bool c2[2];
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            al[h] = fminf(al[h], GS_ALPHA_MAX);
                            // :334 (the pixel still walks: T > 1e-4), :359, :366
                            c2[h] = (To[h] > GS_T_EPS) && (pw[h] <= 0.0f) && !(pw[h] < -4.5f) && !(al[h] < GS_ALPHA_MIN);
                            al[h] = c2[h] ? al[h] : 0.0f;
                        }
                        if (p == 0) { took0 = c2[0]; took1 = c2[1]; }
                        const gs::f32x2 alpha = gs::pk2(al[0], al[1]);
                        const gs::f32x2 w2 = gs::fmul2(alpha, T2[p]);                          // :369
                        Cr2[p] = gs::ffma2(cr, w2, Cr2[p]); Cg2[p] = gs::ffma2(cg, w2, Cg2[p]); Cb2[p] = gs::ffma2(cb, w2, Cb2[p]);
                        T2[p] = gs::fmul2(T2[p], gs::fsub2(gs::bc2(1.0f), alpha));            // :370
                        gs::up2(T2[p], Tn[0], Tn[1]);
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int j = 2 * p + h;
                            Tp[j] = c2[h] ? To[h] : Tp[j];
                            last[j] = c2[h] ? idx : last[j];                                  // :372
                            if (COUNT) nC += c2[h] ? 1u : 0u;
                            alive |= Tn[h] > GS_T_EPS;                                        // :334 loop condition
                        }
                    }
                    if (rec_vis) {
                        const uint32_t m0 = __ballot_sync(0xFFFFFFFFu, took0), m1 = __ballot_sync(0xFFFFFFFFu, took1);
                        // every lane stores the same word to the same address: one shared-memory transaction, no lane predicate
                        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(vbase + (wd * 32u + kb) * (uint32_t)(NW * sizeof(uint2))), "r"(m0), "r"(m1) : "memory");
                    }
                    if (!__any_sync(0xFFFFFFFFu, alive)) { wlive = false; break; }
                }
            }
        }
    }
    if (rec_vis && __syncthreads_or(prev_total != 0u)) {   // the last round's words (the loop ended by exhausting the list)
        for (uint32_t k = threadIdx.x; k < prev_total; k += NT) {
#pragma unroll
            for (int w = 0; w < NW; w++) vis[(size_t)w * vstride + s_e[k].idx] = s_vis[k][w];
        }
    }
    float Tf[PPT], Cr[PPT], Cg[PPT], Cb[PPT];
#pragma unroll
    for (int p = 0; p < NP; p++) {
        gs::up2(T2[p], Tf[2 * p], Tf[2 * p + 1]); gs::up2(Cr2[p], Cr[2 * p], Cr[2 * p + 1]);
        gs::up2(Cg2[p], Cg[2 * p], Cg[2 * p + 1]); gs::up2(Cb2[p], Cb[2 * p], Cb[2 * p + 1]);
    }
    if (COUNT) {
        // E = list entries the reference's loop visits: up to and including the entry that ended the walk, else all
        unsigned long long e = 0, c = nC;
#pragma unroll
        for (int j = 0; j < PPT; j++)
            if (px[j] < W && py[j] < H) e += (!(Tf[j] > GS_T_EPS)) ? (last[j] - range.x + 1u) : range.y;
        for (int d = 16; d >= 1; d >>= 1) { e += __shfl_xor_sync(0xFFFFFFFFu, e, d); c += __shfl_xor_sync(0xFFFFFFFFu, c, d); }
        if (lane == 0) { atomicAdd(&counters[0], e); atomicAdd(&counters[1], c); }
        return;
    }
#pragma unroll
    for (int j = 0; j < PPT; j++) {
        if (px[j] < W && py[j] < H) {
            const size_t pix = (size_t)py[j] * W + px[j];
            image[pix] = make_float4(gs::fadd(Cr[j], Tf[j]), gs::fadd(Cg[j], Tf[j]), gs::fadd(Cb[j], Tf[j]), 1.0f);   // :377-384 white background
            last_idx[pix] = last[j];                                  // :382
            // T_final as tiledBackward's first pass (:430-460) would recompute it: the entry that drives T below
            // 1e-4 is NOT applied there (break before apply), every earlier one is.
            tfinal[pix] = (Tf[j] < GS_T_EPS) ? Tp[j] : Tf[j];
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) tile_flags[tile] = s_degen;
}

// ---------------------------------------------------------------------------------------------------
// K6, warp-autonomous walk (gs_tune "fwd_warp" = 1).  In the kernel above the four warps of a tile stage rounds of 128
// entries together and meet at a barrier per round; a warp whose pixels are saturated, or which has few visits in a
// round, waits for the others (ncu: the barrier is the kernel's largest stall).  Here a warp walks the tile's list on its own,
// 32 entries at a time: lane i reads entry i's pair value (the next chunk's is in flight), gathers its record, tests its
// footprint against the warp's 8x8 pixels and, if it can reach them, puts it into the warp's 1.5 KB of shared memory; the
// warp then visits those entries in list order and stops as soon as its own 64 pixels are done.  The records are read by all
// four warps of the tile (L1 serves the repeats).  VIS: the lane that holds entry kb keeps the visit's two ballots and stores
// the warp's word of its entry after the chunk (zero if the warp did not take it); words exist up to the end of the chunk in
// which the warp stopped, which covers every lastContribIdx of its pixels -- the extent the warp-autonomous backward reads.
// Arithmetic and order of a pixel's walk are those of the kernel above: images and lastContribIdx are bit-identical.
#ifndef GS_FWD_WARP_MINB
#define GS_FWD_WARP_MINB 10
#endif
template <bool VIS>
__global__ void __launch_bounds__(128, GS_FWD_WARP_MINB) blend_forward_warp_kernel(
    const float4* __restrict__ rec, const uint32_t* __restrict__ svals, const uint2* __restrict__ ranges, uint32_t W, uint32_t H,
    uint32_t ntx, float4* __restrict__ image, uint32_t* __restrict__ last_idx, float* __restrict__ tfinal,
    uint32_t* __restrict__ tile_flags, const uint32_t* __restrict__ order, uint2* __restrict__ vis, size_t vstride) {
    __shared__ __align__(16) FwdEnt s_e[4][32];
    const uint32_t tile = order ? order[blockIdx.x] : blockIdx.x;
    const uint32_t tX = tile % ntx, tY = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint2 range = ranges[tile];                                // :320-323
    if (threadIdx.x == 0) tile_flags[tile] = 0u;
    __syncthreads();          // the only barrier: the clear of the tile's flag is ordered before any warp's set (end of kernel)
    uint32_t px[2], py[2], last[2];
    float Tp[2];
    gs::f32x2 X2, Y2, T2, Cr2 = gs::bc2(0.f), Cg2 = gs::bc2(0.f), Cb2 = gs::bc2(0.f);
    bool live = false;
    {
        float xf[2], yf[2], t0[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            uint32_t x, y;
            gs::tile_pixel<2>(warp, lane, h, x, y);
            px[h] = tX * 16u + x; py[h] = tY * 16u + y;
            const bool in = px[h] < W && py[h] < H;
            xf[h] = (float)px[h] + 0.5f; yf[h] = (float)py[h] + 0.5f;   // :328
            t0[h] = in ? 1.0f : 0.0f;                                   // a pixel outside the image never walks (T = 0 <= eps)
            Tp[h] = 1.0f; last[h] = 0xFFFFFFFFu;
            live |= in;
        }
        X2 = gs::pk2(xf[0], xf[1]); Y2 = gs::pk2(yf[0], yf[1]); T2 = gs::pk2(t0[0], t0[1]);
    }
    bool wlive = __any_sync(0xFFFFFFFFu, live);      // warp-uniform: some pixel of this warp still walks
    const uint32_t wu = __shfl_sync(0xFFFFFFFFu, warp, 0);           // the warp index as a value the compiler knows to be uniform
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(s_e) + wu * (32u * (uint32_t)sizeof(FwdEnt));
    FwdEnt* const me = &s_e[wu][lane];
    uint32_t rx, ry;
    float ex, ey;
    gs::warp_region<2>(wu, rx, ry, ex, ey);
    const float bx0 = (float)(tX * 16u + rx) + 0.5f, by0 = (float)(tY * 16u + ry) + 0.5f;
    const uint32_t* const sv = svals + range.x;
    bool degen = false;
    uint32_t v_n = lane < range.y ? sv[lane] : GS_CULL_BIT;
    for (uint32_t base = 0; base < range.y && wlive; base += 32u) {
        const uint32_t v = v_n;
        v_n = base + 32u + lane < range.y ? sv[base + 32u + lane] : GS_CULL_BIT;   // in flight while this chunk is walked
        bool reach = false;
        if (!(v >> 31)) {                                            // a culled pair can never contribute (pairs.cu)
            const float4 r0 = rec[3 * (size_t)v], r2 = rec[3 * (size_t)v + 2];
            float4 r1 = rec[3 * (size_t)v + 1];
            if (__float_as_uint(r2.w) & 1u) { r1.y = 0.f; degen = true; }   // :350-351 conic-magnitude skip (forward only)
            reach = !gs::box_cannot_contribute(gs::cull_rec(r0, r1), bx0, by0, ex, ey);
            if (reach) { me->r0 = r0; me->r1 = r1; me->b = r2.x; me->idx = range.x + base + lane; }
        }
        uint32_t m = __ballot_sync(0xFFFFFFFFu, reach);
        uint32_t w0 = 0u, w1 = 0u;                                   // VIS: the warp's visit word of this lane's entry
        __syncwarp();
        while (m) {                                                  // :334 in list order
            const uint32_t kb = (uint32_t)__ffs(m) - 1u;
            m &= m - 1u;
            const uint32_t addr = ebase + kb * (uint32_t)sizeof(FwdEnt);
            const float4 a = gs::lds128(addr);
            const float4 b = gs::lds128(addr + 16u);
            const float2 c = gs::lds64(addr + 32u);
            const uint32_t idx = __float_as_uint(c.y);
            const gs::f32x2 sx = gs::bc2(a.x), sy = gs::bc2(a.y), Ah = gs::bc2(a.z), Bn = gs::bc2(a.w), Ch = gs::bc2(b.x),
                            op = gs::bc2(b.y), cr = gs::bc2(b.z), cg = gs::bc2(b.w), cb = gs::bc2(c.x);
            const gs::f32x2 dx = gs::fsub2(X2, sx), dy = gs::fsub2(Y2, sy);   // :347
            const gs::f32x2 pw2 = gs::blend_power2(Ah, Bn, Ch, dx, dy);
            const gs::f32x2 G2 = gs::exp_blend2(pw2);                             // :362
            const gs::f32x2 a2 = gs::fmul2(op, G2);                               // :363
            float pw[2], al[2], Tn[2], To[2];
            gs::up2(pw2, pw[0], pw[1]); gs::up2(a2, al[0], al[1]); gs::up2(T2, To[0], To[1]);
            bool c2[2];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                al[h] = fminf(al[h], GS_ALPHA_MAX);
                // :334 (the pixel still walks: T > 1e-4), :359, :366
                c2[h] = (To[h] > GS_T_EPS) && (pw[h] <= 0.0f) && !(pw[h] < -4.5f) && !(al[h] < GS_ALPHA_MIN);
                al[h] = c2[h] ? al[h] : 0.0f;
            }
            const gs::f32x2 alpha = gs::pk2(al[0], al[1]);
            const gs::f32x2 w2 = gs::fmul2(alpha, T2);                          // :369
            Cr2 = gs::ffma2(cr, w2, Cr2); Cg2 = gs::ffma2(cg, w2, Cg2); Cb2 = gs::ffma2(cb, w2, Cb2);
            T2 = gs::fmul2(T2, gs::fsub2(gs::bc2(1.0f), alpha));            // :370
            gs::up2(T2, Tn[0], Tn[1]);
            bool alive = false;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                Tp[h] = c2[h] ? To[h] : Tp[h];
                last[h] = c2[h] ? idx : last[h];                                  // :372
                alive |= Tn[h] > GS_T_EPS;                                        // :334 loop condition
            }
            if (VIS) {
                const uint32_t m0 = __ballot_sync(0xFFFFFFFFu, c2[0]), m1 = __ballot_sync(0xFFFFFFFFu, c2[1]);
                if (lane == kb) { w0 = m0; w1 = m1; }
            }
            if (!__any_sync(0xFFFFFFFFu, alive)) { wlive = false; break; }
        }
        if (VIS) { if (!(v >> 31)) vis[(size_t)wu * vstride + range.x + base + lane] = make_uint2(w0, w1); }   // 256 contiguous bytes per chunk
        __syncwarp();
    }
    if (degen) tile_flags[tile] = 1u;
    float Tf[2], Cr[2], Cg[2], Cb[2];
    gs::up2(T2, Tf[0], Tf[1]); gs::up2(Cr2, Cr[0], Cr[1]); gs::up2(Cg2, Cg[0], Cg[1]); gs::up2(Cb2, Cb[0], Cb[1]);
#pragma unroll
    for (int j = 0; j < 2; j++) {
        if (px[j] < W && py[j] < H) {
            const size_t pix = (size_t)py[j] * W + px[j];
            image[pix] = make_float4(gs::fadd(Cr[j], Tf[j]), gs::fadd(Cg[j], Tf[j]), gs::fadd(Cb[j], Tf[j]), 1.0f);   // :377-384 white background
            last_idx[pix] = last[j];                                  // :382
            tfinal[pix] = (Tf[j] < GS_T_EPS) ? Tp[j] : Tf[j];         // as in blend_forward_kernel
        }
    }
}

}  // namespace

static int g_fwd_ppt = GS_FWD_PPT;
// forward walk: 0 = CTA-cooperative rounds, 1 = warp-autonomous when the frame records visit words (training), cooperative for
// render-only frames, 2 = warp-autonomous always.  Measured on one B200 (profiles/r2_fwd_warp_ab.txt): training forward 227 ->
// 215 us at C2, 447 -> 387 us on a C5 view; render-only 435 -> 409 us on a C5 view and 272 -> 251 us at C3 (dense lists, many
// saturated pixels), but 599 -> 609 us at C4 and 211 -> 214 us at C2 (tiny splats: an entry reaches one warp of four and the
// per-chunk footprint tests weigh more than the barriers did), hence 1.
static int g_fwd_warp = 1;
void gs_tune_fwd_ppt(int v) { if (v == 2 || v == 4) g_fwd_ppt = v; }
void gs_tune_fwd_warp(int v) { if (v >= 0 && v <= 2) g_fwd_warp = v; }
// the visit words of the warp-autonomous forward end with each warp's own walk: only the warp-autonomous backward may read them
bool gs_fwd_words_end_with_the_warp() { return g_fwd_ppt == 2 && g_fwd_warp != 0; }
bool gs_fwd_writes_visits() { return g_fwd_ppt == 2; }

void gs_launch_tile_ranges(const uint64_t* sorted_keys, uint32_t num_pairs, uint32_t num_tiles, uint32_t* /*lb*/,
                           uint2* ranges, cudaStream_t st) {
    if (num_tiles == 0) return;
    tile_ranges_kernel<<<(num_tiles + 127) / 128, 128, 0, st>>>(sorted_keys, num_pairs, num_tiles, ranges);
}

void gs_launch_tile_ranges32(const uint32_t* sorted_tiles, uint32_t max_pairs, const uint32_t* d_n, uint32_t num_tiles,
                             uint2* ranges, cudaStream_t st) {
    if (num_tiles == 0) return;
    tile_starts_kernel<<<148 * 8, 256, 0, st>>>(sorted_tiles, max_pairs, d_n, num_tiles, ranges);
    tile_counts_kernel<<<(num_tiles + 255) / 256, 256, 0, st>>>(max_pairs, d_n, num_tiles, ranges);
}
void gs_launch_rebuild_keys64(const uint32_t* sorted_tiles, const uint32_t* sorted_vals, const uint32_t* dkey, uint32_t n,
                              uint64_t* keys, cudaStream_t st) {
    if (n) rebuild_keys64_kernel<<<(n + 255) / 256, 256, 0, st>>>(sorted_tiles, sorted_vals, dkey, n, keys);
}

void gs_launch_tile_order(const uint2* ranges, uint32_t num_tiles, uint32_t* order, cudaStream_t st) {
    if (num_tiles) tile_order_kernel<<<1, 1024, 0, st>>>(ranges, num_tiles, order);
}

void gs_launch_blend_forward(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                             uint32_t ntx, uint32_t nty, float4* image, uint32_t* last_idx, float* tfinal, int order_ready,
                             uint2* vis, size_t vstride, cudaStream_t st) {
    if (ntx * nty == 0) return;
#define GS_FWD_LAUNCH(PPT, VIS)                                                                                       \
    blend_forward_kernel<PPT, false, VIS><<<ntx * nty, 256 / PPT, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, image, last_idx, \
                                                                            tfinal, O.tile_flags, O.tile_order, nullptr, vis, vstride)
    if (O.tile_order && !order_ready) tile_order_kernel<<<1, 1024, 0, st>>>(ranges, ntx * nty, O.tile_order);
#define GS_FWD_WARP_LAUNCH(VIS)                                                                                       \
    blend_forward_warp_kernel<VIS><<<ntx * nty, 128, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, image, last_idx, tfinal, \
                                                               O.tile_flags, O.tile_order, vis, vstride)
    if (g_fwd_ppt == 4) GS_FWD_LAUNCH(4, false);
    else if (vis && g_fwd_warp) GS_FWD_WARP_LAUNCH(true);
    else if (g_fwd_warp == 2) GS_FWD_WARP_LAUNCH(false);
    else if (vis) GS_FWD_LAUNCH(2, true);
    else GS_FWD_LAUNCH(2, false);
}

// counts E (entries visited) and C (contributing evaluations) of the forward walk; counters[2] zeroed by the caller
void gs_launch_blend_count(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                           uint32_t ntx, uint32_t nty, unsigned long long* counters, cudaStream_t st) {
    if (ntx * nty == 0) return;
    blend_forward_kernel<2, true, false><<<ntx * nty, 128, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, nullptr, nullptr, nullptr,
                                                             O.tile_flags, nullptr, counters, nullptr, 0);
}
