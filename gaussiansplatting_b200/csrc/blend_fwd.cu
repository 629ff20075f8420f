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

constexpr int FB = 256;   // staged entries per round

template <bool COUNT>
__global__ void __launch_bounds__(256) blend_forward_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ svals,
                                                            const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                            float4* __restrict__ image, uint32_t* __restrict__ last_idx,
                                                            float* __restrict__ tfinal, uint32_t* __restrict__ tile_flags,
                                                            unsigned long long* __restrict__ counters) {
    __shared__ float4 s_r0[FB], s_r1[FB];
    __shared__ float s_b[FB];
    __shared__ uint32_t s_degen;
    const uint32_t tile = blockIdx.x;
    const uint32_t tX = tile % ntx, tY = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t px = tX * 16u + (warp & 1u) * 8u + (lane & 7u);
    const uint32_t py = tY * 16u + (warp >> 1) * 4u + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px + 0.5f, pyf = (float)py + 0.5f;      // :328
    const uint2 range = ranges[tile];                                // :320-323
    if (threadIdx.x == 0) s_degen = 0;

    float Cr = 0.f, Cg = 0.f, Cb = 0.f, T = 1.0f, Tp = 1.0f;
    uint32_t nE = 0, nC = 0;   // COUNT only: list entries visited / contributing evaluations (roofline accounting)
    uint32_t last = 0xFFFFFFFFu;
    bool done = !inside;
    for (uint32_t base = 0; base < range.y; base += FB) {
        if (__syncthreads_and(done)) break;
        const uint32_t j = base + threadIdx.x;
        if (j < range.y) {
            const uint32_t g = svals[range.x + j];
            const float4 r0 = rec[3 * (size_t)g], r2 = rec[3 * (size_t)g + 2];
            float4 r1 = rec[3 * (size_t)g + 1];
            if (__float_as_uint(r2.w) & 1u) { r1.y = 0.f; s_degen = 1u; }   // :350-351 conic-magnitude skip (forward only)
            s_r0[threadIdx.x] = r0; s_r1[threadIdx.x] = r1; s_b[threadIdx.x] = r2.x;
        }
        __syncthreads();
        if (!done) {
            const uint32_t cnt = min((uint32_t)FB, range.y - base);
            for (uint32_t k = 0; k < cnt; k++) {                     // :334
                const float4 a = s_r0[k];
                const float4 b = s_r1[k];
                if (COUNT) nE++;
                const float dx = gs::fsub(pxf, a.x), dy = gs::fsub(pyf, a.y);   // :347
                const float power = gs::blend_power(a.z, a.w, b.x, dx, dy);
                if (!(power <= 0.0f) || power < -4.5f) continue;      // :359
                const float G = gs::exp_blend(power);                 // :362
                const float alpha = fminf(gs::fmul(b.y, G), GS_ALPHA_MAX);   // :363
                if (alpha < GS_ALPHA_MIN) continue;                   // :366
                const float w = gs::fmul(alpha, T);                   // :369
                Cr = gs::ffma(b.z, w, Cr); Cg = gs::ffma(b.w, w, Cg); Cb = gs::ffma(s_b[k], w, Cb);
                Tp = T;
                if (COUNT) nC++;
                T = gs::fmul(T, gs::fsub(1.0f, alpha));               // :370
                last = range.x + base + k;                            // :372
                if (!(T > GS_T_EPS)) { done = true; break; }          // :334 loop condition
            }
        }
    }
    if (COUNT) {
        unsigned long long e = nE, c = nC;
        for (int d = 16; d >= 1; d >>= 1) { e += __shfl_xor_sync(0xFFFFFFFFu, e, d); c += __shfl_xor_sync(0xFFFFFFFFu, c, d); }
        if (lane == 0) { atomicAdd(&counters[0], e); atomicAdd(&counters[1], c); }
        return;
    }
    if (inside) {
        const size_t pix = (size_t)py * W + px;
        image[pix] = make_float4(gs::fadd(Cr, T), gs::fadd(Cg, T), gs::fadd(Cb, T), 1.0f);   // :377-384 white background
        last_idx[pix] = last;                                         // :382
        // T_final as tiledBackward's first pass (:430-460) would recompute it: the entry that drives T below
        // 1e-4 is NOT applied there (break before apply), every earlier one is.
        tfinal[pix] = (T < GS_T_EPS) ? Tp : T;
    }
    __syncthreads();
    if (threadIdx.x == 0) tile_flags[tile] = s_degen;
}

}  // namespace

void gs_launch_tile_ranges(const uint64_t* sorted_keys, uint32_t num_pairs, uint32_t num_tiles, uint32_t* /*lb*/,
                           uint2* ranges, cudaStream_t st) {
    if (num_tiles == 0) return;
    tile_ranges_kernel<<<(num_tiles + 127) / 128, 128, 0, st>>>(sorted_keys, num_pairs, num_tiles, ranges);
}

void gs_launch_blend_forward(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                             uint32_t ntx, uint32_t nty, float4* image, uint32_t* last_idx, float* tfinal, cudaStream_t st) {
    if (ntx * nty == 0) return;
    blend_forward_kernel<false><<<ntx * nty, 256, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, image, last_idx, tfinal,
                                                           O.tile_flags, nullptr);
}

// counts E (entries visited) and C (contributing evaluations) of the forward walk; counters[2] zeroed by the caller
void gs_launch_blend_count(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                           uint32_t ntx, uint32_t nty, unsigned long long* counters, cudaStream_t st) {
    if (ntx * nty == 0) return;
    blend_forward_kernel<true><<<ntx * nty, 256, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, nullptr, nullptr, nullptr,
                                                          O.tile_flags, counters);
}
