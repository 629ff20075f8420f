// gs_common.cuh — shared device helpers: pinned fp32 arithmetic, constants, error plumbing.
//
// "Pinned arithmetic" (DESIGN.md): the kernels that decide sort keys, tile lists, tile ranges,
// lastContribIdx and the forward image spell every rounding explicitly with __f*_rn intrinsics
// (never contracted by nvcc) so they reproduce the CPU oracle bit for bit.  The exp() used there is
// gs_exp2_poly(): exp2(x*log2e) by a degree-5 polynomial built from + * fma and one integer add.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GS_TILE 16u
#define GS_SH_C0 0.28209479177387814f      // tiled_shaders.metal:83
#define GS_MAX_RADIUS 512.0f               // tiled_shaders.metal:85
#define GS_MAX_SCALE 5.0f                  // tiled_shaders.metal:87
#define GS_MAX_SCALE_TRAIN 4.0f            // shaders.metal:55
#define GS_MIN_OPACITY 0.005f              // tiled_shaders.metal:742
#define GS_MAX_TILES_PER_GAUSSIAN 256u     // tiled_shaders.metal:743
#define GS_ALPHA_MIN (1.0f / 255.0f)       // tiled_shaders.metal:366
#define GS_ALPHA_MAX 0.99f                 // tiled_shaders.metal:363
#define GS_T_EPS 0.0001f                   // tiled_shaders.metal:334,458
#define GS_CULL_BIT 0x80000000u            // pair value bit 31: this (tile, Gaussian) pair can never contribute (pairs.cu)
#define GS_ID_MASK 0x7FFFFFFFu

#define GS_LOG2E 1.44269504088896340736f
#define GS_EXP2_C1 0x1.62e42ap-1f
#define GS_EXP2_C2 0x1.ebf9bcp-3f
#define GS_EXP2_C3 0x1.c6b752p-5f
#define GS_EXP2_C4 0x1.3cea88p-7f
#define GS_EXP2_C5 0x1.5bba14p-10f
#define GS_EXP2_MAGIC 12582912.0f          // 1.5 * 2^23

namespace gs {

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }
// MSL clamp == min(max(x, lo), hi); fminf/fmaxf return the non-NaN operand, like C.
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// 2^t, |t| < 126: r's low mantissa bits hold round(t); f = t - round(t) in [-0.5, 0.5].
__device__ __forceinline__ float exp2_poly(float t) {
    float r = fadd(t, GS_EXP2_MAGIC);
    float nf = fsub(r, GS_EXP2_MAGIC);
    float f = fsub(t, nf);
    float p = GS_EXP2_C5;
    p = ffma(p, f, GS_EXP2_C4);
    p = ffma(p, f, GS_EXP2_C3);
    p = ffma(p, f, GS_EXP2_C2);
    p = ffma(p, f, GS_EXP2_C1);
    p = ffma(p, f, 1.0f);
    return __uint_as_float(__float_as_uint(p) + (__float_as_uint(r) << 23));
}
// blend exponent already known to lie in [-4.5, 0]
__device__ __forceinline__ float exp_blend(float x) { return exp2_poly(fmul(x, GS_LOG2E)); }
// general exp, exponent clamped to [-125, 125] (NaN -> -125)
__device__ __forceinline__ float exp_pinned(float x) {
    return exp2_poly(fminf(fmaxf(fmul(x, GS_LOG2E), -125.0f), 125.0f));
}

// pinned blend exponent: power = fma(dx, fma(Ah, dx, Bn*dy), (Ch*dy)*dy); Ah=-A/2, Bn=-B, Ch=-C/2
__device__ __forceinline__ float blend_power(float Ah, float Bn, float Ch, float dx, float dy) {
    float u = ffma(Ah, dx, fmul(Bn, dy));
    float v = fmul(fmul(Ch, dy), dy);
    return ffma(dx, u, v);
}

// ---- packed fp32 pairs (sm_100 FFMA2): two IEEE round-to-nearest fp32 operations per issued instruction.  The blend
// kernels keep TWO pixels of a thread in one pair, which halves the FP32 issue slots of the walk.  Only fma.rn.f32x2 is
// used: ptxas contracts mul.rn.f32x2 followed by add.rn.f32x2 into one FFMA2 (measured: tools/ubench/f32x2.cu), which
// would break the pinned rounding sequence, so mul/add/sub are spelled as exact FMAs (a*b + -0, a*1 + b, b*-1 + a).
// fma.rn.f32x2 is bit-identical to __fmaf_rn per half (same micro-check).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ f32x2 bc2(float x) { return pk2(x, x); }
__device__ __forceinline__ void up2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 fmul2(f32x2 a, f32x2 b) { return ffma2(a, b, pk2(-0.0f, -0.0f)); }
__device__ __forceinline__ f32x2 fadd2(f32x2 a, f32x2 b) { return ffma2(a, pk2(1.0f, 1.0f), b); }
__device__ __forceinline__ f32x2 fsub2(f32x2 a, f32x2 b) { return ffma2(b, pk2(-1.0f, -1.0f), a); }
// exp_blend() on a pair (same operation sequence per half)
__device__ __forceinline__ f32x2 exp_blend2(f32x2 x) {
    const f32x2 t = fmul2(x, bc2(GS_LOG2E));
    const f32x2 r = fadd2(t, bc2(GS_EXP2_MAGIC));
    const f32x2 nf = fadd2(r, bc2(-GS_EXP2_MAGIC));
    const f32x2 f = fsub2(t, nf);
    f32x2 p = bc2(GS_EXP2_C5);
    p = ffma2(p, f, bc2(GS_EXP2_C4));
    p = ffma2(p, f, bc2(GS_EXP2_C3));
    p = ffma2(p, f, bc2(GS_EXP2_C2));
    p = ffma2(p, f, bc2(GS_EXP2_C1));
    p = ffma2(p, f, bc2(1.0f));
    float p0, p1, r0, r1;
    up2(p, p0, p1); up2(r, r0, r1);
    return pk2(__uint_as_float(__float_as_uint(p0) + (__float_as_uint(r0) << 23)),
               __uint_as_float(__float_as_uint(p1) + (__float_as_uint(r1) << 23)));
}
// pinned blend exponent on a pair (blend_power per half); the per-entry coefficients are broadcast
__device__ __forceinline__ f32x2 blend_power2(f32x2 Ah, f32x2 Bn, f32x2 Ch, f32x2 dx, f32x2 dy) {
    const f32x2 u = ffma2(Ah, dx, fmul2(Bn, dy));
    const f32x2 v = fmul2(fmul2(Ch, dy), dy);
    return ffma2(dx, u, v);
}
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr)); return v;
}
__device__ __forceinline__ float2 lds64(uint32_t addr) {
    float2 v; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr)); return v;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Pixel (within the 16x16 tile) of slot j of a thread, for PPT pixels per thread (256/PPT threads per tile):
//   PPT 1: 8 warps, each an 8x4 patch;  PPT 2: 4 warps, each an 8x8 block, a lane owns a vertical pair;
//   PPT 4: 2 warps, each a 16x8 strip, a lane owns a 2x2 quad.
template <int PPT>
__device__ __forceinline__ void tile_pixel(uint32_t warp, uint32_t lane, int j, uint32_t& x, uint32_t& y) {
    if (PPT == 1) { x = (warp & 1u) * 8u + (lane & 7u); y = (warp >> 1) * 4u + (lane >> 3); }
    else if (PPT == 2) { x = (warp & 1u) * 8u + (lane & 7u); y = (warp >> 1) * 8u + (lane >> 3) * 2u + (uint32_t)j; }
    else { x = (lane & 7u) * 2u + ((uint32_t)j & 1u); y = warp * 8u + (lane >> 3) * 2u + ((uint32_t)j >> 1); }
}

// ---- conservative footprint test (pairs.cu explains it): can a splat contribute to ANY pixel centre of the box whose first
// pixel centre is (px0,py0) and which extends ex, ey further?  Used per (tile, Gaussian) at emit and per (warp region,
// staged entry) inside the blend kernels.
struct CullRec { float sx, sy, A, B, C, qthr, iA, iC; };
__device__ __forceinline__ CullRec cull_rec(const float4 r0, const float4 r1) {
    CullRec c;
    c.sx = r0.x; c.sy = r0.y; c.A = -2.0f * r0.z; c.B = -r0.w; c.C = -2.0f * r1.x;
    c.qthr = fminf(9.0f, 2.0f * __logf(255.0f * r1.y)) + 2e-3f;
    if (!(c.qthr == c.qthr)) c.qthr = 9.002f;
    c.iA = 1.0f / c.A; c.iC = 1.0f / c.C;       // only steer where on an edge q is sampled; the margin absorbs their rounding
    return c;
}
__device__ __forceinline__ float edge_min_q(float A, float B, float C, float iC, float cx, float ylo, float yhi) {
    // min over dy in [ylo,yhi] of A cx^2 + 2 B cx dy + C dy^2
    const float dy = fminf(fmaxf(-B * cx * iC, ylo), yhi);
    return A * cx * cx + 2.0f * B * cx * dy + C * dy * dy;
}
__device__ __forceinline__ bool box_cannot_contribute(const CullRec& c, float px0, float py0, float ex, float ey) {
    const float x0 = px0 - c.sx, x1 = x0 + ex;                                 // dx range of the pixel centres
    const float y0 = py0 - c.sy, y1 = y0 + ey;
    const bool inx = x0 <= 0.f && x1 >= 0.f, iny = y0 <= 0.f && y1 >= 0.f;
    if (inx && iny) return false;                                              // centre inside the box: q_min = 0
    if (!(c.A > 0.f) || !(c.C > 0.f)) return false;                            // degenerate / non-finite conic: keep
    float q = 3.0e38f;
    if (!inx) q = fminf(q, edge_min_q(c.A, c.B, c.C, c.iC, x0 > 0.f ? x0 : x1, y0, y1));
    if (!iny) q = fminf(q, edge_min_q(c.C, c.B, c.A, c.iA, y0 > 0.f ? y0 : y1, x0, x1));
    const float Dx = fmaxf(fabsf(x0), fabsf(x1)), Dy = fmaxf(fabsf(y0), fabsf(y1));
    const float noise = 4e-6f * (c.A * Dx * Dx + 2.0f * fabsf(c.B) * Dx * Dy + c.C * Dy * Dy);   // ~32 ulp of the largest term
    return q > c.qthr + noise;                                                 // NaN compares false: keep
}
// the pixel rectangle (within the tile) that warp `warp` owns under tile_pixel<PPT>: first pixel and extent
template <int PPT>
__device__ __forceinline__ void warp_region(uint32_t warp, uint32_t& rx, uint32_t& ry, float& ex, float& ey) {
    if (PPT == 1) { rx = (warp & 1u) * 8u; ry = (warp >> 1) * 4u; ex = 7.f; ey = 3.f; }
    else if (PPT == 2) { rx = (warp & 1u) * 8u; ry = (warp >> 1) * 8u; ex = 7.f; ey = 7.f; }
    else { rx = 0u; ry = warp * 8u; ex = 15.f; ey = 7.f; }
}

// exclusive scan of one small count per thread over a CTA of NW warps; returns the exclusive prefix, *total = CTA sum.
// s_w: NW words of shared scratch; contains one __syncthreads().
template <int NW>
__device__ __forceinline__ uint32_t cta_excl_scan(uint32_t x, uint32_t* s_w, uint32_t* total) {
    uint32_t inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane_id() >= (uint32_t)d) inc += t; }
    if (lane_id() == 31u) s_w[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) { const uint32_t c = s_w[w]; if (w < (int)(threadIdx.x >> 5)) off += c; tot += c; }
    *total = tot;
    return off + inc - x;
}

// ---- block-sum look-back in two dependent round trips, whatever the tile count.  Every tile (ticket order) publishes its
// sum in status[tile] and adds it to its group's word (32 tiles per group; the contributor count rides in bits 40+).  A
// tile's exclusive prefix = the earlier tiles of its own group + the completed earlier groups: it never waits for another
// tile's PREFIX, only for earlier tiles to have published their own sums.  Called by one full warp; all words start at 0.
__device__ __forceinline__ uint32_t lookback_excl_warp(uint32_t tile, uint32_t block_sum, unsigned long long* status,
                                                       unsigned long long* group_sum, uint32_t lane) {
    if (lane == 0) {
        const unsigned long long mine = (1ull << 40) | (unsigned long long)block_sum;
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(status + tile), "l"(mine) : "memory");
        atomicAdd(group_sum + (tile >> 5), mine);
    }
    const uint32_t grp = tile >> 5, first = grp << 5;
    uint32_t v = 0;
    if (first + lane < tile) {
        unsigned long long s;
        do { asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(s) : "l"(status + first + lane) : "memory"); } while ((s >> 40) == 0ull);
        v = (uint32_t)s;
    }
    for (uint32_t g0 = 0; g0 < grp; g0 += 32u) {
        if (g0 + lane < grp) {
            unsigned long long s;
            do { asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(s) : "l"(group_sum + g0 + lane) : "memory"); } while ((s >> 40) != 32ull);
            v += (uint32_t)s;
        }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    return v;
}

// relaxed gpu-scope load/store for look-back status words
__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Launch order of the per-tile blend kernels: order[b] = the tile CTA b takes, tiles by descending list length (256
// buckets of the frame's longest list, counting sort by ONE CTA of NT threads), so the long tiles start first and the
// last wave is made of short ones.  Results do not depend on the order (gradient atomics are unordered either way).
template <int NT>
__device__ __forceinline__ void tile_order_cta(const uint2* __restrict__ ranges, uint32_t nt, uint32_t* __restrict__ order) {
    static_assert(NT >= 256 && NT % 32 == 0, "one thread per bucket");
    __shared__ uint32_t s_hist[256], s_ws[8], s_max;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (tid < 256) s_hist[tid] = 0;
    if (tid == 0) s_max = 0;
    __syncthreads();
    uint32_t m = 0;
    for (uint32_t i = tid; i < nt; i += NT) m = max(m, ranges[i].y);
    m = __reduce_max_sync(0xFFFFFFFFu, m);
    if (lane == 0 && m) atomicMax(&s_max, m);
    __syncthreads();
    const float sc = 255.0f / (float)max(s_max, 1u);
    for (uint32_t i = tid; i < nt; i += 4 * NT) {                 // four loads in flight: one CTA walks every tile
        uint32_t c[4];
#pragma unroll
        for (int u = 0; u < 4; u++) c[u] = (i + u * NT < nt) ? ranges[i + u * NT].y : 0u;
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (i + u * NT < nt) atomicAdd(&s_hist[255u - min(255u, (uint32_t)((float)c[u] * sc))], 1u);
    }
    __syncthreads();
    uint32_t v = 0, inc = 0;
    if (tid < 256) {
        v = s_hist[tid]; inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane >= (uint32_t)d) inc += t; }
        if (lane == 31) s_ws[warp] = inc;
    }
    __syncthreads();
    if (tid < 256) {
        uint32_t pre = 0;
        for (uint32_t w = 0; w < warp; w++) pre += s_ws[w];
        s_hist[tid] = pre + inc - v;
    }
    __syncthreads();
    for (uint32_t i = tid; i < nt; i += 4 * NT) {
        uint32_t c[4];
#pragma unroll
        for (int u = 0; u < 4; u++) c[u] = (i + u * NT < nt) ? ranges[i + u * NT].y : 0u;
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (i + u * NT < nt) order[atomicAdd(&s_hist[255u - min(255u, (uint32_t)((float)c[u] * sc))], 1u)] = i + u * NT;
    }
}

}  // namespace gs

// ---------------------------------------------------------------------------------------------
// Device-resident layouts (DESIGN.md "Data layout in HBM")
// ---------------------------------------------------------------------------------------------
// Model state, SoA with plane stride = capacity (floats):
//   pos[3] scale[3] opacity[1] sh[12] planes, and rotation as float4[capacity] (16-B vector loads).
struct GsParams {
    float* pos;      // 3 planes
    float* scale;    // 3 planes (log scale)
    float4* rot;     // (w,x,y,z)
    float* opacity;  // raw
    float* sh;       // 12 planes
    uint32_t cap;    // plane stride
};

// Per-Gaussian projection record.  rec: 3 x float4 per Gaussian (48-B gather unit for the blend kernels)
//   rec[3i+0] = {sx, sy, Ah, Bn}   rec[3i+1] = {Ch, opacity, r, g}   rec[3i+2] = {b, depth, radius, flags}
// flags bit0: |A|+|B|+|C| < 1e-4 (forward-only skip, tiled_shaders.metal:350-351)
struct GsProjected {
    float4* rec;
    uint4* bbox;     // tileMinX, tileMinY, tileMaxX, tileMaxY
    float4* aux;     // viewPos.x, viewPos.y, cov2D.a, cov2D.b
    float* covc;     // cov2D.c
    uint32_t* cnt;   // tiles this Gaussian emits (0 = culled / fails generateTilePairs' admission tests)
    uint32_t* ccnt;  // coarse bins (8x8 tiles) its bounding box touches, 0 when cnt == 0 (binning.cu)
    uint32_t* dkey;  // sortable depth bits (tiled_shaders.metal:773-774); 0xFFFFFFFF when cnt == 0
    uint32_t* offs;  // exclusive scan of cnt in emission order (Gaussian order, or depth order in hoisted mode); [n] = total
    uint32_t* tile_flags;  // per TILE: 1 if the forward walked a conic-degenerate entry (backward then redoes pass 1)
    uint32_t* tile_order;  // per CTA of the blend kernels: the tile it takes (longest lists first), or nullptr = raster order
};

// Packed gradient buffer: 16 planes of stride `stride` floats:
//   0-2 position, 3 opacity, 4-6 scale, 7-10 rotation, 11-13 SH-DC (R,G,B), 14-15 view-space x,y
//   (or, in per-view-stat mode, 14 = sum of clamped |viewspace grad|, 15 = number of views counted).
struct GsGradPack {
    float* g;
    uint32_t stride;
};
