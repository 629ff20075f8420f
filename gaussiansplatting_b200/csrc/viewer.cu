// viewer.cu — the reference's interactive splat renderer as CUDA compute into a headless image.
//
// Replaces vertexShader / fragmentShader (shaders.metal:160-317) and the fixed-function raster + blend state of
// MTLEngine::render (mtl_engine.mm:354-362, 575-605).  The reference draws one instanced quad per Gaussian in the
// back-to-front order of GPURadixSort32 and lets the hardware blend (ONE, ONE_MINUS_SRC_ALPHA) over a white clear.  Here:
//   V1 view_splat_kernel   per Gaussian: the vertex stage (validity, clip, cov3D with the +-8 clamp and the RAW quaternion,
//                          cov2D, radius / conic with the viewer's guards, colour, opacity) -> a 48-B splat record, the
//                          16x16-tile box of the quad and its tile count
//   V2 scan of the counts THROUGH the draw order + view_emit_kernel: (tile id, Gaussian) pairs in draw order
//   V3 stable onesweep passes over the tile-id bits + run boundaries  -> per-tile instance lists in draw order
//   V4 view_blend_kernel   one CTA per tile, one pixel per thread: walk the tile's list in draw order, shade the pixel
//                          centre exactly as the rasteriser would (top-left fill rule on the quad, interpolated quadOffset,
//                          the two discards) and blend.  Back-to-front order has no early exit: every entry is visited.
// Every rounding is spelled out in the oracle's order (oracle/viewer_oracle.cpp), so the image is bit-identical to it.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

struct ViewUni {
    float V[16];
    float VP[16];
    float sw, sh, fx, fy;
    uint32_t n;
};

__device__ __forceinline__ void mulp(const float* m, float x, float y, float z, float out[4]) {
#pragma unroll
    for (int r = 0; r < 4; r++)
        out[r] = gs::fadd(gs::fadd(gs::fadd(gs::fmul(m[r], x), gs::fmul(m[4 + r], y)), gs::fmul(m[8 + r], z)), m[12 + r]);
}
__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return gs::fadd(gs::fadd(gs::fmul(a0, b0), gs::fmul(a1, b1)), gs::fmul(a2, b2));
}

__global__ void __launch_bounds__(256) view_splat_kernel(GsParams P, const __grid_constant__ ViewUni u, float4* __restrict__ rec,
                                                         uint4* __restrict__ bbox, uint32_t* __restrict__ cnt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= u.n) return;
    float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0;
    uint4 bb = make_uint4(0u, 0u, 0u, 0u);
    uint32_t count = 0;
    const uint32_t cap = P.cap;
    const float px = P.pos[i], py = P.pos[cap + i], pz = P.pos[2 * cap + i];
    do {
        if (isnan(px) || isnan(py) || isnan(pz) || fabsf(px) > 1e6f || fabsf(py) > 1e6f || fabsf(pz) > 1e6f) break;   // :175-180
        float cp[4], t[4];
        mulp(u.VP, px, py, pz, cp);                                        // :183
        if (cp[3] <= 0.0f) break;                                          // :186-190
        if (!(cp[2] >= 0.0f && cp[2] <= cp[3])) break;                     // the quad shares z, w: clipped as a whole
        const float4 q = P.rot[i];
        const float w = q.x, x = q.y, y = q.z, z = q.w;                    // raw quaternion, :65-76
        const float R00 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(y, y), gs::fmul(z, z))));
        const float R10 = gs::fmul(2.0f, gs::fadd(gs::fmul(x, y), gs::fmul(w, z)));
        const float R20 = gs::fmul(2.0f, gs::fsub(gs::fmul(x, z), gs::fmul(w, y)));
        const float R01 = gs::fmul(2.0f, gs::fsub(gs::fmul(x, y), gs::fmul(w, z)));
        const float R11 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(x, x), gs::fmul(z, z))));
        const float R21 = gs::fmul(2.0f, gs::fadd(gs::fmul(y, z), gs::fmul(w, x)));
        const float R02 = gs::fmul(2.0f, gs::fadd(gs::fmul(x, z), gs::fmul(w, y)));
        const float R12 = gs::fmul(2.0f, gs::fsub(gs::fmul(y, z), gs::fmul(w, x)));
        const float R22 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(x, x), gs::fmul(y, y))));
        const float s0 = gs::exp_pinned(gs::clampf(P.scale[i], -8.0f, 8.0f));             // :83
        const float s1 = gs::exp_pinned(gs::clampf(P.scale[cap + i], -8.0f, 8.0f));
        const float s2 = gs::exp_pinned(gs::clampf(P.scale[2 * cap + i], -8.0f, 8.0f));
        const float M00 = gs::fmul(R00, s0), M01 = gs::fmul(R01, s1), M02 = gs::fmul(R02, s2);
        const float M10 = gs::fmul(R10, s0), M11 = gs::fmul(R11, s1), M12 = gs::fmul(R12, s2);
        const float M20 = gs::fmul(R20, s0), M21 = gs::fmul(R21, s1), M22 = gs::fmul(R22, s2);
        const float S00 = dot3(M00, M00, M01, M01, M02, M02), S01 = dot3(M00, M10, M01, M11, M02, M12), S02 = dot3(M00, M20, M01, M21, M02, M22);
        const float S10 = dot3(M10, M00, M11, M01, M12, M02), S11 = dot3(M10, M10, M11, M11, M12, M12), S12 = dot3(M10, M20, M11, M21, M12, M22);
        const float S20 = dot3(M20, M00, M21, M01, M22, M02), S21 = dot3(M20, M10, M21, M11, M22, M12), S22 = dot3(M20, M20, M21, M21, M22, M22);

        mulp(u.V, px, py, pz, t);                                          // :97
        float tz = t[2];
        if (tz < 0.001f) tz = 0.001f;                                      // :102
        const float limx = gs::fdiv(gs::fmul(1.3f, u.fx), tz), limy = gs::fdiv(gs::fmul(1.3f, u.fy), tz);
        const float txtz = gs::clampf(gs::fdiv(t[0], tz), -limx, limx);
        const float tytz = gs::clampf(gs::fdiv(t[1], tz), -limy, limy);
        const float J00 = gs::fdiv(u.fx, tz), J02 = gs::fdiv(gs::fmul(-u.fx, txtz), tz);
        const float J11 = gs::fdiv(u.fy, tz), J12 = gs::fdiv(gs::fmul(-u.fy, tytz), tz);
        float T0[3], T1[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            T0[c] = gs::fadd(gs::fmul(J00, u.V[c * 4 + 0]), gs::fmul(J02, u.V[c * 4 + 2]));
            T1[c] = gs::fadd(gs::fmul(J11, u.V[c * 4 + 1]), gs::fmul(J12, u.V[c * 4 + 2]));
        }
        const float U00 = dot3(T0[0], S00, T0[1], S10, T0[2], S20), U01 = dot3(T0[0], S01, T0[1], S11, T0[2], S21), U02 = dot3(T0[0], S02, T0[1], S12, T0[2], S22);
        const float U10 = dot3(T1[0], S00, T1[1], S10, T1[2], S20), U11 = dot3(T1[0], S01, T1[1], S11, T1[2], S21), U12 = dot3(T1[0], S02, T1[1], S12, T1[2], S22);
        float ca = dot3(U00, T0[0], U01, T0[1], U02, T0[2]);
        const float cb = dot3(U00, T1[0], U01, T1[1], U02, T1[2]);
        float cc = dot3(U10, T1[0], U11, T1[1], U12, T1[2]);
        ca = gs::fadd(ca, 0.3f); cc = gs::fadd(cc, 0.3f);                  // :130-131

        const float a = fmaxf(ca, 0.3f), c = fmaxf(cc, 0.3f);              // :201-205
        const float maxB = gs::fmul(gs::fsqrt(gs::fmul(a, c)), 0.99f);
        const float b = gs::clampf(cb, -maxB, maxB);
        const float det = gs::fsub(gs::fmul(a, c), gs::fmul(b, b));
        float radius, k0, k1, k2;
        if (det > 0.0001f) {                                               // :213-238
            const float mid = gs::fmul(0.5f, gs::fadd(a, c));
            const float disc = fmaxf(0.0f, gs::fsub(gs::fmul(mid, mid), det));
            const float lambda = gs::fadd(mid, gs::fsqrt(disc));
            radius = ceilf(gs::fmul(3.0f, gs::fsqrt(fmaxf(lambda, 0.1f))));
            radius = gs::clampf(radius, 1.0f, 512.0f);
            const float invDet = gs::fdiv(1.0f, det);
            k0 = gs::fmul(c, invDet); k1 = gs::fmul(-b, invDet); k2 = gs::fmul(a, invDet);
            const float expected = gs::fdiv(9.0f, gs::fmul(radius, radius));
            const float actual = gs::fmul(0.5f, gs::fadd(k0, k2));
            if (actual < gs::fmul(expected, 0.01f)) {
                const float f = gs::fsqrt(gs::fdiv(expected, fmaxf(actual, 1e-10f)));
                k0 = gs::fmul(k0, f); k1 = gs::fmul(k1, f); k2 = gs::fmul(k2, f);
            }
        } else { radius = 15.0f; k0 = 0.04f; k1 = 0.0f; k2 = 0.04f; }      // :239-246
        const float ndcx = gs::fdiv(cp[0], cp[3]), ndcy = gs::fdiv(cp[1], cp[3]);
        const float cx = gs::fmul(gs::fadd(gs::fmul(ndcx, 0.5f), 0.5f), u.sw);            // viewport: NDC y up, framebuffer y down
        const float cy = gs::fmul(gs::fsub(0.5f, gs::fmul(ndcy, 0.5f)), u.sh);
        if (!(cx == cx) || !(cy == cy)) break;
        const float x0 = fmaxf(floorf(gs::fsub(cx, radius)), 0.0f), x1 = fminf(floorf(gs::fadd(cx, radius)), gs::fsub(u.sw, 1.0f));
        const float y0 = fmaxf(floorf(gs::fsub(cy, radius)), 0.0f), y1 = fminf(floorf(gs::fadd(cy, radius)), gs::fsub(u.sh, 1.0f));
        if (!(x0 <= x1) || !(y0 <= y1)) break;
        const float op = gs::fdiv(1.0f, gs::fadd(1.0f, gs::exp_pinned(-gs::clampf(P.opacity[i], -8.0f, 8.0f))));   // :276
        r0 = make_float4(cx, cy, k0, k1);
        r1 = make_float4(k2, radius, op, gs::clampf(gs::fadd(gs::fmul(GS_SH_C0, P.sh[i]), 0.5f), 0.f, 1.f));
        r2 = make_float4(gs::clampf(gs::fadd(gs::fmul(GS_SH_C0, P.sh[4 * cap + i]), 0.5f), 0.f, 1.f),
                         gs::clampf(gs::fadd(gs::fmul(GS_SH_C0, P.sh[8 * cap + i]), 0.5f), 0.f, 1.f), __uint_as_float(1u), 0.f);
        bb = make_uint4((uint32_t)x0 / GS_TILE, (uint32_t)y0 / GS_TILE, (uint32_t)x1 / GS_TILE, (uint32_t)y1 / GS_TILE);
        count = (bb.z - bb.x + 1u) * (bb.w - bb.y + 1u);
    } while (0);
    rec[3 * (size_t)i + 0] = r0; rec[3 * (size_t)i + 1] = r1; rec[3 * (size_t)i + 2] = r2;
    bbox[i] = bb;
    cnt[i] = count;
}

// draw position j = Gaussian order[j]; its tiles go to offs[j] ... in ty-major, tx-minor order.  One thread per draw position
// writes a small quad's pairs itself; quads above 16 tiles (they reach 65 x 65 at radius 512) are handed to the whole warp, one
// after the other, so that a huge splat does not serialise a lane.
__global__ void __launch_bounds__(256) view_emit_kernel(const uint4* __restrict__ bbox, const uint32_t* __restrict__ cnt,
                                                        const uint32_t* __restrict__ order, const uint32_t* __restrict__ offs,
                                                        uint32_t n, uint32_t ntx, uint32_t* __restrict__ tiles,
                                                        uint32_t* __restrict__ vals, uint32_t cap_pairs) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i = 0, c = 0, off = 0;
    uint4 bb = make_uint4(0u, 0u, 0u, 0u);
    if (j < n) {
        i = order[j];
        if (i < n) { c = cnt[i]; bb = bbox[i]; off = offs[j]; }
    }
    const bool big = c > 16u;
    if (!big && c) {
        const uint32_t wx = bb.z - bb.x + 1u;
        for (uint32_t k = 0; k < c; k++) {
            const uint32_t p = off + k;
            if (p >= cap_pairs) break;
            tiles[p] = (bb.y + k / wx) * ntx + bb.x + k % wx;
            vals[p] = i;
        }
    }
    uint32_t m = __ballot_sync(0xFFFFFFFFu, big);
    while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const uint32_t bi = __shfl_sync(0xFFFFFFFFu, i, src), bc = __shfl_sync(0xFFFFFFFFu, c, src), bo = __shfl_sync(0xFFFFFFFFu, off, src);
        const uint32_t bx = __shfl_sync(0xFFFFFFFFu, bb.x, src), by = __shfl_sync(0xFFFFFFFFu, bb.y, src), bz = __shfl_sync(0xFFFFFFFFu, bb.z, src);
        const uint32_t wx = bz - bx + 1u;
        for (uint32_t k = gs::lane_id(); k < bc; k += 32u) {
            const uint32_t p = bo + k;
            if (p >= cap_pairs) break;
            tiles[p] = (by + k / wx) * ntx + bx + k % wx;
            vals[p] = bi;
        }
    }
}

__device__ __forceinline__ float quant8(float v) { return gs::fdiv(rintf(gs::fmul(fminf(fmaxf(v, 0.0f), 1.0f), 255.0f)), 255.0f); }

// V4: one CTA per tile, one pixel per thread (a warp owns an 8 x 4 pixel patch), entries staged 256 at a time (3 x float4 each)
// through shared memory.  Before a batch is walked, thread t tests staged entry t against the eight warp patches: the quad
// must reach the patch (the rasteriser's own interval test on the patch's extreme pixel centres: exact, rounding is monotone)
// and the splat must be able to pass the two discards somewhere in it (the conservative footprint test of gs_common.cuh:
// power < -4.5 <=> q > 9, alpha < 1/255 <=> q > 2 ln(255 opacity), with its rounding margin).  A warp then walks only the
// entries whose bit is set in its mask word, in draw order.  COUNT (fragment statistics for the parity tests) keeps every
// entry whose quad reaches the patch, so that "shaded" counts what the rasteriser would shade.
template <int QUANT, int COUNT>
__global__ void __launch_bounds__(256) view_blend_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ vals,
                                                         const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                         float4* __restrict__ image, unsigned long long* __restrict__ counters,
                                                         const uint32_t* __restrict__ order) {
    __shared__ float4 s_rec[256 * 3];
    __shared__ uint32_t s_mask[8][8];
    const uint32_t tile = order ? order[blockIdx.x] : blockIdx.x;
    const uint32_t tx = tile % ntx, ty = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t lx = (lane & 7u) + (warp & 1u) * 8u, ly = (lane >> 3) + (warp >> 1) * 4u;
    const uint32_t px = tx * GS_TILE + lx, py = ty * GS_TILE + ly;
    const bool inside = px < W && py < H;
    const float fx = gs::fadd((float)px, 0.5f), fy = gs::fadd((float)py, 0.5f);
    float cr = 1.0f, cg = 1.0f, cb = 1.0f, ca = 1.0f;                               // clear colour, mtl_engine.mm:583
    uint32_t shaded = 0, blended = 0;
    const uint2 rg = ranges[tile];
    for (uint32_t base = 0; base < rg.y; base += 256u) {
        const uint32_t m = min(256u, rg.y - base);
        __syncthreads();
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        if (threadIdx.x < m) {
            const uint32_t id = vals[rg.x + base + threadIdx.x];
            a0 = rec[3 * (size_t)id + 0]; a1 = rec[3 * (size_t)id + 1];
            s_rec[threadIdx.x * 3 + 0] = a0; s_rec[threadIdx.x * 3 + 1] = a1; s_rec[threadIdx.x * 3 + 2] = rec[3 * (size_t)id + 2];
        }
        {   // patch masks: lane tests its staged entry against every warp's 8 x 4 patch
            gs::CullRec c;
            c.sx = a0.x; c.sy = a0.y; c.A = a0.z; c.B = -a0.w; c.C = a1.x;           // pixel y runs down, the conic's y up: B flips
            c.qthr = fminf(9.0f, 2.0f * __logf(255.0f * a1.z)) + 2e-3f;
            if (!(c.qthr == c.qthr)) c.qthr = 9.002f;
            c.iA = 1.0f / c.A; c.iC = 1.0f / c.C;
            const float r = a1.y;
#pragma unroll
            for (int w = 0; w < 8; w++) {
                const float x_lo = gs::fadd((float)(tx * GS_TILE + (uint32_t)(w & 1) * 8u), 0.5f), x_hi = gs::fadd(x_lo, 7.0f);
                const float y_lo = gs::fadd((float)(ty * GS_TILE + (uint32_t)(w >> 1) * 4u), 0.5f), y_hi = gs::fadd(y_lo, 3.0f);
                // some pixel centre of the patch inside the quad?  (dx >= -r && dx < r && dy > -r && dy <= r with dx = x - cx, dy = cy - y)
                bool keep = threadIdx.x < m && gs::fsub(x_hi, a0.x) >= -r && gs::fsub(x_lo, a0.x) < r &&
                            gs::fsub(a0.y, y_lo) > -r && gs::fsub(a0.y, y_hi) <= r;
                if (!COUNT) keep = keep && !gs::box_cannot_contribute(c, x_lo, y_lo, 7.0f, 3.0f);
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, keep);
                if (lane == 0) s_mask[w][warp] = bal;
            }
        }
        __syncthreads();
        if (!inside) continue;
        const uint32_t nwords = (m + 31u) >> 5;
        for (uint32_t wd = 0; wd < nwords; wd++) {
            uint32_t mk = s_mask[warp][wd];
            while (mk) {
                const uint32_t e = wd * 32u + (uint32_t)__ffs(mk) - 1u;
                mk &= mk - 1u;
                const float4 e0 = s_rec[e * 3 + 0], e1 = s_rec[e * 3 + 1];
                const float dx = gs::fsub(fx, e0.x), dy = gs::fsub(e0.y, fy), r = e1.y;
                if (!(dx >= -r && dx < r && dy > -r && dy <= r)) continue;          // top-left fill rule on the quad
                if (COUNT) shaded++;
                const float t1 = gs::fmul(gs::fmul(e0.z, dx), dx), t2 = gs::fmul(gs::fmul(gs::fmul(2.0f, e0.w), dx), dy);
                const float t3 = gs::fmul(gs::fmul(e1.x, dy), dy);
                const float power = gs::fmul(-0.5f, gs::fadd(gs::fadd(t1, t2), t3));    // shaders.metal:296-298
                if (!(power <= 0.0f && power >= -4.5f)) continue;                       // :301
                const float alpha = gs::fmul(e1.z, gs::exp_blend(power));               // :304-305
                if (alpha < GS_ALPHA_MIN) continue;                                     // :308
                if (COUNT) blended++;
                const float4 e2 = s_rec[e * 3 + 2];
                const float om = gs::fsub(1.0f, alpha);
                cr = gs::fadd(gs::fmul(e1.w, alpha), gs::fmul(cr, om));                 // ONE, ONE_MINUS_SRC_ALPHA
                cg = gs::fadd(gs::fmul(e2.x, alpha), gs::fmul(cg, om));
                cb = gs::fadd(gs::fmul(e2.y, alpha), gs::fmul(cb, om));
                ca = gs::fadd(alpha, gs::fmul(ca, om));
                if (QUANT) { cr = quant8(cr); cg = quant8(cg); cb = quant8(cb); ca = quant8(ca); }
            }
        }
    }
    if (inside) image[(size_t)py * W + px] = make_float4(cr, cg, cb, ca);
    if (COUNT) {
        shaded = __reduce_add_sync(0xFFFFFFFFu, shaded);
        blended = __reduce_add_sync(0xFFFFFFFFu, blended);
        if (lane == 0) { atomicAdd(&counters[0], (unsigned long long)shaded); atomicAdd(&counters[1], (unsigned long long)blended); }
    }
}

__global__ void view_splats_to_wire_kernel(const float4* __restrict__ rec, const uint4* __restrict__ bbox, uint32_t n, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 r0 = rec[3 * (size_t)i], r1 = rec[3 * (size_t)i + 1], r2 = rec[3 * (size_t)i + 2];
    const uint4 bb = bbox[i];
    const uint32_t valid = __float_as_uint(r2.z);
    uint32_t* o = out + 16 * (size_t)i;
    o[0] = __float_as_uint(r0.x); o[1] = __float_as_uint(r0.y); o[2] = __float_as_uint(r0.z); o[3] = __float_as_uint(r0.w);
    o[4] = __float_as_uint(r1.x); o[5] = __float_as_uint(r1.y); o[6] = __float_as_uint(r1.z); o[7] = __float_as_uint(r1.w);
    o[8] = __float_as_uint(r2.x); o[9] = __float_as_uint(r2.y); o[10] = valid;
    o[11] = bb.x; o[12] = bb.y; o[13] = bb.z; o[14] = bb.w; o[15] = 0u;
}

}  // namespace

void gs_launch_view_splats(const GsParams& P, const float* V, const float* VP, float sw, float sh, float fx, float fy, uint32_t n,
                           float4* rec, uint4* bbox, uint32_t* cnt, cudaStream_t st) {
    if (n == 0) return;
    ViewUni u;
    for (int k = 0; k < 16; k++) { u.V[k] = V[k]; u.VP[k] = VP[k]; }
    u.sw = sw; u.sh = sh; u.fx = fx; u.fy = fy; u.n = n;
    view_splat_kernel<<<(n + 255) / 256, 256, 0, st>>>(P, u, rec, bbox, cnt);
}
void gs_launch_view_emit(const uint4* bbox, const uint32_t* cnt, const uint32_t* order, const uint32_t* offs, uint32_t n,
                         uint32_t ntx, uint32_t* tiles, uint32_t* vals, uint32_t cap_pairs, cudaStream_t st) {
    if (n == 0) return;
    view_emit_kernel<<<(n + 255) / 256, 256, 0, st>>>(bbox, cnt, order, offs, n, ntx, tiles, vals, cap_pairs);
}
void gs_launch_view_blend(const float4* rec, const uint32_t* vals, const uint2* ranges, uint32_t W, uint32_t H, uint32_t ntx,
                          uint32_t nty, int quantize, float4* image, unsigned long long* counters, const uint32_t* order,
                          cudaStream_t st) {
    const uint32_t g = ntx * nty;
    if (counters) {
        if (quantize) view_blend_kernel<1, 1><<<g, 256, 0, st>>>(rec, vals, ranges, W, H, ntx, image, counters, order);
        else view_blend_kernel<0, 1><<<g, 256, 0, st>>>(rec, vals, ranges, W, H, ntx, image, counters, order);
    } else {
        if (quantize) view_blend_kernel<1, 0><<<g, 256, 0, st>>>(rec, vals, ranges, W, H, ntx, image, counters, order);
        else view_blend_kernel<0, 0><<<g, 256, 0, st>>>(rec, vals, ranges, W, H, ntx, image, counters, order);
    }
}
void gs_launch_view_splats_to_wire(const float4* rec, const uint4* bbox, uint32_t n, void* out64, cudaStream_t st) {
    if (n == 0) return;
    view_splats_to_wire_kernel<<<(n + 255) / 256, 256, 0, st>>>(rec, bbox, n, (uint32_t*)out64);
}
