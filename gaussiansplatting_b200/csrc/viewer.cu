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

// draw position j = Gaussian order[j]; its tiles go to offs[j] ... in ty-major, tx-minor order.  One warp per draw position:
// viewer quads reach 65 x 65 tiles (radius <= 512), a thread per Gaussian would serialise on them.
__global__ void __launch_bounds__(256) view_emit_kernel(const uint4* __restrict__ bbox, const uint32_t* __restrict__ cnt,
                                                        const uint32_t* __restrict__ order, const uint32_t* __restrict__ offs,
                                                        uint32_t n, uint32_t ntx, uint32_t* __restrict__ tiles,
                                                        uint32_t* __restrict__ vals, uint32_t cap_pairs) {
    const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (j >= n) return;
    const uint32_t i = order[j];
    if (i >= n) return;
    const uint32_t c = cnt[i];
    if (c == 0) return;
    const uint4 bb = bbox[i];
    const uint32_t off = offs[j], wx = bb.z - bb.x + 1u;
    for (uint32_t k = gs::lane_id(); k < c; k += 32u) {
        const uint32_t p = off + k;
        if (p >= cap_pairs) break;
        tiles[p] = (bb.y + k / wx) * ntx + bb.x + k % wx;
        vals[p] = i;
    }
}

__device__ __forceinline__ float quant8(float v) { return gs::fdiv(rintf(gs::fmul(fminf(fmaxf(v, 0.0f), 1.0f), 255.0f)), 255.0f); }

// V4: one CTA per tile, one pixel per thread, entries staged 256 at a time (3 x float4 each) through shared memory
template <int QUANT>
__global__ void __launch_bounds__(256) view_blend_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ vals,
                                                         const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                         float4* __restrict__ image, unsigned long long* __restrict__ counters) {
    __shared__ float4 s_rec[256 * 3];
    const uint32_t tile = blockIdx.x;
    const uint32_t tx = tile % ntx, ty = tile / ntx;
    const uint32_t lx = (threadIdx.x & 7u) + ((threadIdx.x >> 5) & 1u) * 8u;       // a warp owns an 8 x 4 pixel patch
    const uint32_t ly = ((threadIdx.x >> 3) & 3u) + (threadIdx.x >> 6) * 4u;
    const uint32_t px = tx * GS_TILE + lx, py = ty * GS_TILE + ly;
    const bool inside = px < W && py < H;
    const float fx = gs::fadd((float)px, 0.5f), fy = gs::fadd((float)py, 0.5f);
    float cr = 1.0f, cg = 1.0f, cb = 1.0f, ca = 1.0f;                               // clear colour, mtl_engine.mm:583
    uint32_t shaded = 0, blended = 0;
    const uint2 rg = ranges[tile];
    for (uint32_t base = 0; base < rg.y; base += 256u) {
        const uint32_t m = min(256u, rg.y - base);
        __syncthreads();
        if (threadIdx.x < m) {
            const uint32_t id = vals[rg.x + base + threadIdx.x];
            s_rec[threadIdx.x * 3 + 0] = rec[3 * (size_t)id + 0];
            s_rec[threadIdx.x * 3 + 1] = rec[3 * (size_t)id + 1];
            s_rec[threadIdx.x * 3 + 2] = rec[3 * (size_t)id + 2];
        }
        __syncthreads();
        if (!inside) continue;
        for (uint32_t e = 0; e < m; e++) {
            const float4 a0 = s_rec[e * 3 + 0], a1 = s_rec[e * 3 + 1];
            const float dx = gs::fsub(fx, a0.x), dy = gs::fsub(a0.y, fy), r = a1.y;
            if (!(dx >= -r && dx < r && dy > -r && dy <= r)) continue;              // top-left fill rule on the quad
            shaded++;
            const float t1 = gs::fmul(gs::fmul(a0.z, dx), dx), t2 = gs::fmul(gs::fmul(gs::fmul(2.0f, a0.w), dx), dy);
            const float t3 = gs::fmul(gs::fmul(a1.x, dy), dy);
            const float power = gs::fmul(-0.5f, gs::fadd(gs::fadd(t1, t2), t3));    // shaders.metal:296-298
            if (!(power <= 0.0f && power >= -4.5f)) continue;                       // :301
            const float alpha = gs::fmul(a1.z, gs::exp_blend(power));               // :304-305
            if (alpha < GS_ALPHA_MIN) continue;                                     // :308
            blended++;
            const float4 a2 = s_rec[e * 3 + 2];
            const float om = gs::fsub(1.0f, alpha);
            cr = gs::fadd(gs::fmul(a1.w, alpha), gs::fmul(cr, om));                 // ONE, ONE_MINUS_SRC_ALPHA
            cg = gs::fadd(gs::fmul(a2.x, alpha), gs::fmul(cg, om));
            cb = gs::fadd(gs::fmul(a2.y, alpha), gs::fmul(cb, om));
            ca = gs::fadd(alpha, gs::fmul(ca, om));
            if (QUANT) { cr = quant8(cr); cg = quant8(cg); cb = quant8(cb); ca = quant8(ca); }
        }
    }
    if (inside) image[(size_t)py * W + px] = make_float4(cr, cg, cb, ca);
    if (counters) {
        shaded = __reduce_add_sync(0xFFFFFFFFu, shaded);
        blended = __reduce_add_sync(0xFFFFFFFFu, blended);
        if (gs::lane_id() == 0) { atomicAdd(&counters[0], (unsigned long long)shaded); atomicAdd(&counters[1], (unsigned long long)blended); }
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
    const uint64_t threads = (uint64_t)n * 32u;
    view_emit_kernel<<<(uint32_t)((threads + 255) / 256), 256, 0, st>>>(bbox, cnt, order, offs, n, ntx, tiles, vals, cap_pairs);
}
void gs_launch_view_blend(const float4* rec, const uint32_t* vals, const uint2* ranges, uint32_t W, uint32_t H, uint32_t ntx,
                          uint32_t nty, int quantize, float4* image, unsigned long long* counters, cudaStream_t st) {
    if (quantize) view_blend_kernel<1><<<ntx * nty, 256, 0, st>>>(rec, vals, ranges, W, H, ntx, image, counters);
    else view_blend_kernel<0><<<ntx * nty, 256, 0, st>>>(rec, vals, ranges, W, H, ntx, image, counters);
}
void gs_launch_view_splats_to_wire(const float4* rec, const uint4* bbox, uint32_t n, void* out64, cudaStream_t st) {
    if (n == 0) return;
    view_splats_to_wire_kernel<<<(n + 255) / 256, 256, 0, st>>>(rec, bbox, n, (uint32_t*)out64);
}
