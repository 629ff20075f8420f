// preprocess.cu — K1: per-Gaussian EWA projection (replaces projectGaussians,
// tiled_shaders.metal:102-304) plus the per-Gaussian tile count that generateTilePairs
// (tiled_shaders.metal:745-770) would emit, and the OR/AND of the depth keys (sort-pass pruning).
//
// HBM-bound: reads 56 B/Gaussian from SoA planes (coalesced 4-B and 16-B lanes), writes the 48-B blend
// record + bbox/aux planes.  Every rounding is explicit (gs::fmul/fadd/...), in the oracle's order.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

struct Uni {
    float V[16];
    float VP[16];
    float sw, sh, fx, fy;
    uint32_t ntx, nty, n;
};

__device__ __forceinline__ void mulp(const float* m, float x, float y, float z, float out[4]) {
#pragma unroll
    for (int r = 0; r < 4; r++)
        out[r] = gs::fadd(gs::fadd(gs::fadd(gs::fmul(m[r], x), gs::fmul(m[4 + r], y)), gs::fmul(m[8 + r], z)), m[12 + r]);
}
__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return gs::fadd(gs::fadd(gs::fmul(a0, b0), gs::fmul(a1, b1)), gs::fmul(a2, b2));
}

// dhist (optional): the digit histograms of the coming depth sort (4 passes over bits 0..30 of dkey, gs_sort_plan(0, 31)):
// hist[p * 256 + d] += number of keys whose digit p is d.  The grid is persistent (a CTA strides over 256-Gaussian chunks), so a
// CTA keeps its four 256-bin histograms in shared memory over all its chunks and flushes them once: the sort then runs with
// GS_SORT_HINT_PREHIST, without its own clear + histogram kernel (one launch and one pass over the keys less per frame).
__global__ void __launch_bounds__(256) preprocess_kernel(GsParams P, GsProjected O, const __grid_constant__ Uni u,
                                                         uint32_t* __restrict__ meta, uint32_t* __restrict__ dhist) {
    __shared__ uint32_t s_hist[4][256];
    if (dhist) {
#pragma unroll
        for (int p = 0; p < 4; p++) s_hist[p][threadIdx.x] = 0u;
        __syncthreads();
    }
    uint32_t acc_or = 0u, acc_and = 0xFFFFFFFFu;
  for (uint32_t base = blockIdx.x * blockDim.x; base < u.n; base += gridDim.x * blockDim.x) {
    const uint32_t i = base + threadIdx.x;
    const bool live = i < u.n;
    uint32_t count = 0, dkey = 0;
    if (live) {
        float4 r0 = make_float4(0.f, 0.f, -0.f, -0.f), r1 = make_float4(-0.f, 0.f, 0.f, 0.f), r2 = make_float4(0.f, 0.f, 0.f, 0.f);
        uint4 bb = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);       // :114-117
        float4 ax = make_float4(0.f, 0.f, 0.f, 0.f);
        float cc = 0.f;
        const uint32_t cap = P.cap;
        const float px = P.pos[i], py = P.pos[cap + i], pz = P.pos[2 * cap + i];
        const float l0 = P.scale[i], l1s = P.scale[cap + i], l2 = P.scale[2 * cap + i];
        do {
            if (isnan(px) || isnan(py) || isnan(pz) || isnan(l0) || isnan(l1s) || isnan(l2) ||
                fabsf(px) > 1e6f || fabsf(py) > 1e6f || fabsf(pz) > 1e6f) break;   // :120-125
            float vp[4], cp[4];
            mulp(u.V, px, py, pz, vp);                                 // :129
            mulp(u.VP, px, py, pz, cp);                                // :130
            if (cp[3] <= 0.1f || vp[2] <= 0.1f) break;                 // :135
            const float ndcx = gs::fdiv(cp[0], cp[3]), ndcy = gs::fdiv(cp[1], cp[3]);   // :141
            if (fabsf(ndcx) > 1.2f || fabsf(ndcy) > 1.2f) break;       // :144
            const float sx = gs::fmul(gs::fadd(gs::fmul(ndcx, 0.5f), 0.5f), u.sw);   // :150-153
            const float sy = gs::fmul(gs::fadd(gs::fmul(ndcy, 0.5f), 0.5f), u.sh);
            r0.x = sx; r0.y = sy;
            r2.y = vp[2];                                              // depth :156
            ax.x = vp[0]; ax.y = vp[1];                                // :157

            float s0 = gs::exp_pinned(gs::clampf(l0, -GS_MAX_SCALE, GS_MAX_SCALE));   // :160-161
            float s1 = gs::exp_pinned(gs::clampf(l1s, -GS_MAX_SCALE, GS_MAX_SCALE));
            float s2 = gs::exp_pinned(gs::clampf(l2, -GS_MAX_SCALE, GS_MAX_SCALE));
            const float maxS = fmaxf(fmaxf(s0, s1), s2), minS = fminf(fminf(s0, s1), s2);
            if (maxS > gs::fmul(20.0f, minS)) {                        // :166-170
                const float k = gs::fdiv(gs::fmul(20.0f, minS), maxS);
                s0 = gs::fmul(s0, k); s1 = gs::fmul(s1, k); s2 = gs::fmul(s2, k);
            }
            const float4 qq = P.rot[i];
            float qw = qq.x, qx = qq.y, qy = qq.z, qz = qq.w;
            const float qLen = gs::fsqrt(gs::fadd(gs::fadd(gs::fadd(gs::fmul(qw, qw), gs::fmul(qx, qx)), gs::fmul(qy, qy)), gs::fmul(qz, qz)));
            if (qLen > 0.001f) { qw = gs::fdiv(qw, qLen); qx = gs::fdiv(qx, qLen); qy = gs::fdiv(qy, qLen); qz = gs::fdiv(qz, qLen); }
            else { qw = 1.f; qx = 0.f; qy = 0.f; qz = 0.f; }           // :173-175
            // R(row,col), tiled_shaders.metal:91-99
            const float R00 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(qy, qy), gs::fmul(qz, qz))));
            const float R10 = gs::fmul(2.0f, gs::fadd(gs::fmul(qx, qy), gs::fmul(qw, qz)));
            const float R20 = gs::fmul(2.0f, gs::fsub(gs::fmul(qx, qz), gs::fmul(qw, qy)));
            const float R01 = gs::fmul(2.0f, gs::fsub(gs::fmul(qx, qy), gs::fmul(qw, qz)));
            const float R11 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(qx, qx), gs::fmul(qz, qz))));
            const float R21 = gs::fmul(2.0f, gs::fadd(gs::fmul(qy, qz), gs::fmul(qw, qx)));
            const float R02 = gs::fmul(2.0f, gs::fadd(gs::fmul(qx, qz), gs::fmul(qw, qy)));
            const float R12 = gs::fmul(2.0f, gs::fsub(gs::fmul(qy, qz), gs::fmul(qw, qx)));
            const float R22 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(qx, qx), gs::fmul(qy, qy))));
            // M = R*S (:188), Sigma = M*M^T (:190)
            const float M00 = gs::fmul(R00, s0), M01 = gs::fmul(R01, s1), M02 = gs::fmul(R02, s2);
            const float M10 = gs::fmul(R10, s0), M11 = gs::fmul(R11, s1), M12 = gs::fmul(R12, s2);
            const float M20 = gs::fmul(R20, s0), M21 = gs::fmul(R21, s1), M22 = gs::fmul(R22, s2);
            const float S00 = dot3(M00, M00, M01, M01, M02, M02), S01 = dot3(M00, M10, M01, M11, M02, M12), S02 = dot3(M00, M20, M01, M21, M02, M22);
            const float S10 = dot3(M10, M00, M11, M01, M12, M02), S11 = dot3(M10, M10, M11, M11, M12, M12), S12 = dot3(M10, M20, M11, M21, M12, M22);
            const float S20 = dot3(M20, M00, M21, M01, M22, M02), S21 = dot3(M20, M10, M21, M11, M22, M12), S22 = dot3(M20, M20, M21, M21, M22, M22);

            const float z = vp[2];
            const float limx = gs::fdiv(gs::fmul(1.3f, u.fx), z), limy = gs::fdiv(gs::fmul(1.3f, u.fy), z);   // :198-201
            const float txtz = gs::clampf(gs::fdiv(vp[0], z), -limx, limx);
            const float tytz = gs::clampf(gs::fdiv(vp[1], z), -limy, limy);
            const float J00 = gs::fdiv(u.fx, z), J02 = gs::fdiv(gs::fmul(-u.fx, txtz), z);   // :205-208
            const float J11 = gs::fdiv(u.fy, z), J12 = gs::fdiv(gs::fmul(-u.fy, tytz), z);
            float T0[3], T1[3];                                         // T = J*W, W(r,c) = V[c*4+r]  :218-223
#pragma unroll
            for (int c = 0; c < 3; c++) {
                T0[c] = gs::fadd(gs::fmul(J00, u.V[c * 4 + 0]), gs::fmul(J02, u.V[c * 4 + 2]));
                T1[c] = gs::fadd(gs::fmul(J11, u.V[c * 4 + 1]), gs::fmul(J12, u.V[c * 4 + 2]));
            }
            const float U00 = dot3(T0[0], S00, T0[1], S10, T0[2], S20), U01 = dot3(T0[0], S01, T0[1], S11, T0[2], S21), U02 = dot3(T0[0], S02, T0[1], S12, T0[2], S22);
            const float U10 = dot3(T1[0], S00, T1[1], S10, T1[2], S20), U11 = dot3(T1[0], S01, T1[1], S11, T1[2], S21), U12 = dot3(T1[0], S02, T1[1], S12, T1[2], S22);
            float a = dot3(U00, T0[0], U01, T0[1], U02, T0[2]);         // :225-230
            const float b = dot3(U00, T1[0], U01, T1[1], U02, T1[2]);
            float c = dot3(U10, T1[0], U11, T1[1], U12, T1[2]);
            a = gs::fadd(a, 0.3f); c = gs::fadd(c, 0.3f);               // :233-234
            ax.z = a; ax.w = b; cc = c;                                 // :237
            const float det = gs::fsub(gs::fmul(a, c), gs::fmul(b, b)); // :240
            if (det < 0.0001f) break;
            const float inv = gs::fdiv(1.0f, det);                      // :247-248
            const float cA = gs::fmul(c, inv), cB = gs::fmul(-b, inv), cC = gs::fmul(a, inv);
            r0.z = gs::fmul(-0.5f, cA); r0.w = -cB; r1.x = gs::fmul(-0.5f, cC);
            if (gs::fadd(gs::fadd(fabsf(cA), fabsf(cB)), fabsf(cC)) < 0.0001f) r2.w = __uint_as_float(1u);   // :350-351
            const float mid = gs::fmul(0.5f, gs::fadd(a, c));           // :251-255
            const float disc = gs::fsub(gs::fmul(mid, mid), det);
            const float lam = gs::fadd(mid, gs::fsqrt(fmaxf(0.1f, disc)));
            const float radius = fminf(ceilf(gs::fmul(3.0f, gs::fsqrt(lam))), GS_MAX_RADIUS);
            r2.z = radius;
            if (radius <= 0.f) break;
            const int minX = max(0, (int)gs::fsub(sx, radius)), minY = max(0, (int)gs::fsub(sy, radius));   // :264-268
            const int maxX = min((int)u.sw - 1, (int)gs::fadd(sx, radius)), maxY = min((int)u.sh - 1, (int)gs::fadd(sy, radius));
            if (minX > maxX || minY > maxY) { r2.z = 0.f; break; }      // :271-275
            bb.x = (uint32_t)minX / GS_TILE; bb.y = (uint32_t)minY / GS_TILE;   // :278-281
            bb.z = min((uint32_t)maxX / GS_TILE, u.ntx - 1); bb.w = min((uint32_t)maxY / GS_TILE, u.nty - 1);
            const uint32_t tiles = (bb.z - bb.x + 1) * (bb.w - bb.y + 1);
            if (tiles > 256u) { r2.z = 0.f; break; }                    // :284-290
            const float raw = gs::clampf(P.opacity[i], -8.0f, 8.0f);    // :293-294
            const float op = gs::fdiv(1.0f, gs::fadd(1.0f, gs::exp_pinned(-raw)));
            r1.y = op;
            r1.z = gs::clampf(gs::fadd(gs::fmul(GS_SH_C0, P.sh[i]), 0.5f), 0.f, 1.f);             // :297-301
            r1.w = gs::clampf(gs::fadd(gs::fmul(GS_SH_C0, P.sh[4 * cap + i]), 0.5f), 0.f, 1.f);
            r2.x = gs::clampf(gs::fadd(gs::fmul(GS_SH_C0, P.sh[8 * cap + i]), 0.5f), 0.f, 1.f);
            // generateTilePairs' admission tests, tiled_shaders.metal:760-770
            if (bb.x > bb.z || bb.y > bb.w) break;
            if (op < GS_MIN_OPACITY) break;
            if (bb.x > 10000u || bb.z > 10000u || bb.y > 10000u || bb.w > 10000u) break;
            if (tiles > GS_MAX_TILES_PER_GAUSSIAN) break;
            count = tiles;
            uint32_t k = __float_as_uint(vp[2]);                        // :773-774
            dkey = (k & 0x80000000u) ? ~k : (k | 0x80000000u);
        } while (0);
        O.rec[3 * (size_t)i + 0] = r0; O.rec[3 * (size_t)i + 1] = r1; O.rec[3 * (size_t)i + 2] = r2;
        O.bbox[i] = bb; O.aux[i] = ax; O.covc[i] = cc;
        O.cnt[i] = count;
        O.ccnt[i] = count ? ((bb.z >> 3) - (bb.x >> 3) + 1u) * ((bb.w >> 3) - (bb.y >> 3) + 1u) : 0u;
        const uint32_t dk = count ? dkey : 0xFFFFFFFFu;
        O.dkey[i] = dk;
        if (dhist) {
            atomicAdd(&s_hist[0][dk & 255u], 1u); atomicAdd(&s_hist[1][(dk >> 8) & 255u], 1u);
            atomicAdd(&s_hist[2][(dk >> 16) & 255u], 1u);
        }
    }
    if (dhist) {
        // top digit (sign + exponent bits): a handful of distinct values per frame, so 256 same-address shared atomics per chunk
        // would serialise; one atomic per distinct value and warp instead
        const uint32_t d3 = live ? (((count ? dkey : 0xFFFFFFFFu) >> 24) & 127u) : 128u;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d3);
        if (d3 < 128u && (uint32_t)(__ffs(peers) - 1) == gs::lane_id()) atomicAdd(&s_hist[3][d3], (uint32_t)__popc(peers));
    }
    if (count) { acc_or |= dkey; acc_and &= dkey; }
  }
    if (dhist) {
        __syncthreads();
#pragma unroll
        for (int p = 0; p < 4; p++) { const uint32_t h = s_hist[p][threadIdx.x]; if (h) atomicAdd(&dhist[p * 256 + threadIdx.x], h); }
    }
    // depth-key OR / AND over Gaussians that emit pairs: digits whose bits never vary need no sort pass (only the literal
    // 64-bit pipeline reads them; two same-address atomics per warp are not free, so the other pipelines skip them)
    if (meta == nullptr) return;
    __shared__ uint32_t s_or[8], s_and[8];
    const uint32_t kor = __reduce_or_sync(0xFFFFFFFFu, acc_or);
    const uint32_t kand = __reduce_and_sync(0xFFFFFFFFu, acc_and);
    if (gs::lane_id() == 0) { s_or[threadIdx.x >> 5] = kor; s_and[threadIdx.x >> 5] = kand; }
    __syncthreads();
    if (threadIdx.x == 0) {                                  // one pair of atomics per CTA (same-address atomics serialise)
        uint32_t o = 0u, a = 0xFFFFFFFFu;
#pragma unroll
        for (int w = 0; w < 8; w++) { o |= s_or[w]; a &= s_and[w]; }
        if (o) atomicOr(&meta[GS_META_KEY_OR], o);
        if (a != 0xFFFFFFFFu) atomicAnd(&meta[GS_META_KEY_AND], a);
    }
}

}  // namespace

void gs_launch_preprocess(const GsParams& P, const GsProjected& O, const float* V, const float* VP, float sw, float sh,
                          float fx, float fy, uint32_t ntx, uint32_t nty, uint32_t n, uint32_t* meta, uint32_t* depth_hist, cudaStream_t st) {
    if (n == 0) return;
    Uni u;
    for (int k = 0; k < 16; k++) { u.V[k] = V[k]; u.VP[k] = VP[k]; }
    u.sw = sw; u.sh = sh; u.fx = fx; u.fy = fy; u.ntx = ntx; u.nty = nty; u.n = n;
    // persistent grid when the histograms ride along (one flush of 4 x 256 bins per CTA); one chunk per CTA otherwise.  The grid is
    // ONE balanced wave: with R resident CTAs on the device every CTA takes k = ceil(chunks / R) chunks and ceil(chunks / k) CTAs
    // are launched, so no CTA waits for a second wave and none does a chunk more than its neighbours.
    const uint32_t chunks = (n + 255u) / 256u;
    uint32_t grid = chunks;
    if (depth_hist) {
        int per_sm = 0, dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, preprocess_kernel, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 2;
        const uint32_t R = (uint32_t)per_sm * (uint32_t)sms;
        const uint32_t k = (chunks + R - 1u) / R;
        grid = k ? (chunks + k - 1u) / k : 1u;
    }
    preprocess_kernel<<<grid, 256, 0, st>>>(P, O, u, meta, depth_hist);
}
