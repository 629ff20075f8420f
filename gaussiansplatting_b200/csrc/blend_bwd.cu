// blend_bwd.cu — K7: reverse-order per-tile backward blend (tiledBackward's per-pixel part,
// tiled_shaders.metal:388-536, 567-573) and K8: the per-Gaussian linear tail (:538-736).
//
// The reference does the whole chain per (pixel, Gaussian) and issues 16 float atomics each.  Everything
// after dL/dalpha is LINEAR in nine per-(pixel,Gaussian) quantities with per-Gaussian coefficients, so K7 only
// accumulates those nine moments
//     m0..2 = sum dLdp_c * w      m3 = sum q      m4,5 = sum q*dx, q*dy      m6,7,8 = sum q*dx^2, q*dx*dy, q*dy^2
// (w = alpha*T, q = dL/dalpha * G), reduced across the warp by a transposed butterfly (14 shuffles for 9
// values); K8 then applies the chain once per Gaussian.  Three walks share that arithmetic:
//   blend_backward_warp_kernel     the default from a rasterizer's second backward on: every warp walks its own extent of the
//                                  list on the visit words the forward recorded (its decisions per pixel slot), the moments of
//                                  a visit go to the per-Gaussian rows as one global float reduction instruction
//   blend_backward_kernel<2,true>  the same words, CTA-cooperative rounds (gs_tune "bwd_warp" = 0)
//   blend_backward_kernel<.,false> the pinned walk: re-derives the decisions (power, G, alpha tests in the pinned arithmetic, so
//                                  it walks exactly the entries the oracle walks), sums the warps' moments in shared memory and
//                                  flushes one atomic per (tile, Gaussian, moment); a rasterizer's first backward, and through
//                                  blend_backward_flagged_kernel the tiles the forward flagged (conic-degenerate entries)
#include "gs_common.cuh"
#include "gs_kernels.h"

#ifdef GS_BWD_MINB         // optional: minimum resident CTAs per SM asked of ptxas (register budget knob; measured: no gain)
#define GS_BWD_BOUNDS(nt, masked) __launch_bounds__(nt, GS_BWD_MINB)
#else
#ifndef GS_BWD_MASKED_MINB
#define GS_BWD_MASKED_MINB 0   // 0 = unconstrained (66 registers, 7 CTAs per SM: 412 us at C2).  Measured: 8 -> 64 registers, 423 us;
                               // 8 with rounds of 120 entries so that eight CTAs' shared memory fits as well: 420 us
#endif
#define GS_BWD_BOUNDS(nt, masked) __launch_bounds__(nt, (masked) ? GS_BWD_MASKED_MINB : 0)
#endif

namespace {

#ifndef GS_BWD_PPT
#define GS_BWD_PPT 2
#endif
#ifndef GS_BWD_BB
#define GS_BWD_BB 128
#endif
constexpr int BB = GS_BWD_BB;   // raw list entries staged per round
#ifndef GS_BWD_BB_MASKED
#define GS_BWD_BB_MASKED 128
#endif
constexpr int BBM = GS_BWD_BB_MASKED;   // the same for the masked variant, whose rounds also hold 32 B of visit words per entry

// sum of v[0..7] over the warp; afterwards lane L holds (in v[0]) the total of value ((L>>2)&7 bit-reversed order below)
__device__ __forceinline__ float warp_reduce8_transposed(float v[8], uint32_t lane) {
    const bool b4 = lane & 16u, b3 = lane & 8u, b2 = lane & 4u;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float send = b4 ? v[i] : v[i + 4], keep = b4 ? v[i + 4] : v[i];
        v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float send = b3 ? v[i] : v[i + 2], keep = b3 ? v[i + 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 8);
    }
    {
        const float send = b2 ? v[0] : v[1], keep = b2 ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 4);
    }
    v[0] += __shfl_xor_sync(0xFFFFFFFFu, v[0], 2);
    v[0] += __shfl_xor_sync(0xFFFFFFFFu, v[0], 1);
    return v[0];
}
// which of the 8 values a lane ends up holding
__device__ __forceinline__ uint32_t reduce8_index(uint32_t lane) {
    return ((lane >> 4) & 1u) * 4u + ((lane >> 3) & 1u) * 2u + ((lane >> 2) & 1u);
}

// PPT (2 or 4) pixels per thread, NT = 256/PPT threads (NW warps) per tile.  Each round stages BB raw list entries (those
// flagged GS_CULL_BIT by pairs.cu are dropped, order kept) and marks per warp the survivors that can reach the warp's
// own pixel rectangle.  A warp reduces the nine moments of an entry over its lanes and stores them in ITS OWN row of
// shared memory (a warp meets an entry once per round, so no atomics and no conflicts); the rows are summed and
// flushed with one global float atomic per (tile, Gaussian, moment).
// The walk is branch-free and keeps the thread's pixels two to a packed fp32 pair (FFMA2, gs_common.cuh): a pixel that
// does not take an entry runs the same instructions with alpha and G forced to 0, which leaves its T and accumulated
// colour unchanged (T * rcp(1), 0*c + 1*acc) and adds exact zeros to the moments.
struct BwdEnt { float4 r0, r1; float b; uint32_t idx, gid, pad; };   // 48 B: one base address, three LDS.128
// MASKED (PPT 2): the forward stored, for every (entry, warp) visit, which of the warp's 64 pixel slots took the entry
// (blend_fwd.cu, `vis`).  The backward's decisions (:489, :495, the extent of the walk) are the forward's, so this variant takes
// them from that word instead of re-deriving them: no pinned exp on the path (G by ex2.approx: gradients are graded at 1e-3),
// no range / threshold compares, no vote, and a visit in which no slot took the entry is skipped before its record is loaded.
// Tiles in which the forward skipped a conic-degenerate entry that the reference's backward does not skip (note I, tile_flags)
// are left to the pinned variant (blend_backward_flagged_kernel, launched behind this one).
__device__ __forceinline__ float ex2_approx(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// one tile's walk by one CTA of 256 / PPT threads (the body of both kernels below); redo = the tile is flagged
template <int PPT, bool MASKED>
__device__ __forceinline__ void blend_backward_tile(const uint32_t tile, const bool redo, const float4* __restrict__ rec,
                                                    const uint32_t* __restrict__ svals, const uint2* __restrict__ ranges, uint32_t W,
                                                    uint32_t H, uint32_t ntx, const float4* __restrict__ rendered,
                                                    const float4* __restrict__ gt, int quantize, const uint32_t* __restrict__ last_idx,
                                                    const float* __restrict__ tfinal, float* __restrict__ acc,
                                                    const uint2* __restrict__ vis, size_t vstride) {
    constexpr int RB = MASKED ? BBM : BB;
    constexpr int NT = 256 / PPT, NW = NT / 32, K = (RB + NT - 1) / NT, NP = PPT / 2;
    static_assert(!MASKED || PPT == 2, "the forward records its visits for two pixels per thread only");
    __shared__ BwdEnt s_e[RB];
    __shared__ uint2 s_vis[MASKED ? RB : 1][NW];
    __shared__ float s_acc[NW][RB * 9];
    __shared__ uint32_t s_w[NW];
    __shared__ uint32_t s_wmax[NW];
    __shared__ unsigned char s_reach[RB];              // per staged entry: which warp regions it can reach at all (bit w)
    const uint32_t tX = tile % ntx, tY = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint2 range = ranges[tile];                                // :410-413
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(s_e);

    gs::f32x2 X2[NP], Y2[NP], T2[NP], dr2[NP], dg2[NP], db2[NP], ar2[NP], ag2[NP], ab2[NP];
    uint32_t e[PPT];                                                 // entries [0, e) of the tile list concern pixel j
    uint32_t emax = 0;
#pragma unroll
    for (int p = 0; p < NP; p++) {
        float pxf[2], pyf[2], T[2], dr[2], dg[2], db[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int j = 2 * p + h;
            uint32_t x, y;
            gs::tile_pixel<PPT>(warp, lane, j, x, y);
            const uint32_t px = tX * 16u + x, py = tY * 16u + y;
            pxf[h] = (float)px + 0.5f; pyf[h] = (float)py + 0.5f;        // :416
            e[j] = 0; T[h] = 1.0f; dr[h] = 0.f; dg[h] = 0.f; db[h] = 0.f;
            if (px < W && py < H) {
                const size_t pix = (size_t)py * W + px;
                const uint32_t li = last_idx[pix];                       // :404-407
                if (li != 0xFFFFFFFFu) {
                    e[j] = min(li + 1u, range.x + range.y) - range.x;    // :427
                    const float4 r = rendered[pix], t = gt[pix];         // :418-423
                    float rr[3] = {r.x, r.y, r.z};
                    const float tt[3] = {t.x, t.y, t.z};
                    float s[3];
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        if (quantize) rr[c] = floorf(gs::clampf(rr[c], 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
                        const float diff = rr[c] - tt[c];
                        s[c] = (diff > 0.f) ? (1.0f / 3.0f) : ((diff < 0.f) ? (-1.0f / 3.0f) : 0.f);
                    }
                    dr[h] = s[0]; dg[h] = s[1]; db[h] = s[2];
                    T[h] = tfinal[pix];
                    if (redo) {
                        // Rare path (a conic-degenerate entry was walked): redo the reference's first pass (:430-460), which,
                        // unlike the forward, does not skip such entries.
                        float Tf = 1.0f;
                        for (uint32_t k = 0; k < e[j]; k++) {
                            const uint32_t g = svals[range.x + k] & GS_ID_MASK;
                            const float4 a = rec[3 * (size_t)g], b = rec[3 * (size_t)g + 1];
                            const float dx = gs::fsub(pxf[h], a.x), dy = gs::fsub(pyf[h], a.y);
                            const float power = gs::blend_power(a.z, a.w, b.x, dx, dy);
                            if (!(power <= 0.0f) || power < -4.5f) continue;
                            const float alpha = fminf(gs::fmul(b.y, gs::exp_blend(power)), GS_ALPHA_MAX);
                            if (alpha < GS_ALPHA_MIN) continue;
                            const float test_T = gs::fmul(Tf, gs::fsub(1.0f, alpha));
                            if (test_T < GS_T_EPS) break;                // :458
                            Tf = test_T;
                        }
                        T[h] = Tf;
                    }
                }
            }
            emax = max(emax, e[j]);
        }
        X2[p] = gs::pk2(pxf[0], pxf[1]); Y2[p] = gs::pk2(pyf[0], pyf[1]); T2[p] = gs::pk2(T[0], T[1]);
        dr2[p] = gs::pk2(dr[0], dr[1]); dg2[p] = gs::pk2(dg[0], dg[1]); db2[p] = gs::pk2(db[0], db[1]);
        ar2[p] = gs::bc2(1.0f); ag2[p] = gs::bc2(1.0f); ab2[p] = gs::bc2(1.0f);   // accum_rec starts at the white background :468-470
    }
    // CTA-wide and warp-wide extent of the walk
    const uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, emax);
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    uint32_t bmax = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) bmax = max(bmax, s_wmax[w]);

    float* my_acc = s_acc[warp];
    const gs::f32x2 one2 = gs::bc2(1.0f);
    for (uint32_t hi = bmax; hi > 0;) {
        const uint32_t lo = hi > (uint32_t)RB ? hi - RB : 0u;
        const uint32_t cntraw = hi - lo;
        // ---- stage (compacting out culled pairs, list order kept) and clear this warp's moment rows
        uint32_t v[K], nl = 0;
#pragma unroll
        for (int q = 0; q < K; q++) {
            const uint32_t j = threadIdx.x * K + q;
            v[q] = (j < cntraw) ? svals[range.x + lo + j] : GS_CULL_BIT;
            nl += (v[q] >> 31) ^ 1u;
        }
        uint32_t total;
        uint32_t slot = gs::cta_excl_scan<NW>(nl, s_w, &total);
#pragma unroll
        for (int q = 0; q < K; q++) {
            if (v[q] >> 31) continue;
            const uint32_t g = v[q];
            const float4 r0 = rec[3 * (size_t)g], r1 = rec[3 * (size_t)g + 1];
            s_e[slot].r0 = r0; s_e[slot].r1 = r1;
            s_e[slot].b = rec[3 * (size_t)g + 2].x; s_e[slot].idx = lo + threadIdx.x * K + q; s_e[slot].gid = g;
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
            if (MASKED) {                                            // the four warps' words of this entry: 32 contiguous bytes
#pragma unroll
                for (int w = 0; w < NW; w++) s_vis[slot][w] = vis[(size_t)w * vstride + range.x + lo + threadIdx.x * K + q];
            }
            slot++;
        }
        for (uint32_t i = lane; i < total * 9u; i += 32u) my_acc[i] = 0.f;
        __syncthreads();
        if (MASKED || wmax > lo) {                                   // MASKED: a zero visit word says "not this warp's entry"
            // staged list indices ascend: entries at or beyond this warp's extent are of no interest to it
            uint32_t kend = total;
            if (!MASKED && wmax < hi) {
                uint32_t a = 0, b = total;
                while (a < b) { const uint32_t m = (a + b) >> 1; if (s_e[m].idx < wmax) a = m + 1; else b = m; }
                kend = a;
            }
            for (int wd = ((int)kend - 1) >> 5; wd >= 0; wd--) {      // :473 back to front, reachable entries only
                const uint32_t sl = (uint32_t)wd * 32u + lane;
                bool want = sl < kend && ((s_reach[sl] >> warp) & 1u);
                if (MASKED) { if (want) { const uint2 vm = s_vis[sl][warp]; want = (vm.x | vm.y) != 0u; } }
                uint32_t mk = __ballot_sync(0xFFFFFFFFu, want);
                const uint32_t wbase = ebase + (uint32_t)wd * 32u * (uint32_t)sizeof(BwdEnt);
                while (mk) {
                    int kb;
                    if (MASKED) {                                    // bfind = the position of the highest set bit in one instruction
                        asm("bfind.u32 %0, %1;" : "=r"(kb) : "r"(mk));
                        mk ^= 1u << kb;
                    } else {
                        kb = 31 - __clz(mk);
                        mk &= ~(1u << kb);
                    }
                    const uint32_t addr = wbase + (uint32_t)kb * (uint32_t)sizeof(BwdEnt);
                    const float4 a = gs::lds128(addr);
                    const float4 b = gs::lds128(addr + 16u);
                    const float2 c = gs::lds64(addr + 32u);
                    const uint32_t idx = __float_as_uint(c.y);
                    const gs::f32x2 sx = gs::bc2(a.x), sy = gs::bc2(a.y), Ah = gs::bc2(a.z), Bn = gs::bc2(a.w), Ch = gs::bc2(b.x),
                                    op = gs::bc2(b.y), cr = gs::bc2(b.z), cg = gs::bc2(b.w), cb = gs::bc2(c.x);
                    gs::f32x2 m2[9];
                    bool anyc = false;
#pragma unroll
                    for (int p = 0; p < NP; p++) {
                        const gs::f32x2 dx = gs::fsub2(X2[p], sx), dy = gs::fsub2(Y2[p], sy);   // :483
                        const gs::f32x2 pw2 = gs::blend_power2(Ah, Bn, Ch, dx, dy);
                        float pw[2], al[2], G[2];
                        if (MASKED) {
                            const uint2 vm = s_vis[wd * 32 + kb][warp];
                            float t0, t1;
                            gs::up2(gs::fmul2(pw2, gs::bc2(GS_LOG2E)), t0, t1);
                            G[0] = ex2_approx(t0); G[1] = ex2_approx(t1);                      // :492
                            const bool c0 = (vm.x >> lane) & 1u, c1 = (vm.y >> lane) & 1u;     // the forward's decision for this slot
                            al[0] = c0 ? fminf(b.y * G[0], GS_ALPHA_MAX) : 0.0f; al[1] = c1 ? fminf(b.y * G[1], GS_ALPHA_MAX) : 0.0f;
                            G[0] = c0 ? G[0] : 0.0f; G[1] = c1 ? G[1] : 0.0f;
                            anyc = true;
                        } else {
                        const gs::f32x2 G2 = gs::exp_blend2(pw2);                             // :492
                        const gs::f32x2 a2 = gs::fmul2(op, G2);
                        gs::up2(pw2, pw[0], pw[1]); gs::up2(a2, al[0], al[1]); gs::up2(G2, G[0], G[1]);
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            al[h] = fminf(al[h], GS_ALPHA_MAX);
                            const bool ch = (idx < e[2 * p + h]) && (pw[h] <= 0.0f) && !(pw[h] < -4.5f) && !(al[h] < GS_ALPHA_MIN);   // :489, :495
                            al[h] = ch ? al[h] : 0.0f;
                            G[h] = ch ? G[h] : 0.0f;
                            anyc |= ch;
                        }
                        }
                        const gs::f32x2 alpha = gs::pk2(al[0], al[1]), Gq = gs::pk2(G[0], G[1]);
                        const gs::f32x2 om = gs::fsub2(one2, alpha);
                        float om0, om1;
                        gs::up2(om, om0, om1);
                        const gs::f32x2 Tn = gs::fmul2(T2[p], gs::pk2(gs::rcp_approx(fmaxf(om0, 0.0001f)), gs::rcp_approx(fmaxf(om1, 0.0001f))));   // :498
                        T2[p] = Tn;
                        const gs::f32x2 w2 = gs::fmul2(alpha, Tn);                             // :500
                        gs::f32x2 s2 = gs::fmul2(dr2[p], gs::fsub2(cr, ar2[p]));
                        s2 = gs::ffma2(dg2[p], gs::fsub2(cg, ag2[p]), s2);
                        s2 = gs::ffma2(db2[p], gs::fsub2(cb, ab2[p]), s2);
                        const gs::f32x2 dLdA = gs::fmul2(Tn, s2);                              // :510
                        ar2[p] = gs::ffma2(alpha, cr, gs::fmul2(om, ar2[p]));                  // :514
                        ag2[p] = gs::ffma2(alpha, cg, gs::fmul2(om, ag2[p]));
                        ab2[p] = gs::ffma2(alpha, cb, gs::fmul2(om, ab2[p]));
                        const gs::f32x2 q = gs::fmul2(dLdA, Gq);
                        const gs::f32x2 qx = gs::fmul2(q, dx), qy = gs::fmul2(q, dy);
                        if (p == 0) {
                            m2[0] = gs::fmul2(dr2[p], w2); m2[1] = gs::fmul2(dg2[p], w2); m2[2] = gs::fmul2(db2[p], w2); m2[3] = q;
                            m2[4] = qx; m2[5] = qy; m2[6] = gs::fmul2(qx, dx); m2[7] = gs::fmul2(qx, dy); m2[8] = gs::fmul2(qy, dy);
                        } else {
                            m2[0] = gs::ffma2(dr2[p], w2, m2[0]); m2[1] = gs::ffma2(dg2[p], w2, m2[1]); m2[2] = gs::ffma2(db2[p], w2, m2[2]);
                            m2[3] = gs::fadd2(m2[3], q); m2[4] = gs::fadd2(m2[4], qx); m2[5] = gs::fadd2(m2[5], qy);
                            m2[6] = gs::ffma2(qx, dx, m2[6]); m2[7] = gs::ffma2(qx, dy, m2[7]); m2[8] = gs::ffma2(qy, dy, m2[8]);
                        }
                    }
                    if (!MASKED) { if (!__any_sync(0xFFFFFFFFu, anyc)) continue; }
                    float v8[8], v9;
#pragma unroll
                    for (int i = 0; i < 9; i++) {
                        float l, h;
                        gs::up2(m2[i], l, h);
                        if (i < 8) v8[i] = l + h; else v9 = l + h;
                    }
                    const float tot = warp_reduce8_transposed(v8, lane);
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) v9 += __shfl_xor_sync(0xFFFFFFFFu, v9, d);
                    const int k = wd * 32 + kb;
                    if (MASKED) {
                        // the four lanes of a group hold the same sum, all lanes the same v9: unpredicated stores to the same
                        // words (one transaction) leave the predicate registers to the lane bits of the butterfly
                        my_acc[k * 9 + reduce8_index(lane)] = tot;
                        my_acc[k * 9 + 8] = v9;
                    } else {
                        if ((lane & 3u) == 0u) my_acc[k * 9 + reduce8_index(lane)] = tot;
                        if (lane == 1u) my_acc[k * 9 + 8] = v9;
                    }
                }
            }
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < total * 9u; i += NT) {
            float x = 0.f;
#pragma unroll
            for (int w = 0; w < NW; w++) x += s_acc[w][i];
            if (x != 0.f) atomicAdd(&acc[(size_t)s_e[i / 9u].gid * 12u + (i % 9u)], x);
        }
        __syncthreads();
        hi = lo;
    }
}

template <int PPT, bool MASKED>
__global__ void GS_BWD_BOUNDS(256 / PPT, MASKED) blend_backward_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ svals,
                                                             const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                             const float4* __restrict__ rendered, const float4* __restrict__ gt,
                                                             int quantize, const uint32_t* __restrict__ last_idx,
                                                             const float* __restrict__ tfinal, const uint32_t* __restrict__ tile_flags,
                                                             const uint32_t* __restrict__ order, float* __restrict__ acc,
                                                             const uint2* __restrict__ vis, size_t vstride) {
    const uint32_t tile = order ? order[blockIdx.x] : blockIdx.x;
    const bool redo = tile_flags[tile] != 0u;
    if (MASKED && redo) return;                                      // CTA-uniform: left to blend_backward_flagged_kernel
    blend_backward_tile<PPT, MASKED>(tile, redo, rec, svals, ranges, W, H, ntx, rendered, gt, quantize, last_idx, tfinal, acc, vis, vstride);
}

// The flagged tiles behind a masked launch, by the pinned walk.  They are rare (none in most frames), so the grid is not one
// CTA per tile: a CTA reads the flags of 256 / PPT consecutive tiles with one coalesced load, leaves at once when none is set
// (3 us for the whole launch instead of the 10 us that 8160 CTAs need to start and return), and otherwise walks its flagged
// tiles one after another.
template <int PPT>
__global__ void __launch_bounds__(256 / PPT) blend_backward_flagged_kernel(const float4* __restrict__ rec, const uint32_t* __restrict__ svals,
                                                             const uint2* __restrict__ ranges, uint32_t W, uint32_t H, uint32_t ntx,
                                                             uint32_t ntiles, const float4* __restrict__ rendered,
                                                             const float4* __restrict__ gt, int quantize,
                                                             const uint32_t* __restrict__ last_idx, const float* __restrict__ tfinal,
                                                             const uint32_t* __restrict__ tile_flags, float* __restrict__ acc) {
    constexpr int NT = 256 / PPT;
    __shared__ unsigned char s_flag[NT];
    const uint32_t t = blockIdx.x * NT + threadIdx.x;
    const bool f = t < ntiles && tile_flags[t] != 0u;
    s_flag[threadIdx.x] = f ? 1 : 0;
    if (!__syncthreads_or(f)) return;
    for (uint32_t i = 0; i < (uint32_t)NT; i++) {
        if (!s_flag[i]) continue;                                    // CTA-uniform
        blend_backward_tile<PPT, false>(blockIdx.x * NT + i, true, rec, svals, ranges, W, H, ntx, rendered, gt, quantize, last_idx, tfinal,
                                        acc, nullptr, 0);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// K7, warp-autonomous walk on the forward's visit words (gs_tune "bwd_warp" = 1).
// In the CTA-cooperative kernels above the four warps of a tile meet at three barriers per round of 128 entries, and the
// warp with the most visits of a round keeps the other three waiting (ncu: the barrier is the largest stall of the kernel,
// 22 % of the warps' time).  With the visit words a warp does not need the CTA: word (entry, warp) says whether the entry
// concerns the warp at all.  Each warp walks its own extent of the tile's list back to front in chunks of 32 entries: lane i
// reads entry i's pair value and the warp's word of it (the next chunk's are in flight while this one is walked), the lanes
// whose word is not zero gather their records into the warp's 1.5 KB of shared memory, and the warp visits them highest
// first.  The nine moments of a visit go to the per-Gaussian rows with global float reductions straight from the registers
// (no shared accumulator, no flush pass): about as many reductions as the cooperative kernel issues, since an entry is
// visited by two warps on average and the cooperative kernel flushes every staged entry's nine moments per tile.  No
// __syncthreads anywhere: a warp whose pixels took nothing leaves at once.
struct WEnt { float4 r0, r1; float b; uint32_t gid, vx, vy; };   // 48 B: three LDS.128 per visit (record, id, visit word)
#ifndef GS_BWD_WARP_MINB
#define GS_BWD_WARP_MINB 8   // 60 registers; measured flat from 7 to 10 CTAs per SM (385-395 us at C2): the kernel is issue-bound
#endif
__global__ void __launch_bounds__(128, GS_BWD_WARP_MINB) blend_backward_warp_kernel(
    const float4* __restrict__ rec, const uint32_t* __restrict__ svals, const uint2* __restrict__ ranges, uint32_t W, uint32_t H,
    uint32_t ntx, const float4* __restrict__ rendered, const float4* __restrict__ gt, int quantize,
    const uint32_t* __restrict__ last_idx, const float* __restrict__ tfinal, const uint32_t* __restrict__ tile_flags,
    const uint32_t* __restrict__ order, float* __restrict__ acc, const uint2* __restrict__ vis, size_t vstride) {
    __shared__ __align__(16) WEnt s_e[4][32];
    const uint32_t tile = order ? order[blockIdx.x] : blockIdx.x;
    if (tile_flags[tile] != 0u) return;                              // left to blend_backward_flagged_kernel
    const uint32_t tX = tile % ntx, tY = tile / ntx;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint2 range = ranges[tile];                                // :410-413
    gs::f32x2 X2, Y2, T2, dr2, dg2, db2, ar2, ag2, ab2;
    uint32_t emax = 0;
    {
        float pxf[2], pyf[2], T[2], dr[2], dg[2], db[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            uint32_t x, y;
            gs::tile_pixel<2>(warp, lane, h, x, y);
            const uint32_t px = tX * 16u + x, py = tY * 16u + y;
            pxf[h] = (float)px + 0.5f; pyf[h] = (float)py + 0.5f;        // :416
            T[h] = 1.0f; dr[h] = 0.f; dg[h] = 0.f; db[h] = 0.f;
            if (px < W && py < H) {
                const size_t pix = (size_t)py * W + px;
                const uint32_t li = last_idx[pix];                       // :404-407
                if (li != 0xFFFFFFFFu) {
                    emax = max(emax, min(li + 1u, range.x + range.y) - range.x);    // :427
                    const float4 r = rendered[pix], t = gt[pix];         // :418-423
                    float rr[3] = {r.x, r.y, r.z};
                    const float tt[3] = {t.x, t.y, t.z};
                    float sg[3];
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        if (quantize) rr[c] = floorf(gs::clampf(rr[c], 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
                        const float diff = rr[c] - tt[c];
                        sg[c] = (diff > 0.f) ? (1.0f / 3.0f) : ((diff < 0.f) ? (-1.0f / 3.0f) : 0.f);
                    }
                    dr[h] = sg[0]; dg[h] = sg[1]; db[h] = sg[2];
                    T[h] = tfinal[pix];
                }
            }
        }
        X2 = gs::pk2(pxf[0], pxf[1]); Y2 = gs::pk2(pyf[0], pyf[1]); T2 = gs::pk2(T[0], T[1]);
        dr2 = gs::pk2(dr[0], dr[1]); dg2 = gs::pk2(dg[0], dg[1]); db2 = gs::pk2(db[0], db[1]);
        ar2 = gs::bc2(1.0f); ag2 = gs::bc2(1.0f); ab2 = gs::bc2(1.0f);   // accum_rec starts at the white background :468-470
    }
    const uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, emax);      // this warp's extent of the list
    if (wmax == 0u) return;
    const uint32_t wu = __shfl_sync(0xFFFFFFFFu, warp, 0);           // the warp index as a value the compiler knows to be uniform
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(s_e) + wu * (32u * (uint32_t)sizeof(WEnt));
    WEnt* const me = &s_e[wu][lane];
    const uint32_t* const sv = svals + range.x;
    const uint2* const vw = vis + (size_t)wu * vstride + range.x;     // this warp's plane: its word of list entry i is vw[i]
    const uint32_t lanebit = 1u << lane, ridx = lane == 1u ? 8u : reduce8_index(lane);
    const bool writer = (lane & 3u) == 0u || lane == 1u;
    const gs::f32x2 one2 = gs::bc2(1.0f);
    // chunks of 32 list entries, back to front, cut at multiples of 32 so that the 128-byte loads of the pair values are aligned
    uint32_t lo = (wmax - 1u) & ~31u;
    uint32_t v_n = GS_CULL_BIT;
    uint2 w_n = make_uint2(0u, 0u);
    if (lo + lane < wmax) { v_n = sv[lo + lane]; w_n = vw[lo + lane]; }
    for (;;) {
        const uint32_t v = v_n;
        const uint2 w = w_n;
        const bool more = lo != 0u;
        if (more) { lo -= 32u; v_n = sv[lo + lane]; w_n = vw[lo + lane]; }   // in flight while this chunk is walked
        // a culled pair was never staged by the forward and has no word (GS_CULL_BIT, pairs.cu); a zero word = not this warp's entry
        const bool has = !(v >> 31) && (w.x | w.y) != 0u;
        uint32_t mk = __ballot_sync(0xFFFFFFFFu, has);
        if (mk) {
            if (has) {
                const uint32_t g = v & GS_ID_MASK;
                me->r0 = rec[3 * (size_t)g]; me->r1 = rec[3 * (size_t)g + 1];
                me->b = rec[3 * (size_t)g + 2].x; me->gid = g; me->vx = w.x; me->vy = w.y;
            }
            __syncwarp();
            while (mk) {                                             // :473 back to front
                int kb;
                asm("bfind.u32 %0, %1;" : "=r"(kb) : "r"(mk));
                mk ^= 1u << kb;
                const uint32_t addr = ebase + (uint32_t)kb * (uint32_t)sizeof(WEnt);
                const float4 a = gs::lds128(addr);
                const float4 b = gs::lds128(addr + 16u);
                const float4 c = gs::lds128(addr + 32u);
                const gs::f32x2 sx = gs::bc2(a.x), sy = gs::bc2(a.y), Ah = gs::bc2(a.z), Bn = gs::bc2(a.w), Ch = gs::bc2(b.x),
                                cr = gs::bc2(b.z), cg = gs::bc2(b.w), cb = gs::bc2(c.x);
                const gs::f32x2 dx = gs::fsub2(X2, sx), dy = gs::fsub2(Y2, sy);   // :483
                const gs::f32x2 pw2 = gs::blend_power2(Ah, Bn, Ch, dx, dy);
                float t0, t1, al[2], G[2];
                gs::up2(gs::fmul2(pw2, gs::bc2(GS_LOG2E)), t0, t1);
                G[0] = ex2_approx(t0); G[1] = ex2_approx(t1);                      // :492
                const bool c0 = (__float_as_uint(c.z) & lanebit) != 0u, c1 = (__float_as_uint(c.w) & lanebit) != 0u;   // the forward's decisions
                al[0] = c0 ? fminf(b.y * G[0], GS_ALPHA_MAX) : 0.0f; al[1] = c1 ? fminf(b.y * G[1], GS_ALPHA_MAX) : 0.0f;
                G[0] = c0 ? G[0] : 0.0f; G[1] = c1 ? G[1] : 0.0f;
                const gs::f32x2 alpha = gs::pk2(al[0], al[1]), Gq = gs::pk2(G[0], G[1]);
                const gs::f32x2 om = gs::fsub2(one2, alpha);
                float om0, om1;
                gs::up2(om, om0, om1);
                const gs::f32x2 Tn = gs::fmul2(T2, gs::pk2(gs::rcp_approx(fmaxf(om0, 0.0001f)), gs::rcp_approx(fmaxf(om1, 0.0001f))));   // :498
                T2 = Tn;
                const gs::f32x2 w2 = gs::fmul2(alpha, Tn);                             // :500
                gs::f32x2 s2 = gs::fmul2(dr2, gs::fsub2(cr, ar2));
                s2 = gs::ffma2(dg2, gs::fsub2(cg, ag2), s2);
                s2 = gs::ffma2(db2, gs::fsub2(cb, ab2), s2);
                const gs::f32x2 dLdA = gs::fmul2(Tn, s2);                              // :510
                ar2 = gs::ffma2(alpha, cr, gs::fmul2(om, ar2));                        // :514
                ag2 = gs::ffma2(alpha, cg, gs::fmul2(om, ag2));
                ab2 = gs::ffma2(alpha, cb, gs::fmul2(om, ab2));
                const gs::f32x2 q = gs::fmul2(dLdA, Gq);
                const gs::f32x2 qx = gs::fmul2(q, dx), qy = gs::fmul2(q, dy);
                gs::f32x2 m2[9];
                m2[0] = gs::fmul2(dr2, w2); m2[1] = gs::fmul2(dg2, w2); m2[2] = gs::fmul2(db2, w2); m2[3] = q;
                m2[4] = qx; m2[5] = qy; m2[6] = gs::fmul2(qx, dx); m2[7] = gs::fmul2(qx, dy); m2[8] = gs::fmul2(qy, dy);
                float v8[8], v9;
#pragma unroll
                for (int i = 0; i < 9; i++) {
                    float l, h;
                    gs::up2(m2[i], l, h);
                    if (i < 8) v8[i] = l + h; else v9 = l + h;
                }
                const float tot = warp_reduce8_transposed(v8, lane);
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) v9 += __shfl_xor_sync(0xFFFFFFFFu, v9, d);
                // one reduction instruction for all nine moments: lanes 0, 4, .., 28 carry the eight transposed sums, lane 1 the ninth
                float* const row = acc + (size_t)__float_as_uint(c.y) * 12u;
                const float val = lane == 1u ? v9 : tot;
                if (writer && val != 0.f) atomicAdd(row + ridx, val);
            }
            __syncwarp();
        }
        if (!more) break;
    }
}

// ---------------------------------------------------------------------------------------------------
// K8: per-Gaussian tail of tiledBackward (:517-736) applied to the nine moments.  HBM-bound (196 B/Gaussian).
// Not pinned arithmetic: gradients are graded at 1e-3 relative.
// ---------------------------------------------------------------------------------------------------
struct V9 { float V[9]; };   // upper 3x3 of the view matrix, V[c*3+r] = viewMatrix[c][r]

// The moment rows are consumed here: a row that held anything is cleared again, so the buffer is all-zero between
// frames and K7 needs no memset.  ACCUMULATE = false writes the gradient planes (zeros for a Gaussian no pixel touched)
// instead of adding to them, which saves the caller's zeroing pass and the read of the planes.
template <bool ACCUMULATE>
__global__ void __launch_bounds__(256) preprocess_backward_kernel(GsParams P, GsProjected O, float* __restrict__ acc,
                                                                  const __grid_constant__ V9 vm, float fx, float fy, uint32_t n,
                                                                  GsGradPack Gp, int stat_mode) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 m0 = *reinterpret_cast<const float4*>(acc + (size_t)i * 12);
    const float4 m1 = *reinterpret_cast<const float4*>(acc + (size_t)i * 12 + 4);
    const float m8 = acc[(size_t)i * 12 + 8];
    const bool any = (m0.x != 0.f) | (m0.y != 0.f) | (m0.z != 0.f) | (m0.w != 0.f) | (m1.x != 0.f) | (m1.y != 0.f) |
                     (m1.z != 0.f) | (m1.w != 0.f) | (m8 != 0.f);
    const uint32_t st = Gp.stride;
    if (!any) {                                         // no pixel contributed: every gradient term is +0
        if (!ACCUMULATE) {
#pragma unroll
            for (int k = 0; k < 16; k++) Gp.g[(size_t)k * st + i] = 0.f;
        }
        return;
    }
    *reinterpret_cast<float4*>(acc + (size_t)i * 12) = make_float4(0.f, 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(acc + (size_t)i * 12 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    acc[(size_t)i * 12 + 8] = 0.f;
    const float4 r0 = O.rec[3 * (size_t)i], r1 = O.rec[3 * (size_t)i + 1], r2 = O.rec[3 * (size_t)i + 2];
    const float4 ax = O.aux[i];
    const float A = -2.0f * r0.z, B = -r0.w, C = -2.0f * r1.x;       // conic
    const float sig = r1.y;
    const float col[3] = {r1.z, r1.w, r2.x};
    float g[16];
    // colour (:503-507, 699-704): gate per channel, then SH_C0
    const float mc[3] = {m0.x, m0.y, m0.z};
#pragma unroll
    for (int c = 0; c < 3; c++) g[11 + c] = (col[c] <= 0.01f || col[c] >= 0.99f) ? 0.f : mc[c] * GS_SH_C0;
    g[3] = sig * (1.0f - sig) * m0.w;                                // :517-519
    // dL/dScreenPos = sig * (A*Sqdx + B*Sqdy, C*Sqdy + B*Sqdx)  (:523-536)
    const float sx = sig * (A * m1.x + B * m1.y), sy = sig * (C * m1.y + B * m1.x);
    const float z = r2.y, vx = ax.x, vy = ax.y;
    const float txtz = vx / z, tytz = vy / z;                        // unclamped (note I) :546-547
    const float dv0 = sx * fx / z, dv1 = sy * fy / z;                // :551-554
    const float dv2 = -sx * fx * txtz / z - sy * fy * tytz / z;
    // dL/dWorld = W^T * dL/dView, W(r,c) = V[c*3+r]  (:558-565)
#pragma unroll
    for (int c = 0; c < 3; c++) g[c] = (vm.V[c * 3 + 0] * dv0 + vm.V[c * 3 + 1] * dv1) + vm.V[c * 3 + 2] * dv2;
    // dL/dConic (:570-573)
    const float k5 = -0.5f * sig;
    const float dA = k5 * m1.z, dB = k5 * 2.0f * m1.w, dC = k5 * m8;
    // conic -> cov2D (:577-596)
    const float ca = ax.z, cb = ax.w, cc = O.covc[i];
    const float denom = ca * cc - cb * cb;
    const float d2i = 1.0f / ((denom * denom) + 0.0000001f);
    const float Ga = d2i * (-cc * cc * dA + 2.0f * cb * cc * dB + (denom - ca * cc) * dC);
    const float Gc = d2i * (-ca * ca * dC + 2.0f * ca * cb * dB + (denom - ca * cc) * dA);
    const float Gb = d2i * 2.0f * (cb * cc * dA - (denom + 2.0f * cb * cb) * dB + ca * cb * dC);
    // T = J*W (rows 0,1; row 2 is zero), J from the unclamped ratios (:602-620)
    const float J00 = fx / z, J02 = -fx * txtz / z, J11 = fy / z, J12 = -fy * tytz / z;
    float T0[3], T1[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        T0[c] = J00 * vm.V[c * 3 + 0] + J02 * vm.V[c * 3 + 2];
        T1[c] = J11 * vm.V[c * 3 + 1] + J12 * vm.V[c * 3 + 2];
    }
    // dL/dSigma = T^T * [[Ga,Gb],[Gb,Gc]] * T  (:624-631; Gb on BOTH off-diagonals, as the reference)
    float dS[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float u0 = T0[r] * Ga + T1[r] * Gb, u1 = T0[r] * Gb + T1[r] * Gc;
#pragma unroll
        for (int c = 0; c < 3; c++) dS[r][c] = u0 * T0[c] + u1 * T1[c];
    }
    // scale / rotation (:635-696): exp(clamp(logs,+-5)) WITHOUT the 20:1 clamp, RAW quaternion
    const uint32_t cap = P.cap;
    const float s[3] = {gs::exp_pinned(gs::clampf(P.scale[i], -GS_MAX_SCALE, GS_MAX_SCALE)),
                        gs::exp_pinned(gs::clampf(P.scale[cap + i], -GS_MAX_SCALE, GS_MAX_SCALE)),
                        gs::exp_pinned(gs::clampf(P.scale[2 * cap + i], -GS_MAX_SCALE, GS_MAX_SCALE))};
    const float4 qq = P.rot[i];
    const float qr = qq.x, qx = qq.y, qy = qq.z, qz = qq.w;
    float R[3][3];   // R[row][col]
    R[0][0] = 1.0f - 2.0f * (qy * qy + qz * qz); R[1][0] = 2.0f * (qx * qy + qr * qz); R[2][0] = 2.0f * (qx * qz - qr * qy);
    R[0][1] = 2.0f * (qx * qy - qr * qz); R[1][1] = 1.0f - 2.0f * (qx * qx + qz * qz); R[2][1] = 2.0f * (qy * qz + qr * qx);
    R[0][2] = 2.0f * (qx * qz + qr * qy); R[1][2] = 2.0f * (qy * qz - qr * qx); R[2][2] = 1.0f - 2.0f * (qx * qx + qy * qy);
    // dL/dM = 2 * dL/dSigma * M, M = R*S  (:655-658)
    float dM[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++)
            dM[r][c] = 2.0f * ((dS[r][0] * R[0][c] + dS[r][1] * R[1][c]) + dS[r][2] * R[2][c]) * s[c];
    // dL/ds_k = (R^T dM)_kk ; dL/dlog s = dL/ds * s  (:661-666)
#pragma unroll
    for (int k = 0; k < 3; k++) g[4 + k] = ((R[0][k] * dM[0][k] + R[1][k] * dM[1][k]) + R[2][k] * dM[2][k]) * s[k];
    // D(i,j) = dL_dMt_scaled[i][j] = dL/dR(row i, col j) = dM[i][j]*s[j]  (:669-675)
    float D[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) D[r][c] = dM[r][c] * s[c];
    g[7] = 2.0f * (qz * (D[0][1] - D[1][0]) + qy * (D[2][0] - D[0][2]) + qx * (D[1][2] - D[2][1]));   // :679-696
    g[8] = 2.0f * (qy * (D[1][0] + D[0][1]) + qz * (D[2][0] + D[0][2]) + qr * (D[1][2] - D[2][1]) - 2.0f * qx * (D[2][2] + D[1][1]));
    g[9] = 2.0f * (qx * (D[1][0] + D[0][1]) + qr * (D[2][0] - D[0][2]) + qz * (D[1][2] + D[2][1]) - 2.0f * qy * (D[2][2] + D[0][0]));
    g[10] = 2.0f * (qr * (D[0][1] - D[1][0]) + qx * (D[2][0] + D[0][2]) + qy * (D[1][2] + D[2][1]) - 2.0f * qz * (D[1][1] + D[0][0]));
    if (stat_mode) {   // per-view densify statistic (density_control.mm:157-168), so it survives a sum over views
        float mag = fminf(sqrtf(sx * sx + sy * sy), 1.0f);
        const bool ok = !isnan(mag) && !isinf(mag) && mag > 0.f;
        g[14] = ok ? mag : 0.f; g[15] = ok ? 1.0f : 0.f;
    } else { g[14] = sx; g[15] = sy; }
#pragma unroll
    for (int k = 0; k < 16; k++) {
        if (ACCUMULATE) Gp.g[(size_t)k * st + i] += g[k];
        else Gp.g[(size_t)k * st + i] = g[k];
    }
}

}  // namespace

static int g_bwd_ppt = GS_BWD_PPT;
static int g_bwd_warp = 1;   // masked walk: 1 = warp-autonomous kernel, 0 = CTA-cooperative rounds
void gs_tune_bwd_ppt(int v) { if (v == 2 || v == 4) g_bwd_ppt = v; }
void gs_tune_bwd_warp(int v) { g_bwd_warp = v ? 1 : 0; }

void gs_launch_blend_backward(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                              uint32_t ntx, uint32_t nty, const float4* rendered, const float4* gt, int quantize,
                              const uint32_t* last_idx, const float* tfinal, float* acc, const uint2* vis, size_t vstride, int vis_warp_extent,
                              cudaStream_t st) {
    if (ntx * nty == 0) return;
#define GS_BWD_LAUNCH(PPT, MASKED)                                                                                  \
    blend_backward_kernel<PPT, MASKED><<<ntx * nty, 256 / PPT, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, rendered, gt, quantize, \
                                                                         last_idx, tfinal, O.tile_flags, O.tile_order, acc, vis, vstride)
    if (g_bwd_ppt == 4) GS_BWD_LAUNCH(4, false);
    else if (vis) {                                  // recorded visits; flagged tiles by the pinned variant
        if (g_bwd_warp || vis_warp_extent)
            blend_backward_warp_kernel<<<ntx * nty, 128, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, rendered, gt, quantize, last_idx, tfinal,
                                                                  O.tile_flags, O.tile_order, acc, vis, vstride);
        else GS_BWD_LAUNCH(2, true);
        blend_backward_flagged_kernel<2><<<(ntx * nty + 127u) / 128u, 128, 0, st>>>(O.rec, sorted_vals, ranges, W, H, ntx, ntx * nty, rendered, gt,
                                                                                    quantize, last_idx, tfinal, O.tile_flags, acc);
    } else GS_BWD_LAUNCH(2, false);
}
bool gs_bwd_uses_visits() { return g_bwd_ppt == 2; }

void gs_launch_preprocess_backward(const GsParams& P, const GsProjected& O, float* acc, const float* V, float fx,
                                   float fy, uint32_t n, GsGradPack G, int stat_mode, int accumulate, cudaStream_t st) {
    if (n == 0) return;
    V9 vm;
    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) vm.V[c * 3 + r] = V[c * 4 + r];
    if (accumulate) preprocess_backward_kernel<true><<<(n + 255) / 256, 256, 0, st>>>(P, O, acc, vm, fx, fy, n, G, stat_mode);
    else preprocess_backward_kernel<false><<<(n + 255) / 256, 256, 0, st>>>(P, O, acc, vm, fx, fy, n, G, stat_mode);
}
