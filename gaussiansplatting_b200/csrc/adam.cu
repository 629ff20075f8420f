// adam.cu — K9: fused Adam over the five parameter groups (replaces adamStep, shaders.metal:536-713; host side
// AdamOptimizer::step, optimizer.mm:228-296).  Pure streaming, HBM-bound: SoA planes read and written in 16-B lanes.  Arithmetic is pinned (no contraction,
// IEEE div/sqrt) so parameters and moments match the oracle bit for bit.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

struct AdamK { float lr[5]; float b1, b2, omb1, omb2, eps, bc1, bc2; };

// A parameter whose gradient and both moments are zero (nine of the twelve SH planes always, every parameter of a Gaussian
// no pixel touched until it first receives a gradient) keeps m = v = +0 and gets the update +0: exactly what the
// arithmetic below yields (b1*0 + omb1*(+-0) = +0, (lr * 0/bc1) / (sqrt(0/bc2) + eps) = +0), without sending zeros
// through the IEEE division's slow path.
__device__ __forceinline__ bool adam_idle(float gr, float m, float v) {
    return gr == 0.0f && __float_as_uint(m) == 0u && __float_as_uint(v) == 0u;
}
__device__ __forceinline__ float adam_update(float gr, float& m, float& v, const AdamK& k, float lr) {
    if (adam_idle(gr, m, v)) return 0.0f;
    m = gs::fadd(gs::fmul(k.b1, m), gs::fmul(k.omb1, gr));
    v = gs::fadd(gs::fmul(k.b2, v), gs::fmul(gs::fmul(k.omb2, gr), gr));
    const float m_hat = gs::fdiv(m, k.bc1), v_hat = gs::fdiv(v, k.bc2);
    return gs::fdiv(gs::fmul(lr, m_hat), gs::fadd(gs::fsqrt(v_hat), k.eps));
}

// Plane-streaming layout of the step.  Apart from three couplings the update of every scalar parameter is independent,
// so the kernel runs as 18 JOBS (blockIdx.y) that each stream a handful of planes with 16-B vector lanes, four
// consecutive Gaussians per thread: job 0 = position (3 planes together: the joint 0.1 magnitude clip), job 1 = rotation
// (float4 planes + renormalisation), jobs 2-4 scale, 5 opacity, 6-17 the SH planes.  A CTA then touches 7-13 streams of
// 4 KB each instead of 152 streams of 1 KB (one-thread-per-Gaussian version: 2.0 TB/s), which is what DRAM pages want.
// The third coupling is the reference's early return (a Gaussian with a non-finite gradient or position is skipped as a
// whole, shaders.metal:567-576): adam_skip_kernel evaluates that predicate once per step into a byte per Gaussian.
struct Skip4 { bool s[4]; };
__device__ __forceinline__ Skip4 adam_skip4(const GsParams& P, const GsGradPack& G, size_t i4) {
    const float4 g0 = *reinterpret_cast<const float4*>(G.g + i4), g3 = *reinterpret_cast<const float4*>(G.g + 3 * (size_t)G.stride + i4),
                 g11 = *reinterpret_cast<const float4*>(G.g + 11 * (size_t)G.stride + i4), p0 = *reinterpret_cast<const float4*>(P.pos + i4);
    const float a[4] = {g0.x, g0.y, g0.z, g0.w}, b[4] = {g3.x, g3.y, g3.z, g3.w}, c[4] = {g11.x, g11.y, g11.z, g11.w},
                d[4] = {p0.x, p0.y, p0.z, p0.w};
    Skip4 r;
#pragma unroll
    for (int e = 0; e < 4; e++)
        r.s[e] = isnan(a[e]) || isnan(b[e]) || isnan(c[e]) || isinf(a[e]) || isinf(b[e]) ||      // :567-570
                 isnan(d[e]) || isinf(d[e]) || fabsf(d[e]) > 1e6f;                                // :573-576
    return r;
}
// The predicate is taken ONCE per step, from a snapshot of the four planes it reads, into one byte per Gaussian (a 32-bit word
// per thread of the main kernel): the 18 jobs then read 4 B instead of 64 B each, and no job can see a position that job 0 of
// the same launch has already updated.
__global__ void __launch_bounds__(256) adam_skip_kernel(GsParams P, GsGradPack G, uint32_t n, uint32_t* __restrict__ skipw) {
    const size_t i4 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 >= n) return;
    const Skip4 r = adam_skip4(P, G, i4);
    skipw[i4 >> 2] = (r.s[0] ? 1u : 0u) | (r.s[1] ? 0x100u : 0u) | (r.s[2] ? 0x10000u : 0u) | (r.s[3] ? 0x1000000u : 0u);
}
__device__ __forceinline__ void ld4(const float* p, float (&v)[4]) { const float4 t = *reinterpret_cast<const float4*>(p); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }

// clean[0] (read): 1 = every non-DC SH value of [0,n) is inside the clamp and its moments are zero, so those nine jobs are the
// identity and return at once.  clean[1] (preset to 1 by the host, read by the NEXT step): cleared by a job that had to
// leave an out-of-range value behind because its Gaussian is skipped.
__global__ void __launch_bounds__(256, 4) adam_kernel(GsParams P, GsGradPack G, GsAdamState S, uint32_t n, const __grid_constant__ AdamK k,
                                                      const uint32_t* __restrict__ clean_in, uint32_t* __restrict__ clean_out,
                                                      const uint32_t* __restrict__ skipw) {
    const size_t i4 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 >= n) return;
    const uint32_t job = blockIdx.y;
    const bool sh_rest = job >= 6 && ((job - 6) % 4) != 0;           // SH planes other than the three DC ones: no gradient plane
    if (sh_rest && *clean_in) return;
    const size_t gs_ = G.stride, cap = P.cap, sc = S.cap;
    const float clip = 0.5f;                                          // shaders.metal:582
    const uint32_t skw = skipw[i4 >> 2];
    Skip4 sk;
#pragma unroll
    for (int e = 0; e < 4; e++) sk.s[e] = ((skw >> (8 * e)) & 0xFFu) != 0u;
    const bool full = i4 + 4 <= n;                                    // a partial last vector is handled element-wise on store
    bool live[4];
#pragma unroll
    for (int e = 0; e < 4; e++) live[e] = !sk.s[e] && (i4 + e < n);

    if (job == 0) {   // position :586-625
        float g[3][4], p[3][4], m[3][4], v[3][4];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            ld4(G.g + c * gs_ + i4, g[c]); ld4(P.pos + c * cap + i4, p[c]);
            ld4(S.m_pos + c * sc + i4, m[c]); ld4(S.v_pos + c * sc + i4, v[c]);
        }
#pragma unroll
        for (int e = 0; e < 4; e++) {
            if (!live[e]) continue;
            float up[3];
#pragma unroll
            for (int c = 0; c < 3; c++) up[c] = adam_update(gs::clampf(g[c][e], -clip, clip), m[c][e], v[c][e], k, k.lr[0]);
            const float mag = gs::fsqrt(gs::fadd(gs::fadd(gs::fmul(up[0], up[0]), gs::fmul(up[1], up[1])), gs::fmul(up[2], up[2])));
            if (mag > 0.1f) { const float s = gs::fdiv(0.1f, mag); up[0] = gs::fmul(up[0], s); up[1] = gs::fmul(up[1], s); up[2] = gs::fmul(up[2], s); }
            const float n0 = gs::fsub(p[0][e], up[0]), n1 = gs::fsub(p[1][e], up[1]), n2 = gs::fsub(p[2][e], up[2]);
            if (!isnan(n0) && !isnan(n1) && !isnan(n2) && fabsf(n0) < 1e6f && fabsf(n1) < 1e6f && fabsf(n2) < 1e6f) {
                p[0][e] = n0; p[1][e] = n1; p[2][e] = n2;
            }
        }
#pragma unroll
        for (int c = 0; c < 3; c++) {
            if (full) { st4(P.pos + c * cap + i4, p[c]); st4(S.m_pos + c * sc + i4, m[c]); st4(S.v_pos + c * sc + i4, v[c]); }
            else for (int e = 0; e < 4; e++) if (i4 + e < n) { P.pos[c * cap + i4 + e] = p[c][e]; S.m_pos[c * sc + i4 + e] = m[c][e]; S.v_pos[c * sc + i4 + e] = v[c][e]; }
        }
        return;
    }
    if (job == 1) {   // rotation :659-675
        float g[4][4];
#pragma unroll
        for (int c = 0; c < 4; c++) ld4(G.g + (7 + c) * gs_ + i4, g[c]);
#pragma unroll
        for (int e = 0; e < 4; e++) {
            if (!live[e]) continue;
            const float4 q = P.rot[i4 + e], m4 = S.m_rot[i4 + e], v4 = S.v_rot[i4 + e];
            float m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
            const float qo[4] = {q.x, q.y, q.z, q.w};
            float nr[4];
#pragma unroll
            for (int c = 0; c < 4; c++) nr[c] = gs::fsub(qo[c], adam_update(gs::clampf(g[c][e], -clip, clip), m[c], v[c], k, k.lr[2]));
            S.m_rot[i4 + e] = make_float4(m[0], m[1], m[2], m[3]); S.v_rot[i4 + e] = make_float4(v[0], v[1], v[2], v[3]);
            const float len = gs::fsqrt(gs::fadd(gs::fadd(gs::fadd(gs::fmul(nr[0], nr[0]), gs::fmul(nr[1], nr[1])), gs::fmul(nr[2], nr[2])), gs::fmul(nr[3], nr[3])));
            P.rot[i4 + e] = (len > 0.001f) ? make_float4(gs::fdiv(nr[0], len), gs::fdiv(nr[1], len), gs::fdiv(nr[2], len), gs::fdiv(nr[3], len))
                                           : make_float4(1.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    // scalar planes: scale (jobs 2-4) :630-656, opacity (5) :678-692, SH (6-17) :695-712
    float *pp, *pm, *pv;
    const float* pg = nullptr;
    float lr, lo, hi;
    if (job < 5) { const uint32_t c = job - 2; pp = P.scale + c * cap; pm = S.m_scale + c * sc; pv = S.v_scale + c * sc; pg = G.g + (4 + c) * gs_; lr = k.lr[1]; lo = -GS_MAX_SCALE_TRAIN; hi = GS_MAX_SCALE_TRAIN; }
    else if (job == 5) { pp = P.opacity; pm = S.m_op; pv = S.v_op; pg = G.g + 3 * gs_; lr = k.lr[3]; lo = -8.0f; hi = 8.0f; }
    else {   // gradients of the non-DC coefficients are identically zero
        const uint32_t c = job - 6; pp = P.sh + c * cap; pm = S.m_sh + c * sc; pv = S.v_sh + c * sc;
        if (c % 4 == 0) pg = G.g + (11 + c / 4) * gs_;
        lr = k.lr[4]; lo = -2.0f; hi = 2.0f;
    }
    float g[4] = {0.f, 0.f, 0.f, 0.f}, p[4], m[4], v[4];
    if (pg) ld4(pg + i4, g);
    ld4(pp + i4, p); ld4(pm + i4, m); ld4(pv + i4, v);
    bool moved = false, pchanged = false, dirty = false;             // moments / parameter different from what was loaded
#pragma unroll
    for (int e = 0; e < 4; e++) {
        if (!live[e]) { dirty |= (i4 + e < n) && !(p[e] >= lo && p[e] <= hi); continue; }
        const float gc = gs::clampf(g[e], -clip, clip);
        moved |= !adam_idle(gc, m[e], v[e]);
        const float u = adam_update(gc, m[e], v[e], k, lr);
        const float pn = gs::clampf(gs::fsub(p[e], u), lo, hi);
        pchanged |= __float_as_uint(pn) != __float_as_uint(p[e]);
        p[e] = pn;
    }
    if (sh_rest && (dirty || moved)) *clean_out = 0u;                // an unclamped value of a skipped Gaussian, or live moments
    // an idle vector (all four parameters idle and inside the clamp) writes nothing: 3 of its 6-7 streams disappear
    if (full) {
        if (pchanged) st4(pp + i4, p);
        if (moved) { st4(pm + i4, m); st4(pv + i4, v); }
    } else if (pchanged || moved) {
        for (int e = 0; e < 4; e++) if (i4 + e < n) { pp[i4 + e] = p[e]; pm[i4 + e] = m[e]; pv[i4 + e] = v[e]; }
    }
}

// A13: DensityController::accumulateGradients (density_control.mm:121-185) on the packed buffer
__global__ void accumulate_stats_kernel(GsGradPack G, int stat_mode, uint32_t n, float* __restrict__ accum, uint32_t* __restrict__ count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float a = G.g[14 * (size_t)G.stride + i], b = G.g[15 * (size_t)G.stride + i];
    if (stat_mode) {              // columns already hold sum of per-view clamped norms and the number of views counted
        if (b > 0.f) { accum[i] = gs::fadd(accum[i], a); count[i] += (uint32_t)b; }
        return;
    }
    float mag = gs::fsqrt(gs::fadd(gs::fmul(a, a), gs::fmul(b, b)));   // :157-158
    mag = fminf(mag, 1.0f);                                            // :163 (NaN stays out: fminf(NaN,1)=1 — see below)
    if (isnan(a) || isnan(b)) return;                                  // std::min(NaN,1) = NaN in the reference -> not accumulated
    if (!isinf(mag) && mag > 0.0f) { accum[i] = gs::fadd(accum[i], mag); count[i] += 1u; }   // :166-168
}

}  // namespace

void gs_launch_adam(GsParams P, GsGradPack G, GsAdamState S, uint32_t n, const float lrs[5], float beta1, float beta2,
                    float eps, float bc1, float bc2, const uint32_t* clean_in, uint32_t* clean_out, uint32_t* skip_words, cudaStream_t st) {
    if (n == 0) return;
    AdamK k;
    for (int i = 0; i < 5; i++) k.lr[i] = lrs[i];
    k.b1 = beta1; k.b2 = beta2; k.omb1 = 1.0f - beta1; k.omb2 = 1.0f - beta2; k.eps = eps; k.bc1 = bc1; k.bc2 = bc2;
    adam_skip_kernel<<<(n + 1023) / 1024, 256, 0, st>>>(P, G, n, skip_words);
    adam_kernel<<<dim3((n + 1023) / 1024, 18), 256, 0, st>>>(P, G, S, n, k, clean_in, clean_out, skip_words);
}

void gs_launch_accumulate_stats(GsGradPack G, int stat_mode, uint32_t n, float* accum, uint32_t* count, cudaStream_t st) {
    if (n) accumulate_stats_kernel<<<(n + 255) / 256, 256, 0, st>>>(G, stat_mode, n, accum, count);
}
