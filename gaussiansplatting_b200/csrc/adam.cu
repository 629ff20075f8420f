// adam.cu — K9: fused Adam over the five parameter groups (replaces adamStep, shaders.metal:536-713; host side
// AdamOptimizer::step, optimizer.mm:228-296).  Pure streaming, HBM-bound: SoA planes, one thread per Gaussian,
// every load/store a coalesced 4-B (or 16-B for the quaternion) lane.  Arithmetic is pinned (no contraction,
// IEEE div/sqrt) so parameters and moments match the oracle bit for bit.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

struct AdamK { float lr[5]; float b1, b2, omb1, omb2, eps, bc1, bc2; };

__device__ __forceinline__ float adam_update(float gr, float& m, float& v, const AdamK& k, float lr) {
    m = gs::fadd(gs::fmul(k.b1, m), gs::fmul(k.omb1, gr));
    v = gs::fadd(gs::fmul(k.b2, v), gs::fmul(gs::fmul(k.omb2, gr), gr));
    const float m_hat = gs::fdiv(m, k.bc1), v_hat = gs::fdiv(v, k.bc2);
    return gs::fdiv(gs::fmul(lr, m_hat), gs::fadd(gs::fsqrt(v_hat), k.eps));
}

// All 83 inputs of a Gaussian are loaded before anything is stored: the planes are not declared restrict (parameters,
// moments and gradients are separate allocations but the compiler cannot know), so a load issued after a store would
// have to wait for it; front-loading keeps ~80 independent 4-B loads in flight per thread.
__global__ void __launch_bounds__(256) adam_kernel(GsParams P, GsGradPack G, GsAdamState S, uint32_t n, const __grid_constant__ AdamK k) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t gs_ = G.stride, cap = P.cap, sc = S.cap;
    const float clip = 0.5f;                                          // shaders.metal:582
    float g[14];
#pragma unroll
    for (int c = 0; c < 14; c++) g[c] = G.g[(size_t)c * gs_ + i];    // pos 0-2, opacity 3, scale 4-6, rot 7-10, SH-DC 11-13
    float p_pos[3], p_sc[3], p_sh[12], m_pos[3], v_pos[3], m_sc[3], v_sc[3], m_sh[12], v_sh[12];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        p_pos[c] = P.pos[c * cap + i]; p_sc[c] = P.scale[c * cap + i];
        m_pos[c] = S.m_pos[c * sc + i]; v_pos[c] = S.v_pos[c * sc + i];
        m_sc[c] = S.m_scale[c * sc + i]; v_sc[c] = S.v_scale[c * sc + i];
    }
#pragma unroll
    for (int c = 0; c < 12; c++) { p_sh[c] = P.sh[c * cap + i]; m_sh[c] = S.m_sh[c * sc + i]; v_sh[c] = S.v_sh[c * sc + i]; }
    const float4 q = P.rot[i];
    float4 m4 = S.m_rot[i], v4 = S.v_rot[i];
    float p_op = P.opacity[i], m_op = S.m_op[i], v_op = S.v_op[i];

    if (isnan(g[0]) || isnan(g[3]) || isnan(g[11]) || isinf(g[0]) || isinf(g[3])) return;   // :567-570
    if (isnan(p_pos[0]) || isinf(p_pos[0]) || fabsf(p_pos[0]) > 1e6f) return;                // :573-576
    {   // position :586-625
        float up[3];
#pragma unroll
        for (int c = 0; c < 3; c++) up[c] = adam_update(gs::clampf(g[c], -clip, clip), m_pos[c], v_pos[c], k, k.lr[0]);
        const float mag = gs::fsqrt(gs::fadd(gs::fadd(gs::fmul(up[0], up[0]), gs::fmul(up[1], up[1])), gs::fmul(up[2], up[2])));
        if (mag > 0.1f) { const float s = gs::fdiv(0.1f, mag); up[0] = gs::fmul(up[0], s); up[1] = gs::fmul(up[1], s); up[2] = gs::fmul(up[2], s); }
        const float n0 = gs::fsub(p_pos[0], up[0]), n1 = gs::fsub(p_pos[1], up[1]), n2 = gs::fsub(p_pos[2], up[2]);
        if (!isnan(n0) && !isnan(n1) && !isnan(n2) && fabsf(n0) < 1e6f && fabsf(n1) < 1e6f && fabsf(n2) < 1e6f) {
            P.pos[i] = n0; P.pos[cap + i] = n1; P.pos[2 * cap + i] = n2;
        }
#pragma unroll
        for (int c = 0; c < 3; c++) { S.m_pos[c * sc + i] = m_pos[c]; S.v_pos[c * sc + i] = v_pos[c]; }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {   // scale :630-656
        const float u = adam_update(gs::clampf(g[4 + c], -clip, clip), m_sc[c], v_sc[c], k, k.lr[1]);
        S.m_scale[c * sc + i] = m_sc[c]; S.v_scale[c * sc + i] = v_sc[c];
        P.scale[c * cap + i] = gs::clampf(gs::fsub(p_sc[c], u), -GS_MAX_SCALE_TRAIN, GS_MAX_SCALE_TRAIN);
    }
    {   // rotation :659-675
        float m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
        const float qo[4] = {q.x, q.y, q.z, q.w};
        float nr[4];
#pragma unroll
        for (int c = 0; c < 4; c++) nr[c] = gs::fsub(qo[c], adam_update(gs::clampf(g[7 + c], -clip, clip), m[c], v[c], k, k.lr[2]));
        S.m_rot[i] = make_float4(m[0], m[1], m[2], m[3]); S.v_rot[i] = make_float4(v[0], v[1], v[2], v[3]);
        const float len = gs::fsqrt(gs::fadd(gs::fadd(gs::fadd(gs::fmul(nr[0], nr[0]), gs::fmul(nr[1], nr[1])), gs::fmul(nr[2], nr[2])), gs::fmul(nr[3], nr[3])));
        P.rot[i] = (len > 0.001f) ? make_float4(gs::fdiv(nr[0], len), gs::fdiv(nr[1], len), gs::fdiv(nr[2], len), gs::fdiv(nr[3], len))
                                  : make_float4(1.f, 0.f, 0.f, 0.f);
    }
    {   // opacity :678-692
        const float u = adam_update(gs::clampf(g[3], -clip, clip), m_op, v_op, k, k.lr[3]);
        S.m_op[i] = m_op; S.v_op[i] = v_op;
        P.opacity[i] = gs::clampf(gs::fsub(p_op, u), -8.0f, 8.0f);
    }
#pragma unroll
    for (int c = 0; c < 12; c++) {   // SH :695-712 (gradients of the non-DC coefficients are identically zero)
        const float graw = (c % 4 == 0) ? g[11 + c / 4] : 0.f;
        const float u = adam_update(gs::clampf(graw, -clip, clip), m_sh[c], v_sh[c], k, k.lr[4]);
        S.m_sh[c * sc + i] = m_sh[c]; S.v_sh[c * sc + i] = v_sh[c];
        P.sh[c * cap + i] = gs::clampf(gs::fsub(p_sh[c], u), -2.0f, 2.0f);
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
                    float eps, float bc1, float bc2, cudaStream_t st) {
    if (n == 0) return;
    AdamK k;
    for (int i = 0; i < 5; i++) k.lr[i] = lrs[i];
    k.b1 = beta1; k.b2 = beta2; k.omb1 = 1.0f - beta1; k.omb2 = 1.0f - beta2; k.eps = eps; k.bc1 = bc1; k.bc2 = bc2;
    adam_kernel<<<(n + 255) / 256, 256, 0, st>>>(P, G, S, n, k);
}

void gs_launch_accumulate_stats(GsGradPack G, int stat_mode, uint32_t n, float* accum, uint32_t* count, cudaStream_t st) {
    if (n) accumulate_stats_kernel<<<(n + 255) / 256, 256, 0, st>>>(G, stat_mode, n, accum, count);
}
