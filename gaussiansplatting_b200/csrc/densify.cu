// densify.cu — K10/K11: density control as mark -> scan -> scatter on the device (replaces the CPU-only
// DensityController::apply, density_control.mm:188-501: 8 GCD threads for the decision pass, a serial rebuild
// loop, and libc rand()).  Output order is the reference's: keep -> 1, clone -> original + identical copy
// adjacent, split -> two children REPLACING the parent, prune -> 0 (:394-483).  rand() is replaced by a
// counter-based generator keyed on (seed, iteration, parent, component) (SURVEY note L); exp by gs::exp_pinned.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

__device__ __forceinline__ uint32_t split_rand_u32(uint32_t seed, uint32_t iteration, uint32_t parent, uint32_t comp) {
    unsigned long long z = (((unsigned long long)seed << 32) | iteration) +
                           0x9E3779B97F4A7C15ull * ((unsigned long long)parent * 3ull + comp + 1ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (uint32_t)(z >> 32);
}

__device__ __forceinline__ float max_scale(const GsParams& P, uint32_t i) {   // :280-283
    const float a = gs::exp_pinned(gs::clampf(P.scale[i], -4.0f, 4.0f));
    const float b = gs::exp_pinned(gs::clampf(P.scale[P.cap + i], -4.0f, 4.0f));
    const float c = gs::exp_pinned(gs::clampf(P.scale[2 * (size_t)P.cap + i], -4.0f, 4.0f));
    return fmaxf(fmaxf(a, b), c);
}

__global__ void __launch_bounds__(256) densify_mark_kernel(GsParams P, const float* __restrict__ accum, const uint32_t* __restrict__ count,
                                                           GsDensifyArgs a, uint32_t* __restrict__ markers, uint32_t* __restrict__ counts3) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t marker = 0;
    if (i < a.n) {
        const bool canDensify = a.iteration > 500u && a.iteration < 15000u;              // :213
        const bool screenPrune = a.iteration > 3000u;                                     // :233
        const float opacity = gs::fdiv(1.0f, gs::fadd(1.0f, gs::exp_pinned(-P.opacity[i])));   // :274
        const uint32_t c = count[i];
        const float avgGrad = c > 0u ? gs::fdiv(accum[i], (float)c) : 0.0f;               // :277
        const float ms = max_scale(P, i);
        bool prune = opacity < 0.005f;                                                    // :289
        if (screenPrune) {                                                                // :295-307
            if (ms > gs::fmul(0.1f, a.scene_extent)) prune = true;
            const float safeDepth = fmaxf(a.avg_depth, 0.1f);                             // :61-62
            const float screenRadius = gs::fdiv(gs::fmul(gs::fmul(a.focal, ms), 3.0f), safeDepth);   // :72
            const float px = gs::fmul(gs::fdiv(screenRadius, a.image_width), a.image_width);         // :75, :303
            if (px > 40.0f) prune = true;
        }
        if (prune) marker = 1;
        else if (canDensify && avgGrad > 0.0002f) marker = (ms > gs::fmul(0.01f, a.scene_extent)) ? 3u : 2u;   // :313-337
        markers[i] = marker;
    }
    const uint32_t m1 = __popc(__ballot_sync(0xFFFFFFFFu, marker == 1u));
    const uint32_t m2 = __popc(__ballot_sync(0xFFFFFFFFu, marker == 2u));
    const uint32_t m3 = __popc(__ballot_sync(0xFFFFFFFFu, marker == 3u));
    if (gs::lane_id() == 0) {
        if (m1) atomicAdd(&counts3[0], m1);
        if (m2) atomicAdd(&counts3[1], m2);
        if (m3) atomicAdd(&counts3[2], m3);
    }
}

// excess over the cap, split between clones (first) and splits (then)  (:358-382)
__device__ __forceinline__ void cap_excess(const uint32_t* counts3, uint32_t n, uint32_t maxG, uint32_t& exClone, uint32_t& exSplit) {
    const unsigned long long newCount = (unsigned long long)n - counts3[0] + counts3[1] + counts3[2];
    exClone = exSplit = 0;
    if (newCount > maxG) {
        unsigned long long excess = newCount - maxG;
        exClone = (uint32_t)min(excess, (unsigned long long)counts3[1]);
        excess -= exClone;
        exSplit = (uint32_t)min(excess, (unsigned long long)counts3[2]);
    }
}

__global__ void flag_marker_kernel(const uint32_t* __restrict__ markers, uint32_t n, uint32_t which, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = markers[i] == which ? 1u : 0u;
}
// un-mark the first `excess` Gaussians (index order) carrying marker `which`
__global__ void unmark_kernel(uint32_t* __restrict__ markers, uint32_t n, uint32_t maxG, const uint32_t* __restrict__ counts3,
                              const uint32_t* __restrict__ rank, uint32_t which) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t exClone, exSplit;
    cap_excess(counts3, n, maxG, exClone, exSplit);
    const uint32_t ex = which == 2u ? exClone : exSplit;
    if (markers[i] == which && rank[i] < ex) markers[i] = 0u;
}
__global__ void out_count_kernel(const uint32_t* __restrict__ markers, uint32_t n, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const uint32_t m = markers[i]; out[i] = m == 0u ? 1u : (m == 1u ? 0u : 2u); }
}
__global__ void densify_stats_kernel(const uint32_t* counts3, uint32_t n, uint32_t maxG, const uint32_t* out_off, uint32_t* stats4) {
    uint32_t exClone, exSplit;
    cap_excess(counts3, n, maxG, exClone, exSplit);
    stats4[0] = counts3[0]; stats4[1] = counts3[1] - exClone; stats4[2] = counts3[2] - exSplit; stats4[3] = out_off[n];
}

__device__ __forceinline__ void copy_gaussian(const GsParams& A, uint32_t i, const GsParams& B, uint32_t o) {
#pragma unroll
    for (int k = 0; k < 3; k++) { B.pos[(size_t)k * B.cap + o] = A.pos[(size_t)k * A.cap + i]; B.scale[(size_t)k * B.cap + o] = A.scale[(size_t)k * A.cap + i]; }
    B.rot[o] = A.rot[i];
    B.opacity[o] = A.opacity[i];
#pragma unroll
    for (int k = 0; k < 12; k++) B.sh[(size_t)k * B.cap + o] = A.sh[(size_t)k * A.cap + i];
}

__global__ void __launch_bounds__(256) densify_scatter_kernel(GsParams A, GsParams B, const uint32_t* __restrict__ markers,
                                                              const uint32_t* __restrict__ out_off, GsDensifyArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const uint32_t m = markers[i];
    if (m == 1u) return;
    const uint32_t o = out_off[i];
    copy_gaussian(A, i, B, o);
    if (m == 0u) return;
    copy_gaussian(A, i, B, o + 1u);                                   // clone: identical, adjacent (:408-421)
    if (m == 2u) return;
    // split (:422-482)
    float sc[3], r[3], p[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        sc[k] = gs::exp_pinned(gs::clampf(A.scale[(size_t)k * A.cap + i], -4.0f, 4.0f));   // :429-433
        const uint32_t h = split_rand_u32(a.seed, a.iteration, i, (uint32_t)k);           // :440-442
        r[k] = gs::fmul(gs::fsub(gs::fmul((float)(h >> 8), 1.0f / 16777216.0f), 0.5f), 2.0f);
        p[k] = A.pos[(size_t)k * A.cap + i];
    }
    const float rn = gs::fsqrt(gs::fadd(gs::fadd(gs::fmul(r[0], r[0]), gs::fmul(r[1], r[1])), gs::fmul(r[2], r[2])));   // :443-446
    if (rn > 0.001f) { r[0] = gs::fdiv(r[0], rn); r[1] = gs::fdiv(r[1], rn); r[2] = gs::fdiv(r[2], rn); }
    const float o0 = gs::fmul(r[0], sc[0]), o1 = gs::fmul(r[1], sc[1]), o2 = gs::fmul(r[2], sc[2]);   // :448
    const float4 q = A.rot[i];                                        // raw quaternion (:452-458)
    const float w = q.x, x = q.y, y = q.z, z = q.w;
    const float R00 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(y, y), gs::fmul(z, z))));
    const float R10 = gs::fmul(2.0f, gs::fadd(gs::fmul(x, y), gs::fmul(w, z)));
    const float R20 = gs::fmul(2.0f, gs::fsub(gs::fmul(x, z), gs::fmul(w, y)));
    const float R01 = gs::fmul(2.0f, gs::fsub(gs::fmul(x, y), gs::fmul(w, z)));
    const float R11 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(x, x), gs::fmul(z, z))));
    const float R21 = gs::fmul(2.0f, gs::fadd(gs::fmul(y, z), gs::fmul(w, x)));
    const float R02 = gs::fmul(2.0f, gs::fadd(gs::fmul(x, z), gs::fmul(w, y)));
    const float R12 = gs::fmul(2.0f, gs::fsub(gs::fmul(y, z), gs::fmul(w, x)));
    const float R22 = gs::fsub(1.0f, gs::fmul(2.0f, gs::fadd(gs::fmul(x, x), gs::fmul(y, y))));
    float ro[3];                                                      // :459 R * offset
    ro[0] = gs::fadd(gs::fadd(gs::fmul(R00, o0), gs::fmul(R01, o1)), gs::fmul(R02, o2));
    ro[1] = gs::fadd(gs::fadd(gs::fmul(R10, o0), gs::fmul(R11, o1)), gs::fmul(R12, o2));
    ro[2] = gs::fadd(gs::fadd(gs::fmul(R20, o0), gs::fmul(R21, o1)), gs::fmul(R22, o2));
    const float lsf = -0.47000363f;                                   // logf(1/1.6) :425-426
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float ns = gs::fadd(A.scale[(size_t)k * A.cap + i], lsf);                   // :465-469
        B.pos[(size_t)k * B.cap + o] = gs::fadd(p[k], ro[k]);        // :463
        B.pos[(size_t)k * B.cap + o + 1u] = gs::fsub(p[k], ro[k]);   // :477
        B.scale[(size_t)k * B.cap + o] = ns; B.scale[(size_t)k * B.cap + o + 1u] = ns;
    }
}

// corrected optimizer-state mode (SURVEY note D / 8f-4): carry the Adam moments of every surviving Gaussian to its new index.
// A kept Gaussian and the first copy of a clone keep their moments; the second copy of a clone and both halves of a split start
// from zero (the destination buffer is zero-filled by the caller).  46 floats per Gaussian: 38 planes + two float4 rows.
__global__ void __launch_bounds__(256) permute_moments_kernel(const float* __restrict__ src, uint32_t src_cap, float* __restrict__ dst,
                                                              uint32_t dst_cap, const uint32_t* __restrict__ markers,
                                                              const uint32_t* __restrict__ out_off, uint32_t n, uint32_t new_n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t m = markers[i];
    if (m == 1u || m == 3u) return;
    const uint32_t o = out_off[i];
    if (o >= new_n) return;
#pragma unroll 2
    for (int k = 0; k < 38; k++) dst[(size_t)k * dst_cap + o] = src[(size_t)k * src_cap + i];
    const float4* sr = reinterpret_cast<const float4*>(src + 38 * (size_t)src_cap);
    float4* dr = reinterpret_cast<float4*>(dst + 38 * (size_t)dst_cap);
    dr[o] = sr[i];
    dr[(size_t)dst_cap + o] = sr[(size_t)src_cap + i];
}

}  // namespace

void gs_launch_permute_moments(const float* src, uint32_t src_cap, float* dst, uint32_t dst_cap, const uint32_t* markers,
                               const uint32_t* out_off, uint32_t n, uint32_t new_n, cudaStream_t st) {
    if (n) permute_moments_kernel<<<(n + 255) / 256, 256, 0, st>>>(src, src_cap, dst, dst_cap, markers, out_off, n, new_n);
}

void gs_launch_densify_mark(GsParams P, const float* accum, const uint32_t* count, GsDensifyArgs a, uint32_t* markers,
                            uint32_t* counts3, cudaStream_t st) {
    cudaMemsetAsync(counts3, 0, 3 * sizeof(uint32_t), st);
    if (a.n) densify_mark_kernel<<<(a.n + 255) / 256, 256, 0, st>>>(P, accum, count, a, markers, counts3);
}

void gs_launch_densify_plan(uint32_t* markers, uint32_t n, uint32_t max_gaussians, uint32_t* counts3, uint32_t* rankbuf,
                            uint32_t* out_off, uint32_t* stats4, void* scan_scratch, cudaStream_t st) {
    const uint32_t blocks = (n + 255) / 256;
    if (n) {
        for (uint32_t which = 2; which <= 3; which++) {               // clones first, then splits (:364-379)
            flag_marker_kernel<<<blocks, 256, 0, st>>>(markers, n, which, rankbuf);
            gs_launch_exclusive_scan(rankbuf, nullptr, rankbuf, n, stats4 + 3, scan_scratch, st);
            unmark_kernel<<<blocks, 256, 0, st>>>(markers, n, max_gaussians, counts3, rankbuf, which);
        }
        out_count_kernel<<<blocks, 256, 0, st>>>(markers, n, out_off);
    }
    gs_launch_exclusive_scan(out_off, nullptr, out_off, n, stats4 + 3, scan_scratch, st);
    densify_stats_kernel<<<1, 1, 0, st>>>(counts3, n, max_gaussians, out_off, stats4);
}

void gs_launch_densify_scatter(GsParams Pin, GsParams Pout, const uint32_t* markers, const uint32_t* out_off,
                               GsDensifyArgs a, cudaStream_t st) {
    if (a.n) densify_scatter_kernel<<<(a.n + 255) / 256, 256, 0, st>>>(Pin, Pout, markers, out_off, a);
}
