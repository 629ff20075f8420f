// layout.cu — conversion between the reference's AoS wire structs (Gaussian 112 B, GaussianGradients 112 B,
// ProjectedGaussian 88 B; ply_loader.hpp:14-20, gradients.hpp:11-31, tiled_rasterizer.hpp:24-39) and the SoA
// planes the kernels stream.  Used at upload / download / inspection time only, never inside a training step.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

// A warp moves 32 consecutive 112-B structs (3584 B = 224 float4) through shared memory so that both the AoS
// side (16-B lanes, contiguous) and the SoA side (4-B lanes, contiguous) are coalesced.
constexpr int STRUCT_F = 28;   // floats per 112-B struct

__global__ void __launch_bounds__(256) params_from_aos_kernel(const float4* __restrict__ aos, uint32_t n, GsParams P) {
    __shared__ __align__(16) float s[8][32 * STRUCT_F + 4];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t g0 = (blockIdx.x * 8u + warp) * 32u;
    if (g0 >= n) return;
    const uint32_t cnt = min(32u, n - g0);
    float4* s4 = reinterpret_cast<float4*>(s[warp]);
    for (uint32_t k = lane; k < cnt * 7u; k += 32u) s4[k] = aos[(size_t)g0 * 7u + k];
    __syncwarp();
    if (lane < cnt) {
        const float* f = s[warp] + lane * STRUCT_F;
        const uint32_t i = g0 + lane, cap = P.cap;
        P.pos[i] = f[0]; P.pos[cap + i] = f[1]; P.pos[2 * cap + i] = f[2];
        P.scale[i] = f[4]; P.scale[cap + i] = f[5]; P.scale[2 * cap + i] = f[6];
        P.rot[i] = make_float4(f[8], f[9], f[10], f[11]);
        P.opacity[i] = f[12];
#pragma unroll
        for (int k = 0; k < 12; k++) P.sh[(size_t)k * cap + i] = f[13 + k];
    }
}

__global__ void __launch_bounds__(256) params_to_aos_kernel(GsParams P, uint32_t n, float4* __restrict__ aos) {
    __shared__ __align__(16) float s[8][32 * STRUCT_F + 4];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t g0 = (blockIdx.x * 8u + warp) * 32u;
    if (g0 >= n) return;
    const uint32_t cnt = min(32u, n - g0);
    if (lane < cnt) {
        float* f = s[warp] + lane * STRUCT_F;
        const uint32_t i = g0 + lane, cap = P.cap;
        f[0] = P.pos[i]; f[1] = P.pos[cap + i]; f[2] = P.pos[2 * cap + i]; f[3] = 0.f;
        f[4] = P.scale[i]; f[5] = P.scale[cap + i]; f[6] = P.scale[2 * cap + i]; f[7] = 0.f;
        const float4 q = P.rot[i];
        f[8] = q.x; f[9] = q.y; f[10] = q.z; f[11] = q.w;
        f[12] = P.opacity[i];
#pragma unroll
        for (int k = 0; k < 12; k++) f[13 + k] = P.sh[(size_t)k * cap + i];
        f[25] = 0.f; f[26] = 0.f; f[27] = 0.f;
    }
    __syncwarp();
    const float4* s4 = reinterpret_cast<const float4*>(s[warp]);
    for (uint32_t k = lane; k < cnt * 7u; k += 32u) aos[(size_t)g0 * 7u + k] = s4[k];
}

// GaussianGradients: 0-2 position, 3 opacity, 4-6 scale, 7 pad, 8-11 rotation, 12-23 sh[12], 24-25 viewspace, 26-27 pad
__global__ void __launch_bounds__(256) grads_to_aos_kernel(GsGradPack G, uint32_t n, float* __restrict__ aos) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float g[16];
#pragma unroll
    for (int k = 0; k < 16; k++) g[k] = G.g[(size_t)k * G.stride + i];
    float f[STRUCT_F];
#pragma unroll
    for (int k = 0; k < STRUCT_F; k++) f[k] = 0.f;
    f[0] = g[0]; f[1] = g[1]; f[2] = g[2]; f[3] = g[3]; f[4] = g[4]; f[5] = g[5]; f[6] = g[6];
    f[8] = g[7]; f[9] = g[8]; f[10] = g[9]; f[11] = g[10];
    f[12] = g[11]; f[16] = g[12]; f[20] = g[13];
    f[24] = g[14]; f[25] = g[15];
    float4* o = reinterpret_cast<float4*>(aos + (size_t)i * STRUCT_F);
#pragma unroll
    for (int k = 0; k < 7; k++) o[k] = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
}
// The packed buffer carries the 14 live gradients (+2 view-space); SH-rest gradients (sh 1-3,5-7,9-11) are zero
// by construction in the reference (tiled_shaders.metal:699-704) and are not stored.
__global__ void __launch_bounds__(256) grads_from_aos_kernel(const float* __restrict__ aos, uint32_t n, GsGradPack G) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* f = aos + (size_t)i * STRUCT_F;
    const int src[16] = {0, 1, 2, 3, 4, 5, 6, 8, 9, 10, 11, 12, 16, 20, 24, 25};
#pragma unroll
    for (int k = 0; k < 16; k++) G.g[(size_t)k * G.stride + i] = f[src[k]];
}

// ProjectedGaussian, 22 floats: 0-1 screenPos, 2-4 conic, 5 depth, 6 opacity, 7-9 colour, 10 radius, 11-14 tile
// bbox (u32), 15 pad, 16-17 viewPos_xy, 18-20 cov2D, 21 pad
__global__ void __launch_bounds__(256) projected_to_aos_kernel(GsProjected O, uint32_t n, uint32_t* __restrict__ aos) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 r0 = O.rec[3 * (size_t)i], r1 = O.rec[3 * (size_t)i + 1], r2 = O.rec[3 * (size_t)i + 2];
    const uint4 bb = O.bbox[i];
    const float4 ax = O.aux[i];
    uint32_t* o = aos + (size_t)i * 22;
    o[0] = __float_as_uint(r0.x); o[1] = __float_as_uint(r0.y);
    o[2] = __float_as_uint(gs::fmul(-2.0f, r0.z)); o[3] = __float_as_uint(-r0.w); o[4] = __float_as_uint(gs::fmul(-2.0f, r1.x));
    o[5] = __float_as_uint(r2.y); o[6] = __float_as_uint(r1.y);
    o[7] = __float_as_uint(r1.z); o[8] = __float_as_uint(r1.w); o[9] = __float_as_uint(r2.x);
    o[10] = __float_as_uint(r2.z);
    o[11] = bb.x; o[12] = bb.y; o[13] = bb.z; o[14] = bb.w; o[15] = 0u;
    o[16] = __float_as_uint(ax.x); o[17] = __float_as_uint(ax.y);
    o[18] = __float_as_uint(ax.z); o[19] = __float_as_uint(ax.w); o[20] = __float_as_uint(O.covc[i]); o[21] = 0u;
}

// RGBA8Unorm quantisation of the render target (SURVEY note G; mtl_engine.mm:722)
__global__ void rgba32f_to_rgba8_kernel(const float4* __restrict__ img, uint32_t npix, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const float4 p = img[i];
    const uint32_t r = (uint32_t)floorf(gs::clampf(p.x, 0.f, 1.f) * 255.0f + 0.5f), g = (uint32_t)floorf(gs::clampf(p.y, 0.f, 1.f) * 255.0f + 0.5f);
    const uint32_t b = (uint32_t)floorf(gs::clampf(p.z, 0.f, 1.f) * 255.0f + 0.5f), a = (uint32_t)floorf(gs::clampf(p.w, 0.f, 1.f) * 255.0f + 0.5f);
    out[i] = r | (g << 8) | (b << 16) | (a << 24);
}
__global__ void rgba8_to_rgba32f_kernel(const uint32_t* __restrict__ img, uint32_t npix, float4* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const uint32_t v = img[i];
    out[i] = make_float4((float)(v & 255u) / 255.0f, (float)((v >> 8) & 255u) / 255.0f, (float)((v >> 16) & 255u) / 255.0f,
                         (float)(v >> 24) / 255.0f);
}
// opacity reset, mtl_engine.mm:1179-1185: opacity = min(opacity, max_raw)
__global__ void clamp_opacity_kernel(GsParams P, uint32_t n, float max_raw) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && P.opacity[i] > max_raw) P.opacity[i] = max_raw;
}

// xyz -> float4 positions for the viewer sort (updatePositionBuffer, mtl_engine.mm:1028-1036)
__global__ void positions_kernel(GsParams P, uint32_t n, float4* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = make_float4(P.pos[i], P.pos[P.cap + i], P.pos[2 * (size_t)P.cap + i], 0.f);
}

}  // namespace

void gs_launch_positions(GsParams P, uint32_t n, float4* out, cudaStream_t st) {
    if (n) positions_kernel<<<(n + 255) / 256, 256, 0, st>>>(P, n, out);
}
void gs_launch_params_from_aos(const void* aos112, uint32_t n, GsParams P, cudaStream_t st) {
    if (n) params_from_aos_kernel<<<(n + 255) / 256, 256, 0, st>>>((const float4*)aos112, n, P);
}
void gs_launch_params_to_aos(GsParams P, uint32_t n, void* aos112, cudaStream_t st) {
    if (n) params_to_aos_kernel<<<(n + 255) / 256, 256, 0, st>>>(P, n, (float4*)aos112);
}
void gs_launch_grads_to_aos(GsGradPack G, uint32_t n, void* aos112, cudaStream_t st) {
    if (n) grads_to_aos_kernel<<<(n + 255) / 256, 256, 0, st>>>(G, n, (float*)aos112);
}
void gs_launch_grads_from_aos(const void* aos112, uint32_t n, GsGradPack G, cudaStream_t st) {
    if (n) grads_from_aos_kernel<<<(n + 255) / 256, 256, 0, st>>>((const float*)aos112, n, G);
}
void gs_launch_projected_to_aos(const GsProjected& O, uint32_t n, void* aos88, cudaStream_t st) {
    if (n) projected_to_aos_kernel<<<(n + 255) / 256, 256, 0, st>>>(O, n, (uint32_t*)aos88);
}
void gs_launch_rgba32f_to_rgba8(const float4* img, uint32_t npix, uint32_t* out, cudaStream_t st) {
    if (npix) rgba32f_to_rgba8_kernel<<<(npix + 255) / 256, 256, 0, st>>>(img, npix, out);
}
void gs_launch_rgba8_to_rgba32f(const uint32_t* img, uint32_t npix, float4* out, cudaStream_t st) {
    if (npix) rgba8_to_rgba32f_kernel<<<(npix + 255) / 256, 256, 0, st>>>(img, npix, out);
}
void gs_launch_clamp_opacity(GsParams P, uint32_t n, float max_raw, cudaStream_t st) {
    if (n) clamp_opacity_kernel<<<(n + 255) / 256, 256, 0, st>>>(P, n, max_raw);
}
