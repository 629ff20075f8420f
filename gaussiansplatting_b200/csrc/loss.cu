// loss.cu — the logging loss of MTLEngine::computeLoss (mtl_engine.mm:769-853): computeL1Loss, computeSSIM,
// computeCombinedLoss and reduceLoss (shaders.metal:319-511) fused into one pass.  Not on the gradient path (the
// backward's loss gradient is sign(rendered - gt)/3, tiled_shaders.metal:418-423); it exists so that trainStep can
// report the number the reference prints.
//
// One CTA per 16x16 pixel tile.  The mean-RGB of both images over the tile plus the 5-pixel halo (clamp to edge, :391-393)
// is staged once in shared memory; every thread then runs the reference's two 11x11 passes (weighted means, then weighted
// variances about those means — the two-pass form, not E[x^2]-mu^2) out of shared memory, adds the L1 term, and the tile's
// sum goes to one double atomic.  The 121 window weights exp(-(dx^2+dy^2)/(2*1.5^2)) are computed once per CTA.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

constexpr int LR = 5, LW = 11, LT = 16, LS = LT + 2 * LR;   // window radius / size, tile, staged side

// PARTIAL = false: the tile's sum goes to one double atomic on *total (gs_loss_compute).  PARTIAL = true: it is stored in
// total[tile] instead (gs_loss_compute_noalloc below: per-tile partial sums in static scratch, summed in a fixed order).
template <bool PARTIAL>
__global__ void __launch_bounds__(LT * LT) loss_kernel(const float4* __restrict__ rendered, const float4* __restrict__ gt,
                                                       uint32_t W, uint32_t H, float lambda, int quantize, double* __restrict__ total) {
    __shared__ float s_x[LS][LS + 1], s_y[LS][LS + 1], s_w[LW * LW];
    __shared__ double s_red[LT * LT / 32];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int ox = blockIdx.x * LT, oy = blockIdx.y * LT;
    if (threadIdx.x < LW * LW) {
        const int dx = (int)threadIdx.x % LW - LR, dy = (int)threadIdx.x / LW - LR;
        s_w[threadIdx.x] = expf(-(float)(dx * dx + dy * dy) / (2.0f * 1.5f * 1.5f));     // :385-386, :397-398
    }
    for (int i = threadIdx.x; i < LS * LS; i += LT * LT) {
        const int lx = i % LS, ly = i / LS;
        const int px = min(max(ox + lx - LR, 0), (int)W - 1), py = min(max(oy + ly - LR, 0), (int)H - 1);   // :391-393
        float4 r = rendered[(size_t)py * W + px];
        const float4 g = gt[(size_t)py * W + px];
        if (quantize) {                                            // the reference's render target is RGBA8Unorm (note G)
            r.x = floorf(gs::clampf(r.x, 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
            r.y = floorf(gs::clampf(r.y, 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
            r.z = floorf(gs::clampf(r.z, 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
        }
        s_x[ly][lx] = (r.x + r.y + r.z) / 3.0f;                    // :405-406 mean RGB
        s_y[ly][lx] = (g.x + g.y + g.z) / 3.0f;
    }
    __syncthreads();
    double mine = 0.0;
    const int gx = ox + tx, gy = oy + ty;
    if (gx < (int)W && gy < (int)H) {
        float mu_x = 0.f, mu_y = 0.f, wsum = 0.f;
        for (int dy = 0; dy < LW; dy++)
#pragma unroll
            for (int dx = 0; dx < LW; dx++) {
                const float w = s_w[dy * LW + dx];
                wsum += w;
                mu_x += w * s_x[ty + dy][tx + dx]; mu_y += w * s_y[ty + dy][tx + dx];
            }
        mu_x /= wsum; mu_y /= wsum;                                // :413-414
        float sxx = 0.f, syy = 0.f, sxy = 0.f;
        for (int dy = 0; dy < LW; dy++)
#pragma unroll
            for (int dx = 0; dx < LW; dx++) {
                const float w = s_w[dy * LW + dx];
                const float a = s_x[ty + dy][tx + dx] - mu_x, b = s_y[ty + dy][tx + dx] - mu_y;
                sxx += w * a * a; syy += w * b * b; sxy += w * a * b;   // :444-446
            }
        sxx /= wsum; syy /= wsum; sxy /= wsum;
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;       // :359-361
        const float num = (2.0f * mu_x * mu_y + C1) * (2.0f * sxy + C2);
        const float den = (mu_x * mu_x + mu_y * mu_y + C1) * (sxx + syy + C2);
        const float dssim = gs::clampf((1.0f - num / den) / 2.0f, 0.0f, 1.0f);   // :461
        float4 r = rendered[(size_t)gy * W + gx];
        const float4 g = gt[(size_t)gy * W + gx];
        if (quantize) {
            r.x = floorf(gs::clampf(r.x, 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
            r.y = floorf(gs::clampf(r.y, 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
            r.z = floorf(gs::clampf(r.z, 0.f, 1.f) * 255.0f + 0.5f) / 255.0f;
        }
        const float l1 = (fabsf(r.x - g.x) + fabsf(r.y - g.y) + fabsf(r.z - g.z)) / 3.0f;   // :335
        mine = (double)((1.0f - lambda) * l1 + lambda * dssim);    // :497
    }
    for (int d = 16; d >= 1; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < LT * LT / 32; w++) s += s_red[w];
        if (PARTIAL) total[blockIdx.y * gridDim.x + blockIdx.x] = s;
        else atomicAdd(total, s);                                  // reduceLoss: :344-358 (double, so the order does not show)
    }
}

}  // namespace

void gs_launch_loss(const float4* rendered, const float4* gt, uint32_t W, uint32_t H, float lambda, int quantize, double* total,
                    cudaStream_t st) {
    if (W == 0 || H == 0) return;
    loss_kernel<false><<<dim3((W + LT - 1) / LT, (H + LT - 1) / LT), LT * LT, 0, st>>>(rendered, gt, W, H, lambda, quantize, total);
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant (gs_loss_compute_noalloc, include/gs_abi.h).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The scratch — one double per 16x16 tile — lives in a file-scope device array, so every GPU
// that loads this library holds GS_LOSS_NOALLOC_MAX_TILES * 8 B = 256 KiB until exit.  Every call shares the array and it
// keeps the last call's contents; nothing has to start at zero because each call overwrites the ntiles slots it then
// reads.  Two calls that overlap on the GPU (different streams) race on it.
// In host code `g_loss_partials` is only a host-side stand-in for the symbol, never a device address: the address comes
// from cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_loss_partials[GS_LOSS_NOALLOC_MAX_TILES];

namespace {
// one CTA: the partial sums in a fixed order (thread t takes slots t, t + 1024, ...; then a fixed tree), so the result does
// not depend on the order in which the tiles finished; mean over the pixels, as MTLEngine::computeLoss returns it
__global__ void __launch_bounds__(1024) loss_finish_kernel(const double* __restrict__ partials, uint32_t ntiles, double inv_pixels,
                                                           float* __restrict__ out) {
    __shared__ double s_sum[1024];
    double s = 0.0;
    for (uint32_t i = threadIdx.x; i < ntiles; i += 1024u) s += partials[i];
    s_sum[threadIdx.x] = s;
    __syncthreads();
    for (uint32_t d = 512u; d >= 1u; d >>= 1) {
        if (threadIdx.x < d) s_sum[threadIdx.x] += s_sum[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = (float)(s_sum[0] * inv_pixels);
}
constexpr int MAX_DEVICES = 64;
double* g_partials_addr[MAX_DEVICES] = {};      // device address of g_loss_partials on each device, resolved on first use
}  // namespace

// Enqueues the two kernels on `st`; allocates nothing.  `device` must be the current device.  On failure returns a negative
// GS_ERR_* code and points *err at a static message.
int gs_launch_loss_noalloc(int device, const float4* rendered, const float4* gt, uint32_t W, uint32_t H, float lambda, int quantize,
                           float* d_loss_out, cudaStream_t st, const char** err) {
    if (!rendered || !gt || !d_loss_out) { *err = "bad argument: null pointer"; return -3; }
    if (W == 0 || H == 0) { *err = "bad argument: empty image"; return -3; }
    const uint32_t ntx = (W + LT - 1) / LT, nty = (H + LT - 1) / LT;
    if ((uint64_t)ntx * nty > (uint64_t)GS_LOSS_NOALLOC_MAX_TILES) {
        *err = "gs_loss_compute_noalloc: image exceeds the static scratch (GS_LOSS_NOALLOC_MAX_TILES 16x16 tiles); use gs_loss_compute";
        return -5;
    }
    if (device < 0 || device >= MAX_DEVICES) { *err = "bad argument: device ordinal beyond 63"; return -3; }
    if (!g_partials_addr[device]) {
        double* p = nullptr;
        if (cudaGetSymbolAddress((void**)&p, g_loss_partials) != cudaSuccess || !p) { *err = "cudaGetSymbolAddress(g_loss_partials) failed"; return -1; }
        g_partials_addr[device] = p;
    }
    double* partials = g_partials_addr[device];
    loss_kernel<true><<<dim3(ntx, nty), LT * LT, 0, st>>>(rendered, gt, W, H, lambda, quantize, partials);
    loss_finish_kernel<<<1, 1024, 0, st>>>(partials, ntx * nty, 1.0 / ((double)W * (double)H), d_loss_out);
    if (cudaGetLastError() != cudaSuccess) { *err = "gs_loss_compute_noalloc: kernel launch failed"; return -1; }
    return 0;
}
