// pairs.cu — K2: single-pass exclusive scan of the per-Gaussian tile counts (decoupled look-back), and
// K3: key duplication (replaces generateTilePairs' global atomic, tiled_shaders.metal:776-793, with scanned
// offsets so the pair order is deterministic: Gaussian index, then ty-major / tx-minor — SURVEY note K).
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

constexpr int SCAN_THREADS = 256;
#ifndef GS_SCAN_IPT
#define GS_SCAN_IPT 8
#endif
constexpr int SCAN_IPT = GS_SCAN_IPT;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;


// scratch layout: [0] dynamic tile counter (u32, + pad to 8 B), one u64 status word per tile, one u64 word per 32 tiles
// out[j] = sum_{j'<j} in[perm ? perm[j'] : j'] ; out[n] = *total = the sum.  in == out (perm == nullptr) is allowed.
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const uint32_t* in, const uint32_t* __restrict__ perm, uint32_t* out,
                                                            uint32_t n, uint32_t* __restrict__ total,
                                                            uint32_t* counter, unsigned long long* status, unsigned long long* group_sum) {
    __shared__ uint32_t s_tile, s_warp[SCAN_THREADS / 32], s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(counter, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t base = tile * SCAN_TILE + threadIdx.x * SCAN_IPT;
    uint32_t x[SCAN_IPT];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) { x[k] = (base + k < n) ? in[perm ? perm[base + k] : base + k] : 0u; sum += x[k]; }
    // block exclusive scan of per-thread sums
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (gs::lane_id() >= d) inc += t; }
    const uint32_t warp = threadIdx.x >> 5;
    if (gs::lane_id() == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t warp_off = 0, block_sum = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) { if (w < (int)warp) warp_off += s_warp[w]; block_sum += s_warp[w]; }
    if (warp == 0) {
        const uint32_t excl = gs::lookback_excl_warp(tile, block_sum, status, group_sum, gs::lane_id());
        if (gs::lane_id() == 0) {
            s_prefix = excl;
            if ((uint64_t)(tile + 1) * SCAN_TILE >= n) { out[n] = excl + block_sum; *total = excl + block_sum; }
        }
    }
    __syncthreads();
    uint32_t run = s_prefix + warp_off + (inc - sum);
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) { if (base + k < n) out[base + k] = run; run += x[k]; }
}

// Exact-conservative footprint test (not in the reference): can Gaussian g contribute to ANY pixel centre of tile
// (tx,ty)?  tiledForward/tiledBackward drop an entry when power < -4.5 or alpha = sigma*exp(power) < 1/255
// (tiled_shaders.metal:359-366,489-495), i.e. when q = A dx^2 + 2B dx dy + C dy^2 = -2 power exceeds
// qthr = min(9, 2 ln(255 sigma)).  q is a convex quadratic centred on the splat, so its minimum over the tile's
// pixel-centre box lies on an edge FACING the centre (<= 2 edges, each a clamped 1-D minimum).  A pair whose minimum
// exceeds qthr by more than the rounding noise of the pinned power arithmetic can never contribute: it stays in the
// list (so lists, ranges and lastContribIdx remain those of the reference) but is emitted with bit 31 of its value
// set, and the blend kernels drop it while staging.
// (CullRec / cull_rec / box_cannot_contribute live in gs_common.cuh: the blend kernels reuse the test per warp region)
__device__ __forceinline__ bool tile_cannot_contribute(const gs::CullRec& c, uint32_t tx, uint32_t ty) {
    return gs::box_cannot_contribute(c, (float)(tx * GS_TILE) + 0.5f, (float)(ty * GS_TILE) + 0.5f, 15.0f, 15.0f);
}

// K3.  One block = 256 consecutive list positions (Gaussian order, or depth order through `perm`); the block's pairs
// form one contiguous output run, written cooperatively (coalesced lanes); the owner of an output slot is found by
// binary search in smem.  KeyT = u64: (tile<<32 | depthKey) keys of the reference; KeyT = u32: the tile id alone
// (hoisted-depth mode, see api.cu).
template <typename KeyT>
__global__ void __launch_bounds__(256) emit_pairs_kernel(GsProjected O, const uint32_t* __restrict__ perm,
                                                         const uint32_t* __restrict__ offs, uint32_t n, uint32_t ntx,
                                                         KeyT* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t cap_pairs) {
    __shared__ uint32_t s_off[512];                           // [cnt+1 .. 511] = 0xFFFFFFFF: the search needs no bound test
    __shared__ uint32_t s_minx[256], s_miny[256], s_tx[256], s_dkey[256], s_gid[256];
    __shared__ float s_rtx[256];
    __shared__ uint32_t s_end;
    __shared__ gs::CullRec s_cull[256];
    const uint32_t g0 = blockIdx.x * 256u;
    const uint32_t i = g0 + threadIdx.x;
    const uint32_t last = min(g0 + 256u, n);
    const uint32_t cnt = last - g0;
    s_off[256 + threadIdx.x] = 0xFFFFFFFFu;
    if (i < n) {
        const uint32_t g = perm ? perm[i] : i;
        s_gid[threadIdx.x] = g;
        s_off[threadIdx.x] = offs[i];
        const uint4 bb = O.bbox[g];
        const uint32_t tw = bb.z - bb.x + 1u;
        s_minx[threadIdx.x] = bb.x; s_miny[threadIdx.x] = bb.y; s_tx[threadIdx.x] = tw;
        s_rtx[threadIdx.x] = 1.0f / (float)tw;
        const float4 r0 = O.rec[3 * (size_t)g], r1 = O.rec[3 * (size_t)g + 1];
        const uint32_t k = __float_as_uint(O.rec[3 * (size_t)g + 2].y);        // depth
        s_dkey[threadIdx.x] = (k & 0x80000000u) ? ~k : (k | 0x80000000u);       // tiled_shaders.metal:773-774
        s_cull[threadIdx.x] = gs::cull_rec(r0, r1);
    } else s_off[threadIdx.x] = 0xFFFFFFFFu;
    if (threadIdx.x == 0) s_end = offs[last];                 // offs[n] holds the total
    __syncthreads();
    const uint32_t begin = s_off[0], stop = min(s_end, cap_pairs);
#pragma unroll 2
    for (uint32_t j = begin + threadIdx.x; j < stop; j += 256u) {
        // largest g in [0,cnt) with s_off[g] <= j   (offsets are non-decreasing; empty Gaussians are skipped)
        uint32_t lo = 0;
#pragma unroll
        for (uint32_t sft = 128u; sft >= 1u; sft >>= 1) { if (s_off[lo + sft] <= j) lo += sft; }
        const uint32_t k = j - s_off[lo];                     // < GS_MAX_TILES_PER_GAUSSIAN (256)
        const uint32_t tw = s_tx[lo];
        // k / tw for k, tw <= 256: (k + 0.5) / tw is at least 1/512 away from an integer, so the float product truncates exactly
        const uint32_t row = (uint32_t)(((float)k + 0.5f) * s_rtx[lo]);
        const uint32_t ty = s_miny[lo] + row, tx = s_minx[lo] + (k - row * tw);   // :784-786
        if (sizeof(KeyT) == 8) keys[j] = (KeyT)(((uint64_t)(ty * ntx + tx) << 32) | (uint64_t)s_dkey[lo]);   // :787
        else keys[j] = (KeyT)(ty * ntx + tx);
        vals[j] = s_gid[lo] | (tile_cannot_contribute(s_cull[lo], tx, ty) ? GS_CULL_BIT : 0u);
    }
}

}  // namespace

size_t gs_scan_scratch_bytes(uint32_t n) {
    size_t tiles = ((size_t)n + SCAN_TILE - 1) / SCAN_TILE + 1;
    return 8 + tiles * 8 + (tiles / 32 + 1) * 8;
}

void gs_launch_exclusive_scan(const uint32_t* in, const uint32_t* perm, uint32_t* out, uint32_t n, uint32_t* total,
                              void* scratch, cudaStream_t st) {
    if (n == 0) { cudaMemsetAsync(out, 0, 4, st); cudaMemsetAsync(total, 0, 4, st); return; }
    uint32_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    cudaMemsetAsync(scratch, 0, 8 + (size_t)tiles * 8 + ((size_t)tiles / 32 + 1) * 8, st);
    unsigned long long* status = (unsigned long long*)((char*)scratch + 8);
    scan_kernel<<<tiles, SCAN_THREADS, 0, st>>>(in, perm, out, n, total, (uint32_t*)scratch, status, status + tiles);
}

void gs_launch_emit_pairs(const GsProjected& O, const uint32_t* offs, uint32_t n, uint32_t ntx, uint64_t* keys, uint32_t* vals,
                          uint32_t cap_pairs, cudaStream_t st) {
    if (n == 0) return;
    emit_pairs_kernel<uint64_t><<<(n + 255) / 256, 256, 0, st>>>(O, nullptr, offs, n, ntx, keys, vals, cap_pairs);
}
void gs_launch_emit_tile_pairs(const GsProjected& O, const uint32_t* perm, const uint32_t* offs, uint32_t n, uint32_t ntx,
                               uint32_t* tile_keys, uint32_t* vals, uint32_t cap_pairs, cudaStream_t st) {
    if (n == 0) return;
    emit_pairs_kernel<uint32_t><<<(n + 255) / 256, 256, 0, st>>>(O, perm, offs, n, ntx, tile_keys, vals, cap_pairs);
}
