// pairs.cu — K2: single-pass exclusive scan of the per-Gaussian tile counts (decoupled look-back), and
// K3: key duplication (replaces generateTilePairs' global atomic, tiled_shaders.metal:776-793, with scanned
// offsets so the pair order is deterministic: Gaussian index, then ty-major / tx-minor — SURVEY note K).
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_IPT = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;
constexpr unsigned long long FLAG_AGG = 1ull << 32, FLAG_INC = 2ull << 32;

__device__ __forceinline__ unsigned long long ld_relaxed64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// scratch layout: [0] dynamic tile counter (u32, + pad to 8 B), then one u64 status word per tile
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(uint32_t* __restrict__ v, uint32_t n, uint32_t* __restrict__ total,
                                                            uint32_t* counter, unsigned long long* status) {
    __shared__ uint32_t s_tile, s_warp[SCAN_THREADS / 32], s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(counter, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t base = tile * SCAN_TILE + threadIdx.x * SCAN_IPT;
    uint32_t x[SCAN_IPT];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) { x[k] = (base + k < n) ? v[base + k] : 0u; sum += x[k]; }
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
    if (threadIdx.x == 0) {
        uint32_t excl = 0;
        if (tile == 0) {
            st_relaxed64(&status[0], FLAG_INC | block_sum);
        } else {
            st_relaxed64(&status[tile], FLAG_AGG | block_sum);
            int t = (int)tile - 1;
            while (true) {
                unsigned long long s;
                do { s = ld_relaxed64(&status[t]); } while ((s >> 32) == 0ull);
                excl += (uint32_t)s;
                if (s & FLAG_INC) break;
                t--;
            }
            st_relaxed64(&status[tile], FLAG_INC | (unsigned long long)(excl + block_sum));
        }
        s_prefix = excl;
        if ((uint64_t)(tile + 1) * SCAN_TILE >= n) { v[n] = excl + block_sum; *total = excl + block_sum; }
    }
    __syncthreads();
    uint32_t run = s_prefix + warp_off + (inc - sum);
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) { if (base + k < n) v[base + k] = run; run += x[k]; }
}

// K3.  One block = 256 consecutive Gaussians; the block's pairs form one contiguous output run, written
// cooperatively (coalesced 8-B / 4-B lanes); the owner of an output slot is found by binary search in smem.
__global__ void __launch_bounds__(256) emit_pairs_kernel(GsProjected O, uint32_t n, uint32_t ntx, uint64_t* __restrict__ keys,
                                                         uint32_t* __restrict__ vals) {
    __shared__ uint32_t s_off[257];
    __shared__ uint32_t s_minx[256], s_miny[256], s_tx[256], s_dkey[256];
    const uint32_t g0 = blockIdx.x * 256u;
    const uint32_t i = g0 + threadIdx.x;
    if (i < n) {
        s_off[threadIdx.x] = O.offs[i];
        const uint4 bb = O.bbox[i];
        s_minx[threadIdx.x] = bb.x; s_miny[threadIdx.x] = bb.y; s_tx[threadIdx.x] = bb.z - bb.x + 1u;
        const uint32_t k = __float_as_uint(O.rec[3 * (size_t)i + 2].y);        // depth
        s_dkey[threadIdx.x] = (k & 0x80000000u) ? ~k : (k | 0x80000000u);       // tiled_shaders.metal:773-774
    }
    const uint32_t last = min(g0 + 256u, n);
    const uint32_t cnt = last - g0;
    if (threadIdx.x == 0) s_off[cnt] = O.offs[last];          // offs[n] holds the total
    __syncthreads();
    const uint32_t begin = s_off[0], stop = s_off[cnt];
    for (uint32_t j = begin + threadIdx.x; j < stop; j += 256u) {
        // largest g in [0,cnt) with s_off[g] <= j   (offsets are non-decreasing; empty Gaussians are skipped)
        uint32_t lo = 0, hi = cnt;
        while (hi - lo > 1u) { uint32_t mid = (lo + hi) >> 1; if (s_off[mid] <= j) lo = mid; else hi = mid; }
        const uint32_t k = j - s_off[lo];
        const uint32_t tw = s_tx[lo];
        const uint32_t ty = s_miny[lo] + k / tw, tx = s_minx[lo] + k % tw;        // :784-786
        keys[j] = ((uint64_t)(ty * ntx + tx) << 32) | (uint64_t)s_dkey[lo];       // :787
        vals[j] = g0 + lo;
    }
}

}  // namespace

size_t gs_scan_scratch_bytes(uint32_t n) {
    size_t tiles = ((size_t)n + SCAN_TILE - 1) / SCAN_TILE + 1;
    return 8 + tiles * 8;
}

void gs_launch_exclusive_scan(uint32_t* v, uint32_t n, uint32_t* total, void* scratch, cudaStream_t st) {
    if (n == 0) { cudaMemsetAsync(v, 0, 4, st); cudaMemsetAsync(total, 0, 4, st); return; }
    uint32_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    cudaMemsetAsync(scratch, 0, 8 + (size_t)tiles * 8, st);
    scan_kernel<<<tiles, SCAN_THREADS, 0, st>>>(v, n, total, (uint32_t*)scratch, (unsigned long long*)((char*)scratch + 8));
}

void gs_launch_emit_pairs(const GsProjected& O, uint32_t n, uint32_t ntx, uint64_t* keys, uint32_t* vals, cudaStream_t st) {
    if (n == 0) return;
    emit_pairs_kernel<<<(n + 255) / 256, 256, 0, st>>>(O, n, ntx, keys, vals);
}
