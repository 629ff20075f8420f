// radix_sort.cu — K4: onesweep-style least-significant-digit radix sort of (key, u32 value) pairs.
//
// Replaces what the reference ships for the tile|depth sort — the 8-thread CPU LSD sort
// (tiled_rasterizer.mm:27-102) — and the disabled Metal path behind GPURadixSort64::sort
// (gpu_sort.mm:456-676; histogram64 + scatter64WithAtomicRank, sort.metal:222-270, whose atomic rank is not
// order preserving).  Same contract: stable ascending sort of u64 keys carrying u32 values; also
// instantiated for u32 keys (GPURadixSort32, gpu_sort.mm:196-338).
//
// Design (HBM-bound; per pass every pair is read once and written once, 12 B + 12 B for u64 keys):
//   * one up-front kernel builds the digit histograms of ALL passes in a single read of the keys;
//   * one kernel per pass: a CTA stages a tile of 384 x 12 pairs into shared memory with two TMA bulk copies
//     (cp.async.bulk + mbarrier: the whole 54 KB is in flight from one instruction, no registers), ranks them with
//     warp-level match_any (stable: warp, row, lane order), publishes its per-digit counts, resolves its
//     global digit offsets by decoupled look-back over the preceding tiles' status words, reorders the tile
//     through shared memory so global stores are digit-contiguous, and scatters keys then values;
//   * tile ids are handed out by an atomic ticket so a tile only ever waits on tiles that already started.
#include <climits>
#include <atomic>
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

// CTA shape of the onesweep passes.  Measured on B200 (tools/sort_bench.py, bench.py stage times): 384 threads x 12 items, 3 CTAs / SM
// (36 warps resident, 4608-pair tiles) beats 256 x 16 x 4 (32 warps) by 3-11 % on 2^20 .. 2^26 pairs: the passes are latency bound.
#ifndef GS_RS_THREADS
#define GS_RS_THREADS 384
#endif
#ifndef GS_RS_CTAS
#define GS_RS_CTAS 3
#endif
#ifndef GS_RS_BIG_IPT
#define GS_RS_BIG_IPT 12
#endif
constexpr int RS_THREADS = GS_RS_THREADS;   // >= 256: threads 0..255 also own one digit each
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr uint32_t ST_AGG = 0x40000000u, ST_INC = 0x80000000u, ST_VAL = 0x3FFFFFFFu;

constexpr int RS_MAX_PASSES = 11;   // 8 LSD passes + up to 3 prefix passes of the hybrid plan
struct PassList { int n; int shift[RS_MAX_PASSES]; uint32_t mask[RS_MAX_PASSES]; };

template <typename KeyT>
__global__ void __launch_bounds__(256) radix_hist_kernel(const KeyT* __restrict__ keys, uint64_t n_max,
                                                         const uint32_t* __restrict__ d_n, const __grid_constant__ PassList pl, int p0, int np,
                                                         uint32_t* __restrict__ ghist, const uint32_t* __restrict__ gate) {
    __shared__ uint32_t sh[8][256];
    if (gate && *gate == 0u) return;                              // hybrid plan succeeded: the LSD histograms are not needed
    const uint64_t n = d_n ? min((uint64_t)*d_n, n_max) : n_max;   // element count may live on the device (no host sync)
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x) (&sh[0][0])[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    // four keys in flight per thread: on a one-wave input the kernel lasts as long as a thread's chain of dependent loads
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 4 * stride) {
        KeyT k[4];
#pragma unroll
        for (int u = 0; u < 4; u++) k[u] = (i + u * stride < n) ? keys[i + u * stride] : (KeyT)0;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (i + u * stride >= n) break;
#pragma unroll
            for (int p = 0; p < 8; p++)
                if (p < np) atomicAdd(&sh[p][(uint32_t)(k[u] >> pl.shift[p]) & pl.mask[p]], 1u);   // pl starts at pass p0 (static indices: no local copy)
        }
    }
    __syncthreads();
    for (int p = 0; p < np; p++)
        for (int d = threadIdx.x; d < 256; d += blockDim.x)
            if (sh[p][d]) atomicAdd(&ghist[(p0 + p) * 256 + d], sh[p][d]);
}

// exclusive scan of one value per thread over the CTA (only the first 256 threads carry non-zero values)
__device__ __forceinline__ uint32_t block_excl_scan256(uint32_t x, uint32_t* s_w /*[8]*/) {
    uint32_t inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (gs::lane_id() >= d) inc += t; }
    if (gs::lane_id() == 31) s_w[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) if (w < (int)(threadIdx.x >> 5)) off += s_w[w];
    __syncthreads();
    return off + inc - x;
}

// ---- TMA 1-D bulk copy + mbarrier (sm_90+/sm_100a): one thread stages the whole tile, no registers involved ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}

constexpr int RS_LOOKBACK = 8;   // predecessor status words fetched per round (independent loads)
constexpr int RS_GROUP = 32;     // tiles per second-level group
constexpr int RS_GLOOK = 4;      // groups inspected per round
__device__ __forceinline__ unsigned long long ld_relaxed64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

template <typename KeyT, int RS_IPT>
struct OnesweepSmem {
    static constexpr int RS_TILE = RS_THREADS * RS_IPT;
    alignas(128) KeyT keys[RS_TILE];          // TMA-staged input keys, then the tile-sorted keys
    alignas(128) uint32_t vals[RS_TILE];      // TMA-staged input values, then the tile-sorted values
    uint16_t wh[RS_WARPS][256];               // per-warp digit counts, then per-warp offsets within the digit
    uint32_t binstart[256];                   // first slot of digit d in the tile-local sorted order
    uint32_t gbase[256];                      // global position of slot i holding digit d = gbase[d] + i
    uint32_t scan[RS_WARPS];
    uint32_t tile;
    alignas(8) uint64_t bar;
};

// RS_IPT items per thread (tile = RS_THREADS * RS_IPT pairs: 4608 for large inputs, 3072 when the whole input then fits in one
// resident wave); BALLOT selects the peer-finding primitive of the ranking step.
template <typename KeyT, int RS_IPT, bool BALLOT, bool GATED>
__global__ void __launch_bounds__(RS_THREADS, GS_RS_CTAS) onesweep_kernel(const KeyT* __restrict__ kin, const uint32_t* __restrict__ vin,
                                                                 KeyT* __restrict__ kout, uint32_t* __restrict__ vout, uint64_t n_max,
                                                                 const uint32_t* __restrict__ d_n,
                                                                 int shift, uint32_t mask, int nbits, const uint32_t* __restrict__ ghist,
                                                                 uint32_t* counter, uint32_t* status, unsigned long long* agg2,
                                                                 int aligned16, const uint32_t* __restrict__ gate) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int RS_TILE = RS_THREADS * RS_IPT;
    OnesweepSmem<KeyT, RS_IPT>& S = *reinterpret_cast<OnesweepSmem<KeyT, RS_IPT>*>(smem_raw);
    const uint64_t n = d_n ? min((uint64_t)*d_n, n_max) : n_max;

    // gate (hybrid sort): a pass of the plan that is not being run takes a tile id beyond the input and leaves through the
    // bounds check below — no extra control flow in the hot path
    if (threadIdx.x == 0) { S.tile = (GATED && *gate == 0u) ? 0x7FFFFFFFu : atomicAdd(counter, 1u); mbar_init(&S.bar, 1); }
    for (int i = threadIdx.x; i < RS_WARPS * 256 / 2; i += RS_THREADS) reinterpret_cast<uint32_t*>(&S.wh[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = S.tile;
    const uint32_t warp = threadIdx.x >> 5, lane = gs::lane_id();
    const uint64_t tile_base = (uint64_t)tile * RS_TILE;
    if (tile_base >= n) return;                              // grid was sized from a host-side bound
    const uint32_t valid = (uint32_t)min((uint64_t)RS_TILE, n - tile_base);
    const uint32_t wbase = warp * (32 * RS_IPT) + lane;    // warp-striped: item j at wbase + 32*j
    const bool bulk = aligned16 && valid == RS_TILE;
    if (bulk) {
        if (threadIdx.x == 0) {
            mbar_expect_tx(&S.bar, RS_TILE * (sizeof(KeyT) + 4));
            bulk_g2s(S.keys, kin + tile_base, RS_TILE * sizeof(KeyT), &S.bar);
            bulk_g2s(S.vals, vin + tile_base, RS_TILE * 4, &S.bar);
        }
        mbar_wait(&S.bar, 0);
    } else {   // ragged last tile / unaligned caller buffers: plain coalesced loads, padded with all-ones keys
#pragma unroll
        for (int k = 0; k < RS_IPT; k++) {
            const uint32_t i = threadIdx.x + k * RS_THREADS;
            S.keys[i] = (i < valid) ? kin[tile_base + i] : (KeyT)~(KeyT)0;
            S.vals[i] = (i < valid) ? vin[tile_base + i] : 0u;
        }
        __syncthreads();
    }
    KeyT key[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) key[j] = S.keys[wbase + 32 * j];
    // Stable ranking inside the warp's chunk.  All match_any of the chunk are issued first (they depend only on the
    // keys, so they pipeline instead of heading a 16-long dependent chain); then, row by row, every lane reads the
    // warp's running count of its digit and the highest lane of each group stores count + group size.  Shared-memory
    // accesses of one warp complete in order, so row j+1's reads see row j's stores without a shuffle.
    uint32_t rank[RS_IPT];
    uint16_t* wh = S.wh[warp];
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t mm[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) {
        const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
        if (BALLOT) {   // one ballot per digit bit: constant cost, wins when the digits of a warp row are mostly distinct
            uint32_t m = 0xFFFFFFFFu;
#pragma unroll
            for (int b = 0; b < 8; b++) {   // bits at or above the digit width are 0 in every lane: those ballots are no-ops
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, (d >> b) & 1u);
                m &= ((d >> b) & 1u) ? bal : ~bal;
            }
            mm[j] = m;
        } else mm[j] = __match_any_sync(0xFFFFFFFFu, d);   // cost grows with the distinct values: wins on runs of equal digits
    }
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) {
        const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
        const uint32_t m = mm[j];
        const uint32_t prev = wh[d];
        rank[j] = prev + __popc(m & lt);
        __syncwarp();
        if ((m >> lane) == 1u) wh[d] = (uint16_t)(prev + __popc(m));   // highest lane of its group
        __syncwarp();
    }
    __syncthreads();
    // thread d: per-warp exclusive offsets of digit d, and the tile's count
    const uint32_t d_own = threadIdx.x & 255u;
    const bool owner = threadIdx.x < 256u;
    uint32_t cnt = 0;
    if (owner) {
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) { uint32_t c = S.wh[w][d_own]; S.wh[w][d_own] = (uint16_t)cnt; cnt += c; }
        if (tile == 0) gs::st_relaxed(&status[d_own], ST_INC | cnt);
        else gs::st_relaxed(&status[(size_t)tile * 256 + d_own], ST_AGG | cnt);
        // second level: the tile adds its count to its group's aggregate; the upper bits count arrivals, so one
        // 64-bit load tells a reader both the sum and whether all RS_GROUP tiles of the group are in it
        atomicAdd(&agg2[(size_t)(tile / RS_GROUP) * 256 + d_own], (1ull << 40) | (unsigned long long)cnt);
    }
    const uint32_t binstart = block_excl_scan256(cnt, S.scan);
    const uint32_t gstart = block_excl_scan256(owner ? ghist[d_own] : 0u, S.scan);
    // Two-level decoupled look-back.  With ~600 tiles in flight a flat look-back walks hundreds of predecessors that
    // have only published their own counts (measured: the walk, not the ranking, set the pass time).  Here a tile
    // sums the counts of the earlier tiles of ITS group (< RS_GROUP words), then steps back a whole group at a time:
    // the inclusive prefix of the group's last tile if that is out already, else the group aggregate.
    uint32_t excl = 0;
    if (tile > 0 && owner) {
        bool done = false;
        const int g0 = (int)(tile / RS_GROUP);
        int t = (int)tile - 1;
        const int lo = g0 * RS_GROUP;
        while (!done && t >= lo) {
            uint32_t s[RS_LOOKBACK];
#pragma unroll
            for (int k = 0; k < RS_LOOKBACK; k++)
                s[k] = (t - k >= lo) ? gs::ld_relaxed(&status[(size_t)(t - k) * 256 + d_own]) : ST_AGG;
            int k = 0;
#pragma unroll
            for (; k < RS_LOOKBACK; k++) {
                if ((s[k] & (ST_AGG | ST_INC)) == 0u) break;           // not published yet: retry from here
                excl += s[k] & ST_VAL;
                if (s[k] & ST_INC) { done = true; break; }
            }
            t -= k;
        }
        int gg = g0 - 1;
        while (!done && gg >= 0) {
            uint32_t s[RS_GLOOK];
            unsigned long long a[RS_GLOOK];
#pragma unroll
            for (int k = 0; k < RS_GLOOK; k++) {
                const int q = gg - k;
                s[k] = (q >= 0) ? gs::ld_relaxed(&status[((size_t)q * RS_GROUP + RS_GROUP - 1) * 256 + d_own]) : ST_INC;
                a[k] = (q >= 0) ? ld_relaxed64(&agg2[(size_t)q * 256 + d_own]) : 0ull;
            }
            int k = 0;
#pragma unroll
            for (; k < RS_GLOOK; k++) {
                if (s[k] & ST_INC) { excl += s[k] & ST_VAL; done = true; break; }
                if ((a[k] >> 40) != (unsigned long long)RS_GROUP) break;   // group not complete yet: retry from here
                excl += (uint32_t)(a[k] & 0xFFFFFFFFFFull);
            }
            gg -= k;
        }
        gs::st_relaxed(&status[(size_t)tile * 256 + d_own], ST_INC | (excl + cnt));
    }
    if (owner) { S.binstart[d_own] = binstart; S.gbase[d_own] = gstart + excl - binstart; }
    // the keys are not kept in registers across the look-back (they would spill): fetch them again before the
    // barrier after which the tile is permuted in place
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) key[j] = S.keys[wbase + 32 * j];
    __syncthreads();
    // tile-local reorder: keys into their sorted slots (every thread holds its keys in registers by now)
    uint32_t slot[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) {
        const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
        slot[j] = S.binstart[d] + wh[d] + rank[j];
    }
    uint32_t v[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) { S.keys[slot[j]] = key[j]; v[j] = S.vals[wbase + 32 * j]; }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RS_IPT; j++) S.vals[slot[j]] = v[j];
    __syncthreads();
    // digit-contiguous global stores
#pragma unroll
    for (int k = 0; k < RS_IPT; k++) {
        const uint32_t i = threadIdx.x + k * RS_THREADS;
        if (i < valid) {
            const KeyT kk = S.keys[i];
            const uint32_t pos = S.gbase[(uint32_t)(kk >> shift) & mask] + i;
            kout[pos] = kk;
            vout[pos] = S.vals[i];
        }
    }
}

__global__ void depth_keys32_kernel(const float4* __restrict__ pos, float cx, float cy, float cz, uint32_t n,
                                    uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = pos[i];
    vals[i] = i;
    uint32_t key = 0xFFFFFFFFu;                                       // sort.metal:50-54, 63-67
    if (!(isnan(p.x) || isnan(p.y) || isnan(p.z) || isinf(p.x) || isinf(p.y) || isinf(p.z))) {
        const float dx = gs::fsub(p.x, cx), dy = gs::fsub(p.y, cy), dz = gs::fsub(p.z, cz);
        const float depth = gs::fadd(gs::fadd(gs::fmul(dx, dx), gs::fmul(dy, dy)), gs::fmul(dz, dz));   // :58-60
        if (!(isnan(depth) || isinf(depth) || depth < 0.0f)) {
            const uint32_t u = __float_as_uint(depth);
            const uint32_t m = (uint32_t)(-(int32_t)(u >> 31)) | 0x80000000u;   // :28-33
            key = ~(u ^ m);                                           // :70
        }
    }
    keys[i] = key;
}


// ------------------------------------------------------------------------------------------------------------------
// Hybrid plan for full 64-bit sorts (GPURadixSort64 on caller keys): the onesweep passes above are bound by per-SM work
// (ranking, shared-memory reorder, look-back), not by HBM, so the way to go faster is FEWER passes over the data.
//   1. two (n <= 2^25) or three stable onesweep passes over the TOP T bits only -> the data is grouped into 2^T
//      segments (avg ~256 elements), each holding its elements in input order;
//   2. ONE local pass: a CTA takes the segments that start in its 2048-element chunk, bins every element by its next B
//      key bits (avg 1-2 elements per bin, shared-memory counters), and finishes each bin with a tiny insertion sort on
//      (full key, original position) — stable, in place.
// That is 3-4 trips through HBM instead of 9.  It only works when the keys spread over the top bits; everything else is
// caught on the device without a host round trip: a pre-check on the top digit's histogram, and in the local pass a
// segment that does not fit, a bin index out of range or a bin with more than LS_KMAX elements raise a flag, after
// which the plain 8-pass LSD plan runs from the untouched input (its launches return at once when the flag is clear).
// ------------------------------------------------------------------------------------------------------------------
constexpr int LS_THREADS = 256, LS_IPT = 12, LS_LOAD = LS_THREADS * LS_IPT, LS_CH = 2048, LS_NBINS = 4096, LS_KMAX = 24;   // window = chunk + 1024: the longest segment that fits
enum { RS_FLAG_HYBRID = 12, RS_FLAG_FAIL = 13, RS_FLAG_LSD = 14 };   // words of the counter block

__global__ void hybrid_precheck_kernel(const uint32_t* __restrict__ top_hist, int bins, uint64_t n, uint32_t* __restrict__ ctr) {
    // keys must spread over the most significant prefix digit: no value of it may hold more than 4x its fair share
    __shared__ uint32_t s_max;
    if (threadIdx.x == 0) s_max = 0;
    __syncthreads();
    uint32_t m = 0;
    for (int d = threadIdx.x; d < bins; d += blockDim.x) m = max(m, top_hist[d]);
    atomicMax(&s_max, m);
    __syncthreads();
    if (threadIdx.x == 0) {
        const bool ok = (uint64_t)s_max <= 4ull * (n / (uint64_t)bins) + 1024ull;
        ctr[RS_FLAG_HYBRID] = ok ? 1u : 0u;
        ctr[RS_FLAG_LSD] = ok ? 0u : 1u;
    }
}
__global__ void hybrid_decide_kernel(uint32_t* __restrict__ ctr) {
    if (threadIdx.x == 0) ctr[RS_FLAG_LSD] = (ctr[RS_FLAG_HYBRID] == 0u || ctr[RS_FLAG_FAIL] != 0u) ? 1u : 0u;
}

__device__ __forceinline__ uint32_t ls_pad(uint32_t bin) { return bin + (bin >> 4); }   // thread-contiguous runs of 16 bins without bank conflicts
struct LocalSortSmem {
    uint64_t keys[LS_LOAD];
    uint32_t vals[LS_LOAD];
    uint32_t cnt[LS_NBINS + LS_NBINS / 16 + 2];   // per-bin counts, then exclusive starts; index through ls_pad (one pad word per 16 bins)
    uint16_t pos[LS_LOAD];           // original position inside the region (the stable tie-break)
    uint32_t scan[LS_THREADS / 32];
    int first, last_end;
    uint32_t fail;
    uint64_t prev_key;
};

__global__ void __launch_bounds__(LS_THREADS, 3) local_sort_kernel(uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint64_t n,
                                                                   int T, int B, uint32_t* __restrict__ ctr) {
    extern __shared__ __align__(16) unsigned char ls_raw[];
    LocalSortSmem& S = *reinterpret_cast<LocalSortSmem*>(ls_raw);
    if (ctr[RS_FLAG_HYBRID] == 0u) return;
    const uint64_t base = (uint64_t)blockIdx.x * LS_CH;
    if (base >= n) return;
    const int avail = (int)min((uint64_t)LS_LOAD, n - base);
    const int tsh = 64 - T, rsh = 64 - T - B;
    if (threadIdx.x == 0) { S.first = INT_MAX; S.last_end = INT_MAX; S.fail = 0u; S.prev_key = base ? keys[base - 1] : 0ull; }
    for (int i = threadIdx.x; i < LS_NBINS + LS_NBINS / 16 + 2; i += LS_THREADS) S.cnt[i] = 0u;
#pragma unroll
    for (int k = 0; k < LS_IPT; k++) {
        const int i = threadIdx.x + k * LS_THREADS;
        if (i < avail) { S.keys[i] = keys[base + i]; S.vals[i] = vals[base + i]; }
    }
    __syncthreads();
    // segment boundaries: the first one inside the chunk starts the region this CTA owns, the first one at or after the
    // chunk's end closes it (a segment belongs to the CTA in whose chunk it starts)
    int my_first = INT_MAX, my_end = INT_MAX;
#pragma unroll
    for (int k = 0; k < LS_IPT; k++) {
        const int i = threadIdx.x + k * LS_THREADS;
        if (i < avail) {
            const bool boundary = (i == 0) ? (base == 0 || (S.keys[0] >> tsh) != (S.prev_key >> tsh)) : ((S.keys[i] >> tsh) != (S.keys[i - 1] >> tsh));
            if (boundary) { if (i < LS_CH) my_first = min(my_first, i); else my_end = min(my_end, i); }
        }
    }
    if (my_first != INT_MAX) atomicMin(&S.first, my_first);
    if (my_end != INT_MAX) atomicMin(&S.last_end, my_end);
    __syncthreads();
    const int first = S.first;
    if (first == INT_MAX) return;                                   // no segment starts here (it belongs to an earlier CTA)
    int le = S.last_end;
    if (le == INT_MAX) {
        if (base + (uint64_t)avail == n) le = avail;               // the data ends inside the loaded window
        else { if (threadIdx.x == 0) atomicExch(&ctr[RS_FLAG_FAIL], 1u); return; }   // a segment longer than the window
    }
    const uint64_t bin0 = (S.keys[first] >> tsh) << B;
    // bin every owned element by its top T + B bits (relative to the region's first segment); rank = arrival order
    uint64_t kk[LS_IPT]; uint32_t vv[LS_IPT]; uint32_t br[LS_IPT];
#pragma unroll
    for (int k = 0; k < LS_IPT; k++) {
        const int i = threadIdx.x + k * LS_THREADS;
        br[k] = 0xFFFFFFFFu;
        if (i >= first && i < le) {
            kk[k] = S.keys[i]; vv[k] = S.vals[i];
            const uint64_t bin = (kk[k] >> rsh) - bin0;
            if (bin >= (uint64_t)LS_NBINS) S.fail = 1u;
            else br[k] = (uint32_t)bin | (atomicAdd(&S.cnt[ls_pad((uint32_t)bin)], 1u) << 12);
        }
    }
    __syncthreads();
    if (S.fail) { if (threadIdx.x == 0) atomicExch(&ctr[RS_FLAG_FAIL], 1u); return; }
    // exclusive scan of the bin counts (16 consecutive bins per thread); a crowded bin means the keys do not spread: give up
    uint32_t c[LS_NBINS / LS_THREADS], tot = 0;
    bool crowded = false;
#pragma unroll
    for (int k = 0; k < LS_NBINS / LS_THREADS; k++) { c[k] = S.cnt[ls_pad(threadIdx.x * (LS_NBINS / LS_THREADS) + k)]; tot += c[k]; crowded |= c[k] > (uint32_t)LS_KMAX; }
    if (crowded) S.fail = 1u;
    uint32_t inc = tot;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (gs::lane_id() >= (uint32_t)d) inc += t; }
    if (gs::lane_id() == 31u) S.scan[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (S.fail) { if (threadIdx.x == 0) atomicExch(&ctr[RS_FLAG_FAIL], 1u); return; }
    uint32_t off = inc - tot;
#pragma unroll
    for (int w = 0; w < LS_THREADS / 32; w++) if (w < (int)(threadIdx.x >> 5)) off += S.scan[w];
#pragma unroll
    for (int k = 0; k < LS_NBINS / LS_THREADS; k++) { S.cnt[ls_pad(threadIdx.x * (LS_NBINS / LS_THREADS) + k)] = off; off += c[k]; }
    __syncthreads();
    // scatter into bin order (every thread holds its elements in registers, the staging area is free to overwrite)
#pragma unroll
    for (int k = 0; k < LS_IPT; k++) {
        if (br[k] != 0xFFFFFFFFu) {
            const int i = threadIdx.x + k * LS_THREADS;
            const uint32_t dst = (uint32_t)first + S.cnt[ls_pad(br[k] & 0xFFFu)] + (br[k] >> 12);
            S.keys[dst] = kk[k]; S.vals[dst] = vv[k]; S.pos[dst] = (uint16_t)i;
        }
    }
    __syncthreads();
    // finish each bin by counting: an element's place inside its bin is the number of bin mates that order before it on
    // (key, original position).  Bins hold 1-2 elements on average (LS_KMAX at most), every lane works on its own element
    // (no serial per-bin loop), and neighbouring lanes read neighbouring shared-memory words.
    const uint32_t total = (uint32_t)(le - first);
    uint32_t dstp[LS_IPT];
#pragma unroll
    for (int k = 0; k < LS_IPT; k++) {
        const uint32_t e = threadIdx.x + (uint32_t)k * LS_THREADS;
        dstp[k] = 0xFFFFFFFFu;
        if (e < total) {
            const uint32_t i = (uint32_t)first + e;
            const uint64_t xk = S.keys[i];
            const uint32_t b = (uint32_t)((xk >> rsh) - bin0);
            const uint32_t s0 = S.cnt[ls_pad(b)], e0 = (b + 1 < (uint32_t)LS_NBINS) ? S.cnt[ls_pad(b + 1)] : total;
            uint32_t r = e - s0;                                   // alone in its bin: it stays where it is
            if (e0 - s0 > 1u) {
                const uint16_t xp = S.pos[i];
                r = 0;
                for (uint32_t j = s0; j < e0; j++) {
                    const uint64_t yk = S.keys[first + j];
                    r += (yk < xk || (yk == xk && S.pos[first + j] < xp)) ? 1u : 0u;
                }
            }
            kk[k] = xk; vv[k] = S.vals[i]; dstp[k] = (uint32_t)first + s0 + r;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < LS_IPT; k++)
        if (dstp[k] != 0xFFFFFFFFu) { S.keys[dstp[k]] = kk[k]; S.vals[dstp[k]] = vv[k]; }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < LS_IPT; k++) {
        const int i = threadIdx.x + k * LS_THREADS;
        if (i >= first && i < le) { keys[base + i] = S.keys[i]; vals[base + i] = S.vals[i]; }
    }
}

constexpr uint32_t RS_BIG_TILE = RS_THREADS * GS_RS_BIG_IPT, RS_SMALL_TILE = RS_THREADS * 8, RS_TINY_TILE = RS_THREADS * 4;
constexpr size_t RS_HIST_BYTES = RS_MAX_PASSES * 256 * 4;
constexpr size_t RS_CTR_BYTES = 64;

// bit of the current device in the per-device "attribute already set" masks (an ordinal beyond 63 has no bit: it sets the
// attribute on every launch, which is cheap and harmless)
static inline uint64_t rs_device_bit() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < 64) ? (1ull << dev) : 0ull;
}

template <typename KeyT, int IPT, bool BALLOT>
void launch_passes(const KeyT* keys_in, const uint32_t* vals_in, KeyT* kbuf[2], uint32_t* vbuf[2], uint64_t n, const uint32_t* d_n,
                   const GsSortPlan& plan, const PassList& pl, uint32_t tiles, uint32_t* ghist, uint32_t* counters,
                   uint32_t* status, unsigned long long* agg2, size_t groups, cudaStream_t st, const uint32_t* gate = nullptr,
                   int slot0 = 0, int dst0 = 0) {
    const size_t smem = sizeof(OnesweepSmem<KeyT, IPT>);
    // cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute: one bit per device ordinal, per instantiation
    static std::atomic<uint64_t> attr_set{0};
    const uint64_t dev_bit = rs_device_bit();
    if (!(attr_set.load(std::memory_order_acquire) & dev_bit)) {
        cudaFuncSetAttribute(onesweep_kernel<KeyT, IPT, BALLOT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(onesweep_kernel<KeyT, IPT, BALLOT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set.fetch_or(dev_bit, std::memory_order_release);        // set after the calls: a racing thread repeats them (idempotent)
    }
    const KeyT* kin = keys_in;
    const uint32_t* vin = vals_in;
    int dst = dst0;
    for (int p = 0; p < plan.num_passes; p++) {
        const int aligned16 = (((uintptr_t)kin | (uintptr_t)vin) & 15) == 0;
        const int q = slot0 + p;                                   // scratch slot (histogram row, ticket, status block) of this pass
        if (gate)
            onesweep_kernel<KeyT, IPT, BALLOT, true><<<tiles, RS_THREADS, smem, st>>>(
                kin, vin, kbuf[dst], vbuf[dst], n, d_n, pl.shift[q], pl.mask[q], plan.bits[p], ghist + q * 256, counters + q,
                status + (size_t)q * tiles * 256, agg2 + (size_t)q * groups * 256, aligned16, gate);
        else
            onesweep_kernel<KeyT, IPT, BALLOT, false><<<tiles, RS_THREADS, smem, st>>>(
                kin, vin, kbuf[dst], vbuf[dst], n, d_n, pl.shift[q], pl.mask[q], plan.bits[p], ghist + q * 256, counters + q,
                status + (size_t)q * tiles * 256, agg2 + (size_t)q * groups * 256, aligned16, gate);
        kin = kbuf[dst]; vin = vbuf[dst];
        dst ^= 1;
    }
}

static int g_rs_tiny = 0;   // experiment knob: 1024-pair tiles for inputs around 1 M

static inline PassList pass_window(const PassList& pl, int p0, int np) {
    PassList w; w.n = np;
    for (int p = 0; p < RS_MAX_PASSES; p++) { const bool in = p < np; w.shift[p] = in ? pl.shift[p0 + p] : 0; w.mask[p] = in ? pl.mask[p0 + p] : 0u; }
    return w;
}
static int g_rs_hybrid = 1;   // knob: 0 = always the plain LSD plan

static inline int ceil_log2_u64(uint64_t n) { int l = 0; while ((1ull << l) < n) l++; return l; }

static inline bool rs_small(uint64_t n) { return n <= (uint64_t)RS_SMALL_TILE * 148 * GS_RS_CTAS; }
static inline bool rs_tiny(uint64_t n) { return g_rs_tiny && n <= (uint64_t)RS_TINY_TILE * 148 * GS_RS_CTAS * 2; }
static int g_rs_even = 1;   // knob: 0 = 3072-pair tiles for every one-wave input
// Items per thread.  An input whose tiles are all resident at once (one wave) takes as long as the SM that holds the most
// elements: among the compiled tile sizes take the one with the smallest (CTAs on the fullest SM) x (tile size), e.g.
// 1 M pairs: 326 tiles of 3072 put three CTAs on 30 SMs (9216 elements), 290 tiles of 3456 put two on each (6912).
static inline int rs_ipt(uint64_t n) {
    if (rs_tiny(n)) return 4;
    if (!rs_small(n)) return GS_RS_BIG_IPT;
    int best = 8;
    if (g_rs_even) {
        uint64_t best_cost = ~0ull;
        for (int ipt : {8, 9, 10, 12}) {     // measured: 435 tiles of 2304 (three per SM, same product) are slower: 79 against 72 us
            const uint64_t tiles = (n + (uint64_t)RS_THREADS * ipt - 1) / ((uint64_t)RS_THREADS * ipt);
            if (tiles > 148ull * GS_RS_CTAS) continue;
            const uint64_t cost = ((tiles + 147) / 148) * ipt;
            if (cost < best_cost) { best_cost = cost; best = ipt; }
        }
    }
    return best;
}
static inline uint32_t rs_tile_elems(uint64_t n) { return (uint32_t)RS_THREADS * (uint32_t)rs_ipt(n); }

template <typename KeyT>
int sort_impl(const KeyT* keys_in, const uint32_t* vals_in, KeyT* kbuf[2], uint32_t* vbuf[2], uint64_t n,
              const uint32_t* d_n, const GsSortPlan& plan, int hints, void* scratch, cudaStream_t st) {
    if (n == 0) return 0;
    if (plan.num_passes == 0) {
        cudaMemcpyAsync(kbuf[0], keys_in, n * sizeof(KeyT), cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(vbuf[0], vals_in, n * 4, cudaMemcpyDeviceToDevice, st);
        return 0;
    }
    // small inputs: 2048-pair tiles so that every tile is resident at once (one wave) instead of 1.x waves of 4096
    const int ipt = rs_ipt(n);
    const uint32_t tile_elems = rs_tile_elems(n);
    const uint32_t tiles = (uint32_t)((n + tile_elems - 1) / tile_elems);
    const size_t status_bytes = (size_t)tiles * 256 * 4;
    uint32_t* ghist = (uint32_t*)scratch;
    uint32_t* counters = (uint32_t*)((char*)scratch + RS_HIST_BYTES);
    uint32_t* status = (uint32_t*)((char*)scratch + RS_HIST_BYTES + RS_CTR_BYTES);
    const size_t groups = (tiles + RS_GROUP - 1) / RS_GROUP;
    const size_t agg_bytes = groups * 256 * 8;
    // hybrid plan: full 64-bit sorts of caller keys (all eight 8-bit digits), element count known on the host, 2^20 .. 2^28
    bool hybrid = g_rs_hybrid && sizeof(KeyT) == 8 && d_n == nullptr && plan.num_passes == 8 && n >= (1ull << 20) && n <= (1ull << 28);
    for (int p = 0; hybrid && p < 8; p++) hybrid = plan.shifts[p] == 8 * p && plan.bits[p] == 8;
    int T = 0, B = 0;
    GsSortPlan pre; pre.num_passes = 0;
    if (hybrid) {
        const int lg = ceil_log2_u64(n);
        T = lg <= 25 ? (lg - 8 < 12 ? 12 : (lg - 8 > 16 ? 16 : lg - 8)) : lg - 8;      // avg segment 256 .. 512 elements
        B = lg - T - 1; if (B < 2) B = 2;                                            // avg 1 .. 2 elements per bin
        pre = gs_sort_plan_even(64 - T, 64);
    }
    const int slots = plan.num_passes + pre.num_passes;
    unsigned long long* agg2 = (unsigned long long*)((char*)status + status_bytes * slots);
    // GS_SORT_HINT_PREHIST: the caller cleared the scratch (gs_sort_prepare) and filled the digit histograms itself
    const bool prehist = (hints & GS_SORT_HINT_PREHIST) != 0 && !hybrid;
    if (!prehist) cudaMemsetAsync(scratch, 0, RS_HIST_BYTES + RS_CTR_BYTES + (status_bytes + agg_bytes) * slots, st);
    PassList pl;
    pl.n = slots;
    for (int p = 0; p < RS_MAX_PASSES; p++) {
        if (p < plan.num_passes) { pl.shift[p] = plan.shifts[p]; pl.mask[p] = (1u << plan.bits[p]) - 1u; }
        else if (p < slots) { pl.shift[p] = pre.shifts[p - plan.num_passes]; pl.mask[p] = (1u << pre.bits[p - plan.num_passes]) - 1u; }
        else { pl.shift[p] = 0; pl.mask[p] = 0u; }
    }
    uint64_t hb = (n + 2047) / 2048; int hist_blocks = (int)(hb < 148 * 8 ? hb : 148 * 8);
    const bool ballot = (hints & GS_SORT_HINT_DISTINCT) != 0;
#define GS_RS_GO(IPT, B_, PLAN, GATE, SLOT0, DST0) launch_passes<KeyT, IPT, B_>(keys_in, vals_in, kbuf, vbuf, n, d_n, PLAN, pl, tiles, ghist, counters, status, agg2, groups, st, GATE, SLOT0, DST0)
#define GS_RS_RUN(PLAN, GATE, SLOT0, DST0)                                                                    \
    do {                                                                                                      \
        if (ipt == 4) { if (ballot) GS_RS_GO(4, true, PLAN, GATE, SLOT0, DST0); else GS_RS_GO(4, false, PLAN, GATE, SLOT0, DST0); }      \
        else if (ipt == 8) { if (ballot) GS_RS_GO(8, true, PLAN, GATE, SLOT0, DST0); else GS_RS_GO(8, false, PLAN, GATE, SLOT0, DST0); } \
        else if (ipt == 9) { if (ballot) GS_RS_GO(9, true, PLAN, GATE, SLOT0, DST0); else GS_RS_GO(9, false, PLAN, GATE, SLOT0, DST0); } \
        else if (ipt == 10) { if (ballot) GS_RS_GO(10, true, PLAN, GATE, SLOT0, DST0); else GS_RS_GO(10, false, PLAN, GATE, SLOT0, DST0); } \
        else { if (ballot) GS_RS_GO(GS_RS_BIG_IPT, true, PLAN, GATE, SLOT0, DST0); else GS_RS_GO(GS_RS_BIG_IPT, false, PLAN, GATE, SLOT0, DST0); } \
    } while (0)
    if (!hybrid) {
        if (!prehist) radix_hist_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys_in, n, d_n, pass_window(pl, 0, plan.num_passes), 0, plan.num_passes, ghist, nullptr);
        GS_RS_RUN(plan, nullptr, 0, 0);
        return (plan.num_passes & 1) ? 0 : 1;
    }
    // ---- hybrid: prefix passes end in buffer 1 (2 passes: 0 then 1; 3 passes: start at 1 -> 1, 0, 1), the local pass works in
    // place there, and the 8-pass LSD fallback ends in buffer 1 as well, so the caller's result index does not depend on the flags
    radix_hist_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys_in, n, d_n, pass_window(pl, plan.num_passes, pre.num_passes), plan.num_passes, pre.num_passes, ghist, nullptr);
    const int top = plan.num_passes + pre.num_passes - 1;
    hybrid_precheck_kernel<<<1, 256, 0, st>>>(ghist + top * 256, 1 << pre.bits[pre.num_passes - 1], n, counters);
    GS_RS_RUN(pre, nullptr, plan.num_passes, (pre.num_passes & 1) ? 1 : 0);   // ungated (a gated CTA pays one more global round trip before its ticket)
    if constexpr (sizeof(KeyT) == 8) {
        static std::atomic<uint64_t> ls_attr{0};                      // per device, as above
        const uint64_t dev_bit = rs_device_bit();
        if (!(ls_attr.load(std::memory_order_acquire) & dev_bit)) {
            cudaFuncSetAttribute(local_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LocalSortSmem));
            ls_attr.fetch_or(dev_bit, std::memory_order_release);
        }
        const uint32_t chunks = (uint32_t)((n + LS_CH - 1) / LS_CH);
        local_sort_kernel<<<chunks, LS_THREADS, sizeof(LocalSortSmem), st>>>(reinterpret_cast<uint64_t*>(kbuf[1]), vbuf[1], n, T, B, counters);
    }
    hybrid_decide_kernel<<<1, 32, 0, st>>>(counters);
    radix_hist_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys_in, n, d_n, pass_window(pl, 0, plan.num_passes), 0, plan.num_passes, ghist, counters + RS_FLAG_LSD);
    GS_RS_RUN(plan, counters + RS_FLAG_LSD, 0, 0);
#undef GS_RS_RUN
#undef GS_RS_GO
    return 1;
}

}  // namespace

void gs_tune_sort_tiny(int v) { g_rs_tiny = v ? 1 : 0; }
void gs_tune_sort_hybrid(int v) { g_rs_hybrid = v ? 1 : 0; }
void gs_tune_sort_even(int v) { g_rs_even = v ? 1 : 0; }
const uint32_t* gs_sort_flags(const void* scratch) { return (const uint32_t*)((const char*)scratch + RS_HIST_BYTES) + RS_FLAG_HYBRID; }
const uint32_t* gs_sort_pass_histogram(const void* scratch, int pass) { return (const uint32_t*)scratch + (size_t)pass * 256; }

GsSortPlan gs_sort_plan(int begin_bit, int end_bit, uint64_t varying_mask) {
    GsSortPlan pl;
    pl.num_passes = 0;
    for (int s = begin_bit; s < end_bit; s += 8) {
        int bits = (end_bit - s) < 8 ? (end_bit - s) : 8;
        uint64_t m = ((bits == 64 ? ~0ull : ((1ull << bits) - 1ull)) << s);
        if ((varying_mask & m) == 0) continue;
        pl.shifts[pl.num_passes] = s; pl.bits[pl.num_passes] = bits; pl.num_passes++;
    }
    return pl;
}

size_t gs_sort_scratch_bytes(uint64_t n, int /*key_bytes*/) {
    size_t tiles = (size_t)((n + RS_TINY_TILE - 1) / RS_TINY_TILE) + 1;
    return RS_HIST_BYTES + RS_CTR_BYTES + (tiles * 256 * 4 + (tiles / RS_GROUP + 1) * 256 * 8) * RS_MAX_PASSES;
}

// Clears the scratch of a coming gs_launch_sort32(..., GS_SORT_HINT_PREHIST) over at most n elements and returns the
// histogram block: the caller's producer kernel adds the count of digit d of pass p (plan order) to hist[p * 256 + d].
uint32_t* gs_sort_prepare(uint64_t n, const GsSortPlan& plan, void* scratch, cudaStream_t st) {
    if (n == 0 || plan.num_passes == 0) return (uint32_t*)scratch;
    const uint32_t tile_elems = rs_tile_elems(n);
    const size_t tiles = (size_t)((n + tile_elems - 1) / tile_elems), groups = (tiles + RS_GROUP - 1) / RS_GROUP;
    cudaMemsetAsync(scratch, 0, RS_HIST_BYTES + RS_CTR_BYTES + (tiles * 256 * 4 + groups * 256 * 8) * plan.num_passes, st);   // as sort_impl
    return (uint32_t*)scratch;
}

int gs_launch_sort64(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* kbuf[2], uint32_t* vbuf[2], uint64_t n,
                     const GsSortPlan& plan, int hints, void* scratch, cudaStream_t st) {
    return sort_impl<uint64_t>(keys_in, vals_in, kbuf, vbuf, n, nullptr, plan, hints, scratch, st);
}
int gs_launch_sort32(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* kbuf[2], uint32_t* vbuf[2], uint64_t n,
                     const uint32_t* d_n, const GsSortPlan& plan, int hints, void* scratch, cudaStream_t st) {
    return sort_impl<uint32_t>(keys_in, vals_in, kbuf, vbuf, n, d_n, plan, hints, scratch, st);
}
GsSortPlan gs_sort_plan_even(int begin_bit, int end_bit) {
    GsSortPlan pl;
    const int bits = end_bit - begin_bit;
    pl.num_passes = (bits + 7) / 8;
    int s = begin_bit;
    for (int p = 0; p < pl.num_passes; p++) {
        const int b = (bits - (s - begin_bit) + (pl.num_passes - p) - 1) / (pl.num_passes - p);
        pl.shifts[p] = s; pl.bits[p] = b; s += b;
    }
    return pl;
}
void gs_launch_depth_keys32(const float4* positions, float cx, float cy, float cz, uint32_t n, uint32_t* keys,
                            uint32_t* vals, cudaStream_t st) {
    if (n == 0) return;
    depth_keys32_kernel<<<(n + 255) / 256, 256, 0, st>>>(positions, cx, cy, cz, n, keys, vals);
}
