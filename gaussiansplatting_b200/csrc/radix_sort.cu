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
//   * one kernel per pass: a CTA stages a tile of 256 x 16 pairs into shared memory with two TMA bulk copies
//     (cp.async.bulk + mbarrier: the whole 48 KB is in flight from one instruction, no registers), ranks them with
//     warp-level match_any (stable: warp, row, lane order), publishes its per-digit counts, resolves its
//     global digit offsets by decoupled look-back over the preceding tiles' status words, reorders the tile
//     through shared memory so global stores are digit-contiguous, and scatters keys then values;
//   * tile ids are handed out by an atomic ticket so a tile only ever waits on tiles that already started.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

#ifndef GS_RS_THREADS
#define GS_RS_THREADS 256
#endif
#ifndef GS_RS_CTAS
#define GS_RS_CTAS 4
#endif
constexpr int RS_THREADS = GS_RS_THREADS;   // >= 256: threads 0..255 also own one digit each
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr uint32_t ST_AGG = 0x40000000u, ST_INC = 0x80000000u, ST_VAL = 0x3FFFFFFFu;

struct PassList { int n; int shift[8]; uint32_t mask[8]; };

template <typename KeyT>
__global__ void __launch_bounds__(256) radix_hist_kernel(const KeyT* __restrict__ keys, uint64_t n_max,
                                                         const uint32_t* __restrict__ d_n, PassList pl,
                                                         uint32_t* __restrict__ ghist) {
    __shared__ uint32_t sh[8][256];
    const uint64_t n = d_n ? min((uint64_t)*d_n, n_max) : n_max;   // element count may live on the device (no host sync)
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x) (&sh[0][0])[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const KeyT k = keys[i];
#pragma unroll
        for (int p = 0; p < 8; p++)
            if (p < pl.n) atomicAdd(&sh[p][(uint32_t)(k >> pl.shift[p]) & pl.mask[p]], 1u);
    }
    __syncthreads();
    for (int p = 0; p < pl.n; p++)
        for (int d = threadIdx.x; d < 256; d += blockDim.x)
            if (sh[p][d]) atomicAdd(&ghist[p * 256 + d], sh[p][d]);
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

// RS_IPT items per thread (tile = 256 * RS_IPT pairs: 4096 for large inputs, 2048 when the whole input then fits in one
// resident wave); BALLOT selects the peer-finding primitive of the ranking step.
template <typename KeyT, int RS_IPT, bool BALLOT>
__global__ void __launch_bounds__(RS_THREADS, GS_RS_CTAS) onesweep_kernel(const KeyT* __restrict__ kin, const uint32_t* __restrict__ vin,
                                                                 KeyT* __restrict__ kout, uint32_t* __restrict__ vout, uint64_t n_max,
                                                                 const uint32_t* __restrict__ d_n,
                                                                 int shift, uint32_t mask, int nbits, const uint32_t* __restrict__ ghist,
                                                                 uint32_t* counter, uint32_t* status, unsigned long long* agg2,
                                                                 int aligned16) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int RS_TILE = RS_THREADS * RS_IPT;
    OnesweepSmem<KeyT, RS_IPT>& S = *reinterpret_cast<OnesweepSmem<KeyT, RS_IPT>*>(smem_raw);
    const uint64_t n = d_n ? min((uint64_t)*d_n, n_max) : n_max;

    if (threadIdx.x == 0) { S.tile = atomicAdd(counter, 1u); mbar_init(&S.bar, 1); }
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

constexpr uint32_t RS_BIG_TILE = RS_THREADS * 16, RS_SMALL_TILE = RS_THREADS * 8, RS_TINY_TILE = RS_THREADS * 4;
constexpr size_t RS_HIST_BYTES = 8 * 256 * 4;
constexpr size_t RS_CTR_BYTES = 64;

template <typename KeyT, int IPT, bool BALLOT>
void launch_passes(const KeyT* keys_in, const uint32_t* vals_in, KeyT* kbuf[2], uint32_t* vbuf[2], uint64_t n, const uint32_t* d_n,
                   const GsSortPlan& plan, const PassList& pl, uint32_t tiles, uint32_t* ghist, uint32_t* counters,
                   uint32_t* status, unsigned long long* agg2, size_t groups, cudaStream_t st) {
    const size_t smem = sizeof(OnesweepSmem<KeyT, IPT>);
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(onesweep_kernel<KeyT, IPT, BALLOT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set = true;
    }
    const KeyT* kin = keys_in;
    const uint32_t* vin = vals_in;
    int dst = 0;
    for (int p = 0; p < plan.num_passes; p++) {
        const int aligned16 = (((uintptr_t)kin | (uintptr_t)vin) & 15) == 0;
        onesweep_kernel<KeyT, IPT, BALLOT><<<tiles, RS_THREADS, smem, st>>>(
            kin, vin, kbuf[dst], vbuf[dst], n, d_n, pl.shift[p], pl.mask[p], plan.bits[p], ghist + p * 256, counters + p,
            status + (size_t)p * tiles * 256, agg2 + (size_t)p * groups * 256, aligned16);
        kin = kbuf[dst]; vin = vbuf[dst];
        dst ^= 1;
    }
}

static int g_rs_tiny = 0;   // experiment knob: 1024-pair tiles for inputs around 1 M

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
    const bool small = n <= (uint64_t)RS_SMALL_TILE * 148 * GS_RS_CTAS;
    const bool tiny = g_rs_tiny && n <= (uint64_t)RS_TINY_TILE * 148 * GS_RS_CTAS * 2;
    const uint32_t tile_elems = tiny ? RS_TINY_TILE : small ? RS_SMALL_TILE : RS_BIG_TILE;
    const uint32_t tiles = (uint32_t)((n + tile_elems - 1) / tile_elems);
    const size_t status_bytes = (size_t)tiles * 256 * 4;
    uint32_t* ghist = (uint32_t*)scratch;
    uint32_t* counters = (uint32_t*)((char*)scratch + RS_HIST_BYTES);
    uint32_t* status = (uint32_t*)((char*)scratch + RS_HIST_BYTES + RS_CTR_BYTES);
    const size_t groups = (tiles + RS_GROUP - 1) / RS_GROUP;
    const size_t agg_bytes = groups * 256 * 8;
    unsigned long long* agg2 = (unsigned long long*)((char*)status + status_bytes * plan.num_passes);
    cudaMemsetAsync(scratch, 0, RS_HIST_BYTES + RS_CTR_BYTES + (status_bytes + agg_bytes) * plan.num_passes, st);
    PassList pl;
    pl.n = plan.num_passes;
    for (int p = 0; p < 8; p++) { pl.shift[p] = p < pl.n ? plan.shifts[p] : 0; pl.mask[p] = p < pl.n ? ((1u << plan.bits[p]) - 1u) : 0u; }
    uint64_t hb = (n + 2047) / 2048; int hist_blocks = (int)(hb < 148 * 8 ? hb : 148 * 8);
    radix_hist_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys_in, n, d_n, pl, ghist);
    const bool ballot = (hints & GS_SORT_HINT_DISTINCT) != 0;
#define GS_RS_GO(IPT, B) launch_passes<KeyT, IPT, B>(keys_in, vals_in, kbuf, vbuf, n, d_n, plan, pl, tiles, ghist, counters, status, agg2, groups, st)
    if (tiny) { if (ballot) GS_RS_GO(4, true); else GS_RS_GO(4, false); }
    else if (small) { if (ballot) GS_RS_GO(8, true); else GS_RS_GO(8, false); }
    else { if (ballot) GS_RS_GO(16, true); else GS_RS_GO(16, false); }
#undef GS_RS_GO
    return (plan.num_passes & 1) ? 0 : 1;
}

}  // namespace

void gs_tune_sort_tiny(int v) { g_rs_tiny = v ? 1 : 0; }
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
    return RS_HIST_BYTES + RS_CTR_BYTES + (tiles * 256 * 4 + (tiles / RS_GROUP + 1) * 256 * 8) * 8;
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
