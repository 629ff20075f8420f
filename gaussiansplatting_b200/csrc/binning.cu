// binning.cu — per-tile lists by two-level binning (default pair pipeline, `sort_mode` 0).
//
// What it replaces: generateTilePairs + the 64-bit (tile|depth) sort + buildTileRanges (tiled_shaders.metal:745-794,
// tiled_rasterizer.mm:27-102 / gpu_sort.mm:456-676, sort.metal:553-589).  The result is the same, element for element:
// for every tile the Gaussians that overlap it in (depth key, Gaussian index) order, and the tile ranges.
//
// The Gaussians are already in depth order (K4a, on N keys).  Instead of duplicating every Gaussian into its ~7 tiles and
// radix-sorting P pairs by tile id, each Gaussian is duplicated into the COARSE bins (8x8 tiles) its bounding box
// touches (~1.8 per Gaussian at 1080p), the coarse pairs are stably sorted by bin id (ONE onesweep pass over P/4
// pairs when there are <= 256 bins), and each bin is then expanded locally:
//   chunk map   a bin's sorted run is cut into chunks of 256 coarse entries; chunk -> (bin, offset) through a prefix table
//   count       per chunk, per tile of the bin (64 of them): how many of the chunk's entries cover the tile (warp ballots)
//   bin scan    per bin, per tile: exclusive prefix of those counts over the bin's chunks, and the tile's total
//   tile scan   exclusive prefix of the tile totals = TileRange.start; the grand total is the pair count P
//   fill        per chunk again: every (entry, covered tile) gets slot start[tile] + chunk prefix + rank inside the chunk
//               (ballot word + popc), evaluates the footprint test of pairs.cu for the cull bit, and stores the id
// Every step is order preserving, so a tile's list comes out in depth order without a sort over the P pairs.
#include "gs_common.cuh"
#include "gs_kernels.h"

namespace {

constexpr uint32_t BT = 8;            // tiles per bin side: a bin's tiles fit one 64-bit mask
constexpr uint32_t CH = 256;          // coarse entries per chunk = threads per CTA
constexpr uint32_t NWARP = CH / 32;

// K3c: coarse key duplication.  One block = 256 consecutive Gaussians of the depth order; the block's coarse pairs form one
// contiguous output run (offsets from the scan of ccnt), owner of an output slot by fixed-step search in smem.
__global__ void __launch_bounds__(256) emit_coarse_kernel(GsProjected O, const uint32_t* __restrict__ perm,
                                                          const uint32_t* __restrict__ offs, uint32_t n, uint32_t nbx,
                                                          uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t cap) {
    __shared__ uint32_t s_off[512];
    __shared__ uint32_t s_bx[256], s_by[256], s_bw[256], s_gid[256];
    __shared__ float s_rbw[256];
    __shared__ uint32_t s_end;
    const uint32_t g0 = blockIdx.x * 256u;
    const uint32_t i = g0 + threadIdx.x;
    const uint32_t last = min(g0 + 256u, n);
    s_off[256 + threadIdx.x] = 0xFFFFFFFFu;
    if (i < n) {
        const uint32_t g = perm[i];
        s_gid[threadIdx.x] = g;
        s_off[threadIdx.x] = offs[i];
        const uint4 bb = O.bbox[g];
        const uint32_t bw = (bb.z / BT) - (bb.x / BT) + 1u;
        s_bx[threadIdx.x] = bb.x / BT; s_by[threadIdx.x] = bb.y / BT; s_bw[threadIdx.x] = bw;
        s_rbw[threadIdx.x] = 1.0f / (float)bw;
    } else s_off[threadIdx.x] = 0xFFFFFFFFu;
    if (threadIdx.x == 0) s_end = offs[last];
    __syncthreads();
    const uint32_t begin = s_off[0], stop = min(s_end, cap);
    for (uint32_t j = begin + threadIdx.x; j < stop; j += 256u) {
        uint32_t lo = 0;
#pragma unroll
        for (uint32_t sft = 128u; sft >= 1u; sft >>= 1) { if (s_off[lo + sft] <= j) lo += sft; }
        const uint32_t k = j - s_off[lo];                     // < 290 coarse bins per Gaussian (256 tiles in a strip)
        const uint32_t bw = s_bw[lo];
        const uint32_t row = (uint32_t)(((float)k + 0.5f) * s_rbw[lo]);
        keys[j] = (s_by[lo] + row) * nbx + s_bx[lo] + (k - row * bw);
        vals[j] = s_gid[lo];
    }
}

// K2 + K3c + the bin sort's histogram kernel in one: the exclusive scan of the coarse counts in depth order runs as a
// look-back over the blocks (ticket order, gs::lookback_excl_warp), each block then
// emits its run as emit_coarse_kernel does and counts the digits of the coming radix passes (plan order, at most two).
struct EmitDigits { int np; int shift[2]; uint32_t mask[2]; };
constexpr uint32_t EC_IPT = 4, EC_RUN = 256u * EC_IPT;   // Gaussians per thread / per block: one look-back per 1024 Gaussians
__global__ void __launch_bounds__(256, 6) emit_coarse_scan_kernel(GsProjected O, const uint32_t* __restrict__ perm, uint32_t n, uint32_t nbx,
                                                               uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t cap,
                                                               uint32_t* __restrict__ total, uint32_t* counter, unsigned long long* status,
                                                               unsigned long long* group_sum, uint32_t* __restrict__ ghist, const EmitDigits dg) {
    __shared__ uint32_t s_off[2 * EC_RUN];                 // first output slot of every Gaussian of the run; [EC_RUN..) = sentinel
    __shared__ uint32_t s_pk[EC_RUN], s_gid[EC_RUN];       // bin x | bin y << 10 | width in bins << 20 ; Gaussian index
    __shared__ uint32_t s_hist[2][256];
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_tile, s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(counter, 1u);
    s_hist[0][threadIdx.x] = 0u; s_hist[1][threadIdx.x] = 0u;
#pragma unroll
    for (uint32_t q = 0; q < EC_IPT; q++) s_off[EC_RUN + q * 256u + threadIdx.x] = 0xFFFFFFFFu;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t i0 = tile * EC_RUN + threadIdx.x * EC_IPT;      // a thread owns EC_IPT consecutive Gaussians of the depth order
    uint32_t g[EC_IPT], c[EC_IPT], pk[EC_IPT];
    if (i0 + EC_IPT <= n) {
        const uint4 p4 = *reinterpret_cast<const uint4*>(perm + i0);
        g[0] = p4.x; g[1] = p4.y; g[2] = p4.z; g[3] = p4.w;
    } else {
#pragma unroll
        for (uint32_t q = 0; q < EC_IPT; q++) g[q] = (i0 + q < n) ? perm[i0 + q] : 0u;
    }
    uint32_t sum = 0;
#pragma unroll
    for (uint32_t q = 0; q < EC_IPT; q++) {
        c[q] = (i0 + q < n) ? O.ccnt[g[q]] : 0u;
        const uint4 bb = O.bbox[g[q]];                       // not made to wait for the count: both gathers fly together
        pk[q] = (bb.x / BT) | ((bb.y / BT) << 10) | (((bb.z / BT) - (bb.x / BT) + 1u) << 20);
        sum += c[q];
    }
    uint32_t block_sum;
    uint32_t ex = gs::cta_excl_scan<8>(sum, s_w, &block_sum);
    if (threadIdx.x < 32u) {
        const uint32_t excl = gs::lookback_excl_warp(tile, block_sum, status, group_sum, threadIdx.x);
        if (threadIdx.x == 0) {
            s_prefix = excl;
            if ((uint64_t)(tile + 1u) * EC_RUN >= n) *total = excl + block_sum;
        }
    }
#pragma unroll
    for (uint32_t q = 0; q < EC_IPT; q++) {                  // block-relative for now: the prefix is added when a slot is looked up
        s_off[threadIdx.x * EC_IPT + q] = (i0 + q < n) ? ex : 0xFFFFFFFFu;
        s_pk[threadIdx.x * EC_IPT + q] = pk[q]; s_gid[threadIdx.x * EC_IPT + q] = g[q];
        ex += c[q];
    }
    __syncthreads();
    const uint32_t begin = s_prefix;
    const uint32_t count = begin < cap ? min(block_sum, cap - begin) : 0u;
    for (uint32_t j = threadIdx.x; j < count; j += 256u) {
        uint32_t lo = 0;
#pragma unroll
        for (uint32_t sft = EC_RUN / 2; sft >= 1u; sft >>= 1) { if (s_off[lo + sft] <= j) lo += sft; }
        const uint32_t k = j - s_off[lo];                     // < 290 coarse bins per Gaussian (256 tiles in a strip)
        const uint32_t p = s_pk[lo], bw = p >> 20;
        const uint32_t row = (uint32_t)(((float)k + 0.5f) * (1.0f / (float)bw));
        const uint32_t key = (((p >> 10) & 1023u) + row) * nbx + (p & 1023u) + (k - row * bw);
        keys[begin + j] = key;
        vals[begin + j] = s_gid[lo];
        atomicAdd(&s_hist[0][(key >> dg.shift[0]) & dg.mask[0]], 1u);
        if (dg.np > 1) atomicAdd(&s_hist[1][(key >> dg.shift[1]) & dg.mask[1]], 1u);
    }
    __syncthreads();
    for (int p = 0; p < dg.np; p++) {
        const uint32_t h = s_hist[p][threadIdx.x];
        if (h) atomicAdd(&ghist[p * 256 + threadIdx.x], h);
    }
}

// chunk map: cstart[b] = number of chunks of the bins before b; cstart[nbins] = total; chunk_bin[c] = the bin of chunk c.  One CTA.
__global__ void __launch_bounds__(1024) chunk_map_kernel(const uint2* __restrict__ bin_ranges, uint32_t nbins, uint32_t* __restrict__ cstart,
                                                         uint32_t* __restrict__ chunk_bin) {
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nbins; base += 1024u) {
        const uint32_t b = base + threadIdx.x;
        const uint32_t c = b < nbins ? (bin_ranges[b].y + CH - 1u) / CH : 0u;
        uint32_t total;
        const uint32_t ex = gs::cta_excl_scan<32>(c, s_w, &total);
        const uint32_t carry = s_carry;
        if (b < nbins) {
            cstart[b] = carry + ex;
            for (uint32_t k = 0; k < c; k++) chunk_bin[carry + ex + k] = b;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) cstart[nbins] = s_carry;
}

// single-pass coarse sort: the sort's digit histogram IS the bin histogram, so the bins' runs and the chunk map come from
// one scan of it (replaces the run-boundary kernels on the sorted bin ids).  One CTA, nbins <= 256.
__global__ void __launch_bounds__(256) bins_from_hist_kernel(const uint32_t* __restrict__ hist, uint32_t nbins, uint2* __restrict__ bin_ranges,
                                                             uint32_t* __restrict__ cstart, uint32_t* __restrict__ chunk_bin) {
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_cs[257];
    const uint32_t b = threadIdx.x;
    const uint32_t cnt = b < nbins ? hist[b] : 0u;
    uint32_t tot;
    const uint32_t start = gs::cta_excl_scan<8>(cnt, s_w, &tot);
    __syncthreads();
    const uint32_t cs = gs::cta_excl_scan<8>((cnt + CH - 1u) / CH, s_w, &tot);
    if (b < nbins) { bin_ranges[b] = make_uint2(start, cnt); cstart[b] = cs; }
    s_cs[b] = b < nbins ? cs : tot;
    if (b == 0) { cstart[nbins] = tot; s_cs[256] = tot; }
    __syncthreads();
    // chunk table, coalesced: chunk c belongs to the last bin whose first chunk is <= c
    for (uint32_t c = threadIdx.x; c < tot; c += 256u) {
        uint32_t lo = 0;
#pragma unroll
        for (uint32_t sft = 128u; sft >= 1u; sft >>= 1) { if (s_cs[lo + sft] <= c) lo += sft; }
        chunk_bin[c] = lo;
    }
}

struct ChunkCtx { uint32_t bin, first, count; };   // the chunk's bin, first coarse entry, entries in the chunk
__device__ __forceinline__ ChunkCtx locate_chunk(uint32_t chunk, const uint32_t* __restrict__ chunk_bin, const uint32_t* __restrict__ cstart,
                                                 const uint2* __restrict__ bin_ranges) {
    const uint32_t b = chunk_bin[chunk];
    const uint2 br = bin_ranges[b];
    const uint32_t off = (chunk - cstart[b]) * CH;
    ChunkCtx c; c.bin = b; c.first = br.x + off; c.count = min(CH, br.y - off);
    return c;
}

// bit-matrix transpose across the warp: bit i of the result in lane t = bit t of x in lane i (five exchange rounds)
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, uint32_t lane) {
    uint32_t m = 0x0000FFFFu;
#pragma unroll
    for (uint32_t j = 16u; j >= 1u; j >>= 1) {
        const uint32_t o = __shfl_xor_sync(0xFFFFFFFFu, x, j);
        if (lane & j) x ^= (x ^ (o >> j)) & m;
        else x ^= ((o ^ (x >> j)) & m) << j;
        m ^= m << (j >> 1);
    }
    return x;
}

// rectangle of local tiles (within bin bx,by) that bbox bb covers
struct BinRect { uint32_t x0, y0, w, h; };
__device__ __forceinline__ BinRect bin_rect(const uint4 bb, uint32_t bx, uint32_t by) {
    const uint32_t ox = bx * BT, oy = by * BT;
    BinRect r;
    r.x0 = max(bb.x, ox) - ox; r.y0 = max(bb.y, oy) - oy;
    r.w = min(bb.z, ox + BT - 1u) - ox - r.x0 + 1u; r.h = min(bb.w, oy + BT - 1u) - oy - r.y0 + 1u;
    return r;
}

// count: chunk_counts[chunk][t] = entries of the chunk covering local tile t, and chunk_words[chunk][warp][t] = WHICH of the
// warp's 32 entries cover it.  Every entry covers a rectangle of the bin's 8x8 tiles = a 64-bit mask; two 32x32 bit-matrix
// transposes across the warp turn the 32 masks into the 64 words "which lanes cover tile t" (five shfl.bfly rounds each).  The
// words are stored (2 KB per chunk, coalesced) so that the fill kernel neither gathers the rectangles again nor searches for a
// pair's owner; the counts are their popcounts summed over the chunk's eight warps.  Grid-stride over the chunks.
__global__ void __launch_bounds__(CH) bin_count_kernel(GsProjected O, const uint32_t* __restrict__ cvals, const uint2* __restrict__ bin_ranges,
                                                       const uint32_t* __restrict__ cstart, const uint32_t* __restrict__ chunk_bin,
                                                       uint32_t nbins, uint32_t nbx, uint32_t* __restrict__ chunk_counts,
                                                       uint32_t* __restrict__ chunk_words) {
    __shared__ uint32_t s_cnt[NWARP][64];
    const uint32_t total = cstart[nbins];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    for (uint32_t chunk = blockIdx.x; chunk < total; chunk += gridDim.x) {
        const ChunkCtx cc = locate_chunk(chunk, chunk_bin, cstart, bin_ranges);
        unsigned long long mask = 0ull;
        if (threadIdx.x < cc.count) {
            const uint32_t g = cvals[cc.first + threadIdx.x];
            const BinRect r = bin_rect(O.bbox[g], cc.bin % nbx, cc.bin / nbx);
            const unsigned long long row = (unsigned long long)(((1u << r.w) - 1u) << r.x0);               // < 256: one byte per tile row
            mask = row * ((0x0101010101010101ull >> (8u * (8u - r.h))) << (8u * r.y0));                     // rows y0 .. y0 + h - 1
        }
        const uint32_t w0 = warp_transpose32((uint32_t)mask, lane), w1 = warp_transpose32((uint32_t)(mask >> 32), lane);
        uint32_t* wd = chunk_words + ((size_t)chunk * NWARP + warp) * 64u;
        wd[lane] = w0; wd[lane + 32u] = w1;
        s_cnt[warp][lane] = __popc(w0); s_cnt[warp][lane + 32u] = __popc(w1);
        __syncthreads();
        if (threadIdx.x < 64u) {
            uint32_t sum = 0;
#pragma unroll
            for (uint32_t w = 0; w < NWARP; w++) sum += s_cnt[w][threadIdx.x];
            chunk_counts[(size_t)chunk * 64u + threadIdx.x] = sum;
        }
        __syncthreads();
    }
}

// tile scan: ranges[t] = (exclusive prefix, count); *total = pair count.  One CTA of 1024 threads (the last bin-scan CTA).
// Ranges are clipped to the list capacity: when the pair buffers turn out too small the frame is redone by the host, and
// the kernels of the clipped attempt must stay inside the buffers.
__device__ __forceinline__ void tile_scan_cta(const uint32_t* __restrict__ tile_count, uint32_t ntiles, uint2* __restrict__ ranges,
                                              uint32_t* __restrict__ total, uint32_t cap_pairs, uint32_t* s_w, uint32_t* s_carry) {
    if (threadIdx.x == 0) *s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < ntiles; base += 8192u) {
        const uint32_t t0 = base + threadIdx.x * 8u;
        uint32_t c[8], sum = 0;
#pragma unroll
        for (uint32_t q = 0; q < 8u; q++) { c[q] = (t0 + q < ntiles) ? __ldcg(tile_count + t0 + q) : 0u; sum += c[q]; }
        uint32_t tot;
        uint32_t ex = gs::cta_excl_scan<32>(sum, s_w, &tot) + *s_carry;
#pragma unroll
        for (uint32_t q = 0; q < 8u; q++) {
            if (t0 + q < ntiles) { const uint32_t a = min(ex, cap_pairs); ranges[t0 + q] = make_uint2(a, min(c[q], cap_pairs - a)); }
            ex += c[q];
        }
        __syncthreads();
        if (threadIdx.x == 0) *s_carry += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = *s_carry;
}

// bin scan: one CTA per bin, 16 groups x 64 tiles.  chunk_counts[c][t] becomes the exclusive prefix over the bin's chunks;
// tile_count[global tile] = the tile's total (0 for a bin without entries; tiles outside the image are skipped).
// The CTA that finishes last (ticket in *done, which it resets) then scans the tile totals into the tile ranges.
__global__ void __launch_bounds__(1024) bin_scan_kernel(const uint32_t* __restrict__ cstart, uint32_t nbx, uint32_t ntx, uint32_t nty,
                                                        uint32_t* __restrict__ chunk_counts, uint32_t* __restrict__ tile_count,
                                                        uint32_t* done, uint2* __restrict__ ranges, uint32_t* __restrict__ total, uint32_t cap_pairs) {
    __shared__ uint32_t s_part[16][64];
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry, s_last;
    const uint32_t b = blockIdx.x;
    const uint32_t t = threadIdx.x & 63u, grp = threadIdx.x >> 6;
    const uint32_t c0 = cstart[b], c1 = cstart[b + 1];
    const uint32_t nch = c1 - c0, per = (nch + 15u) / 16u;
    const uint32_t a = min(c0 + grp * per, c1), e = min(a + per, c1);
    uint32_t s = 0;
    for (uint32_t c = a; c < e; c++) s += chunk_counts[(size_t)c * 64u + t];
    s_part[grp][t] = s;
    __syncthreads();
    uint32_t run = 0, tot = 0;
#pragma unroll
    for (uint32_t g = 0; g < 16u; g++) { const uint32_t v = s_part[g][t]; if (g < grp) run += v; tot += v; }
    for (uint32_t c = a; c < e; c++) {
        const uint32_t v = chunk_counts[(size_t)c * 64u + t];
        chunk_counts[(size_t)c * 64u + t] = run;
        run += v;
    }
    if (grp == 0) {
        const uint32_t tx = (b % nbx) * BT + (t & 7u), ty = (b / nbx) * BT + (t >> 3);
        if (tx < ntx && ty < nty) tile_count[ty * ntx + tx] = tot;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(done, 1u) == gridDim.x - 1u) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) *done = 0u;
    __threadfence();
    tile_scan_cta(tile_count, ntx * nty, ranges, total, cap_pairs, s_w, &s_carry);
}

// fill: every (entry, covered tile) of the chunk is stored at start[tile] + prefix over earlier chunks + rank in the chunk.
// bin_count left the chunk's 8 x 64 coverage words ("which of warp w's 32 entries cover tile t").  Lane l of warp w takes the
// words of tiles l and l + 32 for ITS warp's entries and walks their set bits in ascending order (entry order = depth order),
// writing (Gaussian id, destination slot) into a shared-memory image of the chunk's output: tile runs back to back, inside a
// run the warps' entries in warp order (offsets from the popcounts).  Then the CTA copies the image out, thread k the k-th
// pair: consecutive threads hit consecutive slots of a tile's list -- coalesced stores, five instructions per pair.  Round 1 assigned one thread per pair (a search for the owner entry
// and a per-pair cull test: 168 instructions per pair) and every lane of a store instruction hit a different tile's list --
// 18 sectors per instruction, 56 M four-byte sector writes at C4, which is what bounded the kernel once the cull test had
// gone.  A chunk with more than FILL_CAP pairs (large splats) writes straight to the lists instead.  CULL keeps the cull bit
// (gs_tune "pair_cull").
constexpr uint32_t FILL_CAP = 4096;
template <bool CULL>
__global__ void __launch_bounds__(CH) bin_fill_kernel(GsProjected O, const uint32_t* __restrict__ cvals, const uint2* __restrict__ bin_ranges,
                                                      const uint32_t* __restrict__ cstart, const uint32_t* __restrict__ chunk_bin,
                                                      uint32_t nbins, uint32_t nbx, uint32_t ntx,
                                                      uint32_t nty, const uint32_t* __restrict__ chunk_prefix,
                                                      const uint32_t* __restrict__ chunk_words,
                                                      const uint2* __restrict__ ranges, uint32_t* __restrict__ out_vals, uint32_t cap_pairs) {
    __shared__ uint32_t s_cnt[NWARP][64];                      // popcounts, then their exclusive prefix over the warps
    __shared__ uint32_t s_toff[65];                            // chunk-local start of each tile's run; [64] = pairs of the chunk
    __shared__ uint32_t s_tcnt[64], s_gbase[64];
    __shared__ uint32_t s_gid[CH];
    __shared__ uint2 s_out[FILL_CAP];                          // (Gaussian id, destination slot) of the chunk's pairs, tile runs back to back
    __shared__ gs::CullRec s_cull[CULL ? CH : 1];
    // with a tile order asked for, CTA 0 (first wave) computes it from the finished ranges while the others fill
    const bool with_order = O.tile_order != nullptr;
    if (with_order && blockIdx.x == 0) { gs::tile_order_cta<CH>(ranges, ntx * nty, O.tile_order); return; }
    const uint32_t bid = with_order ? blockIdx.x - 1u : blockIdx.x, nblk = with_order ? gridDim.x - 1u : gridDim.x;
    const uint32_t total = cstart[nbins];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    for (uint32_t chunk = bid; chunk < total; chunk += nblk) {
        const ChunkCtx cc = locate_chunk(chunk, chunk_bin, cstart, bin_ranges);
        const uint32_t bx = cc.bin % nbx, by = cc.bin / nbx;
        const uint32_t* wd = chunk_words + ((size_t)chunk * NWARP + warp) * 64u;
        const uint32_t word[2] = {wd[lane], wd[lane + 32u]};
        if (threadIdx.x < cc.count) {
            const uint32_t g = cvals[cc.first + threadIdx.x];
            s_gid[threadIdx.x] = g;
            if (CULL) s_cull[threadIdx.x] = gs::cull_rec(O.rec[3 * (size_t)g], O.rec[3 * (size_t)g + 1]);
        }
        s_cnt[warp][lane] = __popc(word[0]); s_cnt[warp][lane + 32u] = __popc(word[1]);
        __syncthreads();
        if (threadIdx.x < 64u) {                               // two warps: per tile the warps' prefix, then the scan over the tiles
            const uint32_t t = threadIdx.x;
            uint32_t run = 0;
#pragma unroll
            for (uint32_t w = 0; w < NWARP; w++) { const uint32_t c = s_cnt[w][t]; s_cnt[w][t] = run; run += c; }
            s_tcnt[t] = run;
            const uint32_t tx = bx * BT + (t & 7u), ty = by * BT + (t >> 3);
            s_gbase[t] = (tx < ntx && ty < nty) ? ranges[ty * ntx + tx].x + chunk_prefix[(size_t)chunk * 64u + t] : 0u;
            uint32_t inc = run;
#pragma unroll
            for (uint32_t d = 1; d < 32u; d <<= 1) { const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane >= d) inc += o; }
            if (t == 31u) s_toff[64] = inc;                    // first half's total, parked
            s_toff[t] = inc - run;
        }
        __syncthreads();
        if (threadIdx.x >= 32u && threadIdx.x < 64u) s_toff[threadIdx.x] += s_toff[64];
        __syncthreads();
        if (threadIdx.x == 63u) s_toff[64] = s_toff[63] + s_tcnt[63];
        __syncthreads();
        const bool staged = !CULL && s_toff[64] <= FILL_CAP;   // CTA-uniform
#pragma unroll
        for (uint32_t hh = 0; hh < 2u; hh++) {
            const uint32_t t = lane + 32u * hh;
            uint32_t wv = word[hh];
            if (wv == 0u) continue;
            if (staged) {
                uint32_t pos = s_toff[t] + s_cnt[warp][t], dst = s_gbase[t] + s_cnt[warp][t];
                while (wv) {
                    const uint32_t e = warp * 32u + (uint32_t)__ffs(wv) - 1u;
                    wv &= wv - 1u;
                    s_out[pos++] = make_uint2(s_gid[e], dst++);
                }
            } else {
                const uint32_t tx = bx * BT + (t & 7u), ty = by * BT + (t >> 3);
                uint32_t slot = s_gbase[t] + s_cnt[warp][t];
                while (wv) {
                    const uint32_t e = warp * 32u + (uint32_t)__ffs(wv) - 1u;
                    wv &= wv - 1u;
                    bool cull = false;
                    if (CULL) cull = gs::box_cannot_contribute(s_cull[e], (float)(tx * GS_TILE) + 0.5f, (float)(ty * GS_TILE) + 0.5f, 15.0f, 15.0f);
                    if (slot < cap_pairs) out_vals[slot] = s_gid[e] | (cull ? GS_CULL_BIT : 0u);
                    slot++;
                }
            }
        }
        __syncthreads();
        if (staged) {                                          // consecutive threads = consecutive slots of a tile's run (but at the run ends)
            const uint32_t npairs = s_toff[64];
            for (uint32_t k = threadIdx.x; k < npairs; k += CH) {
                const uint2 o = s_out[k];
                if (o.y < cap_pairs) out_vals[o.y] = o.x;
            }
            __syncthreads();
        }
    }
}

// sorted 64-bit keys of the reference for inspection: slot j belongs to the last tile whose start is <= j
__global__ void __launch_bounds__(256) rebuild_keys64_ranges_kernel(const uint2* __restrict__ ranges, uint32_t ntiles,
                                                                    const uint32_t* __restrict__ vals, const uint32_t* __restrict__ dkey,
                                                                    uint32_t n, uint64_t* __restrict__ keys) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint32_t lo = 0, hi = ntiles;
    while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (ranges[mid].x <= j) lo = mid; else hi = mid; }
    keys[j] = ((uint64_t)lo << 32) | (uint64_t)dkey[vals[j] & GS_ID_MASK];
}

}  // namespace

// 1 = the fill evaluates the per-(tile, Gaussian) cull bit (GS_CULL_BIT); 0 = no pair is ever marked
static int g_pair_cull = 0;
void gs_tune_pair_cull(int v) { g_pair_cull = v ? 1 : 0; }

uint32_t gs_bin_count(uint32_t ntx, uint32_t nty) { return ((ntx + BT - 1u) / BT) * ((nty + BT - 1u) / BT); }
uint32_t gs_bin_nbx(uint32_t ntx) { return (ntx + BT - 1u) / BT; }
// chunks needed for `coarse_cap` coarse entries spread over `nbins` bins (every bin may end with a partial chunk)
size_t gs_bin_max_chunks(uint64_t coarse_cap, uint32_t nbins) { return (size_t)(coarse_cap / CH) + nbins + 1; }

void gs_launch_emit_coarse(const GsProjected& O, const uint32_t* perm, const uint32_t* offs, uint32_t n, uint32_t ntx,
                           uint32_t* bin_keys, uint32_t* vals, uint32_t cap, cudaStream_t st) {
    if (n == 0) return;
    emit_coarse_kernel<<<(n + 255) / 256, 256, 0, st>>>(O, perm, offs, n, gs_bin_nbx(ntx), bin_keys, vals, cap);
}

size_t gs_emit_coarse_scratch_bytes(uint32_t n) { const size_t b = (size_t)n / EC_RUN + 2u; return 8 + b * 8 + (b / 32 + 1) * 8; }
// fused scan + coarse emit + digit histograms; `plan` is the coming sort's (1 or 2 passes), `ghist` its prepared block
void gs_launch_emit_coarse_scan(const GsProjected& O, const uint32_t* perm, uint32_t n, uint32_t ntx, uint32_t* bin_keys, uint32_t* vals,
                                uint32_t cap, uint32_t* total, void* scratch, const GsSortPlan& plan, uint32_t* ghist, cudaStream_t st) {
    if (n == 0) { cudaMemsetAsync(total, 0, 4, st); return; }
    const uint32_t blocks = (n + EC_RUN - 1u) / EC_RUN;
    cudaMemsetAsync(scratch, 0, 8 + (size_t)blocks * 8 + ((size_t)blocks / 32 + 1) * 8, st);
    unsigned long long* status = (unsigned long long*)((char*)scratch + 8);
    EmitDigits dg;
    dg.np = plan.num_passes;
    for (int p = 0; p < 2; p++) { dg.shift[p] = p < plan.num_passes ? plan.shifts[p] : 0; dg.mask[p] = p < plan.num_passes ? (1u << plan.bits[p]) - 1u : 0u; }
    emit_coarse_scan_kernel<<<blocks, 256, 0, st>>>(O, perm, n, gs_bin_nbx(ntx), bin_keys, vals, cap, total, (uint32_t*)scratch,
                                                    status, status + blocks, ghist, dg);
}

// the bins' runs + chunk map from the coarse sort's digit histogram, as a launch of its own: it depends only on the histogram
// (final once the emit kernel is done), so the caller runs it on a side stream while the bin-sort pass is in flight
void gs_launch_bins_from_hist(const uint32_t* bin_hist, uint32_t ntx, uint32_t nty, uint2* bin_ranges, uint32_t* cstart, uint32_t* chunk_bin,
                              cudaStream_t st) {
    const uint32_t nbins = gs_bin_count(ntx, nty);
    if (nbins) bins_from_hist_kernel<<<1, 256, 0, st>>>(bin_hist, nbins, bin_ranges, cstart, chunk_bin);
}

void gs_launch_bin_expand(const GsProjected& O, const uint32_t* cvals, const uint32_t* bin_hist, uint2* bin_ranges, uint32_t ntx, uint32_t nty,
                          uint32_t* cstart, uint32_t* chunk_counts, uint32_t* tile_count, uint2* ranges, uint32_t* total_pairs,
                          uint32_t* out_vals, uint32_t cap_pairs, uint32_t grid_chunks, uint32_t* chunk_bin, uint32_t* chunk_words, uint32_t* done,
                          cudaStream_t st) {
    const uint32_t nbx = gs_bin_nbx(ntx), nbins = gs_bin_count(ntx, nty);
    if (nbins == 0) return;
    // The number of chunks lives on the device; the chunk kernels stride over it, so any grid is correct.  The caller
    // sizes the grid from the previous frame's coarse count (one chunk per CTA: the chunk's dependent gathers of many
    // CTAs overlap), falling back to a persistent grid.
    const uint32_t grid = grid_chunks > 148u * 4u ? grid_chunks : 148u * 4u;
    if (bin_hist == GS_BINS_READY) {}                          // gs_launch_bins_from_hist already ran (side stream, under the bin sort)
    else if (bin_hist) bins_from_hist_kernel<<<1, 256, 0, st>>>(bin_hist, nbins, bin_ranges, cstart, chunk_bin);
    else chunk_map_kernel<<<1, 1024, 0, st>>>(bin_ranges, nbins, cstart, chunk_bin);
    bin_count_kernel<<<grid, CH, 0, st>>>(O, cvals, bin_ranges, cstart, chunk_bin, nbins, nbx, chunk_counts, chunk_words);
    bin_scan_kernel<<<nbins, 1024, 0, st>>>(cstart, nbx, ntx, nty, chunk_counts, tile_count, done, ranges, total_pairs, cap_pairs);
    if (g_pair_cull) bin_fill_kernel<true><<<grid + (O.tile_order ? 1u : 0u), CH, 0, st>>>(O, cvals, bin_ranges, cstart, chunk_bin, nbins, nbx, ntx, nty, chunk_counts, chunk_words, ranges, out_vals, cap_pairs);
    else bin_fill_kernel<false><<<grid + (O.tile_order ? 1u : 0u), CH, 0, st>>>(O, cvals, bin_ranges, cstart, chunk_bin, nbins, nbx, ntx, nty, chunk_counts, chunk_words, ranges, out_vals, cap_pairs);
}

void gs_launch_rebuild_keys64_ranges(const uint2* ranges, uint32_t ntiles, const uint32_t* sorted_vals, const uint32_t* dkey,
                                     uint32_t n, uint64_t* keys, cudaStream_t st) {
    if (n) rebuild_keys64_ranges_kernel<<<(n + 255) / 256, 256, 0, st>>>(ranges, ntiles, sorted_vals, dkey, n, keys);
}
