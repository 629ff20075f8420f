// gs_kernels.h — internal launcher interface between the kernel translation units and the C ABI (api.cu).
// Every launcher enqueues on the given stream and returns; none synchronises.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gs_common.cuh"

// meta words written by preprocess / scan (one small device array per rasterizer)
enum { GS_META_TOTAL_PAIRS = 0, GS_META_KEY_OR = 1, GS_META_KEY_AND = 2, GS_META_COARSE_PAIRS = 3, GS_META_COUNT = 5 };

// K1 ------------------------------------------------------------------------------------------------
void gs_launch_preprocess(const GsParams& P, const GsProjected& O, const float* V, const float* VP, float sw, float sh,
                          float fx, float fy, uint32_t ntx, uint32_t nty, uint32_t n, uint32_t* meta, uint32_t* depth_hist, cudaStream_t st);

// K2: exclusive scan: out[j] = sum_{j'<j} in[perm ? perm[j'] : j']; out[n] and *total receive the sum (in == out allowed
// when perm is null).  scratch: gs_scan_scratch_bytes(n).
size_t gs_scan_scratch_bytes(uint32_t n);
void gs_launch_exclusive_scan(const uint32_t* in, const uint32_t* perm, uint32_t* out, uint32_t n, uint32_t* total,
                              void* scratch, cudaStream_t st);

// K3: emit (tile<<32 | depthKey, gaussian) pairs at the scanned offsets O.offs (Gaussian order, ty-major, tx-minor).
// Bit 31 of a value (GS_CULL_BIT) marks a pair that can never contribute; writes are bounded by cap_pairs.
void gs_launch_emit_pairs(const GsProjected& O, const uint32_t* offs, uint32_t n, uint32_t ntx, uint64_t* keys, uint32_t* vals,
                          uint32_t cap_pairs, cudaStream_t st);
// hoisted-depth mode: list position j is Gaussian perm[j] (depth order), offs[j] its scanned offset; key = tile id only
void gs_launch_emit_tile_pairs(const GsProjected& O, const uint32_t* perm, const uint32_t* offs, uint32_t n, uint32_t ntx,
                               uint32_t* tile_keys, uint32_t* vals, uint32_t cap_pairs, cudaStream_t st);

// K4: onesweep LSD radix sort -------------------------------------------------------------------------
struct GsSortPlan {
    int num_passes;
    int shifts[8];
    int bits[8];
};
// passes over [begin_bit, end_bit) in 8-bit digits, skipping digits in which `varying_mask` has no bit set
GsSortPlan gs_sort_plan(int begin_bit, int end_bit, uint64_t varying_mask);
size_t gs_sort_scratch_bytes(uint64_t n, int key_bytes);
// Sorts (keys_in, vals_in) stably.  keys_in/vals_in are not modified; the result lands in buffer
// (plan.num_passes odd ? 0 : 1) of {kbuf,vbuf}; with 0 passes the input is copied to buffer 0.
// Returns the index (0/1) of the buffer holding the result.
// hints: GS_SORT_HINT_DISTINCT = the digits met in a warp row (32 consecutive elements) are mostly distinct (random keys,
// tile ids of a Gaussian's neighbouring tiles): rank with ballots; otherwise with match.any (runs of equal digits).
// GS_SORT_HINT_PREHIST (32-bit keys) = the scratch was cleared by gs_sort_prepare and the producer of the keys has
// already counted the digits of every pass into the block it returned: no clear and no histogram kernel here.
enum { GS_SORT_HINT_DISTINCT = 1, GS_SORT_HINT_PREHIST = 2 };
uint32_t* gs_sort_prepare(uint64_t n, const GsSortPlan& plan, void* scratch, cudaStream_t st);
int gs_launch_sort64(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* kbuf[2], uint32_t* vbuf[2], uint64_t n,
                     const GsSortPlan& plan, int hints, void* scratch, cudaStream_t st);
// n is a host-side BOUND (grid, scratch); when d_n is non-null the true count is read from the device (min(*d_n, n)),
// so the caller needs no host synchronisation to learn it.
int gs_launch_sort32(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* kbuf[2], uint32_t* vbuf[2], uint64_t n,
                     const uint32_t* d_n, const GsSortPlan& plan, int hints, void* scratch, cudaStream_t st);
// passes over [begin_bit, end_bit) with the bits split evenly over ceil(bits/8) digits (13 bits -> 7 + 6)
GsSortPlan gs_sort_plan_even(int begin_bit, int end_bit);
// viewer depth keys (computeDepths, sort.metal:36-72)
void gs_launch_depth_keys32(const float4* positions, float cx, float cy, float cz, uint32_t n, uint32_t* keys,
                            uint32_t* vals, cudaStream_t st);

// K5: tile ranges from sorted keys (buildTileRanges semantics, sort.metal:553-589).  lb: numTiles+1 scratch words.
void gs_launch_tile_ranges(const uint64_t* sorted_keys, uint32_t num_pairs, uint32_t num_tiles, uint32_t* lb,
                           uint2* ranges, cudaStream_t st);
// same result from sorted u32 tile ids by run-boundary detection (O(P) coalesced reads); the pair count is read from
// the device (min(*d_n, max_pairs)); the grid is sized from max_pairs
void gs_launch_tile_ranges32(const uint32_t* sorted_tiles, uint32_t max_pairs, const uint32_t* d_n, uint32_t num_tiles,
                             uint2* ranges, cudaStream_t st);
// inspection: rebuild the reference's 64-bit sorted keys (tile<<32 | depthKey) from sorted tile ids + values
void gs_launch_rebuild_keys64(const uint32_t* sorted_tiles, const uint32_t* sorted_vals, const uint32_t* dkey, uint32_t n,
                              uint64_t* keys, cudaStream_t st);

// K6: forward blend.  image: W*H float4 RGBA; last_idx/tfinal: W*H.
void gs_launch_blend_forward(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                             uint32_t ntx, uint32_t nty, float4* image, uint32_t* last_idx, float* tfinal, int order_ready,
                             uint2* vis, size_t vstride, cudaStream_t st);

void gs_launch_blend_count(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                           uint32_t ntx, uint32_t nty, unsigned long long* counters, cudaStream_t st);

// K7: backward blend -> per-Gaussian 9-moment accumulators acc[12*n] (zeroed by the caller)
//   acc[12i+0..2] = sum dLdp_c * w            acc[12i+3] = sum q          (q = dL/dalpha * G)
//   acc[12i+4..5] = sum q*dx, q*dy            acc[12i+6..8] = sum q*dx^2, q*dx*dy, q*dy^2
void gs_launch_blend_backward(const GsProjected& O, const uint32_t* sorted_vals, const uint2* ranges, uint32_t W, uint32_t H,
                              uint32_t ntx, uint32_t nty, const float4* rendered, const float4* gt, int quantize,
                              const uint32_t* last_idx, const float* tfinal, float* acc, const uint2* vis, size_t vstride, int vis_warp_extent,
                              cudaStream_t st);

// K8: per-Gaussian chain: moments -> parameter gradients, written to (accumulate = 0) or added into the packed gradient
// planes; clears the moment rows it consumed
void gs_launch_preprocess_backward(const GsParams& P, const GsProjected& O, float* acc, const float* V, float fx,
                                   float fy, uint32_t n, GsGradPack G, int stat_mode, int accumulate, cudaStream_t st);

// two-level binning (binning.cu): coarse (8x8-tile bin, Gaussian) pairs in depth order, then per-bin expansion into the
// per-tile lists and tile ranges
uint32_t gs_bin_count(uint32_t ntx, uint32_t nty);
size_t gs_bin_max_chunks(uint64_t coarse_cap, uint32_t nbins);
void gs_launch_emit_coarse(const GsProjected& O, const uint32_t* perm, const uint32_t* offs, uint32_t n, uint32_t ntx,
                           uint32_t* bin_keys, uint32_t* vals, uint32_t cap, cudaStream_t st);
// the same pairs from ONE kernel that also scans the coarse counts (no offsets array) and counts the bin sort's digits
size_t gs_emit_coarse_scratch_bytes(uint32_t n);
void gs_launch_emit_coarse_scan(const GsProjected& O, const uint32_t* perm, uint32_t n, uint32_t ntx, uint32_t* bin_keys, uint32_t* vals,
                                uint32_t cap, uint32_t* total, void* scratch, const GsSortPlan& plan, uint32_t* ghist, cudaStream_t st);
// bin_hist: per-bin entry counts when the coarse sort was a single pass (its digit histogram, gs_sort_pass_histogram), else
// nullptr and bin_ranges must already hold the bins' runs
const uint32_t* gs_sort_pass_histogram(const void* scratch, int pass);
void gs_tune_sort_tiny(int v);
const uint32_t* gs_sort_flags(const void* scratch);   // three words: hybrid plan admitted, local pass gave up, LSD plan ran
void gs_tune_sort_hybrid(int v);
void gs_tune_sort_even(int v);   // 1 (default): full 64-bit sorts may take the prefix-passes + local-pass plan (radix_sort.cu)
#define GS_BINS_READY (reinterpret_cast<const uint32_t*>(uintptr_t(1)))   // bin_hist value: gs_launch_bins_from_hist already ran
void gs_launch_bins_from_hist(const uint32_t* bin_hist, uint32_t ntx, uint32_t nty, uint2* bin_ranges, uint32_t* cstart, uint32_t* chunk_bin,
                              cudaStream_t st);
void gs_launch_bin_expand(const GsProjected& O, const uint32_t* cvals, const uint32_t* bin_hist, uint2* bin_ranges, uint32_t ntx, uint32_t nty,
                          uint32_t* cstart, uint32_t* chunk_counts, uint32_t* tile_count, uint2* ranges, uint32_t* total_pairs,
                          uint32_t* out_vals, uint32_t cap_pairs, uint32_t grid_chunks, uint32_t* chunk_bin, uint32_t* chunk_words, uint32_t* done,
                          cudaStream_t st);
void gs_launch_rebuild_keys64_ranges(const uint2* ranges, uint32_t ntiles, const uint32_t* sorted_vals, const uint32_t* dkey,
                                     uint32_t n, uint64_t* keys, cudaStream_t st);

// viewer splat path (viewer.cu; vertexShader / fragmentShader + blend state of MTLEngine::render as compute).  rec: 3 float4 per
// Gaussian {cx, cy, k0, k1}{k2, radius, opacity, r}{g, b, valid, 0}; bbox: tile box of the quad; cnt: its tile count (0 = not drawn)
void gs_launch_view_splats(const GsParams& P, const float* V, const float* VP, float sw, float sh, float fx, float fy, uint32_t n,
                           float4* rec, uint4* bbox, uint32_t* cnt, cudaStream_t st);
// draw position j is Gaussian order[j], its pairs start at offs[j]; writes are bounded by cap_pairs
void gs_launch_view_emit(const uint4* bbox, const uint32_t* cnt, const uint32_t* order, const uint32_t* offs, uint32_t n,
                         uint32_t ntx, uint32_t* tiles, uint32_t* vals, uint32_t cap_pairs, cudaStream_t st);
// order[b] = the tile CTA b of a per-tile blend kernel takes: tiles by descending list length (blend_fwd.cu)
void gs_launch_tile_order(const uint2* ranges, uint32_t num_tiles, uint32_t* order, cudaStream_t st);
// counters (may be null): [0] += fragments shaded, [1] += fragments blended; order (may be null): gs_launch_tile_order's
void gs_launch_view_blend(const float4* rec, const uint32_t* vals, const uint2* ranges, uint32_t W, uint32_t H, uint32_t ntx,
                          uint32_t nty, int quantize, float4* image, unsigned long long* counters, const uint32_t* order,
                          cudaStream_t st);
void gs_launch_view_splats_to_wire(const float4* rec, const uint4* bbox, uint32_t n, void* out64, cudaStream_t st);

// logging loss (0.8 L1 + 0.2 D-SSIM, mean over pixels): adds the sum over pixels to *total (zeroed by the caller)
void gs_launch_loss(const float4* rendered, const float4* gt, uint32_t W, uint32_t H, float lambda, int quantize, double* total,
                    cudaStream_t st);
// allocation-free variant: per-tile partial sums in a static device array of GS_LOSS_NOALLOC_MAX_TILES doubles (loss.cu)
#define GS_LOSS_NOALLOC_MAX_TILES 32768
int gs_launch_loss_noalloc(int device, const float4* rendered, const float4* gt, uint32_t W, uint32_t H, float lambda, int quantize,
                           float* d_loss_out, cudaStream_t st, const char** err);

// layout conversion between the reference's AoS wire structs and the device SoA
void gs_launch_params_from_aos(const void* aos112, uint32_t n, GsParams P, cudaStream_t st);
void gs_launch_params_to_aos(GsParams P, uint32_t n, void* aos112, cudaStream_t st);
void gs_launch_grads_to_aos(GsGradPack G, uint32_t n, void* aos112, cudaStream_t st);
void gs_launch_grads_from_aos(const void* aos112, uint32_t n, GsGradPack G, cudaStream_t st);
void gs_launch_projected_to_aos(const GsProjected& O, uint32_t n, void* aos88, cudaStream_t st);
void gs_launch_rgba32f_to_rgba8(const float4* img, uint32_t npix, uint32_t* out, cudaStream_t st);
void gs_launch_rgba8_to_rgba32f(const uint32_t* img, uint32_t npix, float4* out, cudaStream_t st);
void gs_launch_clamp_opacity(GsParams P, uint32_t n, float max_raw, cudaStream_t st);
void gs_launch_positions(GsParams P, uint32_t n, float4* out, cudaStream_t st);

// K9: Adam -------------------------------------------------------------------------------------------
struct GsAdamState {   // the reference's 10 moment buffers (optimizer.mm:34-60), SoA planes of stride cap
    float *m_pos, *m_scale, *m_op, *m_sh, *v_pos, *v_scale, *v_op, *v_sh;
    float4 *m_rot, *v_rot;
    uint32_t cap;
};
void gs_launch_adam(GsParams P, GsGradPack G, GsAdamState S, uint32_t n, const float lrs[5], float beta1, float beta2,
                    float eps, float bc1, float bc2, const uint32_t* clean_in, uint32_t* clean_out, uint32_t* skip_words, cudaStream_t st);

// A13 / K10-K11: density control ----------------------------------------------------------------------
void gs_launch_accumulate_stats(GsGradPack G, int stat_mode, uint32_t n, float* accum, uint32_t* count, cudaStream_t st);
struct GsDensifyArgs {
    uint32_t n, iteration, seed, max_gaussians;
    float scene_extent, focal, image_width, avg_depth;
};
// mark: markers[n] in {0 keep,1 prune,2 clone,3 split}; counts[0..2] = pruned, cloned, split (pre-cap)
void gs_launch_densify_mark(GsParams P, const float* accum, const uint32_t* count, GsDensifyArgs a, uint32_t* markers,
                            uint32_t* counts3, cudaStream_t st);
// cap (density_control.mm:360-382) + output offsets: out_off[n+1]; stats[0..2] final pruned/cloned/split, stats[3]=new count
void gs_launch_densify_plan(uint32_t* markers, uint32_t n, uint32_t max_gaussians, uint32_t* counts3, uint32_t* rankbuf,
                            uint32_t* out_off, uint32_t* stats4, void* scan_scratch, cudaStream_t st);
void gs_launch_densify_scatter(GsParams Pin, GsParams Pout, const uint32_t* markers, const uint32_t* out_off,
                               GsDensifyArgs a, cudaStream_t st);

// corrected optimizer-state mode: moments of survivors move with them (src/dst: 46 floats per Gaussian, dst zero-filled)
void gs_launch_permute_moments(const float* src, uint32_t src_cap, float* dst, uint32_t dst_cap, const uint32_t* markers,
                               const uint32_t* out_off, uint32_t n, uint32_t new_n, cudaStream_t st);

// view-parallel gradient sum over NVLink peer memory (peer_reduce.cu).  bufs[q] / flags[q]: rank q's packed gradient buffer and
// its 3 x GS_MAX_PEERS flag words, mapped into this process (q == rank: local memory)
#define GS_MAX_PEERS 8
struct GsPeerSet { float* bufs[GS_MAX_PEERS]; uint32_t* flags[GS_MAX_PEERS]; int rank, world; };
// mc = multicast (NVLS) address of the ranks' buffers or null (plain peer loads / stores, sums in rank order)
void gs_launch_peer_allreduce(const GsPeerSet& P, float* mc, uint32_t n, uint32_t stride, uint32_t epoch, uint32_t* err, uint32_t* host_err,
                              cudaStream_t st);

void gs_tune_dp_mc(int which, int v);   // in-switch reduce: 0 grid, 1 threads per CTA, 2 vectors in flight per thread
// tuning knobs (gs_tune in the C ABI): pixels per thread of the blend kernels (1, 2 or 4)
void gs_tune_fwd_ppt(int v);
bool gs_fwd_writes_visits();   // the forward variant in use records its (entry, warp) visits (PPT 2)
void gs_tune_pair_cull(int v);  // binned pipeline: 1 = bin_fill evaluates the tile-level cull bit, 0 = never set
void gs_tune_fwd_warp(int v);  // forward blend: 1 = warp-autonomous walk, 0 = CTA-cooperative rounds
bool gs_fwd_words_end_with_the_warp();
void gs_tune_bwd_warp(int v);  // masked backward: 1 = warp-autonomous walk, 0 = CTA-cooperative rounds
bool gs_bwd_uses_visits();     // the backward variant in use can consume them (PPT 2)
void gs_tune_bwd_ppt(int v);
