/*
 * gs_abi.h — C ABI of the B200-native tiled Gaussian-splatting hot path (libgs_b200.so).
 *
 * This is the drop-in boundary: it replaces mtl_engine.mm's device plumbing and the metal-cpp dependency of
 * ctaylo41/GaussianSplatting.  The reference's host classes (TiledRasterizer, GPURadixSort32/64, AdamOptimizer,
 * DensityController) keep their names and methods and forward to these entry points
 * (gaussiansplatting_b200/host/).  Reference paths below are relative to
 * /root/reference/GuassianSplatting/.
 *
 * Conventions
 *   - every call returns int: 0 = GS_OK, negative = error; text via gs_last_error().  No exceptions, no aborts.
 *   - one CUDA stream per gs_ctx; calls ENQUEUE and return unless they hand data back to the host
 *     (download / get / apply / anything documented as synchronising).
 *   - host pointers are never retained.  Handles are created/destroyed by the caller; a handle is not
 *     thread-safe; different gs_ctx are independent (one per GPU / rank).
 *   - wire structs are the reference's, byte for byte (static_asserts below).  On the device the library keeps
 *     its own SoA layouts; conversion happens only in upload/download calls.
 */
#ifndef GS_ABI_H
#define GS_ABI_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { GS_OK = 0, GS_ERR_CUDA = -1, GS_ERR_OOM = -2, GS_ERR_ARG = -3, GS_ERR_NCCL = -4, GS_ERR_CAPACITY = -5 };

/* ---- wire structs ------------------------------------------------------------------------------------ */
typedef struct {               /* Gaussian — ply_loader.hpp:14-20, tiled_shaders.metal:11-22 */
    float position[3]; float _pad0;
    float scale[3];    float _pad1;          /* LOG scale */
    float rotation[4];                       /* (w,x,y,z) */
    float opacity;                           /* raw, pre-sigmoid */
    float sh[12];                            /* R 0-3, G 4-7, B 8-11 (DC at 0,4,8) */
    float _pad2[3];
} gs_gaussian;

typedef struct {               /* GaussianGradients — gradients.hpp:11-31 */
    float position_x, position_y, position_z, opacity;
    float scale_x, scale_y, scale_z, _pad1;
    float rotation[4];
    float sh[12];
    float viewspace_grad_x, viewspace_grad_y, _pad2, _pad3;
} gs_gradients;

typedef struct {               /* ProjectedGaussian — tiled_rasterizer.hpp:24-39 */
    float screenPos[2]; float conic[3]; float depth; float opacity; float color[3]; float radius;
    uint32_t tileMinX, tileMinY, tileMaxX, tileMaxY; float _pad1;
    float viewPos_xy[2]; float cov2D[3]; float _pad2;
} gs_projected;

typedef struct { uint32_t start, count; } gs_tile_range;     /* TileRange — tiled_rasterizer.hpp:16-19 */

typedef struct {               /* TiledUniforms — tiled_rasterizer.hpp:42-53 (column-major 4x4, m[c*4+r]) */
    float viewMatrix[16], projectionMatrix[16], viewProjectionMatrix[16];
    float screenSize[2], focalLength[2];
    float cameraPos[4];
    uint32_t numTilesX, numTilesY, numGaussians, _pad2;
} gs_uniforms;

typedef struct {               /* Uniforms — mtl_engine.hpp:30-38 (host layout: simd_float3 occupies 16 B, struct rounded to 240 B) */
    float viewMatrix[16], projectionMatrix[16], viewProjectionMatrix[16];
    float screenSize[2], focalLength[2];
    float cameraPos[4];
    float _pad[4];
} gs_view_uniforms;

typedef struct {               /* what the restated vertex stage hands to the rasteriser, per instance (inspection only) */
    float center[2];           /* quad centre in framebuffer pixels */
    float conic[3]; float radius; float opacity; float color[3];
    uint32_t valid;            /* 0: the instance draws nothing (invalid, behind the camera, depth-clipped, off screen) */
    uint32_t tileMinX, tileMinY, tileMaxX, tileMaxY; uint32_t _pad;
} gs_view_splat;

typedef struct { uint32_t numPruned, numCloned, numSplit; } gs_density_stats;   /* DensityStats — density_control.hpp:13-17 */

#if defined(__cplusplus)
static_assert(sizeof(gs_gaussian) == 112 && sizeof(gs_gradients) == 112 && sizeof(gs_projected) == 88, "wire layout");
static_assert(sizeof(gs_tile_range) == 8 && sizeof(gs_uniforms) == 240, "wire layout");
static_assert(sizeof(gs_view_uniforms) == 240 && sizeof(gs_view_splat) == 64, "wire layout");
#endif

typedef struct gs_ctx gs_ctx;
typedef struct gs_gaussians gs_gaussians;   /* device-resident model state (the reference's gaussianBuffer) */
typedef struct gs_grads gs_grads;           /* device-resident gradient buffer (the reference's gaussianGradients) */
typedef struct gs_image gs_image;           /* device-resident RGBA image (render target / ground-truth texture) */
typedef struct gs_rast gs_rast;             /* TiledRasterizer */
typedef struct gs_sort64 gs_sort64;         /* GPURadixSort64 */
typedef struct gs_sort32 gs_sort32;         /* GPURadixSort32 */
typedef struct gs_adam gs_adam;             /* AdamOptimizer */
typedef struct gs_dens gs_dens;             /* DensityController */
typedef struct gs_viewer gs_viewer;         /* the draw of MTLEngine::render as CUDA compute */
typedef struct gs_dp gs_dp;                 /* view-parallel gradient sum over NVLink peer memory (one per rank) */

/* ---- context (replaces MTL::Device + MTL::CommandQueue, mtl_engine.mm:79-86) ----------------------------- */
int gs_ctx_create(int device, gs_ctx** out);
/* share an existing CUDA stream (e.g. torch's current stream) instead of creating one */
int gs_ctx_create_on_stream(int device, void* cuda_stream, gs_ctx** out);
int gs_ctx_destroy(gs_ctx*);
int gs_ctx_sync(gs_ctx*);                               /* waitUntilCompleted */
const char* gs_last_error(gs_ctx*);                     /* ctx may be NULL: last creation error */
void* gs_ctx_stream(gs_ctx*);
int gs_version(void);
/* process-wide knobs for benchmarking kernel variants; no knob changes a result, only how it is computed:
 *   "fwd_ppt" / "bwd_ppt"  pixels per thread of the blend kernels (2, 4)
 *   "sort_mode"            pair pipeline: 0 two-level binning (default), 1 literal 64-bit pair sort, 2 hoisted radix
 *   "emit_fused"           1 (default) scan + coarse emit + bin histogram as one kernel
 *   "tile_order"           1 (default) blend / viewer-draw CTAs take the tiles longest list first, 0 raster order
 *   "sort_even"            1 (default) one-wave radix passes pick the tile size that evens out the CTAs per SM
 *   "sort_hybrid"          1 (default) hybrid plan for full 64-bit sorts, 0 always the 8-pass LSD plan
 *   "sort_tiny_tiles", "tile_sort_ballot", "sort_api_ballot"   ranking / tile-size experiments (off / as shipped)
 *   "depth_prehist"        1 (default) the projection kernel fills the depth sort's digit histograms
 *   "side_bins"            1 (default) the bins-from-histogram kernel runs on a side stream under the bin pass
 *   "bwd_masks"            1 (default) a training forward records its visit words and the backward walks on them
 *   "bwd_warp"             1 (default) that backward is the warp-autonomous kernel, 0 the CTA-cooperative one
 *   "fwd_warp"             forward blend: 0 CTA-cooperative rounds, 1 (default) warp-autonomous when the frame records
 *                          visit words, 2 warp-autonomous always
 *   "pair_cull"            binned pipeline: 1 the expansion marks (tile, Gaussian) pairs that cannot contribute (bit 31 of the
 *                          list value), 0 (default) it does not: the blend kernels' per-warp footprint test does the work
 *   "dp_mc_grid", "dp_mc_threads", "dp_mc_unroll"   launch shape of the in-switch gradient sum
 * unknown names return GS_ERR_ARG */
int gs_tune(const char* name, int value);

/* raw device memory (MTL::Buffer with no typed role: sort inputs, positions) */
int gs_dev_alloc(gs_ctx*, size_t bytes, void** dptr);
int gs_dev_free(gs_ctx*, void* dptr);
int gs_dev_upload(gs_ctx*, void* dptr, const void* host, size_t bytes);      /* synchronises */
int gs_dev_download(gs_ctx*, void* host, const void* dptr, size_t bytes);    /* synchronises */
int gs_dev_memset(gs_ctx*, void* dptr, int value, size_t bytes);

/* ---- model state --------------------------------------------------------------------------------------- */
int gs_gaussians_create(gs_ctx*, uint32_t capacity, gs_gaussians** out);
int gs_gaussians_destroy(gs_gaussians*);
int gs_gaussians_upload(gs_gaussians*, const gs_gaussian* host, uint32_t n);   /* sets count = n; grows capacity */
int gs_gaussians_download(gs_gaussians*, gs_gaussian* host, uint32_t n);       /* synchronises */
uint32_t gs_gaussians_count(const gs_gaussians*);
uint32_t gs_gaussians_capacity(const gs_gaussians*);
/* opacity reset, mtl_engine.mm:1179-1185: opacity = min(opacity, max_raw) */
int gs_gaussians_clamp_opacity(gs_gaussians*, float max_raw);
/* xyz + pad float4 positions for the viewer sort (updatePositionBuffer, mtl_engine.mm:1028-1036); dptr: n*16 B */
int gs_gaussians_positions(gs_gaussians*, void* dptr_float4);

/* ---- gradient buffer ------------------------------------------------------------------------------------ */
int gs_grads_create(gs_ctx*, uint32_t capacity, gs_grads** out);
/* The gradient buffer in CALLER-OWNED device memory: 16 planes of `capacity` floats at d_mem (capacity a multiple of 64, d_mem
 * 16-byte aligned); zeroed here, never freed or regrown by the library (GS_ERR_CAPACITY where gs_grads_create's buffer would
 * grow).  No counterpart in the reference (one device): it lets view-parallel training keep the all-reduce payload in a
 * symmetric, multicast-mapped allocation that the NVSwitch can reduce in place (gs_dp_connect_ptrs). */
int gs_grads_create_external(gs_ctx*, uint32_t capacity, void* d_mem, gs_grads** out);
int gs_grads_destroy(gs_grads*);
int gs_grads_zero(gs_grads*, uint32_t n);
int gs_grads_upload(gs_grads*, const gs_gradients* host, uint32_t n);     /* SH-rest gradients must be 0 (never produced) */
int gs_grads_download(gs_grads*, gs_gradients* host, uint32_t n);         /* synchronises */
/* 0 (default): columns 14-15 hold raw summed view-space x,y (reference semantics, one view per buffer).
 * 1: columns 14-15 hold sum over views of min(|viewspace|,1) and the number of views counted, so that the
 *    statistic survives accumulation over several views / an all-reduce over ranks. */
int gs_grads_set_stat_mode(gs_grads*, int per_view_stats);
/* the packed fp32 device buffer (16 planes of `stride` floats) — hand this to ncclAllReduce / torch.distributed */
int gs_grads_packed(gs_grads*, void** dptr, uint64_t* num_floats, uint32_t* stride);

/* ---- view-parallel training: the per-step gradient sum over the ranks of one node ----------------
 * The reference trains on one device (mtl_engine.mm:1085-1093); this is the only exchange step the multi-GPU path adds.  It runs
 * as this library's own kernels over NVLink peer memory (csrc/peer_reduce.cu): every rank maps the others' packed gradient
 * buffers through CUDA IPC, reduces its slice in rank order and stores the sum into every rank's buffer, so all ranks end with
 * bit-identical sums.
 *   1. every rank: gs_dp_create, gs_dp_export(grads, blob)          (blob: GS_DP_BLOB_BYTES)
 *   2. exchange the blobs by any means (torch.distributed.all_gather_object, MPI, a file) — rank order
 *   3. every rank: gs_dp_connect(all blobs)
 *   4. per step, on every rank, after the backward calls: gs_dp_allreduce_grads(grads, n)   (enqueues; planes [0, n) are summed)
 * The gradient buffer must not be reallocated after export (create it at its final capacity).  gs_dp_check synchronises and
 * reports GS_ERR_NCCL if a peer did not arrive within 5 s (the kernels give up instead of hanging the device). */
#define GS_DP_BLOB_BYTES 192
#define GS_DP_MAX_RANKS 8
int gs_dp_create(gs_ctx*, int rank, int world, gs_dp** out);
int gs_dp_destroy(gs_dp*);
int gs_dp_export(gs_dp*, gs_grads*, void* blob_out);
int gs_dp_connect(gs_dp*, const void* blobs_in_rank_order);
/* The same connection from pointers already valid in this process (a symmetric-memory rendezvous done by the caller, e.g.
 * torch.distributed._symmetric_memory; replaces steps 2-3): peer_bufs[q] = rank q's gradient buffer (peer_bufs[rank] must be
 * the buffer of `grads`), peer_flags[q] = 24 u32 flag words in rank q's memory (this rank's are zeroed here; run a host
 * barrier over the ranks before the first all-reduce), multicast_buf = the multicast (NVLS) address mapping ALL ranks'
 * buffers, or NULL.  With a multicast address gs_dp_allreduce_grads reduces IN THE NVSWITCH: rank r issues multimem.ld_reduce
 * over slice r and multimem.st broadcasts the sum, so a GPU moves 1x the payload per direction instead of 2 (g-1)/g x; every
 * rank still ends with bit-identical sums (one reduction per element, broadcast), in the switch's order of addition. */
int gs_dp_connect_ptrs(gs_dp*, gs_grads* grads, void* const* peer_bufs, void* const* peer_flags, void* multicast_buf);
int gs_dp_allreduce_grads(gs_dp*, gs_grads*, uint32_t n);
int gs_dp_check(gs_dp*);

/* ---- images ------------------------------------------------------------------------------------------- */
int gs_image_create(gs_ctx*, uint32_t width, uint32_t height, gs_image** out);
int gs_image_destroy(gs_image*);
int gs_image_upload_rgba32f(gs_image*, const float* host);                /* enqueue on the ctx stream (host should be pinned) */
int gs_image_upload_rgba8(gs_image*, const uint8_t* host);
int gs_image_download_rgba32f(gs_image*, float* host);                    /* synchronises */
int gs_image_download_rgba8(gs_image*, uint8_t* host);                    /* synchronises; RGBA8Unorm rounding */
/* Asynchronous RGBA8 transfers on the context's copy streams (the reference's textures are device-visible shared
 * memory, mtl_engine.mm:715-727 / image_loader.mm:24; a discrete GPU has to copy, and these let the copy of the NEXT
 * view's ground truth and of the LAST render overlap the blend kernels).  `host` must be pinned and stay valid until
 * gs_image_wait().  Compute calls that touch the image are ordered after a pending upload on the device; a compute
 * call that overwrites it is also ordered after a pending download.  One transfer in flight per image. */
int gs_image_upload_rgba8_async(gs_image*, const uint8_t* host);
int gs_image_download_rgba8_async(gs_image*, uint8_t* host);
int gs_image_wait(gs_image*);                                             /* host blocks until the transfer in flight is done */
void* gs_image_devptr(gs_image*);                                          /* float4[width*height] */

/* ---- logging loss: MTLEngine::computeLoss (mtl_engine.mm:769-853; computeL1Loss / computeSSIM / computeCombinedLoss /
 * reduceLoss, shaders.metal:319-511).  (1 - lambda_dssim) * L1 + lambda_dssim * D-SSIM (11x11 Gaussian window, sigma 1.5,
 * mean RGB, clamp to edge), mean over pixels.  Off the gradient path.  flags: GS_BWD_QUANTIZE_RGBA8 reads the rendered
 * image as the reference's RGBA8Unorm texture would hold it.  Synchronises. */
int gs_loss_compute(gs_ctx*, gs_image* rendered, gs_image* ground_truth, float lambda_dssim, int flags, float* loss);
/* The same loss for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Calls no cudaMalloc /
 * cudaFree / allocating routine and does not synchronise: it enqueues two kernels on the context's stream and the mean loss
 * is written to *d_loss_out, ONE float of caller-owned DEVICE memory (e.g. from gs_dev_alloc).  Its scratch — one double per
 * 16x16 tile, summed in a fixed order, so the result is run-to-run deterministic — is a static device array of
 * GS_LOSS_NOALLOC_MAX_TILES (32768) entries = 256 KiB that every GPU loading the library holds until exit: images of more
 * tiles (beyond e.g. 4096x2048 or 3840x2160) are rejected with GS_ERR_CAPACITY.  All calls on a device share that array and it
 * keeps the last call's contents (each call overwrites every slot it reads, nothing needs clearing): calls that overlap on the
 * GPU, i.e. on different contexts / streams of one device, race.  The first call on a device resolves the array's address
 * (cudaGetSymbolAddress): make it once before starting a graph capture. */
#define GS_LOSS_NOALLOC_MAX_TILES 32768
int gs_loss_compute_noalloc(gs_ctx*, gs_image* rendered, gs_image* ground_truth, float lambda_dssim, int flags, float* d_loss_out);

/* ---- TiledRasterizer (tiled_rasterizer.hpp:56-124) ------------------------------------------------------ */
int gs_rast_create(gs_ctx*, uint32_t max_gaussians, gs_rast** out);
int gs_rast_destroy(gs_rast*);
/* forward(queue, gaussianBuffer, gaussianCount, uniforms, outputTexture) — tiled_rasterizer.mm:275-672.
 * numTilesX/Y/numGaussians in *u are overwritten from the image size and the live count (:308-311).
 * Synchronises once internally (pair count read-back, as the reference does at :379-380,454). */
int gs_rast_forward(gs_rast*, gs_gaussians*, const gs_uniforms* u, gs_image* out);
/* backward(queue, gaussianBuffer, gradientBuffer, gaussianCount, uniforms, rendered, groundTruth) —
 * tiled_rasterizer.mm:675-722.  Zeroes the gradient buffer first (:687) unless GS_BWD_ACCUMULATE is set.
 * Must follow a forward with the same uniforms (tiled_rasterizer.hpp:94-106). */
enum { GS_BWD_QUANTIZE_RGBA8 = 1, GS_BWD_ACCUMULATE = 2 };
int gs_rast_backward(gs_rast*, gs_gaussians*, gs_grads*, const gs_uniforms* u, gs_image* rendered, gs_image* gt, int flags);
/* inspection of the state forward leaves behind (parity tests); all synchronise and copy to host */
enum { GS_RAST_PROJECTED = 0,   /* gs_projected[n] */
       GS_RAST_KEYS = 1,        /* uint64[pairs]  pre-sort   */
       GS_RAST_VALUES = 2,      /* uint32[pairs]  pre-sort   */
       GS_RAST_SORTED_KEYS = 3, /* uint64[pairs] */
       GS_RAST_SORTED_VALUES = 4,/* uint32[pairs] */
       GS_RAST_TILE_RANGES = 5, /* gs_tile_range[tiles] */
       GS_RAST_LAST_IDX = 6,    /* uint32[w*h] */
       GS_RAST_TFINAL = 7 };    /* float[w*h] */
int gs_rast_get(gs_rast*, int which, void* host, size_t bytes);
/* per-stage CUDA-event timing of the last forward/backward (replaces the chrono prints of tiled_rasterizer.mm:639-671):
 * ms9 = preprocess, scan, pair-count read-back, emit, sort, tile ranges, blend fwd, blend bwd (+memsets), per-Gaussian bwd */
int gs_rast_set_profiling(gs_rast*, int on);
int gs_rast_get_timings(gs_rast*, float* ms9, int include_backward);
/* E = list entries the forward walk visits, C = contributing (pixel, Gaussian) evaluations of the last forward */
int gs_rast_count_evals(gs_rast*, uint64_t* E, uint64_t* C);
uint64_t gs_rast_pair_count(const gs_rast*);
uint64_t gs_rast_coarse_pair_count(const gs_rast*);                      /* (8x8-tile bin, Gaussian) pairs of the last forward in the default binned pipeline; 0 in the sort pipelines */
int gs_rast_sort_passes(const gs_rast*);                /* digit passes the last forward ran */
/* Opt-in: the stages before the blend run on a high-priority stream of this rasterizer, so that a caller alternating two
 * rasterizers over independent frames gets frame i+1's ordering stage under frame i's blend.  Writes to the Gaussians queued on
 * the context stream after this rasterizer's last call are NOT waited for: call again after such work (re-arms the wait), or
 * leave it off (default).  No counterpart in the reference (one command queue). */
int gs_rast_set_overlap(gs_rast*, int on);
/* 1 if the last backward took its per-pixel decisions from the visit words the forward recorded (every backward of a
 * rasterizer but its first, gs_tune "bwd_masks" = 1), 0 if it re-derived them (diagnostic of this library; no counterpart) */
int gs_rast_backward_used_visits(const gs_rast*);

/* ---- GPURadixSort64 / GPURadixSort32 (gpu_sort.hpp:14-107) ---------------------------------------------- */
int gs_sort64_create(gs_ctx*, uint64_t max_elements, gs_sort64** out);
int gs_sort64_destroy(gs_sort64*);
/* sort(queue, keysIn, valuesIn, numElements): device pointers; inputs are left untouched; stable; bits [begin,end) */
int gs_sort64_sort(gs_sort64*, const uint64_t* d_keys, const uint32_t* d_vals, uint64_t n, int begin_bit, int end_bit);
const uint64_t* gs_sort64_sorted_keys(gs_sort64*);      /* borrowed, valid until the next sort */
const uint32_t* gs_sort64_sorted_values(gs_sort64*);
/* which plan the last sort ran on: 1 = prefix passes + local pass (radix_sort.cu, full 64-bit sorts of spread keys), 0 = the
 * 8-pass LSD plan (always for sub-ranges and small inputs; the device-side fallback otherwise).  Synchronises. */
int gs_sort64_last_plan(gs_sort64*, int* hybrid);
int gs_sort32_create(gs_ctx*, uint64_t max_elements, gs_sort32** out);
int gs_sort32_destroy(gs_sort32*);
/* sort(queue, positionBuffer, cameraPos, n): float4 positions -> back-to-front indices (computeDepths, sort.metal:36-72) */
int gs_sort32_sort_depth(gs_sort32*, const void* d_positions_float4, const float cam[3], uint64_t n);
int gs_sort32_sort(gs_sort32*, const uint32_t* d_keys, const uint32_t* d_vals, uint64_t n);
const uint32_t* gs_sort32_sorted_keys(gs_sort32*);
const uint32_t* gs_sort32_sorted_indices(gs_sort32*);

/* ---- viewer draw: MTLEngine::render (mtl_engine.mm:401-609) without the window ---------------------------
 * vertexShader / fragmentShader (shaders.metal:160-317) and the render pipeline's blend state (mtl_engine.mm:354-362:
 * source ONE, destination ONE_MINUS_SRC_ALPHA, clear colour white :583) as CUDA compute into a headless gs_image: one
 * quad per Gaussian, drawn in the order of d_sorted_indices (the buffer GPURadixSort32::sort returns, far first), each
 * pixel centre inside the quad shaded and blended in that order.  uniforms->screenSize must equal the image size.
 * GS_VIEW_QUANTIZE_RGBA8 rounds the target to RGBA8Unorm after every blend, as the reference's attachment stores it
 * (mtl_engine.mm:351); default is fp32 accumulation.  Synchronises once (pair count), like gs_rast_forward. */
int gs_viewer_create(gs_ctx*, uint32_t max_gaussians, gs_viewer** out);
int gs_viewer_destroy(gs_viewer*);
enum { GS_VIEW_QUANTIZE_RGBA8 = 1, GS_VIEW_COUNT_FRAGMENTS = 2 };
int gs_viewer_render(gs_viewer*, gs_gaussians*, const uint32_t* d_sorted_indices, const gs_view_uniforms* u, gs_image* out, int flags);
enum { GS_VIEW_SPLATS = 0,        /* gs_view_splat[n] */
       GS_VIEW_SORTED_TILES = 1,  /* uint32[pairs] tile id of every (tile, instance) pair, tile-major, draw order inside a tile */
       GS_VIEW_SORTED_VALUES = 2, /* uint32[pairs] Gaussian index of the pair */
       GS_VIEW_TILE_RANGES = 3 }; /* gs_tile_range[tiles] */
int gs_viewer_get(gs_viewer*, int which, void* host, size_t bytes);      /* synchronises */
uint64_t gs_viewer_pair_count(const gs_viewer*);
/* fragments shaded / blended by the last render that had GS_VIEW_COUNT_FRAGMENTS set; synchronises */
int gs_viewer_fragments(gs_viewer*, uint64_t* shaded, uint64_t* blended);

/* ---- AdamOptimizer (optimizer.hpp:23-95) ---------------------------------------------------------------- */
int gs_adam_create(gs_ctx*, uint32_t num_gaussians, gs_adam** out);
int gs_adam_destroy(gs_adam*);
/* step(queue, gaussians, gradients, lr_position, lr_scale, lr_rotation, lr_opacity, lr_sh): ++timestep, then the
 * fused update over the LIVE count (SURVEY note A), beta1 .9, beta2 .999, eps 1e-8 (optimizer.mm:228-296) */
int gs_adam_step(gs_adam*, gs_gaussians*, gs_grads*, const float lrs[5]);
int gs_adam_reset(gs_adam*);
int gs_adam_resize_if_needed(gs_adam*, uint32_t new_num_gaussians);
int gs_adam_reset_opacity_momentum(gs_adam*);
int gs_adam_reset_scale_momentum(gs_adam*);
int gs_adam_reset_state_for_new(gs_adam*, uint32_t start_idx);
uint32_t gs_adam_timestep(const gs_adam*);
int gs_adam_set_timestep(gs_adam*, uint32_t t);          /* restoring a checkpoint (the reference keeps `timestep` as a plain member, optimizer.hpp:64) */
/* moments in the reference's layout: group 0 position [n*3], 1 scale [n*3], 2 rotation [n*4], 3 opacity [n], 4 sh [n*12];
 * which_moment 0 = m, 1 = v.  Synchronise. */
int gs_adam_download(gs_adam*, int group, int which_moment, float* host, uint32_t n);
int gs_adam_upload(gs_adam*, int group, int which_moment, const float* host, uint32_t n);

/* ---- DensityController (density_control.hpp:20-62) ------------------------------------------------------ */
int gs_dens_create(gs_ctx*, uint32_t max_gaussians, gs_dens** out);      /* reference: 1,500,000 */
int gs_dens_destroy(gs_dens*);
int gs_dens_set_scene_extent(gs_dens*, float extent);
int gs_dens_set_seed(gs_dens*, uint32_t seed);
int gs_dens_accumulate(gs_dens*, gs_grads*, uint32_t n);
int gs_dens_reset_accumulator(gs_dens*, uint32_t n);
/* apply(...): may replace the device storage inside *g and changes its count; synchronises (new count read-back) */
int gs_dens_apply(gs_dens*, gs_gaussians* g, uint32_t iteration, float focal_length, float image_width,
                  float avg_depth, gs_density_stats* stats);
/* Opt-in correction of SURVEY note D (the reference re-orders the Gaussians at densify but leaves the Adam moments where they
 * were, mtl_engine.mm:1159-1167): call right after gs_dens_apply INSTEAD of resize_if_needed + reset_state_for_new.  Moments of
 * kept Gaussians and of the first copy of a clone follow them to their new index; the second copy of a clone, both halves of a
 * split and nothing else start from zero.  old_n = the count before the apply; the new count is read from the model. */
int gs_adam_permute_after_densify(gs_adam*, gs_dens*, gs_gaussians*, uint32_t old_n);
int gs_dens_download(gs_dens*, float* accum, uint32_t* count, uint32_t* markers, uint32_t n);   /* any may be NULL */
int gs_dens_upload(gs_dens*, const float* accum, const uint32_t* count, uint32_t n);

#ifdef __cplusplus
}
#endif
#endif /* GS_ABI_H */
