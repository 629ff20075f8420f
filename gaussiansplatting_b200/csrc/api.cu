// api.cu — the C ABI of include/gs_abi.h: handles, device memory, and the host-side orchestration that the
// reference keeps in tiled_rasterizer.mm / gpu_sort.mm / optimizer.mm / density_control.mm.
#include "../../include/gs_abi.h"
#include "gs_kernels.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

struct gs_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    uint32_t* h_pinned = nullptr;   // small pinned read-back area
    cudaStream_t xfer_in = nullptr, xfer_out = nullptr;   // copy streams of the *_async image transfers (created on first use)
    double* d_loss = nullptr;                              // accumulator of gs_loss_compute
    cudaStream_t xfer_meta = nullptr;                     // small read-backs (pair count) that must not queue behind an image
};
static std::string g_create_err;
static int g_tile_sort_hint = 0;                    // in-frame tile-id passes: match.any measured faster
static int g_tile_order = 1;                        // blend CTAs take the tiles longest list first (tile_order_kernel)
static int g_bwd_masks = 1;          // backward takes its per-slot decisions from the forward's visit words (0: re-derives them)
static int g_side_bins = 1;          // bins-from-histogram kernel on the side stream, under the bin-sort pass
static int g_depth_prehist = 1;      // K1 fills the depth sort's histograms (0: the sort runs its own histogram kernel)
static int g_emit_fused = 1;                        // binned pipeline: scan + coarse emit + bin histogram as one kernel
static int g_sort_api_hint = GS_SORT_HINT_DISTINCT; // GPURadixSort32/64 on caller keys: assume no structure
static void gs_set_sort_mode(int v);
#define GS_PROF_EVENTS 11
#define GS_MAX_BINS 4096u                            // binned mode up to 512x512 tiles (8192x8192 pixels); larger frames use the hoisted radix sort

#define GS_SETDEV(c) cudaSetDevice((c)->device)
#define GS_CUDA(c, expr)                                                                          \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            (c)->err = std::string(#expr) + ": " + cudaGetErrorString(e__);                       \
            return e__ == cudaErrorMemoryAllocation ? GS_ERR_OOM : GS_ERR_CUDA;                   \
        }                                                                                         \
    } while (0)
#define GS_ARG(c, cond, msg)                                                                      \
    do { if (!(cond)) { (c)->err = std::string("bad argument: ") + msg; return GS_ERR_ARG; } } while (0)
#define GS_LAUNCH_CHECK(c) GS_CUDA(c, cudaGetLastError())

template <typename T>
static int dalloc(gs_ctx* c, T** p, size_t count) {
    *p = nullptr;
    if (count == 0) count = 1;
    GS_CUDA(c, cudaMalloc((void**)p, count * sizeof(T)));
    return GS_OK;
}
#define GS_TRY(expr) do { int r__ = (expr); if (r__ != GS_OK) return r__; } while (0)

extern "C" {

int gs_version(void) { return 100; }

int gs_tune(const char* name, int value) {
    if (!name) return GS_ERR_ARG;
    if (!strcmp(name, "fwd_ppt")) { gs_tune_fwd_ppt(value); return GS_OK; }
    if (!strcmp(name, "bwd_ppt")) { gs_tune_bwd_ppt(value); return GS_OK; }
    if (!strcmp(name, "tile_order")) { g_tile_order = value ? 1 : 0; return GS_OK; }
    if (!strcmp(name, "sort_mode")) { gs_set_sort_mode(value); return GS_OK; }
    if (!strcmp(name, "sort_tiny_tiles")) { gs_tune_sort_tiny(value); return GS_OK; }
    if (!strcmp(name, "sort_hybrid")) { gs_tune_sort_hybrid(value); return GS_OK; }
    if (!strcmp(name, "sort_even")) { gs_tune_sort_even(value); return GS_OK; }
    if (!strcmp(name, "emit_fused")) { g_emit_fused = value ? 1 : 0; return GS_OK; }
    if (!strcmp(name, "dp_mc_grid")) { gs_tune_dp_mc(0, value); return GS_OK; }
    if (!strcmp(name, "dp_mc_threads")) { gs_tune_dp_mc(1, value); return GS_OK; }
    if (!strcmp(name, "dp_mc_unroll")) { gs_tune_dp_mc(2, value); return GS_OK; }
    if (!strcmp(name, "bwd_masks")) { g_bwd_masks = value ? 1 : 0; return GS_OK; }
    if (!strcmp(name, "bwd_warp")) { gs_tune_bwd_warp(value); return GS_OK; }
    if (!strcmp(name, "fwd_warp")) { gs_tune_fwd_warp(value); return GS_OK; }
    if (!strcmp(name, "pair_cull")) { gs_tune_pair_cull(value); return GS_OK; }
    if (!strcmp(name, "side_bins")) { g_side_bins = value ? 1 : 0; return GS_OK; }
    if (!strcmp(name, "depth_prehist")) { g_depth_prehist = value ? 1 : 0; return GS_OK; }
    if (!strcmp(name, "tile_sort_ballot")) { g_tile_sort_hint = value ? GS_SORT_HINT_DISTINCT : 0; return GS_OK; }
    if (!strcmp(name, "sort_api_ballot")) { g_sort_api_hint = value ? GS_SORT_HINT_DISTINCT : 0; return GS_OK; }
    return GS_ERR_ARG;
}

static int ctx_create_common(int device, cudaStream_t st, bool own, gs_ctx** out) {
    if (!out) return GS_ERR_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || device < 0 || device >= count) {
        g_create_err = e != cudaSuccess ? std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e)
                                        : "no such CUDA device";
        return GS_ERR_CUDA;
    }
    gs_ctx* c = new (std::nothrow) gs_ctx();
    if (!c) return GS_ERR_OOM;
    c->device = device;
    cudaSetDevice(device);
    if (own) {
        e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { g_create_err = cudaGetErrorString(e); delete c; return GS_ERR_CUDA; }
        c->own_stream = true;
    } else c->stream = st;
    e = cudaMallocHost((void**)&c->h_pinned, 256);
    if (e != cudaSuccess) { g_create_err = cudaGetErrorString(e); delete c; return GS_ERR_CUDA; }
    *out = c;
    return GS_OK;
}
int gs_ctx_create(int device, gs_ctx** out) { return ctx_create_common(device, nullptr, true, out); }
int gs_ctx_create_on_stream(int device, void* s, gs_ctx** out) { return ctx_create_common(device, (cudaStream_t)s, false, out); }
int gs_ctx_destroy(gs_ctx* c) {
    if (!c) return GS_OK;
    GS_SETDEV(c);
    cudaStreamSynchronize(c->stream);
    if (c->xfer_in) { cudaStreamSynchronize(c->xfer_in); cudaStreamDestroy(c->xfer_in); }
    if (c->xfer_out) { cudaStreamSynchronize(c->xfer_out); cudaStreamDestroy(c->xfer_out); }
    if (c->xfer_meta) { cudaStreamSynchronize(c->xfer_meta); cudaStreamDestroy(c->xfer_meta); }
    if (c->own_stream) cudaStreamDestroy(c->stream);
    cudaFreeHost(c->h_pinned);
    cudaFree(c->d_loss);
    delete c;
    return GS_OK;
}
int gs_ctx_sync(gs_ctx* c) {
    GS_SETDEV(c);
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->xfer_in) GS_CUDA(c, cudaStreamSynchronize(c->xfer_in));
    if (c->xfer_out) GS_CUDA(c, cudaStreamSynchronize(c->xfer_out));
    if (c->xfer_meta) GS_CUDA(c, cudaStreamSynchronize(c->xfer_meta));
    return GS_OK;
}
const char* gs_last_error(gs_ctx* c) { return c ? c->err.c_str() : g_create_err.c_str(); }
void* gs_ctx_stream(gs_ctx* c) { return (void*)c->stream; }

int gs_dev_alloc(gs_ctx* c, size_t bytes, void** p) { GS_SETDEV(c); GS_CUDA(c, cudaMalloc(p, bytes ? bytes : 1)); return GS_OK; }
int gs_dev_free(gs_ctx* c, void* p) { GS_SETDEV(c); GS_CUDA(c, cudaStreamSynchronize(c->stream)); GS_CUDA(c, cudaFree(p)); return GS_OK; }
int gs_dev_upload(gs_ctx* c, void* d, const void* h, size_t bytes) {
    GS_SETDEV(c);
    GS_CUDA(c, cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_dev_download(gs_ctx* c, void* h, const void* d, size_t bytes) {
    GS_SETDEV(c);
    GS_CUDA(c, cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_dev_memset(gs_ctx* c, void* d, int v, size_t bytes) { GS_SETDEV(c); GS_CUDA(c, cudaMemsetAsync(d, v, bytes, c->stream)); return GS_OK; }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// model state
// ------------------------------------------------------------------------------------------------------
struct gs_gaussians {
    gs_ctx* c;
    float* base = nullptr;      // 19 float planes (pos 3, scale 3, opacity 1, sh 12) of stride cap, then rot float4[cap]
    uint32_t cap = 0, n = 0;
    void* staging = nullptr;    // AoS staging for upload/download
    size_t staging_bytes = 0;
    bool sh_rest_dirty = true;  // the nine non-DC SH planes may hold values outside Adam's [-2,2] clamp (set by every writer)
    GsParams view() const {
        GsParams P;
        P.cap = cap;
        P.pos = base; P.scale = base + 3 * (size_t)cap; P.opacity = base + 6 * (size_t)cap; P.sh = base + 7 * (size_t)cap;
        P.rot = reinterpret_cast<float4*>(base + 19 * (size_t)cap);
        return P;
    }
};
static int params_alloc(gs_ctx* c, uint32_t cap, float** base) {
    cap = (cap + 63u) & ~63u;   // keeps the float4 plane 16-B aligned
    return dalloc(c, base, (size_t)cap * 23);
}
static int staging_reserve(gs_ctx* c, void** p, size_t* have, size_t need) {
    if (*have >= need) return GS_OK;
    if (*p) { GS_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(*p); *p = nullptr; *have = 0; }
    GS_CUDA(c, cudaMalloc(p, need));
    *have = need;
    return GS_OK;
}

extern "C" {

int gs_gaussians_create(gs_ctx* c, uint32_t capacity, gs_gaussians** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_gaussians* g = new (std::nothrow) gs_gaussians();
    if (!g) return GS_ERR_OOM;
    g->c = c;
    g->cap = (std::max(capacity, 64u) + 63u) & ~63u;
    int r = params_alloc(c, g->cap, &g->base);
    if (r != GS_OK) { delete g; return r; }
    *out = g;
    return GS_OK;
}
int gs_gaussians_destroy(gs_gaussians* g) {
    if (!g) return GS_OK;
    cudaSetDevice(g->c->device);
    cudaStreamSynchronize(g->c->stream);
    cudaFree(g->base); cudaFree(g->staging);
    delete g;
    return GS_OK;
}
static int gaussians_reserve(gs_gaussians* g, uint32_t n) {   // contents are NOT preserved
    if (n <= g->cap) return GS_OK;
    gs_ctx* c = g->c;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(g->base); g->base = nullptr;
    g->cap = ((uint32_t)(n * 1.25) + 63u) & ~63u;
    return params_alloc(c, g->cap, &g->base);
}
int gs_gaussians_upload(gs_gaussians* g, const gs_gaussian* host, uint32_t n) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    GS_ARG(c, host || n == 0, "host");
    GS_TRY(gaussians_reserve(g, n));
    GS_TRY(staging_reserve(c, &g->staging, &g->staging_bytes, (size_t)n * sizeof(gs_gaussian)));
    GS_CUDA(c, cudaMemcpyAsync(g->staging, host, (size_t)n * sizeof(gs_gaussian), cudaMemcpyHostToDevice, c->stream));
    gs_launch_params_from_aos(g->staging, n, g->view(), c->stream);
    g->sh_rest_dirty = true;
    GS_LAUNCH_CHECK(c);
    g->n = n;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_gaussians_download(gs_gaussians* g, gs_gaussian* host, uint32_t n) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    GS_ARG(c, n <= g->n, "n exceeds the live count");
    GS_TRY(staging_reserve(c, &g->staging, &g->staging_bytes, (size_t)n * sizeof(gs_gaussian)));
    gs_launch_params_to_aos(g->view(), n, g->staging, c->stream);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaMemcpyAsync(host, g->staging, (size_t)n * sizeof(gs_gaussian), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
uint32_t gs_gaussians_count(const gs_gaussians* g) { return g->n; }
uint32_t gs_gaussians_capacity(const gs_gaussians* g) { return g->cap; }
int gs_gaussians_clamp_opacity(gs_gaussians* g, float max_raw) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    gs_launch_clamp_opacity(g->view(), g->n, max_raw, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// gradient buffer, images
// ------------------------------------------------------------------------------------------------------
struct gs_grads {
    gs_ctx* c;
    float* g = nullptr;         // 16 planes of `stride` floats
    uint32_t cap = 0, stride = 0;
    int stat_mode = 0;
    bool external = false;      // g points into caller-owned memory (gs_grads_create_external): never freed, never regrown
    void* staging = nullptr;
    size_t staging_bytes = 0;
    GsGradPack view() const { GsGradPack G; G.g = g; G.stride = stride; return G; }
};
struct gs_image {
    gs_ctx* c;
    uint32_t w = 0, h = 0;
    float4* px = nullptr;
    uint32_t* rgba8 = nullptr;  // staging for 8-bit transfers
    // *_async transfers run on the ctx's copy streams: ev_use = last compute-stream use of the image, ev_xfer = end of the
    // transfer in flight; pending: 0 none, 1 upload (readers and writers on the compute stream wait), 2 download (writers wait)
    cudaEvent_t ev_use = nullptr, ev_xfer = nullptr;
    int pending = 0;
};
static int image_before_read(gs_image* im, cudaStream_t st) {
    if (im->pending == 1) { GS_CUDA(im->c, cudaStreamWaitEvent(st, im->ev_xfer, 0)); im->pending = 0; }
    return GS_OK;
}
static int image_before_write(gs_image* im, cudaStream_t st) {
    if (im->pending) { GS_CUDA(im->c, cudaStreamWaitEvent(st, im->ev_xfer, 0)); im->pending = 0; }
    return GS_OK;
}
static int image_after_use(gs_image* im, cudaStream_t st) {
    if (im->ev_use) GS_CUDA(im->c, cudaEventRecord(im->ev_use, st));
    return GS_OK;
}
// first async call on an image: create its events; everything enqueued on the compute stream so far counts as a use
static int image_async_begin(gs_image* im, cudaStream_t* xfer) {
    gs_ctx* c = im->c;
    if (!*xfer) GS_CUDA(c, cudaStreamCreateWithFlags(xfer, cudaStreamNonBlocking));
    if (!im->ev_use) {
        GS_CUDA(c, cudaEventCreateWithFlags(&im->ev_use, cudaEventDisableTiming));
        GS_CUDA(c, cudaEventCreateWithFlags(&im->ev_xfer, cudaEventDisableTiming));
        GS_CUDA(c, cudaEventRecord(im->ev_use, c->stream));
    }
    GS_CUDA(c, cudaStreamWaitEvent(*xfer, im->ev_use, 0));
    if (im->pending) GS_CUDA(c, cudaStreamWaitEvent(*xfer, im->ev_xfer, 0));   // one staging buffer per image
    return GS_OK;
}

// contents are NOT preserved on growth.  The caller's capacity is honoured exactly at creation (the 16 planes of `stride`
// floats are the all-reduce payload of view-parallel training: slack there is NVLink traffic); later growth adds 25 %.
static int grads_reserve(gs_grads* g, uint32_t n, bool exact = false) {
    if (n <= g->cap) return GS_OK;
    gs_ctx* c = g->c;
    if (g->external) { c->err = "gradient buffer in caller-provided memory is too small (it cannot grow)"; return GS_ERR_CAPACITY; }
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(g->g); g->g = nullptr;
    g->cap = ((exact ? n : (uint32_t)(n * 1.25)) + 63u) & ~63u;
    g->stride = g->cap;
    return dalloc(c, &g->g, (size_t)g->cap * 16);
}

extern "C" {

int gs_gaussians_positions(gs_gaussians* g, void* dptr) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    GS_ARG(c, dptr, "dptr");
    gs_launch_positions(g->view(), g->n, (float4*)dptr, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}

int gs_grads_create(gs_ctx* c, uint32_t capacity, gs_grads** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_grads* g = new (std::nothrow) gs_grads();
    if (!g) return GS_ERR_OOM;
    g->c = c;
    int r = grads_reserve(g, std::max(capacity, 64u), true);
    if (r != GS_OK) { delete g; return r; }
    *out = g;
    return GS_OK;
}
int gs_grads_destroy(gs_grads* g) {
    if (!g) return GS_OK;
    cudaSetDevice(g->c->device);
    cudaStreamSynchronize(g->c->stream);
    if (!g->external) cudaFree(g->g);
    cudaFree(g->staging);
    delete g;
    return GS_OK;
}
// The gradient buffer in memory the CALLER owns: 16 planes of `capacity` floats at d_mem (capacity a multiple of 64, d_mem
// 16-byte aligned).  For view-parallel training the caller passes a symmetric allocation that every rank of the node has
// mapped and that has a multicast (NVLS) address, so that the per-Gaussian chain kernel writes the all-reduce payload in
// place (gs_dp_connect_ptrs).  No counterpart in the reference (single device).
int gs_grads_create_external(gs_ctx* c, uint32_t capacity, void* d_mem, gs_grads** out) {
    GS_SETDEV(c);
    GS_ARG(c, out && d_mem, "null argument");
    GS_ARG(c, capacity >= 64u && capacity % 64u == 0u && ((uintptr_t)d_mem & 15u) == 0u, "capacity must be a multiple of 64 and the memory 16-byte aligned");
    gs_grads* g = new (std::nothrow) gs_grads();
    if (!g) return GS_ERR_OOM;
    g->c = c; g->g = static_cast<float*>(d_mem); g->cap = capacity; g->stride = capacity; g->external = true;
    if (cudaMemsetAsync(g->g, 0, (size_t)capacity * 16 * sizeof(float), c->stream) != cudaSuccess) { delete g; c->err = "cudaMemsetAsync on the caller's memory"; return GS_ERR_CUDA; }
    *out = g;
    return GS_OK;
}
int gs_grads_zero(gs_grads* g, uint32_t n) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    GS_TRY(grads_reserve(g, n));
    if (n) GS_CUDA(c, cudaMemset2DAsync(g->g, (size_t)g->stride * sizeof(float), 0, (size_t)n * sizeof(float), 16, c->stream));   // the live rows only
    return GS_OK;
}
int gs_grads_upload(gs_grads* g, const gs_gradients* host, uint32_t n) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    GS_TRY(grads_reserve(g, n));
    GS_TRY(staging_reserve(c, &g->staging, &g->staging_bytes, (size_t)n * sizeof(gs_gradients)));
    GS_CUDA(c, cudaMemcpyAsync(g->staging, host, (size_t)n * sizeof(gs_gradients), cudaMemcpyHostToDevice, c->stream));
    gs_launch_grads_from_aos(g->staging, n, g->view(), c->stream);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_grads_download(gs_grads* g, gs_gradients* host, uint32_t n) {
    gs_ctx* c = g->c; GS_SETDEV(c);
    GS_ARG(c, n <= g->cap, "n exceeds capacity");
    GS_TRY(staging_reserve(c, &g->staging, &g->staging_bytes, (size_t)n * sizeof(gs_gradients)));
    gs_launch_grads_to_aos(g->view(), n, g->staging, c->stream);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaMemcpyAsync(host, g->staging, (size_t)n * sizeof(gs_gradients), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_grads_set_stat_mode(gs_grads* g, int m) { g->stat_mode = m ? 1 : 0; return GS_OK; }
int gs_grads_packed(gs_grads* g, void** dptr, uint64_t* num_floats, uint32_t* stride) {
    if (dptr) *dptr = g->g;
    if (num_floats) *num_floats = (uint64_t)g->stride * 16;
    if (stride) *stride = g->stride;
    return GS_OK;
}

int gs_image_create(gs_ctx* c, uint32_t w, uint32_t h, gs_image** out) {
    GS_SETDEV(c);
    GS_ARG(c, out && w && h, "image size");
    gs_image* im = new (std::nothrow) gs_image();
    if (!im) return GS_ERR_OOM;
    im->c = c; im->w = w; im->h = h;
    int r = dalloc(c, &im->px, (size_t)w * h);
    if (r == GS_OK) r = dalloc(c, &im->rgba8, (size_t)w * h);
    if (r != GS_OK) { cudaFree(im->px); delete im; return r; }
    *out = im;
    return GS_OK;
}
int gs_image_destroy(gs_image* im) {
    if (!im) return GS_OK;
    cudaSetDevice(im->c->device);
    cudaStreamSynchronize(im->c->stream);
    if (im->pending) cudaEventSynchronize(im->ev_xfer);
    if (im->ev_use) { cudaEventDestroy(im->ev_use); cudaEventDestroy(im->ev_xfer); }
    cudaFree(im->px); cudaFree(im->rgba8);
    delete im;
    return GS_OK;
}
int gs_image_upload_rgba32f(gs_image* im, const float* host) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    GS_TRY(image_before_write(im, c->stream));
    GS_CUDA(c, cudaMemcpyAsync(im->px, host, (size_t)im->w * im->h * 16, cudaMemcpyHostToDevice, c->stream));
    return GS_OK;
}
int gs_image_upload_rgba8(gs_image* im, const uint8_t* host) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    GS_TRY(image_before_write(im, c->stream));
    GS_CUDA(c, cudaMemcpyAsync(im->rgba8, host, (size_t)im->w * im->h * 4, cudaMemcpyHostToDevice, c->stream));
    gs_launch_rgba8_to_rgba32f(im->rgba8, im->w * im->h, im->px, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}
int gs_image_download_rgba32f(gs_image* im, float* host) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    GS_TRY(image_before_read(im, c->stream));
    GS_CUDA(c, cudaMemcpyAsync(host, im->px, (size_t)im->w * im->h * 16, cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_image_download_rgba8(gs_image* im, uint8_t* host) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    GS_TRY(image_before_write(im, c->stream));      // the 8-bit staging buffer is shared with a transfer in flight
    gs_launch_rgba32f_to_rgba8(im->px, im->w * im->h, im->rgba8, c->stream);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaMemcpyAsync(host, im->rgba8, (size_t)im->w * im->h * 4, cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
// Asynchronous RGBA8 transfers on the context's copy streams, ordered against the compute stream by events, so that the
// next view's ground truth and the last render cross PCIe while the blend kernels run.  `host` must be pinned and stay
// valid until gs_image_wait().  A compute call that touches the image waits (on the device) for a pending upload; a
// compute call that overwrites it also waits for a pending download.
int gs_image_upload_rgba8_async(gs_image* im, const uint8_t* host) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    GS_ARG(c, host, "null host buffer");
    GS_TRY(image_async_begin(im, &c->xfer_in));
    GS_CUDA(c, cudaMemcpyAsync(im->rgba8, host, (size_t)im->w * im->h * 4, cudaMemcpyHostToDevice, c->xfer_in));
    gs_launch_rgba8_to_rgba32f(im->rgba8, im->w * im->h, im->px, c->xfer_in);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaEventRecord(im->ev_xfer, c->xfer_in));
    im->pending = 1;
    return GS_OK;
}
int gs_image_download_rgba8_async(gs_image* im, uint8_t* host) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    GS_ARG(c, host, "null host buffer");
    GS_TRY(image_async_begin(im, &c->xfer_out));
    gs_launch_rgba32f_to_rgba8(im->px, im->w * im->h, im->rgba8, c->xfer_out);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaMemcpyAsync(host, im->rgba8, (size_t)im->w * im->h * 4, cudaMemcpyDeviceToHost, c->xfer_out));
    GS_CUDA(c, cudaEventRecord(im->ev_xfer, c->xfer_out));
    im->pending = 2;
    return GS_OK;
}
// block the host until the image's transfer in flight (if any) has completed
int gs_image_wait(gs_image* im) {
    gs_ctx* c = im->c; GS_SETDEV(c);
    if (im->pending) { GS_CUDA(c, cudaEventSynchronize(im->ev_xfer)); im->pending = 0; }
    return GS_OK;
}
void* gs_image_devptr(gs_image* im) { return im->px; }

// MTLEngine::computeLoss (mtl_engine.mm:769-853): (1 - lambda) * L1 + lambda * D-SSIM, mean over the pixels.  Synchronises,
// as the reference does (waitUntilCompleted + read of the shared buffer).
int gs_loss_compute(gs_ctx* c, gs_image* rendered, gs_image* gt, float lambda_dssim, int flags, float* loss) {
    GS_SETDEV(c);
    GS_ARG(c, rendered && gt && loss, "null argument");
    GS_ARG(c, rendered->w == gt->w && rendered->h == gt->h, "image sizes differ");
    GS_TRY(image_before_read(rendered, c->stream));
    GS_TRY(image_before_read(gt, c->stream));
    if (!c->d_loss) GS_TRY(dalloc(c, &c->d_loss, 1));
    GS_CUDA(c, cudaMemsetAsync(c->d_loss, 0, sizeof(double), c->stream));
    gs_launch_loss(rendered->px, gt->px, rendered->w, rendered->h, lambda_dssim, (flags & GS_BWD_QUANTIZE_RGBA8) ? 1 : 0, c->d_loss, c->stream);
    GS_LAUNCH_CHECK(c);
    GS_TRY(image_after_use(rendered, c->stream));
    GS_TRY(image_after_use(gt, c->stream));
    double total = 0.0;
    GS_CUDA(c, cudaMemcpyAsync(&total, c->d_loss, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    *loss = (float)(total / ((double)rendered->w * rendered->h));
    return GS_OK;
}

// The same loss without any allocation (no counterpart in the reference; added for
// callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).  Enqueues two
// kernels on the context's stream and returns; the mean loss lands in *d_loss_out, ONE float of caller-owned DEVICE memory.
// Scratch: the static device array of csrc/loss.cu (GS_LOSS_NOALLOC_MAX_TILES tiles of 16x16 pixels).
int gs_loss_compute_noalloc(gs_ctx* c, gs_image* rendered, gs_image* gt, float lambda_dssim, int flags, float* d_loss_out) {
    GS_SETDEV(c);
    GS_ARG(c, rendered && gt && d_loss_out, "null argument");
    GS_ARG(c, rendered->w == gt->w && rendered->h == gt->h, "image sizes differ");
    GS_TRY(image_before_read(rendered, c->stream));
    GS_TRY(image_before_read(gt, c->stream));
    const char* msg = nullptr;
    const int rc = gs_launch_loss_noalloc(c->device, rendered->px, gt->px, rendered->w, rendered->h, lambda_dssim,
                                          (flags & GS_BWD_QUANTIZE_RGBA8) ? 1 : 0, d_loss_out, c->stream, &msg);
    if (rc != GS_OK) { c->err = msg ? msg : "gs_loss_compute_noalloc failed"; return rc; }
    GS_TRY(image_after_use(rendered, c->stream));
    GS_TRY(image_after_use(gt, c->stream));
    return GS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// TiledRasterizer — host orchestration of tiled_rasterizer.mm:275-722
// ------------------------------------------------------------------------------------------------------
struct gs_rast {
    gs_ctx* c;
    uint32_t cap_g = 0, cap_tiles = 0, cap_pix = 0;
    uint64_t cap_pairs = 0;
    GsProjected proj{};
    uint32_t* meta = nullptr;
    void* scan_scratch = nullptr; size_t scan_scratch_bytes = 0;
    void* emit_scratch = nullptr;       // ticket + look-back status words of the fused scan / coarse emit
    // pair buffers.  64-bit mode: keys_in/kbuf hold u64 (tile|depth) keys.  Hoisted mode: the same allocations hold
    // u32 tile ids (first half used).
    uint64_t *keys_in = nullptr, *kbuf[2] = {nullptr, nullptr};
    uint32_t *vals_in = nullptr, *vbuf[2] = {nullptr, nullptr};
    void* sort_scratch = nullptr;
    // hoisted mode: depth order of the Gaussians (the low-word passes of the LSD sort, run on N keys instead of P)
    uint32_t *iota = nullptr, *dk_buf[2] = {nullptr, nullptr}, *perm_buf[2] = {nullptr, nullptr};
    void* dsort_scratch = nullptr;
    uint2* ranges = nullptr;
    // binned mode (binning.cu): per-tile totals, the bins' runs in the coarse-sorted list, chunk prefix per bin, per-chunk tile counts
    uint32_t* tile_count = nullptr; uint2* bin_ranges = nullptr; uint32_t* cstart = nullptr; uint32_t* chunk_counts = nullptr;
    uint32_t* last_idx = nullptr;
    float* tfinal = nullptr;
    float* acc = nullptr;
    void* staging = nullptr; size_t staging_bytes = 0;
    // state left behind by forward for backward / inspection (tiled_rasterizer.hpp:94-106)
    uint32_t W = 0, H = 0, ntx = 0, nty = 0, n = 0;
    uint64_t pairs = 0;
    uint64_t coarse_hint = 0;           // coarse pair count of the previous frame (sizes the chunk grids)
    int mode = 0;                       // pair pipeline of the last forward: 0 binned, 1 full 64-bit pair sort, 2 hoisted radix
    int result = 0, passes = 0;
    const uint32_t* sorted_vals = nullptr;   // per-tile lists (bit 31 = GS_CULL_BIT)
    const uint32_t* sorted_tiles = nullptr;  // hoisted mode: sorted u32 tile ids
    const uint64_t* sorted_keys = nullptr;   // 64-bit mode: sorted keys
    cudaEvent_t ev_pairs = nullptr;     // recorded after the pair count landed in pinned memory
    cudaEvent_t ev_scan = nullptr;      // recorded on the compute stream after the scan (the copy-out stream waits on it)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;   // side-stream fork / join around the bins-from-histogram kernel
    // optional per-stage CUDA-event timing (replaces the reference's chrono prints, tiled_rasterizer.mm:639-671)
    bool prof = false;
    bool acc_clean = false;       // the moment buffer is all-zero (K8 clears what K7 wrote)
    // forward's visit words for the backward (blend_fwd.cu / blend_bwd.cu): four planes (one per warp) of vis_cap uint2 words.  Allocated by the first
    // backward call (render-only users never pay), regrown with the pair buffers; vis_valid = the last forward filled it.
    uint2* vis = nullptr; uint64_t vis_cap = 0; bool want_vis = false, vis_valid = false, bwd_used_vis = false, vis_warp_extent = false;
    bool depth_prehist = false;   // this frame's K1 filled the depth sort's digit histograms (GS_SORT_HINT_PREHIST)
    cudaEvent_t ev[GS_PROF_EVENTS] = {};
    unsigned long long* counters = nullptr;
    // gs_rast_set_overlap: the stages before the blend run on `ord` (a high-priority stream of this rasterizer) instead of the context
    // stream.  ev_ord = "ordering done" (the context stream waits on it before the blend), ev_free = position of the context stream
    // at the end of this rasterizer's last forward / backward call: from there on its buffers may be overwritten.
    cudaStream_t ord = nullptr;
    cudaEvent_t ev_ord = nullptr, ev_free = nullptr;
    bool free_valid = false;
    cudaStream_t ord_stream() const { return (ord && mode == 0 && !prof) ? ord : c->stream; }
    void mark(int i) { if (prof) cudaEventRecord(ev[i], c->stream); }
};

__global__ void iota_kernel(uint32_t* v, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = i;
}

static void rast_free_gaussian_arrays(gs_rast* r) {
    cudaFree(r->proj.rec); cudaFree(r->proj.bbox); cudaFree(r->proj.aux); cudaFree(r->proj.covc); cudaFree(r->proj.offs);
    cudaFree(r->proj.cnt); cudaFree(r->proj.dkey); cudaFree(r->proj.ccnt);
    cudaFree(r->acc); cudaFree(r->scan_scratch); cudaFree(r->emit_scratch); cudaFree(r->iota); cudaFree(r->dsort_scratch);
    for (int i = 0; i < 2; i++) { cudaFree(r->dk_buf[i]); cudaFree(r->perm_buf[i]); r->dk_buf[i] = nullptr; r->perm_buf[i] = nullptr; }
    r->proj.rec = nullptr; r->proj.bbox = nullptr; r->proj.aux = nullptr; r->proj.covc = nullptr; r->proj.offs = nullptr;
    r->proj.cnt = nullptr; r->proj.dkey = nullptr; r->proj.ccnt = nullptr;
    r->acc = nullptr; r->scan_scratch = nullptr; r->emit_scratch = nullptr; r->iota = nullptr; r->dsort_scratch = nullptr;
}
static int rast_reserve_gaussians(gs_rast* r, uint32_t n) {
    if (n <= r->cap_g) return GS_OK;
    gs_ctx* c = r->c;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    rast_free_gaussian_arrays(r);
    const uint32_t cap = ((uint32_t)(n * 1.25) + 63u) & ~63u;
    GS_TRY(dalloc(c, &r->proj.rec, (size_t)cap * 3));
    GS_TRY(dalloc(c, &r->proj.bbox, cap));
    GS_TRY(dalloc(c, &r->proj.aux, cap));
    GS_TRY(dalloc(c, &r->proj.covc, cap));
    GS_TRY(dalloc(c, &r->proj.offs, (size_t)cap + 1));
    GS_TRY(dalloc(c, &r->proj.cnt, cap));
    GS_TRY(dalloc(c, &r->proj.dkey, cap));
    GS_TRY(dalloc(c, &r->proj.ccnt, cap));
    GS_TRY(dalloc(c, &r->acc, (size_t)cap * 12));
    r->acc_clean = false;
    GS_TRY(dalloc(c, &r->iota, cap));
    for (int i = 0; i < 2; i++) { GS_TRY(dalloc(c, &r->dk_buf[i], cap)); GS_TRY(dalloc(c, &r->perm_buf[i], cap)); }
    GS_CUDA(c, cudaMalloc(&r->dsort_scratch, gs_sort_scratch_bytes(cap, 4)));
    iota_kernel<<<(cap + 255) / 256, 256, 0, c->stream>>>(r->iota, cap);
    r->scan_scratch_bytes = gs_scan_scratch_bytes(cap);
    GS_CUDA(c, cudaMalloc(&r->scan_scratch, r->scan_scratch_bytes));
    GS_CUDA(c, cudaMalloc(&r->emit_scratch, gs_emit_coarse_scratch_bytes(cap)));
    r->cap_g = cap;
    return GS_OK;
}
static int rast_reserve_image(gs_rast* r, uint32_t W, uint32_t H) {   // ensureBufferCapacity, tiled_rasterizer.mm:210-239
    gs_ctx* c = r->c;
    const uint32_t ntx = (W + GS_TILE - 1) / GS_TILE, nty = (H + GS_TILE - 1) / GS_TILE;
    if (ntx * nty > r->cap_tiles) {
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(r->ranges); cudaFree(r->proj.tile_flags); cudaFree(r->tile_count); cudaFree(r->bin_ranges); cudaFree(r->cstart);
        r->cap_tiles = ntx * nty;
        GS_TRY(dalloc(c, &r->ranges, r->cap_tiles));
        r->proj.tile_order = nullptr;
        GS_TRY(dalloc(c, &r->proj.tile_flags, (size_t)r->cap_tiles * 2));   // flags, then the blend CTAs' tile order
        GS_TRY(dalloc(c, &r->tile_count, (size_t)r->cap_tiles + 1));     // + the bin-scan CTAs' completion ticket (self-resetting)
        GS_CUDA(c, cudaMemsetAsync(r->tile_count + r->cap_tiles, 0, sizeof(uint32_t), c->stream));
        GS_TRY(dalloc(c, &r->bin_ranges, (size_t)r->cap_tiles + 2));      // bins <= tiles
        GS_TRY(dalloc(c, &r->cstart, (size_t)r->cap_tiles + 2));
    }
    if (W * H > r->cap_pix) {
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(r->last_idx); cudaFree(r->tfinal);
        r->cap_pix = W * H;
        GS_TRY(dalloc(c, &r->last_idx, r->cap_pix));
        GS_TRY(dalloc(c, &r->tfinal, r->cap_pix));
    }
    r->W = W; r->H = H; r->ntx = ntx; r->nty = nty;
    return GS_OK;
}
// the buffer the coming forward blend writes its visits to (null: none); marks them valid for the backward of this frame
static uint2* rast_vis_for_forward(gs_rast* r) {
    r->vis_valid = r->vis != nullptr && g_bwd_masks && gs_fwd_writes_visits() && r->vis_cap >= r->cap_pairs;
    r->vis_warp_extent = gs_fwd_words_end_with_the_warp();
    return r->vis_valid ? r->vis : nullptr;
}
static int rast_reserve_vis(gs_rast* r) {      // visit words for cap_pairs pair slots x 4 warps (32 B per pair slot)
    if (r->vis && r->vis_cap >= r->cap_pairs) return GS_OK;
    gs_ctx* c = r->c;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(r->vis); r->vis = nullptr; r->vis_valid = false;
    r->vis_cap = r->cap_pairs;
    return dalloc(c, &r->vis, (size_t)r->vis_cap * 4);
}
static int rast_reserve_pairs(gs_rast* r, uint64_t need) {   // ensurePairsCapacity, tiled_rasterizer.mm:242-272 (no 100 M clamp: note C)
    if (need <= r->cap_pairs) return GS_OK;
    gs_ctx* c = r->c;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(r->keys_in); cudaFree(r->vals_in); cudaFree(r->sort_scratch); cudaFree(r->chunk_counts);
    for (int i = 0; i < 2; i++) { cudaFree(r->kbuf[i]); cudaFree(r->vbuf[i]); r->kbuf[i] = nullptr; r->vbuf[i] = nullptr; }
    r->keys_in = nullptr; r->vals_in = nullptr; r->sort_scratch = nullptr; r->chunk_counts = nullptr;
    // an eighth of headroom over the request: the next view of an orbit rarely has exactly as many pairs, and every growth event
    // costs a stream drain, a dozen reallocations and a redone frame
    const uint64_t cap = std::max<uint64_t>(need + need / 8, (uint64_t)(r->cap_pairs * 1.5));
    GS_TRY(dalloc(c, &r->keys_in, cap));
    GS_TRY(dalloc(c, &r->vals_in, cap));
    for (int i = 0; i < 2; i++) { GS_TRY(dalloc(c, &r->kbuf[i], cap)); GS_TRY(dalloc(c, &r->vbuf[i], cap)); }
    GS_CUDA(c, cudaMalloc(&r->sort_scratch, gs_sort_scratch_bytes(cap, 8)));
    GS_TRY(dalloc(c, &r->chunk_counts, gs_bin_max_chunks(cap, GS_MAX_BINS) * (65 + 512)));    // 64 counts per chunk, then the chunk -> bin table, then 8 x 64 coverage words per chunk
    r->cap_pairs = cap;
    if (r->want_vis) GS_TRY(rast_reserve_vis(r));
    return GS_OK;
}

extern "C" {

int gs_rast_create(gs_ctx* c, uint32_t max_gaussians, gs_rast** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_rast* r = new (std::nothrow) gs_rast();
    if (!r) return GS_ERR_OOM;
    r->c = c;
    int rc = dalloc(c, &r->meta, GS_META_COUNT);
    if (rc == GS_OK) rc = rast_reserve_gaussians(r, std::max(max_gaussians, 64u));
    if (rc == GS_OK) rc = rast_reserve_pairs(r, (uint64_t)std::max(max_gaussians, 64u) * 8);   // AVG_TILES_PER_GAUSSIAN, tiled_rasterizer.hpp:80
    if (rc == GS_OK && cudaEventCreateWithFlags(&r->ev_pairs, cudaEventDisableTiming) != cudaSuccess) rc = GS_ERR_CUDA;
    if (rc == GS_OK && cudaEventCreateWithFlags(&r->ev_scan, cudaEventDisableTiming) != cudaSuccess) rc = GS_ERR_CUDA;
    if (rc != GS_OK) { gs_rast_destroy(r); return rc; }
    *out = r;
    return GS_OK;
}
int gs_rast_destroy(gs_rast* r) {
    if (!r) return GS_OK;
    cudaSetDevice(r->c->device);
    cudaStreamSynchronize(r->c->stream);
    rast_free_gaussian_arrays(r);
    cudaFree(r->meta); cudaFree(r->keys_in); cudaFree(r->vals_in); cudaFree(r->sort_scratch);
    for (int i = 0; i < 2; i++) { cudaFree(r->kbuf[i]); cudaFree(r->vbuf[i]); }
    cudaFree(r->ranges); cudaFree(r->proj.tile_flags); cudaFree(r->last_idx); cudaFree(r->tfinal); cudaFree(r->staging);
    cudaFree(r->tile_count); cudaFree(r->bin_ranges); cudaFree(r->cstart); cudaFree(r->chunk_counts);
    cudaFree(r->counters); cudaFree(r->vis);
    if (r->ev_pairs) cudaEventDestroy(r->ev_pairs);
    if (r->ev_scan) cudaEventDestroy(r->ev_scan);
    if (r->ev_fork) cudaEventDestroy(r->ev_fork);
    if (r->ev_join) cudaEventDestroy(r->ev_join);
    if (r->ev_ord) cudaEventDestroy(r->ev_ord);
    if (r->ev_free) cudaEventDestroy(r->ev_free);
    if (r->ord) cudaStreamDestroy(r->ord);
    for (int i = 0; i < GS_PROF_EVENTS; i++) if (r->ev[i]) cudaEventDestroy(r->ev[i]);
    delete r;
    return GS_OK;
}

// Full 64-bit pair sort (the reference's structure: duplicate keys, sort P pairs on tile|depth).  Needs the pair count on
// the host before the pair kernels can be sized, i.e. one mid-frame synchronisation like the reference's (:379-380,454).
static int forward_pairs64(gs_rast* r, gs_gaussians* g, const gs_uniforms* u, gs_image* out) {
    gs_ctx* c = r->c;
    cudaStream_t st = c->stream;
    const uint32_t n = r->n, W = r->W, H = r->H;
    r->mark(2);
    gs_launch_exclusive_scan(r->proj.cnt, nullptr, r->proj.offs, n, r->meta + GS_META_TOTAL_PAIRS, r->scan_scratch, st);
    r->mark(3);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaMemcpyAsync(c->h_pinned, r->meta, GS_META_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    GS_CUDA(c, cudaStreamSynchronize(st));                            // the reference's read of pairCounter (:379-380,454)
    const uint64_t P = c->h_pinned[GS_META_TOTAL_PAIRS];
    const uint32_t kor = c->h_pinned[GS_META_KEY_OR], kand = c->h_pinned[GS_META_KEY_AND];
    if (P >= (1ull << 30)) { c->err = "pair count exceeds 2^30"; return GS_ERR_CAPACITY; }
    GS_TRY(rast_reserve_pairs(r, P));
    r->pairs = P;
    const uint32_t tiles = r->ntx * r->nty;
    r->sorted_keys = r->kbuf[0]; r->sorted_vals = r->vbuf[0]; r->sorted_tiles = nullptr;
    if (P > 0) {
        gs_launch_emit_pairs(r->proj, r->proj.offs, n, r->ntx, r->keys_in, r->vals_in, (uint32_t)r->cap_pairs, st);
        r->mark(4);
        int tile_bits = 0;
        while ((1u << tile_bits) < tiles) tile_bits++;
        const uint64_t varying = (uint64_t)(kor ^ kand) | (tile_bits ? (((1ull << tile_bits) - 1ull) << 32) : 0ull);
        const GsSortPlan plan = gs_sort_plan(0, 32 + tile_bits, varying);
        r->passes = plan.num_passes;
        r->result = gs_launch_sort64(r->keys_in, r->vals_in, r->kbuf, r->vbuf, P, plan, 0, r->sort_scratch, st);
        r->sorted_keys = r->kbuf[r->result]; r->sorted_vals = r->vbuf[r->result];
    } else r->mark(4);
    r->mark(5);
    gs_launch_tile_ranges(r->sorted_keys, (uint32_t)P, tiles, nullptr, r->ranges, st);
    r->mark(6);
    r->proj.tile_order = g_tile_order ? r->proj.tile_flags + r->cap_tiles : nullptr;
    gs_launch_blend_forward(r->proj, r->sorted_vals, r->ranges, W, H, r->ntx, r->nty, out->px, r->last_idx, r->tfinal, 0, rast_vis_for_forward(r), (size_t)r->vis_cap, st);
    r->mark(7);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}

// Hoisted mode (default).  The LSD radix sort on (tile<<32 | depth) is: the low-word (depth) passes, then the
// high-word (tile) passes, each stable.  Every duplicate of a Gaussian carries the same depth word, so the low-word
// passes are run on the N Gaussians BEFORE duplication (N keys instead of P ~ 7N), pairs are emitted in that order, and
// only the ceil(log2 tiles) tile bits are sorted on the P pairs (2 passes of u32+u32 instead of 6 of u64+u32).  The
// resulting order is, element for element, the stable sort of the reference's 64-bit keys.  Nothing waits for the
// host: pair kernels read the pair count from device memory and their grids are sized from the buffer capacity; the
// host blocks only on the event after the scan (while the GPU keeps working) to validate the capacity.
static int forward_hoisted(gs_rast* r, gs_gaussians* g, const gs_uniforms* u, gs_image* out) {
    gs_ctx* c = r->c;
    cudaStream_t st = c->stream;
    const uint32_t n = r->n, W = r->W, H = r->H;
    const uint32_t tiles = r->ntx * r->nty;
    int tile_bits = 0;
    while ((1u << tile_bits) < tiles) tile_bits++;
    // depth order of the Gaussians: keys = depth bits (bit 31 is always set: depth > 0.1), values = index
    const GsSortPlan dplan = gs_sort_plan(0, 31, ~0ull);
    const int dres = gs_launch_sort32(r->proj.dkey, r->iota, r->dk_buf, r->perm_buf, n, nullptr, dplan,
                                      GS_SORT_HINT_DISTINCT | (r->depth_prehist ? GS_SORT_HINT_PREHIST : 0), r->dsort_scratch, st);
    const uint32_t* perm = r->perm_buf[dres];
    r->mark(2);
    gs_launch_exclusive_scan(r->proj.cnt, perm, r->proj.offs, n, r->meta + GS_META_TOTAL_PAIRS, r->scan_scratch, st);
    // the pair count travels to the host on the copy-out stream, so the compute stream never waits for a D2H copy
    if (!c->xfer_meta) GS_CUDA(c, cudaStreamCreateWithFlags(&c->xfer_meta, cudaStreamNonBlocking));
    GS_CUDA(c, cudaEventRecord(r->ev_scan, st));
    GS_CUDA(c, cudaStreamWaitEvent(c->xfer_meta, r->ev_scan, 0));
    GS_CUDA(c, cudaMemcpyAsync(c->h_pinned, r->meta, GS_META_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->xfer_meta));
    GS_CUDA(c, cudaEventRecord(r->ev_pairs, c->xfer_meta));
    r->mark(3);
    const GsSortPlan tplan = gs_sort_plan_even(0, tile_bits);
    r->passes = dplan.num_passes + tplan.num_passes;
    for (int attempt = 0; attempt < 2; attempt++) {
        uint32_t* tk_in = reinterpret_cast<uint32_t*>(r->keys_in);
        uint32_t* tk[2] = {reinterpret_cast<uint32_t*>(r->kbuf[0]), reinterpret_cast<uint32_t*>(r->kbuf[1])};
        const uint32_t cap = (uint32_t)std::min<uint64_t>(r->cap_pairs, 0x3FFFFFFFull);
        const uint32_t* d_total = r->meta + GS_META_TOTAL_PAIRS;
        gs_launch_emit_tile_pairs(r->proj, perm, r->proj.offs, n, r->ntx, tk_in, r->vals_in, cap, st);
        r->mark(4);
        r->result = gs_launch_sort32(tk_in, r->vals_in, tk, r->vbuf, cap, d_total, tplan, g_tile_sort_hint, r->sort_scratch, st);
        r->sorted_tiles = tplan.num_passes ? tk[r->result] : tk_in;
        r->sorted_vals = tplan.num_passes ? r->vbuf[r->result] : r->vals_in;
        r->sorted_keys = nullptr;
        r->mark(5);
        gs_launch_tile_ranges32(r->sorted_tiles, cap, d_total, tiles, r->ranges, st);
        r->mark(6);
        r->proj.tile_order = g_tile_order ? r->proj.tile_flags + r->cap_tiles : nullptr;
        gs_launch_blend_forward(r->proj, r->sorted_vals, r->ranges, W, H, r->ntx, r->nty, out->px, r->last_idx, r->tfinal, 0, rast_vis_for_forward(r), (size_t)r->vis_cap, st);
        r->mark(7);
        GS_LAUNCH_CHECK(c);
        if (attempt == 0) GS_CUDA(c, cudaEventSynchronize(r->ev_pairs));   // the GPU is still busy with the kernels above
        const uint64_t P = c->h_pinned[GS_META_TOTAL_PAIRS];
        if (P >= (1ull << 30)) { c->err = "pair count exceeds 2^30"; return GS_ERR_CAPACITY; }
        r->pairs = P;
        if (P <= r->cap_pairs) break;
        GS_TRY(rast_reserve_pairs(r, P));                              // rare: grow (synchronises) and redo the pair stages
    }
    return GS_OK;
}


// Binned mode (default): depth order on the N Gaussians, coarse (bin, Gaussian) pairs in that order, ONE stable sort of
// the coarse pairs by bin id, then per-bin expansion straight into the per-tile lists and ranges (binning.cu).  Same
// lists and ranges as the other two modes; no sort over the P pairs.
static int forward_binned(gs_rast* r, gs_gaussians* g, const gs_uniforms* u, gs_image* out) {
    gs_ctx* c = r->c;
    cudaStream_t blend_st = c->stream;
    cudaStream_t st = r->ord_stream();                                // everything up to the expansion; the blend stays on the context stream
    const uint32_t n = r->n, W = r->W, H = r->H;
    const uint32_t tiles = r->ntx * r->nty, nbins = gs_bin_count(r->ntx, r->nty);
    int bin_bits = 0;
    while ((1u << bin_bits) < nbins) bin_bits++;
    const GsSortPlan dplan = gs_sort_plan(0, 31, ~0ull);
    const int dres = gs_launch_sort32(r->proj.dkey, r->iota, r->dk_buf, r->perm_buf, n, nullptr, dplan,
                                      GS_SORT_HINT_DISTINCT | (r->depth_prehist ? GS_SORT_HINT_PREHIST : 0), r->dsort_scratch, st);
    const uint32_t* perm = r->perm_buf[dres];
    r->mark(2);
    uint32_t* d_coarse = r->meta + GS_META_COARSE_PAIRS;
    uint32_t* d_total = r->meta + GS_META_TOTAL_PAIRS;
    const GsSortPlan bplan = gs_sort_plan_even(0, bin_bits);
    // fused: the scan, the coarse emit and the bin sort's digit histograms are one kernel (binning.cu)
    const bool fused = g_emit_fused && bplan.num_passes >= 1 && bplan.num_passes <= 2;
    if (!fused) gs_launch_exclusive_scan(r->proj.ccnt, perm, r->proj.offs, n, d_coarse, r->scan_scratch, st);
    r->mark(3);
    r->passes = dplan.num_passes + bplan.num_passes;
    if (!c->xfer_meta) GS_CUDA(c, cudaStreamCreateWithFlags(&c->xfer_meta, cudaStreamNonBlocking));
    for (int attempt = 0; attempt < 4; attempt++) {
        uint32_t* bk_in = reinterpret_cast<uint32_t*>(r->keys_in);
        uint32_t* bk[2] = {reinterpret_cast<uint32_t*>(r->kbuf[0]), reinterpret_cast<uint32_t*>(r->kbuf[1])};
        const uint32_t cap = (uint32_t)std::min<uint64_t>(r->cap_pairs, 0x3FFFFFFFull);
        if (fused) {
            uint32_t* ghist = gs_sort_prepare(cap, bplan, r->sort_scratch, st);
            gs_launch_emit_coarse_scan(r->proj, perm, n, r->ntx, bk_in, r->vals_in, cap, d_coarse, r->emit_scratch, bplan, ghist, st);
        } else gs_launch_emit_coarse(r->proj, perm, r->proj.offs, n, r->ntx, bk_in, r->vals_in, cap, st);
        r->mark(4);
        // one-pass bin sort: its digit histogram IS the bin histogram and is final now, so the bins' runs and the chunk map are
        // computed on the side stream while the sort pass runs (a one-CTA kernel off the critical path)
        const bool side_bins = fused && bplan.num_passes == 1 && g_side_bins;
        uint32_t* chunk_bin = r->chunk_counts + gs_bin_max_chunks(r->cap_pairs, GS_MAX_BINS) * 64;
        if (side_bins) {
            if (!r->ev_fork) { GS_CUDA(c, cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming)); GS_CUDA(c, cudaEventCreateWithFlags(&r->ev_join, cudaEventDisableTiming)); }
            GS_CUDA(c, cudaEventRecord(r->ev_fork, st));
            GS_CUDA(c, cudaStreamWaitEvent(c->xfer_meta, r->ev_fork, 0));
            gs_launch_bins_from_hist(gs_sort_pass_histogram(r->sort_scratch, 0), r->ntx, r->nty, r->bin_ranges, r->cstart, chunk_bin, c->xfer_meta);
            GS_CUDA(c, cudaEventRecord(r->ev_join, c->xfer_meta));
        }
        const int res = gs_launch_sort32(bk_in, r->vals_in, bk, r->vbuf, cap, d_coarse, bplan, g_tile_sort_hint | (fused ? GS_SORT_HINT_PREHIST : 0),
                                         r->sort_scratch, st);
        if (side_bins) GS_CUDA(c, cudaStreamWaitEvent(st, r->ev_join, 0));
        const uint32_t* sorted_bins = bplan.num_passes ? bk[res] : bk_in;
        const uint32_t* cvals = bplan.num_passes ? r->vbuf[res] : r->vals_in;
        uint32_t* lists = r->vbuf[bplan.num_passes ? (res ^ 1) : 0];
        r->mark(5);
        const uint32_t* bin_hist = side_bins ? GS_BINS_READY : bplan.num_passes == 1 ? gs_sort_pass_histogram(r->sort_scratch, 0) : nullptr;
        r->proj.tile_order = g_tile_order ? r->proj.tile_flags + r->cap_tiles : nullptr;   // written by bin_fill's CTA 0
        if (!bin_hist) gs_launch_tile_ranges32(sorted_bins, cap, d_coarse, nbins, r->bin_ranges, st);
        gs_launch_bin_expand(r->proj, cvals, bin_hist, r->bin_ranges, r->ntx, r->nty, r->cstart, r->chunk_counts, r->tile_count, r->ranges,
                             d_total, lists, cap, r->coarse_hint ? (uint32_t)(r->coarse_hint / 256 * 9 / 8) + nbins + 1u : 0u,
                             chunk_bin, r->chunk_counts + gs_bin_max_chunks(r->cap_pairs, GS_MAX_BINS) * 65, r->tile_count + r->cap_tiles, st);
        r->result = 0;
        r->sorted_tiles = nullptr; r->sorted_keys = nullptr; r->sorted_vals = lists;
        // the pair counts travel to the host on their own copy stream; the compute stream carries on with the blend
        GS_CUDA(c, cudaEventRecord(r->ev_scan, st));
        GS_CUDA(c, cudaStreamWaitEvent(c->xfer_meta, r->ev_scan, 0));
        GS_CUDA(c, cudaMemcpyAsync(c->h_pinned, r->meta, GS_META_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->xfer_meta));
        GS_CUDA(c, cudaEventRecord(r->ev_pairs, c->xfer_meta));
        r->mark(6);
        r->proj.tile_order = g_tile_order ? r->proj.tile_flags + r->cap_tiles : nullptr;
        if (st != blend_st) {                                         // the lists are final: hand over to the context stream
            GS_CUDA(c, cudaEventRecord(r->ev_ord, st));
            GS_CUDA(c, cudaStreamWaitEvent(blend_st, r->ev_ord, 0));
        }
        gs_launch_blend_forward(r->proj, r->sorted_vals, r->ranges, W, H, r->ntx, r->nty, out->px, r->last_idx, r->tfinal, 1, rast_vis_for_forward(r), (size_t)r->vis_cap, blend_st);
        r->mark(7);
        GS_LAUNCH_CHECK(c);
        GS_CUDA(c, cudaEventSynchronize(r->ev_pairs));                   // the GPU is still busy with the kernels above
        const uint64_t Pc = c->h_pinned[GS_META_COARSE_PAIRS], P = c->h_pinned[GS_META_TOTAL_PAIRS];
        if (Pc >= (1ull << 30) || P >= (1ull << 30)) { c->err = "pair count exceeds 2^30"; return GS_ERR_CAPACITY; }
        r->pairs = P;
        r->coarse_hint = Pc;
        if (Pc <= r->cap_pairs && P <= r->cap_pairs) break;
        // rare: grow (synchronises) and redo the pair stages; a clipped coarse list under-counts P, so grow generously then
        GS_TRY(rast_reserve_pairs(r, Pc > r->cap_pairs ? std::max<uint64_t>(Pc * 4, r->cap_pairs * 2) : P));
    }
    (void)u; (void)g; (void)tiles;
    return GS_OK;
}

static int g_sort_mode = 0;
}  // extern "C"
static void gs_set_sort_mode(int v) { g_sort_mode = (v == 1 || v == 2) ? v : 0; }
extern "C" {

int gs_rast_forward(gs_rast* r, gs_gaussians* g, const gs_uniforms* u, gs_image* out) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    GS_ARG(c, g && u && out, "null handle");
    cudaStream_t st = c->stream;
    const uint32_t n = g->n, W = out->w, H = out->h;
    GS_TRY(image_before_write(out, st));
    GS_TRY(rast_reserve_gaussians(r, n));
    GS_TRY(rast_reserve_image(r, W, H));
    r->n = n; r->pairs = 0; r->passes = 0; r->mode = g_sort_mode;
    if (r->mode == 0 && gs_bin_count(r->ntx, r->nty) > GS_MAX_BINS) r->mode = 2;
    cudaStream_t blend_st = st;
    st = r->ord_stream();
    if (st != blend_st && r->free_valid) GS_CUDA(c, cudaStreamWaitEvent(st, r->ev_free, 0));   // this rasterizer's buffers are free from there on
    if (r->mode == 1) {   // meta: key OR = 0, key AND = ~0 (the totals are plain stores of the scans)
        GS_CUDA(c, cudaMemsetAsync(r->meta, 0, 2 * sizeof(uint32_t), st));
        GS_CUDA(c, cudaMemsetAsync(r->meta + GS_META_KEY_AND, 0xFF, sizeof(uint32_t), st));
    }
    r->mark(0);
    // the reference overrides numTilesX/Y/numGaussians from the texture size and live count (:308-311)
    // binned / hoisted pipelines: K1 also fills the digit histograms of the depth sort that follows (its scratch is cleared here)
    r->depth_prehist = (r->mode != 1 && g_depth_prehist && n > 0);
    uint32_t* dhist = r->depth_prehist ? gs_sort_prepare(n, gs_sort_plan(0, 31, ~0ull), r->dsort_scratch, st) : nullptr;
    gs_launch_preprocess(g->view(), r->proj, u->viewMatrix, u->viewProjectionMatrix, u->screenSize[0], u->screenSize[1],
                         u->focalLength[0], u->focalLength[1], r->ntx, r->nty, n, r->mode == 1 ? r->meta : nullptr, dhist, st);
    r->mark(1);
    GS_TRY(r->mode == 1 ? forward_pairs64(r, g, u, out) : r->mode == 2 ? forward_hoisted(r, g, u, out) : forward_binned(r, g, u, out));
    if (r->ord) { GS_CUDA(c, cudaEventRecord(r->ev_free, blend_st)); r->free_valid = true; }
    return image_after_use(out, blend_st);
}

// Opt-in: run this rasterizer's stages BEFORE the blend (projection, depth sort, coarse pairs, bin sort, expansion) on a
// high-priority stream of its own.  They are latency-bound one-wave kernels; the blend kernels are issue-bound and fill the
// GPU.  A caller that alternates TWO rasterizers over independent frames (render-only views, or fwd+bwd frames of one parameter
// set) then gets frame i+1's ordering stage under frame i's blend: the side stream only waits for the end of the SAME
// rasterizer's previous call, the context stream waits for the side stream before the blend.  Contract: work queued on the
// context stream that WRITES the Gaussians (upload, Adam, densify) after this rasterizer's last forward / backward is not
// waited for -- call gs_rast_set_overlap(r, 1) again after such work (it re-arms the wait at the stream's current position),
// or leave overlap off (the default: one stream, the order of the calls).  No counterpart in the reference (one command queue).
int gs_rast_set_overlap(gs_rast* r, int on) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    if (on) {
        if (!r->ord) {
            int lo = 0, hi = 0;
            GS_CUDA(c, cudaDeviceGetStreamPriorityRange(&lo, &hi));
            GS_CUDA(c, cudaStreamCreateWithPriority(&r->ord, cudaStreamNonBlocking, hi));
            GS_CUDA(c, cudaEventCreateWithFlags(&r->ev_ord, cudaEventDisableTiming));
            GS_CUDA(c, cudaEventCreateWithFlags(&r->ev_free, cudaEventDisableTiming));
        }
        GS_CUDA(c, cudaEventRecord(r->ev_free, c->stream));
        r->free_valid = true;
    } else if (r->ord) {
        GS_CUDA(c, cudaStreamSynchronize(r->ord));
        GS_CUDA(c, cudaStreamDestroy(r->ord)); r->ord = nullptr;
        cudaEventDestroy(r->ev_ord); cudaEventDestroy(r->ev_free); r->ev_ord = r->ev_free = nullptr; r->free_valid = false;
    }
    return GS_OK;
}

int gs_rast_backward(gs_rast* r, gs_gaussians* g, gs_grads* gr, const gs_uniforms* u, gs_image* rendered, gs_image* gt, int flags) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    GS_ARG(c, g && gr && u && rendered && gt, "null handle");
    GS_ARG(c, rendered->w == r->W && rendered->h == r->H && gt->w == r->W && gt->h == r->H, "image size differs from the last forward");
    GS_ARG(c, g->n == r->n, "gaussian count differs from the last forward");
    cudaStream_t st = c->stream;
    const uint32_t n = r->n;
    GS_TRY(image_before_read(rendered, st));
    GS_TRY(image_before_read(gt, st));
    r->mark(8);
    const int accumulate = (flags & GS_BWD_ACCUMULATE) ? 1 : 0;
    if (!accumulate) GS_TRY(grads_reserve(gr, n));                    // the CPU memset of :687 is folded into K8's store
    else GS_ARG(c, n <= gr->cap, "gradient buffer too small to accumulate into");
    if (!r->acc_clean) GS_CUDA(c, cudaMemsetAsync(r->acc, 0, (size_t)r->cap_g * 12 * sizeof(float), st));
    r->acc_clean = false;
    const uint32_t* svals = r->sorted_vals;
    const bool masked = r->vis_valid && g_bwd_masks && gs_bwd_uses_visits();
    r->bwd_used_vis = masked;
    gs_launch_blend_backward(r->proj, svals, r->ranges, r->W, r->H, r->ntx, r->nty, rendered->px, gt->px,
                             (flags & GS_BWD_QUANTIZE_RGBA8) ? 1 : 0, r->last_idx, r->tfinal, r->acc, masked ? r->vis : nullptr, (size_t)r->vis_cap, r->vis_warp_extent ? 1 : 0, st);
    if (!r->want_vis && g_bwd_masks) { r->want_vis = true; GS_TRY(rast_reserve_vis(r)); }   // a training caller: record visits from the next forward on
    r->mark(9);
    gs_launch_preprocess_backward(g->view(), r->proj, r->acc, u->viewMatrix, u->focalLength[0], u->focalLength[1], n,
                                  gr->view(), gr->stat_mode, accumulate, st);
    r->acc_clean = true;                                              // K8 cleared every row K7 touched
    r->mark(10);
    GS_LAUNCH_CHECK(c);
    if (r->ord) { GS_CUDA(c, cudaEventRecord(r->ev_free, st)); r->free_valid = true; }
    GS_TRY(image_after_use(rendered, st));
    GS_TRY(image_after_use(gt, st));
    return GS_OK;
}

int gs_rast_set_profiling(gs_rast* r, int on) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    if (on) for (int i = 0; i < GS_PROF_EVENTS; i++) if (!r->ev[i]) GS_CUDA(c, cudaEventCreate(&r->ev[i]));
    r->prof = on != 0;
    return GS_OK;
}
// ms of the last forward (+ backward): [0] preprocess [1] depth sort of the Gaussians (hoisted mode; 0 otherwise) [2] scan
// [3] emit (64-bit mode: includes the host's pair-count read-back) [4] pair sort (all passes) [5] tile ranges
// [6] blend forward [7] backward memsets + blend backward [8] per-Gaussian backward
int gs_rast_get_timings(gs_rast* r, float* ms9, int include_backward) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    GS_ARG(c, r->prof && ms9, "profiling is off");
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    static const int from[9] = {0, 1, 2, 3, 4, 5, 6, 8, 9}, to[9] = {1, 2, 3, 4, 5, 6, 7, 9, 10};
    for (int i = 0; i < 9; i++) {
        ms9[i] = 0.f;
        if (i >= 7 && !include_backward) continue;
        GS_CUDA(c, cudaEventElapsedTime(&ms9[i], r->ev[from[i]], r->ev[to[i]]));
    }
    return GS_OK;
}
// E = list entries the forward walk visits, C = contributing evaluations, for the last forward
int gs_rast_count_evals(gs_rast* r, uint64_t* E, uint64_t* C) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    if (!r->counters) GS_TRY(dalloc(c, &r->counters, 2));
    GS_CUDA(c, cudaMemsetAsync(r->counters, 0, 16, c->stream));
    gs_launch_blend_count(r->proj, r->sorted_vals, r->ranges, r->W, r->H, r->ntx, r->nty, r->counters, c->stream);
    GS_LAUNCH_CHECK(c);
    unsigned long long h[2];
    GS_CUDA(c, cudaMemcpyAsync(h, r->counters, 16, cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (E) *E = h[0];
    if (C) *C = h[1];
    return GS_OK;
}

int gs_rast_get(gs_rast* r, int which, void* host, size_t bytes) {
    gs_ctx* c = r->c; GS_SETDEV(c);
    cudaStream_t st = c->stream;
    const void* src = nullptr; size_t need = 0;
    void* tmp[3] = {nullptr, nullptr, nullptr};
    int rc = GS_OK;
    switch (which) {
        case GS_RAST_PROJECTED:
            need = (size_t)r->n * sizeof(gs_projected);
            GS_TRY(staging_reserve(c, &r->staging, &r->staging_bytes, need));
            gs_launch_projected_to_aos(r->proj, r->n, r->staging, st);
            GS_LAUNCH_CHECK(c);
            src = r->staging; break;
        case GS_RAST_KEYS:
        case GS_RAST_VALUES:
            if (r->mode == 1) { src = which == GS_RAST_KEYS ? (const void*)r->keys_in : (const void*)r->vals_in; }
            else if (r->pairs) {
                // hoisted mode never materialises the reference's pre-sort list: rebuild it (Gaussian order, u64 keys) on demand
                if (cudaMalloc(&tmp[0], ((size_t)r->n + 1) * 4) != cudaSuccess) { cudaGetLastError(); c->err = "cudaMalloc (inspection buffer)"; rc = GS_ERR_OOM; break; }
                if (cudaMalloc(&tmp[1], r->pairs * 8) != cudaSuccess) { cudaGetLastError(); c->err = "cudaMalloc (inspection buffer)"; rc = GS_ERR_OOM; break; }
                if (cudaMalloc(&tmp[2], r->pairs * 4) != cudaSuccess) { cudaGetLastError(); c->err = "cudaMalloc (inspection buffer)"; rc = GS_ERR_OOM; break; }
                gs_launch_exclusive_scan(r->proj.cnt, nullptr, (uint32_t*)tmp[0], r->n, r->meta + GS_META_COUNT - 1, r->scan_scratch, st);
                gs_launch_emit_pairs(r->proj, (const uint32_t*)tmp[0], r->n, r->ntx, (uint64_t*)tmp[1], (uint32_t*)tmp[2], (uint32_t)r->pairs, st);
                src = which == GS_RAST_KEYS ? tmp[1] : tmp[2];
            }
            need = r->pairs * (which == GS_RAST_KEYS ? 8 : 4); break;
        case GS_RAST_SORTED_KEYS:
            if (r->mode == 1) src = r->sorted_keys;
            else if (r->pairs) {
                if (cudaMalloc(&tmp[0], r->pairs * 8) != cudaSuccess) { cudaGetLastError(); c->err = "cudaMalloc (inspection buffer)"; rc = GS_ERR_OOM; break; }
                if (r->mode == 0) gs_launch_rebuild_keys64_ranges(r->ranges, r->ntx * r->nty, r->sorted_vals, r->proj.dkey, (uint32_t)r->pairs, (uint64_t*)tmp[0], st);
                else gs_launch_rebuild_keys64(r->sorted_tiles, r->sorted_vals, r->proj.dkey, (uint32_t)r->pairs, (uint64_t*)tmp[0], st);
                src = tmp[0];
            }
            need = r->pairs * 8; break;
        case GS_RAST_SORTED_VALUES: src = r->sorted_vals; need = r->pairs * 4; break;
        case GS_RAST_TILE_RANGES: src = r->ranges; need = (size_t)r->ntx * r->nty * 8; break;
        case GS_RAST_LAST_IDX: src = r->last_idx; need = (size_t)r->W * r->H * 4; break;
        case GS_RAST_TFINAL: src = r->tfinal; need = (size_t)r->W * r->H * 4; break;
        default: GS_ARG(c, false, "which");
    }
    if (rc == GS_OK && bytes < need) { c->err = "bad argument: host buffer too small"; rc = GS_ERR_ARG; }
    if (rc == GS_OK && need && cudaMemcpyAsync(host, src, need, cudaMemcpyDeviceToHost, st) != cudaSuccess) { c->err = "cudaMemcpyAsync"; rc = GS_ERR_CUDA; }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == GS_OK) { c->err = "cudaStreamSynchronize"; rc = GS_ERR_CUDA; }
    for (void* t : tmp) cudaFree(t);
    if (rc == GS_OK && (which == GS_RAST_VALUES || which == GS_RAST_SORTED_VALUES)) {   // bit 31 is the library's "cannot contribute" mark (pairs.cu)
        uint32_t* v = (uint32_t*)host;
        for (uint64_t i = 0; i < r->pairs; i++) v[i] &= GS_ID_MASK;
    }
    return rc;
}
uint64_t gs_rast_pair_count(const gs_rast* r) { return r->pairs; }
uint64_t gs_rast_coarse_pair_count(const gs_rast* r) { return r->mode == 0 ? r->coarse_hint : 0; }
int gs_rast_sort_passes(const gs_rast* r) { return r->passes; }
int gs_rast_backward_used_visits(const gs_rast* r) { return r->bwd_used_vis ? 1 : 0; }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// GPURadixSort64 / GPURadixSort32 (gpu_sort.hpp:14-107)
// ------------------------------------------------------------------------------------------------------
template <typename KeyT>
struct sort_obj {
    gs_ctx* c;
    uint64_t cap = 0;
    KeyT* kbuf[2] = {nullptr, nullptr};
    uint32_t* vbuf[2] = {nullptr, nullptr};
    KeyT* kin = nullptr;            // sort32_depth: generated keys
    uint32_t* vin = nullptr;
    void* scratch = nullptr;
    int result = 0;
};
struct gs_sort64 : sort_obj<uint64_t> {};
struct gs_sort32 : sort_obj<uint32_t> {};
// getSortedKeys / getSortedValues hand out the sorter's own ping-pong buffers as borrowed pointers: fed straight back into sort()
// (re-sorting a result) pass 0 would scatter into the buffer it reads and the hybrid plan's LSD fallback would start from
// overwritten input.  true = [p, p + bytes) overlaps one of the sorter's internal buffers.
template <typename S>
static bool sort_input_aliases(const S* s, const void* p, size_t bytes, size_t key_bytes) {
    if (!p || !bytes) return false;
    const char* a = static_cast<const char*>(p);
    for (int i = 0; i < 2; i++) {
        const char* k = reinterpret_cast<const char*>(s->kbuf[i]);
        const char* v = reinterpret_cast<const char*>(s->vbuf[i]);
        if (k && a < k + (size_t)s->cap * key_bytes && k < a + bytes) return true;
        if (v && a < v + (size_t)s->cap * 4 && v < a + bytes) return true;
    }
    return false;
}

template <typename S>
static void sort_free(S* s) {
    for (int i = 0; i < 2; i++) { cudaFree(s->kbuf[i]); cudaFree(s->vbuf[i]); s->kbuf[i] = nullptr; s->vbuf[i] = nullptr; }
    cudaFree(s->kin); cudaFree(s->vin); cudaFree(s->scratch);
    s->kin = nullptr; s->vin = nullptr; s->scratch = nullptr;
}
template <typename S>
static int sort_reserve(S* s, uint64_t n, bool gen_inputs) {   // ensureCapacity, gpu_sort.mm
    gs_ctx* c = s->c;
    if (n <= s->cap && (!gen_inputs || s->kin)) return GS_OK;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint64_t cap = std::max<uint64_t>(std::max<uint64_t>(n, s->cap), 1024);
    sort_free(s);
    for (int i = 0; i < 2; i++) { GS_TRY(dalloc(c, &s->kbuf[i], cap)); GS_TRY(dalloc(c, &s->vbuf[i], cap)); }
    if (gen_inputs) { GS_TRY(dalloc(c, &s->kin, cap)); GS_TRY(dalloc(c, &s->vin, cap)); }
    GS_CUDA(c, cudaMalloc(&s->scratch, gs_sort_scratch_bytes(cap, sizeof(*s->kbuf[0]))));
    s->cap = cap;
    return GS_OK;
}

extern "C" {

int gs_sort64_create(gs_ctx* c, uint64_t max_elements, gs_sort64** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_sort64* s = new (std::nothrow) gs_sort64();
    if (!s) return GS_ERR_OOM;
    s->c = c;
    int r = sort_reserve(s, max_elements, false);
    if (r != GS_OK) { sort_free(s); delete s; return r; }
    *out = s;
    return GS_OK;
}
int gs_sort64_destroy(gs_sort64* s) {
    if (!s) return GS_OK;
    cudaSetDevice(s->c->device); cudaStreamSynchronize(s->c->stream);
    sort_free(s); delete s;
    return GS_OK;
}
int gs_sort64_sort(gs_sort64* s, const uint64_t* d_keys, const uint32_t* d_vals, uint64_t n, int begin_bit, int end_bit) {
    gs_ctx* c = s->c; GS_SETDEV(c);
    GS_ARG(c, begin_bit >= 0 && end_bit <= 64 && begin_bit <= end_bit, "bit range");
    GS_ARG(c, (d_keys && d_vals) || n == 0, "null input");
    if (n >= (1ull << 30)) { c->err = "element count exceeds 2^30"; return GS_ERR_CAPACITY; }
    GS_TRY(sort_reserve(s, n, false));
    GS_ARG(c, !sort_input_aliases(s, d_keys, n * 8, 8) && !sort_input_aliases(s, d_vals, n * 4, 8),
           "input aliases the sorter's own result buffers (copy a previous result out before sorting it again)");
    const GsSortPlan plan = gs_sort_plan(begin_bit, end_bit, ~0ull);
    s->result = gs_launch_sort64(d_keys, d_vals, s->kbuf, s->vbuf, n, plan, g_sort_api_hint, s->scratch, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}
int gs_sort64_last_plan(gs_sort64* s, int* hybrid) {
    gs_ctx* c = s->c; GS_SETDEV(c);
    uint32_t f[3] = {0, 0, 0};
    GS_CUDA(c, cudaMemcpyAsync(f, gs_sort_flags(s->scratch), sizeof(f), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (hybrid) *hybrid = (f[0] != 0u && f[1] == 0u) ? 1 : 0;
    return GS_OK;
}
const uint64_t* gs_sort64_sorted_keys(gs_sort64* s) { return s->kbuf[s->result]; }
const uint32_t* gs_sort64_sorted_values(gs_sort64* s) { return s->vbuf[s->result]; }

int gs_sort32_create(gs_ctx* c, uint64_t max_elements, gs_sort32** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_sort32* s = new (std::nothrow) gs_sort32();
    if (!s) return GS_ERR_OOM;
    s->c = c;
    int r = sort_reserve(s, max_elements, true);
    if (r != GS_OK) { sort_free(s); delete s; return r; }
    *out = s;
    return GS_OK;
}
int gs_sort32_destroy(gs_sort32* s) {
    if (!s) return GS_OK;
    cudaSetDevice(s->c->device); cudaStreamSynchronize(s->c->stream);
    sort_free(s); delete s;
    return GS_OK;
}
int gs_sort32_sort(gs_sort32* s, const uint32_t* d_keys, const uint32_t* d_vals, uint64_t n) {
    gs_ctx* c = s->c; GS_SETDEV(c);
    if (n >= (1ull << 30)) { c->err = "element count exceeds 2^30"; return GS_ERR_CAPACITY; }
    GS_TRY(sort_reserve(s, n, true));
    GS_ARG(c, !sort_input_aliases(s, d_keys, n * 4, 4) && !sort_input_aliases(s, d_vals, n * 4, 4),
           "input aliases the sorter's own result buffers (copy a previous result out before sorting it again)");
    const GsSortPlan plan = gs_sort_plan(0, 32, ~0ull);               // 4 passes x 8 bits, gpu_sort.hpp:34
    s->result = gs_launch_sort32(d_keys, d_vals, s->kbuf, s->vbuf, n, nullptr, plan, g_sort_api_hint, s->scratch, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}
int gs_sort32_sort_depth(gs_sort32* s, const void* d_pos4, const float cam[3], uint64_t n) {
    gs_ctx* c = s->c; GS_SETDEV(c);
    GS_ARG(c, n < (1ull << 30), "too many elements");
    GS_TRY(sort_reserve(s, n, true));
    gs_launch_depth_keys32((const float4*)d_pos4, cam[0], cam[1], cam[2], (uint32_t)n, s->kin, s->vin, c->stream);
    return gs_sort32_sort(s, s->kin, s->vin, n);
}
const uint32_t* gs_sort32_sorted_keys(gs_sort32* s) { return s->kbuf[s->result]; }
const uint32_t* gs_sort32_sorted_indices(gs_sort32* s) { return s->vbuf[s->result]; }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// View-parallel gradient sum over NVLink peer memory (peer_reduce.cu)
// ------------------------------------------------------------------------------------------------------
struct gs_dp_blob {             // what a rank publishes: IPC handles of its gradient buffer and flag words + the buffer's geometry
    cudaIpcMemHandle_t buf, flags;
    uint32_t stride, cap, rank, magic;
    unsigned char pad[GS_DP_BLOB_BYTES - 2 * sizeof(cudaIpcMemHandle_t) - 16];
};
static_assert(sizeof(gs_dp_blob) == GS_DP_BLOB_BYTES && GS_DP_MAX_RANKS == GS_MAX_PEERS, "blob layout");
struct gs_dp {
    gs_ctx* c;
    int rank = 0, world = 1;
    gs_grads* bound = nullptr;
    float* bound_ptr = nullptr;
    uint32_t* my_flags = nullptr;       // 3 barriers x GS_MAX_PEERS words, written by the peers
    uint32_t* d_err = nullptr;          // device word: set by a barrier that timed out, read by the reduce kernels (sticky)
    uint32_t* h_err = nullptr;          // the same flag in mapped pinned host memory (polled without a stream synchronisation)
    uint32_t* h_err_dev = nullptr;      // its device-side address
    uint32_t epoch = 0;
    bool connected = false;
    float* mc_buf = nullptr;            // multicast (NVLS) address of the ranks' gradient buffers, or null: plain peer loads / stores
    void* opened[2 * GS_MAX_PEERS] = {};
    GsPeerSet peers{};
};

extern "C" {

int gs_dp_create(gs_ctx* c, int rank, int world, gs_dp** out) {
    GS_SETDEV(c);
    GS_ARG(c, out && world >= 1 && world <= GS_DP_MAX_RANKS && rank >= 0 && rank < world, "rank / world (one node, at most 8 ranks)");
    gs_dp* d = new (std::nothrow) gs_dp();
    if (!d) return GS_ERR_OOM;
    d->c = c; d->rank = rank; d->world = world;
    int rc = dalloc(c, &d->my_flags, 3 * GS_MAX_PEERS);
    if (rc == GS_OK) rc = dalloc(c, &d->d_err, 1);
    if (rc == GS_OK && cudaMemset(d->my_flags, 0, 3 * GS_MAX_PEERS * sizeof(uint32_t)) != cudaSuccess) rc = GS_ERR_CUDA;
    if (rc == GS_OK && cudaMemset(d->d_err, 0, sizeof(uint32_t)) != cudaSuccess) rc = GS_ERR_CUDA;
    if (rc == GS_OK && cudaHostAlloc((void**)&d->h_err, sizeof(uint32_t), cudaHostAllocMapped) != cudaSuccess) rc = GS_ERR_OOM;
    if (rc == GS_OK) { *d->h_err = 0; if (cudaHostGetDevicePointer((void**)&d->h_err_dev, d->h_err, 0) != cudaSuccess) rc = GS_ERR_CUDA; }
    if (rc != GS_OK) { gs_dp_destroy(d); return rc; }
    *out = d;
    return GS_OK;
}
int gs_dp_destroy(gs_dp* d) {
    if (!d) return GS_OK;
    cudaSetDevice(d->c->device); cudaStreamSynchronize(d->c->stream);
    for (void* p : d->opened) if (p) cudaIpcCloseMemHandle(p);
    cudaFree(d->my_flags); cudaFree(d->d_err);
    if (d->h_err) cudaFreeHost(d->h_err);
    delete d;
    return GS_OK;
}
int gs_dp_export(gs_dp* d, gs_grads* g, void* blob_out) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_ARG(c, g && blob_out && g->g, "gradient buffer");
    gs_dp_blob b;
    memset(&b, 0, sizeof(b));
    GS_CUDA(c, cudaIpcGetMemHandle(&b.buf, g->g));
    GS_CUDA(c, cudaIpcGetMemHandle(&b.flags, d->my_flags));
    b.stride = g->stride; b.cap = g->cap; b.rank = (uint32_t)d->rank; b.magic = 0x47534450u;
    memcpy(blob_out, &b, sizeof(b));
    d->bound = g; d->bound_ptr = g->g;
    return GS_OK;
}
int gs_dp_connect(gs_dp* d, const void* blobs) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_ARG(c, blobs && d->bound, "export first, then connect with every rank's blob");
    const gs_dp_blob* B = static_cast<const gs_dp_blob*>(blobs);
    d->peers.rank = d->rank; d->peers.world = d->world;
    for (int q = 0; q < d->world; q++) {
        GS_ARG(c, B[q].magic == 0x47534450u && (int)B[q].rank == q, "blob order / contents");
        GS_ARG(c, B[q].stride == d->bound->stride, "every rank's gradient buffer must have the same capacity");
        if (q == d->rank) { d->peers.bufs[q] = d->bound_ptr; d->peers.flags[q] = d->my_flags; continue; }
        void *pb = nullptr, *pf = nullptr;
        GS_CUDA(c, cudaIpcOpenMemHandle(&pb, B[q].buf, cudaIpcMemLazyEnablePeerAccess));
        d->opened[2 * q] = pb;
        GS_CUDA(c, cudaIpcOpenMemHandle(&pf, B[q].flags, cudaIpcMemLazyEnablePeerAccess));
        d->opened[2 * q + 1] = pf;
        d->peers.bufs[q] = static_cast<float*>(pb); d->peers.flags[q] = static_cast<uint32_t*>(pf);
    }
    d->connected = true;
    return GS_OK;
}
// The same connection from pointers that are already valid in this process (a symmetric-memory rendezvous done by the caller,
// e.g. torch.distributed._symmetric_memory): peer_bufs[q] = rank q's gradient buffer, peer_flags[q] = 3 * 8 zero-initialised
// u32 flag words in rank q's memory, multicast_buf = the multicast address that maps ALL ranks' buffers (null = none).  With a
// multicast address the sum runs in the NVSwitch: rank r issues multimem.ld_reduce over slice r (the switch adds the g copies)
// and multimem.st broadcasts the result; per-GPU traffic drops from 2 (g-1)/g of the payload per direction to 1x.
int gs_dp_connect_ptrs(gs_dp* d, gs_grads* g, void* const* peer_bufs, void* const* peer_flags, void* multicast_buf) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_ARG(c, g && g->g && peer_bufs && peer_flags, "null argument");
    GS_ARG(c, peer_bufs[d->rank] == (void*)g->g, "peer_bufs[rank] must be this rank's gradient buffer");
    d->peers.rank = d->rank; d->peers.world = d->world;
    for (int q = 0; q < d->world; q++) {
        GS_ARG(c, peer_bufs[q] && peer_flags[q], "null peer pointer");
        d->peers.bufs[q] = static_cast<float*>(peer_bufs[q]); d->peers.flags[q] = static_cast<uint32_t*>(peer_flags[q]);
    }
    GS_CUDA(c, cudaMemsetAsync(peer_flags[d->rank], 0, 3 * GS_MAX_PEERS * sizeof(uint32_t), c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));      // the caller runs a host barrier over the ranks before the first all-reduce
    d->bound = g; d->bound_ptr = g->g;
    d->mc_buf = static_cast<float*>(multicast_buf);
    d->connected = true;
    return GS_OK;
}
int gs_dp_allreduce_grads(gs_dp* d, gs_grads* g, uint32_t n) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    if (d->world == 1) return GS_OK;
    GS_ARG(c, d->connected, "gs_dp_connect first");
    GS_ARG(c, g == d->bound && g->g == d->bound_ptr, "the gradient buffer was reallocated after gs_dp_export");
    GS_ARG(c, n <= g->cap, "n");
    if (n == 0) return GS_OK;
    // a barrier of an EARLIER call timed out (a rank fell more than 5 s behind or is gone): that call's sum was withheld and the
    // step poisoned (peer_reduce.cu); report it before another step is built on replicas that no longer agree
    if (*(volatile uint32_t*)d->h_err) { c->err = "peer all-reduce: a rank did not arrive within 5 s (an earlier gradient sum was withheld)"; return GS_ERR_NCCL; }
    d->epoch++;
    gs_launch_peer_allreduce(d->peers, d->mc_buf, n, g->stride, d->epoch, d->d_err, d->h_err_dev, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}
int gs_dp_check(gs_dp* d) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    uint32_t e = 0;
    GS_CUDA(c, cudaMemcpyAsync(&e, d->d_err, sizeof(e), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (e) { c->err = "peer all-reduce: a rank did not arrive within 5 s"; return GS_ERR_NCCL; }
    return GS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// Viewer draw (MTLEngine::render, mtl_engine.mm:401-609, without the window): viewer.cu
// ------------------------------------------------------------------------------------------------------
struct gs_viewer {
    gs_ctx* c;
    uint32_t cap_g = 0, cap_tiles = 0;
    uint64_t cap_pairs = 0;
    float4* rec = nullptr; uint4* bbox = nullptr; uint32_t* cnt = nullptr; uint32_t* offs = nullptr;
    uint32_t* meta = nullptr;                       // [0] = pair count of the frame
    void* scan_scratch = nullptr;
    uint32_t *tiles_in = nullptr, *vals_in = nullptr, *tbuf[2] = {nullptr, nullptr}, *vbuf[2] = {nullptr, nullptr};
    void* sort_scratch = nullptr;
    uint2* ranges = nullptr;
    unsigned long long* counters = nullptr;
    void* staging = nullptr; size_t staging_bytes = 0;
    uint32_t W = 0, H = 0, ntx = 0, nty = 0, n = 0;
    uint64_t pairs = 0;
    const uint32_t* sorted_tiles = nullptr; const uint32_t* sorted_vals = nullptr;
};
static void viewer_free_pairs(gs_viewer* v) {
    cudaFree(v->tiles_in); cudaFree(v->vals_in); cudaFree(v->sort_scratch);
    for (int i = 0; i < 2; i++) { cudaFree(v->tbuf[i]); cudaFree(v->vbuf[i]); v->tbuf[i] = nullptr; v->vbuf[i] = nullptr; }
    v->tiles_in = nullptr; v->vals_in = nullptr; v->sort_scratch = nullptr;
}
static int viewer_reserve_pairs(gs_viewer* v, uint64_t need) {
    if (need <= v->cap_pairs) return GS_OK;
    gs_ctx* c = v->c;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    viewer_free_pairs(v);
    const uint64_t cap = std::max<uint64_t>(need, (uint64_t)(v->cap_pairs * 1.5));
    v->cap_pairs = 0;
    GS_TRY(dalloc(c, &v->tiles_in, cap)); GS_TRY(dalloc(c, &v->vals_in, cap));
    for (int i = 0; i < 2; i++) { GS_TRY(dalloc(c, &v->tbuf[i], cap)); GS_TRY(dalloc(c, &v->vbuf[i], cap)); }
    GS_CUDA(c, cudaMalloc(&v->sort_scratch, gs_sort_scratch_bytes(cap, 4)));
    v->cap_pairs = cap;
    return GS_OK;
}
static void viewer_free_gaussians(gs_viewer* v) {
    cudaFree(v->rec); cudaFree(v->bbox); cudaFree(v->cnt); cudaFree(v->offs); cudaFree(v->scan_scratch);
    v->rec = nullptr; v->bbox = nullptr; v->cnt = nullptr; v->offs = nullptr; v->scan_scratch = nullptr;
}
static int viewer_reserve_gaussians(gs_viewer* v, uint32_t n) {
    if (n <= v->cap_g) return GS_OK;
    gs_ctx* c = v->c;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    viewer_free_gaussians(v);
    v->cap_g = 0;
    const uint32_t cap = ((uint32_t)(n * 1.25) + 63u) & ~63u;
    GS_TRY(dalloc(c, &v->rec, (size_t)cap * 3)); GS_TRY(dalloc(c, &v->bbox, cap)); GS_TRY(dalloc(c, &v->cnt, cap));
    GS_TRY(dalloc(c, &v->offs, (size_t)cap + 1));
    GS_CUDA(c, cudaMalloc(&v->scan_scratch, gs_scan_scratch_bytes(cap)));
    v->cap_g = cap;
    return GS_OK;
}

extern "C" {

int gs_viewer_create(gs_ctx* c, uint32_t max_gaussians, gs_viewer** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_viewer* v = new (std::nothrow) gs_viewer();
    if (!v) return GS_ERR_OOM;
    v->c = c;
    int rc = dalloc(c, &v->meta, 4);
    if (rc == GS_OK) rc = dalloc(c, &v->counters, 2);
    if (rc == GS_OK) rc = viewer_reserve_gaussians(v, std::max(max_gaussians, 64u));
    if (rc == GS_OK) rc = viewer_reserve_pairs(v, (uint64_t)std::max(max_gaussians, 64u) * 8);
    if (rc != GS_OK) { gs_viewer_destroy(v); return rc; }
    *out = v;
    return GS_OK;
}
int gs_viewer_destroy(gs_viewer* v) {
    if (!v) return GS_OK;
    cudaSetDevice(v->c->device);
    cudaStreamSynchronize(v->c->stream);
    viewer_free_gaussians(v); viewer_free_pairs(v);
    cudaFree(v->meta); cudaFree(v->counters); cudaFree(v->ranges); cudaFree(v->staging);
    delete v;
    return GS_OK;
}
int gs_viewer_render(gs_viewer* v, gs_gaussians* g, const uint32_t* d_order, const gs_view_uniforms* u, gs_image* out, int flags) {
    gs_ctx* c = v->c; GS_SETDEV(c);
    cudaStream_t st = c->stream;
    GS_ARG(c, g && u && out, "null argument");
    GS_ARG(c, d_order || g->n == 0, "sorted indices (GPURadixSort32::sort) are required");
    const uint32_t W = out->w, H = out->h, n = g->n;
    GS_ARG(c, (float)W == u->screenSize[0] && (float)H == u->screenSize[1], "uniforms.screenSize must equal the render target size");
    const uint32_t ntx = (W + GS_TILE - 1) / GS_TILE, nty = (H + GS_TILE - 1) / GS_TILE, tiles = ntx * nty;
    GS_TRY(viewer_reserve_gaussians(v, n));
    if (tiles > v->cap_tiles) {
        GS_CUDA(c, cudaStreamSynchronize(st));
        cudaFree(v->ranges); v->ranges = nullptr; v->cap_tiles = 0;
        GS_TRY(dalloc(c, &v->ranges, (size_t)tiles + (tiles + 1) / 2));   // ranges, then the draw CTAs' tile order (u32 each)
        v->cap_tiles = tiles;
    }
    v->W = W; v->H = H; v->ntx = ntx; v->nty = nty; v->n = n;
    GS_TRY(image_before_write(out, st));
    gs_launch_view_splats(g->view(), u->viewMatrix, u->viewProjectionMatrix, u->screenSize[0], u->screenSize[1], u->focalLength[0],
                          u->focalLength[1], n, v->rec, v->bbox, v->cnt, st);
    GS_CUDA(c, cudaMemsetAsync(v->meta, 0, 4 * sizeof(uint32_t), st));
    if (n) gs_launch_exclusive_scan(v->cnt, d_order, v->offs, n, v->meta, v->scan_scratch, st);
    GS_CUDA(c, cudaMemcpyAsync(c->h_pinned, v->meta, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    GS_CUDA(c, cudaStreamSynchronize(st));                              // pair count (sizes the pair buffers), as the tiled forward does
    const uint64_t P = c->h_pinned[0];
    if (P >= (1ull << 30)) { c->err = "pair count exceeds 2^30"; return GS_ERR_CAPACITY; }
    GS_TRY(viewer_reserve_pairs(v, P));
    v->pairs = P;
    int tile_bits = 0;
    while ((1u << tile_bits) < tiles) tile_bits++;
    const GsSortPlan plan = gs_sort_plan_even(0, tile_bits);
    const uint32_t cap = (uint32_t)std::min<uint64_t>(v->cap_pairs, 0x3FFFFFFFull);
    gs_launch_view_emit(v->bbox, v->cnt, d_order, v->offs, n, ntx, v->tiles_in, v->vals_in, cap, st);
    // instances of one tile arrive in draw order; match.any ranking suits the long runs of equal tile ids of large quads
    const int res = gs_launch_sort32(v->tiles_in, v->vals_in, v->tbuf, v->vbuf, cap, v->meta, plan, g_tile_sort_hint, v->sort_scratch, st);
    v->sorted_tiles = plan.num_passes ? v->tbuf[res] : v->tiles_in;
    v->sorted_vals = plan.num_passes ? v->vbuf[res] : v->vals_in;
    gs_launch_tile_ranges32(v->sorted_tiles, cap, v->meta, tiles, v->ranges, st);
    const bool count = (flags & GS_VIEW_COUNT_FRAGMENTS) != 0;
    if (count) GS_CUDA(c, cudaMemsetAsync(v->counters, 0, 2 * sizeof(unsigned long long), st));
    uint32_t* order = g_tile_order ? reinterpret_cast<uint32_t*>(v->ranges + v->cap_tiles) : nullptr;
    if (order) gs_launch_tile_order(v->ranges, tiles, order, st);
    gs_launch_view_blend(v->rec, v->sorted_vals, v->ranges, W, H, ntx, nty, (flags & GS_VIEW_QUANTIZE_RGBA8) ? 1 : 0, out->px,
                         count ? v->counters : nullptr, order, st);
    GS_LAUNCH_CHECK(c);
    GS_TRY(image_after_use(out, st));
    return GS_OK;
}
int gs_viewer_get(gs_viewer* v, int which, void* host, size_t bytes) {
    gs_ctx* c = v->c; GS_SETDEV(c);
    cudaStream_t st = c->stream;
    GS_ARG(c, host || bytes == 0, "host");
    const void* src = nullptr; size_t need = 0;
    switch (which) {
    case GS_VIEW_SPLATS:
        need = (size_t)v->n * sizeof(gs_view_splat);
        GS_TRY(staging_reserve(c, &v->staging, &v->staging_bytes, need));
        gs_launch_view_splats_to_wire(v->rec, v->bbox, v->n, v->staging, st);
        src = v->staging; break;
    case GS_VIEW_SORTED_TILES: need = v->pairs * 4; src = v->sorted_tiles; break;
    case GS_VIEW_SORTED_VALUES: need = v->pairs * 4; src = v->sorted_vals; break;
    case GS_VIEW_TILE_RANGES: need = (size_t)v->ntx * v->nty * sizeof(gs_tile_range); src = v->ranges; break;
    default: GS_ARG(c, false, "which");
    }
    GS_ARG(c, bytes >= need, "host buffer too small");
    if (need) GS_CUDA(c, cudaMemcpyAsync(host, src, need, cudaMemcpyDeviceToHost, st));
    GS_CUDA(c, cudaStreamSynchronize(st));
    return GS_OK;
}
uint64_t gs_viewer_pair_count(const gs_viewer* v) { return v->pairs; }
int gs_viewer_fragments(gs_viewer* v, uint64_t* shaded, uint64_t* blended) {
    gs_ctx* c = v->c; GS_SETDEV(c);
    unsigned long long h[2] = {0, 0};
    GS_CUDA(c, cudaMemcpyAsync(h, v->counters, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (shaded) *shaded = h[0];
    if (blended) *blended = h[1];
    return GS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// AdamOptimizer (optimizer.hpp:23-95, optimizer.mm)
// ------------------------------------------------------------------------------------------------------
struct gs_adam {
    gs_ctx* c;
    float* base = nullptr;      // 38 float planes of stride cap (m: pos3 scale3 op1 sh12, v: same), then m_rot, v_rot float4[cap]
    uint32_t cap = 0;           // allocation stride
    uint32_t num = 0;           // the reference's numGaussians (state covers [0, num))
    uint32_t timestep = 0;
    uint32_t* clean_flags = nullptr;   // device: two words, see gs_adam_step
    uint32_t* skip_words = nullptr;    // one byte per Gaussian: the step's whole-Gaussian skip predicate (adam_skip_kernel)
    uint32_t skip_cap = 0;
    uint32_t sh_clamped_n = 0;  // leading Gaussians whose non-DC SH planes went through the clamp since they were last written
    bool sh_rest_zero = true;   // moments of the nine non-DC SH planes are all +0 (they never receive a gradient; only an upload can change that)
    GsAdamState view() const {
        GsAdamState S;
        const size_t s = cap;
        S.cap = cap;
        S.m_pos = base; S.m_scale = base + 3 * s; S.m_op = base + 6 * s; S.m_sh = base + 7 * s;
        S.v_pos = base + 19 * s; S.v_scale = base + 22 * s; S.v_op = base + 25 * s; S.v_sh = base + 26 * s;
        S.m_rot = reinterpret_cast<float4*>(base + 38 * s); S.v_rot = reinterpret_cast<float4*>(base + 42 * s);
        return S;
    }
};
// plane index / width (floats per Gaussian) of a moment group inside the float-plane block; rotation handled apart
static const int kAdamPlane[5] = {0, 3, -1, 6, 7};
static const int kAdamWidth[5] = {3, 3, 4, 1, 12};

static int adam_zero_range(gs_adam* a, uint32_t lo, uint32_t hi, int group_mask) {   // zero [lo,hi) of the selected groups (m and v)
    gs_ctx* c = a->c;
    if (hi <= lo) return GS_OK;
    for (int g = 0; g < 5; g++) {
        if (!(group_mask & (1 << g))) continue;
        if (g == 2) {
            GS_CUDA(c, cudaMemsetAsync(a->base + 38 * (size_t)a->cap + 4 * (size_t)lo, 0, (size_t)(hi - lo) * 16, c->stream));
            GS_CUDA(c, cudaMemsetAsync(a->base + 42 * (size_t)a->cap + 4 * (size_t)lo, 0, (size_t)(hi - lo) * 16, c->stream));
            continue;
        }
        for (int mv = 0; mv < 2; mv++)
            GS_CUDA(c, cudaMemset2DAsync(a->base + (size_t)(kAdamPlane[g] + 19 * mv) * a->cap + lo, (size_t)a->cap * 4, 0,
                                         (size_t)(hi - lo) * 4, kAdamWidth[g], c->stream));
    }
    return GS_OK;
}

extern "C" {

int gs_adam_create(gs_ctx* c, uint32_t num, gs_adam** out) {
    GS_SETDEV(c);
    GS_ARG(c, out, "out");
    gs_adam* a = new (std::nothrow) gs_adam();
    if (!a) return GS_ERR_OOM;
    a->c = c;
    a->num = num;
    a->cap = (std::max(num, 64u) + 63u) & ~63u;
    int r = dalloc(c, &a->base, (size_t)a->cap * 46);
    if (r != GS_OK) { delete a; return r; }
    *out = a;
    return gs_adam_reset(a);
}
int gs_adam_destroy(gs_adam* a) {
    if (!a) return GS_OK;
    cudaSetDevice(a->c->device); cudaStreamSynchronize(a->c->stream);
    cudaFree(a->base); cudaFree(a->clean_flags); cudaFree(a->skip_words); delete a;
    return GS_OK;
}
int gs_adam_reset(gs_adam* a) {                                        // optimizer.mm:80-93
    gs_ctx* c = a->c; GS_SETDEV(c);
    a->timestep = 0;
    GS_CUDA(c, cudaMemsetAsync(a->base, 0, (size_t)a->cap * 46 * sizeof(float), c->stream));
    a->sh_rest_zero = true;
    a->sh_clamped_n = 0;                                               // the next step runs every job again
    return GS_OK;
}
int gs_adam_resize_if_needed(gs_adam* a, uint32_t n) {                 // optimizer.mm:95-134: grow only, zero-filled tail
    gs_ctx* c = a->c; GS_SETDEV(c);
    if (n <= a->num) return GS_OK;
    if (n > a->cap) {
        const uint32_t ncap = ((uint32_t)(n * 1.25) + 63u) & ~63u;
        float* nb = nullptr;
        GS_TRY(dalloc(c, &nb, (size_t)ncap * 46));
        GS_CUDA(c, cudaMemsetAsync(nb, 0, (size_t)ncap * 46 * sizeof(float), c->stream));
        if (a->num) {
            GS_CUDA(c, cudaMemcpy2DAsync(nb, (size_t)ncap * 4, a->base, (size_t)a->cap * 4, (size_t)a->num * 4, 38, cudaMemcpyDeviceToDevice, c->stream));
            GS_CUDA(c, cudaMemcpyAsync(nb + 38 * (size_t)ncap, a->base + 38 * (size_t)a->cap, (size_t)a->num * 16, cudaMemcpyDeviceToDevice, c->stream));
            GS_CUDA(c, cudaMemcpyAsync(nb + 42 * (size_t)ncap, a->base + 42 * (size_t)a->cap, (size_t)a->num * 16, cudaMemcpyDeviceToDevice, c->stream));
        }
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(a->base);
        a->base = nb; a->cap = ncap;
    } else {
        GS_TRY(adam_zero_range(a, a->num, n, 31));
    }
    a->num = n;
    return GS_OK;
}
int gs_adam_reset_opacity_momentum(gs_adam* a) { GS_SETDEV(a->c); return adam_zero_range(a, 0, a->num, 1 << 3); }   // :137-140
int gs_adam_reset_scale_momentum(gs_adam* a) { GS_SETDEV(a->c); return adam_zero_range(a, 0, a->num, 1 << 1); }     // :143-147
int gs_adam_reset_state_for_new(gs_adam* a, uint32_t start) {          // :150-188
    GS_SETDEV(a->c);
    if (start >= a->num) return GS_OK;
    return adam_zero_range(a, start, a->num, 31);
}
uint32_t gs_adam_timestep(const gs_adam* a) { return a->timestep; }
int gs_adam_set_timestep(gs_adam* a, uint32_t t) { a->timestep = t; return GS_OK; }

int gs_adam_step(gs_adam* a, gs_gaussians* g, gs_grads* gr, const float lrs[5]) {
    gs_ctx* c = a->c; GS_SETDEV(c);
    GS_ARG(c, g && gr && lrs, "null argument");
    a->timestep++;                                                     // optimizer.mm:237
    const uint32_t n = std::min(g->n, a->num);                         // live count, bounded by the state size (note A)
    GS_ARG(c, n <= gr->cap, "gradient buffer smaller than the gaussian count");
    const float beta1 = 0.9f, beta2 = 0.999f, eps = 1e-8f;             // optimizer.mm:272-274
    const float bc1 = 1.0f - powf(beta1, (float)a->timestep);          // shaders.metal:579-580
    const float bc2 = 1.0f - powf(beta2, (float)a->timestep);
    // The non-DC SH planes have no gradient plane at all (gradients.hpp: sh[12] with only the DC terms ever written), so
    // with zero moments their step is "clamp to [-2,2]": once that has run on the current parameter values it is the
    // identity, and those nine jobs return at once (108 B / Gaussian not streamed) until someone writes the planes.  The
    // "clean" word lives on the device (a skipped Gaussian may have to keep an out-of-range value, which only the kernel
    // sees): flags[t & 1] is read by step t, flags[(t + 1) & 1] is preset to 1 and cleared by step t if it found work left.
    if (!a->clean_flags) { GS_TRY(dalloc(c, &a->clean_flags, 2)); GS_CUDA(c, cudaMemsetAsync(a->clean_flags, 0, 8, c->stream)); }
    uint32_t* cin = a->clean_flags + (a->timestep & 1u);
    uint32_t* cout = a->clean_flags + ((a->timestep + 1u) & 1u);
    const bool known = a->sh_rest_zero && !g->sh_rest_dirty && n <= a->sh_clamped_n;
    if (!known) GS_CUDA(c, cudaMemsetAsync(cin, 0, 4, c->stream));      // host-visible writers: force the nine jobs to run
    GS_CUDA(c, cudaMemsetAsync(cout, 1, 4, c->stream));                   // any non-zero word means clean
    if (n > a->skip_cap) {
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(a->skip_words); a->skip_words = nullptr;
        a->skip_cap = std::max(a->cap, (n + 1023u) & ~1023u);
        GS_TRY(dalloc(c, &a->skip_words, (size_t)a->skip_cap / 4 + 1));
    }
    gs_launch_adam(g->view(), gr->view(), a->view(), n, lrs, beta1, beta2, eps, bc1, bc2, cin, cout, a->skip_words, c->stream);
    GS_LAUNCH_CHECK(c);
    a->sh_clamped_n = std::max(a->sh_clamped_n, n);
    g->sh_rest_dirty = false;
    return GS_OK;
}

static int adam_xfer(gs_adam* a, int group, int mv, float* host, const float* src_host, uint32_t n) {
    gs_ctx* c = a->c; GS_SETDEV(c);
    GS_ARG(c, group >= 0 && group < 5 && (mv == 0 || mv == 1) && n <= a->cap, "group / moment / n");
    const int w = kAdamWidth[group];
    if (group == 2) {
        float* d = a->base + (size_t)(38 + 4 * mv) * a->cap;
        if (host) GS_CUDA(c, cudaMemcpyAsync(host, d, (size_t)n * 16, cudaMemcpyDeviceToHost, c->stream));
        else GS_CUDA(c, cudaMemcpyAsync(d, src_host, (size_t)n * 16, cudaMemcpyHostToDevice, c->stream));
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
        return GS_OK;
    }
    float* d = a->base + (size_t)(kAdamPlane[group] + 19 * mv) * a->cap;
    std::vector<float> tmp((size_t)n * w);
    if (host) {     // planes -> the reference's packed [n][w]
        GS_CUDA(c, cudaMemcpy2DAsync(tmp.data(), (size_t)n * 4, d, (size_t)a->cap * 4, (size_t)n * 4, w, cudaMemcpyDeviceToHost, c->stream));
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
        for (uint32_t i = 0; i < n; i++) for (int k = 0; k < w; k++) host[(size_t)i * w + k] = tmp[(size_t)k * n + i];
    } else {
        if (group == 4) a->sh_rest_zero = false;   // caller-supplied SH moments: no longer known to be zero
        for (uint32_t i = 0; i < n; i++) for (int k = 0; k < w; k++) tmp[(size_t)k * n + i] = src_host[(size_t)i * w + k];
        GS_CUDA(c, cudaMemcpy2DAsync(d, (size_t)a->cap * 4, tmp.data(), (size_t)n * 4, (size_t)n * 4, w, cudaMemcpyHostToDevice, c->stream));
        GS_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return GS_OK;
}
int gs_adam_download(gs_adam* a, int group, int mv, float* host, uint32_t n) { return adam_xfer(a, group, mv, host, nullptr, n); }
int gs_adam_upload(gs_adam* a, int group, int mv, const float* host, uint32_t n) { return adam_xfer(a, group, mv, nullptr, host, n); }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// DensityController (density_control.hpp:20-62, density_control.mm)
// ------------------------------------------------------------------------------------------------------
struct gs_dens {
    gs_ctx* c;
    uint32_t max_gaussians = 1500000;   // MAX_GAUSSIANS, density_control.mm:27
    float scene_extent = 1.0f;          // :41
    uint32_t seed = 0x5EEDu;
    uint32_t cap = 0;                   // per-Gaussian arrays
    float* accum = nullptr;
    uint32_t *count = nullptr, *markers = nullptr, *rank = nullptr, *out_off = nullptr, *small = nullptr;   // small: counts3 + stats4
    void* scan_scratch = nullptr;
    float* spare = nullptr; uint32_t spare_cap = 0;   // ping-pong target for the rebuilt model state
    uint32_t applied_n = 0;             // count the last apply marked (markers / out_off describe [0, applied_n)); 0 = no plan
};
static int dens_reserve(gs_dens* d, uint32_t n) {
    if (n <= d->cap) return GS_OK;
    gs_ctx* c = d->c;
    const uint32_t cap = ((uint32_t)(n * 1.25) + 63u) & ~63u;
    float* na = nullptr; uint32_t* nc = nullptr;
    GS_TRY(dalloc(c, &na, cap)); GS_TRY(dalloc(c, &nc, cap));
    GS_CUDA(c, cudaMemsetAsync(na, 0, (size_t)cap * 4, c->stream));
    GS_CUDA(c, cudaMemsetAsync(nc, 0, (size_t)cap * 4, c->stream));
    if (d->cap) {
        GS_CUDA(c, cudaMemcpyAsync(na, d->accum, (size_t)d->cap * 4, cudaMemcpyDeviceToDevice, c->stream));
        GS_CUDA(c, cudaMemcpyAsync(nc, d->count, (size_t)d->cap * 4, cudaMemcpyDeviceToDevice, c->stream));
    }
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(d->accum); cudaFree(d->count); cudaFree(d->markers); cudaFree(d->rank); cudaFree(d->out_off); cudaFree(d->scan_scratch);
    d->accum = na; d->count = nc;
    GS_TRY(dalloc(c, &d->markers, cap)); GS_TRY(dalloc(c, &d->rank, (size_t)cap + 1)); GS_TRY(dalloc(c, &d->out_off, (size_t)cap + 1));
    GS_CUDA(c, cudaMalloc(&d->scan_scratch, gs_scan_scratch_bytes(cap)));
    d->cap = cap;
    return GS_OK;
}

extern "C" {

int gs_dens_create(gs_ctx* c, uint32_t max_gaussians, gs_dens** out) {
    GS_SETDEV(c);
    GS_ARG(c, out && max_gaussians > 0, "out / max_gaussians");
    gs_dens* d = new (std::nothrow) gs_dens();
    if (!d) return GS_ERR_OOM;
    d->c = c; d->max_gaussians = max_gaussians;
    int r = dalloc(c, &d->small, 8);
    if (r == GS_OK) r = dens_reserve(d, std::min<uint32_t>(max_gaussians, 4u << 20));
    if (r != GS_OK) { gs_dens_destroy(d); return r; }
    *out = d;
    return GS_OK;
}
int gs_dens_destroy(gs_dens* d) {
    if (!d) return GS_OK;
    cudaSetDevice(d->c->device); cudaStreamSynchronize(d->c->stream);
    cudaFree(d->accum); cudaFree(d->count); cudaFree(d->markers); cudaFree(d->rank); cudaFree(d->out_off);
    cudaFree(d->small); cudaFree(d->scan_scratch); cudaFree(d->spare);
    delete d;
    return GS_OK;
}
int gs_dens_set_scene_extent(gs_dens* d, float e) { d->scene_extent = e; return GS_OK; }   // density_control.mm:79-84
int gs_dens_set_seed(gs_dens* d, uint32_t s) { d->seed = s; return GS_OK; }
int gs_dens_reset_accumulator(gs_dens* d, uint32_t n) {                // :113-118
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_TRY(dens_reserve(d, n));
    GS_CUDA(c, cudaMemsetAsync(d->accum, 0, (size_t)n * 4, c->stream));
    GS_CUDA(c, cudaMemsetAsync(d->count, 0, (size_t)n * 4, c->stream));
    return GS_OK;
}
int gs_dens_accumulate(gs_dens* d, gs_grads* gr, uint32_t n) {         // :121-185
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_ARG(c, gr && n <= gr->cap, "gradient buffer");
    GS_TRY(dens_reserve(d, n));
    gs_launch_accumulate_stats(gr->view(), gr->stat_mode, n, d->accum, d->count, c->stream);
    GS_LAUNCH_CHECK(c);
    return GS_OK;
}
int gs_dens_apply(gs_dens* d, gs_gaussians* g, uint32_t iteration, float focal, float image_width, float avg_depth,
                  gs_density_stats* stats) {                           // :188-501
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_ARG(c, g, "gaussians");
    gs_density_stats s = {0, 0, 0};
    if (stats) *stats = s;
    const uint32_t n = g->n;
    GS_TRY(dens_reserve(d, n));
    d->applied_n = 0;
    if (iteration >= 15000u) return gs_dens_reset_accumulator(d, n);   // :216-220
    cudaStream_t st = c->stream;
    GsDensifyArgs a;
    a.n = n; a.iteration = iteration; a.seed = d->seed; a.max_gaussians = d->max_gaussians;
    a.scene_extent = d->scene_extent; a.focal = focal; a.image_width = image_width; a.avg_depth = avg_depth;
    const uint64_t bound = std::min<uint64_t>(2ull * n, std::max<uint64_t>(n, d->max_gaussians));
    if (bound > d->spare_cap) {
        GS_CUDA(c, cudaStreamSynchronize(st));
        cudaFree(d->spare); d->spare = nullptr;
        // room for the cap from the start (the model can never exceed max(n, max_gaussians)): no reallocation at later events
        d->spare_cap = ((uint32_t)std::min<uint64_t>(std::max<uint64_t>(bound, d->max_gaussians), 0xFFFFFF00ull) + 63u) & ~63u;
        GS_TRY(dalloc(c, &d->spare, (size_t)d->spare_cap * 23));
    }
    gs_launch_densify_mark(g->view(), d->accum, d->count, a, d->markers, d->small, st);
    gs_launch_densify_plan(d->markers, n, d->max_gaussians, d->small, d->rank, d->out_off, d->small + 4, d->scan_scratch, st);
    gs_gaussians tmp = *g;
    tmp.base = d->spare; tmp.cap = d->spare_cap;
    gs_launch_densify_scatter(g->view(), tmp.view(), d->markers, d->out_off, a, st);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaMemcpyAsync(c->h_pinned, d->small + 4, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    GS_CUDA(c, cudaStreamSynchronize(st));
    s.numPruned = c->h_pinned[0]; s.numCloned = c->h_pinned[1]; s.numSplit = c->h_pinned[2];
    const uint32_t newCount = c->h_pinned[3];
    std::swap(g->base, d->spare); std::swap(g->cap, d->spare_cap);      // :486-490
    g->sh_rest_dirty = true;
    g->n = newCount;
    d->applied_n = n;
    if (stats) *stats = s;
    return gs_dens_reset_accumulator(d, newCount);                      // :493
}
int gs_adam_permute_after_densify(gs_adam* a, gs_dens* d, gs_gaussians* g, uint32_t old_n) {
    gs_ctx* c = a->c; GS_SETDEV(c);
    GS_ARG(c, d && g, "null argument");
    const uint32_t new_n = g->n;
    if (d->applied_n == 0) return gs_adam_resize_if_needed(a, new_n);      // the apply changed nothing (past DENSIFY_UNTIL_ITER)
    GS_ARG(c, d->applied_n == old_n, "old_n does not match the last gs_dens_apply");
    GS_ARG(c, old_n <= a->num, "optimizer state smaller than the model");
    const uint32_t ncap = std::max<uint32_t>(a->cap, ((uint32_t)(new_n * 1.25) + 63u) & ~63u);
    float* nb = nullptr;
    GS_TRY(dalloc(c, &nb, (size_t)ncap * 46));
    GS_CUDA(c, cudaMemsetAsync(nb, 0, (size_t)ncap * 46 * sizeof(float), c->stream));
    gs_launch_permute_moments(a->base, a->cap, nb, ncap, d->markers, d->out_off, old_n, new_n, c->stream);
    GS_LAUNCH_CHECK(c);
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(a->base);
    a->base = nb; a->cap = ncap; a->num = std::max(a->num, new_n);
    return GS_OK;
}
int gs_dens_download(gs_dens* d, float* accum, uint32_t* count, uint32_t* markers, uint32_t n) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_ARG(c, n <= d->cap, "n");
    if (accum) GS_CUDA(c, cudaMemcpyAsync(accum, d->accum, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (count) GS_CUDA(c, cudaMemcpyAsync(count, d->count, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (markers) GS_CUDA(c, cudaMemcpyAsync(markers, d->markers, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}
int gs_dens_upload(gs_dens* d, const float* accum, const uint32_t* count, uint32_t n) {
    gs_ctx* c = d->c; GS_SETDEV(c);
    GS_TRY(dens_reserve(d, n));
    if (accum) GS_CUDA(c, cudaMemcpyAsync(d->accum, accum, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    if (count) GS_CUDA(c, cudaMemcpyAsync(d->count, count, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return GS_OK;
}

}  // extern "C"
