// gs_host.hpp — the reference's host-side operator classes re-bodied over the C ABI (libgs_b200.so).
//
// Same class names, public methods and argument lists as ctaylo41/GaussianSplatting's
//   tiled_rasterizer.hpp:56-75   TiledRasterizer            gpu_sort.hpp:14-107   GPURadixSort32 / GPURadixSort64
//   optimizer.hpp:23-61          AdamOptimizer              density_control.hpp:20-48  DensityController
//   gradients.hpp:11-31          GaussianGradients          ply_loader.hpp:14-20  Gaussian
// so that MTLEngine::trainStep / train (mtl_engine.mm:856-1221) call them unchanged.  Each body is a forwarding shell:
// the pipeline states, scratch buffers and dispatch code of the .mm files live inside the library.  Errors: as in the
// reference, the methods return void and a failing device call is logged to std::cerr (with gs_last_error()'s text) and
// remembered in MTL::lastError(); MTL::setThrowOnError(true) turns that into std::runtime_error (metal_shim.hpp).
#pragma once
#include <vector>
#include "metal_shim.hpp"

// ---- wire structs (byte-identical to the reference; include/gs_abi.h carries the static_asserts) -------------------
struct Gaussian {                 // ply_loader.hpp:14-20
    simd_float3 position;
    simd_float3 scale;            // LOG scale
    simd_float4 rotation;         // (w,x,y,z)
    float opacity;                // raw (pre-sigmoid)
    float sh[12];
};
struct GaussianGradients {        // gradients.hpp:11-31
    float position_x, position_y, position_z, opacity;
    float scale_x, scale_y, scale_z, _pad1;
    simd_float4 rotation;
    float sh[12];
    float viewspace_grad_x, viewspace_grad_y, _pad2, _pad3;
};
struct TileRange { uint32_t start, count; };                          // tiled_rasterizer.hpp:16-19
struct ProjectedGaussian {        // tiled_rasterizer.hpp:24-39
    simd_float2 screenPos; float conic[3]; float depth; float opacity; float color[3]; float radius;
    uint32_t tileMinX, tileMinY, tileMaxX, tileMaxY; float _pad1; simd_float2 viewPos_xy; float cov2D[3]; float _pad2;
};
struct TiledUniforms {            // tiled_rasterizer.hpp:42-53
    simd_float4x4 viewMatrix, projectionMatrix, viewProjectionMatrix;
    simd_float2 screenSize, focalLength;
    simd_float3 cameraPos;
    uint32_t numTilesX, numTilesY, numGaussians, _pad2;
};
static_assert(sizeof(Gaussian) == sizeof(gs_gaussian) && sizeof(GaussianGradients) == sizeof(gs_gradients), "wire layout");
static_assert(sizeof(ProjectedGaussian) == sizeof(gs_projected) && sizeof(TileRange) == sizeof(gs_tile_range), "wire layout");
static_assert(sizeof(TiledUniforms) == sizeof(gs_uniforms), "wire layout");
struct Uniforms {                 // mtl_engine.hpp:30-38 (the viewer's uniform block)
    simd_float4x4 viewMatrix, projectionMatrix, viewProjectionMatrix;
    simd_float2 screenSize, focalLength;
    simd_float3 cameraPos;
    float _pad;
};
static_assert(sizeof(Uniforms) == sizeof(gs_view_uniforms), "wire layout");

// ---- GPURadixSort64 (gpu_sort.hpp:64-107) --------------------------------------------------------------------------
class GPURadixSort64 {
public:
    GPURadixSort64(MTL::Device* device, MTL::Library*, size_t maxElements) : device_(device) {
        MTL::gs_check(device->ctx(), gs_sort64_create(device->ctx(), maxElements, &s_), "GPURadixSort64");
    }
    ~GPURadixSort64() { if (keys_) keys_->release(); if (vals_) vals_->release(); gs_sort64_destroy(s_); }
    // sort 64-bit keys with 32-bit values (stable, all 64 bits); keysIn / valuesIn are raw device buffers
    void sort(MTL::CommandQueue*, MTL::Buffer* keysIn, MTL::Buffer* valuesIn, size_t numElements) {
        MTL::gs_check(device_->ctx(), gs_sort64_sort(s_, (const uint64_t*)keysIn->devptr(), (const uint32_t*)valuesIn->devptr(),
                                                      numElements, 0, 64), "GPURadixSort64::sort");
        n_ = numElements;
    }
    // gpu_sort.hpp:76-77: the sorter's internal result buffers, borrowed, valid until the next sort
    MTL::Buffer* getSortedKeys() { return rewrap(keys_, gs_sort64_sorted_keys(s_), n_ * 8); }
    MTL::Buffer* getSortedValues() { return rewrap(vals_, gs_sort64_sorted_values(s_), n_ * 4); }
private:
    MTL::Buffer* rewrap(MTL::Buffer*& slot, const void* p, size_t bytes) {
        if (slot) slot->release();
        return slot = device_->wrapBuffer(p, bytes);
    }
    MTL::Device* device_;
    gs_sort64* s_ = nullptr;
    size_t n_ = 0;
    MTL::Buffer *keys_ = nullptr, *vals_ = nullptr;
};

// ---- GPURadixSort32 (gpu_sort.hpp:14-62): viewer depth order, far to near -----------------------------------------
class GPURadixSort32 {
public:
    GPURadixSort32(MTL::Device* device, MTL::Library*, size_t maxElements) : device_(device) {
        MTL::gs_check(device->ctx(), gs_sort32_create(device->ctx(), maxElements, &s_), "GPURadixSort32");
    }
    ~GPURadixSort32() { if (idx_) idx_->release(); gs_sort32_destroy(s_); }
    // positionBuffer: float4 per Gaussian (updatePositionBuffer, mtl_engine.mm:1028-1036).  Returns the sorted indices as the
    // reference does (gpu_sort.hpp:19-22): the sorter's internal buffer, borrowed, valid until the next sort
    MTL::Buffer* sort(MTL::CommandQueue*, MTL::Buffer* positionBuffer, simd_float3 cameraPos, size_t numElements) {
        const float cam[3] = {cameraPos.x, cameraPos.y, cameraPos.z};
        MTL::gs_check(device_->ctx(), gs_sort32_sort_depth(s_, positionBuffer->devptr(), cam, numElements), "GPURadixSort32::sort");
        n_ = numElements;
        return getSortedIndices();
    }
    MTL::Buffer* getSortedIndices() {                                        // gpu_sort.hpp:26
        if (idx_) idx_->release();
        return idx_ = device_->wrapBuffer(gs_sort32_sorted_indices(s_), n_ * 4);
    }
private:
    MTL::Device* device_;
    gs_sort32* s_ = nullptr;
    size_t n_ = 0;
    MTL::Buffer* idx_ = nullptr;
};

// ---- the viewer's draw call (MTLEngine::createPipeline + the render encoder of MTLEngine::render, mtl_engine.mm:337-375,575-605):
// vertexShader / fragmentShader + blend state as CUDA compute into a headless texture; no reference class of its own
class SplatRenderPipeline {
public:
    SplatRenderPipeline(MTL::Device* device, MTL::Library*, uint32_t maxGaussians) : device_(device) {
        MTL::gs_check(device->ctx(), gs_viewer_create(device->ctx(), maxGaussians, &v_), "SplatRenderPipeline");
    }
    ~SplatRenderPipeline() { gs_viewer_destroy(v_); }
    // drawPrimitives(TriangleStrip, 0, 4, gaussianCount) with vertex buffers (gaussians, uniforms, sortedIndices), :592-597
    void draw(MTL::CommandQueue*, MTL::Buffer* gaussianBuffer, const Uniforms& uniforms, MTL::Buffer* sortedIndices, MTL::Texture* target) {
        MTL::gs_check(device_->ctx(), gs_viewer_render(v_, gaussianBuffer->gaussians(), (const uint32_t*)sortedIndices->devptr(),
                                                       reinterpret_cast<const gs_view_uniforms*>(&uniforms), target->image(),
                                                       quantize_ ? GS_VIEW_QUANTIZE_RGBA8 : 0), "SplatRenderPipeline::draw");
    }
    void setRGBA8Target(bool on) { quantize_ = on; }                           // colour attachment RGBA8Unorm, :351
    uint64_t pairCount() const { return gs_viewer_pair_count(v_); }
private:
    MTL::Device* device_;
    gs_viewer* v_ = nullptr;
    bool quantize_ = false;
};

// ---- TiledRasterizer (tiled_rasterizer.hpp:56-75) ------------------------------------------------------------------
class TiledRasterizer {
public:
    TiledRasterizer(MTL::Device* device, MTL::Library*, uint32_t maxGaussians) : device_(device) {
        MTL::gs_check(device->ctx(), gs_rast_create(device->ctx(), maxGaussians, &r_), "TiledRasterizer");
    }
    ~TiledRasterizer() { gs_rast_destroy(r_); }
    void forward(MTL::CommandQueue*, MTL::Buffer* gaussianBuffer, size_t /*gaussianCount: the buffer knows*/,
                 const TiledUniforms& uniforms, MTL::Texture* outputTexture) {
        MTL::gs_check(device_->ctx(), gs_rast_forward(r_, gaussianBuffer->gaussians(), reinterpret_cast<const gs_uniforms*>(&uniforms),
                                                      outputTexture->image()), "TiledRasterizer::forward");
    }
    void backward(MTL::CommandQueue*, MTL::Buffer* gaussianBuffer, MTL::Buffer* gradientBuffer, size_t,
                  const TiledUniforms& uniforms, MTL::Texture* renderedTexture, MTL::Texture* groundTruthTexture) {
        MTL::gs_check(device_->ctx(), gs_rast_backward(r_, gaussianBuffer->gaussians(), gradientBuffer->grads(),
                                                       reinterpret_cast<const gs_uniforms*>(&uniforms), renderedTexture->image(),
                                                       groundTruthTexture->image(), quantize_ ? GS_BWD_QUANTIZE_RGBA8 : 0),
                      "TiledRasterizer::backward");
    }
    // the reference's render target is RGBA8Unorm, so its backward reads an 8-bit-quantised image (SURVEY note G)
    void setQuantizeRenderedRGBA8(bool on) { quantize_ = on; }
    uint64_t pairCount() const { return gs_rast_pair_count(r_); }
    gs_rast* handle() const { return r_; }
private:
    MTL::Device* device_;
    gs_rast* r_ = nullptr;
    bool quantize_ = true;
};

// ---- AdamOptimizer (optimizer.hpp:23-61) ---------------------------------------------------------------------------
class AdamOptimizer {
public:
    AdamOptimizer(MTL::Device* device, MTL::Library*, size_t numGaussians) : device_(device) {
        MTL::gs_check(device->ctx(), gs_adam_create(device->ctx(), (uint32_t)numGaussians, &a_), "AdamOptimizer");
    }
    ~AdamOptimizer() { gs_adam_destroy(a_); }
    void step(MTL::CommandQueue*, MTL::Buffer* gaussians, MTL::Buffer* gradients, float lr_position = 0.00016f,
              float lr_scale = 0.005f, float lr_rotation = 0.001f, float lr_opacity = 0.05f, float lr_sh = 0.0025f) {
        const float lrs[5] = {lr_position, lr_scale, lr_rotation, lr_opacity, lr_sh};
        MTL::gs_check(device_->ctx(), gs_adam_step(a_, gaussians->gaussians(), gradients->grads(), lrs), "AdamOptimizer::step");
    }
    void reset() { MTL::gs_check(device_->ctx(), gs_adam_reset(a_), "reset"); }
    void resizeIfNeeded(size_t n) { MTL::gs_check(device_->ctx(), gs_adam_resize_if_needed(a_, (uint32_t)n), "resizeIfNeeded"); }
    void resetOpacityMomentum() { MTL::gs_check(device_->ctx(), gs_adam_reset_opacity_momentum(a_), "resetOpacityMomentum"); }
    void resetScaleMomentum() { MTL::gs_check(device_->ctx(), gs_adam_reset_scale_momentum(a_), "resetScaleMomentum"); }
    void resetStateForNewGaussians(size_t startIdx) {
        MTL::gs_check(device_->ctx(), gs_adam_reset_state_for_new(a_, (uint32_t)startIdx), "resetStateForNewGaussians");
    }
    uint32_t getTimestep() const { return gs_adam_timestep(a_); }
    gs_adam* handle() const { return a_; }
    // opt-in correction of SURVEY note D: moments follow the Gaussians through DensityController::apply (declared after it)
    void permuteStateForDensify(class DensityController& dc, MTL::Buffer* gaussians, size_t oldCount);
private:
    MTL::Device* device_;
    gs_adam* a_ = nullptr;
};

// ---- DensityController (density_control.hpp:20-48) -----------------------------------------------------------------
struct DensityStats { uint32_t numPruned, numCloned, numSplit; };

class DensityController {
public:
    DensityController(MTL::Device* device, MTL::Library*) : device_(device) {
        MTL::gs_check(device->ctx(), gs_dens_create(device->ctx(), 1500000u /* MAX_GAUSSIANS, density_control.mm:27 */, &d_), "DensityController");
        gs_dens_set_scene_extent(d_, sceneExtent());
    }
    ~DensityController() { gs_dens_destroy(d_); }
    // gaussianBuffer's storage may be replaced inside the same handle and gaussianCount is updated (the reference
    // swaps the MTL::Buffer* through the reference parameter, density_control.mm:486-490).  positionBuffer,
    // gradientAccum, gradThreshold, minOpacity and maxScale are accepted and ignored exactly as the reference ignores
    // them (SURVEY notes E/F: file-scope constants win).
    DensityStats apply(MTL::CommandQueue*, MTL::Buffer*& gaussianBuffer, MTL::Buffer*& /*positionBuffer*/, MTL::Buffer* /*gradientAccum*/,
                       size_t& gaussianCount, size_t iteration, float /*gradThreshold*/ = 0.0002f, float /*minOpacity*/ = 0.005f,
                       float /*maxScale*/ = 0.5f, float focalLength = 500.0f, float imageWidth = 800.0f, float avgDepth = 5.0f) {
        gs_density_stats st = {0, 0, 0};
        gs_dens_set_scene_extent(d_, sceneExtent());
        MTL::gs_check(device_->ctx(), gs_dens_apply(d_, gaussianBuffer->gaussians(), (uint32_t)iteration, focalLength, imageWidth,
                                                    avgDepth, &st), "DensityController::apply");
        gaussianCount = gs_gaussians_count(gaussianBuffer->gaussians());
        return DensityStats{st.numPruned, st.numCloned, st.numSplit};
    }
    void accumulateGradients(MTL::CommandQueue*, MTL::Buffer* gradients, size_t gaussianCount) {
        MTL::gs_check(device_->ctx(), gs_dens_accumulate(d_, gradients->grads(), (uint32_t)gaussianCount), "accumulateGradients");
    }
    void resetAccumulator(size_t gaussianCount) {
        MTL::gs_check(device_->ctx(), gs_dens_reset_accumulator(d_, (uint32_t)gaussianCount), "resetAccumulator");
    }
    static void setSceneExtent(float extent) { sceneExtent() = extent; }          // density_control.mm:79-84 (file-scope static)
    gs_dens* handle() const { return d_; }
    void setSeed(uint32_t s) { gs_dens_set_seed(d_, s); }                          // counter-based split RNG (SURVEY note L)
private:
    static float& sceneExtent() { static float e = 1.0f; return e; }
    MTL::Device* device_;
    gs_dens* d_ = nullptr;
};
inline void AdamOptimizer::permuteStateForDensify(DensityController& dc, MTL::Buffer* gaussians, size_t oldCount) {
    MTL::gs_check(device_->ctx(), gs_adam_permute_after_densify(a_, dc.handle(), gaussians->gaussians(), (uint32_t)oldCount), "permuteStateForDensify");
}
