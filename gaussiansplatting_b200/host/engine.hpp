// engine.hpp — headless training driver: the parts of MTLEngine (mtl_engine.hpp / mtl_engine.mm) that call the hot path.
//
//   loadGaussians   mtl_engine.mm:278-330   upload + construct the four operator objects
//   trainStep       mtl_engine.mm:856-1025  uniforms from the COLMAP pose -> forward -> backward -> accumulateGradients -> Adam
//   train           mtl_engine.mm:1047-1221 LR schedule, densify every 100 iterations in (500, 15000), opacity reset every 3000
//   viewMatrixFromColmap / projectionFromColmap  mtl_engine.mm:637-682;  exponentialLRDecay  :1039-1045
//   render          mtl_engine.mm:401-609   viewer frame: GPURadixSort32 draw order + the instanced-quad draw, into a headless texture
// Window, GLFW, swapchain and input callbacks are not here (DESIGN.md "Out of scope").
#pragma once
#include <chrono>
#include <cmath>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>
#include "gs_host.hpp"
#include "ply_io.hpp"

struct TrainingView {              // TrainingImage + ColmapImage/ColmapCamera of the reference, reduced to what trainStep reads
    float qw, qx, qy, qz;          // world -> camera rotation (COLMAP)
    float tx, ty, tz;              // translation
    float fx, fy, cx, cy;          // intrinsics in pixels of `texture`
    MTL::Texture* texture;         // ground truth
};

struct TrainSchedule {             // the constants of MTLEngine::train (mtl_engine.mm:1053-1068, mtl_engine.hpp:167)
    size_t startIteration = 0;     // resuming from a checkpoint: iterations already done (LR decay, densify and reset schedule continue)
    size_t totalIterations = 0;    // 0 = numEpochs * views (the reference); set when a run is split over several train() calls
    size_t opacityResetInterval = 3000, densifyFromIter = 500, densifyUntilIter = 15000, densityControlInterval = 100;
    float opacityResetValue = -4.6f;
    float positionLRInit = 0.00016f, positionLRFinal = 0.0000016f, scaleLR = 0.005f, rotationLR = 0.001f,
          opacityLR = 0.025f, shLR = 0.0025f;
};

struct TrainStats {
    size_t iterations = 0, densifyEvents = 0, opacityResets = 0;
    uint64_t pruned = 0, cloned = 0, split = 0;
    double seconds = 0, densifySeconds = 0;
    size_t finalCount = 0;
};

class Engine {
public:
    static constexpr size_t kOperatorCapacity = 2000000;
    explicit Engine(int deviceOrdinal = 0) {                                    // initHeadless, mtl_engine.mm:79-86
        device = MTL::Device::CreateSystemDefaultDevice(deviceOrdinal);
        if (!device) throw std::runtime_error("no CUDA device");
        commandQueue = device->newCommandQueue();
        library = device->newDefaultLibrary();
    }
    ~Engine() {
        delete tiledRasterizer; delete optimizer; delete densityController; delete gpuSort; delete renderPipeline;
        if (positionBuffer) positionBuffer->release();
        if (gaussianBuffer) gaussianBuffer->release();
        if (gaussianGradients) gaussianGradients->release();
        if (renderTarget) renderTarget->release();
        commandQueue->release(); delete library; device->release();
    }

    void loadGaussians(const std::vector<Gaussian>& gaussians, float extent) {   // mtl_engine.mm:278-330
        sceneExtent = extent;
        gaussianCount = gaussians.size();
        // the reference constructs every operator with capacity 2,000,000 (mtl_engine.mm:317-327) so that densify never
        // reallocates them; same here (the density controller caps the model at 1,500,000)
        const size_t capacity = std::max<size_t>(gaussianCount, kOperatorCapacity);
        gaussianBuffer = device->newGaussianBuffer(reinterpret_cast<const gs_gaussian*>(gaussians.data()), gaussianCount, capacity);
        gaussianGradients = device->newGradientBuffer(capacity);
        optimizer = new AdamOptimizer(device, library, capacity);
        DensityController::setSceneExtent(extent);
        densityController = new DensityController(device, library);
        tiledRasterizer = new TiledRasterizer(device, library, (uint32_t)capacity);
    }

    // updatePositionBuffer, mtl_engine.mm:1028-1036: float4 position per Gaussian for the viewer's depth sort (device-side copy)
    void updatePositionBuffer() {
        if (!positionBuffer || positionCapacity < gaussianCount) {
            if (positionBuffer) positionBuffer->release();
            positionCapacity = std::max<size_t>(gaussianCount, kOperatorCapacity);
            positionBuffer = device->newBuffer(positionCapacity * 16);
        }
        MTL::gs_check(device->ctx(), gs_gaussians_positions(gaussianBuffer->gaussians(), positionBuffer->devptr()), "updatePositionBuffer");
    }
    // the free-camera branch of MTLEngine::render (mtl_engine.mm:537-548): 45 degree vertical field of view, fx = fy
    static Uniforms viewerUniforms(const simd_float4x4& view, const simd_float4x4& projection, simd_float3 cameraPos, uint32_t width, uint32_t height) {
        Uniforms u{};
        u.viewMatrix = view; u.projectionMatrix = projection;
        for (int c = 0; c < 4; c++) {
            const float* vc = &view.columns[c].x;
            float out[4] = {0, 0, 0, 0};
            for (int k = 0; k < 4; k++) { const float* pk = &projection.columns[k].x; for (int r = 0; r < 4; r++) out[r] += pk[r] * vc[k]; }
            u.viewProjectionMatrix.columns[c] = {out[0], out[1], out[2], out[3]};
        }
        const float fy = (float)height / (2.0f * std::tan(45.0f * 3.14159265358979323846f / 180.0f / 2.0f));
        u.screenSize = {(float)width, (float)height};
        u.focalLength = {fy, fy};
        u.cameraPos = cameraPos;
        return u;
    }
    // the training-view branch (mtl_engine.mm:409-440): the COLMAP pose and intrinsics scaled to the target size
    Uniforms viewerUniformsFor(const TrainingView& v, uint32_t width, uint32_t height) const {
        const float sx = (float)width / (float)v.texture->width(), sy = (float)height / (float)v.texture->height();
        TrainingView s = v;
        s.fx = v.fx * sx; s.fy = v.fy * sy; s.cx = v.cx * sx; s.cy = v.cy * sy;
        const simd_float4x4 V = viewMatrixFromColmap(s), P = projectionFromColmap(s, (float)width, (float)height);
        const simd_float4* R = V.columns;
        const simd_float3 cam = {-(R[0].x * v.tx + R[0].y * v.ty + R[0].z * v.tz), -(R[1].x * v.tx + R[1].y * v.ty + R[1].z * v.tz),
                                 -(R[2].x * v.tx + R[2].y * v.ty + R[2].z * v.tz), 0};
        Uniforms u = viewerUniforms(V, P, cam, width, height);
        u.focalLength = {s.fx, s.fy};
        return u;
    }
    // MTLEngine::render, mtl_engine.mm:401-609, into a headless texture instead of the CAMetalLayer drawable: back-to-front
    // order from GPURadixSort32 (:403), then the instanced-quad draw with blending over a white clear (:575-605)
    void render(const Uniforms& uniforms, MTL::Texture* target) {
        if (!gpuSort) gpuSort = new GPURadixSort32(device, library, kOperatorCapacity);               // mtl_engine.mm:317
        if (!renderPipeline) renderPipeline = new SplatRenderPipeline(device, library, (uint32_t)kOperatorCapacity);
        if (!positionBuffer || positionsStale) { updatePositionBuffer(); positionsStale = false; }
        MTL::Buffer* sortedIndices = gpuSort->sort(commandQueue, positionBuffer, uniforms.cameraPos, gaussianCount);   // mtl_engine.mm:403
        renderPipeline->draw(commandQueue, gaussianBuffer, uniforms, sortedIndices, target);
    }

    static simd_float4x4 viewMatrixFromColmap(const TrainingView& v) {           // mtl_engine.mm:637-660: V = [R(q) | t]
        const float w = v.qw, x = v.qx, y = v.qy, z = v.qz;
        simd_float4x4 m{};
        m.columns[0] = {1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y), 0};
        m.columns[1] = {2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x), 0};
        m.columns[2] = {2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y), 0};
        m.columns[3] = {v.tx, v.ty, v.tz, 1};
        return m;
    }
    static simd_float4x4 projectionFromColmap(const TrainingView& v, float width, float height, float nearZ = 0.1f, float farZ = 1000.0f) {
        simd_float4x4 p{};                                                       // mtl_engine.mm:662-682
        p.columns[0] = {2 * v.fx / width, 0, 0, 0};
        p.columns[1] = {0, 2 * v.fy / height, 0, 0};
        p.columns[2] = {2 * v.cx / width - 1, 2 * v.cy / height - 1, farZ / (farZ - nearZ), 1};
        p.columns[3] = {0, 0, -(farZ * nearZ) / (farZ - nearZ), 0};
        return p;
    }
    static float exponentialLRDecay(float lrInit, float lrFinal, size_t iter, size_t maxIter) {   // mtl_engine.mm:1039-1045
        if (maxIter == 0) return lrInit;
        const float t = std::min(1.0f, (float)iter / (float)maxIter);
        return lrInit * std::pow(lrFinal / lrInit, t);
    }

    TiledUniforms uniformsFor(const TrainingView& v) const {                     // mtl_engine.mm:912-924
        const float W = (float)v.texture->width(), H = (float)v.texture->height();
        TiledUniforms u{};
        u.viewMatrix = viewMatrixFromColmap(v);
        u.projectionMatrix = projectionFromColmap(v, W, H);
        for (int c = 0; c < 4; c++) {                                            // VP = P * V (column-major)
            const float* vc = &u.viewMatrix.columns[c].x;
            float out[4] = {0, 0, 0, 0};
            for (int k = 0; k < 4; k++) { const float* pk = &u.projectionMatrix.columns[k].x; for (int r = 0; r < 4; r++) out[r] += pk[r] * vc[k]; }
            u.viewProjectionMatrix.columns[c] = {out[0], out[1], out[2], out[3]};
        }
        u.screenSize = {W, H};
        u.focalLength = {v.fx, v.fy};
        const simd_float4* R = u.viewMatrix.columns;                             // cameraPos = -R^T t
        u.cameraPos = {-(R[0].x * v.tx + R[0].y * v.ty + R[0].z * v.tz), -(R[1].x * v.tx + R[1].y * v.ty + R[1].z * v.tz),
                       -(R[2].x * v.tx + R[2].y * v.ty + R[2].z * v.tz), 0};
        u.numTilesX = (v.texture->width() + 15) / 16; u.numTilesY = (v.texture->height() + 15) / 16;
        u.numGaussians = (uint32_t)gaussianCount;
        return u;
    }

    // computeLoss, mtl_engine.mm:769-853: (1 - lambdaDSSIM) * L1 + lambdaDSSIM * D-SSIM, mean over pixels (synchronises)
    float computeLoss(MTL::Texture* rendered, MTL::Texture* groundTruth) {
        float loss = 0.0f;
        MTL::gs_check(device->ctx(), gs_loss_compute(device->ctx(), rendered->image(), groundTruth->image(), lambdaDSSIM,
                                                     GS_BWD_QUANTIZE_RGBA8 /* RGBA8Unorm render target */, &loss), "computeLoss");
        return loss;
    }
    // saveTextureToPPM, mtl_engine.mm:19-63 (binary P6 from the RGBA8 contents of a headless device image)
    static bool saveTextureToPPM(MTL::Texture* texture, const std::string& path) {
        const uint32_t w = texture->width(), h = texture->height();
        std::vector<uint8_t> rgba((size_t)w * h * 4), rgb((size_t)w * h * 3);
        texture->getBytesRGBA8(rgba.data());
        for (size_t i = 0; i < (size_t)w * h; i++) { rgb[3 * i] = rgba[4 * i]; rgb[3 * i + 1] = rgba[4 * i + 1]; rgb[3 * i + 2] = rgba[4 * i + 2]; }
        std::ofstream f(path, std::ios::binary);
        if (!f) return false;
        f << "P6\n" << w << " " << h << "\n255\n";
        f.write(reinterpret_cast<const char*>(rgb.data()), (std::streamsize)rgb.size());
        return (bool)f;
    }
    // exportTrainingViews, mtl_engine.mm:1223-1306: forward of every training view at its own resolution -> <dir>/view_NNN.ppm
    size_t exportTrainingViews(const std::string& dir) {
        size_t written = 0;
        for (size_t i = 0; i < trainingImages.size(); i++) {
            const TrainingView& v = trainingImages[i];
            MTL::Texture* t = device->newTexture(v.texture->width(), v.texture->height());
            tiledRasterizer->forward(commandQueue, gaussianBuffer, gaussianCount, uniformsFor(v), t);
            char name[64];
            std::snprintf(name, sizeof name, "/view_%03zu.ppm", i);
            written += saveTextureToPPM(t, dir + name) ? 1 : 0;
            t->release();
        }
        return written;
    }

    // one training iteration on view imgIdx.  The reference computes the logging loss in every step with a full device
    // sync (:991); here that is opt-in (logLossEveryStep) because the loss is off the gradient path; lastLoss keeps it.
    float trainStep(size_t imgIdx, float lrPos, float lrScale, float lrRot, float lrOpacity, float lrSH) {
        const TrainingView& v = trainingImages[imgIdx];
        if (!renderTarget || renderTarget->width() != v.texture->width() || renderTarget->height() != v.texture->height()) {
            if (renderTarget) renderTarget->release();
            renderTarget = device->newTexture(v.texture->width(), v.texture->height());   // mtl_engine.mm:715-727
        }
        const TiledUniforms u = uniformsFor(v);
        tiledRasterizer->forward(commandQueue, gaussianBuffer, gaussianCount, u, renderTarget);                       // :973
        if (logLossEveryStep) lastLoss = computeLoss(renderTarget, v.texture);                                        // :991
        tiledRasterizer->backward(commandQueue, gaussianBuffer, gaussianGradients, gaussianCount, u, renderTarget, v.texture);   // :994
        densityController->accumulateGradients(commandQueue, gaussianGradients, gaussianCount);                       // :998
        optimizer->step(commandQueue, gaussianBuffer, gaussianGradients, lrPos, lrScale, lrRot, lrOpacity, lrSH);     // :1001
        positionsStale = true;                                       // the reference refreshes positionBuffer once per epoch (:1198)
        return lastLoss;
    }

    TrainStats train(size_t numEpochs, const TrainSchedule& S = TrainSchedule()) {   // mtl_engine.mm:1047-1221
        TrainStats st;
        const size_t totalExpectedIters = S.totalIterations ? S.totalIterations : S.startIteration + numEpochs * trainingImages.size();
        size_t totalIterations = S.startIteration;
        device->waitUntilCompleted();
        const auto t0 = std::chrono::high_resolution_clock::now();
        for (size_t epoch = 0; epoch < numEpochs; epoch++) {
            for (size_t imgIdx = 0; imgIdx < trainingImages.size(); imgIdx++) {
                const float lrPos = exponentialLRDecay(S.positionLRInit, S.positionLRFinal, totalIterations, totalExpectedIters);
                trainStep(imgIdx, lrPos, S.scaleLR, S.rotationLR, S.opacityLR, S.shLR);
                totalIterations++;
                const bool shouldDensify = totalIterations > S.densifyFromIter && totalIterations < S.densifyUntilIter &&
                                           totalIterations % S.densityControlInterval == 0;                   // :1112-1114
                if (shouldDensify) {
                    // apply() ends with a read-back of the new count (a stream drain, as the reference's CPU apply needs one): drain
                    // the training steps still queued FIRST, so that densifySeconds is the event's own cost and not theirs
                    device->waitUntilCompleted();
                    const auto d0 = std::chrono::high_resolution_clock::now();
                    const TrainingView& v = trainingImages[imgIdx];
                    const size_t oldCount = gaussianCount;
                    MTL::Buffer* positionBuffer = nullptr;
                    const DensityStats ds = densityController->apply(commandQueue, gaussianBuffer, positionBuffer, nullptr, gaussianCount,
                                                                     totalIterations, 0.0002f, 0.005f, 0.1f * sceneExtent, v.fx,
                                                                     (float)v.texture->width(), 2.0f * sceneExtent);   // :1142-1149
                    if (permuteMoments) {
                        optimizer->permuteStateForDensify(*densityController, gaussianBuffer, oldCount);       // corrected mode (note D), opt-in
                    } else {
                        optimizer->resizeIfNeeded(gaussianCount);                                              // :1160
                        if (gaussianCount > oldCount) optimizer->resetStateForNewGaussians(oldCount);          // :1164-1166 (note D)
                    }
                    st.densifyEvents++; st.pruned += ds.numPruned; st.cloned += ds.numCloned; st.split += ds.numSplit;
                    st.densifySeconds += std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - d0).count();
                }
                if (totalIterations % S.opacityResetInterval == 0 && totalIterations > 0 && totalIterations < S.densifyUntilIter) {
                    MTL::gs_check(device->ctx(), gs_gaussians_clamp_opacity(gaussianBuffer->gaussians(), S.opacityResetValue), "opacity reset");   // :1179-1185
                    optimizer->resetOpacityMomentum();                                                         // :1188
                    optimizer->resetScaleMomentum();
                    densityController->resetAccumulator(gaussianCount);                                        // :1191
                    st.opacityResets++;
                }
            }
        }
        device->waitUntilCompleted();
        st.seconds = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
        st.iterations = totalIterations - S.startIteration;
        st.finalCount = gaussianCount;
        return st;
    }

    // mtl_engine.hpp:51-53, signature and body as the reference: Buffer::contents() is a host mirror refreshed from the device by
    // the call (metal_shim.hpp), valid until the next contents() on the buffer or until densify replaces the buffer
    const Gaussian* getGaussians() const {
        return gaussianBuffer ? (const Gaussian*)gaussianBuffer->contents() : nullptr;
    }
    size_t getGaussianCount() const { return gaussianCount; }                    // mtl_engine.hpp:54-56

    // ---- on-disk state ---------------------------------------------------------------------------------
    // main.mm:483-489, verbatim: PLYExporter::exportPLY(path, engine.getGaussians(), engine.getGaussianCount())
    bool exportPLY(const std::string& path) const {
        return PLYExporter::exportPLY(path, getGaussians(), getGaussianCount());
    }
    // main.mm:383-389 (the --ply path): load_ply + loadGaussians; returns the number of Gaussians loaded (0 = failure)
    size_t loadPLY(const std::string& path, float extent) {
        const std::vector<Gaussian> g = load_ply(path);
        if (g.empty()) return 0;
        releaseModel();
        loadGaussians(g, extent);
        return g.size();
    }
    // Raw checkpoint for resumable runs (no reference counterpart: the reference can only restart from a PLY, which loses the
    // optimizer): parameters, the ten Adam moment buffers in the reference's layout (optimizer.mm:34-60), the timestep and the
    // density controller's accumulators.  Layout: "GSCK", u32 version, u64 count, u32 timestep, u32 0, Gaussian[count],
    // for group in (position, scale, rotation, opacity, sh): m then v as float[count*width], float accum[count], u32 views[count].
    bool saveCheckpoint(const std::string& path) const {
        std::ofstream f(path, std::ios::binary);
        if (!f) return false;
        const uint64_t n = gaussianCount;
        const uint32_t version = 1, t = optimizer->getTimestep(), zero = 0;
        f.write("GSCK", 4); f.write((const char*)&version, 4); f.write((const char*)&n, 8); f.write((const char*)&t, 4); f.write((const char*)&zero, 4);
        const Gaussian* g = getGaussians();
        if (n) f.write((const char*)g, (std::streamsize)(n * sizeof(Gaussian)));
        static const int width[5] = {3, 3, 4, 1, 12};
        std::vector<float> buf;
        for (int group = 0; group < 5; group++)
            for (int mv = 0; mv < 2; mv++) {
                buf.assign(n * width[group], 0.0f);
                if (n) MTL::gs_check(device->ctx(), gs_adam_download(optimizer->handle(), group, mv, buf.data(), (uint32_t)n), "checkpoint: moments");
                f.write((const char*)buf.data(), (std::streamsize)(buf.size() * 4));
            }
        std::vector<float> accum(n, 0.0f); std::vector<uint32_t> views(n, 0);
        if (n) MTL::gs_check(device->ctx(), gs_dens_download(densityController->handle(), accum.data(), views.data(), nullptr, (uint32_t)n), "checkpoint: accumulators");
        f.write((const char*)accum.data(), (std::streamsize)(n * 4)); f.write((const char*)views.data(), (std::streamsize)(n * 4));
        return (bool)f;
    }
    bool loadCheckpoint(const std::string& path, float extent) {
        std::ifstream f(path, std::ios::binary);
        char magic[4]; uint32_t version = 0, t = 0, zero = 0; uint64_t n = 0;
        if (!f || !f.read(magic, 4) || std::memcmp(magic, "GSCK", 4) != 0) return false;
        f.read((char*)&version, 4); f.read((char*)&n, 8); f.read((char*)&t, 4); f.read((char*)&zero, 4);
        if (!f || version != 1 || n > (1ull << 30)) return false;
        std::vector<Gaussian> g(n);
        f.read((char*)g.data(), (std::streamsize)(n * sizeof(Gaussian)));
        if (!f) return false;
        releaseModel();
        loadGaussians(g, extent);
        static const int width[5] = {3, 3, 4, 1, 12};
        std::vector<float> buf;
        for (int group = 0; group < 5; group++)
            for (int mv = 0; mv < 2; mv++) {
                buf.resize(n * width[group]);
                f.read((char*)buf.data(), (std::streamsize)(buf.size() * 4));
                if (!f) return false;
                if (n) MTL::gs_check(device->ctx(), gs_adam_upload(optimizer->handle(), group, mv, buf.data(), (uint32_t)n), "checkpoint: moments");
            }
        MTL::gs_check(device->ctx(), gs_adam_set_timestep(optimizer->handle(), t), "checkpoint: timestep");
        std::vector<float> accum(n); std::vector<uint32_t> views(n);
        f.read((char*)accum.data(), (std::streamsize)(n * 4)); f.read((char*)views.data(), (std::streamsize)(n * 4));
        if (!f) return false;
        if (n) MTL::gs_check(device->ctx(), gs_dens_upload(densityController->handle(), accum.data(), views.data(), (uint32_t)n), "checkpoint: accumulators");
        return true;
    }
    void releaseModel() {                  // drop the model and its operator objects (loadGaussians constructs fresh ones)
        delete tiledRasterizer; delete optimizer; delete densityController;
        tiledRasterizer = nullptr; optimizer = nullptr; densityController = nullptr;
        if (gaussianBuffer) gaussianBuffer->release();
        if (gaussianGradients) gaussianGradients->release();
        gaussianBuffer = nullptr; gaussianGradients = nullptr;
        gaussianCount = 0; positionsStale = true;
    }

    MTL::Device* device = nullptr;
    MTL::CommandQueue* commandQueue = nullptr;
    MTL::Library* library = nullptr;
    MTL::Buffer* gaussianBuffer = nullptr;
    MTL::Buffer* gaussianGradients = nullptr;
    MTL::Texture* renderTarget = nullptr;
    MTL::Buffer* positionBuffer = nullptr;
    size_t positionCapacity = 0;
    bool positionsStale = true;
    GPURadixSort32* gpuSort = nullptr;
    SplatRenderPipeline* renderPipeline = nullptr;
    TiledRasterizer* tiledRasterizer = nullptr;
    AdamOptimizer* optimizer = nullptr;
    DensityController* densityController = nullptr;
    std::vector<TrainingView> trainingImages;
    size_t gaussianCount = 0;
    float sceneExtent = 1.0f;
    float lambdaDSSIM = 0.2f;          // mtl_engine.hpp:147
    bool logLossEveryStep = false;
    bool permuteMoments = false;       // false = the reference's behaviour (moments stay at their old index through densify, SURVEY note D)
    float lastLoss = 0.0f;
};
