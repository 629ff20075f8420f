// gs_host_capi.cpp — C entry points of libgs_host.so: the C++ host layer (gs_host.hpp / engine.hpp, the reference's class
// interfaces over libgs_b200.so) driven from bench.py and the tests through ctypes.
#include "engine.hpp"
#include <cstdio>

extern "C" {

typedef struct {
    uint32_t width, height, num_views, epochs;
    float fx, fy, cx, cy, scene_extent;
    uint32_t densify_from, densify_until, densify_interval, opacity_reset_interval;   // 0 = the reference's constants
    uint32_t seed;
    int quantize_rgba8;       // 1 = backward reads the 8-bit-quantised render (what the reference's RGBA8 target does)
    int log_loss_every_step;  // 1 = computeLoss in every trainStep with a device sync, as the reference does (mtl_engine.mm:991)
    const char* export_dir;   // non-NULL: exportTrainingViews(dir) after training (view_NNN.ppm)
    const char* load_checkpoint;   // non-NULL: Engine::loadCheckpoint replaces the model + optimizer state before training
    const char* save_checkpoint;   // non-NULL: Engine::saveCheckpoint after training
    const char* export_ply;        // non-NULL: Engine::exportPLY after training (PLYExporter, main.mm:483-489)
    uint32_t start_iteration, total_iterations;   // TrainSchedule::startIteration / totalIterations (resumed runs)
    int permute_moments;           // 1 = Engine::permuteMoments (corrected optimizer state through densify, SURVEY note D)
} gs_host_train_config;

typedef struct {
    double seconds, densify_seconds;
    uint64_t iterations, densify_events, opacity_resets, pruned, cloned, split;
    uint32_t final_count, adam_timestep;
    uint64_t last_pairs;
    float first_loss, last_loss;   // logging loss of the first / last iteration (0 unless log_loss_every_step)
    uint32_t views_exported;
} gs_host_train_result;

static char g_err[512];
// this file's C entry points report failures as a status code + gs_host_last_error(): they run the shells in exception mode
// (the shells' default is the reference's: log to std::cerr and carry on, metal_shim.hpp)
static const bool g_throw_mode = (MTL::setThrowOnError(true), true);
const char* gs_host_last_error(void) { return g_err; }

// Trains `gaussians` on num_views views (poses: quats[4*v] wxyz + trans[3*v]); the ground truth of view v is the render
// of `gt_gaussians` from the same pose, produced on the device at load time (all datasets are synthetic).  The trained
// Gaussians are written to out (capacity out_cap) when out != NULL.
int gs_host_train(int device, const gs_gaussian* gaussians, uint32_t n, const gs_gaussian* gt_gaussians, uint32_t n_gt,
                  const float* quats, const float* trans, const gs_host_train_config* cfg, gs_host_train_result* res,
                  gs_gaussian* out, uint32_t out_cap) {
    try {
        Engine eng(device);
        std::vector<Gaussian> init(n);
        std::memcpy(init.data(), gaussians, (size_t)n * sizeof(Gaussian));
        // ground truth: loadTrainingData (mtl_engine.mm:630-634) replaced by device renders of a second scene
        {
            std::vector<Gaussian> gtv(n_gt);
            std::memcpy(gtv.data(), gt_gaussians, (size_t)n_gt * sizeof(Gaussian));
            Engine* e = &eng;
            e->loadGaussians(gtv, cfg->scene_extent);
            for (uint32_t v = 0; v < cfg->num_views; v++) {
                TrainingView tv{quats[4 * v], quats[4 * v + 1], quats[4 * v + 2], quats[4 * v + 3], trans[3 * v], trans[3 * v + 1],
                                trans[3 * v + 2], cfg->fx, cfg->fy, cfg->cx, cfg->cy, e->device->newTexture(cfg->width, cfg->height)};
                const TiledUniforms u = e->uniformsFor(tv);
                e->tiledRasterizer->forward(e->commandQueue, e->gaussianBuffer, e->gaussianCount, u, tv.texture);
                e->trainingImages.push_back(tv);
            }
            e->device->waitUntilCompleted();
            // swap in the model to train (fresh operator objects, as loadGaussians constructs them)
            delete e->tiledRasterizer; delete e->optimizer; delete e->densityController;
            e->gaussianBuffer->release(); e->gaussianGradients->release();
            e->gaussianBuffer = nullptr; e->gaussianGradients = nullptr;
            e->tiledRasterizer = nullptr; e->optimizer = nullptr; e->densityController = nullptr;
            e->loadGaussians(init, cfg->scene_extent);
        }
        if (cfg->load_checkpoint && !eng.loadCheckpoint(cfg->load_checkpoint, cfg->scene_extent))
            throw std::runtime_error(std::string("cannot load checkpoint ") + cfg->load_checkpoint);
        eng.tiledRasterizer->setQuantizeRenderedRGBA8(cfg->quantize_rgba8 != 0);
        eng.densityController->setSeed(cfg->seed);
        TrainSchedule S;
        S.startIteration = cfg->start_iteration; S.totalIterations = cfg->total_iterations;
        if (cfg->densify_from) S.densifyFromIter = cfg->densify_from;
        if (cfg->densify_until) S.densifyUntilIter = cfg->densify_until;
        if (cfg->densify_interval) S.densityControlInterval = cfg->densify_interval;
        if (cfg->opacity_reset_interval) S.opacityResetInterval = cfg->opacity_reset_interval;
        eng.logLossEveryStep = cfg->log_loss_every_step != 0;
        eng.permuteMoments = cfg->permute_moments != 0;
        float first_loss = 0.0f;
        if (eng.logLossEveryStep && cfg->epochs) {     // the loss of iteration 1, before any update
            TiledUniforms u0 = eng.uniformsFor(eng.trainingImages[0]);
            MTL::Texture* t0 = eng.device->newTexture(cfg->width, cfg->height);
            eng.tiledRasterizer->forward(eng.commandQueue, eng.gaussianBuffer, eng.gaussianCount, u0, t0);
            first_loss = eng.computeLoss(t0, eng.trainingImages[0].texture);
            t0->release();
        }
        const TrainStats st = eng.train(cfg->epochs, S);
        if (res) {
            res->seconds = st.seconds; res->densify_seconds = st.densifySeconds; res->iterations = st.iterations;
            res->densify_events = st.densifyEvents; res->opacity_resets = st.opacityResets;
            res->pruned = st.pruned; res->cloned = st.cloned; res->split = st.split;
            res->final_count = (uint32_t)st.finalCount; res->adam_timestep = eng.optimizer->getTimestep();
            res->last_pairs = eng.tiledRasterizer->pairCount();
            res->first_loss = first_loss; res->last_loss = eng.lastLoss; res->views_exported = 0;
        }
        if (cfg->save_checkpoint && !eng.saveCheckpoint(cfg->save_checkpoint)) throw std::runtime_error("cannot write checkpoint");
        if (cfg->export_ply && !eng.exportPLY(cfg->export_ply)) throw std::runtime_error("cannot write PLY");
        if (cfg->export_dir) {
            const size_t w = eng.exportTrainingViews(cfg->export_dir);
            if (res) res->views_exported = (uint32_t)w;
        }
        if (out) {
            const Gaussian* g = eng.getGaussians();
            if (g) std::memcpy(out, g, std::min<size_t>(eng.getGaussianCount(), out_cap) * sizeof(Gaussian));
        }
        for (auto& tv : eng.trainingImages) tv.texture->release();
        return 0;
    } catch (const std::exception& e) {
        std::snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

// GPURadixSort64 through the C++ class (host arrays in, host arrays out) — used by the tests of the host layer
// One viewer frame through the C++ host layer: Engine::loadGaussians + Engine::render (MTLEngine::render, mtl_engine.mm:401-609)
// with the training-view uniforms of pose (quat wxyz, trans) at width x height; rgba_out: width*height*4 floats.
int gs_host_render_view(int device, const gs_gaussian* gaussians, uint32_t n, const float quat[4], const float trans[3], float fx, float fy,
                        float cx, float cy, uint32_t width, uint32_t height, int rgba8_target, float* rgba_out, uint64_t* pairs) {
    try {
        Engine eng(device);
        std::vector<Gaussian> init(n);
        std::memcpy(init.data(), gaussians, (size_t)n * sizeof(Gaussian));
        eng.loadGaussians(init, 1.0f);
        MTL::Texture* target = eng.device->newTexture(width, height);
        TrainingView tv{quat[0], quat[1], quat[2], quat[3], trans[0], trans[1], trans[2], fx, fy, cx, cy, target};
        const Uniforms u = eng.viewerUniformsFor(tv, width, height);
        eng.render(u, target);
        if (rgba8_target) { eng.renderPipeline->setRGBA8Target(true); eng.render(u, target); }
        MTL::gs_check(eng.device->ctx(), gs_image_download_rgba32f(target->image(), rgba_out), "download");
        if (pairs) *pairs = eng.renderPipeline->pairCount();
        target->release();
        return 0;
    } catch (const std::exception& e) {
        std::snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

// PLYExporter::exportPLY / load_ply (ply_exporter.hpp:17-162, ply_loader.cpp:54-290): host-only, no device needed
int gs_host_ply_export(const char* path, const gs_gaussian* gaussians, uint64_t n) {
    return PLYExporter::exportPLY(path, reinterpret_cast<const Gaussian*>(gaussians), n) ? 0 : -1;
}
// out may be NULL to query the count; info4 = {vertices in the file, skipped, linear scales detected, has f_rest}
int gs_host_ply_load(const char* path, gs_gaussian* out, uint64_t cap, uint64_t* n_loaded, uint64_t* info4) {
    PlyLoadInfo I;
    const std::vector<Gaussian> g = load_ply(path, &I);
    if (n_loaded) *n_loaded = g.size();
    if (info4) { info4[0] = I.vertexCount; info4[1] = I.skipped; info4[2] = I.linearScales; info4[3] = I.hasShRest; }
    if (!I.error.empty()) { std::snprintf(g_err, sizeof(g_err), "%s", I.error.c_str()); return -1; }
    if (out) std::memcpy(out, g.data(), std::min<uint64_t>(g.size(), cap) * sizeof(Gaussian));
    return 0;
}
// loads a checkpoint into a fresh Engine (device upload of parameters, moments, accumulators) and writes it back from the device
int gs_host_checkpoint_resave(int device, const char* in_path, const char* out_path, float scene_extent) {
    try {
        Engine eng(device);
        if (!eng.loadCheckpoint(in_path, scene_extent)) throw std::runtime_error(std::string("cannot load checkpoint ") + in_path);
        if (!eng.saveCheckpoint(out_path)) throw std::runtime_error(std::string("cannot write checkpoint ") + out_path);
        return 0;
    } catch (const std::exception& e) {
        std::snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}
// load_ply -> Engine::loadGaussians -> Engine::exportPLY: a PLY through the device model and back
int gs_host_ply_device_roundtrip(int device, const char* in_path, const char* out_path) {
    try {
        Engine eng(device);
        if (eng.loadPLY(in_path, 1.0f) == 0) throw std::runtime_error(std::string("cannot load ") + in_path);
        if (!eng.exportPLY(out_path)) throw std::runtime_error(std::string("cannot write ") + out_path);
        return 0;
    } catch (const std::exception& e) {
        std::snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

int gs_host_sort64(int device, const uint64_t* keys, const uint32_t* vals, uint64_t n, uint64_t* keys_out, uint32_t* vals_out) {
    try {
        MTL::Device* dev = MTL::Device::CreateSystemDefaultDevice(device);
        if (!dev) throw std::runtime_error("no CUDA device");
        {
            MTL::CommandQueue* q = dev->newCommandQueue();
            MTL::Library lib;
            MTL::Buffer* k = dev->newBuffer(n * 8 + 8); MTL::Buffer* v = dev->newBuffer(n * 4 + 4);
            if (n) { k->upload(keys, n * 8); v->upload(vals, n * 4); }
            GPURadixSort64 sorter(dev, &lib, n ? n : 1);
            sorter.sort(q, k, v, n);
            dev->waitUntilCompleted();
            if (n) {
                MTL::gs_check(dev->ctx(), gs_dev_download(dev->ctx(), keys_out, sorter.getSortedKeys()->devptr(), n * 8), "download");
                MTL::gs_check(dev->ctx(), gs_dev_download(dev->ctx(), vals_out, sorter.getSortedValues()->devptr(), n * 4), "download");
            }
            k->release(); v->release(); q->release();
        }
        dev->release();
        return 0;
    } catch (const std::exception& e) {
        std::snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

}  // extern "C"
