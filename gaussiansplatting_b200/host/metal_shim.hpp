// metal_shim.hpp — the handle types the reference's host classes are written against, backed by the C ABI.
//
// The reference (ctaylo41/GaussianSplatting) spells its device plumbing with metal-cpp: MTL::Device, MTL::CommandQueue,
// MTL::Library, MTL::Buffer, MTL::Texture (mtl_engine.mm:79-86, 278-330, 715-727).  This header provides types with the
// SAME NAMES AND ROLES over libgs_b200.so so that the operator classes in gs_host.hpp keep the reference's method
// signatures verbatim (tiled_rasterizer.hpp:59-75, gpu_sort.hpp:16-77, optimizer.hpp:26-61, density_control.hpp:22-48).
// What differs, because a CUDA device is not unified memory:
//   * a Buffer has a ROLE (raw bytes, the Gaussian model, the gradient buffer): the library keeps its own SoA layout on
//     the device.  contents() (102 call sites in the reference, e.g. getGaussians(), mtl_engine.hpp:51-53) returns a HOST
//     MIRROR in the reference's wire layout, refreshed from the device by every call (a synchronising download); host
//     writes reach the device only through didModifyRange() (metal-cpp's call for managed storage) or upload();
//   * a Texture is an RGBA image on the device (float4 per pixel; RGBA8 transfers convert on the device);
//   * Library carries nothing: the kernels are compiled into libgs_b200.so.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/gs_abi.h"

// ---- <simd/simd.h> stand-ins (size / alignment of Apple's types: float2 8/8, float3 16/16, float4 16/16, float4x4 64/16)
struct alignas(8) simd_float2 { float x, y; };
struct alignas(16) simd_float3 { float x, y, z, _w; };
struct alignas(16) simd_float4 { float x, y, z, w; };
struct alignas(16) simd_float4x4 { simd_float4 columns[4]; };
static_assert(sizeof(simd_float2) == 8 && sizeof(simd_float3) == 16 && sizeof(simd_float4x4) == 64, "simd shim layout");

namespace MTL {

// Error convention: the reference's host methods return void; on a failing device call they print to std::cerr
// and carry on (mtl_engine.mm:178-181, gpu_sort.mm:167-176).  The shells do the same: a non-zero gs_* status is logged with
// gs_last_error()'s text, remembered (lastError()) and the call returns.  Callers that prefer an exception (this repo's own C
// entry points in gs_host_capi.cpp, which turn it into a status code for Python) switch it on with setThrowOnError(true).
inline bool& throwOnErrorFlag() { static bool f = false; return f; }
inline std::string& lastErrorText() { static std::string s; return s; }
inline void setThrowOnError(bool on) { throwOnErrorFlag() = on; }
inline const std::string& lastError() { return lastErrorText(); }
inline bool gs_check(gs_ctx* c, int rc, const char* what) {
    if (rc == GS_OK) return true;
    lastErrorText() = std::string(what) + ": " + gs_last_error(c);
    if (throwOnErrorFlag()) throw std::runtime_error(lastErrorText());
    std::cerr << lastErrorText() << std::endl;
    return false;
}

enum ResourceOptions { ResourceStorageModeShared = 0, ResourceStorageModePrivate = 1 };
enum PixelFormat { PixelFormatRGBA8Unorm = 70, PixelFormatRGBA32Float = 125 };

class Device;

class Library {};          // kernels live in libgs_b200.so; kept so constructors keep the reference's argument list

class CommandQueue {       // one in-order stream per device context (mtl_engine.mm:84)
public:
    explicit CommandQueue(Device* d) : device_(d) {}
    Device* device() const { return device_; }
    void release() { delete this; }
private:
    Device* device_;
};

class Buffer {
public:
    enum Role { Raw, Gaussians, Gradients };
    Role role() const { return role_; }
    size_t length() const { return bytes_; }
    gs_gaussians* gaussians() const { return g_; }
    gs_grads* grads() const { return gr_; }
    void* devptr() const { return raw_; }
    // host <-> device with the reference's wire structs (Gaussian 112 B, GaussianGradients 112 B) or raw bytes
    void upload(const void* host, size_t bytes);
    void download(void* host, size_t bytes) const;
    // MTL::Buffer::contents(): pointer to a host mirror of the buffer in the reference's layout (Gaussian[count] for the
    // model, GaussianGradients[count] for gradients, the bytes of a raw buffer), refreshed from the device on every call.
    void* contents();
    // host edits made through contents() -> device (the whole buffer; Metal's managed-storage call takes a range)
    void didModifyRange(size_t /*location*/ = 0, size_t /*length*/ = 0) { if (!mirror_.empty()) upload(mirror_.data(), mirror_.size()); }
    void release();
private:
    friend class Device;
    Buffer() = default;
    Device* dev_ = nullptr;
    Role role_ = Raw;
    size_t bytes_ = 0;
    void* raw_ = nullptr;
    gs_gaussians* g_ = nullptr;
    gs_grads* gr_ = nullptr;
    bool borrowed_ = false;            // wraps a device pointer owned by somebody else (e.g. a sorter's internal result buffer)
    std::vector<unsigned char> mirror_;
};

class Texture {            // render target / ground-truth image (mtl_engine.mm:715-727, image_loader.mm:24)
public:
    uint32_t width() const { return w_; }
    uint32_t height() const { return h_; }
    gs_image* image() const { return im_; }
    void replaceRegionRGBA8(const uint8_t* rgba);        // texture->replaceRegion(...) of the reference's loader
    void getBytesRGBA8(uint8_t* rgba) const;             // texture->getBytes(...) (saveTextureToPPM, mtl_engine.mm:19-63)
    void release();
private:
    friend class Device;
    Texture() = default;
    Device* dev_ = nullptr;
    uint32_t w_ = 0, h_ = 0;
    gs_image* im_ = nullptr;
};

class Device {
public:
    static Device* CreateSystemDefaultDevice(int ordinal = 0) {      // MTL::CreateSystemDefaultDevice, mtl_engine.mm:80
        gs_ctx* c = nullptr;
        if (gs_ctx_create(ordinal, &c) != GS_OK) {
            std::fprintf(stderr, "no CUDA device: %s\n", gs_last_error(nullptr));   // the reference exits here (:178-181)
            return nullptr;
        }
        Device* d = new Device();
        d->ctx_ = c;
        return d;
    }
    gs_ctx* ctx() const { return ctx_; }
    CommandQueue* newCommandQueue() { return new CommandQueue(this); }
    Library* newDefaultLibrary() { return new Library(); }
    Buffer* newBuffer(size_t bytes, ResourceOptions = ResourceStorageModeShared) {           // raw device bytes
        Buffer* b = new Buffer(); b->dev_ = this; b->bytes_ = bytes;
        gs_check(ctx_, gs_dev_alloc(ctx_, bytes, &b->raw_), "newBuffer");
        return b;
    }
    // a Buffer handle over device memory that somebody else owns (gpu_sort.hpp:19-26,76-77 return the sorter's internal
    // MTL::Buffers: borrowed, valid until the next sort); release() frees the handle only
    Buffer* wrapBuffer(const void* devptr, size_t bytes) {
        Buffer* b = new Buffer(); b->dev_ = this; b->bytes_ = bytes; b->raw_ = const_cast<void*>(devptr); b->borrowed_ = true;
        return b;
    }
    Buffer* newGaussianBuffer(const gs_gaussian* host, size_t count, size_t capacity = 0) {    // mtl_engine.mm:296-300
        Buffer* b = new Buffer(); b->dev_ = this; b->role_ = Buffer::Gaussians; b->bytes_ = count * sizeof(gs_gaussian);
        gs_check(ctx_, gs_gaussians_create(ctx_, (uint32_t)(capacity > count ? capacity : count), &b->g_), "newGaussianBuffer");
        if (host && count) gs_check(ctx_, gs_gaussians_upload(b->g_, host, (uint32_t)count), "upload gaussians");
        return b;
    }
    Buffer* newGradientBuffer(size_t count) {                                                  // mtl_engine.mm:305-309
        Buffer* b = new Buffer(); b->dev_ = this; b->role_ = Buffer::Gradients; b->bytes_ = count * sizeof(gs_gradients);
        gs_check(ctx_, gs_grads_create(ctx_, (uint32_t)count, &b->gr_), "newGradientBuffer");
        return b;
    }
    Texture* newTexture(uint32_t w, uint32_t h, PixelFormat = PixelFormatRGBA8Unorm) {
        Texture* t = new Texture(); t->dev_ = this; t->w_ = w; t->h_ = h;
        gs_check(ctx_, gs_image_create(ctx_, w, h, &t->im_), "newTexture");
        return t;
    }
    void waitUntilCompleted() { gs_check(ctx_, gs_ctx_sync(ctx_), "sync"); }                   // commit(); waitUntilCompleted()
    void release() { gs_ctx_destroy(ctx_); delete this; }
private:
    Device() = default;
    gs_ctx* ctx_ = nullptr;
};

inline void Buffer::upload(const void* host, size_t bytes) {
    gs_ctx* c = dev_->ctx();
    if (role_ == Gaussians) gs_check(c, gs_gaussians_upload(g_, (const gs_gaussian*)host, (uint32_t)(bytes / sizeof(gs_gaussian))), "upload");
    else if (role_ == Gradients) gs_check(c, gs_grads_upload(gr_, (const gs_gradients*)host, (uint32_t)(bytes / sizeof(gs_gradients))), "upload");
    else gs_check(c, gs_dev_upload(c, raw_, host, bytes), "upload");
}
inline void Buffer::download(void* host, size_t bytes) const {
    gs_ctx* c = dev_->ctx();
    if (role_ == Gaussians) gs_check(c, gs_gaussians_download(g_, (gs_gaussian*)host, (uint32_t)(bytes / sizeof(gs_gaussian))), "download");
    else if (role_ == Gradients) gs_check(c, gs_grads_download(gr_, (gs_gradients*)host, (uint32_t)(bytes / sizeof(gs_gradients))), "download");
    else gs_check(c, gs_dev_download(c, host, raw_, bytes), "download");
}
inline void* Buffer::contents() {
    size_t bytes = bytes_;
    if (role_ == Gaussians) bytes = (size_t)gs_gaussians_count(g_) * sizeof(gs_gaussian);
    mirror_.resize(bytes);
    if (bytes) download(mirror_.data(), bytes);
    return mirror_.data();
}
inline void Buffer::release() {
    if (g_) gs_gaussians_destroy(g_);
    if (gr_) gs_grads_destroy(gr_);
    if (raw_ && !borrowed_) gs_dev_free(dev_->ctx(), raw_);
    delete this;
}
inline void Texture::replaceRegionRGBA8(const uint8_t* rgba) {
    gs_check(dev_->ctx(), gs_image_upload_rgba8(im_, rgba), "replaceRegion");
    gs_check(dev_->ctx(), gs_ctx_sync(dev_->ctx()), "sync");
}
inline void Texture::getBytesRGBA8(uint8_t* rgba) const { gs_check(dev_->ctx(), gs_image_download_rgba8(im_, rgba), "getBytes"); }
inline void Texture::release() { gs_image_destroy(im_); delete this; }

}  // namespace MTL
