/* <Metal/Metal.hpp> stand-in: just enough of metal-cpp's spelling for the CPU halves of the reference's hot path
 * (shared-storage buffers read through contents()) to compile with g++ (oracle/_ref; TEST INFRASTRUCTURE ONLY). */
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <cstdint>
namespace MTL {
enum ResourceOptions { ResourceStorageModeShared = 0 };
class Buffer {
public:
    explicit Buffer(size_t n) : p_(calloc(n ? n : 1, 1)), n_(n) {}
    void* contents() { return p_; }
    size_t length() const { return n_; }
    void release() { free(p_); delete this; }
private:
    void* p_; size_t n_;
};
class Device { public: Buffer* newBuffer(size_t n, int /*options*/) { return new Buffer(n); } };
class Library {};
class CommandQueue {};
class ComputePipelineState {};
class Texture {};
}  // namespace MTL
