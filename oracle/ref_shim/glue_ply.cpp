/* compiled together with the reference's ply_loader.cpp + tinyply.cpp (unchanged) by oracle/Makefile: C entry points over the
 * reference's own load_ply (ply_loader.cpp:54-290) and PLYExporter::exportPLY (ply_exporter.hpp:17-162) */
#include "ply_loader.hpp"
#include "ply_exporter.hpp"
#include <cstring>
#include <cstdint>
extern "C" {
int ref_ply_export(const char* path, const void* gaussians112, uint64_t n) {
    return PLYExporter::exportPLY(path, (const Gaussian*)gaussians112, (size_t)n) ? 0 : 1;
}
/* two-call protocol: out == nullptr returns the count */
int64_t ref_ply_load(const char* path, void* out112, uint64_t cap) {
    try {
        std::vector<Gaussian> g = load_ply(path);
        if (out112) memcpy(out112, g.data(), std::min<uint64_t>(cap, g.size()) * sizeof(Gaussian));
        return (int64_t)g.size();
    } catch (...) { return -1; }
}
}
