// ply_io.hpp — the on-disk formats either side of the hot path, host C++ only (no device code).
//
//   PLYExporter::exportPLY   ply_exporter.hpp:17-162   binary little-endian PLY with the 3DGS property names
//                                                      (x y z nx ny nz f_dc_0..2 f_rest_0..8 opacity scale_0..2 rot_0..3)
//   load_ply                 ply_loader.cpp:54-290     the loader: linear-vs-log scale detection on the first 1000 vertices
//   detectLinearScales       ply_loader.cpp:17-52      (:17-52), invalid positions skipped, log-scale clamp +-8, quaternion
//                                                      normalisation, raw opacity, SH rest de-interleaved (coefficient-major -> per channel)
// Same function names, arguments and behaviour as the reference.  The reference parses with tinyply 2.3.4 (vendored there,
// not copied here): this file has its own small PLY header parser — properties are located BY NAME in whatever order the
// file declares them, binary little-endian / big-endian and ascii bodies, elements before or after "vertex" skipped.  One
// stated difference: tinyply hands back the file's raw property type and the reference reinterprets it as float; here a
// property declared double / int / uchar ... is converted to float instead of being misread.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#include "gs_host.hpp"

namespace plyio {

struct Property { std::string name; int type = -1; bool is_list = false; int list_count_type = -1; size_t offset = 0; };
struct Element { std::string name; size_t count = 0; std::vector<Property> props; size_t row_bytes = 0; bool has_list = false; };
enum Format { ASCII = 0, BINARY_LE = 1, BINARY_BE = 2 };

// PLY scalar types: index -> byte size
inline int type_index(const std::string& t) {
    static const char* names[][2] = {{"char", "int8"}, {"uchar", "uint8"}, {"short", "int16"}, {"ushort", "uint16"},
                                     {"int", "int32"}, {"uint", "uint32"}, {"float", "float32"}, {"double", "float64"}};
    for (int i = 0; i < 8; i++) if (t == names[i][0] || t == names[i][1]) return i;
    return -1;
}
inline size_t type_size(int t) { static const size_t s[8] = {1, 1, 2, 2, 4, 4, 4, 8}; return s[t]; }

inline double read_scalar(const unsigned char* p, int type, bool swap) {
    unsigned char b[8];
    const size_t n = type_size(type);
    if (swap) for (size_t i = 0; i < n; i++) b[i] = p[n - 1 - i]; else std::memcpy(b, p, n);
    switch (type) {
    case 0: { int8_t v; std::memcpy(&v, b, 1); return v; }
    case 1: { uint8_t v; std::memcpy(&v, b, 1); return v; }
    case 2: { int16_t v; std::memcpy(&v, b, 2); return v; }
    case 3: { uint16_t v; std::memcpy(&v, b, 2); return v; }
    case 4: { int32_t v; std::memcpy(&v, b, 4); return v; }
    case 5: { uint32_t v; std::memcpy(&v, b, 4); return v; }
    case 6: { float v; std::memcpy(&v, b, 4); return v; }
    default: { double v; std::memcpy(&v, b, 8); return v; }
    }
}

struct Header { Format format = BINARY_LE; std::vector<Element> elements; };

inline bool parse_header(std::istream& in, Header& h, std::string& err) {
    std::string line;
    if (!std::getline(in, line) || line.substr(0, 3) != "ply") { err = "not a PLY file"; return false; }
    bool got_format = false;
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ls(line);
        std::string tok;
        if (!(ls >> tok)) continue;
        if (tok == "end_header") {
            if (!got_format) { err = "no format line"; return false; }
            for (Element& e : h.elements) {
                size_t off = 0;
                for (Property& p : e.props) { p.offset = off; if (p.is_list) e.has_list = true; else off += type_size(p.type); }
                e.row_bytes = off;
            }
            return true;
        }
        if (tok == "comment" || tok == "obj_info") continue;
        if (tok == "format") {
            std::string f; ls >> f;
            if (f == "ascii") h.format = ASCII; else if (f == "binary_little_endian") h.format = BINARY_LE;
            else if (f == "binary_big_endian") h.format = BINARY_BE; else { err = "unknown format " + f; return false; }
            got_format = true;
        } else if (tok == "element") {
            Element e; ls >> e.name >> e.count; h.elements.push_back(e);
        } else if (tok == "property") {
            if (h.elements.empty()) { err = "property before element"; return false; }
            Property p; std::string t; ls >> t;
            if (t == "list") { std::string ct, it; ls >> ct >> it >> p.name; p.is_list = true; p.list_count_type = type_index(ct); p.type = type_index(it); }
            else { p.type = type_index(t); ls >> p.name; }
            if (p.type < 0 || (p.is_list && p.list_count_type < 0)) { err = "unknown property type in: " + line; return false; }
            h.elements.back().props.push_back(p);
        }
    }
    err = "missing end_header";
    return false;
}

// reads element `e` (scalar properties only) into rows of doubles-as-floats: out[row * want.size() + k] = property want[k]
inline bool read_vertex_table(std::istream& in, const Header& h, const Element& e, const std::vector<const Property*>& want,
                              std::vector<float>& out, std::string& err) {
    out.assign(e.count * want.size(), 0.0f);
    if (h.format == ASCII) {
        std::vector<int> slot(e.props.size(), -1);
        for (size_t k = 0; k < want.size(); k++) slot[(size_t)(want[k] - e.props.data())] = (int)k;
        for (size_t r = 0; r < e.count; r++)
            for (size_t p = 0; p < e.props.size(); p++) {
                double v;
                if (!(in >> v)) { err = "truncated ascii body"; return false; }
                if (slot[p] >= 0) out[r * want.size() + (size_t)slot[p]] = (float)v;
            }
        return true;
    }
    const bool swap = h.format == BINARY_BE;
    const size_t chunk_rows = 1 << 16;
    std::vector<unsigned char> buf(chunk_rows * e.row_bytes);
    for (size_t r0 = 0; r0 < e.count; r0 += chunk_rows) {
        const size_t rows = std::min(chunk_rows, e.count - r0);
        in.read(reinterpret_cast<char*>(buf.data()), (std::streamsize)(rows * e.row_bytes));
        if ((size_t)in.gcount() != rows * e.row_bytes) { err = "truncated binary body"; return false; }
        for (size_t r = 0; r < rows; r++) {
            const unsigned char* row = buf.data() + r * e.row_bytes;
            for (size_t k = 0; k < want.size(); k++) {
                const Property* p = want[k];
                float v;
                if (p->type == 6 && !swap) std::memcpy(&v, row + p->offset, 4); else v = (float)read_scalar(row + p->offset, p->type, swap);
                out[(r0 + r) * want.size() + k] = v;
            }
        }
    }
    return true;
}

// skip an element that precedes "vertex" in the body
inline bool skip_element(std::istream& in, const Header& h, const Element& e, std::string& err) {
    if (h.format == ASCII) {
        std::string line;
        for (size_t r = 0; r < e.count;) { if (!std::getline(in, line)) { err = "truncated body"; return false; } if (!line.empty() && line != "\r") r++; }
        return true;
    }
    if (!e.has_list) { in.seekg((std::streamoff)(e.count * e.row_bytes), std::ios::cur); return (bool)in; }
    const bool swap = h.format == BINARY_BE;
    for (size_t r = 0; r < e.count; r++)
        for (const Property& p : e.props) {
            if (!p.is_list) { in.seekg((std::streamoff)type_size(p.type), std::ios::cur); continue; }
            unsigned char cb[8];
            in.read(reinterpret_cast<char*>(cb), (std::streamsize)type_size(p.list_count_type));
            const size_t cnt = (size_t)read_scalar(cb, p.list_count_type, swap);
            in.seekg((std::streamoff)(cnt * type_size(p.type)), std::ios::cur);
        }
    if (!in) { err = "truncated body"; return false; }
    return true;
}

}  // namespace plyio

// ply_loader.cpp:17-52: linear if no negative value and every sampled value lies in (0, 1]; log space otherwise
inline bool detectLinearScales(const float* scale_data, size_t count, float& outMinVal, float& outMaxVal) {
    if (count == 0) return false;
    int negativeCount = 0;
    float maxVal = -FLT_MAX, minVal = FLT_MAX;
    const size_t sampleSize = std::min(count, (size_t)1000);
    for (size_t i = 0; i < sampleSize * 3; i++) {
        const float val = scale_data[i];
        if (val < 0) negativeCount++;
        maxVal = std::max(maxVal, val);
        minVal = std::min(minVal, val);
    }
    outMinVal = minVal; outMaxVal = maxVal;
    if (negativeCount > 0) return false;
    return maxVal <= 1.0f && minVal > 0.0f;
}

struct PlyLoadInfo { size_t vertexCount = 0, skipped = 0; bool linearScales = false, hasShRest = false; std::string error; };

// load_ply, ply_loader.cpp:54-290.  Returns an empty vector on any failure (the reference logs and returns what it has).
inline std::vector<Gaussian> load_ply(const std::string& file_path, PlyLoadInfo* info = nullptr) {
    std::vector<Gaussian> gaussians;
    PlyLoadInfo local;
    PlyLoadInfo& I = info ? *info : local;
    I = PlyLoadInfo();
    std::ifstream file(file_path, std::ios::binary);
    if (!file.is_open()) { I.error = "Failed to open file: " + file_path; std::cerr << I.error << std::endl; return gaussians; }
    plyio::Header h;
    if (!plyio::parse_header(file, h, I.error)) { std::cerr << "Error loading ply: " << I.error << std::endl; return gaussians; }
    const plyio::Element* ve = nullptr;
    for (const plyio::Element& e : h.elements) {
        if (e.name == "vertex") { ve = &e; break; }
        if (!plyio::skip_element(file, h, e, I.error)) { std::cerr << "Error loading ply: " << I.error << std::endl; return gaussians; }
    }
    if (!ve) { I.error = "no vertex element"; std::cerr << "Error loading ply: " << I.error << std::endl; return gaussians; }
    if (ve->has_list) { I.error = "list property in the vertex element"; std::cerr << "Error loading ply: " << I.error << std::endl; return gaussians; }
    auto find = [&](const std::string& n) -> const plyio::Property* {
        for (const plyio::Property& p : ve->props) if (p.name == n) return &p;
        return nullptr;
    };
    // required groups, in the order the reference requests them (:66-104)
    static const char* required[] = {"x", "y", "z", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3", "opacity",
                                     "f_dc_0", "f_dc_1", "f_dc_2"};
    static const char* group_of[] = {"positions", "positions", "positions", "scales", "scales", "scales", "rotations", "rotations",
                                     "rotations", "rotations", "opacities", "SH DC coefficients", "SH DC coefficients", "SH DC coefficients"};
    std::vector<const plyio::Property*> want;
    for (int k = 0; k < 14; k++) {
        const plyio::Property* p = find(required[k]);
        if (!p) { I.error = std::string("Failed to load ") + group_of[k]; std::cerr << I.error << std::endl; return gaussians; }
        want.push_back(p);
    }
    bool rest = true;                                                           // all nine or none (:107-119)
    for (int k = 0; k < 9; k++) if (!find("f_rest_" + std::to_string(k))) rest = false;
    if (rest) for (int k = 0; k < 9; k++) want.push_back(find("f_rest_" + std::to_string(k)));
    I.hasShRest = rest;
    std::vector<float> tab;
    if (!plyio::read_vertex_table(file, h, *ve, want, tab, I.error)) { std::cerr << "Error loading ply: " << I.error << std::endl; return gaussians; }
    const size_t n = ve->count, stride = want.size();
    I.vertexCount = n;
    std::vector<float> scales(n * 3);
    for (size_t i = 0; i < n; i++) for (int j = 0; j < 3; j++) scales[i * 3 + j] = tab[i * stride + 3 + j];
    float scaleMin, scaleMax;
    const bool isLinearScale = detectLinearScales(scales.data(), n, scaleMin, scaleMax);   // :139-149
    I.linearScales = isLinearScale;
    gaussians.reserve(n);
    for (size_t i = 0; i < n; i++) {
        const float* r = &tab[i * stride];
        Gaussian g;
        std::memset(&g, 0, sizeof(g));
        g.position = {r[0], r[1], r[2], 0};
        if (std::isnan(r[0]) || std::isnan(r[1]) || std::isnan(r[2]) || std::isinf(r[0]) || std::isinf(r[1]) || std::isinf(r[2]) ||
            std::abs(r[0]) > 1e6 || std::abs(r[1]) > 1e6 || std::abs(r[2]) > 1e6) { I.skipped++; continue; }   // :160-166
        float s[3] = {r[3], r[4], r[5]};
        for (int j = 0; j < 3; j++) {
            if (isLinearScale) s[j] = std::log(std::max(s[j], 1e-8f));          // :174-180
            s[j] = std::clamp(s[j], -8.0f, 8.0f);                              // MAX_LOG_SCALE, :184-187
        }
        g.scale = {s[0], s[1], s[2], 0};
        float qw = r[6], qx = r[7], qy = r[8], qz = r[9];
        const float q_len = std::sqrt(qw * qw + qx * qx + qy * qy + qz * qz);   // :198-203
        if (q_len > 0.0001f) { qw /= q_len; qx /= q_len; qy /= q_len; qz /= q_len; }
        else { qw = 1.0f; qx = 0.0f; qy = 0.0f; qz = 0.0f; }
        g.rotation = {qw, qx, qy, qz};
        g.opacity = r[10];                                                      // raw, :209
        g.sh[0] = r[11]; g.sh[4] = r[12]; g.sh[8] = r[13];                      // :212-214
        if (rest) {                                                             // coefficient-major triples -> per channel, :217-226
            const float* f = r + 14;
            g.sh[1] = f[0]; g.sh[5] = f[1]; g.sh[9] = f[2];
            g.sh[2] = f[3]; g.sh[6] = f[4]; g.sh[10] = f[5];
            g.sh[3] = f[6]; g.sh[7] = f[7]; g.sh[11] = f[8];
        }
        gaussians.push_back(g);
    }
    return gaussians;
}

class PLYExporter {
public:
    // ply_exporter.hpp:17-162.  A Gaussian is written iff its position.x is finite and |x| <= 1e6 (the reference's write loop, :74-77;
    // its header count uses |x| < 1e6, :31-32 — the two disagree only at |x| == 1e6 exactly, where this file follows the write loop
    // so that header and body always agree).
    static bool exportPLY(const std::string& filename, const Gaussian* gaussians, size_t count) {
        std::ofstream file(filename, std::ios::binary);
        if (!file.is_open()) { std::cerr << "Failed to open file for writing: " << filename << std::endl; return false; }
        auto valid = [](const Gaussian& g) { return !(std::isnan(g.position.x) || std::isinf(g.position.x) || std::abs(g.position.x) > 1e6); };
        size_t validCount = 0;
        for (size_t i = 0; i < count; i++) validCount += valid(gaussians[i]) ? 1 : 0;
        file << "ply\nformat binary_little_endian 1.0\nelement vertex " << validCount << "\n";
        static const char* names[] = {"x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"};
        for (const char* nm : names) file << "property float " << nm << "\n";
        for (int i = 0; i < 9; i++) file << "property float f_rest_" << i << "\n";
        static const char* tail[] = {"opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"};
        for (const char* nm : tail) file << "property float " << nm << "\n";
        file << "end_header\n";
        std::vector<float> rows;
        rows.reserve(std::min<size_t>(validCount, 1 << 16) * 26);
        for (size_t i = 0; i < count; i++) {
            const Gaussian& g = gaussians[i];
            if (!valid(g)) continue;
            const float row[26] = {g.position.x, g.position.y, g.position.z, 0.0f, 0.0f, 0.0f, g.sh[0], g.sh[4], g.sh[8],
                                   g.sh[1], g.sh[5], g.sh[9], g.sh[2], g.sh[6], g.sh[10], g.sh[3], g.sh[7], g.sh[11],   // :104-113
                                   g.opacity, g.scale.x, g.scale.y, g.scale.z,
                                   g.rotation.x, g.rotation.y, g.rotation.z, g.rotation.w};                              // w, x, y, z (:131-138)
            rows.insert(rows.end(), row, row + 26);
            if (rows.size() >= (size_t)26 << 16) { file.write(reinterpret_cast<const char*>(rows.data()), (std::streamsize)(rows.size() * 4)); rows.clear(); }
        }
        if (!rows.empty()) file.write(reinterpret_cast<const char*>(rows.data()), (std::streamsize)(rows.size() * 4));
        file.close();
        return (bool)file;
    }
};
