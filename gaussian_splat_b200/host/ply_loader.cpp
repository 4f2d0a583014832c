// PLY reader for 3DGS scenes behind the reference's PLYLoader::load interface
// (src/ply_loader.cpp:22-248). Table-driven: the header is compiled once into a list of
// (destination float offset, activation) pairs, the body is read in one go and decoded by
// all host threads.
#include "ply_loader.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <thread>

namespace {

enum Act : int { COPY = 0, SIGMOID = 1, EXP = 2 };
struct Slot { int dst; Act act; };   // dst = float index inside PointData, -1 = ignored property

Slot slot_for(const std::string& name) {
    static const struct { const char* n; int dst; Act act; } table[] = {
        {"x", 0, COPY}, {"y", 1, COPY}, {"z", 2, COPY}, {"nx", 3, COPY}, {"ny", 4, COPY}, {"nz", 5, COPY},
        {"f_dc_0", 6, COPY}, {"f_dc_1", 7, COPY}, {"f_dc_2", 8, COPY}, {"opacity", 9, SIGMOID},
        {"scale_0", 10, EXP}, {"scale_1", 11, EXP}, {"scale_2", 12, EXP},
        {"rot_0", 13, COPY}, {"rot_1", 14, COPY}, {"rot_2", 15, COPY}, {"rot_3", 16, COPY}};
    for (const auto& t : table) if (name == t.n) return {t.dst, t.act};
    if (name.rfind("f_rest_", 0) == 0) {
        char* end = nullptr;
        const long k = std::strtol(name.c_str() + 7, &end, 10);
        if (end != name.c_str() + 7 && k >= 0 && k < 45) return {17 + (int)k, COPY};
    }
    return {-1, COPY};
}

inline float activate(float v, Act a) {
    switch (a) {
        case SIGMOID: return 1.0f / (1.0f + std::exp(-v));   // src/ply_loader.cpp:116
        case EXP: return std::exp(v);                        // :117-119
        default: return v;
    }
}

// DC spherical-harmonic coefficient -> colour (src/ply_loader.cpp:9-20,133-139).
inline void dc_to_rgb(PointData& p) {
    if (p.r == 0.0f && p.g == 0.0f && p.b == 0.0f) return;
    const float c0 = 0.28209479177387814f;
    float* c = &p.r;
    for (int k = 0; k < 3; ++k) c[k] = std::max(0.0f, std::min(1.0f, 0.5f + c0 * c[k]));
}

struct Header { long vertices = 0; bool binary = false; std::vector<Slot> slots; };

bool read_header(std::istream& in, Header& h) {
    std::string line;
    if (!std::getline(in, line) || line != "ply") return false;
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string key; ss >> key;
        if (key == "end_header") break;
        if (key == "format") { std::string f; ss >> f; h.binary = f.rfind("binary_", 0) == 0; }
        else if (key == "element") { std::string what; ss >> what; if (what == "vertex") ss >> h.vertices; }
        else if (key == "property") { std::string type, name; ss >> type >> name; h.slots.push_back(slot_for(name)); }
    }
    return h.vertices > 0 && !h.slots.empty();
}

void decode_rows(const float* rows, size_t first, size_t last, const std::vector<Slot>& slots, PointData* out) {
    const size_t stride = slots.size();
    for (size_t i = first; i < last; ++i) {
        float* dst = reinterpret_cast<float*>(out + i);
        const float* src = rows + i * stride;
        for (size_t j = 0; j < stride; ++j)
            if (slots[j].dst >= 0) dst[slots[j].dst] = activate(src[j], slots[j].act);
        dc_to_rgb(out[i]);
    }
}

}  // namespace

bool PLYLoader::load(const std::string& filepath, std::vector<PointData>& points) {
    std::ifstream in(filepath, std::ios::binary);
    if (!in.is_open()) { std::cerr << "Failed to open PLY file: " << filepath << std::endl; return false; }
    Header h;
    if (!read_header(in, h)) { std::cerr << "Failed to parse PLY header" << std::endl; return false; }
    points.clear();
    if (h.binary) {
        const size_t stride = h.slots.size(), n = (size_t)h.vertices;
        std::vector<float> body(n * stride, 0.0f);
        in.read(reinterpret_cast<char*>(body.data()), (std::streamsize)(body.size() * sizeof(float)));
        points.resize(n);                                  // short files leave defaulted/zero-fed tails, as the reference does
        unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), (unsigned)(n / 65536 + 1)));
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < nt; ++t)
            pool.emplace_back(decode_rows, body.data(), n * t / nt, n * (t + 1) / nt, std::cref(h.slots), points.data());
        for (auto& th : pool) th.join();
    } else {
        std::string line;
        std::vector<float> row(h.slots.size());
        for (long i = 0; i < h.vertices && std::getline(in, line); ++i) {
            std::istringstream ss(line);
            for (auto& v : row) { v = 0.0f; ss >> v; }
            points.emplace_back();
            decode_rows(row.data(), 0, 1, h.slots, &points.back());
        }
    }
    return !points.empty();
}

bool PLYLoader::loadRows(const std::string& filepath, Rows& out) {
    std::ifstream in(filepath, std::ios::binary);
    if (!in.is_open()) { std::cerr << "Failed to open PLY file: " << filepath << std::endl; return false; }
    Header h;
    if (!read_header(in, h)) { std::cerr << "Failed to parse PLY header" << std::endl; return false; }
    out = Rows{};
    out.stride = (unsigned)h.slots.size();
    for (const Slot& s : h.slots) out.slot_dst.push_back(s.dst);
    const size_t stride = h.slots.size();
    if (h.binary) {
        out.n = (size_t)h.vertices;
        out.rows.assign(out.n * stride, 0.0f);             // a short file leaves a zero-fed tail, as in load()
        in.read(reinterpret_cast<char*>(out.rows.data()), (std::streamsize)(out.rows.size() * sizeof(float)));
    } else {
        std::string line;
        for (long i = 0; i < h.vertices && std::getline(in, line); ++i) {
            std::istringstream ss(line);
            const size_t at = out.rows.size();
            out.rows.resize(at + stride, 0.0f);
            for (size_t j = 0; j < stride; ++j) { float v = 0.0f; ss >> v; out.rows[at + j] = v; }
            ++out.n;
        }
    }
    return out.n > 0;
}
